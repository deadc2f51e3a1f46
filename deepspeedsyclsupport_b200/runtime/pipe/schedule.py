"""Pipeline schedules as instruction streams; the pipeline engine is an interpreter for them and the tests check them as
plain data.

Timing model.  Stage s of S works on a global clock of ticks t = 0, 1, ...  A tick is a forward slot when (t - s) is
even and a backward slot otherwise, which interleaves neighbours so that what stage s sends in tick t is what stage
s +- 1 receives in tick t + 1:
      forward  slot: micro-batch  (t - s) / 2
      backward slot: micro-batch  (t + s + 1) / 2 - S
Slots whose micro-batch id falls outside [0, M) are bubbles.  2 (M + S - 1) ticks drain the pipeline (1F1B); the last tick
also carries the gradient reductions and the optimizer step.  Inference uses the forward half only with two
ping-pong buffers.
API parity: deepspeed/runtime/pipe/schedule.py (PipeSchedule, InferenceSchedule, TrainSchedule, DataParallelSchedule and
the instruction classes)."""
import abc


def call_to_str(base, *args, **kwargs):
    """'Name(arg, key=value)' rendering used by instruction reprs and LayerSpec."""
    parts = [repr(a) for a in args] + [f"{k}={v!r}" for k, v in kwargs.items()]
    return f"{base}({', '.join(parts)})"


# ---- instructions -----------------------------------------------------------------------------------------------------
class PipeInstruction(object):
    """One engine command.  Keyword arguments become attributes; equality is structural."""

    def __init__(self, **kwargs):
        self.name = type(self).__name__
        self.kwargs = kwargs
        self.__dict__.update(kwargs)

    def __repr__(self):
        return call_to_str(self.name, **self.kwargs)

    def __eq__(self, other):
        return type(self) is type(other) and self.kwargs == other.kwargs

    def __hash__(self):
        return hash((self.name, tuple(sorted(self.kwargs.items()))))


class BufferOpInstruction(PipeInstruction):
    """A command that operates on one pipeline buffer slot."""

    def __init__(self, buffer_id, **kwargs):
        super().__init__(buffer_id=buffer_id, **kwargs)


def _define(name, base, doc):
    cls = type(name, (base, ), {"__doc__": doc})
    cls.__module__ = __name__
    return cls


OptimizerStep = _define("OptimizerStep", PipeInstruction, "Apply the optimizer (after all reductions).")
ReduceGrads = _define("ReduceGrads", PipeInstruction, "Reduce gradients across the data-parallel group.")
ReduceTiedGrads = _define("ReduceTiedGrads", PipeInstruction, "Reduce gradients of weights tied across stages.")
LoadMicroBatch = _define("LoadMicroBatch", BufferOpInstruction, "Load the next micro-batch into buffer_id.")
ForwardPass = _define("ForwardPass", BufferOpInstruction, "Run the stage forward on buffer_id.")
BackwardPass = _define("BackwardPass", BufferOpInstruction, "Run the stage backward on buffer_id.")
SendActivation = _define("SendActivation", BufferOpInstruction, "Send buffer_id's outputs to the next stage.")
RecvActivation = _define("RecvActivation", BufferOpInstruction, "Receive inputs for buffer_id from the previous stage.")
SendGrad = _define("SendGrad", BufferOpInstruction, "Send input gradients of buffer_id to the previous stage.")
RecvGrad = _define("RecvGrad", BufferOpInstruction, "Receive output gradients for buffer_id from the next stage.")


# ---- schedules --------------------------------------------------------------------------------------------------------
class PipeSchedule(abc.ABC):
    """Iterating a schedule yields one list of instructions per tick."""

    def __init__(self, micro_batches, stages, stage_id):
        self.micro_batches = micro_batches
        self.stages = stages
        self.stage_id = stage_id
        self.prev_stage = stage_id - 1
        self.next_stage = stage_id + 1

    @abc.abstractmethod
    def steps(self):
        ...

    def num_pipe_buffers(self):
        return self.micro_batches

    stage = property(lambda self: self.stage_id)
    num_stages = property(lambda self: self.stages)
    num_micro_batches = property(lambda self: self.micro_batches)
    is_first_stage = property(lambda self: self.stage_id == 0)
    is_last_stage = property(lambda self: self.stage_id == self.stages - 1)

    def _valid_micro_batch(self, micro_batch_id):
        return 0 <= micro_batch_id < self.micro_batches

    def _valid_stage(self, stage_id):
        return 0 <= stage_id < self.stages

    def _buffer_idx(self, micro_batch_id):
        assert self._valid_micro_batch(micro_batch_id)
        return micro_batch_id % self.num_pipe_buffers()

    def __iter__(self):
        self.it = None
        return self

    def __next__(self):
        if self.it is None:
            self.it = self.steps()
        return next(self.it)


class InferenceSchedule(PipeSchedule):
    """Forward-only pipelining with two alternating buffers per stage."""

    def num_pipe_buffers(self):
        return 2

    def steps(self):
        has_prev, has_next = self._valid_stage(self.prev_stage), self._valid_stage(self.next_stage)
        even_stage = self.stage_id % 2 == 0
        for t in range(self.micro_batches + self.stages - 1):
            mb = t - self.stage_id
            # even stages receive into buffer t%2 and send the other one; odd stages the opposite, so a sender's
            # buffer always meets the matching receiver buffer
            recv_buf = t % 2 if even_stage else (t + 1) % 2
            send_buf = 1 - recv_buf
            live, prev_live = self._valid_micro_batch(mb), self._valid_micro_batch(mb - 1)
            cmds = []
            if live and (self.is_first_stage or self.is_last_stage):
                cmds.append(LoadMicroBatch(recv_buf))
            send = [SendActivation(send_buf)] if (has_next and prev_live) else []
            recv = [RecvActivation(recv_buf)] if (has_prev and live) else []
            cmds += (send + recv) if even_stage else (recv + send)  # opposite order on neighbours: no deadlock
            if live:
                cmds.append(ForwardPass(recv_buf))
            yield cmds


class TrainSchedule(PipeSchedule):
    """1F1B training schedule (see the timing model in the module docstring)."""

    def num_pipe_buffers(self):
        # a stage holds at most (stages - stage_id) micro-batches in flight; two buffers minimum for send/recv overlap
        return max(2, min(self.stages - self.stage_id, self.micro_batches))

    def _step_to_micro_batch(self, step_id):
        is_forward = (step_id - self.stage_id) % 2 == 0
        if is_forward:
            return (step_id - self.stage_id) // 2, True
        return (step_id + self.stage_id + 1) // 2 - self.stages, False

    def steps(self):
        has_prev, has_next = self._valid_stage(self.prev_stage), self._valid_stage(self.next_stage)
        total = 2 * (self.micro_batches + self.stages - 1)
        last_mb = -1
        for t in range(total):
            mb, fwd = self._step_to_micro_batch(t)
            live, last_live = self._valid_micro_batch(mb), self._valid_micro_batch(last_mb)
            buf = self._buffer_idx(mb) if live else None
            last_buf = self._buffer_idx(last_mb) if last_live else None
            cmds = []
            # communication first: hand over the result of the previous tick, then fetch this tick's input
            if fwd:
                if last_live and has_prev:
                    cmds.append(SendGrad(last_buf))
                if live and has_prev:
                    cmds.append(RecvActivation(buf))
            else:
                if live and has_next:
                    cmds.append(RecvGrad(buf))
                if last_live and has_next:
                    cmds.append(SendActivation(last_buf))
            if live:
                if fwd and (self.is_first_stage or self.is_last_stage):
                    cmds.append(LoadMicroBatch(buf))
                cmds.append(ForwardPass(buf) if fwd else BackwardPass(buf))
            if t == total - 1:
                cmds += [ReduceTiedGrads(), ReduceGrads(), OptimizerStep()]
            last_mb = mb
            yield cmds


class DataParallelSchedule(PipeSchedule):
    """Gradient accumulation without pipelining, written as a schedule (one stage, one buffer)."""

    def num_pipe_buffers(self):
        return 1

    def steps(self):
        for mb in range(self.micro_batches):
            cmds = [LoadMicroBatch(buffer_id=0), ForwardPass(buffer_id=0), BackwardPass(buffer_id=0)]
            if mb == self.micro_batches - 1:
                cmds += [ReduceGrads(), OptimizerStep()]
            yield cmds
