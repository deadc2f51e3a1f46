"""Backend object protocol (reference: deepspeed/comm/backend.py:25).  This framework has one communication backend, so the
module-level functions of `comm.comm` are the implementation and `TorchBackend` (comm/torch.py) is the object form of them
for code that talks to `deepspeed.comm.comm.cdb`."""


class Backend(object):
    """State every backend object carries: its name, this process's place in the world, the groups it created."""

    def __init__(self, name="backend", rank=0, size=1):
        self.name = name
        self.world_rank, self.world_size = rank, size
        self.world_group = None
        self.process_groups = []
        self.initialized = False

    def is_initialized(self):
        return self.initialized

    def init_process_group(self):
        """Subclasses bring their transport up, then call this to mark the object usable."""
        self.initialized = True

    def new_group(self):
        raise NotImplementedError
