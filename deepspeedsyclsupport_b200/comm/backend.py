"""Backend object protocol (reference: deepspeed/comm/backend.py:25).  This framework has one communication backend, so the
module-level functions of `comm.comm` are the implementation and `TorchBackend` (comm/torch.py) is the object form of them
for code that talks to `deepspeed.comm.comm.cdb`."""


class Backend(object):

    def __init__(self, name="backend", rank=0, size=1):
        self.name = name
        self.world_group = None
        self.world_size = size
        self.world_rank = rank
        self.process_groups = []
        self.initialized = False

    def is_initialized(self):
        return self.initialized

    def new_group(self):
        raise NotImplementedError

    def init_process_group(self):
        self.initialized = True
