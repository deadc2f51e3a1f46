"""DeepSpeedDataLoader + RepeatingLoader (parity: deepspeed/runtime/dataloader.py:17-162)."""
from torch.utils.data import DataLoader, RandomSampler
from torch.utils.data.distributed import DistributedSampler

from ..accelerator import get_accelerator


class RepeatingLoader:
    """Wraps an iterator to restart automatically when exhausted."""

    def __init__(self, loader):
        self.loader = loader
        self.data_iter = iter(self.loader)

    def __iter__(self):
        return self

    def __next__(self):
        try:
            batch = next(self.data_iter)
        except StopIteration:
            self.data_iter = iter(self.loader)
            batch = next(self.data_iter)
        return batch


class DeepSpeedDataLoader:

    def __init__(self, dataset, batch_size, pin_memory, local_rank, tput_timer, collate_fn=None, num_local_io_workers=None,
                 data_sampler=None, data_parallel_world_size=None, data_parallel_rank=None, dataloader_drop_last=False,
                 deepspeed_dataloader_config={}):
        self.deepspeed_dataloader_config = deepspeed_dataloader_config
        self.tput_timer = tput_timer
        self.batch_size = batch_size
        self.curriculum_learning_enabled = bool(deepspeed_dataloader_config.get("curriculum_learning_enabled", False))
        if self.curriculum_learning_enabled:
            # curriculum learning: a batch sampler that draws every global batch from the currently eligible data clusters
            from .data_pipeline.data_sampling.data_sampler import DeepSpeedDataSampler
            c = deepspeed_dataloader_config
            data_sampler = DeepSpeedDataSampler(c["data_efficiency_config"], len(dataset), batch_size, data_parallel_rank, data_parallel_world_size,
                                                c.get("data_parallel_group"), c.get("gradient_accumulation_steps", 1), c.get("global_rank", 0),
                                                drop_last=dataloader_drop_last)
        if local_rank >= 0:
            if data_sampler is None:
                data_sampler = DistributedSampler(dataset=dataset, num_replicas=data_parallel_world_size, rank=data_parallel_rank)
            device_count = 1
        else:
            if data_sampler is None:
                data_sampler = RandomSampler(dataset)
            device_count = get_accelerator().device_count()
            batch_size *= device_count
        if num_local_io_workers is None:
            num_local_io_workers = 2 * device_count
        self.num_local_io_workers = num_local_io_workers
        self.data_sampler = data_sampler
        self.dataset = dataset
        self.collate_fn = collate_fn
        self.device_count = device_count
        self.pin_memory = pin_memory and get_accelerator().device_name() == "cuda"
        self.data = None
        self.dataloader_drop_last = dataloader_drop_last
        self.post_process_func = None
        self.len = len(self.data_sampler) // self.batch_size if self.dataloader_drop_last else \
            -(-len(self.data_sampler) // self.batch_size)

    def __iter__(self):
        self._create_dataloader()
        return self

    def __len__(self):
        return self.len

    def __next__(self):
        if self.tput_timer:
            self.tput_timer.start()
        return next(self.data)

    def _create_dataloader(self):
        if self.curriculum_learning_enabled:
            kw = dict(pin_memory=self.pin_memory, batch_sampler=self.data_sampler, num_workers=self.num_local_io_workers)
        else:
            kw = dict(batch_size=self.batch_size, pin_memory=self.pin_memory, sampler=self.data_sampler,
                      num_workers=self.num_local_io_workers, drop_last=self.dataloader_drop_last)
        if self.collate_fn is not None:
            kw["collate_fn"] = self.collate_fn
        self.dataloader = DataLoader(self.dataset, **kw)
        self.data = (x for x in self.dataloader)
        return self.dataloader
