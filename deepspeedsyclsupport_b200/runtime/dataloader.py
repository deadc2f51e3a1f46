"""Data loading glue of the engine: `DeepSpeedDataLoader` (what `engine.deepspeed_io` returns) and `RepeatingLoader`.

`DeepSpeedDataLoader` picks the sampler the job needs -- the curriculum batch sampler when data-efficiency sampling is on,
the caller's sampler, a `DistributedSampler` over the data-parallel group for launcher-started ranks (local_rank >= 0), or
a `RandomSampler` with the batch scaled by the visible device count for single-process use -- and builds the torch
`DataLoader` lazily on `iter()`.  Constructor signature and attributes follow deepspeed/runtime/dataloader.py:17-162.
"""
from torch.utils.data import DataLoader, RandomSampler
from torch.utils.data.distributed import DistributedSampler

from ..accelerator import get_accelerator


class RepeatingLoader:
    """Endless iterator over a loader: starts it again each time it runs out."""

    def __init__(self, loader):
        self.loader = loader
        self.data_iter = iter(loader)

    def __iter__(self):
        return self

    def __next__(self):
        for attempt in (0, 1):
            try:
                return next(self.data_iter)
            except StopIteration:
                if attempt:
                    raise  # an empty loader stays empty
                self.data_iter = iter(self.loader)


class DeepSpeedDataLoader:

    def __init__(self, dataset, batch_size, pin_memory, local_rank, tput_timer, collate_fn=None, num_local_io_workers=None,
                 data_sampler=None, data_parallel_world_size=None, data_parallel_rank=None, dataloader_drop_last=False,
                 deepspeed_dataloader_config={}):
        cfg = self.deepspeed_dataloader_config = deepspeed_dataloader_config
        self.dataset, self.collate_fn, self.tput_timer = dataset, collate_fn, tput_timer
        self.batch_size = batch_size
        self.dataloader_drop_last = dataloader_drop_last
        self.post_process_func = None
        self.data = None
        self.curriculum_learning_enabled = bool(cfg.get("curriculum_learning_enabled", False))
        launched = local_rank >= 0
        self.device_count = 1 if launched else get_accelerator().device_count()
        if self.curriculum_learning_enabled:
            from .data_pipeline.data_sampling.data_sampler import DeepSpeedDataSampler
            data_sampler = DeepSpeedDataSampler(cfg["data_efficiency_config"], len(dataset), batch_size, data_parallel_rank,
                                                data_parallel_world_size, cfg.get("data_parallel_group"),
                                                cfg.get("gradient_accumulation_steps", 1), cfg.get("global_rank", 0),
                                                drop_last=dataloader_drop_last)
        elif data_sampler is None and launched:
            data_sampler = DistributedSampler(dataset=dataset, num_replicas=data_parallel_world_size, rank=data_parallel_rank)
        elif data_sampler is None:
            data_sampler = RandomSampler(dataset)
        self.data_sampler = data_sampler
        self._loader_batch = batch_size * self.device_count  # one process feeding several devices takes a batch for each
        self.num_local_io_workers = 2 * self.device_count if num_local_io_workers is None else num_local_io_workers
        self.pin_memory = bool(pin_memory) and get_accelerator().device_name() == "cuda"
        samples = len(data_sampler)
        self.len = samples // batch_size if dataloader_drop_last else (samples + batch_size - 1) // batch_size

    def __len__(self):
        return self.len

    def __iter__(self):
        self._create_dataloader()
        return self

    def __next__(self):
        if self.tput_timer:
            self.tput_timer.start()
        return next(self.data)

    def _create_dataloader(self):
        options = {"pin_memory": self.pin_memory, "num_workers": self.num_local_io_workers}
        if self.curriculum_learning_enabled:
            options["batch_sampler"] = self.data_sampler  # yields whole micro-batches of sample ids
        else:
            options.update(batch_size=self._loader_batch, sampler=self.data_sampler, drop_last=self.dataloader_drop_last)
        if self.collate_fn is not None:
            options["collate_fn"] = self.collate_fn
        self.dataloader = DataLoader(self.dataset, **options)
        self.data = iter(self.dataloader)
        return self.dataloader
