"""BaseDataLoader (reference: base/base_data_loader.py:6-28): a torch DataLoader that takes an
optional sampler (the DistributedSampler of multi-GPU runs) and remembers its kwargs."""
from torch.utils.data import DataLoader
from torch.utils.data.dataloader import default_collate


class BaseDataLoader(DataLoader):
    def __init__(self, dataset, batch_size, shuffle, num_workers, collate_fn=default_collate, sampler=None,
                 pin_memory=False):
        self.init_kwargs = dict(dataset=dataset, batch_size=batch_size, collate_fn=collate_fn,
                                num_workers=num_workers, pin_memory=pin_memory,
                                shuffle=False if sampler is not None else shuffle)
        if sampler is not None:
            super().__init__(sampler=sampler, **self.init_kwargs)
        else:
            super().__init__(**self.init_kwargs)
