"""Device-resident batch loader: what TensorDataset + DistributedSampler + DataLoader +
`.to(device)` do per step in the reference (train.py:65-74,104-106), with the dataset
uploaded once and each batch gathered by one kernel (b200ds_gather_batch).

The sample ORDER is produced by torch's own sampler objects, so it is identical to the
reference's by construction, including the draws DataLoader makes from the global CPU
generator each epoch (torch/utils/data/dataloader.py:706-710, sampler.py:163).
"""
from __future__ import annotations

import torch
from torch.utils.data import RandomSampler, SequentialSampler

from . import _lib


class DeviceBatchLoader:
    def __init__(self, x: torch.Tensor, y: torch.Tensor, batch_size: int, device, sampler=None, shuffle: bool = False):
        self.lib = _lib.load()
        self.device = torch.device(device)
        self.n, self.F = x.shape
        self.batch_size = int(batch_size)
        self.x = x.detach().to(self.device, torch.float32).contiguous()
        self.y = y.detach().to(self.device, torch.float32).reshape(-1).contiguous()
        self._y_shape_tail = tuple(y.shape[1:])
        if sampler is None:
            sampler = RandomSampler(range(self.n)) if shuffle else SequentialSampler(range(self.n))
        self.sampler = sampler
        self.order = None            # int64 device tensor of this epoch's sample order

    def __len__(self):
        return (len(self.sampler) + self.batch_size - 1) // self.batch_size

    def begin_epoch(self) -> torch.Tensor:
        """Draw this epoch's order (same global-RNG consumption as DataLoader.__iter__)."""
        torch.empty((), dtype=torch.int64).random_()          # DataLoader's _base_seed draw
        idx = torch.tensor(list(iter(self.sampler)), dtype=torch.int64)
        self.order = idx.to(self.device)
        return self.order

    def __iter__(self):
        order = self.begin_epoch()
        n = order.numel()
        B = self.batch_size
        stream = _lib.current_stream()
        for s in range(0, n, B):
            nb = min(B, n - s)
            bx = torch.empty((nb, self.F), dtype=torch.float32, device=self.device)
            by = torch.empty((nb,) + self._y_shape_tail, dtype=torch.float32, device=self.device)
            _lib.check(self.lib.b200ds_gather_batch(self.x.data_ptr(), self.y.data_ptr(), order.data_ptr() + 8 * s, nb,
                                                    self.F, bx.data_ptr(), by.data_ptr(), stream), "gather_batch")
            yield bx, by
