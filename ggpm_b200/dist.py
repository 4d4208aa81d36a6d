"""Data-parallel plumbing for the encoder: molecules are independent (a batch is a disjoint union of
graphs, ggpm/mol_graph.py:247-251), so ranks encode their own batches with no communication on the
forward/backward path; only the encoder-weight gradients are summed across ranks, as ONE flat bucket
(NCCL all-reduce over NVLink on GPUs; gloo in the CPU tests)."""
from __future__ import annotations

from typing import Iterable, List, Sequence

import torch
import torch.distributed as dist


def shard_batches(n_batches: int, rank: int, world_size: int) -> List[int]:
    """rank r takes batches r, r+W, r+2W, ... of the stream"""
    return list(range(rank, n_batches, world_size))


class GradBucket:
    """Flat fp32 view over the gradients of `params` (registration order; parameters without a gradient,
    e.g. tree_encoder.W_o when only hroot is consumed, contribute zeros)."""

    def __init__(self, params: Iterable[torch.nn.Parameter]):
        self.params = [p for p in params if p.requires_grad]
        seen, uniq = set(), []
        for p in self.params:             # tied embeddings appear once
            if id(p) not in seen:
                seen.add(id(p))
                uniq.append(p)
        self.params = uniq
        self.numel = sum(p.numel() for p in self.params)
        self.flat = None

    def _views(self):
        dev = self.params[0].device
        if self.flat is None or self.flat.device != dev:
            self.flat = torch.zeros(self.numel, dtype=torch.float32, device=dev)
            self.views, off = [], 0
            for p in self.params:
                self.views.append(self.flat[off:off + p.numel()].view_as(p))
                off += p.numel()
        return self.views

    def pack(self) -> torch.Tensor:
        """gradients -> flat bucket: one multi-tensor copy (not one kernel per parameter)"""
        views = self._views()
        have = [(v, p.grad) for v, p in zip(views, self.params) if p.grad is not None]
        none = [v for v, p in zip(views, self.params) if p.grad is None]
        if have:
            torch._foreach_copy_([v for v, _ in have], [g for _, g in have])
        if none:
            torch._foreach_zero_(none)
        return self.flat

    def unpack(self, scale: float = 1.0):
        views = self._views()
        if scale != 1.0:
            self.flat.mul_(scale)
        have = [(p.grad, v) for v, p in zip(views, self.params) if p.grad is not None]
        if have:
            torch._foreach_copy_([g for g, _ in have], [v for _, v in have])
        for v, p in zip(views, self.params):
            if p.grad is None:
                p.grad = v.clone()

    def all_reduce_mean(self, weight: float = 1.0, group=None):
        """sum of (weight * grad) over ranks divided by the sum of weights; weight = molecules in the
        local batch makes uneven shards exact."""
        flat = self.pack()
        multi = dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1
        if weight == 1.0:
            # equal shards: the divisor is the world size, no second collective and no host read-back
            if multi:
                dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=group)
            self.unpack(1.0 / (dist.get_world_size(group) if multi else 1))
            return
        flat.mul_(weight)
        total = torch.tensor([weight], dtype=torch.float64, device=flat.device)
        if multi:
            dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=group)
            dist.all_reduce(total, op=dist.ReduceOp.SUM, group=group)
        self.unpack(1.0 / float(total.item()))

    def all_reduce_sum(self, group=None):
        flat = self.pack()
        if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
            dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=group)
        self.unpack(1.0)
