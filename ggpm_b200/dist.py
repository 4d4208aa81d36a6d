"""Data-parallel plumbing for the encoder: molecules are independent (a batch is a disjoint union of
graphs, ggpm/mol_graph.py:247-251), so ranks encode their own batches with no communication on the
forward/backward path; only the encoder-weight gradients are summed across ranks (NCCL all-reduce over NVLink on
GPUs; gloo in the CPU tests).

`GradBucket` keeps ONE flat fp32 buffer over the gradients (registration order).  Per step:

  * the gradients autograd produced are gathered into the buffer with one multi-tensor copy per group, and
    `p.grad` is re-pointed at the buffer's views — so the reduced values are read by the optimizer in place and
    there is no copy back;
  * the buffer is reduced in GROUPS, each as one collective on a side stream.  With `split_at` (see
    `for_encoder`) the first group — embeddings, W_c / W_i / W_root, the motif and attachment layers: 3.4 of the
    3.8 M floats at hidden 250 — is complete as soon as the backward pass reaches the atom layer, which is the
    longest stretch of the step; post-accumulate-grad hooks notice that moment and launch the group's all-reduce
    while the atom layer's backward is still running.  Which parameters receive a gradient at all
    (`tree_encoder.W_o` does not under the root-only loss) is learned from the previous step; the first step,
    and any step whose set differs, falls back to reducing at `all_reduce_mean()`;
  * `all_reduce_mean()` reduces what is left (the atom layer's 0.4 M floats) and makes the compute stream wait
    for the side stream.
"""
from __future__ import annotations

from typing import Iterable, List, Optional

import torch
import torch.distributed as dist


def shard_batches(n_batches: int, rank: int, world_size: int) -> List[int]:
    """rank r takes batches r, r+W, r+2W, ... of the stream"""
    return list(range(rank, n_batches, world_size))


def _multi(group=None) -> bool:
    return dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1


class GradBucket:
    """Flat fp32 view over the gradients of `params` (registration order; parameters without a gradient,
    e.g. tree_encoder.W_o when only hroot is consumed, contribute zeros).

    split_at: index (into the de-duplicated parameter list) of the first parameter of the LATE group, i.e. of the
    parameters whose gradients are produced last in the backward pass; None = one group.
    overlap: launch the early group's collective from autograd hooks while the backward pass is still running."""

    def __init__(self, params: Iterable[torch.nn.Parameter], split_at: Optional[int] = None, overlap: bool = True,
                 group=None):
        seen, uniq = set(), []
        for p in params:                  # tied embeddings appear once
            if p.requires_grad and id(p) not in seen:
                seen.add(id(p))
                uniq.append(p)
        self.params = uniq
        self.numel = sum(p.numel() for p in self.params)
        self.flat = None
        self.group = group
        n = len(self.params)
        split_at = n if split_at is None else max(0, min(int(split_at), n))
        self.bounds = [(0, split_at), (split_at, n)] if 0 < split_at < n else [(0, n)]
        self.overlap = overlap and len(self.bounds) > 1
        self._group_of = {}
        for g, (a, b) in enumerate(self.bounds):
            for p in self.params[a:b]:
                self._group_of[id(p)] = g
        self._expected = [None] * len(self.bounds)      # per group: ids that fired last step (None: not learned yet)
        self._fired = [set() for _ in self.bounds]
        self._launched = [False] * len(self.bounds)
        self._works = []
        self._late_hook = False
        self._armed = False
        self._comm = None
        self._hooks = []
        if self.overlap:
            for p in self.params:
                self._hooks.append(p.register_post_accumulate_grad_hook(self._on_grad))

    @classmethod
    def for_encoder(cls, model: torch.nn.Module, extra: Iterable[torch.nn.Parameter] = (), **kw) -> "GradBucket":
        """bucket over model.parameters() (+ `extra`, e.g. a latent head kept beside the encoder) whose late group is
        the atom layer (`graph_encoder.*`, the last gradients of the backward pass, ggpm/encoder.py:140-157 in
        reverse); encoders without an atom layer (MotifEncoder) get one group"""
        named = [(k, p) for k, p in model.named_parameters() if p.requires_grad]
        early = [p for k, p in named if not k.startswith("graph_encoder.")] + list(extra)
        late = [p for k, p in named if k.startswith("graph_encoder.")]
        ids = set()
        n_early = 0
        for p in early:
            if id(p) not in ids:
                ids.add(id(p))
                n_early += 1
        return cls(early + late, split_at=n_early if late else None, **kw)

    def arm(self):
        """call right before loss.backward() of a step that ends in all_reduce_mean() (equal shards): lets the hooks
        launch the early group's collective during the backward pass.  Without it (e.g. a backward pass that is not
        followed by a reduction on every rank) the hooks only record which parameters received gradients."""
        self._armed = True
        self._fired = [set() for _ in self.bounds]       # anything recorded by un-armed backward passes is stale
        self._late_hook = False

    # -- flat buffer ------------------------------------------------------------------------------------
    def _views(self):
        dev = self.params[0].device
        if self.flat is None or self.flat.device != dev:
            self.flat = torch.zeros(self.numel, dtype=torch.float32, device=dev)
            self.views, self.offsets, off = [], [], 0
            for p in self.params:
                self.views.append(self.flat[off:off + p.numel()].view_as(p))
                self.offsets.append(off)
                off += p.numel()
            self.offsets.append(off)
        return self.views

    def _slice(self, g):
        self._views()
        a, b = self.bounds[g]
        return self.flat[self.offsets[a]:self.offsets[b]]

    def _gather_group(self, g):
        """gradients of group g -> their slots in the flat buffer (one multi-tensor copy), p.grad -> the slots"""
        views = self._views()
        a, b = self.bounds[g]
        src, dst, none = [], [], []
        for v, p in zip(views[a:b], self.params[a:b]):
            if p.grad is None:
                none.append(v)
            elif p.grad.data_ptr() != v.data_ptr():
                dst.append(v)
                src.append(p.grad)
        if src:
            torch._foreach_copy_(dst, src)
        if none:
            torch._foreach_zero_(none)
        for v, p in zip(views[a:b], self.params[a:b]):
            p.grad = v

    def pack(self) -> torch.Tensor:
        """all gradients -> flat buffer (p.grad become views of it); returns the buffer"""
        for g in range(len(self.bounds)):
            if not self._launched[g]:
                self._gather_group(g)
        return self.flat

    # -- collectives --------------------------------------------------------------------------------------
    def _reduce(self, tensor, avg: bool):
        """sum (or mean) of `tensor` over the ranks, in place, enqueued behind the CURRENT stream's work"""
        if not _multi(self.group):
            if avg:
                pass                                            # mean over one rank
            return None
        if avg and dist.get_backend(self.group) == "nccl":
            return dist.all_reduce(tensor, op=dist.ReduceOp.AVG, group=self.group, async_op=True)
        w = dist.all_reduce(tensor, op=dist.ReduceOp.SUM, group=self.group, async_op=True)
        if avg:
            w.wait()
            tensor.mul_(1.0 / dist.get_world_size(self.group))
            return None
        return w

    def _launch_group(self, g, avg: bool):
        self._gather_group(g)
        self._launched[g] = True
        sl = self._slice(g)
        if not _multi(self.group):
            return
        if sl.is_cuda and self.overlap:
            if self._comm is None:
                self._comm = torch.cuda.Stream(device=sl.device)
            cur = torch.cuda.current_stream(sl.device)
            self._comm.wait_stream(cur)                       # the gathered gradients
            with torch.cuda.stream(self._comm):
                w = self._reduce(sl, avg)
        else:
            w = self._reduce(sl, avg)
        if w is not None:
            self._works.append(w)

    def _on_grad(self, p):
        g = self._group_of.get(id(p))
        if g is None:
            return
        if self._launched[g]:
            self._late_hook = True                             # a gradient arrived after its group was reduced
            return
        self._fired[g].add(id(p))
        last = len(self.bounds) - 1
        if (self._armed and g != last and self._expected[g] is not None and self._fired[g] >= self._expected[g]
                and _multi(self.group)):
            self._launch_group(g, avg=True)

    def _finish(self):
        for w in self._works:
            w.wait()                                           # current stream waits for the collective
        if self._comm is not None:
            torch.cuda.current_stream(self._comm.device).wait_stream(self._comm)
        self._works = []
        self._armed = False
        late = self._late_hook
        for g in range(len(self.bounds)):
            self._expected[g] = set(self._fired[g]) if not late else None
            self._fired[g] = set()
            self._launched[g] = False
        self._late_hook = False
        if late:
            raise RuntimeError("GradBucket: a gradient arrived after its group had been all-reduced (the set of parameters "
                               "that receive gradients changed between steps); re-run the step — the schedule was reset")

    def all_reduce_mean(self, weight: float = 1.0, group=None):
        """sum of (weight * grad) over ranks divided by the sum of weights; weight = molecules in the
        local batch makes uneven shards exact.  After the call p.grad of every parameter is a view of the flat
        buffer holding the reduced value."""
        if group is not None:
            self.group = group
        if weight == 1.0:
            # equal shards: the divisor is the world size (NCCL's AVG), no second collective and no host read-back
            for g in range(len(self.bounds)):
                if not self._launched[g]:
                    self._launch_group(g, avg=True)
            self._finish()
            return
        if any(self._launched):
            raise RuntimeError("GradBucket: weighted reduction cannot follow an early equal-weight launch")
        flat = self.pack()
        flat.mul_(weight)
        total = torch.tensor([weight], dtype=torch.float64, device=flat.device)
        if _multi(self.group):
            dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=self.group)
            dist.all_reduce(total, op=dist.ReduceOp.SUM, group=self.group)
        flat.mul_(1.0 / float(total.item()))
        self._armed = False
        for g in range(len(self.bounds)):
            self._fired[g] = set()
            self._expected[g] = None                           # weighted steps never launch early

    def all_reduce_sum(self, group=None):
        if group is not None:
            self.group = group
        flat = self.pack()
        if _multi(self.group):
            dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=self.group)
        self._armed = False
        for g in range(len(self.bounds)):
            self._fired[g] = set()

    def unpack(self, scale: float = 1.0):
        """kept for callers of the round-1 API: gradients already live in the buffer (p.grad are views of it)"""
        if scale != 1.0:
            self.flat.mul_(scale)
