"""Batch sharding of the rollout over the GPUs of one box: one process per GPU, weights replicated,
rank r owns samples [r*B/n, (r+1)*B/n).  The reference itself is single-process (main.py:82); these helpers are
what the build adds.

  * euler / rk4: no forward communication at all.
  * dopri5: torchdiffeq's error norm is an RMS over the WHOLE batch tensor, so the step size is
    shared by the batch.  `shard_solver` installs the hook that all-reduces the controller's four fp64 reduction
    slots {sum of squares, #non-finite, aux, element count} between the norm kernel and the controller kernel;
    every rank then computes sqrt(sum/count) of the GLOBAL batch and takes the step sequence a single process
    would take ("sum" mode, the parity mode).  "max" mode is north_star's wording: every rank uses the largest
    per-rank mean square -- rank-consistent, more conservative, and NOT the single-process sequence.
  * training: one all-reduce(SUM)/n over the flat gradient bucket (about 1.04 M parameters = 4.2 MB).

All collectives go through torch.distributed on the current stream (NCCL on GPUs; gloo in the CPU tests).
"""
from __future__ import annotations

from typing import Iterable, List, Optional, Sequence, Tuple

import torch
import torch.distributed as dist


def shard_range(global_batch: int, rank: int, world: int) -> Tuple[int, int]:
    """[lo, hi) of the samples rank owns; the remainder goes to the low ranks (ragged batches allowed)."""
    if global_batch < 0 or world < 1 or not (0 <= rank < world):
        raise ValueError("shard_range(%d, rank=%d, world=%d)" % (global_batch, rank, world))
    base, rem = divmod(global_batch, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def make_norm_reducer(mode: str = "sum", group=None):
    """reduce_fn(red) for DiffEqSolver.reduce_fn.  red = fp64[4] on the device: {S0, S1, S2, count}.
    sum: all slots summed -> RMS of the global batch.
    max: the mean squares S0/count, S1/count are maxed over ranks and re-expressed over the rank's own count, so the
         controller kernel (which forms sqrt(S/count)) sees the largest per-rank norm; the non-finite counter S1 of the
         error-norm phase stays a sum-like flag because max(>0) is still > 0."""
    if mode not in ("sum", "max"):
        raise ValueError("mode must be 'sum' or 'max'")

    def reduce_sum(red: torch.Tensor) -> None:
        dist.all_reduce(red, op=dist.ReduceOp.SUM, group=group)

    def reduce_max(red: torch.Tensor) -> None:
        cnt = red[3:4].clone()
        ms = red[:3] / cnt
        dist.all_reduce(ms, op=dist.ReduceOp.MAX, group=group)
        red[:3] = ms * cnt

    return reduce_sum if mode == "sum" else reduce_max


def shard_solver(solver, mode: str = "sum", group=None):
    """Make a DiffEqSolver('dopri5') take batch-shared steps across ranks.  No-op for fixed-step methods."""
    if getattr(solver, "ode_method", None) == "dopri5" and dist.is_available() and dist.is_initialized() \
            and dist.get_world_size(group) > 1:
        solver.reduce_fn = make_norm_reducer(mode, group)
        solver.reduce_mode = mode
    return solver


class GradBucket:
    """Flat fp32 bucket(s) over the parameters' gradients, averaged over ranks.

    Plain use: one all-reduce per step (`allreduce()` after the backward).
    `stages`: parameter groups in the order their gradients become FINAL during the backward (for ODEConvGRU: the decoder side
    -- ConvDecode + the solver's ODE function -- then the encoder cell, then ConvEncode).  Each group gets its own flat bucket,
    and its copy-in / all-reduce / copy-out is launched from a post-accumulate-grad hook as soon as the group's last gradient
    has landed, on a side stream that forks from the backward's stream at that point: the decoder side's all-reduce runs under
    the encoder recurrence's backward instead of after it.  `allreduce()` then only reduces what is left and joins the side
    stream.  The launch order is the same on every rank (the backward is the same program), so the collectives match up; the
    fork / join are stream waits, which a CUDA graph capture records as edges."""

    def __init__(self, params: Iterable[torch.nn.Parameter], group=None,
                 stages: Optional[Sequence[Iterable[torch.nn.Parameter]]] = None):
        seen = set()
        self.params: List[torch.nn.Parameter] = []
        for p in params:
            if p.requires_grad and id(p) not in seen:
                seen.add(id(p))
                self.params.append(p)
        self.group = group
        self.flat: dict = {}
        self.stages: List[List[torch.nn.Parameter]] = []
        self._count: List[int] = []
        self._launched: List[bool] = []
        self._side = None
        self._handles = []
        staged = set()
        for st in (stages or []):
            ps = []
            for p in st:
                if p.requires_grad and id(p) in seen and id(p) not in staged:
                    staged.add(id(p))
                    ps.append(p)
            if ps:
                k = len(self.stages)
                self.stages.append(ps)
                self._count.append(0)
                self._launched.append(False)
                for p in ps:
                    self._handles.append(p.register_post_accumulate_grad_hook(lambda _p, k=k: self._on_grad(k)))
        self._rest = [p for p in self.params if id(p) not in staged]

    def _on_grad(self, k: int) -> None:
        self._count[k] += 1
        if self._count[k] == len(self.stages[k]) and not self._launched[k] and dist.is_initialized():
            ps = self.stages[k]
            if ps[0].grad is not None and ps[0].grad.is_cuda:
                if self._side is None:
                    self._side = torch.cuda.Stream(device=ps[0].grad.device)
                self._side.wait_stream(torch.cuda.current_stream())
                with torch.cuda.stream(self._side):
                    self._reduce(ps, k)
            else:
                self._reduce(ps, k)
            self._launched[k] = True

    def _reduce(self, params: List[torch.nn.Parameter], key) -> int:
        ps = [p for p in params if p.grad is not None]
        if not ps:
            return 0
        n = sum(p.grad.numel() for p in ps)
        flat = self.flat.get(key)
        if flat is None or flat.numel() != n or flat.device != ps[0].grad.device:
            flat = torch.empty(n, dtype=torch.float32, device=ps[0].grad.device)
            self.flat[key] = flat
        views = []
        o = 0
        for p in ps:
            views.append(flat[o:o + p.grad.numel()].view_as(p.grad))
            o += p.grad.numel()
        torch._foreach_copy_(views, [p.grad for p in ps])
        world = dist.get_world_size(self.group)
        dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=self.group)
        flat.div_(world)
        torch._foreach_copy_([p.grad for p in ps], views)
        return n * 4

    def allreduce(self) -> int:
        """Averages .grad over the ranks in place (whatever the hooks have not reduced yet); returns the bytes reduced in total."""
        rest = list(self._rest)
        done = 0
        for k, ps in enumerate(self.stages):
            if self._launched[k]:
                done += sum(p.grad.numel() for p in ps if p.grad is not None) * 4
            else:
                rest += ps
        nbytes = self._reduce(rest, "rest") + done
        if self._side is not None and any(self._launched):
            torch.cuda.current_stream().wait_stream(self._side)
        for k in range(len(self.stages)):
            self._count[k] = 0
            self._launched[k] = False
        return nbytes
