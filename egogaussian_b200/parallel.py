"""View-data-parallel training step plumbing.

The path shards across camera views: every rank holds the full Gaussian parameter set, renders
and back-propagates a different view, and the only exchange is the sum of parameter gradients.
Gradients are written by egs_backward straight into views of ONE flat fp32 buffer, so the
exchange is a single all-reduce (NCCL over NVLink/NVSwitch on GPUs, gloo in the CPU tests); the
replicated fused Adam step then keeps all ranks bit-identical without a broadcast.
"""
from __future__ import annotations

from typing import Dict, List, Sequence, Tuple

import torch


class FlatGrads:
    """One contiguous buffer holding all parameter gradients, with named views into it.

    Order and sizes follow the reference's parameter groups (scene/gaussian_model.py:189-197):
    xyz 3, f_dc+f_rest 3*M, opacity 1, scaling 3, rotation 4 floats per Gaussian; means2D (the
    densification statistic's input, gaussian_model.py:735-740) rides along at the end.
    """

    def __init__(self, P: int, M: int, device, with_means2D: bool = True):
        self.layout: List[Tuple[str, Tuple[int, ...]]] = [
            ("means3D", (P, 3)), ("sh", (P, M, 3)), ("opacity", (P, 1)), ("scales", (P, 3)), ("rotations", (P, 4))]
        if with_means2D:
            self.layout.append(("means2D", (P, 3)))
        sizes = [int(torch.Size(s).numel()) for _, s in self.layout]
        # keep every view 16-byte aligned so the fused Adam kernel can use float4 accesses
        self.offsets, off = [], 0
        for n in sizes:
            self.offsets.append(off)
            off += (n + 3) // 4 * 4
        self.flat = torch.zeros(off, dtype=torch.float32, device=device)
        self.views: Dict[str, torch.Tensor] = {
            name: self.flat[o:o + n].view(shape) for (name, shape), o, n in zip(self.layout, self.offsets, sizes)}

    def as_backward_out(self) -> Dict[str, torch.Tensor]:
        out = dict(self.views)
        out.setdefault("means2D", None)
        return out

    def nbytes(self) -> int:
        return self.flat.numel() * 4


def allreduce_gradients(flat: torch.Tensor, average: bool = False, group=None) -> torch.Tensor:
    """Sum (or mean) the flat gradient buffer over all ranks, in place. No-op without a group."""
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return flat
    dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=group)
    if average:
        flat.div_(dist.get_world_size(group))
    return flat


def views_for_rank(n_views: int, rank: int, world: int) -> Sequence[int]:
    """GPU r of n renders views {r, r+n, ...} of the batch (SURVEY.md 8e partitioning)."""
    return list(range(rank, n_views, world))
