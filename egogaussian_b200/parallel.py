"""View-data-parallel training step plumbing.

The path shards across camera views: every rank holds the full Gaussian parameter set, renders
and back-propagates a different view, and the only exchange is the sum of parameter gradients.
Gradients are written by egs_backward straight into views of ONE flat fp32 buffer (FlatGrads). On
the GPUs of one NVLink domain that buffer is peer-mapped and the exchange is one kernel per rank
(PeerExchange: egs_exchange_sum, or egs_exchange_adam_step with the optimiser inside it); a single
library all-reduce (allreduce_gradients: NCCL, gloo in the CPU tests) followed by the replicated fused
Adam step remains for groups that cannot map each other's memory. Either way all ranks stay
bit-identical without a parameter broadcast.
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

    def __init__(self, P: int, M: int, device, with_means2D: bool = True, allocate=None, factored: bool = False,
                 geometry: str = "scale_rot", rgb_slot: torch.Tensor = None):
        """`allocate(n_floats)` supplies the flat buffer (PeerExchange.allocate for peer-mapped memory).

        `factored`: the SH gradient leaves the flat buffer. Per view it has rank one (basis of the view
        direction x dL/dRGB), so backward writes 3 floats per Gaussian (and the camera centre behind them)
        into `rgb_slot`, this view's slot of a gather buffer (ViewBatch; a private tensor when none is
        given), and the exchange rebuilds the summed [P,M,3] tensor in `sh`, a plain local tensor: 17
        instead of 14 + 3 M floats per Gaussian to move."""
        self.factored = bool(factored)
        self.P, self.M = P, M
        self.layout: List[Tuple[str, Tuple[int, ...]]] = [("means3D", (P, 3))]
        if not self.factored:
            self.layout.append(("sh", (P, M, 3)))
        # geometry "cov": the rasterizer ran on cov3D_precomp (the reference's production mode, train.py:49)
        # and returns dL/dcov3D; the chain to scale / rotation is the model's backward (egs_model_backward)
        self.layout.append(("opacity", (P, 1)))
        self.layout += [("scales", (P, 3)), ("rotations", (P, 4))] if geometry == "scale_rot" else [("cov3D", (P, 6))]
        if with_means2D:
            self.layout.append(("means2D", (P, 3)))
        sizes = [int(torch.Size(s).numel()) for _, s in self.layout]
        # keep every view 16-byte aligned so the fused Adam kernel can use float4 accesses
        self.offsets, off = [], 0
        for n in sizes:
            self.offsets.append(off)
            off += (n + 3) // 4 * 4
        self.flat = torch.zeros(off, dtype=torch.float32, device=device) if allocate is None else allocate(off)
        assert self.flat.numel() == off and self.flat.dtype == torch.float32
        self.sizes = sizes
        self.views: Dict[str, torch.Tensor] = {
            name: self.flat[o:o + n].view(shape) for (name, shape), o, n in zip(self.layout, self.offsets, sizes)}
        if self.factored:
            from . import cabi
            n_slot = cabi.drgb_slot_floats(P)
            self.rgb_slot = rgb_slot if rgb_slot is not None else torch.zeros(n_slot, dtype=torch.float32, device=device)
            assert self.rgb_slot.numel() == n_slot and self.rgb_slot.data_ptr() % 16 == 0
            self.views["rgb"] = self.rgb_slot[:3 * P].view(P, 3)      # the slot's rows; the camera centre follows
            self.views["sh"] = torch.zeros((P, M, 3), dtype=torch.float32, device=device)

    def as_backward_out(self) -> Dict[str, torch.Tensor]:
        """`out` of cabi.backward: in factored mode "rgb" replaces "sh" (which only the exchange writes)."""
        out = dict(self.views)
        out.setdefault("means2D", None)
        if self.factored:
            out["sh"] = None
        return out

    def nbytes(self) -> int:
        return self.flat.numel() * 4

    def adam_segments(self, lrs: Dict[str, float]):
        """The parameter groups of the reference's optimiser (scene/gaussian_model.py:189-197) as
        segments of this layout for PeerExchange.adam_step: (begin, end, lr_a, lr_b, period, split) in
        floats. `lrs` has xyz, f_dc, f_rest, opacity, scaling, rotation. The SH rows hold features_dc
        (first 3 floats) and features_rest, which differ in learning rate only."""
        names = {"means3D": "xyz", "opacity": "opacity", "scales": "scaling", "rotations": "rotation"}
        out = []
        for (name, shape), o, n in zip(self.layout, self.offsets, self.sizes):
            end = o + (n + 3) // 4 * 4
            if name == "sh":
                out.append((o, end, lrs["f_dc"], lrs["f_rest"], 3 * shape[1], 3))
            elif name in names:
                out.append((o, end, lrs[names[name]], 0.0, 0, 0))
        return out


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


class PeerExchange:
    """The gradient exchange over peer-mapped memory: one kernel per rank instead of a library
    all-reduce (csrc/exchange.cu; include/egs_rasterizer.h egs_exchange_sum / egs_exchange_adam_step).

    torch.distributed's symmetric memory is the plumbing: it allocates the flat buffers, exchanges the
    handles and maps every rank's buffer into this process. GPUs of one NVLink domain only; raises
    when the buffers cannot be mapped (the caller then stays on `allreduce_gradients`)."""

    def __init__(self, device, group=None, multicast=None):
        """`multicast`: form the sums inside the NVLink switch when the buffers have a multicast
        mapping (NVLS); otherwise, or when False, peer loads and stores. Default: above four ranks,
        where it halves the bytes per link (measured on B200: 2 ranks 0.38 ms peer / 0.65 ms switch,
        4 ranks 0.57 / 0.58, 8 ranks 0.71 / 0.60 for 248 MB; NCCL 0.49 / 0.63 / 0.66)."""
        import torch.distributed as dist
        import torch.distributed._symmetric_memory as symm
        from . import cabi
        self._symm, self._cabi, self._dist = symm, cabi, dist
        self.device = torch.device(device)
        self.group = group if group is not None else dist.group.WORLD
        self.world, self.rank = dist.get_world_size(self.group), dist.get_rank(self.group)
        if self.world > cabi.EGS_MAX_PEERS:
            raise ValueError(f"PeerExchange spans one node: at most {cabi.EGS_MAX_PEERS} ranks")
        self._handles = []
        self.use_multicast = self.world > 4 if multicast is None else bool(multicast)
        self.epoch = 0
        self.signals = self.allocate(cabi.EGS_EXCHANGE_SIGNAL_WORDS)   # float zeros == u32 zeros
        self._signal_ptrs = self.peer_pointers(self.signals)
        self._groups = {}

    def allocate(self, n_floats: int) -> torch.Tensor:
        """A zeroed fp32 buffer that every rank of the group can address (collective call)."""
        t = self._symm.empty(int(n_floats), dtype=torch.float32, device=self.device)
        t.zero_()
        hdl = self._symm.rendezvous(t, self.group)
        self._handles.append((t, hdl))
        torch.cuda.synchronize(self.device)
        self._dist.barrier(self.group)
        return t

    def _handle_of(self, t: torch.Tensor):
        for buf, hdl in self._handles:
            if buf.data_ptr() == t.data_ptr():
                return hdl
        raise ValueError("tensor was not allocated by this PeerExchange")

    def peer_pointers(self, t: torch.Tensor):
        hdl = self._handle_of(t)
        off = t.data_ptr() - int(hdl.buffer_ptrs[self.rank])   # the tensor may sit inside a larger block
        return [int(p) + off for p in hdl.buffer_ptrs]

    def multicast_pointer(self, t: torch.Tensor) -> int:
        """Address of the switch's multicast mapping of `t`, 0 when there is none."""
        hdl = self._handle_of(t)
        if not self.use_multicast or self.world < 2 or not getattr(hdl, "has_multicast_support", lambda *a: False):
            return 0
        base = int(getattr(hdl, "multicast_ptr", 0) or 0)
        return base + (t.data_ptr() - int(hdl.buffer_ptrs[self.rank])) if base else 0

    def _group_for(self, grads: torch.Tensor):
        key = grads.data_ptr()
        if key not in self._groups:
            self._groups[key] = self._cabi.peer_group(self.world, self.rank, self.peer_pointers(grads),
                                                      self._signal_ptrs, self.multicast_pointer(grads))
        return self._groups[key]

    def sum(self, grads: torch.Tensor, average: bool = False, max_ctas: int = 0) -> torch.Tensor:
        """In place on all ranks: grads <- sum (or mean) over ranks. Stream-ordered, no host sync."""
        self.epoch += 1
        self._cabi.exchange_sum(self._group_for(grads), grads.numel(), self.epoch,
                                scale=1.0 / self.world if average else 1.0, max_ctas=max_ctas)
        return grads

    def moment_shards(self, n_floats: int):
        """This rank's zeroed Adam moments for a flat buffer of n_floats (1/world of it)."""
        n = int(self._cabi.load().egs_exchange_shard_floats(n_floats, self.world))
        return (torch.zeros(n, dtype=torch.float32, device=self.device),
                torch.zeros(n, dtype=torch.float32, device=self.device))

    def adam_step(self, grads: torch.Tensor, params: torch.Tensor, exp_avg, exp_avg_sq, segments, step: int,
                  betas=(0.9, 0.999), eps: float = 1e-15, average: bool = False, max_ctas: int = 0) -> None:
        """Sum the gradients, update this rank's slice of `params` and publish it to all ranks."""
        self.epoch += 1
        self._cabi.exchange_adam_step(self._group_for(grads), self.peer_pointers(params), exp_avg, exp_avg_sq,
                                      segments, grads.numel(), self.epoch, step, betas=betas, eps=eps,
                                      scale=1.0 / self.world if average else 1.0, max_ctas=max_ctas,
                                      params_multicast=self.multicast_pointer(params))

    def failed(self) -> bool:
        """True when a kernel gave up waiting for a peer (synchronises)."""
        torch.cuda.synchronize(self.device)
        return bool(self.signals.view(torch.int32)[self._cabi.EGS_EXCHANGE_SIGNAL_WORDS - 1].item())


class ViewBatch:
    """One training batch of `n_views` camera views spread over the ranks (BASELINE configs[3]; SURVEY.md
    8e: rank r renders views r, r + world, ...; n_views = world is the one-view-per-rank step), with ONE
    exchange per batch and the SH gradient in factored form.

    Every rank holds a gather buffer with one slot per view (12 bytes per Gaussian + the camera centre).
    `layouts[j]` is the FlatGrads(factored=True) of the rank's j-th view; `backward_args(j)` are the extra
    arguments of cabi.backward that make its kernel store the view's slot into every rank's gather buffer
    while it runs. `exchange()` then sums the small gradients over all views of all ranks into
    `result.flat` on every rank and rebuilds the summed SH gradient in `result.views["sh"]` from the local
    slots, both in the global view order, so the numbers do not depend on the number of GPUs. With one
    rank (or without torch.distributed) it is the single-GPU accumulation of the batch through the same
    kernel."""

    def __init__(self, n_views: int, P: int, M: int, device, exchange: "PeerExchange" = None,
                 geometry: str = "scale_rot"):
        from . import cabi
        self._cabi = cabi
        self.ex = exchange
        self.world = exchange.world if exchange else 1
        self.rank = exchange.rank if exchange else 0
        if n_views > cabi.EGS_MAX_PEERS or n_views % self.world:
            raise ValueError(f"a batch holds at most {cabi.EGS_MAX_PEERS} views, a multiple of the number of ranks")
        self.n_views, self.device = n_views, torch.device(device)
        self.local_views = list(views_for_rank(n_views, self.rank, self.world))
        alloc = exchange.allocate if exchange else None
        slot = cabi.drgb_slot_floats(P)
        self.gather = (alloc(n_views * slot) if alloc else torch.zeros(n_views * slot, dtype=torch.float32, device=device))
        self.layouts = [FlatGrads(P, M, device, allocate=alloc, factored=True, geometry=geometry,
                                  rgb_slot=self.gather[v * slot:(v + 1) * slot]) for v in self.local_views]
        self.result = self.layouts[0]            # the sums land in the first local view's buffers
        self._slot_ptrs = [self.gather.data_ptr() + 4 * v * slot for v in range(n_views)]
        self._push = []
        if exchange:
            flat_ptrs = [exchange.peer_pointers(l.flat) for l in self.layouts]        # [local j][rank]
            gather_ptrs = exchange.peer_pointers(self.gather)
            mc = exchange.multicast_pointer(self.gather)
            for v in self.local_views:
                self._push.append(dict(rgb_remote=[g + 4 * v * slot for q, g in enumerate(gather_ptrs) if q != self.rank],
                                       rgb_multicast=mc + 4 * v * slot if mc else 0))
            self._signals = None
            self._group = exchange._group_for(self.result.flat)
        else:
            flat_ptrs = [[l.flat.data_ptr()] for l in self.layouts]
            self._push = [dict(rgb_remote=[], rgb_multicast=0) for _ in self.local_views]
            self._signals = torch.zeros(cabi.EGS_EXCHANGE_SIGNAL_WORDS, dtype=torch.int32, device=device)
            self._group = cabi.peer_group(1, 0, [self.result.flat.data_ptr()], [self._signals.data_ptr()])
            self._epoch = 0
        self._view_flat = [flat_ptrs[v // self.world][v % self.world] for v in range(n_views)]
        #: how the small gradients are summed: inside the NVLink switch (one view per rank and a multicast
        #: mapping of the flat buffers) or by loads from the peers
        self.path = ("switch multicast" if (exchange and n_views == self.world and exchange.multicast_pointer(self.result.flat))
                     else "peer loads/stores")

    def backward_args(self, j: int) -> Dict:
        """Keyword arguments for cabi.backward of the rank's j-th view: its output tensors and the
        addresses of its slot on the other ranks."""
        return dict(out=self.layouts[j].as_backward_out(), **self._push[j])

    def exchange(self, means3D: torch.Tensor, sh_degree: int, average: bool = False, max_ctas: int = 0) -> FlatGrads:
        """`means3D`: the positions backward ran on (the same on every rank)."""
        if self.ex:
            self.ex.epoch += 1
            epoch = self.ex.epoch
        else:
            self._epoch += 1
            epoch = self._epoch
        self._cabi.exchange_sum_factored(self._group, self.result.flat.numel(), epoch, means3D=means3D,
                                         drgb_ptrs=self._slot_ptrs, dL_dsh=self.result.views["sh"], sh_degree=sh_degree,
                                         scale=1.0 / self.n_views if average else 1.0, max_ctas=max_ctas,
                                         view_flat_ptrs=self._view_flat)
        return self.result
