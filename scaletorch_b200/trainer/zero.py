"""ZeRO stage 1 on top of the flat parameter store: optimiser state and the AdamW update are sharded over the
gradient-synchronisation group, parameters stay replicated.

The reference has no optimiser sharding (its only sharded path is the upstream FSDP2 example).  With the
flat store this is a few lines: after the bucket all-reduce every rank holds the complete reduced gradient, so rank ``r``
of ``N`` updates only the contiguous slice ``[r·chunk, (r+1)·chunk)`` of each segment (dense parameters over the dense
sync group, expert parameters over the expert sync group) and the updated bf16 slices are broadcast back in place.

What it buys on B200: m and v shrink from 8 B/param to 8/N B/param (Qwen3-8B DP8: 64 GB → 8 GB per GPU), and the AdamW
kernel, which sits at the HBM roofline (24 B/param per step), touches 1/N of the elements; the price is one in-place
all-gather of the parameters (2 B/param over NVLink).  Enabled with ``--zero_stage 1``.

``state_dict`` gathers the shards into the same full-size tensors the unsharded optimiser saves, so checkpoints are
interchangeable between the two and across data-parallel sizes.
"""
from __future__ import annotations

import torch
import torch.distributed as dist

from ..ops.optim import fused_adamw_step
from ..runtime.flat_params import ALIGN, FlatParamStore
from .optimizer import FusedAdamW


class _Shard:
    """This rank's slice of one segment ``[lo, hi)`` of the flat buffers, split over ``group``."""

    def __init__(self, lo: int, hi: int, group, device):
        self.lo, self.hi, self.group = lo, hi, group
        self.world = dist.get_world_size(group) if (group is not None and dist.is_initialized()) else 1
        self.rank = dist.get_rank(group) if self.world > 1 else 0
        n = hi - lo
        self.chunk = (n + self.world - 1) // self.world
        self.chunk = (self.chunk + ALIGN - 1) // ALIGN * ALIGN      # keeps every slice aligned for the vector kernels
        self.own = self.bounds(self.rank)
        size = max(0, self.own[1] - self.own[0])
        self.exp_avg = torch.zeros(size, dtype=torch.float32, device=device)
        self.exp_avg_sq = torch.zeros(size, dtype=torch.float32, device=device)

    def bounds(self, r: int) -> tuple[int, int]:
        a = min(self.hi, self.lo + r * self.chunk)
        return a, min(self.hi, a + self.chunk)


class ZeroFusedAdamW(FusedAdamW):
    """ZeRO-1 on c10d collectives: every rank updates its contiguous slice of the dense (and expert) segment and broadcasts the updated parameters; the CPU / gloo counterpart of ``PeerShardedAdamW``."""
    def __init__(self, store: FlatParamStore, dense_group=None, expert_group=None, **kw):
        super().__init__(store, _alloc_state=False, **kw)
        self.shards = [_Shard(0, store.dense_numel, dense_group, store.device)]
        if store.expert_numel > 0:
            self.shards.append(_Shard(store.dense_numel, store.total_numel, expert_group, store.device))

    # ------------------------------------------------------------------------------------------ update
    def _apply_update(self, g: dict, coef, grad_scale: float) -> None:
        st = self.store
        for i, sh in enumerate(self.shards):
            lo, hi = sh.own
            gs = grad_scale * (self.expert_grad_scale if i == 1 else 1.0)
            if hi > lo:
                fused_adamw_step(st.param_flat[lo:hi], st.grad_flat[lo:hi], sh.exp_avg, sh.exp_avg_sq, lr=g["lr"],
                                 betas=g["betas"], eps=g["eps"], weight_decay=g["weight_decay"],
                                 step=max(self.step_count, 1), clip_coef=coef, grad_scale=gs, hyper=self._hyper_dev)
            if sh.world > 1:      # in-place all-gather of the updated parameter slices
                for r in range(sh.world):
                    a, b = sh.bounds(r)
                    if b > a:
                        dist.broadcast(st.param_flat[a:b], dist.get_global_rank(sh.group, r), group=sh.group)

    # ------------------------------------------------------------------------------------------ state
    def _gather(self, which: str) -> torch.Tensor:
        """Full-size (CPU) optimiser state assembled from the shards of every rank."""
        full = torch.zeros(self.store.total_numel, dtype=torch.float32)
        for sh in self.shards:
            mine = getattr(sh, which)
            for r in range(sh.world):
                a, b = sh.bounds(r)
                if b <= a:
                    continue
                if sh.world == 1:
                    full[a:b] = mine.cpu()
                    continue
                buf = mine.clone() if r == sh.rank else torch.empty(b - a, dtype=torch.float32, device=mine.device)
                dist.broadcast(buf, dist.get_global_rank(sh.group, r), group=sh.group)
                full[a:b] = buf.cpu()
        return full

    def state_dict(self):
        return {"step": self.step_count, "exp_avg": self._gather("exp_avg"), "exp_avg_sq": self._gather("exp_avg_sq"),
                "param_groups": [{k: v for k, v in g.items() if k != "params"} for g in self.param_groups]}

    def load_state_dict(self, sd) -> None:
        self.step_count = int(sd["step"])
        for sh in self.shards:
            a, b = sh.own
            if b > a:
                sh.exp_avg.copy_(sd["exp_avg"][a:b])
                sh.exp_avg_sq.copy_(sd["exp_avg_sq"][a:b])
        for g, s in zip(self.param_groups, sd.get("param_groups", [])):
            g.update(s)

    def state_bytes(self) -> int:
        return sum(sh.exp_avg.numel() * 8 for sh in self.shards)


class PeerShardedAdamW(FusedAdamW):
    """ZeRO-1 whose collectives are the framework's own peer-memory kernels -- the default optimiser whenever the
    gradient buffer and the bf16 parameters live in symmetric memory (data / context parallel ranks on one NVLink
    domain, no pipeline split).

    Per step and rank, for the dense segment of ``P`` parameters over ``N`` ranks:

    * buckets are **reduce-scattered** during backward (``stb_reduce_scatter_f32``): the rank that owns a range sums it
      through the switch (``multimem.ld_reduce``) and keeps it; nobody receives gradients it will not use;
    * Σg² of the own shard → ``stb_allreduce_tiny_f32`` over the group → clip coefficient (device scalars);
    * ``stb_adamw_shard_bcast`` updates the shard (m, v exist only for it) and writes the new bf16 parameters into every
      rank's copy (``multimem.st``), bracketed by device barriers.

    Against all-reduce + replicated AdamW this removes ``4P·(N-1)/N`` bytes of NVLink ingress and ``4P`` bytes of HBM
    writes from backward, cuts the optimiser's HBM traffic from ``28P`` to ``28P/N`` bytes, and adds ``2P·(N-1)/N`` bytes
    of NVLink ingress for the parameters.  No c10d call, so the step stays capturable in a CUDA graph.  Expert
    parameters (their own, smaller sync group) keep the all-reduce + replicated update.
    ``state_dict`` gathers the shards into the full-size tensors the unsharded optimiser saves.
    """

    SLOT = 3          # flag slot of the dense group's pages used on the main stream (fused_ops.SLOT_AR)
    SLOT_COMM = 4     # flag slot used on the communication stream (shared with the bucket reduce-scatters: SLOT_DP)

    def __init__(self, store: FlatParamStore, bucket_manager, dense_group, sm_budget: int = 0, **kw):
        super().__init__(store, _alloc_state=False, **kw)
        assert store.param_window is not None and store.grad_window is not None
        self.pw, self.gw = store.param_window, store.grad_window
        self.group = dense_group
        self.world, self.rank = self.pw.g.world, self.pw.g.rank
        n = store.dense_numel
        chunk = (n + self.world - 1) // self.world
        self.chunk = (chunk + ALIGN - 1) // ALIGN * ALIGN
        self.own = self.bounds(self.rank)
        lo, hi = self.own
        dev = store.device
        self.exp_avg = torch.zeros(max(0, hi - lo), dtype=torch.float32, device=dev)
        self.exp_avg_sq = torch.zeros(max(0, hi - lo), dtype=torch.float32, device=dev)
        self.exp_avg_e = self.exp_avg_sq_e = None
        if store.expert_numel > 0:
            self.exp_avg_e = torch.zeros(store.expert_numel, dtype=torch.float32, device=dev)
            self.exp_avg_sq_e = torch.zeros(store.expert_numel, dtype=torch.float32, device=dev)
        self.sm_budget = sm_budget
        for b in bucket_manager.buckets:
            if b.grad_window is not None:
                b.own_range = self.own
        # Deferred parameter gather: the update writes only the local shard; the shards are broadcast one owner after
        # the other on the communication stream (stb_param_bcast) and every decoder layer waits only for the owners of
        # ITS parameters, so the forward pass starts once the first shards are in and the rest of the gather (2 bytes
        # per parameter of NVLink ingress) hides behind it.  Falls back to broadcasting from inside the AdamW kernel
        # when the model has no layer list to gate.
        import os

        self.comm_stream = bucket_manager.comm_stream
        self._events: list = [None] * self.world
        self._stale = False
        # Off by default: measured (profiles/dp_sharded_r2.md) the owner-by-owner gather is slower than broadcasting from
        # inside the AdamW kernel, where all owners send at once (N=2: 29.02 vs 27.96 ms/step, N=8: 29.26 vs 28.80).
        self.deferred = (os.environ.get("SCALETORCH_B200_DP_DEFER_GATHER", "0") == "1" and self.comm_stream is not None
                         and self._install_param_gates(store.module))

    # ------------------------------------------------------------------------------------------ deferred gather
    def _owner_of(self, elem: int) -> int:
        return min(self.world - 1, elem // self.chunk)

    def _install_param_gates(self, root) -> bool:
        """Forward pre-hooks that make the current stream wait for the shards a module's parameters live in.  Gated
        units: the root's direct children, with module lists expanded (decoder layers).  The last element of the last
        module list waits for every shard, because what follows it (final norm, LM head) may read weights without
        going through ``Module.__call__``."""
        units, last_list = [], None
        for child in root.children():
            if isinstance(child, torch.nn.ModuleList):
                elems = [m for m in child]
                units.extend(elems)
                if elems:
                    last_list = elems[-1]
            else:
                units.append(child)
        if last_list is None:
            return False
        dense_hi = self.store.dense_numel
        for m in units:
            offs = [p._stb_slot.offset + p._stb_slot.numel - 1 for p in m.parameters()
                    if hasattr(p, "_stb_slot") and p._stb_slot.offset < dense_hi]
            if not offs:
                continue
            k = self.world - 1 if m is last_list else self._owner_of(max(offs))
            m.register_forward_pre_hook(lambda mod, args, k=k: self.wait_shard(k))
        return True

    def wait_shard(self, k: int) -> None:
        """Orders the current stream after the broadcast of shards ``0..k`` of the pending gather (no host blocking)."""
        ev = self._events[k]
        if ev is not None:
            torch.cuda.current_stream().wait_event(ev)

    def gather_params(self, force: bool = False) -> None:
        """Launches the all-gather of the updated parameter shards on the communication stream (asynchronous)."""
        if not self.deferred or not (self._stale or force):
            return
        from .. import _native as N

        sg = self.pw.g
        main = torch.cuda.current_stream()
        fork = torch.cuda.Event()
        fork.record(main)
        self.comm_stream.wait_event(fork)
        with torch.cuda.stream(self.comm_stream):
            for r in range(self.world):
                a, b = self.bounds(r)
                if b > a:
                    N.call("stb_param_bcast", self.pw.bufs_dev, self.pw.mc_ptr or None, sg.flag_pages_dev, a, b - a, r,
                           self.rank, self.world, self.SLOT_COMM, sg.take_epochs(self.SLOT_COMM, 2), 32, N.stream_ptr())
                ev = torch.cuda.Event()
                ev.record(self.comm_stream)
                self._events[r] = ev
        self._stale = False

    def sync_params(self) -> None:
        """Every rank's parameter copy is complete when the current stream reaches this point (call before reading
        parameters outside a forward pass: checkpoints, evaluation with foreign code)."""
        self.gather_params()
        self.wait_shard(self.world - 1)

    def bounds(self, r: int) -> tuple[int, int]:
        n = self.store.dense_numel
        a = min(n, r * self.chunk)
        return a, min(n, a + self.chunk)

    # ------------------------------------------------------------------------------------------ norm
    def _local_sumsq(self) -> torch.Tensor:
        """Same accounting as :meth:`FusedAdamW._local_sumsq`, computed from the shard: partial sums over this rank's
        range are added up across the dense group by ``stb_allreduce_tiny_f32``; after that every rank holds the Σg²
        of the complete, reduced dense gradient."""
        from ..ops.optim import sum_of_squares
        from ..parallel.process_group import process_group_manager as pgm

        st = self.store
        lo, hi = self.own
        g = st.grad_flat
        tp = pgm.tp_world_size if pgm else 1
        ep = pgm.ep_world_size if pgm else 1
        scratch = self.pw.scratch
        scratch[0:1].copy_(sum_of_squares(g[lo:hi])) if hi > lo else scratch[0:1].zero_()
        n = 1
        rep = None
        if self.norm_group is not None and tp > 1:
            rep = getattr(self, "_rep_ranges", None)
            if rep is None:
                from ..utils.misc import is_tp_sharded_name

                rep = self._rep_ranges = [(max(s.offset, lo), min(s.offset + s.numel, hi)) for s in st.slots
                                          if s.segment == "dense" and not is_tp_sharded_name(s.name)]
                rep = self._rep_ranges = [(a, b) for a, b in rep if b > a]
            if rep:
                scratch[1:2].copy_(torch.stack([g[a:b].pow(2).sum() for a, b in rep]).sum().reshape(1))
            else:
                scratch[1:2].zero_()
            n = 2
        tot = self.pw.all_reduce_scalars_(n)
        total = tot[0:1] / ep
        if n == 2:
            total = total - tot[1:2] * ((1.0 - 1.0 / tp) / ep)
        if st.expert_numel > 0:
            ss_exp = sum_of_squares(g[st.dense_numel:st.total_numel])
            total = total + ss_exp * (self.expert_grad_scale ** 2 / tp)
        return total.clone()

    # ------------------------------------------------------------------------------------------ update
    def _apply_update(self, g: dict, coef, grad_scale: float) -> None:
        from .. import _native as N

        st = self.store
        lo, hi = self.own
        sg = self.pw.g
        b1, b2 = g["betas"]
        grad = st.grad_flat[lo:hi]
        if self.deferred:
            # nobody may still be reading the parameters I am about to overwrite locally: only my own forward/backward
            # read my copy, and they precede this launch in stream order
            if hi > lo:
                fused_adamw_step(st.param_flat[lo:hi], grad, self.exp_avg, self.exp_avg_sq, lr=g["lr"], betas=g["betas"],
                                 eps=g["eps"], weight_decay=g["weight_decay"], step=max(self.step_count, 1),
                                 clip_coef=coef, grad_scale=grad_scale, hyper=self._hyper_dev)
            self._stale = True
            if not torch.cuda.is_current_stream_capturing():
                self.gather_params()      # eager: start the gather now; a captured step gathers at its beginning
        else:
            self._bcast_update(g, coef, grad_scale)
        if st.expert_numel > 0:
            a, b = st.dense_numel, st.total_numel
            fused_adamw_step(st.param_flat[a:b], st.grad_flat[a:b], self.exp_avg_e, self.exp_avg_sq_e, lr=g["lr"],
                             betas=g["betas"], eps=g["eps"], weight_decay=g["weight_decay"], step=max(self.step_count, 1),
                             clip_coef=coef, grad_scale=grad_scale * self.expert_grad_scale, hyper=self._hyper_dev)

    def _bcast_update(self, g: dict, coef, grad_scale: float) -> None:
        from .. import _native as N

        st = self.store
        lo, hi = self.own
        sg = self.pw.g
        b1, b2 = g["betas"]
        grad = st.grad_flat[lo:hi]
        N.call("stb_adamw_shard_bcast", self.pw.bufs_dev, self.pw.mc_ptr or None, sg.flag_pages_dev, lo,
               grad.data_ptr() if hi > lo else None, self.exp_avg.data_ptr() if hi > lo else None,
               self.exp_avg_sq.data_ptr() if hi > lo else None, max(0, hi - lo), float(b1), float(b2), float(g["eps"]),
               float(g["weight_decay"]), N.ptr(coef), float(grad_scale), self._hyper_dev.data_ptr(), self.rank, self.world,
               self.SLOT, sg.take_epochs(self.SLOT, 2), self.sm_budget, N.stream_ptr())

    # ------------------------------------------------------------------------------------------ state
    def _gather(self, mine: torch.Tensor, mine_e) -> torch.Tensor:
        full = torch.zeros(self.store.total_numel, dtype=torch.float32)
        for r in range(self.world):
            a, b = self.bounds(r)
            if b <= a:
                continue
            buf = mine.clone() if r == self.rank else torch.empty(b - a, dtype=torch.float32, device=mine.device)
            dist.broadcast(buf, dist.get_global_rank(self.group, r), group=self.group)
            full[a:b] = buf.cpu()
        if mine_e is not None:
            full[self.store.dense_numel:] = mine_e.cpu()
        return full

    def state_dict(self):
        self.sync_params()          # checkpoints read the model right after this: every copy must be complete
        return {"step": self.step_count, "exp_avg": self._gather(self.exp_avg, self.exp_avg_e),
                "exp_avg_sq": self._gather(self.exp_avg_sq, self.exp_avg_sq_e),
                "param_groups": [{k: v for k, v in g.items() if k != "params"} for g in self.param_groups]}

    def load_state_dict(self, sd) -> None:
        self.step_count = int(sd["step"])
        a, b = self.own
        if b > a:
            self.exp_avg.copy_(sd["exp_avg"][a:b])
            self.exp_avg_sq.copy_(sd["exp_avg_sq"][a:b])
        if self.exp_avg_e is not None:
            self.exp_avg_e.copy_(sd["exp_avg"][self.store.dense_numel:])
            self.exp_avg_sq_e.copy_(sd["exp_avg_sq"][self.store.dense_numel:])
        for g, s in zip(self.param_groups, sd.get("param_groups", [])):
            g.update(s)

    def state_bytes(self) -> int:
        return self.exp_avg.numel() * 8 + (self.exp_avg_e.numel() * 8 if self.exp_avg_e is not None else 0)
