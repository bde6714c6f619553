"""Native (C++) micro-batch producer over a flat token array (``csrc/runtime/token_loader.cpp``) and its pure-Python
oracle.

Worker threads inside ``libscaletorch_b200.so`` cut ``(input, target)`` windows out of the token array, apply the
context-parallel slice (contiguous or zig-zag) and write them into a ring of pinned host buffers owned by this
object; the training loop only waits for a slot and issues the H2D copy.  Reference: torch ``DataLoader`` +
``DistributedSampler`` + Python collate with a one-batch prefetch (``scaletorch/data/dataloader.py:107-292``) over the
``concat_chunk`` dataset (``scaletorch/data/dataset.py:63-86``).

``NativeTokenLoader`` yields the same dict as ``MicroBatchDataLoader`` (``input_ids``, ``target_ids``, ``position_ids``,
``hidden_states``), so it drops into ``train_step``.  ``reference_batches`` re-implements the sampling in plain Python /
torch; tests compare the two bit for bit.
"""
from __future__ import annotations

import ctypes

import torch

from .. import _native as N

_MASK = (1 << 64) - 1


def _splitmix64(s: int) -> tuple[int, int]:
    s = (s + 0x9E3779B97F4A7C15) & _MASK
    z = s
    z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & _MASK
    z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & _MASK
    return s, z ^ (z >> 31)


def epoch_permutation(num_samples: int, seed: int, epoch: int, shuffle: bool) -> list[int]:
    """The permutation the native loader uses for ``epoch`` (Fisher-Yates driven by splitmix64)."""
    perm = list(range(num_samples))
    if shuffle:
        s = (seed + epoch * 0x632BE59BD9B4E019) & _MASK
        for i in range(num_samples - 1, 0, -1):
            s, r = _splitmix64(s)
            j = r % (i + 1)
            perm[i], perm[j] = perm[j], perm[i]
    return perm


def _cp_slice(sample: torch.Tensor, S: int, cp_rank: int, cp_size: int, zigzag: bool, shift: int) -> torch.Tensor:
    s_local = S // cp_size
    if not zigzag:
        lo = cp_rank * s_local + shift
        return sample[lo:lo + s_local]
    c = s_local // 2
    a, b = cp_rank * c + shift, (2 * cp_size - 1 - cp_rank) * c + shift
    return torch.cat([sample[a:a + c], sample[b:b + c]])


def reference_batches(tokens: torch.Tensor, seq_len: int, mbs: int, n_batches: int, cp_rank: int = 0, cp_size: int = 1,
                      zigzag: bool = False, data_rank: int = 0, data_world: int = 1, shuffle: bool = True, seed: int = 0):
    """Pure-Python oracle: the first ``n_batches`` ``(input_ids, target_ids)`` pairs the native loader produces."""
    S = seq_len
    num_samples = tokens.numel() // (S + 1)
    per_epoch = (num_samples // data_world) // mbs
    out, perms = [], {}
    for b in range(n_batches):
        epoch, in_epoch = divmod(b, per_epoch)
        if epoch not in perms:
            perms[epoch] = epoch_permutation(num_samples, seed, epoch, shuffle)
        xs, ys = [], []
        for m in range(mbs):
            w = perms[epoch][data_rank + (in_epoch * mbs + m) * data_world]
            sample = tokens[w * (S + 1):(w + 1) * (S + 1)].to(torch.int64)
            xs.append(_cp_slice(sample, S, cp_rank, cp_size, zigzag, 0))
            ys.append(_cp_slice(sample, S, cp_rank, cp_size, zigzag, 1))
        out.append((torch.stack(xs), torch.stack(ys)))
    return out


_P, _I, _L, _U64 = ctypes.c_void_p, ctypes.c_int, ctypes.c_long, ctypes.c_uint64
N.register("stb_loader_create", [_P, _L, _I, _I, _I, _I, _I, _I, _I, _I, _U64, _I, _P, _I], restype=_P)
N.register("stb_loader_batches_per_epoch", [_P], restype=_L)
N.register("stb_loader_num_samples", [_P], restype=_L)
N.register("stb_loader_next", [_P, ctypes.POINTER(_L)], restype=_I)
N.register("stb_loader_release", [_P, _I], restype=_I)
N.register("stb_loader_destroy", [_P], restype=None)


def native_loader_available() -> bool:
    """True when the in-tree library exports the C++ token loader (``stb_loader_*``)."""
    lib = N.load()
    return lib is not None and hasattr(lib, "stb_loader_create")


class NativeTokenLoader:
    """Iterator of micro-batches produced by the C++ worker threads.

    ``tokens``: 1-D int32 tensor (kept alive by this object).  The ring of ``n_slots`` host buffers is pinned when CUDA
    is available.  ``__next__`` returns views into the ring slot of the batch; the slot is handed back to the workers
    on the following ``__next__`` (or ``close``), i.e. after the caller has enqueued its H2D copies -- pass
    ``device=...`` to let the loader do those copies itself and return device tensors.
    """

    def __init__(self, tokens: torch.Tensor, seq_len: int, micro_batch_size: int, cp_rank: int = 0, cp_size: int = 1,
                 zigzag: bool = False, data_rank: int = 0, data_world: int = 1, shuffle: bool = True, seed: int = 0,
                 n_slots: int = 4, n_threads: int = 2, gradient_accumulation_steps: int = 1, device=None):
        lib = N.load(required=True)
        if not hasattr(lib, "stb_loader_create"):
            raise RuntimeError("libscaletorch_b200.so was built without csrc/runtime/token_loader.cpp")
        self.tokens = tokens.to(torch.int32).contiguous()
        self.seq_len, self.mbs = seq_len, micro_batch_size
        self.cp_rank, self.cp_size, self.zigzag = cp_rank, cp_size, bool(zigzag)
        self.s_local = seq_len // cp_size
        self.gradient_accumulation_steps = gradient_accumulation_steps
        self.device = torch.device(device) if device is not None else None
        pin = torch.cuda.is_available()
        self.ring = torch.zeros((n_slots, 2, micro_batch_size, self.s_local), dtype=torch.int64, pin_memory=pin)
        self._lib = lib
        self._h = lib.stb_loader_create(self.tokens.data_ptr(), self.tokens.numel(), seq_len, micro_batch_size, cp_rank,
                                        cp_size, int(self.zigzag), data_rank, data_world, int(shuffle), seed & _MASK,
                                        n_slots, self.ring.data_ptr(), n_threads)
        if not self._h:
            raise ValueError("stb_loader_create rejected the configuration (sequence length not divisible by the context-"
                             "parallel layout, or fewer samples than one batch per data rank)")
        self.batches_per_epoch = int(lib.stb_loader_batches_per_epoch(self._h))
        self.num_samples = int(lib.stb_loader_num_samples(self._h))
        self._held: int | None = None
        self._copied: torch.cuda.Event | None = None
        if self.zigzag:
            from ..parallel.context_parallel.zigzag import local_positions

            pos = local_positions(self.s_local, cp_size, cp_rank)
        else:
            pos = torch.arange(cp_rank * self.s_local, (cp_rank + 1) * self.s_local)
        self._positions = pos.unsqueeze(0).expand(micro_batch_size, -1).contiguous()
        self.batches_served = 0

    def __iter__(self):
        return self

    def _release_held(self) -> None:
        if self._held is not None:
            if self._copied is not None:
                self._copied.synchronize()          # the H2D copies out of the slot have completed
                self._copied = None
            self._lib.stb_loader_release(self._h, self._held)
            self._held = None

    def __next__(self) -> dict:
        self._release_held()
        batch = _L(0)
        slot = int(self._lib.stb_loader_next(self._h, ctypes.byref(batch)))
        self._held = slot
        x, y = self.ring[slot, 0], self.ring[slot, 1]
        pos = self._positions
        if self.device is not None and self.device.type == "cuda":
            x, y = x.to(self.device, non_blocking=True), y.to(self.device, non_blocking=True)
            pos = pos.to(self.device, non_blocking=True)
            self._copied = torch.cuda.Event()
            self._copied.record()
        self.batches_served += 1
        return {"input_ids": x, "target_ids": y, "position_ids": pos, "hidden_states": None}

    def close(self) -> None:
        if getattr(self, "_h", None):
            self._release_held()
            self._lib.stb_loader_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
