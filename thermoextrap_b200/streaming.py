"""Streamed ingestion: reduce a sample set that never fits in (or never sits in) device memory.

BASELINE config 5 / reference ``volume.factory_extrapmodel`` (volume.py:137-209) with ``VolumeDataCallback``
(:92-134): order-1 volume extrapolation needs ``<x>, <W>, <x W>`` (raw moments of ``(xv, uv = W)``) and the mean of
``dxdqv``.  Here the three per-sample streams arrive as HOST chunks (an iterator) and go through a double-buffered
H2D ring: chunk ``i + 1`` is copied (``cudaMemcpyAsync`` on a copy stream, from pinned memory; pageable chunks are
staged through two pinned buffers) while chunk ``i`` is reduced.  Every chunk is reduced about ITS OWN pivot
(strided-subsample means, a microsecond kernel) and turned into a comoment block; the blocks are merged with the
Chan/Pebay kernel (K4, ``txb_reduce_data_f64``) at the end.  That is the drift-proof branch of SURVEY.md 7.3 -- a
common pivot taken from the first chunk loses digits when later chunk means drift by many sigma; per-chunk pivots
plus an exact merge cost two tiny launches per chunk and have no such failure mode.
"""

from __future__ import annotations

import time
from collections.abc import Iterable

import numpy as np
import torch

from . import kernels as K


class _Ring:
    """Two device slots fed by a copy stream; pageable host chunks are staged through two pinned buffers."""

    def __init__(self, dev) -> None:
        self.dev = dev
        self.copy = torch.cuda.Stream(device=dev)
        self.slots: list[dict] = [{}, {}]
        self.ready = [torch.cuda.Event(), torch.cuda.Event()]
        self.free = [torch.cuda.Event(), torch.cuda.Event()]
        self.staged = [torch.cuda.Event(), torch.cuda.Event()]
        self.pinned: list[dict] = [{}, {}]
        self.h2d_bytes = 0
        self.used = [False, False]

    def _buf(self, store: dict, key, shape, pin: bool) -> torch.Tensor:
        t = store.get(key)
        if t is None or t.shape != shape:
            if pin:
                t = torch.empty(shape, dtype=torch.float64).pin_memory()
            else:
                t = torch.empty(shape, dtype=torch.float64, device=self.dev)
            store[key] = t
        return t

    def put(self, i: int, arrays: list[np.ndarray]) -> list[torch.Tensor]:
        """Enqueue the H2D copies of chunk ``i`` on the copy stream; returns the device views (valid once
        ``acquire(i)`` has made the compute stream wait for them)."""
        s = i % 2
        out = []
        with torch.cuda.stream(self.copy):
            if self.used[s]:
                self.copy.wait_event(self.free[s])   # the reduce of chunk i - 2 is done with this slot
            for k, a in enumerate(arrays):
                h = torch.from_numpy(np.ascontiguousarray(a, dtype=np.float64))
                d = self._buf(self.slots[s], (k, tuple(h.shape)), tuple(h.shape), pin=False)
                if not h.is_pinned():
                    p = self._buf(self.pinned[s], (k, tuple(h.shape)), tuple(h.shape), pin=True)
                    if self.used[s]:
                        self.staged[s].synchronize()  # the previous copy out of this pinned buffer has finished
                    p.copy_(h)
                    h = p
                d.copy_(h, non_blocking=True)
                self.h2d_bytes += h.numel() * 8
                out.append(d)
            self.staged[s].record(self.copy)
            self.ready[s].record(self.copy)
        self.used[s] = True
        return out

    def acquire(self, i: int) -> None:
        torch.cuda.current_stream(self.dev).wait_event(self.ready[i % 2])

    def release(self, i: int) -> None:
        self.free[i % 2].record(torch.cuda.current_stream(self.dev))


def volume_moments_stream(chunks: Iterable, order: int = 1, timing: dict | None = None):
    """Reduce an iterator of host chunks ``(x[n(, *val)], dxdq[n(, *val)], W[n])``.

    Returns ``(comom[*val, 2, order+1], dxdq_mean[*val], nsamp)`` as device tensors: the comoments of ``(x, W)`` in
    the cmomy layout and the mean of ``dxdq``.  Scalar observables go through the planar kernel (each stream read
    once, 24 B / sample); observables with value dims through two wide K1 passes per chunk."""
    dev = K.device()
    ring = _Ring(dev)
    it = iter(chunks)
    blocks_x, blocks_q = [], []
    nsamp = 0
    t0 = time.perf_counter()
    nxt = next(it, None)
    pending = None if nxt is None else ring.put(0, list(nxt))
    i = 0
    val_shape: tuple = ()
    while pending is not None:
        cur = pending
        nxt = next(it, None)
        pending = None if nxt is None else ring.put(i + 1, list(nxt))   # overlaps with the reduce of chunk i
        ring.acquire(i)
        x, q, w = cur
        nsamp += int(w.shape[0])
        if x.ndim == 1:
            pivot, S0, S1 = K.reduce_sums_planar([x, q], w, order)
            com = K.finalize_sums(pivot, S0, S1, order, val_shape=(2,))  # [2, 2, order+1]: plane 0 = x, plane 1 = dxdq
            blocks_x.append(com[0])
            blocks_q.append(com[1])
        else:
            val_shape = tuple(x.shape[1:])
            blocks_x.append(K.reduce_vals(x, w, order))
            blocks_q.append(K.reduce_vals(q, w, 1))
        ring.release(i)
        i += 1
    if not blocks_x:
        msg = "no chunks"
        raise ValueError(msg)
    comx = K.reduce_data(torch.stack(blocks_x)) if len(blocks_x) > 1 else blocks_x[0]
    comq = K.reduce_data(torch.stack(blocks_q)) if len(blocks_q) > 1 else blocks_q[0]
    dxdq_mean = comq[..., 1, 0]
    if timing is not None:
        torch.cuda.synchronize(dev)
        timing.update(seconds=time.perf_counter() - t0, h2d_bytes=ring.h2d_bytes, chunks=i)
    return comx.reshape(*val_shape, 2, order + 1), dxdq_mean.reshape(val_shape), nsamp


def bench_volume_stream(total: int, chunk: int, dev, peaks: dict) -> dict:
    """BASELINE config 5 record: ``total`` samples streamed from pinned host memory in ``chunk``-sample pieces
    (a pool of 4 distinct pinned chunks is cycled -- 1e9 samples x 24 B would need 24 GB of pinned host memory),
    end to end (PCIe-bound) and the reduce kernel alone on a resident chunk (HBM-bound)."""
    import ctypes as C

    from . import _lib

    lib = _lib.load()
    pool = []
    g = torch.Generator(device=dev).manual_seed(5)
    for _ in range(4):
        w = torch.empty(chunk, dtype=torch.float64, device=dev).normal_(0.0, 1.0, generator=g)
        x = torch.empty(chunk, dtype=torch.float64, device=dev).uniform_(0.0, 1.0, generator=g).add_(w, alpha=0.1)
        pool.append(tuple(t.cpu().pin_memory().numpy() for t in (x, x.clone(), w)))
        del w, x
    nchunks = -(-total // chunk)

    def gen():
        left = total
        for i in range(nchunks):
            m = min(chunk, left)
            left -= m
            x, q, w = pool[i % len(pool)]
            yield x[:m], q[:m], w[:m]

    volume_moments_stream(gen(), 1)   # warm-up (allocates the ring)
    timing: dict = {}
    com, dq, n = volume_moments_stream(gen(), 1, timing=timing)
    assert n == total and float(com[0, 0]) == float(total)
    # kernel alone on resident chunks
    xd, qd, wd = (torch.from_numpy(a).to(dev) for a in pool[0])
    K.reduce_sums_planar([xd, qd], wd, 1)
    lib.txb_profile_begin(b"reduce_planar")
    for _ in range(10):
        K.reduce_sums_planar([xd, qd], wd, 1)
    ms, nl = C.c_double(), C.c_int()
    lib.txb_profile_end(C.byref(ms), C.byref(nl))
    k_ms = ms.value / max(nl.value, 1)
    gbs = 24.0 * chunk / k_ms / 1e6
    e2e_s = timing["seconds"]
    return {"workload": "C5: volume extrapolation order 1, 1e9 samples streamed from pinned host memory in 2^26-sample "
                        "chunks (x, dxdq, W) through a double-buffered H2D ring; per-chunk pivot + Chan/Pebay merge",
            "samples": total, "chunks": timing["chunks"], "e2e_seconds": e2e_s, "e2e_samples_per_s": total / e2e_s,
            "h2d_GBps": timing["h2d_bytes"] / e2e_s / 1e9, "h2d_bytes": timing["h2d_bytes"],
            "kernel": "k_reduce_planar<1, 2> (cp.async.bulk ring, every stream read once)", "kernel_ms_per_chunk": k_ms,
            "kernel_samples_per_s": chunk / k_ms * 1e3, "algorithmic_bytes_per_sample": 24, "kernel_GBps": gbs,
            "kernel_frac_hbm": gbs / peaks["hbm_gbs"], "bound_e2e": "PCIe (H2D)", "bound_kernel": "hbm",
            "mean_x": float(com[1, 0]), "mean_dxdq": float(dq)}
