"""Multi-GPU forms of the hot path: one process per GPU, ``torch.distributed`` for the plumbing.

Two shardings, each with exactly one small exchange:

* samples sharded (K1): every rank reduces its own rows about a COMMON pivot; the additive pivot
  sums are combined with ``all_reduce(SUM)`` (``2 * nchan * (order+1)`` doubles) and re-centred.
* replicates sharded (K6 + K2): all ranks hold the samples and ONE numpy table
  ``rng.choice(ndat, (nrep, nsamp))`` is split by replicate.  The table is never built: the
  counting pass of the device sampler is split by thread block, the per-block counts (a few
  thousand int64) are summed with one ``all_reduce``, and each rank regenerates + bootstraps only
  its own replicate window.  The slabs are concatenated with ``all_gather``.

The results are identical (integers) / equal to rounding (fp64) to the single-GPU call and hence to
the reference, for any world size.
"""

from __future__ import annotations

import numpy as np
import torch
import torch.distributed as dist

from . import kernels as K


def shard_range(n: int, rank: int, world: int) -> tuple[int, int]:
    """Contiguous, balanced ``[lo, hi)`` slice of ``n`` units for ``rank``."""
    return n * rank // world, n * (rank + 1) // world


def reduce_vals_sharded(x_local, u_local, order: int, w_local=None, group=None) -> torch.Tensor:
    """K1 over samples split across the ranks of ``group``; every rank returns the full comoments."""
    if x_local is None:
        msg = "reduce_vals_sharded needs x (use x=u for x_is_u data)"
        raise ValueError(msg)
    # common pivot: rank 0's estimate (any value near the mean works; it only conditions the sums)
    pivot, _, _, _ = K.reduce_sums(x_local[:1024], u_local[:1024], 1)
    dist.broadcast(pivot, src=dist.get_global_rank(group, 0) if group is not None else 0, group=group)
    _, S0, S1, val_shape = K.reduce_sums(x_local, u_local, order, w=w_local, pivot=pivot)
    packed = torch.cat([S0.reshape(-1), S1.reshape(-1)])
    dist.all_reduce(packed, group=group)
    n0 = S0.numel()
    return K.finalize_sums(pivot, packed[:n0].view_as(S0), packed[n0:].view_as(S1), order, val_shape=val_shape)


def bootstrap_sharded(rng: np.random.Generator, x, u, nrep: int, order: int, w=None, x_is_u: bool = False,
                      group=None, gather: bool = True) -> torch.Tensor:
    """``.resample(sampler={"nrep": nrep, "rng": rng})`` with the replicates split over ranks (scalar x: fused
    sampler + bucketed bootstrap; wide ``x[rec, *val]``: windowed frequency table + tensor-core contraction).

    Every rank must pass a generator in the SAME state (e.g. ``default_rng(seed)``); it is advanced
    exactly as numpy would advance it for the whole table.  Returns ``[nrep, 2, order+1]`` on every
    rank (``gather=False``: only this rank's ``[hi - lo, 2, order+1]`` slab).
    """
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    n = int(u.shape[0])
    plan = K.SamplerPlan(rng, nrep, n, n, group=group)
    lo, hi = shard_range(nrep, rank, world)
    wide = (u if x_is_u else x).ndim > 1
    val_shape = tuple((u if x_is_u else x).shape[1:]) if wide else ()
    if hi <= lo:
        slab = None
    elif wide:
        slab = K.sampler_resample_wide(plan, x, u, order=order, w=w, x_is_u=x_is_u, reps=(lo, hi))
    else:
        slab = K.sampler_resample(plan, x, u, order=order, w=w, x_is_u=x_is_u, reps=(lo, hi))
    if not gather:
        return slab
    tail = (*val_shape, 2, order + 1)
    sizes = [shard_range(nrep, r, world) for r in range(world)]
    width = max(b - a for a, b in sizes)
    buf = torch.zeros((width, *tail), dtype=torch.float64, device=K.device())
    if slab is not None:
        buf[: hi - lo] = slab.reshape(hi - lo, *tail)
    parts = [torch.empty_like(buf) for _ in range(world)]
    dist.all_gather(parts, buf, group=group)
    return torch.cat([p[: b - a] for p, (a, b) in zip(parts, sizes)])
