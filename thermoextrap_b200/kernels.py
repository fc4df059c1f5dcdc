"""Device-level wrappers over the C ABI (torch CUDA tensors in, torch CUDA tensors out).

torch is plumbing here: device memory, streams, pinned host staging.  All
arithmetic happens in ``libtxb.so``.
"""

from __future__ import annotations

import ctypes as C
import os

import numpy as np
import torch

from . import _lib
from ._lib import MODE_COMOMENTS, MODE_X_IS_U, TXB_I32, TXB_I64, TXB_U8, TXB_U16, check

_FREQ_DTYPES = {
    torch.int64: TXB_I64,
    torch.int32: TXB_I32,
    torch.uint16: TXB_U16,
    torch.uint8: TXB_U8,
}

_workspaces: dict[tuple[int, int], torch.Tensor] = {}

#: incremented by every kernel-launching ABI call (for bench.py's gpu_launches)
launch_count = 0


def device() -> torch.device:
    """CUDA device of this process (one process per GPU: LOCAL_RANK)."""
    if not torch.cuda.is_available():
        msg = "thermoextrap_b200 needs a CUDA device (B200, sm_100a); there is no CPU fallback."
        raise RuntimeError(msg)
    return torch.device("cuda", int(os.environ.get("LOCAL_RANK", "0")))


def _stream_ptr() -> int:
    return torch.cuda.current_stream().cuda_stream


def _workspace(nbytes: int, dev: torch.device) -> torch.Tensor:
    key = (dev.index or 0, torch.cuda.current_stream().cuda_stream)
    ws = _workspaces.get(key)
    if ws is None or ws.numel() < nbytes:
        ws = torch.empty(max(nbytes, 1 << 20), dtype=torch.uint8, device=dev)
        _workspaces[key] = ws
    return ws


def _ptr(t: torch.Tensor | None) -> int | None:
    return None if t is None else t.data_ptr()


# ---------------------------------------------------------------- uploads in flight
# A large sample array travels host -> device in row chunks on a dedicated copy stream, one event per chunk
# (backend.to_device).  The device tensor is handed out at once; every wrapper of this module passes its inputs
# through ``_as_f64``, which makes the launching stream wait for the last chunk.  The wide bootstrap
# (``sampler_resample_wide``) does better: its integer kernels do not read the samples at all, and its first
# replicate window contracts chunk after chunk as they land, so the upload hides behind work the job does anyway.
ASYNC_UPLOAD_MIN_BYTES = 1 << 30
UPLOAD_CHUNKS = 16
UPLOAD_TEST_DELAY_CYCLES = 0   # tests: hold the copy stream back this many SM cycles so that consumers find it in flight
follow_windows = 0             # replicate windows contracted chunk by chunk behind an upload (tests, bench)
# how many windows follow an upload: enough contraction work to cover the copy (measured on B200: pinned host -> HBM
# 55 GB/s over PCIe gen5 x16; contraction 32.8 TFLOP/s fp64), within half of the free memory for the extra tables
FOLLOW_H2D_BYTES_PER_S = 50e9
FOLLOW_FLOPS_PER_S = 30e12
FOLLOW_MAX_WINDOWS = 8
upload_trace: dict | None = None   # tools/e2e_probe.py: set to {} to collect timed events of one upload + follow


class _Upload:
    __slots__ = ("tensor", "src", "bounds", "events", "pivot_rows", "pivot_idx", "pivot_landed", "_keep", "pivot",
                 "eta_bytes")

    def __init__(self, tensor, src) -> None:
        self.tensor, self.src = tensor, src   # both kept alive until the last chunk has landed
        self.bounds: list = []                # row ranges in arrival order ...
        self.events: list = []                # ... and the event that marks each one landed
        self.pivot = None                     # a ready-made pivot (parallel.py's gathered arrivals), else from pivot_rows
        self.pivot_landed = None
        self.eta_bytes = None                 # bytes still to cross PCIe into this GPU (None: the whole tensor)


def pivot_row_index(nsamp: int, dev=None) -> torch.Tensor:
    """The <= 256 strided rows whose means txb_pivot_f64 centres the sums on (reduce.cu, k_colmean_sample)."""
    nrows = min(int(nsamp), 256)
    return (torch.arange(nrows, dtype=torch.int64, device=dev) * int(nsamp)) // nrows


_uploads: dict[int, _Upload] = {}
_copy_streams: dict[int, torch.cuda.Stream] = {}


def _storage_key(t: torch.Tensor) -> int:
    return t.untyped_storage().data_ptr()


def purge_uploads() -> None:
    """Forget the uploads whose last chunk has landed (releases the references they hold)."""
    for key in [k for k, up in _uploads.items() if up.events[-1].query()]:
        del _uploads[key]


def upload_async(src: torch.Tensor, dev: torch.device) -> torch.Tensor:
    """Start the chunked copy of the pinned host tensor ``src[rows, ...]`` and return the device tensor."""
    purge_uploads()
    t = torch.empty(src.shape, dtype=src.dtype, device=dev)
    main = torch.cuda.current_stream(dev)
    cs = _copy_streams.get(dev.index or 0)
    if cs is None:
        cs = _copy_streams[dev.index or 0] = torch.cuda.Stream(dev)
    cs.wait_stream(main)   # the block may be recycled memory that queued work of this stream still reads
    rows = src.shape[0]
    step = -(-(-(-rows // UPLOAD_CHUNKS)) // 4096) * 4096
    up = _Upload(t, src)
    # the pivot rows go first (128 KB): a consumer that follows the chunks needs them before any chunk
    up.pivot_idx = pivot_row_index(rows)
    up._keep = src.reshape(rows, -1)[up.pivot_idx].pin_memory()
    up.pivot_rows = torch.empty(up._keep.shape, dtype=src.dtype, device=dev)
    timed = upload_trace is not None
    with torch.cuda.stream(cs):
        if UPLOAD_TEST_DELAY_CYCLES:
            torch.cuda._sleep(int(UPLOAD_TEST_DELAY_CYCLES))
        up.pivot_rows.copy_(up._keep, non_blocking=True)
        up.pivot_landed = torch.cuda.Event(enable_timing=timed)
        up.pivot_landed.record(cs)
        for a in range(0, rows, step):
            b = min(rows, a + step)
            t[a:b].copy_(src[a:b], non_blocking=True)
            ev = torch.cuda.Event(enable_timing=timed)
            ev.record(cs)
            up.bounds.append((a, b))
            up.events.append(ev)
    _uploads[_storage_key(t)] = up
    return t


def pending_upload(t) -> _Upload | None:
    """The upload still in flight into the storage of ``t``, or None."""
    if not _uploads or not isinstance(t, torch.Tensor) or not t.is_cuda:
        return None
    key = _storage_key(t)
    up = _uploads.get(key)
    if up is not None and up.events[-1].query():
        del _uploads[key]
        return None
    return up


def wait_upload(t) -> None:
    """Order the current stream after the upload into ``t`` (no host wait)."""
    up = pending_upload(t)
    if up is not None:
        torch.cuda.current_stream(t.device).wait_event(up.events[-1])


def _as_f64(t, dev, wait: bool = True) -> torch.Tensor:
    if isinstance(t, np.ndarray):
        t = torch.from_numpy(np.ascontiguousarray(t, dtype=np.float64))
    if _uploads and wait:
        wait_upload(t)
    if t.dtype != torch.float64:
        t = t.to(torch.float64)
    if t.device != dev:
        t = t.to(dev, non_blocking=True)
    return t.contiguous()


def _layout(x, u, w, dev, wait: bool = True):
    """Normalise (x, u, w) to x[nsamp, nchan] | None, u[nsamp] | u[nsamp, nchan], w[nsamp] | None."""
    u = _as_f64(u, dev, wait)
    nsamp = u.shape[0]
    ldu = 0
    if u.ndim > 1:
        u = u.reshape(nsamp, -1)
        ldu = u.shape[1]
    val_shape: tuple[int, ...] = ()
    nchan = 1
    if x is not None:
        x = _as_f64(x, dev, wait)
        if x.shape[0] != nsamp:
            msg = f"x has {x.shape[0]} samples but u has {nsamp}"
            raise ValueError(msg)
        val_shape = tuple(x.shape[1:])
        x = x.reshape(nsamp, -1)
        nchan = x.shape[1]
        if ldu not in (0, nchan):
            msg = f"u channels {ldu} do not match x channels {nchan}"
            raise ValueError(msg)
    elif ldu:
        nchan = ldu
        val_shape = tuple(u.shape[1:])
    if w is not None:
        w = _as_f64(w, dev)
        if w.shape != (nsamp,):
            msg = f"weight must have shape ({nsamp},), got {tuple(w.shape)}"
            raise ValueError(msg)
    return x, u, w, nsamp, nchan, ldu, val_shape


def reduce_vals(x, u, order: int, w=None, x_is_u: bool = False) -> torch.Tensor:
    """K1.  Comoments ``[*val, 2, order+1]`` (cmomy layout) of samples along axis 0."""
    global launch_count
    lib = _lib.load()
    dev = device()
    if x_is_u:
        x = None
    x, u, w, nsamp, nchan, ldu, val_shape = _layout(x, u, w, dev)
    if x is None and not x_is_u:
        msg = "x is None requires x_is_u=True"
        raise ValueError(msg)
    if nsamp == 0:
        msg = "cannot reduce zero samples"
        raise ValueError(msg)
    out = torch.empty((nchan, 2, order + 1), dtype=torch.float64, device=dev)
    nb = lib.txb_reduce_workspace_bytes(nsamp, nchan, order)
    ws = _workspace(nb, dev)
    mode = MODE_X_IS_U if x_is_u else MODE_COMOMENTS
    with torch.cuda.device(dev):
        st = lib.txb_reduce_vals_f64(
            _ptr(x), _ptr(u), _ptr(w), nsamp, nchan, nchan, ldu, order, mode,
            out.data_ptr(), ws.data_ptr(), ws.numel(), _stream_ptr(),
        )
    check(st, "txb_reduce_vals_f64")
    launch_count += 6  # pivot x2, sums, sum_parts x2, finalize (scalar x; wide x adds one u-only sum)
    return out.reshape(*val_shape, 2, order + 1)


def reduce_sums(x, u, order: int, w=None, pivot=None):
    """K1 without the final re-centring: additive pivot sums ``(pivot, S0[nu, order+1], S1[nchan, order+1])``.

    Sums about a COMMON ``pivot`` (``[cx[nchan], cu[nu]]``) add across sample shards -- streamed
    chunks or GPUs (``all_reduce(SUM)``) -- and :func:`finalize_sums` turns the total into comoments.
    ``pivot=None`` estimates one from this shard (txb_pivot_f64).  Also returns ``val_shape``.
    """
    global launch_count
    lib = _lib.load()
    dev = device()
    x, u, w, nsamp, nchan, ldu, val_shape = _layout(x, u, w, dev)
    nu = nchan if ldu else 1
    ncx = nchan if x is not None else 0
    with torch.cuda.device(dev):
        if pivot is None:
            pivot = torch.empty(ncx + nu, dtype=torch.float64, device=dev)
            check(lib.txb_pivot_f64(_ptr(x), _ptr(u), nsamp, nchan, nchan, ldu, pivot.data_ptr(), _stream_ptr()),
                  "txb_pivot_f64")
        S0 = torch.empty((nu, order + 1), dtype=torch.float64, device=dev)
        S1 = torch.empty((nchan, order + 1), dtype=torch.float64, device=dev) if x is not None else None
        ws = _workspace(lib.txb_reduce_workspace_bytes(nsamp, nchan, order), dev)
        st = lib.txb_reduce_sums_f64(_ptr(x), _ptr(u), _ptr(w), nsamp, nchan, nchan, ldu, order, pivot.data_ptr(),
                                     S0.data_ptr(), _ptr(S1), ws.data_ptr(), ws.numel(), _stream_ptr())
    check(st, "txb_reduce_sums_f64")
    launch_count += 1
    return pivot, S0, S1, val_shape


def reduce_sums_planar(planes, u, order: int, pivot=None):
    """K1 pivot sums of PLANAR channels: ``planes`` = 1..4 separate ``x_c[nsamp]`` device arrays next to
    ``u[nsamp]``; every stream is read once (the volume path's ``(x, dxdq, W)``: 24 B / sample).  Returns
    ``(pivot[C + 1], S0[1, order+1], S1[C, order+1])`` for :func:`finalize_sums`."""
    global launch_count
    lib = _lib.load()
    dev = device()
    planes = [_as_f64(p, dev) for p in planes]
    u = _as_f64(u, dev)
    n, c = u.shape[0], len(planes)
    if not 1 <= c <= 4 or any(p.shape != (n,) for p in planes):
        msg = f"reduce_sums_planar needs 1..4 planes of shape ({n},)"
        raise ValueError(msg)
    with torch.cuda.device(dev):
        if pivot is None:
            pivot = torch.empty(c + 1, dtype=torch.float64, device=dev)
            tmp = torch.empty(2, dtype=torch.float64, device=dev)
            for i, p in enumerate(planes):   # strided-subsample means: [cx_i, cu]
                check(lib.txb_pivot_f64(p.data_ptr(), u.data_ptr(), n, 1, 1, 0, tmp.data_ptr(), _stream_ptr()),
                      "txb_pivot_f64")
                pivot[i : i + 1].copy_(tmp[:1])
            pivot[c:].copy_(tmp[1:])
        S0 = torch.empty((1, order + 1), dtype=torch.float64, device=dev)
        S1 = torch.empty((c, order + 1), dtype=torch.float64, device=dev)
        ptrs = (C.c_void_p * c)(*[p.data_ptr() for p in planes])
        ws = _workspace(lib.txb_reduce_planar_workspace_bytes(c, order), dev)
        st = lib.txb_reduce_sums_planar_f64(ptrs, c, u.data_ptr(), n, order, pivot.data_ptr(), S0.data_ptr(),
                                            S1.data_ptr(), ws.data_ptr(), ws.numel(), _stream_ptr())
    check(st, "txb_reduce_sums_planar_f64")
    launch_count += 3 + 2 * c
    return pivot, S0, S1


def finalize_sums(pivot, S0, S1, order: int, x_is_u: bool = False, val_shape=()) -> torch.Tensor:
    """Pivot sums -> comoments ``[(nrep,) *val, 2, order+1]`` (``S0`` carries order+1 powers when ``x_is_u``).
    ``S0[(nrep,) nu, ks+1]``, ``S1[(nrep,) nchan, order+1]``: a leading replicate axis is carried through."""
    global launch_count
    lib = _lib.load()
    dev = device()
    lead = tuple(S0.shape[:-2])
    nrep = int(np.prod(lead, dtype=np.int64)) if lead else 1
    nu = S0.shape[-2]
    nchan = S1.shape[-2] if S1 is not None else nu
    S0 = S0.contiguous()
    S1 = None if S1 is None else S1.contiguous()
    out = torch.empty((nrep, nchan, 2, order + 1), dtype=torch.float64, device=dev)
    mode = MODE_X_IS_U if x_is_u else MODE_COMOMENTS
    with torch.cuda.device(dev):
        st = lib.txb_finalize_f64(S0.data_ptr(), _ptr(S1), pivot.data_ptr(), nrep, nchan, nu, order, mode,
                                  out.data_ptr(), _stream_ptr())
    check(st, "txb_finalize_f64")
    launch_count += 1
    return out.reshape(*lead, *val_shape, 2, order + 1)


def indices_to_freq(indices, ndat: int | None = None, dtype: torch.dtype = torch.int64) -> torch.Tensor:
    """K6.  ``freq[r, i] = #{j: indices[r, j] == i}`` on the device (bit-exact integers)."""
    global launch_count
    lib = _lib.load()
    dev = device()
    if isinstance(indices, np.ndarray):
        indices = torch.from_numpy(np.ascontiguousarray(indices, dtype=np.int64))
    indices = indices.to(device=dev, dtype=torch.int64).contiguous()
    if indices.ndim != 2:
        msg = "indices must be 2-D [nrep, nsamp]"
        raise ValueError(msg)
    nrep, nsamp = indices.shape
    ndat = nsamp if ndat is None else int(ndat)
    es = torch.empty((), dtype=dtype).element_size()
    ldf = (ndat * es + 15) // 16 * 16 // es  # rows 16-byte aligned for the cp.async path
    buf = torch.empty((nrep, ldf), dtype=dtype, device=dev)
    status = torch.zeros(1, dtype=torch.int32, device=dev)
    with torch.cuda.device(dev):
        st = lib.txb_indices_to_freq(
            indices.data_ptr(), nrep, nsamp, ndat, buf.data_ptr(), _FREQ_DTYPES[dtype], ldf,
            status.data_ptr(), _stream_ptr(),
        )
    check(st, "txb_indices_to_freq")
    launch_count += 1
    flag = int(status.item())
    if flag & 1:
        msg = f"indices out of range [0, {ndat})"
        raise ValueError(msg)
    if flag & 2:
        msg = f"frequency overflow for dtype {dtype}"
        raise OverflowError(msg)
    return buf[:, :ndat]


def resample_vals(x, u, freq: torch.Tensor, order: int, w=None, x_is_u: bool = False) -> torch.Tensor:
    """K2.  ``out[rep, *val, 2, order+1]`` = comoments reweighted by ``freq[rep, samp]``."""
    global launch_count
    lib = _lib.load()
    dev = device()
    if x_is_u:
        x = None
    x, u, w, nsamp, nchan, ldu, val_shape = _layout(x, u, w, dev)
    if isinstance(freq, np.ndarray):
        freq = torch.from_numpy(freq)
    if freq.dtype not in _FREQ_DTYPES:
        msg = f"freq dtype {freq.dtype} not supported (int64, int32, uint16, uint8)"
        raise TypeError(msg)
    if freq.device != dev:
        freq = freq.to(dev)
    if freq.ndim != 2 or freq.shape[1] != nsamp:
        msg = f"freq must be [nrep, {nsamp}], got {tuple(freq.shape)}"
        raise ValueError(msg)
    if freq.stride(1) != 1:
        freq = freq.contiguous()
    nrep, ldf = freq.shape[0], freq.stride(0)
    out = torch.empty((nrep, nchan, 2, order + 1), dtype=torch.float64, device=dev)
    nb = lib.txb_resample_workspace_bytes(nsamp, nchan, order, nrep)
    ws = _workspace(nb, dev)
    mode = MODE_X_IS_U if x_is_u else MODE_COMOMENTS
    with torch.cuda.device(dev):
        st = lib.txb_resample_vals_f64(
            _ptr(x), _ptr(u), _ptr(w), nsamp, nchan, nchan, ldu, order, mode,
            freq.data_ptr(), _FREQ_DTYPES[freq.dtype], nrep, ldf,
            out.data_ptr(), ws.data_ptr(), ws.numel(), _stream_ptr(),
        )
    check(st, "txb_resample_vals_f64")
    launch_count += 1
    return out.reshape(nrep, *val_shape, 2, order + 1)


def resample_indices(x, u, indices, order: int, w=None, x_is_u: bool = False) -> torch.Tensor:
    """K2, index form (scalar x): ``out[rep] = comoments(x[idx[rep]], u[idx[rep]])``."""
    global launch_count
    lib = _lib.load()
    dev = device()
    if x_is_u:
        x = None
    x, u, w, ndat, nchan, ldu, val_shape = _layout(x, u, w, dev)
    if nchan != 1 or ldu != 0:
        msg = "resample_indices handles scalar x only; use indices_to_freq + resample_vals"
        raise ValueError(msg)
    if isinstance(indices, np.ndarray):
        indices = torch.from_numpy(np.ascontiguousarray(indices, dtype=np.int64))
    indices = indices.to(device=dev, dtype=torch.int64)
    if indices.ndim != 2:
        msg = "indices must be 2-D [nrep, nsamp]"
        raise ValueError(msg)
    if indices.stride(1) != 1:
        indices = indices.contiguous()
    nrep_total, ndraw = indices.shape
    out = torch.empty((nrep_total, 1, 2, order + 1), dtype=torch.float64, device=dev)
    status = torch.zeros(1, dtype=torch.int32, device=dev)
    mode = MODE_X_IS_U if x_is_u else MODE_COMOMENTS
    flag = 0
    with torch.cuda.device(dev):
        for r0 in range(0, nrep_total, 32768):
            sub = indices[r0 : r0 + 32768]
            nrep = sub.shape[0]
            nb = lib.txb_resample_indices_workspace_bytes(ndat, order, nrep, ndraw)
            ws = _workspace(nb, dev)
            st = lib.txb_resample_indices_f64(
                _ptr(x), _ptr(u), _ptr(w), ndat, 1, order, mode, sub.data_ptr(), nrep, ndraw,
                sub.stride(0), out[r0:].data_ptr(), status.data_ptr(), ws.data_ptr(), ws.numel(),
                _stream_ptr(),
            )
            check(st, "txb_resample_indices_f64")
            launch_count += 1
            flag |= int(status.item())
    if flag:
        msg = f"indices out of range [0, {ndat})"
        raise ValueError(msg)
    return out.reshape(nrep_total, *val_shape, 2, order + 1)


# ------------------------------------------------------------------ device sampler (K6)
def pcg64_state(rng: np.random.Generator) -> _lib.Pcg64State:
    """Snapshot a numpy ``Generator(PCG64)`` as the C ABI's ``txb_pcg64``."""
    st = rng.bit_generator.state
    if st.get("bit_generator") != "PCG64":
        msg = f"the device sampler reproduces PCG64 only, got {st.get('bit_generator')!r}"
        raise ValueError(msg)
    s, inc = st["state"]["state"], st["state"]["inc"]
    m = (1 << 64) - 1
    return _lib.Pcg64State(s >> 64, s & m, inc >> 64, inc & m, int(st["has_uint32"]), int(st["uinteger"]))


def pcg64_commit(rng: np.random.Generator, state: _lib.Pcg64State, raw_used: int) -> None:
    """Move ``rng`` past ``raw_used`` raw 32-bit draws, exactly as numpy itself would have."""
    lib = _lib.load()
    check(lib.txb_pcg64_advance(C.byref(state), int(raw_used)), "txb_pcg64_advance")
    st = rng.bit_generator.state
    st["state"]["state"] = (state.state_hi << 64) | state.state_lo
    st["has_uint32"] = int(state.has_uint32)
    st["uinteger"] = int(state.uinteger)
    rng.bit_generator.state = st


def _sampler_flags(status: torch.Tensor, what: str) -> None:
    flag = int(status.item())
    if flag & 2:
        raise OverflowError(f"{what}: frequency count exceeds the table's integer type")
    if flag & 4:
        raise _lib.TxbError(f"{what}: raw-draw range estimate too short (report this)")


class SamplerPlan:
    """A pending ``rng.choice(ndat, (nrep, nsamp))`` table that lives only as generator state.

    Creating the plan runs the device counting pass, learns how many raw draws the table
    consumes and moves ``rng`` past them immediately -- exactly what numpy does when the
    reference builds the table eagerly -- so later use of ``rng`` stays bit-identical.  The
    table itself (80 GB at BASELINE config 2) is regenerated inside whichever kernel needs it.
    """

    def __init__(self, rng: np.random.Generator, nrep: int, nsamp: int, ndat: int | None = None, group=None) -> None:
        """``group``: a ``torch.distributed`` process group whose ranks all build the SAME plan (same
        ``rng`` state); the counting pass is then split across them and combined with one small
        all-reduce of the per-block counts."""
        global launch_count
        lib = _lib.load()
        dev = device()
        self.nrep, self.nsamp = int(nrep), int(nsamp)
        self.ndat = self.nsamp if ndat is None else int(ndat)
        self.state = pcg64_state(rng)
        count = self.nrep * self.nsamp
        nhead = lib.txb_sampler_workspace_bytes(self.ndat, count)
        if nhead == 0:
            check(1, "txb_sampler_workspace_bytes")
        self.head = torch.zeros(nhead, dtype=torch.uint8, device=dev)
        aux = torch.zeros(2, dtype=torch.int64, device=dev)
        status = aux[1:].view(torch.int32)
        args = (C.byref(self.state), self.ndat, count)
        with torch.cuda.device(dev):
            if group is None:
                st = lib.txb_sampler_prepare(*args, aux.data_ptr(), status.data_ptr(), self.head.data_ptr(),
                                             self.head.numel(), _stream_ptr())
                check(st, "txb_sampler_prepare")
            else:
                import torch.distributed as dist

                world, rank = dist.get_world_size(group), dist.get_rank(group)
                nb = lib.txb_sampler_nblocks(*args)
                if nb < 0:
                    check(1, "txb_sampler_nblocks")
                lo, hi = nb * rank // world, nb * (rank + 1) // world
                check(lib.txb_sampler_count(*args, lo, hi, self.head.data_ptr(), self.head.numel(), _stream_ptr()),
                      "txb_sampler_count")
                lay = (C.c_int64 * 4)()
                check(lib.txb_sampler_layout(self.ndat, count, lay), "txb_sampler_layout")
                # ranks counted disjoint block ranges: integer SUM of the two count arrays
                dist.all_reduce(self.head.view(torch.int64)[8 : 8 + lay[0]], group=group)
                dist.all_reduce(self.head[lay[2] : lay[2] + 4 * lay[1]].view(torch.int32), group=group)
                check(lib.txb_sampler_scan(*args, aux.data_ptr(), status.data_ptr(), self.head.data_ptr(),
                                           self.head.numel(), _stream_ptr()), "txb_sampler_scan")
        launch_count += 3
        _sampler_flags(status[:1], "txb_sampler_prepare")
        self.raw_used = int(aux[0].item())
        after = _lib.Pcg64State.from_buffer_copy(self.state)
        pcg64_commit(rng, after, self.raw_used)

    def workspace(self, nbytes: int, dev) -> torch.Tensor:
        ws = _workspace(max(nbytes, self.head.numel()), dev)
        ws[: self.head.numel()].copy_(self.head)
        return ws


def _plan_of(src, nrep, nsamp, ndat) -> SamplerPlan:
    if isinstance(src, SamplerPlan):
        return src
    return SamplerPlan(src, nrep, nsamp, ndat)


def _window(plan: SamplerPlan, reps) -> tuple[int, int]:
    lo, hi = (0, plan.nrep) if reps is None else (int(reps[0]), int(reps[1]))
    if not 0 <= lo < hi <= plan.nrep:
        msg = f"replicate window [{lo}, {hi}) outside [0, {plan.nrep})"
        raise ValueError(msg)
    return lo, hi


def sampler_indices(src, nrep: int | None = None, nsamp: int | None = None, ndat: int | None = None,
                    reps=None) -> torch.Tensor:
    """``rng.choice(ndat, (nrep, nsamp), replace=True)`` generated on the device, bit-exact (int64).

    ``src`` is a numpy ``Generator(PCG64)`` (advanced as numpy would) or a :class:`SamplerPlan`.
    """
    global launch_count
    lib = _lib.load()
    dev = device()
    if not isinstance(src, SamplerPlan) and (nsamp if ndat is None else ndat) == 1:
        return torch.zeros((nrep, nsamp), dtype=torch.int64, device=dev)  # numpy draws nothing here
    plan = _plan_of(src, nrep, nsamp, ndat)
    lo, hi = _window(plan, reps)
    out = torch.empty((hi - lo, plan.nsamp), dtype=torch.int64, device=dev)
    status = torch.zeros(1, dtype=torch.int32, device=dev)
    with torch.cuda.device(dev):
        ws = plan.workspace(0, dev)
        st = lib.txb_sampler_indices(C.byref(plan.state), plan.ndat, plan.nrep * plan.nsamp, lo * plan.nsamp,
                                     hi * plan.nsamp, out.data_ptr(), None, status.data_ptr(), 1, ws.data_ptr(),
                                     ws.numel(), _stream_ptr())
    check(st, "txb_sampler_indices")
    launch_count += 1
    return out


def sampler_freq(src, nrep: int | None = None, nsamp: int | None = None, ndat: int | None = None,
                 dtype: torch.dtype = torch.uint8, reps=None, buckets: bool = True,
                 scratch_bytes: int | None = None, out: torch.Tensor | None = None,
                 status_out: list | None = None) -> torch.Tensor:
    """Frequency table of the same stream, histogrammed on the device without the index table.

    uint8 / uint16 tables are built by partitioning the draws by sample bucket and counting in shared
    memory when scratch allows (``scratch_bytes``, default up to a third of the free memory; replicates
    go in as many windows as that needs); ``buckets=False`` forces global atomics.  ``out`` (optional,
    ``[>= hi - lo, ldf]`` of ``dtype`` with 16-byte rows) is filled instead of a new allocation.
    ``status_out`` (a list) defers the status check: the device status word is appended instead of being
    read back (no host sync); the caller inspects it later (bit 2: a count overflowed ``dtype``; bits 8 | 16:
    a bucket list / ring overflowed) and redoes the window if needed.
    """
    global launch_count
    lib = _lib.load()
    dev = device()
    plan = _plan_of(src, nrep, nsamp, ndat)
    esz = torch.empty((), dtype=dtype).element_size()
    ldf = -(-plan.ndat * esz // 16) * 16 // esz  # rows 16-byte aligned for the contraction's cp.async path
    lo, hi = _window(plan, reps)
    if out is None:
        buf = torch.empty((hi - lo, ldf), dtype=dtype, device=dev)
    else:
        if out.dtype != dtype or out.ndim != 2 or out.shape[0] < hi - lo or out.shape[1] != ldf or not out.is_contiguous():
            msg = f"out must be a contiguous [>= {hi - lo}, {ldf}] tensor of {dtype}"
            raise ValueError(msg)
        buf = out[: hi - lo]
    status = torch.zeros(1, dtype=torch.int32, device=dev)
    with torch.cuda.device(dev):
        nb = 0
        if buckets and esz < 4:
            want = lib.txb_sampler_freq_workspace_bytes(plan.ndat, plan.nrep, plan.nsamp, hi - lo)
            held = _workspaces.get((dev.index or 0, torch.cuda.current_stream().cuda_stream))
            free = torch.cuda.mem_get_info(dev)[0] + (held.numel() if held is not None else 0)
            nb = min(want, free // 3 if scratch_bytes is None else plan.head.numel() + scratch_bytes)
        ws = plan.workspace(nb, dev)
        flags = _lib.SAMPLER_PREPARED | (0 if buckets else _lib.SAMPLER_NO_BUCKETS)
        # the cached workspace only grows: an explicit scratch budget is honoured through the size we declare
        ws_bytes = ws.numel() if scratch_bytes is None else min(ws.numel(), max(nb, plan.head.numel()))
        st = lib.txb_sampler_freq(C.byref(plan.state), plan.ndat, plan.nrep, plan.nsamp, lo, hi, buf.data_ptr(),
                                  _FREQ_DTYPES[dtype], ldf, None, status.data_ptr(), flags, ws.data_ptr(),
                                  ws_bytes, _stream_ptr())
    check(st, "txb_sampler_freq")
    launch_count += 2
    if status_out is not None:
        status_out.append(status)
        return buf[:, : plan.ndat]
    if buckets and int(status.item()) & 24:  # a bucket list / ring overflowed its margin: redo with atomics
        return sampler_freq(plan, dtype=dtype, reps=reps, buckets=False, out=out)
    _sampler_flags(status, "txb_sampler_freq")
    return buf[:, : plan.ndat]


FREQ_BUCKET = 65520  # TXB_FREQ_BUCKET: column ranges of txb_sampler_freq_cols start at multiples of this


def sampler_freq_cols(plan: SamplerPlan, col_lo: int, col_hi: int, reps=None, dtype: torch.dtype = torch.uint8,
                      out: torch.Tensor | None = None, sum_cols: int | None = None):
    """Columns ``[col_lo, col_hi)`` of the plan's frequency table (one rank's sample shard) for the replicate
    window ``reps``: ``(freq[hi - lo, col_hi - col_lo], rowsum[hi - lo] int64)`` where ``rowsum`` covers the first
    ``sum_cols`` columns (default all).  ``col_lo`` must be a multiple of :data:`FREQ_BUCKET`; tables too small
    for the bucketed histogram are built whole and sliced."""
    global launch_count
    lib = _lib.load()
    dev = device()
    lo, hi = _window(plan, reps)
    if col_lo % FREQ_BUCKET or not 0 <= col_lo < col_hi <= plan.ndat:
        msg = f"column range [{col_lo}, {col_hi}) must start at a multiple of {FREQ_BUCKET} inside [0, {plan.ndat})"
        raise ValueError(msg)
    esz = torch.empty((), dtype=dtype).element_size()
    ncol = col_hi - col_lo
    sum_cols = ncol if sum_cols is None else int(sum_cols)
    if plan.ndat < 8 * FREQ_BUCKET or esz >= 4:
        full = sampler_freq(plan, dtype=dtype, reps=(lo, hi))
        sub = full[:, col_lo:col_hi]
        return sub, sub[:, :sum_cols].sum(dim=1, dtype=torch.int64)
    ldf = -(-ncol * esz // 16) * 16 // esz
    buf = torch.empty((hi - lo, ldf), dtype=dtype, device=dev) if out is None else out[: hi - lo]
    if buf.shape[1] != ldf or not buf.is_contiguous() or buf.dtype != dtype:
        msg = f"out must be a contiguous [>= {hi - lo}, {ldf}] tensor of {dtype}"
        raise ValueError(msg)
    rowsum = torch.zeros(hi - lo, dtype=torch.int64, device=dev)
    status = torch.zeros(1, dtype=torch.int32, device=dev)
    with torch.cuda.device(dev):
        want = lib.txb_sampler_freq_workspace_bytes(plan.ndat, plan.nrep, plan.nsamp, hi - lo)
        held = _workspaces.get((dev.index or 0, torch.cuda.current_stream().cuda_stream))
        free = torch.cuda.mem_get_info(dev)[0] + (held.numel() if held is not None else 0)
        ws = plan.workspace(min(want, free // 3), dev)
        st = lib.txb_sampler_freq_cols(C.byref(plan.state), plan.ndat, plan.nrep, plan.nsamp, lo, hi, col_lo, col_hi,
                                       buf.data_ptr(), _FREQ_DTYPES[dtype], ldf, rowsum.data_ptr(), sum_cols,
                                       status.data_ptr(), _lib.SAMPLER_PREPARED, ws.data_ptr(), ws.numel(),
                                       _stream_ptr())
    if st == 2:  # TXB_ERR_UNSUPPORTED: this table does not qualify for the bucketed histogram -> whole table, sliced
        full = sampler_freq(plan, dtype=dtype, reps=(lo, hi))
        sub = full[:, col_lo:col_hi]
        return sub, sub[:, :sum_cols].sum(dim=1, dtype=torch.int64)
    check(st, "txb_sampler_freq_cols")
    launch_count += 3
    flag = int(status.item())
    if flag & 24:
        raise _lib.TxbError("txb_sampler_freq_cols: a bucket list / ring overflowed its margin (report this)")
    return buf[:, :ncol], rowsum


def resample_sums(x, u, freq: torch.Tensor, order: int, pivot: torch.Tensor, w=None, wait: bool = True):
    """K2 without the re-centring: additive pivot sums ``(S0[nrep, 1, order+1], S1[nrep, nchan, order+1])`` of
    ``x[nsamp, nchan]``, ``u[nsamp]`` about ``pivot = [cx[nchan], cu]`` reweighted by ``freq[nrep, nsamp]``.  Sums
    about a common pivot add across sample shards (``all_reduce``); :func:`finalize_sums` re-centres the total."""
    global launch_count
    lib = _lib.load()
    dev = device()
    x, u, w, nsamp, nchan, ldu, val_shape = _layout(x, u, w, dev, wait)
    if x is None or ldu != 0:
        msg = "resample_sums needs x[nsamp, *val] and u[nsamp]"
        raise ValueError(msg)
    if freq.ndim != 2 or freq.shape[1] != nsamp or freq.stride(1) != 1 or freq.dtype not in _FREQ_DTYPES:
        msg = f"freq must be an integer [nrep, {nsamp}] table with unit column stride"
        raise ValueError(msg)
    nrep, ldf = freq.shape[0], freq.stride(0)
    S0 = torch.empty((nrep, 1, order + 1), dtype=torch.float64, device=dev)
    S1 = torch.empty((nrep, nchan, order + 1), dtype=torch.float64, device=dev)
    ws = _workspace(lib.txb_resample_workspace_bytes(nsamp, nchan, order, nrep), dev)
    with torch.cuda.device(dev):
        st = lib.txb_resample_sums_f64(_ptr(x), _ptr(u), _ptr(w), nsamp, nchan, nchan, 0, order, pivot.data_ptr(),
                                       freq.data_ptr(), _FREQ_DTYPES[freq.dtype], nrep, ldf, S0.data_ptr(),
                                       S1.data_ptr(), ws.data_ptr(), ws.numel(), _stream_ptr())
    check(st, "txb_resample_sums_f64")
    launch_count += 2
    return S0, S1, val_shape


def sampler_resample_sums_rows(plan: SamplerPlan, x_local, u_local, order: int, row_lo: int, pivot, w=None,
                               table_bytes: int | None = None, next_row_lo: int | None = None):
    """Sample-sharded K6 + K2: this rank holds rows ``[row_lo, row_lo + n_local)`` of the plan's sample set.
    For every replicate window: frequency columns of those rows (:func:`sampler_freq_cols`), contraction against
    the local rows about the common ``pivot``.  Returns ``(S0[nrep, 1, k+1], S1[nrep, nchan, k+1], val_shape,
    rowsum[nrep])`` of this shard; the caller ``all_reduce``s the sums, and the ``rowsum``s (int64) too: they
    must total ``plan.nsamp`` in every replicate (the narrow table's overflow check).  ``next_row_lo`` is the first
    row of the next rank (``None``: this is the last rank): the table of a rank starts at the bucket boundary at or
    below its first row, so consecutive ranks overlap there and each counts columns up to the next one's table."""
    dev = device()
    xs, us = _as_f64(x_local, dev), _as_f64(u_local, dev)
    ws_ = None if w is None else _as_f64(w, dev)
    n_local = us.shape[0]
    row_hi = row_lo + n_local
    col_lo = row_lo // FREQ_BUCKET * FREQ_BUCKET
    ncol = row_hi - col_lo
    next_col = plan.ndat if next_row_lo is None else next_row_lo // FREQ_BUCKET * FREQ_BUCKET
    sum_cols = max(0, min(ncol, next_col - col_lo))
    if table_bytes is None:
        table_bytes = free_table_bytes()
    ldf = -(-ncol // 16) * 16
    win = max(128, int(table_bytes // ldf) // 128 * 128)
    table = torch.empty((min(win, plan.nrep), ldf), dtype=torch.uint8, device=dev)
    nchan = int(np.prod(xs.shape[1:], dtype=np.int64)) if xs.ndim > 1 else 1
    S0 = torch.empty((plan.nrep, 1, order + 1), dtype=torch.float64, device=dev)
    S1 = torch.empty((plan.nrep, nchan, order + 1), dtype=torch.float64, device=dev)
    rowsum = torch.empty(plan.nrep, dtype=torch.int64, device=dev)
    val_shape: tuple = ()
    for lo in range(0, plan.nrep, win):
        hi = min(plan.nrep, lo + win)
        freq, rs = sampler_freq_cols(plan, col_lo, row_hi, reps=(lo, hi), out=table, sum_cols=sum_cols)
        rowsum[lo:hi] = rs
        s0, s1, val_shape = resample_sums(xs, us, freq[:, row_lo - col_lo :], order, pivot, w=ws_)
        S0[lo:hi], S1[lo:hi] = s0, s1
    return S0, S1, val_shape, rowsum


def check_deferred(statuses: list) -> None:
    """Verify the status words collected by ``sampler_resample(..., status_out=list)`` (one sync for all)."""
    global launch_count
    for st in statuses:
        flag = int(st.item())
        launch_count += 6 if flag & 32 else 2
        if flag & 24:
            msg = "a bucket list / ring overflowed in a deferred-status bootstrap: rerun it with buckets=False"
            raise _lib.TxbError(msg)
        if flag & 4:
            raise _lib.TxbError("txb_sampler_resample_f64: raw-draw range estimate too short (report this)")
    statuses.clear()


def sampler_resample(src, x, u, nrep: int | None = None, order: int = 1, w=None, x_is_u: bool = False,
                     nsamp: int | None = None, sync_rng: bool = True, reps=None, buckets: bool = True,
                     scratch_bytes: int | None = None, status_out: list | None = None) -> torch.Tensor:
    """K2 fused with the generator (scalar x): bootstrap comoments for ``{"nrep": nrep, "rng": rng}``.

    Neither the index nor the frequency table is materialised.  With a raw ``Generator`` and
    ``sync_rng=False`` the counting pass runs inside the call and ``rng`` is left untouched
    (no device->host read at all: the benchmark's steady-state form).  ``status_out`` (a list)
    defers the status check: the device status word is appended instead of being read back, and the
    caller verifies a whole batch of calls with :func:`check_deferred` -- no host sync per call.
    """
    global launch_count
    lib = _lib.load()
    dev = device()
    if x_is_u:
        x = None
    x, u, w, ndat, nchan, ldu, val_shape = _layout(x, u, w, dev)
    if nchan != 1 or ldu != 0:
        msg = "sampler_resample handles scalar x only; use sampler_freq + resample_vals"
        raise ValueError(msg)
    nsamp = ndat if nsamp is None else nsamp
    if isinstance(src, SamplerPlan) or sync_rng:
        plan = _plan_of(src, nrep, nsamp, ndat)
        if plan.ndat != ndat:
            msg = f"sampler ndat={plan.ndat} does not match {ndat} samples"
            raise ValueError(msg)
        state, nrep, nsamp, prepared = plan.state, plan.nrep, plan.nsamp, 1
    else:
        plan, state, prepared = None, pcg64_state(src), 0
    lo, hi = (0, nrep) if reps is None else (int(reps[0]), int(reps[1]))
    if not 0 <= lo < hi <= nrep:
        msg = f"replicate window [{lo}, {hi}) outside [0, {nrep})"
        raise ValueError(msg)
    out = torch.empty((hi - lo, 1, 2, order + 1), dtype=torch.float64, device=dev)
    mode = MODE_X_IS_U if x_is_u else MODE_COMOMENTS
    status = torch.zeros(1, dtype=torch.int32, device=dev)
    flags = (_lib.SAMPLER_PREPARED if prepared else 0) | (0 if buckets else _lib.SAMPLER_NO_BUCKETS)
    with torch.cuda.device(dev):
        nb = lib.txb_sampler_resample_workspace_bytes(ndat, order, nrep, nsamp, hi - lo)
        if nb == 0:
            check(1, "txb_sampler_resample_workspace_bytes")
        if buckets:
            # scratch for the bucketed path: all replicates in one window if memory allows,
            # otherwise as many as fit in a third of what is free (the library windows the rest)
            fast = lib.txb_sampler_resample_fast_workspace_bytes(ndat, order, mode, nrep, nsamp, hi - lo)
            held = _workspaces.get((dev.index or 0, torch.cuda.current_stream().cuda_stream))
            if scratch_bytes is not None:
                nb = max(nb, min(fast, nb + scratch_bytes))
            elif held is not None and held.numel() >= fast:
                nb = max(nb, fast)  # steady state: the cached workspace already covers it (no driver query)
            else:
                free = torch.cuda.mem_get_info(dev)[0] + (held.numel() if held is not None else 0)
                nb = max(nb, min(fast, free // 3))
        ws = plan.workspace(nb, dev) if plan is not None else _workspace(nb, dev)
        st = lib.txb_sampler_resample_f64(
            C.byref(state), _ptr(x), _ptr(u), _ptr(w), ndat, 1, order, mode, nrep, nsamp, lo, hi, out.data_ptr(),
            None, status.data_ptr(), flags, ws.data_ptr(), ws.numel(), _stream_ptr(),
        )
    check(st, "txb_sampler_resample_f64")
    if status_out is not None:
        status_out.append(status)
        launch_count += 3 + (0 if prepared else 4)  # check_deferred adds the path-dependent rest
        return out.reshape(hi - lo, *val_shape, 2, order + 1)
    flag = int(status.item())
    # kernels launched: pivot x2, pack, [count, scan, iter_scan, locate], then
    # gather + finalize, or scatter + consume + sum_parts x2 + finalize + status note (per window)
    launch_count += 3 + (0 if prepared else 4) + (6 if flag & 32 else 2)
    if flag & 4:  # the in-call counting pass found fewer accepted draws than nrep * nsamp: sums are incomplete
        raise _lib.TxbError("txb_sampler_resample_f64: raw-draw range estimate too short (report this)")
    if buckets and flag & 24:  # a bucket list / ring overflowed its margin: redo by gathering
        return sampler_resample(src if plan is None else plan, x, u, nrep, order, w=w, x_is_u=x_is_u, nsamp=nsamp,
                                sync_rng=False, reps=reps, buckets=False)
    return out.reshape(hi - lo, *val_shape, 2, order + 1)


def free_table_bytes() -> int:
    """Default byte budget of one frequency-table window: a third of the free device memory."""
    return torch.cuda.mem_get_info(device())[0] // 3


def sampler_resample_wide(src, x, u, nrep: int | None = None, order: int = 1, w=None, x_is_u: bool = False,
                          nsamp: int | None = None, table_bytes: int | None = None, timing: dict | None = None,
                          reps=None) -> torch.Tensor:
    """K6 + K2 for wide ``x[nsamp, *val]``: the sampler's frequency table is histogrammed on the device one
    replicate WINDOW at a time and contracted against the samples in place, so a table that does not fit in
    HBM (BASELINE config 3: 1e8 samples x 1e4 replicates = 1 TB even as uint8) never exists as a whole.

    ``table_bytes`` bounds one window's table (default: a third of the free device memory); windows are
    multiples of 128 replicates, the contraction kernel's replicate tile.  ``timing`` (optional dict)
    receives the device milliseconds spent in the histogram and contraction legs.  ``reps=(lo, hi)``
    bootstraps only that slice of the plan's replicates (a rank's share, parallel.bootstrap_sharded).
    """
    dev = device()
    xs = None if x_is_u else _as_f64(x, dev, wait=False)
    us = _as_f64(u, dev)
    ndat = us.shape[0]
    plan = _plan_of(src, nrep, ndat if nsamp is None else nsamp, ndat)
    if plan.ndat != ndat:
        msg = f"sampler ndat={plan.ndat} does not match {ndat} samples"
        raise ValueError(msg)
    # samples still on their way (backend.to_device's chunked upload)?  The sampler kernels do not read them, and
    # the first window's contraction can follow the chunks as they land
    arriving = None if xs is None else pending_upload(xs)
    if arriving is not None and not (timing is None and w is None and us.ndim == 1 and xs.ndim >= 2 and xs.data_ptr() == arriving.tensor.data_ptr()
                                     and xs.shape == arriving.tensor.shape):
        wait_upload(xs)
        arriving = None
    follow_pivot = None
    if arriving is not None and arriving.pivot is not None:
        follow_pivot = arriving.pivot
    elif arriving is not None:
        torch.cuda.current_stream(dev).wait_event(arriving.pivot_landed)
        follow_pivot = sample_pivot(arriving.pivot_rows, us[pivot_row_index(ndat, dev)])   # no H2D copy here
    if table_bytes is None:
        table_bytes = free_table_bytes()
    ldf = -(-ndat // 16) * 16
    first, last = _window(plan, reps)
    ev = [] if timing is not None else None
    win = max(128, int(table_bytes // ldf) // 128 * 128)
    # one uint8 table serves every window.  The windows are issued back to back with NO host synchronisation:
    # each window's status word stays on the device and all of them are read once at the end.  A window whose
    # status reports an overflowed count (a sample drawn > 255 times) or an overflowed bucket list is then redone
    # on the careful path (int32 table / global atomics) -- a < 1e-12 event at bootstrap sizes.
    table = torch.empty((min(win, last - first), ldf), dtype=torch.uint8, device=dev)
    # following an upload: as many windows as it takes to keep the tensor pipe busy until the last chunk has landed
    # get a table each and are contracted chunk-major (every chunk against all of them as it lands)
    extra: list = []
    if arriving is not None:
        nchan = int(np.prod(xs.shape[1:], dtype=np.int64))
        upload_s = (arriving.eta_bytes or xs.numel() * 8) / FOLLOW_H2D_BYTES_PER_S
        window_s = 2.0 * (order + 1) * (nchan + 1) * ndat * table.shape[0] / FOLLOW_FLOPS_PER_S
        want_tables = min(-(-(last - first) // win), max(1, int(np.ceil(upload_s / window_s))), FOLLOW_MAX_WINDOWS)
        # free = what the driver reports + what torch's allocator holds cached (last step's tables come back from there)
        room = (torch.cuda.mem_get_info(dev)[0] + torch.cuda.memory_reserved(dev) - torch.cuda.memory_allocated(dev)) // 2
        while len(extra) + 1 < want_tables and table.numel() * (len(extra) + 1) <= room:
            try:
                extra.append(torch.empty_like(table))
            except torch.OutOfMemoryError:
                break
    # scratch of the bucketed histogram, sized once (no driver query per window)
    lib = _lib.load()
    want = lib.txb_sampler_freq_workspace_bytes(plan.ndat, plan.nrep, plan.nsamp, table.shape[0])
    held = _workspaces.get((dev.index or 0, torch.cuda.current_stream().cuda_stream))
    free = torch.cuda.mem_get_info(dev)[0] + (held.numel() if held is not None else 0)
    scratch = max(0, min(want, free // 2) - plan.head.numel())
    out = None
    pending: list = []
    spans: list = []
    lo = first
    if arriving is not None:
        freqs = []
        for tb in [table, *extra]:
            if lo >= last:
                break
            hi = min(last, lo + win)
            freqs.append(sampler_freq(plan, dtype=torch.uint8, reps=(lo, hi), out=tb, status_out=pending,
                                      scratch_bytes=scratch))
            spans.append((lo, hi))
            lo = hi
        parts = _resample_windows_following(arriving, xs, us, freqs, order, follow_pivot)
        if lo < last or len(parts) > 1:
            out = torch.empty((last - first, *parts[0].shape[1:]), dtype=torch.float64, device=dev)
            for (a, b), part in zip(spans, parts):
                out[a - first : b - first] = part
        else:
            out = parts[0]
        del freqs, parts, extra   # the later windows find the whole array resident (stream order) and share `table`
    while lo < last:
        hi = min(last, lo + win)
        marks = [torch.cuda.Event(enable_timing=True) for _ in range(3)] if ev is not None else None
        if marks:
            marks[0].record()
        freq = sampler_freq(plan, dtype=torch.uint8, reps=(lo, hi), out=table, status_out=pending,
                            scratch_bytes=scratch)
        if marks:
            marks[1].record()
        part = resample_vals(xs, us, freq, order, w=w, x_is_u=x_is_u)
        if marks:
            marks[2].record()
            ev.append(marks)
        spans.append((lo, hi))
        if lo == first and hi == last:
            out = part
        else:
            if out is None:
                out = torch.empty((last - first, *part.shape[1:]), dtype=torch.float64, device=dev)
            out[lo - first : hi - first] = part
        lo = hi
    flags = torch.cat(pending).cpu().tolist()   # the one synchronisation of the call
    del table
    for (lo, hi), flag in zip(spans, flags):
        if flag & 4:
            raise _lib.TxbError("txb_sampler_freq: raw-draw range estimate too short (report this)")
        if flag & (2 | 24):
            part = _resample_window_careful(plan, xs, us, order, w, x_is_u, lo, hi, overflow=bool(flag & 2))
            out[lo - first : hi - first] = part
    if ev is not None:
        torch.cuda.synchronize(dev)
        timing["freq_ms"] = sum(m[0].elapsed_time(m[1]) for m in ev)
        timing["contract_ms"] = sum(m[1].elapsed_time(m[2]) for m in ev)
        timing["windows"] = len(ev)
    return out


def sample_pivot(x, u) -> torch.Tensor:
    """``[cx[nchan], cu]``: the strided-subsample means the kernels centre their sums on (txb_pivot_f64)."""
    lib = _lib.load()
    dev = device()
    x, u, _, nsamp, nchan, ldu, _ = _layout(x, u, None, dev)
    pivot = torch.empty(nchan + (nchan if ldu else 1), dtype=torch.float64, device=dev)
    with torch.cuda.device(dev):
        check(lib.txb_pivot_f64(_ptr(x), _ptr(u), nsamp, nchan, nchan, ldu, pivot.data_ptr(), _stream_ptr()),
              "txb_pivot_f64")
    return pivot


def _resample_windows_following(up: _Upload, xs, us, freqs: list, order: int, pivot) -> list:
    """Replicate windows of :func:`sampler_resample_wide` contracted WHILE the samples upload: pivot sums are
    additive over samples, so every row chunk is contracted against each window's table as soon as its copy has
    landed (the stream waits on that chunk's event only) and the windows are re-centred once at the end.  The pivot
    is the one the resident path computes -- the means of the same <= 256 strided rows, uploaded ahead of the
    chunks -- so the two paths differ only in the order the partial sums are added."""
    global follow_windows
    follow_windows += len(freqs)
    main = torch.cuda.current_stream(xs.device)
    sums: list = [None] * len(freqs)
    trace = upload_trace
    if trace is not None:
        trace.update(t0=up.pivot_landed or up.events[0], landed=list(up.events), done=[],
                     start=torch.cuda.Event(enable_timing=True), windows=len(freqs))
        trace["start"].record(main)
    for (a, b), landed in zip(up.bounds, up.events):
        main.wait_event(landed)
        for i, freq in enumerate(freqs):
            s0, s1, _ = resample_sums(xs[a:b], us[a:b], freq[:, a:b], order, pivot, wait=False)
            if sums[i] is None:
                sums[i] = (s0, s1)
            else:
                sums[i][0].add_(s0)
                sums[i][1].add_(s1)
        if trace is not None:
            trace["done"].append(torch.cuda.Event(enable_timing=True))
            trace["done"][-1].record(main)
    return [finalize_sums(pivot, s0, s1, order, x_is_u=False, val_shape=tuple(xs.shape[1:])) for s0, s1 in sums]


def _resample_window_careful(plan, xs, us, order, w, x_is_u, lo, hi, overflow: bool) -> torch.Tensor:
    """Redo one replicate window of :func:`sampler_resample_wide` whose fast pass flagged a problem: an int32
    table when a uint8 count overflowed, global atomics when a bucket list / ring overflowed."""
    dtype = torch.int32 if overflow else torch.uint8
    step = max(1, min(hi - lo, int(free_table_bytes() // (plan.ndat * (4 if overflow else 1) + 64))))
    parts = []
    for a in range(lo, hi, step):
        b = min(hi, a + step)
        try:
            freq = sampler_freq(plan, dtype=dtype, reps=(a, b), buckets=False)
        except OverflowError:
            if dtype == torch.int32:
                raise
            freq = sampler_freq(plan, dtype=torch.int32, reps=(a, b), buckets=False)
        parts.append(resample_vals(xs, us, freq, order, w=w, x_is_u=x_is_u))
        del freq
    return torch.cat(parts)


def _tails(data, dev):
    data = _as_f64(data, dev)
    if data.ndim < 2 or data.shape[-2] != 2:
        msg = f"comoment array must end in (2, order+1), got {tuple(data.shape)}"
        raise ValueError(msg)
    return data, data.shape[-1] - 1


def reduce_data(data) -> torch.Tensor:
    """K4.  Merge blocks along axis 0: ``[nblock, *val, 2, k+1] -> [*val, 2, k+1]``."""
    global launch_count
    lib = _lib.load()
    dev = device()
    data, order = _tails(data, dev)
    nblock = data.shape[0]
    val_shape = tuple(data.shape[1:-2])
    nval = int(np.prod(val_shape, dtype=np.int64))
    out = torch.empty((*val_shape, 2, order + 1), dtype=torch.float64, device=dev)
    with torch.cuda.device(dev):
        st = lib.txb_reduce_data_f64(data.data_ptr(), nblock, nval, order, out.data_ptr(), _stream_ptr())
    check(st, "txb_reduce_data_f64")
    launch_count += 1
    return out


def resample_data(data, freq) -> torch.Tensor:
    """K3.  ``out[rep] = merge(data[block] * freq[rep, block])``."""
    global launch_count
    lib = _lib.load()
    dev = device()
    data, order = _tails(data, dev)
    if isinstance(freq, np.ndarray):
        freq = torch.from_numpy(freq)
    freq = freq.to(device=dev, dtype=torch.int64).contiguous()
    nblock = data.shape[0]
    if freq.ndim != 2 or freq.shape[1] != nblock:
        msg = f"freq must be [nrep, {nblock}], got {tuple(freq.shape)}"
        raise ValueError(msg)
    nrep = freq.shape[0]
    val_shape = tuple(data.shape[1:-2])
    nval = int(np.prod(val_shape, dtype=np.int64))
    out = torch.empty((nrep, *val_shape, 2, order + 1), dtype=torch.float64, device=dev)
    with torch.cuda.device(dev):
        st = lib.txb_resample_data_f64(
            data.data_ptr(), freq.data_ptr(), nrep, nblock, nval, order, out.data_ptr(), _stream_ptr()
        )
    check(st, "txb_resample_data_f64")
    launch_count += 1
    return out


CONVERT_CENTRAL_TO_RAW, CONVERT_RAW_TO_CENTRAL, CONVERT_CENTRAL_TO_CMOM = 0, 1, 2


def convert(data, kind: int) -> torch.Tensor:
    """K5 on ``[..., 2, k+1]`` tails."""
    global launch_count
    lib = _lib.load()
    dev = device()
    data, order = _tails(data, dev)
    out = torch.empty_like(data)
    nelem = data.numel() // (2 * (order + 1))
    with torch.cuda.device(dev):
        st = lib.txb_convert_f64(data.data_ptr(), out.data_ptr(), nelem, order, kind, _stream_ptr())
    check(st, "txb_convert_f64")
    launch_count += 1
    return out


def eval_exprs(vars_, term_ptr, coef, fac_ptr, fac_var, fac_pow) -> torch.Tensor:
    """K7.  ``out[nexpr, nelem]`` from the monomial tables (see ``include/txb.h``)."""
    global launch_count
    lib = _lib.load()
    dev = device()
    v = _as_f64(vars_, dev)
    nvar, nelem = v.shape
    nexpr = len(term_ptr) - 1

    def i32(a):
        return torch.as_tensor(np.asarray(a, dtype=np.int32), device=dev)

    tp, fp, fv, fw = i32(term_ptr), i32(fac_ptr), i32(fac_var), i32(fac_pow)
    cf = torch.as_tensor(np.asarray(coef, dtype=np.float64), device=dev)
    out = torch.empty((nexpr, nelem), dtype=torch.float64, device=dev)
    with torch.cuda.device(dev):
        st = lib.txb_eval_exprs_f64(
            v.data_ptr(), nelem, nexpr, tp.data_ptr(), cf.data_ptr(), fp.data_ptr(), fv.data_ptr(),
            fw.data_ptr(), out.data_ptr(), _stream_ptr(),
        )
    check(st, "txb_eval_exprs_f64")
    launch_count += 1
    return out


PREDICT_SUM, PREDICT_CUMSUM, PREDICT_TERMS = 0, 1, 2


def poly_predict(coefs, dalpha, mode: int = PREDICT_SUM) -> torch.Tensor:
    """``sum_p coefs[p, e] * dalpha[a]**p`` -> ``[nalpha, nelem]`` (or ``[nalpha, np1, nelem]``)."""
    global launch_count
    lib = _lib.load()
    dev = device()
    c = _as_f64(coefs, dev)
    d = _as_f64(np.atleast_1d(np.asarray(dalpha, dtype=np.float64)), dev)
    np1, nelem = c.shape
    na = d.shape[0]
    shape = (na, nelem) if mode == PREDICT_SUM else (na, np1, nelem)
    out = torch.empty(shape, dtype=torch.float64, device=dev)
    with torch.cuda.device(dev):
        st = lib.txb_poly_predict_f64(c.data_ptr(), np1, nelem, d.data_ptr(), na, mode, out.data_ptr(), _stream_ptr())
    check(st, "txb_poly_predict_f64")
    launch_count += 1
    return out


def perturb_predict(x, u, dalpha) -> torch.Tensor:
    """``PerturbModel.predict`` (models.py:1009-1046): ``out[alpha, *val]``, reweighting ``x[rec, *val]`` with
    ``exp(-dalpha * u[rec])`` (max-shifted), one streaming pass per four alphas."""
    global launch_count
    lib = _lib.load()
    dev = device()
    x = _as_f64(x, dev)
    u = _as_f64(u, dev)
    if u.ndim != 1 or x.shape[0] != u.shape[0]:
        msg = f"perturb_predict needs u[rec] and x[rec, ...]; got {tuple(u.shape)} and {tuple(x.shape)}"
        raise ValueError(msg)
    nsamp = u.shape[0]
    val_shape = tuple(x.shape[1:])
    x = x.reshape(nsamp, -1)
    nchan = x.shape[1]
    da = _as_f64(torch.as_tensor(np.atleast_1d(np.asarray(dalpha, dtype=np.float64))), dev)
    nalpha = da.numel()
    out = torch.empty((nalpha, nchan), dtype=torch.float64, device=dev)
    nb = lib.txb_perturb_workspace_bytes(nsamp, nchan, nalpha)
    ws = _workspace(nb, dev)
    with torch.cuda.device(dev):
        st = lib.txb_perturb_predict_f64(x.data_ptr(), u.data_ptr(), nsamp, nchan, nchan, da.data_ptr(), nalpha,
                                         out.data_ptr(), ws.data_ptr(), ws.numel(), _stream_ptr())
    check(st, "txb_perturb_predict_f64")
    launch_count += 4
    return out.reshape(nalpha, *val_shape)


# ------------------------------------------------------------------ K7 device-resident (derivs_fused.cu)
#: (name, central) -> family index of txb_beta_predict_f64 (the four numeric beta families)
BETA_FAMILIES = {("x_ave", True): 0, ("u_ave", True): 1, ("x_ave", False): 2, ("u_ave", False): 3}
BETA_MAX_ORDER = 10
PREDICT_DERIVS = 3


def beta_predict(comom, order: int, family: int, dalpha=None, minus_log: bool = False, mode: int = PREDICT_SUM,
                 norm: bool = False) -> torch.Tensor:
    """Comoment records ``comom[..., 2, K+1]`` (device) -> beta derivatives -> Taylor prediction, one launch.

    ``mode`` ``PREDICT_SUM``: ``[nalpha, nelem]``; ``PREDICT_CUMSUM`` / ``PREDICT_TERMS``: ``[nalpha, order+1, nelem]``;
    ``PREDICT_DERIVS``: the derivatives ``[order+1, nelem]`` (divided by ``i!`` when ``norm``).  ``nelem`` is the
    product of the leading dims of ``comom``.  Generated device functions: orders <= ``BETA_MAX_ORDER``."""
    global launch_count
    lib = _lib.load()
    dev = device()
    c = _as_f64(comom, dev)
    if c.ndim < 2 or c.shape[-2] != 2 or not c.is_contiguous():
        msg = f"comom must be a contiguous [..., 2, K+1] array; got {tuple(c.shape)}"
        raise ValueError(msg)
    kdim = c.shape[-1]
    nelem = c.numel() // (2 * kdim)
    flags = (1 if minus_log else 0) | (2 if norm else 0)
    if mode == PREDICT_DERIVS:
        d, na = None, 0
        shape = (order + 1, nelem)
    else:
        d = _as_f64(np.atleast_1d(np.asarray(dalpha, dtype=np.float64)), dev)
        na = d.shape[0]
        shape = (na, nelem) if mode == PREDICT_SUM else (na, order + 1, nelem)
    out = torch.empty(shape, dtype=torch.float64, device=dev)
    with torch.cuda.device(dev):
        st = lib.txb_beta_predict_f64(c.data_ptr(), nelem, kdim, family, order, flags, mode, _ptr(d), na,
                                      out.data_ptr(), _stream_ptr())
    check(st, "txb_beta_predict_f64")
    launch_count += 1
    return out


#: family indices of the lnPi forms (central, raw): evaluated by :func:`lnpi_predict`
LNPI_FAMILIES = {True: 4, False: 5}


def lnpi_predict(comom, order: int, central: bool, lnpi0, mudotn, dalpha=None, mode: int = PREDICT_SUM,
                 norm: bool = False) -> torch.Tensor:
    """The lnPi family on x_is_u comoment records ``comom[..., 2, K+1]``: ``d_0 = lnPi0``, ``d_1 = mu.N - <u>``,
    ``d_n = -d^(n-1)<u>/dbeta^(n-1)`` (lnpi.py:42-116), fused with ``/ i!`` and the Taylor sum as :func:`beta_predict`.
    ``lnpi0`` / ``mudotn`` hold ``ninner`` values each and broadcast over the leading dims of ``comom``."""
    global launch_count
    lib = _lib.load()
    dev = device()
    c = _as_f64(comom, dev)
    if c.ndim < 2 or c.shape[-2] != 2 or not c.is_contiguous():
        msg = f"comom must be a contiguous [..., 2, K+1] array; got {tuple(c.shape)}"
        raise ValueError(msg)
    kdim = c.shape[-1]
    nelem = c.numel() // (2 * kdim)
    l0 = _as_f64(np.ascontiguousarray(lnpi0, dtype=np.float64).reshape(-1) if isinstance(lnpi0, np.ndarray) else lnpi0, dev)
    mn = _as_f64(np.ascontiguousarray(mudotn, dtype=np.float64).reshape(-1) if isinstance(mudotn, np.ndarray) else mudotn, dev)
    ninner = l0.numel()
    if mn.numel() != ninner or ninner < 1 or nelem % ninner:
        msg = f"lnpi0 ({l0.numel()}) and mudotn ({mn.numel()}) must have equal sizes dividing {nelem} records"
        raise ValueError(msg)
    if mode == PREDICT_DERIVS:
        d, na = None, 0
        shape = (order + 1, nelem)
    else:
        d = _as_f64(np.atleast_1d(np.asarray(dalpha, dtype=np.float64)), dev)
        na = d.shape[0]
        shape = (na, nelem) if mode == PREDICT_SUM else (na, order + 1, nelem)
    out = torch.empty(shape, dtype=torch.float64, device=dev)
    with torch.cuda.device(dev):
        st = lib.txb_lnpi_predict_f64(c.data_ptr(), nelem, kdim, 1 if central else 0, order, 2 if norm else 0, mode,
                                      _ptr(d), na, l0.data_ptr(), mn.data_ptr(), ninner, out.data_ptr(), _stream_ptr())
    check(st, "txb_lnpi_predict_f64")
    launch_count += 1
    return out


def combine(W, D, out: torch.Tensor | None = None) -> torch.Tensor:
    """``out[a, e] (+)= sum_t W[a, t] D[t, e]`` on the device; accumulates into ``out`` when it is given."""
    global launch_count
    lib = _lib.load()
    dev = device()
    Wd = _as_f64(np.ascontiguousarray(W, dtype=np.float64) if isinstance(W, np.ndarray) else W, dev).contiguous()
    Dd = _as_f64(D, dev).contiguous()
    na, nt = Wd.shape
    if Dd.shape[0] != nt:
        msg = f"W[{na}, {nt}] does not match D{tuple(Dd.shape)}"
        raise ValueError(msg)
    nelem = Dd.shape[1]
    acc = out is not None
    if out is None:
        out = torch.empty((na, nelem), dtype=torch.float64, device=dev)
    with torch.cuda.device(dev):
        st = lib.txb_combine_f64(Wd.data_ptr(), na, nt, Dd.data_ptr(), nelem, 1 if acc else 0, out.data_ptr(),
                                 _stream_ptr())
    check(st, "txb_combine_f64")
    launch_count += 1
    return out


def mean_var(arr, ddof: int = 0):
    """``(mean[A, B], var[A, B])`` over the middle axis of ``arr[A, R, B]`` (device), one pass (Welford + Chan)."""
    global launch_count
    lib = _lib.load()
    dev = device()
    a = _as_f64(arr, dev).contiguous()
    if a.ndim != 3:
        msg = f"mean_var needs a 3-D [A, R, B] array; got {tuple(a.shape)}"
        raise ValueError(msg)
    A, R, B = a.shape
    mean = torch.empty((A, B), dtype=torch.float64, device=dev)
    var = torch.empty((A, B), dtype=torch.float64, device=dev)
    with torch.cuda.device(dev):
        st = lib.txb_mean_var_f64(a.data_ptr(), A, R, B, int(ddof), mean.data_ptr(), var.data_ptr(), _stream_ptr())
    check(st, "txb_mean_var_f64")
    launch_count += 1
    return mean, var
