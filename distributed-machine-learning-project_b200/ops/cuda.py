"""sm_100a ops (ctypes over libdmlp_knn.so) on torch CUDA tensors.

Each wrapper launches on the current torch CUDA stream and never falls back to
PyTorch math: if the library is missing or a launch fails, it raises.
"""
import ctypes as C

import torch

from .._native import knn_lib

STATS_BYTES = 32
KP = 32          # keys kept per (query, chunk) by the fused epilogue
FUSED_MAX_K = 1024        # largest k the fused path takes at all (needs n >= 32 k beyond 96, see fused_supported)
FUSED_MAX_K_GATHER = 96   # largest k for launches that gather their operands from peers inside the kernel


def fused_supported(kmax, n, gather=False):
    """Does the filter + certified re-rank path take this k on a shard of n rows?  Each (query,
    chunk) list keeps 32 keys and k gets one list per 8; beyond 256 the filter runs as several
    launches over row ranges of the data (dmlp_fused_supported)."""
    lib = knn_lib()
    lib.dmlp_fused_supported.argtypes = [C.c_int, C.c_int64]
    return lib.dmlp_fused_supported(int(kmax), int(n)) >= (2 if gather else 1)

_ready = False
_vp, _i, _i64, _u64, _sz, _dbl = C.c_void_p, C.c_int, C.c_int64, C.c_uint64, C.c_size_t, C.c_double


def _lib():
    global _ready
    lib = knn_lib()
    if not _ready:
        lib.dmlp_last_error.restype = C.c_char_p
        lib.dmlp_device_info.argtypes = [C.POINTER(_i)] * 3
        lib.dmlp_stats_reset.argtypes = [_vp, _vp]
        lib.dmlp_stats_update.argtypes = [_vp, _i64, _vp, _vp]
        lib.dmlp_pack.argtypes = [_vp, _i64, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp]
        lib.dmlp_topk_plan.argtypes = [_i, _i, _i, _i, _i, C.POINTER(_i * 8)]
        lib.dmlp_dist_topk.argtypes = [_vp, _vp, _vp, _i, _i, _i, _i, _i, _vp, _vp]
        lib.dmlp_rerank.argtypes = [_vp, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _i,
                                    _vp, _vp, _vp, _vp, _vp, _vp, _sz, _vp]
        lib.dmlp_rerank_workspace_bytes.argtypes = [_i, _i]
        lib.dmlp_rerank_workspace_bytes.restype = _sz
        lib.dmlp_exact_workspace_bytes.argtypes = [_i, _i, _i]
        lib.dmlp_exact_workspace_bytes.restype = _sz
        lib.dmlp_exact.argtypes = [_vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _vp, _vp, _vp, _vp, _sz, _vp]
        lib.dmlp_finalize.argtypes = [_vp, _vp, _vp, _i64, _i, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp, _sz, _vp]
        lib.dmlp_finalize_workspace_bytes.argtypes = [_i, _i, _i]
        lib.dmlp_finalize_workspace_bytes.restype = _sz
        lib.dmlp_synth_attrs.argtypes = [_vp, _i64, _i, _i64, _dbl, _dbl, _u64, _u64, _vp]
        lib.dmlp_synth_labels.argtypes = [_vp, _i64, _i64, _i, _u64, _vp]
        lib.dmlp_synth_k.argtypes = [_vp, _i64, _i64, _i, _i, _u64, _vp]
        lib.dmlp_knn_local_workspace_bytes.argtypes = [_i, _i, _i, _i, _i]
        lib.dmlp_knn_local_workspace_bytes.restype = _sz
        lib.dmlp_knn_local.argtypes = [_vp, _vp, _i, _vp, _vp, _i, _i, _i, _i, _vp, _vp, _sz, _vp, _vp, _vp, _vp,
                                       _vp, _vp]
        _ready = True
    return lib


def _check(rc):
    if rc != 0:
        raise RuntimeError("dmlp_knn: " + _lib().dmlp_last_error().decode())


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _p(t):
    return C.c_void_p(0 if t is None else t.data_ptr())


def pad64(a):
    return (max(a, 1) + 63) // 64 * 64


_sm_cache = {}


def sm_count():
    dev = torch.cuda.current_device()
    if dev not in _sm_cache:
        sm, ma, mi = _i(), _i(), _i()
        _check(_lib().dmlp_device_info(C.byref(sm), C.byref(ma), C.byref(mi)))
        _sm_cache[dev] = sm.value
    return _sm_cache[dev]


def new_stats(device=None):
    """A zeroed 32-byte stats block (as 4 float64 words), reset on the device."""
    st = torch.empty(4, dtype=torch.float64, device=device or "cuda")
    _check(_lib().dmlp_stats_reset(_p(st), _stream()))
    return st


def stats_reset(st):
    _check(_lib().dmlp_stats_reset(_p(st), _stream()))


def stats_update(src, st):
    _check(_lib().dmlp_stats_update(_p(src), src.numel(), _p(st), _stream()))


def stats_set(st, lo, hi):
    """Install a host-computed value range into a stats block."""
    lib = _lib()
    lib.dmlp_stats_set.argtypes = [_vp, _dbl, _dbl, _vp]
    _check(lib.dmlp_stats_set(_p(st), float(lo), float(hi), _stream()))


def encode_decimal6(src):
    """fp64 tensor -> (int32 tensor, ok): the decimal6 wire form, ok=False if any value does not round-trip."""
    lib = _lib()
    lib.dmlp_encode_decimal6.argtypes = [_vp, _vp, _i64, _vp, _vp]
    dst = torch.empty(src.shape, dtype=torch.int32, device=src.device)
    bad = torch.zeros(1, dtype=torch.int32, device=src.device)
    _check(lib.dmlp_encode_decimal6(_p(src), _p(dst), src.numel(), _p(bad), _stream()))
    return dst, int(bad.item()) == 0


def expand_decimal6(src, dst):
    lib = _lib()
    lib.dmlp_expand_decimal6.argtypes = [_vp, _vp, _i64, _vp]
    _check(lib.dmlp_expand_decimal6(_p(src), _p(dst), src.numel(), _stream()))
    return dst


def pack(src, rng, maxima=None, out=None):
    """fp64 [rows, a] -> (fp16 [rows, a_pad], norm2 [rows], errn [rows])."""
    rows, a = src.shape
    a_pad = pad64(a)
    if out is None:
        out = (torch.empty((rows, a_pad), dtype=torch.float16, device=src.device),
               torch.empty(rows, dtype=torch.float32, device=src.device),
               torch.empty(rows, dtype=torch.float32, device=src.device))
    h, n2, en = out
    _check(_lib().dmlp_pack(_p(src), rows, a, a_pad, _p(rng), _p(h), _p(n2), _p(en), _p(maxima), _stream()))
    return h, n2, en


def topk_plan(q, n, a_pad, sms=None, kmax=0):
    arr = (_i * 8)()
    _check(_lib().dmlp_topk_plan(q, n, a_pad, sms or sm_count(), kmax, C.byref(arr)))
    keys = ["m_blocks", "n_tiles", "group_m", "chunks", "k_blocks", "grid", "kp", "cluster"]
    return dict(zip(keys, list(arr)[:8]))


def dist_topk(qh, xh, xnorm2, cand=None, sms=None, kmax=0):
    """Filter GEMM + fused top-k.  Returns uint64-as-int64 keys [q, chunks, kp]."""
    q, a_pad = qh.shape
    n = xh.shape[0]
    plan = topk_plan(q, n, a_pad, sms, kmax)
    if cand is None:
        cand = torch.empty((q, plan["chunks"], plan["kp"]), dtype=torch.int64, device=qh.device)
    _check(_lib().dmlp_dist_topk(_p(qh), _p(xh), _p(xnorm2), q, n, a_pad, sms or sm_count(), kmax, _p(cand),
                                 _stream()))
    return cand


_rerank_ws = {}


def rerank_workspace(q, kmax, device):
    """Cached scratch (survivor pair list) for the re-rank kernels."""
    key = (str(device), q, kmax)
    if key not in _rerank_ws:
        _rerank_ws[key] = torch.empty(int(_lib().dmlp_rerank_workspace_bytes(q, kmax)), dtype=torch.uint8,
                                      device=device)
    return _rerank_ws[key]


def rerank(cand, q_attrs, x_attrs, x_labels, k, q_norm2, q_errn, rng, x_maxima, id_base, kmax, out=None):
    q, chunks, kp = cand.shape
    n, a = x_attrs.shape
    dev = cand.device
    if out is None:
        out = (torch.empty((q, kmax), dtype=torch.float64, device=dev),
               torch.empty((q, kmax), dtype=torch.int32, device=dev),
               torch.empty((q, kmax), dtype=torch.int32, device=dev),
               torch.empty(q, dtype=torch.int32, device=dev),
               torch.zeros(1, dtype=torch.int32, device=dev))
    od, oi, ol, flags, nflag = out
    ws = rerank_workspace(q, kmax, dev)
    _check(_lib().dmlp_rerank(_p(cand), chunks, kp, _p(q_attrs), _p(x_attrs), _p(x_labels), _p(k), _p(q_norm2),
                              _p(q_errn), _p(rng), _p(x_maxima), q, n, a, pad64(a), id_base, kmax,
                              _p(od), _p(oi), _p(ol), _p(flags), _p(nflag), _p(ws), ws.numel(), _stream()))
    return od, oi, ol, flags, nflag


def exact_workspace_bytes(n, count, kmax):
    return int(_lib().dmlp_exact_workspace_bytes(n, count, kmax))


def exact(q_attrs, x_attrs, x_labels, k, kmax, id_base=0, query_rows=None, out=None, workspace=None, x_i32=None):
    """Direct fp64 brute force (streaming bounded lists, any k) for all query rows, or the subset
    `query_rows` (int32)."""
    q, a = q_attrs.shape
    n = (x_attrs if x_attrs is not None else x_i32).shape[0]
    dev = q_attrs.device
    count = q if query_rows is None else int(query_rows.numel())
    if out is None:
        out = (torch.empty((q, kmax), dtype=torch.float64, device=dev),
               torch.empty((q, kmax), dtype=torch.int32, device=dev),
               torch.empty((q, kmax), dtype=torch.int32, device=dev))
    if workspace is None:
        workspace = torch.empty(exact_workspace_bytes(n, count, kmax), dtype=torch.uint8, device=dev)
    od, oi, ol = out
    if x_i32 is not None:      # decimal6 rows
        lib = _lib()
        lib.dmlp_exact_ex.argtypes = [C.POINTER(ExactArgs), _vp]
        g = ExactArgs()
        g.q_attrs, g.x_attrs, g.x_labels, g.k, g.query_rows = _ptr(q_attrs), 0, _ptr(x_labels), _ptr(k), _ptr(query_rows)
        g.count, g.n, g.a, g.id_base, g.kmax = count, n, a, id_base, kmax
        g.out_dist, g.out_id, g.out_label = _ptr(od), _ptr(oi), _ptr(ol)
        g.workspace, g.workspace_bytes, g.x_attrs_i32 = _ptr(workspace), workspace.numel(), _ptr(x_i32)
        _check(lib.dmlp_exact_ex(C.byref(g), _stream()))
        return od, oi, ol
    _check(_lib().dmlp_exact(_p(q_attrs), _p(x_attrs), _p(x_labels), _p(k), _p(query_rows), count, n, a, id_base,
                             kmax, _p(od), _p(oi), _p(ol), _p(workspace), workspace.numel(), _stream()))
    return od, oi, ol


def finalize(dist, ids, labels, k, want_lists=False):
    """Merge [sources, q, kmax] exact lists, vote, checksum."""
    sources, q, kmax = dist.shape
    dev = dist.device
    out_label = torch.empty(q, dtype=torch.int32, device=dev)
    out_sum = torch.empty(q, dtype=torch.int64, device=dev)  # uint64 bit pattern
    out_id = torch.empty((q, kmax), dtype=torch.int32, device=dev) if want_lists else None
    out_dist = torch.empty((q, kmax), dtype=torch.float64, device=dev) if want_lists else None
    wsb = int(_lib().dmlp_finalize_workspace_bytes(sources, q, kmax))
    ws = torch.empty(wsb, dtype=torch.uint8, device=dev) if wsb else None
    _check(_lib().dmlp_finalize(_p(dist), _p(ids), _p(labels), dist.stride(0), sources, q, kmax, _p(k), _p(out_label),
                                _p(out_sum), _p(out_id), _p(out_dist), _p(ws), wsb, _stream()))
    return out_label, out_sum, out_id, out_dist


def synth_attrs(rows, a, row_base, lo, hi, seed, stream_id, device="cuda", out=None):
    if out is None:
        out = torch.empty((rows, a), dtype=torch.float64, device=device)
    _check(_lib().dmlp_synth_attrs(_p(out), rows, a, row_base, lo, hi, seed, stream_id, _stream()))
    return out


def synth_labels(rows, row_base, num_labels, seed, device="cuda"):
    out = torch.empty(rows, dtype=torch.int32, device=device)
    _check(_lib().dmlp_synth_labels(_p(out), rows, row_base, num_labels, seed, _stream()))
    return out


def synth_k(rows, row_base, min_k, max_k, seed, device="cuda"):
    out = torch.empty(rows, dtype=torch.int32, device=device)
    _check(_lib().dmlp_synth_k(_p(out), rows, row_base, min_k, max_k, seed, _stream()))
    return out


class GatherArgs(C.Structure):
    _fields_ = [("world", _i), ("rank", _i), ("slice_rows", _i), ("a", _i), ("epoch", C.c_uint32),
                ("peer_base", _vp * 8), ("off_qh", _u64), ("off_q64", _u64), ("off_qn", _u64), ("off_qe", _u64),
                ("off_k", _u64), ("off_pack_flags", _u64), ("rb_count", _vp), ("status", _vp),
                ("timeout_seconds", _dbl)]


class ScatterArgs(C.Structure):
    _fields_ = [("world", _i), ("rank", _i), ("slice_rows", _i), ("peer_base", _vp * 8), ("off_dist", _u64),
                ("off_id", _u64), ("off_label", _u64)]


class PeerGatherArgs(C.Structure):
    _fields_ = [("world", _i), ("rank", _i), ("epoch", C.c_uint32), ("peer_base", _vp * 8), ("off_flags", _u64),
                ("part_off", _u64 * 8), ("part_len", _u64 * 8), ("status", _vp), ("timeout_seconds", _dbl)]


def peer_gather(args):
    """Pull every other rank's part of a symmetric region over NVLink (in-kernel all-gather)."""
    lib = _lib()
    lib.dmlp_peer_gather.argtypes = [C.POINTER(PeerGatherArgs), _vp]
    _check(lib.dmlp_peer_gather(C.byref(args), _stream()))


def dist_topk_part(qh, xh_part, xnorm2_part, cand, chunk_base, row_base, kmax=0):
    """Filter GEMM on a slice of the data shard, writing lists [chunk_base, ...) of `cand`
    ([q, cand_chunks, kp]) with shard-wide row ids."""
    lib = _lib()
    lib.dmlp_dist_topk_part.argtypes = [_vp, _vp, _vp, _i, _i, _i, _i, _i, _vp, _i, _i, _i, _vp]
    q, a_pad = qh.shape
    _check(lib.dmlp_dist_topk_part(_p(qh), _p(xh_part), _p(xnorm2_part), q, xh_part.shape[0], a_pad, sm_count(), kmax,
                                   _p(cand), cand.shape[1], chunk_base, row_base, _stream()))
    return cand


class DataGatherArgs(C.Structure):
    _fields_ = [("world", _i), ("rank", _i), ("a", _i), ("epoch", C.c_uint32), ("peer_base", _vp * 8),
                ("off_xh", _u64), ("off_xn", _u64), ("off_xe", _u64), ("off_x64", _u64), ("off_labels", _u64),
                ("off_pack_flags", _u64), ("part_begin", _i * 9), ("tile_count", _vp), ("status", _vp),
                ("timeout_seconds", _dbl), ("done_signal", _vp), ("x_i32", _i)]


def dist_topk_rows(qh, n, cand, gather, kmax=0):
    """Filter GEMM whose data operand is all-gathered from the row's peers inside the kernel."""
    lib = _lib()
    lib.dmlp_dist_topk_rows.argtypes = [_vp, _i, _i, _i, _i, _i, _vp, C.POINTER(DataGatherArgs), _vp]
    q, a_pad = qh.shape
    _check(lib.dmlp_dist_topk_rows(_p(qh), q, n, a_pad, sm_count(), kmax, _p(cand), C.byref(gather), _stream()))
    return cand


def maxima_from_norms(norm2, errn, st):
    lib = _lib()
    lib.dmlp_maxima_from_norms.argtypes = [_vp, _vp, _i64, _vp, _vp]
    _check(lib.dmlp_maxima_from_norms(_p(norm2), _p(errn), norm2.numel(), _p(st), _stream()))


def dist_topk_gather(xh, xnorm2, cand, gather, kmax=0):
    """K4: filter GEMM whose query operand is all-gathered from the peers inside the kernel."""
    lib = _lib()
    lib.dmlp_dist_topk_gather.argtypes = [_vp, _vp, _i, _i, _i, _i, _vp, C.POINTER(GatherArgs), _vp]
    n, a_pad = xh.shape
    _check(lib.dmlp_dist_topk_gather(_p(xh), _p(xnorm2), n, a_pad, sm_count(), kmax, _p(cand), C.byref(gather),
                                     _stream()))
    return cand


def rerank_scatter(cand, q_attrs, x_attrs, x_labels, k, q_norm2, q_errn, rng, x_maxima, id_base, kmax, flags, nflag,
                   scatter):
    """K5: exact re-rank whose k-lists are stored straight into the owners' receive buffers."""
    lib = _lib()
    lib.dmlp_rerank_scatter.argtypes = [_vp, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _i,
                                        _vp, _vp, C.POINTER(ScatterArgs), _vp, _sz, _vp]
    q, chunks, kp = cand.shape
    n, a = x_attrs.shape
    ws = rerank_workspace(q, kmax, cand.device)
    _check(lib.dmlp_rerank_scatter(_p(cand), chunks, kp, _p(q_attrs), _p(x_attrs), _p(x_labels), _p(k), _p(q_norm2),
                                   _p(q_errn), _p(rng), _p(x_maxima), q, n, a, pad64(a), id_base, kmax, _p(flags),
                                   _p(nflag), C.byref(scatter), _p(ws), ws.numel(), _stream()))


def knn_local_workspace_bytes(q, n, a, kmax=0):
    return int(_lib().dmlp_knn_local_workspace_bytes(q, n, a, sm_count(), kmax))


def knn_local(q_attrs, k, x_attrs, x_labels, kmax, id_base=0, rng=None, workspace=None, out=None):
    """stats -> pack -> dist_topk -> rerank in one call on caller-provided scratch."""
    q, a = q_attrs.shape
    n = x_attrs.shape[0]
    dev = q_attrs.device
    if workspace is None:
        workspace = torch.empty(knn_local_workspace_bytes(q, n, a, kmax), dtype=torch.uint8, device=dev)
    if out is None:
        out = (torch.empty((q, kmax), dtype=torch.float64, device=dev),
               torch.empty((q, kmax), dtype=torch.int32, device=dev),
               torch.empty((q, kmax), dtype=torch.int32, device=dev),
               torch.empty(q, dtype=torch.int32, device=dev),
               torch.zeros(1, dtype=torch.int32, device=dev))
    od, oi, ol, flags, nflag = out
    _check(_lib().dmlp_knn_local(_p(q_attrs), _p(k), q, _p(x_attrs), _p(x_labels), n, a, id_base, kmax, _p(rng),
                                 _p(workspace), workspace.numel(), _p(od), _p(oi), _p(ol), _p(flags), _p(nflag),
                                 _stream()))
    return od, oi, ol, flags, nflag


# ---- struct-based entry points: folded vote, device-side repair, in-kernel K5 signalling --------
class SignalArgs(C.Structure):
    _fields_ = [("peer_flags", _vp * 8), ("world", _i), ("rank", _i), ("value", C.c_uint32), ("done_counter", _vp),
                ("mc_flags", _vp)]   # NVSwitch multicast mapping of the flag arrays (None: one store per peer)


class WaitArgs(C.Structure):
    _fields_ = [("flags", _vp), ("world", _i), ("value", C.c_uint32), ("timeout_seconds", _dbl), ("status", _vp)]


class RerankArgs(C.Structure):
    _fields_ = [("cand", _vp), ("chunks", _i), ("kp", _i), ("q_attrs", _vp), ("x_attrs", _vp), ("x_labels", _vp),
                ("k", _vp), ("q_norm2", _vp), ("q_errn", _vp), ("range", _vp), ("x_maxima", _vp), ("q", _i), ("n", _i),
                ("a", _i), ("a_pad", _i), ("id_base", _i), ("kmax", _i), ("out_dist", _vp), ("out_id", _vp),
                ("out_label", _vp), ("flags", _vp), ("num_flagged", _vp), ("flagged_rows", _vp), ("flagged_cap", _i),
                ("final_label", _vp), ("final_checksum", _vp), ("scatter", C.POINTER(ScatterArgs)),
                ("signal", C.POINTER(SignalArgs)), ("workspace", _vp), ("workspace_bytes", _sz), ("x_attrs_i32", _vp)]


class RepairArgs(C.Structure):
    _fields_ = [("q_attrs", _vp), ("x_attrs", _vp), ("x_labels", _vp), ("k", _vp), ("flagged_rows", _vp),
                ("num_flagged", _vp), ("cap", _i), ("q", _i), ("n", _i), ("a", _i), ("id_base", _i), ("kmax", _i),
                ("out_dist", _vp), ("out_id", _vp), ("out_label", _vp), ("final_label", _vp), ("final_checksum", _vp),
                ("scatter", C.POINTER(ScatterArgs)), ("signal", C.POINTER(SignalArgs)), ("workspace", _vp),
                ("workspace_bytes", _sz), ("x_attrs_i32", _vp)]


class ExactArgs(C.Structure):
    _fields_ = [("q_attrs", _vp), ("x_attrs", _vp), ("x_labels", _vp), ("k", _vp), ("query_rows", _vp), ("count", _i),
                ("n", _i), ("a", _i), ("id_base", _i), ("kmax", _i), ("out_dist", _vp), ("out_id", _vp),
                ("out_label", _vp), ("scatter", C.POINTER(ScatterArgs)), ("signal", C.POINTER(SignalArgs)),
                ("workspace", _vp), ("workspace_bytes", _sz), ("x_attrs_i32", _vp)]


class FinalizeArgs(C.Structure):
    _fields_ = [("dist", _vp), ("id", _vp), ("label", _vp), ("source_stride", _i64), ("sources", _i), ("q", _i),
                ("kmax", _i), ("k", _vp), ("out_label", _vp), ("out_checksum", _vp), ("out_id", _vp), ("out_dist", _vp),
                ("query_rows", _vp), ("count", _i), ("count_dev", _vp), ("wait", C.POINTER(WaitArgs)),
                ("workspace", _vp), ("workspace_bytes", _sz)]


REPAIR_CAP = 64   # flagged queries a step repairs without asking the host (more: the caller's end-of-step check)


def _ptr(t):
    return 0 if t is None else t.data_ptr()


def rerank_ex(cand, q_attrs, x_attrs, x_labels, k, q_norm2, q_errn, rng, x_maxima, id_base, kmax, flags, nflag,
              out=None, final=None, flagged_rows=None, scatter=None, signal=None, x_i32=None):
    """dmlp_rerank_ex: re-rank with the optional folded vote (`final` = (label, checksum) tensors, single
    data shard), flagged-row list, K5 scatter and in-kernel completion signal."""
    lib = _lib()
    lib.dmlp_rerank_ex.argtypes = [C.POINTER(RerankArgs), _vp]
    q, chunks, kp = cand.shape
    n, a = (x_attrs if x_attrs is not None else x_i32).shape
    ws = rerank_workspace(q, kmax, cand.device)
    g = RerankArgs()
    g.cand, g.chunks, g.kp = _ptr(cand), chunks, kp
    g.q_attrs, g.x_attrs, g.x_labels, g.k = _ptr(q_attrs), _ptr(x_attrs), _ptr(x_labels), _ptr(k)
    g.x_attrs_i32 = _ptr(x_i32)   # decimal6 rows: read instead of x_attrs (which may then be None)
    g.q_norm2, g.q_errn, g.range, g.x_maxima = _ptr(q_norm2), _ptr(q_errn), _ptr(rng), _ptr(x_maxima)
    g.q, g.n, g.a, g.a_pad, g.id_base, g.kmax = q, n, a, pad64(a), id_base, kmax
    if out is not None:
        g.out_dist, g.out_id, g.out_label = _ptr(out[0]), _ptr(out[1]), _ptr(out[2])
    g.flags, g.num_flagged = _ptr(flags), _ptr(nflag)
    if flagged_rows is not None:
        g.flagged_rows, g.flagged_cap = _ptr(flagged_rows), flagged_rows.numel()
    if final is not None:
        g.final_label, g.final_checksum = _ptr(final[0]), _ptr(final[1])
    if scatter is not None:
        g.scatter = C.pointer(scatter)
    if signal is not None:
        g.signal = C.pointer(signal)
    g.workspace, g.workspace_bytes = _ptr(ws), ws.numel()
    _check(lib.dmlp_rerank_ex(C.byref(g), _stream()))


_repair_ws = {}


def repair_flagged(q_attrs, x_attrs, x_labels, k, flagged_rows, nflag, id_base, kmax, out=None, final=None,
                   scatter=None, signal=None, x_i32=None):
    """dmlp_repair_flagged: direct fp64 kernel over the first min(*nflag, cap) flagged rows; the count
    stays on the device, so the call is enqueued unconditionally."""
    lib = _lib()
    lib.dmlp_repair_flagged.argtypes = [C.POINTER(RepairArgs), _vp]
    q, a = q_attrs.shape
    n = (x_attrs if x_attrs is not None else x_i32).shape[0]
    cap = flagged_rows.numel()
    key = (str(q_attrs.device), n, cap, kmax)
    if key not in _repair_ws:
        _repair_ws[key] = torch.empty(exact_workspace_bytes(n, cap, kmax), dtype=torch.uint8, device=q_attrs.device)
    ws = _repair_ws[key]
    g = RepairArgs()
    g.q_attrs, g.x_attrs, g.x_labels, g.k = _ptr(q_attrs), _ptr(x_attrs), _ptr(x_labels), _ptr(k)
    g.flagged_rows, g.num_flagged, g.cap = _ptr(flagged_rows), _ptr(nflag), cap
    g.x_attrs_i32 = _ptr(x_i32)
    g.q, g.n, g.a, g.id_base, g.kmax = q, n, a, id_base, kmax
    if out is not None:
        g.out_dist, g.out_id, g.out_label = _ptr(out[0]), _ptr(out[1]), _ptr(out[2])
    if final is not None:
        g.final_label, g.final_checksum = _ptr(final[0]), _ptr(final[1])
    if scatter is not None:
        g.scatter = C.pointer(scatter)
    if signal is not None:
        g.signal = C.pointer(signal)
    g.workspace, g.workspace_bytes = _ptr(ws), ws.numel()
    _check(lib.dmlp_repair_flagged(C.byref(g), _stream()))


def finalize_ex(dist, ids, labels, k, out=None, query_rows=None, count_dev=None, wait=None):
    """dmlp_finalize_ex: finalize with an in-kernel wait for the peers' lists and / or a row subset."""
    lib = _lib()
    lib.dmlp_finalize_ex.argtypes = [C.POINTER(FinalizeArgs), _vp]
    sources, q, kmax = dist.shape
    dev = dist.device
    if out is None:
        out = (torch.empty(q, dtype=torch.int32, device=dev), torch.empty(q, dtype=torch.int64, device=dev))
    wsb = int(lib.dmlp_finalize_workspace_bytes(sources, q, kmax))
    ws = torch.empty(wsb, dtype=torch.uint8, device=dev) if wsb else None
    g = FinalizeArgs()
    g.dist, g.id, g.label, g.source_stride = _ptr(dist), _ptr(ids), _ptr(labels), dist.stride(0)
    g.sources, g.q, g.kmax, g.k = sources, q, kmax, _ptr(k)
    g.out_label, g.out_checksum = _ptr(out[0]), _ptr(out[1])
    if query_rows is not None:
        g.query_rows, g.count, g.count_dev = _ptr(query_rows), query_rows.numel(), _ptr(count_dev)
    if wait is not None:
        g.wait = C.pointer(wait)
    g.workspace, g.workspace_bytes = _ptr(ws), wsb
    _check(lib.dmlp_finalize_ex(C.byref(g), _stream()))
    return out


def pack_ex(src, rng, maxima=None, out=None, wait=None, signal=None):
    """dmlp_pack_ex: pack that first waits for the peers to release the rows it overwrites (`wait`)
    and whose last block announces the new rows (`signal`)."""
    lib = _lib()
    lib.dmlp_pack_ex.argtypes = [_vp, _i64, _i, _i, _vp, _vp, _vp, _vp, _vp, C.POINTER(WaitArgs), C.POINTER(SignalArgs),
                                 _vp]
    rows, a = src.shape
    h, n2, en = out
    _check(lib.dmlp_pack_ex(_p(src), rows, a, pad64(a), _p(rng), _p(h), _p(n2), _p(en), _p(maxima),
                            C.byref(wait) if wait is not None else None,
                            C.byref(signal) if signal is not None else None, _stream()))
    return h, n2, en


def zero_(t):
    """cudaMemsetAsync on the current stream (no library kernel)."""
    lib = _lib()
    lib.dmlp_zero.argtypes = [_vp, _sz, _vp]
    _check(lib.dmlp_zero(_p(t), t.numel() * t.element_size(), _stream()))
    return t


def topk_sync_timeouts():
    """Re-alignment waits of the filter kernel that timed out on the current device so far."""
    lib = _lib()
    lib.dmlp_topk_sync_timeouts.argtypes = [C.POINTER(C.c_uint)]
    out = C.c_uint(0)
    _check(lib.dmlp_topk_sync_timeouts(C.byref(out)))
    return int(out.value)


def pack_i32(src, rng, maxima=None, out=None, wait=None, signal=None):
    """dmlp_pack_i32: pack rows kept in decimal6 form (int32 i, value i / 1e6)."""
    lib = _lib()
    lib.dmlp_pack_i32.argtypes = [_vp, _i64, _i, _i, _vp, _vp, _vp, _vp, _vp, C.POINTER(WaitArgs), C.POINTER(SignalArgs),
                                  _vp]
    rows, a = src.shape
    if out is None:
        out = (torch.empty((rows, pad64(a)), dtype=torch.float16, device=src.device),
               torch.empty(rows, dtype=torch.float32, device=src.device),
               torch.empty(rows, dtype=torch.float32, device=src.device))
    h, n2, en = out
    _check(lib.dmlp_pack_i32(_p(src), rows, a, pad64(a), _p(rng), _p(h), _p(n2), _p(en), _p(maxima),
                             C.byref(wait) if wait is not None else None,
                             C.byref(signal) if signal is not None else None, _stream()))
    return h, n2, en


def wire_words(a, bits):
    """32-bit words per row of `a` values at `bits` bits each."""
    return (a * bits + 31) // 32


def pack_bits(src_i32, bits, out=None):
    """dmlp_pack_bits: int32 [rows, a] -> int32-typed words [rows, wire_words(a, bits)]."""
    lib = _lib()
    lib.dmlp_pack_bits.argtypes = [_vp, _i64, _i, _i, _i, _vp, _vp]
    rows, a = src_i32.shape
    words = wire_words(a, bits)
    if out is None:
        out = torch.empty((rows, words), dtype=torch.int32, device=src_i32.device)
    _check(lib.dmlp_pack_bits(_p(src_i32), rows, a, bits, words, _p(out), _stream()))
    return out


def unpack_bits(src_words, a, bits, dst):
    """dmlp_unpack_bits: words [rows, wire_words(a, bits)] -> int32 or float64 [rows, a] (by dst's dtype)."""
    lib = _lib()
    lib.dmlp_unpack_bits.argtypes = [_vp, _i64, _i, _i, _i, _vp, _vp, _vp]
    rows, words = src_words.shape
    assert words == wire_words(a, bits) and tuple(dst.shape) == (rows, a) and dst.is_contiguous()
    assert src_words.is_contiguous()
    is_f64 = dst.dtype == torch.float64
    assert is_f64 or dst.dtype == torch.int32
    _check(lib.dmlp_unpack_bits(_p(src_words), rows, a, bits, words, None if is_f64 else _p(dst),
                                _p(dst) if is_f64 else None, _stream()))
    return dst


def encode_decimal6_into(src, dst):
    """Re-encode fp64 values into an existing int32 tensor (no check read back: the caller knows
    the values are representable)."""
    lib = _lib()
    lib.dmlp_encode_decimal6.argtypes = [_vp, _vp, _i64, _vp, _vp]
    bad = torch.zeros(1, dtype=torch.int32, device=src.device)
    _check(lib.dmlp_encode_decimal6(_p(src), _p(dst), src.numel(), _p(bad), _stream()))
    return dst
