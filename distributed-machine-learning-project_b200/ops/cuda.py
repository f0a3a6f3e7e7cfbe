"""sm_100a ops (ctypes over libdmlp_knn.so) on torch CUDA tensors.

Each wrapper launches on the current torch CUDA stream and never falls back to
PyTorch math: if the library is missing or a launch fails, it raises.
"""
import ctypes as C

import torch

from .._native import knn_lib

STATS_BYTES = 32
KP = 32          # keys kept per (query, chunk) by the fused epilogue
FUSED_MAX_K = 96  # largest k sent to the fused path: each chunk list keeps 32 keys, so larger k is
#                   given more chunks (one per 8 of k); the re-rank recomputes up to 256 survivors

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


def exact(q_attrs, x_attrs, x_labels, k, kmax, id_base=0, query_rows=None, out=None, workspace=None):
    """Direct fp64 brute force (streaming bounded lists, any k) for all query rows, or the subset
    `query_rows` (int32)."""
    q, a = q_attrs.shape
    n = x_attrs.shape[0]
    dev = q_attrs.device
    count = q if query_rows is None else int(query_rows.numel())
    if out is None:
        out = (torch.empty((q, kmax), dtype=torch.float64, device=dev),
               torch.empty((q, kmax), dtype=torch.int32, device=dev),
               torch.empty((q, kmax), dtype=torch.int32, device=dev))
    if workspace is None:
        workspace = torch.empty(exact_workspace_bytes(n, count, kmax), dtype=torch.uint8, device=dev)
    od, oi, ol = out
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
                ("timeout_seconds", _dbl)]


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
