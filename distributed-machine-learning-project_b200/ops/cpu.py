"""CPU ops (ctypes over libdmlp_cpu.so): oracle, strategies, parser, generator."""
import ctypes as C

import numpy as np

from .._native import cpu_lib

_d = np.ctypeslib.ndpointer(dtype=np.float64, flags="C_CONTIGUOUS")
_i = np.ctypeslib.ndpointer(dtype=np.int32, flags="C_CONTIGUOUS")
_u = np.ctypeslib.ndpointer(dtype=np.uint64, flags="C_CONTIGUOUS")

_ready = False


def _lib():
    global _ready
    lib = cpu_lib()
    if not _ready:
        lib.dmlp_cpu_last_error.restype = C.c_char_p
        lib.dmlp_cpu_knn.argtypes = [_d, _i, C.c_int64, _d, _i, C.c_int64, C.c_int, C.c_char_p,
                                     C.c_int, C.c_int, C.c_int, C.c_int, _i, _u, _i, _d]
        lib.dmlp_exact_sqdist.argtypes = [_d, _d, C.c_int]
        lib.dmlp_exact_sqdist.restype = C.c_double
        lib.dmlp_result_checksum.argtypes = [C.c_int, _i, C.c_int64]
        lib.dmlp_result_checksum.restype = C.c_uint64
        lib.dmlp_merge_topk.argtypes = [_d, _i, _d, _i, C.c_int, _d, _i]
        lib.dmlp_cpu_local_lists.argtypes = [_d, _i, C.c_int64, C.c_int, _d, _i, C.c_int64, C.c_int, C.c_int, C.c_int,
                                             _d, _i, _i]
        lib.dmlp_cpu_finalize_lists.argtypes = [_d, _i, _i, C.c_int, C.c_int64, C.c_int, _i, _i, _u]
        lib.dmlp_parse_header.argtypes = [C.c_char_p, C.c_size_t, C.POINTER(C.c_int64),
                                          C.POINTER(C.c_int64), C.POINTER(C.c_int)]
        lib.dmlp_parse_text.argtypes = [C.c_char_p, C.c_size_t, C.c_int, _d, _i, _d, _i]
        lib.dmlp_parse_update.argtypes = [C.c_char_p, C.c_size_t, C.POINTER(C.c_int), _d, C.c_int,
                                          C.POINTER(C.c_int)]
        lib.dmlp_synth_fill.argtypes = [C.c_int64, C.c_int64, C.c_int, C.c_double, C.c_double, C.c_int,
                                        C.c_int, C.c_int, C.c_uint64, _d, _i, _d, _i]
        lib.dmlp_choose_grid.argtypes = [C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int)]
        lib.dmlp_block_begin.argtypes = [C.c_int64, C.c_int, C.c_int]
        lib.dmlp_block_begin.restype = C.c_int64
        lib.dmlp_block_count.argtypes = [C.c_int64, C.c_int, C.c_int]
        lib.dmlp_block_count.restype = C.c_int64
        _ready = True
    return lib


def _check(rc):
    if rc != 0:
        raise RuntimeError(_lib().dmlp_cpu_last_error().decode())


def knn(data, labels, queries, k, strategy="serial", ranks=1, rows=0, cols=0, threads=0):
    """Run the whole job on the CPU.  Returns (label[q], checksum[q], ids, dist, offsets)."""
    data = np.ascontiguousarray(data, dtype=np.float64)
    queries = np.ascontiguousarray(queries, dtype=np.float64)
    labels = np.ascontiguousarray(labels, dtype=np.int32)
    k = np.ascontiguousarray(k, dtype=np.int32)
    n, q = data.shape[0], queries.shape[0]
    a = data.shape[1] if data.ndim == 2 else queries.shape[1]
    off = np.zeros(q + 1, dtype=np.int64)
    np.cumsum(np.maximum(k, 0), out=off[1:])
    out_label = np.empty(q, dtype=np.int32)
    out_sum = np.empty(q, dtype=np.uint64)
    out_ids = np.empty(int(off[-1]), dtype=np.int32)
    out_dist = np.empty(int(off[-1]), dtype=np.float64)
    _check(_lib().dmlp_cpu_knn(data.reshape(-1), labels, n, queries.reshape(-1), k, q, a,
                               strategy.encode(), ranks, rows, cols, threads,
                               out_label, out_sum, out_ids, out_dist))
    return out_label, out_sum, out_ids, out_dist, off


def local_lists(data, labels, id_base, queries, k, kmax, threads=0):
    """Per-shard exact k-lists ([q, kmax] dist / id / label, sentinel padded, global ids)."""
    data = np.ascontiguousarray(data, dtype=np.float64)
    queries = np.ascontiguousarray(queries, dtype=np.float64)
    labels = np.ascontiguousarray(labels, dtype=np.int32)
    k = np.ascontiguousarray(k, dtype=np.int32)
    q, a = queries.shape
    od = np.empty((q, kmax), np.float64)
    oi = np.empty((q, kmax), np.int32)
    ol = np.empty((q, kmax), np.int32)
    _check(_lib().dmlp_cpu_local_lists(data.reshape(-1), labels, data.shape[0], id_base, queries.reshape(-1), k, q, a,
                                       kmax, threads, od.reshape(-1), oi.reshape(-1), ol.reshape(-1)))
    return od, oi, ol


def finalize_lists(dist, ids, labels, k):
    """Merge [sources, q, kmax] lists with the merge operator, vote, checksum."""
    dist = np.ascontiguousarray(dist, dtype=np.float64)
    ids = np.ascontiguousarray(ids, dtype=np.int32)
    labels = np.ascontiguousarray(labels, dtype=np.int32)
    k = np.ascontiguousarray(k, dtype=np.int32)
    sources, q, kmax = dist.shape
    out_label = np.empty(q, np.int32)
    out_sum = np.empty(q, np.uint64)
    _check(_lib().dmlp_cpu_finalize_lists(dist.reshape(-1), ids.reshape(-1), labels.reshape(-1), sources, q, kmax, k,
                                          out_label, out_sum))
    return out_label, out_sum


def exact_sqdist(a, b):
    a = np.ascontiguousarray(a, dtype=np.float64)
    b = np.ascontiguousarray(b, dtype=np.float64)
    return _lib().dmlp_exact_sqdist(a, b, a.shape[0])


def result_checksum(label, ids):
    ids = np.ascontiguousarray(ids, dtype=np.int32)
    return int(_lib().dmlp_result_checksum(int(label), ids, ids.shape[0]))


def merge_topk(xd, xi, yd, yi):
    k = len(xd)
    od, oi = np.empty(k, np.float64), np.empty(k, np.int32)
    _lib().dmlp_merge_topk(np.ascontiguousarray(xd, np.float64), np.ascontiguousarray(xi, np.int32),
                           np.ascontiguousarray(yd, np.float64), np.ascontiguousarray(yi, np.int32), k, od, oi)
    return od, oi


def parse_text(text, threads=0):
    """Parse the reference text grammar.  Returns (data, labels, queries, k)."""
    if isinstance(text, str):
        text = text.encode()
    n, q, a = C.c_int64(), C.c_int64(), C.c_int()
    _check(_lib().dmlp_parse_header(text, len(text), C.byref(n), C.byref(q), C.byref(a)))
    data = np.empty((n.value, a.value), np.float64)
    labels = np.empty(n.value, np.int32)
    queries = np.empty((q.value, a.value), np.float64)
    k = np.empty(q.value, np.int32)
    _check(_lib().dmlp_parse_text(text, len(text), threads, data.reshape(-1), labels,
                                  queries.reshape(-1), k))
    return data, labels, queries, k


def encode_decimal6(values, threads=0):
    """Six-decimal values as int32 (value == i / 1e6 exactly), or None when some value is not representable."""
    v = np.ascontiguousarray(values, dtype=np.float64)
    out = np.empty(v.shape, np.int32)
    ok = C.c_int(0)
    lib = _lib()
    lib.dmlp_encode_decimal6_host.argtypes = [_d, C.c_int64, C.c_int, _i, C.POINTER(C.c_int)]
    _check(lib.dmlp_encode_decimal6_host(v.reshape(-1), v.size, threads, out.reshape(-1), C.byref(ok)))
    return out if ok.value else None


_u32 = np.ctypeslib.ndpointer(dtype=np.uint32, flags="C_CONTIGUOUS")


def decimal6_bits(values_i32):
    """Smallest field width (bits per value) of the bit-packed wire format for these integers."""
    v = np.ascontiguousarray(values_i32, dtype=np.int32)
    bits = C.c_int(0)
    lib = _lib()
    lib.dmlp_decimal6_bits_host.argtypes = [_i, C.c_int64, C.POINTER(C.c_int)]
    _check(lib.dmlp_decimal6_bits_host(v.reshape(-1), v.size, C.byref(bits)))
    return bits.value


def pack_bits(rows_i32, bits):
    """int32 [rows, a] -> uint32 [rows, ceil(a * bits / 32)]: the host side of K.pack_bits / K.unpack_bits."""
    v = np.ascontiguousarray(rows_i32, dtype=np.int32)
    rows, a = v.shape
    words = (a * bits + 31) // 32
    out = np.empty((rows, words), np.uint32)
    lib = _lib()
    lib.dmlp_pack_bits_host.argtypes = [_i, C.c_int64, C.c_int, C.c_int, C.c_int, _u32]
    _check(lib.dmlp_pack_bits_host(v.reshape(-1), rows, a, bits, words, out.reshape(-1)))
    return out


def unpack_bits(words_u32, a, bits):
    w = np.ascontiguousarray(words_u32, dtype=np.uint32)
    rows, words = w.shape
    out = np.empty((rows, a), np.int32)
    lib = _lib()
    lib.dmlp_unpack_bits_host.argtypes = [_u32, C.c_int64, C.c_int, C.c_int, C.c_int, _i]
    _check(lib.dmlp_unpack_bits_host(w.reshape(-1), rows, a, bits, words, out.reshape(-1)))
    return out


def parse_update(line, max_attrs=1 << 16):
    """Parse one update record 'id v0 v1 ...' (reference: common.cpp:46-55).  Returns (id, values)."""
    if isinstance(line, str):
        line = line.encode()
    ident, count = C.c_int(), C.c_int()
    buf = np.empty(max_attrs, np.float64)
    _check(_lib().dmlp_parse_update(line, len(line), C.byref(ident), buf, max_attrs, C.byref(count)))
    return ident.value, buf[:min(count.value, max_attrs)].copy()


def synth(n, q, a, lo=-100.0, hi=100.0, min_k=1, max_k=1, num_labels=2, seed=42):
    """Host twin of the device fill kernel.  Returns (data, labels, queries, k)."""
    data = np.empty((n, a), np.float64)
    labels = np.empty(n, np.int32)
    queries = np.empty((q, a), np.float64)
    k = np.empty(q, np.int32)
    _check(_lib().dmlp_synth_fill(n, q, a, lo, hi, min_k, max_k, num_labels, seed,
                                  data.reshape(-1), labels, queries.reshape(-1), k))
    return data, labels, queries, k


def choose_grid(ranks):
    r, c = C.c_int(), C.c_int()
    _lib().dmlp_choose_grid(ranks, C.byref(r), C.byref(c))
    return r.value, c.value


def block_range(total, parts, p):
    lib = _lib()
    b = lib.dmlp_block_begin(total, parts, p)
    return int(b), int(b + lib.dmlp_block_count(total, parts, p))
