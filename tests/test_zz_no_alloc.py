"""dmlp_knn_local_static, called straight through the C ABI.

This entry point serves hosts that may not allocate GPU memory while running
(embedded in someone else's CUDA context, real-time loops, fixed memory
reservations): its scratch is a 64 MiB array built into libdmlp_knn.so.  The test
drives it with ctypes, twice in a row on different problems (the second call sees
the first call's leftovers in the shared scratch), and requires exactly the
results of the existing workspace-taking entry point dmlp_knn_local -- fp64
distances compared with ==, like the other GPU tests.  It also checks that an
oversized problem is refused with a clear message rather than run, that an empty
data shard after a normal call yields padding (not the previous call's leftovers),
and that a caller living in its own CUDA context (not the runtime's primary one)
gets that context's copy of the scratch.

Kept in a file that sorts last so a `pytest -x` run has already exercised every
other test before this one can stop it.
"""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu

STATIC_BYTES = 64 << 20
_vp, _i, _sz = C.c_void_p, C.c_int, C.c_size_t


def _abi():
    from dmlp_b200._native import knn_lib
    lib = knn_lib()
    lib.dmlp_last_error.restype = C.c_char_p
    lib.dmlp_knn_local_workspace_bytes.argtypes = [_i, _i, _i, _i, _i]
    lib.dmlp_knn_local_workspace_bytes.restype = _sz
    lib.dmlp_knn_local.argtypes = [_vp, _vp, _i, _vp, _vp, _i, _i, _i, _i, _vp, _vp, _sz, _vp, _vp, _vp, _vp, _vp, _vp]
    lib.dmlp_knn_local.restype = _i
    lib.dmlp_knn_local_static.argtypes = [_vp, _vp, _i, _vp, _vp, _i, _i, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp]
    lib.dmlp_knn_local_static.restype = _i
    return lib


def _outputs(q, kmax):
    return (torch.full((q, kmax), -7.0, dtype=torch.float64, device="cuda"),
            torch.full((q, kmax), -7, dtype=torch.int32, device="cuda"),
            torch.full((q, kmax), -7, dtype=torch.int32, device="cuda"),
            torch.full((q,), -7, dtype=torch.int32, device="cuda"),
            torch.full((1,), -7, dtype=torch.int32, device="cuda"))


def _problem(n, q, a, kmax, seed):
    g = torch.Generator(device="cpu").manual_seed(seed)
    x = (torch.rand((n, a), generator=g, dtype=torch.float64) * 200 - 100).cuda()
    qq = (torch.rand((q, a), generator=g, dtype=torch.float64) * 200 - 100).cuda()
    labels = torch.randint(0, 9, (n,), generator=g, dtype=torch.int32).cuda()
    k = torch.randint(1, kmax + 1, (q,), generator=g, dtype=torch.int32).cuda()
    return x, labels, qq, k


def _run_both(lib, n, q, a, kmax, seed, id_base):
    x, labels, qq, k = _problem(n, q, a, kmax, seed)
    sms = torch.cuda.get_device_properties(0).multi_processor_count
    need = lib.dmlp_knn_local_workspace_bytes(q, n, a, sms, kmax)
    assert need <= STATIC_BYTES, "test problem must fit the built-in scratch"
    ws = torch.empty(need, dtype=torch.uint8, device="cuda")
    want, got = _outputs(q, kmax), _outputs(q, kmax)
    p = lambda t: C.c_void_p(t.data_ptr())  # noqa: E731
    rc = lib.dmlp_knn_local(p(qq), p(k), q, p(x), p(labels), n, a, id_base, kmax, None, p(ws), need,
                            *(p(t) for t in want), None)
    assert rc == 0, lib.dmlp_last_error().decode()
    rc = lib.dmlp_knn_local_static(p(qq), p(k), q, p(x), p(labels), n, a, id_base, kmax, *(p(t) for t in got), None)
    assert rc == 0, lib.dmlp_last_error().decode()
    torch.cuda.synchronize()
    for name, w, g in zip(("dist", "id", "label", "flags", "num_flagged"), want, got):
        assert torch.equal(w, g), f"{name} differs between dmlp_knn_local and dmlp_knn_local_static"
    assert int(got[4].item()) == 0 and (got[1][:, 0] >= id_base).all()


def test_static_scratch_entry_matches_workspace_entry_twice():
    torch.cuda.set_device(0)
    lib = _abi()
    _run_both(lib, n=6000, q=700, a=96, kmax=12, seed=1, id_base=0)
    # second call: smaller, differently shaped problem on the same (now dirty) scratch
    _run_both(lib, n=1500, q=260, a=40, kmax=5, seed=2, id_base=1000)


def test_static_scratch_entry_with_empty_shard_after_a_normal_call():
    """n == 0 right after a real problem: the scratch is full of that problem's candidate keys,
    none of which may leak into the answer."""
    torch.cuda.set_device(0)
    lib = _abi()
    _run_both(lib, n=3000, q=300, a=64, kmax=8, seed=3, id_base=0)
    q, a, kmax = 300, 64, 8
    _, _, qq, k = _problem(16, q, a, kmax, seed=4)
    x = torch.empty((0, a), dtype=torch.float64, device="cuda")
    labels = torch.empty(0, dtype=torch.int32, device="cuda")
    got = _outputs(q, kmax)
    p = lambda t: C.c_void_p(t.data_ptr())  # noqa: E731
    rc = lib.dmlp_knn_local_static(p(qq), p(k), q, p(x), p(labels), 0, a, 0, kmax, *(p(t) for t in got), None)
    assert rc == 0, lib.dmlp_last_error().decode()
    torch.cuda.synchronize()
    assert (got[0] == 1e18).all() and (got[1] == -1).all() and (got[2] == -1).all()
    assert (got[3] == 0).all() and int(got[4].item()) == 0


def test_static_scratch_entry_from_a_second_context():
    """A plug-in inside another program's CUDA context: the `__device__` scratch exists once per
    context, so the address looked up in the primary context must not be reused there.  The
    second context is made with the driver API; all of its buffers are its own allocations."""
    torch.cuda.set_device(0)
    lib = _abi()
    n, q, a, kmax, id_base = 2500, 260, 48, 6, 7
    x, labels, qq, k = _problem(n, q, a, kmax, seed=11)
    want = _outputs(q, kmax)
    p = lambda t: C.c_void_p(t.data_ptr())  # noqa: E731
    rc = lib.dmlp_knn_local_static(p(qq), p(k), q, p(x), p(labels), n, a, id_base, kmax, *(p(t) for t in want), None)
    assert rc == 0, lib.dmlp_last_error().decode()
    torch.cuda.synchronize()
    host_in = [t.cpu().contiguous() for t in (qq, k, x, labels)]
    host_want = [t.cpu() for t in want]

    cu = C.CDLL("libcuda.so.1")
    cu.cuCtxCreate_v2.argtypes = [C.POINTER(_vp), C.c_uint, _i]
    cu.cuMemAlloc_v2.argtypes = [C.POINTER(C.c_uint64), _sz]
    cu.cuMemcpyHtoD_v2.argtypes = [C.c_uint64, _vp, _sz]
    cu.cuMemcpyDtoH_v2.argtypes = [_vp, C.c_uint64, _sz]
    cu.cuMemFree_v2.argtypes = [C.c_uint64]
    cu.cuCtxPopCurrent_v2.argtypes = [C.POINTER(_vp)]
    cu.cuCtxDestroy_v2.argtypes = [_vp]
    assert cu.cuInit(0) == 0
    dev = _i()
    assert cu.cuDeviceGet(C.byref(dev), 0) == 0
    ctx = _vp()
    rc = cu.cuCtxCreate_v2(C.byref(ctx), 0, dev.value)   # becomes current on this thread
    if rc != 0:
        pytest.skip(f"cuCtxCreate failed with {rc} (exclusive-process compute mode?)")
    bufs = []
    try:
        def dalloc(nbytes):
            d = C.c_uint64()
            assert cu.cuMemAlloc_v2(C.byref(d), max(nbytes, 256)) == 0
            bufs.append(d)
            return d

        d_in = []
        for t in host_in:
            d = dalloc(t.numel() * t.element_size())
            assert cu.cuMemcpyHtoD_v2(d, t.data_ptr(), t.numel() * t.element_size()) == 0
            d_in.append(d)
        d_out = [dalloc(t.numel() * t.element_size()) for t in host_want]
        dq, dk, dx, dl = (C.c_void_p(d.value) for d in d_in)
        rc = lib.dmlp_knn_local_static(dq, dk, q, dx, dl, n, a, id_base, kmax,
                                       *(C.c_void_p(d.value) for d in d_out), None)
        assert rc == 0, lib.dmlp_last_error().decode()
        assert cu.cuCtxSynchronize() == 0
        for name, w, d in zip(("dist", "id", "label", "flags", "num_flagged"), host_want, d_out):
            got = torch.empty_like(w)
            assert cu.cuMemcpyDtoH_v2(got.data_ptr(), d, w.numel() * w.element_size()) == 0
            assert torch.equal(w, got), f"{name} differs between the primary and the second context"
    finally:
        for d in bufs:
            cu.cuMemFree_v2(d)
        old = _vp()
        cu.cuCtxPopCurrent_v2(C.byref(old))
        cu.cuCtxDestroy_v2(ctx)
    # the primary context still works afterwards
    again = _outputs(q, kmax)
    rc = lib.dmlp_knn_local_static(p(qq), p(k), q, p(x), p(labels), n, a, id_base, kmax, *(p(t) for t in again), None)
    assert rc == 0, lib.dmlp_last_error().decode()
    torch.cuda.synchronize()
    assert torch.equal(again[0], want[0]) and torch.equal(again[1], want[1])


def test_static_scratch_entry_refuses_oversized_problem():
    torch.cuda.set_device(0)
    lib = _abi()
    n, q, a = 200000, 4096, 1024     # ~420 MB of packed rows: far beyond 64 MiB
    sms = torch.cuda.get_device_properties(0).multi_processor_count
    assert lib.dmlp_knn_local_workspace_bytes(q, n, a, sms, 4) > STATIC_BYTES
    dummy = torch.zeros(16, dtype=torch.float64, device="cuda")
    p = C.c_void_p(dummy.data_ptr())
    rc = lib.dmlp_knn_local_static(p, p, q, p, p, n, a, 0, 4, p, p, p, p, p, None)
    assert rc != 0
    msg = lib.dmlp_last_error().decode()
    assert "built-in scratch" in msg and str(STATIC_BYTES) in msg
