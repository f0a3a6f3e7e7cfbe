"""NCCL + cuBLAS baseline variant (SURVEY.md §5.9 last row, §7.2 step 4).

Same job, same exact answer, but the hot operations go through libraries: NCCL
all-gathers the query operand, cuBLAS (torch.matmul) forms the score matrix in
row blocks, torch.topk selects candidates, NCCL all-to-all moves the k-lists.
The fp16 packing, the exact fp64 re-rank and the vote/checksum are shared with
the fused engine (they are what makes the answer bit-exact).  This is the number
the fused path has to beat, not the product.
"""
import torch

from ..ops import cuda as K

_QB = 4096      # query rows per cuBLAS call (score block = _QB x n fp32)
_HALVES = 2     # candidate lists per query (column ranges), 32 keys each, at least


def _keys_from_topk(vals, idx, base):
    """(score, column) -> the engine's 64-bit ordered keys (int64 bit pattern)."""
    bits = vals.contiguous().view(torch.int32).to(torch.int64) & 0xFFFFFFFF
    neg = bits >> 31
    ordered = torch.where(neg == 1, (~bits) & 0xFFFFFFFF, bits | 0x80000000)
    return (ordered << 32) | (idx.to(torch.int64) + base)


def run_baseline(eng):
    q_all, k_all = eng._gather_queries_nccl()                       # NCCL all-gather of operands
    K.pack(eng.x, eng.stats, eng.stats, out=(eng.xh, eng.xn, eng.xe))
    K.pack(q_all, eng.stats, eng.q_stats, out=(eng.qh, eng.qn, eng.qe_))
    eng.launch_count += 2
    n = eng.x.shape[0]
    qp = eng.qp
    halves = max(_HALVES, -(-eng.kmax // 8))      # same rule as the fused planner: one list per 8 of k
    halves = max(1, min(halves, n // 32)) if n >= 64 else 1
    if not hasattr(eng, "_bl_cand") or eng._bl_cand.shape[1] != halves:
        eng._bl_cand = torch.empty((qp, halves, 32), dtype=torch.int64, device=eng.device)
    cand = eng._bl_cand
    bounds = [n * h // halves for h in range(halves + 1)]
    xh_t = eng.xh.t()
    for b in range(0, qp, _QB):
        e = min(b + _QB, qp)
        dot = _mm_f32(eng.qh[b:e], xh_t)                                # cuBLAS GEMM
        scores = torch.addcmul(eng.xn[None, :], dot, dot.new_full((1,), -2.0))
        for h in range(halves):
            lo, hi = bounds[h], bounds[h + 1]
            kk = min(32, hi - lo)
            vals, idx = torch.topk(scores[:, lo:hi], kk, dim=1, largest=False, sorted=True)   # at::topk
            cand[b:e, h, :kk] = _keys_from_topk(vals, idx, lo)
            if kk < 32:
                cand[b:e, h, kk:] = -1
    od, oi, ol, flags, nflag = eng.lists
    nflag.zero_()
    K.rerank(cand, q_all, eng.x, eng.labels, k_all, eng.qn, eng.qe_, eng.stats, eng.stats, eng.id_base, eng.kmax,
             out=eng.lists)
    eng.launch_count += 3
    flagged = int(nflag.item())
    if flagged:
        rows = torch.nonzero(flags, as_tuple=False).flatten().to(torch.int32)
        K.exact(q_all, eng.x, eng.labels, k_all, eng.kmax, eng.id_base, rows, out=(od, oi, ol))
        eng.launch_count += 3
    eng.last_flagged = flagged
    return eng._reduce_and_finalize_nccl(od, oi, ol)                 # NCCL all-to-all + merge kernel


def _mm_f32(a, b_t):
    """fp16 x fp16 -> fp32 through cuBLAS without rounding the result to fp16."""
    try:
        return torch.mm(a, b_t, out_dtype=torch.float32)
    except TypeError:
        return torch.mm(a.float(), b_t.float())
