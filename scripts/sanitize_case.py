#!/usr/bin/env python3
"""Small end-to-end case that touches every kernel, for `make sanitize-gpu` (compute-sanitizer)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from dmlp_b200.models import knn  # noqa: E402
from dmlp_b200.ops import cpu  # noqa: E402
from dmlp_b200.ops import cuda as K  # noqa: E402

data, labels, queries, k = cpu.synth(1500, 300, 70, -100.0, 100.0, 1, 12, 5, seed=2)
_, want, *_ = cpu.knn(data, labels, queries, k)
for mode in ("fused", "exact"):
    got = knn(data, labels, queries, k, mode=mode, want_lists=True).checksums_u64()
    assert (got == want).all(), mode
x = K.synth_attrs(64, 9, 0, 0.0, 1.0, 1, 0)
enc, ok = K.encode_decimal6(x)
assert ok and torch.equal(K.expand_decimal6(enc, torch.empty_like(x)), x)
torch.cuda.synchronize()
print("sanitize case ok")

# the engine's resident step: in-kernel vote, device-side repair (forced on three rows), streamed end-to-end call
from dmlp_b200.parallel import DistributedKNN  # noqa: E402

eng = DistributedKNN(device=torch.device("cuda", 0))
eng.synth_problem(n_total=3000, q_total=600, a=70, lo=-100.0, hi=100.0, min_k=1, max_k=12, num_labels=5, seed=3)
lab, chk = eng.run_resident()
ref = chk.clone()
assert eng.ensure_repaired() == 0
od, oi, ol, flags, nflag = eng.lists
rows = torch.tensor([0, 5, 7] + [0] * (K.REPAIR_CAP - 3), dtype=torch.int32, device="cuda")
nflag.fill_(3)
oi[[0, 5, 7]] = -9
K.repair_flagged(eng.q_slice, eng.x, eng.labels, eng.k_slice, rows, nflag, eng.id_base, eng.kmax, out=(od, oi, ol),
                 final=eng.final)
torch.cuda.synchronize()
assert torch.equal(eng.final[1][:eng.nq], ref) and int((oi[[0, 5, 7], 0] >= 0).sum()) == 3
eng.stage_host_inputs()
assert (eng.run_from_host()[1].numpy() == ref.cpu().numpy()).all()
assert eng.verify_sampled(n_exact=8, n_cpu=2) == 10
# long rows: the wide tile (lock step, and the opt-in skewed schedule)
d2, l2, q2, k2 = cpu.synth(700, 300, 4096, -100.0, 100.0, 1, 6, 3, seed=5)
_, want2, *_ = cpu.knn(d2, l2, q2, k2)
for gap in ("-1", "6"):
    os.environ["DMLP_TOPK_GAP"] = gap
    assert (knn(d2, l2, q2, k2, mode="fused").checksums_u64() == want2).all(), gap
del os.environ["DMLP_TOPK_GAP"]
# many neighbours: streaming exact kernel with list compaction, rank merge
d3, l3, q3, k3 = cpu.synth(5000, 20, 9, -5.0, 5.0, 200, 300, 3, seed=7)
_, want3, *_ = cpu.knn(d3, l3, q3, k3)
assert (knn(d3, l3, q3, k3).checksums_u64() == want3).all()
torch.cuda.synchronize()
print("sanitize case 2 ok")
