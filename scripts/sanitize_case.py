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
