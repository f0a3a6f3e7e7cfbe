"""The k-NN classifier: the engine's one "model" (reference: Engine::KNN, engine.h:6-12).

``KNNClassifier`` runs the whole reference job on one B200: host (pinned)
arrays -> async H2D -> pack -> tcgen05 filter GEMM with fused top-k -> exact
fp64 re-rank -> vote + checksum -> D2H.  Queries the filter cannot certify, and
problems outside the fused path's envelope (k > 1024, k above ~3 % of the data, tiny
shapes: ``K.fused_supported``), run on the direct fp64 kernel instead; the answer is
bit-identical either way.
"""
from dataclasses import dataclass

import numpy as np
import torch

from ..ops import cuda as K


@dataclass
class KNNResult:
    label: torch.Tensor      # [q] int32 predicted label
    checksum: torch.Tensor   # [q] int64 holding the uint64 bit pattern
    ids: torch.Tensor        # [q, kmax] int32 neighbour ids (-1 padded) or None
    dist: torch.Tensor       # [q, kmax] float64 distances or None
    flagged: int = 0         # queries that took the exact fallback
    path: str = ""

    def checksums_u64(self):
        return self.checksum.cpu().numpy().view(np.uint64)


def _as_device(x, dtype, device, non_blocking=True):
    if isinstance(x, np.ndarray):
        x = torch.from_numpy(np.ascontiguousarray(x))
    if x.dtype != dtype:
        x = x.to(dtype)
    if x.device != device:
        x = x.to(device, non_blocking=non_blocking)
    return x.contiguous()


def choose_path(n, q, a, kmax, mode="auto"):
    """fused | exact.  The fused path needs kmax within the epilogue's list capacity."""
    if mode in ("fused", "exact"):
        if mode == "fused" and not K.fused_supported(kmax, max(n, 1)):
            raise ValueError(f"fused path supports k <= {K.FUSED_MAX_K_GATHER}, or k <= {K.FUSED_MAX_K} with "
                             f"n >= about 32 k; got k = {kmax}, n = {n}")
        return mode
    if n < 1 or q < 1 or not K.fused_supported(kmax, n):
        return "exact"
    # tiny jobs: one tensor-core tile is mostly padding, go direct
    if n * q * max(a, 1) < (1 << 22):
        return "exact"
    return "fused"


class KNNClassifier:
    def __init__(self, device=None, mode="auto", want_lists=False):
        self.device = torch.device(device if device is not None else "cuda")
        self.mode = mode
        self.want_lists = want_lists
        self.data = self.labels = None

    def fit(self, data, labels):
        """Make the labelled points resident on the device (fp64, row-major)."""
        self.data = _as_device(data, torch.float64, self.device)
        self.labels = _as_device(labels, torch.int32, self.device)
        return self

    def update(self, ids, new_attrs, new_labels=None):
        """Apply update records to the resident points: row ids[i] gets new_attrs[i] (and
        new_labels[i]).  The reference declares this record (`Update{id, new_attrs}`,
        common.h:35-38, with its parser common.cpp:46-55) but never applies one; here it is
        a device-side row replacement, and the next classify() sees the new points."""
        if self.data is None:
            raise RuntimeError("update() before fit()")
        idx = _as_device(np.atleast_1d(np.asarray(ids, dtype=np.int64)), torch.int64, self.device)
        rows = _as_device(new_attrs, torch.float64, self.device).reshape(idx.numel(), -1)
        if rows.shape[1] != self.data.shape[1]:
            raise ValueError(f"update rows have {rows.shape[1]} attributes, the dataset has {self.data.shape[1]}")
        if idx.numel() and (int(idx.min()) < 0 or int(idx.max()) >= self.data.shape[0]):
            raise IndexError("update id out of range")
        self.data.index_copy_(0, idx, rows)
        if new_labels is not None:
            lab = _as_device(np.atleast_1d(np.asarray(new_labels, dtype=np.int32)), torch.int32, self.device)
            self.labels.index_copy_(0, idx, lab)
        return self

    def classify(self, queries, k, kmax=None, id_base=0):
        """Classify `queries` ([q, a]) asking k[i] neighbours each."""
        with torch.cuda.device(self.device):
            qd = _as_device(queries, torch.float64, self.device)
            if kmax is None:
                kh = k if isinstance(k, np.ndarray) else (k.cpu().numpy() if torch.is_tensor(k) else np.asarray(k))
                kmax = int(max(1, kh.max())) if kh.size else 1
            kd = _as_device(k, torch.int32, self.device)
            n, a = self.data.shape
            q = qd.shape[0]
            if q == 0:
                e = torch.empty(0, dtype=torch.int32, device=self.device)
                return KNNResult(e, torch.empty(0, dtype=torch.int64, device=self.device), None, None, 0, "empty")
            path = choose_path(n, q, a, kmax, self.mode)
            flagged = 0
            if path == "fused":
                od, oi, ol, flags, nflag = K.knn_local(qd, kd, self.data, self.labels, kmax, id_base)
                label, checksum, ids, dist = K.finalize(od[None], oi[None], ol[None], kd, self.want_lists)
                flagged = int(nflag.item())  # the only host sync of the call
                if flagged:
                    rows = torch.nonzero(flags, as_tuple=False).flatten().to(torch.int32)
                    K.exact(qd, self.data, self.labels, kd, kmax, id_base, rows, out=(od, oi, ol))
                    label, checksum, ids, dist = K.finalize(od[None], oi[None], ol[None], kd, self.want_lists)
            else:
                od, oi, ol = K.exact(qd, self.data, self.labels, kd, kmax, id_base)
                label, checksum, ids, dist = K.finalize(od[None], oi[None], ol[None], kd, self.want_lists)
            return KNNResult(label, checksum, ids, dist, flagged, path)


def knn(data, labels, queries, k, device=None, mode="auto", want_lists=False):
    """One-shot job, the call the reference driver makes (common.cpp:126)."""
    return KNNClassifier(device, mode, want_lists).fit(data, labels).classify(queries, k)
