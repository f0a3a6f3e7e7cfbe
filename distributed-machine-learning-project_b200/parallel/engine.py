"""Distributed k-NN engine: one process per GPU (torch.distributed, NCCL for plumbing).

Layout (the reference's grid with C = 1, engine.cpp:40-57): rank r holds data shard
r and *owns* query slice r.  Every rank scores all queries against its shard, then
the per-query k-lists are reduced onto the slice owner (a reduce-scatter of
k-lists over the data axis, reference: engine.cpp:283-308), which votes and
checksums.

impl = "fused"        hand-written sm_100a kernels; cross-GPU steps move data
                      through peer-mapped symmetric buffers from inside the
                      kernels (see symm.py); NCCL only for tiny control values.
impl = "nccl_cublas"  baseline variant: NCCL all-gather of operands, cuBLAS
                      GEMM (torch.matmul), torch.topk, NCCL exchange of lists.
"""
import numpy as np
import torch
import torch.distributed as dist

from ..ops import cuda as K
from .grid import even_slice


class DistributedKNN:
    def __init__(self, device=None, impl="fused", group=None):
        self.device = torch.device(device if device is not None else "cuda")
        self.impl = impl
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.launch_count = 0
        self.last_flagged = 0
        self.host = None
        self._symm = None

    # ------------------------------------------------------------------ problem
    def set_local_problem(self, x, labels, id_base, q_slice, k_slice, q_total, kmax):
        """Install this rank's resident inputs: its data shard (fp64 [n, a]) with the
        global id of row 0, and the query slice it owns (rows of the even partition)."""
        dev = self.device
        self.x = x.to(dev, torch.float64).contiguous()
        self.labels = labels.to(dev, torch.int32).contiguous()
        self.id_base = int(id_base)
        self.q_total = int(q_total)
        self.qb, self.qe, self.qs = even_slice(self.q_total, self.world, self.rank)
        self.q_slice = q_slice.to(dev, torch.float64).contiguous()
        self.k_slice = k_slice.to(dev, torch.int32).contiguous()
        assert self.q_slice.shape[0] == self.qe - self.qb
        self.a = int(self.x.shape[1])
        self.kmax = int(kmax)
        self._alloc()

    def synth_problem(self, n_total, q_total, a, lo, hi, min_k, max_k, num_labels, seed):
        """Generate this rank's part of the synthetic problem on the device (K8)."""
        with torch.cuda.device(self.device):
            nb, ne, _ = even_slice(n_total, self.world, self.rank)
            qb, qe, _ = even_slice(q_total, self.world, self.rank)
            x = K.synth_attrs(ne - nb, a, nb, lo, hi, seed, 0, self.device)
            labels = K.synth_labels(ne - nb, nb, num_labels, seed, self.device)
            qs = K.synth_attrs(qe - qb, a, qb, lo, hi, seed, 2, self.device)
            kmax_eff = min(max_k, n_total)
            ks = K.synth_k(qe - qb, qb, min(min_k, kmax_eff), kmax_eff, seed, self.device)
            self.synth_spec = dict(n=n_total, q=q_total, a=a, lo=lo, hi=hi, min_k=min_k, max_k=max_k,
                                   num_labels=num_labels, seed=seed)
            self.set_local_problem(x, labels, nb, qs, ks, q_total, max(1, kmax_eff))

    def _alloc(self):
        dev, W = self.device, self.world
        n, a, Q, qs = self.x.shape[0], self.a, self.q_total, self.qs
        self.a_pad = K.pad64(a)
        self.q_pad_total = qs * W  # padded so every slice has qs rows
        with torch.cuda.device(dev):
            self.stats = K.new_stats(dev)
            self.q_stats = K.new_stats(dev)
            self.fused_ok = self.kmax <= K.FUSED_MAX_K and n > 0
            self.xh = torch.empty((n, self.a_pad), dtype=torch.float16, device=dev)
            self.xn = torch.empty(n, dtype=torch.float32, device=dev)
            self.xe = torch.empty(n, dtype=torch.float32, device=dev)
            # all queries (gathered): fp64 rows, k, packed rows + norms
            if W > 1:
                self.q_all = torch.zeros((self.q_pad_total, a), dtype=torch.float64, device=dev)
                self.k_all = torch.zeros(self.q_pad_total, dtype=torch.int32, device=dev)
            self.qh = torch.zeros((self.q_pad_total, self.a_pad), dtype=torch.float16, device=dev)
            self.qn = torch.zeros(self.q_pad_total, dtype=torch.float32, device=dev)
            self.qe_ = torch.zeros(self.q_pad_total, dtype=torch.float32, device=dev)
            plan = K.topk_plan(self.q_pad_total, max(n, 1), self.a_pad)
            self.cand = torch.empty((self.q_pad_total, plan["chunks"], plan["kp"]), dtype=torch.int64, device=dev)
            km = self.kmax
            self.lists = (torch.empty((self.q_pad_total, km), dtype=torch.float64, device=dev),
                          torch.empty((self.q_pad_total, km), dtype=torch.int32, device=dev),
                          torch.empty((self.q_pad_total, km), dtype=torch.int32, device=dev),
                          torch.empty(self.q_pad_total, dtype=torch.int32, device=dev),
                          torch.zeros(1, dtype=torch.int32, device=dev))
            if W > 1:
                self.recv = (torch.empty((W, qs, km), dtype=torch.float64, device=dev),
                             torch.empty((W, qs, km), dtype=torch.int32, device=dev),
                             torch.empty((W, qs, km), dtype=torch.int32, device=dev))

    # --------------------------------------------------------------------- steps
    def _gather_queries(self):
        """All ranks end up with every query slice (fp64 rows and k)."""
        if self.world == 1:
            return self.q_slice, self.k_slice
        own = self.q_all[self.rank * self.qs: self.rank * self.qs + self.q_slice.shape[0]]
        own.copy_(self.q_slice)
        self.k_all[self.rank * self.qs: self.rank * self.qs + self.k_slice.shape[0]].copy_(self.k_slice)
        dist.all_gather_into_tensor(self.q_all, self.q_all[self.rank * self.qs:(self.rank + 1) * self.qs].clone(),
                                    group=self.group)
        dist.all_gather_into_tensor(self.k_all, self.k_all[self.rank * self.qs:(self.rank + 1) * self.qs].clone(),
                                    group=self.group)
        return self.q_all, self.k_all

    def run_resident(self):
        """One full job with inputs resident on the device.  Returns (label, checksum)
        for the query slice this rank owns (device tensors)."""
        with torch.cuda.device(self.device):
            if self.impl == "nccl_cublas":
                from .baseline import run_baseline
                return run_baseline(self)
            return self._run_fused()

    def _run_fused(self):
        W = self.world
        q_all, k_all = self._gather_queries()
        # value range over every rank's data and queries -> common centring / scaling
        K.stats_reset(self.stats)
        K.stats_reset(self.q_stats)
        K.stats_update(self.x, self.stats)
        K.stats_update(self.q_slice, self.stats)
        self.launch_count += 4
        if W > 1:
            dist.all_reduce(self.stats[:2], op=dist.ReduceOp.MAX, group=self.group)
        if not self.fused_ok:
            return self._run_exact(q_all, k_all)
        K.pack(self.x, self.stats, self.stats, out=(self.xh, self.xn, self.xe))
        K.pack(q_all, self.stats, self.q_stats, out=(self.qh, self.qn, self.qe_))
        K.dist_topk(self.qh, self.xh, self.xn, cand=self.cand)
        self.lists[4].zero_()
        od, oi, ol, flags, nflag = K.rerank(self.cand, q_all, self.x, self.labels, k_all, self.qn, self.qe_,
                                            self.stats, self.stats, self.id_base, self.kmax, out=self.lists)
        self.launch_count += 4
        flagged = int(nflag.item())
        if flagged:
            rows = torch.nonzero(flags[: self.q_pad_total], as_tuple=False).flatten().to(torch.int32)
            K.exact(q_all, self.x, self.labels, k_all, self.kmax, self.id_base, rows, out=(od, oi, ol))
            self.launch_count += 3
        self.last_flagged = flagged
        return self._reduce_and_finalize(od, oi, ol)

    def _run_exact(self, q_all, k_all):
        od, oi, ol = K.exact(q_all, self.x, self.labels, k_all, self.kmax, self.id_base,
                             out=(self.lists[0], self.lists[1], self.lists[2]))
        self.launch_count += 3
        return self._reduce_and_finalize(od, oi, ol)

    def _reduce_and_finalize(self, od, oi, ol):
        """k-lists of every data shard -> owner of the query slice -> merge, vote, checksum."""
        W, qs, km = self.world, self.qs, self.kmax
        if W == 1:
            label, checksum, _, _ = K.finalize(od[None], oi[None], ol[None], self.k_slice)
            self.launch_count += 1
            return label, checksum
        for src, dst in zip((od, oi, ol), self.recv):
            dist.all_to_all_single(dst.view(W * qs, km), src.view(W * qs, km), group=self.group)
        mine = self.k_all[self.rank * qs:(self.rank + 1) * qs]
        label, checksum, _, _ = K.finalize(self.recv[0], self.recv[1], self.recv[2], mine)
        self.launch_count += 1
        nq = self.qe - self.qb
        return label[:nq], checksum[:nq]

    # ---------------------------------------------------------------- end to end
    def stage_host_inputs(self):
        """Copy this rank's inputs into pinned host memory (the loader's output)."""
        self.host = {
            "x": self.x.cpu().pin_memory(), "labels": self.labels.cpu().pin_memory(),
            "q": self.q_slice.cpu().pin_memory(), "k": self.k_slice.cpu().pin_memory(),
        }
        nq = self.qe - self.qb
        self.host_out = (torch.empty(nq, dtype=torch.int32).pin_memory(),
                         torch.empty(nq, dtype=torch.int64).pin_memory())
        self.h2d_bytes_per_step = sum(t.numel() * t.element_size() for t in self.host.values())
        self.d2h_bytes_per_step = sum(t.numel() * t.element_size() for t in self.host_out)

    def run_from_host(self):
        """Public end-to-end call: pinned host inputs -> device -> job -> host results."""
        h = self.host
        self.x.copy_(h["x"], non_blocking=True)
        self.labels.copy_(h["labels"], non_blocking=True)
        self.q_slice.copy_(h["q"], non_blocking=True)
        self.k_slice.copy_(h["k"], non_blocking=True)
        label, checksum = self.run_resident()
        self.host_out[0].copy_(label, non_blocking=True)
        self.host_out[1].copy_(checksum, non_blocking=True)
        torch.cuda.current_stream().synchronize()
        return self.host_out

    # -------------------------------------------------------------- verification
    def verify_against_oracle(self, count):
        """Re-derive the first `count` owned queries on the CPU oracle (needs the
        whole dataset on the host: only sensible for small problems)."""
        from ..ops import cpu
        label, checksum = self.run_resident()
        if self.world > 1:
            xs = [torch.empty_like(self.x) for _ in range(self.world)]
            ls = [torch.empty_like(self.labels) for _ in range(self.world)]
            dist.all_gather(xs, self.x, group=self.group)
            dist.all_gather(ls, self.labels, group=self.group)
            x, lab = torch.cat(xs), torch.cat(ls)
        else:
            x, lab = self.x, self.labels
        c = min(count, self.qe - self.qb)
        _, want, *_ = cpu.knn(x.cpu().numpy(), lab.cpu().numpy(), self.q_slice[:c].cpu().numpy(),
                              self.k_slice[:c].cpu().numpy())
        got = checksum[:c].cpu().numpy().view(np.uint64)
        if not (got == want).all():
            raise AssertionError(f"rank {self.rank}: {(got != want).sum()} of {c} checksums differ from the oracle")
        return c
