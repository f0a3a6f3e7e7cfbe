"""Distributed k-NN engine: one process per GPU (torch.distributed, NCCL for plumbing).

Layout (the reference's grid with C = 1, engine.cpp:40-57): rank r holds data shard
r and *owns* query slice r.  Every rank scores all queries against its shard, then
the per-query k-lists are reduced onto the slice owner (a reduce-scatter of
k-lists over the data axis, reference: engine.cpp:283-308), which votes and
checksums.

impl = "fused"        hand-written sm_100a kernels.  Cross-GPU data moves through
                      peer-mapped symmetric buffers from inside the kernels: the
                      filter GEMM all-gathers its query operand while it runs (K4),
                      the re-rank stores its k-lists straight into the owner's
                      receive buffer (K5), flags replace barriers.  NCCL carries
                      one 16-byte all-reduce (value range) per step.
impl = "nccl_cublas"  baseline variant: NCCL all-gather of operands, cuBLAS
                      GEMM (torch.matmul), torch.topk, NCCL exchange of lists.
"""
import numpy as np
import torch
import torch.distributed as dist

from ..ops import cuda as K
from .grid import even_slice
from .symm import Layout, SymmetricBuffer


def _round_up(v, m):
    return (v + m - 1) // m * m


class DistributedKNN:
    def __init__(self, device=None, impl="fused", group=None):
        self.device = torch.device(device if device is not None else "cuda")
        self.impl = impl
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.launch_count = 0
        self.last_flagged = 0
        self.epoch = 0
        self.host = None
        self.sym = None

    # ------------------------------------------------------------------ problem
    def set_local_problem(self, x, labels, id_base, q_slice, k_slice, q_total, kmax):
        """Install this rank's resident inputs: its data shard (fp64 [n, a]) with the
        global id of row 0, and the query slice it owns (rows of the even partition)."""
        dev = self.device
        self.x = x.to(dev, torch.float64).contiguous()
        self.labels = labels.to(dev, torch.int32).contiguous()
        self.id_base = int(id_base)
        self.q_total = int(q_total)
        self.qb, self.qe, self.q_even = even_slice(self.q_total, self.world, self.rank)
        self.nq = self.qe - self.qb
        assert q_slice.shape[0] == self.nq
        self.a = int(self.x.shape[1])
        self.kmax = int(kmax)
        self._alloc(q_slice.to(dev, torch.float64), k_slice.to(dev, torch.int32))

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
            self.set_local_problem(x, labels, nb, qs, ks, q_total, max(1, kmax_eff))

    def _alloc(self, q_slice, k_slice):
        dev, W = self.device, self.world
        n, a, km = self.x.shape[0], self.a, self.kmax
        self.a_pad = K.pad64(a)
        self.fused_ok = km <= K.FUSED_MAX_K and n > 0
        self.peer_fused = W > 1 and self.impl == "fused" and self.fused_ok
        with torch.cuda.device(dev):
            self.stats = K.new_stats(dev)
            self.q_stats = K.new_stats(dev)
            self.xh = torch.empty((n, self.a_pad), dtype=torch.float16, device=dev)
            self.xn = torch.empty(n, dtype=torch.float32, device=dev)
            self.xe = torch.empty(n, dtype=torch.float32, device=dev)
            if self.peer_fused:
                self._alloc_symmetric(q_slice, k_slice)
                return
            # single GPU, or the NCCL-plumbed paths: every slice padded to q_even rows
            self.qs = self.q_even
            self.qp = self.qs * W
            self.q_slice, self.k_slice = q_slice.contiguous(), k_slice.contiguous()
            if W > 1:
                self.q_all = torch.zeros((self.qp, a), dtype=torch.float64, device=dev)
                self.k_all = torch.zeros(self.qp, dtype=torch.int32, device=dev)
                self.q_pad = torch.zeros((self.qs, a), dtype=torch.float64, device=dev)
                self.k_pad = torch.zeros(self.qs, dtype=torch.int32, device=dev)
                self.recv = (torch.empty((W, self.qs, km), dtype=torch.float64, device=dev),
                             torch.empty((W, self.qs, km), dtype=torch.int32, device=dev),
                             torch.empty((W, self.qs, km), dtype=torch.int32, device=dev))
            self.qh = torch.zeros((self.qp, self.a_pad), dtype=torch.float16, device=dev)
            self.qn = torch.zeros(self.qp, dtype=torch.float32, device=dev)
            self.qe_ = torch.zeros(self.qp, dtype=torch.float32, device=dev)
            plan = K.topk_plan(self.qp, max(n, 1), self.a_pad)
            self.cand = torch.empty((self.qp, plan["chunks"], plan["kp"]), dtype=torch.int64, device=dev)
            self.lists = (torch.empty((self.qp, km), dtype=torch.float64, device=dev),
                          torch.empty((self.qp, km), dtype=torch.int32, device=dev),
                          torch.empty((self.qp, km), dtype=torch.int32, device=dev),
                          torch.empty(self.qp, dtype=torch.int32, device=dev),
                          torch.zeros(1, dtype=torch.int32, device=dev))

    def _alloc_symmetric(self, q_slice, k_slice):
        """Symmetric buffer: packed and fp64 queries of every slice, norms, k, the
        receive area for k-lists, and the flag words."""
        dev, W, a, km = self.device, self.world, self.a, self.kmax
        self.qs = _round_up(self.q_even, 256)   # slice stride in the padded query index space
        self.qp = self.qs * W
        lay = Layout()
        lay.add("pack_flags", 128)
        lay.add("list_flags", 128)
        lay.add("qh", self.qp * self.a_pad * 2)
        lay.add("q64", self.qp * a * 8)
        lay.add("qn", self.qp * 4)
        lay.add("qe", self.qp * 4)
        lay.add("k", self.qp * 4)
        lay.add("recv_dist", W * self.qs * km * 8)
        lay.add("recv_id", W * self.qs * km * 4)
        lay.add("recv_label", W * self.qs * km * 4)
        self.sym = sym = SymmetricBuffer(lay, dev, self.group)
        r0 = self.rank * self.qs
        self.q_all = sym.view("q64", torch.float64, (self.qp, a))
        self.k_all = sym.view("k", torch.int32)
        self.qh = sym.view("qh", torch.float16, (self.qp, self.a_pad))
        self.qn = sym.view("qn", torch.float32)
        self.qe_ = sym.view("qe", torch.float32)
        # the resident query slice lives inside the symmetric buffer: peers read it in place
        self.q_slice = self.q_all[r0:r0 + self.nq]
        self.k_slice = self.k_all[r0:r0 + self.nq]
        self.q_slice.copy_(q_slice)
        self.k_slice.copy_(k_slice)
        self.recv = (sym.view("recv_dist", torch.float64, (W, self.qs, km)),
                     sym.view("recv_id", torch.int32, (W, self.qs, km)),
                     sym.view("recv_label", torch.int32, (W, self.qs, km)))
        self.pack_table = sym.pointer_table("pack_flags")
        self.list_table = sym.pointer_table("list_flags")
        self.rb_count = torch.zeros(self.qp // 128, dtype=torch.int32, device=dev)
        self.status = torch.zeros(1, dtype=torch.int32, device=dev)
        n = self.x.shape[0]
        plan = K.topk_plan(self.qp, max(n, 1), self.a_pad)
        self.cand = torch.empty((self.qp, plan["chunks"], plan["kp"]), dtype=torch.int64, device=dev)
        self.flags = torch.empty(self.qp, dtype=torch.int32, device=dev)
        self.nflag = torch.zeros(1, dtype=torch.int32, device=dev)
        g = K.GatherArgs()
        g.world, g.rank, g.slice_rows, g.a = W, self.rank, self.qs, a
        for r in range(W):
            g.peer_base[r] = sym.peer_ptrs[r]
        g.off_qh, g.off_q64 = lay.off["qh"], lay.off["q64"]
        g.off_qn, g.off_qe, g.off_k = lay.off["qn"], lay.off["qe"], lay.off["k"]
        g.off_pack_flags = lay.off["pack_flags"]
        g.rb_count, g.status, g.timeout_seconds = self.rb_count.data_ptr(), self.status.data_ptr(), 20.0
        self.gather = g
        s = K.ScatterArgs()
        s.world, s.rank, s.slice_rows = W, self.rank, self.qs
        for r in range(W):
            s.peer_base[r] = sym.peer_ptrs[r]
        s.off_dist, s.off_id, s.off_label = lay.off["recv_dist"], lay.off["recv_id"], lay.off["recv_label"]
        self.scatter = s
        torch.cuda.synchronize(dev)
        dist.barrier(group=self.group)

    # --------------------------------------------------------------------- steps
    def run_resident(self):
        """One full job with inputs resident on the device.  Returns (label, checksum)
        for the query slice this rank owns (device tensors)."""
        with torch.cuda.device(self.device):
            if self.impl == "nccl_cublas":
                from .baseline import run_baseline
                return run_baseline(self)
            if self.peer_fused:
                return self._run_peer_fused()
            return self._run_local_or_nccl()

    def _range_stats(self):
        """Value range over every rank's data and queries -> common centring / scaling."""
        K.stats_reset(self.stats)
        K.stats_reset(self.q_stats)
        K.stats_update(self.x, self.stats)
        K.stats_update(self.q_slice, self.stats)
        self.launch_count += 4
        if self.world > 1:
            dist.all_reduce(self.stats[:2], op=dist.ReduceOp.MAX, group=self.group)

    def _run_peer_fused(self):
        W, r0 = self.world, self.rank * self.qs
        self.epoch += 1
        self._range_stats()
        K.pack(self.x, self.stats, self.stats, out=(self.xh, self.xn, self.xe))
        K.pack(self.q_slice, self.stats, self.q_stats,
               out=(self.qh[r0:r0 + self.nq], self.qn[r0:r0 + self.nq], self.qe_[r0:r0 + self.nq]))
        self.sym.signal(self.pack_table, self.epoch)          # "my packed slice is ready"
        self.gather.epoch = self.epoch
        K.dist_topk_gather(self.xh, self.xn, self.cand, self.gather)   # K4
        self.nflag.zero_()
        K.rerank_scatter(self.cand, self.q_all, self.x, self.labels, self.k_all, self.qn, self.qe_, self.stats,
                         self.stats, self.id_base, self.kmax, self.flags, self.nflag, self.scatter)   # K5
        self.launch_count += 7   # pack x2, dist_topk, rerank x3, signal
        flagged = int(self.nflag.item())
        if flagged:
            self._repair_flagged()
        self.last_flagged = flagged
        self.sym.signal(self.list_table, self.epoch)          # "my k-lists have landed everywhere"
        self.sym.wait("list_flags", self.epoch, self.status)
        label, checksum, _, _ = K.finalize(self.recv[0], self.recv[1], self.recv[2], self.k_all[r0:r0 + self.qs])
        self.launch_count += 3
        return label[:self.nq], checksum[:self.nq]

    def _repair_flagged(self):
        """Queries the filter could not certify: exact kernel, then store their lists
        into the owners' receive buffers through the peer mappings."""
        rows = torch.nonzero(self.flags, as_tuple=False).flatten()
        km = self.kmax
        tmp = (torch.empty((self.qp, km), dtype=torch.float64, device=self.device),
               torch.empty((self.qp, km), dtype=torch.int32, device=self.device),
               torch.empty((self.qp, km), dtype=torch.int32, device=self.device))
        K.exact(self.q_all, self.x, self.labels, self.k_all, km, self.id_base, rows.to(torch.int32), out=tmp)
        self.launch_count += 3
        owners = rows // self.qs
        for o in torch.unique(owners).tolist():
            sel = rows[owners == o]
            ql = sel - o * self.qs
            for name, dtype, src in (("recv_dist", torch.float64, tmp[0]), ("recv_id", torch.int32, tmp[1]),
                                     ("recv_label", torch.int32, tmp[2])):
                dst = self.sym.view(name, dtype, (self.world, self.qs, km), rank=o)
                dst[self.rank, ql] = src[sel]

    # -- single GPU, and multi-GPU with NCCL plumbing (large k, or impl without peer memory)
    def _gather_queries_nccl(self):
        if self.world == 1:
            return self.q_slice, self.k_slice
        self.q_pad[: self.nq].copy_(self.q_slice)
        self.k_pad[: self.nq].copy_(self.k_slice)
        dist.all_gather_into_tensor(self.q_all, self.q_pad, group=self.group)
        dist.all_gather_into_tensor(self.k_all, self.k_pad, group=self.group)
        return self.q_all, self.k_all

    def _run_local_or_nccl(self):
        q_all, k_all = self._gather_queries_nccl()
        self._range_stats()
        od, oi, ol, flags, nflag = self.lists
        if self.fused_ok:
            K.pack(self.x, self.stats, self.stats, out=(self.xh, self.xn, self.xe))
            K.pack(q_all, self.stats, self.q_stats, out=(self.qh, self.qn, self.qe_))
            K.dist_topk(self.qh, self.xh, self.xn, cand=self.cand)
            nflag.zero_()
            K.rerank(self.cand, q_all, self.x, self.labels, k_all, self.qn, self.qe_, self.stats, self.stats,
                     self.id_base, self.kmax, out=self.lists)
            self.launch_count += 6   # pack x2, dist_topk, rerank x3
            flagged = int(nflag.item())
            if flagged:
                rows = torch.nonzero(flags, as_tuple=False).flatten().to(torch.int32)
                K.exact(q_all, self.x, self.labels, k_all, self.kmax, self.id_base, rows, out=(od, oi, ol))
                self.launch_count += 3
            self.last_flagged = flagged
        else:
            K.exact(q_all, self.x, self.labels, k_all, self.kmax, self.id_base, out=(od, oi, ol))
            self.launch_count += 3
        return self._reduce_and_finalize_nccl(od, oi, ol)

    def _reduce_and_finalize_nccl(self, od, oi, ol):
        """k-lists of every data shard -> owner of the query slice -> merge, vote, checksum."""
        W, qs, km = self.world, self.qs, self.kmax
        if W == 1:
            label, checksum, _, _ = K.finalize(od[None], oi[None], ol[None], self.k_slice)
            self.launch_count += 1
            return label, checksum
        for src, dst in zip((od, oi, ol), self.recv):
            dist.all_to_all_single(dst.view(W * qs, km), src.view(W * qs, km), group=self.group)
        mine = self.k_all[self.rank * qs:(self.rank + 1) * qs]
        if W * km <= 1024:
            label, checksum, _, _ = K.finalize(self.recv[0], self.recv[1], self.recv[2], mine)
            self.launch_count += 1
        else:
            from ..models.knn import _finalize_host_merge
            label, checksum = _finalize_host_merge(self.recv, mine, self.device)
        return label[:self.nq], checksum[:self.nq]

    # ---------------------------------------------------------------- end to end
    def stage_host_inputs(self):
        """Copy this rank's inputs into pinned host memory (the loader's output)."""
        self.host = {
            "x": self.x.cpu().pin_memory(), "labels": self.labels.cpu().pin_memory(),
            "q": self.q_slice.cpu().pin_memory(), "k": self.k_slice.cpu().pin_memory(),
        }
        self.host_out = (torch.empty(self.nq, dtype=torch.int32).pin_memory(),
                         torch.empty(self.nq, dtype=torch.int64).pin_memory())
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

    def check_watchdog(self):
        """Non-zero means a peer signal timed out somewhere."""
        return int(self.status.item()) if self.sym is not None else 0

    def close(self):
        if self.sym is not None:
            self.sym.close()
            self.sym = None

    # -------------------------------------------------------------- verification
    def verify_against_oracle(self, count):
        """Re-derive the first `count` owned queries on the CPU oracle (needs the
        whole dataset on the host: only sensible for small problems)."""
        from ..ops import cpu
        label, checksum = self.run_resident()
        if self.world > 1:
            parts = [None] * self.world     # shards may differ in size: exchange as objects
            dist.all_gather_object(parts, (self.x.cpu().numpy(), self.labels.cpu().numpy()), group=self.group)
            x = torch.from_numpy(np.concatenate([p[0] for p in parts]))
            lab = torch.from_numpy(np.concatenate([p[1] for p in parts]))
        else:
            x, lab = self.x, self.labels
        c = min(count, self.nq)
        want_label, want, *_ = cpu.knn(x.cpu().numpy(), lab.cpu().numpy(), self.q_slice[:c].cpu().numpy(),
                                       self.k_slice[:c].cpu().numpy())
        got = checksum[:c].cpu().numpy().view(np.uint64)
        if not (got == want).all():
            raise AssertionError(f"rank {self.rank}: {(got != want).sum()} of {c} checksums differ from the oracle")
        if not (label[:c].cpu().numpy() == want_label).all():
            raise AssertionError(f"rank {self.rank}: labels differ from the oracle")
        return c
