"""Distributed k-NN engine: one process per GPU (torch.distributed, NCCL for plumbing).

Grid (the reference's process grid, engine.cpp:40-57): R rows x C columns,
rank = r * C + c.  Row r holds data shard r, column c holds query shard c.

  * along a row (same data shard, C ranks): every rank loads 1/C of the shard and
    the parts are all-gathered over NVLink by a pull kernel (the reference's row
    MPI_Bcast of the data shard, engine.cpp:129-131);
  * along a column (same queries, R ranks): every rank owns 1/R of the column's
    queries; the filter GEMM all-gathers the packed queries from inside the kernel
    (K4) and the re-rank stores each query's k-list straight into its owner's
    receive buffer (K5, the reduce-scatter of k-lists of engine.cpp:283-308).

R x 1 is pure data sharding, 1 x C pure query sharding with a replicated dataset
(the data-parallel layout: every GPU sees the whole dataset, the batch is split).

impl = "fused"        the above, hand-written sm_100a kernels, peer-mapped symmetric
                      buffers, flags instead of barriers.  NCCL carries one 16-byte
                      all-reduce per step, and only when R > 1.
impl = "nccl_cublas"  baseline variant: NCCL all-gathers, cuBLAS GEMM (torch.mm),
                      torch.topk, NCCL all-to-all of the lists.
"""
import ctypes as C

import numpy as np
import torch
import torch.distributed as dist

import os

from ..ops import cuda as K
from ..utils.timing import PhaseTracer
from .grid import Grid, even_slice
from .symm import Layout, SymmetricBuffer


def _round_up(v, m):
    return (v + m - 1) // m * m


def simulate_copy_order(order, q_rows, x_rows, a, bytes_per_value=4, pcie_gbs=52.0, gemm_pflops=1.2, hbm_tbs=5.5,
                        pairs_per_query=16.0):
    """Makespan (ms) of one end-to-end step for a given copy order of query chunks and data parts.

    One copy stream moves the pieces back to back; the compute stream runs, in the order the
    step enqueues them, the expand + pack of an arrived piece, the filter GEMM of every (query
    chunk, data part) pair as soon as both are there, and a chunk's re-rank once its last pair
    is done.  Costs are first-order: bytes over PCIe, 2 q x a flops at the sustained GEMM rate,
    HBM bytes for pack and for the re-rank's row fetches."""
    t_copy = t_gpu = 0.0
    have_q, have_x, done, reranked = [], [], set(), set()
    for kind, idx in order:
        rows = q_rows[idx] if kind == "q" else x_rows[idx]
        t_copy += rows * a * bytes_per_value / (pcie_gbs * 1e6)
        # expand (decimal6 only) + pack: read wire, write + read fp64, write fp16
        t_gpu = max(t_gpu, t_copy) + rows * a * (bytes_per_value + 8 + 8 + 2) / (hbm_tbs * 1e9)
        (have_q if kind == "q" else have_x).append(idx)
        for i in sorted(have_q):
            for p_ in sorted(have_x):
                if (i, p_) not in done:
                    done.add((i, p_))
                    t_gpu += 2.0 * q_rows[i] * x_rows[p_] * a / (gemm_pflops * 1e12)
            if i not in reranked and all((i, p_) in done for p_ in range(len(x_rows))):
                reranked.add(i)
                t_gpu += q_rows[i] * pairs_per_query * a * 8 / (hbm_tbs * 1e9)
    return t_gpu


def best_copy_order(q_rows, x_rows, a, bytes_per_value=4, limit=40000, **model):
    """The interleaving of query chunks and data parts (each kind in index order) with the
    smallest simulated makespan; exhaustive while the number of interleavings stays small."""
    from itertools import combinations
    from math import comb
    nq, nx = len(q_rows), len(x_rows)

    def alternate():
        order, xi, qi = [("q", 0), ("x", 0)], 1, 1
        while xi < nx or qi < nq:
            if xi < nx:
                order.append(("x", xi))
                xi += 1
            if qi < nq:
                order.append(("q", qi))
                qi += 1
        return order
    if nq == 0 or nx == 0 or comb(nq + nx, nx) > limit:
        return alternate()
    best, best_t = None, float("inf")
    for slots in combinations(range(nq + nx), nx):       # positions of the data parts
        taken, order, qi, xi = set(slots), [], 0, 0
        for pos in range(nq + nx):
            if pos in taken:
                order.append(("x", xi))
                xi += 1
            else:
                order.append(("q", qi))
                qi += 1
        t = simulate_copy_order(order, q_rows, x_rows, a, bytes_per_value, **model)
        if t < best_t:
            best, best_t = order, t
    return best


class DistributedKNN:
    def __init__(self, device=None, impl="fused", grid=None):
        self.device = torch.device(device if device is not None else "cuda")
        self.impl = impl
        self.world = dist.get_world_size() if dist.is_initialized() else 1
        self.rank = dist.get_rank() if dist.is_initialized() else 0
        if grid is None:
            grid = (self.world, 1)
        self.grid = Grid.for_world(self.world, *grid)
        self.R, self.C = self.grid.rows, self.grid.cols
        self.r, self.c = self.grid.coords(self.rank)
        self.col_group = self.row_group = None
        if self.world > 1:
            # every rank must create every group, in the same order
            for c in range(self.C):
                g = dist.new_group(self.grid.column_ranks(c))
                if c == self.c:
                    self.col_group = g
            for r in range(self.R):
                g = dist.new_group(self.grid.row_ranks(r))
                if r == self.r:
                    self.row_group = g
        self.launch_count = 0
        self._x32_step = None
        self._flagged_host = None
        self.epoch = 0
        self.host = None
        self.sym = self.row_sym = None
        self.tracer = PhaseTracer(os.environ.get("DMLP_TRACE", "0") not in ("", "0"))

    # ------------------------------------------------------------------ problem
    def partition(self, n_total, q_total):
        """(data part, data shard, query slice, query shard) as global [begin, end) ranges."""
        sb, se, _ = even_slice(n_total, self.R, self.r)          # data shard of this row
        pb, pe, _ = even_slice(se - sb, self.C, self.c)          # the part of it this rank loads
        cb, ce, _ = even_slice(q_total, self.C, self.c)          # query shard of this column
        ob, oe, _ = even_slice(ce - cb, self.R, self.r)          # the slice of it this rank owns
        return (sb + pb, sb + pe), (sb, se), (cb + ob, cb + oe), (cb, ce)

    def set_local_problem(self, x_part, labels_part, q_slice, k_slice, n_total, q_total, kmax):
        """Install this rank's resident inputs: the part of its data shard it loaded and
        the query slice it owns (see partition())."""
        dev = self.device
        (pb, pe), (sb, se), (ob, oe), (cb, ce) = self.partition(n_total, q_total)
        assert x_part.shape[0] == pe - pb and q_slice.shape[0] == oe - ob
        self.n_total, self.q_total = int(n_total), int(q_total)
        self.part_rows, self.shard_rows = (pb - sb, pe - sb), se - sb
        self.id_base = sb
        self.q_col = ce - cb                      # queries of this column
        self.q_even = even_slice(self.q_col, self.R, self.r)[2]
        self.nq = oe - ob
        self.a = int(x_part.shape[1])
        self.kmax = int(kmax)
        self._alloc(x_part.to(dev, torch.float64), labels_part.to(dev, torch.int32),
                    q_slice.to(dev, torch.float64), k_slice.to(dev, torch.int32))

    def synth_problem(self, n_total, q_total, a, lo, hi, min_k, max_k, num_labels, seed):
        """Generate this rank's part of the synthetic problem on the device (K8)."""
        with torch.cuda.device(self.device):
            (pb, pe), _, (ob, oe), _ = self.partition(n_total, q_total)
            x = K.synth_attrs(pe - pb, a, pb, lo, hi, seed, 0, self.device)
            labels = K.synth_labels(pe - pb, pb, num_labels, seed, self.device)
            qs = K.synth_attrs(oe - ob, a, ob, lo, hi, seed, 2, self.device)
            kmax_eff = min(max_k, n_total)
            ks = K.synth_k(oe - ob, ob, min(min_k, kmax_eff), kmax_eff, seed, self.device)
            self.set_local_problem(x, labels, qs, ks, n_total, q_total, max(1, kmax_eff))

    # --------------------------------------------------------------- allocation
    def _alloc(self, x_part, labels_part, q_slice, k_slice):
        dev, R, C = self.device, self.R, self.C
        n, km = self.shard_rows, self.kmax
        self.a_pad = K.pad64(self.a)
        # k <= 96: every layout; 96 < k <= 1024 (and n >= 32 k): the filter runs as several launches over
        # row ranges of the shard, which the in-kernel gathers do not do -- a single-rank column then
        # gathers the shard with the peer_gather kernel first, layouts with R > 1 take the direct path
        self.fused_ok = n > 0 and K.fused_supported(km, n, gather=R > 1)
        self.peer_fused = R > 1 and self.impl == "fused" and self.fused_ok
        self.peer_rows = C > 1 and self.impl == "fused"
        # data operand gathered inside the GEMM (needs the CTA-pair kernel: >= 256 queries per column)
        self.rows_fused = (self.peer_rows and R == 1 and self.fused_ok and self.nq >= 256
                           and K.fused_supported(km, n, gather=True))
        with torch.cuda.device(dev):
            self.status = torch.zeros(1, dtype=torch.int32, device=dev)
            self.stats = K.new_stats(dev)
            self.q_stats = K.new_stats(dev)
            # Resident decimal6 rows: when every data value is i / 1e6 for a 32-bit i (the text grammar's
            # "%.6f"), the hot path reads the data as int32 -- half the bytes of pack and of the exact
            # re-rank, the pass that is bound by exactly those bytes -- and rebuilds the doubles bit for
            # bit in the kernels.  The fp64 rows stay resident for the direct kernels and the API.
            self.dec6, x_enc = False, None
            if (self.impl == "fused" and self.fused_ok and self.a % 2 == 0 and (C == 1 or self.rows_fused)
                    and os.environ.get("DMLP_DEC6", "1") != "0"):
                x_enc, ok = K.encode_decimal6(x_part)
                okt = torch.tensor([1 if ok else 0], dtype=torch.int32, device=dev)
                if self.world > 1:
                    dist.all_reduce(okt, op=dist.ReduceOp.MIN)
                self.dec6 = bool(okt.item())
            self._x_enc = x_enc if self.dec6 else None
            self._alloc_shard(x_part, labels_part)
            self._x_enc = None
            if not self.rows_fused:
                self.xh = torch.empty((n, self.a_pad), dtype=torch.float16, device=dev)
                self.xn = torch.empty(n, dtype=torch.float32, device=dev)
                self.xe = torch.empty(n, dtype=torch.float32, device=dev)
            if self.peer_fused:
                self._alloc_symmetric(q_slice, k_slice)
            else:
                self._alloc_plain(q_slice, k_slice)
            # the repair pass of a step works on at most REPAIR_CAP flagged queries without the host
            self.flagged_rows = torch.zeros(K.REPAIR_CAP, dtype=torch.int32, device=dev)
            self.final = (torch.empty(max(self.qs, 1), dtype=torch.int32, device=dev),
                          torch.empty(max(self.qs, 1), dtype=torch.int64, device=dev))
            self._load_range()
            torch.cuda.synchronize(dev)
        if self.world > 1:
            dist.barrier()

    def _load_range(self):
        """Value range of everything that is resident -> common centring / scaling of the packed
        rows on every rank.  Part of loading (the end-to-end path gets the range from the
        loader the same way), not of a step: a step starts at the pack kernels."""
        K.stats_reset(self.stats)
        K.stats_reset(self.q_stats)
        K.stats_update(self.x_part, self.stats)
        K.stats_update(self.q_slice, self.stats)
        if self.world > 1:
            dist.all_reduce(self.stats[:2], op=dist.ReduceOp.MAX)

    def _alloc_shard(self, x_part, labels_part):
        """The data shard of this row: local when C == 1, otherwise a row-symmetric buffer
        in which every rank of the row fills its own part and pulls the others'."""
        dev, C, n, a = self.device, self.C, self.shard_rows, self.a
        b, e = self.part_rows
        self.x32 = self.x32_part = None
        if C == 1:
            self.x, self.labels = x_part.contiguous(), labels_part.contiguous()
            self.x_part, self.labels_part = self.x, self.labels
            if self.dec6:
                self.x32 = self.x32_part = self._x_enc.contiguous()
            return
        if self.peer_rows:
            lay = Layout()
            lay.add("part_flags", 128)
            lay.add("done_flags", 128)
            lay.add("x64", n * a * 8)
            lay.add("labels", n * 4)
            if self.rows_fused:
                lay.add("pack_flags", 128)
                lay.add("xh", n * self.a_pad * 2)
                lay.add("xn", n * 4)
                lay.add("xe", n * 4)
            if self.dec6:
                lay.add("x32", n * a * 4)
            self.row_sym = sym = SymmetricBuffer(lay, dev, self.row_group)
            self.x = sym.view("x64", torch.float64, (n, a))
            self.labels = sym.view("labels", torch.int32)
            self.part_table = sym.pointer_table("part_flags")
            self.done_table = sym.pointer_table("done_flags")
            if self.rows_fused:
                self.xh = sym.view("xh", torch.float16, (n, self.a_pad))
                self.xn = sym.view("xn", torch.float32)
                self.xe = sym.view("xe", torch.float32)
                self.row_pack_table = sym.pointer_table("pack_flags")
                self.tile_count = torch.zeros((n + 255) // 256, dtype=torch.int32, device=dev)
                dg = K.DataGatherArgs()
                dg.world, dg.rank, dg.a = C, self.c, a
                for p in range(C):
                    dg.peer_base[p] = sym.peer_ptrs[p]
                    dg.part_begin[p] = even_slice(n, C, p)[0]
                dg.part_begin[C] = n
                dg.off_xh, dg.off_xn, dg.off_xe = lay.off["xh"], lay.off["xn"], lay.off["xe"]
                dg.off_x64, dg.off_labels, dg.off_pack_flags = lay.off["x64"], lay.off["labels"], lay.off["pack_flags"]
                if self.dec6:    # the re-rank reads decimal6 rows: those are what the GEMM's communication warps pull
                    dg.off_x64, dg.x_i32 = lay.off["x32"], 1
                dg.tile_count, dg.status, dg.timeout_seconds = self.tile_count.data_ptr(), self.status.data_ptr(), 20.0
                self.data_gather = dg
            self._row_gather_args = []
            for name, row_bytes in (("x64", a * 8), ("labels", 4)):
                g = K.PeerGatherArgs()
                g.world, g.rank = C, self.c
                for p in range(C):
                    g.peer_base[p] = sym.peer_ptrs[p]
                    pb, pe, _ = even_slice(n, C, p)
                    g.part_off[p] = lay.off[name] + pb * row_bytes
                    g.part_len[p] = (pe - pb) * row_bytes
                g.off_flags = lay.off["part_flags"]
                g.status, g.timeout_seconds = self.status.data_ptr(), 20.0
                self._row_gather_args.append(g)
        else:
            self.x = torch.empty((n, a), dtype=torch.float64, device=dev)
            self.labels = torch.empty(n, dtype=torch.int32, device=dev)
            self._row_sizes = [even_slice(n, C, p)[1] - even_slice(n, C, p)[0] for p in range(C)]
        self.x_part, self.labels_part = self.x[b:e], self.labels[b:e]
        self.x_part.copy_(x_part)
        self.labels_part.copy_(labels_part)
        if self.dec6:
            self.x32 = self.row_sym.view("x32", torch.int32, (n, a))
            self.x32_part = self.x32[b:e]
            self.x32_part.copy_(self._x_enc)

    def _alloc_plain(self, q_slice, k_slice):
        """Column buffers for a single-rank column or for the NCCL-plumbed paths."""
        dev, R, a, km, n = self.device, self.R, self.a, self.kmax, self.shard_rows
        self.qs = self.q_even
        self.qp = self.qs * R
        self.q_slice, self.k_slice = q_slice.contiguous(), k_slice.contiguous()
        if R > 1:
            self.q_all = torch.zeros((self.qp, a), dtype=torch.float64, device=dev)
            self.k_all = torch.zeros(self.qp, dtype=torch.int32, device=dev)
            self.q_pad = torch.zeros((self.qs, a), dtype=torch.float64, device=dev)
            self.k_pad = torch.zeros(self.qs, dtype=torch.int32, device=dev)
            self.recv = (torch.empty((R, self.qs, km), dtype=torch.float64, device=dev),
                         torch.empty((R, self.qs, km), dtype=torch.int32, device=dev),
                         torch.empty((R, self.qs, km), dtype=torch.int32, device=dev))
        self.qh = torch.zeros((self.qp, self.a_pad), dtype=torch.float16, device=dev)
        self.qn = torch.zeros(self.qp, dtype=torch.float32, device=dev)
        self.qe_ = torch.zeros(self.qp, dtype=torch.float32, device=dev)
        plan = K.topk_plan(self.qp, max(n, 1), self.a_pad, kmax=km)
        self.cand = torch.empty((self.qp, plan["chunks"], plan["kp"]), dtype=torch.int64, device=dev)
        self.lists = (torch.empty((self.qp, km), dtype=torch.float64, device=dev),
                      torch.empty((self.qp, km), dtype=torch.int32, device=dev),
                      torch.empty((self.qp, km), dtype=torch.int32, device=dev),
                      torch.empty(self.qp, dtype=torch.int32, device=dev),
                      torch.zeros(1, dtype=torch.int32, device=dev))

    def _alloc_symmetric(self, q_slice, k_slice):
        """Column-symmetric buffer: packed and fp64 queries of every slice, norms, k, the
        receive area for k-lists, and the flag words."""
        dev, R, a, km = self.device, self.R, self.a, self.kmax
        self.qs = _round_up(self.q_even, 256)   # slice stride in the padded query index space
        self.qp = self.qs * R
        lay = Layout()
        lay.add("pack_flags", 128)
        lay.add("list_flags", 128)
        lay.add("qh", self.qp * self.a_pad * 2)
        lay.add("q64", self.qp * a * 8)
        lay.add("qn", self.qp * 4)
        lay.add("qe", self.qp * 4)
        lay.add("k", self.qp * 4)
        lay.add("recv_dist", R * self.qs * km * 8)
        lay.add("recv_id", R * self.qs * km * 4)
        lay.add("recv_label", R * self.qs * km * 4)
        self.sym = sym = SymmetricBuffer(lay, dev, self.col_group)
        r0 = self.r * self.qs
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
        self.recv = (sym.view("recv_dist", torch.float64, (R, self.qs, km)),
                     sym.view("recv_id", torch.int32, (R, self.qs, km)),
                     sym.view("recv_label", torch.int32, (R, self.qs, km)))
        self.pack_table = sym.pointer_table("pack_flags")
        self.list_table = sym.pointer_table("list_flags")
        self.rb_count = torch.zeros(self.qp // 128, dtype=torch.int32, device=dev)
        plan = K.topk_plan(self.qp, max(self.shard_rows, 1), self.a_pad, kmax=km)
        self.cand = torch.empty((self.qp, plan["chunks"], plan["kp"]), dtype=torch.int64, device=dev)
        self.flags = torch.empty(self.qp, dtype=torch.int32, device=dev)
        self.nflag = torch.zeros(1, dtype=torch.int32, device=dev)
        g = K.GatherArgs()
        g.world, g.rank, g.slice_rows, g.a = R, self.r, self.qs, a
        for r in range(R):
            g.peer_base[r] = sym.peer_ptrs[r]
        g.off_qh, g.off_q64 = lay.off["qh"], lay.off["q64"]
        g.off_qn, g.off_qe, g.off_k = lay.off["qn"], lay.off["qe"], lay.off["k"]
        g.off_pack_flags = lay.off["pack_flags"]
        g.rb_count, g.status, g.timeout_seconds = self.rb_count.data_ptr(), self.status.data_ptr(), 20.0
        self.gather = g
        s = K.ScatterArgs()
        s.world, s.rank, s.slice_rows = R, self.r, self.qs
        for r in range(R):
            s.peer_base[r] = sym.peer_ptrs[r]
        s.off_dist, s.off_id, s.off_label = lay.off["recv_dist"], lay.off["recv_id"], lay.off["recv_label"]
        self.scatter = s

    # --------------------------------------------------------------------- steps
    def run_resident(self):
        """One full job with inputs resident on the device.  Returns (label, checksum)
        for the query slice this rank owns (device tensors)."""
        with torch.cuda.device(self.device):
            self.epoch += 1
            self._flagged_host = None
            self.tracer.begin_step()
            if self.rows_fused:
                return self._run_rows_fused()
            with self.tracer.phase("row_gather"):
                self._gather_shard()
            if self.impl == "nccl_cublas":
                from .baseline import run_baseline
                return run_baseline(self)
            if self.peer_fused:
                return self._run_peer_fused()
            return self._run_local_or_nccl()

    def _gather_shard(self):
        """Row all-gather: after this, self.x / self.labels hold the whole data shard."""
        if self.C == 1:
            return
        if self.peer_rows:
            self.row_sym.signal(self.part_table, self.epoch)     # "my part is in place"
            for g in self._row_gather_args:
                g.epoch = self.epoch
                K.peer_gather(g)
            self.row_sym.signal(self.done_table, self.epoch)     # "I have finished reading your parts"
            self.launch_count += 4
        elif len(set(self._row_sizes)) == 1:
            dist.all_gather_into_tensor(self.x, self.x_part.contiguous(), group=self.row_group)
            dist.all_gather_into_tensor(self.labels, self.labels_part.contiguous(), group=self.row_group)
        else:
            for full, part in ((self.x, self.x_part), (self.labels, self.labels_part)):
                outs, b = [], 0
                for sz in self._row_sizes:
                    outs.append(full[b:b + sz])
                    b += sz
                dist.all_gather(outs, part.contiguous(), group=self.row_group)

    def _pack_part_and_signal(self):
        """rows_fused: pack this rank's part of the data shard into the row-symmetric buffer; the
        kernel itself first waits until every peer has finished pulling the previous step's copy
        and its last block tells the row that the new rows are ready."""
        b, e = self.part_rows
        sym, ep = self.row_sym, self.epoch
        (K.pack_i32 if self.dec6 else K.pack_ex)(
            self.x32_part if self.dec6 else self.x_part, self.stats, None, out=(self.xh[b:e], self.xn[b:e], self.xe[b:e]),
            wait=sym.wait_args("done_flags", ep - 1, self.status), signal=sym.signal_args("pack_flags", ep))
        self.data_gather.epoch = ep
        self._done_signal = sym.signal_args("done_flags", ep)       # sent by the GEMM's last block
        self.data_gather.done_signal = C.cast(C.pointer(self._done_signal), C.c_void_p)
        self.launch_count += 1

    def _run_rows_fused(self):
        """Grid 1 x C on resident inputs: the row all-gather of the data shard runs inside the
        filter GEMM (packed rows just ahead of the tiles that need them, fp64 rows after).  Seven
        launches, every flag handled inside them, no NCCL call and no host synchronisation."""
        tr = self.tracer
        od, oi, ol, flags, nflag = self.lists
        with tr.phase("pack"):
            self._pack_part_and_signal()
            K.pack(self.q_slice, self.stats, self.q_stats, out=(self.qh, self.qn, self.qe_))
        with tr.phase("dist_topk+row_allgather"):
            K.dist_topk_rows(self.qh, self.shard_rows, self.cand, self.data_gather, kmax=self.kmax)
            K.maxima_from_norms(self.xn, self.xe, self.stats)
        K.zero_(nflag)
        with tr.phase("rerank+vote"):
            K.rerank_ex(self.cand, self.q_slice, self.x, self.labels, self.k_slice, self.qn, self.qe_, self.stats,
                        self.stats, self.id_base, self.kmax, flags, nflag, out=(od, oi, ol), final=self.final,
                        flagged_rows=self.flagged_rows, x_i32=self.x32)
        with tr.phase("repair"):
            K.repair_flagged(self.q_slice, self.x, self.labels, self.k_slice, self.flagged_rows, nflag, self.id_base,
                             self.kmax, out=(od, oi, ol), final=self.final, x_i32=self.x32)
        self.launch_count += 9   # pack, dist_topk, maxima, rerank x3, repair x3
        return self.final[0][:self.nq], self.final[1][:self.nq]

    def _run_peer_fused(self):
        """R > 1: K4 (the GEMM all-gathers the packed queries) + K5 (k-lists stored into the owners'
        receive buffers; the last kernel that writes them signals, the owner's finalize waits)."""
        r0, tr, ep = self.r * self.qs, self.tracer, self.epoch
        with tr.phase("pack"):
            self._pack_shard()
            K.pack_ex(self.q_slice, self.stats, self.q_stats,
                      out=(self.qh[r0:r0 + self.nq], self.qn[r0:r0 + self.nq], self.qe_[r0:r0 + self.nq]),
                      signal=self.sym.signal_args("pack_flags", ep))   # "my packed slice is ready"
        self.gather.epoch = ep
        with tr.phase("dist_topk+allgather"):
            K.dist_topk_gather(self.xh, self.xn, self.cand, self.gather, kmax=self.kmax)   # K4
        K.zero_(self.nflag)
        with tr.phase("rerank+scatter"):
            K.rerank_ex(self.cand, self.q_all, self.x, self.labels, self.k_all, self.qn, self.qe_, self.stats,
                        self.stats, self.id_base, self.kmax, self.flags, self.nflag, scatter=self.scatter,
                        flagged_rows=self.flagged_rows, x_i32=self.x32)   # K5 store side
        with tr.phase("repair+signal"):
            K.repair_flagged(self.q_all, self.x, self.labels, self.k_all, self.flagged_rows, self.nflag,
                             self.id_base, self.kmax, scatter=self.scatter, x_i32=self.x32,
                             signal=self.sym.signal_args("list_flags", ep))   # "my k-lists have landed everywhere"
        with tr.phase("wait_lists+finalize"):
            K.finalize_ex(self.recv[0], self.recv[1], self.recv[2], self.k_all[r0:r0 + self.qs], out=self.final,
                          wait=self.sym.wait_args("list_flags", ep, self.status))
        self.launch_count += 9   # pack x2, dist_topk, rerank x3, repair x2, finalize
        return self.final[0][:self.nq], self.final[1][:self.nq]

    def _pack_shard(self):
        """Pack the whole (local) data shard: from the decimal6 rows when they are resident."""
        if self.dec6:
            K.pack_i32(self.x32, self.stats, self.stats, out=(self.xh, self.xn, self.xe))
        else:
            K.pack(self.x, self.stats, self.stats, out=(self.xh, self.xn, self.xe))

    def _nflag_tensor(self):
        return self.nflag if self.peer_fused else self.lists[4]

    @property
    def last_flagged(self):
        """Queries of the last step the filter could not certify.  The fused steps keep the count
        on the device (this reads it, which synchronises); the library-plumbed paths know it."""
        if self._flagged_host is not None:
            return self._flagged_host
        if not getattr(self, "fused_ok", False):
            return 0
        return int(self._nflag_tensor().sum().item())

    @last_flagged.setter
    def last_flagged(self, value):
        self._flagged_host = value

    def ensure_repaired(self, force=False):
        """End-of-step check: up to REPAIR_CAP flagged queries are repaired on the device by the step
        itself; only beyond that (never seen on the generator's distributions) the host steps in,
        re-runs every flagged query on the direct kernel and finalizes again.  Returns the count."""
        flagged = self.last_flagged
        worst = flagged
        if self.peer_fused:   # the ranks of a column repair together or not at all
            t = torch.tensor([flagged], dtype=torch.int32, device=self.device)
            dist.all_reduce(t, op=dist.ReduceOp.MAX, group=self.col_group)
            worst = int(t.item())
        if worst <= K.REPAIR_CAP and not force:
            return flagged
        if self.peer_fused:
            self._repair_flagged_host()
            torch.cuda.synchronize(self.device)
            dist.barrier(group=self.col_group)
            r0 = self.r * self.qs
            K.finalize_ex(self.recv[0], self.recv[1], self.recv[2], self.k_all[r0:r0 + self.qs], out=self.final)
        else:
            od, oi, ol, flags, _ = self.lists
            rows = torch.nonzero(flags[: self.nq], as_tuple=False).flatten().to(torch.int32)
            K.exact(self.q_slice, self.x, self.labels, self.k_slice, self.kmax, self.id_base, rows, out=(od, oi, ol),
                    x_i32=self.x32)
            K.finalize_ex(od[None], oi[None], ol[None], self.k_slice, out=self.final)
        return flagged

    def _repair_flagged_host(self):
        """More flagged queries than the step repairs by itself: exact kernel over all of them, then
        their lists go into the owners' receive buffers through the peer mappings."""
        rows = torch.nonzero(self.flags, as_tuple=False).flatten()
        km = self.kmax
        tmp = (torch.empty((self.qp, km), dtype=torch.float64, device=self.device),
               torch.empty((self.qp, km), dtype=torch.int32, device=self.device),
               torch.empty((self.qp, km), dtype=torch.int32, device=self.device))
        K.exact(self.q_all, self.x, self.labels, self.k_all, km, self.id_base, rows.to(torch.int32), out=tmp,
                x_i32=self.x32)
        owners = rows // self.qs
        for o in torch.unique(owners).tolist():
            sel = rows[owners == o]
            ql = sel - o * self.qs
            for name, dtype, src in (("recv_dist", torch.float64, tmp[0]), ("recv_id", torch.int32, tmp[1]),
                                     ("recv_label", torch.int32, tmp[2])):
                dst = self.sym.view(name, dtype, (self.R, self.qs, km), rank=o)
                dst[self.r, ql] = src[sel]

    # -- single-rank columns, and columns plumbed with NCCL (large k, or the baseline)
    def _gather_queries_nccl(self):
        if self.R == 1:
            return self.q_slice, self.k_slice
        self.q_pad[: self.nq].copy_(self.q_slice)
        self.k_pad[: self.nq].copy_(self.k_slice)
        dist.all_gather_into_tensor(self.q_all, self.q_pad, group=self.col_group)
        dist.all_gather_into_tensor(self.k_all, self.k_pad, group=self.col_group)
        return self.q_all, self.k_all

    def _run_local_or_nccl(self):
        tr = self.tracer
        od, oi, ol, flags, nflag = self.lists
        if self.fused_ok and self.R == 1:
            # one data shard: pack, filter, re-rank with the vote folded in, device-side repair --
            # eight launches, nothing for the host to wait on
            with tr.phase("pack"):
                self._pack_shard()
                K.pack(self.q_slice, self.stats, self.q_stats, out=(self.qh, self.qn, self.qe_))
            with tr.phase("dist_topk"):
                K.dist_topk(self.qh, self.xh, self.xn, cand=self.cand, kmax=self.kmax)
            K.zero_(nflag)
            with tr.phase("rerank+vote"):
                K.rerank_ex(self.cand, self.q_slice, self.x, self.labels, self.k_slice, self.qn, self.qe_, self.stats,
                            self.stats, self.id_base, self.kmax, flags, nflag, out=(od, oi, ol), final=self.final,
                            flagged_rows=self.flagged_rows, x_i32=self.x32)
            with tr.phase("repair"):
                K.repair_flagged(self.q_slice, self.x, self.labels, self.k_slice, self.flagged_rows, nflag,
                                 self.id_base, self.kmax, out=(od, oi, ol), final=self.final, x_i32=self.x32)
            self.launch_count += 9   # pack x2, dist_topk, rerank x3, repair x3
            return self.final[0][:self.nq], self.final[1][:self.nq]
        with tr.phase("query_gather"):
            q_all, k_all = self._gather_queries_nccl()
        if self.fused_ok:
            with tr.phase("pack"):
                K.pack(self.x, self.stats, self.stats, out=(self.xh, self.xn, self.xe))
                K.pack(q_all, self.stats, self.q_stats, out=(self.qh, self.qn, self.qe_))
            with tr.phase("dist_topk"):
                K.dist_topk(self.qh, self.xh, self.xn, cand=self.cand, kmax=self.kmax)
            K.zero_(nflag)
            with tr.phase("rerank"):
                K.rerank(self.cand, q_all, self.x, self.labels, k_all, self.qn, self.qe_, self.stats, self.stats,
                         self.id_base, self.kmax, out=self.lists)
            self.launch_count += 6   # pack x2, dist_topk, rerank x3
            flagged = int(nflag.item())
            if flagged:
                rows = torch.nonzero(flags, as_tuple=False).flatten().to(torch.int32)
                K.exact(q_all, self.x, self.labels, k_all, self.kmax, self.id_base, rows, out=(od, oi, ol))
                self.launch_count += 2
            self.last_flagged = flagged
        else:
            K.exact(q_all, self.x, self.labels, k_all, self.kmax, self.id_base, out=(od, oi, ol))
            self.launch_count += 2
        with tr.phase("reduce+finalize"):
            return self._reduce_and_finalize_nccl(od, oi, ol)

    def _reduce_and_finalize_nccl(self, od, oi, ol):
        """k-lists of every data shard -> owner of the query slice -> merge, vote, checksum."""
        R, qs, km = self.R, self.qs, self.kmax
        if R == 1:
            K.finalize_ex(od[None], oi[None], ol[None], self.k_slice, out=self.final)
            self.launch_count += 1
            return self.final[0][:self.nq], self.final[1][:self.nq]
        for src, dst in zip((od, oi, ol), self.recv):
            dist.all_to_all_single(dst.view(R * qs, km), src.view(R * qs, km), group=self.col_group)
        mine = self.k_all[self.r * qs:(self.r + 1) * qs]
        K.finalize_ex(self.recv[0], self.recv[1], self.recv[2], mine, out=self.final)
        self.launch_count += 1
        return self.final[0][:self.nq], self.final[1][:self.nq]

    # ---------------------------------------------------------------- end to end
    E2E_CHUNK_ROWS = 9472    # 37 cluster rows of 256 queries x 2 chunks: one whole wave of the filter GEMM

    def stage_host_inputs(self, wire="auto"):
        """The loader's output: this rank's inputs in pinned host memory, plus the value
        range (so no device pass over the data is needed before packing can start).

        wire = "fp64" keeps doubles on the wire; "decimal6" sends the six-decimal values of
        the text grammar as 32-bit integers (half the PCIe bytes, expanded bit-exactly on
        the device); "auto" picks decimal6 when every value on every rank round-trips.  When the
        integers need fewer than 32 bits (values in [-100, 100]: 28) the rows travel bit-packed
        (`wire_bits` per value, K.pack_bits / K.unpack_bits; DMLP_WIRE_BITS=0 keeps int32)."""
        dev = self.device
        with torch.cuda.device(dev):
            use_dec = False
            if wire in ("auto", "decimal6"):
                xi, ok_x = K.encode_decimal6(self.x_part)
                qi, ok_q = K.encode_decimal6(self.q_slice)
                ok = torch.tensor([1 if (ok_x and ok_q) else 0], dtype=torch.int32, device=dev)
                if self.world > 1:
                    dist.all_reduce(ok, op=dist.ReduceOp.MIN)
                use_dec = bool(ok.item())
                if wire == "decimal6" and not use_dec:
                    raise ValueError("inputs are not representable in the decimal6 wire format")
            lo = min(self.x_part.min().item() if self.x_part.numel() else 0.0,
                     self.q_slice.min().item() if self.q_slice.numel() else 0.0)
            hi = max(self.x_part.max().item() if self.x_part.numel() else 0.0,
                     self.q_slice.max().item() if self.q_slice.numel() else 0.0)
            rng = torch.tensor([-lo, hi], dtype=torch.float64, device=dev)
            if self.world > 1:
                dist.all_reduce(rng, op=dist.ReduceOp.MAX)
            self.host_range = (-rng[0].item(), rng[1].item())
            self.wire = "decimal6" if use_dec else "fp64"
            self.wire_bits = 0
            if use_dec and os.environ.get("DMLP_WIRE_BITS", "auto") != "0":
                big = torch.tensor([max(int(xi.abs().max().item()) if xi.numel() else 0,
                                        int(qi.abs().max().item()) if qi.numel() else 0)], dtype=torch.int64, device=dev)
                if self.world > 1:
                    dist.all_reduce(big, op=dist.ReduceOp.MAX)
                bits = max(int(big.item()).bit_length() + 1, 2)       # i + 2^(bits-1) in [0, 2^bits)
                forced = os.environ.get("DMLP_WIRE_BITS", "auto")
                if forced != "auto":
                    bits = max(bits, int(forced))
                if bits < 32 and K.wire_words(self.a, bits) < self.a:
                    self.wire_bits = bits
                    xi, qi = K.pack_bits(xi, bits), K.pack_bits(qi, bits)
            self.wire_cols = K.wire_words(self.a, self.wire_bits) if self.wire_bits else self.a
            self.host = {
                "x": (xi if use_dec else self.x_part).cpu().pin_memory(),
                "labels": self.labels_part.cpu().pin_memory(),
                "q": (qi if use_dec else self.q_slice).cpu().pin_memory(),
                "k": self.k_slice.cpu().pin_memory(),
            }
            self.host_out = (torch.empty(self.nq, dtype=torch.int32).pin_memory(),
                             torch.empty(self.nq, dtype=torch.int64).pin_memory())
            self.h2d_bytes_per_step = sum(t.numel() * t.element_size() for t in self.host.values())
            self.d2h_bytes_per_step = sum(t.numel() * t.element_size() for t in self.host_out)
            self.streamed = self.R == 1 and self.fused_ok and self.impl == "fused"
            self.copy_stream = torch.cuda.Stream(device=dev)
            rows = min(int(os.environ.get("DMLP_E2E_QROWS", self.E2E_CHUNK_ROWS)), max(self.nq, 1))
            self.chunk_rows = rows
            if use_dec:
                # two complete sets of wire buffers: the copies of step s + 1 run while step s computes
                self.x_wire = [torch.empty((self.x_part.shape[0], self.wire_cols), dtype=torch.int32, device=dev)
                               for _ in range(2)]
                self.q_wire = [torch.empty((max(self.nq, 1), self.wire_cols), dtype=torch.int32, device=dev)
                               for _ in range(2)]
                self.l_wire = [torch.empty_like(self.labels_part) for _ in range(2)]
                self.k_wire = [torch.empty_like(self.k_slice) for _ in range(2)]
            self._e2e_step, self._prefetched = 0, None
            nchunks = max(1, -(-max(self.nq, 1) // rows))
            self.nflag_chunks = torch.zeros(nchunks, dtype=torch.int32, device=dev)
            self.flagged_rows_chunks = torch.zeros((nchunks, K.REPAIR_CAP), dtype=torch.int32, device=dev)
            self.host_nflag = torch.zeros(nchunks, dtype=torch.int32).pin_memory()
            self.d2h_bytes_per_step += self.host_nflag.numel() * 4
            self._cand_by_rows = {}
            self.x_parts = max(1, int(os.environ.get("DMLP_E2E_XPARTS", "8")))
            while self.x_parts > 1 and self.shard_rows < 4096 * self.x_parts:
                self.x_parts //= 2
            self._tiled = None
            torch.cuda.synchronize(dev)

    def run_from_host(self, prefetch_next=False):
        """Public end-to-end call: pinned host inputs -> device -> job -> host results.
        With a single-rank column the copies overlap the kernels (two streams).  With
        `prefetch_next` (decimal6 wire only) the call also starts the host-to-device copies of the
        NEXT call's inputs (the staged host buffers) into the second set of wire buffers before it
        waits for its own results -- the input pipeline of a serving loop: PCIe stays busy while
        the GPU finishes this step's tail."""
        with torch.cuda.device(self.device):
            self._flagged_host = None
            self._want_prefetch = bool(prefetch_next) and self.wire == "decimal6" and getattr(self, "streamed", False)
            streamed = getattr(self, "streamed", False)
            pipelined = streamed and self._prefetched is not None and self._prefetched[0] == (self._e2e_step & 1)
            if pipelined:
                # this step's inputs were copied (or are being copied) into wire set `slot` during the
                # previous call: start the next call's copies first, so PCIe never idles, then run the
                # whole-size kernels on what has arrived -- a two-stage pipeline, copy | compute
                slot, pieces, events = self._prefetched
                nxt = (slot ^ 1, pieces, self._copy_pieces(pieces, slot ^ 1)) if self._want_prefetch else None
                label, checksum = self._run_from_wire(slot, events)
                self._e2e_step += 1
                self._prefetched = nxt
            else:
                if streamed and self.world == 1 and self.x_parts > 1:
                    label, checksum = self._run_tiled()
                elif streamed:
                    label, checksum = self._run_streamed()
                else:
                    label, checksum = self._run_copy_then_compute()
                if streamed:
                    self._e2e_step += 1
                    self._prefetched = None
                    if self._want_prefetch:
                        slot = self._e2e_step & 1
                        self._prefetched = (slot, self._pieces, self._copy_pieces(self._pieces, slot))
            self.host_out[0].copy_(label, non_blocking=True)
            self.host_out[1].copy_(checksum, non_blocking=True)
            streamed = streamed and not pipelined      # (which flagged-query counter this step used)
            counted = self._flagged_host is None and self.fused_ok and self.impl == "fused"
            if counted and streamed:
                self.host_nflag.copy_(self.nflag_chunks, non_blocking=True)
            elif counted:
                self.host_nflag[:1].copy_(self._nflag_tensor(), non_blocking=True)
            torch.cuda.current_stream().synchronize()
            if counted:
                worst = int(self.host_nflag.max()) if streamed else int(self.host_nflag[0])
                self._flagged_host = int(self.host_nflag.sum()) if streamed else worst
                if worst > K.REPAIR_CAP:   # beyond what the step repairs by itself (not seen in practice)
                    self.ensure_repaired(force=True)
                    self.host_out[0].copy_(self.final[0][:self.nq])
                    self.host_out[1].copy_(self.final[1][:self.nq])
            return self.host_out

    def _run_from_wire(self, slot, events):
        """A step whose inputs already sit (decimal6) in wire set `slot`: expand them and run the
        resident pipeline at full size."""
        main = torch.cuda.current_stream()
        main.wait_event(events[-1])            # one copy stream, in order: the last piece closes the set
        self._wait_part_readers()              # the row's peers may still be pulling the previous step's rows
        self._install_small(slot, events)
        self._install_x(self.x_wire[slot])
        self._wire_to_f64(self.q_wire[slot][: self.nq], self.q_slice)
        K.stats_set(self.stats, *self.host_range)
        self.launch_count += 2
        return self.run_resident()

    def _refresh_x32(self, b=None, e=None):
        """fp64 rows just arrived in x_part (fp64 wire) while the hot path reads decimal6 rows: re-encode."""
        if self.dec6:
            src = self.x_part if b is None else self.x_part[b:e]
            K.encode_decimal6_into(src, self.x32_part if b is None else self.x32_part[b:e])

    def _install_x(self, wire):
        """This rank's part of the data shard has arrived in decimal6 form: with resident decimal6
        rows it goes where the hot path (and, on a row of GPUs, the peers) read it -- one small
        kernel, no expansion; the fp64 copy of the part is then NOT refreshed (the direct kernels read
        the decimal6 rows too).  Otherwise it is expanded to fp64."""
        if self.dec6:
            self._wire_to_i32(wire, self.x32_part)
        else:
            self._wire_to_f64(wire, self.x_part)

    def _wire_to_f64(self, wire, dst):
        """Rows that arrived in the decimal6 wire form (int32, or bit-packed words) -> fp64."""
        if self.wire_bits:
            K.unpack_bits(wire, self.a, self.wire_bits, dst)
        else:
            K.expand_decimal6(wire, dst)

    def _wire_to_i32(self, wire, dst):
        """Rows that arrived in the decimal6 wire form -> the resident int32 rows."""
        if self.wire_bits:
            K.unpack_bits(wire, self.a, self.wire_bits, dst)
        else:
            torch.add(wire, 0, out=dst)

    def _rerank_chunk(self, i, b, e, cand):
        """Re-rank + folded vote + device-side repair of query rows [b, e) (chunk i of a streamed step)."""
        od, oi, ol, flags, _ = self.lists
        nf, rows = self.nflag_chunks[i:i + 1], self.flagged_rows_chunks[i]
        out = (od[b:e], oi[b:e], ol[b:e])
        fin = (self.final[0][b:e], self.final[1][b:e])
        K.rerank_ex(cand, self.q_slice[b:e], self.x, self.labels, self.k_slice[b:e], self.qn[b:e], self.qe_[b:e],
                    self.stats, self.stats, self.id_base, self.kmax, flags[b:e], nf, out=out, final=fin,
                    flagged_rows=rows, x_i32=self._x32_step)
        K.repair_flagged(self.q_slice[b:e], self.x, self.labels, self.k_slice[b:e], rows, nf, self.id_base, self.kmax,
                         out=out, final=fin, x_i32=self._x32_step)
        self.launch_count += 6

    def _wait_part_readers(self, upto=None):
        """Before this rank's part of the data shard is overwritten: every rank of the row
        must have finished pulling the previous step's copy."""
        step = self.epoch if upto is None else upto
        if self.peer_rows and step > 0:
            self.row_sym.wait("done_flags", step, self.status)

    def _run_copy_then_compute(self):
        h = self.host
        self._wait_part_readers()
        if self.wire == "decimal6":
            self.x_wire[0].copy_(h["x"], non_blocking=True)
            self._install_x(self.x_wire[0])
            self.q_wire[0][: self.nq].copy_(h["q"], non_blocking=True)
            self._wire_to_f64(self.q_wire[0][: self.nq], self.q_slice)
        else:
            self.x_part.copy_(h["x"], non_blocking=True)
            self.q_slice.copy_(h["q"], non_blocking=True)
            self._refresh_x32()
        self.labels_part.copy_(h["labels"], non_blocking=True)
        self.k_slice.copy_(h["k"], non_blocking=True)
        return self.run_resident()

    def _tiled_schedule(self):
        """Static schedule of the single-GPU end-to-end step: the data shard is cut into
        `x_parts` row ranges and the queries into chunks; pieces are copied in an order that
        lets work start after the first two pieces, and each (query chunk, data part) pair is
        filtered as soon as both have arrived."""
        if self._tiled is not None:
            return self._tiled
        n, nq, rows = self.shard_rows, self.nq, self.chunk_rows
        part = _round_up(-(-n // self.x_parts), 256)
        xparts = [(b, min(b + part, n)) for b in range(0, n, part)]
        qch = [(b, min(b + rows, nq)) for b in range(0, nq, rows)]
        cands, bases = [], []
        for (b, e) in qch:
            base, acc = [], 0
            for (pb, pe) in xparts:
                base.append(acc)
                acc += K.topk_plan(e - b, pe - pb, self.a_pad, kmax=self.kmax)["chunks"]
            bases.append(base)
            cands.append(torch.empty((e - b, acc, K.KP), dtype=torch.int64, device=self.device))
        # copy order: the interleaving with the smallest simulated makespan (the work is a product
        # of queries x data, so what arrives last decides how much is left to do after the copies)
        order = best_copy_order([e - b for b, e in qch], [pe - pb for pb, pe in xparts], self.a,
                                (self.wire_bits / 8.0 if self.wire_bits else 4) if self.wire == "decimal6" else 8)
        self._tiled = (xparts, qch, cands, bases, order)
        return self._tiled

    def _copy_pieces(self, pieces, slot):
        """Enqueue on the copy stream, in order, the host-to-device copies of `pieces` -- (kind, begin,
        end) row ranges of the data part ("x") or of the queries ("q") -- into wire set `slot`
        (decimal6) or straight into the compute buffers (fp64 wire).  One event per piece."""
        h, cs, dec = self.host, self.copy_stream, self.wire == "decimal6"
        events = []
        with torch.cuda.stream(cs):
            for kind, b, e in pieces:
                if kind == "x":
                    (self.x_wire[slot] if dec else self.x_part)[b:e].copy_(h["x"][b:e], non_blocking=True)
                elif kind == "q":
                    (self.q_wire[slot] if dec else self.q_slice)[b:e].copy_(h["q"][b:e], non_blocking=True)
                elif kind == "l":     # labels and k travel first, in the same queue as everything else
                    (self.l_wire[slot] if dec else self.labels_part).copy_(h["labels"], non_blocking=True)
                else:
                    (self.k_wire[slot] if dec else self.k_slice).copy_(h["k"], non_blocking=True)
                ev = torch.cuda.Event()
                ev.record(cs)
                events.append(ev)
        return events

    def _step_copies(self, pieces):
        """Events of this step's copies: the ones a previous call prefetched, or enqueued now."""
        slot = self._e2e_step & 1
        self._pieces = pieces
        if self._prefetched is not None and self._prefetched[0] == slot and self._prefetched[1] == pieces:
            return slot, self._prefetched[2]
        if self.wire != "decimal6":
            self.copy_stream.wait_stream(torch.cuda.current_stream())   # the compute buffers themselves are the target
        return slot, self._copy_pieces(pieces, slot)

    def _install_small(self, slot, events):
        """Labels and k of this step: they were copied first (pieces 0 and 1); on the decimal6 path
        they sit in the wire set and are moved into place by two small kernels (not by copies: a
        copy would queue behind the large transfers the copy engine is working through)."""
        main = torch.cuda.current_stream()
        main.wait_event(events[1])
        if self.wire == "decimal6":
            torch.add(self.l_wire[slot], 0, out=self.labels_part)
            torch.add(self.k_wire[slot], 0, out=self.k_slice)

    def _small_pieces(self):
        return [("l", 0, int(self.labels_part.numel())), ("k", 0, int(self.k_slice.numel()))]

    def _run_tiled(self):
        main = torch.cuda.current_stream()
        dec = self.wire == "decimal6"
        xparts, qch, cands, bases, order = self._tiled_schedule()
        pieces = self._small_pieces() + [("x",) + tuple(xparts[i]) if kind == "x" else ("q",) + tuple(qch[i])
                                         for kind, i in order]
        self.epoch += 1
        slot, events = self._step_copies(pieces)
        self._install_small(slot, events)
        # int32 wire: pack and re-rank read the wire set itself (no expansion at all); bit-packed wire:
        # each arriving part is unpacked into the resident int32 rows first
        direct = dec and self.dec6 and not self.wire_bits
        self._x32_step = (self.x_wire[slot] if direct else self.x32) if self.dec6 else None
        K.stats_reset(self.stats)
        K.stats_reset(self.q_stats)
        K.stats_set(self.stats, *self.host_range)
        K.zero_(self.nflag_chunks)
        self.launch_count += 2
        have_x, have_q = set(), set()
        done, reranked = set(), set()
        for (kind, idx), ev in zip(order, events[2:]):
            main.wait_event(ev)
            if kind == "x":
                pb, pe = xparts[idx]
                out = (self.xh[pb:pe], self.xn[pb:pe], self.xe[pb:pe])
                if dec and self.dec6:
                    if not direct:
                        self._wire_to_i32(self.x_wire[slot][pb:pe], self.x32[pb:pe])
                        self.launch_count += 1
                    K.pack_i32(self._x32_step[pb:pe], self.stats, self.stats, out=out)
                else:
                    if dec:
                        self._wire_to_f64(self.x_wire[slot][pb:pe], self.x_part[pb:pe])
                    else:
                        self._refresh_x32(pb, pe)
                    K.pack(self.x[pb:pe], self.stats, self.stats, out=out)
                have_x.add(idx)
            else:
                b, e = qch[idx]
                if dec:
                    self._wire_to_f64(self.q_wire[slot][b:e], self.q_slice[b:e])
                K.pack(self.q_slice[b:e], self.stats, self.q_stats, out=(self.qh[b:e], self.qn[b:e], self.qe_[b:e]))
                have_q.add(idx)
            self.launch_count += 1 + (1 if dec else 0)
            for i in sorted(have_q):
                b, e = qch[i]
                for p_ in sorted(have_x):
                    if (i, p_) in done:
                        continue
                    pb, pe = xparts[p_]
                    K.dist_topk_part(self.qh[b:e], self.xh[pb:pe], self.xn[pb:pe], cands[i], bases[i][p_], pb,
                                     kmax=self.kmax)
                    done.add((i, p_))
                    self.launch_count += 1
                if i not in reranked and all((i, p_) in done for p_ in range(len(xparts))):
                    self._rerank_chunk(i, b, e, cands[i])
                    reranked.add(i)
        return self.final[0][:self.nq], self.final[1][:self.nq]

    def _run_streamed(self):
        main = torch.cuda.current_stream()
        dec = self.wire == "decimal6"
        nq, rows = self.nq, self.chunk_rows
        chunks = [(b, min(b + rows, nq)) for b in range(0, nq, rows)]
        n_part = self.part_rows[1] - self.part_rows[0]
        pieces = self._small_pieces() + [("x", 0, n_part)] + [("q", b, e) for b, e in chunks]
        # the rows, labels and packed rows of this rank's part are what the row's peers pull from: nothing
        # may overwrite them before every peer has finished with the previous step's copy
        if not dec:
            self._wait_part_readers()
        slot, events = self._step_copies(pieces)
        self.epoch += 1
        main.wait_event(events[2])
        if dec:
            self._wait_part_readers(self.epoch - 1)
        self._install_small(slot, events)
        # data shard: expand (or install the decimal6 rows), all-gather along the row, pack
        if dec:
            self._install_x(self.x_wire[slot])
        else:
            self._refresh_x32()
        self._x32_step = self.x32
        K.stats_reset(self.stats)
        K.stats_reset(self.q_stats)
        K.stats_set(self.stats, *self.host_range)
        gather_pending = self.rows_fused
        if gather_pending:
            self._pack_part_and_signal()
        else:
            self._gather_shard()
            self._pack_shard()
        K.zero_(self.nflag_chunks)
        self.launch_count += 4 + (1 if dec else 0)
        for i, (b, e) in enumerate(chunks):
            main.wait_event(events[3 + i])
            if dec:
                self._wire_to_f64(self.q_wire[slot][b:e], self.q_slice[b:e])
            qh, qn, qe = self.qh[b:e], self.qn[b:e], self.qe_[b:e]
            K.pack(self.q_slice[b:e], self.stats, self.q_stats, out=(qh, qn, qe))
            key = e - b
            if key not in self._cand_by_rows:
                plan = K.topk_plan(key, max(self.shard_rows, 1), self.a_pad, kmax=self.kmax)
                self._cand_by_rows[key] = torch.empty((key, plan["chunks"], plan["kp"]), dtype=torch.int64,
                                                      device=self.device)
            cand = self._cand_by_rows[key]
            if gather_pending:      # the first GEMM of the step also gathers the data shard
                K.dist_topk_rows(qh, self.shard_rows, cand, self.data_gather, kmax=self.kmax)
                K.maxima_from_norms(self.xn, self.xe, self.stats)
                gather_pending = False
                self.launch_count += 1
            else:
                K.dist_topk(qh, self.xh, self.xn, cand=cand, kmax=self.kmax)
            self._rerank_chunk(i, b, e, cand)
            self.launch_count += 2 + (1 if dec else 0)
        return self.final[0][:self.nq], self.final[1][:self.nq]

    def check_watchdog(self):
        """Non-zero means a device-side wait gave up somewhere: a positive code names
        the peer signal that never came, a negative one counts re-alignment waits of the filter
        kernel that timed out (results stay correct, the sweep lost its L2 sharing)."""
        st = int(self.status.item())
        return st if st else -K.topk_sync_timeouts()

    def close(self):
        for s in (self.sym, self.row_sym):
            if s is not None:
                s.close()
        self.sym = self.row_sym = None

    # -------------------------------------------------------------- verification
    def verify_sampled(self, n_exact=64, n_cpu=4, seed=1234):
        """Check the results of the last run_resident() on sampled queries, at any problem size.

        * `n_exact` queries per rank against the independent direct fp64 kernel (exact.cu) on this
          rank's data shard; with several data shards the sampled queries' local lists are
          all-gathered along the column (a few KB) and merged, and each owner compares its own.
        * `n_cpu` queries against the CPU oracle, driven from rank 0's column only: every rank
          of that column runs the CPU twin on its own shard (one device-to-host copy of the
          shard it already holds), rank 0 merges, votes and compares.  Nothing gathers the
          whole dataset anywhere.
        Raises AssertionError on any mismatch; returns the number of distinct queries that passed,
        summed over the job (n_exact per grid column -- every rank of the column computes the
        direct kernel on its shard for all of them -- plus n_cpu)."""
        from ..ops import cpu
        self.ensure_repaired()
        torch.cuda.synchronize(self.device)
        dev, R, km = self.device, self.R, self.kmax
        fin_label, fin_sum = self.final[0][:self.nq].cpu().numpy(), self.final[1][:self.nq].cpu().numpy()
        gen = np.random.default_rng(seed + 7919 * self.c)            # the ranks of a column draw the same sample
        counts = [even_slice(self.q_col, R, r)[1] - even_slice(self.q_col, R, r)[0] for r in range(R)]

        def sample(count):
            owners = gen.choice([r for r in range(R) if counts[r] > 0], size=count)
            local = np.array([gen.integers(0, counts[o]) for o in owners], dtype=np.int64)
            return owners, local

        passed = 0
        if self.q_col > 0 and n_exact > 0:
            owners, local = sample(n_exact)
            if R == 1:
                q_rows, k_rows = self.q_slice, self.k_slice
                rows = torch.from_numpy(local.astype(np.int32)).to(dev)
            else:   # padded column index space: owner * qs + local row
                q_rows, k_rows = self.q_all, self.k_all
                rows = torch.from_numpy((owners * self.qs + local).astype(np.int32)).to(dev)
            kk = k_rows[rows.long()].contiguous()
            qsel = q_rows[rows.long()].contiguous()
            od, oi, ol = K.exact(qsel, self.x, self.labels, kk, km, self.id_base, x_i32=self.x32)
            if R > 1:
                parts = [[torch.empty_like(t) for _ in range(R)] for t in (od, oi, ol)]
                for t, outs in zip((od, oi, ol), parts):
                    dist.all_gather(outs, t, group=self.col_group)
                od, oi, ol = (torch.stack(p) for p in parts)
            else:
                od, oi, ol = od[None], oi[None], ol[None]
            label, checksum, _, _ = K.finalize(od, oi, ol, kk)
            label, checksum = label.cpu().numpy(), checksum.cpu().numpy()
            for j in range(len(local)):
                if owners[j] != self.r:
                    continue
                if label[j] != fin_label[local[j]] or checksum[j] != fin_sum[local[j]]:
                    raise AssertionError(f"rank {self.rank}: query {local[j]} of this rank differs from the direct "
                                         f"fp64 kernel (label {fin_label[local[j]]} vs {label[j]})")
                passed += 1
        if self.c == 0 and self.q_col > 0 and n_cpu > 0:
            owners, local = sample(n_cpu)
            if R == 1:
                qsel, ksel = self.q_slice[local].cpu().numpy(), self.k_slice[local].cpu().numpy()
            else:
                idx = torch.from_numpy(owners * self.qs + local).to(dev)
                qsel, ksel = self.q_all[idx].cpu().numpy(), self.k_all[idx].cpu().numpy()
            # (with resident decimal6 rows the fp64 copy only holds this rank's own part)
            x_host = (K.expand_decimal6(self.x32, torch.empty(self.x32.shape, dtype=torch.float64, device=dev))
                      if self.dec6 else self.x).cpu().numpy()
            lists = cpu.local_lists(x_host, self.labels.cpu().numpy(), self.id_base, qsel, ksel, km)
            mine = {int(j): (int(fin_label[local[j]]), int(fin_sum[local[j]])) for j in range(n_cpu)
                    if owners[j] == self.r}
            if R > 1:
                gathered = [None] * R if self.r == 0 else None
                dist.gather_object((lists, mine), gathered, dst=self.grid.column_ranks(0)[0], group=self.col_group)
            else:
                gathered = [(lists, mine)]
            if self.r == 0:
                d, i, l = (np.stack([g[0][t] for g in gathered]) for t in range(3))
                want_label, want_sum = cpu.finalize_lists(d, i, l, ksel)
                got = {}
                for g in gathered:
                    got.update(g[1])
                for j in range(n_cpu):
                    if got[j] != (int(want_label[j]), int(want_sum[j].view(np.int64))):
                        raise AssertionError(f"query {local[j]} of row {owners[j]} differs from the CPU oracle: "
                                             f"{got[j]} vs {(int(want_label[j]), int(want_sum[j].view(np.int64)))}")
                    passed += 1
        t = torch.tensor([passed], dtype=torch.int64, device=dev)
        if self.world > 1:
            dist.all_reduce(t)
        return int(t.item())

    def verify_against_oracle(self, count):
        """Re-derive the first `count` owned queries on the CPU oracle (gathers the whole
        dataset on the host: only sensible for small problems)."""
        from ..ops import cpu
        label, checksum = self.run_resident()
        self.ensure_repaired()
        torch.cuda.synchronize(self.device)
        if self.world > 1:
            parts = [None] * self.world     # parts differ in size: exchange as objects
            dist.all_gather_object(parts, (self.x_part.cpu().numpy(), self.labels_part.cpu().numpy()))
            # global row order = shard by shard, part by part = rank order r * C + c
            x = np.concatenate([p[0] for p in parts])
            lab = np.concatenate([p[1] for p in parts])
        else:
            x, lab = self.x.cpu().numpy(), self.labels.cpu().numpy()
        c = min(count, self.nq)
        want_label, want, *_ = cpu.knn(x, lab, self.q_slice[:c].cpu().numpy(), self.k_slice[:c].cpu().numpy())
        got = checksum[:c].cpu().numpy().view(np.uint64)
        if not (got == want).all():
            raise AssertionError(f"rank {self.rank}: {(got != want).sum()} of {c} checksums differ from the oracle")
        if not (label[:c].cpu().numpy() == want_label).all():
            raise AssertionError(f"rank {self.rank}: labels differ from the oracle")
        return c
