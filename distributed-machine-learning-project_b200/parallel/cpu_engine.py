"""The distributed engine's data flow on the CPU (gloo): same grid, same partitions,
same reduce-scatter of k-lists, with the exact CPU kernels in place of the sm_100a
ones.  It exists so the multi-process plumbing -- who loads what, who owns which
queries, how lists travel -- is testable without GPUs, and it
is the multi-process counterpart of the reference's MPI run."""
import numpy as np
import torch.distributed as dist

from ..ops import cpu
from .grid import Grid, even_slice


class DistributedKNNCpu:
    def __init__(self, grid=None):
        self.world = dist.get_world_size() if dist.is_initialized() else 1
        self.rank = dist.get_rank() if dist.is_initialized() else 0
        self.grid = Grid.for_world(self.world, *(grid or (self.world, 1)))
        self.R, self.C = self.grid.rows, self.grid.cols
        self.r, self.c = self.grid.coords(self.rank)
        self.col_group = self.row_group = None
        if self.world > 1:
            for c in range(self.C):
                g = dist.new_group(self.grid.column_ranks(c))
                if c == self.c:
                    self.col_group = g
            for r in range(self.R):
                g = dist.new_group(self.grid.row_ranks(r))
                if r == self.r:
                    self.row_group = g

    def partition(self, n_total, q_total):
        sb, se, _ = even_slice(n_total, self.R, self.r)
        pb, pe, _ = even_slice(se - sb, self.C, self.c)
        cb, ce, _ = even_slice(q_total, self.C, self.c)
        ob, oe, _ = even_slice(ce - cb, self.R, self.r)
        return (sb + pb, sb + pe), (sb, se), (cb + ob, cb + oe), (cb, ce)

    def run(self, x_part, labels_part, q_slice, k_slice, n_total, q_total, kmax):
        """Returns (label, checksum) numpy arrays for the query slice this rank owns."""
        (pb, pe), (sb, se), (ob, oe), (cb, ce) = self.partition(n_total, q_total)
        # row all-gather of the data-shard parts (reference: row Bcast, engine.cpp:129-131)
        if self.C > 1:
            parts = [None] * self.C
            dist.all_gather_object(parts, (x_part, labels_part), group=self.row_group)
            x = np.concatenate([p[0] for p in parts])
            labels = np.concatenate([p[1] for p in parts])
        else:
            x, labels = x_part, labels_part
        # column all-gather of the owned query slices (reference: column Bcast, engine.cpp:207-209)
        if self.R > 1:
            qs = [None] * self.R
            dist.all_gather_object(qs, (q_slice, k_slice), group=self.col_group)
            sizes = [len(s[1]) for s in qs]
            q_all = np.concatenate([s[0] for s in qs])
            k_all = np.concatenate([s[1] for s in qs])
        else:
            sizes, q_all, k_all = [len(k_slice)], q_slice, k_slice
        od, oi, ol = cpu.local_lists(x, labels, sb, q_all, k_all, kmax)
        # reduce-scatter of k-lists onto the slice owners (reference: engine.cpp:283-308)
        if self.R > 1:
            offs = np.concatenate([[0], np.cumsum(sizes)])
            send = [(od[offs[o]:offs[o + 1]], oi[offs[o]:offs[o + 1]], ol[offs[o]:offs[o + 1]]) for o in range(self.R)]
            recv = [None] * self.R
            for o in range(self.R):
                got = [None] * self.R if self.r == o else None
                dist.gather_object(send[o], got, dst=self.grid.rank_of(o, self.c), group=self.col_group)
                if self.r == o:
                    recv = got
            dist_s = np.stack([g[0] for g in recv])
            id_s = np.stack([g[1] for g in recv])
            lab_s = np.stack([g[2] for g in recv])
        else:
            dist_s, id_s, lab_s = od[None], oi[None], ol[None]
        return cpu.finalize_lists(dist_s, id_s, lab_s, k_slice)
