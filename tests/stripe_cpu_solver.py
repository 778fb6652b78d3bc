"""TEST INFRASTRUCTURE: a CPU stand-in for one stripe solver of mazetd_b200/striped.py, so the stripe protocol (boundary
row swap, halo merge, termination, global next-hop ids) can be exercised on gloo without a GPU. It is a plain
label-correcting relaxation in numpy over the oracle's edge weights; the checker of the final result is the oracle's
Dijkstra over the WHOLE grid (sources/map/paths.cpp:24-78)."""
import numpy as np
import torch

M = 0xFFFFFFFF
INF = np.int64(M)
F_BLOCKING = 1 | 16384  # Invalid | Building (sources/map/paths.cpp:65)


class CpuStripe:
    def __init__(self, oracle, W, H, y0, y1, elev, flags, sources):
        """elev / flags: FULL host arrays [H*W] (a test may look at everything; only rows y0-1..y1 are used)"""
        self.W, self.H, self.y0, self.y1 = W, H, y0, y1
        r, d = oracle.edge_weights(W, H, elev)  # weight of the edge to the right / down neighbour of every tile
        self.wr = r.reshape(H, W).astype(np.int64)[y0:y1]
        self.wd_full = d.reshape(H, W).astype(np.int64)
        self.blocked = (flags.reshape(H, W)[y0:y1] & F_BLOCKING) != 0
        self.sources = np.asarray(sources, np.int64)
        self.src_mask_full = np.zeros((H, W), bool)
        self.src_mask_full[self.sources // W, self.sources % W] = True
        self.halo_up = np.full(W, INF)
        self.halo_down = np.full(W, INF)
        self.dist = None

    def _relax(self):
        W, y0, y1 = self.W, self.y0, self.y1
        d = self.dist
        src = self.src_mask_full[y0:y1]
        dead = self.blocked & ~src  # a blocked tile is never entered; a blocked SOURCE still expands (paths.cpp:63-67)
        wd = self.wd_full[y0:y1]  # wd[r] = edge (y0 + r) -> (y0 + r + 1)
        while True:
            old = d.copy()
            c = np.full_like(d, INF)
            c[:, 1:] = np.minimum(c[:, 1:], d[:, :-1] + self.wr[:, :-1])  # from the left
            c[:, :-1] = np.minimum(c[:, :-1], d[:, 1:] + self.wr[:, :-1])  # from the right
            c[1:] = np.minimum(c[1:], d[:-1] + wd[:-1])  # from above
            c[:-1] = np.minimum(c[:-1], d[1:] + wd[:-1])  # from below
            if y0 > 0:
                c[0] = np.minimum(c[0], self.halo_up + self.wd_full[y0 - 1])
            if y1 < self.H:
                c[-1] = np.minimum(c[-1], self.halo_down + wd[-1])
            d = np.minimum(d, c)
            d[dead] = INF
            d[d > INF] = INF
            if np.array_equal(d, old):
                break
        self.dist = d

    def begin(self):
        d = np.full((self.y1 - self.y0, self.W), INF)
        d[self.src_mask_full[self.y0:self.y1]] = 0
        # sources on the halo rows start at 0 on their owner; tell the local relaxation right away
        if self.y0 > 0:
            self.halo_up[self.src_mask_full[self.y0 - 1]] = 0
        if self.y1 < self.H:
            self.halo_down[self.src_mask_full[self.y1]] = 0
        self.dist = d
        self._relax()

    def boundary_rows(self):
        t = torch.from_numpy(self.dist)
        return t[0], t[-1]

    def new_row(self):
        return torch.empty(self.W, dtype=torch.int64)

    def exchange(self, row_above, row_below):
        improved = 0
        if row_above is not None:
            a = row_above.numpy()
            improved += int((a < self.halo_up).sum())
            self.halo_up = np.minimum(self.halo_up, a)
        if row_below is not None:
            b = row_below.numpy()
            improved += int((b < self.halo_down).sum())
            self.halo_down = np.minimum(self.halo_down, b)
        if improved:
            self._relax()
        return improved

    def finalize(self):
        """next hop = neighbour u with d[u] + w(u, v) == d[v] minimising (d[u], u) (closed form of the heap order)"""
        W, H, y0, y1 = self.W, self.H, self.y0, self.y1
        ext = np.full((y1 - y0 + 2, W + 2), INF)
        ext[1:-1, 1:-1] = self.dist
        if y0 > 0:
            ext[0, 1:-1] = self.halo_up
        if y1 < H:
            ext[-1, 1:-1] = self.halo_down
        d = self.dist
        rows = np.arange(y0, y1)[:, None]
        cols = np.arange(W)[None, :]
        g = rows * W + cols
        wd = self.wd_full[y0:y1]
        wup = np.zeros_like(d)
        wup[1:] = wd[:-1]
        if y0 > 0:
            wup[0] = self.wd_full[y0 - 1]
        wl = np.zeros_like(d)
        wl[:, 1:] = self.wr[:, :-1]
        cands = [  # (neighbour distance, weight, neighbour global id), any order: the minimum of (d[u], u) is taken
            (ext[:-2, 1:-1], wup, g - W),
            (ext[1:-1, :-2], wl, g - 1),
            (ext[1:-1, 2:], self.wr, g + 1),
            (ext[2:, 1:-1], wd, g + W),
        ]
        best_d = np.full_like(d, INF)
        best_u = np.full_like(d, INF)
        for du, w, u in cands:
            ok = (du < INF) & (du + w == d) & (d < INF) & (d > 0)
            better = ok & ((du < best_d) | ((du == best_d) & (u < best_u)))
            best_d = np.where(better, du, best_d)
            best_u = np.where(better, u, best_u)
        return torch.from_numpy(best_u.astype(np.uint32).view(np.int32)), torch.from_numpy(d.astype(np.uint32).view(np.int32))
