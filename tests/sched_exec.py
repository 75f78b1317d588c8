"""Replays the engine's per-step schedule (dlafb_schedule_step, the SAME host code the GPU scheduler runs) on the
CPU with the oracle's tile arithmetic. Test infrastructure: lets the CPU suite check ownership, panel routing and
update lists of the distributed algorithm without a GPU, in one process (simulated ranks) or over a real
torch.distributed (gloo) process group."""
import numpy as np

import dlaf_oracle as orc


class Rank:
    """State of one rank: its local block in 2D block-cyclic layout plus the two panel workspaces."""

    def __init__(self, pkg, L, a_global, nb, grid, coords, src):
        self.pkg, self.L = pkg, L
        self.n, self.nb = a_global.shape[0], nb
        (self.P, self.Q), (self.pr, self.pc), (self.sr, self.sc) = grid, coords, src
        self.dt = a_global.dtype
        self.rt = L.dlafb_dist_local_nr_tiles(self.n, nb, self.P, self.pr, self.sr)
        self.ct = L.dlafb_dist_local_nr_tiles(self.n, nb, self.Q, self.pc, self.sc)
        self.grow = [L.dlafb_dist_global_tile_from_local(l, self.P, self.pr, self.sr) for l in range(self.rt)]
        self.gcol = [L.dlafb_dist_global_tile_from_local(l, self.Q, self.pc, self.sc) for l in range(self.ct)]
        self.mloc = L.dlafb_dist_local_size(self.n, nb, self.P, self.pr, self.sr)
        self.nloc = L.dlafb_dist_local_size(self.n, nb, self.Q, self.pc, self.sc)
        rows = np.concatenate([np.arange(g * nb, min(self.n, (g + 1) * nb)) for g in self.grow] + [np.arange(0)])
        cols = np.concatenate([np.arange(g * nb, min(self.n, (g + 1) * nb)) for g in self.gcol] + [np.arange(0)])
        self.rows, self.cols = rows.astype(int), cols.astype(int)
        self.a = np.asfortranarray(a_global[np.ix_(self.rows, self.cols)])
        assert self.a.shape == (self.mloc, self.nloc)
        self.Pr, self.Pc, self.diag = {}, {}, None

    def step(self, k):
        return self.pkg.schedule_step(self.L, self.n, self.nb, self.P, self.Q, self.pr, self.pc, self.sr, self.sc, k)

    def tile(self, il, jl):
        nb = self.nb
        return self.a[il * nb:min(self.mloc, (il + 1) * nb), jl * nb:min(self.nloc, (jl + 1) * nb)]

    # ---- phases of one step, in the order the hp stream issues them ----
    def potrf(self, st):
        if st.own_diag:
            t = self.tile(st.lk_row, st.lk_col)
            info = orc.tile_potrf("L", t)
            assert info == 0
            self.diag = np.array(t, order="F")

    def solve_and_pack(self, st):
        """Panel solve of the local part of column k and copy into the row panel slots il0.."""
        self.Pr, self.Pc = {}, {}
        if st.in_col:
            for il in range(st.il0, st.il_end):
                t = self.tile(il, st.lk_col)
                orc.tile_trsm_L(self.diag, t)
                self.Pr[il] = np.array(t, order="F")

    def update(self, items):
        nb = self.nb
        for row, col, diag in items:
            il, jl = row // nb, col // nb
            c = self.tile(il, jl)
            if diag:
                orc.tile_herk_L(self.Pr[il], c)
            else:
                orc.tile_gemm_L(self.Pr[il], self.Pc[jl], c)


def gather(ranks, n, dt):
    out = np.zeros((n, n), dtype=dt, order="F")
    for r in ranks:
        out[np.ix_(r.rows, r.cols)] = r.a
    return out


def run_simulated(pkg, L, a_global, nb, grid, src=(0, 0)):
    """All ranks in one process; broadcasts are copies from the root's buffers."""
    P, Q = grid
    n = a_global.shape[0]
    R = {(p, q): Rank(pkg, L, a_global, nb, grid, (p, q), src) for p in range(P) for q in range(Q)}
    nt = L.dlafb_dist_nr_tiles(n, nb)
    for k in range(nt):
        S = {rc: r.step(k) for rc, r in R.items()}
        for rc, r in R.items():
            r.potrf(S[rc][0])
        if k == nt - 1:
            break
        for rc, r in R.items():                      # diagonal tile down the grid column that owns column k
            st = S[rc][0]
            if st.in_col:
                r.diag = R[(st.diag_root, rc[1])].diag
                assert r.diag is not None
        for rc, r in R.items():
            r.solve_and_pack(S[rc][0])
        for rc, r in R.items():                      # row panel along the grid row
            st = S[rc][0]
            root = R[(rc[0], st.row_root)]
            for il in range(st.il0, st.il_end):
                r.Pr[il] = root.Pr[il]
        for rc, r in R.items():                      # transposed panel down the grid column
            for jl, g, root, src_il in S[rc][1]:
                holder = R[(root, rc[1])]
                assert holder.grow[src_il] == g and r.gcol[jl] == g
                r.Pc[jl] = holder.Pr[src_il]
        for rc, r in R.items():
            _, _, hp, np_ = S[rc]
            k1 = [(row, col) for row, col, _ in hp]
            assert all(r.gcol[col // nb] == k + 1 for _, col in k1)          # look-ahead list = global column k+1
            assert all(r.gcol[col // nb] > k + 1 for _, col, _ in np_)
            r.update(hp)
            r.update(np_)
    return gather(R.values(), n, a_global.dtype)


def run_process_group(pkg, L, a_global, nb, grid, src, rank, dist, row_group, col_group, row_ranks, col_ranks):
    """One rank of a real process group (torch.distributed, gloo on CPU). row_ranks[c] / col_ranks[r] = world rank of
    the member with that grid coordinate in this rank's row / column."""
    import torch
    P, Q = grid
    coords = (rank // Q, rank % Q)
    r = Rank(pkg, L, a_global, nb, grid, coords, src)
    n = a_global.shape[0]
    nt = L.dlafb_dist_nr_tiles(n, nb)

    def bcast(arr, shape, root_world, group):
        t = torch.from_numpy(np.ascontiguousarray(arr if arr is not None else np.zeros(shape, dtype=r.dt)))
        if np.dtype(r.dt).kind == "c":
            t = torch.view_as_real(t)
        dist.broadcast(t, root_world, group=group)
        out = t.numpy()
        if np.dtype(r.dt).kind == "c":
            out = out.view(r.dt).reshape(shape)
        return np.asfortranarray(out)

    for k in range(nt):
        st, colb, hp, np_ = r.step(k)
        r.potrf(st)
        if k == nt - 1:
            break
        if st.in_col and P > 1:
            r.diag = bcast(r.diag if st.own_diag else None, (nb, nb), col_ranks[st.diag_root], col_group)
        r.solve_and_pack(st)
        if Q > 1:
            for il in range(st.il0, st.il_end):
                h = min(nb, r.mloc - il * nb)
                r.Pr[il] = bcast(r.Pr.get(il), (h, nb), row_ranks[st.row_root], row_group)
        for jl, g, root, src_il in colb:
            h = min(nb, n - g * nb)
            if P > 1:
                r.Pc[jl] = bcast(r.Pr.get(src_il) if root == coords[0] else None, (h, nb), col_ranks[root], col_group)
            else:
                r.Pc[jl] = r.Pr[src_il]
        r.update(hp)
        r.update(np_)
    return r
