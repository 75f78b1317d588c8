"""CPU tests (-m "not gpu") of the host logic behind the C ABI: the shared library loads and exports every declared
symbol, the 2D block-cyclic index algebra and the per-step schedule (ownership, panel routing, update lists) match
the oracle's restatement of the reference, and the N > 1 path runs over a real world_size-2 gloo process group.

Reference tests mirrored: test/unit/matrix/test_util_distribution.cpp, test/unit/communication/
test_broadcast_panel.cpp (routing), test/unit/factorization/test_cholesky.cpp:54-139 (distributed cases).
"""
import os
import re

import numpy as np
import pytest

import dlaf_oracle as orc
import sched_exec

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_SIZES = [(0, 2), (5, 8), (34, 34), (4, 3), (16, 10), (34, 13), (32, 5)]
GRIDS = [(3, 2), (2, 3), (3, 1), (1, 2), (1, 1)]


def test_library_exports_every_declared_symbol(pkg, L):
    missing = [s for s in pkg.declared_symbols() if not hasattr(L, s)]
    assert not missing


def test_headers_and_bindings_agree(pkg, L):
    """Every function the public headers declare is exported by the library (and known to the binding)."""
    names = set()
    hdrs = ["include/dlafb_engine.h", "include/dlafb_solver.h", "include/dlafb_gen_to_std.h", "include/dlafb_matrix.h",
            "include/dlaf_c/init.h", "include/dlaf_c/grid.h", "include/dlaf_c/utils.h",
            "include/dlaf_c/factorization/cholesky.h", "include/dlafb_mpi_shim/mpi.h"]
    for h in hdrs:
        src = open(os.path.join(ROOT, h)).read()
        src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
        names |= set(re.findall(r"\b((?:dlafb?_|MPI_Comm_|grid_ordering|make_dlaf_descriptor)\w*)\s*\(", src))
    src = open(os.path.join(ROOT, "include/dlafb_kernels.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    for stem in re.findall(r"\b(dlafb_\w+_)##X\s*\(", src):
        names |= {stem + x for x in "sdcz"}
    names |= set(re.findall(r"\b(dlafb_\w+)\s*\(char type", src))
    names -= {"dlafb_mpi_comm", "dlafb_step_info"}
    assert len(names) > 60
    for nme in sorted(names):
        assert hasattr(L, nme), nme
    assert set(pkg.declared_symbols()) >= names


@pytest.mark.parametrize("P,src", [(1, 0), (2, 0), (2, 1), (3, 2), (4, 1)])
def test_index_algebra_matches_oracle(L, P, src):
    for n, nb in [(0, 3), (1, 1), (13, 5), (100, 10), (95, 10), (64, 64)]:
        for rank in range(P):
            d = orc.Distribution(n, nb, grid=(P, 1), rank=(rank, 0), src=(src, 0))
            assert L.dlafb_dist_nr_tiles(n, nb) == d.nt
            assert L.dlafb_dist_local_nr_tiles(n, nb, P, rank, src) == d.local_nr_tiles(0)
            assert L.dlafb_dist_local_size(n, nb, P, rank, src) == d.local_size(0)
            for g in range(d.nt + 1):
                assert L.dlafb_dist_rank_global_tile(g, P, src) == d.rank_global_tile(g, 0)
                assert L.dlafb_dist_local_tile_from_global(g, P) == d.local_tile_from_global(g, 0)
                assert L.dlafb_dist_next_local_tile_from_global(g, P, rank, src) == d.next_local_tile_from_global(g, 0)
            for l in range(d.local_nr_tiles(0)):
                assert L.dlafb_dist_global_tile_from_local(l, P, rank, src) == d.global_tile_from_local(l, 0)


@pytest.mark.parametrize("grid", GRIDS + [(2, 4), (4, 2)])
def test_panel_routing_matches_oracle(pkg, L, grid):
    """Row-panel slots and transposed-panel broadcasts (who roots which tile) against the oracle's restatement of
    comm::broadcast(root, panel, panelT, row_chain, col_chain), broadcast_panel.h:123-213."""
    P, Q = grid
    src = (max(0, P - 1), min(1, Q - 1))
    for n, nb in [(34, 5), (64, 8), (50, 7)]:
        for pr in range(P):
            for pc in range(Q):
                d = orc.Distribution(n, nb, grid, (pr, pc), src)
                for k in range(d.nt - 1):
                    st, colb, hp, np_ = pkg.schedule_step(L, n, nb, P, Q, pr, pc, src[0], src[1], k)
                    step1, step2 = orc.panel_broadcast_plan(d, k, "col")
                    assert st.own_diag == int(d.rank_global_tile(k, 0) == pr and d.rank_global_tile(k, 1) == pc)
                    assert st.in_col == int(d.rank_global_tile(k, 1) == pc)
                    assert [d.global_tile_from_local(il, 0) for il in range(st.il0, st.il_end)] == [g for g, _ in step1]
                    assert all(root == st.row_root for _, root in step1)
                    assert [(g, root) for _, g, root, _ in colb] == step2
                    for jl, g, root, src_il in colb:
                        assert d.global_tile_from_local(jl, 1) == g
                        assert src_il == d.local_tile_from_global(g, 0)
                    # update lists: exactly the local tiles (gi >= gj > k), look-ahead column first
                    want = {(il, jl) for il in range(d.local_nr_tiles(0)) for jl in range(d.local_nr_tiles(1))
                            if d.global_tile_from_local(il, 0) >= d.global_tile_from_local(jl, 1) > k}
                    got = [(r // nb, c // nb) for r, c, _ in hp + np_]
                    assert len(got) == len(set(got)) and set(got) == want
                    for r, c, dg in hp + np_:
                        assert dg == int(d.global_tile_from_local(r // nb, 0) == d.global_tile_from_local(c // nb, 1))


@pytest.mark.parametrize("dt", [np.float32, np.float64, np.complex64, np.complex128])
@pytest.mark.parametrize("grid", GRIDS)
@pytest.mark.parametrize("m,mb", REF_SIZES)
def test_schedule_replay_matches_reference_known_answer(pkg, L, dt, grid, m, mb):
    """The engine's schedule, replayed with oracle tile arithmetic over simulated ranks, reproduces the reference's
    analytical factor on the reference's distributed test cases (non-zero source rank included)."""
    P, Q = grid
    src = (max(0, P - 1), min(1, Q - 1))
    a, t = orc.cholesky_setters(m, "L", dt)
    got = sched_exec.run_simulated(pkg, L, a, mb, grid, src)
    tol = 4 * (m + 1) * orc.type_error(dt)
    assert orc.check_near(got, t, tol, tol) == 0


def _gloo_worker(rank, world, grid, port, n, nb, src, q):
    import sys
    import torch.distributed as dist
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    from conftest import load_package
    pkg = load_package()
    L = pkg.lib()
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    P, Q = grid
    pr, pc = rank // Q, rank % Q
    row_ranks = [pr * Q + c for c in range(Q)]
    col_ranks = [r * Q + pc for r in range(P)]
    row_group = col_group = None
    for r_ in range(P):                                  # every rank must create every group, in the same order
        g = dist.new_group([r_ * Q + c for c in range(Q)])
        if r_ == pr:
            row_group = g
    for c_ in range(Q):
        g = dist.new_group([r * Q + c_ for r in range(P)])
        if c_ == pc:
            col_group = g
    a, t = orc.cholesky_setters(n, "L", np.float64)
    res = sched_exec.run_process_group(pkg, L, a, nb, grid, src, rank, dist, row_group, col_group, row_ranks,
                                       col_ranks)
    tol = 4 * (n + 1) * orc.type_error(np.float64)
    bad = orc.check_near(res.a, t[np.ix_(res.rows, res.cols)], tol, tol)
    q.put((rank, bad))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("grid,src", [((1, 2), (0, 1)), ((2, 1), (1, 0))])
def test_world_size_2_gloo(grid, src):
    """N > 1 path over a real process group on the CPU: two ranks, broadcasts through gloo."""
    import socket
    import torch.multiprocessing as mp
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_gloo_worker, args=(r, 2, grid, port, 34, 5, src, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(timeout=60)
    assert res == [(0, 0), (1, 0)]
