"""Property tests (hypothesis) of the HOT PATH's host logic: the factorization engine's per-step schedule
(dlafb_schedule_step: the code the GPU scheduler runs, dla-future_b200/src/schedule.h) for random grids, source ranks,
sizes and tile sizes, against the oracle's restatement of the reference (owner maps util_distribution.h:85-175, panel
routing broadcast_panel.h:123-213, loop bounds cholesky/impl.h:191-313) - and the schedule replayed with oracle tile
arithmetic over simulated ranks must factor a random HPD matrix."""
import os

import numpy as np
from hypothesis import HealthCheck, given, settings, strategies as st

import dlaf_oracle as orc
import sched_exec

N_EXAMPLES = int(os.environ.get("DLAFB_PROPERTY_EXAMPLES", "40"))
COMMON = dict(deadline=None, max_examples=N_EXAMPLES,
              suppress_health_check=[HealthCheck.function_scoped_fixture, HealthCheck.too_slow])
grids = st.tuples(st.integers(1, 4), st.integers(1, 4)).filter(lambda g: g[0] * g[1] <= 8)


@settings(**COMMON)
@given(grid=grids, n=st.integers(1, 60), nb=st.integers(1, 12), seed=st.integers(0, 10 ** 6))
def test_schedule_matches_oracle_routing(pkg, L, grid, n, nb, seed):
    P, Q = grid
    rng = np.random.default_rng(seed)
    src = (int(rng.integers(0, P)), int(rng.integers(0, Q)))
    for pr in range(P):
        for pc in range(Q):
            d = orc.Distribution(n, nb, grid, (pr, pc), src)
            seen = set()
            for k in range(d.nt):
                stp, colb, hp, np_ = pkg.schedule_step(L, n, nb, P, Q, pr, pc, src[0], src[1], k)
                assert stp.own_diag == int(d.rank_global_tile(k, 0) == pr and d.rank_global_tile(k, 1) == pc)
                assert stp.in_col == int(d.rank_global_tile(k, 1) == pc)
                assert stp.diag_root == d.rank_global_tile(k, 0) and stp.row_root == d.rank_global_tile(k, 1)
                assert stp.il0 == d.next_local_tile_from_global(k + 1, 0) and stp.jl0 == d.next_local_tile_from_global(k + 1, 1)
                if k == d.nt - 1:
                    assert not colb and not hp and not np_
                    continue
                step1, step2 = orc.panel_broadcast_plan(d, k, "col")
                assert [d.global_tile_from_local(il, 0) for il in range(stp.il0, stp.il_end)] == [g for g, _ in step1]
                assert [(g, root) for _, g, root, _ in colb] == step2
                want = {(il, jl) for il in range(d.local_nr_tiles(0)) for jl in range(d.local_nr_tiles(1))
                        if d.global_tile_from_local(il, 0) >= d.global_tile_from_local(jl, 1) > k}
                got = [(r // nb, c // nb) for r, c, _ in hp + np_]
                assert len(got) == len(set(got)) and set(got) == want
                # the look-ahead list is exactly global column k + 1
                assert {c // nb for _, c, _ in hp} <= {d.local_tile_from_global(k + 1, 1)}
                assert all(d.global_tile_from_local(c // nb, 1) != k + 1 for _, c, _ in np_)
                seen |= {(il, jl, k) for il, jl in got}
            # every local tile (gi >= gj) is updated once by every step k < gj and never otherwise
            for il in range(d.local_nr_tiles(0)):
                for jl in range(d.local_nr_tiles(1)):
                    gi, gj = d.global_tile_from_local(il, 0), d.global_tile_from_local(jl, 1)
                    ks = {k for (a, b, k) in seen if (a, b) == (il, jl)}
                    assert ks == (set(range(gj)) if gi >= gj else set())


@settings(**{**COMMON, "max_examples": max(1, N_EXAMPLES // 2)})
@given(grid=grids, n=st.integers(1, 48), nb=st.integers(1, 10), seed=st.integers(0, 10 ** 6))
def test_schedule_replay_factors_a_random_hpd_matrix(pkg, L, grid, n, nb, seed):
    P, Q = grid
    rng = np.random.default_rng(seed)
    src = (int(rng.integers(0, P)), int(rng.integers(0, Q)))
    g = rng.standard_normal((n, n))
    a = np.asfortranarray(g @ g.T + n * np.eye(n))
    got = sched_exec.run_simulated(pkg, L, a, nb, grid, src)
    want = np.linalg.cholesky(a)
    assert np.abs(np.tril(got) - want).max() <= 1e-10 * np.abs(want).max() * n
