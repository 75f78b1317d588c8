"""Index algebra behind the ScaLAPACK sub-matrix arguments of dlaf_p?potrf / DLAF_descriptor.i, .j (SURVEY.md §8(f)
rank 4; the reference stops at ia = ja = 1, src/c_api/utils.cpp:25-33). dla-future_b200/src/c_api.cu treats a
tile-aligned sub-matrix of a 2D block-cyclic matrix as a 2D block-cyclic matrix of its own: source rank advanced by
the tiles skipped, local block starting after the local tiles that precede the offset. Checked here against the
oracle's restatement of the reference Distribution (explicit owner lists), through the library's exported index
functions (host logic only, no GPU)."""
import numpy as np
import pytest

import dlaf_oracle as orc


@pytest.mark.parametrize("grid", [(1, 1), (2, 1), (1, 3), (2, 3), (3, 2), (2, 4), (4, 2)])
def test_submatrix_is_block_cyclic_with_shifted_source(L, grid):
    P, Q = grid
    rng = np.random.default_rng(P * 10 + Q)
    for m, nb in [(23, 4), (40, 5), (17, 17), (64, 8)]:
        parent = rng.standard_normal((m, m))
        nt = -(-m // nb)
        for src in {(0, 0), (P - 1, min(1, Q - 1)), (P // 2, Q - 1)}:
            for i0 in range(nt):
                for j0 in {0, i0, nt - 1}:
                    n_sub = min(m - i0 * nb, m - j0 * nb)
                    for n in {n_sub, max(1, n_sub - 3)}:
                        sub = parent[i0 * nb:i0 * nb + n, j0 * nb:j0 * nb + n]
                        sub_src = ((src[0] + i0) % P, (src[1] + j0) % Q)
                        for pr in range(P):
                            for pc in range(Q):
                                whole, _, _ = orc.local_block(parent, nb, nb, grid, (pr, pc), src)
                                want, _, _ = orc.local_block(sub, nb, nb, grid, (pr, pc), sub_src)
                                r0 = L.dlafb_dist_next_local_tile_from_global(i0, P, pr, src[0]) * nb
                                c0 = L.dlafb_dist_next_local_tile_from_global(j0, Q, pc, src[1]) * nb
                                got = whole[r0:r0 + want.shape[0], c0:c0 + want.shape[1]]
                                assert got.shape == want.shape, (m, nb, src, i0, j0, n, pr, pc)
                                assert np.array_equal(got, want), (m, nb, src, i0, j0, n, pr, pc)
                                # and the library's own local size of the sub-matrix agrees
                                assert L.dlafb_dist_local_size(n, nb, P, pr, sub_src[0]) == want.shape[0]
                                assert L.dlafb_dist_local_size(n, nb, Q, pc, sub_src[1]) == want.shape[1]
