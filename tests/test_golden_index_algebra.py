"""Replays the reference's own known-answer tables for the 2D block-cyclic index algebra against
the oracle (oracle/dlaf_oracle.py: Distribution) and against the product's exported index functions (dlafb_dist_*,
dla-future_b200/src/schedule.h: Dist1D - the code the engine, the solvers and the C API's sub-matrix arguments use).

Fixtures: tests/golden/*.json, extracted from the reference's test sources by tests/golden/extract_reference_tables.py
  util_distribution_index_conversion.json   test/unit/matrix/test_util_distribution.cpp:43-62
  distribution_constructor.json             test/unit/matrix/test_distribution.cpp:40-104
Rows outside what this path supports are counted and skipped: more than one tile per block (the factorization's
precondition is single_tile_per_block, factorization/cholesky.h:43,71) and offsets that are not tile aligned (§13 of
DESIGN.md: tile-aligned offsets are supported as a shifted source rank, which is how they are replayed here).
"""
import json
import os

import dlaf_oracle as orc

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def load(name):
    with open(os.path.join(GOLDEN, name)) as f:
        d = json.load(f)
    return [dict(zip(d["keys"], r)) for r in d["rows"]]


def test_index_conversion_table(L):
    rows = load("util_distribution_index_conversion.json")
    used = 0
    for t in rows:
        if t["tiles_per_block"] != 1 or t["tile_offset"] != 0 or t["tile_element_offset"] != 0:
            continue
        used += 1
        nb, P, rank, src = t["tile_size"], t["grid_size"], t["rank"], t["src_rank"]
        g = t["global_tile"]
        # element <-> tile
        assert t["global_element"] // nb == g and t["global_element"] % nb == t["tile_element"]
        # product
        assert L.dlafb_dist_rank_global_tile(g, P, src) == t["rank_tile"]
        assert L.dlafb_dist_next_local_tile_from_global(g, P, rank, src) == t["local_tile_next"]
        if t["local_tile"] >= 0:
            assert t["rank_tile"] == rank
            assert L.dlafb_dist_local_tile_from_global(g, P) == t["local_tile"]
            assert L.dlafb_dist_global_tile_from_local(t["local_tile"], P, rank, src) == g
        else:
            assert t["rank_tile"] != rank
        # oracle (axis 0 of a P x 1 grid)
        d = orc.Distribution((g + 1) * nb, nb, grid=(P, 1), rank=(rank, 0), src=(src, 0))
        assert d.rank_global_tile(g, 0) == t["rank_tile"]
        assert d.next_local_tile_from_global(g, 0) == t["local_tile_next"]
        if t["local_tile"] >= 0:
            assert d.local_tile_from_global(g, 0) == t["local_tile"]
            assert d.global_tile_from_local(t["local_tile"], 0) == g
    assert used >= 10, used


def test_distribution_constructor_table(L):
    rows = load("distribution_constructor.json")
    used = 0
    for t in rows:
        if t["block_size"] != t["tile_size"]:
            continue
        if t["offset"][0] % t["block_size"][0] or t["offset"][1] % t["block_size"][1]:
            continue
        used += 1
        for ax in (0, 1):
            n, nb = t["size"][ax], t["block_size"][ax]
            P, rank = t["grid_size"][ax], t["rank"][ax]
            src = (t["src_rank"][ax] + t["offset"][ax] // nb) % P   # the rule of check_constructor (:106-114)
            assert L.dlafb_dist_nr_tiles(n, nb) == t["global_tiles"][ax], t
            assert L.dlafb_dist_local_nr_tiles(n, nb, P, rank, src) == t["local_tiles"][ax], t
            assert L.dlafb_dist_local_size(n, nb, P, rank, src) == t["local_size"][ax], t
            if n > 0:
                grid = (P, 1) if ax == 0 else (1, P)
                rk = (rank, 0) if ax == 0 else (0, rank)
                sr = (src, 0) if ax == 0 else (0, src)
                d = orc.Distribution(n, nb, grid=grid, rank=rk, src=sr)
                assert d.nt == t["global_tiles"][ax]
                assert d.local_nr_tiles(ax) == t["local_tiles"][ax], t
                assert d.local_size(ax) == t["local_size"][ax], t
    assert used >= 16, used
