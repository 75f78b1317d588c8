"""Extracts the known-answer tables of the reference's own index-algebra unit tests into JSON fixtures (run where
/root/reference exists; the fixtures travel, the reference does not):

  test/unit/matrix/test_util_distribution.cpp:43-62   `tests` of UtilDistribution.IndexConversion
  test/unit/matrix/test_distribution.cpp:40-104       `tests_constructor` of DistributionTest.*
  test/unit/matrix/test_layout_info.cpp:23-34, :91-99 `values` of LayoutInfoTest.Constructor, `comp_values` of .ComparisonOperator

Only the numbers are taken (test DATA, the golden vectors of SURVEY.md §8c item 4), no code.
    python tests/golden/extract_reference_tables.py [/root/reference]
"""
import json
import os
import re
import sys

HERE = os.path.dirname(os.path.abspath(__file__))


def table(text, start_marker):
    """All the `{...}` rows (possibly nested one level) of the initializer list that follows start_marker."""
    text = re.sub(r"//[^\n]*", "", text)  # the tables carry commented-out header rows with braces
    i = text.index(start_marker)
    i = text.index("{", i)
    depth, rows, row_start = 0, [], None
    for j in range(i, len(text)):
        c = text[j]
        if c == "{":
            depth += 1
            if depth == 2:
                row_start = j
        elif c == "}":
            if depth == 2:
                rows.append(text[row_start:j + 1])
            depth -= 1
            if depth == 0:
                break
    return [[int(v) for v in re.findall(r"-?\d+", r)] for r in rows]


def main():
    ref = sys.argv[1] if len(sys.argv) > 1 else "/root/reference"
    src = open(os.path.join(ref, "test/unit/matrix/test_util_distribution.cpp")).read()
    rows = table(src, "std::vector<Parameters> tests")
    keys = ["tile_size", "tiles_per_block", "rank", "grid_size", "src_rank", "global_element", "global_tile", "rank_tile",
            "local_tile", "local_tile_next", "tile_element", "tile_offset", "tile_element_offset"]
    assert all(len(r) == len(keys) for r in rows) and len(rows) >= 30
    with open(os.path.join(HERE, "util_distribution_index_conversion.json"), "w") as f:
        json.dump({"source": "test/unit/matrix/test_util_distribution.cpp:43-62 (DLA-Future v0.6.0)", "keys": keys,
                   "rows": rows}, f, indent=0)
    src = open(os.path.join(ref, "test/unit/matrix/test_distribution.cpp")).read()
    rows = table(src, "std::vector<ParametersConstructor> tests_constructor")
    keys = ["size", "block_size", "tile_size", "rank", "grid_size", "src_rank", "offset", "global_tiles", "local_tiles",
            "local_size"]
    assert all(len(r) == 2 * len(keys) for r in rows) and len(rows) >= 50
    rows = [[r[2 * i:2 * i + 2] for i in range(len(keys))] for r in rows]
    with open(os.path.join(HERE, "distribution_constructor.json"), "w") as f:
        json.dump({"source": "test/unit/matrix/test_distribution.cpp:40-104 (DLA-Future v0.6.0)", "keys": keys,
                   "rows": rows}, f, indent=0)
    src = open(os.path.join(ref, "test/unit/matrix/test_layout_info.cpp")).read()
    rows = table(src, "\n    values(")
    assert all(len(r) == 8 for r in rows) and len(rows) >= 11
    comp = table(src.replace("true", "1").replace("false", "0"), "comp_values(")
    assert all(len(r) == 8 for r in comp) and len(comp) >= 6
    with open(os.path.join(HERE, "layout_info.json"), "w") as f:
        json.dump({"source": "test/unit/matrix/test_layout_info.cpp:23-34, :91-99 (DLA-Future v0.6.0)",
                   "keys": ["rows", "cols", "block_rows", "block_cols", "ld", "row_offset", "col_offset", "min_memory"],
                   "rows": rows,
                   "comparison_keys": ["rows", "cols", "block_rows", "block_cols", "ld", "row_offset", "col_offset",
                                       "equal_to_25_25_5_5_50_8_1000"],
                   "comparison_rows": comp}, f, indent=0)
    print("wrote 3 fixtures")


if __name__ == "__main__":
    main()
