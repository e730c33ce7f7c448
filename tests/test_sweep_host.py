"""Host logic of the parameter sweep (tools/sweep.py, the job of the reference's autotune/*/tune.cpp): the base-case rule it
turns bc_mult_dim into is the reference's (cholinv.hpp:15-18), checked against the oracle's restatement and against the bcdim
the unmodified reference printed when the cholinv fixtures were generated."""
import os
import re
import sys

import pytest

from conftest import ROOT, golden

sys.path.insert(0, os.path.join(ROOT, "tools"))
import capital_oracle as orc  # noqa: E402


@pytest.mark.parametrize("c,d", [(1, 1), (2, 1), (1, 2), (2, 2), (1, 4), (4, 4)])
@pytest.mark.parametrize("n", [64, 1000, 8192, 32768])
@pytest.mark.parametrize("bc", [3, 1, 0, -1, -2, -3, -4, -9])
def test_bc_global_matches_reference_rule(c, d, n, bc):
    import sweep
    l = -(-n // d)
    assert sweep.bc_global_of(l, c, d, bc) == orc.cholinv_bc_dimension(n, c, d, bc)


@pytest.mark.parametrize("tag", ["p1_ci1", "p1_ci0", "p8_ci1", "p8_ci0"])
def test_bc_global_matches_what_the_reference_printed(tag):
    import sweep
    q = golden("cholinv")
    ranks, n, ci, bc, d = [int(v) for v in q[tag + "_meta"]]
    c = 1 if ranks == 1 else 2
    m = re.search(r"bcdim=(\d+)", str(q[tag + "_line"]))
    assert m, str(q[tag + "_line"])
    # the reference stores the LOCAL base-case size it stops at (cholinv.hpp:16-18); ours is the global size d * l / bc
    l = n // d
    assert sweep.bc_global_of(l, c, d, bc) in (int(m.group(1)), int(m.group(1)) * d)
