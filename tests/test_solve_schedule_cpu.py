"""The stream / event schedule of the grid triangular solve (csrc/potrf_dist.cu: solve_panels_multi -- chain on the
compute stream, bulk updates on the update stream, panel prefetch and inverse blocks on the communication stream)
orders every pair of conflicting buffer accesses; tools/sim_solve_streams.py restates the issue order and checks
happens-before.  Leaving out any one of its waits must be reported: the check is not vacuous."""
import itertools
import os
import sys

import pytest

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools"))
from sim_solve_streams import solve_schedule  # noqa: E402

GRIDS = [(1, 1), (1, 2), (2, 1), (2, 2), (2, 4), (3, 2)]


@pytest.mark.parametrize("P,Q", GRIDS)
def test_solve_schedule_has_no_unordered_conflicts(P, Q):
    for mt, lower, cols in itertools.product([1, 2, 3, 7, 12], [True, False], [0, 1, 2]):
        for prow in range(P):
            for pcol in range(Q):
                r = solve_schedule(P, Q, prow, pcol, mt, lower, cols).races()
                assert not r, (P, Q, prow, pcol, mt, lower, cols, r[0])


def _races(drop, mt=9):
    out = []
    for (P, Q), lower in itertools.product([(2, 2), (1, 2), (2, 1)], [True, False]):
        for prow in range(P):
            for pcol in range(Q):
                out += solve_schedule(P, Q, prow, pcol, mt, lower, 2, drop=drop).races()
    return out


@pytest.mark.parametrize("drop", ["lookahead_after_bulk", "panel_free", "x_ready"])
def test_each_wait_is_needed(drop):
    assert _races({drop}), drop


def test_receive_buffer_wait_is_implied_by_the_panel_chain():
    """the wait before the solved-row buffer of a parity is rewritten (bulk update of step kk-2 done) is also implied:
    the compute stream waits for the panel of step kk, whose broadcast waited for the release of the panel set, which
    the update stream records after the bulk update of step kk-2.  Without BOTH the buffer is overwritten under a reader."""
    assert not _races({"xrow_reuse"})
    both = _races({"xrow_reuse", "panel_free"})
    assert any(any(b.startswith("xrow") for b in bufs) for _, _, bufs in both)
