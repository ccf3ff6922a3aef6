"""Parity comparators shared by the tests.

ids must match position by position, except where the two candidates' truth distances tie within
1e-5 relative; distances must match within 1e-4 relative (plus a tiny absolute floor for values
near zero, e.g. a vector searched against itself)."""
import numpy as np

ID_TIE_RTOL = 1e-5
DIST_RTOL = 1e-4
DIST_ATOL = 1e-6


def assert_topk_parity(got_labels, got_dist, want_labels, want_dist, truth_of=None, scale=None, what=""):
    """truth_of(label) -> fp64 truth distance of that label for this query (needed only when ids differ).
    scale: magnitude used for the relative tie test when distances are near zero (defaults to |want|)."""
    got_labels, want_labels = np.asarray(got_labels), np.asarray(want_labels)
    got_dist, want_dist = np.asarray(got_dist, np.float64), np.asarray(want_dist, np.float64)
    assert len(got_labels) == len(want_labels), f"{what}: count {len(got_labels)} != {len(want_labels)}"
    for i, (g, w) in enumerate(zip(got_labels, want_labels)):
        if g != w:
            assert truth_of is not None, f"{what}: id mismatch at rank {i}: {g} != {w}"
            tg, tw = truth_of(int(g)), truth_of(int(w))
            mag = max(abs(tw), abs(tg), scale or 0.0, 1e-30)
            assert abs(tg - tw) <= ID_TIE_RTOL * mag, (
                f"{what}: rank {i}: got id {g} (truth {tg!r}) but oracle has {w} ({tw!r}): not a tie")
    # a swapped pair still has to carry the right distances, rank by rank
    np.testing.assert_allclose(got_dist, want_dist, rtol=DIST_RTOL, atol=DIST_ATOL * max(scale or 1.0, 1.0),
                               err_msg=f"{what}: distances")
    # and the same label may not appear twice
    assert len(set(got_labels.tolist())) == len(got_labels), f"{what}: duplicate labels"
