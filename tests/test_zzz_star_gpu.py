"""The star net of test_bruteforce.star_case on the GPU: views that fold three variables into their pattern tables. The resident
kernel reads such a view's fold list from the op's descriptor (an item carries two pairs, csrc/resident.h); the flat, fused and
level kernels see 64 patterns per entry, the default cap. Every kernel against the brute-force enumeration of the joint.
Collected after the other files on purpose (file name): written when the round's GPU time was spent, so it ran on the CPU
emulators only (tests/test_bruteforce.py) before the round's end."""
import os

import numpy as np
import pytest

from fastbn_b200 import capi
from test_bruteforce import star_case

pytestmark = pytest.mark.gpu


def test_three_folded_variables_on_every_kernel(tmp_path):
    tree, ev_off, ev_vars, ev_vals, want = star_case(tmp_path, n_cases=500)
    ok = ~np.isnan(want)
    for kw in (dict(), dict(flat=True), dict(team_warps=4), dict(team_warps=8, mode=capi.MODE_LEVEL)):  # resident, flat, fused, level
        if kw.pop("flat", False):
            os.environ["FBN_NO_RESIDENT"] = "1"
        try:
            p = capi.Plan(tree, device=0, **kw)
        finally:
            os.environ.pop("FBN_NO_RESIDENT", None)
        labels, post, _ = p.run(ev_off, ev_vars, ev_vals)
        p.close()
        assert np.array_equal(np.isnan(post), np.isnan(want)), kw
        assert np.max(np.abs(post[ok] - want[ok])) <= 1e-12, kw
