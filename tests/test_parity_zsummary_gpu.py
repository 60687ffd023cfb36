"""Aggregate of the arbiter-decided delta comparisons of this run (collected after test_parity_gpu, test_fullsize_gpu and
test_edge_gpu): the device must not be SYSTEMATICALLY further from the exact solution than the reference."""
import numpy as np
import pytest

import conftest

pytestmark = pytest.mark.gpu


def test_device_is_not_systematically_further_from_the_exact_solution_than_the_reference():
    r = conftest.arbiter_ratios
    if len(r) < 3:
        pytest.skip("fewer than three arbiter-decided cases in this run (a filtered selection)")
    gm = float(np.exp(np.mean(np.log(list(r.values())))))
    conftest.record_parity(f"arbiter-decided cases: {len(r)}, geometric mean of |d_gpu-d*| / |d_ref-d*| = {gm:.2f} "
                           f"(min {min(r.values()):.2f}, max {max(r.values()):.2f}); bar {conftest.ARBITER_MEAN}")
    assert gm <= conftest.ARBITER_MEAN, r
