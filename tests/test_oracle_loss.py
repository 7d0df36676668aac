"""CPU: the numpy restatement of the reference's image losses (oracle/loss_ref.py) against the golden fixture that
tests/golden/make_golden_loss.py produced by importing the reference's utils/loss_utils.py."""
import os

import numpy as np
import pytest

from oracle import loss_ref

G = np.load(os.path.join(os.path.dirname(__file__), "golden", "loss_ssim.npz"))


@pytest.mark.parametrize("case", ["a", "b", "c"])
def test_loss_oracle_matches_reference(case):
    x, y = G[case + "_img"].astype(np.float64), G[case + "_gt"].astype(np.float64)
    assert abs(loss_ref.l1_loss(x, y) - G[case + "_l1"]) < 1e-6
    assert abs(loss_ref.ssim(x, y) - G[case + "_ssim"]) < 1e-5
    assert abs(loss_ref.training_loss(x, y, 0.2) - G[case + "_loss"]) < 1e-5
    gs = loss_ref.ssim_grad(x, y)
    ref = G[case + "_g_ssim"]
    assert np.abs(gs - ref).max() <= 1e-3 * np.abs(ref).max()          # analytic vs the reference's autograd
    gl = loss_ref.training_loss_grad(x, y, 0.2)
    ref = G[case + "_g_loss"]
    assert np.abs(gl - ref).max() <= 1e-3 * np.abs(ref).max()
    assert np.array_equal(np.sign(x - y) / x.size != 0, G[case + "_g_l1"] != 0)   # sign(0) = 0 region preserved
