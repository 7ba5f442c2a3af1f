"""give_evaluate / give_evaluate_g (preintegration.hpp:137-294) as the device kernel the LI optimizers use (k_li_imu)
against the oracle's restatement: cost, J^T cov^-1 J and J^T cov^-1 r of a chain of factors.  FP64, <= 1e-10 relative (the device
contracts multiply-adds, the oracle does not)."""
import numpy as np
import pytest

import synth

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("gravity", [False, True])
def test_imu_chain_device_matches_oracle(pkg, ob, ctx, gravity):
    W = 12
    states, segs = synth.imu_trajectory(4242, W, gyr_sigma=1e-3, acc_sigma=1e-2)
    nz = pkg.imu_noise(0.01, 0.1, 1e-4, 1e-3)
    imus = np.concatenate([pkg.imu_pre_new(s, nz, states[i]["bg"], states[i]["ba"]) for i, s in enumerate(segs)])
    rng = np.random.Generator(np.random.PCG64(7))
    x = states.copy()
    x["p"] += rng.normal(0, 0.02, (W, 3)); x["v"] += rng.normal(0, 0.02, (W, 3)); x["bg"] += rng.normal(0, 1e-3, (W, 3))
    imus["dbg"] = rng.normal(0, 1e-3, (W - 1, 3)); imus["dba"] = rng.normal(0, 1e-2, (W - 1, 3))
    cost, jtj, gg = pkg.imu_pre_evaluate_device(ctx, imus, x, with_gravity=gravity)
    cost_only, _, _ = pkg.imu_pre_evaluate_device(ctx, imus, x, with_gravity=gravity, jac=False)
    assert np.abs(cost - cost_only).max() <= 1e-12 * np.abs(cost).max()      # two instantiations, contracted differently
    for i in range(W - 1):
        c, h, g = ob.imu_pre_evaluate(imus[i:i + 1], x[i:i + 1], x[i + 1:i + 2], with_gravity=gravity)
        assert abs(cost[i] - c) <= 1e-10 * abs(c)
        assert np.abs(jtj[i] - h).max() <= 1e-10 * np.abs(h).max()
        assert np.abs(gg[i] - g).max() <= 1e-10 * np.abs(g).max()
        assert np.abs(h).max() > 1.0
