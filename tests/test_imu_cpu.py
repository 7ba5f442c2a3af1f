"""B6 IMU_PRE (preintegration.hpp:11-303): the host-side factor of the product library against the oracle, plus
self-consistency KATs of the oracle itself (the reference ships no vectors for it: parity unpinned).  Needs no GPU —
these entry points take no vx_ctx."""
import numpy as np
import pytest

import synth


@pytest.fixture(scope="module")
def traj():
    return synth.imu_trajectory(31, 6, gyr_sigma=1e-3, acc_sigma=1e-2)


def _noise(pkg):
    return pkg.imu_noise(0.01, 0.1, 1e-4, 1e-3)


def _rel(a, b):
    return np.abs(a - b).max() / max(np.abs(b).max(), 1e-300)


def test_preintegration_matches_oracle(pkg, ob, traj):
    states, segs = traj
    nz = _noise(pkg)
    for i, s in enumerate(segs):
        a = pkg.imu_pre_new(s, nz, states[i]["bg"], states[i]["ba"])
        b = ob.imu_pre_new(s, nz, states[i]["bg"], states[i]["ba"])
        for name in a.dtype.names:
            assert _rel(a[name], b[name]) < 1e-13, name
        assert abs(a["dtime"][0] - 0.1) < 1e-12


def test_evaluate_matches_oracle(pkg, ob, traj):
    states, segs = traj
    nz = _noise(pkg)
    rng = np.random.Generator(np.random.PCG64(5))
    for i, s in enumerate(segs):
        pre = pkg.imu_pre_new(s, nz, states[i]["bg"], states[i]["ba"])
        pre["dbg"] = rng.normal(0, 1e-3, 3); pre["dba"] = rng.normal(0, 1e-2, 3)
        s1 = states[i:i + 1].copy(); s2 = states[i + 1:i + 2].copy()
        s2["p"] += rng.normal(0, 0.01, 3); s2["v"] += rng.normal(0, 0.01, 3)
        s2["R"] = (synth.rodrigues(rng.normal(0, 0.01, 3)).T @ s2["R"].reshape(3, 3)).reshape(9)   # col-major: (R E)^T = E^T R^T
        for grav in (False, True):
            c1, h1, g1 = pkg.imu_pre_evaluate(pre, s1, s2, grav)
            c2, h2, g2 = ob.imu_pre_evaluate(pre, s1, s2, grav)
            assert abs(c1 - c2) <= 1e-12 * abs(c2)
            assert _rel(h1, h2) < 1e-12 and _rel(g1, g2) < 1e-12
            c3, _, _ = pkg.imu_pre_evaluate(pre, s1, s2, grav, jac=False)
            assert c3 == c1


def test_noise_free_preintegration_is_consistent(ob, pkg):
    # KAT: exact IMU of an analytic trajectory => deltas equal the true relative motion up to the integrator's
    # first-order discretisation error (R_delta of the sub-step start, preintegration.hpp:132-134), which must shrink with rate
    nz = _noise(pkg)
    errs = []
    for rate in (200, 2000):
        states, segs = synth.imu_trajectory(32, 5, rate=rate)
        worst = 0.0
        for i, s in enumerate(segs):
            pre = ob.imu_pre_new(s, nz, states[i]["bg"], states[i]["ba"])
            R1 = states[i]["R"].reshape(3, 3).T; R2 = states[i + 1]["R"].reshape(3, 3).T
            dt = pre["dtime"][0]; g = states[i]["g"]
            dv = R1.T @ (states[i + 1]["v"] - states[i]["v"] - g * dt)
            dp = R1.T @ (states[i + 1]["p"] - states[i]["p"] - states[i]["v"] * dt - 0.5 * g * dt * dt)
            worst = max(worst, np.abs(pre["R_delta"][0].reshape(3, 3).T - R1.T @ R2).max(), np.abs(pre["v_delta"][0] - dv).max(),
                        np.abs(pre["p_delta"][0] - dp).max())
        errs.append(worst)
    assert errs[0] < 1e-3 and errs[1] < 0.2 * errs[0]


def test_gradient_is_half_derivative_of_cost(ob, pkg):
    # KAT: gg = J^T cov^-1 r  ==  0.5 d(cost)/d(xi) for the additive coordinates (p, v, bg, ba of state 2 and g)
    states, segs = synth.imu_trajectory(33, 3, gyr_sigma=1e-3, acc_sigma=1e-2)
    nz = _noise(pkg)
    pre = ob.imu_pre_new(segs[0], nz, states[0]["bg"], states[0]["ba"])
    s1 = states[0:1].copy(); s2 = states[1:2].copy()
    s2["p"] += 0.01; s2["v"] -= 0.02
    c0, H, g = ob.imu_pre_evaluate(pre, s1, s2, True)
    eps = 1e-6
    for name, col in (("p", 15 + 3), ("v", 15 + 6), ("bg", 15 + 9), ("ba", 15 + 12)):
        for k in range(3):
            sp = s2.copy(); sm = s2.copy()
            sp[name][0, k] += eps; sm[name][0, k] -= eps
            cp, _, _ = ob.imu_pre_evaluate(pre, s1, sp, True, jac=False)
            cm, _, _ = ob.imu_pre_evaluate(pre, s1, sm, True, jac=False)
            fd = 0.5 * (cp - cm) / (2 * eps)
            assert abs(fd - g[col + k]) <= 1e-5 * max(abs(g[col + k]), np.abs(g).max() * 1e-3), (name, k)
    for k in range(3):   # gravity column (33-wide form)
        sp = s1.copy(); sm = s1.copy()
        sp["g"][0, k] += eps; sm["g"][0, k] -= eps
        cp, _, _ = ob.imu_pre_evaluate(pre, sp, s2, True, jac=False)
        cm, _, _ = ob.imu_pre_evaluate(pre, sm, s2, True, jac=False)
        fd = 0.5 * (cp - cm) / (2 * eps)
        assert abs(fd - g[30 + k]) <= 1e-5 * max(abs(g[30 + k]), np.abs(g).max() * 1e-3)
    assert np.abs(H - H.T).max() <= 1e-9 * np.abs(H).max()
