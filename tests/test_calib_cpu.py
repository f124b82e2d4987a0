"""The host part of the pose refinement: pcp_se3_bfgs = ceres' line-search BFGS over Sophus::Manifold<SE3>, driven here by
analytic host costs (no GPU).  Ceres / Sophus are not in the reference tree, so there is nothing to pin the iterates to: these tests check
the optimiser against known minima and against scipy's BFGS on the same function."""
import numpy as np
import pytest


@pytest.fixture(scope="module")
def api(built_lib):
    from pointcloudprocessor_b200 import api as a
    return a


def _rot(q):
    """Eigen::Quaterniond::toRotationMatrix on {x,y,z,w} (no normalisation), the map Sophus::SE3 applies to its ambient parameters."""
    x, y, z, w = q
    tx, ty, tz = 2 * x, 2 * y, 2 * z
    return np.array([[1 - (ty * y + tz * z), ty * x - tz * w, tz * x + ty * w],
                     [ty * x + tz * w, 1 - (tx * x + tz * z), tz * y - tx * w],
                     [tz * x - ty * w, tz * y + tx * w, 1 - (tx * x + ty * y)]])


def _alignment_cost(P, Q):
    def value(x):
        return float(np.sum((P @ _rot(x[:4]).T + x[4:] - Q) ** 2))

    def fn(x):
        g = np.empty(7)
        for i in range(7):
            h = np.zeros(7)
            h[i] = 1e-6
            g[i] = (value(x + h) - value(x - h)) / 2e-6
        return value(x), g, True
    return value, fn


def _target():
    from scipy.spatial.transform import Rotation
    rng = np.random.default_rng(5)
    P = rng.normal(0, 1, (40, 3))
    R = Rotation.from_rotvec([0.05, -0.08, 0.03])
    t = np.array([0.1, -0.05, 0.02])
    return P, R.apply(P) + t, R, t


def test_bfgs_recovers_a_rigid_alignment(api):
    from scipy.spatial.transform import Rotation
    P, Q, R, t = _target()
    value, fn = _alignment_cost(P, Q)
    x, s = api.se3_bfgs(fn, [0, 0, 0, 1, 0, 0, 0], api.bfgs_options(function_tolerance=1e-14))
    assert s["termination"] == 0, s
    assert s["initial_cost"] == pytest.approx(value(np.array([0, 0, 0, 1, 0, 0, 0.0])))
    assert s["final_cost"] < 1e-9 and s["final_cost"] == pytest.approx(value(x), abs=1e-15)
    assert abs(np.linalg.norm(x[:4]) - 1) < 1e-12  # the iterates stay on the manifold
    assert np.allclose(x[4:], t, atol=1e-5)
    assert (Rotation.from_quat(x[:4]) * R.inv()).magnitude() < 1e-5
    assert 3 <= s["iterations"] <= 50 and s["num_cost_evaluations"] >= s["iterations"]


def test_bfgs_default_tolerances_match_scipy_minimum(api):
    """Same function through scipy's BFGS over a local 6-vector (rotation vector + translation): same minimum value to the tolerance."""
    from scipy.optimize import minimize
    from scipy.spatial.transform import Rotation
    P, Q, _, _ = _target()
    Q = Q + np.random.default_rng(6).normal(0, 0.01, Q.shape)  # noisy: the minimum is not zero
    value, fn = _alignment_cost(P, Q)
    x, s = api.se3_bfgs(fn, [0, 0, 0, 1, 0, 0, 0])
    ref = minimize(lambda v: float(np.sum((Rotation.from_rotvec(v[:3]).apply(P) + v[3:] - Q) ** 2)), np.zeros(6), method="BFGS", options={"gtol": 1e-9})
    assert s["termination"] == 0
    assert s["final_cost"] == pytest.approx(ref.fun, rel=1e-4)
    assert s["final_cost"] < s["initial_cost"]


def test_bfgs_respects_invalid_evaluations(api):
    """A functor that returns false outside a ball (the pose gate of MultiNIDCost): the search halves its step and stays inside."""
    P, Q, _, t = _target()
    value, fn = _alignment_cost(P, Q)
    calls = {"invalid": 0}

    def gated(x):
        if np.linalg.norm(x[4:]) > 0.06:
            calls["invalid"] += 1
            return 0.0, np.zeros(7), False
        return fn(x)
    x, s = api.se3_bfgs(gated, [0, 0, 0, 1, 0, 0, 0])
    assert calls["invalid"] > 0
    assert np.linalg.norm(x[4:]) <= 0.06
    assert s["final_cost"] < 0.5 * s["initial_cost"]


def test_bfgs_failure_modes(api):
    x, s = api.se3_bfgs(lambda x: (0.0, np.zeros(7), False), [0, 0, 0, 1, 0, 0, 0])
    assert s["termination"] == 2 and "Initial" in s["message"] and np.array_equal(x, [0, 0, 0, 1, 0, 0, 0])
    x, s = api.se3_bfgs(lambda x: (1.0, np.zeros(7), True), [0, 0, 0, 1, 1, 2, 3])
    assert s["termination"] == 0 and s["iterations"] == 0 and np.array_equal(x, [0, 0, 0, 1, 1, 2, 3])
    x, s = api.se3_bfgs(_alignment_cost(*_target()[:2])[1], [0, 0, 0, 1, 0, 0, 0], api.bfgs_options(max_num_iterations=2, function_tolerance=0.0))
    assert s["termination"] == 1 and s["iterations"] == 2
