"""Pose refinement on the device: pcp_calibrate = VisualCameraCalibration::calibrate over the device NID costs
(src/vlcal/calib/visual_camera_calibration.cpp:49-252), and PointCloudProcessor::process() with enableNIDOptimize."""
import numpy as np
import pytest

import calib_scene as cs

pytestmark = pytest.mark.gpu


def _se3(T):
    from scipy.spatial.transform import Rotation
    return np.concatenate([Rotation.from_matrix(T[:3, :3]).as_quat(), T[:3, 3]])


def _mat(x):
    from scipy.spatial.transform import Rotation
    T = np.eye(4)
    T[:3, :3] = Rotation.from_quat(x[:4]).as_matrix()
    T[:3, 3] = x[4:]
    return T


def test_calibrate_one_solve_against_the_oracle_driven_solver(oracle, built_lib):
    """One BFGS solve with the device cost against the same solver driven by the oracle's Jet7 cost: the two costs agree to 1e-5 relative
    (test_nid_gradient), so both runs must end in the same minimum."""
    from pointcloudprocessor_b200 import api
    frames = cs.make(20_000, 2)
    cam_o = oracle.camera(cs.INTR, cs.DIST)
    T0 = cs.perturbation()
    x0 = _se3(T0)

    def fn(x):
        if not oracle.nid_gate(x0, x):
            return 0.0, np.zeros(7), False
        v, g, good = 0.0, np.zeros(7), True
        for pts, img in frames:
            ok, vv, gg = oracle.nid_grad(cam_o, img, pts, x)
            v, g, good = v + vv, g + gg, good and ok
        return v, g, good
    x_ref, s_ref = api.se3_bfgs(fn, x0)
    with api.Colorizer(api.create_camera("pinhole", cs.INTR, cs.DIST), api.default_params(image_size=(cs.W, cs.H))) as gc:
        costs = [api.NIDCost(gc, img, pts, 16) for pts, img in frames]
        n0 = gc.launch_count()
        T, s = api.calibrate(costs, T0, max_outer_iterations=1)
        assert gc.launch_count() > n0  # the evaluations ran on the device
        v_dev, _, ok = api.multi_nid_with_gradient(costs, x0, _se3(T))
        for c in costs:
            c.close()
    assert s["outer_iterations"] == 1 and s["last_inner_termination"] == 0
    assert s["initial_cost"] == pytest.approx(s_ref["initial_cost"], rel=1e-5)
    assert s["final_cost"] == pytest.approx(s_ref["final_cost"], rel=2e-4)
    assert ok[0] and v_dev[0] == pytest.approx(s["final_cost"], rel=1e-9)
    assert s["final_cost"] < s["initial_cost"] - 0.05
    dt, dr = cs.pose_error(np.linalg.inv(_mat(x_ref)) @ T)
    assert dt < 1.5e-3 and dr < 6e-4, (dt, dr)


def test_calibrate_outer_loop_pulls_the_extrinsic_back(built_lib):
    from pointcloudprocessor_b200 import api
    frames = cs.make(60_000, 3)
    T0 = cs.perturbation()
    t0, r0 = cs.pose_error(T0)
    with api.Colorizer(api.create_camera("pinhole", cs.INTR, cs.DIST), api.default_params(image_size=(cs.W, cs.H))) as gc:
        costs = [api.NIDCost(gc, img, pts, 16) for pts, img in frames]
        T, s = api.calibrate(costs, T0, delta_trans_thresh=1e-3, delta_rot_thresh=np.deg2rad(0.05))
        T_ref, s_ref = api.calibrate(costs, T0)  # reference thresholds (1 cm, 1 degree): the 8 mm start converges after one solve
        at_truth, _, _ = api.multi_nid_with_gradient(costs, _se3(np.eye(4)), _se3(np.eye(4)))
        for c in costs:
            c.close()
    assert s_ref["converged"] == 1 and s_ref["outer_iterations"] == 1 and s_ref["delta_t"] < 0.01 and s_ref["delta_r"] < np.deg2rad(1.0)
    assert s["converged"] == 1 and 2 <= s["outer_iterations"] <= 10  # the last solve moves less than 1 mm / 0.05 degrees
    assert s["delta_t"] < 1e-3 and s["delta_r"] < np.deg2rad(0.05)
    assert s["final_cost"] <= s_ref["final_cost"] + 1e-9
    assert s["final_cost"] < s["initial_cost"] - 0.1
    assert s["final_cost"] <= at_truth[0] + 1e-3
    t1, r1 = cs.pose_error(T)
    assert t1 < 0.5 * t0 and r1 < 0.5 * r0, (t0, r0, t1, r1)
    assert abs(np.linalg.det(T[:3, :3]) - 1) < 1e-12 and np.allclose(T[3], [0, 0, 0, 1])


def test_process_with_nid_optimisation(built_lib, tmp_path):
    """enableNIDOptimize: applyNIDBasedPoseOptimization between the cull pass and the colour pass (cpp:1019-1022).  The extrinsic it
    finds equals pcp_calibrate over NID costs rebuilt here from the files process() wrote (<ts>_beforeNID.pcd, read back as the reference
    does) and the image planes as NIDCost sees them; the colour pass then equals a run that is handed that extrinsic."""
    from test_process_gpu import _dataset
    from pointcloudprocessor_b200 import api
    W, H = 330, 240
    sc, frames, img_dir, mask_dir, out_dir = _dataset(tmp_path, n=20_000, f=3, W=W, H=H, masks=False)
    cam = api.create_camera("pinhole", sc["intr"], sc["dist"])
    proc = api.PointCloudProcessor(tmp_path / "scans.pcd", tmp_path / "odom.txt", str(img_dir) + "/", "", str(out_dir) + "/", enableNIDOptimize=True,
                                   camera=cam, cull_size=(W, H))
    st = proc.process()
    assert st["nid_outer_iterations"] >= 1 and st["nid_cost_evaluations"] >= 1 and st["seconds_nid"] > 0
    assert st["nid_final_cost"] <= st["nid_initial_cost"]
    T = np.array(list(st["T_camera_lidar"])).reshape(4, 4).T
    assert abs(np.linalg.det(T[:3, :3]) - 1) < 1e-9 and np.allclose(T[3], [0, 0, 0, 1])
    coloured = (out_dir / "cloudInWorldWithRGB.pcd").read_bytes()

    with api.Colorizer(cam, api.default_params(image_size=(W, H))) as gc:
        costs = []
        for fr in frames:
            pts = api.load_pcd_xyzi(out_dir / "filtered_pcd" / (fr["name"] + "_beforeNID.pcd"))
            # CV_64FC3 read through at<double>(y, x): the first W interleaved B,G,R values of every row (visual_camera_calibration.cpp:171-173)
            plane = fr["raw"].reshape(H, W * 3)[:, :W].astype(np.float64) * (1.0 / 255.0)
            costs.append(api.NIDCost(gc, plane, pts, 16))
        T_py, s = api.calibrate(costs)
        ident = np.array([0, 0, 0, 1, 0, 0, 0.0])
        v0, _, _ = api.multi_nid_with_gradient(costs, ident, ident)
        for c in costs:
            c.close()
    assert st["nid_initial_cost"] == pytest.approx(v0[0], rel=1e-9)  # same clouds, same planes, same start
    assert st["nid_final_cost"] == pytest.approx(s["final_cost"], rel=1e-6)
    assert np.allclose(T, T_py, atol=1e-6)

    out2 = tmp_path / "out2"
    out2.mkdir()
    proc2 = api.PointCloudProcessor(tmp_path / "scans.pcd", tmp_path / "odom.txt", str(img_dir) + "/", "", str(out2) + "/", camera=cam, cull_size=(W, H),
                                    T_camera_lidar_optimized=T)
    proc2.process()
    assert (out2 / "cloudInWorldWithRGB.pcd").read_bytes() == coloured
