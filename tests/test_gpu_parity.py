"""Parity of the CUDA path (through the C ABI) against the CPU oracle on the same seeded inputs.
Bar: bit-exact visibility masks, pixel indices and integer accumulators; colours equal in SUM mode, <=1/255 in TOP5."""
import numpy as np
import pytest

from conftest import make_scene

pytestmark = pytest.mark.gpu


def _run_pair(oracle, sc, accum="sum", cull_size=None, T_opt=None, ds=14, tol=0.05, masks=False, bilinear=False, flight=0):
    from pointcloudprocessor_b200 import api
    cam_o = oracle.camera(sc["intr"], sc["dist"])
    cam_g = api.create_camera("pinhole", sc["intr"], sc["dist"])
    W, H = sc["W"], sc["H"]
    p = api.default_params(image_size=(W, H), cull_size=cull_size, downsample=ds, depth_tolerance=tol,
                           accum_mode=api.ACCUM_TOP5 if accum == "top5" else api.ACCUM_SUM,
                           sample_mode=api.SAMPLE_BILINEAR if bilinear else api.SAMPLE_NEAREST, frames_in_flight=flight)
    oc = oracle.Colorizer(cam_o, W, H, sc["cloud"], cull_size=cull_size, ds=ds, tol=tol, keep_obs=(accum == "top5"), bilinear=bilinear)
    with api.Colorizer(cam_g, p) as gc:
        gc.set_cloud(sc["cloud"])
        for i, pose in enumerate(sc["poses"]):
            mk = sc["masks"][i] if masks else None
            ro = oc.frame(pose, sc["images"][i], mask=mk, T_opt=T_opt, want_vis=True, want_pix=True)
            rg = gc.submit_frame(pose, sc["images"][i], mask=mk, T_opt=T_opt, want_vis=True, want_pix=True)
            assert rg["n_candidates"] == ro["n_candidates"], f"frame {i} candidates"
            assert rg["n_visible"] == ro["n_visible"], f"frame {i} visible count"
            assert rg["n_colored"] == ro["n_colored"], f"frame {i} coloured count"
            assert np.array_equal(rg["vis"], ro["vis"]), f"frame {i} visibility mask"
            assert np.array_equal(rg["pix"], ro["pix"]), f"frame {i} pixel indices"
        acc_g, mh_g, mm_g = gc.accumulators()
        acc_o, mh_o, mm_o = oc.accumulators()
        assert np.array_equal(acc_g, acc_o)
        if masks:
            assert np.array_equal(mh_g, mh_o) and np.array_equal(mm_g, mm_o)
        rgb_g, cnt_g = gc.finalize()
        if accum == "sum":
            assert np.array_equal(rgb_g, oc.finalize_sum())
            assert np.array_equal(cnt_g, acc_o[:, 3])
        else:
            rgb_o = oc.finalize_top5()
            assert np.abs(rgb_g.astype(int) - rgb_o.astype(int)).max() <= 1  # tolerance 1/255 per channel
        assert acc_o[:, 3].sum() > 0
    return acc_o


def test_tiny_sum(oracle, built_lib):
    _run_pair(oracle, make_scene(20_000, 3, 320, 240))


def test_tiny_bilinear_sum(oracle, built_lib):
    """PCP_SAMPLE_BILINEAR (north_star (c)): 8.8 fixed-point sums and rounded means, bit-exact against the restated oracle."""
    _run_pair(oracle, make_scene(20_000, 3, 320, 240), bilinear=True)


def test_bilinear_top5(oracle, built_lib):
    _run_pair(oracle, make_scene(30_000, 6, 320, 240), accum="top5", bilinear=True)


def test_bilinear_close_to_nearest(oracle, built_lib):
    """On the smooth synthetic images the bilinear colour stays within the local pixel range of the nearest one."""
    from pointcloudprocessor_b200 import api
    sc = make_scene(50_000, 4, 640, 480)
    out = []
    for mode in (api.SAMPLE_NEAREST, api.SAMPLE_BILINEAR):
        with api.Colorizer(api.create_camera("pinhole", sc["intr"], sc["dist"]), api.default_params(image_size=(640, 480), sample_mode=mode)) as gc:
            gc.set_cloud(sc["cloud"])
            for i, pose in enumerate(sc["poses"]):
                gc.submit_frame(pose, sc["images"][i])
            out.append(gc.finalize())
    (rgb_n, cnt_n), (rgb_b, cnt_b) = out
    assert np.array_equal(cnt_n, cnt_b)  # same coloured set
    assert cnt_n.sum() > 0
    assert np.abs(rgb_n.astype(int) - rgb_b.astype(int)).mean() < 8.0


@pytest.mark.parametrize("flight", [1, 2, 3])
def test_frames_in_flight_invariant(oracle, built_lib, flight):
    """Results do not depend on how many frames run concurrently on the internal streams."""
    _run_pair(oracle, make_scene(200_000, 8, 640, 480), flight=flight)


def test_async_pipeline_matches_oracle(oracle, built_lib):
    """Frames submitted without per-frame outputs (fully asynchronous, 3 in flight) give the oracle's accumulators."""
    from pointcloudprocessor_b200 import api
    sc = make_scene(300_000, 12, 640, 480, masks=True)
    oc = oracle.Colorizer(oracle.camera(sc["intr"], sc["dist"]), 640, 480, sc["cloud"])
    with api.Colorizer(api.create_camera("pinhole", sc["intr"], sc["dist"]), api.default_params(image_size=(640, 480), frames_in_flight=3)) as gc:
        gc.set_cloud(sc["cloud"])
        for rep in range(2):  # the second round re-uses every frame slot and staging buffer
            gc.reset_accumulators()
            for i, pose in enumerate(sc["poses"]):
                gc.submit_frame(pose, sc["images"][i], mask=sc["masks"][i])
        for i, pose in enumerate(sc["poses"]):
            oc.frame(pose, sc["images"][i], mask=sc["masks"][i], threads=4)
        acc_g, mh_g, mm_g = gc.accumulators()
        acc_o, mh_o, mm_o = oc.accumulators()
        assert np.array_equal(acc_g, acc_o) and np.array_equal(mh_g, mh_o) and np.array_equal(mm_g, mm_o)


def test_frames_in_flight_are_all_joined(built_lib):
    """A small cloud with several frames in flight and no per-frame outputs: SUM frames on different slots are independent, so
    reading the accumulators must wait for every slot, not only for the last submitted frame (bench C1 caught that race)."""
    from pointcloudprocessor_b200 import api
    W, H = 1280, 720
    sc = make_scene(1_000_000, 12, W, H)
    cam = api.create_camera("pinhole", sc["intr"], sc["dist"])
    accs = {}
    for flight in (1, 3):
        with api.Colorizer(cam, api.default_params(image_size=(W, H), frames_in_flight=flight)) as gc:
            gc.set_cloud(sc["cloud"])
            for rep in range(4):
                gc.reset_accumulators()
                for i, pose in enumerate(sc["poses"]):
                    gc.submit_frame(pose, sc["images"][i])
                acc, _, _ = gc.accumulators()
                rgb, cnt = gc.finalize()
                if flight == 1 and rep == 0:
                    accs["acc"], accs["rgb"], accs["cnt"] = acc, rgb, cnt
                assert np.array_equal(acc, accs["acc"]) and np.array_equal(rgb, accs["rgb"]) and np.array_equal(cnt, accs["cnt"]), (flight, rep)
    assert accs["cnt"].sum() > 10_000


@pytest.mark.parametrize("n", [20_003, 20_000, 7])
@pytest.mark.parametrize("bilinear", [False, True])
def test_finalize_forms_agree_with_oracle(oracle, built_lib, n, bilinear):
    """pcp_finalize_device has three kernels behind it (pack + unpack, the four-points-per-thread gather, the byte-store gather for
    unaligned output pointers): every one of them, with colours only / counts only / both and a point count that is not a multiple
    of four, must produce the oracle's integer means (smoothColors, cpp:604-631)."""
    import torch
    from pointcloudprocessor_b200 import _capi, api
    W, H = 320, 240
    sc = make_scene(n, 3, W, H)
    cam_o = oracle.camera(sc["intr"], sc["dist"])
    oc = oracle.Colorizer(cam_o, W, H, sc["cloud"], bilinear=bilinear)
    for i, pose in enumerate(sc["poses"]):
        oc.frame(pose, sc["images"][i])
    rgb_o, cnt_o = oc.finalize_sum(), oc.accumulators()[0][:, 3]
    cam = api.create_camera("pinhole", sc["intr"], sc["dist"])
    sm = api.SAMPLE_BILINEAR if bilinear else api.SAMPLE_NEAREST
    for flags in (0, _capi.DEBUG_FINALIZE_ONEPASS, _capi.DEBUG_FINALIZE_TWOPASS):
        with api.Colorizer(cam, api.default_params(image_size=(W, H), sample_mode=sm, debug_flags=flags)) as gc:
            gc.set_cloud(sc["cloud"])
            for i, pose in enumerate(sc["poses"]):
                gc.submit_frame(pose, sc["images"][i])
            for off in (0, 1):  # 1: an output pointer that is not 4-byte aligned
                d_rgb = torch.full((n * 3 + 16,), 77, dtype=torch.uint8, device="cuda")
                d_cnt = torch.full((n + 4,), -1, dtype=torch.int32, device="cuda")
                for want_rgb, want_cnt in ((True, True), (True, False), (False, True)):
                    d_rgb.fill_(77)
                    d_cnt.fill_(-1)
                    gc.finalize_device(d_rgb.data_ptr() + off if want_rgb else 0, d_cnt.data_ptr() if want_cnt else 0)
                    gc.synchronize()
                    r, c = d_rgb.cpu().numpy(), d_cnt.cpu().numpy()
                    if want_rgb:
                        assert np.array_equal(r[off:off + 3 * n].reshape(n, 3), rgb_o), (flags, off)
                        assert (r[:off] == 77).all() and (r[off + 3 * n:] == 77).all(), "wrote outside the colour array"
                    else:
                        assert (r == 77).all()
                    if want_cnt:
                        assert np.array_equal(c[:n].view(np.uint32), cnt_o) and (c[n:] == -1).all(), (flags, off)
                    else:
                        assert (c == -1).all()
    assert cnt_o.sum() > 0 or n < 100


def test_finalize_picks_gather_by_point_order(oracle, built_lib):
    """The same 1e6 points in a shuffled and in a voxel-grid order: the locality measured at pre-binning tells them apart, and the
    colours (whatever gather pcp_finalize picks) equal the oracle's in the caller's order both times."""
    from pointcloudprocessor_b200 import api
    W, H = 320, 240
    sc = make_scene(1_000_000, 3, W, H)
    cloud = sc["cloud"]
    cell = np.floor(cloud[:, :3] / 0.5).astype(np.int64)
    cell -= cell.min(axis=0)
    ext = cell.max(axis=0) + 1
    voxel = np.ascontiguousarray(cloud[np.argsort((cell[:, 2] * ext[1] + cell[:, 1]) * ext[0] + cell[:, 0], kind="stable")])
    cam_o = oracle.camera(sc["intr"], sc["dist"])
    cam = api.create_camera("pinhole", sc["intr"], sc["dist"])
    loc = {}
    for name, pts in (("shuffled", cloud), ("voxel", voxel)):
        oc = oracle.Colorizer(cam_o, W, H, pts)
        with api.Colorizer(cam, api.default_params(image_size=(W, H))) as gc:
            gc.set_cloud(pts)
            for i, pose in enumerate(sc["poses"]):
                oc.frame(pose, sc["images"][i])
                gc.submit_frame(pose, sc["images"][i])
            rgb, cnt = gc.finalize()
            loc[name] = gc.debug_counters()["order_locality"]
        assert np.array_equal(rgb, oc.finalize_sum()) and np.array_equal(cnt, oc.accumulators()[0][:, 3]), name
        assert cnt.sum() > 0
    assert loc["shuffled"] < 0.5 < loc["voxel"], loc


@pytest.mark.parametrize("flight,group", [(0, 0), (12, 4), (6, 2), (4, 4), (6, 4), (3, 1), (8, 1)])
def test_frames_per_launch_invariant(oracle, built_lib, flight, group):
    """pcp_submit_frames_device may put 2 or 4 consecutive frames into one launch of each kernel (small clouds: the launches bound the
    frame rate).  Whatever the grouping, the slot count, the way the calls cut the frame list (groups that do not line up with the
    slots fall back to single frames) and single host-image frames in between: the accumulators, mask accumulators included, are the
    oracle's, bit for bit."""
    import torch
    from pointcloudprocessor_b200 import api
    W, H = 320, 240
    f = 23
    sc = make_scene(150_000, f, W, H, masks=True)
    cam_o = oracle.camera(sc["intr"], sc["dist"])
    oc = oracle.Colorizer(cam_o, W, H, sc["cloud"])
    for i, pose in enumerate(sc["poses"]):
        oc.frame(pose, sc["images"][i], mask=sc["masks"][i] if i >= 9 else None)
    acc_o, mh_o, mm_o = oc.accumulators()
    d_img = torch.from_numpy(np.stack(sc["images"])).cuda()
    d_msk = torch.from_numpy(np.stack(sc["masks"])).cuda()
    iptr = np.array([d_img.data_ptr() + i * H * W * 3 for i in range(f)], np.uint64)
    mptr = np.array([d_msk.data_ptr() + i * H * W for i in range(f)], np.uint64)
    cam = api.create_camera("pinhole", sc["intr"], sc["dist"])
    prm = api.default_params(image_size=(W, H), frames_in_flight=flight, frames_per_launch=group)
    with api.Colorizer(cam, prm) as gc:
        gc.set_cloud(sc["cloud"])
        for rep in range(2):
            gc.reset_accumulators()
            poses = np.asarray(sc["poses"], np.float64)
            gc.submit_frames_device(poses[0:5], iptr[0:5], W * 3)                       # 4 + 1 (or 2 + 2 + 1)
            gc.submit_frame(sc["poses"][5], sc["images"][5])                            # a single frame from a host image in between
            gc.submit_frames_device(poses[6:9], iptr[6:9], W * 3)                       # starts on a slot that is not a group boundary
            gc.submit_frames_device(poses[9:22], iptr[9:22], W * 3, d_mask_ptrs=mptr[9:22], mask_pitch=W)   # masks begin mid-way
            gc.submit_frame(sc["poses"][22], sc["images"][22], mask=sc["masks"][22])
            acc_g, mh_g, mm_g = gc.accumulators()
            assert np.array_equal(acc_g, acc_o), (flight, group, rep)
            assert np.array_equal(mh_g, mh_o) and np.array_equal(mm_g, mm_o), (flight, group, rep)
            rgb, cnt = gc.finalize()
            assert np.array_equal(rgb, oc.finalize_sum()) and np.array_equal(cnt, acc_o[:, 3])
    assert acc_o[:, 3].sum() > 0 and mh_o.sum() > 0


@pytest.mark.parametrize("n", [50_000, 31])
def test_frame_mask_cloud_on_device(oracle, built_lib, n):
    """pcp_frame_mask_cloud (generateSegmentMap, cpp:769-816, gathered on the device) against the same records built from the frame's
    pixel indices with the oracle's transform: positions bit-exact, colours, mask values, red paint for mask 255, the map's order;
    with the pixel indices copied out and with PCP_OUT_PIXELS_ON_DEVICE; and the calls that must be refused."""
    from pointcloudprocessor_b200 import api
    W, H = 320, 240
    sc = make_scene(n, 3, W, H, masks=True)
    cloud = sc["cloud"]
    cam = api.create_camera("pinhole", sc["intr"], sc["dist"])
    with api.Colorizer(cam, api.default_params(image_size=(W, H))) as gc:
        gc.set_cloud(cloud)
        with pytest.raises(api.PcpError):
            gc.frame_mask_cloud()                                     # no frame yet
        for i, pose in enumerate(sc["poses"]):
            img, msk = sc["images"][i], sc["masks"][i].copy()
            msk[::3, ::5] = 255                                       # some pixels of the class that is painted red
            r = gc.submit_frame(pose, img, mask=msk, want_pix=True)
            got = gc.frame_mask_cloud()
            w2c, _ = oracle.pose_matrices(pose)
            sel = np.nonzero(r["pix"] >= 0)[0]
            pix = r["pix"][sel]
            assert len(got) == len(sel) == r["n_colored"]
            want_xyz = oracle.transform(cloud[sel, :3].copy(), w2c)
            assert np.array_equal(np.stack([got["x"], got["y"], got["z"]], axis=1).view(np.uint32), want_xyz.view(np.uint32))
            bgr = img.reshape(-1, 3)[pix]
            mv = msk.reshape(-1)[pix]
            red = mv == 255
            assert np.array_equal(got["mask"], mv.astype(np.uint16)) and (got["a"] == 255).all() and (got["pad"] == 0).all()
            assert np.array_equal(got["r"], np.where(red, 255, bgr[:, 2])) and np.array_equal(got["g"], np.where(red, 0, bgr[:, 1]))
            assert np.array_equal(got["b"], np.where(red, 0, bgr[:, 0]))
            if n > 1000:
                assert red.any() and (~red).any()
            # the same frame again with the indices left on the device: same records; and again after a call that shares its index
            # scratch (the compacted final cloud)
            gc.submit_frame(pose, img, mask=msk, keep_pix_on_device=True)
            assert np.array_equal(gc.frame_mask_cloud(), got)
            gc.finalize_compact()                                     # shares the index scratch; not a frame
            assert np.array_equal(gc.frame_mask_cloud(), got)
            gc.cull_pose(sc["poses"][(i + 1) % len(sc["poses"])])     # a frame of its own (no colours): the kept indices are gone
            with pytest.raises(api.PcpError):
                gc.frame_mask_cloud()
        gc.submit_frame(sc["poses"][0], sc["images"][0], mask=sc["masks"][0])          # pixel indices not kept
        with pytest.raises(api.PcpError):
            gc.frame_mask_cloud()
        gc.submit_frame(sc["poses"][0], sc["images"][0], want_pix=True)                 # no mask
        with pytest.raises(api.PcpError):
            gc.frame_mask_cloud()
        gc.submit_frame(sc["poses"][0], sc["images"][0], mask=sc["masks"][0], want_pix=True)
        assert len(gc.frame_mask_cloud()) > 0 or n < 1000
        gc.set_cloud(cloud)                                                             # a new map: the kept indices are stale
        with pytest.raises(api.PcpError):
            gc.frame_mask_cloud()


def _big_shape_parity(oracle, n, f, W, H, n_images, flight):
    """Accumulators of f frames over an n-point cloud against the oracle (all host threads), frames submitted without per-frame
    outputs so that `flight` of them are in flight; images are cycled (the kernels see the same work)."""
    import os
    from pointcloudprocessor_b200 import api, synth
    intr, dist = synth.make_camera(W, H)
    cloud = synth.make_cloud(n)
    poses, _ = synth.make_poses(f)
    images = [synth.make_image(i, W, H) for i in range(n_images)]
    threads = max(1, len(os.sched_getaffinity(0)))
    oc = oracle.Colorizer(oracle.camera(intr, dist), W, H, cloud)
    with api.Colorizer(api.create_camera("pinhole", intr, dist), api.default_params(image_size=(W, H), frames_in_flight=flight)) as gc:
        gc.set_cloud(cloud)
        for i in range(f):
            gc.submit_frame(poses[i], images[i % n_images])
        for i in range(f):
            oc.frame(poses[i], images[i % n_images], threads=threads)
        acc_g, _, _ = gc.accumulators()
        rgb_g, cnt_g = gc.finalize()
    acc_o = oc.accumulators()[0]
    assert np.array_equal(acc_g, acc_o), int((acc_g != acc_o).any(axis=1).sum())
    assert np.array_equal(rgb_g, oc.finalize_sum()) and int(cnt_g.astype(bool).sum()) > n // 50
    oc.close()


def test_c5_shape_200_small_frames(oracle, built_lib):
    """BASELINE config 5's regime (tiny frames over a large cloud, launch / occlusion dominated): 5e6 points x 200 frames @ 640x480,
    three frames in flight, bit-exact accumulators."""
    _big_shape_parity(oracle, 5_000_000, 200, 640, 480, n_images=8, flight=3)


def test_c4_shape_large_cloud(oracle, built_lib):
    """BASELINE config 4's regime (a 1e8-class map): 2e7 points x 6 frames @ 1920x1080, bit-exact accumulators."""
    _big_shape_parity(oracle, 20_000_000, 6, 1920, 1080, n_images=3, flight=3)


def test_tiny_masks(oracle, built_lib):
    _run_pair(oracle, make_scene(20_000, 3, 320, 240, masks=True), masks=True)


def test_tiny_top5(oracle, built_lib):
    _run_pair(oracle, make_scene(30_000, 6, 320, 240), accum="top5")


def test_c1_shape(oracle, built_lib):
    _run_pair(oracle, make_scene(1_000_000, 4, 1280, 720))


def test_c1_shape_fast_variant(oracle, built_lib):
    """The kernel variant that skips the n_candidates count (what the asynchronous path and the bench run)."""
    from pointcloudprocessor_b200 import api
    sc = make_scene(1_000_000, 4, 1280, 720)
    oc = oracle.Colorizer(oracle.camera(sc["intr"], sc["dist"]), 1280, 720, sc["cloud"])
    with api.Colorizer(api.create_camera("pinhole", sc["intr"], sc["dist"]), api.default_params(image_size=(1280, 720))) as gc:
        gc.set_cloud(sc["cloud"])
        for i, pose in enumerate(sc["poses"]):
            ro = oc.frame(pose, sc["images"][i], want_vis=True, want_pix=True, threads=8)
            rg = gc.submit_frame(pose, sc["images"][i], want_vis=True, want_pix=True, count_candidates=False)
            assert rg["n_visible"] == ro["n_visible"] and rg["n_colored"] == ro["n_colored"]
            assert np.array_equal(rg["vis"], ro["vis"]) and np.array_equal(rg["pix"], ro["pix"])
        assert np.array_equal(gc.accumulators()[0], oc.accumulators()[0])


def test_c3_shape_masks_top5_and_nid_refinement(oracle, built_lib):
    """BASELINE config 3: the C1 shape plus segmentation masks, the reference's TOP5 accumulation, and the NID cost / pose refinement over
    the culled cloud of a keyframe (values against the oracle's doubles; the refinement must not leave the gate and must not raise the cost)."""
    from pointcloudprocessor_b200 import api, synth
    W, H = 1280, 720
    sc = make_scene(1_000_000, 3, W, H, masks=True)
    _run_pair(oracle, sc, accum="top5", masks=True)
    cam_o = oracle.camera(sc["intr"], sc["dist"])
    w2c, _ = oracle.pose_matrices(sc["poses"][0])
    pts_cam = oracle.transform(sc["cloud"], w2c)
    keep = oracle.view_culling(cam_o, W, H, pts_cam)
    culled = np.concatenate([pts_cam[keep], sc["cloud"][keep, 3:]], 1)
    grey = synth.grey_plane(sc["images"][0])
    cands = synth.make_nid_candidates(4)
    with api.Colorizer(api.create_camera("pinhole", sc["intr"], sc["dist"]), api.default_params(image_size=(W, H))) as gc:
        cost = api.NIDCost(gc, grey, culled, 16)
        val, grad, ok = cost.evaluate_with_gradient(cands)
        for b in range(len(cands)):
            rok, rv, rg = oracle.nid_grad(cam_o, grey, culled, cands[b])
            assert ok[b] == rok and abs(val[b] - rv) <= 1e-5 * abs(rv)
            assert np.abs(grad[b] - rg).max() <= 1e-5 * np.abs(rg).max()
        T, s = api.calibrate([cost])
        ident = np.array([0, 0, 0, 1, 0, 0, 0.0])
        assert s["final_cost"] <= s["initial_cost"] and s["initial_cost"] == pytest.approx(oracle.nid(cam_o, grey, culled, ident)[0], rel=1e-5)
        assert np.linalg.norm(T[:3, 3]) <= 0.2 * s["outer_iterations"] + 1e-9
        cost.close()


def test_reference_cull_size(oracle, built_lib):
    # the reference passes a hard-coded 4096x3000 to ViewCulling whatever the image (cpp:525)
    _run_pair(oracle, make_scene(100_000, 2, 640, 480), cull_size=(4096, 3000))


def test_extrinsic_branch(oracle, built_lib):
    # enableNIDOptimize / manual guess: general float affine inverse (cpp:504-519)
    T = np.eye(4)
    a = np.deg2rad(0.7)
    T[:3, :3] = np.array([[np.cos(a), -np.sin(a), 0], [np.sin(a), np.cos(a), 0], [0, 0, 1]])
    T[:3, 3] = (0.03, -0.02, 0.05)
    _run_pair(oracle, make_scene(50_000, 3, 320, 240), T_opt=T)


def test_ragged_and_empty(oracle, built_lib):
    from pointcloudprocessor_b200 import api
    for n in (0, 1, 31, 33, 257):
        sc = make_scene(max(n, 1), 2, 320, 240)
        sc["cloud"] = sc["cloud"][:n]
        cam_o = oracle.camera(sc["intr"], sc["dist"])
        cam_g = api.create_camera("pinhole", sc["intr"], sc["dist"])
        oc = oracle.Colorizer(cam_o, 320, 240, sc["cloud"].reshape(-1, 4))
        with api.Colorizer(cam_g, api.default_params(image_size=(320, 240))) as gc:
            gc.set_cloud(sc["cloud"].reshape(-1, 4))
            for i in range(2):
                ro = oc.frame(sc["poses"][i], sc["images"][i], want_vis=True, want_pix=True)
                rg = gc.submit_frame(sc["poses"][i], sc["images"][i], want_vis=True, want_pix=True)
                assert np.array_equal(rg["vis"], ro["vis"]) and np.array_equal(rg["pix"], ro["pix"])
            assert np.array_equal(gc.accumulators()[0], oc.accumulators()[0])


def test_edge_points(oracle, built_lib):
    """z == 0, points behind the camera, coincident points, exact depth ties, non-finite coordinates, negative-u
    truncation into cell 0, border cells."""
    from pointcloudprocessor_b200 import api
    W, H = 320, 240
    sc = make_scene(5000, 1, W, H)
    rng = np.random.default_rng(7)
    pose = np.array([0, 0, 0, 1, 0, 0, 0], np.float64)  # identity: cloud coordinates are camera coordinates
    intr = sc["intr"]
    pts = []
    # dense fan in front of the camera, including just outside each image border
    u = rng.uniform(-30, W + 30, 4000)
    v = rng.uniform(-30, H + 30, 4000)
    z = rng.uniform(0.5, 6.0, 4000)
    pts.append(np.stack([(u - intr[2]) / intr[0] * z, (v - intr[3]) / intr[1] * z, z], 1))
    pts.append(np.array([[0, 0, 0], [1, 1, 0], [0.1, 0.1, -1], [0, 0, 1], [0, 0, 1], [0, 0, 1.00000012], [np.nan, 0, 1], [0, np.inf, 1],
                         [0, 0, np.inf], [1e30, 1e30, 1e-30], [-1e-3, 0, 1e-38], [3e38, 3e38, 3e38]], np.float64))
    # exact ties: same direction and distance, different indices
    pts.append(np.tile(np.array([[0.2, 0.1, 2.0]]), (5, 1)))
    xyz = np.concatenate(pts).astype(np.float32)
    cloud = np.concatenate([xyz, rng.random((xyz.shape[0], 1), dtype=np.float32)], 1)
    cam_o = oracle.camera(sc["intr"], sc["dist"])
    cam_g = api.create_camera("pinhole", sc["intr"], sc["dist"])
    oc = oracle.Colorizer(cam_o, W, H, cloud)
    with api.Colorizer(cam_g, api.default_params(image_size=(W, H))) as gc:
        gc.set_cloud(cloud)
        ro = oc.frame(pose, sc["images"][0], want_vis=True, want_pix=True)
        rg = gc.submit_frame(pose, sc["images"][0], want_vis=True, want_pix=True)
        assert ro["n_visible"] > 100
        assert rg["n_candidates"] == ro["n_candidates"]
        assert np.array_equal(rg["vis"], ro["vis"]) and np.array_equal(rg["pix"], ro["pix"])
        assert np.array_equal(gc.accumulators()[0], oc.accumulators()[0])


def test_cull_dropin(oracle, built_lib):
    """vlcal::ViewCulling::cull: ascending index list, with a double T_camera_lidar and without depth culling."""
    from pointcloudprocessor_b200 import api
    sc = make_scene(60_000, 2, 640, 480)
    cam_o = oracle.camera(sc["intr"], sc["dist"])
    cam_g = api.create_camera("pinhole", sc["intr"], sc["dist"])
    w2c, _ = oracle.pose_matrices(sc["poses"][1])
    pts_cam = oracle.transform(sc["cloud"], w2c)
    cloud_cam = np.concatenate([pts_cam, sc["cloud"][:, 3:]], 1)
    T = np.eye(4)
    T[:3, 3] = (0.01, 0.02, -0.03)
    b = np.deg2rad(1.0)
    T[:3, :3] = np.array([[1, 0, 0], [0, np.cos(b), -np.sin(b)], [0, np.sin(b), np.cos(b)]])
    for depth in (True, False):
        for Tm in (None, T):
            ref = oracle.view_culling(cam_o, 640, 480, pts_cam, T=Tm, depth_culling=depth)
            vc = api.ViewCulling(cam_g, (640, 480), enable_depth_buffer_culling=depth)
            got = vc.cull(cloud_cam, Tm, return_indices=True)
            vc.close()
            assert np.array_equal(got, ref), (depth, Tm is not None)
    # pose variant (loop body of viewCullingAndSaveFilteredPcds)
    with api.Colorizer(cam_g, api.default_params(image_size=(640, 480))) as gc:
        gc.set_cloud(sc["cloud"])
        got = gc.cull_pose(sc["poses"][1])
        assert np.array_equal(got, oracle.view_culling(cam_o, 640, 480, pts_cam))


def test_pcl_record_layout(oracle, built_lib):
    """pcl::PointXYZI records: 32 bytes, intensity at byte 16."""
    from pointcloudprocessor_b200 import api
    sc = make_scene(10_000, 1, 320, 240)
    rec = np.zeros((10_000, 8), np.float32)
    rec[:, :3] = sc["cloud"][:, :3]
    rec[:, 3] = 1.0
    rec[:, 4] = sc["cloud"][:, 3]
    cam_o = oracle.camera(sc["intr"], sc["dist"])
    cam_g = api.create_camera("pinhole", sc["intr"], sc["dist"])
    oc = oracle.Colorizer(cam_o, 320, 240, sc["cloud"])
    with api.Colorizer(cam_g, api.default_params(image_size=(320, 240))) as gc:
        gc.set_cloud(rec)
        ro = oc.frame(sc["poses"][0], sc["images"][0], want_vis=True)
        rg = gc.submit_frame(sc["poses"][0], sc["images"][0], want_vis=True)
        assert np.array_equal(rg["vis"], ro["vis"])


def test_nid(oracle, built_lib):
    """NID within 1e-5 relative of the oracle's doubles (north_star tolerance)."""
    from pointcloudprocessor_b200 import api, synth
    W, H = 640, 480
    sc = make_scene(200_000, 1, W, H)
    cam_o = oracle.camera(sc["intr"], sc["dist"])
    cam_g = api.create_camera("pinhole", sc["intr"], sc["dist"])
    w2c, _ = oracle.pose_matrices(sc["poses"][0])
    pts_cam = oracle.transform(sc["cloud"], w2c)
    keep = oracle.view_culling(cam_o, W, H, pts_cam)
    culled = np.concatenate([pts_cam[keep], sc["cloud"][keep, 3:]], 1)
    grey = synth.grey_plane(sc["images"][0])
    cands = synth.make_nid_candidates(16)
    with api.Colorizer(cam_g, api.default_params(image_size=(W, H))) as gc:
        cost = api.NIDCost(gc, grey, culled, 16)
        got, ok = cost(cands)
        for b in range(cands.shape[0]):
            ref, rok = oracle.nid(cam_o, grey, culled, cands[b])
            assert ok[b] == rok
            assert abs(got[b] - ref) <= 1e-5 * abs(ref), (b, got[b], ref)
        # the 8-bit image itself (pcp_nid_create_u8: conversion to [0,1] and binning on the device): the same doubles, bit for bit
        grey_u8 = np.rint(grey * 255.0).astype(np.uint8)
        assert np.array_equal(grey_u8.astype(np.float64) * (1.0 / 255.0), grey)
        cost_u8 = api.NIDCost(gc, grey_u8, culled, 16)
        got_u8, ok_u8 = cost_u8(cands)
        assert np.array_equal(got_u8, got) and np.array_equal(ok_u8, ok)
        cost_u8.close()
        # ... and the reference's actual read of a 3-channel image: the first W interleaved B,G,R values of every row (nid_cost.hpp:61, :88)
        bgr = sc["images"][0]
        cost_q = api.NIDCost(gc, bgr, culled, 16)
        plane_q = np.ascontiguousarray(bgr.reshape(H, W * 3)[:, :W]).astype(np.float64) * (1.0 / 255.0)
        got_q, ok_q = cost_q(cands[:4])
        for b in range(4):
            ref, rok = oracle.nid(cam_o, plane_q, culled, cands[b])
            assert ok_q[b] == rok and abs(got_q[b] - ref) <= 1e-5 * abs(ref)
        cost_q.close()
        # MultiNIDCost: gate + sum
        far = cands[1].copy()
        far[4] += 0.5
        s, okm = api.multi_nid([cost, cost], cands[0], np.stack([cands[2], far]))
        r2, _ = oracle.nid(cam_o, grey, culled, cands[2])
        assert okm[0] and not okm[1]
        assert abs(s[0] - 2 * r2) <= 2e-5 * abs(r2)
        assert oracle.nid_gate(cands[0], cands[2]) and not oracle.nid_gate(cands[0], far)
        cost.close()


def test_nid_gradient(oracle, built_lib):
    """NIDCost::operator()<ceres::Jet<double,7>>: value and the 7 ambient partials within 1e-5 relative of the oracle's Jet7."""
    from pointcloudprocessor_b200 import api, synth
    W, H = 640, 480
    sc = make_scene(120_000, 1, W, H)
    cam_o = oracle.camera(sc["intr"], sc["dist"])
    w2c, _ = oracle.pose_matrices(sc["poses"][0])
    pts_cam = oracle.transform(sc["cloud"], w2c)
    keep = oracle.view_culling(cam_o, W, H, pts_cam)
    culled = np.concatenate([pts_cam[keep], sc["cloud"][keep, 3:]], 1)
    grey = synth.grey_plane(sc["images"][0])
    cands = synth.make_nid_candidates(6)
    with api.Colorizer(api.create_camera("pinhole", sc["intr"], sc["dist"]), api.default_params(image_size=(W, H))) as gc:
        cost = api.NIDCost(gc, grey, culled, 16)
        val, grad, ok = cost.evaluate_with_gradient(cands)
        val0, ok0 = cost(cands)
        # the value kernel accumulates its histogram in 2^-31 fixed point (order-independent), the gradient kernel in doubles: they
        # agree far inside the 1e-5 of north_star, not to the last bits
        assert np.allclose(val, val0, rtol=1e-8, atol=0) and np.array_equal(ok, ok0)
        for b in range(cands.shape[0]):
            rok, rv, rg = oracle.nid_grad(cam_o, grey, culled, cands[b])
            assert ok[b] == rok
            assert abs(val[b] - rv) <= 1e-5 * abs(rv)
            assert np.abs(grad[b] - rg).max() <= 1e-5 * np.abs(rg).max(), (b, grad[b], rg)
        s, g, okm = api.multi_nid_with_gradient([cost, cost], cands[0], cands[1:3])
        assert okm.all() and np.allclose(s, 2 * val[1:3], rtol=1e-12) and np.allclose(g, 2 * grad[1:3], rtol=1e-12)
        cost.close()


@pytest.mark.parametrize("width", [4096, 4099, 33])
def test_hsv_roundtrip_device(oracle, built_lib, width):
    """hsv_roundtrip_kernel against the oracle on every 24-bit colour (and against cv2 when it is importable on the box)."""
    from pointcloudprocessor_b200 import api, synth
    v = np.arange(1 << 24, dtype=np.uint32)
    bgr = np.stack([(v & 255), (v >> 8) & 255, (v >> 16) & 255], axis=1).astype(np.uint8)
    rows = (1 << 24) // width
    img = np.ascontiguousarray(bgr[: rows * width].reshape(rows, width, 3))
    intr, dist = synth.make_camera(64, 64)
    with api.Colorizer(api.create_camera("pinhole", intr, dist), api.default_params(image_size=(64, 64))) as gc:
        for sat, val in ((1.0, 1.0), (1.2, 0.9)):
            out = gc.hsv_roundtrip(img, sat, val)
            assert np.array_equal(out, oracle.hsv_roundtrip(img, sat, val))
            try:
                import cv2
            except ImportError:
                continue
            hsv = cv2.cvtColor(img, cv2.COLOR_BGR2HSV)
            hsv[..., 1] = np.clip(np.rint(hsv[..., 1].astype(np.float32) * np.float32(sat)), 0, 255).astype(np.uint8)
            hsv[..., 2] = np.clip(np.rint(hsv[..., 2].astype(np.float32) * np.float32(val)), 0, 255).astype(np.uint8)
            assert np.array_equal(out, cv2.cvtColor(hsv, cv2.COLOR_HSV2BGR))


def test_hsv_adjust_in_pipeline(oracle, built_lib):
    """hsv_adjust = 1: raw images go in, the device applies the round trip; equals the oracle fed with adjusted images."""
    from pointcloudprocessor_b200 import api
    sc = make_scene(50_000, 4, 330, 240)  # 330 = 10 * 32 + 10: vector body and scalar tail
    rng = np.random.default_rng(5)
    raw = [rng.integers(0, 256, (240, 330, 3), dtype=np.uint8) for _ in sc["poses"]]
    oc = oracle.Colorizer(oracle.camera(sc["intr"], sc["dist"]), 330, 240, sc["cloud"])
    with api.Colorizer(api.create_camera("pinhole", sc["intr"], sc["dist"]), api.default_params(image_size=(330, 240), hsv_adjust=1)) as gc:
        gc.set_cloud(sc["cloud"])
        for i, pose in enumerate(sc["poses"]):
            gc.submit_frame(pose, raw[i])
            oc.frame(pose, oracle.hsv_roundtrip(raw[i]))
        assert np.array_equal(gc.accumulators()[0], oc.accumulators()[0])
        assert oc.accumulators()[0][:, 3].sum() > 0


@pytest.mark.parametrize("shape", [(60_000, 640, 480, True), (20_000, 160, 120, False), (200_000, 320, 240, True)])
def test_distance_masks_equal_oracle(oracle, built_lib, shape):
    """pcp_splat_masks_pose (scripts/genNormAndDistanceMask.py:233-265): distance, 3-D point and winner-index masks of
    the frame's culled cloud, bit for bit the oracle's file-order loop (the last point on a pixel wins = the largest index)."""
    from pointcloudprocessor_b200 import api
    n, W, H, depth_culling = shape
    sc = make_scene(n, 3, W, H)
    cam_o = oracle.camera(sc["intr"], sc["dist"])
    cam_g = api.create_camera("pinhole", sc["intr"], sc["dist"])
    p = api.default_params(image_size=(W, H))
    p.enable_depth_buffer_culling = 1 if depth_culling else 0
    with api.Colorizer(cam_g, p) as gc:
        gc.set_cloud(sc["cloud"])
        for pose in sc["poses"]:
            w2c, _ = oracle.pose_matrices(pose)
            pts = oracle.transform(sc["cloud"][:, :3], w2c)
            keep = oracle.view_culling(cam_o, W, H, pts, depth_culling=depth_culling)
            d_o, xyz_o, idx_o = oracle.distance_masks(cam_o, W, H, pts[keep])
            idx_o = np.where(idx_o >= 0, keep[np.maximum(idx_o, 0)], -1).astype(np.int32)  # file position -> map index
            d_g, xyz_g, idx_g, filled = gc.splat_masks_pose(pose)
            assert filled == int((idx_o >= 0).sum()) and filled > 50
            assert np.array_equal(idx_g, idx_o)
            assert np.array_equal(d_g.view(np.uint32), d_o.view(np.uint32))
            assert np.array_equal(xyz_g.view(np.uint32), xyz_o.view(np.uint32))
            # optional outputs may be left out
            d2, x2, i2, f2 = gc.splat_masks_pose(pose, want_xyz=False, want_index=False)
            assert x2 is None and i2 is None and f2 == filled and np.array_equal(d2, d_g)
        # the masks are read-only products: nothing was accumulated
        acc, _, _ = gc.accumulators()
        assert acc.sum() == 0
