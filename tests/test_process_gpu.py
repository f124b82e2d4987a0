"""PointCloudProcessor::process() drop-in (pcp_process): files in, files out, against the oracle driven by a Python restatement of the
reference's host control flow (src/PointCloudProcessor.cpp:965-1075)."""
import os
import struct

import numpy as np
import pytest

from conftest import make_scene

pytestmark = pytest.mark.gpu


def _f8(a):
    """What survives pcl::PCDWriter::writeASCII's 8 significant digits."""
    return np.array([np.float32(float("%.8g" % float(v))) for v in np.asarray(a, np.float32).ravel()], np.float32).reshape(np.shape(a))


def _write_binary_pcd(path, c):
    n = len(c)
    hdr = (f"# .PCD v0.7 - Point Cloud Data file format\nVERSION 0.7\nFIELDS x y z intensity\nSIZE 4 4 4 4\nTYPE F F F F\nCOUNT 1 1 1 1\n"
           f"WIDTH {n}\nHEIGHT 1\nVIEWPOINT 0 0 0 1 0 0 0\nPOINTS {n}\nDATA binary\n")
    with open(path, "wb") as f:
        f.write(hdr.encode() + np.ascontiguousarray(c, np.float32).tobytes())


def _write_ppm(path, bgr):
    with open(path, "wb") as f:
        f.write(b"P6\n%d %d\n255\n" % (bgr.shape[1], bgr.shape[0]) + np.ascontiguousarray(bgr[:, :, ::-1]).tobytes())


def _read_ascii_pcd(path):
    with open(path) as f:
        lines = f.read().splitlines()
    k = lines.index("DATA ascii")
    fields = [l for l in lines if l.startswith("FIELDS")][0].split()[1:]
    rows = [l.split() for l in lines[k + 1:]]
    return fields, rows


def _dataset(tmp_path, n=30_000, f=6, W=330, H=240, masks=True, jpeg=False):
    import cv2
    from pointcloudprocessor_b200 import synth
    sc = make_scene(n, f, W, H, masks=masks)
    img_dir, mask_dir, out_dir = tmp_path / "img", tmp_path / "mask", tmp_path / "out"
    for d in (img_dir, mask_dir, out_dir):
        d.mkdir()
    _write_binary_pcd(tmp_path / "scans.pcd", sc["cloud"])
    rng = np.random.default_rng(11)
    lines, frames = [], []
    for i, pose in enumerate(sc["poses"]):
        ts = 1700000000.0 + 0.1 * i
        raw = rng.integers(0, 256, (H, W, 3), dtype=np.uint8) if not jpeg else sc["images"][i]
        name = "%f" % ts
        if jpeg:
            cv2.imwrite(str(img_dir / (name + ".jpg")), raw, [cv2.IMWRITE_JPEG_QUALITY, 95])
        else:
            _write_ppm(img_dir / (name + ".ppm"), raw)
        have_mask = masks and i != 2  # one keyframe without a mask file: logged and skipped (cpp:776-781)
        if have_mask:
            cv2.imwrite(str(mask_dir / (name + ".png")), sc["masks"][i])
        lines.append("%.6f %.9f %.9f %.9f %.12f %.12f %.12f %.12f" % ((ts,) + tuple(pose)))
        frames.append(dict(ts=ts, pose=np.array([float(x) for x in lines[-1].split()[1:]]), raw=raw, mask=sc["masks"][i] if have_mask else None, name=name))
        # a second pose 3 cm further: not a keyframe (< 0.1 m from the last keyframe), with an image
        p2 = np.array(frames[-1]["pose"]); p2[0] += 0.03
        ts2 = ts + 0.05
        _write_ppm(img_dir / ("%f.ppm" % ts2), raw)
        lines.append("%.6f %.9f %.9f %.9f %.12f %.12f %.12f %.12f" % ((ts2,) + tuple(p2)))
    lines.insert(3, "%.6f 0 0 0 1 0 0 0" % 1600000000.0)  # a pose without an image: skipped (cpp:984)
    lines.append("this line is malformed and ends the file")  # cpp:974-977
    lines.append("%.6f 9 9 9 1 0 0 0" % 1800000000.0)
    (tmp_path / "odom.txt").write_text("\n".join(lines) + "\n")
    return sc, frames, img_dir, mask_dir, out_dir


@pytest.mark.parametrize("masks", [False, True])
def test_process_files(oracle, built_lib, tmp_path, masks):
    from pointcloudprocessor_b200 import api
    W, H = 330, 240
    sc, frames, img_dir, mask_dir, out_dir = _dataset(tmp_path, W=W, H=H, masks=masks)
    cam = api.create_camera("pinhole", sc["intr"], sc["dist"])
    proc = api.PointCloudProcessor(tmp_path / "scans.pcd", tmp_path / "odom.txt", str(img_dir) + "/", (str(mask_dir) + "/") if masks else "",
                                   str(out_dir) + "/", camera=cam, cull_size=(W, H))
    st = proc.process()
    cloud = sc["cloud"]
    assert st["n_points"] == len(cloud) and st["n_frames"] == 2 * len(frames) and st["n_keyframes"] == len(frames)

    oc = oracle.Colorizer(oracle.camera(sc["intr"], sc["dist"]), W, H, cloud, keep_obs=True)
    world_mask_rows = []
    for fr in frames:
        adj = oracle.hsv_roundtrip(fr["raw"])
        r = oc.frame(fr["pose"], adj, mask=fr["mask"], want_vis=True, want_pix=True)
        w2c, c2w = oracle.pose_matrices(fr["pose"])
        # filtered_pcd/<ts>_beforeNID.pcd: the culled cloud in the camera frame (cpp:196-221)
        sel = np.nonzero(r["vis"])[0]
        got = api.load_pcd_xyzi(out_dir / "filtered_pcd" / (fr["name"] + "_beforeNID.pcd"))
        want = np.concatenate([oracle.transform(cloud[sel, :3].copy(), w2c), cloud[sel, 3:4]], axis=1)
        assert got.shape == want.shape and np.array_equal(got, _f8(want))
        mpath = out_dir / "filtered_pcd" / (fr["name"] + "_rgb-mask.pcd")
        if fr["mask"] is None:
            assert not mpath.exists()
            continue
        # filtered_pcd/<ts>_rgb-mask.pcd (cpp:533-546, :783-815)
        csel = np.nonzero(r["pix"] >= 0)[0]
        pix = r["pix"][csel]
        pc = oracle.transform(cloud[csel, :3].copy(), w2c)
        bgr = adj.reshape(-1, 3)[pix].astype(np.uint32)
        mv = fr["mask"].reshape(-1)[pix].astype(np.uint32)
        red = mv == 255
        rr = np.where(red, 255, bgr[:, 2]); gg = np.where(red, 0, bgr[:, 1]); bb = np.where(red, 0, bgr[:, 0])
        packed = (255 << 24) | (rr << 16) | (gg << 8) | bb
        fields, rows = _read_ascii_pcd(mpath)
        assert fields == ["x", "y", "z", "rgb", "segmentMask"] and len(rows) == len(csel)
        arr = np.array(rows)
        assert np.array_equal(arr[:, :3].astype(np.float32), _f8(pc))
        assert np.array_equal(arr[:, 3].astype(np.uint64), packed.astype(np.uint64)) and np.array_equal(arr[:, 4].astype(np.uint32), mv)
        pw = oracle.transform(pc, c2w)
        world_mask_rows.append(np.concatenate([_f8(pw), packed[:, None].astype(np.float64), mv[:, None].astype(np.float64)], axis=1))

    # cloudInWorldWithRGB.pcd: smoothColors + removePointsWithNoColor (cpp:604-631, hpp:238-252)
    rgb_o = oc.finalize_top5()
    keep = np.nonzero(rgb_o.any(axis=1))[0]
    fields, rows = _read_ascii_pcd(out_dir / "cloudInWorldWithRGB.pcd")
    assert fields == ["x", "y", "z", "rgb"]
    arr = np.array(rows)
    assert st["n_colored_points"] == len(rows)
    assert len(rows) == len(keep)
    assert np.array_equal(arr[:, :3].astype(np.float32), _f8(cloud[keep, :3]))
    u = arr[:, 3].astype(np.uint64)
    got_rgb = np.stack([(u >> 16) & 255, (u >> 8) & 255, u & 255], axis=1).astype(int)
    assert np.abs(got_rgb - rgb_o[keep].astype(int)).max() <= 1  # TOP5 weighted mean in float: tolerance 1/255 per channel
    assert np.all((u >> 24) == 255)
    if masks:
        fields, rows = _read_ascii_pcd(out_dir / "cloudInWorldWithRGBandMask.pcd")
        want = np.concatenate(world_mask_rows, axis=0)
        arr = np.array(rows).astype(np.float64)
        assert arr.shape == want.shape
        assert np.array_equal(arr[:, :3].astype(np.float32), want[:, :3].astype(np.float32)) and np.array_equal(arr[:, 3:], want[:, 3:])
    # scans-crop.pcd: the trajectory's bounding box +- 2 m (cpp:94-135)
    P = np.array([fr["pose"][:3] for fr in frames] + [np.array(fr["pose"][:3]) + [0.03, 0, 0] for fr in frames])
    lo = np.minimum(P.min(0), 1e300) - 2.0
    hi = np.maximum(P.max(0), np.finfo(np.float64).tiny) + 2.0
    inside = np.all((cloud[:, :3] >= lo.astype(np.float32)) & (cloud[:, :3] <= hi.astype(np.float32)), axis=1)
    assert st["n_cropped"] == int(inside.sum())
    assert len(api.load_pcd_xyzi(out_dir / "scans-crop.pcd")) == int(inside.sum())


def test_process_jpeg_images(oracle, built_lib, tmp_path):
    """.jpg keyframes go through nvJPEG; everything after the decoded buffer is exact, the decoder itself is compared with cv2."""
    cv2 = pytest.importorskip("cv2")
    from pointcloudprocessor_b200 import api
    W, H = 320, 240
    sc, frames, img_dir, mask_dir, out_dir = _dataset(tmp_path, W=W, H=H, masks=False, jpeg=True)
    cam = api.create_camera("pinhole", sc["intr"], sc["dist"])
    st = api.PointCloudProcessor(tmp_path / "scans.pcd", tmp_path / "odom.txt", str(img_dir) + "/", "", str(out_dir) + "/", camera=cam,
                                 cull_size=(W, H)).process()
    oc = oracle.Colorizer(oracle.camera(sc["intr"], sc["dist"]), W, H, sc["cloud"], keep_obs=True)
    worst = 0
    for fr in frames:
        path = img_dir / (fr["name"] + ".jpg")
        dec = api.imread(path)
        ref = cv2.imread(str(path))
        assert dec.shape == ref.shape
        worst = max(worst, int(np.abs(dec.astype(int) - ref.astype(int)).max()))
        assert np.abs(dec.astype(int) - ref.astype(int)).mean() < 2.0  # nvJPEG vs libjpeg-turbo (chroma up-sampling, IDCT): third-party decoders, ~1 level apart on average
        oc.frame(fr["pose"], oracle.hsv_roundtrip(dec))
    rgb_o = oc.finalize_top5()
    keep = np.nonzero(rgb_o.any(axis=1))[0]
    fields, rows = _read_ascii_pcd(out_dir / "cloudInWorldWithRGB.pcd")
    assert len(rows) == len(keep) == st["n_colored_points"]
    u = np.array(rows)[:, 3].astype(np.uint64)
    got_rgb = np.stack([(u >> 16) & 255, (u >> 8) & 255, u & 255], axis=1).astype(int)
    assert np.abs(got_rgb - rgb_o[keep].astype(int)).max() <= 1
    print("nvJPEG vs cv2 max abs diff:", worst)


def test_process_errors(built_lib, tmp_path):
    from pointcloudprocessor_b200 import api
    (tmp_path / "odom.txt").write_text("1.0 0 0 0 1 0 0 0\n")
    with pytest.raises(api.PcpError, match="Couldn't read point cloud file"):
        api.PointCloudProcessor(tmp_path / "missing.pcd", tmp_path / "odom.txt", str(tmp_path) + "/", "", str(tmp_path) + "/").process()
    with pytest.raises(api.PcpError, match="enableMLS"):
        api.PointCloudProcessor(tmp_path / "missing.pcd", tmp_path / "odom.txt", str(tmp_path) + "/", "", str(tmp_path) + "/", enableMLS=True).process()


def test_cli_matches_library(built_lib, tmp_path):
    """pcp_colorize (src/main.cpp's options) writes the same files as the library call."""
    import subprocess
    from pointcloudprocessor_b200 import api
    W, H = 330, 240
    sc, frames, img_dir, mask_dir, out_dir = _dataset(tmp_path, W=W, H=H, masks=True)
    cam = api.create_camera("pinhole", sc["intr"], sc["dist"])
    api.PointCloudProcessor(tmp_path / "scans.pcd", tmp_path / "odom.txt", str(img_dir) + "/", str(mask_dir) + "/", str(out_dir) + "/", camera=cam,
                            cull_size=(W, H)).process()
    out2 = tmp_path / "out2"
    out2.mkdir()
    exe = os.path.join(os.path.dirname(built_lib), "pcp_colorize")
    r = subprocess.run([exe, "-p", str(tmp_path / "scans.pcd"), "-o", str(tmp_path / "odom.txt"), "-i", str(img_dir) + "/", "-m", str(mask_dir) + "/",
                        "-t", str(out2) + "/", "--cull_size", f"{W}x{H}", "--intrinsics", ",".join(repr(float(x)) for x in sc["intr"]),
                        "--distortion", ",".join(repr(float(x)) for x in sc["dist"])], capture_output=True, text=True)
    assert r.returncode == 0 and "Processing completed successfully." in r.stdout, r.stdout + r.stderr
    names = ["scans-crop.pcd", "cloudInWorldWithRGB.pcd", "cloudInWorldWithRGBandMask.pcd"] + ["filtered_pcd/" + p.name for p in (out_dir / "filtered_pcd").iterdir()]
    assert len(names) > 8
    for nme in names:
        assert (out_dir / nme).read_bytes() == (out2 / nme).read_bytes(), nme


def test_binary_output_and_device_compaction(built_lib, tmp_path):
    """cloudInWorldWithRGB.pcd as DATA binary (pcl::PCDWriter::writeBinary's layout of pcl::PointXYZRGB, opt-in) holds the points the
    ASCII file holds (positions to the 8 digits the text keeps, colours exactly), and pcp_finalize_compact == finalize + the host-side
    removePointsWithNoColor (hpp:238-252)."""
    from pointcloudprocessor_b200 import api
    W, H = 330, 240
    sc, frames, img_dir, mask_dir, out_dir = _dataset(tmp_path, W=W, H=H, masks=False)
    cam = api.create_camera("pinhole", sc["intr"], sc["dist"])
    args = (tmp_path / "scans.pcd", tmp_path / "odom.txt", str(img_dir) + "/", "")
    api.PointCloudProcessor(*args, str(out_dir) + "/", camera=cam, cull_size=(W, H)).process()
    out2 = tmp_path / "out_bin"
    out2.mkdir()
    st = api.PointCloudProcessor(*args, str(out2) + "/", camera=cam, cull_size=(W, H), binary_output=True).process()
    raw = (out2 / "cloudInWorldWithRGB.pcd").read_bytes()
    k = raw.index(b"DATA binary\n") + len(b"DATA binary\n")
    hdr = raw[:k].decode()
    assert "FIELDS x y z rgb\nSIZE 4 4 4 4\nTYPE F F F F\nCOUNT 1 1 1 1\n" in hdr and f"POINTS {st['n_colored_points']}\n" in hdr
    rec = np.frombuffer(raw[k:], np.dtype([("xyz", "<f4", 3), ("bgra", "u1", 4)]))
    assert len(rec) == st["n_colored_points"] > 100
    fields, rows = _read_ascii_pcd(out_dir / "cloudInWorldWithRGB.pcd")
    arr = np.array(rows)
    assert len(rows) == len(rec)
    assert np.array_equal(arr[:, :3].astype(np.float32), _f8(rec["xyz"]))
    u = arr[:, 3].astype(np.uint64)
    assert np.array_equal(np.stack([u & 255, (u >> 8) & 255, (u >> 16) & 255, u >> 24], axis=1).astype(np.uint8), rec["bgra"])
    # the compaction itself, against finalize + a host filter
    with api.Colorizer(cam, api.default_params(image_size=(W, H))) as gc:
        gc.set_cloud(sc["cloud"])
        for i, pose in enumerate(sc["poses"]):
            gc.submit_frame(pose, sc["images"][i])
        rgb, _ = gc.finalize()
        pts = gc.finalize_compact()
    keep = np.nonzero(rgb.any(axis=1))[0]
    assert len(pts) == len(keep) > 100
    assert np.array_equal(np.stack([pts["x"], pts["y"], pts["z"]], axis=1), sc["cloud"][keep, :3])
    assert np.array_equal(np.stack([pts["r"], pts["g"], pts["b"]], axis=1), rgb[keep]) and np.all(pts["a"] == 255)
