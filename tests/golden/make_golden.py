#!/usr/bin/env python
"""Generates the committed golden fixtures under tests/golden/ from the CPU oracle (oracle/pcp_oracle.cpp).

    python tests/golden/make_golden.py

The reference ships no golden vectors and cannot be built here, so these fixtures pin the ORACLE (and
through it the CUDA path), not the reference: parity stays "unpinned" in the sense of the task statement.  They protect
against silent drift of either implementation and travel to the GPU box, where /root/reference does not exist.

Inputs are regenerated from seeds (pointcloudprocessor_b200.synth); each fixture stores a SHA-256 of every input array
so that a generator drift is reported as such, plus the outputs:
  colorize_*.npz : per frame packed visibility bits, pixel indices, counts; final u32 accumulators, mask accumulators,
                   SUM colours and TOP5 colours
  edge_*.npz     : hand-made edge-case cloud (stored explicitly) with its outputs
  nid_*.npz      : NID values, ok flags and joint histograms for a set of candidate poses
  distmask_*.npz : per frame the filled pixels of the distance / 3-D point masks with winner index, range and point
"""
from __future__ import annotations

import hashlib
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
if HERE not in sys.path:
    sys.path.insert(0, HERE)

from oracle import pyoracle as O  # noqa: E402
from pointcloudprocessor_b200 import synth  # noqa: E402


def sha(a: np.ndarray) -> str:
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def colorize_case(name, n, f, W, H, masks=False, cull_size=None, T_opt=None, ds=14, tol=0.05):
    intr, dist = synth.make_camera(W, H)
    cloud = synth.make_cloud(n)
    poses, _ = synth.make_poses(f)
    cam = O.camera(intr, dist)
    oc = O.Colorizer(cam, W, H, cloud, cull_size=cull_size, ds=ds, tol=tol, keep_obs=True)
    out = {"n": n, "f": f, "W": W, "H": H, "masks": int(masks), "ds": ds, "tol": tol,
           "cull_size": np.array(cull_size if cull_size else (W, H)),
           "T_opt": np.asarray(T_opt, np.float64) if T_opt is not None else np.zeros((0, 0)),
           "sha_cloud": sha(cloud), "sha_poses": sha(poses)}
    img_sha, counts = [], []
    for i in range(f):
        img = synth.make_image(i, W, H)
        mk = synth.make_mask(i, W, H) if masks else None
        img_sha.append(sha(img) + (sha(mk) if masks else ""))
        r = oc.frame(poses[i], img, mask=mk, T_opt=T_opt, want_vis=True, want_pix=True)
        out[f"vis_{i}"] = np.packbits(r["vis"], bitorder="little")
        out[f"pix_{i}"] = r["pix"]
        counts.append((r["n_candidates"], r["n_visible"], r["n_colored"]))
    out["sha_images"] = np.array(img_sha)
    out["counts"] = np.array(counts, np.int64)
    acc, mh, mm = oc.accumulators()
    out["acc"] = acc
    out["mask_hits"] = mh
    out["mask_max"] = mm
    out["rgb_sum"] = oc.finalize_sum()
    out["rgb_top5"] = oc.finalize_top5()
    np.savez_compressed(os.path.join(HERE, f"colorize_{name}.npz"), **out)
    print(name, "visible per frame:", [c[1] for c in counts])


def edge_cloud(W, H, intr):
    rng = np.random.default_rng(7)
    pts = []
    u = rng.uniform(-30, W + 30, 4000)
    v = rng.uniform(-30, H + 30, 4000)
    z = rng.uniform(0.5, 6.0, 4000)
    pts.append(np.stack([(u - intr[2]) / intr[0] * z, (v - intr[3]) / intr[1] * z, z], 1))
    pts.append(np.array([[0, 0, 0], [1, 1, 0], [0.1, 0.1, -1], [0, 0, 1], [0, 0, 1], [0, 0, 1.00000012], [np.nan, 0, 1], [0, np.inf, 1],
                         [0, 0, np.inf], [1e30, 1e30, 1e-30], [-1e-3, 0, 1e-38], [3e38, 3e38, 3e38], [np.inf, 0, -5], [0, 0, -np.inf]], np.float64))
    pts.append(np.tile(np.array([[0.2, 0.1, 2.0]]), (5, 1)))
    with np.errstate(over="ignore"):
        xyz = np.concatenate(pts).astype(np.float32)
    return np.concatenate([xyz, rng.random((xyz.shape[0], 1), dtype=np.float32)], 1)


def edge_case():
    W, H = 320, 240
    intr, dist = synth.make_camera(W, H)
    cloud = edge_cloud(W, H, intr)
    cam = O.camera(intr, dist)
    pose = np.array([0, 0, 0, 1, 0, 0, 0], np.float64)
    img = synth.make_image(0, W, H)
    oc = O.Colorizer(cam, W, H, cloud)
    r = oc.frame(pose, img, want_vis=True, want_pix=True)
    np.savez_compressed(os.path.join(HERE, "edge_points.npz"), cloud=cloud, pose=pose, W=W, H=H, sha_image=sha(img),
                        vis=np.packbits(r["vis"], bitorder="little"), pix=r["pix"],
                        counts=np.array([r["n_candidates"], r["n_visible"], r["n_colored"]]), acc=oc.accumulators()[0])
    print("edge: visible", r["n_visible"], "candidates", r["n_candidates"])


def nid_case():
    W, H = 320, 240
    n = 30_000
    intr, dist = synth.make_camera(W, H)
    cloud = synth.make_cloud(n)
    poses, _ = synth.make_poses(2)
    cam = O.camera(intr, dist)
    w2c, _ = O.pose_matrices(poses[1])
    pts_cam = O.transform(cloud, w2c)
    keep = O.view_culling(cam, W, H, pts_cam)
    culled = np.concatenate([pts_cam[keep], cloud[keep, 3:]], 1)
    img = synth.make_image(1, W, H)
    grey = synth.grey_plane(img)
    cands = synth.make_nid_candidates(12)
    # two candidates far outside the image / behind the camera exercise the outlier and non-finite branches
    far = cands[1].copy(); far[4:] = (50.0, 0.0, 0.0)
    cands = np.concatenate([cands, far[None]])
    vals, oks, hists = [], [], []
    for c in cands:
        v, ok, h, cnt = O.nid(cam, grey, culled, c, want_hist=True)
        vals.append(v); oks.append(ok); hists.append(np.concatenate([h.reshape(-1), cnt]))
    np.savez_compressed(os.path.join(HERE, "nid_tiny.npz"), W=W, H=H, n=n, sha_cloud=sha(cloud), sha_image=sha(img), keep=keep.astype(np.int32),
                        cands=cands, nid=np.array(vals), ok=np.array(oks), hist=np.array(hists))
    print("nid:", np.array(vals)[:4], "ok", np.array(oks).astype(int))


def distmask_case():
    """generate_distance_masks (scripts/genNormAndDistanceMask.py:233-265) on the culled cloud of each frame; stored sparse.

    The masks are produced by the REFERENCE'S OWN SOURCE (tests/golden/ref_extract.py cuts the method out of the module with `ast`
    and runs it): this fixture pins the oracle (and through it the CUDA path) to the reference, not to a restatement.  The winner
    index per pixel is an extension the reference does not compute; it comes from the oracle and is checked to describe the same
    filled pixels."""
    import ref_extract as R
    if not R.available():
        raise SystemExit("distmask fixture: /root/reference is not present; this fixture can only be generated next to the reference tree")
    n, f, W, H = 60_000, 3, 640, 480
    intr, dist = synth.make_camera(W, H)
    cloud = synth.make_cloud(n)
    poses, _ = synth.make_poses(f)
    cam = O.camera(intr, dist)
    out = {"n": n, "f": f, "W": W, "H": H, "sha_cloud": sha(cloud), "sha_poses": sha(poses)}
    for i in range(f):
        w2c, _ = O.pose_matrices(poses[i])
        pts = O.transform(cloud[:, :3], w2c)
        keep = O.view_culling(cam, W, H, pts)
        rd, rxyz, lines = R.reference_distance_masks(pts[keep], W, H, intr, dist)   # the reference's statements
        d, xyz, idx = O.distance_masks(cam, W, H, pts[keep])                          # the restatement
        pix = np.flatnonzero(idx.reshape(-1) >= 0).astype(np.int32)
        assert np.array_equal(np.flatnonzero(rd.reshape(-1) > 0), pix), "oracle and reference fill different pixels"
        assert np.array_equal(rd, d) and np.array_equal(rxyz, xyz), "oracle differs from the reference's generate_distance_masks"
        out[f"pix_{i}"] = pix                                            # filled pixels, ascending
        out[f"index_{i}"] = keep[idx.reshape(-1)[pix]].astype(np.int32)  # winning point, as map index (extension, from the oracle)
        out[f"dist_{i}"] = rd.reshape(-1)[pix]                           # from the reference
        out[f"xyz_{i}"] = rxyz.reshape(-1, 3)[pix]                       # from the reference
        print("distmask frame", i, "culled", len(keep), "pixels filled", len(pix), "reference lines", lines)
    out["source"] = np.array(f"dist_*/xyz_* computed by {R.REFERENCE_SCRIPT}:{lines[0]}-{lines[1]} (method cut out with ast and executed); "
                             "pix_*/index_* from oracle/pcp_oracle.cpp orc_distance_masks, checked to fill the same pixels")
    np.savez_compressed(os.path.join(HERE, "distmask_tiny.npz"), **out)


if __name__ == "__main__":
    O.build()
    if "--only-distmask" in sys.argv:
        distmask_case()
        sys.exit(0)
    colorize_case("tiny", 20_000, 3, 320, 240)
    colorize_case("tiny_masks", 20_000, 3, 320, 240, masks=True)
    T = np.eye(4)
    a = np.deg2rad(0.7)
    T[:3, :3] = np.array([[np.cos(a), -np.sin(a), 0], [np.sin(a), np.cos(a), 0], [0, 0, 1]])
    T[:3, 3] = (0.03, -0.02, 0.05)
    colorize_case("tiny_extrinsic", 20_000, 2, 320, 240, T_opt=T)
    colorize_case("refcull", 30_000, 2, 640, 480, cull_size=(4096, 3000))
    edge_case()
    nid_case()
    distmask_case()
