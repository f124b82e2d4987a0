"""Host-side file formats of the drop-in boundary: PCD read (ascii / binary / binary_compressed) and ASCII write in
PCL's layout, PNG / PGM / PPM decoding against cv2.  No GPU needed: these entry points of libpcp_b200.so are pure host code."""
import struct

import numpy as np
import pytest


@pytest.fixture(scope="module")
def api(built_lib):
    from pointcloudprocessor_b200 import api as a
    return a


def _cloud(n=257, seed=3):
    rng = np.random.default_rng(seed)
    c = rng.normal(0, 10, (n, 4)).astype(np.float32)
    c[:, 3] = rng.random(n, dtype=np.float32)
    return c


def test_ascii_pcd_roundtrip_and_header(api, tmp_path):
    c = _cloud()
    c[5, 0] = np.nan
    p = tmp_path / "a.pcd"
    api.save_pcd_xyzi_ascii(p, c)
    lines = p.read_text().splitlines()
    assert lines[:11] == ["# .PCD v0.7 - Point Cloud Data file format", "VERSION 0.7", "FIELDS x y z intensity", "SIZE 4 4 4 4", "TYPE F F F F",
                          "COUNT 1 1 1 1", f"WIDTH {len(c)}", "HEIGHT 1", "VIEWPOINT 0 0 0 1 0 0 0", f"POINTS {len(c)}", "DATA ascii"]
    assert lines[11] == " ".join("%.8g" % float(v) for v in c[0])  # pcl::PCDWriter::writeASCII, precision 8
    back = api.load_pcd_xyzi(p)
    want = np.array([[np.float32(float("%.8g" % float(v))) for v in row] for row in c], np.float32)
    assert np.array_equal(back, want, equal_nan=True)


def _binary_pcd(path, c, compressed=False):
    """x y z intensity with an extra u16 'ring' field in the middle and a double 'time' field (field order shuffled)."""
    n = len(c)
    hdr = ("# .PCD v0.7 - Point Cloud Data file format\nVERSION 0.7\nFIELDS y ring x time intensity z\nSIZE 4 2 4 8 4 4\nTYPE F U F F F F\n"
           f"COUNT 1 1 1 1 1 1\nWIDTH {n}\nHEIGHT 1\nVIEWPOINT 0 0 0 1 0 0 0\nPOINTS {n}\nDATA {'binary_compressed' if compressed else 'binary'}\n")
    ring = (np.arange(n) % 64).astype(np.uint16)
    t = np.linspace(0, 1, n)
    if not compressed:
        body = b"".join(struct.pack("<fHfdff", c[i, 1], ring[i], c[i, 0], t[i], c[i, 3], c[i, 2]) for i in range(n))
    else:
        soa = c[:, 1].tobytes() + ring.tobytes() + c[:, 0].tobytes() + t.tobytes() + c[:, 3].tobytes() + c[:, 2].tobytes()
        # a valid LZF stream: literal runs of <= 32 bytes, plus one back-reference (the last 8 bytes repeat the 8 before them)
        soa = soa[:-8] + soa[-16:-8]
        comp = bytearray()
        lit = soa[:-8]
        for o in range(0, len(lit), 32):
            run = lit[o:o + 32]
            comp.append(len(run) - 1)
            comp += run
        comp += bytes([((8 - 2) << 5) | 0, 8 - 1])  # len 8, offset 8
        body = struct.pack("<II", len(comp), len(soa)) + bytes(comp)
        z = c[:, 2].copy()
        z[-2:] = np.frombuffer(soa[-8:], np.float32)
        c = c.copy()
        c[:, 2] = z
    with open(path, "wb") as f:
        f.write(hdr.encode() + body)
    return c


def test_binary_pcd_with_extra_fields(api, tmp_path):
    c = _cloud(1000)
    _binary_pcd(tmp_path / "b.pcd", c)
    assert np.array_equal(api.load_pcd_xyzi(tmp_path / "b.pcd"), c)


def test_binary_compressed_pcd(api, tmp_path):
    c = _cloud(333)
    want = _binary_pcd(tmp_path / "c.pcd", c, compressed=True)
    assert np.array_equal(api.load_pcd_xyzi(tmp_path / "c.pcd"), want)


def test_missing_pcd_raises_reference_message(api, tmp_path):
    with pytest.raises(api.PcpError, match="Couldn't read point cloud file"):
        api.load_pcd_xyzi(tmp_path / "nope.pcd")


def test_png_pgm_ppm_against_cv2(api, tmp_path):
    cv2 = pytest.importorskip("cv2")
    rng = np.random.default_rng(1)
    grey = rng.integers(0, 256, (37, 53), dtype=np.uint8)
    bgr = rng.integers(0, 256, (37, 53, 3), dtype=np.uint8)
    smooth = np.add.outer(np.arange(64), np.arange(96)).astype(np.uint8)  # exercises the Sub / Up / Paeth filters
    for name, img in (("g.png", grey), ("c.png", bgr), ("s.png", smooth), ("g.pgm", grey), ("c.ppm", bgr)):
        path = str(tmp_path / name)
        assert cv2.imwrite(path, img)
        for grayscale in (False, True):
            got = api.imread(path, grayscale)
            want = cv2.imread(path, cv2.IMREAD_GRAYSCALE if grayscale else cv2.IMREAD_COLOR)
            assert got.shape == want.shape, (name, grayscale)
            assert np.array_equal(got, want), (name, grayscale)
    rgba = np.dstack([bgr, rng.integers(0, 256, (37, 53), dtype=np.uint8)])
    assert cv2.imwrite(str(tmp_path / "a.png"), rgba)
    assert np.array_equal(api.imread(tmp_path / "a.png"), cv2.imread(str(tmp_path / "a.png")))


def test_cli_exit_codes(built_lib, tmp_path):
    """pcp_colorize keeps the exit codes of the reference's main (src/main.cpp:28-68): 1 help, -1 missing arguments, -2 exception."""
    import os
    import subprocess
    exe = os.path.join(os.path.dirname(built_lib), "pcp_colorize")
    assert os.path.exists(exe)
    assert subprocess.run([exe, "--help"], capture_output=True).returncode == 1
    r = subprocess.run([exe, "-p", "x.pcd"], capture_output=True, text=True)
    assert r.returncode == 255 and "Missing required arguments" in r.stderr
    r = subprocess.run([exe, "-p", str(tmp_path / "no.pcd"), "-o", str(tmp_path / "no.txt"), "-i", str(tmp_path) + "/"], capture_output=True, text=True)
    assert r.returncode == 254 and "Couldn't read point cloud file." in r.stderr
