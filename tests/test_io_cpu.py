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


def _write_png(path, w, h, ctype, depth, samples, interlace=0, palette=None, filter_type=None):
    """Hand-made PNG (zlib + struct): `samples` = (h, w, spp) integer array of `depth`-bit values."""
    import struct
    import zlib
    spp = {0: 1, 2: 3, 3: 1, 4: 2, 6: 4}[ctype]
    samples = np.asarray(samples).reshape(h, w, spp)

    def chunk(t, d):
        return struct.pack(">I", len(d)) + t + d + struct.pack(">I", zlib.crc32(t + d) & 0xffffffff)

    def pack_rows(sub):  # sub: (ph, pw, spp) -> filtered scanlines
        ph, pw = sub.shape[:2]
        out = bytearray()
        prev = None
        bpp = max(1, spp * depth // 8)
        for y in range(ph):
            row = sub[y].reshape(-1)
            if depth == 8:
                b = bytes(row.astype(np.uint8))
            elif depth == 16:
                b = row.astype(">u2").tobytes()
            else:
                bits = "".join(format(int(v), f"0{depth}b") for v in row)
                bits += "0" * (-len(bits) % 8)
                b = bytes(int(bits[i:i + 8], 2) for i in range(0, len(bits), 8))
            ft = (y % 5) if filter_type is None else filter_type
            cur = np.frombuffer(b, np.uint8).astype(int)
            up = np.zeros_like(cur) if prev is None else prev
            left = np.concatenate([np.zeros(bpp, int), cur[:-bpp]]) if len(cur) > bpp else np.zeros_like(cur)
            ul = np.concatenate([np.zeros(bpp, int), up[:-bpp]]) if len(cur) > bpp else np.zeros_like(cur)
            if ft == 0: f = cur
            elif ft == 1: f = cur - left
            elif ft == 2: f = cur - up
            elif ft == 3: f = cur - ((left + up) >> 1)
            else:
                pp = left + up - ul
                pa, pb, pc = np.abs(pp - left), np.abs(pp - up), np.abs(pp - ul)
                pred = np.where((pa <= pb) & (pa <= pc), left, np.where(pb <= pc, up, ul))
                f = cur - pred
            out.append(ft)
            out += bytes((f % 256).astype(np.uint8))
            prev = cur
        return bytes(out)

    if interlace:
        passes = [(0, 0, 8, 8), (4, 0, 8, 8), (0, 4, 4, 8), (2, 0, 4, 4), (0, 2, 2, 4), (1, 0, 2, 2), (0, 1, 1, 2)]
        raw = b"".join(pack_rows(samples[y0::dy, x0::dx]) for x0, y0, dx, dy in passes if samples[y0::dy, x0::dx].size)
    else:
        raw = pack_rows(samples)
    data = b"\x89PNG\r\n\x1a\n" + chunk(b"IHDR", struct.pack(">IIBBBBB", w, h, depth, ctype, 0, 0, interlace))
    if palette is not None:
        data += chunk(b"PLTE", bytes(np.asarray(palette, np.uint8).reshape(-1)))
    data += chunk(b"IDAT", zlib.compress(raw)) + chunk(b"IEND", b"")
    open(path, "wb").write(data)


def test_png_every_bit_depth_colour_type_and_adam7_against_cv2(api, tmp_path):
    """cv::imread reads every PNG flavour (segmentation masks are often 1-bit or palette files): bit depths 1 / 2 / 4 / 8 / 16, colour
    types grey / RGB / palette / grey+alpha / RGBA, non-interlaced and Adam7, all five row filters.  The files are written by hand so
    that flavours cv2.imwrite cannot produce are covered; cv2.imread (libpng) is the reference reader."""
    cv2 = pytest.importorskip("cv2")
    rng = np.random.default_rng(7)
    w, h = 37, 21
    pal = rng.integers(0, 256, (256, 3), dtype=np.uint8)
    cases = []
    for depth in (1, 2, 4, 8, 16):
        cases.append((0, depth))
    for depth in (8, 16):
        cases += [(2, depth), (4, depth), (6, depth)]
    for depth in (1, 2, 4, 8):
        cases.append((3, depth))
    k = 0
    for ctype, depth in cases:
        spp = {0: 1, 2: 3, 3: 1, 4: 2, 6: 4}[ctype]
        smp = rng.integers(0, 1 << depth, (h, w, spp))
        for interlace in (0, 1):
            path = str(tmp_path / f"p{k}.png")
            k += 1
            _write_png(path, w, h, ctype, depth, smp, interlace, palette=pal[: 1 << depth] if ctype == 3 else None)
            for grayscale in (False, True):
                want = cv2.imread(path, cv2.IMREAD_GRAYSCALE if grayscale else cv2.IMREAD_COLOR)
                assert want is not None, (ctype, depth, interlace)
                got = api.imread(path, grayscale)
                assert got.shape == want.shape, (ctype, depth, interlace, grayscale)
                assert np.array_equal(got, want), (ctype, depth, interlace, grayscale, int((got != want).sum()))
    # cv2.imwrite's own bilevel writer, as a cross-check of the hand-made files
    bi = (rng.integers(0, 2, (33, 50)) * 255).astype(np.uint8)
    assert cv2.imwrite(str(tmp_path / "bi.png"), bi, [cv2.IMWRITE_PNG_BILEVEL, 1])
    assert np.array_equal(api.imread(tmp_path / "bi.png", True), cv2.imread(str(tmp_path / "bi.png"), cv2.IMREAD_GRAYSCALE))


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
