"""RgbdFileB200Provider (modular-slam_b200/host/rgbd_file_b200_provider.*) against the reference provider's
contract (src/lib/modular_slam/rgbd_file_provider.cpp:42-145): RgbFrame::data == cv::imread(path) bytes (B, G, R),
DepthFrame::data == cv::imread(path, IMREAD_ANYDEPTH) u16, TUM association parsing, end-of-sequence behaviour.

Golden pixels: tests/golden/bundled.npz / depth.npz hold what cv2.imread returned for the reference's bundled
data/rgb/0000.png and data/depth/0000.png (tests/golden/make_golden.py); the two PNG files themselves are committed under
tests/golden/tum/ (data fixtures, byte-identical copies), so the decoder is exercised on real libpng output.  Synthetic
PNGs written here cover every row filter, colour type and bit depth the decoder accepts.  Nothing reads /root/reference."""
import os
import struct
import subprocess
import zlib

import numpy as np
import pytest

import __graft_entry__ as graft
from conftest import GOLDEN

LIB = os.path.join(graft.PKG_DIR, "lib")
DRIVER = os.path.join(LIB, "plugin_host_test")


@pytest.fixture(scope="module")
def built():
    if not os.path.exists(DRIVER):
        graft.build()
    return DRIVER


def _paeth(a, b, c):
    p = a + b - c
    pa, pb, pc = abs(p - a), abs(p - b), abs(p - c)
    return a if (pa <= pb and pa <= pc) else (b if pb <= pc else c)


def write_png(path, rows, width, height, bit_depth, color_type, filters, palette=None, idat_split=1):
    """rows: [height] bytes objects of unfiltered scanlines; filters: per-row filter type"""
    ch = {0: 1, 2: 3, 3: 1, 4: 2, 6: 4}[color_type]
    bpp = max(1, ch * bit_depth // 8)
    raw = bytearray()
    prev = bytes(len(rows[0]))
    for y, row in enumerate(rows):
        f = filters[y % len(filters)]
        out = bytearray(len(row))
        for i, v in enumerate(row):
            a = row[i - bpp] if i >= bpp else 0
            b = prev[i]
            c = prev[i - bpp] if i >= bpp else 0
            pred = (0, a, b, (a + b) >> 1, _paeth(a, b, c))[f]
            out[i] = (v - pred) & 255
        raw.append(f)
        raw += out
        prev = row

    def chunk(t, d):
        return struct.pack(">I", len(d)) + t + d + struct.pack(">I", zlib.crc32(t + d) & 0xffffffff)
    z = zlib.compress(bytes(raw), 6)
    parts = [z[i * len(z) // idat_split:(i + 1) * len(z) // idat_split] for i in range(idat_split)]
    with open(path, "wb") as fh:
        fh.write(b"\x89PNG\r\n\x1a\n")
        fh.write(chunk(b"IHDR", struct.pack(">IIBBBBB", width, height, bit_depth, color_type, 0, 0, 0)))
        if palette is not None:
            fh.write(chunk(b"PLTE", bytes(palette)))
        for p in parts:
            fh.write(chunk(b"IDAT", p))
        fh.write(chunk(b"IEND", b""))


def run_provider(built, assoc, tmp_path, batch=0):
    outp = tmp_path / f"prov{batch}.bin"
    out = subprocess.run([built, "provider", str(assoc), str(outp)] + ([str(batch)] if batch else []), capture_output=True, text=True)
    assert out.returncode == 0, out.stdout + out.stderr
    raw = outp.read_bytes()
    frames, pos = [], 0
    while pos < len(raw):
        w, h, ts, factor = np.frombuffer(raw, np.float64, 4, pos)
        w, h = int(w), int(h)
        pos += 32
        rgb = np.frombuffer(raw, np.uint8, w * h * 3, pos).reshape(h, w, 3)
        pos += w * h * 3
        depth = np.frombuffer(raw, np.uint16, w * h, pos).reshape(h, w)
        pos += w * h * 2
        frames.append((ts, factor, rgb, depth))
    return frames, out.stdout


def test_bundled_png_files_decode_to_what_cv2_imread_returned(built, tmp_path, bundled):
    depth = np.load(os.path.join(GOLDEN, "depth.npz"))
    tum = os.path.join(GOLDEN, "tum")
    assoc = tmp_path / "assoc.txt"
    # relative paths resolve against the association file's directory (rgbd_file_provider.cpp:114,128-129)
    os.symlink(os.path.join(tum, "rgb"), tmp_path / "rgb")
    os.symlink(os.path.join(tum, "depth"), tmp_path / "depth")
    assoc.write_text("# comment lines fail the parse and are skipped\n"
                     "1305031102.175304 rgb/0000.png 1305031102.160407 depth/0000.png\n"
                     "\n"
                     f"1305031102.211214 {tum}/rgb/0000.png 1305031102.194330 {tum}/depth/0000.png\n")
    frames, text = run_provider(built, assoc, tmp_path)
    assert "frames=2 position=2" in text
    for ts, factor, rgb, d in frames:
        assert (rgb == bundled["bgr0"]).all()
        assert (d == depth["depth0"]).all()
        assert np.float32(factor) == np.float32(1.0) / np.float32(5000.0)  # tumRgbdCameraParams (:136-145)
    assert frames[0][0] == 1305031102.175304 and frames[1][0] == 1305031102.211214


def test_every_filter_colour_type_and_bit_depth(built, tmp_path):
    rng = np.random.default_rng(4)
    W, H = 37, 23
    (tmp_path / "d").mkdir()
    cases = []
    # depth map shared by every case: 16-bit gray, all five filters in rotation, IDAT split in three
    dz = rng.integers(0, 65536, (H, W)).astype(np.uint16)
    write_png(tmp_path / "d" / "depth.png", [dz[y].astype(">u2").tobytes() for y in range(H)], W, H, 16, 0, [0, 1, 2, 3, 4], idat_split=3)
    rgb8 = rng.integers(0, 256, (H, W, 3), dtype=np.uint8)
    for f in range(5):
        write_png(tmp_path / "d" / f"rgb_f{f}.png", [rgb8[y].tobytes() for y in range(H)], W, H, 8, 2, [f])
        cases.append((f"rgb_f{f}.png", rgb8[..., ::-1]))
    rgba = rng.integers(0, 256, (H, W, 4), dtype=np.uint8)
    write_png(tmp_path / "d" / "rgba.png", [rgba[y].tobytes() for y in range(H)], W, H, 8, 6, [4, 3, 1])
    cases.append(("rgba.png", rgba[..., 2::-1]))
    rgb16 = rng.integers(0, 65536, (H, W, 3)).astype(np.uint16)
    write_png(tmp_path / "d" / "rgb16.png", [rgb16[y].astype(">u2").tobytes() for y in range(H)], W, H, 16, 2, [1, 4])
    cases.append(("rgb16.png", (rgb16 >> 8).astype(np.uint8)[..., ::-1]))  # png_set_strip_16
    g8 = rng.integers(0, 256, (H, W), dtype=np.uint8)
    write_png(tmp_path / "d" / "gray8.png", [g8[y].tobytes() for y in range(H)], W, H, 8, 0, [2, 3])
    cases.append(("gray8.png", np.repeat(g8[:, :, None], 3, 2)))
    ga = rng.integers(0, 256, (H, W, 2), dtype=np.uint8)
    write_png(tmp_path / "d" / "graya.png", [ga[y].tobytes() for y in range(H)], W, H, 8, 4, [4])
    cases.append(("graya.png", np.repeat(ga[:, :, :1], 3, 2)))
    g4 = rng.integers(0, 16, (H, W), dtype=np.uint8)
    packed = [bytes(((int(g4[y, x]) << 4) | (int(g4[y, x + 1]) if x + 1 < W else 0)) for x in range(0, W, 2)) for y in range(H)]
    write_png(tmp_path / "d" / "gray4.png", packed, W, H, 4, 0, [0, 1, 2])
    cases.append(("gray4.png", np.repeat((g4 * 17)[:, :, None], 3, 2).astype(np.uint8)))
    pal = rng.integers(0, 256, (256, 3), dtype=np.uint8)
    idx = rng.integers(0, 256, (H, W), dtype=np.uint8)
    write_png(tmp_path / "d" / "pal8.png", [idx[y].tobytes() for y in range(H)], W, H, 8, 3, [0, 4], palette=pal.reshape(-1))
    cases.append(("pal8.png", pal[idx][..., ::-1]))
    assoc = tmp_path / "d" / "assoc.txt"
    assoc.write_text("".join(f"{i}.5 {name} {i}.25 depth.png\n" for i, (name, _) in enumerate(cases)))
    frames, text = run_provider(built, assoc, tmp_path)
    assert f"frames={len(cases)}" in text
    for i, ((name, want), (ts, _, rgb, d)) in enumerate(zip(cases, frames)):
        assert ts == i + 0.5
        assert (rgb == want).all(), name
        assert (d == dz).all(), name


def test_unreadable_frames_end_the_sequence_like_the_reference(built, tmp_path):
    (tmp_path / "e").mkdir()
    g = np.zeros((8, 8, 3), np.uint8)
    z = np.arange(64, dtype=np.uint16).reshape(8, 8)
    write_png(tmp_path / "e" / "a.png", [g[y].tobytes() for y in range(8)], 8, 8, 8, 2, [0])
    write_png(tmp_path / "e" / "z.png", [z[y].astype(">u2").tobytes() for y in range(8)], 8, 8, 16, 0, [0])
    (tmp_path / "e" / "broken.png").write_bytes(b"\x89PNG\r\n\x1a\n" + b"garbage" * 20)
    assoc = tmp_path / "e" / "assoc.txt"
    assoc.write_text("1 a.png 1 z.png\n2 broken.png 2 z.png\n3 a.png 3 z.png\n")
    frames, text = run_provider(built, assoc, tmp_path)
    assert "frames=1 position=1" in text  # fetch() returns false at the unreadable frame and does not advance (:69-73)
    assoc.write_text("")  # no frames: init() is false (:50-53)
    out = subprocess.run([built, "provider", str(assoc), str(tmp_path / "x.bin")], capture_output=True, text=True)
    assert out.returncode == 7 and "init failed" in out.stdout


@pytest.mark.gpu
def test_batched_fetch_into_pinned_memory_feeds_the_detector(built, tmp_path, bundled, pkg, oracle):
    depth = np.load(os.path.join(GOLDEN, "depth.npz"))
    tum = os.path.join(GOLDEN, "tum")
    assoc = tmp_path / "assoc.txt"
    assoc.write_text("".join(f"{i}.0 {tum}/rgb/0000.png {i}.0 {tum}/depth/0000.png\n" for i in range(5)))
    frames, text = run_provider(built, assoc, tmp_path, batch=3)  # 3 + 2 frames, decoded on a thread pool into pinned memory
    assert "frames=5 position=5" in text
    for i, (ts, _, rgb, d) in enumerate(frames):
        assert ts == float(i) and (rgb == bundled["bgr0"]).all() and (d == depth["depth0"]).all()
    # and the decoded frame drives the detector to the oracle's answer (configs[0]: bundled data through the provider)
    with pkg.Context(max_width=640, max_height=480, nfeatures=1000) as ctx:
        kp, desc = ctx.detect(frames[0][2])
    okp, odesc = oracle.orb_detect_rgb(bundled["bgr0"])
    assert kp.tobytes() == okp.tobytes() and (desc == odesc).all()
