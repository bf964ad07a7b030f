#!/usr/bin/env python
"""How selective are cheap pre-tests of FAST-9 on the bench's synthetic frames and on the bundled TUM frame?  (design aid for k_fast:
fraction of pixels / aligned 4-pixel words / 128-pixel warp rows that survive the compass test, and that are corners)"""
import importlib
import sys

import numpy as np

sys.path.insert(0, ".")
synth = importlib.import_module("modular-slam_b200.synth")

OFF = [(0, 3), (1, 3), (2, 2), (3, 1), (3, 0), (3, -1), (2, -2), (1, -3), (0, -3), (-1, -3), (-2, -2), (-3, -1), (-3, 0), (-3, 1), (-2, 2), (-1, 3)]


def gray(bgr):
    f = bgr.astype(np.float32)
    return np.minimum(255.0, f[..., 0] * np.float32(.299) + f[..., 1] * np.float32(.587) + f[..., 2] * np.float32(.114)).astype(np.uint8)


def stats(g, t=20):
    h, w = g.shape
    c = g[3:h - 3, 3:w - 3].astype(np.int16)
    br, dk = [], []
    for dx, dy in OFF:
        r = g[3 + dy:h - 3 + dy, 3 + dx:w - 3 + dx].astype(np.int16)
        br.append(r > c + t)
        dk.append(r < c - t)
    br, dk = np.array(br), np.array(dk)

    def run9(m):
        out = np.zeros(m.shape[1:], bool)
        for s in range(16):
            a = np.ones(m.shape[1:], bool)
            for k in range(9):
                a &= m[(s + k) & 15]
            out |= a
        return out

    def compass(m):
        return (m[0] & m[4]) | (m[4] & m[8]) | (m[8] & m[12]) | (m[12] & m[0])

    def opp(m):  # weaker: (0 or 8) and (4 or 12)
        return (m[0] | m[8]) & (m[4] | m[12])

    corner = run9(br) | run9(dk)
    cand = compass(br) | compass(dk)
    cand2 = opp(br) | opp(dk)
    ww = (corner.shape[1] // 4) * 4
    res = {}
    for name, m in (("corner", corner), ("compass", cand), ("opposite", cand2)):
        words = m[:, :ww].reshape(m.shape[0], -1, 4).any(-1)
        w128 = (m.shape[1] // 128) * 128
        rows = m[:, :w128].reshape(m.shape[0], -1, 128).any(-1)
        res[name] = (round(float(m.mean()), 4), round(float(words.mean()), 4), round(float(rows.mean()), 4))
    return res


if __name__ == "__main__":
    fr, _ = synth.frame_batch(640, 480, 2, 1)
    print("synthetic 640x480 (px, 4-px words, 128-px rows):", stats(gray(fr[0])))
    try:
        import cv2
        img = cv2.imread("tests/golden/tum/rgb/0000.png")
        print("bundled TUM frame:", stats(gray(img)))
    except Exception as e:  # noqa: BLE001
        print("no cv2 / frame:", e)
