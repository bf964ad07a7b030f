"""Seeded synthetic inputs of SURVEY.md section 8(d): textured frames with true inter-frame matches, and
descriptor databases with planted near-duplicates.  numpy only (runs on the GPU box without the reference)."""
import numpy as np


def _box_blur3(img):
    p = np.pad(img.astype(np.float32), 1, mode="edge")
    acc = np.zeros_like(img, np.float32)
    for dy in range(3):
        for dx in range(3):
            acc += p[dy:dy + img.shape[0], dx:dx + img.shape[1]]
    return acc / 9.0


def base_texture(width, height, seed, margin=96, n_shapes=200):
    """gray texture (float32) of size (height + margin, width + margin): rectangles / discs of random gray level
    plus lightly blurred noise, so that FAST finds >> 2 n_l corners on every level."""
    rng = np.random.default_rng(seed)
    H, W = height + margin, width + margin
    img = np.full((H, W), 110.0, np.float32)
    scale = max(1.0, width / 640.0)
    n = int(n_shapes * scale * scale)
    yy, xx = np.mgrid[0:H, 0:W]
    for _ in range(n):
        g = float(rng.integers(20, 236))
        cx, cy = int(rng.integers(0, W)), int(rng.integers(0, H))
        if rng.random() < 0.6:
            w, h = int(rng.integers(8, 70 * scale)), int(rng.integers(8, 70 * scale))
            img[max(cy - h // 2, 0):cy + h // 2, max(cx - w // 2, 0):cx + w // 2] = g
        else:
            r = int(rng.integers(5, 36 * scale))
            y0, y1, x0, x1 = max(cy - r, 0), min(cy + r + 1, H), max(cx - r, 0), min(cx + r + 1, W)
            m = (yy[y0:y1, x0:x1] - cy) ** 2 + (xx[y0:y1, x0:x1] - cx) ** 2 <= r * r
            img[y0:y1, x0:x1][m] = g
    img += _box_blur3(rng.uniform(-40, 40, (H, W)).astype(np.float32))
    # blocky noise at several scales: keeps the corner density high on the coarse pyramid levels too
    for cell, amp in ((3, 30.0), (6, 30.0), (12, 25.0)):
        g = rng.uniform(-amp, amp, (H // cell + 1, W // cell + 1)).astype(np.float32)
        img += np.kron(g, np.ones((cell, cell), np.float32))[:H, :W]
    return img


def frame_from_base(base, width, height, t, seed, noise=12):
    """frame t of a sequence: the base translated by (2, 1) px per frame + i.i.d. uniform noise; 3 channels u8."""
    rng = np.random.default_rng((seed * 1000003 + t) & 0x7fffffff)
    ox, oy = (2 * t) % (base.shape[1] - width), (1 * t) % (base.shape[0] - height)
    g = base[oy:oy + height, ox:ox + width]
    out = np.empty((height, width, 3), np.float32)
    tint = (0.96, 1.0, 1.04)
    for c in range(3):
        out[..., c] = g * tint[c] + rng.uniform(-noise, noise, g.shape).astype(np.float32)
    return np.clip(np.rint(out), 0, 255).astype(np.uint8)


def frame_batch(width, height, n_seq, t, seed=1234, bases=None):
    """frame t of n_seq independent sequences -> [n_seq, H, W, 3] u8.  `bases` caches the textures."""
    if bases is None:
        bases = [base_texture(width, height, seed + 7919 * s) for s in range(n_seq)]
    return np.stack([frame_from_base(bases[s], width, height, t, seed + s) for s in range(n_seq)]), bases


def descriptor_db(n_keyframes, per_frame, seed=1):
    rng = np.random.default_rng(seed)
    return rng.integers(0, 256, (n_keyframes, per_frame, 32), dtype=np.uint8)


def plant_copies(db, queries, frac=0.01, flip_p=0.05, seed=3):
    """replace ~frac of the keyframes by noisy copies (each bit flipped w.p. flip_p) of query frames.
    Returns {keyframe index: query index}."""
    rng = np.random.default_rng(seed)
    K, Q = db.shape[0], queries.shape[0]
    n = max(1, int(round(K * frac)))
    planted = {}
    for kf in rng.choice(K, size=n, replace=False):
        q = int(rng.integers(0, Q))
        bits = np.unpackbits(queries[q], axis=-1)
        flips = rng.random(bits.shape) < flip_p
        db[kf] = np.packbits(bits ^ flips, axis=-1)
        planted[int(kf)] = q
    return planted


def _rodrigues(w):
    th = float(np.sqrt(w @ w))
    K = np.array([[0.0, -w[2], w[1]], [w[2], 0.0, -w[0]], [-w[1], w[0], 0.0]])
    if th < 1e-12:
        return np.eye(3) + K
    return np.eye(3) + (np.sin(th) / th) * K + ((1.0 - np.cos(th)) / (th * th)) * (K @ K)


TUM_K = (525.0, 525.0, 319.5, 239.5)  # fx, fy, cx, cy of the reference's TUM camera (rgbd_file_provider.cpp tumRgbdCameraParams)


def pnp_scene(seed, n=300, outlier_frac=0.3, noise=0.5, width=640, height=480, K=TUM_K):
    """Landmark / pixel correspondences of one tracked frame (the input of OpenCvRansacPnp::solvePnp): n world points seen
    by a camera at a random pose, pixel noise ~ N(0, noise), and outlier_frac of the pixels replaced by uniform ones.
    -> obj [n, 3] f32, img [n, 2] f32, rvec [3], tvec [3] (world -> camera), outlier mask [n] bool"""
    rng = np.random.default_rng(seed)
    fx, fy, cx, cy = K
    rv, tv = rng.normal(0, 0.15, 3), rng.normal(0, 0.3, 3)
    R = _rodrigues(rv)
    u, v, z = rng.uniform(20, width - 20, n), rng.uniform(20, height - 20, n), rng.uniform(0.8, 6.0, n)
    Xc = np.stack([(u - cx) / fx * z, (v - cy) / fy * z, z], 1)
    X = (Xc - tv) @ R
    x = np.stack([u, v], 1) + rng.normal(0, noise, (n, 2))
    no = int(outlier_frac * n)
    out = np.zeros(n, bool)
    if no:
        oi = rng.choice(n, no, replace=False)
        x[oi] = np.stack([rng.uniform(0, width, no), rng.uniform(0, height, no)], 1)
        out[oi] = True
    return X.astype(np.float32), x.astype(np.float32), rv, tv, out
