"""CPU restatement of the RANSAC PnP step -- TEST INFRASTRUCTURE ONLY.

Only tests/ and the bench's cpu_baseline leg import this module; the product (modular-slam_b200) never does.

Reference: OpenCvRansacPnp::solvePnp, src/lib/modular_slam/cv_ransac_pnp.cpp:15-85 -- object / image points and the camera
matrix are narrowed to float32 (:20-41,:52-53), then
    cv::solvePnPRansac(obj, img, K, noArray(), rvec, tvec, useExtrinsicGuess=true, 100, 5.0, 0.99, inliers)   (:56-57)
and the result is a world->camera (rvec, tvec) plus an inlier bit per landmark (:79-82).

The arithmetic lives in OpenCV 4.8.1 calib3d (solvepnp.cpp / ptsetreg.cpp), which is not in /root/reference: its RANSAC
draws 5-point subsets from cv::RNG, solves each with EPnP, keeps the hypothesis with most points whose SQUARED reprojection
error is <= 5^2, and finishes with a Levenberg-Marquardt fit (SOLVEPNP_ITERATIVE) over the inliers; the inlier list is the
best hypothesis's, not re-evaluated after the fit.  This module restates that *scheme* with a different minimal solver
and a different random stream, so parity with OpenCV is by tolerance, not by bits -- **parity: partial / statistical**
(DESIGN.md section 4e):

  * hypothesis h = 0 .. iters-1: 4 distinct indices from a counter-based hash (`draw`), Grunert's P3P on the first three
    (quartic in v = s3 / s1 built by polynomial products, closed-form roots + 2 Newton steps), each real solution turned
    into (R, t) by aligning the two point triangles, the 4th point picks the solution;
  * score = number of points with squared reprojection error <= thr^2 (same test as PnPRansacCallback::computeError +
    findInliers, written without the division: both sides times z^2), best = most inliers, lowest h on ties, at least
    5 inliers required (cv: goodCount > modelPoints - 1);
  * Levenberg-Marquardt over the inliers of the best hypothesis, started from it, in float64, to convergence;
  * outputs: rvec = log(R) (Rodrigues vector), tvec, inlier mask of the best hypothesis.

It is pinned two ways (tests/test_pnp.py): against `cv2.solvePnPRansac` (opencv-python-headless 4.13.0) golden outputs
in tests/golden/pnp.npz made by tests/golden/make_golden_pnp.py (identical inlier sets, pose within 1e-4), and against the
generating ground-truth poses.  The CUDA path (csrc/pnp_kernels.cu) follows THIS restatement step for step (same draws,
same formulas), so GPU-vs-oracle agreement is ~1e-9 with identical inlier masks and identical winning hypothesis.
"""
import math

import numpy as np

M32 = 0xFFFFFFFF


def lowbias32(x):
    x &= M32
    x ^= x >> 16
    x = (x * 0x7FEB352D) & M32
    x ^= x >> 15
    x = (x * 0x846CA68B) & M32
    x ^= x >> 16
    return x


def draw(seed, frame, hyp, j, n):
    """j-th draw of hypothesis hyp of frame `frame`: uniform index in [0, n)"""
    r = lowbias32((seed ^ lowbias32((frame * 0x9E3779B9 + hyp * 0x85EBCA6B + j * 0xC2B2AE35 + 0x27D4EB2F) & M32)) & M32)
    return (r * n) >> 32


def sample4(seed, frame, hyp, n):
    idx = []
    j = 0
    while len(idx) < 4 and j < 64:
        i = draw(seed, frame, hyp, j, n)
        j += 1
        if i not in idx:
            idx.append(i)
    return idx if len(idx) == 4 else None


def _cubic_largest_real_root(p, q, r):
    """largest real root of y^3 + p y^2 + q y + r"""
    Q = (p * p - 3.0 * q) / 9.0
    R = (2.0 * p * p * p - 9.0 * p * q + 27.0 * r) / 54.0
    Q3 = Q * Q * Q
    if R * R < Q3:
        th = math.acos(max(-1.0, min(1.0, R / math.sqrt(Q3))))
        sq = -2.0 * math.sqrt(Q)
        y0 = sq * math.cos(th / 3.0)
        y1 = sq * math.cos((th + 2.0 * math.pi) / 3.0)
        y2 = sq * math.cos((th - 2.0 * math.pi) / 3.0)
        return max(y0, y1, y2) - p / 3.0
    A = -math.copysign(1.0, R) * (abs(R) + math.sqrt(max(R * R - Q3, 0.0))) ** (1.0 / 3.0)
    B = Q / A if A != 0.0 else 0.0
    return A + B - p / 3.0


def quartic_real_roots(c4, c3, c2, c1, c0):
    """real roots of c4 x^4 + ... + c0 (Ferrari via the resolvent cubic), each polished by two Newton steps"""
    if c4 == 0.0 or not math.isfinite(c4):
        return []
    a, b, c, d = c3 / c4, c2 / c4, c1 / c4, c0 / c4
    if not (math.isfinite(a) and math.isfinite(b) and math.isfinite(c) and math.isfinite(d)):
        return []
    a2 = a * a
    y = _cubic_largest_real_root(-b, a * c - 4.0 * d, 4.0 * b * d - a2 * d - c * c)
    R2 = 0.25 * a2 - b + y
    scale = abs(0.25 * a2) + abs(b) + abs(y) + 1e-300
    if R2 < -1e-12 * scale:
        return []
    R = math.sqrt(max(R2, 0.0))
    if R > 1e-8 * math.sqrt(scale):
        t = 0.25 * (4.0 * a * b - 8.0 * c - a2 * a) / R
        D2 = 0.75 * a2 - R * R - 2.0 * b + t
        E2 = 0.75 * a2 - R * R - 2.0 * b - t
    else:
        R = 0.0
        w = y * y - 4.0 * d
        if w < 0.0:
            w = 0.0
        D2 = 0.75 * a2 - 2.0 * b + 2.0 * math.sqrt(w)
        E2 = 0.75 * a2 - 2.0 * b - 2.0 * math.sqrt(w)
    tol = 1e-10 * (abs(0.75 * a2) + abs(2.0 * b) + 1e-300)
    roots = []
    if D2 >= -tol:
        D = math.sqrt(max(D2, 0.0))
        roots += [-0.25 * a + 0.5 * R + 0.5 * D, -0.25 * a + 0.5 * R - 0.5 * D]
    if E2 >= -tol:
        E = math.sqrt(max(E2, 0.0))
        roots += [-0.25 * a - 0.5 * R + 0.5 * E, -0.25 * a - 0.5 * R - 0.5 * E]
    out = []
    for x in roots:
        for _ in range(2):
            f = (((x + a) * x + b) * x + c) * x + d
            fp = ((4.0 * x + 3.0 * a) * x + 2.0 * b) * x + c
            if fp != 0.0:
                x = x - f / fp
        out.append(x)
    return out


def _triad(p0, p1, p2):
    e1 = p1 - p0
    n1 = math.sqrt(float(e1 @ e1))
    if n1 == 0.0:
        return None
    e1 = e1 / n1
    e3 = np.cross(e1, p2 - p0)
    n3 = math.sqrt(float(e3 @ e3))
    if n3 == 0.0:
        return None
    e3 = e3 / n3
    e2 = np.cross(e3, e1)
    return np.stack([e1, e2, e3], axis=1)  # columns


def p3p(P, f):
    """P [3,3] world points (rows), f [3,3] unit bearing vectors (rows) -> list of (R, t) with Xc = R X + t"""
    a2 = float((P[1] - P[2]) @ (P[1] - P[2]))
    b2 = float((P[0] - P[2]) @ (P[0] - P[2]))
    c2 = float((P[0] - P[1]) @ (P[0] - P[1]))
    if a2 == 0.0 or b2 == 0.0 or c2 == 0.0:
        return []
    ca, cb, cg = float(f[1] @ f[2]), float(f[0] @ f[2]), float(f[0] @ f[1])
    k1 = (a2 - c2) / b2
    k2 = c2 / b2
    # u = N(v) / D(v):  N = n2 v^2 + n1 v + n0,  D = d1 v + d0
    n2, n1, n0 = k1 - 1.0, -2.0 * k1 * cb, 1.0 + k1
    d1, d0 = -2.0 * ca, 2.0 * cg
    # N^2 - 2 cg N D + D^2 (-k2 v^2 + 2 k2 cb v + 1 - k2) = 0
    NN = [n0 * n0, 2.0 * n0 * n1, n1 * n1 + 2.0 * n0 * n2, 2.0 * n1 * n2, n2 * n2]
    ND = [n0 * d0, n0 * d1 + n1 * d0, n1 * d1 + n2 * d0, n2 * d1, 0.0]
    DD = [d0 * d0, 2.0 * d0 * d1, d1 * d1]
    q0, q1, q2 = 1.0 - k2, 2.0 * k2 * cb, -k2
    DQ = [DD[0] * q0, DD[0] * q1 + DD[1] * q0, DD[0] * q2 + DD[1] * q1 + DD[2] * q0, DD[1] * q2 + DD[2] * q1, DD[2] * q2]
    co = [NN[i] - 2.0 * cg * ND[i] + DQ[i] for i in range(5)]
    sols = []
    Bw = _triad(P[0], P[1], P[2])
    if Bw is None:
        return []
    for v in quartic_real_roots(co[4], co[3], co[2], co[1], co[0]):
        if not (v > 0.0) or not math.isfinite(v):
            continue
        den = d1 * v + d0
        if abs(den) < 1e-12:
            continue
        u = ((n2 * v + n1) * v + n0) / den
        if not (u > 0.0):
            continue
        w = 1.0 + v * v - 2.0 * v * cb
        if not (w > 0.0):
            continue
        s1 = math.sqrt(b2 / w)
        C0, C1, C2 = f[0] * s1, f[1] * (u * s1), f[2] * (v * s1)
        Bc = _triad(C0, C1, C2)
        if Bc is None:
            continue
        R = Bc @ Bw.T
        t = C0 - R @ P[0]
        sols.append((R, t))
    return sols


def project_err2(R, t, X, x, fx, fy, cx, cy):
    Xc = X @ R.T + t
    with np.errstate(divide="ignore", invalid="ignore"):
        u = fx * (Xc[:, 0] / Xc[:, 2]) + cx
        v = fy * (Xc[:, 1] / Xc[:, 2]) + cy
    du, dv = u - x[:, 0], v - x[:, 1]
    return du * du + dv * dv


def inlier_mask(R, t, X, x, fx, fy, cx, cy, thr2):
    """|proj(X) - x|^2 <= thr2, both sides multiplied by z^2 (no division; z == 0 is never an inlier) -- the same operation
    sequence as is_inlier() in csrc/pnp_kernels.cu"""
    Xc = X @ R.T + t
    z = Xc[:, 2]
    a = fx * Xc[:, 0] + (cx - x[:, 0]) * z
    b = fy * Xc[:, 1] + (cy - x[:, 1]) * z
    return (z != 0.0) & (a * a + b * b <= thr2 * (z * z))


def rodrigues_exp(w):
    th = math.sqrt(float(w @ w))
    K = np.array([[0.0, -w[2], w[1]], [w[2], 0.0, -w[0]], [-w[1], w[0], 0.0]])
    if th < 1e-12:
        return np.eye(3) + K
    return np.eye(3) + (math.sin(th) / th) * K + ((1.0 - math.cos(th)) / (th * th)) * (K @ K)


def rodrigues_log(R):
    c = 0.5 * (R[0, 0] + R[1, 1] + R[2, 2] - 1.0)
    c = max(-1.0, min(1.0, c))
    th = math.acos(c)
    r = np.array([R[2, 1] - R[1, 2], R[0, 2] - R[2, 0], R[1, 0] - R[0, 1]])
    s = 0.5 * math.sqrt(float(r @ r))  # sin(th)
    if s < 1e-10:
        if c > 0.0:
            return 0.5 * r
        # th ~ pi: axis from the diagonal of (R + I) / 2
        d = np.sqrt(np.maximum((np.diag(R) + 1.0) * 0.5, 0.0))
        if R[0, 1] < 0.0:
            d[1] = -d[1]
        if R[0, 2] < 0.0:
            d[2] = -d[2]
        n = math.sqrt(float(d @ d))
        return d * (th / n) if n > 0 else np.zeros(3)
    return r * (0.5 * th / s)


def refine(R, t, X, x, fx, fy, cx, cy, iters=30, trace=None):
    """Levenberg-Marquardt on the reprojection error; update Xc' = exp(w) Xc + dt.
    trace (a list) receives one (iteration, trial, accepted) tuple per damping trial."""
    def cost_of(R, t):
        e = project_err2(R, t, X, x, fx, fy, cx, cy)
        return float(e.sum())

    lam = 1e-3
    cost = cost_of(R, t)
    for it in range(iters):
        Xc = X @ R.T + t
        iz = 1.0 / Xc[:, 2]
        px, py = Xc[:, 0] * iz, Xc[:, 1] * iz
        ru = fx * px + cx - x[:, 0]
        rv = fy * py + cy - x[:, 1]
        # d(u,v)/dXc
        Ju = np.stack([fx * iz, np.zeros_like(iz), -fx * px * iz], axis=1)
        Jv = np.stack([np.zeros_like(iz), fy * iz, -fy * py * iz], axis=1)
        # dXc/dw = -[Xc]x  ->  row . (-[Xc]x) = (Xc x row)
        Jwu, Jwv = np.cross(Xc, Ju), np.cross(Xc, Jv)
        J = np.concatenate([np.concatenate([Jwu, Ju], axis=1), np.concatenate([Jwv, Jv], axis=1)], axis=0)
        r = np.concatenate([ru, rv])
        H = J.T @ J
        g = J.T @ r
        improved = False
        for _try in range(8):
            A = H + lam * np.diag(np.diag(H)) + 1e-12 * np.eye(6)
            try:
                dx = -np.linalg.solve(A, g)
            except np.linalg.LinAlgError:
                lam *= 10.0
                continue
            dR = rodrigues_exp(dx[:3])
            Rn, tn = dR @ R, dR @ t + dx[3:]
            cn = cost_of(Rn, tn)
            if trace is not None:
                trace.append((it, _try, bool(math.isfinite(cn) and cn <= cost)))
            if math.isfinite(cn) and cn <= cost:
                R, t = Rn, tn
                step = float(np.abs(dx).max())
                drop = cost - cn
                cost = cn
                lam = max(lam * 0.1, 1e-12)
                improved = True
                break
            lam *= 10.0
        # quadratic convergence: a step below 1e-10 is the last one that matters in float64 (same rule as the kernel)
        if not improved or step < 1e-10 or drop <= 1e-15 * max(cost, 1e-300):
            break
    return R, t


def solve_pnp_ransac(obj, img, fx, fy, cx, cy, iters=100, reproj=5.0, seed=0, frame=0, refine_iters=30, trace=None):
    """-> (ok, rvec [3], tvec [3], inlier mask [n] u8, best hypothesis index, R, t)"""
    X = np.asarray(obj, np.float32).astype(np.float64).reshape(-1, 3)  # cv_ransac_pnp.cpp:20-30 narrows to float
    x = np.asarray(img, np.float32).astype(np.float64).reshape(-1, 2)  # :32-41
    fx, fy, cx, cy = (float(np.float32(v)) for v in (fx, fy, cx, cy))  # :52-53
    n = len(X)
    mask = np.zeros(n, np.uint8)
    fail = (False, np.zeros(3), np.zeros(3), mask, -1, np.eye(3), np.zeros(3))
    if n < 4:
        return fail
    need = 5 if n >= 5 else 4
    thr2 = float(reproj) * float(reproj)
    bear = np.stack([(x[:, 0] - cx) / fx, (x[:, 1] - cy) / fy, np.ones(n)], axis=1)
    bear /= np.sqrt((bear * bear).sum(axis=1))[:, None]
    best = (-1, -1, None, None)
    for h in range(iters):
        idx = sample4(seed, frame, h, n)
        if idx is None:
            continue
        sols = p3p(X[idx[:3]], bear[idx[:3]])
        pick, pick_e = None, math.inf
        for R, t in sols:
            e = float(project_err2(R, t, X[idx[3]:idx[3] + 1], x[idx[3]:idx[3] + 1], fx, fy, cx, cy)[0])
            zc = float(R[2] @ X[idx[3]] + t[2])
            if zc > 0.0 and e < pick_e:
                pick, pick_e = (R, t), e
        if pick is None:
            continue
        cnt = int(inlier_mask(pick[0], pick[1], X, x, fx, fy, cx, cy, thr2).sum())
        if cnt > best[0]:
            best = (cnt, h, pick[0], pick[1])
    if best[0] < need:
        return fail
    _, h, R, t = best
    mask = inlier_mask(R, t, X, x, fx, fy, cx, cy, thr2).astype(np.uint8)
    sel = mask.astype(bool)
    R, t = refine(R, t, X[sel], x[sel], fx, fy, cx, cy, refine_iters, trace)
    return True, rodrigues_log(R), t.copy(), mask, h, R, t
