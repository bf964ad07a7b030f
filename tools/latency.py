"""single-frame latency through the host entry points the plugin calls once per frame (detect, then match vs the previous frame)"""
import json
import sys
import time

import numpy as np

sys.path.insert(0, ".")
import __graft_entry__ as g

pkg = g.load_package()
W, H = 640, 480
fr0, bases = pkg.synth.frame_batch(W, H, 1, 0, seed=1234)
fr1, _ = pkg.synth.frame_batch(W, H, 1, 1, seed=1234, bases=bases)
out = {}
with pkg.Context(max_width=W, max_height=H, max_batch=1, nfeatures=1000) as ctx:
    pf = [pkg.pinned_empty(fr0[0].shape, np.uint8), pkg.pinned_empty(fr0[0].shape, np.uint8)]
    pf[0][...] = fr0[0]
    pf[1][...] = fr1[0]
    k0, d0 = ctx.detect(pf[0])
    for name, frames in (("pinned", pf), ("pageable", [fr0[0], fr1[0]])):
        for _ in range(20):
            ctx.detect(frames[1])
        t = []
        for i in range(200):
            t0 = time.perf_counter()
            k1, d1 = ctx.detect(frames[i & 1])
            t.append(time.perf_counter() - t0)
        out[f"orb_detect_ms_{name}"] = float(np.median(t)) * 1e3
    t = []
    for i in range(200):
        t0 = time.perf_counter()
        m = ctx.match_knn_ratio(d0, d1)
        t.append(time.perf_counter() - t0)
    out["orb_match_ms"] = float(np.median(t)) * 1e3
    out["orb_launches_per_detect"] = None
    l0 = ctx.launch_count
    ctx.detect(pf[0])
    out["orb_launches_per_detect"] = ctx.launch_count - l0
with pkg.DistContext(max_width=W, max_height=H, max_batch=1) as d:
    for _ in range(20):
        d.detect(pf[1])
    t = []
    for i in range(200):
        t0 = time.perf_counter()
        d.detect(pf[i & 1])
        t.append(time.perf_counter() - t0)
    out["dist_detect_ms_pinned"] = float(np.median(t)) * 1e3
print(json.dumps(out))
