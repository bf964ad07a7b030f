"""single-frame calls of the second detector only (for an ncu launch list)"""
import sys
import numpy as np
sys.path.insert(0, ".")
import __graft_entry__ as g
pkg = g.load_package()
W, H = 640, 480
fr0, bases = pkg.synth.frame_batch(W, H, 1, 0, seed=1234)
with pkg.DistContext(max_width=W, max_height=H, max_batch=1) as d:
    pf = pkg.pinned_empty(fr0[0].shape, np.uint8)
    pf[...] = fr0[0]
    for _ in range(12):
        k, _ = d.detect(pf)
    print(len(k))
