"""per-stage device time of the distributed-ORB detector on a 64-frame batch (device-resident inputs)"""
import json
import sys

import numpy as np
import torch

sys.path.insert(0, ".")
import __graft_entry__ as g

pkg = g.load_package()
for (W, H, B) in [(640, 480, 64), (1280, 720, 32), (1920, 1080, 16)]:
    frames, _ = pkg.synth.frame_batch(W, H, B, 1, seed=1234)
    with pkg.DistContext(max_width=W, max_height=H, max_batch=B) as ctx:
        cap = ctx.kp_cap
        d_bgr = torch.from_numpy(frames).cuda()
        d_kp = torch.zeros((B, cap, 6), dtype=torch.float32, device="cuda")
        d_desc = torch.zeros((B, cap, 32), dtype=torch.uint8, device="cuda")
        d_n = torch.zeros(B, dtype=torch.int32, device="cuda")
        torch.cuda.synchronize()
        for _ in range(3):
            ctx.detect_batch_dev(d_bgr, B, W, H, d_kp, d_desc, cap, d_n)
        ctx.synchronize()
        ctx.stage_times(True)
        K = 10
        for _ in range(K):
            ctx.detect_batch_dev(d_bgr, B, W, H, d_kp, d_desc, cap, d_n)
        t = ctx.stage_times(False)
        tot = sum(t.values()) / K
        print(json.dumps({"W": W, "H": H, "batch": B, "kp_cap": cap, "kp_mean": float(d_n.float().mean()),
                          "ms_per_batch": tot, "frames_per_s": B / tot * 1e3,
                          "stages_ms": {k: v / K for k, v in t.items()}}))
