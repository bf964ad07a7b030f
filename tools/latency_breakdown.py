"""where the 0.19 ms of a single-frame detect goes: pieces timed separately through the same entry points
(host wall clock around blocking calls, medians of 300)"""
import ctypes
import json
import sys
import time

import numpy as np
import torch

sys.path.insert(0, ".")
import __graft_entry__ as g

pkg = g.load_package()
W, H = 640, 480
fr0, bases = pkg.synth.frame_batch(W, H, 1, 0, seed=1234)


def med(fn, n=300, warm=30):
    for _ in range(warm):
        fn()
    t = []
    for _ in range(n):
        t0 = time.perf_counter()
        fn()
        t.append(time.perf_counter() - t0)
    return round(float(np.median(t)) * 1e3, 4)


out = {}
with pkg.Context(max_width=W, max_height=H, max_batch=1, nfeatures=1000) as ctx:
    cap = ctx.kp_cap if hasattr(ctx, "kp_cap") else 1340
    pf = pkg.pinned_empty(fr0[0].shape, np.uint8)
    pf[...] = fr0[0]
    k, d = ctx.detect(pf)
    cap = max(cap, len(k))
    out["detect_host_pinned_ms"] = med(lambda: ctx.detect(pf))
    # (1) kernels only: frame resident on the device, outputs stay there
    d_bgr = torch.from_numpy(fr0[0].copy()).cuda().contiguous()
    d_kp = torch.empty(cap * 24, dtype=torch.uint8, device="cuda")
    d_desc = torch.empty(cap * 32, dtype=torch.uint8, device="cuda")
    d_cnt = torch.zeros(4, dtype=torch.int32, device="cuda")
    ctx.set_stream(torch.cuda.current_stream().cuda_stream)

    def dev():
        ctx.detect_batch_dev(d_bgr.data_ptr(), 1, W, H, d_kp.data_ptr(), d_desc.data_ptr(), cap, d_cnt.data_ptr())
        torch.cuda.synchronize()

    out["detect_dev_plus_sync_ms"] = med(dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ts = []
    for _ in range(100):
        e0.record()
        ctx.detect_batch_dev(d_bgr.data_ptr(), 1, W, H, d_kp.data_ptr(), d_desc.data_ptr(), cap, d_cnt.data_ptr())
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    out["detect_dev_gpu_span_ms"] = round(float(np.median(ts)), 4)

    def enqueue_only():
        ctx.detect_batch_dev(d_bgr.data_ptr(), 1, W, H, d_kp.data_ptr(), d_desc.data_ptr(), cap, d_cnt.data_ptr())

    t = []
    for _ in range(100):
        t0 = time.perf_counter()
        enqueue_only()
        t.append(time.perf_counter() - t0)
        torch.cuda.synchronize()
    out["detect_dev_enqueue_host_ms"] = round(float(np.median(t)) * 1e3, 4)
    # (2) raw copies
    hp = torch.from_numpy(np.asarray(pf)) if False else torch.empty(fr0[0].size, dtype=torch.uint8).pin_memory()
    dd = torch.empty(fr0[0].size, dtype=torch.uint8, device="cuda")

    def h2d():
        dd.copy_(hp, non_blocking=True)
        torch.cuda.synchronize()

    out["h2d_921KB_plus_sync_ms"] = med(h2d)
    ho = torch.empty(cap * 56, dtype=torch.uint8).pin_memory()
    do = torch.empty(cap * 56, dtype=torch.uint8, device="cuda")

    def d2h():
        ho.copy_(do, non_blocking=True)
        torch.cuda.synchronize()

    out["d2h_75KB_plus_sync_ms"] = med(d2h)
    out["empty_sync_ms"] = med(torch.cuda.synchronize)
print(json.dumps(out))
