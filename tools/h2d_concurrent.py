#!/usr/bin/env python
"""Concurrent pinned host->device copy ceiling of one box: one process per GPU (torchrun), the c2 batch size (59 MB of BGR frames
per copy), n = 1, 2, 4, 8 ranks copying AT THE SAME TIME, with and without binding each rank to the CPUs / memory of its GPU's
NUMA node before the pinned pool is allocated.  Answers VERDICT r1 item 7: is the end-to-end scaling limit of the frame path the
box's host side, or the way the ranks use it?

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 tools/h2d_concurrent.py > profiles/...
"""
import json
import os
import time

import torch
import torch.distributed as dist

BYTES = 64 * 640 * 480 * 3


def gpu_locality(local):
    """(numa node, cpu list) of the GPU's PCIe device, from sysfs"""
    bdf = torch.cuda.get_device_properties(local).pci_bus_id if hasattr(torch.cuda.get_device_properties(local), "pci_bus_id") else None
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(local)
        bdf = pynvml.nvmlDeviceGetPciInfo(h).busId
        if isinstance(bdf, bytes):
            bdf = bdf.decode()
    except Exception:  # noqa: BLE001
        pass
    if not bdf:
        return None, None, None
    bdf = bdf.lower()
    if len(bdf.split(":")[0]) == 8:
        bdf = bdf[4:]
    base = f"/sys/bus/pci/devices/{bdf}"
    try:
        node = int(open(f"{base}/numa_node").read())
        cpus = open(f"{base}/local_cpulist").read().strip()
    except OSError:
        return bdf, None, None
    return bdf, node, cpus


def parse_cpulist(s):
    out = set()
    for part in s.split(","):
        if "-" in part:
            a, b = part.split("-")
            out.update(range(int(a), int(b) + 1))
        elif part:
            out.add(int(part))
    return out


def measure(local, seconds=1.5):
    x = torch.empty(BYTES, dtype=torch.uint8).pin_memory()
    x.fill_(7)
    d = torch.empty(BYTES, dtype=torch.uint8, device=f"cuda:{local}")
    for _ in range(3):
        d.copy_(x, non_blocking=True)
    torch.cuda.synchronize()
    return x, d


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    bdf, node, cpus = gpu_locality(local)
    aff0 = sorted(os.sched_getaffinity(0))
    rows = []
    for mode in ("as launched", "bound to the GPU's NUMA node"):
        if mode != "as launched":
            if cpus:
                try:
                    os.sched_setaffinity(0, parse_cpulist(cpus) & set(aff0) or set(aff0))
                except OSError:
                    pass
        x, d = measure(local)
        for n in (1, 2, 4, 8):
            if n > world:
                continue
            torch.cuda.synchronize()
            dist.barrier()
            gbs = 0.0
            if rank < n:
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                reps = 40
                e0.record()
                for _ in range(reps):
                    d.copy_(x, non_blocking=True)
                e1.record()
                torch.cuda.synchronize()
                gbs = BYTES * reps / (e0.elapsed_time(e1) * 1e-3) / 1e9
            dist.barrier()
            allg = [None] * world
            dist.all_gather_object(allg, gbs)
            if rank == 0:
                act = allg[:n]
                rows.append({"mode": mode, "concurrent_ranks": n, "gbs_per_rank": [round(v, 1) for v in act], "sum_gbs": round(sum(act), 1),
                             "min_gbs": round(min(act), 1)})
        del x, d
    info = [None] * world
    dist.all_gather_object(info, {"rank": rank, "gpu_bdf": bdf, "gpu_numa_node": node, "gpu_local_cpulist": cpus,
                                  "affinity_as_launched": f"{aff0[0]}-{aff0[-1]} ({len(aff0)} cpus)" if aff0 else None})
    if rank == 0:
        numa = {}
        try:
            for nd in sorted(os.listdir("/sys/devices/system/node")):
                if nd.startswith("node"):
                    numa[nd] = open(f"/sys/devices/system/node/{nd}/cpulist").read().strip()
        except OSError:
            pass
        print(json.dumps({"bytes_per_copy": BYTES, "host_cpus": os.cpu_count(), "numa_nodes": numa, "ranks": info, "table": rows}, indent=1))
    dist.destroy_process_group()


if __name__ == "__main__":
    t0 = time.time()
    main()
