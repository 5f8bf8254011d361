#!/usr/bin/env python
"""Where the end-to-end time of a C3 host-buffer call goes: verdict only / full hit list / truncated hit list (no list copy),
each with the library's own gpu / enqueue / total times, and the box's pinned copy rates.  Run on a GPU box:
PYTHONPATH=. python scripts/e2e_dissect.py   (this is what showed the hit-list shares queued behind the next pass's kernels)"""
import time, numpy as np, torch
from mplib_b200 import _lib
from mplib_b200 import workloads as W
wl = W.c3()
from mplib_b200._lib import DeviceScene
dev = DeviceScene(wl["desc"], 0)
N = wl["batch"]
q = torch.from_numpy(wl["make_q"](0, N).astype(np.float32)).pin_memory()
vw = (N + 31) // 32
verdict_h = torch.zeros(vw, dtype=torch.int32).pin_memory()
cap = 12 * N
hits_h = torch.zeros((cap, 2), dtype=torch.int32).pin_memory()
nh = torch.zeros(1, dtype=torch.int32).pin_memory()
def run(name, f, k=4):
    f(); torch.cuda.synchronize()
    t = time.perf_counter()
    for _ in range(k): st = f()
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t) / k * 1e3
    print(f"{name:28s} {dt:7.2f} ms  gpu {st['gpu_ms']:.2f} enq {st['host_ms_enqueue']:.2f} total {st['host_ms_total']:.2f}", flush=True)
run("verdict only", lambda: dev.collide_raw(q, N, verdict_h, None, _lib.MPB_HOST, None))
run("hits full cap", lambda: dev.collide_hits_raw(q, N, verdict_h, hits_h, cap, nh, _lib.MPB_HOST, None))
run("hits cap 1024 (truncated)", lambda: dev.collide_hits_raw(q, N, verdict_h, hits_h, 1024, nh, _lib.MPB_HOST, None))
# raw D2H / H2D rate of this box
d = torch.empty(80 << 20, dtype=torch.uint8, device="cuda"); h = torch.empty(80 << 20, dtype=torch.uint8).pin_memory()
for name, a, b in (("D2H 80MB", h, d), ("H2D 80MB", d, h)):
    a.copy_(b, non_blocking=True); torch.cuda.synchronize()
    t = time.perf_counter()
    for _ in range(5): a.copy_(b, non_blocking=True)
    torch.cuda.synchronize()
    print(name, "%.1f GB/s" % (5 * 80 * 2**20 / (time.perf_counter() - t) / 1e9))
