"""GPU-box probe: breakdown of the host-buffer (e2e) call of mpb_collide_batch on C2."""
import os, sys, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mplib_b200 import _lib
from tests import workloads as WL

wl = WL.get("c2", 1 << 20)
N, D = wl["batch"], wl["desc"].state_dim
dev = _lib.DeviceScene(wl["desc"], 0)
qh = [torch.from_numpy(wl["make_q"](k, N)).pin_memory() for k in range(4)]
vh = torch.zeros((N + 31) // 32, dtype=torch.int32).pin_memory()
ph = torch.zeros((N, dev.pair_words), dtype=torch.int32).pin_memory()
qd = qh[0].cuda()
for pairs in (True, False):
    for k in range(8):
        t0 = time.perf_counter()
        st = dev.collide_raw(qh[k % 4], N, vh, ph if pairs else None, _lib.MPB_HOST, None)
        dt = (time.perf_counter() - t0) * 1e3
        if k >= 5:
            print(f"chunk={os.environ.get('MPB_HOST_CHUNK','default')} pairs={pairs} wall={dt:.3f} ms enqueue={st['host_ms_enqueue']:.3f} total={st['host_ms_total']:.3f} "
                  f"gpu_cull={st['gpu_ms_cull']:.3f} gpu_narrow={st['gpu_ms_narrow']:.3f} launches={st['kernel_launches']}")
# raw copy rates
torch.cuda.synchronize()
for _ in range(3):
    t0 = time.perf_counter(); qd.copy_(qh[1], non_blocking=True); torch.cuda.synchronize(); t1 = time.perf_counter()
print(f"H2D {N*D*4/1e6:.1f} MB in {(t1-t0)*1e3:.3f} ms = {N*D*4/(t1-t0)/1e9:.1f} GB/s")
pd = torch.zeros((N, dev.pair_words), dtype=torch.int32, device="cuda")
for _ in range(3):
    t0 = time.perf_counter(); ph.copy_(pd, non_blocking=True); torch.cuda.synchronize(); t1 = time.perf_counter()
print(f"D2H {N*4/1e6:.1f} MB in {(t1-t0)*1e3:.3f} ms")
