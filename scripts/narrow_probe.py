"""GPU-box probe: device-resident C2 pass at several batch sizes; prints cull / narrow kernel times."""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mplib_b200 import _lib
from tests import workloads as WL

wl = WL.get(sys.argv[1] if len(sys.argv) > 1 else "c2", 1 << 20)
dev = _lib.DeviceScene(wl["desc"], 0)
for N in [int(a) for a in sys.argv[2:]] or [1 << 14, 1 << 17, 1 << 20]:
    qd = torch.from_numpy(wl["make_q"](0, N)).cuda()
    vd = torch.zeros((N + 31) // 32, dtype=torch.int32, device="cuda")
    pd = torch.zeros((N, dev.pair_words), dtype=torch.int32, device="cuda")
    for k in range(6):
        st = dev.collide_raw(qd, N, vd, pd, _lib.MPB_DEVICE, None)
    print(f"factor={os.environ.get('MPB_UNIT_FACTOR','1')} N={N} cull={st['gpu_ms_cull']:.3f} narrow={st['gpu_ms_narrow']:.3f} total_host={st['host_ms_total']:.3f} "
          f"items={st['pairs_after_cull']} evals={st['narrow_iterations']}")
