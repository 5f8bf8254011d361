import sys, time; sys.path.insert(0,'.')
import numpy as np
from mplib_b200 import _lib, workloads as WL
import ctypes as C
lib=_lib.load()
for name in ("c2","c3"):
    wl=WL.get(name, 1024); dev=_lib.DeviceScene(wl["desc"],0)
    for n in (1,32):
        q=np.ascontiguousarray(wl["make_q"](3,n).astype(np.float32)); vb=np.zeros((n+31)//32,dtype=np.uint32)
        ts=[]
        for it in range(60):
            t0=time.perf_counter(); rc=lib.mpb_collide_batch(dev._h,q.ctypes.data,n,vb.ctypes.data,None,None,0,None); ts.append(time.perf_counter()-t0)
        print(name,"n",n,"median us",np.median(ts[10:])*1e6, "min", min(ts)*1e6)
