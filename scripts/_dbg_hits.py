import sys; sys.path.insert(0,'.')
import numpy as np
from tests.common import random_q
from tests.workloads import c3_world
from oracle import oracle as O
w = c3_world(n_each=8, mesh=True, cloud_points=5000)
dev = w._sync()["dev"]; desc=dev.desc
n=70000
q = random_q(n, 300 + n)
res=[dev.collide(q, want_pairs=True)[1] for _ in range(4)]
for k in range(1,4):
    d=np.argwhere(res[0]!=res[k])
    print("run",k,"diff",len(d))
    for i,p in d[:6]:
        a,b=desc.pair_a[p],desc.pair_b[p]
        sd=O.pair_signed_distance(desc,q[i].astype(np.float32).astype(np.float64),int(p))
        print("   cfg",i,"pair",p,desc.pair_kinds[p],desc.object_names[a],desc.object_names[b],"types",desc.obj_type[a],desc.obj_type[b],"r0",res[0][i,p],"rk",res[k][i,p],"oracle sd",sd)
vh,h = dev.collide_hits(q)
d = np.zeros_like(res[0]); d[h[:,0],h[:,1]]=True
print("hits vs dense0 diff", (d!=res[0]).sum())
# smaller world without BVH
w2 = c3_world(n_each=8, mesh=False, cloud_points=0); dev2=w2._sync()["dev"]
r2=[dev2.collide(q, want_pairs=True)[1] for _ in range(3)]
print("no-BVH world diffs", (r2[0]!=r2[1]).sum(), (r2[0]!=r2[2]).sum())
w3 = c3_world(n_each=0, primitives=False, mesh=True, cloud_points=0, attach=False); dev3=w3._sync()["dev"]
r3=[dev3.collide(q, want_pairs=True)[1] for _ in range(3)]
print("mesh-only world diffs", (r3[0]!=r3[1]).sum(), (r3[0]!=r3[2]).sum())
w4 = c3_world(n_each=0, primitives=False, mesh=False, cloud_points=5000, attach=False); dev4=w4._sync()["dev"]
r4=[dev4.collide(q, want_pairs=True)[1] for _ in range(3)]
print("cloud-only world diffs", (r4[0]!=r4[1]).sum(), (r4[0]!=r4[2]).sum())
