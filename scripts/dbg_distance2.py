import sys, numpy as np
sys.path.insert(0, '.')
from tests.common import *
from tests.workloads import add_c3_primitives, attach_c3_box
w = make_panda_world(); add_c3_primitives(w, 4, seed=3); attach_c3_box(w)
c = w._sync(); dev = c["dev"]; desc = dev.desc
q = random_q(1500, 72).astype(np.float32)
for rep in range(3):
    md, am = dev.distance(q)
    print("full", rep, md[961], am[961], (md < 0).sum())
md, am = dev.distance(q[961:962]); print("single", md, am)
md, am = dev.distance(q[900:1000]); print("100", md[61], am[61])
md, am = dev.distance(q[960:964]); print("4", md, am)
pa, pb = desc.pair_a.copy(), desc.pair_b.copy()
for p in (120, 121, 122):
    dev.set_pairs([pa[p]], [pb[p]], ["x"])
    md, am = dev.distance(q[900:1000]); print("pair", p, md[61], am[61], (md < 0).sum())
    md, am = dev.distance(q[961:962]); print("pair", p, "single", md, am)
dev.set_pairs(pa[100:140], pb[100:140], ["x"] * 40)
md, am = dev.distance(q[900:1000]); print("pairs 100..140", md[61], am[61])
