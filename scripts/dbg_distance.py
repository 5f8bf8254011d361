import sys, numpy as np
sys.path.insert(0, '.')
from tests.common import *
from tests.workloads import add_c3_primitives, attach_c3_box
from oracle import oracle as O
w = make_panda_world(); add_c3_primitives(w, 4, seed=3); attach_c3_box(w)
c = w._sync(); dev = c["dev"]; desc = dev.desc
q = random_q(1500, 72).astype(np.float32)
qd = q.astype(np.float64)
pa, pb = desc.pair_a.copy(), desc.pair_b.copy()
names = desc.object_names; types = desc.obj_type
bad = 0
for p in ([int(sys.argv[1])] if len(sys.argv) > 1 else range(len(pa))):
    dev.set_pairs([pa[p]], [pb[p]], ["x"])
    md, am = dev.distance(q[900:1000])
    d1 = dev.desc
    for i in range(100):
        o = O.pair_signed_distance(d1, qd[900 + i], 0)
        g = md[i]
        if (g < 0) != (o < 0) and abs(o) > 1e-4 or (g >= 0 and o >= 0 and abs(g - o) > 1e-4):
            bad += 1
            if bad < 40:
                print("state", 900 + i, "pair", p, names[pa[p]], types[pa[p]], names[pb[p]], types[pb[p]], "gpu", g, "oracle", o)
print("bad", bad)
