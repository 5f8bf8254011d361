"""The bounding-sphere phase has three code paths (cull table in parameter space with the centres staged in shared
memory; global table, one tile per warp; global table, register-blocked over 2 / 4 tiles).  All of them are conservative
culls in front of the same exact tests, so verdict words and pair masks must be bit-identical whichever one runs.  The
path is chosen per process (environment), hence the subprocesses.  MPB_LAZY_FK=1 selects, for parameter-space scenes,
the fused FK + sphere kernel that writes OBB rows on demand (fk_sphere_kernel): the same bits again."""
import hashlib
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

SCRIPT = r"""
import hashlib, sys
import numpy as np
sys.path.insert(0, %r)
from tests.common import make_panda_world, random_q
from tests.workloads import c3_world, c5_world
h = hashlib.sha256()
for world, n, seed in ((make_panda_world(), 4099, 31), (c3_world(cloud_points=20000), 1031, 32), (c5_world(), 4507, 33)):
    dev = world._sync()["dev"]
    D = dev.desc.state_dim
    q = random_q(n, seed) if D == 7 else np.concatenate([random_q(n, seed), random_q(n, seed + 100)], axis=1)
    v, ph = dev.collide(q[:, :D], want_pairs=True)
    h.update(np.packbits(v).tobytes()); h.update(np.packbits(ph).tobytes())
    print(int(v.sum()), int(ph.sum()), file=sys.stderr)
print(h.hexdigest())
""" % ROOT


def _digest(env_extra):
    env = dict(os.environ)
    env.update(env_extra)
    out = subprocess.run([sys.executable, "-c", SCRIPT], env=env, capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stderr[-2000:]
    return out.stdout.strip().splitlines()[-1], out.stderr


@pytest.mark.gpu
def test_sphere_phase_variants_agree_bit_for_bit():
    ref, err = _digest({})
    assert len(ref) == 64 and hashlib.sha256().digest_size == 32
    for variant in ({"MPB_LAZY_FK": "1"}, {"MPB_CULL_PARAM": "0"}, {"MPB_SPHERE_CB": "1"}, {"MPB_SPHERE_CB": "2"}, {"MPB_CULL_PARAM": "0", "MPB_SPHERE_CB": "1"}):
        got, err2 = _digest(variant)
        assert got == ref, (variant, err, err2)
