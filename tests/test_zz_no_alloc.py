"""mpb_fk_batch_noalloc (include/mpb.h): the FK call that allocates nothing -- its joint-frame table lives in a fixed
`static __device__` array of the library.  Checked through the C-ABI against mpb_fk_batch on the same device buffers, on
two calls in a row (the array keeps the first call's contents).  Collected last on purpose:
a failure here hides no other test."""
import numpy as np
import pytest

from tests.common import make_panda_world, random_q

pytestmark = pytest.mark.gpu

TOL = 1e-5  # the FK tolerance of tests/test_gpu_parity.py::test_fk_parity; the two kernels share their arithmetic


def _fk_both(dev, q_host, stream):
    import torch

    from mplib_b200 import _lib

    n = q_host.shape[0]
    L = dev.n_links
    qd = torch.from_numpy(np.ascontiguousarray(q_host.astype(np.float32))).cuda()
    want = torch.zeros((n, L, 7), dtype=torch.float32, device="cuda")
    got = torch.full((n, L, 7), float("nan"), dtype=torch.float32, device="cuda")
    torch.cuda.synchronize()
    dev.fk_raw(qd.data_ptr(), n, want.data_ptr(), _lib.MPB_DEVICE, stream.cuda_stream)
    dev.fk_noalloc_raw(qd.data_ptr(), n, got.data_ptr(), stream.cuda_stream)
    stream.synchronize()
    return got.cpu().numpy(), want.cpu().numpy()


def test_fk_noalloc_matches_fk_batch_twice_in_a_row():
    import torch

    dev = make_panda_world()._sync()["dev"]
    stream = torch.cuda.Stream()
    # second call: smaller and different, so stale words of the first call sit beside the live ones
    for n, seed in ((20000, 901), (3001, 902)):
        q = random_q(n, seed)
        got, want = _fk_both(dev, q, stream)
        assert np.isfinite(got).all()
        assert np.abs(got - want).max() <= TOL
    # and the result is the one the host-pointer API returns
    q = random_q(257, 903)
    got, _ = _fk_both(dev, q, stream)
    assert np.abs(got - dev.fk(q)).max() <= TOL


def test_fk_noalloc_rejects_batches_beyond_the_fixed_array():
    import torch

    from mplib_b200 import _lib

    dev = make_panda_world()._sync()["dev"]
    n = 1 << 20  # 2^20 x 10 frames x 48 B = 480 MiB > MPB_NOALLOC_SCRATCH_BYTES (16 MiB)
    qd = torch.zeros((n, dev.state_dim), dtype=torch.float32, device="cuda")
    out = torch.zeros((8, dev.n_links, 7), dtype=torch.float32, device="cuda")  # never written: the call must refuse
    with pytest.raises(_lib.MpbError, match="exceeds the fixed scratch array"):
        dev.fk_noalloc_raw(qd.data_ptr(), n, out.data_ptr(), None)
    assert float(out.abs().max()) == 0.0
