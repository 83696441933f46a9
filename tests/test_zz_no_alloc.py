"""dvf_update_v_noalloc: the allocation-free variant of dvf_update_v (scratch in a file-scope device array, fit.cu).
No counterpart in the reference.  Called through the C ABI, compared
with the existing entry point on two calls in a row (the array keeps the previous call's contents), and the size limits
must be refused with a clear error.  Collected last (file name), self-contained (own inputs, own process state)."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _call(fn, UY, Cm, P, minP, N, D, M):
    import torch
    from dynamo_release_b200 import _lib

    V = torch.empty((N, D), dtype=torch.float64, device="cuda")
    r2 = torch.empty((N,), dtype=torch.float64, device="cuda")
    st = torch.empty((1,), dtype=torch.float64, device="cuda")
    _lib.check(fn(UY.data_ptr(), Cm.data_ptr(), P.data_ptr(), float(minP), N, D, M, V.data_ptr(), r2.data_ptr(),
                  st.data_ptr(), torch.cuda.current_stream().cuda_stream))
    torch.cuda.synchronize()
    return V.cpu().numpy(), r2.cpu().numpy(), float(st.item())


def test_update_v_noalloc_matches_update_v_twice_in_a_row():
    import torch
    from dynamo_release_b200 import _lib, kernels

    lib = _lib.load()
    rng = np.random.default_rng(17)
    minP = 1e-5
    # two different shapes back to back: the second call finds the first call's (larger) leftovers in the shared array
    for (N, D, M) in ((40_001, 30, 500), (3_333, 12, 64)):
        X = rng.normal(size=(N, D))
        Y = rng.normal(size=(N, D)) * 0.1
        Xc = X[rng.choice(N, M, replace=False)]
        UY = kernels.build_uy(X, Y, Xc, 0.05)
        Cm = torch.from_numpy(rng.normal(size=(M, D))).cuda()
        P = torch.from_numpy(rng.uniform(0, 1, size=N + 2)).cuda()
        ref = _call(lib.dvf_update_v, UY, Cm, P, minP, N, D, M)
        for attempt in range(2):
            got = _call(lib.dvf_update_v_noalloc, UY, Cm, P, minP, N, D, M)
            # same kernels on the same inputs, fixed summation order: the tolerance of the teacher-forced tests (1e-12)
            # is met with equality
            for a, b in zip(got[:2], ref[:2]):
                assert np.max(np.abs(a - b)) <= 1e-12 * np.max(np.abs(b)), (N, D, M, attempt)
                assert np.array_equal(a, b), (N, D, M, attempt)
            assert abs(got[2] - ref[2]) <= 1e-12 * abs(ref[2]) and got[2] == ref[2]
        # and against plain float64 arithmetic on the host
        U = UY[:, :M].cpu().numpy()
        Vh = U @ Cm.cpu().numpy()
        assert np.max(np.abs(got[0] - Vh)) <= 1e-12 * np.max(np.abs(Vh))


def test_update_v_noalloc_refuses_sizes_beyond_its_scratch():
    import torch
    from dynamo_release_b200 import _lib

    lib = _lib.load()
    d = torch.zeros(8, dtype=torch.float64, device="cuda")
    p, s = d.data_ptr(), torch.cuda.current_stream().cuda_stream
    for (N, D, M) in ((4194304 + 1, 30, 100), (1000, 30, 1025)):
        rc = lib.dvf_update_v_noalloc(p, p, p, 1e-5, N, D, M, p, p, p, s)      # refused before anything is touched
        assert rc == _lib.DVF_ERR_WORKSPACE
        msg = lib.dvf_last_error().decode()
        assert "dvf_update_v_noalloc" in msg and "M <= 1024" in msg and "N <= 4194304" in msg
