"""rslqr_cuda_solve_static (include/rslqr_cuda.h): the scan-path solve with its scratch in a static device array, for
callers that cannot allocate. Checked through the C-ABI against the existing entry point (a handle's
rslqr_cuda_solve_async) on two calls in a row. Collected last on purpose."""
import ctypes as C

import numpy as np
import pytest

import rslqr_b200 as R
from helpers import rel_err_classes

pytestmark = pytest.mark.gpu
TOL = 1e-10
E_ARG = -1


def _by_handle(n, m, N, kn, x0):
    s = R.CudaSolver(n, m, N, batch=x0.shape[0], flags=0)
    assert s.mode == R.MODE_SCAN
    s.upload(kn, x0)
    s.solve()
    z, fl = s.solution().copy(), s.flags().copy()
    s.close()
    return z, fl


def _static(torch, n, m, N, kn, x0):
    """one call of the new entry point on caller-owned device buffers; soln / flags start as garbage"""
    B = x0.shape[0]
    dev = torch.device("cuda", 0)
    d_kn = torch.from_numpy(np.ascontiguousarray(kn)).to(dev)
    d_x0 = torch.from_numpy(np.ascontiguousarray(x0)).to(dev)
    d_z = torch.full((B, N * (2 * n + m)), float("nan"), dtype=torch.float64, device=dev)
    d_fl = torch.full((B,), 7, dtype=torch.int32, device=dev)
    st = torch.cuda.current_stream(dev)
    torch.cuda.synchronize(dev)
    free0 = torch.cuda.mem_get_info(dev)[0]
    rc = R.LIB.rslqr_cuda_solve_static(n, m, N, B, 0, d_kn.data_ptr(), d_x0.data_ptr(), d_z.data_ptr(), d_fl.data_ptr(),
                                       C.c_void_p(st.cuda_stream))
    assert rc == 0, R.LIB.rslqr_cuda_last_error().decode()
    torch.cuda.synchronize(dev)
    free1 = torch.cuda.mem_get_info(dev)[0]
    return d_z.cpu().numpy(), d_fl.cpu().numpy(), free0 - free1


def test_solve_static_matches_the_handle_solve_on_two_calls_in_a_row():
    import torch
    n, m = 12, 4
    # call 1: 20 problems of N = 256 (19.6 MB of the 32 MiB array); call 2: other problems, another horizon and
    # batch, over the first call's leftovers; then the first ones again
    cases = [(256, 20, 5), (64, 9, 6), (16, 100, 7), (256, 20, 5)]  # (batch >= 64 takes the other tier arity)
    grew = []
    for N, B, seed in cases:
        kn, x0 = R.synth_problems(n, m, N, seed, 0, B)
        want, wflags = _by_handle(n, m, N, kn, x0)
        z, fl, d = _static(torch, n, m, N, kn, x0)
        grew.append(d)
        assert not fl.any() and not wflags.any()
        for p in range(B):
            assert max(rel_err_classes(z[p], want[p], n, m, N)) < TOL
    # the first call may load the kernels' code lazily; after that the entry point must not take device memory
    assert grew[-1] == 0


def test_solve_static_other_shape_and_failure_flag():
    import torch
    n, m, N, B = 6, 3, 128, 5
    kn, x0 = R.synth_problems(n, m, N, 8, 0, B)
    kn = kn.copy()
    kn.reshape(B, N, -1)[3, 17, 2] = -1.0  # Q(2) of knot 17 of problem 3: the reference's Cholesky fails there
    want, wflags = _by_handle(n, m, N, kn, x0)
    for _ in range(2):
        z, fl, _ = _static(torch, n, m, N, kn, x0)
        assert np.array_equal(fl != 0, wflags != 0) and fl[3] != 0
        for p in range(B):
            if not fl[p]:
                assert max(rel_err_classes(z[p], want[p], n, m, N)) < TOL


def test_solve_static_rejects_what_does_not_fit():
    import torch
    dev = torch.device("cuda", 0)
    one = torch.zeros(16, dtype=torch.float64, device=dev)
    fl = torch.zeros(4, dtype=torch.int32, device=dev)
    # 64 problems of n=12, m=4, N=256 need ~62 MB: beyond RSLQR_STATIC_SCRATCH_BYTES; rejected before any launch
    rc = R.LIB.rslqr_cuda_solve_static(12, 4, 256, 64, 0, one.data_ptr(), one.data_ptr(), one.data_ptr(), fl.data_ptr(), None)
    assert rc == E_ARG
    msg = R.LIB.rslqr_cuda_last_error().decode()
    assert "RSLQR_STATIC_SCRATCH_BYTES" in msg and "bytes of scratch" in msg
    # (20, 7) has no instance of the scan kernels (RSLQR_DUAL_SHAPES): rejected before any launch
    assert R.LIB.rslqr_cuda_solve_static(20, 7, 256, 1, 0, one.data_ptr(), one.data_ptr(), one.data_ptr(), fl.data_ptr(), None) == -5
    assert R.LIB.rslqr_cuda_solve_static(12, 4, 200, 1, 0, one.data_ptr(), one.data_ptr(), one.data_ptr(), fl.data_ptr(), None) == E_ARG
    assert R.LIB.rslqr_cuda_solve_static(12, 4, 256, 1, 0, None, one.data_ptr(), one.data_ptr(), fl.data_ptr(), None) == E_ARG
