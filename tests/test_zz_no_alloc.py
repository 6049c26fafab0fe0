"""GPU test of the no-allocation entry point mo_qp_nullspace_solve_batch_noalloc (static device scratch, no cudaMalloc).
Collected last on purpose (file name). Passes alone in a fresh process and after the
other tests: it only needs the library and a device, and allocates its inputs / outputs BEFORE the calls under test."""
import ctypes as C

import numpy as np
import pytest

from mini_opt_b200 import abi, synth

pytestmark = pytest.mark.gpu


def test_nullspace_noalloc_matches_existing_entry_point_twice():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from mini_opt_b200 import lib, qp
    L = lib.load()
    dev = torch.device("cuda:0")
    stream = torch.cuda.current_stream(dev)
    # two different shapes in a row through the SAME static scratch: the second call sees the first call's leftovers
    for n, k, B in ((12, 3, 4000), (64, 16, 600), (8, 2, 5000)):
        arrs = qp.to_device(synth.box_qp_batch_numpy(n, k, B, seed=900 + n, explicit_constraints=False), dev)
        want_x, want_st = qp.nullspace_solve_device(arrs, dev)  # existing entry point (shared-memory workspace)
        x = torch.full((B, n), float("nan"), dtype=torch.float64, device=dev)
        st = torch.full((B,), -7, dtype=torch.int32, device=dev)
        b = arrs.to_c()
        torch.cuda.synchronize()
        for _ in range(2):  # two calls in a row, nothing allocated in between
            lib.check(L.mo_qp_nullspace_solve_batch_noalloc(C.byref(b), x.data_ptr(), st.data_ptr(), 0, stream.cuda_stream))
            torch.cuda.synchronize()
            assert np.array_equal(st.cpu().numpy(), want_st.cpu().numpy())
            # same device function, same arithmetic order: the tests' tolerance for x is relative 1e-9
            np.testing.assert_allclose(x.cpu().numpy(), want_x.cpu().numpy(), rtol=1e-9, atol=1e-12)
            x.fill_(float("nan"))
            st.fill_(-7)
    # beyond the stated maximum: a clear error, no launch
    big = abi.QPBatchArrays(n=abi.MO_NOALLOC_MAX_N + 8, k=2, m=0, batch=1, G=x, c=x, A_eq=x, b_eq=x)
    bb = big.to_c()
    rc = L.mo_qp_nullspace_solve_batch_noalloc(C.byref(bb), x.data_ptr(), st.data_ptr(), 0, stream.cuda_stream)
    assert rc == abi.MO_ERR_UNSUPPORTED_SHAPE
    assert b"MO_NOALLOC_MAX_N" in L.mo_last_error_string()
