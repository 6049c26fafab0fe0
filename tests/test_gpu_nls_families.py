"""Device-side residual families (mo_nls_solve_family_batch, csrc/nls_families.cuh: SURVEY 8(f) row 2) against the UNMODIFIED
reference: the reference's own nonlinear test problems (test/nonlinear_test.cc:375-500 Rosenbrock / Rosenbrock LM /
inequality constrained Rosenbrock, :668-718 Himmelblau in a quadrant, :745-830 sphere with product equality constraints),
every start point of a test as ONE device batch through the whole SQP + LM loop, compared start point by start point with
tests/golden/nls_host_reference.txt (written by tools/make_golden_nls_host.sh from the reference itself)."""
from pathlib import Path

import numpy as np
import pytest

from mini_opt_b200 import abi

pytestmark = pytest.mark.gpu
ROOT = Path(__file__).resolve().parents[1]
ROSENBROCK_STARTS = [(-5, -3), (10, 8), (-20, 3), (0, -5), (4, 0), (100, 50), (-35, 40), (1000, -50), (0.8, -0.3)]


def golden(name):
    ref = {}
    for line in (ROOT / "tests" / "golden" / "nls_host_reference.txt").read_text().splitlines():
        f = line.split()
        if f[0] == name:
            ref[int(f[1])] = (int(f[2]), int(f[3]), int(f[4]), float(f[5]), float(f[6]))
    return ref


def solve(family, x0, params, ineq=None, data=None):
    import torch
    from mini_opt_b200 import nls
    dev = torch.device("cuda:0")
    n, _, _ = nls.family_dims(family)
    x0 = np.ascontiguousarray(np.asarray(x0, dtype=np.float64).reshape(-1, n))
    out = nls.alloc_nls_out_torch(x0.shape[0], n, dev)
    dineq = None
    if ineq is not None:
        dineq = (torch.tensor(ineq[0], dtype=torch.int32, device=dev), torch.tensor(ineq[1], dtype=torch.float64, device=dev),
                 torch.tensor(ineq[2], dtype=torch.float64, device=dev))
    ddata = None if data is None else torch.from_numpy(np.ascontiguousarray(data)).to(dev)
    nls.solve_family_device(family, torch.from_numpy(x0).to(dev), params, out, dev, ineq=dineq, data=ddata)
    torch.cuda.synchronize()
    return nls.nls_out_to_numpy(out)


def rows(o):
    return {i: (int(o.termination[i]), int(o.iterations[i]), int(o.qp_iterations[i]), float(o.x[i, 0]), float(o.x[i, 1]))
            for i in range(o.x.shape[0])}


def test_family_dims():
    from mini_opt_b200 import lib, nls
    assert nls.family_dims(abi.MO_NLS_FAMILY_ROSENBROCK) == (2, 2, 0)
    assert nls.family_dims(abi.MO_NLS_FAMILY_HIMMELBLAU) == (2, 2, 0)
    assert nls.family_dims(abi.MO_NLS_FAMILY_SPHERE_PRODUCT) == (6, 6, 2)
    with pytest.raises(Exception):
        nls.family_dims(99)


@pytest.mark.parametrize("lm", [False, True])
def test_rosenbrock_on_device(lm):
    """nonlinear_test.cc:391-460. No constraints at all: the inner QP is one Newton step (m = 0, k = 0)."""
    p = abi.NLSParams(max_iterations=10 if lm else 5, max_qp_iterations=1)
    if lm:
        p.absolute_first_derivative_tol = 1e-12
        p.max_line_search_iterations = 0
    got, ref = rows(solve(abi.MO_NLS_FAMILY_ROSENBROCK, ROSENBROCK_STARTS, p)), golden("rosenbrock_lm" if lm else "rosenbrock")
    assert len(ref) == 9
    for i, r in ref.items():
        g = got[i]
        assert g[:3] == r[:3], (i, g, r)  # termination, outer iterations, QP iterations: identical
        assert abs(g[3] - r[3]) <= 1e-7 and abs(g[4] - r[4]) <= 1e-7, (i, g, r)
        assert abs(g[3] - 1.0) < 1e-6 and abs(g[4] - 1.0) < 1e-6


def test_rosenbrock_per_problem_data():
    """`data` reaches the functor: (a, b) = (1, 100) given explicitly is bit-identical to the defaults; (2, 100) moves the optimum to (2, 4)."""
    p = abi.NLSParams(max_iterations=30, max_qp_iterations=1)
    base = solve(abi.MO_NLS_FAMILY_ROSENBROCK, ROSENBROCK_STARTS, p)
    same = solve(abi.MO_NLS_FAMILY_ROSENBROCK, ROSENBROCK_STARTS, p, data=np.tile([1.0, 100.0], (9, 1)))
    assert np.array_equal(base.x, same.x) and np.array_equal(base.iterations, same.iterations)
    moved = solve(abi.MO_NLS_FAMILY_ROSENBROCK, ROSENBROCK_STARTS, p, data=np.tile([2.0, 100.0], (9, 1)))
    assert np.all(moved.termination != abi.MO_NLS_MAX_ITERATIONS)
    assert np.allclose(moved.x, [2.0, 4.0], atol=1e-5)


def test_inequality_constrained_rosenbrock_on_device():
    """nonlinear_test.cc:463-500: Var(0) >= 1.2, Var(1) <= 0.5 (one-sided, not box pairs)."""
    p = abi.NLSParams(max_iterations=10, max_qp_iterations=10)
    starts = [(12, -5), (100.0, -20.0), (1423.0, -400.0), (-20.0, 10.0), (-120.0, 35.0), (-50.0, 0.5)]
    # LinearInequalityConstraint (qp.hpp:28-66): a x + b >= 0.  x0 >= 1.2 -> (0, 1, -1.2);  x1 <= 0.5 -> (1, -1, 0.5)
    got = rows(solve(abi.MO_NLS_FAMILY_ROSENBROCK, starts, p, ineq=([0, 1], [1.0, -1.0], [-1.2, 0.5])))
    ref = golden("rosenbrock_ineq")
    assert len(ref) == 6
    for i, r in ref.items():
        g = got[i]
        assert g[:2] == r[:2], (i, g, r)
        assert abs(g[2] - r[2]) <= 2, (i, g, r)  # as in the host test: the G ~ 8e8 start crosses kkt_tol two steps early
        assert abs(g[3] - 1.2) < 1e-6 and abs(g[4] - 0.5) < 1e-6 and abs(g[3] - r[3]) <= 1e-7 and abs(g[4] - r[4]) <= 1e-7


def test_himmelblau_quadrant_on_device():
    """nonlinear_test.cc:668-718 on the reference's own 23 x 23 grid of start points, box 0.1 <= x, y <= 5."""
    import sys
    sys.path.insert(0, str(ROOT / "tests"))
    from test_cpp_host_api import compare_himmelblau
    p = abi.NLSParams(max_iterations=20, max_qp_iterations=10, relative_exit_tol=1e-12, absolute_first_derivative_tol=1e-8,
                      termination_kkt_tolerance=1e-6)
    starts = []
    x = 0.2
    while x <= 4.8:  # the reference's accumulating loops
        y = 0.2
        while y <= 4.8:
            starts.append((x, y))
            y += 0.2
        x += 0.2
    assert len(starts) == 529
    o = solve(abi.MO_NLS_FAMILY_HIMMELBLAU, starts, p, ineq=([0, 0, 1, 1], [1.0, -1.0, 1.0, -1.0], [-0.1, 5.0, -0.1, 5.0]))
    got = {("himmelblau_quadrant", i): v for i, v in rows(o).items()}
    ref = {("himmelblau_quadrant", i): v for i, v in golden("himmelblau_quadrant").items()}
    # measured on B200 (round 2): 529 / 529 identical termination, outer and QP iteration counts
    compare_himmelblau(got, ref, floors={"termination": 0.995, "termination_and_outer": 0.995, "all_counts": 0.995}, tag="device")


class MT19937:
    """std::mt19937 (the start points of the sphere test come from std::mt19937{7} + std::uniform_real_distribution)."""

    def __init__(self, seed):
        self.mt = [0] * 624
        self.mt[0] = seed & 0xFFFFFFFF
        for i in range(1, 624):
            self.mt[i] = (1812433253 * (self.mt[i - 1] ^ (self.mt[i - 1] >> 30)) + i) & 0xFFFFFFFF
        self.idx = 624

    def __call__(self):
        if self.idx >= 624:
            mt = self.mt
            for i in range(624):
                y = (mt[i] & 0x80000000) | (mt[(i + 1) % 624] & 0x7FFFFFFF)
                mt[i] = mt[(i + 397) % 624] ^ (y >> 1) ^ (0x9908B0DF if y & 1 else 0)
            self.idx = 0
        y = self.mt[self.idx]
        self.idx += 1
        y ^= y >> 11
        y ^= (y << 7) & 0x9D2C5680
        y ^= (y << 15) & 0xEFC60000
        return y ^ (y >> 18)

    def uniform(self, lo, hi):  # libstdc++ generate_canonical<double, 53>: two draws, low word first
        s = float(self()) + float(self()) * 4294967296.0
        r = s / 18446744073709551616.0
        if r >= 1.0:
            r = np.nextafter(1.0, 0.0)
        return r * (hi - lo) + lo


def test_sphere_with_product_equalities_on_device():
    """nonlinear_test.cc:745-830: equality rows and no inequalities -> the null-space solver and the (18.36) penalty, like
    the reference. 100 seeded start points; identical termination, outer and QP counts; x to 2e-6 (first-order exit)."""
    eng = MT19937(7)
    starts = [[eng.uniform(-30.0, 30.0) for _ in range(6)] for _ in range(100)]
    p = abi.NLSParams(max_iterations=100, max_qp_iterations=1, relative_exit_tol=1e-12, absolute_first_derivative_tol=1e-9,
                      termination_kkt_tolerance=1e-6, lambda_initial=0.001)
    o = solve(abi.MO_NLS_FAMILY_SPHERE_PRODUCT, starts, p)
    got, ref = rows(o), golden("sphere")
    assert len(ref) == 100
    for i, r in ref.items():
        g = got[i]
        assert g[:3] == r[:3], (i, g, r)
        assert abs(g[3] - r[3]) <= 2e-6 * max(1.0, abs(r[3])) and abs(g[4] - r[4]) <= 2e-6 * max(1.0, abs(r[4])), (i, g, r)
    x = o.x
    assert np.allclose(np.abs(x[:, 0] * x[:, 1]), 4.0, atol=1e-5) and np.allclose(np.abs(x[:, 2] * x[:, 3]), 9.0, atol=1e-5)
    assert np.allclose(x[:, 4:], 0.0, atol=1e-5)
