"""CPU tests of the boundary: the product library loads (no GPU needed for that) and exports every function that
include/mini_opt_b200.h declares; struct layouts of the ctypes mirror match the header's field order; parameter
defaults equal the reference's; host-side argument validation runs before any CUDA call."""
import ctypes as C
import re
from pathlib import Path

import numpy as np
import pytest

from mini_opt_b200 import abi, lib, synth

ROOT = Path(__file__).resolve().parents[1]


def declared_functions():
    text = (ROOT / "include" / "mini_opt_b200.h").read_text()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(mo_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    L = lib.load()
    names = declared_functions()
    assert len(names) >= 16
    for name in names:
        assert hasattr(L, name), f"{name} declared in include/mini_opt_b200.h but not exported"
    assert set(lib.EXPORTS) == set(names)


def test_header_struct_fields_match_ctypes():
    text = (ROOT / "include" / "mini_opt_b200.h").read_text()
    for cname, cls in (("mo_qp_batch", abi.MoQpBatch), ("mo_qp_params", abi.MoQpParams), ("mo_qp_out", abi.MoQpOut),
                       ("mo_nls_params", abi.MoNlsParams), ("mo_nls_linearize_args", abi.MoNlsLinearizeArgs),
                       ("mo_chain_ik_batch", abi.MoChainIkBatch), ("mo_nls_out", abi.MoNlsOut)):
        body = re.search(r"typedef struct %s \{(.*?)\} %s;" % (cname, cname), text, flags=re.S).group(1)
        body = re.sub(r"/\*.*?\*/", "", body, flags=re.S)
        fields = [re.split(r"[\s\*]+", d.strip())[-1] for d in body.split(";") if d.strip()]
        mirror = [f[0] for f in cls._fields_]
        mirror = ["lambda" if f == "lam" else f for f in mirror]
        assert fields == mirror, (cname, fields, mirror)


def test_param_defaults_match_reference():
    L = lib.load()
    p = abi.MoQpParams()
    L.mo_qp_params_default(C.byref(p))
    d = abi.QPParams().to_c()
    for f, _ in abi.MoQpParams._fields_:
        assert getattr(p, f) == getattr(d, f), f
    assert (p.initial_mu, p.sigma, p.termination_kkt_tol, p.max_iterations) == (1.0, 0.5, 1e-9, 10)  # qp.hpp:136-163
    q = abi.MoNlsParams()
    L.mo_nls_params_default(C.byref(q))
    e = abi.NLSParams().to_c()
    for f, _ in abi.MoNlsParams._fields_:
        assert getattr(q, f) == getattr(e, f), f
    assert L.mo_qp_vars_stride(8, 2, 16) == 42 and L.mo_abi_version() == abi.MO_ABI_VERSION


def test_argument_validation_precedes_cuda():
    """Bad arguments are rejected on the host (no device needed)."""
    from mini_opt_b200 import qp
    arrs = synth.box_qp_batch_numpy(8, 2, 4, seed=1, explicit_constraints=False)
    out = abi.alloc_out_numpy(4, 8, 2, 16)
    with pytest.raises(lib.MiniOptError) as e:
        qp.solve_batch_host(arrs, abi.QPParams(sigma=2.0), out, 0)
    assert e.value.code == abi.MO_ERR_INVALID_PARAMS
    arrs.G = None
    with pytest.raises(lib.MiniOptError) as e:
        qp.solve_batch_host(arrs, abi.QPParams(), out, 0)
    assert e.value.code == abi.MO_ERR_INVALID_ARGUMENT
    assert lib.kernel_variant(8, 2, 16, abi.MO_QP_FLAG_BOX_PAIRS) == "tpp"
    assert lib.kernel_variant(7, 2, 14, abi.MO_QP_FLAG_BOX_PAIRS) == "warp16"  # every n + k <= 16, m <= 32 shape
    assert lib.kernel_variant(12, 3, 24, 0) == "warp16"
    assert lib.kernel_variant(20, 4, 40, abi.MO_QP_FLAG_BOX_PAIRS).startswith("generic")


def test_no_cpu_fallback_without_device():
    """On a box without a GPU the compute entry points fail loudly instead of computing on the host."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from mini_opt_b200 import qp
    arrs = synth.box_qp_batch_numpy(8, 2, 4, seed=1, explicit_constraints=False)
    out = abi.alloc_out_numpy(4, 8, 2, 16)
    with pytest.raises(lib.MiniOptError) as e:
        qp.solve_batch_host(arrs, synth.solve_params(), out, 0)
    assert e.value.code in (abi.MO_ERR_NO_DEVICE, abi.MO_ERR_CUDA)
    assert (out.status == -1).all() and not np.any(out.vars)
