"""Diagnostic (CPU only): which arithmetic difference makes the CUDA kernels' status differ from the reference on the
degenerate box QPs? Runs the C oracle in its diagnostic solve modes (oracle/qp_ip_oracle.c: oracle_set_solve_mode)
against the unmodified reference on identical bytes."""
import sys, json
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
sys.path.insert(0, str(Path(__file__).resolve().parent))
import numpy as np
from mini_opt_b200 import abi, synth
from oracle import oracle_py as O
import parity

B = int(sys.argv[1]) if len(sys.argv) > 1 else 200000
arrs = synth.box_qp_batch_numpy(8, 2, B, seed=1234, explicit_constraints=False)
prm = synth.solve_params()
ref, _ = O.qp_solve("ref", arrs, prm, nthreads=8)
lib = O.oracle_lib()
for mode, variant in ((5, 0), (5, 1), (5, 2), (5, 3), (5, 7), (2, 3), (2, 7)):
    lib.oracle_set_solve_mode(mode)
    lib.oracle_set_variant(variant)
    got, _ = O.qp_solve("oracle", arrs, prm)
    r = parity.compare_qp(got, ref, 8)
    print(mode, variant, json.dumps({k: r[k] for k in ("status_equal", "iters_equal", "converged_gpu", "converged_ref", "status_mismatch_pairs") if k in r}))
lib.oracle_set_solve_mode(0)
