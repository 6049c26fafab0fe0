// Instantiations of the device SQP + LM kernel for the functor residual families (nls_families.cuh) and their dispatch.
// Own translation unit: one kernel instantiation per family, compiled next to nls_launch.cu's.
#include "nls_families.cuh"
#include "nls_solve_launch.cuh"

// (MO_NLS_FAMILY_* id, functor struct)
#define MO_NLS_FUNCTOR_FAMILIES(X)                     \
  X(MO_NLS_FAMILY_ROSENBROCK, RosenbrockResidual)      \
  X(MO_NLS_FAMILY_HIMMELBLAU, HimmelblauResidual)      \
  X(MO_NLS_FAMILY_SPHERE_PRODUCT, SphereProductResidual)

namespace mo {

int nls_family_dims(int family, int* n, int* n_cost, int* cost_block, int* n_eq) {
  switch (family) {
#define X(ID, FN) case ID: *n = FN::N; *n_cost = FN::R_COST; *cost_block = FN::COST_BLOCK; *n_eq = FN::R_EQ; return MO_OK;
    MO_NLS_FUNCTOR_FAMILIES(X)
#undef X
    default: return MO_ERR_INVALID_ARGUMENT;
  }
}

int nls_launch_family(const mo_nls_family_batch* b, const mo_nls_params* params, const mo_nls_out* out,
                      const NlsLaunchInfo& dev, cudaStream_t stream, std::string* err) {
  FunctorArgs a{};
  int rc = nls_family_dims(b->family, &a.n, &a.n_cost, &a.cost_block, &a.n_eq);
  if (rc) { *err = "unknown residual family"; return rc; }
  a.m = b->m; a.batch = b->batch; a.ineq_var = b->ineq_var; a.ineq_a = b->ineq_a; a.ineq_b = b->ineq_b; a.x0 = b->x0;
  a.data = b->data; a.data_stride = b->data_stride;
  switch (b->family) {
#define X(ID, FN) case ID: return nls_solve_launch<FunctorFamily<FN>, 16, 1>(a, params, out, dev, stream, 3, err);
    MO_NLS_FUNCTOR_FAMILIES(X)
#undef X
    default: return MO_ERR_INVALID_ARGUMENT;
  }
}

}  // namespace mo
