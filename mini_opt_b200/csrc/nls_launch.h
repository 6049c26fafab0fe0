// Interface between capi.cu (argument checking, device selection) and nls_launch.cu (the nonlinear least squares kernels:
// linearization, chain FK, the device SQP + LM loop). Own translation unit: the IK kernel is under active tuning and this
// keeps its rebuild short.
#pragma once
#include <cuda_runtime.h>

#include <string>

#include "mini_opt_b200.h"

namespace mo {

constexpr int kMaxChainLinks = 32;
constexpr int kMaxChainRows = 8;  // cost / equality rows per problem

struct NlsLaunchInfo {
  int sms;
  size_t max_smem_optin;
};

// Each returns MO_OK, MO_ERR_UNSUPPORTED_SHAPE or MO_ERR_CUDA (message in *err); arguments are already validated and
// the target device is current. One kernel launch each on `stream`.
int nls_launch_linearize(const mo_nls_linearize_args* g, const NlsLaunchInfo& dev, cudaStream_t stream, std::string* err);
int nls_launch_chain_fk(const mo_chain_ik_batch* chain, const double* params, double* translation, double* rotation_q,
                        double* t_D_params, double* r_D_params, const NlsLaunchInfo& dev, cudaStream_t stream,
                        std::string* err);
int nls_launch_chain_ik(const mo_chain_ik_batch* problems, const mo_nls_params* params, const mo_nls_out* out,
                        const NlsLaunchInfo& dev, cudaStream_t stream, std::string* err);

// functor residual families (nls_families.cu)
int nls_family_dims(int family, int* n, int* n_cost, int* cost_block, int* n_eq);
int nls_launch_family(const mo_nls_family_batch* problems, const mo_nls_params* params, const mo_nls_out* out,
                      const NlsLaunchInfo& dev, cudaStream_t stream, std::string* err);

}  // namespace mo
