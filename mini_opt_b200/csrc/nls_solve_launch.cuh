// Host-side launch of nls_solve_kernel<Family, W, MINB> (nls_chain.cuh), shared by the translation units that instantiate
// the kernel for their residual families (nls_launch.cu: transform chains; nls_families.cu: functor families).
#pragma once
#include <atomic>
#include <string>

#include "nls_chain.cuh"
#include "nls_launch.h"

namespace mo {

inline int nls_cuda_err(cudaError_t e, const char* what, std::string* err) {
  *err = std::string(what) + ": " + cudaGetErrorString(e);
  return MO_ERR_CUDA;
}
#define NLS_CUDA(call)                                          \
  do {                                                          \
    cudaError_t e_ = (call);                                    \
    if (e_ != cudaSuccess) return ::mo::nls_cuda_err(e_, #call, err); \
  } while (0)

template <typename K>
int nls_block_grid(K kernel, int threads, size_t smem, long long blocks_wanted, const NlsLaunchInfo& dev, int* grid,
                   std::string* err) {
  if (smem > dev.max_smem_optin) {
    *err = "problem too large for shared memory: needs " + std::to_string(smem) + " bytes";
    return MO_ERR_UNSUPPORTED_SHAPE;
  }
  int bps = 0;
  NLS_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  NLS_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, kernel, threads, smem));
  const long long cap = (long long)dev.sms * (bps > 0 ? bps : 1);
  *grid = (int)(blocks_wanted < cap ? blocks_wanted : cap);
  return MO_OK;
}

inline NlsParams nls_make_params(const mo_nls_params* params) {
  return NlsParams{params->max_iterations, params->max_qp_iterations, params->termination_kkt_tolerance,
                   params->absolute_exit_tol, params->relative_exit_tol, params->absolute_first_derivative_tol,
                   params->max_line_search_iterations, params->line_search_strategy, params->armijo_search_tau,
                   params->equality_penalty_initial, params->equality_penalty_scale_factor,
                   params->equality_penalty_rho, params->lambda_initial, params->lambda_failure_init,
                   params->lambda_decrease_on_success, params->lambda_decrease_on_restore, params->max_lambda,
                   params->min_lambda, params->retraction, params->inner_qp_solver};
}

// Per-warp shared-memory doubles of a family's problem shape (even, so that every warp region stays 16-byte aligned).
template <class F>
long long nls_per_warp_doubles(const typename F::Args& a) {
  return (long long)((nls_smem_doubles(F::smem_doubles(a), a.n, a.n_cost, a.n_eq, a.m) + 1) & ~size_t(1));
}

// One launch of the lock-step SQP kernel with W warps per CTA (W x MINB = 16 resident warps per SM).
template <class F, int W, int MINB>
int nls_solve_launch(const typename F::Args& a, const mo_nls_params* params, const mo_nls_out* out, const NlsLaunchInfo& dev,
                     cudaStream_t stream, int sync_mask, std::string* err) {
  const NlsParams p = nls_make_params(params);
  const NlsOut o{out->x, out->termination, out->iterations, out->qp_iterations, out->ls_steps, out->errors};
  // ticket counter of this launch: a slot of a small ring in device memory, zeroed on the launch stream
  static std::atomic<unsigned> ring{0};
  unsigned long long* tickets = nullptr;
  NLS_CUDA(cudaGetSymbolAddress(reinterpret_cast<void**>(&tickets), g_ik_tickets));
  tickets += ring.fetch_add(1) % kIkTicketSlots;
  NLS_CUDA(cudaMemsetAsync(tickets, 0, sizeof(unsigned long long), stream));
  const long long per_warp = nls_per_warp_doubles<F>(a);
  const size_t smem = sizeof(double) * (size_t)per_warp * W;
  auto kernel = nls_solve_kernel<F, W, MINB>;
  int grid = 0;
  const int rc = nls_block_grid(kernel, 32 * W, smem, (a.batch + W - 1) / W, dev, &grid, err);
  if (rc) return rc;
  kernel<<<grid, 32 * W, smem, stream>>>(a, p, o, tickets, per_warp, sync_mask);
  NLS_CUDA(cudaGetLastError());
  return MO_OK;
}

}  // namespace mo
