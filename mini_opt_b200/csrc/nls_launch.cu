// Launchers of the nonlinear least squares kernels (nls_chain.cuh, nls_linearize_dmma.cuh). Called from capi.cu.
#include <atomic>

#include "nls_solve_launch.cuh"
#include "nls_linearize_dmma.cuh"

namespace mo {
namespace {

#ifndef MO_IK_WARPS
#define MO_IK_WARPS 24
#endif
std::atomic<int> g_ik_warps{MO_IK_WARPS};  // mo_debug_set_ik_warps
#ifndef MO_IK_SYNC
#define MO_IK_SYNC 1
#endif
std::atomic<int> g_ik_sync{MO_IK_SYNC};    // mo_debug_set_ik_sync: CTA barriers after phase A (1) and phase B (2)

ChainArgs make_chain_args(const mo_chain_ik_batch* b) {
  ChainArgs a{};
  a.links = b->links; a.n = b->n; a.n_cost = b->n_cost; a.cost_block = b->cost_block; a.n_eq = b->n_eq; a.m = b->m;
  a.batch = b->batch; a.link_pose = b->link_pose; a.link_mask = b->link_mask; a.cost_w = b->cost_w;
  a.cost_target = b->cost_target; a.eq_w = b->eq_w; a.eq_target = b->eq_target; a.ineq_var = b->ineq_var;
  a.ineq_a = b->ineq_a; a.ineq_b = b->ineq_b; a.x0 = b->x0;
  return a;
}

}  // namespace

int nls_launch_linearize(const mo_nls_linearize_args* g, const NlsLaunchInfo& dev, cudaStream_t stream, std::string* err) {
  LinearizeArgs a{g->n, g->r_cost, g->cost_block, g->r_eq, g->m, g->batch, g->x, g->J_cost, g->r_cost_v, g->J_eq,
                  g->r_eq_v, g->lambda, g->ineq_var, g->ineq_a, g->ineq_b, g->G, g->c, g->A_eq, g->b_eq,
                  g->ineq_b_shifted, g->errors};
  // n = 32 .. 64 (multiple of 8), even row count, 16-byte aligned J / r: the DMMA kernel (nls_linearize_dmma.cuh)
  auto aligned16 = [](const void* ptr) { return (reinterpret_cast<uintptr_t>(ptr) & 15u) == 0; };
  if (g->n % 8 == 0 && g->n >= 32 && g->n <= 64 && g->r_cost >= 8 && g->r_cost % 2 == 0 && aligned16(g->J_cost) &&
      aligned16(g->r_cost_v)) {
    auto launch = [&](auto kernel) -> int {
      int bps = 0;
      NLS_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, kernel, 32, 0));
      const long long cap = (long long)dev.sms * (bps > 0 ? bps : 1);
      kernel<<<(int)(g->batch < cap ? g->batch : cap), 32, 0, stream>>>(a);
      NLS_CUDA(cudaGetLastError());
      return MO_OK;
    };
    switch (g->n / 8) {
      case 4: return launch(nls_linearize_dmma_kernel<4>);
      case 5: return launch(nls_linearize_dmma_kernel<5>);
      case 6: return launch(nls_linearize_dmma_kernel<6>);
      case 7: return launch(nls_linearize_dmma_kernel<7>);
      default: return launch(nls_linearize_dmma_kernel<8>);
    }
  }
  const long long cap = (long long)dev.sms * 32;
  nls_linearize_kernel<<<(int)(g->batch < cap ? g->batch : cap), 32, 0, stream>>>(a);
  NLS_CUDA(cudaGetLastError());
  return MO_OK;
}

int nls_launch_chain_fk(const mo_chain_ik_batch* chain, const double* params, double* translation, double* rotation_q,
                        double* t_D_params, double* r_D_params, const NlsLaunchInfo& dev, cudaStream_t stream,
                        std::string* err) {
  const size_t smem = sizeof(double) * chain_smem_doubles(chain->links, chain->n);
  int grid = 0;
  const int rc = nls_block_grid(chain_fk_kernel, 32, smem, chain->batch, dev, &grid, err);
  if (rc) return rc;
  chain_fk_kernel<<<grid, 32, smem, stream>>>(make_chain_args(chain), params, translation, rotation_q, t_D_params, r_D_params);
  NLS_CUDA(cudaGetLastError());
  return MO_OK;
}

int nls_launch_chain_ik(const mo_chain_ik_batch* problems, const mo_nls_params* params, const mo_nls_out* out,
                        const NlsLaunchInfo& dev, cudaStream_t stream, std::string* err) {
  const ChainArgs ca = make_chain_args(problems);
  const int mask = g_ik_sync.load();
  // warps per CTA marching in lock-step (nls_chain.cuh), one CTA per SM. Measured (profiles/r02_ik_lockstep_ab.jsonl, 6 links,
  // 100K problems): 16 warps (128 registers) 90.3 ms, 24 (80 registers, spills) 79.2, 32 (64 registers) 79.9, two CTAs of 16: 87.4
  int w = g_ik_warps.load();
  if (w == 24) {    // one lock-step CTA of 24 warps per SM (85 registers per thread)
    if (sizeof(double) * (size_t)nls_per_warp_doubles<ChainFamily>(ca) * 24 <= dev.max_smem_optin)
      return nls_solve_launch<ChainFamily, 24, 1>(ca, params, out, dev, stream, mask, err);
    w = 16;
  }
  while (w > 1 && sizeof(double) * (size_t)nls_per_warp_doubles<ChainFamily>(ca) * w > dev.max_smem_optin) w /= 2;
  switch (w) {
    case 16: return nls_solve_launch<ChainFamily, 16, 1>(ca, params, out, dev, stream, mask, err);
    default: return nls_solve_launch<ChainFamily, 1, 16>(ca, params, out, dev, stream, mask, err);
  }
}

}  // namespace mo

// Measurement hook: warps per CTA of the lock-step chain-IK kernel (1, 16 or 24; 1 = every warp on its own, the round-1
// schedule; 24 = the default, 80 registers per thread; shapes whose 24 regions exceed shared memory run with 16). Returns the previous value. Results do not depend on it (bit-identical), only the time does.
extern "C" int mo_debug_set_ik_warps(int warps) {
  if (warps != 1 && warps != 16 && warps != 24) return -1;
  return mo::g_ik_warps.exchange(warps);
}

// Measurement hook: which of the two optional CTA barriers of a lock-step pass are taken (bit 0: after the linearization,
// bit 1: after the inner QP); the barrier at the top of a pass always is. Results do not depend on it.
extern "C" int mo_debug_set_ik_sync(int mask) { return mo::g_ik_sync.exchange(mask & 15); }
