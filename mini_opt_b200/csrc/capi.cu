// C ABI of mini_opt_b200 (include/mini_opt_b200.h): argument checking, kernel selection, launches,
// and the host-buffer convenience entry point. No torch, no Eigen: plain CUDA runtime.
#include <nvtx3/nvToolsExt.h>

#include <sys/syscall.h>
#include <unistd.h>

#include <atomic>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "qp_dmma.cuh"
#include "wd_launch.h"
#include "qp_generic.cuh"
#include "qp_tpp.cuh"
#include "qp_w16.cuh"

namespace {

thread_local std::string g_last_error;
std::atomic<long long> g_launches{0};
#ifndef MO_TPP_ASYNC
#define MO_TPP_ASYNC 1
#endif
// refill of the tpp kernel: 1 = cp.async into the idle lane's column one round ahead (default), 0 = LDG/STS (A/B, mo_debug_set_tpp_async)
static std::atomic<int> g_tpp_async{MO_TPP_ASYNC};
// staging mode of the warp16 kernel: 1 = TMA bulk copies (default), 0 = lane loads (A/B measurements, mo_debug_set_tma)
std::atomic<int> g_w16_tma{1};

int cuda_fail(cudaError_t e, const char* what) {
  g_last_error = std::string(what) + ": " + cudaGetErrorString(e);
  return MO_ERR_CUDA;
}
#define MO_CUDA(call)                                  \
  do {                                                 \
    cudaError_t e_ = (call);                           \
    if (e_ != cudaSuccess) return cuda_fail(e_, #call); \
  } while (0)

// NVTX range per entry point: the batched stand-in for the reference's opt-in MINI_OPT_FUNCTION_TRACE scoped timers
// (include/mini_opt/tracing.hpp:67-83; call sites source/nonlinear.cc:77,173,218,280,355,459). Header-only NVTX v3: a
// no-op unless a tool (nsys, ncu --nvtx) is attached.
struct NvtxRange {
  explicit NvtxRange(const char* name) { nvtxRangePushA(name); }
  ~NvtxRange() { nvtxRangePop(); }
};
#define MO_TRACE() NvtxRange mo_nvtx_range_(__func__)

struct DeviceGuard {
  int prev = -1;
  bool ok = false;
  explicit DeviceGuard(int dev) {
    if (cudaGetDevice(&prev) != cudaSuccess) return;
    ok = (prev == dev) || cudaSetDevice(dev) == cudaSuccess;
    if (prev == dev) prev = -1;
  }
  ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
};

struct DeviceInfo {
  bool queried = false;
  int sms = 0, cc_major = 0, max_smem_optin = 0;
};
DeviceInfo g_dev[64];
std::mutex g_dev_mu;

int device_info(int dev, DeviceInfo* out) {
  if (dev < 0 || dev >= 64) { g_last_error = "device index out of range"; return MO_ERR_NO_DEVICE; }
  std::lock_guard<std::mutex> lk(g_dev_mu);
  DeviceInfo& d = g_dev[dev];
  if (!d.queried) {
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || dev >= count) {
      cudaGetLastError();
      g_last_error = "no CUDA device " + std::to_string(dev);
      return MO_ERR_NO_DEVICE;
    }
    MO_CUDA(cudaDeviceGetAttribute(&d.sms, cudaDevAttrMultiProcessorCount, dev));
    MO_CUDA(cudaDeviceGetAttribute(&d.cc_major, cudaDevAttrComputeCapabilityMajor, dev));
    MO_CUDA(cudaDeviceGetAttribute(&d.max_smem_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev));
    d.queried = true;
  }
  if (d.cc_major != 10) {
    g_last_error = "mini_opt_b200 kernels are built for sm_100a only; device has compute capability major " +
                   std::to_string(d.cc_major);
    return MO_ERR_NO_DEVICE;
  }
  *out = d;
  return MO_OK;
}

// CheckParams, qp.cc:76-82
int check_params(const mo_qp_params* p) {
  if (!(p->initial_mu > 0) || !(p->sigma > 0) || !(p->sigma <= 1.0) || !(p->termination_kkt_tol > 0) ||
      !(p->max_iterations > 0))
    return MO_ERR_INVALID_PARAMS;
  if (p->barrier_strategy < 0 || p->barrier_strategy > 2 || p->initial_guess_method < 0 || p->initial_guess_method > 2)
    return MO_ERR_INVALID_PARAMS;
  return MO_OK;
}

int check_batch(const mo_qp_batch* b) {
  if (!b || b->n <= 0 || b->k < 0 || b->m < 0 || b->batch < 0) return MO_ERR_INVALID_ARGUMENT;
  if (b->batch == 0) return MO_OK;
  if (!b->G || !b->c) return MO_ERR_INVALID_ARGUMENT;
  if (b->k > 0 && (!b->A_eq || !b->b_eq)) return MO_ERR_INVALID_ARGUMENT;
  if (b->m > 0) {
    if (!b->ineq_b) return MO_ERR_INVALID_ARGUMENT;
    if (b->flags & MO_QP_FLAG_BOX_PAIRS) {
      if (b->m != 2 * b->n) return MO_ERR_INVALID_ARGUMENT;
    } else if (!b->ineq_var || !b->ineq_a) {
      return MO_ERR_INVALID_ARGUMENT;
    }
  }
  if ((b->flags & MO_QP_FLAG_UNIFORM_BOX) && !(b->flags & MO_QP_FLAG_BOX_PAIRS)) return MO_ERR_INVALID_ARGUMENT;
  return MO_OK;
}

constexpr int kHostOnlyFlags = MO_QP_FLAG_PACKED_G | MO_QP_FLAG_UNIFORM_BOX;

mo::QPArgs make_args(const mo_qp_batch* b, const mo_qp_params* p, const mo_qp_out* o) {
  mo::QPArgs a{};
  a.n = b->n; a.k = b->k; a.m = b->m; a.flags = b->flags; a.batch = b->batch;
  a.G = b->G; a.c = b->c; a.A = b->A_eq; a.b = b->b_eq;
  a.ivar = b->ineq_var; a.ia = b->ineq_a; a.ib = b->ineq_b; a.icount = b->ineq_count;
  if (p) {
    a.initial_mu = p->initial_mu; a.sigma = p->sigma; a.kkt_tol = p->termination_kkt_tol;
    a.comp_tol = p->termination_complementarity_tol; a.max_iterations = p->max_iterations;
    a.strategy = p->barrier_strategy; a.decrease_mu_only_on_small_error = p->decrease_mu_only_on_small_error;
    a.guess = p->initial_guess_method; a.init_mu_with_comp = p->initialize_mu_with_complementarity;
  }
  if (o) {
    a.vars = o->vars; a.status = o->status; a.iters = o->iters; a.kkt_final = o->kkt_final;
    a.lagrange = o->lagrange; a.iter_log = o->iter_log;
  }
  return a;
}

}  // namespace

// ---------------------------------------------------------------------------------------------
// kernel selection
// ---------------------------------------------------------------------------------------------
namespace {

constexpr int kTppWarpsPerBlock = 4;
constexpr int kTppMinBlocks = 2;

// (N, K) shapes with a thread-per-problem instantiation (box pairs, m == 2n, uniform count).
#define MO_TPP_SHAPES(X) X(8, 2) X(8, 0) X(4, 1)

template <int N, int K, bool ASYNC>
int launch_tpp_impl(const mo::QPArgs& a, const DeviceInfo& dev, cudaStream_t stream) {
  constexpr int threads = 32 * kTppWarpsPerBlock;
  auto guess = mo::qp_tpp_guess_kernel<N, K>;
  auto iterate = mo::qp_tpp_iterate_kernel<N, K, kTppWarpsPerBlock, kTppMinBlocks, ASYNC>;
  const size_t smem_guess = 0;
  const size_t smem_iter = mo::tpp_iterate_smem_bytes<N, K, kTppWarpsPerBlock>(ASYNC);
  static thread_local int configured_dev = -1;
  static thread_local int bps_guess = 0, bps_iter = 0;
  int cur = 0;
  MO_CUDA(cudaGetDevice(&cur));
  if (configured_dev != cur) {
    MO_CUDA(cudaFuncSetAttribute(iterate, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_iter));
    MO_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps_guess, guess, threads, smem_guess));
    MO_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps_iter, iterate, threads, smem_iter));
    if (bps_guess < 1 || bps_iter < 1) { g_last_error = "tpp kernels do not fit on an SM"; return MO_ERR_UNSUPPORTED_SHAPE; }
    configured_dev = cur;
  }
  const long long want = (a.batch + threads - 1) / threads;
  const bool user_guess = a.guess == MO_GUESS_USER_PROVIDED;
  if (!user_guess) {  // ComputeInitialGuess -> start state in vars
    const long long cap = (long long)dev.sms * bps_guess;
    guess<<<(int)(want < cap ? want : cap), threads, smem_guess, stream>>>(a);
    g_launches.fetch_add(1);
    MO_CUDA(cudaGetLastError());
  }
  const long long cap = (long long)dev.sms * bps_iter;
  iterate<<<(int)(want < cap ? want : cap), threads, smem_iter, stream>>>(a, user_guess ? 0 : 1);
  g_launches.fetch_add(1);
  MO_CUDA(cudaGetLastError());
  return MO_OK;
}

template <int N, int K>
int launch_tpp(const mo::QPArgs& a, const DeviceInfo& dev, cudaStream_t stream) {
  return g_tpp_async.load() ? launch_tpp_impl<N, K, true>(a, dev, stream) : launch_tpp_impl<N, K, false>(a, dev, stream);
}

bool tpp_eligible(int n, int k, int m, int flags, bool ragged, int strategy) {
  if (!(flags & MO_QP_FLAG_BOX_PAIRS) || ragged || m != 2 * n) return false;
  if (strategy == MO_BARRIER_PREDICTOR_CORRECTOR) return false;
#define X(N_, K_) if (n == N_ && k == K_) return true;
  MO_TPP_SHAPES(X)
#undef X
  return false;
}

// (N, K, warps per CTA, min CTAs per SM) shapes with a DMMA CTA-per-problem instantiation (box pairs, m == 2n).
#ifndef MO_DM64_W
#define MO_DM64_W 4     /* measured on B200, 64K QPs n=64: W=4/MINB=4 35.7 ms, W=4/MINB=5 41.9, W=8/MINB=4 61.7 */
#define MO_DM64_MINB 4
#endif
#define MO_DM_SHAPES(X) X(64, 16, MO_DM64_W, MO_DM64_MINB) X(32, 8, 2, 8) X(16, 8, 1, 16)

template <int N, int K, int W, int MINB>
int launch_dm(const mo::QPArgs& a, const DeviceInfo& dev, cudaStream_t stream) {
  using L = mo::DmLayout<N, K, W>;
  auto kernel = mo::qp_dm_kernel<N, K, W, MINB>;
  static thread_local int configured_dev = -1;
  static thread_local int bps = 0;
  int cur = 0;
  MO_CUDA(cudaGetDevice(&cur));
  if (configured_dev != cur) {
    MO_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)L::SMEM_BYTES));
    MO_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, kernel, L::THREADS, L::SMEM_BYTES));
    if (bps < 1) { g_last_error = "dm kernel does not fit on an SM"; return MO_ERR_UNSUPPORTED_SHAPE; }
    configured_dev = cur;
  }
  const long long cap = (long long)dev.sms * bps;
  kernel<<<(int)(a.batch < cap ? a.batch : cap), L::THREADS, L::SMEM_BYTES, stream>>>(a);
  g_launches.fetch_add(1);
  MO_CUDA(cudaGetLastError());
  return MO_OK;
}

// mo_debug_set_dm_kernel(1) selects the CTA-per-problem kernel (qp_dmma.cuh), the cross-check partner of the default
// warp-per-problem kernel, for the variant-agreement test and A/B measurements. No environment variables are read.
std::atomic<int> g_dm_kernel{0};
bool dm_use_cta_kernel() { return g_dm_kernel.load() == 1; }

bool dm_eligible(int n, int k, int m, int flags, bool ragged, int strategy) {
  if (!(flags & MO_QP_FLAG_BOX_PAIRS) || ragged || m != 2 * n) return false;
  if (strategy == MO_BARRIER_PREDICTOR_CORRECTOR) return false;
#define X(N_, K_, W_, B_) if (n == N_ && k == K_) return true;
  MO_DM_SHAPES(X)
#undef X
  return false;
}

int generic_threads(int n, int k) {
  const int p = n + k;
  return p <= 32 ? 32 : (p <= 64 ? 128 : 256);
}

template <int THREADS>
int launch_generic(const mo::QPArgs& a, const mo::StepArgs* step, const DeviceInfo& dev, cudaStream_t stream) {
  const size_t smem = sizeof(double) * mo::gen_smem_doubles(a.n, a.k, a.m, THREADS);
  if (smem > (size_t)dev.max_smem_optin) {
    g_last_error = "problem too large for shared memory: needs " + std::to_string(smem) + " bytes";
    return MO_ERR_UNSUPPORTED_SHAPE;
  }
  int bps = 0;
  if (step) {
    auto kernel = mo::qp_ip_generic_step_kernel<THREADS>;
    MO_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    MO_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, kernel, THREADS, smem));
    const long long cap = (long long)dev.sms * (bps > 0 ? bps : 1);
    const int grid = (int)(a.batch < cap ? a.batch : cap);
    kernel<<<grid, THREADS, smem, stream>>>(a, *step);
  } else {
    auto kernel = mo::qp_ip_generic_kernel<THREADS>;
    MO_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    MO_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, kernel, THREADS, smem));
    const long long cap = (long long)dev.sms * (bps > 0 ? bps : 1);
    const int grid = (int)(a.batch < cap ? a.batch : cap);
    kernel<<<grid, THREADS, smem, stream>>>(a);
  }
  g_launches.fetch_add(1);
  MO_CUDA(cudaGetLastError());
  return MO_OK;
}

int dispatch_generic(const mo::QPArgs& a, const mo::StepArgs* step, const DeviceInfo& dev, cudaStream_t stream) {
  switch (generic_threads(a.n, a.k)) {
    case 32: return launch_generic<32>(a, step, dev, stream);
    case 128: return launch_generic<128>(a, step, dev, stream);
    default: return launch_generic<256>(a, step, dev, stream);
  }
}

// warp16 (qp_w16.cuh): every small shape without a thread-per-problem instantiation.
bool w16_eligible(int n, int k, int m) { return mo::wip_shape_ok(n, k, m); }

template <int PMAX>
int launch_w16(const mo::QPArgs& a, const DeviceInfo& dev, cudaStream_t stream) {
  const bool box = (a.flags & MO_QP_FLAG_BOX_PAIRS) != 0;
  const size_t smem = mo::w16_smem_bytes(a.n, a.k, a.m, box);
  if (smem > (size_t)dev.max_smem_optin) {
    g_last_error = "problem too large for shared memory: needs " + std::to_string(smem) + " bytes";
    return MO_ERR_UNSUPPORTED_SHAPE;
  }
  auto kernel = mo::qp_warp16_kernel<PMAX>;
  int bps = 0;
  MO_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  MO_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, kernel, 32, smem));
  const long long cap = (long long)dev.sms * (bps > 0 ? bps : 1);
  static std::atomic<unsigned> ring{0};
  unsigned long long* tickets = nullptr;
  MO_CUDA(cudaGetSymbolAddress(reinterpret_cast<void**>(&tickets), mo::g_w16_tickets));
  tickets += ring.fetch_add(1) % mo::kW16TicketSlots;
  MO_CUDA(cudaMemsetAsync(tickets, 0, sizeof(unsigned long long), stream));
  kernel<<<(int)(a.batch < cap ? a.batch : cap), 32, smem, stream>>>(a, tickets, g_w16_tma.load() ? 1 : 0);
  g_launches.fetch_add(1);
  MO_CUDA(cudaGetLastError());
  return MO_OK;
}

int dispatch_w16(const mo::QPArgs& a, const DeviceInfo& dev, cudaStream_t stream) {
  const int p = a.n + a.k;
  if (p <= 4) return launch_w16<4>(a, dev, stream);
  if (p <= 8) return launch_w16<8>(a, dev, stream);
  if (p <= 12) return launch_w16<12>(a, dev, stream);
  return launch_w16<16>(a, dev, stream);
}

int dispatch_solve(const mo::QPArgs& a, const DeviceInfo& dev, cudaStream_t stream) {
  auto aligned16 = [](const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; };
  const bool vec_ok = aligned16(a.G) && aligned16(a.c) && aligned16(a.A) && aligned16(a.b) && aligned16(a.ib) &&
                      aligned16(a.vars);  // the tpp kernels use 16-byte vector requests on even-length runs
  if (vec_ok && tpp_eligible(a.n, a.k, a.m, a.flags, a.icount != nullptr, a.strategy)) {
#define X(N_, K_) if (a.n == N_ && a.k == K_) return launch_tpp<N_, K_>(a, dev, stream);
    MO_TPP_SHAPES(X)
#undef X
  }
  if (vec_ok && dm_eligible(a.n, a.k, a.m, a.flags, a.icount != nullptr, a.strategy) && !dm_use_cta_kernel()) {
    if (mo::wd_has_shape(a.n, a.k)) {
      MO_CUDA(mo::wd_launch(a, dev.sms, stream));
      g_launches.fetch_add(1);
      return MO_OK;
    }
  }
  if (vec_ok && dm_eligible(a.n, a.k, a.m, a.flags, a.icount != nullptr, a.strategy)) {
#define X(N_, K_, W_, B_) if (a.n == N_ && a.k == K_) return launch_dm<N_, K_, W_, B_>(a, dev, stream);
    MO_DM_SHAPES(X)
#undef X
  }
  if (w16_eligible(a.n, a.k, a.m)) return dispatch_w16(a, dev, stream);
  return dispatch_generic(a, nullptr, dev, stream);
}

}  // namespace

// ---------------------------------------------------------------------------------------------
// exported C ABI
// ---------------------------------------------------------------------------------------------
extern "C" {

void mo_qp_params_default(mo_qp_params* p) {  // qp.hpp:136-163
  if (!p) return;
  p->initial_mu = 1.0;
  p->sigma = 0.5;
  p->termination_kkt_tol = 1.0e-9;
  p->termination_complementarity_tol = 1.0e-6;
  p->max_iterations = 10;
  p->barrier_strategy = MO_BARRIER_COMPLEMENTARITY;
  p->decrease_mu_only_on_small_error = 0;
  p->initial_guess_method = MO_GUESS_NAIVE;
  p->initialize_mu_with_complementarity = 0;
  p->reserved = 0;
}

int64_t mo_qp_vars_stride(int32_t n, int32_t k, int32_t m) { return (int64_t)n + 2 * (int64_t)m + k; }

int mo_abi_version(void) { return MO_ABI_VERSION; }

int mo_debug_set_tma(int enabled) { return g_w16_tma.exchange(enabled ? 1 : 0); }
int mo_debug_set_tpp_async(int enabled) { return g_tpp_async.exchange(enabled ? 1 : 0); }

int mo_debug_set_dm_kernel(int which) { return g_dm_kernel.exchange(which == 1 ? 1 : 0); }

int64_t mo_kernel_launch_count(void) { return g_launches.load(); }

const char* mo_last_error_string(void) { return g_last_error.c_str(); }

const char* mo_qp_kernel_variant(int32_t n, int32_t k, int32_t m, int32_t flags) {
  if (tpp_eligible(n, k, m, flags, false, MO_BARRIER_COMPLEMENTARITY)) return "tpp";
  if (dm_eligible(n, k, m, flags, false, MO_BARRIER_COMPLEMENTARITY))
    return (!dm_use_cta_kernel() && mo::wd_has_shape(n, k)) ? "dmma_warp" : "dmma_cta";
  if (w16_eligible(n, k, m)) return "warp16";
  switch (generic_threads(n, k)) {
    case 32: return "generic32";
    case 128: return "generic128";
    default: return "generic256";
  }
}

int mo_qp_ip_solve_batch(const mo_qp_batch* problems, const mo_qp_params* params, const mo_qp_out* out, int device,
                         void* cuda_stream) {
  MO_TRACE();
  if (!problems || !params || !out) return MO_ERR_INVALID_ARGUMENT;
  int rc = check_batch(problems);
  if (rc) return rc;
  rc = check_params(params);
  if (rc) return rc;
  if (problems->batch == 0) return MO_OK;
  if (!out->vars || !out->status || !out->iters) return MO_ERR_INVALID_ARGUMENT;
  if (problems->flags & kHostOnlyFlags) {
    g_last_error = "MO_QP_FLAG_PACKED_G / MO_QP_FLAG_UNIFORM_BOX are input formats of the *_host entry points";
    return MO_ERR_INVALID_ARGUMENT;
  }
  DeviceInfo dev;
  rc = device_info(device, &dev);
  if (rc) return rc;
  DeviceGuard guard(device);
  if (!guard.ok) { g_last_error = "cudaSetDevice failed"; return MO_ERR_CUDA; }
  const mo::QPArgs a = make_args(problems, params, out);
  return dispatch_solve(a, dev, static_cast<cudaStream_t>(cuda_stream));
}

int mo_qp_ip_newton_step_batch(const mo_qp_batch* problems, double mu, int32_t barrier_strategy, double* vars,
                               double* delta_out, double* ip_out, double* kkt_out, int32_t* status, int device,
                               void* cuda_stream) {
  MO_TRACE();
  if (!problems || !vars || !status) return MO_ERR_INVALID_ARGUMENT;
  int rc = check_batch(problems);
  if (rc) return rc;
  if (barrier_strategy < 0 || barrier_strategy > 2) return MO_ERR_INVALID_PARAMS;
  if (problems->batch == 0) return MO_OK;
  DeviceInfo dev;
  rc = device_info(device, &dev);
  if (rc) return rc;
  DeviceGuard guard(device);
  if (!guard.ok) { g_last_error = "cudaSetDevice failed"; return MO_ERR_CUDA; }
  mo::QPArgs a = make_args(problems, nullptr, nullptr);
  a.strategy = barrier_strategy;
  a.vars = vars;
  a.status = status;
  mo::StepArgs sa{mu, delta_out, ip_out, kkt_out};
  return dispatch_generic(a, &sa, dev, static_cast<cudaStream_t>(cuda_stream));
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------
// host-buffer entry point: chunked H2D -> solve -> D2H pipeline over three stream slots
// ---------------------------------------------------------------------------------------------
namespace {

struct Slot {
  cudaStream_t stream = nullptr;
  char* dbuf = nullptr;
  size_t cap = 0;
};
struct HostCtx {
  Slot slots[3];
  std::mutex mu;  // one host pipeline per device at a time; different devices run concurrently (_host_multi)
};
HostCtx g_host[64];

inline size_t align256(size_t v) { return (v + 255) & ~size_t(255); }

struct ChunkLayout {
  size_t G, c, A, b, ivar, ia, ib, icount, vars, status, iters, kkt, lag, log, Gpacked, xonly, total;
};

ChunkLayout chunk_layout(const mo_qp_batch& B, const mo_qp_params& P, const mo_qp_out& O, long long C) {
  ChunkLayout L{};
  size_t off = 0;
  auto take = [&](size_t bytes) { size_t o = off; off = align256(off + bytes); return o; };
  const size_t n = B.n, k = B.k, m = B.m, d = sizeof(double);
  L.G = take(C * n * n * d);
  L.c = take(C * n * d);
  L.A = take(C * k * n * d);
  L.b = take(C * k * d);
  L.ivar = take(B.ineq_var ? C * m * sizeof(int32_t) : 0);
  L.ia = take(B.ineq_a ? C * m * d : 0);
  L.ib = take(C * m * d);
  L.icount = take(B.ineq_count ? C * sizeof(int32_t) : 0);
  L.vars = take(C * (n + 2 * m + k) * d);
  L.status = take(C * sizeof(int32_t));
  L.iters = take(C * sizeof(int32_t));
  L.kkt = take(O.kkt_final ? C * 4 * d : 0);
  L.lag = take(O.lagrange ? C * 2 * d : 0);
  L.log = take(O.iter_log ? C * (size_t)P.max_iterations * MO_QP_ITER_LOG_STRIDE * d : 0);
  L.Gpacked = take((B.flags & MO_QP_FLAG_PACKED_G) ? C * (n * (n + 1) / 2) * d : 0);
  L.xonly = take(O.x ? C * n * d : 0);
  L.total = off;
  return L;
}

// Device-side expansion of the compact host formats (MO_QP_FLAG_PACKED_G / MO_QP_FLAG_UNIFORM_BOX) into the layout the
// solve kernels read, and compaction of x_block() for the x-only output: HBM-speed passes that replace PCIe bytes.
__global__ void __launch_bounds__(256) unpack_g_kernel(const double* __restrict__ packed, double* __restrict__ G, int n,
                                                       long long count) {
  const long long per = (long long)n * (n + 1) / 2, total = count * per;
  for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (long long)gridDim.x * blockDim.x) {
    const long long p = idx / per;
    int e = (int)(idx - p * per), j = 0;
    while (e >= n - j) { e -= n - j; ++j; }  // column j holds rows j..n-1
    G[p * n * n + (long long)j * n + (j + e)] = packed[idx];
  }
}
__global__ void __launch_bounds__(256) broadcast_box_kernel(const double* __restrict__ box, double* __restrict__ ib, int m,
                                                            long long count) {
  const long long total = count * m;
  for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (long long)gridDim.x * blockDim.x)
    ib[idx] = box[idx % m];
}
__global__ void __launch_bounds__(256) gather_x_kernel(const double* __restrict__ vars, double* __restrict__ x, int n,
                                                       long long stride, long long count) {
  const long long total = count * n;
  for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (long long)gridDim.x * blockDim.x) {
    const long long p = idx / n;
    x[idx] = vars[p * stride + (idx - p * n)];
  }
}
inline int grid_for(long long total, int sms) {
  const long long want = (total + 255) / 256, cap = (long long)sms * 8;
  return (int)(want < cap ? (want > 0 ? want : 1) : cap);
}

int solve_host_range(const mo_qp_batch& B, const mo_qp_params& P, const mo_qp_out& O, long long begin, long long end,
                     int device) {
  DeviceInfo dev;
  int rc = device_info(device, &dev);
  if (rc) return rc;
  DeviceGuard guard(device);
  if (!guard.ok) { g_last_error = "cudaSetDevice failed"; return MO_ERR_CUDA; }
  HostCtx& ctx = g_host[device];
  std::lock_guard<std::mutex> lk(ctx.mu);

  const size_t n = B.n, k = B.k, m = B.m, d = sizeof(double), vs = n + 2 * m + k;
  const bool packed = (B.flags & MO_QP_FLAG_PACKED_G) != 0, ubox = (B.flags & MO_QP_FLAG_UNIFORM_BOX) != 0;
  const size_t gpack = n * (n + 1) / 2;
  const size_t per_problem = (n * n + n + k * n + k + 2 * m + vs + 8) * d +
                             (O.iter_log ? (size_t)P.max_iterations * MO_QP_ITER_LOG_STRIDE * d : 0);
  const long long total = end - begin;
  long long C = (long long)((size_t(96) << 20) / per_problem);
  if (C < 4096) C = 4096;
  if (C > total) C = total;
  const ChunkLayout L = chunk_layout(B, P, O, C);
  for (Slot& s : ctx.slots) {
    if (!s.stream) MO_CUDA(cudaStreamCreateWithFlags(&s.stream, cudaStreamNonBlocking));
    if (s.cap < L.total) {
      if (s.dbuf) MO_CUDA(cudaFree(s.dbuf));
      s.dbuf = nullptr; s.cap = 0;
      MO_CUDA(cudaMalloc(&s.dbuf, L.total));
      s.cap = L.total;
    }
  }
  const bool user_guess = P.initial_guess_method == MO_GUESS_USER_PROVIDED;
  int slot_i = 0;
  for (long long lo = begin; lo < end; lo += C, slot_i = (slot_i + 1) % 3) {
    const long long cnt = (end - lo < C) ? end - lo : C;
    Slot& s = ctx.slots[slot_i];
    MO_CUDA(cudaStreamSynchronize(s.stream));  // previous use of this slot has drained
    char* D = s.dbuf;
    auto h2d = [&](size_t off, const void* src, size_t stride_bytes) -> cudaError_t {
      if (!src || !stride_bytes) return cudaSuccess;
      return cudaMemcpyAsync(D + off, (const char*)src + lo * stride_bytes, cnt * stride_bytes,
                             cudaMemcpyHostToDevice, s.stream);
    };
    if (packed) {
      MO_CUDA(h2d(L.Gpacked, B.G, gpack * d));
      unpack_g_kernel<<<grid_for(cnt * (long long)gpack, dev.sms), 256, 0, s.stream>>>(
          (const double*)(D + L.Gpacked), (double*)(D + L.G), (int)n, cnt);
      g_launches.fetch_add(1);
    } else {
      MO_CUDA(h2d(L.G, B.G, n * n * d));
    }
    MO_CUDA(h2d(L.c, B.c, n * d));
    MO_CUDA(h2d(L.A, B.A_eq, k * n * d));
    MO_CUDA(h2d(L.b, B.b_eq, k * d));
    MO_CUDA(h2d(L.ivar, B.ineq_var, m * sizeof(int32_t)));
    MO_CUDA(h2d(L.ia, B.ineq_a, m * d));
    if (ubox) {  // one [2n] vector for the batch: staged in the (still unused) vars area, broadcast on the device
      MO_CUDA(cudaMemcpyAsync(D + L.vars, B.ineq_b, m * d, cudaMemcpyHostToDevice, s.stream));
      broadcast_box_kernel<<<grid_for(cnt * (long long)m, dev.sms), 256, 0, s.stream>>>(
          (const double*)(D + L.vars), (double*)(D + L.ib), (int)m, cnt);
      g_launches.fetch_add(1);
    } else {
      MO_CUDA(h2d(L.ib, B.ineq_b, m * d));
    }
    MO_CUDA(h2d(L.icount, B.ineq_count, sizeof(int32_t)));
    if (user_guess) MO_CUDA(h2d(L.vars, O.vars, vs * d));
    mo_qp_batch db = B;
    db.flags = B.flags & ~kHostOnlyFlags;
    db.batch = cnt;
    db.G = (const double*)(D + L.G); db.c = (const double*)(D + L.c);
    db.A_eq = k ? (const double*)(D + L.A) : nullptr; db.b_eq = k ? (const double*)(D + L.b) : nullptr;
    db.ineq_var = B.ineq_var ? (const int32_t*)(D + L.ivar) : nullptr;
    db.ineq_a = B.ineq_a ? (const double*)(D + L.ia) : nullptr;
    db.ineq_b = m ? (const double*)(D + L.ib) : nullptr;
    db.ineq_count = B.ineq_count ? (const int32_t*)(D + L.icount) : nullptr;
    mo_qp_out dout{};
    dout.vars = (double*)(D + L.vars); dout.status = (int32_t*)(D + L.status); dout.iters = (int32_t*)(D + L.iters);
    dout.kkt_final = O.kkt_final ? (double*)(D + L.kkt) : nullptr;
    dout.lagrange = O.lagrange ? (double*)(D + L.lag) : nullptr;
    dout.iter_log = O.iter_log ? (double*)(D + L.log) : nullptr;
    if (dout.iter_log)
      MO_CUDA(cudaMemsetAsync(dout.iter_log, 0xff, cnt * (size_t)P.max_iterations * MO_QP_ITER_LOG_STRIDE * d, s.stream));
    const mo::QPArgs a = make_args(&db, &P, &dout);
    rc = dispatch_solve(a, dev, s.stream);
    if (rc) return rc;
    auto d2h = [&](void* dst, size_t off, size_t stride_bytes) -> cudaError_t {
      if (!dst || !stride_bytes) return cudaSuccess;
      return cudaMemcpyAsync((char*)dst + lo * stride_bytes, D + off, cnt * stride_bytes, cudaMemcpyDeviceToHost,
                             s.stream);
    };
    if (O.x) {
      gather_x_kernel<<<grid_for(cnt * (long long)n, dev.sms), 256, 0, s.stream>>>(
          (const double*)(D + L.vars), (double*)(D + L.xonly), (int)n, (long long)vs, cnt);
      g_launches.fetch_add(1);
      MO_CUDA(d2h(O.x, L.xonly, n * d));
    }
    MO_CUDA(d2h(O.vars, L.vars, vs * d));
    MO_CUDA(d2h(O.status, L.status, sizeof(int32_t)));
    MO_CUDA(d2h(O.iters, L.iters, sizeof(int32_t)));
    MO_CUDA(d2h(O.kkt_final, L.kkt, 4 * d));
    MO_CUDA(d2h(O.lagrange, L.lag, 2 * d));
    MO_CUDA(d2h(O.iter_log, L.log, (size_t)P.max_iterations * MO_QP_ITER_LOG_STRIDE * d));
    MO_CUDA(cudaGetLastError());
  }
  for (Slot& s : ctx.slots) MO_CUDA(cudaStreamSynchronize(s.stream));
  return MO_OK;
}

int check_host_call(const mo_qp_batch* problems, const mo_qp_params* params, const mo_qp_out* out) {
  if (!problems || !params || !out) return MO_ERR_INVALID_ARGUMENT;
  int rc = check_batch(problems);
  if (rc) return rc;
  rc = check_params(params);
  if (rc) return rc;
  if (problems->batch == 0) return MO_OK;
  if ((!out->vars && !out->x) || !out->status || !out->iters) return MO_ERR_INVALID_ARGUMENT;
  if (params->initial_guess_method == MO_GUESS_USER_PROVIDED && !out->vars) return MO_ERR_INVALID_ARGUMENT;
  return MO_OK;
}

}  // namespace

extern "C" {

void* mo_host_alloc(size_t bytes) {
  void* p = nullptr;
  if (cudaHostAlloc(&p, bytes ? bytes : 1, cudaHostAllocPortable) != cudaSuccess) {
    g_last_error = std::string("cudaHostAlloc: ") + cudaGetErrorString(cudaGetLastError());
    return nullptr;
  }
  return p;
}

void mo_host_free(void* p) {
  if (p) cudaFreeHost(p);
}

// Pinned memory whose pages are spread round-robin over all NUMA nodes of the host (set_mempolicy(MPOL_INTERLEAVE) around
// the allocation, raw syscall: no libnuma in the image). For batches that feed SEVERAL GPUs from one process: a plain
// allocation lands on the node of the allocating thread and every GPU's DMA then drains that one node's memory controllers
// (measured: 114 GB/s aggregate over 4 GPUs against 190 GB/s for four processes with their own buffers). Falls back to
// mo_host_alloc when the policy cannot be set (single node, restricted container).
void* mo_host_alloc_interleaved(size_t bytes) {
#if defined(__linux__) && defined(SYS_set_mempolicy)
  unsigned long mask = 0;
  if (FILE* f = std::fopen("/sys/devices/system/node/online", "r")) {  // e.g. "0-1" or "0"
    int lo = 0, hi = -1;
    const int got = std::fscanf(f, "%d-%d", &lo, &hi);
    std::fclose(f);
    if (got == 1) hi = lo;
    for (int i = lo; i <= hi && i < 64 && got >= 1; ++i) mask |= 1UL << i;
  }
  const bool multi = (mask & (mask - 1)) != 0;
  const bool set = multi && syscall(SYS_set_mempolicy, 3 /* MPOL_INTERLEAVE */, &mask, sizeof(mask) * 8 + 1) == 0;
  void* p = mo_host_alloc(bytes);
  if (p && set) std::memset(p, 0, bytes);  // touch under the policy (cudaHostAlloc normally populates; this makes sure)
  if (set) syscall(SYS_set_mempolicy, 0 /* MPOL_DEFAULT */, nullptr, 0);
  return p;
#else
  return mo_host_alloc(bytes);
#endif
}

int mo_qp_ip_solve_batch_host(const mo_qp_batch* problems, const mo_qp_params* params, const mo_qp_out* out,
                              int device) {
  MO_TRACE();
  const int rc = check_host_call(problems, params, out);
  if (rc || problems->batch == 0) return rc;
  return solve_host_range(*problems, *params, *out, 0, problems->batch, device);
}

int mo_qp_ip_solve_batch_host_multi(const mo_qp_batch* problems, const mo_qp_params* params, const mo_qp_out* out,
                                    const int* devices, int n_devices) {
  MO_TRACE();
  if (!devices || n_devices < 1 || n_devices > 64) return MO_ERR_INVALID_ARGUMENT;
  const int rc = check_host_call(problems, params, out);
  if (rc || problems->batch == 0) return rc;
  if (n_devices == 1) return solve_host_range(*problems, *params, *out, 0, problems->batch, devices[0]);
  // contiguous shards, one host thread (and one 3-stream ring) per device, no exchange between them
  const long long B = problems->batch, shard = (B + n_devices - 1) / n_devices;
  std::vector<int> rcs(n_devices, MO_OK);
  std::vector<std::string> errs(n_devices);
  std::vector<std::thread> pool;
  for (int i = 0; i < n_devices; ++i) {
    const long long lo = i * shard, hi = lo + shard < B ? lo + shard : B;
    if (lo >= hi) break;
    pool.emplace_back([&, i, lo, hi]() {
      rcs[i] = solve_host_range(*problems, *params, *out, lo, hi, devices[i]);
      if (rcs[i]) errs[i] = g_last_error;  // g_last_error is thread local
    });
  }
  for (std::thread& t : pool) t.join();
  for (int i = 0; i < n_devices; ++i)
    if (rcs[i]) { g_last_error = errs[i]; return rcs[i]; }
  return MO_OK;
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------
// nonlinear least squares entry points
// ---------------------------------------------------------------------------------------------
#include "nls_launch.h"
#include "qp_nullspace.cuh"

namespace {

// CheckParams, nonlinear.cc:49-73
int check_nls_params(const mo_nls_params* p) {
  const bool ok =
      p->max_iterations >= 0 && p->max_qp_iterations >= 1 && p->termination_kkt_tolerance > 0 &&
      p->absolute_exit_tol > 0 && p->max_line_search_iterations >= 0 && p->relative_exit_tol >= 0 &&
      p->relative_exit_tol <= 1 && p->absolute_first_derivative_tol >= 0 && p->armijo_search_tau > 0 &&
      p->armijo_search_tau < 1 && p->equality_penalty_initial >= 0 && p->equality_penalty_scale_factor >= 1.0 &&
      p->equality_penalty_rho >= 0 && p->equality_penalty_rho < 1 && p->max_lambda >= 0 &&
      p->min_lambda <= p->max_lambda && p->lambda_initial >= p->min_lambda && p->lambda_initial <= p->max_lambda &&
      p->lambda_failure_init >= 0 && p->lambda_decrease_on_success >= 0 && p->lambda_decrease_on_success < 1.0 &&
      p->lambda_decrease_on_restore >= 0 && p->lambda_decrease_on_restore < 1.0 &&
      (p->line_search_strategy == MO_LS_ARMIJO_BACKTRACK || p->line_search_strategy == MO_LS_POLYNOMIAL_APPROXIMATION) &&
      (p->retraction == MO_RETRACT_ADD || p->retraction == MO_RETRACT_MOD_PI) && p->max_line_search_iterations < 40 &&
      (p->inner_qp_solver == MO_NLS_QP_AUTO || p->inner_qp_solver == MO_NLS_QP_GENERIC);
  return ok ? MO_OK : MO_ERR_INVALID_PARAMS;
}

int check_chain(const mo_chain_ik_batch* b, bool fk_only) {
  if (!b || b->links <= 0 || b->links > mo::kMaxChainLinks || b->n <= 0 || b->n > 6 * b->links || b->batch < 0)
    return MO_ERR_INVALID_ARGUMENT;
  if (!b->link_pose || !b->link_mask) return MO_ERR_INVALID_ARGUMENT;
  if (fk_only) return MO_OK;
  if (b->n_cost < 0 || b->n_cost > mo::kMaxChainRows || b->n_eq < 0 || b->n_eq > mo::kMaxChainRows || b->m < 0)
    return MO_ERR_INVALID_ARGUMENT;
  if (b->cost_block < 0 || (b->cost_block > 0 && b->n_cost % b->cost_block != 0)) return MO_ERR_INVALID_ARGUMENT;
  if (b->batch == 0) return MO_OK;
  if ((b->n_cost > 0 && (!b->cost_w || !b->cost_target)) || (b->n_eq > 0 && (!b->eq_w || !b->eq_target)) ||
      (b->m > 0 && (!b->ineq_var || !b->ineq_a || !b->ineq_b)) || !b->x0)
    return MO_ERR_INVALID_ARGUMENT;
  return MO_OK;
}

template <typename K>
int warp_grid(K kernel, size_t smem, long long batch, const DeviceInfo& dev, int* grid) {
  if (smem > (size_t)dev.max_smem_optin) {
    g_last_error = "problem too large for shared memory: needs " + std::to_string(smem) + " bytes";
    return MO_ERR_UNSUPPORTED_SHAPE;
  }
  int bps = 0;
  MO_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  MO_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, kernel, 32, smem));
  const long long cap = (long long)dev.sms * (bps > 0 ? bps : 1);
  *grid = (int)(batch < cap ? batch : cap);
  return MO_OK;
}

}  // namespace

extern "C" {

void mo_nls_params_default(mo_nls_params* p) {  // nonlinear.hpp:37-124
  if (!p) return;
  p->max_iterations = 10; p->max_qp_iterations = 10; p->termination_kkt_tolerance = 1.0e-6;
  p->absolute_exit_tol = 1.0e-12; p->relative_exit_tol = 1.0e-5; p->absolute_first_derivative_tol = 1.0e-6;
  p->max_line_search_iterations = 2; p->line_search_strategy = MO_LS_POLYNOMIAL_APPROXIMATION;
  p->armijo_search_tau = 0.8; p->equality_penalty_initial = 1.0; p->equality_penalty_scale_factor = 1.01;
  p->equality_penalty_rho = 0.1; p->lambda_initial = 0.0; p->lambda_failure_init = 1.0e-2;
  p->lambda_decrease_on_success = 0.1; p->lambda_decrease_on_restore = 0.8; p->max_lambda = 1.0; p->min_lambda = 0.0;
  p->retraction = MO_RETRACT_ADD; p->inner_qp_solver = MO_NLS_QP_AUTO;
}

int mo_qp_nullspace_solve_batch(const mo_qp_batch* problems, double* x, int32_t* status, int device, void* cuda_stream) {
  MO_TRACE();
  if (!problems || problems->n <= 0 || problems->k <= 0 || problems->batch < 0) return MO_ERR_INVALID_ARGUMENT;
  if (problems->batch == 0) return MO_OK;
  if (!problems->G || !problems->c || !problems->A_eq || !problems->b_eq || !x || !status) return MO_ERR_INVALID_ARGUMENT;
  if (problems->flags & kHostOnlyFlags) return MO_ERR_INVALID_ARGUMENT;
  DeviceInfo dev;
  int rc = device_info(device, &dev);
  if (rc) return rc;
  DeviceGuard guard(device);
  if (!guard.ok) { g_last_error = "cudaSetDevice failed"; return MO_ERR_CUDA; }
  const size_t smem = sizeof(double) * mo::ns_smem_doubles(problems->n, problems->k);
  int grid = 0;
  rc = warp_grid(mo::qp_nullspace_kernel, smem, problems->batch, dev, &grid);
  if (rc) return rc;
  const mo::NullSpaceArgs a{problems->n, problems->k, problems->batch, problems->G, problems->c, problems->A_eq,
                            problems->b_eq, x, status};
  mo::qp_nullspace_kernel<<<grid, 32, smem, static_cast<cudaStream_t>(cuda_stream)>>>(a);
  g_launches.fetch_add(1);
  MO_CUDA(cudaGetLastError());
  return MO_OK;
}

int mo_nls_linearize_batch(const mo_nls_linearize_args* g, int device, void* cuda_stream) {
  MO_TRACE();
  if (!g || g->n <= 0 || g->r_cost < 0 || g->r_eq < 0 || g->m < 0 || g->batch < 0) return MO_ERR_INVALID_ARGUMENT;
  if (g->cost_block < 0 || (g->cost_block > 0 && g->r_cost % g->cost_block != 0)) return MO_ERR_INVALID_ARGUMENT;
  if (g->batch == 0) return MO_OK;
  if (!g->x || !g->lambda || !g->G || !g->c || !g->errors || (g->r_cost > 0 && (!g->J_cost || !g->r_cost_v)) ||
      (g->r_eq > 0 && (!g->J_eq || !g->r_eq_v || !g->A_eq || !g->b_eq)) ||
      (g->m > 0 && (!g->ineq_var || !g->ineq_a || !g->ineq_b || !g->ineq_b_shifted)))
    return MO_ERR_INVALID_ARGUMENT;
  DeviceInfo dev;
  int rc = device_info(device, &dev);
  if (rc) return rc;
  DeviceGuard guard(device);
  if (!guard.ok) { g_last_error = "cudaSetDevice failed"; return MO_ERR_CUDA; }
  const mo::NlsLaunchInfo info{dev.sms, (size_t)dev.max_smem_optin};
  rc = mo::nls_launch_linearize(g, info, static_cast<cudaStream_t>(cuda_stream), &g_last_error);
  if (rc == MO_OK) g_launches.fetch_add(1);
  return rc;
}

int mo_chain_fk_batch(const mo_chain_ik_batch* chain, const double* params, double* translation, double* rotation_q,
                      double* t_D_params, double* r_D_params, int device, void* cuda_stream) {
  MO_TRACE();
  int rc = check_chain(chain, true);
  if (rc) return rc;
  if (chain->batch == 0) return MO_OK;
  if (!params) return MO_ERR_INVALID_ARGUMENT;
  DeviceInfo dev;
  rc = device_info(device, &dev);
  if (rc) return rc;
  DeviceGuard guard(device);
  if (!guard.ok) { g_last_error = "cudaSetDevice failed"; return MO_ERR_CUDA; }
  const mo::NlsLaunchInfo info{dev.sms, (size_t)dev.max_smem_optin};
  rc = mo::nls_launch_chain_fk(chain, params, translation, rotation_q, t_D_params, r_D_params, info,
                               static_cast<cudaStream_t>(cuda_stream), &g_last_error);
  if (rc == MO_OK) g_launches.fetch_add(1);
  return rc;
}

int mo_nls_solve_chain_ik_batch(const mo_chain_ik_batch* problems, const mo_nls_params* params, const mo_nls_out* out,
                                int device, void* cuda_stream) {
  MO_TRACE();
  if (!params || !out) return MO_ERR_INVALID_ARGUMENT;
  int rc = check_chain(problems, false);
  if (rc) return rc;
  rc = check_nls_params(params);
  if (rc) return rc;
  if (problems->batch == 0) return MO_OK;
  if (!out->x || !out->termination || !out->iterations) return MO_ERR_INVALID_ARGUMENT;
  DeviceInfo dev;
  rc = device_info(device, &dev);
  if (rc) return rc;
  DeviceGuard guard(device);
  if (!guard.ok) { g_last_error = "cudaSetDevice failed"; return MO_ERR_CUDA; }
  const mo::NlsLaunchInfo info{dev.sms, (size_t)dev.max_smem_optin};
  rc = mo::nls_launch_chain_ik(problems, params, out, info, static_cast<cudaStream_t>(cuda_stream), &g_last_error);
  if (rc == MO_OK) g_launches.fetch_add(1);
  return rc;
}

int mo_nls_family_dims(int32_t family, int32_t* n, int32_t* n_cost, int32_t* n_eq) {
  int vn = 0, vc = 0, vb = 0, ve = 0;
  const int rc = mo::nls_family_dims(family, &vn, &vc, &vb, &ve);
  if (rc) { g_last_error = "unknown residual family"; return rc; }
  if (n) *n = vn;
  if (n_cost) *n_cost = vc;
  if (n_eq) *n_eq = ve;
  return MO_OK;
}

int mo_nls_solve_family_batch(const mo_nls_family_batch* problems, const mo_nls_params* params, const mo_nls_out* out,
                              int device, void* cuda_stream) {
  MO_TRACE();
  if (!problems || !params || !out || problems->batch < 0 || problems->m < 0) return MO_ERR_INVALID_ARGUMENT;
  int n = 0;
  int rc = mo_nls_family_dims(problems->family, &n, nullptr, nullptr);
  if (rc) return rc;
  if (problems->m > 0 && (!problems->ineq_var || !problems->ineq_a || !problems->ineq_b)) return MO_ERR_INVALID_ARGUMENT;
  if (problems->data && problems->data_stride <= 0) return MO_ERR_INVALID_ARGUMENT;
  rc = check_nls_params(params);
  if (rc) return rc;
  if (problems->batch == 0) return MO_OK;
  if (!problems->x0 || !out->x || !out->termination || !out->iterations) return MO_ERR_INVALID_ARGUMENT;
  DeviceInfo dev;
  rc = device_info(device, &dev);
  if (rc) return rc;
  DeviceGuard guard(device);
  if (!guard.ok) { g_last_error = "cudaSetDevice failed"; return MO_ERR_CUDA; }
  const mo::NlsLaunchInfo info{dev.sms, (size_t)dev.max_smem_optin};
  rc = mo::nls_launch_family(problems, params, out, info, static_cast<cudaStream_t>(cuda_stream), &g_last_error);
  if (rc == MO_OK) g_launches.fetch_add(1);
  return rc;
}

}  // extern "C"

#ifdef MO_DM_TRACE
extern "C" int mo_debug_read_trace(long long* out, int cap) {
  int n = 0;
  cudaMemcpyFromSymbol(&n, mo::g_dm_trace_n, sizeof(int));
  if (n > cap) n = cap;
  cudaMemcpyFromSymbol(out, mo::g_dm_trace, sizeof(long long) * n);
  int zero = 0;
  cudaMemcpyToSymbol(mo::g_dm_trace_n, &zero, sizeof(int));
  return n;
}
#endif

// Host-pointer variant of mo_qp_nullspace_solve_batch (the call the C++ mirror's QPNullSpaceSolver makes): H2D, solve, D2H,
// synchronous. Plain cudaMalloc / cudaFree per call: equality-only QPs are the small side path of the NLS driver.
extern "C" int mo_qp_nullspace_solve_batch_host(const mo_qp_batch* problems, double* x, int32_t* status, int device) {
  MO_TRACE();
  if (!problems || problems->n <= 0 || problems->k <= 0 || problems->batch < 0) return MO_ERR_INVALID_ARGUMENT;
  if (problems->batch == 0) return MO_OK;
  if (!problems->G || !problems->c || !problems->A_eq || !problems->b_eq || !x || !status) return MO_ERR_INVALID_ARGUMENT;
  if (problems->flags & kHostOnlyFlags) return MO_ERR_INVALID_ARGUMENT;
  DeviceInfo dev;
  int rc = device_info(device, &dev);
  if (rc) return rc;
  DeviceGuard guard(device);
  if (!guard.ok) { g_last_error = "cudaSetDevice failed"; return MO_ERR_CUDA; }
  const size_t n = problems->n, k = problems->k, B = (size_t)problems->batch, d = sizeof(double);
  const size_t oG = 0, oc = align256(oG + B * n * n * d), oA = align256(oc + B * n * d), ob = align256(oA + B * n * k * d),
               ox = align256(ob + B * k * d), os = align256(ox + B * n * d), total = align256(os + B * sizeof(int32_t));
  char* D = nullptr;
  MO_CUDA(cudaMalloc(&D, total));
  auto fail = [&](cudaError_t e, const char* what) { cudaFree(D); return cuda_fail(e, what); };
  cudaError_t e;
  if ((e = cudaMemcpy(D + oG, problems->G, B * n * n * d, cudaMemcpyHostToDevice)) != cudaSuccess) return fail(e, "H2D G");
  if ((e = cudaMemcpy(D + oc, problems->c, B * n * d, cudaMemcpyHostToDevice)) != cudaSuccess) return fail(e, "H2D c");
  if ((e = cudaMemcpy(D + oA, problems->A_eq, B * n * k * d, cudaMemcpyHostToDevice)) != cudaSuccess) return fail(e, "H2D A_eq");
  if ((e = cudaMemcpy(D + ob, problems->b_eq, B * k * d, cudaMemcpyHostToDevice)) != cudaSuccess) return fail(e, "H2D b_eq");
  mo_qp_batch db = *problems;
  db.G = (const double*)(D + oG); db.c = (const double*)(D + oc); db.A_eq = (const double*)(D + oA); db.b_eq = (const double*)(D + ob);
  rc = mo_qp_nullspace_solve_batch(&db, (double*)(D + ox), (int32_t*)(D + os), device, nullptr);
  if (rc) { cudaFree(D); return rc; }
  if ((e = cudaMemcpy(x, D + ox, B * n * d, cudaMemcpyDeviceToHost)) != cudaSuccess) return fail(e, "D2H x");
  if ((e = cudaMemcpy(status, D + os, B * sizeof(int32_t), cudaMemcpyDeviceToHost)) != cudaSuccess) return fail(e, "D2H status");
  MO_CUDA(cudaFree(D));
  return MO_OK;
}

// ---------------------------------------------------------------------------------------------
// No-allocation entry point (no counterpart in the reference).
// Variant of qp_nullspace_kernel whose per-warp workspace (QR, Q, G Z, Z'G Z, vectors: ns_smem_doubles(n, k) doubles)
// lives in a file-scope device array instead of dynamic shared memory: callers under CUDA graph capture or in loops
// that may not call cudaMalloc get a fixed footprint, and problems whose workspace exceeds the 227 KB of shared memory
// (n > ~100) can be solved at all. Every GPU that loads the library holds the array (32 MiB) until exit.
// ---------------------------------------------------------------------------------------------
namespace mo {
constexpr size_t kNoAllocScratchBytes = size_t(32) << 20;
static __device__ __align__(256) unsigned char g_noalloc_scratch[kNoAllocScratchBytes];

// `scratch`: the DEVICE address of g_noalloc_scratch obtained with cudaGetSymbolAddress (the host-side name of the array
// is only a stand-in); block b uses the slice [b * per_warp, (b + 1) * per_warp) doubles.
__global__ void __launch_bounds__(32) qp_nullspace_noalloc_kernel(const NullSpaceArgs a, double* __restrict__ scratch,
                                                                  const long long per_warp) {
  const int n = a.n, k = a.k;
  double* work = scratch + (long long)blockIdx.x * per_warp;
  for (long long p = blockIdx.x; p < a.batch; p += gridDim.x) {
    __syncwarp();
    const int st = ns_solve_warp(n, k, a.G + p * (long long)n * n, a.c + p * n, a.A + p * (long long)n * k, a.b + p * k,
                                 a.x + p * n, work);
    if (threadIdx.x == 0) a.status[p] = st;
  }
}
}  // namespace mo

namespace {
double* g_noalloc_ptr[64] = {};  // device address of the scratch array per device, resolved once
std::mutex g_noalloc_mu;
}  // namespace

extern "C" int mo_qp_nullspace_solve_batch_noalloc(const mo_qp_batch* problems, double* x, int32_t* status, int device,
                                                   void* cuda_stream) {
  MO_TRACE();
  if (!problems || problems->n <= 0 || problems->k <= 0 || problems->k > problems->n || problems->batch < 0)
    return MO_ERR_INVALID_ARGUMENT;
  if (problems->batch == 0) return MO_OK;
  if (!problems->G || !problems->c || !problems->A_eq || !problems->b_eq || !x || !status) return MO_ERR_INVALID_ARGUMENT;
  if (problems->flags & kHostOnlyFlags) return MO_ERR_INVALID_ARGUMENT;
  if (problems->n > MO_NOALLOC_MAX_N) {
    g_last_error = "mo_qp_nullspace_solve_batch_noalloc: n = " + std::to_string(problems->n) + " exceeds MO_NOALLOC_MAX_N = " +
                   std::to_string(MO_NOALLOC_MAX_N) + " (static scratch of " + std::to_string(mo::kNoAllocScratchBytes >> 20) +
                   " MiB); use mo_qp_nullspace_solve_batch";
    return MO_ERR_UNSUPPORTED_SHAPE;
  }
  DeviceInfo dev;
  int rc = device_info(device, &dev);
  if (rc) return rc;
  DeviceGuard guard(device);  // the symbol address is per device: resolve it with THIS device current
  if (!guard.ok) { g_last_error = "cudaSetDevice failed"; return MO_ERR_CUDA; }
  double* scratch = nullptr;
  {
    std::lock_guard<std::mutex> lk(g_noalloc_mu);
    if (!g_noalloc_ptr[device]) {
      void* p = nullptr;
      MO_CUDA(cudaGetSymbolAddress(&p, mo::g_noalloc_scratch));
      g_noalloc_ptr[device] = static_cast<double*>(p);
    }
    scratch = g_noalloc_ptr[device];
  }
  // per-warp slice rounded up to 256 bytes; as many resident warps as the array holds (at least 1 by the n limit)
  const long long per_warp = (long long)((mo::ns_smem_doubles(problems->n, problems->k) + 31) & ~size_t(31));
  long long grid = (long long)(mo::kNoAllocScratchBytes / sizeof(double)) / per_warp;
  const long long cap = (long long)dev.sms * 16;
  if (grid > cap) grid = cap;
  if (grid > problems->batch) grid = problems->batch;
  if (grid < 1) { g_last_error = "mo_qp_nullspace_solve_batch_noalloc: workspace of one problem exceeds the static scratch"; return MO_ERR_UNSUPPORTED_SHAPE; }
  const mo::NullSpaceArgs a{problems->n, problems->k, problems->batch, problems->G, problems->c, problems->A_eq,
                            problems->b_eq, x, status};
  mo::qp_nullspace_noalloc_kernel<<<(int)grid, 32, 0, static_cast<cudaStream_t>(cuda_stream)>>>(a, scratch, per_warp);
  g_launches.fetch_add(1);
  MO_CUDA(cudaGetLastError());
  return MO_OK;
}
