/*
 * kernels_cta.cuh -- CTA-per-problem kernels for medium shapes (compile-time NX, NU):
 *   k_csqp_direction_stream  SolverCSQP::computeDirection (src/csqp.cpp:454-517): DMMA full
 *                            sweeps (kernels_ric.cuh) + streaming light / forward sweeps
 *                            (kernels_stream.cuh), the whole ADMM loop in one launch
 *   k_sqp_direction_cta      SolverSQP::computeDirection (src/sqp.cpp:224-242): DMMA backward
 *                            sweep + forward pass / KKT
 *   k_fddp_direction_cta     SolverFDDP::computeDirection (src/fddp.cpp:189-204)
 */
#pragma once
#include <cstdio>
#include <cstdlib>
#include "launchers.h"
#include "kernels_ric.cuh"
#include "kernels_stream.cuh"

namespace mimgpu {

/* ------------------------------------------------------------------ the kernel */
/* out-of-line copies of the shape-generic pieces: keeps the register allocation of the
 * streaming sweeps independent of the (much larger) full Riccati sweep */
template <int MODE>
static __device__ __noinline__ int riccati_backward_ool(const Prob& p, double* sm, const Smem& s, const RicPar& rp,
                                                 int stage_fx) {
  return riccati_backward<MODE>(p, as_shared(sm), s, rp, stage_fx);
}
static __device__ __noinline__ void csqp_forward_eq_ool(const Prob& p, double* sm, const Smem& s) { csqp_forward_eq(p, as_shared(sm), s); }
static __device__ __noinline__ void csqp_kkt_ool(const Prob& p, double* sm, const Smem& s) { csqp_kkt(p, as_shared(sm), s); }

struct StreamKPar {
  mimgpu_csqp_params q;
  int dbg;
  int with_kkt, ns, generic_total; /* generic_total: doubles of the union region the non-streaming code needs */
};

/* SolverCSQP::computeDirection (src/csqp.cpp:454-517): same control flow as
 * csqp_direction_body (kernels_generic.cuh) with the light sweep, the gain solve and the
 * forward/update sweep replaced by the streaming versions above. */
/* CTA size cap and resident CTAs per SM (register budget) per shape */
template <int NX, int NU> struct StreamTune {
  static constexpr int MAXT = NX <= 16 ? 128 : 256, MINB = NX <= 16 ? 6 : 2;
};
#ifndef MIMGPU_STREAM_MAX_STAGES
#define MIMGPU_STREAM_MAX_STAGES 3   /* the pipeline has barriers for 4; 4 stages not measured yet */
#endif
#ifndef MIMGPU_TUNE_C3_MINB
#define MIMGPU_TUNE_C3_MINB 2
#endif
#ifndef MIMGPU_TUNE_C1_MINB
#define MIMGPU_TUNE_C1_MINB 8
#endif
#ifndef MIMGPU_TUNE_C1_MAXT
#define MIMGPU_TUNE_C1_MAXT 96
#endif
#ifndef MIMGPU_C1_SPLIT
#define MIMGPU_C1_SPLIT 4   /* lanes per dot product of the 12/6 sweeps */
#endif
template <> struct StreamTune<36, 12> { static constexpr int MAXT = 192, MINB = MIMGPU_TUNE_C3_MINB; };
template <> struct StreamTune<12, 6> { static constexpr int MAXT = MIMGPU_TUNE_C1_MAXT, MINB = MIMGPU_TUNE_C1_MINB; };
template <> struct StreamTune<32, 16> { static constexpr int MAXT = 192, MINB = 2; };
/* DIRECT shapes: one CTA per SM (the Riccati layout of the full sweeps fills the shared memory) */
template <> struct StreamTune<64, 32> { static constexpr int MAXT = 384, MINB = 1; };
template <> struct StreamTune<76, 32> { static constexpr int MAXT = 384, MINB = 1; };
template <> struct StreamTune<48, 24> { static constexpr int MAXT = 256, MINB = 1; };

template <int NX, int NU, int SPLIT, bool DIRECT = false>
__global__ void __launch_bounds__(StreamTune<NX, NU>::MAXT, StreamTune<NX, NU>::MINB)
k_csqp_direction_stream(KArgs a, StreamKPar kp) {
  static_assert(NX % 2 == 0 && NU % 2 == 0, "TMA bulk copies need 16-byte blocks");
  extern __shared__ __align__(16) double sm_stream[];
  if ((int)blockIdx.x >= a.n_active) return;
  const int b = a.active ? a.active[blockIdx.x] : (int)blockIdx.x + a.first;
  Prob p = bind(a, b);
  const StreamSmem ss = stream_layout<NX, NU>(p.ng, p.ngT, kp.ns, kp.generic_total);
  double* sm = sm_stream;
  double* gsm = sm + ss.uni; /* generic Riccati layout lives in the union region */
  const Smem gs = smem_layout(NX, NU, p.ng, p.ngT, a.stage_fx);
  /* DIRECT shapes: the constrained DMMA Riccati layout does not fit the shared memory, their full
   * sweeps run the shape-generic code */
  constexpr bool USE_MMA = (NX % 4 == 0) && !DIRECT;
  const RicSmem rs = ric_layout<NX, NU>(p.ng, p.ngT, true);
  uint64_t* ric_bars = reinterpret_cast<uint64_t*>(sm + ss.bars) + 4;
  uint32_t ric_seq = 0u;
  Pipe<NX, NU> pipe;
  pipe.bars = reinterpret_cast<uint64_t*>(sm + ss.bars);
  pipe.stages = sm + ss.uni;
  pipe.stage = ss.stage; pipe.ns = kp.ns; pipe.ng = p.ng;
  pipe.n_issue = 0; pipe.n_wait = 0; pipe.islot = 0; pipe.wslot = 0; pipe.wphase = 0u;
  pipe.no_prefetch = (kp.dbg >> 3) & 1;
  if (threadIdx.x == 0) {
    for (int i = 0; i < kp.ns; ++i) mbar_init(pipe.bars + i, 1);
    mbar_init(ric_bars, 1);
    mbar_init(ric_bars + 1, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  const mimgpu_csqp_params& q = kp.q;
  const size_t B = p.B;
  const int nc = p.T * p.ng + p.ngT;
  double rho_sparse = p.scal[MIMGPU_S_RHO_SPARSE * B + b];
  RicPar rp;
  rp.preg = a.preg_arr ? a.preg_arr[b] : q.preg;
  rp.dreg = a.dreg_arr ? a.dreg_arr[b] : q.dreg;
  rp.sigma = q.sigma;
  rp.is_feasible = 0;
  rp.dbg = kp.dbg >> 4;
  if (q.reset_rho) { rho_sparse = q.rho_sparse_base; csqp_apply_rho(p, rho_sparse, q.rho_min); }
  for (int e = threadIdx.x; e < nc; e += blockDim.x) { p.z[e] = 0.0; if (q.reset_y) p.y[e] = 0.0; }
  __syncthreads();
  int st = 0, converged = 0, it = 0;
  AdmmNorms n;
  n.primal = p.scal[MIMGPU_S_NORM_PRIMAL * B + b]; n.dual = p.scal[MIMGPU_S_NORM_DUAL * B + b];
  n.primal_rel = p.scal[MIMGPU_S_NORM_PRIMAL_REL * B + b]; n.dual_rel = p.scal[MIMGPU_S_NORM_DUAL_REL * B + b];
  if (q.equality_qp_initial_guess) {
    if constexpr (USE_MMA) st = ric_backward_mma<MODE_FULL, NX, NU>(p, gsm, rs, rp, ric_bars, ric_seq, true);
    else st = riccati_backward_ool<MODE_EQ>(p, gsm, gs, rp, a.stage_fx);
    if (kp.dbg & 1) return;
    if (!st && !(kp.dbg & 2)) { __syncthreads(); csqp_forward_eq_ool(p, gsm, gs); }
    if (kp.dbg & 4) return;
  }
  const int interval = q.rho_update_interval;
  if (!st) {
    for (it = 1; it < q.max_qp_iters + 1; ++it) {
      if (it % interval == 1 || interval == 1) {
        if constexpr (USE_MMA) st = ric_backward_mma<MODE_FULL, NX, NU>(p, gsm, rs, rp, ric_bars, ric_seq);
        else st = riccati_backward_ool<MODE_FULL>(p, gsm, gs, rp, a.stage_fx);
        if (st) break;
        fence_proxy_async(); /* K written through the generic proxy, read back by TMA */
      } else {
        stream_backward_light<NX, NU, SPLIT, DIRECT>(p, sm, ss, pipe, q.sigma);
        stream_solve_gains<NU>(p);
      }
      __syncthreads();
      stream_forward_update<NX, NU, SPLIT, DIRECT>(p, sm, ss, pipe, q, it, n);
      if (it % interval == 0) {
        const double scale = sqrt((n.primal * n.dual_rel) / (n.dual * n.primal_rel));
        const double rho_est = std_min(std_max(scale * rho_sparse, q.rho_min), q.rho_max);
        if (rho_est > rho_sparse * q.adaptive_rho_tolerance || rho_est < rho_sparse / q.adaptive_rho_tolerance) {
          rho_sparse = rho_est;
          csqp_apply_rho(p, rho_sparse, q.rho_min);
          __syncthreads();
        }
        double tol_p, tol_d;
        if (fabs(q.eps_rel) <= 2.220446049250313e-16) {
          tol_p = q.eps_abs; tol_d = q.eps_abs;
        } else {
          tol_p = q.eps_abs + q.eps_rel * n.primal_rel;
          tol_d = q.eps_abs + q.eps_rel * n.dual_rel;
        }
        if (n.primal <= tol_p && n.dual <= tol_d) { converged = 1; break; }
      }
    }
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    a.status[b] = st; a.qp_iters[b] = it; a.converged[b] = converged;
    p.scal[MIMGPU_S_RHO_SPARSE * B + b] = rho_sparse;
    p.scal[MIMGPU_S_NORM_PRIMAL * B + b] = n.primal; p.scal[MIMGPU_S_NORM_DUAL * B + b] = n.dual;
    p.scal[MIMGPU_S_NORM_PRIMAL_REL * B + b] = n.primal_rel; p.scal[MIMGPU_S_NORM_DUAL_REL * B + b] = n.dual_rel;
  }
  if (!st && kp.with_kkt) csqp_kkt_ool(p, gsm, gs);
}

/* ------------------------------------------------------------------- dispatch */
#define MIMGPU_STREAM_SHAPES(X) X(12, 6, MIMGPU_C1_SPLIT) X(36, 12, 4) X(16, 8, 4) X(8, 4, 4) X(32, 16, 4) X(4, 2, 2) X(6, 2, 2) X(10, 4, 2)


/* MIMGPU_STREAM_NT=<32|64|...>: CTA size of the non-DIRECT streaming kernel (tuning switch, read when the
 * plan is made at mimgpu_create).  The sweeps are written for any multiple of 32 threads and split their dot
 * products by SPLIT / half-warps, not by the CTA size, so the results do not depend on it. */
static inline int stream_nt_override(int nt, int maxt) {
  const char* e = getenv("MIMGPU_STREAM_NT");
  if (!e) return nt;
  const int v = atoi(e);
  return (v >= 32 && v <= maxt && v % 32 == 0) ? v : nt;
}

template <int NX, int NU, int SPLIT>
static StreamPlan stream_plan_t(const mimgpu_shape& sh, int stage_fx, int max_smem) {
  StreamPlan pl{};
  const Smem gs = smem_layout(NX, NU, sh.ng, sh.ngT, stage_fx);
  pl.generic_total = gs.total;
  if (NX % 4 == 0) {
    const RicSmem rs = ric_layout<NX, NU>(sh.ng, sh.ngT, true);
    if (rs.total > pl.generic_total) pl.generic_total = rs.total;
  }
  const int ngm = sh.ng > sh.ngT ? sh.ng : sh.ngT;
  int nt = SPLIT * (NX + NU);
  if (2 * (NX + ngm) > nt) nt = 2 * (NX + ngm);
  nt = ((nt + 31) / 32) * 32;
  if (nt > StreamTune<NX, NU>::MAXT) nt = StreamTune<NX, NU>::MAXT;
  nt = stream_nt_override(nt, StreamTune<NX, NU>::MAXT);
  pl.nt = nt;
  const int nvmax = StreamCfg<NX, NU>::VPT * nt;
  if (StreamCfg<NX, NU>::nv_bwd(sh.ng) > nvmax || StreamCfg<NX, NU>::nv_fwd(ngm) > nvmax) return pl;
  for (int ns = MIMGPU_STREAM_MAX_STAGES; ns >= 2; --ns) {
    const StreamSmem ss = stream_layout<NX, NU>(sh.ng, sh.ngT, ns, pl.generic_total);
    /* more than two stages only while the shape's target of resident CTAs still fits an SM */
    const int bytes = ss.total * 8;
    if (bytes > max_smem) continue;
    if (ns >= 3 && StreamTune<NX, NU>::MINB * (bytes + 1024) > 233472) continue;
    pl.ns = ns; pl.smem_bytes = bytes;
    if (cudaFuncSetAttribute(k_csqp_direction_stream<NX, NU, SPLIT>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                             bytes) != cudaSuccess) {
      cudaGetLastError();
      return pl;
    }
    pl.ok = 1;
    return pl;
  }
  return pl;
}

/* shapes whose knot does not fit two stages: no stage area, sweeps read global memory (L2-prefetched) */
template <int NX, int NU, int SPLIT>
static StreamPlan stream_plan_direct_t(const mimgpu_shape& sh, int stage_fx, int max_smem) {
  StreamPlan pl{};
  const Smem gs = smem_layout(NX, NU, sh.ng, sh.ngT, stage_fx);
  pl.generic_total = gs.total;
  const int ngm = sh.ng > sh.ngT ? sh.ng : sh.ngT;
  int nt = SPLIT * (NX + NU);
  if (2 * (NX + ngm) > nt) nt = 2 * (NX + ngm);
  nt = ((nt + 31) / 32) * 32;
  if (nt > StreamTune<NX, NU>::MAXT) nt = StreamTune<NX, NU>::MAXT;
  pl.nt = nt;
  const int nvmax = StreamCfg<NX, NU>::VPT * nt;
  if (StreamCfg<NX, NU>::nv_bwd(sh.ng) > nvmax || StreamCfg<NX, NU>::nv_fwd(ngm) > nvmax) return pl;
  const StreamSmem ss = stream_layout<NX, NU>(sh.ng, sh.ngT, 0, pl.generic_total);
  const int bytes = ss.total * 8;
  if (bytes > max_smem) return pl;
  pl.ns = 0; pl.smem_bytes = bytes; pl.direct = 1;
  if (cudaFuncSetAttribute(k_csqp_direction_stream<NX, NU, SPLIT, true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                           bytes) != cudaSuccess) {
    cudaGetLastError();
    return pl;
  }
  pl.ok = 1;
  return pl;
}

#define MIMGPU_DIRECT_SHAPES(X) X(64, 32, 4) X(76, 32, 4) X(48, 24, 4)

static inline StreamPlan stream_plan(const mimgpu_shape& sh, int stage_fx, int max_smem) {
#define X(NX_, NU_, SP_) \
  if (sh.nx == NX_ && sh.nu == NU_) return stream_plan_t<NX_, NU_, SP_>(sh, stage_fx, max_smem);
  MIMGPU_STREAM_SHAPES(X)
#undef X
#define X(NX_, NU_, SP_) \
  if (sh.nx == NX_ && sh.nu == NU_) return stream_plan_direct_t<NX_, NU_, SP_>(sh, stage_fx, max_smem);
  MIMGPU_DIRECT_SHAPES(X)
#undef X
  return StreamPlan{};
}

static inline int stream_csqp_launch(const KArgs& a, const mimgpu_csqp_params& q, int with_kkt, const StreamPlan& pl,
                                     cudaStream_t st) {
  StreamKPar kp;
  kp.q = q; kp.with_kkt = with_kkt; kp.dbg = getenv("MIMGPU_DBG") ? atoi(getenv("MIMGPU_DBG")) : 0; kp.ns = pl.ns; kp.generic_total = pl.generic_total;
#define X(NX_, NU_, SP_)                                                                              \
  if (a.sh.nx == NX_ && a.sh.nu == NU_) {                                                             \
    k_csqp_direction_stream<NX_, NU_, SP_><<<a.n_active, pl.nt, pl.smem_bytes, st>>>(a, kp);          \
    return 0;                                                                                         \
  }
  MIMGPU_STREAM_SHAPES(X)
#undef X
#define X(NX_, NU_, SP_)                                                                              \
  if (a.sh.nx == NX_ && a.sh.nu == NU_) {                                                             \
    k_csqp_direction_stream<NX_, NU_, SP_, true><<<a.n_active, pl.nt, pl.smem_bytes, st>>>(a, kp);    \
    return 0;                                                                                         \
  }
  MIMGPU_DIRECT_SHAPES(X)
#undef X
  return -1;
}

/* ============================================================ SQP / FDDP CTA kernels */

static __device__ __noinline__ void sqp_forward_kkt_ool(const Prob& p, double* sm, const Smem& s, double mu, int stage_fx) {
  sqp_forward_kkt(p, as_shared(sm), s, mu, stage_fx);
}
static __device__ __noinline__ void fddp_expected_improvement_ool(const Prob& p, double* sm, const Smem& s, int feas) {
  fddp_expected_improvement(p, as_shared(sm), s, feas);
}

/* SolverSQP::computeDirection (src/sqp.cpp:224-242) / SolverFDDP::computeDirection +
 * updateExpectedImprovement (src/fddp.cpp:189-204, 256-279): DMMA backward sweep, then the
 * shape-generic forward pass / KKT (SQP) or expected-improvement reductions (FDDP). */
template <int NX, int NU>
struct CtaTune { /* 4 lanes per right-hand side of the gain solve; MINB = resident CTAs per SM aimed at */
  static constexpr int MAXT = ((4 * (NX + 1) + 31) / 32) * 32 > 512 ? 512 : ((4 * (NX + 1) + 31) / 32) * 32;
  static constexpr int MINB = NX <= 16 ? 6 : NX <= 24 ? 4 : NX <= 36 ? 3 : 1;
};
/* per-shape overrides (build-time switches for tuning runs, scripts/build_variant.sh) */
#ifdef MIMGPU_CTA_N64_MAXT
template <> struct CtaTune<64, 32> { static constexpr int MAXT = MIMGPU_CTA_N64_MAXT, MINB = 1; };
#endif
#ifdef MIMGPU_CTA_N76_MAXT
template <> struct CtaTune<76, 32> { static constexpr int MAXT = MIMGPU_CTA_N76_MAXT, MINB = 1; };
#endif
#ifdef MIMGPU_CTA_N32_MAXT
#ifndef MIMGPU_CTA_N32_MINB
#define MIMGPU_CTA_N32_MINB 3
#endif
template <> struct CtaTune<32, 16> { static constexpr int MAXT = MIMGPU_CTA_N32_MAXT, MINB = MIMGPU_CTA_N32_MINB; };
#endif

template <int NX, int NU, bool FDDP>
__global__ void __launch_bounds__(CtaTune<NX, NU>::MAXT, CtaTune<NX, NU>::MINB) k_sqp_direction_cta(KArgs a, SqpKPar kp, int nbuf, int ns) {
  extern __shared__ __align__(16) double sm_cta[];
  if ((int)blockIdx.x >= a.n_active) return;
  const int b = a.active ? a.active[blockIdx.x] : (int)blockIdx.x + a.first;
  Prob p = bind(a, b);
  uint64_t* bars = reinterpret_cast<uint64_t*>(sm_cta); /* [0,1]: backward sweep, [2..5): forward pipeline */
  double* um = sm_cta + 8;
  const RicSmem rs = ric_layout<NX, NU>(0, 0, false, nbuf);
  const Smem gs = smem_layout(NX, NU, 0, 0, a.stage_fx);
  uint32_t seq = 0u;
  if (threadIdx.x == 0) {
    for (int i = 0; i < 5; ++i) mbar_init(bars + i, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  RicPar rp;
  rp.preg = a.preg_arr ? a.preg_arr[b] : kp.preg;
  rp.dreg = a.dreg_arr ? a.dreg_arr[b] : kp.dreg;
  rp.sigma = 0.0;
  rp.is_feasible = kp.is_feasible;
  if (FDDP) {
    const double gap = gap_norm_all(p, um + gs.red);
    if (threadIdx.x == 0) p.scal[MIMGPU_S_GAP_NORM * (size_t)p.B + b] = gap;
  }
#ifdef MIMGPU_RIC_TIMING
  const long long tk0_ = clock64();
#endif
  const int st = ric_backward_mma<MODE_DDP, NX, NU>(p, um, rs, rp, bars, seq);
#ifdef MIMGPU_RIC_TIMING
  const long long tk1_ = clock64();
#endif
  if (threadIdx.x == 0) { a.status[b] = st; a.qp_iters[b] = 0; }
  if (st) {
    if (!FDDP) { /* gap norm / merit are computed before the backward pass in the reference */
      const double gap = gap_norm_all(p, um + gs.red);
      if (threadIdx.x == 0) {
        p.scal[MIMGPU_S_GAP_NORM * (size_t)p.B + b] = gap;
        p.scal[MIMGPU_S_MERIT * (size_t)p.B + b] = p.cost + kp.mu_dynamic * gap;
      }
    }
    return;
  }
  __syncthreads();
  if (FDDP) {
    fddp_expected_improvement_ool(p, um, gs, kp.is_feasible);
  } else if (ns >= 2) {
    fence_proxy_async(); /* K, Vxx written through the generic proxy, read back by TMA */
    __syncthreads();
    PipeSqp<NX, NU> pipe;
    pipe.bars = bars + 2;
    pipe.stages = um + SqpFwdSmem<NX, NU>::stages;
    pipe.ns = ns; pipe.n_issue = 0; pipe.n_wait = 0;
    stream_sqp_forward<NX, NU, 4>(p, um, pipe, kp.mu_dynamic);
  } else {
    sqp_forward_kkt_ool(p, um, gs, kp.mu_dynamic, a.stage_fx);
  }
#ifdef MIMGPU_RIC_TIMING
  __syncthreads();
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    const long long tk2_ = clock64();
    printf("ric timing nx=%d nu=%d T=%d (cycles, CTA 0): load+pack %lld  V+B %lld  D+C %lld  LLT %lld  solve %lld  F %lld  "
           "sym %lld  close+spill %lld | backward %lld forward %lld\n", NX, NU, p.T, g_ric_t[0], g_ric_t[1], g_ric_t[2],
           g_ric_t[3], g_ric_t[4], g_ric_t[5], g_ric_t[6], g_ric_t[7], tk1_ - tk0_, tk2_ - tk1_);
    for (int i = 0; i < 16; ++i) g_ric_t[i] = 0;
  }
#endif
}

#define MIMGPU_CTA_SHAPES(X) X(36, 12) X(64, 32) X(76, 32) X(32, 16) X(16, 8) X(24, 12) X(48, 24)

template <int NX, int NU>
static CtaPlan cta_plan_t(const mimgpu_shape& sh, int stage_fx, int max_smem) {
  CtaPlan pl{};
  const Smem gs = smem_layout(NX, NU, 0, 0, stage_fx);
  pl.nt = CtaTune<NX, NU>::MAXT;
  /* first pass: the largest staging that still lets MINB CTAs share an SM; second pass: whatever fits */
  const int budget = 233472 / CtaTune<NX, NU>::MINB - 1024;
  for (int pass = 0; pass < 2; ++pass) {
    const int limit = (pass == 0 && budget < max_smem) ? budget : max_smem;
    for (int nbuf = 2; nbuf >= 1; --nbuf) {
      const RicSmem rs = ric_layout<NX, NU>(0, 0, false, nbuf);
      int tot = rs.total > gs.total ? rs.total : gs.total;
      if ((8 + tot) * 8 > limit) continue;
      /* streaming forward pass: 3 or 2 stages next to its vectors, inside the same union */
      pl.ns = 0;
      for (int ns = 3; ns >= 2; --ns) {
        const int f = SqpFwdSmem<NX, NU>::stages + ns * PipeSqp<NX, NU>::STAGE;
        if ((8 + f) * 8 <= limit) { pl.ns = ns; if (f > tot) tot = f; break; }
      }
      pl.nbuf = nbuf; pl.smem_bytes = (8 + tot) * 8;
      if (getenv("MIMGPU_DBG_PLAN"))
        fprintf(stderr, "cta_plan<%d,%d>: nt %d nbuf %d ns %d smem %d B (pass %d, MINB %d)\n", NX, NU, pl.nt, pl.nbuf,
                pl.ns, pl.smem_bytes, pass, CtaTune<NX, NU>::MINB);
      if (cudaFuncSetAttribute(k_sqp_direction_cta<NX, NU, false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                               pl.smem_bytes) != cudaSuccess ||
          cudaFuncSetAttribute(k_sqp_direction_cta<NX, NU, true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                               pl.smem_bytes) != cudaSuccess) {
        cudaGetLastError();
        return pl;
      }
      pl.ok = 1;
      return pl;
    }
  }
  return pl;
}
static inline CtaPlan cta_plan(const mimgpu_shape& sh, int stage_fx, int max_smem) {
#define X(NX_, NU_) \
  if (sh.nx == NX_ && sh.nu == NU_) return cta_plan_t<NX_, NU_>(sh, stage_fx, max_smem);
  MIMGPU_CTA_SHAPES(X)
#undef X
  return CtaPlan{};
}
static inline int cta_sqp_launch(const KArgs& a, const SqpKPar& kp, bool fddp, const CtaPlan& pl, cudaStream_t st) {
#define X(NX_, NU_)                                                                                      \
  if (a.sh.nx == NX_ && a.sh.nu == NU_) {                                                                \
    if (fddp) k_sqp_direction_cta<NX_, NU_, true><<<a.n_active, pl.nt, pl.smem_bytes, st>>>(a, kp, pl.nbuf, pl.ns);  \
    else k_sqp_direction_cta<NX_, NU_, false><<<a.n_active, pl.nt, pl.smem_bytes, st>>>(a, kp, pl.nbuf, pl.ns);      \
    return 0;                                                                                            \
  }
  MIMGPU_CTA_SHAPES(X)
#undef X
  return -1;
}

}  // namespace mimgpu
