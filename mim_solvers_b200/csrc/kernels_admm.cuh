/*
 * kernels_admm.cuh -- the ADMM loop of SolverCSQP::computeDirection (src/csqp.cpp:454-517) with
 * the light iterations CONDENSED on the closed-loop dynamics.
 *
 * Between two rho updates (rho_update_interval - 1 = 24 of every 25 iterations) the reference
 * runs backwardPass_without_rho_update (src/csqp.cpp:1025-1108), forwardPass (:630-647) and
 * update_lagrangian_parameters (:1196-1340) with K_t, LLT(Quu_t), Vxx_t frozen.  Written out,
 * those two sweeps are LINEAR recurrences driven by the closed-loop matrix
 *
 *     Acl_t = Fx_t - Fu_t K_t                                   (frozen with K_t)
 *
 *   backward  Vx_t  = Qx_t - K_t^T Qu_t                         (src/csqp.cpp:1101-1104)
 *                   = r_t + Acl_t^T tmp_t ,   tmp_t = Vx_{t+1} + Vxx_fs_t          (:1060)
 *             r_t   = qx_t - K_t^T qu_t
 *             qx_t  = Lx_t - sigma dx_t + Gx_t^T d_t   (t > 0),  d_t = y_t - rho_t o z_t  (:1068-1078)
 *             qu_t  = Lu_t - sigma du_t + Gu_t^T d_t                               (:1083-1092)
 *   gains     Qu_t  = qu_t + Fu_t^T tmp_t ,   k_t = Quu_t^-1 Qu_t                  (:1093-1097)
 *   forward   dx~_{t+1} = b_t + Acl_t dx~_t ,  b_t = fs_{t+1} - Fu_t k_t           (:636-644)
 *             du~_t = -k_t - K_t dx~_t
 *
 * so the only work that is SEQUENTIAL in t is one NX x NX mat-vec per knot and sweep (the two
 * "chains"); everything else (r_t, qu_t, Qu_t, k_t, b_t, du~_t, the constraint rows c_t =
 * Gx dx~ + Gu du~, the z / y projection and the residual norms) is independent across knots and
 * runs warp-per-knot, streaming each of K, Gx, Gu, Fu, LLT once per iteration straight from
 * global memory (coalesced, no shared-memory staging, many knots in flight per SM).  Per
 * iteration and knot the kernel reads 2 NX^2 + 2 NX NU + NU^2 + ng (NX + NU) doubles instead of
 * the 2 (NX^2 + 2 NX NU + ng (NX + NU)) + NU^2 of the sweep-by-sweep form, and the latency
 * chain per knot drops from five dependent CTA-wide phases to one.  The arithmetic is the
 * reference's up to the order of the floating-point sums (scripts/proto_condensed_admm.py:
 * 1e-14 relative after 25 iterations on every BASELINE shape; parity bar 1e-9).
 *
 * Full sweeps (iteration k = 1 mod interval, src/csqp.cpp:477) keep the Riccati code of
 * kernels_ric.cuh (DMMA) / kernels_generic.cuh; right after one, Acl_t and b_t are rebuilt from the
 * fresh K_t, k_t.
 */
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <type_traits>

#include "launchers.h"
#include "kernels_ric.cuh"

namespace mimgpu {

#define MG_FULLMASK 0xffffffffu

/* ------------------------------------------------------------------ shared memory */
struct AdmmSmem {
  int bars, wbars, cbars, red, uni, vb, ws0, wsn, stg, total;
};
/* doubles of one warp's TMA stage in the knot-parallel phases: update = [K | Gx Gu (if stage_g)],
 * gains = [Fu | LLT] */
template <int NX, int NU>
__host__ __device__ inline int admm_stage_doubles(int ng, bool staged, bool stage_g) {
  if (!staged) return 0;
  const int a = NX * NU + (stage_g ? ng * (NX + NU) : 0), b = NX * NU + NU * NU;
  return a > b ? a : b;
}
template <int NX, int NU>
__host__ __device__ inline AdmmSmem admm_layout(int ng, int ngT, int nwarps, int ric_total, int stage_doubles) {
  AdmmSmem s;
  const int ngm = ng > ngT ? ng : ngT;
  s.bars = 0;                         /* two mbarriers of the Riccati sweep */
  s.wbars = 8;                        /* three mbarriers per warp (pipeline of the knot-parallel phases) */
  s.cbars = 8 + 48;                   /* full[8] + empty[8] mbarriers of the chains' TMA ring */
  s.red = 8 + 48 + 16;
  s.uni = 8 + 48 + 16 + 64;
  s.vb = s.uni;                       /* chain vector, double buffered */
  s.ws0 = s.uni + 2 * NX;             /* per-warp scratch + stage of the knot-parallel phases */
  s.stg = (2 * NX + 2 * NU + 3 * ngm + 1) & ~1;
  s.wsn = s.stg + ((stage_doubles + 1) & ~1);
  const int a = 2 * NX + nwarps * s.wsn;
  s.total = s.uni + (a > ric_total ? a : ric_total);
  return s;
}
/* per-warp scratch offsets */
template <int NX, int NU>
struct AdmmWs {
  static constexpr int xs = 0, us = NX, qx = NX + NU, qu = 2 * NX + NU, ds = 2 * NX + 2 * NU;
};

/* --------------------------------------------------------------- warp primitives */
/* y[o] = sum_k M[o * NIN + k] v[k], o < nrows: rows of M contiguous (16-byte aligned, NIN even),
 * four lanes per row, eight rows per pass, 16-byte loads.  f(o, sum) runs on one lane per row. */
template <int NIN, class F>
__device__ __forceinline__ void wdot4(const double* M, int nrows, const double* v, int lane, F f) {
  static_assert(NIN % 2 == 0, "16-byte row loads");
  constexpr int KP = (NIN + 7) / 8;
  const int g = lane >> 2, q = lane & 3;
  for (int o0 = 0; o0 < nrows; o0 += 8) {
    const int o = o0 + g;
    const bool valid = o < nrows;
    const double* row = M + (size_t)(valid ? o : 0) * NIN;
    double a0 = 0.0, a1 = 0.0;
#pragma unroll
    for (int m = 0; m < KP; ++m) {
      const int k = 2 * q + 8 * m;
      if (k < NIN) {
        const double2 mv = *reinterpret_cast<const double2*>(row + k);
        const double2 vv = *reinterpret_cast<const double2*>(v + k);
        a0 = fma(mv.x, vv.x, a0);
        a1 = fma(mv.y, vv.y, a1);
      }
    }
    double s = a0 + a1;
    s += __shfl_xor_sync(MG_FULLMASK, s, 1);
    s += __shfl_xor_sync(MG_FULLMASK, s, 2);
    if (valid && q == 0) f(o, s);
  }
}

/* Columns of a run-time-tall matrix against up to three vectors at once:
 * s_e[o] = sum_r M[r + nin * o] v_e[r], o < nrows (column o of a col-major nin x nrows block).
 * U0: v0 takes part; U12: v1 and v2 take part (check iterations).  f(o, s0, s1, s2) runs on one
 * lane per column. */
template <bool U0, bool U12, class F>
__device__ __forceinline__ void wdot4_rtv(const double* M, int nrows, int nin, const double* v0, const double* v1,
                                          const double* v2, int lane, F f) {
  const int g = lane >> 2, q = lane & 3;
  for (int o0 = 0; o0 < nrows; o0 += 8) {
    const int o = o0 + g;
    const bool valid = o < nrows;
    const double* col = M + (size_t)(valid ? o : 0) * nin;
    double s0 = 0.0, s1 = 0.0, s2 = 0.0;
    /* four independent loads in flight per lane before the first FMA (global-memory operands: L2 latency) */
    int r = q;
    for (; r + 12 < nin; r += 16) {
      const double m0 = col[r], m1 = col[r + 4], m2 = col[r + 8], m3 = col[r + 12];
      if (U0) { s0 = fma(m0, v0[r], s0); s0 = fma(m1, v0[r + 4], s0); s0 = fma(m2, v0[r + 8], s0); s0 = fma(m3, v0[r + 12], s0); }
      if (U12) {
        s1 = fma(m0, v1[r], s1); s1 = fma(m1, v1[r + 4], s1); s1 = fma(m2, v1[r + 8], s1); s1 = fma(m3, v1[r + 12], s1);
        s2 = fma(m0, v2[r], s2); s2 = fma(m1, v2[r + 4], s2); s2 = fma(m2, v2[r + 8], s2); s2 = fma(m3, v2[r + 12], s2);
      }
    }
    for (; r < nin; r += 4) {
      const double mv = col[r];
      if (U0) s0 = fma(mv, v0[r], s0);
      if (U12) { s1 = fma(mv, v1[r], s1); s2 = fma(mv, v2[r], s2); }
    }
#pragma unroll
    for (int off = 1; off <= 2; off <<= 1) {
      if (U0) s0 += __shfl_xor_sync(MG_FULLMASK, s0, off);
      if (U12) { s1 += __shfl_xor_sync(MG_FULLMASK, s1, off); s2 += __shfl_xor_sync(MG_FULLMASK, s2, off); }
    }
    if (valid && q == 0) f(o, s0, s1, s2);
  }
}
template <class F>
__device__ __forceinline__ void wdot4_rt3(const double* M, int nrows, int nin, const double* v0, const double* v1,
                                          const double* v2, int lane, F f) {
  if (v0 && v1) wdot4_rtv<true, true>(M, nrows, nin, v0, v1, v2, lane, f);
  else if (v0) wdot4_rtv<true, false>(M, nrows, nin, v0, v1, v2, lane, f);
  else if (v1) wdot4_rtv<false, true>(M, nrows, nin, v0, v1, v2, lane, f);
  else wdot4_rtv<false, false>(M, nrows, nin, v0, v1, v2, lane, f);
}

/* y[i] = sum_j M[i + NI * j] w[j], i < NI, j < NJ (col-major NI x NJ, lanes walk consecutive i:
 * coalesced).  Shapes with NI <= 16 split the columns over 32 / NP lane groups.  f(i, sum) runs on
 * one lane per row. */
template <int NI, int NJ, class F>
__device__ __forceinline__ void waxpy(const double* M, const double* w, int lane, F f) {
  if constexpr (NI <= 16) {
    constexpr int NP = NI <= 4 ? 4 : NI <= 8 ? 8 : 16, PH = 32 / NP;
    const int i = lane % NP, ph = lane / NP;
    const bool valid = i < NI;
    const int ic = valid ? i : 0;
    double a0 = 0.0;
#pragma unroll
    for (int j0 = 0; j0 < NJ; j0 += PH) {
      const int j = j0 + ph;
      if (j < NJ) a0 = fma(M[ic + NI * j], w[j], a0);
    }
#pragma unroll
    for (int off = NP; off < 32; off <<= 1) a0 += __shfl_xor_sync(MG_FULLMASK, a0, off);
    if (valid && ph == 0) f(i, a0);
  } else {
#pragma unroll
    for (int i0 = 0; i0 < NI; i0 += 32) {
      const int i = i0 + lane;
      const bool valid = i < NI;
      const int ic = valid ? i : 0;
      double a0 = 0.0, a1 = 0.0;
#pragma unroll 8
      for (int j = 0; j + 1 < NJ; j += 2) {
        a0 = fma(M[ic + NI * j], w[j], a0);
        a1 = fma(M[ic + NI * (j + 1)], w[j + 1], a1);
      }
      if (NJ & 1) a0 = fma(M[ic + NI * (NJ - 1)], w[NJ - 1], a0);
      if (valid) f(i, a0 + a1);
    }
  }
}

/* x := (L L^T)^-1 x for one right-hand side held one element per lane (NU <= 32), L = lower
 * Cholesky factor, col-major NU x NU in global memory.  Column-oriented substitution: every
 * x[i] subtracts its terms in the reference's order (Eigen LLT::solveInPlace, src/csqp.cpp:1096). */
template <int NU>
__device__ __forceinline__ double wsolve_llt(const double* L, double xi, int lane) {
  static_assert(NU <= 32, "one lane per control");
  const bool in = lane < NU;
  const int lc = in ? lane : 0;
  /* 1 / L_jj once per knot (lane j), multiplied in: one division per lane instead of one per step */
  const double rdiag = 1.0 / (in ? L[lc + NU * lc] : 1.0);
  double lv[NU];
#pragma unroll
  for (int j = 0; j < NU; ++j) lv[j] = (in && lane > j) ? L[lc + NU * j] : 0.0;   /* column entries of my row */
#pragma unroll
  for (int j = 0; j < NU; ++j) {
    const double xj = __shfl_sync(MG_FULLMASK, xi * rdiag, j);
    if (lane == j) xi = xj;
    else xi = fma(-lv[j], xj, xi);   /* lv[j] = 0 for lanes <= j */
  }
#pragma unroll
  for (int j = 0; j < NU; ++j) lv[j] = (in && lane < j) ? L[j + NU * lc] : 0.0;   /* row j, my column */
#pragma unroll
  for (int j = NU - 1; j >= 0; --j) {
    const double xj = __shfl_sync(MG_FULLMASK, xi * rdiag, j);
    if (lane == j) xi = xj;
    else xi = fma(-lv[j], xj, xi);
  }
  return xi;
}

__device__ __forceinline__ double warp_nmax_all(double v) {
#pragma unroll
  for (int o = 16; o; o >>= 1) v = nmax(v, __shfl_xor_sync(MG_FULLMASK, v, o));
  return v;
}
__device__ __forceinline__ void named_bar_sync(int id, int nthreads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

/* L2 prefetch: the knot-parallel phases and the chains pull the NEXT knots' blocks into L2 while the
 * current ones are processed, so that the dependent steps of a knot pay L2 latency, not DRAM's. */
__device__ __forceinline__ void pf_line(const void* p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }
/* contiguous block, 16-byte aligned, even number of doubles: one bulk prefetch by one lane */
__device__ __forceinline__ void pf_block(const double* p, int ndoubles, int lane) {
  if (lane == 0 && ndoubles > 0) l2_prefetch(p, (uint32_t)ndoubles * 8u);
}
/* short vector: one 128-byte line per lane */
__device__ __forceinline__ void pf_vec(const double* p, int ndoubles, int lane) {
  if (lane >= 0 && lane * 16 < ndoubles) pf_line(p + lane * 16);
}

/* ----------------------------------------------------------- per-problem scratch */
struct AdmmScr {
  double* acl;   /* [T][NX*NX] Acl_t col-major */
  double* rv;    /* [T][NX]    r_t */
  double* tv;    /* [T][NX]    tmp_t = Vx_{t+1} + Vxx_fs_t */
};

/* ------------------------------------------------------ after a full Riccati sweep */
/* Acl_t = Fx_t - Fu_t K_t and b_t = fs_{t+1} - Fu_t k_t (stored in the
 * dx~_{t+1} slot, the forward chain turns it into dx~_{t+1}); one warp per knot. */
template <int NX, int NU>
static __device__ __noinline__ void admm_post_full(const Prob& p, const AdmmScr& sc, double* sm, const AdmmSmem& s) {
  extern __shared__ __align__(16) double sm_admm[];
  sm = sm_admm;
  using W = AdmmWs<NX, NU>;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
  double* ws = sm + s.ws0 + warp * s.wsn;
  for (int t = warp; t < p.T; t += nw) {
    const double* Fx = p.Fx + (size_t)t * NX * NX;
    const double* Fu = p.Fu + (size_t)t * NX * NU;
    const double* K = p.K + (size_t)t * NX * NU;
    double* A = sc.acl + (size_t)t * NX * NX;
    /* lanes over rows i: Fu[i][.] in registers, K[.][j] broadcast */
    for (int i0 = 0; i0 < NX; i0 += 32) {
      const int i = i0 + lane;
      const bool valid = i < NX;
      const int ic = valid ? i : 0;
      double fu[NU];
#pragma unroll
      for (int a = 0; a < NU; ++a) fu[a] = Fu[ic + NX * a];
      for (int j = 0; j < NX; ++j) {
        double acc = 0.0;
#pragma unroll
        for (int a = 0; a < NU; ++a) acc = fma(fu[a], K[a * NX + j], acc);
        if (valid) A[i + NX * j] = Fx[i + NX * j] - acc;
      }
    }
    /* b_t */
    for (int a = lane; a < NU; a += 32) ws[W::us + a] = p.kff[(size_t)t * NU + a];
    __syncwarp();
    waxpy<NX, NU>(Fu, ws + W::us, lane, [&](int i, double acc) {
      p.dxt[(size_t)(t + 1) * NX + i] = p.fs[(size_t)(t + 1) * NX + i] - acc;
    });
    __syncwarp();
  }
  fence_proxy_async(); /* Acl, b_t (and the sweep's K, LLT) were written through the generic proxy; TMA reads them */
}

/* --------------------------------------------------------------------- the chains */
/* BWD: Vx_t = r_t + Acl_t^T tmp_t, tmp_{t-1} = Vx_t + Vxx_fs_{t-1}, t = T-1 .. 0
 * FWD: dx~_{t+1} = b_t + Acl_t dx~_t, t = 0 .. T-1
 * Four lanes per output row, the warps holding threads < nchain take part (named barrier 1), each
 * thread prefetches its slice of the next two knots' matrix through registers. */
struct ChainSeq {
  int slot;        /* ring slot of the next load */
  uint32_t phase;  /* mbarrier phase parity of that slot */
};
template <int NX>
__host__ __device__ constexpr int admm_chain_stage() { return NX * NX + 2 * NX; } /* [M | add0 | add1] */

__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}

/* Warps [0, nchain / 32) compute (named barrier 1); the NEXT warp is the producer: its lane 0 streams
 * step st's operands into ring slot st % ns as soon as the consumers have released that slot
 * (full[slot] = TMA completion, empty[slot] = one arrive by consumer thread 0 after the step's
 * barrier).  The bulk-copy issue is off the consumers' critical path. */
template <int NX, bool BWD>
static __device__ __noinline__ void admm_chain(const Prob& p, const AdmmScr& sc, double* sm, const AdmmSmem& s, int nchain,
                                               int ns, ChainSeq& cs) {
  extern __shared__ __align__(16) double sm_admm[];
  sm = sm_admm;
  const int tid = threadIdx.x;
  constexpr int KP = (NX + 7) / 8, SD = admm_chain_stage<NX>();
  const int T = p.T;
  const double* M = sc.acl;    /* col-major: BWD walks columns (rows of Acl^T), FWD walks rows with stride NX */
  double* ring = sm + s.ws0;   /* the knot-parallel phases' per-warp area is idle during a chain */
  uint64_t* full = reinterpret_cast<uint64_t*>(sm + s.cbars);
  uint64_t* empty = full + 8;
  auto knot_of = [&](int st) { return BWD ? T - 1 - st : st; };
  /* ring position after this call (every thread of the CTA keeps the same bookkeeping) */
  const int slot0 = cs.slot;
  const uint32_t phase0 = cs.phase;
  {
    const int adv = slot0 + T;
    cs.slot = adv % ns;
    cs.phase = phase0 ^ (uint32_t)((adv / ns) & 1);
  }
  if (tid >= nchain) {
    if (tid == nchain) { /* producer */
      int slot = slot0;
      uint32_t ephase = phase0 ^ 1u; /* a fresh / fully drained slot: waiting on the PREVIOUS phase returns at once */
      for (int st = 0; st < T; ++st) {
        mbar_wait(empty + slot, ephase);
        const int t = knot_of(st);
        double* dst = ring + (size_t)slot * SD;
        uint64_t* bar = full + slot;
        const bool two = BWD && t > 0;
        mbar_expect_tx(bar, (uint32_t)(NX * NX + NX + (two ? NX : 0)) * 8u);
        tma_bulk_g2s(dst, M + (size_t)t * NX * NX, NX * NX * 8, bar);
        if (BWD) {
          tma_bulk_g2s(dst + NX * NX, sc.rv + (size_t)t * NX, NX * 8, bar);
          if (two) tma_bulk_g2s(dst + NX * NX + NX, p.Vxx_fs + (size_t)(t - 1) * NX, NX * 8, bar);
        } else {
          tma_bulk_g2s(dst + NX * NX, p.dxt + (size_t)(t + 1) * NX, NX * 8, bar);
        }
        if (++slot == ns) { slot = 0; ephase ^= 1u; }
      }
    }
    return;
  }
  const int i = tid >> 2, q = tid & 3;
  const bool valid = i < NX;
  const int ic = valid ? i : 0;
  double* vb = sm + s.vb;
  if (tid < NX) vb[tid] = BWD ? sc.tv[(size_t)(T - 1) * NX + tid] : p.dxt[tid];
  named_bar_sync(1, nchain);
  int wslot = slot0;
  uint32_t wphase = phase0;
  for (int st = 0; st < T; ++st) {
    const int t = knot_of(st);
    const double* v = vb + (st & 1) * NX;
    double* vn = vb + ((st + 1) & 1) * NX;
    mbar_wait(full + wslot, wphase);
    const double* stg = ring + (size_t)wslot * SD;
    double a0 = 0.0, a1 = 0.0;
    if (BWD) { /* (Acl^T tmp)_i: column i of Acl is contiguous */
      const double* row = stg + ic * NX;
#pragma unroll
      for (int m = 0; m < KP; ++m) {
        const int k = 2 * q + 8 * m;
        if (k < NX) {
          const double2 mv = *reinterpret_cast<const double2*>(row + k);
          const double2 vv = *reinterpret_cast<const double2*>(v + k);
          a0 = fma(mv.x, vv.x, a0);
          a1 = fma(mv.y, vv.y, a1);
        }
      }
    } else {   /* (Acl x)_i: row i, stride NX; lane q takes columns q, q + 4, ... */
#pragma unroll
      for (int m = 0; m < (NX + 7) / 8; ++m) {
        const int j0 = q + 8 * m, j1 = j0 + 4;
        if (j0 < NX) a0 = fma(stg[ic + NX * j0], v[j0], a0);
        if (j1 < NX) a1 = fma(stg[ic + NX * j1], v[j1], a1);
      }
    }
    double acc = a0 + a1;
    acc += __shfl_xor_sync(MG_FULLMASK, acc, 1);
    acc += __shfl_xor_sync(MG_FULLMASK, acc, 2);
    double out0 = 0.0, out1 = 0.0;
    if (valid && q == 0) {
      out0 = stg[NX * NX + i] + acc;                            /* Vx_t | dx~_{t+1} */
      if (BWD && t > 0) { out1 = out0 + stg[NX * NX + NX + i]; vn[i] = out1; }   /* tmp_{t-1} */
      if (!BWD) vn[i] = out0;
    }
    named_bar_sync(1, nchain);   /* step st complete: its slot is free, v_{st+1} visible */
    if (tid == 0) mbar_arrive(empty + wslot);
    if (valid && q == 0) {       /* the global stores are off the chain's critical path */
      if (BWD) {
        p.Vx[(size_t)t * NX + i] = out0;
        if (t > 0) sc.tv[(size_t)(t - 1) * NX + i] = out1;
      } else {
        p.dxt[(size_t)(t + 1) * NX + i] = out0;
      }
    }
    if (++wslot == ns) { wslot = 0; wphase ^= 1u; }
  }
}

/* --------------------------------------------- gains and b_t after the backward chain */
/* Per-warp TMA stage of the knot-parallel phases (STAGED shapes): lane 0 pulls the knot's matrix
 * blocks into the warp's shared-memory stage with bulk copies that complete on the warp's
 * mbarrier; `seq` counts the loads of this warp (phase parity), it persists across phases. */
struct WarpStage {
  uint64_t* bar;
  double* buf;
  uint32_t seq;
  __device__ __forceinline__ void begin(uint32_t bytes, int lane) { if (lane == 0) mbar_expect_tx(bar, bytes); }
  __device__ __forceinline__ void copy(int off, const double* src, int ndoubles, int lane) {
    if (lane == 0 && ndoubles > 0) tma_bulk_g2s(buf + off, src, (uint32_t)ndoubles * 8u, bar);
  }
  __device__ __forceinline__ void wait() { mbar_wait(bar, seq & 1u); ++seq; }
};

/* Qu_t = qu_t + Fu_t^T tmp_t, k_t = Quu_t^-1 Qu_t (src/csqp.cpp:1083-1097), b_t = fs_{t+1} - Fu_t k_t;
 * one warp per knot.  qu_t was left in the Qu_ slot by admm_update. */
template <int NX, int NU, bool STAGED>
static __device__ __noinline__ void admm_gains(const Prob& p, const AdmmScr& sc, double* sm, const AdmmSmem& s,
                                               uint32_t& wseq) {
  extern __shared__ __align__(16) double sm_admm[];
  sm = sm_admm;
  using W = AdmmWs<NX, NU>;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
  double* ws = sm + s.ws0 + warp * s.wsn;
  double *xs = ws + W::xs, *us = ws + W::us, *qu = ws + W::qu;
  WarpStage stg{reinterpret_cast<uint64_t*>(sm + s.wbars) + warp, ws + s.stg, wseq};
  auto pull = [&](int t) { /* knot t's blocks and vectors -> L2 */
    if (t >= p.T) return;
    pf_block(p.Fu + (size_t)t * NX * NU, NX * NU, lane);
    pf_block(p.L + (size_t)t * NU * NU, NU * NU, lane - 1);
    pf_vec(sc.tv + (size_t)t * NX, NX, lane - 2);
    pf_vec(p.fs + (size_t)(t + 1) * NX, NX, lane - 10);
    pf_vec(p.Qu + (size_t)t * NU, NU, lane - 18);
  };
  pull(warp + nw);
  for (int t = warp; t < p.T; t += nw) {
    const double* Fu = p.Fu + (size_t)t * NX * NU;
    const double* L = p.L + (size_t)t * NU * NU;
    if constexpr (STAGED) {
      stg.begin((uint32_t)(NX * NU + NU * NU) * 8u, lane);
      stg.copy(0, Fu, NX * NU, lane);
      stg.copy(NX * NU, L, NU * NU, lane);
    }
    pull(t + 2 * nw);
    /* the knot's vectors: all loads in flight at once */
    double tv0 = 0.0, tv1 = 0.0, fs0 = 0.0, fs1 = 0.0, quv = 0.0;
    if (lane < NX) { tv0 = sc.tv[(size_t)t * NX + lane]; fs0 = p.fs[(size_t)(t + 1) * NX + lane]; }
    if (NX > 32 && lane + 32 < NX) { tv1 = sc.tv[(size_t)t * NX + lane + 32]; fs1 = p.fs[(size_t)(t + 1) * NX + lane + 32]; }
    if (lane < NU) quv = p.Qu[(size_t)t * NU + lane];
    if (lane < NX) xs[lane] = tv0;
    if (NX > 32 && lane + 32 < NX) xs[lane + 32] = tv1;
    if (NX > 64) for (int i = lane + 64; i < NX; i += 32) xs[i] = sc.tv[(size_t)t * NX + i];
    if (lane < NU) qu[lane] = quv;
    if constexpr (STAGED) stg.wait();
    __syncwarp();
    auto body = [&](const double* mFu, const double* mL) {
      wdot4<NX>(mFu, NU, xs, lane, [&](int a, double acc) {
        double v = qu[a];
        v += acc;
        us[a] = v;                      /* Qu_t (us doubles as its buffer until the solve) */
        p.Qu[(size_t)t * NU + a] = v;
      });
      __syncwarp();
      double ki = (lane < NU) ? us[lane] : 0.0;
      ki = wsolve_llt<NU>(mL, ki, lane);
      __syncwarp();
      if (lane < NU) { us[lane] = ki; p.kff[(size_t)t * NU + lane] = ki; }
      __syncwarp();
      waxpy<NX, NU>(mFu, us, lane, [&](int i, double acc) {
        /* the lane that owns row i also loaded fs_{t+1}[i] */
        const double f = (i < 32) ? fs0 : (i < 64) ? fs1 : p.fs[(size_t)(t + 1) * NX + i];
        p.dxt[(size_t)(t + 1) * NX + i] = f - acc;
      });
    };
    if constexpr (STAGED) body(stg.buf, stg.buf + NX * NU);
    else body(Fu, L);
    __syncwarp();
  }
  wseq = stg.seq;
  fence_proxy_async(); /* b_t -> the forward chain's TMA */
}

/* ------------------------------- du~, update_lagrangian_parameters, r_t of the next sweep */
/* One knot: du~_t = -k_t - K_t dx~_t (src/csqp.cpp:636-637); dx_ := dx~, du_ := du~ (:1222-1226,
 * 1241-1242); c = Gx dx~ + Gu du~, z_relaxed, z, y (:1229-1239); the residual norms on check
 * iterations (:1244-1278); and, when the next backward sweep is a light one, its knot-local terms
 * qu_t (left in the Qu_ slot) and r_t = qx_t - K_t^T qu_t (src/csqp.cpp:1068-1104 regrouped). */
template <int NX, int NU, bool STAGED>
static __device__ __noinline__ void admm_update(const Prob& p, const AdmmScr& sc, double* sm, const AdmmSmem& s,
                                                const mimgpu_csqp_params& q_, int iter, bool need_r, bool stage_g,
                                                AdmmNorms& n, uint32_t& wseq) {
  extern __shared__ __align__(16) double sm_admm[];
  sm = sm_admm;
  using W = AdmmWs<NX, NU>;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
  const int T = p.T, ng = p.ng, ngT = p.ngT;
  const int ngm = ng > ngT ? ng : ngT;
  double* ws = sm + s.ws0 + warp * s.wsn;
  double *xs = ws + W::xs, *us = ws + W::us, *sqx = ws + W::qx, *squ = ws + W::qu, *sd = ws + W::ds, *se1 = sd + ngm,
         *se2 = se1 + ngm;
  WarpStage stg{reinterpret_cast<uint64_t*>(sm + s.wbars) + warp, ws + s.stg, wseq};
  constexpr bool DEEP = !(NX == 36 && NU == 12);
  const bool check = (iter % q_.rho_update_interval) == 0;
  const bool heur = q_.update_rho_with_heuristic != 0;
  const bool dualg = check && !heur;
  const double alpha = q_.alpha, sigma = q_.sigma;
  double np_ = -CUDART_INF, nd_ = -CUDART_INF, npr = -CUDART_INF, ndr = -CUDART_INF; /* src/csqp.cpp:1201-1204 */

  /* row r of one knot: c given, z / y projection and the row-local norms (src/csqp.cpp:1231-1239) */
  auto project = [&](size_t o, int r, double c, double zp, double rho, double yo, double lo, double hi) {
    double zr = alpha * c;
    zr += (1. - alpha) * zp;
    const double irho = 1. / rho;
    double z = zr + yo * irho;
    z = (z < lo) ? lo : z;
    z = (hi < z) ? hi : z;
    const double yn = yo + rho * (zr - z);
    p.z[o + r] = z;
    p.y[o + r] = yn;
    double d = yn;
    d -= rho * z;
    sd[r] = d;
    if (check) {
      se1[r] = rho * (z - zp); se2[r] = yn;
      np_ = nmax(np_, fabs(c - z));
      npr = nmax(npr, fabs(c));
      npr = nmax(npr, fabs(z));
      if (heur) {
        nd_ = nmax(nd_, fabs(rho * (z - zp)));
        ndr = nmax(ndr, fabs(yn));
      }
    }
  };
  auto pull = [&](int t) { /* running knot t's blocks and vectors -> L2 */
    if (t >= T) return;
    pf_block(p.K + (size_t)t * NX * NU, NX * NU, lane);
    if (ng) {
      pf_block(p.Gx + (size_t)t * ng * NX, ng * NX, lane - 1);
      pf_block(p.Gu + (size_t)t * ng * NU, ng * NU, lane - 2);
      const size_t o = (size_t)t * ng;
      pf_vec(p.y + o, ng, lane - 3); pf_vec(p.z + o, ng, lane - 3); pf_vec(p.rho + o, ng, lane - 3);
      pf_vec(p.g + o, ng, lane - 3); pf_vec(p.glb + o, ng, lane - 3); pf_vec(p.gub + o, ng, lane - 3);
    }
    pf_vec(p.dxt + (size_t)t * NX, NX, lane - 11);
    pf_vec(p.kff + (size_t)t * NU, NU, lane - 19);
    if (need_r) { pf_vec(p.Lx + (size_t)t * NX, NX, lane - 11); pf_vec(p.Lu + (size_t)t * NU, NU, lane - 19); }
  };
  pull(warp + nw);
  for (int t = warp; t <= T; t += nw) {
    const bool last = (t == T);
    const int nc = last ? ngT : ng;
    const size_t o = (size_t)t * ng;
    const double* Kg = p.K + (size_t)(last ? 0 : t) * NX * NU;
    const double* Gxg = last ? p.GxT : (ng ? p.Gx + (size_t)t * ng * NX : nullptr);
    const double* Gug = (last || !ng) ? nullptr : p.Gu + (size_t)t * ng * NU;
    const bool sg = STAGED && stage_g && !last && ng > 0;
    if constexpr (STAGED) {
      if (!last) {
        stg.begin((uint32_t)(NX * NU + (sg ? ng * (NX + NU) : 0)) * 8u, lane);
        stg.copy(0, Kg, NX * NU, lane);
        if (sg) { stg.copy(NX * NU, Gxg, ng * NX, lane); stg.copy(NX * NU + ng * NX, Gug, ng * NU, lane); }
      }
    }
    pull(t + 2 * nw);
    /* the knot's vectors: every load in flight before the first use */
    double x0 = 0.0, x1 = 0.0, kf = 0.0, lx0 = 0.0, lx1 = 0.0, luv = 0.0;
    double zp = 0.0, rho = 1.0, yo = 0.0, lo = 0.0, hi = 0.0;
    const bool rowreg = nc > 0 && nc <= 32;   /* lane r holds row r's vectors */
    if (lane < NX) x0 = p.dxt[(size_t)t * NX + lane];
    if (NX > 32 && lane + 32 < NX) x1 = p.dxt[(size_t)t * NX + lane + 32];
    if (!last && lane < NU) kf = p.kff[(size_t)t * NU + lane];
    if (rowreg && lane < nc) {
      const double* g = last ? p.gT : p.g + o;
      const double* lb = last ? p.glbT : p.glb + o;
      const double* ub = last ? p.gubT : p.gub + o;
      zp = p.z[o + lane]; rho = p.rho[o + lane]; yo = p.y[o + lane];
      const double gv = g[lane];
      lo = lb[lane] - gv; hi = ub[lane] - gv;
    }
    if (need_r && !last) {
      if (lane < NX) lx0 = p.Lx[(size_t)t * NX + lane];
      if (NX > 32 && lane + 32 < NX) lx1 = p.Lx[(size_t)t * NX + lane + 32];
      if (lane < NU) luv = p.Lu[(size_t)t * NU + lane];
    }
    if (lane < NX) { xs[lane] = x0; p.dx[(size_t)t * NX + lane] = x0; }
    if (NX > 32 && lane + 32 < NX) { xs[lane + 32] = x1; p.dx[(size_t)t * NX + lane + 32] = x1; }
    if (NX > 64)
      for (int i = lane + 64; i < NX; i += 32) { const double x = p.dxt[(size_t)t * NX + i]; xs[i] = x; p.dx[(size_t)t * NX + i] = x; }
    if (need_r && !last) { /* qx_t / qu_t start from Lx - sigma dx~ and Lu (the G^T d and -sigma du~ terms follow) */
      if (lane < NX) sqx[lane] = lx0;
      if (NX > 32 && lane + 32 < NX) sqx[lane + 32] = lx1;
      if (NX > 64) for (int i = lane + 64; i < NX; i += 32) sqx[i] = p.Lx[(size_t)t * NX + i];
      if (lane < NU) squ[lane] = luv;
    }
    if (!last && lane < NU) us[lane] = kf;    /* k_t, replaced by du~_t below */
    if constexpr (STAGED) { if (!last) stg.wait(); }
    __syncwarp();
    const double* K = Kg;
    const double* Gx = Gxg;
    const double* Gu = Gug;
    if constexpr (STAGED) {
      if (!last) K = stg.buf;
      if (sg) { Gx = stg.buf + NX * NU; Gu = Gx + ng * NX; }
    }
    if (!last) { /* du~ = -k - K dx~ (src/csqp.cpp:636-637); us[a] holds k_t[a] and is replaced by du~_t[a] */
      wdot4<NX>(K, NU, xs, lane, [&](int a, double acc) {
        double u = -us[a];
        u -= acc;
        us[a] = u;
        p.dut[(size_t)t * NU + a] = u;
        p.du[(size_t)t * NU + a] = u;
      });
      __syncwarp();
    }
    if (nc) {
      /* c_r = sum_j Gx[r][j] dx~[j] + sum_a Gu[r][a] du~[a]: lanes over rows, columns split over lane groups */
      const int np2 = nc <= 4 ? 4 : nc <= 8 ? 8 : nc <= 16 ? 16 : 32, ph_n = 32 / np2;
      const int rl = lane % np2, ph = lane / np2;
      for (int r0 = 0; r0 < nc; r0 += np2) {
        const int r = r0 + rl;
        const bool valid = r < nc;
        const int rc = valid ? r : 0;
        double a1 = 0.0, a2 = 0.0, b1 = 0.0, b2 = 0.0;
        int j = ph;
        /* DEEP: eight independent loads in flight per lane where Gx / Gu come from global memory (half as many L2
         * round trips; same accumulation order as two passes of the four-load loop below).  Measured on B200
         * (profiles/variants_r02_v4.log): Talos +4 %, constrained LQ nx=64 +10 %, nx=32 +12 %; Solo12, whose G is
         * staged in shared memory, loses 2 % to the extra code, so 36/12 is compiled without it. */
        for (; DEEP && !sg && j + 7 * ph_n < NX; j += 8 * ph_n) {
          const double g0 = Gx[rc + nc * j], g1 = Gx[rc + nc * (j + ph_n)], g2 = Gx[rc + nc * (j + 2 * ph_n)],
                       g3 = Gx[rc + nc * (j + 3 * ph_n)], g4 = Gx[rc + nc * (j + 4 * ph_n)],
                       g5 = Gx[rc + nc * (j + 5 * ph_n)], g6 = Gx[rc + nc * (j + 6 * ph_n)],
                       g7 = Gx[rc + nc * (j + 7 * ph_n)];
          a1 = fma(g0, xs[j], a1);
          b1 = fma(g1, xs[j + ph_n], b1);
          a1 = fma(g2, xs[j + 2 * ph_n], a1);
          b1 = fma(g3, xs[j + 3 * ph_n], b1);
          a1 = fma(g4, xs[j + 4 * ph_n], a1);
          b1 = fma(g5, xs[j + 5 * ph_n], b1);
          a1 = fma(g6, xs[j + 6 * ph_n], a1);
          b1 = fma(g7, xs[j + 7 * ph_n], b1);
        }
        for (; j + 3 * ph_n < NX; j += 4 * ph_n) { /* four independent loads in flight per lane */
          const double g0 = Gx[rc + nc * j], g1 = Gx[rc + nc * (j + ph_n)], g2 = Gx[rc + nc * (j + 2 * ph_n)],
                       g3 = Gx[rc + nc * (j + 3 * ph_n)];
          a1 = fma(g0, xs[j], a1);
          b1 = fma(g1, xs[j + ph_n], b1);
          a1 = fma(g2, xs[j + 2 * ph_n], a1);
          b1 = fma(g3, xs[j + 3 * ph_n], b1);
        }
        for (; j < NX; j += ph_n) a1 = fma(Gx[rc + nc * j], xs[j], a1);
        if (Gu) {
          int a = ph;
          for (; DEEP && !sg && a + 7 * ph_n < NU; a += 8 * ph_n) {
            const double g0 = Gu[rc + nc * a], g1 = Gu[rc + nc * (a + ph_n)], g2 = Gu[rc + nc * (a + 2 * ph_n)],
                         g3 = Gu[rc + nc * (a + 3 * ph_n)], g4 = Gu[rc + nc * (a + 4 * ph_n)],
                         g5 = Gu[rc + nc * (a + 5 * ph_n)], g6 = Gu[rc + nc * (a + 6 * ph_n)],
                         g7 = Gu[rc + nc * (a + 7 * ph_n)];
            a2 = fma(g0, us[a], a2);
            b2 = fma(g1, us[a + ph_n], b2);
            a2 = fma(g2, us[a + 2 * ph_n], a2);
            b2 = fma(g3, us[a + 3 * ph_n], b2);
            a2 = fma(g4, us[a + 4 * ph_n], a2);
            b2 = fma(g5, us[a + 5 * ph_n], b2);
            a2 = fma(g6, us[a + 6 * ph_n], a2);
            b2 = fma(g7, us[a + 7 * ph_n], b2);
          }
          for (; a + 3 * ph_n < NU; a += 4 * ph_n) {
            const double g0 = Gu[rc + nc * a], g1 = Gu[rc + nc * (a + ph_n)], g2 = Gu[rc + nc * (a + 2 * ph_n)],
                         g3 = Gu[rc + nc * (a + 3 * ph_n)];
            a2 = fma(g0, us[a], a2);
            b2 = fma(g1, us[a + ph_n], b2);
            a2 = fma(g2, us[a + 2 * ph_n], a2);
            b2 = fma(g3, us[a + 3 * ph_n], b2);
          }
          for (; a < NU; a += ph_n) a2 = fma(Gu[rc + nc * a], us[a], a2);
        }
        a1 += b1; a2 += b2;
        for (int off = np2; off < 32; off <<= 1) {
          a1 += __shfl_xor_sync(MG_FULLMASK, a1, off);
          a2 += __shfl_xor_sync(MG_FULLMASK, a2, off);
        }
        if (valid && ph == 0) {
          double c = 0.0;
          c += a1;
          c += a2;
          if (rowreg) {
            project(o, r, c, zp, rho, yo, lo, hi);
          } else {
            const double* g = last ? p.gT : p.g + o;
            const double* lb = last ? p.glbT : p.glb + o;
            const double* ub = last ? p.gubT : p.gub + o;
            project(o, r, c, p.z[o + r], p.rho[o + r], p.y[o + r], lb[r] - g[r], ub[r] - g[r]);
          }
        }
      }
      __syncwarp();
    }
    /* knot-local terms of the next light backward sweep, and the G^T products of the dual norms */
    if (last) {
      if (need_r || (dualg && nc)) {
        /* column i of GxT (ngT contiguous) against d, rho dz, y */
        wdot4_rt3(p.GxT, NX, nc, (need_r && nc) ? sd : nullptr, (dualg && nc) ? se1 : nullptr,
                  (dualg && nc) ? se2 : nullptr, lane, [&](int i, double s0, double s1, double s2) {
                    if (dualg && nc) { nd_ = nmax(nd_, fabs(s1)); ndr = nmax(ndr, fabs(s2)); }
                    if (need_r) { /* Vx_T (src/csqp.cpp:1045-1052) and tmp_{T-1} */
                      double v = p.LxT[i];
                      v -= sigma * xs[i];
                      if (nc) v += s0;
                      p.Vx[(size_t)T * NX + i] = v;
                      sc.tv[(size_t)(T - 1) * NX + i] = v + p.Vxx_fs[(size_t)(T - 1) * NX + i];
                    }
                  });
      }
      __syncwarp();
      continue;
    }
    if (need_r || (dualg && nc)) {
      const bool gx_d = need_r && nc && t > 0;   /* x0 is fixed: no Gx^T d at t = 0 (src/csqp.cpp:1073) */
      if (nc) {
        wdot4_rt3(Gx, NX, nc, gx_d ? sd : nullptr, dualg ? se1 : nullptr, dualg ? se2 : nullptr, lane,
                  [&](int i, double s0, double s1, double s2) {
                    if (dualg) { nd_ = nmax(nd_, fabs(s1)); ndr = nmax(ndr, fabs(s2)); }
                    if (need_r) {
                      double v = sqx[i];
                      v -= sigma * xs[i];
                      if (gx_d) v += s0;
                      sqx[i] = v;
                    }
                  });
        wdot4_rt3(Gu, NU, nc, need_r ? sd : nullptr, dualg ? se1 : nullptr, dualg ? se2 : nullptr, lane,
                  [&](int a, double s0, double s1, double s2) {
                    if (dualg) { nd_ = nmax(nd_, fabs(s1)); ndr = nmax(ndr, fabs(s2)); }
                    if (need_r) {
                      double v = squ[a];
                      v -= sigma * us[a];
                      v += s0;
                      squ[a] = v;
                      p.Qu[(size_t)t * NU + a] = v;
                    }
                  });
      } else if (need_r) {
        for (int i = lane; i < NX; i += 32) sqx[i] -= sigma * xs[i];
        for (int a = lane; a < NU; a += 32) {
          const double v = squ[a] - sigma * us[a];
          squ[a] = v;
          p.Qu[(size_t)t * NU + a] = v;
        }
      }
      __syncwarp();
      if (need_r) { /* r_t = qx_t - K_t^T qu_t; K row-major NU x NX = col-major NX x NU */
        waxpy<NX, NU>(K, squ, lane, [&](int i, double acc) { sc.rv[(size_t)t * NX + i] = sqx[i] - acc; });
      }
    }
    __syncwarp();
  }
  wseq = stg.seq;
  fence_proxy_async(); /* r_t -> the backward chain's TMA */
  /* src/csqp.cpp:1201-1204: the four norms are -inf unless this is a check iteration */
  n.primal = n.dual = n.primal_rel = n.dual_rel = -CUDART_INF;
  if (check) {
    double* red = sm + s.red;
    np_ = warp_nmax_all(np_); nd_ = warp_nmax_all(nd_); npr = warp_nmax_all(npr); ndr = warp_nmax_all(ndr);
    __syncthreads();
    if (lane == 0) { red[warp] = np_; red[16 + warp] = nd_; red[32 + warp] = npr; red[48 + warp] = ndr; }
    __syncthreads();
    for (int w = 0; w < nw; ++w) {
      n.primal = nmax(n.primal, red[w]); n.dual = nmax(n.dual, red[16 + w]);
      n.primal_rel = nmax(n.primal_rel, red[32 + w]); n.dual_rel = nmax(n.dual_rel, red[48 + w]);
    }
  }
  __syncthreads();
}

/* ------------------------------------------------------------------ the kernel */
struct AdmmKPar {
  mimgpu_csqp_params q;
  int with_kkt, ric_total, nchain, stage_g, chain_ns;
  double *acl, *rv, *tv; /* [B][T][...] scratch of the condensed iterations */
  long long* timing;            /* debug (MIMGPU_ADMM_TIMING): clock64 cycles of CTA 0 per phase, or null */
};

/* shapes whose knot blocks go through a per-warp TMA stage in the knot-parallel phases (the
 * larger ones read global memory directly: their blocks do not fit one stage per warp) */
template <int NX, int NU>
__host__ __device__ constexpr bool admm_staged() { return NX <= 36; }

/* MAXT: CTA size; MINB: resident CTAs per SM the register allocation aims at; RBUF: F / G buffers of
 * the DMMA Riccati sweep (2 = next knot streams in while this one is processed, 1 = smaller CTA) */
#ifndef MIMGPU_ADMM_C3_MINB
#define MIMGPU_ADMM_C3_MINB 2
#endif
#ifndef MIMGPU_ADMM_C3_RBUF
#define MIMGPU_ADMM_C3_RBUF 2
#endif
#ifndef MIMGPU_ADMM_C3_MAXT
#define MIMGPU_ADMM_C3_MAXT 192
#endif
#ifndef MIMGPU_ADMM_C5_MAXT   /* the compact-layout shapes: 64/32, 76/32 */
#define MIMGPU_ADMM_C5_MAXT 384
#endif
#ifndef MIMGPU_ADMM_N32_MAXT
#define MIMGPU_ADMM_N32_MAXT 192
#endif
template <int NX, int NU> struct AdmmTune { static constexpr int MAXT = NX <= 16 ? 128 : 256, MINB = NX <= 16 ? 6 : 2, RBUF = 2; };
template <> struct AdmmTune<36, 12> { static constexpr int MAXT = MIMGPU_ADMM_C3_MAXT, MINB = MIMGPU_ADMM_C3_MINB, RBUF = MIMGPU_ADMM_C3_RBUF; };
template <> struct AdmmTune<12, 6> { static constexpr int MAXT = 96, MINB = 8, RBUF = 2; };
template <> struct AdmmTune<32, 16> { static constexpr int MAXT = MIMGPU_ADMM_N32_MAXT, MINB = 2, RBUF = 2; };
template <> struct AdmmTune<64, 32> { static constexpr int MAXT = MIMGPU_ADMM_C5_MAXT, MINB = 1, RBUF = 2; };
template <> struct AdmmTune<76, 32> { static constexpr int MAXT = MIMGPU_ADMM_C5_MAXT, MINB = 1, RBUF = 2; };
template <> struct AdmmTune<48, 24> { static constexpr int MAXT = 256, MINB = 1, RBUF = 2; };

template <int MODE>
static __device__ __noinline__ int admm_riccati_generic(const Prob& p, double* sm, const Smem& s, const RicPar& rp,
                                                        int stage_fx) {
  return riccati_backward<MODE>(p, as_shared(sm), s, rp, stage_fx);
}
static __device__ __noinline__ void admm_forward_eq(const Prob& p, double* sm, const Smem& s) { csqp_forward_eq(p, as_shared(sm), s); }
static __device__ __noinline__ void admm_kkt(const Prob& p, double* sm, const Smem& s) { csqp_kkt(p, as_shared(sm), s); }

/* SolverCSQP::computeDirection (src/csqp.cpp:454-517): control flow of csqp_direction_body
 * (kernels_generic.cuh); USE_MMA selects the DMMA Riccati sweep for the full iterations. */
/* RIC: full Riccati sweeps by 0 = the shape-generic code, 1 = DMMA (double-buffered F / G), 2 = DMMA on the
 * compact shared-memory layout (large shapes: Talos 76/32, 64/32, 48/24) */
template <int NX, int NU, int RIC>
__global__ void __launch_bounds__(AdmmTune<NX, NU>::MAXT, AdmmTune<NX, NU>::MINB)
k_csqp_direction_admm(KArgs a, AdmmKPar kp) {
  constexpr bool USE_MMA = RIC != 0;
  static_assert(NX % 2 == 0 && NU % 2 == 0 && NU <= 32, "16-byte rows, one lane per control in the gain solve");
  extern __shared__ __align__(16) double sm_admm[];
  if ((int)blockIdx.x >= a.n_active) return;
  const int b = a.active ? a.active[blockIdx.x] : (int)blockIdx.x + a.first;
  Prob p = bind(a, b);
  const int nw = blockDim.x >> 5;
  constexpr bool STAGED = admm_staged<NX, NU>();
  const AdmmSmem ss = admm_layout<NX, NU>(p.ng, p.ngT, nw, kp.ric_total,
                                          admm_stage_doubles<NX, NU>(p.ng, STAGED, kp.stage_g != 0));
  double* sm = sm_admm;
  double* gsm = sm + ss.uni; /* Riccati layouts live in the union region */
  const Smem gs = smem_layout(NX, NU, p.ng, p.ngT, a.stage_fx);
  RicSmem rs{};
  if constexpr (RIC == 1) rs = ric_layout<NX, NU>(p.ng, p.ngT, true, AdmmTune<NX, NU>::RBUF);
  if constexpr (RIC == 2) rs = ric_layout<NX, NU>(p.ng, p.ngT, true, 1, true);
  uint64_t* ric_bars = reinterpret_cast<uint64_t*>(sm + ss.bars);
  uint32_t ric_seq = 0u;
  if (threadIdx.x == 0) {
    mbar_init(ric_bars, 1);
    mbar_init(ric_bars + 1, 1);
    for (int w = 0; w < 3 * nw; ++w) mbar_init(reinterpret_cast<uint64_t*>(sm + ss.wbars) + w, 1);
    for (int w = 0; w < 16; ++w) mbar_init(reinterpret_cast<uint64_t*>(sm + ss.cbars) + w, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  uint32_t wseq = 0u; /* loads this warp has put through its TMA stage */
  ChainSeq cseq{0, 0u};  /* ring position of the chains' TMA pipeline */
  AdmmScr sc;
  {
    const size_t sb = (size_t)b * p.T;
    sc.acl = kp.acl + sb * NX * NX;
    sc.rv = kp.rv + sb * NX; sc.tv = kp.tv + sb * NX;
  }
  const mimgpu_csqp_params& q = kp.q;
  const size_t B = p.B;
  const int nc = p.T * p.ng + p.ngT;
  double rho_sparse = p.scal[MIMGPU_S_RHO_SPARSE * B + b];
  RicPar rp;
  rp.preg = a.preg_arr ? a.preg_arr[b] : q.preg;
  rp.dreg = a.dreg_arr ? a.dreg_arr[b] : q.dreg;
  rp.sigma = q.sigma;
  rp.is_feasible = 0;
  /* reset_params (src/csqp.cpp:170-193) */
  if (q.reset_rho) { rho_sparse = q.rho_sparse_base; csqp_apply_rho(p, rho_sparse, q.rho_min); }
  for (int e = threadIdx.x; e < nc; e += blockDim.x) { p.z[e] = 0.0; if (q.reset_y) p.y[e] = 0.0; }
  __syncthreads();
  int st = 0, converged = 0, it = 0;
  AdmmNorms n;
  n.primal = p.scal[MIMGPU_S_NORM_PRIMAL * B + b]; n.dual = p.scal[MIMGPU_S_NORM_DUAL * B + b];
  n.primal_rel = p.scal[MIMGPU_S_NORM_PRIMAL_REL * B + b]; n.dual_rel = p.scal[MIMGPU_S_NORM_DUAL_REL * B + b];
  if (q.equality_qp_initial_guess) { /* src/csqp.cpp:461-464 */
    if constexpr (USE_MMA) st = ric_backward_mma<MODE_FULL, NX, NU, RIC == 2>(p, gsm, rs, rp, ric_bars, ric_seq, true);
    else st = admm_riccati_generic<MODE_EQ>(p, gsm, gs, rp, a.stage_fx);
    if (!st) { __syncthreads(); admm_forward_eq(p, gsm, gs); }
  }
  const int interval = q.rho_update_interval;
  auto is_full = [&](int k) { return k % interval == 1 || interval == 1; };
  long long tk0 = 0;
  const bool timed = kp.timing != nullptr && blockIdx.x == 0 && threadIdx.x == 0;
  auto tick = [&](int slot) { /* phase timer of CTA 0 (debug switch; one predicated branch per phase otherwise) */
    if (timed) { const long long c = clock64(); kp.timing[slot] += c - tk0; tk0 = c; }
  };
  if (timed) tk0 = clock64();
  if (!st) {
    for (it = 1; it < q.max_qp_iters + 1; ++it) { /* src/csqp.cpp:470 */
      if (is_full(it)) {
        if constexpr (USE_MMA) st = ric_backward_mma<MODE_FULL, NX, NU, RIC == 2>(p, gsm, rs, rp, ric_bars, ric_seq);
        else st = admm_riccati_generic<MODE_FULL>(p, gsm, gs, rp, a.stage_fx);
        if (st) break;
        __syncthreads();
        tick(0);
        admm_post_full<NX, NU>(p, sc, sm, ss);
        __syncthreads();
        tick(1);
      } else {
        admm_chain<NX, true>(p, sc, sm, ss, kp.nchain, kp.chain_ns, cseq);
        __syncthreads();
        tick(2);
        admm_gains<NX, NU, STAGED>(p, sc, sm, ss, wseq);
        __syncthreads();
        tick(3);
      }
      admm_chain<NX, false>(p, sc, sm, ss, kp.nchain, kp.chain_ns, cseq);
      __syncthreads();
      tick(4);
      const bool need_r = (it + 1 < q.max_qp_iters + 1) && !is_full(it + 1);
      admm_update<NX, NU, STAGED>(p, sc, sm, ss, q, it, need_r, kp.stage_g != 0, n, wseq);
      tick(5);
      if (it % interval == 0) {
        /* update_rho_vec (src/csqp.cpp:519-532) */
        const double scale = sqrt((n.primal * n.dual_rel) / (n.dual * n.primal_rel));
        const double rho_est = std_min(std_max(scale * rho_sparse, q.rho_min), q.rho_max);
        if (rho_est > rho_sparse * q.adaptive_rho_tolerance || rho_est < rho_sparse / q.adaptive_rho_tolerance) {
          rho_sparse = rho_est;
          csqp_apply_rho(p, rho_sparse, q.rho_min);
          __syncthreads();
        }
        double tol_p, tol_d; /* src/csqp.cpp:497-512 */
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
  if (!st && kp.with_kkt) admm_kkt(p, gsm, gs);
}

/* ------------------------------------------------------------------- dispatch */

template <int NX, int NU, int MMA>   /* MMA: 0 none, 1 standard DMMA layout, 2 compact DMMA layout */
static AdmmPlan admm_plan_t(const mimgpu_shape& sh, int stage_fx, int max_smem) {
  AdmmPlan pl{};
  const Smem gs = smem_layout(NX, NU, sh.ng, sh.ngT, stage_fx);
  pl.nt = AdmmTune<NX, NU>::MAXT;
  pl.nchain = ((4 * NX + 31) / 32) * 32;
  if (pl.nchain + 32 > pl.nt) return pl; /* compute warps + one producer warp */
  for (int pass = (MMA ? 0 : 1); pass < 2; ++pass) {
    const int mma = pass == 0 ? MMA : 0;
    int ric_total = gs.total;
    if constexpr (MMA == 1) {
      if (mma) { const RicSmem rs = ric_layout<NX, NU>(sh.ng, sh.ngT, true, AdmmTune<NX, NU>::RBUF); if (rs.total > ric_total) ric_total = rs.total; }
    }
    if constexpr (MMA == 2) {
      if (mma) {
        const RicSmem rs = ric_layout<NX, NU>(sh.ng, sh.ngT, true, 1, true);
        if (rs.total > ric_total) ric_total = rs.total;   /* (csqp_kkt / forward_eq still use the generic layout's vectors) */
        if ((NX + NU + 7) / 8 > 2 * (pl.nt / 32)) continue;    /* two tile rows per warp at most */
      }
    }
    /* stage Gx / Gu next to K only while MINB CTAs of this size still share an SM */
    constexpr bool STAGED = admm_staged<NX, NU>();
    int stage_g = STAGED ? 1 : 0;
    AdmmSmem ss = admm_layout<NX, NU>(sh.ng, sh.ngT, pl.nt / 32, ric_total, admm_stage_doubles<NX, NU>(sh.ng, STAGED, true));
    if (stage_g && AdmmTune<NX, NU>::MINB * (ss.total * 8 + 1024) > 233472) {
      stage_g = 0;
      ss = admm_layout<NX, NU>(sh.ng, sh.ngT, pl.nt / 32, ric_total, admm_stage_doubles<NX, NU>(sh.ng, STAGED, false));
    }
    /* TMA ring of the chains: as many [Acl | add0 | add1] stages as the union region holds */
    auto ring_ns = [&](const AdmmSmem& L) { int n = (L.total - L.ws0) / admm_chain_stage<NX>(); return n > 8 ? 8 : n; };
    if (ring_ns(ss) < 3) { /* grow the union region to three stages */
      const int need = 2 * NX + 3 * admm_chain_stage<NX>();
      if (need > ric_total) ric_total = need;
      ss = admm_layout<NX, NU>(sh.ng, sh.ngT, pl.nt / 32, ric_total, admm_stage_doubles<NX, NU>(sh.ng, STAGED, stage_g != 0));
    }
    const int bytes = ss.total * 8;
    if (bytes > max_smem) continue;
    pl.stage_g = stage_g;
    pl.chain_ns = ring_ns(ss);
    cudaError_t e;
    if (mma == 1) e = cudaFuncSetAttribute(k_csqp_direction_admm<NX, NU, (MMA == 1 ? 1 : 0)>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
    else if (mma == 2) e = cudaFuncSetAttribute(k_csqp_direction_admm<NX, NU, (MMA == 2 ? 2 : 0)>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
    else e = cudaFuncSetAttribute(k_csqp_direction_admm<NX, NU, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
    if (e != cudaSuccess) { cudaGetLastError(); continue; }
    pl.ok = 1; pl.use_mma = mma; pl.ric_total = ric_total; pl.smem_bytes = bytes;
    return pl;
  }
  return pl;
}

/* shapes with the standard DMMA Riccati sweep | large shapes on the compact DMMA layout | shape-generic sweeps */
#define MIMGPU_ADMM_MMA_SHAPES(X) X(12, 6) X(36, 12) X(16, 8) X(8, 4) X(32, 16) X(4, 2)
#define MIMGPU_ADMM_CMP_SHAPES(X) X(64, 32) X(76, 32) X(48, 24) X(24, 12)
#define MIMGPU_ADMM_GEN_SHAPES(X) X(6, 2) X(10, 4)

static inline AdmmPlan admm_plan(const mimgpu_shape& sh, int stage_fx, int max_smem) {
#define X(NX_, NU_) \
  if (sh.nx == NX_ && sh.nu == NU_) return admm_plan_t<NX_, NU_, 1>(sh, stage_fx, max_smem);
  MIMGPU_ADMM_MMA_SHAPES(X)
#undef X
#define X(NX_, NU_) \
  if (sh.nx == NX_ && sh.nu == NU_) return admm_plan_t<NX_, NU_, 2>(sh, stage_fx, max_smem);
  MIMGPU_ADMM_CMP_SHAPES(X)
#undef X
#define X(NX_, NU_) \
  if (sh.nx == NX_ && sh.nu == NU_) return admm_plan_t<NX_, NU_, 0>(sh, stage_fx, max_smem);
  MIMGPU_ADMM_GEN_SHAPES(X)
#undef X
  return AdmmPlan{};
}

static inline int admm_csqp_launch(const KArgs& a, const mimgpu_csqp_params& q, int with_kkt, const AdmmPlan& pl,
                                   double* const scratch[3], cudaStream_t st, long long* timing = nullptr) {
  AdmmKPar kp;
  kp.timing = timing;
  kp.q = q; kp.with_kkt = with_kkt; kp.ric_total = pl.ric_total; kp.nchain = pl.nchain; kp.stage_g = pl.stage_g; kp.chain_ns = pl.chain_ns;
  kp.acl = scratch[0]; kp.rv = scratch[1]; kp.tv = scratch[2];
#define X(NX_, NU_)                                                                                      \
  if (a.sh.nx == NX_ && a.sh.nu == NU_) {                                                                \
    if (pl.use_mma) k_csqp_direction_admm<NX_, NU_, 1><<<a.n_active, pl.nt, pl.smem_bytes, st>>>(a, kp); \
    else k_csqp_direction_admm<NX_, NU_, 0><<<a.n_active, pl.nt, pl.smem_bytes, st>>>(a, kp);            \
    return 0;                                                                                            \
  }
  MIMGPU_ADMM_MMA_SHAPES(X)
#undef X
#define X(NX_, NU_)                                                                                      \
  if (a.sh.nx == NX_ && a.sh.nu == NU_) {                                                                \
    if (pl.use_mma) k_csqp_direction_admm<NX_, NU_, 2><<<a.n_active, pl.nt, pl.smem_bytes, st>>>(a, kp); \
    else k_csqp_direction_admm<NX_, NU_, 0><<<a.n_active, pl.nt, pl.smem_bytes, st>>>(a, kp);            \
    return 0;                                                                                            \
  }
  MIMGPU_ADMM_CMP_SHAPES(X)
#undef X
#define X(NX_, NU_)                                                                                      \
  if (a.sh.nx == NX_ && a.sh.nu == NU_) {                                                                \
    k_csqp_direction_admm<NX_, NU_, 0><<<a.n_active, pl.nt, pl.smem_bytes, st>>>(a, kp);                 \
    return 0;                                                                                            \
  }
  MIMGPU_ADMM_GEN_SHAPES(X)
#undef X
  return -1;
}

}  // namespace mimgpu
