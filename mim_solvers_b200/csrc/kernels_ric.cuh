/*
 * kernels_ric.cuh -- Riccati backward sweep for medium shapes with the block products on the
 * FP64 tensor pipe (mma.sync.m8n8k4.f64, SASS DMMA); one CTA owns one problem.
 *
 * Restates, per knot (compile-time NX, NU with NX % 4 == 0, NU even; run-time ng):
 *   MODE_DDP  SolverDDP::backwardPass + computeGains              src/ddp.cpp:265-348, 401-429
 *   MODE_EQ   SolverCSQP::backwardPass_without_constraints         src/csqp.cpp:804-898
 *   MODE_FULL SolverCSQP::backwardPass                             src/csqp.cpp:671-802
 *
 * Data flow of one knot (N = NX + NU, F = [Fx Fu], all operands "contraction index contiguous"
 * so that every DMMA fragment load is a conflict-free 8-byte shared-memory access):
 *   stage   F_t (and [Gx Gu]_t) arrive by TMA bulk copy, issued one knot ahead (double buffer);
 *           the knot's vectors are prefetched through registers one knot ahead
 *   V       Vxx_fs = Vxx' fs, tmp = Vxx_fs + Vx'                       (split dot products)
 *   B       W^T = F^T Vxx'            A = F^T tile rows (kept in registers for step C)
 *   D       [Qx; Qu] = l - sigma [dx; du] + G^T d + F^T tmp           (split dot products)
 *   C       Q = H + F^T W + G^T diag(rho) G   upper tile triangle; accumulators start from H
 *           read straight from global memory; Qxx goes to the Vxx buffer, Qxu / Quu to theirs
 *   E       LLT(Quu) by one warp (lane = row, right-looking, rsqrt pivots)
 *   K       K = Quu^-1 Qxu^T, k = Quu^-1 Qu   one right-hand side per thread, in registers
 *   F       M = Qxx - Qxu K on the upper tiles (DMMA), Vx = Qx - K^T Qu
 *   S       Vxx = (M + M^T)/2 + preg, spill Vxx, Vx, K, k, LLT, Qu to HBM
 * The strictly-lower tiles of Qxx / M are never computed: both are symmetric up to rounding, the
 * symmetrisation takes the upper-tile value for them.
 */
#pragma once
#include <cuda_runtime.h>

#include "kernels_stream.cuh"

namespace mimgpu {

__device__ __forceinline__ void dmma(double& c0, double& c1, double a, double b) {
  asm("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
      : "+d"(c0), "+d"(c1)
      : "d"(a), "d"(b));
}

template <int NX, int NU>
struct RicCfg {
  static constexpr int N = NX + NU, XT = (NX + 7) / 8, NT = (N + 7) / 8, KT = NX / 4, UT = (NU + 3) / 4;
  static constexpr int VPT = 3;
  static_assert(NX % 4 == 0 && NU % 2 == 0, "DMMA k-steps of 4, 16-byte TMA blocks");
};

struct RicSmem {
  int V, F0, F1, G0, G1, W, Qxu, Quu, L, K, Vx, tmp, Qv, kff, fs, d, rho, linv, col, vec, red, total;
};
/* compact = true (large shapes, FULL / EQ sweeps only): one F / G buffer, the Cholesky factor overwrites Quu in
 * place, and Qxu lives in the G buffer (it is written after every warp is done reading G: ric_backward_mma
 * <.., COMPACT = true> defers those stores) -- Talos (76, 32, ng = 32) then fits the 227 KB of one SM. */
template <int NX, int NU>
__host__ __device__ inline RicSmem ric_layout(int ng, int ngT, bool with_g, int nbuf = 2, bool compact = false) {
  RicSmem s;
  constexpr int N = NX + NU;
  const int ngm = ng > ngT ? ng : ngT;
  int o = 0;
#define RC_TAKE(n) (o += (((n) + 1) & ~1), o - (((n) + 1) & ~1))
  s.V = RC_TAKE(NX * NX);
  s.F0 = RC_TAKE(NX * N); s.F1 = (nbuf == 2) ? RC_TAKE(NX * N) : s.F0;
  s.G0 = RC_TAKE(with_g ? ng * N : 0); s.G1 = (nbuf == 2) ? RC_TAKE(with_g ? ng * N : 0) : s.G0;
  s.W = RC_TAKE(NX * N);
  s.K = s.W; /* K (nu x nx) is built after the last read of W and consumed before the next write */
  if (compact) {
    const int gsz = with_g ? ng * N : 0;
    if (gsz >= NX * NU) s.Qxu = s.G0; else s.Qxu = RC_TAKE(NX * NU);
    s.Quu = RC_TAKE(NU * NU); s.L = s.Quu;
  } else {
    s.Qxu = RC_TAKE(NX * NU); s.Quu = RC_TAKE(NU * NU); s.L = RC_TAKE(NU * NU);
  }
  s.Vx = RC_TAKE(NX); s.tmp = RC_TAKE(NX); s.Qv = RC_TAKE(N); s.kff = RC_TAKE(NU); s.fs = RC_TAKE(NX);
  s.d = RC_TAKE(ngm); s.rho = RC_TAKE(ngm); s.linv = RC_TAKE(NU); s.col = RC_TAKE(2 * NU);
  s.vec = RC_TAKE(3 * NX + 2 * NU + 3 * ngm);
  s.red = RC_TAKE(64);
  /* slack: fragment loads of padded tile rows / columns read (and discard) up to 8 rows past an array */
  o += 8 * NX;
#undef RC_TAKE
  s.total = o;
  return s;
}

/* one warp: Cholesky of the NU x NU matrix in sQuu (full symmetric, col-major) -> lower factor
 * in sL (upper part zero), reciprocal pivots in linv.  Eigen::LLT semantics: fails iff a pivot
 * is <= 0 (src/ddp.cpp:413-419).  Returns 1 on failure (uniform over the warp). */
/* Pivot K of the right-looking factorisation as a template recursion: every index into the per-lane row `a` is a
 * compile-time constant, so the row stays in registers.  (Written as a `#pragma unroll` loop over k, nvcc kept the
 * loop rolled for NU = 32 and put `a` into local memory: 58 k cycles per knot in the Talos kernel, 39 % of its
 * backward sweep -- profiles/ric_step_timing_r02_v5.txt.)  Same operations in the same order as before. */
template <int K, int NU, int P>
struct LltStep {
  static __device__ __forceinline__ void run(double (&a)[P][NU], double* linv, double* col, int lane, int& fail) {
    /* pivot lives in lane K % 32, register row K / 32 */
    const double piv = __shfl_sync(0xffffffffu, a[K / 32][K], K % 32);
    if (piv <= 0.0) fail = 1;
    const double r = rsqrt(piv);
    double* cb = col + (K & 1) * NU;
#pragma unroll
    for (int rr = 0; rr < P; ++rr) {
      const int i = lane + 32 * rr;
      const double lik = (i == K) ? piv * r : a[rr][K] * r;
      a[rr][K] = lik;
      if (i < NU) cb[i] = lik;
    }
    if (lane == 0) linv[K] = r;
    __syncwarp();
#pragma unroll
    for (int j = K + 1; j < NU; ++j) {
      const double ljk = cb[j];
#pragma unroll
      for (int rr = 0; rr < P; ++rr) a[rr][j] = fma(-a[rr][K], ljk, a[rr][j]);
    }
    LltStep<K + 1, NU, P>::run(a, linv, col, lane, fail);
  }
};
template <int NU, int P>
struct LltStep<NU, NU, P> {
  static __device__ __forceinline__ void run(double (&)[P][NU], double*, double*, int, int&) {}
};

template <int NU>
__device__ __forceinline__ int warp_llt(const double* sQuu, double* sL, double* linv, double* col, int lane) {
  constexpr int P = (NU + 31) / 32;          /* rows per lane */
  double a[P][NU];
  int fail = 0;
#pragma unroll
  for (int r = 0; r < P; ++r) {
    const int i = lane + 32 * r;
#pragma unroll
    for (int j = 0; j < NU; ++j) a[r][j] = (i < NU) ? sQuu[i + NU * j] : 0.0;
  }
  LltStep<0, NU, P>::run(a, linv, col, lane, fail);
#pragma unroll
  for (int rr = 0; rr < P; ++rr) {
    const int i = lane + 32 * rr;
    if (i < NU) {
#pragma unroll
      for (int j = 0; j < NU; ++j) sL[i + NU * j] = (j <= i) ? a[rr][j] : 0.0;
    }
  }
  return fail;
}

/* x := Quu^-1 x by forward / backward substitution, column oriented (every x[i] subtracts its
 * terms in the reference's order), reciprocal pivots. */
template <int NU>
__device__ __forceinline__ void llt_solve_regs(const double* sL, const double* linv, double (&x)[NU]) {
#pragma unroll
  for (int j = 0; j < NU; ++j) {
    x[j] *= linv[j];
#pragma unroll
    for (int i = j + 1; i < NU; ++i) x[i] = fma(-sL[i + NU * j], x[j], x[i]);
  }
#pragma unroll
  for (int j = NU - 1; j >= 0; --j) {
    x[j] *= linv[j];
#pragma unroll
    for (int i = 0; i < j; ++i) x[i] = fma(-sL[j + NU * i], x[j], x[i]);
  }
}

/* Per-step clock64 timer of CTA 0 (build-time debug switch MIMGPU_RIC_TIMING, scripts/build_variant.sh): cycles of
 * thread 0 between the step barriers of ric_backward_mma, printed by the calling kernel. */
#ifdef MIMGPU_RIC_TIMING
static __device__ long long g_ric_t[16];
#define RIC_T0() long long rt0_ = clock64()
#define RIC_T(slot)                                                     \
  do {                                                                  \
    if (blockIdx.x == 0 && threadIdx.x == 0) {                          \
      const long long c_ = clock64();                                   \
      g_ric_t[slot] += c_ - rt0_;                                       \
      rt0_ = c_;                                                        \
    }                                                                   \
  } while (0)
#else
#define RIC_T0() do {} while (0)
#define RIC_T(slot) do {} while (0)
#endif

/* Elimination steps of the gain solve (step K of ric_backward_mma) as template recursions, for the same reason as
 * LltStep: the per-lane slice x[RL] of a right-hand side must stay in registers.  QS lanes share one right-hand side,
 * lane q owns rows i = q (mod QS). */
template <int J, int NU, int QS, int RL>
struct GainSolve {
  static __device__ __forceinline__ void forward(double (&x)[RL], const double* sL, const double* linv, int lane, int qs) {
    if constexpr (J < NU) {
      const double xj = __shfl_sync(0xffffffffu, x[J / QS] * linv[J], (lane & ~(QS - 1)) | (J % QS));
      if (qs == J % QS) x[J / QS] = xj;
#pragma unroll
      for (int r = J / QS; r < RL; ++r) {
        const int i = r * QS + qs;
        if (i > J && i < NU) x[r] = fma(-sL[i + NU * J], xj, x[r]);
      }
      GainSolve<J + 1, NU, QS, RL>::forward(x, sL, linv, lane, qs);
    }
  }
  static __device__ __forceinline__ void backward(double (&x)[RL], const double* sL, const double* linv, int lane, int qs) {
    if constexpr (J >= 0) {
      const double xj = __shfl_sync(0xffffffffu, x[J / QS] * linv[J], (lane & ~(QS - 1)) | (J % QS));
      if (qs == J % QS) x[J / QS] = xj;
#pragma unroll
      for (int r = 0; r <= J / QS; ++r) {
        const int i = r * QS + qs;
        if (i < J) x[r] = fma(-sL[J + NU * i], xj, x[r]);
      }
      GainSolve<J - 1, NU, QS, RL>::backward(x, sL, linv, lane, qs);
    }
  }
};

/* ------------------------------------------------------------ the backward sweep */
/* Returns 0 or the status word (uniform over the CTA).  `sm` is the base of a RicSmem
 * carve-up; `bars` are two initialised mbarriers (count 1) outside it; `seq` counts the TMA
 * loads this kernel has consumed on them so far (load n uses barrier and buffer n & 1 with
 * phase parity (n >> 1) & 1; the count carries over between sweeps of the same kernel). */
template <int MODE, int NX, int NU, bool COMPACT = false>
static __device__ __noinline__ int ric_backward_mma(const Prob& p, double* sm, const RicSmem& s, const RicPar& rp,
                                             uint64_t* bars, uint32_t& seq, bool eq_run = false) {
  static_assert(!COMPACT || MODE != MODE_DDP, "the compact layout overwrites Quu with its Cholesky factor");
  using C = RicCfg<NX, NU>;
  sm = as_shared(sm); /* keep LDS / STS addressing inside this out-of-line function */
  constexpr int N = C::N, XT = C::XT, NT = C::NT, KT = C::KT, UT = C::UT;
  constexpr bool FULL = (MODE == MODE_FULL), DDP = (MODE == MODE_DDP);
  /* eq_run: MODE_FULL code executing backwardPass_without_constraints (no constraint rows, no
   * sigma, no Vxx_fs spill) -- same arithmetic as MODE_EQ, one instantiation less */
  const int T = p.T, ng = (FULL && !eq_run) ? p.ng : 0, ngT = (FULL && !eq_run) ? p.ngT : 0;
  const int tid = threadIdx.x, nth = blockDim.x, lane = tid & 31, warp = tid >> 5, nw = nth >> 5;
  const int g = lane >> 2, tq = lane & 3;
  const double preg = rp.preg, dreg = rp.dreg, sigma = (FULL && !eq_run) ? rp.sigma : 0.0;
  double *V = sm + s.V, *W = sm + s.W, *sQxu = sm + s.Qxu, *sQuu = sm + s.Quu, *sL = sm + s.L, *sK = sm + s.K;
  double *sVx = sm + s.Vx, *tmp = sm + s.tmp, *sQv = sm + s.Qv, *skff = sm + s.kff, *sfs = sm + s.fs;
  double *sd = sm + s.d, *srho = sm + s.rho, *linv = sm + s.linv, *vec = sm + s.vec;
  const bool dbl = (s.F1 != s.F0);
  const uint32_t seq0 = seq; /* load number of knot T-1; knot t is load seq0 + (T-1-t) */
  auto issue = [&](int t) { /* thread 0: TMA of knot t's matrices into buffer (load number) & 1 */
    const uint32_t n = seq0 + (uint32_t)(T - 1 - t);
    if (tid == 0) {
      double* Fd = sm + ((n & 1u) ? s.F1 : s.F0);
      double* Gd = sm + ((n & 1u) ? s.G1 : s.G0);
      uint64_t* bar = bars + (n & 1u);
      const uint32_t bytes = (uint32_t)(NX * N + ng * N) * 8u;
      mbar_expect_tx(bar, bytes);
      tma_bulk_g2s(Fd, p.Fx + (size_t)t * NX * NX, NX * NX * 8, bar);
      tma_bulk_g2s(Fd + NX * NX, p.Fu + (size_t)t * NX * NU, NX * NU * 8, bar);
      if (ng) {
        tma_bulk_g2s(Gd, p.Gx + (size_t)t * ng * NX, (uint32_t)ng * NX * 8, bar);
        tma_bulk_g2s(Gd + ng * NX, p.Gu + (size_t)t * ng * NU, (uint32_t)ng * NU * 8, bar);
      }
    }
  };
  auto drain = [&](int t) { /* early exit inside knot t: the load of knot t-1 may be in flight */
    const uint32_t n = seq0 + (uint32_t)(T - 1 - t);
    seq = n + 1;
    if (t > 0 && dbl) {
      mbar_wait(bars + ((n + 1) & 1u), ((n + 1) >> 1) & 1u);
      seq = n + 2;
    }
    __syncthreads();
  };
  /* vector pack of knot t: [Lx NX][Lu NU][fs NX] (+ FULL: [dx NX][du NU][y ng][z ng][rho ng]) */
  const int oLx = 0, oLu = NX, oFs = NX + NU, oDx = 2 * NX + NU, oDu = 3 * NX + NU, oY = 3 * NX + 2 * NU, oZ = oY + ng,
            oR = oZ + ng, nv = FULL ? oR + ng : oDx;
  VecSrc vs[C::VPT];
  double vr[C::VPT];
#pragma unroll
  for (int j = 0; j < C::VPT; ++j) {
    const int e = tid + j * nth;
    vs[j].base = nullptr; vs[j].stride = 0;
    if (e < oLu) { vs[j].base = p.Lx + e; vs[j].stride = NX; }
    else if (e < oFs) { vs[j].base = p.Lu + (e - oLu); vs[j].stride = NU; }
    else if (e < oDx) { vs[j].base = p.fs + (DDP ? 0 : NX) + (e - oFs); vs[j].stride = NX; }
    else if (e < nv) {
      if (e < oDu) { vs[j].base = p.dx + (e - oDx); vs[j].stride = NX; }
      else if (e < oY) { vs[j].base = p.du + (e - oDu); vs[j].stride = NU; }
      else if (e < oZ) { vs[j].base = p.y + (e - oY); vs[j].stride = ng; }
      else if (e < oR) { vs[j].base = p.z + (e - oZ); vs[j].stride = ng; }
      else { vs[j].base = p.rho + (e - oR); vs[j].stride = ng; }
    }
  }
  __syncthreads(); /* the whole carve-up is free */
  issue(T - 1);
#pragma unroll
  for (int j = 0; j < C::VPT; ++j) vr[j] = vs[j].base ? vs[j].base[(size_t)(T - 1) * vs[j].stride] : 0.0;

  /* ---- terminal knot: src/ddp.cpp:268-277, src/csqp.cpp:827-832, :693-707 */
  if (FULL && ngT) {
    for (int r = tid; r < ngT; r += nth) {
      const double rho = p.rho[(size_t)T * p.ng + r];
      srho[r] = rho;
      double d = p.y[(size_t)T * p.ng + r];
      d -= rho * p.z[(size_t)T * p.ng + r];
      sd[r] = d;
    }
  }
  for (int i = tid; i < NX; i += nth) sfs[i] = p.fs[(size_t)T * NX + i];
  __syncthreads();
  for (int e = tid; e < NX * NX; e += nth) {
    const int i = e % NX, j = e / NX;
    double v = p.LxxT[e];
    if (FULL && !eq_run) {
      if (i == j) v += sigma;
      if (ngT) {
        double acc = 0.0;
        for (int r = 0; r < ngT; ++r) acc += p.GxT[r + ngT * i] * (srho[r] * p.GxT[r + ngT * j]);
        v += acc;
      }
    }
    if (i == j && !isnan(preg)) v += preg;
    V[e] = v;
    p.Vxx[(size_t)T * NX * NX + e] = v;
  }
  __syncthreads();
  for (int i = tid; i < NX; i += nth) {
    double v = p.LxT[i];
    if (DDP) {
      if (!rp.is_feasible) {
        double acc = 0.0;
        for (int j = 0; j < NX; ++j) acc += V[i + NX * j] * sfs[j];
        v += acc;
      }
    } else if (FULL && !eq_run) {
      v -= sigma * p.dx[(size_t)T * NX + i];
      if (ngT) {
        double acc = 0.0;
        for (int r = 0; r < ngT; ++r) acc += p.GxT[r + ngT * i] * sd[r];
        v += acc;
      }
    }
    sVx[i] = v;
    p.Vx[(size_t)T * NX + i] = v;
  }

  const int grp = tid >> 2, q4 = tid & 3, ngrp = nth >> 2;
  RIC_T0();
  for (int t = T - 1; t >= 0; --t) {
    const uint32_t nload = seq0 + (uint32_t)(T - 1 - t);
    __syncthreads(); /* S0: knot t+1 is complete (V, Vx final; every buffer of it consumed) */
    if (dbl && t > 0) issue(t - 1);
#pragma unroll
    for (int j = 0; j < C::VPT; ++j) {
      const int e = tid + j * nth;
      if (vs[j].base) vec[e] = vr[j];
      if (vs[j].base && t > 0) vr[j] = vs[j].base[(size_t)(t - 1) * vs[j].stride];
    }
    mbar_wait(bars + (nload & 1u), (nload >> 1) & 1u);
    __syncthreads(); /* S1: vector pack visible */
    RIC_T(0);
    const double* F = sm + ((nload & 1u) ? s.F1 : s.F0);
    const double* G = sm + ((nload & 1u) ? s.G1 : s.G0);
    const bool gx_on = FULL && ng && t > 0; /* x0 is fixed: src/csqp.cpp:734,745,777 */
    const int gcol0 = (t > 0) ? 0 : NX;     /* first column of [Gx Gu] that takes part */

    /* ---- step V: Vxx_fs, tmp, d */
    if (DDP) {
      for (int i = tid; i < NX; i += nth) tmp[i] = sVx[i];
    } else {
      for (int o0 = 0; o0 < NX; o0 += ngrp) {
        const int i = o0 + grp;
        const int ic = i < NX ? i : 0;
        double acc = 0.0;
#pragma unroll
        for (int m = 0; m < KT; ++m) acc = fma(V[ic + NX * (4 * m + q4)], vec[oFs + 4 * m + q4], acc);
        acc = split_reduce<4>(acc);
        if (i < NX && q4 == 0) {
          if (FULL && !eq_run) p.Vxx_fs[(size_t)t * NX + i] = acc;
          tmp[i] = acc + sVx[i];
        }
      }
    }
    if (FULL) {
      for (int r = tid; r < ng; r += nth) {
        const double rho = vec[oR + r];
        srho[r] = rho;
        double d = vec[oY + r];
        d -= rho * vec[oZ + r];
        sd[r] = d;
      }
    }
    /* ---- step B: W^T = F^T Vxx' for this warp's row tiles; A fragments stay for step C */
    for (int a = warp; a < NT; a += nw) {
      double fa[KT];
#pragma unroll
      for (int kt = 0; kt < KT; ++kt) fa[kt] = F[(8 * a + g) * NX + 4 * kt + tq];
      double acc[XT][2];
#pragma unroll
      for (int it = 0; it < XT; ++it) { acc[it][0] = 0.0; acc[it][1] = 0.0; }
#pragma unroll
      for (int kt = 0; kt < KT; ++kt) {
#pragma unroll
        for (int it = 0; it < XT; ++it) dmma(acc[it][0], acc[it][1], fa[kt], V[(8 * it + g) * NX + 4 * kt + tq]);
      }
      if (8 * a + g < N) {
#pragma unroll
        for (int it = 0; it < XT; ++it)
          if (8 * it + 2 * tq < NX)
            *reinterpret_cast<double2*>(W + (8 * a + g) * NX + 8 * it + 2 * tq) = make_double2(acc[it][0], acc[it][1]);
      }
    }
    __syncthreads(); /* S2: W, tmp, d complete */
    RIC_T(1);
    /* ---- step D: [Qx; Qu] (src/ddp.cpp:291-292,301-302; src/csqp.cpp:727-739,755-759) */
    for (int o0 = 0; o0 < N; o0 += ngrp) {
      const int n = o0 + grp;
      const bool valid = n < N;
      const int nc = valid ? n : 0;
      double acc = 0.0;
#pragma unroll
      for (int m = 0; m < KT; ++m) acc = fma(F[nc * NX + 4 * m + q4], tmp[4 * m + q4], acc);
      double accg = 0.0;
      const bool gon = FULL && ng && nc >= gcol0;
      if (gon)
        for (int r = q4; r < ng; r += 4) accg = fma(G[r + ng * nc], sd[r], accg);
      acc = split_reduce<4>(acc);
      accg = split_reduce<4>(accg);
      if (valid && q4 == 0) {
        double v = vec[n]; /* [Lx; Lu] */
        if (FULL && !eq_run) v -= sigma * vec[oDx + n]; /* [dx; du] are adjacent in the pack */
        if (gon) v += accg;
        v += acc;
        sQv[n] = v;
        if (n >= NX) p.Qu[(size_t)t * NU + n - NX] = v;
      }
    }
    /* ---- step C: Q = H + F^T W (+ G^T rho G), upper tile triangle.  COMPACT: the Qxu block shares its buffer
     * with G, so its tiles wait in registers (qd) until every warp is done reading G. */
    constexpr int QT0 = NX / 8, NQ = NT - QT0;   /* column tiles that can hold Qxu entries */
    auto tile_row = [&](int a, double (*qd)[2]) {
      double fa[KT];
#pragma unroll
      for (int kt = 0; kt < KT; ++kt) fa[kt] = F[(8 * a + g) * NX + 4 * kt + tq];
      double acc[NT][2];
      const int r = 8 * a + g;
#pragma unroll
      for (int nt = 0; nt < NT; ++nt) {
        acc[nt][0] = 0.0; acc[nt][1] = 0.0;
        if (nt >= a) {
          const int c = 8 * nt + 2 * tq;
          if (r < NX) {
            if (c < NX) {
              acc[nt][0] = p.Lxx[(size_t)t * NX * NX + r + NX * c];
              acc[nt][1] = p.Lxx[(size_t)t * NX * NX + r + NX * (c + 1)];
            } else if (c < N) {
              acc[nt][0] = p.Lxu[(size_t)t * NX * NU + r + NX * (c - NX)];
              acc[nt][1] = p.Lxu[(size_t)t * NX * NU + r + NX * (c + 1 - NX)];
            }
          } else if (r < N && c >= NX && c < N) {
            acc[nt][0] = p.Luu[(size_t)t * NU * NU + (r - NX) + NU * (c - NX)];
            acc[nt][1] = p.Luu[(size_t)t * NU * NU + (r - NX) + NU * (c + 1 - NX)];
          }
        }
      }
#pragma unroll
      for (int kt = 0; kt < KT; ++kt) {
#pragma unroll
        for (int nt = 0; nt < NT; ++nt)
          if (nt >= a) dmma(acc[nt][0], acc[nt][1], fa[kt], W[(8 * nt + g) * NX + 4 * kt + tq]);
      }
      if (FULL && ng) {
        for (int r0 = 0; r0 < ng; r0 += 4) {
          const int rr = r0 + tq;
          const bool rok = rr < ng;
          const int ca = 8 * a + g;
          const double ga = (rok && ca < N && ca >= gcol0) ? G[rr + ng * ca] * srho[rr] : 0.0;
#pragma unroll
          for (int nt = 0; nt < NT; ++nt)
            if (nt >= a) {
              const int cb = 8 * nt + g;
              const double gb = (rok && cb < N && cb >= gcol0) ? G[rr + ng * cb] : 0.0;
              dmma(acc[nt][0], acc[nt][1], ga, gb);
            }
        }
      }
#pragma unroll
      for (int nt = 0; nt < NT; ++nt)
        if (nt >= a) {
          const int c = 8 * nt + 2 * tq;
          double v0 = acc[nt][0], v1 = acc[nt][1];
          if (FULL && !eq_run) { if (r == c) v0 += sigma; if (r == c + 1) v1 += sigma; }
          if (r >= NX && !isnan(dreg)) { if (r == c) v0 += dreg; if (r == c + 1) v1 += dreg; }
          if (r < NX) {
            if (c < NX) {
              *reinterpret_cast<double2*>(V + r * NX + c) = make_double2(v0, v1);
            } else if (c < N) {
              if constexpr (COMPACT) {
                if (nt >= QT0) { qd[nt - QT0][0] = v0; qd[nt - QT0][1] = v1; }
              } else {
                sQxu[r + NX * (c - NX)] = v0;
                sQxu[r + NX * (c + 1 - NX)] = v1;
              }
            }
          } else if (r < N && c >= NX && c < N) {
            const int ru = r - NX, cu = c - NX;
            if (cu >= ru) { sQuu[ru + NU * cu] = v0; sQuu[cu + NU * ru] = v0; }
            if (cu + 1 >= ru) { sQuu[ru + NU * (cu + 1)] = v1; sQuu[cu + 1 + NU * ru] = v1; }
          }
        }
    };
    if constexpr (COMPACT) {
      double qd[2][NQ > 0 ? NQ : 1][2];   /* at most two tile rows per warp (host: NT <= 2 * warps) */
#pragma unroll
      for (int ai = 0; ai < 2; ++ai) {
        const int a = warp + ai * nw;
        if (a < NT) tile_row(a, qd[ai]);
      }
      __syncthreads(); /* every warp is done with G: Qxu may land in its buffer */
#pragma unroll
      for (int ai = 0; ai < 2; ++ai) {
        const int a = warp + ai * nw, r = 8 * a + g;
        if (a < NT && r < NX) {
#pragma unroll
          for (int nt = QT0; nt < NT; ++nt) {
            const int c = 8 * nt + 2 * tq;
            if (nt >= a && c >= NX && c < N) {
              sQxu[r + NX * (c - NX)] = qd[ai][nt - QT0][0];
              sQxu[r + NX * (c + 1 - NX)] = qd[ai][nt - QT0][1];
            }
          }
        }
      }
    } else {
      for (int a = warp; a < NT; a += nw) tile_row(a, nullptr);
    }
    __syncthreads(); /* S3: Q, [Qx; Qu] complete */
    RIC_T(2);
    if (!COMPACT && !dbl && t > 0) issue(t - 1); /* single buffer: F, G of knot t are consumed */
    /* ---- step E: computeGains (src/ddp.cpp:401-429) */
    int fail = 0;
    if (warp == 0) fail = warp_llt<NU>(sQuu, sL, linv, sm + s.col, lane);
    if (__syncthreads_or(fail)) { /* S4 */
      if (!COMPACT && !dbl && t > 0) { mbar_wait(bars + ((nload + 1) & 1u), ((nload + 1) >> 1) & 1u); seq = nload + 2; __syncthreads(); }
      else drain(t);
      return MIMGPU_STATUS_LLT | (t << 8);
    }
    RIC_T(3);
    /* K = Quu^-1 Qxu^T (nx right-hand sides) and k = Quu^-1 Qu: QS lanes share one right-hand
     * side, lane q owns rows i = q (mod QS); per elimination step the owner of row j scales it
     * and broadcasts it by shuffle, every lane updates its own rows (independent FMAs).  Each
     * x[i] still subtracts its terms in the reference's order (j ascending / descending). */
    {
      constexpr int QS = 4, RL = (NU + QS - 1) / QS; /* rows per lane */
      const int qs = tid & (QS - 1), ngq = nth / QS;
      for (int c0 = 0; c0 < NX + 1; c0 += ngq) {
        const int c = c0 + tid / QS;
        const bool valid = c < NX + 1;
        double x[RL];
#pragma unroll
        for (int r = 0; r < RL; ++r) {
          const int i = r * QS + qs;
          x[r] = (valid && i < NU) ? ((c < NX) ? sQxu[c + NX * i] : sQv[NX + i]) : 0.0;
        }
        GainSolve<0, NU, QS, RL>::forward(x, sL, linv, lane, qs);   /* L y = b */
        GainSolve<NU - 1, NU, QS, RL>::backward(x, sL, linv, lane, qs); /* L^T x = y */
        if (valid) {
#pragma unroll
          for (int r = 0; r < RL; ++r) {
            const int i = r * QS + qs;
            if (i < NU) {
              if (c < NX) sK[i * NX + c] = x[r];
              else { skff[i] = x[r]; p.kff[(size_t)t * NU + i] = x[r]; }
            }
          }
        }
      }
    }
    for (int e = tid; e < NU * NU; e += nth) p.L[(size_t)t * NU * NU + e] = sL[e];
    __syncthreads(); /* S5: K, k complete */
    RIC_T(4);
    for (int e = tid; e < NU * NX; e += nth) p.K[(size_t)t * NX * NU + e] = sK[e];
    /* ---- step F: M = Qxx - Qxu K on the upper tiles; Vx = Qx - K^T Qu; Quuk */
    {
      int idx = 0;
      for (int it = 0; it < XT; ++it)
        for (int jt = it; jt < XT; ++jt, ++idx) {
          if (idx % nw != warp) continue;
          const int i = 8 * it + g, j = 8 * jt + 2 * tq;
          double m0 = 0.0, m1 = 0.0;
          if (i < NX && j < NX) {
            const double2 m2 = *reinterpret_cast<const double2*>(V + i * NX + j);
            m0 = m2.x; m1 = m2.y;
          }
#pragma unroll
          for (int ut = 0; ut < UT; ++ut) {
            const int a4 = 4 * ut + tq;
            const bool ok = a4 < NU;
            const double av = ok ? -sQxu[i + NX * (ok ? a4 : 0)] : 0.0;
            const double bv = ok ? sK[(ok ? a4 : 0) * NX + 8 * jt + g] : 0.0;
            dmma(m0, m1, av, bv);
          }
          if (i < NX && j < NX) *reinterpret_cast<double2*>(V + i * NX + j) = make_double2(m0, m1);
        }
    }
    for (int i = tid; i < NX; i += nth) {
      double acc = 0.0;
#pragma unroll
      for (int a = 0; a < NU; ++a) acc = fma(sK[a * NX + i], sQv[NX + a], acc);
      sVx[i] = sQv[i] - acc;
    }
    if (DDP) {
      for (int a = tid; a < NU; a += nth) {
        double acc = 0.0;
#pragma unroll
        for (int c = 0; c < NU; ++c) acc = fma(sQuu[a + NU * c], skff[c], acc);
        p.Quuk[(size_t)t * NU + a] = acc;
      }
    }
    __syncthreads(); /* S6: M complete */
    RIC_T(5);
    if (COMPACT && t > 0) issue(t - 1); /* compact layout: Qxu (in the G buffer) has been consumed by steps K and F */
    /* ---- step S: Vxx = (M + M^T)/2 + preg  (src/ddp.cpp:328-333) */
    int bad = 0;
    for (int e = tid; e < NX * NX; e += nth) {
      const int i = e / NX, j = e - i * NX;
      if (i <= j) {
        double v = V[i * NX + j];
        if ((i >> 3) == (j >> 3)) v = 0.5 * (v + V[j * NX + i]);
        if (i == j && !isnan(preg)) v += preg;
        V[i * NX + j] = v;
        V[j * NX + i] = v;
        if (DDP) bad |= raise_if_nan(fabs(v));
      }
    }
    __syncthreads(); /* S7: Vxx complete */
    RIC_T(6);
    if (DDP && !rp.is_feasible) { /* Vx += Vxx fs_t  (src/ddp.cpp:336-338) */
      for (int o0 = 0; o0 < NX; o0 += ngrp) {
        const int i = o0 + grp;
        const int ic = i < NX ? i : 0;
        double acc = 0.0;
#pragma unroll
        for (int m = 0; m < KT; ++m) acc = fma(V[ic + NX * (4 * m + q4)], vec[oFs + 4 * m + q4], acc);
        acc = split_reduce<4>(acc);
        if (i < NX && q4 == 0) sVx[i] += acc;
      }
      __syncthreads();
    }
    for (int e = tid; e < NX * NX; e += nth) p.Vxx[(size_t)t * NX * NX + e] = V[e];
    for (int i = tid; i < NX; i += nth) {
      const double vx = sVx[i];
      p.Vx[(size_t)t * NX + i] = vx;
      if (DDP) bad |= raise_if_nan(fabs(vx));
    }
    if (DDP) {
      if (__syncthreads_or(bad)) { /* src/ddp.cpp:340-345 */
        if (!dbl && t > 0) { mbar_wait(bars + ((nload + 1) & 1u), ((nload + 1) >> 1) & 1u); seq = nload + 2; __syncthreads(); }
        else drain(t);
        return MIMGPU_STATUS_NONFINITE | (t << 8);
      }
    }
    RIC_T(7);
  }
  seq = seq0 + (uint32_t)T;
  __syncthreads();
  return 0;
}

}  // namespace mimgpu
