/*
 * kernels_small.cuh -- warp-per-problem SQP direction kernel for small shapes
 * (UR5 class: nx = 12, nu = 6; any even NX, NU with NX + 2 NU <= 32).
 *
 * Replaces, in ONE launch: SolverDDP::backwardPass + computeGains (src/ddp.cpp:265-348,
 * 401-429), SolverSQP::forwardPass (src/sqp.cpp:270-305), checkKKTConditions
 * (src/sqp.cpp:244-268) and the gap norm / merit of computeDirection (src/sqp.cpp:229-236).
 *
 * Mapping.  One warp owns one problem; all sizes are compile-time so every matrix lives
 * in registers as "one column per lane":
 *   lanes [0, NX)        column j of Fx   -> W = Vxx' Fx,  Qxx column j, Vxx column j, K column j
 *   lanes [NX, N)        column a of Fu   -> W = Vxx' Fu,  Quu column a
 *   lanes [N, N + NU)    column a of Fu   -> (duplicate W)  Qxu column a
 * The operand every lane needs in common (Vxx', [Fx Fu], Qxu) is read from shared memory
 * as a warp-wide broadcast (one wavefront per 16-byte load).  The 6x6 Cholesky and the
 * triangular solves are done redundantly per lane in registers (no cross-lane traffic).
 * Each knot's derivative blocks are staged global -> shared with cp.async one knot ahead
 * (double buffer), so the HBM latency of knot t-1 hides behind the arithmetic of knot t.
 */
#pragma once
#include <cuda_runtime.h>

#include "kernels_generic.cuh"
#include "kernels_stream.cuh" /* mbarrier / TMA bulk helpers */

namespace mimgpu {

template <int NX, int NU>
struct SmallCfg {
  static constexpr int N = NX + NU;
  static constexpr int LDQ = NX + 2;  /* leading dim of the shared Qxu (+ Qu row) block */
  /* backward stage: [Fx|Fu] [Lxx|Lxu] Luu [Lx|Lu] fs */
  static constexpr int bF = 0, bH = NX * N, bLuu = 2 * NX * N, bL = bLuu + NU * NU, bFs = bL + N,
                       STAGE_B = bFs + NX;
  /* forward stage: Vxx+ Vx+ K k [Fx|Fu] fs+ [Lx|Lu] */
  static constexpr int fV = 0, fVx = NX * NX, fK = fVx + NX, fk = fK + NX * NU, fF = fk + NU, fFs = fF + NX * N,
                       fL = fFs + NX, STAGE_F = fL + N;
  static constexpr int STAGE = (STAGE_B > STAGE_F ? STAGE_B : STAGE_F);
  /* per-warp shared memory (doubles) */
  /* DMMA tiling: NT tiles of 8 over the N = NX + NU columns of [Fx Fu], XT over the NX states */
  static constexpr int NT = (N + 7) / 8, XT = (NX + 7) / 8, KT = NX / 4, KPT = (NX + 1 + 7) / 8;
  static constexpr int PC0 = 8 * (NX / 8);          /* first published column of Q */
  static constexpr int LDP = 8 * NT - PC0 + 2;      /* leading dim of the published Q block */
  static constexpr int LDK = 24;                    /* leading dim of the shared K (NU x NX) */
  static constexpr int oS0 = 0, oS1 = STAGE, oV = 2 * STAGE, oT = oV + NX * NX, oVx = oT + NX * NX,
                       oK = oVx + NX, oX = oK + NU, oU = oX + NX, oXn = oU + NU, oLam = oXn + NX,
                       oUnion = oLam + NX + 2,
                       /* DFMA variant: Qxu (+ Qu row), Quu */
                       oQxu = oUnion, oQuu = oQxu + LDQ * NU + 32, endA = oQuu + NU * NU,
                       /* DMMA variant: published Q columns, [Qx; Qu], K */
                       oQp = oUnion, oQv = oQp + 8 * NT * LDP, oKm = oQv + 8 * NT, endB = oKm + NU * LDK + 2,
                       WARP = ((endA > endB ? endA : endB) + 1) & ~1;
  static_assert(NX % 2 == 0 && NU % 2 == 0, "16-byte staging needs even block sizes");
  static_assert(N + NU <= 32, "one column per lane");
};

__device__ __forceinline__ void cp_async16(double* smem_dst, const double* gsrc) {
  const unsigned s = (unsigned)__cvta_generic_to_shared(smem_dst);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(s), "l"(gsrc) : "memory");
}
/* warp-cooperative copy of NDBL doubles (even, 16-byte aligned on both sides) */
template <int NDBL>
__device__ __forceinline__ void warp_stage(double* dst, const double* src, int lane) {
  static_assert(NDBL % 2 == 0, "even");
#pragma unroll
  for (int c = 0; c < (NDBL / 2 + 31) / 32; ++c) {
    const int i = lane + 32 * c;
    if (i < NDBL / 2) cp_async16(dst + 2 * i, src + 2 * i);
  }
}
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ double warp_nmax(double v) {
#pragma unroll
  for (int o = 16; o; o >>= 1) v = nmax(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

/* ------------------------------------------------------------------ backward */
/* Returns the status word (uniform over the warp). */
template <int NX, int NU>
__device__ __forceinline__ int small_backward(const Prob& p, double* sw, int lane, double preg, double dreg,
                                              int is_feasible) {
  using C = SmallCfg<NX, NU>;
  constexpr int N = C::N, LDQ = C::LDQ;
  const int T = p.T;
  double* sV = sw + C::oV;
  double* sT = sw + C::oT;
  double* sQxu = sw + C::oQxu;
  double* sQuu = sw + C::oQuu;
  double* sVx = sw + C::oVx;
  double* sK = sw + C::oK;
  const bool gxx = lane < NX, guu = lane >= NX && lane < N, gxu = lane >= N && lane < N + NU;
  const int col = lane < N ? lane : (gxu ? lane - NU : 0);
  const int ua = guu ? lane - NX : (gxu ? lane - N : 0);
  const int rr = lane <= NX ? lane : 0;   /* rhs row of the gain solve */
  const int lx = gxx ? lane : 0;

  auto prefetch = [&](int t, double* st) {
    warp_stage<NX * NX>(st + C::bF, p.Fx + (size_t)t * NX * NX, lane);
    warp_stage<NX * NU>(st + C::bF + NX * NX, p.Fu + (size_t)t * NX * NU, lane);
    warp_stage<NX * NX>(st + C::bH, p.Lxx + (size_t)t * NX * NX, lane);
    warp_stage<NX * NU>(st + C::bH + NX * NX, p.Lxu + (size_t)t * NX * NU, lane);
    warp_stage<NU * NU>(st + C::bLuu, p.Luu + (size_t)t * NU * NU, lane);
    warp_stage<NX>(st + C::bL, p.Lx + (size_t)t * NX, lane);
    warp_stage<NU>(st + C::bL + NX, p.Lu + (size_t)t * NU, lane);
    warp_stage<NX>(st + C::bFs, p.fs + (size_t)t * NX, lane);
    cp_async_commit();
  };
  prefetch(T - 1, sw + C::oS0);

  /* ---- terminal knot (src/ddp.cpp:268-277).  sV holds Vxx_T TRANSPOSED so that the
   * first sweep uses Vxx_T[m,k] exactly as the reference does even if Lxx_T is not
   * bitwise symmetric; every later Vxx is symmetrised. */
  {
    double v[NX];
#pragma unroll
    for (int i = 0; i < NX; i += 2) {
      const double2 t2 = *reinterpret_cast<const double2*>(p.LxxT + NX * lx + i);
      v[i] = t2.x; v[i + 1] = t2.y;
    }
    if (!isnan(preg)) {
#pragma unroll
      for (int i = 0; i < NX; ++i) if (i == lx) v[i] += preg;
    }
    if (gxx) {
#pragma unroll
      for (int i = 0; i < NX; ++i) sV[lane + NX * i] = v[i];       /* sV[j + NX i] = Vxx_T[i, j] */
#pragma unroll
      for (int i = 0; i < NX; i += 2)
        *reinterpret_cast<double2*>(p.Vxx + (size_t)T * NX * NX + NX * lane + i) = make_double2(v[i], v[i + 1]);
    }
    __syncwarp();
    double vx = p.LxT[lx];
    if (!is_feasible) {
      double acc = 0.0;
#pragma unroll
      for (int j = 0; j < NX; ++j) acc = fma(sV[j + NX * lx], p.fs[(size_t)T * NX + j], acc); /* Vxx_T[lx, j] */
      vx += acc;
    }
    if (gxx) { sVx[lane] = vx; p.Vx[(size_t)T * NX + lane] = vx; }
    __syncwarp();
  }

  int s = 0;
  for (int t = T - 1; t >= 0; --t) {
    double* st = sw + (s ? C::oS1 : C::oS0);
    if (t > 0) {
      prefetch(t - 1, sw + (s ? C::oS0 : C::oS1));
      cp_async_wait<1>();
    } else {
      cp_async_wait<0>();
    }
    __syncwarp();

    /* ---- step 1: w = Vxx' F[:, col]  (= row `col` of FxTVxx / FuTVxx, src/ddp.cpp:289,299)
     *      and qv = [Lx; Lu][col] + F[:, col] . Vx'  (src/ddp.cpp:291-292, 301-302) */
    double f[NX], w[NX];
#pragma unroll
    for (int i = 0; i < NX; i += 2) {
      const double2 t2 = *reinterpret_cast<const double2*>(st + C::bF + NX * col + i);
      f[i] = t2.x; f[i + 1] = t2.y;
      w[i] = 0.0; w[i + 1] = 0.0;
    }
    double qv = 0.0;
#pragma unroll
    for (int m = 0; m < NX; ++m) {
#pragma unroll
      for (int k = 0; k < NX; k += 2) {
        const double2 v2 = *reinterpret_cast<const double2*>(sV + NX * m + k);
        w[k] = fma(v2.x, f[m], w[k]);
        w[k + 1] = fma(v2.y, f[m], w[k + 1]);
      }
      qv = fma(f[m], sVx[m], qv);
    }
    qv = st[C::bL + col] + qv;

    /* ---- step 2: Q[i, col] = H[i, col] + sum_k F[k, i] w[k]  (src/ddp.cpp:295-314) */
    double q[NX];
    {
      const double* fb = st + C::bF + (guu ? NX * NX : 0);
      const double* hp = guu ? st + C::bLuu + NU * ua : st + C::bH + NX * col;
#pragma unroll
      for (int r = 0; r < NX; r += 2) {
        double a0 = 0.0, a1 = 0.0;
#pragma unroll
        for (int k = 0; k < NX; k += 2) {
          const double2 f0 = *reinterpret_cast<const double2*>(fb + NX * r + k);
          const double2 f1 = *reinterpret_cast<const double2*>(fb + NX * (r + 1) + k);
          a0 = fma(f0.x, w[k], a0); a0 = fma(f0.y, w[k + 1], a0);
          a1 = fma(f1.x, w[k], a1); a1 = fma(f1.y, w[k + 1], a1);
        }
        const double2 h2 = *reinterpret_cast<const double2*>(hp + r);
        q[r] = h2.x + a0; q[r + 1] = h2.y + a1;
      }
    }
    if (guu && !isnan(dreg)) {
#pragma unroll
      for (int r = 0; r < NU; ++r) if (r == ua) q[r] += dreg;
    }
    /* ---- publish Quu, Qxu, Qu */
    if (guu) {
#pragma unroll
      for (int r = 0; r < NU; ++r) sQuu[r + NU * ua] = q[r];
      sQxu[NX + LDQ * ua] = qv;                       /* Qu[ua] */
      p.Qu[(size_t)t * NU + ua] = qv;
    }
    if (gxu) {
#pragma unroll
      for (int r = 0; r < NX; ++r) sQxu[r + LDQ * ua] = q[r];
    }
    __syncwarp();

    /* ---- computeGains (src/ddp.cpp:401-429): Cholesky redundantly in every lane */
    double Lm[NU][NU], inv[NU];
#pragma unroll
    for (int j = 0; j < NU; ++j)
#pragma unroll
      for (int i = j; i < NU; ++i) Lm[i][j] = sQuu[i + NU * j];
    bool llt_fail = false;
#pragma unroll
    for (int k = 0; k < NU; ++k) {
      double x = Lm[k][k];
#pragma unroll
      for (int j = 0; j < k; ++j) x = fma(-Lm[k][j], Lm[k][j], x);
      if (x <= 0.0) llt_fail = true;
      const double ri = rsqrt(x);
      Lm[k][k] = x * ri;
      inv[k] = ri;
#pragma unroll
      for (int i = k + 1; i < NU; ++i) {
        double v = Lm[i][k];
#pragma unroll
        for (int j = 0; j < k; ++j) v = fma(-Lm[i][j], Lm[k][j], v);
        Lm[i][k] = v * ri;
      }
    }
    if (llt_fail) return MIMGPU_STATUS_LLT | (t << 8);
    double kc[NU];                                    /* K[:, lane] (lane < NX) or k (lane == NX) */
#pragma unroll
    for (int a = 0; a < NU; ++a) kc[a] = sQxu[rr + LDQ * a];
#pragma unroll
    for (int i = 0; i < NU; ++i) {
      double v = kc[i];
#pragma unroll
      for (int j = 0; j < i; ++j) v = fma(-Lm[i][j], kc[j], v);
      kc[i] = v * inv[i];
    }
#pragma unroll
    for (int i = NU - 1; i >= 0; --i) {
      double v = kc[i];
#pragma unroll
      for (int j = i + 1; j < NU; ++j) v = fma(-Lm[j][i], kc[j], v);
      kc[i] = v * inv[i];
    }
    if (gxx) {
#pragma unroll
      for (int a = 0; a < NU; ++a) p.K[(size_t)t * NX * NU + a * NX + lane] = kc[a];
    }
    if (lane == NX) {
#pragma unroll
      for (int a = 0; a < NU; ++a) { sK[a] = kc[a]; p.kff[(size_t)t * NU + a] = kc[a]; }
    }
    if (lane == NX + 1) {                              /* Quu_llt_[t]: lower factor, upper part zero */
#pragma unroll
      for (int c = 0; c < NU; ++c)
#pragma unroll
        for (int r = 0; r < NU; r += 2)
          *reinterpret_cast<double2*>(p.L + (size_t)t * NU * NU + NU * c + r) =
              make_double2(r >= c ? Lm[r][c] : 0.0, r + 1 >= c ? Lm[r + 1][c] : 0.0);
    }
    __syncwarp();
    if (lane < NU) {                                  /* Quuk = Quu k  (src/ddp.cpp:322) */
      double acc = 0.0;
#pragma unroll
      for (int c = 0; c < NU; ++c) acc = fma(sQuu[lane + NU * c], sK[c], acc);
      p.Quuk[(size_t)t * NU + lane] = acc;
    }
    /* ---- value function (src/ddp.cpp:319-338) */
    double vx;
    {
      double acc = 0.0;
#pragma unroll
      for (int a = 0; a < NU; ++a) acc = fma(kc[a], sQxu[NX + LDQ * a], acc);
      vx = qv - acc;
    }
#pragma unroll
    for (int i = 0; i < NX; i += 2) {
      double a0 = 0.0, a1 = 0.0;
#pragma unroll
      for (int a = 0; a < NU; ++a) {
        const double2 x2 = *reinterpret_cast<const double2*>(sQxu + LDQ * a + i);
        a0 = fma(x2.x, kc[a], a0);
        a1 = fma(x2.y, kc[a], a1);
      }
      q[i] -= a0; q[i + 1] -= a1;
    }
    if (gxx) {
#pragma unroll
      for (int i = 0; i < NX; i += 2) *reinterpret_cast<double2*>(sT + NX * lane + i) = make_double2(q[i], q[i + 1]);
    }
    __syncwarp();
    bool bad = false;
    {
      double acc = 0.0;
#pragma unroll
      for (int i = 0; i < NX; ++i) {
        double v = 0.5 * (q[i] + sT[lx + NX * i]);
        if (i == lx && !isnan(preg)) v += preg;
        q[i] = v;
        acc = fma(v, st[C::bFs + i], acc);
        bad |= raise_if_nan(fabs(v));
      }
      if (!is_feasible) vx += acc;
      bad |= raise_if_nan(fabs(vx));
    }
    if (gxx) {
#pragma unroll
      for (int i = 0; i < NX; i += 2) {
        const double2 v2 = make_double2(q[i], q[i + 1]);
        *reinterpret_cast<double2*>(sV + NX * lane + i) = v2;
        *reinterpret_cast<double2*>(p.Vxx + (size_t)t * NX * NX + NX * lane + i) = v2;
      }
      sVx[lane] = vx;
      p.Vx[(size_t)t * NX + lane] = vx;
    }
    if (__any_sync(0xffffffffu, bad && gxx)) return MIMGPU_STATUS_NONFINITE | (t << 8);
    __syncwarp();
    s ^= 1;
  }
  return 0;
}

}  // namespace mimgpu
#include "kernels_small_mma.cuh"
#include "kernels_small_gj.cuh"
namespace mimgpu {

/* ------------------------------------------------------------------- forward */
/* SolverSQP::forwardPass + checkKKTConditions + gap norm (src/sqp.cpp:270-305, 244-268,
 * 229-236).  Lane i < NX owns state row i, lane a < NU control row a, lane c < N column c
 * of [Fx Fu] for the stationarity residual. */
template <int NX, int NU, class C = SmallCfg<NX, NU>>
__device__ __forceinline__ void small_forward(const Prob& p, double* sw, int lane, double mu_dynamic) {
  constexpr int N = C::N;
  const int T = p.T;
  double* sX = sw + C::oX;
  double* sU = sw + C::oU;
  double* sXn = sw + C::oXn;
  double* sLam = sw + C::oLam;
  const bool gx = lane < NX, gu = lane < NU, gc = lane < N;
  const int lx = gx ? lane : 0, lu = gu ? lane : 0, lc = gc ? lane : 0;

  auto prefetch = [&](int t, double* st) {
    warp_stage<NX * NX>(st + C::fV, p.Vxx + (size_t)(t + 1) * NX * NX, lane);
    warp_stage<NX>(st + C::fVx, p.Vx + (size_t)(t + 1) * NX, lane);
    warp_stage<NX * NU>(st + C::fK, p.K + (size_t)t * NX * NU, lane);
    warp_stage<NU>(st + C::fk, p.kff + (size_t)t * NU, lane);
    warp_stage<NX * NX>(st + C::fF, p.Fx + (size_t)t * NX * NX, lane);
    warp_stage<NX * NU>(st + C::fF + NX * NX, p.Fu + (size_t)t * NX * NU, lane);
    warp_stage<NX>(st + C::fFs, p.fs + (size_t)(t + 1) * NX, lane);
    warp_stage<NX>(st + C::fL, p.Lx + (size_t)t * NX, lane);
    warp_stage<NU>(st + C::fL + NX, p.Lu + (size_t)t * NU, lane);
    cp_async_commit();
  };
  __syncwarp();                 /* the backward pass' global stores are ordered before these reads */
  prefetch(0, sw + C::oS0);

  double xg = 0.0, ug = 0.0, gap = 0.0, kkt = 0.0;
  /* knot 0: dx_0 (member, stays 0), lag_mul_0 = Vxx_0 (dx_0 - fs_0) + Vx_0 */
  double fs_cur = p.fs[lx];
  double x_cur = p.dx[lx];
  if (gx) sX[lane] = x_cur - fs_cur;
  __syncwarp();
  double lam;
  {
    double acc = 0.0;
#pragma unroll
    for (int j = 0; j < NX; ++j) acc = fma(p.Vxx[lx + NX * j], sX[j], acc);
    lam = acc + p.Vx[lx];
    if (gx) p.lag[lane] = lam;
  }
  __syncwarp();
  if (gx) sX[lane] = x_cur;
  __syncwarp();

  int s = 0;
  for (int t = 0; t < T; ++t) {
    const double* st = sw + (s ? C::oS1 : C::oS0);
    if (t + 1 < T) {
      prefetch(t + 1, sw + (s ? C::oS0 : C::oS1));
      cp_async_wait<1>();
    } else {
      cp_async_wait<0>();
    }
    __syncwarp();
    if (gx) { xg += fabs(x_cur); gap += fabs(fs_cur); kkt = nmax(kkt, fabs(fs_cur)); }
    /* du = -K dx - k  (src/sqp.cpp:285-286) */
    {
      double acc = 0.0;
#pragma unroll
      for (int i = 0; i < NX; i += 2) {
        const double2 k2 = *reinterpret_cast<const double2*>(st + C::fK + NX * lu + i);
        acc = fma(k2.x, sX[i], acc);
        acc = fma(k2.y, sX[i + 1], acc);
      }
      double v = -acc;
      v -= st[C::fk + lu];
      if (gu) { sU[lane] = v; p.du[(size_t)t * NU + lane] = v; ug += fabs(v); }
    }
    __syncwarp();
    /* dx+ = fs+ + Fu du + Fx dx  (src/sqp.cpp:287-289) */
    const double fs_n = st[C::fFs + lx];
    double x_n;
    {
      double acc = 0.0;
#pragma unroll
      for (int a = 0; a < NU; ++a) acc = fma(st[C::fF + NX * NX + lx + NX * a], sU[a], acc);
      x_n = fs_n + acc;
      acc = 0.0;
#pragma unroll
      for (int j = 0; j < NX; ++j) acc = fma(st[C::fF + lx + NX * j], sX[j], acc);
      x_n += acc;
      if (gx) { p.dx[(size_t)(t + 1) * NX + lane] = x_n; sXn[lane] = x_n - fs_n; }
    }
    __syncwarp();
    /* lag_mul+ = Vxx+ (dx+ - fs+) + Vx+  (src/sqp.cpp:282-284, 296-298) */
    double lam_n;
    {
      double acc = 0.0;
#pragma unroll
      for (int j = 0; j < NX; ++j) acc = fma(st[C::fV + lx + NX * j], sXn[j], acc);
      lam_n = acc + st[C::fVx + lx];
      if (gx) { p.lag[(size_t)(t + 1) * NX + lane] = lam_n; sLam[lane] = lam_n; }
    }
    __syncwarp();
    /* stationarity rows of knot t  (src/sqp.cpp:249-258) */
    {
      double acc = 0.0;
#pragma unroll
      for (int k = 0; k < NX; k += 2) {
        const double2 f2 = *reinterpret_cast<const double2*>(st + C::fF + NX * lc + k);
        acc = fma(f2.x, sLam[k], acc);
        acc = fma(f2.y, sLam[k + 1], acc);
      }
      double v = st[C::fL + lc] + acc;
      if (gx) v -= lam;
      if (gc) kkt = nmax(kkt, fabs(v));
    }
    if (gx) sX[lane] = x_n;
    x_cur = x_n; fs_cur = fs_n; lam = lam_n;
    __syncwarp();
    s ^= 1;
  }
  if (gx) {
    xg += fabs(x_cur); gap += fabs(fs_cur);
    kkt = nmax(kkt, fabs(fs_cur));
    kkt = nmax(kkt, fabs(p.LxT[lane] - lam));          /* src/sqp.cpp:259-262 */
  }
  xg = warp_sum(xg); ug = warp_sum(ug); gap = warp_sum(gap); kkt = warp_nmax(kkt);
  if (lane == 0) {
    const size_t B = p.B, b = p.b;
    p.scal[MIMGPU_S_KKT * B + b] = kkt;
    p.scal[MIMGPU_S_GAP_NORM * B + b] = gap;
    p.scal[MIMGPU_S_MERIT * B + b] = p.cost + mu_dynamic * gap;
    p.scal[MIMGPU_S_XGRAD_NORM * B + b] = xg / (double)(T + 1);
    p.scal[MIMGPU_S_UGRAD_NORM * B + b] = ug / (double)T;
  }
}

/* -------------------------------------------------------------------- kernel */
/* Launch shape per specialisation: warps per CTA and resident CTAs per SM.  The UR5 shape
 * runs 2 x 7 = 14 problems per SM (register budget 144, 15.7 KB of shared memory per warp):
 * 148 SMs x 14 = 2072 slots, so a 4096-problem batch is two full waves. */
template <int NX, int NU> struct SmallTune { static constexpr int WPB = 4, MINB = 1; };
template <> struct SmallTune<12, 6> { static constexpr int WPB = 2, MINB = 7; };

template <int NX, int NU, int WPB, bool NOMMA = false>
__global__ void __launch_bounds__(WPB * 32, SmallTune<NX, NU>::MINB) k_sqp_direction_small(KArgs a, SqpKPar kp) {
  extern __shared__ double smem_small[];
  using C = SmallCfg<NX, NU>;
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  const int slot = blockIdx.x * WPB + wib;
  if (slot >= a.n_active) return;
  const int b = a.active ? a.active[slot] : slot;
  double* sw = smem_small + (size_t)wib * C::WARP;
  Prob p = bind(a, b);
  const double preg = a.preg_arr ? a.preg_arr[b] : kp.preg;
  const double dreg = a.dreg_arr ? a.dreg_arr[b] : kp.dreg;
  int st;
  if constexpr (NX % 4 == 0 && !NOMMA)
    st = small_backward_mma<NX, NU>(p, sw, lane, preg, dreg, kp.is_feasible);
  else
    st = small_backward<NX, NU>(p, sw, lane, preg, dreg, kp.is_feasible);
  /* pad_ bit 0: members-only rerun (the C ABI asks for Qu_ / Quuk_ / Quu_llt_ after a fast launch):
   * the direction, the scalars and the status word of the fast launch stay as they are */
  if (a.pad_ & 1) { cp_async_wait<0>(); return; }
  if (lane == 0) { a.status[b] = st; a.qp_iters[b] = 0; }
  if (st) {
    cp_async_wait<0>();
    double gap = 0.0;
    for (int e = lane; e < (p.T + 1) * NX; e += 32) gap += fabs(p.fs[e]);
    gap = warp_sum(gap);
    if (lane == 0) {
      p.scal[MIMGPU_S_GAP_NORM * (size_t)p.B + b] = gap;
      p.scal[MIMGPU_S_MERIT * (size_t)p.B + b] = p.cost + kp.mu_dynamic * gap;
    }
    return;
  }
  small_forward<NX, NU>(p, sw, lane, kp.mu_dynamic);
}

/* Fast variant (kernels_small_gj.cuh): register-resident backward sweep, K k Vxx Vx spilled for
 * the forward pass only.  8.3 KB of shared memory per warp. */
template <int NX, int NU> struct GjTune { static constexpr int WPB = 2, MINB = 7; };

template <int NX, int NU, int WPB>
__global__ void __launch_bounds__(WPB * 32, GjTune<NX, NU>::MINB) k_sqp_direction_small_gj(KArgs a, SqpKPar kp) {
  extern __shared__ double smem_small[];
  using C = GjCfg<NX, NU>;
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  const int slot = blockIdx.x * WPB + wib;
  if (slot >= a.n_active) return;
  const int b = a.active ? a.active[slot] : slot + a.first;
  double* sw = smem_small + (size_t)wib * C::WARP;
  Prob p = bind(a, b);
  const int T = p.T;
  const double* pk = a.pack + (size_t)b * C::pack_doubles(T);
  double* sp = a.spill + (size_t)b * C::spill_doubles(T);
  const double preg = a.preg_arr ? a.preg_arr[b] : kp.preg;
  const double dreg = a.dreg_arr ? a.dreg_arr[b] : kp.dreg;
  const int st = small_backward_gj<NX, NU>(pk, sp, T, sw, lane, preg, dreg, kp.is_feasible);
  if (lane == 0) { a.status[b] = st; a.qp_iters[b] = 0; }
  if (st) {
    double gap = 0.0;
    for (int e = lane; e < (T + 1) * NX; e += 32) gap += fabs(pk[(size_t)(e / NX) * C::REC + C::bFs + e % NX]); /* fs_t heads every record */
    gap = warp_sum(gap);
    if (lane == 0) {
      p.scal[MIMGPU_S_GAP_NORM * (size_t)p.B + b] = gap;
      p.scal[MIMGPU_S_MERIT * (size_t)p.B + b] = p.cost + kp.mu_dynamic * gap;
    }
    return;
  }
  small_forward_gj<NX, NU>(p, pk, sp, sw, lane, kp.mu_dynamic);
}

/* Upload-time repack of the field-major C-ABI arrays into knot records (one CTA per knot). */
template <int NX, int NU>
__global__ void __launch_bounds__(128) k_pack_knots(KArgs a, double* pack, int b0) {
  using C = GjCfg<NX, NU>;
  constexpr int N = C::N;
  const int T = a.sh.T;
  const size_t b = (size_t)b0 + blockIdx.x / (T + 1);
  const int t = blockIdx.x % (T + 1);
  double* dst = pack + b * C::pack_doubles(T) + (size_t)t * C::REC;
  auto copy = [&](double* d, const double* s, int n) {
    for (int e = threadIdx.x; e < n; e += blockDim.x) d[e] = s[e];
  };
  /* symmetric block -> its packed lower triangle (GjCfg::sym_off); the upper triangle of the input is not read */
  auto copy_sym = [&](double* d, const double* s, int n) {
    for (int e = threadIdx.x; e < n * n; e += blockDim.x) {
      const int i = e % n, j = e / n;
      if (i >= j) d[C::sym_off(n, i, j)] = s[e];
    }
  };
  copy(dst + C::bFs, a.fs + (b * (T + 1) + t) * NX, NX);
  if (t == T) {
    copy_sym(dst + C::tLxx, a.LxxT + b * NX * NX, NX);
    copy(dst + C::tLx, a.LxT + b * NX, NX);
    return;
  }
  const size_t k = b * T + t;
  copy_sym(dst + C::bH, a.Lxx + k * NX * NX, NX);
  copy(dst + C::bLxu, a.Lxu + k * NX * NU, NX * NU);
  copy_sym(dst + C::bLuu, a.Luu + k * NU * NU, NU);
  copy(dst + C::bL, a.Lx + k * NX, NX);
  copy(dst + C::bL + NX, a.Lu + k * NU, NU);
  copy(dst + C::bF, a.Fx + k * NX * NX, NX * NX);
  copy(dst + C::bF + NX * NX, a.Fu + k * NX * NU, NX * NU);
  (void)N;
}

/* Inverse of k_pack_knots: knot records -> the field-major C-ABI arrays (out.Fx ... are written).
 * Needed only when a caller that uploaded RECORDS later asks for something the record kernel does
 * not serve (Qu_/Quuk_/Quu_llt_ downloads, FDDP or CSQP on the same ctx). */
struct UnpackDst {
  double *Fx, *Fu, *Lxx, *Lxu, *Luu, *Lx, *Lu, *fs, *LxxT, *LxT;
};
template <int NX, int NU>
__global__ void __launch_bounds__(128) k_unpack_records(const double* pack, UnpackDst o, int T) {
  using C = GjCfg<NX, NU>;
  const size_t b = blockIdx.x / (T + 1);
  const int t = blockIdx.x % (T + 1);
  const double* src = pack + b * C::pack_doubles(T) + (size_t)t * C::REC;
  auto copy = [&](double* d, const double* s, int n) {
    for (int e = threadIdx.x; e < n; e += blockDim.x) d[e] = s[e];
  };
  auto copy_sym = [&](double* d, const double* s, int n) {   /* packed lower triangle -> full symmetric block */
    for (int e = threadIdx.x; e < n * n; e += blockDim.x) d[e] = s[C::sym_off(n, e % n, e / n)];
  };
  copy(o.fs + (b * (T + 1) + t) * NX, src + C::bFs, NX);
  if (t == T) {
    copy_sym(o.LxxT + b * NX * NX, src + C::tLxx, NX);
    copy(o.LxT + b * NX, src + C::tLx, NX);
    return;
  }
  const size_t k = b * T + t;
  copy_sym(o.Lxx + k * NX * NX, src + C::bH, NX);
  copy(o.Lxu + k * NX * NU, src + C::bLxu, NX * NU);
  copy_sym(o.Luu + k * NU * NU, src + C::bLuu, NU);
  copy(o.Lx + k * NX, src + C::bL, NX);
  copy(o.Lu + k * NU, src + C::bL + NX, NU);
  copy(o.Fx + k * NX * NX, src + C::bF, NX * NX);
  copy(o.Fu + k * NX * NU, src + C::bF + NX * NX, NX * NU);
}

/* On-demand copy of the spill records into the solver-member arrays Vxx_ Vx_ K_ k_ of the C ABI. */
template <int NX, int NU>
__global__ void __launch_bounds__(128) k_unpack_spill(KArgs a) {
  using C = GjCfg<NX, NU>;
  const int T = a.sh.T;
  const size_t b = blockIdx.x / (T + 1);
  const int t = blockIdx.x % (T + 1);
  const double* src = a.spill + b * C::spill_doubles(T) + (size_t)t * C::SREC;
  auto copy = [&](double* d, const double* s, int n) {
    for (int e = threadIdx.x; e < n; e += blockDim.x) d[e] = s[e];
  };
  /* Vxx_ from the spill: dense rows, or both halves from the packed upper triangle (GjCfg::TRI, vrow) */
  {
    double* V = a.Vxx + (b * (T + 1) + t) * NX * NX;
    for (int e = threadIdx.x; e < NX * NX; e += blockDim.x) {
      const int r = e / NX, c = e % NX;
      const int i = (C::TRI && r > c) ? c : r, n = (C::TRI && r > c) ? r : c;     /* TRI: (i, n), i <= n */
      V[e] = src[C::sV + C::vrow(i) + n - C::vstart(i)];
    }
  }
  copy(a.Vx + (b * (T + 1) + t) * NX, src + C::sVx, NX);
  if (t < T) {
    copy(a.K + (b * T + t) * NX * NU, src + C::sK, NX * NU);
    copy(a.kff + (b * T + t) * NU, src + C::sk, NU);
  }
}

/* ------------------------------------------------------------------ dispatch */
/* MIMGPU_SMALL_NOMMA=1 selects the DFMA (CUDA-core) variant of the full-spill backward sweep and
 * MIMGPU_SMALL_FULL=1 makes the full-spill kernel the default: A/B switches for profiling, all
 * variants are parity-tested. */
static inline bool small_nomma() {
  static const bool v = getenv("MIMGPU_SMALL_NOMMA") != nullptr;
  return v;
}
static inline bool small_force_full() {
  static const bool v = getenv("MIMGPU_SMALL_FULL") != nullptr || getenv("MIMGPU_SMALL_NOMMA") != nullptr;
  return v;
}

template <int NX, int NU, int WPB>
static int small_launch_t(const KArgs& a, const SqpKPar& kp, cudaStream_t st, bool configure_only) {
  using C = SmallCfg<NX, NU>;
  const int smem = WPB * C::WARP * (int)sizeof(double);
  auto kern = small_nomma() ? k_sqp_direction_small<NX, NU, WPB, true> : k_sqp_direction_small<NX, NU, WPB, false>;
  if (configure_only) {
    if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem) != cudaSuccess) return 1;
    return 0;
  }
  const int grid = (a.n_active + WPB - 1) / WPB;
  kern<<<grid, WPB * 32, smem, st>>>(a, kp);
  return 0;
}
template <int NX, int NU, int WPB>
static int small_launch_gj_t(const KArgs& a, const SqpKPar& kp, cudaStream_t st, bool configure_only) {
  if constexpr (NX % 4 == 0) {
    using C = GjCfg<NX, NU>;
    const int smem = WPB * C::WARP * (int)sizeof(double);
    auto kern = k_sqp_direction_small_gj<NX, NU, WPB>;
    if (configure_only) {
      if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem) != cudaSuccess) return 1;
      cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, 100);
      return 0;
    }
    const int grid = (a.n_active + WPB - 1) / WPB;
    kern<<<grid, WPB * 32, smem, st>>>(a, kp);
    return 0;
  } else {
    return -1;
  }
}

#define MIMGPU_SMALL_SHAPES(X) X(12, 6) X(8, 4) X(16, 8) X(4, 2) X(6, 2) X(10, 4)

/* bit 0: a full-spill kernel exists for this shape, bit 1: the fast variant too */
static inline int small_kernels_configure(const mimgpu_shape& s) {
  KArgs dummy{};
  SqpKPar kp{};
#define X(NX_, NU_)                                                                                         \
  if (s.nx == NX_ && s.nu == NU_) {                                                                         \
    int r = small_launch_t<NX_, NU_, SmallTune<NX_, NU_>::WPB>(dummy, kp, nullptr, true) == 0 ? 1 : 0;      \
    if (r && small_launch_gj_t<NX_, NU_, GjTune<NX_, NU_>::WPB>(dummy, kp, nullptr, true) == 0) r |= 2;     \
    return r;                                                                                               \
  }
  MIMGPU_SMALL_SHAPES(X)
#undef X
  return 0;
}

/* variant 0: fast kernel when it exists (else full), 1: full-spill kernel.  Returns 0 when the
 * fast kernel was launched, 1 when the full-spill kernel was launched, < 0 when no
 * specialisation applies (caller uses the generic kernel). */
static inline int small_sqp_direction_launch(const KArgs& a, const SqpKPar& kp, int variant, cudaStream_t st) {
  /* cp.async needs 16-byte aligned sources: all blocks have even sizes, so only the base
   * pointers matter */
  const uintptr_t al = (uintptr_t)a.Fx | (uintptr_t)a.Fu | (uintptr_t)a.Lxx | (uintptr_t)a.Lxu | (uintptr_t)a.Luu |
                       (uintptr_t)a.Lx | (uintptr_t)a.Lu | (uintptr_t)a.fs | (uintptr_t)a.LxxT;
  if (al & 15) return -1;
#define X(NX_, NU_)                                                                                          \
  if (a.sh.nx == NX_ && a.sh.nu == NU_) {                                                                    \
    if (variant == 0 && !small_force_full() && a.pack && a.spill &&                                         \
        small_launch_gj_t<NX_, NU_, GjTune<NX_, NU_>::WPB>(a, kp, st, false) == 0)                           \
      return 0;                                                                                              \
    return small_launch_t<NX_, NU_, SmallTune<NX_, NU_>::WPB>(a, kp, st, false) == 0 ? 1 : -1;              \
  }
  MIMGPU_SMALL_SHAPES(X)
#undef X
  return -1;
}

/* knot-packed buffers of the fast variant: doubles per problem (0: no fast variant for the shape) */
static inline int small_pack_layout(const mimgpu_shape& s, size_t* pack_doubles, size_t* spill_doubles) {
  *pack_doubles = 0; *spill_doubles = 0;
#define X(NX_, NU_)                                                             \
  if (s.nx == NX_ && s.nu == NU_) {                                             \
    if constexpr (NX_ % 4 == 0) {                                               \
      *pack_doubles = GjCfg<NX_, NU_>::pack_doubles(s.T);                       \
      *spill_doubles = GjCfg<NX_, NU_>::spill_doubles(s.T);                     \
      return 0;                                                                 \
    }                                                                           \
    return -1;                                                                  \
  }
  MIMGPU_SMALL_SHAPES(X)
#undef X
  return -1;
}
/* field-major inputs (a.Fx ...) -> knot records in `pack`, problems [b0, b0 + nb) */
static inline int small_pack_launch(const KArgs& a, double* pack, int b0, int nb, cudaStream_t st) {
#define X(NX_, NU_)                                                                             \
  if (a.sh.nx == NX_ && a.sh.nu == NU_) {                                                       \
    if constexpr (NX_ % 4 == 0) {                                                               \
      k_pack_knots<NX_, NU_><<<(unsigned)nb * (a.sh.T + 1), 128, 0, st>>>(a, pack, b0);         \
      return 0;                                                                                 \
    }                                                                                           \
    return -1;                                                                                  \
  }
  MIMGPU_SMALL_SHAPES(X)
#undef X
  return -1;
}
/* knot records -> field-major arrays (a.Fx ... must point at writable storage), all B problems */
static inline int small_unpack_inputs_launch(const KArgs& a, const double* pack, cudaStream_t st) {
  UnpackDst o{const_cast<double*>(a.Fx), const_cast<double*>(a.Fu), const_cast<double*>(a.Lxx), const_cast<double*>(a.Lxu),
              const_cast<double*>(a.Luu), const_cast<double*>(a.Lx), const_cast<double*>(a.Lu), const_cast<double*>(a.fs),
              const_cast<double*>(a.LxxT), const_cast<double*>(a.LxT)};
#define X(NX_, NU_)                                                                                   \
  if (a.sh.nx == NX_ && a.sh.nu == NU_) {                                                             \
    if constexpr (NX_ % 4 == 0) {                                                                     \
      k_unpack_records<NX_, NU_><<<(unsigned)a.sh.B * (a.sh.T + 1), 128, 0, st>>>(pack, o, a.sh.T);   \
      return 0;                                                                                       \
    }                                                                                                 \
    return -1;                                                                                        \
  }
  MIMGPU_SMALL_SHAPES(X)
#undef X
  return -1;
}
/* offsets (in doubles) of every block inside a knot record / the terminal record */
static inline int small_record_offsets(const mimgpu_shape& s, int out[16]) {
#define X(NX_, NU_)                                                                                   \
  if (s.nx == NX_ && s.nu == NU_) {                                                                   \
    if constexpr (NX_ % 4 == 0) {                                                                     \
      using C = GjCfg<NX_, NU_>;                                                                      \
      const int v[16] = {C::REC, C::TREC, (int)C::pack_doubles(s.T), C::bFs, C::bH, C::bLxu, C::bLuu, C::bL,          \
                         C::bL + NX_, C::bF, C::bF + NX_ * NX_, C::tFs, C::tLxx, C::tLx, C::SYM ? 1 : 0, 0};       \
      for (int i = 0; i < 16; ++i) out[i] = v[i];                                                     \
      return 0;                                                                                       \
    }                                                                                                 \
    return -1;                                                                                        \
  }
  MIMGPU_SMALL_SHAPES(X)
#undef X
  return -1;
}
/* spill records -> Vxx_ Vx_ K_ k_ arrays of the C ABI, all B problems */
static inline int small_unpack_launch(const KArgs& a, cudaStream_t st) {
#define X(NX_, NU_)                                                                             \
  if (a.sh.nx == NX_ && a.sh.nu == NU_) {                                                       \
    if constexpr (NX_ % 4 == 0) {                                                               \
      k_unpack_spill<NX_, NU_><<<(unsigned)a.sh.B * (a.sh.T + 1), 128, 0, st>>>(a);             \
      return 0;                                                                                 \
    }                                                                                           \
    return -1;                                                                                  \
  }
  MIMGPU_SMALL_SHAPES(X)
#undef X
  return -1;
}

}  // namespace mimgpu
