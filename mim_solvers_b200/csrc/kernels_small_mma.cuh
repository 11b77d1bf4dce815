/*
 * kernels_small_mma.cuh -- Riccati backward sweep of the warp-per-problem kernel with the
 * three block products on the FP64 tensor pipe (mma.sync.m8n8k4.f64, SASS DMMA).
 * Included by kernels_small.cuh (needs SmallCfg, warp_stage, cp_async_*).
 *
 * Why DMMA at nx = 12: peak FP64 rate is the same as DFMA on B200 (measured 36.8 vs 37.1
 * TFLOP/s), but a DFMA column-per-lane product needs one shared-memory broadcast per FMA
 * instruction and the first version of this kernel was bound by the shared-memory pipe
 * (82 % busy, FP64 pipe 31 %; profiles/ncu_full_r01_v2_small_sqp.csv).  DMMA takes its
 * operands as register fragments spread over the warp: 2 loads per 256 FMAs.
 *
 * Restates SolverDDP::backwardPass + computeGains (src/ddp.cpp:265-348, 401-429):
 *   step 1  Wt = [Fx Fu]^T Vxx'          (FxTVxx, FuTVxx; extra column Vx' gives Qx, Qu)
 *   step 2  Q  = H + [Fx Fu]^T W         (Qxx, Qxu, Quu)
 *   gains   LLT(Quu) redundantly per lane, K / k by substitution (one column per lane)
 *   step 4  Vxx = Qxx - Qxu K, symmetrise, + preg;  Vx = Qx - K^T Qu + Vxx fs
 * Fragment notation: g = lane >> 2, t = lane & 3.  A (8x4): A[g][t];  B (4x8): B[t][g];
 * C (8x8): C[g][2t], C[g][2t+1].  The contraction index of steps 1 and 2 is visited in
 * the order kperm() so that the C fragments of step 1 are directly the B fragments of
 * step 2 (no shared-memory round trip).
 */
#pragma once

namespace mimgpu {

__device__ __forceinline__ void dmma884(double& c0, double& c1, double a, double b) {
  asm("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
      : "+d"(c0), "+d"(c1)
      : "d"(a), "d"(b));
}

/* contraction index handled by lane-in-group t in k-tile kt (NX % 4 == 0) */
template <int NX>
__device__ __forceinline__ int kperm(int kt, int t) {
  const int b8 = kt >> 1;
  if (8 * b8 + 8 <= NX) return 8 * b8 + 2 * t + (kt & 1);
  return 8 * b8 + (((t & 1) << 1) | (t >> 1));   /* trailing 4-block: 0,2,1,3 */
}

template <int NX, int NU>
__device__ __forceinline__ int small_backward_mma(const Prob& p, double* sw, int lane, double preg, double dreg,
                                                  int is_feasible) {
  using C = SmallCfg<NX, NU>;
  static_assert(NX % 4 == 0, "k-tiles of 4");
  constexpr int N = C::N, NT = C::NT, XT = C::XT, KT = C::KT, KPT = C::KPT, LDP = C::LDP, PC0 = C::PC0,
                LDK = C::LDK;
  constexpr int KQ = NX / 8, TQ = (NX % 8) / 2;     /* Wt tile column / lane-in-group holding k' = NX */
  const int T = p.T;
  const int g = lane >> 2, tq = lane & 3;
  double* sV = sw + C::oV;     /* Vxx' (symmetric), [m * NX + k] */
  double* sT = sw + C::oT;     /* raw Vxx for the symmetrisation, row-major NX x NX */
  double* sVx = sw + C::oVx;
  double* sK = sw + C::oK;     /* k */
  double* sQp = sw + C::oQp;   /* published Q columns >= PC0, row-major [i * LDP + n - PC0] */
  double* sQv = sw + C::oQv;   /* [Qx; Qu] */
  double* sKm = sw + C::oKm;   /* K row-major [a * LDK + j] */
  const bool gxx = lane < NX;
  const int lx = gxx ? lane : 0;
  const int rr = lane <= NX ? lane : 0;

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

  /* ---- terminal knot (src/ddp.cpp:268-277): sV[m * NX + k] = Vxx_T[m][k] */
  {
    double v[NX];
#pragma unroll
    for (int i = 0; i < NX; i += 2) {
      const double2 t2 = *reinterpret_cast<const double2*>(p.LxxT + NX * lx + i);
      v[i] = t2.x; v[i + 1] = t2.y;                      /* column lx of Lxx_T */
    }
    if (!isnan(preg)) {
#pragma unroll
      for (int i = 0; i < NX; ++i) if (i == lx) v[i] += preg;
    }
    if (gxx) {
#pragma unroll
      for (int i = 0; i < NX; ++i) sV[i * NX + lane] = v[i];
#pragma unroll
      for (int i = 0; i < NX; i += 2)
        *reinterpret_cast<double2*>(p.Vxx + (size_t)T * NX * NX + NX * lane + i) = make_double2(v[i], v[i + 1]);
    }
    __syncwarp();
    double vx = p.LxT[lx];
    if (!is_feasible) {
      double acc = 0.0;
#pragma unroll
      for (int j = 0; j < NX; ++j) acc = fma(sV[lx * NX + j], p.fs[(size_t)T * NX + j], acc);
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

    /* ---- A fragments of [Fx Fu]^T, shared by steps 1 and 2: F[kperm(kt)][8 nt + g] */
    double fa[NT][KT];
#pragma unroll
    for (int nt = 0; nt < NT; ++nt) {
      const int n = 8 * nt + g;
#pragma unroll
      for (int kt = 0; kt < KT; ++kt) fa[nt][kt] = (n < N) ? st[C::bF + kperm<NX>(kt, tq) + NX * n] : 0.0;
    }
    /* ---- step 1: Wt[n][k'] = sum_m F[m][n] Vp[m][k'],  Vp = [Vxx' | Vx' | 0] */
    double wt[NT][KPT][2];
#pragma unroll
    for (int nt = 0; nt < NT; ++nt)
#pragma unroll
      for (int kp = 0; kp < KPT; ++kp) { wt[nt][kp][0] = 0.0; wt[nt][kp][1] = 0.0; }
#pragma unroll
    for (int kt = 0; kt < KT; ++kt) {
      const int m = kperm<NX>(kt, tq);
#pragma unroll
      for (int kp = 0; kp < KPT; ++kp) {
        const int kk = 8 * kp + g;
        const double vb = (kk < NX) ? sV[m * NX + kk] : (kk == NX ? sVx[m] : 0.0);
#pragma unroll
        for (int nt = 0; nt < NT; ++nt) dmma884(wt[nt][kp][0], wt[nt][kp][1], fa[nt][kt], vb);
      }
    }
    /* [Qx; Qu][n] = [Lx; Lu][n] + (F^T Vx')[n]  (src/ddp.cpp:291-292, 301-302) */
    if (tq == TQ) {
#pragma unroll
      for (int nt = 0; nt < NT; ++nt) {
        const int n = 8 * nt + g;
        if (n < N) {
          const double qv = st[C::bL + n] + wt[nt][KQ][0];
          sQv[n] = qv;
          if (n >= NX) p.Qu[(size_t)t * NU + n - NX] = qv;
        }
      }
    }
    /* ---- step 2: Q[i][n] = H[i][n] + sum_k F[k][i] W[k][n]  (src/ddp.cpp:295-314) */
    auto hval = [&](int i, int n) -> double {
      if (i < NX && n < N) return st[C::bH + i + NX * n];                       /* Lxx | Lxu */
      if (i >= NX && i < N && n >= NX && n < N) {
        double v = st[C::bLuu + (i - NX) + NU * (n - NX)];                      /* Luu (+ dreg) */
        if (i == n && !isnan(dreg)) v += dreg;
        return v;
      }
      return 0.0;
    };
    double q[NT][NT][2];
#pragma unroll
    for (int it = 0; it < NT; ++it)
#pragma unroll
      for (int nt = 0; nt < NT; ++nt) {
        if (it < XT || nt >= it) {
          const int i = 8 * it + g, n = 8 * nt + 2 * tq;
          q[it][nt][0] = hval(i, n);
          q[it][nt][1] = hval(i, n + 1);
        }
      }
#pragma unroll
    for (int kt = 0; kt < KT; ++kt) {
      const int b8 = kt >> 1;
      const bool full = (8 * b8 + 8 <= NX);
#pragma unroll
      for (int nt = 0; nt < NT; ++nt) {
        double wb;
        if (full) {
          wb = wt[nt][b8][kt & 1];
        } else {  /* trailing 4-block: k = 8 b8 + {0,2,1,3}[t] */
          const double up = __shfl_sync(0xffffffffu, wt[nt][b8][1], (lane - 2) & 31);
          wb = (tq < 2) ? wt[nt][b8][0] : up;
        }
#pragma unroll
        for (int it = 0; it < NT; ++it)
          if (it < XT || nt >= it) dmma884(q[it][nt][0], q[it][nt][1], fa[it][kt], wb);
      }
    }
    /* ---- publish the Q columns >= PC0 (Qxu, Quu) */
#pragma unroll
    for (int it = 0; it < NT; ++it)
#pragma unroll
      for (int nt = NX / 8; nt < NT; ++nt)
        if (it < XT || nt >= it)
          *reinterpret_cast<double2*>(sQp + (8 * it + g) * LDP + 8 * nt + 2 * tq - PC0) =
              make_double2(q[it][nt][0], q[it][nt][1]);
    __syncwarp();

    /* ---- computeGains (src/ddp.cpp:401-429): Cholesky redundantly in every lane */
    double Lm[NU][NU], inv[NU];
#pragma unroll
    for (int c = 0; c < NU; ++c)
#pragma unroll
      for (int a = c; a < NU; ++a) {
        const int r = NX + a, n = NX + c;
        const bool direct = (r / 8 < XT) || (n / 8 >= r / 8);
        Lm[a][c] = direct ? sQp[r * LDP + n - PC0] : sQp[n * LDP + r - PC0];
      }
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
    double kc[NU];                                      /* K[:, lane] (lane < NX) or k (lane == NX) */
#pragma unroll
    for (int a = 0; a < NU; ++a) kc[a] = (lane == NX) ? sQv[NX + a] : sQp[rr * LDP + NX + a - PC0];
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
      for (int a = 0; a < NU; ++a) { p.K[(size_t)t * NX * NU + a * NX + lane] = kc[a]; sKm[a * LDK + lane] = kc[a]; }
    } else if (lane < 8 * XT) {
#pragma unroll
      for (int a = 0; a < NU; ++a) sKm[a * LDK + lane] = 0.0;
    }
    if (lane == NX) {
#pragma unroll
      for (int a = 0; a < NU; ++a) { sK[a] = kc[a]; p.kff[(size_t)t * NU + a] = kc[a]; }
    }
    if (lane == NX + 1) {                                /* Quu_llt_[t]: lower factor, upper part zero */
#pragma unroll
      for (int c = 0; c < NU; ++c)
#pragma unroll
        for (int r = 0; r < NU; r += 2)
          *reinterpret_cast<double2*>(p.L + (size_t)t * NU * NU + NU * c + r) =
              make_double2(r >= c ? Lm[r][c] : 0.0, r + 1 >= c ? Lm[r + 1][c] : 0.0);
    }
    __syncwarp();
    if (lane < NU) {                                    /* Quuk = Quu k  (src/ddp.cpp:322) */
      double acc = 0.0;
#pragma unroll
      for (int c = 0; c < NU; ++c) {
        const int r = NX + lane, n = NX + c;
        /* element (lane, c): published directly unless its tile was skipped */
        const bool direct = (r / 8 < XT) || (n / 8 >= r / 8);
        const double quu = direct ? sQp[r * LDP + n - PC0] : sQp[n * LDP + r - PC0];
        acc = fma(quu, sK[c], acc);
      }
      p.Quuk[(size_t)t * NU + lane] = acc;
    }
    /* ---- step 4: Vxx = Qxx - Qxu K  (src/ddp.cpp:319-327) on the Qxx accumulators */
    {
      constexpr int AT = (NU + 3) / 4;
#pragma unroll
      for (int at = 0; at < AT; ++at) {
        const int a = 4 * at + tq;
        double xa[XT], kb[XT];
#pragma unroll
        for (int it = 0; it < XT; ++it) {
          const int i = 8 * it + g;
          xa[it] = (a < NU && i < NX) ? sQp[i * LDP + NX + a - PC0] : 0.0;
          kb[it] = (a < NU) ? -sKm[a * LDK + i] : 0.0;
        }
#pragma unroll
        for (int it = 0; it < XT; ++it)
#pragma unroll
          for (int jt = 0; jt < XT; ++jt) dmma884(q[it][jt][0], q[it][jt][1], xa[it], kb[jt]);
      }
    }
    double vx;                                           /* Vx = Qx - K^T Qu */
    {
      double acc = 0.0;
#pragma unroll
      for (int a = 0; a < NU; ++a) acc = fma(kc[a], sQv[NX + a], acc);
      vx = sQv[lx] - acc;
    }
    /* ---- symmetrise, + preg  (src/ddp.cpp:328-333) */
#pragma unroll
    for (int it = 0; it < XT; ++it)
#pragma unroll
      for (int jt = 0; jt < XT; ++jt) {
        const int i = 8 * it + g, n = 8 * jt + 2 * tq;
        if (i < NX && n < NX)
          *reinterpret_cast<double2*>(sT + i * NX + n) = make_double2(q[it][jt][0], q[it][jt][1]);
      }
    __syncwarp();
    bool bad = false;
#pragma unroll
    for (int it = 0; it < XT; ++it)
#pragma unroll
      for (int jt = 0; jt < XT; ++jt) {
        const int i = 8 * it + g, n = 8 * jt + 2 * tq;
        if (i < NX && n < NX) {
          double v0 = 0.5 * (q[it][jt][0] + sT[n * NX + i]);
          double v1 = 0.5 * (q[it][jt][1] + sT[(n + 1) * NX + i]);
          if (!isnan(preg)) { if (i == n) v0 += preg; if (i == n + 1) v1 += preg; }
          bad |= raise_if_nan(fabs(v0)) | raise_if_nan(fabs(v1));
          const double2 v2 = make_double2(v0, v1);
          *reinterpret_cast<double2*>(sV + i * NX + n) = v2;
          *reinterpret_cast<double2*>(p.Vxx + (size_t)t * NX * NX + i * NX + n) = v2;
        }
      }
    __syncwarp();
    if (!is_feasible) {                                  /* Vx += Vxx fs_t  (src/ddp.cpp:336-338) */
      double acc = 0.0;
#pragma unroll
      for (int j = 0; j < NX; j += 2) {
        const double2 v2 = *reinterpret_cast<const double2*>(sV + lx * NX + j);
        acc = fma(v2.x, st[C::bFs + j], acc);
        acc = fma(v2.y, st[C::bFs + j + 1], acc);
      }
      vx += acc;
    }
    bad |= gxx && raise_if_nan(fabs(vx));
    if (gxx) { sVx[lane] = vx; p.Vx[(size_t)t * NX + lane] = vx; }
    if (__any_sync(0xffffffffu, bad)) return MIMGPU_STATUS_NONFINITE | (t << 8);
    __syncwarp();
    s ^= 1;
  }
  return 0;
}

}  // namespace mimgpu
