/*
 * kernels_small_gj.cuh -- register-resident Riccati backward sweep of the warp-per-problem
 * SQP kernel ("fast" variant).  Included by kernels_small.cuh.
 *
 * Restates SolverDDP::backwardPass + computeGains (src/ddp.cpp:265-348, 401-429) with the
 * whole knot in DMMA fragments -- nothing but the staged inputs goes through shared memory:
 *   step 1  Wt = [Fx Fu]^T [Vxx' | Vx']   DMMA; the B operand is the previous knot's Vxx,
 *           which already sits in C-fragment registers: Vxx is symmetric, and element
 *           (8 kp + g, kperm(kt, t)) of a C fragment is exactly B[t][g] of k-tile kt.
 *   step 2  M  = [H | l] + [Fx Fu]^T [W | .]   DMMA on all tiles of the (N x N+1) matrix
 *           M = [[Qxx Qxu | Qx], [Qux Quu | Qu]]   (N = NX + NU, last column = gradient).
 *   gains   NU Gauss-Jordan pivots on the u-rows of M, in the order 0..NU-1.  The pivots are
 *           the squares of the diagonal of the Cholesky factor the reference computes, so
 *           "pivot <= 0" is Eigen's LLT failure test (src/ddp.cpp:413-419); afterwards
 *           the u-rows hold [K | k] and the x-rows hold the Schur complement
 *           [Qxx - Qxu K | Qx - K^T Qu] = [Vxx | Vx] (src/ddp.cpp:319-327).  A pivot costs
 *           ten 64-bit warp shuffles and one FMA per matrix element; no Cholesky factor,
 *           no substitution, no shared-memory round trip.
 *   close   + preg, Vx += Vxx fs (src/ddp.cpp:331-338), raiseIfNaN, spill K k Vxx Vx for the
 *           forward pass.
 * Differences from the reference that stay inside the 1e-9 parity bar: the elimination order
 * of the floating-point sums, and Vxx is symmetric only to rounding (the reference
 * averages Vxx and its transpose; here both triangles come out of the same symmetric
 * update).  Qu, Quuk and the LLT factor are NOT spilled by this variant: the C ABI
 * recomputes them on demand with the full-spill kernel (mimgpu_download, mimgpu.cu).
 */
#pragma once

namespace mimgpu {

/* Knot-packed HBM layout of the fast variant (built once per upload by k_pack_knots):
 *   input record of knot t  (REC doubles):  [fs_t | tril(Lxx) | Lxu | tril(Luu) | Lx Lu | Fx Fu]   (GjCfg::sym_off)
 *   terminal record         (TREC doubles): [fs_T | tril(Lxx_T) | Lx_T]
 *   spill record of knot t  (SREC doubles): [Vxx_t or triu(Vxx_t) | Vx_t | K_t | k_t]   (GjCfg::TRI, vrow)
 * so that the backward sweep stages a knot with ONE bulk copy (the whole record) and the
 * forward sweep with TWO: [Lx Lu Fx Fu](t) ++ [fs](t+1) is contiguous across two input
 * records, and [K k](t) ++ [Vxx Vx](t+1) is contiguous across two spill records. */
template <int NX, int NU>
struct GjCfg {
  static constexpr int N = NX + NU;
  static constexpr int NTC = (N + 1 + 7) / 8;             /* column tiles incl. the gradient column N */
  static constexpr int RCT = N / 8, RTQ = (N % 8) / 2;    /* tile / lane-in-group of column N */
  /* The symmetric blocks Lxx, Luu, Lxx_T travel as their LOWER triangle, packed column by column (column j =
   * rows j .. n-1: contiguous in the Eigen column-major image, one memcpy per column), each block padded to an
   * even number of doubles: 417 instead of 498 doubles per UR5 knot -- 16 % fewer bytes over PCIe and from HBM.
   * Element (r, c), r >= c, of an n x n block sits at sym_off(n, r, c). */
#ifndef MIMGPU_GJ_SYMREC
#define MIMGPU_GJ_SYMREC 1
#endif
  static constexpr bool SYM = MIMGPU_GJ_SYMREC != 0;
  __host__ __device__ static constexpr int sym_size(int n) { return SYM ? ((n * (n + 1) / 2 + 1) & ~1) : n * n; }
  __host__ __device__ static constexpr int sym_off(int n, int i, int j) {   /* element (i, j) of a symmetric block */
    if (!SYM) return i + n * j;
    const int r = i > j ? i : j, c = i > j ? j : i;
    return c * n - c * (c - 1) / 2 + (r - c);
  }
  /* input record = backward stage */
  static constexpr int bFs = 0, bH = NX, bLxu = bH + sym_size(NX), bLuu = bLxu + NX * NU, bL = bLuu + sym_size(NU),
                       bF = bL + N, REC = bF + NX * N;
  static constexpr int TREC = NX + sym_size(NX) + NX, tFs = 0, tLxx = NX, tLx = NX + sym_size(NX);
  /* Vxx is symmetric: from NX = MIMGPU_GJ_TRI_MIN_NX up the spill keeps its upper triangle only, row i from the
   * even column i & ~1 on (every row segment starts 16-byte aligned and has an even length; 152 instead of 256
   * doubles for NX = 16): element (i, n), n >= vstart(i), sits at vrow(i) + n - vstart(i).  Measured on B200
   * (profiles/variants_r02_v2.log): a gain where the kernel is closer to the DRAM limit (NX = 16: 0.830 -> 0.766 ms),
   * neutral to slightly negative for the 12/6 and 8/4 shapes, which are issue / LSU-bound and pay for the
   * scalar triangle reads of the forward sweep -- those keep the dense rows. */
#ifndef MIMGPU_GJ_TRI_MIN_NX
#define MIMGPU_GJ_TRI_MIN_NX 16
#endif
  static constexpr bool TRI = NX >= MIMGPU_GJ_TRI_MIN_NX;
  __host__ __device__ static constexpr int vstart(int i) { return TRI ? (i & ~1) : 0; }
  __host__ __device__ static constexpr int vrow(int i) {
    if (!TRI) return i * NX;
    const int e = i & ~1;
    return e * NX - e * e / 2 + e + ((i & 1) ? NX - e : 0);
  }
  static constexpr int VTRI = vrow(NX);
  /* spill record */
  static constexpr int sV = 0, sVx = VTRI, sK = sVx + NX, sk = sK + NX * NU, SREC = sk + NU;
  /* forward stage: [Lx Lu | Fx Fu | fs+] ++ [K k | Vxx+ Vx+] */
  static constexpr int fL = 0, fF = N, fFs = fF + NX * N, FA = fFs + NX,
                       fK = FA, fk = fK + NX * NU, fV = fk + NU, fVx = fV + VTRI, STAGE_F = fVx + NX;
  /* (Vxx y)[i]: dense rows, or from the packed triangle -- columns n >= (i & ~1) come from row i, the others from
   * row n (element (n, i)); two partial sums over even / odd n in both cases */
  __device__ static __forceinline__ double tri_row_dot(const double* V, const double* y, int i) {
    double a0 = 0.0, a1 = 0.0;
    if constexpr (!TRI) {
#pragma unroll
      for (int n = 0; n < NX; n += 2) {
        const double2 v2 = *reinterpret_cast<const double2*>(V + NX * i + n);
        const double2 y2 = *reinterpret_cast<const double2*>(y + n);
        a0 = fma(v2.x, y2.x, a0);
        a1 = fma(v2.y, y2.y, a1);
      }
    } else {
      const int e = i & ~1, rb = vrow(i) - e;
#pragma unroll
      for (int n = 0; n < NX; n += 2) {
        const int cb = vrow(n) - n;                 /* n even: n & ~1 == n */
        const int cb1 = vrow(n + 1) - n;
        const double v0 = V[(n >= e) ? rb + n : cb + i];
        const double v1 = V[(n >= e) ? rb + n + 1 : cb1 + i];
        a0 = fma(v0, y[n], a0);
        a1 = fma(v1, y[n + 1], a1);
      }
    }
    return a0 + a1;
  }
  static constexpr int STAGE_B = REC;
  static constexpr int STAGE = (STAGE_B > STAGE_F ? STAGE_B : STAGE_F);
  /* per-warp shared memory: two stages, the forward pass' vectors, two mbarriers */
  static constexpr int oS0 = 0, oS1 = STAGE, oX = 2 * STAGE, oU = oX + NX, oXn = oU + NU, oLam = oXn + NX,
                       oBar = (oLam + NX + 1) & ~1, WARP = oBar + 2;
  static_assert(NX % 2 == 0 && NU % 2 == 0, "16-byte bulk copies need even block sizes");
  __host__ __device__ static size_t pack_doubles(int T) { return (size_t)T * REC + TREC; }
  __host__ __device__ static size_t spill_doubles(int T) { return (size_t)(T + 1) * SREC; }
};

/* 1/x for a well-scaled positive pivot: hardware seed + two Newton steps (no slow path; a
 * non-positive or non-finite pivot is reported through the status word, not used) */
__device__ __forceinline__ double rcp_pivot(double x) {
  double r;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x));
  double e = fma(-x, r, 1.0);
  r = fma(r, e, r);
  e = fma(-x, r, 1.0);
  return fma(r, e, r);
}

template <int NX, int NU>
__device__ __forceinline__ int small_backward_gj(const double* pk, double* sp, int T, double* sw, int lane,
                                                 double preg, double dreg, int is_feasible) {
  using C = GjCfg<NX, NU>;
  static_assert(NX % 4 == 0 && NU % 2 == 0, "k-tiles of 4, 16-byte staging");
  constexpr int N = C::N, NT = (N + 7) / 8, NTC = C::NTC, XT = (NX + 7) / 8, KT = NX / 4, KPT = (NX + 1 + 7) / 8;
  constexpr int KQ = NX / 8, TQ = (NX % 8) / 2;          /* Wt tile column / lane-in-group holding k' = NX */
  constexpr int RCT = C::RCT, RTQ = C::RTQ;
  constexpr unsigned FULL = 0xffffffffu;
  const int g = lane >> 2, tq = lane & 3;
  const bool use_preg = !isnan(preg), use_dreg = !isnan(dreg);

  /* one elected lane stages a knot with one TMA bulk copy that completes on the stage's mbarrier */
  uint64_t* bars = reinterpret_cast<uint64_t*>(sw + C::oBar);
  auto prefetch = [&](int t, int slot) {
    if (lane == 0) {
      uint64_t* bar = bars + slot;
      mbar_expect_tx(bar, (uint32_t)C::REC * 8u);
      tma_bulk_g2s(sw + (slot ? C::oS1 : C::oS0), pk + (size_t)t * C::REC, C::REC * 8, bar);
    }
  };
  /* a failing problem leaves with no bulk copy in flight */
  auto drain = [&](int t) {
    const int j = T - 1 - t;
    if (t > 0) mbar_wait(bars + ((j + 1) & 1), (uint32_t)((j + 1) >> 1) & 1u);
  };
  if (lane == 0) {
    mbar_init(bars, 1);
    mbar_init(bars + 1, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncwarp();
  prefetch(T - 1, 0);

  /* m[it][nt][e] = M[8 it + g][8 nt + 2 tq + e]; after a knot its (it, nt < XT) tiles hold Vxx,
   * vxg[it] = Vx[8 it + g] (replicated over the four lanes of a group) */
  double m[NT][NTC][2];
  double vxg[XT];

  /* V.fs + gradient column -> vxg (all lanes of a group), + preg, raiseIfNaN; fsp: gap vector */
  auto close_knot = [&](const double* fsp, bool& ok) {
#pragma unroll
    for (int it = 0; it < XT; ++it) {
      double part = 0.0;
#pragma unroll
      for (int nt = 0; nt < XT; ++nt) {
        const int i = 8 * it + g, n = 8 * nt + 2 * tq;
        if (i < NX && n < NX) {
          double v0 = m[it][nt][0], v1 = m[it][nt][1];
          if (use_preg) { if (i == n) v0 += preg; if (i == n + 1) v1 += preg; }
          m[it][nt][0] = v0; m[it][nt][1] = v1;
          ok = ok && (fabs(v0) < 1e30) && (fabs(v1) < 1e30);   /* false for NaN / inf too (raiseIfNaN) */
          if (!is_feasible) {
            const double2 f2 = *reinterpret_cast<const double2*>(fsp + n);
            part = fma(v0, f2.x, part);
            part = fma(v1, f2.y, part);
          }
        }
      }
      if (tq == RTQ) part += m[it][RCT][0];
      part += __shfl_xor_sync(FULL, part, 1);
      part += __shfl_xor_sync(FULL, part, 2);
      vxg[it] = part;
      ok = ok && ((fabs(part) < 1e30) || (8 * it + g >= NX));
    }
  };
  auto spill_value = [&](int t) {
#pragma unroll
    for (int it = 0; it < XT; ++it) {
      const int i = 8 * it + g;
      if (i < NX) {
#pragma unroll
        const int e = C::vstart(i), rb = C::vrow(i) - e;
#pragma unroll
        for (int nt = 0; nt < XT; ++nt) {
          const int n = 8 * nt + 2 * tq;
          if (n < NX && n >= e)                      /* TRI: upper triangle (from the even column i & ~1) only */
            *reinterpret_cast<double2*>(sp + (size_t)t * C::SREC + C::sV + rb + n) =
                make_double2(m[it][nt][0], m[it][nt][1]);
        }
        if (tq == 0) sp[(size_t)t * C::SREC + C::sVx + i] = vxg[it];
      }
    }
  };

  /* ---- terminal knot (src/ddp.cpp:268-277): Vxx_T = Lxx_T (+ preg), Vx_T = Lx_T + Vxx_T fs_T.
   * The fragment holds the TRANSPOSE of Lxx_T, because step 1 reads Vxx'[m][k'] from element
   * (k', m): the first sweep then uses Lxx_T exactly as the reference does. */
  {
#pragma unroll
    for (int it = 0; it < NT; ++it)
#pragma unroll
      for (int nt = 0; nt < NTC; ++nt) { m[it][nt][0] = 0.0; m[it][nt][1] = 0.0; }
#pragma unroll
    for (int it = 0; it < XT; ++it)
#pragma unroll
      for (int nt = 0; nt < XT; ++nt) {
        const int i = 8 * it + g, n = 8 * nt + 2 * tq;
        if (i < NX && n < NX) {
          const double* LT = pk + (size_t)T * C::REC + C::tLxx;
          m[it][nt][0] = LT[C::sym_off(NX, n, i)]; m[it][nt][1] = LT[C::sym_off(NX, n + 1, i)];
        }
      }
#pragma unroll
    for (int it = 0; it < XT; ++it) {
      const int i = 8 * it + g;
      if (NTC > RCT) m[it][RCT][0] = (tq == RTQ && i < NX) ? pk[(size_t)T * C::REC + C::tLx + i] : m[it][RCT][0];
    }
    bool ok = true;
    close_knot(pk + (size_t)T * C::REC + C::tFs, ok);
    spill_value(T);
  }

  /* Per-lane table of where each element of the M fragments sits in a staged input record
   * (knot-invariant): 10-bit offsets packed three to a word, one validity bit per element.
   * Element (i, n) of [[Lxx Lxu | Lx], [Lxu^T Luu | Lu]]: rows / columns >= N (+1) are padding. */
  static_assert(C::REC < 1024, "10-bit record offsets");
  static_assert(NT * NTC * 2 <= 32, "one validity bit per fragment element");
  uint32_t hoff[(NT * NTC * 2 + 2) / 3];
  uint32_t hmask = 0;
#pragma unroll
  for (int w = 0; w < (NT * NTC * 2 + 2) / 3; ++w) hoff[w] = 0;
#pragma unroll
  for (int it = 0; it < NT; ++it)
#pragma unroll
    for (int nt = 0; nt < NTC; ++nt)
#pragma unroll
      for (int e = 0; e < 2; ++e) {
        const int k = (it * NTC + nt) * 2 + e;
        const int i = 8 * it + g, n = 8 * nt + 2 * tq + e;
        int off = 0;
        bool valid = false;
        if (i < NX && n < NX) { off = C::bH + C::sym_off(NX, i, n); valid = true; }                   /* Lxx */
        else if (i < NX && n < N) { off = C::bLxu + i + NX * (n - NX); valid = true; }               /* Lxu */
        else if (i < N && n < NX) { off = C::bLxu + n + NX * (i - NX); valid = true; }               /* Lxu^T */
        else if (i < N && n < N) { off = C::bLuu + C::sym_off(NU, i - NX, n - NX); valid = true; }   /* Luu */
        else if (i < N && n == N) { off = C::bL + i; valid = true; }                                 /* Lx | Lu */
        hoff[k / 3] |= (uint32_t)off << (10 * (k % 3));
        hmask |= (valid ? 1u : 0u) << k;
      }

  for (int t = T - 1; t >= 0; --t) {
    const int j = T - 1 - t, s = j & 1;          /* load sequence number -> stage, mbarrier phase */
    double* st = sw + (s ? C::oS1 : C::oS0);
    if (t > 0) prefetch(t - 1, s ^ 1);
    mbar_wait(bars + s, (uint32_t)(j >> 1) & 1u);

    /* ---- A fragments of [Fx Fu]^T, shared by steps 1 and 2: F[kperm(kt)][8 nt + g] */
    double fa[NT][KT];
#pragma unroll
    for (int nt = 0; nt < NT; ++nt) {
      const int n = 8 * nt + g;
#pragma unroll
      for (int kt = 0; kt < KT; ++kt) fa[nt][kt] = (n < N) ? st[C::bF + kperm<NX>(kt, tq) + NX * n] : 0.0;
    }
    /* ---- step 1: Wt[n][k'] = sum_mm F[mm][n] Vp[mm][k'],  Vp = [Vxx' | Vx' | 0] */
    double wt[NT][KPT][2];
#pragma unroll
    for (int nt = 0; nt < NT; ++nt)
#pragma unroll
      for (int kp = 0; kp < KPT; ++kp) { wt[nt][kp][0] = 0.0; wt[nt][kp][1] = 0.0; }
#pragma unroll
    for (int kt = 0; kt < KT; ++kt) {
      const int b8 = kt >> 1;
      const bool full = (8 * b8 + 8 <= NX);
      /* Vx'[kperm(kt, tq)]: rows of a tile live at g, so fetch from the lane group that owns it */
      const int mm = kperm<NX>(kt, tq);
      const double vxm = __shfl_sync(FULL, vxg[b8], 4 * (mm & 7) + tq);
#pragma unroll
      for (int kp = 0; kp < KPT; ++kp) {
        const int kk = 8 * kp + g;
        double vb = 0.0;
        if (kp < XT) {
          if (full) {
            vb = m[kp][b8][kt & 1];
          } else {  /* trailing 4-block: column 8 b8 + {0,2,1,3}[tq] */
            const double up = __shfl_sync(FULL, m[kp][b8][1], (lane - 2) & 31);
            vb = (tq < 2) ? m[kp][b8][0] : up;
          }
        }
        vb = (kk < NX) ? vb : (kk == NX ? vxm : 0.0);
#pragma unroll
        for (int nt = 0; nt < NT; ++nt) dmma884(wt[nt][kp][0], wt[nt][kp][1], fa[nt][kt], vb);
      }
    }
    /* ---- M = [H | l]: Lxx Lxu / Lxu^T Luu (+ dreg), gradient column = [Lx; Lu] + F^T Vx'.
     * One predicated load per element through the lane's packed offset table (no divergence). */
#pragma unroll
    for (int it = 0; it < NT; ++it) {
      const int i = 8 * it + g;
      const double qv = __shfl_sync(FULL, wt[it][KQ][0], 4 * g + TQ);
#pragma unroll
      for (int nt = 0; nt < NTC; ++nt)
#pragma unroll
        for (int e = 0; e < 2; ++e) {
          const int k = (it * NTC + nt) * 2 + e;
          const int n = 8 * nt + 2 * tq + e;
          const int off = (int)((hoff[k / 3] >> (10 * (k % 3))) & 1023u);
          double v = ((hmask >> k) & 1u) ? st[off] : 0.0;
          if (it == nt && it >= NX / 8 && use_dreg && i == n && i >= NX && i < N) v += dreg;   /* Luu + dreg */
          if (nt == RCT && n == N && i < N) v += qv;                                           /* Qx | Qu */
          m[it][nt][e] = v;
        }
    }
    /* ---- step 2: M[i][n] += sum_k F[k][i] W[k][n]  (src/ddp.cpp:295-314) */
#pragma unroll
    for (int kt = 0; kt < KT; ++kt) {
      const int b8 = kt >> 1;
      const bool full = (8 * b8 + 8 <= NX);
#pragma unroll
      for (int nt = 0; nt < NT; ++nt) {
        double wb;
        if (full) {
          wb = wt[nt][b8][kt & 1];
        } else {
          const double up = __shfl_sync(FULL, wt[nt][b8][1], (lane - 2) & 31);
          wb = (tq < 2) ? wt[nt][b8][0] : up;
        }
#pragma unroll
        for (int it = 0; it < NT; ++it) dmma884(m[it][nt][0], m[it][nt][1], fa[it][kt], wb);
      }
    }
    /* ---- gains: Gauss-Jordan on the u-rows (computeGains, src/ddp.cpp:401-429) */
    /* The pivot row itself is left unscaled while the later pivots run (row scaling commutes
     * with the remaining eliminations) and is multiplied by 1/pivot once at the end. */
    double rinv[NT];
#pragma unroll
    for (int it = 0; it < NT; ++it) rinv[it] = 1.0;
#pragma unroll
    for (int a = NX; a < N; ++a) {
      const int ta = a >> 3, ga = a & 7, tqa = (a & 7) >> 1, ea = a & 1;
      const double piv = __shfl_sync(FULL, m[ta][ta][ea], 4 * ga + tqa);
      if (piv <= 0.0) { drain(t); return MIMGPU_STATUS_LLT | (t << 8); }   /* warp-uniform: piv is a broadcast */
      const double inv = rcp_pivot(piv);
      double r[NTC][2];
#pragma unroll
      for (int nt = 0; nt < NTC; ++nt) {
        r[nt][0] = __shfl_sync(FULL, m[ta][nt][0], 4 * ga + tq);
        r[nt][1] = __shfl_sync(FULL, m[ta][nt][1], 4 * ga + tq);
      }
#pragma unroll
      for (int it = 0; it < NT; ++it) {
        double f = __shfl_sync(FULL, m[it][ta][ea], 4 * g + tqa) * inv;
        if (it == ta) {
          const bool prow = (g == ga);
          f = prow ? 0.0 : f;
          rinv[it] = prow ? inv : rinv[it];
        }
#pragma unroll
        for (int nt = 0; nt < NTC; ++nt) {
          m[it][nt][0] = fma(-f, r[nt][0], m[it][nt][0]);
          m[it][nt][1] = fma(-f, r[nt][1], m[it][nt][1]);
        }
      }
    }
    /* ---- spill K (row-major NU x NX) and k from the u-rows */
#pragma unroll
    for (int it = NX / 8; it < NT; ++it) {
      const int i = 8 * it + g;
      if (i >= NX && i < N) {
#pragma unroll
        for (int nt = 0; nt < XT; ++nt) {
          const int n = 8 * nt + 2 * tq;
          if (n < NX)
            *reinterpret_cast<double2*>(sp + (size_t)t * C::SREC + C::sK + (i - NX) * NX + n) =
                make_double2(m[it][nt][0] * rinv[it], m[it][nt][1] * rinv[it]);
        }
        if (tq == RTQ) sp[(size_t)t * C::SREC + C::sk + i - NX] = m[it][RCT][0] * rinv[it];
      }
    }
    /* ---- Vxx + preg, Vx = Qx - K^T Qu + Vxx fs_t, raiseIfNaN  (src/ddp.cpp:319-345) */
    bool ok = true;
    close_knot(st + C::bFs, ok);
    spill_value(t);
    if (__any_sync(FULL, !ok)) { drain(t); return MIMGPU_STATUS_NONFINITE | (t << 8); }
    __syncwarp();                                /* every lane is done reading stage s */
  }
  return 0;
}

/* ------------------------------------------------------------------- forward */
/* SolverSQP::forwardPass + checkKKTConditions + gap norm (src/sqp.cpp:270-305, 244-268,
 * 229-236) of the fast variant: same arithmetic as small_forward (kernels_small.cuh), knots
 * staged by TMA bulk copies on the two mbarriers the backward sweep used (load sequence
 * numbers continue at T).  Lane i < NX owns state row i, lane a < NU control row a, lane
 * c < N column c of [Fx Fu] for the stationarity residual. */
template <int NX, int NU>
__device__ __forceinline__ void small_forward_gj(const Prob& p, const double* pk, const double* sp, double* sw,
                                                 int lane, double mu_dynamic) {
  using C = GjCfg<NX, NU>;
  constexpr int N = C::N;
  const int T = p.T;
  double* sX = sw + C::oX;
  double* sU = sw + C::oU;
  double* sXn = sw + C::oXn;
  double* sLam = sw + C::oLam;
  uint64_t* bars = reinterpret_cast<uint64_t*>(sw + C::oBar);
  const bool gx = lane < NX, gu = lane < NU, gc = lane < N;
  const int lx = gx ? lane : 0, lu = gu ? lane : 0, lc = gc ? lane : 0;

  auto prefetch = [&](int t, int slot) {
    if (lane == 0) {
      double* st = sw + (slot ? C::oS1 : C::oS0);
      uint64_t* bar = bars + slot;
      mbar_expect_tx(bar, (uint32_t)C::STAGE_F * 8u);
      tma_bulk_g2s(st + C::fL, pk + (size_t)t * C::REC + C::bL, C::FA * 8, bar);
      tma_bulk_g2s(st + C::fK, sp + (size_t)t * C::SREC + C::sK, C::SREC * 8, bar);
    }
  };
  /* the backward sweep's generic-proxy stores of Vxx Vx K k -> TMA (async proxy) reads */
  fence_proxy_async();
  __syncwarp();
  prefetch(0, T & 1);

  double xg = 0.0, ug = 0.0, gap = 0.0, kkt = 0.0;
  /* knot 0: dx_0 (member, stays 0), lag_mul_0 = Vxx_0 (dx_0 - fs_0) + Vx_0 */
  double fs_cur = pk[C::bFs + lx];
  double x_cur = p.dx[lx];
  if (gx) sX[lane] = x_cur - fs_cur;
  __syncwarp();
  double lam;
  {
    lam = C::tri_row_dot(sp + C::sV, sX, lx) + sp[C::sVx + lx];
    if (gx) p.lag[lane] = lam;
  }
  __syncwarp();
  if (gx) sX[lane] = x_cur;
  __syncwarp();

  for (int t = 0; t < T; ++t) {
    const int j = T + t, s = j & 1;
    const double* st = sw + (s ? C::oS1 : C::oS0);
    if (t + 1 < T) prefetch(t + 1, s ^ 1);
    mbar_wait(bars + s, (uint32_t)(j >> 1) & 1u);
    if (gx) { xg += fabs(x_cur); gap += fabs(fs_cur); kkt = nmax(kkt, fabs(fs_cur)); }
    /* du = -K dx - k  (src/sqp.cpp:285-286); idle lanes are predicated off (no shared-memory wavefronts) */
    if (gu) {
      double a0 = 0.0, a1 = 0.0;
#pragma unroll
      for (int i = 0; i < NX; i += 2) {
        const double2 k2 = *reinterpret_cast<const double2*>(st + C::fK + NX * lu + i);
        const double2 x2 = *reinterpret_cast<const double2*>(sX + i);
        a0 = fma(k2.x, x2.x, a0);
        a1 = fma(k2.y, x2.y, a1);
      }
      const double v = -(a0 + a1) - st[C::fk + lu];
      sU[lane] = v; p.du[(size_t)t * NU + lane] = v; ug += fabs(v);
    }
    __syncwarp();
    /* dx+ = fs+ + Fu du + Fx dx  (src/sqp.cpp:287-289) */
    double fs_n = 0.0, x_n = 0.0;
    if (gx) {
      fs_n = st[C::fFs + lx];
      double a0 = 0.0, a1 = 0.0;
#pragma unroll
      for (int a = 0; a < NU; a += 2) {
        const double2 u2 = *reinterpret_cast<const double2*>(sU + a);
        a0 = fma(st[C::fF + NX * NX + lx + NX * a], u2.x, a0);
        a1 = fma(st[C::fF + NX * NX + lx + NX * (a + 1)], u2.y, a1);
      }
      double b0 = 0.0, b1 = 0.0;
#pragma unroll
      for (int jj = 0; jj < NX; jj += 2) {
        const double2 x2 = *reinterpret_cast<const double2*>(sX + jj);
        b0 = fma(st[C::fF + lx + NX * jj], x2.x, b0);
        b1 = fma(st[C::fF + lx + NX * (jj + 1)], x2.y, b1);
      }
      x_n = (fs_n + (a0 + a1)) + (b0 + b1);
      p.dx[(size_t)(t + 1) * NX + lane] = x_n; sXn[lane] = x_n - fs_n;
    }
    __syncwarp();
    /* lag_mul+ = Vxx+ (dx+ - fs+) + Vx+  (src/sqp.cpp:282-284, 296-298); Vxx+ arrives as its upper
     * triangle (symmetric: the reference averages Vxx and its transpose, src/ddp.cpp:329) */
    double lam_n = 0.0;
    if (gx) {
      lam_n = C::tri_row_dot(st + C::fV, sXn, lx) + st[C::fVx + lx];
      p.lag[(size_t)(t + 1) * NX + lane] = lam_n; sLam[lane] = lam_n;
    }
    __syncwarp();
    /* stationarity rows of knot t  (src/sqp.cpp:249-258) */
    if (gc) {
      double a0 = 0.0, a1 = 0.0;
#pragma unroll
      for (int k = 0; k < NX; k += 2) {
        const double2 f2 = *reinterpret_cast<const double2*>(st + C::fF + NX * lc + k);
        const double2 l2 = *reinterpret_cast<const double2*>(sLam + k);
        a0 = fma(f2.x, l2.x, a0);
        a1 = fma(f2.y, l2.y, a1);
      }
      double v = st[C::fL + lc] + (a0 + a1);
      if (gx) v -= lam;
      kkt = nmax(kkt, fabs(v));
    }
    if (gx) sX[lane] = x_n;
    x_cur = x_n; fs_cur = fs_n; lam = lam_n;
    __syncwarp();
  }
  if (gx) {
    xg += fabs(x_cur); gap += fabs(fs_cur);
    kkt = nmax(kkt, fabs(fs_cur));
    kkt = nmax(kkt, fabs(pk[(size_t)T * C::REC + C::tLx + lane] - lam));          /* src/sqp.cpp:259-262 */
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

}  // namespace mimgpu
