/*
 * kernels_generic.cuh -- shape-generic CUDA kernels of the batched OCP-QP engine.
 *
 * One CTA owns one problem and sweeps the horizon inside the kernel; Vxx, Vx and
 * the per-knot Q blocks live in shared memory, the per-knot outputs that the
 * reference keeps as solver members (Vxx_, Vx_, K_, k_, Qu_, Quuk_, Quu_llt_,
 * Vxx_fs_) are written to HBM once per knot.  Works for every (T, nx, nu, ng, ngT);
 * the hot shapes additionally get specialised kernels (kernels_small.cuh).
 *
 * The bodies are written as sequences of data-parallel phases (MG_PAR) separated by
 * CTA barriers (MG_SYNC).  With -DMIMGPU_EMU the same source compiles for the host
 * with the phases run as serial loops: tests/emu uses that to debug index logic in
 * the GPU-less build container.  The emulation is TEST INFRASTRUCTURE, it is never
 * linked into libmimgpu.so and there is no CPU fallback in the product.
 *
 * Reference functions restated (paths relative to the reference tree):
 *   SolverDDP::backwardPass / computeGains      src/ddp.cpp:265-348, 401-429
 *   SolverSQP::computeDirection / forwardPass   src/sqp.cpp:224-242, 270-305
 *   SolverSQP::checkKKTConditions               src/sqp.cpp:244-268
 *   SolverFDDP::updateExpectedImprovement       src/fddp.cpp:256-279
 *   SolverCSQP::computeDirection and callees    src/csqp.cpp:454-517, 170-193, 519-582,
 *                                               630-669, 671-898, 1025-1108, 1196-1340
 *   SolverCSQP::checkKKTConditions              src/csqp.cpp:584-628
 *   SolverCSQP::calc (norms)                    src/csqp.cpp:417-451
 */
#pragma once
#include <math.h>
#include <stdint.h>

#include "../../include/mimgpu.h"

#ifdef MIMGPU_EMU
#define MG_DEV inline
#define MG_PAR(e, n) for (int e = 0; e < (n); ++e)
#define MG_SYNC() ((void)0)
#define MG_SINGLE if (true)
#define MG_INF INFINITY
#else
#define MG_DEV __device__ __forceinline__
#define MG_PAR(e, n) for (int e = threadIdx.x; e < (n); e += blockDim.x)
#define MG_SYNC() __syncthreads()
#define MG_SINGLE if (threadIdx.x == 0)
#define MG_INF CUDART_INF
#include <math_constants.h>
#endif

namespace mimgpu {

/* Everything a kernel needs; device pointers, problem-major SoA (include/mimgpu.h). */
struct KArgs {
  mimgpu_shape sh;
  const double *Fx, *Fu, *Lxx, *Lxu, *Luu, *Lx, *Lu, *fs, *LxxT, *LxT;
  const double *Gx, *Gu, *g, *glb, *gub, *GxT, *gT, *glbT, *gubT, *cost;
  double *dx, *du, *lag, *dxt, *dut, *y, *z, *rho;
  double *Vxx, *Vx, *K, *kff, *Qu, *Quuk, *L, *Vxx_fs;
  double* scal;                 /* [MIMGPU_S_COUNT][B] */
  int32_t *qp_iters, *status, *converged;
  const double *preg_arr, *dreg_arr; /* optional per-problem regularisation */
  const int32_t* active;        /* optional subset: block -> problem */
  int32_t n_active;             /* number of blocks that have a problem */
  int32_t stage_fx;             /* 1: Fx_t is staged in shared memory */
  const double* pack;           /* knot-packed copy of the inputs (kernels_small_gj.cuh), or null */
  double* spill;                /* knot-packed Vxx Vx K k of the fast SQP kernel, or null */
  int32_t first;                /* first problem of the slice when `active` is null (chunked host path) */
  int32_t pad_;
};

/* One problem's view. */
struct Prob {
  int b, B, T, nx, nu, ng, ngT;
  const double *Fx, *Fu, *Lxx, *Lxu, *Luu, *Lx, *Lu, *fs, *LxxT, *LxT;
  const double *Gx, *Gu, *g, *glb, *gub, *GxT, *gT, *glbT, *gubT;
  double cost;
  double *dx, *du, *lag, *dxt, *dut, *y, *z, *rho;
  double *Vxx, *Vx, *K, *kff, *Qu, *Quuk, *L, *Vxx_fs;
  double* scal;
};

MG_DEV Prob bind(const KArgs& a, int b) {
  Prob p;
  const size_t T = a.sh.T, nx = a.sh.nx, nu = a.sh.nu, ng = a.sh.ng, ngT = a.sh.ngT;
  const size_t nc = T * ng + ngT, sb = (size_t)b;
  p.b = b; p.B = a.sh.B; p.T = (int)T; p.nx = (int)nx; p.nu = (int)nu; p.ng = (int)ng; p.ngT = (int)ngT;
  p.Fx = a.Fx + sb * T * nx * nx;   p.Fu = a.Fu + sb * T * nx * nu;
  p.Lxx = a.Lxx + sb * T * nx * nx; p.Lxu = a.Lxu + sb * T * nx * nu;
  p.Luu = a.Luu + sb * T * nu * nu; p.Lx = a.Lx + sb * T * nx; p.Lu = a.Lu + sb * T * nu;
  p.fs = a.fs + sb * (T + 1) * nx;  p.LxxT = a.LxxT + sb * nx * nx; p.LxT = a.LxT + sb * nx;
  p.Gx = ng ? a.Gx + sb * T * ng * nx : nullptr; p.Gu = ng ? a.Gu + sb * T * ng * nu : nullptr;
  p.g = ng ? a.g + sb * T * ng : nullptr; p.glb = ng ? a.glb + sb * T * ng : nullptr;
  p.gub = ng ? a.gub + sb * T * ng : nullptr;
  p.GxT = ngT ? a.GxT + sb * ngT * nx : nullptr; p.gT = ngT ? a.gT + sb * ngT : nullptr;
  p.glbT = ngT ? a.glbT + sb * ngT : nullptr; p.gubT = ngT ? a.gubT + sb * ngT : nullptr;
  p.cost = a.cost ? a.cost[b] : 0.0;
  p.dx = a.dx + sb * (T + 1) * nx; p.du = a.du + sb * T * nu; p.lag = a.lag + sb * (T + 1) * nx;
  p.dxt = a.dxt + sb * (T + 1) * nx; p.dut = a.dut + sb * T * nu;
  p.y = a.y + sb * nc; p.z = a.z + sb * nc; p.rho = a.rho + sb * nc;
  p.Vxx = a.Vxx + sb * (T + 1) * nx * nx; p.Vx = a.Vx + sb * (T + 1) * nx;
  p.K = a.K + sb * T * nu * nx; p.kff = a.kff + sb * T * nu; p.Qu = a.Qu + sb * T * nu;
  p.Quuk = a.Quuk + sb * T * nu; p.L = a.L + sb * T * nu * nu; p.Vxx_fs = a.Vxx_fs + sb * T * nx;
  p.scal = a.scal;
  return p;
}

/* Shared-memory carve-up (in doubles); same function on host (sizing) and device. */
struct Smem {
  int ldx;  /* leading dimension of the staged Fx / Fu copies (odd: conflict-free transposed reads) */
  int V, W, FuTV, Qxu, Quu, L, K, Fx, Fu;
  int Vx, tv, Qx, Qu, kff, fs, x, u, lam, lamn, d, rho, c, dz, red;
  int total;
};
#ifndef MIMGPU_EMU
__host__ __device__
#endif
inline Smem smem_layout(int nx, int nu, int ng, int ngT, int stage_fx) {
  Smem s;
  int o = 0;
  const int ngm = ng > ngT ? ng : ngT;
  s.ldx = nx | 1;
#define MG_TAKE(n) (o += (((n) + 1) & ~1), o - (((n) + 1) & ~1)) /* 16-byte aligned slices */
  s.V = MG_TAKE(nx * nx); s.W = MG_TAKE(nx * nx); s.FuTV = MG_TAKE(nu * nx); s.Qxu = MG_TAKE(nx * nu);
  s.Quu = MG_TAKE(nu * nu); s.L = MG_TAKE(nu * nu); s.K = MG_TAKE(nu * nx);
  s.Fx = MG_TAKE(stage_fx ? s.ldx * nx : 0); s.Fu = MG_TAKE(s.ldx * nu);
  s.Vx = MG_TAKE(nx); s.tv = MG_TAKE(nx); s.Qx = MG_TAKE(nx); s.Qu = MG_TAKE(nu); s.kff = MG_TAKE(nu + 1);
  s.fs = MG_TAKE(nx); s.x = MG_TAKE(nx); s.u = MG_TAKE(nu); s.lam = MG_TAKE(nx); s.lamn = MG_TAKE(nx);
  s.d = MG_TAKE(ngm); s.rho = MG_TAKE(ngm); s.c = MG_TAKE(ngm); s.dz = MG_TAKE(ngm); s.red = MG_TAKE(64);
#undef MG_TAKE
  s.total = o;
  return s;
}

/* ----------------------------------------------------------------- scalars */
MG_DEV bool raise_if_nan(double v) { return isnan(v) || isinf(v) || v >= 1e30; }
/* max with sticky NaN (any NaN entry makes the norm NaN, as in the oracle) */
MG_DEV double nmax(double m, double a) { return (a > m || isnan(a)) ? a : m; }
/* std::max / std::min with their exact NaN behaviour (src/csqp.cpp:522-523) */
MG_DEV double std_max(double a, double b) { return (a < b) ? b : a; }
MG_DEV double std_min(double a, double b) { return (b < a) ? b : a; }

#ifdef MIMGPU_EMU
MG_DEV double cta_max(double v, double*) { return v; }
MG_DEV double cta_sum(double v, double*) { return v; }
MG_DEV int cta_or(int v, double*) { return v; }
#else
/* CTA-wide reductions; every thread gets the result. `red` = 64 doubles of smem. */
__device__ __forceinline__ double cta_max(double v, double* red) {
  for (int o = 16; o; o >>= 1) v = nmax(v, __shfl_xor_sync(0xffffffffu, v, o));
  if (blockDim.x > 32) {
    const int w = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
    __syncthreads();
    if ((threadIdx.x & 31) == 0) red[w] = v;
    __syncthreads();
    v = red[0];
    for (int i = 1; i < nw; ++i) v = nmax(v, red[i]);
  }
  return v;
}
__device__ __forceinline__ double cta_sum(double v, double* red) {
  for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  if (blockDim.x > 32) {
    const int w = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
    __syncthreads();
    if ((threadIdx.x & 31) == 0) red[w] = v;
    __syncthreads();
    v = red[0];
    for (int i = 1; i < nw; ++i) v += red[i];
  }
  return v;
}
__device__ __forceinline__ int cta_or(int v, double*) { return __syncthreads_or(v); }
#endif

/* ------------------------------------------------------- Cholesky + solves */
/* Eigen::LLT<MatrixXd, Lower> semantics (src/ddp.cpp:413-419): lower triangle only,
 * fails iff a pivot is <= 0 (a NaN pivot does not fail).  L (nu x nu, col-major,
 * shared memory) holds a copy of Quu on entry.  Left-looking, one row per thread.
 * Returns 0 or 1 (uniform over the CTA). */
MG_DEV int llt_cta(double* L, int n) {
  for (int k = 0; k < n; ++k) {
    MG_PAR(i0, n - k) {
      const int i = k + i0;
      double v = L[i + n * k];
      for (int j = 0; j < k; ++j) v -= L[i + n * j] * L[k + n * j];
      L[i + n * k] = v;
    }
    MG_SYNC();
    const double x = L[k + n * k];
    MG_SYNC();
    if (x <= 0.0) return 1;
    const double r = sqrt(x);
    MG_PAR(i0, n - k) {
      const int i = k + i0;
      L[i + n * k] = (i == k) ? r : L[i + n * k] / r;
    }
    MG_SYNC();
  }
  return 0;
}

/* LLT::solveInPlace on `nrhs` right-hand sides, rhs c = b[c + inc*row], one rhs per
 * thread (serial substitution in the reference's order). */
MG_DEV void llt_solve_multi(const double* L, int n, double* b, int nrhs, int inc) {
  MG_PAR(c, nrhs) {
    double* x = b + c;
    for (int i = 0; i < n; ++i) {
      double v = x[i * inc];
      for (int j = 0; j < i; ++j) v -= L[i + n * j] * x[j * inc];
      x[i * inc] = v / L[i + n * i];
    }
    for (int i = n - 1; i >= 0; --i) {
      double v = x[i * inc];
      for (int j = i + 1; j < n; ++j) v -= L[j + n * i] * x[j * inc];
      x[i * inc] = v / L[i + n * i];
    }
  }
}

/* C(i, j) = sum_k A[k + lda * i] * B[k + ldb * j]  (i < ni, j < nj), handed to store(i, j, sum).
 * Each thread owns the 2 x 4 elements (i | i + hi) x (j + q hj, q < 4): six loads feed eight FMAs
 * instead of two feeding one, lanes still walk consecutive i (same bank pattern as one element per
 * thread) and the B operands are warp broadcasts.  Every element keeps its own k-ascending sum,
 * so results are bit-identical to the plain loop. */
#define MG_TJ 8
template <class Store>
MG_DEV void tiled_tn(const double* A, int lda, int ni, const double* B, int ldb, int nj, int nk, Store store) {
  const int hi = (ni + 1) / 2, hj = (nj + MG_TJ - 1) / MG_TJ;
  MG_PAR(te, hi * hj) {
    const int i0 = te % hi, j0 = te / hi, i1 = i0 + hi;
    const bool ri = i1 < ni;
    const double *a0 = A + (size_t)lda * i0, *a1 = A + (size_t)lda * (ri ? i1 : i0);
    const double* b[MG_TJ];
    double c0[MG_TJ], c1[MG_TJ];
#pragma unroll
    for (int q = 0; q < MG_TJ; ++q) {
      const int j = j0 + q * hj;
      b[q] = B + (size_t)ldb * (j < nj ? j : j0);
      c0[q] = 0.0; c1[q] = 0.0;
    }
    for (int k = 0; k < nk; ++k) {
      const double x0 = a0[k], x1 = a1[k];
#pragma unroll
      for (int q = 0; q < MG_TJ; ++q) { const double y = b[q][k]; c0[q] += x0 * y; c1[q] += x1 * y; }
    }
#pragma unroll
    for (int q = 0; q < MG_TJ; ++q) {
      const int j = j0 + q * hj;
      if (j < nj) { store(i0, j, c0[q]); if (ri) store(i1, j, c1[q]); }
    }
  }
}

/* C(i, j) = base(i, j) + [sum_r G[r + ng * i] * (rho[r] * H[r + ng * j])] + sum_k A[i + lda * k] * B[k + ldb * j]
 * with the same 2 x 4 ownership; the bracket is skipped when G is null.  finish(i, j, gsum, sum)
 * combines and stores.  Summation orders per element are those of the plain loops. */
template <class Finish>
MG_DEV void tiled_nn_g(const double* A, int lda, int ni, const double* B, int ldb, int nj, int nk,
                       const double* G, const double* H, const double* rho, int ng, Finish finish) {
  const int hi = (ni + 1) / 2, hj = (nj + MG_TJ - 1) / MG_TJ;
  MG_PAR(te, hi * hj) {
    const int i0 = te % hi, j0 = te / hi, i1 = i0 + hi;
    const bool ri = i1 < ni;
    const int i1c = ri ? i1 : i0;
    int jc[MG_TJ];
    double g0[MG_TJ], g1[MG_TJ], c0[MG_TJ], c1[MG_TJ];
#pragma unroll
    for (int q = 0; q < MG_TJ; ++q) {
      const int j = j0 + q * hj;
      jc[q] = j < nj ? j : j0;
      g0[q] = 0.0; g1[q] = 0.0; c0[q] = 0.0; c1[q] = 0.0;
    }
    if (G) {
      const double *p0 = G + (size_t)ng * i0, *p1 = G + (size_t)ng * i1c;
      for (int r = 0; r < ng; ++r) {
        const double x0 = p0[r], x1 = p1[r], rr = rho[r];
#pragma unroll
        for (int q = 0; q < MG_TJ; ++q) { const double y = rr * H[r + (size_t)ng * jc[q]]; g0[q] += x0 * y; g1[q] += x1 * y; }
      }
    }
    for (int k = 0; k < nk; ++k) {
      const double x0 = A[i0 + (size_t)lda * k], x1 = A[i1c + (size_t)lda * k];
#pragma unroll
      for (int q = 0; q < MG_TJ; ++q) { const double y = B[k + (size_t)ldb * jc[q]]; c0[q] += x0 * y; c1[q] += x1 * y; }
    }
#pragma unroll
    for (int q = 0; q < MG_TJ; ++q) {
      const int j = j0 + q * hj;
      if (j < nj) { finish(i0, j, g0[q], c0[q]); if (ri) finish(i1, j, g1[q], c1[q]); }
    }
  }
}

/* One right-hand side, column-oriented so that the CTA shares the work; every x[i]
 * still subtracts its terms in the reference's order (j ascending / descending). */
MG_DEV void llt_solve_one(const double* L, int n, double* x) {
#ifndef MIMGPU_EMU
  if (n <= 32) { /* one warp, x in registers: same operations per element, no CTA barrier per column */
    if (threadIdx.x < 32) {
      const int i = threadIdx.x;
      double xi = (i < n) ? x[i] : 0.0;
      for (int j = 0; j < n; ++j) {
        const double xj = __shfl_sync(0xffffffffu, xi, j) / L[j + n * j];
        if (i == j) xi = xj;
        if (i > j && i < n) xi -= L[i + n * j] * xj;
      }
      for (int j = n - 1; j >= 0; --j) {
        const double xj = __shfl_sync(0xffffffffu, xi, j) / L[j + n * j];
        if (i == j) xi = xj;
        if (i < j) xi -= L[j + n * i] * xj;
      }
      if (i < n) x[i] = xi;
    }
    MG_SYNC();
    return;
  }
#endif
  for (int j = 0; j < n; ++j) {
    MG_SINGLE { x[j] = x[j] / L[j + n * j]; }
    MG_SYNC();
    const double xj = x[j];
    MG_PAR(i0, n - j - 1) { const int i = j + 1 + i0; x[i] -= L[i + n * j] * xj; }
    MG_SYNC();
  }
  for (int j = n - 1; j >= 0; --j) {
    MG_SINGLE { x[j] = x[j] / L[j + n * j]; }
    MG_SYNC();
    const double xj = x[j];
    MG_PAR(i, j) { x[i] -= L[j + n * i] * xj; }
    MG_SYNC();
  }
}

/* ------------------------------------------------ Riccati backward sweeps */
enum { MODE_DDP = 0, MODE_EQ = 1, MODE_FULL = 2 };

struct RicPar {
  double preg, dreg, sigma;
  int is_feasible;
  int dbg = 0; /* debug checkpoints of kernels_ric.cuh (0 = off) */
};

/* MODE_DDP : SolverDDP::backwardPass (src/ddp.cpp:265-348) -- SQP and FDDP.
 * MODE_EQ  : SolverCSQP::backwardPass_without_constraints (src/csqp.cpp:804-898).
 * MODE_FULL: SolverCSQP::backwardPass (src/csqp.cpp:671-802).
 * Returns 0 or the status word (uniform over the CTA). */
template <int MODE>
MG_DEV int riccati_backward(const Prob& p, double* sm, const Smem& s, const RicPar& rp, int stage_fx) {
  const int T = p.T, nx = p.nx, nu = p.nu, ng = p.ng, ngT = p.ngT;
  const int nxx = nx * nx, nxu = nx * nu, nuu = nu * nu;
  const double preg = rp.preg, dreg = rp.dreg, sigma = rp.sigma;
  double *V = sm + s.V, *W = sm + s.W, *FuTV = sm + s.FuTV, *Qxu = sm + s.Qxu, *Quu = sm + s.Quu;
  double *L = sm + s.L, *K = sm + s.K, *sFx = sm + s.Fx, *sFu = sm + s.Fu;
  double *Vx = sm + s.Vx, *tv = sm + s.tv, *Qx = sm + s.Qx, *Qu = sm + s.Qu, *kff = sm + s.kff;
  double *sfs = sm + s.fs, *sd = sm + s.d, *srho = sm + s.rho;
  const int ldx = s.ldx;

  /* ---- terminal knot: src/ddp.cpp:268-277, src/csqp.cpp:827-832, :693-707 */
  if (MODE == MODE_FULL && ngT) {
    MG_PAR(r, ngT) {
      const double rho = p.rho[(size_t)T * ng + r];
      srho[r] = rho;
      double d = p.y[(size_t)T * ng + r];
      d -= rho * p.z[(size_t)T * ng + r];
      sd[r] = d;
    }
  }
  MG_PAR(i, nx) sfs[i] = p.fs[(size_t)T * nx + i];
  MG_SYNC();
  MG_PAR(e, nxx) {
    const int i = e % nx, j = e / nx;
    double v = p.LxxT[e];
    if (MODE == MODE_FULL) {
      if (i == j) v += sigma;
      if (ngT) {
        double acc = 0.0;
        for (int r = 0; r < ngT; ++r) acc += p.GxT[r + ngT * i] * (srho[r] * p.GxT[r + ngT * j]);
        v += acc;
      }
    }
    if (i == j && !isnan(preg)) v += preg;
    V[e] = v;
  }
  MG_SYNC();
  MG_PAR(i, nx) {
    double v = p.LxT[i];
    if (MODE == MODE_DDP) {
      if (!rp.is_feasible) {
        double acc = 0.0;
        for (int j = 0; j < nx; ++j) acc += V[i + nx * j] * sfs[j];
        v += acc;
      }
    } else if (MODE == MODE_FULL) {
      v -= sigma * p.dx[(size_t)T * nx + i];
      if (ngT) {
        double acc = 0.0;
        for (int r = 0; r < ngT; ++r) acc += p.GxT[r + ngT * i] * sd[r];
        v += acc;
      }
    }
    Vx[i] = v;
    p.Vx[(size_t)T * nx + i] = v;
  }
  MG_PAR(e, nxx) p.Vxx[(size_t)T * nxx + e] = V[e];
  MG_SYNC();

  for (int t = T - 1; t >= 0; --t) {
    const double* gFx = p.Fx + (size_t)t * nxx;
    const double* gFu = p.Fu + (size_t)t * nxu;
    const double* Gx = ng ? p.Gx + (size_t)t * ng * nx : nullptr;
    const double* Gu = ng ? p.Gu + (size_t)t * ng * nu : nullptr;
    const bool gx_on = (MODE == MODE_FULL) && ng && t > 0;   /* x0 is fixed: src/csqp.cpp:734,745,777 */
    const bool gu_on = (MODE == MODE_FULL) && ng;
    /* ---- stage this knot's dynamics and vectors */
    if (stage_fx) { MG_PAR(e, nxx) sFx[(e % nx) + ldx * (e / nx)] = gFx[e]; }
    MG_PAR(e, nxu) sFu[(e % nx) + ldx * (e / nx)] = gFu[e];
    if (MODE == MODE_DDP) {
      MG_PAR(i, nx) { tv[i] = Vx[i]; sfs[i] = p.fs[(size_t)t * nx + i]; }
    } else {
      MG_PAR(i, nx) sfs[i] = p.fs[(size_t)(t + 1) * nx + i];
    }
    if (MODE == MODE_FULL && ng) {
      MG_PAR(r, ng) {
        const double rho = p.rho[(size_t)t * ng + r];
        srho[r] = rho;
        double d = p.y[(size_t)t * ng + r];
        d -= rho * p.z[(size_t)t * ng + r];
        sd[r] = d;
      }
    }
    MG_SYNC();
    const double* Fx = stage_fx ? sFx : gFx;
    const int lfx = stage_fx ? ldx : nx;
    const double* Fu = sFu;
    if (MODE != MODE_DDP) { /* tmp_Vx = Vxx' fs_{t+1} + Vx'  (src/csqp.cpp:844-845, :720-721) */
      MG_PAR(i, nx) {
        double acc = 0.0;
        for (int j = 0; j < nx; ++j) acc += V[i + nx * j] * sfs[j];
        if (MODE == MODE_FULL) p.Vxx_fs[(size_t)t * nx + i] = acc;
        tv[i] = acc + Vx[i];
      }
      MG_SYNC();
    }
    /* ---- phase A: FxTVxx, FuTVxx, Qx, Qu  (src/ddp.cpp:289-302) */
    tiled_tn(Fx, lfx, nx, V, nx, nx, nx, [&](int i, int j, double acc) { W[i + nx * j] = acc; });
    tiled_tn(Fu, ldx, nu, V, nx, nx, nx, [&](int a, int j, double acc) { FuTV[a + nu * j] = acc; });
    MG_PAR(i, nx) {
      double v = p.Lx[(size_t)t * nx + i];
      if (MODE == MODE_FULL) {
        v -= sigma * p.dx[(size_t)t * nx + i];
        if (gx_on) {
          double acc = 0.0;
          for (int r = 0; r < ng; ++r) acc += Gx[r + ng * i] * sd[r];
          v += acc;
        }
      }
      double acc = 0.0;
      for (int k = 0; k < nx; ++k) acc += Fx[k + lfx * i] * tv[k];
      Qx[i] = v + acc;
    }
    MG_PAR(a, nu) {
      double v = p.Lu[(size_t)t * nu + a];
      if (MODE == MODE_FULL) {
        v -= sigma * p.du[(size_t)t * nu + a];
        if (gu_on) {
          double acc = 0.0;
          for (int r = 0; r < ng; ++r) acc += Gu[r + ng * a] * sd[r];
          v += acc;
        }
      }
      double acc = 0.0;
      for (int k = 0; k < nx; ++k) acc += Fu[k + ldx * a] * tv[k];
      Qu[a] = v + acc;
    }
    MG_SYNC();
    /* ---- phase B: Qxx (into V), Quu, Qxu  (src/ddp.cpp:295-314, src/csqp.cpp:743-780) */
    {
      const double* lxx = p.Lxx + (size_t)t * nxx;
      tiled_nn_g(W, nx, nx, Fx, lfx, nx, nx, gx_on ? Gx : nullptr, Gx, srho, ng,
                 [&](int i, int j, double gacc, double acc) {
                   double v = lxx[i + nx * j];
                   if (MODE == MODE_FULL) {
                     if (i == j) v += sigma;
                     if (gx_on) v += gacc;
                   }
                   V[i + nx * j] = v + acc;
                 });
    }
    MG_PAR(e, nuu) {
      const int a = e % nu, c = e / nu;
      double v = p.Luu[(size_t)t * nuu + e];
      if (MODE == MODE_FULL && a == c) v += sigma;
      double acc = 0.0;
      for (int k = 0; k < nx; ++k) acc += FuTV[a + nu * k] * Fu[k + ldx * c];
      v += acc;
      if (gu_on) {
        double acc2 = 0.0;
        for (int r = 0; r < ng; ++r) acc2 += Gu[r + ng * a] * (srho[r] * Gu[r + ng * c]);
        v += acc2;
      }
      if (a == c && !isnan(dreg)) v += dreg;
      Quu[e] = v;
      L[e] = v;
    }
    {
      const double* lxu = p.Lxu + (size_t)t * nxu;
      tiled_nn_g(W, nx, nx, Fu, ldx, nu, nx, gx_on ? Gx : nullptr, Gu, srho, ng,
                 [&](int i, int a, double gacc, double acc) {
                   double v = lxu[i + nx * a];
                   if (gx_on) v += gacc;
                   Qxu[i + nx * a] = v + acc;
                 });
    }
    MG_SYNC();
    /* ---- computeGains (src/ddp.cpp:401-429) */
    if (nu > 0) {
      if (llt_cta(L, nu)) return MIMGPU_STATUS_LLT | (t << 8);
      MG_PAR(e, nxu) { const int i = e % nx, a = e / nx; K[a * nx + i] = Qxu[e]; } /* K = Qxu^T, row-major */
      MG_PAR(a, nu) kff[a] = Qu[a];
      MG_SYNC();
      llt_solve_multi(L, nu, K, nx, nx);
      llt_solve_one(L, nu, kff);
      MG_SYNC();
    }
    /* ---- value function (src/ddp.cpp:319-333, src/csqp.cpp:784-799) */
    if (MODE == MODE_DDP) {
      MG_PAR(a, nu) {
        double acc = 0.0;
        for (int c = 0; c < nu; ++c) acc += Quu[a + nu * c] * kff[c];
        p.Quuk[(size_t)t * nu + a] = acc;
      }
    }
    MG_PAR(i, nx) {
      double acc = 0.0;
      for (int a = 0; a < nu; ++a) acc += K[a * nx + i] * Qu[a];
      Vx[i] = Qx[i] - acc;
    }
    MG_PAR(e, nxx) {
      const int i = e % nx, j = e / nx;
      double acc = 0.0;
      for (int a = 0; a < nu; ++a) acc += Qxu[i + nx * a] * K[a * nx + j];
      V[e] -= acc;
    }
    MG_SYNC();
    MG_PAR(e, nxx) { /* Vxx = 0.5 (Vxx + Vxx^T), then + preg */
      const int i = e % nx, j = e / nx;
      if (i > j) {
        const double v = 0.5 * (V[i + nx * j] + V[j + nx * i]);
        V[i + nx * j] = v;
        V[j + nx * i] = v;
      } else if (i == j) {
        double v = 0.5 * (V[e] + V[e]);
        if (!isnan(preg)) v += preg;
        V[e] = v;
      }
    }
    MG_SYNC();
    int bad = 0;
    if (MODE == MODE_DDP) {
      if (!rp.is_feasible) { /* Vx += Vxx fs_t  (src/ddp.cpp:336-338) */
        MG_PAR(i, nx) {
          double acc = 0.0;
          for (int j = 0; j < nx; ++j) acc += V[i + nx * j] * sfs[j];
          Vx[i] += acc;
        }
      }
      MG_SYNC();
      MG_PAR(i, nx) bad |= raise_if_nan(fabs(Vx[i]));
      MG_PAR(e, nxx) bad |= raise_if_nan(fabs(V[e]));
    }
    /* ---- spill the knot's solver members */
    MG_PAR(e, nxx) p.Vxx[(size_t)t * nxx + e] = V[e];
    MG_PAR(i, nx) p.Vx[(size_t)t * nx + i] = Vx[i];
    MG_PAR(e, nxu) p.K[(size_t)t * nxu + e] = K[e];
    MG_PAR(e, nuu) p.L[(size_t)t * nuu + e] = L[e];
    MG_PAR(a, nu) { p.kff[(size_t)t * nu + a] = kff[a]; p.Qu[(size_t)t * nu + a] = Qu[a]; }
    if (MODE == MODE_DDP) {
      if (cta_or(bad, sm + s.red)) return MIMGPU_STATUS_NONFINITE | (t << 8); /* src/ddp.cpp:340-345 */
    } else {
      MG_SYNC();
    }
  }
  return 0;
}

/* ------------------------------------------------------------ SQP forward */
/* SolverSQP::forwardPass (src/sqp.cpp:270-305) fused with checkKKTConditions
 * (src/sqp.cpp:244-268) and the gap norm of computeDirection (src/sqp.cpp:229-236).
 * Writes the per-problem scalars. */
MG_DEV void sqp_forward_kkt(const Prob& p, double* sm, const Smem& s, double mu_dynamic, int stage_fx) {
  const int T = p.T, nx = p.nx, nu = p.nu, nxx = nx * nx, nxu = nx * nu;
  double *sFx = sm + s.Fx, *sFu = sm + s.Fu, *x = sm + s.x, *u = sm + s.u, *lam = sm + s.lam,
         *lamn = sm + s.lamn, *sfs = sm + s.fs, *tv = sm + s.tv, *red = sm + s.red;
  const int ldx = s.ldx;
  double xg = 0.0, ug = 0.0, gap = 0.0, kkt = 0.0;
  /* lag_mul_0 = Vxx_0 (dx_0 - fs_0) + Vx_0 ; dx_0 is whatever the member holds (0) */
  MG_PAR(i, nx) { x[i] = p.dx[i]; sfs[i] = p.fs[i]; }
  MG_SYNC();
  MG_PAR(i, nx) {
    double acc = 0.0;
    for (int j = 0; j < nx; ++j) acc += p.Vxx[i + nx * j] * (x[j] - sfs[j]);
    const double v = acc + p.Vx[i];
    lam[i] = v;
    p.lag[i] = v;
  }
  for (int t = 0; t < T; ++t) {
    const double* gFx = p.Fx + (size_t)t * nxx;
    const double* gFu = p.Fu + (size_t)t * nxu;
    const double* K = p.K + (size_t)t * nxu;
    if (stage_fx) { MG_PAR(e, nxx) sFx[(e % nx) + ldx * (e / nx)] = gFx[e]; }
    MG_PAR(e, nxu) sFu[(e % nx) + ldx * (e / nx)] = gFu[e];
    MG_PAR(a, nu) { /* du = -K dx - k  (src/sqp.cpp:285-286) */
      double acc = 0.0;
      for (int i = 0; i < nx; ++i) acc += K[a * nx + i] * x[i];
      double v = -acc;
      v -= p.kff[(size_t)t * nu + a];
      u[a] = v;
      p.du[(size_t)t * nu + a] = v;
    }
    {
      double a1 = 0.0, a2 = 0.0, a3 = 0.0;
      MG_PAR(i, nx) { a1 += fabs(x[i]); a3 += fabs(sfs[i]); kkt = nmax(kkt, fabs(sfs[i])); }
      MG_SYNC();
      MG_PAR(a, nu) a2 += fabs(u[a]);
      xg += cta_sum(a1, red); ug += cta_sum(a2, red); gap += cta_sum(a3, red);
    }
    const double* Fx = stage_fx ? sFx : gFx;
    const int lfx = stage_fx ? ldx : nx;
    MG_PAR(i, nx) { /* dx+ = fs+ + Fu du + Fx dx  (src/sqp.cpp:287-289) */
      double v = p.fs[(size_t)(t + 1) * nx + i];
      double acc = 0.0;
      for (int a = 0; a < nu; ++a) acc += sFu[i + ldx * a] * u[a];
      v += acc;
      acc = 0.0;
      for (int j = 0; j < nx; ++j) acc += Fx[i + lfx * j] * x[j];
      v += acc;
      tv[i] = v;                     /* dx_{t+1} */
      p.dx[(size_t)(t + 1) * nx + i] = v;
    }
    MG_SYNC();
    MG_PAR(i, nx) { x[i] = tv[i]; sfs[i] = p.fs[(size_t)(t + 1) * nx + i]; }
    MG_SYNC();
    { /* lag_mul_{t+1} = Vxx_{t+1} (dx_{t+1} - fs_{t+1}) + Vx_{t+1}  (src/sqp.cpp:282-284, 296-298) */
      const double* Vxx = p.Vxx + (size_t)(t + 1) * nxx;
      MG_PAR(i, nx) {
        double acc = 0.0;
        for (int j = 0; j < nx; ++j) acc += Vxx[i + nx * j] * (x[j] - sfs[j]);
        const double v = acc + p.Vx[(size_t)(t + 1) * nx + i];
        lamn[i] = v;
        p.lag[(size_t)(t + 1) * nx + i] = v;
      }
    }
    MG_SYNC();
    /* KKT rows of knot t  (src/sqp.cpp:249-258) */
    MG_PAR(i, nx) {
      double acc = 0.0;
      for (int k = 0; k < nx; ++k) acc += Fx[k + lfx * i] * lamn[k];
      const double v = p.Lx[(size_t)t * nx + i] + acc - lam[i];
      kkt = nmax(kkt, fabs(v));
    }
    MG_PAR(a, nu) {
      double acc = 0.0;
      for (int k = 0; k < nx; ++k) acc += sFu[k + ldx * a] * lamn[k];
      const double v = p.Lu[(size_t)t * nu + a] + acc;
      kkt = nmax(kkt, fabs(v));
    }
    MG_SYNC();
    MG_PAR(i, nx) lam[i] = lamn[i];
    MG_SYNC();
  }
  {
    double a1 = 0.0, a3 = 0.0;
    MG_PAR(i, nx) {
      a1 += fabs(x[i]); a3 += fabs(sfs[i]);
      kkt = nmax(kkt, fabs(sfs[i]));
      kkt = nmax(kkt, fabs(p.LxT[i] - lam[i]));   /* src/sqp.cpp:259-262 */
    }
    xg += cta_sum(a1, red); gap += cta_sum(a3, red);
  }
  kkt = cta_max(kkt, red);
  MG_SINGLE {
    const size_t B = p.B, b = p.b;
    p.scal[MIMGPU_S_KKT * B + b] = kkt;
    p.scal[MIMGPU_S_GAP_NORM * B + b] = gap;
    p.scal[MIMGPU_S_MERIT * B + b] = p.cost + mu_dynamic * gap;
    p.scal[MIMGPU_S_XGRAD_NORM * B + b] = xg / (double)(T + 1);
    p.scal[MIMGPU_S_UGRAD_NORM * B + b] = ug / (double)T;
  }
}

/* gap_norm_ = sum_{t=0..T} |fs_t|_1  (src/sqp.cpp:229-234, src/fddp.cpp:195-200) */
MG_DEV double gap_norm_all(const Prob& p, double* red) {
  double acc = 0.0;
  MG_PAR(e, (p.T + 1) * p.nx) acc += fabs(p.fs[e]);
  return cta_sum(acc, red);
}

/* SolverFDDP::updateExpectedImprovement (src/fddp.cpp:256-279) from the spilled members. */
MG_DEV void fddp_expected_improvement(const Prob& p, double* sm, const Smem& s, int is_feasible) {
  const int T = p.T, nx = p.nx, nu = p.nu;
  double dg = 0.0, dq = 0.0;
  MG_PAR(e, T * nu) {
    dg += p.Qu[e] * p.kff[e];
    dq -= p.kff[e] * p.Quuk[e];
  }
  if (!is_feasible) {
    MG_PAR(e, (T + 1) * nx) {
      const int t = e / nx, i = e % nx;
      const double* Vxx = p.Vxx + (size_t)t * nx * nx;
      const double* fs = p.fs + (size_t)t * nx;
      dg -= p.Vx[e] * fs[i];
      double acc = 0.0;
      for (int j = 0; j < nx; ++j) acc += Vxx[i + nx * j] * fs[j];
      dq += fs[i] * acc;
    }
  }
  dg = cta_sum(dg, sm + s.red);
  dq = cta_sum(dq, sm + s.red);
  MG_SINGLE {
    p.scal[MIMGPU_S_DG * (size_t)p.B + p.b] = dg;
    p.scal[MIMGPU_S_DQ * (size_t)p.B + p.b] = dq;
  }
}

/* ----------------------------------------------------------------- bodies */
struct SqpKPar {
  double preg, dreg, mu_dynamic;
  int is_feasible;
};

MG_DEV void sqp_direction_body(const KArgs& a, int b, double* sm, const SqpKPar& kp) {
  Prob p = bind(a, b);
  const Smem s = smem_layout(p.nx, p.nu, p.ng, p.ngT, a.stage_fx);
  RicPar rp;
  rp.preg = a.preg_arr ? a.preg_arr[b] : kp.preg;
  rp.dreg = a.dreg_arr ? a.dreg_arr[b] : kp.dreg;
  rp.sigma = 0.0;
  rp.is_feasible = kp.is_feasible;
  const int st = riccati_backward<MODE_DDP>(p, sm, s, rp, a.stage_fx);
  MG_SINGLE { a.status[b] = st; a.qp_iters[b] = 0; }
  if (st) { /* gap norm / merit are computed before the backward pass in the reference */
    const double gap = gap_norm_all(p, sm + s.red);
    MG_SINGLE {
      p.scal[MIMGPU_S_GAP_NORM * (size_t)p.B + b] = gap;
      p.scal[MIMGPU_S_MERIT * (size_t)p.B + b] = p.cost + kp.mu_dynamic * gap;
    }
    return;
  }
  MG_SYNC();
  sqp_forward_kkt(p, sm, s, kp.mu_dynamic, a.stage_fx);
}

MG_DEV void fddp_direction_body(const KArgs& a, int b, double* sm, const SqpKPar& kp) {
  Prob p = bind(a, b);
  const Smem s = smem_layout(p.nx, p.nu, p.ng, p.ngT, a.stage_fx);
  RicPar rp;
  rp.preg = a.preg_arr ? a.preg_arr[b] : kp.preg;
  rp.dreg = a.dreg_arr ? a.dreg_arr[b] : kp.dreg;
  rp.sigma = 0.0;
  rp.is_feasible = kp.is_feasible;
  const double gap = gap_norm_all(p, sm + s.red);
  MG_SINGLE { p.scal[MIMGPU_S_GAP_NORM * (size_t)p.B + b] = gap; }
  const int st = riccati_backward<MODE_DDP>(p, sm, s, rp, a.stage_fx);
  MG_SINGLE { a.status[b] = st; a.qp_iters[b] = 0; }
  if (st) return;
  MG_SYNC();
  fddp_expected_improvement(p, sm, s, kp.is_feasible);
}

/* ------------------------------------------------------------------- CSQP */
struct CsqpKPar {
  mimgpu_csqp_params q;
  int with_kkt;
};

/* SolverCSQP::apply_rho_update (src/csqp.cpp:539-582) */
MG_DEV void csqp_apply_rho(const Prob& p, double rho_sparse, double rho_min) {
  const int nc = p.T * p.ng + p.ngT;
  MG_PAR(e, nc) {
    const bool term = e >= p.T * p.ng;
    const double lb = term ? p.glbT[e - p.T * p.ng] : p.glb[e];
    const double ub = term ? p.gubT[e - p.T * p.ng] : p.gub[e];
    if (lb == -MG_INF && ub == MG_INF) p.rho[e] = rho_min;
    else if (fabs(lb - ub) <= 1e-6) p.rho[e] = 1e3 * rho_sparse;
    else if (lb < ub) p.rho[e] = rho_sparse;
  }
}

/* SolverCSQP::forwardPass_without_constraints (src/csqp.cpp:649-669): writes dx_, du_. */
MG_DEV void csqp_forward_eq(const Prob& p, double* sm, const Smem& s) {
  const int T = p.T, nx = p.nx, nu = p.nu, nxx = nx * nx, nxu = nx * nu;
  double *x = sm + s.x, *u = sm + s.u, *tv = sm + s.tv;
  MG_PAR(i, nx) x[i] = p.dx[i];
  MG_SYNC();
  for (int t = 0; t < T; ++t) {
    const double* Fx = p.Fx + (size_t)t * nxx;
    const double* Fu = p.Fu + (size_t)t * nxu;
    const double* K = p.K + (size_t)t * nxu;
    MG_PAR(a, nu) {
      double acc = 0.0;
      for (int i = 0; i < nx; ++i) acc += K[a * nx + i] * x[i];
      double v = -p.kff[(size_t)t * nu + a];
      v -= acc;
      u[a] = v;
      p.du[(size_t)t * nu + a] = v;
    }
    MG_SYNC();
    MG_PAR(i, nx) {
      double v = p.fs[(size_t)(t + 1) * nx + i];
      double acc = 0.0;
      for (int j = 0; j < nx; ++j) acc += Fx[i + nx * j] * x[j];
      v += acc;
      acc = 0.0;
      for (int a = 0; a < nu; ++a) acc += Fu[i + nx * a] * u[a];
      v += acc;
      tv[i] = v;
      p.dx[(size_t)(t + 1) * nx + i] = v;
    }
    MG_SYNC();
    MG_PAR(i, nx) x[i] = tv[i];
    MG_SYNC();
  }
}

/* SolverCSQP::backwardPass_without_rho_update (src/csqp.cpp:1025-1108). */
MG_DEV void csqp_backward_light(const Prob& p, double* sm, const Smem& s, double sigma) {
  const int T = p.T, nx = p.nx, nu = p.nu, ng = p.ng, ngT = p.ngT, nxx = nx * nx, nxu = nx * nu, nuu = nu * nu;
  double *Vx = sm + s.Vx, *tv = sm + s.tv, *Qx = sm + s.Qx, *Qu = sm + s.Qu, *kff = sm + s.kff, *sd = sm + s.d,
         *L = sm + s.L;
  if (ngT) {
    MG_PAR(r, ngT) {
      double d = p.y[(size_t)T * ng + r];
      d -= p.rho[(size_t)T * ng + r] * p.z[(size_t)T * ng + r];
      sd[r] = d;
    }
    MG_SYNC();
  }
  MG_PAR(i, nx) {
    double v = p.LxT[i];
    v -= sigma * p.dx[(size_t)T * nx + i];
    if (ngT) {
      double acc = 0.0;
      for (int r = 0; r < ngT; ++r) acc += p.GxT[r + ngT * i] * sd[r];
      v += acc;
    }
    Vx[i] = v;
    p.Vx[(size_t)T * nx + i] = v;
  }
  MG_SYNC();
  for (int t = T - 1; t >= 0; --t) {
    const double* Fx = p.Fx + (size_t)t * nxx;
    const double* Fu = p.Fu + (size_t)t * nxu;
    const double* Gx = ng ? p.Gx + (size_t)t * ng * nx : nullptr;
    const double* Gu = ng ? p.Gu + (size_t)t * ng * nu : nullptr;
    const double* K = p.K + (size_t)t * nxu;
    MG_PAR(i, nx) tv[i] = p.Vxx_fs[(size_t)t * nx + i] + Vx[i];
    MG_PAR(r, ng) {
      double d = p.y[(size_t)t * ng + r];
      d -= p.rho[(size_t)t * ng + r] * p.z[(size_t)t * ng + r];
      sd[r] = d;
    }
    MG_PAR(e, nuu) L[e] = p.L[(size_t)t * nuu + e];
    MG_SYNC();
    MG_PAR(i, nx) {
      double v = p.Lx[(size_t)t * nx + i];
      v -= sigma * p.dx[(size_t)t * nx + i];
      if (ng && t > 0) {
        double acc = 0.0;
        for (int r = 0; r < ng; ++r) acc += Gx[r + ng * i] * sd[r];
        v += acc;
      }
      double acc = 0.0;
      for (int k = 0; k < nx; ++k) acc += Fx[k + nx * i] * tv[k];
      Qx[i] = v + acc;
    }
    MG_PAR(a, nu) {
      double v = p.Lu[(size_t)t * nu + a];
      v -= sigma * p.du[(size_t)t * nu + a];
      if (ng) {
        double acc = 0.0;
        for (int r = 0; r < ng; ++r) acc += Gu[r + ng * a] * sd[r];
        v += acc;
      }
      double acc = 0.0;
      for (int k = 0; k < nx; ++k) acc += Fu[k + nx * a] * tv[k];
      v += acc;
      Qu[a] = v;
      kff[a] = v;
      p.Qu[(size_t)t * nu + a] = v;
    }
    MG_SYNC();
    llt_solve_one(L, nu, kff);
    MG_PAR(i, nx) {
      double acc = 0.0;
      for (int a = 0; a < nu; ++a) acc += K[a * nx + i] * Qu[a];
      const double v = Qx[i] - acc;
      Vx[i] = v;
      p.Vx[(size_t)t * nx + i] = v;
    }
    MG_PAR(a, nu) p.kff[(size_t)t * nu + a] = kff[a];
    MG_SYNC();
  }
}

struct AdmmNorms {
  double primal, dual, primal_rel, dual_rel;
};

/* SolverCSQP::update_lagrangian_parameters for one knot (src/csqp.cpp:1221-1281,
 * 1283-1338); x = dxtilde_t, u = dutilde_t in shared memory. */
MG_DEV void csqp_update_knot(const Prob& p, double* sm, const Smem& s, int t, const double* x, const double* u,
                             const mimgpu_csqp_params& q, bool check, AdmmNorms& n) {
  const int T = p.T, nx = p.nx, nu = p.nu;
  const bool last = (t == T);
  const int nc = last ? p.ngT : p.ng;
  if (nc == 0) return;
  const size_t o = (size_t)t * p.ng;
  const double* Gx = last ? p.GxT : p.Gx + (size_t)t * nc * nx;
  const double* Gu = last ? nullptr : p.Gu + (size_t)t * nc * nu;
  const double* g = last ? p.gT : p.g + o;
  const double* lb = last ? p.glbT : p.glb + o;
  const double* ub = last ? p.gubT : p.gub + o;
  double *sc = sm + s.c, *sdz = sm + s.dz, *sy = sm + s.d, *sz = sm + s.rho;
  MG_PAR(r, nc) {
    const double zp = p.z[o + r];                       /* z_prev = z */
    double c = 0.0;
    {
      double acc = 0.0;
      for (int j = 0; j < nx; ++j) acc += Gx[r + nc * j] * x[j];
      c += acc;
      if (!last) {
        acc = 0.0;
        for (int a = 0; a < nu; ++a) acc += Gu[r + nc * a] * u[a];
        c += acc;
      }
    }
    double zr = q.alpha * c;                            /* z_relaxed */
    zr += (1. - q.alpha) * zp;
    const double rho = p.rho[o + r];
    const double irho = 1. / rho;                       /* inv_rho_vec_ */
    const double yo = p.y[o + r];
    double v = zr + yo * irho;
    const double lo = lb[r] - g[r], hi = ub[r] - g[r];
    v = (v < lo) ? lo : v;                              /* cwiseMax(lb - g) */
    v = (hi < v) ? hi : v;                              /* cwiseMin(ub - g) */
    const double yn = yo + rho * (zr - v);
    p.z[o + r] = v;
    p.y[o + r] = yn;
    if (check) {
      sc[r] = c; sdz[r] = rho * (v - zp); sy[r] = yn; sz[r] = v;
      n.primal = nmax(n.primal, fabs(c - v));
      n.primal_rel = nmax(n.primal_rel, fabs(c));
      n.primal_rel = nmax(n.primal_rel, fabs(v));
      if (q.update_rho_with_heuristic) {
        n.dual = nmax(n.dual, fabs(rho * (v - zp)));
        n.dual_rel = nmax(n.dual_rel, fabs(yn));
      }
    }
  }
  if (check && !q.update_rho_with_heuristic) {
    MG_SYNC();
    MG_PAR(j, nx) {
      double a1 = 0.0, a2 = 0.0;
      for (int r = 0; r < nc; ++r) { a1 += Gx[r + nc * j] * sdz[r]; a2 += Gx[r + nc * j] * sy[r]; }
      n.dual = nmax(n.dual, fabs(a1));
      n.dual_rel = nmax(n.dual_rel, fabs(a2));
    }
    if (!last) {
      MG_PAR(a, nu) {
        double a1 = 0.0, a2 = 0.0;
        for (int r = 0; r < nc; ++r) { a1 += Gu[r + nc * a] * sdz[r]; a2 += Gu[r + nc * a] * sy[r]; }
        n.dual = nmax(n.dual, fabs(a1));
        n.dual_rel = nmax(n.dual_rel, fabs(a2));
      }
    }
    MG_SYNC();
  }
}

/* SolverCSQP::forwardPass (src/csqp.cpp:630-647) fused with
 * update_lagrangian_parameters (src/csqp.cpp:1196-1340). */
MG_DEV void csqp_forward_update(const Prob& p, double* sm, const Smem& s, const mimgpu_csqp_params& q, int iter,
                                AdmmNorms& n) {
  const int T = p.T, nx = p.nx, nu = p.nu, nxx = nx * nx, nxu = nx * nu;
  double *x = sm + s.x, *u = sm + s.u, *tv = sm + s.tv, *red = sm + s.red;
  const bool check = (iter % q.rho_update_interval) == 0;
  n.primal = n.dual = n.primal_rel = n.dual_rel = -MG_INF;      /* src/csqp.cpp:1201-1204 */
  MG_PAR(i, nx) x[i] = p.dxt[i];
  MG_SYNC();
  for (int t = 0; t < T; ++t) {
    const double* Fx = p.Fx + (size_t)t * nxx;
    const double* Fu = p.Fu + (size_t)t * nxu;
    const double* K = p.K + (size_t)t * nxu;
    MG_PAR(a, nu) {
      double acc = 0.0;
      for (int i = 0; i < nx; ++i) acc += K[a * nx + i] * x[i];
      double v = -p.kff[(size_t)t * nu + a];
      v -= acc;
      u[a] = v;
      p.dut[(size_t)t * nu + a] = v;
      p.du[(size_t)t * nu + a] = v;                     /* du_ = dutilde_  (:1226,:1242) */
    }
    MG_PAR(i, nx) p.dx[(size_t)t * nx + i] = x[i];      /* dx_ = dxtilde_ */
    MG_SYNC();
    MG_PAR(i, nx) {
      double v = p.fs[(size_t)(t + 1) * nx + i];
      double acc = 0.0;
      for (int j = 0; j < nx; ++j) acc += Fx[i + nx * j] * x[j];
      v += acc;
      acc = 0.0;
      for (int a = 0; a < nu; ++a) acc += Fu[i + nx * a] * u[a];
      v += acc;
      tv[i] = v;
      p.dxt[(size_t)(t + 1) * nx + i] = v;
    }
    csqp_update_knot(p, sm, s, t, x, u, q, check, n);
    MG_SYNC();
    MG_PAR(i, nx) x[i] = tv[i];
    MG_SYNC();
  }
  MG_PAR(i, nx) p.dx[(size_t)T * nx + i] = x[i];
  csqp_update_knot(p, sm, s, T, x, u, q, check, n);
  if (check) {
    n.primal = cta_max(n.primal, red);
    n.dual = cta_max(n.dual, red);
    n.primal_rel = cta_max(n.primal_rel, red);
    n.dual_rel = cta_max(n.dual_rel, red);
  }
  MG_SYNC();
}

/* SolverCSQP::checkKKTConditions (src/csqp.cpp:584-628). */
MG_DEV void csqp_kkt(const Prob& p, double* sm, const Smem& s) {
  const int T = p.T, nx = p.nx, nu = p.nu, ng = p.ng, ngT = p.ngT, nxx = nx * nx, nxu = nx * nu;
  double* red = sm + s.red;
  double xg = 0.0, ug = 0.0, kkt = 0.0;
  MG_PAR(e, (T + 1) * nx) {
    const int t = e / nx, i = e % nx;
    const double* Vxx = p.Vxx + (size_t)t * nxx;
    const double* x = p.dxt + (size_t)t * nx;
    double acc = 0.0;
    for (int j = 0; j < nx; ++j) acc += Vxx[i + nx * j] * x[j];
    p.lag[e] = p.Vx[e] + acc;
    xg += fabs(x[i]);
    kkt = nmax(kkt, fabs(p.fs[e]));
  }
  MG_PAR(e, T * nu) ug += fabs(p.dut[e]);
  MG_SYNC();
  MG_PAR(e, T * nx) {
    const int t = e / nx, i = e % nx;
    const double* Fx = p.Fx + (size_t)t * nxx;
    const double* lamn = p.lag + (size_t)(t + 1) * nx;
    double acc = 0.0;
    for (int k = 0; k < nx; ++k) acc += Fx[k + nx * i] * lamn[k];
    double v = p.Lx[e] + acc - p.lag[e];
    if (t > 0 && ng) {
      const double* Gx = p.Gx + (size_t)t * ng * nx;
      double a2 = 0.0;
      for (int r = 0; r < ng; ++r) a2 += Gx[r + ng * i] * p.y[(size_t)t * ng + r];
      v += a2;
    }
    kkt = nmax(kkt, fabs(v));
  }
  MG_PAR(e, T * nu) {
    const int t = e / nu, a = e % nu;
    const double* Fu = p.Fu + (size_t)t * nxu;
    const double* lamn = p.lag + (size_t)(t + 1) * nx;
    double acc = 0.0;
    for (int k = 0; k < nx; ++k) acc += Fu[k + nx * a] * lamn[k];
    double v = p.Lu[e] + acc;
    if (ng) {
      const double* Gu = p.Gu + (size_t)t * ng * nu;
      double a2 = 0.0;
      for (int r = 0; r < ng; ++r) a2 += Gu[r + ng * a] * p.y[(size_t)t * ng + r];
      v += a2;
    }
    kkt = nmax(kkt, fabs(v));
  }
  MG_PAR(i, nx) {
    double v = p.LxT[i] - p.lag[(size_t)T * nx + i];
    if (ngT) {
      double a2 = 0.0;
      for (int r = 0; r < ngT; ++r) a2 += p.GxT[r + ngT * i] * p.y[(size_t)T * ng + r];
      v += a2;
    }
    kkt = nmax(kkt, fabs(v));
  }
  kkt = cta_max(kkt, red);
  xg = cta_sum(xg, red);
  ug = cta_sum(ug, red);
  MG_SINGLE {
    const size_t B = p.B, b = p.b;
    kkt = std_max(kkt, p.scal[MIMGPU_S_CONSTRAINT_NORM * B + b]);
    p.scal[MIMGPU_S_KKT * B + b] = kkt;
    p.scal[MIMGPU_S_XGRAD_NORM * B + b] = xg / (double)(T + 1);
    p.scal[MIMGPU_S_UGRAD_NORM * B + b] = ug / (double)T;
  }
}

/* SolverCSQP::computeDirection (src/csqp.cpp:454-517), whole ADMM loop. */
MG_DEV void csqp_direction_body(const KArgs& a, int b, double* sm, const CsqpKPar& kp) {
  Prob p = bind(a, b);
  const Smem s = smem_layout(p.nx, p.nu, p.ng, p.ngT, a.stage_fx);
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
  MG_PAR(e, nc) { p.z[e] = 0.0; if (q.reset_y) p.y[e] = 0.0; }
  MG_SYNC();
  int st = 0, converged = 0, it = 0;
  AdmmNorms n;
  n.primal = p.scal[MIMGPU_S_NORM_PRIMAL * B + b]; n.dual = p.scal[MIMGPU_S_NORM_DUAL * B + b];
  n.primal_rel = p.scal[MIMGPU_S_NORM_PRIMAL_REL * B + b]; n.dual_rel = p.scal[MIMGPU_S_NORM_DUAL_REL * B + b];
  if (q.equality_qp_initial_guess) { /* src/csqp.cpp:461-464 */
    st = riccati_backward<MODE_EQ>(p, sm, s, rp, a.stage_fx);
    if (!st) { MG_SYNC(); csqp_forward_eq(p, sm, s); }
  }
  const int interval = q.rho_update_interval;
  if (!st) {
    for (it = 1; it < q.max_qp_iters + 1; ++it) { /* src/csqp.cpp:470 */
      if (it % interval == 1 || interval == 1) {
        st = riccati_backward<MODE_FULL>(p, sm, s, rp, a.stage_fx);
        if (st) break;
      } else {
        csqp_backward_light(p, sm, s, q.sigma);
      }
      MG_SYNC();
      csqp_forward_update(p, sm, s, q, it, n);
      if (it % interval == 0) {
        /* update_rho_vec (src/csqp.cpp:519-532); on other iterations it cannot change anything */
        const double scale = sqrt((n.primal * n.dual_rel) / (n.dual * n.primal_rel));
        const double rho_est = std_min(std_max(scale * rho_sparse, q.rho_min), q.rho_max);
        if (rho_est > rho_sparse * q.adaptive_rho_tolerance || rho_est < rho_sparse / q.adaptive_rho_tolerance) {
          rho_sparse = rho_est;
          csqp_apply_rho(p, rho_sparse, q.rho_min);
          MG_SYNC();
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
  MG_SYNC();
  MG_SINGLE {
    a.status[b] = st; a.qp_iters[b] = it; a.converged[b] = converged;
    p.scal[MIMGPU_S_RHO_SPARSE * B + b] = rho_sparse;
    p.scal[MIMGPU_S_NORM_PRIMAL * B + b] = n.primal; p.scal[MIMGPU_S_NORM_DUAL * B + b] = n.dual;
    p.scal[MIMGPU_S_NORM_PRIMAL_REL * B + b] = n.primal_rel; p.scal[MIMGPU_S_NORM_DUAL_REL * B + b] = n.dual_rel;
  }
  if (!st && kp.with_kkt) csqp_kkt(p, sm, s);
}

/* SolverCSQP::reset_rho_vec (src/csqp.cpp:534-537). */
MG_DEV void csqp_reset_rho_body(const KArgs& a, int b, double rho_sparse_base, double rho_min) {
  Prob p = bind(a, b);
  csqp_apply_rho(p, rho_sparse_base, rho_min);
  MG_SINGLE { p.scal[MIMGPU_S_RHO_SPARSE * (size_t)p.B + b] = rho_sparse_base; }
}

/* SolverCSQP::calc, recalc = false part (src/csqp.cpp:417-451): gap_norm_ over
 * fs_[1..T] and constraint_norm_, accumulated knot by knot as in the reference. */
MG_DEV void csqp_calc_norms_body(const KArgs& a, int b, double* red) {
  Prob p = bind(a, b);
  const int T = p.T, nx = p.nx, ng = p.ng, ngT = p.ngT;
  double gap = 0.0, con = 0.0;
  MG_PAR(e, T * nx) gap += fabs(p.fs[nx + e]);
  MG_PAR(e, T * ng) {
    const double lo = p.glb[e] - p.g[e], hi = p.g[e] - p.gub[e];
    con += fabs(lo > 0. ? lo : 0.);
    con += fabs(hi > 0. ? hi : 0.);
  }
  MG_PAR(e, ngT) {
    const double lo = p.glbT[e] - p.gT[e], hi = p.gT[e] - p.gubT[e];
    con += fabs(lo > 0. ? lo : 0.);
    con += fabs(hi > 0. ? hi : 0.);
  }
  gap = cta_sum(gap, red);
  con = cta_sum(con, red);
  MG_SINGLE {
    p.scal[MIMGPU_S_GAP_NORM * (size_t)p.B + b] = gap;
    p.scal[MIMGPU_S_CONSTRAINT_NORM * (size_t)p.B + b] = con;
  }
}

}  // namespace mimgpu
