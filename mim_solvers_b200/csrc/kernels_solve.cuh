/*
 * kernels_solve.cuh -- device side of a batched SolverSQP::solve() for ANALYTIC models
 * (SURVEY 8f row n1): model evaluation, merit / filter line search, acceptance and the
 * regularisation schedule all run on the GPU, so an SQP iteration is
 *     k_model_calcdiff -> direction kernel -> k_sqp_post_direction -> k_sqp_linesearch
 * with no derivative upload and no download of dx / du; the host only reads two counters per
 * iteration (how many problems need a regularised retry, how many are still running).
 *
 * Reference code restated per problem (paths relative to the reference tree):
 *   SolverSQP::solve                       src/sqp.cpp:75-222   (loop, KKT test, filter, reg schedule)
 *   SolverSQP::tryStep                     src/sqp.cpp:307-374  (candidates, cost_try, gap_norm_try, merit_try)
 *   SolverDDP::increase/decreaseRegularization  src/ddp.cpp:431-445
 *   filter ring buffers                    src/sqp.cpp:34-35,130-131,148-157 (boost::circular_buffer,
 *                                          resized to filter_size_, zero-initialised, never cleared)
 * The model evaluations (m->calc / calcDiff, crocoddyl's job in the reference) are those of the
 * analytic family below -- LQ plus separable polynomial terms, which makes the problem genuinely
 * nonlinear (alpha < 1, several iterations) while staying evaluable element-wise:
 *     x+  = A_t x + B_t u + f_t + c3 x^3                                (x^3 element-wise)
 *     l_t = 1/2 [x;u]^T [[Q,S],[S^T,R]] [x;u] + q.x + r.u + w4/4 sum_i x_i^4
 *     l_T = 1/2 x^T Q_T x + q_T.x + w4/4 sum_i x_i^4
 * All sums that feed a comparison (cost, gap norm) are accumulated in increasing t by one lane
 * (SURVEY Appendix B.15).
 */
#pragma once
#include <cuda_runtime.h>
#include <math_constants.h>
#include <stdint.h>

#include "launchers.h"

namespace mimgpu {

__device__ __forceinline__ double wsum_fixed(double v) { /* butterfly: every lane gets the same, order-fixed sum */
#pragma unroll
  for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

/* One running knot of the model at (x, u) held in shared memory by the calling warp:
 * lanes own rows.  Returns the knot's cost (same value in every lane); writes, when the pointers
 * are non-null, xnext[nx], Lx[nx], Lu[nu]. */
__device__ __forceinline__ double poly_knot(const PolyModelDev& m, int nx, int nu, size_t kt, const double* x,
                                            const double* u, int lane, double* xnext, double* Lx, double* Lu) {
  const double* A = m.A + kt * nx * nx;
  const double* Bm = m.Bm + kt * nx * nu;
  const double* Q = m.Q + kt * nx * nx;
  const double* S = m.S + kt * nx * nu;
  const double* R = m.R + kt * nu * nu;
  double part = 0.0;
  for (int i = lane; i < nx; i += 32) {
    double hx = 0.0, xn = 0.0;
    for (int j = 0; j < nx; ++j) { hx = fma(Q[i + nx * j], x[j], hx); xn = fma(A[i + nx * j], x[j], xn); }
    for (int a = 0; a < nu; ++a) { hx = fma(S[i + nx * a], u[a], hx); xn = fma(Bm[i + nx * a], u[a], xn); }
    const double xi = x[i], x2 = xi * xi;
    if (xnext) xnext[i] = xn + m.f[kt * nx + i] + m.c3 * x2 * xi;
    const double qi = m.q[kt * nx + i];
    if (Lx) Lx[i] = hx + qi + m.w4 * x2 * xi;
    part += xi * (0.5 * hx + qi) + 0.25 * m.w4 * x2 * x2;
  }
  for (int a = lane; a < nu; a += 32) {
    double hu = 0.0;
    for (int i = 0; i < nx; ++i) hu = fma(S[i + nx * a], x[i], hu);
    for (int c = 0; c < nu; ++c) hu = fma(R[a + nu * c], u[c], hu);
    const double ra = m.r[kt * nu + a];
    if (Lu) Lu[a] = hu + ra;
    part += u[a] * (0.5 * hu + ra);
  }
  return wsum_fixed(part);
}
__device__ __forceinline__ double poly_terminal(const PolyModelDev& m, int nx, size_t b, const double* x, int lane,
                                                double* LxT) {
  const double* QT = m.QT + b * nx * nx;
  double part = 0.0;
  for (int i = lane; i < nx; i += 32) {
    double hx = 0.0;
    for (int j = 0; j < nx; ++j) hx = fma(QT[i + nx * j], x[j], hx);
    const double xi = x[i], x2 = xi * xi, qi = m.qT[b * nx + i];
    if (LxT) LxT[i] = hx + qi + m.w4 * x2 * xi;
    part += xi * (0.5 * hx + qi) + 0.25 * m.w4 * x2 * x2;
  }
  return wsum_fixed(part);
}

/* problem_->calc + calcDiff + computeDynamicFeasibility (src/ddp.cpp:253-263) of the analytic
 * model at (xs_, us_): writes Lx Lu LxT fs cost, and Fx Lxx LxxT when the polynomial terms are on
 * (they are the constant A, Q, QT otherwise, copied once at attach time).  One CTA per problem,
 * one warp per knot; the per-knot costs are summed in increasing t by one thread. */
__global__ void __launch_bounds__(256) k_model_calcdiff(KArgs a, PolyModelDev m, SolveDev s, const int32_t* list, int n) {
  extern __shared__ double sm_solve[];
  if ((int)blockIdx.x >= n) return;
  const int b = list ? list[blockIdx.x] : (int)blockIdx.x;
  const int T = a.sh.T, nx = a.sh.nx, nu = a.sh.nu;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
  const size_t sb = (size_t)b;
  double* costs = sm_solve;                                  /* [T+1] */
  double* wsx = sm_solve + (T + 2) + warp * (2 * nx + nu);   /* per warp: x | u | xnext */
  double* x = wsx; double* u = wsx + nx; double* xn = u + nu;
  const double* xs = s.xs + sb * (T + 1) * nx;
  const double* us = s.us + sb * T * nu;
  double* Fx = const_cast<double*>(a.Fx) + sb * T * nx * nx;
  double* Lxx = const_cast<double*>(a.Lxx) + sb * T * nx * nx;
  double* Lx = const_cast<double*>(a.Lx) + sb * T * nx;
  double* Lu = const_cast<double*>(a.Lu) + sb * T * nu;
  double* fs = const_cast<double*>(a.fs) + sb * (T + 1) * nx;
  const bool nonlin = (m.c3 != 0.0) || (m.w4 != 0.0);
  for (int t = warp; t <= T; t += nw) {
    for (int i = lane; i < nx; i += 32) x[i] = xs[(size_t)t * nx + i];
    if (t < T) for (int c = lane; c < nu; c += 32) u[c] = us[(size_t)t * nu + c];
    __syncwarp();
    if (t < T) {
      const size_t kt = sb * T + t;
      const double c = poly_knot(m, nx, nu, kt, x, u, lane, xn, Lx + (size_t)t * nx, Lu + (size_t)t * nu);
      __syncwarp();
      if (lane == 0) costs[t] = c;
      for (int i = lane; i < nx; i += 32) fs[(size_t)(t + 1) * nx + i] = xn[i] - xs[(size_t)(t + 1) * nx + i];
      if (nonlin) {
        const double* A = m.A + kt * nx * nx;
        const double* Q = m.Q + kt * nx * nx;
        for (int e = lane; e < nx * nx; e += 32) {
          const int i = e % nx, j = e / nx;
          const double x2 = x[i] * x[i];
          Fx[(size_t)t * nx * nx + e] = A[e] + (i == j ? 3.0 * m.c3 * x2 : 0.0);
          Lxx[(size_t)t * nx * nx + e] = Q[e] + (i == j ? 3.0 * m.w4 * x2 : 0.0);
        }
      }
    } else {
      const double c = poly_terminal(m, nx, sb, x, lane, const_cast<double*>(a.LxT) + sb * nx);
      if (lane == 0) costs[T] = c;
      if (nonlin) {
        double* LxxT = const_cast<double*>(a.LxxT) + sb * nx * nx;
        for (int e = lane; e < nx * nx; e += 32) {
          const int i = e % nx, j = e / nx;
          LxxT[e] = m.QT[sb * nx * nx + e] + (i == j ? 3.0 * m.w4 * x[i] * x[i] : 0.0);
        }
      }
    }
    __syncwarp();
  }
  for (int i = threadIdx.x; i < nx; i += blockDim.x) fs[i] = m.x0[sb * nx + i] - xs[i]; /* fs_0 = x0 (-) xs_0 */
  __syncthreads();
  if (threadIdx.x == 0) {
    double c = 0.0;
    for (int t = 0; t <= T; ++t) c += costs[t];   /* cost_ = sum of the knots' costs, increasing t */
    s.cost[b] = c;
  }
}

/* After a direction launch on the problems of `list`: the catch block of solve() for the problems
 * whose backward pass failed (increaseRegularization, give up at reg_max: src/sqp.cpp:103-116), the
 * KKT test (:128-131) and the filter push (:133-134) for the others.  counts[0] = problems that
 * need the direction again (retry_out, a list other than `list`), counts[1] = problems that go to the line search (list_b). */
__global__ void k_sqp_post_direction(KArgs a, SolveDev s, SolvePar sp, const int32_t* list, int n, int32_t* retry_out) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= n) return;
  const int b = list ? list[k] : k;
  const size_t B = a.sh.B;
  if (a.status[b] != 0) {
    double pr = s.preg[b] * sp.reg_inc;
    if (pr > sp.reg_max) pr = sp.reg_max;
    s.preg[b] = pr; s.dreg[b] = pr;
    if (pr == sp.reg_max) { s.done[b] = 2; return; }
    retry_out[atomicAdd(&s.counts[0], 1)] = b;
    return;
  }
  if (a.scal[MIMGPU_S_KKT * B + b] <= sp.termination_tol) { s.done[b] = 1; return; }
  double* cl = s.cost_list + (size_t)b * sp.filter_size;
  double* gl = s.gap_list + (size_t)b * sp.filter_size;
  for (int i = 0; i + 1 < sp.filter_size; ++i) { cl[i] = cl[i + 1]; gl[i] = gl[i + 1]; }   /* push_back on a full ring */
  cl[sp.filter_size - 1] = s.cost[b];
  gl[sp.filter_size - 1] = a.scal[MIMGPU_S_GAP_NORM * B + b];
  s.list_b[atomicAdd(&s.counts[1], 1)] = b;
}

/* The line search of one SQP iteration (src/sqp.cpp:136-183): warp n evaluates tryStep(2^-n) on the
 * model (src/sqp.cpp:307-374), thread 0 walks the step lengths in order and applies the filter /
 * merit rule, then the CTA applies the accepted step and the regularisation schedule. */
__global__ void __launch_bounds__(32 * SOLVE_NALPHA) k_sqp_linesearch(KArgs a, PolyModelDev m, SolveDev s, SolvePar sp,
                                                                      const int32_t* list, int n) {
  extern __shared__ double sm_solve[];
  if ((int)blockIdx.x >= n) return;
  const int b = list[blockIdx.x];
  const int T = a.sh.T, nx = a.sh.nx, nu = a.sh.nu;
  const size_t sb = (size_t)b, B = a.sh.B;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  double* res = sm_solve;                                          /* [NALPHA][2]: cost_try, gap_norm_try */
  double* wsx = sm_solve + 2 * SOLVE_NALPHA + 2 + warp * (3 * nx + nu);
  double *x = wsx, *u = x + nx, *xn = u + nu, *xprev_next = xn + nx; /* x_try_t | u_try_t | xnext_t | xnext_{t-1} */
  double* xs = s.xs + sb * (T + 1) * nx;
  double* us = s.us + sb * T * nu;
  const double* dx = a.dx + sb * (T + 1) * nx;
  const double* du = a.du + sb * T * nu;
  const double alpha = 1.0 / (double)(1 << warp);
  {
    double cost = 0.0, gap = 0.0;
    for (int t = 0; t <= T; ++t) {
      for (int i = lane; i < nx; i += 32) x[i] = xs[(size_t)t * nx + i] + alpha * dx[(size_t)t * nx + i];
      if (t < T) for (int c = lane; c < nu; c += 32) { double v = us[(size_t)t * nu + c]; v += alpha * du[(size_t)t * nu + c]; u[c] = v; }
      __syncwarp();
      if (t > 0) { /* fs_try_{t-1} = xnext_{t-1} (-) xs_try_t */
        double g = 0.0;
        for (int i = lane; i < nx; i += 32) g += fabs(xprev_next[i] - x[i]);
        gap += wsum_fixed(g);
      }
      if (t < T) cost += poly_knot(m, nx, nu, sb * T + t, x, u, lane, xn, nullptr, nullptr);
      else cost += poly_terminal(m, nx, sb, x, lane, nullptr);
      __syncwarp();
      if (t < T) for (int i = lane; i < nx; i += 32) xprev_next[i] = xn[i];
      __syncwarp();
    }
    if (lane == 0) { res[2 * warp] = cost; res[2 * warp + 1] = gap; }
  }
  __syncthreads();
  __shared__ double s_alpha;
  __shared__ int s_accept;
  if (threadIdx.x == 0) {
    const double* cl = s.cost_list + sb * sp.filter_size;
    const double* gl = s.gap_list + sb * sp.filter_size;
    const int iter = s.iter[b];
    const double merit = a.scal[MIMGPU_S_MERIT * B + b];
    double step = 1.0;
    int acc = 0;
    for (int k = 0; k < SOLVE_NALPHA; ++k) {
      step = 1.0 / (double)(1 << k);
      const double ct = res[2 * k], gt = res[2 * k + 1];
      if (raise_if_nan(ct)) continue;                                   /* "step_error": next alpha */
      if (sp.use_filter) {
        bool worse = false;
        int count = 0;
        while (count < sp.filter_size && !worse && count <= iter) {
          worse = cl[sp.filter_size - 1 - count] <= ct && gl[sp.filter_size - 1 - count] <= gt;
          ++count;
        }
        if (!worse) { acc = 1; break; }
      } else {
        const double mt = ct + sp.mu_dynamic * gt;
        if (merit > mt) { acc = 1; break; }
      }
    }
    s_alpha = step; s_accept = acc;
    s.steplength[b] = step;
    /* regularisation schedule and iteration count (src/sqp.cpp:174-183, :100) */
    double pr = s.preg[b];
    if (step > sp.th_stepdec) { pr /= sp.reg_dec; if (pr < sp.reg_min) pr = sp.reg_min; }
    int fail = 0;
    if (step <= sp.th_stepinc) { pr *= sp.reg_inc; if (pr > sp.reg_max) pr = sp.reg_max; if (pr == sp.reg_max) fail = 1; }
    s.preg[b] = pr; s.dreg[b] = pr;
    if (fail) s.done[b] = 2; else s.iter[b] = iter + 1;
  }
  __syncthreads();
  if (s_accept) { /* setCandidate(xs_try_, us_try_) */
    const double al = s_alpha;
    for (int e = threadIdx.x; e < (T + 1) * nx; e += blockDim.x) xs[e] = xs[e] + al * dx[e];
    for (int e = threadIdx.x; e < T * nu; e += blockDim.x) { double v = us[e]; v += al * du[e]; us[e] = v; }
  }
}

/* problems with done == 0 -> list_a, counts[2] */
__global__ void k_solve_running(SolveDev s, int B) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= B) return;
  if (s.done[b] == 0) s.list_a[atomicAdd(&s.counts[2], 1)] = b;
}

__global__ void k_solve_init(SolveDev s, PolyModelDev m, const double* xs0, const double* us0, int B, int T, int nx, int nu,
                             double reg0) {
  const size_t nxs = (size_t)B * (T + 1) * nx, nus = (size_t)B * T * nu;
  for (size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x; e < nxs; e += (size_t)gridDim.x * blockDim.x) {
    const size_t b = e / ((size_t)(T + 1) * nx), r = e % ((size_t)(T + 1) * nx);
    s.xs[e] = (r < (size_t)nx) ? m.x0[b * nx + r] : (xs0 ? xs0[e] : 0.0);   /* xs_[0] = x0 (src/sqp.cpp:86) */
  }
  for (size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x; e < nus; e += (size_t)gridDim.x * blockDim.x)
    s.us[e] = us0 ? us0[e] : 0.0;
  for (size_t b = blockIdx.x * (size_t)blockDim.x + threadIdx.x; b < (size_t)B; b += (size_t)gridDim.x * blockDim.x) {
    s.preg[b] = reg0; s.dreg[b] = reg0; s.iter[b] = 0; s.done[b] = 0; s.steplength[b] = 1.0; s.cost[b] = 0.0;
  }
}

}  // namespace mimgpu
