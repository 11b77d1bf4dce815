/*
 * kernels_rec.cuh -- CSQP direction kernel for the SMALL shapes (UR5 class, nx <= 16): the ADMM loop of
 * SolverCSQP::computeDirection (src/csqp.cpp:454-517) on per-knot RECORDS with the iterates on chip.
 *
 * Why.  On these shapes the sweep-by-sweep streaming kernel (kernels_stream.cuh) moves a knot as five bulk
 * copies of 144 ... 1152 bytes plus ~90 eight-byte asynchronous copies out of eight different arrays, and
 * writes six small vectors back per knot.  Measured on UR5 (profiles/variants_r02_v6_stream_nt.log,
 * profiles/ncu_full_r02_v7_csqp_stream_ur5.csv): 93-108 k QP/s whether a problem has one, two or three warps
 * and whether 8 or 15 problems share an SM, 990 warp instructions per knot step, 1.4 TB/s -- it is bound by
 * the instructions it spends on moving small pieces, not by occupancy, barriers, DRAM or arithmetic.
 *
 * What changes in the multi-warp form (arithmetic and its order are those of the streaming sweeps, results
 * are bit-identical; the one-warp form further down keeps less on chip, see there):
 *   - everything a light iteration reads that does NOT change between two rho updates sits in one record
 *     per knot in a scratch array, [Fx | Fu | K | Gx | Gu | Lx | Lu | Vxx_fs | fs_{t+1} | rho | g | lb | ub],
 *     written by the CTA after every full Riccati sweep: a sweep moves a knot with ONE bulk copy (3.2 KB
 *     for UR5) through the TMA + mbarrier ring;
 *   - everything that DOES change per iteration (dx~ du~ y z Vx Qu k of the whole horizon) lives in shared
 *     memory for the whole ADMM loop and goes to the member arrays before a full sweep reads them and once
 *     at the end: the light iterations issue no other global load or store.
 * Full sweeps, warm start and KKT are the code the streaming kernel runs (kernels_ric.cuh / kernels_generic.cuh).
 * Default CSQP kernel of the shapes in MIMGPU_REC_SHAPES, in its one-warp-per-problem form (below);
 * MIMGPU_CSQP_REC=0 / MIMGPU_CSQP_STREAM=1 give the streaming kernel, MIMGPU_REC_W1=0 the multi-warp form.
 *
 * Reference functions restated: backwardPass_without_rho_update (src/csqp.cpp:1025-1108), forwardPass
 * (src/csqp.cpp:630-647), update_lagrangian_parameters (src/csqp.cpp:1196-1340), update_rho_vec /
 * apply_rho_update (src/csqp.cpp:519-582).
 */
#pragma once
#include "kernels_cta.cuh"

namespace mimgpu {

/* offsets (doubles) inside one knot record */
struct RecLay {
  int Fx, Fu, K, Gx, Gu, Lx, Lu, Vf, fs, rho, g, lb, ub, size;
};
template <int NX, int NU>
__host__ __device__ inline RecLay rec_layout(int ng) {
  RecLay L;
  L.Fx = 0; L.Fu = NX * NX; L.K = L.Fu + NX * NU; L.Gx = L.K + NX * NU; L.Gu = L.Gx + ng * NX;
  L.Lx = L.Gu + ng * NU; L.Lu = L.Lx + NX; L.Vf = L.Lu + NU; L.fs = L.Vf + NX; L.rho = L.fs + NX;
  L.g = L.rho + ng; L.lb = L.g + ng; L.ub = L.lb + ng;
  L.size = (L.ub + ng + 1) & ~1; /* 16-byte multiple: one bulk copy */
  return L;
}

struct RecSmem {
  int bars, hdx, hdu, hy, hz, hQu, hk, hVx, tmp, d, Qx, x0, x1, u, dz, yy, red, uni, total;
};
template <int NX, int NU>
__host__ __device__ inline RecSmem rec_smem(int T, int ng, int ngT, int ns, int rec_size, int generic_total,
                                            bool w1 = false) {
  RecSmem s;
  const int ngm = ng > ngT ? ng : ngT;
  const int nc = T * ng + ngT;
  int o = 0;
#define RC_TAKE(n) (o += (((n) + 1) & ~1), o - (((n) + 1) & ~1))
  s.bars = RC_TAKE(16); /* 8 ring barriers + 2 of the full sweeps */
  /* one warp per problem: only y, z and k / Qu (one array, solved in place) stay on chip; dx~ du~ stay in the
   * member arrays (read one knot ahead into a register), Vx is carried in a register -- 13.7 KB instead of
   * 27.6 KB for UR5 T=40, 15 problems per SM instead of 8 */
  s.hdx = RC_TAKE(w1 ? 0 : (T + 1) * NX); s.hdu = RC_TAKE(w1 ? 0 : T * NU); s.hy = RC_TAKE(nc); s.hz = RC_TAKE(nc);
  s.hQu = RC_TAKE(w1 ? 0 : T * NU); s.hk = RC_TAKE(T * NU); s.hVx = RC_TAKE(w1 ? 0 : (T + 1) * NX);
  s.tmp = RC_TAKE(NX); s.d = RC_TAKE(ngm); s.Qx = RC_TAKE(NX); s.x0 = RC_TAKE(NX); s.x1 = RC_TAKE(NX);
  s.u = RC_TAKE(NU); s.dz = RC_TAKE(ngm); s.yy = RC_TAKE(ngm); s.red = RC_TAKE(w1 ? 0 : 64); /* one warp reduces by shuffles */
  s.uni = o; /* union: Riccati layout of the full sweeps | ns record stages */
  const int a = ns * rec_size;
  o += (a > generic_total ? a : generic_total);
#undef RC_TAKE
  s.total = o;
  return s;
}

/* ring of `ns` record stages fed by one elected thread (lane 0 of the last warp, as in kernels_stream.cuh) */
struct RecPipe {
  uint64_t* bars;
  double* stages;
  const double* rec; /* the problem's records, knot-major */
  int stage, ns;     /* doubles per record, ring depth */
  int islot, wslot;
  uint32_t wphase;
  __device__ __forceinline__ void issue(int t) {
    if (threadIdx.x == blockDim.x - 32) {
      uint64_t* bar = bars + islot;
      mbar_expect_tx(bar, (uint32_t)stage * 8u);
      tma_bulk_g2s(stages + (size_t)islot * stage, rec + (size_t)t * stage, (uint32_t)stage * 8u, bar);
    }
    if (++islot == ns) islot = 0;
  }
  __device__ __forceinline__ int acquire_slot() {
    mbar_wait(bars + wslot, wphase);
    const int r = wslot;
    if (++wslot == ns) { wslot = 0; wphase ^= 1u; }
    return r;
  }
};

/* member arrays -> records.  all = false: only what a full sweep changes (K, Vxx_fs, rho). */
template <int NX, int NU>
static __device__ __noinline__ void rec_build(const Prob& p, double* rec, const RecLay& L, bool all) {
  const int T = p.T, ng = p.ng, tid = threadIdx.x, nth = blockDim.x;
  auto blk = [&](const double* src, int n, int off) { /* src[t][n] -> rec[t][off .. off + n) */
    for (int e = tid; e < T * n; e += nth) {
      const int t = e / n, i = e - t * n;
      rec[(size_t)t * L.size + off + i] = src[e];
    }
  };
  blk(p.K, NX * NU, L.K);
  blk(p.Vxx_fs, NX, L.Vf);
  if (ng) blk(p.rho, ng, L.rho);
  if (all) {
    blk(p.Fx, NX * NX, L.Fx);
    blk(p.Fu, NX * NU, L.Fu);
    blk(p.Lx, NX, L.Lx);
    blk(p.Lu, NU, L.Lu);
    blk(p.fs + NX, NX, L.fs); /* fs_{t+1} */
    if (ng) {
      blk(p.Gx, ng * NX, L.Gx);
      blk(p.Gu, ng * NU, L.Gu);
      blk(p.g, ng, L.g);
      blk(p.glb, ng, L.lb);
      blk(p.gub, ng, L.ub);
    }
  }
  fence_proxy_async(); /* generic-proxy writes, read back by TMA after the next CTA barrier */
}

/* iterates: member arrays <-> shared memory */
template <int NX, int NU>
static __device__ __noinline__ void rec_load(const Prob& p, const RecSmem& s, bool all) {
  extern __shared__ __align__(16) double sm_stream[];
  double* sm = sm_stream;
  const int T = p.T, nc = p.T * p.ng + p.ngT, tid = threadIdx.x, nth = blockDim.x;
  for (int e = tid; e < (T + 1) * NX; e += nth) sm[s.hVx + e] = p.Vx[e];
  for (int e = tid; e < T * NU; e += nth) { sm[s.hQu + e] = p.Qu[e]; sm[s.hk + e] = p.kff[e]; }
  if (all) {
    for (int e = tid; e < (T + 1) * NX; e += nth) sm[s.hdx + e] = p.dx[e];
    for (int e = tid; e < T * NU; e += nth) sm[s.hdu + e] = p.du[e];
    for (int e = tid; e < nc; e += nth) { sm[s.hy + e] = p.y[e]; sm[s.hz + e] = p.z[e]; }
  }
}
template <int NX, int NU>
static __device__ __noinline__ void rec_flush(const Prob& p, const RecSmem& s, bool fwd_ran, bool all) {
  extern __shared__ __align__(16) double sm_stream[];
  const double* sm = sm_stream;
  const int T = p.T, nc = p.T * p.ng + p.ngT, tid = threadIdx.x, nth = blockDim.x;
  for (int e = tid; e < (T + 1) * NX; e += nth) {
    const double v = sm[s.hdx + e];
    p.dx[e] = v;
    if (fwd_ran) p.dxt[e] = v; /* dx_ = dxtilde_ after update_lagrangian_parameters (src/csqp.cpp:1222-1226) */
  }
  for (int e = tid; e < T * NU; e += nth) {
    const double v = sm[s.hdu + e];
    p.du[e] = v;
    if (fwd_ran) p.dut[e] = v;
  }
  for (int e = tid; e < nc; e += nth) { p.y[e] = sm[s.hy + e]; p.z[e] = sm[s.hz + e]; }
  if (all) {
    for (int e = tid; e < (T + 1) * NX; e += nth) p.Vx[e] = sm[s.hVx + e];
    for (int e = tid; e < T * NU; e += nth) { p.Qu[e] = sm[s.hQu + e]; p.kff[e] = sm[s.hk + e]; }
  }
}

/* one warp per problem: y, z, k only */
static __device__ __noinline__ void rec_load_w1(const Prob& p, const RecSmem& s, bool all) {
  extern __shared__ __align__(16) double sm_stream[];
  double* sm = sm_stream;
  const int T = p.T, nc = p.T * p.ng + p.ngT, NU = p.nu, lane = threadIdx.x;
  for (int e = lane; e < T * NU; e += 32) sm[s.hk + e] = p.kff[e];
  if (all)
    for (int e = lane; e < nc; e += 32) { sm[s.hy + e] = p.y[e]; sm[s.hz + e] = p.z[e]; }
}
/* final: also k, and dxtilde_ / dutilde_ := dx_ / du_ (what update_lagrangian_parameters leaves, src/csqp.cpp:1222-1226) */
static __device__ __noinline__ void rec_flush_w1(const Prob& p, const RecSmem& s, bool fwd_ran, bool final, bool with_k) {
  extern __shared__ __align__(16) double sm_stream[];
  const double* sm = sm_stream;
  const int T = p.T, nc = p.T * p.ng + p.ngT, NX = p.nx, NU = p.nu, lane = threadIdx.x;
  for (int e = lane; e < nc; e += 32) { p.y[e] = sm[s.hy + e]; p.z[e] = sm[s.hz + e]; }
  if (final) {
    if (with_k)
      for (int e = lane; e < T * NU; e += 32) p.kff[e] = sm[s.hk + e];
    if (fwd_ran) {
      for (int e = lane; e < (T + 1) * NX; e += 32) p.dxt[e] = p.dx[e];
      for (int e = lane; e < T * NU; e += 32) p.dut[e] = p.du[e];
    }
  }
}

/* ------------------------------------------------------- light backward sweep */
/* SolverCSQP::backwardPass_without_rho_update (src/csqp.cpp:1025-1108); the gains k_t = Quu_t^-1 Qu_t
 * (src/csqp.cpp:1096-1097) for all knots after the sweep, one thread per knot. */
template <int NX, int NU, int SPLIT>
static __device__ __noinline__ void rec_backward_light(const Prob& p, const RecSmem& s_, const RecLay& L_, RecPipe& pipe_,
                                                       double sigma) {
  extern __shared__ __align__(16) double sm_stream[];
  double* sm = sm_stream;
  /* by-value copies: a struct reached through a reference lives in local memory, and every inline-asm
   * "memory" clobber (mbarrier, TMA) would force its fields to be re-read there (1081 LDL/STL, 8.9
   * long-scoreboard stall slots per issue in the first capture) */
  const RecSmem s = s_;
  const RecLay L = L_;
  RecPipe pipe = pipe_;
  const int T = p.T, ng = p.ng, ngT = p.ngT, tid = threadIdx.x, nth = blockDim.x;
  double *tmp = sm + s.tmp, *sd = sm + s.d, *sQx = sm + s.Qx;
  double *hdx = sm + s.hdx, *hdu = sm + s.hdu, *hy = sm + s.hy, *hz = sm + s.hz, *hQu = sm + s.hQu, *hVx = sm + s.hVx,
         *hk = sm + s.hk;
  __syncthreads(); /* every stage is free, the iterates of the forward sweep are visible */
  for (int j = 0; j < pipe.ns - 1; ++j)
    if (T - 1 - j >= 0) pipe.issue(T - 1 - j);
  /* terminal knot (src/csqp.cpp:1045-1052) */
  if (ngT) {
    for (int r = tid; r < ngT; r += nth) {
      double d = hy[(size_t)T * ng + r];
      d -= p.rho[(size_t)T * ng + r] * hz[(size_t)T * ng + r];
      sd[r] = d;
    }
    __syncthreads();
  }
  for (int i = tid; i < NX; i += nth) {
    double v = p.LxT[i];
    v -= sigma * hdx[T * NX + i];
    if (ngT) {
      double acc = 0.0;
      for (int r = 0; r < ngT; ++r) acc += p.GxT[r + ngT * i] * sd[r];
      v += acc;
    }
    hVx[T * NX + i] = v;
  }
  __syncthreads(); /* sd (terminal) consumed */
  const int grp = tid / SPLIT, q = tid % SPLIT, ngrp = nth / SPLIT;
  for (int t = T - 1; t >= 0; --t) {
    const double* st = sm + s.uni + pipe.acquire_slot() * L.size;
    const double *sFx = st + L.Fx, *sFu = st + L.Fu, *sK = st + L.K, *sGx = st + L.Gx, *sGu = st + L.Gu;
    /* P0 (src/csqp.cpp:1056-1066): Vx' + Vxx' fs', y - rho z.  hVx[t + 1] was written by this same thread. */
    for (int i = tid; i < NX; i += nth) tmp[i] = st[L.Vf + i] + hVx[(t + 1) * NX + i];
    for (int r = tid; r < ng; r += nth) {
      double d = hy[t * ng + r];
      d -= st[L.rho + r] * hz[t * ng + r];
      sd[r] = d;
    }
    __syncthreads(); /* A: tmp, d visible; knot t+1 completely consumed */
    /* P1: Qx, Qu (src/csqp.cpp:1068-1093) */
    for (int o0 = 0; o0 < NX + NU; o0 += ngrp) {
      const int o = o0 + grp;
      const bool valid = o < NX + NU;
      const int oc = valid ? o : 0;
      const bool isx = oc < NX;
      const double* Fc = isx ? sFx + NX * oc : sFu + NX * (oc - NX);
      double acc = 0.0;
#pragma unroll
      for (int k = 0; k < (NX + SPLIT - 1) / SPLIT; ++k) {
        const int kk = k * SPLIT + q;
        if (kk < NX) acc = fma(Fc[kk], tmp[kk], acc);
      }
      double accg = 0.0;
      const bool gon = ng && (isx ? t > 0 : true);
      if (gon) {
        const double* Gc = isx ? sGx + ng * oc : sGu + ng * (oc - NX);
        for (int r = q; r < ng; r += SPLIT) accg = fma(Gc[r], sd[r], accg);
      }
      acc = split_reduce<SPLIT>(acc);
      accg = split_reduce<SPLIT>(accg);
      if (valid && q == 0) {
        if (isx) {
          double x = st[L.Lx + o];
          x -= sigma * hdx[t * NX + o];
          if (gon) x += accg;
          sQx[o] = x + acc;
        } else {
          const int a = o - NX;
          double x = st[L.Lu + a];
          x -= sigma * hdu[t * NU + a];
          if (gon) x += accg;
          x += acc;
          hQu[t * NU + a] = x;
        }
      }
    }
    __syncthreads(); /* B */
    if (t - (pipe.ns - 1) >= 0) pipe.issue(t - (pipe.ns - 1)); /* into the stage of knot t+1, free since barrier A */
    /* P2: Vx = Qx - K^T Qu (src/csqp.cpp:1101-1104) */
    for (int i = tid; i < NX; i += nth) {
      double acc = 0.0;
#pragma unroll
      for (int a = 0; a < NU; ++a) acc = fma(sK[a * NX + i], hQu[t * NU + a], acc);
      hVx[t * NX + i] = sQx[i] - acc;
    }
  }
  __syncthreads();
  /* gains: serial substitution in the reference's order, L = LLT(Quu_t) of the last full sweep */
  for (int t = tid; t < T; t += nth) {
    const double* Lt = p.L + (size_t)t * NU * NU;
    double x[NU];
#pragma unroll
    for (int i = 0; i < NU; ++i) x[i] = hQu[t * NU + i];
#pragma unroll
    for (int i = 0; i < NU; ++i) {
      double v = x[i];
#pragma unroll
      for (int j = 0; j < i; ++j) v -= Lt[i + NU * j] * x[j];
      x[i] = v / Lt[i + NU * i];
    }
#pragma unroll
    for (int i = NU - 1; i >= 0; --i) {
      double v = x[i];
#pragma unroll
      for (int j = i + 1; j < NU; ++j) v -= Lt[j + NU * i] * x[j];
      x[i] = v / Lt[i + NU * i];
    }
#pragma unroll
    for (int i = 0; i < NU; ++i) hk[t * NU + i] = x[i];
  }
  pipe_ = pipe;
  __syncthreads();
}

/* ------------------------------------------------- forward pass + ADMM update */
/* SolverCSQP::forwardPass (src/csqp.cpp:630-647) fused with update_lagrangian_parameters
 * (src/csqp.cpp:1196-1340): dx~, du~, z, y, residual norms. */
template <int NX, int NU, int SPLIT>
static __device__ __noinline__ void rec_forward_update(const Prob& p, const RecSmem& s_, const RecLay& L_, RecPipe& pipe_,
                                                       const mimgpu_csqp_params& q_, int iter, AdmmNorms& n_) {
  extern __shared__ __align__(16) double sm_stream[];
  double* sm = sm_stream;
  const RecSmem s = s_;
  const RecLay L = L_;
  RecPipe pipe = pipe_;
  AdmmNorms n;
  const int T = p.T, ng = p.ng, ngT = p.ngT, tid = threadIdx.x, nth = blockDim.x, lane = tid & 31;
  double *su = sm + s.u, *sdz = sm + s.dz, *sy = sm + s.yy, *red = sm + s.red;
  double *hdx = sm + s.hdx, *hdu = sm + s.hdu, *hy = sm + s.hy, *hz = sm + s.hz, *hk = sm + s.hk;
  double* xs[2] = {sm + s.x0, sm + s.x1};
  const bool check = (iter % q_.rho_update_interval) == 0;
  const bool heur = q_.update_rho_with_heuristic != 0;
  const double alpha = q_.alpha;
  n.primal = n.dual = n.primal_rel = n.dual_rel = -CUDART_INF; /* src/csqp.cpp:1201-1204 */
  __syncthreads();
  for (int j = 0; j < pipe.ns - 1; ++j)
    if (j < T) pipe.issue(j);
  for (int i = tid; i < NX; i += nth) xs[0][i] = p.dxt[i];
  __syncthreads();
  const int grp = tid / SPLIT, q = tid % SPLIT, ngrp = nth / SPLIT;
  const int h = lane >> 4, l16 = lane & 15, w16 = (tid >> 5) * 16, n16 = (nth >> 5) * 16;
  for (int t = 0; t < T; ++t) {
    const double* st = sm + s.uni + pipe.acquire_slot() * L.size;
    const double *sFx = st + L.Fx, *sFu = st + L.Fu, *sK = st + L.K, *sGx = st + L.Gx, *sGu = st + L.Gu;
    const double* x = xs[t & 1];
    double* xn = xs[(t + 1) & 1];
    /* F1: du~ = -k - K dx~ (src/csqp.cpp:636-637); dx_ := dx~ (src/csqp.cpp:1222-1226, 1241-1242) */
    for (int o0 = 0; o0 < NU; o0 += ngrp) {
      const int a = o0 + grp;
      const bool valid = a < NU;
      const double* Kr = sK + NX * (valid ? a : 0);
      double acc = 0.0;
#pragma unroll
      for (int k = 0; k < (NX + SPLIT - 1) / SPLIT; ++k) {
        const int kk = k * SPLIT + q;
        if (kk < NX) acc = fma(Kr[kk], x[kk], acc);
      }
      acc = split_reduce<SPLIT>(acc);
      if (valid && q == 0) {
        double u = -hk[t * NU + a];
        u -= acc;
        su[a] = u;
        hdu[t * NU + a] = u;
      }
    }
    for (int i = tid; i < NX; i += nth) hdx[t * NX + i] = x[i];
    __syncthreads(); /* 1: du~ visible; knot t-1 completely consumed */
    if (t + pipe.ns - 1 < T) pipe.issue(t + pipe.ns - 1);
    /* F2: dx~+ = fs+ + Fx dx~ + Fu du~ (src/csqp.cpp:638-644) and the constraint rows of knot t
     * (src/csqp.cpp:1229-1239); two half-warps share one output (even / odd contraction index) */
    for (int o0 = 0; o0 < NX + ng; o0 += n16) {
      const int o = o0 + w16 + l16;
      const bool valid = o < NX + ng;
      const int oc = valid ? o : 0;
      const bool isx = oc < NX;
      const int ld = isx ? NX : ng, row = isx ? oc : oc - NX;
      const double* A1 = (isx ? sFx : sGx) + row;
      const double* A2 = (isx ? sFu : sGu) + row;
      double a1 = 0.0, a2 = 0.0;
#pragma unroll 6
      for (int j = h; j < NX; j += 2) a1 = fma(A1[ld * j], x[j], a1);
#pragma unroll 4
      for (int a = h; a < NU; a += 2) a2 = fma(A2[ld * a], su[a], a2);
      a1 += __shfl_xor_sync(0xffffffffu, a1, 16);
      a2 += __shfl_xor_sync(0xffffffffu, a2, 16);
      if (valid && h == 0) {
        if (isx) {
          double xv = st[L.fs + o];
          xv += a1;
          xv += a2;
          xn[o] = xv;
        } else {
          const int r = o - NX;
          const double zp = hz[t * ng + r];
          double c = 0.0;
          c += a1;
          c += a2;
          double zr = alpha * c;
          zr += (1. - alpha) * zp;
          const double rho = st[L.rho + r], irho = 1. / rho, yo = hy[t * ng + r];
          double z = zr + yo * irho;
          const double lo = st[L.lb + r] - st[L.g + r], hi = st[L.ub + r] - st[L.g + r];
          z = (z < lo) ? lo : z;
          z = (hi < z) ? hi : z;
          const double yn = yo + rho * (zr - z);
          hz[t * ng + r] = z;
          hy[t * ng + r] = yn;
          if (check) {
            sdz[r] = rho * (z - zp); sy[r] = yn;
            n.primal = nmax(n.primal, fabs(c - z));
            n.primal_rel = nmax(n.primal_rel, fabs(c));
            n.primal_rel = nmax(n.primal_rel, fabs(z));
            if (heur) {
              n.dual = nmax(n.dual, fabs(rho * (z - zp)));
              n.dual_rel = nmax(n.dual_rel, fabs(yn));
            }
          }
        }
      }
    }
    __syncthreads(); /* 2: dx~+ visible */
    if (check && !heur && ng) { /* F3: |Gx^T rho dz|, |Gu^T rho dz|, |Gx^T y|, |Gu^T y| (src/csqp.cpp:1259-1278) */
      for (int o = tid; o < NX + NU; o += nth) {
        const double* Gc = (o < NX) ? sGx + ng * o : sGu + ng * (o - NX);
        double a1 = 0.0, a2 = 0.0;
        for (int r = 0; r < ng; ++r) { a1 = fma(Gc[r], sdz[r], a1); a2 = fma(Gc[r], sy[r], a2); }
        n.dual = nmax(n.dual, fabs(a1));
        n.dual_rel = nmax(n.dual_rel, fabs(a2));
      }
    }
  }
  __syncthreads(); /* F3 of the last knot is done with sdz / sy */
  /* terminal knot (src/csqp.cpp:1283-1338) */
  const double* x = xs[T & 1];
  for (int i = tid; i < NX; i += nth) hdx[T * NX + i] = x[i];
  if (ngT) {
    const size_t o = (size_t)T * ng;
    for (int r = tid; r < ngT; r += nth) {
      const double zp = hz[o + r];
      double acc = 0.0;
      for (int j = 0; j < NX; ++j) acc += p.GxT[r + ngT * j] * x[j];
      double c = 0.0;
      c += acc;
      double zr = alpha * c;
      zr += (1. - alpha) * zp;
      const double rho = p.rho[o + r], irho = 1. / rho, yo = hy[o + r];
      double z = zr + yo * irho;
      const double lo = p.glbT[r] - p.gT[r], hi = p.gubT[r] - p.gT[r];
      z = (z < lo) ? lo : z;
      z = (hi < z) ? hi : z;
      const double yn = yo + rho * (zr - z);
      hz[o + r] = z;
      hy[o + r] = yn;
      if (check) {
        sdz[r] = rho * (z - zp); sy[r] = yn;
        n.primal = nmax(n.primal, fabs(c - z));
        n.primal_rel = nmax(n.primal_rel, fabs(c));
        n.primal_rel = nmax(n.primal_rel, fabs(z));
        if (heur) {
          n.dual = nmax(n.dual, fabs(rho * (z - zp)));
          n.dual_rel = nmax(n.dual_rel, fabs(yn));
        }
      }
    }
    if (check && !heur) {
      __syncthreads();
      for (int j = tid; j < NX; j += nth) {
        double a1 = 0.0, a2 = 0.0;
        for (int r = 0; r < ngT; ++r) { a1 += p.GxT[r + ngT * j] * sdz[r]; a2 += p.GxT[r + ngT * j] * sy[r]; }
        n.dual = nmax(n.dual, fabs(a1));
        n.dual_rel = nmax(n.dual_rel, fabs(a2));
      }
    }
  }
  if (check) {
    n.primal = cta_max(n.primal, red);
    n.dual = cta_max(n.dual, red);
    n.primal_rel = cta_max(n.primal_rel, red);
    n.dual_rel = cta_max(n.dual_rel, red);
  }
  n_ = n;
  pipe_ = pipe;
  __syncthreads();
}

/* ======================================================= one WARP per problem */
/* The same two sweeps for a 32-thread CTA (MIMGPU_REC_W1): no CTA barrier (a warp barrier orders the
 * shared-memory hand-offs), one lane per output in the backward mat-vecs (a column of [Fx Fu] / [Gx Gu] is
 * contiguous in the record: 16-byte loads, no shuffle reduction), Vx carried in a register from knot to
 * knot.  A third of the instructions of the multi-warp sweeps per knot; sums run over the contraction
 * index in two interleaved partial sums (the reference: one), inside the 1e-9 bar. */
template <int NX, int NU>
static __device__ __noinline__ void rec_backward_light_w1(const Prob& p, const RecSmem& s_, const RecLay& L_, RecPipe& pipe_,
                                                          double sigma) {
  static_assert(NX + NU <= 32 && NX % 2 == 0, "one lane per row of [Qx; Qu]");
  extern __shared__ __align__(16) double sm_stream[];
  double* sm = sm_stream;
  const RecSmem s = s_;
  const RecLay L = L_;
  RecPipe pipe = pipe_;
  constexpr int N = NX + NU;
  const int T = p.T, ng = p.ng, ngT = p.ngT, lane = threadIdx.x;
  double *tmp = sm + s.tmp, *sd = sm + s.d;
  double *hy = sm + s.hy, *hz = sm + s.hz, *hk = sm + s.hk; /* hk holds Qu_t until the gains are solved in place */
  __syncwarp();
  for (int j = 0; j < pipe.ns - 1; ++j)
    if (T - 1 - j >= 0) pipe.issue(T - 1 - j);
  /* terminal knot (src/csqp.cpp:1045-1052) */
  if (ngT) {
    for (int r = lane; r < ngT; r += 32) {
      double d = hy[T * ng + r];
      d -= p.rho[(size_t)T * ng + r] * hz[T * ng + r];
      sd[r] = d;
    }
    __syncwarp();
  }
  double vx = 0.0; /* lane i < NX: Vx_{t+1}[i] */
  if (lane < NX) {
    double v = p.LxT[lane];
    v -= sigma * p.dx[(size_t)T * NX + lane];
    if (ngT) {
      double acc = 0.0;
      for (int r = 0; r < ngT; ++r) acc += p.GxT[r + ngT * lane] * sd[r];
      v += acc;
    }
    p.Vx[(size_t)T * NX + lane] = v;
    vx = v;
  }
  __syncwarp(); /* sd (terminal) consumed */
  const bool isx = lane < NX;
  /* dx~_t / du~_t of the previous forward sweep, read from the member arrays one knot ahead */
  const double* dsrc = isx ? p.dx + lane : p.du + (lane - NX);
  const int dstride = isx ? NX : NU;
  double dcur = (lane < N && T > 0) ? dsrc[(size_t)(T - 1) * dstride] : 0.0;
  for (int t = T - 1; t >= 0; --t) {
    const double dnext = (lane < N && t > 0) ? dsrc[(size_t)(t - 1) * dstride] : 0.0;
    const double* st = sm + s.uni + pipe.acquire_slot() * L.size;
    /* P0 (src/csqp.cpp:1056-1066) */
    if (isx) tmp[lane] = st[L.Vf + lane] + vx;
    for (int r = lane; r < ng; r += 32) {
      double d = hy[t * ng + r];
      d -= st[L.rho + r] * hz[t * ng + r];
      sd[r] = d;
    }
    __syncwarp(); /* A: tmp, d visible; every lane is done with knot t+1 */
    if (t - (pipe.ns - 1) >= 0) pipe.issue(t - (pipe.ns - 1)); /* into the stage of knot t+1 */
    /* P1: row `lane` of [Qx; Qu] (src/csqp.cpp:1068-1093) */
    double qv = 0.0;
    if (lane < N) {
      const double2* Fc = reinterpret_cast<const double2*>(st + NX * lane); /* [Fx | Fu]: columns of NX */
      const double2* tv = reinterpret_cast<const double2*>(tmp);
      double a0 = 0.0, a1 = 0.0;
#pragma unroll
      for (int k = 0; k < NX / 2; ++k) {
        const double2 f = Fc[k], w = tv[k];
        a0 = fma(f.x, w.x, a0);
        a1 = fma(f.y, w.y, a1);
      }
      const bool gon = ng && (isx ? t > 0 : true);
      double accg = 0.0;
      if (gon) {
        const double* Gc = st + L.Gx + ng * lane; /* [Gx | Gu]: columns of ng */
        for (int r = 0; r < ng; ++r) accg = fma(Gc[r], sd[r], accg);
      }
      double x = st[L.Lx + lane]; /* [Lx | Lu] */
      x -= sigma * dcur;
      if (gon) x += accg;
      x += a0 + a1;
      qv = x;
      if (!isx) { hk[t * NU + lane - NX] = x; p.Qu[(size_t)t * NU + lane - NX] = x; }
    }
    dcur = dnext;
    __syncwarp(); /* B: Qu visible */
    /* P2: Vx = Qx - K^T Qu (src/csqp.cpp:1101-1104) */
    if (isx) {
      const double* sK = st + L.K;
      double acc = 0.0;
#pragma unroll
      for (int a = 0; a < NU; ++a) acc = fma(sK[a * NX + lane], hk[t * NU + a], acc);
      vx = qv - acc;
      p.Vx[(size_t)t * NX + lane] = vx;
    }
  }
  __syncwarp();
  /* gains (src/csqp.cpp:1096-1097): serial substitution in the reference's order, one lane per knot */
  for (int t = lane; t < T; t += 32) {
    const double* Lt = p.L + (size_t)t * NU * NU;
    double x[NU];
#pragma unroll
    for (int i = 0; i < NU; ++i) x[i] = hk[t * NU + i];
#pragma unroll
    for (int i = 0; i < NU; ++i) {
      double v = x[i];
#pragma unroll
      for (int j = 0; j < i; ++j) v -= Lt[i + NU * j] * x[j];
      x[i] = v / Lt[i + NU * i];
    }
#pragma unroll
    for (int i = NU - 1; i >= 0; --i) {
      double v = x[i];
#pragma unroll
      for (int j = i + 1; j < NU; ++j) v -= Lt[j + NU * i] * x[j];
      x[i] = v / Lt[i + NU * i];
    }
#pragma unroll
    for (int i = 0; i < NU; ++i) hk[t * NU + i] = x[i];
  }
  pipe_ = pipe;
  __syncwarp();
}

template <int NX, int NU>
static __device__ __noinline__ void rec_forward_update_w1(const Prob& p, const RecSmem& s_, const RecLay& L_, RecPipe& pipe_,
                                                          const mimgpu_csqp_params& q_, int iter, AdmmNorms& n_) {
  static_assert(NU <= 16, "F1 gives a half-warp lane to every control row");
  extern __shared__ __align__(16) double sm_stream[];
  double* sm = sm_stream;
  const RecSmem s = s_;
  const RecLay L = L_;
  RecPipe pipe = pipe_;
  AdmmNorms n;
  const int T = p.T, ng = p.ng, ngT = p.ngT, lane = threadIdx.x;
  double *su = sm + s.u, *sdz = sm + s.dz, *sy = sm + s.yy;
  double *hy = sm + s.hy, *hz = sm + s.hz, *hk = sm + s.hk;
  double* xs[2] = {sm + s.x0, sm + s.x1};
  const bool check = (iter % q_.rho_update_interval) == 0;
  const bool heur = q_.update_rho_with_heuristic != 0;
  const double alpha = q_.alpha;
  n.primal = n.dual = n.primal_rel = n.dual_rel = -CUDART_INF; /* src/csqp.cpp:1201-1204 */
  __syncwarp();
  for (int j = 0; j < pipe.ns - 1; ++j)
    if (j < T) pipe.issue(j);
  if (lane < NX) xs[0][lane] = p.dxt[lane];
  __syncwarp();
  const int h = lane >> 4, l16 = lane & 15;
  for (int t = 0; t < T; ++t) {
    const double* st = sm + s.uni + pipe.acquire_slot() * L.size;
    const double *sFx = st + L.Fx, *sFu = st + L.Fu, *sK = st + L.K, *sGx = st + L.Gx, *sGu = st + L.Gu;
    const double* x = xs[t & 1];
    double* xn = xs[(t + 1) & 1];
    /* F1: du~ = -k - K dx~ (src/csqp.cpp:636-637): row l16, even / odd contraction index per half-warp */
    {
      const bool valid = l16 < NU;
      const double* Kr = sK + NX * (valid ? l16 : 0);
      double acc = 0.0;
#pragma unroll
      for (int k = 0; k < NX / 2; ++k) acc = fma(Kr[2 * k + h], x[2 * k + h], acc);
      acc += __shfl_xor_sync(0xffffffffu, acc, 16);
      if (valid && h == 0) {
        double u = -hk[t * NU + l16];
        u -= acc;
        su[l16] = u;
        p.du[(size_t)t * NU + l16] = u;
      }
    }
    if (lane < NX) p.dx[(size_t)t * NX + lane] = x[lane]; /* dx_ := dx~ (src/csqp.cpp:1222-1226, 1241-1242) */
    __syncwarp(); /* 1: du~ visible; every lane is done with knot t-1 */
    if (t + pipe.ns - 1 < T) pipe.issue(t + pipe.ns - 1);
    /* F2: dx~+ = fs+ + Fx dx~ + Fu du~ (src/csqp.cpp:638-644) and the constraint rows of knot t
     * (src/csqp.cpp:1229-1239); two half-warps share one output */
    for (int o0 = 0; o0 < NX + ng; o0 += 16) {
      const int o = o0 + l16;
      const bool valid = o < NX + ng;
      const int oc = valid ? o : 0;
      const bool isx = oc < NX;
      const int ld = isx ? NX : ng, row = isx ? oc : oc - NX;
      const double* A1 = (isx ? sFx : sGx) + row;
      const double* A2 = (isx ? sFu : sGu) + row;
      double a1 = 0.0, a2 = 0.0;
#pragma unroll
      for (int j = 0; j < NX / 2; ++j) a1 = fma(A1[ld * (2 * j + h)], x[2 * j + h], a1);
#pragma unroll
      for (int a = 0; a < NU / 2; ++a) a2 = fma(A2[ld * (2 * a + h)], su[2 * a + h], a2);
      a1 += __shfl_xor_sync(0xffffffffu, a1, 16);
      a2 += __shfl_xor_sync(0xffffffffu, a2, 16);
      if (valid && h == 0) {
        if (isx) {
          double xv = st[L.fs + o];
          xv += a1;
          xv += a2;
          xn[o] = xv;
        } else {
          const int r = o - NX;
          const double zp = hz[t * ng + r];
          double c = 0.0;
          c += a1;
          c += a2;
          double zr = alpha * c;
          zr += (1. - alpha) * zp;
          const double rho = st[L.rho + r], irho = 1. / rho, yo = hy[t * ng + r];
          double z = zr + yo * irho;
          const double lo = st[L.lb + r] - st[L.g + r], hi = st[L.ub + r] - st[L.g + r];
          z = (z < lo) ? lo : z;
          z = (hi < z) ? hi : z;
          const double yn = yo + rho * (zr - z);
          hz[t * ng + r] = z;
          hy[t * ng + r] = yn;
          if (check) {
            sdz[r] = rho * (z - zp); sy[r] = yn;
            n.primal = nmax(n.primal, fabs(c - z));
            n.primal_rel = nmax(n.primal_rel, fabs(c));
            n.primal_rel = nmax(n.primal_rel, fabs(z));
            if (heur) {
              n.dual = nmax(n.dual, fabs(rho * (z - zp)));
              n.dual_rel = nmax(n.dual_rel, fabs(yn));
            }
          }
        }
      }
    }
    __syncwarp(); /* 2: dx~+ visible */
    if (check && !heur && ng) { /* F3 (src/csqp.cpp:1259-1278) */
      for (int o = lane; o < NX + NU; o += 32) {
        const double* Gc = sGx + ng * o; /* [Gx | Gu]: columns of ng */
        double a1 = 0.0, a2 = 0.0;
        for (int r = 0; r < ng; ++r) { a1 = fma(Gc[r], sdz[r], a1); a2 = fma(Gc[r], sy[r], a2); }
        n.dual = nmax(n.dual, fabs(a1));
        n.dual_rel = nmax(n.dual_rel, fabs(a2));
      }
    }
  }
  __syncwarp(); /* F3 of the last knot is done with sdz / sy */
  /* terminal knot (src/csqp.cpp:1283-1338) */
  const double* x = xs[T & 1];
  if (lane < NX) p.dx[(size_t)T * NX + lane] = x[lane];
  if (ngT) {
    const size_t o = (size_t)T * ng;
    for (int r = lane; r < ngT; r += 32) {
      const double zp = hz[o + r];
      double acc = 0.0;
      for (int j = 0; j < NX; ++j) acc += p.GxT[r + ngT * j] * x[j];
      double c = 0.0;
      c += acc;
      double zr = alpha * c;
      zr += (1. - alpha) * zp;
      const double rho = p.rho[o + r], irho = 1. / rho, yo = hy[o + r];
      double z = zr + yo * irho;
      const double lo = p.glbT[r] - p.gT[r], hi = p.gubT[r] - p.gT[r];
      z = (z < lo) ? lo : z;
      z = (hi < z) ? hi : z;
      const double yn = yo + rho * (zr - z);
      hz[o + r] = z;
      hy[o + r] = yn;
      if (check) {
        sdz[r] = rho * (z - zp); sy[r] = yn;
        n.primal = nmax(n.primal, fabs(c - z));
        n.primal_rel = nmax(n.primal_rel, fabs(c));
        n.primal_rel = nmax(n.primal_rel, fabs(z));
        if (heur) {
          n.dual = nmax(n.dual, fabs(rho * (z - zp)));
          n.dual_rel = nmax(n.dual_rel, fabs(yn));
        }
      }
    }
    if (check && !heur) {
      __syncwarp();
      for (int j = lane; j < NX; j += 32) {
        double a1 = 0.0, a2 = 0.0;
        for (int r = 0; r < ngT; ++r) { a1 += p.GxT[r + ngT * j] * sdz[r]; a2 += p.GxT[r + ngT * j] * sy[r]; }
        n.dual = nmax(n.dual, fabs(a1));
        n.dual_rel = nmax(n.dual_rel, fabs(a2));
      }
    }
  }
  if (check) { /* one warp: shuffles only */
    for (int o = 16; o; o >>= 1) {
      n.primal = nmax(n.primal, __shfl_xor_sync(0xffffffffu, n.primal, o));
      n.dual = nmax(n.dual, __shfl_xor_sync(0xffffffffu, n.dual, o));
      n.primal_rel = nmax(n.primal_rel, __shfl_xor_sync(0xffffffffu, n.primal_rel, o));
      n.dual_rel = nmax(n.dual_rel, __shfl_xor_sync(0xffffffffu, n.dual_rel, o));
    }
  }
  n_ = n;
  pipe_ = pipe;
  __syncwarp();
}

/* ------------------------------------------------------------------ the kernel */
struct RecKPar {
  mimgpu_csqp_params q;
  int with_kkt, ns, generic_total;
  double* rec;          /* [B][T][RecLay::size] */
};

/* SolverCSQP::computeDirection (src/csqp.cpp:454-517): control flow of k_csqp_direction_stream. */
#ifndef MIMGPU_REC_W1_MINB
#define MIMGPU_REC_W1_MINB 16
#endif
template <int NX, int NU, int SPLIT, bool W1 = false> /* W1: one warp per problem (32-thread CTA) */
__global__ void __launch_bounds__(W1 ? 32 : StreamTune<NX, NU>::MAXT, W1 ? MIMGPU_REC_W1_MINB : StreamTune<NX, NU>::MINB)
k_csqp_direction_rec(KArgs a, RecKPar kp) {
  static_assert(NX % 2 == 0 && NU % 2 == 0, "TMA bulk copies need 16-byte blocks");
  extern __shared__ __align__(16) double sm_stream[];
  if ((int)blockIdx.x >= a.n_active) return;
  const int b = a.active ? a.active[blockIdx.x] : (int)blockIdx.x + a.first;
  Prob p = bind(a, b);
  const RecLay L = rec_layout<NX, NU>(p.ng);
  const RecSmem ss = rec_smem<NX, NU>(p.T, p.ng, p.ngT, kp.ns, L.size, kp.generic_total, W1);
  double* sm = sm_stream;
  double* gsm = sm + ss.uni;
  const Smem gs = smem_layout(NX, NU, p.ng, p.ngT, a.stage_fx);
  constexpr bool USE_MMA = (NX % 4 == 0);
  const RicSmem rs = ric_layout<NX, NU>(p.ng, p.ngT, true, W1 ? 1 : 2); /* W1: single-buffered full sweeps */
  uint64_t* ric_bars = reinterpret_cast<uint64_t*>(sm + ss.bars) + 8;
  uint32_t ric_seq = 0u;
  double* rec = kp.rec + (size_t)b * p.T * L.size;
  RecPipe pipe;
  pipe.bars = reinterpret_cast<uint64_t*>(sm + ss.bars);
  pipe.stages = sm + ss.uni;
  pipe.rec = rec;
  pipe.stage = L.size; pipe.ns = kp.ns;
  pipe.islot = 0; pipe.wslot = 0; pipe.wphase = 0u;
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
    if (!st) { __syncthreads(); csqp_forward_eq_ool(p, gsm, gs); }
  }
  const int interval = q.rho_update_interval;
  bool in_smem = false;   /* the iterates of the ADMM loop are in shared memory (members stale) */
  bool fwd_ran = false;   /* dx~ / du~ have been produced by this call */
  bool rec_made = false;  /* the static part of the records has been written */
  if (!st) {
    for (it = 1; it < q.max_qp_iters + 1; ++it) {
      if (it % interval == 1 || interval == 1) {
        if (in_smem) { /* the full sweep reads dx_ du_ y z from the member arrays */
          __syncthreads();
          if constexpr (W1) rec_flush_w1(p, ss, fwd_ran, false, false);
          else rec_flush<NX, NU>(p, ss, fwd_ran, false);
          __syncthreads();
        }
        if constexpr (USE_MMA) st = ric_backward_mma<MODE_FULL, NX, NU>(p, gsm, rs, rp, ric_bars, ric_seq);
        else st = riccati_backward_ool<MODE_FULL>(p, gsm, gs, rp, a.stage_fx);
        if (st) break;
        __syncthreads(); /* K, Vxx_fs, k, Vx, Qu of the sweep are visible to the whole CTA */
        rec_build<NX, NU>(p, rec, L, !rec_made);
        rec_made = true;
        if constexpr (W1) rec_load_w1(p, ss, !in_smem);
        else rec_load<NX, NU>(p, ss, !in_smem);
        in_smem = true;
      } else {
        if constexpr (W1) rec_backward_light_w1<NX, NU>(p, ss, L, pipe, q.sigma);
        else rec_backward_light<NX, NU, SPLIT>(p, ss, L, pipe, q.sigma);
      }
      __syncthreads();
      if constexpr (W1) rec_forward_update_w1<NX, NU>(p, ss, L, pipe, q, it, n);
      else rec_forward_update<NX, NU, SPLIT>(p, ss, L, pipe, q, it, n);
      fwd_ran = true;
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
  if (in_smem) { /* after a failed full sweep Vx / Qu / k stay as that sweep left them, as in the streaming kernel */
    if constexpr (W1) rec_flush_w1(p, ss, fwd_ran, true, st == 0);
    else rec_flush<NX, NU>(p, ss, fwd_ran, st == 0);
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    a.status[b] = st; a.qp_iters[b] = it; a.converged[b] = converged;
    p.scal[MIMGPU_S_RHO_SPARSE * B + b] = rho_sparse;
    p.scal[MIMGPU_S_NORM_PRIMAL * B + b] = n.primal; p.scal[MIMGPU_S_NORM_DUAL * B + b] = n.dual;
    p.scal[MIMGPU_S_NORM_PRIMAL_REL * B + b] = n.primal_rel; p.scal[MIMGPU_S_NORM_DUAL_REL * B + b] = n.dual_rel;
  }
  if (!st && kp.with_kkt) csqp_kkt_ool(p, gsm, gs);
}

/* ------------------------------------------------------------------- dispatch */
#define MIMGPU_REC_SHAPES(X) X(12, 6, MIMGPU_C1_SPLIT) X(16, 8, 4) X(8, 4, 4) X(4, 2, 2) X(6, 2, 2) X(10, 4, 2)

template <int NX, int NU, int SPLIT>
static RecPlan rec_plan_w(const mimgpu_shape& sh, int stage_fx, int max_smem, int w1) {
  RecPlan pl{};
  pl.w1 = w1;
  const Smem gs = smem_layout(NX, NU, sh.ng, sh.ngT, stage_fx);
  pl.generic_total = gs.total;
  if (NX % 4 == 0) {
    const RicSmem rs = ric_layout<NX, NU>(sh.ng, sh.ngT, true, pl.w1 ? 1 : 2);
    if (rs.total > pl.generic_total) pl.generic_total = rs.total;
  }
  const int ngm = sh.ng > sh.ngT ? sh.ng : sh.ngT;
  int nt = SPLIT * (NX + NU);
  if (2 * (NX + ngm) > nt) nt = 2 * (NX + ngm);
  nt = ((nt + 31) / 32) * 32;
  if (nt > StreamTune<NX, NU>::MAXT) nt = StreamTune<NX, NU>::MAXT;
  nt = stream_nt_override(nt, StreamTune<NX, NU>::MAXT);
  if (pl.w1) nt = 32;
  pl.nt = nt;
  /* the full sweeps keep their vector packs: same limit as the streaming kernel */
  const int nvmax = StreamCfg<NX, NU>::VPT * nt;
  if (StreamCfg<NX, NU>::nv_bwd(sh.ng) > nvmax) return pl;
  const RecLay L = rec_layout<NX, NU>(sh.ng);
  pl.rec_doubles = (size_t)sh.T * L.size;
  /* at least four problems per SM, else the streaming kernel (iterates in HBM) serves the shape */
  int ns_max = 3; /* MIMGPU_REC_NS=2..8: depth of the record ring */
  if (const char* e = getenv("MIMGPU_REC_NS")) { const int v = atoi(e); if (v >= 2 && v <= 8) ns_max = v; }
  /* one warp per problem: resident problems count at least as much as ring depth (measured on UR5: 3 stages x 15 per
   * SM 232.5 k, 2 stages x 16 per SM 229.8 k, 4 stages x 12 per SM 192 k QP/s), so a shallower ring is taken when it
   * lets one more problem in */
  if (pl.w1 && ns_max > 2 && !getenv("MIMGPU_REC_NS")) {
    auto per_sm = [&](int ns) {
      const int b = rec_smem<NX, NU>(sh.T, sh.ng, sh.ngT, ns, L.size, pl.generic_total, true).total * 8 + 1024;
      const int n = 233472 / b;
      return n > MIMGPU_REC_W1_MINB ? MIMGPU_REC_W1_MINB : n;
    };
    while (ns_max > 2 && per_sm(ns_max - 1) > per_sm(ns_max)) --ns_max;
  }
  for (int ns = ns_max; ns >= 2; --ns) {
    const RecSmem ss = rec_smem<NX, NU>(sh.T, sh.ng, sh.ngT, ns, L.size, pl.generic_total, pl.w1 != 0);
    int bytes = ss.total * 8;
    if (bytes > max_smem || 4 * (bytes + 1024) > 233472) continue;
    /* MIMGPU_REC_MAXCTA=n: at most n resident problems per SM (the launch asks for 1/n of the shared memory),
     * so that the records of the resident wave (n x 148 x T x record) can be sized against the L2 */
    if (const char* e = getenv("MIMGPU_REC_MAXCTA")) {
      const int ncta = atoi(e);
      if (ncta >= 1 && ncta <= 32) {
        const int want = (233472 / ncta - 1024) & ~15;
        if (want > bytes && want <= max_smem) bytes = want;
      }
    }
    pl.ns = ns; pl.smem_bytes = bytes;
    const cudaError_t e1 = pl.w1 ? cudaFuncSetAttribute(k_csqp_direction_rec<NX, NU, SPLIT, true>,
                                                        cudaFuncAttributeMaxDynamicSharedMemorySize, bytes)
                                 : cudaFuncSetAttribute(k_csqp_direction_rec<NX, NU, SPLIT, false>,
                                                        cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
    if (e1 != cudaSuccess) {
      cudaGetLastError();
      return pl;
    }
    pl.ok = 1;
    return pl;
  }
  return pl;
}

/* one warp per problem (the default; MIMGPU_REC_W1=0 turns it off) when the shape allows it and the vector pack of
 * the full sweeps fits 32 threads, else the multi-warp CTA */
template <int NX, int NU, int SPLIT>
static RecPlan rec_plan_t(const mimgpu_shape& sh, int stage_fx, int max_smem) {
  bool want_w1 = NX + NU <= 32 && NU <= 16;
  if (const char* e = getenv("MIMGPU_REC_W1")) want_w1 = want_w1 && atoi(e) != 0;
  if (want_w1) {
    const RecPlan pl = rec_plan_w<NX, NU, SPLIT>(sh, stage_fx, max_smem, 1);
    if (pl.ok) return pl;
  }
  return rec_plan_w<NX, NU, SPLIT>(sh, stage_fx, max_smem, 0);
}

static inline RecPlan rec_plan(const mimgpu_shape& sh, int stage_fx, int max_smem) {
#define X(NX_, NU_, SP_) \
  if (sh.nx == NX_ && sh.nu == NU_) return rec_plan_t<NX_, NU_, SP_>(sh, stage_fx, max_smem);
  MIMGPU_REC_SHAPES(X)
#undef X
  return RecPlan{};
}

static inline int rec_csqp_launch(const KArgs& a, const mimgpu_csqp_params& q, int with_kkt, const RecPlan& pl, double* rec,
                                  cudaStream_t st) {
  RecKPar kp;
  kp.q = q; kp.with_kkt = with_kkt; kp.ns = pl.ns; kp.generic_total = pl.generic_total; kp.rec = rec;
#define X(NX_, NU_, SP_)                                                                        \
  if (a.sh.nx == NX_ && a.sh.nu == NU_) {                                                       \
    if (pl.w1) k_csqp_direction_rec<NX_, NU_, SP_, true><<<a.n_active, 32, pl.smem_bytes, st>>>(a, kp);   \
    else k_csqp_direction_rec<NX_, NU_, SP_, false><<<a.n_active, pl.nt, pl.smem_bytes, st>>>(a, kp);   \
    return 0;                                                                                   \
  }
  MIMGPU_REC_SHAPES(X)
#undef X
  return -1;
}

}  // namespace mimgpu
