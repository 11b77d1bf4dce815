/*
 * kernels_stream.cuh -- streaming ADMM sweeps of the CSQP direction kernel for medium
 * shapes (compile-time NX, NU; run-time ng, ngT).
 *
 * Of the 50..1000 ADMM iterations of one SolverCSQP::computeDirection (src/csqp.cpp:454-517)
 * all but one in `rho_update_interval` run the light backward sweep
 * (backwardPass_without_rho_update, src/csqp.cpp:1025-1108), and every iteration runs
 * forwardPass (src/csqp.cpp:630-647) + update_lagrangian_parameters (src/csqp.cpp:1196-1340).
 * Both are chains of T dependent mat-vec steps whose operands (Fx, Fu, K, Gx, Gu of the knot)
 * are read exactly once per sweep: the sweeps are bound by how fast a problem's knots can be
 * streamed to its CTA, not by arithmetic.  So:
 *   - one CTA owns one problem and keeps all sweep vectors in shared memory;
 *   - one elected thread streams the knot's matrices HBM/L2 -> shared memory with TMA bulk
 *     copies (cp.async.bulk ... mbarrier::complete_tx), NS knots ahead of the consumer;
 *   - the knot's small vectors are software-prefetched through registers one knot ahead;
 *   - every dot product is split over SPLIT lanes (conflict-free shared-memory walks, 4x
 *     shorter dependent FMA chains) and reduced with warp shuffles;
 *   - the gain solve k = Quu^-1 Qu is off the recursion's critical path (Vx = Qx - K^T Qu
 *     does not need k): it is done for all knots at once after the sweep, one thread per knot.
 * The full Riccati sweeps (1 in `rho_update_interval` iterations) still use the shape-generic
 * code of kernels_generic.cuh.
 */
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "kernels_generic.cuh"

namespace mimgpu {

/* ------------------------------------------------------------ TMA / mbarrier */
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred P1;\n"
      "LAB_WAIT:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n"
      "@P1 bra DONE;\n"
      "bra LAB_WAIT;\n"
      "DONE:\n"
      "}\n" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}
__device__ __forceinline__ void tma_bulk_g2s(double* dst, const double* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   smem_u32(dst)),
               "l"(src), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
/* A `double*` that reaches a __noinline__ function as a parameter is a GENERIC pointer: the compiler
 * then emits LD.E with 64-bit address arithmetic instead of LDS.  Round-tripping it through the
 * shared window tells the address-space inference what it points to. */
template <class T>
__device__ __forceinline__ T* as_shared(T* p) {
  extern __shared__ __align__(16) unsigned char mg_dyn_smem_base[]; /* aliases the kernel's dynamic array */
  const unsigned off = (unsigned)__cvta_generic_to_shared(const_cast<void*>(static_cast<const void*>(p))) -
                       (unsigned)__cvta_generic_to_shared(mg_dyn_smem_base);
  return reinterpret_cast<T*>(mg_dyn_smem_base + off);
}
/* Ampere-style asynchronous copies for the small per-knot vectors (8 bytes per element) */
__device__ __forceinline__ void cp_async8(double* smem_dst, const double* gsrc) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(smem_u32(smem_dst)), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int NPEND>
__device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;" ::"n"(NPEND) : "memory");
}
/* L2 prefetch of a contiguous global range (16-byte aligned, multiple of 16 bytes) */
__device__ __forceinline__ void l2_prefetch(const double* src, uint32_t bytes) {
  asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(src), "r"(bytes) : "memory");
}
/* generic-proxy global writes (K, Quu_llt of a full sweep) -> later TMA reads of them */
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async;" ::: "memory"); }

/* --------------------------------------------------------------- smem layout */
template <int NX, int NU>
struct StreamCfg {
  static constexpr int VPT = 3;  /* vector-pack elements per thread (upper bound) */
  static constexpr int VD = 4;   /* ring depth of the vector packs (cp.async, VD - 1 knots ahead) */
  /* stage = the knot's matrices, in consumption order */
  __host__ __device__ static int stage_doubles(int ng) { return NX * NX + 2 * NX * NU + ng * (NX + NU); }
  __host__ __device__ static int nv_bwd(int ng) { return 3 * NX + 2 * NU + 3 * ng; }
  __host__ __device__ static int nv_fwd(int ng) { return NX + NU + 6 * ng; }
};

struct StreamSmem {
  int bars, vec0, nvs, tmp, d, Qx, Qu, Vx, x0, x1, u, dz, yy, red, uni, stage, total;
};
template <int NX, int NU>
__host__ __device__ inline StreamSmem stream_layout(int ng, int ngT, int ns, int generic_total) {
  StreamSmem s;
  const int ngm = ng > ngT ? ng : ngT;
  const int nvb = StreamCfg<NX, NU>::nv_bwd(ng), nvf = StreamCfg<NX, NU>::nv_fwd(ngm);
  const int nv = ((nvb > nvf ? nvb : nvf) + 1) & ~1;
  int o = 0;
#define ST_TAKE(n) (o += (((n) + 1) & ~1), o - (((n) + 1) & ~1))
  s.bars = ST_TAKE(8);
  constexpr int vd_ = StreamCfg<NX, NU>::VD;
  s.nvs = nv; s.vec0 = ST_TAKE(vd_ * nv);
  s.tmp = ST_TAKE(NX); s.d = ST_TAKE(ngm); s.Qx = ST_TAKE(NX); s.Qu = ST_TAKE(NU); s.Vx = ST_TAKE(NX);
  s.x0 = ST_TAKE(NX); s.x1 = ST_TAKE(NX); s.u = ST_TAKE(NU); s.dz = ST_TAKE(ngm); s.yy = ST_TAKE(ngm);
  s.red = ST_TAKE(64);
  s.uni = o;  /* union: generic Riccati layout | NS streaming stages */
  s.stage = StreamCfg<NX, NU>::stage_doubles(ng);
  const int a = ns * s.stage;
  o += (a > generic_total ? a : generic_total);
#undef ST_TAKE
  s.total = o;
  return s;
}

/* --------------------------------------------------------------- the pipeline */
template <int NX, int NU>
struct Pipe {
  uint64_t* bars;
  double* stages;
  int stage, ns, ng;
  uint32_t n_issue, n_wait; /* load sequence numbers since the last reset (kept for the callers that set them) */
  /* slot / phase of the next issue and of the next wait, advanced without integer division */
  int islot, wslot;
  uint32_t wphase;
  int no_prefetch = 0; /* debug switch (MIMGPU_DBG bit 3) */

  /* The elected producer is lane 0 of the LAST warp: in the phases where the sweeps issue, that
   * warp has no outputs of its own (the mat-vec outputs are dealt to the first warps), so the
   * ~40 instructions of the five bulk copies are off the other warps' critical path. */
  __device__ __forceinline__ void issue(const Prob& p, int t) {
    if (threadIdx.x == blockDim.x - 32) {
      double* dst = stages + (size_t)islot * stage;
      uint64_t* bar = bars + islot;
      mbar_expect_tx(bar, (uint32_t)stage * 8u);
      const size_t tt = (size_t)t;
      tma_bulk_g2s(dst, p.Fx + tt * (NX * NX), NX * NX * 8, bar);
      tma_bulk_g2s(dst + NX * NX, p.Fu + tt * (NX * NU), NX * NU * 8, bar);
      tma_bulk_g2s(dst + NX * NX + NX * NU, p.K + tt * (NX * NU), NX * NU * 8, bar);
      if (ng) {
        tma_bulk_g2s(dst + NX * NX + 2 * NX * NU, p.Gx + tt * (size_t)(ng * NX), (uint32_t)ng * NX * 8, bar);
        tma_bulk_g2s(dst + NX * NX + 2 * NX * NU + ng * NX, p.Gu + tt * (size_t)(ng * NU), (uint32_t)ng * NU * 8, bar);
      }
    }
    if (++islot == ns) islot = 0;
    ++n_issue;
  }
  /* DIRECT mode (shapes whose knot does not fit two shared-memory stages): the sweeps read the
   * knot's matrices straight from global memory with coalesced accesses; the producer only pulls
   * the knot into L2 a few knots ahead so that those reads are L2 hits. */
  __device__ __forceinline__ void prefetch(const Prob& p, int t) {
    if (threadIdx.x == blockDim.x - 32 && !no_prefetch) {
      const size_t tt = (size_t)t;
      l2_prefetch(p.Fx + tt * (NX * NX), NX * NX * 8);
      l2_prefetch(p.Fu + tt * (NX * NU), NX * NU * 8);
      l2_prefetch(p.K + tt * (NX * NU), NX * NU * 8);
      if (ng) {
        l2_prefetch(p.Gx + tt * (size_t)(ng * NX), (uint32_t)ng * NX * 8);
        l2_prefetch(p.Gu + tt * (size_t)(ng * NU), (uint32_t)ng * NU * 8);
      }
    }
  }
  /* same, returning the slot index (callers that address the stage from the shared array itself) */
  __device__ __forceinline__ int acquire_slot() {
    mbar_wait(bars + wslot, wphase);
    const int r = wslot;
    if (++wslot == ns) { wslot = 0; wphase ^= 1u; }
    ++n_wait;
    return r;
  }
  /* all threads: wait for the next stage in consumption order, return its base */
  __device__ __forceinline__ const double* acquire() {
    mbar_wait(bars + wslot, wphase);
    const double* r = stages + (size_t)wslot * stage;
    if (++wslot == ns) { wslot = 0; wphase ^= 1u; }
    ++n_wait;
    return r;
  }
};

/* vector pack: which global array / stride / offset each element of the pack comes from */
struct VecSrc {
  const double* base; /* address of the element at knot 0 (nullptr: no element) */
  int stride;         /* doubles per knot */
};

template <int SPLIT>
__device__ __forceinline__ double split_reduce(double v) {
#pragma unroll
  for (int o = SPLIT / 2; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

/* ------------------------------------------------------- light backward sweep */
/* SolverCSQP::backwardPass_without_rho_update (src/csqp.cpp:1025-1108), gains deferred to
 * stream_solve_gains.  Writes Vx_, Qu_. */
template <int NX, int NU, int SPLIT, bool DIRECT = false>
static __device__ __noinline__ void stream_backward_light(const Prob& p, double* sm, const StreamSmem& s, Pipe<NX, NU>& pipe,
                                      double sigma) {
  using C = StreamCfg<NX, NU>;
  extern __shared__ __align__(16) double sm_stream[]; /* the same dynamic array the kernel declares: keeps LDS/STS */
  sm = sm_stream;
  const int T = p.T, ng = p.ng, ngT = p.ngT, tid = threadIdx.x, nth = blockDim.x;
  double *tmp = sm + s.tmp, *sd = sm + s.d, *sQx = sm + s.Qx, *sQu = sm + s.Qu, *sVx = sm + s.Vx;
  constexpr int VD = C::VD;
  double* vecb = sm + s.vec0;
  const int nvs = s.nvs;
  auto vecof = [&](int t) -> double* { return vecb + (size_t)(t % VD) * nvs; };
  /* pack of knot t: [Lx NX][dx NX][Vxx_fs NX][Lu NU][du NU][y ng][z ng][rho ng] */
  const int oLx = 0, oDx = NX, oVf = 2 * NX, oLu = 3 * NX, oDu = 3 * NX + NU, oY = 3 * NX + 2 * NU, oZ = oY + ng,
            oR = oZ + ng, nv = oR + ng;
  VecSrc vs[C::VPT];
#pragma unroll
  for (int j = 0; j < C::VPT; ++j) {
    const int e = tid + j * nth;
    vs[j].base = nullptr; vs[j].stride = 0;
    if (e < oDx) { vs[j].base = p.Lx + e; vs[j].stride = NX; }
    else if (e < oVf) { vs[j].base = p.dx + (e - oDx); vs[j].stride = NX; }
    else if (e < oLu) { vs[j].base = p.Vxx_fs + (e - oVf); vs[j].stride = NX; }
    else if (e < oDu) { vs[j].base = p.Lu + (e - oLu); vs[j].stride = NU; }
    else if (e < oY) { vs[j].base = p.du + (e - oDu); vs[j].stride = NU; }
    else if (e < oZ) { vs[j].base = p.y + (e - oY); vs[j].stride = ng; }
    else if (e < oR) { vs[j].base = p.z + (e - oZ); vs[j].stride = ng; }
    else if (e < nv) { vs[j].base = p.rho + (e - oR); vs[j].stride = ng; }
  }
  /* the pack of knot t goes global -> shared memory with 8-byte cp.async, VD - 1 knots ahead of
   * its consumer (one commit group per knot, empty groups keep the count uniform) */
  auto issue_pack = [&](int t) {
    if (t >= 0) {
      double* dst = vecof(t);
#pragma unroll
      for (int j = 0; j < C::VPT; ++j)
        if (vs[j].base) cp_async8(dst + tid + j * nth, vs[j].base + (size_t)t * vs[j].stride);
    }
    cp_async_commit();
  };
  constexpr int PD = 3; /* DIRECT: L2 prefetch distance in knots */
  __syncthreads(); /* every stage and vector buffer is free */
  if (DIRECT) {
    for (int j = 0; j < PD; ++j)
      if (T - 1 - j >= 0) pipe.prefetch(p, T - 1 - j);
  } else {
    for (int j = 0; j < pipe.ns - 1; ++j)
      if (T - 1 - j >= 0) pipe.issue(p, T - 1 - j);
  }
  for (int j = 0; j < VD - 1; ++j) issue_pack(T - 1 - j);
  cp_async_wait<VD - 2>(); /* pack of knot T-1 has landed (made visible by the barrier below) */
  /* terminal knot (src/csqp.cpp:1045-1052) */
  if (ngT) {
    for (int r = tid; r < ngT; r += nth) {
      double d = p.y[(size_t)T * ng + r];
      d -= p.rho[(size_t)T * ng + r] * p.z[(size_t)T * ng + r];
      sd[r] = d;
    }
    __syncthreads();
  }
  for (int i = tid; i < NX; i += nth) {
    double v = p.LxT[i];
    v -= sigma * p.dx[(size_t)T * NX + i];
    if (ngT) {
      double acc = 0.0;
      for (int r = 0; r < ngT; ++r) acc += p.GxT[r + ngT * i] * sd[r];
      v += acc;
    }
    sVx[i] = v;
    p.Vx[(size_t)T * NX + i] = v;
  }
  __syncthreads(); /* sd (terminal) consumed, pack of knot T-1 visible */
  /* P0 of knot T-1 */
  for (int i = tid; i < NX; i += nth) tmp[i] = vecof(T - 1)[oVf + i] + sVx[i];
  for (int r = tid; r < ng; r += nth) {
    const double* v = vecof(T - 1);
    double d = v[oY + r];
    d -= v[oR + r] * v[oZ + r];
    sd[r] = d;
  }
  const int grp = tid / SPLIT, q = tid % SPLIT, ngrp = nth / SPLIT;
  for (int t = T - 1; t >= 0; --t) {
    const double* st = DIRECT ? nullptr : sm + s.uni + pipe.acquire_slot() * s.stage;
    __syncthreads(); /* A: tmp, d visible; knot t+1 completely consumed */
    issue_pack(t - (VD - 1));   /* into the slot of knot t+1, consumed before barrier A */
    cp_async_wait<VD - 2>();    /* this thread's share of pack t-1 has landed; barrier B publishes it */
    const double *sFx = st, *sFu = st + NX * NX, *sK = st + NX * NX + NX * NU, *sGx = sK + NX * NU,
                 *sGu = sGx + ng * NX;
    if (DIRECT) {
      sFx = p.Fx + (size_t)t * NX * NX; sFu = p.Fu + (size_t)t * NX * NU; sK = p.K + (size_t)t * NX * NU;
      sGx = p.Gx + (size_t)t * ng * NX; sGu = p.Gu + (size_t)t * ng * NU;
    }
    const double* v = vecof(t);
    /* P1: Qx, Qu (src/csqp.cpp:1068-1093) */
    for (int o0 = 0; o0 < NX + NU; o0 += ngrp) {
      const int o = o0 + grp;
      const bool valid = o < NX + NU;
      const int oc = valid ? o : 0;
      const bool isx = oc < NX;   /* idle groups read column 0 of Fx: never a negative offset */
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
          double x = v[oLx + o];
          x -= sigma * v[oDx + o];
          if (gon) x += accg;
          sQx[o] = x + acc;
        } else {
          const int a = o - NX;
          double x = v[oLu + a];
          x -= sigma * v[oDu + a];
          if (gon) x += accg;
          x += acc;
          sQu[a] = x;
          p.Qu[(size_t)t * NU + a] = x;
        }
      }
    }
    __syncthreads(); /* B */
    /* refill the stage of knot t+1 (free since barrier A) while the last warp has no P2 rows */
    if (DIRECT) {
      if (t - PD >= 0) pipe.prefetch(p, t - PD);
    } else if (t - (pipe.ns - 1) >= 0) {
      pipe.issue(p, t - (pipe.ns - 1));
    }
    /* P2: Vx = Qx - K^T Qu (src/csqp.cpp:1101-1104); P0 of knot t-1 */
    for (int i = tid; i < NX; i += nth) {
      double acc = 0.0;
#pragma unroll
      for (int a = 0; a < NU; ++a) acc = fma(sK[a * NX + i], sQu[a], acc);
      const double vx = sQx[i] - acc;
      p.Vx[(size_t)t * NX + i] = vx;
      if (t > 0) tmp[i] = vecof(t - 1)[oVf + i] + vx;
    }
    if (t > 0) {
      const double* vn = vecof(t - 1);
      for (int r = tid; r < ng; r += nth) {
        double d = vn[oY + r];
        d -= vn[oR + r] * vn[oZ + r];
        sd[r] = d;
      }
    }
  }
  __syncthreads();
}

/* k_t = Quu_t^-1 Qu_t for all knots (src/csqp.cpp:1096-1097), one thread per knot, serial
 * substitution in the reference's order. */
template <int NU>
static __device__ __noinline__ void stream_solve_gains(const Prob& p) {
  for (int t = threadIdx.x; t < p.T; t += blockDim.x) {
    const double* L = p.L + (size_t)t * NU * NU;
    double x[NU];
#pragma unroll
    for (int i = 0; i < NU; ++i) x[i] = p.Qu[(size_t)t * NU + i];
#pragma unroll
    for (int i = 0; i < NU; ++i) {
      double v = x[i];
#pragma unroll
      for (int j = 0; j < i; ++j) v -= L[i + NU * j] * x[j];
      x[i] = v / L[i + NU * i];
    }
#pragma unroll
    for (int i = NU - 1; i >= 0; --i) {
      double v = x[i];
#pragma unroll
      for (int j = i + 1; j < NU; ++j) v -= L[j + NU * i] * x[j];
      x[i] = v / L[i + NU * i];
    }
#pragma unroll
    for (int i = 0; i < NU; ++i) p.kff[(size_t)t * NU + i] = x[i];
  }
  __syncthreads();
}

/* ------------------------------------------------- forward pass + ADMM update */
/* SolverCSQP::forwardPass (src/csqp.cpp:630-647) fused with update_lagrangian_parameters
 * (src/csqp.cpp:1196-1340): dx~, du~, z, y, dx_ := dx~, du_ := du~, residual norms. */
template <int NX, int NU, int SPLIT, bool DIRECT = false>
static __device__ __noinline__ void stream_forward_update(const Prob& p, double* sm, const StreamSmem& s, Pipe<NX, NU>& pipe,
                                      const mimgpu_csqp_params& q_, int iter, AdmmNorms& n) {
  using C = StreamCfg<NX, NU>;
  extern __shared__ __align__(16) double sm_stream[]; /* the same dynamic array the kernel declares: keeps LDS/STS */
  sm = sm_stream;
  const int T = p.T, ng = p.ng, ngT = p.ngT, tid = threadIdx.x, nth = blockDim.x, lane = tid & 31;
  double *su = sm + s.u, *sdz = sm + s.dz, *sy = sm + s.yy, *red = sm + s.red;
  double* xs[2] = {sm + s.x0, sm + s.x1};
  constexpr int VD = C::VD;
  double* vecb = sm + s.vec0;
  const int nvs = s.nvs;
  auto vecof = [&](int t) -> double* { return vecb + (size_t)(t % VD) * nvs; };
  const bool check = (iter % q_.rho_update_interval) == 0;
  const bool heur = q_.update_rho_with_heuristic != 0;
  const double alpha = q_.alpha;
  n.primal = n.dual = n.primal_rel = n.dual_rel = -CUDART_INF; /* src/csqp.cpp:1201-1204 */
  /* pack of knot t: [k NU][fs_{t+1} NX][y ng][z ng][rho ng][g ng][lb ng][ub ng] */
  const int oK = 0, oFs = NU, oY = NU + NX, oZ = oY + ng, oR = oZ + ng, oG = oR + ng, oLb = oG + ng, oUb = oLb + ng,
            nv = oUb + ng;
  VecSrc vs[C::VPT];
#pragma unroll
  for (int j = 0; j < C::VPT; ++j) {
    const int e = tid + j * nth;
    vs[j].base = nullptr; vs[j].stride = 0;
    if (e < oFs) { vs[j].base = p.kff + e; vs[j].stride = NU; }
    else if (e < oY) { vs[j].base = p.fs + NX + (e - oFs); vs[j].stride = NX; }
    else if (e < oZ) { vs[j].base = p.y + (e - oY); vs[j].stride = ng; }
    else if (e < oR) { vs[j].base = p.z + (e - oZ); vs[j].stride = ng; }
    else if (e < oG) { vs[j].base = p.rho + (e - oR); vs[j].stride = ng; }
    else if (e < oLb) { vs[j].base = p.g + (e - oG); vs[j].stride = ng; }
    else if (e < oUb) { vs[j].base = p.glb + (e - oLb); vs[j].stride = ng; }
    else if (e < nv) { vs[j].base = p.gub + (e - oUb); vs[j].stride = ng; }
  }
  auto issue_pack = [&](int t) { /* 8-byte cp.async, one commit group per knot (see the backward sweep) */
    if (t < T) {
      double* dst = vecof(t);
#pragma unroll
      for (int j = 0; j < C::VPT; ++j)
        if (vs[j].base) cp_async8(dst + tid + j * nth, vs[j].base + (size_t)t * vs[j].stride);
    }
    cp_async_commit();
  };
  constexpr int PD = 3;
  __syncthreads();
  if (DIRECT) {
    for (int j = 0; j < PD; ++j)
      if (j < T) pipe.prefetch(p, j);
  } else {
    for (int j = 0; j < pipe.ns - 1; ++j)
      if (j < T) pipe.issue(p, j);
  }
  for (int j = 0; j < VD - 1; ++j) issue_pack(j);
  for (int i = tid; i < NX; i += nth) xs[0][i] = p.dxt[i];
  cp_async_wait<VD - 2>(); /* pack of knot 0 */
  __syncthreads();
  const int grp = tid / SPLIT, q = tid % SPLIT, ngrp = nth / SPLIT;
  const int h = lane >> 4, l16 = lane & 15, w16 = (tid >> 5) * 16, n16 = (nth >> 5) * 16;
  for (int t = 0; t < T; ++t) {
    const double* st = DIRECT ? nullptr : sm + s.uni + pipe.acquire_slot() * s.stage;
    const double *sFx = st, *sFu = st + NX * NX, *sK = st + NX * NX + NX * NU, *sGx = sK + NX * NU,
                 *sGu = sGx + ng * NX;
    if (DIRECT) {
      sFx = p.Fx + (size_t)t * NX * NX; sFu = p.Fu + (size_t)t * NX * NU; sK = p.K + (size_t)t * NX * NU;
      sGx = p.Gx + (size_t)t * ng * NX; sGu = p.Gu + (size_t)t * ng * NU;
    }
    const double* x = xs[t & 1];
    double* xn = xs[(t + 1) & 1];
    const double* v = vecof(t);
    /* F1: du~ = -k - K dx~  (src/csqp.cpp:636-637); dx_ := dx~ (src/csqp.cpp:1222-1226, 1241-1242) */
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
        double u = -v[oK + a];
        u -= acc;
        su[a] = u;
        p.dut[(size_t)t * NU + a] = u;
        p.du[(size_t)t * NU + a] = u;
      }
    }
    for (int i = tid; i < NX; i += nth) p.dx[(size_t)t * NX + i] = x[i];
    __syncthreads(); /* 1: du~ visible; knot t-1 completely consumed */
    if (DIRECT) {
      if (t + PD < T) pipe.prefetch(p, t + PD);
    } else if (t + pipe.ns - 1 < T) {
      pipe.issue(p, t + pipe.ns - 1);
    }
    issue_pack(t + VD - 1);   /* into the slot of knot t-1 */
    cp_async_wait<VD - 2>();  /* this thread's share of pack t+1 has landed; barrier 2 publishes it */
    /* F2: dx~+ = fs+ + Fx dx~ + Fu du~ (src/csqp.cpp:638-644) and the constraint rows of knot t
     * (src/csqp.cpp:1229-1239); outputs o < NX are states, NX <= o < NX + ng constraint rows;
     * two half-warps share one output (even / odd contraction index) */
    for (int o0 = 0; o0 < NX + ng; o0 += n16) {
      const int o = o0 + w16 + l16;
      const bool valid = o < NX + ng;
      const int oc = valid ? o : 0;
      const bool isx = oc < NX;   /* idle lanes read row 0 of Fx: never a negative offset */
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
          double xv = v[oFs + o];
          xv += a1;
          xv += a2;
          xn[o] = xv;
          p.dxt[(size_t)(t + 1) * NX + o] = xv;
        } else {
          const int r = o - NX;
          const double zp = v[oZ + r];
          double c = 0.0;
          c += a1;
          c += a2;
          double zr = alpha * c;
          zr += (1. - alpha) * zp;
          const double rho = v[oR + r], irho = 1. / rho, yo = v[oY + r];
          double z = zr + yo * irho;
          const double lo = v[oLb + r] - v[oG + r], hi = v[oUb + r] - v[oG + r];
          z = (z < lo) ? lo : z;
          z = (hi < z) ? hi : z;
          const double yn = yo + rho * (zr - z);
          p.z[(size_t)t * ng + r] = z;
          p.y[(size_t)t * ng + r] = yn;
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
  for (int i = tid; i < NX; i += nth) p.dx[(size_t)T * NX + i] = x[i];
  if (ngT) {
    const size_t o = (size_t)T * ng;
    for (int r = tid; r < ngT; r += nth) {
      const double zp = p.z[o + r];
      double acc = 0.0;
      for (int j = 0; j < NX; ++j) acc += p.GxT[r + ngT * j] * x[j];
      double c = 0.0;
      c += acc;
      double zr = alpha * c;
      zr += (1. - alpha) * zp;
      const double rho = p.rho[o + r], irho = 1. / rho, yo = p.y[o + r];
      double z = zr + yo * irho;
      const double lo = p.glbT[r] - p.gT[r], hi = p.gubT[r] - p.gT[r];
      z = (z < lo) ? lo : z;
      z = (hi < z) ? hi : z;
      const double yn = yo + rho * (zr - z);
      p.z[o + r] = z;
      p.y[o + r] = yn;
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
  __syncthreads();
}

/* ------------------------------------------------------------ SQP forward pass */
/* stage of knot t: [Fx][Fu][K][Vxx_{t+1}] */
template <int NX, int NU>
struct PipeSqp {
  uint64_t* bars;
  double* stages;
  int ns;
  uint32_t n_issue, n_wait;
  int islot = 0, wslot = 0;     /* slot of the next issue / wait, advanced without integer division */
  uint32_t wphase = 0u;
  static constexpr int STAGE = 2 * NX * NX + 2 * NX * NU;
  __device__ __forceinline__ void issue(const Prob& p, int t) {
    const int slot = islot;
    if (++islot == ns) islot = 0;
    if (threadIdx.x == 0) {
      double* dst = stages + (size_t)slot * STAGE;
      uint64_t* bar = bars + slot;
      mbar_expect_tx(bar, (uint32_t)STAGE * 8u);
      tma_bulk_g2s(dst, p.Fx + (size_t)t * NX * NX, NX * NX * 8, bar);
      tma_bulk_g2s(dst + NX * NX, p.Fu + (size_t)t * NX * NU, NX * NU * 8, bar);
      tma_bulk_g2s(dst + NX * NX + NX * NU, p.K + (size_t)t * NX * NU, NX * NU * 8, bar);
      tma_bulk_g2s(dst + NX * NX + 2 * NX * NU, p.Vxx + (size_t)(t + 1) * NX * NX, NX * NX * 8, bar);
    }
    ++n_issue;
  }
  __device__ __forceinline__ const double* acquire() {
    const int slot = wslot;
    mbar_wait(bars + slot, wphase);
    if (++wslot == ns) { wslot = 0; wphase ^= 1u; }
    ++n_wait;
    return stages + (size_t)slot * STAGE;
  }
};

/* shared memory (doubles) of stream_sqp_forward next to the stages */
template <int NX, int NU>
struct SqpFwdSmem {
  static constexpr int NV = 3 * NX + 2 * NU;
  static constexpr int vec0 = 0, vec1 = NV, x0 = 2 * NV, x1 = x0 + NX, u = x1 + NX, dm = u + NU, lam0 = dm + NX,
                       lam1 = lam0 + NX, red = lam1 + NX, stages = red + 64;
};

/* SolverSQP::forwardPass (src/sqp.cpp:270-305) fused with checkKKTConditions
 * (src/sqp.cpp:244-268) and the gap norm / merit of computeDirection (src/sqp.cpp:229-236):
 * the knot's Fx, Fu, K, Vxx+ are streamed by TMA NS knots ahead, vectors through registers. */
template <int NX, int NU, int SPLIT>
static __device__ __noinline__ void stream_sqp_forward(const Prob& p, double* sm, PipeSqp<NX, NU>& pipe, double mu_dynamic) {
  sm = as_shared(sm);
  using S = SqpFwdSmem<NX, NU>;
  constexpr int N = NX + NU, VPT = 3;
  const int T = p.T, tid = threadIdx.x, nth = blockDim.x, lane = tid & 31;
  double *su = sm + S::u, *sdm = sm + S::dm, *red = sm + S::red;
  /* pack of knot t: [k NU][fs_{t+1} NX][Vx_{t+1} NX][Lx NX][Lu NU] */
  constexpr int oK = 0, oFs = NU, oVx = NU + NX, oLx = NU + 2 * NX, nv = S::NV;
  VecSrc vs[VPT];
  double vr[VPT];
#pragma unroll
  for (int j = 0; j < VPT; ++j) {
    const int e = tid + j * nth;
    vs[j].base = nullptr; vs[j].stride = 0;
    if (e < oFs) { vs[j].base = p.kff + e; vs[j].stride = NU; }
    else if (e < oVx) { vs[j].base = p.fs + NX + (e - oFs); vs[j].stride = NX; }
    else if (e < oLx) { vs[j].base = p.Vx + NX + (e - oVx); vs[j].stride = NX; }
    else if (e < oLx + NX) { vs[j].base = p.Lx + (e - oLx); vs[j].stride = NX; }
    else if (e < nv) { vs[j].base = p.Lu + (e - oLx - NX); vs[j].stride = NU; }
  }
  __syncthreads();
  for (int j = 0; j < pipe.ns - 1; ++j)
    if (j < T) pipe.issue(p, j);
#pragma unroll
  for (int j = 0; j < VPT; ++j) {
    const int e = tid + j * nth;
    if (vs[j].base) (sm + S::vec0)[e] = vs[j].base[0];
    vr[j] = (vs[j].base && T >= 2) ? vs[j].base[(size_t)vs[j].stride] : 0.0;
  }
  double xg = 0.0, ug = 0.0, gap = 0.0, kkt = 0.0;
  /* knot 0: dx_0 is whatever the member holds (0); lag_mul_0 = Vxx_0 (dx_0 - fs_0) + Vx_0 */
  for (int i = tid; i < NX; i += nth) {
    const double x0 = p.dx[i], f0 = p.fs[i];
    (sm + S::x0)[i] = x0;
    sdm[i] = x0 - f0;
    xg += fabs(x0); gap += fabs(f0); kkt = nmax(kkt, fabs(f0));
  }
  __syncthreads();
  for (int i = tid; i < NX; i += nth) {
    double acc = 0.0;
    for (int j = 0; j < NX; ++j) acc += p.Vxx[i + NX * j] * sdm[j];
    const double v = acc + p.Vx[i];
    (sm + S::lam0)[i] = v;
    p.lag[i] = v;
  }
  __syncthreads();
  const int grp = tid / SPLIT, q = tid % SPLIT, ngrp = nth / SPLIT;
  const int h = lane >> 4, l16 = lane & 15, w16 = (tid >> 5) * 16, n16 = (nth >> 5) * 16;
  for (int t = 0; t < T; ++t) {
    const double* st = as_shared(pipe.acquire());
    const double *sFx = st, *sFu = st + NX * NX, *sK = st + NX * NX + NX * NU, *sV = sK + NX * NU;
    const double* x = sm + ((t & 1) ? S::x1 : S::x0);
    double* xn = sm + ((t & 1) ? S::x0 : S::x1);
    const double* lam = sm + ((t & 1) ? S::lam1 : S::lam0);
    double* lamn = sm + ((t & 1) ? S::lam0 : S::lam1);
    const double* v = sm + ((t & 1) ? S::vec1 : S::vec0);
    /* F1: du = -K dx - k  (src/sqp.cpp:285-286) */
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
        double u = -acc;
        u -= v[oK + a];
        su[a] = u;
        p.du[(size_t)t * NU + a] = u;
        ug += fabs(u);
      }
    }
    __syncthreads(); /* 1: du visible; knot t-1 completely consumed */
    if (t + pipe.ns - 1 < T) pipe.issue(p, t + pipe.ns - 1);
    if (t + 1 < T) {
      double* vn = sm + ((t & 1) ? S::vec0 : S::vec1);
#pragma unroll
      for (int j = 0; j < VPT; ++j) {
        const int e = tid + j * nth;
        if (vs[j].base) vn[e] = vr[j];
      }
      if (t + 2 < T) {
#pragma unroll
        for (int j = 0; j < VPT; ++j)
          if (vs[j].base) vr[j] = vs[j].base[(size_t)(t + 2) * vs[j].stride];
      }
    }
    /* F2: dx+ = fs+ + Fu du + Fx dx  (src/sqp.cpp:287-289); two half-warps share a row */
    for (int o0 = 0; o0 < NX; o0 += n16) {
      const int i = o0 + w16 + l16;
      const int ic = i < NX ? i : 0;
      double a1 = 0.0, a2 = 0.0;
#pragma unroll
      for (int a = 0; a < NU / 2; ++a) a2 = fma(sFu[ic + NX * (2 * a + h)], su[2 * a + h], a2);
#pragma unroll
      for (int j = 0; j < NX / 2; ++j) a1 = fma(sFx[ic + NX * (2 * j + h)], x[2 * j + h], a1);
      a1 += __shfl_xor_sync(0xffffffffu, a1, 16);
      a2 += __shfl_xor_sync(0xffffffffu, a2, 16);
      if (i < NX && h == 0) {
        const double fsn = v[oFs + i];
        double xv = fsn;
        xv += a2;
        xv += a1;
        xn[i] = xv;
        sdm[i] = xv - fsn;
        p.dx[(size_t)(t + 1) * NX + i] = xv;
        xg += fabs(xv); gap += fabs(fsn); kkt = nmax(kkt, fabs(fsn));
      }
    }
    __syncthreads(); /* 2: dx+ - fs+ visible */
    /* F3: lag_mul+ = Vxx+ (dx+ - fs+) + Vx+  (src/sqp.cpp:282-284, 296-298) */
    for (int o0 = 0; o0 < NX; o0 += n16) {
      const int i = o0 + w16 + l16;
      const int ic = i < NX ? i : 0;
      double a1 = 0.0;
#pragma unroll
      for (int j = 0; j < NX / 2; ++j) a1 = fma(sV[ic + NX * (2 * j + h)], sdm[2 * j + h], a1);
      a1 += __shfl_xor_sync(0xffffffffu, a1, 16);
      if (i < NX && h == 0) {
        const double lv = a1 + v[oVx + i];
        lamn[i] = lv;
        p.lag[(size_t)(t + 1) * NX + i] = lv;
      }
    }
    __syncthreads(); /* 3: lag_mul+ visible */
    /* F4: stationarity rows of knot t  (src/sqp.cpp:249-258) */
    for (int o0 = 0; o0 < N; o0 += ngrp) {
      const int o = o0 + grp;
      const bool valid = o < N;
      const double* Fc = sFx + NX * (valid ? o : 0); /* [Fx Fu] are adjacent in the stage */
      double acc = 0.0;
#pragma unroll
      for (int k = 0; k < (NX + SPLIT - 1) / SPLIT; ++k) {
        const int kk = k * SPLIT + q;
        if (kk < NX) acc = fma(Fc[kk], lamn[kk], acc);
      }
      acc = split_reduce<SPLIT>(acc);
      if (valid && q == 0) {
        double r = v[oLx + o] + acc; /* [Lx; Lu] are adjacent in the pack */
        if (o < NX) r -= lam[o];
        kkt = nmax(kkt, fabs(r));
      }
    }
  }
  __syncthreads();
  {
    const double* lam = sm + ((T & 1) ? S::lam1 : S::lam0);
    for (int i = tid; i < NX; i += nth) kkt = nmax(kkt, fabs(p.LxT[i] - lam[i])); /* src/sqp.cpp:259-262 */
  }
  kkt = cta_max(kkt, red);
  xg = cta_sum(xg, red); ug = cta_sum(ug, red); gap = cta_sum(gap, red);
  if (tid == 0) {
    const size_t B = p.B, b = p.b;
    p.scal[MIMGPU_S_KKT * B + b] = kkt;
    p.scal[MIMGPU_S_GAP_NORM * B + b] = gap;
    p.scal[MIMGPU_S_MERIT * B + b] = p.cost + mu_dynamic * gap;
    p.scal[MIMGPU_S_XGRAD_NORM * B + b] = xg / (double)(T + 1);
    p.scal[MIMGPU_S_UGRAD_NORM * B + b] = ug / (double)T;
  }
}

}  // namespace mimgpu
