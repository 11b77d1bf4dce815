/*
 * mimgpu.cu -- C ABI (include/mimgpu.h) of the B200-native batched OCP-QP engine.
 *
 * Thin host layer: owns the device buffers of one shape bucket, copies the
 * derivative tensors in, launches the direction kernels on the caller's stream and
 * copies solver members out.  No CPU fallback: without a CUDA device every entry
 * point returns MIMGPU_ERR_NODEVICE / MIMGPU_ERR_CUDA.
 */
#include <cuda_runtime.h>

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>
#include <string>
#include <utility>
#include <vector>

#include "launchers.h"

using namespace mimgpu;

/* ------------------------------------------------------------------ errors */
static thread_local std::string g_err;
static int fail(int code, const std::string& msg) {
  g_err = msg;
  return code;
}
namespace mimgpu {
/* for the entry points that live in other translation units (noalloc.cu) */
void set_last_error(const std::string& msg) { g_err = msg; }
}  // namespace mimgpu
#define CU(call)                                                                                   \
  do {                                                                                             \
    cudaError_t e_ = (call);                                                                       \
    if (e_ != cudaSuccess)                                                                         \
      return fail(MIMGPU_ERR_CUDA, std::string(#call) + ": " + cudaGetErrorString(e_));            \
  } while (0)

/* --------------------------------------------------------------------- ctx */
struct mimgpu_ctx {
  int device = 0;
  mimgpu_shape sh{};
  KArgs k{};                       /* device pointers handed to the kernels */
  double* in_buf = nullptr;        /* owned copy of the derivatives (MEM_HOST uploads) */
  size_t in_doubles = 0;
  bool have_inputs = false;
  double* state_buf = nullptr;     /* all solver members, one allocation */
  size_t state_doubles = 0;
  double* field[MIMGPU_F_COUNT] = {};
  size_t field_n[MIMGPU_F_COUNT] = {};
  int32_t* ibuf = nullptr;         /* qp_iters, status, converged, active */
  double* reg_buf = nullptr;       /* preg[B], dreg[B] */
  void* summary_dev = nullptr;
  int64_t launches = 0;
  int nt = 32;                     /* CTA size of the generic kernels */
  int smem_bytes = 0;
  int sm_count = 148;
  int use_small = 0;               /* specialised small-shape kernels: bit 0 full-spill, bit 1 fast */
  bool extras_stale = false;       /* Qu / Quuk / LLT not written by the last (fast) SQP direction */
  bool spill_packed = false;       /* Vxx / Vx / K / k of the last direction live in the spill records */
  double* pack_buf = nullptr;      /* knot-packed copy of the inputs (fast small-shape kernel) */
  double* spill_buf = nullptr;     /* knot-packed Vxx Vx K k */
  size_t pack_pp = 0, spill_pp = 0; /* doubles per problem */
  cudaStream_t pipe_st[2] = {nullptr, nullptr}; /* internal streams of the chunked host path */
  SqpKPar last_sqp{};              /* its parameters, for the on-demand recomputation */
  CtaPlan cplan{};                 /* DMMA CTA-per-problem SQP / FDDP kernel (kernels_cta.cuh) */
  StreamPlan splan{};              /* streaming CSQP kernel (kernels_stream.cuh), ok = 0: not available */
  RecPlan rplan{};                 /* record-based CSQP kernel of the small shapes (kernels_rec.cuh) */
  double* rec_scr = nullptr;       /* its knot records [B][T][record]: allocated at the first CSQP call that uses it */
  bool csqp_rec = true;            /* MIMGPU_CSQP_REC=0: streaming kernel instead */
  AdmmPlan aplan{};                /* condensed-chain CSQP kernel (kernels_admm.cuh): the default CSQP path */
  double* admm_scr[3] = {nullptr, nullptr, nullptr}; /* Acl, r, tmp of the condensed ADMM iterations: allocated at the first CSQP call */
  /* debug / A-B switches, read ONCE at mimgpu_create (never on the hot path) */
  bool force_generic = false;      /* MIMGPU_FORCE_GENERIC: shape-generic kernels only */
  bool csqp_stream = false;        /* MIMGPU_CSQP_STREAM: sweep-by-sweep streaming CSQP kernel for every shape */
  bool csqp_admm = false;          /* MIMGPU_CSQP_ADMM: condensed-chain CSQP kernel for every shape it is compiled for */
  bool small_alt = false;          /* MIMGPU_SMALL_FULL / MIMGPU_SMALL_NOMMA: full-spill small kernel as default */
  bool warned_generic = false;
  long long* admm_timing = nullptr; /* MIMGPU_ADMM_TIMING: device counters of the per-phase debug timer */
  int last_kind = 0;               /* 0 none, 1 SQP, 2 FDDP, 3 CSQP: what the last direction call was */
  bool pack_dirty = false;         /* field-major inputs changed since the knot records were built */
  bool attached = false;           /* inputs are caller-owned device memory: re-read at EVERY direction call */
  bool records_mode = false;       /* inputs arrived as knot records (mimgpu_upload_records) */
  bool fields_valid = true;        /* records mode: the field-major arrays have been rebuilt from the records */
  double* cost_buf = nullptr;      /* [B] cost_ of host-side record uploads */
  cudaStream_t last_stream = nullptr; /* stream of the last asynchronous call (ordering of the chunked host path) */
  /* device-side solve() (mimgpu_model_attach / mimgpu_sqp_solve) */
  PolyModelDev model{};
  double* model_buf = nullptr;     /* owned copy of a host-side model */
  bool have_model = false;
  SolveDev sv{};
  void* solve_buf = nullptr;       /* iter, done, flag, two lists, counters, filter memory */
  int solve_filter_size = 0;
};

static size_t deriv_sizes(const mimgpu_shape& s, size_t n[20]) {
  const size_t B = s.B, T = s.T, nx = s.nx, nu = s.nu, ng = s.ng, ngT = s.ngT;
  size_t i = 0, tot = 0;
  n[i++] = B * T * nx * nx;  /* Fx */
  n[i++] = B * T * nx * nu;  /* Fu */
  n[i++] = B * T * nx * nx;  /* Lxx */
  n[i++] = B * T * nx * nu;  /* Lxu */
  n[i++] = B * T * nu * nu;  /* Luu */
  n[i++] = B * T * nx;       /* Lx */
  n[i++] = B * T * nu;       /* Lu */
  n[i++] = B * (T + 1) * nx; /* fs */
  n[i++] = B * nx * nx;      /* LxxT */
  n[i++] = B * nx;           /* LxT */
  n[i++] = B * T * ng * nx;  /* Gx */
  n[i++] = B * T * ng * nu;  /* Gu */
  n[i++] = B * T * ng;       /* g */
  n[i++] = B * T * ng;       /* g_lb */
  n[i++] = B * T * ng;       /* g_ub */
  n[i++] = B * ngT * nx;     /* GxT */
  n[i++] = B * ngT;          /* gT */
  n[i++] = B * ngT;          /* g_lbT */
  n[i++] = B * ngT;          /* g_ubT */
  n[i++] = B;                /* cost */
  for (size_t j = 0; j < 20; ++j) tot += (n[j] + 1) & ~(size_t)1;
  return tot;
}

static const double** deriv_slots(KArgs& k, const double** out[20]) {
  const double** p[20] = {&k.Fx, &k.Fu, &k.Lxx, &k.Lxu, &k.Luu, &k.Lx, &k.Lu, &k.fs, &k.LxxT, &k.LxT,
                          &k.Gx, &k.Gu, &k.g, &k.glb, &k.gub, &k.GxT, &k.gT, &k.glbT, &k.gubT, &k.cost};
  for (int i = 0; i < 20; ++i) out[i] = p[i];
  return nullptr;
}

static void field_sizes(const mimgpu_shape& s, size_t n[MIMGPU_F_COUNT]) {
  const size_t B = s.B, T = s.T, nx = s.nx, nu = s.nu, nc = T * s.ng + s.ngT;
  n[MIMGPU_F_DX] = B * (T + 1) * nx;
  n[MIMGPU_F_DU] = B * T * nu;
  n[MIMGPU_F_LAG_MUL] = B * (T + 1) * nx;
  n[MIMGPU_F_DX_TILDE] = B * (T + 1) * nx;
  n[MIMGPU_F_DU_TILDE] = B * T * nu;
  n[MIMGPU_F_Y] = B * nc;
  n[MIMGPU_F_Z] = B * nc;
  n[MIMGPU_F_RHO_VEC] = B * nc;
  n[MIMGPU_F_VXX] = B * (T + 1) * nx * nx;
  n[MIMGPU_F_VX] = B * (T + 1) * nx;
  n[MIMGPU_F_K] = B * T * nu * nx;
  n[MIMGPU_F_KFF] = B * T * nu;
  n[MIMGPU_F_QU] = B * T * nu;
  n[MIMGPU_F_QUUK] = B * T * nu;
  n[MIMGPU_F_QUU_LLT] = B * T * nu * nu;
  n[MIMGPU_F_XS] = B * (T + 1) * nx;
  n[MIMGPU_F_US] = B * T * nu;
}

/* ------------------------------------------------------------------ kernels */
__global__ void __launch_bounds__(512) k_sqp_direction_generic(KArgs a, SqpKPar kp) {
  extern __shared__ double sm[];
  if ((int)blockIdx.x >= a.n_active) return;
  const int b = a.active ? a.active[blockIdx.x] : (int)blockIdx.x + a.first;
  sqp_direction_body(a, b, sm, kp);
}
__global__ void __launch_bounds__(512) k_fddp_direction_generic(KArgs a, SqpKPar kp) {
  extern __shared__ double sm[];
  if ((int)blockIdx.x >= a.n_active) return;
  const int b = a.active ? a.active[blockIdx.x] : (int)blockIdx.x + a.first;
  fddp_direction_body(a, b, sm, kp);
}
__global__ void __launch_bounds__(512) k_csqp_direction_generic(KArgs a, CsqpKPar kp) {
  extern __shared__ double sm[];
  if ((int)blockIdx.x >= a.n_active) return;
  const int b = a.active ? a.active[blockIdx.x] : (int)blockIdx.x + a.first;
  csqp_direction_body(a, b, sm, kp);
}
__global__ void k_csqp_reset_rho(KArgs a, double base, double rho_min) {
  if ((int)blockIdx.x >= a.n_active) return;
  const int b = a.active ? a.active[blockIdx.x] : (int)blockIdx.x + a.first;
  csqp_reset_rho_body(a, b, base, rho_min);
}
__global__ void k_csqp_calc_norms(KArgs a) {
  __shared__ double red[64];
  if ((int)blockIdx.x >= a.n_active) return;
  const int b = a.active ? a.active[blockIdx.x] : (int)blockIdx.x + a.first;
  csqp_calc_norms_body(a, b, red);
}
__global__ void k_fill(double* p, size_t n, double v) {
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) p[i] = v;
}
struct SummaryDev {
  double max_kkt;
  long long sum_it, max_it, n_conv, n_fail;
};
__global__ void k_summary(const double* kkt, const int32_t* it, const int32_t* st, const int32_t* conv, int B,
                          double tol, int csqp, SummaryDev* out) {
  __shared__ double smax[32];
  __shared__ long long ssum[32], smx[32], scv[32], sfl[32];
  double m = 0.0;
  long long s = 0, mx = 0, cv = 0, fl = 0;
  for (int b = threadIdx.x; b < B; b += blockDim.x) {
    m = nmax(m, kkt[b]);
    s += it[b];
    mx = it[b] > mx ? it[b] : mx;
    cv += csqp ? (conv[b] != 0) : (kkt[b] <= tol);
    fl += (st[b] != 0);
  }
  for (int o = 16; o; o >>= 1) {
    m = nmax(m, __shfl_xor_sync(0xffffffffu, m, o));
    s += __shfl_xor_sync(0xffffffffu, s, o);
    long long t = __shfl_xor_sync(0xffffffffu, mx, o);
    mx = t > mx ? t : mx;
    cv += __shfl_xor_sync(0xffffffffu, cv, o);
    fl += __shfl_xor_sync(0xffffffffu, fl, o);
  }
  const int w = threadIdx.x >> 5;
  if ((threadIdx.x & 31) == 0) { smax[w] = m; ssum[w] = s; smx[w] = mx; scv[w] = cv; sfl[w] = fl; }
  __syncthreads();
  if (threadIdx.x == 0) {
    const int nw = blockDim.x >> 5;
    for (int i = 1; i < nw; ++i) {
      m = nmax(m, smax[i]); s += ssum[i]; mx = smx[i] > mx ? smx[i] : mx; cv += scv[i]; fl += sfl[i];
    }
    out->max_kkt = m; out->sum_it = s; out->max_it = mx; out->n_conv = cv; out->n_fail = fl;
  }
}

/* ---------------------------------------------------------------- lifecycle */
extern "C" {

int mimgpu_abi_version(void) { return MIMGPU_ABI_VERSION; }

const char* mimgpu_last_error(void) { return g_err.c_str(); }

int mimgpu_device_count(int* n) {
  if (!n) return fail(MIMGPU_ERR_INVALID, "null pointer");
  int c = 0;
  cudaError_t e = cudaGetDeviceCount(&c);
  if (e != cudaSuccess) {
    *n = 0;
    cudaGetLastError();
    return fail(MIMGPU_ERR_NODEVICE, std::string("cudaGetDeviceCount: ") + cudaGetErrorString(e));
  }
  *n = c;
  return MIMGPU_OK;
}

static int configure_launch(mimgpu_ctx* c) {
  const mimgpu_shape& s = c->sh;
  cudaDeviceProp prop;
  CU(cudaGetDeviceProperties(&prop, c->device));
  c->sm_count = prop.multiProcessorCount;
  const int max_smem = (int)prop.sharedMemPerBlockOptin;
  int stage = 1;
  Smem L = smem_layout(s.nx, s.nu, s.ng, s.ngT, stage);
  if (L.total * 8 > max_smem) {
    stage = 0;
    L = smem_layout(s.nx, s.nu, s.ng, s.ngT, stage);
  }
  if (L.total * 8 > max_smem)
    return fail(MIMGPU_ERR_UNSUPPORTED, "shape needs " + std::to_string(L.total * 8) + " B of shared memory per CTA");
  c->k.stage_fx = stage;
  c->smem_bytes = L.total * 8;
  int nt = ((s.nx * s.nx / 8 + 31) / 32) * 32;
  if (nt < 32) nt = 32;
  if (nt > 512) nt = 512;
  c->nt = nt;
  CU(cudaFuncSetAttribute(k_sqp_direction_generic, cudaFuncAttributeMaxDynamicSharedMemorySize, max_smem));
  CU(cudaFuncSetAttribute(k_fddp_direction_generic, cudaFuncAttributeMaxDynamicSharedMemorySize, max_smem));
  CU(cudaFuncSetAttribute(k_csqp_direction_generic, cudaFuncAttributeMaxDynamicSharedMemorySize, max_smem));
  c->use_small = small_kernels_configure_x(s);
  c->splan = stream_plan_x(s, stage, max_smem);
  c->aplan = admm_plan_x(s, stage, max_smem);
  c->rplan = rec_plan_x(s, stage, max_smem);
  { const char* e = getenv("MIMGPU_CSQP_REC"); c->csqp_rec = e ? atoi(e) != 0 : true; }
  c->cplan = cta_plan_x(s, stage, max_smem);
  c->force_generic = getenv("MIMGPU_FORCE_GENERIC") != nullptr;
  c->csqp_stream = getenv("MIMGPU_CSQP_STREAM") != nullptr;
  c->csqp_admm = getenv("MIMGPU_CSQP_ADMM") != nullptr;
  c->small_alt = getenv("MIMGPU_SMALL_FULL") != nullptr || getenv("MIMGPU_SMALL_NOMMA") != nullptr;
  if (getenv("MIMGPU_ADMM_TIMING") != nullptr) CU(cudaMalloc(&c->admm_timing, 8 * sizeof(long long)));
  return MIMGPU_OK;
}

int mimgpu_create(mimgpu_ctx** out, int device, const mimgpu_shape* shape) {
  if (!out || !shape) return fail(MIMGPU_ERR_INVALID, "null pointer");
  *out = nullptr;
  const mimgpu_shape& s = *shape;
  if (s.B < 1 || s.T < 1 || s.nx < 1 || s.nu < 0 || s.ng < 0 || s.ngT < 0)
    return fail(MIMGPU_ERR_INVALID, "bad shape");
  int ndev = 0;
  int rc = mimgpu_device_count(&ndev);
  if (rc != MIMGPU_OK) return rc;
  if (ndev == 0) return fail(MIMGPU_ERR_NODEVICE, "no CUDA device: there is no CPU fallback");
  if (device < 0 || device >= ndev) return fail(MIMGPU_ERR_INVALID, "bad device index");
  CU(cudaSetDevice(device));
  mimgpu_ctx* c = new (std::nothrow) mimgpu_ctx();
  if (!c) return fail(MIMGPU_ERR_INVALID, "out of host memory");
  c->device = device;
  c->sh = s;
  c->k.sh = s;
  rc = configure_launch(c);
  if (rc != MIMGPU_OK) { delete c; return rc; }
  /* solver members: zero-initialised as in src/ddp.cpp:447-502, src/csqp.cpp:24-168 */
  field_sizes(s, c->field_n);
  size_t tot = 0;
  for (int f = 0; f < MIMGPU_F_COUNT; ++f) tot += (c->field_n[f] + 1) & ~(size_t)1;
  const size_t n_vfs = ((size_t)s.B * s.T * s.nx + 1) & ~(size_t)1;
  const size_t n_scal = ((size_t)MIMGPU_S_COUNT * s.B + 1) & ~(size_t)1;
  tot += n_vfs + n_scal;
  c->state_doubles = tot;
  cudaError_t e = cudaMalloc(&c->state_buf, tot * sizeof(double));
  if (e != cudaSuccess) { delete c; return fail(MIMGPU_ERR_CUDA, std::string("cudaMalloc(state): ") + cudaGetErrorString(e)); }
  cudaMemset(c->state_buf, 0, tot * sizeof(double));
  double* p = c->state_buf;
  for (int f = 0; f < MIMGPU_F_COUNT; ++f) { c->field[f] = p; p += (c->field_n[f] + 1) & ~(size_t)1; }
  c->k.Vxx_fs = p; p += n_vfs;
  c->k.scal = p;
  c->k.dx = c->field[MIMGPU_F_DX]; c->k.du = c->field[MIMGPU_F_DU]; c->k.lag = c->field[MIMGPU_F_LAG_MUL];
  c->k.dxt = c->field[MIMGPU_F_DX_TILDE]; c->k.dut = c->field[MIMGPU_F_DU_TILDE];
  c->k.y = c->field[MIMGPU_F_Y]; c->k.z = c->field[MIMGPU_F_Z]; c->k.rho = c->field[MIMGPU_F_RHO_VEC];
  c->k.Vxx = c->field[MIMGPU_F_VXX]; c->k.Vx = c->field[MIMGPU_F_VX]; c->k.K = c->field[MIMGPU_F_K];
  c->k.kff = c->field[MIMGPU_F_KFF]; c->k.Qu = c->field[MIMGPU_F_QU]; c->k.Quuk = c->field[MIMGPU_F_QUUK];
  c->k.L = c->field[MIMGPU_F_QUU_LLT];
  /* KKT_ = +inf (include/mim_solvers/sqp.hpp:182-183), rho_sparse_ = 1e-1 (csqp.hpp:323) */
  k_fill<<<(s.B + 255) / 256, 256>>>(c->k.scal + (size_t)MIMGPU_S_KKT * s.B, s.B, INFINITY);
  k_fill<<<(s.B + 255) / 256, 256>>>(c->k.scal + (size_t)MIMGPU_S_RHO_SPARSE * s.B, s.B, 1e-1);
  e = cudaMalloc(&c->ibuf, sizeof(int32_t) * 4 * (size_t)s.B);
  if (e == cudaSuccess) e = cudaMemset(c->ibuf, 0, sizeof(int32_t) * 4 * (size_t)s.B);
  if (e == cudaSuccess) e = cudaMalloc(&c->reg_buf, sizeof(double) * 2 * (size_t)s.B);
  if (e == cudaSuccess) e = cudaMalloc(&c->summary_dev, sizeof(SummaryDev));
  if (e == cudaSuccess) e = cudaMalloc(&c->cost_buf, sizeof(double) * (size_t)s.B);
  if (e == cudaSuccess && (c->use_small & 2) &&
      small_pack_layout_x(s, &c->pack_pp, &c->spill_pp) == 0) {
    e = cudaMalloc(&c->pack_buf, c->pack_pp * (size_t)s.B * sizeof(double));
    if (e == cudaSuccess) e = cudaMalloc(&c->spill_buf, c->spill_pp * (size_t)s.B * sizeof(double));
    if (e == cudaSuccess) e = cudaMemset(c->spill_buf, 0, c->spill_pp * (size_t)s.B * sizeof(double));
  }
  if (e == cudaSuccess) e = cudaDeviceSynchronize();
  if (e != cudaSuccess) {
    mimgpu_destroy(c);
    return fail(MIMGPU_ERR_CUDA, std::string("cudaMalloc(aux): ") + cudaGetErrorString(e));
  }
  c->k.qp_iters = c->ibuf; c->k.status = c->ibuf + s.B; c->k.converged = c->ibuf + 2 * (size_t)s.B;
  c->k.active = nullptr; c->k.n_active = s.B;
  c->k.preg_arr = nullptr; c->k.dreg_arr = nullptr;
  c->k.pack = nullptr; c->k.spill = c->spill_buf; c->k.first = 0; c->k.pad_ = 0;
  *out = c;
  return MIMGPU_OK;
}

void mimgpu_destroy(mimgpu_ctx* c) {
  if (!c) return;
  cudaSetDevice(c->device);
  if (c->pipe_st[0]) cudaStreamDestroy(c->pipe_st[0]);
  if (c->pipe_st[1]) cudaStreamDestroy(c->pipe_st[1]);
  for (int i = 0; i < 3; ++i) cudaFree(c->admm_scr[i]);
  cudaFree(c->rec_scr);
  cudaFree(c->cost_buf); cudaFree(c->admm_timing); cudaFree(c->model_buf); cudaFree(c->solve_buf);
  cudaFree(c->pack_buf); cudaFree(c->spill_buf); cudaFree(c->in_buf); cudaFree(c->state_buf); cudaFree(c->ibuf); cudaFree(c->reg_buf); cudaFree(c->summary_dev);
  delete c;
}

int mimgpu_get_shape(const mimgpu_ctx* c, mimgpu_shape* shape) {
  if (!c || !shape) return fail(MIMGPU_ERR_INVALID, "null pointer");
  *shape = c->sh;
  return MIMGPU_OK;
}

/* ------------------------------------------------------------------ data in */
/* Shapes served by the fast warp-per-problem SQP kernel read KNOT RECORDS (one record per knot,
 * kernels_small_gj.cuh).  A caller can hand the records in directly (mimgpu_upload_records: no
 * re-layout at all); field-major inputs are re-laid-out by one streaming pass (k_pack_knots, read +
 * write of the inputs at HBM speed) as part of the direction call that consumes them. */
static int pack_inputs(mimgpu_ctx* c, cudaStream_t st) {
  if (small_pack_launch_x(c->k, c->pack_buf, 0, c->sh.B, st) != 0) return fail(MIMGPU_ERR_CUDA, "k_pack_knots launch failed");
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return fail(MIMGPU_ERR_CUDA, std::string("k_pack_knots: ") + cudaGetErrorString(e));
  c->launches++;
  c->k.pack = c->pack_buf;
  c->pack_dirty = false;
  return MIMGPU_OK;
}

static int ensure_in_buf(mimgpu_ctx* c, size_t n[20]) {
  const size_t tot = deriv_sizes(c->sh, n);
  if (!c->in_buf) {
    CU(cudaMalloc(&c->in_buf, tot * sizeof(double)));
    c->in_doubles = tot;
  }
  return MIMGPU_OK;
}

/* Records mode: rebuild the field-major arrays (into the ctx's own buffer) for the paths the record
 * kernel does not serve.  Caller-owned records are re-read every time (they may have changed). */
static int ensure_fields(mimgpu_ctx* c, cudaStream_t st) {
  if (!c->records_mode || (c->fields_valid && !c->attached)) return MIMGPU_OK;
  size_t n[20];
  int rc = ensure_in_buf(c, n);
  if (rc) return rc;
  const double** slot[20];
  deriv_slots(c->k, slot);
  double* q = c->in_buf;
  for (int i = 0; i < 19; ++i) { /* slot 19 (cost) keeps the pointer given with the records */
    *slot[i] = n[i] ? q : nullptr;
    q += (n[i] + 1) & ~(size_t)1;
  }
  if (small_unpack_inputs_launch_x(c->k, c->k.pack, st) != 0) return fail(MIMGPU_ERR_CUDA, "k_unpack_records launch failed");
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return fail(MIMGPU_ERR_CUDA, std::string("k_unpack_records: ") + cudaGetErrorString(e));
  c->launches++;
  c->fields_valid = true;
  return MIMGPU_OK;
}

static void inputs_changed(mimgpu_ctx* c) {
  c->have_inputs = true;
  c->extras_stale = false; /* lazily materialised members of an earlier direction belong to the old inputs */
  c->spill_packed = false;
}

int mimgpu_upload(mimgpu_ctx* c, const mimgpu_derivs* d, int memkind, void* stream) {
  if (!c || !d) return fail(MIMGPU_ERR_INVALID, "null pointer");
  cudaStream_t st = (cudaStream_t)stream;
  CU(cudaSetDevice(c->device));
  size_t n[20];
  deriv_sizes(c->sh, n);
  const double* src[20] = {d->Fx, d->Fu, d->Lxx, d->Lxu, d->Luu, d->Lx, d->Lu, d->fs, d->LxxT, d->LxT,
                           d->Gx, d->Gu, d->g, d->g_lb, d->g_ub, d->GxT, d->gT, d->g_lbT, d->g_ubT, d->cost};
  for (int i = 0; i < 19; ++i)
    if (n[i] && !src[i]) return fail(MIMGPU_ERR_INVALID, "missing derivative array #" + std::to_string(i));
  const double** slot[20];
  deriv_slots(c->k, slot);
  c->last_stream = st;
  if (memkind == MIMGPU_MEM_DEVICE) {
    for (int i = 0; i < 20; ++i) *slot[i] = n[i] ? src[i] : nullptr;
    c->attached = true;
  } else if (memkind == MIMGPU_MEM_HOST) {
    int rc = ensure_in_buf(c, n);
    if (rc) return rc;
    double* p = c->in_buf;
    for (int i = 0; i < 20; ++i) {
      if (n[i] && src[i]) {
        CU(cudaMemcpyAsync(p, src[i], n[i] * sizeof(double), cudaMemcpyHostToDevice, st));
        *slot[i] = p;
      } else {
        *slot[i] = nullptr;
      }
      p += (n[i] + 1) & ~(size_t)1;
    }
    c->attached = false;
  } else {
    return fail(MIMGPU_ERR_INVALID, "bad memkind");
  }
  c->records_mode = false;
  c->fields_valid = true;
  c->k.pack = nullptr;
  c->pack_dirty = true;
  inputs_changed(c);
  return MIMGPU_OK;
}

int mimgpu_get_record_layout(const mimgpu_ctx* c, mimgpu_record_layout* out) {
  if (!c || !out) return fail(MIMGPU_ERR_INVALID, "null pointer");
  int v[16];
  if (!(c->use_small & 2) || small_record_offsets_x(c->sh, v) != 0)
    return fail(MIMGPU_ERR_UNSUPPORTED, "shape has no knot-record kernel");
  out->rec_doubles = v[0]; out->term_doubles = v[1]; out->problem_doubles = v[2];
  out->off_fs = v[3]; out->off_Lxx = v[4]; out->off_Lxu = v[5]; out->off_Luu = v[6]; out->off_Lx = v[7];
  out->off_Lu = v[8]; out->off_Fx = v[9]; out->off_Fu = v[10];
  out->toff_fs = v[11]; out->toff_Lxx = v[12]; out->toff_Lx = v[13];
  out->sym_packed = v[14];
  out->reserved = 0;
  return MIMGPU_OK;
}

int mimgpu_upload_records(mimgpu_ctx* c, const double* records, const double* cost, int memkind, void* stream) {
  if (!c || !records) return fail(MIMGPU_ERR_INVALID, "null pointer");
  if (!(c->use_small & 2) || !c->pack_buf) return fail(MIMGPU_ERR_UNSUPPORTED, "shape has no knot-record kernel");
  if (c->sh.ng || c->sh.ngT) return fail(MIMGPU_ERR_UNSUPPORTED, "knot records carry no constraint blocks (ng = ngT = 0 only)");
  cudaStream_t st = (cudaStream_t)stream;
  CU(cudaSetDevice(c->device));
  c->last_stream = st;
  if (memkind == MIMGPU_MEM_DEVICE) {
    if ((uintptr_t)records & 15) return fail(MIMGPU_ERR_INVALID, "records must be 16-byte aligned");
    c->k.pack = records;
    c->k.cost = cost;
    c->attached = true;
  } else if (memkind == MIMGPU_MEM_HOST) {
    CU(cudaMemcpyAsync(c->pack_buf, records, c->pack_pp * (size_t)c->sh.B * sizeof(double), cudaMemcpyHostToDevice, st));
    if (cost) CU(cudaMemcpyAsync(c->cost_buf, cost, (size_t)c->sh.B * sizeof(double), cudaMemcpyHostToDevice, st));
    c->k.pack = c->pack_buf;
    c->k.cost = cost ? c->cost_buf : nullptr;
    c->attached = false;
  } else {
    return fail(MIMGPU_ERR_INVALID, "bad memkind");
  }
  c->records_mode = true;
  c->fields_valid = false;
  c->pack_dirty = false;
  inputs_changed(c);
  return MIMGPU_OK;
}

int mimgpu_set_reg(mimgpu_ctx* c, const double* preg, const double* dreg, void* stream) {
  if (!c) return fail(MIMGPU_ERR_INVALID, "null pointer");
  cudaStream_t st = (cudaStream_t)stream;
  CU(cudaSetDevice(c->device));
  const size_t B = c->sh.B;
  if (preg) {
    CU(cudaMemcpyAsync(c->reg_buf, preg, B * sizeof(double), cudaMemcpyHostToDevice, st));
    c->k.preg_arr = c->reg_buf;
  } else {
    c->k.preg_arr = nullptr;
  }
  if (dreg) {
    CU(cudaMemcpyAsync(c->reg_buf + B, dreg, B * sizeof(double), cudaMemcpyHostToDevice, st));
    c->k.dreg_arr = c->reg_buf + B;
  } else {
    c->k.dreg_arr = nullptr;
  }
  return MIMGPU_OK;
}

int mimgpu_set_active(mimgpu_ctx* c, const int32_t* idx, int32_t n, void* stream) {
  if (!c) return fail(MIMGPU_ERR_INVALID, "null pointer");
  cudaStream_t st = (cudaStream_t)stream;
  CU(cudaSetDevice(c->device));
  if (!idx || n < 0) {
    c->k.active = nullptr;
    c->k.n_active = c->sh.B;
    return MIMGPU_OK;
  }
  if (n > c->sh.B) return fail(MIMGPU_ERR_INVALID, "more active problems than B");
  for (int i = 0; i < n; ++i)
    if (idx[i] < 0 || idx[i] >= c->sh.B) return fail(MIMGPU_ERR_INVALID, "active index out of range");
  int32_t* dst = c->ibuf + 3 * (size_t)c->sh.B;
  if (n) CU(cudaMemcpyAsync(dst, idx, n * sizeof(int32_t), cudaMemcpyHostToDevice, st));
  c->k.active = dst;
  c->k.n_active = n;
  return MIMGPU_OK;
}

int mimgpu_set_field(mimgpu_ctx* c, int field, const double* src, void* stream) {
  if (!c || !src) return fail(MIMGPU_ERR_INVALID, "null pointer");
  if (field < 0 || field >= MIMGPU_F_COUNT) return fail(MIMGPU_ERR_INVALID, "bad field id");
  CU(cudaSetDevice(c->device));
  if (c->field_n[field])
    CU(cudaMemcpyAsync(c->field[field], src, c->field_n[field] * sizeof(double), cudaMemcpyHostToDevice,
                       (cudaStream_t)stream));
  /* a member written by the caller is no longer "to be materialised from the last fast SQP launch" */
  if (field == MIMGPU_F_VXX || field == MIMGPU_F_VX || field == MIMGPU_F_K || field == MIMGPU_F_KFF) c->spill_packed = false;
  if (field == MIMGPU_F_QU || field == MIMGPU_F_QUUK || field == MIMGPU_F_QUU_LLT) c->extras_stale = false;
  c->last_stream = (cudaStream_t)stream;
  return MIMGPU_OK;
}

/* ------------------------------------------------------------- the hot path */
static int pre_launch(mimgpu_ctx* c) {
  if (!c) return fail(MIMGPU_ERR_INVALID, "null ctx");
  if (!c->have_inputs) return fail(MIMGPU_ERR_STATE, "direction requested before mimgpu_upload");
  cudaError_t e = cudaSetDevice(c->device);
  if (e != cudaSuccess) return fail(MIMGPU_ERR_CUDA, cudaGetErrorString(e));
  return MIMGPU_OK;
}
/* Shapes outside the template lists run the shape-generic CTA kernels (correct, but far from the
 * specialised ones): say so once per ctx instead of silently. */
static void warn_generic(mimgpu_ctx* c, const char* what) {
  if (c->warned_generic || c->force_generic) return;
  c->warned_generic = true;
  fprintf(stderr, "mimgpu: %s direction for shape (nx=%d, nu=%d, ng=%d) has no specialised kernel: running the "
                  "shape-generic CTA kernel\n", what, c->sh.nx, c->sh.nu, c->sh.ng);
}
static int post_launch(mimgpu_ctx* c, const char* what) {
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return fail(MIMGPU_ERR_CUDA, std::string(what) + ": " + cudaGetErrorString(e));
  c->launches++;
  return MIMGPU_OK;
}

int mimgpu_sqp_direction(mimgpu_ctx* c, const mimgpu_sqp_params* p, void* stream) {
  int rc = pre_launch(c);
  if (rc) return rc;
  if (!p) return fail(MIMGPU_ERR_INVALID, "null params");
  if (c->k.n_active == 0) return MIMGPU_OK;
  SqpKPar kp{p->preg, p->dreg, p->mu_dynamic, p->is_feasible};
  cudaStream_t st = (cudaStream_t)stream;
  c->last_stream = st;
  c->last_kind = 1;
  const bool fast = (c->use_small & 2) && c->pack_buf && !c->force_generic && !c->small_alt;
  if (fast && !c->records_mode && (c->pack_dirty || c->attached)) {
    /* field-major inputs: (re)build the knot records.  Caller-owned device inputs are re-read at every
     * call -- every computeDirection follows a fresh calcDiff (src/sqp.cpp:224-228) */
    rc = pack_inputs(c, st);
    if (rc) return rc;
  }
  if (!fast) {
    rc = ensure_fields(c, st);
    if (rc) return rc;
  }
  if (c->use_small && !c->force_generic) {
    rc = small_sqp_direction_launch_x(c->k, kp, fast ? 0 : 1, st);
    if (rc >= 0) {
      c->extras_stale = (rc == 0);
      c->spill_packed = (rc == 0);
      c->last_sqp = kp;
      return post_launch(c, rc == 0 ? "k_sqp_direction_small_gj" : "k_sqp_direction_small");
    }
  }
  c->extras_stale = false;
  c->spill_packed = false;
  if (c->cplan.ok && !c->use_small && !c->force_generic &&
      (((uintptr_t)c->k.Fx | (uintptr_t)c->k.Fu) & 15) == 0 && cta_sqp_launch_x(c->k, kp, false, c->cplan, st) == 0)
    return post_launch(c, "k_sqp_direction_cta");
  warn_generic(c, "SQP");
  k_sqp_direction_generic<<<c->k.n_active, c->nt, c->smem_bytes, st>>>(c->k, kp);
  return post_launch(c, "k_sqp_direction_generic");
}

/* SolverSQP::computeDirection for HOST-resident inputs, end to end: the batch is cut into
 * `n_chunks` slices; slice i is uploaded, (re-laid-out,) solved and its dx_ / du_ / lag_mul_ / KKT_
 * downloaded on one of two internal streams, so the PCIe copies of both directions overlap each
 * other and the kernels of the neighbouring slices.  Inputs are either the field-major arrays (d) or
 * knot records (records, cost).  Returns after everything has landed in the host arrays. */
static int sqp_direction_host_impl(mimgpu_ctx* c, const mimgpu_derivs* d, const double* records, const double* cost,
                                   const mimgpu_sqp_params* p, double* dx, double* du, double* lag_mul, double* kkt,
                                   int32_t n_chunks) {
  CU(cudaSetDevice(c->device));
  const mimgpu_shape& sh = c->sh;
  const size_t B = sh.B, T = sh.T, nx = sh.nx, nu = sh.nu;
  const bool sliceable = c->pack_buf && (c->use_small & 2) && !c->force_generic && !c->small_alt && !c->k.active;
  if (!sliceable || n_chunks <= 1) {
    int rc = records ? mimgpu_upload_records(c, records, cost, MIMGPU_MEM_HOST, nullptr)
                     : mimgpu_upload(c, d, MIMGPU_MEM_HOST, nullptr);
    if (rc == MIMGPU_OK) rc = mimgpu_sqp_direction(c, p, nullptr);
    if (rc == MIMGPU_OK && dx) rc = mimgpu_download(c, MIMGPU_F_DX, dx, nullptr);
    if (rc == MIMGPU_OK && du) rc = mimgpu_download(c, MIMGPU_F_DU, du, nullptr);
    if (rc == MIMGPU_OK && lag_mul) rc = mimgpu_download(c, MIMGPU_F_LAG_MUL, lag_mul, nullptr);
    if (rc == MIMGPU_OK && kkt) rc = mimgpu_scalar(c, MIMGPU_S_KKT, kkt, nullptr);
    if (rc == MIMGPU_OK) CU(cudaStreamSynchronize(nullptr));
    return rc;
  }
  if ((size_t)n_chunks > B) n_chunks = (int32_t)B;
  for (int i = 0; i < 2; ++i)
    if (!c->pipe_st[i]) CU(cudaStreamCreateWithFlags(&c->pipe_st[i], cudaStreamNonBlocking));
  size_t n[20];
  const double* src[20] = {};
  double* dst[20] = {};
  const double** slot[20];
  deriv_slots(c->k, slot);
  if (!records) {
    int rc = ensure_in_buf(c, n);
    if (rc) return rc;
    const double* s2[20] = {d->Fx, d->Fu, d->Lxx, d->Lxu, d->Luu, d->Lx, d->Lu, d->fs, d->LxxT, d->LxT,
                            d->Gx, d->Gu, d->g, d->g_lb, d->g_ub, d->GxT, d->gT, d->g_lbT, d->g_ubT, d->cost};
    for (int i = 0; i < 20; ++i) src[i] = s2[i];
    for (int i = 0; i < 19; ++i)
      if (n[i] && !src[i]) return fail(MIMGPU_ERR_INVALID, "missing derivative array #" + std::to_string(i));
    double* q = c->in_buf;
    for (int i = 0; i < 20; ++i) {
      dst[i] = q;
      *slot[i] = (n[i] && src[i]) ? q : nullptr;
      q += (n[i] + 1) & ~(size_t)1;
    }
  } else {
    if (sh.ng || sh.ngT) return fail(MIMGPU_ERR_UNSUPPORTED, "knot records carry no constraint blocks");
    c->k.cost = cost ? c->cost_buf : nullptr;
  }
  /* earlier asynchronous work of this ctx (on the caller's stream) is done with the buffers */
  if (c->last_stream != c->pipe_st[0] && c->last_stream != c->pipe_st[1]) CU(cudaStreamSynchronize(c->last_stream));
  SqpKPar kp{p->preg, p->dreg, p->mu_dynamic, p->is_feasible};
  KArgs ka = c->k;
  ka.pack = c->pack_buf;
  const size_t per = (B + n_chunks - 1) / n_chunks;
  for (size_t b0 = 0, ci = 0; b0 < B; b0 += per, ++ci) {
    const size_t nb = (b0 + per <= B) ? per : B - b0;
    cudaStream_t st = c->pipe_st[ci & 1];
    if (records) {
      CU(cudaMemcpyAsync(c->pack_buf + b0 * c->pack_pp, records + b0 * c->pack_pp, nb * c->pack_pp * sizeof(double),
                         cudaMemcpyHostToDevice, st));
      if (cost) CU(cudaMemcpyAsync(c->cost_buf + b0, cost + b0, nb * sizeof(double), cudaMemcpyHostToDevice, st));
    } else {
      for (int i = 0; i < 20; ++i) {
        if (!(n[i] && src[i])) continue;
        const size_t pb = n[i] / B; /* doubles per problem of this field */
        CU(cudaMemcpyAsync(dst[i] + b0 * pb, src[i] + b0 * pb, nb * pb * sizeof(double), cudaMemcpyHostToDevice, st));
      }
      if (small_pack_launch_x(ka, c->pack_buf, (int)b0, (int)nb, st) != 0)
        return fail(MIMGPU_ERR_CUDA, "k_pack_knots launch failed");
      c->launches++;
    }
    ka.first = (int32_t)b0;
    ka.n_active = (int32_t)nb;
    if (small_sqp_direction_launch_x(ka, kp, 0, st) != 0) return fail(MIMGPU_ERR_CUDA, "k_sqp_direction_small_gj launch failed");
    int rc = post_launch(c, "k_sqp_direction_small_gj");
    if (rc) return rc;
    if (dx) CU(cudaMemcpyAsync(dx + b0 * (T + 1) * nx, c->field[MIMGPU_F_DX] + b0 * (T + 1) * nx,
                               nb * (T + 1) * nx * sizeof(double), cudaMemcpyDeviceToHost, st));
    if (du) CU(cudaMemcpyAsync(du + b0 * T * nu, c->field[MIMGPU_F_DU] + b0 * T * nu, nb * T * nu * sizeof(double),
                               cudaMemcpyDeviceToHost, st));
    if (lag_mul) CU(cudaMemcpyAsync(lag_mul + b0 * (T + 1) * nx, c->field[MIMGPU_F_LAG_MUL] + b0 * (T + 1) * nx,
                                    nb * (T + 1) * nx * sizeof(double), cudaMemcpyDeviceToHost, st));
    if (kkt) CU(cudaMemcpyAsync(kkt + b0, c->k.scal + (size_t)MIMGPU_S_KKT * B + b0, nb * sizeof(double),
                                cudaMemcpyDeviceToHost, st));
  }
  CU(cudaStreamSynchronize(c->pipe_st[0]));
  CU(cudaStreamSynchronize(c->pipe_st[1]));
  c->k.pack = c->pack_buf;
  c->attached = false;
  c->records_mode = records != nullptr;
  c->fields_valid = !c->records_mode;
  c->pack_dirty = false;
  inputs_changed(c);
  c->extras_stale = true;
  c->spill_packed = true;
  c->last_sqp = kp;
  c->last_kind = 1;
  c->last_stream = c->pipe_st[0];
  return MIMGPU_OK;
}

int mimgpu_sqp_direction_host(mimgpu_ctx* c, const mimgpu_derivs* d, const mimgpu_sqp_params* p, double* dx,
                              double* du, double* lag_mul, double* kkt, int32_t n_chunks) {
  if (!c || !d || !p) return fail(MIMGPU_ERR_INVALID, "null pointer");
  return sqp_direction_host_impl(c, d, nullptr, nullptr, p, dx, du, lag_mul, kkt, n_chunks);
}

int mimgpu_sqp_direction_host_records(mimgpu_ctx* c, const double* records, const double* cost,
                                      const mimgpu_sqp_params* p, double* dx, double* du, double* lag_mul,
                                      double* kkt, int32_t n_chunks) {
  if (!c || !records || !p) return fail(MIMGPU_ERR_INVALID, "null pointer");
  if (!(c->use_small & 2) || !c->pack_buf) return fail(MIMGPU_ERR_UNSUPPORTED, "shape has no knot-record kernel");
  return sqp_direction_host_impl(c, nullptr, records, cost, p, dx, du, lag_mul, kkt, n_chunks);
}

int mimgpu_fddp_direction(mimgpu_ctx* c, const mimgpu_sqp_params* p, void* stream) {
  int rc = pre_launch(c);
  if (rc) return rc;
  if (!p) return fail(MIMGPU_ERR_INVALID, "null params");
  if (c->k.n_active == 0) return MIMGPU_OK;
  SqpKPar kp{p->preg, p->dreg, p->mu_dynamic, p->is_feasible};
  c->extras_stale = false;
  c->spill_packed = false;
  c->last_stream = (cudaStream_t)stream;
  c->last_kind = 2;
  rc = ensure_fields(c, (cudaStream_t)stream);
  if (rc) return rc;
  if (c->cplan.ok && !c->force_generic &&
      (((uintptr_t)c->k.Fx | (uintptr_t)c->k.Fu) & 15) == 0 &&
      cta_sqp_launch_x(c->k, kp, true, c->cplan, (cudaStream_t)stream) == 0)
    return post_launch(c, "k_fddp_direction_cta");
  if (!c->use_small) warn_generic(c, "FDDP");
  k_fddp_direction_generic<<<c->k.n_active, c->nt, c->smem_bytes, (cudaStream_t)stream>>>(c->k, kp);
  return post_launch(c, "k_fddp_direction_generic");
}

int mimgpu_csqp_calc_norms(mimgpu_ctx* c, void* stream) {
  int rc = pre_launch(c);
  if (rc) return rc;
  if (c->k.n_active == 0) return MIMGPU_OK;
  c->last_stream = (cudaStream_t)stream;
  rc = ensure_fields(c, (cudaStream_t)stream);
  if (rc) return rc;
  k_csqp_calc_norms<<<c->k.n_active, 128, 0, (cudaStream_t)stream>>>(c->k);
  return post_launch(c, "k_csqp_calc_norms");
}

int mimgpu_csqp_reset_rho(mimgpu_ctx* c, const mimgpu_csqp_params* p, void* stream) {
  int rc = pre_launch(c);
  if (rc) return rc;
  if (!p) return fail(MIMGPU_ERR_INVALID, "null params");
  if (c->k.n_active == 0) return MIMGPU_OK;
  c->last_stream = (cudaStream_t)stream;
  k_csqp_reset_rho<<<c->k.n_active, 128, 0, (cudaStream_t)stream>>>(c->k, p->rho_sparse_base, p->rho_min);
  return post_launch(c, "k_csqp_reset_rho");
}

int mimgpu_csqp_direction(mimgpu_ctx* c, const mimgpu_csqp_params* p, int with_kkt, void* stream) {
  int rc = pre_launch(c);
  if (rc) return rc;
  if (!p) return fail(MIMGPU_ERR_INVALID, "null params");
  if (p->rho_update_interval < 1 || p->max_qp_iters < 0) return fail(MIMGPU_ERR_INVALID, "bad CSQP parameters");
  if (c->k.n_active == 0) return MIMGPU_OK;
  CsqpKPar kp;
  kp.q = *p;
  kp.with_kkt = with_kkt;
  cudaStream_t st = (cudaStream_t)stream;
  c->last_kind = 3;
  c->last_stream = st;
  rc = ensure_fields(c, st);
  if (rc) return rc;
  /* the fast SQP kernel's lazily materialised members are overwritten by this call */
  c->extras_stale = false;
  c->spill_packed = false;
  /* 16-byte rows / TMA bulk copies need 16-byte aligned sources (block sizes are even by construction) */
  const uintptr_t al = (uintptr_t)c->k.Fx | (uintptr_t)c->k.Fu | (uintptr_t)c->k.Gx | (uintptr_t)c->k.Gu;
  /* Measured on B200 (profiles/variants_r02_v1.log): the condensed-chain kernel wins from nx = 32 up (32/16:
   * 13.5 k vs 11.3 k QP/s; Solo12, 48/24, 64/32, Talos); below that a knot is too little work for a
   * warp-per-knot phase: the small shapes go to the record kernel, one warp per problem (UR5: 232 k QP/s against
   * 108 k of the sweep-by-sweep streaming kernel and 80 k of the condensed one, profiles/rec_w1_r02_v10.log). */
  const bool want_admm = c->csqp_admm || (c->sh.nx >= 32 && !c->csqp_stream) || !c->splan.ok;
  if (c->aplan.ok && !c->force_generic && want_admm && (al & 15) == 0) {
    if (!c->admm_scr[0]) {
      const size_t BT = (size_t)c->sh.B * c->sh.T, nx = c->sh.nx;
      const size_t bytes[3] = {BT * nx * nx * 8, BT * nx * 8, BT * nx * 8};
      for (int i = 0; i < 3; ++i) CU(cudaMalloc(&c->admm_scr[i], bytes[i]));
    }
    if (c->admm_timing) { /* debug: per-phase clock64 cycles of CTA 0, printed after the launch (synchronises) */
      CU(cudaMemsetAsync(c->admm_timing, 0, 8 * sizeof(long long), st));
    }
    if (admm_csqp_launch_x(c->k, *p, with_kkt, c->aplan, c->admm_scr, st, c->admm_timing) == 0) {
      rc = post_launch(c, "k_csqp_direction_admm");
      if (rc == MIMGPU_OK && c->admm_timing) {
        long long h[8];
        CU(cudaMemcpyAsync(h, c->admm_timing, sizeof(h), cudaMemcpyDeviceToHost, st));
        CU(cudaStreamSynchronize(st));
        fprintf(stderr, "mimgpu admm timing (cycles, CTA 0): ric %lld post_full %lld chain_bwd %lld gains %lld chain_fwd %lld update %lld\n",
                h[0], h[1], h[2], h[3], h[4], h[5]);
      }
      return rc;
    }
  }
  if (c->rplan.ok && c->csqp_rec && !c->csqp_stream && !c->force_generic && (al & 15) == 0) {
    if (!c->rec_scr && cudaMalloc(&c->rec_scr, (size_t)c->sh.B * c->rplan.rec_doubles * 8) != cudaSuccess) {
      /* no room for the records: the streaming kernel reads the field-major arrays and needs no scratch */
      cudaGetLastError();
      c->rec_scr = nullptr;
      c->csqp_rec = false;
      fprintf(stderr, "mimgpu: no memory for the CSQP knot records (%zu MB), using the streaming kernel\n",
              ((size_t)c->sh.B * c->rplan.rec_doubles * 8) >> 20);
    }
    if (c->rec_scr && rec_csqp_launch_x(c->k, *p, with_kkt, c->rplan, c->rec_scr, st) == 0)
      return post_launch(c, "k_csqp_direction_rec");
  }
  if (c->splan.ok && !c->force_generic && (al & 15) == 0 &&
      stream_csqp_launch_x(c->k, *p, with_kkt, c->splan, st) == 0)
    return post_launch(c, "k_csqp_direction_stream");
  warn_generic(c, "CSQP");
  k_csqp_direction_generic<<<c->k.n_active, c->nt, c->smem_bytes, st>>>(c->k, kp);
  return post_launch(c, "k_csqp_direction_generic");
}

/* SolverCSQP::computeDirection for HOST-resident derivatives, end to end (row n3): the batch is cut into
 * `n_chunks` slices; slice i is uploaded, its calc norms / rho reset (optional) and its whole ADMM loop
 * run, and dx~ / du~ / lag_mul / y / z / KKT are downloaded on one of two internal streams, so the PCIe
 * traffic of one slice overlaps the kernels of its neighbours.  Same results as mimgpu_upload +
 * mimgpu_csqp_calc_norms + mimgpu_csqp_reset_rho + mimgpu_csqp_direction + mimgpu_download. */
int mimgpu_csqp_direction_host(mimgpu_ctx* c, const mimgpu_derivs* d, const mimgpu_csqp_params* p, int with_kkt,
                               int calc_norms, int reset_rho, double* dx_tilde, double* du_tilde, double* lag_mul,
                               double* y, double* z, double* kkt, int32_t* qp_iters, int32_t n_chunks) {
  if (!c || !d || !p) return fail(MIMGPU_ERR_INVALID, "null pointer");
  if (p->rho_update_interval < 1 || p->max_qp_iters < 0) return fail(MIMGPU_ERR_INVALID, "bad CSQP parameters");
  CU(cudaSetDevice(c->device));
  const mimgpu_shape& sh = c->sh;
  const size_t B = sh.B, T = sh.T, nx = sh.nx, nu = sh.nu, nc = T * sh.ng + sh.ngT;
  if (c->k.active) return fail(MIMGPU_ERR_STATE, "mimgpu_csqp_direction_host with an active subset");
  if (n_chunks < 1) n_chunks = 1;
  if ((size_t)n_chunks > B) n_chunks = (int32_t)B;
  for (int i = 0; i < 2; ++i)
    if (!c->pipe_st[i]) CU(cudaStreamCreateWithFlags(&c->pipe_st[i], cudaStreamNonBlocking));
  size_t n[20];
  int rc = ensure_in_buf(c, n);
  if (rc) return rc;
  const double* src[20] = {d->Fx, d->Fu, d->Lxx, d->Lxu, d->Luu, d->Lx, d->Lu, d->fs, d->LxxT, d->LxT,
                           d->Gx, d->Gu, d->g, d->g_lb, d->g_ub, d->GxT, d->gT, d->g_lbT, d->g_ubT, d->cost};
  for (int i = 0; i < 19; ++i)
    if (n[i] && !src[i]) return fail(MIMGPU_ERR_INVALID, "missing derivative array #" + std::to_string(i));
  const double** slot[20];
  deriv_slots(c->k, slot);
  double* dst[20];
  {
    double* q = c->in_buf;
    for (int i = 0; i < 20; ++i) {
      dst[i] = q;
      *slot[i] = (n[i] && src[i]) ? q : nullptr;
      q += (n[i] + 1) & ~(size_t)1;
    }
  }
  if (c->last_stream != c->pipe_st[0] && c->last_stream != c->pipe_st[1]) CU(cudaStreamSynchronize(c->last_stream));
  c->attached = false; c->records_mode = false; c->fields_valid = true; c->k.pack = nullptr; c->pack_dirty = true;
  inputs_changed(c);
  const size_t per = (B + n_chunks - 1) / n_chunks;
  for (size_t b0 = 0, ci = 0; b0 < B; b0 += per, ++ci) {
    const size_t nb = (b0 + per <= B) ? per : B - b0;
    cudaStream_t st = c->pipe_st[ci & 1];
    for (int i = 0; i < 20; ++i) {
      if (!(n[i] && src[i])) continue;
      const size_t pb = n[i] / B;
      CU(cudaMemcpyAsync(dst[i] + b0 * pb, src[i] + b0 * pb, nb * pb * sizeof(double), cudaMemcpyHostToDevice, st));
    }
    c->k.first = (int32_t)b0;
    c->k.n_active = (int32_t)nb;
    if (calc_norms) rc = mimgpu_csqp_calc_norms(c, st);
    if (rc == MIMGPU_OK && reset_rho) rc = mimgpu_csqp_reset_rho(c, p, st);
    if (rc == MIMGPU_OK) rc = mimgpu_csqp_direction(c, p, with_kkt, st);
    c->k.first = 0;
    c->k.n_active = (int32_t)B;
    if (rc) return rc;
    auto out = [&](double* h, int f, size_t pp) -> cudaError_t {
      return (h && pp) ? cudaMemcpyAsync(h + b0 * pp, c->field[f] + b0 * pp, nb * pp * sizeof(double), cudaMemcpyDeviceToHost, st)
                       : cudaSuccess;
    };
    CU(out(dx_tilde, MIMGPU_F_DX_TILDE, (T + 1) * nx));
    CU(out(du_tilde, MIMGPU_F_DU_TILDE, T * nu));
    CU(out(lag_mul, MIMGPU_F_LAG_MUL, (T + 1) * nx));
    CU(out(y, MIMGPU_F_Y, nc));
    CU(out(z, MIMGPU_F_Z, nc));
    if (kkt) CU(cudaMemcpyAsync(kkt + b0, c->k.scal + (size_t)MIMGPU_S_KKT * B + b0, nb * sizeof(double), cudaMemcpyDeviceToHost, st));
    if (qp_iters) CU(cudaMemcpyAsync(qp_iters + b0, c->k.qp_iters + b0, nb * sizeof(int32_t), cudaMemcpyDeviceToHost, st));
  }
  CU(cudaStreamSynchronize(c->pipe_st[0]));
  CU(cudaStreamSynchronize(c->pipe_st[1]));
  c->last_stream = c->pipe_st[0];
  return MIMGPU_OK;
}

/* ------------------------------------------------------ solve() on the device */
int mimgpu_model_attach(mimgpu_ctx* c, const mimgpu_poly_model* m, int memkind, void* stream) {
  if (!c || !m) return fail(MIMGPU_ERR_INVALID, "null pointer");
  if (c->sh.ng || c->sh.ngT) return fail(MIMGPU_ERR_UNSUPPORTED, "device-side solve(): unconstrained models only (ng = ngT = 0)");
  cudaStream_t st = (cudaStream_t)stream;
  CU(cudaSetDevice(c->device));
  const size_t B = c->sh.B, T = c->sh.T, nx = c->sh.nx, nu = c->sh.nu;
  const size_t n[11] = {B * T * nx * nx, B * T * nx * nu, B * T * nx, B * T * nx * nx, B * T * nx * nu, B * T * nu * nu,
                        B * T * nx, B * T * nu, B * nx * nx, B * nx, B * nx};
  const double* src[11] = {m->A, m->Bm, m->f, m->Q, m->S, m->R, m->q, m->r, m->QT, m->qT, m->x0};
  for (int i = 0; i < 11; ++i)
    if (n[i] && !src[i]) return fail(MIMGPU_ERR_INVALID, "missing model array #" + std::to_string(i));
  const double* dev[11];
  if (memkind == MIMGPU_MEM_DEVICE) {
    for (int i = 0; i < 11; ++i) dev[i] = src[i];
  } else if (memkind == MIMGPU_MEM_HOST) {
    size_t tot = 0;
    for (int i = 0; i < 11; ++i) tot += (n[i] + 1) & ~(size_t)1;
    if (!c->model_buf) CU(cudaMalloc(&c->model_buf, tot * sizeof(double)));
    double* q = c->model_buf;
    for (int i = 0; i < 11; ++i) {
      if (n[i]) CU(cudaMemcpyAsync(q, src[i], n[i] * sizeof(double), cudaMemcpyHostToDevice, st));
      dev[i] = q;
      q += (n[i] + 1) & ~(size_t)1;
    }
  } else {
    return fail(MIMGPU_ERR_INVALID, "bad memkind");
  }
  PolyModelDev& d = c->model;
  d.A = dev[0]; d.Bm = dev[1]; d.f = dev[2]; d.Q = dev[3]; d.S = dev[4]; d.R = dev[5]; d.q = dev[6]; d.r = dev[7];
  d.QT = dev[8]; d.qT = dev[9]; d.x0 = dev[10]; d.c3 = m->c3; d.w4 = m->w4;
  /* the derivative arrays the direction kernels read live in the ctx's own buffer: the constant blocks
   * (Fx = A, Fu = B, Lxx = Q, Lxu = S, Luu = R, LxxT = QT) are copied once, k_model_calcdiff fills the rest */
  size_t dn[20];
  int rc = ensure_in_buf(c, dn);
  if (rc) return rc;
  const double** slot[20];
  deriv_slots(c->k, slot);
  double* q = c->in_buf;
  double* at[20];
  for (int i = 0; i < 20; ++i) { at[i] = q; *slot[i] = dn[i] ? q : nullptr; q += (dn[i] + 1) & ~(size_t)1; }
  const int map[6][2] = {{0, 0}, {1, 1}, {2, 3}, {3, 4}, {4, 5}, {8, 8}}; /* deriv slot <- model array */
  for (auto& mm : map)
    if (dn[mm[0]]) CU(cudaMemcpyAsync(at[mm[0]], dev[mm[1]], dn[mm[0]] * sizeof(double), cudaMemcpyDeviceToDevice, st));
  c->have_model = true;
  c->attached = false; c->records_mode = false; c->fields_valid = true; c->k.pack = nullptr; c->pack_dirty = true;
  c->last_stream = st;
  inputs_changed(c);
  return MIMGPU_OK;
}

static int solve_alloc(mimgpu_ctx* c, int filter_size) {
  const size_t B = c->sh.B;
  if (c->solve_buf && c->solve_filter_size == filter_size) return MIMGPU_OK;
  cudaFree(c->solve_buf);
  c->solve_buf = nullptr;
  const size_t bytes = (5 * B + 4) * sizeof(int32_t) + 8 + 2 * B * (size_t)filter_size * sizeof(double);
  CU(cudaMalloc(&c->solve_buf, bytes));
  CU(cudaMemset(c->solve_buf, 0, bytes)); /* the filter ring buffers start at 0 and persist across solve() calls */
  c->solve_filter_size = filter_size;
  SolveDev& s = c->sv;
  double* dl = (double*)c->solve_buf;
  s.cost_list = dl; s.gap_list = dl + B * filter_size;
  int32_t* ip = (int32_t*)(dl + 2 * B * filter_size);
  s.iter = ip; s.done = ip + B; s.flag = ip + 2 * B; s.list_a = ip + 3 * B; s.list_b = ip + 4 * B; s.counts = ip + 5 * B;
  s.xs = c->field[MIMGPU_F_XS]; s.us = c->field[MIMGPU_F_US];
  s.cost = c->k.scal + (size_t)MIMGPU_S_COST * B;
  s.steplength = c->k.scal + (size_t)MIMGPU_S_STEPLENGTH * B;
  s.preg = c->k.scal + (size_t)MIMGPU_S_PREG * B;
  s.dreg = s.preg; /* dreg_ = preg_ throughout (src/ddp.cpp:431-445) */
  return MIMGPU_OK;
}

int mimgpu_sqp_solve(mimgpu_ctx* c, const mimgpu_sqp_solve_params* p, const double* xs0, const double* us0,
                     int32_t* n_sqp_iters, void* stream) {
  if (!c || !p) return fail(MIMGPU_ERR_INVALID, "null pointer");
  if (!c->have_model) return fail(MIMGPU_ERR_STATE, "mimgpu_sqp_solve before mimgpu_model_attach");
  if (p->filter_size < 1 || p->filter_size > SOLVE_MAXFILTER || p->maxiter < 0)
    return fail(MIMGPU_ERR_INVALID, "bad solve parameters");
  cudaStream_t st = (cudaStream_t)stream;
  CU(cudaSetDevice(c->device));
  int rc = solve_alloc(c, p->filter_size);
  if (rc) return rc;
  const mimgpu_shape& sh = c->sh;
  const int B = sh.B;
  SolveDev& s = c->sv;
  const size_t nxs = c->field_n[MIMGPU_F_XS], nus = c->field_n[MIMGPU_F_US];
  if (xs0) CU(cudaMemcpyAsync(s.xs, xs0, nxs * sizeof(double), cudaMemcpyHostToDevice, st));
  else CU(cudaMemsetAsync(s.xs, 0, nxs * sizeof(double), st));
  if (nus) {
    if (us0) CU(cudaMemcpyAsync(s.us, us0, nus * sizeof(double), cudaMemcpyHostToDevice, st));
    else CU(cudaMemsetAsync(s.us, 0, nus * sizeof(double), st));
  }
  const double reg0 = std::isnan(p->reg_init) ? p->reg_min : p->reg_init;
  solve_init_launch_x(s, c->model, s.xs, s.us, sh, reg0, st); /* xs_[0] = x0, preg_ = dreg_ = reg0, iter_ = 0 */
  c->launches++;
  SolvePar sp{p->reg_min, p->reg_max, p->reg_incfactor, p->reg_decfactor, p->th_stepdec, p->th_stepinc,
              p->termination_tol, p->mu_dynamic, p->use_filter_line_search, p->filter_size};
  mimgpu_sqp_params dp{reg0, reg0, p->mu_dynamic, 0, 0};
  /* per-problem regularisation and cost live in the solve state */
  const double* save_cost = c->k.cost;
  c->k.cost = s.cost;
  c->k.preg_arr = s.preg;
  c->k.dreg_arr = s.dreg;
  int32_t counts[4];
  auto read_counts = [&]() -> int {
    CU(cudaMemcpyAsync(counts, s.counts, sizeof(counts), cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    return MIMGPU_OK;
  };
  auto kerr = [&](const char* what) -> int {
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return fail(MIMGPU_ERR_CUDA, std::string(what) + ": " + cudaGetErrorString(e));
    c->launches++;
    return MIMGPU_OK;
  };
  /* one pass of "computeDirection with the regularisation retry + KKT test" on the running problems
   * (list_a, n_run of them); leaves the line-search candidates in list_b, their number in counts[1] */
  auto direction_pass = [&](int n_run) -> int {
    if (solve_calcdiff_launch_x(c->k, c->model, s, s.list_a, n_run, st) != 0)
      return fail(MIMGPU_ERR_UNSUPPORTED, "model kernels: shape needs more than 48 KB of shared memory");
    if ((rc = kerr("k_model_calcdiff"))) return rc;
    c->pack_dirty = true;
    CU(cudaMemsetAsync(s.counts, 0, 2 * sizeof(int32_t), st));
    int n_dir = n_run;
    int n_ls = 0;
    int32_t *cur = s.list_a, *nxt = s.flag;   /* the retry list ping-pongs between two buffers */
    while (n_dir > 0) {
      c->k.active = cur; c->k.n_active = n_dir;
      rc = mimgpu_sqp_direction(c, &dp, st);
      if (rc) return rc;
      solve_post_direction_launch_x(c->k, s, sp, cur, n_dir, nxt, st);
      if ((rc = kerr("k_sqp_post_direction"))) return rc;
      if ((rc = read_counts())) return rc;
      n_ls = counts[1];
      n_dir = counts[0];                        /* problems whose backward pass failed: again, with more regularisation */
      if (n_dir > 0) CU(cudaMemsetAsync(s.counts, 0, sizeof(int32_t), st));
      std::swap(cur, nxt);
    }
    counts[1] = n_ls;
    return MIMGPU_OK;
  };
  auto running = [&](int* n_run) -> int {
    CU(cudaMemsetAsync(s.counts + 2, 0, sizeof(int32_t), st));
    solve_running_launch_x(s, B, st);
    if ((rc = kerr("k_solve_running"))) return rc;
    if ((rc = read_counts())) return rc;
    *n_run = counts[2];
    return MIMGPU_OK;
  };
  int n_run = 0, it = 0;
  rc = running(&n_run);
  for (; rc == MIMGPU_OK && it < p->maxiter && n_run > 0; ++it) {
    if ((rc = direction_pass(n_run))) break;
    if (counts[1] > 0) {
      if (solve_linesearch_launch_x(c->k, c->model, s, sp, s.list_b, counts[1], st) != 0) {
        rc = fail(MIMGPU_ERR_UNSUPPORTED, "line-search kernel: shape needs more than 48 KB of shared memory");
        break;
      }
      if ((rc = kerr("k_sqp_linesearch"))) break;
    }
    if ((rc = running(&n_run))) break;
  }
  if (rc == MIMGPU_OK && p->extra_iteration_for_last_kkt && n_run > 0) rc = direction_pass(n_run);
  c->k.active = nullptr; c->k.n_active = B;
  c->k.cost = s.cost; (void)save_cost;
  c->last_stream = st;
  if (n_sqp_iters) *n_sqp_iters = it;
  return rc;
}

int mimgpu_solve_info(mimgpu_ctx* c, int32_t* iters, int32_t* outcome, void* stream) {
  if (!c) return fail(MIMGPU_ERR_INVALID, "null pointer");
  if (!c->solve_buf) return fail(MIMGPU_ERR_STATE, "no mimgpu_sqp_solve has run on this ctx");
  CU(cudaSetDevice(c->device));
  const size_t B = c->sh.B;
  if (iters) CU(cudaMemcpyAsync(iters, c->sv.iter, B * sizeof(int32_t), cudaMemcpyDeviceToHost, (cudaStream_t)stream));
  if (outcome) CU(cudaMemcpyAsync(outcome, c->sv.done, B * sizeof(int32_t), cudaMemcpyDeviceToHost, (cudaStream_t)stream));
  CU(cudaStreamSynchronize((cudaStream_t)stream));
  return MIMGPU_OK;
}

/* ----------------------------------------------------------------- data out */
/* Qu_, Quuk_ and Quu_llt_ are solver members the reference keeps per knot but that neither the
 * SQP forward pass nor its outputs use; the fast small-shape kernel does not spill them.  When
 * a caller asks for one of them the direction is recomputed by the full-spill kernel (same
 * inputs, same parameters), once. */
static int refresh_extras(mimgpu_ctx* c, int field, cudaStream_t st) {
  if (c->spill_packed && (field == MIMGPU_F_VXX || field == MIMGPU_F_VX || field == MIMGPU_F_K ||
                          field == MIMGPU_F_KFF)) {
    /* Vxx_ Vx_ K_ k_ of the fast kernel live in its knot-packed spill records: copy them out once */
    if (small_unpack_launch_x(c->k, st) != 0) return fail(MIMGPU_ERR_CUDA, "k_unpack_spill launch failed");
    int rc = post_launch(c, "k_unpack_spill");
    if (rc) return rc;
    c->spill_packed = false;
    return MIMGPU_OK;
  }
  if (!c->extras_stale) return MIMGPU_OK;
  if (field != MIMGPU_F_QU && field != MIMGPU_F_QUUK && field != MIMGPU_F_QUU_LLT) return MIMGPU_OK;
  int rc0 = ensure_fields(c, st);
  if (rc0) return rc0;
  if (c->k.n_active > 0) {
    KArgs ka = c->k;
    ka.pad_ = 1; /* members only: dx_ / du_ / lag_mul_ / scalars / status of the fast launch are left alone */
    if (small_sqp_direction_launch_x(ka, c->last_sqp, 1, st) != 1)
      return fail(MIMGPU_ERR_CUDA, "full-spill kernel launch failed");
    int rc = post_launch(c, "k_sqp_direction_small");
    if (rc) return rc;
  }
  c->extras_stale = false;
  c->spill_packed = false;
  return MIMGPU_OK;
}

int mimgpu_download(mimgpu_ctx* c, int field, double* dst, void* stream) {
  if (!c || !dst) return fail(MIMGPU_ERR_INVALID, "null pointer");
  if (field < 0 || field >= MIMGPU_F_COUNT) return fail(MIMGPU_ERR_INVALID, "bad field id");
  CU(cudaSetDevice(c->device));
  int rc = refresh_extras(c, field, (cudaStream_t)stream);
  if (rc) return rc;
  if (c->field_n[field])
    CU(cudaMemcpyAsync(dst, c->field[field], c->field_n[field] * sizeof(double), cudaMemcpyDeviceToHost,
                       (cudaStream_t)stream));
  return MIMGPU_OK;
}

int mimgpu_field_ptr(mimgpu_ctx* c, int field, double** dev_ptr, int64_t* n_doubles, void* stream) {
  if (!c) return fail(MIMGPU_ERR_INVALID, "null pointer");
  if (field < 0 || field >= MIMGPU_F_COUNT) return fail(MIMGPU_ERR_INVALID, "bad field id");
  if (c->extras_stale || c->spill_packed) { /* lazily materialised members: enqueued on the caller's stream */
    CU(cudaSetDevice(c->device));
    int rc = refresh_extras(c, field, (cudaStream_t)stream);
    if (rc) return rc;
  }
  if (dev_ptr) *dev_ptr = c->field[field];
  if (n_doubles) *n_doubles = (int64_t)c->field_n[field];
  return MIMGPU_OK;
}

int mimgpu_scalar(mimgpu_ctx* c, int which, double* dst, void* stream) {
  if (!c || !dst) return fail(MIMGPU_ERR_INVALID, "null pointer");
  if (which < 0 || which >= MIMGPU_S_COUNT) return fail(MIMGPU_ERR_INVALID, "bad scalar id");
  CU(cudaSetDevice(c->device));
  CU(cudaMemcpyAsync(dst, c->k.scal + (size_t)which * c->sh.B, c->sh.B * sizeof(double), cudaMemcpyDeviceToHost,
                     (cudaStream_t)stream));
  return MIMGPU_OK;
}

int mimgpu_iters_status(mimgpu_ctx* c, int32_t* qp_iters, int32_t* status, void* stream) {
  if (!c) return fail(MIMGPU_ERR_INVALID, "null pointer");
  CU(cudaSetDevice(c->device));
  const size_t B = c->sh.B;
  if (qp_iters)
    CU(cudaMemcpyAsync(qp_iters, c->k.qp_iters, B * sizeof(int32_t), cudaMemcpyDeviceToHost, (cudaStream_t)stream));
  if (status)
    CU(cudaMemcpyAsync(status, c->k.status, B * sizeof(int32_t), cudaMemcpyDeviceToHost, (cudaStream_t)stream));
  return MIMGPU_OK;
}

int mimgpu_converged(mimgpu_ctx* c, int32_t* converged, void* stream) {
  if (!c || !converged) return fail(MIMGPU_ERR_INVALID, "null pointer");
  CU(cudaSetDevice(c->device));
  CU(cudaMemcpyAsync(converged, c->k.converged, c->sh.B * sizeof(int32_t), cudaMemcpyDeviceToHost,
                     (cudaStream_t)stream));
  return MIMGPU_OK;
}

int mimgpu_summary_reduce(mimgpu_ctx* c, double termination_tol, mimgpu_summary* out, void* stream) {
  if (!c || !out) return fail(MIMGPU_ERR_INVALID, "null pointer");
  CU(cudaSetDevice(c->device));
  cudaStream_t st = (cudaStream_t)stream;
  const int csqp_kind = (c->last_kind == 3); /* n_converged counts ADMM breaks for CSQP, KKT <= tol otherwise */
  k_summary<<<1, 1024, 0, st>>>(c->k.scal + (size_t)MIMGPU_S_KKT * c->sh.B, c->k.qp_iters, c->k.status,
                                c->k.converged, c->sh.B, termination_tol, csqp_kind, (SummaryDev*)c->summary_dev);
  int rc = post_launch(c, "k_summary");
  if (rc) return rc;
  SummaryDev h;
  CU(cudaMemcpyAsync(&h, c->summary_dev, sizeof(h), cudaMemcpyDeviceToHost, st));
  CU(cudaStreamSynchronize(st));
  out->max_kkt = h.max_kkt; out->sum_qp_iters = h.sum_it; out->max_qp_iters = h.max_it;
  out->n_converged = h.n_conv; out->n_failed = h.n_fail;
  return MIMGPU_OK;
}

int mimgpu_stream_sync(mimgpu_ctx* c, void* stream) {
  if (!c) return fail(MIMGPU_ERR_INVALID, "null pointer");
  CU(cudaSetDevice(c->device));
  CU(cudaStreamSynchronize((cudaStream_t)stream));
  return MIMGPU_OK;
}

int64_t mimgpu_launch_count(const mimgpu_ctx* c) { return c ? c->launches : 0; }

} /* extern "C" */
