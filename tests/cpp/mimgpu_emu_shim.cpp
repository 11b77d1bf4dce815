/*
 * mimgpu_emu_shim.cpp -- TEST INFRASTRUCTURE ONLY.  The subset of the C ABI of include/mimgpu.h
 * that the C++ host layer (include/mim_solvers_b200/batch_solvers.hpp) calls, implemented on
 * the CUDA kernel bodies of mim_solvers_b200/csrc/kernels_generic.cuh compiled for the host
 * (-DMIMGPU_EMU, as tests/emu/emu.cpp does), so that the host logic of the C++ solve() loops
 * can be exercised in the GPU-less build container.  Never linked into libmimgpu.so, never
 * loaded by the product; the GPU build of the same test links libmimgpu.so instead.
 */
#define MIMGPU_EMU 1
#include <cmath>
#include <cstring>
#include <limits>
#include <vector>

#include "../../mim_solvers_b200/csrc/kernels_generic.cuh"

using namespace mimgpu;

struct mimgpu_ctx {
  mimgpu_shape sh;
  KArgs k;
  std::vector<double> f[MIMGPU_F_COUNT], vxx_fs, scal, preg, dreg;
  std::vector<int32_t> qp_iters, status, converged, active;
  int64_t launches = 0;
};

static size_t field_size(const mimgpu_shape& s, int f) {
  const size_t B = s.B, T = s.T, nx = s.nx, nu = s.nu, nc = T * s.ng + s.ngT;
  switch (f) {
    case MIMGPU_F_DX: case MIMGPU_F_LAG_MUL: case MIMGPU_F_DX_TILDE: case MIMGPU_F_VX: return B * (T + 1) * nx;
    case MIMGPU_F_DU: case MIMGPU_F_DU_TILDE: case MIMGPU_F_KFF: case MIMGPU_F_QU: case MIMGPU_F_QUUK: return B * T * nu;
    case MIMGPU_F_Y: case MIMGPU_F_Z: case MIMGPU_F_RHO_VEC: return B * nc;
    case MIMGPU_F_VXX: return B * (T + 1) * nx * nx;
    case MIMGPU_F_K: return B * T * nu * nx;
    case MIMGPU_F_QUU_LLT: return B * T * nu * nu;
  }
  return 0;
}

template <class F>
static void for_active(mimgpu_ctx* c, F body) {
  const KArgs& a = c->k;
  for (int i = 0; i < a.n_active; ++i) body(a.active ? a.active[i] : i);
  ++c->launches;
}
static std::vector<double> smem_for(const KArgs& a) {
  Smem s = smem_layout(a.sh.nx, a.sh.nu, a.sh.ng, a.sh.ngT, a.stage_fx);
  return std::vector<double>((size_t)s.total + 8, 0.0);
}

extern "C" {

int mimgpu_abi_version(void) { return MIMGPU_ABI_VERSION; }
int mimgpu_device_count(int* n) { *n = 0; return MIMGPU_OK; }
const char* mimgpu_last_error(void) { return "emulator shim"; }

int mimgpu_create(mimgpu_ctx** out, int, const mimgpu_shape* shape) {
  auto* c = new mimgpu_ctx();
  c->sh = *shape;
  std::memset(&c->k, 0, sizeof c->k);
  c->k.sh = *shape;
  for (int f = 0; f < MIMGPU_F_COUNT; ++f) c->f[f].assign(field_size(*shape, f) + 1, 0.0);
  c->vxx_fs.assign((size_t)shape->B * shape->T * shape->nx + 1, 0.0);
  c->scal.assign((size_t)MIMGPU_S_COUNT * shape->B, 0.0);
  for (int b = 0; b < shape->B; ++b) {
    c->scal[(size_t)MIMGPU_S_KKT * shape->B + b] = std::numeric_limits<double>::infinity();
    c->scal[(size_t)MIMGPU_S_RHO_SPARSE * shape->B + b] = 1e-1;
  }
  c->qp_iters.assign(shape->B, 0); c->status.assign(shape->B, 0); c->converged.assign(shape->B, 0);
  KArgs& k = c->k;
  k.dx = c->f[MIMGPU_F_DX].data(); k.du = c->f[MIMGPU_F_DU].data(); k.lag = c->f[MIMGPU_F_LAG_MUL].data();
  k.dxt = c->f[MIMGPU_F_DX_TILDE].data(); k.dut = c->f[MIMGPU_F_DU_TILDE].data();
  k.y = c->f[MIMGPU_F_Y].data(); k.z = c->f[MIMGPU_F_Z].data(); k.rho = c->f[MIMGPU_F_RHO_VEC].data();
  k.Vxx = c->f[MIMGPU_F_VXX].data(); k.Vx = c->f[MIMGPU_F_VX].data(); k.K = c->f[MIMGPU_F_K].data();
  k.kff = c->f[MIMGPU_F_KFF].data(); k.Qu = c->f[MIMGPU_F_QU].data(); k.Quuk = c->f[MIMGPU_F_QUUK].data();
  k.L = c->f[MIMGPU_F_QUU_LLT].data(); k.Vxx_fs = c->vxx_fs.data();
  k.scal = c->scal.data(); k.qp_iters = c->qp_iters.data(); k.status = c->status.data(); k.converged = c->converged.data();
  k.n_active = shape->B; k.stage_fx = 1;
  *out = c;
  return MIMGPU_OK;
}
void mimgpu_destroy(mimgpu_ctx* c) { delete c; }
int mimgpu_get_shape(const mimgpu_ctx* c, mimgpu_shape* s) { *s = c->sh; return MIMGPU_OK; }

int mimgpu_upload(mimgpu_ctx* c, const mimgpu_derivs* d, int, void*) {
  KArgs& k = c->k;   /* host pointers attached: the caller keeps them alive (HostDerivs member) */
  k.Fx = d->Fx; k.Fu = d->Fu; k.Lxx = d->Lxx; k.Lxu = d->Lxu; k.Luu = d->Luu; k.Lx = d->Lx; k.Lu = d->Lu;
  k.fs = d->fs; k.LxxT = d->LxxT; k.LxT = d->LxT; k.Gx = d->Gx; k.Gu = d->Gu; k.g = d->g; k.glb = d->g_lb;
  k.gub = d->g_ub; k.GxT = d->GxT; k.gT = d->gT; k.glbT = d->g_lbT; k.gubT = d->g_ubT; k.cost = d->cost;
  return MIMGPU_OK;
}
int mimgpu_set_reg(mimgpu_ctx* c, const double* preg, const double* dreg, void*) {
  if (preg) c->preg.assign(preg, preg + c->sh.B);
  if (dreg) c->dreg.assign(dreg, dreg + c->sh.B);
  c->k.preg_arr = preg ? c->preg.data() : nullptr;
  c->k.dreg_arr = dreg ? c->dreg.data() : nullptr;
  return MIMGPU_OK;
}
int mimgpu_set_active(mimgpu_ctx* c, const int32_t* idx, int32_t n, void*) {
  if (!idx || n < 0) { c->k.active = nullptr; c->k.n_active = c->sh.B; return MIMGPU_OK; }
  c->active.assign(idx, idx + n);
  c->k.active = c->active.data(); c->k.n_active = n;
  return MIMGPU_OK;
}
int mimgpu_set_field(mimgpu_ctx* c, int f, const double* src, void*) {
  if (f < 0 || f >= MIMGPU_F_COUNT) return MIMGPU_ERR_INVALID;
  std::memcpy(c->f[f].data(), src, field_size(c->sh, f) * sizeof(double));
  return MIMGPU_OK;
}
int mimgpu_sqp_direction(mimgpu_ctx* c, const mimgpu_sqp_params* p, void*) {
  if (!c->k.Fx) return MIMGPU_ERR_STATE;
  std::vector<double> sm = smem_for(c->k);
  SqpKPar kp{p->preg, p->dreg, p->mu_dynamic, p->is_feasible};
  for_active(c, [&](int b) { sqp_direction_body(c->k, b, sm.data(), kp); });
  return MIMGPU_OK;
}
int mimgpu_fddp_direction(mimgpu_ctx* c, const mimgpu_sqp_params* p, void*) {
  if (!c->k.Fx) return MIMGPU_ERR_STATE;
  std::vector<double> sm = smem_for(c->k);
  SqpKPar kp{p->preg, p->dreg, p->mu_dynamic, p->is_feasible};
  for_active(c, [&](int b) { fddp_direction_body(c->k, b, sm.data(), kp); });
  return MIMGPU_OK;
}
int mimgpu_csqp_calc_norms(mimgpu_ctx* c, void*) {
  double red[64];
  for_active(c, [&](int b) { csqp_calc_norms_body(c->k, b, red); });
  return MIMGPU_OK;
}
int mimgpu_csqp_reset_rho(mimgpu_ctx* c, const mimgpu_csqp_params* p, void*) {
  for_active(c, [&](int b) { csqp_reset_rho_body(c->k, b, p->rho_sparse_base, p->rho_min); });
  return MIMGPU_OK;
}
int mimgpu_csqp_direction(mimgpu_ctx* c, const mimgpu_csqp_params* p, int with_kkt, void*) {
  if (!c->k.Fx) return MIMGPU_ERR_STATE;
  std::vector<double> sm = smem_for(c->k);
  CsqpKPar kp;
  kp.q = *p;
  kp.with_kkt = with_kkt;
  for_active(c, [&](int b) { csqp_direction_body(c->k, b, sm.data(), kp); });
  return MIMGPU_OK;
}
int mimgpu_download(mimgpu_ctx* c, int f, double* dst, void*) {
  if (f < 0 || f >= MIMGPU_F_COUNT) return MIMGPU_ERR_INVALID;
  std::memcpy(dst, c->f[f].data(), field_size(c->sh, f) * sizeof(double));
  return MIMGPU_OK;
}
int mimgpu_scalar(mimgpu_ctx* c, int which, double* dst, void*) {
  if (which < 0 || which >= MIMGPU_S_COUNT) return MIMGPU_ERR_INVALID;
  std::memcpy(dst, c->scal.data() + (size_t)which * c->sh.B, c->sh.B * sizeof(double));
  return MIMGPU_OK;
}
int mimgpu_iters_status(mimgpu_ctx* c, int32_t* qp, int32_t* st, void*) {
  if (qp) std::memcpy(qp, c->qp_iters.data(), c->sh.B * sizeof(int32_t));
  if (st) std::memcpy(st, c->status.data(), c->sh.B * sizeof(int32_t));
  return MIMGPU_OK;
}
int mimgpu_converged(mimgpu_ctx* c, int32_t* cv, void*) {
  std::memcpy(cv, c->converged.data(), c->sh.B * sizeof(int32_t));
  return MIMGPU_OK;
}
int mimgpu_summary_reduce(mimgpu_ctx* c, double, mimgpu_summary* out, void*) {
  std::memset(out, 0, sizeof *out);
  for (int b = 0; b < c->sh.B; ++b) {
    out->max_kkt = std::fmax(out->max_kkt, c->scal[(size_t)MIMGPU_S_KKT * c->sh.B + b]);
    out->sum_qp_iters += c->qp_iters[b];
    out->n_converged += c->converged[b];
    out->n_failed += c->status[b] != 0;
  }
  return MIMGPU_OK;
}
int mimgpu_stream_sync(mimgpu_ctx*, void*) { return MIMGPU_OK; }
int64_t mimgpu_launch_count(const mimgpu_ctx* c) { return c->launches; }
}
