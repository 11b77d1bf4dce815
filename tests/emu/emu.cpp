/*
 * emu.cpp -- TEST INFRASTRUCTURE ONLY.  Compiles the CUDA kernel bodies of
 * mim_solvers_b200/csrc/kernels_generic.cuh for the host (-DMIMGPU_EMU: every CTA
 * phase becomes a serial loop) so that index logic can be debugged in the GPU-less
 * build container.  Never linked into libmimgpu.so, never loaded by the product.
 */
#define MIMGPU_EMU 1
#include <cstdlib>
#include <vector>

#include "../../mim_solvers_b200/csrc/kernels_generic.cuh"

using namespace mimgpu;

extern "C" {

int emu_sizeof_kargs(void) { return (int)sizeof(KArgs); }

static std::vector<double> smem_for(const KArgs* a) {
  Smem s = smem_layout(a->sh.nx, a->sh.nu, a->sh.ng, a->sh.ngT, a->stage_fx);
  return std::vector<double>((size_t)s.total + 8, 0.0);
}

int emu_sqp_direction(const KArgs* a, const mimgpu_sqp_params* p) {
  std::vector<double> sm = smem_for(a);
  SqpKPar kp{p->preg, p->dreg, p->mu_dynamic, p->is_feasible};
  for (int i_ = 0, b = 0; i_ < a->n_active && ((b = a->active ? a->active[i_] : i_), true); ++i_) sqp_direction_body(*a, b, sm.data(), kp);
  return 0;
}
int emu_fddp_direction(const KArgs* a, const mimgpu_sqp_params* p) {
  std::vector<double> sm = smem_for(a);
  SqpKPar kp{p->preg, p->dreg, p->mu_dynamic, p->is_feasible};
  for (int i_ = 0, b = 0; i_ < a->n_active && ((b = a->active ? a->active[i_] : i_), true); ++i_) fddp_direction_body(*a, b, sm.data(), kp);
  return 0;
}
int emu_csqp_direction(const KArgs* a, const mimgpu_csqp_params* p, int with_kkt) {
  std::vector<double> sm = smem_for(a);
  CsqpKPar kp;
  kp.q = *p;
  kp.with_kkt = with_kkt;
  for (int i_ = 0, b = 0; i_ < a->n_active && ((b = a->active ? a->active[i_] : i_), true); ++i_) csqp_direction_body(*a, b, sm.data(), kp);
  return 0;
}
int emu_csqp_reset_rho(const KArgs* a, const mimgpu_csqp_params* p) {
  for (int i_ = 0, b = 0; i_ < a->n_active && ((b = a->active ? a->active[i_] : i_), true); ++i_) csqp_reset_rho_body(*a, b, p->rho_sparse_base, p->rho_min);
  return 0;
}
int emu_csqp_calc_norms(const KArgs* a) {
  double red[64];
  for (int i_ = 0, b = 0; i_ < a->n_active && ((b = a->active ? a->active[i_] : i_), true); ++i_) csqp_calc_norms_body(*a, b, red);
  return 0;
}
}
