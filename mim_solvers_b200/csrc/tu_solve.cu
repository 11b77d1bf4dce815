/* translation unit of the device-side solve() kernels (kernels_solve.cuh) */
#include "launchers.h"
#include "kernels_solve.cuh"

namespace mimgpu {

int solve_init_launch_x(const SolveDev& s, const PolyModelDev& m, const double* xs0, const double* us0,
                        const mimgpu_shape& sh, double reg0, cudaStream_t st) {
  k_solve_init<<<296, 256, 0, st>>>(s, m, xs0, us0, sh.B, sh.T, sh.nx, sh.nu, reg0);
  return 0;
}
int solve_calcdiff_launch_x(const KArgs& a, const PolyModelDev& m, const SolveDev& s, const int32_t* list, int n,
                            cudaStream_t st) {
  const int nt = 256, nw = nt / 32;
  const size_t smem = ((size_t)a.sh.T + 2 + (size_t)nw * (2 * a.sh.nx + a.sh.nu)) * sizeof(double);
  if (smem > 48 * 1024) return -1;
  k_model_calcdiff<<<n, nt, smem, st>>>(a, m, s, list, n);
  return 0;
}
int solve_post_direction_launch_x(const KArgs& a, const SolveDev& s, const SolvePar& sp, const int32_t* list, int n,
                                  int32_t* retry_out, cudaStream_t st) {
  k_sqp_post_direction<<<(n + 127) / 128, 128, 0, st>>>(a, s, sp, list, n, retry_out);
  return 0;
}
int solve_linesearch_launch_x(const KArgs& a, const PolyModelDev& m, const SolveDev& s, const SolvePar& sp,
                              const int32_t* list, int n, cudaStream_t st) {
  const size_t smem = (2 * SOLVE_NALPHA + 2 + (size_t)SOLVE_NALPHA * (3 * a.sh.nx + a.sh.nu)) * sizeof(double);
  if (smem > 48 * 1024) return -1;
  k_sqp_linesearch<<<n, 32 * SOLVE_NALPHA, smem, st>>>(a, m, s, sp, list, n);
  return 0;
}
int solve_running_launch_x(const SolveDev& s, int B, cudaStream_t st) {
  k_solve_running<<<(B + 255) / 256, 256, 0, st>>>(s, B);
  return 0;
}

}  // namespace mimgpu
