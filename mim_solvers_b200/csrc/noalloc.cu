/*
 * noalloc.cu -- mimgpu_sqp_direction_records_noalloc: the warp-per-problem SQP direction kernel
 * (k_sqp_direction_small_gj, kernels_small_gj.cuh) without a ctx and without ANY allocation.
 *
 * No counterpart in the reference; added for callers that
 * cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid it).  The caller passes device inputs (knot records) and device outputs; the kernel's
 * scratch -- the spill records [Vxx_t | Vx_t | K_t | k_t] that its backward sweep writes and its
 * forward sweep reads back, plus the qp_iters word it clears -- lives in the file-scope device
 * array below (MIMGPU_NOALLOC_SCRATCH_BYTES; every GPU that loads the library holds it until exit).
 *
 * g_noalloc_scratch is a HOST-SIDE STAND-IN in host code, not a device address: the address the
 * kernel gets is fetched once per device, with that device current, by cudaGetSymbolAddress.
 *
 * What must start at zero: dx_[0] of every problem in the caller's dx buffer (the kernel reads it as the
 * member that stays zero) -- cleared here by a strided cudaMemset2DAsync on the caller's stream.
 * Nothing in the scratch has to start at zero: the backward sweep writes every spill record of
 * every problem before the forward sweep of the same warp reads it, and problems whose backward
 * sweep fails (status != 0) never read theirs.  All calls share the array, which keeps the last
 * call's contents: calls that overlap on the GPU (two streams, two host threads) RACE.
 */
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <mutex>
#include <string>

#include "../../include/mimgpu.h"
#include "launchers.h"

namespace mimgpu {
void set_last_error(const std::string& msg); /* mimgpu.cu: the text mimgpu_last_error returns */
}

static __device__ __align__(256) unsigned char g_noalloc_scratch[MIMGPU_NOALLOC_SCRATCH_BYTES];

namespace {

constexpr int kMaxDevices = 64;
struct DeviceSetup {
  unsigned char* scratch = nullptr; /* device address of g_noalloc_scratch on this device */
  int configured[8][2] = {};        /* (nx, nu) pairs whose kernel attributes are set on this device */
  int n_configured = 0;
};
std::mutex g_mu;
DeviceSetup g_dev[kMaxDevices];

int fail(int code, const std::string& msg) {
  mimgpu::set_last_error(msg);
  return code;
}

size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

/* bytes of scratch one call needs: spill records of all problems, then qp_iters [B] */
int scratch_layout(const mimgpu_shape& s, size_t* spill_bytes, size_t* total) {
  size_t pack = 0, spill = 0;
  if (mimgpu::small_pack_layout_x(s, &pack, &spill) != 0 || spill == 0) return -1;
  *spill_bytes = align_up((size_t)s.B * spill * sizeof(double), 256);
  *total = *spill_bytes + align_up((size_t)s.B * sizeof(int32_t), 256);
  return 0;
}

}  // namespace

extern "C" int64_t mimgpu_noalloc_scratch_needed(const mimgpu_shape* shape) {
  size_t spill = 0, total = 0;
  if (!shape || shape->B < 0 || shape->T < 1 || scratch_layout(*shape, &spill, &total) != 0) return -1;
  return (int64_t)total;
}

extern "C" int mimgpu_sqp_direction_records_noalloc(const mimgpu_shape* shape, const double* records,
                                                    const double* cost, const mimgpu_sqp_params* p, double* dx,
                                                    double* du, double* lag_mul, double* scalars, int32_t* status,
                                                    void* cuda_stream) {
  using namespace mimgpu;
  if (!shape || !p) return fail(MIMGPU_ERR_INVALID, "null shape / params");
  const mimgpu_shape s = *shape;
  if (s.B < 0 || s.T < 1 || s.nx < 1 || s.nu < 1) return fail(MIMGPU_ERR_INVALID, "bad shape");
  if (s.ng != 0 || s.ngT != 0)
    return fail(MIMGPU_ERR_UNSUPPORTED, "mimgpu_sqp_direction_records_noalloc: unconstrained shapes only (ng = ngT = 0)");
  if (s.B == 0) return MIMGPU_OK;
  if (!records || !dx || !du || !lag_mul || !scalars || !status)
    return fail(MIMGPU_ERR_INVALID, "null device pointer (records, dx, du, lag_mul, scalars, status are required)");
  if (((uintptr_t)records & 15) != 0) return fail(MIMGPU_ERR_INVALID, "records must be 16-byte aligned");
  if (getenv("MIMGPU_SMALL_FULL") || getenv("MIMGPU_SMALL_NOMMA"))
    return fail(MIMGPU_ERR_UNSUPPORTED, "the record kernel is switched off by MIMGPU_SMALL_FULL / MIMGPU_SMALL_NOMMA");
  size_t spill_bytes = 0, total = 0;
  if (scratch_layout(s, &spill_bytes, &total) != 0)
    return fail(MIMGPU_ERR_UNSUPPORTED, "shape (nx=" + std::to_string(s.nx) + ", nu=" + std::to_string(s.nu) +
                                            ") is not served by the knot-record kernel");
  if (total > (size_t)MIMGPU_NOALLOC_SCRATCH_BYTES)
    return fail(MIMGPU_ERR_INVALID,
                "mimgpu_sqp_direction_records_noalloc: B=" + std::to_string(s.B) + ", T=" + std::to_string(s.T) +
                    " needs " + std::to_string(total) + " bytes of scratch, the static array holds " +
                    std::to_string((size_t)MIMGPU_NOALLOC_SCRATCH_BYTES) +
                    " (MIMGPU_NOALLOC_SCRATCH_BYTES): split the batch or use a ctx (mimgpu_create)");
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return fail(MIMGPU_ERR_NODEVICE, std::string("cudaGetDevice: ") + cudaGetErrorString(e));
  if (dev < 0 || dev >= kMaxDevices) return fail(MIMGPU_ERR_INVALID, "device ordinal out of range");
  unsigned char* base = nullptr;
  {
    /* set-up, once per device, with that device current (it is: dev = cudaGetDevice) */
    std::lock_guard<std::mutex> lk(g_mu);
    DeviceSetup& d = g_dev[dev];
    if (!d.scratch) {
      void* ptr = nullptr;
      e = cudaGetSymbolAddress(&ptr, g_noalloc_scratch);
      if (e != cudaSuccess || !ptr)
        return fail(MIMGPU_ERR_CUDA, std::string("cudaGetSymbolAddress(g_noalloc_scratch): ") + cudaGetErrorString(e));
      d.scratch = (unsigned char*)ptr;
    }
    bool done = false;
    for (int i = 0; i < d.n_configured; ++i) done |= d.configured[i][0] == s.nx && d.configured[i][1] == s.nu;
    if (!done) {
      /* dynamic shared-memory attribute of the kernel instantiation (per device; allocates nothing) */
      if (!(small_kernels_configure_x(s) & 2))
        return fail(MIMGPU_ERR_UNSUPPORTED, "no knot-record kernel for this shape on this device");
      if (d.n_configured < 8) {
        d.configured[d.n_configured][0] = s.nx;
        d.configured[d.n_configured][1] = s.nu;
        ++d.n_configured;
      }
    }
    base = d.scratch;
  }
  KArgs a{};
  a.sh = s;
  a.cost = cost;
  a.dx = dx;
  a.du = du;
  a.lag = lag_mul;
  a.scal = scalars;
  a.status = status;
  a.qp_iters = (int32_t*)(base + spill_bytes);
  a.pack = records;
  a.spill = (double*)base;
  a.n_active = s.B;
  a.active = nullptr;
  a.first = 0;
  SqpKPar kp{p->preg, p->dreg, p->mu_dynamic, p->is_feasible};
  /* dx_[0] is a solver member that computeDirection reads and never writes (x0 is fixed: it stays at the zero
   * the reference's allocateData gave it, src/ddp.cpp:467); a ctx zero-initialises it at create, here the
   * caller's buffer holds anything: clear row 0 of every problem (strided memset, no allocation) */
  e = cudaMemset2DAsync(dx, (size_t)(s.T + 1) * s.nx * sizeof(double), 0, (size_t)s.nx * sizeof(double), (size_t)s.B,
                        (cudaStream_t)cuda_stream);
  if (e != cudaSuccess) return fail(MIMGPU_ERR_CUDA, std::string("cudaMemset2DAsync(dx_0): ") + cudaGetErrorString(e));
  const int rc = small_sqp_direction_launch_x(a, kp, 0, (cudaStream_t)cuda_stream);
  if (rc != 0) return fail(MIMGPU_ERR_UNSUPPORTED, "knot-record kernel not launched (rc " + std::to_string(rc) + ")");
  e = cudaGetLastError();
  if (e != cudaSuccess) return fail(MIMGPU_ERR_CUDA, std::string("k_sqp_direction_small_gj launch: ") + cudaGetErrorString(e));
  return MIMGPU_OK;
}
