/*
 * peaks.cu -- mimgpu_measure_peaks: the FP64 roofline denominators that
 * MEASURED_PEAKS.json lacks (it only holds an HBM copy and a bf16 GEMM figure).
 * Register-resident DFMA loop, mma.sync.m8n8k4.f64 (DMMA) loop and a streaming read.
 */
#include <cuda_runtime.h>

#include "../../include/mimgpu.h"

namespace {

constexpr int kIters = 4096;

__global__ void __launch_bounds__(256) k_dfma(double* out, double a, double b) {
  double x0 = threadIdx.x, x1 = x0 + 1, x2 = x0 + 2, x3 = x0 + 3, x4 = x0 + 4, x5 = x0 + 5, x6 = x0 + 6, x7 = x0 + 7;
#pragma unroll 8
  for (int i = 0; i < kIters; ++i) {
    x0 = fma(x0, a, b); x1 = fma(x1, a, b); x2 = fma(x2, a, b); x3 = fma(x3, a, b);
    x4 = fma(x4, a, b); x5 = fma(x5, a, b); x6 = fma(x6, a, b); x7 = fma(x7, a, b);
  }
  const double s = x0 + x1 + x2 + x3 + x4 + x5 + x6 + x7;
  if (s == 123.456) out[0] = s;
}

__global__ void __launch_bounds__(256) k_dmma(double* out, double a, double b) {
  double c0[2] = {0, 0}, c1[2] = {0, 0}, c2[2] = {0, 0}, c3[2] = {0, 0};
  const double fa = a + threadIdx.x * 1e-9, fb = b;
#pragma unroll 4
  for (int i = 0; i < kIters; ++i) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                 : "+d"(c0[0]), "+d"(c0[1]) : "d"(fa), "d"(fb));
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                 : "+d"(c1[0]), "+d"(c1[1]) : "d"(fa), "d"(fb));
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                 : "+d"(c2[0]), "+d"(c2[1]) : "d"(fa), "d"(fb));
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                 : "+d"(c3[0]), "+d"(c3[1]) : "d"(fa), "d"(fb));
  }
  const double s = c0[0] + c0[1] + c1[0] + c1[1] + c2[0] + c2[1] + c3[0] + c3[1];
  if (s == 123.456) out[0] = s;
}

__global__ void __launch_bounds__(256) k_read(const double2* __restrict__ p, size_t n, double* out) {
  double acc = 0.0;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    const double2 v = p[i];
    acc += v.x + v.y;
  }
  if (acc == 123.456) out[0] = acc;
}

template <typename F>
float best_ms(F launch, int reps) {
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  float best = 1e30f;
  for (int r = 0; r < reps; ++r) {
    cudaEventRecord(e0);
    launch();
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    if (r > 0 && ms < best) best = ms;
  }
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  return best;
}

}  // namespace

extern "C" int mimgpu_measure_peaks(int device, double* dfma_tflops, double* dmma_tflops, double* hbm_read_gbs) {
  if (cudaSetDevice(device) != cudaSuccess) return MIMGPU_ERR_NODEVICE;
  cudaDeviceProp prop;
  if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) return MIMGPU_ERR_CUDA;
  const int sms = prop.multiProcessorCount;
  double* out = nullptr;
  if (cudaMalloc(&out, 64) != cudaSuccess) return MIMGPU_ERR_CUDA;
  const int blocks = sms * 8, threads = 256;
  if (dfma_tflops) {
    const float ms = best_ms([&] { k_dfma<<<blocks, threads>>>(out, 1.0000001, 1e-9); }, 6);
    *dfma_tflops = 2.0 * 8 * kIters * (double)blocks * threads / (ms * 1e-3) / 1e12;
  }
  if (dmma_tflops) {
    const float ms = best_ms([&] { k_dmma<<<blocks, threads>>>(out, 1.0000001, 1e-9); }, 6);
    *dmma_tflops = 2.0 * 8 * 8 * 4 * 4.0 * kIters * (double)blocks * (threads / 32) / (ms * 1e-3) / 1e12;
  }
  if (hbm_read_gbs) {
    const size_t bytes = (size_t)2 << 30;
    double2* buf = nullptr;
    if (cudaMalloc(&buf, bytes) != cudaSuccess) { cudaFree(out); return MIMGPU_ERR_CUDA; }
    cudaMemset(buf, 0, bytes);
    const float ms = best_ms([&] { k_read<<<sms * 16, 256>>>(buf, bytes / sizeof(double2), out); }, 6);
    *hbm_read_gbs = (double)bytes / (ms * 1e-3) / 1e9;
    cudaFree(buf);
  }
  cudaFree(out);
  return cudaDeviceSynchronize() == cudaSuccess ? MIMGPU_OK : MIMGPU_ERR_CUDA;
}
