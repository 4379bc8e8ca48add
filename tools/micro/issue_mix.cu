// Micro-benchmark: how do FP64-pipe instructions (DFMA/DADD/DMUL, one warp instruction per 2 cycles per SM sub-partition)
// share the issue port with other instructions on sm_100a?  Each thread runs NCH independent chains of DFMA and, between
// them, K independent integer (LOP3/IADD) or FP32 (FFMA) instructions per DFMA.  If an FP64 instruction held the issue
// port for both of its pipe cycles, time would grow as 2F + O; if the port is free in the second cycle, as max(2F, F + O).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o build/micro/issue_mix tools/micro/issue_mix.cu && build/micro/issue_mix
#include <cstdio>
#include <cuda_runtime.h>

template <int K, int KIND>
__global__ void __launch_bounds__(256) k_mix(double* out, unsigned* iout, int iters, double a, double b, unsigned m) {
  double x[8];
  unsigned u[8];
  float f[8];
#pragma unroll
  for (int c = 0; c < 8; ++c) { x[c] = a + c + threadIdx.x; u[c] = m + c * 77u + threadIdx.x; f[c] = (float)(c + threadIdx.x); }
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int c = 0; c < 8; ++c) {
      asm volatile("fma.rn.f64 %0, %0, %1, %2;" : "+d"(x[c]) : "d"(a), "d"(b));
#pragma unroll
      for (int k = 0; k < K; ++k) {
        if (KIND == 0) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(u[(c + k) & 7]) : "r"(m), "r"(u[(c + k + 1) & 7]));
        else asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(f[(c + k) & 7]) : "f"((float)a), "f"((float)b));
      }
    }
  }
  double s = 0; unsigned t = 0; float g = 0;
#pragma unroll
  for (int c = 0; c < 8; ++c) { s += x[c]; t ^= u[c]; g += f[c]; }
  out[blockIdx.x * blockDim.x + threadIdx.x] = s + g;
  iout[blockIdx.x * blockDim.x + threadIdx.x] = t;
}

template <int K, int KIND>
static void run(const char* what, double* out, unsigned* iout, int sms) {
  const int iters = 4096, blocks = sms * 8;
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  k_mix<K, KIND><<<blocks, 256>>>(out, iout, 64, 1.0000001, 1e-9, 0x9e3779b9u);
  cudaEventRecord(e0);
  k_mix<K, KIND><<<blocks, 256>>>(out, iout, iters, 1.0000001, 1e-9, 0x9e3779b9u);
  cudaEventRecord(e1);
  cudaEventSynchronize(e1);
  float ms;
  cudaEventElapsedTime(&ms, e0, e1);
  int khz = 0;
  cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
  // per SM sub-partition: 8 CTAs x 8 warps / 4 = 16 warps, each iters x 8 DFMA and iters x 8 x K others
  const double warps = 16.0, F = warps * iters * 8.0, O = F * K;
  const double cycles = ms * 1e-3 * khz * 1e3;
  printf("%-28s K=%d  %.3f ms  cycles/SMSP %.3e  F %.3e  O %.3e  cycles/(2F) %.3f  cycles/(2F+O) %.3f  cycles/(F+O) %.3f\n", what, K, ms,
         cycles, F, O, cycles / (2 * F), cycles / (2 * F + O), cycles / (F + O));
}

int main() {
  int sms = 0;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  double* out; unsigned* iout;
  cudaMalloc(&out, sizeof(double) * sms * 8 * 256);
  cudaMalloc(&iout, sizeof(unsigned) * sms * 8 * 256);
  run<0, 0>("DFMA only", out, iout, sms);
  run<1, 0>("DFMA + 1 LOP3", out, iout, sms);
  run<2, 0>("DFMA + 2 LOP3", out, iout, sms);
  run<3, 0>("DFMA + 3 LOP3", out, iout, sms);
  run<1, 1>("DFMA + 1 FFMA", out, iout, sms);
  run<2, 1>("DFMA + 2 FFMA", out, iout, sms);
  run<3, 1>("DFMA + 3 FFMA", out, iout, sms);
  return 0;
}
