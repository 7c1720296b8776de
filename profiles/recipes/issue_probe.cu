// Does a non-FP64 instruction issue "for free" in the shadow of an FP64 instruction on sm_100a?
// Each thread runs 8 independent DFMA chains; variant K adds K independent integer ops (IMAD/LOP3)
// per DFMA. Reports cycles per DFMA warp-instruction per SMSP for a few resident-warp counts.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /root/repo/scratch/issue_probe scratch/issue_probe.cu
#include <cstdio>
#include <cuda_runtime.h>

template <int K>
__global__ void probe(double* out, int* iout, int iters, double a, double b, int m) {
  double x0 = threadIdx.x, x1 = x0 + 1, x2 = x0 + 2, x3 = x0 + 3, x4 = x0 + 4, x5 = x0 + 5, x6 = x0 + 6, x7 = x0 + 7;
  int i0 = threadIdx.x, i1 = i0 + 1, i2 = i0 + 2, i3 = i0 + 3, i4 = i0 + 5, i5 = i0 + 7, i6 = i0 + 11, i7 = i0 + 13;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      x0 = fma(x0, a, b); if (K > 0) i0 = i0 * m + 3; if (K > 1) i4 = (i4 ^ m) + it; if (K > 2) i0 = i0 * m + i4;
      x1 = fma(x1, a, b); if (K > 0) i1 = i1 * m + 3; if (K > 1) i5 = (i5 ^ m) + it; if (K > 2) i1 = i1 * m + i5;
      x2 = fma(x2, a, b); if (K > 0) i2 = i2 * m + 3; if (K > 1) i6 = (i6 ^ m) + it; if (K > 2) i2 = i2 * m + i6;
      x3 = fma(x3, a, b); if (K > 0) i3 = i3 * m + 3; if (K > 1) i7 = (i7 ^ m) + it; if (K > 2) i3 = i3 * m + i7;
      x4 = fma(x4, a, b); if (K > 0) i0 = i0 * m + 5; if (K > 1) i4 = (i4 ^ m) + it; if (K > 2) i0 = i0 * m + i4;
      x5 = fma(x5, a, b); if (K > 0) i1 = i1 * m + 5; if (K > 1) i5 = (i5 ^ m) + it; if (K > 2) i1 = i1 * m + i5;
      x6 = fma(x6, a, b); if (K > 0) i2 = i2 * m + 5; if (K > 1) i6 = (i6 ^ m) + it; if (K > 2) i2 = i2 * m + i6;
      x7 = fma(x7, a, b); if (K > 0) i3 = i3 * m + 5; if (K > 1) i7 = (i7 ^ m) + it; if (K > 2) i3 = i3 * m + i7;
    }
  }
  out[blockIdx.x * blockDim.x + threadIdx.x] = x0 + x1 + x2 + x3 + x4 + x5 + x6 + x7;
  iout[blockIdx.x * blockDim.x + threadIdx.x] = i0 + i1 + i2 + i3 + i4 + i5 + i6 + i7;
}

// dependent-chain latency: ONE chain per thread, one warp per SMSP
__global__ void chain(double* out, int iters, double a, double b) {
  double x = threadIdx.x;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int u = 0; u < 32; ++u) x = fma(x, a, b);
  }
  out[blockIdx.x * blockDim.x + threadIdx.x] = x;
}

template <int K>
void run(int warps_per_smsp, double* out, int* iout, double ghz, int sms) {
  const int iters = 4096;
  const int block = 128 * warps_per_smsp;  // 4 SMSPs
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  probe<K><<<sms, block>>>(out, iout, 64, 1.0000001, 1e-9, 3);
  cudaEventRecord(e0);
  probe<K><<<sms, block>>>(out, iout, iters, 1.0000001, 1e-9, 3);
  cudaEventRecord(e1); cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  const double cyc = ms * 1e-3 * ghz * 1e9;
  const double dfma_per_smsp = (double)iters * 32 * warps_per_smsp;
  printf("K=%d int/DFMA  warps/SMSP=%d  %.3f ms  cycles per DFMA warp-instr = %.3f\n", K, warps_per_smsp, ms, cyc / dfma_per_smsp);
}

int main() {
  cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
  int khz = 0; cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
  const double ghz = khz * 1e-6;
  const int sms = p.multiProcessorCount;
  printf("%s, %d SMs, %.3f GHz nominal\n", p.name, sms, ghz);
  double* out; int* iout;
  cudaMalloc(&out, sizeof(double) * sms * 1024); cudaMalloc(&iout, sizeof(int) * sms * 1024);
  for (int w : {1, 2, 4, 8}) { run<0>(w, out, iout, ghz, sms); run<1>(w, out, iout, ghz, sms); run<2>(w, out, iout, ghz, sms); run<3>(w, out, iout, ghz, sms); }
  // latency
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  chain<<<sms, 128>>>(out, 64, 1.0000001, 1e-9);
  cudaEventRecord(e0);
  chain<<<sms, 128>>>(out, 8192, 1.0000001, 1e-9);
  cudaEventRecord(e1); cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  printf("dependent DFMA latency = %.2f cycles\n", ms * 1e-3 * ghz * 1e9 / (8192.0 * 32));
  return 0;
}
