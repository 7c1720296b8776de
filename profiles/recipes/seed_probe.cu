// accuracy of the fp64 MUFU seeds and of short refinements (relative error, max over 2^22 operands)
#include <cstdio>
#include <cmath>
#include <cuda_runtime.h>
__global__ void k(const double* b, double* out, int n) {
  int i = blockIdx.x * blockDim.x + threadIdx.x; if (i >= n) return;
  double v = b[i], r, y;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(v));
  asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(v));
  double e = fma(-v, r, 1.0);
  double r1 = fma(r, e, r);                       // 2nd order
  double e2 = fma(e, e, e); double r2 = fma(r, e2, r);  // 3rd order
  double e3 = fma(-v, r2, 1.0); double r3 = fma(r2, e3, r2);
  double t = y * y, f = fma(v, -t, 1.0), p = fma(f, 0.375, 0.5), s = y * f, y1 = fma(p, s, y);
  double y1b = fma(0.5 * f, y, y);  // 2nd order
  out[i] = r; out[n + i] = r1; out[2 * n + i] = r2; out[3 * n + i] = r3; out[4 * n + i] = y; out[5 * n + i] = y1; out[6*n+i] = y1b;
}
int main() {
  const int n = 1 << 22; double* hb = new double[n]; double* ho = new double[7 * n];
  unsigned long long s = 88172645463325252ull;
  for (int i = 0; i < n; ++i) { s ^= s << 13; s ^= s >> 7; s ^= s << 17; double u = (s >> 11) * (1.0 / 9007199254740992.0); hb[i] = ldexp(1.0 + u, (int)(s % 41) - 20); }
  double *b, *o; cudaMalloc(&b, n * 8); cudaMalloc(&o, 7 * n * 8);
  cudaMemcpy(b, hb, n * 8, cudaMemcpyHostToDevice);
  k<<<n / 256, 256>>>(b, o, n); cudaMemcpy(ho, o, 7 * n * 8, cudaMemcpyDeviceToHost);
  const char* nm[7] = {"rcp seed", "rcp 2nd-order", "rcp 3rd-order", "rcp 3rd+Newton", "rsqrt seed", "rsqrt 3rd-order", "rsqrt 2nd-order"};
  for (int q = 0; q < 7; ++q) { long double mx = 0; for (int i = 0; i < n; ++i) { long double ex = q < 4 ? 1.0L / hb[i] : 1.0L / sqrtl((long double)hb[i]); long double e = fabsl((ho[q * n + i] - ex) / ex); if (e > mx) mx = e; } printf("%-18s max rel err = %.3Le  (2^%.1f)\n", nm[q], mx, (double)log2l(mx)); }
  return 0;
}
