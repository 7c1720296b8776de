// Strict build: compiled with -fmad=false so that every + - * / sqrt rounds exactly where the
// reference's NumPy float64 expression rounds (bit-identical vignetting masks).
#define OLRT_NS olrt_strict
#include "trace_launch.cuh"
#include "olrt_internal.h"

namespace olrt_strict {
int fwd_f64(const olrt_sys* sys, const double* presc, const int32_t* ops, long long n_rays,
            const double* const in[8], const double* opd_in, double* const out[8], double* record,
            double* dir0, unsigned long long* nr_resid, uint32_t* status, uint32_t* fixup, cudaStream_t stream) {
  return launch_fwd<double>(sys, presc, ops, n_rays, in, opd_in, out, record, dir0, nr_resid, status, fixup, stream);
}

__global__ void selftest_math_kernel(const double* a, const double* b, double* q, double* r, long long n) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    q[i] = m_div<MM_GUARD>(a[i], b[i]);
    r[i] = m_sqrt<MM_GUARD>(a[i]);
  }
}
int selftest_math(const double* a, const double* b, double* q, double* r, long long n, cudaStream_t stream) {
  selftest_math_kernel<<<sm_count() * 8, 256, 0, stream>>>(a, b, q, r, n);
  return cudaGetLastError() == cudaSuccess ? OLRT_OK : OLRT_E_LAUNCH;
}
}  // namespace olrt_strict
