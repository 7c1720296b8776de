#define OLRT_NS olrt_strict
#include "trace_fwd.cuh"
using namespace olrt_strict;
template<int KIND>
__global__ void one(const double* in, double* out, const int32_t* ops, const double* presc, const double* der) {
  __shared__ double s_presc[64]; __shared__ double s_der[8]; __shared__ int32_t s_ops[32];
  if (threadIdx.x < 64) s_presc[threadIdx.x] = presc[threadIdx.x];
  if (threadIdx.x < 8) s_der[threadIdx.x] = der[threadIdx.x];
  if (threadIdx.x < 32) s_ops[threadIdx.x] = ops[threadIdx.x];
  __syncthreads();
  int i = blockIdx.x*blockDim.x+threadIdx.x; int n = gridDim.x*blockDim.x;
  RayState<double> r; r.x=in[i]; r.y=in[i+n]; r.z=in[i+2*n]; r.L=in[i+3*n]; r.M=in[i+4*n]; r.N=in[i+5*n]; r.i=in[i+6*n]; r.opd=0;
  SurfParams<double> sp; load_surface<double, false>(sp, s_ops, s_presc, s_der, 0, 1, 0);
  uint32_t st=0; double L0,M0,N0;
  surface_step<double, KIND, MM_FAST>(r, sp, 0, 1, 0.5, 12.5, true, nullptr, st, L0, M0, N0);
  out[i]=r.x; out[i+n]=r.y; out[i+2*n]=r.z; out[i+3*n]=r.L; out[i+4*n]=r.M; out[i+5*n]=r.N; out[i+6*n]=r.i; out[i+7*n]=r.opd;
}
template __global__ void one<KIND_CONIC>(const double*, double*, const int32_t*, const double*, const double*);
template __global__ void one<KIND_PLANE>(const double*, double*, const int32_t*, const double*, const double*);
template __global__ void one<KIND_EVEN>(const double*, double*, const int32_t*, const double*, const double*);
