// olrt_api.cu -- the extern "C" boundary (include/olrt.h), the spot epilogue and the FMA-peak probe.
#include <cuda_runtime.h>
#include <math.h>
#include <mutex>
#include "olrt_internal.h"

// ---------------------------------------------------------------------------------------------
// argument checks shared by fwd / bwd
// ---------------------------------------------------------------------------------------------
static int check_sys(const olrt_sys* sys, const void* presc, const void* ops, long long n_rays) {
  if (!sys || !presc || !ops || n_rays < 0) return OLRT_E_ARG;
  if (sys->n_surfaces < 1 || sys->n_surfaces > OLRT_MAX_SURFACES) return OLRT_E_LIMIT;
  if (sys->n_wavelengths < 1 || sys->n_wavelengths > OLRT_MAX_WAVELENGTHS) return OLRT_E_LIMIT;
  if (sys->n_presc < 1 || sys->presc_stride < 1) return OLRT_E_ARG;
  if (sys->ops_len < sys->n_surfaces * OLRT_OPS_STRIDE) return OLRT_E_ARG;
  if (sys->n_presc > 1) {
    if (sys->rays_per_presc < 1) return OLRT_E_ARG;
    if ((long long)sys->n_presc * sys->rays_per_presc < n_rays) return OLRT_E_ARG;
  }
  return OLRT_OK;
}

template <typename R>
static int check_rays(const olrt_sys* sys, const R* const in[8], R* const out[8]) {
  if (!in || !out) return OLRT_E_ARG;
  for (int q = 0; q < 7; ++q)
    if (!in[q]) return OLRT_E_ARG;
  if (sys->n_wavelengths > 1 && !in[7]) return OLRT_E_ARG;
  for (int q = 0; q < 8; ++q)
    if (!out[q]) return OLRT_E_ARG;
  return OLRT_OK;
}

extern "C" int olrt_version(void) { return OLRT_ABI_VERSION; }

extern "C" const char* olrt_build_info(void) {
  return "libolrt sm_100a; fp64 strict(-fmad=false)+fast, fp32 fast; built " __DATE__ " " __TIME__;
}

extern "C" int olrt_trace_fwd_f64(const olrt_sys* sys, const double* presc, const int32_t* ops,
                                  int64_t n_rays, const double* const in[8], const double* opd_in,
                                  double* const out[8], double* record, double* dir0,
                                  unsigned long long* nr_resid, uint32_t* status, uint32_t* fixup,
                                  int flags, void* stream) {
  int rc = check_sys(sys, presc, ops, n_rays);
  if (rc) return rc;
  if (n_rays == 0) return OLRT_OK;  // an empty batch has no arrays to check (their pointers may be null)
  rc = check_rays<double>(sys, in, out);
  if (rc) return rc;
  if (sys->n_record > 0 && !record) return OLRT_E_ARG;
  if (!fixup || n_rays > 0xffffffffLL) return OLRT_E_ARG;
  cudaStream_t st = (cudaStream_t)stream;
  if ((flags & OLRT_FWD_CKPT_BLOCKED) && !(flags & OLRT_FWD_FAST)) return OLRT_E_UNSUPPORTED;
  if (flags & OLRT_FWD_FAST)
    return olrt_fast::fwd_f64(sys, presc, ops, n_rays, in, opd_in, out, record, dir0, nr_resid, status, fixup, flags, st);
  return olrt_strict::fwd_f64(sys, presc, ops, n_rays, in, opd_in, out, record, dir0, nr_resid, status, fixup, 0, st);
}

extern "C" int olrt_trace_fwd_f32(const olrt_sys* sys, const float* presc, const int32_t* ops,
                                  int64_t n_rays, const float* const in[8], const float* opd_in,
                                  float* const out[8], float* record, float* dir0,
                                  unsigned long long* nr_resid, uint32_t* status, uint32_t* fixup,
                                  int flags, void* stream) {
  int rc = check_sys(sys, presc, ops, n_rays);
  if (rc) return rc;
  if (n_rays == 0) return OLRT_OK;
  rc = check_rays<float>(sys, in, out);
  if (rc) return rc;
  if (sys->n_record > 0 && !record) return OLRT_E_ARG;
  return olrt_fast::fwd_f32(sys, presc, ops, n_rays, in, opd_in, out, record, dir0, nr_resid, status, fixup, flags,
                            (cudaStream_t)stream);
}

// ---------------------------------------------------------------------------------------------
// Forward trace that allocates nothing (include/olrt.h: olrt_trace_fwd_noalloc_f64). The re-trace list the caller of
// olrt_trace_fwd_f64 passes as `fixup` lives here, in the library's own device image: one copy per GPU that loads
// libolrt.so, held until the process exits, shared by every call (calls that overlap on one GPU race on it).
// In HOST code `g_noalloc_fixup` is only a stand-in for the symbol -- the device address comes from
// cudaGetSymbolAddress, once per device, with that device current.
// ---------------------------------------------------------------------------------------------
static __device__ __align__(256) uint32_t g_noalloc_fixup[1 + OLRT_NOALLOC_MAX_RAYS];

#define OLRT_NOALLOC_MAX_DEVICES 64
static uint32_t* g_noalloc_addr[OLRT_NOALLOC_MAX_DEVICES];  // per-device address of g_noalloc_fixup, 0 = not asked yet
static std::mutex g_noalloc_mutex;

static int noalloc_fixup_address(uint32_t** ptr) {
  int dev = -1;
  if (cudaGetDevice(&dev) != cudaSuccess) return OLRT_E_LAUNCH;
  if (dev < 0 || dev >= OLRT_NOALLOC_MAX_DEVICES) return OLRT_E_LIMIT;
  std::lock_guard<std::mutex> lock(g_noalloc_mutex);
  if (g_noalloc_addr[dev] == nullptr) {
    uint32_t* p = nullptr;
    if (cudaGetSymbolAddress((void**)&p, g_noalloc_fixup) != cudaSuccess || p == nullptr) {
      (void)cudaGetLastError();
      return OLRT_E_LAUNCH;
    }
    g_noalloc_addr[dev] = p;
  }
  *ptr = g_noalloc_addr[dev];
  return OLRT_OK;
}

extern "C" int olrt_trace_fwd_noalloc_f64(const olrt_sys* sys, const double* presc, const int32_t* ops,
                                          int64_t n_rays, const double* const in[8], const double* opd_in,
                                          double* const out[8], double* record, double* dir0,
                                          unsigned long long* nr_resid, uint32_t* status, int flags, void* stream) {
  if (n_rays > OLRT_NOALLOC_MAX_RAYS) return OLRT_E_LIMIT;  // the list holds one entry per ray
  int rc = check_sys(sys, presc, ops, n_rays);
  if (rc) return rc;
  if (n_rays == 0) return OLRT_OK;
  rc = check_rays<double>(sys, in, out);
  if (rc) return rc;
  if (sys->n_record > 0 && !record) return OLRT_E_ARG;
  if ((flags & OLRT_FWD_CKPT_BLOCKED) && !(flags & OLRT_FWD_FAST)) return OLRT_E_UNSUPPORTED;
  uint32_t* fixup = nullptr;
  rc = noalloc_fixup_address(&fixup);
  if (rc) return rc;
  // the launcher clears the list's counter (word 0) on `stream` before the main kernel; entries past the counter
  // are never read, so the previous call's contents need no other clearing
  cudaStream_t st = (cudaStream_t)stream;
  if (flags & OLRT_FWD_FAST)
    return olrt_fast::fwd_f64(sys, presc, ops, n_rays, in, opd_in, out, record, dir0, nr_resid, status, fixup, flags, st);
  return olrt_strict::fwd_f64(sys, presc, ops, n_rays, in, opd_in, out, record, dir0, nr_resid, status, fixup, 0, st);
}

template <typename R>
static int check_pupil(const olrt_sys* sys, const R* presc, const int32_t* ops, const olrt_pupil* q, R* const out[8],
                       const double* spot, const double* shift) {
  if (!q || !q->Px || !q->Py || !q->field_xy) return OLRT_E_ARG;
  if (q->n_fields < 1 || q->n_pupil < 1 || q->p_lo < 0 || q->p_len < 0 || q->p_lo + q->p_len > q->n_pupil)
    return OLRT_E_ARG;
  if (!sys || sys->n_presc < 1) return OLRT_E_ARG;
  olrt_sys one = *sys;   // in pupil mode every prescription traces the same pupil range: rays_per_presc is implied
  one.rays_per_presc = q->p_len * q->n_fields;
  const int rc = check_sys(&one, presc, ops, q->p_len * q->n_fields * sys->n_presc);
  if (rc) return rc;
  if (out) for (int k = 0; k < 8; ++k) if (!out[k]) return OLRT_E_ARG;
  if (spot && !shift) return OLRT_E_ARG;
  if (q->n_pupil * (long long)q->n_fields * sys->n_presc > 0xffffffffLL) return OLRT_E_LIMIT;
  if (!(sys->kernel_hint & OLRT_HINT_FASTKINDS) || sys->n_wavelengths != 1 || sys->n_record != 0)
    return OLRT_E_UNSUPPORTED;
  return OLRT_OK;
}

extern "C" int olrt_trace_pupil_f64(const olrt_sys* sys, const double* presc, const int32_t* ops,
                                    const olrt_pupil* pupil, double* const out[8], double* dir0,
                                    unsigned long long* nr_resid, uint32_t* status, uint32_t* fixup,
                                    double* spot_sums, const double* spot_shift, int spot_flags,
                                    const int32_t* group_k, unsigned long long* group_resid, int flags,
                                    void* stream) {
  const int rc = check_pupil<double>(sys, presc, ops, pupil, out, spot_sums, spot_shift);
  if (rc) return rc;
  if (!fixup) return OLRT_E_ARG;
  if (!(flags & OLRT_FWD_FAST) || (flags & OLRT_FWD_REPLAY)) return OLRT_E_UNSUPPORTED;
  if (pupil->p_len == 0) return OLRT_OK;
  return olrt_fast::pupil_f64(sys, presc, ops, pupil, out, dir0, nr_resid, status, fixup, spot_sums, spot_shift,
                              spot_flags, group_k, group_resid, flags, (cudaStream_t)stream);
}

extern "C" int olrt_trace_pupil_f32(const olrt_sys* sys, const float* presc, const int32_t* ops,
                                    const olrt_pupil* pupil, float* const out[8], float* dir0,
                                    unsigned long long* nr_resid, uint32_t* status, uint32_t* fixup,
                                    double* spot_sums, const double* spot_shift, int spot_flags,
                                    const int32_t* group_k, unsigned long long* group_resid, int flags,
                                    void* stream) {
  const int rc = check_pupil<float>(sys, presc, ops, pupil, out, spot_sums, spot_shift);
  if (rc) return rc;
  if (flags & OLRT_FWD_REPLAY) return OLRT_E_UNSUPPORTED;
  if (pupil->p_len == 0) return OLRT_OK;
  return olrt_fast::pupil_f32(sys, presc, ops, pupil, out, dir0, nr_resid, status, fixup, spot_sums, spot_shift,
                              spot_flags, group_k, group_resid, flags, (cudaStream_t)stream);
}

// ---------------------------------------------------------------------------------------------
// Spot epilogue. One CTA owns a contiguous slice of rays and walks the groups it overlaps;
// per group: register accumulation -> warp shuffles -> shared -> one set of atomics per CTA.
//   RayOperand.rms_spot_size        optiland/optimization/operand/ray.py:336-340
//   SpotDiagram.centroid/rms/geo    optiland/analysis/spot_diagram.py:317-357
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ double warp_sum(double v) {
  for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ double warp_maxd(double v) {
  for (int o = 16; o > 0; o >>= 1) v = fmax(v, __shfl_down_sync(0xffffffffu, v, o));
  return v;
}

template <typename R>
__global__ void __launch_bounds__(256) spot_sums_kernel(const R* __restrict__ x, const R* __restrict__ y,
                                                        const R* __restrict__ inten, long long n,
                                                        long long rpg, int n_groups,
                                                        const double* __restrict__ shift, int flags,
                                                        double* sums) {
  __shared__ double s_part[8][6];
  const long long per = (n + gridDim.x - 1) / gridDim.x;
  const long long lo = (long long)blockIdx.x * per;
  long long hi = lo + per;
  if (hi > n) hi = n;
  if (lo >= hi) return;
  const int g_lo = (int)(lo / rpg), g_hi = (int)((hi - 1) / rpg);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int g = g_lo; g <= g_hi && g < n_groups; ++g) {
    long long a = (long long)g * rpg, b = a + rpg;
    double sx0, sy0;
    if (shift) { sx0 = shift[2 * g]; sy0 = shift[2 * g + 1]; }
    else {
      sx0 = (double)x[a]; sy0 = (double)y[a];
      if (!isfinite(sx0)) sx0 = 0.0;
      if (!isfinite(sy0)) sy0 = 0.0;
    }
    if (a < lo) a = lo;
    if (b > hi) b = hi;
    double cnt = 0, sx = 0, sy = 0, sxx = 0, syy = 0, mx = 0, si = 0;
    bool mnan = false;
    for (long long q = a + threadIdx.x; q < b; q += blockDim.x) {
      const double iv = (double)inten[q];
      if ((flags & OLRT_SPOT_VALID_ONLY) && !(iv > 0.0)) continue;
      const double dx = (double)x[q] - sx0, dy = (double)y[q] - sy0;
      cnt += 1.0; sx += dx; sy += dy; sxx += dx * dx; syy += dy * dy; si += iv;
      const double r2 = dx * dx + dy * dy;
      if (r2 != r2) mnan = true;
      mx = fmax(mx, r2);
    }
    if (mnan) mx = __longlong_as_double(0x7ff8000000000000LL);
    double v[6] = {cnt, sx, sy, sxx, syy, si};
#pragma unroll
    for (int k = 0; k < 6; ++k) v[k] = warp_sum(v[k]);
    // NaN-propagating max: fmax drops NaN, so carry the flag through the count of NaN lanes
    const unsigned nanmask = __ballot_sync(0xffffffffu, mnan);
    double wm = warp_maxd(mnan ? 0.0 : mx);
    __syncthreads();
    if (lane == 0) {
#pragma unroll
      for (int k = 0; k < 6; ++k) s_part[warp][k] = v[k];
    }
    __shared__ double s_max[8];
    __shared__ unsigned s_nan[8];
    if (lane == 0) { s_max[warp] = wm; s_nan[warp] = nanmask; }
    __syncthreads();
    if (threadIdx.x == 0) {
      const int nw = blockDim.x >> 5;
      double t[6] = {0, 0, 0, 0, 0, 0}, m = 0;
      unsigned anynan = 0;
      for (int w = 0; w < nw; ++w) {
        for (int k = 0; k < 6; ++k) t[k] += s_part[w][k];
        m = fmax(m, s_max[w]);
        anynan |= s_nan[w];
      }
      double* o = sums + (long long)g * OLRT_SPOT_SUMS;
      atomicAdd(o + 0, t[0]); atomicAdd(o + 1, t[1]); atomicAdd(o + 2, t[2]);
      atomicAdd(o + 3, t[3]); atomicAdd(o + 4, t[4]); atomicAdd(o + 6, t[5]);
      unsigned long long key = anynan ? 0x7ff8000000000000ULL : (unsigned long long)__double_as_longlong(m);
      atomicMax(reinterpret_cast<unsigned long long*>(o + 5), key);
      o[8] = sx0; o[9] = sy0;  // identical from every CTA
    }
  }
}

template <typename R>
static int launch_spot(const R* x, const R* y, const R* inten, long long n, long long rpg, int n_groups,
                       const double* shift, int flags, double* sums, cudaStream_t st) {
  if (!x || !y || !inten || !sums || n < 0 || n_groups < 1) return OLRT_E_ARG;
  if (rpg < 1) rpg = n > 0 ? n : 1;
  if (n == 0) return OLRT_OK;
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  long long grid = (long long)sms * 8;
  const long long max_grid = (n + 255) / 256;
  if (grid > max_grid) grid = max_grid;
  spot_sums_kernel<R><<<(unsigned)grid, 256, 0, st>>>(x, y, inten, n, rpg, n_groups, shift, flags, sums);
  return cudaGetLastError() == cudaSuccess ? OLRT_OK : OLRT_E_LAUNCH;
}

extern "C" int olrt_spot_sums_f64(const double* x, const double* y, const double* intensity,
                                  int64_t n_rays, int64_t rays_per_group, int32_t n_groups,
                                  const double* shift, int flags, double* sums, void* stream) {
  return launch_spot<double>(x, y, intensity, n_rays, rays_per_group, n_groups, shift, flags, sums,
                             (cudaStream_t)stream);
}
extern "C" int olrt_spot_sums_f32(const float* x, const float* y, const float* intensity,
                                  int64_t n_rays, int64_t rays_per_group, int32_t n_groups,
                                  const double* shift, int flags, double* sums, void* stream) {
  return launch_spot<float>(x, y, intensity, n_rays, rays_per_group, n_groups, shift, flags, sums,
                            (cudaStream_t)stream);
}

// ---------------------------------------------------------------------------------------------
// FMA peak probe: 8 independent chains per thread, 2 flops per FMA.
// flops = blocks * threads * iters * 8 * 2
// ---------------------------------------------------------------------------------------------
template <typename R>
__global__ void fma_peak_kernel(R* sink, long long iters) {
  R a0 = R(threadIdx.x) * R(1e-3), a1 = a0 + R(1), a2 = a0 + R(2), a3 = a0 + R(3);
  R a4 = a0 + R(4), a5 = a0 + R(5), a6 = a0 + R(6), a7 = a0 + R(7);
  const R m = R(0.999999), c = R(1e-7);
  for (long long q = 0; q < iters; ++q) {
    a0 = fma(a0, m, c); a1 = fma(a1, m, c); a2 = fma(a2, m, c); a3 = fma(a3, m, c);
    a4 = fma(a4, m, c); a5 = fma(a5, m, c); a6 = fma(a6, m, c); a7 = fma(a7, m, c);
  }
  const R s = a0 + a1 + a2 + a3 + a4 + a5 + a6 + a7;
  if (s == R(-1)) sink[0] = s;  // never true; keeps the chains alive
}

extern "C" int olrt_fma_peak_f64(double* sink, int64_t iters, int blocks, int threads, void* stream) {
  if (!sink || iters < 1 || blocks < 1 || threads < 1 || threads > 1024) return OLRT_E_ARG;
  fma_peak_kernel<double><<<blocks, threads, 0, (cudaStream_t)stream>>>(sink, iters);
  return cudaGetLastError() == cudaSuccess ? OLRT_OK : OLRT_E_LAUNCH;
}
extern "C" int olrt_fma_peak_f32(float* sink, int64_t iters, int blocks, int threads, void* stream) {
  if (!sink || iters < 1 || blocks < 1 || threads < 1 || threads > 1024) return OLRT_E_ARG;
  fma_peak_kernel<float><<<blocks, threads, 0, (cudaStream_t)stream>>>(sink, iters);
  return cudaGetLastError() == cudaSuccess ? OLRT_OK : OLRT_E_LAUNCH;
}

static int check_generate(const olrt_pupil* q, void* const out[8]) {
  if (!q || !q->Px || !q->Py || !q->field_xy || !out || q->n_pupil < 0 || q->n_fields < 1) return OLRT_E_ARG;
  for (int k = 0; k < 7; ++k)
    if (!out[k]) return OLRT_E_ARG;
  return OLRT_OK;
}
extern "C" int olrt_generate_rays_f64(const olrt_pupil* pupil, double* const out[8], void* stream) {
  int rc = check_generate(pupil, (void* const*)out);
  if (rc) return rc;
  if (pupil->n_pupil == 0) return OLRT_OK;
  return olrt_strict::generate_f64(pupil, out, (cudaStream_t)stream);
}
extern "C" int olrt_generate_rays_f32(const olrt_pupil* pupil, float* const out[8], void* stream) {
  int rc = check_generate(pupil, (void* const*)out);
  if (rc) return rc;
  if (pupil->n_pupil == 0) return OLRT_OK;
  return olrt_strict::generate_f32(pupil, out, (cudaStream_t)stream);
}

extern "C" int olrt_selftest_math_f64(const double* a, const double* b, double* quot, double* root,
                                      int64_t n, void* stream) {
  if (!a || !b || !quot || !root || n < 0) return OLRT_E_ARG;
  if (n == 0) return OLRT_OK;
  return olrt_strict::selftest_math(a, b, quot, root, n, (cudaStream_t)stream);
}

// ---------------------------------------------------------------------------------------------
// backward (adjoint) entry points
// ---------------------------------------------------------------------------------------------
template <typename R>
static int check_bwd(const olrt_sys* sys, const R* presc, const int32_t* ops, long long n_rays,
                     const R* const in[8], const R* record, R* gpresc, int flags) {
  int rc = check_sys(sys, presc, ops, n_rays);
  if (rc) return rc;
  if (!in || !record || !gpresc) return OLRT_E_ARG;
  for (int q = 0; q < 7; ++q)
    if (!in[q]) return OLRT_E_ARG;
  if (sys->n_wavelengths > 1 && !in[7]) return OLRT_E_ARG;
  if (!(flags & OLRT_BWD_CKPT_BLOCKED) && sys->n_record != sys->n_surfaces)
    return OLRT_E_UNSUPPORTED;  // every surface must be recorded (or tile-blocked checkpoints passed)
  return OLRT_OK;
}

extern "C" int olrt_trace_bwd_f64(const olrt_sys* sys, const double* presc, const int32_t* ops,
                                  int64_t n_rays, const double* const in[8], const double* record,
                                  const double* const gout[8], const double* grecord,
                                  double* const gin[8], double* gpresc, uint32_t* status, int flags,
                                  void* stream) {
  int rc = check_bwd<double>(sys, presc, ops, n_rays, in, record, gpresc, flags);
  if (rc) return rc;
  if (n_rays == 0) return OLRT_OK;
  return olrt_fast::bwd_f64(sys, presc, ops, n_rays, in, record, gout, grecord, gin, gpresc, status, flags,
                            (cudaStream_t)stream);
}
extern "C" int olrt_trace_bwd_f32(const olrt_sys* sys, const float* presc, const int32_t* ops,
                                  int64_t n_rays, const float* const in[8], const float* record,
                                  const float* const gout[8], const float* grecord, float* const gin[8],
                                  float* gpresc, uint32_t* status, int flags, void* stream) {
  int rc = check_bwd<float>(sys, presc, ops, n_rays, in, record, gpresc, flags);
  if (rc) return rc;
  if (n_rays == 0) return OLRT_OK;
  return olrt_fast::bwd_f32(sys, presc, ops, n_rays, in, record, gout, grecord, gin, gpresc, status, flags,
                            (cudaStream_t)stream);
}

extern "C" int olrt_upload_pupil(double* Px_dev, double* Py_dev, const double* Px_host, const double* Py_host,
                                 int64_t n_pupil, int64_t ready_chunk, uint32_t* ready, const uint32_t* one_host,
                                 void* copy_stream) {
  if (!Px_dev || !Py_dev || !Px_host || !Py_host || !ready || !one_host || n_pupil < 0) return OLRT_E_ARG;
  if (ready_chunk <= 0 || ready_chunk % 256 != 0) return OLRT_E_ARG;
  cudaStream_t st = (cudaStream_t)copy_stream;
  int64_t c = 0;
  for (int64_t lo = 0; lo < n_pupil; lo += ready_chunk, ++c) {
    const int64_t len = (n_pupil - lo < ready_chunk) ? (n_pupil - lo) : ready_chunk;
    if (cudaMemcpyAsync(Px_dev + lo, Px_host + lo, (size_t)len * sizeof(double), cudaMemcpyHostToDevice, st) != cudaSuccess ||
        cudaMemcpyAsync(Py_dev + lo, Py_host + lo, (size_t)len * sizeof(double), cudaMemcpyHostToDevice, st) != cudaSuccess ||
        cudaMemcpyAsync(ready + c, one_host, sizeof(uint32_t), cudaMemcpyHostToDevice, st) != cudaSuccess)
      return OLRT_E_LAUNCH;
  }
  return OLRT_OK;
}

// ---------------------------------------------------------------------------------------------
// Random pupil on the device: RandomDistribution.generate_points (optiland/distribution.py:132-158) draws
//   r = rng.uniform(size=n); theta = rng.uniform(0, 2 pi, size=n); x = sqrt(r) cos(theta); y = sqrt(r) sin(theta)
// from numpy.random.default_rng(seed) = PCG64: a 128-bit LCG (state = state * MULT + inc) with the XSL-RR output
// rotr64(hi ^ lo, hi >> 58) taken AFTER the step, a double = (u64 >> 11) * 2^-53, uniform(a, b) = a + (b - a) d.
// The stream is random-access: thread t jumps to draw t in O(log t) 128-bit steps (LCG advance), then walks on with
// the one multiplier / increment pair of its fixed stride. r uses draws [0, n), theta draws [n, 2n): the host passes
// the generator's state before its first draw (SeedSequence hashing stays in NumPy, microseconds) and the same state
// advanced by n. The draws are NumPy's bit for bit; x and y differ from NumPy's by the last bit or two of cos / sin
// (libm vs CUDA), tests/test_gpu_pupil_random.py.
// ---------------------------------------------------------------------------------------------
typedef unsigned __int128 u128;
__host__ __device__ static inline u128 mk128(unsigned long long hi, unsigned long long lo) { return ((u128)hi << 64) | lo; }
#define OLRT_PCG_MULT mk128(0x2360ED051FC65DA4ULL, 0x4385DF649FCCF645ULL)

// (acc_mult, acc_plus) of `delta` LCG steps: state' = acc_mult * state + acc_plus
__host__ __device__ static inline void pcg_advance(u128 mult, u128 plus, unsigned long long delta, u128& am, u128& ap) {
  am = 1; ap = 0;
  while (delta > 0) {
    if (delta & 1) { am *= mult; ap = ap * mult + plus; }
    plus = (mult + 1) * plus;
    mult *= mult;
    delta >>= 1;
  }
}
__device__ __forceinline__ double pcg_double(u128 state) {
  const unsigned long long hi = (unsigned long long)(state >> 64), lo = (unsigned long long)state;
  const unsigned long long x = hi ^ lo;
  const unsigned rot = (unsigned)(hi >> 58);
  const unsigned long long o = (x >> rot) | (x << ((64u - rot) & 63u));
  return (double)(o >> 11) * (1.0 / 9007199254740992.0);
}

__global__ void __launch_bounds__(256) pupil_random_kernel(u128 s_r, u128 s_t, u128 inc, u128 a_stride, u128 c_stride,
                                                           long long n, double two_pi, double* Px, double* Py,
                                                           double* raw) {
  const long long T = (long long)gridDim.x * blockDim.x;
  long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  u128 am, ap;
  pcg_advance(OLRT_PCG_MULT, inc, (unsigned long long)i + 1ull, am, ap);   // the output follows the step: draw i = state i + 1
  u128 sr = am * s_r + ap, st = am * s_t + ap;
  for (; i < n; i += T) {
    const double r = pcg_double(sr);
    const double th = 0.0 + two_pi * pcg_double(st);   // Generator.uniform: low + (high - low) * d
    double sn, cs;
    sincos(th, &sn, &cs);
    const double q = sqrt(r);
    Px[i] = q * cs; Py[i] = q * sn;
    if (raw != nullptr) { raw[i] = r; raw[n + i] = th; }
    sr = a_stride * sr + c_stride; st = a_stride * st + c_stride;
  }
}

extern "C" int olrt_pupil_random_f64(const uint64_t state[2], const uint64_t inc[2], int64_t n, double* Px, double* Py,
                                     double* raw, void* stream) {
  if (!state || !inc || !Px || !Py || n < 0) return OLRT_E_ARG;
  if (n == 0) return OLRT_OK;
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  long long grid = (long long)sms * 8;
  const long long need = (n + 255) / 256;
  if (grid > need) grid = need;
  const u128 s0 = mk128(state[0], state[1]), ic = mk128(inc[0], inc[1]);
  u128 am, ap, as, cs;
  pcg_advance(OLRT_PCG_MULT, ic, (unsigned long long)n, am, ap);            // theta's draws start n steps on
  const u128 s_t = am * s0 + ap;
  pcg_advance(OLRT_PCG_MULT, ic, (unsigned long long)(grid * 256), as, cs);  // one stride of the grid
  pupil_random_kernel<<<(unsigned)grid, 256, 0, (cudaStream_t)stream>>>(s0, s_t, ic, as, cs, n, 2.0 * 3.141592653589793,
                                                                        Px, Py, raw);
  return cudaGetLastError() == cudaSuccess ? OLRT_OK : OLRT_E_LAUNCH;
}

// HexagonalDistribution.generate_points (optiland/distribution.py:189-220): the centre, then ring i = 1 .. R with 6 i
// points at radius linspace(0, 1, R + 1)[i] and angles linspace(0, 2 pi, 6 i + 1)[:-1] -- index arithmetic per point.
// NumPy's linspace is arange * (delta / div) with the last element set to the stop value.
__global__ void __launch_bounds__(256) pupil_hexapolar_kernel(int rings, long long n, double two_pi, double* Px,
                                                              double* Py) {
  const long long T = (long long)gridDim.x * blockDim.x;
  const double rstep = 1.0 / (double)rings;
  for (long long p = (long long)blockIdx.x * blockDim.x + threadIdx.x; p < n; p += T) {
    if (p == 0) { Px[0] = 0.0; Py[0] = 0.0; continue; }
    // ring i holds points 1 + 3 (i - 1) i .. 3 i (i + 1): i from the root of 3 i^2 + 3 i = p, corrected exactly
    long long i = (long long)ceil((sqrt(12.0 * (double)p - 3.0) - 3.0) / 6.0);
    if (i < 1) i = 1;
    while (3 * i * (i + 1) < p) ++i;
    while (i > 1 && 3 * (i - 1) * i >= p) --i;
    const long long k = p - 1 - 3 * (i - 1) * i;
    const double r = (i == rings) ? 1.0 : __dmul_rn((double)i, rstep);
    const double tstep = two_pi / (double)(6 * i);
    const double th = __dmul_rn((double)k, tstep);
    double sn, cs;
    sincos(th, &sn, &cs);
    Px[p] = __dmul_rn(r, cs); Py[p] = __dmul_rn(r, sn);
  }
}

extern "C" int olrt_pupil_hexapolar_f64(int32_t num_rings, double* Px, double* Py, void* stream) {
  if (!Px || !Py || num_rings < 0) return OLRT_E_ARG;
  const long long n = 1 + 3LL * num_rings * (num_rings + 1LL);
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  long long grid = (long long)sms * 8;
  const long long need = (n + 255) / 256;
  if (grid > need) grid = need;
  pupil_hexapolar_kernel<<<(unsigned)grid, 256, 0, (cudaStream_t)stream>>>(num_rings, n, 2.0 * 3.141592653589793, Px, Py);
  return cudaGetLastError() == cudaSuccess ? OLRT_OK : OLRT_E_LAUNCH;
}

// UniformDistribution.generate_points (optiland/distribution.py:174-185): x = linspace(-1, 1, n), the n x n grid
// (row j holds y = x[j]), masked to x^2 + y^2 <= 1, rows first. Pass 1 counts the kept points of every row, the caller
// turns the counts into row offsets (an exclusive prefix sum), pass 2 writes the points. NumPy's linspace is
// arange * (delta / div) + start with the last element set to the stop value; x ** 2 is x * x.
__device__ __forceinline__ double uniform_coord(long long i, long long n, double step) {
  return (i == n - 1 && n > 1) ? 1.0 : __dadd_rn(__dmul_rn((double)i, step), -1.0);
}
__global__ void __launch_bounds__(256) pupil_uniform_kernel(long long n, double step, const long long* row_offset,
                                                            long long* row_count, double* Px, double* Py) {
  // one warp per row: count (row_count != null) or write at row_offset[j] in column order
  const long long warps = ((long long)gridDim.x * blockDim.x) >> 5;
  const int lane = threadIdx.x & 31;
  for (long long j = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5; j < n; j += warps) {
    const double y = uniform_coord(j, n, step);
    const double yy = __dmul_rn(y, y);
    long long kept = 0;
    for (long long i0 = 0; i0 < n; i0 += 32) {
      const long long i = i0 + lane;
      bool in = false;
      double x = 0.0;
      if (i < n) {
        x = uniform_coord(i, n, step);
        in = __dadd_rn(__dmul_rn(x, x), yy) <= 1.0;
      }
      const unsigned m = __ballot_sync(0xffffffffu, in);
      if (row_count == nullptr && in) {
        const long long pos = row_offset[j] + kept + __popc(m & ((1u << lane) - 1u));
        Px[pos] = x; Py[pos] = y;
      }
      kept += __popc(m);
    }
    if (row_count != nullptr && lane == 0) row_count[j] = kept;
  }
}

extern "C" int olrt_pupil_uniform_f64(int64_t num_points, const int64_t* row_offset, int64_t* row_count, double* Px,
                                      double* Py, void* stream) {
  if (num_points < 0) return OLRT_E_ARG;
  if ((row_count == nullptr) == (row_offset == nullptr)) return OLRT_E_ARG;   // exactly one of the two passes
  if (row_count == nullptr && (!Px || !Py)) return OLRT_E_ARG;
  if (num_points == 0) return OLRT_OK;
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  long long grid = (long long)sms * 8;
  const long long need = (num_points + 7) / 8;     // 8 warps (rows) per CTA
  if (grid > need) grid = need;
  const double step = (num_points > 1) ? 2.0 / (double)(num_points - 1) : 0.0;
  pupil_uniform_kernel<<<(unsigned)grid, 256, 0, (cudaStream_t)stream>>>(
      (long long)num_points, step, (const long long*)row_offset, (long long*)row_count, Px, Py);
  return cudaGetLastError() == cudaSuccess ? OLRT_OK : OLRT_E_LAUNCH;
}

// ---------------------------------------------------------------------------------------------
// OPD / wavefront epilogue (kernel in trace_strict.cu)
// ---------------------------------------------------------------------------------------------
template <typename R>
static int check_opd(const R* const rays[8], long long n, const olrt_refsphere* s, const R* opd_out) {
  if (!rays || !s || n < 0) return OLRT_E_ARG;
  if (n == 0) return OLRT_OK;
  for (int q = 0; q < 8; ++q) if (q != 6 && !rays[q]) return OLRT_E_ARG;
  return opd_out ? OLRT_OK : OLRT_E_ARG;
}
extern "C" int olrt_opd_epilogue_f64(const double* const rays[8], const double* Px, const double* Py, int64_t n,
                                     const olrt_refsphere* sphere, double* opd_out, double* pupil, int flags,
                                     void* stream) {
  const int rc = check_opd<double>(rays, n, sphere, opd_out);
  if (rc || n == 0) return rc;
  return olrt_strict::opd_f64(rays, Px, Py, n, sphere, opd_out, pupil, flags, (cudaStream_t)stream);
}
extern "C" int olrt_opd_epilogue_f32(const float* const rays[8], const double* Px, const double* Py, int64_t n,
                                     const olrt_refsphere* sphere, float* opd_out, float* pupil, int flags,
                                     void* stream) {
  const int rc = check_opd<float>(rays, n, sphere, opd_out);
  if (rc || n == 0) return rc;
  return olrt_strict::opd_f32(rays, Px, Py, n, sphere, opd_out, pupil, flags, (cudaStream_t)stream);
}
