// Strict build: compiled with -fmad=false so that every + - * / sqrt rounds exactly where the
// reference's NumPy float64 expression rounds (bit-identical vignetting masks).
#include <cstring>
#define OLRT_NS olrt_strict
#include "trace_launch.cuh"
#include "olrt_internal.h"

namespace olrt_strict {
int fwd_f64(const olrt_sys* sys, const double* presc, const int32_t* ops, long long n_rays,
            const double* const in[8], const double* opd_in, double* const out[8], double* record,
            double* dir0, unsigned long long* nr_resid, uint32_t* status, uint32_t* fixup, int flags, cudaStream_t stream) {
  return launch_fwd<double>(sys, presc, ops, n_rays, in, opd_in, out, record, dir0, nr_resid, status, fixup, flags, stream);
}

int fixup_raw_f64(const void* params, size_t bytes, unsigned grid, int block, size_t smem, cudaStream_t stream) {
  FwdParams<double> P;
  if (bytes != sizeof(P)) return OLRT_E_ARG;
  memcpy(&P, params, sizeof(P));
  launch_fixup(P, grid, block, smem, stream);
  return cudaGetLastError() == cudaSuccess ? OLRT_OK : OLRT_E_LAUNCH;
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

// Ray launch (object at infinity or finite, angle / object-height fields, vignetting factors):
//   RayGenerator.generate_rays            optiland/rays/ray_generator.py:28-106
//   AngleField / ObjectHeightField.get_ray_origins   optiland/fields/field_types.py:139-169, :245-272
// The per-field scalars (launch-plane offset or object point, vx, vy) are evaluated on the host; everything per
// ray is in gen_ray (raygen.cuh), in the reference's operation order (this TU is built -fmad=false).
// Rays are field-major: ray f * n_pupil + p. Arithmetic in fp64, stored as R.
template <typename R>
__global__ void __launch_bounds__(256) generate_rays_kernel(const GenParams G, R* ox, R* oy, R* oz, R* oL, R* oM,
                                                            R* oN, R* oi, R* ow) {
  const long long total = G.n_pupil * G.n_fields;
  for (long long r = blockIdx.x * (long long)blockDim.x + threadIdx.x; r < total;
       r += (long long)gridDim.x * blockDim.x) {
    const long long f = r / G.n_pupil, p = r - f * G.n_pupil;
    double x0, y0, z0, dL, dM, dN;
    gen_ray(G, (int)f, p, x0, y0, z0, dL, dM, dN);
    ox[r] = (R)x0; oy[r] = (R)y0; oz[r] = (R)z0;
    oL[r] = (R)dL; oM[r] = (R)dM; oN[r] = (R)dN;
    oi[r] = (R)1.0;
    if (ow != nullptr) ow[r] = (R)G.wavelength;
  }
}
static GenParams to_gen(const olrt_pupil* q) {
  GenParams g;
  g.px = q->Px; g.py = q->Py; g.fxy = q->field_xy; g.vig = q->field_vig; g.fz = q->field_z;
  g.n_pupil = q->n_pupil; g.p_lo = 0; g.len = q->n_pupil;
  g.n_fields = q->n_fields; g.EPD = q->EPD; g.EPL = q->EPL; g.z_start = q->z_start; g.wavelength = q->wavelength;
  g.ready = nullptr; g.ready_chunk = 0;
  return g;
}
template <typename R>
static int launch_generate(const olrt_pupil* pupil, R* const out[8], cudaStream_t stream) {
  const GenParams G = to_gen(pupil);
  const long long total = G.n_pupil * G.n_fields;
  long long grid = (long long)sm_count() * 8;
  const long long need = (total + 255) / 256;
  if (grid > need) grid = need;
  generate_rays_kernel<R><<<(unsigned)grid, 256, 0, stream>>>(G, out[0], out[1], out[2], out[3], out[4], out[5],
                                                              out[6], out[7]);
  return cudaGetLastError() == cudaSuccess ? OLRT_OK : OLRT_E_LAUNCH;
}
int generate_f64(const olrt_pupil* pupil, double* const out[8], cudaStream_t st) {
  return launch_generate<double>(pupil, out, st);
}
int generate_f32(const olrt_pupil* pupil, float* const out[8], cudaStream_t st) {
  return launch_generate<float>(pupil, out, st);
}

// OPD / wavefront epilogue, wavefront/strategy.py:68-116, :118-166, :203-229. This TU is built -fmad=false and
// uses IEEE division / sqrt: for the same traced rays the fp64 results are the reference's bits. Streaming kernel
// (56 B in, 8-32 B out per ray): HBM bound, the arithmetic is free.
template <typename R>
__global__ void __launch_bounds__(256) opd_epilogue_kernel(const R* __restrict__ x, const R* __restrict__ y,
                                                           const R* __restrict__ z, const R* __restrict__ L,
                                                           const R* __restrict__ M, const R* __restrict__ N,
                                                           const R* __restrict__ opd_in, const double* __restrict__ Px,
                                                           const double* __restrict__ Py, long long n,
                                                           const olrt_refsphere S, R* opd_out, R* pupil, int flags) {
  for (long long r = blockIdx.x * (long long)blockDim.x + threadIdx.x; r < n; r += (long long)gridDim.x * blockDim.x) {
    const double xr = (double)x[r], yr = (double)y[r], zr = (double)z[r];
    const double l0 = (double)L[r], m0 = (double)M[r], n0 = (double)N[r];
    const double l = -l0, m = -m0, nn = -n0;
    const double a = l * l + m * m + nn * nn;
    const double b = 2.0 * (l * (xr - S.xc) + m * (yr - S.yc) + nn * (zr - S.zc));
    const double c = xr * xr + yr * yr + zr * zr - 2.0 * (xr * S.xc + yr * S.yc + zr * S.zc) + S.xc * S.xc +
                     S.yc * S.yc + S.zc * S.zc - S.R * S.R;
    double d = b * b - 4.0 * a * c;
    if (d < 0.0) d = 0.0;
    const double sd = sqrt(d);
    double t = (-b - sd) / (2.0 * a);
    if (t < 0.0) t = (-b + sd) / (2.0 * a);
    const double opd_img = S.n_image * t;
    double opd = (double)opd_in[r] - opd_img;
    if (S.tilt_ux != 0.0 || S.tilt_uy != 0.0 || Px != nullptr) {  // AngleField: remove the launch-plane tilt
      const double px = Px ? Px[r] : 0.0, py = Py ? Py[r] : 0.0;
      opd = opd + (S.tilt_ux * (px * S.epd / 2.0) + S.tilt_uy * (py * S.epd / 2.0));
    }
    opd_out[r] = (R)((flags & OLRT_OPD_RAW) ? opd : (S.opd_ref - opd) / (S.wavelength * 1e-3));
    if (pupil != nullptr) {
      const double t2 = opd_img / S.n_image;
      pupil[r] = (R)(xr - t2 * l0);
      pupil[n + r] = (R)(yr - t2 * m0);
      pupil[2 * n + r] = (R)(zr - t2 * n0);
    }
  }
}
template <typename R>
static int launch_opd(const R* const rays[8], const double* Px, const double* Py, long long n, const olrt_refsphere* s,
                      R* opd_out, R* pupil, int flags, cudaStream_t st) {
  long long grid = (long long)sm_count() * 8;
  const long long need = (n + 255) / 256;
  if (grid > need) grid = need;
  opd_epilogue_kernel<R><<<(unsigned)grid, 256, 0, st>>>(rays[0], rays[1], rays[2], rays[3], rays[4], rays[5], rays[7],
                                                         Px, Py, n, *s, opd_out, pupil, flags);
  return cudaGetLastError() == cudaSuccess ? OLRT_OK : OLRT_E_LAUNCH;
}
int opd_f64(const double* const rays[8], const double* Px, const double* Py, long long n, const olrt_refsphere* s,
            double* opd_out, double* pupil, int flags, cudaStream_t st) {
  return launch_opd<double>(rays, Px, Py, n, s, opd_out, pupil, flags, st);
}
int opd_f32(const float* const rays[8], const double* Px, const double* Py, long long n, const olrt_refsphere* s,
            float* opd_out, float* pupil, int flags, cudaStream_t st) {
  return launch_opd<float>(rays, Px, Py, n, s, opd_out, pupil, flags, st);
}
}  // namespace olrt_strict
