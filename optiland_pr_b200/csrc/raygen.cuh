// raygen.cuh -- one launched ray, shared by the ray-launch kernel (trace_strict.cu) and the trace kernels'
// pupil mode (trace_fx.cuh, trace_fwd.cuh fix-up pass):
//   RayGenerator.generate_rays        optiland/rays/ray_generator.py:28-106
//   AngleField.get_ray_origins        optiland/fields/field_types.py:139-169   (object at infinity / finite)
//   ObjectHeightField.get_ray_origins optiland/fields/field_types.py:245-272
//   vx, vy = 1 - FieldGroup.get_vig_factor   optiland/fields/field_group.py:80-109 (per field, from the host)
// Written with the explicitly rounded intrinsics (never contracted into FMAs) and the kernels' own IEEE
// division / square-root sequences (math64.cuh: the same bits as div.rn.f64 / sqrt.rn.f64 for operands of
// ordinary magnitude, one shared reciprocal for the three quotients, no branches), so every translation unit
// produces the same bits as the reference's NumPy expression (tests/test_raygen.py).
#pragma once
#include <cuda_runtime.h>
#include "math64.cuh"

namespace OLRT_NS {

struct GenParams {
  const double* px;   // [n_pupil] normalised pupil coordinates (device); nullptr = rays come from FwdParams::in
  const double* py;
  const double* fxy;  // [n_fields][2]: launch-plane offset (object at infinity) or object point (finite), x then y
  const double* vig;  // nullptr or [n_fields][2]: vx, vy
  const double* fz;   // nullptr (object at infinity) or [n_fields]: z of the object point
  long long n_pupil;  // pupil points per field = distance between fields in the output arrays
  long long p_lo;     // this launch traces pupil points [p_lo, p_lo + len) of every field
  long long len;
  int n_fields;
  double EPD, EPL, z_start, wavelength;
  const unsigned* ready;  // nullptr, or one word per ready_chunk pupil points, set by the copy engine (olrt.h)
  long long ready_chunk;
};

// ray (field f, pupil point p): origin on the launch plane (or the object point), direction towards the pupil point
// (Px EPD vx / 2, Py EPD vy / 2, EPL). Px * EPD / 2 * vx and Px * EPD * vx / 2 are the same double (the halving is
// exact), and vx = 1 multiplies exactly, so one product serves both and the unvignetted bits do not change.
__device__ __forceinline__ void gen_ray(const GenParams& G, int f, long long p, double& x0, double& y0, double& z0,
                                        double& L, double& M, double& N) {
  const double px = __ldcg(G.px + p), py = __ldcg(G.py + p);  // L2 only: read once per field, maybe just copied in
  const double xf = G.fxy[2 * f], yf = G.fxy[2 * f + 1];
  double x1 = __dmul_rn(px, G.EPD), y1 = __dmul_rn(py, G.EPD);
  if (G.vig != nullptr) {
    x1 = __dmul_rn(x1, G.vig[2 * f]);
    y1 = __dmul_rn(y1, G.vig[2 * f + 1]);
  }
  x1 = __dmul_rn(x1, 0.5);
  y1 = __dmul_rn(y1, 0.5);
  if (G.fz != nullptr) {
    x0 = xf; y0 = yf; z0 = G.fz[f];
  } else {
    x0 = __dadd_rn(x1, xf);
    y0 = __dadd_rn(y1, yf);
    z0 = G.z_start;
  }
  const double dx = __dadd_rn(x1, -x0), dy = __dadd_rn(y1, -y0), dz = __dadd_rn(G.EPL, -z0);
  const double rad = __dadd_rn(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)), __dmul_rn(dz, dz));
  double mag = m_sqrt<MM_FAST>(rad);
  const bool is_zero = (mag < 1e-9) | (rad == 0.0);   // the unguarded sequence returns NaN for sqrt(0)
  if (is_zero) mag = 1.0;
  const Recip<double, MM_FAST> rm(mag);
  L = is_zero ? 0.0 : rm.div(dx);
  M = is_zero ? 0.0 : rm.div(dy);
  N = is_zero ? 1.0 : rm.div(dz);
}

}  // namespace OLRT_NS
