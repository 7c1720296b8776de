// Fast build: FMA contraction allowed (-fmad=true). fp64 here is NOT bit-faithful to the
// reference (last-ulp differences); fp32 is only ever built fast (its parity bar is 1e-5).
#define OLRT_NS olrt_fast
#define OLRT_FX 1  // this TU also carries the reduced-algebra kernel (trace_fx.cuh)
#include "trace_launch.cuh"
#include "trace_bx.cuh"
#include "olrt_internal.h"

namespace olrt_fast {
int fwd_f64(const olrt_sys* sys, const double* presc, const int32_t* ops, long long n_rays,
            const double* const in[8], const double* opd_in, double* const out[8], double* record,
            double* dir0, unsigned long long* nr_resid, uint32_t* status, uint32_t* fixup, int flags, cudaStream_t stream) {
  return launch_fwd<double>(sys, presc, ops, n_rays, in, opd_in, out, record, dir0, nr_resid, status, fixup, flags, stream);
}
int fwd_f32(const olrt_sys* sys, const float* presc, const int32_t* ops, long long n_rays,
            const float* const in[8], const float* opd_in, float* const out[8], float* record,
            float* dir0, unsigned long long* nr_resid, uint32_t* status, uint32_t* fixup, int flags, cudaStream_t stream) {
  return launch_fwd<float>(sys, presc, ops, n_rays, in, opd_in, out, record, dir0, nr_resid, status, fixup, flags, stream);
}
static GenParams to_gen(const olrt_pupil* q) {
  GenParams g;
  g.px = q->Px; g.py = q->Py; g.fxy = q->field_xy; g.vig = q->field_vig; g.fz = q->field_z;
  g.n_pupil = q->n_pupil; g.p_lo = q->p_lo; g.len = q->p_len;
  g.ready = q->ready; g.ready_chunk = q->ready_chunk;
  g.n_fields = q->n_fields; g.EPD = q->EPD; g.EPL = q->EPL; g.z_start = q->z_start; g.wavelength = q->wavelength;
  return g;
}
int pupil_f64(const olrt_sys* sys, const double* presc, const int32_t* ops, const olrt_pupil* pupil,
              double* const out[8], double* dir0, unsigned long long* nr_resid, uint32_t* status, uint32_t* fixup,
              double* spot, const double* spot_shift, int spot_flags, const int32_t* group_k,
              unsigned long long* group_resid, int flags, cudaStream_t stream) {
  const GenParams g = to_gen(pupil);
  return launch_fwd<double>(sys, presc, ops, g.len * g.n_fields * sys->n_presc, nullptr, nullptr, out, nullptr, dir0, nr_resid, status,
                            fixup, flags, stream, &g, spot, spot_shift, spot_flags, group_k, group_resid);
}
int pupil_f32(const olrt_sys* sys, const float* presc, const int32_t* ops, const olrt_pupil* pupil,
              float* const out[8], float* dir0, unsigned long long* nr_resid, uint32_t* status, uint32_t* fixup,
              double* spot, const double* spot_shift, int spot_flags, const int32_t* group_k,
              unsigned long long* group_resid, int flags, cudaStream_t stream) {
  const GenParams g = to_gen(pupil);
  return launch_fwd<float>(sys, presc, ops, g.len * g.n_fields * sys->n_presc, nullptr, nullptr, out, nullptr, dir0, nr_resid, status,
                           fixup, flags, stream, &g, spot, spot_shift, spot_flags, group_k, group_resid);
}
int bwd_f64(const olrt_sys* sys, const double* presc, const int32_t* ops, long long n_rays,
            const double* const in[8], const double* record, const double* const gout[8], const double* grecord,
            double* const gin[8], double* gpresc, uint32_t* status, int flags, cudaStream_t stream) {
  return launch_bx<double>(sys, presc, ops, n_rays, in, record, gout, grecord, gin, gpresc, status, flags, stream);
}
int bwd_f32(const olrt_sys* sys, const float* presc, const int32_t* ops, long long n_rays,
            const float* const in[8], const float* record, const float* const gout[8], const float* grecord,
            float* const gin[8], float* gpresc, uint32_t* status, int flags, cudaStream_t stream) {
  return launch_bx<float>(sys, presc, ops, n_rays, in, record, gout, grecord, gin, gpresc, status, flags, stream);
}
}  // namespace olrt_fast
