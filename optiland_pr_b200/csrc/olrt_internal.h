// olrt_internal.h -- launch entry points exported by the two kernel translation units.
#pragma once
#include <cuda_runtime.h>
#include "../../include/olrt.h"

#define OLRT_DECL_FWD(NS, SUF, R)                                                                  \
  namespace NS {                                                                                   \
  int fwd_##SUF(const olrt_sys* sys, const R* presc, const int32_t* ops, long long n_rays,         \
                const R* const in[8], const R* opd_in, R* const out[8], R* record, R* dir0,        \
                unsigned long long* nr_resid, uint32_t* status, uint32_t* fixup, int flags,      \
                cudaStream_t stream);                                                              \
  }
OLRT_DECL_FWD(olrt_strict, f64, double)
OLRT_DECL_FWD(olrt_fast, f64, double)
OLRT_DECL_FWD(olrt_fast, f32, float)

#define OLRT_DECL_PUPIL(SUF, R)                                                                          \
  namespace olrt_fast {                                                                                  \
  int pupil_##SUF(const olrt_sys* sys, const R* presc, const int32_t* ops, const olrt_pupil* pupil,      \
                  R* const out[8], R* dir0, unsigned long long* nr_resid, uint32_t* status,              \
                  uint32_t* fixup, double* spot, const double* spot_shift, int spot_flags,               \
                  const int32_t* group_k, unsigned long long* group_resid, int flags,                    \
                  cudaStream_t stream);                                                                  \
  }
OLRT_DECL_PUPIL(f64, double)
OLRT_DECL_PUPIL(f32, float)

namespace olrt_strict {
// launch the strict build's IEEE fix-up kernel on a FwdParams<double> passed by bytes (trace_launch.cuh)
int fixup_raw_f64(const void* params, size_t bytes, unsigned grid, int block, size_t smem, cudaStream_t stream);
int selftest_math(const double* a, const double* b, double* q, double* r, long long n, cudaStream_t stream);
}

#define OLRT_DECL_BWD(SUF, R)                                                                          \
  namespace olrt_fast {                                                                                \
  int bwd_##SUF(const olrt_sys* sys, const R* presc, const int32_t* ops, long long n_rays,             \
                const R* const in[8], const R* record, const R* const gout[8], const R* grecord,       \
                R* const gin[8], R* gpresc, uint32_t* status, int flags, cudaStream_t stream);         \
  }
OLRT_DECL_BWD(f64, double)
OLRT_DECL_BWD(f32, float)

namespace olrt_strict {
int generate_f64(const olrt_pupil* pupil, double* const out[8], cudaStream_t st);
int generate_f32(const olrt_pupil* pupil, float* const out[8], cudaStream_t st);
}

namespace olrt_strict {
int opd_f64(const double* const rays[8], const double* Px, const double* Py, long long n, const olrt_refsphere* s,
            double* opd_out, double* pupil, int flags, cudaStream_t st);
int opd_f32(const float* const rays[8], const double* Px, const double* Py, long long n, const olrt_refsphere* s,
            float* opd_out, float* pupil, int flags, cudaStream_t st);
}
