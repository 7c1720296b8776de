// trace_launch.cuh -- host-side launchers, instantiated once per translation unit
// (strict / fast). Grid = one wave of persistent CTAs: SM count x resident CTAs per SM.
#pragma once
#include "olrt_internal.h"
#include "trace_fwd.cuh"
#include "trace_fk.cuh"
#ifdef OLRT_FX
#include "trace_fx.cuh"
#endif
#include "trace_bwd.cuh"

namespace OLRT_NS {

inline int sm_count() {
  int dev = 0, n = 148;
  if (cudaGetDevice(&dev) == cudaSuccess)
    cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
  return n > 0 ? n : 148;
}

// block size: 256 unless the prescription batch needs tiles that do not straddle prescriptions
inline int pick_block(const olrt_sys* sys, bool fk, bool fx, bool pupil) {
#ifdef OLRT_FX
  if ((sys->n_presc <= 1 || pupil) && fx) return OLRT_FX_BLOCK;  // pupil mode: tiles never straddle groups
#else
  (void)fx;
#endif
  if (sys->n_presc <= 1) return fk ? OLRT_FK_BLOCK : 256;
  for (int b = 256; b >= 32; b >>= 1)
    if (sys->rays_per_presc % b == 0) return b;
  return 0;
}

// The IEEE re-trace of the listed rays always runs the STRICT build's kernel (trace_strict.cu: -fmad=false, the
// reference's operation order), also when the main pass was a tolerance-mode kernel of the fast build: a ray that
// is listed because one of its decisions fell inside a guard band (trace_fx.cuh) gets the reference's decision.
inline void launch_fixup(const FwdParams<double>& P, unsigned grid, int block, size_t smem, cudaStream_t stream) {
#ifdef OLRT_FX
  olrt_strict::fixup_raw_f64(&P, sizeof(P), grid, block, smem, stream);   // same struct layout in both namespaces
#else
  if (smem > 48 * 1024)
    cudaFuncSetAttribute(trace_fwd_kernel<double, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  trace_fwd_kernel<double, true><<<grid, block, smem, stream>>>(P);
#endif
}
inline void launch_fixup(const FwdParams<float>&, unsigned, int, size_t, cudaStream_t) {}

template <typename R>
int launch_fwd(const olrt_sys* sys, const R* presc, const int32_t* ops,
               long long n_rays, const R* const in[8], const R* opd_in, R* const out[8], R* record,
               R* dir0, unsigned long long* nr_resid, uint32_t* status, uint32_t* fixup, int flags,
               cudaStream_t stream, const GenParams* gen = nullptr, double* spot = nullptr,
               const double* spot_shift = nullptr, int spot_flags = 0, const int32_t* group_k = nullptr,
               unsigned long long* group_resid = nullptr) {
  FwdParams<R> P;
  for (int q = 0; q < 8; ++q) { P.in[q] = in ? in[q] : nullptr; P.out[q] = out ? out[q] : nullptr; }
  P.opd_in = opd_in; P.record = record; P.ckpt = nullptr; P.dir0 = dir0; P.presc = presc; P.ops = ops;
  P.nr_resid = nr_resid; P.status = status; P.n_rays = n_rays;
  P.fixup = (sizeof(R) == 8) ? fixup : nullptr;
  P.rays_per_presc = sys->rays_per_presc > 0 ? sys->rays_per_presc : n_rays;
  P.n_surfaces = sys->n_surfaces; P.n_wavelengths = sys->n_wavelengths; P.n_presc = sys->n_presc;
  P.presc_stride = sys->presc_stride; P.ops_len = sys->ops_len; P.n_record = sys->n_record;
  for (int q = 0; q < OLRT_MAX_WAVELENGTHS; ++q) P.wavelengths[q] = (R)sys->wavelengths[q];
  P.four_pi = (R)(4.0 * 3.141592653589793);
  P.n_total = n_rays;
  P.gen.px = nullptr;
  if (gen != nullptr) {  // pupil mode: n_rays = rays of this launch, arrays span every field's full pupil grid
    P.gen = *gen;
    P.n_total = gen->n_pupil * gen->n_fields * sys->n_presc;
  }
  P.spot = spot; P.spot_shift = spot_shift; P.spot_flags = spot_flags;
  P.group_k = (gen != nullptr) ? group_k : nullptr;
  P.group_resid = (gen != nullptr && group_k != nullptr) ? group_resid : nullptr;

  const bool fk = (sys->kernel_hint & OLRT_HINT_FASTKINDS) != 0;
#ifdef OLRT_FX
  const bool fx = fk && !(flags & OLRT_FWD_REPLAY);
#else
  const bool fx = false;
  (void)flags;
#endif
  const int block = pick_block(sys, fk, fx, gen != nullptr);
  if (block == 0) return OLRT_E_ARG;
  if (flags & OLRT_FWD_CKPT_BLOCKED) {  // `record` is the tile-blocked checkpoint buffer
    if (!fx || gen != nullptr || sys->n_record != 0 || sys->n_presc > 1 || block != OLRT_CKPT_TILE || record == nullptr)
      return OLRT_E_UNSUPPORTED;
    P.ckpt = record; P.record = nullptr;
  }
  int resid_words = nr_resid ? sys->nr_resid_words : 0;
  if (P.group_resid != nullptr && resid_words < sys->n_surfaces * OLRT_GROUP_RESID)
    resid_words = sys->n_surfaces * OLRT_GROUP_RESID;
  const size_t smem = (size_t)((sys->presc_stride + 1) & ~1) * sizeof(R) +
                      (size_t)((sys->n_surfaces * der_stride(sys->n_wavelengths) + 1) & ~1) * sizeof(R) +
                      (size_t)((sys->ops_len + 1) & ~1) * sizeof(int32_t) +
                      (size_t)resid_words * sizeof(unsigned long long);
  P.off_der = ((sys->presc_stride + 1) & ~1) * (int)sizeof(R);
  P.off_ops = P.off_der + ((sys->n_surfaces * der_stride(sys->n_wavelengths) + 1) & ~1) * (int)sizeof(R);
  P.off_resid = P.off_ops + ((sys->ops_len + 1) & ~1) * (int)sizeof(int32_t);
  if (smem > 200 * 1024) return OLRT_E_LIMIT;
  if (smem > 48 * 1024) {
    cudaFuncSetAttribute(trace_fwd_kernel<R, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (sizeof(R) == 8)
      cudaFuncSetAttribute(trace_fwd_kernel<R, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  }
  long long n_tiles = (n_rays + block - 1) / block;
  if (gen != nullptr) {
    if (!fx) return OLRT_E_UNSUPPORTED;  // in-kernel ray launch exists in the reduced-algebra kernel only
    n_tiles = ((gen->len + block - 1) / block) * gen->n_fields * sys->n_presc;
  }
  if (spot != nullptr && (!fx || gen == nullptr || spot_shift == nullptr)) return OLRT_E_UNSUPPORTED;
  if (P.fixup != nullptr) cudaMemsetAsync(P.fixup, 0, sizeof(uint32_t), stream);
  int per_sm = 0;
  if (fx) {
#ifdef OLRT_FX
    const size_t xsmem = fx_smem_bytes<R>(sys->presc_stride, sys->n_surfaces, sys->n_wavelengths, sys->ops_len,
                                          resid_words);
    if (xsmem > 200 * 1024) return OLRT_E_LIMIT;
    // group mode without residual reports still needs the per-group trip counts: the maxima variant carries them
    const bool verify = nr_resid != nullptr || P.group_resid != nullptr;
    // per-group counts: bit reports when the ops table's counts fit one register (nr_resid_words <= 32), else maxima
    const bool grp = P.group_k != nullptr;
    const bool gbits = grp && sys->nr_resid_words <= 32 && P.record == nullptr && P.ckpt == nullptr;
    const bool maxima = grp ? !gbits : resid_words > 32;
    const int mode = ((verify || grp) ? 1 : 0) | ((P.record != nullptr || P.ckpt != nullptr) ? 2 : 0) |
                     (((verify || grp) && maxima) ? 4 : 0) | (gbits ? 16 : 0);
    const bool ord = (sys->kernel_hint & OLRT_HINT_ORDINARY) != 0 && (mode & 4) == 0;
    void (*kern)(const FwdParams<R>) = nullptr;
    switch (mode | (ord ? 8 : 0)) {
      case 0: kern = trace_fx_kernel<R, 0>; break;
      case 1: kern = trace_fx_kernel<R, 1>; break;
      case 2: kern = trace_fx_kernel<R, 2>; break;
      case 3: kern = trace_fx_kernel<R, 3>; break;
      case 5: kern = trace_fx_kernel<R, 5>; break;
      case 7: kern = trace_fx_kernel<R, 7>; break;
      case 8: kern = trace_fx_kernel<R, 8>; break;
      case 9: kern = trace_fx_kernel<R, 9>; break;
      case 10: kern = trace_fx_kernel<R, 10>; break;
      case 11: kern = trace_fx_kernel<R, 11>; break;
      case 17: kern = trace_fx_kernel<R, 17>; break;
      default: kern = trace_fx_kernel<R, 25>; break;
    }
    if (xsmem > 48 * 1024) cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)xsmem);
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, block, xsmem);
    if (per_sm < 1) per_sm = 1;
    long long grid = (long long)sm_count() * per_sm;
    if (grid > n_tiles) grid = n_tiles;
    if (gen != nullptr && gen->ready != nullptr) {  // streamed pupil grid: every CTA keeps one field
      if (sys->n_presc != 1 || gen->p_lo != 0 || gen->ready_chunk <= 0 || gen->ready_chunk % 256 != 0 ||
          (group_k != nullptr && !gbits))   // per-field counts ride on the bit reports only
        return OLRT_E_ARG;
      grid = grid / gen->n_fields * gen->n_fields;
      if (grid < gen->n_fields) grid = gen->n_fields;
    }
    kern<<<(unsigned)grid, block, xsmem, stream>>>(P);
#endif
  } else if (fk) {
    if (smem > 48 * 1024)
      cudaFuncSetAttribute(trace_fk_kernel<R>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, trace_fk_kernel<R>, block, smem);
    if (per_sm < 1) per_sm = 1;
    long long grid = (long long)sm_count() * per_sm;
    if (grid > n_tiles) grid = n_tiles;
    trace_fk_kernel<R><<<(unsigned)grid, block, smem, stream>>>(P);
  } else {
    void (*kern)(const FwdParams<R>) = trace_fwd_kernel<R, false>;
#ifdef OLRT_FX
    if (sys->kernel_hint & OLRT_HINT_ZERNIKE) {   // the variant with the merged Zernike tables (a kernel of its own)
      kern = trace_fwd_kernel<R, false, true>;
      if (smem > 48 * 1024) cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    }
#endif
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, block, smem);
    if (per_sm < 1) per_sm = 1;
    long long grid = (long long)sm_count() * per_sm;
    if (grid > n_tiles) grid = n_tiles;
    kern<<<(unsigned)grid, block, smem, stream>>>(P);
  }
  if (cudaGetLastError() != cudaSuccess) return OLRT_E_LAUNCH;
  if (sizeof(R) == 8 && P.fixup != nullptr) {
    // re-trace the listed rays (non-finite results) with IEEE division / sqrt; the count is read on
    // the device, so the grid is fixed: one CTA per SM walks the list
    long long fgrid = sm_count();
    if (fgrid > n_tiles) fgrid = n_tiles;
    launch_fixup(P, (unsigned)fgrid, block > 256 ? 256 : block, smem, stream);
    if (cudaGetLastError() != cudaSuccess) return OLRT_E_LAUNCH;
  }
  return OLRT_OK;
}

}  // namespace OLRT_NS
