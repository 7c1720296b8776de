// trace_fwd.cuh -- the fused forward kernel: SurfaceGroup.trace for RealRays.
//   surface loop            optiland/surfaces/surface_group.py:232-244
//   one surface             optiland/surfaces/standard_surface.py:186-233
// One thread owns one ray; its state (x y z L M N i opd) is read once from HBM,
// stays in registers across every surface, and is written once (plus the
// recorded surfaces asked for). The prescription lives in shared memory.
#pragma once
#include "trace_geom.cuh"

namespace OLRT_NS {

template <typename R>
struct FwdParams {
  const R* in[8];
  const R* opd_in;
  R* out[8];
  R* record;
  R* dir0;
  const R* presc;
  const int32_t* ops;
  unsigned long long* nr_resid;
  uint32_t* status;
  long long n_rays;
  long long rays_per_presc;
  int n_surfaces, n_wavelengths, n_presc, presc_stride, ops_len, n_record;
  R wavelengths[OLRT_MAX_WAVELENGTHS];
  R four_pi;  // 4*pi rounded once on the host like the reference's Python float product
};

template <typename R>
__device__ __forceinline__ void load_surface(SurfParams<R>& sp, const int32_t* s_ops, const R* s_presc,
                                             int s, int W) {
  const int32_t* op = s_ops + s * OLRT_OPS_STRIDE;
  sp.op = op;
  sp.ops = s_ops;
  const R* blk = s_presc + op[OLRT_OP_DOFF];
  sp.frames = blk;
  const R* g = blk + op[OLRT_OP_NFRAMES] * OLRT_FRAME_REALS;
  sp.radius = g[0];
  sp.conic = g[1];
  sp.tol = g[2];
  sp.norm_radius = g[3];
  sp.media = g + OLRT_GEOM_REALS;
  sp.coef = sp.media + 3 * W;
  sp.aper = blk + op[OLRT_OP_APER_DOFF];
  sp.geom = op[OLRT_OP_GEOM];
  sp.ncoef = op[OLRT_OP_NCOEF];
  sp.aux = op[OLRT_OP_AUX];
}

// Surface.trace for one ray. `resid` (shared, may be null) collects max|f| per Newton iteration.
template <typename R>
__device__ __forceinline__ void surface_step(RayState<R>& r, const SurfParams<R>& sp, int widx, int W,
                                             R wavelength, R four_pi, bool valid,
                                             unsigned long long* resid, uint32_t& status,
                                             R& L0, R& M0, R& N0) {
  const int32_t* op = sp.op;
  const int flags = op[OLRT_OP_FLAGS];
  const int nframes = op[OLRT_OP_NFRAMES], fflags = op[OLRT_OP_FRAMEFLAGS];

  cs_localize(r, sp.frames, nframes, fflags);  // geometry.localize, standard_surface.py:199

  // t = geometry.distance(rays), standard_surface.py:211
  R t;
  if (sp.geom == OLRT_GEOM_PLANE) {
    t = (-r.z) / r.N;  // plane.py:61-77
  } else {
    t = conic_distance(r, sp.radius, sp.conic, (flags & OLRT_F_RADIUS_INF) != 0);
    if (sp.geom >= OLRT_GEOM_EVEN_ASPHERE) {  // newton_raphson.py:119-168
      const int K = op[OLRT_OP_NR_ITERS];
      for (int j = 0; j <= K; ++j) {
        if (j == K && resid == nullptr) break;
        const R xi = r.x + t * r.L, yi = r.y + t * r.M, zi = r.z + t * r.N;
        R sag, zx = R(0), zy = R(0);
        if (j < K) nr_eval<R, true>(sp, xi, yi, sag, zx, zy, status);
        else nr_eval<R, false>(sp, xi, yi, sag, zx, zy, status);
        const R f = sag - zi;
        if (resid != nullptr) {
          const unsigned long long key = valid ? resid_bits((double)r_abs(f)) : 0ull;
          const unsigned long long wmax = warp_max_u64(key);
          if ((threadIdx.x & 31) == 0 && wmax != 0ull) atomicMax(resid + j, wmax);
        }
        if (j == K) break;
        R nx, ny, nz;
        slopes_to_normal(zx, zy, nx, ny, nz);
        const R nzs = (r_abs(nz) > R(1e-14)) ? nz : R(1e-14);
        const R fx = (-nx) / nzs, fy = (-ny) / nzs;
        const R df = fx * r.L + fy * r.M - r.N;
        const R dfs = (r_abs(df) > R(1e-14)) ? df : R(1e-14);
        t = t - f / dfs;
      }
    }
  }

  // material_pre.propagation_model.propagate(rays, t), homogeneous.py:30-57
  r.x = r.x + t * r.L;
  r.y = r.y + t * r.M;
  r.z = r.z + t * r.N;
  if (flags & OLRT_F_ABSORB) {
    const R kk = sp.media[2 * W + widx];
    const R alpha = four_pi * kk / wavelength;
    r.i = r.i * r_exp((-alpha) * t * R(1e3));
  }
  // rays.opd = rays.opd + |t * n_pre|, standard_surface.py:218
  const R n1 = sp.media[widx];
  r.opd = r.opd + r_abs(t * n1);

  // aperture.clip, physical_apertures/base.py:71-82 -> real_rays.py:132-139
  const int aoff = op[OLRT_OP_APER_OFF];
  if (aoff >= 0) {
    const bool inside = aperture_contains(sp.ops + aoff, op[OLRT_OP_APER_LEN], sp.aper, r.x, r.y);
    if (!inside) r.i = R(0);
  }

  // interaction_model.interact_real_rays, refractive_reflective_model.py:32-55
  R nx, ny, nz;
  if (sp.geom == OLRT_GEOM_PLANE) {
    nx = R(0); ny = R(0); nz = R(1);  // plane.py:79-98
  } else if (sp.geom == OLRT_GEOM_STANDARD) {
    conic_normal(r.x, r.y, sp.radius, sp.conic, nx, ny, nz);
  } else {
    R sag, zx, zy;
    uint32_t st2 = 0;
    nr_eval<R, true>(sp, r.x, r.y, sag, zx, zy, st2);
    slopes_to_normal(zx, zy, nx, ny, nz);
  }
  L0 = r.L; M0 = r.M; N0 = r.N;
  if (op[OLRT_OP_INTERACT] == 1) {
    reflect(r, nx, ny, nz);
  } else {
    const R n2 = sp.media[W + widx];
    refract(r, nx, ny, nz, n1, n2);
  }

  cs_globalize(r, sp.frames, nframes, fflags);  // geometry.globalize, standard_surface.py:228
}

template <typename R>
__global__ void __launch_bounds__(256) trace_fwd_kernel(const FwdParams<R> P) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  R* s_presc = reinterpret_cast<R*>(smem_raw);
  const int presc_pad = (P.presc_stride + 1) & ~1;
  int32_t* s_ops = reinterpret_cast<int32_t*>(s_presc + presc_pad);
  const int ops_pad = (P.ops_len + 1) & ~1;
  unsigned long long* s_resid = reinterpret_cast<unsigned long long*>(s_ops + ops_pad);
  __shared__ int s_roff[OLRT_MAX_SURFACES + 1];

  const int S = P.n_surfaces, W = P.n_wavelengths;
  for (int q = threadIdx.x; q < P.ops_len; q += blockDim.x) s_ops[q] = P.ops[q];
  __syncthreads();
  const bool want_resid = (P.nr_resid != nullptr);
  if (threadIdx.x == 0) {
    int off = 0;
    for (int s = 0; s < S; ++s) {
      s_roff[s] = off;
      const int32_t* op = s_ops + s * OLRT_OPS_STRIDE;
      if (want_resid && op[OLRT_OP_GEOM] >= OLRT_GEOM_EVEN_ASPHERE) off += op[OLRT_OP_NR_ITERS] + 1;
    }
    s_roff[S] = off;
  }
  __syncthreads();
  const int n_resid = s_roff[S];
  for (int q = threadIdx.x; q < n_resid; q += blockDim.x) s_resid[q] = 0ull;

  uint32_t status = 0;
  int cur_p = -1;
  const long long n_tiles = (P.n_rays + blockDim.x - 1) / blockDim.x;
  for (long long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    const long long base = tile * blockDim.x;
    const int p = (P.n_presc > 1) ? (int)(base / P.rays_per_presc) : 0;
    if (p != cur_p) {
      __syncthreads();
      const R* src = P.presc + (long long)p * P.presc_stride;
      for (int q = threadIdx.x; q < P.presc_stride; q += blockDim.x) s_presc[q] = src[q];
      cur_p = p;
      __syncthreads();
    }
    const long long idx = base + threadIdx.x;
    const bool valid = idx < P.n_rays;
    const long long li = valid ? idx : (P.n_rays - 1);

    RayState<R> r;
    r.x = P.in[0][li]; r.y = P.in[1][li]; r.z = P.in[2][li];
    r.L = P.in[3][li]; r.M = P.in[4][li]; r.N = P.in[5][li];
    r.i = P.in[6][li];
    r.opd = (P.opd_in != nullptr) ? P.opd_in[li] : R(0);
    int widx = 0;
    R wl = P.wavelengths[0];
    if (W > 1) {
      wl = P.in[7][li];
      widx = -1;
      for (int q = 0; q < W; ++q) if (wl == P.wavelengths[q]) widx = q;
      if (widx < 0) { status |= OLRT_ST_WAVELENGTH; widx = 0; r.i = R(0) / R(0); }
    }

    R L0 = r.L, M0 = r.M, N0 = r.N;
    for (int s = 0; s < S; ++s) {
      SurfParams<R> sp;
      load_surface(sp, s_ops, s_presc, s, W);
      unsigned long long* resid =
          (want_resid && sp.geom >= OLRT_GEOM_EVEN_ASPHERE) ? (s_resid + s_roff[s]) : nullptr;
      surface_step(r, sp, widx, W, wl, P.four_pi, valid, resid, status, L0, M0, N0);
      const int slot = sp.op[OLRT_OP_RECORD];
      if (slot >= 0 && valid && P.record != nullptr) {  // Surface._record, standard_surface.py:266-286
        R* rec = P.record + (long long)slot * 8 * P.n_rays + idx;
        rec[0] = r.x; rec[P.n_rays] = r.y; rec[2 * P.n_rays] = r.z;
        rec[3 * P.n_rays] = r.L; rec[4 * P.n_rays] = r.M; rec[5 * P.n_rays] = r.N;
        rec[6 * P.n_rays] = r.i; rec[7 * P.n_rays] = r.opd;
      }
    }
    if (valid) {
      P.out[0][idx] = r.x; P.out[1][idx] = r.y; P.out[2][idx] = r.z;
      P.out[3][idx] = r.L; P.out[4][idx] = r.M; P.out[5][idx] = r.N;
      P.out[6][idx] = r.i; P.out[7][idx] = r.opd;
      if (P.dir0 != nullptr) {
        P.dir0[idx] = L0; P.dir0[P.n_rays + idx] = M0; P.dir0[2 * P.n_rays + idx] = N0;
      }
    }
  }

  // flush block-level results
  __syncthreads();
  if (want_resid) {
    for (int s = 0; s < S; ++s) {
      const int lo = s_roff[s], hi = s_roff[s + 1];
      for (int q = lo + threadIdx.x; q < hi; q += blockDim.x) {
        const unsigned long long v = s_resid[q];
        if (v != 0ull) atomicMax(P.nr_resid + (long long)s * OLRT_RESID_STRIDE + (q - lo), v);
      }
    }
  }
  status = __reduce_or_sync(0xffffffffu, status);
  if (P.status != nullptr && status != 0 && (threadIdx.x & 31) == 0) atomicOr(P.status, status);
}

}  // namespace OLRT_NS
