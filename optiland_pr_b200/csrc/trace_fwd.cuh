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
  R* ckpt;  // nullptr, or tile-blocked checkpoints [tile][surface][OLRT_CKPT_VALS][OLRT_CKPT_TILE] (olrt.h, OLRT_FWD_CKPT_BLOCKED)
  R* dir0;
  const R* presc;
  const int32_t* ops;
  unsigned long long* nr_resid;
  uint32_t* status;
  uint32_t* fixup;  // [1 + n_rays]: count, then indices of rays to re-trace in IEEE mode (fp64)
  long long n_rays;
  long long rays_per_presc;
  int n_surfaces, n_wavelengths, n_presc, presc_stride, ops_len, n_record;
  R wavelengths[OLRT_MAX_WAVELENGTHS];
  R four_pi;  // 4*pi rounded once on the host like the reference's Python float product
  // length of the out / dir0 / record arrays (ray idx lives at [q][idx], stride n_total); == n_rays unless the
  // rays are launched in the kernel (pupil mode), where one launch covers a pupil range of every field
  long long n_total;
  GenParams gen;             // gen.px != nullptr: pupil mode (olrt_trace_pupil_*), `in` is not read
  // per-group Newton trip counts (pupil mode): the reference traces every (trial, field) on its own, so its exit
  // rule -- max |f| over THAT trace's rays < tol, newton_raphson.py:148 -- yields one K_s per group. group_k
  // [G][S] replaces ops word 6, group_resid [G][S][OLRT_GROUP_RESID] receives each group's residual maxima.
  const int32_t* group_k;
  unsigned long long* group_resid;
  double* spot;              // fused spot epilogue: [n_groups][OLRT_SPOT_SUMS] sums, group = field (pupil mode)
  const double* spot_shift;  // [n_groups][2] shift the sums are taken about (required with spot)
  int spot_flags;            // OLRT_SPOT_*
  // trace_fwd_kernel's dynamic shared memory: byte offsets of the derived constants, the ops table and the residual
  // words behind the prescription row (set by the launcher: the kernel is short of registers and would otherwise
  // recompute them from presc_stride / n_surfaces / ops_len wherever it needs one of the pointers)
  int off_der, off_ops, off_resid;
};

// derived per-surface constants in shared memory: c1, R2, twoR, rR2, then u[W], u2[W], alpha[W]
#define OLRT_DER_BASE 4
__host__ __device__ __forceinline__ int der_stride(int W) { return OLRT_DER_BASE + 3 * W; }

template <typename R> __device__ __forceinline__ R rcp_for_table(R v);
template <> __device__ __forceinline__ double rcp_for_table<double>(double v) { return rcp_refined(v); }
template <> __device__ __forceinline__ float rcp_for_table<float>(float v) { return 1.0f / v; }

// one thread per surface: plain IEEE ops on the prescription (what the reference evaluates per call)
template <typename R>
__device__ __forceinline__ void derive_surface(const int32_t* s_ops, const R* s_presc, R* s_der, int s, int W,
                                               const R* wavelengths, R four_pi) {
  const int32_t* op = s_ops + s * OLRT_OPS_STRIDE;
  const R* g = s_presc + op[OLRT_OP_DOFF] + op[OLRT_OP_NFRAMES] * OLRT_FRAME_REALS;
  R* d = s_der + s * der_stride(W);
  const R radius = g[0], k = g[1];
  d[0] = R(1) + k;
  d[1] = radius * radius;
  d[2] = R(2) * radius;
  d[3] = rcp_for_table<R>(d[1]);
  const R* media = g + OLRT_GEOM_REALS;
  for (int w = 0; w < W; ++w) {
    const R u = media[w] / media[W + w];  // real_rays.py:153
    d[OLRT_DER_BASE + w] = u;
    d[OLRT_DER_BASE + W + w] = u * u;
    d[OLRT_DER_BASE + 2 * W + w] = four_pi * media[2 * W + w] / wavelengths[w];  // homogeneous.py:49
  }
}

// DIRECT: read the prescription where it lies (global memory) and derive the constants on the spot
// (fix-up kernel: every thread may belong to a different prescription of the batch).
template <typename R, bool DIRECT>
__device__ __forceinline__ void load_surface(SurfParams<R>& sp, const int32_t* s_ops, const R* s_presc,
                                             const R* s_der, int s, int W, int widx) {
  const int32_t* op = s_ops + s * OLRT_OPS_STRIDE;
  sp.op = op;
  sp.ops = s_ops;
  const R* blk = s_presc + op[OLRT_OP_DOFF];
  sp.frames = blk;
  const R* g = blk + op[OLRT_OP_NFRAMES] * OLRT_FRAME_REALS;
  sp.cc.radius = g[0];
  sp.cc.k = g[1];
  sp.tol = g[2];
  sp.norm_radius = g[3];
  sp.media = g + OLRT_GEOM_REALS;
  sp.n1 = sp.media[widx];
  if (DIRECT) {
    sp.cc.c1 = R(1) + g[1];
    sp.cc.R2 = g[0] * g[0];
    sp.cc.twoR = R(2) * g[0];
    sp.cc.rR2 = R(0);
    sp.u = sp.n1 / sp.media[W + widx];
    sp.u2 = sp.u * sp.u;
    sp.alpha = R(0);
  } else {
    const R* d = s_der + s * der_stride(W);
    sp.cc.c1 = d[0];
    sp.cc.R2 = d[1];
    sp.cc.twoR = d[2];
    sp.cc.rR2 = d[3];
    sp.u = d[OLRT_DER_BASE + widx];
    sp.u2 = d[OLRT_DER_BASE + W + widx];
    sp.alpha = d[OLRT_DER_BASE + 2 * W + widx];
  }
  sp.coef = sp.media + 3 * W;
  sp.aper = blk + op[OLRT_OP_APER_DOFF];
  sp.geom = op[OLRT_OP_GEOM];
  sp.ncoef = op[OLRT_OP_NCOEF];
  sp.aux = op[OLRT_OP_AUX];
  sp.nr_iters = op[OLRT_OP_NR_ITERS];
  sp.ztab = nullptr;
  sp.zlen = nullptr;
}

#ifdef OLRT_FX
// the CTA's merged Zernike tables (tolerance build, main pass). Function-scope shared arrays: their addresses are
// constants of the kernel, so nothing about them is carried in registers through the surface loop
template <typename R>
__device__ __forceinline__ R* zt_tab() { __shared__ R t[OLRT_ZT_SURF * OLRT_ZT_REALS]; return t; }
__device__ __forceinline__ int* zt_len() { __shared__ int t[OLRT_ZT_SURF * OLRT_ZT_INTS]; return t; }
__device__ __forceinline__ int* zt_slot() { __shared__ int t[OLRT_MAX_SURFACES]; return t; }   // table of a surface, or -1
#endif

// surface kinds: straight-line instantiations for the common cases, the interpreter for the rest
enum { KIND_GENERIC = 0, KIND_PLANE = 1, KIND_CONIC = 2, KIND_EVEN = 3,
       // interpreter only: an x^i y^j polynomial surface on ANY frame chain (tilts, decenters), refracting or reflecting,
       // any aperture program: the geometry is fixed at compile time and its divisions / square roots run unguarded
       // (MM_FAST: the same refinement arithmetic, no slow-path test) -- nothing in that path can hide a NaN behind a
       // select, so a ray outside the fast path's domain comes out non-finite and is re-traced by the IEEE fix-up pass
       KIND_POLY = 4,
       // interpreter only: the generic body (any frame chain, geometry read from the ops table) with the same unguarded
       // arithmetic, for plane / conic / even- and odd-asphere / Zernike surfaces without coating, thin-lens or
       // renormalisation flags. The two square roots whose argument can be exactly zero ON the axis (rho of a Zernike
       // term, r of an odd asphere) keep their guard: their NaN could otherwise disappear behind the reference's own
       // `where(|f'| > 1e-14, ...)` select
       KIND_GFAST = 5,
       // tolerance build only: a KIND_GFAST Zernike surface whose terms are merged into the CTA's shared tables
       // (trace_geom.cuh: zernike_merge / zernike_eval_merged) -- its own instantiation, so the other kinds' code and
       // register allocation stay what they are without it
       KIND_ZMERGED = 6 };

__device__ __forceinline__ int surface_kind(const int32_t* op) {
  const int geom = op[OLRT_OP_GEOM], flags = op[OLRT_OP_FLAGS];
  const bool simple = op[OLRT_OP_NFRAMES] == 1 && op[OLRT_OP_FRAMEFLAGS] == 0 &&
                      !(flags & (OLRT_F_COATING | OLRT_F_THIN_LENS | OLRT_F_RENORM));
  if (!simple) return KIND_GENERIC;
  if (geom == OLRT_GEOM_PLANE) return KIND_PLANE;
  if (geom == OLRT_GEOM_STANDARD && !(flags & OLRT_F_RADIUS_INF)) return KIND_CONIC;
  if (geom == OLRT_GEOM_EVEN_ASPHERE && !(flags & OLRT_F_RADIUS_INF)) return KIND_EVEN;
  return KIND_GENERIC;
}

// Surface.trace for one ray. KIND fixes the geometry and the single untilted frame at compile time.
// `resid` (shared, may be null) collects max|f| per Newton iteration.
template <typename R, int KIND, int MM>
__device__ __forceinline__ void surface_step(RayState<R>& r, const SurfParams<R>& sp, int widx, int W,
                                             R wavelength, R four_pi, bool valid,
                                             unsigned long long* resid, uint32_t& status,
                                             R& L0, R& M0, R& N0, int& guard) {
  const int32_t* op = sp.op;
#ifdef OLRT_FX   // tolerance build, main pass, fp64: decision margins (trace_common.cuh, Margin)
  constexpr bool GD = (MM != MM_IEEE) && (sizeof(R) == 8);
#else
  constexpr bool GD = false;
#endif
  constexpr bool FRAMES = (KIND == KIND_GENERIC || KIND == KIND_POLY || KIND == KIND_GFAST ||
                           KIND == KIND_ZMERGED);   // frame chain with rotations
  const int geom = (KIND == KIND_PLANE) ? OLRT_GEOM_PLANE
                   : (KIND == KIND_CONIC) ? OLRT_GEOM_STANDARD
                   : (KIND == KIND_EVEN) ? OLRT_GEOM_EVEN_ASPHERE
                   : (KIND == KIND_POLY) ? OLRT_GEOM_POLYNOMIAL
                   : (KIND == KIND_ZMERGED) ? OLRT_GEOM_ZERNIKE : sp.geom;
  const int flags = (KIND == KIND_GENERIC) ? op[OLRT_OP_FLAGS]
                    : (KIND == KIND_GFAST || KIND == KIND_ZMERGED) ? (op[OLRT_OP_FLAGS] & (OLRT_F_ABSORB | OLRT_F_RADIUS_INF))
                                           : (op[OLRT_OP_FLAGS] & OLRT_F_ABSORB);
  const int nframes = FRAMES ? op[OLRT_OP_NFRAMES] : 1;
  const int fflags = FRAMES ? op[OLRT_OP_FRAMEFLAGS] : 0;

  // geometry.localize, standard_surface.py:199
  if (FRAMES) cs_localize(r, sp.frames, nframes, fflags);
  else frame_localize(r, sp.frames, 0);

  // t = geometry.distance(rays), standard_surface.py:211
  R t;
  if (geom == OLRT_GEOM_PLANE) {
    t = m_div<MM>(-r.z, r.N);  // plane.py:61-77
  } else {
    t = conic_distance<R, MM, GD>(r, sp.cc, (flags & OLRT_F_RADIUS_INF) != 0, &guard);
    if (geom >= OLRT_GEOM_EVEN_ASPHERE) {  // newton_raphson.py:119-168
      const int K = sp.nr_iters;
      for (int j = 0; j <= K; ++j) {
        if (j == K && resid == nullptr) break;
        const R xi = r.x + t * r.L, yi = r.y + t * r.M, zi = r.z + t * r.N;
        R sag, zx = R(0), zy = R(0);
        if (j < K) {
          if (KIND == KIND_EVEN) even_asphere_eval<R, true, MM>(sp, xi, yi, sag, zx, zy);
          else if (KIND == KIND_POLY) polynomial_eval<R, true, MM, true>(sp, xi, yi, sag, zx, zy);
          else if (KIND == KIND_ZMERGED) zernike_eval_merged<R, true>(sp, xi, yi, sag, zx, zy, status);
          else nr_eval<R, true, MM>(sp, xi, yi, sag, zx, zy, status);
        } else {
          if (KIND == KIND_EVEN) even_asphere_eval<R, false, MM>(sp, xi, yi, sag, zx, zy);
          else if (KIND == KIND_POLY) polynomial_eval<R, false, MM, true>(sp, xi, yi, sag, zx, zy);
          else if (KIND == KIND_ZMERGED) zernike_eval_merged<R, false>(sp, xi, yi, sag, zx, zy, status);
          else nr_eval<R, false, MM>(sp, xi, yi, sag, zx, zy, status);
        }
        const R f = sag - zi;
        if (resid != nullptr) {
          // MM_FAST: a NaN here puts the ray on the fix-up list, whose IEEE re-trace reports it instead
          const unsigned long long key = (valid && !(MM == MM_FAST && f != f)) ? resid_bits((double)r_abs(f)) : 0ull;
          // the IEEE fix-up pass (MM_IEEE) re-traces a handful of listed rays, one per lane: in group mode every lane
          // has its OWN slot pointer and trip count (its ray's group), so no warp-wide aggregation there -- each
          // valid lane raises its own slot
          if (MM == MM_IEEE) { if (key != 0ull) atomicMax(resid + j, key); }
          else report_resid(resid + j, key);
        }
        if (j == K) break;
        R nx, ny, nz;
        slopes_to_normal<R, MM>(zx, zy, nx, ny, nz);
        const R nzs = (r_abs(nz) > R(1e-14)) ? nz : R(1e-14);
        const Recip<R, MM> rn(nzs);
        const R fx = rn.div(-nx), fy = rn.div(-ny);
        const R df = fx * r.L + fy * r.M - r.N;
        const R dfs = (r_abs(df) > R(1e-14)) ? df : R(1e-14);
        Margin<GD>::sw14(guard, nz); Margin<GD>::sw14(guard, df);   // newton_raphson.py:157, :162
        t = t - m_div<MM>(f, dfs);
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
  if (KIND == KIND_GENERIC && (flags & OLRT_F_RENORM)) {  // homogeneous.py:55-57 -> real_rays.py:503-509
    const Recip<R, MM> rm(m_sqrt<MM>(r.L * r.L + r.M * r.M + r.N * r.N));
    r.L = rm.div(r.L); r.M = rm.div(r.M); r.N = rm.div(r.N);
  }
  // rays.opd = rays.opd + |t * n_pre|, standard_surface.py:218
  r.opd = r.opd + r_abs(t * sp.n1);

  // aperture.clip, physical_apertures/base.py:71-82 -> real_rays.py:132-139
  const int aoff = op[OLRT_OP_APER_OFF];
  if (aoff >= 0) {
    const bool inside = aperture_contains<R, GD>(sp.ops + aoff, op[OLRT_OP_APER_LEN], sp.aper, r.x, r.y, &guard);
    if (!inside) r.i = R(0);
  }

  const bool coated = (KIND == KIND_GENERIC) && (flags & OLRT_F_COATING);
  if (KIND == KIND_GENERIC && (flags & OLRT_F_THIN_LENS)) {
    // ThinLensInteractionModel.interact_real_rays, thin_lens_interaction_model.py:55-113
    const R f = sp.aper[coated ? -2 : -1];
    r.opd = r.opd - m_div<MM>(r.x * r.x + r.y * r.y, R(2) * f);
    const R n1 = sp.n1;
    const R n2 = (op[OLRT_OP_INTERACT] == 1) ? -n1 : sp.media[W + widx];
    const Recip<R, MM> rN(r.N);
    const R ux1 = rN.div(r.L), uy1 = rN.div(r.M);
    const R in2 = m_div<MM>(R(1), n2);
    const R ux2 = in2 * (n1 * ux1 - m_div<MM>(r.x, f));
    const R uy2 = in2 * (n1 * uy1 - m_div<MM>(r.y, f));
    L0 = r.L; M0 = r.M; N0 = r.N;
    if (coated) r.i = r.i * sp.aper[-1];
    r.L = ux2; r.M = uy2; r.N = R(1);
  } else {
  // interaction_model.interact_real_rays, refractive_reflective_model.py:32-55
  R nx, ny, nz;
  if (geom == OLRT_GEOM_PLANE) {
    nx = R(0); ny = R(0); nz = R(1);  // plane.py:79-98
  } else if (geom == OLRT_GEOM_STANDARD) {
    conic_normal<R, MM>(r.x, r.y, sp.cc, nx, ny, nz);
  } else {
    R sag, zx, zy;
    uint32_t st2 = 0;
    if (KIND == KIND_EVEN) even_asphere_eval<R, true, MM>(sp, r.x, r.y, sag, zx, zy);
    else if (KIND == KIND_POLY) polynomial_eval<R, true, MM, true>(sp, r.x, r.y, sag, zx, zy);
    else if (KIND == KIND_ZMERGED) zernike_eval_merged<R, true>(sp, r.x, r.y, sag, zx, zy, st2);
    else nr_eval<R, true, MM>(sp, r.x, r.y, sag, zx, zy, st2);
    slopes_to_normal<R, MM>(zx, zy, nx, ny, nz);
  }
  L0 = r.L; M0 = r.M; N0 = r.N;
  if (op[OLRT_OP_INTERACT] == 1) reflect<R, GD>(r, nx, ny, nz, &guard);
  else refract<R, MM, GD>(r, nx, ny, nz, sp.u, sp.u2, &guard);
  if (coated) r.i = r.i * sp.aper[-1];   // BaseCoating.interact -> SimpleCoating.transmit / reflect
  }

  // geometry.globalize, standard_surface.py:228
  if (FRAMES) cs_globalize(r, sp.frames, nframes, fflags);
  else frame_globalize(r, sp.frames, 0);
}

#ifndef OLRT_FWD_MIN_BLOCKS
#define OLRT_FWD_MIN_BLOCKS 2
#endif
// fast-kinds kernel CTA shape (trace_fk.cuh): threads per CTA and resident CTAs per SM the register
// budget is sized for
#ifndef OLRT_FK_BLOCK
#define OLRT_FK_BLOCK 256
#endif
#ifndef OLRT_FK_MIN_BLOCKS
#define OLRT_FK_MIN_BLOCKS 2
#endif

template <typename R> __device__ __forceinline__ bool nonfinite(R v);
template <> __device__ __forceinline__ bool nonfinite<double>(double v) {
  return ((unsigned)__double2hiint(v) << 1) >= 0xffe00000u;  // exponent all ones: inf or NaN
}
template <> __device__ __forceinline__ bool nonfinite<float>(float v) {
  return ((unsigned)__float_as_int(v) << 1) >= 0xff000000u;
}

// all surfaces for one ray; stores the recorded surfaces on the way
template <typename R, bool FIX, bool ZM = false>
__device__ __forceinline__ void trace_ray(const FwdParams<R>& P, RayState<R>& r, const int32_t* s_ops,
                                          const R* presc, const R* s_der, const int* s_kind,
                                          const int* s_roff, unsigned long long* s_resid, int widx, R wl,
                                          bool valid, long long idx, uint32_t& status, R& L0, R& M0, R& N0,
                                          int& guard) {
  const int S = P.n_surfaces, W = P.n_wavelengths;
  const bool want_resid = (P.nr_resid != nullptr);
  for (int s = 0; s < S; ++s) {
    SurfParams<R> sp;
    load_surface<R, FIX>(sp, s_ops, presc, s_der, s, W, widx);
    unsigned long long* resid =
        (want_resid && sp.geom >= OLRT_GEOM_EVEN_ASPHERE) ? (s_resid + s_roff[s]) : nullptr;
    if (FIX && P.group_k != nullptr && P.gen.px != nullptr) {  // group mode: this ray's group decides K and the slots
      const long long g_ = idx / P.gen.n_pupil;
      sp.nr_iters = min(P.group_k[g_ * S + s], OLRT_GROUP_RESID - 1);
      resid = (P.group_resid != nullptr && sp.geom >= OLRT_GEOM_EVEN_ASPHERE)
                  ? P.group_resid + (g_ * S + s) * OLRT_GROUP_RESID : nullptr;
    }
    if (FIX) {
      surface_step<R, KIND_GENERIC, MM_IEEE>(r, sp, widx, W, wl, P.four_pi, valid, resid, status, L0, M0, N0, guard);
    } else {
      switch (s_kind[s]) {  // warp-uniform
        case KIND_PLANE: surface_step<R, KIND_PLANE, MM_FAST>(r, sp, widx, W, wl, P.four_pi, valid, resid, status, L0, M0, N0, guard); break;
        case KIND_CONIC: surface_step<R, KIND_CONIC, MM_FAST>(r, sp, widx, W, wl, P.four_pi, valid, resid, status, L0, M0, N0, guard); break;
        case KIND_EVEN: surface_step<R, KIND_EVEN, MM_FAST>(r, sp, widx, W, wl, P.four_pi, valid, resid, status, L0, M0, N0, guard); break;
        case KIND_POLY: surface_step<R, KIND_POLY, MM_FAST>(r, sp, widx, W, wl, P.four_pi, valid, resid, status, L0, M0, N0, guard); break;
        case KIND_GFAST: surface_step<R, KIND_GFAST, MM_FAST>(r, sp, widx, W, wl, P.four_pi, valid, resid, status, L0, M0, N0, guard); break;
#ifdef OLRT_FX
        case KIND_ZMERGED:
          if (ZM) {
            sp.ztab = zt_tab<R>() + zt_slot()[s] * OLRT_ZT_REALS;
            sp.zlen = zt_len() + zt_slot()[s] * OLRT_ZT_INTS;
            surface_step<R, KIND_ZMERGED, MM_FAST>(r, sp, widx, W, wl, P.four_pi, valid, resid, status, L0, M0, N0, guard);
          }
          break;
#endif
        default: surface_step<R, KIND_GENERIC, MM_GUARD>(r, sp, widx, W, wl, P.four_pi, valid, resid, status, L0, M0, N0, guard); break;
      }
    }
    const int slot = sp.op[OLRT_OP_RECORD];
    if (slot >= 0 && valid && P.record != nullptr) {  // Surface._record, standard_surface.py:266-286
      const long long NT = P.n_total;
      R* rec = P.record + (long long)slot * 8 * NT + idx;
      rec[0] = r.x; rec[NT] = r.y; rec[2 * NT] = r.z;
      rec[3 * NT] = r.L; rec[4 * NT] = r.M; rec[5 * NT] = r.N;
      rec[6 * NT] = r.i; rec[7 * NT] = r.opd;
    }
    if (P.ckpt != nullptr && valid && !(s == S - 1 && s_kind[s] == KIND_PLANE)) {
      // blocked checkpoint of this ray (the fix-up pass repairs single rays; a last plane has none, olrt.h)
      R* ck = P.ckpt + (((idx / OLRT_CKPT_TILE) * S + s) * OLRT_CKPT_VALS) * OLRT_CKPT_TILE + (idx % OLRT_CKPT_TILE);
      ck[0] = r.x; ck[OLRT_CKPT_TILE] = r.y; ck[2 * OLRT_CKPT_TILE] = r.z;
      ck[3 * OLRT_CKPT_TILE] = r.L; ck[4 * OLRT_CKPT_TILE] = r.M; ck[5 * OLRT_CKPT_TILE] = copysign(r.i, r.N);
    }
  }
}

// one ray's contribution to the fused spot sums, straight to global memory (fix-up pass: a handful of rays)
template <typename R>
__device__ __forceinline__ void spot_add_one(const FwdParams<R>& P, int g, const RayState<R>& r) {
  const double iv = (double)r.i;
  if ((P.spot_flags & OLRT_SPOT_VALID_ONLY) && !(iv > 0.0)) return;
  const double dx = (double)r.x - P.spot_shift[2 * g], dy = (double)r.y - P.spot_shift[2 * g + 1];
  double* o = P.spot + (long long)g * OLRT_SPOT_SUMS;
  atomicAdd(o + 0, 1.0); atomicAdd(o + 1, dx); atomicAdd(o + 2, dy);
  atomicAdd(o + 3, dx * dx); atomicAdd(o + 4, dy * dy); atomicAdd(o + 6, iv);
  const double r2 = dx * dx + dy * dy;
  const unsigned long long key = (r2 != r2) ? 0x7ff8000000000000ULL : (unsigned long long)__double_as_longlong(r2);
  atomicMax(reinterpret_cast<unsigned long long*>(o + 5), key);
}

// FIX = false: the main pass over every ray (tiles of blockDim rays, grid-stride).
// FIX = true : the fix-up pass over the rays the main pass listed in P.fixup (fp64 only): IEEE
//              division / square root everywhere, prescription read in place.
// ZM (tolerance build, main pass): the variant that also carries the merged-table Zernike kind (OLRT_HINT_ZERNIKE) --
//              a kernel of its own, so lenses without such a surface run exactly the code they ran before it existed
//              (this kernel waits on instruction fetch: 3 % on the three-mirror system from the extra code alone).
template <typename R, bool FIX, bool ZM = false>
__global__ void __launch_bounds__(256, OLRT_FWD_MIN_BLOCKS) trace_fwd_kernel(const FwdParams<R> P) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int S = P.n_surfaces, W = P.n_wavelengths;
  R* s_presc = reinterpret_cast<R*>(smem_raw);
  R* s_der = reinterpret_cast<R*>(smem_raw + P.off_der);
  int32_t* s_ops = reinterpret_cast<int32_t*>(smem_raw + P.off_ops);
  unsigned long long* s_resid = reinterpret_cast<unsigned long long*>(smem_raw + P.off_resid);
  __shared__ int s_roff[OLRT_MAX_SURFACES + 1];
  __shared__ int s_kind[OLRT_MAX_SURFACES];

  for (int q = threadIdx.x; q < P.ops_len; q += blockDim.x) s_ops[q] = P.ops[q];
  __syncthreads();
  const bool want_resid = (P.nr_resid != nullptr);
  if (threadIdx.x < S) {
    const int32_t* op_ = s_ops + threadIdx.x * OLRT_OPS_STRIDE;
    int k_ = surface_kind(op_);
    if (k_ == KIND_GENERIC && op_[OLRT_OP_GEOM] == OLRT_GEOM_POLYNOMIAL &&
        !(op_[OLRT_OP_FLAGS] & (OLRT_F_RADIUS_INF | OLRT_F_COATING | OLRT_F_THIN_LENS | OLRT_F_RENORM)))
      k_ = KIND_POLY;
    if (k_ == KIND_GENERIC && !(op_[OLRT_OP_FLAGS] & (OLRT_F_COATING | OLRT_F_THIN_LENS | OLRT_F_RENORM))) {
      const int g_ = op_[OLRT_OP_GEOM];
      if (g_ == OLRT_GEOM_PLANE || g_ == OLRT_GEOM_STANDARD || g_ == OLRT_GEOM_EVEN_ASPHERE ||
          g_ == OLRT_GEOM_ODD_ASPHERE || g_ == OLRT_GEOM_ZERNIKE)
        k_ = KIND_GFAST;
    }
    s_kind[threadIdx.x] = k_;
  }
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
#ifdef OLRT_FX
  if (ZM) {   // merged tables for the first OLRT_ZT_SURF Zernike surfaces of the unguarded kind that fit them
    if (threadIdx.x == 0) {
      int used = 0;
      for (int s = 0; s < S; ++s) {
        const int32_t* op = s_ops + s * OLRT_OPS_STRIDE;
        zt_slot()[s] = -1;
        if (s_kind[s] == KIND_GFAST && op[OLRT_OP_GEOM] == OLRT_GEOM_ZERNIKE && used < OLRT_ZT_SURF &&
            zernike_mergeable(s_ops + op[OLRT_OP_AUX], op[OLRT_OP_NCOEF])) {
          zt_slot()[s] = used++;
          s_kind[s] = KIND_ZMERGED;
        }
      }
    }
    __syncthreads();
  }
#endif
  const int n_resid = s_roff[S];
  for (int q = threadIdx.x; q < n_resid; q += blockDim.x) s_resid[q] = 0ull;
  __syncthreads();

  uint32_t status = 0;
  int cur_p = -1;
  long long n_items = P.n_rays;
  if (FIX) n_items = (long long)P.fixup[0];
  const long long n_tiles = (n_items + blockDim.x - 1) / blockDim.x;
  for (long long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    const long long base = tile * blockDim.x;
    if (!FIX) {
      const int p = (P.n_presc > 1) ? (int)(base / P.rays_per_presc) : 0;
      if (p != cur_p) {
        __syncthreads();
        const R* src = P.presc + (long long)p * P.presc_stride;
        for (int q = threadIdx.x; q < P.presc_stride; q += blockDim.x) s_presc[q] = src[q];
        cur_p = p;
        __syncthreads();
        if (threadIdx.x < S) derive_surface<R>(s_ops, s_presc, s_der, threadIdx.x, W, P.wavelengths, P.four_pi);
#ifdef OLRT_FX
        if (ZM && threadIdx.x < S && zt_slot()[threadIdx.x] >= 0) {
          const int32_t* op = s_ops + threadIdx.x * OLRT_OPS_STRIDE;
          const R* coef = s_presc + op[OLRT_OP_DOFF] + op[OLRT_OP_NFRAMES] * OLRT_FRAME_REALS + OLRT_GEOM_REALS + 3 * W;
          const int zs = zt_slot()[threadIdx.x];
          zernike_merge<R>(s_ops + op[OLRT_OP_AUX], coef, op[OLRT_OP_NCOEF], zt_tab<R>() + zs * OLRT_ZT_REALS,
                           zt_len() + zs * OLRT_ZT_INTS);
        }
#endif
        __syncthreads();
      }
    }
    const long long item = base + threadIdx.x;
    const bool valid = item < n_items;
    long long idx = valid ? item : (n_items - 1);
    if (FIX) idx = (long long)P.fixup[1 + idx];
    const R* presc = s_presc;
    if (FIX) {  // pupil mode: idx = (prescription * n_fields + field) * n_pupil + pupil point
      long long pi = 0;
      if (P.n_presc > 1) pi = (P.gen.px != nullptr) ? (idx / P.gen.n_pupil) / P.gen.n_fields : idx / P.rays_per_presc;
      presc = P.presc + pi * P.presc_stride;
    }

    RayState<R> r;
    int widx = 0;
    R wl = P.wavelengths[0];
    if (FIX && P.gen.px != nullptr) {  // pupil mode: launch the listed ray again (idx = f * n_pupil + p)
      const long long g_ = idx / P.gen.n_pupil;
      const int f = (int)(g_ % P.gen.n_fields);
      double x0, y0, z0, dL, dM, dN;
      gen_ray(P.gen, f, idx - g_ * P.gen.n_pupil, x0, y0, z0, dL, dM, dN);
      r.x = (R)x0; r.y = (R)y0; r.z = (R)z0; r.L = (R)dL; r.M = (R)dM; r.N = (R)dN;
      r.i = R(1); r.opd = R(0);
    } else {
      r.x = P.in[0][idx]; r.y = P.in[1][idx]; r.z = P.in[2][idx];
      r.L = P.in[3][idx]; r.M = P.in[4][idx]; r.N = P.in[5][idx];
      r.i = P.in[6][idx];
      r.opd = (P.opd_in != nullptr) ? P.opd_in[idx] : R(0);
    }
    if (W > 1 && !(FIX && P.gen.px != nullptr)) {
      wl = P.in[7][idx];
      widx = -1;
      for (int q = 0; q < W; ++q) if (wl == P.wavelengths[q]) widx = q;
      if (widx < 0) { status |= OLRT_ST_WAVELENGTH; widx = 0; r.i = R(0) / R(0); }
    }

    R L0 = r.L, M0 = r.M, N0 = r.N;
    int guard = 0x7fffffff;   // smallest decision margin of this ray (tolerance build; untouched elsewhere)
    trace_ray<R, FIX, ZM>(P, r, s_ops, presc, s_der, s_kind, s_roff, s_resid, widx, wl, valid, idx, status,
                      L0, M0, N0, guard);
    if (valid) {
      if (P.out[0] != nullptr) {
        P.out[0][idx] = r.x; P.out[1][idx] = r.y; P.out[2][idx] = r.z;
        P.out[3][idx] = r.L; P.out[4][idx] = r.M; P.out[5][idx] = r.N;
        P.out[6][idx] = r.i; P.out[7][idx] = r.opd;
      }
      if (P.dir0 != nullptr) {
        P.dir0[idx] = L0; P.dir0[P.n_total + idx] = M0; P.dir0[2 * P.n_total + idx] = N0;
      }
      if (FIX && P.spot != nullptr)  // the main pass left this ray out of the fused spot sums
        spot_add_one(P, (P.gen.px != nullptr) ? (int)(idx / P.gen.n_pupil) : (int)(idx / P.rays_per_presc), r);
    }
    if (!FIX && sizeof(R) == 8 && P.fixup != nullptr) {
      // anything non-finite (a real miss / TIR, or a value outside the fast path's domain): list the
      // ray for the IEEE re-trace. Warp-aggregated append.
      // ... or a decision taken within 2^-36 of its threshold (tolerance build): the strict build decides
      const bool bad = valid && ((guard < OLRT_GM_HI) | nonfinite(r.x) | nonfinite(r.y) | nonfinite(r.z) | nonfinite(r.L) |
                                 nonfinite(r.M) | nonfinite(r.N) | nonfinite(r.opd) | nonfinite(r.i));
      const unsigned m = __ballot_sync(0xffffffffu, bad);
      if (m != 0u) {
        const int lane = threadIdx.x & 31, leader = __ffs(m) - 1;
        unsigned pos = 0;
        if (lane == leader) pos = atomicAdd(P.fixup, (unsigned)__popc(m));
        pos = __shfl_sync(0xffffffffu, pos, leader);
        if (bad) P.fixup[1 + pos + __popc(m & ((1u << lane) - 1u))] = (uint32_t)idx;
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
