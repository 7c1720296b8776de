// trace_fx.cuh -- the tolerance-mode forward kernel for the fast kinds (plane, sphere / conic, even
// asphere on one untilted frame; radial aperture or none; absorbing media allowed).
//
// The strict kernels (trace_fk.cuh, trace_fwd.cuh) replay the reference's operations one by one so that
// the results are the same bits. This kernel computes the SAME quantities -- the same roots, the same
// Newton iterates, the same clip / TIR decisions -- from algebraically equivalent expressions that need
// far fewer fp64 instructions, and is held to the stated tolerance (1e-10 relative, identical masks)
// instead of to the bit:
//   * 1/b and 1/sqrt(a) come from the MUFU seed (2^-20, measured) and ONE third-order refinement
//     (measured 2^-53 on 2^22 operands, profiles/recipes/seed_probe.cu); a/b = a * (1/b), sqrt(a) = a * rsqrt(a)
//   * sphere normal = (P - C)/R (no square root); conic normal = sgn(R) (x, y, -(R - (1+k) z)) / |.|
//     (standard.py:142-167 evaluates the same vector as (x/(R q), y/(R q), -1)/|.|, q = sqrt(1-(1+k)r^2/R^2),
//     and R q = R - (1+k) z on the sheet the root selection picks)
//   * Newton slopes are used as they are: the reference normalises (zx, zy, -1), then divides nx, ny by nz
//     again (newton_raphson.py:151-160) -- fx = -nx/nz = zx
//   * even-asphere sums in Horner form; Snell as T = u D + n (root - u |dot|) sgn(dot)
// Zero radicands, zero divisors and NaN rays leave this kernel as non-finite outputs and are re-traced by
// the IEEE fix-up pass (trace_fwd.cuh, FIX = true) exactly like in the strict build; fp32 uses the
// hardware's IEEE-safe sqrtf / reciprocal instead (there is no fp32 fix-up pass).
// DECISIONS (fp64). The values above differ from the reference's by rounding (~1e-15 relative), so a yes / no
// decision taken on them -- inside the aperture? does the ray hit the conic (discriminant >= 0)? total internal
// reflection (Snell radicand < 0)? which root? which side of the normal? -- could come out differently from the
// reference's when the deciding quantity sits within that rounding of its threshold. Every such quantity is
// therefore also tested against a GUARD BAND of 2^-36 (1.5e-11 relative, four orders of magnitude wider than the
// rounding differences): a ray inside a band is put on the fix-up list and re-traced by the STRICT build's IEEE
// kernel (-fmad=false, the reference's operation order), so the decision that is stored is always the one the
// reference takes. Masks, NaN patterns and root choices are identical to the reference's for every ray.
#pragma once
#include "trace_fk.cuh"

namespace OLRT_NS {

#ifndef OLRT_FX_BLOCK
#define OLRT_FX_BLOCK 256
#endif
#ifndef OLRT_FX_MIN_BLOCKS
#define OLRT_FX_MIN_BLOCKS 2
#endif
#ifndef OLRT_FX_MIN_BLOCKS_F32  // fp32 rays need half the registers: twice the resident warps
#define OLRT_FX_MIN_BLOCKS_F32 3
#endif

#define OLRT_RESID_FAILED 0x7E37E43C8800759CULL   /* bits of 1e300: "some ray failed |f| < tol" (nr_resid, bits mode) */
#define OLRT_FX_GHI 0x3DB00000   /* high word of 2^-36: the guard band of every decision (see DECISIONS) */
template <typename R> struct FxMath;
template <> struct FxMath<double> {
  static __device__ __forceinline__ double rcp(double b) {
    double r;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(b));
    double e = fma(-b, r, 1.0);
    e = fma(e, e, e);
    return fma(r, e, r);
  }
  static __device__ __forceinline__ double rsqrt(double a) {
    double y;
    asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(a));
    const double e = fma(a, -(y * y), 1.0);
    const double p = fma(e, 0.375, 0.5);
    return fma(p, y * e, y);
  }
  static __device__ __forceinline__ double sqrt_(double a) { return a * rsqrt(a); }
  static __device__ __forceinline__ void sqrt_pair(double a, double& s, double& is) { is = rsqrt(a); s = a * is; }
  static __device__ __forceinline__ double copysign_(double m, double s) { return copysign(m, s); }
  // m * sign(s) for s != 0: flip m's sign bit when s is negative
  static __device__ __forceinline__ double flipsign(double m, double s) {
    return __hiloint2double(__double2hiint(m) ^ (__double2hiint(s) & (int)0x80000000), __double2loint(m));
  }
  // Guard-band tests (see "DECISIONS" above) on the high words -- exponent and top 20 mantissa bits, a monotone proxy
  // of log2 -- in the INTEGER pipe: every test folds one margin into a running minimum `g` (one or two integer
  // instructions, no FP64 issue slot), and the kernel compares that minimum ONCE per ray with OLRT_FX_GHI = the high
  // word of 2^-36. A margin below it (or negative: the quantity itself was negative) lists the ray for the strict pass.
  static __device__ __forceinline__ void g_pos(int& g, double v) { g = min(g, __double2hiint(v)); }   // v > 2^-36 ?
  static __device__ __forceinline__ void g_abs(int& g, double v) { g = min(g, __double2hiint(v) & 0x7fffffff); }
  // a > 2^-18 |b| ?  (the margin is rescaled so that the common threshold applies; flags below 2^-18, not above 2^-16)
  static __device__ __forceinline__ void g_ratio18(int& g, double a, double b) {
    g = min(g, __double2hiint(a) - (__double2hiint(b) & 0x7fffffff) + ((17 << 20) + OLRT_FX_GHI));
  }
  // |v| > 1e-12 ?  (rescaled: hi(1e-12) = 0x3D719799)
  static __device__ __forceinline__ void g_abs12(int& g, double v) {
    g = min(g, (__double2hiint(v) & 0x7fffffff) + (OLRT_FX_GHI - 0x3D719799));
  }
  // |a| > 2^-20 |b| ?  -- conditioning of the conic's quadratic, see fx_conic_t
  static __device__ __forceinline__ void g_ratio20(int& g, double a, double b) {
    g = min(g, (__double2hiint(a) & 0x7fffffff) - (__double2hiint(b) & 0x7fffffff) + ((20 << 20) + OLRT_FX_GHI));
  }
};
// fp32: the hardware's approximate reciprocal / rsqrt / sqrt (1-2 ulp, one MUFU each; 0 -> inf / 0, negative -> NaN
// like the IEEE functions): the fp32 bar is 1e-5 relative, 100x the fp32 rounding unit.
template <> struct FxMath<float> {
  static __device__ __forceinline__ float rcp(float b) {
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(b));
    return r;
  }
  static __device__ __forceinline__ float rsqrt(float a) {
    float r;
    asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(a));
    return r;
  }
  static __device__ __forceinline__ float sqrt_(float a) {
    float r;
    asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(a));
    return r;
  }
  static __device__ __forceinline__ void sqrt_pair(float a, float& s, float& is) { s = sqrt_(a); is = rsqrt(a); }
  static __device__ __forceinline__ float copysign_(float m, float s) { return copysignf(m, s); }
  static __device__ __forceinline__ float flipsign(float m, float s) {
    return __int_as_float(__float_as_int(m) ^ (__float_as_int(s) & (int)0x80000000));
  }
  // fp32 has no fix-up pass and no guard bands (its bar is 1e-5 with the masks of its own arithmetic)
  static __device__ __forceinline__ void g_pos(int&, float) {}
  static __device__ __forceinline__ void g_abs(int&, float) {}
  static __device__ __forceinline__ void g_ratio18(int&, float, float) {}
  static __device__ __forceinline__ void g_abs12(int&, float) {}
  static __device__ __forceinline__ void g_ratio20(int&, float, float) {}
};

// exp() of a Beer-Lambert exponent (homogeneous.py:49: -(4 pi k / lambda) t 1e3). Through optical glass it is a small
// negative number (k ~ 1e-8: |x| ~ 1e-3 per cm), where ten Horner steps of the Taylor series are exact to 3e-18
// (|x| < 1/8) against ~50 instructions and a slow-path call for the library routine, which larger arguments still get.
template <typename R>
__device__ __forceinline__ R fx_exp(R x) {
  if (sizeof(R) == 8 && r_abs(x) < R(0.125)) {
    R p = R(1.0 / 3628800.0);
    p = fma(p, x, R(1.0 / 362880.0)); p = fma(p, x, R(1.0 / 40320.0)); p = fma(p, x, R(1.0 / 5040.0));
    p = fma(p, x, R(1.0 / 720.0)); p = fma(p, x, R(1.0 / 120.0)); p = fma(p, x, R(1.0 / 24.0));
    p = fma(p, x, R(1.0 / 6.0)); p = fma(p, x, R(0.5)); p = fma(p, x, R(1.0));
    return fma(p, x, R(1.0));
  }
  return r_exp(x);
}
template <typename R>
__device__ __forceinline__ R fx_intensity(R i, R att) { return (att != R(0)) ? i * fx_exp(att) : i; }

// per-surface table in shared memory, built once per CTA (per prescription of a Monte-Carlo batch)
template <typename R> struct __align__(16) FxSurf {
  R xc, yc;
  R zc, radius;
  R iR, c1;         // 1/R, 1+k
  R k, c1iR2;       // k, (1+k)/R^2
  R rmax2, rmin2;   // radial aperture (r_max^2, r_min^2)
  R twoR, tol;      // 2 R; Newton exit tolerance (newton_raphson.py:148)
  R gmax, gmin;     // guard bands of the aperture test: 2^-36 r_max^2, 2^-36 r_min^2
};
#define OLRT_FX_GB 1.4551915228366852e-11      /* 2^-36 */
#define OLRT_FX_GB_SQRT 3.814697265625e-06     /* 2^-18 */
template <typename R> struct __align__(16) FxMedium {
  R n1, u, u2, nalpha;  // n_pre, n_pre/n_post, its square, -(4 pi k / lambda) * 1e3
};
struct __align__(16) FxInfo {
  int kind;      // KIND_* | FX_*
  int K;         // Newton updates
  int ncoef;
  int coef_off;  // offset (reals) of this surface's coefficients in the padded table s_coef
  // ncoef above is rounded up to a multiple of 4 there (zero padding leaves the Horner sums unchanged)
  int record;
  int roff;      // first residual slot of this surface
  int pad0, pad1;
};
enum { FX_CENTERED = 16, FX_REFLECT = 32, FX_SPHERE = 64, FX_APER = 128 };

template <typename R>
__host__ __device__ __forceinline__ size_t fx_smem_bytes(int presc_stride, int S, int W, int ops_len, int resid_words) {
  return (size_t)((presc_stride + 3) & ~3) * sizeof(R) + (size_t)((presc_stride + 4 * S + 3) & ~3) * sizeof(R) +
         (size_t)S * sizeof(FxSurf<R>) +
         (size_t)S * W * sizeof(FxMedium<R>) + (size_t)S * sizeof(FxInfo) +
         (size_t)((ops_len + 3) & ~3) * sizeof(int32_t) + (size_t)resid_words * sizeof(unsigned long long);
}

template <typename R>
__device__ __forceinline__ void fx_build_surface(const int32_t* s_ops, const R* s_presc, FxSurf<R>* s_surf,
                                                 FxMedium<R>* s_med, FxInfo* s_info, R* s_coef, int s, int S, int W,
                                                 const R* wavelengths, R four_pi) {
  const int32_t* op = s_ops + s * OLRT_OPS_STRIDE;
  const int doff = op[OLRT_OP_DOFF];
  const R* f = s_presc + doff;
  const R* g = f + OLRT_FRAME_REALS;  // one frame (fast kind)
  const R* media = g + OLRT_GEOM_REALS;
  FxSurf<R>& sf = s_surf[s];
  sf.xc = f[0]; sf.yc = f[1]; sf.zc = f[2];
  const R radius = g[0], k = g[1];
  sf.radius = radius; sf.k = k; sf.c1 = R(1) + k;
  sf.iR = R(1) / radius;
  sf.c1iR2 = (R(1) + k) / (radius * radius);
  sf.twoR = R(2) * radius;
  sf.tol = g[2];
  int kind = s_info[s].kind & 15;
  if (is_pos_zero(f[0]) && is_pos_zero(f[1])) kind |= FX_CENTERED;
  if (op[OLRT_OP_INTERACT] == 1) kind |= FX_REFLECT;
  if (k == R(0)) kind |= FX_SPHERE;
  if (op[OLRT_OP_APER_OFF] >= 0) {
    const R* ap = f + op[OLRT_OP_APER_DOFF];
    sf.rmax2 = ap[0]; sf.rmin2 = ap[1];
    kind |= FX_APER;
  } else {
    sf.rmax2 = R(0); sf.rmin2 = R(0);
  }
  sf.gmax = R(OLRT_FX_GB) * sf.rmax2; sf.gmin = R(OLRT_FX_GB) * sf.rmin2;
  s_info[s].kind = kind;
  // dense variant id for the kernel's switch: 0-2 = plane / conic / even asphere with run-time flags; 3-7 = the same
  // with the flags of an ordinary lens surface (centred, refracting, no aperture) fixed at compile time
  {
    const int k3 = kind & 15;
    int v = (k3 == KIND_PLANE) ? 0 : ((k3 == KIND_CONIC) ? 1 : 2);
    if ((kind & (FX_CENTERED | FX_REFLECT | FX_APER)) == FX_CENTERED) {
      const bool sph = (kind & FX_SPHERE) != 0;
      v = (k3 == KIND_PLANE) ? 3 : ((k3 == KIND_CONIC) ? (sph ? 4 : 5) : (sph ? 6 : 7));
    }
    s_info[s].pad1 = v;
  }
  s_info[s].K = op[OLRT_OP_NR_ITERS];
  // coefficients, zero-padded to a multiple of 4: slot = (reals before this surface) + 4 s bounds the padding
  const int m = ((kind & 15) == KIND_EVEN) ? op[OLRT_OP_NCOEF] : 0;
  const int m4 = (m + 3) & ~3;
  const int coff = (doff + 4 * s + 3) & ~3;
  const R* csrc = media + 3 * W;
  for (int i = 0; i < m4; ++i) s_coef[coff + i] = (i < m) ? csrc[i] : R(0);
  s_info[s].ncoef = m4;
  s_info[s].coef_off = coff;
  s_info[s].record = op[OLRT_OP_RECORD];
  const bool absorb = (op[OLRT_OP_FLAGS] & OLRT_F_ABSORB) != 0;
  for (int w = 0; w < W; ++w) {
    FxMedium<R>& md = s_med[s * W + w];
    const R u = media[w] / media[W + w];
    md.n1 = media[w]; md.u = u; md.u2 = u * u;
    md.nalpha = absorb ? -(four_pi * media[2 * W + w] / wavelengths[w]) * R(1e3) : R(0);
  }
}

// StandardGeometry.distance, standard.py:89-140, for a ray already in the surface frame
template <typename R>
__device__ __forceinline__ R fx_conic_t(R x, R y, R z, R L, R M, R N, const FxSurf<R>& sf, bool sphere, int& guard) {
  typedef FxMath<R> Mx;
  if (sizeof(R) == 4) {
    // fp32: b^2 - 4ac cancels catastrophically for a ray launched far from the vertex, so the
    // origin moves to the ray's closest approach to the vertex first and the roots are taken as q/a and c/q.
    // Root choice is the reference's (standard.py:134).
    const R k = sf.k;
    const R dd = fma(L, L, fma(M, M, N * N));
    const R idd = Mx::rcp(dd);
    const R ts = -fma(x, L, fma(y, M, z * N)) * idd;
    const R xs = fma(ts, L, x), ys = fma(ts, M, y), zs = fma(ts, N, z);
    const R a = fma(k * N, N, dd);
    const R hb = fma(k * N, zs, fma(L, xs, fma(M, ys, N * zs))) - N * sf.radius;
    const R c = fma(k * zs, zs, fma(xs, xs, fma(ys, ys, zs * (zs - sf.twoR))));
    const R sd = Mx::sqrt_(fma(hb, hb, -(a * c)));
    const R q = -(hb + ((hb < R(0)) ? -sd : sd));
    const R t1 = q * Mx::rcp(a);
    const R t2 = (q != R(0)) ? c * Mx::rcp(q) : R(0);
    R t = (r_abs(fma(t1, N, zs)) <= r_abs(fma(t2, N, zs))) ? t1 : t2;
    if (a == R(0)) t = -c * Mx::rcp(R(2) * hb);
    return ts + t;
  }
  const R NN = N * N;
  R a = fma(L, L, fma(M, M, NN));
  R hb = fma(L, x, fma(M, y, N * (z - sf.radius)));
  R c = fma(x, x, fma(y, y, z * (z - sf.twoR)));
  if (!sphere) {
    const R k = sf.k;
    a = fma(k, NN, a);
    hb = fma(k * N, z, hb);
    c = fma(k * z, z, c);
  }
  // Conditioning, not a decision: the REFERENCE's root (-b + sqrt d) / 2a cancels when |b / a| is large -- a -> 0 (k = -1
  // and a ray nearly along the axis, an asymptotic direction of a hyperboloid) or a huge radius (b ~ N R: a "flat"
  // surface entered as R = 1e10) -- and its t then carries rounding noise of ~5 ulp(1) |b / a| ABSOLUTE that no other
  // evaluation order reproduces (measured against the strict build: 3.5 relative at 1e-7 rad from a paraboloid's
  // axis; 1.6e-17 R relative on weak spheres, i.e. 1.7e-7 at R = 1e10). Rays with |a| < 2^-20 |b| take the strict
  // build's value, which is the reference's own
  Mx::g_ratio20(guard, a, hb);
  const R sd = Mx::sqrt_(fma(hb, hb, -(a * c)));  // miss: NaN
  const R ra = Mx::rcp(a);
  const R t1 = (sd - hb) * ra, t2 = (-hb - sd) * ra;
  const R z1 = fma(t1, N, z), z2 = fma(t2, N, z);
  const R az1 = r_abs(z1), az2 = r_abs(z2);
  // guard bands. Hit / miss: sqrt(discriminant) against |b| / 2. Root choice: the two intersections of a sphere (or
  // any conic with k > -1, whose sag has the sign of R) lie on one side of the vertex plane, so |z1| == |z2| needs
  // the discriminant to vanish -- the same band; only a hyperboloid's two sheets need the explicit test.
  Mx::g_ratio18(guard, sd, hb);
  if (!sphere && sf.c1 < R(0) && r_abs(az1 - az2) <= R(OLRT_FX_GB) * (az1 + az2)) guard = -1;
  return (az1 <= az2) ? t1 : t2;  // a == 0 (paraboloid, ray along the axis): non-finite -> fix-up
}

// propagate + Beer-Lambert exponent + OPD + radial clip, with the ray in the surface frame
template <typename R>
__device__ __forceinline__ void fx_arrive(RayState<R>& r, R t, const FxSurf<R>& sf, const FxMedium<R>& md,
                                          int kind, R r2, R& att, int& guard) {
  att = fma(md.nalpha, t, att);
  r.opd = fma(r_abs(t), r_abs(md.n1), r.opd);  // opd + |t n|, standard_surface.py:218
  if (kind & FX_APER) {
    const bool inside = (r2 <= sf.rmax2) & (r2 >= sf.rmin2);
    if (sizeof(R) == 8 && ((r_abs(r2 - sf.rmax2) <= sf.gmax) | (r_abs(r2 - sf.rmin2) < sf.gmin))) guard = -1;
    if (!inside) { r.i = R(0); att = R(0); }
  }
}

// RealRays.refract / reflect, real_rays.py:141-181, with dot = D . n already known
template <typename R>
__device__ __forceinline__ void fx_interact(RayState<R>& r, R nx, R ny, R nz, R dot, const FxMedium<R>& md, int kind,
                                            int& guard) {
  typedef FxMath<R> Mx;
  Mx::g_abs(guard, dot);           // which side of the normal: sign(dot), real_rays.py:529
  if (kind & FX_REFLECT) {
    const R d2 = dot + dot;
    r.L = fma(-d2, nx, r.L); r.M = fma(-d2, ny, r.M); r.N = fma(-d2, nz, r.N);
    return;
  }
  const R v = fma(md.u2, fma(dot, dot, R(-1)), R(1));  // 1 - u^2 (1 - dot^2); < 0: TIR -> NaN
  Mx::g_pos(guard, v);             // refracted or totally reflected
  const R root = Mx::sqrt_(v);
  R g = Mx::flipsign(fma(-md.u, r_abs(dot), root), dot);
  if (dot == R(0)) g = R(0);  // n * sign(0)
  r.L = fma(md.u, r.L, nx * g);
  r.M = fma(md.u, r.M, ny * g);
  r.N = fma(md.u, r.N, nz * g);
}

template <typename R>
__device__ __forceinline__ void fx_localize(RayState<R>& r, const FxSurf<R>& sf, int kind) {
  if (!(kind & FX_CENTERED)) { r.x = r.x - sf.xc; r.y = r.y - sf.yc; }
  r.z = r.z - sf.zc;
}
template <typename R>
__device__ __forceinline__ void fx_globalize(RayState<R>& r, const FxSurf<R>& sf, int kind) {
  if (!(kind & FX_CENTERED)) { r.x = r.x + sf.xc; r.y = r.y + sf.yc; }
  r.z = r.z + sf.zc;
}

template <typename R>
__device__ __forceinline__ void fx_plane(RayState<R>& r, const FxSurf<R>& sf, const FxMedium<R>& md, int kind,
                                         R& att, int& guard) {
  typedef FxMath<R> Mx;
  fx_localize(r, sf, kind);
  const R t = -r.z * Mx::rcp(r.N);  // plane.py:61-77
  r.x = fma(t, r.L, r.x); r.y = fma(t, r.M, r.y); r.z = fma(t, r.N, r.z);
  fx_arrive(r, t, sf, md, kind, fma(r.x, r.x, r.y * r.y), att, guard);
  if (kind & FX_REFLECT) {
    r.N = -r.N;
  } else {  // n = (0, 0, 1): T = (u L, u M, sgn(N) root)
    const R N = r.N;
    const R v = fma(md.u2, fma(N, N, R(-1)), R(1));
    Mx::g_pos(guard, v); Mx::g_abs(guard, N);
    const R root = Mx::sqrt_(v);
    r.L = md.u * r.L; r.M = md.u * r.M;
    r.N = (N == R(0)) ? R(0) * root : Mx::copysign_(root, N);
  }
  fx_globalize(r, sf, kind);
}

template <typename R>
__device__ __forceinline__ void fx_conic(RayState<R>& r, const FxSurf<R>& sf, const FxMedium<R>& md, int kind,
                                         R& att, int& guard) {
  typedef FxMath<R> Mx;
  fx_localize(r, sf, kind);
  const bool sphere = (kind & FX_SPHERE) != 0;
  const R t = fx_conic_t(r.x, r.y, r.z, r.L, r.M, r.N, sf, sphere, guard);
  r.x = fma(t, r.L, r.x); r.y = fma(t, r.M, r.y); r.z = fma(t, r.N, r.z);
  const R r2 = fma(r.x, r.x, r.y * r.y);
  fx_arrive(r, t, sf, md, kind, r2, att, guard);
  R nx, ny, nz;
  if (sphere) {
    nx = r.x * sf.iR; ny = r.y * sf.iR; nz = fma(r.z, sf.iR, R(-1));
  } else {
    const R w = fma(-sf.c1, r.z, sf.radius);
    const R g = Mx::copysign_(Mx::rsqrt(fma(w, w, r2)), sf.radius);
    nx = r.x * g; ny = r.y * g; nz = -w * g;
  }
  const R dot = fma(r.L, nx, fma(r.M, ny, r.N * nz));
  fx_interact(r, nx, ny, nz, dot, md, kind, guard);
  fx_globalize(r, sf, kind);
}

// Even asphere, newton_raphson.py:119-168 + even_asphere.py:82-129.  With s = r^2:
//   sag = s/(R (1 + q)) + s Q(s),  q = sqrt(1 - (1+k) s / R^2),  Q(s) = sum_i C_i s^i
//   dz/dx = x (1/(R q) + 2 (Q + s Q')),  f = sag - z,  f' = zx L + zy M - N
template <typename R>
__device__ __forceinline__ void fx_even(RayState<R>& r, const FxSurf<R>& sf, const FxMedium<R>& md, int kind,
                                        const FxInfo& info, int K, const R* coef, bool valid,
                                        unsigned long long* resid, unsigned* rbits, R& att, int& guard) {
  typedef FxMath<R> Mx;
  fx_localize(r, sf, kind);
  R t = fx_conic_t(r.x, r.y, r.z, r.L, r.M, r.N, sf, (kind & FX_SPHERE) != 0, guard);
  const int m = info.ncoef;
  R xi, yi, zi, zx, zy, fp, r2;
  // residual reports by bits: the tolerance and this surface's first bit, read once (not per iteration)
  const R tol_ = (rbits != nullptr) ? sf.tol : R(0);
  unsigned bit_ = (rbits != nullptr) ? (1u << info.roff) : 0u;
  for (int j = 0;; ++j) {
    xi = fma(t, r.L, r.x); yi = fma(t, r.M, r.y); zi = fma(t, r.N, r.z);
    r2 = fma(xi, xi, yi * yi);
    R q, iq;
    Mx::sqrt_pair(fma(-sf.c1iR2, r2, R(1)), q, iq);
    // Q(s) and Q'(s) by Horner, four coefficients per step (the table is zero-padded to a multiple of 4);
    // the top block starts from Q = c3, Q' = 0
    R Q = R(0), Qp = R(0);
    if (m > 0) {
      const R* cb = coef + (m - 4);
      Q = fma(cb[3], r2, cb[2]); Qp = cb[3];
      Qp = fma(Qp, r2, Q); Q = fma(Q, r2, cb[1]);
      Qp = fma(Qp, r2, Q); Q = fma(Q, r2, cb[0]);
      for (int b = m - 8; b >= 0; b -= 4) {
        cb = coef + b;
        Qp = fma(Qp, r2, Q); Q = fma(Q, r2, cb[3]);
        Qp = fma(Qp, r2, Q); Q = fma(Q, r2, cb[2]);
        Qp = fma(Qp, r2, Q); Q = fma(Q, r2, cb[1]);
        Qp = fma(Qp, r2, Q); Q = fma(Q, r2, cb[0]);
      }
    }
    const R g = fma(R(2), fma(r2, Qp, Q), sf.iR * iq);
    zx = xi * g; zy = yi * g;
    fp = fma(zx, r.L, fma(zy, r.M, -r.N));
    const bool lastj = (j >= K);
    if (lastj && resid == nullptr && rbits == nullptr) break;
    const R f = fma(r2, Q, (r2 * sf.iR) * Mx::rcp(R(1) + q)) - zi;
    if (rbits != nullptr) {
      // the reference's exit test is max_rays |f_j| < tol (newton_raphson.py:148): what the host needs per iteration
      // is whether SOME ray fails it -- one bit per (surface, iteration) kept in a register for the whole launch.
      // NaN compares false here: the fix-up pass (IEEE) reports that ray.
      if (valid && r_abs(f) >= tol_) *rbits |= bit_;
      bit_ <<= 1;
    } else if (resid != nullptr) {  // NaN: the fix-up pass (IEEE) reports it
      const unsigned long long key = (valid && f == f) ? resid_bits((double)r_abs(f)) : 0ull;
      report_resid(resid + j, key);
    }
    if (lastj) break;
    const R dfs = (r_abs(fp) > R(1e-14)) ? fp : R(1e-14);
    Mx::g_abs12(guard, fp);   // near the reference's |f'| > 1e-14 switch, newton_raphson.py:162
    t = fma(-f, Mx::rcp(dfs), t);
  }
  r.x = xi; r.y = yi; r.z = zi;
  fx_arrive(r, t, sf, md, kind, r2, att, guard);
  const R w = Mx::rsqrt(fma(zx, zx, fma(zy, zy, R(1))));
  fx_interact(r, zx * w, zy * w, -w, fp * w, md, kind, guard);
  fx_globalize(r, sf, kind);
}

// group mode: add one group's residual maxima (shared, [S][OLRT_GROUP_RESID]) to its global rows and clear them.
// Called by the whole CTA between two __syncthreads().
__device__ __forceinline__ void fx_group_flush(unsigned long long* group_resid, int g, int S, unsigned long long* s_resid) {
  for (int q = threadIdx.x; q < S * OLRT_GROUP_RESID; q += blockDim.x) {
    const unsigned long long v = s_resid[q];
    if (v != 0ull) atomicMax(group_resid + (long long)g * S * OLRT_GROUP_RESID + q, v);
    s_resid[q] = 0ull;
  }
}

// warp-reduce one thread's spot sums and add them to the group's row (one set of atomics per warp; a CTA
// flushes once per field it visits). Column layout: include/olrt.h, olrt_spot_sums_*.
__device__ __forceinline__ void fx_spot_flush(double* spot, int g, double cnt, double sx, double sy, double sxx,
                                              double syy, double si, double mx, bool mnan) {
  double v[6] = {cnt, sx, sy, sxx, syy, si};
#pragma unroll
  for (int k = 0; k < 6; ++k)
    for (int o = 16; o > 0; o >>= 1) v[k] += __shfl_down_sync(0xffffffffu, v[k], o);
  const unsigned nanmask = __ballot_sync(0xffffffffu, mnan);
  double wm = mnan ? 0.0 : mx;
  for (int o = 16; o > 0; o >>= 1) wm = fmax(wm, __shfl_down_sync(0xffffffffu, wm, o));
  if ((threadIdx.x & 31) == 0) {
    double* o = spot + (long long)g * OLRT_SPOT_SUMS;
    atomicAdd(o + 0, v[0]); atomicAdd(o + 1, v[1]); atomicAdd(o + 2, v[2]);
    atomicAdd(o + 3, v[3]); atomicAdd(o + 4, v[4]); atomicAdd(o + 6, v[5]);
    const unsigned long long key = nanmask ? 0x7ff8000000000000ULL : (unsigned long long)__double_as_longlong(wm);
    atomicMax(reinterpret_cast<unsigned long long*>(o + 5), key);
  }
}

// MODE (compile time, so that a launch carries only the code and registers of what it was asked for):
//   bit 0  Newton verification: the residual reports (nr_resid / group_resid)
//   bit 1  per-surface stores: [S][8][N] records or tile-blocked checkpoints
//   bit 3  every surface is centred, refracting and aperture-free (OLRT_HINT_ORDINARY, verified on the device):
//          only those surface bodies are compiled in
//   bit 2  (with bit 0) the reports are running maxima in shared memory (more than 32 residual slots); without it
//          they are one bit per (surface, iteration) in a register
//   bit 4  (with bit 0, without bit 2) per-group Newton trip counts (FwdParams::group_k) with the bit reports: the
//          register is handed to the group's rows of group_resid whenever a CTA moves on to another group
template <typename R, int MODE>
__global__ void __launch_bounds__(OLRT_FX_BLOCK, (sizeof(R) == 4 ? OLRT_FX_MIN_BLOCKS_F32 : OLRT_FX_MIN_BLOCKS))
trace_fx_kernel(const FwdParams<R> P) {
  constexpr bool M_VERIFY = (MODE & 1) != 0, M_STORES = (MODE & 2) != 0, M_MAXIMA = (MODE & 4) != 0;
  constexpr bool M_ORD = (MODE & 8) != 0, M_GBITS = (MODE & 16) != 0;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int S = P.n_surfaces, W = P.n_wavelengths;
  R* s_presc = reinterpret_cast<R*>(smem_raw);
  R* s_coef = s_presc + ((P.presc_stride + 3) & ~3);
  FxSurf<R>* s_surf = reinterpret_cast<FxSurf<R>*>(s_coef + ((P.presc_stride + 4 * S + 3) & ~3));
  FxMedium<R>* s_med = reinterpret_cast<FxMedium<R>*>(s_surf + S);
  FxInfo* s_info = reinterpret_cast<FxInfo*>(s_med + S * W);
  int32_t* s_ops = reinterpret_cast<int32_t*>(s_info + S);
  unsigned long long* s_resid = reinterpret_cast<unsigned long long*>(s_ops + ((P.ops_len + 3) & ~3));
  __shared__ int s_bad, s_nresid;
  __shared__ unsigned s_ready;
  __shared__ int s_gk[OLRT_MAX_SURFACES];  // group mode: the current group's Newton trip counts

  for (int q = threadIdx.x; q < P.ops_len; q += blockDim.x) s_ops[q] = P.ops[q];
  if (threadIdx.x == 0) s_bad = 0;
  __syncthreads();
  // group mode: per-(trial, field) Newton trip counts and residual maxima (FwdParams::group_k)
  const bool gmode = M_MAXIMA && (P.gen.px != nullptr) && (P.group_k != nullptr);   // launch_fwd picks the variant
  // the same counts with bit reports: slots are laid out by the ops table's counts (the caller stores the maximum
  // over the groups there), a group uses the first K_g + 1 of each surface's
  const bool gbits = M_GBITS && (P.gen.px != nullptr) && (P.group_k != nullptr);
  const bool want_resid = M_VERIFY && ((gmode || gbits) ? (P.group_resid != nullptr) : (P.nr_resid != nullptr));
  if (threadIdx.x < S) {
    const int k = fk_kind(s_ops, threadIdx.x);
    s_info[threadIdx.x].kind = k;
    if (k == KIND_GENERIC) s_bad = 1;
  }
  if (threadIdx.x == 0) {
    int off = 0;
    for (int s = 0; s < S; ++s) {
      s_info[s].roff = off;
      const int32_t* op = s_ops + s * OLRT_OPS_STRIDE;
      if (gmode) { if (want_resid) off += OLRT_GROUP_RESID; }
      else if (want_resid && op[OLRT_OP_GEOM] >= OLRT_GEOM_EVEN_ASPHERE) off += op[OLRT_OP_NR_ITERS] + 1;
    }
    s_nresid = off;
  }
  __syncthreads();
  if (s_bad) {
    if (threadIdx.x == 0 && P.status != nullptr) atomicOr(P.status, OLRT_ST_BAD_HINT);
    return;
  }
  const int n_resid = s_nresid;
  for (int q = threadIdx.x; q < n_resid; q += blockDim.x) s_resid[q] = 0ull;
  __syncthreads();
  // Newton verification by bits (batch-global counts, <= 32 residual slots): a thread ORs "some ray of mine failed
  // |f_j| < tol" into one register; the warps merge at the end of the launch and store OLRT_RESID_FAILED (a huge
  // finite value) in the failing slots of nr_resid -- the host's rule "first j with nr_resid[j] < tol" reads the
  // same counts from it as from the true maxima, which the strict kernels and the fix-up pass store
  const bool bits_mode = M_VERIFY && !M_MAXIMA && want_resid;   // launch_fwd: not group mode, n_resid <= 32
  unsigned rbits = 0u;

  uint32_t status = 0;
  int cur_p = -1, cur_g = -1;
  // pupil mode (olrt_trace_pupil_*): the launch covers pupil points [p_lo, p_lo + len) of every GROUP
  // g = prescription * n_fields + field (Monte-Carlo trials x fields); a tile never straddles two groups, ray
  // (g, p) lives at index g * n_pupil + p of the output arrays and belongs to spot group g. Each CTA walks a
  // contiguous range of tiles, so it stays inside one prescription / one spot group as long as possible.
  const bool pupil = (P.gen.px != nullptr);
  const long long tpf = pupil ? (P.gen.len + blockDim.x - 1) / blockDim.x : 0;  // tiles per group
  const long long n_tiles = pupil ? tpf * P.gen.n_fields * P.n_presc : (P.n_rays + blockDim.x - 1) / blockDim.x;
  const long long NT = P.n_total;
  long long tile_lo = blockIdx.x, tile_hi = n_tiles, tile_step = gridDim.x;
  // streamed pupil grid (olrt_pupil.ready): the grid is still arriving from the host, chunk by chunk in pupil order.
  // CTA b keeps ONE field, b % n_fields, and walks that field's pupil tiles b / n_fields, + gridDim / n_fields, ...
  // so every CTA consumes the pupil in copy order; before a tile it waits for the word of the chunk the tile ends in.
  const bool streamed = pupil && (P.gen.ready != nullptr);
  int ready_c = -1;
  if (streamed) {
    tile_lo = blockIdx.x / P.gen.n_fields;
    tile_hi = tpf;
    tile_step = gridDim.x / P.gen.n_fields;
  } else if (M_GBITS && pupil && P.n_presc == 1) {
    // per-field counts on one prescription: the fields differ in cost, so every CTA takes the same share of each
    // (tiles b, b + grid, ...: still one run of tiles per field)
  } else if (pupil) {
    tile_lo = n_tiles * blockIdx.x / gridDim.x;
    tile_hi = n_tiles * (blockIdx.x + 1) / gridDim.x;
    tile_step = 1;
  }
  // fused spot epilogue: per-thread sums for the current group, flushed when the group changes
  const bool want_spot = (P.spot != nullptr);
  const bool rec_on = M_STORES && (P.record != nullptr);
  const bool ck_on = M_STORES && (P.ckpt != nullptr);   // array mode, 256-thread tiles (launch_fwd checks)
  int sp_g = -1;
  double sp_cnt = 0, sp_x = 0, sp_y = 0, sp_xx = 0, sp_yy = 0, sp_i = 0, sp_max = 0, sp_sx = 0, sp_sy = 0;
  bool sp_nan = false;
  for (long long it = tile_lo; it < tile_hi; it += tile_step) {
    const long long tile = streamed ? it + (long long)(blockIdx.x % P.gen.n_fields) * tpf : it;
    const long long base = tile * blockDim.x;
    if (streamed) {
      long long last = (it + 1) * blockDim.x - 1;
      if (last > P.gen.len - 1) last = P.gen.len - 1;
      const int c = (int)((P.gen.p_lo + last) / P.gen.ready_chunk);
      if (c > ready_c) {  // the copy engine sets the words in order: waiting for the last one covers the tile
        if (threadIdx.x == 0) {  // one poller per CTA; the barrier hands its acquire on to the other threads
          const long long t0 = clock64();
          unsigned v;
          for (;;) {
            asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(P.gen.ready + c) : "memory");
            if (v != 0u || clock64() - t0 > 4000000000LL) break;  // ~2 s: never hang the GPU
            __nanosleep(256);
          }
          s_ready = v;
        }
        __syncthreads();
        const unsigned v = s_ready;
        __syncthreads();
        if (v == 0u) { status |= OLRT_ST_TIMEOUT; break; }  // CTA-uniform
        ready_c = c;
      }
    }
    const int grp = pupil ? (int)(tile / tpf) : 0;
    const int p = (P.n_presc > 1) ? (pupil ? grp / P.gen.n_fields : (int)(base / P.rays_per_presc)) : 0;
    if (p != cur_p) {
      __syncthreads();
      const R* src = P.presc + (long long)p * P.presc_stride;
      for (int q = threadIdx.x; q < P.presc_stride; q += blockDim.x) s_presc[q] = src[q];
      cur_p = p;
      __syncthreads();
      if (threadIdx.x < S)
        fx_build_surface<R>(s_ops, s_presc, s_surf, s_med, s_info, s_coef, threadIdx.x, S, W, P.wavelengths,
                            P.four_pi);
      __syncthreads();
    }
    if (M_GBITS && gbits && grp != cur_g) {  // next (trial, field): hand the bits over, fetch its trip counts
      __syncthreads();
      if (bits_mode && cur_g >= 0) {
        rbits = __reduce_or_sync(0xffffffffu, rbits);
        if ((threadIdx.x & 31) == 0 && rbits != 0u) {
          for (int s = 0; s < S; ++s) {
            const int lo = s_info[s].roff, hi = (s + 1 < S) ? s_info[s + 1].roff : n_resid;
            for (int q = lo; q < hi; ++q)
              if (rbits & (1u << q))
                atomicMax(P.group_resid + ((long long)cur_g * S + s) * OLRT_GROUP_RESID + (q - lo), OLRT_RESID_FAILED);
          }
        }
        rbits = 0u;
      }
      if (threadIdx.x < S) {  // the group's counts replace the table's in s_info: the surface loop is the batch one
        const int k = P.group_k[(long long)grp * S + threadIdx.x];
        const int kmax = s_ops[threadIdx.x * OLRT_OPS_STRIDE + OLRT_OP_NR_ITERS];
        s_info[threadIdx.x].K = k < kmax ? k : kmax;   // never past the slots the ops table's counts laid out
      }
      cur_g = grp;
      __syncthreads();
    }
    if (gmode && grp != cur_g) {  // next (trial, field): hand its residual maxima over, fetch its trip counts
      __syncthreads();
      if (want_resid && cur_g >= 0) fx_group_flush(P.group_resid, cur_g, S, s_resid);
      if (threadIdx.x < S) {
        int k = P.group_k[(long long)grp * S + threadIdx.x];
        s_gk[threadIdx.x] = k < OLRT_GROUP_RESID ? k : OLRT_GROUP_RESID - 1;
      }
      cur_g = grp;
      __syncthreads();
    }
    RayState<R> r;
    bool valid;
    long long idx;
    int group = 0;
    int widx = 0;
    if (pupil) {
      const int f = grp - p * P.gen.n_fields;
      const long long pp = (tile - (long long)grp * tpf) * blockDim.x + threadIdx.x;
      valid = pp < P.gen.len;
      const long long pq = P.gen.p_lo + (valid ? pp : (P.gen.len - 1));
      idx = (long long)grp * P.gen.n_pupil + pq;
      group = grp;
      double x0, y0, z0, dL, dM, dN;
      gen_ray(P.gen, f, pq, x0, y0, z0, dL, dM, dN);
      r.x = (R)x0; r.y = (R)y0; r.z = (R)z0; r.L = (R)dL; r.M = (R)dM; r.N = (R)dN;
      r.i = R(1); r.opd = R(0);
    } else {
      const long long item = base + threadIdx.x;
      valid = item < P.n_rays;
      idx = valid ? item : (P.n_rays - 1);
      r.x = P.in[0][idx]; r.y = P.in[1][idx]; r.z = P.in[2][idx];
      r.L = P.in[3][idx]; r.M = P.in[4][idx]; r.N = P.in[5][idx];
      r.i = P.in[6][idx];
      r.opd = (P.opd_in != nullptr) ? P.opd_in[idx] : R(0);
      if (W > 1) {
        const R wl = P.in[7][idx];
        widx = -1;
        for (int q = 0; q < W; ++q) if (wl == P.wavelengths[q]) widx = q;
        if (widx < 0) { status |= OLRT_ST_WAVELENGTH; widx = 0; r.i = R(0) / R(0); }
      }
    }

    R L0 = r.L, M0 = r.M, N0 = r.N, att = R(0);
    int guard = 0x7fffffff;  // smallest decision margin seen (FxMath::g_*): below OLRT_FX_GHI the strict pass takes the ray
    // this thread's column of the tile's first checkpoint block; surface s sits s blocks further on
    R* const ck_ptr = ck_on ? P.ckpt + (tile * S * OLRT_CKPT_VALS) * OLRT_CKPT_TILE + threadIdx.x : nullptr;
    R i_rec = r.i;          // intensity as recorded, r.i exp(att): exp() only after an absorbing medium
    bool i_stale = false;
    for (int s = 0; s < S; ++s) {
      // direction entering the surface: after the loop, the pre-interaction direction at the last surface
      // (real_rays.py:82-86; the frame of a fast kind has no rotation, so local == global direction)
      L0 = r.L; M0 = r.M; N0 = r.N;
      const FxSurf<R>& sf = s_surf[s];
      const FxMedium<R>& md = s_med[s * W + widx];
      const FxInfo& info = s_info[s];
      const int kind = info.kind;
      const int Ks = gmode ? s_gk[s] : info.K;
      unsigned long long* const rs = (M_MAXIMA && want_resid) ? (s_resid + info.roff) : nullptr;
      unsigned* const rb = bits_mode ? &rbits : nullptr;
      const R* const cf = s_coef + info.coef_off;
      if (M_ORD) {  // ordinary lens: centred, refracting, aperture-free surfaces only (warp-uniform switch)
        switch (info.pad1) {
          case 3: fx_plane<R>(r, sf, md, KIND_PLANE | FX_CENTERED, att, guard); break;
          case 4: fx_conic<R>(r, sf, md, KIND_CONIC | FX_CENTERED | FX_SPHERE, att, guard); break;
          case 5: fx_conic<R>(r, sf, md, KIND_CONIC | FX_CENTERED, att, guard); break;
          case 6: fx_even<R>(r, sf, md, KIND_EVEN | FX_CENTERED | FX_SPHERE, info, Ks, cf, valid, rs, rb, att, guard); break;
          case 7: fx_even<R>(r, sf, md, KIND_EVEN | FX_CENTERED, info, Ks, cf, valid, rs, rb, att, guard); break;
          default: status |= OLRT_ST_BAD_HINT; r.x = R(0) / R(0); break;   // the hint was wrong: nothing usable comes out
        }
      } else {
        switch (info.pad1) {  // warp-uniform, dense (fx_build_surface)
          case 0: fx_plane<R>(r, sf, md, kind, att, guard); break;
          case 1: fx_conic<R>(r, sf, md, kind, att, guard); break;
          case 3: fx_plane<R>(r, sf, md, KIND_PLANE | FX_CENTERED, att, guard); break;
          case 4: fx_conic<R>(r, sf, md, KIND_CONIC | FX_CENTERED | FX_SPHERE, att, guard); break;
          case 5: fx_conic<R>(r, sf, md, KIND_CONIC | FX_CENTERED, att, guard); break;
          case 6: fx_even<R>(r, sf, md, KIND_EVEN | FX_CENTERED | FX_SPHERE, info, Ks, cf, valid, rs, rb, att, guard); break;
          case 7: fx_even<R>(r, sf, md, KIND_EVEN | FX_CENTERED, info, Ks, cf, valid, rs, rb, att, guard); break;
          default: fx_even<R>(r, sf, md, kind, info, Ks, cf, valid, rs, rb, att, guard); break;
        }
      }
      if (!rec_on && !ck_on) continue;
      if (md.nalpha != R(0)) i_stale = true;
      if (ck_on) {  // tile-blocked checkpoint: the state after this surface, immediate offsets from one pointer
        if (i_stale) { i_rec = fx_intensity(r.i, att); i_stale = false; }
        else if (!(r.i != R(0))) i_rec = r.i;
        // a plane as the LAST surface (the image plane) is not checkpointed: nothing downstream needs its exit state
        // and the backward computes the plane's own intensity factor directly (trace_bx.cuh, z0p)
        if (valid && !(s == S - 1 && (kind & 15) == KIND_PLANE)) {
          R* ckp = ck_ptr + (long long)s * (OLRT_CKPT_VALS * OLRT_CKPT_TILE);
          __stcs(ckp, r.x); __stcs(ckp + OLRT_CKPT_TILE, r.y); __stcs(ckp + 2 * OLRT_CKPT_TILE, r.z);
          __stcs(ckp + 3 * OLRT_CKPT_TILE, r.L); __stcs(ckp + 4 * OLRT_CKPT_TILE, r.M);
          __stcs(ckp + 5 * OLRT_CKPT_TILE, FxMath<R>::copysign_(i_rec, r.N));   // N itself is rebuilt by the backward (olrt.h)
        }
        continue;
      }
      const int slot = info.record;
      if (slot >= 0 && valid) {  // Surface._record, standard_surface.py:266-286
        R* rec = P.record + (long long)slot * 8 * NT + idx;  // streaming stores: never re-read by this launch
        __stcs(rec, r.x); __stcs(rec + NT, r.y); __stcs(rec + 2 * NT, r.z);
        __stcs(rec + 3 * NT, r.L); __stcs(rec + 4 * NT, r.M); __stcs(rec + 5 * NT, r.N);
        if (i_stale) { i_rec = fx_intensity(r.i, att); i_stale = false; }
        else if (!(r.i != R(0))) i_rec = r.i;  // clipped at this surface (att was reset with it)
        __stcs(rec + 6 * NT, i_rec); __stcs(rec + 7 * NT, r.opd);
      }
    }
    r.i = fx_intensity(r.i, att);
    if (valid) {
      if (P.out[0] != nullptr) {
        __stcs(P.out[0] + idx, r.x); __stcs(P.out[1] + idx, r.y); __stcs(P.out[2] + idx, r.z);
        __stcs(P.out[3] + idx, r.L); __stcs(P.out[4] + idx, r.M); __stcs(P.out[5] + idx, r.N);
        __stcs(P.out[6] + idx, r.i); __stcs(P.out[7] + idx, r.opd);
      }
      if (P.dir0 != nullptr) {
        P.dir0[idx] = L0; P.dir0[NT + idx] = M0; P.dir0[2 * NT + idx] = N0;
      }
    }
    bool bad = false;
    if (sizeof(R) == 8 && P.fixup != nullptr) {
      bad = valid && ((guard < OLRT_FX_GHI) | nonfinite(r.x) | nonfinite(r.y) | nonfinite(r.z) | nonfinite(r.L) |
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
    if (want_spot) {  // RayOperand.rms_spot_size / SpotDiagram sums, ray.py:336-340, spot_diagram.py:317-357
      if (group != sp_g) {  // warp-uniform (a tile belongs to one field)
        if (sp_g >= 0) fx_spot_flush(P.spot, sp_g, sp_cnt, sp_x, sp_y, sp_xx, sp_yy, sp_i, sp_max, sp_nan);
        sp_g = group;
        sp_cnt = sp_x = sp_y = sp_xx = sp_yy = sp_i = sp_max = 0.0;
        sp_nan = false;
        sp_sx = P.spot_shift[2 * group]; sp_sy = P.spot_shift[2 * group + 1];
      }
      const double iv = (double)r.i;
      // a ray on the fix-up list is added by the fix-up pass with its IEEE result
      if (valid && !bad && !((P.spot_flags & OLRT_SPOT_VALID_ONLY) && !(iv > 0.0))) {
        const double dx = (double)r.x - sp_sx, dy = (double)r.y - sp_sy;
        sp_cnt += 1.0; sp_x += dx; sp_y += dy; sp_xx = fma(dx, dx, sp_xx); sp_yy = fma(dy, dy, sp_yy); sp_i += iv;
        const double r2 = fma(dx, dx, dy * dy);
        if (r2 != r2) sp_nan = true;
        sp_max = fmax(sp_max, r2);
      }
    }
  }
  if (want_spot && sp_g >= 0) fx_spot_flush(P.spot, sp_g, sp_cnt, sp_x, sp_y, sp_xx, sp_yy, sp_i, sp_max, sp_nan);

  __syncthreads();
  if (bits_mode) {
    rbits = __reduce_or_sync(0xffffffffu, rbits);
    if ((threadIdx.x & 31) == 0 && rbits != 0u && !(M_GBITS && gbits && cur_g < 0)) {
      unsigned long long* const dst = (M_GBITS && gbits) ? P.group_resid + (long long)cur_g * S * OLRT_GROUP_RESID
                                                         : P.nr_resid;
      const int stride = (M_GBITS && gbits) ? OLRT_GROUP_RESID : OLRT_RESID_STRIDE;
      for (int s = 0; s < S; ++s) {
        const int lo = s_info[s].roff, hi = (s + 1 < S) ? s_info[s + 1].roff : n_resid;
        for (int q = lo; q < hi; ++q)
          if (rbits & (1u << q)) atomicMax(dst + (long long)s * stride + (q - lo), OLRT_RESID_FAILED);
      }
    }
  } else if (gmode) {
    if (want_resid && cur_g >= 0) fx_group_flush(P.group_resid, cur_g, S, s_resid);
  } else if (want_resid) {
    for (int s = 0; s < S; ++s) {
      const int lo = s_info[s].roff, hi = (s + 1 < S) ? s_info[s + 1].roff : n_resid;
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
