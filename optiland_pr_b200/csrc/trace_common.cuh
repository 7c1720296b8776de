// trace_common.cuh -- device-side building blocks of the fused sequential ray trace.
//
// Every function restates one piece of the reference's per-surface arithmetic
// (file:line cited at each function) for ONE ray held in registers. Operation
// order follows the reference line by line so that the strict build
// (-fmad=false) reproduces NumPy float64 bit-for-bit for + - * / sqrt.
//
// This header is compiled into two translation units:
//   trace_strict.cu  (-fmad=false)  namespace olrt_strict   fp64
//   trace_fast.cu    (-fmad=true)   namespace olrt_fast     fp64 + fp32
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "../../include/olrt.h"
#include "math64.cuh"
#include "raygen.cuh"

namespace OLRT_NS {

template <typename R>
struct RayState {
  R x, y, z, L, M, N, i, opd;
};

template <typename R> __device__ __forceinline__ R r_sqrt(R v);
template <> __device__ __forceinline__ double r_sqrt<double>(double v) { return sqrt(v); }
template <> __device__ __forceinline__ float r_sqrt<float>(float v) { return sqrtf(v); }
template <typename R> __device__ __forceinline__ R r_abs(R v);
template <> __device__ __forceinline__ double r_abs<double>(double v) { return fabs(v); }
template <> __device__ __forceinline__ float r_abs<float>(float v) { return fabsf(v); }
template <typename R> __device__ __forceinline__ R r_exp(R v);
template <> __device__ __forceinline__ double r_exp<double>(double v) { return exp(v); }
template <> __device__ __forceinline__ float r_exp<float>(float v) { return expf(v); }

// np.sign semantics: -1, 0, +1, NaN -> NaN   (real_rays.py:541)
template <typename R>
__device__ __forceinline__ R r_sign(R v) {
  return (v > R(0)) ? R(1) : ((v < R(0)) ? R(-1) : ((v == R(0)) ? R(0) : v));
}

// ---------------------------------------------------------------------------
// CoordinateSystem.localize / globalize     coordinate_system.py:73-107
// RealRays.rotate_x/y/z                     rays/real_rays.py:90-130
// BaseRays.translate                        rays/base.py:28-42
// One frame = x y z cz sz cy sy cx sx ; fl bit0 rz, bit1 ry, bit2 rx truthy.
// ---------------------------------------------------------------------------
template <typename R>
__device__ __forceinline__ void frame_localize(RayState<R>& r, const R* f, int fl) {
  r.x = r.x + (-f[0]);
  r.y = r.y + (-f[1]);
  r.z = r.z + (-f[2]);
  if (fl & 1) {  // rotate_z(-rz): cos(-a)=cos a, sin(-a)=-sin a
    const R c = f[3], s = -f[4];
    const R x = r.x * c - r.y * s, y = r.x * s + r.y * c;
    const R L = r.L * c - r.M * s, M = r.L * s + r.M * c;
    r.x = x; r.y = y; r.L = L; r.M = M;
  }
  if (fl & 2) {  // rotate_y(-ry)
    const R c = f[5], s = -f[6];
    const R x = r.x * c + r.z * s, z = (-r.x) * s + r.z * c;
    const R L = r.L * c + r.N * s, N = (-r.L) * s + r.N * c;
    r.x = x; r.z = z; r.L = L; r.N = N;
  }
  if (fl & 4) {  // rotate_x(-rx)
    const R c = f[7], s = -f[8];
    const R y = r.y * c - r.z * s, z = r.y * s + r.z * c;
    const R M = r.M * c - r.N * s, N = r.M * s + r.N * c;
    r.y = y; r.z = z; r.M = M; r.N = N;
  }
}

template <typename R>
__device__ __forceinline__ void frame_globalize(RayState<R>& r, const R* f, int fl) {
  if (fl & 4) {  // rotate_x(rx)
    const R c = f[7], s = f[8];
    const R y = r.y * c - r.z * s, z = r.y * s + r.z * c;
    const R M = r.M * c - r.N * s, N = r.M * s + r.N * c;
    r.y = y; r.z = z; r.M = M; r.N = N;
  }
  if (fl & 2) {  // rotate_y(ry)
    const R c = f[5], s = f[6];
    const R x = r.x * c + r.z * s, z = (-r.x) * s + r.z * c;
    const R L = r.L * c + r.N * s, N = (-r.L) * s + r.N * c;
    r.x = x; r.z = z; r.L = L; r.N = N;
  }
  if (fl & 1) {  // rotate_z(rz)
    const R c = f[3], s = f[4];
    const R x = r.x * c - r.y * s, y = r.x * s + r.y * c;
    const R L = r.L * c - r.M * s, M = r.L * s + r.M * c;
    r.x = x; r.y = y; r.L = L; r.M = M;
  }
  r.x = r.x + f[0];
  r.y = r.y + f[1];
  r.z = r.z + f[2];
}

// chain: reference_cs first on the way in, last on the way out
template <typename R>
__device__ __forceinline__ void cs_localize(RayState<R>& r, const R* frames, int nframes, int flags) {
  for (int q = 0; q < nframes; ++q)
    frame_localize(r, frames + q * OLRT_FRAME_REALS, (flags >> (3 * q)) & 7);
}
template <typename R>
__device__ __forceinline__ void cs_globalize(RayState<R>& r, const R* frames, int nframes, int flags) {
  for (int q = nframes - 1; q >= 0; --q)
    frame_globalize(r, frames + q * OLRT_FRAME_REALS, (flags >> (3 * q)) & 7);
}

// ---------------------------------------------------------------------------
// Integer powers: NumPy evaluates `a ** e` as a (e = 1), np.square (e = 2) and libm pow() for e >= 3.
// glibc's pow is within 0.53 ulp but not always correctly rounded (measured here: equal to the
// correctly rounded a^3 for 96 % of random doubles), so no device sequence can match it bit for bit;
// the kernels use the chain of rounded products (<= 1 ulp from it). This is why surfaces with
// polynomial terms of degree >= 3 are compared at 1e-10 instead of bit for bit.
// ---------------------------------------------------------------------------
template <typename R> struct PowChain {
  R a, hi;
  __device__ __forceinline__ explicit PowChain(R a_) : a(a_), hi(a_) {}
  __device__ __forceinline__ R value() const { return hi; }
  __device__ __forceinline__ void next() { hi = hi * a; }
};

// ---------------------------------------------------------------------------
// Per-surface constants derived once per CTA (plain IEEE ops on the prescription, so the same
// bits the reference gets when it evaluates them per call): staged in shared memory.
// ---------------------------------------------------------------------------
template <typename R>
struct ConicConst {
  R radius, k;
  R c1;    // 1 + k                 standard.py:84, :155
  R R2;    // radius**2
  R twoR;  // 2 * radius            standard.py:114
  R rR2;   // refined reciprocal of R2 (fp64: shared by every  .../radius**2)
};

template <int MM> __device__ __forceinline__ double div_R2(double a, const ConicConst<double>& cc) {
  return div_with<MM>(a, cc.R2, cc.rR2);
}
template <int MM> __device__ __forceinline__ float div_R2(float a, const ConicConst<float>& cc) {
  return a / cc.R2;
}

// ---------------------------------------------------------------------------
// StandardGeometry.distance                 geometries/standard.py:89-140
// The reference's b = 2(...), d = b^2 - 4ac, t = (-b +- sqrt d)/(2a) is evaluated in the half-b
// form: scaling by 2 and 4 is exact in binary floating point, so b' = b/2, d' = d/4,
// sqrt(d') = sqrt(d)/2 and (-b' +- sqrt d')/a are the SAME bits with fewer multiplications.
// ---------------------------------------------------------------------------
// ---------------------------------------------------------------------------
// Decision margins of the interpreter's tolerance build (G = true; everything below is a no-op otherwise and for
// fp32). Its arithmetic differs from the strict build's by rounding (FMA contraction, Horner sums), so a DECISION --
// the sign of a radicand, an aperture comparison, a root choice, one of the reference's 1e-14 switches -- could fall
// the other way for a ray that sits on the threshold. Each decision folds the high word of its relative distance to
// the threshold into a running minimum in the INTEGER pipe (exponent and top 20 mantissa bits: a monotone proxy of
// log2); a ray whose minimum ends below OLRT_GM_HI (2^-36) is listed for the strict build's re-trace like a
// non-finite one. Same scheme as the reduced-algebra kernel's (trace_fx.cuh, FxMath::g_*).
// ---------------------------------------------------------------------------
#define OLRT_GM_HI 0x3DB00000   /* high word of 2^-36 */
template <bool G> struct Margin {
  // |d| > 2^-36 |b| ?   (d = lhs - rhs of a comparison, b = its scale)
  static __device__ __forceinline__ void rel(int& g, double d, double b) {
    // (the difference is clamped before the bias is added: a scale of exactly 0 -- r_min = 0 -- must not wrap around)
    if (G) g = min(g, min((__double2hiint(d) & 0x7fffffff) - (__double2hiint(b) & 0x7fffffff), 0x3fffffff) +
                          ((36 << 20) + OLRT_GM_HI));
  }
  // |v| > 2^-36 ?   (v is O(1) by construction: a radicand of unit vectors, a cosine)
  static __device__ __forceinline__ void unit(int& g, double v) {
    if (G) g = min(g, __double2hiint(v) & 0x7fffffff);
  }
  // v > 2^-36 ?   (negative: listed as well -- the strict build decides what a negative radicand means)
  static __device__ __forceinline__ void pos(int& g, double v) {
    if (G) g = min(g, __double2hiint(v));
  }
  // | |v| - 1e-14 | > 2^-36 * 1e-14 ?  near one of the reference's `abs(v) > 1e-14` switches (hi(1e-14) = 0x3D06849B)
  static __device__ __forceinline__ void sw14(int& g, double v) {
    if (G) {
      const double d = fabs(v) - 1e-14;
      g = min(g, (__double2hiint(d) & 0x7fffffff) - 0x3D06849B + ((36 << 20) + OLRT_GM_HI));
    }
  }
  static __device__ __forceinline__ void rel(int&, float, float) {}
  static __device__ __forceinline__ void unit(int&, float) {}
  static __device__ __forceinline__ void pos(int&, float) {}
  static __device__ __forceinline__ void sw14(int&, float) {}
};

template <typename R, int MM, bool G = false>
__device__ __forceinline__ R conic_distance(const RayState<R>& r, const ConicConst<R>& cc, bool radius_inf, int* guard = nullptr) {
  int g_ = 0x7fffffff;
  if (radius_inf) {  // standard.py:100-103
    const R Ns = (r_abs(r.N) > R(1e-14)) ? r.N : R(1e-14);
    Margin<G>::sw14(g_, r.N);
    if (G) *guard = min(*guard, g_);
    return m_div<MM>(-r.z, Ns);
  }
  const R radius = cc.radius, k = cc.k;
  if (sizeof(R) == 4) {
    // fp32 only (parity bar is a tolerance): b^2-4ac cancels catastrophically for a
    // ray launched far from the vertex, so move the origin to the ray's closest approach to the
    // vertex first and take the roots as q/a and c/q. Root choice is the reference's (standard.py:134).
    const R dd = r.L * r.L + r.M * r.M + r.N * r.N;
    const R t0 = -(r.x * r.L + r.y * r.M + r.z * r.N) / dd;
    const R x = r.x + t0 * r.L, y = r.y + t0 * r.M, z = r.z + t0 * r.N;
    const R a = k * r.N * r.N + dd;
    const R hb = k * r.N * z + (r.L * x + r.M * y + r.N * z) - r.N * radius;
    const R c = k * z * z - cc.twoR * z + (x * x + y * y + z * z);
    const R sd = r_sqrt(hb * hb - a * c);
    const R q = -(hb + ((hb < R(0)) ? -sd : sd));
    const R t1 = q / a;
    const R t2 = (q != R(0)) ? c / q : R(0);
    R t = (r_abs(z + t1 * r.N) <= r_abs(z + t2 * r.N)) ? t1 : t2;
    if (a == R(0)) t = (-c) / (R(2) * hb);
    return t0 + t;
  }
  const R N2 = r.N * r.N;
  const R z2 = r.z * r.z;
  R a, hb, c;
  if (k == R(0)) {  // 0*v + w == w bit for bit (finite v): the sphere skips the conic terms
    a = r.L * r.L + r.M * r.M + N2;
    hb = r.L * r.x + r.M * r.y - r.N * radius + r.N * r.z;
    c = R(0) - cc.twoR * r.z + r.x * r.x + r.y * r.y + z2;
  } else {
    a = k * N2 + r.L * r.L + r.M * r.M + N2;
    hb = (k * r.N) * r.z + r.L * r.x + r.M * r.y - r.N * radius + r.N * r.z;
    c = k * z2 - cc.twoR * r.z + r.x * r.x + r.y * r.y + z2;
  }
  const R d = hb * hb - a * c;
  const R sd = m_sqrt<MM>(d);
  const Recip<R, MM> ra(a);
  const R t1 = ra.div((-hb) + sd);
  const R t2 = ra.div((-hb) - sd);
  const R z1 = r.z + t1 * r.N;
  const R zz2 = r.z + t2 * r.N;
  R t = (r_abs(z1) <= r_abs(zz2)) ? t1 : t2;
  if (a == R(0)) t = m_div<MM>(-c, R(2) * hb);
  if (G) {   // hit or miss (d against the b^2 it is the small difference of), which root, the a == 0 branch
    Margin<G>::rel(g_, d, hb * hb);
    Margin<G>::rel(g_, r_abs(z1) - r_abs(zz2), z1);
    // conditioning, not a decision: the reference's (-b + sqrt d) / 2a cancels when |b / a| is large -- a -> 0 (k = -1 and
    // a ray nearly along the axis, an asymptotic direction of a hyperboloid) or a huge radius (b ~ N R) -- and its t
    // then carries ~5 ulp(1) |b / a| of absolute noise that depends on the evaluation order. Rays with
    // |a| < 2^-20 |b| take the strict build's (= the reference's) answer. a exactly 0 -- L = M = 0 on a paraboloid -- is
    // 0 in both builds and takes the well-conditioned -c / 2b branch
    if (a != R(0)) Margin<G>::rel(g_, a, hb * R(65536.0));
    *guard = min(*guard, g_);
  }
  return t;
}

// conic part of the sag and of the slope denominator, shared by every NR geometry
// sag:   r2 / (R * (1 + sqrt(1 - (1+k) r2 / R^2)))     standard.py:73-87
// denom: R * sqrt(1 - (1+k) r2 / R^2)                  standard.py:154-156
template <typename R, int MM>
__device__ __forceinline__ void conic_sag_denom(R r2, const ConicConst<R>& cc, R& sag, R& denom) {
  const R s = m_sqrt<MM>(R(1) - div_R2<MM>(cc.c1 * r2, cc));
  sag = m_div<MM>(r2, cc.radius * (R(1) + s));
  denom = cc.radius * s;
}
template <typename R, int MM>
__device__ __forceinline__ R conic_denom(R r2, const ConicConst<R>& cc) {
  return cc.radius * m_sqrt<MM>(R(1) - div_R2<MM>(cc.c1 * r2, cc));
}

// slopes -> unit normal: (zx, zy, -1)/|.|   standard.py:158-165 and every NR _surface_normal
template <typename R, int MM>
__device__ __forceinline__ void slopes_to_normal(R zx, R zy, R& nx, R& ny, R& nz) {
  const Recip<R, MM> rm(m_sqrt<MM>(zx * zx + zy * zy + R(1)));
  nx = rm.div(zx);
  ny = rm.div(zy);
  nz = rm.div(R(-1));
}

// StandardGeometry.surface_normal           geometries/standard.py:142-167
template <typename R, int MM>
__device__ __forceinline__ void conic_normal(R x, R y, const ConicConst<R>& cc, R& nx, R& ny, R& nz) {
  const R r2 = x * x + y * y;
  const Recip<R, MM> rd(conic_denom<R, MM>(r2, cc));
  slopes_to_normal<R, MM>(rd.div(x), rd.div(y), nx, ny, nz);
}

// ---------------------------------------------------------------------------
// RealRays.refract / reflect / _align_surface_normal   rays/real_rays.py:141-181, 511-547
// u = n1/n2 and u2 = u*u are per-(surface, wavelength) constants.
// ---------------------------------------------------------------------------
template <typename R>
__device__ __forceinline__ void align_normal(R& nx, R& ny, R& nz, R& dot) {
  if (dot < R(0)) { nx = -nx; ny = -ny; nz = -nz; }        // n * sign(dot), sign = -1
  else if (!(dot > R(0))) { const R sg = r_sign(dot); nx = nx * sg; ny = ny * sg; nz = nz * sg; }
  dot = r_abs(dot);
}

template <typename R, int MM, bool G = false>
__device__ __forceinline__ void refract(RayState<R>& r, R nx, R ny, R nz, R u, R u2, int* guard = nullptr) {
  const R L0 = r.L, M0 = r.M, N0 = r.N;
  R dot = L0 * nx + M0 * ny + N0 * nz;
  align_normal(nx, ny, nz, dot);
  const R rad = R(1) - u2 * (R(1) - dot * dot);
  if (G) {   // which side of the normal (real_rays.py:529), refracted or totally reflected
    int g_ = *guard;
    Margin<G>::unit(g_, dot);
    Margin<G>::pos(g_, rad);
    *guard = g_;
  }
  const R root = m_sqrt<MM>(rad);  // NaN => TIR
  r.L = u * L0 + nx * root - (u * nx) * dot;
  r.M = u * M0 + ny * root - (u * ny) * dot;
  r.N = u * N0 + nz * root - (u * nz) * dot;
}

template <typename R, bool G = false>
__device__ __forceinline__ void reflect(RayState<R>& r, R nx, R ny, R nz, int* guard = nullptr) {
  R dot = r.L * nx + r.M * ny + r.N * nz;
  if (G) { int g_ = *guard; Margin<G>::unit(g_, dot); *guard = g_; }
  align_normal(nx, ny, nz, dot);
  const R d2 = R(2) * dot;
  r.L = r.L - d2 * nx;
  r.M = r.M - d2 * ny;
  r.N = r.N - d2 * nz;
}

// ---------------------------------------------------------------------------
// Aperture program (reverse Polish) on local x, y     physical_apertures/*.py
// ---------------------------------------------------------------------------
// G: every leaf comparison also reports its margin (any leaf on its edge lists the ray, whatever the tree does with it)
template <typename R, bool G = false>
__device__ __forceinline__ bool aperture_contains(const int32_t* prog, int len, const R* p, R x, R y, int* guard = nullptr) {
  uint32_t stack = 0;  // bit stack, top = bit 0
  int g_ = 0x7fffffff;
  for (int q = 0; q < len; ++q) {
    const int tok = prog[q];
    bool v;
    switch (tok) {
      case OLRT_AP_RADIAL: {  // radial.py:56-70
        const R r2 = x * x + y * y;
        v = (r2 <= p[0]) & (r2 >= p[1]);
        Margin<G>::rel(g_, r2 - p[0], p[0]); Margin<G>::rel(g_, r2 - p[1], p[1]);
        p += 2;
        stack = (stack << 1) | (v ? 1u : 0u);
        break;
      }
      case OLRT_AP_OFFSET_RADIAL: {  // offset_radial.py:48
        const R dx = x - p[2], dy = y - p[3];
        const R r2 = dx * dx + dy * dy;
        v = (r2 <= p[0]) & (r2 >= p[1]);
        Margin<G>::rel(g_, r2 - p[0], p[0]); Margin<G>::rel(g_, r2 - p[1], p[1]);
        p += 4;
        stack = (stack << 1) | (v ? 1u : 0u);
        break;
      }
      case OLRT_AP_ELLIPTICAL: {  // elliptical.py:42
        const R dx = x - p[2], dy = y - p[3];
        const R e = (dx * dx) / p[0] + (dy * dy) / p[1];
        v = e <= R(1);
        Margin<G>::unit(g_, e - R(1));
        p += 4;
        stack = (stack << 1) | (v ? 1u : 0u);
        break;
      }
      case OLRT_AP_RECTANGULAR: {  // rectangular.py:42
        v = (p[0] <= x) & (x <= p[1]) & (p[2] <= y) & (y <= p[3]);
        if (G) {   // the scale of an edge at 0 is the opening's width
          const R wx = p[1] - p[0], wy = p[3] - p[2];
          Margin<G>::rel(g_, x - p[0], wx); Margin<G>::rel(g_, x - p[1], wx);
          Margin<G>::rel(g_, y - p[2], wy); Margin<G>::rel(g_, y - p[3], wy);
        }
        p += 4;
        stack = (stack << 1) | (v ? 1u : 0u);
        break;
      }
      default: {  // boolean nodes, base.py:255-336: operands a (below) and b (top)
        const uint32_t b = stack & 1u, a = (stack >> 1) & 1u;
        uint32_t o;
        if (tok == OLRT_AP_OR) o = a | b;
        else if (tok == OLRT_AP_AND) o = a & b;
        else o = a & (b ^ 1u);
        stack = ((stack >> 2) << 1) | o;
        break;
      }
    }
  }
  if (G) *guard = min(*guard, g_);
  return (stack & 1u) != 0;
}

// ordered-uint64 view of a non-negative double; NaN -> +inf pattern
__device__ __forceinline__ unsigned long long resid_bits(double a) {
  if (!(a == a)) a = __longlong_as_double(0x7ff0000000000000LL);
  return (unsigned long long)__double_as_longlong(a);
}

// warp max of u64 keys with two 32-bit redux ops, result valid on all lanes
__device__ __forceinline__ unsigned long long warp_max_u64(unsigned long long v) {
  const uint32_t hi = (uint32_t)(v >> 32), lo = (uint32_t)v;
  const uint32_t mhi = __reduce_max_sync(0xffffffffu, hi);
  const uint32_t mlo = __reduce_max_sync(0xffffffffu, (hi == mhi) ? lo : 0u);
  return ((unsigned long long)mhi << 32) | mlo;
}

// Fold one warp's keys into a shared-memory running maximum. The slot is read first (one broadcast
// LDS) and the redux + atomic run only when some lane would raise it: after the first few tiles of a
// persistent CTA the slot already holds (nearly) the CTA's maximum and the common case is 5 instructions.
// A stale read can only cause a redundant atomicMax, never a lost update. All 32 lanes must call.
__device__ __forceinline__ void report_resid(unsigned long long* slot, unsigned long long key) {
  const unsigned long long cur = *reinterpret_cast<volatile unsigned long long*>(slot);
  if (__any_sync(0xffffffffu, key > cur)) {
    const unsigned long long wmax = warp_max_u64(key);
    if ((threadIdx.x & 31) == 0) atomicMax(slot, wmax);
  }
}

}  // namespace OLRT_NS
