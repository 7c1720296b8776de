// trace_bwd.cuh -- the adjoint (backward) kernel of the fused trace: what torch autograd computes through
//   Surface.trace                      optiland/surfaces/standard_surface.py:186-233
// on the reference's torch backend, for one ray per thread, surface by surface in REVERSE order.
//
// Checkpoints: the forward launch records the post-interaction global state of every surface
// (Surface._record, standard_surface.py:266-286); record[s-1] is exactly the state that enters
// surface s, so this kernel never re-traces the system: per surface it loads the 7 entry values,
// recomputes that one surface keeping the intermediates in registers, and pulls the incoming
// gradients (d loss / d state after surface s) back to the entry state and to the surface's
// prescription entries. Parameter gradients: per-thread partials -> warp shuffles -> one
// shared-memory atomic per warp -> one global atomicAdd per CTA and entry at the end.
//
// Newton-Raphson surfaces are differentiated the way autograd sees them: the K batch-uniform updates
// t <- t - f/f' actually taken in the forward are unrolled and each is differentiated exactly (with the
// second derivatives of the sag), NOT replaced by the implicit-function gradient of the converged
// root -- the two differ by O(residual), more than the 1e-8 gradient bar when tol = 1e-6.
//
// Differentiable here: plane, sphere / conic, even asphere, odd asphere, POLYNOMIAL (x^i y^j, polynomial.py:93-142) and ZERNIKE
// surfaces on ONE frame that may be decentred AND tilted (coordinate_system.py:73-107: rz, ry, rx), with any
// aperture program, absorption, refraction and reflection; gradients w.r.t. the frame's x y z and the cosines /
// sines of its three angles (the caller chains d cos / d angle: optimization/variable/tilt.py:59), radius, conic,
// even-asphere / polynomial coefficients (polynomial_coeff.py:72) and the input rays x y z L M N intensity opd.
// Anything else returns OLRT_E_UNSUPPORTED on the host side (no silent approximation).
#pragma once
#include "trace_fwd.cuh"

namespace OLRT_NS {

#define OLRT_BWD_MAXC 8  // even-asphere coefficients with a gradient slot in registers (reduced kernel, trace_bx.cuh)
#define OLRT_BWD_SLOTS (5 + OLRT_BWD_MAXC)  // frame x y z, radius, conic, coefficients (reduced kernel)
#define OLRT_BWG_MAXC 32   // this kernel: coefficient gradient slots per surface (a Zernike term takes two: c and c N)
#define OLRT_BWG_DEFC 16   // ... of them in shared memory unless olrt_sys.kernel_hint asks for more (bits 8-15)
#define OLRT_BWG_MAXM 12   // largest |m| of a Zernike term

template <typename R>
struct BwdParams {
  const R* in[8];
  const R* record;    // [S][8][N] forward records, slot s == surface s
  const R* gout[8];   // d loss / d final state (entries may be null)
  const R* grecord;   // null or [S][8][N]: d loss / d recorded state of every surface
  R* gin[8];          // d loss / d x y z L M N intensity opd_in (entries may be null)
  R* gpresc;          // [n_presc][presc_stride], atomically accumulated
  const R* presc;
  const int32_t* ops;
  long long n_rays;
  long long rays_per_presc;
  int n_surfaces, n_wavelengths, n_presc, presc_stride, ops_len;
  int cslots;  // general kernel: coefficient gradient slots per surface in shared memory (<= OLRT_BWG_MAXC)
  R wavelengths[OLRT_MAX_WAVELENGTHS];
  R four_pi;
};

// The adjoint has no bit-parity contract (bar: 1e-8): unguarded refinement sequences, no slow paths.
template <typename R> __device__ __forceinline__ R bdiv(R a, R b) { return m_div<MM_FAST>(a, b); }
template <typename R> __device__ __forceinline__ R bsqrt(R a) { return m_sqrt<MM_FAST>(a); }

template <typename R> __device__ __forceinline__ R brcp(R a);
template <> __device__ __forceinline__ double brcp<double>(double a) {  // one cubic Newton step: ~2^-60 relative
  double r = rcp_seed(a);
  double e = __fma_rn(-a, r, 1.0);
  e = __fma_rn(e, e, e);
  return __fma_rn(r, e, r);
}
template <> __device__ __forceinline__ float brcp<float>(float a) { return 1.0f / a; }

// reciprocal powers of the radius, once per surface
template <typename R>
struct BwdConst {
  R iR, iR2, iR3;
};

// sag S, slope factor G (zx = x G, zy = y G) and dG/dr2 of conic + even-asphere terms at (x, y)
template <typename R>
struct SagEval {
  R r2, s, is, i1s, S, G, Gp;  // radial part: zx = x G, zy = y G; is = 1/s, i1s = 1/(1+s)
  R Px, Py, Pxx, Pxy, Pyy;     // non-radial part of the SLOPES the reference codes: zx = x G + Px, zxx = G + 2 x^2 Gp + Pxx
  R dSx, dSy;                  // non-radial part of the true d S / d x, d y (= Px, Py except for Zernike terms, whose
                               // coded slopes leave the normalisation constant out, zernike.py:222-224 vs base.py:63-68)
};

// One Zernike term in Cartesian form: rho^|m| cos / sin(|m| phi) = Re / Im (xn + i yn)^|m| and R_n^m(rho) =
// rho^|m| Q(rho^2) (n - |m| is even), so Z = Q(u) H(xn, yn) with u = xn^2 + yn^2 is a polynomial in the normalised
// coordinates: value, first and second partials w.r.t. (xn, yn). rc: radial coefficients, highest power first;
// re / im: the powers (xn + i yn)^k.
template <typename R>
struct ZTerm { R T, Tx, Ty, Txx, Txy, Tyy; };

template <typename R>
__device__ __forceinline__ void zernike_term(int n, int m, const R* rc, R xn, R yn, R u, const R* re, const R* im,
                                             ZTerm<R>& z) {
  const int ma = (m < 0) ? -m : m;
  const int smax = ((n - ma) >> 1) + 1;
  R Q = R(0), Q1 = R(0), Q2 = R(0);  // Q, Q', Q'' by Horner in u
  for (int k = 0; k < smax; ++k) { Q2 = Q2 * u + R(2) * Q1; Q1 = Q1 * u + Q; Q = Q * u + rc[k]; }
  R H = R(1), Hx = R(0), Hy = R(0), Hxx = R(0), Hxy = R(0), Hyy = R(0);
  if (ma > 0) {
    const R fm = R(ma), fmm = R(ma * (ma - 1));
    const R r1 = re[ma - 1], i1 = im[ma - 1];
    const R r2 = (ma >= 2) ? re[ma - 2] : R(0), i2 = (ma >= 2) ? im[ma - 2] : R(0);
    if (m > 0) { H = re[ma]; Hx = fm * r1; Hy = -fm * i1; Hxx = fmm * r2; Hxy = -fmm * i2; Hyy = -fmm * r2; }
    else { H = im[ma]; Hx = fm * i1; Hy = fm * r1; Hxx = fmm * i2; Hxy = fmm * r2; Hyy = -fmm * i2; }
  }
  z.T = Q * H;
  z.Tx = R(2) * xn * Q1 * H + Q * Hx;
  z.Ty = R(2) * yn * Q1 * H + Q * Hy;
  z.Txx = R(4) * xn * xn * Q2 * H + R(2) * Q1 * H + R(4) * xn * Q1 * Hx + Q * Hxx;
  z.Txy = R(4) * xn * yn * Q2 * H + R(2) * xn * Q1 * Hy + R(2) * yn * Q1 * Hx + Q * Hxy;
  z.Tyy = R(4) * yn * yn * Q2 * H + R(2) * Q1 * H + R(4) * yn * Q1 * Hy + Q * Hyy;
}

template <typename R>
__device__ __forceinline__ void zernike_powers(R xn, R yn, R* re, R* im) {
  re[0] = R(1); im[0] = R(0);
  for (int k = 1; k <= OLRT_BWG_MAXM; ++k) {
    re[k] = re[k - 1] * xn - im[k - 1] * yn;
    im[k] = re[k - 1] * yn + im[k - 1] * xn;
  }
}

template <typename R>
__device__ __forceinline__ void sag_eval(const SurfParams<R>& sp, const BwdConst<R>& bc, bool plane, R x, R y,
                                         SagEval<R>& e) {
  e.r2 = x * x + y * y;
  e.Px = R(0); e.Py = R(0); e.Pxx = R(0); e.Pxy = R(0); e.Pyy = R(0); e.dSx = R(0); e.dSy = R(0);
  if (plane) { e.s = R(1); e.is = R(1); e.i1s = R(0.5); e.S = R(0); e.G = R(0); e.Gp = R(0); return; }
  const R c1 = sp.cc.c1;
  const R B = R(1) - c1 * e.r2 * bc.iR2;
  e.s = bsqrt(B);
  e.is = brcp(e.s);
  e.i1s = brcp(R(1) + e.s);
  e.S = e.r2 * bc.iR * e.i1s;
  e.G = bc.iR * e.is;
  e.Gp = R(0.5) * c1 * bc.iR3 * e.is * e.is * e.is;
  if (sp.geom == OLRT_GEOM_EVEN_ASPHERE) {
    R pw = R(1), pwm1 = R(0);  // r2^i, r2^(i-1)
    for (int i = 0; i < sp.ncoef; ++i) {
      const R Ci = sp.coef[i];
      const R k2 = R(2 * (i + 1));
      e.G += k2 * Ci * pw;
      e.Gp += k2 * R(i) * Ci * pwm1;
      pwm1 = pw;
      pw = pw * e.r2;
      e.S += Ci * pw;
    }
  } else if (sp.geom == OLRT_GEOM_POLYNOMIAL) {  // sum C[i][j] x^i y^j, polynomial.py:108-111, :129-136
    const int J = sp.aux, I = (J > 0) ? sp.ncoef / J : 0;
    R xi = R(1), xim1 = R(0), xim2 = R(0);  // x^i, x^(i-1), x^(i-2)
    for (int i = 0; i < I; ++i) {
      R yj = R(1), yjm1 = R(0), yjm2 = R(0);
      for (int j = 0; j < J; ++j) {
        const R C = sp.coef[i * J + j];
        e.S += C * xi * yj;
        e.Px += R(i) * C * xim1 * yj;
        e.Py += R(j) * C * xi * yjm1;
        e.Pxx += R(i * (i - 1)) * C * xim2 * yj;
        e.Pxy += R(i * j) * C * xim1 * yjm1;
        e.Pyy += R(j * (j - 1)) * C * xi * yjm2;
        yjm2 = yjm1; yjm1 = yj; yj = yj * y;
      }
      xim2 = xim1; xim1 = xi; xi = xi * x;
    }
    e.dSx = e.Px; e.dSy = e.Py;
  } else if (sp.geom == OLRT_GEOM_ODD_ASPHERE) {  // sum C_i r^(i+1), odd_asphere.py:85-88, :112-121: radial in r
    const R r = bsqrt(e.r2);
    const R ir = (r > R(0)) ? brcp(r) : R(0);     // the reference zeroes the non-finite slope terms at r = 0
    R pm3 = ir * ir * ir, pm1 = ir, pp1 = r;      // r^(i-3), r^(i-1), r^(i+1)
    for (int i = 0; i < sp.ncoef; ++i) {
      const R Ci = sp.coef[i];
      e.S += Ci * pp1;
      e.G += R(i + 1) * Ci * pm1;
      e.Gp += R(0.5) * R((i + 1) * (i - 1)) * Ci * pm3;
      pm3 = pm3 * r; pm1 = pm1 * r; pp1 = pp1 * r;
    }
  } else if (sp.geom == OLRT_GEOM_ZERNIKE) {      // zernike.py:133-224
    const R inr = brcp(sp.norm_radius);
    const R xn = x * inr, yn = y * inr, u = xn * xn + yn * yn;
    R re[OLRT_BWG_MAXM + 1], im[OLRT_BWG_MAXM + 1];
    zernike_powers(xn, yn, re, im);
    const int32_t* nm = sp.ops + sp.aux;
    const R* tc = sp.coef;
    for (int j = 0; j < sp.ncoef; ++j) {
      const int n = nm[2 * j], m = nm[2 * j + 1];
      const int ma = (m < 0) ? -m : m;
      const R c = tc[0], cN = tc[1];
      ZTerm<R> zt;
      zernike_term(n, m, tc + 2, xn, yn, u, re, im, zt);
      tc += 2 + ((n - ma) >> 1) + 1;
      e.S += cN * zt.T;
      e.dSx += cN * zt.Tx * inr; e.dSy += cN * zt.Ty * inr;
      if (c != R(0)) {                            // `if c == 0: continue`, zernike.py:218-219
        e.Px += c * zt.Tx * inr; e.Py += c * zt.Ty * inr;
        e.Pxx += c * zt.Txx * inr * inr; e.Pxy += c * zt.Txy * inr * inr; e.Pyy += c * zt.Tyy * inr * inr;
      }
    }
  }
}

// slopes of the sag at the evaluation point
template <typename R> __device__ __forceinline__ R sag_zx(const SagEval<R>& e, R x) { return x * e.G + e.Px; }
template <typename R> __device__ __forceinline__ R sag_zy(const SagEval<R>& e, R y) { return y * e.G + e.Py; }

// adjoint of (S, zx, zy)(x, y; R, k, C) given (gS, gzx, gzy); accumulates gx, gy and the parameter partials
template <typename R>
__device__ __forceinline__ void sag_adjoint(const SurfParams<R>& sp, const BwdConst<R>& bc, bool plane, R x, R y,
                                            const SagEval<R>& e, R gS, R gzx, R gzy, R& gx, R& gy, R& gR, R& gk,
                                            R* gC) {
  if (plane) return;
  const R q = gzx * x + gzy * y;  // gradient w.r.t. G
  const R Sx = x * e.G + e.dSx, Sy = y * e.G + e.dSy;   // d S / d x, d y
  gx += gS * Sx + gzx * (e.G + e.Pxx) + gzy * e.Pxy + q * e.Gp * (R(2) * x);
  gy += gS * Sy + gzy * (e.G + e.Pyy) + gzx * e.Pxy + q * e.Gp * (R(2) * y);
  const R c1 = sp.cc.c1, r2 = e.r2;
  const R ds_dR = c1 * r2 * bc.iR3 * e.is;
  const R ds_dk = R(-0.5) * r2 * bc.iR2 * e.is;
  const R dSc_ds = -r2 * bc.iR * e.i1s * e.i1s;
  const R dSc_dR = -r2 * bc.iR2 * e.i1s + dSc_ds * ds_dR;
  const R dg0_ds = -bc.iR * e.is * e.is;
  const R dg0_dR = -bc.iR2 * e.is + dg0_ds * ds_dR;
  gR += gS * dSc_dR + q * dg0_dR;
  gk += (gS * dSc_ds + q * dg0_ds) * ds_dk;
  if (sp.geom == OLRT_GEOM_EVEN_ASPHERE) {
    R pw = R(1);
    for (int i = 0; i < OLRT_BWG_MAXC; ++i) {
      if (i < sp.ncoef) {
        const R k2 = R(2 * (i + 1));
        const R pw1 = pw * r2;
        gC[i] += gS * pw1 + q * k2 * pw;
        pw = pw1;
      }
    }
  } else if (sp.geom == OLRT_GEOM_POLYNOMIAL) {
    const int J = sp.aux, I = (J > 0) ? sp.ncoef / J : 0;
    R xi = R(1), xim1 = R(0);
    for (int i = 0; i < I; ++i) {
      R yj = R(1), yjm1 = R(0);
      for (int j = 0; j < J; ++j) {
        gC[i * J + j] += gS * xi * yj + gzx * (R(i) * xim1 * yj) + gzy * (R(j) * xi * yjm1);
        yjm1 = yj; yj = yj * y;
      }
      xim1 = xi; xi = xi * x;
    }
  } else if (sp.geom == OLRT_GEOM_ODD_ASPHERE) {
    const R r = bsqrt(r2);
    const R ir = (r > R(0)) ? brcp(r) : R(0);
    R pm1 = ir, pp1 = r;
    for (int i = 0; i < OLRT_BWG_MAXC; ++i) {
      if (i < sp.ncoef) {
        gC[i] += gS * pp1 + q * R(i + 1) * pm1;
        pm1 = pm1 * r; pp1 = pp1 * r;
      }
    }
  } else if (sp.geom == OLRT_GEOM_ZERNIKE) {  // slot 2 j: d / d c_j (through the coded slopes), 2 j + 1: d / d (c_j N_j) (sag)
    const R inr = brcp(sp.norm_radius);
    const R xn = x * inr, yn = y * inr, u = xn * xn + yn * yn;
    R re[OLRT_BWG_MAXM + 1], im[OLRT_BWG_MAXM + 1];
    zernike_powers(xn, yn, re, im);
    const int32_t* nm = sp.ops + sp.aux;
    const R* tc = sp.coef;
    for (int j = 0; j < sp.ncoef && 2 * j + 1 < OLRT_BWG_MAXC; ++j) {
      const int n = nm[2 * j], m = nm[2 * j + 1];
      const int ma = (m < 0) ? -m : m;
      ZTerm<R> zt;
      zernike_term(n, m, tc + 2, xn, yn, u, re, im, zt);
      if (tc[0] != R(0)) gC[2 * j] += (gzx * zt.Tx + gzy * zt.Ty) * inr;
      gC[2 * j + 1] += gS * zt.T;
      tc += 2 + ((n - ma) >> 1) + 1;
    }
  }
}

// Adjoint of one plane rotation (a', b') = (a c - b s, a s + b c) -- RealRays.rotate_x / y / z, real_rays.py:90-130 --
// walked BACKWARDS: on entry (a, b) is the rotated pair and (ga, gb) the gradient w.r.t. it; on exit they are the pair
// before the rotation (the inverse of an orthogonal map is its transpose: no intermediate state is stored) and the
// gradient w.r.t. that; gc / gs collect d loss / d cos, d sin.
template <typename R>
__device__ __forceinline__ void rot_adjoint(R c, R s, R& a, R& b, R& ga, R& gb, R& gc, R& gs) {
  const R a0 = a * c + b * s, b0 = b * c - a * s;
  gc += a0 * ga + b0 * gb;
  gs += a0 * gb - b0 * ga;
  const R ga0 = c * ga + s * gb, gb0 = c * gb - s * ga;
  a = a0; b = b0; ga = ga0; gb = gb0;
}

// Adjoint of frame_globalize (trace_common.cuh): (P, D) is the GLOBAL state after the frame, g* the gradient w.r.t.
// it; on exit both refer to the local state before it. gF[9]: d loss / d (x y z cz sz cy sy cx sx) of the frame.
template <typename R>
__device__ __forceinline__ void frame_globalize_adjoint(const R* f, int fl, R& x, R& y, R& z, R& L, R& M, R& N,
                                                        R& gx, R& gy, R& gz, R& gL, R& gM, R& gN, R* gF) {
  gF[0] += gx; gF[1] += gy; gF[2] += gz;
  x -= f[0]; y -= f[1]; z -= f[2];
  if (fl & 1) {  // rotate_z(rz): pairs (x, y), (L, M)
    rot_adjoint(f[3], f[4], x, y, gx, gy, gF[3], gF[4]);
    rot_adjoint(f[3], f[4], L, M, gL, gM, gF[3], gF[4]);
  }
  if (fl & 2) {  // rotate_y(ry): z' = z c - x s, x' = z s + x c: pairs (z, x), (N, L)
    rot_adjoint(f[5], f[6], z, x, gz, gx, gF[5], gF[6]);
    rot_adjoint(f[5], f[6], N, L, gN, gL, gF[5], gF[6]);
  }
  if (fl & 4) {  // rotate_x(rx): pairs (y, z), (M, N)
    rot_adjoint(f[7], f[8], y, z, gy, gz, gF[7], gF[8]);
    rot_adjoint(f[7], f[8], M, N, gM, gN, gF[7], gF[8]);
  }
}

// Adjoint of frame_localize: (P, D) is the LOCAL state after the frame (translate by -T, rotate_z(-rz), rotate_y(-ry),
// rotate_x(-rx)); on exit the global state before it. The rotations use sin(-a) = -sin a: the sine gradients flip sign.
template <typename R>
__device__ __forceinline__ void frame_localize_adjoint(const R* f, int fl, R& x, R& y, R& z, R& L, R& M, R& N,
                                                       R& gx, R& gy, R& gz, R& gL, R& gM, R& gN, R* gF) {
  R gs;
  if (fl & 4) {
    gs = R(0);
    rot_adjoint(f[7], -f[8], y, z, gy, gz, gF[7], gs);
    rot_adjoint(f[7], -f[8], M, N, gM, gN, gF[7], gs);
    gF[8] -= gs;
  }
  if (fl & 2) {
    gs = R(0);
    rot_adjoint(f[5], -f[6], z, x, gz, gx, gF[5], gs);
    rot_adjoint(f[5], -f[6], N, L, gN, gL, gF[5], gs);
    gF[6] -= gs;
  }
  if (fl & 1) {
    gs = R(0);
    rot_adjoint(f[3], -f[4], x, y, gx, gy, gF[3], gs);
    rot_adjoint(f[3], -f[4], L, M, gL, gM, gF[3], gs);
    gF[4] -= gs;
  }
  gF[0] -= gx; gF[1] -= gy; gF[2] -= gz;
  x += f[0]; y += f[1]; z += f[2];
}

// closed-form conic intersection in the half-b form (trace_common.cuh); also returns a, hb for the adjoint
template <typename R>
__device__ __forceinline__ R conic_t(const SurfParams<R>& sp, R x, R y, R z, R L, R M, R N, R& a, R& hb) {
  const R Rr = sp.cc.radius, k = sp.cc.k;
  a = k * N * N + L * L + M * M + N * N;
  hb = k * N * z + L * x + M * y - N * Rr + N * z;
  const R c = k * z * z - R(2) * Rr * z + x * x + y * y + z * z;
  const R sd = bsqrt(hb * hb - a * c);
  const R ia = brcp(a);
  const R t1 = (-hb + sd) * ia, t2 = (-hb - sd) * ia;
  R t = (r_abs(z + t1 * N) <= r_abs(z + t2 * N)) ? t1 : t2;
  if (a == R(0)) t = -c / (R(2) * hb);
  return t;
}

// 4- or 8-byte asynchronous global -> shared copy (one element of a checkpoint), cp.async group helpers
template <typename R>
__device__ __forceinline__ void cp_async_elem(R* dst_shared, const R* src_global) {
  const unsigned d = (unsigned)__cvta_generic_to_shared(dst_shared);
  if (sizeof(R) == 8) asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(d), "l"(src_global) : "memory");
  else asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(d), "l"(src_global) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N_> __device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;" ::"n"(N_) : "memory");
}

template <typename R>
__device__ __forceinline__ R warp_sum(R v) {
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// one surface, reverse mode. Entry state e* (global), incoming grads g* (w.r.t. the state after the surface,
// global) are replaced by the grads w.r.t. the entry state. gi: intensity, gopd: opd (passes through).
// gF[9]: d loss / d (frame x y z cz sz cy sy cx sx); gC[OLRT_BWG_MAXC]: coefficients.
template <typename R>
__device__ __forceinline__ void bwd_surface(const SurfParams<R>& sp, int W, int widx, R ex, R ey, R ez, R eL, R eM,
                                            R eN, R ei, R& gx, R& gy, R& gz, R& gL, R& gM, R& gN, R& gi, R gopd,
                                            R* gF, R& gR, R& gk, R* gC) {
  const int32_t* op = sp.op;
  const bool plane = (sp.geom == OLRT_GEOM_PLANE);
  const int flags = op[OLRT_OP_FLAGS];
  const R* f = sp.frames;
  const int fl = op[OLRT_OP_FRAMEFLAGS] & 7;
  // ---- forward recompute (local frame): CoordinateSystem.localize, coordinate_system.py:73-88
  RayState<R> loc;
  loc.x = ex; loc.y = ey; loc.z = ez; loc.L = eL; loc.M = eM; loc.N = eN; loc.i = ei; loc.opd = R(0);
  frame_localize(loc, f, fl);
  const R x = loc.x, y = loc.y, z = loc.z;
  const R L = loc.L, M = loc.M, N = loc.N;
  BwdConst<R> bc;
  bc.iR = plane ? R(0) : brcp(sp.cc.radius);
  bc.iR2 = bc.iR * bc.iR;
  bc.iR3 = bc.iR2 * bc.iR;
  const R iN = plane ? brcp(N) : R(0);
  R a = R(0), hb = R(0), t0;
  if (plane) t0 = -z * iN;
  else t0 = conic_t(sp, x, y, z, L, M, N, a, hb);
  const int K = (sp.geom >= OLRT_GEOM_EVEN_ASPHERE) ? op[OLRT_OP_NR_ITERS] : 0;
  R t = t0;
  for (int j = 0; j < K; ++j) {
    SagEval<R> e;
    const R qx = x + t * L, qy = y + t * M, qz = z + t * N;
    sag_eval(sp, bc, false, qx, qy, e);
    const R h = sag_zx(e, qx) * L + sag_zy(e, qy) * M - N;
    t = t - (e.S - qz) * brcp(h);
  }
  const R Qx = x + t * L, Qy = y + t * M, Qz = z + t * N;
  SagEval<R> eq;
  sag_eval(sp, bc, plane, Qx, Qy, eq);
  R nx, ny, nz, imag = R(1);
  const R zx = sag_zx(eq, Qx), zy = sag_zy(eq, Qy);
  if (plane) { nx = R(0); ny = R(0); nz = R(1); }
  else { imag = brcp(bsqrt(zx * zx + zy * zy + R(1))); nx = zx * imag; ny = zy * imag; nz = -imag; }
  const R dot = L * nx + M * ny + N * nz;
  const R sg = r_sign(dot), c = r_abs(dot);
  const R ax = nx * sg, ay = ny * sg, az = nz * sg;  // aligned normal
  const bool mirror = (op[OLRT_OP_INTERACT] == 1);
  const R u = sp.u, u2 = sp.u2;
  const R root = mirror ? R(0) : bsqrt(R(1) - u2 * (R(1) - c * c));
  const R w = root - u * c;
  // ---- adjoint of globalize (coordinate_system.py:90-107): needs the state it was applied to. Rebuild the local
  // exit state (position Q, new direction), push it through the frame, and walk back
  if (fl != 0) {
    RayState<R> o;
    o.x = Qx; o.y = Qy; o.z = Qz;
    if (mirror) { o.L = L - R(2) * c * ax; o.M = M - R(2) * c * ay; o.N = N - R(2) * c * az; }
    else { o.L = u * L + ax * w; o.M = u * M + ay * w; o.N = u * N + az * w; }
    o.i = R(0); o.opd = R(0);
    frame_globalize(o, f, fl);
    frame_globalize_adjoint(f, fl, o.x, o.y, o.z, o.L, o.M, o.N, gx, gy, gz, gL, gM, gN, gF);
  } else {  // out = local + T
    gF[0] += gx; gF[1] += gy; gF[2] += gz;
  }
  // ---- interaction
  R gDx, gDy, gDz, gax, gay, gaz, gc;
  if (mirror) {  // D' = D - 2 c a
    gDx = gL; gDy = gM; gDz = gN;
    gc = R(-2) * (ax * gL + ay * gM + az * gN);
    gax = R(-2) * c * gL; gay = R(-2) * c * gM; gaz = R(-2) * c * gN;
  } else {  // D' = u D + a (root - u c)
    gDx = u * gL; gDy = u * gM; gDz = u * gN;
    gax = w * gL; gay = w * gM; gaz = w * gN;
    const R gw = ax * gL + ay * gM + az * gN;
    gc = -u * gw + gw * brcp(root) * (u2 * c);
  }
  const R gdot = sg * gc;
  R gnx = sg * gax + gdot * L, gny = sg * gay + gdot * M, gnz = sg * gaz + gdot * N;
  gDx += gdot * nx; gDy += gdot * ny; gDz += gdot * nz;
  // ---- normal = (zx, zy, -1) / mag
  R gzx = R(0), gzy = R(0);
  if (!plane) {
    const R ng = nx * gnx + ny * gny + nz * gnz;
    gzx = (gnx - nx * ng) * imag;
    gzy = (gny - ny * ng) * imag;
  }
  R gQx = gx, gQy = gy, gQz = gz;  // position after propagation (clip and interaction do not move it)
  sag_adjoint(sp, bc, plane, Qx, Qy, eq, R(0), gzx, gzy, gQx, gQy, gR, gk, gC);
  // ---- clip (where(~inside, 0, i)), absorption, opd
  R gt = gopd * ((t * sp.n1 < R(0)) ? -sp.n1 : sp.n1);
  const int aoff = op[OLRT_OP_APER_OFF];
  if (aoff >= 0) {
    if (!aperture_contains(sp.ops + aoff, op[OLRT_OP_APER_LEN], sp.aper, Qx, Qy)) gi = R(0);
  }
  if (flags & OLRT_F_ABSORB) {
    const R alpha = sp.alpha;  // 4 pi k / lambda, homogeneous.py:49
    const R E = r_exp((-alpha) * t * R(1e3));
    gt += gi * ei * E * (-alpha * R(1e3));  // i' = i E
    gi = gi * E;
  }
  // ---- Q = P + t D
  R gPx = gQx, gPy = gQy, gPz = gQz;
  gDx += t * gQx; gDy += t * gQy; gDz += t * gQz;
  R gtj = gt + L * gQx + M * gQy + N * gQz;
  // ---- Newton updates in reverse: t_{j+1} = t_j - f_j / h_j
  for (int j = K - 1; j >= 0; --j) {
    R tj = t0;
    for (int q = 0; q < j; ++q) {  // recompute t_j (K is 1-3 in practice)
      SagEval<R> e;
      const R qx = x + tj * L, qy = y + tj * M, qz = z + tj * N;
      sag_eval(sp, bc, false, qx, qy, e);
      tj = tj - (e.S - qz) * brcp(sag_zx(e, qx) * L + sag_zy(e, qy) * M - N);
    }
    SagEval<R> e;
    const R qx = x + tj * L, qy = y + tj * M, qz = z + tj * N;
    sag_eval(sp, bc, false, qx, qy, e);
    const R sx = sag_zx(e, qx), sy = sag_zy(e, qy);
    const R fj = e.S - qz, ih = brcp(sx * L + sy * M - N);
    const R gf = -gtj * ih, gh = gtj * fj * ih * ih;
    gDx += gh * sx; gDy += gh * sy; gDz -= gh;
    R gqx = R(0), gqy = R(0);
    const R gqz = -gf;
    sag_adjoint(sp, bc, false, qx, qy, e, gf, gh * L, gh * M, gqx, gqy, gR, gk, gC);
    gPx += gqx; gPy += gqy; gPz += gqz;
    gDx += tj * gqx; gDy += tj * gqy; gDz += tj * gqz;
    gtj = gtj + L * gqx + M * gqy + N * gqz;
  }
  // ---- t0: plane or conic (implicit differentiation of the quadratic the root satisfies)
  if (plane) {
    gPz += -gtj * iN;
    gDz += -gtj * t0 * iN;
  } else {
    const R Rr = sp.cc.radius, k = sp.cc.k;
    const R gcc = -gtj * brcp(R(2) * (a * t0 + hb));
    const R ga = gcc * t0 * t0, ghb = gcc * R(2) * t0;
    gk += ga * N * N + ghb * N * z + gcc * z * z;
    gR += -ghb * N - R(2) * gcc * z;
    gDx += ga * R(2) * L + ghb * x;
    gDy += ga * R(2) * M + ghb * y;
    gDz += ga * R(2) * N * (k + R(1)) + ghb * (k * z - Rr + z);
    gPx += ghb * L + R(2) * gcc * x;
    gPy += ghb * M + R(2) * gcc * y;
    gPz += ghb * (k * N + N) + gcc * (R(2) * k * z - R(2) * Rr + R(2) * z);
  }
  // ---- localize: the entry state went through translate(-T) and the three rotations
  gx = gPx; gy = gPy; gz = gPz;
  gL = gDx; gM = gDy; gN = gDz;
  if (fl != 0) {
    R lx = x, ly = y, lz = z, lL = L, lM = M, lN = N;
    frame_localize_adjoint(f, fl, lx, ly, lz, lL, lM, lN, gx, gy, gz, gL, gM, gN, gF);
  } else {  // P = entry - T
    gF[0] -= gPx; gF[1] -= gPy; gF[2] -= gPz;
  }
}


// sum the 32 lane-private partials of every (surface, slot) and add them to the prescription gradient; clears
// the accumulators. Called by the whole CTA after a __syncthreads().
template <typename R>
__device__ __forceinline__ void bwd_flush(R* s_acc, const int32_t* s_ops, int S, int W, int slots, R* gpresc) {
  for (int q = threadIdx.x; q < S * slots; q += blockDim.x) {
    const int s = q / slots, j = q - s * slots;
    R* a = s_acc + (size_t)q * 32;
    R v = R(0);
    for (int l = 0; l < 32; ++l) { v += a[(l + threadIdx.x) & 31]; }   // rotated start: no bank conflicts
    for (int l = 0; l < 32; ++l) a[l] = R(0);
    if (v == R(0)) continue;
    const int32_t* op = s_ops + s * OLRT_OPS_STRIDE;
    const int doff = op[OLRT_OP_DOFF];
    const int goff = doff + op[OLRT_OP_NFRAMES] * OLRT_FRAME_REALS;
    int off;
    if (j < OLRT_FRAME_REALS) off = doff + j;                                  // x y z cz sz cy sy cx sx
    else if (j < OLRT_FRAME_REALS + 2) off = goff + (j - OLRT_FRAME_REALS);    // radius, conic
    else {                                                                     // coefficients
      const int ci = j - OLRT_FRAME_REALS - 2;
      off = goff + OLRT_GEOM_REALS + 3 * W;
      if (op[OLRT_OP_GEOM] == OLRT_GEOM_ZERNIKE) {  // per term: c, c N, radial coefficients; slot 2 t -> c, 2 t + 1 -> c N
        const int32_t* nm = s_ops + op[OLRT_OP_AUX];
        for (int t = 0; t < (ci >> 1); ++t) {
          const int n = nm[2 * t], m = nm[2 * t + 1];
          off += 2 + ((n - (m < 0 ? -m : m)) >> 1) + 1;
        }
        off += ci & 1;
      } else {
        off += ci;
      }
    }
    atomicAdd(gpresc + off, v);
  }
}

template <typename R>
__device__ __forceinline__ bool bwd_supported(const int32_t* s_ops, int s, int cslots) {
  const int32_t* op = s_ops + s * OLRT_OPS_STRIDE;
  const int geom = op[OLRT_OP_GEOM], flags = op[OLRT_OP_FLAGS];
  if (op[OLRT_OP_NFRAMES] != 1) return false;                                 // one frame, tilted or not
  if (flags & (OLRT_F_COATING | OLRT_F_THIN_LENS | OLRT_F_RENORM)) return false;
  if (geom != OLRT_GEOM_PLANE) {
    if (geom != OLRT_GEOM_STANDARD && geom != OLRT_GEOM_EVEN_ASPHERE && geom != OLRT_GEOM_POLYNOMIAL &&
        geom != OLRT_GEOM_ODD_ASPHERE && geom != OLRT_GEOM_ZERNIKE)
      return false;
    if (flags & OLRT_F_RADIUS_INF) return false;
    const int need = (geom == OLRT_GEOM_STANDARD) ? 0 : (geom == OLRT_GEOM_ZERNIKE ? 2 : 1) * op[OLRT_OP_NCOEF];
    if (need > cslots) return false;
    if (geom == OLRT_GEOM_ZERNIKE) {
      const int32_t* nm = s_ops + op[OLRT_OP_AUX];
      for (int t = 0; t < op[OLRT_OP_NCOEF]; ++t) {
        const int m = nm[2 * t + 1];
        if ((m < 0 ? -m : m) > OLRT_BWG_MAXM) return false;
      }
    }
  }
  return op[OLRT_OP_RECORD] == s;
}

template <typename R>
__global__ void __launch_bounds__(256, 2) trace_bwd_kernel(const BwdParams<R> P, uint32_t* status) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int S = P.n_surfaces, W = P.n_wavelengths;
  R* s_presc = reinterpret_cast<R*>(smem_raw);
  const int presc_pad = (P.presc_stride + 1) & ~1;
  R* s_der = s_presc + presc_pad;
  const int der_pad = (S * der_stride(W) + 1) & ~1;
  // d loss / d presc, accumulated per LANE: s_acc[(s * OLRT_BWD_SLOTS + j) * 32 + lane], j = frame x y z, radius,
  // conic, coefficients. A lane adds its own partial with one shared-memory atomic (8 warps share a slot, no
  // two lanes do), so there is no per-ray warp reduction; the 32 lanes are summed once per prescription.
  R* s_acc = s_der + der_pad;
  // entry states (checkpoints) of the next surface, staged by cp.async while the current one is differentiated:
  // s_ck[(buf * 7 + q) * blockDim + tid]
  const int slots = OLRT_FRAME_REALS + 2 + P.cslots;
  R* s_ck = s_acc + (size_t)S * slots * 32;
  int32_t* s_ops = reinterpret_cast<int32_t*>(s_ck + 2 * 7 * blockDim.x);
  __shared__ int s_bad;

  for (int q = threadIdx.x; q < P.ops_len; q += blockDim.x) s_ops[q] = P.ops[q];
  for (int q = threadIdx.x; q < S * slots * 32; q += blockDim.x) s_acc[q] = R(0);
  if (threadIdx.x == 0) s_bad = 0;
  __syncthreads();
  if (threadIdx.x < S && !bwd_supported<R>(s_ops, threadIdx.x, P.cslots)) s_bad = 1;
  __syncthreads();
  if (s_bad) {
    if (threadIdx.x == 0 && status != nullptr) atomicOr(status, OLRT_ST_BAD_HINT);
    return;
  }
  const long long N = P.n_rays;
  const int lane = threadIdx.x & 31;
  int cur_p = -1;
  const long long n_tiles = (N + blockDim.x - 1) / blockDim.x;
  for (long long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    const long long base = tile * blockDim.x;
    const int p = (P.n_presc > 1) ? (int)(base / P.rays_per_presc) : 0;
    if (p != cur_p) {
      __syncthreads();
      if (cur_p >= 0) {  // flush the previous prescription's gradients
        bwd_flush<R>(s_acc, s_ops, S, W, slots, P.gpresc + (long long)cur_p * P.presc_stride);
        __syncthreads();
      }
      const R* src = P.presc + (long long)p * P.presc_stride;
      for (int q = threadIdx.x; q < P.presc_stride; q += blockDim.x) s_presc[q] = src[q];
      cur_p = p;
      __syncthreads();
      if (threadIdx.x < S) derive_surface<R>(s_ops, s_presc, s_der, threadIdx.x, W, P.wavelengths, P.four_pi);
      __syncthreads();
    }
    const long long item = base + threadIdx.x;
    const bool valid = item < N;
    const long long idx = valid ? item : (N - 1);
    int widx = 0;
    if (W > 1) {
      const R wl = P.in[7][idx];
      for (int q = 0; q < W; ++q) if (wl == P.wavelengths[q]) widx = q;
    }
    auto ld = [&](const R* ptr) { return (ptr != nullptr && valid) ? ptr[idx] : R(0); };
    R gx = ld(P.gout[0]), gy = ld(P.gout[1]), gz = ld(P.gout[2]);
    R gL = ld(P.gout[3]), gM = ld(P.gout[4]), gN = ld(P.gout[5]);
    R gi = ld(P.gout[6]), gopd = ld(P.gout[7]);
    // entry state of surface s = record[s-1] (the launch state for s == 0): staged one surface ahead
    auto stage = [&](int s_) {
      R* dst = s_ck + (size_t)((s_ & 1) * 7) * blockDim.x + threadIdx.x;
      if (s_ == 0) {
#pragma unroll
        for (int q = 0; q < 7; ++q) cp_async_elem(dst + (size_t)q * blockDim.x, P.in[q] + idx);
      } else {
        const R* rc = P.record + (long long)(s_ - 1) * 8 * N + idx;
#pragma unroll
        for (int q = 0; q < 7; ++q) cp_async_elem(dst + (size_t)q * blockDim.x, rc + (long long)q * N);
      }
      cp_async_commit();
    };
    stage(S - 1);
    for (int s = S - 1; s >= 0; --s) {
      if (s > 0) { stage(s - 1); cp_async_wait<1>(); } else { cp_async_wait<0>(); }
      if (P.grecord != nullptr && valid) {
        const R* g = P.grecord + (long long)s * 8 * N + idx;
        gx += g[0]; gy += g[N]; gz += g[2 * N]; gL += g[3 * N]; gM += g[4 * N]; gN += g[5 * N];
        gi += g[6 * N]; gopd += g[7 * N];
      }
      const R* ck = s_ck + (size_t)((s & 1) * 7) * blockDim.x + threadIdx.x;
      const R ex = ck[0], ey = ck[blockDim.x], ez = ck[2 * blockDim.x], eL = ck[3 * blockDim.x],
              eM = ck[4 * blockDim.x], eN = ck[5 * blockDim.x], ei = ck[6 * blockDim.x];
      SurfParams<R> sp;
      load_surface<R, false>(sp, s_ops, s_presc, s_der, s, W, widx);
      R gR = R(0), gk = R(0);
      R gF[OLRT_FRAME_REALS];
      R gC[OLRT_BWG_MAXC];
#pragma unroll
      for (int i = 0; i < OLRT_FRAME_REALS; ++i) gF[i] = R(0);
#pragma unroll
      for (int i = 0; i < OLRT_BWG_MAXC; ++i) gC[i] = R(0);
      bwd_surface<R>(sp, W, widx, ex, ey, ez, eL, eM, eN, ei, gx, gy, gz, gL, gM, gN, gi, gopd, gF, gR, gk, gC);
      if (valid) {  // parameter partials: one lane-private shared atomic each
        R* acc = s_acc + (size_t)s * slots * 32 + lane;
        atomicAdd(acc + 0 * 32, gF[0]); atomicAdd(acc + 1 * 32, gF[1]); atomicAdd(acc + 2 * 32, gF[2]);
        if (sp.op[OLRT_OP_FRAMEFLAGS] & 7) {
#pragma unroll
          for (int i = 3; i < OLRT_FRAME_REALS; ++i) atomicAdd(acc + i * 32, gF[i]);
        }
        if (sp.geom != OLRT_GEOM_PLANE) {
          atomicAdd(acc + OLRT_FRAME_REALS * 32, gR); atomicAdd(acc + (OLRT_FRAME_REALS + 1) * 32, gk);
          if (sp.geom != OLRT_GEOM_STANDARD) {
            const int nc = (sp.geom == OLRT_GEOM_ZERNIKE) ? 2 * sp.ncoef : sp.ncoef;
            for (int i = 0; i < OLRT_BWG_MAXC; ++i)
              if (i < nc && gC[i] != R(0)) atomicAdd(acc + (OLRT_FRAME_REALS + 2 + i) * 32, gC[i]);
          }
        }
      }
    }
    if (valid) {
      if (P.gin[0]) P.gin[0][idx] = gx;
      if (P.gin[1]) P.gin[1][idx] = gy;
      if (P.gin[2]) P.gin[2][idx] = gz;
      if (P.gin[3]) P.gin[3][idx] = gL;
      if (P.gin[4]) P.gin[4][idx] = gM;
      if (P.gin[5]) P.gin[5][idx] = gN;
      if (P.gin[6]) P.gin[6][idx] = gi;
      if (P.gin[7]) P.gin[7][idx] = gopd;
    }
  }
  __syncthreads();
  if (cur_p >= 0) bwd_flush<R>(s_acc, s_ops, S, W, slots, P.gpresc + (long long)cur_p * P.presc_stride);
}

template <typename R>
int launch_bwd(const olrt_sys* sys, const R* presc, const int32_t* ops, long long n_rays, const R* const in[8],
               const R* record, const R* const gout[8], const R* grecord, R* const gin[8], R* gpresc,
               uint32_t* status, cudaStream_t stream) {
  BwdParams<R> P;
  for (int q = 0; q < 8; ++q) { P.in[q] = in[q]; P.gout[q] = gout ? gout[q] : nullptr; P.gin[q] = gin ? gin[q] : nullptr; }
  P.record = record; P.grecord = grecord; P.gpresc = gpresc; P.presc = presc; P.ops = ops; P.n_rays = n_rays;
  P.rays_per_presc = sys->rays_per_presc > 0 ? sys->rays_per_presc : n_rays;
  P.n_surfaces = sys->n_surfaces; P.n_wavelengths = sys->n_wavelengths; P.n_presc = sys->n_presc;
  P.presc_stride = sys->presc_stride; P.ops_len = sys->ops_len;
  for (int q = 0; q < OLRT_MAX_WAVELENGTHS; ++q) P.wavelengths[q] = (R)sys->wavelengths[q];
  P.four_pi = (R)(4.0 * 3.141592653589793);
  // coefficient gradient slots per surface: bits 8-15 of the hint (the caller's maximum over the surfaces), else 16
  P.cslots = (sys->kernel_hint >> 8) & 0xFF;
  if (P.cslots == 0) P.cslots = OLRT_BWG_DEFC;
  if (P.cslots > OLRT_BWG_MAXC) return OLRT_E_UNSUPPORTED;
  int block = 256;
  if (sys->n_presc > 1) {
    block = 0;
    for (int b = 256; b >= 32; b >>= 1) if (sys->rays_per_presc % b == 0) { block = b; break; }
    if (block == 0) return OLRT_E_ARG;
  }
  const size_t pp = (size_t)((sys->presc_stride + 1) & ~1);
  const size_t smem = pp * sizeof(R) +
                      (size_t)((sys->n_surfaces * der_stride(sys->n_wavelengths) + 1) & ~1) * sizeof(R) +
                      (size_t)sys->n_surfaces * (OLRT_FRAME_REALS + 2 + P.cslots) * 32 * sizeof(R) +
                      (size_t)2 * 7 * block * sizeof(R) +
                      (size_t)((sys->ops_len + 1) & ~1) * sizeof(int32_t);
  if (smem > 200 * 1024) return OLRT_E_LIMIT;
  if (smem > 48 * 1024)
    cudaFuncSetAttribute(trace_bwd_kernel<R>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  int per_sm = 0, dev = 0, sms = 148;
  cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, trace_bwd_kernel<R>, block, smem);
  if (per_sm < 1) per_sm = 1;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  const long long n_tiles = (n_rays + block - 1) / block;
  long long grid = (long long)sms * per_sm;
  if (grid > n_tiles) grid = n_tiles;
  trace_bwd_kernel<R><<<(unsigned)grid, block, smem, stream>>>(P, status);
  return cudaGetLastError() == cudaSuccess ? OLRT_OK : OLRT_E_LAUNCH;
}

}  // namespace OLRT_NS
