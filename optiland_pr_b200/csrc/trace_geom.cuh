// trace_geom.cuh -- sag f(x,y) and slopes (df/dx, df/dy) of the Newton-Raphson geometries.
// Each evaluator returns the sag and, when WANT_SLOPES, the un-normalised slopes that the
// reference's `_surface_normal` turns into (zx, zy, -1)/|.|.
#pragma once
#include "trace_common.cuh"

namespace OLRT_NS {

template <typename R>
struct SurfParams {      // one surface's view of the prescription staged in shared memory
  const int32_t* op;     // ops row
  const int32_t* ops;    // ops base (aperture programs, zernike index tables)
  const R* frames;       // NFRAMES x 9
  ConicConst<R> cc;      // radius, conic and the constants derived from them
  R tol, norm_radius;
  const R* media;        // n_pre[W], n_post[W], k_pre[W]
  R n1, u, u2, alpha;    // this ray's wavelength: n_pre, u = n_pre/n_post, u*u, 4 pi k_pre / lambda
  const R* coef;         // NCOEF reals
  const R* aper;         // aperture leaf parameters
  int geom, ncoef, aux;
  int nr_iters;          // K_s: Newton updates at this surface (ops word 6, or the ray's group in group mode)
  const R* ztab;         // tolerance build: this Zernike surface's merged tables in shared memory (zernike_merge), or null
  const int* zlen;
};

// Merged Zernike tables (tolerance build only). Every term is rho^|m| Q(u) {cos, sin}(|m| phi) with u = rho^2 and
// Q a polynomial in u, so the terms of one (|m|, cos / sin) pair add up to ONE polynomial in u:
//   Z = sum_m A_m(u) Re(w^m) + B_m(u) Im(w^m),   w = xn + i yn.
// One table per Zernike surface, built once per CTA and prescription: [set][m][cos / sin][power of u], set 0 weighted
// with c (the reference's slopes, zernike.py:176-232, leave the norm constant out), set 1 with c N (the sag).
#define OLRT_ZT_SURF 2   /* Zernike surfaces with a table (the rest keep the term loop) */
#define OLRT_ZT_M 9      /* |m| = 0 .. 8 */
#define OLRT_ZT_L 8      /* (n - |m|) / 2 <= 7 */
#define OLRT_ZT_REALS (2 * OLRT_ZT_M * 2 * OLRT_ZT_L)
#define OLRT_ZT_INTS (OLRT_ZT_M + 2)   /* table length per m, largest m, 1 when both sets coincide */

// can this surface's term list be merged? (ops words only: the same answer for every prescription of a batch)
__device__ __forceinline__ bool zernike_mergeable(const int32_t* nm, int nterms) {
  for (int j = 0; j < nterms; ++j) {
    const int n = nm[2 * j], m = nm[2 * j + 1];
    const int ma = (m < 0) ? -m : m;
    if (ma >= OLRT_ZT_M || ((n - ma) >> 1) >= OLRT_ZT_L) return false;
  }
  return nterms > 0;
}

// one thread builds one surface's tables (terms in list order: the same bits in every CTA of every launch)
template <typename R>
__device__ __forceinline__ void zernike_merge(const int32_t* nm, const R* tc, int nterms, R* tab, int* len) {
  for (int q = 0; q < OLRT_ZT_REALS; ++q) tab[q] = R(0);
  for (int q = 0; q < OLRT_ZT_INTS; ++q) len[q] = 0;
  int same = 1, mmax = 0;
  for (int j = 0; j < nterms; ++j) {
    const int n = nm[2 * j], m = nm[2 * j + 1];
    const int ma = (m < 0) ? -m : m;
    const int d = (n - ma) >> 1;
    const R c = tc[0], cN = tc[1];
    const R* rc = tc + 2;          // radial coefficients of rho^n, rho^(n-2), ..., rho^|m|
    tc += 2 + d + 1;
    if (c != cN) same = 0;
    R* t0 = tab + ((0 * OLRT_ZT_M + ma) * 2 + (m < 0 ? 1 : 0)) * OLRT_ZT_L;
    R* t1 = tab + ((1 * OLRT_ZT_M + ma) * 2 + (m < 0 ? 1 : 0)) * OLRT_ZT_L;
    for (int k = 0; k <= d; ++k) {
      t0[d - k] = t0[d - k] + c * rc[k];
      t1[d - k] = t1[d - k] + cN * rc[k];
    }
    if (d + 1 > len[ma]) len[ma] = d + 1;
    if (ma > mmax) mmax = ma;
  }
  for (int q = 0; q <= mmax; ++q) if (len[q] < 1) len[q] = 1;   // the evaluator seeds its Horner pass with entry L - 1
  len[OLRT_ZT_M] = mmax;
  len[OLRT_ZT_M + 1] = same;
}

template <typename R>
__device__ __forceinline__ R ipow(R b, int e) {  // b**e by repeated multiplication, e >= 0
  R p = R(1);
  for (int q = 0; q < e; ++q) p = p * b;
  return p;
}

// EvenAsphere.sag / _surface_normal         geometries/even_asphere.py:82-129
template <typename R, bool WANT_SLOPES, int MM>
__device__ __forceinline__ void even_asphere_eval(const SurfParams<R>& sp, R x, R y, R& sag, R& zx, R& zy) {
  const R r2 = x * x + y * y;
  R denom;
  conic_sag_denom<R, MM>(r2, sp.cc, sag, denom);
  if (WANT_SLOPES) { const Recip<R, MM> rd(denom); zx = rd.div(x); zy = rd.div(y); }
  R prev = R(1);        // r2**i
  PowChain<R> pw(r2);   // r2**(i+1), correctly rounded like np.power
  for (int i = 0; i < sp.ncoef; ++i) {
    const R Ci = sp.coef[i];
    if (WANT_SLOPES) {
      const R two_i1 = R(2 * (i + 1));
      zx = zx + ((two_i1 * x) * Ci) * prev;  // 2*(i+1)*x*Ci*r2**i
      zy = zy + ((two_i1 * y) * Ci) * prev;
    }
    sag = sag + Ci * pw.value();  // z + Ci*r2**(i+1)
    prev = pw.value();
    if (i + 1 < sp.ncoef) pw.next();
  }
}

// OddAsphere.sag / _surface_normal          geometries/odd_asphere.py:73-129
template <typename R, bool WANT_SLOPES, int MM>
__device__ __forceinline__ void odd_asphere_eval(const SurfParams<R>& sp, R x, R y, R& sag, R& zx, R& zy) {
  const R r2 = x * x + y * y;
  // (r can be exactly 0 on the axis: the unguarded square root would turn it into a NaN that the non-finite filter
  // below and the Newton loop's selects could hide -- this one keeps its guard in every mode)
  const R r = m_sqrt<(MM == MM_FAST ? MM_GUARD : MM)>(r2);
  R denom;
  conic_sag_denom<R, MM>(r2, sp.cc, sag, denom);
  if (WANT_SLOPES) { const Recip<R, MM> rd(denom); zx = rd.div(x); zy = rd.div(y); }
#ifdef OLRT_FX
  // tolerance build: sum_i C_i r^(i+1) = r H(r) and the slope factor G(r) = sum_{i>=1} (i+1) C_i r^(i-1) by Horner;
  // the C_0 term of the slopes, x C_0 / r, keeps the reference's non-finite filter (the only term that needs it)
  if (MM == MM_FAST) {
    const int nc = sp.ncoef;
    if (nc > 0) {
      R H = R(0), G = R(0), k = R(nc);
#pragma unroll 1
      for (int i = nc - 1; i >= 1; --i) {
        const R Ci = sp.coef[i];
        H = fma(H, r, Ci);
        if (WANT_SLOPES) { G = fma(G, r, k * Ci); k = k - R(1); }
      }
      const R C0 = sp.coef[0];
      H = fma(H, r, C0);
      sag = fma(H, r, sag);
      if (WANT_SLOPES) {
        const R rp = m_div<MM>(R(1), r);
        R xt = (x * C0) * rp, yt = (y * C0) * rp;
        if (!isfinite(xt)) xt = R(0);
        if (!isfinite(yt)) yt = R(0);
        zx = fma(x, G, zx + xt);
        zy = fma(y, G, zy + yt);
      }
    }
    return;
  }
#endif
  R pw = R(1);  // r**i
  for (int i = 0; i < sp.ncoef; ++i) {
    const R Ci = sp.coef[i];
    if (WANT_SLOPES) {
      // (i+1)*x*Ci*r**(i-1); non-finite terms are zeroed (odd_asphere.py:117-118)
      const R rp = (i == 0) ? m_div<MM>(R(1), r) : ipow(r, i - 1);
      R xt = ((R(i + 1) * x) * Ci) * rp;
      R yt = ((R(i + 1) * y) * Ci) * rp;
      if (!isfinite(xt)) xt = R(0);
      if (!isfinite(yt)) yt = R(0);
      zx = zx + xt;
      zy = zy + yt;
    }
    pw = pw * r;            // r**(i+1)
    sag = sag + Ci * pw;
  }
}

// PolynomialGeometry.sag / _surface_normal  geometries/polynomial.py:93-142
// coef = C[i][j] row-major, I = ncoef/J rows, J = aux columns
// STRAIGHT44 (the polynomial surface kind's own call sites only: the interpreter waits on instruction fetch, and the
// same lines inlined into every generic body cost the Zernike plate 2.5 %): the 4 x 4 table without loop control
template <typename R, bool WANT_SLOPES, int MM, bool STRAIGHT44 = false>
__device__ __forceinline__ void polynomial_eval(const SurfParams<R>& sp, R x, R y, R& sag, R& zx, R& zy) {
  const R r2 = x * x + y * y;
  R denom;
  conic_sag_denom<R, MM>(r2, sp.cc, sag, denom);
  if (WANT_SLOPES) { const Recip<R, MM> rd(denom); zx = rd.div(x); zy = rd.div(y); }
#ifdef OLRT_FX
  // tolerance build (-fmad=true TU): the same polynomial and partials by bivariate Horner -- per row i the sum
  // A_i(y) = sum_j C_ij y^j and its derivative, then P = sum_i x^i A_i(y), dP/dx, dP/dy by Horner in x: 2 FMAs per
  // coefficient + 3 per row instead of ~11 multiplications per coefficient, no integer division, no int -> double
  {
    const int J = sp.aux;
    R P = R(0), Px = R(0), Py = R(0);
    if (STRAIGHT44 && J == 4 && sp.ncoef == 16) {   // the 4 x 4 table (x^3 y^3 freeforms): no loop control at all
      const R* c = sp.coef;
#pragma unroll
      for (int i = 3; i >= 0; --i) {
        R A = c[4 * i + 3], Ad;
        Ad = A;               A = fma(A, y, c[4 * i + 2]);
        Ad = fma(Ad, y, A);   A = fma(A, y, c[4 * i + 1]);
        Ad = fma(Ad, y, A);   A = fma(A, y, c[4 * i]);
        if (i == 3) { P = A; Py = Ad; }
        else {
          if (WANT_SLOPES) { Px = fma(Px, x, P); Py = fma(Py, x, Ad); }
          P = fma(P, x, A);
        }
      }
    } else if (J > 0) {
      for (int b = sp.ncoef - J; b >= 0; b -= J) {  // rows i = I-1 .. 0
        const R* c = sp.coef + b;
        R A, Ad = R(0);
        if (J == 4) {
          A = c[3];
          Ad = A;               A = fma(A, y, c[2]);
          Ad = fma(Ad, y, A);   A = fma(A, y, c[1]);
          Ad = fma(Ad, y, A);   A = fma(A, y, c[0]);
        } else {
          A = c[J - 1];
          for (int j = J - 2; j >= 0; --j) { Ad = fma(Ad, y, A); A = fma(A, y, c[j]); }
        }
        if (WANT_SLOPES) { Px = fma(Px, x, P); Py = fma(Py, x, Ad); }
        P = fma(P, x, A);
      }
    }
    sag = sag + P;
    if (WANT_SLOPES) { zx = zx + Px; zy = zy + Py; }
    return;
  }
#endif
  const int J = sp.aux, I = (J > 0) ? sp.ncoef / J : 0;
  R xi = R(1), xim1 = R(0);  // x**i, x**(i-1)
  for (int i = 0; i < I; ++i) {
    R yj = R(1);
    for (int j = 0; j < J; ++j) {
      const R C = sp.coef[i * J + j];
      sag = sag + (C * xi) * yj;
      if (WANT_SLOPES) {
        if (i >= 1) zx = zx + ((R(i) * C) * xim1) * yj;
      }
      yj = yj * y;
    }
    xim1 = xi;
    xi = xi * x;
  }
  if (WANT_SLOPES) {  // second pass, same order as the reference's dzdy loop
    R xi2 = R(1);
    for (int i = 0; i < I; ++i) {
      R yjm1 = R(1);  // y**(j-1) starting at j=1
      for (int j = 1; j < J; ++j) {
        const R C = sp.coef[i * J + j];
        zy = zy + ((R(j) * C) * xi2) * yjm1;
        yjm1 = yjm1 * y;
      }
      xi2 = xi2 * x;
    }
  }
}

// ZernikePolynomialGeometry.sag / _surface_normal   geometries/zernike.py:133-232
// BaseZernike.get_term/_radial_term/_azimuthal_term/get_derivative/_radial_derivative
//                                                   zernike/base.py:42-299
// Per term in coef: c, c*N, rc_0 .. rc_{smax-1}  (radial coefficients, highest power first),
// (n, m) pairs in ops at sp.aux. cos/sin(m phi) by angle-addition from cos phi = x/rho.
// The merged form (tolerance build). With rho Z_rho = sum_m (m A_m + 2u A_m') Re(w^m) + (m B_m + 2u B_m') Im(w^m) and
// Z_phi = sum_m m (B_m Re(w^m) - A_m Im(w^m)) the reference's slopes -- INCLUDING its two regularised quotients
// 1 / (rho + 1e-14) and 1 / (rho^2 + 1e-14), zernike.py:196-205 -- are
//   zx += (xn rho Z_rho / (rho (rho + eps)) - yn Z_phi / (u + eps)) / Rn,   zy += (yn ... + xn ...) / Rn:
// no angle, no per-term control flow, one Horner step per table entry.
template <typename R, bool WANT_SLOPES>
__device__ __forceinline__ void zernike_eval_merged(const SurfParams<R>& sp, R x, R y, R& sag, R& zx, R& zy,
                                                    uint32_t& status) {
  const R Rn = sp.norm_radius;
  const Recip<R, MM_FAST> rRn(Rn);
  const R xn = rRn.div(x), yn = rRn.div(y);
  if (r_abs(xn) > R(1) || r_abs(yn) > R(1)) status |= OLRT_ST_ZERNIKE_RANGE;  // zernike.py:234-246
  const R u = xn * xn + yn * yn;
  const R r2 = x * x + y * y;
  R denom;
  conic_sag_denom<R, MM_FAST>(r2, sp.cc, sag, denom);
  if (WANT_SLOPES) { const Recip<R, MM_FAST> rd(denom); zx = rd.div(x); zy = rd.div(y); }
  const int* zl = sp.zlen;
  const int mmax = zl[OLRT_ZT_M];
  const bool same = zl[OLRT_ZT_M + 1] != 0;
  const R two_u = u + u;
  R C = R(1), S = R(0), Z = R(0), RZ = R(0), ZP = R(0);
  const R* ta = sp.ztab;
  for (int m = 0; m <= mmax; ++m, ta += 2 * OLRT_ZT_L) {
    const int L = zl[m];      // >= 1 (an |m| without terms holds one zero)
    const R* tb = ta + OLRT_ZT_L;
    R A = ta[L - 1], B = tb[L - 1], Ad = R(0), Bd = R(0);
#pragma unroll 1            // 0 .. 3 trips as a rule: an unrolled body with its remainder code costs more than it saves
    for (int j = L - 2; j >= 0; --j) {
      if (WANT_SLOPES) { Ad = fma(Ad, u, A); Bd = fma(Bd, u, B); }
      A = fma(A, u, ta[j]);
      B = fma(B, u, tb[j]);
    }
    R An = A, Bn = B;
    if (!same) {
      const R* na = ta + OLRT_ZT_M * 2 * OLRT_ZT_L;
      const R* nb = na + OLRT_ZT_L;
      An = na[L - 1]; Bn = nb[L - 1];
#pragma unroll 1
      for (int j = L - 2; j >= 0; --j) { An = fma(An, u, na[j]); Bn = fma(Bn, u, nb[j]); }
    }
    Z = fma(An, C, fma(Bn, S, Z));
    if (WANT_SLOPES) {
      const R fm = R(m);
      RZ = fma(fma(fm, A, two_u * Ad), C, fma(fma(fm, B, two_u * Bd), S, RZ));
      ZP = fma(fm, fma(B, C, -(A * S)), ZP);
    }
    const R Cn = fma(C, xn, -(S * yn));
    S = fma(S, xn, C * yn);
    C = Cn;
  }
  sag = sag + Z;
  if (WANT_SLOPES) {
    const R eps = R(1e-14);
    const R rho = m_sqrt<MM_GUARD>(u);   // exactly 0 on the axis: guarded like the term loop's
    R q1 = R(0);
    if (rho > R(0)) q1 = m_div<MM_FAST>(R(1), rho * (rho + eps));
    const R q2 = m_div<MM_FAST>(R(1), u + eps);
    const R iRn = rRn.div(R(1));
    const R a = (RZ * q1) * iRn, b = (ZP * q2) * iRn;
    zx = zx + fma(xn, a, -(yn * b));
    zy = zy + fma(yn, a, xn * b);
  }
}

template <typename R, bool WANT_SLOPES, int MM>
__device__ __forceinline__ void zernike_eval(const SurfParams<R>& sp, R x, R y, R& sag, R& zx, R& zy,
                                             uint32_t& status) {
  // every quotient below goes through m_div / Recip (bit-identical to IEEE division inside the fast path's domain; the
  // divisors are a positive norm radius, rho under a `rho > 0` test, and sums with eps > 0). rho itself can be exactly
  // 0 on the axis: its square root keeps the guard in the unguarded mode (a NaN there could hide behind the select)
  const R Rn = sp.norm_radius;
  const Recip<R, MM> rRn(Rn);
  const R xn = rRn.div(x), yn = rRn.div(y);
  if (r_abs(xn) > R(1) || r_abs(yn) > R(1)) status |= OLRT_ST_ZERNIKE_RANGE;  // zernike.py:234-246
  const R rho2 = xn * xn + yn * yn;
  const R rho = m_sqrt<(MM == MM_FAST ? MM_GUARD : MM)>(rho2);
  // phi = arctan2(yn, xn); arctan2(0,0)=0
  R cph = R(1), sph = R(0);
  if (rho > R(0)) { const Recip<R, MM> rrho(rho); cph = rrho.div(xn); sph = rrho.div(yn); }
  const R r2 = x * x + y * y;
  R denom;
  conic_sag_denom<R, MM>(r2, sp.cc, sag, denom);
  R drho_dx = R(0), drho_dy = R(0), dphi_dx = R(0), dphi_dy = R(0);
  if (WANT_SLOPES) {
    { const Recip<R, MM> rd(denom); zx = rd.div(x); zy = rd.div(y); }
    const R eps = R(1e-14);
    const R Rn2 = Rn * Rn;
    const Recip<R, MM> rRn2(Rn2), rre(rho + eps), rr2e(rho2 + eps);
    const R iRn = rRn.div(R(1));
    drho_dx = rre.div(rRn2.div(x));
    drho_dy = rre.div(rRn2.div(y));
    dphi_dx = rr2e.div(-yn) * iRn;
    dphi_dy = rr2e.div(xn) * iRn;
  }
  R zsum = R(0);
  const int32_t* nm = sp.ops + sp.aux;
  const R* tc = sp.coef;
  const int nterms = sp.ncoef;
  // cos(|m| phi), sin(|m| phi), rho^|m| and rho^(|m|-1) are carried from term to term: the angle-addition steps and the
  // power chain from |m'| to |m| > |m'| are exactly the tail of the steps from 0 to |m|, so continuing gives the same
  // bits as starting over (the index schemes visit (n, m), (n, -m) back to back); a smaller |m| starts over.
  int cur = 0;
  R cm = R(1), sm = R(0), rm = R(1), rm1 = R(1);
  for (int j = 0; j < nterms; ++j) {
    const int n = nm[2 * j], m = nm[2 * j + 1];
    const int ma = (m < 0) ? -m : m;
    const int smax = ((n - ma) >> 1) + 1;  // n - |m| is even and >= 0
    const R c = tc[0], cN = tc[1];
    const R* rc = tc + 2;
    tc += 2 + smax;
    if (ma < cur) { cur = 0; cm = R(1); sm = R(0); rm = R(1); rm1 = R(1); }
    for (; cur < ma; ++cur) {
      const R cn = cm * cph - sm * sph;
      sm = sm * cph + cm * sph;
      cm = cn;
      rm1 = rm;
      rm = rm * rho;
    }
    // radial polynomial and derivative (Horner in rho^2); the constant term (m == 0, last k) has no derivative
    R H = R(0), Hd = R(0);
    {
      int e = n;  // power of the k-th radial term
      for (int k = 0; k < smax - 1; ++k, e -= 2) {
        H = H * rho2 + rc[k];
        if (WANT_SLOPES) Hd = Hd * rho2 + rc[k] * R(e);
      }
      H = H * rho2 + rc[smax - 1];
      if (WANT_SLOPES && ma > 0) Hd = Hd * rho2 + rc[smax - 1] * R(e);
    }
    const R Rad = H * rm;
    const R az = (m >= 0) ? cm : sm;
    zsum = zsum + (cN * Rad) * az;
    if (WANT_SLOPES && c != R(0)) {
      const R dRad = (ma > 0) ? Hd * rm1 : Hd * rho;
      R dZdr, dZdp;
      if (m == 0) { dZdr = dRad; dZdp = R(0); }
      else if (m > 0) { dZdr = dRad * cm; dZdp = R(-m) * Rad * sm; }
      else { dZdr = dRad * sm; dZdp = R(ma) * Rad * cm; }
      zx = zx + c * (dZdr * drho_dx + dZdp * dphi_dx);
      zy = zy + c * (dZdr * drho_dy + dZdp * dphi_dy);
    }
  }
  sag = sag + zsum;
}

// BiconicGeometry.sag / _surface_normal     geometries/biconic.py:69-158
// coef = cx, cy, (1+kx) cx^2, (1+ky) cy^2 (host, NumPy fp64: the reference's own scalars)
template <typename R, bool WANT_SLOPES, int MM>
__device__ __forceinline__ void biconic_eval(const SurfParams<R>& sp, R x, R y, R& sag, R& zx, R& zy) {
  const R eps = R(1e-14);
  R z[2];
  const R v[2] = {x, y};
  R sl[2] = {R(0), R(0)};
#pragma unroll
  for (int q = 0; q < 2; ++q) {
    const R c = sp.coef[q], A = sp.coef[2 + q], w = v[q];
    z[q] = R(0);
    if (c == R(0)) continue;                      // be.all(self.cx == 0): no contribution from this profile
    const R val = R(1) - A * (w * w);
    const R den = R(1) + m_sqrt<MM>((val < eps) ? R(0) : val);
    z[q] = m_div<MM>(c * (w * w), (r_abs(den) < eps) ? eps : den);
    if (WANT_SLOPES) {
      const R ds = m_sqrt<MM>((val < eps) ? eps : val);
      sl[q] = m_div<MM>(c * w, (r_abs(ds) < eps) ? eps : ds);
    }
  }
  sag = z[0] + z[1];
  if (WANT_SLOPES) { zx = sl[0]; zy = sl[1]; }
}

// ToroidalGeometry.sag / _surface_normal    geometries/toroidal.py:84-240
// coef = R_rot, c_yz, (1+k) c_yz^2, then aux polynomial coefficients of y^2, y^4, ...
template <typename R, bool WANT_SLOPES, int MM>
__device__ __forceinline__ void toroidal_eval(const SurfParams<R>& sp, R x, R y, R& sag, R& zx, R& zy) {
  const R eps = R(1e-14);
  const R Rrot = sp.coef[0], c = sp.coef[1], A = sp.coef[2];
  const R* poly = sp.coef + 3;
  const int np_ = sp.aux;
  const R y2 = y * y;
  R z_y = R(0), dz = R(0);
  if (c != R(0)) {
    const R val = R(1) - A * y2;
    const R den = R(1) + m_sqrt<MM>((val < R(0)) ? R(0) : val);
    z_y = m_div<MM>(c * y2, (r_abs(den) < eps) ? eps : den);
    if (WANT_SLOPES) {
      const R ds = m_sqrt<MM>((val < eps) ? eps : val);
      dz = m_div<MM>(c * y, (r_abs(ds) < eps) ? eps : ds);
    }
  }
  if (np_ > 0) {
    R pt = R(0), pw = y2, pd = R(0), pwd = y;
    for (int i = 0; i < np_; ++i) {
      const R Ci = poly[i];
      pt = pt + Ci * pw;
      pw = pw * y2;
      if (WANT_SLOPES) {
        pd = pd + (Ci * R(2 * (i + 1))) * pwd;
        pwd = pwd * y2;
      }
    }
    z_y = z_y + pt;
    if (WANT_SLOPES) dz = dz + pd;
  }
  if (isinf(Rrot)) {
    sag = z_y;
    if (WANT_SLOPES) { zx = R(0); zy = dz; }
    return;
  }
  const R d = Rrot - z_y;
  const R term = d * d - x * x;
  sag = (term < R(0)) ? R(NAN) : z_y + (d - r_sign(d) * m_sqrt<MM>(term));
  if (WANT_SLOPES) {
    const bool ok = term >= R(0);
    const R sq = m_sqrt<MM>(ok ? term : eps);
    const R ssq = (r_abs(sq) < eps) ? eps : sq;
    const R sR = r_sign(Rrot);
    zx = ok ? m_div<MM>(sR * x, ssq) : R(0);
    zy = ok ? m_div<MM>((sR * d) * dz, ssq) : R(0);
  }
}

// ChebyshevPolynomialGeometry.sag / _surface_normal   geometries/chebyshev.py:103-217
// coef = norm_x, norm_y, then C[i][j] row-major with aux columns; T_n(x) = cos(n arccos x), the derivative
// n sin(n arccos x) / sqrt(1 - x^2) exactly as the reference writes it (libm vs CUDA libm: 1e-10 bar, not bits).
template <typename R, bool WANT_SLOPES, int MM>
__device__ __forceinline__ void chebyshev_eval(const SurfParams<R>& sp, R x, R y, R& sag, R& zx, R& zy,
                                               uint32_t& status) {
  const R xn = x / sp.coef[0], yn = y / sp.coef[1];
  if (r_abs(xn) > R(1) || r_abs(yn) > R(1)) status |= OLRT_ST_CHEBYSHEV_RANGE;  // chebyshev.py:204-217
  const R r2 = x * x + y * y;
  R denom;
  conic_sag_denom<R, MM>(r2, sp.cc, sag, denom);
  if (WANT_SLOPES) { zx = x / denom; zy = y / denom; }
  const int J = sp.aux, I = (J > 0) ? (sp.ncoef - 2) / J : 0;
  const R* C = sp.coef + 2;
  const R ax = acos(xn), ay = acos(yn);
  const R sx = r_sqrt(R(1) - xn * xn), sy = r_sqrt(R(1) - yn * yn);
  for (int i = 0; i < I; ++i) {
    const R Ti = cos(R(i) * ax);
    const R dTi = WANT_SLOPES ? (R(i) * sin(R(i) * ax)) / sx : R(0);
    for (int j = 0; j < J; ++j) {
      const R Cij = C[i * J + j];
      if (Cij == R(0)) continue;                  // be.argwhere(self.coefficients != 0)
      const R Tj = cos(R(j) * ay);
      sag = sag + (Cij * Ti) * Tj;
      if (WANT_SLOPES) {
        const R dTj = (R(j) * sin(R(j) * ay)) / sy;
        zx = zx + (dTi * Cij) * Tj;
        zy = zy + (dTj * Cij) * Ti;
      }
    }
  }
}

// dispatch: sag (and slopes) of an NR geometry at (x, y)
template <typename R, bool WANT_SLOPES, int MM>
__device__ __forceinline__ void nr_eval(const SurfParams<R>& sp, R x, R y, R& sag, R& zx, R& zy,
                                        uint32_t& status) {
  switch (sp.geom) {
    case OLRT_GEOM_EVEN_ASPHERE: even_asphere_eval<R, WANT_SLOPES, MM>(sp, x, y, sag, zx, zy); break;
    case OLRT_GEOM_ODD_ASPHERE: odd_asphere_eval<R, WANT_SLOPES, MM>(sp, x, y, sag, zx, zy); break;
    case OLRT_GEOM_POLYNOMIAL: polynomial_eval<R, WANT_SLOPES, MM>(sp, x, y, sag, zx, zy); break;
    case OLRT_GEOM_BICONIC: biconic_eval<R, WANT_SLOPES, MM>(sp, x, y, sag, zx, zy); break;
    case OLRT_GEOM_TOROIDAL: toroidal_eval<R, WANT_SLOPES, MM>(sp, x, y, sag, zx, zy); break;
    case OLRT_GEOM_CHEBYSHEV: chebyshev_eval<R, WANT_SLOPES, MM>(sp, x, y, sag, zx, zy, status); break;
    default: zernike_eval<R, WANT_SLOPES, MM>(sp, x, y, sag, zx, zy, status); break;
  }
}

}  // namespace OLRT_NS
