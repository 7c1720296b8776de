// trace_bx.cuh -- reduced-algebra backward (adjoint) kernel for the fast kinds (plane, sphere / conic, even
// asphere on one untilted frame; radial aperture or none; absorbing media allowed): the backward counterpart
// of trace_fx.cuh. Same contract as trace_bwd.cuh (what torch autograd computes through Surface.trace,
// optiland/surfaces/standard_surface.py:186-233, on the reference's torch backend; bar 1e-8 relative), same
// checkpoints (record[s-1] = state entering surface s), far fewer instructions:
//   * a closed-form conic needs no quadratic solve in the backward: the point it reached is the NEXT checkpoint's
//     position, so t = (Q - P).D / D.D, and the conic normal there is (Qx, Qy, -(R - (1+k) Qz)) / |.|
//     (standard.py:142-167 evaluates the same vector from x, y alone; it is differentiated as that function
//     of x, y: G = 1/(R q) = sgn(R) (R^2 - (1+k) r^2)^(-1/2), dG/dR = -R G^3, dG/dk = r^2 G^3 / 2,
//     dG/dr^2 = (1+k) G^3 / 2)
//   * the adjoint of the conic root t (standard.py:89-140) is the implicit derivative of the surface equation
//     F(P + t D) = 0, F = k z^2 - 2 R z + x^2 + y^2 + z^2, whose gradient is parallel to that normal:
//     dt = -(n.(dP + t dD) + g (F_R dR + F_k dk) / 2) / (n.D)  -- identical to differentiating the quadratic
//     formula (either root, and the a == 0 branch, satisfy F = 0)
//   * Newton-Raphson surfaces keep the exact adjoint of the K unrolled updates (newton_raphson.py:119-168; the
//     implicit gradient of the converged root differs by O(residual)); the iterates t_j are kept in registers
//     (K <= 3), the sums are Horner forms with the padded coefficient table of trace_fx.cuh, and the sag
//     derivatives come from dS0/dR = -S0 G0, dS0/dk = S0^2 G0 / 2
//   * intensity: the records already hold i after every surface, so the Beer-Lambert factor of a surface
//     (homogeneous.py:49) is the ratio of two checkpoints instead of an exp(), and the clip (real_rays.py:132-139)
//     is "recorded intensity == 0"
//   * 1/x and 1/sqrt(x) from the MUFU seed + one third-order refinement (FxMath, trace_fx.cuh)
// Parameter partials go to lane-private shared accumulators exactly as in trace_bwd.cuh (bwd_flush).
#pragma once
#include "trace_fx.cuh"
#include "trace_bwd.cuh"

namespace OLRT_NS {

#ifndef OLRT_BX_MIN_BLOCKS
#define OLRT_BX_MIN_BLOCKS 2
#endif
#define OLRT_BX_MAXK 3  // Newton updates whose iterates are kept in registers
#define OLRT_BX_MAX_MEDIA 256  // surfaces x wavelengths of the static media table
#define OLRT_BX_COLS 33        // accumulator columns per slot: 32 lanes + one for the idle lanes of a tail tile

// d(opd)/dt of opd += |t n1| (standard_surface.py:218)
template <typename R>
__device__ __forceinline__ R bx_opd_dt(R gopd, R t, R n1) { return gopd * ((t * n1 < R(0)) ? -n1 : n1); }

// One surface of the intensity chain in reverse: i_rec = clip(ei * exp(nalpha t)). Updates gi (d loss / d i_rec ->
// d loss / d ei) and returns the contribution to d loss / d t.
template <typename R>
__device__ __forceinline__ R bx_intensity(R& gi, R ei, R irec, R t, R r2, const FxSurf<R>& sf,
                                          const FxMedium<R>& md, int kind, bool z0) {
  typedef FxMath<R> Mx;
  R gt = R(0);
  if (z0) {  // ray launched with zero intensity: no checkpoint ratio to read the factors from
    const bool inside = !(kind & FX_APER) || ((r2 <= sf.rmax2) & (r2 >= sf.rmin2));
    if (!inside) gi = R(0);
    else if (md.nalpha != R(0)) {
      const R E = fx_exp(md.nalpha * t);
      gt = gi * ei * E * md.nalpha;
      gi = gi * E;
    }
    return gt;
  }
  if (md.nalpha != R(0)) {
    gt = gi * irec * md.nalpha;
    gi = (ei != R(0)) ? gi * irec * Mx::rcp(ei) : R(0);
  } else if (kind & FX_APER) {
    if (irec == R(0)) gi = R(0);
  }
  return gt;
}

// adjoint of RealRays.refract / reflect (real_rays.py:141-181) for D' = u D + n h(dot), h = sgn(dot) root - u dot,
// or D' = D - 2 dot n. In: entry direction D, unit normal n (not aligned), dot = D.n, g* = d loss / d D'.
// Out: g* = explicit d loss / d D, gn* = d loss / d n.
template <typename R>
__device__ __forceinline__ void bx_interact_adj(const FxMedium<R>& md, int kind, R L, R M, R N, R nx, R ny, R nz,
                                                R dot, R& gL, R& gM, R& gN, R& gnx, R& gny, R& gnz) {
  typedef FxMath<R> Mx;
  const R gw = fma(nx, gL, fma(ny, gM, nz * gN));
  if (kind & FX_REFLECT) {
    const R gdot = R(-2) * gw, d2 = R(-2) * dot;
    gnx = fma(d2, gL, gdot * L); gny = fma(d2, gM, gdot * M); gnz = fma(d2, gN, gdot * N);
    gL = fma(gdot, nx, gL); gM = fma(gdot, ny, gM); gN = fma(gdot, nz, gN);
    return;
  }
  const R c = r_abs(dot);
  const R v = fma(md.u2, fma(dot, dot, R(-1)), R(1));
  const R iroot = Mx::rsqrt(v);
  const R h = Mx::flipsign(fma(-md.u, c, v * iroot), dot);
  const R gdot = gw * fma(md.u2 * c, iroot, -md.u);
  gnx = fma(h, gL, gdot * L); gny = fma(h, gM, gdot * M); gnz = fma(h, gN, gdot * N);
  gL = fma(md.u, gL, gdot * nx); gM = fma(md.u, gM, gdot * ny); gN = fma(md.u, gN, gdot * nz);
}

// Plane, plane.py:61-98. Only the frame's z receives a gradient (x and y enter through the clip alone).
template <typename R>
__device__ __forceinline__ void bx_plane(const FxSurf<R>& sf, const FxMedium<R>& md, int kind, bool z0, R ez, R eL,
                                         R eM, R eN, R ei, R irec, R ex, R ey, R& gx, R& gy, R& gz, R& gL, R& gM,
                                         R& gN, R& gi, R gopd, R& gTz) {
  typedef FxMath<R> Mx;
  const R z = ez - sf.zc;
  const R iN = Mx::rcp(eN);
  const R t = -z * iN;
  R dL, dM, dN;
  if (kind & FX_REFLECT) { dL = gL; dM = gM; dN = -gN; }
  else {  // D' = (u L, u M, sgn(N) root), root = sqrt(1 - u^2 (1 - N^2))
    const R v = fma(md.u2, fma(eN, eN, R(-1)), R(1));
    dL = md.u * gL; dM = md.u * gM;
    dN = gN * (md.u2 * r_abs(eN)) * Mx::rsqrt(v);
  }
  R r2 = R(0);
  if (z0 && (kind & FX_APER)) {
    const R x = (kind & FX_CENTERED) ? ex : ex - sf.xc, y = (kind & FX_CENTERED) ? ey : ey - sf.yc;
    const R Qx = fma(t, eL, x), Qy = fma(t, eM, y);
    r2 = fma(Qx, Qx, Qy * Qy);
  }
  R gt = bx_opd_dt(gopd, t, md.n1) + bx_intensity(gi, ei, irec, t, r2, sf, md, kind, z0);
  gt = fma(eL, gx, fma(eM, gy, fma(eN, gz, gt)));
  dL = fma(t, gx, dL); dM = fma(t, gy, dM); dN = fma(t, gz, dN);
  const R gtn = gt * iN;  // t = -z / N
  dN = fma(-gtn, t, dN);
  gTz = gtn;              // gz(out) - gz(in)
  gz = gz - gtn;
  gL = dL; gM = dM; gN = dN;
}

// Sphere / conic, standard.py:73-167. (Qgx, Qgy, Qgz): the global position the ray reached = next checkpoint.
template <typename R>
__device__ __forceinline__ void bx_conic(const FxSurf<R>& sf, const FxMedium<R>& md, int kind, bool z0, R ex, R ey,
                                         R ez, R eL, R eM, R eN, R ei, R irec, R Qgx, R Qgy, R Qgz, R& gx, R& gy,
                                         R& gz, R& gL, R& gM, R& gN, R& gi, R gopd, R& gTx, R& gTy, R& gTz, R& gR,
                                         R& gk, bool unit_dir = false) {
  typedef FxMath<R> Mx;
  const bool cen = (kind & FX_CENTERED) != 0;
  const R x = cen ? ex : ex - sf.xc, y = cen ? ey : ey - sf.yc, z = ez - sf.zc;
  const R Qx = cen ? Qgx : Qgx - sf.xc, Qy = cen ? Qgy : Qgy - sf.yc, Qz = Qgz - sf.zc;
  // unit_dir: N was rebuilt as sqrt(1 - L^2 - M^2) (blocked checkpoints), so |d|^2 = 1 by construction
  R t = fma(Qx - x, eL, fma(Qy - y, eM, (Qz - z) * eN));
  if (!unit_dir) t *= Mx::rcp(fma(eL, eL, fma(eM, eM, eN * eN)));
  const R r2 = fma(Qx, Qx, Qy * Qy);
  const R w = fma(-sf.c1, Qz, sf.radius);
  const R G = Mx::rcp(w);
  const R g = Mx::copysign_(Mx::rsqrt(fma(w, w, r2)), sf.radius);
  const R nx = Qx * g, ny = Qy * g, nz = -w * g;
  const R dot = fma(eL, nx, fma(eM, ny, eN * nz));
  R gnx, gny, gnz;
  bx_interact_adj(md, kind, eL, eM, eN, nx, ny, nz, dot, gL, gM, gN, gnx, gny, gnz);
  // n = (zx, zy, -1) / |.|, zx = Qx G(r2), zy = Qy G(r2)
  const R ng = fma(nx, gnx, fma(ny, gny, nz * gnz));
  const R hx = fma(-nx, ng, gnx) * g, hy = fma(-ny, ng, gny) * g;  // d loss / d zx * G
  const R q = fma(hx, Qx, hy * Qy) * w;                            // d loss / d G
  const R qG3 = q * G * G * G;
  const R cq = sf.c1 * qG3;
  const R gQx = fma(cq, Qx, gx + hx), gQy = fma(cq, Qy, gy + hy), gQz = gz;
  R pR = -sf.radius * qG3, pk = R(0.5) * r2 * qG3;
  R gt = bx_opd_dt(gopd, t, md.n1) + bx_intensity(gi, ei, irec, t, r2, sf, md, kind, z0);
  gt = fma(eL, gQx, fma(eM, gQy, fma(eN, gQz, gt)));
  gL = fma(t, gQx, gL); gM = fma(t, gQy, gM); gN = fma(t, gQz, gN);
  const R lam = -gt * Mx::rcp(dot);
  const R gPx = fma(lam, nx, gQx), gPy = fma(lam, ny, gQy), gPz = fma(lam, nz, gQz);
  const R lt = lam * t;
  gL = fma(lt, nx, gL); gM = fma(lt, ny, gM); gN = fma(lt, nz, gN);
  const R lgz = lam * g * Qz;
  pR -= lgz;
  pk = fma(R(0.5) * lgz, Qz, pk);
  gTx = gx - gPx; gTy = gy - gPy; gTz = gz - gPz;
  gx = gPx; gy = gPy; gz = gPz;
  gR = pR; gk = pk;
}

// conic + even-asphere sag S, slope factor G (zx = x G), Gp = dG/dr2 at (x, y); G0 / S0 = their conic parts
// (even_asphere.py:82-129). Q(s) = sum_i C_i s^i by Horner together with Q' and Q''/2 (table zero-padded to 4).
template <typename R>
struct BxEval {
  R r2, G0, S0, S, G, Gp;
};
template <typename R, bool WANT_S>
__device__ __forceinline__ void bx_eval(const FxSurf<R>& sf, const R* coef, int m4, R x, R y, BxEval<R>& e) {
  typedef FxMath<R> Mx;
  const R r2 = fma(x, x, y * y);
  e.r2 = r2;
  const R B = fma(-sf.c1iR2, r2, R(1));
  const R is = Mx::rsqrt(B);
  e.G0 = sf.iR * is;
  R Q = R(0), Qp = R(0), Qpp = R(0);
  if (m4 > 0) {  // top block starts from Q = c3, Q' = Q'' = 0
    const R* cb = coef + (m4 - 4);
    Q = fma(cb[3], r2, cb[2]); Qp = cb[3];
    Qpp = Qp; Qp = fma(Qp, r2, Q); Q = fma(Q, r2, cb[1]);
    Qpp = fma(Qpp, r2, Qp); Qp = fma(Qp, r2, Q); Q = fma(Q, r2, cb[0]);
    for (int b = m4 - 8; b >= 0; b -= 4) {
      cb = coef + b;
#pragma unroll
      for (int i = 3; i >= 0; --i) {
        Qpp = fma(Qpp, r2, Qp); Qp = fma(Qp, r2, Q); Q = fma(Q, r2, cb[i]);
      }
    }
  }
  e.G = fma(R(2), fma(r2, Qp, Q), e.G0);
  e.Gp = fma(R(0.5) * sf.c1 * e.G0 * e.G0, e.G0, R(4) * fma(r2, Qpp, Qp));
  if (WANT_S) {
    e.S0 = (r2 * sf.iR) * Mx::rcp(fma(B, is, R(1)));
    e.S = fma(r2, Q, e.S0);
  } else {
    e.S0 = R(0); e.S = R(0);
  }
}

// parameter partials of one evaluation: gS = d loss / d S, qq = d loss / d G
template <typename R, bool WANT_S>
__device__ __forceinline__ void bx_eval_params(const FxSurf<R>& sf, int m4, const BxEval<R>& e, R gS, R qq, R& gR,
                                               R& gk, R* gC) {
  const R qG3 = qq * e.G0 * e.G0 * e.G0;
  gR = fma(-sf.radius, qG3, gR);
  gk = fma(R(0.5) * e.r2, qG3, gk);
  if (WANT_S) {
    const R sg0 = gS * e.S0 * e.G0;
    gR -= sg0;
    gk = fma(R(0.5) * sg0, e.S0, gk);
  }
  // gC_i += gS r2^(i+1) + qq 2 (i+1) r2^i = r2^i (gS r2 + 2 qq (i+1))
  const R a = WANT_S ? gS * e.r2 : R(0), b = R(2) * qq;
  R pw = R(1);
#pragma unroll
  for (int blk = 0; blk < OLRT_BWD_MAXC; blk += 4) {
    if (blk < m4) {
#pragma unroll
      for (int i = blk; i < blk + 4; ++i) {
        gC[i] = fma(pw, fma(b, R(i + 1), a), gC[i]);
        pw = pw * e.r2;
      }
    }
  }
}

// Even asphere: newton_raphson.py:119-168 over even_asphere.py:82-129, K batch-uniform updates from the conic root.
template <typename R>
__device__ __forceinline__ void bx_even(const FxSurf<R>& sf, const FxMedium<R>& md, int kind, int K, int m4,
                                        const R* coef, bool z0, R ex, R ey, R ez, R eL, R eM, R eN, R ei, R irec,
                                        R& gx, R& gy, R& gz, R& gL, R& gM, R& gN, R& gi, R gopd, R& gTx, R& gTy,
                                        R& gTz, R& gR, R& gk, R* gC) {
  typedef FxMath<R> Mx;
  const bool cen = (kind & FX_CENTERED) != 0;
  const R x = cen ? ex : ex - sf.xc, y = cen ? ey : ey - sf.yc, z = ez - sf.zc;
  int unused_guard = 0;
  const R t0 = fx_conic_t(x, y, z, eL, eM, eN, sf, (kind & FX_SPHERE) != 0, unused_guard);
  R t = t0, t1 = t0, t2 = t0;
  BxEval<R> e;
  for (int j = 0; j < K; ++j) {  // forward recompute of the iterates
    const R qx = fma(t, eL, x), qy = fma(t, eM, y), qz = fma(t, eN, z);
    bx_eval<R, true>(sf, coef, m4, qx, qy, e);
    const R h = fma(qx * e.G, eL, fma(qy * e.G, eM, -eN));
    const R hs = (r_abs(h) > R(1e-14)) ? h : R(1e-14);
    t = fma(-(e.S - qz), Mx::rcp(hs), t);
    if (j == 0) t1 = t;
    if (j == 1) t2 = t;
  }
  const R Qx = fma(t, eL, x), Qy = fma(t, eM, y);
  bx_eval<R, false>(sf, coef, m4, Qx, Qy, e);
  const R zx = Qx * e.G, zy = Qy * e.G;
  const R imag = Mx::rsqrt(fma(zx, zx, fma(zy, zy, R(1))));
  const R nx = zx * imag, ny = zy * imag, nz = -imag;
  const R dot = fma(eL, nx, fma(eM, ny, eN * nz));
  R gnx, gny, gnz;
  bx_interact_adj(md, kind, eL, eM, eN, nx, ny, nz, dot, gL, gM, gN, gnx, gny, gnz);
  const R ng = fma(nx, gnx, fma(ny, gny, nz * gnz));
  const R gzx = fma(-nx, ng, gnx) * imag, gzy = fma(-ny, ng, gny) * imag;
  R qq = fma(gzx, Qx, gzy * Qy);
  R pR = R(0), pk = R(0);
  bx_eval_params<R, false>(sf, m4, e, R(0), qq, pR, pk, gC);
  R c2 = R(2) * qq * e.Gp;
  R gPx = fma(c2, Qx, fma(gzx, e.G, gx)), gPy = fma(c2, Qy, fma(gzy, e.G, gy)), gPz = gz;
  R gt = bx_opd_dt(gopd, t, md.n1) + bx_intensity(gi, ei, irec, t, e.r2, sf, md, kind, z0);
  gt = fma(eL, gPx, fma(eM, gPy, fma(eN, gPz, gt)));
  gL = fma(t, gPx, gL); gM = fma(t, gPy, gM); gN = fma(t, gPz, gN);
  for (int j = K - 1; j >= 0; --j) {  // t_{j+1} = t_j - f_j / h_j, reverse
    const R tj = (j == 0) ? t0 : ((j == 1) ? t1 : t2);
    const R qx = fma(tj, eL, x), qy = fma(tj, eM, y), qz = fma(tj, eN, z);
    bx_eval<R, true>(sf, coef, m4, qx, qy, e);
    const R sx = qx * e.G, sy = qy * e.G;
    const R f = e.S - qz;
    const R h = fma(sx, eL, fma(sy, eM, -eN));
    const bool live = r_abs(h) > R(1e-14);
    const R ih = Mx::rcp(live ? h : R(1e-14));
    const R gf = -gt * ih;
    const R gh = live ? gt * f * ih * ih : R(0);
    gL = fma(gh, sx, gL); gM = fma(gh, sy, gM); gN -= gh;
    const R gsx = gh * eL, gsy = gh * eM;
    qq = fma(gsx, qx, gsy * qy);
    bx_eval_params<R, true>(sf, m4, e, gf, qq, pR, pk, gC);
    c2 = R(2) * qq * e.Gp;
    const R gqx = fma(c2, qx, fma(gsx, e.G, gf * sx)), gqy = fma(c2, qy, fma(gsy, e.G, gf * sy)), gqz = -gf;
    gPx += gqx; gPy += gqy; gPz += gqz;
    gL = fma(tj, gqx, gL); gM = fma(tj, gqy, gM); gN = fma(tj, gqz, gN);
    gt = fma(eL, gqx, fma(eM, gqy, fma(eN, gqz, gt)));
  }
  {  // conic root t0: implicit derivative of F(P + t0 D) = 0
    const R ax = fma(t0, eL, x), ay = fma(t0, eM, y), az = fma(t0, eN, z);
    const R w0 = fma(-sf.c1, az, sf.radius);
    const R lam = -gt * Mx::rcp(fma(ax, eL, fma(ay, eM, -w0 * eN)));
    gPx = fma(lam, ax, gPx); gPy = fma(lam, ay, gPy); gPz = fma(-lam, w0, gPz);
    const R lt = lam * t0;
    gL = fma(lt, ax, gL); gM = fma(lt, ay, gM); gN = fma(-lt, w0, gN);
    const R lz = lam * az;
    pR -= lz;
    pk = fma(R(0.5) * lz, az, pk);
  }
  gTx = gx - gPx; gTy = gy - gPy; gTz = gz - gPz;
  gx = gPx; gy = gPy; gz = gPz;
  gR = pR; gk = pk;
}

// dynamic shared memory: checkpoints (offset 0: constant addresses), lane-private gradient accumulators,
// prescription, padded coefficient table, ops; the per-surface FxSurf / FxInfo / FxMedium tables are static
template <typename R>
__host__ __device__ __forceinline__ size_t bx_smem_bytes(int presc_stride, int S, int W, int ops_len, int block,
                                                         int nbuf = 2) {
  return (size_t)nbuf * 7 * block * sizeof(R) + (size_t)((S * OLRT_BWD_SLOTS * OLRT_BX_COLS + 3) & ~3) * sizeof(R) +
         (size_t)((presc_stride + 3) & ~3) * sizeof(R) +
         (size_t)((presc_stride + 4 * S + 3) & ~3) * sizeof(R) + (size_t)((ops_len + 3) & ~3) * sizeof(int32_t);
}

// sum the 32 lane columns of every (surface, slot) into the prescription gradient and clear all 33 (trace_bwd.cuh:
// bwd_flush); coefficient slots beyond the surface's real count (padding of the table) are dropped
template <typename R>
__device__ __forceinline__ void bx_flush(R* s_acc, const int32_t* s_ops, int S, int W, R* gpresc) {
  for (int q = threadIdx.x; q < S * OLRT_BWD_SLOTS; q += blockDim.x) {
    const int s = q / OLRT_BWD_SLOTS, j = q - s * OLRT_BWD_SLOTS;
    R* a = s_acc + (size_t)q * OLRT_BX_COLS;
    R v = R(0);
    for (int l = 0; l < 32; ++l) { v += a[(l + threadIdx.x) & 31]; }
    for (int l = 0; l < OLRT_BX_COLS; ++l) a[l] = R(0);
    const int32_t* op = s_ops + s * OLRT_OPS_STRIDE;
    if (v == R(0) || (j >= 5 && j - 5 >= op[OLRT_OP_NCOEF])) continue;
    const int doff = op[OLRT_OP_DOFF];
    const int goff = doff + op[OLRT_OP_NFRAMES] * OLRT_FRAME_REALS;
    int off;
    if (j < 3) off = doff + j;
    else if (j < 5) off = goff + (j - 3);
    else off = goff + OLRT_GEOM_REALS + 3 * W + (j - 5);
    atomicAdd(gpresc + off, v);
  }
}

// Add NV partials to consecutive accumulator slots of this lane's column. sm_100a has no shared-memory fp64 add: the
// compiler's atomicAdd is a load / add / compare-and-swap loop per value, whose three latencies then run one after
// the other for every partial. Here the NV loads are issued together, then the adds, then the NV compare-and-swaps;
// a value that lost its race (another warp of the CTA updated the same lane's slot in between: rare) falls back to
// the plain loop. fp32 has a native shared-memory add.
template <int NV>
__device__ __forceinline__ void bx_add_n(double* acc, const double* v) {
  unsigned long long o[NV], r[NV];
#pragma unroll
  for (int i = 0; i < NV; ++i)
    o[i] = *reinterpret_cast<volatile unsigned long long*>(acc + i * OLRT_BX_COLS);
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    const unsigned long long n = (unsigned long long)__double_as_longlong(__longlong_as_double((long long)o[i]) + v[i]);
    r[i] = atomicCAS(reinterpret_cast<unsigned long long*>(acc + i * OLRT_BX_COLS), o[i], n);
  }
#pragma unroll
  for (int i = 0; i < NV; ++i)
    if (r[i] != o[i]) atomicAdd(acc + i * OLRT_BX_COLS, v[i]);
}
template <int NV>
__device__ __forceinline__ void bx_add_n(float* acc, const float* v) {
#pragma unroll
  for (int i = 0; i < NV; ++i) atomicAdd(acc + i * OLRT_BX_COLS, v[i]);
}

// 4- or 8-byte cp.async with a 32-bit shared address
template <typename R>
__device__ __forceinline__ void bx_cp_async(unsigned dst, const R* src) {
  if (sizeof(R) == 8) asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(dst), "l"(src) : "memory");
  else asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(dst), "l"(src) : "memory");
}

// the seven components of one checkpoint, STRIDE bytes apart on both sides (tile-blocked checkpoints: 256 reals):
// one asm block with immediate offsets -- no per-copy address arithmetic
template <typename R, int STRIDE>
__device__ __forceinline__ void bx_cp_async7(unsigned dst, const R* src) {
  if (sizeof(R) == 8)
    asm volatile(
        "cp.async.ca.shared.global [%0], [%1], 8;\n\t"
        "cp.async.ca.shared.global [%0+%2], [%1+%2], 8;\n\t"
        "cp.async.ca.shared.global [%0+%3], [%1+%3], 8;\n\t"
        "cp.async.ca.shared.global [%0+%4], [%1+%4], 8;\n\t"
        "cp.async.ca.shared.global [%0+%5], [%1+%5], 8;\n\t"
        "cp.async.ca.shared.global [%0+%6], [%1+%6], 8;\n\t"
        "cp.async.ca.shared.global [%0+%7], [%1+%7], 8;" ::"r"(dst), "l"(src), "n"(STRIDE), "n"(2 * STRIDE),
        "n"(3 * STRIDE), "n"(4 * STRIDE), "n"(5 * STRIDE), "n"(6 * STRIDE)
        : "memory");
  else
    asm volatile(
        "cp.async.ca.shared.global [%0], [%1], 4;\n\t"
        "cp.async.ca.shared.global [%0+%2], [%1+%2], 4;\n\t"
        "cp.async.ca.shared.global [%0+%3], [%1+%3], 4;\n\t"
        "cp.async.ca.shared.global [%0+%4], [%1+%4], 4;\n\t"
        "cp.async.ca.shared.global [%0+%5], [%1+%5], 4;\n\t"
        "cp.async.ca.shared.global [%0+%6], [%1+%6], 4;\n\t"
        "cp.async.ca.shared.global [%0+%7], [%1+%7], 4;" ::"r"(dst), "l"(src), "n"(STRIDE), "n"(2 * STRIDE),
        "n"(3 * STRIDE), "n"(4 * STRIDE), "n"(5 * STRIDE), "n"(6 * STRIDE)
        : "memory");
}

// the six components of one tile-blocked checkpoint (x y z L M copysign(i, N), olrt.h)
template <typename R, int STRIDE>
__device__ __forceinline__ void bx_cp_async6(unsigned dst, const R* src) {
  if (sizeof(R) == 8)
    asm volatile(
        "cp.async.ca.shared.global [%0], [%1], 8;\n\t"
        "cp.async.ca.shared.global [%0+%2], [%1+%2], 8;\n\t"
        "cp.async.ca.shared.global [%0+%3], [%1+%3], 8;\n\t"
        "cp.async.ca.shared.global [%0+%4], [%1+%4], 8;\n\t"
        "cp.async.ca.shared.global [%0+%5], [%1+%5], 8;\n\t"
        "cp.async.ca.shared.global [%0+%6], [%1+%6], 8;" ::"r"(dst), "l"(src), "n"(STRIDE), "n"(2 * STRIDE),
        "n"(3 * STRIDE), "n"(4 * STRIDE), "n"(5 * STRIDE)
        : "memory");
  else
    asm volatile(
        "cp.async.ca.shared.global [%0], [%1], 4;\n\t"
        "cp.async.ca.shared.global [%0+%2], [%1+%2], 4;\n\t"
        "cp.async.ca.shared.global [%0+%3], [%1+%3], 4;\n\t"
        "cp.async.ca.shared.global [%0+%4], [%1+%4], 4;\n\t"
        "cp.async.ca.shared.global [%0+%5], [%1+%5], 4;\n\t"
        "cp.async.ca.shared.global [%0+%6], [%1+%6], 4;" ::"r"(dst), "l"(src), "n"(STRIDE), "n"(2 * STRIDE),
        "n"(3 * STRIDE), "n"(4 * STRIDE), "n"(5 * STRIDE)
        : "memory");
}

// shared load from a 32-bit shared address + element offset
template <typename R> __device__ __forceinline__ R bx_lds(unsigned base, int elem);
template <> __device__ __forceinline__ double bx_lds<double>(unsigned base, int elem) {
  double v;
  asm volatile("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"(base + (unsigned)elem * 8u) : "memory");
  return v;
}
template <> __device__ __forceinline__ float bx_lds<float>(unsigned base, int elem) {
  float v;
  asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(base + (unsigned)elem * 4u) : "memory");
  return v;
}

// BD: block size known at compile time (256), or 0 = blockDim.x (prescription batches with short tiles).
// CK: checkpoints are tile-blocked, [tile][surface][7][256] (olrt.h, OLRT_BWD_CKPT_BLOCKED): P.record is that buffer;
// a thread's seven values of a surface sit at compile-time offsets from ONE pointer that steps back a block per
// surface, so the cp.async copies need no per-copy address arithmetic. (A variant that fetched each 14 KB block with
// one cp.async.bulk per CTA behind mbarriers was measured slower: +9 % instructions for the barrier hand-off and the
// warps of a CTA coupled to its slowest one -- per-thread copies keep the warps independent.)
// NOXY: the caller does not want d loss / d (frame x, frame y) (OLRT_BWD_NO_FRAME_XY: no decenter is a variable):
// two of a curved surface's five shared-memory accumulations less.
// ORD: every surface is centred, refracting and aperture-free (OLRT_HINT_ORDINARY, verified below): only those
// surface bodies are compiled into the variant.
template <typename R, int BD, bool CK, bool NOXY, bool ORD>
__global__ void __launch_bounds__(256, OLRT_BX_MIN_BLOCKS) trace_bx_kernel(const BwdParams<R> P, uint32_t* status) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  __shared__ FxSurf<R> s_surf[OLRT_MAX_SURFACES];
  __shared__ FxInfo s_info[OLRT_MAX_SURFACES];
  __shared__ FxMedium<R> s_med[OLRT_BX_MAX_MEDIA];
  __shared__ int s_bad;
  constexpr int NB = 2;  // checkpoint buffers
  const int S = P.n_surfaces, W = P.n_wavelengths;
  const int bd = BD ? BD : (int)blockDim.x;
  // entry states (checkpoints) staged one surface ahead by cp.async: s_ck[(buf * 7 + q) * bd + tid]
  R* s_ck = reinterpret_cast<R*>(smem_raw);
  // d loss / d presc per LANE (trace_bwd.cuh): s_acc[(s * OLRT_BWD_SLOTS + j) * OLRT_BX_COLS + lane]; the lanes a
  // tail tile leaves idle add (to) column 32, which is never summed: no `valid` branch around the atomics
  R* s_acc = s_ck + NB * 7 * bd;
  R* s_presc = s_acc + ((S * OLRT_BWD_SLOTS * OLRT_BX_COLS + 3) & ~3);  // keep 16 B alignment
  R* s_coef = s_presc + ((P.presc_stride + 3) & ~3);
  int32_t* s_ops = reinterpret_cast<int32_t*>(s_coef + ((P.presc_stride + 4 * S + 3) & ~3));

  for (int q = threadIdx.x; q < P.ops_len; q += bd) s_ops[q] = P.ops[q];
  for (int q = threadIdx.x; q < S * OLRT_BWD_SLOTS * OLRT_BX_COLS; q += bd) s_acc[q] = R(0);
  if (threadIdx.x == 0) s_bad = 0;
  __syncthreads();
  if (threadIdx.x < S) {
    const int32_t* op = s_ops + threadIdx.x * OLRT_OPS_STRIDE;
    const int k = fk_kind(s_ops, threadIdx.x);
    s_info[threadIdx.x].kind = k;
    if (k == KIND_GENERIC || (!CK && op[OLRT_OP_RECORD] != (int)threadIdx.x)) s_bad = 1;
    if (k == KIND_EVEN && (op[OLRT_OP_NCOEF] > OLRT_BWD_MAXC || op[OLRT_OP_NR_ITERS] > OLRT_BX_MAXK)) s_bad = 1;
  }
  __syncthreads();
  if (s_bad) {
    if (threadIdx.x == 0 && status != nullptr) atomicOr(status, OLRT_ST_BAD_HINT);
    return;
  }
  const long long N = P.n_rays;
  const int lane = threadIdx.x & 31;
  // this thread's column of the checkpoint buffers as a 32-bit shared address; the volatile asm keeps the compiler
  // from re-deriving it (cvta + 64-bit arithmetic) at every cp.async
  unsigned ck_a;
  asm volatile("{ .reg .u64 t; cvta.to.shared.u64 t, %1; cvt.u32.u64 %0, t; }" : "=r"(ck_a) : "l"(s_ck + threadIdx.x));
  int cur_p = -1;
  const long long n_tiles = (N + bd - 1) / bd;
  for (long long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    const long long base = tile * bd;
    const int p = (P.n_presc > 1) ? (int)(base / P.rays_per_presc) : 0;
    if (p != cur_p) {
      __syncthreads();
      if (cur_p >= 0) {
        bx_flush<R>(s_acc, s_ops, S, W, P.gpresc + (long long)cur_p * P.presc_stride);
        __syncthreads();
      }
      const R* src = P.presc + (long long)p * P.presc_stride;
      for (int q = threadIdx.x; q < P.presc_stride; q += bd) s_presc[q] = src[q];
      cur_p = p;
      __syncthreads();
      if (threadIdx.x < S) {
        fx_build_surface<R>(s_ops, s_presc, s_surf, s_med, s_info, s_coef, threadIdx.x, S, W, P.wavelengths,
                            P.four_pi);
        s_info[threadIdx.x].pad0 = s_ops[threadIdx.x * OLRT_OPS_STRIDE + OLRT_OP_NCOEF];  // unpadded count
        if (ORD && s_info[threadIdx.x].pad1 < 3) s_bad = 1;   // the hint claimed an ordinary lens
      }
      __syncthreads();
      if (ORD && s_bad) {   // CTA-uniform: nothing is written for a wrong hint
        if (threadIdx.x == 0 && status != nullptr) atomicOr(status, OLRT_ST_BAD_HINT);
        return;
      }
    }
    const long long item = base + threadIdx.x;
    const bool valid = item < N;
    const long long idx = valid ? item : (N - 1);
    const int col = valid ? lane : 32;
    int widx = 0;
    if (W > 1) {
      const R wl = P.in[7][idx];
      for (int q = 0; q < W; ++q) if (wl == P.wavelengths[q]) widx = q;
    }
    // state after the last surface, then the entry state of surface S-1 (= record S-2; the launch state for S == 1)
    const R* rp = CK ? P.record + (tile * S + (S - 1)) * (long long)(OLRT_CKPT_VALS * bd) + threadIdx.x  // block of surface S-1
                     : P.record + (long long)(S - 1) * 8 * N + idx;
    // CK: a plane as the last surface has no checkpoint (olrt.h): bx_plane needs no exit point, and its intensity
    // factor is computed directly (z0p below) instead of from the exit intensity
    const bool last_plane = CK && (s_info[S - 1].kind & 15) == KIND_PLANE;
    R Qgx = R(0), Qgy = R(0), Qgz = R(0), irec = R(0);
    if (CK) {
      if (!last_plane) { Qgx = rp[0]; Qgy = rp[bd]; Qgz = rp[2 * bd]; irec = fabs(rp[5 * bd]); }
    } else {
      Qgx = rp[0]; Qgy = rp[N]; Qgz = rp[2 * N]; irec = rp[6 * N];
    }
    const long long N8 = 8 * N;
    auto stage = [&](int s_) {  // entry state of surface s_ -> buffer s_ & 1; rp walks down the records
      const unsigned dst = ck_a + (unsigned)((s_ & 1) * 7 * bd * (int)sizeof(R));
      if (CK && s_ > 0) {
        rp -= OLRT_CKPT_VALS * bd;  // block of surface s_ - 1 (rp already includes this thread's column)
        bx_cp_async6<R, (BD ? BD : 256) * (int)sizeof(R)>(dst, rp);   // CK implies 256-ray tiles
        cp_async_commit();
        return;
      }
      if (s_ == 0) {
#pragma unroll
        for (int q = 0; q < 7; ++q) bx_cp_async<R>(dst + (unsigned)(q * bd * (int)sizeof(R)), P.in[q] + idx);
      } else {
        rp -= N8;  // record s_ - 1
        const R* pr = rp;
#pragma unroll
        for (int q = 0; q < 7; ++q) { bx_cp_async<R>(dst + (unsigned)(q * bd * (int)sizeof(R)), pr); pr += N; }
      }
      cp_async_commit();
    };
    stage(S - 1);
    // idx is clamped for the idle lanes of a tail tile: they read the last ray's values and add to the spare column
    auto ld = [&](const R* ptr) { return (ptr != nullptr) ? ptr[idx] : R(0); };
    R gx = ld(P.gout[0]), gy = ld(P.gout[1]), gz = ld(P.gout[2]);
    R gL = ld(P.gout[3]), gM = ld(P.gout[4]), gN = ld(P.gout[5]);
    R gi = ld(P.gout[6]), gopd = ld(P.gout[7]);
    const bool z0 = (P.in[6][idx] == R(0));
    for (int s = S - 1; s >= 0; --s) {
      const int cb = s & 1;  // buffer of surface s
      if (s > 0) { stage(s - 1); cp_async_wait<1>(); } else { cp_async_wait<0>(); }
      if (!CK && P.grecord != nullptr && valid) {
        const R* g = P.grecord + (long long)s * 8 * N + idx;
        gx += g[0]; gy += g[N]; gz += g[2 * N]; gL += g[3 * N]; gM += g[4 * N]; gN += g[5 * N];
        gi += g[6 * N]; gopd += g[7 * N];
      }
      const unsigned cka = ck_a + (unsigned)(cb * 7 * bd * (int)sizeof(R));
      const R ex = bx_lds<R>(cka, 0), ey = bx_lds<R>(cka, bd), ez = bx_lds<R>(cka, 2 * bd),
              eL = bx_lds<R>(cka, 3 * bd), eM = bx_lds<R>(cka, 4 * bd);
      R eN = bx_lds<R>(cka, 5 * bd), ei;
      if (CK && s > 0) {   // blocked checkpoints keep copysign(intensity, N) in the sixth slot (olrt.h): rebuild N
        typedef FxMath<R> Mq;
        ei = fabs(eN);
        const R n2 = fma(-eM, eM, fma(-eL, eL, R(1)));
        eN = Mq::copysign_(n2 > R(0) ? n2 * Mq::rsqrt(n2) : n2 * R(0), eN);
      } else {
        ei = bx_lds<R>(cka, 6 * bd);
      }
      const FxSurf<R>& sf = s_surf[s];
      const FxMedium<R>& md = s_med[s * W + widx];
      const FxInfo& info = s_info[s];
      const int kind = info.kind;
      R* acc = s_acc + s * (OLRT_BWD_SLOTS * OLRT_BX_COLS) + col;
      // the per-kind bodies take the flag word as a parameter: the dense variant id of fx_build_surface lets an
      // ordinary lens surface (centred, refracting, no aperture) pass its flags as a compile-time constant
      const bool z0p = z0 || (last_plane && s == S - 1);
      auto do_plane = [&](const int kk) {
        R gTz;
        bx_plane<R>(sf, md, kk, z0p, ez, eL, eM, eN, ei, irec, ex, ey, gx, gy, gz, gL, gM, gN, gi, gopd, gTz);
        atomicAdd(acc + 2 * OLRT_BX_COLS, gTz);
      };
      auto do_conic = [&](const int kk) {
        R gTx, gTy, gTz, gR, gk;
        bx_conic<R>(sf, md, kk, z0, ex, ey, ez, eL, eM, eN, ei, irec, Qgx, Qgy, Qgz, gx, gy, gz, gL, gM, gN, gi,
                    gopd, gTx, gTy, gTz, gR, gk, CK && s > 0);
        if (NOXY) { const R pv[3] = {gTz, gR, gk}; bx_add_n<3>(acc + 2 * OLRT_BX_COLS, pv); }
        else { const R pv[5] = {gTx, gTy, gTz, gR, gk}; bx_add_n<5>(acc, pv); }
      };
      auto do_even = [&](const int kk) {
        R gTx, gTy, gTz, gR, gk;
        R gC[OLRT_BWD_MAXC];
#pragma unroll
        for (int i = 0; i < OLRT_BWD_MAXC; ++i) gC[i] = R(0);
        const int m4 = info.ncoef;
        bx_even<R>(sf, md, kk, info.K, m4, s_coef + info.coef_off, z0, ex, ey, ez, eL, eM, eN, ei, irec, gx, gy,
                   gz, gL, gM, gN, gi, gopd, gTx, gTy, gTz, gR, gk, gC);
        if (NOXY) { const R pv[3] = {gTz, gR, gk}; bx_add_n<3>(acc + 2 * OLRT_BX_COLS, pv); }
        else { const R pv[5] = {gTx, gTy, gTz, gR, gk}; bx_add_n<5>(acc, pv); }
        // coefficient partials in blocks of 4 like the padded table (bx_flush drops the padding's)
        if (m4 > 0) bx_add_n<4>(acc + 5 * OLRT_BX_COLS, gC);
        if (m4 > 4) bx_add_n<4>(acc + 9 * OLRT_BX_COLS, gC + 4);
      };
      if (ORD) {
        switch (info.pad1) {  // warp-uniform
          case 3: do_plane(KIND_PLANE | FX_CENTERED); break;
          case 4:
          case 5: do_conic(KIND_CONIC | FX_CENTERED); break;
          case 6: do_even(KIND_EVEN | FX_CENTERED | FX_SPHERE); break;
          default: do_even(KIND_EVEN | FX_CENTERED); break;
        }
      } else {
        switch (info.pad1) {  // warp-uniform, dense
          case 0: do_plane(kind); break;
          case 1: do_conic(kind); break;
          case 3: do_plane(KIND_PLANE | FX_CENTERED); break;
          case 4:
          case 5: do_conic(KIND_CONIC | FX_CENTERED); break;
          case 6: do_even(KIND_EVEN | FX_CENTERED | FX_SPHERE); break;
          case 7: do_even(KIND_EVEN | FX_CENTERED); break;
          default: do_even(kind); break;
        }
      }
      Qgx = ex; Qgy = ey; Qgz = ez; irec = ei;
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
  if (cur_p >= 0) bx_flush<R>(s_acc, s_ops, S, W, P.gpresc + (long long)cur_p * P.presc_stride);
}

template <typename R, int BD, bool CK, bool NOXY, bool ORD>
int bx_launch_kernel(const BwdParams<R>& P, int block, size_t smem, long long n_tiles, uint32_t* status,
                     cudaStream_t stream) {
  // static tables (16 KB) + dynamic can pass the 48 KB default even when the dynamic part alone does not
  cudaFuncSetAttribute(trace_bx_kernel<R, BD, CK, NOXY, ORD>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  int per_sm = 0, dev = 0, sms = 148;
  cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, trace_bx_kernel<R, BD, CK, NOXY, ORD>, block, smem);
  if (per_sm < 1) per_sm = 1;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  long long grid = (long long)sms * per_sm;
  if (grid > n_tiles) grid = n_tiles;
  trace_bx_kernel<R, BD, CK, NOXY, ORD><<<(unsigned)grid, block, smem, stream>>>(P, status);
  return cudaGetLastError() == cudaSuccess ? OLRT_OK : OLRT_E_LAUNCH;
}

// flags: OLRT_BWD_GENERIC forces the interpreter-style adjoint of trace_bwd.cuh (any aperture program, any Newton count)
template <typename R>
int launch_bx(const olrt_sys* sys, const R* presc, const int32_t* ops, long long n_rays, const R* const in[8],
              const R* record, const R* const gout[8], const R* grecord, R* const gin[8], R* gpresc,
              uint32_t* status, int flags, cudaStream_t stream) {
  if ((flags & OLRT_BWD_CKPT_BLOCKED) && (!(sys->kernel_hint & OLRT_HINT_FASTKINDS) || (flags & OLRT_BWD_GENERIC)))
    return OLRT_E_UNSUPPORTED;
  if (!(sys->kernel_hint & OLRT_HINT_FASTKINDS) || (flags & OLRT_BWD_GENERIC))
    return launch_bwd<R>(sys, presc, ops, n_rays, in, record, gout, grecord, gin, gpresc, status, stream);
  BwdParams<R> P;
  for (int q = 0; q < 8; ++q) { P.in[q] = in[q]; P.gout[q] = gout ? gout[q] : nullptr; P.gin[q] = gin ? gin[q] : nullptr; }
  P.record = record; P.grecord = grecord; P.gpresc = gpresc; P.presc = presc; P.ops = ops; P.n_rays = n_rays;
  P.rays_per_presc = sys->rays_per_presc > 0 ? sys->rays_per_presc : n_rays;
  P.n_surfaces = sys->n_surfaces; P.n_wavelengths = sys->n_wavelengths; P.n_presc = sys->n_presc;
  P.presc_stride = sys->presc_stride; P.ops_len = sys->ops_len;
  for (int q = 0; q < OLRT_MAX_WAVELENGTHS; ++q) P.wavelengths[q] = (R)sys->wavelengths[q];
  P.four_pi = (R)(4.0 * 3.141592653589793);
  int block = 256;
  if (sys->n_presc > 1) {
    block = 0;
    for (int b = 256; b >= 32; b >>= 1) if (sys->rays_per_presc % b == 0) { block = b; break; }
    if (block == 0) return OLRT_E_ARG;
  }
  const size_t smem = bx_smem_bytes<R>(sys->presc_stride, sys->n_surfaces, sys->n_wavelengths, sys->ops_len, block);
  if (smem > 180 * 1024 || sys->n_surfaces * sys->n_wavelengths > OLRT_BX_MAX_MEDIA) return OLRT_E_LIMIT;
  const long long n_tiles = (n_rays + block - 1) / block;
  if (flags & OLRT_BWD_CKPT_BLOCKED) {  // `record` is the tile-blocked checkpoint buffer of the forward
    if (block != OLRT_CKPT_TILE || sys->n_presc > 1 || grecord != nullptr) return OLRT_E_UNSUPPORTED;
    const bool noxy = (flags & OLRT_BWD_NO_FRAME_XY) != 0, ord = (sys->kernel_hint & OLRT_HINT_ORDINARY) != 0;
    if (noxy && ord) return bx_launch_kernel<R, 256, true, true, true>(P, block, smem, n_tiles, status, stream);
    if (noxy) return bx_launch_kernel<R, 256, true, true, false>(P, block, smem, n_tiles, status, stream);
    if (ord) return bx_launch_kernel<R, 256, true, false, true>(P, block, smem, n_tiles, status, stream);
    return bx_launch_kernel<R, 256, true, false, false>(P, block, smem, n_tiles, status, stream);
  }
  if (block == 256) {
    const bool noxy = (flags & OLRT_BWD_NO_FRAME_XY) != 0, ord = (sys->kernel_hint & OLRT_HINT_ORDINARY) != 0;
    if (noxy && ord) return bx_launch_kernel<R, 256, false, true, true>(P, block, smem, n_tiles, status, stream);
    if (noxy) return bx_launch_kernel<R, 256, false, true, false>(P, block, smem, n_tiles, status, stream);
    if (ord) return bx_launch_kernel<R, 256, false, false, true>(P, block, smem, n_tiles, status, stream);
    return bx_launch_kernel<R, 256, false, false, false>(P, block, smem, n_tiles, status, stream);
  }
  return bx_launch_kernel<R, 0, false, false, false>(P, block, smem, n_tiles, status, stream);
}

}  // namespace OLRT_NS
