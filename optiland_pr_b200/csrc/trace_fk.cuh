// trace_fk.cuh -- the "fast kinds" forward kernel: lenses made only of planes, spheres / conics and
// even aspheres on a single untilted frame (the Cooke triplet, the double Gauss and its even-asphere
// variant: BASELINE configs C1, C2, C4, C5). Same arithmetic, same operation order as the reference
// lines cited in trace_common.cuh / trace_fwd.cuh; what differs from the interpreter kernel
// (trace_fwd.cuh, which stays the path for every other lens and the IEEE fix-up pass) is only the
// code shape: no per-operation guards (MM_FAST + fix-up list), one evaluation body per even asphere
// shared by the Newton iterations and the final surface normal, plane-specific Snell algebra, and a
// small instruction footprint (the round-1 profile showed the interpreter stalling on instruction
// fetch and on ~60 % integer / control instructions).
#pragma once
#include "trace_fwd.cuh"

namespace OLRT_NS {

// frame of a fast kind: translation only (surface_kind() guarantees no rotation, one frame).
// x + (-0.0) == x bit for bit, so a centred surface (x = y = +0.0) skips those additions.
template <typename R>
__device__ __forceinline__ void fk_localize(RayState<R>& r, const R* f, bool centered) {
  if (!centered) { r.x = r.x + (-f[0]); r.y = r.y + (-f[1]); }
  r.z = r.z + (-f[2]);
}
template <typename R>
__device__ __forceinline__ void fk_globalize(RayState<R>& r, const R* f, bool centered) {
  if (!centered) { r.x = r.x + f[0]; r.y = r.y + f[1]; }
  r.z = r.z + f[2];
}

// RadialAperture.contains + RealRays.clip     physical_apertures/radial.py:56-70, real_rays.py:132-139
template <typename R>
__device__ __forceinline__ void fk_clip(RayState<R>& r, const SurfParams<R>& sp, R& att) {
  const int aoff = sp.op[OLRT_OP_APER_OFF];
  if (aoff >= 0) {
    const R r2 = r.x * r.x + r.y * r.y;
    const bool inside = (r2 <= sp.aper[0]) & (r2 >= sp.aper[1]);
    if (!inside) { r.i = R(0); att = R(0); }  // where(~inside, 0, i): exactly 0 from here on
  }
}

// Beer-Lambert attenuation in material_pre, homogeneous.py:45-53: i *= exp(-alpha t 1e3). The exponents of
// successive surfaces are SUMMED in `att` and one exp is taken where the intensity is stored (record /
// output): i0 exp(e1) exp(e2) == i0 exp(e1 + e2) to a few ulp, far inside the 1e-10 bar (exp is not
// bit-comparable with libm anyway). A clipped ray has i == 0 and stays 0.
template <typename R>
__device__ __forceinline__ void fk_absorb(const SurfParams<R>& sp, R t, R& att) {
  if (sp.op[OLRT_OP_FLAGS] & OLRT_F_ABSORB) att = att + ((-sp.alpha) * t) * R(1e3);
}
template <typename R>
__device__ __forceinline__ R fk_intensity(R i, R att) { return (att != R(0)) ? i * r_exp(att) : i; }

// Plane: normal (0, 0, 1)   plane.py:79-98.  With nx = ny = 0 the reference's Snell expression
//   T = u D + n root - (u n) dot,  n = (+-0, +-0, sg),  dot = |N0|
// reduces bit for bit to L = u L0, M = u M0, N = (u N0 + sg root) - u N0  ((u sg)|N0| == u N0 exactly).
// TIR (root = NaN) makes N NaN, which sends the ray to the IEEE fix-up pass where L, M become NaN too.
template <typename R, int MM>
__device__ __forceinline__ void fk_interact_plane(RayState<R>& r, const SurfParams<R>& sp) {
  if (sp.op[OLRT_OP_INTERACT] == 1) {  // reflect: D - 2|N0| n  ->  (L0, M0, N0 - 2 N0)
    r.N = r.N - R(2) * r.N;
    return;
  }
  const R N0 = r.N;
  const R root = m_sqrt<MM>(R(1) - sp.u2 * (R(1) - N0 * N0));
  const R uN = sp.u * N0;
  const R sroot = (N0 < R(0)) ? -root : ((N0 > R(0)) ? root : root * r_sign(N0));
  r.L = sp.u * r.L;
  r.M = sp.u * r.M;
  r.N = (uN + sroot) - uN;
}

template <typename R, int MM>
__device__ __forceinline__ void fk_interact(RayState<R>& r, const SurfParams<R>& sp, R nx, R ny, R nz) {
  if (sp.op[OLRT_OP_INTERACT] == 1) reflect(r, nx, ny, nz);
  else refract<R, MM>(r, nx, ny, nz, sp.u, sp.u2);
}

template <typename R, int MM>
__device__ __forceinline__ void fk_plane(RayState<R>& r, const SurfParams<R>& sp, bool centered, bool last,
                                         R& att, R& L0, R& M0, R& N0) {
  fk_localize(r, sp.frames, centered);
  const R t = m_div<MM>(-r.z, r.N);  // plane.py:61-77
  r.x = r.x + t * r.L;
  r.y = r.y + t * r.M;
  r.z = r.z + t * r.N;
  fk_absorb(sp, t, att);
  r.opd = r.opd + r_abs(t * sp.n1);
  fk_clip(r, sp, att);
  if (last) { L0 = r.L; M0 = r.M; N0 = r.N; }
  fk_interact_plane<R, MM>(r, sp);
  fk_globalize(r, sp.frames, centered);
}

template <typename R, int MM>
__device__ __forceinline__ void fk_conic(RayState<R>& r, const SurfParams<R>& sp, bool centered, bool last,
                                         R& att, R& L0, R& M0, R& N0) {
  fk_localize(r, sp.frames, centered);
  const R t = conic_distance<R, MM>(r, sp.cc, false);
  r.x = r.x + t * r.L;
  r.y = r.y + t * r.M;
  r.z = r.z + t * r.N;
  fk_absorb(sp, t, att);
  r.opd = r.opd + r_abs(t * sp.n1);
  fk_clip(r, sp, att);
  R nx, ny, nz;
  conic_normal<R, MM>(r.x, r.y, sp.cc, nx, ny, nz);
  if (last) { L0 = r.L; M0 = r.M; N0 = r.N; }
  fk_interact<R, MM>(r, sp, nx, ny, nz);
  fk_globalize(r, sp.frames, centered);
}

// Even asphere: conic start, K batch-uniform Newton updates (newton_raphson.py:119-168), then the
// surface normal at the propagated point. The propagated point IS P0 + t_K D -- the expression the
// next Newton iteration would evaluate -- so iteration K of the same loop body yields the normal
// (and, when the host verifies K, the residual max|f| after the last update).
template <typename R, int MM>
__device__ __forceinline__ void fk_even(RayState<R>& r, const SurfParams<R>& sp, bool centered, bool last,
                                        bool valid, unsigned long long* resid, R& att, R& L0, R& M0, R& N0) {
  fk_localize(r, sp.frames, centered);
  R t = conic_distance<R, MM>(r, sp.cc, false);
  const int K = sp.op[OLRT_OP_NR_ITERS];
  const int m = sp.ncoef;
  R xi, yi, zi, nx, ny, nz;
  for (int j = 0;; ++j) {
    xi = r.x + t * r.L;
    yi = r.y + t * r.M;
    zi = r.z + t * r.N;
    // EvenAsphere.sag / _surface_normal, even_asphere.py:82-129
    const R r2 = xi * xi + yi * yi;
    R sag, denom;
    conic_sag_denom<R, MM>(r2, sp.cc, sag, denom);
    const Recip<R, MM> rd(denom);
    R zx = rd.div(xi), zy = rd.div(yi);
    R prev = R(1);        // r2**i
    PowChain<R> pw(r2);   // r2**(i+1), correctly rounded like np.power
    for (int i = 0; i < m; ++i) {
      const R Ci = sp.coef[i];
      if (Ci != R(0)) {   // v + 0*w == v: a zero coefficient (the usual r^2 term) costs nothing
        const R two_i1 = R(2 * (i + 1));
        zx = zx + ((two_i1 * xi) * Ci) * prev;
        zy = zy + ((two_i1 * yi) * Ci) * prev;
        sag = sag + Ci * pw.value();
      }
      prev = pw.value();
      if (i + 1 < m) pw.next();
    }
    const R f = sag - zi;
    if (resid != nullptr) {  // NaN: the ray is on its way to the fix-up list, which reports it (IEEE)
      const unsigned long long key = (valid && f == f) ? resid_bits((double)r_abs(f)) : 0ull;
      report_resid(resid + j, key);
    }
    slopes_to_normal<R, MM>(zx, zy, nx, ny, nz);
    if (j >= K) break;
    const R nzs = (r_abs(nz) > R(1e-14)) ? nz : R(1e-14);
    const Recip<R, MM> rn(nzs);
    const R fx = rn.div(-nx), fy = rn.div(-ny);
    const R df = fx * r.L + fy * r.M - r.N;
    const R dfs = (r_abs(df) > R(1e-14)) ? df : R(1e-14);
    t = t - m_div<MM>(f, dfs);
  }
  r.x = xi; r.y = yi; r.z = zi;  // propagate, homogeneous.py:30-57
  fk_absorb(sp, t, att);
  r.opd = r.opd + r_abs(t * sp.n1);
  fk_clip(r, sp, att);
  if (last) { L0 = r.L; M0 = r.M; N0 = r.N; }
  fk_interact<R, MM>(r, sp, nx, ny, nz);
  fk_globalize(r, sp.frames, centered);
}

// fast-kind eligibility of one surface (ops only): kind, or KIND_GENERIC
__device__ __forceinline__ int fk_kind(const int32_t* s_ops, int s) {
  const int32_t* op = s_ops + s * OLRT_OPS_STRIDE;
  int kind = surface_kind(op);
  const int aoff = op[OLRT_OP_APER_OFF];
  if (aoff >= 0 && !(op[OLRT_OP_APER_LEN] == 1 && s_ops[aoff] == OLRT_AP_RADIAL)) kind = KIND_GENERIC;
  return kind;
}

template <typename R>
__device__ __forceinline__ bool is_pos_zero(R v) { return v == R(0) && !signbit(v); }

// Main pass for lenses whose surfaces are ALL fast kinds (olrt_sys.kernel_hint & OLRT_HINT_FASTKINDS,
// verified here: a wrong hint sets OLRT_ST_BAD_HINT and traces nothing).
template <typename R>
__global__ void __launch_bounds__(OLRT_FK_BLOCK, OLRT_FK_MIN_BLOCKS) trace_fk_kernel(const FwdParams<R> P) {
  constexpr int MM = MM_FAST;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int S = P.n_surfaces, W = P.n_wavelengths;
  R* s_presc = reinterpret_cast<R*>(smem_raw);
  const int presc_pad = (P.presc_stride + 1) & ~1;
  R* s_der = s_presc + presc_pad;
  const int der_pad = (S * der_stride(W) + 1) & ~1;
  int32_t* s_ops = reinterpret_cast<int32_t*>(s_der + der_pad);
  const int ops_pad = (P.ops_len + 1) & ~1;
  unsigned long long* s_resid = reinterpret_cast<unsigned long long*>(s_ops + ops_pad);
  __shared__ int s_roff[OLRT_MAX_SURFACES + 1];
  __shared__ int s_kind[OLRT_MAX_SURFACES];  // kind | centered << 4
  __shared__ int s_bad;

  for (int q = threadIdx.x; q < P.ops_len; q += blockDim.x) s_ops[q] = P.ops[q];
  if (threadIdx.x == 0) s_bad = 0;
  __syncthreads();
  const bool want_resid = (P.nr_resid != nullptr);
  if (threadIdx.x < S) {
    const int k = fk_kind(s_ops, threadIdx.x);
    s_kind[threadIdx.x] = k;
    if (k == KIND_GENERIC) s_bad = 1;
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
  if (s_bad) {
    if (threadIdx.x == 0 && P.status != nullptr) atomicOr(P.status, OLRT_ST_BAD_HINT);
    return;
  }
  const int n_resid = s_roff[S];
  for (int q = threadIdx.x; q < n_resid; q += blockDim.x) s_resid[q] = 0ull;
  __syncthreads();

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
      if (threadIdx.x < S) {
        derive_surface<R>(s_ops, s_presc, s_der, threadIdx.x, W, P.wavelengths, P.four_pi);
        const R* f = s_presc + s_ops[threadIdx.x * OLRT_OPS_STRIDE + OLRT_OP_DOFF];
        const int c = (is_pos_zero(f[0]) && is_pos_zero(f[1])) ? 16 : 0;
        s_kind[threadIdx.x] = (s_kind[threadIdx.x] & 15) | c;
      }
      __syncthreads();
    }
    const long long item = base + threadIdx.x;
    const bool valid = item < P.n_rays;
    const long long idx = valid ? item : (P.n_rays - 1);

    RayState<R> r;
    r.x = P.in[0][idx]; r.y = P.in[1][idx]; r.z = P.in[2][idx];
    r.L = P.in[3][idx]; r.M = P.in[4][idx]; r.N = P.in[5][idx];
    r.i = P.in[6][idx];
    r.opd = (P.opd_in != nullptr) ? P.opd_in[idx] : R(0);
    int widx = 0;
    if (W > 1) {
      const R wl = P.in[7][idx];
      widx = -1;
      for (int q = 0; q < W; ++q) if (wl == P.wavelengths[q]) widx = q;
      if (widx < 0) { status |= OLRT_ST_WAVELENGTH; widx = 0; r.i = R(0) / R(0); }
    }

    R L0 = r.L, M0 = r.M, N0 = r.N, att = R(0);
    for (int s = 0; s < S; ++s) {
      SurfParams<R> sp;
      load_surface<R, false>(sp, s_ops, s_presc, s_der, s, W, widx);
      const int kd = s_kind[s];
      const bool centered = (kd & 16) != 0, last = (s == S - 1);
      switch (kd & 15) {  // warp-uniform
        case KIND_PLANE: fk_plane<R, MM>(r, sp, centered, last, att, L0, M0, N0); break;
        case KIND_CONIC: fk_conic<R, MM>(r, sp, centered, last, att, L0, M0, N0); break;
        default:
          fk_even<R, MM>(r, sp, centered, last, valid, want_resid ? (s_resid + s_roff[s]) : nullptr, att, L0, M0, N0);
          break;
      }
      const int slot = sp.op[OLRT_OP_RECORD];
      if (slot >= 0 && valid && P.record != nullptr) {  // Surface._record, standard_surface.py:266-286
        R* rec = P.record + (long long)slot * 8 * P.n_rays + idx;
        rec[0] = r.x; rec[P.n_rays] = r.y; rec[2 * P.n_rays] = r.z;
        rec[3 * P.n_rays] = r.L; rec[4 * P.n_rays] = r.M; rec[5 * P.n_rays] = r.N;
        rec[6 * P.n_rays] = fk_intensity(r.i, att); rec[7 * P.n_rays] = r.opd;
      }
    }
    r.i = fk_intensity(r.i, att);
    if (valid) {
      P.out[0][idx] = r.x; P.out[1][idx] = r.y; P.out[2][idx] = r.z;
      P.out[3][idx] = r.L; P.out[4][idx] = r.M; P.out[5][idx] = r.N;
      P.out[6][idx] = r.i; P.out[7][idx] = r.opd;
      if (P.dir0 != nullptr) {
        P.dir0[idx] = L0; P.dir0[P.n_rays + idx] = M0; P.dir0[2 * P.n_rays + idx] = N0;
      }
    }
    if (sizeof(R) == 8 && P.fixup != nullptr) {
      const bool bad = valid && (nonfinite(r.x) | nonfinite(r.y) | nonfinite(r.z) | nonfinite(r.L) |
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
