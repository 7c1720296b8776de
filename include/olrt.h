/* olrt.h -- C ABI of libolrt.so: fused sequential real-ray trace for NVIDIA B200 (sm_100a).
 *
 * The reference (Optiland fork, pure Python) has no FFI; the seam this library
 * replaces is the method
 *     SurfaceGroup.trace(self, rays, skip=0)      optiland/surfaces/surface_group.py:232-244
 * i.e. `for surface in surfaces: surface.trace(rays)` with
 *     Surface.trace                               optiland/surfaces/standard_surface.py:186-233
 * and, for the optimiser, torch autograd through those same lines, plus the
 * RMS-spot / centroid consumers
 *     RayOperand.rms_spot_size                    optiland/optimization/operand/ray.py:299-340
 *     SpotDiagram.centroid / rms_spot_radius      optiland/analysis/spot_diagram.py:317-357
 * INTEGRATION.md shows the ctypes binding a maintainer adds on the reference side.
 *
 * Conventions
 *  - extern "C", plain pointers and sizes, no torch types.
 *  - All array pointers are DEVICE pointers unless marked [host].
 *  - The caller owns every buffer; the library never allocates, frees or keeps
 *    a pointer past the call. Host structs are copied at enqueue time.
 *  - Every call only ENQUEUES work on `stream` (a cudaStream_t passed as void*).
 *  - Return 0 = enqueued; <0 = OLRT_E_* argument / launch error. Nothing throws.
 *  - Data-dependent conditions are bits in a device `status` word.
 */
#ifndef OLRT_H
#define OLRT_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define OLRT_ABI_VERSION 6   /* 6: 6-value blocked checkpoints, OLRT_HINT_ORDINARY, OLRT_BWD_NO_FRAME_XY, group counts with bit reports, olrt_pupil_random_f64, olrt_trace_fwd_noalloc_f64 */

#define OLRT_MAX_SURFACES 64    /* intersected surfaces per system (object surface excluded) */
#define OLRT_MAX_WAVELENGTHS 8  /* distinct wavelengths per launch */
#define OLRT_MAX_NR_ITERS 128   /* per-surface Newton trip count upper bound (reference default max_iter=100) */
#define OLRT_OPS_STRIDE 16      /* int32 words per surface in the ops table */
#define OLRT_RESID_STRIDE (OLRT_MAX_NR_ITERS + 1)
#define OLRT_SPOT_SUMS 10       /* doubles per spot group, see olrt_spot_sums_* */

/* error codes */
#define OLRT_OK 0
#define OLRT_E_ARG (-1)      /* null pointer / bad size / bad table */
#define OLRT_E_LIMIT (-2)    /* exceeds an OLRT_MAX_* limit or shared memory */
#define OLRT_E_LAUNCH (-3)   /* cudaGetLastError() != cudaSuccess after the launch */
#define OLRT_E_UNSUPPORTED (-4) /* the request is outside what the kernel implements (never approximated) */

/* ops table, one row of OLRT_OPS_STRIDE int32 per intersected surface */
enum {
  OLRT_OP_GEOM = 0,      /* OLRT_GEOM_* */
  OLRT_OP_INTERACT = 1,  /* 0 refract, 1 reflect */
  OLRT_OP_FLAGS = 2,     /* OLRT_F_* */
  OLRT_OP_DOFF = 3,      /* offset (in reals) of this surface's block inside one prescription */
  OLRT_OP_NCOEF = 4,     /* even asphere: m; polynomial: I*J; zernike: number of terms */
  OLRT_OP_AUX = 5,       /* polynomial: J (columns); zernike: offset (ints) of its (n,m) table in ops */
  OLRT_OP_NR_ITERS = 6,  /* K_s: Newton updates to take at this surface (batch-uniform) */
  OLRT_OP_APER_OFF = 7,  /* offset (ints) of the aperture program in ops, or -1 */
  OLRT_OP_APER_LEN = 8,  /* tokens in the aperture program */
  OLRT_OP_RECORD = 9,    /* record slot (>=0) or -1: post-interaction global state is stored */
  OLRT_OP_NFRAMES = 10,  /* coordinate frames in the reference_cs chain, root first (>=1) */
  OLRT_OP_FRAMEFLAGS = 11, /* 3 bits per frame: bit0 rz, bit1 ry, bit2 rx is "truthy" (rotation applied) */
  OLRT_OP_APER_DOFF = 12 /* offset (in reals, inside the surface block) of aperture parameters */
};

enum {
  OLRT_GEOM_PLANE = 0,        /* optiland/geometries/plane.py:61-98 */
  OLRT_GEOM_STANDARD = 1,     /* optiland/geometries/standard.py:89-167 */
  OLRT_GEOM_EVEN_ASPHERE = 2, /* optiland/geometries/even_asphere.py:82-129 + newton_raphson.py:119-168 */
  OLRT_GEOM_POLYNOMIAL = 3,   /* optiland/geometries/polynomial.py:93-142 */
  OLRT_GEOM_ZERNIKE = 4,      /* optiland/geometries/zernike.py:133-232 */
  OLRT_GEOM_ODD_ASPHERE = 5,  /* optiland/geometries/odd_asphere.py:73-129 */
  OLRT_GEOM_BICONIC = 6,      /* optiland/geometries/biconic.py:69-158; coefficients: cx cy (1+kx)cx^2 (1+ky)cy^2;
                                 radius / conic slots = Rx, kx (the base conic of the first Newton guess) */
  OLRT_GEOM_TOROIDAL = 7,     /* optiland/geometries/toroidal.py:84-240; coefficients: R_rot c_yz (1+k)c_yz^2 then
                                 AUX polynomial terms of y^2, y^4 ...; radius / conic slots = R_yz, 0 */
  OLRT_GEOM_CHEBYSHEV = 8     /* optiland/geometries/chebyshev.py:103-217; coefficients: norm_x norm_y then C[i][j]
                                 row-major with AUX columns */
};

enum {
  OLRT_F_RADIUS_INF = 1,  /* StandardGeometry with infinite radius: clamped plane branch, standard.py:100-103 */
  OLRT_F_ABSORB = 2,      /* material_pre has k>0 for some wavelength: homogeneous.py:45-53 */
  /* The next three put a surface on the interpreter kernel (never a fast kind, not differentiable). Their reals sit
   * at the END of the surface's block, just before the aperture parameters: [focal length][coating factor]. */
  OLRT_F_COATING = 4,     /* SimpleCoating: intensity *= transmittance (reflectance at a mirror) after the
                             interaction, coatings.py:156-224 + interactions/base.py:111-128 */
  OLRT_F_THIN_LENS = 8,   /* ThinLensInteractionModel.interact_real_rays (paraxial surface of focal length f),
                             interactions/thin_lens_interaction_model.py:55-113: leaves (L, M, 1) un-normalised */
  OLRT_F_RENORM = 16      /* the previous surface was a thin lens: the direction is normalised at the end of this
                             surface's propagation step, propagation/homogeneous.py:55-57, rays/real_rays.py:503-509 */
};

/* aperture program tokens (reverse Polish), optiland/physical_apertures/ */
enum {
  OLRT_AP_RADIAL = 1,        /* 2 reals: r_max^2, r_min^2            radial.py:56-70 */
  OLRT_AP_OFFSET_RADIAL = 2, /* 4 reals: r_max^2, r_min^2, ox, oy    offset_radial.py:48 */
  OLRT_AP_ELLIPTICAL = 3,    /* 4 reals: a^2, b^2, ox, oy            elliptical.py:42 */
  OLRT_AP_RECTANGULAR = 4,   /* 4 reals: x_min, x_max, y_min, y_max  rectangular.py:42 */
  OLRT_AP_OR = 10,           /* base.py:255-275 */
  OLRT_AP_AND = 11,          /* base.py:278-300 */
  OLRT_AP_DIFF = 12          /* base.py:303-336 */
};

/* Layout of one surface block of reals (double for *_f64, float for *_f32):
 *   frames   : NFRAMES x 9  = x, y, z, cos rz, sin rz, cos ry, sin ry, cos rx, sin rx
 *   geometry : radius, conic, tol, norm_radius
 *   media    : n_pre[W], n_post[W], k_pre[W]            (W = n_wavelengths)
 *   coefs    : NCOEF reals (even asphere C_i; polynomial C_ij row-major; zernike:
 *              per term c_j*N_j followed by its radial-polynomial coefficients)
 *   aperture : parameters of the aperture program leaves, in token order
 */
#define OLRT_FRAME_REALS 9
#define OLRT_GEOM_REALS 4

/* status bits (device word, OR-accumulated) */
#define OLRT_ST_ZERNIKE_RANGE 1u /* |x/R_n|>1 or |y/R_n|>1 at a Zernike surface: zernike.py:234-246 */
#define OLRT_ST_WAVELENGTH 2u    /* a ray wavelength is not in olrt_sys.wavelengths */
#define OLRT_ST_CHEBYSHEV_RANGE 8u /* |x/norm_x|>1 or |y/norm_y|>1 at a Chebyshev surface: chebyshev.py:204-217 */
#define OLRT_ST_TIMEOUT 16u      /* streamed pupil grid (olrt_pupil.ready): a chunk never arrived */
#define OLRT_ST_BAD_HINT 4u      /* olrt_sys.kernel_hint claims something the ops table does not satisfy: nothing traced */

/* kernel_hint bits */
#define OLRT_HINT_FASTKINDS 1 /* every surface is a plane / conic / even asphere on one untilted frame,
                                aperture none or a single RadialAperture: the straight-line kernel (trace_fk.cuh) runs */
#define OLRT_HINT_ORDINARY 2  /* with FASTKINDS: every surface of EVERY prescription of the batch is also centred
                                (frame x = y = +0.0), refracting and has no aperture -- an ordinary lens. The kernels
                                then run variants that carry only those surface bodies (smaller code, fewer registers).
                                Verified on the device: a wrong hint sets OLRT_ST_BAD_HINT and traces nothing */
#define OLRT_HINT_ZERNIKE 4   /* the ops table holds a Zernike surface: tolerance-mode launches (OLRT_FWD_FAST) run the
                                interpreter variant that merges each such surface's terms into one polynomial in rho^2
                                per (|m|, cos / sin) pair, per CTA in shared memory. A hint in the plain
                                sense: without it the same surfaces run the term loop (slower, same tolerance); with
                                it and no Zernike surface nothing changes */

typedef struct olrt_sys {
  int32_t n_surfaces;     /* intersected surfaces */
  int32_t n_wavelengths;  /* W, 1..OLRT_MAX_WAVELENGTHS */
  int32_t n_presc;        /* prescriptions in the batch (Monte Carlo trials); >=1 */
  int32_t presc_stride;   /* reals per prescription */
  int32_t ops_len;        /* int32 words in ops */
  int32_t n_record;       /* record slots */
  int32_t nr_resid_words; /* sum over Newton surfaces of (K_s + 1): shared-memory words for nr_resid */
  int32_t kernel_hint;    /* OLRT_HINT_* (the ops table lives on the device; the host states what it holds) */
  int64_t rays_per_presc; /* ray r uses prescription r / rays_per_presc (ignored if n_presc==1) */
  double wavelengths[OLRT_MAX_WAVELENGTHS]; /* distinct wavelengths, micrometres */
} olrt_sys;

/* ray state, structure of arrays; 8 device pointers each of n_rays elements.
 * in : x y z L M N intensity wavelength   (wavelength may be NULL when W==1)
 * out: x y z L M N intensity opd          (may alias `in` element-wise)        */

int olrt_version(void);
const char* olrt_build_info(void);

/* Forward trace: SurfaceGroup.trace for RealRays (surface_group.py:232-244).
 *  presc      [n_presc][presc_stride] reals
 *  ops        [ops_len] int32 (shared by all prescriptions)
 *  opd_in     initial OPD (RealRays.opd), NULL = zeros
 *  record     NULL or [n_record][8][n_rays]: x y z L M N intensity opd after each recorded surface
 *  dir0       NULL or [3][n_rays]: L0 M0 N0, local pre-interaction direction at the last surface
 *  nr_resid   NULL or [n_surfaces][OLRT_RESID_STRIDE] u64-ordered doubles, zero-initialised by the
 *             caller: a value v_j per Newton iteration j with  v_j < tol  <=>  every ray had |f_j| < tol  (the
 *             reference's exit test, newton_raphson.py:148), and v_j = +inf when a ray's f_j was NaN. The strict
 *             kernels and the fix-up pass store the maximum of |f_j|; the tolerance-mode kernel stores 1e300 where
 *             some ray failed the test and leaves 0 otherwise (one compare per ray instead of a running maximum)
 *  status     NULL or one uint32, OR-accumulated
 *  fixup      [1 + n_rays] uint32 workspace, required by _f64 (ignored by _f32): the main kernel uses
 *             unguarded division / square-root sequences and lists the rays whose result came out
 *             non-finite; a second kernel re-traces exactly those with IEEE div.rn / sqrt.rn, so the
 *             stored result is the IEEE one for every ray. fixup[0] = how many rays were re-traced.
 *             With OLRT_FWD_FAST the list also takes every ray one of whose DECISIONS (hit / miss, root choice,
 *             aperture comparison, side of the normal, total reflection, the reference's 1e-14 switches) was
 *             taken within 2^-36 (relative) of its threshold, and every ray whose conic quadratic a t^2 + 2 b t + c
 *             has |a| < 2^-20 |b| (the reference's root cancels there: a paraboloid near its axis, a "flat" surface
 *             entered with a huge radius): the re-trace is always the strict build's (the reference's operation
 *             order), so those rays carry the reference's decision / value.
 *  flags      OLRT_FWD_*                                                                  */
#define OLRT_FWD_FAST 1 /* tolerance mode (<= 1e-10 relative, identical masks; not bit-faithful): FMA contraction, for
                           lenses of planes / conics / even aspheres the reduced-algebra kernel (csrc/trace_fx.cuh), in
                           the interpreter kernel Horner forms of the polynomial, odd-asphere and (OLRT_HINT_ZERNIKE)
                           Zernike sums */
#define OLRT_FWD_REPLAY 2 /* with OLRT_FWD_FAST (fp32: always): not the reduced-algebra kernel but the straight-line /
                             interpreter kernels of the tolerance build, which replay the reference's operation order
                             for planes, conics and even aspheres with FMA contraction -- the round-1 "fast" path, kept
                             for A/B measurements */

int olrt_trace_fwd_f64(const olrt_sys* sys /*host*/, const double* presc, const int32_t* ops,
                       int64_t n_rays, const double* const in[8] /*host array of device ptrs*/,
                       const double* opd_in, double* const out[8] /*host array of device ptrs*/,
                       double* record, double* dir0, unsigned long long* nr_resid,
                       uint32_t* status, uint32_t* fixup, int flags, void* stream);
int olrt_trace_fwd_f32(const olrt_sys* sys, const float* presc, const int32_t* ops,
                       int64_t n_rays, const float* const in[8], const float* opd_in,
                       float* const out[8], float* record, float* dir0,
                       unsigned long long* nr_resid, uint32_t* status, uint32_t* fixup, int flags,
                       void* stream);

/* Forward trace that ALLOCATES NOTHING. No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid
 * it). Same arguments, kernels and results as olrt_trace_fwd_f64, minus `fixup`: the re-trace list lives in a
 * file-scope `static __device__ __align__(256)` array of 1 + OLRT_NOALLOC_MAX_RAYS uint32 inside the library (4 MiB
 * per GPU that loads libolrt.so, held until the process exits; its device address is taken once per device with
 * cudaGetSymbolAddress). n_rays > OLRT_NOALLOC_MAX_RAYS returns OLRT_E_LIMIT. The call issues no cudaMalloc / cudaFree
 * or allocating library routine: a cudaMemsetAsync of the list's counter and the kernel launches, all on `stream`.
 * EVERY call on a GPU shares the one array (its counter is cleared per call, older entries are never read): calls
 * that overlap on the same GPU (different streams, different host threads) race -- serialise them.              */
#define OLRT_NOALLOC_MAX_RAYS 1048576
int olrt_trace_fwd_noalloc_f64(const olrt_sys* sys /*host*/, const double* presc, const int32_t* ops,
                               int64_t n_rays, const double* const in[8], const double* opd_in,
                               double* const out[8], double* record, double* dir0,
                               unsigned long long* nr_resid, uint32_t* status, int flags, void* stream);

/* Backward (adjoint) of the forward trace: what torch autograd computes through Surface.trace on the
 * reference torch backend (optimization/optimizer/torch/base.py:95-154 -> loss.backward()).
 * Checkpoints = the forward's records: the forward must have recorded EVERY surface (n_record ==
 * n_surfaces, record slot s == surface s); record[s-1] is the state entering surface s.
 *  record  [n_surfaces][8][n_rays] as written by olrt_trace_fwd_*
 *  gout    [8] device ptrs or NULL: d loss / d out x y z L M N intensity opd (NULL entry = zero)
 *  grecord NULL or [n_surfaces][8][n_rays]: d loss / d recorded state (added to gout at the last surface)
 *  gin     [8] device ptrs: d loss / d in x y z L M N intensity opd_in (NULL entry = not wanted)
 *  gpresc  [n_presc][presc_stride] accumulated with atomicAdd (caller zero-initialises): filled for
 *          frame x y z, radius, conic, even-asphere / polynomial coefficients and -- general kernel, tilted frames --
 *          the frame's cos rz, sin rz, cos ry, sin ry, cos rx, sin rx entries (the caller chains them to the angles:
 *          d / d angle = -sin * g_cos + cos * g_sin, optimization/variable/tilt.py:59); other entries stay zero
 *  status  NULL or one uint32: OLRT_ST_BAD_HINT when a surface is outside the differentiable set (plane / conic /
 *          even or odd asphere / x^i y^j polynomial / Zernike surface on ONE frame, decentred and / or tilted; at most
 *          32 coefficient gradient slots per surface -- a Zernike term takes two, for its c and c N entries -- announced
 *          in bits 8-15 of olrt_sys.kernel_hint when more than 16; no coating or thin-lens flag): nothing is written then
 *  flags   0: lenses with olrt_sys.kernel_hint & OLRT_HINT_FASTKINDS, <= 3 Newton updates and <= 8 even-asphere
 *          coefficients per surface run the reduced-algebra adjoint kernel (csrc/trace_bx.cuh), every other
 *          differentiable lens (tilted frames, polynomial freeforms, more updates or coefficients: the caller passes
 *          OLRT_BWD_GENERIC for the last two) the general one (csrc/trace_bwd.cuh); OLRT_BWD_GENERIC forces it   */
#define OLRT_BWD_GENERIC 1
/* Tile-blocked checkpoints (differentiable trace without user-visible records): with OLRT_FWD_CKPT_BLOCKED the
 * forward's `record` argument is a buffer of ceil(n_rays / 256) * n_surfaces * OLRT_CKPT_VALS * 256 reals laid out
 * [tile][surface][x y z L M copysign(intensity, N)][256]: the state after every surface, 48 B per ray and surface, one
 * contiguous 12 KB block per (256-ray tile, surface). N is not stored: the backward rebuilds it as
 * copysign(sqrt(1 - L^2 - M^2), sixth value) -- the direction is a unit vector to a few ulps, the sign rides on the
 * intensity (>= 0, so its sign bit is free) -- a square root per surface for 1/7 of the checkpoint traffic. A plane as the LAST surface (the image plane) is the one
 * exception: its block is never written nor read (nothing needs its exit state). (Leaving ALL the planes out and recomputing their exit states in
 * the backward was measured: -224 B per ray of traffic but +160 instructions per ray in kernels that are
 * issue-bound, not bandwidth-bound: slower. DESIGN.md section 5.)
 * olrt_sys.n_record must be 0 (no [S][8][N] records in the same
 * launch); lenses of planes / conics / even aspheres in tolerance mode only (OLRT_E_UNSUPPORTED otherwise).
 * With OLRT_BWD_CKPT_BLOCKED the backward's `record` argument is that buffer (grecord must be NULL): a thread's
 * six values of a surface sit at fixed offsets from one pointer (cp.async copies without address arithmetic).
 * OLRT_BWD_NO_FRAME_XY: the caller does not want d loss / d (frame x, frame y) (no decenter is a variable): those two
 * columns of gpresc stay zero and the reduced-algebra kernel skips their accumulation (ignored by the general one). */
#define OLRT_FWD_CKPT_BLOCKED 4
#define OLRT_BWD_CKPT_BLOCKED 2
#define OLRT_BWD_NO_FRAME_XY 4
#define OLRT_CKPT_TILE 256
#define OLRT_CKPT_VALS 6
int olrt_trace_bwd_f64(const olrt_sys* sys, const double* presc, const int32_t* ops, int64_t n_rays,
                       const double* const in[8], const double* record, const double* const gout[8],
                       const double* grecord, double* const gin[8], double* gpresc, uint32_t* status,
                       int flags, void* stream);
int olrt_trace_bwd_f32(const olrt_sys* sys, const float* presc, const int32_t* ops, int64_t n_rays,
                       const float* const in[8], const float* record, const float* const gout[8],
                       const float* grecord, float* const gin[8], float* gpresc, uint32_t* status,
                       int flags, void* stream);

/* Spot epilogue: per-group sums over rays (ray r belongs to group r / rays_per_group).
 *  sums [n_groups][OLRT_SPOT_SUMS] doubles, accumulated (caller zero-initialises):
 *    0 count  1 sum dx  2 sum dy  3 sum dx^2  4 sum dy^2  5 max (dx^2+dy^2) (u64-ordered)
 *    6 sum intensity  7 reserved  8 shift x  9 shift y       with dx = x - shift x, dy = y - shift y
 *  shift  NULL or [n_groups][2]: subtracted before summing (pass the centroid to get the centred
 *         second moments and the geometric radius of spot_diagram.py:329-357 in a second call);
 *         NULL = the group's first ray if finite, else 0 (keeps the second moments well conditioned).
 *  flags: OLRT_SPOT_VALID_ONLY = only rays with intensity>0 (SpotDiagram, spot_diagram.py:418-425);
 *         otherwise all rays (RayOperand.rms_spot_size, ray.py:336-340).                   */
#define OLRT_SPOT_VALID_ONLY 1
int olrt_spot_sums_f64(const double* x, const double* y, const double* intensity, int64_t n_rays,
                       int64_t rays_per_group, int32_t n_groups, const double* shift, int flags,
                       double* sums, void* stream);
int olrt_spot_sums_f32(const float* x, const float* y, const float* intensity, int64_t n_rays,
                       int64_t rays_per_group, int32_t n_groups, const double* shift, int flags,
                       double* sums, void* stream);

/* Benchmark helper: dependent-free FMA chains on every SM; writes the measured
 * flop count basis to `sink`. Used by bench.py to measure the FP64 / FP32 FMA peak. */
int olrt_fma_peak_f64(double* sink, int64_t iters, int blocks, int threads, void* stream);
int olrt_fma_peak_f32(float* sink, int64_t iters, int blocks, int threads, void* stream);

/* Ray launch: RayGenerator.generate_rays (optiland/rays/ray_generator.py:28-106, the
 * non-telecentric branch) with the ray origins of AngleField.get_ray_origins (optiland/fields/field_types.py:139-169:
 * object at infinity or finite) and ObjectHeightField.get_ray_origins (:245-272) and the per-field vignetting
 * factors of FieldGroup.get_vig_factor (fields/field_group.py:80-109). Everything per FIELD is a host scalar
 * (a handful of values per field), everything per RAY is in the kernel, in the reference's operation order.
 * Rays are field-major (ray f * n_pupil + p), as RealRayTracer.trace builds them (raytrace/real_ray_tracer.py:67-77).
 *  Px, Py     [n_pupil] doubles (device): normalised pupil coordinates
 *  field_xy   [n_fields][2] doubles (device), x then y.
 *             field_z == NULL (object at infinity): -tan(radians(max_field * H)) * (offset + EPL); the ray starts
 *             at (Px EPD / 2 vx + field_x, Py EPD / 2 vy + field_y, z_start), z_start = positions[1] - offset.
 *             field_z != NULL (finite object): the object POINT of the field, (field_x, field_y, field_z[f]) --
 *             -tan(radians(max_field * H)) * (EPL - z_obj) on the object plane for angle fields, (max_field * H,
 *             sag + z_obj) for object-height fields; every ray of the field starts there; z_start is not read.
 *  field_vig  NULL (no vignetting), or [n_fields][2] doubles (device): vx = 1 - vig_x, vy = 1 - vig_y
 *             (ray_generator.py:50-52); the ray aims at (Px EPD vx / 2, Py EPD vy / 2, EPL).
 *  EPD, EPL   entrance pupil diameter / location (paraxial.py:207-280)
 *  out        [8] device ptrs x y z L M N intensity wavelength (entry 7 may be NULL); fp64 arithmetic, stored
 *             in the output precision. p_lo / p_len are ignored here (all n_pupil points are launched).        */
typedef struct olrt_pupil {
  const double* Px;        /* [n_pupil] normalised pupil coordinates (device) */
  const double* Py;
  const double* field_xy;  /* [n_fields][2] (device) */
  const double* field_vig; /* NULL or [n_fields][2] (device) */
  const double* field_z;   /* NULL or [n_fields] (device) */
  int64_t n_pupil;         /* pupil points per field */
  int64_t p_lo, p_len;     /* olrt_trace_pupil_*: the pupil range this call traces */
  int32_t n_fields;
  int32_t reserved0;
  double EPD, EPL, z_start, wavelength;
  /* olrt_trace_pupil_* only, streamed pupil grid: NULL, or [ceil(n_pupil / ready_chunk)] device words that a copy
   * engine sets to non-zero, in order, as the host->device copy of pupil points [c * ready_chunk, (c+1) *
   * ready_chunk) of Px and Py lands (a 4-byte cudaMemcpyAsync queued behind each chunk's copies on the copy stream).
   * ONE launch then covers the whole grid while it is still arriving: the kernel's CTAs walk the pupil in copy
   * order and wait on the word of the chunk they reach. ready_chunk must be a multiple of 256; needs
   * olrt_sys.n_presc == 1, p_lo == 0. A word that stays zero for ~2 s sets OLRT_ST_TIMEOUT and ends the launch. */
  const uint32_t* ready;
  int64_t ready_chunk;
} olrt_pupil;
int olrt_generate_rays_f64(const olrt_pupil* pupil, double* const out[8], void* stream);
int olrt_generate_rays_f32(const olrt_pupil* pupil, float* const out[8], void* stream);

/* Fused ray launch + trace + spot epilogue ("pupil mode"): what RealRayTracer.trace does per call
 * (raytrace/real_ray_tracer.py:61-82: generate the rays of every field from one pupil grid, then
 * SurfaceGroup.trace) followed, optionally, by the sums RayOperand.rms_spot_size / SpotDiagram need
 * (optimization/operand/ray.py:336-340, analysis/spot_diagram.py:317-357) -- in ONE kernel: the launched rays
 * never exist in HBM, and with out == NULL nothing per ray is written at all (merit-function evaluation).
 * One call traces pupil points [p_lo, p_lo + p_len) of EVERY field, so a host pupil grid can be streamed in
 * chunks; ray (field f, pupil point p) lives at index f * n_pupil + p of the output arrays, the field-major
 * order of real_ray_tracer.py:67-77, and belongs to spot group f.
 * With olrt_sys.n_presc > 1 (Monte-Carlo tolerancing, tolerancing/monte_carlo.py:60-123: every trial re-traces
 * the same fields and pupil grid through a perturbed prescription) the groups are g = trial * n_fields + f:
 * arrays hold n_presc * n_fields * n_pupil rays, spot_sums / spot_shift n_presc * n_fields rows; with
 * out == NULL the whole tolerancing run moves no per-ray data at all.
 * Implemented by the tolerance-mode kernel for lenses of planes / conics / even aspheres (flags must contain
 * OLRT_FWD_FAST for fp64; olrt_sys.kernel_hint & OLRT_HINT_FASTKINDS; one wavelength, no
 * per-surface records): anything else returns OLRT_E_UNSUPPORTED and the caller uses
 * olrt_generate_rays_* + olrt_trace_fwd_* + olrt_spot_sums_*.                                              */
/*  out        NULL, or [8] device ptrs to arrays of n_fields * n_pupil (x y z L M N intensity opd)
 *  dir0       NULL or [3][n_fields * n_pupil]
 *  spot_sums  NULL or [n_fields][OLRT_SPOT_SUMS], zeroed by the caller before the first chunk; columns as for
 *             olrt_spot_sums_* except 8, 9 (the caller knows the shift); spot_shift [n_fields][2] is required
 *  spot_flags OLRT_SPOT_*; flags OLRT_FWD_*; the other arguments as for olrt_trace_fwd_*
 *  group_k    NULL, or [groups][n_surfaces] int32 (device): Newton trip counts PER GROUP instead of ops word 6.
 *             The reference traces every (trial, field) on its own, so its exit rule (max |f| over that trace's
 *             rays < tol, newton_raphson.py:148) gives each group its own K_s (<= OLRT_GROUP_RESID - 1 here).
 *             Ops word 6 must hold the MAXIMUM of the groups' counts for every Newton surface (a group never takes
 *             more than that many updates). Allowed together with a streamed grid when olrt_sys.nr_resid_words <= 32.
 *  group_resid NULL, or [groups][n_surfaces][OLRT_GROUP_RESID] u64 (device, zeroed by the caller): v_j per group,
 *             surface and iteration with "v_j < tol <=> every ray of the group converged at iteration j" (bits of a
 *             non-negative double: the true maximum of |f|, or OLRT_RESID_FAILED for a failing slot when the counts
 *             fit one register -- nr_resid_words <= 32 -- exactly as for nr_resid), for the caller to check K_s with. */
/* Host -> device upload of a pinned pupil grid for the streamed mode above, in ONE call: for every chunk c of
 * `ready_chunk` points, cudaMemcpyAsync of Px and Py followed by a 4-byte copy that sets ready[c] (from `one_host`,
 * a pinned non-zero word), all on `copy_stream` -- the copy engine executes them in order, so ready[c] != 0 implies
 * the chunk's data is in device memory. The caller zeroes `ready` (ordered before this call's stream) and launches
 * olrt_trace_pupil_* with the same ready / ready_chunk on its compute stream AFTER this call returns.           */
int olrt_upload_pupil(double* Px_dev, double* Py_dev, const double* Px_host, const double* Py_host,
                      int64_t n_pupil, int64_t ready_chunk, uint32_t* ready, const uint32_t* one_host,
                      void* copy_stream);
#define OLRT_GROUP_RESID 8
int olrt_trace_pupil_f64(const olrt_sys* sys, const double* presc, const int32_t* ops, const olrt_pupil* pupil,
                         double* const out[8], double* dir0, unsigned long long* nr_resid, uint32_t* status,
                         uint32_t* fixup, double* spot_sums, const double* spot_shift, int spot_flags,
                         const int32_t* group_k, unsigned long long* group_resid, int flags, void* stream);
int olrt_trace_pupil_f32(const olrt_sys* sys, const float* presc, const int32_t* ops, const olrt_pupil* pupil,
                         float* const out[8], float* dir0, unsigned long long* nr_resid, uint32_t* status,
                         uint32_t* fixup, double* spot_sums, const double* spot_shift, int spot_flags,
                         const int32_t* group_k, unsigned long long* group_resid, int flags, void* stream);

/* Random pupil distribution generated on the device: RandomDistribution.generate_points
 * (optiland/distribution.py:132-158) with the NumPy backend's generator, numpy.random.default_rng(seed) = PCG64:
 * r = uniform(size = n), theta = uniform(0, 2 pi, size = n), Px = sqrt(r) cos(theta), Py = sqrt(r) sin(theta).
 * `Optic.trace(Hx, Hy, wavelength, num_rays, "random")` then moves a seed, not 16 bytes per pupil point, to the GPU.
 *  state, inc [host] the generator's 128-bit state and increment BEFORE its first draw, high word first
 *             (numpy: default_rng(seed).bit_generator.state["state"]["state" | "inc"]; the seed hashing stays in NumPy)
 *  Px, Py     [n] doubles (device)
 *  raw        NULL, or [2][n] doubles (device): the draws r and theta themselves (NumPy's, bit for bit)      */
int olrt_pupil_random_f64(const uint64_t state[2] /*host*/, const uint64_t inc[2] /*host*/, int64_t n, double* Px,
                          double* Py, double* raw, void* stream);
/* Hexapolar pupil distribution generated on the device: HexagonalDistribution.generate_points
 * (optiland/distribution.py:189-220) -- the centre, then ring i = 1 .. num_rings with 6 i points at radius
 * linspace(0, 1, num_rings + 1)[i] and angles linspace(0, 2 pi, 6 i + 1)[:-1]; radii and angles are NumPy's bit for bit,
 * Px, Py within an ulp or two (cos / sin).  Px, Py: [1 + 3 num_rings (num_rings + 1)] doubles (device).        */
int olrt_pupil_hexapolar_f64(int32_t num_rings, double* Px, double* Py, void* stream);
/* Uniform pupil distribution generated on the device: UniformDistribution.generate_points (optiland/distribution.py:
 * 174-185) -- x = linspace(-1, 1, num_points), the num_points x num_points grid masked to x^2 + y^2 <= 1, rows (y)
 * first; NumPy's values and mask bit for bit. Two passes, because the number of kept points is data:
 *   row_count != NULL (row_offset NULL): writes the number of kept points of every row, [num_points] int64 (device);
 *   row_offset != NULL (row_count NULL): [num_points] int64 (device), the exclusive prefix sum of those counts, which
 *     the caller forms (it also needs their total to size Px, Py): writes the points.                           */
int olrt_pupil_uniform_f64(int64_t num_points, const int64_t* row_offset, int64_t* row_count, double* Px, double* Py,
                           void* stream);

/* OPD / wavefront epilogue: per ray, the optical path from the image surface back to the
 * reference sphere, the launch-plane tilt correction and the OPD in waves relative to the chief ray --
 * ReferenceStrategy._opd_image_to_xp (optiland/wavefront/strategy.py:68-116), _correct_tilt (:118-166) and steps
 * 4-5 of ChiefRayStrategy.compute_wavefront_data (:203-229), in the reference's operation order (fp64: the same
 * bits for the same traced rays). The reference sphere (centre = chief-ray image point, radius to the paraxial
 * exit pupil, :231-240) and opd_ref are per-field scalars the caller derives from the chief ray with the same
 * kernel (OLRT_OPD_RAW).
 *  rays   [8] device ptrs x y z L M N intensity opd AT THE IMAGE SURFACE (intensity is not read)
 *  Px, Py NULL, or [n] normalised pupil coordinates of the rays (tilt term; a chief ray passes NULL = 0, 0)
 *  opd_out [n]: OPD in waves, (opd_ref - opd) / (wavelength * 1e-3); with OLRT_OPD_RAW the tilt-corrected
 *          opd (mm) itself, which is what opd_ref is for the chief ray
 *  pupil  NULL or [3][n]: intersection with the reference sphere (pupil_x, pupil_y, pupil_z)              */
typedef struct olrt_refsphere {
  double xc, yc, zc, R;    /* reference sphere */
  double n_image;          /* index of the image space */
  double opd_ref;          /* chief-ray OPD (mm), ignored with OLRT_OPD_RAW */
  double wavelength;       /* micrometres */
  double tilt_ux, tilt_uy; /* direction cosines of the field (angle fields), 0 0 otherwise */
  double epd;              /* entrance pupil diameter */
} olrt_refsphere;
#define OLRT_OPD_RAW 1
int olrt_opd_epilogue_f64(const double* const rays[8], const double* Px, const double* Py, int64_t n,
                          const olrt_refsphere* sphere, double* opd_out, double* pupil, int flags, void* stream);
int olrt_opd_epilogue_f32(const float* const rays[8], const double* Px, const double* Py, int64_t n,
                          const olrt_refsphere* sphere, float* opd_out, float* pupil, int flags, void* stream);

/* Self-test of the kernels' own IEEE division / square-root sequences (csrc/math64.cuh):
 * quot[i] = a[i] / b[i], root[i] = sqrt(a[i]) as the trace kernels compute them; the caller compares
 * them bit for bit with an IEEE implementation (tests/test_gpu_math.py). */
int olrt_selftest_math_f64(const double* a, const double* b, double* quot, double* root, int64_t n,
                           void* stream);

#ifdef __cplusplus
}
#endif
#endif /* OLRT_H */
