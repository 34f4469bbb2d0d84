// celeste_b200.cu -- sm_100a kernels + C ABI (include/celeste_b200.h)
//
// Hot path of CelestePy re-designed for B200: render point sources / galaxies
// as PSF-convolved mixtures of Gaussians onto band images, Poisson pixel
// log-likelihood, batched MH proposal scoring.  Reference citations are
// relative to the HIPS/DESI-MCMC repository root.
//
// Kernel plan:
//   prepare_kernel   one warp per (source, image) item: WCS, galaxy affine
//                    transform, PSF (x) profile convolution table, footprint,
//                    flux -- all in float64, table rounded to fp32 as an
//                    integer-coordinate quadratic in log2 units.
//   render_kernel    persistent CTAs pull 1024-pixel strips from a device work
//                    list; thread = 8 consecutive pixels of a row; per
//                    component 3 FMA hoisted per row + (2 FMA, 1 MUFU.EX2,
//                    1 FADD) per pixel.  Epilogues: write patch | Poisson ll
//                    against nelec/background/mask | conditional photon ll.
//   field_reduce     gather: 64x16 pixel tiles sum the overlapping patches in
//                    source order (deterministic), Poisson ll in fp64.
//   score_finalize   delta / accept per proposal, guard-band list for the
//                    float64 re-score.
//
// Compiled with -fmad=false: every fused multiply-add below is explicit so
// the float64 preparation follows the reference's (unfused) NumPy arithmetic.

#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdarg.h>
#include <string.h>
#include <math.h>
#include <atomic>
#include <string>
#include <vector>
#include <algorithm>
#include <map>
#include <mutex>

#include "celeste_b200.h"

// ---------------------------------------------------------------------------
// error plumbing
// ---------------------------------------------------------------------------
static thread_local std::string g_err;
static std::atomic<long long> g_launches{0};
static int g_device = -1;
static int g_sm_count = 0;

static int fail(int code, const char* fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof buf, fmt, ap);
  va_end(ap);
  g_err = buf;
  return code;
}
#define CUDA_TRY(expr)                                                            \
  do {                                                                            \
    cudaError_t _e = (expr);                                                      \
    if (_e != cudaSuccess)                                                        \
      return fail(CG_ECUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), \
                  __FILE__, __LINE__);                                            \
  } while (0)
#define LAUNCH_CHECK()                                                            \
  do {                                                                            \
    g_launches.fetch_add(1, std::memory_order_relaxed);                           \
    cudaError_t _e = cudaGetLastError();                                          \
    if (_e != cudaSuccess)                                                        \
      return fail(CG_ECUDA, "kernel launch failed: %s (%s:%d)",                   \
                  cudaGetErrorString(_e), __FILE__, __LINE__);                    \
  } while (0)
// Checked build (python -m desi_mcmc_b200.build --checked -> libceleste_b200_checked.so, loaded when
// CG_CHECKED=1): every index into a workspace array, a shared-memory table or a box-shaped patch is
// asserted in the kernels; a violation prints its source line and traps (the launch fails with an
// "unspecified launch failure", which the tests see).  compute-sanitizer is closed on this GPU pool
// (profiles/r02_sanitizer_refused.log); this is the bounds evidence that can be produced here.
#ifdef CG_CHECKED
#define CG_CHECK(cond)                                                                       \
  do {                                                                                       \
    if (!(cond)) {                                                                           \
      printf("CG_CHECK failed: %s (%s:%d) block %d thread %d\n", #cond, __FILE__, __LINE__,  \
             (int)blockIdx.x, (int)threadIdx.x);                                             \
      __trap();                                                                              \
    }                                                                                        \
  } while (0)
#else
#define CG_CHECK(cond) ((void)0)
#endif
#define NEED_INIT()                                                               \
  do {                                                                            \
    if (g_device < 0) return fail(CG_ESTATE, "cg_init() has not been called");    \
  } while (0)

// ---------------------------------------------------------------------------
// constants
// ---------------------------------------------------------------------------
#define LOG2PI_D 1.8378770664093453     /* gmm_like_fast.pyx:29 */
#define LOG2E_D 1.4426950408889634
#define LN2_D 0.6931471805599453
#define PI_D 3.141592653589793

// mixture_profiles.py:9-19 (raw, unnormalised; normalised at cg_init exactly
// as the reference does: amp /= sum(amp)).
static const double H_EXP_AMP_RAW[6] = {2.34853813e-03, 3.07995260e-02, 2.23364214e-01,
                                        1.17949102e+00, 4.33873750e+00, 5.99820770e+00};
static const double H_EXP_VAR[6] = {1.20078965e-03, 8.84526493e-03, 3.91463084e-02,
                                    1.39976817e-01, 4.60962500e-01, 1.50159566e+00};
static const double H_DEV_AMP_RAW[8] = {4.26347652e-02, 2.40127183e-01, 6.85907632e-01,
                                        1.51937350e+00, 2.83627243e+00, 4.46467501e+00,
                                        5.72440830e+00, 5.60989349e+00};
static const double H_DEV_VAR[8] = {2.23759216e-04, 1.00220099e-03, 4.18731126e-03,
                                    1.69432589e-02, 6.84850479e-02, 2.87207080e-01,
                                    1.33320254e+00, 8.40215071e+00};

struct ProfileConsts {
  double exp_amp[6], exp_var[6], dev_amp[8], dev_var[8];
};
__constant__ ProfileConsts c_prof;

#define MAX_FILTER_ROWS 1024
struct FilterTables {
  int count[CG_NUM_BANDS];
  int start[CG_NUM_BANDS];
  int loaded;
  double wl[MAX_FILTER_ROWS];    // metres, sorted per band
  double hvec[MAX_FILTER_ROWS];  // planck.py:52-54
  double sens[MAX_FILTER_ROWS];
};
static FilterTables* g_d_filters = nullptr;  // device
struct LogTable;
static LogTable* g_d_logtab = nullptr;       // device, filled at cg_init
static int g_filters_loaded = 0;
static const double* g_flux_table = nullptr;  // device [g_flux_rows][5] nanomaggies (cg_set_flux_table)
static long long g_flux_rows = 0;

// ---------------------------------------------------------------------------
// device-side structures
// ---------------------------------------------------------------------------
struct DevImage {
  int H, W, band, n_psf;
  double epsilon, kappa, calib;
  double psf_w[CG_MAX_PSF];
  double psf_mu[CG_MAX_PSF][2];
  double psf_cov[CG_MAX_PSF][3];
  double rho_n[2], phi_n[2];
  double ups[4], ups_inv[4];
  double R;         // FitsImage.R, fits_image.py:151-155
  double psf_wsum;  // np.sum(fits_img.weights), sources.py:161,174
  const float* nelec;
  const float* mask;
};

struct CgImage {
  DevImage h;     // host copy
  DevImage* d;    // device copy
};

struct ImgPtrs {
  const DevImage* p[CG_MAX_IMAGES];
};

struct ItemHdr {
  int x0, x1, y0, y1;  // footprint (half-open)
  int ox, oy;          // local origin (pixel nearest the source centre)
  int K;               // components (0: renders nothing)
  int img;
  int strip_start, n_strips;   // first partial-sum slot / number of warp tiles
  int aux;
  int flags;
  double flux;         // rate multiplier (1 for unit-flux patches)
  double norm_term;    // band_flux * sum(psf weights), sources.py:174
};

struct Counters {
  int total_strips;    // partial-sum slots handed out
  int work_counter;
  int guard_count;
  int err_flag;        // 1: footprint larger than max_patch_px, 2: guard list overflow, 3: work list overflow, 4: CG_SRC_EIGEN row outside the flux table, 5: CG_SRC_PLANCK without filter curves
  unsigned long long evals;    // algorithmic: footprint pixels x components
  unsigned long long issued;   // lane-evaluations actually issued after culling (warp granular)
  int heavy_strips;    // work units of galaxies, listed from the front of the work list
  int light_strips;    // work units of stars (K <= 4), listed from the back
  int pad[2];
};

// Component table: 12 values per component (three 16-byte loads):
//   A, B, C, b10 | c01, c00, mx, my | qs, kx, ky, sl
//   the quadratic A x^2 + B x y + C y^2 + b10 x + c01 y + c00 in log2 units,
//   footprint-local integer coordinates, followed by the culling record: mean
//   (mx, my), peak value qs, kx = 2 Sxx / log2(e), ky = 2 Syy / log2(e)
//   (half-widths^2 of the level-set bounding box per unit of log2 drop) and
//   sl = log2(e) / (2 lambda_min(Sigma)) (steepest log2 drop per squared pixel).
#define TAB_W 12
#ifndef FT_RECUR
#define FT_RECUR 1          // row recurrence compiled in (run-time switch: CG_OPT_ROW_RECURRENCE)
#endif
#define FT_RECUR_DMAX 96.f   // largest |log2 ratio of neighbouring pixels| the row recurrence is used for
#define FT_RECUR_FLOOR -64.f // log2 of the unit-flux rate below which a value lost to underflow does not count
#define TAB_STRIDE (CG_MAX_COMP * TAB_W)
static_assert(14 * CG_MAX_PSF <= CG_MAX_COMP, "a galaxy has (6 exp + 8 dev) x n_psf components");
static_assert(CG_MAX_COMP < 64, "bin-list entries pack item * 64 + K; cull_components owns components lane and lane + 32");

#define STRIP_THREADS 256
#define PIX_PER_THREAD 4
#ifndef PREP_WARPS
// items (warps) per prepare_kernel CTA.  ONE warp at 64 registers = 2048 registers per CTA: two such CTAs fit
// beside the persistent field_tile_kernel (3 x 256 x 80 of the SM's 65536 registers), so the set-up of the next
// field set runs under the tile kernel of the previous one when the sets are on different streams
// (64 fields: 5.76 -> 5.67 ms; alone the kernel is as fast as with 4 warps per CTA at 96 registers).
#define PREP_WARPS 1
#endif
// Work decomposition of a footprint.  A LANE UNIT is two 4-pixel chunks in
// the same columns of two rows of a 16-row tile band (rows r and r + R/2 of a
// band of R rows); lane units are numbered in TILE-MAJOR order (tiles of
// TILE_CW chunks x TILE_R row pairs = 16 x 16 pixels; edge tiles are narrower /
// shorter and simply hold fewer units, so no lane is spent on padding tiles).
// A WARP TILE is 32 consecutive lane units (256 pixels in a compact block, so
// whole components can be culled for the warp); a work UNIT is up to UNIT_WT
// consecutive warp tiles of one item and is processed by one warp.
#define TILE_CW 4
#define TILE_R 8
#define WT_PIX (32 * 2 * PIX_PER_THREAD)
static_assert(PIX_PER_THREAD == 4, "recur_row4 and the float4 pixel groups assume 4 pixels per lane");
#define UNIT_WT_MAX 8
// a component is skipped for a warp when its largest value over the warp's
// pixels is below 2^-CULL_T of a lower bound of the mixture there
#ifdef CULL_T_OVERRIDE
#define CULL_T CULL_T_OVERRIDE
#else
#define CULL_T 27.0f
#endif
// photon allocation: image tiles of 64 x 16 pixels, per-tile source lists (the field path bins 16 x 16, see BIN_W)
#define TILE_W 64
#define TILE_H 16
#define TILE_LIST_CAP 64
#define FR_WARPS 4          // warps per block of the photon-allocation kernel (one per half tile)

#define CG_MAX_FIELDS 16
struct FieldBatch {   // implicit_mode 3: independent fields, each a cross product of its sources x images
  int n_fields;
  int item_off[CG_MAX_FIELDS + 1], src_off[CG_MAX_FIELDS + 1], img_off[CG_MAX_FIELDS + 1];
};

struct PrepArgs {
  ImgPtrs imgs;
  FieldBatch fb;
  int n_img;
  const CgSource* srcs;
  int n_src;
  const CgItem* items;      // NULL: implicit
  long long n_items;
  int default_flags;
  int implicit_mode;        // 0 cross product; 1 per-image fixed boxes + aux=img; 2 indirect list; 3 field batch
  CgBox img_boxes[CG_MAX_IMAGES];   // by value (implicit_mode 1, 2): no host->device copy, graph-capturable
  const int* ind_list;      // device (implicit_mode 2)
  const int* ind_count;     // device (implicit_mode 2): points at Counters.guard_count of pass 1
  int aux_per_img;          // implicit_mode >= 1: aux patch index = image index (else -1)
  double bound_cap;
  long long max_patch_px;
  ItemHdr* hdr;
  void* tab;                // float or double, n_items * TAB_STRIDE
  int2* worklist;
  int worklist_cap;
  Counters* cnt;
  CgBox* out_boxes;         // optional
  int cap_boxes;            // footprints above max_patch_px are an error (set when box-shaped buffers are written)
  const FilterTables* filters;
  const double* flux_table;  // CG_SRC_EIGEN: [flux_rows][5] nanomaggies resident in HBM
  long long flux_rows;
  // full-field path: per (image, 64x16 tile) list of the items that overlap it
  int unit_wt;              // warp tiles per work unit (same value in RenderArgs)
  int* tile_count;          // [n_img][max_tiles] (NULL: no binning)
  int* tile_list;           // [n_img][max_tiles][bin_cap]
  int max_tiles;
  int bin_w, bin_h, bin_cap;
};

// ---------------------------------------------------------------------------
// small device helpers
// ---------------------------------------------------------------------------
__device__ __forceinline__ float ex2_approx(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float rexp2(float x) { return ex2_approx(x); }
__device__ __forceinline__ double rexp2(double x) { return exp2(x); }
__device__ __forceinline__ float rfma(float a, float b, float c) { return fmaf(a, b, c); }
__device__ __forceinline__ double rfma(double a, double b, double c) { return fma(a, b, c); }
// Row recurrence, underflow side.  ex2.approx.ftz returns 0 below 2^-126, and a lane whose
// first pixel x0 flushed carries 0 to its other three pixels.  Along a row the component is
// 2^(qr - |A| t^2), t = distance to the row's peak, qr <= qs (its global peak).  A flushed
// start means t > D = sqrt((qs + 124) / |A|); the largest value the lane can then lose is at
// x0 + 3: qs - |A| (D - 3)^2 = -124 + 6 |A| D - 9 |A| (the peak itself if D < 3).  A
// component is marked unsafe for the recurrence (prepare_kernel, sign of table entry 9) when
// that exceeds FT_RECUR_FLOOR, i.e. for Gaussians narrower than about 0.8 pixel.
__device__ __forceinline__ bool recur_flush_unsafe(double A, double qs) {
  const double aA = fabs(A);
  const double AD = sqrt(aA * fmax(qs + 124., 0.));               // |A| D
  const double lost = (AD >= 3. * aA) ? (-124. + 6. * AD - 9. * aA) : qs;
  return !(lost <= (double)FT_RECUR_FLOOR);
}
// four consecutive pixels of a row from the first value e, the neighbour ratio g and the
// ratio of ratios h (see "Row recurrence" at ft_eval2r)
template <typename T>
__device__ __forceinline__ void recur_row4(T e, T g, T h, T (&acc)[4]) {
  acc[0] += e;
  e *= g; g *= h;
  acc[1] += e;
  e *= g; g *= h;
  acc[2] += e;
  e *= g;
  acc[3] += e;
}

#ifndef FT_PACKED
#define FT_PACKED 1          // FMUL2 / FADD2 row pairs in the row-recurrence loops (ft_eval2r, render_kernel)
#endif
#ifndef FT_PACKQ
#define FT_PACKQ 1           // ... and FFMA2 for the two rows' quadratics in ft_eval2r
#endif
// packed FP32 pairs (sm_100 FMUL2 / FADD2): one issue slot for the same operation on two
// floats held in an aligned register pair, each half rounded exactly like the scalar op
typedef unsigned long long f32x2;
__device__ __forceinline__ f32x2 pk2(float lo, float hi) {
  f32x2 r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
  return r;
}
__device__ __forceinline__ void unpk2(f32x2 v, float& lo, float& hi) {
  asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ f32x2 fma2(f32x2 a, f32x2 b, f32x2 c) {
  f32x2 r;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
  return r;
}
__device__ __forceinline__ f32x2 mul2_raw(f32x2 a, f32x2 b) {
  f32x2 r;
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}
__device__ __forceinline__ f32x2 add2_raw(f32x2 a, f32x2 b) {
  f32x2 r;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}
#ifndef FT_FMA_ONLY
#define FT_FMA_ONLY 0
#endif
// FT_FMA_ONLY: products and sums issued as FFMA2 (a * b + (-0) and a * 1 + b round exactly like
// a * b and a + b, signed zeros included)
// (the constants live in constant memory so that ptxas cannot fold the FFMA2 back into FMUL2 / FADD2)
__constant__ unsigned long long c_pk_negz = 0x8000000080000000ull, c_pk_one = 0x3f8000003f800000ull;
__device__ __forceinline__ f32x2 mul2(f32x2 a, f32x2 b) {
#if FT_FMA_ONLY
  return fma2(a, b, c_pk_negz);
#else
  return mul2_raw(a, b);
#endif
}
__device__ __forceinline__ f32x2 add2(f32x2 a, f32x2 b) {
#if FT_FMA_ONLY
  return fma2(a, c_pk_one, b);
#else
  return add2_raw(a, b);
#endif
}
// recur_row4 on two rows at once: e = (row r, row r + 8), likewise g and acc; h the same on both
__device__ __forceinline__ void recur_row4x2(f32x2 e, f32x2 g, f32x2 h, f32x2 (&acc)[4]) {
  acc[0] = add2(acc[0], e);
  e = mul2(e, g); g = mul2(g, h);
  acc[1] = add2(acc[1], e);
  e = mul2(e, g); g = mul2(g, h);
  acc[2] = add2(acc[2], e);
  e = mul2(e, g);
  acc[3] = add2(acc[3], e);
}

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ double warp_max(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmax(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// log(lam) for a positive normal float, absolute error ~3e-9: a float64 table
// t[idx] = (e - 127) ln 2 + log(c_i), idx = the float's exponent and top five
// mantissa bits (c_i = centre of the mantissa bin), built once at cg_init and
// L1-resident (a workload touches a few dozen entries), plus a degree-5 log1p
// of the residual r = (m - c_i) / c_i, |r| <= 1/64, in fp32.
struct LogTable {
  double t[8192];
};
__device__ __forceinline__ float rcp_approx(float x) {
  float y;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float log_residual(float lam) {
  const int bits = __float_as_int(lam);
  const float m = __int_as_float((bits & 0x007fffff) | 0x3f800000);   // [1, 2)
  const float c = __int_as_float((bits & 0x007c0000) | 0x3f820000);   // 1 + (2i+1)/64
  const float r = (m - c) * rcp_approx(c);
  float p = fmaf(r, 0.2f, -0.25f);
  p = fmaf(r, p, 0.33333334f);
  p = fmaf(r, p, -0.5f);
  p = fmaf(r, p, 1.0f);
  return r * p;
}
__device__ __forceinline__ double log_table(float lam, const LogTable* t) {
  return __ldg(&t->t[(__float_as_int(lam) >> 18) & 8191]);
}
__device__ __forceinline__ double log_accurate(float lam, const LogTable* t) {
  return __ldg(&t->t[(__float_as_int(lam) >> 18) & 8191]) + (double)log_residual(lam);
}
__device__ __forceinline__ double log_accurate(double lam, const LogTable*) { return log(lam); }

// util/bound/bounding_box.py:13-26 for one component (same operation order).
__host__ __device__ inline double major_axis(double s00, double s11, double s01, double Rsq) {
  double sigma1 = sqrt(s00);
  double sigma2 = sqrt(s11);
  double rho = s01 / (sigma1 * sigma2);
  double A11 = sigma1;
  double A21 = rho * sigma2;
  double A22 = sigma2 * sqrt(1 - rho * rho);
  double An = 1. / Rsq * (1 / (A11 * A11) + (A21 * A21) / (A22 * A22));
  double Bn = 1. / Rsq * (-2 * A21 / (A11 * (A22 * A22)));
  double Cn = 1. / Rsq * 1. / (A22 * A22);
  return 1.0 / sqrt(0.5 * (An + Cn - sqrt(Bn * Bn + (An - Cn) * (An - Cn))));
}

// fits_image.py:176-181
__device__ inline void pixel2equa(const DevImage& im, double x, double y, double* ra, double* dec) {
  double phi1rad = im.phi_n[1] / 180. * PI_D;
  double dx = x - im.rho_n[0], dy = y - im.rho_n[1];
  double i0 = im.ups[0] * dx + im.ups[1] * dy;
  double i1 = im.ups[2] * dx + im.ups[3] * dy;
  *ra = i0 / cos(phi1rad) + im.phi_n[0];
  *dec = i1 + im.phi_n[1];
}

// planck.py:82-92,139-140 for one band; the warp strides over wavelengths.
__device__ double planck_photons_per_joule_warp(const FilterTables* f, double T, int band, int lane) {
  const double h = 6.6260693e-34, k = 1.3806488e-23, c = 299792458.0;
  const double hc_k = (h * c) / k;
  const double hpicc2 = 2 * PI_D * h * c * c;
  const double pi5 = PI_D * PI_D * PI_D * PI_D * PI_D;
  const double sigma = 2 * pi5 * (k * k * k * k) / (15 * (c * c) * (h * h * h));
  if (!(T > 0)) return 0.0;
  double T4 = T * T * T * T;
  double total = sigma * T4;
  int n = f->count[band], s = f->start[band];
  double acc = 0.0;
  for (int i = lane; i < n; i += 32) {
    double w = f->wl[s + i];
    double w5 = w * w * w * w * w;
    double rad = hpicc2 / (w5 * (exp(hc_k / (w * T)) - 1));
    acc += (rad / total) * f->hvec[s + i];
  }
  return warp_sum(acc);  // lane 0 holds the sum
}

// The float64 part of one (source, image) item, by one WARP: pixel position, galaxy transformation,
// the convolved component table (rows written to `tab`: global memory in prepare_kernel, shared
// memory in the on-device samplers) and the footprint radius.  prepare_kernel and
// slice_location_kernel share it, so the tables -- and everything computed from them -- are
// bit-identical on both paths.
struct ItemGeom {
  double px, py, bound, flux;
  int K, ox, oy, err;
  bool star, trunc_box;
};
template <typename real>
__device__ __forceinline__ ItemGeom item_table(const DevImage& im, const CgSource& s, int flags, int lane, real* tab) {
  ItemGeom g;
  g.err = 0;
  const int H = im.H, W = im.W;
  // v_s = image.equa2pixel(u)                       (fits_image.py:166-174)
  const double phi1rad = im.phi_n[1] / 180. * PI_D;
  const double iwc0 = (s.u[0] - im.phi_n[0]) * cos(phi1rad);
  const double iwc1 = (s.u[1] - im.phi_n[1]);
  const double px = (im.ups_inv[0] * iwc0 + im.ups_inv[1] * iwc1) + im.rho_n[0];
  const double py = (im.ups_inv[2] * iwc0 + im.ups_inv[3] * iwc1) + im.rho_n[1];

  const int n_psf = im.n_psf;
  int K = 0;
  double bound = 0.0;
  bool trunc_box = true;
  double Wm[3] = {0, 0, 0};  // W = Tinv Tinv^T : xx, yy, xy
  const bool star = (s.type == 0);
  int n_exp = 0, n_dev = 0;
  double th_exp = 1.0, th_dev = 1.0;

  if (star) {
    // celeste.py:130-134 overlap test (v_s[0] against both dims, as written)
    bool off = (px < -50 || px > 2.0 * H || py < -50 || px > 2.0 * W);
    K = off ? 0 : n_psf;
    bound = im.R;
  } else {
    // cd: FitsImage.cd_at_pixel (fits_image.py:196-216) or header CD
    double cd[4];
    if (flags & CG_ITEM_HEADER_CD) {
      cd[0] = im.ups[0]; cd[1] = im.ups[1]; cd[2] = im.ups[2]; cd[3] = im.ups[3];
    } else {
      double ra0, dec0, rax, decx, ray, decy;
      const double step = 10.;
      pixel2equa(im, px, py, &ra0, &dec0);
      pixel2equa(im, px + step, py, &rax, &decx);
      pixel2equa(im, px, py + step, &ray, &decy);
      double cosd = cos(dec0 * (PI_D / 180.));
      cd[0] = (rax - ra0) / step * cosd; cd[1] = (ray - ra0) / step * cosd;
      cd[2] = (decx - dec0) / step;      cd[3] = (decy - dec0) / step;
    }
    // celeste_galaxy_conditionals.py:90-125
    double phi = (90. - s.phi) * PI_D / 180.;
    double re_deg = fmax(1. / 30, s.sigma) / 3600.;
    double cp = cos(phi), sp = sin(phi);
    double G[4] = {re_deg * cp, re_deg * (sp * s.rho), re_deg * (-sp), re_deg * (cp * s.rho)};
    double detG = G[0] * G[3] - G[1] * G[2];
    double Gi[4] = {G[3] / detG, -G[1] / detG, -G[2] / detG, G[0] / detG};
    double T[4] = {Gi[0] * cd[0] + Gi[1] * cd[2], Gi[0] * cd[1] + Gi[1] * cd[3],
                   Gi[2] * cd[0] + Gi[3] * cd[2], Gi[2] * cd[1] + Gi[3] * cd[3]};
    double detT = T[0] * T[3] - T[1] * T[2];
    double Ti[4] = {T[3] / detT, -T[1] / detT, -T[2] / detT, T[0] / detT};
    Wm[0] = Ti[0] * Ti[0] + Ti[1] * Ti[1];
    Wm[1] = Ti[2] * Ti[2] + Ti[3] * Ti[3];
    Wm[2] = Ti[0] * Ti[2] + Ti[1] * Ti[3];
    if (flags & CG_ITEM_PROF_EXP) { n_exp = 6; n_dev = 0; }
    else if (flags & CG_ITEM_PROF_DEV) { n_exp = 0; n_dev = 8; }
    else { n_exp = 6; n_dev = 8; th_exp = s.theta; th_dev = 1. - s.theta; }
    K = (n_exp + n_dev) * n_psf;
    trunc_box = (flags & CG_ITEM_TRUNC_BOX) != 0;
  }

  const int ox = (int)rint(px), oy = (int)rint(py);
  const double Rsq_gal = 23.025850929940457;  // chi2.ppf(1 - 1e-5, 2)

  double mybound = -1e300;
  for (int c = lane; c < K; c += 32) {
    double w, s00, s11, s01, mux, muy;
    int k;
    if (star) {
      k = c;
      w = im.psf_w[k];
      s00 = im.psf_cov[k][0]; s11 = im.psf_cov[k][1]; s01 = im.psf_cov[k][2];
    } else {
      // profile-major ordering of the live route (mog.py:75-81): c = j*n_psf + k
      int j = c / n_psf;
      k = c - j * n_psf;
      double amp, var;
      if (j < n_exp) { amp = th_exp * c_prof.exp_amp[j]; var = c_prof.exp_var[j]; }
      else { amp = th_dev * c_prof.dev_amp[j - n_exp]; var = c_prof.dev_var[j - n_exp]; }
      w = amp * im.psf_w[k];
      s00 = var * Wm[0] + im.psf_cov[k][0];
      s11 = var * Wm[1] + im.psf_cov[k][1];
      s01 = var * Wm[2] + im.psf_cov[k][2];
      // The footprint radius is the max over components of (major axis + |psf mean|)
      // (bounding_box.py:27-31).  For a fixed PSF component the major axis grows with the profile
      // variance (var W is positive definite), so the maximum sits on the profile component of largest
      // variance -- the LAST one of the included profiles (dev 8.40 > exp 1.50): only those n_psf
      // components pay for the 4 square roots and 7 divisions of major_axis; the value is the one
      // the full loop would return.
      if (j == n_exp + n_dev - 1) {
        double dist = sqrt(im.psf_mu[k][0] * im.psf_mu[k][0] + im.psf_mu[k][1] * im.psf_mu[k][1]);
        mybound = fmax(mybound, major_axis(s00, s11, s01, Rsq_gal) + dist);
      }
    }
    mux = px + im.psf_mu[k][0];
    muy = py + im.psf_mu[k][1];
    // gmm_like_fast.pyx:165-176 folded into log2 units
    double det = s00 * s11 - s01 * s01;
    double i00 = s11 / det, i11 = s00 / det, i01 = -1 * s01 / det;
    double k2 = (-LOG2PI_D - .5 * log(det) + log(w)) * LOG2E_D;
    double A = -.5 * LOG2E_D * i00, B = -LOG2E_D * i01, C = -.5 * LOG2E_D * i11;
    double mx = mux - ox, my = muy - oy;
    real* t = tab + c * TAB_W;
    t[0] = (real)A;
    t[1] = (real)B;
    t[2] = (real)C;
    t[3] = (real)(-2. * A * mx - B * my);
    t[4] = (real)(-B * mx - 2. * C * my);
    t[5] = (real)(A * mx * mx + B * mx * my + C * my * my + k2);
    t[6] = (real)mx;
    t[7] = (real)my;
    t[8] = (real)k2;
    // the sign of t[9] carries recur_flush_unsafe (the value itself is positive)
    t[9] = (real)(recur_flush_unsafe(A, k2) ? -(2. * s00 / LOG2E_D) : (2. * s00 / LOG2E_D));
    t[10] = (real)(2. * s11 / LOG2E_D);
    {
      const double dd = s00 - s11;
      const double lmin = .5 * ((s00 + s11) - sqrt(dd * dd + 4. * s01 * s01));
      t[11] = (real)(.5 * LOG2E_D / fmax(lmin, 1e-300));
    }
  }
  if (!star) {
    bound = warp_max(mybound);
    if (K == 0) bound = 0;
  }
  g.px = px; g.py = py; g.bound = bound; g.K = K; g.ox = ox; g.oy = oy; g.star = star; g.trunc_box = trunc_box;
  g.flux = 0.0;
  return g;
}

// rate multiplier of an item (celeste.py:38-45,80-81,91-92; planck.py:155-158), by one warp
__device__ __forceinline__ double item_flux(const DevImage& im, const CgSource& s, int flags, int lane, bool star,
                                            const FilterTables* filters, const double* flux_table,
                                            long long flux_rows, ItemGeom& g) {
  // flux (celeste.py:38-45,80-81,91-92; planck.py:155-158)
  double flux;
  if (flags & CG_ITEM_FLUX_UNIT) {
    flux = 1.0;
  } else if (s.flags & CG_SRC_EIGEN) {
    // eigen-spectrum route: the band flux was projected on the device (eigen_flux_kernel) and is read
    // from HBM here -- redshift_utils.py:304-323 feeding celeste.py:80-81,91-92
    if (flux_table && s.flux_row >= 0 && s.flux_row < flux_rows) {
      flux = (flux_table[s.flux_row * CG_NUM_BANDS + im.band] / im.calib) * im.kappa;
    } else {
      flux = (double)NAN;
      g.err = 4;
    }
  } else if (star && (s.flags & CG_SRC_PLANCK) && s.t != 0.0 && !filters->loaded) {
    flux = (double)NAN;                 // cg_set_filter_curves has not been called: fail loudly, not with a dark star
    g.err = 5;
  } else if (star && (s.flags & CG_SRC_PLANCK) && s.t != 0.0) {
    double ppj = planck_photons_per_joule_warp(filters, s.t, im.band, lane);
    ppj = __shfl_sync(0xffffffffu, ppj, 0);
    const double c_ = 299792458.0;
    const double m_per_ly = c_ * 31556952.;
    const double lens_area = .75 * PI_D * (1.25 * 1.25);
    double lens_watts = lens_area * s.b * (3.846 * 1e26) / (m_per_ly * m_per_ly);
    flux = ppj * lens_watts * 54.;
  } else {
    flux = (s.flux[im.band] / im.calib) * im.kappa;
  }

  return flux;
}

// ---------------------------------------------------------------------------
// prepare: one warp per item
// ---------------------------------------------------------------------------
#ifndef PREP_MINB
#define PREP_MINB 32        // 64 registers (48 bytes of spills)
#endif
template <typename real>
__global__ void __launch_bounds__(32 * PREP_WARPS, PREP_MINB) prepare_kernel(PrepArgs a) {
  const int lane = threadIdx.x & 31;
  const long long i = (long long)blockIdx.x * PREP_WARPS + (threadIdx.x >> 5);
  if (i >= a.n_items) return;

  CgItem it;
  if (a.items) {
    it = a.items[i];
  } else if (a.implicit_mode == 0) {
    it.src = (int)(i / a.n_img);
    it.img = (int)(i % a.n_img);
    it.flags = a.default_flags;
    it.aux = -1;
    it.box[0] = it.box[1] = it.box[2] = it.box[3] = 0;
  } else if (a.implicit_mode == 3) {
    int f = 0;
    while (f + 1 < a.fb.n_fields && i >= a.fb.item_off[f + 1]) ++f;
    const int local = (int)i - a.fb.item_off[f];
    const int ni = a.fb.img_off[f + 1] - a.fb.img_off[f];
    it.src = a.fb.src_off[f] + local / ni;
    it.img = a.fb.img_off[f] + local % ni;
    it.flags = a.default_flags;
    it.aux = -1;
    it.box[0] = it.box[1] = it.box[2] = it.box[3] = 0;
  } else {
    int g = (int)(i / a.n_img);
    it.img = (int)(i % a.n_img);
    if (a.implicit_mode == 2) {
      int cnt = *a.ind_count;
      if (g > cnt) return;
      it.src = (g == 0) ? 0 : a.ind_list[g - 1];
    } else {
      it.src = g;
    }
    it.flags = a.default_flags | CG_ITEM_FIXED_BOX;
    it.aux = a.aux_per_img ? it.img : -1;
    const CgBox b = a.img_boxes[it.img];
    it.box[0] = b.x0; it.box[1] = b.x1; it.box[2] = b.y0; it.box[3] = b.y1;
  }

  const DevImage& im = *a.imgs.p[it.img];
  const CgSource& s = a.srcs[it.src];
  const int H = im.H, W = im.W;

  real* tab = reinterpret_cast<real*>(a.tab) + i * TAB_STRIDE;
  ItemGeom geo = item_table<real>(im, s, it.flags, lane, tab);
  const double px = geo.px, py = geo.py;
  const int K = geo.K, ox = geo.ox, oy = geo.oy;
  const bool star = geo.star, trunc_box = geo.trunc_box;
  double bound = geo.bound;
  if (a.bound_cap > 0) bound = fmin(bound, a.bound_cap);

  int x0, x1, y0, y1;
  if (it.flags & CG_ITEM_FIXED_BOX) {
    x0 = it.box[0]; x1 = it.box[1]; y0 = it.box[2]; y1 = it.box[3];
    x0 = max(x0, 0); y0 = max(y0, 0); x1 = min(x1, W); y1 = min(y1, H);
  } else if (trunc_box) {
    // celeste.py:139-140 / celeste_galaxy_conditionals.py:161-162
    x0 = max(0, (int)(px - bound)); x1 = min((int)(px + bound + 1), W);
    y0 = max(0, (int)(py - bound)); y1 = min((int)(py + bound + 1), H);
  } else {
    // celeste_galaxy_conditionals.py:208-211
    x0 = (int)fmax(0., floor(px - bound)); x1 = (int)fmin((double)W, ceil(px + bound));
    y0 = (int)fmax(0., floor(py - bound)); y1 = (int)fmin((double)H, ceil(py + bound));
  }
  if (x1 <= x0 || y1 <= y0 || (K == 0 && !(it.flags & CG_ITEM_FIXED_BOX))) { x1 = x0; y1 = y0; }
  long long npx = (long long)(x1 - x0) * (y1 - y0);
  int err = 0;
  // only the entry points that write box-shaped buffers (patches, auxiliary patches) are bounded by
  // max_patch_px; the fused field path (no work list) clips the footprint to the image and is done
  if ((a.worklist != nullptr || a.cap_boxes) && npx > a.max_patch_px) { err = 1; x1 = x0; y1 = y0; npx = 0; }

  const double flux = item_flux(im, s, it.flags, lane, star, a.filters, a.flux_table, a.flux_rows, geo);
  if (geo.err) err = geo.err;

  const int wbox = x1 - x0, hbox = y1 - y0;
  const int cpr = (wbox + PIX_PER_THREAD - 1) / PIX_PER_THREAD;
  // row pairs: full 16-row bands hold 8, the last band ceil(rows / 2)
  const int hp = (hbox / (2 * TILE_R)) * TILE_R + ((hbox % (2 * TILE_R)) + 1) / 2;
  const int n_strips = (int)(((long long)hp * cpr + 31) / 32);               // warp tiles
  const int n_units = (n_strips + a.unit_wt - 1) / a.unit_wt;
  const bool heavy = K > 4;
  int start = 0, wpos = 0;
  const bool listed = a.worklist != nullptr;      // the fused field path renders by image tile, not by work list
  if (lane == 0) {
    if (n_strips && listed) {
      start = atomicAdd(&a.cnt->total_strips, n_strips);
      wpos = atomicAdd(heavy ? &a.cnt->heavy_strips : &a.cnt->light_strips, n_units);
    }
    if (npx) atomicAdd(&a.cnt->evals, (unsigned long long)npx * (unsigned long long)K);
    if (err) atomicExch(&a.cnt->err_flag, err);
    if (listed && start + n_strips > a.worklist_cap) { atomicExch(&a.cnt->err_flag, 3); }
  }
  start = __shfl_sync(0xffffffffu, start, 0);
  wpos = __shfl_sync(0xffffffffu, wpos, 0);
  // heavy strips fill the list from the front, light ones from the back: the
  // render kernel drains the front first (longest-processing-time-first)
  if (listed && start + n_strips <= a.worklist_cap)
    for (int j = lane; j < n_units; j += 32) {
      const int slot = heavy ? (wpos + j) : (a.worklist_cap - 1 - (wpos + j));
      CG_CHECK(slot >= 0 && slot < a.worklist_cap);
      a.worklist[slot] = make_int2((int)i, j * a.unit_wt);   // (item, first warp tile of the unit)
    }
  // bin the footprint into the image tiles it overlaps (full-field gather)
  if (a.tile_count && npx > 0 && K > 0) {
    // bin_w x bin_h bins with lists of bin_cap entries (16 x 16 / 32: field_tile_kernel)
    const int bw = a.bin_w, bh = a.bin_h, cap = a.bin_cap;
    const int tiles_x = (W + bw - 1) / bw;
    const int txa = x0 / bw, txb = (x1 - 1) / bw;
    const int tya = y0 / bh, tyb = (y1 - 1) / bh;
    const int nx = txb - txa + 1, nt = nx * (tyb - tya + 1);
    for (int j = lane; j < nt; j += 32) {
      const int t = (tya + j / nx) * tiles_x + txa + (j % nx);
      const long long slot = (long long)it.img * a.max_tiles + t;
      const int pos = atomicAdd(&a.tile_count[slot], 1);
      // entry = item * 64 + K (K <= CG_MAX_COMP < 64): the tile kernels weigh a bin
      // without touching the item headers
      CG_CHECK(t >= 0 && t < a.max_tiles && K <= CG_MAX_COMP);
      if (pos < cap) a.tile_list[slot * cap + pos] = (int)i * 64 + K;
    }
  }
  if (lane == 0) {
    ItemHdr h;
    h.x0 = x0; h.x1 = x1; h.y0 = y0; h.y1 = y1;
    h.ox = ox; h.oy = oy; h.K = K; h.img = it.img;
    h.strip_start = start; h.n_strips = n_strips;
    h.aux = it.aux; h.flags = it.flags;
    h.flux = flux;
    h.norm_term = flux * im.psf_wsum;
    a.hdr[i] = h;
    if (a.out_boxes) {
      CgBox b; b.x0 = x0; b.x1 = x1; b.y0 = y0; b.y1 = y1;
      a.out_boxes[i] = b;
    }
  }
}

// ---------------------------------------------------------------------------
// render: persistent CTAs over the strip work list
// ---------------------------------------------------------------------------
#define MODE_PATCH 0
#define MODE_POISSON 1
#define MODE_COND 2

struct RenderArgs {
  ImgPtrs imgs;
  const ItemHdr* hdr;
  const void* tab;
  const int2* worklist;
  int worklist_cap;
  Counters* cnt;
  float* patches;             // MODE_PATCH
  long long max_patch_px;
  const float* aux;           // background / photon patches
  const long long* aux_offsets;
  long long aux_off_v[CG_MAX_IMAGES];   // used instead of aux_offsets when that is NULL (cg_score_proposals: by value)
  long long aux_stride;       // > 0 (with aux_offsets NULL): auxiliary patch k lies at aux + k * aux_stride
  double* partials;           // one per strip
  int cull;                   // 0 disables component culling
  int unit_wt;                // warp tiles per work unit
  int cond_sum;               // MODE_COND: subtract the rendered rates (CG_LL_CONDITIONAL_SUM)
  const LogTable* logtab;
};

// q / d for 0 <= q < 2^22, d >= 1 through the fp32 reciprocal (+ one correction step)
__device__ __forceinline__ int fast_div(int q, int d, float rd) {
  int t = (int)((float)q * rd);
  int r = q - t * d;
  if (r < 0) { --t; r += d; }
  if (r >= d) { ++t; }
  return t;
}

// Lane unit q of a (w x h) footprint in tile-major order -> rows (row0, row1)
// and first column.  hp = number of row pairs (prepare_kernel), bands of
// TILE_R row pairs = 2*TILE_R rows.  row1 >= hbox means the unit has one row.
__device__ __forceinline__ void unit_decode(int q, int cpr, int hp, float r_band, int* row0,
                                            int* row1, int* cx) {
  const int band_units = cpr * TILE_R;
  const int tr = fast_div(q, band_units, r_band);
  const int rem = q - tr * band_units;
  const int R = min(TILE_R, hp - tr * TILE_R);          // row pairs in this band
  const int ft = TILE_CW * R;
  const int tc = fast_div(rem, ft, __frcp_rn((float)ft));
  const int rem2 = rem - tc * ft;
  const int tw = min(TILE_CW, cpr - tc * TILE_CW);
  const int r = (tw == TILE_CW) ? (rem2 / TILE_CW) : fast_div(rem2, tw, __frcp_rn((float)tw));
  const int c = rem2 - r * tw;
  *row0 = tr * (2 * TILE_R) + r;
  *row1 = *row0 + R;
  *cx = (tc * TILE_CW + c) * PIX_PER_THREAD;
}

// Per-warp component culling over the warp's pixel rectangle [xa,xb] x [ya,yb]
// (footprint-local coordinates).  Each lane owns components lane and lane+32.
//   lower bound of the mixture  lb = max_k ( qs_k - sl_k * dfar_k^2 ), dfar = distance from the
//                               mean to the farthest corner (Mahalanobis^2 <= d^2 / lambda_min);
//   component k is dropped when even the bounding box of its level set
//   { q_k >= lb - CULL_T } misses the rectangle, i.e. its largest value over
//   the rectangle is below 2^-CULL_T of the mixture everywhere in it.
// Survivors are compacted (component order preserved) into the warp's table
// of 8-float entries; returns their number.
// The row pair (h, k8) = (2^(2A), 2^(8B)) of each surviving component feeds the
// row recurrence of ft_eval2r/ft_eval1r; *recur_ok says whether that recurrence is
// safe for every survivor on this rectangle (see FT_RECUR below).
// LDG: the table is read from global memory through the read-only path (field kernel: L2)
// or from shared memory (render_kernel).  xext: pixels a lane may start left of the rectangle.
template <bool LDG>
__device__ __forceinline__ int cull_components(const float* gtab, int K, int lane, float xa, float xb,
                                               float ya, float yb, float xext, float* wtab, bool* recur_ok) {
  float lb = -INFINITY;
  float4 e0[2], e1[2];
  float qs[2], kx[2], ky[2], dx2[2], dy2[2], dmx[2];
#pragma unroll
  for (int r = 0; r < 2; ++r) {
    const int k = lane + 32 * r;
    qs[r] = -INFINITY; kx[r] = ky[r] = 0.f; dx2[r] = dy2[r] = 0.f; dmx[r] = 0.f;
    if (k < K) {
      const float4* t4 = reinterpret_cast<const float4*>(gtab + k * TAB_W);
      e0[r] = LDG ? __ldg(t4) : t4[0];
      e1[r] = LDG ? __ldg(t4 + 1) : t4[1];
      const float4 e2 = LDG ? __ldg(t4 + 2) : t4[2];
      const float mx = e1[r].z, my = e1[r].w;
      qs[r] = e2.x; kx[r] = fabsf(e2.y); ky[r] = e2.z;
      const float fx = fmaxf(fabsf(xa - mx), fabsf(xb - mx)), fy = fmaxf(fabsf(ya - my), fabsf(yb - my));
      const float l = qs[r] - e2.w * (fx * fx + fy * fy);
      if (l == l) lb = fmaxf(lb, l);
      const float dx = fmaxf(0.f, fmaxf(xa - mx, mx - xb)), dy = fmaxf(0.f, fmaxf(ya - my, my - yb));
      dx2[r] = dx * dx; dy2[r] = dy * dy;
      // bound of |q(x+1,y) - q(x,y)| = |2A(x-mx) + B(y-my) + A| over the rectangle widened
      // by the xext pixels a lane's group of 4 may start to the left of it; 8|B| bounds k8
      dmx[r] = fmaxf(fmaf(fabsf(e0[r].x), 2.f * (fx + xext) + 1.f, fabsf(e0[r].y) * fy), 8.f * fabsf(e0[r].y));
      if (e2.y < 0.f) dmx[r] = INFINITY;                         // recur_flush_unsafe (prepare_kernel)
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) lb = fmaxf(lb, __shfl_xor_sync(0xffffffffu, lb, o));
  const float thr = lb - (CULL_T + 0.5f);
  bool keep[2];
#pragma unroll
  for (int r = 0; r < 2; ++r) {
    const float rho = (qs[r] - thr) * 1.0001f;
    keep[r] = (lane + 32 * r < K) && !(rho < 0.f) && !(dx2[r] > rho * kx[r]) && !(dy2[r] > rho * ky[r]);
  }
  const unsigned m0 = __ballot_sync(0xffffffffu, keep[0]);
  const unsigned m1 = __ballot_sync(0xffffffffu, keep[1]);
  const unsigned lt = (1u << lane) - 1u;
  CG_CHECK(K <= CG_MAX_COMP && __popc(m0) + __popc(m1) <= CG_MAX_COMP);
  if (keep[0]) {
    float* d = wtab + __popc(m0 & lt) * 8;
    *reinterpret_cast<float4*>(d) = e0[0];
    *reinterpret_cast<float4*>(d + 4) = make_float4(e1[0].x, e1[0].y, ex2_approx(2.f * e0[0].x), ex2_approx(8.f * e0[0].y));
  }
  if (keep[1]) {
    float* d = wtab + (__popc(m0) + __popc(m1 & lt)) * 8;
    *reinterpret_cast<float4*>(d) = e0[1];
    *reinterpret_cast<float4*>(d + 4) = make_float4(e1[1].x, e1[1].y, ex2_approx(2.f * e0[1].x), ex2_approx(8.f * e0[1].y));
  }
  *recur_ok = !__any_sync(0xffffffffu, (keep[0] && !(dmx[0] <= FT_RECUR_DMAX)) || (keep[1] && !(dmx[1] <= FT_RECUR_DMAX)));
  __syncwarp();
  return __popc(m0) + __popc(m1);
}

// One WARP TILE (32 lane units = 256 pixels) of an item: cull the item's components against the
// tile, evaluate, and either write the patch (MODE_PATCH) or return the tile's likelihood terms summed
// over the warp (lane 0 holds the sum).  Shared by render_kernel and the on-device samplers.
// where the auxiliary patch of an item starts
__device__ __forceinline__ long long aux_offset_of(const RenderArgs& a, int idx) {
  return a.aux_offsets ? a.aux_offsets[idx] : (a.aux_stride ? (long long)idx * a.aux_stride : a.aux_off_v[idx]);
}
// the few fields render_warp_tile reads, for kernels that do not carry a whole RenderArgs
struct TileArgs {
  int cull, cond_sum;
  float* patches;
  long long max_patch_px;
  const float* aux;
  long long aux_stride;
};
__device__ __forceinline__ long long aux_offset_of(const TileArgs& a, int idx) { return (long long)idx * a.aux_stride; }

#ifndef RW_ROWPAIR
#define RW_ROWPAIR 1        // warp tiles whose lanes' rows are 8 apart take the field kernel's packed loop (ft_eval2r)
#endif
__device__ __forceinline__ void ft_eval2r(const float* tb, int n, float y0l, float y1l,
                                          float x0, float (&acc0)[4], float (&acc1)[4]);

template <typename real, int MODE, typename Args>
__device__ __forceinline__ double render_warp_tile(const Args& a, const ItemHdr& h, long long item,
                                                   const real* s_tab, float* s_wtab, int wt, int lane, int wbox,
                                                   int hbox, int cpr, int hp, int n_lu, float r_band,
                                                   const DevImage* imp, real fl, const LogTable* logt,
                                                   unsigned long long& issued) {
  const int q = wt * 32 + lane;
  const bool active = q < n_lu;
  int row0 = 0, row1 = 0, cx = 0;
  if (active) unit_decode(q, cpr, hp, r_band, &row0, &row1, &cx);
  const bool two = active && row1 < hbox;
  const int npx = active ? min(PIX_PER_THREAD, wbox - cx) : 0;

  int n = h.K;
  const real* tb = s_tab;
  int stride = TAB_W;
  bool recur = false;
  if (sizeof(real) == 4 && (a.cull & 1) && n > 1) {
    const int rmin = __reduce_min_sync(0xffffffffu, active ? row0 : 0x7fffffff);
    const int rmax = __reduce_max_sync(0xffffffffu, active ? (two ? row1 : row0) : -1);
    const int cmin = __reduce_min_sync(0xffffffffu, active ? cx : 0x7fffffff);
    const int cmax = __reduce_max_sync(0xffffffffu, active ? cx + npx - 1 : -1);
    __syncwarp();             // the warp is done reading the culled table of the previous tile
    n = cull_components<false>(reinterpret_cast<const float*>(s_tab), n, lane,
                               (float)(h.x0 + cmin - h.ox), (float)(h.x0 + cmax - h.ox),
                               (float)(h.y0 + rmin - h.oy), (float)(h.y0 + rmax - h.oy), 0.f, s_wtab, &recur);
    recur = recur && FT_RECUR && !(a.cull & 2);
    tb = reinterpret_cast<const real*>(s_wtab);
    stride = 8;
  }
  if (lane == 0) issued += (unsigned long long)n * WT_PIX;

  real acc0[PIX_PER_THREAD], acc1[PIX_PER_THREAD];
#pragma unroll
  for (int p = 0; p < PIX_PER_THREAD; ++p) { acc0[p] = (real)0; acc1[p] = (real)0; }
  {
    const real y0l = (real)(h.y0 + row0 - h.oy), y1l = (real)(h.y0 + row1 - h.oy);
    const real xl0 = (real)(h.x0 + cx - h.ox);
    real xl[PIX_PER_THREAD];
#pragma unroll
    for (int p = 0; p < PIX_PER_THREAD; ++p) xl[p] = xl0 + (real)p;
#if RW_ROWPAIR
    if constexpr (sizeof(real) == 4) {
      // a full band of a footprint: every lane's rows are TILE_R = 8 apart, the geometry of the field
      // kernel's bins -- its loop (3 exp2 and 22 issue slots per component and lane instead of 4 and 36;
      // the second row's ratio is the first's times k8 = 2^(8B), which cull_components already tabulated
      // and range-checked)
      if (recur && __all_sync(0xffffffffu, !active || row1 - row0 == 8)) {
        ft_eval2r(tb, n, y0l, y1l, xl0, acc0, acc1);
        n = 0;
        recur = false;
      }
    }
#endif
    if (sizeof(real) == 4 && recur) {
      // row recurrence: 2 exp2 per row of 4 pixels (the rows of a lane are not a fixed
      // distance apart here, so each row takes its own neighbour ratio)
      const real tx = (real)2 * xl0 + (real)1;
#if FT_PACKED
      f32x2 pacc[PIX_PER_THREAD];         // (row0, row1) pairs: FMUL2 / FADD2, see ft_eval2r
#pragma unroll
      for (int p = 0; p < PIX_PER_THREAD; ++p) pacc[p] = pk2(0.f, 0.f);
#endif
#pragma unroll 2
      for (int k = 0; k < n; ++k, tb += 8) {
        const real A = tb[0], B = tb[1], C = tb[2], b10 = tb[3], c01 = tb[4], c00 = tb[5], hh = tb[6];
        const real b1a = rfma(B, y0l, b10), b1b = rfma(B, y1l, b10);
        const real c1a = rfma(rfma(C, y0l, c01), y0l, c00);
        const real c1b = rfma(rfma(C, y1l, c01), y1l, c00);
        const real ea = rexp2(rfma(rfma(A, xl0, b1a), xl0, c1a)), ga = rexp2(rfma(A, tx, b1a));
        const real eb = rexp2(rfma(rfma(A, xl0, b1b), xl0, c1b)), gb = rexp2(rfma(A, tx, b1b));
#if FT_PACKED
        recur_row4x2(pk2((float)ea, (float)eb), pk2((float)ga, (float)gb), pk2((float)hh, (float)hh), pacc);
#else
        recur_row4(ea, ga, hh, acc0);
        recur_row4(eb, gb, hh, acc1);
#endif
      }
#if FT_PACKED
#pragma unroll
      for (int p = 0; p < PIX_PER_THREAD; ++p) {
        float lo, hi;
        unpk2(pacc[p], lo, hi);
        acc0[p] = (real)lo; acc1[p] = (real)hi;
      }
#endif
      n = 0;
    }
#pragma unroll 2
    for (int k = 0; k < n; ++k, tb += stride) {
      const real A = tb[0], B = tb[1], C = tb[2], b10 = tb[3], c01 = tb[4], c00 = tb[5];
      const real b1a = rfma(B, y0l, b10), b1b = rfma(B, y1l, b10);
      const real c1a = rfma(rfma(C, y0l, c01), y0l, c00);
      const real c1b = rfma(rfma(C, y1l, c01), y1l, c00);
#pragma unroll
      for (int p = 0; p < PIX_PER_THREAD; ++p) {
        acc0[p] += rexp2(rfma(rfma(A, xl[p], b1a), xl[p], c1a));
        acc1[p] += rexp2(rfma(rfma(A, xl[p], b1b), xl[p], c1b));
      }
    }
  }

  if (MODE == MODE_PATCH) {
    if (active) {
      CG_CHECK((long long)(two ? row1 : row0) * wbox + cx + npx <= a.max_patch_px && row0 >= 0 && cx >= 0);
      float* out = a.patches + (long long)item * a.max_patch_px + (long long)row0 * wbox + cx;
#pragma unroll
      for (int p = 0; p < PIX_PER_THREAD; ++p)
        if (p < npx) out[p] = (float)(acc0[p] * fl);
      if (two) {
        out += (long long)(row1 - row0) * wbox;
#pragma unroll
        for (int p = 0; p < PIX_PER_THREAD; ++p)
          if (p < npx) out[p] = (float)(acc1[p] * fl);
      }
    }
  } else {
    double s = 0.0;
    if (active) {
      const DevImage& im = *imp;
      const real eps = (real)im.epsilon;
      const bool use_mask = im.mask && !(h.flags & CG_ITEM_NO_MASK);
#pragma unroll
      for (int rr = 0; rr < 2; ++rr) {
        if (rr == 1 && !two) break;
        const int row = rr ? row1 : row0;
        const long long poff = (long long)row * wbox + cx;
        const float* auxp = (h.aux >= 0) ? a.aux + aux_offset_of(a, h.aux) + poff : nullptr;
        const long long ioff = (long long)(h.y0 + row) * im.W + (h.x0 + cx);
#pragma unroll
        for (int p = 0; p < PIX_PER_THREAD; ++p) {
          if (p < npx) {
            const real v = fl * (rr ? acc1[p] : acc0[p]);
            if (MODE == MODE_POISSON) {
              const real lam = (auxp ? (real)auxp[p] : eps) + v;
              const bool good = (lam > (real)0) && (!use_mask || im.mask[ioff + p] != 0.f);
              if (good) s += (double)im.nelec[ioff + p] * log_accurate(lam, logt) - (double)lam;
            } else {      // MODE_COND, sources.py:171-174
              if (v >= (real)1.17549435e-38f) s += log_accurate(v, logt) * (double)auxp[p];
              if (a.cond_sum) s -= (double)v;     // celeste_galaxy_conditionals.py:41
            }
          }
        }
      }
    }
    s = warp_sum(s);
    return s;
  }
  return 0.0;
}

template <typename real, int MODE>
__global__ void __launch_bounds__(STRIP_THREADS) render_kernel(RenderArgs a) {
  // per-warp staging: the item's component table and the culled, compacted table
  __shared__ __align__(16) real s_tab_all[(STRIP_THREADS / 32) * TAB_STRIDE];
  __shared__ __align__(16) float s_wtab_all[(sizeof(real) == 4) ? (STRIP_THREADS / 32) * (CG_MAX_COMP + 1) * 8 : 4];   // + ft_eval2r's prefetch row
  const LogTable* const logt = a.logtab;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  real* const s_tab = s_tab_all + warp * TAB_STRIDE;
  float* const s_wtab = s_wtab_all + ((sizeof(real) == 4) ? warp * (CG_MAX_COMP + 1) * 8 : 0);
  const int n_heavy = *((volatile int*)&a.cnt->heavy_strips);
  const int total = n_heavy + *((volatile int*)&a.cnt->light_strips);
  unsigned long long issued = 0;

  int wk = 0;
  if (lane == 0) wk = atomicAdd(&a.cnt->work_counter, 1);
  wk = __shfl_sync(0xffffffffu, wk, 0);
  while (wk < total) {
    const int slot = (wk < n_heavy) ? wk : (a.worklist_cap - 1 - (wk - n_heavy));
    CG_CHECK(slot >= 0 && slot < a.worklist_cap);
    const int2 w = a.worklist[slot];
    const int item = w.x, wt0 = w.y;
    int nxt = 0;
    if (lane == 0) nxt = atomicAdd(&a.cnt->work_counter, 1);     // in flight during the unit
    const ItemHdr h = a.hdr[item];
    __syncwarp();                 // the warp is done with the tables of its previous unit
    {
      const float4* g4 = reinterpret_cast<const float4*>(reinterpret_cast<const real*>(a.tab) +
                                                          (long long)item * TAB_STRIDE);
      float4* s4 = reinterpret_cast<float4*>(s_tab);
      const int n4 = h.K * TAB_W * (int)sizeof(real) / 16;
      for (int t = lane; t < n4; t += 32) s4[t] = g4[t];
    }
    __syncwarp();

    const int wbox = h.x1 - h.x0, hbox = h.y1 - h.y0;
    const int cpr = (wbox + PIX_PER_THREAD - 1) / PIX_PER_THREAD;
    const int hp = (hbox / (2 * TILE_R)) * TILE_R + ((hbox % (2 * TILE_R)) + 1) / 2;
    const int n_lu = cpr * hp;
    const float r_band = __frcp_rn((float)(cpr * TILE_R));
    const int wt_end = min(wt0 + a.unit_wt, h.n_strips);
    const DevImage* imp = (MODE != MODE_PATCH) ? a.imgs.p[h.img] : nullptr;
    const real fl = (real)h.flux;

#pragma unroll 1
    for (int wt = wt0; wt < wt_end; ++wt) {
      const double s = render_warp_tile<real, MODE>(a, h, item, s_tab, s_wtab, wt, lane, wbox, hbox, cpr, hp, n_lu,
                                                    r_band, imp, fl, logt, issued);
      CG_CHECK(h.strip_start + wt < a.worklist_cap && wt < h.n_strips);
      if (MODE != MODE_PATCH && lane == 0) a.partials[h.strip_start + wt] = s;
    }
    wk = __shfl_sync(0xffffffffu, nxt, 0);
  }
  if (lane == 0 && issued) atomicAdd(&a.cnt->issued, issued);
}

// per-item sum of strip partials (fixed order -> deterministic)
__global__ void items_finalize_kernel(const ItemHdr* hdr, const double* partials, long long n_items,
                                      int mode, double* item_ll, const int* count_limit, int n_img) {
  long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_items) return;
  if (count_limit && (i / n_img) > *count_limit) return;
  const ItemHdr h = hdr[i];
  double s = 0;
  for (int j = 0; j < h.n_strips; ++j) s += partials[h.strip_start + j];
  if (mode == CG_LL_CONDITIONAL) s -= h.norm_term;
  item_ll[i] = s;
}

// ---------------------------------------------------------------------------
// full-field log-likelihood : lambda = eps + sum_s rate_s ; ll = sum x log lam - lam
// ---------------------------------------------------------------------------
#define FT_BUCKETS 8
struct FieldArgs {
  ImgPtrs imgs;
  int n_img;
  const ItemHdr* hdr;
  const float* tab;        // component tables, n_items * TAB_STRIDE
  const int* tile_count;   // [n_img][max_tiles] sources listed per 16 x 16 bin
  int* tile_list;          // [n_img][max_tiles][BIN_CAP] entries item * 64 + K (put in source order by bin_order_kernel)
  int2* order;             // [FT_BUCKETS][tiles_total] (bin id, list length) by weight class
  int* sched;              // [FT_BUCKETS] class sizes, [FT_BUCKETS] work counter (zero on entry)
  double* block_partials;  // [n_img][max_tiles] per-bin sums
  double* ll;              // [n_img]
  int max_tiles;           // 16 x 16 bins per image (largest image of the call)
  int tiles_total;         // n_img * max_tiles
  // list overflow fallback: the items of image i are item_base[i] + s*item_stride[i], s < n_src[i]
  int item_base[CG_MAX_IMAGES], item_stride[CG_MAX_IMAGES], n_src[CG_MAX_IMAGES];
  float* const* model;     // device array [n_img] of optional model outputs, or NULL
  const LogTable* logtab;  // precomputed at cg_init
  Counters* cnt;
  int cull;
};

// ---------------------------------------------------------------------------
// fused field path: render + gather + Poisson terms by IMAGE TILE
// ---------------------------------------------------------------------------
// One WARP owns a 16 x 16 pixel bin of an image (lane = 4 pixels of rows r and
// r + 8, the lane layout of render_kernel) and is fully autonomous: no block
// barrier anywhere.  It walks the bin's source list in source order (bitwise
// reproducible), culls each source's components against the part of the
// footprint inside its 256 pixels (tables read straight from L2, survivors
// compacted into the warp's shared-memory table), accumulates
// lambda = eps + sum_s rate_s in registers and finishes with the Poisson terms:
// neither lambda nor any patch ever leaves the SM.  Pixel traffic is the 4 B of
// nelec per pixel, read once.
//
// Bins differ in cost by three orders of magnitude (empty sky ... several
// overlapping galaxies), so bin_order_kernel first sorts them into FT_BUCKETS
// weight classes (weight = components in the bin's list) and the persistent
// warps drain the classes heaviest first through one atomic counter, two work
// indices ahead so that neither the atomic nor the descriptor load is waited for.
#define BIN_W 16
#define BIN_H 16
#define BIN_CAP 32
#define BINS_PER_TILE ((TILE_W / BIN_W) * (TILE_H / BIN_H))
#ifndef FT_THREADS
#define FT_THREADS 256
#endif
#ifndef FT_STAGE_D
#define FT_STAGE_D 0        // pixel counts staged in shared memory by cp.async (0: held in registers, 3 CTAs/SM)
#endif
#ifndef FT_LAUNCH_THREADS
#define FT_LAUNCH_THREADS FT_THREADS   // fewer warps per CTA leave registers for the next set's prepare_kernel to co-reside
#endif
#ifndef FT_MINB
#define FT_MINB (FT_STAGE_D ? 4 : 3)
#endif
#define FT_CHUNK 8
#ifndef FT_UNROLL
#define FT_UNROLL 2         // components per trip of the recurrence loops
#endif
#define FT_PRAGMA_(x) _Pragma(#x)
#define FT_PRAGMA(x) FT_PRAGMA_(x)
#ifndef FT_SMALLK
#define FT_SMALLK 4         // mixtures up to this size are evaluated whole (no culling)
#endif
#define FT_LIGHT 2          // classes below this are latency-bound (no galaxy in the bin)
// bins per claimed work unit by weight class: heavy bins one at a time (balance), sky in bulk
__device__ __forceinline__ constexpr int ft_chunk(int cls) {
  return cls >= 5 ? 1 : (cls >= 3 ? 2 : (cls == 2 ? 4 : FT_CHUNK));
}

__device__ __forceinline__ int ft_bucket(int w) {
  if (w <= 0) return 0;
  if (w < 8) return 1;
  if (w < 32) return 2;
  return min(FT_BUCKETS - 1, 3 + (31 - __clz(w >> 5)));     // 32.. -> 3, 64.. -> 4, 128.. -> 5, ...
}

__device__ __forceinline__ int bins_of(const DevImage& im, int* bins_x) {
  *bins_x = (im.W + BIN_W - 1) / BIN_W;
  return *bins_x * ((im.H + BIN_H - 1) / BIN_H);
}

#ifndef BO_THREADS
#define BO_THREADS 256
#endif
__global__ void __launch_bounds__(BO_THREADS) bin_order_kernel(FieldArgs a) {
  const int g = blockIdx.x * BO_THREADS + threadIdx.x;
  int b = -1, cnt = 0;
  if (g < a.tiles_total) {
    const int img = g / a.max_tiles, t = g - img * a.max_tiles;
    int bx;
    if (t < bins_of(*a.imgs.p[img], &bx)) {
      cnt = a.tile_count[g];
      if (cnt > BIN_CAP) {
        b = FT_BUCKETS - 1;
      } else {
        // weigh the bin and put its (atomically appended, hence unordered) list
        // into source order: the tile kernel then takes the entries as they lie
        int* lst = a.tile_list + (long long)g * BIN_CAP;
        int w = 0;
        for (int j = 0; j < cnt; ++j) {
          const int key = lst[j];
          w += key & 63;
          int k = j - 1;
          while (k >= 0 && lst[k] > key) { lst[k + 1] = lst[k]; --k; }
          lst[k + 1] = key;
        }
        b = ft_bucket(w);
      }
    }
  }
  // warp-aggregated append to the class list
  const unsigned lane = threadIdx.x & 31;
  const unsigned peers = __match_any_sync(0xffffffffu, b);
  if (b >= 0) {
    const int leader = __ffs(peers) - 1;
    int base = 0;
    if ((int)lane == leader) base = atomicAdd(&a.sched[b], __popc(peers));
    base = __shfl_sync(peers, base, leader);
    // descriptor: image (8 bits) | bin column (12) | bin row (12), then bin index g (26 bits) |
    // list length clamped to 63 (6 bits; > BIN_CAP = overflow) -- the tile kernel needs no
    // division to place the bin and can fetch the NEXT bin's list before it gets there
    const int img = g / a.max_tiles, t = g - img * a.max_tiles;
    int bx;
    bins_of(*a.imgs.p[img], &bx);
    const int by = t / bx;
    a.order[(long long)b * a.tiles_total + base + __popc(peers & ((1u << lane) - 1u))] =
        make_int2((img << 24) | ((t - by * bx) << 12) | by, (g << 6) | min(cnt, 63));
  }
}

// evaluate the compacted mixture (8 floats per component in the warp's shared
// table: one 16-byte + one 8-byte load) at the lane's pixels: two rows / one row.
// The table entry of component k+1 is loaded before component k is evaluated
// (the table has one spare row, so the last prefetch needs no predicate).
__device__ __forceinline__ void ft_eval2(const float* tb, int n, float y0l, float y1l,
                                         const float (&xl)[4], float (&acc0)[4], float (&acc1)[4]) {
  float4 t0 = *reinterpret_cast<const float4*>(tb);      // A, B, C, b10
  float2 t1 = *reinterpret_cast<const float2*>(tb + 4);  // c01, c00
#pragma unroll 2
  for (int k = 0; k < n; ++k) {
    tb += 8;
    const float4 n0 = *reinterpret_cast<const float4*>(tb);
    const float2 n1 = *reinterpret_cast<const float2*>(tb + 4);
    const float b1a = fmaf(t0.y, y0l, t0.w), b1b = fmaf(t0.y, y1l, t0.w);
    const float c1a = fmaf(fmaf(t0.z, y0l, t1.x), y0l, t1.y);
    const float c1b = fmaf(fmaf(t0.z, y1l, t1.x), y1l, t1.y);
    float q[8];
#pragma unroll
    for (int p = 0; p < 4; ++p) {
      q[p] = fmaf(fmaf(t0.x, xl[p], b1a), xl[p], c1a);
      q[4 + p] = fmaf(fmaf(t0.x, xl[p], b1b), xl[p], c1b);
    }
#pragma unroll
    for (int p = 0; p < 8; ++p) q[p] = ex2_approx(q[p]);
#pragma unroll
    for (int p = 0; p < 4; ++p) { acc0[p] += q[p]; acc1[p] += q[4 + p]; }
    t0 = n0; t1 = n1;
  }
}
__device__ __forceinline__ void ft_eval1(const float* tb, int n, float yl,
                                         const float (&xl)[4], float (&acc)[4]) {
  float4 t0 = *reinterpret_cast<const float4*>(tb);
  float2 t1 = *reinterpret_cast<const float2*>(tb + 4);
#pragma unroll 2
  for (int k = 0; k < n; ++k) {
    tb += 8;
    const float4 n0 = *reinterpret_cast<const float4*>(tb);
    const float2 n1 = *reinterpret_cast<const float2*>(tb + 4);
    const float b1 = fmaf(t0.y, yl, t0.w);
    const float c1 = fmaf(fmaf(t0.z, yl, t1.x), yl, t1.y);
    float q[4];
#pragma unroll
    for (int p = 0; p < 4; ++p) q[p] = fmaf(fmaf(t0.x, xl[p], b1), xl[p], c1);
#pragma unroll
    for (int p = 0; p < 4; ++p) q[p] = ex2_approx(q[p]);
#pragma unroll
    for (int p = 0; p < 4; ++p) acc[p] += q[p];
    t0 = n0; t1 = n1;
  }
}

// Row recurrence (FT_RECUR): along a row the component is exp2 of a quadratic in x, so
// its values at x0..x0+3 are e0, e0*g0, e0*g0*(g0*h), ... with g0 = 2^(q(x0+1)-q(x0))
// and h = 2^(2A): TWO exp2 and five multiplies for four pixels instead of four exp2
// and six FMAs; the second row's g0 is the first row's times k8 = 2^(8B).  3 MUFU.EX2
// per component for the lane's 8 pixels instead of 8: the XU pipe stops being the
// bound and the loop becomes issue-bound (35 instructions per 8 pixels, was 40; 22 with
// the rows packed into FP32x2 instructions, FT_PACKED).
// Rejected: the second row's exponent as q(x0, y) + 8 (B x0 + C (2y + 8) + c01) -- 2 FFMA fewer,
// 1.5 % faster, but the error of the difference is multiplied by 8: worst model-image error
// 1.36e-6 instead of 4.5e-7 against the float64 oracle, linearity test failed at 1e-6.
// Every product stays the true pixel value times roundings (6 roundings + 3 exp2 at
// the fourth pixel, ~1.5e-6 relative); it is taken only when cull_components proved
// |q(x+1)-q(x)| <= FT_RECUR_DMAX on the lanes that count (no overflow of g) and that a
// first pixel flushed to zero cannot hide a value that counts (recur_flush_unsafe, decided per component in prepare_kernel),
// otherwise ft_eval2/1.
__device__ __forceinline__ void ft_eval2r(const float* tb, int n, float y0l, float y1l,
                                          float x0, float (&acc0)[4], float (&acc1)[4]) {
  const float tx = 2.f * x0 + 1.f;
  float4 t0 = *reinterpret_cast<const float4*>(tb);      // A, B, C, b10
  float4 t1 = *reinterpret_cast<const float4*>(tb + 4);  // c01, c00, h, k8
#if FT_PACKED
  // Packed FP32 (FT_PACKED): the two rows of a lane run as (row r, row r + 8) register pairs.
  // They share every ratio update -- 5 FMUL2 + 4 FADD2 instead of 10 FMUL + 8 FADD per
  // component -- and, with FT_PACKQ, the row terms and the first exponent too (5 FFMA2
  // instead of 10 FFMA).  Each half rounds as the scalar operation does (ptxas contracts the
  // last multiply-add of a row into one FFMA2: one rounding fewer, never more).
  f32x2 pacc[4];
#pragma unroll
  for (int p = 0; p < 4; ++p) pacc[p] = pk2(acc0[p], acc1[p]);
  const f32x2 yy = pk2(y0l, y1l), xx = pk2(x0, x0);
#endif
FT_PRAGMA(unroll FT_UNROLL)
  for (int k = 0; k < n; ++k) {
    tb += 8;
    const float4 n0 = *reinterpret_cast<const float4*>(tb);
    const float4 n1 = *reinterpret_cast<const float4*>(tb + 4);
#if FT_PACKED && FT_PACKQ
    const f32x2 b1 = fma2(pk2(t0.y, t0.y), yy, pk2(t0.w, t0.w));
    const f32x2 c1 = fma2(fma2(pk2(t0.z, t0.z), yy, pk2(t1.x, t1.x)), yy, pk2(t1.y, t1.y));
    const f32x2 q = fma2(fma2(pk2(t0.x, t0.x), xx, b1), xx, c1);
    float qa, qb, b1a, b1b;
    unpk2(q, qa, qb);
    unpk2(b1, b1a, b1b);
#else
    const float b1a = fmaf(t0.y, y0l, t0.w), b1b = fmaf(t0.y, y1l, t0.w);
    const float qa = fmaf(fmaf(t0.x, x0, b1a), x0, fmaf(fmaf(t0.z, y0l, t1.x), y0l, t1.y));
    const float qb = fmaf(fmaf(t0.x, x0, b1b), x0, fmaf(fmaf(t0.z, y1l, t1.x), y1l, t1.y));
#endif
    const float ea = ex2_approx(qa);
    const float eb = ex2_approx(qb);
    const float ga = ex2_approx(fmaf(t0.x, tx, b1a));
#if FT_PACKED
    recur_row4x2(pk2(ea, eb), pk2(ga, ga * t1.w), pk2(t1.z, t1.z), pacc);
#else
    recur_row4(ea, ga, t1.z, acc0);
    recur_row4(eb, ga * t1.w, t1.z, acc1);
#endif
    t0 = n0; t1 = n1;
  }
#if FT_PACKED
#pragma unroll
  for (int p = 0; p < 4; ++p) unpk2(pacc[p], acc0[p], acc1[p]);
#endif
}
__device__ __forceinline__ void ft_eval1r(const float* tb, int n, float yl, float x0, float (&acc)[4]) {
  const float tx = 2.f * x0 + 1.f;
  float4 t0 = *reinterpret_cast<const float4*>(tb);
  float4 t1 = *reinterpret_cast<const float4*>(tb + 4);
FT_PRAGMA(unroll FT_UNROLL)
  for (int k = 0; k < n; ++k) {
    tb += 8;
    const float4 n0 = *reinterpret_cast<const float4*>(tb);
    const float4 n1 = *reinterpret_cast<const float4*>(tb + 4);
    const float b1 = fmaf(t0.y, yl, t0.w);
    const float c1 = fmaf(fmaf(t0.z, yl, t1.x), yl, t1.y);
    recur_row4(ex2_approx(fmaf(fmaf(t0.x, x0, b1), x0, c1)), ex2_approx(fmaf(t0.x, tx, b1)), t1.z, acc);
    t0 = n0; t1 = n1;
  }
}

struct FtGeom { int X0, Y0, x, yA, yB; };

// one source on the warp's bin: cull the components against the part of the rectangle bx that lies
// in the bin and evaluate the survivors at the lane's pixels -> UNIT-flux values acc0 (row yA) and
// acc1 (row yB), unmasked (lanes outside bx hold values the caller must not use).  Returns false
// (warp-uniform) when bx misses the bin.
__device__ __forceinline__ bool ft_source_rates(const FieldArgs& a, int item, int K, int4 bx, int2 o,
                                                const FtGeom& q, int lane, float* wtab,
                                                float (&acc0)[4], float (&acc1)[4], unsigned long long& issued) {
  const int ix0 = max(q.X0, bx.x), ix1 = min(q.X0 + BIN_W, bx.y);
  const int iy0 = max(q.Y0, bx.z), iy1 = min(q.Y0 + BIN_H, bx.w);
  if (ix0 >= ix1 || iy0 >= iy1 || K <= 0) return false;          // warp-uniform
  const float* gtab = a.tab + (long long)item * TAB_STRIDE;
  int n = K;
  bool recur = false;
  __syncwarp();                       // the warp is done reading its previous table
  if ((a.cull & 1) && n > 1 && n <= FT_SMALLK) {
    // a star's PSF mixture: culling would cost more than the few evaluations it can save;
    // copy the quadratics, add (h, k8) and the recurrence range check of cull_components
    const float xa = (float)(ix0 - o.x), xb = (float)(ix1 - 1 - o.x), ya = (float)(iy0 - o.y), yb = (float)(iy1 - 1 - o.y);
    float dmx = 0.f;
    if (lane < n) {
      const float4* t4 = reinterpret_cast<const float4*>(gtab + lane * TAB_W);
      const float4 e0 = __ldg(t4), e1 = __ldg(t4 + 1);
      if (__ldg(gtab + lane * TAB_W + 9) < 0.f) dmx = INFINITY;    // recur_flush_unsafe (prepare_kernel)
      const float fx = fmaxf(fabsf(xa - e1.z), fabsf(xb - e1.z)), fy = fmaxf(fabsf(ya - e1.w), fabsf(yb - e1.w));
      dmx = fmaxf(dmx, fmaxf(fmaf(fabsf(e0.x), 2.f * (fx + 3.f) + 1.f, fabsf(e0.y) * fy), 8.f * fabsf(e0.y)));
      *reinterpret_cast<float4*>(wtab + lane * 8) = e0;
      *reinterpret_cast<float4*>(wtab + lane * 8 + 4) = make_float4(e1.x, e1.y, ex2_approx(2.f * e0.x), ex2_approx(8.f * e0.y));
    }
    recur = !__any_sync(0xffffffffu, !(dmx <= FT_RECUR_DMAX)) && FT_RECUR && !(a.cull & 2);
    __syncwarp();
#ifdef FT_ABLATE_CULL       // timing experiment: galaxies evaluate a fixed number of components, no culling
  } else if ((a.cull & 1) && n > 1) {
    for (int k = lane; k < n; k += 32) {
      const float4* t4 = reinterpret_cast<const float4*>(gtab + k * TAB_W);
      const float4 e0 = __ldg(t4), e1 = __ldg(t4 + 1);
      *reinterpret_cast<float4*>(wtab + k * 8) = e0;
      *reinterpret_cast<float4*>(wtab + k * 8 + 4) = make_float4(e1.x, e1.y, 1.f, 1.f);
    }
    n = min(n, FT_ABLATE_CULL);
    recur = true;
    __syncwarp();
#endif
  } else if ((a.cull & 1) && n > 1) {
    n = cull_components<true>(gtab, n, lane, (float)(ix0 - o.x), (float)(ix1 - 1 - o.x), (float)(iy0 - o.y),
                              (float)(iy1 - 1 - o.y), 3.f, wtab, &recur);
    recur = recur && FT_RECUR && !(a.cull & 2);
  } else {
    // no culling: copy the quadratics as they are (the eval loops read shared memory only)
    for (int k = lane; k < n; k += 32) {
      const float4* t4 = reinterpret_cast<const float4*>(gtab + k * TAB_W);
      const float4 e0 = __ldg(t4), e1 = __ldg(t4 + 1);
      *reinterpret_cast<float4*>(wtab + k * 8) = e0;
      *reinterpret_cast<float2*>(wtab + k * 8 + 4) = make_float2(e1.x, e1.y);
    }
    __syncwarp();
  }
  const float* tb = wtab;
  float xl[4];
#pragma unroll
  for (int p = 0; p < 4; ++p) xl[p] = (float)(q.x + p - o.x);
  const bool doA = iy0 < q.Y0 + 8, doB = iy1 > q.Y0 + 8;         // which row halves the footprint reaches
#pragma unroll
  for (int p = 0; p < 4; ++p) { acc0[p] = 0.f; acc1[p] = 0.f; }
#ifdef FT_ABLATE_EVAL       // timing experiment (tools/build_variant.py): everything but the evaluation loops
  if (n >= 0) { acc0[0] = (float)n * 1e-12f + fabsf(tb[0]) * 1e-30f; if (lane == 0) issued += recur; return true; }
#endif
  if (doA && doB) {
    if (recur) ft_eval2r(tb, n, (float)(q.yA - o.y), (float)(q.yB - o.y), xl[0], acc0, acc1);
    else ft_eval2(tb, n, (float)(q.yA - o.y), (float)(q.yB - o.y), xl, acc0, acc1);
    if (lane == 0) issued += (unsigned long long)n * 256;
  } else if (doA) {
    if (recur) ft_eval1r(tb, n, (float)(q.yA - o.y), xl[0], acc0);
    else ft_eval1(tb, n, (float)(q.yA - o.y), xl, acc0);
    if (lane == 0) issued += (unsigned long long)n * 128;
  } else {
    if (recur) ft_eval1r(tb, n, (float)(q.yB - o.y), xl[0], acc1);
    else ft_eval1(tb, n, (float)(q.yB - o.y), xl, acc1);
    if (lane == 0) issued += (unsigned long long)n * 128;
  }
  return true;
}

// ... and add the rates inside the footprint to lambda (the field kernel's use)
__device__ __forceinline__ void ft_source(const FieldArgs& a, int item, int K, int4 bx, int2 o, float fl,
                                          const FtGeom& q, int lane, float* wtab,
                                          float (&lam)[2][4], unsigned long long& issued) {
  float acc0[4], acc1[4];
  if (!ft_source_rates(a, item, K, bx, o, q, lane, wtab, acc0, acc1, issued)) return;
#ifdef FT_ABLATE_LAM        // timing experiment: no masks, no lambda update
  lam[0][0] += acc0[0]; return;
#endif
  if (min(q.X0 + BIN_W, bx.y) - max(q.X0, bx.x) == BIN_W &&
      min(q.Y0 + BIN_H, bx.w) - max(q.Y0, bx.z) == BIN_H) {       // the footprint covers the bin: no masks
#pragma unroll
    for (int p = 0; p < 4; ++p) {
      lam[0][p] += acc0[p] * fl;
      lam[1][p] += acc1[p] * fl;
    }
    return;
  }
  const bool inA = q.yA >= bx.z && q.yA < bx.w, inB = q.yB >= bx.z && q.yB < bx.w;
#pragma unroll
  for (int p = 0; p < 4; ++p) {
    const bool inx = (q.x + p >= bx.x) && (q.x + p < bx.y);
    lam[0][p] += (inA && inx) ? acc0[p] * fl : 0.f;
    lam[1][p] += (inB && inx) ? acc1[p] * fl : 0.f;
  }
}

// 3 CTAs x 8 warps per SM at 80 registers (no spills; 4 CTAs at 64 registers spill the pixel counts and measured slower)
__global__ void __launch_bounds__(FT_THREADS, FT_MINB) field_tile_kernel(FieldArgs a) {
  __shared__ __align__(16) float s_wtab[FT_THREADS / 32][(CG_MAX_COMP + 1) * 8];   // + the prefetch row
#if FT_STAGE_D
  // the bin's pixel counts wait in shared memory (cp.async) until the epilogue: 8 registers fewer
  // through the source loop, which is what lets the kernel run 4 CTAs per SM at 64 registers
  __shared__ __align__(16) float4 s_d[FT_THREADS / 32][2][32];
#endif
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  float* const wtab = s_wtab[warp];
  const LogTable* const logt = a.logtab;

  // class sizes (written by bin_order_kernel, complete before this launch)
  __shared__ int s_csize[FT_BUCKETS];
  if (threadIdx.x < FT_BUCKETS) s_csize[threadIdx.x] = a.sched[threadIdx.x];
  __syncthreads();
  int total = 0;
#pragma unroll
  for (int b = 0; b < FT_BUCKETS; ++b) total += s_csize[b];
  unsigned long long issued = 0;

  // Work is claimed in UNITS of 1 (heavy classes) to FT_CHUNK (sky) bins of one
  // class -- one same-address atomic per unit -- two units ahead: while unit c0
  // is processed the descriptors of unit c1 are loading (lane j holds bin j of
  // the unit) and the atomic for c2 is in flight.
  // Units of the LIGHT classes (sky, stars only: a chain of dependent loads and
  // a handful of exps -- latency-bound) are interleaved evenly with the HEAVY
  // ones (XU-bound), heaviest first within each pool, so that the SM always has
  // arithmetic to issue while the light warps wait on memory.
  int n_light = 0, n_heavy = 0;
#pragma unroll
  for (int b = 0; b < FT_BUCKETS; ++b) {
    const int nu = (s_csize[b] + ft_chunk(b) - 1) / ft_chunk(b);
    if (b < FT_LIGHT) n_light += nu; else n_heavy += nu;
  }
  total = n_light + n_heavy;                                   // units
  auto load_unit = [&](int u) -> int2 {
    // Bresenham split of the unit index into the two pools
    const int lb = (int)(((long long)u * n_light) / total);           // light units before u
    const bool light = (int)(((long long)(u + 1) * n_light) / total) > lb;
    int v = light ? lb : u - lb;
    int cls = light ? FT_LIGHT - 1 : FT_BUCKETS - 1;
    const int lo = light ? 0 : FT_LIGHT;
#pragma unroll
    for (int b = FT_BUCKETS - 1; b > 0; --b) {
      const int nu = (s_csize[b] + ft_chunk(b) - 1) / ft_chunk(b);
      if (cls == b && b > lo && v >= nu) { v -= nu; cls = b - 1; }
    }
    const int first = v * ft_chunk(cls) + lane;
    int2 dsc = make_int2(-1, 0);
    if (lane < ft_chunk(cls) && first < s_csize[cls]) dsc = a.order[(long long)cls * a.tiles_total + first];
    return dsc;
  };
  int c0 = 0, c1 = 0;
  if (lane == 0) {
    c0 = atomicAdd(&a.sched[FT_BUCKETS], 1);
    c1 = atomicAdd(&a.sched[FT_BUCKETS], 1);
  }
  c0 = __shfl_sync(0xffffffffu, c0, 0);
  c1 = __shfl_sync(0xffffffffu, c1, 0);
  int2 dcur = make_int2(-1, 0);
  if (c0 < total) dcur = load_unit(c0);
  // lane's entry of a bin's source list (descriptor word y = g << 6 | length); overflowed lists are scanned instead
  auto list_entry = [&](int gy) -> int {
    const int n = gy & 63;
    return (n <= BIN_CAP && lane < n) ? __ldg(a.tile_list + (long long)(gy >> 6) * BIN_CAP + lane) : 0x7fffffff;
  };
  int e_nx = __shfl_sync(0xffffffffu, dcur.x, 0) >= 0 ? list_entry(__shfl_sync(0xffffffffu, dcur.y, 0)) : 0x7fffffff;

  while (c0 < total) {
    int c2 = 0;
    if (lane == 0) c2 = atomicAdd(&a.sched[FT_BUCKETS], 1);
    int2 dnext = make_int2(-1, 0);
    if (c1 < total) dnext = load_unit(c1);

   for (int cj = 0; cj < FT_CHUNK; ++cj) {
    const int dsc = __shfl_sync(0xffffffffu, dcur.x, cj), gy = __shfl_sync(0xffffffffu, dcur.y, cj);
    if (dsc < 0) break;
    const int cnt = gy & 63, g = gy >> 6;
    const int img = dsc >> 24;
    CG_CHECK(img >= 0 && img < a.n_img && g >= 0 && g < a.tiles_total);
    const DevImage& im = *a.imgs.p[img];
    const int H = im.H, W = im.W;
    FtGeom q;
    q.X0 = ((dsc >> 12) & 0xfff) * BIN_W; q.Y0 = (dsc & 0xfff) * BIN_H;
    // this bin's list entry was fetched while the previous bin was processed; fetch the next
    // bin's now if it belongs to this unit (else before the epilogue, from the next unit)
    const int e_cur = e_nx;
    bool have_nx = false;
    if (cj + 1 < FT_CHUNK && __shfl_sync(0xffffffffu, dcur.x, (cj + 1) & 31) >= 0) {
      e_nx = list_entry(__shfl_sync(0xffffffffu, dcur.y, (cj + 1) & 31));
      have_nx = true;
    }
    q.x = q.X0 + (lane & 3) * 4; q.yA = q.Y0 + (lane >> 2); q.yB = q.yA + 8;

    // pixel counts: independent of the source list, issue the loads first
#if FT_STAGE_D
    {
      const bool vec = ((W & 3) == 0) && (((uintptr_t)im.nelec & 15) == 0);
#pragma unroll
      for (int r = 0; r < 2; ++r) {
        const int y = r ? q.yB : q.yA;
        float4* dst = &s_d[warp][r][lane];
        if (y < H && q.x < W) {
          const float* rp = im.nelec + (long long)y * W + q.x;
          if (vec) {
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((unsigned)__cvta_generic_to_shared(dst)), "l"(rp) : "memory");
          } else {
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            v.x = rp[0];
            if (q.x + 1 < W) v.y = rp[1];
            if (q.x + 2 < W) v.z = rp[2];
            if (q.x + 3 < W) v.w = rp[3];
            *dst = v;
          }
        } else {
          *dst = make_float4(0.f, 0.f, 0.f, 0.f);
        }
      }
      asm volatile("cp.async.commit_group;" ::: "memory");
    }
#define FT_LOAD_D()                                                             \
    float d[2][4];                                                              \
    {                                                                           \
      asm volatile("cp.async.wait_group 0;" ::: "memory");                      \
      const float4 v0 = s_d[warp][0][lane], v1 = s_d[warp][1][lane];            \
      d[0][0] = v0.x; d[0][1] = v0.y; d[0][2] = v0.z; d[0][3] = v0.w;           \
      d[1][0] = v1.x; d[1][1] = v1.y; d[1][2] = v1.z; d[1][3] = v1.w;           \
    }
#else
    float d[2][4];
    {
      const bool vec = ((W & 3) == 0) && (((uintptr_t)im.nelec & 15) == 0);
#pragma unroll
      for (int r = 0; r < 2; ++r) {
        const int y = r ? q.yB : q.yA;
#pragma unroll
        for (int p = 0; p < 4; ++p) d[r][p] = 0.f;
        if (y < H && q.x < W) {
          const float* rp = im.nelec + (long long)y * W + q.x;
          if (vec) {
#ifdef FT_ABLATE_LOAD       // timing experiment: no pixel traffic
            const float4 v = make_float4((float)y, (float)q.x, 1.f, 2.f);
#else
            const float4 v = __ldcs(reinterpret_cast<const float4*>(rp));     // read once: streaming
#endif
            d[r][0] = v.x; d[r][1] = v.y; d[r][2] = v.z; d[r][3] = v.w;
          } else {
#pragma unroll
            for (int p = 0; p < 4; ++p)
              if (q.x + p < W) d[r][p] = rp[p];
          }
        }
      }
    }
#define FT_LOAD_D()
#endif
    const float eps = (float)im.epsilon;
    float* mdl = a.model ? a.model[img] : nullptr;
    double s = 0.0;

    if (cnt == 0 && !mdl) {
      if (!have_nx) e_nx = __shfl_sync(0xffffffffu, dnext.x, 0) >= 0 ? list_entry(__shfl_sync(0xffffffffu, dnext.y, 0)) : 0x7fffffff;
      // no source reaches this bin: lambda = epsilon on every pixel, so its terms
      // are log(eps) * sum(x) - n * eps -- one log per warp instead of one per pixel
      FT_LOAD_D()
      double sd = 0.0;
      int npx = 0;
#pragma unroll
      for (int r = 0; r < 2; ++r) {
        sd += ((double)d[r][0] + (double)d[r][1]) + ((double)d[r][2] + (double)d[r][3]);
        if ((r ? q.yB : q.yA) < H) npx += max(0, min(4, W - q.x));
      }
      s = fma(sd, log_accurate(eps, logt), -(double)npx * (double)eps);
    } else {
      float lam[2][4];
#pragma unroll
      for (int r = 0; r < 2; ++r)
#pragma unroll
        for (int p = 0; p < 4; ++p) lam[r][p] = eps;

      // candidates per round, one per lane: the bin's list (one round; lane j owns
      // entry j) or, when the list overflowed, every source of the image 32 at a
      // time.  They are taken in source order: rank by value / by lane.
      const bool scan = cnt > BIN_CAP;
      const int ns = scan ? a.n_src[img] : 0;
      const int rounds = scan ? (ns + 31) / 32 : 1;
      for (int rd = 0; rd < rounds; ++rd) {
        int e = 0x7fffffff;
        int4 bx = make_int4(0, 0, 0, 0);
        int2 o = make_int2(0, 0);
        float fl = 0.f;
        bool valid = false;
        if (!scan) {
          if (lane < cnt) {
            e = e_cur;
            valid = true;
          }
        } else {
          const int sidx = rd * 32 + lane;
          if (sidx < ns) {
            e = (a.item_base[img] + sidx * a.item_stride[img]) * 64;
            valid = true;
          }
        }
        if (valid) {
          const ItemHdr* hp = a.hdr + (e >> 6);
          bx = *reinterpret_cast<const int4*>(hp);
          o = *reinterpret_cast<const int2*>(&hp->ox);
          fl = (float)hp->flux;
          if (scan) {
            const int K = hp->K;
            e += K;
            valid = (K > 0) && bx.x < q.X0 + BIN_W && bx.y > q.X0 && bx.z < q.Y0 + BIN_H && bx.w > q.Y0;
          }
        }
        const unsigned vm = __ballot_sync(0xffffffffu, valid);
        const int nv = __popc(vm);
        const int rk = __popc(vm & ((1u << lane) - 1u));         // lane order = source order (sorted list / scan)
#ifdef FT_ABLATE_PAIRS      // timing experiment: the per-(bin, source) work removed
        for (int r = 0; r < 0; ++r) {
#else
        for (int r = 0; r < nv; ++r) {
#endif
          const unsigned m = __ballot_sync(0xffffffffu, valid && rk == r);
          const int sl = __ffs(m) - 1;
          const int ee = __shfl_sync(0xffffffffu, e, sl);
          int4 b;
          b.x = __shfl_sync(0xffffffffu, bx.x, sl); b.y = __shfl_sync(0xffffffffu, bx.y, sl);
          b.z = __shfl_sync(0xffffffffu, bx.z, sl); b.w = __shfl_sync(0xffffffffu, bx.w, sl);
          const int2 oo = make_int2(__shfl_sync(0xffffffffu, o.x, sl), __shfl_sync(0xffffffffu, o.y, sl));
          CG_CHECK((ee & 63) <= CG_MAX_COMP && b.x >= 0 && b.y <= im.W && b.z >= 0 && b.w <= im.H);
          ft_source(a, ee >> 6, ee & 63, b, oo, __shfl_sync(0xffffffffu, fl, sl), q, lane, wtab, lam, issued);
        }
      }

      if (!have_nx) e_nx = __shfl_sync(0xffffffffu, dnext.x, 0) >= 0 ? list_entry(__shfl_sync(0xffffffffu, dnext.y, 0)) : 0x7fffffff;
      // Poisson terms.  log(lam) = T[bin] + res (log_accurate): sum x T in float64,
      // sum x res (|res| < 1/60) and the 4-pixel row sums of lam in fp32, then float64
      FT_LOAD_D()
      float sres = 0.f;
      double slam = 0.0;
#ifdef FT_ABLATE_EPI        // timing experiment: no per-pixel logarithm
      if (q.x >= 0) {
#pragma unroll
        for (int r = 0; r < 2; ++r) slam += (double)(((lam[r][0] + lam[r][1]) + (lam[r][2] + lam[r][3])) * ((d[r][0] + d[r][1]) + (d[r][2] + d[r][3])));
      } else
#endif
      if (q.x + 3 < W && q.yB < H) {                          // interior lane: no bounds tests
#pragma unroll
        for (int r = 0; r < 2; ++r) {
#pragma unroll
          for (int p = 0; p < 4; ++p) {
            s = fma((double)d[r][p], log_table(lam[r][p], logt), s);
            sres = fmaf(d[r][p], log_residual(lam[r][p]), sres);
          }
          slam += (double)((lam[r][0] + lam[r][1]) + (lam[r][2] + lam[r][3]));
        }
        if (mdl) {
          float* m0 = mdl + (long long)q.yA * W + q.x;
          float* m1 = mdl + (long long)q.yB * W + q.x;
          if (((W & 3) == 0) && (((uintptr_t)mdl & 15) == 0)) {
            *reinterpret_cast<float4*>(m0) = make_float4(lam[0][0], lam[0][1], lam[0][2], lam[0][3]);
            *reinterpret_cast<float4*>(m1) = make_float4(lam[1][0], lam[1][1], lam[1][2], lam[1][3]);
          } else {
#pragma unroll
            for (int p = 0; p < 4; ++p) { m0[p] = lam[0][p]; m1[p] = lam[1][p]; }
          }
        }
      } else {
#pragma unroll
        for (int r = 0; r < 2; ++r) {
          const int y = r ? q.yB : q.yA;
          if (y < H) {
#pragma unroll
            for (int p = 0; p < 4; ++p)
              if (q.x + p < W) {
                s = fma((double)d[r][p], log_table(lam[r][p], logt), s);
                sres = fmaf(d[r][p], log_residual(lam[r][p]), sres);
                slam += (double)lam[r][p];
                if (mdl) mdl[(long long)y * W + q.x + p] = lam[r][p];
              }
          }
        }
      }
      s += (double)sres - slam;
    }
    s = warp_sum(s);
    if (lane == 0) a.block_partials[g] = s;
   }

    c0 = c1;
    dcur = dnext;
    c1 = __shfl_sync(0xffffffffu, c2, 0);
  }
  if (lane == 0 && issued) atomicAdd(&a.cnt->issued, issued);
}

// one block per image: fixed-shape tree over its bin sums
#ifndef FIN_THREADS
#define FIN_THREADS 1024
#endif
__global__ void __launch_bounds__(FIN_THREADS) field_tile_final_kernel(FieldArgs a) {
  const int img = blockIdx.x;
  int bx;
  const int n_bins = bins_of(*a.imgs.p[img], &bx);
  __shared__ double s_fin[FIN_THREADS];
  double t0 = 0, t1 = 0, t2 = 0, t3 = 0;
  const double* bp = a.block_partials + (long long)img * a.max_tiles;
  int i = threadIdx.x;
  for (; i + 3 * FIN_THREADS < n_bins; i += 4 * FIN_THREADS) {
    t0 += bp[i]; t1 += bp[i + FIN_THREADS]; t2 += bp[i + 2 * FIN_THREADS]; t3 += bp[i + 3 * FIN_THREADS];
  }
  for (; i < n_bins; i += FIN_THREADS) t0 += bp[i];
  s_fin[threadIdx.x] = (t0 + t1) + (t2 + t3);
  __syncthreads();
  for (int o = FIN_THREADS / 2; o > 0; o >>= 1) {
    if (threadIdx.x < o) s_fin[threadIdx.x] += s_fin[threadIdx.x + o];
    __syncthreads();
  }
  // an error raised on the device (Counters.err_flag) poisons the result: callers that never
  // synchronise to read the flag (FieldSet.loglik_device, CUDA-graph replays) still cannot miss it
  if (threadIdx.x == 0) a.ll[img] = a.cnt->err_flag ? (double)NAN : s_fin[0];
}

// ---------------------------------------------------------------------------
// proposal scoring finalize (sources.py:258, :300-306)
// ---------------------------------------------------------------------------
__global__ void score_finalize_kernel(const double* item_ll, int n_img, int P,
                                      const double* log_terms, const double* log_u, double guard,
                                      double* delta, uint8_t* accept, Counters* cnt,
                                      int* guard_list, int guard_cap) {
  int p = blockIdx.x * blockDim.x + threadIdx.x;  // 0..P-1 -> source index p+1
  if (p >= P) return;
  double cur = 0, prop = 0;
  for (int b = 0; b < n_img; ++b) {
    cur += item_ll[b];
    prop += item_ll[(long long)(p + 1) * n_img + b];
  }
  double d = (prop - cur) + (log_terms ? log_terms[p] : 0.0);
  delta[p] = d;
  double lu = log_u[p];
  accept[p] = (lu < d) ? 1 : 0;
  if (guard > 0 && fabs(lu - d) < guard) {
    int pos = atomicAdd(&cnt->guard_count, 1);
    if (pos < guard_cap) guard_list[pos] = p + 1;
    else atomicExch(&cnt->err_flag, 2);
  }
}

__global__ void score_rescore_kernel(const double* item_ll2, int n_img, const Counters* cnt1,
                                     const int* guard_list, int guard_cap,
                                     const double* log_terms, const double* log_u,
                                     double* delta, uint8_t* accept) {
  int g = blockIdx.x * blockDim.x + threadIdx.x;
  int n = min(cnt1->guard_count, guard_cap);
  if (g >= n) return;
  int src = guard_list[g];
  double cur = 0, prop = 0;
  for (int b = 0; b < n_img; ++b) {
    cur += item_ll2[b];
    prop += item_ll2[(long long)(g + 1) * n_img + b];
  }
  double d = (prop - cur) + (log_terms ? log_terms[src - 1] : 0.0);
  delta[src - 1] = d;
  accept[src - 1] = (log_u[src - 1] < d) ? 1 : 0;
}

__global__ void clamp_guard_kernel(Counters* c, int cap) {
  if (c->guard_count > cap) c->guard_count = cap;
}

// ---------------------------------------------------------------------------
// A2: gmm_like_2d on an arbitrary point list
// ---------------------------------------------------------------------------
// Component constants in float64 (log2 units); the quadratic form is evaluated
// in float64 because arbitrary covariances can be ill-conditioned (the KAT in
// CelestePy/test/test_gmm.py draws random 2x2 Wishart matrices); the
// exponential is the fp32 MUFU.EX2.
__global__ void gmm_prep_kernel(const double* ws, const double* mus, const double* sigs, int K,
                                double* tab) {
  int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= K) return;
  double s00 = sigs[k * 4 + 0], s01 = sigs[k * 4 + 1], s10 = sigs[k * 4 + 2], s11 = sigs[k * 4 + 3];
  double det = s00 * s11 - s01 * s10;
  double i00 = s11 / det, i11 = s00 / det, i01 = -1 * s01 / det;
  tab[k * 6 + 0] = mus[k * 2 + 0];
  tab[k * 6 + 1] = mus[k * 2 + 1];
  tab[k * 6 + 2] = -.5 * LOG2E_D * i00;
  tab[k * 6 + 3] = -LOG2E_D * i01;
  tab[k * 6 + 4] = -.5 * LOG2E_D * i11;
  tab[k * 6 + 5] = (-LOG2PI_D - .5 * log(det) + log(ws[k])) * LOG2E_D;
}

__global__ void __launch_bounds__(256) gmm_points_kernel(double* probs, const double* x,
                                                          const double* tab, long long N, int K) {
  extern __shared__ double s_gt[];
  for (int t = threadIdx.x; t < K * 6; t += blockDim.x) s_gt[t] = tab[t];
  __syncthreads();
  for (long long n = (long long)blockIdx.x * blockDim.x + threadIdx.x; n < N;
       n += (long long)gridDim.x * blockDim.x) {
    const double2 pt = *reinterpret_cast<const double2*>(x + 2 * n);
    float acc = 0.f;
    for (int k = 0; k < K; ++k) {
      const double dx = pt.x - s_gt[k * 6 + 0], dy = pt.y - s_gt[k * 6 + 1];
      const double arg = fma(dx, fma(s_gt[k * 6 + 2], dx, s_gt[k * 6 + 3] * dy),
                             fma(s_gt[k * 6 + 4] * dy, dy, s_gt[k * 6 + 5]));
      acc += ex2_approx((float)arg);
    }
    probs[n] = (double)acc;
  }
}

// gmm_like_fast.pyx:34-77
struct MvnArgs { double mean[2], ci[4], logdet; };
__global__ void mvn2d_logpdf_kernel(double* lls, const double* x, MvnArgs m, long long N) {
  long long n = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= N) return;
  double x0 = x[2 * n], x1 = x[2 * n + 1];
  double mean_sig = m.mean[0] * m.mean[0] * m.ci[0] + m.mean[1] * m.mean[1] * m.ci[3] +
                    2. * m.mean[0] * m.mean[1] * m.ci[1];
  double xs0 = x0 * m.ci[0] + x1 * m.ci[2];
  double xs1 = x0 * m.ci[1] + x1 * m.ci[3];
  double quad = (x0 * xs0 + x1 * xs1) - 2. * (m.mean[0] * xs0 + m.mean[1] * xs1) + mean_sig;
  lls[n] = -.5 * 2 * LOG2PI_D - .5 * m.logdet - .5 * quad;
}

// ---------------------------------------------------------------------------
// A16 / A17
// ---------------------------------------------------------------------------
__global__ void planck_kernel(const double* T, const double* b, long long n, double* out,
                              const FilterTables* f) {
  long long w = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  int lane = threadIdx.x & 31;
  if (w >= n * CG_NUM_BANDS) return;
  long long i = w / CG_NUM_BANDS;
  int band = (int)(w % CG_NUM_BANDS);
  double ppj = planck_photons_per_joule_warp(f, T[i], band, lane);
  if (lane == 0) {
    const double c_ = 299792458.0;
    const double m_per_ly = c_ * 31556952.;
    const double lens_area = .75 * PI_D * (1.25 * 1.25);
    double lens_watts = lens_area * b[i] * (3.846 * 1e26) / (m_per_ly * m_per_ly);
    out[w] = ppj * lens_watts * 54.;
  }
}

// np.interp semantics (clamped ends)
__device__ inline double interp_clamped(double xq, const double* xp_m, const double* fp, int n) {
  // xp = wavelength_lookup*(10**10)   (redshift_utils.py:308)
  const double sc = 1e10;
  if (xq <= xp_m[0] * sc) return fp[0];
  if (xq >= xp_m[n - 1] * sc) return fp[n - 1];
  int lo = 0, hi = n - 1;
  while (hi - lo > 1) {
    int mid = (lo + hi) >> 1;
    if (xp_m[mid] * sc <= xq) lo = mid; else hi = mid;
  }
  double x0 = xp_m[lo] * sc, x1 = xp_m[lo + 1] * sc;
  double slope = (fp[lo + 1] - fp[lo]) / (x1 - x0);
  return slope * (xq - x0) + fp[lo];
}

__global__ void __launch_bounds__(256) eigen_flux_kernel(const double* z, const double* w,
                                                          const double* m, const double* B,
                                                          const double* lam0, int K, int V,
                                                          const double* fluxes, const double* ivar,
                                                          double* mu, double* ll,
                                                          const FilterTables* f) {
  const int c = blockIdx.x, tid = threadIdx.x;
  const double zz = z[c];
  double acc[CG_NUM_BANDS], nrm[CG_NUM_BANDS];
#pragma unroll
  for (int b = 0; b < CG_NUM_BANDS; ++b) { acc[b] = 0; nrm[b] = 0; }
  for (int v = tid; v < V; v += 256) {
    double spec = 0;
    for (int k = 0; k < K; ++k) spec += w[(long long)c * K + k] * B[(long long)k * V + v];
    double lam = lam0[v] * (1. + zz);
    double f2 = lam * lam / 2.99792e18;
#pragma unroll
    for (int b = 0; b < CG_NUM_BANDS; ++b) {
      double sv = interp_clamped(lam, f->wl + f->start[b], f->sens + f->start[b], f->count[b]);
      nrm[b] += sv;
      acc[b] += sv * spec * f2;
    }
  }
  __shared__ double s_red[8][2 * CG_NUM_BANDS];
  const int lane = tid & 31, warp = tid >> 5;
#pragma unroll
  for (int b = 0; b < CG_NUM_BANDS; ++b) {
    double a1 = warp_sum(acc[b]), n1 = warp_sum(nrm[b]);
    if (lane == 0) { s_red[warp][b] = a1; s_red[warp][CG_NUM_BANDS + b] = n1; }
  }
  __syncthreads();
  if (tid == 0) {
    double l = 0;
    for (int b = 0; b < CG_NUM_BANDS; ++b) {
      double a1 = 0, n1 = 0;
      for (int wv = 0; wv < 8; ++wv) { a1 += s_red[wv][b]; n1 += s_red[wv][CG_NUM_BANDS + b]; }
      double fl = (n1 <= 0.) ? 0. : (a1 / n1) * 275422870333.81744384765625;
      double mub = fl * m[c] / (1. + zz);
      mu[(long long)c * CG_NUM_BANDS + b] = mub;
      if (fluxes) {
        double d = fluxes[(long long)c * CG_NUM_BANDS + b] - mub;
        l += ivar[(long long)c * CG_NUM_BANDS + b] * d * d;
      }
    }
    if (fluxes && ll) ll[c] = isinf(m[c]) ? -INFINITY : -0.5 * l;
  }
}

// ---------------------------------------------------------------------------
// sources.py:6-12 on arbitrary arrays (float64 in, float64 arithmetic)
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(256) poisson_partial_kernel(const double* data, const double* model,
                                                               const double* mask, long long n,
                                                               double* partials) {
  double s = 0;
  for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < n; i += (long long)gridDim.x * 256) {
    const double m = model[i];
    if (m > 0. && (!mask || mask[i] != 0.)) s += log(m) * data[i] - m;
  }
  __shared__ double s_red[8];
  s = warp_sum(s);
  if ((threadIdx.x & 31) == 0) s_red[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x == 0) {
    double t = 0;
    for (int w = 0; w < 8; ++w) t += s_red[w];
    partials[blockIdx.x] = t;
  }
}
__global__ void __launch_bounds__(256) sum_partials_kernel(const double* partials, int n, double* out) {
  __shared__ double s_fin[256];
  double t = 0;
  for (int i = threadIdx.x; i < n; i += 256) t += partials[i];
  s_fin[threadIdx.x] = t;
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) {
    if (threadIdx.x < o) s_fin[threadIdx.x] += s_fin[threadIdx.x + o];
    __syncthreads();
  }
  if (threadIdx.x == 0) *out = s_fin[0];
}

// ---------------------------------------------------------------------------
// N3: empirical priors (util/dists/mog.py:5-21, flux_prior.py) and their
// sufficient statistics
// ---------------------------------------------------------------------------
// one warp per point; lane k (and k + 32, ...) owns mixture component k
__global__ void __launch_bounds__(128) mog_logpdf_kernel(const double* x, long long N, int D,
                                                         const double* means, const double* icovs,
                                                         const double* dets, const double* pis, int K,
                                                         double* logp, double* resp) {
  const int lane = threadIdx.x & 31;
  const long long n = (long long)blockIdx.x * 4 + (threadIdx.x >> 5);
  if (n >= N) return;
  double xv[CG_MOG_MAX_DIM];
  for (int d = 0; d < D; ++d) xv[d] = x[n * D + d];
  double mx = -INFINITY;
  for (int k = lane; k < K; k += 32) {
    double c[CG_MOG_MAX_DIM];
    for (int d = 0; d < D; ++d) c[d] = xv[d] - means[k * D + d];
    double q = 0.0;                       // solved_m = sum_j icov[j][m] c_j ; q = sum_m solved_m c_m (mog.py:15-16)
    for (int m = 0; m < D; ++m) {
      double sm = 0.0;
      for (int j = 0; j < D; ++j) sm += icovs[((long long)k * D + j) * D + m] * c[j];
      q += sm * c[m];
    }
    const double lp = -0.5 * q - LOG2PI_D - 0.5 * log(dets[k]) + log(pis[k]);
    if (resp) resp[n * K + k] = lp;       // finished below
    mx = fmax(mx, lp);
  }
  mx = warp_max(mx);
  double se = 0.0;
  // second pass: exp(lp - max); lp is recomputed from resp when stored, else from scratch
  for (int k = lane; k < K; k += 32) {
    double lp;
    if (resp) {
      lp = resp[n * K + k];
    } else {
      double c[CG_MOG_MAX_DIM];
      for (int d = 0; d < D; ++d) c[d] = xv[d] - means[k * D + d];
      double q = 0.0;
      for (int m = 0; m < D; ++m) {
        double sm = 0.0;
        for (int j = 0; j < D; ++j) sm += icovs[((long long)k * D + j) * D + m] * c[j];
        q += sm * c[m];
      }
      lp = -0.5 * q - LOG2PI_D - 0.5 * log(dets[k]) + log(pis[k]);
    }
    se += (mx == -INFINITY) ? 0.0 : exp(lp - mx);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) se += __shfl_xor_sync(0xffffffffu, se, o);
  const double lse = (mx == -INFINITY) ? -INFINITY : mx + log(se);
  if (lane == 0) logp[n] = lse;
  if (resp)
    for (int k = lane; k < K; k += 32) resp[n * K + k] = exp(resp[n * K + k] - lse);
}

// one block per output entry e of [K | K*D | K*D(D+1)/2]: fixed-shape tree over the points
__global__ void __launch_bounds__(256) prior_stats_kernel(const double* resp, const double* c, long long N,
                                                          int K, int D, double* out, const int* mask = nullptr) {
  const int T = D * (D + 1) / 2;
  const int e = blockIdx.x;
  int k, a = -1, b = -1;
  if (e < K) {
    k = e;
  } else if (e < K + K * D) {
    k = (e - K) / D; a = (e - K) % D;
  } else {
    const int r = e - K - K * D;
    k = r / T;
    int t = r % T;
    a = 0;
    while (t >= D - a) { t -= D - a; ++a; }      // upper triangle, row-major (np.triu_indices)
    b = a + t;
  }
  double s = 0.0;
  for (long long n = threadIdx.x; n < N; n += 256) {
    if (mask && !mask[n]) continue;
    double v = resp[n * K + k];
    if (a >= 0) v *= c[n * D + a];
    if (b >= 0) v *= c[n * D + b];
    s += v;
  }
  __shared__ double s_red[256];
  s_red[threadIdx.x] = s;
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) {
    if (threadIdx.x < o) s_red[threadIdx.x] += s_red[threadIdx.x + o];
    __syncthreads();
  }
  if (threadIdx.x == 0) out[e] = s_red[0];
}

// ---------------------------------------------------------------------------
// N1: photon allocation (celeste_sample_sources.pyx:61-156)
// ---------------------------------------------------------------------------
// Counter-based generator (Philox-4x32, 10 rounds; Salmon et al. 2011): every
// pixel owns the stream keyed by (seed, image, pixel), so the draws do not
// depend on scheduling.  The reference consumes one MT19937 stream serially
// over pixels; that order cannot be reproduced in parallel, so parity for this
// row is distributional (tests: conservation, moments, chi-square vs the
// binomial pmf), as SURVEY.md 8(f) N1 anticipates.
struct Philox {
  uint32_t key[2], ctr[4];
  __device__ __forceinline__ void round_once(uint32_t k0, uint32_t k1) {
    const uint32_t hi0 = __umulhi(0xD2511F53u, ctr[0]), lo0 = 0xD2511F53u * ctr[0];
    const uint32_t hi1 = __umulhi(0xCD9E8D57u, ctr[2]), lo1 = 0xCD9E8D57u * ctr[2];
    const uint32_t c0 = hi1 ^ ctr[1] ^ k0, c2 = hi0 ^ ctr[3] ^ k1;
    ctr[0] = c0; ctr[1] = lo1; ctr[2] = c2; ctr[3] = lo0;
  }
  // uniform in (0, 1) with 53 random bits, from counter value (a, b, c, d)
  __device__ __noinline__ double uniform(uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
    ctr[0] = a; ctr[1] = b; ctr[2] = c; ctr[3] = d;
    uint32_t k0 = key[0], k1 = key[1];
#pragma unroll
    for (int r = 0; r < 10; ++r) {
      round_once(k0, k1);
      k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
    const unsigned long long bits = ((unsigned long long)(ctr[0] >> 5) << 26) | (ctr[1] >> 6);
    return ((double)bits + 0.5) * (1.0 / 9007199254740992.0);
  }
};

// One binomial(n, p) variate from ONE uniform by inversion (exact up to
// float64 rounding of the pmf recurrences): from zero when n*min(p,q) is
// small, else outwards from the mode (Kemp 1986), whose expected number of
// steps is O(sqrt(n p q)).  deps/randomkit/distributions.c:239-449 switches
// between inversion and BTPE rejection; both sample the same law.
// (__noinline__: the allocation kernels draw at 8 unrolled pixel sites per lane; inlined, the body --
// exp, log1p, three lgamma -- is replicated at every site and the kernel stalls on instruction fetch:
// ncu showed 24.8 warps waiting on `no_instruction` per issue before, profiles/r02_prof_alloc_*.)
__device__ __noinline__ double binomial_from_uniform(long long n, double p, double U) {
  if (n <= 0 || !(p > 0.)) return 0.;
  if (p >= 1.) return (double)n;
  const bool flip = p > 0.5;
  const double pp = flip ? 1. - p : p, q = 1. - pp, r = pp / q;
  const double dn = (double)n;
  double X;
  if (dn * pp < 12.) {
    double px = exp(dn * log1p(-pp));
    X = 0.;
    while (U > px && X < dn) {
      U -= px;
      X += 1.;
      px *= (dn - X + 1.) * r / X;
      if (px <= 0.) break;
    }
  } else {
    const double m = floor((dn + 1.) * pp);
    const double fm = exp(lgamma(dn + 1.) - lgamma(m + 1.) - lgamma(dn - m + 1.) + m * log(pp) +
                          (dn - m) * log1p(-pp));
    X = m;
    U -= fm;
    if (U > 0.) {
      double lo = m, hi = m, flo = fm, fhi = fm;
      for (;;) {
        bool moved = false;
        if (hi < dn) {
          fhi *= (dn - hi) / (hi + 1.) * r;
          hi += 1.;
          U -= fhi;
          moved = true;
          if (U <= 0.) { X = hi; break; }
        }
        if (lo > 0.) {
          flo *= lo / (dn - lo + 1.) / r;
          lo -= 1.;
          U -= flo;
          moved = true;
          if (U <= 0.) { X = lo; break; }
        }
        if (!moved || (fhi <= 0. && flo <= 0.)) break;   // rounding residue: stay at the mode
      }
    }
  }
  return flip ? dn - X : X;
}

// instrumentation: count draws of binomial(n, p) for the sampler's unit test
__global__ void binomial_test_kernel(long long n, double p, unsigned long long seed, long long count,
                                     double* out) {
  long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= count) return;
  Philox g;
  g.key[0] = (uint32_t)seed; g.key[1] = (uint32_t)(seed >> 32);
  out[i] = binomial_from_uniform(n, p, g.uniform((uint32_t)i, (uint32_t)(i >> 32), 0u, 0x5eedu));
}

struct AllocArgs {
  ImgPtrs imgs;
  int n_img, n_src;
  const CgBox* boxes;      // [n_src * n_img], item = src * n_img + img
  const float* patches;    // rate patches, item-major, stride max_patch_px
  float* samp;             // sampled photon patches, same layout (zeroed by the caller)
  long long max_patch_px;
  int* tile_count;         // [n_img][max_tiles]
  int* tile_list;          // [n_img][max_tiles][TILE_LIST_CAP]
  double* block_partials;  // [n_img][blocks_per_img]
  double* noise_sum;       // [n_img]
  int max_tiles, blocks_per_img;
  unsigned long long seed;
};

// one warp per item: append the item to the list of every image tile its box overlaps
__global__ void __launch_bounds__(128) bin_boxes_kernel(AllocArgs a) {
  const int lane = threadIdx.x & 31;
  const long long i = (long long)blockIdx.x * 4 + (threadIdx.x >> 5);
  if (i >= (long long)a.n_src * a.n_img) return;
  const int img = (int)(i % a.n_img);
  const CgBox b = a.boxes[i];
  if (b.x1 <= b.x0 || b.y1 <= b.y0) return;
  const int W = a.imgs.p[img]->W;
  const int tiles_x = (W + TILE_W - 1) / TILE_W;
  const int txa = b.x0 / TILE_W, txb = (b.x1 - 1) / TILE_W;
  const int tya = b.y0 / TILE_H, tyb = (b.y1 - 1) / TILE_H;
  const int nx = txb - txa + 1, nt = nx * (tyb - tya + 1);
  for (int j = lane; j < nt; j += 32) {
    const int t = (tya + j / nx) * tiles_x + txa + (j % nx);
    const long long slot = (long long)img * a.max_tiles + t;
    const int pos = atomicAdd(&a.tile_count[slot], 1);
    if (pos < TILE_LIST_CAP) a.tile_list[slot * TILE_LIST_CAP + pos] = (int)i;
  }
}

// One warp per half image tile (64 x 8 pixels; lane = 4 pixels x 4 rows).
// For every pixel: the sources whose box strictly contains it
// (celeste_sample_sources.pyx:53-54: x > x0 and x < x1 and y > y0 and y < y1,
// so the first row and column of a patch never receive photons), then the
// multinomial split of int(nelec) over [rates..., epsilon] by conditional
// binomials in source order (:127-156); the remainder is sky noise (:112).
__global__ void __launch_bounds__(FR_WARPS * 32) sample_counts_kernel(AllocArgs a) {
  const int img = blockIdx.y;
  const DevImage& im = *a.imgs.p[img];
  const int H = im.H, W = im.W;
  const int tiles_x = (W + TILE_W - 1) / TILE_W, tiles_y = (H + TILE_H - 1) / TILE_H;
  const int n_tiles = tiles_x * tiles_y;
  const int n_blocks = (2 * n_tiles + FR_WARPS - 1) / FR_WARPS;
  if ((int)blockIdx.x >= n_blocks) return;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  __shared__ int s_item[FR_WARPS][TILE_LIST_CAP];
  __shared__ int4 s_box[FR_WARPS][TILE_LIST_CAP];
  __shared__ double s_red[FR_WARPS];

  const int ht = blockIdx.x * FR_WARPS + warp;
  const int tile = ht >> 1;
  double noise = 0.0;
  if (tile < n_tiles) {
    const int tx = tile % tiles_x, ty = tile / tiles_x;
    const int X0 = tx * TILE_W, Y0 = ty * TILE_H + (ht & 1) * (TILE_H / 2);
    const int x = X0 + (lane & 15) * 4, ybase = Y0 + (lane >> 4) * 4;
    const long long tslot = (long long)img * a.max_tiles + tile;
    const int cnt_raw = a.tile_count[tslot];
    int cnt = 0;
    // ordered (by item = source order) list of the boxes over this half tile
    if (cnt_raw <= TILE_LIST_CAP) {
      int it0 = 0x7fffffff, it1 = 0x7fffffff;
      if (lane < cnt_raw) it0 = a.tile_list[tslot * TILE_LIST_CAP + lane];
      if (lane + 32 < cnt_raw) it1 = a.tile_list[tslot * TILE_LIST_CAP + lane + 32];
      int r0 = 0, r1 = 0;
      for (int j = 0; j < min(cnt_raw, 32); ++j) {
        const int v = __shfl_sync(0xffffffffu, it0, j);
        r0 += (v < it0); r1 += (v < it1);
      }
      for (int j = 0; j < cnt_raw - 32; ++j) {
        const int v = __shfl_sync(0xffffffffu, it1, j);
        r0 += (v < it0); r1 += (v < it1);
      }
      if (lane < cnt_raw) {
        s_item[warp][r0] = it0;
        const CgBox b = a.boxes[it0];
        s_box[warp][r0] = make_int4(b.x0, b.x1, b.y0, b.y1);
      }
      if (lane + 32 < cnt_raw) {
        s_item[warp][r1] = it1;
        const CgBox b = a.boxes[it1];
        s_box[warp][r1] = make_int4(b.x0, b.x1, b.y0, b.y1);
      }
      cnt = cnt_raw;
      __syncwarp();
    }
    const float eps = (float)im.epsilon;
    Philox g;
    g.key[0] = (uint32_t)a.seed; g.key[1] = (uint32_t)(a.seed >> 32);

    for (int r = 0; r < 4; ++r) {
      const int y = ybase + r;
      if (y >= H) break;
      for (int p = 0; p < 4; ++p) {
        const int xx = x + p;
        if (xx >= W) break;
        const double nph = (double)im.nelec[(long long)y * W + xx];
        // pass 1: total rate of the sources that own this pixel
        double sum_probs = 0.;
        int owners = 0;
        if (cnt_raw <= TILE_LIST_CAP) {
          for (int j = 0; j < cnt; ++j) {
            const int4 b = s_box[warp][j];
            if (xx > b.x && xx < b.y && y > b.z && y < b.w) {
              sum_probs += (double)a.patches[(long long)s_item[warp][j] * a.max_patch_px +
                                             (long long)(y - b.z) * (b.y - b.x) + (xx - b.x)];
              ++owners;
            }
          }
        } else {
          for (int s = 0; s < a.n_src; ++s) {
            const long long item = (long long)s * a.n_img + img;
            const CgBox b = a.boxes[item];
            if (xx > b.x0 && xx < b.x1 && y > b.y0 && y < b.y1) {
              sum_probs += (double)a.patches[item * a.max_patch_px + (long long)(y - b.y0) * (b.x1 - b.x0) +
                                             (xx - b.x0)];
              ++owners;
            }
          }
        }
        if (owners == 0) { noise += nph; continue; }     // :94-96
        sum_probs += (double)eps;
        long long unsampled = (long long)nph;            // int(num_photons_xy), :110
        const uint32_t pix = (uint32_t)((long long)y * W + xx);
        int draw = 0;
        // pass 2: conditional binomials in source order
        const int n_scan = (cnt_raw <= TILE_LIST_CAP) ? cnt : a.n_src;
        for (int j = 0; j < n_scan && unsampled > 0; ++j) {
          long long item;
          int4 b;
          if (cnt_raw <= TILE_LIST_CAP) {
            item = s_item[warp][j];
            b = s_box[warp][j];
          } else {
            item = (long long)j * a.n_img + img;
            const CgBox bb = a.boxes[item];
            b = make_int4(bb.x0, bb.x1, bb.y0, bb.y1);
          }
          if (!(xx > b.x && xx < b.y && y > b.z && y < b.w)) continue;
          const long long off = item * a.max_patch_px + (long long)(y - b.z) * (b.y - b.x) + (xx - b.x);
          const double rate = (double)a.patches[off];
          const double U = g.uniform(pix, (uint32_t)img, (uint32_t)draw++, 0xa110cu);
          const double got = binomial_from_uniform(unsampled, rate / sum_probs, U);
          a.samp[off] = (float)got;
          unsampled -= (long long)got;
          sum_probs -= rate;
        }
        noise += (double)unsampled;                      // samples[S-1] += unsampled, :155
      }
    }
  }
  noise = warp_sum(noise);
  if (lane == 0) s_red[warp] = noise;
  __syncthreads();
  if (tid == 0) {
    double t = 0;
#pragma unroll
    for (int wv = 0; wv < FR_WARPS; ++wv) t += s_red[wv];
    a.block_partials[(long long)img * a.blocks_per_img + blockIdx.x] = t;
  }
}

// per-image sum of the block partials (fixed order)
__global__ void __launch_bounds__(256) alloc_final_kernel(AllocArgs a) {
  const int img = blockIdx.x;
  const DevImage& im = *a.imgs.p[img];
  const int n_tiles = ((im.W + TILE_W - 1) / TILE_W) * ((im.H + TILE_H - 1) / TILE_H);
  const int n_blocks = (2 * n_tiles + FR_WARPS - 1) / FR_WARPS;
  __shared__ double s_fin[256];
  double t = 0;
  const double* bp = a.block_partials + (long long)img * a.blocks_per_img;
  for (int i = threadIdx.x; i < n_blocks; i += 256) t += bp[i];
  s_fin[threadIdx.x] = t;
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) {
    if (threadIdx.x < o) s_fin[threadIdx.x] += s_fin[threadIdx.x + o];
    __syncthreads();
  }
  if (threadIdx.x == 0) a.noise_sum[img] = s_fin[0];
}

// ---------------------------------------------------------------------------
// XU (MUFU.EX2) throughput probe: the roofline denominator of render_kernel,
// measured on the device it runs on.  8 independent ex2 chains per thread.
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(256) xu_probe_kernel(float* out, int iters) {
  float v[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) v[j] = -1.0f - 0.001f * (float)(threadIdx.x + j);
  for (int i = 0; i < iters; ++i) {
#pragma unroll
    for (int j = 0; j < 8; ++j) v[j] = ex2_approx(v[j]) - 1.5f;
  }
  float s = 0.f;
#pragma unroll
  for (int j = 0; j < 8; ++j) s += v[j];
  if (s == 123.456f) out[blockIdx.x * blockDim.x + threadIdx.x] = s;   // never true; keeps the chains live
}

// FP32 FMA throughput probe (secondary pipe of render_kernel).
__global__ void __launch_bounds__(256) fma_probe_kernel(float* out, int iters) {
  float v[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) v[j] = 0.5f + 0.001f * (float)(threadIdx.x + j);
  const float a = 0.999f, b = 0.0007f;
  for (int i = 0; i < iters; ++i) {
#pragma unroll
    for (int j = 0; j < 8; ++j) v[j] = fmaf(v[j], a, b);
  }
  float s = 0.f;
#pragma unroll
  for (int j = 0; j < 8; ++j) s += v[j];
  if (s == 123.456f) out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// which = 7: packed FFMA2 (fma.rn.f32x2), 8 independent pair chains per thread.  Result: scalar
// FMA lane-operations/s (2 per FFMA2 lane-instruction) -- tells whether the packed form adds
// FP32 throughput or only saves issue slots.
__global__ void __launch_bounds__(256) fma2_probe_kernel(float* out, int iters) {
  f32x2 v[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) v[j] = pk2(0.5f + 0.001f * (float)(threadIdx.x + j), 0.25f + 0.001f * (float)(threadIdx.x + j));
  const f32x2 a = pk2(0.999f, 0.998f), b = pk2(0.0007f, 0.0005f);
  for (int i = 0; i < iters; ++i) {
#pragma unroll
    for (int j = 0; j < 8; ++j) v[j] = fma2(v[j], a, b);
  }
  float s = 0.f;
#pragma unroll
  for (int j = 0; j < 8; ++j) { float lo, hi; unpk2(v[j], lo, hi); s += lo + hi; }
  if (s == 123.456f) out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// which = 8 / 9: packed FMUL2 / FADD2 chains, counted like which = 7
template <int OP>
__global__ void __launch_bounds__(256) pk2_probe_kernel(float* out, int iters) {
  f32x2 v[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) v[j] = pk2(0.5f + 0.001f * (float)(threadIdx.x + j), 0.25f + 0.001f * (float)(threadIdx.x + j));
  const f32x2 a = OP == 0 ? pk2(0.9999f, 1.0001f) : pk2(1e-7f, -1e-7f);
  for (int i = 0; i < iters; ++i) {
#pragma unroll
    for (int j = 0; j < 8; ++j) v[j] = OP == 0 ? mul2_raw(v[j], a) : add2_raw(v[j], a);
  }
  float s = 0.f;
#pragma unroll
  for (int j = 0; j < 8; ++j) { float lo, hi; unpk2(v[j], lo, hi); s += lo + hi; }
  if (s == 123.456f) out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// which = 20 / 21 / 22: does a packed FP32x2 instruction hold the issue port for the two cycles it holds
// the FMA pipe?  Per group: 2 FFMA2 (20), 2 FFMA (21) or 4 FFMA (22: the arithmetic of 20 unpacked), each
// with ONE independent integer LOP3 (ALU pipe).  FMA-pipe cycles per group 4 / 2 / 4, issue slots 3 / 3 / 5
// if a packed instruction takes one slot, 5 / 3 / 5 if it takes two.  Result: groups per second per lane
// (scheduler clock / (result / 32 lanes / schedulers) = cycles per group).
template <int VAR>
__global__ void __launch_bounds__(256) issue_probe_kernel(float* out, int iters) {
  f32x2 v[16];
  float w[32];
  unsigned n[8];
#pragma unroll
  for (int j = 0; j < 16; ++j) v[j] = pk2(0.5f + 0.001f * (float)(threadIdx.x + j), 0.25f + 0.001f * (float)(threadIdx.x + j));
#pragma unroll
  for (int j = 0; j < 32; ++j) w[j] = 0.5f + 0.001f * (float)(threadIdx.x + j);
#pragma unroll
  for (int j = 0; j < 8; ++j) n[j] = threadIdx.x * 2654435761u + j;
  const f32x2 a = pk2(0.999f, 0.998f), b = pk2(0.0007f, 0.0005f);
  const float fa = 0.999f, fb = 0.0007f;
  const unsigned m = out ? 0x9e3779b9u : 0u;
  for (int i = 0; i < iters; ++i) {
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      // volatile: exactly these instructions, in these proportions
      if (VAR == 0) {
        asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(v[j]) : "l"(a), "l"(b));
        asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(v[j + 8]) : "l"(a), "l"(b));
      } else {
        asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(w[j]) : "f"(fa), "f"(fb));
        asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(w[j + 8]) : "f"(fa), "f"(fb));
        if (VAR == 2) {
          asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(w[j + 16]) : "f"(fa), "f"(fb));
          asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(w[j + 24]) : "f"(fa), "f"(fb));
        }
      }
      asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(n[j]) : "r"(m), "r"(i));   // n ^= m ^ i: not foldable across trips
    }
  }
  float s = 0.f;
#pragma unroll
  for (int j = 0; j < 16; ++j) { float lo, hi; unpk2(v[j], lo, hi); s += lo + hi + w[j] + w[j + 16]; }
#pragma unroll
  for (int j = 0; j < 8; ++j) s += (float)n[j];
  if (s == 123.456f) out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// Instruction-mix probes: the inner loop of field_tile_kernel / render_kernel in
// isolation (no global memory, every lane busy), to measure what fraction of
// the MUFU peak this MIX can reach on the device.  VAR 2: the full loop (table
// entry from shared memory, 6 row-term FFMA, then 2 FFMA + MUFU.EX2 + FADD per
// pixel, 8 pixels); 3: the same with the table in registers (no LDS); 4: no row
// terms either; 5: 1 FFMA + MUFU + FADD per pixel.  Result: MUFU lane-ops/s.
template <int VAR>
__global__ void __launch_bounds__(256) mix_probe_kernel(float* out, int iters) {
  __shared__ __align__(16) float s_tab[48 * 8];
  for (int i = threadIdx.x; i < 48 * 8; i += 256) s_tab[i] = -1e-3f * (float)(1 + (i % 7));
  __syncthreads();
  const float* tb0 = s_tab + (threadIdx.x >> 5) * 8;
  float acc0[4] = {0.f, 0.f, 0.f, 0.f}, acc1[4] = {0.f, 0.f, 0.f, 0.f};
  float xl[4];
#pragma unroll
  for (int p = 0; p < 4; ++p) xl[p] = 0.25f * (float)((threadIdx.x & 3) * 4 + p);
  const float y0l = 0.125f * (float)((threadIdx.x & 31) >> 2), y1l = y0l + 1.f;
  float4 t0 = *reinterpret_cast<const float4*>(tb0);
  float2 t1 = *reinterpret_cast<const float2*>(tb0 + 4);
  for (int i = 0; i < iters; ++i) {
    const float* tb = tb0;
#pragma unroll 2
    for (int k = 0; k < 32; ++k) {
      float b1a, b1b, c1a, c1b;
      if (VAR == 2) {
        tb += 8;
        const float4 n0 = *reinterpret_cast<const float4*>(tb);
        const float2 n1 = *reinterpret_cast<const float2*>(tb + 4);
        b1a = fmaf(t0.y, y0l, t0.w); b1b = fmaf(t0.y, y1l, t0.w);
        c1a = fmaf(fmaf(t0.z, y0l, t1.x), y0l, t1.y);
        c1b = fmaf(fmaf(t0.z, y1l, t1.x), y1l, t1.y);
        float q[8];
#pragma unroll
        for (int p = 0; p < 4; ++p) {
          q[p] = fmaf(fmaf(t0.x, xl[p], b1a), xl[p], c1a);
          q[4 + p] = fmaf(fmaf(t0.x, xl[p], b1b), xl[p], c1b);
        }
#pragma unroll
        for (int p = 0; p < 8; ++p) q[p] = ex2_approx(q[p]);
#pragma unroll
        for (int p = 0; p < 4; ++p) { acc0[p] += q[p]; acc1[p] += q[4 + p]; }
        t0 = n0; t1 = n1;
      } else {
        if (VAR == 3) {
          b1a = fmaf(t0.y, y0l, t0.w); b1b = fmaf(t0.y, y1l, t0.w);
          c1a = fmaf(fmaf(t0.z, y0l, t1.x), y0l, t1.y);
          c1b = fmaf(fmaf(t0.z, y1l, t1.x), y1l, t1.y);
        } else {
          b1a = t0.y; b1b = t0.w; c1a = t1.x; c1b = t1.y;
        }
        float q[8];
#pragma unroll
        for (int p = 0; p < 4; ++p) {
          if (VAR == 5) {
            q[p] = fmaf(t0.x, xl[p], c1a + acc0[p] * 1e-30f);
            q[4 + p] = fmaf(t0.x, xl[p], c1b + acc1[p] * 1e-30f);
          } else {
            q[p] = fmaf(fmaf(t0.x, xl[p], b1a), xl[p], c1a);
            q[4 + p] = fmaf(fmaf(t0.x, xl[p], b1b), xl[p], c1b);
          }
        }
#pragma unroll
        for (int p = 0; p < 8; ++p) q[p] = ex2_approx(q[p]);
#pragma unroll
        for (int p = 0; p < 4; ++p) { acc0[p] += q[p]; acc1[p] += q[4 + p]; }
        t0.y += 1e-7f;          // keep the row terms loop-variant
      }
    }
  }
  float sacc = 0.f;
#pragma unroll
  for (int p = 0; p < 4; ++p) sacc += acc0[p] + acc1[p];
  if (sacc == 123.456f) out[blockIdx.x * blockDim.x + threadIdx.x] = sacc;
}

// which = 6 (16 at 8 CTAs/SM): the row-recurrence loop of the field kernel (ft_eval2r) on a
// 32-component shared-memory table.  Result: component x pixel evaluations/s.
__global__ void __launch_bounds__(256) recur_probe_kernel(float* out, int iters) {
  __shared__ __align__(16) float s_tab[8][33 * 8];
  float* tb = s_tab[threadIdx.x >> 5];
  for (int i = threadIdx.x & 31; i < 33 * 8; i += 32) {
    const int c = i & 7;
    tb[i] = (c == 6) ? 0.998f : (c == 7) ? 1.001f : -1e-3f * (float)(1 + (i % 5));
  }
  __syncwarp();
  float acc0[4] = {0.f, 0.f, 0.f, 0.f}, acc1[4] = {0.f, 0.f, 0.f, 0.f};
  float x0 = 0.25f * (float)((threadIdx.x & 3) * 4);
  const float y0l = 0.125f * (float)((threadIdx.x & 31) >> 2), y1l = y0l + 1.f;
  for (int i = 0; i < iters; ++i) {
    ft_eval2r(tb, 32, y0l, y1l, x0, acc0, acc1);
    x0 += 1e-7f;            // keep the loop body loop-variant
  }
  float sacc = 0.f;
#pragma unroll
  for (int p = 0; p < 4; ++p) sacc += acc0[p] + acc1[p];
  if (sacc == 123.456f) out[blockIdx.x * blockDim.x + threadIdx.x] = sacc;
}

// per-launch CUDA-event timing of render_kernel (bench.py's roofline leg)
#define PROF_MAX 8192
static bool g_prof_on = false;
static std::vector<cudaEvent_t> g_prof_ev;   // 2 per launch
static int g_prof_n = 0;

// ===========================================================================
// host side
// ===========================================================================
static inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

struct WsLayout {
  size_t counters, hdr, tab, worklist, partials, item_ll, img_boxes, aux_offsets, tiles;
  size_t tile_count, done, tile_list, model_ptrs;   // full-field gather (tile_count and done are contiguous)
  size_t tile_zero_bytes, order;
  // second pass (float64 re-score)
  size_t counters2, hdr2, tab2, worklist2, partials2, item_ll2, guard_list;
  size_t total;
  int worklist_cap, worklist_cap2, guard_cap;
  long long n_items2;
};

#define MAX_TILES_PER_IMG_DEFAULT 0

static WsLayout ws_layout(long long n_items, long long max_patch_px, int with_patches,
                          int n_img_for_guard, long long n_tiles_total, int guard_cap = 0) {
  WsLayout L;
  memset(&L, 0, sizeof L);
  size_t o = 0;
  auto take = [&](size_t bytes) { size_t r = o; o = align_up(o + bytes, 256); return r; };
  long long strips_per_item = (max_patch_px + WT_PIX - 1) / WT_PIX + 64;   // warp tiles
  // rows whose width is not a multiple of 8 waste chunk slots: bound chunks by px/8 + rows
  long long cap = n_items * strips_per_item;
  if (cap > 0x7fffffffLL) cap = 0x7fffffffLL;
  L.worklist_cap = (int)cap;
  L.counters = take(sizeof(Counters));
  L.hdr = take(sizeof(ItemHdr) * n_items);
  L.tab = take(sizeof(float) * TAB_STRIDE * n_items);
  L.worklist = take(sizeof(int2) * (size_t)L.worklist_cap);
  L.partials = take(sizeof(double) * (size_t)L.worklist_cap);
  L.item_ll = take(sizeof(double) * n_items);
  L.img_boxes = take(sizeof(CgBox) * CG_MAX_IMAGES);
  L.aux_offsets = take(sizeof(long long) * CG_MAX_IMAGES);
  L.tiles = take(sizeof(double) * (size_t)n_tiles_total * (with_patches ? BINS_PER_TILE : 1));
  if (with_patches) {
    L.tile_count = take(sizeof(int) * (size_t)n_tiles_total * BINS_PER_TILE);
    L.done = take(sizeof(int) * CG_MAX_IMAGES);
    L.tile_zero_bytes = (L.done + sizeof(int) * CG_MAX_IMAGES) - L.tile_count;
    L.tile_list = take(sizeof(int) * (size_t)n_tiles_total * BINS_PER_TILE * BIN_CAP);
    L.order = take(sizeof(int2) * (size_t)n_tiles_total * BINS_PER_TILE * FT_BUCKETS);
    L.model_ptrs = take(sizeof(float*) * CG_MAX_IMAGES);
  }
  if (n_img_for_guard > 0) {
    // every proposal of the call may fall into the guard band: the float64 re-score list holds them all
    L.guard_cap = std::max(guard_cap, 1);
    L.n_items2 = (long long)(L.guard_cap + 1) * n_img_for_guard;
    long long cap2 = L.n_items2 * strips_per_item;
    if (cap2 > 0x7fffffffLL) cap2 = 0x7fffffffLL;
    L.worklist_cap2 = (int)cap2;
    L.counters2 = take(sizeof(Counters));
    L.hdr2 = take(sizeof(ItemHdr) * L.n_items2);
    L.tab2 = take(sizeof(double) * TAB_STRIDE * L.n_items2);
    L.worklist2 = take(sizeof(int2) * (size_t)L.worklist_cap2);
    L.partials2 = take(sizeof(double) * (size_t)L.worklist_cap2);
    L.item_ll2 = take(sizeof(double) * L.n_items2);
    L.guard_list = take(sizeof(int) * (size_t)L.guard_cap);
  }
  L.total = o;
  return L;
}

static Counters* g_last_counters = nullptr;

// Scratch of the small entry points (A2's component table, the Poisson block partials): one
// block PER STREAM, so that calls enqueued on different streams never share a buffer
// ("thread-compatible, one stream per thread").  Grown blocks are retired, not
// freed: work already enqueued may still read them.
struct StreamScratch { double* p = nullptr; size_t n = 0; };
static std::mutex g_scratch_mu;
static std::map<std::pair<int, cudaStream_t>, StreamScratch> g_scratch_map;
static int stream_scratch(int kind, cudaStream_t st, size_t n_doubles, double** out) {
  std::lock_guard<std::mutex> lk(g_scratch_mu);
  StreamScratch& sc = g_scratch_map[std::make_pair(kind, st)];
  if (sc.n < n_doubles) {
    double* p = nullptr;
    const size_t n = std::max<size_t>(n_doubles, 1024);
    CUDA_TRY(cudaMalloc(&p, sizeof(double) * n));
    sc.p = p; sc.n = n;          // the previous (smaller) block stays allocated for in-flight work
  }
  *out = sc.p;
  return CG_OK;
}

// ---- scratch of cg_poisson_loglike_static: no allocation at call time -------------------------
// The block partials of the Poisson reduction live in a file-scope device array (8 KB that every GPU
// which loads this library holds until the process exits).  In HOST code the name g_static_partials is
// a host-side stand-in, not a device address: its device address is taken once per device at set-up
// (cg_init, with that device current) by cudaGetSymbolAddress and only that pointer is handed to kernels.
static __device__ __align__(256) double g_static_partials[CG_POISSON_STATIC_MAX_BLOCKS];
#define STATIC_MAX_DEVICES 64
static double* g_static_partials_ptr[STATIC_MAX_DEVICES] = {nullptr};
static int static_scratch_setup(int device) {
  if (device < 0 || device >= STATIC_MAX_DEVICES) return CG_OK;   // cg_poisson_loglike_static reports it
  if (g_static_partials_ptr[device]) return CG_OK;
  double* p = nullptr;
  CUDA_TRY(cudaGetSymbolAddress((void**)&p, g_static_partials));
  g_static_partials_ptr[device] = p;
  return CG_OK;
}

extern "C" {

int cg_abi_version(void) { return CG_ABI_VERSION; }

int cg_last_error(char* buf, size_t n) {
  if (buf && n) {
    strncpy(buf, g_err.c_str(), n - 1);
    buf[n - 1] = 0;
  }
  return (int)g_err.size();
}

int cg_init(int device) {
  int count = 0;
  cudaError_t e = cudaGetDeviceCount(&count);
  if (e != cudaSuccess || count == 0)
    return fail(CG_ECUDA, "no CUDA device available (%s); this library has no CPU fallback",
                cudaGetErrorString(e));
  if (device < 0 || device >= count) return fail(CG_EINVAL, "device %d out of range [0,%d)", device, count);
  CUDA_TRY(cudaSetDevice(device));
  cudaDeviceProp prop;
  CUDA_TRY(cudaGetDeviceProperties(&prop, device));
  g_sm_count = prop.multiProcessorCount;
  ProfileConsts pc;
  double se = 0, sd = 0;
  for (int i = 0; i < 6; ++i) se += H_EXP_AMP_RAW[i];
  for (int i = 0; i < 8; ++i) sd += H_DEV_AMP_RAW[i];
  for (int i = 0; i < 6; ++i) { pc.exp_amp[i] = H_EXP_AMP_RAW[i] / se; pc.exp_var[i] = H_EXP_VAR[i]; }
  for (int i = 0; i < 8; ++i) { pc.dev_amp[i] = H_DEV_AMP_RAW[i] / sd; pc.dev_var[i] = H_DEV_VAR[i]; }
  CUDA_TRY(cudaMemcpyToSymbol(c_prof, &pc, sizeof pc));
  if (!g_d_logtab) {
    CUDA_TRY(cudaMalloc(&g_d_logtab, sizeof(LogTable)));
    std::vector<double> lt(8192);
    for (int idx = 0; idx < 8192; ++idx) {
      const long double c = 1.0L + (2.0L * (idx & 31) + 1.0L) / 64.0L;
      lt[idx] = (double)((long double)((idx >> 5) - 127) * 0.693147180559945309417232121458L + logl(c));
    }
    CUDA_TRY(cudaMemcpy(g_d_logtab, lt.data(), sizeof(LogTable), cudaMemcpyHostToDevice));
  }
  if (!g_d_filters) {
    CUDA_TRY(cudaMalloc(&g_d_filters, sizeof(FilterTables)));
    CUDA_TRY(cudaMemset(g_d_filters, 0, sizeof(FilterTables)));
  }
  { int rc = static_scratch_setup(device); if (rc) return rc; }   // device address of g_static_partials
  g_device = device;
  return CG_OK;
}

int cg_shutdown(void) {
  if (g_d_filters) { cudaFree(g_d_filters); g_d_filters = nullptr; }
  g_filters_loaded = 0;
  g_device = -1;
  return CG_OK;
}

int cg_sm_count(void) { return g_sm_count; }

int64_t cg_launch_count(int reset) {
  long long v = g_launches.load();
  if (reset) g_launches.store(0);
  return v;
}

int64_t cg_last_issued_count(void* stream) {
  if (!g_last_counters) return 0;
  Counters c;
  if (cudaStreamSynchronize((cudaStream_t)stream) != cudaSuccess) return -1;
  if (cudaMemcpy(&c, g_last_counters, sizeof c, cudaMemcpyDeviceToHost) != cudaSuccess) return -1;
  return (int64_t)c.issued;
}

int64_t cg_last_eval_count(void* stream) {
  if (!g_last_counters) return 0;
  Counters c;
  if (cudaStreamSynchronize((cudaStream_t)stream) != cudaSuccess) return -1;
  if (cudaMemcpy(&c, g_last_counters, sizeof c, cudaMemcpyDeviceToHost) != cudaSuccess) return -1;
  if (c.err_flag) {
    fail(CG_ENOMEM, c.err_flag == 1 ? "a footprint exceeded max_patch_px"
                                    : (c.err_flag == 2 ? "guard list overflow"
                                    : (c.err_flag == 3 ? "work list overflow"
                                    : (c.err_flag == 4 ? "a CG_SRC_EIGEN source points outside the bound flux table"
                                                       : "a CG_SRC_PLANCK star was rendered before cg_set_filter_curves"))));
    return -(int64_t)c.err_flag;
  }
  return (int64_t)c.evals;
}

// ---- images ----------------------------------------------------------------
double cg_calc_bounding_radius(const double* means, const double* covars, int32_t K, double error,
                               const double* center) {
  // util/bound/bounding_box.py:9-31 ; chi2.ppf(1-error, 2) = -2 log(error)
  double Rsq;
  if (error == 0.001) Rsq = 13.815510557964274;
  else if (error == 1e-5) Rsq = 23.025850929940457;
  else Rsq = -2.0 * log(error);
  double cx = center ? center[0] : 0, cy = center ? center[1] : 0;
  double best = -INFINITY;
  for (int i = 0; i < K; ++i) {
    double maj = major_axis(covars[i * 4 + 0], covars[i * 4 + 3], covars[i * 4 + 1], Rsq);
    double dx = means[i * 2] - cx, dy = means[i * 2 + 1] - cy;
    double v = maj + sqrt(dx * dx + dy * dy);
    if (v > best) best = v;
  }
  return best;
}

int cg_image_create(const CgImageDesc* d, CgImage** out) {
  NEED_INIT();
  if (!d || !out) return fail(CG_EINVAL, "null argument");
  if (d->n_psf < 1 || d->n_psf > CG_MAX_PSF) return fail(CG_EINVAL, "n_psf must be in [1,%d]", CG_MAX_PSF);
  if (d->H <= 0 || d->W <= 0) return fail(CG_EINVAL, "bad image shape %dx%d", d->H, d->W);
  if (d->band < 0 || d->band >= CG_NUM_BANDS) return fail(CG_EINVAL, "band must be 0..4");
  CgImage* im = new CgImage();
  DevImage& h = im->h;
  memset(&h, 0, sizeof h);
  h.H = d->H; h.W = d->W; h.band = d->band; h.n_psf = d->n_psf;
  h.epsilon = d->epsilon; h.kappa = d->kappa; h.calib = d->calib;
  double means[CG_MAX_PSF * 2], covs[CG_MAX_PSF * 4];
  h.psf_wsum = 0;
  for (int k = 0; k < d->n_psf; ++k) {
    h.psf_w[k] = d->psf_w[k];
    h.psf_wsum += d->psf_w[k];
    for (int j = 0; j < 2; ++j) { h.psf_mu[k][j] = d->psf_mu[k][j]; means[k * 2 + j] = d->psf_mu[k][j]; }
    for (int j = 0; j < 3; ++j) h.psf_cov[k][j] = d->psf_cov[k][j];
    covs[k * 4 + 0] = d->psf_cov[k][0]; covs[k * 4 + 3] = d->psf_cov[k][1];
    covs[k * 4 + 1] = covs[k * 4 + 2] = d->psf_cov[k][2];
  }
  for (int j = 0; j < 2; ++j) { h.rho_n[j] = d->rho_n[j]; h.phi_n[j] = d->phi_n[j]; }
  for (int j = 0; j < 4; ++j) h.ups[j] = d->ups[j];
  double det = d->ups[0] * d->ups[3] - d->ups[1] * d->ups[2];
  if (det == 0) { delete im; return fail(CG_EINVAL, "singular CD matrix"); }
  h.ups_inv[0] = d->ups[3] / det; h.ups_inv[1] = -d->ups[1] / det;
  h.ups_inv[2] = -d->ups[2] / det; h.ups_inv[3] = d->ups[0] / det;
  h.R = cg_calc_bounding_radius(means, covs, d->n_psf, 0.001, nullptr);
  h.nelec = d->d_nelec; h.mask = d->d_mask;
  cudaError_t e = cudaMalloc(&im->d, sizeof(DevImage));
  if (e != cudaSuccess) { delete im; return fail(CG_ECUDA, "cudaMalloc: %s", cudaGetErrorString(e)); }
  e = cudaMemcpy(im->d, &h, sizeof h, cudaMemcpyHostToDevice);
  if (e != cudaSuccess) { cudaFree(im->d); delete im; return fail(CG_ECUDA, "cudaMemcpy: %s", cudaGetErrorString(e)); }
  *out = im;
  return CG_OK;
}

int cg_image_destroy(CgImage* img) {
  if (!img) return CG_OK;
  if (img->d) cudaFree(img->d);
  delete img;
  return CG_OK;
}

int cg_image_set_epsilon(CgImage* img, double epsilon) {
  NEED_INIT();
  if (!img) return fail(CG_EINVAL, "null image");
  img->h.epsilon = epsilon;
  CUDA_TRY(cudaMemcpy(img->d, &img->h, sizeof(DevImage), cudaMemcpyHostToDevice));
  return CG_OK;
}

double cg_image_radius(const CgImage* img) { return img ? img->h.R : NAN; }

int cg_set_flux_table(const double* d_flux, int64_t n_rows) {
  NEED_INIT();
  if (d_flux && n_rows <= 0) return fail(CG_EINVAL, "flux table needs at least one row");
  g_flux_table = d_flux;
  g_flux_rows = d_flux ? n_rows : 0;
  return CG_OK;
}

int cg_set_filter_curves(const double* wl, const double* hv, const double* sens, const int32_t* counts) {
  NEED_INIT();
  FilterTables* f = new FilterTables();
  memset(f, 0, sizeof *f);
  int o = 0;
  for (int b = 0; b < CG_NUM_BANDS; ++b) {
    if (counts[b] < 2 || o + counts[b] > MAX_FILTER_ROWS) { delete f; return fail(CG_EINVAL, "bad filter table sizes"); }
    f->count[b] = counts[b];
    f->start[b] = o;
    o += counts[b];
  }
  memcpy(f->wl, wl, sizeof(double) * o);
  memcpy(f->hvec, hv, sizeof(double) * o);
  memcpy(f->sens, sens, sizeof(double) * o);
  f->loaded = 1;
  cudaError_t e = cudaMemcpy(g_d_filters, f, sizeof *f, cudaMemcpyHostToDevice);
  delete f;
  if (e != cudaSuccess) return fail(CG_ECUDA, "cudaMemcpy: %s", cudaGetErrorString(e));
  g_filters_loaded = 1;
  return CG_OK;
}

// ---- A1 (host) ---------------------------------------------------------------
int cg_galaxy_prof_psf_mixture_params(const double* W, const double* v_s, const double* image_ws,
                                      const double* image_means, const double* image_covars,
                                      int32_t K_psf, const double* prof_amp, const double* prof_sigs,
                                      int32_t K_prof, double* weights, double* means, double* covars) {
  // celeste_fast.pyx:123-139 : PSF-major, c = k*K_prof + j
  int c = 0;
  for (int k = 0; k < K_psf; ++k)
    for (int j = 0; j < K_prof; ++j, ++c) {
      weights[c] = image_ws[k] * prof_amp[j];
      means[c * 2 + 0] = v_s[0] + image_means[k * 2 + 0];
      means[c * 2 + 1] = v_s[1] + image_means[k * 2 + 1];
      for (int e = 0; e < 4; ++e) covars[c * 4 + e] = image_covars[k * 4 + e] + prof_sigs[j] * W[e];
    }
  return CG_OK;
}

int cg_galaxy_psf_mixture_params(const double* thetas, const double* W, const double* v_s,
                                 const double* image_ws, const double* image_means,
                                 const double* image_covars, int32_t K_psf, const double* exp_amp,
                                 const double* exp_sigs, int32_t K_exp, const double* dev_amp,
                                 const double* dev_sigs, int32_t K_dev, double* weights,
                                 double* means, double* covars) {
  // celeste_fast.pyx:62-93
  int c = 0;
  for (int k = 0; k < K_psf; ++k)
    for (int i = 0; i < 2; ++i) {
      const double* amp = i == 0 ? exp_amp : dev_amp;
      const double* var = i == 0 ? exp_sigs : dev_sigs;
      int Ki = i == 0 ? K_exp : K_dev;
      for (int j = 0; j < Ki; ++j, ++c) {
        weights[c] = image_ws[k] * thetas[i] * amp[j];
        means[c * 2 + 0] = v_s[0] + image_means[k * 2 + 0];
        means[c * 2 + 1] = v_s[1] + image_means[k * 2 + 1];
        for (int e = 0; e < 4; ++e) covars[c * 4 + e] = image_covars[k * 4 + e] + var[j] * W[e];
      }
    }
  return CG_OK;
}

// ---- A2 ------------------------------------------------------------------------

int cg_gmm_like_2d(double* d_probs, const double* d_x, const double* d_ws, const double* d_mus,
                   const double* d_sigs, int64_t N, int32_t K, void* stream) {
  NEED_INIT();
  if (N < 0 || K < 0) return fail(CG_EINVAL, "negative size");
  if (K > 4096) return fail(CG_EINVAL, "K=%d exceeds 4096 components", K);
  cudaStream_t st = (cudaStream_t)stream;
  if (N == 0) return CG_OK;
  if (K == 0) { CUDA_TRY(cudaMemsetAsync(d_probs, 0, sizeof(double) * N, st)); return CG_OK; }
  double* g_gmm_tab = nullptr;
  { int rc = stream_scratch(0, st, (size_t)6 * K, &g_gmm_tab); if (rc) return rc; }
  gmm_prep_kernel<<<(K + 63) / 64, 64, 0, st>>>(d_ws, d_mus, d_sigs, K, g_gmm_tab);
  LAUNCH_CHECK();
  int blocks = (int)std::min<long long>((N + 255) / 256, (long long)g_sm_count * 8);
  size_t smem = sizeof(double) * 6 * K;
  if (smem > 48 * 1024)
    CUDA_TRY(cudaFuncSetAttribute(gmm_points_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  gmm_points_kernel<<<blocks, 256, smem, st>>>(d_probs, d_x, g_gmm_tab, N, K);
  LAUNCH_CHECK();
  return CG_OK;
}

int cg_mvn2d_logpdf(double* d_lls, const double* d_x, const double* mean, const double* covinv,
                    double logdet, int64_t N, void* stream) {
  NEED_INIT();
  if (N <= 0) return CG_OK;
  MvnArgs m;
  m.mean[0] = mean[0]; m.mean[1] = mean[1];
  for (int i = 0; i < 4; ++i) m.ci[i] = covinv[i];
  m.logdet = (double)(float)logdet;  // the reference declares `float logdet` (gmm_like_fast.pyx:39)
  mvn2d_logpdf_kernel<<<(unsigned)((N + 255) / 256), 256, 0, (cudaStream_t)stream>>>(d_lls, d_x, m, N);
  LAUNCH_CHECK();
  return CG_OK;
}

// ---- engine --------------------------------------------------------------------
size_t cg_workspace_bytes(int32_t kind, int64_t n_items, int32_t n_img, int64_t max_patch_px,
                          int64_t n_tiles) {
  WsLayout L;
  switch (kind) {
    case CG_WS_RENDER: L = ws_layout(n_items, max_patch_px, 0, 0, 0); break;
    case CG_WS_FIELD: L = ws_layout(std::max<long long>(n_items, 1), max_patch_px, 1, 0, n_tiles); break;
    case CG_WS_ITEMS: L = ws_layout(n_items * 2, max_patch_px, 0, 0, 0); break;
    case CG_WS_SCORE: L = ws_layout(n_items, max_patch_px, 0, n_img, 0, n_img > 0 ? (int)(n_items / n_img) - 1 : 0); break;
    default: return 0;
  }
  return L.total + 256;
}

int32_t cg_field_tiles(int32_t H, int32_t W) {
  return ((W + TILE_W - 1) / TILE_W) * ((H + TILE_H - 1) / TILE_H);
}

static int fill_imgs(CgImage* const* imgs, int n_img, ImgPtrs* out) {
  if (n_img < 1 || n_img > CG_MAX_IMAGES) return fail(CG_EINVAL, "n_img must be in [1,%d]", CG_MAX_IMAGES);
  for (int i = 0; i < n_img; ++i) {
    if (!imgs[i]) return fail(CG_EINVAL, "null image handle");
    out->p[i] = imgs[i]->d;
  }
  return CG_OK;
}

}  // extern "C"

static int g_cull = 1;
static int g_recur = 1;   // CG_OPT_ROW_RECURRENCE

// persistent grid: one CTA per resident slot
template <typename real, int MODE>
static int render_grid() {
  static int per_sm = 0;
  if (!per_sm) {
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, render_kernel<real, MODE>, STRIP_THREADS, 0) !=
            cudaSuccess || per_sm < 1)
      per_sm = 4;
  }
  return g_sm_count * per_sm;
}

// warp tiles per work unit: coarse units amortise the table staging, fine ones
// balance a small launch over the resident warps
static int choose_unit_wt(long long n_items) {
  const long long warps = (long long)g_sm_count * 32;
  long long u = (n_items * 16) / (warps * 8);
  return (int)std::min<long long>(std::max<long long>(u, 1), UNIT_WT_MAX);
}

template <typename real>
static int run_prepare(const PrepArgs& pa_in, cudaStream_t st) {
  if (pa_in.n_items == 0) return CG_OK;
  PrepArgs pa = pa_in;
  pa.unit_wt = choose_unit_wt(pa.n_items);
  unsigned blocks = (unsigned)((pa.n_items + PREP_WARPS - 1) / PREP_WARPS);
  prepare_kernel<real><<<blocks, 32 * PREP_WARPS, 0, st>>>(pa);
  LAUNCH_CHECK();
  return CG_OK;
}

// CUDA-event bracket of the dominant kernel's launches (cg_profile_render)
static bool prof_begin(cudaStream_t st) {
  bool prof = g_prof_on && g_prof_n < PROF_MAX;
  if (prof) {      // events of a captured launch cannot be timed: skip them inside a CUDA graph capture
    cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
    if (cudaStreamIsCapturing(st, &cs) != cudaSuccess || cs != cudaStreamCaptureStatusNone) prof = false;
  }
  if (prof) cudaEventRecord(g_prof_ev[2 * g_prof_n], st);
  return prof;
}
static void prof_end(bool prof, cudaStream_t st) {
  if (prof) { cudaEventRecord(g_prof_ev[2 * g_prof_n + 1], st); ++g_prof_n; }
}

template <typename real, int MODE>
static int run_render(const RenderArgs& ra_in, long long n_items, cudaStream_t st) {
  if (n_items == 0) return CG_OK;
  RenderArgs ra = ra_in;
  ra.unit_wt = choose_unit_wt(n_items);
  const bool prof = prof_begin(st);
  render_kernel<real, MODE><<<render_grid<real, MODE>(), STRIP_THREADS, 0, st>>>(ra);
  prof_end(prof, st);
  LAUNCH_CHECK();
  return CG_OK;
}

extern "C" {

int cg_render_patches(CgImage* const* imgs, int32_t n_img, const CgSource* d_srcs, int32_t n_src,
                      const CgItem* d_items, int64_t n_items, int32_t default_flags,
                      double bound_cap, int64_t max_patch_px, float* d_patches, CgBox* d_boxes,
                      void* d_ws, size_t ws_bytes, void* stream) {
  NEED_INIT();
  cudaStream_t st = (cudaStream_t)stream;
  PrepArgs pa;
  memset(&pa, 0, sizeof pa);
  int rc = fill_imgs(imgs, n_img, &pa.imgs);
  if (rc) return rc;
  if (!d_items) n_items = (int64_t)n_src * n_img;
  if (n_items <= 0) return CG_OK;
  WsLayout L = ws_layout(n_items, max_patch_px, 0, 0, 0);
  if (L.total > ws_bytes) return fail(CG_ENOMEM, "workspace too small: need %zu bytes, have %zu", L.total, ws_bytes);
  char* base = (char*)d_ws;
  Counters* cnt = (Counters*)(base + L.counters);
  CUDA_TRY(cudaMemsetAsync(cnt, 0, sizeof(Counters), st));
  pa.n_img = n_img; pa.srcs = d_srcs; pa.n_src = n_src; pa.items = d_items; pa.n_items = n_items;
  pa.default_flags = default_flags; pa.implicit_mode = 0; pa.bound_cap = bound_cap;
  pa.max_patch_px = max_patch_px; pa.hdr = (ItemHdr*)(base + L.hdr); pa.tab = base + L.tab;
  pa.worklist = (int2*)(base + L.worklist); pa.worklist_cap = L.worklist_cap; pa.cnt = cnt;
  pa.out_boxes = d_boxes; pa.filters = g_d_filters; pa.flux_table = g_flux_table; pa.flux_rows = g_flux_rows;
  rc = run_prepare<float>(pa, st);
  if (rc) return rc;
  RenderArgs ra;
  memset(&ra, 0, sizeof ra);
  ra.imgs = pa.imgs; ra.hdr = pa.hdr; ra.tab = pa.tab; ra.worklist = pa.worklist; ra.worklist_cap = pa.worklist_cap; ra.cull = g_cull | (g_recur ? 0 : 2); ra.logtab = g_d_logtab; ra.cnt = cnt;
  ra.patches = d_patches; ra.max_patch_px = max_patch_px;
  rc = run_render<float, MODE_PATCH>(ra, n_items, st);
  g_last_counters = cnt;
  return rc;
}

}  // extern "C"

// Set-up shared by the entry points that walk the 16 x 16 image bins (full-field likelihood, photon
// allocation): prepare_kernel with binning, then bin_order_kernel.  Fills pa / fa / L.
static int field_setup(CgImage* const* imgs, const int32_t* h_img_off, const CgSource* d_srcs,
                       const int32_t* h_src_off, int32_t n_fields, double bound_cap, int64_t max_patch_px,
                       void* d_ws, size_t ws_bytes, size_t extra_bytes, cudaStream_t st, int default_flags,
                       CgBox* d_out_boxes, PrepArgs& pa, FieldArgs& fa, WsLayout& L) {
  if (n_fields < 1 || n_fields > CG_MAX_FIELDS) return fail(CG_EINVAL, "n_fields must be in [1,%d]", CG_MAX_FIELDS);
  if (!h_img_off || !h_src_off) return fail(CG_EINVAL, "null offsets");
  const int n_img = h_img_off[n_fields];
  memset(&pa, 0, sizeof pa);
  int rc = fill_imgs(imgs, n_img, &pa.imgs);
  if (rc) return rc;
  memset(&fa, 0, sizeof fa);
  long long n_items = 0;
  pa.fb.n_fields = n_fields;
  for (int f = 0; f < n_fields; ++f) {
    const int ni = h_img_off[f + 1] - h_img_off[f], ns = h_src_off[f + 1] - h_src_off[f];
    if (ni < 1 || ns < 0) return fail(CG_EINVAL, "field %d: bad image/source offsets", f);
    pa.fb.item_off[f] = (int)n_items;
    pa.fb.src_off[f] = h_src_off[f];
    pa.fb.img_off[f] = h_img_off[f];
    for (int b = 0; b < ni; ++b) {
      const int gi = h_img_off[f] + b;
      fa.item_base[gi] = (int)n_items + b;
      fa.item_stride[gi] = ni;
      fa.n_src[gi] = ns;
    }
    n_items += (long long)ns * ni;
  }
  pa.fb.item_off[n_fields] = (int)n_items;
  pa.fb.src_off[n_fields] = h_src_off[n_fields];
  pa.fb.img_off[n_fields] = n_img;
  if (n_items > 0x7fffffffLL / 64) return fail(CG_EINVAL, "too many (source, image) items");
  int max_tiles = 0;
  for (int i = 0; i < n_img; ++i) {
    max_tiles = std::max(max_tiles, (int)cg_field_tiles(imgs[i]->h.H, imgs[i]->h.W));
    if (!imgs[i]->h.nelec) return fail(CG_EINVAL, "image %d has no nelec buffer", i);
    if (imgs[i]->h.H > 4096 * BIN_H || imgs[i]->h.W > 4096 * BIN_W)       // 12-bit bin coordinates (bin_order_kernel)
      return fail(CG_EINVAL, "image %d is larger than %d x %d pixels", i, 4096 * BIN_W, 4096 * BIN_H);
  }
  L = ws_layout(std::max<long long>(n_items, 1), max_patch_px, 1, 0, (long long)max_tiles * n_img);
  if (L.total + extra_bytes > ws_bytes)
    return fail(CG_ENOMEM, "workspace too small: need %zu bytes, have %zu", L.total + extra_bytes, ws_bytes);
  char* base = (char*)d_ws;
  Counters* cnt = (Counters*)(base + L.counters);
  CUDA_TRY(cudaMemsetAsync(cnt, 0, sizeof(Counters), st));
  CUDA_TRY(cudaMemsetAsync(base + L.tile_count, 0, L.tile_zero_bytes, st));
  pa.n_img = n_img; pa.srcs = d_srcs; pa.n_src = h_src_off[n_fields]; pa.items = nullptr; pa.n_items = n_items;
  pa.default_flags = default_flags; pa.implicit_mode = 3; pa.bound_cap = bound_cap;
  pa.max_patch_px = max_patch_px; pa.hdr = (ItemHdr*)(base + L.hdr); pa.tab = base + L.tab;
  pa.worklist = nullptr; pa.worklist_cap = 0; pa.cnt = cnt;     // rendered by image bin, not by work list
  pa.filters = g_d_filters; pa.flux_table = g_flux_table; pa.flux_rows = g_flux_rows;
  pa.out_boxes = d_out_boxes; pa.cap_boxes = d_out_boxes != nullptr;
  pa.tile_count = (int*)(base + L.tile_count); pa.tile_list = (int*)(base + L.tile_list);
  const int max_bins = max_tiles * BINS_PER_TILE;
  pa.max_tiles = max_bins;
  pa.bin_w = BIN_W; pa.bin_h = BIN_H; pa.bin_cap = BIN_CAP;
  rc = run_prepare<float>(pa, st);
  if (rc) return rc;
  fa.imgs = pa.imgs; fa.n_img = n_img; fa.hdr = pa.hdr; fa.tab = (const float*)pa.tab;
  fa.tile_count = pa.tile_count; fa.tile_list = pa.tile_list;
  fa.order = (int2*)(base + L.order); fa.sched = (int*)(base + L.done);
  fa.block_partials = (double*)(base + L.tiles);
  fa.max_tiles = max_bins; fa.tiles_total = max_bins * n_img;
  if ((long long)max_bins * n_img >= (1LL << 25))
    return fail(CG_EINVAL, "%lld image bins in one call exceed the bin descriptor's 25-bit index", (long long)max_bins * n_img);
  fa.logtab = g_d_logtab; fa.cnt = cnt; fa.cull = g_cull | (g_recur ? 0 : 2);
  bin_order_kernel<<<(fa.tiles_total + BO_THREADS - 1) / BO_THREADS, BO_THREADS, 0, st>>>(fa);
  LAUNCH_CHECK();
  return CG_OK;
}

extern "C" {

int cg_fields_loglik(CgImage* const* imgs, const int32_t* h_img_off, const CgSource* d_srcs,
                     const int32_t* h_src_off, int32_t n_fields, double bound_cap,
                     int64_t max_patch_px, double* d_ll, float* const* h_d_model, void* d_ws,
                     size_t ws_bytes, void* stream) {
  NEED_INIT();
  cudaStream_t st = (cudaStream_t)stream;
  PrepArgs pa;
  FieldArgs fa;
  WsLayout L;
  int rc = field_setup(imgs, h_img_off, d_srcs, h_src_off, n_fields, bound_cap, max_patch_px, d_ws, ws_bytes, 0,
                       st, CG_ITEM_FLUX_BAND, nullptr, pa, fa, L);
  if (rc) return rc;
  char* base = (char*)d_ws;
  const int n_img = pa.n_img;
  fa.ll = d_ll;
  if (h_d_model) {
    bool any = false;
    for (int i = 0; i < n_img; ++i) any = any || h_d_model[i];
    if (any) {
      CUDA_TRY(cudaMemcpyAsync(base + L.model_ptrs, h_d_model, sizeof(float*) * n_img, cudaMemcpyHostToDevice, st));
      fa.model = (float* const*)(base + L.model_ptrs);
    }
  }
  const bool prof = prof_begin(st);
  field_tile_kernel<<<g_sm_count * FT_MINB, FT_LAUNCH_THREADS, 0, st>>>(fa);      // persistent: FT_MINB CTAs per SM (launch bounds)
  prof_end(prof, st);
  LAUNCH_CHECK();
  field_tile_final_kernel<<<n_img, FIN_THREADS, 0, st>>>(fa);
  LAUNCH_CHECK();
  g_last_counters = pa.cnt;
  return CG_OK;
}

int cg_field_loglik(CgImage* const* imgs, int32_t n_img, const CgSource* d_srcs, int32_t n_src,
                    double bound_cap, int64_t max_patch_px, double* d_ll, float* const* h_d_model,
                    void* d_ws, size_t ws_bytes, void* stream) {
  if (n_src < 0) return fail(CG_EINVAL, "negative n_src");
  const int32_t img_off[2] = {0, n_img}, src_off[2] = {0, n_src};
  return cg_fields_loglik(imgs, img_off, d_srcs, src_off, 1, bound_cap, max_patch_px, d_ll, h_d_model,
                          d_ws, ws_bytes, stream);
}

int cg_set_option(int32_t option, int32_t value) {
  switch (option) {
    case CG_OPT_CULL: g_cull = value ? 1 : 0; return CG_OK;
    case CG_OPT_ROW_RECURRENCE: g_recur = value ? 1 : 0; return CG_OK;
    default: return fail(CG_EINVAL, "unknown option %d", option);
  }
}

int cg_items_loglik(CgImage* const* imgs, int32_t n_img, const CgSource* d_srcs, int32_t n_src,
                    const CgItem* d_items, int64_t n_items, int32_t mode, double bound_cap,
                    int64_t max_patch_px, const float* d_aux, const int64_t* d_aux_offsets,
                    double* d_item_ll, CgBox* d_boxes, int32_t use_fp64, void* d_ws,
                    size_t ws_bytes, void* stream) {
  NEED_INIT();
  cudaStream_t st = (cudaStream_t)stream;
  if (mode != CG_LL_POISSON_BG && mode != CG_LL_CONDITIONAL && mode != CG_LL_CONDITIONAL_SUM)
    return fail(CG_EINVAL, "unknown mode %d", mode);
  if (!d_items) return fail(CG_EINVAL, "d_items is required");
  if (mode != CG_LL_POISSON_BG && !d_aux) return fail(CG_EINVAL, "conditional mode needs photon patches");
  PrepArgs pa;
  memset(&pa, 0, sizeof pa);
  int rc = fill_imgs(imgs, n_img, &pa.imgs);
  if (rc) return rc;
  if (n_items <= 0) return CG_OK;
  long long max_px = 0;
  for (int i = 0; i < n_img; ++i) max_px = std::max(max_px, (long long)imgs[i]->h.H * imgs[i]->h.W);
  if (max_patch_px > 0) max_px = std::min<long long>(max_px, max_patch_px);
  // tables are sized for double so either precision fits
  WsLayout L = ws_layout(n_items * 2, max_px, 0, 0, 0);
  if (L.total > ws_bytes) return fail(CG_ENOMEM, "workspace too small: need %zu bytes, have %zu", L.total, ws_bytes);
  char* base = (char*)d_ws;
  Counters* cnt = (Counters*)(base + L.counters);
  CUDA_TRY(cudaMemsetAsync(cnt, 0, sizeof(Counters), st));
  pa.n_img = n_img; pa.srcs = d_srcs; pa.n_src = n_src; pa.items = d_items; pa.n_items = n_items;
  pa.bound_cap = bound_cap; pa.max_patch_px = max_px; pa.hdr = (ItemHdr*)(base + L.hdr);
  pa.tab = base + L.tab; pa.worklist = (int2*)(base + L.worklist); pa.worklist_cap = L.worklist_cap;
  pa.cnt = cnt; pa.out_boxes = d_boxes; pa.filters = g_d_filters; pa.flux_table = g_flux_table; pa.flux_rows = g_flux_rows;
  RenderArgs ra;
  memset(&ra, 0, sizeof ra);
  ra.imgs = pa.imgs; ra.hdr = pa.hdr; ra.tab = pa.tab; ra.worklist = pa.worklist; ra.worklist_cap = pa.worklist_cap; ra.cull = g_cull | (g_recur ? 0 : 2); ra.logtab = g_d_logtab; ra.cnt = cnt;
  ra.aux = d_aux; ra.aux_offsets = (const long long*)d_aux_offsets;
  ra.partials = (double*)(base + L.partials);
  ra.cond_sum = (mode == CG_LL_CONDITIONAL_SUM);
  if (use_fp64) {
    rc = run_prepare<double>(pa, st);
    if (rc) return rc;
    rc = (mode == CG_LL_POISSON_BG) ? run_render<double, MODE_POISSON>(ra, n_items, st)
                                    : run_render<double, MODE_COND>(ra, n_items, st);
  } else {
    rc = run_prepare<float>(pa, st);
    if (rc) return rc;
    rc = (mode == CG_LL_POISSON_BG) ? run_render<float, MODE_POISSON>(ra, n_items, st)
                                    : run_render<float, MODE_COND>(ra, n_items, st);
  }
  if (rc) return rc;
  items_finalize_kernel<<<(unsigned)((n_items + 255) / 256), 256, 0, st>>>(
      pa.hdr, ra.partials, n_items, mode, d_item_ll, nullptr, 1);
  LAUNCH_CHECK();
  g_last_counters = cnt;
  return CG_OK;
}

int cg_score_proposals(CgImage* const* imgs, int32_t n_img, const CgSource* d_srcs, int32_t P,
                       const CgBox* h_boxes, const float* d_bg, const int64_t* h_bg_offsets,
                       const double* d_log_terms, const double* d_log_u, double guard,
                       double* d_delta, uint8_t* d_accept, int32_t* d_n_guard, void* d_ws,
                       size_t ws_bytes, void* stream) {
  NEED_INIT();
  cudaStream_t st = (cudaStream_t)stream;
  PrepArgs pa;
  memset(&pa, 0, sizeof pa);
  int rc = fill_imgs(imgs, n_img, &pa.imgs);
  if (rc) return rc;
  if (P <= 0) return CG_OK;
  if (!h_boxes) return fail(CG_EINVAL, "bounding boxes are required");
  long long max_px = 0;
  for (int i = 0; i < n_img; ++i) {
    long long w = std::max(0, std::min(h_boxes[i].x1, imgs[i]->h.W) - std::max(h_boxes[i].x0, 0));
    long long h = std::max(0, std::min(h_boxes[i].y1, imgs[i]->h.H) - std::max(h_boxes[i].y0, 0));
    max_px = std::max(max_px, w * h);
    if (!imgs[i]->h.nelec) return fail(CG_EINVAL, "image %d has no nelec buffer", i);
  }
  const long long n_items = (long long)(P + 1) * n_img;
  WsLayout L = ws_layout(n_items, max_px, 0, guard > 0 ? n_img : 0, 0, P);
  if (L.total > ws_bytes) return fail(CG_ENOMEM, "workspace too small: need %zu bytes, have %zu", L.total, ws_bytes);
  char* base = (char*)d_ws;
  Counters* cnt = (Counters*)(base + L.counters);
  CUDA_TRY(cudaMemsetAsync(cnt, 0, sizeof(Counters), st));
  if (d_bg && !h_bg_offsets) return fail(CG_EINVAL, "background offsets are required with d_bg");
  // the boxes and offsets (n_img <= 64 small records) travel inside the kernel arguments: no
  // host->device copy from pageable memory, and the call can be captured into a CUDA graph
  for (int i = 0; i < n_img; ++i) pa.img_boxes[i] = h_boxes[i];
  pa.n_img = n_img; pa.srcs = d_srcs; pa.n_src = P + 1; pa.items = nullptr; pa.n_items = n_items;
  pa.default_flags = CG_ITEM_FLUX_BAND; pa.implicit_mode = 1;
  pa.aux_per_img = d_bg ? 1 : 0;
  pa.bound_cap = 0; pa.max_patch_px = max_px; pa.hdr = (ItemHdr*)(base + L.hdr);
  pa.tab = base + L.tab; pa.worklist = (int2*)(base + L.worklist); pa.worklist_cap = L.worklist_cap;
  pa.cnt = cnt; pa.filters = g_d_filters; pa.flux_table = g_flux_table; pa.flux_rows = g_flux_rows;
  rc = run_prepare<float>(pa, st);
  if (rc) return rc;
  RenderArgs ra;
  memset(&ra, 0, sizeof ra);
  ra.imgs = pa.imgs; ra.hdr = pa.hdr; ra.tab = pa.tab; ra.worklist = pa.worklist; ra.worklist_cap = pa.worklist_cap; ra.cull = g_cull | (g_recur ? 0 : 2); ra.logtab = g_d_logtab; ra.cnt = cnt;
  ra.aux = d_bg; ra.aux_offsets = nullptr; ra.partials = (double*)(base + L.partials);
  if (d_bg) for (int i = 0; i < n_img; ++i) ra.aux_off_v[i] = h_bg_offsets[i];
  rc = run_render<float, MODE_POISSON>(ra, n_items, st);
  if (rc) return rc;
  double* item_ll = (double*)(base + L.item_ll);
  items_finalize_kernel<<<(unsigned)((n_items + 255) / 256), 256, 0, st>>>(
      pa.hdr, ra.partials, n_items, CG_LL_POISSON_BG, item_ll, nullptr, 1);
  LAUNCH_CHECK();
  int* guard_list = guard > 0 ? (int*)(base + L.guard_list) : nullptr;
  score_finalize_kernel<<<(P + 255) / 256, 256, 0, st>>>(item_ll, n_img, P, d_log_terms, d_log_u, guard,
                                                        d_delta, d_accept, cnt, guard_list, L.guard_cap);
  LAUNCH_CHECK();
  if (guard > 0) {
    // float64 re-score of the guard-band proposals (+ the current state)
    Counters* cnt2 = (Counters*)(base + L.counters2);
    CUDA_TRY(cudaMemsetAsync(cnt2, 0, sizeof(Counters), st));
    clamp_guard_kernel<<<1, 1, 0, st>>>(cnt, L.guard_cap);
    LAUNCH_CHECK();
    PrepArgs p2 = pa;
    p2.implicit_mode = 2; p2.ind_list = guard_list; p2.ind_count = &cnt->guard_count;
    p2.n_items = L.n_items2; p2.hdr = (ItemHdr*)(base + L.hdr2); p2.tab = base + L.tab2;
    p2.worklist = (int2*)(base + L.worklist2); p2.worklist_cap = L.worklist_cap2; p2.cnt = cnt2;
    rc = run_prepare<double>(p2, st);
    if (rc) return rc;
    RenderArgs r2 = ra;
    r2.hdr = p2.hdr; r2.tab = p2.tab; r2.worklist = p2.worklist; r2.worklist_cap = p2.worklist_cap; r2.cnt = cnt2;
    r2.partials = (double*)(base + L.partials2);
    rc = run_render<double, MODE_POISSON>(r2, L.n_items2, st);
    if (rc) return rc;
    double* item_ll2 = (double*)(base + L.item_ll2);
    items_finalize_kernel<<<(unsigned)((L.n_items2 + 255) / 256), 256, 0, st>>>(
        p2.hdr, r2.partials, L.n_items2, CG_LL_POISSON_BG, item_ll2, &cnt->guard_count, n_img);
    LAUNCH_CHECK();
    score_rescore_kernel<<<(L.guard_cap + 255) / 256, 256, 0, st>>>(
        item_ll2, n_img, cnt, guard_list, L.guard_cap, d_log_terms, d_log_u, d_delta, d_accept);
    LAUNCH_CHECK();
  }
  if (d_n_guard)
    CUDA_TRY(cudaMemcpyAsync(d_n_guard, &cnt->guard_count, sizeof(int), cudaMemcpyDeviceToDevice, st));
  g_last_counters = cnt;
  return CG_OK;
}

// ---- sources.py:6-12 ---------------------------------------------------------------
int cg_poisson_loglike(const double* d_data, const double* d_model, const double* d_mask, int64_t n,
                       double* d_out, void* stream) {
  NEED_INIT();
  if (n < 0) return fail(CG_EINVAL, "negative size");
  cudaStream_t st = (cudaStream_t)stream;
  double* g_scratch = nullptr;        // 1024 doubles of block partials, owned by this stream
  { int rc = stream_scratch(1, st, 1024, &g_scratch); if (rc) return rc; }
  int blocks = (int)std::min<long long>(1024, std::max<long long>(1, (n + 255) / 256));
  poisson_partial_kernel<<<blocks, 256, 0, st>>>(d_data, d_model, d_mask, n, g_scratch);
  LAUNCH_CHECK();
  sum_partials_kernel<<<1, 256, 0, st>>>(g_scratch, blocks, d_out);
  LAUNCH_CHECK();
  return CG_OK;
}

// the same reduction with its scratch in g_static_partials: no cudaMalloc / cudaFree, no allocating
// library call, no lock, no map look-up -- callable under CUDA graph capture
int cg_poisson_loglike_static(const double* d_data, const double* d_model, const double* d_mask, int64_t n,
                              double* d_out, void* stream) {
  NEED_INIT();
  if (n < 0) return fail(CG_EINVAL, "negative size");
  if (n > (int64_t)CG_POISSON_STATIC_MAX_N)
    return fail(CG_EINVAL, "cg_poisson_loglike_static: n = %lld exceeds the static scratch's limit of %lld elements "
                "(CG_POISSON_STATIC_MAX_N); use cg_poisson_loglike", (long long)n, (long long)CG_POISSON_STATIC_MAX_N);
  if (!d_data || !d_model || !d_out) return fail(CG_EINVAL, "null argument");
  double* scratch = (g_device >= 0 && g_device < STATIC_MAX_DEVICES) ? g_static_partials_ptr[g_device] : nullptr;
  if (!scratch) return fail(CG_ECUDA, "the static scratch of device %d was not resolved at cg_init", g_device);
  cudaStream_t st = (cudaStream_t)stream;
  // the launch shape of cg_poisson_loglike: the two entry points return the same bits
  int blocks = (int)std::min<long long>(CG_POISSON_STATIC_MAX_BLOCKS, std::max<long long>(1, (n + 255) / 256));
  // nothing has to start at zero: every one of the `blocks` slots read below is written by its block first
  poisson_partial_kernel<<<blocks, 256, 0, st>>>(d_data, d_model, d_mask, n, scratch);
  LAUNCH_CHECK();
  sum_partials_kernel<<<1, 256, 0, st>>>(scratch, blocks, d_out);
  LAUNCH_CHECK();
  return CG_OK;
}

// ---- N1: celeste_sample_sources.pyx:61-122 -------------------------------------
size_t cg_alloc_workspace_bytes(int32_t n_img, int64_t n_tiles) {
  size_t o = 0;
  auto take = [&](size_t bytes) { size_t r = o; o = align_up(o + bytes, 256); return r; };
  take(sizeof(int) * (size_t)n_tiles);                       // tile_count
  take(sizeof(int) * (size_t)n_tiles * TILE_LIST_CAP);       // tile_list
  take(sizeof(double) * (size_t)n_tiles);                    // block partials
  (void)n_img;
  return o + 256;
}

int cg_sample_source_counts(CgImage* const* imgs, int32_t n_img, int32_t n_src, const CgBox* d_boxes,
                            const float* d_patches, int64_t max_patch_px, uint64_t seed,
                            float* d_samp, double* d_noise_sum, void* d_ws, size_t ws_bytes,
                            void* stream) {
  NEED_INIT();
  cudaStream_t st = (cudaStream_t)stream;
  AllocArgs a;
  memset(&a, 0, sizeof a);
  int rc = fill_imgs(imgs, n_img, &a.imgs);
  if (rc) return rc;
  if (n_src < 0) return fail(CG_EINVAL, "negative n_src");
  int max_tiles = 0;
  for (int i = 0; i < n_img; ++i) {
    max_tiles = std::max(max_tiles, (int)cg_field_tiles(imgs[i]->h.H, imgs[i]->h.W));
    if (!imgs[i]->h.nelec) return fail(CG_EINVAL, "image %d has no nelec buffer", i);
  }
  const long long n_tiles = (long long)max_tiles * n_img;
  if (cg_alloc_workspace_bytes(n_img, n_tiles) > ws_bytes)
    return fail(CG_ENOMEM, "workspace too small: need %zu bytes, have %zu",
                cg_alloc_workspace_bytes(n_img, n_tiles), ws_bytes);
  char* base = (char*)d_ws;
  size_t o = 0;
  auto take = [&](size_t bytes) { size_t r = o; o = align_up(o + bytes, 256); return r; };
  a.tile_count = (int*)(base + take(sizeof(int) * (size_t)n_tiles));
  a.tile_list = (int*)(base + take(sizeof(int) * (size_t)n_tiles * TILE_LIST_CAP));
  a.block_partials = (double*)(base + take(sizeof(double) * (size_t)n_tiles));
  a.n_img = n_img; a.n_src = n_src; a.boxes = d_boxes; a.patches = d_patches; a.samp = d_samp;
  a.max_patch_px = max_patch_px; a.noise_sum = d_noise_sum; a.max_tiles = max_tiles;
  a.blocks_per_img = (2 * max_tiles + FR_WARPS - 1) / FR_WARPS; a.seed = seed;
  CUDA_TRY(cudaMemsetAsync(a.tile_count, 0, sizeof(int) * (size_t)n_tiles, st));
  const long long n_items = (long long)n_src * n_img;
  if (n_items > 0) {
    CUDA_TRY(cudaMemsetAsync(d_samp, 0, sizeof(float) * (size_t)(n_items * max_patch_px), st));
    bin_boxes_kernel<<<(unsigned)((n_items + 3) / 4), 128, 0, st>>>(a);
    LAUNCH_CHECK();
  }
  sample_counts_kernel<<<dim3(a.blocks_per_img, n_img), FR_WARPS * 32, 0, st>>>(a);
  LAUNCH_CHECK();
  alloc_final_kernel<<<n_img, 256, 0, st>>>(a);
  LAUNCH_CHECK();
  return CG_OK;
}

int cg_binomial_draws(int64_t n, double p, uint64_t seed, int64_t count, double* d_out, void* stream) {
  NEED_INIT();
  if (count <= 0) return CG_OK;
  binomial_test_kernel<<<(unsigned)((count + 255) / 256), 256, 0, (cudaStream_t)stream>>>(n, p, seed, count, d_out);
  LAUNCH_CHECK();
  return CG_OK;
}

// ---- A16 / A17 -----------------------------------------------------------------
int cg_planck_fluxes(const double* d_T, const double* d_b, int64_t n, double* d_out, void* stream) {
  NEED_INIT();
  if (!g_filters_loaded) return fail(CG_ESTATE, "cg_set_filter_curves() has not been called");
  if (n <= 0) return CG_OK;
  long long threads = n * CG_NUM_BANDS * 32;
  planck_kernel<<<(unsigned)((threads + 255) / 256), 256, 0, (cudaStream_t)stream>>>(d_T, d_b, n, d_out,
                                                                                     g_d_filters);
  LAUNCH_CHECK();
  return CG_OK;
}

int cg_eigen_fluxes(const double* d_z, const double* d_w, const double* d_m, const double* d_B,
                    const double* d_lam0, int32_t n_chain, int32_t K, int32_t V,
                    const double* d_fluxes, const double* d_fluxes_ivar, double* d_mu, double* d_ll,
                    void* stream) {
  NEED_INIT();
  if (!g_filters_loaded) return fail(CG_ESTATE, "cg_set_filter_curves() has not been called");
  if (n_chain <= 0) return CG_OK;
  if (d_fluxes && !d_fluxes_ivar) return fail(CG_EINVAL, "fluxes_ivar required with fluxes");
  eigen_flux_kernel<<<n_chain, 256, 0, (cudaStream_t)stream>>>(d_z, d_w, d_m, d_B, d_lam0, K, V, d_fluxes,
                                                               d_fluxes_ivar, d_mu, d_ll, g_d_filters);
  LAUNCH_CHECK();
  return CG_OK;
}

// ---- instrumentation -------------------------------------------------------------
int cg_mog_logpdf(const double* d_x, int64_t N, int32_t D, const double* d_means, const double* d_icovs,
                  const double* d_dets, const double* d_pis, int32_t K, double* d_logp, double* d_resp,
                  void* stream) {
  NEED_INIT();
  if (N < 0 || K < 1 || D < 1 || D > CG_MOG_MAX_DIM) return fail(CG_EINVAL, "cg_mog_logpdf: need N >= 0, K >= 1, 1 <= D <= %d", CG_MOG_MAX_DIM);
  if (N == 0) return CG_OK;
  mog_logpdf_kernel<<<(unsigned)((N + 3) / 4), 128, 0, (cudaStream_t)stream>>>(d_x, N, D, d_means, d_icovs, d_dets,
                                                                              d_pis, K, d_logp, d_resp);
  LAUNCH_CHECK();
  return CG_OK;
}

int cg_prior_stats(const double* d_resp, const double* d_c, int64_t N, int32_t K, int32_t D, double* d_out,
                   void* stream) {
  NEED_INIT();
  if (N < 0 || K < 1 || D < 1 || D > CG_MOG_MAX_DIM) return fail(CG_EINVAL, "cg_prior_stats: need N >= 0, K >= 1, 1 <= D <= %d", CG_MOG_MAX_DIM);
  const int entries = K * (1 + D + D * (D + 1) / 2);
  prior_stats_kernel<<<entries, 256, 0, (cudaStream_t)stream>>>(d_resp, d_c, N, K, D, d_out);
  LAUNCH_CHECK();
  return CG_OK;
}

int cg_profile_render(int enable) {
  NEED_INIT();
  if (enable && g_prof_ev.empty()) {
    g_prof_ev.resize(2 * PROF_MAX);
    for (auto& e : g_prof_ev) CUDA_TRY(cudaEventCreate(&e));
  }
  g_prof_on = enable != 0;
  g_prof_n = 0;
  return CG_OK;
}

int cg_profile_render_read(double* total_ms, int64_t* launches) {
  NEED_INIT();
  double t = 0;
  for (int i = 0; i < g_prof_n; ++i) {
    CUDA_TRY(cudaEventSynchronize(g_prof_ev[2 * i + 1]));
    float ms = 0;
    CUDA_TRY(cudaEventElapsedTime(&ms, g_prof_ev[2 * i], g_prof_ev[2 * i + 1]));
    t += ms;
  }
  if (total_ms) *total_ms = t;
  if (launches) *launches = g_prof_n;
  return CG_OK;
}

int cg_pipe_probe(int which, double* ops_per_s, void* stream) {
  NEED_INIT();
  cudaStream_t st = (cudaStream_t)stream;
  // which: 0 MUFU.EX2, 1 FFMA, 7 FFMA2 (packed; counted as 2 FMA per lane); 2..5 instruction-mix probes and 6 the row-recurrence loop at
  // 3 CTAs/SM (24 warps, the tile kernel's residency), 12..16 the same at 8 CTAs/SM (64 warps)
  const int var = which % 10;
  const bool raw = var < 2 || (which >= 7 && which <= 9) || which >= 20;   // single-pipe probes: long chains, 8 CTAs/SM
  const int iters = raw ? 4096 : 128, blocks = g_sm_count * ((which >= 2 && which < 10 && !raw) ? 3 : 8), threads = 256;
  float* d_out = nullptr;
  CUDA_TRY(cudaMalloc(&d_out, sizeof(float) * blocks * threads));
  cudaEvent_t e0, e1;
  CUDA_TRY(cudaEventCreate(&e0));
  CUDA_TRY(cudaEventCreate(&e1));
  float best = 1e30f;
  for (int rep = 0; rep < 6; ++rep) {
    CUDA_TRY(cudaEventRecord(e0, st));
    if (which == 0) xu_probe_kernel<<<blocks, threads, 0, st>>>(d_out, iters);
    else if (which == 1) fma_probe_kernel<<<blocks, threads, 0, st>>>(d_out, iters);
    else if (which == 7) fma2_probe_kernel<<<blocks, threads, 0, st>>>(d_out, iters);
    else if (which == 8) pk2_probe_kernel<0><<<blocks, threads, 0, st>>>(d_out, iters);
    else if (which == 9) pk2_probe_kernel<1><<<blocks, threads, 0, st>>>(d_out, iters);
    else if (which == 20) issue_probe_kernel<0><<<blocks, threads, 0, st>>>(d_out, iters);
    else if (which == 21) issue_probe_kernel<1><<<blocks, threads, 0, st>>>(d_out, iters);
    else if (which == 22) issue_probe_kernel<2><<<blocks, threads, 0, st>>>(d_out, iters);
    else if (var == 2) mix_probe_kernel<2><<<blocks, threads, 0, st>>>(d_out, iters);
    else if (var == 3) mix_probe_kernel<3><<<blocks, threads, 0, st>>>(d_out, iters);
    else if (var == 4) mix_probe_kernel<4><<<blocks, threads, 0, st>>>(d_out, iters);
    else if (var == 5) mix_probe_kernel<5><<<blocks, threads, 0, st>>>(d_out, iters);
    else recur_probe_kernel<<<blocks, threads, 0, st>>>(d_out, iters);
    CUDA_TRY(cudaEventRecord(e1, st));
    LAUNCH_CHECK();
    CUDA_TRY(cudaEventSynchronize(e1));
    float ms = 0;
    CUDA_TRY(cudaEventElapsedTime(&ms, e0, e1));
    if (rep > 0 && ms < best) best = ms;
  }
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  cudaFree(d_out);
  if (ops_per_s) *ops_per_s = (double)blocks * threads * 8.0 * iters * ((which >= 7 && which <= 9) ? 2.0 : (raw ? 1.0 : 32.0)) / (best * 1e-3);
  return CG_OK;
}

}  // extern "C"

#include "celeste_sweep.cuh"
