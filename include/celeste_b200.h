/*
 * celeste_b200.h -- C ABI of libceleste_b200.so
 *
 * B200-native (sm_100a) implementation of the CelestePy per-proposal image
 * log-likelihood: mixture-of-Gaussians rendering of stars / galaxies onto
 * SDSS-style band images, the Poisson pixel log-likelihood, and batched
 * Metropolis-Hastings proposal scoring with on-device accept/reject.
 *
 * The reference (HIPS/DESI-MCMC) has no FFI registry; its boundary is the
 * Python call surface of its Cython modules and the numpy functions around
 * them.  Each entry point below names the reference interface it replaces
 * (paths relative to the reference repository root).  The reference-side
 * binding (a ctypes stub) is shown in INTEGRATION.md.
 *
 * Conventions
 *   - every function returns 0 on success, a negative CG_E* code on failure;
 *     cg_last_error() returns the message of the calling thread's last failure.
 *   - pointers named d_* are DEVICE pointers (owned by the caller, normally
 *     torch tensors); pointers named h_* are HOST pointers; `stream` is a
 *     cudaStream_t passed as void* (NULL = legacy default stream).
 *   - all work is enqueued on `stream`; nothing synchronises unless stated.
 *   - no exceptions cross this boundary; no persistent device allocation is
 *     made by the library except the small CgImage descriptor copy.
 */
#ifndef CELESTE_B200_H
#define CELESTE_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define CG_ABI_VERSION 2

#define CG_MAX_PSF 4      /* PSF mixture components per image (reference: 3, fits_image.py:130-133) */
#define CG_NUM_BANDS 5    /* u g r i z */
#define CG_MAX_COMP 56    /* components per (source, image): (6 exp + 8 dev) x n_psf; 42 with the reference's 3-component PSF */
#define CG_MAX_IMAGES 64  /* images per call */

enum {
  CG_OK = 0,
  CG_EINVAL = -1,   /* bad argument (the Python shim raises ValueError) */
  CG_ECUDA = -2,    /* CUDA runtime error (sticky; RuntimeError)        */
  CG_ENOMEM = -3,   /* workspace too small                              */
  CG_ESTATE = -4    /* library not initialised                          */
};

/* ---- data contract ------------------------------------------------------ */

/* What the hot path reads from FitsImage (CelestePy/fits_image.py:16-155). */
typedef struct CgImageDesc {
  int32_t H, W;                 /* nelec.shape                         (:90-92)  */
  int32_t band;                 /* 0..4 = u g r i z                              */
  int32_t n_psf;                /* PSF components, <= CG_MAX_PSF        (:130)   */
  double epsilon;               /* sky rate, electrons / pixel          (:113)   */
  double kappa;                 /* gain                                 (:112)   */
  double calib;                 /* nanomaggies per count                (:116)   */
  double psf_w[CG_MAX_PSF];     /* weights                              (:130)   */
  double psf_mu[CG_MAX_PSF][2]; /* means (x, y)                         (:131)   */
  double psf_cov[CG_MAX_PSF][3];/* var_x, var_y, cov_xy                 (:132-138) */
  double rho_n[2];              /* pixel reference point (0-based)      (:99)    */
  double phi_n[2];              /* equatorial reference point           (:100)   */
  double ups[4];                /* Ups_n, row-major CD matrix           (:101-102) */
  const float* d_nelec;         /* H*W electrons, row-major [y][x]; may be NULL for render-only */
  const float* d_mask;          /* H*W invvar-style mask (0 = bad) or NULL = all good */
} CgImageDesc;

typedef struct CgImage CgImage;   /* opaque handle */

/* What the hot path reads from SrcParams (CelestePy/celeste_src.py:12-157). */
typedef struct CgSource {
  int32_t type;        /* a: 0 = star, 1 = galaxy                         */
  int32_t flags;       /* CG_SRC_* */
  double u[2];         /* RA, Dec (degrees)                               */
  double flux[CG_NUM_BANDS]; /* nanomaggies, u g r i z                    */
  double theta;        /* exp-profile weight                              */
  double sigma;        /* effective radius, arcsec                        */
  double phi;          /* position angle                                  */
  double rho;          /* axis ratio                                      */
  double t;            /* star temperature (K), used with CG_SRC_PLANCK   */
  double b;            /* star brightness ell/d^2, used with CG_SRC_PLANCK*/
  int64_t flux_row;    /* row of the bound flux table, used with CG_SRC_EIGEN */
} CgSource;

#define CG_SRC_PLANCK 1   /* star flux from black-body route (celeste.py:38-40, planck.py:155) */
#define CG_SRC_EIGEN 2    /* band fluxes (nanomaggies) = row flux_row of the table bound with cg_set_flux_table:
                             the eigen-spectrum route, experiments/redshift/redshift_utils.py:304-323 and
                             quasar_infer_photometry.py:31-46 (mu = project_to_bands(w.B, lam0 (1+z)) m / (1+z)),
                             evaluated by cg_eigen_fluxes into a DEVICE table -- no host hop per band */

/* One unit of render work: a (source, image) pair. */
typedef struct CgItem {
  int32_t src;         /* index into the source array */
  int32_t img;         /* index into the image array  */
  int32_t flags;       /* CG_ITEM_* */
  int32_t box[4];      /* x0, x1, y0, y1 (half-open) when CG_ITEM_FIXED_BOX */
  int32_t aux;         /* index of the per-item auxiliary patch (background / photons), or -1 */
} CgItem;

#define CG_ITEM_FLUX_BAND   0x00  /* rate = unit patch * flux/calib*kappa (celeste.py:80-81,91-92) */
#define CG_ITEM_FLUX_UNIT   0x01  /* unit-flux patch (celeste.py:114, celeste_galaxy_conditionals.py:185) */
#define CG_ITEM_FIXED_BOX   0x04  /* use box[] (the xlim/ylim arguments of the reference) */
#define CG_ITEM_TRUNC_BOX   0x08  /* galaxy footprint by int() truncation (celeste_galaxy_conditionals.py:161-162) */
#define CG_ITEM_PROF_EXP    0x10  /* exponential profile only, no theta weight (gen_galaxy_prof_psf_image 'exp') */
#define CG_ITEM_PROF_DEV    0x20  /* de Vaucouleurs profile only ('dev') */
#define CG_ITEM_HEADER_CD   0x40  /* galaxy transformation from Ups_n instead of cd_at_pixel (:33) */
#define CG_ITEM_NO_MASK     0x80  /* ignore the image mask (sources.py:188-237 has none) */

/* Box of an item after preparation, readable by the host. */
typedef struct CgBox { int32_t x0, x1, y0, y1; } CgBox;

/* ---- life cycle --------------------------------------------------------- */

int cg_abi_version(void);
int cg_init(int device);                     /* select device, query SM count */
int cg_shutdown(void);
int cg_last_error(char* buf, size_t n);      /* copies the message, returns its length */
int cg_sm_count(void);

/* FitsImage constructor replacement (fits_image.py:48-155): copies the
 * descriptor, derives invcovars/logdets (:141-145), R at err=.001 (:151-155). */
int cg_image_create(const CgImageDesc* desc, CgImage** out);
int cg_image_destroy(CgImage* img);
int cg_image_set_epsilon(CgImage* img, double epsilon);    /* models.py:159 resamples epsilon */
double cg_image_radius(const CgImage* img);                /* FitsImage.R */

/* Filter curves for the black-body route (planck.py:31-55): per band,
 * wavelengths (m, sorted) and h_vec = S*lambda/(hc)*dLambda, concatenated. */
int cg_set_filter_curves(const double* h_wavelengths, const double* h_hvecs,
                         const double* h_sens, const int32_t* h_counts /* [5] */);

/* Binds a device table of band fluxes [n_rows][5] (float64 nanomaggies, e.g. the d_mu output of
 * cg_eigen_fluxes) for sources flagged CG_SRC_EIGEN.  The pointer is read when a call is
 * ENQUEUED and travels in that call's kernel arguments; the table contents are read when the
 * kernels run, so an eigen_flux launch earlier on the same stream feeds the render directly.
 * d_flux = NULL unbinds.  A flagged source whose row is out of range poisons the results (NaN). */
int cg_set_flux_table(const double* d_flux, int64_t n_rows);

/* ---- A2: util/like/gmm_like_fast.pyx:130-176 ---------------------------- */
/* probs[n] = sum_k ws[k] N(x[n]; mus[k], sigs[k]).  All device pointers,
 * float64 as in the reference; arithmetic is fp32 in centred coordinates. */
int cg_gmm_like_2d(double* d_probs, const double* d_x, const double* d_ws,
                   const double* d_mus, const double* d_sigs,
                   int64_t N, int32_t K, void* stream);

/* gmm_like_fast.pyx:34-77 */
int cg_mvn2d_logpdf(double* d_lls, const double* d_x, const double* h_mean,
                    const double* h_covinv, double logdet, int64_t N, void* stream);

/* ---- A1: celeste_fast.pyx:29-94, :100-140 (host arrays, tiny) ----------- */
int cg_galaxy_prof_psf_mixture_params(const double* W, const double* v_s,
                                      const double* image_ws, const double* image_means,
                                      const double* image_covars, int32_t K_psf,
                                      const double* prof_amp, const double* prof_sigs,
                                      int32_t K_prof,
                                      double* weights, double* means, double* covars);
int cg_galaxy_psf_mixture_params(const double* thetas, const double* W, const double* v_s,
                                 const double* image_ws, const double* image_means,
                                 const double* image_covars, int32_t K_psf,
                                 const double* exp_amp, const double* exp_sigs, int32_t K_exp,
                                 const double* dev_amp, const double* dev_sigs, int32_t K_dev,
                                 double* weights, double* means, double* covars);

/* A12: util/bound/bounding_box.py:9-31 (host) */
double cg_calc_bounding_radius(const double* means, const double* covars, int32_t K,
                               double error, const double* center);

/* ---- render + likelihood engine ----------------------------------------- */

/* Workspace size (bytes) the caller must provide (a device buffer, e.g. a
 * torch uint8 tensor).  kind selects the entry point; n_items = number of
 * (source, image) items (CG_WS_SCORE: (P+1)*n_img); max_patch_px bounds the
 * pixels of one footprint; n_tiles (CG_WS_FIELD) = n_img * max_i
 * cg_field_tiles(H_i, W_i). */
#define CG_WS_RENDER 0
#define CG_WS_FIELD 1
#define CG_WS_ITEMS 2
#define CG_WS_SCORE 3
size_t cg_workspace_bytes(int32_t kind, int64_t n_items, int32_t n_img, int64_t max_patch_px,
                          int64_t n_tiles);
int32_t cg_field_tiles(int32_t H, int32_t W);

/* A5/A6/A10: render patches.  For item i the rate patch (row-major,
 * (y1-y0) x (x1-x0)) is written to d_patches + i*max_patch_px and its box to
 * d_boxes[i].  h_items may be NULL: items are then the cross product
 * source-major (item = src*n_img + img) with flags = default_flags.
 * Replaces celeste.py:114-176 (star), celeste_galaxy_conditionals.py:185-214
 * (galaxy, live route) and :134-182 (Cython route) and the flux scaling of
 * celeste.py:72-96.  bound_cap > 0 caps the automatic footprint radius. */
int cg_render_patches(CgImage* const* imgs, int32_t n_img,
                      const CgSource* d_srcs, int32_t n_src,
                      const CgItem* d_items, int64_t n_items, int32_t default_flags,
                      double bound_cap, int64_t max_patch_px,
                      float* d_patches, CgBox* d_boxes,
                      void* d_workspace, size_t workspace_bytes, void* stream);

/* A7/A8: celeste.py:203-252, models.py:88-108.  Full-frame model
 * lambda = epsilon + sum_s patch_s and ll_img = sum(nelec*log(lambda) - lambda)
 * for every image; d_ll[n_img] (double).  d_model[i] (device float*, H*W) may
 * be NULL; pass d_model = NULL to skip all model output. */
int cg_field_loglik(CgImage* const* imgs, int32_t n_img,
                    const CgSource* d_srcs, int32_t n_src,
                    double bound_cap, int64_t max_patch_px,
                    double* d_ll, float* const* h_d_model,
                    void* d_workspace, size_t workspace_bytes, void* stream);

/* Batch of independent fields (models.py:75-79 loops the fields of a
 * CelesteBase): field f owns images [h_img_off[f], h_img_off[f+1]) of imgs[]
 * and sources [h_src_off[f], h_src_off[f+1]) of d_srcs; n_fields <= 16 and
 * h_img_off[n_fields] <= CG_MAX_IMAGES.  d_ll[h_img_off[n_fields]] receives
 * the per-image log-likelihoods; one prepare / render / gather launch each. */
int cg_fields_loglik(CgImage* const* imgs, const int32_t* h_img_off,
                     const CgSource* d_srcs, const int32_t* h_src_off, int32_t n_fields,
                     double bound_cap, int64_t max_patch_px,
                     double* d_ll, float* const* h_d_model,
                     void* d_workspace, size_t workspace_bytes, void* stream);

/* Library options.  CG_OPT_CULL (default 1): per-warp skipping of mixture
 * components whose largest value over the warp's pixels is below 2^-27 of a
 * lower bound of the mixture there (bounded relative error K*2^-27 on a rate). */
#define CG_OPT_CULL 1
/* CG_OPT_ROW_RECURRENCE (default 1): the full-frame kernel evaluates a component at
 * the 4 consecutive pixels of a lane from 2 exp2 and the geometric recurrence
 * e(x+1) = e(x) g(x), g(x+1) = g(x) 2^(2A) (about 1.5e-6 relative on a rate) where
 * that is provably safe; 0 = one exp2 per pixel everywhere. */
#define CG_OPT_ROW_RECURRENCE 2
int cg_set_option(int32_t option, int32_t value);

/* Likelihood modes of cg_items_loglik */
#define CG_LL_POISSON_BG 0  /* sources.py:6-12 with model = aux + rate, data = nelec crop, mask = image mask
                               (aux < 0: background = epsilon; sources.py:188-237) */
#define CG_LL_CONDITIONAL 1 /* sources.py:134-183: sum_{rate>0} log(rate)*z - band_flux*sum(psf_w); z = aux patch */
#define CG_LL_CONDITIONAL_SUM 2 /* celeste_galaxy_conditionals.py:15-42: sum z*log(rate) - rate over the box */

/* A13/A14/A15 building block: per-item log-likelihood over its box.  The aux
 * patch of item i (background or photon counts, float, box-shaped row-major)
 * is d_aux + aux_offsets[item.aux].  Output d_item_ll[n_items] (double). */
int cg_items_loglik(CgImage* const* imgs, int32_t n_img,
                    const CgSource* d_srcs, int32_t n_src,
                    const CgItem* d_items, int64_t n_items,
                    int32_t mode, double bound_cap, int64_t max_patch_px,
                    const float* d_aux, const int64_t* d_aux_offsets,
                    double* d_item_ll, CgBox* d_boxes, int32_t use_fp64,
                    void* d_workspace, size_t workspace_bytes, void* stream);

/* A15: Source.calculate_acceptance_logprob + the accept test
 * (sources.py:258, :275-306) for P proposals of one source.
 *   delta[p] = sum_img ll(prop_p) - sum_img ll(cur) + d_log_terms[p]
 *   accept[p] = d_log_u[p] < delta[p]
 * d_srcs[0] is the current state, d_srcs[1..P] the proposals.  h_boxes[n_img]
 * are the cached bounding boxes, d_bg + bg_offsets[img] the cached background
 * patches (sources.py:279-282).  Proposals whose margin |log_u - delta| is
 * below `guard` are re-scored in float64 on device (guard <= 0 disables).
 * d_n_guard (int32, may be NULL) receives the number of re-scored proposals.
 * d_bg may be NULL: the background is then epsilon. */
int cg_score_proposals(CgImage* const* imgs, int32_t n_img,
                       const CgSource* d_srcs, int32_t P,
                       const CgBox* h_boxes,
                       const float* d_bg, const int64_t* h_bg_offsets,
                       const double* d_log_terms, const double* d_log_u,
                       double guard,
                       double* d_delta, uint8_t* d_accept, int32_t* d_n_guard,
                       void* d_workspace, size_t workspace_bytes, void* stream);

/* sources.py:6-12 poisson_loglike(data, model_img, mask) on arbitrary device
 * arrays (float64): sum over {model > 0 and mask != 0} of log(model)*data - model.
 * d_mask may be NULL (all good).  d_out: one double. */
int cg_poisson_loglike(const double* d_data, const double* d_model, const double* d_mask,
                       int64_t n, double* d_out, void* stream);

/* cg_poisson_loglike without any allocation at call time.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph
 * capture, tight serving loops, harnesses that forbid it).  Same arguments, same result bits as
 * cg_poisson_loglike (same kernels, same launch shape); the block partials live in a file-scope
 * __device__ array of CG_POISSON_STATIC_MAX_BLOCKS doubles (8 KB per GPU that loads the library, held
 * until exit) whose device address cg_init resolves once per device (cudaGetSymbolAddress).  The entry
 * point calls no cudaMalloc / cudaFree / allocating routine, takes no lock and touches no host map.
 * Limit: n <= CG_POISSON_STATIC_MAX_N elements (a whole 2048 x 1489 frame is 3.05e6); larger n returns
 * CG_EINVAL with a message.  EVERY call shares the one array, which keeps the last call's partials
 * (nothing needs clearing: each slot is written before it is read): two calls that overlap on the GPU
 * (different streams, or graph branches without an edge between them) RACE -- serialise them, or use
 * cg_poisson_loglike, whose scratch is per stream. */
#define CG_POISSON_STATIC_MAX_BLOCKS 1024
#define CG_POISSON_STATIC_MAX_N (1024LL * 256LL * 64LL)   /* 16 777 216: at most 64 grid-stride trips per thread */
int cg_poisson_loglike_static(const double* d_data, const double* d_model, const double* d_mask,
                              int64_t n, double* d_out, void* stream);

/* ---- N1: celeste_sample_sources.pyx:61-156 ------------------------------- */
/* sample_source_counts for n_img images at once.  Item = src*n_img + img (the
 * layout cg_render_patches produces with d_items = NULL): d_boxes[item] is the
 * source's patch box, d_patches + item*max_patch_px its rate patch.  For every
 * pixel the photons int(nelec) are split over the sources whose box STRICTLY
 * contains the pixel (the reference's test, :53-54) and the sky rate epsilon
 * by conditional binomials in source order; d_samp (same layout as d_patches)
 * receives the per-source counts, d_noise_sum[n_img] the sky photons (:96,:112).
 * Random numbers: Philox-4x32-10 keyed by (seed, image, pixel, draw) -- the
 * reference's serial MT19937 stream cannot be followed in parallel, so parity
 * is distributional.  Workspace: cg_alloc_workspace_bytes(n_img, n_img*max_i
 * cg_field_tiles(H_i, W_i)). */
size_t cg_alloc_workspace_bytes(int32_t n_img, int64_t n_tiles);
int cg_sample_source_counts(CgImage* const* imgs, int32_t n_img, int32_t n_src,
                            const CgBox* d_boxes, const float* d_patches, int64_t max_patch_px,
                            uint64_t seed, float* d_samp, double* d_noise_sum,
                            void* d_workspace, size_t workspace_bytes, void* stream);
/* sample_binomial (:161-166): count draws of binomial(n, p) -> d_out (float64). */
int cg_binomial_draws(int64_t n, double p, uint64_t seed, int64_t count, double* d_out, void* stream);

/* ---- device-resident Gibbs sweep over a crowded field (BASELINE config 4) --- */
/* The per-sweep steps of CelesteBase.resample_model (models.py:75-83) for ALL sources of a field
 * with no host hop: every array below is a device array owned by the caller.  Items are the cross
 * product item = src * n_img + img; box-shaped per-item patches are row-major with stride
 * max_patch_px.  Workspaces: cg_sweep_workspace_bytes(kind, ...), n_tiles as for CG_WS_FIELD. */
#define CG_WS_SWEEP_ALLOC 4
#define CG_WS_SWEEP_SCORE 5
#define CG_WS_SWEEP_BACKGROUND 6
size_t cg_sweep_workspace_bytes(int32_t kind, int32_t n_src, int32_t n_img, int32_t P, int64_t max_patch_px,
                                int64_t n_tiles);

/* Field.resample_photons (models.py:123-160) = sample_source_photons_single_image_cython
 * (celeste_mcmc.py:98-150) + sample_source_counts (celeste_sample_sources.pyx:61-156) for the n_img
 * images of a field in ONE pass over the 16 x 16 image bins: each bin's sources are rendered into
 * registers and the pixel's photons split over them and the sky by conditional binomials -- there
 * is no rate-patch buffer.  Outputs: d_samp [n_items][max_patch_px] photons per source (every pixel
 * of a box is written), d_boxes [n_items] the boxes (footprints, radius capped by bound_cap),
 * d_noise_sum [n_img] sky photons, d_counts [n_items] = sum(d_samp[item]) and d_unit_sums [n_items] =
 * sum of the unit-flux patch over the box (the two statistics Source.resample_fluxes needs,
 * sources.py:340-342; either may be NULL).  Random numbers as cg_sample_source_counts. */
int cg_field_sample_photons(CgImage* const* imgs, int32_t n_img, const CgSource* d_srcs, int32_t n_src,
                            double bound_cap, int64_t max_patch_px, uint64_t seed, float* d_samp,
                            CgBox* d_boxes, double* d_noise_sum, double* d_counts, double* d_unit_sums,
                            void* d_workspace, size_t workspace_bytes, void* stream);

/* Source.resample_fluxes (sources.py:327-350) for every source: flux[band] ~ Gamma(a0 + sum of its
 * photons in the band, 1 / (b0 + sum_psf * kappa / calib)), written into d_srcs in place.  Sources
 * whose fluxes are functions of other parameters (CG_SRC_PLANCK stars, CG_SRC_EIGEN) are skipped.
 * Philox + Marsaglia-Tsang on the device: distributional parity with numpy.random.gamma. */
int cg_resample_fluxes(CgImage* const* imgs, int32_t n_img, CgSource* d_srcs, int32_t n_src,
                       const double* d_counts, const double* d_unit_sums, double a0, double b0, uint64_t seed,
                       void* stream);
/* models.py:156-160: epsilon_img ~ Gamma(a0 + noise_sum, 1 / (b0 + H W)), written into the image
 * descriptor ON THE DEVICE (d_eps_out [n_img] optional); cg_image_sync_epsilon refreshes the host
 * copy (synchronises the stream). */
int cg_resample_epsilon(CgImage* const* imgs, int32_t n_img, const double* d_noise_sum, double a0, double b0,
                        uint64_t seed, double* d_eps_out, void* stream);
int cg_image_sync_epsilon(CgImage* img, void* stream, double* h_epsilon);

/* P Gaussian perturbations of every source's state: d_states [n_src][P + 1], entry 0 = the current
 * state.  h_scales[6] = standard deviations of u (degrees), relative flux, theta, rho, phi
 * (degrees), relative sigma.  (No reference analogue: the reference draws one proposal at a time
 * on the host, experiments/rjump/test_rjmcmc.py.) */
int cg_make_proposals(const CgSource* d_cur, int32_t n_src, int32_t P, const double* h_scales, uint64_t seed,
                      CgSource* d_states, void* stream);

/* A13 / A15 for EVERY source of a field in one call: for source s and proposal p
 *   delta[s*P + p-1] = sum_img ll(states[s][p]) - sum_img ll(states[s][0]) + d_log_terms[s*P + p-1]
 *   accept = d_log_u < delta
 * over the boxes d_boxes[s * n_img + img], with
 *   mode CG_LL_POISSON_BG   ll = poisson_loglike(nelec crop, d_aux patch + rate, mask)  (sources.py:275-306;
 *                           d_aux = cached background patches, NULL = epsilon)
 *   mode CG_LL_CONDITIONAL  ll = Source.log_likelihood given the photon patches d_aux (sources.py:134-183).
 * Proposals within `guard` of their log-uniform are re-scored in float64 (list of
 * cg_sweep_guard_cap(n_src, P) entries; overflow poisons d_delta with NaN). */
int32_t cg_sweep_guard_cap(int32_t n_src, int32_t P);
int cg_score_proposals_multi(CgImage* const* imgs, int32_t n_img, const CgSource* d_states, int32_t n_src,
                             int32_t P, const CgBox* d_boxes, int32_t mode, const float* d_aux,
                             int64_t max_patch_px, const double* d_log_terms, const double* d_log_u,
                             double guard, double* d_delta, uint8_t* d_accept, int32_t* d_n_guard,
                             void* d_workspace, size_t workspace_bytes, void* stream);
/* The accept test of sources.py:258 applied in order: source s moves to its first accepted proposal
 * (exactly that many sequential MH steps: all P were drawn around the current state). */
int cg_select_proposals(const CgSource* d_states, int32_t n_src, int32_t P, const uint8_t* d_accept,
                        CgSource* d_cur, int32_t* d_n_accepted, void* stream);

/* generate_background_patch (sources.py:476-483) for every (source, image): epsilon + the model
 * patches of the active neighbours of (source, image) (CSR lists d_nbr_off [n_src * n_img + 1], d_nbr_idx
 * of SOURCE indices; the intersection test of sources.py:457-472 is the caller's) rendered on the source's cached box
 * d_boxes[s * n_img + img] (sides <= max_box_side) -> d_bg [n_items][max_patch_px]. */
int cg_background_patches(CgImage* const* imgs, int32_t n_img, const CgSource* d_srcs, int32_t n_src,
                          const CgBox* d_boxes, const int32_t* d_nbr_off, const int32_t* d_nbr_idx,
                          int64_t max_patch_px, int32_t max_box_side, float* d_bg,
                          void* d_workspace, size_t workspace_bytes, void* stream);

/* ---- A16: planck.py:94-158 ----------------------------------------------- */
/* out[i*5+band] = photons_expected_brightness(T[i], b[i], band); float64. */
int cg_planck_fluxes(const double* d_T, const double* d_b, int64_t n,
                     double* d_out, void* stream);

/* ---- A17: experiments/redshift/redshift_utils.py:304-323,
 *           quasar_infer_photometry.py:31-46 ------------------------------- */
/* For chain c: spec = w[c] . B (K x V), lam_obs = lam0*(1+z[c]),
 * mu[c] = project_to_bands(spec, lam_obs)*m[c]/(1+z[c]) -> d_mu[c*5+band];
 * if d_fluxes != NULL also ll[c] = -0.5*sum(ivar*(fluxes-mu)^2). */
int cg_eigen_fluxes(const double* d_z, const double* d_w, const double* d_m,
                    const double* d_B, const double* d_lam0,
                    int32_t n_chain, int32_t K, int32_t V,
                    const double* d_fluxes, const double* d_fluxes_ivar,
                    double* d_mu, double* d_ll, void* stream);

/* ---- N3: empirical priors, util/dists/mog.py:5-21,38-67, flux_prior.py:4-93 - */
#define CG_MOG_MAX_DIM 8
/* logp[n] = logsumexp_k( -0.5 (x_n-mu_k)^T icov_k (x_n-mu_k) - log(2 pi)
 *                        - 0.5 log(det_k) + log(pi_k) )      (mog_loglike, as written:
 * the -log(2 pi) term does not scale with D).  x [N][D], means [K][D], icovs
 * [K][D][D], dets [K], pis [K]; D <= CG_MOG_MAX_DIM; float64 device arrays.
 * d_resp (optional, [N][K]) receives the responsibilities exp(logprob_nk - logp_n). */
int cg_mog_logpdf(const double* d_x, int64_t N, int32_t D,
                  const double* d_means, const double* d_icovs, const double* d_dets,
                  const double* d_pis, int32_t K, double* d_logp, double* d_resp, void* stream);
/* Sufficient statistics of a MoG prior over the sources of this rank -- the
 * vector dist.SweepStats all-reduces (the M-step inputs of
 * experiments/empirical_priors/make_source_priors.py:76-110):
 *   out = [ sum_n r_nk (K) | sum_n r_nk c_n (K x D) | sum_n r_nk c_n c_n^T (K x D(D+1)/2, upper triangle) ]
 * resp [N][K], c [N][D].  Fixed-shape float64 tree per entry (bitwise reproducible). */
int cg_prior_stats(const double* d_resp, const double* d_c, int64_t N, int32_t K, int32_t D,
                   double* d_out, void* stream);

/* ---- N2: sampler loops on the device ------------------------------------------- */
/* Source.resample_location (sources.py:308-319: slicesample(u, location_likelihood, step_out=False),
 * util/infer/slicesample.py:89-247) for EVERY source in one launch: one thread block per source runs
 * the sampler's control flow and evaluates each conditional likelihood (Source.log_likelihood,
 * sources.py:134-183, over the n_img <= 8 photon patches d_photons[(s * n_img + img) * max_patch_px]
 * with boxes d_boxes) itself.  Random numbers are host-supplied in the reference's call order:
 * d_order [n_src][2] the shuffled coordinate order, d_uniforms / d_log_uniforms [n_src][n_uniforms]
 * the uniforms the sampler consumes one after the other and their logarithms.  Outputs per source:
 * the new location written into d_srcs (status 0 only), d_ll the likelihood there, d_n_used the
 * uniforms consumed (to advance the host generator by exactly that many), d_n_evals, d_n_refined
 * (optional: comparisons that fell within `guard` of a tie and were decided by float64 evaluations of
 * level and value inside the kernel), d_status: 0 done; 1 ran out of uniforms (redo the source on the
 * host path); 3 NaN likelihood; 4 shrank to zero (the reference raises).
 * Likelihood values, decisions and traces are bit-identical to the lock-step path (cg_items_loglik). */
int cg_slice_sample_locations(CgImage* const* imgs, int32_t n_img, CgSource* d_srcs, int32_t n_src,
                              const CgBox* d_boxes, const float* d_photons, int64_t max_patch_px,
                              const double* d_uniforms, const double* d_log_uniforms, const int32_t* d_order,
                              int32_t n_uniforms, double sigma, double guard, double* d_ll, int32_t* d_n_used,
                              int32_t* d_status, int32_t* d_n_evals, int32_t* d_n_refined, void* stream);

/* One HMC update of (theta, sigma, phi, rho, u, fluxes) of EVERY galaxy in one launch
 * (celeste_mcmc.py:281-323 hmc_sample_galaxy_params around util/infer/hmc.py:16-98, num_iter = 1): one
 * thread block per galaxy runs the n_steps leapfrog steps and all 11 (n_steps + 1) + 2 float64
 * likelihoods of the trajectory (galaxy_source_like / galaxy_source_like_grad,
 * celeste_galaxy_conditionals.py:15-88: theta and the fluxes analytically, sigma / phi / rho / u by
 * central differences of 1e-5) on the FIXED footprints d_boxes of the photon patches
 * d_photons[(s * n_img + img) * max_patch_px] (the reference re-derives the footprint per evaluation and
 * fails when its shape leaves that of Z_s).  Host-supplied randomness in the reference's draw order:
 * d_p_in [n_src][11] the momentum after the refresh (hmc.py:67), d_ke_curr [n_src] = sum(P P / mass) of
 * the momentum before it (hmc.py:64), d_log_u [n_src] the log of the acceptance uniform (hmc.py:73).
 * mass: 11 host doubles.  Outputs: the accepted state written into d_srcs, d_p_out [n_src][11] the
 * momentum to carry (hmc.py:79), d_h_out [n_src][2] = -H at the start / at the proposal, d_accepted,
 * d_status (0 ok, 3 non-finite: state unchanged, rejected).
 * d_sel (optional): n_src indices into d_srcs / d_boxes / d_photons -- the galaxies of a mixed source list;
 * the per-call arrays (d_p_in ... d_status) have one row per selected galaxy.  item_flags:
 * CG_ITEM_HEADER_CD for galaxy_source_like's Cython route (header CD matrix), 0 for the live route
 * (FitsImage.cd_at_pixel), which is what the field path and the Gibbs sweep render with. */
int cg_hmc_galaxy_params(CgImage* const* imgs, int32_t n_img, CgSource* d_srcs, int32_t n_src,
                         const int32_t* d_sel, int32_t item_flags,
                         const CgBox* d_boxes, const float* d_photons, int64_t max_patch_px,
                         const double* d_p_in, const double* d_ke_curr, const double* d_log_u, double step,
                         const double* mass, int32_t n_steps, double* d_p_out, double* d_h_out,
                         int32_t* d_accepted, int32_t* d_status, void* stream);

/* ---- per-sweep statistics for the all-reduce ------------------ */
/* Sufficient statistics of the empirical colour prior over the sources of one type, straight from
 * the device source array: colours = FluxColorMoG.to_colors(log fluxes) (util/dists/flux_prior.py:4-40),
 * responsibilities under the K-component mixture (cg_mog_logpdf), then the cg_prior_stats vector
 * d_out [K (1 + 5 + 15)] -- the M-step inputs of experiments/empirical_priors/make_source_priors.py:76-110.
 * d_logp [n_src] optional.  Scratch: cg_flux_prior_scratch_bytes(n_src, K). */
size_t cg_flux_prior_scratch_bytes(int64_t n_src, int32_t K);
int cg_flux_prior_stats(const CgSource* d_srcs, int64_t n_src, int32_t src_type, const double* d_means,
                        const double* d_icovs, const double* d_dets, const double* d_pis, int32_t K,
                        void* d_scratch, size_t scratch_bytes, double* d_logp, double* d_out, void* stream);
/* Packs the vector a rank contributes to the per-sweep all-reduce:
 *   [0] sum of d_ll[n_ll] (fixed order)  [1:6] sky photons per band (d_noise[i] belongs to band d_img_band[i])
 *   [6] stars  [7] galaxies  [8 : 8 + block] d_prior0  [8 + block : 8 + 2 block] d_prior1
 * written directly into the collective's send buffer d_vec (8 + 2 block doubles): one launch, no
 * host hop (models.py:75-79 loops the fields; the global quantities are sums over ranks). */
int cg_sweep_stats(const double* d_ll, int32_t n_ll, const double* d_noise, const int32_t* d_img_band,
                   int32_t n_noise, const CgSource* d_srcs, int64_t n_src, const double* d_prior0,
                   const double* d_prior1, int32_t block, double* d_vec, void* stream);

/* ---- instrumentation ------------------------------------------------------ */
/* Number of kernels launched by this library since cg_init (or last reset). */
int64_t cg_launch_count(int reset);
/* Component evaluations (pixel x component) of the last cg_field_loglik /
 * cg_render_patches / cg_items_loglik / cg_score_proposals call: reads a
 * device counter, synchronises `stream`. */
int64_t cg_last_eval_count(void* stream);
/* Lane-evaluations the render kernel actually issued for that call (after
 * component culling; counted per warp, idle lanes included). */
int64_t cg_last_issued_count(void* stream);

/* CUDA-event timing of every render_kernel launch (bench.py's roofline leg):
 * enable resets the record; read synchronises on the recorded events. */
int cg_profile_render(int enable);
int cg_profile_render_read(double* h_total_ms, int64_t* h_launches);
/* Measured pipe throughput on this device: which = 0 MUFU.EX2 results/s (the
 * XU roofline of the render kernels), 1 = FP32 FMA instructions/s, 2..5 / 12..15 =
 * MUFU.EX2 results/s of the kernels' inner-loop instruction MIX in isolation at 24 /
 * 64 warps per SM (2 full loop with shared-memory table, 3 table in registers,
 * 4 without row terms, 5 one FFMA per exp); 6 / 16 = component x pixel
 * evaluations/s of the row-recurrence loop (3 MUFU.EX2 + 2 FP32 + 13 packed FP32x2
 * instructions per 8 pixels) in isolation; 7 = FP32 FMA operations/s issued as packed
 * FFMA2 (2 per lane-instruction).  Synchronises. */
int cg_pipe_probe(int which, double* h_ops_per_s, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* CELESTE_B200_H */
