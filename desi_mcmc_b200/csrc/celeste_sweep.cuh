// celeste_sweep.cuh -- device-resident Gibbs sweep over a crowded field (BASELINE config 4).
// Included at the end of celeste_b200.cu (one translation unit: it reuses the field path's bin
// lists, ft_source_rates, the render kernels and the Philox / binomial samplers).
//
//   models.py:123-160 + celeste_mcmc.py:98-150 + celeste_sample_sources.pyx:61-156
//       cg_field_sample_photons   render + multinomial photon allocation per 16 x 16 image bin:
//                                 no rate-patch buffer, the rates live in registers
//   sources.py:327-350, models.py:156-160
//       cg_resample_fluxes / cg_resample_epsilon   gamma conditionals drawn on the device
//   sources.py:476-483
//       cg_background_patches     epsilon + the active neighbours rendered on every source's box
//   sources.py:134-183 / :275-306 for EVERY source at once
//       cg_make_proposals, cg_score_proposals_multi, cg_select_proposals

// ---------------------------------------------------------------------------
// random variates from the counter-based generator (distributional parity, as N1)
// ---------------------------------------------------------------------------
__device__ inline double philox_normal(Philox& g, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  const double u1 = g.uniform(a, b, c, d), u2 = g.uniform(a, b, c, d ^ 0x80000000u);
  return sqrt(-2. * log(u1)) * cospi(2. * u2);
}

// Marsaglia & Tsang (2000) for shape >= 1; shape < 1 through gamma(shape + 1) U^(1 / shape).
// numpy.random.gamma (the reference's draw, sources.py:349, models.py:159) samples the same law
// with the same method from MT19937; the stream cannot be followed in parallel.
__device__ inline double gamma_draw(Philox& g, double shape, uint32_t id0, uint32_t id1, uint32_t tag) {
  if (!(shape > 0.)) return 0.;
  double boost = 1.;
  if (shape < 1.) {
    boost = pow(g.uniform(id0, id1, 0xffffu, tag), 1. / shape);
    shape += 1.;
  }
  const double d = shape - 1. / 3., c = 1. / sqrt(9. * d);
  for (uint32_t it = 0; it < 256; ++it) {
    const double x = philox_normal(g, id0, id1, 2 * it, tag);
    double v = 1. + c * x;
    if (v <= 0.) continue;
    v = v * v * v;
    const double u = g.uniform(id0, id1, 2 * it + 1, tag);
    if (log(u) < .5 * x * x + d - d * v + d * log(v)) return d * v * boost;
  }
  return d * boost;
}

// ---------------------------------------------------------------------------
// photon allocation on the field path's bin lists
// ---------------------------------------------------------------------------
#define UNIT_FX 17592186044416.0        /* 2^44: fixed-point scale of the unit-flux patch sums */
struct AllocTileArgs {
  FieldArgs f;                 // tables, headers, bin lists in source order, class lists (bin_order_kernel)
  float* samp;                 // [n_items][max_patch_px] photons given to each source, box-shaped row-major
  long long max_patch_px;
  double* item_counts;         // [n_items] sum of samp (integers: exact in float64 whatever the order)
  unsigned long long* item_unit_fx;  // [n_items] sum of the unit-flux patch over the box, fixed point 2^-44
  unsigned long long seed;
};

// One WARP per 16 x 16 image bin, lane = 4 pixels of rows r and r + 8 (field_tile_kernel's layout).
// Pass 1 renders every source of the bin's list and sums, per pixel and in source order, the rates
// of the sources whose box STRICTLY contains the pixel (celeste_sample_sources.pyx:53-54) --
// sum_probs of sample_multinomial (:140-142).  Pass 2 renders them again (same instructions, same
// values) and draws the conditional binomials in source order (:144-151); what is left is sky
// (:152).  Pixels no box contains are all sky (:94-96).  The draws are keyed by (seed, image,
// pixel, draw index) exactly as sample_counts_kernel keys them.
__global__ void __launch_bounds__(256, 2) alloc_tile_kernel(AllocTileArgs A) {
  const FieldArgs& a = A.f;
  __shared__ __align__(16) float s_wtab[8][(CG_MAX_COMP + 1) * 8];
  __shared__ int s_csize[FT_BUCKETS];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  float* const wtab = s_wtab[warp];
  if (threadIdx.x < FT_BUCKETS) s_csize[threadIdx.x] = a.sched[threadIdx.x];
  __syncthreads();
  int total = 0;
#pragma unroll
  for (int b = 0; b < FT_BUCKETS; ++b) total += s_csize[b];
  unsigned long long issued = 0;
  Philox g;
  g.key[0] = (uint32_t)A.seed; g.key[1] = (uint32_t)(A.seed >> 32);

  for (;;) {
    int u = 0;
    if (lane == 0) u = atomicAdd(&a.sched[FT_BUCKETS + 1], 1);
    u = __shfl_sync(0xffffffffu, u, 0);
    if (u >= total) break;
    int cls = FT_BUCKETS - 1, v = u;                   // heaviest class first
    while (v >= s_csize[cls]) { v -= s_csize[cls]; --cls; }
    const int2 dd = a.order[(long long)cls * a.tiles_total + v];
    const int dsc = dd.x, gy = dd.y;
    const int cnt = gy & 63, gbin = gy >> 6;
    const int img = dsc >> 24;
    const DevImage& im = *a.imgs.p[img];
    const int H = im.H, W = im.W;
    FtGeom q;
    q.X0 = ((dsc >> 12) & 0xfff) * BIN_W; q.Y0 = (dsc & 0xfff) * BIN_H;
    q.x = q.X0 + (lane & 3) * 4; q.yA = q.Y0 + (lane >> 2); q.yB = q.yA + 8;

    float d[2][4];
    bool inimg[2][4];
#pragma unroll
    for (int r = 0; r < 2; ++r) {
      const int y = r ? q.yB : q.yA;
#pragma unroll
      for (int p = 0; p < 4; ++p) {
        inimg[r][p] = (y < H) && (q.x + p < W);
        d[r][p] = inimg[r][p] ? __ldg(im.nelec + (long long)y * W + q.x + p) : 0.f;
      }
    }
    double noise = 0.0;
    if (cnt == 0) {
#pragma unroll
      for (int r = 0; r < 2; ++r)
#pragma unroll
        for (int p = 0; p < 4; ++p) noise += (double)d[r][p];
    } else {
      // candidates: the bin's list in source order, or (overflowed list) every source of the image
      const bool scan = cnt > BIN_CAP;
      const int ns = scan ? a.n_src[img] : 0;
      const int rounds = scan ? (ns + 31) / 32 : 1;
      double tot[2][4];
      unsigned any = 0u;                       // bit r * 4 + p: some box strictly contains the pixel
      int left[2][4];                          // photons not yet given away (nelec < 2^24: exact in fp32)
      unsigned long long draw = 0ull;          // 8 bits per pixel: owners seen so far = index of the next draw
#pragma unroll
      for (int r = 0; r < 2; ++r)
#pragma unroll
        for (int p = 0; p < 4; ++p) { tot[r][p] = 0.0; left[r][p] = (int)d[r][p]; }

#pragma unroll 1
      for (int pass = 0; pass < 2; ++pass) {
        if (pass == 1) {
          const double eps = im.epsilon;
#pragma unroll
          for (int r = 0; r < 2; ++r)
#pragma unroll
            for (int p = 0; p < 4; ++p) {
              tot[r][p] += eps;                                        // model_fluxes[S-1] = epsilon (:104)
              if (!((any >> (r * 4 + p)) & 1u)) { noise += (double)d[r][p]; left[r][p] = 0; }   // (:94-96)
            }
        }
#pragma unroll 1
        for (int rd = 0; rd < rounds; ++rd) {
          int e = 0x7fffffff;
          bool valid = false;
          if (!scan) {
            if (lane < cnt) { e = __ldg(a.tile_list + (long long)gbin * BIN_CAP + lane); valid = true; }
          } else {
            const int sidx = rd * 32 + lane;
            if (sidx < ns) {
              const int it = a.item_base[img] + sidx * a.item_stride[img];
              const ItemHdr* hp = a.hdr + it;
              const int K = hp->K;
              e = it * 64 + K;
              valid = (K > 0) && hp->x0 < q.X0 + BIN_W && hp->x1 > q.X0 && hp->y0 < q.Y0 + BIN_H && hp->y1 > q.Y0;
            }
          }
          const unsigned vm = __ballot_sync(0xffffffffu, valid);
          const int nv = __popc(vm);
          const int rk = __popc(vm & ((1u << lane) - 1u));
#pragma unroll 1
          for (int r0 = 0; r0 < nv; ++r0) {
            const unsigned m = __ballot_sync(0xffffffffu, valid && rk == r0);
            const int ee = __shfl_sync(0xffffffffu, e, __ffs(m) - 1);
            const int item = ee >> 6;
            const ItemHdr* hp = a.hdr + item;                           // warp-uniform address: one broadcast load
            const int4 bx = *reinterpret_cast<const int4*>(hp);
            const int2 o = *reinterpret_cast<const int2*>(&hp->ox);
            const float fl = (float)hp->flux;
            float acc[2][4];
            if (!ft_source_rates(a, item, ee & 63, bx, o, q, lane, wtab, acc[0], acc[1], issued)) continue;
            const int wbox = bx.y - bx.x;
            double got_sum = 0.0;
            float unit_sum = 0.f;
#pragma unroll
            for (int r = 0; r < 2; ++r) {
              const int y = r ? q.yB : q.yA;
              const bool iny = y >= bx.z && y < bx.w;
#pragma unroll
              for (int p = 0; p < 4; ++p) {
                const int x = q.x + p;
                const bool inbox = iny && x >= bx.x && x < bx.y;
                if (!inbox) continue;
                const bool own = x > bx.x && y > bx.z;                  // strict on the low sides (:53-54)
                const float rate = acc[r][p] * fl;
                if (pass == 0) {
                  if (own) { tot[r][p] += (double)rate; any |= 1u << (r * 4 + p); }
                } else {
                  double got = 0.0;
                  if (own && left[r][p] > 0) {
                    const uint32_t pix = (uint32_t)((long long)y * W + x);
                    const double U = g.uniform(pix, (uint32_t)img, (uint32_t)((draw >> (8 * (r * 4 + p))) & 255ull), 0xa110cu);
                    got = binomial_from_uniform((long long)left[r][p], (double)rate / tot[r][p], U);
                    left[r][p] -= (int)got;
                  }
                  if (own) { draw += 1ull << (8 * (r * 4 + p)); tot[r][p] -= (double)rate; }
                  CG_CHECK((long long)(y - bx.z) * wbox + (x - bx.x) < A.max_patch_px && y >= bx.z && x >= bx.x);
                  A.samp[(long long)item * A.max_patch_px + (long long)(y - bx.z) * wbox + (x - bx.x)] = (float)got;
                  got_sum += got;
                  unit_sum += acc[r][p];
                }
              }
            }
            if (pass == 1) {
              got_sum = warp_sum(got_sum);
#pragma unroll
              for (int o2 = 16; o2 > 0; o2 >>= 1) unit_sum += __shfl_down_sync(0xffffffffu, unit_sum, o2);
              if (lane == 0) {
                if (got_sum != 0.0) atomicAdd(&A.item_counts[item], got_sum);
                // fixed point: integer adds commute, so the sum is bitwise reproducible
                atomicAdd(&A.item_unit_fx[item], (unsigned long long)llrint((double)unit_sum * UNIT_FX));
              }
            }
          }
        }
      }
#pragma unroll
      for (int r = 0; r < 2; ++r)
#pragma unroll
        for (int p = 0; p < 4; ++p) noise += (double)left[r][p];      // samples[S-1] += unsampled (:152)
    }
    noise = warp_sum(noise);
    if (lane == 0) a.block_partials[gbin] = noise;
  }
  if (lane == 0 && issued) atomicAdd(&a.cnt->issued, issued);
}

__global__ void unit_sums_kernel(const unsigned long long* fx, long long n, double* out) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = (double)fx[i] * (1.0 / UNIT_FX);
}

// ---------------------------------------------------------------------------
// gamma conditionals (sources.py:327-350, models.py:156-160)
// ---------------------------------------------------------------------------
__global__ void resample_fluxes_kernel(CgSource* srcs, int n_src, int n_img, ImgPtrs imgs,
                                       const double* item_counts, const double* item_unit, double a0,
                                       double b0, unsigned long long seed) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n_src * CG_NUM_BANDS) return;
  const int s = t / CG_NUM_BANDS, band = t - s * CG_NUM_BANDS;
  CgSource& src = srcs[s];
  // fluxes that are functions of other parameters (temperature, eigen-spectrum weights) are not free
  if ((src.flags & CG_SRC_EIGEN) || (src.type == 0 && (src.flags & CG_SRC_PLANCK) && src.t != 0.0)) return;
  double counts = 0.0, psf = 0.0;
  for (int b = 0; b < n_img; ++b) {
    const DevImage& im = *imgs.p[b];
    if (im.band != band) continue;
    const long long it = (long long)s * n_img + b;
    counts += item_counts[it];                                         // :340
    psf += item_unit[it] * im.kappa / im.calib;                        // :342
  }
  Philox g;
  g.key[0] = (uint32_t)seed; g.key[1] = (uint32_t)(seed >> 32);
  src.flux[band] = gamma_draw(g, a0 + counts, (uint32_t)s, (uint32_t)band, 0xf1u) / (b0 + psf);   // :346-349
}

__global__ void resample_epsilon_kernel(ImgPtrs imgs, int n_img, const double* noise_sum, double a0, double b0,
                                        unsigned long long seed, double* eps_out) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= n_img) return;
  DevImage* im = const_cast<DevImage*>(imgs.p[b]);
  Philox g;
  g.key[0] = (uint32_t)seed; g.key[1] = (uint32_t)(seed >> 32);
  const double a_n = a0 + noise_sum[b], b_n = b0 + (double)im->H * (double)im->W;     // models.py:157-158
  const double e = gamma_draw(g, a_n, (uint32_t)b, 0u, 0xe9u) / b_n;
  im->epsilon = e;
  if (eps_out) eps_out[b] = e;
}

// ---------------------------------------------------------------------------
// proposals for every source
// ---------------------------------------------------------------------------
struct ProposalScales { double u_deg, flux_rel, theta, rho, phi_deg, sigma_rel; };

// states[s][0] = current state, states[s][1..P] = Gaussian perturbations of it
__global__ void make_proposals_kernel(const CgSource* cur, int n_src, int P, ProposalScales sc,
                                      unsigned long long seed, CgSource* states) {
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= (long long)n_src * (P + 1)) return;
  const int s = (int)(t / (P + 1)), p = (int)(t - (long long)s * (P + 1));
  CgSource x = cur[s];
  if (p > 0) {
    Philox g;
    g.key[0] = (uint32_t)seed; g.key[1] = (uint32_t)(seed >> 32);
    uint32_t k = 0;
    auto nrm = [&]() { return philox_normal(g, (uint32_t)s, (uint32_t)p, k++, 0x9e0u); };
    x.u[0] += sc.u_deg * nrm();
    x.u[1] += sc.u_deg * nrm();
    const bool free_flux = !((x.flags & CG_SRC_EIGEN) || (x.type == 0 && (x.flags & CG_SRC_PLANCK) && x.t != 0.0));
    for (int b = 0; b < CG_NUM_BANDS; ++b) {
      const double z = nrm();
      if (free_flux) x.flux[b] *= (1. + sc.flux_rel * z);
    }
    if (x.type == 1) {
      x.theta = fmin(fmax(x.theta + sc.theta * nrm(), .01), .99);
      x.rho = fmin(fmax(x.rho + sc.rho * nrm(), .05), 1.);
      x.phi += sc.phi_deg * nrm();
      x.sigma *= (1. + sc.sigma_rel * nrm());
    }
  }
  states[t] = x;
}

// item (s, p, b) -> index (s (P+1) + p) n_img + b : fixed box of (s, b), auxiliary patch of (s, b)
__global__ void score_items_kernel(const CgBox* boxes, int n_src, int P, int n_img, int flags, int use_aux,
                                   CgItem* items) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const long long n = (long long)n_src * (P + 1) * n_img;
  if (i >= n) return;
  const int b = (int)(i % n_img);
  const long long sp = i / n_img;
  const int s = (int)(sp / (P + 1));
  const CgBox bx = boxes[(long long)s * n_img + b];
  CgItem it;
  it.src = (int)sp; it.img = b; it.flags = flags | CG_ITEM_FIXED_BOX;
  it.box[0] = bx.x0; it.box[1] = bx.x1; it.box[2] = bx.y0; it.box[3] = bx.y1;
  it.aux = use_aux ? s * n_img + b : -1;
  items[i] = it;
}

__global__ void score_multi_finalize_kernel(const double* item_ll, int n_src, int P, int n_img,
                                            const double* log_terms, const double* log_u, double guard,
                                            double* delta, uint8_t* accept, Counters* cnt, int* guard_list,
                                            int guard_cap) {
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;   // (s, p - 1)
  if (t >= (long long)n_src * P) return;
  const int s = (int)(t / P), p = (int)(t - (long long)s * P) + 1;
  const double* c = item_ll + (long long)s * (P + 1) * n_img;
  double cur = 0, prop = 0;
  for (int b = 0; b < n_img; ++b) { cur += c[b]; prop += c[(long long)p * n_img + b]; }
  const double d = (prop - cur) + (log_terms ? log_terms[t] : 0.0);
  const double lu = log_u[t];
  delta[t] = d;
  accept[t] = (lu < d) ? 1 : 0;
  if (guard > 0 && fabs(lu - d) < guard) {
    const int pos = atomicAdd(&cnt->guard_count, 1);
    if (pos < guard_cap) guard_list[pos] = (int)t;
    else atomicExch(&cnt->err_flag, 2);
  }
}

// float64 re-score of the guard band: entry g of the list -> items [g][cur | prop][b]
__global__ void rescore_items_kernel(const CgBox* boxes, const Counters* cnt1, const int* guard_list, int guard_cap,
                                     int P, int n_img, int flags, int use_aux, CgItem* items) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (long long)guard_cap * 2 * n_img) return;
  const int b = (int)(i % n_img);
  const int which = (int)((i / n_img) & 1);
  const int gi = (int)(i / (2 * n_img));
  CgItem it;
  it.src = 0; it.img = b; it.flags = flags | CG_ITEM_FIXED_BOX; it.aux = -1;
  it.box[0] = it.box[1] = it.box[2] = it.box[3] = 0;                   // empty: renders nothing
  if (gi < min(cnt1->guard_count, guard_cap)) {
    const int t = guard_list[gi];
    const int s = t / P, p = t - s * P + 1;
    const CgBox bx = boxes[(long long)s * n_img + b];
    it.src = s * (P + 1) + (which ? p : 0);
    it.box[0] = bx.x0; it.box[1] = bx.x1; it.box[2] = bx.y0; it.box[3] = bx.y1;
    it.aux = use_aux ? s * n_img + b : -1;
  }
  items[i] = it;
}

__global__ void score_multi_rescore_kernel(const double* item_ll2, int n_img, const Counters* cnt1,
                                           const int* guard_list, int guard_cap, const double* log_terms,
                                           const double* log_u, double* delta, uint8_t* accept) {
  const int gi = blockIdx.x * blockDim.x + threadIdx.x;
  if (gi >= min(cnt1->guard_count, guard_cap)) return;
  const int t = guard_list[gi];
  const double* c = item_ll2 + (long long)gi * 2 * n_img;
  double cur = 0, prop = 0;
  for (int b = 0; b < n_img; ++b) { cur += c[b]; prop += c[n_img + b]; }
  const double d = (prop - cur) + (log_terms ? log_terms[t] : 0.0);
  delta[t] = d;
  accept[t] = (log_u[t] < d) ? 1 : 0;
}

// a device error (Counters.err_flag) poisons the scores: no caller can miss it
__global__ void score_poison_kernel(const Counters* c1, const Counters* c2, long long n, double* delta) {
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (t < n && (c1->err_flag || (c2 && c2->err_flag))) delta[t] = (double)NAN;
}

// Speculative MH: the P proposals of a source were all drawn around its current state, so taking
// them in order until the first acceptance is exactly p* sequential MH steps (the proposals after
// p* would have to be re-drawn around the new state and are discarded).
__global__ void select_proposals_kernel(const CgSource* states, int n_src, int P, const uint8_t* accept,
                                        CgSource* cur, int* n_accepted) {
  const int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= n_src) return;
  for (int p = 0; p < P; ++p)
    if (accept[(long long)s * P + p]) {
      cur[s] = states[(long long)s * (P + 1) + p + 1];
      if (n_accepted) atomicAdd(n_accepted, 1);
      return;
    }
}

// ---------------------------------------------------------------------------
// background patches (sources.py:476-483)
// ---------------------------------------------------------------------------
struct BgArgs {
  FieldArgs f;               // hdr / tab of the (source, image) items, item = src * n_img + img
  const CgBox* boxes;        // [n_src * n_img] the cached bounding boxes (sources.py:434-455)
  const int* nbr_off;        // [n_src * n_img + 1] CSR of each (source, image)'s active neighbours (sources.py:457-472)
  const int* nbr_idx;
  float* bg;                 // [n_src * n_img][max_patch_px]
  long long max_patch_px;
  int n_src, n_img, tiles_side;   // tiles_side^2 16 x 16 tiles cover the largest box
};

// One warp per 16 x 16 tile of a target box: the active neighbours are rendered ON THE TARGET'S BOX
// (not on their own footprints: compute_model_patch(image, xlim, ylim) of sources.py:481) and summed
// in list order on top of epsilon.
__global__ void __launch_bounds__(256, 2) bg_tile_kernel(BgArgs A) {
  const FieldArgs& a = A.f;
  __shared__ __align__(16) float s_wtab[8][(CG_MAX_COMP + 1) * 8];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  float* const wtab = s_wtab[warp];
  unsigned long long issued = 0;
  const int tiles_per = A.tiles_side * A.tiles_side;
  const long long n_units = (long long)A.n_src * A.n_img * tiles_per;
  const long long n_warps = (long long)gridDim.x * (blockDim.x >> 5);
  for (long long u = (long long)blockIdx.x * (blockDim.x >> 5) + warp; u < n_units; u += n_warps) {
    const long long target = u / tiles_per;
    const int tile = (int)(u - target * tiles_per);
    const int s = (int)(target / A.n_img), b = (int)(target - (long long)s * A.n_img);
    const CgBox box = A.boxes[target];
    FtGeom q;
    q.X0 = box.x0 + (tile % A.tiles_side) * BIN_W; q.Y0 = box.y0 + (tile / A.tiles_side) * BIN_H;
    if (q.X0 >= box.x1 || q.Y0 >= box.y1) continue;
    q.x = q.X0 + (lane & 3) * 4; q.yA = q.Y0 + (lane >> 2); q.yB = q.yA + 8;
    const float eps = (float)a.imgs.p[b]->epsilon;
    float lam[2][4];
#pragma unroll
    for (int r = 0; r < 2; ++r)
#pragma unroll
      for (int p = 0; p < 4; ++p) lam[r][p] = eps;
    const int4 bx = make_int4(box.x0, box.x1, box.y0, box.y1);
    for (int j = A.nbr_off[target]; j < A.nbr_off[target + 1]; ++j) {
      const int item = A.nbr_idx[j] * A.n_img + b;
      CG_CHECK(A.nbr_idx[j] >= 0 && A.nbr_idx[j] < A.n_src);
      const ItemHdr* hp = a.hdr + item;
      const int2 o = *reinterpret_cast<const int2*>(&hp->ox);
      ft_source(a, item, hp->K, bx, o, (float)hp->flux, q, lane, wtab, lam, issued);
    }
    const int wbox = box.x1 - box.x0;
    CG_CHECK((long long)(box.y1 - box.y0) * wbox <= A.max_patch_px);
    float* out = A.bg + target * A.max_patch_px;
#pragma unroll
    for (int r = 0; r < 2; ++r) {
      const int y = r ? q.yB : q.yA;
      if (y >= box.y1) continue;
#pragma unroll
      for (int p = 0; p < 4; ++p)
        if (q.x + p < box.x1) out[(long long)(y - box.y0) * wbox + (q.x + p - box.x0)] = lam[r][p];
    }
  }
  if (lane == 0 && issued) atomicAdd(&a.cnt->issued, issued);
}

// ---------------------------------------------------------------------------
// N2: the location slice sampler of every source in ONE launch
// ---------------------------------------------------------------------------
// Source.resample_location (sources.py:308-319) = slicesample(u, location_likelihood, step_out=False,
// compwise) (util/infer/slicesample.py:89-247).  One CTA per source, one warp per image: the CTA runs
// the sampler's whole control flow -- per direction: bracket [upper - sigma, upper], slice level
// log(rand) + ll(0), shrink until a point above the level is drawn -- and evaluates every likelihood
// it needs itself: each warp builds its image's component table in shared memory (item_table, the
// float64 code prepare_kernel runs), renders the fixed box against the source's photon patch
// (render_warp_tile<MODE_COND>) and the CTA sums the images in order.  The same device functions, the
// same order of sums as the lock-step path (cg_items_loglik): the likelihood values, hence the
// decisions and the trace, are bit-identical to it.
// Random numbers are HOST-supplied, in the reference's call order: ord[s] = the shuffled coordinate
// order, uni[s][0..B) = the uniforms the sampler draws one after the other (log_uni = their float64
// logarithms, taken on the host so that the slice level matches numpy's).  A source that runs out of
// uniforms or meets a NaN stops with a status: the host redoes that source with the lock-step sampler
// from the same RNG state.  A comparison inside the guard band is decided in the kernel by float64
// evaluations of level and value (item_table<double> / render_warp_tile<double>: the use_fp64 path of
// cg_items_loglik), exactly as compat.slicesample does on the host.
#define SLICE_MAX_IMG 8
struct SliceArgs {
  ImgPtrs imgs;
  int n_img, n_src, B;
  CgSource* srcs;              // in: current state; out: u replaced for status 0
  const CgBox* boxes;          // [n_src * n_img] the photon patches' boxes
  const float* aux;            // photon patches, item-major, stride aux_stride
  long long aux_stride;
  const double* uni;           // [n_src][B]
  const double* log_uni;       // [n_src][B]
  const int* ord;              // [n_src][2]
  double sigma, guard;
  double* ll_out;              // [n_src] likelihood at the accepted point
  int* n_used;                 // [n_src] uniforms consumed
  int* status;                 // [n_src] 0 ok, 1 out of uniforms, 3 NaN, 4 shrank to zero
  int* n_evals;                // [n_src] likelihood evaluations
  int* n_refined;              // [n_src] comparisons decided in float64 (may be NULL)
  const FilterTables* filters;
  const double* flux_table;
  long long flux_rows;
  const LogTable* logtab;
  int cull;
};

#define SLICE_WARP_BYTES (CG_MAX_COMP * TAB_W * 8 + (CG_MAX_COMP + 1) * 8 * 4)   /* float64-sized table + compacted table */
// what one warp (one image of one source) carries from one likelihood evaluation to the next
struct SliceCtx {
  const DevImage* im;
  CgSource src;
  ItemHdr h;
  unsigned char* my;           // this warp's SLICE_WARP_BYTES of shared memory
  double* s_ll;                // [SLICE_MAX_IMG] the images' terms of the current evaluation
  TileArgs ra;
  const FilterTables* filters;
  const double* flux_table;
  long long flux_rows;
  const LogTable* logtab;
  unsigned long long issued;
  int lane, b, n_img, wbox, hbox, cpr, hp, n_lu, n_strips, evals;
  float r_band;
};

// Source.log_likelihood(u = x) summed over the images (sources.py:134-183); real = float: the fp32 path
// of cg_items_loglik; real = double: its use_fp64 path (the guard band's evaluation).  Out of line: the
// float64 instantiation is rare (a few comparisons per hundred sources) and must not set the kernel's
// register count.
template <typename real>
__device__ __noinline__ double slice_loglik(SliceCtx& c, double x0, double x1) {
  real* const s_tab = reinterpret_cast<real*>(c.my);
  float* const s_wtab = reinterpret_cast<float*>(c.my + CG_MAX_COMP * TAB_W * 8);
  const DevImage& im = *c.im;
  c.src.u[0] = x0; c.src.u[1] = x1;
  __syncthreads();                   // every warp is done with the previous evaluation's tables and s_ll
  ItemGeom g = item_table<real>(im, c.src, c.h.flags, c.lane, s_tab);
  const double flux = item_flux(im, c.src, c.h.flags, c.lane, g.star, c.filters, c.flux_table, c.flux_rows, g);
  c.h.ox = g.ox; c.h.oy = g.oy; c.h.K = g.K; c.h.flux = flux; c.h.norm_term = flux * im.psf_wsum;
  __syncwarp();
  double t = 0.0;
  for (int wt = 0; wt < c.n_strips; ++wt)         // the order items_finalize_kernel sums the strips in
    t += render_warp_tile<real, MODE_COND>(c.ra, c.h, 0, s_tab, s_wtab, wt, c.lane, c.wbox, c.hbox, c.cpr, c.hp,
                                           c.n_lu, c.r_band, c.im, (real)flux, c.logtab, c.issued);
  if (c.lane == 0) c.s_ll[c.b] = t - c.h.norm_term;   // CG_LL_CONDITIONAL (sources.py:174)
  __syncthreads();
  double tot = 0.0;
  for (int k = 0; k < c.n_img; ++k) tot += c.s_ll[k];  // image order: the lock-step driver's sum
  ++c.evals;
  return tot;
}

__global__ void __launch_bounds__(32 * SLICE_MAX_IMG, 2) slice_location_kernel(SliceArgs A) {
  extern __shared__ __align__(16) unsigned char s_dyn[];           // n_img x SLICE_WARP_BYTES
  __shared__ double s_ll[SLICE_MAX_IMG];
  const int s = blockIdx.x;
  if (s >= A.n_src) return;
  SliceCtx c;
  c.lane = threadIdx.x & 31; c.b = threadIdx.x >> 5;               // warp b <-> image b
  c.im = A.imgs.p[c.b];
  const DevImage& im = *c.im;
  c.my = s_dyn + (size_t)c.b * SLICE_WARP_BYTES;
  c.s_ll = s_ll;
  c.src = A.srcs[s];
  c.n_img = A.n_img;
  const CgBox box0 = A.boxes[(long long)s * A.n_img + c.b];
  // the fixed box, clipped to the image as prepare_kernel clips CG_ITEM_FIXED_BOX
  ItemHdr& h = c.h;
  h.x0 = max(box0.x0, 0); h.y0 = max(box0.y0, 0); h.x1 = min(box0.x1, im.W); h.y1 = min(box0.y1, im.H);
  if (h.x1 <= h.x0 || h.y1 <= h.y0) { h.x1 = h.x0; h.y1 = h.y0; }
  h.img = c.b; h.aux = s * A.n_img + c.b; h.flags = CG_ITEM_FLUX_BAND | CG_ITEM_FIXED_BOX; h.strip_start = 0;
  c.wbox = h.x1 - h.x0; c.hbox = h.y1 - h.y0;
  c.cpr = (c.wbox + PIX_PER_THREAD - 1) / PIX_PER_THREAD;
  c.hp = (c.hbox / (2 * TILE_R)) * TILE_R + ((c.hbox % (2 * TILE_R)) + 1) / 2;
  c.n_lu = c.cpr * c.hp;
  c.n_strips = (int)(((long long)c.hp * c.cpr + 31) / 32);
  c.r_band = __frcp_rn((float)max(c.cpr * TILE_R, 1));
  c.ra.cull = A.cull; c.ra.cond_sum = 0; c.ra.patches = nullptr; c.ra.max_patch_px = 0;
  c.ra.aux = A.aux; c.ra.aux_stride = A.aux_stride;
  c.filters = A.filters; c.flux_table = A.flux_table; c.flux_rows = A.flux_rows; c.logtab = A.logtab;
  c.issued = 0; c.evals = 0;
  int refined = 0;
  const CgSource& src = c.src;

  const double* U = A.uni + (long long)s * A.B;
  const double* LU = A.log_uni + (long long)s * A.B;
  double x[2] = {src.u[0], src.u[1]};
  int idx = 0, st = 0;
  double ll_new = 0.0;
  for (int dd = 0; dd < 2 && st == 0; ++dd) {
    const int d = A.ord[s * 2 + dd];
    if (idx + 2 > A.B) { st = 1; break; }
    double upper = A.sigma * U[idx++];                      // slicesample.py:155-156
    double lower = upper - A.sigma;
    const double log_u = LU[idx++];
    const double llh_s = log_u + slice_loglik<float>(c, x[0], x[1]);   // :159
    double level64 = 0.0;
    bool have64 = false;
    for (;;) {                                              // :181-198 (step_out = False)
      if (idx >= A.B) { st = 1; break; }
      const double z = (upper - lower) * U[idx++] + lower;
      const double xt0 = (d == 0 ? 1.0 : 0.0) * z + x[0], xt1 = (d == 1 ? 1.0 : 0.0) * z + x[1];
      ll_new = slice_loglik<float>(c, xt0, xt1);
      if (ll_new != ll_new) { st = 3; break; }
      bool above = ll_new > llh_s;
      if (A.guard > 0. && fabs(ll_new - llh_s) < A.guard) {
        // inside the guard band: level and value again in float64 (compat.slicesample's `above`)
        if (!have64) { level64 = log_u + slice_loglik<double>(c, x[0], x[1]); have64 = true; }
        above = slice_loglik<double>(c, xt0, xt1) > level64;
        ++refined;
      }
      if (above) { x[0] = xt0; x[1] = xt1; break; }
      else if (z < 0) lower = z;
      else if (z > 0) upper = z;
      else { st = 4; break; }
    }
  }
  if (threadIdx.x == 0) {
    A.status[s] = st;
    A.n_used[s] = idx;
    A.n_evals[s] = c.evals;
    A.ll_out[s] = ll_new;
    if (A.n_refined) A.n_refined[s] = refined;
    if (st == 0) { A.srcs[s].u[0] = x[0]; A.srcs[s].u[1] = x[1]; }
  }
}

// ---------------------------------------------------------------------------
// N2: one HMC update of every galaxy's (shape, location, fluxes) in ONE launch
// ---------------------------------------------------------------------------
// celeste_mcmc.py:281-323 (hmc_sample_galaxy_params) around util/infer/hmc.py:16-98 with num_iter = 1:
// x = [theta, sigma, phi, rho, u0, u1, b_u .. b_z]; energy = -galaxy_source_like
// (celeste_galaxy_conditionals.py:15-42), gradient = galaxy_source_like_grad (:44-88): theta and the
// fluxes analytically, sigma / phi / rho / u by central differences of 1e-5 -- ten more likelihoods per
// gradient, num_steps + 1 gradients per trajectory.  One thread block per galaxy, one warp per image.
// Every likelihood is evaluated in float64 straight from the float64 component table (item_table<double>,
// the table prepare_kernel builds) on the FIXED footprint of the photon patch: no culling, no recurrence,
// exp2 / log in float64 -- the differences are divided by 2e-5.  The momentum after the refresh and the
// logarithm of the acceptance uniform come from the host generator (hmc.py:62,67,73 draw order).
#define HMC_DIM 11
struct HmcArgs {
  ImgPtrs imgs;
  int n_img, n_src, n_steps;
  int item_flags;              // CG_ITEM_HEADER_CD (the Cython route of galaxy_source_like) or 0 (cd_at_pixel, the live route)
  const int* sel;              // optional [n_src] indices into srcs / boxes / patches (the galaxies of a mixed list)
  CgSource* srcs;              // in: current state; out: the accepted state
  const CgBox* boxes;          // [.. * n_img] footprints of the photon patches
  const float* aux;            // photon patches, item-major, stride aux_stride
  long long aux_stride;
  const double* p_in;          // [n_src][HMC_DIM] momentum after the refresh (hmc.py:67)
  const double* ke_curr;       // [n_src] sum(P P / mass) of the momentum BEFORE the refresh (hmc.py:64 takes ll_curr there)
  const double* log_u;         // [n_src] log of the acceptance uniform (hmc.py:73)
  double eps;
  double mass[HMC_DIM];
  double* p_out;               // [n_src][HMC_DIM] momentum carried to the next call (hmc.py:79)
  double* h_out;               // [n_src][2] -H at the start and at the proposal
  int* accepted;               // [n_src]
  int* status;                 // [n_src] 0 ok, 3 NaN / non-finite energy (state left unchanged)
  const FilterTables* filters;
  const double* flux_table;
  long long flux_rows;
};

struct HmcCtx {
  const DevImage* im;
  CgSource src;
  CgBox box;                   // clipped to the image
  const float* z;              // this (source, image)'s photon patch
  double* tab;                 // this warp's float64 table
  double* s_part;              // [SLICE_MAX_IMG][4] per-image sums of the current evaluation
  const FilterTables* filters;
  const double* flux_table;
  long long flux_rows;
  int lane, b, n_img, item_flags;
};

// galaxy_source_like(th) (and, with parts, the sums the analytic part of the gradient needs) over the
// images of the block; every thread returns the same values.  out[0] = ll; parts: out[1] = d ll / d theta,
// the per-image flux terms stay in s_part[img * 4 + 2] for hmc_grad.
__device__ __noinline__ void hmc_eval(HmcCtx& c, const double* th, bool parts, double* out) {
  const DevImage& im = *c.im;
  c.src.theta = th[0]; c.src.sigma = th[1]; c.src.phi = th[2]; c.src.rho = th[3];
  c.src.u[0] = th[4]; c.src.u[1] = th[5];
#pragma unroll
  for (int k = 0; k < CG_NUM_BANDS; ++k) c.src.flux[k] = th[6 + k];
  __syncthreads();                     // everybody is done with the previous tables and sums
  const int flags = CG_ITEM_TRUNC_BOX | c.item_flags;
  ItemGeom g = item_table<double>(im, c.src, flags, c.lane, c.tab);
  const double flux = item_flux(im, c.src, flags, c.lane, false, c.filters, c.flux_table, c.flux_rows, g);
  const double b_raw = th[6 + im.band];       // the gradient's "bs[img.band]" (:63,66) is the raw flux
  __syncwarp();
  const int K = g.K, Kexp = 6 * im.n_psf;
  const int wbox = c.box.x1 - c.box.x0, npx = wbox * (c.box.y1 - c.box.y0);
  double ll = 0., sf = 0., gt = 0., sz = 0.;
  const double ith = 1. / th[0], ith1 = 1. / (1. - th[0]);
  for (int p = c.lane; p < npx; p += 32) {
    const int yy = p / wbox, xx = p - yy * wbox;
    const double xl = (double)(c.box.x0 + xx - g.ox), yl = (double)(c.box.y0 + yy - g.oy);
    double fe = 0., fd = 0.;
    for (int k = 0; k < K; ++k) {
      const double* t = c.tab + k * TAB_W;
      const double e = exp2(fma(fma(t[0], xl, fma(t[1], yl, t[3])), xl, fma(fma(t[2], yl, t[4]), yl, t[5])));
      if (k < Kexp) fe += e; else fd += e;
    }
    const double f = fe + fd, v = flux * f, z = (double)c.z[p];
    if (v >= 1.17549435e-38) ll += log(v) * z;
    ll -= v;
    if (parts) {
      sf += f; sz += z;
      gt += (z / f - b_raw) * (fe * ith - fd * ith1);         // (Z / f_nms - b) (f_exp - f_dev), :63
    }
  }
  ll = warp_sum(ll);
  if (parts) { sf = warp_sum(sf); gt = warp_sum(gt); sz = warp_sum(sz); }
  if (c.lane == 0) {
    double* sp = c.s_part + c.b * 4;
    sp[0] = ll; sp[1] = gt; sp[2] = sz / b_raw - sf; sp[3] = 0.;
  }
  __syncthreads();
  double tot = 0., tg = 0.;
  for (int k = 0; k < c.n_img; ++k) { tot += c.s_part[k * 4]; tg += c.s_part[k * 4 + 1]; }
  out[0] = tot;
  if (parts) out[1] = tg;
}

// galaxy_source_like_grad (:44-88) -> grad[HMC_DIM]; returns false on a non-finite value
__device__ __noinline__ bool hmc_grad(HmcCtx& c, const ImgPtrs& imgs, const double* th, double* grad) {
  double out[2];
  hmc_eval(c, th, true, out);
  grad[0] = out[1];
#pragma unroll
  for (int k = 0; k < CG_NUM_BANDS; ++k) grad[6 + k] = 0.;
  for (int k = 0; k < c.n_img; ++k) grad[6 + imgs.p[k]->band] += c.s_part[k * 4 + 2];      // :66
  bool ok = isfinite(out[0]);
  double t2[HMC_DIM];
  for (int i = 1; i <= 5; ++i) {                                                          // :69-78
#pragma unroll
    for (int k = 0; k < HMC_DIM; ++k) t2[k] = th[k];
    double lp, lm;
    t2[i] = th[i] + 1e-5;
    hmc_eval(c, t2, false, &lp);
    t2[i] = th[i] - 1e-5;
    hmc_eval(c, t2, false, &lm);
    grad[i] = (lp - lm) / (2. * 1e-5);
    ok = ok && isfinite(lp) && isfinite(lm);
  }
  return ok;
}

#define HMC_WARP_BYTES (CG_MAX_COMP * TAB_W * 8)
#ifndef HMC_MINB
#define HMC_MINB 3
#endif
__global__ void __launch_bounds__(32 * SLICE_MAX_IMG, HMC_MINB) hmc_galaxy_kernel(HmcArgs A) {
  extern __shared__ __align__(16) unsigned char s_dyn[];           // n_img x HMC_WARP_BYTES
  __shared__ double s_part[SLICE_MAX_IMG * 4];
  const int j = blockIdx.x;          // j: row of the per-call arrays; s: the source it updates
  if (j >= A.n_src) return;
  const int s = A.sel ? A.sel[j] : j;
  HmcCtx c;
  c.lane = threadIdx.x & 31; c.b = threadIdx.x >> 5; c.n_img = A.n_img; c.item_flags = A.item_flags;
  c.im = A.imgs.p[c.b];
  c.src = A.srcs[s];
  c.tab = reinterpret_cast<double*>(s_dyn + (size_t)c.b * HMC_WARP_BYTES);
  c.s_part = s_part;
  c.filters = A.filters; c.flux_table = A.flux_table; c.flux_rows = A.flux_rows;
  const CgBox b0 = A.boxes[(long long)s * A.n_img + c.b];
  c.box.x0 = max(b0.x0, 0); c.box.y0 = max(b0.y0, 0); c.box.x1 = min(b0.x1, c.im->W); c.box.y1 = min(b0.y1, c.im->H);
  if (c.box.x1 <= c.box.x0 || c.box.y1 <= c.box.y0) { c.box.x1 = c.box.x0; c.box.y1 = c.box.y0; }
  c.z = A.aux + ((long long)s * A.n_img + c.b) * A.aux_stride;

  double X[HMC_DIM], P[HMC_DIM], im_[HMC_DIM], grad[HMC_DIM];
  X[0] = c.src.theta; X[1] = c.src.sigma; X[2] = c.src.phi; X[3] = c.src.rho; X[4] = c.src.u[0]; X[5] = c.src.u[1];
#pragma unroll
  for (int k = 0; k < CG_NUM_BANDS; ++k) X[6 + k] = c.src.flux[k];
#pragma unroll
  for (int k = 0; k < HMC_DIM; ++k) { P[k] = A.p_in[(long long)j * HMC_DIM + k]; im_[k] = 1. / A.mass[k]; }
  const double eps = A.eps;
  bool ok = true;
  // ll_curr = -hamiltonian(X, P) is taken BEFORE the momentum refresh (hmc.py:64): the host passes that
  // kinetic term
  double ll0;
  hmc_eval(c, X, false, &ll0);
  const double h_curr = ll0 - .5 * A.ke_curr[j];
  // leapstep (hmc.py:52-59); grad_energy = -grad
  double xx[HMC_DIM], pp[HMC_DIM];
#pragma unroll
  for (int k = 0; k < HMC_DIM; ++k) { xx[k] = X[k]; pp[k] = P[k]; }
  ok = hmc_grad(c, A.imgs, xx, grad) && ok;
#pragma unroll
  for (int k = 0; k < HMC_DIM; ++k) pp[k] = pp[k] - .5 * eps * (-grad[k]);
  for (int l = 0; l < A.n_steps && ok; ++l) {
#pragma unroll
    for (int k = 0; k < HMC_DIM; ++k) xx[k] = xx[k] + eps * im_[k] * pp[k];
    const double em = (l == A.n_steps - 1) ? .5 * eps : eps;
    ok = hmc_grad(c, A.imgs, xx, grad) && ok;
#pragma unroll
    for (int k = 0; k < HMC_DIM; ++k) pp[k] = pp[k] - em * (-grad[k]);
  }
  double ll1 = 0., ke1 = 0.;
  if (ok) hmc_eval(c, xx, false, &ll1);
#pragma unroll
  for (int k = 0; k < HMC_DIM; ++k) { pp[k] = -pp[k]; ke1 += im_[k] * pp[k] * pp[k]; }
  const double h_prop = ll1 - .5 * ke1;
  ok = ok && isfinite(h_prop) && isfinite(h_curr);
  const bool acc = ok && (A.log_u[j] < h_prop - h_curr);              // hmc.py:73
  if (threadIdx.x == 0) {
    A.status[j] = ok ? 0 : 3;
    A.accepted[j] = acc ? 1 : 0;
    A.h_out[2 * j] = h_curr; A.h_out[2 * j + 1] = h_prop;
    for (int k = 0; k < HMC_DIM; ++k) A.p_out[(long long)j * HMC_DIM + k] = -(acc ? pp[k] : P[k]);   // :74-79
    if (acc) {
      CgSource& o = A.srcs[s];
      o.theta = xx[0]; o.sigma = xx[1]; o.phi = xx[2]; o.rho = xx[3]; o.u[0] = xx[4]; o.u[1] = xx[5];
      for (int k = 0; k < CG_NUM_BANDS; ++k) o.flux[k] = xx[6 + k];
    }
  }
}

// ---------------------------------------------------------------------------
// per-sweep statistics a rank contributes to the all-reduce
// ---------------------------------------------------------------------------
// colour-space coordinates of the sources of one type (FluxColorMoG, util/dists/flux_prior.py:4-40:
// [lu - lr, lg - lr, li - lr, lz - lr, lr] of the log fluxes); mask = 0 for the others
__global__ void flux_colors_kernel(const CgSource* srcs, long long n, int type, double* colors, int* mask) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const CgSource& s = srcs[i];
  bool ok = (s.type == type);
  double l[CG_NUM_BANDS];
#pragma unroll
  for (int b = 0; b < CG_NUM_BANDS; ++b) {
    ok = ok && (s.flux[b] > 0.);
    l[b] = ok ? log(s.flux[b]) : 0.;
  }
  mask[i] = ok ? 1 : 0;
  double* c = colors + i * CG_NUM_BANDS;
  c[0] = ok ? l[0] - l[2] : 0.;
  c[1] = ok ? l[1] - l[2] : 0.;
  c[2] = ok ? l[3] - l[2] : 0.;
  c[3] = ok ? l[4] - l[2] : 0.;
  c[4] = ok ? l[2] : 0.;
}

// dist.SweepStats layout: [0] sum of the per-image log-likelihoods (fixed order), [1:6] sky photons per
// band, [6] stars, [7] galaxies, then the two prior blocks
__global__ void sweep_stats_kernel(const double* ll, int n_ll, const double* noise, const int* img_band, int n_noise,
                                   const CgSource* srcs, long long n_src, const double* prior0,
                                   const double* prior1, int block, double* vec) {
  __shared__ int s_cnt[2];
  if (threadIdx.x < 2) s_cnt[threadIdx.x] = 0;
  __syncthreads();
  int c0 = 0, c1 = 0;
  for (long long i = threadIdx.x; i < n_src; i += blockDim.x) {
    c0 += srcs[i].type == 0;
    c1 += srcs[i].type == 1;
  }
  atomicAdd(&s_cnt[0], c0);
  atomicAdd(&s_cnt[1], c1);
  for (int i = threadIdx.x; i < block; i += blockDim.x) {
    vec[8 + i] = prior0 ? prior0[i] : 0.;
    vec[8 + block + i] = prior1 ? prior1[i] : 0.;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    double t = 0;
    for (int i = 0; i < n_ll; ++i) t += ll[i];
    vec[0] = t;
    for (int b = 0; b < CG_NUM_BANDS; ++b) vec[1 + b] = 0.;
    for (int i = 0; i < n_noise; ++i) vec[1 + img_band[i]] += noise[i];
    vec[6] = (double)s_cnt[0];
    vec[7] = (double)s_cnt[1];
  }
}

// ===========================================================================
// host side
// ===========================================================================
struct SweepLayout {
  size_t items, items2, unit_fx, total;
};

extern "C" {

size_t cg_sweep_workspace_bytes(int32_t kind, int32_t n_src, int32_t n_img, int32_t P, int64_t max_patch_px,
                                int64_t n_tiles) {
  const long long n_items = (long long)n_src * n_img;
  switch (kind) {
    case CG_WS_SWEEP_ALLOC: {
      WsLayout L = ws_layout(std::max<long long>(n_items, 1), max_patch_px, 1, 0, n_tiles);
      return L.total + align_up(sizeof(unsigned long long) * std::max<long long>(n_items, 1), 256) + 512;
    }
    case CG_WS_SWEEP_SCORE: {
      const long long n1 = n_items * (P + 1);
      const int cap = cg_sweep_guard_cap(n_src, P);
      WsLayout L = ws_layout(std::max<long long>(n1, 1), max_patch_px, 0, 0, 0);
      WsLayout L2 = ws_layout((long long)cap * 2 * n_img * 2, max_patch_px, 0, 0, 0);   // x2: float64 tables
      return L.total + L2.total + align_up(sizeof(CgItem) * n1, 256) +
             align_up(sizeof(CgItem) * (size_t)cap * 2 * n_img, 256) + align_up(sizeof(int) * (size_t)cap, 256) + 1024;
    }
    case CG_WS_SWEEP_BACKGROUND: {
      WsLayout L = ws_layout(std::max<long long>(n_items, 1), max_patch_px, 0, 0, 0);
      return L.total + 512;
    }
    default: return 0;
  }
}

int32_t cg_sweep_guard_cap(int32_t n_src, int32_t P) {
  const long long n = (long long)n_src * P;
  return (int32_t)std::min<long long>(n, std::max<long long>(4096, n / 8));
}

int cg_field_sample_photons(CgImage* const* imgs, int32_t n_img, const CgSource* d_srcs, int32_t n_src,
                            double bound_cap, int64_t max_patch_px, uint64_t seed, float* d_samp,
                            CgBox* d_boxes, double* d_noise_sum, double* d_counts, double* d_unit_sums,
                            void* d_ws, size_t ws_bytes, void* stream) {
  NEED_INIT();
  cudaStream_t st = (cudaStream_t)stream;
  if (n_src < 0) return fail(CG_EINVAL, "negative n_src");
  if (!d_samp || !d_boxes || !d_noise_sum) return fail(CG_EINVAL, "null output");
  const long long n_items = (long long)n_src * n_img;
  const int32_t img_off[2] = {0, n_img}, src_off[2] = {0, n_src};
  PrepArgs pa;
  AllocTileArgs A;
  memset(&A, 0, sizeof A);
  WsLayout L;
  const size_t extra = align_up(sizeof(unsigned long long) * std::max<long long>(n_items, 1), 256);
  int rc = field_setup(imgs, img_off, d_srcs, src_off, 1, bound_cap, max_patch_px, d_ws, ws_bytes, extra, st,
                       CG_ITEM_FLUX_BAND, d_boxes, pa, A.f, L);
  if (rc) return rc;
  char* base = (char*)d_ws;
  A.samp = d_samp; A.max_patch_px = max_patch_px; A.seed = seed;
  A.item_unit_fx = (unsigned long long*)(base + align_up(L.total, 256));
  // the counts are accumulated here whether or not the caller wants them
  double* counts = d_counts ? d_counts : (double*)(base + L.item_ll);
  A.item_counts = counts;
  if (n_items > 0) {
    CUDA_TRY(cudaMemsetAsync(counts, 0, sizeof(double) * n_items, st));
    CUDA_TRY(cudaMemsetAsync(A.item_unit_fx, 0, sizeof(unsigned long long) * n_items, st));
  }
  int per_sm = 0;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, alloc_tile_kernel, 256, 0) != cudaSuccess || per_sm < 1)
    per_sm = 2;
  alloc_tile_kernel<<<g_sm_count * per_sm, 256, 0, st>>>(A);
  LAUNCH_CHECK();
  A.f.ll = d_noise_sum;                       // per-image sum of the per-bin noise photons, fixed order
  field_tile_final_kernel<<<n_img, FIN_THREADS, 0, st>>>(A.f);
  LAUNCH_CHECK();
  if (d_unit_sums && n_items > 0) {
    unit_sums_kernel<<<(unsigned)((n_items + 255) / 256), 256, 0, st>>>(A.item_unit_fx, n_items, d_unit_sums);
    LAUNCH_CHECK();
  }
  g_last_counters = pa.cnt;
  return CG_OK;
}

int cg_resample_fluxes(CgImage* const* imgs, int32_t n_img, CgSource* d_srcs, int32_t n_src,
                       const double* d_counts, const double* d_unit_sums, double a0, double b0, uint64_t seed,
                       void* stream) {
  NEED_INIT();
  ImgPtrs ip;
  int rc = fill_imgs(imgs, n_img, &ip);
  if (rc) return rc;
  if (n_src <= 0) return CG_OK;
  if (!d_counts || !d_unit_sums) return fail(CG_EINVAL, "photon counts and unit-patch sums are required");
  resample_fluxes_kernel<<<(n_src * CG_NUM_BANDS + 255) / 256, 256, 0, (cudaStream_t)stream>>>(
      d_srcs, n_src, n_img, ip, d_counts, d_unit_sums, a0, b0, seed);
  LAUNCH_CHECK();
  return CG_OK;
}

int cg_resample_epsilon(CgImage* const* imgs, int32_t n_img, const double* d_noise_sum, double a0, double b0,
                        uint64_t seed, double* d_eps_out, void* stream) {
  NEED_INIT();
  ImgPtrs ip;
  int rc = fill_imgs(imgs, n_img, &ip);
  if (rc) return rc;
  resample_epsilon_kernel<<<1, 64, 0, (cudaStream_t)stream>>>(ip, n_img, d_noise_sum, a0, b0, seed, d_eps_out);
  LAUNCH_CHECK();
  return CG_OK;
}

int cg_image_sync_epsilon(CgImage* img, void* stream, double* h_epsilon) {
  NEED_INIT();
  if (!img) return fail(CG_EINVAL, "null image");
  CUDA_TRY(cudaStreamSynchronize((cudaStream_t)stream));
  CUDA_TRY(cudaMemcpy(&img->h.epsilon, &img->d->epsilon, sizeof(double), cudaMemcpyDeviceToHost));
  if (h_epsilon) *h_epsilon = img->h.epsilon;
  return CG_OK;
}

int cg_make_proposals(const CgSource* d_cur, int32_t n_src, int32_t P, const double* h_scales, uint64_t seed,
                      CgSource* d_states, void* stream) {
  NEED_INIT();
  if (n_src <= 0) return CG_OK;
  if (P < 0 || !h_scales) return fail(CG_EINVAL, "bad proposal arguments");
  ProposalScales sc = {h_scales[0], h_scales[1], h_scales[2], h_scales[3], h_scales[4], h_scales[5]};
  const long long n = (long long)n_src * (P + 1);
  make_proposals_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(d_cur, n_src, P, sc, seed,
                                                                                     d_states);
  LAUNCH_CHECK();
  return CG_OK;
}

int cg_select_proposals(const CgSource* d_states, int32_t n_src, int32_t P, const uint8_t* d_accept,
                        CgSource* d_cur, int32_t* d_n_accepted, void* stream) {
  NEED_INIT();
  if (n_src <= 0) return CG_OK;
  cudaStream_t st = (cudaStream_t)stream;
  if (d_n_accepted) CUDA_TRY(cudaMemsetAsync(d_n_accepted, 0, sizeof(int), st));
  select_proposals_kernel<<<(n_src + 255) / 256, 256, 0, st>>>(d_states, n_src, P, d_accept, d_cur, d_n_accepted);
  LAUNCH_CHECK();
  return CG_OK;
}

int cg_score_proposals_multi(CgImage* const* imgs, int32_t n_img, const CgSource* d_states, int32_t n_src,
                             int32_t P, const CgBox* d_boxes, int32_t mode, const float* d_aux,
                             int64_t max_patch_px, const double* d_log_terms, const double* d_log_u,
                             double guard, double* d_delta, uint8_t* d_accept, int32_t* d_n_guard, void* d_ws,
                             size_t ws_bytes, void* stream) {
  NEED_INIT();
  cudaStream_t st = (cudaStream_t)stream;
  if (mode != CG_LL_POISSON_BG && mode != CG_LL_CONDITIONAL) return fail(CG_EINVAL, "unknown mode %d", mode);
  if (mode == CG_LL_CONDITIONAL && !d_aux) return fail(CG_EINVAL, "conditional mode needs the photon patches");
  if (n_src <= 0 || P <= 0) return CG_OK;
  if (!d_boxes || !d_log_u || !d_delta || !d_accept) return fail(CG_EINVAL, "null argument");
  PrepArgs pa;
  memset(&pa, 0, sizeof pa);
  int rc = fill_imgs(imgs, n_img, &pa.imgs);
  if (rc) return rc;
  for (int i = 0; i < n_img; ++i)
    if (!imgs[i]->h.nelec) return fail(CG_EINVAL, "image %d has no nelec buffer", i);
  const long long n1 = (long long)n_src * (P + 1) * n_img;
  if (n1 > 0x7fffffffLL) return fail(CG_EINVAL, "too many items");
  const int cap = guard > 0 ? cg_sweep_guard_cap(n_src, P) : 0;
  const long long n2 = (long long)cap * 2 * n_img;
  WsLayout L = ws_layout(n1, max_patch_px, 0, 0, 0);
  WsLayout L2 = ws_layout(std::max<long long>(n2, 1) * 2, max_patch_px, 0, 0, 0);
  size_t o = align_up(L.total, 256);
  const size_t o_l2 = o; o += align_up(L2.total, 256);
  const size_t o_items = o; o += align_up(sizeof(CgItem) * n1, 256);
  const size_t o_items2 = o; o += align_up(sizeof(CgItem) * (size_t)std::max<long long>(n2, 1), 256);
  const size_t o_glist = o; o += align_up(sizeof(int) * (size_t)std::max(cap, 1), 256);
  if (o > ws_bytes) return fail(CG_ENOMEM, "workspace too small: need %zu bytes, have %zu", o, ws_bytes);
  char* base = (char*)d_ws;
  Counters* cnt = (Counters*)(base + L.counters);
  CUDA_TRY(cudaMemsetAsync(cnt, 0, sizeof(Counters), st));
  CgItem* items = (CgItem*)(base + o_items);
  const int use_aux = d_aux != nullptr;
  score_items_kernel<<<(unsigned)((n1 + 255) / 256), 256, 0, st>>>(d_boxes, n_src, P, n_img, CG_ITEM_FLUX_BAND,
                                                                   use_aux, items);
  LAUNCH_CHECK();
  pa.n_img = n_img; pa.srcs = d_states; pa.n_src = n_src * (P + 1); pa.items = items; pa.n_items = n1;
  pa.bound_cap = 0; pa.max_patch_px = max_patch_px; pa.hdr = (ItemHdr*)(base + L.hdr); pa.tab = base + L.tab;
  pa.worklist = (int2*)(base + L.worklist); pa.worklist_cap = L.worklist_cap; pa.cnt = cnt;
  pa.filters = g_d_filters; pa.flux_table = g_flux_table; pa.flux_rows = g_flux_rows;
  rc = run_prepare<float>(pa, st);
  if (rc) return rc;
  RenderArgs ra;
  memset(&ra, 0, sizeof ra);
  ra.imgs = pa.imgs; ra.hdr = pa.hdr; ra.tab = pa.tab; ra.worklist = pa.worklist; ra.worklist_cap = pa.worklist_cap;
  ra.cull = g_cull | (g_recur ? 0 : 2); ra.logtab = g_d_logtab; ra.cnt = cnt;
  ra.aux = d_aux; ra.aux_offsets = nullptr; ra.aux_stride = max_patch_px; ra.partials = (double*)(base + L.partials);
  rc = (mode == CG_LL_POISSON_BG) ? run_render<float, MODE_POISSON>(ra, n1, st) : run_render<float, MODE_COND>(ra, n1, st);
  if (rc) return rc;
  double* item_ll = (double*)(base + L.item_ll);
  items_finalize_kernel<<<(unsigned)((n1 + 255) / 256), 256, 0, st>>>(pa.hdr, ra.partials, n1, mode, item_ll, nullptr, 1);
  LAUNCH_CHECK();
  int* guard_list = (int*)(base + o_glist);
  const long long n_sp = (long long)n_src * P;
  score_multi_finalize_kernel<<<(unsigned)((n_sp + 255) / 256), 256, 0, st>>>(
      item_ll, n_src, P, n_img, d_log_terms, d_log_u, guard, d_delta, d_accept, cnt, guard_list, cap);
  LAUNCH_CHECK();
  Counters* cnt2 = nullptr;
  if (cap > 0) {
    char* b2 = base + o_l2;
    cnt2 = (Counters*)(b2 + L2.counters);
    CUDA_TRY(cudaMemsetAsync(cnt2, 0, sizeof(Counters), st));
    CgItem* items2 = (CgItem*)(base + o_items2);
    rescore_items_kernel<<<(unsigned)((n2 + 255) / 256), 256, 0, st>>>(d_boxes, cnt, guard_list, cap, P, n_img,
                                                                       CG_ITEM_FLUX_BAND, use_aux, items2);
    LAUNCH_CHECK();
    PrepArgs p2 = pa;
    p2.items = items2; p2.n_items = n2; p2.hdr = (ItemHdr*)(b2 + L2.hdr); p2.tab = b2 + L2.tab;
    p2.worklist = (int2*)(b2 + L2.worklist); p2.worklist_cap = L2.worklist_cap; p2.cnt = cnt2;
    rc = run_prepare<double>(p2, st);
    if (rc) return rc;
    RenderArgs r2 = ra;
    r2.hdr = p2.hdr; r2.tab = p2.tab; r2.worklist = p2.worklist; r2.worklist_cap = p2.worklist_cap; r2.cnt = cnt2;
    r2.partials = (double*)(b2 + L2.partials);
    rc = (mode == CG_LL_POISSON_BG) ? run_render<double, MODE_POISSON>(r2, n2, st)
                                    : run_render<double, MODE_COND>(r2, n2, st);
    if (rc) return rc;
    double* item_ll2 = (double*)(b2 + L2.item_ll);
    items_finalize_kernel<<<(unsigned)((n2 + 255) / 256), 256, 0, st>>>(p2.hdr, r2.partials, n2, mode, item_ll2,
                                                                        nullptr, 1);
    LAUNCH_CHECK();
    score_multi_rescore_kernel<<<(cap + 255) / 256, 256, 0, st>>>(item_ll2, n_img, cnt, guard_list, cap, d_log_terms,
                                                                  d_log_u, d_delta, d_accept);
    LAUNCH_CHECK();
  }
  score_poison_kernel<<<(unsigned)((n_sp + 255) / 256), 256, 0, st>>>(cnt, cnt2, n_sp, d_delta);
  LAUNCH_CHECK();
  if (d_n_guard)
    CUDA_TRY(cudaMemcpyAsync(d_n_guard, &cnt->guard_count, sizeof(int), cudaMemcpyDeviceToDevice, st));
  g_last_counters = cnt;
  return CG_OK;
}

int cg_background_patches(CgImage* const* imgs, int32_t n_img, const CgSource* d_srcs, int32_t n_src,
                          const CgBox* d_boxes, const int32_t* d_nbr_off, const int32_t* d_nbr_idx,
                          int64_t max_patch_px, int32_t max_box_side, float* d_bg, void* d_ws, size_t ws_bytes,
                          void* stream) {
  NEED_INIT();
  cudaStream_t st = (cudaStream_t)stream;
  if (n_src <= 0) return CG_OK;
  if (!d_boxes || !d_nbr_off || !d_bg) return fail(CG_EINVAL, "null argument");
  if (max_box_side < 1 || (long long)max_box_side * max_box_side > max_patch_px)
    return fail(CG_EINVAL, "max_box_side^2 must not exceed max_patch_px");
  PrepArgs pa;
  memset(&pa, 0, sizeof pa);
  int rc = fill_imgs(imgs, n_img, &pa.imgs);
  if (rc) return rc;
  const long long n_items = (long long)n_src * n_img;
  WsLayout L = ws_layout(n_items, max_patch_px, 0, 0, 0);
  if (L.total > ws_bytes) return fail(CG_ENOMEM, "workspace too small: need %zu bytes, have %zu", L.total, ws_bytes);
  char* base = (char*)d_ws;
  Counters* cnt = (Counters*)(base + L.counters);
  CUDA_TRY(cudaMemsetAsync(cnt, 0, sizeof(Counters), st));
  // tables of every (source, image) item; the footprints are not used (the neighbours are rendered on
  // the TARGET's box), so nothing is listed, binned or capped
  pa.n_img = n_img; pa.srcs = d_srcs; pa.n_src = n_src; pa.items = nullptr; pa.n_items = n_items;
  pa.default_flags = CG_ITEM_FLUX_BAND; pa.implicit_mode = 0; pa.bound_cap = 0; pa.max_patch_px = max_patch_px;
  pa.hdr = (ItemHdr*)(base + L.hdr); pa.tab = base + L.tab; pa.worklist = nullptr; pa.cnt = cnt;
  pa.filters = g_d_filters; pa.flux_table = g_flux_table; pa.flux_rows = g_flux_rows;
  rc = run_prepare<float>(pa, st);
  if (rc) return rc;
  BgArgs B;
  memset(&B, 0, sizeof B);
  B.f.imgs = pa.imgs; B.f.n_img = n_img; B.f.hdr = pa.hdr; B.f.tab = (const float*)pa.tab; B.f.cnt = cnt;
  B.f.cull = g_cull | (g_recur ? 0 : 2); B.f.logtab = g_d_logtab;
  B.boxes = d_boxes; B.nbr_off = d_nbr_off; B.nbr_idx = d_nbr_idx; B.bg = d_bg; B.max_patch_px = max_patch_px;
  B.n_src = n_src; B.n_img = n_img; B.tiles_side = (max_box_side + BIN_W - 1) / BIN_W;
  int per_sm = 0;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, bg_tile_kernel, 256, 0) != cudaSuccess || per_sm < 1)
    per_sm = 2;
  bg_tile_kernel<<<g_sm_count * per_sm, 256, 0, st>>>(B);
  LAUNCH_CHECK();
  g_last_counters = cnt;
  return CG_OK;
}

int cg_flux_prior_stats(const CgSource* d_srcs, int64_t n_src, int32_t src_type, const double* d_means,
                        const double* d_icovs, const double* d_dets, const double* d_pis, int32_t K,
                        void* d_scratch, size_t scratch_bytes, double* d_logp, double* d_out, void* stream) {
  NEED_INIT();
  cudaStream_t st = (cudaStream_t)stream;
  const int D = CG_NUM_BANDS;
  if (n_src < 0 || K < 1) return fail(CG_EINVAL, "cg_flux_prior_stats: need n_src >= 0, K >= 1");
  const size_t need = align_up(sizeof(double) * n_src * D, 256) + align_up(sizeof(double) * n_src * K, 256) +
                      align_up(sizeof(double) * n_src, 256) + align_up(sizeof(int) * n_src, 256);
  if (need > scratch_bytes) return fail(CG_ENOMEM, "scratch too small: need %zu bytes, have %zu", need, scratch_bytes);
  char* base = (char*)d_scratch;
  double* colors = (double*)base;
  double* resp = (double*)(base + align_up(sizeof(double) * n_src * D, 256));
  double* logp = (double*)((char*)resp + align_up(sizeof(double) * n_src * K, 256));
  int* mask = (int*)((char*)logp + align_up(sizeof(double) * n_src, 256));
  const int entries = K * (1 + D + D * (D + 1) / 2);
  if (n_src == 0) { CUDA_TRY(cudaMemsetAsync(d_out, 0, sizeof(double) * entries, st)); return CG_OK; }
  flux_colors_kernel<<<(unsigned)((n_src + 255) / 256), 256, 0, st>>>(d_srcs, n_src, src_type, colors, mask);
  LAUNCH_CHECK();
  mog_logpdf_kernel<<<(unsigned)((n_src + 3) / 4), 128, 0, st>>>(colors, n_src, D, d_means, d_icovs, d_dets, d_pis, K,
                                                                 d_logp ? d_logp : logp, resp);
  LAUNCH_CHECK();
  prior_stats_kernel<<<entries, 256, 0, st>>>(resp, colors, n_src, K, D, d_out, mask);
  LAUNCH_CHECK();
  return CG_OK;
}

size_t cg_flux_prior_scratch_bytes(int64_t n_src, int32_t K) {
  return align_up(sizeof(double) * n_src * CG_NUM_BANDS, 256) + align_up(sizeof(double) * n_src * K, 256) +
         align_up(sizeof(double) * n_src, 256) + align_up(sizeof(int) * n_src, 256) + 256;
}

int cg_sweep_stats(const double* d_ll, int32_t n_ll, const double* d_noise, const int32_t* d_img_band,
                   int32_t n_noise, const CgSource* d_srcs, int64_t n_src, const double* d_prior0,
                   const double* d_prior1, int32_t block, double* d_vec, void* stream) {
  NEED_INIT();
  if (!d_ll || !d_vec || n_ll < 0 || block < 0) return fail(CG_EINVAL, "bad argument");
  if (n_noise > 0 && (!d_noise || !d_img_band)) return fail(CG_EINVAL, "noise sums need the image bands");
  sweep_stats_kernel<<<1, 256, 0, (cudaStream_t)stream>>>(d_ll, n_ll, d_noise, d_img_band, n_noise, d_srcs, n_src,
                                                         d_prior0, d_prior1, block, d_vec);
  LAUNCH_CHECK();
  return CG_OK;
}

int cg_slice_sample_locations(CgImage* const* imgs, int32_t n_img, CgSource* d_srcs, int32_t n_src,
                              const CgBox* d_boxes, const float* d_photons, int64_t max_patch_px,
                              const double* d_uniforms, const double* d_log_uniforms, const int32_t* d_order,
                              int32_t n_uniforms, double sigma, double guard, double* d_ll, int32_t* d_n_used,
                              int32_t* d_status, int32_t* d_n_evals, int32_t* d_n_refined, void* stream) {
  NEED_INIT();
  SliceArgs A;
  memset(&A, 0, sizeof A);
  int rc = fill_imgs(imgs, n_img, &A.imgs);
  if (rc) return rc;
  if (n_img > SLICE_MAX_IMG) return fail(CG_EINVAL, "at most %d images per source in the on-device sampler", SLICE_MAX_IMG);
  if (n_src <= 0) return CG_OK;
  if (!d_srcs || !d_boxes || !d_photons || !d_uniforms || !d_log_uniforms || !d_order || !d_ll || !d_n_used ||
      !d_status || !d_n_evals)
    return fail(CG_EINVAL, "null argument");
  if (n_uniforms < 4) return fail(CG_EINVAL, "at least 4 uniforms per source");
  A.n_img = n_img; A.n_src = n_src; A.B = n_uniforms; A.srcs = d_srcs; A.boxes = d_boxes; A.aux = d_photons;
  A.aux_stride = max_patch_px; A.uni = d_uniforms; A.log_uni = d_log_uniforms; A.ord = d_order; A.sigma = sigma;
  A.guard = guard; A.ll_out = d_ll; A.n_used = d_n_used; A.status = d_status; A.n_evals = d_n_evals;
  A.filters = g_d_filters; A.flux_table = g_flux_table; A.flux_rows = g_flux_rows; A.logtab = g_d_logtab;
  A.cull = g_cull | (g_recur ? 0 : 2);
  A.n_refined = d_n_refined;
  const size_t smem = (size_t)n_img * SLICE_WARP_BYTES;
  if (smem > 48 * 1024)
    CUDA_TRY(cudaFuncSetAttribute(slice_location_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  slice_location_kernel<<<n_src, 32 * n_img, smem, (cudaStream_t)stream>>>(A);
  LAUNCH_CHECK();
  return CG_OK;
}

int cg_hmc_galaxy_params(CgImage* const* imgs, int32_t n_img, CgSource* d_srcs, int32_t n_src,
                         const int32_t* d_sel, int32_t item_flags,
                         const CgBox* d_boxes, const float* d_photons, int64_t max_patch_px,
                         const double* d_p_in, const double* d_ke_curr, const double* d_log_u, double step,
                         const double* mass, int32_t n_steps, double* d_p_out, double* d_h_out,
                         int32_t* d_accepted, int32_t* d_status, void* stream) {
  NEED_INIT();
  HmcArgs A;
  memset(&A, 0, sizeof A);
  int rc = fill_imgs(imgs, n_img, &A.imgs);
  if (rc) return rc;
  if (n_img > SLICE_MAX_IMG) return fail(CG_EINVAL, "at most %d images per source in the on-device sampler", SLICE_MAX_IMG);
  if (n_src <= 0) return CG_OK;
  if (!d_srcs || !d_boxes || !d_photons || !d_p_in || !d_ke_curr || !d_log_u || !mass || !d_p_out || !d_h_out ||
      !d_accepted || !d_status)
    return fail(CG_EINVAL, "null argument");
  if (n_steps < 1 || !(step > 0.)) return fail(CG_EINVAL, "n_steps >= 1 and step > 0");
  for (int k = 0; k < HMC_DIM; ++k) {
    if (!(mass[k] > 0.)) return fail(CG_EINVAL, "mass[%d] must be positive", k);
    A.mass[k] = mass[k];
  }
  A.item_flags = item_flags & CG_ITEM_HEADER_CD; A.sel = d_sel;
  A.n_img = n_img; A.n_src = n_src; A.n_steps = n_steps; A.srcs = d_srcs; A.boxes = d_boxes; A.aux = d_photons;
  A.aux_stride = max_patch_px; A.p_in = d_p_in; A.ke_curr = d_ke_curr; A.log_u = d_log_u; A.eps = step;
  A.p_out = d_p_out; A.h_out = d_h_out; A.accepted = d_accepted; A.status = d_status;
  A.filters = g_d_filters; A.flux_table = g_flux_table; A.flux_rows = g_flux_rows;
  const size_t smem = (size_t)n_img * HMC_WARP_BYTES;
  if (smem > 48 * 1024)
    CUDA_TRY(cudaFuncSetAttribute(hmc_galaxy_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  hmc_galaxy_kernel<<<n_src, 32 * n_img, smem, (cudaStream_t)stream>>>(A);
  LAUNCH_CHECK();
  return CG_OK;
}

}  // extern "C"
