// Gradient GEMMs of the InfoNCE backward from STORED fp16 weight images (tcgen05.mma.cta_group::2 on CTA pairs).
//
// The weight matrix of an unordered modality pair (x, y) is one matrix:  g_ij = e_ij (u_i + v_j) - 2 T_ij  gives both
// dXhat = G * Yhat and dYhat = G^T * Xhat.  Who writes the images (16 KB each, 128 rows x 64 fp16, 128B-swizzled):
//   mode 2 / 3  tc_sim_kernel<EPI_WEIGHTS> (tc_sim.cu): G[128 i][64 j], one block per unordered pair (all rows local)
//               or per ordered pair (this rank's rows of a multi-GPU run);
//   mode 1      the pair sweep (tc_bwd2.cu), which keeps dXhat in TMEM and streams the spans [128 j][64 i] its grad MMAs
//               read; a stored weight then carries gs(x,y) / |y_j| and the epilogue here undoes that per row.
// One persistent kernel runs every job  out[M rows][d] = A[M][K] * unit_in[K][d]:
//   A = an image as it lies in HBM (raw TMA copy, no swizzle mode): K-major when the image rows are the output rows,
//       MN-major (halves of two images) when they are the reduction index -- G^T * Xhat from the SAME images;
//   B = 64 fp16 unit rows (d contiguous = MN-major), each CTA its half of the accumulator's d columns;
//   accumulators [128 lanes][256 (+128) columns] fp32 in TMEM; epilogue scales a row by gs_pair / gs_stored and writes
//   dxraw for combine_kernel.
// Executed work of the backward on one GPU: 6 N^2 D (weights) + 12 N^2 D (here) instead of 24 N^2 D.
// Warp roles (384 threads): 0 TMA producer, 1 MMA issuer (leader CTA), 2 TMEM allocator, 4..11 epilogue.
#include <cstdlib>

#include "tc_common.cuh"

namespace bsc {

int make_tmap_raw128(CUtensorMap* out, const void* base, int64_t rows, int box_rows);

namespace {

constexpr int BK = 64;
constexpr int A_BYTES = 16384;            // one stored span
constexpr int B_BOX = 8192;               // [64 i][64 d] fp16
constexpr int TMEM_COLS = 512;
constexpr int MAX_STAGES = 6;
// WIDE = 0: 256 d columns per item, one accumulator (N = 256), TMEM double-buffered over items, 6 x 32 KB stages.
// WIDE = 1: 384 d columns per item, two accumulators (N = 256 and N = 128) fed by the SAME A tile -- d = 768 is covered
//           by two items instead of three, i.e. the weight images are read twice per GEMM instead of three times (the
//           three sibling clusters of WIDE = 0 share them through L2 only partly: 25-31 GB of DRAM reads per launch
//           where 2 x 2 x 6.4 GB are needed); one TMEM buffer (384 of 512 columns), 5 x 40 KB stages.
template <int WIDE>
struct Cfg {
  static constexpr int BN = WIDE ? 384 : 256;
  static constexpr int NBOX = WIDE ? 3 : 2;  // 64-column boxes of B per CTA and stage
  static constexpr int STAGE_BYTES = A_BYTES + NBOX * B_BOX;
  static constexpr int STAGES = WIDE ? 5 : 6;
  static constexpr int NBUF = WIDE ? 1 : 2;
  static constexpr size_t SMEM = 1024 + (size_t)STAGES * STAGE_BYTES + 256;
};

struct RevParams {
  CUtensorMap tmap_g128;  // gimg as [rows][64] fp16, box 64 x 128, no swizzle (the bytes are already swizzled)
  CUtensorMap tmap_g64;   // same, box 64 x 64: half an image (MN-major jobs)
  CUtensorMap tmap_u[3];  // unit rows per modality, box 64 d x 64 rows, SWIZZLE_128B
  // job = one gradient GEMM  out[M rows][d] = A[M][K] * unit_in[K][d];  image(pair, r, c) = (pair * img_r + r) * img_c + c
  //   K-major job:  A block (128 rows mblk, 64 K kb)  = image(pair, mblk, kb)
  //   MN-major job: A block (128 rows mblk, 64 K kb)  = rows 64 (kb & 1) .. + 63 of images (pair, kb >> 1, 2 mblk + {0, 1})
  int n_jobs;
  int job_pair[6], job_mn[6], job_out[6], job_in[6], job_slot[6], job_nk[6];
  int job_fnum[6], job_fden[6], job_finv[6];  // epilogue row factor gsbuf[fnum] * gsbuf[fden] (/ inv_out[row] if finv)
  int img_r, img_c;
  int n_mb, nchunk, kparts, num_items, d;
  int wide;  // 384-wide items (two accumulators), d % 384 == 0
  int64_t n;
  const float* inv[3];
  const float* gsbuf;
  float* dxraw;  // [3][2][kparts][n][d]
};

struct __align__(8) RevTail {
  uint64_t full[MAX_STAGES];
  uint64_t empty[MAX_STAGES];
  uint64_t tmem_full[2];
  uint64_t tmem_empty[2];
  uint32_t tmem_base;
  uint32_t pad;
};

static_assert(sizeof(RevTail) <= 256, "tail budget");

__device__ __forceinline__ void rev_item(const RevParams& p, int item, int& job, int& mb, int& part, int& c, int& kb_lo,
                                         int& kb_hi) {
  c = item % p.nchunk;  // the d chunks of one (row block, K part) run on neighbouring clusters: they share A through L2
  int rest = item / p.nchunk;
  part = rest % p.kparts;
  rest /= p.kparts;
  mb = rest % p.n_mb;
  job = rest / p.n_mb;
  const int nk = p.job_nk[job];
  kb_lo = (int)((int64_t)part * nk / p.kparts);
  kb_hi = (int)((int64_t)(part + 1) * nk / p.kparts);
}

template <int WIDE>
__global__ void __launch_bounds__(384, 1) tc_bwd_rev_kernel(const __grid_constant__ RevParams p) {
  constexpr int BN = Cfg<WIDE>::BN;
  constexpr int STAGES = Cfg<WIDE>::STAGES;
  constexpr int STAGE_BYTES = Cfg<WIDE>::STAGE_BYTES;
  constexpr int NBUF = Cfg<WIDE>::NBUF;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw_addr = ptx::smem_u32(smem_raw);
  const uint32_t ring = (raw_addr + 1023u) & ~1023u;
  RevTail* tail = reinterpret_cast<RevTail*>(smem_raw + (ring - raw_addr) + (size_t)STAGES * STAGE_BYTES);
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t crank = ptx::cluster_ctarank();
  const int item0 = (int)ptx::cluster_id_x();
  const int item_step = (int)ptx::n_clusters_x();

  if (threadIdx.x == 0) {
    for (int s = 0; s < STAGES; ++s) {
      ptx::mbar_init(ptx::smem_u32(&tail->full[s]), 1);
      ptx::mbar_init(ptx::smem_u32(&tail->empty[s]), 1);
    }
    for (int b = 0; b < 2; ++b) {
      ptx::mbar_init(ptx::smem_u32(&tail->tmem_full[b]), 1);
      ptx::mbar_init(ptx::smem_u32(&tail->tmem_empty[b]), 16);
    }
    ptx::fence_barrier_init();
  }
  if (warp == 0 && lane == 0) {
    ptx::prefetch_tmap(&p.tmap_g128);
    ptx::prefetch_tmap(&p.tmap_g64);
    for (int a = 0; a < 3; ++a) ptx::prefetch_tmap(&p.tmap_u[a]);
  }
  if (warp == 2) {
    ptx::tmem_alloc_2cta(ptx::smem_u32(&tail->tmem_base), TMEM_COLS);
    ptx::tmem_relinquish_2cta();
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::cluster_sync_all();
  ptx::tc_fence_after();
  const uint32_t tmem_base = tail->tmem_base;

  if (warp == 0) {
    if (ptx::elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      for (int item = item0; item < p.num_items; item += item_step) {
        int job, mb, part, c, kb_lo, kb_hi;
        rev_item(p, item, job, mb, part, c, kb_lo, kb_hi);
        const CUtensorMap* mu = &p.tmap_u[p.job_in[job]];
        const int mblk = 2 * mb + (int)crank;
        const int64_t pair0 = (int64_t)p.job_pair[job] * p.img_r;
        const bool mn = p.job_mn[job] != 0;
        const int d0 = BN * c + 128 * (int)crank;       // this CTA's half of the N = 256 accumulator
        const int d1 = BN * c + 256 + 64 * (int)crank;  // ... and of the N = 128 accumulator (WIDE)
        for (int kb = kb_lo; kb < kb_hi; ++kb) {
          ptx::mbar_wait(ptx::smem_u32(&tail->empty[stage]), phase ^ 1u, 100 + stage);
          const uint32_t full = ptx::smem_u32(&tail->full[stage]);
          const uint32_t full_l = ptx::mapa(full, 0);
          const uint32_t sa = ring + stage * STAGE_BYTES;
          if (crank == 0) ptx::mbar_arrive_expect_tx(full, 2 * STAGE_BYTES);
          if (!mn) {
            const int64_t img = (pair0 + mblk) * p.img_c + kb;
            ptx::tma_load_2d_2cta(sa, &p.tmap_g128, full_l, 0, (int32_t)(img * 128));
          } else {
            const int64_t img = (pair0 + (kb >> 1)) * p.img_c + 2 * mblk;
            const int32_t r0 = (int32_t)(img * 128) + 64 * (kb & 1);
            ptx::tma_load_2d_2cta(sa, &p.tmap_g64, full_l, 0, r0);
            ptx::tma_load_2d_2cta(sa + A_BYTES / 2, &p.tmap_g64, full_l, 0, r0 + 128);
          }
          ptx::tma_load_2d_2cta(sa + A_BYTES, mu, full_l, d0, kb * BK);
          ptx::tma_load_2d_2cta(sa + A_BYTES + B_BOX, mu, full_l, d0 + 64, kb * BK);
          if (WIDE) ptx::tma_load_2d_2cta(sa + A_BYTES + 2 * B_BOX, mu, full_l, d1, kb * BK);
          if (++stage == STAGES) {
            stage = 0;
            phase ^= 1u;
          }
        }
      }
    }
  } else if (warp == 1) {
    if (crank == 0 && ptx::elect_one()) {
      constexpr uint32_t idesc_k = umma_idesc(256, 256, 0, 0, 0, 1);    // A K-major, B MN-major
      constexpr uint32_t idesc_mn = umma_idesc(256, 256, 0, 0, 1, 1);   // A MN-major, B MN-major
      constexpr uint32_t idesc_k2 = umma_idesc(256, 128, 0, 0, 0, 1);   // second accumulator (WIDE): N = 128
      constexpr uint32_t idesc_mn2 = umma_idesc(256, 128, 0, 0, 1, 1);
      const uint64_t desc_ak = ptx::umma_desc_sw128(0, 16, 1024);
      const uint64_t desc_amn = ptx::umma_desc_sw128(0, A_BYTES / 2, 1024);  // two 64-wide M chunks 8 KB apart
      const uint64_t desc_b = ptx::umma_desc_sw128(0, B_BOX, 1024);
      int stage = 0;
      uint32_t phase = 0;
      uint32_t count = 0;
      for (int item = item0; item < p.num_items; item += item_step, ++count) {
        int job, mb, part, c, kb_lo, kb_hi;
        rev_item(p, item, job, mb, part, c, kb_lo, kb_hi);
        const bool mn = p.job_mn[job] != 0;
        const uint32_t idesc = mn ? idesc_mn : idesc_k;
        const uint32_t idesc2 = mn ? idesc_mn2 : idesc_k2;
        const uint64_t desc_a = mn ? desc_amn : desc_ak;
        const uint64_t a_step = mn ? 128u : 2u;  // 16 K rows of 128 bytes / 32 bytes along a K-major row
        const uint32_t buf = NBUF == 2 ? (count & 1u) : 0u;
        const uint32_t use = NBUF == 2 ? (count >> 1) : count;
        ptx::mbar_wait_cluster(ptx::smem_u32(&tail->tmem_empty[buf]), (use & 1u) ^ 1u, 200 + buf);
        ptx::tc_fence_after();
        const uint32_t d_tmem = tmem_base + buf * BN;
        for (int kb = kb_lo; kb < kb_hi; ++kb) {
          ptx::mbar_wait(ptx::smem_u32(&tail->full[stage]), phase, 300 + stage);
          ptx::tc_fence_after();
          const uint32_t sa = ring + stage * STAGE_BYTES;
          const uint64_t da = desc_a + ptx::desc_addr(sa);
          const uint64_t db = desc_b + ptx::desc_addr(sa + A_BYTES);
#pragma unroll
          for (int k = 0; k < BK / 16; ++k) {
            const uint32_t acc = (kb != kb_lo || k != 0) ? 1u : 0u;
            ptx::umma_f16_2cta(d_tmem, da + a_step * k, db + (uint64_t)(k * 128), idesc, acc);
            if (WIDE) {
              const uint64_t db2 = desc_b + ptx::desc_addr(sa + A_BYTES + 2 * B_BOX);
              ptx::umma_f16_2cta(d_tmem + 256, da + a_step * k, db2 + (uint64_t)(k * 128), idesc2, acc);
            }
          }
          ptx::umma_commit_2cta(ptx::smem_u32(&tail->empty[stage]), 3);
          if (kb == kb_hi - 1) ptx::umma_commit_2cta(ptx::smem_u32(&tail->tmem_full[buf]), 3);
          if (++stage == STAGES) {
            stage = 0;
            phase ^= 1u;
          }
        }
      }
    }
  } else if (warp >= 4) {
    const int q = warp & 3;
    const int grp = (warp - 4) >> 2;  // column half of the item's accumulators
    constexpr int CPG = BN / 64;      // 32-column chunks per warp group
    const uint32_t empty_l0 = ptx::mapa(ptx::smem_u32(&tail->tmem_empty[0]), 0);
    uint32_t count = 0;
    for (int item = item0; item < p.num_items; item += item_step, ++count) {
      int job, mb, part, c, kb_lo, kb_hi;
      rev_item(p, item, job, mb, part, c, kb_lo, kb_hi);
      const int mo = p.job_out[job];
      const uint32_t buf = NBUF == 2 ? (count & 1u) : 0u;
      const uint32_t use = NBUF == 2 ? (count >> 1) : count;
      const int64_t row = ((int64_t)(2 * mb + (int)crank)) * 128 + q * 32 + lane;
      float f = 0.f;
      if (row < p.n) {
        f = p.gsbuf[p.job_fnum[job]] * p.gsbuf[p.job_fden[job]];
        if (p.job_finv[job]) {
          const float iv = p.inv[mo][row];
          f = iv > 0.f ? f / iv : 0.f;
        }
      }
      float* out = p.dxraw + (((int64_t)(mo * 2 + p.job_slot[job]) * p.kparts + part) * p.n + row) * p.d + BN * c;
      if (kb_hi > kb_lo) {
        ptx::mbar_wait(ptx::smem_u32(&tail->tmem_full[buf]), use & 1u, 400 + buf);
        ptx::tc_fence_after();
      }
      const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + buf * BN;
#pragma unroll 1
      for (int cc = grp * CPG; cc < grp * CPG + CPG; ++cc) {
        uint32_t raw[32];
        if (kb_hi > kb_lo) {
          ptx::tmem_ld_32x32(taddr + cc * 32, raw);
          ptx::tmem_ld_wait();
        } else {
#pragma unroll
          for (int t = 0; t < 32; ++t) raw[t] = 0u;
        }
        if (row < p.n) {
          float4* o4 = reinterpret_cast<float4*>(out + cc * 32);
#pragma unroll
          for (int t = 0; t < 8; ++t)
            o4[t] = make_float4(f * __uint_as_float(raw[4 * t]), f * __uint_as_float(raw[4 * t + 1]),
                                f * __uint_as_float(raw[4 * t + 2]), f * __uint_as_float(raw[4 * t + 3]));
        }
      }
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0 && kb_hi > kb_lo) ptx::mbar_arrive_remote_nodata(empty_l0 + buf * 8u);
    }
  }

  ptx::tc_fence_before();
  __syncthreads();
  ptx::cluster_sync_all();
  if (warp == 2) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc_2cta(tmem_base, TMEM_COLS);
  }
}

int sm_count3() {
  static int n = 0;
  if (n == 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
    if (n <= 0) n = 148;
  }
  return n;
}

}  // namespace

static int rev_launch(RevParams& p, cudaStream_t st) {
  if (p.num_items == 0) return BSC_OK;
  const bool wide = p.wide != 0;
  static bool attr_done = false;
  if (!attr_done) {
    BSC_CUDA(cudaFuncSetAttribute(tc_bwd_rev_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  (int)Cfg<0>::SMEM));
    BSC_CUDA(cudaFuncSetAttribute(tc_bwd_rev_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  (int)Cfg<1>::SMEM));
    attr_done = true;
  }
  int clusters = sm_count3() / 2;
  if (clusters > p.num_items) clusters = p.num_items;
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3((unsigned)(2 * clusters));
  cfg.blockDim = dim3(384);
  cfg.dynamicSmemBytes = wide ? Cfg<1>::SMEM : Cfg<0>::SMEM;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 2;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  timing_begin(3, st);
  if (wide)
    BSC_CUDA(cudaLaunchKernelEx(&cfg, tc_bwd_rev_kernel<1>, p));
  else
    BSC_CUDA(cudaLaunchKernelEx(&cfg, tc_bwd_rev_kernel<0>, p));
  timing_end(3, st);
  BSC_LAUNCH_CHECK("tc_bwd_rev_kernel");
  return BSC_OK;
}

static int rev_common(RevParams& p, const bsc_infonce_desc* d, const InfoNceWs& L, char* ws,
                      const void* const xunit[3], int64_t img_rows) {
  const int64_t N = d->n_global;
  int rc = make_tmap_raw128(&p.tmap_g128, ws + L.gimg, img_rows, 128);
  if (rc) return rc;
  rc = make_tmap_raw128(&p.tmap_g64, ws + L.gimg, img_rows, 64);
  if (rc) return rc;
  int first = -1;
  for (int a = 0; a < 3; ++a)
    if (xunit[a]) {
      rc = make_tmap_2d(&p.tmap_u[a], xunit[a], BSC_F16, N, d->d, 64);
      if (rc) return rc;
      if (first < 0) first = a;
    }
  BSC_REQUIRE(first >= 0, BSC_EINVAL, "tc_bwd_rev: no unit rows");
  for (int a = 0; a < 3; ++a) {
    if (!xunit[a]) p.tmap_u[a] = p.tmap_u[first];
    p.inv[a] = d->inv[a];
  }
  {
    const char* e = getenv("BSC_BWD_WIDE");  // 0 forces 256-wide items
    p.wide = (d->d % 384 == 0 && !(e && e[0] == '0')) ? 1 : 0;
  }
  p.nchunk = d->d / (p.wide ? 384 : 256);
  p.kparts = L.bwd_jsplit > 0 ? L.bwd_jsplit : 1;
  p.d = d->d;
  p.n = N;
  p.n_mb = (int)((N + 255) / 256);
  p.gsbuf = (const float*)(ws + L.gs);
  p.dxraw = (float*)(ws + L.dxraw);
  return BSC_OK;
}

// slot of partner modality `other` among the present modalities other than `m`, ascending
static int slot_of(const bsc_infonce_desc* d, int m, int other) {
  int slot = 0;
  for (int z = 0; z < other; ++z)
    if (z != m && d->xh[z]) ++slot;
  return slot;
}

// Stored-span mode (L.bwd_gstore == 1): dYhat of every unordered pair from the spans the pair sweep stored,
// gimg[(pair * 2 n_steps + j tile) * n_i64 + i64] = [128 j][64 i].  `pairs_x/y` list the pairs in the sweep's order;
// xunit[a] = fp16 unit rows of modality a (only the x side of a pair is read).
int tc_bwd_rev(const bsc_infonce_desc* d, const InfoNceWs& L, char* ws, const void* const xunit[3], int n_pairs,
               const int* pairs_x, const int* pairs_y, int n_jb, int n_i64, cudaStream_t st) {
  RevParams p{};
  int rc = rev_common(p, d, L, ws, xunit, (int64_t)n_pairs * 2 * n_jb * n_i64 * 128);
  if (rc) return rc;
  p.img_r = 2 * n_jb;
  p.img_c = n_i64;
  p.n_jobs = n_pairs;
  for (int q = 0; q < n_pairs; ++q) {
    const int x = pairs_x[q], y = pairs_y[q];
    p.job_pair[q] = q;
    p.job_mn[q] = 0;
    p.job_out[q] = y;
    p.job_in[q] = x;
    p.job_slot[q] = slot_of(d, y, x);
    p.job_nk[q] = n_i64;
    // a stored weight carries gs(x,y) / |y_j|; combine expects gs(y,x) * sum
    p.job_fnum[q] = 2 * pair_id(y, x);
    p.job_fden[q] = 2 * pair_id(x, y) + 1;
    p.job_finv[q] = 1;
  }
  BSC_REQUIRE(p.kparts <= n_i64, BSC_EUNSUPPORTED, "tc_bwd_rev: column split %d exceeds the %d row spans", p.kparts,
              n_i64);
  p.num_items = p.n_jobs * p.n_mb * p.kparts * p.nchunk;
  return rev_launch(p, st);
}

// K split of the gradient GEMMs (= dxraw parts that combine_kernel sums in a fixed order): the clusters take whole items,
// so a few hundred items on 74 clusters (many ranks: 192 at 8 GPUs) would leave the last round mostly idle.
int tc_bwd_gemm_ksplit(const bsc_infonce_desc* d, int mode) {
  int m = 0;
  for (int a = 0; a < 3; ++a) m += d->xh[a] ? 1 : 0;
  const int64_t jobs = (int64_t)m * (m - 1);  // 2 per unordered pair (mode 2) = 1 per ordered pair (mode 3)
  const int64_t n_mb = (d->n_local + 255) / 256;
  const int64_t nchunk = d->d % 384 == 0 ? d->d / 384 : d->d / 256;
  const int64_t items0 = jobs * n_mb * nchunk;
  const int64_t nk = (d->n_global + 63) / 64;
  const int clusters = sm_count3() / 2;
  (void)mode;
  if (items0 <= 0) return 1;
  int best = 1;
  double best_cost = 1e300;
  for (int s = 1; s <= 8; ++s) {
    if (s > nk) break;
    const double rounds = (double)((items0 * s + clusters - 1) / clusters);
    // cost in K blocks: rounds * (K blocks per item + ~12 blocks' worth of pipeline fill and accumulator drain)
    const double cost = rounds * ((double)nk / s + 12.0);
    if (cost < best_cost * 0.99) {
      best_cost = cost;
      best = s;
    }
  }
  return best;
}

// Weight-matrix modes: the gradient GEMMs from the images gimg[(pair * n_iblk_pad + iblk) * n_j64 + j64] = G[128 i][64 j]
// that tc_bwd_matrix wrote.
//   mode 2 (all rows local, unordered pairs (x, y)):  dXhat = G * Yunit (A K-major: image rows are i, K runs over j) and
//                                                     dYhat = G^T * Xunit (A MN-major from the same images, K over i);
//   mode 3 (ordered pairs (a, b), this rank's rows of a):  dAhat = G_ab * Bunit, K-major only.
int tc_bwd_gemms(const bsc_infonce_desc* d, const InfoNceWs& L, char* ws, const void* const xunit[3], int n_iblk_pad,
                 int n_j64, int n_pairs, const int* pairs_a, const int* pairs_b, int ordered, cudaStream_t st) {
  RevParams p{};
  int rc = rev_common(p, d, L, ws, xunit, (int64_t)n_pairs * n_iblk_pad * n_j64 * 128);
  if (rc) return rc;
  p.img_r = n_iblk_pad;
  p.img_c = n_j64;
  p.n = d->n_local;  // output rows (== n_global in mode 2)
  p.n_mb = (int)((d->n_local + 255) / 256);
  const int nk64 = (int)((d->n_global + 63) / 64);
  int job = 0;
  for (int q = 0; q < n_pairs; ++q) {
    const int x = pairs_a[q], y = pairs_b[q];
    for (int side = 0; side < (ordered ? 1 : 2); ++side, ++job) {
      const int mo = side == 0 ? x : y, mi = side == 0 ? y : x;
      p.job_pair[job] = q;
      p.job_mn[job] = side;
      p.job_out[job] = mo;
      p.job_in[job] = mi;
      p.job_slot[job] = slot_of(d, mo, mi);
      p.job_nk[job] = nk64;  // K blocks of 64 columns j (side 0) / 64 rows i (side 1, where n_local == n_global)
      p.job_fnum[job] = 2 * pair_id(mo, mi);
      // mode 2: weights scaled by gsbuf[16], combine expects gs(mo, mi); mode 3: the weights already carry gs(mo, mi)
      p.job_fden[job] = ordered ? 2 * pair_id(mo, mi) + 1 : 17;
      p.job_finv[job] = 0;
    }
  }
  p.n_jobs = job;
  BSC_REQUIRE(p.kparts >= 1 && p.kparts <= nk64, BSC_EUNSUPPORTED, "tc_bwd_gemms: K split %d of %d blocks", p.kparts,
              nk64);
  p.num_items = p.n_jobs * p.n_mb * p.kparts * p.nchunk;
  return rev_launch(p, st);
}

}  // namespace bsc
