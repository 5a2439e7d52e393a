// Reverse-direction gradients of the InfoNCE backward from STORED weight tiles (single-GPU runs).
//
// The weight matrix of an unordered modality pair (x, y) is one matrix:  g_ij = e_ij (u_i + v_j) - 2 T_ij  gives both
// dXhat = G * Yhat and dYhat = G^T * Xhat.  The pair sweep (tc_bwd2.cu) used to recompute S once per ORDERED pair
// (24 N^2 D executed for 12 algorithmic); with all rows local it now sweeps the three unordered pairs once, keeps dXhat
// in TMEM as before and additionally bulk-copies every fp16 weight span it builds (the very bytes its grad MMAs read,
// [128 j][64 i], 128B-swizzled, i contiguous) to HBM:  gimg[pair][j tile][i64][16 KB].
// This kernel then forms   dYhat[256 j x 256 d] = sum_i  G^T[j][i] * Xunit[i][d]   as a plain tcgen05 GEMM on CTA pairs
// (cta_group::2, M = 256, N = 256, K = 64 per stage):
//   A = a stored span = a K-major SWIZZLE_128B tile [128 j][64 i] as it lies in HBM (raw TMA copy, no swizzle mode),
//   B = 64 rows of Xunit (fp16 unit rows, d contiguous = MN-major), each CTA 128 of the chunk's 256 d columns,
//   accumulator [128 lanes = j][256 columns = d] fp32 in TMEM, double-buffered over work items.
// A stored weight carries gs(x,y) / |y_j| (tc_bwd2.cu folds the operand scale of Y16 into it); the epilogue multiplies
// row j by gs(y,x) |y_j| / gs(x,y) so that combine_kernel sees exactly what the ordered sweep (y -> x) used to write.
// Executed work of the backward: 12 N^2 D (sweep) + 6 N^2 D (this kernel) instead of 24 N^2 D.
// Warp roles (384 threads): 0 TMA producer, 1 MMA issuer (leader CTA), 2 TMEM allocator, 4..11 epilogue.
#include <cstdlib>

#include "tc_common.cuh"

namespace bsc {

int make_tmap_raw128(CUtensorMap* out, const void* base, int64_t rows, int box_rows);

namespace {

constexpr int BK = 64;
constexpr int A_BYTES = 16384;            // one stored span
constexpr int B_BOX = 8192;               // [64 i][64 d] fp16
constexpr int STAGE_BYTES = A_BYTES + 2 * B_BOX;  // 32 KB per CTA
constexpr int STAGES = 6;
constexpr int TMEM_COLS = 512;
constexpr int BN = 256;  // d columns per item

struct RevParams {
  CUtensorMap tmap_g;     // gimg as [rows][64] fp16, box 64 x 128, no swizzle (the bytes are already swizzled)
  CUtensorMap tmap_x[3];  // unit rows per modality, box 64 d x 64 rows, SWIZZLE_128B
  int n_pairs;
  int pair_x[3], pair_y[3], pair_slot[3];  // slot = index of x among y's partner modalities
  int n_jb, n_i64, nchunk, kparts, num_items, d;
  int64_t n;
  const float* inv[3];
  const float* gsbuf;
  float* dxraw;  // [3][2][kparts][n][d]
};

struct __align__(8) RevTail {
  uint64_t full[STAGES];
  uint64_t empty[STAGES];
  uint64_t tmem_full[2];
  uint64_t tmem_empty[2];
  uint32_t tmem_base;
  uint32_t pad;
};

constexpr size_t REV_SMEM_BYTES = 1024 + (size_t)STAGES * STAGE_BYTES + sizeof(RevTail);

__device__ __forceinline__ void rev_item(const RevParams& p, int item, int& pair, int& jb, int& part, int& c, int& kb_lo,
                                         int& kb_hi) {
  c = item % p.nchunk;  // the d chunks of one (j block, K part) run on neighbouring clusters: they share A through L2
  int rest = item / p.nchunk;
  part = rest % p.kparts;
  rest /= p.kparts;
  jb = rest % p.n_jb;
  pair = rest / p.n_jb;
  kb_lo = (int)((int64_t)part * p.n_i64 / p.kparts);
  kb_hi = (int)((int64_t)(part + 1) * p.n_i64 / p.kparts);
}

__global__ void __launch_bounds__(384, 1) tc_bwd_rev_kernel(const __grid_constant__ RevParams p) {
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
    ptx::prefetch_tmap(&p.tmap_g);
    for (int a = 0; a < 3; ++a) ptx::prefetch_tmap(&p.tmap_x[a]);
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
      const uint64_t pol_stream = ptx::l2_policy_evict_first();  // spans are read nchunk times back to back, then never
      int stage = 0;
      uint32_t phase = 0;
      for (int item = item0; item < p.num_items; item += item_step) {
        int pair, jb, part, c, kb_lo, kb_hi;
        rev_item(p, item, pair, jb, part, c, kb_lo, kb_hi);
        const CUtensorMap* mx = &p.tmap_x[p.pair_x[pair]];
        const int64_t span0 = ((int64_t)pair * (2 * p.n_jb) + (2 * jb + (int)crank)) * p.n_i64;
        const int d0 = BN * c + 128 * (int)crank;
        for (int kb = kb_lo; kb < kb_hi; ++kb) {
          ptx::mbar_wait(ptx::smem_u32(&tail->empty[stage]), phase ^ 1u, 100 + stage);
          const uint32_t full = ptx::smem_u32(&tail->full[stage]);
          const uint32_t full_l = ptx::mapa(full, 0);
          const uint32_t sa = ring + stage * STAGE_BYTES;
          if (crank == 0) ptx::mbar_arrive_expect_tx(full, 2 * STAGE_BYTES);
          ptx::tma_load_2d_2cta_hint(sa, &p.tmap_g, full_l, 0, (int32_t)((span0 + kb) * 128), pol_stream);
          ptx::tma_load_2d_2cta(sa + A_BYTES, mx, full_l, d0, kb * BK);
          ptx::tma_load_2d_2cta(sa + A_BYTES + B_BOX, mx, full_l, d0 + 64, kb * BK);
          if (++stage == STAGES) {
            stage = 0;
            phase ^= 1u;
          }
        }
      }
    }
  } else if (warp == 1) {
    if (crank == 0 && ptx::elect_one()) {
      constexpr uint32_t idesc = umma_idesc(256, BN, 0, 0, 0, 1);  // A K-major, B MN-major
      const uint64_t desc_a = ptx::umma_desc_sw128(0, 16, 1024);
      const uint64_t desc_b = ptx::umma_desc_sw128(0, B_BOX, 1024);
      int stage = 0;
      uint32_t phase = 0;
      uint32_t count = 0;
      for (int item = item0; item < p.num_items; item += item_step, ++count) {
        int pair, jb, part, c, kb_lo, kb_hi;
        rev_item(p, item, pair, jb, part, c, kb_lo, kb_hi);
        const uint32_t buf = count & 1u;
        ptx::mbar_wait_cluster(ptx::smem_u32(&tail->tmem_empty[buf]), ((count >> 1) & 1u) ^ 1u, 200 + buf);
        ptx::tc_fence_after();
        const uint32_t d_tmem = tmem_base + buf * BN;
        for (int kb = kb_lo; kb < kb_hi; ++kb) {
          ptx::mbar_wait(ptx::smem_u32(&tail->full[stage]), phase, 300 + stage);
          ptx::tc_fence_after();
          const uint32_t sa = ring + stage * STAGE_BYTES;
          const uint64_t da = desc_a + ptx::desc_addr(sa);
          const uint64_t db = desc_b + ptx::desc_addr(sa + A_BYTES);
#pragma unroll
          for (int k = 0; k < BK / 16; ++k)
            ptx::umma_f16_2cta(d_tmem, da + 2 * k, db + (uint64_t)(k * 128), idesc, (kb != kb_lo || k != 0) ? 1u : 0u);
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
    const int grp = (warp - 4) >> 2;  // 128-column half of the accumulator
    const uint32_t empty_l0 = ptx::mapa(ptx::smem_u32(&tail->tmem_empty[0]), 0);
    uint32_t count = 0;
    for (int item = item0; item < p.num_items; item += item_step, ++count) {
      int pair, jb, part, c, kb_lo, kb_hi;
      rev_item(p, item, pair, jb, part, c, kb_lo, kb_hi);
      const int x = p.pair_x[pair], y = p.pair_y[pair];
      const uint32_t buf = count & 1u;
      const int64_t j = ((int64_t)(2 * jb + (int)crank)) * 128 + q * 32 + lane;
      float f = 0.f;
      if (j < p.n) {
        const float ij = p.inv[y][j];
        if (ij > 0.f) f = p.gsbuf[2 * pair_id(y, x)] * p.gsbuf[2 * pair_id(x, y) + 1] / ij;
      }
      float* out = p.dxraw + (((int64_t)(y * 2 + p.pair_slot[pair]) * p.kparts + part) * p.n + j) * p.d + BN * c;
      if (kb_hi > kb_lo) {
        ptx::mbar_wait(ptx::smem_u32(&tail->tmem_full[buf]), (count >> 1) & 1u, 400 + buf);
        ptx::tc_fence_after();
      }
      const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + buf * BN;
#pragma unroll 1
      for (int cc = grp * 4; cc < grp * 4 + 4; ++cc) {
        uint32_t raw[32];
        if (kb_hi > kb_lo) {
          ptx::tmem_ld_32x32(taddr + cc * 32, raw);
          ptx::tmem_ld_wait();
        } else {
#pragma unroll
          for (int t = 0; t < 32; ++t) raw[t] = 0u;
        }
        if (j < p.n) {
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

// dYhat of every unordered pair from the stored weight spans.  `pairs_x/y` list the unordered pairs in the order the
// sweep stored them; xunit[a] = fp16 unit rows of modality a (only the x side of a pair is read).
int tc_bwd_rev(const bsc_infonce_desc* d, const InfoNceWs& L, char* ws, const void* const xunit[3], int n_pairs,
               const int* pairs_x, const int* pairs_y, int n_jb, int n_i64, cudaStream_t st) {
  RevParams p{};
  const int64_t N = d->n_global;
  int rc = make_tmap_raw128(&p.tmap_g, ws + L.gimg, (int64_t)n_pairs * 2 * n_jb * n_i64 * 128, 128);
  if (rc) return rc;
  int first = -1;
  for (int a = 0; a < 3; ++a)
    if (xunit[a]) {
      rc = make_tmap_2d(&p.tmap_x[a], xunit[a], BSC_F16, N, d->d, 64);
      if (rc) return rc;
      if (first < 0) first = a;
    }
  BSC_REQUIRE(first >= 0, BSC_EINVAL, "tc_bwd_rev: no unit rows");
  for (int a = 0; a < 3; ++a) {
    if (!xunit[a]) p.tmap_x[a] = p.tmap_x[first];
    p.inv[a] = d->inv[a];
  }
  p.n_pairs = n_pairs;
  for (int q = 0; q < n_pairs; ++q) {
    const int x = pairs_x[q], y = pairs_y[q];
    p.pair_x[q] = x;
    p.pair_y[q] = y;
    int slot = 0;  // index of x among the present modalities other than y, in ascending order
    for (int z = 0; z < x; ++z)
      if (z != y && d->xh[z]) ++slot;
    p.pair_slot[q] = slot;
  }
  p.n_jb = n_jb;
  p.n_i64 = n_i64;
  p.nchunk = d->d / BN;
  p.kparts = L.bwd_jsplit > 0 ? L.bwd_jsplit : 1;
  if (p.kparts > n_i64) p.kparts = n_i64;
  p.d = d->d;
  p.n = N;
  p.gsbuf = (const float*)(ws + L.gs);
  p.dxraw = (float*)(ws + L.dxraw);
  p.num_items = n_pairs * n_jb * p.kparts * p.nchunk;
  // parts of the K split that this kernel does not produce (kparts clipped) must still read as zero in combine
  BSC_REQUIRE(p.kparts == (L.bwd_jsplit > 0 ? L.bwd_jsplit : 1), BSC_EUNSUPPORTED,
              "tc_bwd_rev: column split %d exceeds the %d row spans", L.bwd_jsplit, n_i64);
  if (p.num_items == 0) return BSC_OK;

  static bool attr_done = false;
  if (!attr_done) {
    BSC_CUDA(cudaFuncSetAttribute(tc_bwd_rev_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)REV_SMEM_BYTES));
    attr_done = true;
  }
  int clusters = sm_count3() / 2;
  if (clusters > p.num_items) clusters = p.num_items;
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3((unsigned)(2 * clusters));
  cfg.blockDim = dim3(384);
  cfg.dynamicSmemBytes = REV_SMEM_BYTES;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 2;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  timing_begin(3, st);
  BSC_CUDA(cudaLaunchKernelEx(&cfg, tc_bwd_rev_kernel, p));
  timing_end(3, st);
  BSC_LAUNCH_CHECK("tc_bwd_rev_kernel");
  return BSC_OK;
}

}  // namespace bsc
