// tcgen05 similarity sweep: S = A * B^T over 128 x 256 tiles, operands streamed by TMA into a 4-stage
// 128B-swizzled shared-memory ring, accumulators double-buffered in TMEM (2 x 256 columns), epilogue
// applied straight out of TMEM so the N x N tile never reaches HBM.  Two epilogues:
//   EPI_STATS  exp(s*S - s) row sums and column sums            (InfoNCE forward, loss_func.py:46-51)
//   EPI_TOPK   per-row running top-k over the B rows            (IndexFlatIP.search, inference_and_eval.py:421)
// Warp roles (256 threads): 0 = TMA producer, 1 = MMA issuer, 2 = TMEM allocator, 4..7 = epilogue.
#ifdef BSC_EXPERIMENTS
#include <cstdlib>
#endif

#include "tc_common.cuh"

namespace bsc {

int reduce_parts_launch(const float* part, int nparts, int64_t part_stride, int64_t n, float* out, cudaStream_t st);
int rep_range_launch(const int32_t* rep, int64_t n, int tile, int2* out, cudaStream_t st);
int sum_rows_launch(const float* part, int nparts, int64_t part_stride, int64_t n, float scale, float* out,
                    double* out64, cudaStream_t st);
int topk_merge_launch(const float* cs, const int64_t* ci, int parts, int64_t nq, int k, float* os, int64_t* oi,
                      cudaStream_t st);

namespace {

constexpr int BM = kTcBM;    // 128
constexpr int BN = kTcBNF;   // 256
constexpr int BK = 64;       // one 128-byte swizzle span of 16-bit elements
constexpr int STAGES = 4;         // cta_group::1: 4 x 48 KB
constexpr int STAGES2 = 6;        // cta_group::2: 6 x 32 KB (every CTA stages its A rows and HALF of the B tile)
constexpr int A_BYTES = BM * BK * 2;            // 16 KB
constexpr int B_BYTES = BN * BK * 2;            // 32 KB
constexpr int STAGE_BYTES = A_BYTES + B_BYTES;  // 48 KB
constexpr int STAGE2_BYTES = A_BYTES + B_BYTES / 2;  // 32 KB
constexpr int TMEM_COLS = 512;
constexpr int KTOP = 8;  // register-resident top-k capacity of the TOPK epilogue
// Epilogue warp groups: 12 warps (3 per SM sub-partition) share the 8 column chunks (32 columns each) of a tile 3 / 3 / 2.
// With 2 groups (4 chunks per warp) the epilogue of a tile took longer than its MMAs -- the tensor pipe was 71 % (weights)
// and 84 % (statistics) active, the epilogue warps stalled on shared-memory and TMEM latencies with 2 warps per scheduler.
constexpr int NG = kSimNG;
constexpr int NTHREADS = 128 + NG * 128;
__device__ __forceinline__ int chunk_lo(int grp) { return grp == 0 ? 0 : (grp == 1 ? 3 : 6); }
__device__ __forceinline__ int chunk_hi(int grp) { return grp == 0 ? 3 : (grp == 1 ? 6 : 8); }
static_assert(NG == 3 && kTcBNF == 256, "chunk_lo / chunk_hi split 8 chunks of 32 columns over 3 groups");

enum { EPI_STATS = 0, EPI_TOPK = 1, EPI_WEIGHTS = 2, EPI_WEIGHTS_X3 = 3 };
__host__ __device__ constexpr bool epi_weights(int epi) { return epi == EPI_WEIGHTS || epi == EPI_WEIGHTS_X3; }

struct SimParams {
  CUtensorMap tmap[3];  // per operand matrix: [rows][d], box {64, 128}
  // work decomposition: item = (pair, jchunk, iblk), iblk fastest
  int n_pairs;
  int pair_a[6], pair_b[6];  // unordered pairs (forward, single-GPU weights) or ordered pairs (multi-GPU weights)
  int n_iblk, n_jtile, n_jchunk, jpc;
  int n_iblk_items;  // row blocks per work item index: n_iblk (cta_group::1) or n_iblk / 2 (CTA pairs)
  int num_items;
  int kblocks;  // d / 64 (BSC_MODE_TC16X3: 3 d / 64)
  // BSC_MODE_TC16X3 (x3_kt = d / 64, else 0): the operand rows are [hi | lo] of width 2 d and the K loop runs the three
  // terms lo_a * hi_b, hi_a * lo_b, hi_a * hi_b one after the other -- the small terms first: the accumulator seems to be
  // rounded toward zero at every step, which biases a sum that sits at its full size while small terms are still added
  int x3_kt, x3_d;
  int64_t a_rows;    // valid A rows (local rows of the stripe)
  int64_t a_row0;    // first A row inside the A matrix
  int64_t b_rows;    // valid B rows
  // EPI_STATS
  float k1, k2;      // exp(s*x - s) = ex2(x*k1 - k2)
  float* rowpart;    // [pair][jchunk][a_pad]
  float* colpart;    // [pair][iblk][b_pad]
  float* rowpos;     // [pair][jchunk][a_pad]  sum_j T_ij * (S_ij*k1 - k2)   (log2 units)
  const int32_t* rep;   // [b_rows] class representative of every row of the global batch (T_ij = rep_i == rep_j)
  const float* inv[3];  // per operand matrix: inverse row norms (EPI_STATS: cos_ij = <a_i, b_j> * inv_i * inv_j)
  const int2* jrange;   // [n_jtile] {min, max} of rep over the columns of a tile
  int64_t a_pad, b_pad;
  // EPI_TOPK
  int k;
  int64_t base_id;
  // At most 64 queries (spread = 8 or 16, else 0): the queries are laid out `spread` per TMEM lane quarter (query r ->
  // tile row 32 (r / spread) + r % spread; tmap[2] = the query matrix with a box of `spread` rows), so that all four
  // epilogue warps of a column group keep lists instead of one: a warp takes the insert path for the union of its
  // lanes' columns, and 8 queries per warp make that union 4 times sparser than 32 (BSC_SCAN_V2 >= 2).
  int spread;
  float* cand_s;     // [jchunk][a_rows][k]
  int64_t* cand_i;
  // EPI_WEIGHTS: the fp16 weight block of every pair, G_ij = gs (e_ij (u_i + v_j) - 2 T_ij) / |b_j|, written as
  // 128B-swizzled [128 i][64 j] images (K-major A operand over j, MN-major A operand over i: tc_bwd3.cu)
  const float* uv;     // [6][N]
  const float* gsbuf;  // [16] = gs of the stored weights
  uint8_t* gimg;       // [pair][n_iblk_pad][n_j64][16 KB]
  uint8_t* gimg_lo;    // EPI_WEIGHTS_X3: the same images of the fp16 remainders g - fp16(g)
  int n_iblk_pad, n_j64;
  // Fused all-gather (EPI_STATS, several ranks): the operand rows of source rank s (rows [s * src_rows, (s+1) * src_rows)
  // of every matrix, their inverse norms) are complete once arrive[s] >= arrive_target; null = complete on entry.
  // Items are then ordered (arrival position of the column chunk, pair, row block): the sweep starts on the rank's own
  // columns while the peers' rows are still in flight (gather_push_kernel, ordered mode).
  const unsigned long long* arrive;
  unsigned long long arrive_target;
  int src_rows, my_rank, world;
  int chunk_major, cps;  // cps = column chunks per source rank
  // The weights carry 1 / |b_j| and gs of the ordered pair (a, b): the K-major GEMM's B operand is then the scaled fp16
  // rows themselves (exact for bf16 inputs), the MN-major GEMM undoes the column factor per output row (tc_bwd3.cu).
#ifdef BSC_EXPERIMENTS
  int dbg;  // experiment knob BSC_DBG_W: 1 = no stores, 2 = no label compare (results invalid)
#endif
};

struct __align__(8) SimSmemTail {
  uint64_t full[STAGES2];
  uint64_t empty[STAGES2];
  uint64_t tmem_full[2];
  uint64_t tmem_empty[2];
  uint32_t tmem_base;
  uint32_t pad;
  float cs[2][4][BN];  // per-warp column partial sums, double-buffered by tile parity
  float invb[2][BN];   // k1 * inverse norms of the tile's columns, double-buffered by tile parity
  __align__(16) uint8_t wst[4 * NG][2048];  // EPI_WEIGHTS: per-warp staging of 32 half rows (64 bytes each)
};

constexpr size_t SIM_SMEM_BYTES = (size_t)STAGES * STAGE_BYTES + sizeof(SimSmemTail);
static_assert(SIM_SMEM_BYTES <= 232448, "similarity sweep exceeds 227 KB of shared memory");
static_assert(STAGES2 * STAGE2_BYTES == STAGES * STAGE_BYTES, "both ring shapes use the same shared memory");

__device__ __forceinline__ void decode_item(const SimParams& p, int item, int& pair, int& jchunk, int& iblk) {
  iblk = item % p.n_iblk_items;
  const int rest = item / p.n_iblk_items;
  if (p.chunk_major) {
    pair = rest % p.n_pairs;
    const int pos = rest / p.n_pairs;  // arrival position: this rank's own columns first, then rank - 1, rank - 2, ...
    const int src = (p.my_rank - pos / p.cps + p.world) % p.world;
    jchunk = src * p.cps + pos % p.cps;
  } else {
    jchunk = rest % p.n_jchunk;
    pair = rest / p.n_jchunk;
  }
}

// Block until the rows of source rank `src` have landed in this rank's buffer (acquire at system scope pairs with the
// red.release of the sender's last store).  Bounded: a lost sender must surface as a launch failure, not as a hung GPU.
__device__ __forceinline__ void wait_source(const SimParams& p, int src, uint32_t& seen) {
  if (seen & (1u << src)) return;
  const unsigned long long* f = p.arrive + src;
  unsigned long long v;
  asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(f) : "memory");
  if (v < p.arrive_target) {
    const long long t0 = clock64();
    do {
      __nanosleep(64);
      asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(f) : "memory");
      if (clock64() - t0 > 8000000000LL) {
        printf("bsc: rows of rank %d never arrived (counter %llu, expected %llu) block=%d\n", src, v, p.arrive_target,
               (int)blockIdx.x);
        __trap();
      }
    } while (v < p.arrive_target);
  }
  seen |= 1u << src;
}

// Butterfly transpose-reduce: on entry every lane holds 32 column values of its own row; on exit lane l
// holds the sum over the warp's 32 rows of column l.  31 shuffles.
__device__ __forceinline__ float warp_colsum32(float (&v)[32], int lane) {
#pragma unroll
  for (int off = 16; off >= 1; off >>= 1) {
    const bool upper = (lane & off) != 0;
#pragma unroll
    for (int k = 0; k < off; ++k) {
      const float send = upper ? v[k] : v[k + off];
      const float keep = upper ? v[k + off] : v[k];
      v[k] = keep + __shfl_xor_sync(0xffffffffu, send, off);
    }
  }
  return v[0];
}

// EPI_WEIGHTS_X3: 32 rows x 32 columns of fp16 pairs (pk = this lane's row, 16 x half2) -> half rows (64 bytes) of a
// 128B-swizzled [128 i][64 j] image, staged through 2 KB of shared memory per warp so that four lanes write the 64
// contiguous bytes of one row (same scheme as the EPI_WEIGHTS branch of the kernel).
__device__ __forceinline__ void store_image_half_rows(uint8_t* stg, const uint32_t (&pk)[16], uint8_t* img_base, int c,
                                                      int q, int lane) {
  {
    const int r = lane;
#pragma unroll
    for (int u = 0; u < 4; ++u)
      *reinterpret_cast<uint4*>(stg + r * 64 + ((u ^ ((r >> 1) & 3)) << 4)) =
          make_uint4(pk[4 * u], pk[4 * u + 1], pk[4 * u + 2], pk[4 * u + 3]);
  }
  __syncwarp();
  const int t4 = lane & 3;
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const int r = 8 * k + (lane >> 2);
    const int rt = q * 32 + r;
    const int u = t4 ^ (r & 3);
    const uint4 v = *reinterpret_cast<const uint4*>(stg + r * 64 + ((u ^ ((r >> 1) & 3)) << 4));
    const int half_off = (((c & 1) * 4) ^ (rt & 4)) << 4;
    __stcs(reinterpret_cast<uint4*>(img_base + (rt >> 3) * 1024 + (rt & 7) * 128 + half_off + (t4 << 4)), v);
  }
  __syncwarp();
}

// PAIR = 1: the two CTAs of a 2-CTA cluster own two adjacent row blocks of the SAME column chunk and run every MMA as
// one tcgen05.mma.cta_group::2 (M = 256, N = 256): each CTA stages its own 128 A rows and half of the B tile, which
// cuts the shared-memory operand bytes per MMA cycle from 96 to 64 per SM (measured: the cta_group::1 sweep is bound by
// operand fetch at ~62 % tensor-pipe activity).  The epilogue is unchanged: every CTA reads its own 128 x 256 fp32 tile.
template <int EPI, int FMT, int PAIR>
__global__ void __launch_bounds__(NTHREADS, 1) tc_sim_kernel(const __grid_constant__ SimParams p) {
  constexpr int NST = PAIR ? STAGES2 : STAGES;
  constexpr int SB = PAIR ? STAGE2_BYTES : STAGE_BYTES;
  const uint32_t crank = PAIR ? ptx::cluster_ctarank() : 0u;
  const int item0 = PAIR ? (int)ptx::cluster_id_x() : (int)blockIdx.x;
  const int item_step = PAIR ? (int)ptx::n_clusters_x() : (int)gridDim.x;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  const uint32_t ring = ptx::smem_u32(smem_raw);
  if ((ring & 1023u) != 0) {
    if (threadIdx.x == 0) printf("bsc: dynamic shared memory is not 1024-byte aligned (0x%x)\n", ring);
    __trap();
  }
  uint8_t* ring_ptr = smem_raw;
  SimSmemTail* tail = reinterpret_cast<SimSmemTail*>(ring_ptr + (size_t)NST * SB);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    for (int s = 0; s < NST; ++s) {
      ptx::mbar_init(ptx::smem_u32(&tail->full[s]), 1);
      ptx::mbar_init(ptx::smem_u32(&tail->empty[s]), 1);
    }
    for (int b = 0; b < 2; ++b) {
      ptx::mbar_init(ptx::smem_u32(&tail->tmem_full[b]), 1);
      ptx::mbar_init(ptx::smem_u32(&tail->tmem_empty[b]), (PAIR ? 2 : 1) * 4 * NG);
    }
    ptx::fence_barrier_init();
  }
  if (warp == 0 && lane == 0) {
    ptx::prefetch_tmap(&p.tmap[0]);
    ptx::prefetch_tmap(&p.tmap[1]);
    ptx::prefetch_tmap(&p.tmap[2]);
  }
  if (warp == 2) {
    if (PAIR) {
      ptx::tmem_alloc_2cta(ptx::smem_u32(&tail->tmem_base), TMEM_COLS);
      ptx::tmem_relinquish_2cta();
    } else {
      ptx::tmem_alloc(ptx::smem_u32(&tail->tmem_base), TMEM_COLS);
      ptx::tmem_relinquish();
    }
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (PAIR) ptx::cluster_sync_all();  // the peer's barriers and TMEM exist before anything crosses the pair
  ptx::tc_fence_after();
  const uint32_t tmem_base = tail->tmem_base;

  if (warp == 0) {
    // ================= TMA producer (one elected thread) =================
    if (ptx::elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      uint32_t seen = 0;
      for (int item = item0; item < p.num_items; item += item_step) {
        int pair, jchunk, iblk;
        decode_item(p, item, pair, jchunk, iblk);
        if (PAIR) iblk = 2 * iblk + (int)crank;
        const CUtensorMap* ma = &p.tmap[p.pair_a[pair]];
        const CUtensorMap* mb = &p.tmap[p.pair_b[pair]];
        const int arow = (int)(p.a_row0 + (int64_t)iblk * BM);
        const int jt_end = min(p.n_jtile, (jchunk + 1) * p.jpc);
        if (EPI == EPI_STATS && p.arrive != nullptr) {
          // the A rows are this rank's own, the B rows of a chunk all come from one source (chunks do not straddle ranks)
          const uint32_t before = seen;
          wait_source(p, p.my_rank, seen);
          wait_source(p, jchunk / p.cps, seen);
          // the rows were written through the generic proxy (by another GPU); TMA reads them through the async proxy
          if (seen != before) ptx::fence_proxy_async_all();
        }
        for (int jt = jchunk * p.jpc; jt < jt_end; ++jt) {
          const int brow = jt * BN;
          for (int kb = 0; kb < p.kblocks; ++kb) {
            ptx::mbar_wait(ptx::smem_u32(&tail->empty[stage]), phase ^ 1u, 100 + stage);
            const uint32_t full = ptx::smem_u32(&tail->full[stage]);
            const uint32_t sa = ring + stage * SB;
            int ca = kb * BK, cb = kb * BK;  // first column of the A / B box
            if (p.x3_kt > 0) {
              const int term = kb / p.x3_kt, kk = (kb - term * p.x3_kt) * BK;
              ca = kk + (term == 0 ? p.x3_d : 0);  // term 0: lo_a * hi_b
              cb = kk + (term == 1 ? p.x3_d : 0);  // term 1: hi_a * lo_b, term 2: hi_a * hi_b
            }
            if (PAIR) {
              // both CTAs' bytes complete on the leader's barrier; the leader expects the total
              const uint32_t full_l = ptx::mapa(full, 0);
              if (crank == 0) ptx::mbar_arrive_expect_tx(full, 2 * SB);
              ptx::tma_load_2d_2cta(sa, ma, full_l, ca, arow);
              ptx::tma_load_2d_2cta(sa + A_BYTES, mb, full_l, cb, brow + 128 * (int)crank);
            } else {
              if (EPI == EPI_TOPK && p.spread > 0) {
                // `spread` query rows per lane quarter: 8-row atoms (1024 B) 4 g, 4 g + 1 of the A tile; the rows in
                // between are never written and never read back (their accumulator rows belong to no query)
                ptx::mbar_arrive_expect_tx(full, 4u * (uint32_t)p.spread * 128u + B_BYTES);
#pragma unroll
                for (int g = 0; g < 4; ++g) ptx::tma_load_2d(sa + g * 4096, &p.tmap[2], full, ca, arow + g * p.spread);
              } else {
                ptx::mbar_arrive_expect_tx(full, STAGE_BYTES);
                ptx::tma_load_2d(sa, ma, full, ca, arow);
              }
              ptx::tma_load_2d(sa + A_BYTES, mb, full, cb, brow);
              ptx::tma_load_2d(sa + A_BYTES + B_BYTES / 2, mb, full, cb, brow + 128);
            }
            if (++stage == NST) {
              stage = 0;
              phase ^= 1u;
            }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ================= MMA issuer (one elected thread runs the whole role) =================
    if ((!PAIR || crank == 0) && ptx::elect_one()) {
      constexpr uint32_t idesc = umma_idesc(PAIR ? 2 * BM : BM, BN, FMT, FMT, 0, 0);
      // K-major SWIZZLE_128B descriptors differ only in the start-address field: desc(addr) = base + (addr >> 4)
      const uint64_t desc0 = ptx::umma_desc_sw128(0, 16, 1024);
      int stage = 0;
      uint32_t phase = 0;
      uint32_t tile_count = 0;
      for (int item = item0; item < p.num_items; item += item_step) {
        int pair, jchunk, iblk;
        decode_item(p, item, pair, jchunk, iblk);
        const int jt_end = min(p.n_jtile, (jchunk + 1) * p.jpc);
        for (int jt = jchunk * p.jpc; jt < jt_end; ++jt, ++tile_count) {
          const uint32_t buf = tile_count & 1u;
          const uint32_t use = tile_count >> 1;
          if (PAIR)
            ptx::mbar_wait_cluster(ptx::smem_u32(&tail->tmem_empty[buf]), (use & 1u) ^ 1u, 200 + buf);
          else
            ptx::mbar_wait(ptx::smem_u32(&tail->tmem_empty[buf]), (use & 1u) ^ 1u, 200 + buf);
          ptx::tc_fence_after();
          const uint32_t d_tmem = tmem_base + buf * BN;
          for (int kb = 0; kb < p.kblocks; ++kb) {
            ptx::mbar_wait(ptx::smem_u32(&tail->full[stage]), phase, 300 + stage);
            ptx::tc_fence_after();
            const uint32_t sa = ring + stage * SB;
            const uint64_t da = desc0 + ptx::desc_addr(sa);
            const uint64_t db = desc0 + ptx::desc_addr(sa + A_BYTES);
            if (PAIR) {
#pragma unroll
              for (int k = 0; k < BK / 16; ++k)
                ptx::umma_f16_2cta(d_tmem, da + 2 * k, db + 2 * k, idesc, (kb | k) != 0 ? 1u : 0u);
              ptx::umma_commit_2cta(ptx::smem_u32(&tail->empty[stage]), 3);
              if (kb == p.kblocks - 1) ptx::umma_commit_2cta(ptx::smem_u32(&tail->tmem_full[buf]), 3);
            } else {
#pragma unroll
              for (int k = 0; k < BK / 16; ++k)
                ptx::umma_f16(d_tmem, da + 2 * k, db + 2 * k, idesc, (kb | k) != 0 ? 1u : 0u);
              ptx::umma_commit(ptx::smem_u32(&tail->empty[stage]));
              if (kb == p.kblocks - 1) ptx::umma_commit(ptx::smem_u32(&tail->tmem_full[buf]));
            }
            if (++stage == NST) {
              stage = 0;
              phase ^= 1u;
            }
          }
        }
      }
    }
  } else if (warp >= 4) {
    // ================= epilogue (4 warps, thread <-> TMEM lane) =================
    const int q = warp & 3;            // TMEM lane quarter this warp may read
    const int grp = (warp - 4) >> 2;   // column group of the tile
    const int col_lo = 32 * chunk_lo(grp), col_hi = 32 * chunk_hi(grp);  // this group's columns of every tile
    const int row_in_tile = q * 32 + lane;
    const int et = (threadIdx.x - 128) & 127;  // 0..127 inside the group
    uint32_t tile_count = 0;
    uint32_t seen_e = 0;
    const uint32_t empty_l0 = PAIR ? ptx::mapa(ptx::smem_u32(&tail->tmem_empty[0]), 0) : 0u;
    for (int item = item0; item < p.num_items; item += item_step) {
      int pair, jchunk, iblk;
      decode_item(p, item, pair, jchunk, iblk);
      if (PAIR) iblk = 2 * iblk + (int)crank;
      int64_t arow = (int64_t)iblk * BM + row_in_tile;  // row inside the stripe
      bool row_ok = arow < p.a_rows;
      if (EPI == EPI_TOPK && p.spread > 0) {  // queries spread over the lane quarters (iblk == 0)
        arow = q * p.spread + lane;
        row_ok = lane < p.spread && arow < p.a_rows;
      }
      const int jt_end = min(p.n_jtile, (jchunk + 1) * p.jpc);

      float rowsum = 0.f;
      float rowpos = 0.f;
      int rep_i = -2, wmin = 0, wmax = -1;
      float inv_i = 0.f;
      const float* invb_g = nullptr;
      if (EPI == EPI_STATS && p.arrive != nullptr) {
        // the inverse norms of the rows and of the chunk's columns travel with the operand rows
        if (lane == 0) {
          wait_source(p, p.my_rank, seen_e);
          wait_source(p, jchunk / p.cps, seen_e);
        }
        __syncwarp();
      }
      if (EPI == EPI_STATS) {
        if (row_ok) inv_i = p.inv[p.pair_a[pair]][p.a_row0 + arow];
        invb_g = p.inv[p.pair_b[pair]];
        // label class of this thread's row and the class range of the warp's 32 rows
        if (row_ok) rep_i = p.rep[p.a_row0 + arow];
        wmin = row_ok ? rep_i : 0x7fffffff;
        wmax = row_ok ? rep_i : -1;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
          wmin = min(wmin, __shfl_xor_sync(0xffffffffu, wmin, o));
          wmax = max(wmax, __shfl_xor_sync(0xffffffffu, wmax, o));
        }
      }
      float ugs_i = 0.f, gsw = 0.f;
      const float* vcol_g = nullptr;
      if (epi_weights(EPI)) {
        gsw = p.gsbuf[2 * pair_id(p.pair_a[pair], p.pair_b[pair])];
        if (row_ok) {
          inv_i = p.inv[p.pair_a[pair]][p.a_row0 + arow];
          rep_i = p.rep[p.a_row0 + arow];
          ugs_i = gsw * p.uv[(int64_t)pair_id(p.pair_a[pair], p.pair_b[pair]) * p.b_rows + p.a_row0 + arow];
        }
        invb_g = p.inv[p.pair_b[pair]];
        vcol_g = p.uv + (int64_t)pair_id(p.pair_b[pair], p.pair_a[pair]) * p.b_rows;
        wmin = row_ok ? rep_i : 0x7fffffff;
        wmax = row_ok ? rep_i : -1;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
          wmin = min(wmin, __shfl_xor_sync(0xffffffffu, wmin, o));
          wmax = max(wmax, __shfl_xor_sync(0xffffffffu, wmax, o));
        }
      }
      float ts[KTOP];
      int ti[KTOP];
      if (EPI == EPI_TOPK) {
#pragma unroll
        for (int t = 0; t < KTOP; ++t) {
          ts[t] = -3.4028234663852886e38f;
          ti[t] = -1;
        }
      }

      for (int jt = jchunk * p.jpc; jt < jt_end; ++jt, ++tile_count) {
        const uint32_t buf = tile_count & 1u;
        const uint32_t use = tile_count >> 1;
        const int64_t col0 = (int64_t)jt * BN;
        if (EPI == EPI_STATS) {
          // stage k1 / ||b_j|| of the tile's columns (2 per thread) while the MMAs of this tile are in flight
          for (int col = col_lo + et; col < col_hi; col += 128) {
            const int64_t j = col0 + col;
            tail->invb[buf][col] = j < p.b_rows ? p.k1 * invb_g[j] : 0.f;
          }
          ptx::named_bar_sync(2 + 2 * grp, 128);
        }
        if (epi_weights(EPI)) {
          // column facts of the tile: k1 / |b_j|, gs * v_j, label class (the column-sum scratch is free in this mode)
          float* vgb = &tail->cs[0][0][0] + buf * BN;
          float* wfb = &tail->cs[0][2][0] + buf * BN;
          int* repb = reinterpret_cast<int*>(&tail->cs[1][0][0]) + buf * BN;
          for (int col = col_lo + et; col < col_hi; col += 128) {
            const int64_t j = col0 + col;
            const bool ok = j < p.b_rows;
            tail->invb[buf][col] = ok ? p.k1 * invb_g[j] : 0.f;
            const float wf = ok ? invb_g[j] : 0.f;  // column factor of the weights
            wfb[col] = wf;
            vgb[col] = ok ? gsw * vcol_g[j] * wf : 0.f;
            repb[col] = ok ? p.rep[j] : -1;
          }
          ptx::named_bar_sync(2 + 2 * grp, 128);
        }
        ptx::mbar_wait(ptx::smem_u32(&tail->tmem_full[buf]), use & 1u, 400 + buf);
        ptx::tc_fence_after();
        const bool edge = (col0 + BN > p.b_rows) || ((int64_t)iblk * BM + BM > p.a_rows);
        const bool col_edge = col0 + BN > p.b_rows;
        const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + buf * BN;
        float* cs = tail->cs[buf][q];
        const float* ivb = tail->invb[buf];
        bool tile_pos = false;  // can any (row of this warp, column of this tile) pair share a label?
        if (EPI == EPI_STATS || epi_weights(EPI)) {
          const int2 jr = p.jrange[jt];
          tile_pos = (wmax >= jr.x) && (wmin <= jr.y);
        }
#pragma unroll 1
        for (int c = chunk_lo(grp); c < chunk_hi(grp); ++c) {
          uint32_t raw[32];
          ptx::tmem_ld_32x32(taddr + c * 32, raw);
          ptx::tmem_ld_wait();
          if (EPI == EPI_STATS) {
            float e[32];
            if (tile_pos) {
              // positive-pair term from the very logits that enter the exponentials, so that the
              // r_i*log(sumexp_i) - sum_j T_ij a_ij difference cancels like log_softmax does
              const int64_t jcol = col0 + c * 32 + lane;
              const int repv = jcol < p.b_rows ? p.rep[jcol] : -1;
#pragma unroll
              for (int t = 0; t < 32; ++t) {
                const float a = fmaf(__uint_as_float(raw[t]), inv_i * ivb[c * 32 + t], -p.k2);
                const int rj = __shfl_sync(0xffffffffu, repv, t);
                rowpos += (rj == rep_i) ? a : 0.f;
                e[t] = ptx::ex2(a);
              }
            } else {
#pragma unroll
              for (int t = 0; t < 32; ++t)
                e[t] = ptx::ex2(fmaf(__uint_as_float(raw[t]), inv_i * ivb[c * 32 + t], -p.k2));
            }
            if (edge) {
#pragma unroll
              for (int t = 0; t < 32; ++t) {
                const bool ok = row_ok && (col0 + c * 32 + t < p.b_rows);
                e[t] = ok ? e[t] : 0.f;
              }
            }
            float s = 0.f;
#pragma unroll
            for (int t = 0; t < 32; ++t) s += e[t];
            rowsum += s;
            cs[c * 32 + lane] = warp_colsum32(e, lane);
          } else if (EPI == EPI_WEIGHTS) {
            const float* vgb = &tail->cs[0][0][0] + buf * BN + c * 32;
            const float* wfb = &tail->cs[0][2][0] + buf * BN + c * 32;
            const int* repb = reinterpret_cast<const int*>(&tail->cs[1][0][0]) + buf * BN + c * 32;
            const float m2g = -2.0f * gsw;
            uint32_t pk[16];
#ifdef BSC_EXPERIMENTS
            const bool want_pos = tile_pos && p.dbg != 2;
#else
            const bool want_pos = tile_pos;
#endif
            if (want_pos) {  // some (row of this warp, column of this tile) pair may share a label
#pragma unroll
              for (int t = 0; t < 32; t += 2) {
                float g2[2];
#pragma unroll
                for (int z = 0; z < 2; ++z) {
                  const float ex = ptx::ex2(fmaf(__uint_as_float(raw[t + z]), inv_i * ivb[c * 32 + t + z], -p.k2));
                  const float tv = (repb[t + z] == rep_i) ? m2g * wfb[t + z] : 0.f;
                  g2[z] = fmaf(ex, fmaf(ugs_i, wfb[t + z], vgb[t + z]), tv);
                }
                const __half2 hh = __floats2half2_rn(g2[0], g2[1]);
                pk[t >> 1] = *reinterpret_cast<const uint32_t*>(&hh);
              }
            } else {
#pragma unroll
              for (int t = 0; t < 32; t += 2) {
                float g2[2];
#pragma unroll
                for (int z = 0; z < 2; ++z) {
                  const float ex = ptx::ex2(fmaf(__uint_as_float(raw[t + z]), inv_i * ivb[c * 32 + t + z], -p.k2));
                  g2[z] = ex * fmaf(ugs_i, wfb[t + z], vgb[t + z]);
                }
                const __half2 hh = __floats2half2_rn(g2[0], g2[1]);
                pk[t >> 1] = *reinterpret_cast<const uint32_t*>(&hh);
              }
            }
            // Half a row (64 bytes) of image (iblk, 4 jt + c / 2) per lane.  Staged through 2 KB of shared memory per
            // warp so that four lanes write the 64 contiguous bytes of one row: one 16-byte store per lane to 32
            // different lines kept the load/store unit busy for two thirds of a tile's MMA time (5.5 ms, 4.0 without
            // the stores).  Staging slot of (row r, chunk u) = u ^ ((r >> 1) & 3): conflict-free both ways.
            uint8_t* stg = tail->wst[warp - 4];
            {
              const int r = lane;
#pragma unroll
              for (int u = 0; u < 4; ++u)
                *reinterpret_cast<uint4*>(stg + r * 64 + ((u ^ ((r >> 1) & 3)) << 4)) =
                    make_uint4(pk[4 * u], pk[4 * u + 1], pk[4 * u + 2], pk[4 * u + 3]);
            }
            __syncwarp();
            const int64_t img = ((int64_t)pair * p.n_iblk_pad + iblk) * p.n_j64 + (int64_t)jt * 4 + (c >> 1);
            uint8_t* img_base = p.gimg + img * 16384;
            const int t4 = lane & 3;
#pragma unroll
            for (int k = 0; k < 4; ++k) {
              const int r = 8 * k + (lane >> 2);        // row inside the warp's 32
              const int rt = q * 32 + r;                // row inside the tile = row of the image
              const int u = t4 ^ (r & 3);               // the chunk that belongs at 16-byte slot t4 of the row's half
              const uint4 v = *reinterpret_cast<const uint4*>(stg + r * 64 + ((u ^ ((r >> 1) & 3)) << 4));
              const int half_off = (((c & 1) * 4) ^ (rt & 4)) << 4;  // 128B swizzle: slot = chunk ^ (row & 7)
#ifdef BSC_EXPERIMENTS
              if (p.dbg != 1 || v.x == 0x12345678u)
#endif
                __stcs(reinterpret_cast<uint4*>(img_base + (rt >> 3) * 1024 + (rt & 7) * 128 + half_off + (t4 << 4)), v);
            }
            __syncwarp();
          } else if (EPI == EPI_WEIGHTS_X3) {
            // float32-grade weights: g is formed as in EPI_WEIGHTS and stored as fp16(g) (image set gimg) and
            // fp16(g - fp16(g)) (image set gimg_lo); the gradient GEMMs add both (tc_bwd3.cu, three terms)
            const float* vgb = &tail->cs[0][0][0] + buf * BN + c * 32;
            const float* wfb = &tail->cs[0][2][0] + buf * BN + c * 32;
            const int* repb = reinterpret_cast<const int*>(&tail->cs[1][0][0]) + buf * BN + c * 32;
            const float m2g = -2.0f * gsw;
#pragma unroll
            for (int t = 0; t < 32; ++t) {
              const float ex = ptx::ex2(fmaf(__uint_as_float(raw[t]), inv_i * ivb[c * 32 + t], -p.k2));
              const float tv = (tile_pos && repb[t] == rep_i) ? m2g * wfb[t] : 0.f;
              raw[t] = __float_as_uint(fmaf(ex, fmaf(ugs_i, wfb[t], vgb[t]), tv));
            }
            const int64_t img = ((int64_t)pair * p.n_iblk_pad + iblk) * p.n_j64 + (int64_t)jt * 4 + (c >> 1);
#pragma unroll 1
            for (int part = 0; part < 2; ++part) {
              uint32_t pk[16];
#pragma unroll
              for (int t = 0; t < 32; t += 2) {
                const float g0 = __uint_as_float(raw[t]), g1 = __uint_as_float(raw[t + 1]);
                __half2 hh = __floats2half2_rn(g0, g1);
                if (part) {
                  const float2 b = __half22float2(hh);
                  hh = __floats2half2_rn(g0 - b.x, g1 - b.y);
                }
                pk[t >> 1] = *reinterpret_cast<const uint32_t*>(&hh);
              }
              store_image_half_rows(tail->wst[warp - 4], pk, (part ? p.gimg_lo : p.gimg) + img * 16384, c, q, lane);
            }
          } else {
            // Scan path (always executed, kept tiny so it stays in the instruction cache): which of the 32 scores beat
            // this row's current KTOP-th best?  Rare path: the union of those columns over the warp is re-read from
            // TMEM one column at a time (uniform address) and inserted, ascending in id so ties keep the lower id.
            const int jbase = (int)(col0 + c * 32);
            const float thr = ts[KTOP - 1];
            uint32_t m = 0;
#pragma unroll
            for (int t = 0; t < 32; ++t) m |= (__uint_as_float(raw[t]) > thr) ? (1u << t) : 0u;
            if (col_edge) {
              // zero-filled rows past the end of the bank never compete
              const int64_t valid = p.b_rows - jbase;
              m &= valid >= 32 ? 0xffffffffu : (valid <= 0 ? 0u : ((1u << (int)valid) - 1u));
            }
#if BSC_SCAN_V2
            if (!row_ok) m = 0;  // rows past the last query (zero-filled by TMA) never compete
#endif
            uint32_t um = __reduce_or_sync(0xffffffffu, m);
#if BSC_SCAN_V2
            // Four adjacent columns per TMEM read: while the lists of a work item are still filling up nearly every
            // column is in the union, and the latency of the read (not the insertion) is what a column costs.
            while (um) {
              const int t0 = (__ffs(um) - 1) & ~3;
              um &= ~(0xFu << t0);
              uint32_t v4[4];
              ptx::tmem_ld_32x32_x4(taddr + c * 32 + t0, v4);
              ptx::tmem_ld_wait();
#pragma unroll
              for (int w = 0; w < 4; ++w) {
                const int t = t0 + w;
                const float s = __uint_as_float(v4[w]);
                if (((m >> t) & 1u) && s > ts[KTOP - 1]) {
                  const int j = jbase + t;
                  int pos = KTOP - 1;
#pragma unroll
                  for (int u = KTOP - 1; u > 0; --u)
                    if (pos == u && s > ts[u - 1]) pos = u - 1;
#pragma unroll
                  for (int u = KTOP - 1; u > 0; --u)
                    if (u > pos) {
                      ts[u] = ts[u - 1];
                      ti[u] = ti[u - 1];
                    }
#pragma unroll
                  for (int u = 0; u < KTOP; ++u)
                    if (u == pos) {
                      ts[u] = s;
                      ti[u] = j;
                    }
                }
              }
            }
#else
            while (um) {
              const int t = __ffs(um) - 1;
              um &= um - 1;
              const float s = __uint_as_float(ptx::tmem_ld_32x32_x1(taddr + c * 32 + t));
              ptx::tmem_ld_wait();
              if (((m >> t) & 1u) && s > ts[KTOP - 1]) {
                const int j = jbase + t;
                int pos = KTOP - 1;
#pragma unroll
                for (int u = KTOP - 1; u > 0; --u)
                  if (pos == u && s > ts[u - 1]) pos = u - 1;
#pragma unroll
                for (int u = KTOP - 1; u > 0; --u)
                  if (u > pos) {
                    ts[u] = ts[u - 1];
                    ti[u] = ti[u - 1];
                  }
#pragma unroll
                for (int u = 0; u < KTOP; ++u)
                  if (u == pos) {
                    ts[u] = s;
                    ti[u] = j;
                  }
              }
            }
#endif
          }
        }
        // release the accumulator buffer to the MMA warp
        ptx::tc_fence_before();
        __syncwarp();
        if (lane == 0) {
          if (PAIR)
            ptx::mbar_arrive_remote_nodata(empty_l0 + buf * 8u);  // the leader's barrier counts both CTAs' epilogue warps
          else
            ptx::mbar_arrive(ptx::smem_u32(&tail->tmem_empty[buf]));
        }

        if (EPI == EPI_STATS) {
          ptx::named_bar_sync(1 + 2 * grp, 128);
          float* out = p.colpart + ((int64_t)pair * p.n_iblk + iblk) * p.b_pad + col0;
          for (int col = col_lo + et; col < col_hi; col += 128) {
            if (col0 + col < p.b_rows)
              out[col] = (tail->cs[buf][0][col] + tail->cs[buf][1][col]) +
                         (tail->cs[buf][2][col] + tail->cs[buf][3][col]);
          }
        }
      }

      if (EPI == EPI_STATS) {
        if (row_ok) {
          p.rowpart[(((int64_t)pair * p.n_jchunk + jchunk) * NG + grp) * p.a_pad + arow] = rowsum;
          p.rowpos[(((int64_t)pair * p.n_jchunk + jchunk) * NG + grp) * p.a_pad + arow] = rowpos;
        }
      } else if (EPI == EPI_TOPK) {
        if (row_ok) {
          const int64_t o = (((int64_t)jchunk * NG + grp) * p.a_rows + arow) * p.k;
#pragma unroll
          for (int t = 0; t < KTOP; ++t) {
            if (t < p.k) {
              p.cand_s[o + t] = ts[t];
              p.cand_i[o + t] = ti[t] < 0 ? (int64_t)-1 : p.base_id + ti[t];
            }
          }
        }
      }
    }
  }

  ptx::tc_fence_before();
  __syncthreads();
  if (PAIR) ptx::cluster_sync_all();
  if (warp == 2) {
    ptx::tc_fence_after();
    if (PAIR)
      ptx::tmem_dealloc_2cta(tmem_base, TMEM_COLS);
    else
      ptx::tmem_dealloc(tmem_base, TMEM_COLS);
  }
}

inline int num_sms() { return device_sm_count(); }

template <int EPI, int FMT>
int launch_sim(const SimParams& p_in, cudaStream_t st) {
  SimParams p = p_in;
  p.n_iblk_items = p.n_iblk;
  if (p.num_items <= 0) return BSC_OK;
  // CTA pairs (cta_group::2, M = 256): a third less operand traffic per SM (58 -> 39 GB from L2 per sweep), and the same
  // ring covers 50 % more MMA time.  Measured (B200, N = 32768): bank scan +1 %, weights sweep 5.1 -> 4.3 ms.  The
  // forward sweep timed alone after an idle gap was 9 % slower on pairs (3.85 vs 3.54 ms), but inside the step, where
  // every kernel runs at the power cap, it is 5 % faster (4.20 -> 3.97 ms, three alternating runs) -- pairs everywhere.
  const bool pair_ok = true;
  // CTA pairs need an even number of row blocks (the two CTAs of a pair take blocks 2m and 2m+1 of one column chunk)
  if (pair_ok && p.n_iblk >= 2 && (p.n_iblk % 2) == 0 && num_sms() >= 2) {
    auto kern = tc_sim_kernel<EPI, FMT, 1>;
    BSC_SET_SMEM_ONCE(kern, SIM_SMEM_BYTES);
    p.n_iblk_items = p.n_iblk / 2;
    p.num_items = p.num_items / 2;
    int clusters = num_sms() / 2;
    if (clusters > p.num_items) clusters = p.num_items;
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3((unsigned)(2 * clusters));
    cfg.blockDim = dim3(NTHREADS);
    cfg.dynamicSmemBytes = SIM_SMEM_BYTES;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    BSC_CUDA(cudaLaunchKernelEx(&cfg, kern, p));
    BSC_LAUNCH_CHECK("tc_sim_kernel (CTA pairs)");
    return BSC_OK;
  }
  auto kern = tc_sim_kernel<EPI, FMT, 0>;
  BSC_SET_SMEM_ONCE(kern, SIM_SMEM_BYTES);
  const int grid = p.num_items < num_sms() ? p.num_items : num_sms();
  kern<<<grid, NTHREADS, SIM_SMEM_BYTES, st>>>(p);
  BSC_LAUNCH_CHECK("tc_sim_kernel");
  return BSC_OK;
}

}  // namespace

// Choose how many column tiles one work item sweeps so that there are several items per SM but the
// row-partial buffer stays small.
static int pick_jpc(int n_iblk, int n_jtile, int n_pairs) {
  const int target_items = num_sms() * 4;
  int jpc = (int)(((int64_t)n_pairs * n_iblk * n_jtile + target_items - 1) / target_items);
  if (jpc < 1) jpc = 1;
  if (jpc > n_jtile) jpc = n_jtile;
  return jpc;
}

// Several ranks with row blocks that are whole column tiles (n_local % 256 == 0): column chunks never straddle two
// source ranks (what the fused all-gather needs), and the chunk length is the divisor of the tiles per source that
// wastes the least of the last round of items.
bool tc_fwd_source_aligned(int64_t n_local, int64_t n_global, int world) {
  return world > 1 && n_local > 0 && n_local % BN == 0 && n_global == (int64_t)world * n_local;
}

void tc_fwd_tiling(int64_t n_local, int64_t n_global, int n_pairs, int world, int* n_iblk, int* n_jtile, int* jpc,
                   int* n_jchunk) {
  *n_iblk = (int)((n_local + BM - 1) / BM);
  *n_jtile = (int)((n_global + BN - 1) / BN);
  if (tc_fwd_source_aligned(n_local, n_global, world)) {
    const int tps = (int)(n_local / BN);
    const int np = n_pairs > 0 ? n_pairs : 1;
    const int units = (*n_iblk % 2 == 0 && *n_iblk >= 2) ? num_sms() / 2 : num_sms();
    const int iblk_items = (*n_iblk % 2 == 0 && *n_iblk >= 2) ? *n_iblk / 2 : *n_iblk;
    int best = tps;
    double best_cost = 1e300;
    for (int dv = 1; dv <= tps; ++dv) {
      if (tps % dv) continue;
      const int64_t items = (int64_t)np * world * (tps / dv) * iblk_items;
      if (items > 64ll * units) continue;  // keep the partial-sum buffers small
      const double rounds = (double)((items + units - 1) / units);
      const double cost = rounds * (dv + 0.5);  // tiles per item + half a tile of fill / drain
      if (cost < best_cost * 0.999) {
        best_cost = cost;
        best = dv;
      }
    }
    *jpc = best;
    *n_jchunk = world * (tps / best);
    return;
  }
  *jpc = pick_jpc(*n_iblk, *n_jtile, n_pairs > 0 ? n_pairs : 1);
  *n_jchunk = (*n_jtile + *jpc - 1) / *jpc;
}

int tc_fwd_stats(const bsc_infonce_desc* d, const InfoNceFwdWs& L, char* ws, float* stats,
                 const bsc_infonce_exchange* ex, cudaStream_t st) {
  BSC_REQUIRE(d->d % BK == 0, BSC_EUNSUPPORTED, "TC16 mode needs d %% 64 == 0 (got %d); use BSC_MODE_F32", d->d);
  BSC_REQUIRE(d->n_global < (1ll << 31) - 512, BSC_EUNSUPPORTED, "n_global too large for 32-bit TMA coordinates");
  const bool x3 = d->mode == BSC_MODE_TC16X3;
  SimParams p{};
  int np = 0;
  for (int a = 0; a < 3; ++a) {
    if (d->xh[a]) {
      int rc = make_tmap_2d(&p.tmap[a], d->xh[a], d->x16_dtype, d->n_global, (x3 ? 2 : 1) * (int64_t)d->d, 128);
      if (rc) return rc;
    }
    p.inv[a] = d->inv[a];
    for (int b = a + 1; b < 3; ++b)
      if (d->xh[a] && d->xh[b]) {
        p.pair_a[np] = a;
        p.pair_b[np] = b;
        ++np;
      }
  }
  for (int a = 0; a < 3; ++a)
    if (!d->xh[a]) p.tmap[a] = p.tmap[p.pair_a[0]];
  p.n_pairs = np;
  p.n_iblk = L.n_iblk;
  p.n_jtile = L.n_jtile;
  p.n_jchunk = L.n_jchunk;
  p.jpc = L.jtiles_per_chunk;
  p.num_items = np * L.n_jchunk * L.n_iblk;
  p.kblocks = (x3 ? 3 : 1) * (d->d / BK);
  p.x3_kt = x3 ? d->d / BK : 0;
  p.x3_d = d->d;
  p.a_rows = d->n_local;
  p.a_row0 = d->row0;
  p.b_rows = d->n_global;
  const float log2e = 1.4426950408889634f;
  p.k1 = d->logit_scale * log2e;
  p.k2 = d->logit_scale * log2e;
  if (x3) p.k1 = (float)((double)p.k1 * x3_accumulator_gain(d->d));
  p.rowpart = (float*)(ws + L.rowpart);
  p.colpart = (float*)(ws + L.colpart);
  p.rowpos = (float*)(ws + L.rowpos);
  p.rep = d->rep;
  p.jrange = (const int2*)(ws + L.jrange);
  p.a_pad = L.n_local_pad;
  p.b_pad = L.n_global_pad;
  if (ex != nullptr && ex->arrive != nullptr) {
    BSC_REQUIRE(tc_fwd_source_aligned(d->n_local, d->n_global, d->world), BSC_EUNSUPPORTED,
                "the fused all-gather needs n_local %% 256 == 0 and equal row blocks");
    p.arrive = (const unsigned long long*)ex->arrive;
    p.arrive_target = ex->arrive_target;
    p.src_rows = (int)d->n_local;
    p.my_rank = d->rank;
    p.world = d->world;
    p.chunk_major = 1;
    p.cps = L.n_jchunk / d->world;
  }
  int rc = rep_range_launch(d->rep, d->n_global, BN, (int2*)(ws + L.jrange), st);
  if (rc) return rc;
  timing_begin(0, st);
  rc = d->x16_dtype == BSC_BF16 ? launch_sim<EPI_STATS, 1>(p, st) : launch_sim<EPI_STATS, 0>(p, st);
  timing_end(0, st);
  if (rc) return rc;
  // One launch reduces the partial sums into this rank's statistics record -- sum-exp rows of the local rows, partial
  // column sums, the positive term sum_ij T_ij (s*S_ij - s) in natural-log units -- and mirrors it into slot `rank` of
  // every peer's record area (the statistics exchange: no NCCL call)
  ReduceJobs jobs{};
  for (int q = 0; q < np; ++q) {
    const int a = p.pair_a[q], b = p.pair_b[q];
    jobs.job[jobs.count++] = ReduceJob{p.rowpart + (int64_t)q * L.n_jchunk * NG * L.n_local_pad, L.n_jchunk * NG,
                                       L.n_local_pad, d->n_local,
                                       stats + (int64_t)pair_id(a, b) * d->n_global + d->row0};
    jobs.job[jobs.count++] = ReduceJob{p.colpart + (int64_t)q * L.n_iblk * L.n_global_pad, L.n_iblk, L.n_global_pad,
                                       d->n_global, stats + (int64_t)pair_id(b, a) * d->n_global};
  }
  jobs.pos_part = p.rowpos;
  jobs.pos_nparts = np * L.n_jchunk * NG;
  jobs.pos_stride = L.n_local_pad;
  jobs.pos_n = d->n_local;
  jobs.pos_scale = 0.6931471805599453f;
  jobs.pos_out = reinterpret_cast<double*>(stats + stats_pos_offset(d->n_global));
  if (ex != nullptr && (ex->stats_peers != nullptr || ex->stats_multicast != nullptr)) {
    const int64_t slot = (int64_t)d->rank * stats_record_len(d->n_global);
    jobs.local_base = stats;
    if (ex->stats_multicast != nullptr) {
      jobs.mc_base = (float*)ex->stats_multicast + slot;
    } else {
      jobs.n_peers = d->world;
      for (int r = 0; r < d->world; ++r) jobs.peer_base[r] = (float*)ex->stats_peers[r] + slot;
    }
  }
  return reduce_parts_batch_launch(jobs, st);
}

// ---- stored-weight backward: the weight blocks ------------------------------------------------------------
// This file's sweep with the EPI_WEIGHTS epilogue recomputes S[local rows, all columns] per pair (the forward's
// 128 x 256 tiles on CTA pairs) and writes G = gs (e (u + v) - 2 T) / |b_j| as fp16 images; tc_bwd_gemms (tc_bwd3.cu)
// forms the gradient GEMMs from them.  BSC_BWD_SCATTER: one block per UNORDERED pair (a < b) serves both sides;
// BSC_BWD_BLOCKS: one block per ORDERED pair.
void tc_bwd_matrix_dims(const bsc_infonce_desc* d, int* n_iblk_pad, int* n_j64) {
  const int n_iblk = (int)((d->n_local + BM - 1) / BM);
  *n_iblk_pad = 2 * ((n_iblk + 1) / 2);
  *n_j64 = 4 * (int)((d->n_global + BN - 1) / BN);
}

size_t tc_bwd_gimg_bytes(const bsc_infonce_desc* d) {
  int m = 0;
  for (int a = 0; a < 3; ++a) m += d->xh[a] ? 1 : 0;
  const size_t blocks = (size_t)(d->bwd_mode == BSC_BWD_SCATTER ? m * (m - 1) / 2 : m * (m - 1));
  int n_iblk_pad, n_j64;
  tc_bwd_matrix_dims(d, &n_iblk_pad, &n_j64);
  return blocks * (size_t)n_iblk_pad * (size_t)n_j64 * 16384 * (size_t)mode_parts(d->mode);  // X3: hi set, then lo set
}

int tc_bwd_weights(const bsc_infonce_desc* d, const InfoNceBwdWs& L, char* ws, const float* aux, int* n_pairs,
                   int* pairs_a, int* pairs_b, cudaStream_t st) {
  BSC_REQUIRE(d->n_global < (1ll << 31) - 512, BSC_EUNSUPPORTED, "n_global too large for 32-bit TMA coordinates");
  const bool ordered = L.mode == BSC_BWD_BLOCKS;
  const bool x3 = d->mode == BSC_MODE_TC16X3;
  SimParams p{};
  int np = 0, first = -1;
  for (int a = 0; a < 3; ++a) {
    if (d->xh[a]) {
      int rc = make_tmap_2d(&p.tmap[a], d->xh[a], d->x16_dtype, d->n_global, (x3 ? 2 : 1) * (int64_t)d->d, 128);
      if (rc) return rc;
      if (first < 0) first = a;
    }
    p.inv[a] = d->inv[a];
    for (int b = ordered ? 0 : a + 1; b < 3; ++b)
      if (b != a && d->xh[a] && d->xh[b]) {
        p.pair_a[np] = pairs_a[np] = a;
        p.pair_b[np] = pairs_b[np] = b;
        ++np;
      }
  }
  *n_pairs = np;
  for (int a = 0; a < 3; ++a)
    if (!d->xh[a]) p.tmap[a] = p.tmap[first];
  p.n_pairs = np;
  // the forward's tiling was sized for its own pair count; the item count only has to cover np * n_jchunk * n_iblk
  p.n_iblk = L.n_iblk;
  p.n_jtile = L.n_jtile;
  p.n_jchunk = L.n_jchunk;
  p.jpc = L.jtiles_per_chunk;
  p.num_items = np * L.n_jchunk * L.n_iblk;
  p.kblocks = (x3 ? 3 : 1) * (d->d / BK);
  p.x3_kt = x3 ? d->d / BK : 0;
  p.x3_d = d->d;
  p.a_rows = d->n_local;
  p.a_row0 = d->row0;
  p.b_rows = d->n_global;
  const float log2e = 1.4426950408889634f;
  p.k1 = d->logit_scale * log2e;
  p.k2 = d->logit_scale * log2e;
  if (x3) p.k1 = (float)((double)p.k1 * x3_accumulator_gain(d->d));
  p.rep = d->rep;
  p.uv = aux;
  p.gsbuf = aux + 6 * d->n_global;
  p.gimg = (uint8_t*)(ws + L.gimg);
  tc_bwd_matrix_dims(d, &p.n_iblk_pad, &p.n_j64);
  p.gimg_lo = p.gimg + (size_t)np * p.n_iblk_pad * p.n_j64 * 16384;  // used by EPI_WEIGHTS_X3 only
  p.jrange = (const int2*)(ws + L.jrange);
  {
    int rc = rep_range_launch(d->rep, d->n_global, BN, (int2*)(ws + L.jrange), st);
    if (rc) return rc;
  }
#ifdef BSC_EXPERIMENTS
  {
    const char* e = getenv("BSC_DBG_W");
    p.dbg = e ? atoi(e) : 0;
  }
#endif
  timing_begin(1, st);
  int rc = x3 ? launch_sim<EPI_WEIGHTS_X3, 0>(p, st) : launch_sim<EPI_WEIGHTS, 0>(p, st);
  timing_end(1, st);
  return rc;
}

// ---- retrieval ---------------------------------------------------------------------------------------
static void bank_tiling(int64_t nq, int64_t nb, int* n_iblk, int* n_jtile, int* jpc, int* n_jchunk,
                        int min_parts = 0) {
  *n_iblk = (int)((nq + BM - 1) / BM);
  *n_jtile = (int)((nb + BN - 1) / BN);
  // enough items to fill the machine a few times, few enough that the candidate lists stay small
  const int target = num_sms() * 2;
  int chunks = (target + *n_iblk - 1) / *n_iblk;
  if (chunks < 1) chunks = 1;
  if (*n_iblk > target) {
    // many query blocks: whole passes over the bank per work item.  Cut the bank in a few chunks so that the last
    // round of items does not leave most SMs idle (782 blocks on 74 CTA pairs: 5.3 rounds -> 6 without chunks)
    const bool pairs = (*n_iblk % 2) == 0;
    const double units = pairs ? num_sms() / 2 : num_sms();
    const double per_chunk = pairs ? *n_iblk / 2 : *n_iblk;
    double best = 1e30;
    for (int c = 1; c <= 16; ++c) {
      const double rounds = per_chunk * c / units;
      const double waste = (double)(int64_t)(rounds + 0.999999) / rounds;
      if (waste < best * 0.99) {
        best = waste;
        chunks = c;
      }
    }
  }
#if BSC_SCAN_V2
  {
    // Few query blocks (a 32-query pass: one block): ONE work item per SM / CTA pair.  Every item starts with empty
    // top-k lists, and until they have filled up nearly every column takes the insert path (32 queries share a warp, so
    // the union of their insert columns stays dense for the first ~2000 rows): two items per SM pay that twice and
    // hand twice as many lists to the merge.
    const bool pairs = (*n_iblk % 2) == 0;
    const int units = pairs ? num_sms() / 2 : num_sms();
    const int per_chunk = pairs ? *n_iblk / 2 : *n_iblk;
    if (per_chunk * 8 <= units) chunks = units / per_chunk;
  }
#endif
  if (chunks * NG < min_parts) chunks = (min_parts + NG - 1) / NG;  // screening pass: at least min_parts lists
  if (chunks > *n_jtile) chunks = *n_jtile > 0 ? *n_jtile : 1;
  *jpc = (*n_jtile + chunks - 1) / chunks;
  if (*jpc < 1) *jpc = 1;
  *n_jchunk = (*n_jtile + *jpc - 1) / *jpc;
}

// At most 64 queries: 8 or 16 per TMEM lane quarter instead of 32 / 64 in the first one or two (SimParams::spread)
static int spread_queries(SimParams* p, const void* q, int64_t nq, int d) {
  p->spread = 0;
#if BSC_SCAN_V2 >= 2
  if (nq > 64) return BSC_OK;
  p->spread = nq <= 32 ? 8 : 16;
  return make_tmap_2d(&p->tmap[2], q, BSC_BF16, nq, d, p->spread);
#else
  (void)q;
  (void)nq;
  (void)d;
  return BSC_OK;
#endif
}

size_t tc_bank_ws(int64_t nq, int64_t nb, int d, int k) {
  (void)d;
  int n_iblk, n_jtile, jpc, n_jchunk;
  bank_tiling(nq, nb, &n_iblk, &n_jtile, &jpc, &n_jchunk);
  const size_t cand = (size_t)n_jchunk * NG * (size_t)nq * (size_t)k;
  return align_up(cand * sizeof(float), 256) + align_up(cand * sizeof(int64_t), 256);
}

int tc_bank_search(const void* q, int64_t nq, const void* bank, int64_t nb, int d, int k, int64_t base_id,
                   float* out_s, int64_t* out_i, char* ws, size_t ws_bytes, cudaStream_t st) {
  BSC_REQUIRE(d % BK == 0, BSC_EUNSUPPORTED, "TC16 retrieval needs d %% 64 == 0 (got %d); use BSC_MODE_F32", d);
  BSC_REQUIRE(k >= 1 && k <= KTOP, BSC_EUNSUPPORTED, "TC16 retrieval supports k <= %d (got %d); use BSC_MODE_F32",
              KTOP, k);
  BSC_REQUIRE(nb < (1ll << 31) - 512 && nq < (1ll << 31) - 512, BSC_EUNSUPPORTED,
              "shard too large for 32-bit TMA coordinates");
  BSC_REQUIRE(ws_bytes >= tc_bank_ws(nq, nb, d, k), BSC_EWORKSPACE, "bsc_bank_search: workspace too small");
  if (nq == 0) return BSC_OK;
  SimParams p{};
  int rc = make_tmap_2d(&p.tmap[0], q, BSC_BF16, nq, d, 128);
  if (rc) return rc;
  if (nb > 0) {
    rc = make_tmap_2d(&p.tmap[1], bank, BSC_BF16, nb, d, 128);
    if (rc) return rc;
  } else {
    p.tmap[1] = p.tmap[0];
  }
  p.tmap[2] = p.tmap[0];
  p.n_pairs = 1;
  p.pair_a[0] = 0;
  p.pair_b[0] = 1;
  bank_tiling(nq, nb, &p.n_iblk, &p.n_jtile, &p.jpc, &p.n_jchunk);
  p.num_items = nb > 0 ? p.n_jchunk * p.n_iblk : 0;
  rc = spread_queries(&p, q, nq, d);
  if (rc) return rc;
  p.kblocks = d / BK;
  p.a_rows = nq;
  p.a_row0 = 0;
  p.b_rows = nb;
  p.k = k;
  p.base_id = base_id;
  const size_t cand = (size_t)p.n_jchunk * NG * (size_t)nq * (size_t)k;
  p.cand_s = (float*)ws;
  p.cand_i = (int64_t*)(ws + align_up(cand * sizeof(float), 256));
  int parts = p.n_jchunk * NG;
  if (nb == 0) {
    // nothing to scan: the merge of zero parts writes the padding
    parts = 0;
  } else {
    timing_begin(2, st);
    rc = launch_sim<EPI_TOPK, 1>(p, st);
    timing_end(2, st);
    if (rc) return rc;
  }
  return topk_merge_launch(p.cand_s, p.cand_i, parts, nq, k, out_s, out_i, st);
}

// Screening pass of the exact search: the bf16 scan with every (column chunk, epilogue group) candidate list of KTOP
// entries handed out unmerged.  parts = number of lists; a bank row that is in no list scored (in bf16) at most the
// last entry of its own list.
int tc_bank_screen_parts(int64_t nq, int64_t nb, int min_parts) {
  int n_iblk, n_jtile, jpc, n_jchunk;
  bank_tiling(nq, nb, &n_iblk, &n_jtile, &jpc, &n_jchunk, min_parts);
  return n_jchunk * NG;
}

int tc_bank_screen(const void* q, int64_t nq, const void* bank, int64_t nb, int d, int64_t base_id, int min_parts,
                   float* cand_s, int64_t* cand_i, cudaStream_t st) {
  BSC_REQUIRE(d % BK == 0, BSC_EUNSUPPORTED, "screening needs d %% 64 == 0 (got %d)", d);
  BSC_REQUIRE(nb > 0 && nb < (1ll << 31) - 512 && nq < (1ll << 31) - 512, BSC_EUNSUPPORTED,
              "shard size unsupported by the screening pass");
  if (nq == 0) return BSC_OK;
  SimParams p{};
  int rc = make_tmap_2d(&p.tmap[0], q, BSC_BF16, nq, d, 128);
  if (rc) return rc;
  rc = make_tmap_2d(&p.tmap[1], bank, BSC_BF16, nb, d, 128);
  if (rc) return rc;
  p.tmap[2] = p.tmap[0];
  p.n_pairs = 1;
  p.pair_a[0] = 0;
  p.pair_b[0] = 1;
  bank_tiling(nq, nb, &p.n_iblk, &p.n_jtile, &p.jpc, &p.n_jchunk, min_parts);
  p.num_items = p.n_jchunk * p.n_iblk;
  rc = spread_queries(&p, q, nq, d);
  if (rc) return rc;
  p.kblocks = d / BK;
  p.a_rows = nq;
  p.a_row0 = 0;
  p.b_rows = nb;
  p.k = KTOP;
  p.base_id = base_id;
  p.cand_s = cand_s;
  p.cand_i = cand_i;
  timing_begin(2, st);
  rc = launch_sim<EPI_TOPK, 1>(p, st);
  timing_end(2, st);
  return rc;
}

}  // namespace bsc
