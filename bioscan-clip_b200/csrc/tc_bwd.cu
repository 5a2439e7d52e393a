// tcgen05 InfoNCE backward sweep (the autograd of loss_func.py:46-51 without the N x N matrices).
//
// For an ordered modality pair (x, y) and the local row block I (128 rows):
//     dXhat_x[I, :] (+)= sum_J  G[I, J] * Yhat[J, :],    G = gs * (exp(s*S - s) * (u_i + v_j) - 2 T_ij),  S = Xhat_I Yhat_J^T
// with u_i = r_i / sumexp_xy[i], v_j = r_j / sumexp_yx[j] (softmax rows and columns share one exp) and
// T_ij = [label_i == label_j].  The target is subtracted BEFORE the tile is rounded to fp16: for a peaked softmax
// r_i P_ij - T_ij cancels to ~1e-3 of its terms, which a 11-bit operand could not represent after the fact.
//
// One CTA owns (x, y, I, column chunk of D): TMEM holds the chunk accumulator (<= 384 fp32 columns) plus one
// 128 x 128 S tile.  Per column tile J: S is recomputed on the tensor cores (K = D), the epilogue warps turn
// it into the fp16 tile G in 128B-swizzled shared memory, and a second tcgen05.mma chain (A = G K-major,
// B = the Yhat tile read MN-major from the very bytes TMA delivered) accumulates into TMEM.
// MMA issue order is software pipelined -- S(t+1) is queued before grad(t) -- so the tensor pipe executes
// grad(t-1) while the epilogue is reading S(t) out of TMEM.
// Warp roles (256 threads): 0 = TMA producer, 1 = MMA issuer, 2 = TMEM allocator, 4..7 = epilogue.
#include <cstdlib>

#include "tc_common.cuh"

namespace bsc {
namespace {

constexpr int BM = kTcBM;   // 128 rows of I
constexpr int BN = kTcBNB;  // 128 columns of J per tile
constexpr int BK = 64;
constexpr int SLOT = 16384;           // one TMA box: 128 rows x 128 bytes
constexpr int STAGE_BYTES = 2 * SLOT; // S stage: {Xhat k-block, Yhat k-block}; grad stage: 128 d-columns of Yhat_J
constexpr int STAGES = 5;
constexpr int G_BYTES = BM * BN * 2;  // 32 KB fp16 tile, two 64-column swizzle spans
constexpr int TMEM_COLS = 512;
constexpr int ACC_COL0 = 128;         // accumulator columns start after the S tile
constexpr int MAX_GROUPS = 3;         // 3 x 128 accumulator columns per CTA

struct BwdParams {
  CUtensorMap tmap[3];
  int n_xy;
  int xy_x[6], xy_y[6], xy_slot[6];
  int n_iblk, n_jtile, kblocks;
  int n_chunks;          // column chunks of D per row block
  int groups_per_chunk;  // 128-column groups per chunk (last chunk may hold fewer)
  int n_groups;          // d / 128
  int num_items;
  int d;
  int64_t n_global, n_local, row0;
  float k1, k2;
  const float* uv;     // [6][N]
  const int32_t* rep;  // [N] label class representative: T_ij = (rep[i] == rep[j])
  const float* gsbuf;  // [6][2] {gs, 1/gs} per ordered pair
  const float* inv[3]; // inverse row norms per modality
  float* dxraw;        // [3][2][n_local][d]
  int dbg_skip_tma;    // timing experiment only (BSC_DBG_SKIP_TMA=1): stages after the first ring round are not loaded
};

struct __align__(8) BwdSmemTail {
  uint64_t full[STAGES];
  uint64_t empty[STAGES];
  uint64_t s_full, s_free;
  uint64_t g_ready[2], g_free[2];
  uint64_t acc_full, acc_free;
  uint32_t tmem_base;
  uint32_t pad;
  // per-column facts of the current tile (single buffer, two named barriers per tile)
  float sc[BN];   // k1 / ||y_j||
  float vi[BN];   // gs * v_j / ||y_j||
  int repj[BN];   // label class (-1 past the end of the batch)
};

// dynamic shared memory starts 1024-byte aligned (no static __shared__ in this kernel; checked at run time)
constexpr size_t BWD_SMEM_BYTES = (size_t)STAGES * STAGE_BYTES + 2 * (size_t)G_BYTES + sizeof(BwdSmemTail);
static_assert(BWD_SMEM_BYTES <= 232448, "backward kernel exceeds the 227 KB shared-memory limit");

__device__ __forceinline__ void decode_item(const BwdParams& p, int item, int& xy, int& iblk, int& chunk) {
  chunk = item % p.n_chunks;
  const int rest = item / p.n_chunks;
  iblk = rest % p.n_iblk;
  xy = rest / p.n_iblk;
}

__device__ __forceinline__ int chunk_groups(const BwdParams& p, int chunk) {
  const int g0 = chunk * p.groups_per_chunk;
  return min(p.groups_per_chunk, p.n_groups - g0);
}

template <int FMT>
__global__ void __launch_bounds__(256, 1) tc_bwd_kernel(const __grid_constant__ BwdParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  const uint32_t ring = ptx::smem_u32(smem_raw);
  if ((ring & 1023u) != 0) {
    if (threadIdx.x == 0) printf("bsc: dynamic shared memory is not 1024-byte aligned (0x%x)\n", ring);
    __trap();
  }
  uint8_t* ring_ptr = smem_raw;
  const uint32_t gbuf = ring + STAGES * STAGE_BYTES;
  uint8_t* gbuf_ptr = ring_ptr + (size_t)STAGES * STAGE_BYTES;
  BwdSmemTail* tail = reinterpret_cast<BwdSmemTail*>(gbuf_ptr + 2 * (size_t)G_BYTES);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    for (int s = 0; s < STAGES; ++s) {
      ptx::mbar_init(ptx::smem_u32(&tail->full[s]), 1);
      ptx::mbar_init(ptx::smem_u32(&tail->empty[s]), 1);
    }
    ptx::mbar_init(ptx::smem_u32(&tail->s_full), 1);
    ptx::mbar_init(ptx::smem_u32(&tail->s_free), 4);
    for (int b = 0; b < 2; ++b) {
      ptx::mbar_init(ptx::smem_u32(&tail->g_ready[b]), 4);
      ptx::mbar_init(ptx::smem_u32(&tail->g_free[b]), 1);
    }
    ptx::mbar_init(ptx::smem_u32(&tail->acc_full), 1);
    ptx::mbar_init(ptx::smem_u32(&tail->acc_free), 4);
    ptx::fence_barrier_init();
  }
  if (warp == 0 && lane == 0) {
    ptx::prefetch_tmap(&p.tmap[0]);
    ptx::prefetch_tmap(&p.tmap[1]);
    ptx::prefetch_tmap(&p.tmap[2]);
  }
  if (warp == 2) {
    ptx::tmem_alloc(ptx::smem_u32(&tail->tmem_base), TMEM_COLS);
    ptx::tmem_relinquish();
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = tail->tmem_base;
  const int T = p.n_jtile;

  if (warp == 0) {
    // ================= TMA producer: stages in exactly the order the MMA warp consumes them =================
    if (ptx::elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      auto next_stage = [&]() {
        if (++stage == STAGES) {
          stage = 0;
          phase ^= 1u;
        }
      };
      for (int item = blockIdx.x; item < p.num_items; item += gridDim.x) {
        int xy, iblk, chunk;
        decode_item(p, item, xy, iblk, chunk);
        const CUtensorMap* mx = &p.tmap[p.xy_x[xy]];
        const CUtensorMap* my = &p.tmap[p.xy_y[xy]];
        const int xrow = (int)(p.row0 + (int64_t)iblk * BM);
        const int ng = chunk_groups(p, chunk);
        const int dcol0 = chunk * p.groups_per_chunk * 128;
        auto load_s = [&](int t) {
          for (int kb = 0; kb < p.kblocks; ++kb) {
            ptx::mbar_wait(ptx::smem_u32(&tail->empty[stage]), phase ^ 1u, 100 + stage);
            const uint32_t full = ptx::smem_u32(&tail->full[stage]);
            const uint32_t sa = ring + stage * STAGE_BYTES;
            ptx::mbar_arrive_expect_tx(full, STAGE_BYTES);
            ptx::tma_load_2d(sa, mx, full, kb * BK, xrow);
            ptx::tma_load_2d(sa + SLOT, my, full, kb * BK, t * BN);
            next_stage();
          }
        };
        auto load_g = [&](int t) {
          for (int g = 0; g < ng; ++g) {
            ptx::mbar_wait(ptx::smem_u32(&tail->empty[stage]), phase ^ 1u, 110 + stage);
            const uint32_t full = ptx::smem_u32(&tail->full[stage]);
            const uint32_t sa = ring + stage * STAGE_BYTES;
            ptx::mbar_arrive_expect_tx(full, STAGE_BYTES);
            ptx::tma_load_2d(sa, my, full, dcol0 + g * 128, t * BN);
            ptx::tma_load_2d(sa + SLOT, my, full, dcol0 + g * 128 + 64, t * BN);
            next_stage();
          }
        };
        load_s(0);
        for (int t = 0; t < T; ++t) {
          if (t + 1 < T) load_s(t + 1);
          load_g(t);
        }
      }
    }
  } else if (warp == 1) {
    // ================= MMA issuer (one elected thread runs the whole role) =================
    if (ptx::elect_one()) {
      constexpr uint32_t idesc_s = umma_idesc(BM, BN, FMT, FMT, 0, 0);  // S = X_I * Y_J^T, both K-major
      constexpr uint32_t idesc_g = umma_idesc(BM, 128, 0, FMT, 0, 1);  // acc += G (fp16, K-major) * Y_J (MN-major)
      // descriptors differ only in the start-address field: desc(addr) = base + (addr >> 4)
      const uint64_t desc_k = ptx::umma_desc_sw128(0, 16, 1024);     // K-major operands (X, Y for S; G for grad)
      const uint64_t desc_mn = ptx::umma_desc_sw128(0, SLOT, 1024);  // Y tile read MN-major for the grad MMA
      int stage = 0;
      uint32_t phase = 0;
      uint32_t s_count = 0;  // S tiles issued so far (all items)
      uint32_t g_count = 0;  // grad tiles issued so far
      uint32_t item_count = 0;
      auto next_stage = [&]() {
        if (++stage == STAGES) {
          stage = 0;
          phase ^= 1u;
        }
      };
      for (int item = blockIdx.x; item < p.num_items; item += gridDim.x, ++item_count) {
        int xy, iblk, chunk;
        decode_item(p, item, xy, iblk, chunk);
        const int ng = chunk_groups(p, chunk);
        auto issue_s = [&]() {
          // the single S buffer is free once the epilogue has pulled the previous tile into registers
          ptx::mbar_wait(ptx::smem_u32(&tail->s_free), (s_count & 1u) ^ 1u, 200);
          ptx::tc_fence_after();
          for (int kb = 0; kb < p.kblocks; ++kb) {
            ptx::mbar_wait(ptx::smem_u32(&tail->full[stage]), phase, 300 + stage);
            ptx::tc_fence_after();
            const uint32_t sa = ring + stage * STAGE_BYTES;
            const uint64_t da = desc_k + ptx::desc_addr(sa);
            const uint64_t db = desc_k + ptx::desc_addr(sa + SLOT);
#pragma unroll
            for (int k = 0; k < BK / 16; ++k)
              ptx::umma_f16(tmem_base, da + 2 * k, db + 2 * k, idesc_s, (kb | k) != 0 ? 1u : 0u);
            ptx::umma_commit(ptx::smem_u32(&tail->empty[stage]));
            if (kb == p.kblocks - 1) ptx::umma_commit(ptx::smem_u32(&tail->s_full));
            next_stage();
          }
          ++s_count;
        };
        auto issue_g = [&](int t) {
          const uint32_t buf = g_count & 1u;
          const uint32_t use = g_count >> 1;
          if (t == 0) {
            // first accumulation of the item overwrites TMEM: the previous item must have been drained
            ptx::mbar_wait(ptx::smem_u32(&tail->acc_free), (item_count & 1u) ^ 1u, 210);
          }
          ptx::mbar_wait(ptx::smem_u32(&tail->g_ready[buf]), use & 1u, 220 + buf);
          ptx::tc_fence_after();
          for (int g = 0; g < ng; ++g) {
            ptx::mbar_wait(ptx::smem_u32(&tail->full[stage]), phase, 310 + stage);
            ptx::tc_fence_after();
            const uint32_t sb = ring + stage * STAGE_BYTES;
            // A: G[128 x 128] K-major, two 64-column spans of 16 KB, 32 bytes per K=16 step inside a span
            const uint64_t ga = desc_k + ptx::desc_addr(gbuf + buf * G_BYTES);
            // B: Y[j][d] read MN-major: 8-row (j) groups 1024 B apart, 64-column (d) groups one box apart
            const uint64_t yb = desc_mn + ptx::desc_addr(sb);
            const uint32_t d_tmem = tmem_base + ACC_COL0 + g * 128;
#pragma unroll
            for (int kk = 0; kk < BN / 16; ++kk)
              ptx::umma_f16(d_tmem, ga + (uint64_t)(((kk >> 2) * SLOT + (kk & 3) * 32) >> 4), yb + (uint64_t)(kk * 128),
                            idesc_g, (t | kk) != 0 ? 1u : 0u);
            ptx::umma_commit(ptx::smem_u32(&tail->empty[stage]));
            if (g == ng - 1) {
              ptx::umma_commit(ptx::smem_u32(&tail->g_free[buf]));
              if (t == T - 1) ptx::umma_commit(ptx::smem_u32(&tail->acc_full));
            }
            next_stage();
          }
          ++g_count;
        };
        issue_s();
        for (int t = 0; t < T; ++t) {
          if (t + 1 < T) issue_s();
          issue_g(t);
        }
      }
    }
  } else if (warp >= 4) {
    // ================= epilogue: S (TMEM) -> G (fp16, swizzled smem); final accumulator drain =================
    const int q = warp & 3;
    const int row_in_tile = q * 32 + lane;
    const int et = threadIdx.x - 128;
    const float inv_k1 = 1.0f / p.k1;
    uint32_t tile_count = 0;  // == s_count == g_count on this side
    uint32_t item_count = 0;
    for (int item = blockIdx.x; item < p.num_items; item += gridDim.x, ++item_count) {
      int xy, iblk, chunk;
      decode_item(p, item, xy, iblk, chunk);
      const int x = p.xy_x[xy], y = p.xy_y[xy];
      const int64_t arow = (int64_t)iblk * BM + row_in_tile;
      const bool row_ok = arow < p.n_local;
      const float gs = p.gsbuf[2 * pair_id(x, y)];
      // row facts: inverse norm, gs*u_i/k1 and 2*gs/k1 (both later multiplied by sc_j = k1/||y_j||)
      const float inv_i = row_ok ? p.inv[x][p.row0 + arow] : 0.f;
      const float ugk = row_ok ? gs * inv_k1 * p.uv[(int64_t)pair_id(x, y) * p.n_global + p.row0 + arow] : 0.f;
      const int rep_i = row_ok ? p.rep[p.row0 + arow] : -2;
      const float m2gk = -2.0f * gs * inv_k1;
      const float* vrow = p.uv + (int64_t)pair_id(y, x) * p.n_global;
      const float* invy = p.inv[y];
      const int ng = chunk_groups(p, chunk);

      for (int t = 0; t < T; ++t, ++tile_count) {
        const uint32_t buf = tile_count & 1u;
        const uint32_t use = tile_count >> 1;
        // stage the column facts of this tile (one column per thread) for the 4 epilogue warps
        ptx::named_bar_sync(2, 128);  // everyone is done with the previous tile's columns
        {
          const int64_t j = (int64_t)t * BN + et;
          const bool ok = j < p.n_global;
          const float ij = ok ? invy[j] : 0.f;
          tail->sc[et] = p.k1 * ij;
          tail->vi[et] = ok ? gs * vrow[j] * ij : 0.f;
          tail->repj[et] = ok ? p.rep[j] : -1;
        }
        ptx::named_bar_sync(1, 128);
        ptx::mbar_wait(ptx::smem_u32(&tail->s_full), tile_count & 1u, 400);
        ptx::tc_fence_after();
        // G buffer `buf` was last read by the grad MMAs of tile_count - 2
        ptx::mbar_wait(ptx::smem_u32(&tail->g_free[buf]), (use & 1u) ^ 1u, 410 + buf);

        const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16);
        const uint32_t grow = gbuf + buf * G_BYTES + (uint32_t)(row_in_tile >> 3) * 1024u +
                              (uint32_t)(row_in_tile & 7) * 128u;
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          uint32_t r0[32], r1[32];
          ptx::tmem_ld_32x32(taddr + h * 64, r0);
          ptx::tmem_ld_32x32(taddr + h * 64 + 32, r1);
          ptx::tmem_ld_wait();
          if (h == 1) {
            // S is now entirely in registers: hand the TMEM tile back to the MMA warp
            ptx::tc_fence_before();
            __syncwarp();
            if (lane == 0) ptx::mbar_arrive(ptx::smem_u32(&tail->s_free));
          }
          const float* sc = &tail->sc[h * 64];
          const float* vi = &tail->vi[h * 64];
          const int* rj = &tail->repj[h * 64];
#pragma unroll
          for (int u = 0; u < 8; ++u) {  // 16-byte units of 8 columns
            uint32_t pk[4];
#pragma unroll
            for (int w = 0; w < 4; ++w) {
              const int c = u * 8 + w * 2;
              float gq[2];
#pragma unroll
              for (int z = 0; z < 2; ++z) {
                const int cc = c + z;
                const float dot = __uint_as_float(cc < 32 ? r0[cc & 31] : r1[cc & 31]);
                const float scj = sc[cc];
                // e = exp(s*cos - s), cos = dot * inv_i * inv_j;  weight = gs*(u_i + v_j)/||y_j||
                const float e = ptx::ex2(fmaf(dot, inv_i * scj, -p.k2));
                const float wgt = fmaf(ugk, scj, vi[cc]);
                const float tv = (rj[cc] == rep_i) ? m2gk * scj : 0.f;
                gq[z] = fmaf(e, wgt, tv);
              }
              const __half2 hh = __floats2half2_rn(gq[0], gq[1]);
              pk[w] = *reinterpret_cast<const uint32_t*>(&hh);
            }
            const uint32_t addr = grow + h * SLOT + ((uint32_t)(u ^ (row_in_tile & 7)) << 4);
            asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(pk[0]), "r"(pk[1]), "r"(pk[2]),
                         "r"(pk[3])
                         : "memory");
          }
        }
        // generic-proxy writes -> visible to the tensor core's async-proxy reads, then signal
        ptx::fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) ptx::mbar_arrive(ptx::smem_u32(&tail->g_ready[buf]));
      }

      // ---- drain the accumulator chunk of this item ----
      ptx::mbar_wait(ptx::smem_u32(&tail->acc_full), item_count & 1u, 420);
      ptx::tc_fence_after();
      {
        const int slot = p.xy_slot[xy];
        float* out = p.dxraw + (((int64_t)(x * 2 + slot) * p.n_local + arow) * p.d) +
                     (int64_t)chunk * p.groups_per_chunk * 128;
        const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + ACC_COL0;
        for (int c = 0; c < ng * 4; ++c) {
          uint32_t raw[32];
          ptx::tmem_ld_32x32(taddr + c * 32, raw);
          ptx::tmem_ld_wait();
          if (row_ok) {
#pragma unroll
            for (int w = 0; w < 8; ++w) {
              float4 v = make_float4(__uint_as_float(raw[w * 4]), __uint_as_float(raw[w * 4 + 1]),
                                     __uint_as_float(raw[w * 4 + 2]), __uint_as_float(raw[w * 4 + 3]));
              *reinterpret_cast<float4*>(out + c * 32 + w * 4) = v;
            }
          }
        }
      }
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(ptx::smem_u32(&tail->acc_free));
    }
  }

  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc(tmem_base, TMEM_COLS);
  }
}

// ---------------------------------------------------------------------------------------------------------------
// Paired variant (d split in exactly two column chunks): the two CTAs of a 2-CTA cluster own the two chunks of the
// SAME (ordered pair, row block).  S is computed once per column tile -- by CTA (t & 1) -- and the fp16 weight tile
// is written into local shared memory and bulk-copied (cp.async.bulk shared::cta -> shared::cluster) into the
// peer's, so the tensor cores execute
// 2 + 2 instead of 2*2 + 2 GEMM-units per ordered pair.  G buffer b is always written by CTA b.
//   g_ready[b]  CTA b: 8 arrivals per use (its epilogue warps); peer: 8 remote arrive.expect_tx + 8 x 4 KB of bulk-copy bytes
//   g_free[b]   (CTA b only) 2 arrivals per use: the grad MMAs of both CTAs (tcgen05.commit multicast to CTA b)
// ---------------------------------------------------------------------------------------------------------------
template <int FMT>
__global__ void __launch_bounds__(384, 1) tc_bwd_pair_kernel(const __grid_constant__ BwdParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  const uint32_t ring = ptx::smem_u32(smem_raw);
  if ((ring & 1023u) != 0) {
    if (threadIdx.x == 0) printf("bsc: dynamic shared memory is not 1024-byte aligned (0x%x)\n", ring);
    __trap();
  }
  const uint32_t gbuf = ring + STAGES * STAGE_BYTES;
  BwdSmemTail* tail = reinterpret_cast<BwdSmemTail*>(smem_raw + (size_t)STAGES * STAGE_BYTES + 2 * (size_t)G_BYTES);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t crank = ptx::cluster_ctarank();  // 0 / 1 = column chunk of D owned by this CTA
  const uint32_t peer = crank ^ 1u;
  const int cluster = (int)ptx::cluster_id_x();
  const int n_clusters = (int)ptx::n_clusters_x();

  if (threadIdx.x == 0) {
    for (int s = 0; s < STAGES; ++s) {
      ptx::mbar_init(ptx::smem_u32(&tail->full[s]), 1);
      ptx::mbar_init(ptx::smem_u32(&tail->empty[s]), 1);
    }
    ptx::mbar_init(ptx::smem_u32(&tail->s_full), 1);
    ptx::mbar_init(ptx::smem_u32(&tail->s_free), 8);
    for (int b = 0; b < 2; ++b) {
      // own buffer: the 8 local epilogue warps; peer's buffer: 8 remote arrive.expect_tx + the bulk copies' bytes
      ptx::mbar_init(ptx::smem_u32(&tail->g_ready[b]), 8);
      ptx::mbar_init(ptx::smem_u32(&tail->g_free[b]), 2);
    }
    ptx::mbar_init(ptx::smem_u32(&tail->acc_full), 1);
    ptx::mbar_init(ptx::smem_u32(&tail->acc_free), 8);
    ptx::fence_barrier_init();
  }
  if (warp == 0 && lane == 0) {
    ptx::prefetch_tmap(&p.tmap[0]);
    ptx::prefetch_tmap(&p.tmap[1]);
    ptx::prefetch_tmap(&p.tmap[2]);
  }
  if (warp == 2) {
    ptx::tmem_alloc(ptx::smem_u32(&tail->tmem_base), TMEM_COLS);
    ptx::tmem_relinquish();
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::cluster_sync_all();  // the peer's barriers are initialised before anything is signalled across CTAs
  ptx::tc_fence_after();
  const uint32_t tmem_base = tail->tmem_base;
  const int T = p.n_jtile;
  const int chunk = (int)crank;
  const int ng = chunk_groups(p, chunk);
  const int dcol0 = chunk * p.groups_per_chunk * 128;

  if (warp == 0) {
    // ================= TMA producer: stages in exactly the order the MMA thread consumes them =================
    if (ptx::elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      uint32_t nloads = 0;
      auto next_stage = [&]() {
        if (++stage == STAGES) {
          stage = 0;
          phase ^= 1u;
        }
      };
      for (int item = cluster; item < p.num_items; item += n_clusters) {
        const int iblk = item % p.n_iblk;
        const int xy = item / p.n_iblk;
        const CUtensorMap* mx = &p.tmap[p.xy_x[xy]];
        const CUtensorMap* my = &p.tmap[p.xy_y[xy]];
        const int xrow = (int)(p.row0 + (int64_t)iblk * BM);
        auto load_s = [&](int t) {
          for (int kb = 0; kb < p.kblocks; ++kb) {
            ptx::mbar_wait(ptx::smem_u32(&tail->empty[stage]), phase ^ 1u, 100 + stage);
            const uint32_t full = ptx::smem_u32(&tail->full[stage]);
            const uint32_t sa = ring + stage * STAGE_BYTES;
            if (p.dbg_skip_tma && nloads >= STAGES) {
              ptx::mbar_arrive(full);
            } else {
              ptx::mbar_arrive_expect_tx(full, STAGE_BYTES);
              ptx::tma_load_2d(sa, mx, full, kb * BK, xrow);
              ptx::tma_load_2d(sa + SLOT, my, full, kb * BK, t * BN);
            }
            ++nloads;
            next_stage();
          }
        };
        auto load_g = [&](int t) {
          for (int g = 0; g < ng; ++g) {
            ptx::mbar_wait(ptx::smem_u32(&tail->empty[stage]), phase ^ 1u, 110 + stage);
            const uint32_t full = ptx::smem_u32(&tail->full[stage]);
            const uint32_t sa = ring + stage * STAGE_BYTES;
            if (p.dbg_skip_tma && nloads >= STAGES) {
              ptx::mbar_arrive(full);
            } else {
              ptx::mbar_arrive_expect_tx(full, STAGE_BYTES);
              ptx::tma_load_2d(sa, my, full, dcol0 + g * 128, t * BN);
              ptx::tma_load_2d(sa + SLOT, my, full, dcol0 + g * 128 + 64, t * BN);
            }
            ++nloads;
            next_stage();
          }
        };
        if ((int)crank < T) load_s((int)crank);
        for (int t = 0; t < T; ++t) {
          if ((uint32_t)(t & 1) == crank && t + 2 < T) load_s(t + 2);
          load_g(t);
        }
      }
    }
  } else if (warp == 1) {
    // ================= MMA issuer (one elected thread) =================
    if (ptx::elect_one()) {
      constexpr uint32_t idesc_s = umma_idesc(BM, BN, FMT, FMT, 0, 0);
      constexpr uint32_t idesc_g = umma_idesc(BM, 128, 0, FMT, 0, 1);
      const uint64_t desc_k = ptx::umma_desc_sw128(0, 16, 1024);
      const uint64_t desc_mn = ptx::umma_desc_sw128(0, SLOT, 1024);
      int stage = 0;
      uint32_t phase = 0;
      uint32_t s_count = 0;          // S tiles issued by this CTA
      uint32_t gu[2] = {0u, 0u};     // uses of weight-tile buffer 0 / 1 so far
      uint32_t item_count = 0;
      auto next_stage = [&]() {
        if (++stage == STAGES) {
          stage = 0;
          phase ^= 1u;
        }
      };
      for (int item = cluster; item < p.num_items; item += n_clusters, ++item_count) {
        auto issue_s = [&]() {
          ptx::mbar_wait(ptx::smem_u32(&tail->s_free), (s_count & 1u) ^ 1u, 200);
          ptx::tc_fence_after();
          for (int kb = 0; kb < p.kblocks; ++kb) {
            ptx::mbar_wait(ptx::smem_u32(&tail->full[stage]), phase, 300 + stage);
            ptx::tc_fence_after();
            const uint32_t sa = ring + stage * STAGE_BYTES;
            const uint64_t da = desc_k + ptx::desc_addr(sa);
            const uint64_t db = desc_k + ptx::desc_addr(sa + SLOT);
#pragma unroll
            for (int k = 0; k < BK / 16; ++k)
              ptx::umma_f16(tmem_base, da + 2 * k, db + 2 * k, idesc_s, (kb | k) != 0 ? 1u : 0u);
            ptx::umma_commit(ptx::smem_u32(&tail->empty[stage]));
            if (kb == p.kblocks - 1) ptx::umma_commit(ptx::smem_u32(&tail->s_full));
            next_stage();
          }
          ++s_count;
        };
        auto issue_g = [&](int t) {
          const uint32_t buf = (uint32_t)(t & 1);
          if (t == 0) ptx::mbar_wait(ptx::smem_u32(&tail->acc_free), (item_count & 1u) ^ 1u, 210);
          // the weight tile of column tile t was written (here and in the peer) by CTA `buf`
          ptx::mbar_wait_cluster(ptx::smem_u32(&tail->g_ready[buf]), gu[buf] & 1u, 220 + buf);
          ptx::fence_proxy_async_all();  // the tile may have been written by the peer CTA through DSMEM
          ptx::tc_fence_after();
          for (int g = 0; g < ng; ++g) {
            ptx::mbar_wait(ptx::smem_u32(&tail->full[stage]), phase, 310 + stage);
            ptx::tc_fence_after();
            const uint32_t sb = ring + stage * STAGE_BYTES;
            const uint64_t ga = desc_k + ptx::desc_addr(gbuf + buf * G_BYTES);
            const uint64_t yb = desc_mn + ptx::desc_addr(sb);
            const uint32_t d_tmem = tmem_base + ACC_COL0 + g * 128;
#pragma unroll
            for (int kk = 0; kk < BN / 16; ++kk)
              ptx::umma_f16(d_tmem, ga + (uint64_t)(((kk >> 2) * SLOT + (kk & 3) * 32) >> 4), yb + (uint64_t)(kk * 128),
                            idesc_g, (t | kk) != 0 ? 1u : 0u);
            ptx::umma_commit(ptx::smem_u32(&tail->empty[stage]));
            if (g == ng - 1) {
              // tell the writer of this buffer (CTA `buf`) that this CTA's reads of it are done
              ptx::umma_commit_mcast(ptx::smem_u32(&tail->g_free[buf]), (uint16_t)(1u << buf));
              if (t == T - 1) ptx::umma_commit(ptx::smem_u32(&tail->acc_full));
            }
            next_stage();
          }
          ++gu[buf];
        };
        if ((int)crank < T) issue_s();
        for (int t = 0; t < T; ++t) {
          if ((uint32_t)(t & 1) == crank && t + 2 < T) issue_s();
          issue_g(t);
        }
      }
    }
  } else if (warp >= 4) {
    // ================= epilogue (8 warps): S (TMEM) -> weight tile in BOTH CTAs' shared memory; accumulator drain =====
    // warp e = warp - 4: TMEM lane quadrant q = e & 3 (rows 32q..32q+31), column span h = e >> 2 (64 columns).
    // Per duty tile a thread turns 64 dot products into 64 fp16 weights.  The latency chain between the tensor
    // pipes of the two CTAs is kept short: S is pulled out of TMEM and the weights are computed BEFORE waiting for
    // the previous use of the weight buffer to retire, and every warp ships its own 4 KB piece to the peer.
    const int e8 = warp - 4;
    const int q = e8 & 3;
    const int h = e8 >> 2;
    const int row_in_tile = q * 32 + lane;
    const int et = threadIdx.x - 128;  // 0..255; threads < 128 also stage the column facts
    const float inv_k1 = 1.0f / p.k1;
    const uint32_t my_gbuf = gbuf + crank * G_BYTES;          // the buffer this CTA writes (here and in the peer)
    const uint32_t peer_gbuf = ptx::mapa(my_gbuf, peer);
    const uint32_t my_ready = ptx::smem_u32(&tail->g_ready[crank]);
    const uint32_t peer_ready = ptx::mapa(my_ready, peer);
    const uint32_t piece = (uint32_t)h * SLOT + (uint32_t)q * 4096u;  // this warp's 32 rows x 128 B of span h
    uint32_t dcount = 0;  // duty tiles processed so far by this CTA (all items)
    uint32_t item_count = 0;
    for (int item = cluster; item < p.num_items; item += n_clusters, ++item_count) {
      const int iblk = item % p.n_iblk;
      const int xy = item / p.n_iblk;
      const int x = p.xy_x[xy], y = p.xy_y[xy];
      const int64_t arow = (int64_t)iblk * BM + row_in_tile;
      const bool row_ok = arow < p.n_local;
      const float gs = p.gsbuf[2 * pair_id(x, y)];
      const float inv_i = row_ok ? p.inv[x][p.row0 + arow] : 0.f;
      const float ugk = row_ok ? gs * inv_k1 * p.uv[(int64_t)pair_id(x, y) * p.n_global + p.row0 + arow] : 0.f;
      const int rep_i = row_ok ? p.rep[p.row0 + arow] : -2;
      const float m2gk = -2.0f * gs * inv_k1;
      const float* vrow = p.uv + (int64_t)pair_id(y, x) * p.n_global;
      const float* invy = p.inv[y];

      // column facts of a duty tile, fetched one tile ahead into registers (threads et < 128, one column each)
      float f_sc = 0.f, f_vi = 0.f;
      int f_rep = -1;
      auto fetch_facts = [&](int t) {
        if (et < BN && t < T) {
          const int64_t j = (int64_t)t * BN + et;
          const bool ok = j < p.n_global;
          const float ij = ok ? invy[j] : 0.f;
          f_sc = p.k1 * ij;
          f_vi = ok ? gs * vrow[j] * ij : 0.f;
          f_rep = ok ? p.rep[j] : -1;
        }
      };
      auto publish_facts = [&]() {
        if (et < BN) {
          tail->sc[et] = f_sc;
          tail->vi[et] = f_vi;
          tail->repj[et] = f_rep;
        }
      };
      fetch_facts((int)crank);
      ptx::named_bar_sync(2, 256);  // the previous item's last tile is done with the facts
      publish_facts();
      ptx::named_bar_sync(1, 256);

      for (int t = (int)crank; t < T; t += 2, ++dcount) {
        fetch_facts(t + 2);  // global loads in flight while this tile is processed
        ptx::mbar_wait(ptx::smem_u32(&tail->s_full), dcount & 1u, 400);
        ptx::tc_fence_after();
        const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(h * 64);
        uint32_t r0[32], r1[32];
        ptx::tmem_ld_32x32(taddr, r0);
        ptx::tmem_ld_32x32(taddr + 32, r1);
        ptx::tmem_ld_wait();
        // this warp's part of S is in registers: hand the TMEM tile back to the MMA thread
        ptx::tc_fence_before();
        __syncwarp();
        if (lane == 0) ptx::mbar_arrive(ptx::smem_u32(&tail->s_free));

        const float* sc = &tail->sc[h * 64];
        const float* vi = &tail->vi[h * 64];
        const int* rj = &tail->repj[h * 64];
        uint32_t pk[32];
#pragma unroll
        for (int c = 0; c < 64; c += 2) {
          float gq[2];
#pragma unroll
          for (int z = 0; z < 2; ++z) {
            const int cc = c + z;
            const float dot = __uint_as_float(cc < 32 ? r0[cc & 31] : r1[cc & 31]);
            const float scj = sc[cc];
            const float ex = ptx::ex2(fmaf(dot, inv_i * scj, -p.k2));
            const float wgt = fmaf(ugk, scj, vi[cc]);
            const float tv = (rj[cc] == rep_i) ? m2gk * scj : 0.f;
            gq[z] = fmaf(ex, wgt, tv);
          }
          const __half2 hh = __floats2half2_rn(gq[0], gq[1]);
          pk[c >> 1] = *reinterpret_cast<const uint32_t*>(&hh);
        }
        // buffer `crank` was last read by the grad MMAs (of both CTAs) of this CTA's previous duty tile; the barrier
        // is signalled by tcgen05.commit, no data is acquired through it (write-after-read on shared memory only)
        ptx::mbar_wait(ptx::smem_u32(&tail->g_free[crank]), (dcount & 1u) ^ 1u, 410);
        const uint32_t goff = (uint32_t)h * SLOT + (uint32_t)(row_in_tile >> 3) * 1024u + (uint32_t)(row_in_tile & 7) * 128u;
#pragma unroll
        for (int u = 0; u < 8; ++u) {
          const uint32_t off = goff + ((uint32_t)(u ^ (row_in_tile & 7)) << 4);
          asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(my_gbuf + off), "r"(pk[4 * u]), "r"(pk[4 * u + 1]),
                       "r"(pk[4 * u + 2]), "r"(pk[4 * u + 3])
                       : "memory");
        }
        // generic-proxy writes -> async proxy; then the local tensor core may read this warp's rows, and lane 0 ships
        // them to the peer CTA with a shared::cta -> shared::cluster bulk copy that completes (in bytes) on the PEER's
        // barrier (8 warps x {arrive.expect_tx 4096 + 4096 bytes} per use)
        ptx::fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) {
          ptx::mbar_arrive(my_ready);
          ptx::mbar_arrive_expect_tx_remote(peer_ready, 4096u);
          ptx::bulk_copy_to_peer(peer_gbuf + piece, my_gbuf + piece, 4096u, peer_ready);
        }
        // column facts of the next duty tile (off the critical path: the weight tile is already on its way)
        if (t + 2 < T) {
          ptx::named_bar_sync(2, 256);
          publish_facts();
          ptx::named_bar_sync(1, 256);
        }
      }

      // ---- drain this CTA's accumulator chunk: quadrant q, half h of the 32-column groups ----
      ptx::mbar_wait(ptx::smem_u32(&tail->acc_full), item_count & 1u, 420);
      ptx::tc_fence_after();
      {
        const int slot = p.xy_slot[xy];
        float* out = p.dxraw + (((int64_t)(x * 2 + slot) * p.n_local + arow) * p.d) + (int64_t)dcol0;
        const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + ACC_COL0;
        const int nc = ng * 4;
        const int c_lo = h == 0 ? 0 : nc / 2;
        const int c_hi = h == 0 ? nc / 2 : nc;
        for (int c = c_lo; c < c_hi; ++c) {
          uint32_t raw[32];
          ptx::tmem_ld_32x32(taddr + c * 32, raw);
          ptx::tmem_ld_wait();
          if (row_ok) {
#pragma unroll
            for (int w = 0; w < 8; ++w) {
              float4 v = make_float4(__uint_as_float(raw[w * 4]), __uint_as_float(raw[w * 4 + 1]),
                                     __uint_as_float(raw[w * 4 + 2]), __uint_as_float(raw[w * 4 + 3]));
              *reinterpret_cast<float4*>(out + c * 32 + w * 4) = v;
            }
          }
        }
      }
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(ptx::smem_u32(&tail->acc_free));
    }
  }

  ptx::tc_fence_before();
  __syncthreads();
  ptx::cluster_sync_all();  // neither CTA leaves while the other may still write into its shared memory
  if (warp == 2) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc(tmem_base, TMEM_COLS);
  }
}

int sm_count() {
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

int tc_bwd_raw(const bsc_infonce_desc* d, const InfoNceWs& L, char* ws, cudaStream_t st) {
  BSC_REQUIRE(d->d % 128 == 0, BSC_EUNSUPPORTED, "TC16 backward needs d %% 128 == 0 (got %d); use BSC_MODE_F32",
              d->d);
  BSC_REQUIRE(d->n_global < (1ll << 31) - 512, BSC_EUNSUPPORTED, "n_global too large for 32-bit TMA coordinates");
  {
    const int rc2 = tc_bwd_duo(d, L, ws, st);  // cta_group::2 sweep for d in {256, 512, 768}
    if (rc2 <= 0) return rc2;
  }
  BwdParams p{};
  int first = -1;
  for (int a = 0; a < 3; ++a)
    if (d->xh[a]) {
      int rc = make_tmap_2d(&p.tmap[a], d->xh[a], d->x16_dtype, d->n_global, d->d, 128);
      if (rc) return rc;
      p.inv[a] = d->inv[a];
      if (first < 0) first = a;
    }
  for (int a = 0; a < 3; ++a)
    if (!d->xh[a]) p.tmap[a] = p.tmap[first];
  int n = 0;
  for (int x = 0; x < 3; ++x) {
    if (!d->xh[x]) continue;
    int slot = 0;
    for (int y = 0; y < 3; ++y) {
      if (y == x || !d->xh[y]) continue;
      p.xy_x[n] = x;
      p.xy_y[n] = y;
      p.xy_slot[n] = slot++;
      ++n;
    }
  }
  p.n_xy = n;
  p.n_iblk = (int)((d->n_local + BM - 1) / BM);
  p.n_jtile = (int)((d->n_global + BN - 1) / BN);
  p.kblocks = d->d / BK;
  p.n_groups = d->d / 128;
  p.n_chunks = (p.n_groups + MAX_GROUPS - 1) / MAX_GROUPS;
  p.groups_per_chunk = (p.n_groups + p.n_chunks - 1) / p.n_chunks;
  p.num_items = n * p.n_iblk * p.n_chunks;
  p.d = d->d;
  p.n_global = d->n_global;
  p.n_local = d->n_local;
  p.row0 = d->row0;
  const float log2e = 1.4426950408889634f;
  p.k1 = d->logit_scale * log2e;
  p.k2 = d->logit_scale * log2e;
  p.uv = (const float*)(ws + L.uv);
  p.rep = d->rep;
  p.gsbuf = (const float*)(ws + L.gs);
  p.dxraw = (float*)(ws + L.dxraw);
  {
    const char* e = getenv("BSC_DBG_SKIP_TMA");
    p.dbg_skip_tma = (e && e[0] == '1') ? 1 : 0;
  }
  if (p.num_items == 0 || p.n_jtile == 0) return BSC_OK;

  // two column chunks: CTA pairs share the S recompute through DSMEM (BSC_BWD_PAIR=0 forces the unpaired kernel)
  static const bool pair_ok = []() {
    const char* e = getenv("BSC_BWD_PAIR");
    return !(e && e[0] == '0');
  }();
  if (p.n_chunks == 2 && pair_ok && sm_count() >= 2) {
    auto pk = tc_bwd_pair_kernel<0>;
    static bool pair_attr = false;
    if (!pair_attr) {
      BSC_CUDA(cudaFuncSetAttribute(pk, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)BWD_SMEM_BYTES));
      pair_attr = true;
    }
    p.num_items = n * p.n_iblk;  // the chunk is the cluster rank
    int clusters = sm_count() / 2;
    if (clusters > p.num_items) clusters = p.num_items;
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3((unsigned)(2 * clusters));
    cfg.blockDim = dim3(384);
    cfg.dynamicSmemBytes = BWD_SMEM_BYTES;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    timing_begin(1, st);
    BSC_CUDA(cudaLaunchKernelEx(&cfg, pk, p));
    timing_end(1, st);
    BSC_LAUNCH_CHECK("tc_bwd_pair_kernel");
    return BSC_OK;
  }
  auto kern = d->x16_dtype == BSC_BF16 ? tc_bwd_kernel<1> : tc_bwd_kernel<0>;
  static bool attr_done[2] = {false, false};
  const int ki = d->x16_dtype == BSC_BF16 ? 1 : 0;
  if (!attr_done[ki]) {
    BSC_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)BWD_SMEM_BYTES));
    attr_done[ki] = true;
  }
  const int grid = p.num_items < sm_count() ? p.num_items : sm_count();
  timing_begin(1, st);
  kern<<<grid, 256, BWD_SMEM_BYTES, st>>>(p);
  timing_end(1, st);
  BSC_LAUNCH_CHECK("tc_bwd_kernel");
  return BSC_OK;
}

}  // namespace bsc
