// tcgen05 InfoNCE backward sweep on CTA pairs (tcgen05.mma.cta_group::2), transposed formulation.
//
// Why: with cta_group::1 every MMA fetches 128 rows of A and all N rows of B from the issuing SM's shared memory;
// measured on B200 the M128 x N128 MMAs of the unpaired sweep keep the tensor pipe only 50 % busy even with no TMA
// traffic at all (operand fetch, not the epilogue or L2, is the limit).  A pair MMA halves the B bytes per SM.
//
// One 2-CTA cluster owns (ordered modality pair x->y, row block I of 128 rows) and sweeps the columns J two tiles
// (2 x 128) per step; CTA c owns tile 2s+c.  Everything is computed TRANSPOSED so that both GEMMs have M = 256:
//
//   S^T step    D[256 j x 128 i] = Y_{J0 u J1}[256 x K] * X_I[128 x K]^T            (K = d, 64 per stage)
//               A = the CTA's own Y tile (K-major TMA boxes), B = 64 rows of X_I per CTA (N split over the pair)
//               -> each CTA's TMEM columns 0..127 hold S^T of ITS tile: lane = j, column = i.
//   epilogue    thread (lane j) turns 64 dot products into weights  G[j][i] = gs*(exp(s*cos - s)(u_i + v_j) - 2 T_ij)/|y_j|
//               and writes them as rows of a [128 j][64 i] fp16 span (128B-swizzled, i contiguous = MN-major B operand).
//               Span c (i in 64c..64c+63) stays in the producing CTA, the other span is staged and bulk-copied
//               (cp.async.bulk shared::cta -> shared::cluster) into the peer.
//   grad step   per tile t and 256-wide block b of d:  D'[256 d x 128 i] += Y_t[128 j x 256 d]^T * G_t[128 j x 128 i]
//               A = Y_t read MN-major from the very TMA boxes (each CTA its own 128 d columns), B = G_t (each CTA its
//               64-wide span) -> TMEM columns 128 + 128 b + i, lane = d: the accumulator of dXhat^T.
//   drain       lane = d, column = i: a warp stores 32 consecutive d of one row i (coalesced).
//
// Barriers (L = leader CTA 0, the only MMA issuer):
//   full[s]   (L)     1 arrive.expect_tx by L's producer for the bytes of BOTH CTAs; both producers' TMA complete on it
//   empty[s]  (each)  tcgen05.commit multicast 0b11
//   s_full    (each)  commit multicast 0b11          s_free (L) 16 = 8 epilogue warps of each CTA
//   g_ready[0](L)     4 local warps (span 0 written) + 1 forwarded arrival (CTA 1 received span 1 of tile 2s)
//   g_ready[1](L)     4 remote arrivals (CTA 1 wrote span 1) + 4 remote arrive.expect_tx (CTA 1 ships span 0 into L)
//   rx        (CTA 1) 4 remote arrive.expect_tx + bytes: span 1 of tile 2s arriving from L; forwarded to g_ready[0]
//   g_free[c] (CTA c) commit multicast 1<<c after the grad MMAs of tile 2s+c
//   acc_full  (each)  commit multicast 0b11          acc_free (L) 16
// Warp roles (384 threads): 0 TMA producer, 1 MMA issuer (L) / forwarder (CTA 1), 2 TMEM allocator, 4..11 epilogue.
#include "tc_common.cuh"

namespace bsc {
namespace {

constexpr int BM = 128;
constexpr int BN = 128;
constexpr int BK = 64;
constexpr int SLOT = 16384;             // one 128-row TMA box (128 rows x 128 bytes)
constexpr int STAGE_BYTES = 2 * SLOT;   // S stage: {Y box 16 KB, X half box 8 KB}; grad stage: two Y boxes (128 d columns)
constexpr int S_STAGE_TX = SLOT + SLOT / 2;
constexpr int STAGES = 5;
constexpr int GSPAN = BN * 64 * 2;      // [128 j][64 i] fp16 = 16 KB
constexpr int TMEM_COLS = 512;
constexpr int ACC_COL0 = 128;

struct DuoParams {
  CUtensorMap tmap128[3];  // box 64 columns x 128 rows
  CUtensorMap tmap64[3];   // box 64 columns x 64 rows
  int n_xy;
  int xy_x[6], xy_y[6], xy_slot[6];
  int n_iblk, n_jtile, n_steps, kblocks, nblk, num_items, d;
  int jsplit;  // the column sweep of a (pair, row block) is cut into jsplit items (own dxraw slot each) so that the
               // number of items per cluster is not quantised coarsely when there are few row blocks (many ranks)
  int64_t n_global, n_local, row0;
  float k1, k2;
  const float* uv;      // [6][N]
  const int32_t* rep;   // [N]
  const float* gsbuf;   // [6][2]
  const float* inv[3];
  float* dxraw;         // [3][2][jsplit][n_local][d]
};

// item -> (ordered pair, row block, part of the column sweep); steps [s_lo, s_hi) of part `part`
__device__ __forceinline__ void duo_item(const DuoParams& p, int item, int& xy, int& iblk, int& part, int& s_lo,
                                         int& s_hi) {
  part = item % p.jsplit;
  const int rest = item / p.jsplit;
  iblk = rest % p.n_iblk;
  xy = rest / p.n_iblk;
  s_lo = (int)((int64_t)part * p.n_steps / p.jsplit);
  s_hi = (int)((int64_t)(part + 1) * p.n_steps / p.jsplit);
}

struct __align__(8) DuoTail {
  uint64_t full[STAGES];
  uint64_t empty[STAGES];
  uint64_t s_full, s_free;
  uint64_t g_ready[2];
  uint64_t g_free[2];
  uint64_t rx;
  uint64_t acc_full, acc_free;
  uint32_t tmem_base;
  uint32_t pad;
  float inv_i[BM];  // 1/|x_i| of the item's rows (0 past the end)
  float ugk_i[BM];  // gs * u_i / k1
  int rep_i[BM];    // label class (-2 past the end)
};

// ring | GB[0] | GB[1] | OUT | tail
constexpr size_t DUO_SMEM_BYTES = (size_t)STAGES * STAGE_BYTES + 3 * (size_t)GSPAN + sizeof(DuoTail);
static_assert(DUO_SMEM_BYTES <= 232448, "pair backward kernel exceeds the 227 KB shared-memory limit");

__global__ void __launch_bounds__(384, 1) tc_bwd_duo_kernel(const __grid_constant__ DuoParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  const uint32_t ring = ptx::smem_u32(smem_raw);
  if ((ring & 1023u) != 0) {
    if (threadIdx.x == 0) printf("bsc: dynamic shared memory is not 1024-byte aligned (0x%x)\n", ring);
    __trap();
  }
  const uint32_t gb0 = ring + STAGES * STAGE_BYTES;  // GB[p] = gb0 + p * GSPAN, OUT = gb0 + 2 * GSPAN
  DuoTail* tail = reinterpret_cast<DuoTail*>(smem_raw + (size_t)STAGES * STAGE_BYTES + 3 * (size_t)GSPAN);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t crank = ptx::cluster_ctarank();
  const uint32_t peer = crank ^ 1u;
  const bool leader = crank == 0;
  const int cluster = (int)ptx::cluster_id_x();
  const int n_clusters = (int)ptx::n_clusters_x();

  if (threadIdx.x == 0) {
    for (int s = 0; s < STAGES; ++s) {
      ptx::mbar_init(ptx::smem_u32(&tail->full[s]), 1);
      ptx::mbar_init(ptx::smem_u32(&tail->empty[s]), 1);
    }
    ptx::mbar_init(ptx::smem_u32(&tail->s_full), 1);
    ptx::mbar_init(ptx::smem_u32(&tail->s_free), 16);
    ptx::mbar_init(ptx::smem_u32(&tail->g_ready[0]), 5);
    ptx::mbar_init(ptx::smem_u32(&tail->g_ready[1]), 8);
    ptx::mbar_init(ptx::smem_u32(&tail->g_free[0]), 1);
    ptx::mbar_init(ptx::smem_u32(&tail->g_free[1]), 1);
    ptx::mbar_init(ptx::smem_u32(&tail->rx), 4);
    ptx::mbar_init(ptx::smem_u32(&tail->acc_full), 1);
    ptx::mbar_init(ptx::smem_u32(&tail->acc_free), 16);
    ptx::fence_barrier_init();
  }
  if (warp == 0 && lane == 0) {
    for (int a = 0; a < 3; ++a) {
      ptx::prefetch_tmap(&p.tmap128[a]);
      ptx::prefetch_tmap(&p.tmap64[a]);
    }
  }
  if (warp == 2) {
    ptx::tmem_alloc_2cta(ptx::smem_u32(&tail->tmem_base), TMEM_COLS);
    ptx::tmem_relinquish_2cta();
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::cluster_sync_all();  // both CTAs' barriers and TMEM exist before anything crosses the pair
  ptx::tc_fence_after();
  const uint32_t tmem_base = tail->tmem_base;
  const int nblk = p.nblk;

  if (warp == 0) {
    // ================= TMA producer (both CTAs, identical stage sequence) =================
    if (ptx::elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      auto next_stage = [&]() {
        if (++stage == STAGES) {
          stage = 0;
          phase ^= 1u;
        }
      };
      for (int item = cluster; item < p.num_items; item += n_clusters) {
        int xy, iblk, part, s_lo, s_hi;
        duo_item(p, item, xy, iblk, part, s_lo, s_hi);
        const int x = p.xy_x[xy], y = p.xy_y[xy];
        const CUtensorMap* my = &p.tmap128[y];
        const CUtensorMap* mx = &p.tmap64[x];
        const int xrow = (int)(p.row0 + (int64_t)iblk * BM) + 64 * (int)crank;
        auto load_s = [&](int s) {
          const int jrow = (2 * s + (int)crank) * BN;
          for (int kb = 0; kb < p.kblocks; ++kb) {
            ptx::mbar_wait(ptx::smem_u32(&tail->empty[stage]), phase ^ 1u, 100 + stage);
            const uint32_t full_l = ptx::mapa(ptx::smem_u32(&tail->full[stage]), 0);
            const uint32_t sa = ring + stage * STAGE_BYTES;
            if (leader) ptx::mbar_arrive_expect_tx(ptx::smem_u32(&tail->full[stage]), 2 * S_STAGE_TX);
            ptx::tma_load_2d_2cta(sa, my, full_l, kb * BK, jrow);
            ptx::tma_load_2d_2cta(sa + SLOT, mx, full_l, kb * BK, xrow);
            next_stage();
          }
        };
        auto load_g = [&](int t) {
          for (int b = 0; b < nblk; ++b) {
            ptx::mbar_wait(ptx::smem_u32(&tail->empty[stage]), phase ^ 1u, 110 + stage);
            const uint32_t full_l = ptx::mapa(ptx::smem_u32(&tail->full[stage]), 0);
            const uint32_t sa = ring + stage * STAGE_BYTES;
            if (leader) ptx::mbar_arrive_expect_tx(ptx::smem_u32(&tail->full[stage]), 2 * STAGE_BYTES);
            const int dc = 256 * b + 128 * (int)crank;
            ptx::tma_load_2d_2cta(sa, my, full_l, dc, t * BN);
            ptx::tma_load_2d_2cta(sa + SLOT, my, full_l, dc + 64, t * BN);
            next_stage();
          }
        };
        load_s(s_lo);
        for (int s = s_lo; s < s_hi; ++s) {
          if (s + 1 < s_hi) load_s(s + 1);
          load_g(2 * s);
          load_g(2 * s + 1);
        }
      }
    }
  } else if (warp == 1) {
    if (leader) {
      // ================= MMA issuer: one elected thread of the leader CTA drives both tensor cores =================
      if (ptx::elect_one()) {
        // S^T: A = Y (K-major), B = X half (K-major); grad: A = Y (MN-major), B = G span (MN-major)
        constexpr uint32_t idesc_s = umma_idesc(256, BM, 0, 0, 0, 0);
        constexpr uint32_t idesc_g = umma_idesc(256, BM, 0, 0, 1, 1);
        const uint64_t desc_k = ptx::umma_desc_sw128(0, 16, 1024);
        const uint64_t desc_mn = ptx::umma_desc_sw128(0, SLOT, 1024);
        int stage = 0;
        uint32_t phase = 0;
        uint32_t s_count = 0;
        uint32_t gu[2] = {0u, 0u};
        uint32_t item_count = 0;
        auto next_stage = [&]() {
          if (++stage == STAGES) {
            stage = 0;
            phase ^= 1u;
          }
        };
        for (int item = cluster; item < p.num_items; item += n_clusters, ++item_count) {
          int xy, iblk, part, s_lo, s_hi;
          duo_item(p, item, xy, iblk, part, s_lo, s_hi);
          auto issue_s = [&]() {
            ptx::mbar_wait_cluster(ptx::smem_u32(&tail->s_free), (s_count & 1u) ^ 1u, 200);
            ptx::tc_fence_after();
            for (int kb = 0; kb < p.kblocks; ++kb) {
              ptx::mbar_wait(ptx::smem_u32(&tail->full[stage]), phase, 300 + stage);
              ptx::tc_fence_after();
              const uint32_t sa = ring + stage * STAGE_BYTES;
              const uint64_t da = desc_k + ptx::desc_addr(sa);
              const uint64_t db = desc_k + ptx::desc_addr(sa + SLOT);
#pragma unroll
              for (int k = 0; k < BK / 16; ++k)
                ptx::umma_f16_2cta(tmem_base, da + 2 * k, db + 2 * k, idesc_s, (kb | k) != 0 ? 1u : 0u);
              ptx::umma_commit_2cta(ptx::smem_u32(&tail->empty[stage]), 3);
              if (kb == p.kblocks - 1) ptx::umma_commit_2cta(ptx::smem_u32(&tail->s_full), 3);
              next_stage();
            }
            ++s_count;
          };
          auto issue_g = [&](int t, bool first, bool last) {
            const uint32_t pb = (uint32_t)(t & 1);  // producer CTA of this tile's weights
            if (first) ptx::mbar_wait_cluster(ptx::smem_u32(&tail->acc_free), (item_count & 1u) ^ 1u, 210);
            ptx::mbar_wait_cluster(ptx::smem_u32(&tail->g_ready[pb]), gu[pb] & 1u, 220 + pb);
            ptx::fence_proxy_async_all();
            ptx::tc_fence_after();
            const uint64_t gbd = desc_mn + ptx::desc_addr(gb0 + pb * GSPAN);
            for (int b = 0; b < nblk; ++b) {
              ptx::mbar_wait(ptx::smem_u32(&tail->full[stage]), phase, 310 + stage);
              ptx::tc_fence_after();
              const uint32_t sa = ring + stage * STAGE_BYTES;
              const uint64_t ya = desc_mn + ptx::desc_addr(sa);
              const uint32_t d_tmem = tmem_base + ACC_COL0 + b * 128;
#pragma unroll
              for (int kk = 0; kk < BN / 16; ++kk)
                ptx::umma_f16_2cta(d_tmem, ya + (uint64_t)(kk * 128), gbd + (uint64_t)(kk * 128), idesc_g,
                                   (first && kk == 0) ? 0u : 1u);
              ptx::umma_commit_2cta(ptx::smem_u32(&tail->empty[stage]), 3);
              if (b == nblk - 1) {
                ptx::umma_commit_2cta(ptx::smem_u32(&tail->g_free[pb]), (uint16_t)(1u << pb));
                if (last) ptx::umma_commit_2cta(ptx::smem_u32(&tail->acc_full), 3);
              }
              next_stage();
            }
            ++gu[pb];
          };
          issue_s();
          for (int s = s_lo; s < s_hi; ++s) {
            if (s + 1 < s_hi) issue_s();
            issue_g(2 * s, s == s_lo, false);
            issue_g(2 * s + 1, false, s == s_hi - 1);
          }
        }
      }
    } else {
      // ================= forwarder (CTA 1): span 1 of the leader's tile has landed here -> tell the leader =========
      if (ptx::elect_one()) {
        const uint32_t ready0_l = ptx::mapa(ptx::smem_u32(&tail->g_ready[0]), 0);
        uint32_t n = 0;
        for (int item = cluster; item < p.num_items; item += n_clusters) {
          int xy, iblk, part, s_lo, s_hi;
          duo_item(p, item, xy, iblk, part, s_lo, s_hi);
          for (int s = s_lo; s < s_hi; ++s, ++n) {
            ptx::mbar_wait_cluster(ptx::smem_u32(&tail->rx), n & 1u, 500);
            ptx::fence_proxy_async_all();
            ptx::mbar_arrive_remote(ready0_l);
          }
        }
      }
    }
  } else if (warp >= 4) {
    // ================= epilogue (8 warps): S^T (TMEM) -> weight spans; accumulator drain =================
    // warp e = warp - 4: TMEM lane quadrant q = e & 3 (j rows 32q..32q+31 of this CTA's tile), span h = e >> 2
    // (i columns 64h..64h+63).  Span h == crank stays here (GB[crank]); the other one goes through OUT to the peer.
    const int e8 = warp - 4;
    const int q = e8 & 3;
    const int h = e8 >> 2;
    const int jl = q * 32 + lane;      // j row inside the tile
    const int et = threadIdx.x - 128;  // 0..255
    const float inv_k1 = 1.0f / p.k1;
    const bool keep = (uint32_t)h == crank;
    const uint32_t dst_local = keep ? gb0 + crank * GSPAN : gb0 + 2 * GSPAN;
    const uint32_t piece = (uint32_t)q * 4096u;  // this warp's 32 rows x 128 B
    const uint32_t peer_dst = ptx::mapa(gb0 + crank * GSPAN + piece, peer);
    // barrier that tells the leader "this warp's part of the tile is in place"
    const uint32_t ready_l = ptx::mapa(ptx::smem_u32(&tail->g_ready[crank]), 0);
    // barrier on which the shipped bytes complete: CTA 0 ships into CTA 1 (rx), CTA 1 ships into the leader (g_ready[1])
    const uint32_t ship_bar = leader ? ptx::mapa(ptx::smem_u32(&tail->rx), 1) : ready_l;
    const uint32_t s_free_l = ptx::mapa(ptx::smem_u32(&tail->s_free), 0);
    const uint32_t acc_free_l = ptx::mapa(ptx::smem_u32(&tail->acc_free), 0);
    uint32_t dcount = 0;  // steps processed so far (all items)
    uint32_t item_count = 0;
    for (int item = cluster; item < p.num_items; item += n_clusters, ++item_count) {
      int xy, iblk, part, s_lo, s_hi;
      duo_item(p, item, xy, iblk, part, s_lo, s_hi);
      const int x = p.xy_x[xy], y = p.xy_y[xy];
      const float gs = p.gsbuf[2 * pair_id(x, y)];
      const float m2gk = -2.0f * gs * inv_k1;
      const float* vrow = p.uv + (int64_t)pair_id(y, x) * p.n_global;
      const float* invy = p.inv[y];

      // facts of the item's rows i (columns of S^T): staged once per item
      ptx::named_bar_sync(2, 256);  // the previous item is done with them
      if (et < BM) {
        const int64_t arow = (int64_t)iblk * BM + et;
        const bool ok = arow < p.n_local;
        tail->inv_i[et] = ok ? p.inv[x][p.row0 + arow] : 0.f;
        tail->ugk_i[et] = ok ? gs * inv_k1 * p.uv[(int64_t)pair_id(x, y) * p.n_global + p.row0 + arow] : 0.f;
        tail->rep_i[et] = ok ? p.rep[p.row0 + arow] : -2;
      }
      ptx::named_bar_sync(1, 256);

      // facts of this thread's column j (a lane of S^T), fetched one step ahead
      float n_sc = 0.f, n_vi = 0.f;
      int n_rep = -1;
      auto fetch_j = [&](int s) {
        n_sc = 0.f;
        n_vi = 0.f;
        n_rep = -1;
        const int64_t j = (int64_t)(2 * s + (int)crank) * BN + jl;
        if (s < s_hi && j < p.n_global) {
          const float ij = invy[j];
          n_sc = p.k1 * ij;
          n_vi = gs * vrow[j] * ij;
          n_rep = p.rep[j];
        }
      };
      fetch_j(s_lo);

      for (int s = s_lo; s < s_hi; ++s, ++dcount) {
        const float sc_j = n_sc, vi_j = n_vi;
        const int rep_j = n_rep;
        const float tvj = m2gk * sc_j;
        fetch_j(s + 1);
        ptx::mbar_wait(ptx::smem_u32(&tail->s_full), dcount & 1u, 400);
        ptx::tc_fence_after();
        const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(h * 64);
        uint32_t r0[32], r1[32];
        ptx::tmem_ld_32x32(taddr, r0);
        ptx::tmem_ld_32x32(taddr + 32, r1);
        ptx::tmem_ld_wait();
        ptx::tc_fence_before();
        __syncwarp();
        if (lane == 0) ptx::mbar_arrive_remote_nodata(s_free_l);

        const float* fi = &tail->inv_i[h * 64];
        const float* fu = &tail->ugk_i[h * 64];
        const int* fr = &tail->rep_i[h * 64];
        uint32_t pk[32];
#pragma unroll
        for (int c = 0; c < 64; c += 2) {
          float gq[2];
#pragma unroll
          for (int z = 0; z < 2; ++z) {
            const int cc = c + z;
            const float dot = __uint_as_float(cc < 32 ? r0[cc & 31] : r1[cc & 31]);
            const float ex = ptx::ex2(fmaf(dot, fi[cc] * sc_j, -p.k2));
            const float wgt = fmaf(fu[cc], sc_j, vi_j);
            const float tv = (fr[cc] == rep_j) ? tvj : 0.f;
            gq[z] = fmaf(ex, wgt, tv);
          }
          const __half2 hh = __floats2half2_rn(gq[0], gq[1]);
          pk[c >> 1] = *reinterpret_cast<const uint32_t*>(&hh);
        }
        // GB[crank] here and in the peer, and OUT, were last read by the grad MMAs / the bulk copy of the previous step
        ptx::mbar_wait(ptx::smem_u32(&tail->g_free[crank]), (dcount & 1u) ^ 1u, 410);
        const uint32_t goff = (uint32_t)(jl >> 3) * 1024u + (uint32_t)(jl & 7) * 128u;
#pragma unroll
        for (int u = 0; u < 8; ++u) {
          const uint32_t off = goff + ((uint32_t)(u ^ (jl & 7)) << 4);
          asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(dst_local + off), "r"(pk[4 * u]),
                       "r"(pk[4 * u + 1]), "r"(pk[4 * u + 2]), "r"(pk[4 * u + 3])
                       : "memory");
        }
        ptx::fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) {
          if (keep) {
            ptx::mbar_arrive_remote_nodata(ready_l);
          } else {
            ptx::mbar_arrive_expect_tx_remote_cta(ship_bar, 4096u);
            ptx::bulk_copy_to_peer(peer_dst, dst_local + piece, 4096u, ship_bar);
          }
        }
      }

      // ---- drain: lane = d (this CTA's 128 of every 256-block), column = i ----
      ptx::mbar_wait(ptx::smem_u32(&tail->acc_full), item_count & 1u, 420);
      ptx::tc_fence_after();
      {
        const int slot = p.xy_slot[xy];
        float* out = p.dxraw + ((int64_t)((x * 2 + slot) * p.jsplit + part) * p.n_local + (int64_t)iblk * BM) * p.d;
        const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + ACC_COL0;
        const int nc = nblk * 4;  // 32-column chunks
        const int c_lo = h == 0 ? 0 : nc / 2;
        const int c_hi = h == 0 ? nc / 2 : nc;
        const int64_t rows_left = p.n_local - (int64_t)iblk * BM;
        for (int c = c_lo; c < c_hi; ++c) {
          uint32_t raw[32];
          ptx::tmem_ld_32x32(taddr + c * 32, raw);
          ptx::tmem_ld_wait();
          const int b = c >> 2;
          const int i0 = (c & 3) * 32;
          const int dcol = 256 * b + 128 * (int)crank + jl;
#pragma unroll
          for (int w = 0; w < 32; ++w)
            if (i0 + w < rows_left) out[(int64_t)(i0 + w) * p.d + dcol] = __uint_as_float(raw[w]);
        }
      }
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive_remote_nodata(acc_free_l);
    }
  }

  ptx::tc_fence_before();
  __syncthreads();
  ptx::cluster_sync_all();  // neither CTA leaves (or frees TMEM) while the other may still use its memory
  if (warp == 2) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc_2cta(tmem_base, TMEM_COLS);
  }
}

}  // namespace

// The recomputing sweep (BSC_BWD_RECOMPUTE): S is formed once more per ORDERED pair and nothing N^2-sized reaches HBM.
bool tc_bwd_duo_covers(const bsc_infonce_desc* d) {
  return d->mode == BSC_MODE_TC16 && d->x16_dtype == BSC_F16 && d->d % 256 == 0 && d->d <= 768 && device_sm_count() >= 2;
}

// In how many parts to cut every (ordered pair, row block) column sweep: the clusters take whole items, so with few
// row blocks (n_local small: many ranks) ceil(items / clusters) wastes up to a whole sweep.  Cost model per item:
// steps * (cycles per step) + a fixed fill / drain cost; pick the split in {1, 2, 4, 8} with the lowest makespan.
int tc_bwd_duo_jsplit(const bsc_infonce_desc* d) {
  if (!tc_bwd_duo_covers(d)) return 1;
  int m = 0;
  for (int a = 0; a < 3; ++a) m += d->xh[a] ? 1 : 0;
  const int64_t items0 = (int64_t)m * (m - 1) * ((d->n_local + BM - 1) / BM);
  const int64_t steps = ((d->n_global + BN - 1) / BN + 1) / 2;
  const int clusters = device_sm_count() / 2;
  if (items0 <= 0 || steps <= 0) return 1;
  // measured on B200 (1 GPU, N = 8192 and 32768, forced splits 1, 2, 4, 8): T = K0 + rounds * (steps * c + F) fits
  // with c = 5.2 us (9.9 k cycles) per step and F = 10 us (19 k cycles) per item; K0 (86 us per launch) is common
  const double step_cost = 9900.0, item_cost = 19000.0;
  int best = 1;
  double best_t = 1e300;
  for (int s = 1; s <= 8; s *= 2) {
    if (s > steps) break;
    const double rounds = (double)((items0 * s + clusters - 1) / clusters);
    const double t = rounds * ((double)steps / s * step_cost + item_cost);
    if (t < best_t * 0.995) {
      best_t = t;
      best = s;
    }
  }
  return best;
}

int tc_bwd_duo(const bsc_infonce_desc* d, const InfoNceBwdWs& L, char* ws, const float* aux, cudaStream_t st) {
  BSC_REQUIRE(tc_bwd_duo_covers(d), BSC_EUNSUPPORTED, "the recomputing sweep covers d in {256, 512, 768} (got %d)", d->d);
  BSC_REQUIRE(d->n_global < (1ll << 31) - 512, BSC_EUNSUPPORTED, "n_global too large for 32-bit TMA coordinates");
  DuoParams p{};
  int first = -1;
  for (int a = 0; a < 3; ++a)
    if (d->xh[a]) {
      int rc = make_tmap_2d(&p.tmap128[a], d->xh[a], d->x16_dtype, d->n_global, d->d, 128);
      if (rc) return rc;
      rc = make_tmap_2d(&p.tmap64[a], d->xh[a], d->x16_dtype, d->n_global, d->d, 64);
      if (rc) return rc;
      p.inv[a] = d->inv[a];
      if (first < 0) first = a;
    }
  for (int a = 0; a < 3; ++a)
    if (!d->xh[a]) {
      p.tmap128[a] = p.tmap128[first];
      p.tmap64[a] = p.tmap64[first];
    }
  int n = 0;
  for (int x = 0; x < 3; ++x) {
    if (!d->xh[x]) continue;
    int slot = 0;
    for (int y = 0; y < 3; ++y) {
      if (y == x || !d->xh[y]) continue;
      p.xy_x[n] = x;
      p.xy_y[n] = y;
      p.xy_slot[n] = slot;
      ++n;
      ++slot;
    }
  }
  p.n_xy = n;
  p.n_iblk = (int)((d->n_local + BM - 1) / BM);
  p.n_jtile = (int)((d->n_global + BN - 1) / BN);
  p.n_steps = (p.n_jtile + 1) / 2;
  p.kblocks = d->d / BK;
  p.nblk = d->d / 256;
  p.jsplit = L.kparts > 0 ? L.kparts : 1;
  if (p.jsplit > p.n_steps) p.jsplit = p.n_steps > 0 ? p.n_steps : 1;
  p.num_items = n * p.n_iblk * p.jsplit;
  p.d = d->d;
  p.n_global = d->n_global;
  p.n_local = d->n_local;
  p.row0 = d->row0;
  const float log2e = 1.4426950408889634f;
  p.k1 = d->logit_scale * log2e;
  p.k2 = d->logit_scale * log2e;
  p.uv = aux;
  p.rep = d->rep;
  p.gsbuf = aux + 6 * d->n_global;
  p.dxraw = (float*)(ws + L.dxraw);
  if (p.num_items == 0 || p.n_jtile == 0) return BSC_OK;

  BSC_SET_SMEM_ONCE(tc_bwd_duo_kernel, DUO_SMEM_BYTES);
  int clusters = device_sm_count() / 2;
  if (clusters > p.num_items) clusters = p.num_items;
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3((unsigned)(2 * clusters));
  cfg.blockDim = dim3(384);
  cfg.dynamicSmemBytes = DUO_SMEM_BYTES;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 2;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  timing_begin(1, st);
  BSC_CUDA(cudaLaunchKernelEx(&cfg, tc_bwd_duo_kernel, p));
  timing_end(1, st);
  BSC_LAUNCH_CHECK("tc_bwd_duo_kernel");
  return BSC_OK;
}

}  // namespace bsc
