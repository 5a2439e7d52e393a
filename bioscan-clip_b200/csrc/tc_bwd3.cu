// Gradient GEMMs of the InfoNCE backward from STORED fp16 weight blocks (tcgen05.mma.cta_group::2 on CTA pairs), fused
// with the reduce-scatter of the partial gradients over NVLink peer memory.
//
// tc_sim_kernel<EPI_WEIGHTS> (tc_sim.cu) wrote, for every pair q = (a, b), the block G_q[local rows i][all columns j],
//     g_ij = gs(a,b) (e_ij (u_i + v_j) - 2 T_ij) / |b_j|,
// as 128B-swizzled images [128 i][64 j] (16 KB each).  One persistent kernel runs every gradient GEMM from them:
//   K-major job   dA[local i][d]  = sum_j G[i][j] * B16[j][d]      A = an image as it lies in HBM (raw TMA copy), K = j over
//                                   ALL rows of b (the gathered fp16 rows: 1/|b_j| is in the weights, operands exact)
//   MN-major job  dB[all j][d]    = sum_i G[i][j] * Aunit[i][d]    A = halves of two images read MN-major, K = i over the
//                                   LOCAL rows only; B = fp16 unit rows of the local rows of a.  Each rank obtains a
//                                   PARTIAL sum for every row j of the global batch; the epilogue stores row block q of
//                                   the result straight into rank q's exchange buffer (slot = this rank): a GEMM whose
//                                   epilogue is the reduce-scatter.  combine (prep.cu) adds the W slots in rank order.
// BSC_BWD_BLOCKS (no peer memory): ordered pairs, K-major jobs only.  BSC_BWD_SCATTER: unordered pairs, both jobs --
// every rank executes 6 + 12 n N D in the backward, what a single GPU executes for its n = N rows.
//   B operand = 64 rows x {256 | 384} d (MN-major), each CTA its half of the accumulator's d columns;
//   accumulators [128 lanes][256 (+128) columns] fp32 in TMEM.
// Work items are equal in size by construction (the K-major jobs are K-split `world` ways, so that both kinds run n/64
// K blocks) and are dealt to the CTA pairs from a host-built schedule that alternates rounds of scattered and local
// items: the NVLink stores are spread over the whole kernel and over all SMs, and rank r starts with the row blocks
// of rank r+1 so that no two ranks target the same peer at the same time.
// Warp roles (384 threads): 0 TMA producer, 1 MMA issuer (leader CTA), 2 TMEM allocator, 4..11 epilogue.
#include <algorithm>
#include <cstring>
#include <vector>

#include "tc_common.cuh"

namespace bsc {

int make_tmap_raw128(CUtensorMap* out, const void* base, int64_t rows, int box_rows);

namespace {

constexpr int BK = 64;
constexpr int A_BYTES = 16384;            // one stored image
constexpr int B_BOX = 8192;               // [64 rows][64 d] fp16
constexpr int TMEM_COLS = 512;
constexpr int MAX_STAGES = 6;
constexpr int MAX_JOBS = 6;
constexpr int MAX_SCHED = 6144;           // schedule entries carried in the kernel parameters (24 KB)
constexpr int MAX_CLUSTERS = 96;          // CTA pairs of the grid (74 on a B200)
constexpr int STG_BYTES = 2048;           // per epilogue warp: 32 rows x 16 floats staged for coalesced scatter stores
// WIDE = 0: 256 d columns per item, one accumulator (N = 256), TMEM double-buffered over items, 6 x 32 KB stages.
// WIDE = 1: 384 d columns per item, two accumulators (N = 256 and N = 128) fed by the SAME A tile -- d = 768 is covered
//           by two items instead of three, i.e. the weight images are read twice per GEMM instead of three times;
//           one TMEM buffer (384 of 512 columns), 5 x 40 KB stages.
template <int WIDE>
struct Cfg {
  static constexpr int BN = WIDE ? 384 : 256;
  static constexpr int NBOX = WIDE ? 3 : 2;  // 64-column boxes of B per CTA and stage
  static constexpr int STAGE_BYTES = A_BYTES + NBOX * B_BOX;
  static constexpr int STAGES = WIDE ? 5 : 6;
  static constexpr int NBUF = WIDE ? 1 : 2;
  static constexpr size_t SMEM = 1024 + (size_t)STAGES * STAGE_BYTES + 8 * STG_BYTES + 256;
};
static_assert(Cfg<1>::SMEM <= 232448 && Cfg<0>::SMEM <= 232448, "gradient GEMM kernel exceeds 227 KB of shared memory");

struct RevParams {
  CUtensorMap tmap_g128;  // gimg as [rows][64] fp16, box 64 x 128, no swizzle (the bytes are already swizzled)
  CUtensorMap tmap_g64;   // same, box 64 x 64: half an image (MN-major jobs)
  CUtensorMap tmap_in[MAX_JOBS];  // B operand rows of every job, box 64 d x 64 rows, SWIZZLE_128B
  CUtensorMap tmap_in_lo[MAX_JOBS];  // BSC_MODE_TC16X3: the lo parts of the same rows
  // terms = 1: out = G * in.  terms = 3 (BSC_MODE_TC16X3): G = Ghi + Glo (two image sets, the lo set lo_img0 images after
  // the hi set), in = hi + lo; every item runs its K range three times into the SAME accumulator, small terms first:
  //   Glo * in_hi,  Ghi * in_lo,  Ghi * in_hi
  int terms;
  int64_t lo_img0;
  // job = one gradient GEMM  out[M rows][d] = A[M][K] * in[K][d];  image(pair, r, c) = (pair * img_r + r) * img_c + c
  //   K-major job:  A block (128 rows mblk, 64 K kb)  = image(pair, mblk, kb)
  //   MN-major job: A block (128 rows mblk, 64 K kb)  = rows 64 (kb & 1) .. + 63 of images (pair, kb >> 1, 2 mblk + {0, 1})
  int n_jobs;
  int job_pair[MAX_JOBS], job_mn[MAX_JOBS], job_nk[MAX_JOBS], job_kparts[MAX_JOBS];
  int64_t job_rows[MAX_JOBS];      // valid output rows
  float* job_out[MAX_JOBS];        // local jobs: [kparts][rows][d]; scattered jobs: null
  int job_xslot[MAX_JOBS];         // scattered jobs: pair slot inside the exchange buffers
  int job_fnum[MAX_JOBS], job_fden[MAX_JOBS];  // epilogue row factor gsbuf[fnum] * gsbuf[fden] ...
  const float* job_finv[MAX_JOBS];             // ... / finv[row] when not null
  float job_fmul[MAX_JOBS];                    // ... times a constant (undoes the 2^8 of the X3 unit rows)
  int img_r, img_c;
  int num_items, d, n_clusters;
  // scatter destinations: exchange buffer of rank q = float [world * mn_kparts][n_xslots][n_local][d]
  float* peer[kMaxPeers];
  int world, rank, n_xslots;
  int mn_kparts;  // K parts of a scattered job: exchange slot of (source rank, part) = rank * mn_kparts + part
  int64_t n_local;
  const float* gsbuf;
  int sched_off[MAX_CLUSTERS + 1];  // cluster c runs sched[sched_off[c]] .. sched[sched_off[c + 1] - 1], in this order
  uint32_t sched[MAX_SCHED];        // item -> job (3 bits) | d chunk (3) | K part (8) | 256-row block (18)
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

__host__ __device__ inline uint32_t pack_item(int job, int c, int part, int mb) {
  return (uint32_t)job | ((uint32_t)c << 3) | ((uint32_t)part << 6) | ((uint32_t)mb << 14);
}

__device__ __forceinline__ void rev_item(const RevParams& p, int idx, int& job, int& mb, int& part, int& c, int& kb_lo,
                                         int& kb_hi) {
  const uint32_t w = p.sched[idx];
  job = (int)(w & 7u);
  c = (int)((w >> 3) & 7u);
  part = (int)((w >> 6) & 255u);
  mb = (int)(w >> 14);
  const int nk = p.job_nk[job], kp = p.job_kparts[job];
  kb_lo = (int)((int64_t)part * nk / kp);
  kb_hi = (int)((int64_t)(part + 1) * nk / kp);
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
  uint8_t* stg_base = smem_raw + (ring - raw_addr) + (size_t)STAGES * STAGE_BYTES;
  RevTail* tail = reinterpret_cast<RevTail*>(stg_base + 8 * STG_BYTES);
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t crank = ptx::cluster_ctarank();
  const int item0 = p.sched_off[ptx::cluster_id_x()];      // this cluster's slice of the host-built schedule
  const int item_end = p.sched_off[ptx::cluster_id_x() + 1];

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
    for (int a = 0; a < p.n_jobs; ++a) ptx::prefetch_tmap(&p.tmap_in[a]);
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
      for (int item = item0; item < item_end; ++item) {
        int job, mb, part, c, kb_lo, kb_hi;
        rev_item(p, item, job, mb, part, c, kb_lo, kb_hi);
        const CUtensorMap* mu = &p.tmap_in[job];
        const int mblk = 2 * mb + (int)crank;
        const int64_t pair0 = (int64_t)p.job_pair[job] * p.img_r;
        const bool mn = p.job_mn[job] != 0;
        const int d0 = BN * c + 128 * (int)crank;       // this CTA's half of the N = 256 accumulator
        const int d1 = BN * c + 256 + 64 * (int)crank;  // ... and of the N = 128 accumulator (WIDE)
        for (int term = 0; term < p.terms; ++term) {
          if (p.terms == 3) mu = term == 1 ? &p.tmap_in_lo[job] : &p.tmap_in[job];
          const int64_t set0 = (p.terms == 3 && term == 0) ? p.lo_img0 : 0;  // first image of the set this term reads
          for (int kb = kb_lo; kb < kb_hi; ++kb) {
            ptx::mbar_wait(ptx::smem_u32(&tail->empty[stage]), phase ^ 1u, 100 + stage);
            const uint32_t full = ptx::smem_u32(&tail->full[stage]);
            const uint32_t full_l = ptx::mapa(full, 0);
            const uint32_t sa = ring + stage * STAGE_BYTES;
            if (crank == 0) ptx::mbar_arrive_expect_tx(full, 2 * STAGE_BYTES);
            if (!mn) {
              const int64_t img = set0 + (pair0 + mblk) * p.img_c + kb;
              ptx::tma_load_2d_2cta(sa, &p.tmap_g128, full_l, 0, (int32_t)(img * 128));
            } else {
              const int64_t img = set0 + (pair0 + (kb >> 1)) * p.img_c + 2 * mblk;
              const int32_t r0 = (int32_t)(img * 128) + 64 * (kb & 1);
              ptx::tma_load_2d_2cta(sa, &p.tmap_g64, full_l, 0, r0);
              ptx::tma_load_2d_2cta(sa + A_BYTES / 2, &p.tmap_g64, full_l, 0, r0 + 128);
            }
            // columns past d (d % 256 == 128 leaves the odd CTA's half of the last chunk empty) are zero-filled by TMA
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
      for (int item = item0; item < item_end; ++item, ++count) {
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
        const int nsteps = (kb_hi - kb_lo) * p.terms;  // the terms of BSC_MODE_TC16X3 share the accumulator
        for (int step = 0; step < nsteps; ++step) {
          ptx::mbar_wait(ptx::smem_u32(&tail->full[stage]), phase, 300 + stage);
          ptx::tc_fence_after();
          const uint32_t sa = ring + stage * STAGE_BYTES;
          const uint64_t da = desc_a + ptx::desc_addr(sa);
          const uint64_t db = desc_b + ptx::desc_addr(sa + A_BYTES);
#pragma unroll
          for (int k = 0; k < BK / 16; ++k) {
            const uint32_t acc = (step != 0 || k != 0) ? 1u : 0u;
            ptx::umma_f16_2cta(d_tmem, da + a_step * k, db + (uint64_t)(k * 128), idesc, acc);
            if (WIDE) {
              const uint64_t db2 = desc_b + ptx::desc_addr(sa + A_BYTES + 2 * B_BOX);
              ptx::umma_f16_2cta(d_tmem + 256, da + a_step * k, db2 + (uint64_t)(k * 128), idesc2, acc);
            }
          }
          ptx::umma_commit_2cta(ptx::smem_u32(&tail->empty[stage]), 3);
          if (step == nsteps - 1) ptx::umma_commit_2cta(ptx::smem_u32(&tail->tmem_full[buf]), 3);
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
    uint8_t* stg = stg_base + (warp - 4) * STG_BYTES;
    uint32_t count = 0;
    for (int item = item0; item < item_end; ++item, ++count) {
      int job, mb, part, c, kb_lo, kb_hi;
      rev_item(p, item, job, mb, part, c, kb_lo, kb_hi);
      const uint32_t buf = NBUF == 2 ? (count & 1u) : 0u;
      const uint32_t use = NBUF == 2 ? (count >> 1) : count;
      const int64_t rows = p.job_rows[job];
      const int64_t row_w0 = ((int64_t)(2 * mb + (int)crank)) * 128 + q * 32;  // first row of this warp
      const int64_t row = row_w0 + lane;
      float f = 0.f;
      if (row < rows) {
        f = p.gsbuf[p.job_fnum[job]] * p.gsbuf[p.job_fden[job]];
        if (p.job_finv[job] != nullptr) {
          const float iv = p.job_finv[job][row];
          f = iv > 0.f ? f / iv : 0.f;
        }
        f *= p.job_fmul[job];
      }
      ptx::mbar_wait(ptx::smem_u32(&tail->tmem_full[buf]), use & 1u, 400 + buf);
      ptx::tc_fence_after();
      const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + buf * BN;
      float* out_local = p.job_out[job];
      if (out_local != nullptr) {
        // ---- local partial sums: every thread stores its own row, 128 contiguous bytes per chunk ----
        float* out = out_local + ((int64_t)part * rows + row) * p.d + BN * c;
#pragma unroll 1
        for (int cc = grp * CPG; cc < grp * CPG + CPG; ++cc) {
          uint32_t raw[32];
          ptx::tmem_ld_32x32(taddr + cc * 32, raw);
          ptx::tmem_ld_wait();
          if (row < rows && BN * c + cc * 32 < p.d) {
            float4* o4 = reinterpret_cast<float4*>(out + cc * 32);
#pragma unroll
            for (int t = 0; t < 8; ++t)
              o4[t] = make_float4(f * __uint_as_float(raw[4 * t]), f * __uint_as_float(raw[4 * t + 1]),
                                  f * __uint_as_float(raw[4 * t + 2]), f * __uint_as_float(raw[4 * t + 3]));
          }
        }
      } else {
        // ---- scattered partial sums: row j of the global batch belongs to rank j / n_local.  16 columns at a time are
        // staged through 2 KB of shared memory per warp so that four lanes store the 64 contiguous bytes of one row
        // (peer stores travel over NVLink: 16-byte stores to 32 different rows would waste most of every packet).
        // Staging slot of (row r, 16-byte chunk u) = u ^ ((r >> 1) & 3): conflict-free both ways.
        const int64_t slab = p.n_local * (int64_t)p.d;
        const int64_t src_off =
            (((int64_t)p.rank * p.mn_kparts + part) * p.n_xslots + p.job_xslot[job]) * slab + BN * c;
        const int t4 = lane & 3;
#pragma unroll 1
        for (int cc = grp * 2 * CPG; cc < grp * 2 * CPG + 2 * CPG; ++cc) {  // 16-column chunks
          uint32_t raw[16];
          ptx::tmem_ld_32x16(taddr + cc * 16, raw);
          ptx::tmem_ld_wait();
#pragma unroll
          for (int u = 0; u < 4; ++u)
            *reinterpret_cast<float4*>(stg + lane * 64 + ((u ^ ((lane >> 1) & 3)) << 4)) =
                make_float4(f * __uint_as_float(raw[4 * u]), f * __uint_as_float(raw[4 * u + 1]),
                            f * __uint_as_float(raw[4 * u + 2]), f * __uint_as_float(raw[4 * u + 3]));
          __syncwarp();
          if (BN * c + cc * 16 < p.d) {
#pragma unroll
            for (int k = 0; k < 4; ++k) {
              const int r = 8 * k + (lane >> 2);
              const int64_t j = row_w0 + r;
              if (j < rows) {
                const float4 v = *reinterpret_cast<const float4*>(stg + r * 64 + ((t4 ^ ((r >> 1) & 3)) << 4));
                const int dest = (int)(j / p.n_local);
                const int64_t jl = j - (int64_t)dest * p.n_local;
                *reinterpret_cast<float4*>(p.peer[dest] + src_off + jl * p.d + cc * 16 + t4 * 4) = v;
              }
            }
          }
          __syncwarp();
        }
      }
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive_remote_nodata(empty_l0 + buf * 8u);
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

}  // namespace

static int rev_launch(RevParams& p, bool wide, cudaStream_t st) {
  if (p.num_items == 0) return BSC_OK;
  BSC_SET_SMEM_ONCE(tc_bwd_rev_kernel<0>, Cfg<0>::SMEM);
  BSC_SET_SMEM_ONCE(tc_bwd_rev_kernel<1>, Cfg<1>::SMEM);
  const int clusters = p.n_clusters;
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
  if (wide)
    BSC_CUDA(cudaLaunchKernelEx(&cfg, tc_bwd_rev_kernel<1>, p));
  else
    BSC_CUDA(cudaLaunchKernelEx(&cfg, tc_bwd_rev_kernel<0>, p));
  BSC_LAUNCH_CHECK("tc_bwd_rev_kernel");
  return BSC_OK;
}

// slot of partner modality `other` among the present modalities other than `m`, ascending
static int slot_of(const bsc_infonce_desc* d, int m, int other) {
  int slot = 0;
  for (int z = 0; z < other; ++z)
    if (z != m && d->xh[z]) ++slot;
  return slot;
}

static bool rev_wide(const bsc_infonce_desc* d) { return d->d % 384 == 0; }

// K split of the K-major gradient GEMMs (= local partial sums that combine adds in a fixed order).
//   SCATTER over W > 1 ranks: W parts, so that a K-major item (N / W / 64 K blocks) is as long as an MN-major item
//   (n_local / 64 K blocks) and the static round-robin over equal items stays balanced;
//   otherwise: the split that minimises the makespan when a few hundred items meet 74 clusters.
// BSC_MODE_TC16X3: the accumulator is truncated at every MMA instruction (common.cuh, x3_accumulator_gain), an error that
// grows with the length of the accumulation chain; the K-major GEMMs are therefore cut into chains of at most
// kX3ChainCols columns (the parts are added in fp32, round to nearest, by combine).
constexpr int kX3ChainCols = 1024;
static int tc_bwd_gemm_kparts_tc16(const bsc_infonce_desc* d);

int tc_bwd_gemm_kparts(const bsc_infonce_desc* d) {
  int s = tc_bwd_gemm_kparts_tc16(d);
  if (d->mode == BSC_MODE_TC16X3) {
    const int64_t nk = (d->n_global + 63) / 64;
    int64_t want = (d->n_global + kX3ChainCols - 1) / kX3ChainCols;
    if (want > 32) want = 32;  // every part is one more [n_local][d] fp32 slab per job
    if (want > nk) want = nk;
    if (s < want) s = (int)want;
  }
  return s;
}

// K split of the MN-major (scattered) gradient GEMMs, whose K runs over the LOCAL rows: 1 part (one exchange slot per
// source rank), except in BSC_MODE_TC16X3, where the same cap on the accumulation chains applies (measured at N = 32768
// on one GPU with a single 32768-row chain: gradient error 3e-5 .. 6e-5 of max |g| instead of < 1e-5).  Every part is
// one more exchange slot per source rank; combine adds the world * parts slots in (rank, part) order.
int tc_bwd_mn_kparts(const bsc_infonce_desc* d) {
  if (d->mode != BSC_MODE_TC16X3) return 1;
  int64_t want = (d->n_local + kX3ChainCols - 1) / kX3ChainCols;
  const int64_t nk = (d->n_local + 63) / 64;
  int64_t cap = 32 / (d->world > 0 ? d->world : 1);
  if (cap < 1) cap = 1;
  if (want > cap) want = cap;
  if (want > nk) want = nk;
  if (want < 1) want = 1;
  return (int)want;
}

static int tc_bwd_gemm_kparts_tc16(const bsc_infonce_desc* d) {
  int m = 0;
  for (int a = 0; a < 3; ++a) m += d->xh[a] ? 1 : 0;
  const int64_t nk = (d->n_global + 63) / 64;
  // SCATTER over W > 1 ranks: the MN-major items run n_local / 64 K blocks, a K-major item N / 64; the schedule
  // balances unequal items (longest first), so the K-major jobs are only split enough to give every cluster a few
  // long items -- each extra part is one more [n_local][d] fp32 slab written here and read again by combine
  if (d->bwd_mode == BSC_BWD_SCATTER && d->world > 1) {
    const int64_t pairs = (int64_t)m * (m - 1) / 2;
    const int64_t n_mb = (d->n_local + 255) / 256;
    const int64_t nchunk = rev_wide(d) ? d->d / 384 : (d->d + 255) / 256;
    const int64_t items0 = pairs * n_mb * nchunk;
    const int clusters = device_sm_count() / 2;
    int s = 1;
    while (items0 * s < 2 * clusters && s < d->world && 2 * s <= nk) s *= 2;
    return s;
  }
  const int64_t jobs = (int64_t)m * (m - 1);  // 2 per unordered pair (SCATTER, one GPU) = 1 per ordered pair (BLOCKS)
  const int64_t n_mb = (d->n_local + 255) / 256;
  const int64_t nchunk = rev_wide(d) ? d->d / 384 : (d->d + 255) / 256;
  const int64_t items0 = jobs * n_mb * nchunk;
  const int clusters = device_sm_count() / 2;
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

// Gradient GEMMs from the images gimg[(pair * n_iblk_pad + iblk) * n_j64 + j64] = G[128 i][64 j] that tc_bwd_weights wrote
// for the pairs (pairs_a[q], pairs_b[q]): ordered pairs in BSC_BWD_BLOCKS, unordered (a < b) in BSC_BWD_SCATTER.
int tc_bwd_gemms(const bsc_infonce_desc* d, const InfoNceBwdWs& L, char* ws, const float* aux, int n_pairs,
                 const int* pairs_a, const int* pairs_b, const ScatterDst& dst, cudaStream_t st) {
  const bool scatter = L.mode == BSC_BWD_SCATTER;
  const int64_t N = d->n_global, n = d->n_local;
  int n_iblk_pad, n_j64;
  tc_bwd_matrix_dims(d, &n_iblk_pad, &n_j64);
  static thread_local RevParams p;  // 26 KB: kept off the stack
  memset(&p, 0, sizeof(p));
  const bool x3 = d->mode == BSC_MODE_TC16X3;
  p.terms = mode_terms(d->mode);
  p.lo_img0 = (int64_t)n_pairs * n_iblk_pad * n_j64;  // the lo image set follows the hi set (tc_bwd_weights)
  const int64_t img_rows = (int64_t)n_pairs * n_iblk_pad * n_j64 * 128 * mode_parts(d->mode);
  BSC_REQUIRE(img_rows < (1ll << 31), BSC_EUNSUPPORTED, "tc_bwd_gemms: %lld image rows exceed 32-bit TMA coordinates",
              (long long)img_rows);
  int rc = make_tmap_raw128(&p.tmap_g128, ws + L.gimg, img_rows, 128);
  if (rc) return rc;
  rc = make_tmap_raw128(&p.tmap_g64, ws + L.gimg, img_rows, 64);
  if (rc) return rc;
  // B operand rows: [rows][d] fp16, or the hi / lo column halves of [rows][2 d] (BSC_MODE_TC16X3)
  const int64_t ldb = (x3 ? 2 : 1) * (int64_t)d->d;
  const size_t unit_bytes = (size_t)n * ldb * 2;  // unit rows of one modality
  const bool wide = rev_wide(d);
  const int BNW = wide ? 384 : 256;
  const int nchunk = (d->d + BNW - 1) / BNW;
  BSC_REQUIRE(nchunk <= 8, BSC_EUNSUPPORTED, "tc_bwd_gemms: d = %d needs more than 8 column chunks", d->d);
  p.img_r = n_iblk_pad;
  p.img_c = n_j64;
  p.d = d->d;
  p.gsbuf = aux + 6 * d->n_global;
  p.world = dst.world;
  p.rank = dst.rank;
  p.n_xslots = n_pairs;
  p.mn_kparts = tc_bwd_mn_kparts(d);
  p.n_local = n;
  for (int r = 0; r < dst.world && r < kMaxPeers; ++r) p.peer[r] = dst.peer[r];
  const int64_t slab = n * (int64_t)d->d;
  float* dxraw = (float*)(ws + L.dxraw);
  const int nk_all = (int)((N + 63) / 64), nk_loc = (int)((n + 63) / 64);

  // unit rows of the local rows of every modality that is the `a` side of a scattered pair
  if (scatter) {
    const void* src[3] = {nullptr, nullptr, nullptr};
    void* out[3] = {nullptr, nullptr, nullptr};
    for (int q = 0; q < n_pairs; ++q) {
      const int a = pairs_a[q];
      src[a] = d->xh[a];
      out[a] = ws + L.xunit + (size_t)a * unit_bytes;
    }
    rc = x3 ? rows_unit_f16x3_launch3(src, d->inv, d->row0, n, d->d, out, st)
            : rows_unit_f16_launch3(src, d->inv, d->row0, n, d->d, out, st);
    if (rc) return rc;
  }

  int job = 0;
  int job_nmb[MAX_JOBS];
  for (int q = 0; q < n_pairs; ++q) {
    const int a = pairs_a[q], b = pairs_b[q];
    // K-major: dA[local rows] = G_q * B16 (all rows of b); the weights carry gs(a,b) / |b_j|
    p.job_pair[job] = q;
    p.job_mn[job] = 0;
    p.job_nk[job] = nk_all;
    p.job_kparts[job] = L.kparts;
    p.job_rows[job] = n;
    p.job_out[job] = dxraw + (int64_t)(scatter ? q : a * 2 + slot_of(d, a, b)) * L.kparts * slab;
    p.job_fnum[job] = 2 * pair_id(a, b);
    p.job_fden[job] = 2 * pair_id(a, b) + 1;  // factor 1: combine divides by gs(a,b)
    p.job_finv[job] = nullptr;
    p.job_fmul[job] = 1.0f;
    job_nmb[job] = (int)((n + 255) / 256);
    rc = make_tmap_2d_ld(&p.tmap_in[job], d->xh[b], BSC_F16, N, d->d, ldb, 64);
    if (rc) return rc;
    if (x3) {
      rc = make_tmap_2d_ld(&p.tmap_in_lo[job], (const char*)d->xh[b] + (size_t)d->d * 2, BSC_F16, N, d->d, ldb, 64);
      if (rc) return rc;
    }
    ++job;
    if (!scatter) continue;
    // MN-major: partial dB[all rows] = G_q^T * Aunit (local rows of a); undo gs(a,b) / |b_j| per output row and hand
    // combine the sums in units of gs(b,a)
    p.job_pair[job] = q;
    p.job_mn[job] = 1;
    p.job_nk[job] = nk_loc;
    p.job_kparts[job] = p.mn_kparts;
    p.job_rows[job] = N;
    p.job_out[job] = nullptr;
    p.job_xslot[job] = q;
    p.job_fnum[job] = 2 * pair_id(b, a);
    p.job_fden[job] = 2 * pair_id(a, b) + 1;
    p.job_finv[job] = d->inv[b];
    p.job_fmul[job] = x3 ? 1.0f / kX3UnitScale : 1.0f;
    job_nmb[job] = (int)((N + 255) / 256);
    rc = make_tmap_2d_ld(&p.tmap_in[job], ws + L.xunit + (size_t)a * unit_bytes, BSC_F16, n, d->d, ldb, 64);
    if (rc) return rc;
    if (x3) {
      rc = make_tmap_2d_ld(&p.tmap_in_lo[job], ws + L.xunit + (size_t)a * unit_bytes + (size_t)d->d * 2, BSC_F16, n,
                           d->d, ldb, 64);
      if (rc) return rc;
    }
    ++job;
  }
  p.n_jobs = job;
  for (int z = job; z < MAX_JOBS; ++z) p.tmap_in[z] = p.tmap_in[0];
  for (int z = 0; z < MAX_JOBS; ++z)
    if (!x3 || z >= job) p.tmap_in_lo[z] = p.tmap_in[0];
  BSC_REQUIRE(L.kparts >= 1 && L.kparts <= nk_all && L.kparts <= 255, BSC_EUNSUPPORTED,
              "tc_bwd_gemms: K split %d of %d blocks", L.kparts, nk_all);

  // ---- schedule.  Items differ in length (K blocks), so they are dealt to the clusters by longest-processing-time
  // first: sorted by length (stable: the d chunks of one A tile stay adjacent and land on neighbouring clusters, where
  // they share the tile through L2), each item goes to the least-loaded cluster.  Inside a cluster the local and the
  // scattered items are then interleaved, and neighbouring cluster pairs start with opposite kinds, so that at any time
  // about half of the SMs are in a scattered item: all of them at once would push ~29 MB at the NVLink ports in one
  // burst.  A rank starts its scattered list at the row blocks of the next rank (no two ranks target the same peer).
  struct Item {
    uint32_t code;
    int cost, kind;
  };
  // the schedule depends on the shape only: built once per shape and thread, copied on the later calls
  struct SchedKey {
    int64_t n, N;
    int d, pairs, mode, world, rank, kparts, sms, terms;
    bool operator==(const SchedKey& o) const {
      return n == o.n && N == o.N && d == o.d && pairs == o.pairs && mode == o.mode && world == o.world &&
             rank == o.rank && kparts == o.kparts && sms == o.sms && terms == o.terms;
    }
  };
  // One launch carries at most MAX_SCHED items in its parameters; larger problems (BSC_MODE_TC16X3 at N = 32768: 32 K
  // parts per row block) run as several launches over consecutive slices of the item list, each with its own schedule.
  struct Batch {
    int num_items = 0, n_clusters = 0;
    int off[MAX_CLUSTERS + 1];
    std::vector<uint32_t> sched;
  };
  struct SchedCache {
    SchedKey key{};
    bool valid = false;
    std::vector<Batch> batches;
  };
  static thread_local SchedCache cache;
  const SchedKey key{n, N, d->d, n_pairs, L.mode, dst.world, dst.rank, L.kparts, device_sm_count(), p.terms};
  if (!(cache.valid && cache.key == key)) {
    std::vector<Item> items;
    for (int kind = 0; kind < 2; ++kind) {  // 0 = scattered (MN-major), 1 = local (K-major)
      for (int z = 0; z < job; ++z) {
        if ((p.job_out[z] == nullptr) != (kind == 0)) continue;
        const int nmb = job_nmb[z], kp = p.job_kparts[z];
        const int rot =
            (kind == 0 && dst.world > 1) ? (int)(((int64_t)(dst.rank + 1) % dst.world) * nmb / dst.world) : 0;
        for (int mbi = 0; mbi < nmb; ++mbi)
          for (int part = 0; part < kp; ++part) {
            const int nkb = (int)((int64_t)(part + 1) * p.job_nk[z] / kp - (int64_t)part * p.job_nk[z] / kp);
            for (int c = 0; c < nchunk; ++c)
              items.push_back(
                  Item{pack_item(z, c, part, (mbi + rot) % nmb), nkb * p.terms + 6 /* fill + drain */, kind});
          }
      }
    }
    const size_t total = items.size();
    const size_t n_batches = total ? (total + MAX_SCHED - 1) / MAX_SCHED : 1;
    cache.valid = false;
    cache.batches.assign(n_batches, Batch{});
    for (size_t bi = 0; bi < n_batches; ++bi) {
      Batch& B = cache.batches[bi];
      const size_t lo = bi * total / n_batches, hi = (bi + 1) * total / n_batches;
      std::vector<Item> part_items(items.begin() + lo, items.begin() + hi);
      int clusters = device_sm_count() / 2;
      if (clusters > MAX_CLUSTERS) clusters = MAX_CLUSTERS;
      if ((size_t)clusters > part_items.size()) clusters = (int)part_items.size();
      if (clusters < 1) clusters = 1;
      std::stable_sort(part_items.begin(), part_items.end(), [](const Item& x, const Item& y) { return x.cost > y.cost; });
      std::vector<std::vector<Item>> per(clusters);
      std::vector<long long> load(clusters, 0);
      for (const Item& it : part_items) {
        int best = 0;
        for (int cl = 1; cl < clusters; ++cl)
          if (load[cl] < load[best]) best = cl;
        per[best].push_back(it);
        load[best] += it.cost;
      }
      B.sched.reserve(part_items.size());
      for (int cl = 0; cl < clusters; ++cl) {
        B.off[cl] = (int)B.sched.size();
        std::vector<Item> lst[2];
        for (const Item& it : per[cl]) lst[it.kind].push_back(it);
        // proportional interleave of the two kinds, the starting kind alternating between neighbouring cluster pairs
        const int first = (cl >> 1) & 1;
        size_t pos[2] = {0, 0};
        const size_t n0 = lst[first].size(), n1 = lst[first ^ 1].size();
        for (size_t t = 0; t < n0 + n1; ++t) {
          // take from `first` while its share of the emitted items is not ahead of its share of the total
          const bool take_first = pos[1] >= n1 || (pos[0] < n0 && pos[0] * (n0 + n1) <= t * n0);
          const int kind = take_first ? first : first ^ 1;
          B.sched.push_back(lst[kind][pos[take_first ? 0 : 1]++].code);
        }
      }
      for (int cl = clusters; cl <= MAX_CLUSTERS; ++cl) B.off[cl] = (int)B.sched.size();
      B.num_items = (int)part_items.size();
      B.n_clusters = clusters;
    }
    cache.key = key;
    cache.valid = true;
  }
  timing_begin(3, st);  // all launches of the gradient GEMMs (one unless the schedule had to be cut)
  for (const Batch& B : cache.batches) {
    if (B.num_items == 0) continue;
    memcpy(p.sched_off, B.off, sizeof(B.off));
    memcpy(p.sched, B.sched.data(), B.sched.size() * sizeof(uint32_t));
    p.num_items = B.num_items;
    p.n_clusters = B.n_clusters;
    rc = rev_launch(p, wide, st);
    if (rc) return rc;
  }
  timing_end(3, st);
  return BSC_OK;
}

}  // namespace bsc
