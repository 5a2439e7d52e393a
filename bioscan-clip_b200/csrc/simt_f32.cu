// BSC_MODE_F32: the exact-parity path.  Plain fp32 FFMA tiles (64x64x16, 4x4 per thread), expf, fixed
// summation orders.  It exists so that loss and gradients can be compared with the fp32 reference at
// 1e-5; the throughput path is the tcgen05 code in tc_*.cu.  The backward of this mode stages the
// gradient-weight slab G (rows x N, fp32) in the workspace, so it is meant for N up to a few thousand.
#include "common.cuh"

namespace bsc {

int reduce_parts_launch(const float* part, int nparts, int64_t part_stride, int64_t n, float* out, cudaStream_t st);
int sum_rows_launch(const float* part, int nparts, int64_t part_stride, int64_t n, float scale, float* out,
                    double* out64, cudaStream_t st);

namespace {

constexpr int T = kSimtT;  // 64
constexpr int BK = 16;

// acc[r][c] += sum_k A[row0 + ty*4 + r][k] * B[col0 + tx*4 + c][k]   (B given as [n][k] rows)
// or, with B_KN, B[k][col0 + tx*4 + c] (B given as [k][n]).
template <bool B_KN>
__device__ __forceinline__ void tile_gemm(const float* __restrict__ A, int64_t lda, int64_t a_rows, int64_t arow0,
                                          const float* __restrict__ B, int64_t ldb, int64_t b_lim, int64_t bcol0,
                                          int64_t kdim, float (&acc)[4][4], float (*As)[T + 4], float (*Bs)[T + 4]) {
  const int tid = threadIdx.x;
  const int tx = tid & 15, ty = tid >> 4;
#pragma unroll
  for (int r = 0; r < 4; ++r)
#pragma unroll
    for (int c = 0; c < 4; ++c) acc[r][c] = 0.f;

  for (int64_t k0 = 0; k0 < kdim; k0 += BK) {
    // A tile: 64 rows x 16 k -> As[k][row]; thread loads 4 consecutive k of one row
    {
      const int row = tid >> 2, kq = (tid & 3) * 4;
      const int64_t gr = arow0 + row;
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int64_t gk = k0 + kq + u;
        As[kq + u][row] = (gr < a_rows && gk < kdim) ? A[gr * lda + gk] : 0.f;
      }
    }
    if (!B_KN) {
      const int row = tid >> 2, kq = (tid & 3) * 4;
      const int64_t gr = bcol0 + row;
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int64_t gk = k0 + kq + u;
        Bs[kq + u][row] = (gr < b_lim && gk < kdim) ? B[gr * ldb + gk] : 0.f;
      }
    } else {
      // B[k][n]: 16 k rows x 64 n; thread loads 4 consecutive n of one k
      const int kk = tid >> 4, nq = (tid & 15) * 4;
      const int64_t gk = k0 + kk;
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int64_t gn = bcol0 + nq + u;
        Bs[kk][nq + u] = (gk < kdim && gn < b_lim) ? B[gk * ldb + gn] : 0.f;
      }
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < BK; ++k) {
      float a[4], b[4];
#pragma unroll
      for (int r = 0; r < 4; ++r) a[r] = As[k][ty * 4 + r];
#pragma unroll
      for (int c = 0; c < 4; ++c) b[c] = Bs[k][tx * 4 + c];
#pragma unroll
      for (int r = 0; r < 4; ++r)
#pragma unroll
        for (int c = 0; c < 4; ++c) acc[r][c] = fmaf(a[r], b[c], acc[r][c]);
    }
    __syncthreads();
  }
}

struct SimArgs {
  const float* xa;  // [N][d] rows of the row-side modality (global batch)
  const float* xb;  // [N][d] rows of the column-side modality
  int64_t n_global, n_local, row0;
  int d;
  float scale;
  // forward outputs
  float* rowpart;  // [n_jtile][n_local_pad]
  float* colpart;  // [n_iblk][n_global_pad]
  float* rowpos;   // [n_jtile][n_local_pad]  sum_j T_ij (s*S_ij - s) of the tile
  const int32_t* rep;  // [N] label class representative: T_ij = (rep[i] == rep[j])
  int64_t n_local_pad, n_global_pad;
  // G outputs
  const float* u;  // [N] weights of the row side
  const float* v;  // [N] weights of the column side
  float* g;        // [g_rows][N]
  int64_t g_row0;  // first local row of this chunk
  int64_t g_rows;
};

// grid: (n_jtile, n_iblk)
template <bool WRITE_G>
__global__ void __launch_bounds__(256) simt_sim_kernel(SimArgs a) {
  __shared__ float As[BK][T + 4];
  __shared__ float Bs[BK][T + 4];
  __shared__ float cs[16][T];
  const int tid = threadIdx.x;
  const int tx = tid & 15, ty = tid >> 4;
  const int64_t il0 = (WRITE_G ? a.g_row0 : 0) + (int64_t)blockIdx.y * T;  // local row of the tile
  const int64_t j0 = (int64_t)blockIdx.x * T;
  float acc[4][4];
  // rows are addressed in the global batch: row0 + local
  tile_gemm<false>(a.xa, a.d, a.row0 + a.n_local, a.row0 + il0, a.xb, a.d, a.n_global, j0, a.d, acc, As, Bs);

  // label classes of the thread's rows and columns; -2 / -1 never match
  int repr[4], repc[4];
#pragma unroll
  for (int r = 0; r < 4; ++r) {
    const int64_t il = il0 + ty * 4 + r;
    repr[r] = il < a.n_local ? a.rep[a.row0 + il] : -2;
  }
#pragma unroll
  for (int c = 0; c < 4; ++c) {
    const int64_t j = j0 + tx * 4 + c;
    repc[c] = j < a.n_global ? a.rep[j] : -1;
  }
  float e[4][4];
  float pos[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
  for (int r = 0; r < 4; ++r)
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      const int64_t il = il0 + ty * 4 + r, j = j0 + tx * 4 + c;
      const bool ok = il < a.n_local && j < a.n_global;
      const float lg = fmaf(a.scale, acc[r][c], -a.scale);
      e[r][c] = ok ? expf(lg) : 0.f;
      // the positive term uses the very logit that enters the exponential (cancels like log_softmax)
      if (repr[r] == repc[c]) pos[r] += lg;
    }

  if constexpr (WRITE_G) {
#pragma unroll
    for (int r = 0; r < 4; ++r) {
      const int64_t il = il0 + ty * 4 + r;
      if (il >= a.n_local || il - a.g_row0 >= a.g_rows) continue;
      const float ui = a.u[a.row0 + il];
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        const int64_t j = j0 + tx * 4 + c;
        if (j < a.n_global)
          a.g[(il - a.g_row0) * a.n_global + j] = fmaf(e[r][c], ui + a.v[j], repr[r] == repc[c] ? -2.0f : 0.0f);
      }
    }
    return;
  } else {
  // row sums: over the thread's 4 columns, then over the 16 tx lanes of the half-warp (fixed order)
#pragma unroll
  for (int r = 0; r < 4; ++r) {
    float s = (e[r][0] + e[r][1]) + (e[r][2] + e[r][3]);
    float ps = pos[r];
#pragma unroll
    for (int o = 8; o > 0; o >>= 1) {
      s += __shfl_xor_sync(0xffffffffu, s, o);
      ps += __shfl_xor_sync(0xffffffffu, ps, o);
    }
    const int64_t il = il0 + ty * 4 + r;
    if (tx == 0 && il < a.n_local) {
      a.rowpart[(int64_t)blockIdx.x * a.n_local_pad + il] = s;
      a.rowpos[(int64_t)blockIdx.x * a.n_local_pad + il] = ps;
    }
  }
  // column sums: over the thread's 4 rows, then over ty through shared memory
#pragma unroll
  for (int c = 0; c < 4; ++c) cs[ty][tx * 4 + c] = (e[0][c] + e[1][c]) + (e[2][c] + e[3][c]);
  __syncthreads();
  if (tid < T) {
    float s = 0.f;
#pragma unroll
    for (int q = 0; q < 16; ++q) s += cs[q][tid];
    const int64_t j = j0 + tid;
    if (j < a.n_global) a.colpart[(int64_t)blockIdx.y * a.n_global_pad + j] = s;
  }
  }
}

// C[m][n] = sum_k G[m][k] * Y[k][n];  grid (ceil(d/64), ceil(rows/64))
__global__ void __launch_bounds__(256) simt_gy_kernel(const float* __restrict__ g, int64_t rows, int64_t kdim,
                                                      const float* __restrict__ y, int d, float* __restrict__ out) {
  __shared__ float As[BK][T + 4];
  __shared__ float Bs[BK][T + 4];
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  const int64_t m0 = (int64_t)blockIdx.y * T, n0 = (int64_t)blockIdx.x * T;
  float acc[4][4];
  tile_gemm<true>(g, kdim, rows, m0, y, d, d, n0, kdim, acc, As, Bs);
#pragma unroll
  for (int r = 0; r < 4; ++r) {
    const int64_t m = m0 + ty * 4 + r;
    if (m >= rows) continue;
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      const int64_t n = n0 + tx * 4 + c;
      if (n < d) out[m * (int64_t)d + n] = acc[r][c];
    }
  }
}

// scores[q][j] for a block of bank rows, then a per-query running top-k (one thread per query).
__global__ void __launch_bounds__(256) simt_scores_kernel(const float* __restrict__ q, int64_t nq,
                                                          const float* __restrict__ bank, int64_t nb_blk, int d,
                                                          float* __restrict__ scores, int64_t ld) {
  __shared__ float As[BK][T + 4];
  __shared__ float Bs[BK][T + 4];
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  const int64_t m0 = (int64_t)blockIdx.y * T, n0 = (int64_t)blockIdx.x * T;
  float acc[4][4];
  tile_gemm<false>(q, d, nq, m0, bank, d, nb_blk, n0, d, acc, As, Bs);
#pragma unroll
  for (int r = 0; r < 4; ++r) {
    const int64_t m = m0 + ty * 4 + r;
    if (m >= nq) continue;
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      const int64_t n = n0 + tx * 4 + c;
      if (n < nb_blk) scores[m * ld + n] = acc[r][c];
    }
  }
}

constexpr int KMAX = 32;

// One warp per query: lanes stride over the block's scores keeping a per-lane sorted top-k in shared
// memory-free registers would cost too much; instead each lane filters against the current k-th best of
// the query (kept by lane 0 semantics below) -- simple and exact: lane-private lists then a warp merge.
__global__ void __launch_bounds__(128) simt_topk_update_kernel(const float* __restrict__ scores, int64_t ld,
                                                               int64_t nq, int64_t nb_blk, int64_t id0, int k,
                                                               float* __restrict__ best_s,
                                                               int64_t* __restrict__ best_i) {
  const int64_t qi = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (qi >= nq) return;
  float bs[KMAX];
  int64_t bi[KMAX];
  for (int t = 0; t < k; ++t) {
    bs[t] = best_s[qi * k + t];
    bi[t] = best_i[qi * k + t];
  }
  const float* row = scores + qi * ld;
  for (int64_t j = 0; j < nb_blk; ++j) {
    const float s = row[j];
    // ids ascend with j and with the block order, so a tie never displaces an earlier entry
    if (bi[k - 1] >= 0 && !(s > bs[k - 1])) continue;
    int pos = k - 1;
    while (pos > 0 && (bi[pos - 1] < 0 || s > bs[pos - 1])) --pos;
    for (int u = k - 1; u > pos; --u) {
      bs[u] = bs[u - 1];
      bi[u] = bi[u - 1];
    }
    bs[pos] = s;
    bi[pos] = id0 + j;
  }
  for (int t = 0; t < k; ++t) {
    best_s[qi * k + t] = bs[t];
    best_i[qi * k + t] = bi[t];
  }
}

__global__ void topk_init_kernel(float* s, int64_t* i, int64_t n) {
  const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t < n) {
    s[t] = -3.4028234663852886e38f;
    i[t] = -1;
  }
}

}  // namespace

// ---------------------------------------------------------------------------------------------
int simt_fwd_stats(const bsc_infonce_desc* p, const InfoNceFwdWs& L, char* ws, float* stats, cudaStream_t st) {
  const int64_t N = p->n_global;
  int pair = 0;
  int npairs_done = 0;
  for (int a = 0; a < 3; ++a)
    for (int b = a + 1; b < 3; ++b, ++pair) {
      if (!p->xh[a] || !p->xh[b]) continue;
      SimArgs s{};
      s.xa = (const float*)p->xh[a];
      s.xb = (const float*)p->xh[b];
      s.n_global = N;
      s.n_local = p->n_local;
      s.row0 = p->row0;
      s.d = p->d;
      s.scale = p->logit_scale;
      s.n_local_pad = L.n_local_pad;
      s.n_global_pad = L.n_global_pad;
      s.rowpart = (float*)(ws + L.rowpart) + (int64_t)pair * L.n_jchunk * L.n_local_pad;
      s.rowpos = (float*)(ws + L.rowpos) + (int64_t)pair * L.n_jchunk * L.n_local_pad;
      s.rep = p->rep;
      s.colpart = (float*)(ws + L.colpart) + (int64_t)pair * L.n_iblk * L.n_global_pad;
      dim3 grid((unsigned)L.n_jtile, (unsigned)L.n_iblk);
      simt_sim_kernel<false><<<grid, 256, 0, st>>>(s);
      BSC_LAUNCH_CHECK("simt_sim_kernel<fwd>");
      int rc = reduce_parts_launch(s.rowpart, L.n_jchunk, L.n_local_pad, p->n_local,
                                   stats + (int64_t)pair_id(a, b) * N + p->row0, st);
      if (rc) return rc;
      rc = reduce_parts_launch(s.colpart, L.n_iblk, L.n_global_pad, N, stats + (int64_t)pair_id(b, a) * N, st);
      if (rc) return rc;
      // compact the positive partials of this pair into slot `npairs_done` so one sum covers all pairs
      rc = reduce_parts_launch(s.rowpos, L.n_jchunk, L.n_local_pad, p->n_local,
                               (float*)(ws + L.rowpart) + (int64_t)npairs_done * L.n_local_pad, st);
      if (rc) return rc;
      ++npairs_done;
    }
  // the record's positive term: float copy at [6N], the value the loss is computed from as part 0 of the double parts
  return sum_rows_launch((float*)(ws + L.rowpart), npairs_done, L.n_local_pad, p->n_local, 1.0f, stats + 6 * N,
                         reinterpret_cast<double*>(stats + stats_pos_offset(N)), st);
}

int simt_bwd_raw(const bsc_infonce_desc* p, const InfoNceBwdWs& L, char* ws, const float* aux, cudaStream_t st) {
  const int64_t N = p->n_global;
  const float* uv = aux;
  float* g = (float*)(ws + L.gscratch);
  for (int x = 0; x < 3; ++x) {
    if (!p->xh[x]) continue;
    int slot = 0;
    for (int y = 0; y < 3; ++y) {
      if (y == x || !p->xh[y]) continue;
      float* out = (float*)(ws + L.dxraw) + ((int64_t)(x * 2 + slot)) * p->n_local * p->d;
      for (int64_t r0 = 0; r0 < p->n_local; r0 += L.g_rows) {
        const int64_t rows = (p->n_local - r0 < L.g_rows) ? p->n_local - r0 : L.g_rows;
        SimArgs s{};
        s.xa = (const float*)p->xh[x];
        s.xb = (const float*)p->xh[y];
        s.n_global = N;
        s.n_local = p->n_local;
        s.row0 = p->row0;
        s.d = p->d;
        s.scale = p->logit_scale;
        s.rep = p->rep;
        s.u = uv + (int64_t)pair_id(x, y) * N;
        s.v = uv + (int64_t)pair_id(y, x) * N;
        s.g = g;
        s.g_row0 = r0;
        s.g_rows = rows;
        dim3 grid((unsigned)((N + T - 1) / T), (unsigned)((rows + T - 1) / T));
        simt_sim_kernel<true><<<grid, 256, 0, st>>>(s);
        BSC_LAUNCH_CHECK("simt_sim_kernel<g>");
        dim3 grid2((unsigned)((p->d + T - 1) / T), (unsigned)((rows + T - 1) / T));
        simt_gy_kernel<<<grid2, 256, 0, st>>>(g, rows, N, (const float*)p->xh[y], p->d, out + r0 * p->d);
        BSC_LAUNCH_CHECK("simt_gy_kernel");
      }
      ++slot;
    }
  }
  return BSC_OK;
}

// ---------------------------------------------------------------------------------------------
// Exact fp32 retrieval: score blocks of kBankBlk bank rows into the workspace, then update the per-query
// top-k.  Bank rows are visited in ascending id, so ties keep the lowest id.
// ---------------------------------------------------------------------------------------------
constexpr int64_t kBankBlk = 16384;

size_t simt_bank_ws(int64_t nq, int64_t nb, int d, int k) {
  (void)d;
  (void)k;
  const int64_t blk = nb < kBankBlk ? nb : kBankBlk;
  return align_up((size_t)nq * (size_t)(blk > 0 ? blk : 1) * sizeof(float), 256);
}

int simt_bank_search(const float* q, int64_t nq, const float* bank, int64_t nb, int d, int k, int64_t base_id,
                     float* out_s, int64_t* out_i, char* ws, size_t ws_bytes, cudaStream_t st) {
  BSC_REQUIRE(k >= 1 && k <= KMAX, BSC_EUNSUPPORTED, "bsc_bank_search: k=%d outside [1,%d]", k, KMAX);
  BSC_REQUIRE(ws_bytes >= simt_bank_ws(nq, nb, d, k), BSC_EWORKSPACE, "bsc_bank_search: workspace too small");
  if (nq == 0) return BSC_OK;
  topk_init_kernel<<<(unsigned)((nq * k + 255) / 256), 256, 0, st>>>(out_s, out_i, nq * k);
  BSC_LAUNCH_CHECK("topk_init_kernel");
  float* scores = (float*)ws;
  for (int64_t b0 = 0; b0 < nb; b0 += kBankBlk) {
    const int64_t blk = (nb - b0 < kBankBlk) ? nb - b0 : kBankBlk;
    dim3 grid((unsigned)((blk + T - 1) / T), (unsigned)((nq + T - 1) / T));
    simt_scores_kernel<<<grid, 256, 0, st>>>(q, nq, bank + b0 * d, blk, d, scores, blk);
    BSC_LAUNCH_CHECK("simt_scores_kernel");
    simt_topk_update_kernel<<<(unsigned)((nq + 127) / 128), 128, 0, st>>>(scores, blk, nq, blk, base_id + b0, k,
                                                                         out_s, out_i);
    BSC_LAUNCH_CHECK("simt_topk_update_kernel");
  }
  return BSC_OK;
}

}  // namespace bsc
