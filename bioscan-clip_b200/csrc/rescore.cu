// Second stage of the exact inner-product search (IndexFlatIP.search, inference_and_eval.py:418-421): the candidates of
// the bf16 tensor-core screening pass are re-scored with exact float32 dot products and ordered by (score desc, id asc).
// A query's result is PROVEN equal to the exhaustive float32 search when its k-th exact score exceeds, by more than the
// bf16 rounding bound, the best screening score any non-candidate row can have; the others are flagged for the
// exhaustive kernel.  One warp per query.
#include "common.cuh"

namespace bsc {
namespace {

constexpr int KS = 8;  // entries per candidate list (KTOP of tc_sim.cu)

__device__ __forceinline__ float wsum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

__global__ void __launch_bounds__(128) rescore_kernel(const float* __restrict__ q, int64_t nq, const float* __restrict__ bank,
                                                      int64_t nb, int d, int64_t base_id, const float* __restrict__ cs,
                                                      const int64_t* __restrict__ ci, int parts, int k, float eps_scale,
                                                      float* __restrict__ os, int64_t* __restrict__ oi,
                                                      int32_t* __restrict__ unsure, int32_t* __restrict__ n_unsure) {
  const int lane = threadIdx.x & 31;
  const int64_t qi = (int64_t)blockIdx.x * 4 + (threadIdx.x >> 5);
  if (qi >= nq) return;
  constexpr float kLowest = -3.4028234663852886e38f;
  const float* qr = q + qi * d;
  float qn2 = 0.f;
  for (int c = lane; c < d; c += 32) qn2 = fmaf(qr[c], qr[c], qn2);
  qn2 = wsum(qn2);
  // best screening score a row OUTSIDE the lists can have: the last entry of its own (full) list
  float thr = kLowest;
  bool any_full = false;
  for (int p = lane; p < parts; p += 32) {
    const int64_t o = ((int64_t)p * nq + qi) * KS + (KS - 1);
    if (ci[o] >= 0) {
      thr = fmaxf(thr, cs[o]);
      any_full = true;
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) thr = fmaxf(thr, __shfl_xor_sync(0xffffffffu, thr, o));
  any_full = __any_sync(0xffffffffu, any_full);

  float ts[KS];
  int64_t ti[KS];
#pragma unroll
  for (int t = 0; t < KS; ++t) {
    ts[t] = kLowest;
    ti[t] = -1;
  }
  int count = 0;
  const int total = parts * KS;
  for (int c = 0; c < total; ++c) {
    const int64_t o = ((int64_t)(c / KS) * nq + qi) * KS + (c % KS);
    const int64_t id = ci[o];
    if (id < 0) continue;  // uniform across the warp
    const float* br = bank + (id - base_id) * d;
    float acc = 0.f;
    for (int e = lane; e < d; e += 32) acc = fmaf(qr[e], br[e], acc);
    const float s = wsum(acc);
    ++count;
    // insertion, identical in every lane: better = higher score, or equal score and lower id
    const bool beats_last = (ti[KS - 1] < 0) || (s > ts[KS - 1]) || (s == ts[KS - 1] && id < ti[KS - 1]);
    if (beats_last) {
      int pos = KS - 1;
#pragma unroll
      for (int u = KS - 1; u > 0; --u) {
        const bool better = (ti[u - 1] < 0) || (s > ts[u - 1]) || (s == ts[u - 1] && id < ti[u - 1]);
        if (pos == u && better) pos = u - 1;
      }
#pragma unroll
      for (int u = KS - 1; u > 0; --u)
        if (u > pos) {
          ts[u] = ts[u - 1];
          ti[u] = ti[u - 1];
        }
#pragma unroll
      for (int u = 0; u < KS; ++u)
        if (u == pos) {
          ts[u] = s;
          ti[u] = id;
        }
    }
  }
  // |exact - screening score| <= eps for every bank row (operands rounded to bf16, fp32 accumulation)
  const float eps = eps_scale * sqrtf(qn2);
  float kth = kLowest;
#pragma unroll
  for (int u = 0; u < KS; ++u)
    if (u == k - 1) kth = ts[u];
  const bool sure = !any_full || (count >= k && kth > thr + eps);
  if (lane == 0) {
    for (int u = 0; u < k; ++u) {
      float sv = kLowest;
      int64_t iv = -1;
#pragma unroll
      for (int w = 0; w < KS; ++w)
        if (w == u) {
          sv = ts[w];
          iv = ti[w];
        }
      os[qi * k + u] = iv < 0 ? kLowest : sv;
      oi[qi * k + u] = iv;
    }
    unsure[qi] = sure ? 0 : 1;
    if (!sure) atomicAdd(n_unsure, 1);
  }
}

}  // namespace

int rescore_launch(const float* q, int64_t nq, const float* bank, int64_t nb, int d, int64_t base_id, const float* cs,
                   const int64_t* ci, int parts, int k, float eps_scale, float* os, int64_t* oi, int32_t* unsure,
                   int32_t* n_unsure, cudaStream_t st) {
  BSC_REQUIRE(k >= 1 && k <= KS, BSC_EUNSUPPORTED, "bsc_bank_rescore: k=%d (1..%d)", k, KS);
  BSC_CUDA(cudaMemsetAsync(n_unsure, 0, sizeof(int32_t), st));
  if (nq == 0) return BSC_OK;
  rescore_kernel<<<(unsigned)((nq + 3) / 4), 128, 0, st>>>(q, nq, bank, nb, d, base_id, cs, ci, parts, k, eps_scale, os,
                                                            oi, unsure, n_unsure);
  BSC_LAUNCH_CHECK("rescore_kernel");
  return BSC_OK;
}

}  // namespace bsc
