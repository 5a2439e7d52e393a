// Taxonomy bookkeeping of the retrieval output as integer kernels (scripts/inference_and_eval.py:422-433 label
// mapping, :448-464 top_k_micro_accuracy, :467-511 top_k_macro_accuracy).  Labels are int32 ids per level (the host
// wrapper interns the strings); everything here is exact integer arithmetic, the final ratios are formed on the host
// in the reference's summation order.
#include "common.cuh"

namespace bsc {
namespace {

constexpr int kMaxLevels = 8;
constexpr int kMaxK = 8;  // entries of k_list

struct TaxoArgs {
  const int64_t* indices;  // [nq][k] bank ids from the search (-1 = padding)
  int64_t nq, nb;
  int k, levels, nk;
  const int32_t* key_ids;  // [nb][levels]
  const int32_t* gt_ids;   // [nq][levels]
  int32_t k_list[kMaxK];
  int64_t class_off[kMaxLevels];  // first slot of every level inside the per-class arrays
  int32_t* pred_ids;       // [nq][k][levels] or null
  uint8_t* first_hit;      // [nq][levels]: smallest rank whose label equals the ground truth, 255 = none (or null)
  unsigned long long* micro;  // [nk][levels] hits
  int32_t* pred_counts;    // [nk][total_classes]
  int32_t* gt_counts;      // [total_classes]
  int64_t total_classes;
};

__global__ void __launch_bounds__(256) taxo_kernel(const TaxoArgs a) {
  __shared__ unsigned int sh_micro[kMaxK * kMaxLevels];
  for (int t = threadIdx.x; t < a.nk * a.levels; t += blockDim.x) sh_micro[t] = 0u;
  __syncthreads();
  const int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (q < a.nq) {
    for (int l = 0; l < a.levels; ++l) {
      const int gt = a.gt_ids[q * a.levels + l];
      int first = 255;
      for (int t = 0; t < a.k; ++t) {
        const int64_t id = a.indices[q * a.k + t];
        const int lab = (id >= 0 && id < a.nb) ? a.key_ids[id * a.levels + l] : -1;
        if (a.pred_ids) a.pred_ids[(q * a.k + t) * a.levels + l] = lab;
        if (lab == gt && gt >= 0 && first == 255) first = t;
      }
      if (a.first_hit) a.first_hit[q * a.levels + l] = (uint8_t)first;
      if (gt >= 0) {
        atomicAdd(&a.gt_counts[a.class_off[l] + gt], 1);
        for (int kk = 0; kk < a.nk; ++kk)
          if (first < a.k_list[kk]) {
            atomicAdd(&sh_micro[kk * a.levels + l], 1u);
            atomicAdd(&a.pred_counts[(int64_t)kk * a.total_classes + a.class_off[l] + gt], 1);
          }
      }
    }
  }
  __syncthreads();
  for (int t = threadIdx.x; t < a.nk * a.levels; t += blockDim.x)
    if (sh_micro[t]) atomicAdd(&a.micro[t], (unsigned long long)sh_micro[t]);
}

}  // namespace

int taxo_launch(const int64_t* indices, int64_t nq, int k, const int32_t* key_ids, int64_t nb, const int32_t* gt_ids,
                int levels, const int32_t* k_list, int nk, const int64_t* class_off, int64_t total_classes,
                int32_t* pred_ids, uint8_t* first_hit, int64_t* micro, int32_t* pred_counts, int32_t* gt_counts,
                cudaStream_t st) {
  BSC_REQUIRE(levels >= 1 && levels <= kMaxLevels, BSC_EUNSUPPORTED, "bsc_taxonomy_eval: levels=%d (1..%d)", levels,
              kMaxLevels);
  BSC_REQUIRE(nk >= 1 && nk <= kMaxK, BSC_EUNSUPPORTED, "bsc_taxonomy_eval: %d entries in k_list (1..%d)", nk, kMaxK);
  BSC_REQUIRE(k >= 1 && k < 255, BSC_EUNSUPPORTED, "bsc_taxonomy_eval: k=%d", k);
  BSC_CUDA(cudaMemsetAsync(micro, 0, sizeof(int64_t) * nk * levels, st));
  BSC_CUDA(cudaMemsetAsync(pred_counts, 0, sizeof(int32_t) * (size_t)nk * total_classes, st));
  BSC_CUDA(cudaMemsetAsync(gt_counts, 0, sizeof(int32_t) * (size_t)total_classes, st));
  if (nq == 0) return BSC_OK;
  TaxoArgs a{};
  a.indices = indices;
  a.nq = nq;
  a.nb = nb;
  a.k = k;
  a.levels = levels;
  a.nk = nk;
  a.key_ids = key_ids;
  a.gt_ids = gt_ids;
  for (int t = 0; t < nk; ++t) a.k_list[t] = k_list[t];
  for (int l = 0; l < levels; ++l) a.class_off[l] = class_off[l];
  a.pred_ids = pred_ids;
  a.first_hit = first_hit;
  a.micro = (unsigned long long*)micro;
  a.pred_counts = pred_counts;
  a.gt_counts = gt_counts;
  a.total_classes = total_classes;
  taxo_kernel<<<(unsigned)((nq + 255) / 256), 256, 0, st>>>(a);
  BSC_LAUNCH_CHECK("taxo_kernel");
  return BSC_OK;
}

}  // namespace bsc
