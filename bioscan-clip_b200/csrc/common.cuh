// Shared host/device helpers for libbsc_hot.so (sm_100a only).
#pragma once

#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>

#include <cstdarg>
#include <cstdint>
#include <cstdio>

#include "../../include/bsc_hot.h"

namespace bsc {

// ---- error plumbing ---------------------------------------------------------------------------
void set_error(const char* fmt, ...);
int cuda_fail(cudaError_t e, const char* what);
void count_launch(int n = 1);

#define BSC_CUDA(call)                                   \
  do {                                                   \
    cudaError_t e__ = (call);                            \
    if (e__ != cudaSuccess) return ::bsc::cuda_fail(e__, #call); \
  } while (0)

#define BSC_LAUNCH_CHECK(name)                           \
  do {                                                   \
    cudaError_t e__ = cudaGetLastError();                \
    if (e__ != cudaSuccess) return ::bsc::cuda_fail(e__, name); \
    ::bsc::count_launch();                               \
  } while (0)

#define BSC_REQUIRE(cond, code, ...)  \
  do {                                \
    if (!(cond)) {                    \
      ::bsc::set_error(__VA_ARGS__);  \
      return (code);                  \
    }                                 \
  } while (0)

// ---- ordered-pair bookkeeping -------------------------------------------------------------------
// Ordered pair (a,b), a != b, of modality slots 0..2 -> 0..5.
__host__ __device__ inline int pair_id(int a, int b) { return a * 2 + (b > a ? b - 1 : b); }

constexpr int kMaxPeers = 16;  // ranks of one NVLink domain addressed by the fused gather
constexpr int kStatsExtra = 4;  // scalars after the 6*N sum-exp rows; [0] = sum over unordered pairs and local rows of
                                // sum_j T_ij (s*S_ij - s)

inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

// ---- workspace layout of the infonce calls -----------------------------------------------------
struct InfoNceWs {
  size_t uv;        // float [6][N]          r_i / sumexp
  size_t gs;        // float [4]             {gs, 1/gs}
  size_t rowpart;   // float [3][n_jchunk][n_local_pad]   row sum-exp partials
  size_t rowpos;    // float [3][n_jchunk][n_local_pad]   positive-pair partials
  size_t colpart;   // float [3][n_iblk][N_pad]
  size_t jrange;    // int2  [n_jtile]       label-class range of every column tile
  size_t dxraw;     // float [3][2][bwd_jsplit][n_local][D]
  size_t gscratch;  // float [g_rows][N]     fp32 mode only
  size_t xunit;     // half  [3][N][D]       unit rows (stored-weight backward only)
  size_t gimg;      // half  [pairs][2*n_steps][2*n_iblk][128][64]  weight spans of the backward sweep (stored-weight only)
  size_t total;
  // tiling facts shared by host and kernels
  int n_iblk, n_jtile, n_jchunk, jtiles_per_chunk;
  int bwd_jsplit;   // dxraw parts per (modality, slot): > 1 only for the CTA-pair backward sweep (tc_bwd2.cu)
  int bwd_gstore;   // tc_bwd_gstore(): 0 recompute, 1 stored spans of the pair sweep, 2 weight matrix + GEMMs
  int64_t n_local_pad, n_global_pad, g_rows;
};

InfoNceWs infonce_ws_layout(const bsc_infonce_desc* p);


// ---- argument blocks of the O(N*D) kernels in prep.cu ----------------------------------------------
struct CombineArgs {
  const void* x[3];      // original local rows (un-normalised), null when the modality is absent
  void* grad[3];
  const float* raw[3];   // [2][n_local][d] per modality
  const float* gsbuf;
  const float* grad_out;
  int64_t row0, n_local, n_global, ldx;
  int d, n_slots;
  int n_parts;           // partial sums per slot (column-sweep parts of the backward), >= 1
  unsigned present_mask;
  float scale;
};

// ---- sub-launchers implemented in the .cu files ---------------------------------------------------
int simt_fwd_stats(const bsc_infonce_desc* p, const InfoNceWs& L, char* ws, float* stats, cudaStream_t st);
int simt_bwd_raw(const bsc_infonce_desc* p, const InfoNceWs& L, char* ws, cudaStream_t st);
int tc_fwd_stats(const bsc_infonce_desc* p, const InfoNceWs& L, char* ws, float* stats, cudaStream_t st);
int tc_bwd_raw(const bsc_infonce_desc* p, const InfoNceWs& L, char* ws, cudaStream_t st);
// CTA-pair (cta_group::2) sweep; returns 1 when the shape is not covered and the caller must use tc_bwd_raw's own kernels
int tc_bwd_duo(const bsc_infonce_desc* p, const InfoNceWs& L, char* ws, cudaStream_t st);
int tc_bwd_duo_jsplit(const bsc_infonce_desc* p);
// stored-weight backward (all rows local): is it used for this problem, and how many bytes of weight spans it keeps
// 0 = recomputing sweep, 1 = pair sweep per unordered pair + stored spans, 2 = weight matrix + two GEMMs per pair
// (all rows local), 3 = weight blocks per ordered pair + one GEMM each (this rank's rows of a multi-GPU run)
int tc_bwd_gstore(const bsc_infonce_desc* p);
size_t tc_bwd_gimg_bytes(const bsc_infonce_desc* p, int mode);
void tc_bwd_matrix_dims(const bsc_infonce_desc* d, int* n_iblk_pad, int* n_j64);
int tc_bwd_gemm_ksplit(const bsc_infonce_desc* d, int mode);
int tc_bwd_matrix(const bsc_infonce_desc* d, const InfoNceWs& L, char* ws, cudaStream_t st);
int tc_bwd_gemms(const bsc_infonce_desc* d, const InfoNceWs& L, char* ws, const void* const xunit[3], int n_iblk_pad,
                 int n_j64, int n_pairs, const int* pairs_a, const int* pairs_b, int ordered, cudaStream_t st);
int tc_bwd_rev(const bsc_infonce_desc* d, const InfoNceWs& L, char* ws, const void* const xunit[3], int n_pairs,
               const int* pairs_x, const int* pairs_y, int n_jb, int n_i64, cudaStream_t st);
int rows_unit_f16_launch(const void* x16, const float* inv, int64_t n, int d, void* out, cudaStream_t st);
int rows_unit_f16_launch3(const void* const x16[3], const float* const inv[3], int64_t n, int d, void* const out[3],
                          cudaStream_t st);
struct ReduceJob {
  const float* part;
  int nparts;
  int64_t part_stride, n;
  float* out;
};
struct ReduceJobs {
  ReduceJob job[6];
  int count;
};
int reduce_parts_batch_launch(const ReduceJobs& jobs, cudaStream_t st);

int simt_bank_search(const float* q, int64_t nq, const float* bank, int64_t nb, int d, int k, int64_t base_id,
                     float* out_s, int64_t* out_i, char* ws, size_t ws_bytes, cudaStream_t st);
int tc_bank_search(const void* q, int64_t nq, const void* bank, int64_t nb, int d, int k, int64_t base_id,
                   float* out_s, int64_t* out_i, char* ws, size_t ws_bytes, cudaStream_t st);
size_t simt_bank_ws(int64_t nq, int64_t nb, int d, int k);
size_t tc_bank_ws(int64_t nq, int64_t nb, int d, int k);

// tile shapes of the tensor-core kernels (host needs them for the workspace layout)
constexpr int kTcBM = 128;   // rows per CTA tile (TMEM lanes)
constexpr int kTcBNF = 256;  // forward: columns per tile
constexpr int kTcBNB = 128;  // backward: columns per tile
constexpr int kSimtT = 64;   // SIMT tile edge

// timing slots (bench instrumentation)
void timing_begin(int which, cudaStream_t st);
void timing_end(int which, cudaStream_t st);

}  // namespace bsc
