// Shared host/device helpers for libbsc_hot.so (sm_100a only).
#pragma once

#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>

#include <cstdarg>
#include <cstdint>
#include <cstdio>

#include "../../include/bsc_hot.h"

// Bank scan with few queries (32-query passes over a shard).  1: one work item per SM, candidate re-reads in groups of
// four TMEM columns, block-per-query merge of the candidate lists; 2 (default): also at most 64 queries spread over the
// four TMEM lane quarters (SimParams::spread).  0 = the earlier layout (two items per SM, single-column re-reads,
// warp-per-query merge), kept for A/B builds (python build.py --variant v0 -DBSC_SCAN_V2=0).  Measured on one B200,
// 32 queries x a 625k-row shard: scan 305 / 249 / 172 us, whole search 682 / 296 / 211 us for 0 / 1 / 2
// (profiles/r2_scan_v2_ab.txt).
#ifndef BSC_SCAN_V2
#define BSC_SCAN_V2 2
#endif

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

constexpr int kMaxPeers = 16;  // ranks of one NVLink domain addressed by the fused gather / gradient scatter
constexpr int kStatsExtra = 4;  // scalars after the 6*N sum-exp rows; [0] = sum over unordered pairs and local rows of
                                // sum_j T_ij (s*S_ij - s) (float copy, informational in a per-rank record)
constexpr int kPosParts = 64;   // the same sum as kPosParts double partial sums (fixed assignment of rows to parts):
                                // what the loss is computed from, added over parts and ranks in a fixed order
// One statistics record = 6 N floats | kStatsExtra floats | kPosParts doubles
__host__ __device__ inline int64_t stats_record_len(int64_t n) { return (6 * n + kStatsExtra + 2 * kPosParts + 3) / 4 * 4; }
__host__ __device__ inline int64_t stats_pos_offset(int64_t n) { return (6 * n + kStatsExtra + 1) / 2 * 2; }  // float index, 8-byte aligned

inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

// tensor-core modes: BSC_MODE_TC16 (one fp16 part per row) and BSC_MODE_TC16X3 (hi + lo parts, three MMA terms)
inline bool mode_is_tc(int mode) { return mode == BSC_MODE_TC16 || mode == BSC_MODE_TC16X3; }
inline int mode_terms(int mode) { return mode == BSC_MODE_TC16X3 ? 3 : 1; }
inline int mode_parts(int mode) { return mode == BSC_MODE_TC16X3 ? 2 : 1; }  // fp16 parts stored per value
// tcgen05.mma accumulates TOWARD ZERO: every instruction (K = 16) truncates the fp32 accumulator, so a dot product whose
// partial sums grow steadily comes out too small by a data-independent share -- measured on B200 (tools/acc_probe.py,
// profiles/r2_acc_probe.log: always negative, 0.19-0.31 ulp per instruction with bf16 operands; 0.48 ulp per
// instruction of the hi*hi term with the fp16 parts of BSC_MODE_TC16X3, where it showed as a 2.7e-6 bias of every
// cosine at d = 768, i.e. 3e-5 of a peaked-softmax loss).  BSC_MODE_TC16X3 multiplies the similarities by the inverse
// of the expected shrinkage (first-order compensation; the spread around it, ~1e-7, stays).
constexpr double kX3AccUlpPerStep = 0.48;
inline double x3_accumulator_gain(int d) { return 1.0 + kX3AccUlpPerStep * (d / 16) * 1.1920928955078125e-7; }
constexpr float kX3UnitScale = 256.0f;  // BSC_MODE_TC16X3: the unit rows of the MN-major GEMMs are stored times 2^8

// ---- workspace layouts of the infonce calls -----------------------------------------------------
// Forward and backward own separate caller-allocated workspaces (bsc_infonce_plan reports both sizes): the forward
// of an evaluation-only call never pays for the backward's buffers.
struct InfoNceFwdWs {
  size_t rowpart;   // float [pairs][n_jchunk][2][n_local_pad]   row sum-exp partials (2 epilogue warp groups)
  size_t rowpos;    // float [pairs][n_jchunk][2][n_local_pad]   positive-pair partials
  size_t colpart;   // float [pairs][n_iblk][N_pad]
  size_t jrange;    // int2  [n_jtile]       label-class range of every column tile
  size_t total;
  int n_iblk, n_jtile, n_jchunk, jtiles_per_chunk;
  int64_t n_local_pad, n_global_pad;
};

struct InfoNceBwdWs {
  size_t jrange;    // int2  [n_jtile]
  size_t dxraw;     // float [3 K-major jobs... see tc_bwd3.cu][kparts][n_local][D]   local partial gradients
  size_t xunit;     // half  [3][n_local][D]  unit rows of the LOCAL rows (B operand of the MN-major jobs)
  size_t gimg;      // half  weight images [blocks][n_iblk_pad][n_j64][128][64]
  size_t gscratch;  // float [g_rows][N]     fp32 (SIMT) mode only
  size_t total;
  int n_iblk, n_jtile, n_jchunk, jtiles_per_chunk;  // tiling of the weights sweep (the forward's)
  int kparts;       // K split of the K-major gradient GEMMs (BLOCKS / SCATTER) or column split of the recomputing sweep
  int mode;         // resolved BSC_BWD_* (== p->bwd_mode after validation)
  int64_t g_rows;
};

InfoNceFwdWs infonce_fwd_layout(const bsc_infonce_desc* p);
InfoNceBwdWs infonce_bwd_layout(const bsc_infonce_desc* p);

// Once-per-(kernel, device) guard: function attributes (dynamic shared-memory opt-in) are per device, so an index that
// spans several GPUs in one process must set them on every device it launches on.
struct PerDeviceOnce {
  unsigned long long mask[2] = {0ull, 0ull};  // device ordinals 0..127
  // returns true the first time it is asked about the current device (racing threads may both see true: harmless)
  bool first(int dev) {
    if (dev < 0 || dev >= 128) return true;
    const unsigned long long bit = 1ull << (dev & 63);
    unsigned long long& m = mask[dev >> 6];
    if (m & bit) return false;
    m |= bit;
    return true;
  }
};
int current_device();
int device_sm_count();  // SM count of the CURRENT device (cached per device)

#define BSC_SET_SMEM_ONCE(kernel, bytes)                                                                   \
  do {                                                                                                     \
    static ::bsc::PerDeviceOnce once__;                                                                    \
    if (once__.first(::bsc::current_device()))                                                             \
      BSC_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(bytes)));   \
  } while (0)

// ---- argument blocks of the O(N*D) kernels in prep.cu ----------------------------------------------
// Raw gradient sources of one (modality, partner) slot: nparts partial sums [n_local][d], part q at src + q * stride,
// summed in the order q = 0, 1, ... (fixed: the result does not depend on scheduling or on which rank wrote what).
struct CombineSrc {
  const float* src;
  int64_t stride;   // elements between parts
  int nparts;
  int gs_index;     // gsbuf[gs_index] = 1 / gs of the slot's sums
};

struct CombineArgs {
  const void* x[3];      // original local rows (un-normalised), null when the modality is absent
  void* grad[3];         // dense [n_local][d]
  CombineSrc src[3][2];
  int64_t ldx[3];        // row stride of x[m] in elements
  const float* gsbuf;
  const float* grad_out;
  int64_t n_local, n_global;
  int d, n_slots;
  unsigned present_mask;
  float scale;
};

// ---- sub-launchers implemented in the .cu files ---------------------------------------------------
int simt_fwd_stats(const bsc_infonce_desc* p, const InfoNceFwdWs& L, char* ws, float* stats, cudaStream_t st);
// aux = uv[6][N] (r_i / sumexp per ordered pair) followed by gsbuf[32], written by bsc_infonce_loss
int simt_bwd_raw(const bsc_infonce_desc* p, const InfoNceBwdWs& L, char* ws, const float* aux, cudaStream_t st);
int tc_fwd_stats(const bsc_infonce_desc* p, const InfoNceFwdWs& L, char* ws, float* stats,
                 const bsc_infonce_exchange* ex, cudaStream_t st);
// recomputing CTA-pair sweep (BSC_BWD_RECOMPUTE), d in {256, 512, 768}
bool tc_bwd_duo_covers(const bsc_infonce_desc* p);
int tc_bwd_duo(const bsc_infonce_desc* p, const InfoNceBwdWs& L, char* ws, const float* aux, cudaStream_t st);
int tc_bwd_duo_jsplit(const bsc_infonce_desc* p);
// stored-weight backward (BSC_BWD_BLOCKS / BSC_BWD_SCATTER): weight images, then the gradient GEMMs
size_t tc_bwd_gimg_bytes(const bsc_infonce_desc* p);
void tc_bwd_matrix_dims(const bsc_infonce_desc* d, int* n_iblk_pad, int* n_j64);
int tc_bwd_gemm_kparts(const bsc_infonce_desc* d);
int tc_bwd_mn_kparts(const bsc_infonce_desc* d);  // K parts of the scattered GEMMs = exchange slots per source rank
int tc_bwd_weights(const bsc_infonce_desc* d, const InfoNceBwdWs& L, char* ws, const float* aux, int* n_pairs,
                   int* pairs_a, int* pairs_b, cudaStream_t st);
struct ScatterDst {
  float* peer[kMaxPeers];  // exchange buffer of every rank (index = rank), [world src][3][n_local][d] float
  int world, rank;
};
int tc_bwd_gemms(const bsc_infonce_desc* d, const InfoNceBwdWs& L, char* ws, const float* aux, int n_pairs,
                 const int* pairs_a, const int* pairs_b, const ScatterDst& dst, cudaStream_t st);
int rows_unit_f16_launch3(const void* const x16[3], const float* const inv[3], int64_t row0, int64_t n, int d,
                          void* const out[3], cudaStream_t st);
// BSC_MODE_TC16X3: hl16[m] = [rows][2 d] (hi | lo); out[m] = [n][2 d] (hi | lo) of kX3UnitScale * unit rows
int rows_unit_f16x3_launch3(const void* const hl16[3], const float* const inv[3], int64_t row0, int64_t n, int d,
                            void* const out[3], cudaStream_t st);
struct ReduceJob {
  const float* part;
  int nparts;
  int64_t part_stride, n;
  float* out;
};
struct ReduceJobs {
  ReduceJob job[6];
  int count;
  // positive-pair term: pos_out[b] = scale * sum over the elements (p, i) dealt to part b, b < kPosParts (double)
  const float* pos_part;
  int pos_nparts;
  int64_t pos_stride, pos_n;
  float pos_scale;
  double* pos_out;
  // copies of everything written above into the records of other ranks: an address inside [local_base, local_base + len)
  // is mirrored at the same offset from peer_base[r] (r < n_peers) or, when mc_base is set, by ONE multicast store
  const float* local_base;
  float* peer_base[kMaxPeers];
  float* mc_base;
  int n_peers;
};
int reduce_parts_batch_launch(const ReduceJobs& jobs, cudaStream_t st);
struct StatsFinishArgs {
  const float* slots;   // [world][record_len] per-rank records
  int world;
  int64_t n, record_len;
  const float* r;       // [n]
  const float* rmax;    // [1]
  const float* inv[3];  // TC16: inverse row norms (null in fp32 mode)
  unsigned present_mask;
  int use_gs;
  float* stats;         // [record_len] combined record
  float* uv;            // [6][n]
  float* gsbuf;         // [32]
  float* loss;          // [1]
  char* scratch;        // stats_finish_scratch_bytes(), zeroed once by the caller, left zero by every call
};
int stats_finish_launch(const StatsFinishArgs& a, cudaStream_t st);
size_t stats_finish_scratch_bytes();

int simt_bank_search(const float* q, int64_t nq, const float* bank, int64_t nb, int d, int k, int64_t base_id,
                     float* out_s, int64_t* out_i, char* ws, size_t ws_bytes, cudaStream_t st);
int tc_bank_search(const void* q, int64_t nq, const void* bank, int64_t nb, int d, int k, int64_t base_id,
                   float* out_s, int64_t* out_i, char* ws, size_t ws_bytes, cudaStream_t st);
size_t simt_bank_ws(int64_t nq, int64_t nb, int d, int k);
size_t tc_bank_ws(int64_t nq, int64_t nb, int d, int k);

// tile shapes of the tensor-core kernels (host needs them for the workspace layout)
constexpr int kSimNG = 3;    // epilogue warp groups of the similarity sweep (partial-sum slabs per column chunk)
constexpr int kTcBM = 128;   // rows per CTA tile (TMEM lanes)
constexpr int kTcBNF = 256;  // forward: columns per tile
constexpr int kTcBNB = 128;  // backward: columns per tile
constexpr int kSimtT = 64;   // SIMT tile edge

// timing slots (bench instrumentation)
void timing_begin(int which, cudaStream_t st);
void timing_end(int which, cudaStream_t st);

}  // namespace bsc
