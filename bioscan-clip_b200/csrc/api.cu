// C ABI of libbsc_hot.so (see include/bsc_hot.h) and the host-side glue: argument checks, workspace
// layout, tensor-map encoding, launch counting and the optional CUDA-event timing used by bench.py.
#include <atomic>
#include <cstring>
#include <mutex>

#include "tc_common.cuh"

namespace bsc {

// launchers from prep.cu
int l2norm_launch(const void* x, int x_dtype, int64_t n, int d, int64_t ldx, void* y16, int y16_dtype, float* y32,
                  float* inv_norm, int zero_rule, cudaStream_t st);
int rows_to_f16_launch(const void* x, int x_dtype, int64_t n, int d, int64_t ldx, void* y16, float* inv,
                       cudaStream_t st);
int rows_to_f16x3_launch(const void* x, int x_dtype, int64_t n, int d, int64_t ldx, void* hl, float* inv,
                         cudaStream_t st);
size_t label_stats_ws_bytes(int64_t n);
int tc_bank_screen_parts(int64_t nq, int64_t nb, int min_parts);
int tc_bank_screen(const void* q, int64_t nq, const void* bank, int64_t nb, int d, int64_t base_id, int min_parts,
                   float* cand_s, int64_t* cand_i, cudaStream_t st);
int rescore_launch(const float* q, int64_t nq, const float* bank, int64_t nb, int d, int64_t base_id, const float* cs,
                   const int64_t* ci, int parts, int k, float eps_scale, float* os, int64_t* oi, int32_t* unsure,
                   int32_t* n_unsure, cudaStream_t st);
int taxo_launch(const int64_t* indices, int64_t nq, int k, const int32_t* key_ids, int64_t nb, const int32_t* gt_ids,
                int levels, const int32_t* k_list, int nk, const int64_t* class_off, int64_t total_classes,
                int32_t* pred_ids, uint8_t* first_hit, int64_t* micro, int32_t* pred_counts, int32_t* gt_counts,
                cudaStream_t st);
void gather_layout(int64_t n_global, int d, int world, size_t off_xh[3], size_t off_inv[3], size_t* off_labels,
                   size_t* off_stats, size_t* off_flags, size_t* total);
int gather_blocks_per_dest(int64_t n_local);
int gather_ring_launch(void* const* peers, int world, int rank, int64_t n_local, int64_t row0, int64_t n_global, int d,
                       unsigned present_mask, cudaStream_t st);
int gather_labels_launch(const int64_t* labels, int64_t n_local, int64_t row0, int64_t n_global, int d,
                         void* const* peers, void* multicast, int world, cudaStream_t st);
int gather_push_launch(const void* const x[3], int x_dtype, const int64_t ldx[3], const int64_t* labels,
                       int64_t n_local, int64_t row0, int64_t n_global, int d, void* const* peers, void* multicast,
                       int world, int rank, int ordered, uint64_t epoch, cudaStream_t st);
int label_stats_launch(const int64_t* labels, int64_t n, float* r, int32_t* rep, int32_t* next, float* rmax, char* ws,
                       cudaStream_t st);
int rep_range_launch(const int32_t* rep, int64_t n, int tile, int2* out, cudaStream_t st);
int combine_launch(const CombineArgs& a, int x_dtype, cudaStream_t st);
int topk_merge_launch(const float* cs, const int64_t* ci, int parts, int64_t nq, int k, float* os, int64_t* oi,
                      cudaStream_t st);
void tc_fwd_tiling(int64_t n_local, int64_t n_global, int n_pairs, int world, int* n_iblk, int* n_jtile, int* jpc,
                   int* n_jchunk);
bool tc_fwd_source_aligned(int64_t n_local, int64_t n_global, int world);

// ---- thread-local state -----------------------------------------------------------------------
// The error text is per thread.  Launch counter and timing slots are per process: torch runs the backward of
// an autograd function on its own thread, and bench.py reads both from the main thread.
static thread_local char g_err[512] = "";
static std::atomic<int64_t> g_launches{0};
static std::atomic<int> g_bwd_mode{-1};
static std::atomic<int> g_timing{0};
static std::mutex g_ev_mutex;
static cudaEvent_t g_ev[4][2] = {{nullptr, nullptr}, {nullptr, nullptr}, {nullptr, nullptr}, {nullptr, nullptr}};
static bool g_ev_valid[4] = {false, false, false, false};

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int cuda_fail(cudaError_t e, const char* what) {
  set_error("CUDA error %d (%s) in %s", (int)e, cudaGetErrorString(e), what);
  return BSC_ECUDA;
}

void count_launch(int n) { g_launches += n; }

int current_device() {
  int dev = 0;
  cudaGetDevice(&dev);
  return dev;
}

int device_sm_count() {
  static std::atomic<int> cache[128];
  const int dev = current_device();
  if (dev >= 0 && dev < 128) {
    const int c = cache[dev].load();
    if (c > 0) return c;
  }
  int n = 0;
  cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
  if (n <= 0) n = 148;
  if (dev >= 0 && dev < 128) cache[dev].store(n);
  return n;
}

void timing_begin(int which, cudaStream_t st) {
  if (!g_timing) return;
  std::lock_guard<std::mutex> lock(g_ev_mutex);
  if (!g_ev[which][0]) {
    cudaEventCreate(&g_ev[which][0]);
    cudaEventCreate(&g_ev[which][1]);
  }
  cudaEventRecord(g_ev[which][0], st);
}
void timing_end(int which, cudaStream_t st) {
  if (!g_timing) return;
  std::lock_guard<std::mutex> lock(g_ev_mutex);
  cudaEventRecord(g_ev[which][1], st);
  g_ev_valid[which] = true;
}

// ---- TMA tensor maps ------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encoder() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* sym = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &sym, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = (EncodeTiledFn)sym;
  });
  return fn;
}

// cuTensorMapEncodeTiled is a DRIVER call: it needs the device's primary context bound to the calling thread, which the
// runtime does lazily.  torch runs the backward of an autograd function on its own worker thread; when every buffer
// comes out of the caching allocator no runtime call precedes the first tensor map there (CUDA_ERROR_INVALID_CONTEXT,
// seen in the 8-GPU train step).  One cudaFree(0) per (thread, device) binds it.
static void bind_context() {
  static thread_local int bound_dev = -1;
  const int dev = current_device();
  if (dev != bound_dev) {
    cudaFree(nullptr);
    bound_dev = dev;
  }
}

int make_tmap_2d(CUtensorMap* out, const void* base, int dtype, int64_t rows, int64_t cols, int box_rows) {
  return make_tmap_2d_ld(out, base, dtype, rows, cols, cols, box_rows);
}

// [rows][cols] 16-bit elements with a row stride of ld elements (a column slice of a wider matrix), box {64, box_rows}
int make_tmap_2d_ld(CUtensorMap* out, const void* base, int dtype, int64_t rows, int64_t cols, int64_t ld,
                    int box_rows) {
  bind_context();
  EncodeTiledFn enc = get_encoder();
  BSC_REQUIRE(enc != nullptr, BSC_ECUDA, "cuTensorMapEncodeTiled is not available from this driver");
  BSC_REQUIRE(((uintptr_t)base & 15) == 0, BSC_EINVAL, "TMA operand base must be 16-byte aligned");
  BSC_REQUIRE(rows > 0 && cols > 0 && cols % 8 == 0 && ld >= cols && ld % 8 == 0, BSC_EINVAL,
              "TMA operand needs rows > 0, cols %% 8 == 0 and a row stride that is a multiple of 8 elements");
  cuuint64_t gdim[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t gstride[1] = {(cuuint64_t)ld * 2};
  cuuint32_t box[2] = {64u, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1u, 1u};
  CUtensorMapDataType dt = dtype == BSC_BF16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16;
  CUresult r = enc(out, dt, 2, const_cast<void*>(base), gdim, gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  BSC_REQUIRE(r == CUDA_SUCCESS, BSC_ECUDA, "cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
  return BSC_OK;
}

int make_tmap_raw128(CUtensorMap* out, const void* base, int64_t rows, int box_rows) {
  bind_context();
  EncodeTiledFn enc = get_encoder();
  BSC_REQUIRE(enc != nullptr, BSC_ECUDA, "cuTensorMapEncodeTiled is not available from this driver");
  BSC_REQUIRE(((uintptr_t)base & 127) == 0 && rows > 0, BSC_EINVAL, "raw TMA operand must be 128-byte aligned");
  cuuint64_t gdim[2] = {64u, (cuuint64_t)rows};
  cuuint64_t gstride[1] = {128u};
  cuuint32_t box[2] = {64u, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1u, 1u};
  CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, const_cast<void*>(base), gdim, gstride, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  BSC_REQUIRE(r == CUDA_SUCCESS, BSC_ECUDA, "cuTensorMapEncodeTiled (raw) failed with CUresult %d", (int)r);
  return BSC_OK;
}

// ---- workspace layout ---------------------------------------------------------------------------
static int present_count(const bsc_infonce_desc* p, unsigned* mask) {
  int m = 0;
  unsigned k = 0;
  for (int a = 0; a < 3; ++a)
    if (p->xh[a]) {
      ++m;
      k |= 1u << a;
    }
  if (mask) *mask = k;
  return m;
}

static size_t take(size_t& off, size_t bytes) {
  const size_t o = off;
  off += align_up(bytes, 256);
  return o;
}

InfoNceFwdWs infonce_fwd_layout(const bsc_infonce_desc* p) {
  InfoNceFwdWs L{};
  const int64_t N = p->n_global, n = p->n_local;
  const int m = present_count(p, nullptr);
  const int n_pairs = m * (m - 1) / 2;
  if (mode_is_tc(p->mode)) {
    tc_fwd_tiling(n, N, n_pairs, p->world, &L.n_iblk, &L.n_jtile, &L.jtiles_per_chunk, &L.n_jchunk);
  } else {
    L.n_iblk = (int)((n + kSimtT - 1) / kSimtT);
    L.n_jtile = (int)((N + kSimtT - 1) / kSimtT);
    L.jtiles_per_chunk = 1;
    L.n_jchunk = L.n_jtile;
  }
  L.n_local_pad = (int64_t)align_up((size_t)(n > 0 ? n : 1), 64);
  L.n_global_pad = (int64_t)align_up((size_t)(N > 0 ? N : 1), 64);
  size_t off = 0;
  // the tensor-core forward sweep keeps one partial per epilogue warp group (tc_sim.cu NG)
  L.rowpart = take(off, (size_t)3 * kSimNG * (L.n_jchunk > 0 ? L.n_jchunk : 1) * L.n_local_pad * sizeof(float));
  L.rowpos = take(off, (size_t)3 * kSimNG * (L.n_jchunk > 0 ? L.n_jchunk : 1) * L.n_local_pad * sizeof(float));
  L.jrange = take(off, (size_t)(L.n_jtile > 0 ? L.n_jtile : 1) * 2 * sizeof(int));
  L.colpart = take(off, (size_t)3 * (L.n_iblk > 0 ? L.n_iblk : 1) * L.n_global_pad * sizeof(float));
  L.total = off;
  return L;
}

// number of raw-gradient slabs [n_local][d] the strategy keeps in the workspace (before the K / column split)
static int bwd_local_slots(const bsc_infonce_desc* p) {
  const int m = present_count(p, nullptr);
  // SCATTER: one K-major job per present pair; the others address slabs as (modality * 2 + partner slot), modality 0..2
  return (mode_is_tc(p->mode) && p->bwd_mode == BSC_BWD_SCATTER) ? m * (m - 1) / 2 : 6;
}

InfoNceBwdWs infonce_bwd_layout(const bsc_infonce_desc* p) {
  InfoNceBwdWs L{};
  const int64_t N = p->n_global, n = p->n_local;
  const int D = p->d;
  const int m = present_count(p, nullptr);
  const int n_pairs = m * (m - 1) / 2;
  L.mode = mode_is_tc(p->mode) ? p->bwd_mode : BSC_BWD_RECOMPUTE;
  const bool stored = mode_is_tc(p->mode) && (L.mode == BSC_BWD_BLOCKS || L.mode == BSC_BWD_SCATTER);
  if (mode_is_tc(p->mode))
    tc_fwd_tiling(n, N, n_pairs, p->world, &L.n_iblk, &L.n_jtile, &L.jtiles_per_chunk, &L.n_jchunk);
  L.g_rows = n < 2048 ? (n > 0 ? n : 1) : 2048;
  L.kparts = 1;
  if (mode_is_tc(p->mode)) L.kparts = stored ? tc_bwd_gemm_kparts(p) : tc_bwd_duo_jsplit(p);
  size_t off = 0;
  L.jrange = take(off, (size_t)((N + kTcBNF - 1) / kTcBNF + 1) * 2 * sizeof(int));
  L.dxraw = take(off, (size_t)bwd_local_slots(p) * L.kparts * (n > 0 ? n : 1) * D * sizeof(float));
  if (stored) {
    L.xunit = take(off, (size_t)3 * (n > 0 ? n : 1) * D * 2 * mode_parts(p->mode));
    L.gimg = take(off, tc_bwd_gimg_bytes(p));
  } else {
    L.xunit = L.gimg = off;
  }
  L.gscratch = (p->mode == BSC_MODE_F32) ? take(off, (size_t)L.g_rows * N * sizeof(float)) : off;
  L.total = off;
  return L;
}

static size_t exchange_bytes_of(const bsc_infonce_desc* p) {
  if (!mode_is_tc(p->mode) || p->bwd_mode != BSC_BWD_SCATTER) return 0;
  const int m = present_count(p, nullptr);
  // one slot per (source rank, K part of the scattered GEMMs: 1 except in BSC_MODE_TC16X3) and unordered pair
  return (size_t)p->world * (size_t)tc_bwd_mn_kparts(p) * (size_t)(m * (m - 1) / 2) *
         (size_t)(p->n_local > 0 ? p->n_local : 1) * p->d * sizeof(float);
}

static int check_desc(const bsc_infonce_desc* p, const char* who) {
  BSC_REQUIRE(p != nullptr, BSC_EINVAL, "%s: null descriptor", who);
  BSC_REQUIRE(p->mode == BSC_MODE_F32 || mode_is_tc(p->mode), BSC_EINVAL, "%s: unknown mode %d", who, p->mode);
  BSC_REQUIRE(p->n_global >= 1 && p->n_local >= 0 && p->row0 >= 0 && p->row0 + p->n_local <= p->n_global, BSC_EINVAL,
              "%s: bad row partition (n_global=%lld n_local=%lld row0=%lld)", who, (long long)p->n_global,
              (long long)p->n_local, (long long)p->row0);
  BSC_REQUIRE(p->d >= 1, BSC_EINVAL, "%s: d must be positive", who);
  BSC_REQUIRE(present_count(p, nullptr) >= 2, BSC_EINVAL, "%s: Too less element for calculating the contrastive loss.",
              who);
  BSC_REQUIRE(p->r && p->rep && p->next && p->rmax, BSC_EINVAL, "%s: label statistics pointers are null", who);
  BSC_REQUIRE(p->world >= 1 && p->world <= kMaxPeers && p->rank >= 0 && p->rank < p->world, BSC_EINVAL,
              "%s: bad rank %d / world %d (1..%d ranks)", who, p->rank, p->world, kMaxPeers);
  if (mode_is_tc(p->mode)) {
    BSC_REQUIRE(p->d % 128 == 0, BSC_EUNSUPPORTED,
                "%s: the tensor-core modes need d %% 128 == 0 (got %d); use BSC_MODE_F32", who, p->d);
    BSC_REQUIRE(p->x16_dtype == BSC_F16, BSC_EUNSUPPORTED,
                "%s: TC16 mode takes float16 rows prepared by bsc_rows_to_f16 (exact for bfloat16 values); "
                "tcgen05 kind::f16 cannot mix the fp16 weight tile with bfloat16 rows (x16_dtype=%d)",
                who, p->x16_dtype);
    for (int a = 0; a < 3; ++a)
      BSC_REQUIRE(!p->xh[a] || p->inv[a], BSC_EINVAL, "%s: TC16 mode needs the inverse row norms of modality %d", who,
                  a);
  }
  return BSC_OK;
}

}  // namespace bsc

using namespace bsc;

extern "C" {

int bsc_version(void) { return BSC_VERSION; }
const char* bsc_last_error(void) { return g_err; }

int bsc_device_check(void) {
  int dev = 0;
  BSC_CUDA(cudaGetDevice(&dev));
  cudaDeviceProp prop;
  BSC_CUDA(cudaGetDeviceProperties(&prop, dev));
  BSC_REQUIRE(prop.major == 10, BSC_EUNSUPPORTED, "device %d is sm_%d%d; libbsc_hot.so is built for sm_100a only",
              dev, prop.major, prop.minor);
  BSC_REQUIRE(get_encoder() != nullptr, BSC_ECUDA, "cuTensorMapEncodeTiled is not available from this driver");
  return BSC_OK;
}

int bsc_set_timing(int enabled) {
  g_timing = enabled ? 1 : 0;
  return BSC_OK;
}

float bsc_last_kernel_ms(int which) {
  std::lock_guard<std::mutex> lock(g_ev_mutex);
  if (which < 0 || which > 3 || !g_ev_valid[which]) return -1.0f;
  if (cudaEventSynchronize(g_ev[which][1]) != cudaSuccess) return -1.0f;
  float ms = -1.0f;
  if (cudaEventElapsedTime(&ms, g_ev[which][0], g_ev[which][1]) != cudaSuccess) return -1.0f;
  return ms;
}

int bsc_last_bwd_mode(void) { return g_bwd_mode.load(); }

int64_t bsc_launch_count(int reset) {
  return reset ? g_launches.exchange(0) : g_launches.load();
}

int bsc_l2norm_fwd(const void* x, int x_dtype, int64_t n, int d, int64_t ldx, void* y16, int y16_dtype, float* y32,
                   float* inv_norm, int zero_rule, void* stream) {
  BSC_REQUIRE(x != nullptr || n == 0, BSC_EINVAL, "bsc_l2norm_fwd: x is null");
  BSC_REQUIRE(n >= 0 && d >= 1 && ldx >= d, BSC_EINVAL, "bsc_l2norm_fwd: bad shape n=%lld d=%d ldx=%lld", (long long)n,
              d, (long long)ldx);
  BSC_REQUIRE(zero_rule == BSC_NORM_TORCH || zero_rule == BSC_NORM_SKLEARN, BSC_EINVAL,
              "bsc_l2norm_fwd: unknown zero_rule %d", zero_rule);
  return l2norm_launch(x, x_dtype, n, d, ldx, y16, y16_dtype, y32, inv_norm, zero_rule, (cudaStream_t)stream);
}

int bsc_rows_to_f16(const void* x, int x_dtype, int64_t n, int d, int64_t ldx, void* y16, float* inv, void* stream) {
  BSC_REQUIRE((x && y16 && inv) || n == 0, BSC_EINVAL, "bsc_rows_to_f16: null pointer");
  BSC_REQUIRE(n >= 0 && d >= 1 && ldx >= d, BSC_EINVAL, "bsc_rows_to_f16: bad shape n=%lld d=%d ldx=%lld", (long long)n,
              d, (long long)ldx);
  return rows_to_f16_launch(x, x_dtype, n, d, ldx, y16, inv, (cudaStream_t)stream);
}

int bsc_rows_to_f16x3(const void* x, int x_dtype, int64_t n, int d, int64_t ldx, void* hl, float* inv, void* stream) {
  BSC_REQUIRE((x && hl && inv) || n == 0, BSC_EINVAL, "bsc_rows_to_f16x3: null pointer");
  BSC_REQUIRE(n >= 0 && d >= 1 && ldx >= d, BSC_EINVAL, "bsc_rows_to_f16x3: bad shape n=%lld d=%d ldx=%lld",
              (long long)n, d, (long long)ldx);
  return rows_to_f16x3_launch(x, x_dtype, n, d, ldx, hl, inv, (cudaStream_t)stream);
}

int bsc_gather_layout(int64_t n_global, int d, int world, size_t off_xh[3], size_t off_inv[3], size_t* off_labels,
                      size_t* off_stats, size_t* off_flags, size_t* total_bytes) {
  BSC_REQUIRE(off_xh && off_inv && off_labels && off_stats && off_flags && total_bytes, BSC_EINVAL,
              "bsc_gather_layout: null pointer");
  BSC_REQUIRE(n_global >= 0 && d >= 1 && world >= 1 && world <= kMaxPeers, BSC_EINVAL, "bsc_gather_layout: bad shape");
  gather_layout(n_global, d, world, off_xh, off_inv, off_labels, off_stats, off_flags, total_bytes);
  return BSC_OK;
}

static int check_peers(void* const* peer_buffers, int world, const char* who) {
  BSC_REQUIRE(peer_buffers != nullptr, BSC_EINVAL, "%s: null peer list", who);
  BSC_REQUIRE(world >= 1 && world <= kMaxPeers, BSC_EUNSUPPORTED, "%s: world=%d (1..%d supported)", who, world, kMaxPeers);
  for (int r = 0; r < world; ++r) BSC_REQUIRE(peer_buffers[r], BSC_EINVAL, "%s: peer %d is null", who, r);
  return BSC_OK;
}

int bsc_gather_push_labels(const int64_t* labels, int64_t n_local, int64_t row0, int64_t n_global, int d,
                           void* const* peer_buffers, void* multicast_buffer, int world, void* stream) {
  int rc = check_peers(peer_buffers, world, "bsc_gather_push_labels");
  if (rc) return rc;
  BSC_REQUIRE(labels || n_local == 0, BSC_EINVAL, "bsc_gather_push_labels: null labels");
  BSC_REQUIRE(n_local >= 0 && row0 >= 0 && row0 + n_local <= n_global && d >= 1, BSC_EINVAL,
              "bsc_gather_push_labels: bad shape");
  return gather_labels_launch(labels, n_local, row0, n_global, d, peer_buffers, multicast_buffer, world,
                              (cudaStream_t)stream);
}

int bsc_gather_push(const void* const x[3], int x_dtype, const int64_t ldx[3], const int64_t* labels, int64_t n_local,
                    int64_t row0, int64_t n_global, int d, void* const* peer_buffers, void* multicast_buffer, int world,
                    int rank, int ordered, uint64_t epoch, void* stream) {
  BSC_REQUIRE(x && ldx, BSC_EINVAL, "bsc_gather_push: null pointer");
  int rc = check_peers(peer_buffers, world, "bsc_gather_push");
  if (rc) return rc;
  BSC_REQUIRE(rank >= 0 && rank < world, BSC_EINVAL, "bsc_gather_push: rank %d of %d", rank, world);
  BSC_REQUIRE(d >= 8 && d % 8 == 0, BSC_EUNSUPPORTED, "bsc_gather_push: d %% 8 != 0 (got %d)", d);
  BSC_REQUIRE(n_local >= 0 && row0 >= 0 && row0 + n_local <= n_global, BSC_EINVAL,
              "bsc_gather_push: rows [%lld, %lld) outside the global batch of %lld", (long long)row0,
              (long long)(row0 + n_local), (long long)n_global);
  for (int m = 0; m < 3; ++m)
    BSC_REQUIRE(!x[m] || ldx[m] >= d, BSC_EINVAL, "bsc_gather_push: ldx[%d] < d", m);
  BSC_REQUIRE(multicast_buffer == nullptr || ((uintptr_t)multicast_buffer & 15) == 0, BSC_EINVAL,
              "bsc_gather_push: multicast address is not 16-byte aligned");
  BSC_REQUIRE(!ordered || epoch > 0, BSC_EINVAL, "bsc_gather_push: ordered pushes are numbered from 1");
  return gather_push_launch(x, x_dtype, ldx, labels, n_local, row0, n_global, d, peer_buffers, multicast_buffer, world,
                            rank, ordered ? 1 : 0, epoch, (cudaStream_t)stream);
}

int bsc_gather_ring(void* const* peer_buffers, int world, int rank, int64_t n_local, int64_t row0, int64_t n_global,
                    int d, unsigned present_mask, void* stream) {
  int rc = check_peers(peer_buffers, world, "bsc_gather_ring");
  if (rc) return rc;
  BSC_REQUIRE(rank >= 0 && rank < world, BSC_EINVAL, "bsc_gather_ring: rank %d of %d", rank, world);
  BSC_REQUIRE(n_local >= 0 && row0 >= 0 && row0 + n_local <= n_global && d >= 8 && d % 8 == 0, BSC_EINVAL,
              "bsc_gather_ring: bad shape");
  return gather_ring_launch(peer_buffers, world, rank, n_local, row0, n_global, d, present_mask, (cudaStream_t)stream);
}

int bsc_taxonomy_eval(const int64_t* indices, int64_t nq, int k, const int32_t* key_ids, int64_t nb,
                      const int32_t* gt_ids, int levels, const int32_t* k_list, int nk, const int64_t* class_off,
                      int64_t total_classes, int32_t* pred_ids, uint8_t* first_hit, int64_t* micro,
                      int32_t* pred_counts, int32_t* gt_counts, void* stream) {
  BSC_REQUIRE(nq >= 0 && nb >= 0 && total_classes >= 0, BSC_EINVAL, "bsc_taxonomy_eval: negative size");
  BSC_REQUIRE((indices && gt_ids) || nq == 0, BSC_EINVAL, "bsc_taxonomy_eval: null query arrays");
  BSC_REQUIRE(key_ids || nb == 0, BSC_EINVAL, "bsc_taxonomy_eval: null key_ids");
  BSC_REQUIRE(k_list && class_off && micro && pred_counts && gt_counts, BSC_EINVAL, "bsc_taxonomy_eval: null pointer");
  return taxo_launch(indices, nq, k, key_ids, nb, gt_ids, levels, k_list, nk, class_off, total_classes, pred_ids,
                     first_hit, micro, pred_counts, gt_counts, (cudaStream_t)stream);
}

int bsc_label_stats_workspace_bytes(int64_t n, size_t* bytes) {
  BSC_REQUIRE(bytes && n >= 0 && n < (1ll << 30), BSC_EINVAL, "bsc_label_stats_workspace_bytes: bad arguments");
  *bytes = label_stats_ws_bytes(n);
  return BSC_OK;
}

int bsc_label_stats(const int64_t* labels, int64_t n, float* r, int32_t* rep, int32_t* next, float* rmax,
                    void* workspace, size_t workspace_bytes, void* stream) {
  BSC_REQUIRE(labels && r && rep && next && rmax, BSC_EINVAL, "bsc_label_stats: null pointer");
  BSC_REQUIRE(n >= 0 && n < (1ll << 30), BSC_EINVAL, "bsc_label_stats: n out of range");
  BSC_REQUIRE(workspace && workspace_bytes >= label_stats_ws_bytes(n), BSC_EWORKSPACE,
              "bsc_label_stats: workspace too small");
  return label_stats_launch(labels, n, r, rep, next, rmax, (char*)workspace, (cudaStream_t)stream);
}

int64_t bsc_infonce_stats_len(int64_t n_global) { return stats_record_len(n_global); }
int64_t bsc_infonce_aux_len(int64_t n_global) { return 6 * n_global + 32; }

// Is `mode` a strategy this problem can run?  (message in g_err when not)
static int bwd_mode_ok(const bsc_infonce_desc* p, int mode, int peer_memory, const char* who) {
  if (p->mode == BSC_MODE_F32) {
    BSC_REQUIRE(mode == BSC_BWD_RECOMPUTE, BSC_EUNSUPPORTED, "%s: BSC_MODE_F32 has one backward (BSC_BWD_RECOMPUTE)", who);
    return BSC_OK;
  }
  switch (mode) {
    case BSC_BWD_RECOMPUTE:
      BSC_REQUIRE(p->mode != BSC_MODE_TC16X3, BSC_EUNSUPPORTED,
                  "%s: BSC_MODE_TC16X3 runs the stored-weight backwards only (BSC_BWD_BLOCKS / BSC_BWD_SCATTER)", who);
      BSC_REQUIRE(tc_bwd_duo_covers(p), BSC_EUNSUPPORTED,
                  "%s: the recomputing sweep covers d in {256, 512, 768} only (got %d)", who, p->d);
      return BSC_OK;
    case BSC_BWD_BLOCKS:
      return BSC_OK;
    case BSC_BWD_SCATTER:
      BSC_REQUIRE(p->n_global == (int64_t)p->world * p->n_local && p->row0 == (int64_t)p->rank * p->n_local,
                  BSC_EUNSUPPORTED, "%s: BSC_BWD_SCATTER needs equal row blocks (n_global = world * n_local, row0 = "
                  "rank * n_local)", who);
      BSC_REQUIRE(p->world == 1 || peer_memory, BSC_EUNSUPPORTED,
                  "%s: BSC_BWD_SCATTER over %d ranks needs peer-mapped exchange buffers", who, p->world);
      return BSC_OK;
  }
  set_error("%s: unknown backward mode %d", who, mode);
  return BSC_EINVAL;
}

int bsc_infonce_plan(const bsc_infonce_desc* p_in, int requested, int peer_memory, size_t budget,
                     bsc_infonce_plan_t* plan) {
  BSC_REQUIRE(p_in != nullptr && plan != nullptr, BSC_EINVAL, "bsc_infonce_plan: null pointer");
  bsc_infonce_desc p = *p_in;
  p.bwd_mode = BSC_BWD_RECOMPUTE;
  int rc = check_desc(&p, "bsc_infonce_plan");
  if (rc) return rc;
  int mode;
  if (requested != BSC_BWD_AUTO) {
    rc = bwd_mode_ok(&p, requested, peer_memory, "bsc_infonce_plan");
    if (rc) return rc;
    mode = requested;
  } else if (p.mode == BSC_MODE_F32) {
    mode = BSC_BWD_RECOMPUTE;
  } else {
    const bool scatter_ok = p.n_global == (int64_t)p.world * p.n_local && p.row0 == (int64_t)p.rank * p.n_local &&
                            (p.world == 1 || peer_memory);
    mode = scatter_ok ? BSC_BWD_SCATTER : BSC_BWD_BLOCKS;
    if (budget > 0 && tc_bwd_duo_covers(&p)) {
      p.bwd_mode = mode;
      if (infonce_bwd_layout(&p).total + exchange_bytes_of(&p) > budget) mode = BSC_BWD_RECOMPUTE;
    }
  }
  p.bwd_mode = mode;
  plan->bwd_mode = mode;
  plan->reserved = 0;
  plan->fwd_workspace_bytes = infonce_fwd_layout(&p).total;
  plan->bwd_workspace_bytes = infonce_bwd_layout(&p).total;
  plan->exchange_bytes = exchange_bytes_of(&p);
  if (plan->fwd_workspace_bytes == 0) plan->fwd_workspace_bytes = 256;
  return BSC_OK;
}

int bsc_infonce_fwd(const bsc_infonce_desc* p, float* stats, const bsc_infonce_exchange* ex, void* workspace,
                    size_t workspace_bytes, void* stream) {
  int rc = check_desc(p, "bsc_infonce_fwd");
  if (rc) return rc;
  BSC_REQUIRE(stats && workspace, BSC_EINVAL, "bsc_infonce_fwd: null pointer");
  const InfoNceFwdWs L = infonce_fwd_layout(p);
  BSC_REQUIRE(workspace_bytes >= L.total, BSC_EWORKSPACE, "bsc_infonce_fwd: workspace %zu < %zu", workspace_bytes,
              L.total);
  if (ex != nullptr) {
    BSC_REQUIRE(mode_is_tc(p->mode) || (!ex->arrive && !ex->stats_peers && !ex->stats_multicast), BSC_EUNSUPPORTED,
                "bsc_infonce_fwd: the peer-memory exchange is implemented for the tensor-core modes");
    BSC_REQUIRE(p->mode != BSC_MODE_TC16X3 || !ex->arrive, BSC_EUNSUPPORTED,
                "bsc_infonce_fwd: the fused all-gather (arrival counters) carries BSC_MODE_TC16 rows only");
    if (ex->stats_peers)
      for (int r = 0; r < p->world; ++r)
        BSC_REQUIRE(ex->stats_peers[r] != nullptr, BSC_EINVAL, "bsc_infonce_fwd: stats area of rank %d is null", r);
  }
  cudaStream_t st = (cudaStream_t)stream;
  char* ws = (char*)workspace;
  const int64_t N = p->n_global;
  BSC_CUDA(cudaMemsetAsync(stats, 0, (size_t)bsc_infonce_stats_len(N) * sizeof(float), st));
  if (p->n_local == 0) return BSC_OK;
  if (p->mode == BSC_MODE_F32) return simt_fwd_stats(p, L, ws, stats, st);
  return tc_fwd_stats(p, L, ws, stats, ex, st);
}

size_t bsc_infonce_loss_scratch_bytes(void) { return stats_finish_scratch_bytes(); }

int bsc_infonce_loss(const bsc_infonce_desc* p, const float* records, int n_records, float* stats, float* aux,
                     float* loss, void* scratch, void* stream) {
  int rc = check_desc(p, "bsc_infonce_loss");
  if (rc) return rc;
  BSC_REQUIRE(records && stats && aux && loss && scratch, BSC_EINVAL, "bsc_infonce_loss: null pointer");
  BSC_REQUIRE(((uintptr_t)scratch & 15) == 0, BSC_EINVAL, "bsc_infonce_loss: scratch must be 16-byte aligned");
  BSC_REQUIRE(n_records >= 1 && n_records <= kMaxPeers, BSC_EINVAL, "bsc_infonce_loss: %d records", n_records);
  unsigned mask = 0;
  present_count(p, &mask);
  StatsFinishArgs a{};
  a.slots = records;
  a.world = n_records;
  a.n = p->n_global;
  a.record_len = stats_record_len(p->n_global);
  a.r = p->r;
  a.rmax = p->rmax;
  for (int m = 0; m < 3; ++m) a.inv[m] = mode_is_tc(p->mode) ? p->inv[m] : nullptr;
  a.present_mask = mask;
  a.use_gs = mode_is_tc(p->mode) ? 1 : 0;
  a.stats = stats;
  a.uv = aux;
  a.gsbuf = aux + 6 * p->n_global;
  a.loss = loss;
  a.scratch = (char*)scratch;
  return stats_finish_launch(a, (cudaStream_t)stream);
}

int bsc_infonce_bwd_partials(const bsc_infonce_desc* p, const float* aux, void* const* exchange, void* workspace,
                             size_t workspace_bytes, void* stream) {
  int rc = check_desc(p, "bsc_infonce_bwd_partials");
  if (rc) return rc;
  rc = bwd_mode_ok(p, p->mode == BSC_MODE_F32 ? BSC_BWD_RECOMPUTE : p->bwd_mode, 1, "bsc_infonce_bwd_partials");
  if (rc) return rc;
  BSC_REQUIRE(aux && workspace, BSC_EINVAL, "bsc_infonce_bwd_partials: null pointer");
  const InfoNceBwdWs L = infonce_bwd_layout(p);
  BSC_REQUIRE(workspace_bytes >= L.total, BSC_EWORKSPACE, "bsc_infonce_bwd_partials: workspace %zu < %zu",
              workspace_bytes, L.total);
  cudaStream_t st = (cudaStream_t)stream;
  char* ws = (char*)workspace;
  if (p->n_local == 0) return BSC_OK;
  if (p->mode == BSC_MODE_F32) return simt_bwd_raw(p, L, ws, aux, st);
  g_bwd_mode.store(L.mode);
  if (L.mode == BSC_BWD_RECOMPUTE) return tc_bwd_duo(p, L, ws, aux, st);
  ScatterDst dst{};
  dst.world = p->world;
  dst.rank = p->rank;
  if (L.mode == BSC_BWD_SCATTER) {
    BSC_REQUIRE(exchange != nullptr, BSC_EINVAL, "bsc_infonce_bwd_partials: BSC_BWD_SCATTER needs the exchange buffers");
    for (int r = 0; r < p->world; ++r) {
      BSC_REQUIRE(exchange[r] != nullptr && ((uintptr_t)exchange[r] & 15) == 0, BSC_EINVAL,
                  "bsc_infonce_bwd_partials: exchange buffer of rank %d is null or not 16-byte aligned", r);
      dst.peer[r] = (float*)exchange[r];
    }
  }
  int np = 0, pa[6], pb[6];
  rc = tc_bwd_weights(p, L, ws, aux, &np, pa, pb, st);
  if (rc) return rc;
  return tc_bwd_gemms(p, L, ws, aux, np, pa, pb, dst, st);
}

int bsc_infonce_bwd_finish(const bsc_infonce_desc* p, const float* aux, const void* const* x, int x_dtype, const int64_t* ldx,
                           const float* grad_out, void* const* grad, const void* own_exchange, void* workspace,
                           size_t workspace_bytes, void* stream) {
  int rc = check_desc(p, "bsc_infonce_bwd_finish");
  if (rc) return rc;
  BSC_REQUIRE(aux && x && ldx && grad && grad_out && workspace, BSC_EINVAL, "bsc_infonce_bwd_finish: null pointer");
  const InfoNceBwdWs L = infonce_bwd_layout(p);
  BSC_REQUIRE(workspace_bytes >= L.total, BSC_EWORKSPACE, "bsc_infonce_bwd_finish: workspace %zu < %zu",
              workspace_bytes, L.total);
  unsigned mask = 0;
  const int m = present_count(p, &mask);
  for (int a = 0; a < 3; ++a) {
    BSC_REQUIRE(!p->xh[a] || (x[a] && grad[a]), BSC_EINVAL,
                "bsc_infonce_bwd_finish: x/grad pointer missing for modality %d", a);
    BSC_REQUIRE(!p->xh[a] || ldx[a] >= p->d, BSC_EINVAL, "bsc_infonce_bwd_finish: ldx[%d] < d", a);
  }
  const bool scatter = mode_is_tc(p->mode) && L.mode == BSC_BWD_SCATTER;
  BSC_REQUIRE(!scatter || own_exchange, BSC_EINVAL, "bsc_infonce_bwd_finish: BSC_BWD_SCATTER needs own_exchange");
  if (p->n_local == 0) return BSC_OK;
  char* ws = (char*)workspace;
  const int64_t slab = p->n_local * (int64_t)p->d;
  const float* dxraw = (const float*)(ws + L.dxraw);
  CombineArgs ca{};
  int pair_index[3][3];  // unordered pair (a < b) -> its index among the present pairs
  {
    int q = 0;
    for (int a = 0; a < 3; ++a)
      for (int b = a + 1; b < 3; ++b) pair_index[a][b] = pair_index[b][a] = (p->xh[a] && p->xh[b]) ? q++ : -1;
  }
  const int np = m * (m - 1) / 2;
  for (int a = 0; a < 3; ++a) {
    ca.x[a] = p->xh[a] ? x[a] : nullptr;
    ca.grad[a] = p->xh[a] ? grad[a] : nullptr;
    ca.ldx[a] = ldx[a];
    if (!p->xh[a]) continue;
    int slot = 0;
    for (int y = 0; y < 3; ++y) {
      if (y == a || !p->xh[y]) continue;
      CombineSrc& s = ca.src[a][slot];
      s.gs_index = 2 * pair_id(a, y) + 1;
      if (!scatter) {
        // one slab group per (modality, partner slot): RECOMPUTE, BLOCKS and the fp32 path
        s.src = dxraw + (int64_t)(a * 2 + slot) * L.kparts * slab;
        // BLOCKS / RECOMPUTE with two modalities only: slots are still addressed as (a * 2 + slot)
        s.stride = slab;
        s.nparts = L.kparts;
      } else if (a < y) {  // K-major job of pair (a, y): local partial sums
        s.src = dxraw + (int64_t)pair_index[a][y] * L.kparts * slab;
        s.stride = slab;
        s.nparts = L.kparts;
      } else {             // MN-major job of pair (y, a): one partial sum per source rank in the exchange buffer
        s.src = (const float*)own_exchange + (int64_t)pair_index[a][y] * slab;
        s.stride = (int64_t)np * slab;
        s.nparts = p->world * tc_bwd_mn_kparts(p);
      }
      ++slot;
    }
  }
  ca.gsbuf = aux + 6 * p->n_global;
  ca.grad_out = grad_out;
  ca.n_local = p->n_local;
  ca.n_global = p->n_global;
  ca.d = p->d;
  ca.n_slots = m - 1;
  ca.present_mask = mask;
  ca.scale = p->logit_scale;
  return combine_launch(ca, x_dtype, (cudaStream_t)stream);
}

int bsc_bank_search_workspace_bytes(int64_t nq, int64_t nb, int d, int k, int mode, size_t* bytes) {
  BSC_REQUIRE(bytes != nullptr && nq >= 0 && nb >= 0 && d >= 1 && k >= 1, BSC_EINVAL,
              "bsc_bank_search_workspace_bytes: bad argument");
  if (mode == BSC_MODE_F32)
    *bytes = simt_bank_ws(nq, nb, d, k);
  else if (mode == BSC_MODE_TC16)
    *bytes = tc_bank_ws(nq, nb, d, k);
  else {
    set_error("bsc_bank_search_workspace_bytes: unknown mode %d", mode);
    return BSC_EINVAL;
  }
  if (*bytes == 0) *bytes = 256;
  return BSC_OK;
}

int bsc_bank_search(const void* queries, int64_t nq, const void* bank, int64_t nb, int d, int k, int64_t base_id,
                    int mode, float* out_scores, int64_t* out_ids, void* workspace, size_t workspace_bytes,
                    void* stream) {
  BSC_REQUIRE(nq >= 0 && nb >= 0 && d >= 1 && k >= 1, BSC_EINVAL, "bsc_bank_search: bad shape");
  BSC_REQUIRE((queries || nq == 0) && (bank || nb == 0) && (out_scores && out_ids || nq == 0) && workspace, BSC_EINVAL,
              "bsc_bank_search: null pointer");
  cudaStream_t st = (cudaStream_t)stream;
  if (mode == BSC_MODE_F32)
    return simt_bank_search((const float*)queries, nq, (const float*)bank, nb, d, k, base_id, out_scores, out_ids,
                            (char*)workspace, workspace_bytes, st);
  if (mode == BSC_MODE_TC16)
    return tc_bank_search(queries, nq, bank, nb, d, k, base_id, out_scores, out_ids, (char*)workspace,
                          workspace_bytes, st);
  set_error("bsc_bank_search: unknown mode %d", mode);
  return BSC_EINVAL;
}

int bsc_bank_screen_parts(int64_t nq, int64_t nb, int min_parts, int* parts) {
  BSC_REQUIRE(parts && nq >= 0 && nb >= 0 && min_parts >= 0, BSC_EINVAL, "bsc_bank_screen_parts: bad argument");
  *parts = tc_bank_screen_parts(nq, nb, min_parts);
  return BSC_OK;
}

int bsc_bank_screen(const void* queries_bf16, int64_t nq, const void* bank_bf16, int64_t nb, int d, int64_t base_id,
                    int min_parts, float* cand_scores, int64_t* cand_ids, void* stream) {
  BSC_REQUIRE(nq >= 0 && nb >= 1 && d >= 1, BSC_EINVAL, "bsc_bank_screen: bad shape");
  BSC_REQUIRE((queries_bf16 && cand_scores && cand_ids) || nq == 0, BSC_EINVAL, "bsc_bank_screen: null pointer");
  BSC_REQUIRE(bank_bf16, BSC_EINVAL, "bsc_bank_screen: null bank");
  return tc_bank_screen(queries_bf16, nq, bank_bf16, nb, d, base_id, min_parts, cand_scores, cand_ids,
                        (cudaStream_t)stream);
}

int bsc_bank_rescore(const float* queries, int64_t nq, const float* bank, int64_t nb, int d, int64_t base_id,
                     const float* cand_scores, const int64_t* cand_ids, int parts, int k, float eps_scale,
                     float* out_scores, int64_t* out_ids, int32_t* unsure, int32_t* n_unsure, void* stream) {
  BSC_REQUIRE(nq >= 0 && nb >= 1 && d >= 1 && parts >= 1, BSC_EINVAL, "bsc_bank_rescore: bad shape");
  BSC_REQUIRE((queries && cand_scores && cand_ids && out_scores && out_ids && unsure) || nq == 0, BSC_EINVAL,
              "bsc_bank_rescore: null pointer");
  BSC_REQUIRE(bank && n_unsure, BSC_EINVAL, "bsc_bank_rescore: null pointer");
  BSC_REQUIRE(eps_scale >= 0.f, BSC_EINVAL, "bsc_bank_rescore: negative eps_scale");
  return rescore_launch(queries, nq, bank, nb, d, base_id, cand_scores, cand_ids, parts, k, eps_scale, out_scores,
                        out_ids, unsure, n_unsure, (cudaStream_t)stream);
}

int bsc_topk_merge(const float* cand_scores, const int64_t* cand_ids, int parts, int64_t nq, int k, float* out_scores,
                   int64_t* out_ids, void* stream) {
  BSC_REQUIRE(parts >= 0 && nq >= 0 && k >= 1, BSC_EINVAL, "bsc_topk_merge: bad shape");
  BSC_REQUIRE((cand_scores && cand_ids) || parts == 0 || nq == 0, BSC_EINVAL, "bsc_topk_merge: null candidates");
  BSC_REQUIRE((out_scores && out_ids) || nq == 0, BSC_EINVAL, "bsc_topk_merge: null outputs");
  return topk_merge_launch(cand_scores, cand_ids, parts, nq, k, out_scores, out_ids, (cudaStream_t)stream);
}

}  // extern "C"
