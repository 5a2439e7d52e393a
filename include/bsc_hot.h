/*
 * bsc_hot.h -- C ABI of libbsc_hot.so: the B200 (sm_100a) contrastive hot path of BIOSCAN-CLIP.
 *
 * The reference (bioscan-ml/bioscan-clip @ 382d5ff) is pure Python and has no FFI; the entry points
 * below are what a binding for its two leaf computations would bind:
 *
 *   loss       bioscanclip/model/loss_func.py:23-54 (ContrastiveLoss) and :94-165 (ClipLoss),
 *              called from bioscanclip/epoch/train_epoch.py:37-38
 *   retrieval  scripts/inference_and_eval.py:414-421 (make_prediction: normalize + IndexFlatIP
 *              add/search), bioscanclip/epoch/eval_epoch.py:153-154,203, bioscanclip/util/util.py:163-165
 *
 * Conventions
 *   - every function returns 0 on success or a negative BSC_E* code; bsc_last_error() returns a
 *     thread-local message for the last failure on the calling thread;
 *   - the caller owns ALL memory: every pointer is a device pointer obtained by the caller
 *     (torch tensors in the Python host code) unless its name starts with h_; scratch space is one
 *     caller-allocated workspace whose size comes from the matching *_workspace_bytes call;
 *   - every launch goes to the cudaStream_t passed as `stream` (as void*); no function synchronises
 *     the host with the device or allocates device memory.  Nothing a result depends on is global: the
 *     backward strategy travels in the descriptor (bsc_infonce_plan resolves it once), no entry point
 *     reads the environment.  The only process-wide state is instrumentation (launch counter, the event
 *     pairs behind bsc_last_kernel_ms) and per-device caches of immutable facts (SM count, shared-memory
 *     opt-in of each kernel);
 *   - there is no CPU fallback: without an sm_100 device the launches fail and the error is returned.
 *
 * Row-major everywhere; `ld` arguments are row strides in ELEMENTS.
 */
#ifndef BSC_HOT_H
#define BSC_HOT_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define BSC_VERSION 200

/* error codes */
#define BSC_OK 0
#define BSC_EINVAL (-1)       /* bad argument (null pointer, bad shape, bad enum)            */
#define BSC_EUNSUPPORTED (-2) /* legal request this build does not implement (message says)  */
#define BSC_ECUDA (-3)        /* a CUDA runtime / driver call failed (message has the text)  */
#define BSC_EWORKSPACE (-4)   /* workspace too small                                          */

/* element types */
#define BSC_F32 0
#define BSC_BF16 1
#define BSC_F16 2
#define BSC_F64 3 /* input of bsc_l2norm_fwd only: the reference's eval features are float64 numpy arrays */

/* arithmetic modes */
#define BSC_MODE_F32 0  /* SIMT fp32 path: exact-parity mode (1e-5 vs the fp32 reference)          */
#define BSC_MODE_TC16 1 /* tcgen05 path: the caller's 16-bit rows as operands, fp32 accumulate in TMEM,
                           row norms applied as fp32 scales in the epilogue                           */
#define BSC_MODE_TC16X3 2 /* tcgen05 path at float32 grade: every row is split into two fp16 parts (hi + lo = 22
                           significant bits) and every product x*y is formed as lo*hi + hi*lo + hi*hi on the
                           tensor cores with fp32 accumulation (the fp32 sgemm of loss_func.py:46-47 at 3 MMAs
                           per product); the stored gradient weights are split the same way                */

/* zero-norm rules for bsc_l2norm_fwd */
#define BSC_NORM_TORCH 0   /* F.normalize: x / max(||x||, 1e-12)            (loss_func.py:43-44)   */
#define BSC_NORM_SKLEARN 1 /* sklearn normalize: ||x|| < 10*eps64 -> x kept (inference_and_eval.py:416) */

#define BSC_MAX_MODALITIES 3

int bsc_version(void);
const char* bsc_last_error(void);
/* 0 when the current device is sm_100 and the tcgen05 kernels can be launched. */
int bsc_device_check(void);

/* ------------------------------------------------------------------------------------------------
 * K1  row L2 normalisation (F.normalize / sklearn normalize) -- loss_func.py:43-44,150-151,
 *     simple_clip.py:34-49, inference_and_eval.py:416-417.
 *     y16 (BSC_F16 or BSC_BF16, may be NULL) and y32 (float, may be NULL) receive x * inv_norm;
 *     inv_norm[n] (may be NULL) receives the factor itself.
 *     x_dtype BSC_F64: norm and quotient are formed in float64 and rounded ONCE to the output type, as
 *     sklearn normalize(float64 array).astype(float32) does (inference_and_eval.py:416-417).
 * ---------------------------------------------------------------------------------------------- */
int bsc_l2norm_fwd(const void* x, int x_dtype, int64_t n, int d, int64_t ldx,
                   void* y16, int y16_dtype, float* y32, float* inv_norm, int zero_rule, void* stream);

/* TC16 operand preparation: y16[n][d] (fp16) = x * 2^k_row with the row maximum scaled into [1,2) -- exact for
 * bfloat16 / float16 values -- and inv[n] = 2^-k_row / max(||x_row||, 1e-12), so that
 * cos_ij = <y16_i, y16_j> * inv_i * inv_j equals the reference's normalised dot product without re-rounding
 * any operand (F.normalize, loss_func.py:43-44). */
int bsc_rows_to_f16(const void* x, int x_dtype, int64_t n, int d, int64_t ldx, void* y16, float* inv, void* stream);

/* BSC_MODE_TC16X3 operand preparation (same F.normalize lines): hl[n][2 d] (fp16), columns [0, d) = hi and [d, 2 d) = lo
 * with hi = fp16(x * 2^k_row), lo = fp16(x * 2^k_row - hi) -- the row maximum is scaled into [128, 256) so that the lo
 * parts of all but negligible elements stay normal fp16 numbers -- and inv[n] = 2^-k_row / max(||x_row||, 1e-12). */
int bsc_rows_to_f16x3(const void* x, int x_dtype, int64_t n, int d, int64_t ldx, void* hl, float* inv, void* stream);

/* ------------------------------------------------------------------------------------------------
 * K1+gather  TC16 operand preparation fused with the all_gather of gather_features (loss_func.py:58-91) and of the
 * labels (:122) over NVLink peer memory.  Every rank owns a buffer of bsc_gather_layout().total_bytes that is mapped
 * into all ranks of the node (peer_buffers[r] = address of rank r's buffer in THIS process, e.g. torch symmetric memory;
 * multicast_buffer = the NVSwitch multicast address of the same buffer or NULL):
 *     xh[m]  at off_xh[m]: n_global x d fp16     inv[m] at off_inv[m]: n_global fp32     labels at off_labels: int64
 *     statistics records at off_stats: float[world][bsc_infonce_stats_len]     arrival counters at off_flags: u64[16]
 * bsc_gather_push_labels stores the local labels into every rank's buffer (all labels are needed before the label
 * statistics can start; a cross-rank barrier follows).
 * bsc_gather_push computes, for the n_local rows of every present modality, what bsc_rows_to_f16 computes and stores row
 * row0 + l (fp16 row, scale, and the label when `labels` is given) into every rank's buffer.
 *   ordered = 0: all destinations at once (one multimem.st per 16 bytes when multicast_buffer is given); the caller runs a
 *                cross-rank barrier before any rank reads.
 *   ordered = 1: the fused all-gather -> similarity sweep.  The call writes the rows into this rank's OWN buffer only
 *                and then publishes arrive[rank] = epoch there (epoch = number of ordered pushes into this buffer so far,
 *                this one included: 1, 2, ...).  bsc_gather_ring, queued on a SIDE stream after this call has been
 *                reached, lets the COPY ENGINES carry rows, inverse norms and finally the flag to rank+1, rank+2, ...
 *                (cudaMemcpyAsync between the peer-mapped buffers; copies of a stream complete in order, so a peer that
 *                sees the flag sees the rows; a receiver hears from one sender at a time).  No kernel is involved: the
 *                persistent sweep may own every register of every SM.  The consumer (bsc_infonce_fwd with
 *                bsc_infonce_exchange.arrive, arrive_target = epoch) waits per source rank and sweeps the sources in
 *                their order of arrival while the later ones are still in flight.  No barrier.
 * The caller must not push into a buffer a peer may still read (alternate two buffers by step; every step contains at
 * least one cross-rank barrier).  All modalities share x_dtype; x[m] = NULL marks an absent modality.
 * Measured on 8 x B200: the gather delivers (W-1)/W of the global batch into every GPU, 132 MB of NVLink ingress =
 * 0.22 ms whichever way the bytes are sent -- it can be overlapped, not shortened.  SM stores that back up behind a
 * saturated NVLink port stall every memory operation of their SM (co-resident kernels ran up to 4x slower), hence DMA.
 * ---------------------------------------------------------------------------------------------- */
int bsc_gather_layout(int64_t n_global, int d, int world, size_t off_xh[3], size_t off_inv[3], size_t* off_labels,
                      size_t* off_stats, size_t* off_flags, size_t* total_bytes);
int bsc_gather_push_labels(const int64_t* labels, int64_t n_local, int64_t row0, int64_t n_global, int d,
                           void* const* peer_buffers, void* multicast_buffer, int world, void* stream);
int bsc_gather_push(const void* const x[3], int x_dtype, const int64_t ldx[3], const int64_t* labels, int64_t n_local,
                    int64_t row0, int64_t n_global, int d, void* const* peer_buffers, void* multicast_buffer, int world,
                    int rank, int ordered, uint64_t epoch, void* stream);
int bsc_gather_ring(void* const* peer_buffers, int world, int rank, int64_t n_local, int64_t row0, int64_t n_global,
                    int d, unsigned present_mask, void* stream);

/* ------------------------------------------------------------------------------------------------
 * K2  label statistics replacing construct_label_metrix (loss_func.py:18-21): for labels[n]
 *     r[i]    = #{j : labels[j] == labels[i]}      (row sum of the 0/1 label matrix)
 *     rep[i]  = min{j : labels[j] == labels[i]}    (class representative)
 *     next[i] = min{j > i : labels[j] == labels[i]} or -1
 *     rmax[0] = max_i r[i] (float)
 *     workspace: caller-owned device scratch of bsc_label_stats_workspace_bytes(n) bytes (hash table of the labels).
 * ---------------------------------------------------------------------------------------------- */
int bsc_label_stats_workspace_bytes(int64_t n, size_t* bytes);
int bsc_label_stats(const int64_t* labels, int64_t n, float* r, int32_t* rep, int32_t* next, float* rmax,
                    void* workspace, size_t workspace_bytes, void* stream);

/* bsc_label_stats without a workspace argument and without any allocation.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight
 * serving loops, harnesses that forbid it).  Same kernels and same results as bsc_label_stats; the scratch is a
 * file-scope __device__ array inside the library (1.25 MiB, held by every GPU that loads the library until exit), its
 * device address resolved with cudaGetSymbolAddress on the first call per device.  The call reaches no cudaMalloc,
 * cudaFree or allocating library routine.
 *   limit: n <= BSC_LABEL_STATS_STATIC_MAX_N, larger n -> BSC_EUNSUPPORTED (use bsc_label_stats).
 *   every call on a device shares the one array, which keeps the previous call's contents; the call itself clears what
 *   must start empty, on `stream`.  Calls that overlap on the GPU (different streams, no ordering between them) RACE:
 *   serialise them. */
#define BSC_LABEL_STATS_STATIC_MAX_N 65536
int bsc_label_stats_static(const int64_t* labels, int64_t n, float* r, int32_t* rep, int32_t* next, float* rmax,
                           void* stream);

/* ------------------------------------------------------------------------------------------------
 * InfoNCE loss -- ContrastiveLoss.forward / ClipLoss.forward.
 *
 * The problem descriptor.  xh[m] points to the GLOBAL (already gathered) batch of modality m, NULL when the
 * modality is absent; this rank owns rows [row0, row0 + n_local).
 *   BSC_MODE_F32   xh[m] = n_global L2-NORMALISED rows of d floats (bsc_l2norm_fwd); inv[m] is ignored.
 *   BSC_MODE_TC16  xh[m] = n_global rows of d fp16 elements (x16_dtype = BSC_F16) and inv[m][n_global], both
 *                  from bsc_rows_to_f16: the caller's values times an exact power of two per row.  The tensor
 *                  cores multiply those rows; cos_ij = <x_i, y_j> * inv_i * inv_j is formed in fp32, so no
 *                  operand is re-rounded.
 *   BSC_MODE_TC16X3  xh[m] = n_global rows of 2 d fp16 elements (hi | lo) and inv[m][n_global] from bsc_rows_to_f16x3;
 *                  d is still the embedding width.  Backward strategies: BSC_BWD_BLOCKS / BSC_BWD_SCATTER (the weight
 *                  blocks are stored as hi and lo images: twice the bytes of BSC_MODE_TC16).
 * ---------------------------------------------------------------------------------------------- */
typedef struct bsc_infonce_desc {
  int64_t n_global;
  int64_t n_local;
  int64_t row0;
  int32_t d;
  int32_t mode;        /* BSC_MODE_*                                              */
  float logit_scale;   /* s = 1/0.07 in the reference (scripts/train_cl.py:190)   */
  int32_t x16_dtype;   /* BSC_MODE_TC16: element type of xh (BSC_F16)             */
  const void* xh[BSC_MAX_MODALITIES];
  const float* inv[BSC_MAX_MODALITIES]; /* BSC_MODE_TC16: inverse row norms          */
  const float* r;      /* [n_global] from bsc_label_stats                          */
  const int32_t* rep;  /* [n_global]                                               */
  const int32_t* next; /* [n_global]                                               */
  const float* rmax;   /* [1]                                                      */
  int32_t bwd_mode;    /* BSC_BWD_* as resolved by bsc_infonce_plan (never BSC_BWD_AUTO here) */
  int32_t world;       /* ranks sharing the global batch (>= 1)                    */
  int32_t rank;        /* this rank; BSC_BWD_SCATTER needs row0 == rank * n_local and n_global == world * n_local */
  int32_t reserved;
} bsc_infonce_desc;

/* Backward strategies of BSC_MODE_TC16 (BSC_MODE_F32 always recomputes through its fp32 slab).
 *   BSC_BWD_RECOMPUTE  one fused sweep per ORDERED pair that recomputes S tile by tile and keeps the gradient block in
 *                      TMEM: nothing N^2-sized ever reaches HBM (d in {256, 512, 768}); executes 24 n N D.
 *   BSC_BWD_BLOCKS     the fp16 weight block G_ab[local rows, all columns] of every ORDERED pair is written once
 *                      (2 n N bytes each) and one GEMM per block forms dA = G B: no exchange between ranks
 *                      (gather_with_grad=False semantics, loss_func.py:84-89); executes 24 n N D.
 *   BSC_BWD_SCATTER    one weight block per UNORDERED pair serves both sides: dA_local = G B (local rows) and the
 *                      partial dB_all = G^T A_local for ALL rows, whose row blocks are stored straight into their
 *                      owners' exchange buffers (peer memory over NVLink: a fused GEMM + reduce-scatter, the backward
 *                      of the autograd-aware all_gather of loss_func.py:81-82); bsc_infonce_bwd_finish sums the W
 *                      sources in rank order.  Executes 18 n N D -- the single-GPU count on every rank.
 * Stored-weight workspace: (pairs or ordered pairs) * 2 * n_local * n_global bytes -- the one N^2-sized buffer. */
#define BSC_BWD_AUTO (-1)
#define BSC_BWD_RECOMPUTE 0
#define BSC_BWD_BLOCKS 3
#define BSC_BWD_SCATTER 4

typedef struct bsc_infonce_plan_t {
  int32_t bwd_mode;            /* resolved strategy                                                              */
  int32_t reserved;
  size_t fwd_workspace_bytes;  /* bsc_infonce_fwd                                                                */
  size_t bwd_workspace_bytes;  /* bsc_infonce_bwd_partials / _finish (allocate only when a backward will run)    */
  size_t exchange_bytes;       /* BSC_BWD_SCATTER: this rank's exchange buffer float[world * parts][pairs][n_local][d]
                                  (parts = 1; BSC_MODE_TC16X3: one slot per 1024-row accumulation chain of a source
                                  rank's rows), mapped into every peer when world > 1 (torch symmetric memory); 0 for
                                  the other modes                                                                 */
} bsc_infonce_plan_t;

/* A statistics RECORD: bsc_infonce_stats_len(n_global) floats = 6 ordered-pair sum-exp rows of n_global | 4 scalars |
 * 64 double partial sums of the positive-pair term.  bsc_infonce_aux_len floats hold what the backward needs from the
 * combined statistics: uv[6][n_global] = r_i / sum-exp, then 32 scale factors. */
int64_t bsc_infonce_stats_len(int64_t n_global);
int64_t bsc_infonce_aux_len(int64_t n_global);

/* Resolve the backward strategy and size the buffers.  p->bwd_mode is ignored; requested_bwd_mode = BSC_BWD_AUTO picks
 * SCATTER when all rows are local or the ranks can map each other's memory (peer_memory != 0), BLOCKS otherwise, and
 * falls back to RECOMPUTE when the stored weights would exceed memory_budget_bytes (0 = no limit).  An explicit request
 * that cannot be honoured is an error (BSC_EUNSUPPORTED), never a silent substitution. */
int bsc_infonce_plan(const bsc_infonce_desc* p, int requested_bwd_mode, int peer_memory, size_t memory_budget_bytes,
                     bsc_infonce_plan_t* plan);

/* Peer-memory side of the forward (all fields optional; NULL = none). */
typedef struct bsc_infonce_exchange {
  void* const* stats_peers;  /* [world] the statistics-record area (float[world][stats_len]) of every rank's buffer: the
                                forward also stores its record into slot p->rank of each -- the statistics exchange
                                without an NCCL call; the caller runs a barrier before bsc_infonce_loss               */
  void* stats_multicast;     /* multicast address of the same area: one store per value instead of `world`            */
  const uint64_t* arrive;    /* [world] arrival counters in THIS rank's gather buffer (bsc_gather_push ordered = 1):
                                the sweep waits per source rank; NULL: the operands are complete on entry             */
  uint64_t arrive_target;    /* the epoch of the push whose rows this forward reads                                   */
} bsc_infonce_exchange;

/* Forward sweep.  stats (one record, overwritten) receives, for every ordered pair (a,b) of present modalities,
 * sum_j exp(S_ab[i,j] - s) for the rows this rank can complete and the PARTIAL column sums over its local rows, plus the
 * local part of sum over unordered pairs sum_ij T_ij (S_ab[i,j] - s), taken from the same logits as the exponentials. */
int bsc_infonce_fwd(const bsc_infonce_desc* p, float* stats, const bsc_infonce_exchange* exchange, void* workspace,
                    size_t workspace_bytes, void* stream);

/* Statistics finish: records = n_records per-rank records (this rank's own one when n_records == 1), added in rank
 * order (deterministic); stats receives the combined record, aux the backward's factors, loss[0] the scalar of
 * loss_func.py:54 / :163.  scratch: bsc_infonce_loss_scratch_bytes() of device memory that the caller zeroes ONCE; every
 * call leaves it zeroed again (a ticket counter and per-block partial sums; one scratch per stream in flight). */
size_t bsc_infonce_loss_scratch_bytes(void);
int bsc_infonce_loss(const bsc_infonce_desc* p, const float* records, int n_records, float* stats, float* aux,
                     float* loss, void* scratch, void* stream);

/* Backward, phase 1: raw gradient sums.  Writes the local partial sums into the workspace and, in BSC_BWD_SCATTER, the
 * partial sums that belong to rank q's rows into exchange[q] (slot p->rank).  exchange = p->world device pointers in
 * THIS process's address space (exchange[p->rank] = own buffer); NULL for the other strategies.  With world > 1 the
 * caller runs a cross-rank barrier on the stream between this call and bsc_infonce_bwd_finish, and must not start the
 * next step's phase 1 before every rank has finished reading its buffer (the forward's barrier of the next step). */
int bsc_infonce_bwd_partials(const bsc_infonce_desc* p, const float* aux, void* const* exchange, void* workspace,
                             size_t workspace_bytes, void* stream);

/* Backward, phase 2: x[m] are this rank's ORIGINAL (un-normalised) local rows, dtype x_dtype, row stride ldx[m]
 * elements; grad[m] (same dtype, DENSE [n_local][d]) receives grad_out[0] * dLoss/dx[m]: fixed-order sum of the partial
 * sums (workspace and own_exchange), then the chain through F.normalize. */
int bsc_infonce_bwd_finish(const bsc_infonce_desc* p, const float* aux, const void* const* x, int x_dtype, const int64_t* ldx,
                           const float* grad_out, void* const* grad, const void* own_exchange, void* workspace,
                           size_t workspace_bytes, void* stream);

/* ------------------------------------------------------------------------------------------------
 * Retrieval -- IndexFlatIP.search on one bank shard (inference_and_eval.py:418-421).
 *   bank    [nb, d] normalised rows, float (BSC_MODE_F32) or bf16 (BSC_MODE_TC16)
 *   queries [nq, d] same dtype
 *   out_scores [nq, k] float descending, out_ids [nq, k] int64 = base_id + row; equal scores are
 *   ordered by lowest id; missing entries (k > nb) are id -1 / score -3.4028235e38 as faiss does.
 * ---------------------------------------------------------------------------------------------- */
int bsc_bank_search_workspace_bytes(int64_t nq, int64_t nb, int d, int k, int mode, size_t* bytes);
int bsc_bank_search(const void* queries, int64_t nq, const void* bank, int64_t nb, int d, int k,
                    int64_t base_id, int mode, float* out_scores, int64_t* out_ids, void* workspace,
                    size_t workspace_bytes, void* stream);
/* ------------------------------------------------------------------------------------------------
 * Exact float32 search at tensor-core speed (same results as BSC_MODE_F32 of bsc_bank_search, i.e. as
 * IndexFlatIP.search on float32 data): screen + rescore.
 *   bsc_bank_screen   bf16 scan of a bf16 copy of the shard; hands out `parts` unmerged candidate lists of 8 entries
 *                     [parts][nq][8] (score, id; id -1 = empty).  parts = bsc_bank_screen_parts(nq, nb, min_parts)
 *                     >= min_parts.  A row that is in no list has a bf16 score <= the last entry of its own list.
 *   bsc_bank_rescore  exact float32 dot products of every candidate with the float32 query, ordered (score desc,
 *                     id asc) -> out [nq,k], k <= 8.  unsure[q] = 0 when the k-th exact score exceeds the best
 *                     possible score of any non-candidate (its list's last bf16 score + eps_scale * ||q||, eps_scale =
 *                     2^-8 * max bank row norm * slack): then the result provably equals the exhaustive float32 search.
 *                     Queries with unsure[q] = 1 (n_unsure of them) must be re-run with bsc_bank_search(BSC_MODE_F32).
 * ---------------------------------------------------------------------------------------------- */
int bsc_bank_screen_parts(int64_t nq, int64_t nb, int min_parts, int* parts);
int bsc_bank_screen(const void* queries_bf16, int64_t nq, const void* bank_bf16, int64_t nb, int d, int64_t base_id,
                    int min_parts, float* cand_scores, int64_t* cand_ids, void* stream);
int bsc_bank_rescore(const float* queries, int64_t nq, const float* bank, int64_t nb, int d, int64_t base_id,
                     const float* cand_scores, const int64_t* cand_ids, int parts, int k, float eps_scale,
                     float* out_scores, int64_t* out_ids, int32_t* unsure, int32_t* n_unsure, void* stream);
/* k-way merge of `parts` candidate lists [parts][nq][k] (e.g. one per GPU after an all_gather). */
int bsc_topk_merge(const float* cand_scores, const int64_t* cand_ids, int parts, int64_t nq, int k,
                   float* out_scores, int64_t* out_ids, void* stream);

/* ------------------------------------------------------------------------------------------------
 * Taxonomy bookkeeping of a search result (scripts/inference_and_eval.py:422-433 label mapping, :448-464
 * top_k_micro_accuracy, :467-511 top_k_macro_accuracy) on label ids: every taxonomy level's strings are interned
 * to int32 ids 0..classes_l-1 by the caller (-1 = no label); class_off[l] is the first slot of level l in the
 * per-class arrays of total_classes entries.
 *   indices [nq,k] int64 (-1 = padding)      key_ids [nb,levels]      gt_ids [nq,levels]      k_list[nk] (nk <= 8)
 *   pred_ids   [nq,k,levels] label id of every hit (NULL to skip)
 *   first_hit  [nq,levels]   smallest rank t with key_ids[indices[q,t],l] == gt_ids[q,l], 255 = none (NULL to skip)
 *   micro      [nk,levels]   #queries whose ground truth is among the first k_list[kk] hits (int64)
 *   pred_counts[nk,total_classes], gt_counts[total_classes]  the reference's pred_counts / gt_counts per class
 * Outputs are zeroed by the call.  All arithmetic is integer: results are exact.
 * ---------------------------------------------------------------------------------------------- */
int bsc_taxonomy_eval(const int64_t* indices, int64_t nq, int k, const int32_t* key_ids, int64_t nb,
                      const int32_t* gt_ids, int levels, const int32_t* k_list, int nk, const int64_t* class_off,
                      int64_t total_classes, int32_t* pred_ids, uint8_t* first_hit, int64_t* micro,
                      int32_t* pred_counts, int32_t* gt_counts, void* stream);

/* ------------------------------------------------------------------------------------------------
 * Instrumentation for bench.py: CUDA-event time of the dominant kernel of the last call made with
 * bsc_set_timing(1) on this thread (milliseconds, -1 if none).  which: 0 = infonce forward sweep,
 * 1 = infonce backward sweep (stored-weight strategies: the weight-block kernel), 2 = bank scan, 3 = gradient GEMMs
 * from the stored weight blocks.  Reading it synchronises on the recorded events only.
 * bsc_last_bwd_mode: the BSC_BWD_* strategy of the most recent bsc_infonce_bwd_partials in TC16 mode (-1 before the
 * first call).
 * ---------------------------------------------------------------------------------------------- */
int bsc_set_timing(int enabled);
float bsc_last_kernel_ms(int which);
int bsc_last_bwd_mode(void);
/* kernels launched by this library on this thread since the last reset */
int64_t bsc_launch_count(int reset);

#ifdef __cplusplus
}
#endif
#endif /* BSC_HOT_H */
