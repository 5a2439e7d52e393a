// O(N*D) and O(N^2)-integer kernels around the similarity sweeps: row L2 normalisation (K1), label
// statistics (K2), class sums, the positive-pair term, the loss scalar and the normalize backward (K6).
// All of them are HBM- or latency-bound; none touches an N x N float matrix.
#include "common.cuh"

#include <cooperative_groups.h>

#include <atomic>

namespace bsc {

template <typename T>
__device__ __forceinline__ float to_f32(T v);
template <>
__device__ __forceinline__ float to_f32<float>(float v) { return v; }
template <>
__device__ __forceinline__ float to_f32<__nv_bfloat16>(__nv_bfloat16 v) { return __bfloat162float(v); }
template <>
__device__ __forceinline__ float to_f32<__half>(__half v) { return __half2float(v); }

template <typename T>
__device__ __forceinline__ T from_f32(float v);
template <>
__device__ __forceinline__ float from_f32<float>(float v) { return v; }
template <>
__device__ __forceinline__ __nv_bfloat16 from_f32<__nv_bfloat16>(float v) { return __float2bfloat16_rn(v); }
template <>
__device__ __forceinline__ __half from_f32<__half>(float v) { return __float2half_rn(v); }

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// ---------------------------------------------------------------------------------------------
// K1: one warp per row.  Two passes over the row (second pass hits L1/L2).
// ---------------------------------------------------------------------------------------------
template <typename TI, typename TO>
__global__ void __launch_bounds__(256) l2norm_rows_kernel(const TI* __restrict__ x, int64_t n, int d, int64_t ldx,
                                                          TO* __restrict__ y16, float* __restrict__ y32,
                                                          float* __restrict__ inv_norm, int zero_rule) {
  const int lane = threadIdx.x & 31;
  const int64_t row = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= n) return;
  const TI* xr = x + row * ldx;
  float ss = 0.f;
  for (int c = lane; c < d; c += 32) {
    float v = to_f32<TI>(xr[c]);
    ss = fmaf(v, v, ss);
  }
  ss = warp_sum(ss);
  float nrm = sqrtf(ss);
  float inv;
  if (zero_rule == BSC_NORM_SKLEARN) {
    inv = (nrm < 2.220446049250313e-15f) ? 1.0f : 1.0f / nrm;
  } else {
    inv = 1.0f / fmaxf(nrm, 1e-12f);
  }
  if (inv_norm != nullptr && lane == 0) inv_norm[row] = inv;
  for (int c = lane; c < d; c += 32) {
    float v = to_f32<TI>(xr[c]) * inv;
    if (y16 != nullptr) y16[row * (int64_t)d + c] = from_f32<TO>(v);
    if (y32 != nullptr) y32[row * (int64_t)d + c] = v;
  }
}

// TC16 operand preparation: y16[i,:] = fp16(x[i,:] * 2^k_i) with k_i chosen so the largest |element| of the row
// lands in [1, 2), inv[i] = 2^-k_i / max(||x_i||, 1e-12).  A power-of-two scale is exact, and bf16 / fp16 values
// convert to fp16 without rounding (elements below 2^-24 of the row maximum aside), so the tensor cores see the
// caller's numbers verbatim; cos_ij = <y16_i, y16_j> * inv_i * inv_j is then formed in fp32.
template <typename TI>
__global__ void __launch_bounds__(256) rows_to_f16_kernel(const TI* __restrict__ x, int64_t n, int d, int64_t ldx,
                                                          __half* __restrict__ y16, float* __restrict__ inv) {
  const int lane = threadIdx.x & 31;
  const int64_t row = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= n) return;
  const TI* xr = x + row * ldx;
  float ss = 0.f, amax = 0.f;
  for (int c = lane; c < d; c += 32) {
    const float v = to_f32<TI>(xr[c]);
    ss = fmaf(v, v, ss);
    amax = fmaxf(amax, fabsf(v));
  }
  ss = warp_sum(ss);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, o));
  int k = 0;
  if (amax > 0.f && amax < 3.0e38f) k = -ilogbf(amax);
  k = max(-120, min(120, k));
  const float scale = exp2f((float)k);
  const float nrm = sqrtf(ss);
  // ss itself can overflow / underflow for extreme rows; the scaled row never does
  float inv_row;
  if (nrm > 1e-12f && nrm < 3.0e38f) {
    inv_row = exp2f((float)-k) / nrm;
  } else {
    float s2 = 0.f;
    for (int c = lane; c < d; c += 32) {
      const float v = to_f32<TI>(xr[c]) * scale;
      s2 = fmaf(v, v, s2);
    }
    s2 = warp_sum(s2);
    const float n2 = sqrtf(s2);                     // norm of the scaled row = ||x|| * 2^k
    const float true_nrm = n2 * exp2f((float)-k);   // may flush to zero: then F.normalize clamps at 1e-12
    inv_row = n2 > 0.f ? (true_nrm >= 1e-12f ? 1.0f / n2 : exp2f((float)-k) * 1e12f) : 0.f;
  }
  if (lane == 0) inv[row] = inv_row;
  for (int c = lane; c < d; c += 32) y16[row * (int64_t)d + c] = __float2half_rn(to_f32<TI>(xr[c]) * scale);
}

int rows_to_f16_launch(const void* x, int x_dtype, int64_t n, int d, int64_t ldx, void* y16, float* inv,
                       cudaStream_t st) {
  if (n == 0) return BSC_OK;
  dim3 block(256), grid((unsigned)((n + 7) / 8));
  switch (x_dtype) {
    case BSC_F32:
      rows_to_f16_kernel<float><<<grid, block, 0, st>>>((const float*)x, n, d, ldx, (__half*)y16, inv);
      break;
    case BSC_BF16:
      rows_to_f16_kernel<__nv_bfloat16><<<grid, block, 0, st>>>((const __nv_bfloat16*)x, n, d, ldx, (__half*)y16, inv);
      break;
    case BSC_F16:
      rows_to_f16_kernel<__half><<<grid, block, 0, st>>>((const __half*)x, n, d, ldx, (__half*)y16, inv);
      break;
    default:
      set_error("bsc_rows_to_f16: unknown x_dtype %d", x_dtype);
      return BSC_EINVAL;
  }
  BSC_LAUNCH_CHECK("rows_to_f16_kernel");
  return BSC_OK;
}

// BSC_MODE_TC16X3 operand preparation: hl[i] = [hi | lo] (2 d fp16) with hi = fp16(x 2^k), lo = fp16(x 2^k - hi), i.e.
// hi + lo carries 22 significant bits of every element; the row maximum lands in [128, 256) -- fp16 keeps the lo part
// (about 2^-12 of the element) a NORMAL number for every element down to 2^-10 of the row maximum, and products of
// two scaled rows (< 2^16 each, 2^26 per row) are far inside fp32.  inv[i] = 2^-k / max(||x_i||, 1e-12) as in TC16.
constexpr int kX3RowExp = 7;

template <typename TI>
__global__ void __launch_bounds__(256) rows_to_f16x3_kernel(const TI* __restrict__ x, int64_t n, int d, int64_t ldx,
                                                            __half* __restrict__ hl, float* __restrict__ inv) {
  const int lane = threadIdx.x & 31;
  const int64_t row = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= n) return;
  const TI* xr = x + row * ldx;
  float amax = 0.f;
  for (int c = lane; c < d; c += 32) amax = fmaxf(amax, fabsf(to_f32<TI>(xr[c])));
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, o));
  int k = 0;
  if (amax > 0.f && amax < 3.0e38f) k = kX3RowExp - ilogbf(amax);
  k = max(-120, min(120, k));
  const float scale = exp2f((float)k);
  // norm of the SCALED row (never overflows or flushes), sum of squares in double: the inverse norm multiplies every
  // logit of the row, so its rounding is a systematic error of that row
  double s2 = 0.0;
  for (int c = lane; c < d; c += 32) {
    const double v = (double)(to_f32<TI>(xr[c]) * scale);
    s2 = fma(v, v, s2);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) s2 += __shfl_xor_sync(0xffffffffu, s2, o);
  const double n2 = sqrt(s2);                             // ||x|| * 2^k
  const double true_nrm = n2 * exp2((double)-k);
  const float inv_row = n2 > 0.0 ? (float)(true_nrm >= 1e-12 ? 1.0 / n2 : exp2((double)-k) * 1e12) : 0.f;
  if (lane == 0) inv[row] = inv_row;
  __half* out = hl + row * (int64_t)(2 * d);
  for (int c = lane; c < d; c += 32) {
    const float v = to_f32<TI>(xr[c]) * scale;
    const __half h = __float2half_rn(v);
    out[c] = h;
    out[d + c] = __float2half_rn(v - __half2float(h));
  }
}

int rows_to_f16x3_launch(const void* x, int x_dtype, int64_t n, int d, int64_t ldx, void* hl, float* inv,
                         cudaStream_t st) {
  if (n == 0) return BSC_OK;
  dim3 block(256), grid((unsigned)((n + 7) / 8));
  switch (x_dtype) {
    case BSC_F32:
      rows_to_f16x3_kernel<float><<<grid, block, 0, st>>>((const float*)x, n, d, ldx, (__half*)hl, inv);
      break;
    case BSC_BF16:
      rows_to_f16x3_kernel<__nv_bfloat16><<<grid, block, 0, st>>>((const __nv_bfloat16*)x, n, d, ldx, (__half*)hl, inv);
      break;
    case BSC_F16:
      rows_to_f16x3_kernel<__half><<<grid, block, 0, st>>>((const __half*)x, n, d, ldx, (__half*)hl, inv);
      break;
    default:
      set_error("bsc_rows_to_f16x3: unknown x_dtype %d", x_dtype);
      return BSC_EINVAL;
  }
  BSC_LAUNCH_CHECK("rows_to_f16x3_kernel");
  return BSC_OK;
}

// fp16 unit rows of rows [row0, row0 + n) from the scaled fp16 rows: out[l] = half(x16[row0 + l] * inv[row0 + l]).
// B operand of the MN-major gradient GEMMs (tc_bwd3.cu).  blockIdx.y = modality (absent ones have a null source).
// 8 elements (16 bytes) per thread and pass.
struct UnitRowsArgs {
  const uint4* x[3];
  const float* inv[3];
  uint4* out[3];
};

__global__ void __launch_bounds__(256) rows_unit_f16_kernel(const UnitRowsArgs a, int64_t row0, int64_t n, int d8) {
  const uint4* __restrict__ x = a.x[blockIdx.y];
  if (x == nullptr) return;
  const float* __restrict__ inv = a.inv[blockIdx.y] + row0;
  uint4* __restrict__ out = a.out[blockIdx.y];
  x += row0 * d8;
  const int64_t total = n * (int64_t)d8;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const float s = inv[i / d8];
    uint4 v = x[i];
    __half2* h = reinterpret_cast<__half2*>(&v);
#pragma unroll
    for (int t = 0; t < 4; ++t) {
      const float2 f = __half22float2(h[t]);
      h[t] = __floats2half2_rn(f.x * s, f.y * s);
    }
    out[i] = v;
  }
}

int rows_unit_f16_launch3(const void* const x16[3], const float* const inv[3], int64_t row0, int64_t n, int d,
                          void* const out[3], cudaStream_t st) {
  if (n == 0) return BSC_OK;
  BSC_REQUIRE(d % 8 == 0, BSC_EUNSUPPORTED, "rows_unit_f16: d %% 8 != 0");
  UnitRowsArgs a{};
  bool any = false;
  for (int m = 0; m < 3; ++m) {
    a.x[m] = out[m] ? (const uint4*)x16[m] : nullptr;
    a.inv[m] = inv[m];
    a.out[m] = (uint4*)out[m];
    any = any || a.x[m] != nullptr;
  }
  if (!any) return BSC_OK;
  const int64_t total = n * (int64_t)(d / 8);
  const unsigned gx = (unsigned)((total + 255) / 256 > 148 * 8 ? 148 * 8 : (total + 255) / 256);
  rows_unit_f16_kernel<<<dim3(gx, 3), 256, 0, st>>>(a, row0, n, d / 8);
  BSC_LAUNCH_CHECK("rows_unit_f16_kernel");
  return BSC_OK;
}

// BSC_MODE_TC16X3: unit rows of rows [row0, row0 + n) as hi | lo fp16 parts of kX3UnitScale * (hi + lo) * inv, from the
// [rows][2 d] hi | lo operand rows.  One thread per 8 elements.
__global__ void __launch_bounds__(256) rows_unit_f16x3_kernel(const UnitRowsArgs a, int64_t row0, int64_t n, int d8) {
  const uint4* __restrict__ x = a.x[blockIdx.y];
  if (x == nullptr) return;
  const float* __restrict__ inv = a.inv[blockIdx.y] + row0;
  uint4* __restrict__ out = a.out[blockIdx.y];
  x += row0 * 2 * d8;
  const int64_t total = n * (int64_t)d8;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t row = i / d8;
    const int c8 = (int)(i - row * d8);
    const float s = inv[row] * kX3UnitScale;
    uint4 vh = x[row * 2 * d8 + c8], vl = x[row * 2 * d8 + d8 + c8];
    __half2* h = reinterpret_cast<__half2*>(&vh);
    __half2* l = reinterpret_cast<__half2*>(&vl);
#pragma unroll
    for (int t = 0; t < 4; ++t) {
      const float2 fh = __half22float2(h[t]), fl = __half22float2(l[t]);
      const float ux = (fh.x + fl.x) * s, uy = (fh.y + fl.y) * s;  // hi + lo is exact in fp32 (22 bits)
      const __half2 uh = __floats2half2_rn(ux, uy);
      const float2 back = __half22float2(uh);
      h[t] = uh;
      l[t] = __floats2half2_rn(ux - back.x, uy - back.y);
    }
    out[row * 2 * d8 + c8] = vh;
    out[row * 2 * d8 + d8 + c8] = vl;
  }
}

int rows_unit_f16x3_launch3(const void* const hl16[3], const float* const inv[3], int64_t row0, int64_t n, int d,
                            void* const out[3], cudaStream_t st) {
  if (n == 0) return BSC_OK;
  BSC_REQUIRE(d % 8 == 0, BSC_EUNSUPPORTED, "rows_unit_f16x3: d %% 8 != 0");
  UnitRowsArgs a{};
  bool any = false;
  for (int m = 0; m < 3; ++m) {
    a.x[m] = out[m] ? (const uint4*)hl16[m] : nullptr;
    a.inv[m] = inv[m];
    a.out[m] = (uint4*)out[m];
    any = any || a.x[m] != nullptr;
  }
  if (!any) return BSC_OK;
  const int64_t total = n * (int64_t)(d / 8);
  const unsigned gx = (unsigned)((total + 255) / 256 > 148 * 8 ? 148 * 8 : (total + 255) / 256);
  rows_unit_f16x3_kernel<<<dim3(gx, 3), 256, 0, st>>>(a, row0, n, d / 8);
  BSC_LAUNCH_CHECK("rows_unit_f16x3_kernel");
  return BSC_OK;
}

// ---------------------------------------------------------------------------------------------
// Fused TC16 operand preparation + all-gather over NVLink peer memory (gather_features, loss_func.py:58-91, for the
// default gather_with_grad=False where only values travel).  Every rank holds a symmetric buffer
//     xh[3][N][d] fp16 | inv[3][N] fp32 | labels[N] int64            (offsets: gather_layout)
// mapped into all peers.  One warp per (modality, local row) computes what rows_to_f16_kernel computes and stores the
// fp16 row, its inverse-norm scale and (modality 0's warps) the row's label at global row row0 + l of EVERY rank's
// buffer with 16-byte peer stores, so the gather costs one kernel instead of seven NCCL calls.  Consumers on a peer
// may read only after a cross-rank barrier that follows this kernel on every rank's stream.
// ---------------------------------------------------------------------------------------------
struct GatherPushArgs {
  const void* x[3];   // local rows per modality (null = absent)
  int64_t ldx[3];
  const int64_t* labels;  // local labels (null: they travelled with bsc_gather_push_labels)
  char* peer[kMaxPeers];  // base of every rank's symmetric buffer (index = rank)
  char* mc;               // NVSwitch multicast address of the same buffer (one store reaches every rank), or null
  int world, rank;
  int ordered;            // 1: own buffer only (the DMA ring forwards the rows), 0: every rank's buffer
  int blocks_per_dest;
  int64_t n_local, n_global, row0;
  int d;
  size_t off_xh[3], off_inv[3], off_labels, off_flags;
};

// One store to the multicast address is replicated by the NVSwitch into every rank's buffer.  It does not shorten the
// gather of the rows themselves -- every GPU still has to RECEIVE (W-1)/W of the global batch, 132 MB at 8 x 4096 x 768 x
// 3 fp16, 0.22 ms of NVLink ingress however the bytes are sent (measured: the same time as one store per peer) -- but it
// makes the small exchanges (labels, statistics records) one store instead of W.
__device__ __forceinline__ void mc_store16(void* addr, uint4 v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "r"(v.x), "r"(v.y),
               "r"(v.z), "r"(v.w)
               : "memory");
}
__device__ __forceinline__ void mc_store8(void* addr, uint2 v) {
  asm volatile("multimem.st.relaxed.sys.global.v2.f32 [%0], {%1, %2};" ::"l"(addr), "r"(v.x), "r"(v.y) : "memory");
}
__device__ __forceinline__ void mc_store4(void* addr, uint32_t v) {
  asm volatile("multimem.st.relaxed.sys.global.f32 [%0], %1;" ::"l"(addr), "r"(v) : "memory");
}

// One warp per (modality, local row).  self_only = 0: the row goes into EVERY rank's buffer (or once through the
// multicast address); the caller runs a cross-rank barrier before anyone reads.  self_only = 1: the row goes into this
// rank's own buffer only -- the first step of the DMA ring below.
template <typename TI>
__global__ void __launch_bounds__(256) gather_push_kernel(const GatherPushArgs a) {
  const int lane = threadIdx.x & 31;
  const int dest = a.ordered ? a.rank : -1;
  const int64_t w = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int m = (int)(w / a.n_local);
  const int64_t l = w % a.n_local;
  if (m >= 3 || a.x[m] == nullptr) return;
  const int d = a.d;
  const TI* xr = (const TI*)a.x[m] + l * a.ldx[m];
  float ss = 0.f, amax = 0.f;
  for (int c = lane; c < d; c += 32) {
    const float v = to_f32<TI>(xr[c]);
    ss = fmaf(v, v, ss);
    amax = fmaxf(amax, fabsf(v));
  }
  ss = warp_sum(ss);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, o));
  int k = 0;
  if (amax > 0.f && amax < 3.0e38f) k = -ilogbf(amax);
  k = max(-120, min(120, k));
  const float scale = exp2f((float)k);
  const float nrm = sqrtf(ss);
  float inv_row;
  if (nrm > 1e-12f && nrm < 3.0e38f) {
    inv_row = exp2f((float)-k) / nrm;
  } else {
    float s2 = 0.f;
    for (int c = lane; c < d; c += 32) {
      const float v = to_f32<TI>(xr[c]) * scale;
      s2 = fmaf(v, v, s2);
    }
    s2 = warp_sum(s2);
    const float n2 = sqrtf(s2);
    const float true_nrm = n2 * exp2f((float)-k);
    inv_row = n2 > 0.f ? (true_nrm >= 1e-12f ? 1.0f / n2 : exp2f((float)-k) * 1e12f) : 0.f;
  }
  const int64_t g = a.row0 + l;  // global row
  // 8 fp16 per lane per pass (d % 8 == 0 is checked by the launcher): one 16-byte store per destination
  for (int c0 = lane * 8; c0 < d; c0 += 256) {
    __half hv[8];
#pragma unroll
    for (int t = 0; t < 8; ++t) hv[t] = __float2half_rn(to_f32<TI>(xr[c0 + t]) * scale);
    const uint4 pk = *reinterpret_cast<const uint4*>(hv);
    const size_t off = a.off_xh[m] + ((size_t)g * d + c0) * 2;
    if (dest >= 0) {
      *reinterpret_cast<uint4*>(a.peer[dest] + off) = pk;
    } else if (a.mc != nullptr) {
      mc_store16(a.mc + off, pk);
    } else {
      for (int r = 0; r < a.world; ++r) *reinterpret_cast<uint4*>(a.peer[r] + off) = pk;
    }
  }
  // the label of the row travels with the first present modality's warp
  int first = 0;
  while (first < 3 && a.x[first] == nullptr) ++first;
  const bool with_label = m == first && a.labels != nullptr;
  if (dest >= 0) {
    if (lane == 0) *reinterpret_cast<float*>(a.peer[dest] + a.off_inv[m] + (size_t)g * 4) = inv_row;
    if (lane == 1 && with_label) *reinterpret_cast<int64_t*>(a.peer[dest] + a.off_labels + (size_t)g * 8) = a.labels[l];
  } else if (a.mc != nullptr) {
    if (lane == 0) mc_store4(a.mc + a.off_inv[m] + (size_t)g * 4, __float_as_uint(inv_row));
    if (lane == 1 && with_label) {
      const int64_t lab = a.labels[l];
      mc_store8(a.mc + a.off_labels + (size_t)g * 8, make_uint2((uint32_t)lab, (uint32_t)((uint64_t)lab >> 32)));
    }
  } else {
    if (lane < a.world) *reinterpret_cast<float*>(a.peer[lane] + a.off_inv[m] + (size_t)g * 4) = inv_row;
    if (with_label && lane < a.world)
      *reinterpret_cast<int64_t*>(a.peer[lane] + a.off_labels + (size_t)g * 8) = a.labels[l];
  }
}

// Publishes this rank's own rows: arrive[rank] = epoch in its OWN buffer (the value the DMA ring then copies into the
// peers' buffers as their arrival flag).  Runs on the stream of the preparation kernel, i.e. BEFORE the sweep: nothing
// of the ring needs an SM later -- the persistent sweep owns every register of every SM (512 threads x 128 registers),
// a kernel queued beside it would never be scheduled, and a sweep waiting for that kernel's flag would never end.
__global__ void gather_signal_kernel(unsigned long long* flag, unsigned long long epoch) {
  if (threadIdx.x == 0) asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(flag), "l"(epoch) : "memory");
}

// labels only: tiny, every rank needs ALL labels before its label statistics can start
__global__ void __launch_bounds__(256) gather_labels_kernel(const int64_t* __restrict__ labels, int64_t n_local,
                                                            int64_t row0, GatherPushArgs a) {
  const int64_t l = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (l >= n_local) return;
  const int64_t lab = labels[l];
  const size_t off = a.off_labels + (size_t)(row0 + l) * 8;
  if (a.mc != nullptr) {
    mc_store8(a.mc + off, make_uint2((uint32_t)lab, (uint32_t)((uint64_t)lab >> 32)));
  } else {
    for (int r = 0; r < a.world; ++r) *reinterpret_cast<int64_t*>(a.peer[r] + off) = lab;
  }
}

void gather_layout(int64_t n_global, int d, int world, size_t off_xh[3], size_t off_inv[3], size_t* off_labels,
                   size_t* off_stats, size_t* off_flags, size_t* total) {
  const size_t ax = align_up((size_t)n_global * d * 2, 256);
  const size_t ai = align_up((size_t)n_global * 4, 256);
  for (int m = 0; m < 3; ++m) {
    off_xh[m] = (size_t)m * ax;
    off_inv[m] = 3 * ax + (size_t)m * ai;
  }
  *off_labels = 3 * ax + 3 * ai;
  *off_stats = *off_labels + align_up((size_t)n_global * 8, 256);
  *off_flags = *off_stats + align_up((size_t)(world > 0 ? world : 1) * stats_record_len(n_global) * sizeof(float), 256);
  *total = *off_flags + align_up((size_t)kMaxPeers * 8, 256);
}

int gather_blocks_per_dest(int64_t n_local) { return (int)((3 * n_local + 7) / 8); }

static void gather_fill(GatherPushArgs& a, int64_t n_local, int64_t row0, int64_t n_global, int d, void* const* peers,
                        void* multicast, int world, int rank) {
  for (int r = 0; r < world; ++r) a.peer[r] = (char*)peers[r];
  a.mc = (char*)multicast;
  a.world = world;
  a.rank = rank;
  a.n_local = n_local;
  a.n_global = n_global;
  a.row0 = row0;
  a.d = d;
  size_t off_stats, total;
  gather_layout(n_global, d, world, a.off_xh, a.off_inv, &a.off_labels, &off_stats, &a.off_flags, &total);
}

int gather_labels_launch(const int64_t* labels, int64_t n_local, int64_t row0, int64_t n_global, int d,
                         void* const* peers, void* multicast, int world, cudaStream_t st) {
  if (n_local == 0) return BSC_OK;
  GatherPushArgs a{};
  gather_fill(a, n_local, row0, n_global, d, peers, multicast, world, 0);
  gather_labels_kernel<<<(unsigned)((n_local + 255) / 256), 256, 0, st>>>(labels, n_local, row0, a);
  BSC_LAUNCH_CHECK("gather_labels_kernel");
  return BSC_OK;
}

int gather_push_launch(const void* const x[3], int x_dtype, const int64_t ldx[3], const int64_t* labels,
                       int64_t n_local, int64_t row0, int64_t n_global, int d, void* const* peers, void* multicast,
                       int world, int rank, int ordered, uint64_t epoch, cudaStream_t st) {
  if (n_local == 0) return BSC_OK;
  GatherPushArgs a{};
  unsigned mask = 0;
  for (int m = 0; m < 3; ++m) {
    a.x[m] = x[m];
    a.ldx[m] = ldx[m];
    if (x[m]) mask |= 1u << m;
  }
  a.labels = labels;
  gather_fill(a, n_local, row0, n_global, d, peers, multicast, world, rank);
  a.ordered = ordered;
  a.blocks_per_dest = gather_blocks_per_dest(n_local);
  dim3 block(256), grid((unsigned)a.blocks_per_dest);
  switch (x_dtype) {
    case BSC_F32: gather_push_kernel<float><<<grid, block, 0, st>>>(a); break;
    case BSC_BF16: gather_push_kernel<__nv_bfloat16><<<grid, block, 0, st>>>(a); break;
    case BSC_F16: gather_push_kernel<__half><<<grid, block, 0, st>>>(a); break;
    default: set_error("bsc_gather_push: unknown x_dtype %d", x_dtype); return BSC_EINVAL;
  }
  BSC_LAUNCH_CHECK("gather_push_kernel");
  if (!ordered) return BSC_OK;
  unsigned long long* flag = reinterpret_cast<unsigned long long*>(a.peer[rank] + a.off_flags) + rank;
  gather_signal_kernel<<<1, 32, 0, st>>>(flag, (unsigned long long)epoch);
  BSC_LAUNCH_CHECK("gather_signal_kernel");
  return BSC_OK;
}

// The DMA ring of the fused all-gather -> similarity sweep: copy engines only, no kernel.  For destination rank+1,
// rank+2, ...: the fp16 rows of every present modality, their inverse norms, then the arrival flag (8 bytes copied from
// this rank's own flag, which holds the epoch) -- copies of one stream complete in order, so a peer that sees the flag
// sees the rows.  Every receiver hears from one sender at a time, in the order rank-1, rank-2, ... in which its sweep
// takes the sources.  The rows travel by DMA because SM stores that back up behind a saturated NVLink port stall every
// memory operation of their SM (measured on 2 GPUs: label statistics 46 -> 209 us, sweep +4 % beside a store-based push).
int gather_ring_launch(void* const* peers, int world, int rank, int64_t n_local, int64_t row0, int64_t n_global, int d,
                       unsigned present_mask, cudaStream_t st) {
  if (n_local == 0 || world <= 1) return BSC_OK;
  GatherPushArgs a{};
  gather_fill(a, n_local, row0, n_global, d, peers, nullptr, world, rank);
  const char* own = a.peer[rank];
  const size_t row_bytes = (size_t)d * 2;
  for (int k = 1; k < world; ++k) {
    char* dst = a.peer[(rank + k) % world];
    for (int m = 0; m < 3; ++m)
      if ((present_mask >> m) & 1) {
        const size_t off = a.off_xh[m] + (size_t)row0 * row_bytes;
        BSC_CUDA(cudaMemcpyAsync(dst + off, own + off, (size_t)n_local * row_bytes, cudaMemcpyDefault, st));
      }
    for (int m = 0; m < 3; ++m)
      if ((present_mask >> m) & 1) {
        const size_t off = a.off_inv[m] + (size_t)row0 * 4;
        BSC_CUDA(cudaMemcpyAsync(dst + off, own + off, (size_t)n_local * 4, cudaMemcpyDefault, st));
      }
    const size_t foff = a.off_flags + (size_t)rank * 8;
    BSC_CUDA(cudaMemcpyAsync(dst + foff, own + foff, 8, cudaMemcpyDefault, st));
  }
  return BSC_OK;
}

// float64 rows (the reference's eval features, inference_epoch.py:67): sum of squares, square root and the quotient in
// float64, ONE rounding to the output type -- normalize(x).astype(float32) of inference_and_eval.py:416-417.
template <typename TO>
__global__ void __launch_bounds__(256) l2norm_rows_f64_kernel(const double* __restrict__ x, int64_t n, int d,
                                                              int64_t ldx, TO* __restrict__ y16,
                                                              float* __restrict__ y32, float* __restrict__ inv_norm,
                                                              int zero_rule) {
  const int lane = threadIdx.x & 31;
  const int64_t row = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= n) return;
  const double* xr = x + row * ldx;
  double ss = 0.0;
  for (int c = lane; c < d; c += 32) ss = fma(xr[c], xr[c], ss);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, o);
  const double nrm = sqrt(ss);
  double div;
  if (zero_rule == BSC_NORM_SKLEARN)
    div = (nrm < 2.220446049250313e-15) ? 1.0 : nrm;  // 10 * eps(float64): the row is left as it is
  else
    div = fmax(nrm, 1e-12);
  if (inv_norm != nullptr && lane == 0) inv_norm[row] = (float)(1.0 / div);
  for (int c = lane; c < d; c += 32) {
    const float v = (float)(xr[c] / div);
    if (y16 != nullptr) y16[row * (int64_t)d + c] = from_f32<TO>(v);
    if (y32 != nullptr) y32[row * (int64_t)d + c] = v;
  }
}

template <typename TI>
static int l2norm_dispatch_out(const void* x, int64_t n, int d, int64_t ldx, void* y16, int y16_dtype, float* y32,
                               float* inv_norm, int zero_rule, cudaStream_t st) {
  dim3 block(256), grid((unsigned)((n + 7) / 8));
  if (y16 == nullptr || y16_dtype == BSC_F16) {
    l2norm_rows_kernel<TI, __half><<<grid, block, 0, st>>>((const TI*)x, n, d, ldx, (__half*)y16, y32, inv_norm,
                                                          zero_rule);
  } else if (y16_dtype == BSC_BF16) {
    l2norm_rows_kernel<TI, __nv_bfloat16><<<grid, block, 0, st>>>((const TI*)x, n, d, ldx, (__nv_bfloat16*)y16,
                                                                 y32, inv_norm, zero_rule);
  } else {
    set_error("bsc_l2norm_fwd: y16_dtype must be BSC_F16 or BSC_BF16");
    return BSC_EINVAL;
  }
  BSC_LAUNCH_CHECK("l2norm_rows_kernel");
  return BSC_OK;
}

int l2norm_launch(const void* x, int x_dtype, int64_t n, int d, int64_t ldx, void* y16, int y16_dtype, float* y32,
                  float* inv_norm, int zero_rule, cudaStream_t st) {
  if (n == 0) return BSC_OK;
  if (x_dtype == BSC_F64) {
    dim3 block(256), grid((unsigned)((n + 7) / 8));
    if (y16 != nullptr && y16_dtype == BSC_BF16)
      l2norm_rows_f64_kernel<__nv_bfloat16><<<grid, block, 0, st>>>((const double*)x, n, d, ldx, (__nv_bfloat16*)y16,
                                                                    y32, inv_norm, zero_rule);
    else
      l2norm_rows_f64_kernel<__half><<<grid, block, 0, st>>>((const double*)x, n, d, ldx, (__half*)y16, y32, inv_norm,
                                                             zero_rule);
    BSC_LAUNCH_CHECK("l2norm_rows_f64_kernel");
    return BSC_OK;
  }
  switch (x_dtype) {
    case BSC_F32: return l2norm_dispatch_out<float>(x, n, d, ldx, y16, y16_dtype, y32, inv_norm, zero_rule, st);
    case BSC_BF16:
      return l2norm_dispatch_out<__nv_bfloat16>(x, n, d, ldx, y16, y16_dtype, y32, inv_norm, zero_rule, st);
    case BSC_F16: return l2norm_dispatch_out<__half>(x, n, d, ldx, y16, y16_dtype, y32, inv_norm, zero_rule, st);
  }
  set_error("bsc_l2norm_fwd: unknown x_dtype %d", x_dtype);
  return BSC_EINVAL;
}

// ---------------------------------------------------------------------------------------------
// K2: label statistics without an N x N matrix and (for the reference's default labels) without N^2 work.
//   1. every row inserts its label into an open-addressing table (slot = index of the first row that claimed it);
//      the claimer is the class "head".  count[head] += 1, minidx[head] = min(minidx[head], i): both results are
//      independent of the order in which the atomics land.
//   2. r[i] = count[head_i], rep[i] = minidx[head_i], rmax = max r.
//   3. next[i] = min{j > i : label_j == label_i}: only row blocks that contain a row with r > 1 scan the labels
//      (tiles staged in shared memory); with labels = arange (util/dataset.py:147) no block does.
// ---------------------------------------------------------------------------------------------
constexpr int kLabelTile = 2048;

__device__ __forceinline__ uint32_t label_hash(int64_t v) {
  uint64_t x = (uint64_t)v;
  x ^= x >> 33;
  x *= 0xff51afd7ed558ccdull;
  x ^= x >> 33;
  x *= 0xc4ceb9fe1a85ec53ull;
  x ^= x >> 33;
  return (uint32_t)x;
}

__global__ void __launch_bounds__(256) label_insert_kernel(const int64_t* __restrict__ labels, int64_t n, int* slots,
                                                           uint32_t mask, int* __restrict__ head, int* count,
                                                           int* minidx) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int64_t key = labels[i];
  uint32_t h = label_hash(key) & mask;
  int hd;
  for (;;) {
    int cur = ((volatile int*)slots)[h];
    if (cur < 0) {
      const int old = atomicCAS(&slots[h], -1, (int)i);
      if (old < 0) {
        hd = (int)i;
        break;
      }
      cur = old;
    }
    if (labels[cur] == key) {
      hd = cur;
      break;
    }
    h = (h + 1) & mask;
  }
  head[i] = hd;
  atomicAdd(&count[hd], 1);
  atomicMin(&minidx[hd], (int)i);
}

__global__ void __launch_bounds__(256) label_finish_kernel(int64_t n, const int* __restrict__ head,
                                                           const int* __restrict__ count, const int* __restrict__ minidx,
                                                           float* __restrict__ r, int32_t* __restrict__ rep,
                                                           int* __restrict__ rmax_bits) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  float c = 0.f;
  if (i < n) {
    const int hd = head[i];
    c = (float)count[hd];
    r[i] = c;
    rep[i] = minidx[hd];
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) c = fmaxf(c, __shfl_xor_sync(0xffffffffu, c, o));
  // positive floats order like their bit patterns
  if ((threadIdx.x & 31) == 0) atomicMax(rmax_bits, __float_as_int(c));
}

// 32 rows per block, 8 threads per row; a block whose rows are all alone in their class has nothing to look for.
__global__ void __launch_bounds__(256) label_next_kernel(const int64_t* __restrict__ labels, int64_t n,
                                                         const float* __restrict__ r, int32_t* __restrict__ next) {
  __shared__ int64_t tile[kLabelTile];
  const int sub = threadIdx.x & 7;
  const int64_t i = (int64_t)blockIdx.x * 32 + (threadIdx.x >> 3);
  const bool multi = i < n && r[i] > 1.5f;
  if (!__syncthreads_or(multi ? 1 : 0)) {
    if (i < n && sub == 0) next[i] = -1;
    return;
  }
  const int64_t mine = (i < n) ? labels[i] : 0;
  constexpr int kNone = 0x7fffffff;
  int nxt = kNone;
  // members after row i can only live in tiles at or after row i's own tile
  const int64_t base0 = ((int64_t)blockIdx.x * 32 / kLabelTile) * kLabelTile;
  for (int64_t base = base0; base < n; base += kLabelTile) {
    const int len = (int)min((int64_t)kLabelTile, n - base);
    __syncthreads();
    for (int t = threadIdx.x; t < len; t += blockDim.x) tile[t] = labels[base + t];
    __syncthreads();
    if (multi) {
#pragma unroll 8
      for (int t = sub; t < len; t += 8) {
        const int j = (int)(base + t);
        if (tile[t] == mine && j > (int)i) nxt = min(nxt, j);
      }
    }
    // all members found so far are minima of their sub-sequence; a hit in this tile ends the search for the row
    int found = nxt != kNone ? 1 : 0;
#pragma unroll
    for (int o = 1; o < 8; o <<= 1) found |= __shfl_xor_sync(0xffffffffu, found, o);
    if (__syncthreads_and((!multi || found) ? 1 : 0)) break;
  }
#pragma unroll
  for (int o = 1; o < 8; o <<= 1) nxt = min(nxt, __shfl_xor_sync(0xffffffffu, nxt, o));
  if (i < n && sub == 0) next[i] = nxt == kNone ? -1 : nxt;
}

static uint32_t label_table_slots(int64_t n) {
  uint32_t m = 1024;
  while ((int64_t)m < 2 * n) m <<= 1;
  return m;
}

size_t label_stats_ws_bytes(int64_t n) {
  // slots[m] | head[n] | count[n] | minidx[n], all int32
  return align_up((size_t)label_table_slots(n) * 4, 256) + 3 * align_up((size_t)(n > 0 ? n : 1) * 4, 256);
}

int label_stats_launch(const int64_t* labels, int64_t n, float* r, int32_t* rep, int32_t* next, float* rmax,
                       char* ws, cudaStream_t st) {
  BSC_CUDA(cudaMemsetAsync(rmax, 0, sizeof(float), st));
  if (n == 0) return BSC_OK;
  const uint32_t m = label_table_slots(n);
  const size_t nb = align_up((size_t)n * 4, 256);
  int* slots = (int*)ws;
  int* head = (int*)(ws + align_up((size_t)m * 4, 256));
  int* count = (int*)((char*)head + nb);
  int* minidx = (int*)((char*)count + nb);
  BSC_CUDA(cudaMemsetAsync(slots, 0xff, (size_t)m * 4, st));   // -1 = empty
  BSC_CUDA(cudaMemsetAsync(count, 0, (size_t)n * 4, st));
  BSC_CUDA(cudaMemsetAsync(minidx, 0x7f, (size_t)n * 4, st));  // 0x7f7f7f7f > any row index
  const unsigned grid = (unsigned)((n + 255) / 256);
  label_insert_kernel<<<grid, 256, 0, st>>>(labels, n, slots, m - 1, head, count, minidx);
  BSC_LAUNCH_CHECK("label_insert_kernel");
  label_finish_kernel<<<grid, 256, 0, st>>>(n, head, count, minidx, r, rep, (int*)rmax);
  BSC_LAUNCH_CHECK("label_finish_kernel");
  label_next_kernel<<<(unsigned)((n + 31) / 32), 256, 0, st>>>(labels, n, r, next);
  BSC_LAUNCH_CHECK("label_next_kernel");
  return BSC_OK;
}

// ---------------------------------------------------------------------------------------------
// Label-class range of every column tile: out[t] = {min, max} of rep over rows [t*tile, (t+1)*tile).
// The sweeps use it to skip the label compare on tiles that cannot contain a positive pair.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) rep_range_kernel(const int32_t* __restrict__ rep, int64_t n, int tile,
                                                        int2* __restrict__ out) {
  __shared__ int smin[8], smax[8];
  const int64_t base = (int64_t)blockIdx.x * tile;
  int lo = 0x7fffffff, hi = -1;
  for (int t = threadIdx.x; t < tile; t += blockDim.x) {
    const int64_t j = base + t;
    if (j < n) {
      const int v = rep[j];
      lo = min(lo, v);
      hi = max(hi, v);
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    lo = min(lo, __shfl_xor_sync(0xffffffffu, lo, o));
    hi = max(hi, __shfl_xor_sync(0xffffffffu, hi, o));
  }
  if ((threadIdx.x & 31) == 0) {
    smin[threadIdx.x >> 5] = lo;
    smax[threadIdx.x >> 5] = hi;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w = 1; w < (int)(blockDim.x >> 5); ++w) {
      lo = min(lo, smin[w]);
      hi = max(hi, smax[w]);
    }
    out[blockIdx.x] = make_int2(lo, hi);
  }
}

int rep_range_launch(const int32_t* rep, int64_t n, int tile, int2* out, cudaStream_t st) {
  if (n == 0) return BSC_OK;
  rep_range_kernel<<<(unsigned)((n + tile - 1) / tile), 256, 0, st>>>(rep, n, tile, out);
  BSC_LAUNCH_CHECK("rep_range_kernel");
  return BSC_OK;
}

// Scalar reductions run on ONE cluster of 8 CTAs x 1024 threads (a single block spent 34 - 56 us streaming N-sized
// vectors, a fixed cost of every step however many ranks share the batch).  Every thread accumulates a fixed subset in
// double, every CTA tree-reduces in shared memory and stores its partial into CTA 0's shared memory over DSMEM; CTA 0
// adds the 8 partials in rank order: the result does not depend on scheduling.
constexpr int kRedCtas = 8;

__device__ __forceinline__ double cluster_sum_to_cta0(double acc, double* sh /*[1024]*/, double* slots /*[kRedCtas]*/) {
  namespace cg = cooperative_groups;
  cg::cluster_group cluster = cg::this_cluster();
  sh[threadIdx.x] = acc;
  __syncthreads();
  for (int s = 512; s > 0; s >>= 1) {
    if ((int)threadIdx.x < s) sh[threadIdx.x] += sh[threadIdx.x + s];
    __syncthreads();
  }
  if (threadIdx.x == 0) cluster.map_shared_rank(slots, 0)[cluster.block_rank()] = sh[0];
  cluster.sync();
  double total = 0.0;
  if (cluster.block_rank() == 0 && threadIdx.x == 0)
    for (int c = 0; c < kRedCtas; ++c) total += slots[c];
  return total;  // meaningful in thread 0 of CTA 0
}

// out[0] = scale * sum_p sum_{i<n} part[p*stride + i]  (double accumulation, fixed order).
__global__ void __cluster_dims__(kRedCtas, 1, 1) __launch_bounds__(1024)
    sum_rows_kernel(const float* __restrict__ part, int nparts, int64_t part_stride, int64_t n, float scale,
                    float* __restrict__ out, double* __restrict__ out64) {
  __shared__ double sh[1024];
  __shared__ double slots[kRedCtas];
  double acc = 0.0;
  // work items = (part row, 1024-column chunk), dealt round-robin to the CTAs
  const int64_t nchunk = (n + 1023) / 1024;
  const int64_t items = (int64_t)nparts * nchunk;
#pragma unroll 4
  for (int64_t q = blockIdx.x; q < items; q += kRedCtas) {
    const int64_t p = q / nchunk, i = (q - p * nchunk) * 1024 + threadIdx.x;
    if (i < n) acc += (double)part[p * part_stride + i];
  }
  const double total = cluster_sum_to_cta0(acc, sh, slots);
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    out[0] = (float)((double)scale * total);
    if (out64 != nullptr) out64[0] = (double)scale * total;
  }
}

int sum_rows_launch(const float* part, int nparts, int64_t part_stride, int64_t n, float scale, float* out,
                    double* out64, cudaStream_t st) {
  sum_rows_kernel<<<kRedCtas, 1024, 0, st>>>(part, nparts, part_stride, n, scale, out, out64);
  BSC_LAUNCH_CHECK("sum_rows_kernel");
  return BSC_OK;
}

// ---------------------------------------------------------------------------------------------
// Partial-sum reduction of the sweeps: stats[pid(a,b)][row0+i] = sum_chunk rowpart, stats[pid(b,a)][j] =
// sum_iblk colpart, in a fixed order.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) reduce_parts_kernel(const float* __restrict__ part, int nparts,
                                                           int64_t part_stride, int64_t n,
                                                           float* __restrict__ out) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  float acc = 0.f;
  for (int p = 0; p < nparts; ++p) acc += part[p * part_stride + i];
  out[i] = acc;
}

// Up to 6 reductions in one launch (blockIdx.y = job): the forward's row and column sum-exp vectors of every pair.
// blockIdx.y == jobs.count: the positive-pair term as kPosParts double partial sums (rows are dealt to the parts
// round-robin in chunks of 256, so the result does not depend on the grid).  Every value is also stored into the
// records of the other ranks (statistics exchange over NVLink peer memory: one multicast store or one store per peer).
__device__ __forceinline__ void mirror_f32(const ReduceJobs& jobs, float* local_addr, float v) {
  const int64_t off = local_addr - jobs.local_base;
  if (jobs.mc_base != nullptr) {
    asm volatile("multimem.st.relaxed.sys.global.f32 [%0], %1;" ::"l"(jobs.mc_base + off), "f"(v) : "memory");
  } else {
    for (int r = 0; r < jobs.n_peers; ++r) jobs.peer_base[r][off] = v;
  }
}

__global__ void __launch_bounds__(256) reduce_parts_batch_kernel(const ReduceJobs jobs) {
  if ((int)blockIdx.y < jobs.count) {
    const ReduceJob j = jobs.job[blockIdx.y];
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= j.n) return;
    float acc = 0.f;
    for (int p = 0; p < j.nparts; ++p) acc += j.part[p * j.part_stride + i];
    j.out[i] = acc;
    if (jobs.local_base != nullptr) mirror_f32(jobs, j.out + i, acc);
    return;
  }
  if ((int)blockIdx.x >= kPosParts || jobs.pos_out == nullptr) return;
  __shared__ double sh[256];
  double acc = 0.0;
  const int64_t nchunk = (jobs.pos_n + 255) / 256;
  const int64_t items = (int64_t)jobs.pos_nparts * nchunk;
  for (int64_t q = blockIdx.x; q < items; q += kPosParts) {
    const int64_t p = q / nchunk, i = (q - p * nchunk) * 256 + threadIdx.x;
    if (i < jobs.pos_n) acc += (double)jobs.pos_part[p * jobs.pos_stride + i];
  }
  sh[threadIdx.x] = acc;
  __syncthreads();
  for (int s2 = 128; s2 > 0; s2 >>= 1) {
    if ((int)threadIdx.x < s2) sh[threadIdx.x] += sh[threadIdx.x + s2];
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    const double v = (double)jobs.pos_scale * sh[0];
    jobs.pos_out[blockIdx.x] = v;
    if (jobs.local_base != nullptr) {
      float* lo = reinterpret_cast<float*>(jobs.pos_out + blockIdx.x);
      const long long bits = __double_as_longlong(v);
      mirror_f32(jobs, lo, __int_as_float((int)(bits & 0xffffffffll)));
      mirror_f32(jobs, lo + 1, __int_as_float((int)(bits >> 32)));
    }
  }
}

int reduce_parts_batch_launch(const ReduceJobs& jobs, cudaStream_t st) {
  int64_t nmax = jobs.pos_out ? (int64_t)kPosParts * 256 : 0;
  for (int q = 0; q < jobs.count; ++q) nmax = jobs.job[q].n > nmax ? jobs.job[q].n : nmax;
  if ((jobs.count == 0 && !jobs.pos_out) || nmax == 0) return BSC_OK;
  dim3 grid((unsigned)((nmax + 255) / 256), (unsigned)(jobs.count + (jobs.pos_out ? 1 : 0)));
  reduce_parts_batch_kernel<<<grid, 256, 0, st>>>(jobs);
  BSC_LAUNCH_CHECK("reduce_parts_batch_kernel");
  return BSC_OK;
}

int reduce_parts_launch(const float* part, int nparts, int64_t part_stride, int64_t n, float* out, cudaStream_t st) {
  if (n == 0) return BSC_OK;
  reduce_parts_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(part, nparts, part_stride, n, out);
  BSC_LAUNCH_CHECK("reduce_parts_kernel");
  return BSC_OK;
}

// ---------------------------------------------------------------------------------------------
// Statistics finish: one launch turns the per-rank records into everything the rest of the step needs --
//   stats[p][i] = sum over ranks (in rank order) of the records' sum-exp rows          (the all-reduce, deterministic)
//   uv[p][i]    = r_i / stats[p][i]                                                     (backward weights)
//   loss        = c [ sum over ordered pairs sum_i r_i log(stats[p][i]) - 2 pos ],  c = 1 / (N M(M-1))   (loss_func.py:54;
//                 with a_ij = s*S_ij - s the r_i*s terms cancel against the s inside a_ij); pos = sum over ranks and
//                 parts of the double partial sums, fixed order
//   gsbuf       = {gs, 1/gs} per ordered pair (x, y): the power of two that keeps the fp16 weight tile
//                 |gs (e (u + v) - 2 T) / |y_j|| <= 4 gs rmax max_j(1/|y_j|) in range; max inverse norms at [12..14]
// Double accumulation, every thread a fixed subset, block partials added in block order by whichever block finishes
// last (ticket counter): independent of scheduling.
// ---------------------------------------------------------------------------------------------
constexpr int kFinishBlocks = 256;  // grid cap = partial sums kept in the scratch

// scratch layout (caller-zeroed once, left zero): u32 ticket | pad | double part[kFinishBlocks] | float vmax[3][kFinishBlocks]
__global__ void __launch_bounds__(256) stats_finish_kernel(const StatsFinishArgs a) {
  __shared__ double sh[256];
  __shared__ float mx[3][8];
  __shared__ bool last;
  unsigned* ticket = reinterpret_cast<unsigned*>(a.scratch);
  double* part = reinterpret_cast<double*>(a.scratch + 16);
  float* vpart = reinterpret_cast<float*>(a.scratch + 16 + kFinishBlocks * 8);
  const int64_t n = a.n;
  double acc = 0.0;
  float vmax[3] = {0.f, 0.f, 0.f};
  for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < n; i += (int64_t)gridDim.x * 256) {
    const float ri = a.r[i];
    float se[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    for (int w = 0; w < a.world; ++w) {  // rank order: the sum is the same on every rank and in every run
      const float* rec = a.slots + (int64_t)w * a.record_len + i;
#pragma unroll
      for (int p = 0; p < 6; ++p) se[p] += rec[(int64_t)p * n];
    }
#pragma unroll
    for (int p = 0; p < 6; ++p) {
      a.stats[(int64_t)p * n + i] = se[p];
      a.uv[(int64_t)p * n + i] = se[p] > 0.f ? ri / se[p] : 0.f;
      const int x = p >> 1, y = (p & 1) + ((p & 1) >= x ? 1 : 0);  // pair_id(x, y) = p
      if (((a.present_mask >> x) & 1) && ((a.present_mask >> y) & 1)) acc += (double)ri * (double)logf(se[p]);
    }
#pragma unroll
    for (int m = 0; m < 3; ++m)
      if (a.use_gs && a.inv[m] != nullptr) vmax[m] = fmaxf(vmax[m], a.inv[m][i]);
  }
#pragma unroll
  for (int m = 0; m < 3; ++m) {
    float v = vmax[m];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    if ((threadIdx.x & 31) == 0) mx[m][threadIdx.x >> 5] = v;
  }
  sh[threadIdx.x] = acc;
  __syncthreads();
  for (int s2 = 128; s2 > 0; s2 >>= 1) {
    if ((int)threadIdx.x < s2) sh[threadIdx.x] += sh[threadIdx.x + s2];
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    part[blockIdx.x] = sh[0];
    for (int m = 0; m < 3; ++m) {
      float v = mx[m][0];
      for (int w = 1; w < 8; ++w) v = fmaxf(v, mx[m][w]);
      vpart[m * kFinishBlocks + blockIdx.x] = v;
    }
    __threadfence();
    last = atomicAdd(ticket, 1u) == gridDim.x - 1;  // whoever comes last adds the partial sums in BLOCK order
  }
  __syncthreads();
  if (!last) return;
  __threadfence();
  // positive-pair term: world x kPosParts double partial sums, loaded by as many threads, added by one in a fixed order
  __shared__ double pp[kMaxPeers * kPosParts];
  const int64_t pos_off = stats_pos_offset(n);
  const int nparts = a.world * kPosParts;
  for (int t = threadIdx.x; t < nparts; t += 256)
    pp[t] = reinterpret_cast<const double*>(a.slots + (int64_t)(t / kPosParts) * a.record_len + pos_off)[t % kPosParts];
  double* cp = reinterpret_cast<double*>(a.stats + pos_off);
  if (threadIdx.x >= 1 && threadIdx.x < kPosParts) cp[threadIdx.x] = 0.0;
  sh[threadIdx.x] = threadIdx.x < gridDim.x ? __ldcg(part + threadIdx.x) : 0.0;
  __syncthreads();
  if (threadIdx.x != 0) return;
  *ticket = 0u;  // ready for the next call on this stream
  double total = 0.0;
  for (int c = 0; c < (int)gridDim.x; ++c) total += sh[c];
  double pos = 0.0;
  for (int t = 0; t < nparts; ++t) pos += pp[t];
  int m = 0;
  for (int q = 0; q < 3; ++q) m += (a.present_mask >> q) & 1;
  const double c = 1.0 / ((double)n * m * (m - 1));
  a.loss[0] = (float)(c * (total - 2.0 * pos));
  a.stats[6 * n] = (float)pos;
  for (int q = 1; q < kStatsExtra; ++q) a.stats[6 * n + q] = 0.f;
  cp[0] = pos;
  float invmax[3];
  for (int q = 0; q < 3; ++q) {
    float t = 0.f;
    for (int cc = 0; cc < (int)gridDim.x; ++cc) t = fmaxf(t, __ldcg(vpart + q * kFinishBlocks + cc));
    invmax[q] = (a.use_gs && a.inv[q] != nullptr) ? t : 1.0f;
    a.gsbuf[12 + q] = invmax[q];
  }
  const float rm = fmaxf(a.rmax[0], 1.0f);
  for (int x = 0; x < 3; ++x)
    for (int y = 0; y < 3; ++y) {
      if (x == y) continue;
      float gs = 1.0f;
      if (a.use_gs) gs = exp2f(floorf(log2f(8192.0f / (rm * fmaxf(invmax[y], 1e-30f)))));
      a.gsbuf[2 * pair_id(x, y)] = gs;
      a.gsbuf[2 * pair_id(x, y) + 1] = 1.0f / gs;
    }
}

size_t stats_finish_scratch_bytes() { return 16 + (size_t)kFinishBlocks * 8 + 3 * (size_t)kFinishBlocks * 4; }

int stats_finish_launch(const StatsFinishArgs& a, cudaStream_t st) {
  int64_t blocks = (a.n + 255) / 256;
  if (blocks > kFinishBlocks) blocks = kFinishBlocks;
  if (blocks < 1) blocks = 1;
  stats_finish_kernel<<<(unsigned)blocks, 256, 0, st>>>(a);
  BSC_LAUNCH_CHECK("stats_finish_kernel");
  return BSC_OK;
}

// ---------------------------------------------------------------------------------------------
// K6: combine the raw sweep output with the positive term and chain through F.normalize:
//   ghat = c*s*(raw0/gs0 + raw1/gs1)   (the -2*T_ij term is already inside the sweep's weights),
//   dx = gout * (ghat - xhat <xhat,ghat>) / ||x||
// One warp per (row, modality).
// ---------------------------------------------------------------------------------------------

template <typename T>
__global__ void __launch_bounds__(256) combine_kernel(CombineArgs a) {
  const int lane = threadIdx.x & 31;
  const int64_t il = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int m = blockIdx.y;
  if (il >= a.n_local || a.x[m] == nullptr) return;
  int mcount = 0;
  for (int q = 0; q < 3; ++q) mcount += (a.present_mask >> q) & 1;
  const float c = 1.0f / ((float)a.n_global * (float)(mcount * (mcount - 1)));
  const float cs = c * a.scale;
  // slot 0 / 1 hold the partial sums against the first / second other present modality
  const CombineSrc s0src = a.src[m][0], s1src = a.src[m][1];
  const float inv_gs[2] = {a.gsbuf[s0src.gs_index], a.n_slots > 1 ? a.gsbuf[s1src.gs_index] : 0.f};
  const float gout = a.grad_out[0];
  const T* xr = (const T*)a.x[m] + il * a.ldx[m];
  T* gr = (T*)a.grad[m] + il * (int64_t)a.d;
  const float* raw0 = s0src.src + il * (int64_t)a.d;
  const float* raw1 = a.n_slots > 1 ? s1src.src + il * (int64_t)a.d : raw0;

  float ss = 0.f;
  for (int cidx = lane; cidx < a.d; cidx += 32) {
    const float v = to_f32<T>(xr[cidx]);
    ss = fmaf(v, v, ss);
  }
  ss = warp_sum(ss);
  const float nrm = sqrtf(ss);
  const bool clamped = nrm < 1e-12f;
  const float inv = 1.0f / fmaxf(nrm, 1e-12f);

  auto ghat_at = [&](int cidx) -> float {
    float s0 = raw0[cidx];
    for (int q = 1; q < s0src.nparts; ++q) s0 += raw0[q * s0src.stride + cidx];  // fixed order: deterministic
    float raw = s0 * inv_gs[0];
    if (a.n_slots > 1) {
      float s1 = raw1[cidx];
      for (int q = 1; q < s1src.nparts; ++q) s1 += raw1[q * s1src.stride + cidx];
      raw = fmaf(s1, inv_gs[1], raw);
    }
    return cs * raw;
  };

  float dot = 0.f;
  for (int cidx = lane; cidx < a.d; cidx += 32) dot = fmaf(to_f32<T>(xr[cidx]) * inv, ghat_at(cidx), dot);
  dot = warp_sum(dot);
  if (clamped) dot = 0.f;  // y = x / eps: the Jacobian is I / eps
  for (int cidx = lane; cidx < a.d; cidx += 32) {
    const float xh = to_f32<T>(xr[cidx]) * inv;
    gr[cidx] = from_f32<T>(gout * inv * (ghat_at(cidx) - xh * dot));
  }
}

// Four consecutive columns per lane, 16- / 8-byte accesses.
template <typename T>
__device__ __forceinline__ void load4(const T* p, float (&o)[4]);
template <>
__device__ __forceinline__ void load4<float>(const float* p, float (&o)[4]) {
  const float4 v = *reinterpret_cast<const float4*>(p);
  o[0] = v.x, o[1] = v.y, o[2] = v.z, o[3] = v.w;
}
template <>
__device__ __forceinline__ void load4<__nv_bfloat16>(const __nv_bfloat16* p, float (&o)[4]) {
  const uint2 v = *reinterpret_cast<const uint2*>(p);
  o[0] = __uint_as_float(v.x << 16), o[1] = __uint_as_float(v.x & 0xffff0000u);
  o[2] = __uint_as_float(v.y << 16), o[3] = __uint_as_float(v.y & 0xffff0000u);
}
template <>
__device__ __forceinline__ void load4<__half>(const __half* p, float (&o)[4]) {
  const uint2 v = *reinterpret_cast<const uint2*>(p);
  const float2 a = __half22float2(*reinterpret_cast<const __half2*>(&v.x));
  const float2 b = __half22float2(*reinterpret_cast<const __half2*>(&v.y));
  o[0] = a.x, o[1] = a.y, o[2] = b.x, o[3] = b.y;
}
template <typename T>
__device__ __forceinline__ void store4(T* p, const float (&o)[4]);
template <>
__device__ __forceinline__ void store4<float>(float* p, const float (&o)[4]) {
  *reinterpret_cast<float4*>(p) = make_float4(o[0], o[1], o[2], o[3]);
}
template <>
__device__ __forceinline__ void store4<__nv_bfloat16>(__nv_bfloat16* p, const float (&o)[4]) {
  const __nv_bfloat162 a = __floats2bfloat162_rn(o[0], o[1]), b = __floats2bfloat162_rn(o[2], o[3]);
  *reinterpret_cast<uint2*>(p) =
      make_uint2(*reinterpret_cast<const uint32_t*>(&a), *reinterpret_cast<const uint32_t*>(&b));
}
template <>
__device__ __forceinline__ void store4<__half>(__half* p, const float (&o)[4]) {
  const __half2 a = __floats2half2_rn(o[0], o[1]), b = __floats2half2_rn(o[2], o[3]);
  *reinterpret_cast<uint2*>(p) =
      make_uint2(*reinterpret_cast<const uint32_t*>(&a), *reinterpret_cast<const uint32_t*>(&b));
}

// Same arithmetic for d % 128 == 0, d <= 128 * NV and 4-element-aligned rows: the row and its raw gradient stay in
// registers (each read ONCE, all loads of a row in flight together) -- the per-column version above is bound by the
// latency of its 4-byte loads (0.36 ms for 4096 rows x 3 modalities with 4 partial sums; 8 % of a step on 8 GPUs).
template <typename T, int NV>
__global__ void __launch_bounds__(256) combine_rows4_kernel(CombineArgs a) {
  const int lane = threadIdx.x & 31;
  const int64_t il = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int m = blockIdx.y;
  if (il >= a.n_local || a.x[m] == nullptr) return;
  int mcount = 0;
  for (int q = 0; q < 3; ++q) mcount += (a.present_mask >> q) & 1;
  const float c = 1.0f / ((float)a.n_global * (float)(mcount * (mcount - 1)));
  const float cs = c * a.scale;
  const CombineSrc s0src = a.src[m][0], s1src = a.src[m][1];
  const float inv_gs[2] = {a.gsbuf[s0src.gs_index], a.n_slots > 1 ? a.gsbuf[s1src.gs_index] : 0.f};
  const float gout = a.grad_out[0];
  const T* xr = (const T*)a.x[m] + il * a.ldx[m] + lane * 4;
  T* gr = (T*)a.grad[m] + il * (int64_t)a.d + lane * 4;
  const float* raw0 = s0src.src + il * (int64_t)a.d + lane * 4;
  const float* raw1 = a.n_slots > 1 ? s1src.src + il * (int64_t)a.d + lane * 4 : raw0;
  const int nv = a.d >> 7;

  float x[NV][4], g[NV][4];
#pragma unroll
  for (int v = 0; v < NV; ++v)
    if (v < nv) load4<T>(xr + v * 128, x[v]);
#pragma unroll
  for (int v = 0; v < NV; ++v) {
    if (v >= nv) continue;
    float s0[4], s1[4] = {0.f, 0.f, 0.f, 0.f};
    load4<float>(raw0 + v * 128, s0);
    for (int q = 1; q < s0src.nparts; ++q) {  // fixed order: deterministic
      float t[4];
      load4<float>(raw0 + q * s0src.stride + v * 128, t);
#pragma unroll
      for (int e = 0; e < 4; ++e) s0[e] += t[e];
    }
    if (a.n_slots > 1) {
      load4<float>(raw1 + v * 128, s1);
      for (int q = 1; q < s1src.nparts; ++q) {
        float t[4];
        load4<float>(raw1 + q * s1src.stride + v * 128, t);
#pragma unroll
        for (int e = 0; e < 4; ++e) s1[e] += t[e];
      }
    }
#pragma unroll
    for (int e = 0; e < 4; ++e) g[v][e] = cs * fmaf(s1[e], inv_gs[1], s0[e] * inv_gs[0]);
  }
  float ss = 0.f;
#pragma unroll
  for (int v = 0; v < NV; ++v)
    if (v < nv)
#pragma unroll
      for (int e = 0; e < 4; ++e) ss = fmaf(x[v][e], x[v][e], ss);
  ss = warp_sum(ss);
  const float nrm = sqrtf(ss);
  const bool clamped = nrm < 1e-12f;
  const float inv = 1.0f / fmaxf(nrm, 1e-12f);
  float dot = 0.f;
#pragma unroll
  for (int v = 0; v < NV; ++v)
    if (v < nv)
#pragma unroll
      for (int e = 0; e < 4; ++e) dot = fmaf(x[v][e] * inv, g[v][e], dot);
  dot = warp_sum(dot);
  if (clamped) dot = 0.f;
#pragma unroll
  for (int v = 0; v < NV; ++v) {
    if (v >= nv) continue;
    float o[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) o[e] = gout * inv * (g[v][e] - x[v][e] * inv * dot);
    store4<T>(gr + v * 128, o);
  }
}

template <typename T>
static bool combine_rows4_ok(const CombineArgs& a) {
  if (a.d % 128 != 0 || a.d > 1024) return false;
  for (int m = 0; m < 3; ++m) {
    if (a.x[m] == nullptr) continue;
    if (a.ldx[m] % 4 != 0 || ((uintptr_t)a.x[m] | (uintptr_t)a.grad[m]) % (4 * sizeof(T)) != 0) return false;
    for (int sl = 0; sl < a.n_slots; ++sl)
      if ((uintptr_t)a.src[m][sl].src % 16 != 0 || a.src[m][sl].stride % 4 != 0) return false;
  }
  return true;
}

template <typename T>
static void combine_dispatch(const CombineArgs& a, dim3 grid, cudaStream_t st) {
  if (combine_rows4_ok<T>(a)) {
    if (a.d <= 512)
      combine_rows4_kernel<T, 4><<<grid, 256, 0, st>>>(a);
    else if (a.d <= 768)
      combine_rows4_kernel<T, 6><<<grid, 256, 0, st>>>(a);
    else
      combine_rows4_kernel<T, 8><<<grid, 256, 0, st>>>(a);
  } else
    combine_kernel<T><<<grid, 256, 0, st>>>(a);
}

int combine_launch(const CombineArgs& a, int x_dtype, cudaStream_t st) {
  if (a.n_local == 0) return BSC_OK;
  dim3 grid((unsigned)((a.n_local + 7) / 8), 3);
  if (x_dtype == BSC_F32)
    combine_dispatch<float>(a, grid, st);
  else if (x_dtype == BSC_BF16)
    combine_dispatch<__nv_bfloat16>(a, grid, st);
  else if (x_dtype == BSC_F16)
    combine_dispatch<__half>(a, grid, st);
  else {
    set_error("bsc_infonce_bwd: unknown x_dtype %d", x_dtype);
    return BSC_EINVAL;
  }
  BSC_LAUNCH_CHECK("combine_kernel");
  return BSC_OK;
}

// ---------------------------------------------------------------------------------------------
// k-way merge of candidate lists (cross-CTA and cross-GPU): one warp per query, ordering
// (score desc, id asc), ids < 0 are padding.
// ---------------------------------------------------------------------------------------------
// One warp per query.  Round t selects the best candidate that comes strictly after round t-1's pick in the total
// order (score desc, id asc); ids are unique across parts, so no "taken" marks are needed and the result does not
// depend on how the candidates are spread over lanes.
__global__ void __launch_bounds__(128) topk_merge_kernel(const float* __restrict__ cs, const int64_t* __restrict__ ci,
                                                         int parts, int64_t nq, int k, float* __restrict__ os,
                                                         int64_t* __restrict__ oi) {
  const int lane = threadIdx.x & 31;
  const int64_t q = (int64_t)blockIdx.x * 4 + (threadIdx.x >> 5);
  if (q >= nq) return;
  constexpr float kLowest = -3.4028234663852886e38f;
  const int total = parts * k;
  float prev_s = 0.f;
  int64_t prev_i = -1;
  int t = 0;
  for (; t < k; ++t) {
    float best_s = kLowest;
    int64_t best_i = -1;
    for (int c = lane; c < total; c += 32) {
      const int64_t idx = ((int64_t)(c / k) * nq + q) * k + (c % k);
      const int64_t id = ci[idx];
      if (id < 0) continue;
      const float s = cs[idx];
      const bool after_prev = (t == 0) || (s < prev_s) || (s == prev_s && id > prev_i);
      const bool better = (best_i < 0) || (s > best_s) || (s == best_s && id < best_i);
      if (after_prev && better) {
        best_s = s;
        best_i = id;
      }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float s2 = __shfl_xor_sync(0xffffffffu, best_s, o);
      const int64_t i2 = __shfl_xor_sync(0xffffffffu, best_i, o);
      const bool take = (i2 >= 0) && ((best_i < 0) || (s2 > best_s) || (s2 == best_s && i2 < best_i));
      if (take) {
        best_s = s2;
        best_i = i2;
      }
    }
    if (best_i < 0) break;  // only padding is left
    if (lane == 0) {
      os[q * k + t] = best_s;
      oi[q * k + t] = best_i;
    }
    prev_s = best_s;
    prev_i = best_i;
  }
  for (int u = t + lane; u < k; u += 32) {
    os[q * k + u] = kLowest;
    oi[q * k + u] = -1;
  }
}

#if BSC_SCAN_V2
// Few queries, many lists (a 32-query pass over a bank shard leaves 3 lists per SM): one BLOCK per query.  Every thread
// reads its share of the candidates ONCE and keeps its own best KM of them in registers; the 128 short lists meet in
// shared memory, where warp 0 runs the k selection rounds of topk_merge_kernel (same total order: score desc, id asc, so
// the result is the same whatever the split).  The warp-per-query kernel walks all parts * k candidates k times with
// dependent global loads: ~0.2 ms for 888 lists, as long as the scan of a 625k-row shard itself.
constexpr int KM = 8;
__device__ __forceinline__ bool cand_before(float s, int64_t i, float s2, int64_t i2) {
  // (s, i) valid; does it come before (s2, i2) in (score desc, id asc)?  Padding (i2 < 0) comes after everything.
  return (i2 < 0) || (s > s2) || (s == s2 && i < i2);
}
__global__ void __launch_bounds__(128) topk_merge_block_kernel(const float* __restrict__ cs,
                                                               const int64_t* __restrict__ ci, int parts, int64_t nq,
                                                               int k, float* __restrict__ os, int64_t* __restrict__ oi) {
  __shared__ float sh_s[128 * KM];
  __shared__ int64_t sh_i[128 * KM];
  const int64_t q = blockIdx.x;
  constexpr float kLowest = -3.4028234663852886e38f;
  float ts[KM];
  int64_t ti[KM];
#pragma unroll
  for (int u = 0; u < KM; ++u) {
    ts[u] = kLowest;
    ti[u] = -1;
  }
  const int total = parts * k;
  for (int c = threadIdx.x; c < total; c += 128) {
    const int64_t idx = ((int64_t)(c / k) * nq + q) * k + (c % k);
    const int64_t id = ci[idx];
    const float s = cs[idx];
    if (id < 0 || !cand_before(s, id, ts[KM - 1], ti[KM - 1])) continue;
    ts[KM - 1] = s;
    ti[KM - 1] = id;
#pragma unroll
    for (int u = KM - 1; u > 0; --u) {
      if (ti[u] >= 0 && cand_before(ts[u], ti[u], ts[u - 1], ti[u - 1])) {
        const float fs = ts[u];
        ts[u] = ts[u - 1];
        ts[u - 1] = fs;
        const int64_t fi = ti[u];
        ti[u] = ti[u - 1];
        ti[u - 1] = fi;
      }
    }
  }
#pragma unroll
  for (int u = 0; u < KM; ++u) {
    sh_s[u * 128 + threadIdx.x] = ts[u];
    sh_i[u * 128 + threadIdx.x] = ti[u];
  }
  __syncthreads();
  if (threadIdx.x >= 32) return;
  const int lane = threadIdx.x;
  float prev_s = 0.f;
  int64_t prev_i = -1;
  int t = 0;
  for (; t < k; ++t) {
    float best_s = kLowest;
    int64_t best_i = -1;
    for (int c = lane; c < 128 * KM; c += 32) {
      const int64_t id = sh_i[c];
      if (id < 0) continue;
      const float s = sh_s[c];
      const bool after_prev = (t == 0) || (s < prev_s) || (s == prev_s && id > prev_i);
      if (after_prev && cand_before(s, id, best_s, best_i)) {
        best_s = s;
        best_i = id;
      }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float s2 = __shfl_xor_sync(0xffffffffu, best_s, o);
      const int64_t i2 = __shfl_xor_sync(0xffffffffu, best_i, o);
      if (i2 >= 0 && cand_before(s2, i2, best_s, best_i)) {
        best_s = s2;
        best_i = i2;
      }
    }
    if (best_i < 0) break;  // only padding is left
    if (lane == 0) {
      os[q * k + t] = best_s;
      oi[q * k + t] = best_i;
    }
    prev_s = best_s;
    prev_i = best_i;
  }
  for (int u = t + lane; u < k; u += 32) {
    os[q * k + u] = kLowest;
    oi[q * k + u] = -1;
  }
}
#endif

int topk_merge_launch(const float* cs, const int64_t* ci, int parts, int64_t nq, int k, float* os, int64_t* oi,
                      cudaStream_t st) {
  if (nq == 0) return BSC_OK;
  if (k > 32) {
    set_error("bsc_topk_merge: k=%d > 32 unsupported", k);
    return BSC_EUNSUPPORTED;
  }
#if BSC_SCAN_V2
  if (k <= KM && nq <= 2048 && (int64_t)parts * k >= 256) {
    topk_merge_block_kernel<<<(unsigned)nq, 128, 0, st>>>(cs, ci, parts, nq, k, os, oi);
    BSC_LAUNCH_CHECK("topk_merge_block_kernel");
    return BSC_OK;
  }
#endif
  topk_merge_kernel<<<(unsigned)((nq + 3) / 4), 128, 0, st>>>(cs, ci, parts, nq, k, os, oi);
  BSC_LAUNCH_CHECK("topk_merge_kernel");
  return BSC_OK;
}

}  // namespace bsc

// ---------------------------------------------------------------------------------------------
// bsc_label_stats_static: K2 for callers that may not allocate (CUDA graph capture, serving loops, harnesses that
// forbid cudaMalloc).  Same kernels as bsc_label_stats; the hash table / head / count / minidx scratch lives in ONE
// file-scope device array of fixed size (every GPU that loads the library holds these 1.25 MiB until exit), so the
// call takes inputs and outputs only and reaches no cudaMalloc / cudaFree / allocating library routine.
//
// g_label_static_scratch is a HOST-SIDE STAND-IN in host code, not a device address: its address on the current
// device is resolved with cudaGetSymbolAddress once per device (below, in this translation unit) and that pointer is
// what the kernels receive.
// ---------------------------------------------------------------------------------------------
namespace bsc {

constexpr int64_t kLabelStaticMaxN = BSC_LABEL_STATS_STATIC_MAX_N;
// slots[2 * max n] | head[max n] | count[max n] | minidx[max n], int32 each (label_stats_ws_bytes of the maximum)
constexpr size_t kLabelStaticBytes = (size_t)(2 * kLabelStaticMaxN + 3 * kLabelStaticMaxN) * 4;
static_assert((kLabelStaticMaxN & (kLabelStaticMaxN - 1)) == 0 && kLabelStaticMaxN >= 1024,
              "the table of the maximum size must be exactly 2 * max n slots");

static __device__ __align__(256) unsigned char g_label_static_scratch[kLabelStaticBytes];

// Device address of the array on the CURRENT device (resolved on the first call per device, then cached).
static int label_static_scratch(char** out) {
  static std::atomic<char*> cache[128];
  const int dev = current_device();
  if (dev >= 0 && dev < 128) {
    char* p = cache[dev].load(std::memory_order_acquire);
    if (p) {
      *out = p;
      return BSC_OK;
    }
  }
  char* p = nullptr;
  BSC_CUDA(cudaGetSymbolAddress((void**)&p, g_label_static_scratch));
  BSC_REQUIRE(p != nullptr, BSC_ECUDA, "bsc_label_stats_static: cudaGetSymbolAddress returned a null address");
  if (dev >= 0 && dev < 128) cache[dev].store(p, std::memory_order_release);
  *out = p;
  return BSC_OK;
}

}  // namespace bsc

extern "C" int bsc_label_stats_static(const int64_t* labels, int64_t n, float* r, int32_t* rep, int32_t* next,
                                      float* rmax, void* stream) {
  using namespace bsc;
  BSC_REQUIRE(labels && r && rep && next && rmax, BSC_EINVAL, "bsc_label_stats_static: null pointer");
  BSC_REQUIRE(n >= 0, BSC_EINVAL, "bsc_label_stats_static: negative n");
  BSC_REQUIRE(n <= kLabelStaticMaxN && label_stats_ws_bytes(n) <= kLabelStaticBytes, BSC_EUNSUPPORTED,
              "bsc_label_stats_static: n = %lld exceeds the static scratch (at most %lld labels per call); use "
              "bsc_label_stats with a caller-owned workspace",
              (long long)n, (long long)kLabelStaticMaxN);
  char* scratch = nullptr;
  const int rc = label_static_scratch(&scratch);
  if (rc != BSC_OK) return rc;
  // label_stats_launch clears what must start empty (table slots, counts, minima, rmax) on `stream` before the
  // kernels; the array keeps the previous call's contents otherwise.
  return label_stats_launch(labels, n, r, rep, next, rmax, scratch, (cudaStream_t)stream);
}
