// preprocess.cu — K-pre: fused batch load + standardise + categorical gather (+ label / index gather).
//
// Replaces, for one batch, the reference's sampler/loader batching (task.py:260-263), the Normalize
// transform (task.py:250-251: tensor.sub_(mean).div_(std)), the H2D copies (task.py:305,307) and the
// embedding lookup + concat at the front of the tabular model.  Results are bit-exact
// with the CPU oracle: IEEE subtract then divide (no FMA contraction, no reciprocal), RNE to bf16.
//
// Column layout of the output row (an internal choice; the first Linear's weight columns are stored in the
// same order, see b200step/model.py):   [ 26 x dim embedding columns | num_dense standardised columns | zero pad ].
// With the embeddings first, float4 k of a sample's gathered table rows lands at element 4k of the row: every lane
// writes one aligned 8-byte (bf16) / 16-byte (fp32) chunk and a warp writes 256 / 512 contiguous bytes per
// instruction — no shared-memory staging, and the embedding backward reads the gradient back the same way.
//
// Persistent warps, one sample per warp at a time.  The per-sample work is a chain of three dependent gathers
// (sampler index -> dataset row -> 26 table rows); the chain of sample k+1 is started before sample k is
// finished (software pipelining in registers), so a warp pays one gather latency per sample instead of three.
#include <stdlib.h>

#include "common.cuh"

namespace b200 {

constexpr int PRE_THREADS = 256;   // default shape: 256-thread CTAs, four resident per SM
// Budgeted shape (b200_set_sm_budget below the device's SM count): ONE 1024-thread CTA per SM, pinned there by a
// shared-memory request nothing else fits beside, on `budget` SMs only.  The other SMs stay empty for the collective
// this launch is meant to run under; with the default shape the block scheduler would spread the CTAs over every SM.
constexpr int PRE_FAT_THREADS = 1024;
constexpr int PRE_FAT_SMEM = 160 * 1024;

// four fp32 values -> one aligned store in the output dtype
__device__ __forceinline__ void store4(float* dst, uint4 raw) { stg_stream(dst, raw); }
__device__ __forceinline__ void store4(__nv_bfloat16* dst, uint4 raw) {
  const uint2 o = make_uint2(pack_bf16x2(__uint_as_float(raw.x), __uint_as_float(raw.y)),
                             pack_bf16x2(__uint_as_float(raw.z), __uint_as_float(raw.w)));
  asm volatile("st.global.L1::no_allocate.v2.u32 [%0], {%1,%2};" ::"l"(dst), "r"(o.x), "r"(o.y) : "memory");
}
__device__ __forceinline__ void store1(float* dst, float v) { *dst = v; }
__device__ __forceinline__ void store1(__nv_bfloat16* dst, float v) { *dst = __float2bfloat16_rn(v); }

// A categorical id as a 32-bit table row.  nn.Embedding raises on ids outside [0, num_embeddings); a kernel cannot,
// so the id is clamped to row 0 (every later access stays inside its table, here and in the embedding backward,
// which reuses the stored ids) and the status word tells the host.
template <typename IdT>
__device__ __forceinline__ int checked_id(IdT raw, long long table_rows, uint32_t* status) {
  if ((unsigned long long)raw >= (unsigned long long)table_rows) {
    flag_status(status, B200_STATUS_BAD_CATEGORICAL_ID);
    return 0;
  }
  return (int)raw;
}

template <typename OutT, typename IdT, int kThreads, int kMinBlocks>
__global__ void __launch_bounds__(kThreads, kMinBlocks)
preprocess_kernel(const float* __restrict__ ds_dense, const IdT* __restrict__ ds_cat,
                  const int64_t* __restrict__ ds_label, const int64_t* __restrict__ row_idx,
                  const int32_t* __restrict__ cursor, const float* __restrict__ mean,
                  const float* __restrict__ stdv, const float* __restrict__ tables,
                  OutT* __restrict__ x_out, int64_t* __restrict__ y_out, int32_t* __restrict__ idx_out,
                  long long batch, int num_dense, int num_cat, long long table_rows, int embed_dim,
                  int vpr_shift, long long ldx, long long cursor_stride, uint32_t* __restrict__ status,
                  int32_t* __restrict__ label_stats, long long ignore_index) {
  pdl_launch_dependents();
  pdl_wait();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  int n_counted = 0;  // lane 0: labels != ignore_index among this warp's samples
  const long long cur = cursor ? (long long)(*cursor) : 0;
  const int64_t* ridx = row_idx + cur * cursor_stride;
  const int vec_per_row = embed_dim >> 2;        // float4 chunks per table row
  const int emb_cols = num_cat * embed_dim;      // dense columns start here
  const int tail = (int)ldx - emb_cols;          // dense + zero padding
  constexpr int PRE_WARPS = kThreads / 32;
  const long long stride = (long long)gridDim.x * PRE_WARPS;
  const float mu = lane < num_dense ? __ldg(mean + lane) : 0.f;  // num_dense <= 32 on the fast path
  const float sd = lane < num_dense ? __ldg(stdv + lane) : 1.f;

  const long long b_first = (long long)blockIdx.x * PRE_WARPS + warp;
  // The sampler indices of this warp's next 32 samples are fetched with ONE load (lane l holds the index of the
  // l-th upcoming sample) and handed out by shuffles; the dataset row of sample k+1 (its dense value, its first 32
  // categorical ids, its label) is requested while sample k is being gathered.
  long long b = b_first;
  int k = 0;  // position inside the current group of 32 samples
  long long rgrp = (b + (long long)lane * stride) < batch ? __ldg(ridx + b + (long long)lane * stride) : 0;
  long long r_b = __shfl_sync(0xffffffffu, rgrp, 0);
  float dense_b = 0.f;
  IdT id_b = 0;   // raw id: range-checked when it is USED, one iteration later, so the load stays a prefetch
  long long label_b = 0;
  if (b < batch) {
    if (lane < num_dense) dense_b = __ldg(ds_dense + r_b * num_dense + lane);
    if (lane < num_cat) id_b = __ldg(ds_cat + r_b * num_cat + lane);
    if (lane == 0) label_b = __ldg(ds_label + r_b);
  }
  for (; b < batch; b += stride) {
    const long long bn = b + stride;
    const long long r_cur = r_b;
    const float dense_cur = dense_b;
    const int id_cur = lane < num_cat ? checked_id(id_b, table_rows, status) : 0;
    const long long label_cur = label_b;
    if (++k == 32) {  // next group of sampler indices (warp-uniform branch)
      k = 0;
      rgrp = (bn + (long long)lane * stride) < batch ? __ldg(ridx + bn + (long long)lane * stride) : 0;
    }
    r_b = __shfl_sync(0xffffffffu, rgrp, k);
    if (bn < batch) {
      if (lane < num_dense) dense_b = __ldg(ds_dense + r_b * num_dense + lane);
      if (lane < num_cat) id_b = __ldg(ds_cat + r_b * num_cat + lane);
      if (lane == 0) label_b = __ldg(ds_label + r_b);
    }

    OutT* xr = x_out + b * ldx;
    // categorical ids: lane t holds the id of table tb + t
    for (int tb = 0; tb < num_cat; tb += 32) {
      const int t = tb + lane;
      int id = id_cur;
      if (tb > 0) id = t < num_cat ? checked_id(__ldg(ds_cat + r_cur * num_cat + t), table_rows, status) : 0;
      if (t < num_cat) idx_out[b * num_cat + t] = id;
      const int tcount = min(32, num_cat - tb);
      const int nvec = tcount * vec_per_row;
      OutT* xe = xr + tb * embed_dim;
      // gather: chunk i -> table tb + i / vec_per_row, float4 i % vec_per_row of that row -> row elements 4i..4i+3.
      // Four independent 16-byte loads per lane are issued before the first store, so one L2/HBM latency is paid
      // per 128 chunks (a whole T0 sample: 26 x 16 floats = 104 chunks) instead of one per 32.  32-bit offsets:
      // the host checks that the tables hold fewer than 2^31 elements.
      for (int i0 = 0; i0 < nvec; i0 += 128) {
        uint4 raw[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const int i = i0 + u * 32 + lane;
          const int tt = vpr_shift >= 0 ? (i >> vpr_shift) : (i / vec_per_row);
          const int qv = i - tt * vec_per_row;
          const uint32_t idt = (uint32_t)__shfl_sync(0xffffffffu, id, tt & 31);
          if (i < nvec) {
            const uint32_t off = ((uint32_t)(tb + tt) * (uint32_t)table_rows + idt) * (uint32_t)embed_dim + qv * 4;
            raw[u] = ldg_stream(tables + off);
          }
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const int i = i0 + u * 32 + lane;
          if (i < nvec) store4(xe + 4 * i, raw[u]);
        }
      }
    }
    // dense columns (x - mean) / std in IEEE fp32, then the zero K-padding: one contiguous tail
    {
      const float v = lane < num_dense ? __fdiv_rn(__fsub_rn(dense_cur, mu), sd) : 0.f;
      if (lane < tail) store1(xr + emb_cols + lane, v);
      for (int e = 32 + lane; e < tail; e += 32) {  // > 32 tail columns: plain path
        float w = 0.f;
        if (e < num_dense)
          w = __fdiv_rn(__fsub_rn(__ldg(ds_dense + r_cur * num_dense + e), __ldg(mean + e)), __ldg(stdv + e));
        store1(xr + emb_cols + e, w);
      }
    }
    if (lane == 0) {
      y_out[b] = label_cur;
      n_counted += (label_cur != ignore_index);
    }
  }
  // Rows that count towards the mean of the loss (nn.CrossEntropyLoss divides by the number of targets that are not
  // ignore_index): every warp adds its count, the LAST block to finish (ticket) publishes the total for K-loss and
  // re-arms the accumulator, so a captured graph can be replayed without a memset.
  if (label_stats != nullptr) {
    if (lane == 0 && n_counted) atomicAdd(label_stats + 1, n_counted);
    __syncthreads();
    if (threadIdx.x == 0) {
      __threadfence();
      if (atomicAdd(reinterpret_cast<unsigned*>(label_stats + 2), 1u) == gridDim.x - 1) {
        __threadfence();
        label_stats[0] = atomicExch(label_stats + 1, 0);
        label_stats[2] = 0;
      }
    }
  }
}

// ---- fast path -----------------------------------------------------------------------------------------------------
// The general kernel above spends ~340 warp instructions per sample, most of them loop control, 64-bit index
// arithmetic and constant loads of shapes it only learns at run time (ncu: 22.5 M instructions per 65536-row batch,
// issue-bound at 41 us).  The shapes the step actually uses — at most 32 tables, at most 32 dense columns, a
// power-of-two embedding width up to 128 — let the mapping "float4 chunk -> (table, float4 of its row)" be fixed per
// lane: with VPR = 2^VS chunks per table row, round u of a sample gives lane l the chunk u*32 + l, i.e. table
// u*(32 >> VS) + (l >> VS) and float4 (l & (VPR-1)) of that table's row.  Per round that leaves one shuffle (the
// table's id), one multiply-add, one 128-bit gather and one aligned store.
template <typename OutT, typename IdT, int VS>
__global__ void __launch_bounds__(PRE_THREADS, 4)
preprocess_fast_kernel(const float* __restrict__ ds_dense, const IdT* __restrict__ ds_cat,
                       const int64_t* __restrict__ ds_label, const int64_t* __restrict__ row_idx,
                       const int32_t* __restrict__ cursor, const float* __restrict__ mean,
                       const float* __restrict__ stdv, const float* __restrict__ tables,
                       OutT* __restrict__ x_out, int64_t* __restrict__ y_out, int32_t* __restrict__ idx_out,
                       int batch, int num_dense, int num_cat, int table_rows, int ldx, long long cursor_stride,
                       uint32_t* __restrict__ status, int32_t* __restrict__ label_stats, long long ignore_index) {
  pdl_launch_dependents();
  pdl_wait();
  constexpr int VPR = 1 << VS;          // float4 chunks per table row
  constexpr int DIM = 4 << VS;          // embedding width
  constexpr int TPR = 32 >> VS;         // tables covered by one round of 32 chunks
  constexpr int PRE_WARPS = PRE_THREADS / 32;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long cur = cursor ? (long long)(*cursor) : 0;
  const int64_t* ridx = row_idx + cur * cursor_stride;
  const int nvec = num_cat << VS;
  const int rounds = (nvec + 31) >> 5;
  const int emb_cols = nvec * 4;
  const int tail = ldx - emb_cols;      // dense columns + zero padding
  const int stride = gridDim.x * PRE_WARPS;
  const float mu = lane < num_dense ? __ldg(mean + lane) : 0.f;
  const float sd = lane < num_dense ? __ldg(stdv + lane) : 1.f;
  const int tl = lane >> VS;                                    // table of this lane's chunk in round 0
  const uint32_t lane_off = (uint32_t)(lane & (VPR - 1)) * 4u;  // float offset of the chunk inside a table row
  const uint32_t tstep = (uint32_t)TPR * (uint32_t)table_rows * (uint32_t)DIM;
  const uint32_t tbase0 = (uint32_t)tl * (uint32_t)table_rows * (uint32_t)DIM + lane_off;
  int n_counted = 0;

  // Two samples ahead: the dataset rows are random DRAM reads (~1.5 us under load) while one sample's gather takes
  // well under a microsecond, so a one-deep prefetch would leave every iteration waiting for DRAM.
  int b = blockIdx.x * PRE_WARPS + warp;
  int bl = b;       // next sample whose dataset row is requested
  int kl = 32;      // its position in the current group of 32 sampler indices (lane l holds the l-th upcoming one)
  long long rgrp = 0;
  float dense_s[2] = {0.f, 0.f};
  IdT id_s[2] = {0, 0};   // raw ids: range-checked when they are USED, so the loads stay prefetches
  long long label_s[2] = {0, 0};
  auto request = [&](float& dn, IdT& idr, long long& lb) {
    if (kl == 32) {  // next group of sampler indices (warp-uniform)
      kl = 0;
      rgrp = (long long)bl + (long long)lane * stride < batch ? __ldg(ridx + bl + (long long)lane * stride) : 0;
    }
    const long long r = __shfl_sync(0xffffffffu, rgrp, kl);
    if (bl < batch) {
      if (lane < num_dense) dn = __ldg(ds_dense + r * num_dense + lane);
      if (lane < num_cat) idr = __ldg(ds_cat + r * num_cat + lane);
      if (lane == 0) lb = __ldg(ds_label + r);
    }
    bl += stride;
    ++kl;
  };
  request(dense_s[0], id_s[0], label_s[0]);
  request(dense_s[1], id_s[1], label_s[1]);
  for (; b < batch; b += stride) {
    const float dense_cur = dense_s[0];
    const int id = lane < num_cat ? checked_id(id_s[0], (long long)table_rows, status) : 0;
    const long long label_cur = label_s[0];
    dense_s[0] = dense_s[1];
    id_s[0] = id_s[1];
    label_s[0] = label_s[1];
    request(dense_s[1], id_s[1], label_s[1]);
    OutT* xr = x_out + (long long)b * ldx;
    if (lane < num_cat) idx_out[(long long)b * num_cat + lane] = id;
    // gather: four independent 16-byte loads per lane in flight before the first store
    OutT* outp = xr + lane * 4;
    int tt = tl;
    uint32_t tbase = tbase0;
    for (int u0 = 0; u0 < rounds; u0 += 4) {
      uint4 raw[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int t = tt + u * TPR;
        const uint32_t idt = (uint32_t)__shfl_sync(0xffffffffu, id, t & 31);
        if (t < num_cat) raw[u] = ldg_stream(tables + (tbase + (uint32_t)u * tstep + idt * (uint32_t)DIM));
      }
#pragma unroll
      for (int u = 0; u < 4; ++u)
        if (tt + u * TPR < num_cat) store4(outp + u * 128, raw[u]);
      tt += 4 * TPR;
      tbase += 4 * tstep;
      outp += 4 * 128;
    }
    // dense columns (x - mean) / std in IEEE fp32 (subtract, then true divide), then the zero K-padding
    {
      const float v = lane < num_dense ? __fdiv_rn(__fsub_rn(dense_cur, mu), sd) : 0.f;
      OutT* xt = xr + emb_cols;
      if (lane < tail) store1(xt + lane, v);
      for (int e = 32 + lane; e < tail; e += 32) store1(xt + e, 0.f);
    }
    if (lane == 0) {
      y_out[b] = label_cur;
      n_counted += (label_cur != ignore_index);
    }
  }
  if (label_stats != nullptr) {  // same protocol as the general kernel
    if (lane == 0 && n_counted) atomicAdd(label_stats + 1, n_counted);
    __syncthreads();
    if (threadIdx.x == 0) {
      __threadfence();
      if (atomicAdd(reinterpret_cast<unsigned*>(label_stats + 2), 1u) == gridDim.x - 1) {
        __threadfence();
        label_stats[0] = atomicExch(label_stats + 1, 0);
        label_stats[2] = 0;
      }
    }
  }
}

}  // namespace b200

using namespace b200;

extern "C" int b200_preprocess(const float* ds_dense, const void* ds_cat, int cat_id_bytes, const int64_t* ds_label,
                               const int64_t* row_idx, const int32_t* cursor, const float* mean,
                               const float* stdv, const float* tables, void* x_out, int64_t* y_out,
                               int32_t* idx_out, int64_t batch, int num_dense, int num_cat,
                               int64_t table_rows, int embed_dim, int64_t ldx, int x_dtype,
                               int64_t cursor_stride, int32_t* label_stats, int64_t ignore_index, int max_ctas,
                               b200_stream_t stream) {
  if (cursor_stride <= 0) cursor_stride = batch;
  uint32_t* status = status_word();
  B200_REQUIRE(x_dtype == B200_F32 || x_dtype == B200_BF16, "unknown dtype code %d", x_dtype);
  B200_REQUIRE(batch > 0, "batch must be positive");
  B200_REQUIRE(num_dense >= 0 && num_cat >= 0 && embed_dim >= 0, "negative feature counts");
  B200_REQUIRE(num_cat == 0 || (embed_dim % 4 == 0 && embed_dim > 0),
               "embed_dim=%d must be a positive multiple of 4 (128-bit table reads)", embed_dim);
  B200_REQUIRE(ldx >= num_dense + (int64_t)num_cat * embed_dim, "ldx=%lld smaller than the feature count",
               (long long)ldx);
  const size_t es = x_dtype == B200_F32 ? 4 : 2;
  B200_REQUIRE((ldx * es) % 16 == 0, "ldx*elemsize must be a multiple of 16 bytes");
  B200_REQUIRE((reinterpret_cast<uintptr_t>(x_out) & 15) == 0 && (reinterpret_cast<uintptr_t>(tables) & 15) == 0,
               "x_out and tables must be 16-byte aligned");
  B200_REQUIRE((int64_t)num_cat * table_rows * embed_dim < (1ll << 31),
               "embedding tables hold 2^31 or more elements: 32-bit gather offsets would overflow");
  const int vpr = embed_dim / 4;
  int vpr_shift = -1;  // log2(float4 chunks per table row) when that is a power of two
  if (vpr > 0 && (vpr & (vpr - 1)) == 0) {
    vpr_shift = 0;
    while ((1 << vpr_shift) < vpr) ++vpr_shift;
  }
  const bool fat = sm_count() < sm_count_physical();
  const int warps = (fat ? PRE_FAT_THREADS : PRE_THREADS) / 32;
  long long blocks = (batch + warps - 1) / warps;
  long long cap = fat ? (long long)sm_count() : (long long)sm_count() * 8;
  if (!fat && max_ctas > 0 && max_ctas < cap) cap = max_ctas;   // a launch meant to share the SMs with another grid
  const unsigned grid = (unsigned)(blocks < cap ? blocks : cap);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  B200_REQUIRE(cat_id_bytes == 8 || cat_id_bytes == 4, "categorical ids must be int64 or int32 (cat_id_bytes=%d)",
               cat_id_bytes);
  // the shapes the training step uses take the specialised kernel
  const bool fast = !fat && num_cat >= 1 && num_cat <= 32 && num_dense <= 32 && vpr_shift >= 0 && vpr_shift <= 4 &&
                    batch < (1ll << 30) && ldx < (1ll << 30) && table_rows < (1ll << 30) &&
                    getenv("B200STEP_PRE_GENERAL") == nullptr;
#define B200_PRE_FAST(OUT_T, ID_T, VS_)                                                                           \
  B200_CHECK_CUDA(launch_pdl(preprocess_fast_kernel<OUT_T, ID_T, VS_>, dim3(grid), dim3(PRE_THREADS), 0, st,     \
                             ds_dense, static_cast<const ID_T*>(ds_cat), ds_label, row_idx, cursor, mean, stdv,   \
                             tables, static_cast<OUT_T*>(x_out), y_out, idx_out, (int)batch, num_dense, num_cat,  \
                             (int)table_rows, (int)ldx, (long long)cursor_stride, status, label_stats,            \
                             (long long)ignore_index))
#define B200_PRE_LAUNCH(OUT_T, ID_T)                                                                              \
  do {                                                                                                            \
    if (fast) {                                                                                                   \
      switch (vpr_shift) {                                                                                        \
        case 0: B200_PRE_FAST(OUT_T, ID_T, 0); break;                                                             \
        case 1: B200_PRE_FAST(OUT_T, ID_T, 1); break;                                                             \
        case 2: B200_PRE_FAST(OUT_T, ID_T, 2); break;                                                             \
        case 3: B200_PRE_FAST(OUT_T, ID_T, 3); break;                                                             \
        default: B200_PRE_FAST(OUT_T, ID_T, 4); break;                                                            \
      }                                                                                                           \
    } else if (fat) {                                                                                                    \
      auto kern = preprocess_kernel<OUT_T, ID_T, PRE_FAT_THREADS, 1>;                                             \
      static bool configured = false;                                                                             \
      if (!configured) {                                                                                          \
        B200_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, PRE_FAT_SMEM));   \
        configured = true;                                                                                        \
      }                                                                                                           \
      B200_CHECK_CUDA(launch_pdl(kern, dim3(grid), dim3(PRE_FAT_THREADS), PRE_FAT_SMEM, st, ds_dense,            \
                                 static_cast<const ID_T*>(ds_cat), ds_label, row_idx, cursor, mean, stdv, tables, \
                                 static_cast<OUT_T*>(x_out), y_out, idx_out, batch, num_dense, num_cat,           \
                                 table_rows, embed_dim, vpr_shift, ldx, (long long)cursor_stride, status,         \
                                 label_stats, (long long)ignore_index));                                          \
    } else {                                                                                                      \
      B200_CHECK_CUDA(launch_pdl(preprocess_kernel<OUT_T, ID_T, PRE_THREADS, 4>, dim3(grid), dim3(PRE_THREADS),   \
                                 0, st, ds_dense, static_cast<const ID_T*>(ds_cat), ds_label, row_idx, cursor,    \
                                 mean, stdv, tables, static_cast<OUT_T*>(x_out), y_out, idx_out, batch,           \
                                 num_dense, num_cat, table_rows, embed_dim, vpr_shift, ldx,                       \
                                 (long long)cursor_stride, status, label_stats, (long long)ignore_index));        \
    }                                                                                                             \
  } while (0)
  if (x_dtype == B200_F32) {
    if (cat_id_bytes == 8) B200_PRE_LAUNCH(float, int64_t);
    else B200_PRE_LAUNCH(float, int32_t);
  } else {
    if (cat_id_bytes == 8) B200_PRE_LAUNCH(__nv_bfloat16, int64_t);
    else B200_PRE_LAUNCH(__nv_bfloat16, int32_t);
  }
#undef B200_PRE_LAUNCH
#undef B200_PRE_FAST
  B200_CHECK_CUDA(cudaGetLastError());
  return 0;
}
