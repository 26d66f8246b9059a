// loss.cu — K-loss: fused softmax cross-entropy loss + gradient (+ last-layer bias gradient), and the
// column-sum kernel that produces the hidden layers' bias gradients.
//
// Replaces nn.CrossEntropyLoss (task.py:211,310; reduction='mean') and its backward (task.py:311).
// One warp owns one logits row at a time: the row is read once with 128-bit loads into registers,
// max / sum-exp / label pick are warp-shuffle reductions, and the gradient row is written once.
// Per-column sums of the written gradient (the bias gradient of the output layer) ride along in
// registers; per-CTA partials go to a scratch array and a second tiny kernel reduces them in a fixed
// order, so loss and bias gradient are run-to-run deterministic (no float atomics).
#include "common.cuh"

namespace b200 {

constexpr int CE_THREADS = 512;
constexpr int CE_WARPS = CE_THREADS / 32;
constexpr int CE_MAX_BLOCKS = 1024;
constexpr int CE_MAX_CLASSES = 1024;  // 32 lanes x 32 register-resident values
constexpr int CE_BUFS = 3;            // load(i+1) | compute(i) | store(i-1)

// Persistent CTAs walk tiles of 32 (bf16) / 16 (fp32) logits rows.  A tile is a contiguous block of global memory,
// so it is fetched with 1-D TMA bulk copies into shared memory while the previous tile is being processed and the
// one before that is being written back: the copy engine, not the register file, holds the bytes in flight.
// Each warp then reads its rows from shared memory (conflict-free 128-bit accesses), does the softmax
// arithmetic in registers, and overwrites the row in place with the gradient; the finished tile leaves through a
// bulk store.  Per-column sums of the stored gradient (output-layer bias gradient) and the loss ride along in
// registers; one partial row per CTA, reduced in a fixed order by reduce_partials_kernel (deterministic).
template <bool kBf16>
__global__ void __launch_bounds__(CE_THREADS, 1)
ce_loss_grad_kernel(const void* __restrict__ logits_, long long ldl, const int64_t* __restrict__ labels,
                    void* __restrict__ dlogits_, long long lddl, float* __restrict__ partials,
                    long long batch, int C, float scale, int want_dbias, long long ignore_index,
                    const int32_t* __restrict__ n_valid, uint32_t* __restrict__ status) {
  constexpr int VEC = kBf16 ? 8 : 4;          // elements per 16-byte vector
  constexpr int ITERS = 32 / VEC;             // vectors per lane (covers 1024 columns)
  constexpr int ES = kBf16 ? 2 : 4;
  constexpr int TILE_ROWS = kBf16 ? 32 : 16;  // 64*C bytes per buffer either way
  extern __shared__ __align__(128) uint8_t ce_smem[];
  __shared__ uint64_t full_bar[CE_BUFS];
  __shared__ float red_loss[CE_WARPS];

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nvec = C / VEC;
  const uint32_t row_bytes = (uint32_t)C * ES;
  const uint32_t buf_bytes = TILE_ROWS * row_bytes;
  const long long num_tiles = (batch + TILE_ROWS - 1) / TILE_ROWS;
  const uint8_t* gin = reinterpret_cast<const uint8_t*>(logits_);
  uint8_t* gout = reinterpret_cast<uint8_t*>(dlogits_);

  pdl_launch_dependents();
  if (threadIdx.x == 0) {
    for (int i = 0; i < CE_BUFS; ++i) mbar_init(&full_bar[i], 1);
    fence_mbar_init();
  }
  __syncthreads();
  pdl_wait();  // the logits come from the previous kernel
  if (n_valid != nullptr) {  // mean over the rows that are not ignore_index (counted by K-pre) instead of all rows
    const int nv = *n_valid;
    scale = nv > 0 ? scale * (float)batch / (float)nv : 0.f;
  }

  auto issue_load = [&](long long tile, int buf) {  // one thread
    const long long r0 = tile * TILE_ROWS;
    const int rows = (int)min((long long)TILE_ROWS, batch - r0);
    mbar_expect_tx(&full_bar[buf], rows * row_bytes);
    uint8_t* dst = ce_smem + (size_t)buf * buf_bytes;
    if (ldl == C) {
      bulk_load_1d(dst, gin + r0 * ldl * ES, rows * row_bytes, &full_bar[buf]);
    } else {
      for (int r = 0; r < rows; ++r)
        bulk_load_1d(dst + r * row_bytes, gin + (r0 + r) * ldl * ES, row_bytes, &full_bar[buf]);
    }
  };

  float dbacc[32];
#pragma unroll
  for (int i = 0; i < 32; ++i) dbacc[i] = 0.f;
  float loss_acc = 0.f;

  if (threadIdx.x == 0 && blockIdx.x < num_tiles) issue_load(blockIdx.x, 0);
  int it = 0;
  for (long long tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++it) {
    const int b = it % CE_BUFS;
    if (threadIdx.x == 0) {
      const long long tn = tile + gridDim.x;
      if (tn < num_tiles) {
        bulk_wait_read<1>();  // the store that last read buffer (it+1)%3 (two tiles ago) has drained it
        issue_load(tn, (it + 1) % CE_BUFS);
      }
    }
    mbar_wait(&full_bar[b], (it / CE_BUFS) & 1);
    const long long r0 = tile * TILE_ROWS;
    const int rows = (int)min((long long)TILE_ROWS, batch - r0);
    const uint32_t sbuf = smem_u32(ce_smem + (size_t)b * buf_bytes);
    for (int rr = warp; rr < rows; rr += CE_WARPS) {
      const long long row = r0 + rr;
      const uint32_t srow = sbuf + rr * row_bytes;
      const long long y64 = __ldg(labels + row);
      if (y64 == ignore_index || y64 < 0 || y64 >= C) {  // warp-uniform
        // nn.CrossEntropyLoss: a target equal to ignore_index contributes neither loss nor gradient; any other
        // target outside [0, C) makes torch raise — here the row is dropped the same way and the host is told
        if (y64 != ignore_index && lane == 0) flag_status(status, B200_STATUS_BAD_LABEL);
#pragma unroll
        for (int i = 0; i < ITERS; ++i) {
          const int v = lane + 32 * i;
          if (v < nvec) st_shared_v4(srow + v * 16, 0u, 0u, 0u, 0u);
        }
        continue;
      }
      const int y = (int)y64;
      const int vy = y / VEC, jy = y - vy * VEC;  // 16-byte vector / slot that holds the label's logit
      float z[32];
      float zmax = -INFINITY;
#pragma unroll
      for (int i = 0; i < ITERS; ++i) {
        const int v = lane + 32 * i;
        if (v < nvec) {
          const uint4 raw = ld_shared_v4(srow + v * 16);
          if constexpr (kBf16) {
            const uint32_t w[4] = {raw.x, raw.y, raw.z, raw.w};
#pragma unroll
            for (int t = 0; t < 4; ++t) {
              z[i * 8 + 2 * t] = bf16_lo(w[t]);
              z[i * 8 + 2 * t + 1] = bf16_hi(w[t]);
            }
          } else {
            z[i * 4 + 0] = __uint_as_float(raw.x);
            z[i * 4 + 1] = __uint_as_float(raw.y);
            z[i * 4 + 2] = __uint_as_float(raw.z);
            z[i * 4 + 3] = __uint_as_float(raw.w);
          }
#pragma unroll
          for (int j = 0; j < VEC; ++j) zmax = fmaxf(zmax, z[i * VEC + j]);
        } else {
#pragma unroll
          for (int j = 0; j < VEC; ++j) z[i * VEC + j] = -INFINITY;
        }
      }
      // the label's logit: one scalar shared-memory read instead of a compare per element
      float zy;
      if constexpr (kBf16) zy = __uint_as_float((uint32_t)ld_shared_u16(srow + y * 2) << 16);
      else zy = __uint_as_float(ld_shared_u32(srow + y * 4));
      zmax = warp_max(zmax);
      const float zmax_l2 = zmax * 1.4426950408889634f;
      float sum = 0.f;
#pragma unroll
      for (int i = 0; i < 32; ++i) {
        // exp(z - zmax) = 2^(z*log2e - zmax*log2e): one FFMA + MUFU.EX2 (rel. error ~1e-6, inside the 1e-5 loss
        // bar); the -inf tail gives 0
        z[i] = ex2_approx(fmaf(z[i], 1.4426950408889634f, -zmax_l2));
        sum += z[i];
      }
      sum = warp_sum(sum);
      loss_acc += (zmax + logf(sum)) - zy;  // identical in every lane
      const float inv = 1.0f / sum;
      const float inv_scale = inv * scale;
#pragma unroll
      for (int i = 0; i < ITERS; ++i) {
        const int v = lane + 32 * i;
        if (v < nvec) {
          float d[VEC];
#pragma unroll
          for (int j = 0; j < VEC; ++j) d[j] = z[i * VEC + j] * inv_scale;  // softmax * grad_scale
          if (v == vy) {  // exactly one lane, once per row: subtract the one-hot
#pragma unroll
            for (int j = 0; j < VEC; ++j)
              if (j == jy) d[j] = (z[i * VEC + j] * inv - 1.0f) * scale;
          }
          if constexpr (kBf16) {
            const uint32_t w[4] = {pack_bf16x2(d[0], d[1]), pack_bf16x2(d[2], d[3]), pack_bf16x2(d[4], d[5]),
                                   pack_bf16x2(d[6], d[7])};
            st_shared_v4(srow + v * 16, w[0], w[1], w[2], w[3]);
#pragma unroll
            for (int t = 0; t < 4; ++t) {  // bias gradient sums what is stored (bf16-rounded)
              dbacc[i * 8 + 2 * t] += bf16_lo(w[t]);
              dbacc[i * 8 + 2 * t + 1] += bf16_hi(w[t]);
            }
          } else {
            st_shared_v4(srow + v * 16, __float_as_uint(d[0]), __float_as_uint(d[1]), __float_as_uint(d[2]),
                         __float_as_uint(d[3]));
#pragma unroll
            for (int j = 0; j < 4; ++j) dbacc[i * 4 + j] += d[j];
          }
        }
      }
    }
    fence_proxy_async_smem();  // in-place gradient rows -> visible to the bulk store
    __syncthreads();
    if (threadIdx.x == 0) {
      const uint8_t* src = ce_smem + (size_t)b * buf_bytes;
      if (lddl == C) {
        bulk_store_1d(gout + r0 * lddl * ES, src, rows * row_bytes);
      } else {
        for (int r = 0; r < rows; ++r) bulk_store_1d(gout + (r0 + r) * lddl * ES, src + r * row_bytes, row_bytes);
      }
      bulk_commit();
    }
  }
  if (threadIdx.x == 0) bulk_wait_all();
  __syncthreads();

  // CTA-level reduction in a fixed order (buffer 0 is free again: 16 warps x C floats = 64*C bytes), then one
  // partial row per CTA
  float* red = reinterpret_cast<float*>(ce_smem);  // [CE_WARPS][C]
  float* prow = partials + (long long)blockIdx.x * (C + 1);
  if (want_dbias) {
#pragma unroll
    for (int i = 0; i < ITERS; ++i) {
      const int v = lane + 32 * i;
      if (v < nvec) {
#pragma unroll
        for (int j = 0; j < VEC; ++j) red[warp * C + v * VEC + j] = dbacc[i * VEC + j];
      }
    }
  }
  if (lane == 0) red_loss[warp] = loss_acc;
  __syncthreads();
  if (want_dbias) {
    for (int c = threadIdx.x; c < C; c += CE_THREADS) {
      float s = 0.f;
#pragma unroll
      for (int w = 0; w < CE_WARPS; ++w) s += red[w * C + c];
      prow[c] = s;
    }
  }
  if (threadIdx.x == 0) {
    float s = 0.f;
#pragma unroll
    for (int w = 0; w < CE_WARPS; ++w) s += red_loss[w];
    prow[C] = s;
  }
}

// out[c] (+)= sum over partial rows in a fixed order.  Columns [loss_col, loss_col + loss_ncols) (if loss_col >= 0;
// they must sit inside one 32-column block) are loss partials instead: their totals are added up in column order,
// scaled, and written/added to loss_out.
// One CTA per 32 columns: warp w walks rows w, w+8, ... (128 contiguous bytes per warp and row, eight
// independent loads in flight), then the 8 per-warp sums are combined in a fixed order -> deterministic.
constexpr int RED_THREADS = 256;
__global__ void __launch_bounds__(RED_THREADS)
reduce_partials_kernel(const float* __restrict__ partials, int nrows, int row_stride, int ncols,
                       float* __restrict__ out, int loss_col, int loss_ncols, float loss_scale,
                       float* __restrict__ loss_out, int accumulate, const int32_t* __restrict__ n_valid,
                       float batch_rows) {
  pdl_launch_dependents();
  pdl_wait();
  // mean over the non-ignored rows: 1/batch -> 1/n_valid (no valid row at all gives 0 * inf = NaN, like torch)
  if (n_valid != nullptr) loss_scale = loss_scale * batch_rows / (float)(*n_valid);
  __shared__ float part[RED_THREADS / 32][33];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int c = blockIdx.x * 32 + lane;
  const bool is_loss = loss_col >= 0 && c >= loss_col && c < loss_col + loss_ncols;
  const bool live = c < row_stride && (c < ncols || is_loss) && (is_loss || out != nullptr);
  float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
  if (live) {
    const float* col = partials + c;
    int r = warp;
    for (; r + 56 < nrows; r += 64) {  // eight independent loads in flight per lane
      const float a0 = col[(long long)(r + 0) * row_stride], a1 = col[(long long)(r + 8) * row_stride];
      const float a2 = col[(long long)(r + 16) * row_stride], a3 = col[(long long)(r + 24) * row_stride];
      const float a4 = col[(long long)(r + 32) * row_stride], a5 = col[(long long)(r + 40) * row_stride];
      const float a6 = col[(long long)(r + 48) * row_stride], a7 = col[(long long)(r + 56) * row_stride];
      s0 += a0; s1 += a1; s2 += a2; s3 += a3;
      s0 += a4; s1 += a5; s2 += a6; s3 += a7;
    }
    for (; r < nrows; r += 8) s0 += col[(long long)r * row_stride];
  }
  part[warp][lane] = (s0 + s1) + (s2 + s3);
  __syncthreads();
  if (warp == 0) {
    float s = 0.f;
    if (live) {
#pragma unroll
      for (int w = 0; w < RED_THREADS / 32; ++w) s += part[w][lane];
    }
    if (__any_sync(0xffffffffu, is_loss)) {  // the block that holds the loss columns
      float total = 0.f;
      for (int l = 0; l < 32; ++l) {  // column order
        const float v = __shfl_sync(0xffffffffu, is_loss ? s : 0.f, l);
        total += v;
      }
      if (lane == 0 && loss_out) loss_out[0] = (accumulate ? loss_out[0] : 0.f) + total * loss_scale;
    }
    if (live && !is_loss) out[c] = (accumulate ? out[c] : 0.f) + s;
  }
}

int launch_reduce_partials_loss(const float* partials, int nrows, int row_stride, int ncols, float* out, int loss_col,
                                int loss_ncols, float loss_scale, float* loss_out, int accumulate,
                                const int32_t* n_valid, float batch_rows, cudaStream_t st) {
  B200_REQUIRE(loss_col < 0 || (loss_col % 32) + loss_ncols <= 32, "loss columns must sit inside one 32-column block");
  B200_CHECK_CUDA(launch_pdl(reduce_partials_kernel, dim3((row_stride + 31) / 32), dim3(RED_THREADS), 0, st, partials,
                             nrows, row_stride, ncols, out, loss_col, loss_ncols, loss_scale, loss_out, accumulate,
                             n_valid, batch_rows));
  B200_CHECK_CUDA(cudaGetLastError());
  return 0;
}

int launch_reduce_partials(const float* partials, int nrows, int row_stride, int ncols, float* out,
                           cudaStream_t st) {
  return launch_reduce_partials_loss(partials, nrows, row_stride, ncols, out, -1, 0, 0.f, nullptr, 0, nullptr, 0.f, st);
}

// ---- column sum (hidden-layer bias gradients) -------------------------------------------------------
constexpr int CS_THREADS = 256;
constexpr int CS_MAX_BLOCKS = 1024;

// Each CTA sums a contiguous slab of rows; thread t owns one 16-byte column vector (VEC columns) of a
// 256-vector-wide column panel, so every warp reads 512 contiguous bytes per row.
template <bool kBf16>
__global__ void __launch_bounds__(CS_THREADS)
colsum_kernel(const void* __restrict__ dy_, long long ld, float* __restrict__ partials, long long M, int N,
              int rows_per_block, int panels) {
  pdl_launch_dependents();
  pdl_wait();
  constexpr int VEC = kBf16 ? 8 : 4;
  const int panel = blockIdx.x % panels;
  const int slab = blockIdx.x / panels;
  const int vcol = panel * CS_THREADS + threadIdx.x;  // vector column
  const int nvec = N / VEC;
  const long long r0 = (long long)slab * rows_per_block;
  const long long r1 = min(M, r0 + rows_per_block);
  float acc[VEC];
#pragma unroll
  for (int j = 0; j < VEC; ++j) acc[j] = 0.f;
  if (vcol < nvec) {
    long long r = r0;
    for (; r + 4 <= r1; r += 4) {  // 4 independent loads in flight
      uint4 raw[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        if constexpr (kBf16)
          raw[u] = ldg_stream(reinterpret_cast<const uint16_t*>(dy_) + (r + u) * ld + vcol * 8);
        else
          raw[u] = ldg_stream(reinterpret_cast<const float*>(dy_) + (r + u) * ld + vcol * 4);
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const uint32_t w[4] = {raw[u].x, raw[u].y, raw[u].z, raw[u].w};
        if constexpr (kBf16) {
#pragma unroll
          for (int t = 0; t < 4; ++t) {
            acc[2 * t] += bf16_lo(w[t]);
            acc[2 * t + 1] += bf16_hi(w[t]);
          }
        } else {
#pragma unroll
          for (int t = 0; t < 4; ++t) acc[t] += __uint_as_float(w[t]);
        }
      }
    }
    for (; r < r1; ++r) {
      uint4 raw;
      if constexpr (kBf16)
        raw = ldg_stream(reinterpret_cast<const uint16_t*>(dy_) + r * ld + vcol * 8);
      else
        raw = ldg_stream(reinterpret_cast<const float*>(dy_) + r * ld + vcol * 4);
      const uint32_t w[4] = {raw.x, raw.y, raw.z, raw.w};
      if constexpr (kBf16) {
#pragma unroll
        for (int t = 0; t < 4; ++t) {
          acc[2 * t] += bf16_lo(w[t]);
          acc[2 * t + 1] += bf16_hi(w[t]);
        }
      } else {
#pragma unroll
        for (int t = 0; t < 4; ++t) acc[t] += __uint_as_float(w[t]);
      }
    }
    float* prow = partials + (long long)slab * N + vcol * VEC;
#pragma unroll
    for (int j = 0; j < VEC; ++j) prow[j] = acc[j];
  }
}

}  // namespace b200

using namespace b200;

extern "C" int64_t b200_ce_scratch_floats(int num_classes) {
  return (int64_t)CE_MAX_BLOCKS * (num_classes + 1);
}

extern "C" int b200_ce_loss_grad(const void* logits, int64_t ldl, const int64_t* labels, void* dlogits,
                                 int64_t lddl, float* loss_out, float* dbias, float* scratch,
                                 int64_t scratch_floats, int64_t batch, int num_classes, float grad_scale,
                                 float loss_scale, int accumulate, int64_t ignore_index, const int32_t* n_valid,
                                 int dtype, b200_stream_t stream) {
  B200_REQUIRE(dtype == B200_F32 || dtype == B200_BF16, "unknown dtype code %d", dtype);
  const int vec = dtype == B200_F32 ? 4 : 8;
  B200_REQUIRE(num_classes > 0 && num_classes <= CE_MAX_CLASSES && num_classes % vec == 0,
               "num_classes=%d must be a multiple of %d and <= %d", num_classes, vec, CE_MAX_CLASSES);
  B200_REQUIRE(ldl % vec == 0 && lddl % vec == 0, "logit row strides must keep rows 16-byte aligned");
  B200_REQUIRE(batch > 0, "batch must be positive");
  B200_REQUIRE(scratch != nullptr, "scratch must not be NULL");
  const int tile_rows = dtype == B200_BF16 ? 32 : 16;
  const long long num_tiles = (batch + tile_rows - 1) / tile_rows;
  int grid = (int)(num_tiles < sm_count() ? num_tiles : sm_count());
  B200_REQUIRE(grid <= CE_MAX_BLOCKS, "device has more SMs than the scratch sizing assumes");
  B200_REQUIRE(scratch_floats >= (int64_t)grid * (num_classes + 1),
               "scratch too small: %lld floats, need %lld", (long long)scratch_floats,
               (long long)grid * (num_classes + 1));
  B200_REQUIRE(((reinterpret_cast<uintptr_t>(logits) | reinterpret_cast<uintptr_t>(dlogits)) & 15) == 0,
               "logits / dlogits must be 16-byte aligned");
  const size_t smem = (size_t)CE_BUFS * 64 * num_classes;  // 3 tiles of 64*C bytes
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (dtype == B200_BF16) {
    static bool cfg = false;
    if (!cfg) {
      B200_CHECK_CUDA(cudaFuncSetAttribute(ce_loss_grad_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, CE_BUFS * 64 * CE_MAX_CLASSES));
      cfg = true;
    }
    B200_CHECK_CUDA(launch_pdl(ce_loss_grad_kernel<true>, dim3(grid), dim3(CE_THREADS), smem, st, logits, ldl, labels,
                               dlogits, lddl, scratch, batch, num_classes, grad_scale, (int)(dbias != nullptr),
                               (long long)ignore_index, n_valid, status_word()));
  } else {
    static bool cfg = false;
    if (!cfg) {
      B200_CHECK_CUDA(cudaFuncSetAttribute(ce_loss_grad_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, CE_BUFS * 64 * CE_MAX_CLASSES));
      cfg = true;
    }
    B200_CHECK_CUDA(launch_pdl(ce_loss_grad_kernel<false>, dim3(grid), dim3(CE_THREADS), smem, st, logits, ldl, labels,
                               dlogits, lddl, scratch, batch, num_classes, grad_scale, (int)(dbias != nullptr),
                               (long long)ignore_index, n_valid, status_word()));
  }
  B200_CHECK_CUDA(cudaGetLastError());
  const int stride = num_classes + 1;
  return launch_reduce_partials_loss(scratch, grid, stride, num_classes, dbias, num_classes, 1,
                                     loss_scale > 0.f ? loss_scale : 1.0f / (float)batch, loss_out, accumulate, n_valid,
                                     (float)batch, st);
}

extern "C" int64_t b200_colsum_scratch_floats(int64_t N) { return (int64_t)CS_MAX_BLOCKS * N; }

extern "C" int b200_colsum(const void* dy, int64_t lddy, float* out, float* scratch, int64_t scratch_floats,
                           int64_t M, int64_t N, int dtype, b200_stream_t stream) {
  B200_REQUIRE(dtype == B200_F32 || dtype == B200_BF16, "unknown dtype code %d", dtype);
  const int vec = dtype == B200_F32 ? 4 : 8;
  B200_REQUIRE(M > 0 && N > 0 && N % vec == 0 && lddy % vec == 0, "colsum: N and lddy must be multiples of %d", vec);
  B200_REQUIRE(N < (1 << 30), "colsum: N too large");
  B200_REQUIRE(dy != nullptr && out != nullptr && (reinterpret_cast<uintptr_t>(dy) & 15) == 0,
               "colsum: dy and out must not be NULL and dy must be 16-byte aligned (128-bit loads)");
  const int nvec = (int)(N / vec);
  const int panels = (nvec + CS_THREADS - 1) / CS_THREADS;
  int slabs = (sm_count() * 4 + panels - 1) / panels;
  if (slabs > CS_MAX_BLOCKS / panels) slabs = CS_MAX_BLOCKS / panels;
  if (slabs < 1) slabs = 1;
  if (slabs > M) slabs = (int)M;
  const int rows_per_block = (int)((M + slabs - 1) / slabs);
  slabs = (int)((M + rows_per_block - 1) / rows_per_block);
  B200_REQUIRE(scratch != nullptr && scratch_floats >= (int64_t)slabs * N, "colsum scratch too small");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (dtype == B200_BF16)
    B200_CHECK_CUDA(launch_pdl(colsum_kernel<true>, dim3(slabs * panels), dim3(CS_THREADS), 0, st, dy, lddy, scratch, M, (int)N, rows_per_block, panels));
  else
    B200_CHECK_CUDA(launch_pdl(colsum_kernel<false>, dim3(slabs * panels), dim3(CS_THREADS), 0, st, dy, lddy, scratch, M, (int)N, rows_per_block, panels));
  B200_CHECK_CUDA(cudaGetLastError());
  return launch_reduce_partials(scratch, slabs, (int)N, (int)N, out, st);
}

// ---- column sum without any allocation --------------------------------------------------------------------
// Some hosts of this library cannot hand it scratch memory and must not see it allocate either: a thin C
// binding that only forwards the tensors it was given has no allocator to call; a process whose framework pool
// (PyTorch's caching allocator, for instance) has already reserved the card's HBM can see a raw cudaMalloc fail
// or fragment that pool; and cudaMalloc / cudaFree synchronise the whole device, which stalls the gradient
// all-reduce running on the side stream.  For them the partial sums of the column sum live in this file-scope
// device array instead: 1 MiB, part of the loaded module, held by every GPU that loads the library until the
// process exits.
constexpr long long CS_STATIC_FLOATS = 256 * 1024;  // 1 MiB of fp32 partials
constexpr int CS_STATIC_MAX_N = 4096;               // widest row the static variant accepts
constexpr int CS_STATIC_MAX_DEVICES = 64;
static __device__ __align__(256) float g_colsum_static_scratch[CS_STATIC_FLOATS];

// The array's name is only a host-side handle in host code; the address a kernel can use has to be asked for,
// per device, with that device current.
static int colsum_static_scratch_ptr(float** out) {
  static float* ptrs[CS_STATIC_MAX_DEVICES] = {nullptr};
  int dev = -1;
  B200_CHECK_CUDA(cudaGetDevice(&dev));
  B200_REQUIRE(dev >= 0 && dev < CS_STATIC_MAX_DEVICES, "device ordinal %d out of range", dev);
  if (ptrs[dev] == nullptr) {
    float* p = nullptr;
    B200_CHECK_CUDA(cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_colsum_static_scratch));
    B200_REQUIRE(p != nullptr, "cudaGetSymbolAddress returned a null scratch address");
    ptrs[dev] = p;
  }
  *out = ptrs[dev];
  return 0;
}

extern "C" int b200_colsum_static(const void* dy, int64_t lddy, float* out, int64_t M, int64_t N, int dtype,
                                  b200_stream_t stream) {
  B200_REQUIRE(dtype == B200_F32 || dtype == B200_BF16, "unknown dtype code %d", dtype);
  const int vec = dtype == B200_F32 ? 4 : 8;
  B200_REQUIRE(M > 0 && N > 0 && N % vec == 0 && lddy % vec == 0, "colsum: N and lddy must be multiples of %d", vec);
  B200_REQUIRE(dy != nullptr && out != nullptr && (reinterpret_cast<uintptr_t>(dy) & 15) == 0,
               "colsum: dy and out must not be NULL and dy must be 16-byte aligned (128-bit loads)");
  B200_REQUIRE(N <= CS_STATIC_MAX_N,
               "b200_colsum_static: N=%lld exceeds the static scratch limit of %d columns; use b200_colsum with "
               "caller-provided scratch", (long long)N, CS_STATIC_MAX_N);
  float* scratch = nullptr;
  if (int r = colsum_static_scratch_ptr(&scratch)) return r;  // first call on this device does the set-up
  const int nvec = (int)(N / vec);
  const int panels = (nvec + CS_THREADS - 1) / CS_THREADS;
  long long slabs = ((long long)sm_count() * 4 + panels - 1) / panels;
  if (slabs > CS_STATIC_FLOATS / N) slabs = CS_STATIC_FLOATS / N;  // partial rows must fit the static array
  if (slabs > M) slabs = M;
  if (slabs < 1) slabs = 1;
  const int rows_per_block = (int)((M + slabs - 1) / slabs);
  slabs = (M + rows_per_block - 1) / rows_per_block;
  B200_REQUIRE(slabs * N <= CS_STATIC_FLOATS, "b200_colsum_static: internal sizing error");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  // every partial row that the second kernel reads is written by the first one in this same call, so nothing
  // in the shared array has to start at zero
  if (dtype == B200_BF16)
    B200_CHECK_CUDA(launch_pdl(colsum_kernel<true>, dim3((int)slabs * panels), dim3(CS_THREADS), 0, st, dy, lddy, scratch, M, (int)N, rows_per_block, panels));
  else
    B200_CHECK_CUDA(launch_pdl(colsum_kernel<false>, dim3((int)slabs * panels), dim3(CS_THREADS), 0, st, dy, lddy, scratch, M, (int)N, rows_per_block, panels));
  B200_CHECK_CUDA(cudaGetLastError());
  return launch_reduce_partials(scratch, (int)slabs, (int)N, (int)N, out, st);
}
