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

constexpr int CE_THREADS = 256;
constexpr int CE_WARPS = CE_THREADS / 32;
constexpr int CE_MAX_BLOCKS = 1024;
constexpr int CE_MAX_CLASSES = 1024;  // 32 lanes x 32 register-resident values

template <bool kBf16>
__global__ void __launch_bounds__(CE_THREADS)
ce_loss_grad_kernel(const void* __restrict__ logits_, long long ldl, const int64_t* __restrict__ labels,
                    void* __restrict__ dlogits_, long long lddl, float* __restrict__ partials,
                    long long batch, int C, float scale, int want_dbias) {
  constexpr int VEC = kBf16 ? 8 : 4;         // elements per 16-byte vector
  constexpr int ITERS = 32 / VEC;            // vectors per lane (covers 1024 columns)
  __shared__ float red[CE_WARPS][CE_MAX_CLASSES];
  __shared__ float red_loss[CE_WARPS];

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nvec = C / VEC;
  float dbacc[32];
#pragma unroll
  for (int i = 0; i < 32; ++i) dbacc[i] = 0.f;
  float loss_acc = 0.f;

  for (long long row = (long long)blockIdx.x * CE_WARPS + warp; row < batch;
       row += (long long)gridDim.x * CE_WARPS) {
    const int y = (int)__ldg(labels + row);
    float z[32];
    float zmax = -INFINITY, zy = 0.f;
#pragma unroll
    for (int it = 0; it < ITERS; ++it) {
      const int v = lane + 32 * it;
      if (v < nvec) {
        if constexpr (kBf16) {
          const uint4 raw = ldg_stream(reinterpret_cast<const uint16_t*>(logits_) + row * ldl + v * 8);
          const uint32_t w[4] = {raw.x, raw.y, raw.z, raw.w};
#pragma unroll
          for (int t = 0; t < 4; ++t) {
            z[it * 8 + 2 * t] = bf16_lo(w[t]);
            z[it * 8 + 2 * t + 1] = bf16_hi(w[t]);
          }
        } else {
          const uint4 raw = ldg_stream(reinterpret_cast<const float*>(logits_) + row * ldl + v * 4);
          z[it * 4 + 0] = __uint_as_float(raw.x);
          z[it * 4 + 1] = __uint_as_float(raw.y);
          z[it * 4 + 2] = __uint_as_float(raw.z);
          z[it * 4 + 3] = __uint_as_float(raw.w);
        }
#pragma unroll
        for (int j = 0; j < VEC; ++j) {
          zmax = fmaxf(zmax, z[it * VEC + j]);
          if (v * VEC + j == y) zy = z[it * VEC + j];
        }
      } else {
#pragma unroll
        for (int j = 0; j < VEC; ++j) z[it * VEC + j] = -INFINITY;
      }
    }
    zmax = warp_max(zmax);
    zy = warp_sum(zy);
    float sum = 0.f;
#pragma unroll
    for (int i = 0; i < 32; ++i) {
      z[i] = expf(z[i] - zmax);  // exp(-inf) = 0 for the tail
      sum += z[i];
    }
    sum = warp_sum(sum);
    loss_acc += (zmax + logf(sum)) - zy;  // identical in every lane
    const float inv = 1.0f / sum;
#pragma unroll
    for (int it = 0; it < ITERS; ++it) {
      const int v = lane + 32 * it;
      if (v < nvec) {
        float d[VEC];
#pragma unroll
        for (int j = 0; j < VEC; ++j) {
          const float pj = z[it * VEC + j] * inv;
          d[j] = (pj - ((v * VEC + j == y) ? 1.f : 0.f)) * scale;
        }
        if constexpr (kBf16) {
          uint4 o;
          o.x = pack_bf16x2(d[0], d[1]);
          o.y = pack_bf16x2(d[2], d[3]);
          o.z = pack_bf16x2(d[4], d[5]);
          o.w = pack_bf16x2(d[6], d[7]);
          stg_stream(reinterpret_cast<uint16_t*>(dlogits_) + row * lddl + v * 8, o);
          const uint32_t w[4] = {o.x, o.y, o.z, o.w};
#pragma unroll
          for (int t = 0; t < 4; ++t) {  // bias gradient sums what was stored (bf16-rounded)
            dbacc[it * 8 + 2 * t] += bf16_lo(w[t]);
            dbacc[it * 8 + 2 * t + 1] += bf16_hi(w[t]);
          }
        } else {
          stg_stream(reinterpret_cast<float*>(dlogits_) + row * lddl + v * 4,
                     make_uint4(__float_as_uint(d[0]), __float_as_uint(d[1]), __float_as_uint(d[2]),
                                __float_as_uint(d[3])));
#pragma unroll
          for (int j = 0; j < 4; ++j) dbacc[it * 4 + j] += d[j];
        }
      }
    }
  }

  // CTA-level reduction in a fixed order, then one partial row per CTA
  float* prow = partials + (long long)blockIdx.x * (C + 1);
  if (want_dbias) {
#pragma unroll
    for (int it = 0; it < ITERS; ++it) {
      const int v = lane + 32 * it;
      if (v < nvec) {
#pragma unroll
        for (int j = 0; j < VEC; ++j) red[warp][v * VEC + j] = dbacc[it * VEC + j];
      }
    }
  }
  if (lane == 0) red_loss[warp] = loss_acc;
  __syncthreads();
  if (want_dbias) {
    for (int c = threadIdx.x; c < C; c += CE_THREADS) {
      float s = 0.f;
#pragma unroll
      for (int w = 0; w < CE_WARPS; ++w) s += red[w][c];
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

// out[c] = sum over partial rows in a fixed order; column `loss_col` (if >= 0) is scaled and written to loss_out.
// One CTA per 32 columns: warp w walks rows w, w+8, ... (128 contiguous bytes per warp and row, four
// independent loads in flight), then the 8 per-warp sums are combined in a fixed order -> deterministic.
constexpr int RED_THREADS = 256;
__global__ void __launch_bounds__(RED_THREADS)
reduce_partials_kernel(const float* __restrict__ partials, int nrows, int row_stride, int ncols,
                       float* __restrict__ out, int loss_col, float loss_scale, float* __restrict__ loss_out) {
  __shared__ float part[RED_THREADS / 32][33];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int c = blockIdx.x * 32 + lane;
  const bool is_loss = (c == loss_col);
  const bool live = c < row_stride && (c < ncols || is_loss) && (is_loss || out != nullptr);
  float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
  if (live) {
    const float* col = partials + c;
    int r = warp;
    for (; r + 24 < nrows; r += 32) {
      s0 += col[(long long)(r + 0) * row_stride];
      s1 += col[(long long)(r + 8) * row_stride];
      s2 += col[(long long)(r + 16) * row_stride];
      s3 += col[(long long)(r + 24) * row_stride];
    }
    for (; r < nrows; r += 8) s0 += col[(long long)r * row_stride];
  }
  part[warp][lane] = (s0 + s1) + (s2 + s3);
  __syncthreads();
  if (warp == 0 && live) {
    float s = 0.f;
#pragma unroll
    for (int w = 0; w < RED_THREADS / 32; ++w) s += part[w][lane];
    if (is_loss) {
      if (loss_out) loss_out[0] = s * loss_scale;
    } else {
      out[c] = s;
    }
  }
}

int launch_reduce_partials(const float* partials, int nrows, int row_stride, int ncols, float* out,
                           cudaStream_t st) {
  reduce_partials_kernel<<<(row_stride + 31) / 32, RED_THREADS, 0, st>>>(partials, nrows, row_stride, ncols, out, -1, 0.f,
                                                                 nullptr);
  B200_CHECK_CUDA(cudaGetLastError());
  return 0;
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
                                 int dtype, b200_stream_t stream) {
  B200_REQUIRE(dtype == B200_F32 || dtype == B200_BF16, "unknown dtype code %d", dtype);
  const int vec = dtype == B200_F32 ? 4 : 8;
  B200_REQUIRE(num_classes > 0 && num_classes <= CE_MAX_CLASSES && num_classes % vec == 0,
               "num_classes=%d must be a multiple of %d and <= %d", num_classes, vec, CE_MAX_CLASSES);
  B200_REQUIRE(ldl % vec == 0 && lddl % vec == 0, "logit row strides must keep rows 16-byte aligned");
  B200_REQUIRE(batch > 0, "batch must be positive");
  B200_REQUIRE(scratch != nullptr, "scratch must not be NULL");
  int grid = (int)((batch + CE_WARPS - 1) / CE_WARPS);
  const int cap = sm_count() * 4 < CE_MAX_BLOCKS ? sm_count() * 4 : CE_MAX_BLOCKS;
  if (grid > cap) grid = cap;
  B200_REQUIRE(scratch_floats >= (int64_t)grid * (num_classes + 1),
               "scratch too small: %lld floats, need %lld", (long long)scratch_floats,
               (long long)grid * (num_classes + 1));
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (dtype == B200_BF16)
    ce_loss_grad_kernel<true><<<grid, CE_THREADS, 0, st>>>(logits, ldl, labels, dlogits, lddl, scratch, batch,
                                                           num_classes, grad_scale, dbias != nullptr);
  else
    ce_loss_grad_kernel<false><<<grid, CE_THREADS, 0, st>>>(logits, ldl, labels, dlogits, lddl, scratch, batch,
                                                            num_classes, grad_scale, dbias != nullptr);
  B200_CHECK_CUDA(cudaGetLastError());
  const int stride = num_classes + 1;
  reduce_partials_kernel<<<(stride + 31) / 32, RED_THREADS, 0, st>>>(scratch, grid, stride, num_classes, dbias,
                                                             num_classes, 1.0f / (float)batch, loss_out);
  B200_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int64_t b200_colsum_scratch_floats(int64_t N) { return (int64_t)CS_MAX_BLOCKS * N; }

extern "C" int b200_colsum(const void* dy, int64_t lddy, float* out, float* scratch, int64_t scratch_floats,
                           int64_t M, int64_t N, int dtype, b200_stream_t stream) {
  B200_REQUIRE(dtype == B200_F32 || dtype == B200_BF16, "unknown dtype code %d", dtype);
  const int vec = dtype == B200_F32 ? 4 : 8;
  B200_REQUIRE(M > 0 && N > 0 && N % vec == 0 && lddy % vec == 0, "colsum: N and lddy must be multiples of %d", vec);
  B200_REQUIRE(N < (1 << 30), "colsum: N too large");
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
    colsum_kernel<true><<<slabs * panels, CS_THREADS, 0, st>>>(dy, lddy, scratch, M, (int)N, rows_per_block, panels);
  else
    colsum_kernel<false><<<slabs * panels, CS_THREADS, 0, st>>>(dy, lddy, scratch, M, (int)N, rows_per_block, panels);
  B200_CHECK_CUDA(cudaGetLastError());
  reduce_partials_kernel<<<(int)((N + 31) / 32), RED_THREADS, 0, st>>>(scratch, slabs, (int)N, (int)N, out, -1, 0.f, nullptr);
  B200_CHECK_CUDA(cudaGetLastError());
  return 0;
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
    colsum_kernel<true><<<(int)slabs * panels, CS_THREADS, 0, st>>>(dy, lddy, scratch, M, (int)N, rows_per_block, panels);
  else
    colsum_kernel<false><<<(int)slabs * panels, CS_THREADS, 0, st>>>(dy, lddy, scratch, M, (int)N, rows_per_block, panels);
  B200_CHECK_CUDA(cudaGetLastError());
  reduce_partials_kernel<<<(int)((N + 31) / 32), RED_THREADS, 0, st>>>(scratch, (int)slabs, (int)N, (int)N, out, -1, 0.f, nullptr);
  B200_CHECK_CUDA(cudaGetLastError());
  return 0;
}
