// preprocess.cu — K-pre: fused batch load + standardise + categorical gather (+ label / index gather).
//
// Replaces, for one batch, the reference's sampler/loader batching (task.py:260-263), the Normalize
// transform (task.py:250-251: tensor.sub_(mean).div_(std)), the H2D copies (task.py:305,307) and the
// embedding lookup + concat at the front of the tabular model.  Results are bit-exact
// with the CPU oracle: IEEE subtract then divide (no FMA contraction, no reciprocal), RNE to bf16.
//
// Layout: one CTA assembles up to 32 consecutive output rows in shared memory (embedding columns start at
// element num_dense, which is not 16-byte aligned, so the row is built with scalar shared stores), then the
// whole tile — a contiguous TILE*ldx block of x_out — leaves through one 1-D TMA bulk store.  Table rows are
// read with 128-bit streaming loads, one float4 per lane.
#include "common.cuh"

namespace b200 {

constexpr int PRE_THREADS = 256;
constexpr int PRE_MAX_TILE = 32;           // samples per CTA (fewer when rows are wide)
constexpr int PRE_SMEM_TARGET = 96 * 1024;  // keeps >= 2 CTAs per SM resident

template <typename OutT>
__device__ __forceinline__ OutT cvt_out(float v);
template <>
__device__ __forceinline__ float cvt_out<float>(float v) { return v; }
template <>
__device__ __forceinline__ __nv_bfloat16 cvt_out<__nv_bfloat16>(float v) { return __float2bfloat16_rn(v); }

template <typename OutT>
__global__ void __launch_bounds__(PRE_THREADS)
preprocess_kernel(const float* __restrict__ ds_dense, const int64_t* __restrict__ ds_cat,
                  const int64_t* __restrict__ ds_label, const int64_t* __restrict__ row_idx,
                  const int32_t* __restrict__ cursor, const float* __restrict__ mean,
                  const float* __restrict__ stdv, const float* __restrict__ tables,
                  OutT* __restrict__ x_out, int64_t* __restrict__ y_out, int32_t* __restrict__ idx_out,
                  long long batch, int num_dense, int num_cat, long long table_rows, int embed_dim,
                  long long ldx, int tile_rows) {
  extern __shared__ __align__(128) uint8_t pre_smem[];
  OutT* tile = reinterpret_cast<OutT*>(pre_smem);  // [tile_rows][ldx]

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long b0 = (long long)blockIdx.x * tile_rows;
  const long long cur = cursor ? (long long)(*cursor) : 0;
  const int rows_here = (int)min((long long)tile_rows, batch - b0);
  const int vec_per_row = embed_dim >> 2;           // float4 chunks per table row
  const int vec_total = num_cat * vec_per_row;      // float4 chunks per sample
  const int used = num_dense + num_cat * embed_dim; // first padding column

  for (int s = warp; s < rows_here; s += PRE_THREADS / 32) {
    const long long b = b0 + s;
    const long long r = __ldg(row_idx + cur * batch + b);
    OutT* xr = tile + (long long)s * ldx;
    // dense columns: (x - mean) / std, IEEE fp32
    for (int j = lane; j < num_dense; j += 32) {
      const float x = __ldg(ds_dense + r * num_dense + j);
      xr[j] = cvt_out<OutT>(__fdiv_rn(__fsub_rn(x, __ldg(mean + j)), __ldg(stdv + j)));
    }
    // K padding
    for (int j = used + lane; j < ldx; j += 32) xr[j] = cvt_out<OutT>(0.0f);
    // categorical ids: lane t holds id of table t (+32, +64 ... for > 32 tables handled by the loop below)
    for (int tb = 0; tb < num_cat; tb += 32) {
      const int t = tb + lane;
      long long id = 0;
      if (t < num_cat) {
        id = __ldg(ds_cat + r * num_cat + t);
        idx_out[b * num_cat + t] = (int32_t)id;
      }
      const int tcount = min(32, num_cat - tb);
      // gather: chunk i -> table tb + i / vec_per_row, float4 i % vec_per_row of that row
      for (int i0 = 0; i0 < tcount * vec_per_row; i0 += 32) {
        const int i = i0 + lane;
        const int tt = i / vec_per_row;
        const int qv = i - tt * vec_per_row;
        const long long idt = __shfl_sync(0xffffffffu, id, tt < tcount ? tt : 0);
        if (i < tcount * vec_per_row) {
          const float* src = tables + ((long long)(tb + tt) * table_rows + idt) * embed_dim + qv * 4;
          const uint4 raw = ldg_stream(src);
          OutT* dst = xr + num_dense + (tb + tt) * embed_dim + qv * 4;
          dst[0] = cvt_out<OutT>(__uint_as_float(raw.x));
          dst[1] = cvt_out<OutT>(__uint_as_float(raw.y));
          dst[2] = cvt_out<OutT>(__uint_as_float(raw.z));
          dst[3] = cvt_out<OutT>(__uint_as_float(raw.w));
        }
      }
    }
    if (lane == 0) y_out[b] = __ldg(ds_label + r);
  }
  (void)vec_total;
  // hand the finished tile to the TMA engine: generic-proxy writes -> async-proxy read
  fence_proxy_async_smem();
  __syncthreads();
  if (threadIdx.x == 0) {
    bulk_store_1d(x_out + b0 * ldx, tile, (uint32_t)(rows_here * ldx * sizeof(OutT)));
    bulk_commit();
    bulk_wait_read0();  // shared memory must outlive the read
  }
}

}  // namespace b200

using namespace b200;

extern "C" int b200_preprocess(const float* ds_dense, const int64_t* ds_cat, const int64_t* ds_label,
                               const int64_t* row_idx, const int32_t* cursor, const float* mean,
                               const float* stdv, const float* tables, void* x_out, int64_t* y_out,
                               int32_t* idx_out, int64_t batch, int num_dense, int num_cat,
                               int64_t table_rows, int embed_dim, int64_t ldx, int x_dtype,
                               b200_stream_t stream) {
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
  const size_t row_bytes = (size_t)ldx * es;
  B200_REQUIRE(row_bytes <= 200 * 1024, "row of %lld elements does not fit in shared memory", (long long)ldx);
  int tile_rows = (int)(PRE_SMEM_TARGET / row_bytes);
  if (tile_rows > PRE_MAX_TILE) tile_rows = PRE_MAX_TILE;
  if (tile_rows < 1) tile_rows = 1;
  const size_t smem = (size_t)tile_rows * row_bytes;
  const unsigned grid = (unsigned)((batch + tile_rows - 1) / tile_rows);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (x_dtype == B200_F32) {
    static bool cfg = false;
    if (!cfg) {
      B200_CHECK_CUDA(cudaFuncSetAttribute(preprocess_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
      cfg = true;
    }
    preprocess_kernel<float><<<grid, PRE_THREADS, smem, st>>>(
        ds_dense, ds_cat, ds_label, row_idx, cursor, mean, stdv, tables, static_cast<float*>(x_out), y_out,
        idx_out, batch, num_dense, num_cat, table_rows, embed_dim, ldx, tile_rows);
  } else {
    static bool cfg = false;
    if (!cfg) {
      B200_CHECK_CUDA(cudaFuncSetAttribute(preprocess_kernel<__nv_bfloat16>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
      cfg = true;
    }
    preprocess_kernel<__nv_bfloat16><<<grid, PRE_THREADS, smem, st>>>(
        ds_dense, ds_cat, ds_label, row_idx, cursor, mean, stdv, tables,
        static_cast<__nv_bfloat16*>(x_out), y_out, idx_out, batch, num_dense, num_cat, table_rows,
        embed_dim, ldx, tile_rows);
  }
  B200_CHECK_CUDA(cudaGetLastError());
  return 0;
}
