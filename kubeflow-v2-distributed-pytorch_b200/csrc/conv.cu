// conv.cu — the memory-bound kernels of the convolutional path (ResNet-18 on 3x32x32, task.py:171,247-257).
//
// Layout: activations are NHWC, i.e. a row-major matrix [B*H*W, C] — exactly what the tcgen05 GEMMs of
// gemm_tcgen05.cu read and write, so a convolution is
//     im2col (this file)  ->  b200_linear_fwd   (col[M, kh*kw*C] . Wg[Cout, kh*kw*C]^T, no bias)
//     b200_linear_dgrad   ->  col2im (this file, gather form: deterministic, no atomics)
//     b200_linear_wgrad   (dY^T . col, fp32 into the flat gradient buffer)
// with the filter stored as Wg[Cout][ky][kx][c] (the permutation of torch's OIHW that makes a patch contiguous).
// BatchNorm2d in training mode (batch statistics, running-stat update, torch/nn/modules/batchnorm.py) is two passes
// over the GEMM output: statistics (fixed-order partials, last block finalises) and apply (+residual, +ReLU).
// Every kernel here is HBM-bound: 8- or 16-byte vectors of 4 channels per thread, coalesced along C.
#include <algorithm>

#include "common.cuh"

namespace b200 {

// ---- 4-channel vector access for both storage types ------------------------------------------------------------
template <typename T> struct Vec4;
template <> struct Vec4<float> {
  static __device__ __forceinline__ float4 load(const float* p) { return *reinterpret_cast<const float4*>(p); }
  static __device__ __forceinline__ void store(float* p, float4 v) { *reinterpret_cast<float4*>(p) = v; }
};
template <> struct Vec4<__nv_bfloat16> {
  static __device__ __forceinline__ float4 load(const __nv_bfloat16* p) {
    const uint2 r = *reinterpret_cast<const uint2*>(p);
    return make_float4(bf16_lo(r.x), bf16_hi(r.x), bf16_lo(r.y), bf16_hi(r.y));
  }
  static __device__ __forceinline__ void store(__nv_bfloat16* p, float4 v) {
    *reinterpret_cast<uint2*>(p) = make_uint2(pack_bf16x2(v.x, v.y), pack_bf16x2(v.z, v.w));
  }
};
template <typename T> __device__ __forceinline__ T from_float(float v);
template <> __device__ __forceinline__ float from_float<float>(float v) { return v; }
template <> __device__ __forceinline__ __nv_bfloat16 from_float<__nv_bfloat16>(float v) { return __float2bfloat16_rn(v); }
template <typename T> __device__ __forceinline__ float to_float(T v);
template <> __device__ __forceinline__ float to_float<float>(float v) { return v; }
template <> __device__ __forceinline__ float to_float<__nv_bfloat16>(__nv_bfloat16 v) { return __bfloat162float(v); }

static inline int grid_for(int64_t work_items, int threads, int per_sm = 8) {
  const int64_t blocks = (work_items + threads - 1) / threads;
  const int64_t cap = (int64_t)sm_count() * per_sm;
  return (int)(blocks < 1 ? 1 : (blocks > cap ? cap : blocks));
}

// ---- K-aug: RandomCrop(32, padding=4) + RandomHorizontalFlip + ToTensor + Normalize (task.py:250-255) -----------
// One thread per output pixel.  draws[p] = {top, left, flip, -}: the three numbers torchvision draws per sample
// (transforms.py RandomCrop.get_params: randint(0, 9) twice; RandomHorizontalFlip: rand(1) < 0.5), drawn on the host
// with torch's generator in the same order.  Arithmetic is ToTensor's u8 -> fp32 `/255` and Normalize's
// `(x - mean) / std`, each correctly rounded (no FMA contraction), so the fp32 result is bit-identical to torchvision.
template <typename T>
__global__ void __launch_bounds__(256)
augment_kernel(const uint8_t* __restrict__ images, const int64_t* __restrict__ labels, const int64_t* __restrict__ row_idx,
               const int32_t* __restrict__ cursor, int64_t cursor_stride, const int8_t* __restrict__ draws,
               int64_t n_rows, float m0, float m1, float m2, float s0, float s1, float s2, int pad,
               T* __restrict__ x, int64_t* __restrict__ y, int B, int H, int W, uint32_t* status) {
  pdl_launch_dependents();
  pdl_wait();
  const int64_t base = cursor != nullptr ? (int64_t)cursor[0] * cursor_stride : 0;
  const int64_t total = (int64_t)B * H * W;
  for (int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
    const int w = (int)(t % W), h = (int)((t / W) % H);
    const int64_t b = t / ((int64_t)W * H);
    const int64_t pos = base + b;
    int64_t r = row_idx != nullptr ? row_idx[pos] : pos;
    if (r < 0 || r >= n_rows) { flag_status(status, B200_STATUS_BAD_ROW_INDEX); r = 0; }
    const int top = draws != nullptr ? draws[pos * 4 + 0] : pad;
    const int left = draws != nullptr ? draws[pos * 4 + 1] : pad;
    const int flip = draws != nullptr ? draws[pos * 4 + 2] : 0;
    // output pixel (h, w) of the flipped crop = pixel (top + h, left + (flip ? W-1-w : w)) of the padded image
    const int sh = top + h - pad, sw = left + (flip ? W - 1 - w : w) - pad;
    float p0 = 0.f, p1 = 0.f, p2 = 0.f;     // RandomCrop pads with fill = 0 BEFORE ToTensor / Normalize
    if (sh >= 0 && sh < H && sw >= 0 && sw < W) {
      const uint8_t* px = images + ((r * H + sh) * (int64_t)W + sw) * 3;
      p0 = __fdiv_rn((float)px[0], 255.f); p1 = __fdiv_rn((float)px[1], 255.f); p2 = __fdiv_rn((float)px[2], 255.f);
    }
    T* o = x + t * 3;
    o[0] = from_float<T>(__fdiv_rn(__fsub_rn(p0, m0), s0));
    o[1] = from_float<T>(__fdiv_rn(__fsub_rn(p1, m1), s1));
    o[2] = from_float<T>(__fdiv_rn(__fsub_rn(p2, m2), s2));
    if (h == 0 && w == 0 && y != nullptr) y[b] = labels[r];
  }
}

// ---- im2col ------------------------------------------------------------------------------------------------------
// col[m, (ky*kw + kx)*C + c] = x[b, ho*stride - pad + ky, wo*stride - pad + kx, c]   (0 outside the image),
// m = (b*Ho + ho)*Wo + wo; columns [kh*kw*C, ldcol) are zero-filled (the GEMM's K padding).
// VEC = 4: one thread moves 4 channels (C % 4 == 0); VEC = 1: scalar (the 3-channel input of conv1).
template <typename T, int VEC>
__global__ void __launch_bounds__(256)
im2col_kernel(const T* __restrict__ x, T* __restrict__ col, int B, int H, int W, int C, int Ho, int Wo, int kh, int kw,
              int stride, int pad, int64_t ldcol, int kcols) {
  pdl_launch_dependents();
  pdl_wait();
  const int K = kh * kw * C;
  const int vrow = kcols / VEC;                       // vectors per col row (padding included)
  const int64_t total = (int64_t)B * Ho * Wo * vrow;
  for (int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
    const int k = (int)(t % vrow) * VEC;
    const int64_t m = t / vrow;
    const int wo = (int)(m % Wo), ho = (int)((m / Wo) % Ho);
    const int64_t b = m / ((int64_t)Wo * Ho);
    T* dst = col + m * ldcol + k;
    bool inside = false;
    const T* src = nullptr;
    if (k < K) {
      const int tap = k / C, c = k - tap * C;
      const int ky = tap / kw, kx = tap - ky * kw;
      const int h = ho * stride - pad + ky, w = wo * stride - pad + kx;
      inside = h >= 0 && h < H && w >= 0 && w < W;
      src = x + ((b * H + h) * (int64_t)W + w) * C + c;
    }
    if (VEC == 4) {
      if (sizeof(T) == 4) {
        *reinterpret_cast<float4*>(dst) = inside ? *reinterpret_cast<const float4*>(src) : make_float4(0, 0, 0, 0);
      } else {
        *reinterpret_cast<uint2*>(dst) = inside ? *reinterpret_cast<const uint2*>(src) : make_uint2(0, 0);
      }
    } else {
      *dst = inside ? *src : from_float<T>(0.f);
    }
  }
}

// Fast path, C / VEC a power of two (every ResNet layer but conv1): one warp per output row, lanes stride over the
// row's 16-byte vectors; tap and channel of a vector come from a shift / mask and one multiply-shift division.
template <typename T>
__global__ void __launch_bounds__(256)
im2col_vec_kernel(const T* __restrict__ x, T* __restrict__ col, int B, int H, int W, int C, int Ho, int Wo, int kh, int kw,
                  int stride, int pad, int64_t ldcol, int cshift /*log2(C / VEC)*/) {
  pdl_launch_dependents();
  pdl_wait();
  constexpr int VEC = 16 / (int)sizeof(T);
  const int taps = kh * kw;
  const int vrow = (int)(ldcol / VEC);
  const int vk = taps << cshift;                       // vectors that carry data; the rest of the row is padding
  const uint32_t inv_kw = 65536u / (uint32_t)kw + 1u;  // tap / kw for tap < 64: exact
  const int64_t rows = (int64_t)B * Ho * Wo;
  const int lane = threadIdx.x & 31;
  const int64_t warp0 = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int64_t nwarps = (int64_t)gridDim.x * (blockDim.x >> 5);
  for (int64_t m = warp0; m < rows; m += nwarps) {
    const int wo = (int)(m % Wo), ho = (int)((m / Wo) % Ho);
    const int64_t b = m / ((int64_t)Wo * Ho);
    const int h0 = ho * stride - pad, w0 = wo * stride - pad;
    const T* img = x + b * H * (int64_t)W * C;
    uint4* dst = reinterpret_cast<uint4*>(col + m * ldcol);
#pragma unroll 4
    for (int v = lane; v < vrow; v += 32) {
      uint4 val = make_uint4(0, 0, 0, 0);
      if (v < vk) {
        const int tap = v >> cshift, cv = v & ((1 << cshift) - 1);
        const int ky = (int)(((uint32_t)tap * inv_kw) >> 16), kx = tap - ky * kw;
        const int h = h0 + ky, w = w0 + kx;
        if (h >= 0 && h < H && w >= 0 && w < W)
          val = *reinterpret_cast<const uint4*>(img + ((int64_t)h * W + w) * C + cv * VEC);
      }
      dst[v] = val;
    }
  }
}

// Small channel counts (conv1: C = 3, K = 147): a block builds IM_ROWS col rows in shared memory — one (row, tap)
// pair per thread step, C scalar copies each — and writes them out as coalesced 16-byte vectors.
constexpr int IM_ROWS = 32;
template <typename T>
__global__ void __launch_bounds__(256)
im2col_smallc_kernel(const T* __restrict__ x, T* __restrict__ col, int B, int H, int W, int C, int Ho, int Wo, int kh,
                     int kw, int stride, int pad, int64_t ldcol) {
  pdl_launch_dependents();
  pdl_wait();
  extern __shared__ __align__(16) unsigned char im_smem[];
  T* tile = reinterpret_cast<T*>(im_smem);              // [IM_ROWS][ldcol]
  constexpr int VEC = 16 / (int)sizeof(T);
  const int taps = kh * kw, K = taps * C;
  const int64_t rows = (int64_t)B * Ho * Wo;
  const int kcols = (int)ldcol;
  const uint32_t inv_kw = 65536u / (uint32_t)kw + 1u;     // tap / kw for tap < 64: exact (the host checks kh*kw < 64)
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarp = blockDim.x >> 5;
  for (int64_t m0 = (int64_t)blockIdx.x * IM_ROWS; m0 < rows; m0 += (int64_t)gridDim.x * IM_ROWS) {
    const int nr = (int)min((int64_t)IM_ROWS, rows - m0);
    // a warp per col row: the row's pixel coordinates are computed once, the lanes walk its taps
    for (int r = warp; r < nr; r += nwarp) {
      const int64_t m = m0 + r;
      const int wo = (int)(m % Wo), ho = (int)((m / Wo) % Ho);
      const int64_t b = m / ((int64_t)Wo * Ho);
      const int h0 = ho * stride - pad, w0 = wo * stride - pad;
      const T* img = x + b * H * (int64_t)W * C;
      T* drow = tile + r * kcols;
      for (int tap = lane; tap < taps; tap += 32) {
        const int ky = (int)(((uint32_t)tap * inv_kw) >> 16), kx = tap - ky * kw;
        const int h = h0 + ky, w = w0 + kx;
        const bool inside = h >= 0 && h < H && w >= 0 && w < W;
        const T* src = img + ((int64_t)h * W + w) * C;
        T* d = drow + tap * C;
        for (int c = 0; c < C; ++c) d[c] = inside ? src[c] : from_float<T>(0.f);
      }
    }
    for (int i = threadIdx.x; i < nr * (kcols - K); i += blockDim.x) {     // the GEMM's K padding
      const int r = i / (kcols - K);
      tile[r * kcols + K + (i - r * (kcols - K))] = from_float<T>(0.f);
    }
    __syncthreads();
    const int nvec = nr * kcols / VEC;                   // rows are contiguous in col when ldcol == kcols
    uint4* dst = reinterpret_cast<uint4*>(col + m0 * ldcol);
    const uint4* srcv = reinterpret_cast<const uint4*>(tile);
    for (int i = threadIdx.x; i < nvec; i += blockDim.x) dst[i] = srcv[i];
    __syncthreads();
  }
}

// ---- col2im (gather form) -------------------------------------------------------------------------------------------
// dx[b,h,w,c] = sum over taps (ky,kx) with (h + pad - ky) = ho*stride, (w + pad - kx) = wo*stride in range of
//               dcol[(b,ho,wo), (ky*kw+kx)*C + c]   (+ add[b,h,w,c] when `add` is given: the residual branch's gradient)
// Every dcol element is read exactly once; sums are taken in tap order in fp32.
template <typename T>
__global__ void __launch_bounds__(256)
col2im_kernel(const T* __restrict__ dcol, int64_t ldcol, const T* __restrict__ add, T* __restrict__ dx, int B, int H,
              int W, int C, int Ho, int Wo, int kh, int kw, int stride, int pad) {
  pdl_launch_dependents();
  pdl_wait();
  const int vc = C / 4;
  const int64_t total = (int64_t)B * H * W * vc;
  for (int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
    const int c = (int)(t % vc) * 4;
    const int64_t pix = t / vc;
    const int w = (int)(pix % W), h = (int)((pix / W) % H);
    const int64_t b = pix / ((int64_t)W * H);
    float4 acc = add != nullptr ? Vec4<T>::load(add + pix * C + c) : make_float4(0, 0, 0, 0);
    // only the windows that contain this pixel: ho in [ceil((h + pad - kh + 1) / stride), floor((h + pad) / stride)]
    const int hn = h + pad - kh + 1, wn = w + pad - kw + 1;
    const int hlo = hn <= 0 ? 0 : (hn + stride - 1) / stride, wlo = wn <= 0 ? 0 : (wn + stride - 1) / stride;
    const int hhi = min(Ho - 1, (h + pad) / stride), whi = min(Wo - 1, (w + pad) / stride);
    for (int ho = hhi; ho >= hlo; --ho) {                 // descending ho = ascending tap: the order of the sums is kept
      const int ky = h + pad - ho * stride;
      for (int wo = whi; wo >= wlo; --wo) {
        const int kx = w + pad - wo * stride;
        const int64_t m = (b * Ho + ho) * (int64_t)Wo + wo;
        const float4 v = Vec4<T>::load(dcol + m * ldcol + (int64_t)(ky * kw + kx) * C + c);
        acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
      }
    }
    Vec4<T>::store(dx + pix * C + c, acc);
  }
}

// ---- BatchNorm2d, training mode ------------------------------------------------------------------------------------
// Column statistics of y[M, C] (M = B*H*W).  256 threads = (C/4 channel vectors) x (rows in flight); block j owns a
// contiguous row range, reduces it in registers, then across its row lanes in shared memory in a fixed order, and
// writes one partial row.  The last block to finish (ticket) folds the partial rows IN ORDER in fp64, so the result does
// not depend on scheduling.  `Fin` turns the two column sums into the kernel's outputs.
constexpr int BN_THREADS = 256;
constexpr int BN_MAX_C = 1024;

template <typename Acc>
__device__ __forceinline__ void bn_block_reduce(Acc a, Acc b, int vc, float* partial_row, int C) {
  // a, b: this thread's two float4 column sums for channel vector (tid % vc)
  __shared__ float4 sh[2][BN_THREADS];
  const int tid = threadIdx.x;
  sh[0][tid] = a;
  sh[1][tid] = b;
  __syncthreads();
  if (tid < vc) {
    float4 s0 = sh[0][tid], s1 = sh[1][tid];
    for (int r = tid + vc; r < BN_THREADS; r += vc) {
      const float4 u = sh[0][r], v = sh[1][r];
      s0.x += u.x; s0.y += u.y; s0.z += u.z; s0.w += u.w;
      s1.x += v.x; s1.y += v.y; s1.z += v.z; s1.w += v.w;
    }
    *reinterpret_cast<float4*>(partial_row + tid * 4) = s0;
    *reinterpret_cast<float4*>(partial_row + C + tid * 4) = s1;
  }
}

// true in exactly one block: the one that arrives last (all partial rows are then visible to it).  ONE thread per block
// takes the ticket with an acq_rel atomic at gpu scope: bar.sync orders the block's partial-row stores before it
// (release is cumulative) and its acquire before the other threads' reads, which go to L2 (ld.cg).  A __threadfence()
// in every thread instead compiles to MEMBAR.SC.GPU + CCTL.IVALL x 256 and cost ~20 us per launch.
__device__ __forceinline__ bool bn_last_block(uint32_t* ticket) {
  __shared__ uint32_t s_last;
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t t;
    asm volatile("atom.add.acq_rel.gpu.global.u32 %0, [%1], 1;" : "=r"(t) : "l"(ticket) : "memory");
    s_last = (t == gridDim.x - 1);
    if (s_last) *ticket = 0;                 // ready for the next launch (graph replay)
  }
  __syncthreads();
  return s_last != 0;
}

// The last block folds the `nb` partial rows [nb][2*C] into totals[2*C] (shared memory, fp64).  All 256 threads take
// part: thread (v, jg) sums rows jg, jg + lanes, ... of the 4 columns 4v..4v+3 with independent 16-byte loads in flight,
// then the `lanes` sub-sums of a column are added in a fixed order — the result is the same on every run.
// The running sums are fp32 with a Neumaier compensation term (sum + comp carries ~48 bits) instead of fp64 adds.
// Plain (weak) 16-byte global load of a partial row written by another block of this grid.  It is ordered after the
// ticket's acquire (bar.sync) and the L1 was invalidated by that acquire, so it cannot see stale data; `__ldcg` here
// compiled to LDG.STRONG.GPU, which ptxas would not batch: 64 serialised L2 round trips = 20 us on one SM.
__device__ __forceinline__ float4 ld_partial_v4(const float* p) {
  float4 r;
  asm volatile("ld.global.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p));
  return r;
}
__device__ __forceinline__ void neumaier_add(float& sum, float& comp, float x) {
  const float t = sum + x;
  comp += (fabsf(sum) >= fabsf(x)) ? ((sum - t) + x) : ((x - t) + sum);
  sum = t;
}
__device__ __forceinline__ void bn_fold_partials(const float* __restrict__ partials, int nb, int C, double* totals) {
  __shared__ float2 stage[BN_THREADS][4];
  const int vcols = 2 * C / 4;
  for (int v0 = 0; v0 < vcols; v0 += BN_THREADS) {
    const int width = min(vcols - v0, BN_THREADS);       // vector columns handled in this pass
    const int lanes = BN_THREADS / width;                // row lanes per column (>= 1)
    const int v = threadIdx.x % width, jg = threadIdx.x / width;
    float s0 = 0, s1 = 0, s2 = 0, s3 = 0, c0 = 0, c1 = 0, c2 = 0, c3 = 0;
    if (jg < lanes) {
      const float* col = partials + (v0 + v) * 4;
      for (int j = jg; j < nb; j += 8 * lanes) {          // 8 loads in flight, then the (ordered) additions
        float4 t[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) {
          const int jj = j + u * lanes;
          t[u] = jj < nb ? ld_partial_v4(col + (int64_t)jj * 2 * C) : make_float4(0.f, 0.f, 0.f, 0.f);
        }
#pragma unroll
        for (int u = 0; u < 8; ++u) {
          neumaier_add(s0, c0, t[u].x); neumaier_add(s1, c1, t[u].y);
          neumaier_add(s2, c2, t[u].z); neumaier_add(s3, c3, t[u].w);
        }
      }
    }
    stage[threadIdx.x][0] = make_float2(s0, c0); stage[threadIdx.x][1] = make_float2(s1, c1);
    stage[threadIdx.x][2] = make_float2(s2, c2); stage[threadIdx.x][3] = make_float2(s3, c3);
    __syncthreads();
    if (threadIdx.x < width) {
      double t[4] = {0, 0, 0, 0};
      for (int g = 0; g < lanes; ++g)
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const float2 p = stage[g * width + threadIdx.x][q];
          t[q] += (double)p.x + (double)p.y;
        }
      double* o = totals + (v0 + threadIdx.x) * 4;
      o[0] = t[0]; o[1] = t[1]; o[2] = t[2]; o[3] = t[3];
    }
    __syncthreads();
  }
}

template <typename T>
__global__ void __launch_bounds__(BN_THREADS)
bn_stats_kernel(const T* __restrict__ y, int64_t M, int C, float* __restrict__ partials, uint32_t* ticket, float eps,
                float momentum, float* __restrict__ mean, float* __restrict__ invstd, float* __restrict__ running_mean,
                float* __restrict__ running_var, int64_t* __restrict__ num_batches) {
  pdl_launch_dependents();
  pdl_wait();
  const int vc = C / 4, lanes = BN_THREADS / vc;
  const int cv = threadIdx.x % vc, ry = threadIdx.x / vc;
  const int64_t chunk = (M + gridDim.x - 1) / gridDim.x;
  const int64_t r0 = blockIdx.x * chunk, r1 = min(M, r0 + chunk);
  float4 s = make_float4(0, 0, 0, 0), q = make_float4(0, 0, 0, 0);
  if (ry < lanes)
#pragma unroll 8
    for (int64_t r = r0 + ry; r < r1; r += lanes) {
      const float4 v = Vec4<T>::load(y + r * C + cv * 4);
      s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
      q.x = fmaf(v.x, v.x, q.x); q.y = fmaf(v.y, v.y, q.y); q.z = fmaf(v.z, v.z, q.z); q.w = fmaf(v.w, v.w, q.w);
    }
  bn_block_reduce(s, q, vc, partials + (int64_t)blockIdx.x * 2 * C, C);
  if (!bn_last_block(ticket)) return;
  __shared__ double totals[2 * BN_MAX_C];
  bn_fold_partials(partials, (int)gridDim.x, C, totals);
  for (int c = threadIdx.x; c < C; c += BN_THREADS) {
    const double ds = totals[c], dq = totals[C + c];
    const double mu = ds / (double)M;
    double var = dq / (double)M - mu * mu;             // biased: what normalises the batch
    if (var < 0.0) var = 0.0;
    mean[c] = (float)mu;
    invstd[c] = (float)(1.0 / sqrt(var + (double)eps));
    if (running_mean != nullptr) {                       // batchnorm.py: running = (1 - momentum)*running + momentum*new,
      const double unbiased = M > 1 ? var * (double)M / (double)(M - 1) : var;   // with the UNBIASED variance
      running_mean[c] = (float)((1.0 - momentum) * (double)running_mean[c] + momentum * mu);
      running_var[c] = (float)((1.0 - momentum) * (double)running_var[c] + momentum * unbiased);
    }
  }
  if (threadIdx.x == 0 && num_batches != nullptr) num_batches[0] += 1;
}

// out = act(y*scale + shift (+ res)),  scale = gamma*invstd, shift = beta - mean*scale
template <typename T>
__global__ void __launch_bounds__(256)
bn_apply_kernel(const T* __restrict__ y, const T* __restrict__ res, T* __restrict__ out, int64_t M, int C,
                const float* __restrict__ mean, const float* __restrict__ invstd, const float* __restrict__ gamma,
                const float* __restrict__ beta, int relu) {
  pdl_launch_dependents();
  pdl_wait();
  const int vc = C / 4;
  const int64_t total = M * vc;
  for (int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
    const int c = (int)(t % vc) * 4;
    const int64_t off = (t / vc) * C + c;
    const float4 mu = *reinterpret_cast<const float4*>(mean + c), is = *reinterpret_cast<const float4*>(invstd + c);
    const float4 g = *reinterpret_cast<const float4*>(gamma + c), be = *reinterpret_cast<const float4*>(beta + c);
    const float4 v = Vec4<T>::load(y + off);
    float4 o;
    o.x = fmaf((v.x - mu.x) * is.x, g.x, be.x); o.y = fmaf((v.y - mu.y) * is.y, g.y, be.y);
    o.z = fmaf((v.z - mu.z) * is.z, g.z, be.z); o.w = fmaf((v.w - mu.w) * is.w, g.w, be.w);
    if (res != nullptr) {
      const float4 r = Vec4<T>::load(res + off);
      o.x += r.x; o.y += r.y; o.z += r.z; o.w += r.w;
    }
    if (relu) { o.x = fmaxf(o.x, 0.f); o.y = fmaxf(o.y, 0.f); o.z = fmaxf(o.z, 0.f); o.w = fmaxf(o.w, 0.f); }
    Vec4<T>::store(out + off, o);
  }
}

// backward, pass 1: dbeta[c] = sum_m dz, dgamma[c] = sum_m dz * xhat,  dz = dout * (out > 0) when `out` is given
template <typename T>
__global__ void __launch_bounds__(BN_THREADS)
bn_bwd_reduce_kernel(const T* __restrict__ dout, const T* __restrict__ dout2, const T* __restrict__ out,
                     const T* __restrict__ y, int64_t M, int C,
                     const float* __restrict__ mean, const float* __restrict__ invstd, float* __restrict__ partials,
                     uint32_t* ticket, float* __restrict__ dgamma, float* __restrict__ dbeta) {
  pdl_launch_dependents();
  pdl_wait();
  const int vc = C / 4, lanes = BN_THREADS / vc;
  const int cv = threadIdx.x % vc, ry = threadIdx.x / vc;
  const int64_t chunk = (M + gridDim.x - 1) / gridDim.x;
  const int64_t r0 = blockIdx.x * chunk, r1 = min(M, r0 + chunk);
  const float4 mu = *reinterpret_cast<const float4*>(mean + cv * 4), is = *reinterpret_cast<const float4*>(invstd + cv * 4);
  float4 sb = make_float4(0, 0, 0, 0), sg = make_float4(0, 0, 0, 0);
  if (ry < lanes)
#pragma unroll 4
    for (int64_t r = r0 + ry; r < r1; r += lanes) {
      const int64_t off = r * C + cv * 4;
      float4 d = Vec4<T>::load(dout + off);
      if (dout2 != nullptr) {                       // the incoming gradient is the sum of two branches
        const float4 e = Vec4<T>::load(dout2 + off);
        d.x += e.x; d.y += e.y; d.z += e.z; d.w += e.w;
      }
      if (out != nullptr) {
        const float4 o = Vec4<T>::load(out + off);
        d.x = o.x > 0.f ? d.x : 0.f; d.y = o.y > 0.f ? d.y : 0.f; d.z = o.z > 0.f ? d.z : 0.f; d.w = o.w > 0.f ? d.w : 0.f;
      }
      const float4 v = Vec4<T>::load(y + off);
      sb.x += d.x; sb.y += d.y; sb.z += d.z; sb.w += d.w;
      sg.x = fmaf(d.x, (v.x - mu.x) * is.x, sg.x); sg.y = fmaf(d.y, (v.y - mu.y) * is.y, sg.y);
      sg.z = fmaf(d.z, (v.z - mu.z) * is.z, sg.z); sg.w = fmaf(d.w, (v.w - mu.w) * is.w, sg.w);
    }
  bn_block_reduce(sb, sg, vc, partials + (int64_t)blockIdx.x * 2 * C, C);
  if (!bn_last_block(ticket)) return;
  __shared__ double totals[2 * BN_MAX_C];
  bn_fold_partials(partials, (int)gridDim.x, C, totals);
  for (int c = threadIdx.x; c < C; c += BN_THREADS) {
    dbeta[c] = (float)totals[c];
    dgamma[c] = (float)totals[C + c];
  }
}

// backward, pass 2: dy = gamma*invstd * (dz - dbeta/M - xhat*dgamma/M);  dz_out (nullable) = dz for the residual branch
template <typename T>
__global__ void __launch_bounds__(256)
bn_bwd_apply_kernel(const T* __restrict__ dout, const T* __restrict__ dout2, const T* __restrict__ out,
                    const T* __restrict__ y, T* __restrict__ dy, T* __restrict__ dz_out, int64_t M, int C, const float* __restrict__ mean,
                    const float* __restrict__ invstd, const float* __restrict__ gamma, const float* __restrict__ dgamma,
                    const float* __restrict__ dbeta) {
  pdl_launch_dependents();
  pdl_wait();
  const int vc = C / 4;
  const int64_t total = M * vc;
  const float inv_m = 1.f / (float)M;
  for (int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
    const int c = (int)(t % vc) * 4;
    const int64_t off = (t / vc) * C + c;
    const float4 mu = *reinterpret_cast<const float4*>(mean + c), is = *reinterpret_cast<const float4*>(invstd + c);
    const float4 g = *reinterpret_cast<const float4*>(gamma + c);
    const float4 dg = *reinterpret_cast<const float4*>(dgamma + c), db = *reinterpret_cast<const float4*>(dbeta + c);
    float4 d = Vec4<T>::load(dout + off);
    if (dout2 != nullptr) {
      const float4 e = Vec4<T>::load(dout2 + off);
      d.x += e.x; d.y += e.y; d.z += e.z; d.w += e.w;
    }
    if (out != nullptr) {
      const float4 o = Vec4<T>::load(out + off);
      d.x = o.x > 0.f ? d.x : 0.f; d.y = o.y > 0.f ? d.y : 0.f; d.z = o.z > 0.f ? d.z : 0.f; d.w = o.w > 0.f ? d.w : 0.f;
    }
    if (dz_out != nullptr) Vec4<T>::store(dz_out + off, d);
    const float4 v = Vec4<T>::load(y + off);
    float4 r;
    r.x = g.x * is.x * (d.x - db.x * inv_m - (v.x - mu.x) * is.x * dg.x * inv_m);
    r.y = g.y * is.y * (d.y - db.y * inv_m - (v.y - mu.y) * is.y * dg.y * inv_m);
    r.z = g.z * is.z * (d.z - db.z * inv_m - (v.z - mu.z) * is.z * dg.z * inv_m);
    r.w = g.w * is.w * (d.w - db.w * inv_m - (v.w - mu.w) * is.w * dg.w * inv_m);
    Vec4<T>::store(dy + off, r);
  }
}

// ---- fused variants: statistics AND apply in one launch ------------------------------------------------------------
// Same first phase as above.  Then, instead of ending the grid and launching the apply kernel, every block waits for
// the last block's result (an epoch word next to the ticket: read before arriving, bumped with a release store after the
// per-channel results are written) and applies it to ITS OWN rows, which it read a few microseconds earlier and
// which are still in L1 / L2.  One launch and one DRAM pass instead of two.  All blocks are co-resident (the grid is at
// most two blocks per SM and no other kernel waits on this one while it runs), so the wait cannot deadlock.
__device__ __forceinline__ uint32_t ld_acquire_u32(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_u32(uint32_t* p, uint32_t v) {
  asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void bn_wait_epoch(const uint32_t* epoch, uint32_t e0) {
  if (threadIdx.x == 0)
    while (ld_acquire_u32(epoch) == e0) __nanosleep(40);
  __syncthreads();
}
__device__ __forceinline__ float4 ldcg4(const float* p) { return __ldcg(reinterpret_cast<const float4*>(p)); }

template <typename T>
__global__ void __launch_bounds__(BN_THREADS)
bn_fwd_fused_kernel(const T* __restrict__ y, const T* __restrict__ res, T* __restrict__ out, int64_t M, int C,
                    float* __restrict__ partials, uint32_t* ticket, uint32_t* epoch, float eps, float momentum,
                    float* mean, float* invstd, const float* __restrict__ gamma, const float* __restrict__ beta, int relu,
                    float* __restrict__ running_mean, float* __restrict__ running_var, int64_t* __restrict__ num_batches) {
  pdl_launch_dependents();
  pdl_wait();
  const int vc = C / 4, lanes = BN_THREADS / vc;
  const int cv = threadIdx.x % vc, ry = threadIdx.x / vc;
  const int64_t chunk = (M + gridDim.x - 1) / gridDim.x;
  const int64_t r0 = blockIdx.x * chunk, r1 = min(M, r0 + chunk);
  __shared__ uint32_t s_e0;
  if (threadIdx.x == 0) s_e0 = ld_acquire_u32(epoch);          // before this block arrives at the ticket
  float4 s = make_float4(0, 0, 0, 0), q = make_float4(0, 0, 0, 0);
  if (ry < lanes)
#pragma unroll 8
    for (int64_t r = r0 + ry; r < r1; r += lanes) {
      const float4 v = Vec4<T>::load(y + r * C + cv * 4);
      s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
      q.x = fmaf(v.x, v.x, q.x); q.y = fmaf(v.y, v.y, q.y); q.z = fmaf(v.z, v.z, q.z); q.w = fmaf(v.w, v.w, q.w);
    }
  bn_block_reduce(s, q, vc, partials + (int64_t)blockIdx.x * 2 * C, C);
  const uint32_t e0 = s_e0;
  if (bn_last_block(ticket)) {
    __shared__ double totals[2 * BN_MAX_C];
    bn_fold_partials(partials, (int)gridDim.x, C, totals);
    for (int c = threadIdx.x; c < C; c += BN_THREADS) {
      const double ds = totals[c], dq = totals[C + c];
      const double mu = ds / (double)M;
      double var = dq / (double)M - mu * mu;
      if (var < 0.0) var = 0.0;
      mean[c] = (float)mu;
      invstd[c] = (float)(1.0 / sqrt(var + (double)eps));
      if (running_mean != nullptr) {
        const double unbiased = M > 1 ? var * (double)M / (double)(M - 1) : var;
        running_mean[c] = (float)((1.0 - momentum) * (double)running_mean[c] + momentum * mu);
        running_var[c] = (float)((1.0 - momentum) * (double)running_var[c] + momentum * unbiased);
      }
    }
    if (threadIdx.x == 0 && num_batches != nullptr) num_batches[0] += 1;
    __syncthreads();
    if (threadIdx.x == 0) st_release_u32(epoch, e0 + 1);
  }
  bn_wait_epoch(epoch, e0);
  if (ry >= lanes) return;
  const float4 mu = ldcg4(mean + cv * 4), is = ldcg4(invstd + cv * 4);
  const float4 g = *reinterpret_cast<const float4*>(gamma + cv * 4), be = *reinterpret_cast<const float4*>(beta + cv * 4);
#pragma unroll 4
  for (int64_t r = r0 + ry; r < r1; r += lanes) {
    const int64_t off = r * C + cv * 4;
    const float4 v = Vec4<T>::load(y + off);
    float4 o;
    o.x = fmaf((v.x - mu.x) * is.x, g.x, be.x); o.y = fmaf((v.y - mu.y) * is.y, g.y, be.y);
    o.z = fmaf((v.z - mu.z) * is.z, g.z, be.z); o.w = fmaf((v.w - mu.w) * is.w, g.w, be.w);
    if (res != nullptr) {
      const float4 rr = Vec4<T>::load(res + off);
      o.x += rr.x; o.y += rr.y; o.z += rr.z; o.w += rr.w;
    }
    if (relu) { o.x = fmaxf(o.x, 0.f); o.y = fmaxf(o.y, 0.f); o.z = fmaxf(o.z, 0.f); o.w = fmaxf(o.w, 0.f); }
    Vec4<T>::store(out + off, o);
  }
}

template <typename T>
__global__ void __launch_bounds__(BN_THREADS)
bn_bwd_fused_kernel(const T* __restrict__ dout, const T* __restrict__ dout2, const T* __restrict__ out,
                    const T* __restrict__ y, T* __restrict__ dy, T* __restrict__ dz_out, int64_t M, int C,
                    const float* __restrict__ mean, const float* __restrict__ invstd, const float* __restrict__ gamma,
                    float* __restrict__ partials, uint32_t* ticket, uint32_t* epoch, float* dgamma, float* dbeta) {
  pdl_launch_dependents();
  pdl_wait();
  const int vc = C / 4, lanes = BN_THREADS / vc;
  const int cv = threadIdx.x % vc, ry = threadIdx.x / vc;
  const int64_t chunk = (M + gridDim.x - 1) / gridDim.x;
  const int64_t r0 = blockIdx.x * chunk, r1 = min(M, r0 + chunk);
  __shared__ uint32_t s_e0;
  if (threadIdx.x == 0) s_e0 = ld_acquire_u32(epoch);
  const float4 mu = *reinterpret_cast<const float4*>(mean + cv * 4), is = *reinterpret_cast<const float4*>(invstd + cv * 4);
  float4 sb = make_float4(0, 0, 0, 0), sg = make_float4(0, 0, 0, 0);
  if (ry < lanes)
#pragma unroll 4
    for (int64_t r = r0 + ry; r < r1; r += lanes) {
      const int64_t off = r * C + cv * 4;
      float4 d = Vec4<T>::load(dout + off);
      if (dout2 != nullptr) {
        const float4 e = Vec4<T>::load(dout2 + off);
        d.x += e.x; d.y += e.y; d.z += e.z; d.w += e.w;
      }
      if (out != nullptr) {
        const float4 o = Vec4<T>::load(out + off);
        d.x = o.x > 0.f ? d.x : 0.f; d.y = o.y > 0.f ? d.y : 0.f; d.z = o.z > 0.f ? d.z : 0.f; d.w = o.w > 0.f ? d.w : 0.f;
      }
      const float4 v = Vec4<T>::load(y + off);
      sb.x += d.x; sb.y += d.y; sb.z += d.z; sb.w += d.w;
      sg.x = fmaf(d.x, (v.x - mu.x) * is.x, sg.x); sg.y = fmaf(d.y, (v.y - mu.y) * is.y, sg.y);
      sg.z = fmaf(d.z, (v.z - mu.z) * is.z, sg.z); sg.w = fmaf(d.w, (v.w - mu.w) * is.w, sg.w);
    }
  bn_block_reduce(sb, sg, vc, partials + (int64_t)blockIdx.x * 2 * C, C);
  const uint32_t e0 = s_e0;
  if (bn_last_block(ticket)) {
    __shared__ double totals[2 * BN_MAX_C];
    bn_fold_partials(partials, (int)gridDim.x, C, totals);
    for (int c = threadIdx.x; c < C; c += BN_THREADS) {
      dbeta[c] = (float)totals[c];
      dgamma[c] = (float)totals[C + c];
    }
    __syncthreads();
    if (threadIdx.x == 0) st_release_u32(epoch, e0 + 1);
  }
  bn_wait_epoch(epoch, e0);
  if (ry >= lanes) return;
  const float4 g = *reinterpret_cast<const float4*>(gamma + cv * 4);
  const float4 dg = ldcg4(dgamma + cv * 4), db = ldcg4(dbeta + cv * 4);
  const float inv_m = 1.f / (float)M;
#pragma unroll 2
  for (int64_t r = r0 + ry; r < r1; r += lanes) {
    const int64_t off = r * C + cv * 4;
    float4 d = Vec4<T>::load(dout + off);
    if (dout2 != nullptr) {
      const float4 e = Vec4<T>::load(dout2 + off);
      d.x += e.x; d.y += e.y; d.z += e.z; d.w += e.w;
    }
    if (out != nullptr) {
      const float4 o = Vec4<T>::load(out + off);
      d.x = o.x > 0.f ? d.x : 0.f; d.y = o.y > 0.f ? d.y : 0.f; d.z = o.z > 0.f ? d.z : 0.f; d.w = o.w > 0.f ? d.w : 0.f;
    }
    if (dz_out != nullptr) Vec4<T>::store(dz_out + off, d);
    const float4 v = Vec4<T>::load(y + off);
    float4 rr;
    rr.x = g.x * is.x * (d.x - db.x * inv_m - (v.x - mu.x) * is.x * dg.x * inv_m);
    rr.y = g.y * is.y * (d.y - db.y * inv_m - (v.y - mu.y) * is.y * dg.y * inv_m);
    rr.z = g.z * is.z * (d.z - db.z * inv_m - (v.z - mu.z) * is.z * dg.z * inv_m);
    rr.w = g.w * is.w * (d.w - db.w * inv_m - (v.w - mu.w) * is.w * dg.w * inv_m);
    Vec4<T>::store(dy + off, rr);
  }
}

// ---- MaxPool2d(kernel 3, stride 2, padding 1) (torchvision resnet.py: self.maxpool) ------------------------------------
// fwd keeps the winning tap (first maximum in (ky, kx) order, NaN wins: aten's `val > max || isnan(val)`) as one byte
// per output element; bwd is the gather over the <= 4 windows that contain an input pixel.
template <typename T>
__global__ void __launch_bounds__(256)
maxpool_fwd_kernel(const T* __restrict__ x, T* __restrict__ out, uint8_t* __restrict__ tap_out, int B, int H, int W,
                   int C, int Ho, int Wo, int k, int stride, int pad) {
  pdl_launch_dependents();
  pdl_wait();
  const int vc = C / 4;
  const int64_t total = (int64_t)B * Ho * Wo * vc;
  for (int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
    const int c = (int)(t % vc) * 4;
    const int64_t m = t / vc;
    const int wo = (int)(m % Wo), ho = (int)((m / Wo) % Ho);
    const int64_t b = m / ((int64_t)Wo * Ho);
    float best[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
    int tap[4] = {255, 255, 255, 255};
    for (int ky = 0; ky < k; ++ky) {
      const int h = ho * stride - pad + ky;
      if (h < 0 || h >= H) continue;
      for (int kx = 0; kx < k; ++kx) {
        const int w = wo * stride - pad + kx;
        if (w < 0 || w >= W) continue;
        const float4 v4 = Vec4<T>::load(x + ((b * H + h) * (int64_t)W + w) * C + c);
        const float v[4] = {v4.x, v4.y, v4.z, v4.w};
#pragma unroll
        for (int j = 0; j < 4; ++j)
          if (tap[j] == 255 || v[j] > best[j] || v[j] != v[j]) { best[j] = v[j]; tap[j] = ky * k + kx; }
      }
    }
    Vec4<T>::store(out + m * C + c, make_float4(best[0], best[1], best[2], best[3]));
    *reinterpret_cast<uchar4*>(tap_out + m * C + c) = make_uchar4(tap[0], tap[1], tap[2], tap[3]);
  }
}

template <typename T>
__global__ void __launch_bounds__(256)
maxpool_bwd_kernel(const T* __restrict__ dout, const T* __restrict__ dout2, const uint8_t* __restrict__ tap_in,
                   T* __restrict__ dx, int B, int H, int W,
                   int C, int Ho, int Wo, int k, int stride, int pad) {
  pdl_launch_dependents();
  pdl_wait();
  const int vc = C / 4;
  const int64_t total = (int64_t)B * H * W * vc;
  for (int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
    const int c = (int)(t % vc) * 4;
    const int64_t pix = t / vc;
    const int w = (int)(pix % W), h = (int)((pix / W) % H);
    const int64_t b = pix / ((int64_t)W * H);
    float4 acc = make_float4(0, 0, 0, 0);
    // only the windows that contain this pixel: ho in [ceil((h + pad - k + 1) / stride), floor((h + pad) / stride)]
    // (at most two per dimension for the ResNet stem's 3x3 / stride 2), clipped to the output
    const bool s2 = stride == 2;
    const int hn = h + pad - k + 1, wn = w + pad - k + 1;
    const int hlo = hn <= 0 ? 0 : (s2 ? (hn + 1) >> 1 : (hn + stride - 1) / stride);
    const int wlo = wn <= 0 ? 0 : (s2 ? (wn + 1) >> 1 : (wn + stride - 1) / stride);
    const int hhi = min(Ho - 1, s2 ? (h + pad) >> 1 : (h + pad) / stride);
    const int whi = min(Wo - 1, s2 ? (w + pad) >> 1 : (w + pad) / stride);
    for (int hq = hlo; hq <= hhi; ++hq) {
      const int ky = h + pad - hq * stride;
      for (int wq = wlo; wq <= whi; ++wq) {
        const int kx = w + pad - wq * stride;
        const int64_t m = (b * Ho + hq) * (int64_t)Wo + wq;
        const uchar4 tp = *reinterpret_cast<const uchar4*>(tap_in + m * C + c);
        float4 d = Vec4<T>::load(dout + m * C + c);
        if (dout2 != nullptr) {
          const float4 e = Vec4<T>::load(dout2 + m * C + c);
          d.x += e.x; d.y += e.y; d.z += e.z; d.w += e.w;
        }
        const int me = ky * k + kx;
        if (tp.x == me) acc.x += d.x;
        if (tp.y == me) acc.y += d.y;
        if (tp.z == me) acc.z += d.z;
        if (tp.w == me) acc.w += d.w;
      }
    }
    Vec4<T>::store(dx + pix * C + c, acc);
  }
}

// ---- AdaptiveAvgPool2d((1,1)): out[b, c] = mean over the HW pixels; bwd spreads dout / HW --------------------------------
template <typename T>
__global__ void __launch_bounds__(256)
avgpool_fwd_kernel(const T* __restrict__ x, T* __restrict__ out, int B, int HW, int C) {
  pdl_launch_dependents();
  pdl_wait();
  const int vc = C / 4;
  const int64_t total = (int64_t)B * vc;
  for (int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
    const int c = (int)(t % vc) * 4;
    const int64_t b = t / vc;
    float4 s = make_float4(0, 0, 0, 0);
    for (int p = 0; p < HW; ++p) {
      const float4 v = Vec4<T>::load(x + (b * HW + p) * C + c);
      s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
    }
    const float inv = 1.f / (float)HW;
    Vec4<T>::store(out + b * C + c, make_float4(s.x * inv, s.y * inv, s.z * inv, s.w * inv));
  }
}
template <typename T>
__global__ void __launch_bounds__(256)
avgpool_bwd_kernel(const T* __restrict__ dout, T* __restrict__ dx, int B, int HW, int C) {
  pdl_launch_dependents();
  pdl_wait();
  const int vc = C / 4;
  const int64_t total = (int64_t)B * HW * vc;
  const float inv = 1.f / (float)HW;
  for (int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
    const int c = (int)(t % vc) * 4;
    const int64_t pix = t / vc;
    const float4 d = Vec4<T>::load(dout + (pix / HW) * C + c);
    Vec4<T>::store(dx + pix * C + c, make_float4(d.x * inv, d.y * inv, d.z * inv, d.w * inv));
  }
}

}  // namespace b200

// ---- C ABI ----------------------------------------------------------------------------------------------------------
using namespace b200;
typedef __nv_bfloat16 bf16_t;

#define CONV_COMMON(name, ptr_ok)                                                                              \
  B200_REQUIRE(dtype == B200_F32 || dtype == B200_BF16, name ": unknown dtype code %d", dtype);                \
  B200_REQUIRE(ptr_ok, name ": NULL or misaligned (16 bytes) tensor pointer");                                 \
  cudaStream_t st = static_cast<cudaStream_t>(stream)
#define ALIGNED16(p) ((p) != nullptr && (reinterpret_cast<uintptr_t>(p) & 15) == 0)
#define ALIGNED16_OR_NULL(p) ((reinterpret_cast<uintptr_t>(p) & 15) == 0)
static inline int conv_out(int in, int k, int stride, int pad) { return (in + 2 * pad - k) / stride + 1; }

extern "C" int b200_image_augment(const uint8_t* images, const int64_t* labels, const int64_t* row_idx,
                                  const int32_t* cursor, int64_t cursor_stride, const int8_t* draws, int64_t n_rows,
                                  const float* mean3, const float* std3, int pad, void* x_out, int64_t* y_out,
                                  int64_t batch, int H, int W, int dtype, b200_stream_t stream) {
  CONV_COMMON("image_augment", images != nullptr && x_out != nullptr && mean3 != nullptr && std3 != nullptr);
  B200_REQUIRE(batch > 0 && H > 0 && W > 0 && n_rows > 0 && batch * H * W < (1ll << 31) && pad >= 0 && pad < 64,
               "image_augment: bad sizes (batch=%lld H=%d W=%d rows=%lld pad=%d)", (long long)batch, H, W,
               (long long)n_rows, pad);
  B200_REQUIRE(y_out == nullptr || labels != nullptr, "image_augment: y_out without labels");
  const int grid = grid_for(batch * H * W, 256);
  const void* xo = x_out;
  if (dtype == B200_BF16)
    B200_CHECK_CUDA(launch_pdl(augment_kernel<bf16_t>, dim3(grid), dim3(256), 0, st, images, labels, row_idx, cursor,
                               cursor_stride, draws, n_rows, mean3[0], mean3[1], mean3[2], std3[0], std3[1], std3[2],
                               pad, (bf16_t*)xo, y_out, (int)batch, H, W, status_word()));
  else
    B200_CHECK_CUDA(launch_pdl(augment_kernel<float>, dim3(grid), dim3(256), 0, st, images, labels, row_idx, cursor,
                               cursor_stride, draws, n_rows, mean3[0], mean3[1], mean3[2], std3[0], std3[1], std3[2],
                               pad, (float*)xo, y_out, (int)batch, H, W, status_word()));
  B200_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int b200_im2col(const void* x, void* col, int64_t ldcol, int64_t B, int H, int W, int C, int kh, int kw,
                           int stride, int pad, int dtype, b200_stream_t stream) {
  CONV_COMMON("im2col", ALIGNED16(col) && x != nullptr);
  B200_REQUIRE(B > 0 && H > 0 && W > 0 && C > 0 && kh > 0 && kw > 0 && stride > 0 && pad >= 0 && kh <= H + 2 * pad &&
                   kw <= W + 2 * pad, "im2col: bad geometry");
  const int Ho = conv_out(H, kh, stride, pad), Wo = conv_out(W, kw, stride, pad);
  const int K = kh * kw * C;
  const int es = dtype == B200_F32 ? 4 : 2;
  B200_REQUIRE(ldcol >= K && (ldcol * es) % 16 == 0, "im2col: ldcol=%lld must be >= %d and a multiple of 16 bytes",
               (long long)ldcol, K);
  B200_REQUIRE(B * Ho * Wo * ldcol < (1ll << 40), "im2col: col matrix too large");
  const int vecn = 16 / es;                               // elements per 16-byte vector
  const bool aligned = (reinterpret_cast<uintptr_t>(x) & 15) == 0;
  int cshift = -1;
  if (aligned && C % vecn == 0 && ldcol % vecn == 0 && kh * kw < 64)
    for (int sft = 0; sft < 16; ++sft)
      if ((C / vecn) == (1 << sft)) cshift = sft;
  const int64_t rows = B * Ho * Wo;
  if (cshift >= 0) {                                       // warp per row, 16-byte vectors
    const int grid = grid_for(rows * 32, 256, 16);
    if (dtype == B200_BF16) B200_CHECK_CUDA(launch_pdl(im2col_vec_kernel<bf16_t>, dim3(grid), dim3(256), 0, st, (const bf16_t*)x, (bf16_t*)col, (int)B, H, W, C, Ho, Wo, kh, kw, stride, pad, ldcol, cshift));
    else B200_CHECK_CUDA(launch_pdl(im2col_vec_kernel<float>, dim3(grid), dim3(256), 0, st, (const float*)x, (float*)col, (int)B, H, W, C, Ho, Wo, kh, kw, stride, pad, ldcol, cshift));
  } else if (C <= 8 && kh * kw < 64 && ldcol % vecn == 0 && ldcol * es * IM_ROWS <= 48 * 1024) {   // conv1: rows staged in shared memory
    const int grid = grid_for((rows + IM_ROWS - 1) / IM_ROWS * 256, 256, 8);
    const size_t smem = (size_t)ldcol * es * IM_ROWS;
    if (dtype == B200_BF16) B200_CHECK_CUDA(launch_pdl(im2col_smallc_kernel<bf16_t>, dim3(grid), dim3(256), smem, st, (const bf16_t*)x, (bf16_t*)col, (int)B, H, W, C, Ho, Wo, kh, kw, stride, pad, ldcol));
    else B200_CHECK_CUDA(launch_pdl(im2col_smallc_kernel<float>, dim3(grid), dim3(256), smem, st, (const float*)x, (float*)col, (int)B, H, W, C, Ho, Wo, kh, kw, stride, pad, ldcol));
  } else {                                                 // any other geometry: one thread per 4 channels / per element
    const bool vec = C % 4 == 0 && aligned;
    const int64_t items = rows * (vec ? ldcol / 4 : ldcol);
    const int grid = grid_for(items, 256, 16);
    const int kcols = (int)ldcol;
    if (dtype == B200_BF16) {
      if (vec) B200_CHECK_CUDA(launch_pdl(im2col_kernel<bf16_t, 4>, dim3(grid), dim3(256), 0, st, (const bf16_t*)x, (bf16_t*)col, (int)B, H, W, C, Ho, Wo, kh, kw, stride, pad, ldcol, kcols));
      else B200_CHECK_CUDA(launch_pdl(im2col_kernel<bf16_t, 1>, dim3(grid), dim3(256), 0, st, (const bf16_t*)x, (bf16_t*)col, (int)B, H, W, C, Ho, Wo, kh, kw, stride, pad, ldcol, kcols));
    } else {
      if (vec) B200_CHECK_CUDA(launch_pdl(im2col_kernel<float, 4>, dim3(grid), dim3(256), 0, st, (const float*)x, (float*)col, (int)B, H, W, C, Ho, Wo, kh, kw, stride, pad, ldcol, kcols));
      else B200_CHECK_CUDA(launch_pdl(im2col_kernel<float, 1>, dim3(grid), dim3(256), 0, st, (const float*)x, (float*)col, (int)B, H, W, C, Ho, Wo, kh, kw, stride, pad, ldcol, kcols));
    }
  }
  B200_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int b200_col2im(const void* dcol, int64_t ldcol, const void* add, void* dx, int64_t B, int H, int W, int C,
                           int kh, int kw, int stride, int pad, int dtype, b200_stream_t stream) {
  CONV_COMMON("col2im", ALIGNED16(dcol) && ALIGNED16(dx) && ALIGNED16_OR_NULL(add));
  B200_REQUIRE(B > 0 && H > 0 && W > 0 && C > 0 && C % 4 == 0 && kh > 0 && kw > 0 && stride > 0 && pad >= 0,
               "col2im: bad geometry (C must be a multiple of 4)");
  const int es = dtype == B200_F32 ? 4 : 2;
  B200_REQUIRE(ldcol >= (int64_t)kh * kw * C && (ldcol * es) % 16 == 0, "col2im: bad ldcol=%lld", (long long)ldcol);
  const int Ho = conv_out(H, kh, stride, pad), Wo = conv_out(W, kw, stride, pad);
  const int grid = grid_for(B * H * W * (C / 4), 256, 16);
  if (dtype == B200_BF16)
    B200_CHECK_CUDA(launch_pdl(col2im_kernel<bf16_t>, dim3(grid), dim3(256), 0, st, (const bf16_t*)dcol, ldcol, (const bf16_t*)add, (bf16_t*)dx, (int)B, H, W, C, Ho, Wo, kh, kw, stride, pad));
  else
    B200_CHECK_CUDA(launch_pdl(col2im_kernel<float>, dim3(grid), dim3(256), 0, st, (const float*)dcol, ldcol, (const float*)add, (float*)dx, (int)B, H, W, C, Ho, Wo, kh, kw, stride, pad));
  B200_CHECK_CUDA(cudaGetLastError());
  return 0;
}

// ---- BatchNorm entry points ---------------------------------------------------------------------------------------
constexpr int BN_MAX_BLOCKS = 296;                   // two row blocks per SM
constexpr int BN_SCRATCH_HEAD = 4;                   // floats in front of the partial rows: the ticket word + padding

extern "C" int64_t b200_bn_scratch_floats(int64_t C) { return BN_SCRATCH_HEAD + (int64_t)BN_MAX_BLOCKS * 2 * C; }

static int bn_check(const char* who, int64_t M, int64_t C, const float* scratch, int64_t scratch_floats) {
  B200_REQUIRE(M > 0 && M < (1ll << 40) && C >= 4 && C <= 1024 && C % 4 == 0 && BN_THREADS % (C / 4) == 0,
               "%s: C=%lld must be 4 * (a divisor of 256), M=%lld positive", who, (long long)C, (long long)M);
  B200_REQUIRE(scratch == nullptr || (ALIGNED16(scratch) && scratch_floats >= b200_bn_scratch_floats(C)),
               "%s: scratch must be 16-byte aligned and hold b200_bn_scratch_floats(C) = %lld floats", who,
               (long long)b200_bn_scratch_floats(C));
  return 0;
}
static inline int bn_blocks(int64_t M, int64_t C) {
  const int64_t lanes = BN_THREADS / (C / 4);
  const int64_t want = (M + lanes * 16 - 1) / (lanes * 16);          // >= 16 rows per thread before splitting further
  const int64_t cap = std::min<int64_t>(BN_MAX_BLOCKS, (int64_t)sm_count() * 2);
  return (int)std::max<int64_t>(1, std::min(want, cap));
}

template <typename T>
static int bn_stats_t(const void* y, int64_t M, int C, float* scratch, float eps, float momentum, float* mean,
                      float* invstd, float* rm, float* rv, int64_t* nb, cudaStream_t st) {
  B200_CHECK_CUDA(launch_pdl(bn_stats_kernel<T>, dim3(bn_blocks(M, C)), dim3(BN_THREADS), 0, st, (const T*)y, M, C,
                             scratch + BN_SCRATCH_HEAD, reinterpret_cast<uint32_t*>(scratch), eps, momentum, mean,
                             invstd, rm, rv, nb));
  B200_CHECK_CUDA(cudaGetLastError());
  return 0;
}
extern "C" int b200_bn_fwd_stats(const void* y, int64_t M, int64_t C, float eps, float momentum, float* mean,
                                 float* invstd, float* running_mean, float* running_var, int64_t* num_batches,
                                 float* scratch, int64_t scratch_floats, int dtype, b200_stream_t stream) {
  CONV_COMMON("bn_fwd_stats", ALIGNED16(y) && mean != nullptr && invstd != nullptr && scratch != nullptr);
  if (int r = bn_check("bn_fwd_stats", M, C, scratch, scratch_floats)) return r;
  B200_REQUIRE((running_mean == nullptr) == (running_var == nullptr), "bn_fwd_stats: running_mean / running_var go together");
  return dtype == B200_BF16 ? bn_stats_t<bf16_t>(y, M, (int)C, scratch, eps, momentum, mean, invstd, running_mean, running_var, num_batches, st)
                            : bn_stats_t<float>(y, M, (int)C, scratch, eps, momentum, mean, invstd, running_mean, running_var, num_batches, st);
}

template <typename T>
static int bn_apply_t(const void* y, const void* res, void* out, int64_t M, int C, const float* mean, const float* invstd,
                      const float* gamma, const float* beta, int relu, cudaStream_t st) {
  B200_CHECK_CUDA(launch_pdl(bn_apply_kernel<T>, dim3(grid_for(M * (C / 4), 256, 16)), dim3(256), 0, st, (const T*)y,
                             (const T*)res, (T*)out, M, C, mean, invstd, gamma, beta, relu));
  B200_CHECK_CUDA(cudaGetLastError());
  return 0;
}
extern "C" int b200_bn_fwd_apply(const void* y, const void* res, void* out, int64_t M, int64_t C, const float* mean,
                                 const float* invstd, const float* gamma, const float* beta, int relu, int dtype,
                                 b200_stream_t stream) {
  CONV_COMMON("bn_fwd_apply", ALIGNED16(y) && ALIGNED16(out) && ALIGNED16_OR_NULL(res) && ALIGNED16(mean) &&
                                  ALIGNED16(invstd) && ALIGNED16(gamma) && ALIGNED16(beta));
  if (int r = bn_check("bn_fwd_apply", M, C, nullptr, 0)) return r;
  return dtype == B200_BF16 ? bn_apply_t<bf16_t>(y, res, out, M, (int)C, mean, invstd, gamma, beta, relu, st)
                            : bn_apply_t<float>(y, res, out, M, (int)C, mean, invstd, gamma, beta, relu, st);
}

template <typename T>
static int bn_bwd_reduce_t(const void* dout, const void* dout2, const void* out, const void* y, int64_t M, int C, const float* mean,
                           const float* invstd, float* scratch, float* dgamma, float* dbeta, cudaStream_t st) {
  B200_CHECK_CUDA(launch_pdl(bn_bwd_reduce_kernel<T>, dim3(bn_blocks(M, C)), dim3(BN_THREADS), 0, st, (const T*)dout,
                             (const T*)dout2, (const T*)out, (const T*)y, M, C, mean, invstd, scratch + BN_SCRATCH_HEAD,
                             reinterpret_cast<uint32_t*>(scratch), dgamma, dbeta));
  B200_CHECK_CUDA(cudaGetLastError());
  return 0;
}
extern "C" int b200_bn_bwd_reduce(const void* dout, const void* dout2, const void* out, const void* y, int64_t M, int64_t C,
                                  const float* mean, const float* invstd, float* dgamma, float* dbeta, float* scratch,
                                  int64_t scratch_floats, int dtype, b200_stream_t stream) {
  CONV_COMMON("bn_bwd_reduce", ALIGNED16(dout) && ALIGNED16_OR_NULL(dout2) && ALIGNED16(y) && ALIGNED16_OR_NULL(out) && ALIGNED16(mean) &&
                                   ALIGNED16(invstd) && dgamma != nullptr && dbeta != nullptr && scratch != nullptr);
  if (int r = bn_check("bn_bwd_reduce", M, C, scratch, scratch_floats)) return r;
  return dtype == B200_BF16 ? bn_bwd_reduce_t<bf16_t>(dout, dout2, out, y, M, (int)C, mean, invstd, scratch, dgamma, dbeta, st)
                            : bn_bwd_reduce_t<float>(dout, dout2, out, y, M, (int)C, mean, invstd, scratch, dgamma, dbeta, st);
}

template <typename T>
static int bn_bwd_apply_t(const void* dout, const void* dout2, const void* out, const void* y, void* dy, void* dz, int64_t M, int C,
                          const float* mean, const float* invstd, const float* gamma, const float* dgamma,
                          const float* dbeta, cudaStream_t st) {
  B200_CHECK_CUDA(launch_pdl(bn_bwd_apply_kernel<T>, dim3(grid_for(M * (C / 4), 256, 16)), dim3(256), 0, st,
                             (const T*)dout, (const T*)dout2, (const T*)out, (const T*)y, (T*)dy, (T*)dz, M, C, mean, invstd, gamma,
                             dgamma, dbeta));
  B200_CHECK_CUDA(cudaGetLastError());
  return 0;
}
extern "C" int b200_bn_bwd_apply(const void* dout, const void* dout2, const void* out, const void* y, void* dy, void* dz_out, int64_t M,
                                 int64_t C, const float* mean, const float* invstd, const float* gamma,
                                 const float* dgamma, const float* dbeta, int dtype, b200_stream_t stream) {
  CONV_COMMON("bn_bwd_apply", ALIGNED16(dout) && ALIGNED16_OR_NULL(dout2) && ALIGNED16(y) && ALIGNED16(dy) && ALIGNED16_OR_NULL(out) &&
                                  ALIGNED16_OR_NULL(dz_out) && ALIGNED16(mean) && ALIGNED16(invstd) && ALIGNED16(gamma) &&
                                  ALIGNED16(dgamma) && ALIGNED16(dbeta));
  if (int r = bn_check("bn_bwd_apply", M, C, nullptr, 0)) return r;
  return dtype == B200_BF16
             ? bn_bwd_apply_t<bf16_t>(dout, dout2, out, y, dy, dz_out, M, (int)C, mean, invstd, gamma, dgamma, dbeta, st)
             : bn_bwd_apply_t<float>(dout, dout2, out, y, dy, dz_out, M, (int)C, mean, invstd, gamma, dgamma, dbeta, st);
}

// ---- fused BatchNorm entry points ----------------------------------------------------------------------------------
// the grid must be co-resident (blocks wait for the last one): clamp it to what the device can hold at once
template <typename K>
static int bn_fused_blocks(K kern, int64_t M, int64_t C) {
  int per_sm = 0;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, BN_THREADS, 0) != cudaSuccess || per_sm < 1) return 0;
  const int64_t cap = (int64_t)std::min(per_sm, 2) * sm_count();
  return (int)std::min<int64_t>(bn_blocks(M, C), cap);
}
template <typename T>
static int bn_fwd_fused_t(const void* y, const void* res, void* out, int64_t M, int C, float* scratch, float eps,
                          float momentum, float* mean, float* invstd, const float* gamma, const float* beta, int relu,
                          float* rm, float* rv, int64_t* nb, cudaStream_t st) {
  const int blocks = bn_fused_blocks(bn_fwd_fused_kernel<T>, M, C);
  B200_REQUIRE(blocks >= 1, "bn_fwd: the fused kernel cannot be resident on this device");
  uint32_t* words = reinterpret_cast<uint32_t*>(scratch);
  B200_CHECK_CUDA(launch_pdl(bn_fwd_fused_kernel<T>, dim3(blocks), dim3(BN_THREADS), 0, st, (const T*)y, (const T*)res,
                             (T*)out, M, C, scratch + BN_SCRATCH_HEAD, words, words + 1, eps, momentum, mean, invstd,
                             gamma, beta, relu, rm, rv, nb));
  B200_CHECK_CUDA(cudaGetLastError());
  return 0;
}
extern "C" int b200_bn_fwd(const void* y, const void* res, void* out, int64_t M, int64_t C, float eps, float momentum,
                           float* mean, float* invstd, const float* gamma, const float* beta, int relu,
                           float* running_mean, float* running_var, int64_t* num_batches, float* scratch,
                           int64_t scratch_floats, int dtype, b200_stream_t stream) {
  CONV_COMMON("bn_fwd", ALIGNED16(y) && ALIGNED16(out) && ALIGNED16_OR_NULL(res) && ALIGNED16(mean) && ALIGNED16(invstd) &&
                            ALIGNED16(gamma) && ALIGNED16(beta) && scratch != nullptr);
  if (int r = bn_check("bn_fwd", M, C, scratch, scratch_floats)) return r;
  B200_REQUIRE((running_mean == nullptr) == (running_var == nullptr), "bn_fwd: running_mean / running_var go together");
  return dtype == B200_BF16
             ? bn_fwd_fused_t<bf16_t>(y, res, out, M, (int)C, scratch, eps, momentum, mean, invstd, gamma, beta, relu,
                                      running_mean, running_var, num_batches, st)
             : bn_fwd_fused_t<float>(y, res, out, M, (int)C, scratch, eps, momentum, mean, invstd, gamma, beta, relu,
                                     running_mean, running_var, num_batches, st);
}
template <typename T>
static int bn_bwd_fused_t(const void* dout, const void* dout2, const void* out, const void* y, void* dy, void* dz, int64_t M,
                          int C, const float* mean, const float* invstd, const float* gamma, float* scratch,
                          float* dgamma, float* dbeta, cudaStream_t st) {
  const int blocks = bn_fused_blocks(bn_bwd_fused_kernel<T>, M, C);
  B200_REQUIRE(blocks >= 1, "bn_bwd: the fused kernel cannot be resident on this device");
  uint32_t* words = reinterpret_cast<uint32_t*>(scratch);
  B200_CHECK_CUDA(launch_pdl(bn_bwd_fused_kernel<T>, dim3(blocks), dim3(BN_THREADS), 0, st, (const T*)dout,
                             (const T*)dout2, (const T*)out, (const T*)y, (T*)dy, (T*)dz, M, C, mean, invstd, gamma,
                             scratch + BN_SCRATCH_HEAD, words, words + 1, dgamma, dbeta));
  B200_CHECK_CUDA(cudaGetLastError());
  return 0;
}
extern "C" int b200_bn_bwd(const void* dout, const void* dout2, const void* out, const void* y, void* dy, void* dz_out,
                           int64_t M, int64_t C, const float* mean, const float* invstd, const float* gamma,
                           float* dgamma, float* dbeta, float* scratch, int64_t scratch_floats, int dtype,
                           b200_stream_t stream) {
  CONV_COMMON("bn_bwd", ALIGNED16(dout) && ALIGNED16_OR_NULL(dout2) && ALIGNED16(y) && ALIGNED16(dy) &&
                            ALIGNED16_OR_NULL(out) && ALIGNED16_OR_NULL(dz_out) && ALIGNED16(mean) && ALIGNED16(invstd) &&
                            ALIGNED16(gamma) && ALIGNED16(dgamma) && ALIGNED16(dbeta) && scratch != nullptr);
  if (int r = bn_check("bn_bwd", M, C, scratch, scratch_floats)) return r;
  return dtype == B200_BF16 ? bn_bwd_fused_t<bf16_t>(dout, dout2, out, y, dy, dz_out, M, (int)C, mean, invstd, gamma, scratch,
                                                     dgamma, dbeta, st)
                            : bn_bwd_fused_t<float>(dout, dout2, out, y, dy, dz_out, M, (int)C, mean, invstd, gamma, scratch,
                                                    dgamma, dbeta, st);
}

// ---- pooling entry points -------------------------------------------------------------------------------------------
template <typename T>
static int maxpool_t(bool fwd, const void* in, const void* in2, void* out, uint8_t* taps, int B, int H, int W, int C, int Ho, int Wo, int k,
                     int stride, int pad, cudaStream_t st) {
  if (fwd)
    B200_CHECK_CUDA(launch_pdl(maxpool_fwd_kernel<T>, dim3(grid_for((int64_t)B * Ho * Wo * (C / 4), 256, 16)), dim3(256),
                               0, st, (const T*)in, (T*)out, taps, B, H, W, C, Ho, Wo, k, stride, pad));
  else
    B200_CHECK_CUDA(launch_pdl(maxpool_bwd_kernel<T>, dim3(grid_for((int64_t)B * H * W * (C / 4), 256, 16)), dim3(256), 0,
                               st, (const T*)in, (const T*)in2, (const uint8_t*)taps, (T*)out, B, H, W, C, Ho, Wo, k, stride, pad));
  B200_CHECK_CUDA(cudaGetLastError());
  return 0;
}
static int maxpool_check(int64_t B, int H, int W, int C, int k, int stride, int pad) {
  B200_REQUIRE(B > 0 && H > 0 && W > 0 && C > 0 && C % 4 == 0 && k > 0 && k <= 15 && stride > 0 && pad >= 0 &&
                   2 * pad <= k && B * H * W < (1ll << 31),
               "maxpool: bad geometry (C %% 4 == 0, kernel <= 15, padding <= kernel / 2)");
  return 0;
}
extern "C" int b200_maxpool_fwd(const void* x, void* out, uint8_t* taps, int64_t B, int H, int W, int C, int k,
                                int stride, int pad, int dtype, b200_stream_t stream) {
  CONV_COMMON("maxpool_fwd", ALIGNED16(x) && ALIGNED16(out) && ALIGNED16(taps));
  if (int r = maxpool_check(B, H, W, C, k, stride, pad)) return r;
  const int Ho = conv_out(H, k, stride, pad), Wo = conv_out(W, k, stride, pad);
  return dtype == B200_BF16 ? maxpool_t<bf16_t>(true, x, nullptr, out, taps, (int)B, H, W, C, Ho, Wo, k, stride, pad, st)
                            : maxpool_t<float>(true, x, nullptr, out, taps, (int)B, H, W, C, Ho, Wo, k, stride, pad, st);
}
extern "C" int b200_maxpool_bwd(const void* dout, const void* dout2, const uint8_t* taps, void* dx, int64_t B, int H, int W, int C, int k,
                                int stride, int pad, int dtype, b200_stream_t stream) {
  CONV_COMMON("maxpool_bwd", ALIGNED16(dout) && ALIGNED16_OR_NULL(dout2) && ALIGNED16(dx) && ALIGNED16(taps));
  if (int r = maxpool_check(B, H, W, C, k, stride, pad)) return r;
  const int Ho = conv_out(H, k, stride, pad), Wo = conv_out(W, k, stride, pad);
  return dtype == B200_BF16
             ? maxpool_t<bf16_t>(false, dout, dout2, dx, const_cast<uint8_t*>(taps), (int)B, H, W, C, Ho, Wo, k, stride, pad, st)
             : maxpool_t<float>(false, dout, dout2, dx, const_cast<uint8_t*>(taps), (int)B, H, W, C, Ho, Wo, k, stride, pad, st);
}

template <typename T>
static int avgpool_t(bool fwd, const void* in, void* out, int B, int HW, int C, cudaStream_t st) {
  if (fwd)
    B200_CHECK_CUDA(launch_pdl(avgpool_fwd_kernel<T>, dim3(grid_for((int64_t)B * (C / 4), 256)), dim3(256), 0, st,
                               (const T*)in, (T*)out, B, HW, C));
  else
    B200_CHECK_CUDA(launch_pdl(avgpool_bwd_kernel<T>, dim3(grid_for((int64_t)B * HW * (C / 4), 256)), dim3(256), 0, st,
                               (const T*)in, (T*)out, B, HW, C));
  B200_CHECK_CUDA(cudaGetLastError());
  return 0;
}
extern "C" int b200_avgpool_fwd(const void* x, void* out, int64_t B, int HW, int C, int dtype, b200_stream_t stream) {
  CONV_COMMON("avgpool_fwd", ALIGNED16(x) && ALIGNED16(out));
  B200_REQUIRE(B > 0 && HW > 0 && C > 0 && C % 4 == 0 && B * HW < (1ll << 31), "avgpool: bad sizes (C %% 4 == 0)");
  return dtype == B200_BF16 ? avgpool_t<bf16_t>(true, x, out, (int)B, HW, C, st) : avgpool_t<float>(true, x, out, (int)B, HW, C, st);
}
extern "C" int b200_avgpool_bwd(const void* dout, void* dx, int64_t B, int HW, int C, int dtype, b200_stream_t stream) {
  CONV_COMMON("avgpool_bwd", ALIGNED16(dout) && ALIGNED16(dx));
  B200_REQUIRE(B > 0 && HW > 0 && C > 0 && C % 4 == 0 && B * HW < (1ll << 31), "avgpool: bad sizes (C %% 4 == 0)");
  return dtype == B200_BF16 ? avgpool_t<bf16_t>(false, dout, dx, (int)B, HW, C, st) : avgpool_t<float>(false, dout, dx, (int)B, HW, C, st);
}
