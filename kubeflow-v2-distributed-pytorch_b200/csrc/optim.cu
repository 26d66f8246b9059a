// optim.cu — K-opt: fused SGD-momentum / Adam over the flat fp32 parameter buffer, the embedding
// gradient scatter-add (K-embbwd), and the small device-side step bookkeeping kernels.
//
// SGD follows torch.optim.SGD as the reference configures it (task.py:212-214: lr, momentum, coupled L2
// weight decay, dampening 0, no nesterov) and is applied once per step where the reference calls
// optimizer.step() (task.py:312).  All of these are pure streaming kernels: 128-bit loads/stores,
// grid sized to a multiple of the SM count, 20 B/param (SGD) or 28 B/param (Adam) of HBM traffic
// plus 2 B/param for the bf16 operand copy where one is kept.
#include <stdlib.h>

#include "common.cuh"

namespace b200 {

constexpr int OPT_THREADS = 256;

__device__ __forceinline__ void store_shadow4(void* shadow, long long i, float a, float b, float c, float d) {
  uint2 o = make_uint2(pack_bf16x2(a, b), pack_bf16x2(c, d));
  *reinterpret_cast<uint2*>(reinterpret_cast<uint16_t*>(shadow) + i) = o;
}

__global__ void __launch_bounds__(OPT_THREADS)
sgd_kernel(float* __restrict__ p, float* __restrict__ g, float* __restrict__ buf,
           void* __restrict__ shadow, long long n, long long n_shadow, float lr, float mu, float wd, int zero_grad,
           const StepAdvance adv) {
  pdl_launch_dependents();
  pdl_wait();
  const long long nv = n >> 2;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long v = (long long)blockIdx.x * blockDim.x + threadIdx.x; v < nv; v += stride) {
    const long long i = v << 2;
    float4 pp = *reinterpret_cast<const float4*>(p + i);
    const float4 gg = *reinterpret_cast<const float4*>(g + i);
    float4 bb = *reinterpret_cast<const float4*>(buf + i);
    // grad.add(param, alpha=wd) -> buf.mul_(mu).add_(grad) -> param.add_(buf, alpha=-lr)
    bb.x = __fadd_rn(__fmul_rn(bb.x, mu), __fmaf_rn(pp.x, wd, gg.x));
    bb.y = __fadd_rn(__fmul_rn(bb.y, mu), __fmaf_rn(pp.y, wd, gg.y));
    bb.z = __fadd_rn(__fmul_rn(bb.z, mu), __fmaf_rn(pp.z, wd, gg.z));
    bb.w = __fadd_rn(__fmul_rn(bb.w, mu), __fmaf_rn(pp.w, wd, gg.w));
    pp.x = __fmaf_rn(bb.x, -lr, pp.x);
    pp.y = __fmaf_rn(bb.y, -lr, pp.y);
    pp.z = __fmaf_rn(bb.z, -lr, pp.z);
    pp.w = __fmaf_rn(bb.w, -lr, pp.w);
    *reinterpret_cast<float4*>(buf + i) = bb;
    *reinterpret_cast<float4*>(p + i) = pp;
    if (i < n_shadow) store_shadow4(shadow, i, pp.x, pp.y, pp.z, pp.w);
    if (zero_grad) *reinterpret_cast<float4*>(g + i) = make_float4(0.f, 0.f, 0.f, 0.f);  // optimizer.zero_grad()
  }
  // scalar tail (n % 4)
  for (long long i = (nv << 2) + (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const float gi = __fmaf_rn(p[i], wd, g[i]);
    const float bi = __fadd_rn(__fmul_rn(buf[i], mu), gi);
    const float pi = __fmaf_rn(bi, -lr, p[i]);
    buf[i] = bi;
    p[i] = pi;
    if (i < n_shadow) reinterpret_cast<__nv_bfloat16*>(shadow)[i] = __float2bfloat16_rn(pi);
    if (zero_grad) g[i] = 0.f;
  }
  advance_if_last_block(adv);
}

__global__ void __launch_bounds__(OPT_THREADS)
adam_kernel(float* __restrict__ p, float* __restrict__ g, float* __restrict__ m, float* __restrict__ v,
            void* __restrict__ shadow, long long n, long long n_shadow, double lr, double b1, double b2, float eps,
            float wd, const int* __restrict__ step, int zero_grad, const StepAdvance adv) {
  pdl_launch_dependents();
  pdl_wait();
  // bias corrections in double like torch (python floats): 1 - beta**t, t = completed steps + 1
  const double t = (double)(*step) + 1.0;
  const float step_size = (float)(lr / (1.0 - pow(b1, t)));
  const float bc2_sqrt = (float)sqrt(1.0 - pow(b2, t));
  const float w1 = (float)(1.0 - b1), w2 = (float)(1.0 - b2), b2f = (float)b2;
  const long long nv = n >> 2;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long q = (long long)blockIdx.x * blockDim.x + threadIdx.x; q < nv; q += stride) {
    const long long i = q << 2;
    float4 pp = *reinterpret_cast<const float4*>(p + i);
    const float4 gg = *reinterpret_cast<const float4*>(g + i);
    float4 mm = *reinterpret_cast<const float4*>(m + i);
    float4 vv = *reinterpret_cast<const float4*>(v + i);
    adam_one(pp.x, gg.x, mm.x, vv.x, wd, w1, b2f, w2, eps, step_size, bc2_sqrt);
    adam_one(pp.y, gg.y, mm.y, vv.y, wd, w1, b2f, w2, eps, step_size, bc2_sqrt);
    adam_one(pp.z, gg.z, mm.z, vv.z, wd, w1, b2f, w2, eps, step_size, bc2_sqrt);
    adam_one(pp.w, gg.w, mm.w, vv.w, wd, w1, b2f, w2, eps, step_size, bc2_sqrt);
    *reinterpret_cast<float4*>(m + i) = mm;
    *reinterpret_cast<float4*>(v + i) = vv;
    *reinterpret_cast<float4*>(p + i) = pp;
    if (i < n_shadow) store_shadow4(shadow, i, pp.x, pp.y, pp.z, pp.w);
    if (zero_grad) *reinterpret_cast<float4*>(g + i) = make_float4(0.f, 0.f, 0.f, 0.f);
  }
  for (long long i = (nv << 2) + (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    float pi = p[i], mi = m[i], vi = v[i];
    adam_one(pi, g[i], mi, vi, wd, w1, b2f, w2, eps, step_size, bc2_sqrt);
    p[i] = pi; m[i] = mi; v[i] = vi;
    if (i < n_shadow) reinterpret_cast<__nv_bfloat16*>(shadow)[i] = __float2bfloat16_rn(pi);
    if (zero_grad) g[i] = 0.f;
  }
  advance_if_last_block(adv);
}

__global__ void __launch_bounds__(OPT_THREADS)
cast_bf16_kernel(const float* __restrict__ src, void* __restrict__ dst, long long n) {
  pdl_launch_dependents();
  pdl_wait();
  const long long nv = n >> 2;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long q = (long long)blockIdx.x * blockDim.x + threadIdx.x; q < nv; q += stride) {
    const float4 s = *reinterpret_cast<const float4*>(src + (q << 2));
    store_shadow4(dst, q << 2, s.x, s.y, s.z, s.w);
  }
  for (long long i = (nv << 2) + (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
    reinterpret_cast<__nv_bfloat16*>(dst)[i] = __float2bfloat16_rn(src[i]);
}

// ---- fp32 operands for error-compensated TF32 products ("3xTF32") ---------------------------------------------------
// The tensor cores have no fp32 multiply.  With x = hi + lo, hi = tf32(x) and lo = x - hi (exact in fp32, then itself
// within 2^-11 of a tf32 number),  a.b = a_hi.b_hi + a_hi.b_lo + a_lo.b_hi  up to 2^-22 relative: three TF32
// products accumulated in fp32.  Laid side by side along the contraction dimension they are ONE ordinary GEMM:
//     [A_hi | A_hi | A_lo] . [B_hi | B_lo | B_hi]^T      (contraction 3x as long)
// so the TF32 kernels stay as they are and this kernel prepares the operands: block j of the output holds hi or lo
// according to bit j of the mask, as three column blocks (`kcat`, for an operand whose contraction dimension is
// its row) and / or three row blocks (`rstack`, contraction over rows).
__device__ __forceinline__ float tf32_hi(float x) {
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
  return __uint_as_float(r);
}
__global__ void __launch_bounds__(OPT_THREADS)
split_tf32_kernel(const float* __restrict__ src, long long ld, long long R, int C, float* __restrict__ kcat,
                  int kmask, float* __restrict__ rstack, int rmask) {
  pdl_launch_dependents();
  pdl_wait();
  const int cv = C >> 2;
  const long long total = R * cv;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long q = (long long)blockIdx.x * blockDim.x + threadIdx.x; q < total; q += stride) {
    const long long r = q / cv;
    const int c = (int)(q - r * cv) << 2;
    const float4 x = *reinterpret_cast<const float4*>(src + r * ld + c);
    const float4 hi = make_float4(tf32_hi(x.x), tf32_hi(x.y), tf32_hi(x.z), tf32_hi(x.w));
    const float4 lo = make_float4(x.x - hi.x, x.y - hi.y, x.z - hi.z, x.w - hi.w);
#pragma unroll
    for (int j = 0; j < 3; ++j) {
      if (kcat) *reinterpret_cast<float4*>(kcat + r * (3ll * C) + (long long)j * C + c) = ((kmask >> j) & 1) ? lo : hi;
      if (rstack) *reinterpret_cast<float4*>(rstack + ((long long)j * R + r) * C + c) = ((rmask >> j) & 1) ? lo : hi;
    }
  }
}

__global__ void advance_step_kernel(int* cursor, int* step, int steps_per_epoch) {
  pdl_launch_dependents();
  pdl_wait();
  if (threadIdx.x == 0 && blockIdx.x == 0) {
    if (cursor) {
      const int c = *cursor + 1;
      *cursor = (c >= steps_per_epoch) ? 0 : c;
    }
    if (step) *step = *step + 1;
  }
}

// ---- K-embbwd ----------------------------------------------------------------------------------------
// One thread per (sample, table, float4 of the embedding row): one aligned 8-byte (bf16) / 16-byte (fp32) load of
// the gradient row (embedding columns start at emb_col0, a multiple of 4 — column 0 in the step's layout) and one
// 128-bit red.global.add.v4.f32 into the table gradient.  Consecutive lanes read consecutive chunks (coalesced);
// the scatter side is index-driven by nature and resolves in L2.
// Measured alternatives (profiles/r02_emb_bwd_experiments.md): a per-table shared-memory copy filled with atomicAdd
// is 2.7x SLOWER (fp32 shared-memory atomics are compare-and-swap loops on sm_100: 90 us against 33 us), and issuing
// the same red.adds from the layer-0 dgrad epilogue (b200_linear_dgrad_emb) takes as long as GEMM + scatter: the
// 6.8 M L2 atomics per step are the floor either way.
template <bool kBf16>
__global__ void __launch_bounds__(OPT_THREADS)
emb_bwd_kernel(const void* __restrict__ dx_, long long lddx, const int32_t* __restrict__ idx,
               float* __restrict__ dtables, long long batch, int emb_col0, int num_cat, long long table_rows,
               int embed_dim, int vpr_shift) {
  pdl_launch_dependents();
  pdl_wait();
  const int vec_per_row = embed_dim >> 2;
  const int vec_per_sample = num_cat * vec_per_row;
  const long long total = batch * vec_per_sample;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long w = (long long)blockIdx.x * blockDim.x + threadIdx.x; w < total; w += stride) {
    const long long b = w / vec_per_sample;
    const int rem = (int)(w - b * vec_per_sample);
    const int t = vpr_shift >= 0 ? (rem >> vpr_shift) : (rem / vec_per_row);
    const int qv = rem - t * vec_per_row;
    float d0, d1, d2, d3;
    if constexpr (kBf16) {
      const uint2 raw = __ldg(reinterpret_cast<const uint2*>(reinterpret_cast<const uint16_t*>(dx_) + b * lddx +
                                                              emb_col0 + 4 * rem));
      d0 = bf16_lo(raw.x); d1 = bf16_hi(raw.x); d2 = bf16_lo(raw.y); d3 = bf16_hi(raw.y);
    } else {
      const float4 raw = __ldg(reinterpret_cast<const float4*>(reinterpret_cast<const float*>(dx_) + b * lddx +
                                                               emb_col0 + 4 * rem));
      d0 = raw.x; d1 = raw.y; d2 = raw.z; d3 = raw.w;
    }
    const long long id = __ldg(idx + b * num_cat + t);
    float* dst = dtables + ((long long)t * table_rows + id) * embed_dim + qv * 4;
    red_add_v4(dst, d0, d1, d2, d3);
  }
}

static int stream_grid(long long work_items) {
  long long blocks = (work_items + OPT_THREADS - 1) / OPT_THREADS;
  const long long cap = (long long)sm_count() * 8;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  return (int)blocks;
}

}  // namespace b200

using namespace b200;

extern "C" int b200_sgd_step(float* p, float* g, float* momentum_buf, void* shadow_bf16, int64_t n,
                             int64_t n_shadow, double lr, double momentum, double weight_decay, int zero_grad,
                             int32_t* adv_cursor, int32_t* adv_step, int32_t steps_per_epoch, uint32_t* adv_ticket,
                             b200_stream_t stream) {
  const StepAdvance adv{adv_cursor, adv_step, steps_per_epoch, adv_ticket};
  B200_REQUIRE(n > 0 && n_shadow >= 0 && n_shadow <= n, "bad sizes n=%lld n_shadow=%lld", (long long)n, (long long)n_shadow);
  B200_REQUIRE(n_shadow % 4 == 0 || n_shadow == n, "n_shadow must be a multiple of 4");
  B200_REQUIRE(((reinterpret_cast<uintptr_t>(p) | reinterpret_cast<uintptr_t>(g) |
                 reinterpret_cast<uintptr_t>(momentum_buf)) & 15) == 0, "p/g/buf must be 16-byte aligned");
  if (!shadow_bf16) n_shadow = 0;
  B200_CHECK_CUDA(launch_pdl(sgd_kernel, dim3(stream_grid(n / 4 + 1)), dim3(OPT_THREADS), 0, static_cast<cudaStream_t>(stream), 
      p, g, momentum_buf, shadow_bf16, n, n_shadow, (float)lr, (float)momentum, (float)weight_decay, zero_grad, adv));
  B200_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int b200_adam_step(float* p, float* g, float* exp_avg, float* exp_avg_sq, void* shadow_bf16,
                              int64_t n, int64_t n_shadow, double lr, double beta1, double beta2, double eps,
                              double weight_decay, const int32_t* step, int zero_grad, int32_t* adv_cursor,
                              int32_t* adv_step, int32_t steps_per_epoch, uint32_t* adv_ticket, b200_stream_t stream) {
  const StepAdvance adv{adv_cursor, adv_step, steps_per_epoch, adv_ticket};
  B200_REQUIRE(n > 0 && n_shadow >= 0 && n_shadow <= n, "bad sizes n=%lld n_shadow=%lld", (long long)n, (long long)n_shadow);
  B200_REQUIRE(n_shadow % 4 == 0 || n_shadow == n, "n_shadow must be a multiple of 4");
  B200_REQUIRE(step != nullptr, "adam needs the device step counter");
  B200_REQUIRE(((reinterpret_cast<uintptr_t>(p) | reinterpret_cast<uintptr_t>(g) | reinterpret_cast<uintptr_t>(exp_avg) |
                 reinterpret_cast<uintptr_t>(exp_avg_sq)) & 15) == 0, "p/g/m/v must be 16-byte aligned");
  if (!shadow_bf16) n_shadow = 0;
  B200_CHECK_CUDA(launch_pdl(adam_kernel, dim3(stream_grid(n / 4 + 1)), dim3(OPT_THREADS), 0, static_cast<cudaStream_t>(stream), 
      p, g, exp_avg, exp_avg_sq, shadow_bf16, n, n_shadow, lr, beta1, beta2, (float)eps, (float)weight_decay, step,
      zero_grad, adv));
  B200_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int b200_cast_bf16(const float* src, void* dst_bf16, int64_t n, b200_stream_t stream) {
  B200_REQUIRE(n > 0, "n must be positive");
  B200_CHECK_CUDA(launch_pdl(cast_bf16_kernel, dim3(stream_grid(n / 4 + 1)), dim3(OPT_THREADS), 0, static_cast<cudaStream_t>(stream), src, dst_bf16, n));
  B200_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int b200_split_tf32(const float* src, int64_t ld, int64_t rows, int64_t cols, float* kcat, int kmask,
                               float* rstack, int rmask, b200_stream_t stream) {
  B200_REQUIRE(src != nullptr && rows > 0 && cols > 0 && cols % 4 == 0 && ld % 4 == 0 && ld >= cols,
               "b200_split_tf32: cols and ld must be positive multiples of 4");
  B200_REQUIRE(kcat != nullptr || rstack != nullptr, "b200_split_tf32: no output given");
  B200_REQUIRE(((reinterpret_cast<uintptr_t>(src) | reinterpret_cast<uintptr_t>(kcat) |
                 reinterpret_cast<uintptr_t>(rstack)) & 15) == 0, "b200_split_tf32: 16-byte aligned buffers needed");
  B200_REQUIRE(kmask >= 0 && kmask < 8 && rmask >= 0 && rmask < 8, "masks are 3 bits");
  B200_CHECK_CUDA(launch_pdl(split_tf32_kernel, dim3(stream_grid(rows * cols / 4)), dim3(OPT_THREADS), 0,
                             static_cast<cudaStream_t>(stream), src, (long long)ld, (long long)rows, (int)cols, kcat,
                             kmask, rstack, rmask));
  B200_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int b200_advance_step(int32_t* cursor, int32_t* step, int32_t steps_per_epoch, b200_stream_t stream) {
  B200_CHECK_CUDA(launch_pdl(advance_step_kernel, dim3(1), dim3(32), 0, static_cast<cudaStream_t>(stream), cursor, step, steps_per_epoch));
  B200_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int b200_emb_bwd(const void* dx, int64_t lddx, const int32_t* idx, float* dtables, int64_t batch,
                            int emb_col0, int num_cat, int64_t table_rows, int embed_dim, int dtype,
                            b200_stream_t stream) {
  B200_REQUIRE(dtype == B200_F32 || dtype == B200_BF16, "unknown dtype code %d", dtype);
  if (num_cat == 0) return 0;
  B200_REQUIRE(embed_dim > 0 && embed_dim % 4 == 0, "embed_dim=%d must be a positive multiple of 4", embed_dim);
  B200_REQUIRE((reinterpret_cast<uintptr_t>(dtables) & 15) == 0, "dtables must be 16-byte aligned");
  B200_REQUIRE(emb_col0 >= 0 && emb_col0 % 4 == 0 && lddx % 4 == 0 && (reinterpret_cast<uintptr_t>(dx) & 15) == 0,
               "emb_bwd: dx, lddx and emb_col0 must keep every 4-element chunk aligned");
  const int vpr = embed_dim / 4;
  int vpr_shift = -1;
  if ((vpr & (vpr - 1)) == 0) {
    vpr_shift = 0;
    while ((1 << vpr_shift) < vpr) ++vpr_shift;
  }
  const long long total = (long long)batch * num_cat * (embed_dim / 4);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (dtype == B200_BF16)
    B200_CHECK_CUDA(launch_pdl(emb_bwd_kernel<true>, dim3(stream_grid(total)), dim3(OPT_THREADS), 0, st, dx, lddx, idx, dtables, batch, emb_col0,
                                                                     num_cat, table_rows, embed_dim, vpr_shift));
  else
    B200_CHECK_CUDA(launch_pdl(emb_bwd_kernel<false>, dim3(stream_grid(total)), dim3(OPT_THREADS), 0, st, dx, lddx, idx, dtables, batch, emb_col0,
                                                                      num_cat, table_rows, embed_dim, vpr_shift));
  B200_CHECK_CUDA(cudaGetLastError());
  return 0;
}
