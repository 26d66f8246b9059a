// gemm_common.cuh — pieces shared by the tcgen05 GEMM kernels (gemm_tcgen05.cu, gemm_ce.cu): tile constants, UMMA
// shared-memory / instruction descriptors, and the host-side tensor-map builder.
#pragma once
#include <stdlib.h>

#include "common.cuh"

namespace b200 {

constexpr int BLOCK_M = 128;
constexpr int BLOCK_N = 256;
constexpr int GEMM_THREADS = 384;  // 4 control warps + 8 epilogue warps
constexpr int EPI_WARPS = 8;
constexpr int A_TILE_BYTES = BLOCK_M * 128;  // 128 rows (or 128-byte swizzle rows) per stage and CTA
constexpr int PIPE_BYTES = 192 * 1024;       // operand ring: 4 x 48 KiB (single CTA) or 6 x 32 KiB (CTA pair)
constexpr int NUM_ACC = 2;                   // TMEM accumulator double buffer (2 x 256 columns)
constexpr int EPI_BUF_BYTES = 32 * 128;                       // one epilogue slab: 32 rows x 128 B, one per warp
constexpr int EPI_STAGING_BYTES = EPI_WARPS * EPI_BUF_BYTES;  // 32 KiB
constexpr int GEMM_SMEM_BYTES = PIPE_BYTES + EPI_STAGING_BYTES + 1024 /*align slack*/ + 256 /*barriers*/;

// ---- UMMA descriptors (bit layout: PTX ISA "tcgen05 matrix/instruction descriptor") ------------
// shared-memory matrix descriptor, 128-byte swizzle, version 1 (Blackwell)
// layout_type 2 = SWIZZLE_128B (16-byte atoms); 1 = SWIZZLE_128B_BASE32B (32-byte atoms), the only layout the
// tensor core accepts for MN-major 32-bit (tf32) operands.
__device__ __forceinline__ uint64_t make_smem_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes,
                                                   uint32_t layout_type = 2) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr & 0x3FFFF) >> 4);
  d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16;
  d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32;
  d |= (uint64_t)1 << 46;  // descriptor version
  d |= (uint64_t)layout_type << 61;
  return d;
}
// instruction descriptor: fp32 accumulate, A/B format bf16(1) or tf32(2), majors, N>>3, M>>4
__host__ __device__ constexpr uint32_t make_idesc(int fmt, int a_mn, int b_mn, int m, int n) {
  return (1u << 4) | ((uint32_t)fmt << 7) | ((uint32_t)fmt << 10) | ((uint32_t)a_mn << 15) |
         ((uint32_t)b_mn << 16) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(m >> 4) << 24);
}

// ---- host side ------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                                  CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                  CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (fn) return fn;
  void* sym = nullptr;
  cudaDriverEntryPointQueryResult qres;
  cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &sym, cudaEnableDefault, &qres);
  if (e != cudaSuccess || qres != cudaDriverEntryPointSuccess || !sym) {
    set_error("cudaGetDriverEntryPoint(cuTensorMapEncodeTiled) failed: %s", cudaGetErrorString(e));
    return nullptr;
  }
  fn = reinterpret_cast<EncodeTiledFn>(sym);
  return fn;
}

// 2-D row-major tensor [outer, inner] with row stride ld (elements); box = box_inner x box_outer, 128B swizzle.
// mn_major operands in fp32 use the 32-byte-atom flavour of the 128B swizzle (see make_smem_desc).
static int make_tmap(CUtensorMap* m, int tf32, const void* ptr, uint64_t inner, uint64_t outer, uint64_t ld,
                     uint32_t box_inner, uint32_t box_outer, bool mn_major = false, bool is_output = false) {
  EncodeTiledFn fn = get_encode_fn();
  if (!fn) return 1;
  const uint64_t es = tf32 ? 4 : 2;
  B200_REQUIRE((reinterpret_cast<uintptr_t>(ptr) & 15) == 0, "GEMM operand %p is not 16-byte aligned", ptr);
  B200_REQUIRE((ld * es) % 16 == 0, "GEMM operand row stride %llu elements is not a multiple of 16 bytes",
               (unsigned long long)ld);
  cuuint64_t dims[2] = {inner, outer};
  cuuint64_t strides[1] = {ld * es};
  cuuint32_t box[2] = {box_inner, box_outer};
  cuuint32_t estr[2] = {1, 1};
  // fp32 operands: the TFLOAT32 tensor-map type rounds to tf32 (nearest) on the way into shared memory; with
  // plain FLOAT32 the tensor core would truncate the low 13 mantissa bits, a biased error that accumulates.
#ifdef B200STEP_EXPERIMENTS
  static const bool tf32_truncate = getenv("B200STEP_TF32_TRUNCATE") != nullptr;  // debugging knob
#else
  constexpr bool tf32_truncate = false;
#endif
  const CUtensorMapDataType dt = !tf32 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16
                                 : ((tf32_truncate || is_output) ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32
                                                                 : CU_TENSOR_MAP_DATA_TYPE_TFLOAT32);
  CUresult r = fn(m, dt, 2,
                  const_cast<void*>(ptr), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                  (tf32 && mn_major) ? CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B : CU_TENSOR_MAP_SWIZZLE_128B,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  B200_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled failed with CUresult %d (inner=%llu outer=%llu ld=%llu)",
               (int)r, (unsigned long long)inner, (unsigned long long)outer, (unsigned long long)ld);
  return 0;
}

// ---- im2col-mode tensor map: the A operand of a convolution read straight from the NHWC activation ---------------
// One TMA instruction then delivers `pixels` consecutive OUTPUT pixels (row-major over Wo, Ho, N — the GEMM's row
// order) x `channels` input channels of ONE filter tap into a 128B-swizzled tile, zero-filling the padding halo: the
// `col` matrix never exists.  Bounding box as in cuTensorMapEncodeIm2col: lower corner = -pad, upper corner =
// pad - (k - 1); traversal stride = the convolution's stride.  Instruction coordinates: (c, w, h, n) = (channel offset,
// q*stride - pad, p*stride - pad, image) and offsets (kx, ky).
typedef CUresult (*EncodeIm2colFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                   const cuuint64_t*, const int*, const int*, cuuint32_t, cuuint32_t, const cuuint32_t*,
                                   CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                   CUtensorMapFloatOOBfill);
static EncodeIm2colFn get_encode_im2col_fn() {
  static EncodeIm2colFn fn = nullptr;
  if (fn) return fn;
  void* sym = nullptr;
  cudaDriverEntryPointQueryResult qres;
  cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeIm2col", &sym, cudaEnableDefault, &qres);
  if (e != cudaSuccess || qres != cudaDriverEntryPointSuccess || !sym) {
    set_error("cudaGetDriverEntryPoint(cuTensorMapEncodeIm2col) failed: %s", cudaGetErrorString(e));
    return nullptr;
  }
  fn = reinterpret_cast<EncodeIm2colFn>(sym);
  return fn;
}
static int make_tmap_im2col(CUtensorMap* m, int tf32, const void* x, int64_t B, int H, int W, int C, int kh, int kw,
                            int stride, int pad, uint32_t channels, uint32_t pixels, bool mn_major = false) {
  EncodeIm2colFn fn = get_encode_im2col_fn();
  if (!fn) return 1;
  const uint64_t es = tf32 ? 4 : 2;
  B200_REQUIRE((reinterpret_cast<uintptr_t>(x) & 15) == 0 && (C * es) % 16 == 0,
               "conv operand %p must be 16-byte aligned with C*elemsize a multiple of 16", x);
  cuuint64_t dims[4] = {(cuuint64_t)C, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)B};
  cuuint64_t strides[3] = {C * es, (cuuint64_t)W * C * es, (cuuint64_t)H * W * C * es};
  int lower[2] = {-pad, -pad};
  int upper[2] = {pad - (kw - 1), pad - (kh - 1)};
  cuuint32_t estr[4] = {1, (cuuint32_t)stride, (cuuint32_t)stride, 1};
  CUresult r = fn(m, tf32 ? CU_TENSOR_MAP_DATA_TYPE_TFLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4,
                  const_cast<void*>(x), dims, strides, lower, upper, channels, pixels, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE,
                  (tf32 && mn_major) ? CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B : CU_TENSOR_MAP_SWIZZLE_128B,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  B200_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeIm2col failed with CUresult %d (B=%lld H=%d W=%d C=%d k=%dx%d s=%d p=%d)",
               (int)r, (long long)B, H, W, C, kh, kw, stride, pad);
  // drivers up to 13.1 set a descriptor bit that is wrong for tensors under 128 KiB (same correction as CUTLASS's
  // make_im2col_tma_copy_desc)
  int drv = 0;
  if (cudaDriverGetVersion(&drv) == cudaSuccess && drv <= 13010 && (uint64_t)B * H * W * C * es < 131072)
    reinterpret_cast<uint64_t*>(m)[1] &= ~(1ull << 21);
  return 0;
}

static int check_dims(int64_t M, int64_t N, int64_t K, int dtype) {
  B200_REQUIRE(dtype == B200_F32 || dtype == B200_BF16, "unknown dtype code %d", dtype);
  B200_REQUIRE(M > 0 && N > 0 && K > 0, "GEMM dims must be positive (M=%lld N=%lld K=%lld)", (long long)M,
               (long long)N, (long long)K);
  B200_REQUIRE(M < (1ll << 31) && N < (1ll << 31) && K < (1ll << 31), "GEMM dims must fit int32");
  const int v = dtype == B200_F32 ? 4 : 8;
  B200_REQUIRE(N % v == 0 && K % v == 0, "N=%lld and K=%lld must be multiples of %d for dtype %d", (long long)N,
               (long long)K, v, dtype);
  return 0;
}


}  // namespace b200
