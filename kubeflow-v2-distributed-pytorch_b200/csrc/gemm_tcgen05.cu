// gemm_tcgen05.cu — K-gemm: the linear layers of the step as tcgen05/TMEM GEMMs (sm_100a only).
//
// One persistent, warp-specialised kernel template covers the three contractions of nn.Linear
// (reference call sites: forward task.py:309, backward task.py:311):
//   fwd   D[batch,N]   = X[batch,K] · W[N,K]^T      A K-major,  B K-major   epilogue: +bias, ReLU, dropout
//   dgrad D[batch,K]   = dY[batch,N] · W[N,K]       A K-major,  B MN-major  epilogue: *dscale, ReLU mask
//   wgrad D[N,K]       = dY[batch,N]^T · X[batch,K] A MN-major, B MN-major  epilogue: fp32 store / split-K red.add
// Operands are staged by TMA into 128B-swizzled shared-memory tiles, multiplied by single-thread
// tcgen05.mma (UMMA 128 x 256 x {16 bf16 | 8 tf32}) into double-buffered TMEM accumulators, and drained
// by four epilogue warps with tcgen05.ld while the next tile's MMAs run.
//
// Roles (384 threads): warp0 = TMA producer, warp1 = MMA issuer, warp2 = TMEM allocator, warps4-11 = epilogue
// (two warps per TMEM lane quarter, each draining half of the tile's columns).
#include "gemm_common.cuh"

namespace b200 {

enum { EPI_FWD = 0, EPI_DGRAD = 1, EPI_WGRAD = 2, EPI_EMB = 3 };

struct GemmParams {
  int M, N;          // D is M x N
  int k_blocks;      // ceil(contraction / BLOCK_K)
  int m_tiles, n_tiles, splits, kb_per_split;
  void* out;
  long long ldo;
  const float* bias;
  const void* act;
  long long ldact;
  uint32_t* mask_out;        // fwd: bit j of word [row, col/32] = (output[row, col] > 0)
  const uint32_t* mask_in;   // dgrad: same layout, replaces reading `act`
  long long ldmask;          // row stride of the bit mask in 32-bit words
  float* cs_partials;        // dgrad: per (CTA, lane quarter) column sums of the stored output (deterministic mode)
  float* cs_atomic;          // dgrad: or add each CTA's column sums straight onto this zeroed [N] vector
  float dscale;
  int flags;
  float dropout_p;
  unsigned long long seed;
  const int* step;
  int layer_id;
  long long row0;            // fwd dropout: row index of this call's first row inside the whole batch
  // EPI_EMB: the output columns are the concatenated embedding rows of a sample; they are not stored but added
  // onto the table gradients, out = dtables[num_tables][emb_rows][1 << emb_shift]
  const int32_t* emb_idx;    // [M, ld_idx] table row of sample m in table t
  long long ld_idx;
  long long emb_rows;
  int emb_shift;             // log2(embedding width)
  // convolution as implicit GEMM: the K-major A operand comes from an im2col-mode tensor map over the NHWC input
  // (k-block kb = filter tap kb / cv_cpb, channels (kb % cv_cpb) * BLOCK_K ...); 0 = plain matrix A
  int cv_on, cv_Ho, cv_Wo, cv_stride, cv_pad, cv_kw, cv_cpb;
  int n_eff;                 // UMMA N of this launch: 256, or 128 / 64 when the whole output is that narrow (one column
                             // tile): the MMAs then skip the unused accumulator columns (0 = choose in launch_gemm)
  int cv_C, cv_K;            // wgrad (cv_on 1, MN-major B from the input): channels and kh*kw*C;
                             // dgrad (cv_on 2): input channels and the number of filter taps
};

// kCtas = 2: the two CTAs of a cluster form a pair (tcgen05 cta_group::2).  The pair owns a 256 x 256 tile: each CTA
// loads its own 128 rows of A and HALF of the B tile (so L2->smem operand traffic per FLOP drops by a third), the
// leader CTA's single MMA thread issues M=256 UMMAs that read both CTAs' shared memory and write both CTAs'
// TMEM, and each CTA drains its own 128 accumulator rows.
template <int kTf32, bool kAMN, bool kBMN, int kEpi, int kCtas>
__global__ void __launch_bounds__(GEMM_THREADS, 1)
gemm_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
            const __grid_constant__ CUtensorMap tmC, const GemmParams p) {
  constexpr int ESIZE = kTf32 ? 4 : 2;
  constexpr int BLOCK_K = 128 / ESIZE;          // one 128-byte swizzle row of K (K-major) / BLOCK_K rows (MN-major)
  constexpr int UMMA_K = 32 / ESIZE;            // 16 (bf16) or 8 (tf32)
  constexpr int EPR = 128 / ESIZE;              // elements per 128-byte row
  constexpr int MN_BOX_BYTES = BLOCK_K * 128;   // one TMA box of an MN-major operand: BLOCK_K rows x 128 B
  constexpr uint32_t MN_SBO = kTf32 ? 512 : 1024;   // swizzle atom height: 4 rows (32B atoms) / 8 rows (16B atoms)
  constexpr uint32_t MN_LAYOUT = kTf32 ? 1 : 2;
  constexpr uint32_t IDESC = make_idesc(kTf32 ? 2 : 1, kAMN, kBMN, BLOCK_M * kCtas, BLOCK_N);
  constexpr int B_ROWS = BLOCK_N / kCtas;             // B-tile rows (output columns) staged by this CTA
  constexpr int STAGE_BYTES = A_TILE_BYTES + B_ROWS * 128;
  constexpr int NUM_STAGES = PIPE_BYTES / STAGE_BYTES;
  constexpr int TILE_M = BLOCK_M * kCtas;             // output rows per work item (per CTA pair)

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + PIPE_BYTES + EPI_STAGING_BYTES);
  uint64_t* full_bar = bars;                       // [NUM_STAGES]
  uint64_t* empty_bar = bars + NUM_STAGES;         // [NUM_STAGES]
  uint64_t* acc_full = bars + 2 * NUM_STAGES;      // [NUM_ACC]
  uint64_t* acc_empty = acc_full + NUM_ACC;        // [NUM_ACC]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_empty + NUM_ACC);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int rank = kCtas == 2 ? (int)cluster_ctarank() : 0;   // 0 = leader of the pair
  const int unit = blockIdx.x / kCtas, nunits = gridDim.x / kCtas;

  pdl_launch_dependents();  // the next kernel may start its own prologue
  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    if constexpr (kEpi != EPI_WGRAD && kEpi != EPI_EMB) tma_prefetch_desc(&tmC);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < NUM_STAGES; ++i) {
      mbar_init(&full_bar[i], 1);
      mbar_init(&empty_bar[i], 1);
    }
    for (int i = 0; i < NUM_ACC; ++i) {
      mbar_init(&acc_full[i], 1);
      mbar_init(&acc_empty[i], EPI_WARPS * kCtas);  // one arrive per epilogue warp (of both CTAs)
    }
    fence_mbar_init();
  }
  if (warp == 2) {
    if constexpr (kCtas == 2) tmem_alloc_pair(tmem_slot, NUM_ACC * BLOCK_N);
    else tmem_alloc(tmem_slot, NUM_ACC * BLOCK_N);
  }
  tc_fence_before();
  if constexpr (kCtas == 2) cluster_sync_all();  // peer barriers must be initialised before any remote arrive
  else __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_wait();  // barriers, TMEM and descriptors are set up; operands come from the previous kernel

  const int tiles = p.m_tiles * p.n_tiles;
  const int total = tiles * p.splits;

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (lane == 0) {
      int s = 0;
      uint32_t ph = 0;
      // B rows this CTA stages: its half of the columns the MMAs compute.  Narrow outputs (n_eff = 64 / 128) stage a
      // quarter / half of the B tile: fewer TMA row requests per stage (the tensor map's box, built by the entry point
      // from the same n_eff, is that small for K-major B; MN-major B simply issues fewer 128-byte column boxes)
      const int brows = p.n_eff / kCtas;
      const int b_boxes = brows >= EPR ? brows / EPR : 1;
      const uint32_t stage_tx = A_TILE_BYTES + (kBMN ? b_boxes * MN_BOX_BYTES : brows * 128);
      for (int w = unit; w < total; w += nunits) {
        const int tile = w % tiles, split = w / tiles;
        int mt = tile / p.n_tiles;
        if (p.flags & B200_FLAG_REVERSE_M) mt = p.m_tiles - 1 - mt;       // walk the row tiles from the last one down
        const int m0 = mt * TILE_M + rank * BLOCK_M;                      // this CTA's rows of A
        // this CTA's share of the B tile: half of the columns the MMA really computes (p.n_eff <= BLOCK_N)
        const int n0 = (tile % p.n_tiles) * BLOCK_N + rank * (p.n_eff / kCtas);
        const int kb0 = split * p.kb_per_split;
        const int kb1 = min(kb0 + p.kb_per_split, p.k_blocks);
        int cv_w = 0, cv_h = 0, cv_n = 0;       // implicit GEMM: window origin of this CTA's first output pixel
        if (p.cv_on) {
          const int q = m0 % p.cv_Wo, t = m0 / p.cv_Wo;
          cv_w = q * p.cv_stride - p.cv_pad;
          cv_h = (t % p.cv_Ho) * p.cv_stride - p.cv_pad;
          cv_n = t / p.cv_Ho;
        }
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(&empty_bar[s], ph ^ 1);
          uint8_t* sa = smem + s * STAGE_BYTES;
          uint8_t* sb = sa + A_TILE_BYTES;
          if constexpr (kCtas == 2) {
            // both CTAs' bytes are credited to the LEADER's full barrier (the MMA thread waits there)
            if (rank == 0) mbar_expect_tx(&full_bar[s], 2 * stage_tx);
            const uint32_t fb = smem_u32(&full_bar[s]) & 0xFEFFFFFFu;  // same offset in the leader CTA
            if constexpr (!kAMN) {
              if (p.cv_on) {
                const int tap = kb / p.cv_cpb, ky = tap / p.cv_kw;
                tma_load_im2col_pair(sa, &tmA, fb, (kb - tap * p.cv_cpb) * BLOCK_K, cv_w, cv_h, cv_n,
                                     (uint16_t)(tap - ky * p.cv_kw), (uint16_t)ky);
              } else {
                tma_load_2d_pair(sa, &tmA, fb, kb * BLOCK_K, m0);
              }
            } else {
#pragma unroll
              for (int j = 0; j < BLOCK_M / EPR; ++j)
                tma_load_2d_pair(sa + j * MN_BOX_BYTES, &tmA, fb, m0 + j * EPR, kb * BLOCK_K);
            }
            if constexpr (!kBMN) {
              tma_load_2d_pair(sb, &tmB, fb, kb * BLOCK_K, n0);
            } else if (p.cv_on == 2) {
              // dgrad of a stride-1 convolution = a convolution of dy with the flipped filter: k-block kb is
              // (tap', Cout slice); B(n' = ci, k' = co) is the [co][ci] block of Wg at the FLIPPED tap's columns
              const int tap = kb / p.cv_cpb, co0 = (kb - tap * p.cv_cpb) * BLOCK_K;
              const int wc0 = (p.cv_K - 1 - tap) * p.cv_C + n0;
#pragma unroll
              for (int j = 0; j < B_ROWS / EPR; ++j) if (j < b_boxes)
                tma_load_2d_pair(sb + j * MN_BOX_BYTES, &tmB, fb, wc0 + j * EPR, co0);
            } else if (p.cv_on) {
              // wgrad of a convolution: B(n' = filter column, k' = pixel) straight from the NHWC input; this k-block
              // is BLOCK_K consecutive output pixels, each 128-byte column group one (tap, channel slice)
              const int pix = kb * BLOCK_K, q = pix % p.cv_Wo, t = pix / p.cv_Wo;
              const int bw = q * p.cv_stride - p.cv_pad, bh = (t % p.cv_Ho) * p.cv_stride - p.cv_pad, bn = t / p.cv_Ho;
#pragma unroll
              for (int j = 0; j < B_ROWS / EPR; ++j) if (j < b_boxes) {
                const int col0 = n0 + j * EPR;
                const bool in = col0 < p.cv_K;
                const int tap = in ? col0 / p.cv_C : 0, ky = tap / p.cv_kw;
                tma_load_im2col_pair(sb + j * MN_BOX_BYTES, &tmB, fb, in ? col0 - tap * p.cv_C : p.cv_C, bw, bh, bn,
                                     (uint16_t)(tap - ky * p.cv_kw), (uint16_t)ky);
              }
            } else {
#pragma unroll
              for (int j = 0; j < B_ROWS / EPR; ++j) if (j < b_boxes)
                tma_load_2d_pair(sb + j * MN_BOX_BYTES, &tmB, fb, n0 + j * EPR, kb * BLOCK_K);
            }
          } else {
            mbar_expect_tx(&full_bar[s], stage_tx);
            if constexpr (!kAMN) {
              if (p.cv_on) {
                const int tap = kb / p.cv_cpb, ky = tap / p.cv_kw;
                tma_load_im2col(sa, &tmA, &full_bar[s], (kb - tap * p.cv_cpb) * BLOCK_K, cv_w, cv_h, cv_n,
                                (uint16_t)(tap - ky * p.cv_kw), (uint16_t)ky);
              } else {
                tma_load_2d(sa, &tmA, &full_bar[s], kb * BLOCK_K, m0);
              }
            } else {
#pragma unroll
              for (int j = 0; j < BLOCK_M / EPR; ++j)
                tma_load_2d(sa + j * MN_BOX_BYTES, &tmA, &full_bar[s], m0 + j * EPR, kb * BLOCK_K);
            }
            if constexpr (!kBMN) {
              tma_load_2d(sb, &tmB, &full_bar[s], kb * BLOCK_K, n0);
            } else if (p.cv_on == 2) {
              const int tap = kb / p.cv_cpb, co0 = (kb - tap * p.cv_cpb) * BLOCK_K;
              const int wc0 = (p.cv_K - 1 - tap) * p.cv_C + n0;
#pragma unroll
              for (int j = 0; j < B_ROWS / EPR; ++j) if (j < b_boxes)
                tma_load_2d(sb + j * MN_BOX_BYTES, &tmB, &full_bar[s], wc0 + j * EPR, co0);
            } else if (p.cv_on) {
              const int pix = kb * BLOCK_K, q = pix % p.cv_Wo, t = pix / p.cv_Wo;
              const int bw = q * p.cv_stride - p.cv_pad, bh = (t % p.cv_Ho) * p.cv_stride - p.cv_pad, bn = t / p.cv_Ho;
#pragma unroll
              for (int j = 0; j < B_ROWS / EPR; ++j) if (j < b_boxes) {
                const int col0 = n0 + j * EPR;
                const bool in = col0 < p.cv_K;
                const int tap = in ? col0 / p.cv_C : 0, ky = tap / p.cv_kw;
                tma_load_im2col(sb + j * MN_BOX_BYTES, &tmB, &full_bar[s], in ? col0 - tap * p.cv_C : p.cv_C, bw, bh, bn,
                                (uint16_t)(tap - ky * p.cv_kw), (uint16_t)ky);
              }
            } else {
#pragma unroll
              for (int j = 0; j < B_ROWS / EPR; ++j) if (j < b_boxes)
                tma_load_2d(sb + j * MN_BOX_BYTES, &tmB, &full_bar[s], n0 + j * EPR, kb * BLOCK_K);
            }
          }
          if (++s == NUM_STAGES) { s = 0; ph ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer (one thread) =====================
    if (lane == 0 && rank == 0) {
      int s = 0;
      uint32_t ph = 0;
      int it = 0;
      // narrow outputs (N <= 64 / 128 in a single column tile): issue N = 64 / 128 MMAs instead of 256-wide ones
      const uint32_t idesc = p.n_eff == BLOCK_N ? IDESC : make_idesc(kTf32 ? 2 : 1, kAMN, kBMN, BLOCK_M * kCtas, p.n_eff);
      for (int w = unit; w < total; w += nunits, ++it) {
        const int split = w / tiles;
        const int kb0 = split * p.kb_per_split;
        const int kb1 = min(kb0 + p.kb_per_split, p.k_blocks);
        const int a = it & 1;
        const uint32_t aph = (it >> 1) & 1;
        mbar_wait(&acc_empty[a], aph ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + a * BLOCK_N;
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(&full_bar[s], ph);
          tc_fence_after();
          const uint32_t sa = smem_u32(smem + s * STAGE_BYTES);
          const uint32_t sb = sa + A_TILE_BYTES;
#pragma unroll
          for (int k = 0; k < BLOCK_K / UMMA_K; ++k) {
            // K-major: advance 32 B inside the swizzle row; MN-major: advance UMMA_K rows of 128 B
            // (MN-major: LBO = next 128-byte MN atom = next TMA box, SBO = next group of 8 (bf16) / 4 (tf32) K rows)
            const uint64_t ad = kAMN ? make_smem_desc(sa + k * UMMA_K * 128, MN_BOX_BYTES, MN_SBO, MN_LAYOUT)
                                     : make_smem_desc(sa + k * 32, 16, 1024);
            const uint64_t bd = kBMN ? make_smem_desc(sb + k * UMMA_K * 128, MN_BOX_BYTES, MN_SBO, MN_LAYOUT)
                                     : make_smem_desc(sb + k * 32, 16, 1024);
#ifdef B200STEP_EXPERIMENTS
            if (p.flags & 0x800) continue;  // timing experiment: stream the operands, skip the MMAs
#endif
            if constexpr (kCtas == 2) umma_ss_pair<kTf32>(d_tmem, ad, bd, idesc, (kb > kb0 || k > 0) ? 1u : 0u);
            else umma_ss<kTf32>(d_tmem, ad, bd, idesc, (kb > kb0 || k > 0) ? 1u : 0u);
          }
          // frees the smem stage (in both CTAs of a pair) once these MMAs retire
          if constexpr (kCtas == 2) umma_commit_pair(&empty_bar[s], 3);
          else umma_commit(&empty_bar[s]);
          if (++s == NUM_STAGES) { s = 0; ph ^= 1; }
        }
        // accumulator complete -> epilogue warps (of both CTAs)
        if constexpr (kCtas == 2) umma_commit_pair(&acc_full[a], 3);
        else umma_commit(&acc_full[a]);
      }
    }
  } else if (warp >= 4) {
    // ===================== epilogue: TMEM -> registers -> (swizzled smem -> TMA store | red.add) =============
    // Two warps share each TMEM lane quarter (a warp may only read lanes 32*(warp%4)..+31) and split the tile's
    // columns: with a single warp per scheduler the epilogue ALU work could not hide its own issue latency.
    const int q = warp & 3;             // TMEM lane quarter
    const int chalf = (warp - 4) >> 2;  // column half of the tile this warp drains
    int it = 0;
    uint8_t* sbuf_ptr = smem + PIPE_BYTES + (warp - 4) * EPI_BUF_BYTES;
    // dgrad: running column sums of everything this warp stores (lane i owns column chunk*32 + i); the host sizes
    // the grid to a multiple of n_tiles, so a CTA sees the same n_tile in every tile it processes
    float csacc[4] = {0.f, 0.f, 0.f, 0.f};
    for (int w = unit; w < total; w += nunits, ++it) {
      const int tile = w % tiles;
      int mt = tile / p.n_tiles;
      if (p.flags & B200_FLAG_REVERSE_M) mt = p.m_tiles - 1 - mt;
      const int m0 = mt * TILE_M + rank * BLOCK_M, n0 = (tile % p.n_tiles) * BLOCK_N;
      const int a = it & 1;
      const uint32_t aph = (it >> 1) & 1;
      const int row0 = m0 + q * 32;
      const int row = row0 + lane;
      const bool row_ok = row < p.M;
      // this row's 128 ReLU-mask bits for the warp's column half.  dgrad: fetched while the MMAs still run;
      // fwd: collected chunk by chunk and written once per tile
      uint32_t mrow[4] = {0u, 0u, 0u, 0u};
      if constexpr (kEpi == EPI_DGRAD) {
        if ((p.flags & B200_FLAG_RELU) && p.mask_in != nullptr && row_ok) {
          const long long wbase = (n0 >> 5) + chalf * 4;  // ldmask % 4 == 0: the uint4 is fully in or fully out
          if (wbase + 4 <= p.ldmask) {
            const uint4 m4 = __ldg(reinterpret_cast<const uint4*>(p.mask_in + (long long)row * p.ldmask + wbase));
            mrow[0] = m4.x; mrow[1] = m4.y; mrow[2] = m4.z; mrow[3] = m4.w;
          }
        }
      }
      mbar_wait(&acc_full[a], aph);
      tc_fence_after();
      const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + a * BLOCK_N;
      if constexpr (kEpi == EPI_WGRAD) {
#pragma unroll 1
        for (int c = chalf * (BLOCK_N / 64); c < (chalf + 1) * (BLOCK_N / 64); ++c) {
          const int col0 = n0 + c * 32;
          if (col0 >= p.N) break;  // warp-uniform
          uint32_t v[32];
          tmem_ld32(taddr + c * 32, v);
          tmem_ld_wait();
          if (!row_ok) continue;
          float* o = reinterpret_cast<float*>(p.out) + (long long)row * p.ldo + col0;
          const bool atomic = (p.splits > 1) || (p.flags & B200_FLAG_ACCUMULATE);
#pragma unroll
          for (int j = 0; j < 32; j += 4) {
            if (col0 + j < p.N) {
              if (atomic) {
                red_add_v4(o + j, __uint_as_float(v[j]), __uint_as_float(v[j + 1]),
                           __uint_as_float(v[j + 2]), __uint_as_float(v[j + 3]));
              } else {
                *reinterpret_cast<uint4*>(o + j) = make_uint4(v[j], v[j + 1], v[j + 2], v[j + 3]);
              }
            }
          }
        }
      } else if constexpr (kEpi == EPI_EMB) {
        // Layer-0 dgrad whose output row is [26 x dim embedding gradients]: nothing is stored.  Every 16 bytes of
        // accumulator go straight onto the gradient of the table row the sample looked up (fp32 red.add, no bf16
        // rounding on the way): the dx0 matrix, its HBM round trip and the separate scatter kernel disappear.
        const int dim = 1 << p.emb_shift;
#pragma unroll 1
        for (int c = chalf * (BLOCK_N / 64); c < (chalf + 1) * (BLOCK_N / 64); ++c) {
          const int col0 = n0 + c * 32;
          if (col0 >= p.N) break;  // warp-uniform
          uint32_t v[32];
          tmem_ld32(taddr + c * 32, v);
          tmem_ld_wait();
          if (!row_ok) continue;
          const int32_t* ids = p.emb_idx + (long long)row * p.ld_idx;
          float* dst = nullptr;
#pragma unroll
          for (int j = 0; j < 32; j += 4) {
            const int col = col0 + j;
            if (col < p.N) {  // N = num_tables * dim, a multiple of 4
              const int off = col & (dim - 1);
              if (j == 0 || off == 0) {  // first chunk of this thread's slice of a table row
                const int t = col >> p.emb_shift;
                dst = reinterpret_cast<float*>(p.out) + (((long long)t * p.emb_rows + __ldg(ids + t)) << p.emb_shift) + off;
              }
              red_add_v4(dst, __uint_as_float(v[j]), __uint_as_float(v[j + 1]), __uint_as_float(v[j + 2]),
                         __uint_as_float(v[j + 3]));
              dst += 4;
            }
          }
        }
      } else {
        // Output leaves in groups of one 128-byte row per thread (64 bf16 / 32 fp32 columns): each warp builds a
        // 32-row x 128-byte slab in 128B-swizzled shared memory and one lane hands it to the TMA engine, which
        // writes full lines and clips rows >= M / columns >= N.
        constexpr int GROUP_COLS = EPR;
        constexpr int LDS_PER_GROUP = GROUP_COLS / 32;
        constexpr int GROUPS_PER_WARP = BLOCK_N / GROUP_COLS / 2;
        static_assert(GROUPS_PER_WARP * LDS_PER_GROUP == 4, "csacc holds four 32-column chunks per warp");
#pragma unroll
        for (int gi = 0; gi < GROUPS_PER_WARP; ++gi) {
          const int g = chalf * GROUPS_PER_WARP + gi;
          const int gcol0 = n0 + g * GROUP_COLS;
          if (gcol0 >= p.N) break;  // warp-uniform
#ifdef B200STEP_EXPERIMENTS
          if (p.flags & 0x200) break;  // timing experiment
#endif
          uint32_t vv[LDS_PER_GROUP][32];  // the whole group's accumulators: one TMEM round trip, not two
#pragma unroll
          for (int h = 0; h < LDS_PER_GROUP; ++h) tmem_ld32(taddr + g * GROUP_COLS + h * 32, vv[h]);
          tmem_ld_wait();
#pragma unroll
          for (int h = 0; h < LDS_PER_GROUP; ++h) {
            const int col0 = gcol0 + h * 32;
            float f[32];
#pragma unroll
            for (int j = 0; j < 32; ++j) f[j] = __uint_as_float(vv[h][j]);
            if constexpr (kEpi == EPI_FWD) {
              if (p.bias != nullptr) {
#pragma unroll
                for (int j = 0; j < 32; j += 4) {
                  if (col0 + j < p.N) {  // N % 4 == 0: a float4 never straddles the edge
                    const float4 b4 = __ldg(reinterpret_cast<const float4*>(p.bias + col0 + j));
                    f[j] += b4.x; f[j + 1] += b4.y; f[j + 2] += b4.z; f[j + 3] += b4.w;
                  }
                }
              }
              if (p.flags & B200_FLAG_RELU) {
#pragma unroll
                for (int j = 0; j < 32; ++j) f[j] = fmaxf(f[j], 0.0f);
              }
              if (p.dropout_p > 0.0f) {
                // keep iff u32 >= p * 2^32; counter = linear element index / 4, key = (seed, layer, step)
                const uint32_t thresh = (uint32_t)fminf(p.dropout_p * 4294967296.0f, 4294967040.0f);
                const float inv_keep = 1.0f / (1.0f - p.dropout_p);
                const uint32_t stepv = p.step ? (uint32_t)(*p.step) : 0u;
                const uint2 key = make_uint2((uint32_t)p.seed, (uint32_t)(p.seed >> 32) ^ stepv);
#pragma unroll
                for (int j = 0; j < 32; j += 4) {
                  const unsigned long long e = (unsigned long long)(row + p.row0) * (unsigned long long)p.N + (col0 + j);
                  const uint4 r = philox4x32_10(
                      make_uint4((uint32_t)(e >> 2), (uint32_t)(e >> 34), (uint32_t)p.layer_id, 0u), key);
                  f[j + 0] = (r.x >= thresh) ? f[j + 0] * inv_keep : 0.0f;
                  f[j + 1] = (r.y >= thresh) ? f[j + 1] * inv_keep : 0.0f;
                  f[j + 2] = (r.z >= thresh) ? f[j + 2] * inv_keep : 0.0f;
                  f[j + 3] = (r.w >= thresh) ? f[j + 3] * inv_keep : 0.0f;
                }
              }
              if (p.mask_out != nullptr) {  // 1 bit per element for the backward ReLU(/dropout) mask
                uint32_t mbits = 0;
#pragma unroll
                for (int j = 0; j < 32; ++j) mbits |= (f[j] > 0.0f) ? (1u << j) : 0u;
                mrow[gi * LDS_PER_GROUP + h] = mbits;
              }
            } else {  // EPI_DGRAD
              if (p.dscale != 1.0f) {
#pragma unroll
                for (int j = 0; j < 32; ++j) f[j] *= p.dscale;
              }
              if ((p.flags & B200_FLAG_RELU) && p.mask_in != nullptr) {
                const uint32_t mbits = mrow[gi * LDS_PER_GROUP + h];
#pragma unroll
                for (int j = 0; j < 32; ++j) f[j] = (mbits & (1u << j)) ? f[j] : 0.0f;
              } else if ((p.flags & B200_FLAG_RELU) && p.act != nullptr && row_ok) {
                if constexpr (kTf32) {
                  const float* ar = reinterpret_cast<const float*>(p.act) + (long long)row * p.ldact + col0;
#pragma unroll
                  for (int j = 0; j < 32; j += 4) {
                    if (col0 + j < p.N) {
                      const float4 hh = *reinterpret_cast<const float4*>(ar + j);
                      f[j + 0] = hh.x > 0.f ? f[j + 0] : 0.f;
                      f[j + 1] = hh.y > 0.f ? f[j + 1] : 0.f;
                      f[j + 2] = hh.z > 0.f ? f[j + 2] : 0.f;
                      f[j + 3] = hh.w > 0.f ? f[j + 3] : 0.f;
                    }
                  }
                } else {
                  const uint16_t* ar = reinterpret_cast<const uint16_t*>(p.act) + (long long)row * p.ldact + col0;
#pragma unroll
                  for (int j = 0; j < 32; j += 8) {
                    if (col0 + j < p.N) {
                      const uint4 hh = *reinterpret_cast<const uint4*>(ar + j);
                      const uint32_t hw[4] = {hh.x, hh.y, hh.z, hh.w};
#pragma unroll
                      for (int t = 0; t < 4; ++t) {
                        f[j + 2 * t + 0] = bf16_lo(hw[t]) > 0.f ? f[j + 2 * t + 0] : 0.f;
                        f[j + 2 * t + 1] = bf16_hi(hw[t]) > 0.f ? f[j + 2 * t + 1] : 0.f;
                      }
                    }
                  }
                }
              }
            }
            if constexpr (kEpi == EPI_DGRAD) {
              if (p.cs_partials != nullptr || p.cs_atomic != nullptr) {
                // column sums over this warp's 32 rows of what is stored (rounded to the output dtype):
                // halving butterfly, 31 shuffles, lane i ends with the sum of column col0 + i
                float r[32];
#pragma unroll
                for (int j = 0; j < 32; ++j)
                  r[j] = kTf32 ? f[j] : __bfloat162float(__float2bfloat16_rn(f[j]));
#pragma unroll
                for (int half = 16; half >= 1; half >>= 1) {
                  const bool upper = (lane & half) != 0;
#pragma unroll
                  for (int j = 0; j < half; ++j) {
                    const float send = upper ? r[j] : r[j + half];
                    const float keep = upper ? r[j + half] : r[j];
                    r[j] = keep + __shfl_xor_sync(0xffffffffu, send, half);
                  }
                }
                csacc[gi * LDS_PER_GROUP + h] += r[0];
              }
            }
            if (h == 0) {  // the previous store of this warp must have finished READING the slab before we refill it
              if (lane == 0) bulk_wait_read<0>();
              __syncwarp();
            }
            // registers -> swizzled slab: 16-byte chunk c of row `lane` lives at chunk (c ^ (lane & 7))
            const uint32_t srow = smem_u32(sbuf_ptr) + lane * 128;
            if constexpr (kTf32) {
#pragma unroll
              for (int c = 0; c < 8; ++c)
                st_shared_v4(srow + ((c ^ (lane & 7)) << 4), __float_as_uint(f[4 * c]), __float_as_uint(f[4 * c + 1]),
                             __float_as_uint(f[4 * c + 2]), __float_as_uint(f[4 * c + 3]));
            } else {
#pragma unroll
              for (int c = 0; c < 4; ++c)
                st_shared_v4(srow + (((h * 4 + c) ^ (lane & 7)) << 4), pack_bf16x2(f[8 * c], f[8 * c + 1]),
                             pack_bf16x2(f[8 * c + 2], f[8 * c + 3]), pack_bf16x2(f[8 * c + 4], f[8 * c + 5]),
                             pack_bf16x2(f[8 * c + 6], f[8 * c + 7]));
            }
          }
          fence_proxy_async_smem();  // generic-proxy writes -> visible to the TMA (async proxy)
          __syncwarp();
#ifdef B200STEP_EXPERIMENTS
          const bool skip_store = (p.flags & 0x100) != 0;  // timing experiment
#else
          constexpr bool skip_store = false;
#endif
          if (lane == 0 && row0 < p.M && !skip_store) {
            tma_store_2d(&tmC, sbuf_ptr, gcol0, row0);
            bulk_commit();
          }
        }
        if constexpr (kEpi == EPI_FWD) {
          // the row's 128 mask bits of this column half leave as one 16-byte store
#ifdef B200STEP_EXPERIMENTS
          const bool skip_mask = (p.flags & 0x400) != 0;  // timing experiment
#else
          constexpr bool skip_mask = false;
#endif
          if (p.mask_out != nullptr && row_ok && !skip_mask) {
            const long long wbase = (n0 >> 5) + chalf * 4;
            if (wbase + 4 <= p.ldmask)
              *reinterpret_cast<uint4*>(p.mask_out + (long long)row * p.ldmask + wbase) =
                  make_uint4(mrow[0], mrow[1], mrow[2], mrow[3]);
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {  // the accumulator stage may be overwritten: tell the (leader's) MMA thread
        if constexpr (kCtas == 2) mbar_arrive_cluster(&acc_empty[a], 0);
        else mbar_arrive(&acc_empty[a]);
      }
    }
    if constexpr (kEpi == EPI_DGRAD) {
      if (p.cs_partials != nullptr) {
        // one partial row per (CTA group, lane quarter); row stride = n_tiles * BLOCK_N
        const long long prow = ((long long)(unit / p.n_tiles) * kCtas + rank) * 4 + q;
        float* dst = p.cs_partials + prow * ((long long)p.n_tiles * BLOCK_N) +
                     (unit % p.n_tiles) * BLOCK_N + chalf * (BLOCK_N / 2) + lane;
#pragma unroll
        for (int k = 0; k < 4; ++k) dst[k * 32] = csacc[k];
      } else if (p.cs_atomic != nullptr) {
        // one coalesced red.add per 32 columns and warp: ~300 adds per address over the whole kernel
        const int c0 = (unit % p.n_tiles) * BLOCK_N + chalf * (BLOCK_N / 2) + lane;
#pragma unroll
        for (int k = 0; k < 4; ++k)
          if (c0 + k * 32 < p.N) atomicAdd(p.cs_atomic + c0 + k * 32, csacc[k]);
      }
    }
    if (lane == 0) bulk_wait_all();  // all output writes issued by this warp have completed
  }

  tc_fence_before();
  if constexpr (kCtas == 2) cluster_sync_all();  // neither CTA may leave while its peer can still touch it
  else __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    if constexpr (kCtas == 2) tmem_dealloc_pair(tmem_base, NUM_ACC * BLOCK_N);
    else tmem_dealloc(tmem_base, NUM_ACC * BLOCK_N);
  }
}

// ---- grouped wgrad: the weight gradients of ALL layers in one launch -----------------------------------------
// The per-layer wgrad launches each give a CTA pair exactly one work unit (one output tile x one slice of the
// batch), so prologue, pipeline fill, the red.add epilogue and the ~6 us fixed cost of a launch are paid per
// layer and nothing overlaps them.  Here the units of every layer form one list; a pair walks its share of it
// (unit, unit + npairs, ...), so the epilogue of one unit drains under the MMAs of the next and there is one
// launch for the whole wgrad phase.  Same tiles, same pipeline, same epilogue as gemm_kernel<MN, MN, WGRAD, 2>.
constexpr int WG_MAX_PROBLEMS = 8;

struct WgradProblem {
  int M, N;             // dW is M x N (= out_features x in_features)
  int k_blocks;         // ceil(batch / BLOCK_K)
  int m_tiles, n_tiles, splits, kb_per_split;
  int unit_begin;       // first unit of this problem in the global unit list
  float* out;
  long long ldo;
};

struct WgradGroup {
  CUtensorMap tmA[WG_MAX_PROBLEMS];  // dY^T (MN-major)
  CUtensorMap tmB[WG_MAX_PROBLEMS];  // X^T  (MN-major)
  WgradProblem prob[WG_MAX_PROBLEMS];
  int count;
  int total_units;
};

template <int kTf32>
__global__ void __launch_bounds__(GEMM_THREADS, 1)
wgrad_grouped_kernel(const __grid_constant__ WgradGroup g) {
  constexpr int ESIZE = kTf32 ? 4 : 2;
  constexpr int BLOCK_K = 128 / ESIZE;
  constexpr int UMMA_K = 32 / ESIZE;
  constexpr int EPR = 128 / ESIZE;
  constexpr int MN_BOX_BYTES = BLOCK_K * 128;
  constexpr uint32_t MN_SBO = kTf32 ? 512 : 1024;
  constexpr uint32_t MN_LAYOUT = kTf32 ? 1 : 2;
  constexpr uint32_t IDESC = make_idesc(kTf32 ? 2 : 1, 1, 1, BLOCK_M * 2, BLOCK_N);
  constexpr int B_ROWS = BLOCK_N / 2;
  constexpr int STAGE_BYTES = A_TILE_BYTES + B_ROWS * 128;
  constexpr int NUM_STAGES = PIPE_BYTES / STAGE_BYTES;
  constexpr int TILE_M = BLOCK_M * 2;

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + PIPE_BYTES + EPI_STAGING_BYTES);
  uint64_t* full_bar = bars;
  uint64_t* empty_bar = bars + NUM_STAGES;
  uint64_t* acc_full = bars + 2 * NUM_STAGES;
  uint64_t* acc_empty = acc_full + NUM_ACC;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_empty + NUM_ACC);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int rank = (int)cluster_ctarank();
  const int unit0 = blockIdx.x / 2, nunits = gridDim.x / 2;

  pdl_launch_dependents();
  if (warp == 0 && lane == 0) {
    for (int j = 0; j < g.count; ++j) {
      tma_prefetch_desc(&g.tmA[j]);
      tma_prefetch_desc(&g.tmB[j]);
    }
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < NUM_STAGES; ++i) {
      mbar_init(&full_bar[i], 1);
      mbar_init(&empty_bar[i], 1);
    }
    for (int i = 0; i < NUM_ACC; ++i) {
      mbar_init(&acc_full[i], 1);
      mbar_init(&acc_empty[i], EPI_WARPS * 2);
    }
    fence_mbar_init();
  }
  if (warp == 2) tmem_alloc_pair(tmem_slot, NUM_ACC * BLOCK_N);
  tc_fence_before();
  cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_wait();

  // unit -> (problem, tile, split): problems are few, a linear scan is cheapest
  auto locate = [&](int w, int& j, int& m0, int& n0, int& kb0, int& kb1) {
    j = 0;
    while (j + 1 < g.count && w >= g.prob[j + 1].unit_begin) ++j;
    const WgradProblem& q = g.prob[j];
    const int lw = w - q.unit_begin;
    const int tiles = q.m_tiles * q.n_tiles;
    const int tile = lw % tiles, split = lw / tiles;
    m0 = (tile / q.n_tiles) * TILE_M;
    n0 = (tile % q.n_tiles) * BLOCK_N;
    kb0 = split * q.kb_per_split;
    kb1 = min(kb0 + q.kb_per_split, q.k_blocks);
  };

  if (warp == 0) {
    if (lane == 0) {  // ---- TMA producer
      int s = 0;
      uint32_t ph = 0;
      for (int w = unit0; w < g.total_units; w += nunits) {
        int j, m0, n0, kb0, kb1;
        locate(w, j, m0, n0, kb0, kb1);
        m0 += rank * BLOCK_M;
        n0 += rank * B_ROWS;
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(&empty_bar[s], ph ^ 1);
          uint8_t* sa = smem + s * STAGE_BYTES;
          uint8_t* sb = sa + A_TILE_BYTES;
          if (rank == 0) mbar_expect_tx(&full_bar[s], 2 * STAGE_BYTES);
          const uint32_t fb = smem_u32(&full_bar[s]) & 0xFEFFFFFFu;
#pragma unroll
          for (int i = 0; i < BLOCK_M / EPR; ++i)
            tma_load_2d_pair(sa + i * MN_BOX_BYTES, &g.tmA[j], fb, m0 + i * EPR, kb * BLOCK_K);
#pragma unroll
          for (int i = 0; i < B_ROWS / EPR; ++i)
            tma_load_2d_pair(sb + i * MN_BOX_BYTES, &g.tmB[j], fb, n0 + i * EPR, kb * BLOCK_K);
          if (++s == NUM_STAGES) { s = 0; ph ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0 && rank == 0) {  // ---- MMA issuer
      int s = 0;
      uint32_t ph = 0;
      int it = 0;
      for (int w = unit0; w < g.total_units; w += nunits, ++it) {
        int j, m0, n0, kb0, kb1;
        locate(w, j, m0, n0, kb0, kb1);
        const int a = it & 1;
        const uint32_t aph = (it >> 1) & 1;
        mbar_wait(&acc_empty[a], aph ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + a * BLOCK_N;
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(&full_bar[s], ph);
          tc_fence_after();
          const uint32_t sa = smem_u32(smem + s * STAGE_BYTES);
          const uint32_t sb = sa + A_TILE_BYTES;
#pragma unroll
          for (int k = 0; k < BLOCK_K / UMMA_K; ++k) {
            const uint64_t ad = make_smem_desc(sa + k * UMMA_K * 128, MN_BOX_BYTES, MN_SBO, MN_LAYOUT);
            const uint64_t bd = make_smem_desc(sb + k * UMMA_K * 128, MN_BOX_BYTES, MN_SBO, MN_LAYOUT);
            umma_ss_pair<kTf32>(d_tmem, ad, bd, IDESC, (kb > kb0 || k > 0) ? 1u : 0u);
          }
          umma_commit_pair(&empty_bar[s], 3);
          if (++s == NUM_STAGES) { s = 0; ph ^= 1; }
        }
        umma_commit_pair(&acc_full[a], 3);
      }
    }
  } else if (warp >= 4) {  // ---- epilogue: TMEM -> registers -> red.add onto the (zeroed) gradient
    const int q4 = warp & 3;
    const int chalf = (warp - 4) >> 2;
    int it = 0;
    for (int w = unit0; w < g.total_units; w += nunits, ++it) {
      int j, m0, n0, kb0, kb1;
      locate(w, j, m0, n0, kb0, kb1);
      const WgradProblem& q = g.prob[j];
      const int a = it & 1;
      const uint32_t aph = (it >> 1) & 1;
      const int row = m0 + rank * BLOCK_M + q4 * 32 + lane;
      const bool row_ok = row < q.M;
      mbar_wait(&acc_full[a], aph);
      tc_fence_after();
      const uint32_t taddr = tmem_base + ((uint32_t)(q4 * 32) << 16) + a * BLOCK_N;
#pragma unroll 1
      for (int c = chalf * (BLOCK_N / 64); c < (chalf + 1) * (BLOCK_N / 64); ++c) {
        const int col0 = n0 + c * 32;
        if (col0 >= q.N) break;  // warp-uniform
        uint32_t v[32];
        tmem_ld32(taddr + c * 32, v);
        tmem_ld_wait();
        if (!row_ok) continue;
        float* o = q.out + (long long)row * q.ldo + col0;
#pragma unroll
        for (int i = 0; i < 32; i += 4)
          if (col0 + i < q.N)
            red_add_v4(o + i, __uint_as_float(v[i]), __uint_as_float(v[i + 1]), __uint_as_float(v[i + 2]),
                       __uint_as_float(v[i + 3]));
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_cluster(&acc_empty[a], 0);
    }
  }

  tc_fence_before();
  cluster_sync_all();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc_pair(tmem_base, NUM_ACC * BLOCK_N);
  }
}

// CTAs per work item: 2 = CTA pairs (default), 1 = single-CTA kernel (B200STEP_GEMM_CTAS=1, for A/B runs)
static int gemm_ctas() {
  static int v = 0;
  if (v == 0) {
    const char* e = getenv("B200STEP_GEMM_CTAS");
    v = (e && atoi(e) == 1) ? 1 : 2;
  }
  return v;
}

// UMMA N (and staged B rows) of a launch: 64 / 128 when the whole output is that narrow, else the full 256-wide tile
static int pick_n_eff(int64_t N, int n_tiles) {
  static const bool narrow = getenv("B200STEP_NARROW_MMA") == nullptr || atoi(getenv("B200STEP_NARROW_MMA")) != 0;
  if (!narrow || n_tiles != 1) return BLOCK_N;
  return N <= 64 ? 64 : (N <= 128 ? 128 : BLOCK_N);
}

// Launch on min(work items, available units) units, rounded down to a multiple of `unit_multiple`.
template <int kTf32, bool kAMN, bool kBMN, int kEpi, int kCtas>
static int launch_gemm_n(const CUtensorMap& ta, const CUtensorMap& tb, const CUtensorMap& tc, const GemmParams& p,
                         cudaStream_t st, int unit_multiple, int* units_out) {
  auto kern = gemm_kernel<kTf32, kAMN, kBMN, kEpi, kCtas>;
  static bool configured = false;  // per instantiation
  if (!configured) {
    B200_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, GEMM_SMEM_BYTES));
    configured = true;
  }
  const int total = p.m_tiles * p.n_tiles * p.splits;
  int units = sm_count() / kCtas;
  if (units > total) units = total;
  if (unit_multiple > 1) units = units / unit_multiple * unit_multiple;
  B200_REQUIRE(units >= 1, "GEMM: no schedulable work units (tiles=%d, multiple=%d)", total, unit_multiple);
  if (units_out) *units_out = units;
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(units * kCtas);
  cfg.blockDim = dim3(GEMM_THREADS);
  cfg.dynamicSmemBytes = GEMM_SMEM_BYTES;
  cfg.stream = st;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = kCtas;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[1].val.programmaticStreamSerializationAllowed = pdl_enabled() ? 1 : 0;
  cfg.attrs = attr;
  cfg.numAttrs = 2;
  B200_CHECK_CUDA(cudaLaunchKernelEx(&cfg, kern, ta, tb, tc, p));
  return 0;
}

// Timing experiments (results are wrong), compiled only with -DB200STEP_EXPERIMENTS (build.py --experiments), never
// into the shipped library: bit 8 = do not issue the output TMA stores, bit 9 = skip the epilogue math and staging,
// bit 10 = skip the ReLU bit-mask store, bit 11 = skip the MMAs (operand streaming only)
static int gemm_debug_flags() {
#ifdef B200STEP_EXPERIMENTS
  static int v = -1;
  if (v < 0) v = getenv("B200STEP_GEMM_DEBUG") ? (atoi(getenv("B200STEP_GEMM_DEBUG")) << 8) : 0;
  return v;
#else
  return 0;
#endif
}

template <int kTf32, bool kAMN, bool kBMN, int kEpi>
static int launch_gemm(const CUtensorMap& ta, const CUtensorMap& tb, const CUtensorMap& tc, const GemmParams& p_in,
                       cudaStream_t st, int unit_multiple = 1, int* units_out = nullptr) {
  GemmParams p = p_in;
  p.flags |= gemm_debug_flags();
  if (p.n_eff == 0) p.n_eff = pick_n_eff(p.N, p.n_tiles);
  return gemm_ctas() == 2
             ? launch_gemm_n<kTf32, kAMN, kBMN, kEpi, 2>(ta, tb, tc, p, st, unit_multiple, units_out)
             : launch_gemm_n<kTf32, kAMN, kBMN, kEpi, 1>(ta, tb, tc, p, st, unit_multiple, units_out);
}

}  // namespace b200

using namespace b200;

extern "C" int b200_linear_fwd(const void* x, int64_t ldx, const void* w, int64_t ldw, const float* bias,
                               void* y, int64_t ldy, int64_t M, int64_t N, int64_t K, int dtype, int flags,
                               float dropout_p, uint64_t seed, const int32_t* step, int layer_id,
                               int64_t dropout_row0, uint32_t* relu_mask, int64_t ldmask, b200_stream_t stream) {
  if (int r = check_dims(M, N, K, dtype)) return r;
  B200_REQUIRE(relu_mask == nullptr || (ldmask * 32 >= N && ldmask % 4 == 0 &&
                                        (reinterpret_cast<uintptr_t>(relu_mask) & 15) == 0),
               "relu_mask needs a 16-byte aligned base and a row stride (%lld words) that is a multiple of 4 and covers N=%lld",
               (long long)ldmask, (long long)N);
  B200_REQUIRE(dropout_p >= 0.f && dropout_p < 1.f, "dropout_p=%f out of [0,1)", dropout_p);
  const int tf32 = dtype == B200_F32;
  const int bk = tf32 ? 32 : 64;
  CUtensorMap ta, tb;
  const int n_eff = pick_n_eff(N, (int)((N + BLOCK_N - 1) / BLOCK_N));
  if (int r = make_tmap(&ta, tf32, x, K, M, ldx, bk, BLOCK_M)) return r;
  if (int r = make_tmap(&tb, tf32, w, K, N, ldw, bk, n_eff / gemm_ctas())) return r;  // a pair splits the B tile
  CUtensorMap tc;  // output slabs: one 128-byte row x 32 rows per TMA store
  if (int r = make_tmap(&tc, tf32, y, N, M, ldy, bk, 32, false, true)) return r;
  GemmParams p{};
  p.M = (int)M; p.N = (int)N;
  const int tile_m = BLOCK_M * gemm_ctas();
  p.k_blocks = (int)((K + bk - 1) / bk);
  p.m_tiles = (int)((M + tile_m - 1) / tile_m);
  p.n_tiles = (int)((N + BLOCK_N - 1) / BLOCK_N);
  p.splits = 1; p.kb_per_split = p.k_blocks;
  p.out = y; p.ldo = ldy; p.bias = bias; p.flags = flags; p.dscale = 1.f; p.n_eff = n_eff;
  p.dropout_p = dropout_p; p.seed = seed; p.step = step; p.layer_id = layer_id;
  p.mask_out = relu_mask; p.ldmask = ldmask; p.row0 = dropout_row0;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  return tf32 ? launch_gemm<1, false, false, EPI_FWD>(ta, tb, tc, p, st)
              : launch_gemm<0, false, false, EPI_FWD>(ta, tb, tc, p, st);
}

// Small-M convolutions (layer3/4 of a ResNet on 32x32 images: 1024-4096 output pixels) have 8-16 output tiles for 74 CTA
// pairs while their contraction is 36-72 k-blocks long.  With a caller-provided fp32 workspace the contraction is split
// across CTA pairs: every split adds its partial tile onto the zeroed workspace (red.global.add.v4.f32, the wgrad
// epilogue), and one cast pass writes the bf16 result (fp32 outputs are accumulated in place, no workspace needed).
// Not used with B200_FLAG_DETERMINISTIC (the sum order of the splits is not fixed).
static int conv_splits(const GemmParams& p, int flags) {
  if (flags & B200_FLAG_DETERMINISTIC) return 1;
  const int tiles = p.m_tiles * p.n_tiles, units = sm_count() / gemm_ctas();
  if (tiles * 2 > units || p.k_blocks < 8) return 1;
  int splits = units / tiles;
  if (splits > p.k_blocks / 4) splits = p.k_blocks / 4;       // at least 4 k-blocks per split
  return splits < 2 ? 1 : splits;
}
extern "C" int b200_cast_bf16(const float* src, void* dst, int64_t n, b200_stream_t stream);

// Convolution forward as an implicit GEMM: y[B*Ho*Wo, Cout] = unfold(x)[.., kh*kw*C] . w[Cout, kh*kw*C]^T without the
// unfold ever being stored — the A tiles come from an im2col-mode tensor map over the NHWC input.
extern "C" int b200_conv_fwd(const void* x, const void* w, int64_t ldw, void* y, int64_t ldy, int64_t B, int H, int W,
                             int C, int kh, int kw, int stride, int pad, int64_t Cout, int dtype, int flags,
                             float* ws, int64_t ws_floats, b200_stream_t stream) {
  B200_REQUIRE(dtype == B200_F32 || dtype == B200_BF16, "unknown dtype code %d", dtype);
  const int tf32 = dtype == B200_F32;
  const int bk = tf32 ? 32 : 64;
  B200_REQUIRE(B > 0 && H > 0 && W > 0 && kh > 0 && kw > 0 && kh * kw < 256 && stride > 0 && pad >= 0 && C > 0 &&
                   C % bk == 0, "conv_fwd: C=%d must be a multiple of %d (one 128-byte k-block per tap slice)", C, bk);
  const int Ho = (H + 2 * pad - kh) / stride + 1, Wo = (W + 2 * pad - kw) / stride + 1;
  B200_REQUIRE(Ho > 0 && Wo > 0, "conv_fwd: empty output");
  const int64_t M = B * Ho * Wo, K = (int64_t)kh * kw * C;
  if (int r = check_dims(M, Cout, K, dtype)) return r;
  B200_REQUIRE(ldw >= K, "conv_fwd: filter row stride %lld < kh*kw*C = %lld", (long long)ldw, (long long)K);
  CUtensorMap ta, tb, tc;
  const int n_eff = pick_n_eff(Cout, (int)((Cout + BLOCK_N - 1) / BLOCK_N));
  if (int r = make_tmap_im2col(&ta, tf32, x, B, H, W, C, kh, kw, stride, pad, bk, BLOCK_M)) return r;
  if (int r = make_tmap(&tb, tf32, w, K, Cout, ldw, bk, n_eff / gemm_ctas())) return r;
  if (int r = make_tmap(&tc, tf32, y, Cout, M, ldy, bk, 32, false, true)) return r;
  GemmParams p{};
  p.M = (int)M; p.N = (int)Cout;
  const int tile_m = BLOCK_M * gemm_ctas();
  p.k_blocks = (int)(K / bk);
  p.m_tiles = (int)((M + tile_m - 1) / tile_m);
  p.n_tiles = (int)((Cout + BLOCK_N - 1) / BLOCK_N);
  p.splits = 1; p.kb_per_split = p.k_blocks;
  p.out = y; p.ldo = ldy; p.flags = flags & (B200_FLAG_RELU | B200_FLAG_REVERSE_M); p.dscale = 1.f;
  p.cv_on = 1; p.cv_Ho = Ho; p.cv_Wo = Wo; p.cv_stride = stride; p.cv_pad = pad; p.cv_kw = kw; p.cv_cpb = C / bk;
  p.n_eff = n_eff;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int splits = (flags & B200_FLAG_RELU) ? 1 : conv_splits(p, flags);
  const bool ws_ok = tf32 ? (ldy % 4 == 0) : (ws != nullptr && ws_floats >= M * Cout && ldy == Cout &&
                                              (reinterpret_cast<uintptr_t>(ws) & 15) == 0);
  if (splits > 1 && ws_ok) {
    float* acc = tf32 ? static_cast<float*>(y) : ws;
    const int64_t ldacc = tf32 ? ldy : Cout;
    p.kb_per_split = (p.k_blocks + splits - 1) / splits;
    p.splits = (p.k_blocks + p.kb_per_split - 1) / p.kb_per_split;
    p.out = acc; p.ldo = ldacc; p.flags = 0;
    B200_CHECK_CUDA(cudaMemset2DAsync(acc, ldacc * sizeof(float), 0, Cout * sizeof(float), M, st));
    if (int r = tf32 ? launch_gemm<1, false, false, EPI_WGRAD>(ta, tb, ta /*unused*/, p, st)
                     : launch_gemm<0, false, false, EPI_WGRAD>(ta, tb, ta /*unused*/, p, st))
      return r;
    return tf32 ? 0 : b200_cast_bf16(ws, y, M * Cout, stream);
  }
  return tf32 ? launch_gemm<1, false, false, EPI_FWD>(ta, tb, tc, p, st)
              : launch_gemm<0, false, false, EPI_FWD>(ta, tb, tc, p, st);
}

extern "C" int b200_linear_dgrad(const void* dy, int64_t lddy, const void* w, int64_t ldw, const void* act,
                                 int64_t ldact, void* dx, int64_t lddx, int64_t M, int64_t N, int64_t K,
                                 int dtype, int flags, float dscale, const uint32_t* relu_mask, int64_t ldmask,
                                 float* dbias, float* scratch, int64_t scratch_floats, b200_stream_t stream) {
  if (int r = check_dims(M, N, K, dtype)) return r;
  B200_REQUIRE(relu_mask == nullptr || (ldmask * 32 >= K && ldmask % 4 == 0 &&
                                        (reinterpret_cast<uintptr_t>(relu_mask) & 15) == 0),
               "relu_mask needs a 16-byte aligned base and a row stride (%lld words) that is a multiple of 4 and covers K=%lld",
               (long long)ldmask, (long long)K);
  B200_REQUIRE(dbias == nullptr || scratch == nullptr || scratch_floats >= b200_dgrad_scratch_floats(M, K),
               "dgrad bias-gradient scratch too small: need %lld floats", (long long)b200_dgrad_scratch_floats(M, K));
  const int tf32 = dtype == B200_F32;
  const int bk = tf32 ? 32 : 64;   // contraction (over N) elements per stage
  const int epr = tf32 ? 32 : 64;  // elements per 128-byte row
  CUtensorMap ta, tb;
  // A = dY [M, N], contraction dim N contiguous (K-major)
  if (int r = make_tmap(&ta, tf32, dy, N, M, lddy, bk, BLOCK_M)) return r;
  // B(n'=k, k'=n) = W[n, k]: output dim contiguous (MN-major); box = 128 B of k x bk rows of n
  if (int r = make_tmap(&tb, tf32, w, K, N, ldw, epr, bk, true)) return r;
  CUtensorMap tc;
  if (int r = make_tmap(&tc, tf32, dx, K, M, lddx, epr, 32, false, true)) return r;
  GemmParams p{};
  p.M = (int)M; p.N = (int)K;
  const int tile_m = BLOCK_M * gemm_ctas();
  p.k_blocks = (int)((N + bk - 1) / bk);
  p.m_tiles = (int)((M + tile_m - 1) / tile_m);
  p.n_tiles = (int)((K + BLOCK_N - 1) / BLOCK_N);
  p.splits = 1; p.kb_per_split = p.k_blocks;
  p.out = dx; p.ldo = lddx; p.act = act; p.ldact = ldact; p.flags = flags; p.dscale = dscale;
  p.mask_in = relu_mask; p.ldmask = ldmask;
  p.cs_partials = dbias ? scratch : nullptr;
  p.cs_atomic = (dbias && !scratch) ? dbias : nullptr;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  // with bias-gradient accumulation every work unit must keep one n_tile: units = multiple of n_tiles
  int units = 0;
  if (int r = tf32 ? launch_gemm<1, false, true, EPI_DGRAD>(ta, tb, tc, p, st, dbias ? p.n_tiles : 1, &units)
                   : launch_gemm<0, false, true, EPI_DGRAD>(ta, tb, tc, p, st, dbias ? p.n_tiles : 1, &units))
    return r;
  if (dbias && scratch)
    return launch_reduce_partials(scratch, (units / p.n_tiles) * gemm_ctas() * 4, p.n_tiles * BLOCK_N, (int)K,
                                  dbias, st);
  return 0;
}

// one partial row per (CTA group, lane quarter), row stride n_tiles*BLOCK_N: at most sm_count * 4 * BLOCK_N floats
extern "C" int64_t b200_dgrad_scratch_floats(int64_t M, int64_t K) {
  (void)M;
  (void)K;
  return (int64_t)sm_count_physical() * 4 * BLOCK_N;
}

// Layer-0 dgrad fused with the embedding backward: dtables[t][idx[m,t]][:] += (dy[M,N] . w[N, :num_tables*dim])[m, t*dim : (t+1)*dim]
extern "C" int b200_linear_dgrad_emb(const void* dy, int64_t lddy, const void* w, int64_t ldw, const int32_t* idx,
                                     int64_t ld_idx, float* dtables, int64_t M, int64_t N, int num_tables,
                                     int64_t table_rows, int embed_dim, int dtype, b200_stream_t stream) {
  const int64_t K = (int64_t)num_tables * embed_dim;
  if (int r = check_dims(M, N, K, dtype)) return r;
  B200_REQUIRE(embed_dim >= 4 && (embed_dim & (embed_dim - 1)) == 0,
               "b200_linear_dgrad_emb needs a power-of-two embedding width >= 4 (got %d); use b200_linear_dgrad + "
               "b200_emb_bwd", embed_dim);
  B200_REQUIRE(idx != nullptr && dtables != nullptr && ld_idx >= num_tables, "idx / dtables missing or ld_idx too small");
  B200_REQUIRE((reinterpret_cast<uintptr_t>(dtables) & 15) == 0, "dtables must be 16-byte aligned");
  const int tf32 = dtype == B200_F32;
  const int bk = tf32 ? 32 : 64;
  const int epr = tf32 ? 32 : 64;
  CUtensorMap ta, tb;
  if (int r = make_tmap(&ta, tf32, dy, N, M, lddy, bk, BLOCK_M)) return r;
  // only the embedding columns of W are read: the tensor map ends at K, boxes past it are zero-filled
  if (int r = make_tmap(&tb, tf32, w, K, N, ldw, epr, bk, true)) return r;
  GemmParams p{};
  p.M = (int)M; p.N = (int)K;
  const int tile_m = BLOCK_M * gemm_ctas();
  p.k_blocks = (int)((N + bk - 1) / bk);
  p.m_tiles = (int)((M + tile_m - 1) / tile_m);
  p.n_tiles = (int)((K + BLOCK_N - 1) / BLOCK_N);
  p.splits = 1; p.kb_per_split = p.k_blocks;
  p.out = dtables; p.dscale = 1.f;
  p.emb_idx = idx; p.ld_idx = ld_idx; p.emb_rows = table_rows;
  p.emb_shift = 0;
  while ((1 << p.emb_shift) < embed_dim) ++p.emb_shift;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  return tf32 ? launch_gemm<1, false, true, EPI_EMB>(ta, tb, ta /*unused*/, p, st)
              : launch_gemm<0, false, true, EPI_EMB>(ta, tb, ta /*unused*/, p, st);
}

extern "C" int b200_linear_wgrad(const void* dy, int64_t lddy, const void* x, int64_t ldx, float* dw,
                                 int64_t lddw, int64_t M, int64_t N, int64_t K, int dtype, int flags,
                                 b200_stream_t stream) {
  if (int r = check_dims(M, N, K, dtype)) return r;
  B200_REQUIRE(lddw % 4 == 0 && (reinterpret_cast<uintptr_t>(dw) & 15) == 0,
               "wgrad output must be 16-byte aligned with lddw %% 4 == 0");
  const int tf32 = dtype == B200_F32;
  const int bk = tf32 ? 32 : 64;   // batch rows per stage
  const int epr = tf32 ? 32 : 64;
  CUtensorMap ta, tb;
  // A(m'=n, k'=b) = dY[b, n]; B(n'=k, k'=b) = X[b, k]: both MN-major
  if (int r = make_tmap(&ta, tf32, dy, N, M, lddy, epr, bk, true)) return r;
  if (int r = make_tmap(&tb, tf32, x, K, M, ldx, epr, bk, true)) return r;
  GemmParams p{};
  p.M = (int)N; p.N = (int)K;
  const int tile_m = BLOCK_M * gemm_ctas();
  p.k_blocks = (int)((M + bk - 1) / bk);
  p.m_tiles = (int)((N + tile_m - 1) / tile_m);
  p.n_tiles = (int)((K + BLOCK_N - 1) / BLOCK_N);
  const int tiles = p.m_tiles * p.n_tiles;
  // two batch slices per CTA (pair): the epilogue of the first overlaps the MMAs of the second
  static const int wgrad_waves = getenv("B200STEP_WGRAD_WAVES") ? atoi(getenv("B200STEP_WGRAD_WAVES")) : 1;
  int splits = wgrad_waves * (sm_count() / gemm_ctas()) / tiles;
  if (splits < 1 || (flags & B200_FLAG_DETERMINISTIC)) splits = 1;   // one CTA pair sums a tile's whole batch in order
  if (splits > p.k_blocks) splits = p.k_blocks;
  p.kb_per_split = (p.k_blocks + splits - 1) / splits;
  p.splits = (p.k_blocks + p.kb_per_split - 1) / p.kb_per_split;
  p.out = dw; p.ldo = lddw; p.flags = flags; p.dscale = 1.f;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (p.splits > 1 && !(flags & B200_FLAG_ACCUMULATE)) {
    // split-K partial tiles are summed with red.add: start from zero
    B200_CHECK_CUDA(cudaMemset2DAsync(dw, lddw * sizeof(float), 0, K * sizeof(float), N, st));
  }
  return tf32 ? launch_gemm<1, true, true, EPI_WGRAD>(ta, tb, ta /*unused*/, p, st)
              : launch_gemm<0, true, true, EPI_WGRAD>(ta, tb, ta /*unused*/, p, st);
}

// Input gradient of a STRIDE-1 convolution as an implicit GEMM: dx = conv2d(dy, flipped filter) with padding k-1-pad.
// A tiles: im2col-mode TMA over dy [B, Ho, Wo, Cout]; B tiles: the [co][ci] block of Wg at the flipped tap, read as an
// MN-major operand from the SAME filter matrix the forward pass uses (no transposed copy); dx is written once — the
// 9x larger dcol matrix and the col2im pass of the explicit path do not exist.
extern "C" int b200_conv_dgrad(const void* dy, const void* w, int64_t ldw, void* dx, int64_t lddx, int64_t B, int H, int W,
                               int C, int kh, int kw, int pad, int64_t Cout, int dtype, int flags, float* ws,
                               int64_t ws_floats, b200_stream_t stream) {
  B200_REQUIRE(dtype == B200_F32 || dtype == B200_BF16, "unknown dtype code %d", dtype);
  const int tf32 = dtype == B200_F32;
  const int bk = tf32 ? 32 : 64, epr = bk;
  B200_REQUIRE(B > 0 && H > 0 && W > 0 && kh > 0 && kw > 0 && kh * kw < 256 && pad >= 0 && pad < kh && pad < kw &&
                   C > 0 && C % 8 == 0 && Cout > 0 && Cout % bk == 0,
               "conv_dgrad: Cout=%lld must be a multiple of %d, C=%d of 8, padding < kernel", (long long)Cout, bk, C);
  const int Ho = H + 2 * pad - kh + 1, Wo = W + 2 * pad - kw + 1;      // stride 1
  B200_REQUIRE(Ho > 0 && Wo > 0, "conv_dgrad: empty output");
  const int64_t M = B * H * W, K = (int64_t)kh * kw * C, Kc = (int64_t)kh * kw * Cout;
  if (int r = check_dims(M, Cout, C, dtype)) return r;
  B200_REQUIRE(ldw >= K, "conv_dgrad: filter row stride %lld < kh*kw*C = %lld", (long long)ldw, (long long)K);
  CUtensorMap ta, tb, tc;
  // "full" correlation of dy: a kh x kw window over dy padded by k-1-pad gives back H x W positions
  if (int r = make_tmap_im2col(&ta, tf32, dy, B, Ho, Wo, (int)Cout, kh, kw, 1, kh - 1 - pad, bk, BLOCK_M)) return r;
  if (int r = make_tmap(&tb, tf32, w, K, Cout, ldw, epr, bk, true)) return r;
  if (int r = make_tmap(&tc, tf32, dx, C, M, lddx, epr, 32, false, true)) return r;
  B200_REQUIRE(kh == kw, "conv_dgrad: square filters only");
  GemmParams p{};
  p.M = (int)M; p.N = C;
  const int tile_m = BLOCK_M * gemm_ctas();
  p.k_blocks = (int)(Kc / bk);
  p.m_tiles = (int)((M + tile_m - 1) / tile_m);
  p.n_tiles = (C + BLOCK_N - 1) / BLOCK_N;
  p.splits = 1; p.kb_per_split = p.k_blocks;
  p.out = dx; p.ldo = lddx; p.flags = 0; p.dscale = 1.f;
  p.cv_on = 2; p.cv_Ho = H; p.cv_Wo = W; p.cv_stride = 1; p.cv_pad = kh - 1 - pad; p.cv_kw = kw;
  p.cv_cpb = (int)(Cout / bk); p.cv_C = C; p.cv_K = kh * kw;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int splits = conv_splits(p, flags);
  const bool ws_ok = tf32 ? (lddx % 4 == 0) : (ws != nullptr && ws_floats >= M * C && lddx == C &&
                                               (reinterpret_cast<uintptr_t>(ws) & 15) == 0);
  if (splits > 1 && ws_ok) {
    float* acc = tf32 ? static_cast<float*>(dx) : ws;
    const int64_t ldacc = tf32 ? lddx : C;
    p.kb_per_split = (p.k_blocks + splits - 1) / splits;
    p.splits = (p.k_blocks + p.kb_per_split - 1) / p.kb_per_split;
    p.out = acc; p.ldo = ldacc;
    B200_CHECK_CUDA(cudaMemset2DAsync(acc, ldacc * sizeof(float), 0, C * sizeof(float), M, st));
    if (int r = tf32 ? launch_gemm<1, false, true, EPI_WGRAD>(ta, tb, ta /*unused*/, p, st)
                     : launch_gemm<0, false, true, EPI_WGRAD>(ta, tb, ta /*unused*/, p, st))
      return r;
    return tf32 ? 0 : b200_cast_bf16(ws, dx, M * C, stream);
  }
  return tf32 ? launch_gemm<1, false, true, EPI_DGRAD>(ta, tb, tc, p, st)
              : launch_gemm<0, false, true, EPI_DGRAD>(ta, tb, tc, p, st);
}

// Filter gradient of a convolution as an implicit GEMM: dw[Cout, kh*kw*C] (+)= dy[B*Ho*Wo, Cout]^T . unfold(x), with the
// unfold read by im2col-mode TMA from the NHWC input (the col matrix of the forward pass is not needed either).
extern "C" int b200_conv_wgrad(const void* dy, int64_t lddy, const void* x, float* dw, int64_t lddw, int64_t B, int H,
                               int W, int C, int kh, int kw, int stride, int pad, int64_t Cout, int dtype, int flags,
                               b200_stream_t stream) {
  B200_REQUIRE(dtype == B200_F32 || dtype == B200_BF16, "unknown dtype code %d", dtype);
  const int tf32 = dtype == B200_F32;
  const int bk = tf32 ? 32 : 64;   // output pixels per stage
  const int epr = tf32 ? 32 : 64;  // filter columns per 128-byte group
  B200_REQUIRE(B > 0 && H > 0 && W > 0 && kh > 0 && kw > 0 && kh * kw < 256 && stride > 0 && pad >= 0 && C > 0 &&
                   C % epr == 0, "conv_wgrad: C=%d must be a multiple of %d", C, epr);
  const int Ho = (H + 2 * pad - kh) / stride + 1, Wo = (W + 2 * pad - kw) / stride + 1;
  B200_REQUIRE(Ho > 0 && Wo > 0, "conv_wgrad: empty output");
  const int64_t M = B * Ho * Wo, K = (int64_t)kh * kw * C, N = Cout;
  if (int r = check_dims(M, N, K, dtype)) return r;
  B200_REQUIRE(lddw >= K && lddw % 4 == 0 && (reinterpret_cast<uintptr_t>(dw) & 15) == 0,
               "conv_wgrad output must be 16-byte aligned with lddw %% 4 == 0 and lddw >= kh*kw*C");
  CUtensorMap ta, tb;
  if (int r = make_tmap(&ta, tf32, dy, N, M, lddy, epr, bk, true)) return r;
  if (int r = make_tmap_im2col(&tb, tf32, x, B, H, W, C, kh, kw, stride, pad, epr, bk, true)) return r;
  GemmParams p{};
  p.M = (int)N; p.N = (int)K;
  const int tile_m = BLOCK_M * gemm_ctas();
  p.k_blocks = (int)((M + bk - 1) / bk);
  p.m_tiles = (int)((N + tile_m - 1) / tile_m);
  p.n_tiles = (int)((K + BLOCK_N - 1) / BLOCK_N);
  const int tiles = p.m_tiles * p.n_tiles;
  int splits = (sm_count() / gemm_ctas()) / tiles;
  if (splits < 1 || (flags & B200_FLAG_DETERMINISTIC)) splits = 1;
  if (splits > p.k_blocks) splits = p.k_blocks;
  p.kb_per_split = (p.k_blocks + splits - 1) / splits;
  p.splits = (p.k_blocks + p.kb_per_split - 1) / p.kb_per_split;
  p.out = dw; p.ldo = lddw; p.flags = flags; p.dscale = 1.f;
  p.cv_on = 1; p.cv_Ho = Ho; p.cv_Wo = Wo; p.cv_stride = stride; p.cv_pad = pad; p.cv_kw = kw; p.cv_cpb = C / bk;
  p.cv_C = C; p.cv_K = (int)K;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (p.splits > 1 && !(flags & B200_FLAG_ACCUMULATE))
    B200_CHECK_CUDA(cudaMemset2DAsync(dw, lddw * sizeof(float), 0, K * sizeof(float), N, st));
  return tf32 ? launch_gemm<1, true, true, EPI_WGRAD>(ta, tb, ta /*unused*/, p, st)
              : launch_gemm<0, true, true, EPI_WGRAD>(ta, tb, ta /*unused*/, p, st);
}

// All layers' weight gradients in one launch: dw[j][N_j, K_j] += dy[j][M, N_j]^T . x[j][M, K_j]  (fp32, red.add: the
// outputs must hold zeros or the value to accumulate onto).
extern "C" int b200_linear_wgrad_grouped(int count, const void* const* dy, const int64_t* lddy, const void* const* x,
                                         const int64_t* ldx, float* const* dw, const int64_t* lddw, int64_t M,
                                         const int64_t* N, const int64_t* K, int dtype, b200_stream_t stream) {
  B200_REQUIRE(count >= 1 && count <= WG_MAX_PROBLEMS, "grouped wgrad takes 1..%d problems, got %d", WG_MAX_PROBLEMS,
               count);
  const int tf32 = dtype == B200_F32;
  const int bk = tf32 ? 32 : 64;
  const int epr = tf32 ? 32 : 64;
  static WgradGroup g;  // host staging copy (one caller thread per process, see b200step.h)
  g.count = count;
  const int k_blocks = (int)((M + bk - 1) / bk);
  int total_tiles = 0;
  for (int j = 0; j < count; ++j) {
    if (int r = check_dims(M, N[j], K[j], dtype)) return r;
    B200_REQUIRE(lddw[j] % 4 == 0 && (reinterpret_cast<uintptr_t>(dw[j]) & 15) == 0,
                 "wgrad output %d must be 16-byte aligned with lddw %% 4 == 0", j);
    if (int r = make_tmap(&g.tmA[j], tf32, dy[j], N[j], M, lddy[j], epr, bk, true)) return r;
    if (int r = make_tmap(&g.tmB[j], tf32, x[j], K[j], M, ldx[j], epr, bk, true)) return r;
    WgradProblem& q = g.prob[j];
    q.M = (int)N[j]; q.N = (int)K[j];
    q.k_blocks = k_blocks;
    q.m_tiles = (int)((N[j] + 2 * BLOCK_M - 1) / (2 * BLOCK_M));
    q.n_tiles = (int)((K[j] + BLOCK_N - 1) / BLOCK_N);
    q.out = dw[j]; q.ldo = lddw[j];
    total_tiles += q.m_tiles * q.n_tiles;
  }
  // every tile is cut into the same number of batch slices, sized so that each CTA pair gets ~2 units (measured on T0: 113 us at 2, 123 us at 4, 147 us at 8 units per pair)
  const int pairs = sm_count() / 2;
  static const int rounds = getenv("B200STEP_WGRAD_ROUNDS") ? atoi(getenv("B200STEP_WGRAD_ROUNDS")) : 2;
  int splits = rounds * pairs / total_tiles;
  if (splits < 1) splits = 1;
  if (splits > k_blocks) splits = k_blocks;
  const int kbps = (k_blocks + splits - 1) / splits;
  splits = (k_blocks + kbps - 1) / kbps;
  int units = 0;
  for (int j = 0; j < count; ++j) {
    WgradProblem& q = g.prob[j];
    q.splits = splits; q.kb_per_split = kbps;
    q.unit_begin = units;
    units += q.m_tiles * q.n_tiles * splits;
  }
  g.total_units = units;
  auto launch = [&](auto kern) -> int {
    static bool configured[2] = {false, false};
    if (!configured[tf32]) {
      B200_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, GEMM_SMEM_BYTES));
      configured[tf32] = true;
    }
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3((units < pairs ? units : pairs) * 2);
    cfg.blockDim = dim3(GEMM_THREADS);
    cfg.dynamicSmemBytes = GEMM_SMEM_BYTES;
    cfg.stream = static_cast<cudaStream_t>(stream);
    cudaLaunchAttribute attr[2];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[1].val.programmaticStreamSerializationAllowed = pdl_enabled() ? 1 : 0;
    cfg.attrs = attr; cfg.numAttrs = 2;
    B200_CHECK_CUDA(cudaLaunchKernelEx(&cfg, kern, g));
    return 0;
  };
  return tf32 ? launch(wgrad_grouped_kernel<1>) : launch(wgrad_grouped_kernel<0>);
}
