// gemm_ce.cu — K-gemm-ce: the output layer and the cross-entropy loss as ONE kernel (sm_100a only).
//
// Replaces, for the last nn.Linear of the model, `outputs = model(inputs)` + `loss = criterion(outputs, labels)` +
// the start of `loss.backward()` (task.py:309-311): logits = H·Wᵀ + b, loss = mean(logsumexp(z) − z[y]),
// dlogits = (softmax(z) − onehot(y))·grad_scale, db = Σ_rows dlogits.  The logits never travel to HBM: they are
// produced in TMEM, reduced there, and only dlogits (which the dgrad / wgrad GEMMs need) is written — for a
// 65536 x 1000 bf16 batch that removes one 131 MB write, one 131 MB read and a whole memory-bound kernel.
//
// Shape of the computation.  A softmax row spans up to 1024 classes = four 256-column accumulator tiles, more than
// one SM's TMEM can hold next to a second (double-buffer) accumulator.  So a CLUSTER owns a 256-row slab: CTA pairs
// (tcgen05 cta_group::2, as in gemm_tcgen05.cu: each CTA stages its 128 rows of A and half of the pair's 256-wide
// B tile, the leader issues 256x256 UMMAs) sit side by side along N, pair j computing classes [256j, 256j+256).
// Per tile, every epilogue thread owns one row of its CTA's 128 x 256 accumulator (tcgen05.ld 32x32b: lane = row):
//   pass 1  online softmax over the thread's 128 columns: z = acc + bias (rounded to the storage dtype, which is
//           what a separate loss kernel would have read back), running max m and Σ exp(z − m); the exponentials
//           are written back INTO the accumulator columns (tcgen05.st), each 32-column chunk with the running max it
//           was formed against, so pass 2 needs no second exp per element;
//   xchg    (m, Σ) of every (CTA, column half) go to the CTAs that hold the same rows — distributed shared memory
//           stores + one cluster-scope mbarrier per accumulator stage — and every thread combines the 2·pairs
//           partial pairs of its row into the row's max M and Z = Σ exp(z − M);
//   pass 2  dlogits = e · exp(m_chunk − M) · grad_scale / Z, minus grad_scale at the label; rounded, staged in
//           swizzled shared memory and written by TMA; column sums of what is stored accumulate in registers
//           (output-layer bias gradient); the row's loss M + ln Z − z[y] is summed per thread.
// Partials (bias gradient, loss) leave as one row per (cluster, CTA, lane quarter) and are summed in a fixed order by
// reduce_partials_kernel, so both are run-to-run deterministic.
//
// Roles (384 threads, as gemm_tcgen05.cu): warp0 = TMA producer, warp1 = MMA issuer, warp2 = TMEM allocator,
// warps4-11 = epilogue.
#include "gemm_common.cuh"

namespace b200 {

constexpr int CE_STAGES = 5;                                          // operand ring: 5 x 32 KiB
constexpr int CE_STAGE_BYTES = A_TILE_BYTES + (BLOCK_N / 2) * 128;
constexpr int CE_PIPE_BYTES = CE_STAGES * CE_STAGE_BYTES;
constexpr int CE_MAX_PAIRS = 4;                                       // CTA pairs along N: up to 1024 classes
constexpr int CE_STATS_BYTES = 2 * CE_MAX_PAIRS * 2 * BLOCK_M * 8;    // [stage][pair][column half][row] float2
constexpr int CE_SMEM_BYTES = CE_PIPE_BYTES + EPI_STAGING_BYTES + CE_STATS_BYTES + 1024 /*align*/ + 256 /*barriers*/;
constexpr int CE_LOSS_COLS = 32;                                      // loss partial columns behind the class columns
constexpr float L2E = 1.4426950408889634f;

struct CeParams {
  int M, N;               // batch rows, classes
  int k_blocks;
  int m_tiles;            // 256-row slabs
  int pairs;              // CTA pairs along N in one cluster (1, 2 or 4)
  const float* bias;
  const int64_t* labels;
  float* partials;        // [clusters * 8][pstride]
  int pstride;            // pairs * 256 + CE_LOSS_COLS
  float scale;            // 1 / (batch * world)
  float batch;            // rows of the whole batch (n_valid rescaling)
  long long ignore_index;
  const int32_t* n_valid;
  uint32_t* status;
};

// ---- helpers this kernel adds to common.cuh's set ---------------------------------------------------------------
__device__ __forceinline__ void tmem_st32(uint32_t taddr, const uint32_t (&v)[32]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
      "{%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,"
      "%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31,%32};" ::"r"(taddr),
      "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]), "r"(v[8]), "r"(v[9]),
      "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15]), "r"(v[16]), "r"(v[17]), "r"(v[18]),
      "r"(v[19]), "r"(v[20]), "r"(v[21]), "r"(v[22]), "r"(v[23]), "r"(v[24]), "r"(v[25]), "r"(v[26]), "r"(v[27]),
      "r"(v[28]), "r"(v[29]), "r"(v[30]), "r"(v[31])
      : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
// address of the same shared-memory location in CTA `cta` of the cluster
__device__ __forceinline__ uint32_t mapa_shared(uint32_t saddr, uint32_t cta) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(saddr), "r"(cta));
  return r;
}
__device__ __forceinline__ void st_cluster_v2f(uint32_t raddr, float a, float b) {
  asm volatile("st.shared::cluster.v2.f32 [%0], {%1, %2};" ::"r"(raddr), "f"(a), "f"(b) : "memory");
}
__device__ __forceinline__ float2 ld_shared_v2f(uint32_t saddr) {
  float2 r;
  asm volatile("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(r.x), "=f"(r.y) : "r"(saddr));
  return r;
}
__device__ __forceinline__ void fence_acq_rel_cluster() { asm volatile("fence.acq_rel.cluster;" ::: "memory"); }
__device__ __forceinline__ void mbar_arrive_release_cluster(uint32_t raddr) {
  asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(raddr) : "memory");
}
__device__ __forceinline__ void mbar_wait_acquire_cluster(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred P1;\n"
      "LAB_WAIT_CL:\n"
      "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 P1, [%0], %1;\n"
      "@P1 bra DONE_CL;\n"
      "bra LAB_WAIT_CL;\n"
      "DONE_CL:\n"
      "}\n" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}

template <int kTf32>
__global__ void __launch_bounds__(GEMM_THREADS, 1)
linear_ce_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                 const __grid_constant__ CUtensorMap tmC, const CeParams p) {
  constexpr int ESIZE = kTf32 ? 4 : 2;
  constexpr int BLOCK_K = 128 / ESIZE;
  constexpr int UMMA_K = 32 / ESIZE;
  constexpr int EPR = 128 / ESIZE;                     // elements per 128-byte row = columns per output slab
  constexpr uint32_t IDESC = make_idesc(kTf32 ? 2 : 1, 0, 0, BLOCK_M * 2, BLOCK_N);
  constexpr int B_ROWS = BLOCK_N / 2;
  constexpr int TILE_M = BLOCK_M * 2;
  constexpr int CHUNKS = BLOCK_N / 2 / 32;             // 32-column chunks per epilogue warp (4)
  constexpr int LDS_PER_GROUP = EPR / 32;              // chunks per output slab: 2 (bf16) / 1 (fp32)
  constexpr int GROUPS_PER_WARP = CHUNKS / LDS_PER_GROUP;

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* stats = smem + CE_PIPE_BYTES + EPI_STAGING_BYTES;
  uint64_t* bars = reinterpret_cast<uint64_t*>(stats + CE_STATS_BYTES);
  uint64_t* full_bar = bars;                       // [CE_STAGES]
  uint64_t* empty_bar = bars + CE_STAGES;          // [CE_STAGES]
  uint64_t* acc_full = bars + 2 * CE_STAGES;       // [NUM_ACC]
  uint64_t* acc_empty = acc_full + NUM_ACC;        // [NUM_ACC]
  uint64_t* stat_bar = acc_empty + NUM_ACC;        // [NUM_ACC]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(stat_bar + NUM_ACC);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t ctarank = cluster_ctarank();
  const int rank = (int)(ctarank & 1u);            // 0 = leader of the pair
  const int nj = (int)(ctarank >> 1);              // which 256-class slice this pair computes
  const uint32_t leader = ctarank & ~1u;
  const uint16_t pair_mask = (uint16_t)(3u << (2 * nj));
  const int cluster_ctas = 2 * p.pairs;
  const int unit = blockIdx.x / cluster_ctas, nunits = gridDim.x / cluster_ctas;
  const int n0 = nj * BLOCK_N;

  pdl_launch_dependents();
  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    tma_prefetch_desc(&tmC);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < CE_STAGES; ++i) {
      mbar_init(&full_bar[i], 1);
      mbar_init(&empty_bar[i], 1);
    }
    for (int i = 0; i < NUM_ACC; ++i) {
      mbar_init(&acc_full[i], 1);
      mbar_init(&acc_empty[i], EPI_WARPS * 2);
      mbar_init(&stat_bar[i], EPI_WARPS * p.pairs);   // one arrive per epilogue warp of every CTA that holds my rows
    }
    fence_mbar_init();
  }
  if (warp == 2) tmem_alloc_pair(tmem_slot, NUM_ACC * BLOCK_N);
  tc_fence_before();
  cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_wait();

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (lane == 0) {
      int s = 0;
      uint32_t ph = 0;
      for (int w = unit; w < p.m_tiles; w += nunits) {
        const int m0 = w * TILE_M + rank * BLOCK_M;
        const int nb0 = n0 + rank * B_ROWS;
        for (int kb = 0; kb < p.k_blocks; ++kb) {
          mbar_wait(&empty_bar[s], ph ^ 1);
          uint8_t* sa = smem + s * CE_STAGE_BYTES;
          uint8_t* sb = sa + A_TILE_BYTES;
          if (rank == 0) mbar_expect_tx(&full_bar[s], 2 * CE_STAGE_BYTES);
          const uint32_t fb = smem_u32(&full_bar[s]) & 0xFEFFFFFFu;  // same offset in the pair's leader CTA
          tma_load_2d_pair(sa, &tmA, fb, kb * BLOCK_K, m0);
          tma_load_2d_pair(sb, &tmB, fb, kb * BLOCK_K, nb0);
          if (++s == CE_STAGES) { s = 0; ph ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer (one thread of the pair's leader) =====================
    if (lane == 0 && rank == 0) {
      int s = 0;
      uint32_t ph = 0;
      int it = 0;
      for (int w = unit; w < p.m_tiles; w += nunits, ++it) {
        const int a = it & 1;
        const uint32_t aph = (it >> 1) & 1;
        mbar_wait(&acc_empty[a], aph ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + a * BLOCK_N;
        for (int kb = 0; kb < p.k_blocks; ++kb) {
          mbar_wait(&full_bar[s], ph);
          tc_fence_after();
          const uint32_t sa = smem_u32(smem + s * CE_STAGE_BYTES);
          const uint32_t sb = sa + A_TILE_BYTES;
#pragma unroll
          for (int k = 0; k < BLOCK_K / UMMA_K; ++k) {
            const uint64_t ad = make_smem_desc(sa + k * 32, 16, 1024);
            const uint64_t bd = make_smem_desc(sb + k * 32, 16, 1024);
            umma_ss_pair<kTf32>(d_tmem, ad, bd, IDESC, (kb > 0 || k > 0) ? 1u : 0u);
          }
          umma_commit_pair(&empty_bar[s], pair_mask);
          if (++s == CE_STAGES) { s = 0; ph ^= 1; }
        }
        umma_commit_pair(&acc_full[a], pair_mask);
      }
    }
  } else if (warp >= 4) {
    // ===================== epilogue: softmax statistics, exchange, gradient =====================
    const int q = warp & 3;             // TMEM lane quarter -> rows q*32 .. q*32+31 of this CTA's 128
    const int chalf = (warp - 4) >> 2;  // column half of the CTA's 256 classes
    uint8_t* sbuf_ptr = smem + CE_PIPE_BYTES + (warp - 4) * EPI_BUF_BYTES;
    float scale = p.scale;
    if (p.n_valid != nullptr) {  // mean over the rows whose label is not ignore_index
      const int nv = *p.n_valid;
      scale = nv > 0 ? scale * p.batch / (float)nv : 0.f;
    }
    float csacc[CHUNKS];
#pragma unroll
    for (int c = 0; c < CHUNKS; ++c) csacc[c] = 0.f;
    float loss_acc = 0.f;
    const int wcol0 = n0 + chalf * (BLOCK_N / 2);   // first class of this warp's 128 columns
    int it = 0;
    for (int w = unit; w < p.m_tiles; w += nunits, ++it) {
      const int a = it & 1;
      const uint32_t aph = (it >> 1) & 1;
      const int row0 = w * TILE_M + rank * BLOCK_M + q * 32;
      const int row = row0 + lane;
      const bool row_ok = row < p.M;
      // the label: fetched while the MMAs still run
      int y = -1;
      bool valid = false;
      if (row_ok) {
        const long long y64 = __ldg(p.labels + row);
        if (y64 >= 0 && y64 < p.N && y64 != p.ignore_index) {
          y = (int)y64;
          valid = true;
        } else if (y64 != p.ignore_index && nj == 0 && chalf == 0) {
          flag_status(p.status, B200_STATUS_BAD_LABEL);  // torch raises; here the row is dropped like an ignored one
        }
      }
      mbar_wait(&acc_full[a], aph);
      tc_fence_after();
      const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + a * BLOCK_N + chalf * (BLOCK_N / 2);

      // ---- pass 1: online max / sum-exp over this thread's columns; exponentials go back into TMEM
      float m_run = -INFINITY, s_run = 0.f, zy = 0.f;
      float mref[CHUNKS];
#pragma unroll
      for (int c = 0; c < CHUNKS; ++c) {
        mref[c] = 0.f;
        const int col0 = wcol0 + c * 32;
        if (col0 < p.N) {  // warp-uniform
          uint32_t v[32];
          tmem_ld32(taddr + c * 32, v);
          tmem_ld_wait();
          const bool full = col0 + 32 <= p.N;
          float z[32];
#pragma unroll
          for (int j = 0; j < 32; j += 4) {
            if (full || col0 + j < p.N) {  // N % 4 == 0: a float4 of biases is all in or all out
              const float4 b4 = __ldg(reinterpret_cast<const float4*>(p.bias + col0 + j));
              z[j + 0] = __uint_as_float(v[j + 0]) + b4.x;
              z[j + 1] = __uint_as_float(v[j + 1]) + b4.y;
              z[j + 2] = __uint_as_float(v[j + 2]) + b4.z;
              z[j + 3] = __uint_as_float(v[j + 3]) + b4.w;
            } else {
              z[j + 0] = z[j + 1] = z[j + 2] = z[j + 3] = -INFINITY;  // classes that do not exist
            }
          }
          if constexpr (!kTf32) {  // the logits as they would have been stored (bf16, RNE)
#pragma unroll
            for (int j = 0; j < 32; j += 2) {
              const uint32_t pk = pack_bf16x2(z[j], z[j + 1]);
              z[j] = bf16_lo(pk);
              z[j + 1] = bf16_hi(pk);
            }
          }
          float cm[8];
#pragma unroll
          for (int j = 0; j < 8; ++j) cm[j] = fmaxf(fmaxf(z[j], z[j + 8]), fmaxf(z[j + 16], z[j + 24]));
          const float cmax = fmaxf(fmaxf(fmaxf(cm[0], cm[1]), fmaxf(cm[2], cm[3])),
                                   fmaxf(fmaxf(cm[4], cm[5]), fmaxf(cm[6], cm[7])));
          const float m_new = fmaxf(m_run, cmax);   // finite: the chunk holds at least one real class
          s_run *= ex2_approx((m_run - m_new) * L2E);  // first chunk: 0 * 2^-inf = 0
          const float mnl = m_new * L2E;
          if (__any_sync(0xffffffffu, (unsigned)(y - col0) < 32u)) {  // some row's label lives in this chunk
            const int yrel = y - col0;
#pragma unroll
            for (int j = 0; j < 32; ++j) zy = (j == yrel) ? z[j] : zy;
          }
          float cs[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            z[j] = ex2_approx(fmaf(z[j], L2E, -mnl));  // exp(z - m_new); missing classes give 2^-inf = 0
            cs[j & 3] += z[j];
            v[j] = __float_as_uint(z[j]);
          }
          s_run += (cs[0] + cs[1]) + (cs[2] + cs[3]);
          tmem_st32(taddr + c * 32, v);
          mref[c] = m_new;
          m_run = m_new;
        }
      }
      tmem_st_wait();

      // ---- exchange (m, s) with the CTAs that hold the same rows, then the row's global max and normaliser
      {
        const uint32_t slot = smem_u32(stats) +
                              ((((uint32_t)a * CE_MAX_PAIRS + nj) * 2 + chalf) * BLOCK_M + q * 32 + lane) * 8;
        for (int k = 0; k < p.pairs; ++k) st_cluster_v2f(mapa_shared(slot, 2 * k + rank), m_run, s_run);
        fence_acq_rel_cluster();
        __syncwarp();
        if (lane == 0) {
          const uint32_t sb = smem_u32(&stat_bar[a]);
          for (int k = 0; k < p.pairs; ++k) mbar_arrive_release_cluster(mapa_shared(sb, 2 * k + rank));
        }
        mbar_wait_acquire_cluster(&stat_bar[a], aph);
      }
      float Mx = -INFINITY, Z = 0.f;
      {
        const uint32_t base = smem_u32(stats) + (((uint32_t)a * CE_MAX_PAIRS * 2) * BLOCK_M + q * 32 + lane) * 8;
        float2 st[2 * CE_MAX_PAIRS];
#pragma unroll
        for (int k = 0; k < 2 * CE_MAX_PAIRS; ++k) {
          st[k] = make_float2(-INFINITY, 0.f);
          if (k < 2 * p.pairs) st[k] = ld_shared_v2f(base + k * BLOCK_M * 8);
          Mx = fmaxf(Mx, st[k].x);
        }
#pragma unroll
        for (int k = 0; k < 2 * CE_MAX_PAIRS; ++k) Z += st[k].y * ex2_approx((st[k].x - Mx) * L2E);
      }
      if (valid) {
        if (nj == 0 && chalf == 0) loss_acc += Mx + logf(Z);          // logsumexp, once per row
        if ((unsigned)(y - wcol0) < (unsigned)(BLOCK_N / 2)) loss_acc -= zy;  // minus the label's logit, by its owner
      }
      const float inv = valid ? scale / Z : 0.f;   // ignored / out-of-range / padding rows get a zero gradient
      const float mxl = Mx * L2E;

      // ---- pass 2: gradient from the cached exponentials, staged in swizzled shared memory, written by TMA
#pragma unroll
      for (int gi = 0; gi < GROUPS_PER_WARP; ++gi) {
        const int gcol0 = wcol0 + gi * EPR;
        if (gcol0 >= p.N) break;  // warp-uniform
        uint32_t vv[LDS_PER_GROUP][32];
#pragma unroll
        for (int h = 0; h < LDS_PER_GROUP; ++h) tmem_ld32(taddr + (gi * LDS_PER_GROUP + h) * 32, vv[h]);
        tmem_ld_wait();
#pragma unroll
        for (int h = 0; h < LDS_PER_GROUP; ++h) {
          const int c = gi * LDS_PER_GROUP + h;
          const int col0 = gcol0 + h * 32;
          const bool live = col0 < p.N;  // warp-uniform; a dead chunk was never written in pass 1
          const float fc = live ? ex2_approx(fmaf(mref[c], L2E, -mxl)) * inv : 0.f;
          float f[32];
#pragma unroll
          for (int j = 0; j < 32; ++j) f[j] = live ? __uint_as_float(vv[h][j]) * fc : 0.f;
          if (__any_sync(0xffffffffu, valid && (unsigned)(y - col0) < 32u)) {
            const int yrel = valid ? y - col0 : -1;
#pragma unroll
            for (int j = 0; j < 32; ++j) f[j] = (j == yrel) ? f[j] - scale : f[j];   // minus the one-hot
          }
          {
            // column sums over this warp's 32 rows of what is stored (rounded to the output dtype): halving
            // butterfly, 31 shuffles, lane i ends with the sum of column col0 + i
            float r[32];
#pragma unroll
            for (int j = 0; j < 32; ++j) r[j] = kTf32 ? f[j] : __bfloat162float(__float2bfloat16_rn(f[j]));
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
            csacc[c] += r[0];
          }
          if (h == 0) {  // the previous store of this warp must have finished READING the slab before we refill it
            if (lane == 0) bulk_wait_read<0>();
            __syncwarp();
          }
          const uint32_t srow = smem_u32(sbuf_ptr) + lane * 128;
          if constexpr (kTf32) {
#pragma unroll
            for (int cc = 0; cc < 8; ++cc)
              st_shared_v4(srow + ((cc ^ (lane & 7)) << 4), __float_as_uint(f[4 * cc]), __float_as_uint(f[4 * cc + 1]),
                           __float_as_uint(f[4 * cc + 2]), __float_as_uint(f[4 * cc + 3]));
          } else {
#pragma unroll
            for (int cc = 0; cc < 4; ++cc)
              st_shared_v4(srow + (((h * 4 + cc) ^ (lane & 7)) << 4), pack_bf16x2(f[8 * cc], f[8 * cc + 1]),
                           pack_bf16x2(f[8 * cc + 2], f[8 * cc + 3]), pack_bf16x2(f[8 * cc + 4], f[8 * cc + 5]),
                           pack_bf16x2(f[8 * cc + 6], f[8 * cc + 7]));
          }
        }
        fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0 && row0 < p.M) {
          tma_store_2d(&tmC, sbuf_ptr, gcol0, row0);
          bulk_commit();
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_cluster(&acc_empty[a], leader);  // the accumulator stage may be overwritten
    }
    // ---- partials: one row per (cluster, CTA of the pair, lane quarter); every cell is written exactly once
    {
      const long long prow = ((long long)unit * 2 + rank) * 4 + q;
      float* dst = p.partials + prow * p.pstride;
#pragma unroll
      for (int c = 0; c < CHUNKS; ++c) dst[wcol0 + c * 32 + lane] = csacc[c];
      const float wl = warp_sum(loss_acc);
      if (lane == 0) dst[p.pairs * BLOCK_N + nj * 2 + chalf] = wl;
    }
    if (lane == 0) bulk_wait_all();
  }

  tc_fence_before();
  cluster_sync_all();  // no CTA may leave while a peer can still write its shared memory or read its TMEM
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc_pair(tmem_base, NUM_ACC * BLOCK_N);
  }
}

// fixed-order sum of the per-CTA partial rows (defined in loss.cu)
int launch_reduce_partials_loss(const float* partials, int nrows, int row_stride, int ncols, float* out, int loss_col,
                                int loss_ncols, float loss_scale, float* loss_out, int accumulate,
                                const int32_t* n_valid, float batch_rows, cudaStream_t st);

static int ce_pairs_for(int64_t N) {
  const int n_tiles = (int)((N + BLOCK_N - 1) / BLOCK_N);
  return n_tiles <= 1 ? 1 : (n_tiles == 2 ? 2 : 4);
}

template <int kTf32>
static int ce_max_clusters(int pairs, int* out) {
  static int cached[CE_MAX_PAIRS + 1] = {0, 0, 0, 0, 0};
  if (cached[pairs] == 0) {
    auto kern = linear_ce_kernel<kTf32>;
    B200_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, CE_SMEM_BYTES));
    if (2 * pairs > 8)
      B200_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(2 * pairs * 64);
    cfg.blockDim = dim3(GEMM_THREADS);
    cfg.dynamicSmemBytes = CE_SMEM_BYTES;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2 * pairs;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    int n = 0;
    B200_CHECK_CUDA(cudaOccupancyMaxActiveClusters(&n, kern, &cfg));
    B200_REQUIRE(n >= 1, "no cluster of %d CTAs with %d bytes of shared memory fits on this device", 2 * pairs,
                 CE_SMEM_BYTES);
    cached[pairs] = n;
  }
  *out = cached[pairs];
  return 0;
}

}  // namespace b200

using namespace b200;

// one partial row per (cluster, CTA of the pair, lane quarter): at most sm_count * 4 rows
extern "C" int64_t b200_linear_ce_scratch_floats(int64_t N) {
  return (int64_t)sm_count_physical() * 4 * (ce_pairs_for(N) * BLOCK_N + CE_LOSS_COLS);
}

extern "C" int b200_linear_ce(const void* x, int64_t ldx, const void* w, int64_t ldw, const float* bias,
                              const int64_t* labels, void* dlogits, int64_t lddl, float* loss_out, float* dbias,
                              float* scratch, int64_t scratch_floats, int64_t M, int64_t N, int64_t K, int dtype,
                              float grad_scale, float loss_scale, int accumulate, int64_t ignore_index,
                              const int32_t* n_valid, b200_stream_t stream) {
  if (int r = check_dims(M, N, K, dtype)) return r;
  B200_REQUIRE(N <= CE_MAX_PAIRS * BLOCK_N, "b200_linear_ce handles up to %d classes (got %lld); use b200_linear_fwd + "
               "b200_ce_loss_grad", CE_MAX_PAIRS * BLOCK_N, (long long)N);
  B200_REQUIRE(bias != nullptr && labels != nullptr && scratch != nullptr, "bias, labels and scratch must not be NULL");
  B200_REQUIRE((reinterpret_cast<uintptr_t>(bias) & 15) == 0, "bias must be 16-byte aligned");
  const int tf32 = dtype == B200_F32;
  const int bk = tf32 ? 32 : 64;
  const int pairs = ce_pairs_for(N);
  int max_clusters = 0;
  if (int r = tf32 ? ce_max_clusters<1>(pairs, &max_clusters) : ce_max_clusters<0>(pairs, &max_clusters)) return r;
  CUtensorMap ta, tb, tc;
  if (int r = make_tmap(&ta, tf32, x, K, M, ldx, bk, BLOCK_M)) return r;
  if (int r = make_tmap(&tb, tf32, w, K, N, ldw, bk, BLOCK_N / 2)) return r;
  if (int r = make_tmap(&tc, tf32, dlogits, N, M, lddl, bk, 32, false, true)) return r;
  CeParams p{};
  p.M = (int)M; p.N = (int)N;
  p.k_blocks = (int)((K + bk - 1) / bk);
  p.m_tiles = (int)((M + 2 * BLOCK_M - 1) / (2 * BLOCK_M));
  p.pairs = pairs;
  p.bias = bias; p.labels = labels; p.partials = scratch;
  p.pstride = pairs * BLOCK_N + CE_LOSS_COLS;
  p.scale = grad_scale; p.batch = (float)M;
  p.ignore_index = ignore_index; p.n_valid = n_valid; p.status = status_word();
  int clusters = sm_count() / (2 * pairs);
  if (clusters > max_clusters) clusters = max_clusters;
  if (clusters > p.m_tiles) clusters = p.m_tiles;
  B200_REQUIRE(clusters >= 1, "b200_linear_ce: the SM budget (%d) does not hold one cluster of %d CTAs", sm_count(),
               2 * pairs);
  const int nrows = clusters * 8;
  B200_REQUIRE(scratch_floats >= (int64_t)nrows * p.pstride, "b200_linear_ce scratch too small: need %lld floats",
               (long long)nrows * p.pstride);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(clusters * 2 * pairs);
  cfg.blockDim = dim3(GEMM_THREADS);
  cfg.dynamicSmemBytes = CE_SMEM_BYTES;
  cfg.stream = st;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 2 * pairs;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[1].val.programmaticStreamSerializationAllowed = pdl_enabled() ? 1 : 0;
  cfg.attrs = attr;
  cfg.numAttrs = 2;
  if (tf32) B200_CHECK_CUDA(cudaLaunchKernelEx(&cfg, linear_ce_kernel<1>, ta, tb, tc, p));
  else B200_CHECK_CUDA(cudaLaunchKernelEx(&cfg, linear_ce_kernel<0>, ta, tb, tc, p));
  // bias gradient (class columns) and loss (the 2 * pairs loss columns behind them), fixed summation order
  return launch_reduce_partials_loss(scratch, nrows, p.pstride, (int)N, dbias, pairs * BLOCK_N, 2 * pairs,
                                     loss_scale > 0.f ? loss_scale : 1.0f / (float)M, loss_out, accumulate, n_valid,
                                     (float)M, st);
}
