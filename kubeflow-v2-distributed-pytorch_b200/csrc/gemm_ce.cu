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
//   pass 1  online softmax over the thread's 128 columns: z = acc + bias in fp32 (the logits are never rounded to the
//           storage dtype, because they are never stored), running max m and Σ exp(z − m); the exponentials are
//           written back INTO the accumulator columns (tcgen05.st), each 32-column chunk with the running max it
//           was formed against, so pass 2 needs no second exp per element;
//   xchg    the four column-quarter warps of a row fold their (m, Σ) in shared memory; the CTA's pair goes to the CTAs
//           that hold the same rows with st.async (distributed shared memory; the bytes are credited to an mbarrier of
//           the receiving CTA, so no fences are needed), and every thread combines the `pairs` entries of its row
//           into the row's max M and Z = Σ exp(z − M);
//   pass 2  dlogits = e · exp(m_chunk − M) · grad_scale / Z, minus grad_scale at the label; rounded, staged in
//           swizzled shared memory and written by TMA; column sums of what is stored accumulate in registers
//           (output-layer bias gradient); the row's loss M + ln Z − z[y] is summed per thread.
// Partials (bias gradient, loss) leave as one row per (cluster, CTA, lane quarter) and are summed in a fixed order by
// reduce_partials_kernel, so both are run-to-run deterministic.
//
// Roles (640 threads): warp0 = TMA producer, warp1 = MMA issuer, warp2 = TMEM allocator, warps4-19 = epilogue
// (4 TMEM lane quarters x 4 column quarters of 64 classes: the epilogue is the long pole of this kernel and needs the
// warps to hide its TMEM / MUFU / shared-memory latencies).
#include "gemm_common.cuh"

namespace b200 {

constexpr int CE_THREADS = 640;                                       // 4 control warps + 16 epilogue warps
constexpr int CE_EPI_WARPS = 16;                                      // 4 lane quarters x 4 column quarters
constexpr int CE_WARP_COLS = BLOCK_N / 4;                             // 64 classes per epilogue warp
constexpr int CE_STAGES = 4;                                          // operand ring: 4 x 32 KiB
constexpr int CE_STAGE_BYTES = A_TILE_BYTES + (BLOCK_N / 2) * 128;
constexpr int CE_PIPE_BYTES = CE_STAGES * CE_STAGE_BYTES;
constexpr int CE_STAGING_BYTES = CE_EPI_WARPS * EPI_BUF_BYTES;        // one 32-row x 128-byte slab per epilogue warp
constexpr int CE_MAX_PAIRS = 4;                                       // CTA pairs along N: up to 1024 classes
constexpr int CE_STATS_BYTES = 2 * CE_MAX_PAIRS * BLOCK_M * 8;        // [stage][pair][row] float2, filled by the peers
constexpr int CE_LSTATS_BYTES = 4 * BLOCK_M * 8;                      // [column quarter][row] float2, this CTA only
constexpr int CE_BIAS_BYTES = BLOCK_N * 4;                            // this CTA's 256 biases
constexpr int CE_SMEM_BYTES = CE_PIPE_BYTES + CE_STAGING_BYTES + CE_STATS_BYTES + CE_LSTATS_BYTES + CE_BIAS_BYTES +
                              1024 /*align*/ + 256 /*barriers*/;
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
// 8 bytes into a peer CTA's shared memory; their arrival is counted (complete_tx) on an mbarrier of THAT CTA, so the
// reader needs no fence: once its barrier phase completes the bytes are there
__device__ __forceinline__ void st_async_v2f(uint32_t raddr, float a, float b, uint32_t rbar) {
  asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.v2.f32 [%0], {%1, %2}, [%3];" ::"r"(raddr),
               "f"(a), "f"(b), "r"(rbar)
               : "memory");
}
__device__ __forceinline__ float2 ld_shared_v2f(uint32_t saddr) {
  float2 r;
  asm volatile("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(r.x), "=f"(r.y) : "r"(saddr));
  return r;
}
__device__ __forceinline__ void st_shared_v2f(uint32_t saddr, float a, float b) {
  asm volatile("st.shared.v2.f32 [%0], {%1, %2};" ::"r"(saddr), "f"(a), "f"(b) : "memory");
}

template <int kTf32>
__global__ void __launch_bounds__(CE_THREADS, 1)
linear_ce_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                 const __grid_constant__ CUtensorMap tmC, const CeParams p) {
  constexpr int ESIZE = kTf32 ? 4 : 2;
  constexpr int BLOCK_K = 128 / ESIZE;
  constexpr int UMMA_K = 32 / ESIZE;
  constexpr int EPR = 128 / ESIZE;                     // elements per 128-byte row = columns per output slab
  constexpr uint32_t IDESC = make_idesc(kTf32 ? 2 : 1, 0, 0, BLOCK_M * 2, BLOCK_N);
  constexpr int B_ROWS = BLOCK_N / 2;
  constexpr int TILE_M = BLOCK_M * 2;
  constexpr int CHUNKS = CE_WARP_COLS / 32;            // 32-column chunks per epilogue warp (2)
  constexpr int CHUNKS_PER_SLAB = EPR / 32;            // chunks per output slab: 2 (bf16) / 1 (fp32)

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* stats = smem + CE_PIPE_BYTES + CE_STAGING_BYTES;
  uint8_t* lstats = stats + CE_STATS_BYTES;
  float* bias_smem = reinterpret_cast<float*>(lstats + CE_LSTATS_BYTES);
  uint64_t* bars = reinterpret_cast<uint64_t*>(lstats + CE_LSTATS_BYTES + CE_BIAS_BYTES);
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
      mbar_init(&acc_empty[i], CE_EPI_WARPS * 2);   // one arrive per epilogue warp of both CTAs of the pair
      mbar_init(&stat_bar[i], 1);                   // one local arrive (expect_tx); the peers' bytes complete it
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
    const int cq = (warp - 4) >> 2;     // column quarter of the CTA's 256 classes
    uint8_t* sbuf_ptr = smem + CE_PIPE_BYTES + (warp - 4) * EPI_BUF_BYTES;
    const uint32_t sbuf = smem_u32(sbuf_ptr);
    float scale = p.scale;
    if (p.n_valid != nullptr) {  // mean over the rows whose label is not ignore_index
      const int nv = *p.n_valid;
      scale = nv > 0 ? scale * p.batch / (float)nv : 0.f;
    }
    const int wcol0 = n0 + cq * CE_WARP_COLS;   // first class of this warp's 64 columns
    // the CTA's 256 biases -> shared memory once (broadcast 128-bit reads later); one warp per column quarter
    const uint32_t sbias = smem_u32(bias_smem) + cq * CE_WARP_COLS * 4;
    if (q == 0) {
#pragma unroll
      for (int k = 0; k < CHUNKS; ++k) {
        const int col = wcol0 + k * 32 + lane;
        bias_smem[cq * CE_WARP_COLS + k * 32 + lane] = col < p.N ? __ldg(p.bias + col) : 0.f;
      }
    }
    asm volatile("bar.sync 5, 512;" ::: "memory");   // the sixteen epilogue warps
    // read-back addresses for the column sums of a stored slab: lane owns 4 bytes (a bf16 column pair / one fp32
    // column) of every row; 16-byte chunk c of row r sits at chunk c ^ (r & 7)
    uint32_t rb_addr[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) rb_addr[k] = sbuf + ((((uint32_t)lane >> 2) ^ (uint32_t)k) << 4) + (lane & 3) * 4;
    float csacc[2] = {0.f, 0.f};   // bf16: columns (2*lane, 2*lane+1) of the warp's 64;  fp32: columns lane, 32+lane
    float loss_acc = 0.f;
    // Column sums of a slab AS STORED (output-layer bias gradient): every lane walks the 32 rows of its own 4 bytes —
    // conflict-free under the swizzle.  bf16: one slab per warp and tile, so the walk is deferred into the next tile's
    // wait for the statistics exchange, where the warp would otherwise idle (the slab stays intact until pass 2).
    auto colsum_slab_bf16 = [&]() {
      float lo0 = 0.f, hi0 = 0.f, lo1 = 0.f, hi1 = 0.f;
#pragma unroll
      for (int r = 0; r < 32; r += 2) {
        const uint32_t w0 = ld_shared_u32(rb_addr[r & 7] + r * 128);
        const uint32_t w1 = ld_shared_u32(rb_addr[(r + 1) & 7] + (r + 1) * 128);
        fadd2(lo0, hi0, bf16_lo(w0), bf16_hi(w0));
        fadd2(lo1, hi1, bf16_lo(w1), bf16_hi(w1));
      }
      csacc[0] += lo0 + lo1;
      csacc[1] += hi0 + hi1;
    };
    const bool warp_live = wcol0 < p.N;   // this warp's 64 columns hold at least one real class
    const uint32_t my_lstat = smem_u32(lstats) + (uint32_t)((cq * BLOCK_M + q * 32 + lane) * 8);
    int it = 0;
    long long y64 = (unit < p.m_tiles && unit * TILE_M + rank * BLOCK_M + q * 32 + lane < p.M)
                        ? __ldg(p.labels + unit * TILE_M + rank * BLOCK_M + q * 32 + lane) : p.ignore_index;
    for (int w = unit; w < p.m_tiles; w += nunits, ++it) {
      const int a = it & 1;
      const uint32_t aph = (it >> 1) & 1;
      const int row0 = w * TILE_M + rank * BLOCK_M + q * 32;
      const int row = row0 + lane;
      const bool row_ok = row < p.M;
      int y = -1;
      bool valid = false;
      if (row_ok) {
        if (y64 >= 0 && y64 < p.N && y64 != p.ignore_index) {
          y = (int)y64;
          valid = true;
        } else if (y64 != p.ignore_index && nj == 0 && cq == 0) {
          flag_status(p.status, B200_STATUS_BAD_LABEL);  // torch raises; here the row is dropped like an ignored one
        }
      }
      {  // the next tile's label: in flight while this tile is processed
        const long long nrow = (long long)(w + nunits) * TILE_M + rank * BLOCK_M + q * 32 + lane;
        y64 = (w + nunits < p.m_tiles && nrow < p.M) ? __ldg(p.labels + nrow) : p.ignore_index;
      }
      mbar_wait(&acc_full[a], aph);
      tc_fence_after();
      const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + a * BLOCK_N + cq * CE_WARP_COLS;

      // ---- pass 1: online max / sum-exp over this thread's 64 columns; exponentials go back into TMEM
      float m_run = -INFINITY, s_run = 0.f, zy = 0.f;
      float mref[CHUNKS];
#pragma unroll
      for (int c = 0; c < CHUNKS; ++c) {
        mref[c] = 0.f;
        const int col0 = wcol0 + c * 32;
        if (col0 < p.N) {  // warp-uniform
          uint32_t v[32];
          tmem_ld32(taddr + c * 32, v);
          float z[32];
#pragma unroll
          for (int j = 0; j < 32; j += 4) {
            const uint4 b4 = ld_shared_v4(sbias + (c * 32 + j) * 4);
            z[j + 0] = __uint_as_float(b4.x); z[j + 1] = __uint_as_float(b4.y);
            z[j + 2] = __uint_as_float(b4.z); z[j + 3] = __uint_as_float(b4.w);
          }
          tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < 32; j += 2) fadd2(z[j], z[j + 1], __uint_as_float(v[j]), __uint_as_float(v[j + 1]));
          if (col0 + 32 > p.N) {  // ragged last chunk: classes that do not exist
#pragma unroll
            for (int j = 0; j < 32; ++j) z[j] = (col0 + j < p.N) ? z[j] : -INFINITY;
          }
          float cm[8];
#pragma unroll
          for (int j = 0; j < 8; ++j) cm[j] = fmaxf(fmaxf(z[j], z[j + 8]), fmaxf(z[j + 16], z[j + 24]));
          const float cmax = fmaxf(fmaxf(fmaxf(cm[0], cm[1]), fmaxf(cm[2], cm[3])),
                                   fmaxf(fmaxf(cm[4], cm[5]), fmaxf(cm[6], cm[7])));
          const float m_new = fmaxf(m_run, cmax);   // finite: the chunk holds at least one real class
          s_run *= ex2_approx((m_run - m_new) * L2E);  // first chunk: 0 * 2^-inf = 0
          const float nmnl = -m_new * L2E;
          if (__any_sync(0xffffffffu, (unsigned)(y - col0) < 32u)) {  // some row's label lives in this chunk
            const int yrel = y - col0;
#pragma unroll
            for (int j = 0; j < 32; ++j) zy = (j == yrel) ? z[j] : zy;
          }
          float cs[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
          for (int j = 0; j < 32; j += 4) {
            // exp(z - m_new) = 2^(z*log2e - m_new*log2e); missing classes give 2^-inf = 0
            ffma2(z[j], z[j + 1], z[j], z[j + 1], L2E, L2E, nmnl, nmnl);
            ffma2(z[j + 2], z[j + 3], z[j + 2], z[j + 3], L2E, L2E, nmnl, nmnl);
            z[j] = ex2_approx(z[j]); z[j + 1] = ex2_approx(z[j + 1]);
            z[j + 2] = ex2_approx(z[j + 2]); z[j + 3] = ex2_approx(z[j + 3]);
            fadd2(cs[0], cs[1], z[j], z[j + 1]);
            fadd2(cs[2], cs[3], z[j + 2], z[j + 3]);
          }
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] = __float_as_uint(z[j]);
          s_run += (cs[0] + cs[1]) + (cs[2] + cs[3]);
          tmem_st32(taddr + c * 32, v);
          mref[c] = m_new;
          m_run = m_new;
        }
      }

      // ---- (m, s) of the four column quarters of a row meet in this CTA's shared memory; the quarter-0 warp folds them
      // and sends the CTA's pair to every CTA that holds the same rows (st.async: data + barrier credit travel
      // together, so neither side needs a fence); then everybody reads the `pairs` entries of its row
      st_shared_v2f(my_lstat, m_run, s_run);
      asm volatile("bar.sync %0, 128;" ::"r"(1 + q) : "memory");   // the four warps that share these 32 rows
      if (cq == 0) {
        if (q == 0 && lane == 0) mbar_expect_tx(&stat_bar[a], (uint32_t)p.pairs * BLOCK_M * 8);
        float2 e4[4];
        float mc = -INFINITY;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          e4[k] = ld_shared_v2f(smem_u32(lstats) + (uint32_t)((k * BLOCK_M + q * 32 + lane) * 8));
          mc = fmaxf(mc, e4[k].x);
        }
        const float mcs = (mc == -INFINITY) ? 0.f : mc;   // a CTA with no real class at all sends (-inf, 0), not NaN
        float sc = 0.f;
#pragma unroll
        for (int k = 0; k < 4; ++k) sc += e4[k].y * ex2_approx((e4[k].x - mcs) * L2E);  // empty quarter: 0 * 2^-inf
        const uint32_t slot = smem_u32(stats) + (uint32_t)((((a * CE_MAX_PAIRS) + nj) * BLOCK_M + q * 32 + lane) * 8);
        const uint32_t sb = smem_u32(&stat_bar[a]);
        for (int k = 0; k < p.pairs; ++k) st_async_v2f(mapa_shared(slot, 2 * k + rank), mc, sc, mapa_shared(sb, 2 * k + rank));
      }
      if constexpr (!kTf32) {
        if (it > 0 && warp_live) colsum_slab_bf16();   // the previous tile's slab, while the exchange is in flight
      }
      mbar_wait(&stat_bar[a], aph);
      float Mx = -INFINITY, Z = 0.f;
      {
        const uint32_t base = smem_u32(stats) + (uint32_t)(((a * CE_MAX_PAIRS) * BLOCK_M + q * 32 + lane) * 8);
        float2 st[CE_MAX_PAIRS];
#pragma unroll
        for (int k = 0; k < CE_MAX_PAIRS; ++k) {
          st[k] = make_float2(-INFINITY, 0.f);
          if (k < p.pairs) st[k] = ld_shared_v2f(base + k * BLOCK_M * 8);
          Mx = fmaxf(Mx, st[k].x);
        }
#pragma unroll
        for (int k = 0; k < CE_MAX_PAIRS; ++k) Z += st[k].y * ex2_approx((st[k].x - Mx) * L2E);
      }
      if (valid) {
        if (nj == 0 && cq == 0) loss_acc += Mx + logf(Z);                       // logsumexp, once per row
        if ((unsigned)(y - wcol0) < (unsigned)CE_WARP_COLS) loss_acc -= zy;     // minus the label's logit, by its owner
      }
      const float inv = valid ? scale / Z : 0.f;   // ignored / out-of-range / padding rows get a zero gradient
      const float mxl = Mx * L2E;
      tmem_st_wait();

      // ---- pass 2: gradient from the cached exponentials, staged in swizzled shared memory, written by TMA
#pragma unroll
      for (int c = 0; c < CHUNKS; ++c) {
        const int col0 = wcol0 + c * 32;
        const int h = c % CHUNKS_PER_SLAB;                     // position of the chunk inside its 128-byte slab row
        const int slab_col0 = col0 - h * 32;
        if (slab_col0 >= p.N) break;  // warp-uniform
        uint32_t vv[32];
        if (col0 < p.N) tmem_ld32(taddr + c * 32, vv);
        if (h == 0) {
          if (lane == 0) bulk_wait_read<0>();  // the previous store of this warp has finished READING the slab
          __syncwarp();
        }
        float f[32];
        if (col0 < p.N) {  // warp-uniform
          const float fc = ex2_approx(fmaf(mref[c], L2E, -mxl)) * inv;
          tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < 32; j += 2) {
            f[j] = __uint_as_float(vv[j]);
            f[j + 1] = __uint_as_float(vv[j + 1]);
            fmul2(f[j], f[j + 1], fc, fc);
          }
          if (__any_sync(0xffffffffu, valid && (unsigned)(y - col0) < 32u)) {
            const int yrel = valid ? y - col0 : -1;
#pragma unroll
            for (int j = 0; j < 32; ++j) f[j] = (j == yrel) ? f[j] - scale : f[j];   // minus the one-hot
          }
        } else {  // a chunk past the last class was never written in pass 1 (TMA clips it anyway)
#pragma unroll
          for (int j = 0; j < 32; ++j) f[j] = 0.f;
        }
        const uint32_t srow = sbuf + lane * 128;
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
        if (h == CHUNKS_PER_SLAB - 1) {  // slab complete
          fence_proxy_async_smem();
          __syncwarp();
          if (lane == 0 && row0 < p.M) {
            tma_store_2d(&tmC, sbuf_ptr, slab_col0, row0);
            bulk_commit();
          }
          if constexpr (kTf32) {  // fp32: two slabs per tile share the buffer, so each is summed right away
            float s0 = 0.f, s1 = 0.f;
#pragma unroll
            for (int r = 0; r < 32; r += 2) {
              s0 += __uint_as_float(ld_shared_u32(rb_addr[r & 7] + r * 128));
              s1 += __uint_as_float(ld_shared_u32(rb_addr[(r + 1) & 7] + (r + 1) * 128));
            }
            csacc[c] += s0 + s1;
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_cluster(&acc_empty[a], leader);  // the accumulator stage may be overwritten
    }
    if constexpr (!kTf32) {
      if (it > 0 && warp_live) colsum_slab_bf16();   // the last tile's slab
    }
    // ---- partials: one row per (cluster, CTA of the pair, lane quarter); every cell is written exactly once
    {
      const long long prow = ((long long)unit * 2 + rank) * 4 + q;
      float* dst = p.partials + prow * p.pstride;
      if constexpr (kTf32) {
        dst[wcol0 + lane] = csacc[0];
        dst[wcol0 + 32 + lane] = csacc[1];
      } else {
        *reinterpret_cast<float2*>(dst + wcol0 + 2 * lane) = make_float2(csacc[0], csacc[1]);
      }
      const float wl = warp_sum(loss_acc);
      if (lane == 0) dst[p.pairs * BLOCK_N + nj * 4 + cq] = wl;
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
    cfg.blockDim = dim3(CE_THREADS);
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
    if (getenv("B200STEP_VERBOSE"))
      fprintf(stderr, "[b200step] linear_ce: clusters of %d CTAs, %d bytes smem: %d co-resident\n", 2 * pairs,
              CE_SMEM_BYTES, n);
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
  cfg.blockDim = dim3(CE_THREADS);
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
  return launch_reduce_partials_loss(scratch, nrows, p.pstride, (int)N, dbias, pairs * BLOCK_N, 4 * pairs,
                                     loss_scale > 0.f ? loss_scale : 1.0f / (float)M, loss_out, accumulate, n_valid,
                                     (float)M, st);
}
