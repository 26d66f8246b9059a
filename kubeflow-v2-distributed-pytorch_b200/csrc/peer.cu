// peer.cu — K-opt-peer: DDP's gradient all-reduce (task.py:189) and optimizer.step() (task.py:312) as ONE
// kernel per bucket over NVLink peer memory, instead of an NCCL all-reduce followed by the optimizer.
//
// Every rank maps every other rank's gradient / parameter / flag buffers (CUDA IPC).  For a bucket [start, end)
// rank r owns the r-th of `world` equal slices.  Its kernel
//   1. waits until every rank has flagged "my gradients of this bucket are complete" (flags live in the WAITER's
//      memory; the producer stores them remotely, so polling never crosses NVLink),
//   2. sums the `world` gradient slices in rank order (peer loads over NVLink; one owner per element, so every
//      rank ends up with bit-identical parameters), applies the optimizer to its slice of the fp32 master copy,
//   3. stores the new parameters (and their bf16 operand copy) into EVERY rank's buffers (peer stores),
//   4. flags "slice done" on every rank.
// A finishing kernel on the main stream waits for the done-flags of all buckets from all ranks, zeroes the local
// gradient buffer (every reader of it has finished by then) and advances the step bookkeeping.
// Traffic per rank and step: (W-1)/W of the gradient bytes in, (W-1)/W of the parameter bytes out — what a ring
// all-reduce moves — but the optimizer touches 1/W of the parameters, there is no NCCL kernel, and the 256-thread,
// register-light CTAs fit on an SM NEXT TO a persistent wgrad GEMM CTA (60 registers x 384 threads) instead of
// displacing it, which is what exposes NCCL's all-reduce under those grids (profiles/r01_comm_overlap_experiments.md).
//
// Ordering argument (no double buffering needed):
//   * a rank's gradient slice is read by its owner only after that rank's ready-flag (set after its wgrad) and is
//     overwritten (zeroed, then accumulated onto by the next step) only after the owner's done-flag was seen by the
//     finishing kernel;
//   * a rank's parameters are overwritten by the owner only after that rank's ready-flag, and in the step's schedule
//     every read of the weights (forward, dgrad chain, K-pre) precedes the first ready-flag; the next forward starts
//     after the finishing kernel.
// Flags carry the step number (+1) and only ever grow, so no reset races; every wait has a time-out that records an
// error word instead of hanging the device (120 s by default).
#include "common.cuh"

namespace b200 {

constexpr int PEER_MAX_WORLD = 8;
constexpr int PEER_SLOTS = 64;          // flag slots per rank (two per bucket)
constexpr int PEER_THREADS = 256;
// How long a kernel polls a flag before it records an error and goes on: long enough for rank 0's evaluation +
// checkpoint write while the other ranks already wait for it (task.py:279-297), B200STEP_PEER_TIMEOUT_S overrides.
static unsigned long long peer_timeout_ns() {
  static unsigned long long v = 0;
  if (v == 0) {
    const char* e = getenv("B200STEP_PEER_TIMEOUT_S");
    const double s = (e && atof(e) > 0.0) ? atof(e) : 120.0;
    v = (unsigned long long)(s * 1e9);
  }
  return v;
}

struct PeerTable {
  float* g[PEER_MAX_WORLD];
  float* p[PEER_MAX_WORLD];
  void* shadow[PEER_MAX_WORLD];
  unsigned* flags[PEER_MAX_WORLD];  // [PEER_SLOTS][PEER_MAX_WORLD] per rank
  int world, rank;
};

struct PeerOptArgs {
  float* s1;            // local momentum buffer | Adam exp_avg
  float* s2;            // Adam exp_avg_sq
  long long start, end, n_shadow;
  int slot_ready, slot_done;
  const int* step;
  float lr, mu, wd;     // SGD
  double alr, b1, b2;   // Adam
  float eps;
  unsigned* ticket;
  unsigned* err;
  unsigned long long timeout_ns;
};

struct PeerSlots {
  int slot[PEER_SLOTS];
  int count;
};

__device__ __forceinline__ unsigned ld_acquire_sys(const unsigned* p) {
  unsigned v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_sys(unsigned* p, unsigned v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned long long globaltimer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
// peer gradients: never through the non-coherent path, never cached in L1
__device__ __forceinline__ float4 ld_peer4(const float* p) {
  float4 v;
  asm volatile("ld.volatile.global.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
  return v;
}

// thread `t` of the block polls flag `f` until it reaches `want` (wrap-safe); records the slot on time-out
__device__ __forceinline__ void spin_until(const unsigned* f, unsigned want, unsigned* err, unsigned code,
                                           unsigned long long timeout_ns) {
  const unsigned long long t0 = globaltimer_ns();
  while ((int)(ld_acquire_sys(f) - want) < 0) {
    if (globaltimer_ns() - t0 > timeout_ns) {
      atomicExch(err, code);
      break;
    }
    __nanosleep(200);
  }
}

__global__ void peer_signal_kernel(const PeerTable t, int slot, const int* step) {
  // plain launch: everything earlier on this stream (the event wait on the wgrad / scatter-add that produced the
  // bucket) is complete when this kernel starts
  if (threadIdx.x < t.world) {
    __threadfence_system();
    st_release_sys(t.flags[threadIdx.x] + slot * PEER_MAX_WORLD + t.rank, (unsigned)(*step) + 1u);
  }
}

template <bool kAdam>
__global__ void __launch_bounds__(PEER_THREADS)
peer_opt_kernel(const PeerTable t, const PeerOptArgs a) {
  const unsigned want = (unsigned)(*a.step) + 1u;
  if (threadIdx.x < t.world)
    spin_until(t.flags[t.rank] + a.slot_ready * PEER_MAX_WORLD + threadIdx.x, want, a.err, 1u + a.slot_ready,
               a.timeout_ns);
  __syncthreads();

  const long long n = a.end - a.start;
  const long long per = (((n + t.world - 1) / t.world) + 3) & ~3ll;   // bucket ranges are multiples of 64 floats
  const long long lo = a.start + per * t.rank;
  long long hi = lo + per;
  if (hi > a.end) hi = a.end;
  const long long nv = hi > lo ? (hi - lo) >> 2 : 0;
  const long long stride = (long long)gridDim.x * blockDim.x;
  float* p_local = t.p[t.rank];

  float step_size = 0.f, bc2_sqrt = 1.f, w1 = 0.f, w2 = 0.f, b2f = 0.f;
  if constexpr (kAdam) {
    const double tt = (double)(*a.step) + 1.0;   // bias corrections in double like torch
    step_size = (float)(a.alr / (1.0 - pow(a.b1, tt)));
    bc2_sqrt = (float)sqrt(1.0 - pow(a.b2, tt));
    w1 = (float)(1.0 - a.b1);
    w2 = (float)(1.0 - a.b2);
    b2f = (float)a.b2;
  }

  for (long long v = (long long)blockIdx.x * blockDim.x + threadIdx.x; v < nv; v += stride) {
    const long long i = lo + (v << 2);
    // all peer loads are issued before the first add (one NVLink round trip per element, not `world` of them);
    // the sum runs in rank order 0..W-1 whoever owns the element
    float4 gq[PEER_MAX_WORLD];
#pragma unroll
    for (int q = 0; q < PEER_MAX_WORLD; ++q)
      if (q < t.world) gq[q] = ld_peer4(t.g[q] + i);
    float4 gs = gq[0];
#pragma unroll
    for (int q = 1; q < PEER_MAX_WORLD; ++q) {
      if (q < t.world) {
        gs.x = __fadd_rn(gs.x, gq[q].x);
        gs.y = __fadd_rn(gs.y, gq[q].y);
        gs.z = __fadd_rn(gs.z, gq[q].z);
        gs.w = __fadd_rn(gs.w, gq[q].w);
      }
    }
    float4 pp = *reinterpret_cast<const float4*>(p_local + i);
    float4 m1 = *reinterpret_cast<const float4*>(a.s1 + i);
    if constexpr (kAdam) {
      float4 m2 = *reinterpret_cast<const float4*>(a.s2 + i);
      adam_one(pp.x, gs.x, m1.x, m2.x, a.wd, w1, b2f, w2, a.eps, step_size, bc2_sqrt);
      adam_one(pp.y, gs.y, m1.y, m2.y, a.wd, w1, b2f, w2, a.eps, step_size, bc2_sqrt);
      adam_one(pp.z, gs.z, m1.z, m2.z, a.wd, w1, b2f, w2, a.eps, step_size, bc2_sqrt);
      adam_one(pp.w, gs.w, m1.w, m2.w, a.wd, w1, b2f, w2, a.eps, step_size, bc2_sqrt);
      *reinterpret_cast<float4*>(a.s2 + i) = m2;
    } else {
      // grad.add(param, alpha=wd) -> buf.mul_(mu).add_(grad) -> param.add_(buf, alpha=-lr)
      m1.x = __fadd_rn(__fmul_rn(m1.x, a.mu), __fmaf_rn(pp.x, a.wd, gs.x));
      m1.y = __fadd_rn(__fmul_rn(m1.y, a.mu), __fmaf_rn(pp.y, a.wd, gs.y));
      m1.z = __fadd_rn(__fmul_rn(m1.z, a.mu), __fmaf_rn(pp.z, a.wd, gs.z));
      m1.w = __fadd_rn(__fmul_rn(m1.w, a.mu), __fmaf_rn(pp.w, a.wd, gs.w));
      pp.x = __fmaf_rn(m1.x, -a.lr, pp.x);
      pp.y = __fmaf_rn(m1.y, -a.lr, pp.y);
      pp.z = __fmaf_rn(m1.z, -a.lr, pp.z);
      pp.w = __fmaf_rn(m1.w, -a.lr, pp.w);
    }
    *reinterpret_cast<float4*>(a.s1 + i) = m1;
    const uint2 sh = make_uint2(pack_bf16x2(pp.x, pp.y), pack_bf16x2(pp.z, pp.w));
#pragma unroll
    for (int q = 0; q < PEER_MAX_WORLD; ++q) {    // the new slice goes to every rank (peer stores, fire and forget)
      if (q < t.world) {
        *reinterpret_cast<float4*>(t.p[q] + i) = pp;
        if (i < a.n_shadow) *reinterpret_cast<uint2*>(reinterpret_cast<uint16_t*>(t.shadow[q]) + i) = sh;
      }
    }
  }

  // "slice done": the last block to finish tells every rank (all earlier blocks fenced their peer stores first)
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence_system();
    const unsigned tk = atomicAdd(a.ticket, 1u);
    if (tk == gridDim.x - 1) {
      *a.ticket = 0u;
      __threadfence_system();
      for (int q = 0; q < t.world; ++q) st_release_sys(t.flags[q] + a.slot_done * PEER_MAX_WORLD + t.rank, want);
    }
  }
}

__global__ void __launch_bounds__(PEER_THREADS)
peer_finish_kernel(const unsigned* flags_local, int world, const PeerSlots slots, const int* step, float* g,
                   long long n, const StepAdvance adv, unsigned* err, unsigned long long timeout_ns) {
  const unsigned want = (unsigned)(*step) + 1u;
  for (int w = threadIdx.x; w < slots.count * world; w += blockDim.x) {
    const int s = slots.slot[w / world];
    spin_until(flags_local + s * PEER_MAX_WORLD + (w % world), want, err, 1u + s, timeout_ns);
  }
  __syncthreads();
  // every owner has read this rank's gradients: zero them for the next step's accumulation (optimizer.zero_grad())
  const long long nv = n >> 2;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long v = (long long)blockIdx.x * blockDim.x + threadIdx.x; v < nv; v += stride)
    *reinterpret_cast<float4*>(g + (v << 2)) = make_float4(0.f, 0.f, 0.f, 0.f);
  for (long long i = (nv << 2) + (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) g[i] = 0.f;
  advance_if_last_block(adv);
}

static int fill_table(PeerTable* t, void* const* g, void* const* p, void* const* shadow, void* const* flags,
                      int world, int rank) {
  B200_REQUIRE(world >= 1 && world <= PEER_MAX_WORLD, "peer optimizer supports 1..%d ranks, got %d", PEER_MAX_WORLD,
               world);
  B200_REQUIRE(rank >= 0 && rank < world, "rank %d outside [0, %d)", rank, world);
  memset(t, 0, sizeof(*t));
  for (int q = 0; q < world; ++q) {
    B200_REQUIRE(flags[q] != nullptr, "flags[%d] is NULL", q);
    t->flags[q] = static_cast<unsigned*>(flags[q]);
    if (g) t->g[q] = static_cast<float*>(g[q]);
    if (p) t->p[q] = static_cast<float*>(p[q]);
    if (shadow) t->shadow[q] = shadow[q];
  }
  t->world = world;
  t->rank = rank;
  return 0;
}

static int peer_opt_launch(bool adam, void* const* g, void* const* p, void* const* shadow, void* const* flags,
                           int world, int rank, PeerOptArgs a, cudaStream_t st) {
  PeerTable t;
  if (int r = fill_table(&t, g, p, shadow, flags, world, rank)) return r;
  B200_REQUIRE(g != nullptr && p != nullptr, "gradient / parameter pointer tables are NULL");
  for (int q = 0; q < world; ++q) {
    B200_REQUIRE(t.g[q] != nullptr && t.p[q] != nullptr, "g[%d] / p[%d] is NULL", q, q);
    B200_REQUIRE(a.n_shadow <= 0 || t.shadow[q] != nullptr, "shadow[%d] is NULL but n_shadow > 0", q);
    B200_REQUIRE(((reinterpret_cast<uintptr_t>(t.g[q]) | reinterpret_cast<uintptr_t>(t.p[q]) |
                   reinterpret_cast<uintptr_t>(t.shadow[q])) & 15) == 0, "peer buffers must be 16-byte aligned");
  }
  B200_REQUIRE(a.start >= 0 && a.end > a.start && a.start % 4 == 0 && (a.end - a.start) % 4 == 0,
               "bucket [%lld, %lld) must start and end on multiples of 4 elements", a.start, a.end);
  B200_REQUIRE(a.slot_ready >= 0 && a.slot_ready < PEER_SLOTS && a.slot_done >= 0 && a.slot_done < PEER_SLOTS,
               "flag slots must be in [0, %d)", PEER_SLOTS);
  B200_REQUIRE(a.s1 != nullptr && a.step != nullptr && a.ticket != nullptr && a.err != nullptr &&
                   (!adam || a.s2 != nullptr), "optimizer state / step / ticket / err pointers must not be NULL");
  a.timeout_ns = peer_timeout_ns();
  const long long n = a.end - a.start;
  const long long per = (((n + world - 1) / world) + 3) & ~3ll;
  long long blocks = ((per >> 2) + PEER_THREADS - 1) / PEER_THREADS;
  const long long cap = sm_count();
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  if (adam) peer_opt_kernel<true><<<(unsigned)blocks, PEER_THREADS, 0, st>>>(t, a);
  else peer_opt_kernel<false><<<(unsigned)blocks, PEER_THREADS, 0, st>>>(t, a);
  B200_CHECK_CUDA(cudaGetLastError());
  return 0;
}

}  // namespace b200

using namespace b200;

extern "C" int b200_peer_alloc(int64_t bytes, void** ptr_out, void* handle_out) {
  B200_REQUIRE(bytes > 0 && ptr_out != nullptr && handle_out != nullptr, "bad arguments to b200_peer_alloc");
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
  void* ptr = nullptr;
  B200_CHECK_CUDA(cudaMalloc(&ptr, (size_t)bytes));
  cudaError_t e = cudaMemset(ptr, 0, (size_t)bytes);
  cudaIpcMemHandle_t h;
  if (e == cudaSuccess) e = cudaIpcGetMemHandle(&h, ptr);
  if (e != cudaSuccess) {
    cudaFree(ptr);
    B200_CHECK_CUDA(e);
  }
  memcpy(handle_out, &h, sizeof(h));
  *ptr_out = ptr;
  return 0;
}

extern "C" int b200_peer_open(const void* handle, void** ptr_out) {
  B200_REQUIRE(handle != nullptr && ptr_out != nullptr, "bad arguments to b200_peer_open");
  cudaIpcMemHandle_t h;
  memcpy(&h, handle, sizeof(h));
  B200_CHECK_CUDA(cudaIpcOpenMemHandle(ptr_out, h, cudaIpcMemLazyEnablePeerAccess));
  return 0;
}

extern "C" int b200_peer_enable_access(int peer_device) {
  int cur = 0, can = 0;
  B200_CHECK_CUDA(cudaGetDevice(&cur));
  if (peer_device == cur) return 0;
  B200_CHECK_CUDA(cudaDeviceCanAccessPeer(&can, cur, peer_device));
  B200_REQUIRE(can, "device %d cannot access device %d over NVLink/PCIe peer-to-peer", cur, peer_device);
  const cudaError_t e = cudaDeviceEnablePeerAccess(peer_device, 0);
  if (e == cudaErrorPeerAccessAlreadyEnabled) {
    (void)cudaGetLastError();  // clear the sticky-less error state
    return 0;
  }
  B200_CHECK_CUDA(e);
  return 0;
}

extern "C" int b200_peer_close(void* ptr) {
  if (ptr) B200_CHECK_CUDA(cudaIpcCloseMemHandle(ptr));
  return 0;
}

extern "C" int b200_peer_free(void* ptr) {
  if (ptr) B200_CHECK_CUDA(cudaFree(ptr));
  return 0;
}

extern "C" int b200_peer_flag_words(void) { return PEER_SLOTS * PEER_MAX_WORLD; }

extern "C" int b200_peer_signal(void* const* flags, int world, int rank, int slot, const int32_t* step,
                                b200_stream_t stream) {
  PeerTable t;
  if (int r = fill_table(&t, nullptr, nullptr, nullptr, flags, world, rank)) return r;
  B200_REQUIRE(slot >= 0 && slot < PEER_SLOTS && step != nullptr, "bad slot / step");
  peer_signal_kernel<<<1, 32, 0, static_cast<cudaStream_t>(stream)>>>(t, slot, static_cast<const int*>(step));
  B200_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int b200_peer_sgd_step(void* const* g, void* const* p, void* const* shadow, void* const* flags, int world,
                                  int rank, float* buf, int64_t start, int64_t end, int64_t n_shadow, int slot_ready,
                                  int slot_done, const int32_t* step, float lr, float momentum, float weight_decay,
                                  uint32_t* ticket, uint32_t* err, b200_stream_t stream) {
  PeerOptArgs a{};
  a.s1 = buf; a.start = start; a.end = end; a.n_shadow = n_shadow; a.slot_ready = slot_ready; a.slot_done = slot_done;
  a.step = step; a.lr = lr; a.mu = momentum; a.wd = weight_decay; a.ticket = ticket; a.err = err;
  return peer_opt_launch(false, g, p, shadow, flags, world, rank, a, static_cast<cudaStream_t>(stream));
}

extern "C" int b200_peer_adam_step(void* const* g, void* const* p, void* const* shadow, void* const* flags, int world,
                                   int rank, float* m, float* v, int64_t start, int64_t end, int64_t n_shadow,
                                   int slot_ready, int slot_done, const int32_t* step, double lr, double beta1,
                                   double beta2, float eps, float weight_decay, uint32_t* ticket, uint32_t* err,
                                   b200_stream_t stream) {
  PeerOptArgs a{};
  a.s1 = m; a.s2 = v; a.start = start; a.end = end; a.n_shadow = n_shadow; a.slot_ready = slot_ready;
  a.slot_done = slot_done; a.step = step; a.alr = lr; a.b1 = beta1; a.b2 = beta2; a.eps = eps; a.wd = weight_decay;
  a.ticket = ticket; a.err = err;
  return peer_opt_launch(true, g, p, shadow, flags, world, rank, a, static_cast<cudaStream_t>(stream));
}

extern "C" int b200_peer_finish(const void* flags_local, int world, const int32_t* slots, int n_slots, float* g,
                                int64_t n, int32_t* cursor, int32_t* step, int steps_per_epoch, uint32_t* ticket,
                                uint32_t* err, b200_stream_t stream) {
  B200_REQUIRE(flags_local != nullptr && world >= 1 && world <= PEER_MAX_WORLD, "bad flags / world");
  B200_REQUIRE(slots != nullptr && n_slots >= 1 && n_slots <= PEER_SLOTS, "1..%d slots", PEER_SLOTS);
  B200_REQUIRE(g != nullptr && n >= 0 && step != nullptr && ticket != nullptr && err != nullptr,
               "gradient / step / ticket / err pointers must not be NULL");
  B200_REQUIRE((reinterpret_cast<uintptr_t>(g) & 15) == 0, "gradient buffer must be 16-byte aligned");
  PeerSlots s{};
  for (int i = 0; i < n_slots; ++i) {
    B200_REQUIRE(slots[i] >= 0 && slots[i] < PEER_SLOTS, "slot %d out of range", slots[i]);
    s.slot[i] = slots[i];
  }
  s.count = n_slots;
  StepAdvance adv{cursor, step, steps_per_epoch, ticket};
  long long blocks = ((n >> 2) + PEER_THREADS - 1) / PEER_THREADS;
  const long long cap = (long long)sm_count() * 8;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  peer_finish_kernel<<<(unsigned)blocks, PEER_THREADS, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const unsigned*>(flags_local), world, s, step, g, n, adv, err, peer_timeout_ns());
  B200_CHECK_CUDA(cudaGetLastError());
  return 0;
}
