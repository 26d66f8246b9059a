// comm.cu — K-comm: the gradient all-reduce of the data-parallel step, NCCL over NVLink 5 / NVSwitch.
//
// Replaces what DistributedDataParallel does inside loss.backward() (task.py:189/194, task.py:311):
// per-bucket all-reduce of the gradients, result = mean over ranks.  Here the 1/world factor is folded
// into the loss-gradient scale, so the collective is a plain SUM, in place, on a contiguous bucket of the
// flat fp32 gradient buffer, issued on whatever (side) stream the host passes.
//
// NCCL is bound at run time from the libnccl.so.2 that PyTorch has already loaded into the process
// (same version torch.distributed uses), so libb200step.so has no link-time NCCL dependency.
#include <dlfcn.h>
#include <nccl.h>
#include <stdlib.h>
#include <string.h>

#include "common.cuh"

namespace b200 {

struct NcclApi {
  void* handle = nullptr;
  ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*Broadcast)(const void*, void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  const char* (*GetErrorString)(ncclResult_t) = nullptr;
  // user-buffer registration (NCCL >= 2.19; optional: resolved lazily, absent symbols just disable the feature)
  ncclResult_t (*MemAlloc)(void**, size_t) = nullptr;
  ncclResult_t (*MemFree)(void*) = nullptr;
  ncclResult_t (*CommRegister)(const ncclComm_t, void*, size_t, void**) = nullptr;
  ncclResult_t (*CommDeregister)(const ncclComm_t, void*) = nullptr;
  // symmetric-memory windows (NCCL >= 2.27): collectives on a window use NCCL's low-latency symmetric kernels
  ncclResult_t (*CommWindowRegister)(ncclComm_t, void*, size_t, ncclWindow_t*, int) = nullptr;
  ncclResult_t (*CommWindowDeregister)(ncclComm_t, ncclWindow_t) = nullptr;
};

static NcclApi g_nccl;
static ncclComm_t g_comm = nullptr;
static int g_world = 0;
constexpr int MAX_REG = 16;
static void* g_reg_handles[MAX_REG];
static int g_num_reg = 0;
static ncclWindow_t g_windows[MAX_REG];
static int g_num_win = 0;

static int load_nccl() {
  if (g_nccl.handle) return 0;
  const char* env = getenv("B200STEP_NCCL_LIB");
  void* h = nullptr;
  if (env && *env) h = dlopen(env, RTLD_NOW | RTLD_GLOBAL);
  if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD | RTLD_GLOBAL);  // the copy torch loaded
  if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
  B200_REQUIRE(h != nullptr, "cannot load libnccl.so.2 (import torch first or set B200STEP_NCCL_LIB): %s", dlerror());
#define B200_NCCL_SYM(field, name)                                              \
  *reinterpret_cast<void**>(&g_nccl.field) = dlsym(h, name);                    \
  B200_REQUIRE(g_nccl.field != nullptr, "libnccl.so.2 lacks symbol %s", name)
  B200_NCCL_SYM(GetUniqueId, "ncclGetUniqueId");
  B200_NCCL_SYM(CommInitRank, "ncclCommInitRank");
  B200_NCCL_SYM(CommDestroy, "ncclCommDestroy");
  B200_NCCL_SYM(AllReduce, "ncclAllReduce");
  B200_NCCL_SYM(Broadcast, "ncclBroadcast");
  B200_NCCL_SYM(GetErrorString, "ncclGetErrorString");
#undef B200_NCCL_SYM
  *reinterpret_cast<void**>(&g_nccl.MemAlloc) = dlsym(h, "ncclMemAlloc");
  *reinterpret_cast<void**>(&g_nccl.MemFree) = dlsym(h, "ncclMemFree");
  *reinterpret_cast<void**>(&g_nccl.CommRegister) = dlsym(h, "ncclCommRegister");
  *reinterpret_cast<void**>(&g_nccl.CommDeregister) = dlsym(h, "ncclCommDeregister");
  *reinterpret_cast<void**>(&g_nccl.CommWindowRegister) = dlsym(h, "ncclCommWindowRegister");
  *reinterpret_cast<void**>(&g_nccl.CommWindowDeregister) = dlsym(h, "ncclCommWindowDeregister");
  g_nccl.handle = h;
  return 0;
}

#define B200_CHECK_NCCL(expr)                                                                    \
  do {                                                                                           \
    ncclResult_t _r = (expr);                                                                    \
    if (_r != ncclSuccess) {                                                                     \
      ::b200::set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #expr, g_nccl.GetErrorString(_r)); \
      return 3;                                                                                  \
    }                                                                                            \
  } while (0)

}  // namespace b200

using namespace b200;

extern "C" int b200_comm_unique_id(void* out_host_128) {
  B200_REQUIRE(out_host_128 != nullptr, "out buffer is NULL");
  if (int r = load_nccl()) return r;
  static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId is expected to be 128 bytes");
  ncclUniqueId id;
  B200_CHECK_NCCL(g_nccl.GetUniqueId(&id));
  memcpy(out_host_128, &id, sizeof(id));
  return 0;
}

extern "C" int b200_comm_init(const void* unique_id_host_128, int rank, int world) {
  B200_REQUIRE(unique_id_host_128 != nullptr, "unique id is NULL");
  B200_REQUIRE(world >= 1 && rank >= 0 && rank < world, "bad rank %d / world %d", rank, world);
  B200_REQUIRE(g_comm == nullptr, "communicator already initialised");
  if (int r = load_nccl()) return r;
  ncclUniqueId id;
  memcpy(&id, unique_id_host_128, sizeof(id));
  B200_CHECK_NCCL(g_nccl.CommInitRank(&g_comm, world, id, rank));
  g_world = world;
  return 0;
}

extern "C" int b200_comm_allreduce_sum(float* buf, int64_t count, b200_stream_t stream) {
  B200_REQUIRE(g_comm != nullptr, "b200_comm_init has not been called");
  B200_REQUIRE(count > 0, "count must be positive");
  B200_CHECK_NCCL(g_nccl.AllReduce(buf, buf, (size_t)count, ncclFloat32, ncclSum, g_comm,
                                   static_cast<cudaStream_t>(stream)));
  return 0;
}

extern "C" int b200_comm_broadcast(float* buf, int64_t count, int root, b200_stream_t stream) {
  B200_REQUIRE(g_comm != nullptr, "b200_comm_init has not been called");
  B200_REQUIRE(count > 0 && root >= 0 && root < g_world, "bad broadcast arguments");
  B200_CHECK_NCCL(g_nccl.Broadcast(buf, buf, (size_t)count, ncclFloat32, root, g_comm,
                                   static_cast<cudaStream_t>(stream)));
  return 0;
}

// Memory NCCL can use in place (ncclMemAlloc: a cuMem allocation that NVSwitch multicast / in-switch reduction can
// map) and its registration with the communicator: an all-reduce on a registered buffer skips NCCL's staging copies
// and, on an NVSwitch node, may run as an NVLS (in-network) reduction.
extern "C" int b200_comm_mem_alloc(int64_t bytes, void** out_host) {
  B200_REQUIRE(out_host != nullptr && bytes > 0, "bad arguments to b200_comm_mem_alloc");
  if (int r = load_nccl()) return r;
  B200_REQUIRE(g_nccl.MemAlloc != nullptr, "this libnccl has no ncclMemAlloc");
  B200_CHECK_NCCL(g_nccl.MemAlloc(out_host, (size_t)bytes));
  return 0;
}

extern "C" int b200_comm_mem_free(void* ptr) {
  if (ptr == nullptr) return 0;
  B200_REQUIRE(g_nccl.MemFree != nullptr, "this libnccl has no ncclMemFree");
  B200_CHECK_NCCL(g_nccl.MemFree(ptr));
  return 0;
}

extern "C" int b200_comm_register(void* buf, int64_t bytes) {
  B200_REQUIRE(g_comm != nullptr, "b200_comm_init has not been called");
  B200_REQUIRE(buf != nullptr && bytes > 0, "bad arguments to b200_comm_register");
  B200_REQUIRE(g_nccl.CommRegister != nullptr, "this libnccl has no ncclCommRegister");
  B200_REQUIRE(g_num_reg < MAX_REG, "too many registered buffers");
  B200_CHECK_NCCL(g_nccl.CommRegister(g_comm, buf, (size_t)bytes, &g_reg_handles[g_num_reg]));
  ++g_num_reg;
  // Also as a symmetric window (a collective call: every rank registers its buffer of the same size, in the same
  // order): small all-reduces — T0's gradient buckets are 1-2 MB — then take NCCL's symmetric low-latency path.
  // Only for small buffers: measured on 8 B200s, T0 (5.65 MB of gradients) 0.560 -> 0.555 ms/step, but T1's 460 MB
  // buffer 3.37 -> 5.64 ms/step (the symmetric kernels lose to the ring / NVLS path at that size).
  // B200STEP_NCCL_WINDOW=0 skips it, =1 forces it; a library without the call, or a refusal, leaves the plain
  // registration.
  const char* e = getenv("B200STEP_NCCL_WINDOW");
  const bool want = e ? (e[0] != '0') : (bytes <= (32ll << 20));
  if (g_nccl.CommWindowRegister != nullptr && want && g_num_win < MAX_REG) {
    ncclWindow_t win = nullptr;
    if (g_nccl.CommWindowRegister(g_comm, buf, (size_t)bytes, &win, NCCL_WIN_COLL_SYMMETRIC) == ncclSuccess && win)
      g_windows[g_num_win++] = win;
  }
  return 0;
}

extern "C" int b200_comm_destroy(void) {
  if (g_comm) {
    for (int i = 0; i < g_num_win; ++i)
      if (g_nccl.CommWindowDeregister) g_nccl.CommWindowDeregister(g_comm, g_windows[i]);
    g_num_win = 0;
    for (int i = 0; i < g_num_reg; ++i)
      if (g_nccl.CommDeregister) g_nccl.CommDeregister(g_comm, g_reg_handles[i]);
    g_num_reg = 0;
    B200_CHECK_NCCL(g_nccl.CommDestroy(g_comm));
    g_comm = nullptr;
    g_world = 0;
  }
  return 0;
}
