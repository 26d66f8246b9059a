// abi.cu — error text, device queries and the stream / event / graph plumbing of the C ABI.
#include <stdarg.h>
#include <stdlib.h>
#include <string.h>

#include "common.cuh"

namespace b200 {

static thread_local char g_err[1024] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int sm_count_physical() {
  static int cached = 0;
  if (cached > 0) return cached;
  int dev = 0, n = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 148;
  if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) return 148;
  cached = n;
  return n;
}

// SMs the persistent grids are sized for: the physical count unless b200_set_sm_budget lowered it.
static int g_sm_budget = 0;
int sm_count() {
  const int n = sm_count_physical();
  return (g_sm_budget > 0 && g_sm_budget < n) ? g_sm_budget : n;
}

static uint32_t* g_status_word = nullptr;
uint32_t* status_word() { return g_status_word; }

bool pdl_enabled() {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("B200STEP_PDL");
    v = (e && e[0] == '0') ? 0 : 1;
  }
  return v == 1;
}

}  // namespace b200

using namespace b200;

extern "C" const char* b200_last_error(void) { return g_err; }
extern "C" int b200_abi_version(void) { return 1; }

extern "C" int b200_device_sm_count(int* out_host) {
  B200_REQUIRE(out_host != nullptr, "out_host is NULL");
  int dev = 0, n = 0;
  B200_CHECK_CUDA(cudaGetDevice(&dev));
  B200_CHECK_CUDA(cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev));
  *out_host = n;
  return 0;
}

extern "C" int b200_set_sm_budget(int sms) {
  B200_REQUIRE(sms >= 0, "SM budget must be >= 0 (0 = every SM of the device)");
  g_sm_budget = (sms > 0 && sms < 2) ? 2 : sms;  // a CTA pair needs two
  return 0;
}

extern "C" int b200_set_status_word(uint32_t* device_word) {
  g_status_word = device_word;
  return 0;
}

extern "C" int b200_zero(void* ptr, int64_t bytes, b200_stream_t stream) {
  B200_REQUIRE(ptr != nullptr && bytes >= 0, "bad arguments to b200_zero");
  if (bytes == 0) return 0;
  B200_CHECK_CUDA(cudaMemsetAsync(ptr, 0, (size_t)bytes, static_cast<cudaStream_t>(stream)));
  return 0;
}

extern "C" int b200_stream_create(b200_stream_t* out_host, int high_priority) {
  B200_REQUIRE(out_host != nullptr, "out_host is NULL");
  int lo = 0, hi = 0;
  B200_CHECK_CUDA(cudaDeviceGetStreamPriorityRange(&lo, &hi));
  cudaStream_t s;
  B200_CHECK_CUDA(cudaStreamCreateWithPriority(&s, cudaStreamNonBlocking, high_priority ? hi : lo));
  *out_host = s;
  return 0;
}
extern "C" int b200_stream_destroy(b200_stream_t s) {
  B200_CHECK_CUDA(cudaStreamDestroy(static_cast<cudaStream_t>(s)));
  return 0;
}
extern "C" int b200_stream_sync(b200_stream_t s) {
  B200_CHECK_CUDA(cudaStreamSynchronize(static_cast<cudaStream_t>(s)));
  return 0;
}
extern "C" int b200_event_create(void** out_host, int timing) {
  B200_REQUIRE(out_host != nullptr, "out_host is NULL");
  cudaEvent_t e;
  B200_CHECK_CUDA(cudaEventCreateWithFlags(&e, timing ? cudaEventDefault : cudaEventDisableTiming));
  *out_host = e;
  return 0;
}
extern "C" int b200_event_destroy(void* ev) {
  B200_CHECK_CUDA(cudaEventDestroy(static_cast<cudaEvent_t>(ev)));
  return 0;
}
extern "C" int b200_event_record(void* ev, b200_stream_t s) {
  B200_CHECK_CUDA(cudaEventRecord(static_cast<cudaEvent_t>(ev), static_cast<cudaStream_t>(s)));
  return 0;
}
extern "C" int b200_stream_wait_event(b200_stream_t s, void* ev) {
  B200_CHECK_CUDA(cudaStreamWaitEvent(static_cast<cudaStream_t>(s), static_cast<cudaEvent_t>(ev), 0));
  return 0;
}
extern "C" int b200_event_elapsed_ms(void* start, void* stop, float* out_host) {
  B200_REQUIRE(out_host != nullptr, "out_host is NULL");
  B200_CHECK_CUDA(cudaEventSynchronize(static_cast<cudaEvent_t>(stop)));
  B200_CHECK_CUDA(cudaEventElapsedTime(out_host, static_cast<cudaEvent_t>(start), static_cast<cudaEvent_t>(stop)));
  return 0;
}

extern "C" int b200_graph_begin(b200_stream_t s) {
  B200_CHECK_CUDA(cudaStreamBeginCapture(static_cast<cudaStream_t>(s), cudaStreamCaptureModeRelaxed));
  return 0;
}
extern "C" int b200_graph_end(b200_stream_t s, void** out_exec_host) {
  B200_REQUIRE(out_exec_host != nullptr, "out_exec_host is NULL");
  cudaGraph_t graph = nullptr;
  B200_CHECK_CUDA(cudaStreamEndCapture(static_cast<cudaStream_t>(s), &graph));
  cudaGraphExec_t exec = nullptr;
  cudaError_t e = cudaGraphInstantiate(&exec, graph, 0);
  cudaGraphDestroy(graph);
  if (e != cudaSuccess) {
    set_error("cudaGraphInstantiate failed: %s", cudaGetErrorString(e));
    return 1;
  }
  *out_exec_host = exec;
  return 0;
}
extern "C" int b200_graph_launch(void* exec, b200_stream_t s) {
  B200_CHECK_CUDA(cudaGraphLaunch(static_cast<cudaGraphExec_t>(exec), static_cast<cudaStream_t>(s)));
  return 0;
}
extern "C" int b200_graph_destroy(void* exec) {
  B200_CHECK_CUDA(cudaGraphExecDestroy(static_cast<cudaGraphExec_t>(exec)));
  return 0;
}
