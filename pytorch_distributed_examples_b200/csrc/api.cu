// Misc entry points of libddp_step.so: versioning, error text, device binding, layout
// queries and the elastic-state snapshot copies.
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <mutex>
#include "common.cuh"

namespace ddp {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int cuda_rc(cudaError_t e, const char* what) {
  set_error("%s: CUDA error %d: %s", what, int(e), cudaGetErrorString(e));
  return int(e) > 0 ? int(e) : DDP_EINVAL;
}

bool pdl_enabled() {
  static int v = -1;
  if (v < 0) { const char* e = getenv("DDP_NO_PDL"); v = (e && e[0] == '1') ? 0 : 1; }
  return v == 1;
}

bool wgrad2_tc() {
  static int v = -1;
  if (v < 0) { const char* e = getenv("DDP_W2_TC"); v = (e && e[0] == '1') ? 1 : 0; }
  return v == 1;
}

bool bwd_dense() {
  static int v = -1;
  if (v < 0) { const char* e = getenv("DDP_BWD_DENSE"); v = (e && e[0] == '1') ? 1 : 0; }
  return v == 1;
}

int sm_count() {
  static int cached[64] = {0};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  if (cached[dev] == 0) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
    cached[dev] = n;
  }
  return cached[dev];
}

// events: 0 fwd start, 1 conv1_fwd, 2 conv2_fwd, 3 head_fwd, 4 bwd start, 5 head_bwd, 6 dgrad, 7 wgrad2, 8 wgrad1, 9 reduce
static bool g_prof_on = false;
static cudaEvent_t g_prof_ev[10];
static bool g_prof_made = false;
void prof_mark(int idx, cudaStream_t stream) {
  if (!g_prof_on) return;
  if (!g_prof_made) {
    for (auto& e : g_prof_ev) cudaEventCreate(&e);
    g_prof_made = true;
  }
  cudaEventRecord(g_prof_ev[idx], stream);
}

bool prof_enabled() { return g_prof_on; }

// Side stream + event pairs for the small-batch fork in ddp_net_backward (conv2 weight gradient beside the data
// gradient).  Library-owned, created on first use per device (outside any stream capture: the host layer runs eager
// warm-up steps before it captures); a ring of event pairs under a mutex keeps record/wait pairs of concurrent callers apart.
static std::mutex g_side_mu;
static cudaStream_t g_side_stream[64] = {nullptr};
static cudaEvent_t g_side_ev[64][8][2];
static unsigned g_side_next[64] = {0};
int side_fork(cudaStream_t main, cudaStream_t* side, cudaEvent_t* join) {
  int dev = 0;
  DDP_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64) { set_error("side_fork: device index out of range"); return DDP_EINVAL; }
  std::lock_guard<std::mutex> lock(g_side_mu);
  if (!g_side_stream[dev]) {
    cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
    DDP_CUDA(cudaStreamIsCapturing(main, &cap));
    if (cap != cudaStreamCaptureStatusNone) { *side = nullptr; *join = nullptr; return 0; }   // cannot create resources now: stay serial
    DDP_CUDA(cudaStreamCreateWithFlags(&g_side_stream[dev], cudaStreamNonBlocking));
    for (int i = 0; i < 8; ++i)
      for (int j = 0; j < 2; ++j) DDP_CUDA(cudaEventCreateWithFlags(&g_side_ev[dev][i][j], cudaEventDisableTiming));
  }
  const unsigned slot = g_side_next[dev]++ & 7u;
  DDP_CUDA(cudaEventRecord(g_side_ev[dev][slot][0], main));
  DDP_CUDA(cudaStreamWaitEvent(g_side_stream[dev], g_side_ev[dev][slot][0], 0));
  *side = g_side_stream[dev];
  *join = g_side_ev[dev][slot][1];
  return 0;
}

}  // namespace ddp

using namespace ddp;

extern "C" int ddp_prof_enable(int on) { g_prof_on = on != 0; return 0; }
extern "C" int ddp_prof_read(float* ms_out) {
  DDP_REQUIRE(ms_out, "ms_out is NULL");
  DDP_REQUIRE(g_prof_made, "no profiled launch yet");
  DDP_CUDA(cudaEventSynchronize(g_prof_ev[9]));
  const int pairs[DDP_PROF_KERNELS][2] = {{0, 1}, {1, 2}, {2, 3}, {4, 5}, {5, 6}, {6, 7}, {7, 8}, {8, 9}};
  for (int i = 0; i < DDP_PROF_KERNELS; ++i)
    DDP_CUDA(cudaEventElapsedTime(&ms_out[i], g_prof_ev[pairs[i][0]], g_prof_ev[pairs[i][1]]));
  return 0;
}

extern "C" int ddp_abi_version(void) { return DDP_ABI_VERSION; }
extern "C" const char* ddp_last_error(void) { return g_err; }

extern "C" int ddp_set_device(int device) {
  DDP_CUDA(cudaSetDevice(device));
  return 0;
}

extern "C" int ddp_net_param_layout(int64_t* offsets, int64_t* sizes) {
  DDP_REQUIRE(offsets && sizes, "null output");
  const int64_t off[DDP_NET_NUM_TENSORS] = {OFF_C1W, OFF_C1B, OFF_C2W, OFF_C2B, OFF_F1W, OFF_F1B, OFF_F2W, OFF_F2B};
  const int64_t sz[DDP_NET_NUM_TENSORS] = {250, 10, 5000, 20, 16000, 50, 500, 10};
  for (int i = 0; i < DDP_NET_NUM_TENSORS; ++i) { offsets[i] = off[i]; sizes[i] = sz[i]; }
  return 0;
}

extern "C" int64_t ddp_net_workspace_bytes(int64_t batch) { return batch < 1 ? -1 : workspace_bytes(batch); }
extern "C" int64_t ddp_net_scratch_bytes(int64_t batch) { return batch < 1 ? -1 : scratch_bytes(batch); }

extern "C" int ddp_state_snapshot(const void* device_arena, void* pinned_host, int64_t nbytes, void* stream) {
  DDP_REQUIRE(nbytes >= 0 && (nbytes == 0 || (device_arena && pinned_host)), "bad arguments");
  if (nbytes == 0) return 0;
  DDP_CUDA(cudaMemcpyAsync(pinned_host, device_arena, size_t(nbytes), cudaMemcpyDeviceToHost,
                           static_cast<cudaStream_t>(stream)));
  return 0;
}

extern "C" int ddp_state_restore(void* device_arena, const void* pinned_host, int64_t nbytes, void* stream) {
  DDP_REQUIRE(nbytes >= 0 && (nbytes == 0 || (device_arena && pinned_host)), "bad arguments");
  if (nbytes == 0) return 0;
  DDP_CUDA(cudaMemcpyAsync(device_arena, pinned_host, size_t(nbytes), cudaMemcpyHostToDevice,
                           static_cast<cudaStream_t>(stream)));
  return 0;
}
