// Shared constants and helpers for libddp_step.so (sm_100a).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "../../include/ddp_step.h"

namespace ddp {

// ---- geometry of Net (ref: horovod/mnist_horovod.py:9-25) -------------------
constexpr int IMG = 28;
constexpr int IMG_PIX = 784;
constexpr int C1 = 10;        // conv1 out channels
constexpr int P1 = 12;        // pooled conv1 side
constexpr int A1_PER = 1440;  // 10*12*12 pooled+ReLU activations per sample
constexpr int C2 = 20;        // conv2 out channels
constexpr int O2 = 8;         // conv2 output side
constexpr int A2_PER = 320;   // 20*4*4 = fc1 fan-in
constexpr int F1 = 50;        // fc1 out
constexpr int NC = 10;        // classes

// ---- arena offsets (floats); every tensor padded to 4 floats ----------------
constexpr int OFF_C1W = 0;      // 250
constexpr int OFF_C1B = 252;    // 10
constexpr int OFF_C2W = 264;    // 5000
constexpr int OFF_C2B = 5264;   // 20
constexpr int OFF_F1W = 5284;   // 16000
constexpr int OFF_F1B = 21284;  // 50
constexpr int OFF_F2W = 21336;  // 500
constexpr int OFF_F2B = 21836;  // 10
constexpr int P_ARENA = DDP_NET_ARENA_ELEMS;

// ---- saved-activation workspace layout (bytes, all 256-aligned) -------------
struct Workspace {
  float* a1;      // [B,10,12,12] relu(pool(conv1))
  uint8_t* idx1;  // [B,10,12,12] argmax in 2x2 window (0..3)
  float* a2;      // [B,320]      relu(pool(drop2d(conv2)))
  uint8_t* idx2;  // [B,320]
  float* h;       // [B,50]       dropout(relu(fc1))  (= fc2 input)
};
__host__ __device__ inline int64_t align256(int64_t v) { return (v + 255) & ~int64_t(255); }
inline int64_t workspace_bytes(int64_t B) {
  return align256(B * A1_PER * 4) + align256(B * A1_PER) + align256(B * A2_PER * 4) +
         align256(B * A2_PER) + align256(B * F1 * 4);
}
inline Workspace carve_workspace(void* base, int64_t B) {
  char* p = static_cast<char*>(base);
  Workspace w;
  w.a1 = reinterpret_cast<float*>(p);   p += align256(B * A1_PER * 4);
  w.idx1 = reinterpret_cast<uint8_t*>(p); p += align256(B * A1_PER);
  w.a2 = reinterpret_cast<float*>(p);   p += align256(B * A2_PER * 4);
  w.idx2 = reinterpret_cast<uint8_t*>(p); p += align256(B * A2_PER);
  w.h = reinterpret_cast<float*>(p);
  return w;
}

// ---- scratch layout -----------------------------------------------------------
// [0,256)            : counters (u32 ticket for the loss reduction)
// loss_part          : MAX_PARTS floats
// partials           : MAX_PARTS x P_ARENA floats (per-CTA partial weight gradients)
// da2                : [B,320] f32  dL/d(a2)
// dp1                : [B,1440] f32 dL/d(pooled conv1, pre-ReLU)
constexpr int MAX_PARTS = 592;  // up to 4 CTAs per SM on 148 SMs
struct Scratch {
  unsigned int* counters;
  float* loss_part;
  float* partials;
  float* da2;
  float* dp1;
};
inline int64_t scratch_bytes(int64_t B) {
  return 256 + align256(MAX_PARTS * 4) + align256(int64_t(MAX_PARTS) * P_ARENA * 4) +
         align256(B * A2_PER * 4) + align256(B * A1_PER * 4);
}
inline Scratch carve_scratch(void* base, int64_t B) {
  char* p = static_cast<char*>(base);
  Scratch s;
  s.counters = reinterpret_cast<unsigned int*>(p); p += 256;
  s.loss_part = reinterpret_cast<float*>(p); p += align256(MAX_PARTS * 4);
  s.partials = reinterpret_cast<float*>(p);  p += align256(int64_t(MAX_PARTS) * P_ARENA * 4);
  s.da2 = reinterpret_cast<float*>(p);       p += align256(B * A2_PER * 4);
  s.dp1 = reinterpret_cast<float*>(p);
  return s;
}

// ---- error plumbing -----------------------------------------------------------
void set_error(const char* fmt, ...);
int cuda_rc(cudaError_t e, const char* what);
#define DDP_CUDA(call)                                   \
  do {                                                   \
    cudaError_t e__ = (call);                            \
    if (e__ != cudaSuccess) return ::ddp::cuda_rc(e__, #call); \
  } while (0)
#define DDP_REQUIRE(cond, msg)                           \
  do {                                                   \
    if (!(cond)) { ::ddp::set_error("%s: %s", __func__, msg); return DDP_EINVAL; } \
  } while (0)
inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }
int sm_count();
// tensor-core conv2 forward (conv2_fwd_tc.cu)
int launch_conv2_fwd_tc(const float* params, const float* a1, const float* mask2d, uint64_t seed, uint64_t offset,
                        const int64_t* offset_dev, int training, float* a2, uint8_t* idx2, int B, cudaStream_t stream);
// launch helpers shared by ddp_net_forward / ddp_net_backward and the fused training entry point (head_fused.cu)
int launch_fwd_convs(const float* params, const float* x, const float* mask2d, uint64_t seed, uint64_t offset,
                     const int64_t* offset_dev, int training, const Workspace& ws, int B, cudaStream_t stream);
// Optional tail of the backward: instead of reducing the partial gradient rows into the bucket, reduce them straight into
// the NVLink peer-memory exchange (peer.cu) -- the fc / conv2 part while conv1_wgrad still runs -- and finish with the
// rank-ordered sum + optimizer update.  `peer` is a ddp_peer_create context.
struct PeerExchange {
  void* peer;
  int opt;
  float* params;
  float* state1;
  float* state2;
  const ddp_hyper_t* hyper;
  int64_t step;
  const int64_t* step_dev;
};
int launch_peer_early(const PeerExchange& px, const float* partials, int parts_c2, int parts_head, cudaStream_t stream);
int launch_peer_final(const PeerExchange& px, const float* partials, int parts_c1, bool have_early, int parts_c2, int parts_head,
                      float* grad_bucket, cudaStream_t stream);
int launch_bwd_after_head(const float* params, const float* x, int training, int B, const Workspace& ws, const Scratch& sc,
                          int parts_head, float* grad_bucket, cudaStream_t stream, const PeerExchange* px = nullptr);
int launch_head_fwd(const float* params, const Workspace& ws, const int64_t* target, const float* mask1d, uint64_t seed,
                    uint64_t offset, const int64_t* offset_dev, int training, float* out_logp, float* out_loss,
                    int64_t* out_argmax, int64_t* out_correct, const Scratch& sc, int B, cudaStream_t stream);
int launch_head_bwd(const float* params, const Workspace& ws, const float* logp, const float* dlogp, const int64_t* target,
                    float nll_coef, float drop_scale, const Scratch& sc, int B, int* parts_head, cudaStream_t stream);
// sparse conv2 backward (net_bwd_sparse.cu): one visit per live (channel, pool cell) entry instead of dense 8x8 planes
int launch_conv2_wgrad_sp(const float* da2, const float* a2, const uint8_t* idx2, const float* a1, float drop_scale,
                          float* partials, int B, int* parts_out, cudaStream_t stream, bool pdl);
int launch_conv2_dgrad_sp(const float* params, const float* da2, const float* a2, const uint8_t* idx2, const float* a1,
                          float drop_scale, float* dp1, int B, cudaStream_t stream);
int launch_conv2_wgrad_tc(const float* da2, const float* a2, const uint8_t* idx2, const float* a1, float drop_scale,
                          float* partials, int B, int* parts_out, cudaStream_t stream, bool pdl);   // conv2_wgrad_tc.cu
// conv1 weight gradient on packed channel pairs (conv1_wgrad_col.cu); `parts` CTAs, one partial row each
int launch_conv1_wgrad_col(const float* dp1, const uint8_t* idx1, const float* x, float* partials, int B, int parts,
                          cudaStream_t stream, bool pdl);
bool wgrad2_tc();                        // api.cu: true when DDP_W2_TC=1 (A/B switch: the tcgen05 conv2 weight gradient; default is the sparse FFMA2 kernel, which is faster)
bool bwd_dense();                        // api.cu: true when DDP_BWD_DENSE=1 (A/B switch: the dense conv2 backward kernels)
// per-kernel event timing (api.cu); mark(i) records event i on `stream` when profiling is on
void prof_mark(int idx, cudaStream_t stream);
bool prof_enabled();
// fork: records an event on `main`, makes the library's side stream wait for it, returns the side stream and the
// event the caller must record on it and wait for on `main` to join (api.cu)
int side_fork(cudaStream_t main, cudaStream_t* side, cudaEvent_t* join);

// Warm the SM's L1 with a read-only global array that every CTA on the SM will stream through
// (fc1.weight, 64 KB): all lines are requested at once instead of being missed one loop iteration at a time.
__device__ __forceinline__ void prefetch_l1_range(const float* base, int nfloats, int tid, int nthreads) {
  for (int i = tid * 32; i < nfloats; i += nthreads * 32)
    asm volatile("prefetch.global.L1 [%0];" :: "l"(base + i));
}

// ---- programmatic dependent launch (PDL) ---------------------------------------------------------------------------
// The step is a chain of ~9 kernels, each with a prologue that only touches parameters and shared memory (weight
// staging, barrier / TMEM set-up).  Launched with the programmatic-serialization attribute, kernel N+1 may start that
// prologue while kernel N is still running; pdl_wait() (griddepcontrol.wait) then blocks until kernel N has completed
// and its writes are visible.  RULE: nothing written by an earlier kernel of the step may be read, and nothing an
// earlier kernel reads may be written, before pdl_wait().  pdl_trigger() lets the next kernel be scheduled as soon as
// every CTA of this one has issued it.  Both are no-ops in a kernel launched the ordinary way.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
bool pdl_enabled();                      // api.cu: false when DDP_NO_PDL=1 (A/B switch for measurements)
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream, bool pdl,
                              Args&&... args) {
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = (pdl && pdl_enabled()) ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...);
}

// ---- mbarrier + bulk-copy (TMA engine, 1-D) helpers -----------------------------------------------------
// cp.async.bulk moves a contiguous block global -> shared without touching registers and signals an
// mbarrier with the byte count (complete_tx); one thread issues, everybody waits on the barrier.
__device__ __forceinline__ uint32_t smem_addr_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbarrier_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(smem_addr_u32(bar)), "r"(count));
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbarrier_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("{\n\t.reg .b64 st;\n\tmbarrier.arrive.expect_tx.shared::cta.b64 st, [%0], %1;\n\t}"
               :: "r"(smem_addr_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_copy_g2s(void* dst_smem, const void* src_gmem, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               :: "r"(smem_addr_u32(dst_smem)), "l"(src_gmem), "r"(bytes), "r"(smem_addr_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbarrier_wait(uint64_t* bar, uint32_t parity) {
  uint32_t done = 0, spins = 0;
  while (!done) {
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.b32 %0, 1, 0, p;\n\t}"
                 : "=r"(done) : "r"(smem_addr_u32(bar)), "r"(parity) : "memory");
    if (++spins > (1u << 26)) __trap();
  }
}

// ---- packed fp32 pairs: fma.rn.f32x2 does two IEEE fp32 FMAs in ONE issue slot (same FMA pipe throughput) -------------
__device__ __forceinline__ void fma2(unsigned long long& d, unsigned long long a, unsigned long long b) {
  asm("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(d) : "l"(a), "l"(b));
}
__device__ __forceinline__ unsigned long long pack2(float a, float b) {
  return (unsigned long long)__float_as_uint(a) | ((unsigned long long)__float_as_uint(b) << 32);
}
__device__ __forceinline__ float lo_f(unsigned long long v) { return __uint_as_float(unsigned(v)); }
__device__ __forceinline__ float hi_f(unsigned long long v) { return __uint_as_float(unsigned(v >> 32)); }

// ---- Philox4x32-10 (counter-based; same constants as Random123 / curand) ------
struct Philox4 { uint32_t x, y, z, w; };
__host__ __device__ inline Philox4 philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3,
                                                 uint32_t k0, uint32_t k1) {
  const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    uint64_t p0 = uint64_t(M0) * c0;
    uint64_t p1 = uint64_t(M1) * c2;
    uint32_t hi0 = uint32_t(p0 >> 32), lo0 = uint32_t(p0);
    uint32_t hi1 = uint32_t(p1 >> 32), lo1 = uint32_t(p1);
    uint32_t n0 = hi1 ^ c1 ^ k0;
    uint32_t n2 = hi0 ^ c3 ^ k1;
    c0 = n0; c1 = lo1; c2 = n2; c3 = lo0;
    k0 += W0; k1 += W1;
  }
  return Philox4{c0, c1, c2, c3};
}
// Bernoulli(0.5) keep decision for dropout element `elem` of stream `which`.
__device__ inline float philox_keep(uint64_t seed, uint64_t offset, uint32_t which, uint64_t elem) {
  Philox4 r = philox4x32_10(uint32_t(elem), uint32_t(elem >> 32), uint32_t(offset), which ^ (uint32_t(offset >> 32) << 8),
                            uint32_t(seed), uint32_t(seed >> 32));
  return (r.x & 0x80000000u) ? 0.f : 1.f;
}

}  // namespace ddp
