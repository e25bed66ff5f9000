// One-shot allreduce over NVLink peer memory, fused with the (1/N)*g + optimizer update.
//
// The CNN's gradient bucket is 87 KB: ncclAllReduce on it is pure latency (~45 us at 8 ranks on NVSwitch), an eighth
// of the whole training step.  Here every rank PUSHES its bucket into a slot of every peer's receive buffer with plain
// 128-bit stores through the NVLink-mapped (cudaIpc) address, publishes a per-source epoch flag with a system-scope
// release, waits for the W flags addressed to it, then sums the W slots in RANK ORDER (so all replicas stay
// bit-identical) and applies the optimizer rule — one kernel, no intermediate round trip, ~10 us.
// Replaces, for small buckets on one node, the allreduce DDP / Horovod issue per step
// (ref: pytorch_elastic/mnist_ddp_elastic.py:58 ; horovod/mnist_horovod.py:53,64).
//
// Protocol per launch (epoch e, parity p = e & 1 selects one of two slot sets, so a fast rank's next step cannot
// overwrite data a slow rank is still reading):
//   1. all CTAs: copy bucket -> slot[p][my_rank] on every rank (own copy included); __threadfence_system()
//   2. last CTA to finish (device-local ticket): st.release.sys flag[p][my_rank] = e on every rank; epoch <- e
//   3. all CTAs: spin (ld.acquire.sys) until flag[p][src] == e for every src
//   4. all CTAs: g = sum_src slot[p][src][i] in rank order; fused update
// All CTAs of the grid must be co-resident (they wait on each other): grid <= 16 CTAs.  The kernels of different
// ranks wait on one another across GPUs, never on the same GPU.  A spin cap traps instead of hanging.
// The receive buffers are communicator resources (like NCCL's own): cudaMalloc'ed once by ddp_peer_create,
// exported with cudaIpcGetMemHandle, exchanged by the host layer through the rendezvous store.
#include <string.h>
#include "common.cuh"
#include "update_rules.cuh"

namespace ddp {

constexpr int PEER_MAX_WORLD = 8;
constexpr int PEER_FLAG_STRIDE = 32;   // uint32 per flag slot (128 B apart)
constexpr int PEER_CTAS = 8;
constexpr int PEER_THREADS = 512;

struct PeerCtx {
  int world, rank, device;
  int64_t max_elems;
  char* local_base;                       // [slots: 2*W*max_elems floats][flags: 2*W*PEER_FLAG_STRIDE u32][epoch u32][ticket u32]
  size_t bytes;
  char* peer_base[PEER_MAX_WORLD];        // peer_base[rank] == local_base
  bool opened[PEER_MAX_WORLD];
};

struct PeerDev {                          // passed by value to the kernel
  int world, rank;
  int64_t max_elems;
  float* slots[PEER_MAX_WORLD];           // each: [2][W][max_elems]
  unsigned int* flags[PEER_MAX_WORLD];    // each: [2][W][PEER_FLAG_STRIDE]
  unsigned int* epoch;                    // local
  unsigned int* ticket;                   // local
  unsigned long long* ll[PEER_MAX_WORLD]; // each: [2][W][max_elems] (value, epoch) words of the flag-in-data protocol
};

__device__ __forceinline__ void st_release_sys(unsigned int* p, unsigned int v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" :: "l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned int ld_acquire_sys(const unsigned int* p) {
  unsigned int v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

// MODE 0: allreduce(sum) only, result written to `out`; MODE 1: + fused optimizer update of params/state.
// Where this rank's contribution comes from.  mode 0: `bucket` (n floats).  mode 1: the per-CTA partial gradient rows of
// the backward kernels (net_bwd.cu), summed here in row order -- the separate reduce_partials launch disappears.
// mode 2: like 1, but the fc / conv2 part [OFF_C2W, P_ARENA) was already reduced and pushed by peer_early_kernel while
// conv1_wgrad was running: only conv1's columns are left.
struct PartialSrc {
  int mode;
  const float* partials;
  int parts_c1, parts_c2, parts_head;
};
constexpr int PEER_C1_COLS = OFF_C2W / 4;          // float4 columns of conv1.weight + conv1.bias (incl. padding)
__device__ __forceinline__ bool peer_is_pad(int e) {
  return (e >= 250 && e < OFF_C1B) || (e >= OFF_C1B + 10 && e < OFF_C2W) || (e >= OFF_C2B + 20 && e < OFF_F1W) ||
         (e >= OFF_F1B + 50 && e < OFF_F2W) || (e >= OFF_F2B + 10);
}
__device__ __forceinline__ float4 peer_zero_pads(float4 v, int e) {
  if (peer_is_pad(e)) v.x = 0.f;
  if (peer_is_pad(e + 1)) v.y = 0.f;
  if (peer_is_pad(e + 2)) v.z = 0.f;
  if (peer_is_pad(e + 3)) v.w = 0.f;
  return v;
}

template <int OPT, bool HAS1, bool HAS2, bool UPDATE>
__global__ void __launch_bounds__(PEER_THREADS, 1)
peer_allreduce_kernel(PeerDev pd, const float* __restrict__ bucket, float* __restrict__ out, int64_t n,
                      float* __restrict__ params, float* __restrict__ state1, float* __restrict__ state2,
                      UpdateConsts kc, int64_t step, const int64_t* __restrict__ step_dev, PartialSrc src) {
  __shared__ UpdateConsts ks;
  __shared__ unsigned int epoch_s;
  __shared__ float4 wsum_s[PEER_THREADS / 32];
  const int tid = threadIdx.x;
  if (tid == 0) {
    epoch_s = *((volatile unsigned int*)pd.epoch) + 1u;   // every CTA reads it before the ticket winner bumps it
    if (UPDATE) {
      UpdateConsts k = kc;
      finish_consts(k, OPT, step_dev ? (*step_dev + 1) : step);
      ks = k;
    }
  }
  __syncthreads();
  // launched as a programmatic dependent of the kernel that produces the bucket / the partial rows: everything above read
  // only state of the PREVIOUS exchange.  No launch_dependents here: the next kernel of the step reads the parameters.
  pdl_wait();
  const unsigned int e = epoch_s;
  const int par = int(e & 1u);
  const int W = pd.world, R = pd.rank;
  const int64_t nvec = (n + 3) >> 2;                       // slots are padded to float4; bucket is 16-byte aligned, n%4==0 enforced by host
  const int64_t gtid = int64_t(blockIdx.x) * PEER_THREADS + tid, gthreads = int64_t(gridDim.x) * PEER_THREADS;
  const int64_t slot_off = (int64_t(par) * W + R) * pd.max_elems;

  // 1. push my contribution into slot[par][R] of every rank
  if (src.mode == 0) {
    for (int64_t i = gtid; i < nvec; i += gthreads) {
      const float4 v = reinterpret_cast<const float4*>(bucket)[i];
#pragma unroll
      for (int p = 0; p < PEER_MAX_WORLD; ++p)
        if (p < W) reinterpret_cast<float4*>(pd.slots[p] + slot_off)[i] = v;
    }
  } else if (src.mode == 1) {
    // every column of the arena: 8 lanes share a column (lane j sums the partial rows j, j+8, ..., four loads in flight),
    // then a fixed-order shuffle tree over the 8; a warp covers 4 columns at a time
    const int lane = tid & 31, sub = lane & 7;
    const int64_t gwarp = gtid >> 5, nwarps = gthreads >> 5;
    for (int64_t c0 = gwarp * 4; c0 < nvec; c0 += nwarps * 4) {
      const int64_t i = c0 + (lane >> 3);
      float4 a = make_float4(0.f, 0.f, 0.f, 0.f);
      const bool live = i < nvec;
      const int e = live ? int(i) * 4 : 0;
      if (live) {
        const int parts = e < OFF_C2W ? src.parts_c1 : (e < OFF_F1W ? src.parts_c2 : src.parts_head);
        const float* p0 = src.partials + e;
        for (int r0 = sub; r0 < parts; r0 += 32) {
          float4 v[4];
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            const int r = r0 + 8 * q;
            v[q] = r < parts ? __ldcg(reinterpret_cast<const float4*>(p0 + size_t(r) * P_ARENA)) : make_float4(0.f, 0.f, 0.f, 0.f);
          }
#pragma unroll
          for (int q = 0; q < 4; ++q) { a.x += v[q].x; a.y += v[q].y; a.z += v[q].z; a.w += v[q].w; }
        }
      }
#pragma unroll
      for (int o = 4; o > 0; o >>= 1) {
        a.x += __shfl_xor_sync(0xffffffffu, a.x, o); a.y += __shfl_xor_sync(0xffffffffu, a.y, o);
        a.z += __shfl_xor_sync(0xffffffffu, a.z, o); a.w += __shfl_xor_sync(0xffffffffu, a.w, o);
      }
      if (live && sub == 0) {
        a = peer_zero_pads(a, e);
#pragma unroll
        for (int p = 0; p < PEER_MAX_WORLD; ++p)
          if (p < W) reinterpret_cast<float4*>(pd.slots[p] + slot_off)[i] = a;
      }
    }
  } else {
    // conv1's columns only: the whole CTA sums one column's rows (thread t: rows t, t+512), fixed-order shuffle tree,
    // then the 16 warp sums in order
    for (int c = blockIdx.x; c < PEER_C1_COLS; c += gridDim.x) {
      const float* p0 = src.partials + c * 4;
      float4 a = make_float4(0.f, 0.f, 0.f, 0.f);
      for (int r = tid; r < src.parts_c1; r += PEER_THREADS) {
        const float4 v = __ldcg(reinterpret_cast<const float4*>(p0 + size_t(r) * P_ARENA));
        a.x += v.x; a.y += v.y; a.z += v.z; a.w += v.w;
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        a.x += __shfl_xor_sync(0xffffffffu, a.x, o); a.y += __shfl_xor_sync(0xffffffffu, a.y, o);
        a.z += __shfl_xor_sync(0xffffffffu, a.z, o); a.w += __shfl_xor_sync(0xffffffffu, a.w, o);
      }
      if ((tid & 31) == 0) wsum_s[tid >> 5] = a;
      __syncthreads();
      if (tid == 0) {
        float4 t = wsum_s[0];
        for (int w = 1; w < PEER_THREADS / 32; ++w) { t.x += wsum_s[w].x; t.y += wsum_s[w].y; t.z += wsum_s[w].z; t.w += wsum_s[w].w; }
        t = peer_zero_pads(t, c * 4);
        for (int p = 0; p < W; ++p) reinterpret_cast<float4*>(pd.slots[p] + slot_off)[c] = t;
      }
      __syncthreads();
    }
  }
  __threadfence_system();
  __syncthreads();
  // 2. the last CTA to get here publishes the flags
  if (tid == 0) {
    const unsigned int t = atomicAdd(pd.ticket, 1u);
    if (t == gridDim.x - 1) {
      __threadfence_system();
      for (int p = 0; p < W; ++p) st_release_sys(pd.flags[p] + (par * W + R) * PEER_FLAG_STRIDE, e);
      *pd.ticket = 0u;
      *pd.epoch = e;
      if (UPDATE && kc.bump_step && step_dev) *const_cast<int64_t*>(step_dev) += 1;   // all CTAs read it at their start
    }
  }
  // 3. wait for every source's flag for this epoch
  if (tid < W) {
    const unsigned int* f = pd.flags[R] + (par * W + tid) * PEER_FLAG_STRIDE;
    unsigned int spins = 0;
    while (ld_acquire_sys(f) != e) {
      if (++spins > (1u << 28)) __trap();   // a lost peer must abort the launch, not hang the GPU
    }
  }
  __syncthreads();
  // 4. rank-ordered sum + (optional) update
  const float* base = pd.slots[R] + int64_t(par) * W * pd.max_elems;
  for (int64_t i = gtid; i < nvec; i += gthreads) {
    float4 acc = __ldcg(reinterpret_cast<const float4*>(base) + i);
#pragma unroll
    for (int s = 1; s < PEER_MAX_WORLD; ++s)
      if (s < W) {
        const float4 v = __ldcg(reinterpret_cast<const float4*>(base + int64_t(s) * pd.max_elems) + i);
        acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
      }
    if (!UPDATE) {
      reinterpret_cast<float4*>(out)[i] = acc;
    } else {
      const UpdateConsts k = ks;
      float4 pv = ld4(params + i * 4);
      float4 s1v = HAS1 ? ld4(state1 + i * 4) : make_float4(0.f, 0.f, 0.f, 0.f);
      float4 s2v = HAS2 ? ld4(state2 + i * 4) : make_float4(0.f, 0.f, 0.f, 0.f);
      float g[4] = {acc.x, acc.y, acc.z, acc.w}, p[4] = {pv.x, pv.y, pv.z, pv.w};
      float s1[4] = {s1v.x, s1v.y, s1v.z, s1v.w}, s2[4] = {s2v.x, s2v.y, s2v.z, s2v.w}, go[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) update_one<OPT>(g[j], p[j], s1[j], s2[j], go[j], k);
      st4(params + i * 4, make_float4(p[0], p[1], p[2], p[3]));
      if (HAS1) st4(state1 + i * 4, make_float4(s1[0], s1[1], s1[2], s1[3]));
      if (HAS2) st4(state2 + i * 4, make_float4(s2[0], s2[1], s2[2], s2[3]));
      if (out) reinterpret_cast<float4*>(out)[i] = make_float4(go[0], go[1], go[2], go[3]);   // averaged gradient
    }
  }
}

// ---------------------------------------------------------------------------------------------------------------------
// Flag-in-data variant of the one-shot exchange (default for the bucket-fed call; DDP_PEER_LL=0 keeps the kernel above).
// Every 8-byte word a rank pushes carries (value, epoch): the receiver polls the DATA words themselves, so there is no
// separate flag, no system-scope fence and no "last CTA publishes" ticket between the stores and the wait -- the exposed
// cost is ONE one-way NVLink store latency instead of store -> fence.sys -> ticket -> fence.sys -> flag -> poll
// (three serialized round trips in the kernel above).  64-bit scalar stores are single-copy atomic, the epoch alternates
// two slot sets exactly as above, a rank's own contribution never goes through memory, and the sum runs in rank order,
// so the result is bit-identical on every rank.  Twice the bytes on the wire (175 KB per peer for the CNN): irrelevant.
// ---------------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ void st_ll2(unsigned long long* p, unsigned long long a, unsigned long long b) {
  asm volatile("st.volatile.global.v2.u64 [%0], {%1, %2};" :: "l"(p), "l"(a), "l"(b) : "memory");
}
__device__ __forceinline__ void ld_ll2(const unsigned long long* p, unsigned long long& a, unsigned long long& b) {
  asm volatile("ld.volatile.global.v2.u64 {%0, %1}, [%2];" : "=l"(a), "=l"(b) : "l"(p) : "memory");
}
template <int OPT, bool HAS1, bool HAS2, bool UPDATE>
__global__ void __launch_bounds__(PEER_THREADS, 1)
peer_ll_kernel(PeerDev pd, const float* __restrict__ bucket, float* __restrict__ out, int64_t n,
               float* __restrict__ params, float* __restrict__ state1, float* __restrict__ state2,
               UpdateConsts kc, int64_t step, const int64_t* __restrict__ step_dev) {
  __shared__ UpdateConsts ks;
  __shared__ unsigned int epoch_s;
  const int tid = threadIdx.x;
  if (tid == 0) {
    epoch_s = *((volatile unsigned int*)pd.epoch) + 1u;   // bumped by the last CTA to finish, after every CTA has read it
    if (UPDATE) {
      UpdateConsts k = kc;
      finish_consts(k, OPT, step_dev ? (*step_dev + 1) : step);
      ks = k;
    }
  }
  __syncthreads();
  pdl_wait();                                             // the bucket is the previous kernel's output
  const unsigned int e = epoch_s;
  const unsigned long long etag = (unsigned long long)e << 32;
  const int par = int(e & 1u);
  const int W = pd.world, R = pd.rank;
  const int64_t nvec = (n + 3) >> 2;
  const int64_t gtid = int64_t(blockIdx.x) * PEER_THREADS + tid, gthreads = int64_t(gridDim.x) * PEER_THREADS;
  const int64_t my_off = (int64_t(par) * W + R) * pd.max_elems;
  // 1. push (value, epoch) words into slot[par][R] of every OTHER rank
  for (int64_t i = gtid; i < nvec; i += gthreads) {
    const float4 v = reinterpret_cast<const float4*>(bucket)[i];
    const unsigned long long q0 = etag | __float_as_uint(v.x), q1 = etag | __float_as_uint(v.y);
    const unsigned long long q2 = etag | __float_as_uint(v.z), q3 = etag | __float_as_uint(v.w);
#pragma unroll
    for (int p = 0; p < PEER_MAX_WORLD; ++p)
      if (p < W && p != R) {
        unsigned long long* dst = pd.ll[p] + my_off + 4 * i;
        st_ll2(dst, q0, q1);
        st_ll2(dst + 2, q2, q3);
      }
  }
  // 2. rank-ordered sum of the W contributions as they arrive + (optional) update
  const unsigned long long* base = pd.ll[R] + int64_t(par) * W * pd.max_elems;
  for (int64_t i = gtid; i < nvec; i += gthreads) {
    const float4 mine = reinterpret_cast<const float4*>(bucket)[i];
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
    for (int s = 0; s < PEER_MAX_WORLD; ++s)
      if (s < W) {
        float4 v = mine;
        if (s != R) {
          const unsigned long long* src = base + int64_t(s) * pd.max_elems + 4 * i;
          unsigned long long a, b, c, d;
          unsigned int spins = 0;
          for (;;) {
            ld_ll2(src, a, b);
            ld_ll2(src + 2, c, d);
            if (unsigned(a >> 32) == e && unsigned(b >> 32) == e && unsigned(c >> 32) == e && unsigned(d >> 32) == e) break;
            if (++spins > (1u << 26)) __trap();           // a lost peer must abort the launch, not hang the GPU
          }
          v = make_float4(__uint_as_float(unsigned(a)), __uint_as_float(unsigned(b)), __uint_as_float(unsigned(c)),
                          __uint_as_float(unsigned(d)));
        }
        if (s == 0) acc = v; else { acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w; }
      }
    if (!UPDATE) {
      reinterpret_cast<float4*>(out)[i] = acc;
    } else {
      const UpdateConsts k = ks;
      float4 pv = ld4(params + i * 4);
      float4 s1v = HAS1 ? ld4(state1 + i * 4) : make_float4(0.f, 0.f, 0.f, 0.f);
      float4 s2v = HAS2 ? ld4(state2 + i * 4) : make_float4(0.f, 0.f, 0.f, 0.f);
      float g[4] = {acc.x, acc.y, acc.z, acc.w}, p[4] = {pv.x, pv.y, pv.z, pv.w};
      float s1[4] = {s1v.x, s1v.y, s1v.z, s1v.w}, s2[4] = {s2v.x, s2v.y, s2v.z, s2v.w}, go[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) update_one<OPT>(g[j], p[j], s1[j], s2[j], go[j], k);
      st4(params + i * 4, make_float4(p[0], p[1], p[2], p[3]));
      if (HAS1) st4(state1 + i * 4, make_float4(s1[0], s1[1], s1[2], s1[3]));
      if (HAS2) st4(state2 + i * 4, make_float4(s2[0], s2[1], s2[2], s2[3]));
      if (out) reinterpret_cast<float4*>(out)[i] = make_float4(go[0], go[1], go[2], go[3]);   // averaged gradient
    }
  }
  // 3. off the critical path: the last CTA to finish bumps the epoch (and the step counter); every CTA read both at its start
  __syncthreads();
  if (tid == 0) {
    __threadfence();
    if (atomicAdd(pd.ticket, 1u) == gridDim.x - 1) {
      *pd.ticket = 0u;
      *pd.epoch = e;
      if (UPDATE && kc.bump_step && step_dev) *const_cast<int64_t*>(step_dev) += 1;
    }
  }
}

// The early half of the exchange (PartialSrc mode 2): reduce the fc / conv2 partial rows (columns PEER_C1_COLS .. end) in
// fixed order and push the sums into slot[par][R] of every rank.  Runs on a side stream beside conv1_wgrad; publishes
// nothing -- the flag is raised by the final kernel, which is ordered after this one on every rank.
constexpr int PEER_EARLY_CTAS = 32;     // small CTAs (256 threads, <= 64 registers): two of them fit beside ONE conv1_wgrad CTA
constexpr int PEER_EARLY_WARPS = 8;
__global__ void __launch_bounds__(PEER_EARLY_WARPS * 32, 4)
peer_early_kernel(PeerDev pd, const float* __restrict__ partials, int parts_c2, int parts_head) {
  __shared__ float4 red[PEER_EARLY_WARPS][33];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const unsigned int e = *((volatile unsigned int*)pd.epoch) + 1u;      // the final kernel of this step bumps it
  const int par = int(e & 1u), W = pd.world, R = pd.rank;
  const int64_t slot_off = (int64_t(par) * W + R) * pd.max_elems;
  const int ncols = P_ARENA / 4 - PEER_C1_COLS;
  for (int blk = blockIdx.x; blk * 32 < ncols; blk += gridDim.x) {
    const int col = PEER_C1_COLS + blk * 32 + lane;
    const int el = col * 4;
    float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
    if (el < P_ARENA) {
      const int parts = el < OFF_F1W ? parts_c2 : parts_head;
      const float* p0 = partials + el;
      for (int r0 = warp; r0 < parts; r0 += 4 * PEER_EARLY_WARPS) {      // rows warp, warp+32, ...: four loads in flight
        float4 v[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const int r = r0 + q * PEER_EARLY_WARPS;
          v[q] = r < parts ? __ldcg(reinterpret_cast<const float4*>(p0 + size_t(r) * P_ARENA)) : make_float4(0.f, 0.f, 0.f, 0.f);
        }
#pragma unroll
        for (int q = 0; q < 4; ++q) { s.x += v[q].x; s.y += v[q].y; s.z += v[q].z; s.w += v[q].w; }
      }
    }
    red[warp][lane] = s;
    __syncthreads();
    if (warp == 0 && el < P_ARENA) {
      float4 t = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
      for (int w = 0; w < PEER_EARLY_WARPS; ++w) { const float4 r = red[w][lane]; t.x += r.x; t.y += r.y; t.z += r.z; t.w += r.w; }
      t = peer_zero_pads(t, el);
#pragma unroll
      for (int p = 0; p < PEER_MAX_WORLD; ++p)
        if (p < W) reinterpret_cast<float4*>(pd.slots[p] + slot_off)[col] = t;
    }
    __syncthreads();
  }
  __threadfence_system();
}

static size_t peer_legacy_bytes(int world, int64_t max_elems) {
  return (size_t(2) * world * max_elems * 4 + size_t(2) * world * PEER_FLAG_STRIDE * 4 + 256 + 255) & ~size_t(255);
}
static size_t peer_bytes(int world, int64_t max_elems) {
  return peer_legacy_bytes(world, max_elems) + size_t(2) * world * max_elems * 8;     // + the flag-in-data slots
}
static bool peer_ll_enabled() {            // A/B switch: DDP_PEER_LL=0 = separate flags + fences (the first protocol)
  static int v = -1;
  if (v < 0) { const char* e = getenv("DDP_PEER_LL"); v = (e && e[0] == '0') ? 0 : 1; }
  return v == 1;
}
static PeerDev make_dev(const PeerCtx* c) {
  PeerDev d{};
  d.world = c->world; d.rank = c->rank; d.max_elems = c->max_elems;
  const size_t slots_bytes = size_t(2) * c->world * c->max_elems * 4;
  for (int p = 0; p < c->world; ++p) {
    d.slots[p] = reinterpret_cast<float*>(c->peer_base[p]);
    d.flags[p] = reinterpret_cast<unsigned int*>(c->peer_base[p] + slots_bytes);
  }
  char* tail = c->local_base + slots_bytes + size_t(2) * c->world * PEER_FLAG_STRIDE * 4;
  d.epoch = reinterpret_cast<unsigned int*>(tail);
  d.ticket = reinterpret_cast<unsigned int*>(tail + 128);
  for (int p = 0; p < c->world; ++p)
    d.ll[p] = reinterpret_cast<unsigned long long*>(c->peer_base[p] + peer_legacy_bytes(c->world, c->max_elems));
  return d;
}

}  // namespace ddp

using namespace ddp;

extern "C" int ddp_peer_create(void** peer_out, int world, int rank, int64_t max_elems) {
  DDP_REQUIRE(peer_out, "peer_out is NULL");
  DDP_REQUIRE(world >= 1 && world <= PEER_MAX_WORLD && rank >= 0 && rank < world, "world must be 1..8, rank in range");
  DDP_REQUIRE(max_elems >= 4 && max_elems <= (int64_t(1) << 22) && (max_elems & 3) == 0, "max_elems must be a multiple of 4, <= 4M");
  PeerCtx* c = new PeerCtx();
  memset(c, 0, sizeof(*c));
  c->world = world; c->rank = rank; c->max_elems = max_elems;
  c->bytes = peer_bytes(world, max_elems);
  cudaError_t e = cudaGetDevice(&c->device);
  if (e == cudaSuccess) e = cudaMalloc(reinterpret_cast<void**>(&c->local_base), c->bytes);
  if (e == cudaSuccess) e = cudaMemset(c->local_base, 0, c->bytes);
  if (e != cudaSuccess) { delete c; return cuda_rc(e, "ddp_peer_create: cudaMalloc"); }
  c->peer_base[rank] = c->local_base;
  *peer_out = c;
  return 0;
}

extern "C" int ddp_peer_local_handle(void* peer, void* handle_out) {
  DDP_REQUIRE(peer && handle_out, "null argument");
  static_assert(sizeof(cudaIpcMemHandle_t) == DDP_IPC_HANDLE_BYTES, "cudaIpcMemHandle_t size");
  PeerCtx* c = static_cast<PeerCtx*>(peer);
  cudaIpcMemHandle_t h;
  DDP_CUDA(cudaIpcGetMemHandle(&h, c->local_base));
  memcpy(handle_out, &h, sizeof(h));
  return 0;
}

extern "C" int ddp_peer_connect(void* peer, const void* all_handles) {
  DDP_REQUIRE(peer && all_handles, "null argument");
  PeerCtx* c = static_cast<PeerCtx*>(peer);
  for (int p = 0; p < c->world; ++p) {
    if (p == c->rank) continue;
    cudaIpcMemHandle_t h;
    memcpy(&h, static_cast<const char*>(all_handles) + size_t(p) * DDP_IPC_HANDLE_BYTES, sizeof(h));
    void* ptr = nullptr;
    DDP_CUDA(cudaIpcOpenMemHandle(&ptr, h, cudaIpcMemLazyEnablePeerAccess));
    c->peer_base[p] = static_cast<char*>(ptr);
    c->opened[p] = true;
  }
  return 0;
}

extern "C" int ddp_peer_destroy(void* peer) {
  if (!peer) return 0;
  PeerCtx* c = static_cast<PeerCtx*>(peer);
  for (int p = 0; p < c->world; ++p)
    if (c->opened[p]) cudaIpcCloseMemHandle(c->peer_base[p]);
  cudaFree(c->local_base);
  delete c;
  return 0;
}

static int peer_check(const PeerCtx* c, const void* bucket, int64_t n) {
  DDP_REQUIRE(c && bucket, "null peer context or bucket");
  DDP_REQUIRE(n >= 1 && n <= c->max_elems && (n & 3) == 0, "n must be a multiple of 4 and <= max_elems");
  DDP_REQUIRE(aligned16(bucket), "bucket must be 16-byte aligned");
  for (int p = 0; p < c->world; ++p) DDP_REQUIRE(c->peer_base[p] != nullptr, "ddp_peer_connect has not been called");
  return 0;
}

extern "C" int ddp_peer_allreduce(void* peer, float* bucket, int64_t n, void* stream_) {
  PeerCtx* c = static_cast<PeerCtx*>(peer);
  if (int rc = peer_check(c, bucket, n)) return rc;
  UpdateConsts k{};
  if (peer_ll_enabled())
    peer_ll_kernel<0, false, false, false><<<PEER_CTAS, PEER_THREADS, 0, static_cast<cudaStream_t>(stream_)>>>(
        make_dev(c), bucket, bucket, n, nullptr, nullptr, nullptr, k, 1, nullptr);
  else
  peer_allreduce_kernel<0, false, false, false><<<PEER_CTAS, PEER_THREADS, 0, static_cast<cudaStream_t>(stream_)>>>(
      make_dev(c), bucket, bucket, n, nullptr, nullptr, nullptr, k, 1, nullptr, PartialSrc{0, nullptr, 0, 0, 0});
  DDP_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int ddp_peer_allreduce_update(void* peer, float* bucket, int64_t n, int opt, float* params, float* state1,
                                         float* state2, const ddp_hyper_t* h, int64_t step, const int64_t* step_dev,
                                         void* stream_) {
  PeerCtx* c = static_cast<PeerCtx*>(peer);
  if (int rc = peer_check(c, bucket, n)) return rc;
  DDP_REQUIRE(params && h && aligned16(params), "params (16-byte aligned) and hyper are required");
  DDP_REQUIRE(step >= 1 || step_dev, "step must be >= 1 unless step_dev is given");
  const bool need1 = (opt == DDP_OPT_SGD) ? (h->momentum != 0.0) : true;
  const bool need2 = (opt != DDP_OPT_SGD);
  DDP_REQUIRE(!need1 || (state1 && aligned16(state1)), "state1 required for this optimizer");
  DDP_REQUIRE(!need2 || (state2 && aligned16(state2)), "state2 required for this optimizer");
  const UpdateConsts k = make_update_consts(h, 1.0f / float(c->world));
  float* wb = h->write_back_grad ? bucket : nullptr;
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  const PeerDev d = make_dev(c);
  const ddp::PartialSrc nosrc{0, nullptr, 0, 0, 0};
  const float* bucket_c = bucket;
  const bool ll = peer_ll_enabled();
  // flag-in-data kernel: one float4 per thread when the bucket allows it (the CNN's 5,462 vectors: 11 CTAs), at most 32 CTAs
  int ll_ctas = int(((n + 3) / 4 + PEER_THREADS - 1) / PEER_THREADS);
  if (ll_ctas < 1) ll_ctas = 1;
  if (ll_ctas > 32) ll_ctas = 32;
#define LAUNCH(OPT, H1, H2) do { if (ll) DDP_CUDA(ddp::launch_pdl(peer_ll_kernel<OPT, H1, H2, true>, dim3(ll_ctas), dim3(PEER_THREADS), 0, \
      stream, true, d, bucket_c, wb, n, params, state1, state2, k, step, step_dev)); \
    else DDP_CUDA(ddp::launch_pdl(peer_allreduce_kernel<OPT, H1, H2, true>, dim3(PEER_CTAS), dim3(PEER_THREADS), 0, \
      stream, true, d, bucket_c, wb, n, params, state1, state2, k, step, step_dev, nosrc)); } while (0)
  switch (opt) {
    case DDP_OPT_SGD: if (need1) LAUNCH(DDP_OPT_SGD, true, false); else LAUNCH(DDP_OPT_SGD, false, false); break;
    case DDP_OPT_ADAM: LAUNCH(DDP_OPT_ADAM, true, true); break;
    case DDP_OPT_ADAMW: LAUNCH(DDP_OPT_ADAMW, true, true); break;
    case DDP_OPT_ADADELTA: LAUNCH(DDP_OPT_ADADELTA, true, true); break;
    default: set_error("ddp_peer_allreduce_update: unknown optimizer kind %d", opt); return DDP_EINVAL;
  }
#undef LAUNCH
  DDP_CUDA(cudaGetLastError());
  return 0;
}

// ---- the exchange fed directly by the backward's partial gradient rows (common.cuh: PeerExchange) ------------------
namespace ddp {
int launch_peer_early(const PeerExchange& px, const float* partials, int parts_c2, int parts_head, cudaStream_t stream) {
  PeerCtx* c = static_cast<PeerCtx*>(px.peer);
  DDP_REQUIRE(c && partials, "null peer context or partials");
  DDP_REQUIRE(c->max_elems >= P_ARENA, "peer context too small for the Net arena");
  for (int p = 0; p < c->world; ++p) DDP_REQUIRE(c->peer_base[p] != nullptr, "ddp_peer_connect has not been called");
  peer_early_kernel<<<PEER_EARLY_CTAS, PEER_EARLY_WARPS * 32, 0, stream>>>(make_dev(c), partials, parts_c2, parts_head);
  DDP_CUDA(cudaGetLastError());
  return 0;
}

int launch_peer_final(const PeerExchange& px, const float* partials, int parts_c1, bool have_early, int parts_c2, int parts_head,
                      float* grad_bucket, cudaStream_t stream) {
  PeerCtx* c = static_cast<PeerCtx*>(px.peer);
  DDP_REQUIRE(c && partials && px.params && px.hyper, "null peer context, partials, params or hyper");
  DDP_REQUIRE(c->max_elems >= P_ARENA, "peer context too small for the Net arena");
  for (int p = 0; p < c->world; ++p) DDP_REQUIRE(c->peer_base[p] != nullptr, "ddp_peer_connect has not been called");
  const ddp_hyper_t* h = px.hyper;
  const int opt = px.opt;
  const bool need1 = (opt == DDP_OPT_SGD) ? (h->momentum != 0.0) : true;
  const bool need2 = (opt != DDP_OPT_SGD);
  DDP_REQUIRE(!need1 || (px.state1 && aligned16(px.state1)), "state1 required for this optimizer");
  DDP_REQUIRE(!need2 || (px.state2 && aligned16(px.state2)), "state2 required for this optimizer");
  const UpdateConsts k = make_update_consts(h, 1.0f / float(c->world));
  float* wb = h->write_back_grad ? grad_bucket : nullptr;
  const PeerDev d = make_dev(c);
  const PartialSrc src{have_early ? 2 : 1, partials, parts_c1, parts_c2, parts_head};
  const int64_t n = P_ARENA;
  // fed by partial rows the kernel also does the reduction: 64 co-resident CTAs (they wait on each other only through
  // the device-local ticket; 64 <= number of SMs) instead of 8
  const int ctas = have_early ? PEER_CTAS : 64;
  const float* no_bucket = nullptr;
#define LAUNCH(OPT, H1, H2) DDP_CUDA(launch_pdl(peer_allreduce_kernel<OPT, H1, H2, true>, dim3(ctas), dim3(PEER_THREADS), 0, \
      stream, true, d, no_bucket, wb, n, px.params, px.state1, px.state2, k, px.step, px.step_dev, src))
  switch (opt) {
    case DDP_OPT_SGD: if (need1) LAUNCH(DDP_OPT_SGD, true, false); else LAUNCH(DDP_OPT_SGD, false, false); break;
    case DDP_OPT_ADAM: LAUNCH(DDP_OPT_ADAM, true, true); break;
    case DDP_OPT_ADAMW: LAUNCH(DDP_OPT_ADAMW, true, true); break;
    case DDP_OPT_ADADELTA: LAUNCH(DDP_OPT_ADADELTA, true, true); break;
    default: set_error("launch_peer_final: unknown optimizer kind %d", opt); return DDP_EINVAL;
  }
#undef LAUNCH
  DDP_CUDA(cudaGetLastError());
  return 0;
}
}  // namespace ddp
