// Two-shot allreduce for LARGE gradient buckets over NVLink 5 / NVSwitch, in place in a symmetric buffer.
//
// The ResNet-50 bucket is 102 MB (fp32) / 51 MB (bf16): bandwidth, not latency (ref: the per-bucket allreduce
// DDP issues, pytorch_elastic/mnist_ddp_elastic.py:58; SURVEY §8(d) M3 puts the bar at 80 % of 900 GB/s bus
// bandwidth at 8 GPUs).  Every rank owns one W-th of the bucket ("its chunk"):
//   multicast mode  (NVSwitch SHARP): the owner issues multimem.ld_reduce on the chunk's multicast address -- the
//                   SWITCH fetches the W replicas and adds them -- scales by 1/W and multimem.st's the result,
//                   which the switch broadcasts into every replica.  Each byte crosses a GPU's links once per
//                   direction (≈1.1 S in, 1.1 S out) instead of 1.75 S for a ring or a load/store two-shot.
//   peer mode       (no multicast object): the owner loads the chunk from every replica through the peer-mapped
//                   addresses, adds in RANK ORDER, and stores the result into every replica.
// Either way each element is reduced by exactly one rank and then copied, so all replicas end bit-identical.
//
// The buffer is symmetric memory (same size and layout on every rank, each rank's copy mapped on all peers and,
// for multicast mode, bound to one multicast object).  The host layer obtains it from torch's symmetric-memory
// allocator (plumbing) and passes raw addresses; this file owns the kernel and the synchronisation.
// Flags live in the last DDP_SYM_FLAG_BYTES of the buffer:  u32 flag[2 phases][W sources] (128 B apart), epoch, ticket.
// Protocol per launch (epoch e = previous + 1, kept on the device so graph replays stay in step):
//   0. CTA 0 publishes "my replica holds this step's gradients" (st.release.sys flag[0][me] = e on every rank);
//      every CTA waits until all W sources have published e.
//   1. all CTAs: reduce + broadcast my chunk (grid-stride, 16-byte vectors, 2-4 in flight per thread).
//   2. the last CTA to finish publishes flag[1][me] = e everywhere, then waits for all W sources' flag[1] == e:
//      when the kernel ends, every chunk has landed in this replica and nobody is still reading it.
// CTAs never wait on CTAs of the same GPU, only on other GPUs, so the grid need not be co-resident.
// Spin loops trap after a cap: a lost peer aborts the launch instead of hanging the GPU.
#include "common.cuh"

namespace ddp {

constexpr int SYM_MAX_WORLD = 8;
constexpr int SYM_THREADS = 512;
constexpr int SYM_FLAG_STRIDE = 32;                       // u32 per flag (128 B)
constexpr int SYM_EPOCH_OFF = 2 * SYM_MAX_WORLD * SYM_FLAG_STRIDE * 4;   // 2048
constexpr int SYM_TICKET_OFF = SYM_EPOCH_OFF + 128;

struct SymDev {
  int world, rank;
  char* buf[SYM_MAX_WORLD];        // replica of every rank as mapped here; buf[rank] is the local one
  char* mc;                        // multicast address of the replicas (nullptr in peer mode)
  int64_t flag_off;                // byte offset of the flag block inside each replica
};

__device__ __forceinline__ void sym_st_release(unsigned int* p, unsigned int v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" :: "l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned int sym_ld_acquire(const unsigned int* p) {
  unsigned int v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ uint4 ld_peer(const char* p) {          // remote replicas are never cached in L1
  uint4 v;
  asm volatile("ld.relaxed.sys.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_peer(char* p, uint4 v) {
  asm volatile("st.relaxed.sys.global.v4.u32 [%0], {%1,%2,%3,%4};" :: "l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
template <bool BF16>
__device__ __forceinline__ uint4 mc_ld_reduce(const char* p) {
  uint4 v;
  if (BF16)
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0,%1,%2,%3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
  else
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
  return v;
}
template <bool BF16>
__device__ __forceinline__ void mc_st(char* p, uint4 v) {
  if (BF16)
    asm volatile("multimem.st.relaxed.sys.global.v4.bf16x2 [%0], {%1,%2,%3,%4};" :: "l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
  else
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" :: "l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

__device__ __forceinline__ float bf_lo(uint32_t u) { return __uint_as_float(u << 16); }
__device__ __forceinline__ float bf_hi(uint32_t u) { return __uint_as_float(u & 0xffff0000u); }
__device__ __forceinline__ uint32_t bf_pack(float lo, float hi) {   // round to nearest even, like torch's bf16 cast
  uint32_t r;
  asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
  return r;
}
template <bool BF16>
__device__ __forceinline__ uint4 scale_vec(uint4 v, float s) {
  if (BF16) {
    v.x = bf_pack(bf_lo(v.x) * s, bf_hi(v.x) * s); v.y = bf_pack(bf_lo(v.y) * s, bf_hi(v.y) * s);
    v.z = bf_pack(bf_lo(v.z) * s, bf_hi(v.z) * s); v.w = bf_pack(bf_lo(v.w) * s, bf_hi(v.w) * s);
  } else {
    v.x = __float_as_uint(__uint_as_float(v.x) * s); v.y = __float_as_uint(__uint_as_float(v.y) * s);
    v.z = __float_as_uint(__uint_as_float(v.z) * s); v.w = __float_as_uint(__uint_as_float(v.w) * s);
  }
  return v;
}
// rank-ordered sum of W replicas of one 16-byte vector, fp32 accumulation
template <bool BF16>
__device__ __forceinline__ uint4 sum_vecs(const uint4 (&v)[SYM_MAX_WORLD], int W, float s) {
  if (BF16) {
    float a[8];
    a[0] = bf_lo(v[0].x); a[1] = bf_hi(v[0].x); a[2] = bf_lo(v[0].y); a[3] = bf_hi(v[0].y);
    a[4] = bf_lo(v[0].z); a[5] = bf_hi(v[0].z); a[6] = bf_lo(v[0].w); a[7] = bf_hi(v[0].w);
#pragma unroll
    for (int q = 1; q < SYM_MAX_WORLD; ++q)
      if (q < W) {
        a[0] += bf_lo(v[q].x); a[1] += bf_hi(v[q].x); a[2] += bf_lo(v[q].y); a[3] += bf_hi(v[q].y);
        a[4] += bf_lo(v[q].z); a[5] += bf_hi(v[q].z); a[6] += bf_lo(v[q].w); a[7] += bf_hi(v[q].w);
      }
    return make_uint4(bf_pack(a[0] * s, a[1] * s), bf_pack(a[2] * s, a[3] * s), bf_pack(a[4] * s, a[5] * s), bf_pack(a[6] * s, a[7] * s));
  } else {
    float a0 = __uint_as_float(v[0].x), a1 = __uint_as_float(v[0].y), a2 = __uint_as_float(v[0].z), a3 = __uint_as_float(v[0].w);
#pragma unroll
    for (int q = 1; q < SYM_MAX_WORLD; ++q)
      if (q < W) {
        a0 += __uint_as_float(v[q].x); a1 += __uint_as_float(v[q].y); a2 += __uint_as_float(v[q].z); a3 += __uint_as_float(v[q].w);
      }
    return make_uint4(__float_as_uint(a0 * s), __float_as_uint(a1 * s), __float_as_uint(a2 * s), __float_as_uint(a3 * s));
  }
}

template <bool MC, bool BF16, int UNROLL>
__global__ void __launch_bounds__(SYM_THREADS, 1)
sym_allreduce_kernel(SymDev sd, int64_t byte_off, int64_t nvec, float scale) {
  __shared__ unsigned int epoch_s;
  __shared__ int last_s;
  const int tid = threadIdx.x, W = sd.world, R = sd.rank;
  char* local = sd.buf[R];
  unsigned int* lflags = reinterpret_cast<unsigned int*>(local + sd.flag_off);
  unsigned int* epoch_p = reinterpret_cast<unsigned int*>(local + sd.flag_off + SYM_EPOCH_OFF);
  unsigned int* ticket_p = reinterpret_cast<unsigned int*>(local + sd.flag_off + SYM_TICKET_OFF);
  if (tid == 0) epoch_s = *((volatile unsigned int*)epoch_p) + 1u;
  __syncthreads();
  const unsigned int e = epoch_s;

  // 0. start barrier
  if (blockIdx.x == 0 && tid < W) {
    __threadfence_system();
    sym_st_release(reinterpret_cast<unsigned int*>(sd.buf[tid] + sd.flag_off) + (0 * SYM_MAX_WORLD + R) * SYM_FLAG_STRIDE, e);
  }
  if (tid < W) {
    const unsigned int* f = lflags + (0 * SYM_MAX_WORLD + tid) * SYM_FLAG_STRIDE;
    unsigned int spins = 0;
    while (sym_ld_acquire(f) != e)
      if (++spins > (1u << 28)) __trap();
  }
  __syncthreads();

  // 1. reduce + broadcast my chunk
  const int64_t per = (nvec + W - 1) / W;
  const int64_t lo = per * R, hi = (lo + per < nvec) ? lo + per : nvec;
  const int64_t gthreads = int64_t(gridDim.x) * SYM_THREADS;
  for (int64_t i0 = lo + int64_t(blockIdx.x) * SYM_THREADS + tid; i0 < hi; i0 += gthreads * UNROLL) {
    if (MC) {
      uint4 v[UNROLL];
#pragma unroll
      for (int u = 0; u < UNROLL; ++u) {
        const int64_t i = i0 + u * gthreads;
        if (i < hi) v[u] = mc_ld_reduce<BF16>(sd.mc + byte_off + i * 16);
      }
#pragma unroll
      for (int u = 0; u < UNROLL; ++u) {
        const int64_t i = i0 + u * gthreads;
        if (i < hi) mc_st<BF16>(sd.mc + byte_off + i * 16, scale_vec<BF16>(v[u], scale));
      }
    } else {
#pragma unroll
      for (int u = 0; u < UNROLL; u += 2) {           // 2 vectors x W replicas in flight per thread
        uint4 a[SYM_MAX_WORLD], b[SYM_MAX_WORLD];
        const int64_t ia = i0 + u * gthreads, ib = (u + 1 < UNROLL) ? ia + gthreads : hi;
#pragma unroll
        for (int q = 0; q < SYM_MAX_WORLD; ++q)
          if (q < W) {
            if (ia < hi) a[q] = ld_peer(sd.buf[q] + byte_off + ia * 16);
            if (ib < hi) b[q] = ld_peer(sd.buf[q] + byte_off + ib * 16);
          }
        if (ia < hi) {
          const uint4 r = sum_vecs<BF16>(a, W, scale);
#pragma unroll
          for (int q = 0; q < SYM_MAX_WORLD; ++q)
            if (q < W) st_peer(sd.buf[q] + byte_off + ia * 16, r);
        }
        if (ib < hi) {
          const uint4 r = sum_vecs<BF16>(b, W, scale);
#pragma unroll
          for (int q = 0; q < SYM_MAX_WORLD; ++q)
            if (q < W) st_peer(sd.buf[q] + byte_off + ib * 16, r);
        }
      }
    }
  }
  __threadfence_system();
  __syncthreads();

  // 2. end barrier, driven by the last CTA of this rank
  if (tid == 0) {
    const unsigned int t = atomicAdd(ticket_p, 1u);
    last_s = (t == gridDim.x - 1);
  }
  __syncthreads();
  if (!last_s) return;
  if (tid < W) {
    __threadfence_system();
    sym_st_release(reinterpret_cast<unsigned int*>(sd.buf[tid] + sd.flag_off) + (1 * SYM_MAX_WORLD + R) * SYM_FLAG_STRIDE, e);
    const unsigned int* f = lflags + (1 * SYM_MAX_WORLD + tid) * SYM_FLAG_STRIDE;
    unsigned int spins = 0;
    while (sym_ld_acquire(f) != e)
      if (++spins > (1u << 28)) __trap();
  }
  __syncthreads();
  if (tid == 0) {
    *ticket_p = 0u;
    *epoch_p = e;
  }
}

}  // namespace ddp

using namespace ddp;

extern "C" int64_t ddp_sym_flag_bytes(void) { return DDP_SYM_FLAG_BYTES; }

extern "C" int ddp_sym_allreduce(const uint64_t* replicas, uint64_t multicast, int64_t flag_offset, int world, int rank,
                                 int64_t byte_offset, int64_t n, int dtype, float scale, int max_ctas, int unroll,
                                 void* stream_) {
  static_assert(SYM_TICKET_OFF + 4 <= DDP_SYM_FLAG_BYTES, "flag block");
  DDP_REQUIRE(replicas, "replicas is NULL");
  DDP_REQUIRE(world >= 2 && world <= SYM_MAX_WORLD && rank >= 0 && rank < world, "world must be 2..8, rank in range");
  DDP_REQUIRE(dtype == 0 || dtype == 1, "dtype must be 0 (f32) or 1 (bf16)");
  const int64_t esz = dtype == 0 ? 4 : 2;
  DDP_REQUIRE(n >= 1 && (n * esz) % 16 == 0, "n * element size must be a multiple of 16 bytes");
  DDP_REQUIRE(byte_offset >= 0 && (byte_offset & 15) == 0, "byte_offset must be a multiple of 16");
  DDP_REQUIRE(byte_offset + n * esz <= flag_offset && (flag_offset & 127) == 0,
              "the reduced range must end before the flag block (flag_offset, 128-byte aligned)");
  SymDev sd{};
  sd.world = world; sd.rank = rank; sd.flag_off = flag_offset;
  sd.mc = reinterpret_cast<char*>(multicast);
  for (int q = 0; q < world; ++q) {
    DDP_REQUIRE(replicas[q] != 0 && (replicas[q] & 15) == 0, "replica addresses must be non-null and 16-byte aligned");
    sd.buf[q] = reinterpret_cast<char*>(replicas[q]);
  }
  DDP_REQUIRE((multicast & 15) == 0, "multicast address must be 16-byte aligned");
  const int64_t nvec = n * esz / 16;
  const int64_t per = (nvec + world - 1) / world;
  int dev = 0, sms = 148;
  DDP_CUDA(cudaGetDevice(&dev));
  DDP_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  const bool mc = multicast != 0, bf = dtype == 1;
  // Measured on 8 x B200 (tools/sym_probe.py): the switch path is fastest with FEWER requests in flight
  // (64 CTAs x 2 vectors per thread); the load/store path wants every SM and 4 vectors per thread.
  const int un = unroll > 0 ? unroll : (mc ? 2 : 4);
  DDP_REQUIRE(un == 1 || un == 2 || un == 4 || un == 8, "unroll must be 0 (default), 1, 2, 4 or 8");
  int64_t want = (per + int64_t(SYM_THREADS) * un - 1) / (int64_t(SYM_THREADS) * un);
  int cap = max_ctas > 0 ? max_ctas : (mc ? 64 : sms);
  int grid = int(want < 1 ? 1 : (want > cap ? cap : want));
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
#define SYM_LAUNCH(MC, BF, UN) sym_allreduce_kernel<MC, BF, UN><<<grid, SYM_THREADS, 0, stream>>>(sd, byte_offset, nvec, scale)
#define SYM_PICK(UN) do { if (mc && bf) SYM_LAUNCH(true, true, UN); else if (mc) SYM_LAUNCH(true, false, UN); \
                          else if (bf) SYM_LAUNCH(false, true, UN); else SYM_LAUNCH(false, false, UN); } while (0)
  if (un == 1) SYM_PICK(1); else if (un == 2) SYM_PICK(2); else if (un == 4) SYM_PICK(4); else SYM_PICK(8);
#undef SYM_PICK
#undef SYM_LAUNCH
  DDP_CUDA(cudaGetLastError());
  return 0;
}
