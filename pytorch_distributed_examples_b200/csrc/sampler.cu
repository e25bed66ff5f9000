// DistributedSampler shard-index generation, bit-exact with torch.
//   torch: utils/data/distributed.py:102-134  (g = Generator().manual_seed(seed+epoch);
//          indices = randperm(len, generator=g); pad/truncate; indices[rank::W])
//   torch randperm on CPU for n < 2^32/20: r = arange(n); for i in 0..n-2:
//          z = mt19937() % (n-i); swap(r[i], r[i+z])       (ATen TensorFactories.cpp)
//   ref call sites: horovod/mnist_horovod.py:41-42, horovod_mnist_elastic.py:50-51,
//          pytorch_elastic/mnist_ddp_elastic.py:188 (+ set_epoch :89)
// The MT19937 stream is generated 624 words at a time by 7 warps (three parallel twist
// phases) while one lane of warp 0 replays the inherently sequential swap chain of the
// previous block out of shared memory.
#include "common.cuh"

namespace ddp {

constexpr int MT_N = 624, MT_M = 397;
constexpr int RP_THREADS = 256;
constexpr int RP_GEN = RP_THREADS - 32;  // generator threads (warps 1..7)
constexpr int64_t RP_SMEM_MAX_N = 65536; // permutation kept as u16 in shared memory

__device__ __forceinline__ uint32_t mt_twist(uint32_t u, uint32_t v) {
  return (((u & 0x80000000u) | (v & 0x7fffffffu)) >> 1) ^ ((v & 1u) ? 0x9908b0dfu : 0u);
}
__device__ __forceinline__ uint32_t mt_temper(uint32_t y) {
  y ^= y >> 11;
  y ^= (y << 7) & 0x9d2c5680u;
  y ^= (y << 15) & 0xefc60000u;
  y ^= y >> 18;
  return y;
}
__device__ __forceinline__ void gen_barrier() { asm volatile("bar.sync 1, %0;" :: "n"(RP_GEN) : "memory"); }

// Generator threads (gt in [0,RP_GEN)): advance the 624-word state one generation in
// place and emit z[i] = temper(state[i]) % (n - (base+i)) for the draws of this block.
__device__ void mt_next_block(uint32_t* st, uint32_t* zout, int gt, int64_t base, int64_t n) {
  // phase 1: i in [0,227): new[i] = old[i+397] ^ twist(old[i], old[i+1])
  // phase 2: i in [227,454): new[i] = new[i-227] ^ twist(old[i], old[i+1])
  // phase 3: i in [454,623): new[i] = new[i-227] ^ twist(old[i], old[i+1]);  i=623 uses new[0]
  const int lo[3] = {0, MT_N - MT_M, 2 * (MT_N - MT_M)};
  const int hi[3] = {MT_N - MT_M, 2 * (MT_N - MT_M), MT_N};
#pragma unroll
  for (int ph = 0; ph < 3; ++ph) {
    uint32_t val[2];
    int cnt = 0;
    for (int i = lo[ph] + gt; i < hi[ph]; i += RP_GEN, ++cnt) {
      uint32_t a = st[i];
      uint32_t b = (i == MT_N - 1) ? st[0] : st[i + 1];  // st[0] is already new in phase 3
      uint32_t c = (i < MT_N - MT_M) ? st[i + MT_M] : st[i - (MT_N - MT_M)];
      val[cnt] = c ^ mt_twist(a, b);
    }
    gen_barrier();
    cnt = 0;
    for (int i = lo[ph] + gt; i < hi[ph]; i += RP_GEN, ++cnt) st[i] = val[cnt];
    gen_barrier();
  }
  for (int i = gt; i < MT_N; i += RP_GEN) {
    int64_t k = base + i;                 // draw index = Fisher-Yates position
    uint32_t y = mt_temper(st[i]);
    zout[i] = (k < n - 1) ? uint32_t(y % uint32_t(n - k)) : 0u;
  }
}

template <bool SMEM_PERM>
__global__ void __launch_bounds__(RP_THREADS, 1)
randperm_kernel(int64_t n, uint32_t seed, int32_t* __restrict__ perm) {
  extern __shared__ __align__(16) unsigned char rp_smem[];
  uint32_t* st = reinterpret_cast<uint32_t*>(rp_smem);   // [624]
  uint32_t* zb = st + MT_N;                               // [2][624]
  uint16_t* r16 = reinterpret_cast<uint16_t*>(zb + 2 * MT_N);
  const int tid = threadIdx.x;
  if (tid == 0) {
    uint32_t s = seed;
    st[0] = s;
    for (int j = 1; j < MT_N; ++j) { s = 1812433253u * (s ^ (s >> 30)) + uint32_t(j); st[j] = s; }
  }
  if (SMEM_PERM) { for (int64_t i = tid; i < n; i += RP_THREADS) r16[i] = uint16_t(i); }
  else           { for (int64_t i = tid; i < n; i += RP_THREADS) perm[i] = int32_t(i); }
  __threadfence_block();
  __syncthreads();
  const int64_t ndraw = n - 1;
  const int64_t nblocks = (ndraw + MT_N - 1) / MT_N;
  const int gt = tid - 32;
  if (gt >= 0 && nblocks > 0) mt_next_block(st, zb, gt, 0, n);
  __syncthreads();
  for (int64_t blk = 0; blk < nblocks; ++blk) {
    if (gt >= 0) {
      if (blk + 1 < nblocks) mt_next_block(st, zb + ((blk + 1) & 1) * MT_N, gt, (blk + 1) * MT_N, n);
    } else if (tid == 0) {
      const uint32_t* z = zb + (blk & 1) * MT_N;
      const int64_t base = blk * MT_N;
      const int cnt = int(ndraw - base < MT_N ? ndraw - base : MT_N);
      if (SMEM_PERM) {
        for (int i = 0; i < cnt; ++i) {
          const uint32_t a = uint32_t(base) + i, b = a + z[i];
          uint16_t t = r16[a]; r16[a] = r16[b]; r16[b] = t;
        }
      } else {
        for (int i = 0; i < cnt; ++i) {
          const int64_t a = base + i, b = a + z[i];
          int32_t t = perm[a]; perm[a] = perm[b]; perm[b] = t;
        }
      }
    }
    __syncthreads();
  }
  if (SMEM_PERM) for (int64_t i = tid; i < n; i += RP_THREADS) perm[i] = int32_t(r16[i]);
}

__global__ void __launch_bounds__(256)
shard_kernel(const int32_t* __restrict__ perm, int64_t n, int world, int rank, int64_t num_samples,
             int shuffle, int64_t* __restrict__ out) {
  const int64_t j = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  if (j >= num_samples) return;
  const int64_t k = (int64_t(rank) + j * world) % n;   // wrap-around padding == list repetition
  out[j] = shuffle ? int64_t(perm[k]) : k;
}

}  // namespace ddp

using namespace ddp;

extern "C" int64_t ddp_sampler_num_samples(int64_t n, int world, int drop_last) {
  if (n < 0 || world <= 0) return -1;
  if (drop_last && n % world != 0) return (n - 1) / world;  // == ceil((n - world) / world) == floor(n / world)
  return (n + world - 1) / world;
}

extern "C" int ddp_sampler_indices(int64_t n, int world, int rank, int shuffle, uint64_t seed,
                                   uint64_t epoch, int drop_last, int32_t* perm_scratch, int64_t* out,
                                   void* stream_) {
  DDP_REQUIRE(n >= 1 && n < (int64_t(1) << 31) / 20 * 2, "dataset length out of range");
  DDP_REQUIRE(world >= 1 && rank >= 0 && rank < world, "bad world/rank");
  DDP_REQUIRE(out && (perm_scratch || !shuffle), "null output / perm_scratch");
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  const int64_t num_samples = ddp_sampler_num_samples(n, world, drop_last);
  if (shuffle) {
    const uint32_t seed32 = uint32_t((seed + epoch) & 0xffffffffu);
    if (n <= RP_SMEM_MAX_N) {
      static bool attr_done = false;
      const int max_smem = int(3 * MT_N * 4 + RP_SMEM_MAX_N * 2);
      if (!attr_done) {
        DDP_CUDA(cudaFuncSetAttribute(randperm_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, max_smem));
        attr_done = true;
      }
      const int smem = int(3 * MT_N * 4 + ((n * 2 + 15) & ~int64_t(15)));
      randperm_kernel<true><<<1, RP_THREADS, smem, stream>>>(n, seed32, perm_scratch);
    } else {
      randperm_kernel<false><<<1, RP_THREADS, 3 * MT_N * 4, stream>>>(n, seed32, perm_scratch);
    }
  }
  if (num_samples > 0)
    shard_kernel<<<int((num_samples + 255) / 256), 256, 0, stream>>>(perm_scratch, n, world, rank, num_samples,
                                                                   shuffle, out);
  DDP_CUDA(cudaGetLastError());
  return 0;
}

// ---------------------------------------------------------------------------------------------------
// ddp_sampler_indices_static: the same shard-index computation for callers that cannot allocate.
//
// ddp_sampler_indices needs a caller-provided permutation buffer of n int32.  Some hosts have no way to
// give it one: a data-feeding thread that runs after the application has handed every byte of device
// memory to its own fixed pool, a plain-C program that links this library without bringing an allocator,
// or a loop that must not pay the device-wide synchronisation cudaMalloc/cudaFree imply.  For them the
// permutation lives in the array below, which is part of the library image: every GPU that loads the
// library carries these 256 KiB until the process exits, and nothing is allocated or freed here or in
// anything this function calls.
//
// The array name is only a host-side handle.  Its device address is looked up once per device with
// cudaGetSymbolAddress while that device is current (the first call on each device does this, so make
// that first call before starting any stream capture) and the kernels are handed that address.
// All calls share the one array and it keeps whatever the previous call left in it; the permutation
// kernel rewrites all n entries before reading any, so nothing needs clearing — but two calls whose GPU
// work overlaps (different streams, or different host threads) race on it: serialise them.
// ---------------------------------------------------------------------------------------------------
static __device__ __align__(256) int32_t g_static_perm[DDP_SAMPLER_STATIC_MAX_N];

extern "C" int ddp_sampler_indices_static(int64_t n, int world, int rank, int shuffle, uint64_t seed,
                                          uint64_t epoch, int drop_last, int64_t* out, void* stream_) {
  DDP_REQUIRE(n >= 1, "dataset length must be >= 1");
  if (n > DDP_SAMPLER_STATIC_MAX_N) {
    set_error("ddp_sampler_indices_static: dataset length %lld exceeds the built-in scratch of %d entries; "
              "use ddp_sampler_indices with a caller-provided perm_scratch",
              (long long)n, DDP_SAMPLER_STATIC_MAX_N);
    return DDP_EINVAL;
  }
  DDP_REQUIRE(world >= 1 && rank >= 0 && rank < world, "bad world/rank");
  DDP_REQUIRE(out, "null output");
  static_assert(DDP_SAMPLER_STATIC_MAX_N <= RP_SMEM_MAX_N, "static scratch must fit the shared-memory path");
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);

  // per-device set-up: resolve the device address of the scratch array, opt the kernel into large smem
  static int32_t* dev_ptr[64] = {nullptr};
  int dev = -1;
  DDP_CUDA(cudaGetDevice(&dev));
  DDP_REQUIRE(dev >= 0 && dev < 64, "device ordinal out of range");
  if (dev_ptr[dev] == nullptr) {
    int32_t* p = nullptr;
    DDP_CUDA(cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_static_perm));
    DDP_REQUIRE(p != nullptr, "cudaGetSymbolAddress returned a null address");
    DDP_CUDA(cudaFuncSetAttribute(randperm_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  int(3 * MT_N * 4 + RP_SMEM_MAX_N * 2)));
    dev_ptr[dev] = p;
  }
  int32_t* perm = dev_ptr[dev];

  const int64_t num_samples = ddp_sampler_num_samples(n, world, drop_last);
  if (shuffle) {
    const uint32_t seed32 = uint32_t((seed + epoch) & 0xffffffffu);
    const int smem = int(3 * MT_N * 4 + ((n * 2 + 15) & ~int64_t(15)));
    randperm_kernel<true><<<1, RP_THREADS, smem, stream>>>(n, seed32, perm);
  }
  if (num_samples > 0)
    shard_kernel<<<int((num_samples + 255) / 256), 256, 0, stream>>>(perm, n, world, rank, num_samples, shuffle, out);
  DDP_CUDA(cudaGetLastError());
  return 0;
}
