// Development probe (not part of the product library): which half of the NVSwitch two-shot bounds it?
// Variants of the chunk loop of csrc/sym_allreduce.cu without any cross-rank synchronisation (the caller
// barriers on the host): mode 0 = ld_reduce + st, 1 = ld_reduce only, 2 = multimem.st only,
// 3 = ld_reduce + st with a contiguous slab per CTA, 4 = plain peer loads only, 5 = plain peer stores only.
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ uint4 mc_ld(const char* p) {
  uint4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void mc_st(char* p, uint4 v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" :: "l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ uint4 ld_peer(const char* p) {
  uint4 v;
  asm volatile("ld.relaxed.sys.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_peer(char* p, uint4 v) {
  asm volatile("st.relaxed.sys.global.v4.u32 [%0], {%1,%2,%3,%4};" :: "l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

struct Peers { char* buf[8]; };

template <int UNROLL>
__global__ void __launch_bounds__(512, 1)
probe_kernel(char* mc, Peers peers, int world, int rank, int64_t lo, int64_t hi, int mode, float* sink) {
  const int tid = threadIdx.x;
  const int64_t gthreads = int64_t(gridDim.x) * blockDim.x;
  uint4 acc = make_uint4(0, 0, 0, 0);
  if (mode == 3) {            // contiguous slab per CTA
    const int64_t per = (hi - lo + gridDim.x - 1) / gridDim.x;
    const int64_t a = lo + per * blockIdx.x, b = a + per < hi ? a + per : hi;
    for (int64_t i0 = a + tid; i0 < b; i0 += int64_t(blockDim.x) * UNROLL) {
      uint4 v[UNROLL];
#pragma unroll
      for (int u = 0; u < UNROLL; ++u) { int64_t i = i0 + u * blockDim.x; if (i < b) v[u] = mc_ld(mc + i * 16); }
#pragma unroll
      for (int u = 0; u < UNROLL; ++u) { int64_t i = i0 + u * blockDim.x; if (i < b) mc_st(mc + i * 16, v[u]); }
    }
    return;
  }
  for (int64_t i0 = lo + int64_t(blockIdx.x) * blockDim.x + tid; i0 < hi; i0 += gthreads * UNROLL) {
    uint4 v[UNROLL];
#pragma unroll
    for (int u = 0; u < UNROLL; ++u) {
      const int64_t i = i0 + u * gthreads;
      if (i < hi) {
        if (mode == 0 || mode == 1) v[u] = mc_ld(mc + i * 16);
        else if (mode == 4) { v[u] = ld_peer(peers.buf[(rank + 1 + (u % (world - 1))) % world] + i * 16); }
        else v[u] = make_uint4(i, 1, 2, 3);
      }
    }
#pragma unroll
    for (int u = 0; u < UNROLL; ++u) {
      const int64_t i = i0 + u * gthreads;
      if (i < hi) {
        if (mode == 0 || mode == 2) mc_st(mc + i * 16, v[u]);
        else if (mode == 5) st_peer(peers.buf[(rank + 1 + (u % (world - 1))) % world] + i * 16, v[u]);
        else { acc.x ^= v[u].x; acc.y ^= v[u].y; acc.z ^= v[u].z; acc.w ^= v[u].w; }
      }
    }
  }
  if (acc.x == 0x12345678u && acc.y == 0x9abcdef0u) sink[0] = 1.f;   // keep the loads alive
}

extern "C" int mc_probe_launch(uint64_t mc, const uint64_t* replicas, int world, int rank, int64_t nvec, int mode,
                               int ctas, int unroll, int whole, float* sink, void* stream) {
  Peers p{};
  for (int q = 0; q < world; ++q) p.buf[q] = reinterpret_cast<char*>(replicas[q]);
  const int64_t per = (nvec + world - 1) / world;
  int64_t lo = per * rank, hi = lo + per < nvec ? lo + per : nvec;
  if (whole) { lo = 0; hi = nvec; }
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (unroll == 2) probe_kernel<2><<<ctas, 512, 0, s>>>(reinterpret_cast<char*>(mc), p, world, rank, lo, hi, mode, sink);
  else if (unroll == 8) probe_kernel<8><<<ctas, 512, 0, s>>>(reinterpret_cast<char*>(mc), p, world, rank, lo, hi, mode, sink);
  else probe_kernel<4><<<ctas, 512, 0, s>>>(reinterpret_cast<char*>(mc), p, world, rank, lo, hi, mode, sink);
  return int(cudaGetLastError());
}
