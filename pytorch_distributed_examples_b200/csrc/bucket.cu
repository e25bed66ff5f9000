// Flatten-into-bucket / unflatten: one launch gathers every gradient tensor of a
// bucket into the flat allreduce buffer (what DDP's C++ Reducer does with one
// copy_ per parameter; torch: nn/parallel/distributed.py:1255-1280, and what
// Horovod's fusion buffer does; ref: horovod/mnist_horovod.py:53).
// Pure byte movement, HBM-bound: 2 x bucket bytes per launch, 128-bit accesses.
#include "common.cuh"

namespace ddp {

__device__ __forceinline__ uint4 ld_stream16(const void* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
  return r;
}
__device__ __forceinline__ void st_stream16(void* p, uint4 v) {
  asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};"
               :: "l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

// One CTA per chunk of at most DDP_BUCKET_CHUNK_BYTES bytes of one tensor.
__global__ void __launch_bounds__(256)
bucket_copy_kernel(const uint64_t* __restrict__ ptrs, const int64_t* __restrict__ offsets,
                   const int64_t* __restrict__ nbytes, const int32_t* __restrict__ chunk_item,
                   const int64_t* __restrict__ chunk_start, char* __restrict__ bucket, int direction) {
  const int item = chunk_item[blockIdx.x];
  const int64_t start = chunk_start[blockIdx.x];
  int64_t len = nbytes[item] - start;
  if (len > DDP_BUCKET_CHUNK_BYTES) len = DDP_BUCKET_CHUNK_BYTES;
  char* tensor = reinterpret_cast<char*>(ptrs[item]) + start;
  char* slot = bucket + offsets[item] + start;
  const char* src = direction == 0 ? tensor : slot;
  char* dst = direction == 0 ? slot : tensor;
  const int tid = threadIdx.x;
  if (((reinterpret_cast<uintptr_t>(src) | reinterpret_cast<uintptr_t>(dst)) & 15) == 0) {
    const int64_t nvec = len >> 4;
    // 4 independent 128-bit loads in flight per thread
    int64_t i = tid;
    for (; i + 768 < nvec; i += 1024) {
      uint4 v0 = ld_stream16(src + (i << 4));
      uint4 v1 = ld_stream16(src + ((i + 256) << 4));
      uint4 v2 = ld_stream16(src + ((i + 512) << 4));
      uint4 v3 = ld_stream16(src + ((i + 768) << 4));
      st_stream16(dst + (i << 4), v0);
      st_stream16(dst + ((i + 256) << 4), v1);
      st_stream16(dst + ((i + 512) << 4), v2);
      st_stream16(dst + ((i + 768) << 4), v3);
    }
    for (; i < nvec; i += 256) st_stream16(dst + (i << 4), ld_stream16(src + (i << 4)));
    for (int64_t j = (nvec << 4) + tid; j < len; j += 256) dst[j] = src[j];
  } else if (((reinterpret_cast<uintptr_t>(src) | reinterpret_cast<uintptr_t>(dst)) & 3) == 0) {
    const int64_t nw = len >> 2;
    for (int64_t i = tid; i < nw; i += 256)
      reinterpret_cast<uint32_t*>(dst)[i] = reinterpret_cast<const uint32_t*>(src)[i];
    for (int64_t j = (nw << 2) + tid; j < len; j += 256) dst[j] = src[j];
  } else {
    for (int64_t j = tid; j < len; j += 256) dst[j] = src[j];
  }
}

}  // namespace ddp

using namespace ddp;

extern "C" int ddp_bucket_copy(const uint64_t* ptrs, const int64_t* offsets, const int64_t* nbytes,
                               const int32_t* chunk_item, const int64_t* chunk_start, int64_t nchunks,
                               void* bucket, int direction, void* stream_) {
  DDP_REQUIRE(nchunks >= 0 && nchunks < (int64_t(1) << 31), "nchunks out of range");
  if (nchunks == 0) return 0;
  DDP_REQUIRE(ptrs && offsets && nbytes && chunk_item && chunk_start && bucket, "null table or bucket");
  DDP_REQUIRE(direction == 0 || direction == 1, "direction must be 0 (flatten) or 1 (unflatten)");
  bucket_copy_kernel<<<int(nchunks), 256, 0, static_cast<cudaStream_t>(stream_)>>>(
      ptrs, offsets, nbytes, chunk_item, chunk_start, static_cast<char*>(bucket), direction);
  DDP_CUDA(cudaGetLastError());
  return 0;
}
