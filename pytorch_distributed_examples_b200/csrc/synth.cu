// On-device synthetic 28x28 data (replaces ref: horovod/mnist_horovod.py:34-38, the
// torchvision MNIST download + ToTensor + Normalize((0.1307,), (0.3081,)) pipeline, which
// needs a network).  Sample `i` of the virtual dataset is a pure function of (seed, i):
//   pixels  u8[784]: bytes of Philox4x32-10(counter=(i_lo, i_hi, blk, 0), key=seed), blk=0..48
//   x = (float(u8) / 255 - 0.1307) / 0.3081      (IEEE fp32, same op order as torchvision)
//   label   Philox(counter=(i_lo, i_hi, 49, 0)).x % 10
// HBM-bound writer: 3,136 + 8 bytes per sample, 64 contiguous bytes per thread.
#include "common.cuh"

namespace ddp {

// the 16 pixels of block `blk` of sample `idx`, packed little-endian in 4 words; label returned via ref
__device__ __forceinline__ void synth_block(uint64_t seed, uint64_t idx, int blk, uint32_t w[4], uint32_t& label) {
  const uint32_t k0 = uint32_t(seed), k1 = uint32_t(seed >> 32);
  label = philox4x32_10(uint32_t(idx), uint32_t(idx >> 32), 49u, 0u, k0, k1).x % 10u;
  if (blk >= 49) return;
  const Philox4 n = philox4x32_10(uint32_t(idx), uint32_t(idx >> 32), uint32_t(blk), 0u, k0, k1);
  const uint32_t stroke = philox4x32_10(label, uint32_t(blk), 0x7E3Du, 0x5EEDu, k0, k1).x;
  const uint32_t nz[4] = {n.x, n.y, n.z, n.w};
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    uint32_t out = 0u;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const uint32_t noise = (nz[q] >> (8 * k)) & 0xffu;
      const uint32_t px = (noise >> 1) + (((stroke >> (4 * q + k)) & 1u) ? 112u : 0u);
      out |= px << (8 * k);
    }
    w[q] = out;
  }
}

__global__ void __launch_bounds__(256)
synth_batch_kernel(uint64_t seed, const int64_t* __restrict__ indices, int64_t base_index, int64_t batch,
                   float* __restrict__ out_x, int64_t* __restrict__ out_y) {
  const int64_t t = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  const int64_t b = t / 50;
  const int blk = int(t % 50);
  if (b >= batch) return;
  const uint64_t idx = uint64_t(indices ? indices[b] : base_index + b);
  uint32_t w[4], label;
  synth_block(seed, idx, blk, w, label);
  if (blk == 49) {
    if (out_y) out_y[b] = int64_t(label);
    return;
  }
  float4* o = reinterpret_cast<float4*>(out_x + b * IMG_PIX + blk * 16);
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    float v[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      float u = float((w[q] >> (8 * k)) & 0xffu);
      v[k] = (u / 255.0f - 0.1307f) / 0.3081f;
    }
    o[q] = make_float4(v[0], v[1], v[2], v[3]);
  }
}

// Raw-pixel variant: the 784 bytes of each sample as the dataset stores them (MNIST is uint8 on disk).
__global__ void __launch_bounds__(256)
synth_batch_u8_kernel(uint64_t seed, const int64_t* __restrict__ indices, int64_t base_index, int64_t batch,
                      uint8_t* __restrict__ out_px, int64_t* __restrict__ out_y) {
  const int64_t t = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  const int64_t b = t / 50;
  const int blk = int(t % 50);
  if (b >= batch) return;
  const uint64_t idx = uint64_t(indices ? indices[b] : base_index + b);
  uint32_t w[4], label;
  synth_block(seed, idx, blk, w, label);
  if (blk == 49) {
    if (out_y) out_y[b] = int64_t(label);
    return;
  }
  *reinterpret_cast<uint4*>(out_px + b * IMG_PIX + blk * 16) = make_uint4(w[0], w[1], w[2], w[3]);
}

// ToTensor + Normalize on the device (ref: horovod/mnist_horovod.py:34-37): out = (u8/255 - mean)/std,
// IEEE fp32 in torchvision's op order.  16 pixels per thread: one 128-bit load, four 128-bit stores.
__global__ void __launch_bounds__(256)
normalize_u8_kernel(const uint8_t* __restrict__ in, float* __restrict__ out, int64_t n, float mean, float stdv) {
  const int64_t nvec = n >> 4;
  const int64_t stride = int64_t(gridDim.x) * blockDim.x;
  for (int64_t i = int64_t(blockIdx.x) * blockDim.x + threadIdx.x; i < nvec; i += stride) {
    const uint4 v = *reinterpret_cast<const uint4*>(in + (i << 4));
    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
    float4* o = reinterpret_cast<float4*>(out + (i << 4));
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      float f[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) f[k] = (float((w[q] >> (8 * k)) & 0xffu) / 255.0f - mean) / stdv;
      o[q] = make_float4(f[0], f[1], f[2], f[3]);
    }
  }
  if (blockIdx.x == 0 && threadIdx.x < (n & 15)) {
    const int64_t i = (nvec << 4) + threadIdx.x;
    out[i] = (float(in[i]) / 255.0f - mean) / stdv;
  }
}

}  // namespace ddp

using namespace ddp;

extern "C" int ddp_synth_batch_u8(uint64_t seed, const int64_t* indices, int64_t base_index, int64_t batch,
                                  uint8_t* out_px, int64_t* out_y, void* stream_) {
  DDP_REQUIRE(batch >= 0, "negative batch");
  if (batch == 0) return 0;
  DDP_REQUIRE(out_px && aligned16(out_px), "out_px must be non-null and 16-byte aligned");
  const int64_t threads = batch * 50;
  synth_batch_u8_kernel<<<int((threads + 255) / 256), 256, 0, static_cast<cudaStream_t>(stream_)>>>(
      seed, indices, base_index, batch, out_px, out_y);
  DDP_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int ddp_normalize_u8(const uint8_t* in, float* out, int64_t n, float mean, float stdv, void* stream_) {
  DDP_REQUIRE(n >= 0, "negative count");
  if (n == 0) return 0;
  DDP_REQUIRE(in && out && aligned16(in) && aligned16(out), "in/out must be non-null and 16-byte aligned");
  DDP_REQUIRE(stdv != 0.f, "std must be non-zero");
  int64_t want = ((n >> 4) + 255) / 256;
  int cap = sm_count() * 8;
  int grid = int(want < 1 ? 1 : (want > cap ? cap : want));
  normalize_u8_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream_)>>>(in, out, n, mean, stdv);
  DDP_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int ddp_synth_batch(uint64_t seed, const int64_t* indices, int64_t base_index, int64_t batch,
                               float* out_x, int64_t* out_y, void* stream_) {
  DDP_REQUIRE(batch >= 0, "negative batch");
  if (batch == 0) return 0;
  DDP_REQUIRE(out_x && aligned16(out_x), "out_x must be non-null and 16-byte aligned");
  const int64_t threads = batch * 50;
  synth_batch_kernel<<<int((threads + 255) / 256), 256, 0, static_cast<cudaStream_t>(stream_)>>>(
      seed, indices, base_index, batch, out_x, out_y);
  DDP_CUDA(cudaGetLastError());
  return 0;
}
