// On-device synthetic 28x28 data (replaces ref: horovod/mnist_horovod.py:34-38, the
// torchvision MNIST download + ToTensor + Normalize((0.1307,), (0.3081,)) pipeline, which
// needs a network).  Sample `i` of the virtual dataset is a pure function of (seed, i):
//   pixels  u8[784]: bytes of Philox4x32-10(counter=(i_lo, i_hi, blk, 0), key=seed), blk=0..48
//   x = (float(u8) / 255 - 0.1307) / 0.3081      (IEEE fp32, same op order as torchvision)
//   label   Philox(counter=(i_lo, i_hi, 49, 0)).x % 10
// HBM-bound writer: 3,136 + 8 bytes per sample, 64 contiguous bytes per thread and loop trip.
#include "common.cuh"

namespace ddp {

// One CTA generates groups of SY_S samples.  What does not depend on the pixel is computed once and shared:
//   * the stroke masks -- a function of (label, block) only: a 10 x 49 table per CTA (490 Philox calls, once);
//   * the sample's label (one Philox call per sample instead of one per 64 bytes);
//   * the ToTensor + Normalize value of each of the 256 possible pixel bytes (two IEEE divisions each, once),
// so the streaming part is ONE Philox call, 16 table look-ups and four 128-bit stores per 64 bytes written.
constexpr int SY_S = 32;                  // samples per group
constexpr int SY_THREADS = 256;
template <bool F32>
__global__ void __launch_bounds__(SY_THREADS)
synth_kernel(uint64_t seed, const int64_t* __restrict__ indices, int64_t base_index, int64_t batch,
             float* __restrict__ out_x, uint8_t* __restrict__ out_px, int64_t* __restrict__ out_y) {
  __shared__ uint4 stroke_s[10 * 49];                  // per (label, block): the 16 stroke offsets (0 or 112) as 16 bytes
  __shared__ float lut_s[256];
  __shared__ uint32_t label_s[SY_S];
  const uint32_t k0 = uint32_t(seed), k1 = uint32_t(seed >> 32);
  const int tid = threadIdx.x;
  for (int i = tid; i < 10 * 49; i += SY_THREADS) {
    const uint32_t bits = philox4x32_10(uint32_t(i / 49), uint32_t(i % 49), 0x7E3Du, 0x5EEDu, k0, k1).x;
    uint32_t a[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      a[q] = 0u;
#pragma unroll
      for (int k = 0; k < 4; ++k) a[q] |= (((bits >> (4 * q + k)) & 1u) ? 112u : 0u) << (8 * k);
    }
    stroke_s[i] = make_uint4(a[0], a[1], a[2], a[3]);
  }
  if (F32) lut_s[tid] = (float(tid) / 255.0f - 0.1307f) / 0.3081f;      // torchvision's op order, IEEE fp32
  const int64_t ngroups = (batch + SY_S - 1) / SY_S;
  for (int64_t grp = blockIdx.x; grp < ngroups; grp += gridDim.x) {
    const int64_t b0 = grp * SY_S;
    const int nvalid = int(batch - b0 < SY_S ? batch - b0 : SY_S);
    __syncthreads();                                  // tables ready / the previous group's labels consumed
    if (tid < nvalid) {
      const uint64_t idx = uint64_t(indices ? indices[b0 + tid] : base_index + b0 + tid);
      const uint32_t label = philox4x32_10(uint32_t(idx), uint32_t(idx >> 32), 49u, 0u, k0, k1).x % 10u;
      label_s[tid] = label;
      if (out_y) out_y[b0 + tid] = int64_t(label);
    }
    __syncthreads();
    const int nwork = nvalid * 49;
    for (int t0 = 0; t0 < nwork; t0 += SY_THREADS) {         // warp-uniform trip count: the f32 store path shuffles
      const int t = min(t0 + tid, nwork - 1);                 // tail lanes redo the last block (same bytes, harmless)
      const int sl = t / 49, blk = t % 49;
      const int64_t b = b0 + sl;
      const uint64_t idx = uint64_t(indices ? indices[b] : base_index + b);
      const Philox4 n = philox4x32_10(uint32_t(idx), uint32_t(idx >> 32), uint32_t(blk), 0u, k0, k1);
      // pixel = (noise byte >> 1) + stroke offset, four pixels per 32-bit word (127 + 112 < 256: no carry between bytes)
      const uint4 add = stroke_s[label_s[sl] * 49 + blk];
      uint32_t w[4];
      w[0] = ((n.x >> 1) & 0x7f7f7f7fu) + add.x;
      w[1] = ((n.y >> 1) & 0x7f7f7f7fu) + add.y;
      w[2] = ((n.z >> 1) & 0x7f7f7f7fu) + add.z;
      w[3] = ((n.w >> 1) & 0x7f7f7f7fu) + add.w;
      if (F32) {
        // 64 output bytes per thread: stored directly, the lanes of one 128-bit store would sit 64 bytes apart (every
        // 32-byte sector written in two halves by different instructions).  Instead the 4 lanes of a quad take turns:
        // in turn s they write the four float4 of lane s's block, i.e. 64 contiguous bytes per quad and instruction.
        const int lane = tid & 31, j = lane & 3, qbase = lane & ~3;
        float* ob = out_x + b * IMG_PIX + blk * 16;
#pragma unroll
        for (int sx = 0; sx < 4; ++sx) {
          const uint32_t x0 = __shfl_sync(0xffffffffu, w[0], qbase + sx), x1 = __shfl_sync(0xffffffffu, w[1], qbase + sx);
          const uint32_t x2 = __shfl_sync(0xffffffffu, w[2], qbase + sx), x3 = __shfl_sync(0xffffffffu, w[3], qbase + sx);
          const unsigned long long p = (unsigned long long)__shfl_sync(0xffffffffu, (unsigned long long)(uintptr_t)ob, qbase + sx);
          const uint32_t wj = j == 0 ? x0 : (j == 1 ? x1 : (j == 2 ? x2 : x3));
          reinterpret_cast<float4*>(reinterpret_cast<float*>(uintptr_t(p)))[j] =
              make_float4(lut_s[wj & 0xffu], lut_s[(wj >> 8) & 0xffu], lut_s[(wj >> 16) & 0xffu], lut_s[wj >> 24]);
        }
      } else {
        *reinterpret_cast<uint4*>(out_px + b * IMG_PIX + blk * 16) = make_uint4(w[0], w[1], w[2], w[3]);
      }
    }
  }
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
  const int64_t ngroups = (batch + SY_S - 1) / SY_S;
  const int cap = sm_count() * 8;
  synth_kernel<false><<<int(ngroups < cap ? ngroups : cap), SY_THREADS, 0, static_cast<cudaStream_t>(stream_)>>>(
      seed, indices, base_index, batch, nullptr, out_px, out_y);
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
  const int64_t ngroups = (batch + SY_S - 1) / SY_S;
  const int cap = sm_count() * 8;
  synth_kernel<true><<<int(ngroups < cap ? ngroups : cap), SY_THREADS, 0, static_cast<cudaStream_t>(stream_)>>>(
      seed, indices, base_index, batch, out_x, nullptr, out_y);
  DDP_CUDA(cudaGetLastError());
  return 0;
}
