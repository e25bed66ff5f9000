// K_B1 conv1 weight / bias gradient on packed channel pairs (ref: horovod/mnist_horovod.py:12 conv1, autograd of :19).
//   dW[c][kh][kw] = sum_b sum_oy,ox dy1[b][c][oy][ox] * x[b][oy+kh][ox+kw]        db[c] = sum dy1[b][c][.][.]
// dy1 (24x24 per channel, one non-zero per 2x2 pool window) is rebuilt dense in shared memory from dp1 / idx1, but with
// the channels of a pair (2cp, 2cp+1) interleaved, so that one 64-bit shared-memory word feeds one fma.rn.f32x2: the
// issue slots, not the FMA pipe, limited the scalar kernel (1.9 k instructions per 1.2 k FFMA).
//   thread <-> (output row oy, kh, half row h): holds x[oy+kh][12h .. 12h+15] as 16 broadcast pairs and 25 packed
//   accumulators (5 channel pairs x 5 kw); per channel pair it reads its 12 dy pairs (6 LDS.128) and issues 60 FFMA2.
// One sample per stage; raw inputs (x, dp1, idx1) arrive through a 3-deep cp.async.bulk ring, the dense dy1 is double
// buffered, so there is ONE block barrier per sample.  The bias gradient is summed where dp1 is read (scatter phase).
#include "common.cuh"

namespace ddp {

constexpr int WP_THREADS = 256;
constexpr int WP_ACTIVE = 240;                              // (oy 24, half 2, kh 5)
constexpr int WP_DY = 5 * 576;                              // u64 per dy buffer: [cp][oy][ox]
constexpr int WP_RAW = IMG_PIX * 4 + A1_PER * 4 + A1_PER;   // bytes per raw slot: x | dp1 | idx1  (10336, 16-byte multiple)
constexpr int WP_RING = 3;
constexpr int WP_SMEM = 2 * WP_DY * 8 + WP_RING * WP_RAW + 64;
static_assert(WP_RAW % 16 == 0, "raw slot must keep 16-byte alignment");
static_assert(2 * WP_DY * 8 >= WP_ACTIVE * 15 * 8 && 2 * WP_DY * 8 >= WP_THREADS * 6 * 4, "final reductions reuse the dy buffers");

__global__ void __launch_bounds__(WP_THREADS, 2)
conv1_wgrad_pk_kernel(const float* __restrict__ dp1, const uint8_t* __restrict__ idx1, const float* __restrict__ x,
                      float* __restrict__ partials, int B) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  unsigned long long* dy_s = reinterpret_cast<unsigned long long*>(smem_raw);          // [2][5][24][24] channel pairs
  unsigned char* raw_s = smem_raw + 2 * WP_DY * 8;                                      // [3][x | dp1 | idx1]
  uint64_t* bar = reinterpret_cast<uint64_t*>(raw_s + WP_RING * WP_RAW);                // [3]
  const int tid = threadIdx.x;
  pdl_trigger();
  pdl_wait();
  const bool active = tid < WP_ACTIVE;
  const int kh = tid % 5, t5 = tid / 5, h = t5 & 1, oy = t5 >> 1;
  const int G = gridDim.x;
  auto issue = [&](int b, int slot) {                        // thread 0: prefetch one sample's x, dp1, idx1
    unsigned char* r = raw_s + slot * WP_RAW;
    mbarrier_expect_tx(&bar[slot], WP_RAW);
    bulk_copy_g2s(r, x + size_t(b) * IMG_PIX, IMG_PIX * 4, &bar[slot]);
    bulk_copy_g2s(r + IMG_PIX * 4, dp1 + size_t(b) * A1_PER, A1_PER * 4, &bar[slot]);
    bulk_copy_g2s(r + IMG_PIX * 4 + A1_PER * 4, idx1 + size_t(b) * A1_PER, A1_PER, &bar[slot]);
  };
  if (tid == 0) {
#pragma unroll
    for (int s = 0; s < WP_RING; ++s) mbarrier_init(&bar[s], 1);
    if (int(blockIdx.x) < B) issue(blockIdx.x, 0);
    if (int(blockIdx.x) + G < B) issue(blockIdx.x + G, 1);
  }
  __syncthreads();
  unsigned long long acc[5][5];
#pragma unroll
  for (int cp = 0; cp < 5; ++cp)
#pragma unroll
    for (int k = 0; k < 5; ++k) acc[cp][k] = 0ull;
  float bacc[3][2] = {{0.f, 0.f}, {0.f, 0.f}, {0.f, 0.f}};   // bias share of the (channel pair, cell) entries tid, tid+256, tid+512
  int src_off[3], dst_off[3];
#pragma unroll
  for (int q = 0; q < 3; ++q) {
    const int e = tid + q * WP_THREADS, cp = e / 144, cell = e % 144;
    src_off[q] = 2 * cp * 144 + cell;
    dst_off[q] = cp * 576 + (2 * (cell / 12)) * 24 + 2 * (cell % 12);
  }
  int it = 0;
  for (int b = blockIdx.x; b < B; b += G, ++it) {
    const int slot = it % WP_RING;
    mbarrier_wait(&bar[slot], (it / WP_RING) & 1);
    const unsigned char* r = raw_s + slot * WP_RAW;
    const float* x_s = reinterpret_cast<const float*>(r);
    const float* dp_s = reinterpret_cast<const float*>(r + IMG_PIX * 4);
    const uint8_t* ix_s = r + IMG_PIX * 4 + A1_PER * 4;
    unsigned long long* dy = dy_s + (it & 1) * WP_DY;
#pragma unroll
    for (int q = 0; q < 3; ++q) {
      if (tid + q * WP_THREADS < 720) {
        const float g0 = dp_s[src_off[q]], g1 = dp_s[src_off[q] + 144];
        const int i0 = ix_s[src_off[q]], i1 = ix_s[src_off[q] + 144];
        bacc[q][0] += g0; bacc[q][1] += g1;
        ulonglong2 top, bot;
        top.x = pack2(i0 == 0 ? g0 : 0.f, i1 == 0 ? g1 : 0.f);
        top.y = pack2(i0 == 1 ? g0 : 0.f, i1 == 1 ? g1 : 0.f);
        bot.x = pack2(i0 == 2 ? g0 : 0.f, i1 == 2 ? g1 : 0.f);
        bot.y = pack2(i0 == 3 ? g0 : 0.f, i1 == 3 ? g1 : 0.f);
        *reinterpret_cast<ulonglong2*>(dy + dst_off[q]) = top;
        *reinterpret_cast<ulonglong2*>(dy + dst_off[q] + 24) = bot;
      }
    }
    __syncthreads();               // dy[it&1] complete; everybody is done with sample it-1 (its raw slot and dy buffer)
    if (tid == 0 && b + 2 * G < B) issue(b + 2 * G, (it + 2) % WP_RING);
    if (active) {
      unsigned long long xx[16];
      const float* xp = x_s + (oy + kh) * IMG + 12 * h;
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const float4 v = *reinterpret_cast<const float4*>(xp + 4 * q);
        xx[4 * q] = pack2(v.x, v.x); xx[4 * q + 1] = pack2(v.y, v.y);
        xx[4 * q + 2] = pack2(v.z, v.z); xx[4 * q + 3] = pack2(v.w, v.w);
      }
      const unsigned long long* dbase = dy + oy * 24 + 12 * h;
#pragma unroll
      for (int cp = 0; cp < 5; ++cp) {
#pragma unroll
        for (int hf = 0; hf < 2; ++hf) {
          unsigned long long d[6];
#pragma unroll
          for (int q = 0; q < 3; ++q) {
            const ulonglong2 v = *reinterpret_cast<const ulonglong2*>(dbase + cp * 576 + hf * 6 + 2 * q);
            d[2 * q] = v.x; d[2 * q + 1] = v.y;
          }
#pragma unroll
          for (int j = 0; j < 6; ++j)
#pragma unroll
            for (int k = 0; k < 5; ++k) fma2(acc[cp][k], d[j], xx[hf * 6 + j + k]);
        }
      }
    }
  }
  // fixed-order reduction over the 48 (oy, half) threads of every (channel pair, kh, kw): channel pairs 0-2, then 3-4
  float* part = partials + size_t(blockIdx.x) * P_ARENA;
  unsigned long long* red = dy_s;                            // [240][15]
#pragma unroll
  for (int pass = 0; pass < 2; ++pass) {
    const int cp0 = pass * 3, ncp = pass == 0 ? 3 : 2;
    __syncthreads();
    if (active) {
#pragma unroll
      for (int c = 0; c < 3; ++c)
        if (c < ncp)
#pragma unroll
          for (int k = 0; k < 5; ++k) red[tid * 15 + c * 5 + k] = acc[cp0 + c][k];
    }
    __syncthreads();
    if (tid < ncp * 25) {                                    // (local channel pair, kh, kw)
      const int cl = tid / 25, kh2 = (tid % 25) / 5, kw = tid % 5;
      float s0 = 0.f, s1 = 0.f;
      for (int t = 0; t < 48; ++t) {
        const unsigned long long v = red[(t * 5 + kh2) * 15 + cl * 5 + kw];
        s0 += lo_f(v); s1 += hi_f(v);
      }
      part[OFF_C1W + (2 * (cp0 + cl)) * 25 + kh2 * 5 + kw] = s0;
      part[OFF_C1W + (2 * (cp0 + cl) + 1) * 25 + kh2 * 5 + kw] = s1;
    }
  }
  __syncthreads();
  float* redb = reinterpret_cast<float*>(dy_s);              // [768 entries][2]
#pragma unroll
  for (int q = 0; q < 3; ++q) { redb[(tid + q * WP_THREADS) * 2] = bacc[q][0]; redb[(tid + q * WP_THREADS) * 2 + 1] = bacc[q][1]; }
  __syncthreads();
  if (tid < C1) {                                            // channel c = entries [144*(c/2), 144*(c/2)+144), component c&1
    float sum = 0.f;
    for (int e = 0; e < 144; ++e) sum += redb[((tid >> 1) * 144 + e) * 2 + (tid & 1)];
    part[OFF_C1B + tid] = sum;
  }
}

int launch_conv1_wgrad_pk(const float* dp1, const uint8_t* idx1, const float* x, float* partials, int B, int parts,
                          cudaStream_t stream, bool pdl) {
  static bool attr_done[64] = {false};
  int dev = 0;
  DDP_CUDA(cudaGetDevice(&dev));
  if (dev >= 0 && dev < 64 && !attr_done[dev]) {
    DDP_CUDA(cudaFuncSetAttribute(conv1_wgrad_pk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, WP_SMEM));
    attr_done[dev] = true;
  }
  DDP_CUDA(launch_pdl(conv1_wgrad_pk_kernel, dim3(parts), dim3(WP_THREADS), WP_SMEM, stream, pdl, dp1, idx1, x, partials, B));
  return 0;
}

}  // namespace ddp
