// K_B1 conv1 weight / bias gradient, column walker on packed pairs (ref: horovod/mnist_horovod.py:12 conv1, autograd of :19).
//   dW[c][kh][kw] = sum_b sum_oy,ox dy1[b][c][oy][ox] * x[b][oy+kh][ox+kw]        db[c] = sum dy1[b][c][.][.]
// dy1 has one non-zero per 2x2 pool window (dp1 = its value, idx1 = its position).  The scalar kernel rebuilt dy1 dense
// in shared memory and was bound by issue slots; a first packed version (channel pairs, dense dy1 in smem) was bound by
// the shared-memory pipe (1.8 k wavefronts per sample against 1.2 k FMA-pipe cycles).  This one needs neither:
//   thread <-> (channel c, pool column px) of one sample, walking down the 24 output rows.
//   * the two horizontally adjacent outputs (oy, 2px) and (oy, 2px+1) are ONE packed operand p = (dy[oy][2px], dy[oy][2px+1]),
//     built in registers from (dp1, idx1) of the thread's own pool cell -- no dense dy1, no scatter phase;
//   * the x window x[oy .. oy+4][2px .. 2px+5] lives in registers as 5 rows x 5 overlapping pairs (x[j], x[j+1]); going down
//     one row loads ONE new row and reuses four -- 25 fma.rn.f32x2 per 5 LDS.64.  The odd pairs (x[2px+1], x[2px+2]) come
//     from a copy of the image shifted by one pixel, written once per sample (assembling them from the even pairs cost two
//     MOVs in front of nearly every FFMA2 that used them: ptxas re-materialises instead of keeping the pair);
//   * acc[kh][kw] holds (sum over even ox, sum over odd ox) and is folded at the end.
// 128 threads (120 used) per sample; a CTA runs two such groups, each with its own double-buffered cp.async.bulk ring and
// its own named barrier, so there is no block-wide barrier in the loop.  Fixed-order reductions: deterministic.
#include "common.cuh"

namespace ddp {

constexpr int WC_GROUPS = 2;
constexpr int WC_GT = 128;                                  // threads per group (one sample in flight)
constexpr int WC_THREADS = WC_GROUPS * WC_GT;
constexpr int WC_USED = C1 * 12;                            // (c, px)
constexpr int WC_RAW = IMG_PIX * 4 + A1_PER * 4 + A1_PER;   // bytes per raw slot: x | dp1 | idx1  (10336, 16-byte multiple)
constexpr int WC_X1 = IMG_PIX * 4;                            // bytes of the shifted image copy of one group
constexpr int WC_SMEM = WC_GROUPS * 2 * (WC_RAW + WC_X1) + 64;
static_assert(WC_RAW % 16 == 0, "raw slot must keep 16-byte alignment");
static_assert(WC_GROUPS * 2 * WC_RAW >= WC_GROUPS * WC_USED * 26 * 4, "the final reduction reuses the raw slots");

__device__ __forceinline__ void group_sync(int g) { asm volatile("bar.sync %0, %1;" :: "r"(1 + g), "n"(WC_GT) : "memory"); }

__global__ void __launch_bounds__(WC_THREADS, 2)
conv1_wgrad_col_kernel(const float* __restrict__ dp1, const uint8_t* __restrict__ idx1, const float* __restrict__ x,
                       float* __restrict__ partials, int B) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  float* x1_g = reinterpret_cast<float*>(smem_raw + WC_GROUPS * 2 * WC_RAW) + (threadIdx.x / WC_GT) * 2 * IMG_PIX;   // [2] x1[i] = x[i+1]
  uint64_t* bar = reinterpret_cast<uint64_t*>(smem_raw + WC_GROUPS * 2 * (WC_RAW + WC_X1));     // [group][2]
  const int tid = threadIdx.x, g = tid / WC_GT, gt = tid % WC_GT;
  unsigned char* raw_g = smem_raw + g * 2 * WC_RAW;
  pdl_trigger();
  pdl_wait();
  const bool active = gt < WC_USED;
  const int c = active ? gt / 12 : 0, px = gt % 12;
  const int stride = gridDim.x * WC_GROUPS;                  // samples between two visits of this group
  const int b_first = blockIdx.x * WC_GROUPS + g;
  auto issue = [&](int b, int slot) {                        // group thread 0: prefetch one sample's x, dp1, idx1
    unsigned char* r = raw_g + slot * WC_RAW;
    mbarrier_expect_tx(&bar[g * 2 + slot], WC_RAW);
    bulk_copy_g2s(r, x + size_t(b) * IMG_PIX, IMG_PIX * 4, &bar[g * 2 + slot]);
    bulk_copy_g2s(r + IMG_PIX * 4, dp1 + size_t(b) * A1_PER, A1_PER * 4, &bar[g * 2 + slot]);
    bulk_copy_g2s(r + IMG_PIX * 4 + A1_PER * 4, idx1 + size_t(b) * A1_PER, A1_PER, &bar[g * 2 + slot]);
  };
  if (gt == 0) {
    mbarrier_init(&bar[g * 2], 1); mbarrier_init(&bar[g * 2 + 1], 1);
    if (b_first < B) issue(b_first, 0);
    if (b_first + stride < B) issue(b_first + stride, 1);
  }
  __syncthreads();
  unsigned long long acc[5][5];
#pragma unroll
  for (int kh = 0; kh < 5; ++kh)
#pragma unroll
    for (int k = 0; k < 5; ++k) acc[kh][k] = 0ull;
  float bsum = 0.f;
  // shifted copy of the image in raw slot `slot` (waits for the slot's bulk copies first)
  auto shift_copy = [&](int slot, uint32_t parity) {
    mbarrier_wait(&bar[g * 2 + slot], parity);
    const float* xs = reinterpret_cast<const float*>(raw_g + slot * WC_RAW);
    float* x1 = x1_g + slot * IMG_PIX;
#pragma unroll
    for (int q = 0; q < (IMG_PIX + WC_GT - 1) / WC_GT; ++q) {
      const int i = gt + q * WC_GT;
      if (i < IMG_PIX - 1) x1[i] = xs[i + 1];
    }
  };
  if (b_first < B) shift_copy(0, 0);
  group_sync(g);
  int it = 0;
  for (int b = b_first; b < B; b += stride, ++it) {
    const int slot = it & 1;
    if (active) {
      const unsigned char* r = raw_g + slot * WC_RAW;
      const float* xcol = reinterpret_cast<const float*>(r) + 2 * px;
      const float* x1col = x1_g + slot * IMG_PIX + 2 * px;
      const float* dpc = reinterpret_cast<const float*>(r + IMG_PIX * 4) + c * 144 + px;
      const uint8_t* ixc = r + IMG_PIX * 4 + A1_PER * 4 + c * 144 + px;
      unsigned long long W[5][5];                            // W[row % 5][k] = (x[row][2px+k], x[row][2px+k+1])
      auto load_row = [&](int row, unsigned long long (&w)[5]) {
        const unsigned long long* p = reinterpret_cast<const unsigned long long*>(xcol + row * IMG);
        const unsigned long long* p1 = reinterpret_cast<const unsigned long long*>(x1col + row * IMG);
        w[0] = p[0]; w[2] = p[1]; w[4] = p[2];
        w[1] = p1[0]; w[3] = p1[1];
      };
#pragma unroll
      for (int row = 0; row < 4; ++row) load_row(row, W[row]);
#pragma unroll
      for (int py = 0; py < 12; ++py) {
        const float gv = dpc[py * 12];
        const int id = ixc[py * 12];
        bsum += gv;
#pragma unroll
        for (int idy = 0; idy < 2; ++idy) {
          const int oy = 2 * py + idy;
          load_row(oy + 4, W[(oy + 4) % 5]);
          const unsigned long long p = pack2(id == 2 * idy ? gv : 0.f, id == 2 * idy + 1 ? gv : 0.f);
#pragma unroll
          for (int kh = 0; kh < 5; ++kh)
#pragma unroll
            for (int k = 0; k < 5; ++k) fma2(acc[kh][k], p, W[(oy + kh) % 5][k]);
        }
      }
    }
    // the next sample's slot was filled a whole sample ago: its shifted copy is written here, so the ONE group barrier per
    // sample covers both "done with this slot" and "next shifted copy complete"
    if (b + stride < B) shift_copy(slot ^ 1, ((it + 1) >> 1) & 1);
    group_sync(g);
    if (gt == 0 && b + 2 * stride < B) issue(b + 2 * stride, slot);
  }
  // fixed-order reduction over (group, px) for every (c, kh, kw); the raw slots are free once every group has left its loop
  __syncthreads();
  float* red = reinterpret_cast<float*>(smem_raw);           // [group][120][26]: 25 weights + bias share
  if (active) {
    float* rr = red + (g * WC_USED + gt) * 26;
#pragma unroll
    for (int kh = 0; kh < 5; ++kh)
#pragma unroll
      for (int k = 0; k < 5; ++k) rr[kh * 5 + k] = lo_f(acc[kh][k]) + hi_f(acc[kh][k]);
    rr[25] = bsum;
  }
  __syncthreads();
  float* part = partials + size_t(blockIdx.x) * P_ARENA;
  for (int o = tid; o < C1 * 26; o += WC_THREADS) {          // o = (channel, tap or bias)
    const int cc = o / 26, j = o % 26;
    float s = 0.f;
#pragma unroll
    for (int gg = 0; gg < WC_GROUPS; ++gg)
#pragma unroll
      for (int p = 0; p < 12; ++p) s += red[(gg * WC_USED + cc * 12 + p) * 26 + j];
    if (j < 25) part[OFF_C1W + cc * 25 + j] = s; else part[OFF_C1B + cc] = s;
  }
}

int launch_conv1_wgrad_col(const float* dp1, const uint8_t* idx1, const float* x, float* partials, int B, int parts,
                          cudaStream_t stream, bool pdl) {
  static bool attr_done[64] = {false};
  int dev = 0;
  DDP_CUDA(cudaGetDevice(&dev));
  if (dev >= 0 && dev < 64 && !attr_done[dev]) {
    DDP_CUDA(cudaFuncSetAttribute(conv1_wgrad_col_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, WC_SMEM));
    attr_done[dev] = true;
  }
  DDP_CUDA(launch_pdl(conv1_wgrad_col_kernel, dim3(parts), dim3(WC_THREADS), WC_SMEM, stream, pdl, dp1, idx1, x, partials, B));
  return 0;
}

}  // namespace ddp
