// Sparse forms of the two conv2 backward contractions (autograd of ref: horovod/mnist_horovod.py:20,63).
//
// The gradient that reaches conv2's output has passed through ReLU, a 2x2 max-pool and Dropout2d, so of the
// 20 x 8 x 8 values of dy2 at most ONE per pool window is non-zero (16 per channel), none at all in a channel
// that Dropout2d dropped, and none where the pooled activation was clipped by ReLU.  The dense kernels in
// net_bwd.cu rebuild the 8x8 planes and multiply the zeros; here every live (channel, pool cell) pair -- an
// "entry": gradient g, position (oy, ox) = (2*py + idy, 2*px + idx) -- is visited once and costs exactly its
// 250 multiply-adds:
//   wgrad  dW[co][ci][kh][kw] += g * a1[ci][oy+kh][ox+kw]          (a 10x5x5 window of a1 per entry)
//   dgrad  da1[ci][oy+kh][ox+kw] += g * W[co][ci][kh][kw]          (a 10x5x5 stamp of W per entry)
// One operand of every multiply-add comes from shared memory at a data-dependent address (one 8-byte load feeds one
// packed FFMA2), so the kernels are bound by instruction issue / the shared-memory pipe instead of the FMA pipe --
// but they do a twelfth of the dense work in training mode (1/4 from the pool, ~1/2 from Dropout2d, the rest ReLU).
// All layouts below are chosen so that the active lanes of every LDS.64 hit distinct banks.
// Summation orders are fixed (sample order per CTA, entry order per sample): reruns are bit-identical.
#include <stdlib.h>
#include "common.cuh"

namespace ddp {

__device__ __forceinline__ void mbarrier_arrive(uint64_t* bar) {
  asm volatile("{\n\t.reg .b64 st;\n\tmbarrier.arrive.shared::cta.b64 st, [%0];\n\t}" :: "r"(smem_addr_u32(bar)) : "memory");
}

// ---------------------------------------------------------------------------
// K_B2w (sparse) conv2 weight/bias gradient.
// CTA = 20 consumer warps (warp <-> output channel co) + 2 producer warps, persistent over the samples b = blockIdx,
// blockIdx + grid, ...  Consumer lane = cp*5 + kw (cp = input-channel pair, 25 active lanes) holds the 10
// accumulators dW[co][2cp+c][kh][kw] (c < 2 packed in one 64-bit register, kh < 5) for ALL samples the CTA sees.
// Per live entry: 5 LDS.64 + 5 packed FFMA2; entries are taken two at a time so that 10 loads are in flight.
// Shared memory per sample ("slot", ring of SW_SLOTS): a1 as channel-pair-interleaved planes [cp][12x12][c] with a
// pair stride of 149 eight-byte words (== 5 mod 16: the lanes of a half-warp read distinct 8-byte banks), and per
// channel the COMPACTED list of its live entries (g, position oy*12+ox), zero-filled to 16.  The raw tiles (a1, a2,
// da2, idx2) arrive by cp.async.bulk into a 4-deep raw ring; the producer warps (even / odd samples) turn them into
// slots.  There is NO block-wide barrier in the loop: Dropout2d kills half of the (sample, channel) pairs, so the
// consumer warps have very different amounts of work per sample -- full/empty mbarriers per slot let them drift
// apart by up to the ring depth.
// ---------------------------------------------------------------------------
constexpr int SW_CONS = C2;                        // consumer warps
constexpr int SW_THREADS = (SW_CONS + 2) * 32;     // 704
constexpr int SW_SLOTS = 6;
constexpr int SW_RAW = 4;                          // raw stages (even: a producer warp always meets its own stages)
constexpr int SW_PS2 = 149;                        // channel-pair stride in 8-byte words
constexpr int SW_A1W = 5 * SW_PS2 + 3;             // 748 words per slot
constexpr int SW_SLOT_BYTES = SW_A1W * 8 + C2 * 16 * 8 + 32;              // a1p | list[20][16] (g, pos) | cnt[20] u8 = 8576
constexpr int SW_RAW_BYTES = A1_PER * 4 + 2 * A2_PER * 4 + A2_PER;        // a1 | a2 | da2 | idx2 = 8640
constexpr int SW_SMEM = SW_SLOTS * SW_SLOT_BYTES + SW_RAW * SW_RAW_BYTES + (2 * SW_SLOTS + SW_RAW) * 8;
__global__ void __launch_bounds__(SW_THREADS, 2)
conv2_wgrad_sp_kernel(const float* __restrict__ da2, const float* __restrict__ a2,
                      const uint8_t* __restrict__ idx2, const float* __restrict__ a1,
                      float drop_scale, float* __restrict__ partials, int B) {
  extern __shared__ __align__(128) unsigned char smem_b[];
  unsigned char* slots = smem_b;
  unsigned char* raws = slots + SW_SLOTS * SW_SLOT_BYTES;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(raws + SW_RAW * SW_RAW_BYTES);   // [SLOTS] producer -> consumers
  uint64_t* empty_bar = full_bar + SW_SLOTS;                                         // [SLOTS] 20 consumers -> producer
  uint64_t* raw_bar = empty_bar + SW_SLOTS;                                          // [RAW]   bulk copies landed
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  pdl_trigger();
  if (tid == 0) {
    for (int i = 0; i < SW_SLOTS; ++i) { mbarrier_init(&full_bar[i], 1); mbarrier_init(&empty_bar[i], SW_CONS); }
    for (int i = 0; i < SW_RAW; ++i) mbarrier_init(&raw_bar[i], 1);
  }
  for (int i = tid; i < SW_SLOTS * SW_SLOT_BYTES / 4; i += SW_THREADS) reinterpret_cast<float*>(slots)[i] = 0.f;   // pads stay finite
  __syncthreads();
  pdl_wait();                                          // da2 (head_bwd) is read from here on
  const int n = (B - int(blockIdx.x) + int(gridDim.x) - 1) / int(gridDim.x);   // samples of this CTA
  if (warp >= SW_CONS) {
    // ===================== producers: warp 20 takes the even, warp 21 the odd samples =====================
    const int p = warp - SW_CONS;
    auto issue = [&](int k) {                          // lane 0: raw tiles of the CTA's k-th sample
      const size_t b = size_t(blockIdx.x) + size_t(k) * gridDim.x;
      unsigned char* r = raws + (k % SW_RAW) * SW_RAW_BYTES;
      uint64_t* bar = &raw_bar[k % SW_RAW];
      mbarrier_expect_tx(bar, SW_RAW_BYTES);
      bulk_copy_g2s(r, a1 + b * A1_PER, A1_PER * 4, bar);
      bulk_copy_g2s(r + A1_PER * 4, a2 + b * A2_PER, A2_PER * 4, bar);
      bulk_copy_g2s(r + A1_PER * 4 + A2_PER * 4, da2 + b * A2_PER, A2_PER * 4, bar);
      bulk_copy_g2s(r + A1_PER * 4 + 2 * A2_PER * 4, idx2 + b * A2_PER, A2_PER, bar);
    };
    if (lane == 0) {
      if (p < n) issue(p);
      if (p + 2 < n) issue(p + 2);
    }
    const unsigned half_lt = (1u << (lane & 15)) - 1u;   // lanes below this one inside its half-warp
    for (int k = p; k < n; k += 2) {
      const int slot = k % SW_SLOTS, rs = k % SW_RAW;
      mbarrier_wait(&raw_bar[rs], (k / SW_RAW) & 1);
      if (k >= SW_SLOTS) mbarrier_wait(&empty_bar[slot], ((k / SW_SLOTS) - 1) & 1);
      const float* ra = reinterpret_cast<const float*>(raws + rs * SW_RAW_BYTES);
      unsigned long long* dst = reinterpret_cast<unsigned long long*>(slots + slot * SW_SLOT_BYTES);
      // a1 [ci][144] -> [cp][144][c]: lanes walk consecutive positions (conflict-free LDS and STS.64)
#pragma unroll
      for (int cpi = 0; cpi < 5; ++cpi) {
        const float* r0 = ra + (2 * cpi) * 144;
        unsigned long long* d = dst + cpi * SW_PS2;
#pragma unroll
        for (int q = 0; q < 5; ++q) {
          const int pos = lane + 32 * q;
          if (pos < 144) d[pos] = pack2(r0[pos], r0[pos + 144]);
        }
      }
      // entries: a half-warp owns one channel (16 pool cells); live ones (g != 0) are packed to the front of the
      // channel's list in cell order, dead ones become zero entries behind them
      {
        const float* e_a2 = ra + A1_PER;
        const float* e_da2 = e_a2 + A2_PER;
        const uint8_t* e_id = reinterpret_cast<const uint8_t*>(e_da2 + A2_PER);
        unsigned long long* list = dst + SW_A1W;
        uint8_t* cnt = reinterpret_cast<uint8_t*>(list + C2 * 16);
#pragma unroll
        for (int q = 0; q < 10; ++q) {
          const int e = lane + 32 * q;               // = co*16 + pp
          const float gv = e_a2[e] > 0.f ? e_da2[e] * drop_scale : 0.f;
          const int id = e_id[e], pp = e & 15, co = e >> 4;
          const unsigned pos = unsigned((2 * (pp >> 2) + (id >> 1)) * P1 + 2 * (pp & 3) + (id & 1));
          const bool on = gv != 0.f;
          const unsigned m = (__ballot_sync(0xffffffffu, on) >> (lane & 16)) & 0xffffu;
          const int nlive = __popc(m);
          const int at = on ? __popc(m & half_lt) : nlive + __popc(~m & half_lt);
          list[co * 16 + at] = on ? ((unsigned long long)__float_as_uint(gv) | ((unsigned long long)pos << 32)) : 0ull;
          if ((lane & 15) == 0) cnt[co] = uint8_t(nlive);
        }
      }
      __syncwarp();
      if (lane == 0) {
        mbarrier_arrive(&full_bar[slot]);              // release: the slot is complete
        if (k + SW_RAW < n) {                          // this raw stage is consumed: refill it
          asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
          issue(k + SW_RAW);
        }
      }
    }
    return;
  }
  // ===================== consumers =====================
  const int co = warp;
  const int lt = lane < 25 ? lane : lane - 9;          // idle lanes shadow lanes 16..22 of their own half-warp (broadcast)
  const int cp = lt / 5, kw = lt % 5;
  const int lane_w = cp * SW_PS2 + kw;
  unsigned long long acc[5];
#pragma unroll
  for (int k = 0; k < 5; ++k) acc[k] = 0ull;
  float bacc = 0.f;                                    // bias gradient: every lane adds the same g's in list order
  for (int k = 0; k < n; ++k) {
    const int slot = k % SW_SLOTS;
    mbarrier_wait(&full_bar[slot], (k / SW_SLOTS) & 1);
    const unsigned long long* a1w = reinterpret_cast<const unsigned long long*>(slots + slot * SW_SLOT_BYTES);
    const ulonglong2* list = reinterpret_cast<const ulonglong2*>(a1w + SW_A1W + co * 16);
    const int cnt = reinterpret_cast<const uint8_t*>(a1w + SW_A1W + C2 * 16)[co];
    const unsigned long long* base = a1w + lane_w;
    for (int i = 0; i < cnt; i += 2) {                 // two entries per trip (an odd count meets a zero entry)
      const ulonglong2 e = list[i >> 1];
      const float g0 = lo_f(e.x), g1 = lo_f(e.y);
      const unsigned long long* q0 = base + unsigned(e.x >> 32);
      const unsigned long long* q1 = base + unsigned(e.y >> 32);
      unsigned long long v0[5], v1[5];
#pragma unroll
      for (int kh = 0; kh < 5; ++kh) { v0[kh] = q0[kh * P1]; v1[kh] = q1[kh * P1]; }
      const unsigned long long g20 = pack2(g0, g0), g21 = pack2(g1, g1);
#pragma unroll
      for (int kh = 0; kh < 5; ++kh) fma2(acc[kh], g20, v0[kh]);
#pragma unroll
      for (int kh = 0; kh < 5; ++kh) fma2(acc[kh], g21, v1[kh]);
      bacc += g0;
      bacc += g1;
    }
    __syncwarp();
    if (lane == 0) mbarrier_arrive(&empty_bar[slot]);
  }
  float* part = partials + size_t(blockIdx.x) * P_ARENA;
  if (lane < 25) {
#pragma unroll
    for (int kh = 0; kh < 5; ++kh) {
      part[OFF_C2W + co * 250 + (2 * cp) * 25 + kh * 5 + kw] = lo_f(acc[kh]);
      part[OFF_C2W + co * 250 + (2 * cp + 1) * 25 + kh * 5 + kw] = hi_f(acc[kh]);
    }
  }
  if (lane == 0) part[OFF_C2B + co] = bacc;
}

// ---------------------------------------------------------------------------
// K_B2d (sparse) conv2 data gradient.  ONE warp per sample.
// The 12 output rows are produced in 4 bands: band py (the entries of pool row py) touches rows 2py .. 2py+5.
// lane = ry*5 + cp (ry = row inside the band, cp = input-channel pair; 30 active lanes) holds the band's
// accumulators acc[x 0..11] as packed pairs (channels 2cp, 2cp+1).  An entry (co, py, px; g, idy, idx) adds
//   g * W[co][2cp+c][ry-idy][kw]   to   acc[2px+idx+kw][c],   kw = 0..4
// W sits in shared memory as [co][kw][kh+1 (7)][cp][c] with ZERO rows at kh+1 in {0, 6}: rows outside the 5x5 window
// read zeros (no predicate), the entry's idy only moves the warp-uniform base address, and the lanes of one LDS.64
// read consecutive 8-byte words.  Per band the live entries of each pool cell are first COMPACTED into two lists in
// shared memory, one per idx, so the inner loops have static accumulator indices, no branch and no shuffle:
// per live entry 5 LDS.64 + 5 packed FFMA2 (fma.rn.f32x2), two entries per trip.  After a band, rows 2py and 2py+1
// are final and are written out; the rest moves down two rows with shuffles.  Entry data of band py+1 is fetched
// while band py runs.  Output dp1 = da1 * (a1 > 0): gradient w.r.t. the pooled conv1 map before ReLU.
// ---------------------------------------------------------------------------
constexpr int SD_MAXW = 10;                            // warps per CTA (launch picks 4..10)
constexpr int SD_CO2 = 175;                            // 8-byte words per output channel: 5 x 7 x 5
constexpr int SD_W2 = C2 * SD_CO2 + 8;                 // + over-read of the two idle lanes
constexpr int SD_LIST = 22;                            // entries per list: 20 + zero terminator, even
__device__ __forceinline__ void sd_emit(const unsigned long long (&acc)[12], int lane, int y, bool on,
                                        const float* __restrict__ a1b, float* __restrict__ dpb) {
  if (!on) return;
  const int cp = lane % 5;
  const int o0 = (2 * cp) * 144 + y * P1;
  float4 m0[3], m1[3];
#pragma unroll
  for (int q = 0; q < 3; ++q) {
    m0[q] = reinterpret_cast<const float4*>(a1b + o0)[q];
    m1[q] = reinterpret_cast<const float4*>(a1b + o0 + 144)[q];
  }
#pragma unroll
  for (int q = 0; q < 3; ++q) {
    float4 v, u;
    v.x = m0[q].x > 0.f ? lo_f(acc[4 * q]) : 0.f;      u.x = m1[q].x > 0.f ? hi_f(acc[4 * q]) : 0.f;
    v.y = m0[q].y > 0.f ? lo_f(acc[4 * q + 1]) : 0.f;  u.y = m1[q].y > 0.f ? hi_f(acc[4 * q + 1]) : 0.f;
    v.z = m0[q].z > 0.f ? lo_f(acc[4 * q + 2]) : 0.f;  u.z = m1[q].z > 0.f ? hi_f(acc[4 * q + 2]) : 0.f;
    v.w = m0[q].w > 0.f ? lo_f(acc[4 * q + 3]) : 0.f;  u.w = m1[q].w > 0.f ? hi_f(acc[4 * q + 3]) : 0.f;
    reinterpret_cast<float4*>(dpb + o0)[q] = v;
    reinterpret_cast<float4*>(dpb + o0 + 144)[q] = u;
  }
}
// all entries of one list: acc[X0 + kw] += g * W-row, two entries per trip
template <int X0>
__device__ __forceinline__ void sd_run_list(unsigned long long (&acc)[12], const ulonglong2* __restrict__ list, int cnt,
                                            const unsigned long long* __restrict__ wl) {
  for (int i = 0; i < cnt; i += 2) {
    const ulonglong2 e = list[i >> 1];
    const unsigned long long* w0 = wl + unsigned(e.x >> 32);
    const unsigned long long* w1 = wl + unsigned(e.y >> 32);
    unsigned long long v0[5], v1[5];
#pragma unroll
    for (int k = 0; k < 5; ++k) { v0[k] = w0[k * 35]; v1[k] = w1[k * 35]; }
    const unsigned long long g0 = pack2(lo_f(e.x), lo_f(e.x)), g1 = pack2(lo_f(e.y), lo_f(e.y));
#pragma unroll
    for (int k = 0; k < 5; ++k) fma2(acc[X0 + k], g0, v0[k]);
#pragma unroll
    for (int k = 0; k < 5; ++k) fma2(acc[X0 + k], g1, v1[k]);
  }
}
__global__ void __launch_bounds__(SD_MAXW * 32, 3)
conv2_dgrad_sp_kernel(const float* __restrict__ params, const float* __restrict__ da2,
                      const float* __restrict__ a2, const uint8_t* __restrict__ idx2,
                      const float* __restrict__ a1, float drop_scale, float* __restrict__ dp1, int B) {
  __shared__ __align__(16) unsigned long long w_s[SD_W2];
  __shared__ __align__(16) unsigned long long list_s[SD_MAXW][8][SD_LIST];     // per warp: [px*2 + idx][entry]
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int nthreads = blockDim.x, wpc = nthreads >> 5;
  pdl_trigger();
  for (int i = tid; i < SD_W2; i += nthreads) {
    float v0 = 0.f, v1 = 0.f;
    if (i < C2 * SD_CO2) {
      const int co = i / SD_CO2, r = i % SD_CO2, kw = r / 35, khp = (r % 35) / 5, cp = r % 5;
      if (khp >= 1 && khp <= 5) {
        const float* w = params + OFF_C2W + co * 250 + (2 * cp) * 25 + (khp - 1) * 5 + kw;
        v0 = w[0]; v1 = w[25];
      }
    }
    w_s[i] = pack2(v0, v1);
  }
  __syncthreads();
  pdl_wait();                                          // da2 (head_bwd) is read from here on
  const int ry = lane / 5;                             // 0..5 (6 for the two idle lanes)
  const unsigned long long* wl = w_s + lane;
  unsigned long long (*lists)[SD_LIST] = list_s[warp];
  const unsigned lt_mask = (1u << lane) - 1u;
  const int nwarps = gridDim.x * wpc;
  const int lco = lane < C2 ? lane : 0;                // lane co < 20 fetches its channel's entries
  for (int b = blockIdx.x * wpc + warp; b < B; b += nwarps) {
    const float4* a4 = reinterpret_cast<const float4*>(a2 + size_t(b) * A2_PER + lco * 16);
    const float4* d4 = reinterpret_cast<const float4*>(da2 + size_t(b) * A2_PER + lco * 16);
    const unsigned* i4 = reinterpret_cast<const unsigned*>(idx2 + size_t(b) * A2_PER + lco * 16);
    const float* a1b = a1 + size_t(b) * A1_PER;
    float* dpb = dp1 + size_t(b) * A1_PER;
    float4 na = a4[0], nd = d4[0];
    unsigned ni = i4[0];
    unsigned long long acc[12];
#pragma unroll
    for (int x = 0; x < 12; ++x) acc[x] = 0ull;
#pragma unroll 1
    for (int py = 0; py < 4; ++py) {                   // NOT unrolled: 8 list loops per band are 1/4 of the code this way
      // this band's 4 entries of channel `lane`
      float g[4];
      g[0] = na.x > 0.f ? nd.x * drop_scale : 0.f;
      g[1] = na.y > 0.f ? nd.y * drop_scale : 0.f;
      g[2] = na.z > 0.f ? nd.z * drop_scale : 0.f;
      g[3] = na.w > 0.f ? nd.w * drop_scale : 0.f;
      if (lane >= C2) { g[0] = 0.f; g[1] = 0.f; g[2] = 0.f; g[3] = 0.f; }
      const unsigned ids = ni;
      if (py < 3) { na = a4[py + 1]; nd = d4[py + 1]; ni = i4[py + 1]; }
      int cnt[8];
      __syncwarp();                                    // the previous band's lists are consumed
#pragma unroll
      for (int px = 0; px < 4; ++px) {
        const unsigned id = (ids >> (8 * px)) & 3u;    // idy = id >> 1, idx = id & 1
        const bool on = g[px] != 0.f;
        const unsigned m0 = __ballot_sync(0xffffffffu, on && !(id & 1u));
        const unsigned m1 = __ballot_sync(0xffffffffu, on && (id & 1u));
        cnt[2 * px] = __popc(m0);
        cnt[2 * px + 1] = __popc(m1);
        // entry = (g, word offset of W[co][kw=0][kh+1 = ry - idy + 1][cp] relative to this lane's base)
        const unsigned long long ent = (unsigned long long)__float_as_uint(g[px]) |
                                       ((unsigned long long)(unsigned(lane) * SD_CO2 + 5u - 5u * (id >> 1)) << 32);
        if (on) lists[2 * px + (id & 1u)][__popc(((id & 1u) ? m1 : m0) & lt_mask)] = ent;
        if (lane == 0) { lists[2 * px][cnt[2 * px]] = 0ull; lists[2 * px + 1][cnt[2 * px + 1]] = 0ull; }
      }
      __syncwarp();
      sd_run_list<0>(acc, reinterpret_cast<const ulonglong2*>(lists[0]), cnt[0], wl);
      sd_run_list<1>(acc, reinterpret_cast<const ulonglong2*>(lists[1]), cnt[1], wl);
      sd_run_list<2>(acc, reinterpret_cast<const ulonglong2*>(lists[2]), cnt[2], wl);
      sd_run_list<3>(acc, reinterpret_cast<const ulonglong2*>(lists[3]), cnt[3], wl);
      sd_run_list<4>(acc, reinterpret_cast<const ulonglong2*>(lists[4]), cnt[4], wl);
      sd_run_list<5>(acc, reinterpret_cast<const ulonglong2*>(lists[5]), cnt[5], wl);
      sd_run_list<6>(acc, reinterpret_cast<const ulonglong2*>(lists[6]), cnt[6], wl);
      sd_run_list<7>(acc, reinterpret_cast<const ulonglong2*>(lists[7]), cnt[7], wl);
      // rows 2py, 2py+1 are complete (after the last band: all six rows)
      sd_emit(acc, lane, 2 * py + ry, lane < (py < 3 ? 10 : 30), a1b, dpb);
#pragma unroll
      for (int x = 0; x < 12; ++x) {
        const unsigned long long v = __shfl_down_sync(0xffffffffu, acc[x], 10);
        acc[x] = lane < 20 ? v : 0ull;
      }
    }
  }
}

// ---------------------------------------------------------------------------
// K_B2d (sparse, channel-stationary).  Same lane mapping, bands and W layout as conv2_dgrad_sp_kernel, other loop order:
// per band the warp walks the LIVE CHANNELS (ballot over "any of my 4 cells has gradient"), loads the channel's W rows
// for both vertical phases ONCE (10 LDS.64: the 5 kw taps x {idy = 0, idy = 1}) and then applies the channel's up to four
// entries of the band from registers -- the entry's (px, phase) only selects one of 16 static FFMA2 groups through
// warp-uniform branches.  A live channel has ~3.7 live entries per band, so the shared-memory traffic per entry drops
// from 5 LDS.64 to ~2.7 and there is no list building at all.
// ---------------------------------------------------------------------------
template <int X0>
__device__ __forceinline__ void sd_apply(unsigned long long (&acc)[12], unsigned long long g2, const unsigned long long (&w)[5]) {
#pragma unroll
  for (int k = 0; k < 5; ++k) fma2(acc[X0 + k], g2, w[k]);
}
template <int PX>
__device__ __forceinline__ void sd_entry(unsigned long long (&acc)[12], float gb, unsigned id, const unsigned long long (&w0)[5],
                                         const unsigned long long (&w1)[5]) {
  if (gb == 0.f) return;                               // warp-uniform: gb was broadcast
  const unsigned long long g2 = pack2(gb, gb);
  if (id & 2u) {
    if (id & 1u) sd_apply<2 * PX + 1>(acc, g2, w1); else sd_apply<2 * PX>(acc, g2, w1);
  } else {
    if (id & 1u) sd_apply<2 * PX + 1>(acc, g2, w0); else sd_apply<2 * PX>(acc, g2, w0);
  }
}
__global__ void __launch_bounds__(SD_MAXW * 32, 3)
conv2_dgrad_cs_kernel(const float* __restrict__ params, const float* __restrict__ da2,
                      const float* __restrict__ a2, const uint8_t* __restrict__ idx2,
                      const float* __restrict__ a1, float drop_scale, float* __restrict__ dp1, int B, int wpc) {
  // wpc = warps per CTA that take samples; at small batches the CTA carries extra warps that only help with the staging
  __shared__ __align__(16) unsigned long long w_s[SD_W2];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int nthreads = blockDim.x;
  pdl_trigger();
  // conv2.weight -> [co][kw][kh + 1][cp][c] with zero rows 0 and 6: zero fill, then ONE coalesced pass over the 5,000
  // weights (a gather per shared-memory word cost 2 uncoalesced loads per item: ~10 us of prologue with 128 threads)
  for (int i = tid; i < SD_W2; i += nthreads) w_s[i] = 0ull;
  __syncthreads();
  {
    float* wf = reinterpret_cast<float*>(w_s);
#pragma unroll 4
    for (int j = tid; j < C2 * 250; j += nthreads) {
      const float v = params[OFF_C2W + j];
      const int co = j / 250, r = j % 250, ci = r / 25, t = r % 25, kh = t / 5, kw = t % 5;
      wf[(co * SD_CO2 + kw * 35 + (kh + 1) * 5 + (ci >> 1)) * 2 + (ci & 1)] = v;
    }
  }
  __syncthreads();
  pdl_wait();                                          // da2 (head_bwd) is read from here on
  const int ry = lane / 5;                             // 0..5 (6 for the two idle lanes)
  const unsigned long long* wl = w_s + lane;
  const int nwarps = gridDim.x * wpc;
  const int lco = lane < C2 ? lane : 0;                // lane co < 20 fetches its channel's entries
  if (warp >= wpc) return;
  for (int b = blockIdx.x * wpc + warp; b < B; b += nwarps) {
    const float4* a4 = reinterpret_cast<const float4*>(a2 + size_t(b) * A2_PER + lco * 16);
    const float4* d4 = reinterpret_cast<const float4*>(da2 + size_t(b) * A2_PER + lco * 16);
    const unsigned* i4 = reinterpret_cast<const unsigned*>(idx2 + size_t(b) * A2_PER + lco * 16);
    const float* a1b = a1 + size_t(b) * A1_PER;
    float* dpb = dp1 + size_t(b) * A1_PER;
    float4 na = a4[0], nd = d4[0];
    unsigned ni = i4[0];
    unsigned long long acc[12];
#pragma unroll
    for (int x = 0; x < 12; ++x) acc[x] = 0ull;
#pragma unroll 1
    for (int py = 0; py < 4; ++py) {
      float g[4];                                      // this band's 4 entries of channel `lane`
      g[0] = na.x > 0.f ? nd.x * drop_scale : 0.f;
      g[1] = na.y > 0.f ? nd.y * drop_scale : 0.f;
      g[2] = na.z > 0.f ? nd.z * drop_scale : 0.f;
      g[3] = na.w > 0.f ? nd.w * drop_scale : 0.f;
      if (lane >= C2) { g[0] = 0.f; g[1] = 0.f; g[2] = 0.f; g[3] = 0.f; }
      const unsigned ids = ni;
      if (py < 3) { na = a4[py + 1]; nd = d4[py + 1]; ni = i4[py + 1]; }
      unsigned live = __ballot_sync(0xffffffffu, (g[0] != 0.f) | (g[1] != 0.f) | (g[2] != 0.f) | (g[3] != 0.f));
      while (live) {
        const int co = __ffs(live) - 1;
        live &= live - 1;
        const unsigned long long* wp = wl + co * SD_CO2;
        unsigned long long w0[5], w1[5];               // rows kh = ry (idy = 0) and kh = ry - 1 (idy = 1); zero rows outside
#pragma unroll
        for (int k = 0; k < 5; ++k) { w0[k] = wp[k * 35 + 5]; w1[k] = wp[k * 35]; }
        const unsigned idb = __shfl_sync(0xffffffffu, ids, co);
        const float g0 = __shfl_sync(0xffffffffu, g[0], co), g1 = __shfl_sync(0xffffffffu, g[1], co);
        const float g2 = __shfl_sync(0xffffffffu, g[2], co), g3 = __shfl_sync(0xffffffffu, g[3], co);
        sd_entry<0>(acc, g0, idb & 3u, w0, w1);
        sd_entry<1>(acc, g1, (idb >> 8) & 3u, w0, w1);
        sd_entry<2>(acc, g2, (idb >> 16) & 3u, w0, w1);
        sd_entry<3>(acc, g3, (idb >> 24) & 3u, w0, w1);
      }
      // rows 2py, 2py+1 are complete (after the last band: all six rows)
      sd_emit(acc, lane, 2 * py + ry, lane < (py < 3 ? 10 : 30), a1b, dpb);
#pragma unroll
      for (int x = 0; x < 12; ++x) {
        const unsigned long long v = __shfl_down_sync(0xffffffffu, acc[x], 10);
        acc[x] = lane < 20 ? v : 0ull;
      }
    }
  }
}

// ---------------------------------------------------------------------------
int launch_conv2_wgrad_sp(const float* da2, const float* a2, const uint8_t* idx2, const float* a1, float drop_scale,
                          float* partials, int B, int* parts_out, cudaStream_t stream, bool pdl) {
  static bool done[64] = {false};
  int dev = 0;
  DDP_CUDA(cudaGetDevice(&dev));
  if (dev >= 0 && dev < 64 && !done[dev]) {
    DDP_CUDA(cudaFuncSetAttribute(conv2_wgrad_sp_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SW_SMEM));
    done[dev] = true;
  }
  static int mult = -1;
  if (mult < 0) { const char* e = getenv("DDP_W2_CTAS_PER_SM"); mult = (e && e[0] == '1') ? 1 : 2; }
  int grid = sm_count() * mult;
  static int min_spc = -1;                            // samples per CTA at small batches (latency vs partial rows to reduce)
  if (min_spc < 0) { const char* e = getenv("DDP_W2_MIN_SPC"); min_spc = e ? atoi(e) : 0; if (min_spc < 0) min_spc = 0; }
  // 8 samples per CTA keep the number of partial rows down; in the small-batch latency mode (B <= 1024, this kernel
  // runs beside conv2_dgrad -> conv1_wgrad on the side stream) 4 per CTA finish sooner (measured: 118.8 -> 112.6 us at B = 768)
  const int spc = min_spc > 0 ? min_spc : (B <= 1024 ? 4 : 8);
  if (grid > (B + spc - 1) / spc) grid = (B + spc - 1) / spc;
  if (grid < 1) grid = 1;
  if (grid > MAX_PARTS) grid = MAX_PARTS;
  *parts_out = grid;
  DDP_CUDA(launch_pdl(conv2_wgrad_sp_kernel, dim3(grid), dim3(SW_THREADS), SW_SMEM, stream, pdl, da2, a2, idx2, a1,
                      drop_scale, partials, B));
  return 0;
}

int launch_conv2_dgrad_sp(const float* params, const float* da2, const float* a2, const uint8_t* idx2, const float* a1,
                          float drop_scale, float* dp1, int B, cudaStream_t stream) {
  // one warp per sample; warps per CTA chosen so that the samples spread evenly over 3 CTAs per SM
  // (B = 4096: 28 per SM -> 8 warps per CTA, two rounds for a few CTAs); larger batches loop with 8 warps per CTA
  const int sms = sm_count();
  const int per_sm = (B + sms - 1) / sms;
  int wpc = (per_sm + 2) / 3;
  static int lists = -1;                               // A/B switch: DDP_D2_LISTS=1 = the entry-list kernel
  if (lists < 0) { const char* e = getenv("DDP_D2_LISTS"); lists = (e && e[0] == '1') ? 1 : 0; }
  if (wpc < (lists ? 4 : 1)) wpc = lists ? 4 : 1;
  if (wpc > SD_MAXW) wpc = SD_MAXW;
  // up to 4 samples per SM: ONE CTA per SM (every CTA stages the 20 KB of weights; measured at B = 512: 128 CTAs x 4
  // sample warps 90 us/step, 256 CTAs x 2 sample warps 94 us)
  if (!lists && B <= sms * 4) wpc = per_sm;
  const int nblk = (B + wpc - 1) / wpc;
  int grid = sms * 3;
  if (grid > nblk) grid = nblk;
  if (lists)
    DDP_CUDA(launch_pdl(conv2_dgrad_sp_kernel, dim3(grid), dim3(wpc * 32), 0, stream, true, params, da2, a2, idx2, a1,
                        drop_scale, dp1, B));
  else
    DDP_CUDA(launch_pdl(conv2_dgrad_cs_kernel, dim3(grid), dim3((wpc < 8 ? 8 : wpc) * 32), 0, stream, true, params, da2, a2, idx2, a1,
                        drop_scale, dp1, B, wpc));               // >= 8 warps stage the weights
  return 0;
}

}  // namespace ddp
