// Backward kernels of `Net` (autograd of ref: horovod/mnist_horovod.py:18-25,62-63):
//   K_B3  head: log_softmax/NLL -> fc2 -> dropout/ReLU -> fc1      -> partial dW/db, da2
//   K_B2d conv2 dgrad (through ReLU/pool/Dropout2d)                 da2 -> dp1
//   K_B2w conv2 wgrad + bgrad                                       da2,a1 -> partials
//   K_B1  conv1 wgrad + bgrad (through ReLU/pool)                   dp1,x -> partials
//   K_R   fixed-order reduction of the per-CTA partials into the gradient bucket
// Weight gradients are accumulated in registers across all samples a CTA sees and
// written once per CTA; there are no floating-point atomics anywhere.
#include <stdlib.h>
#include "common.cuh"

namespace ddp {

// ---------------------------------------------------------------------------
// K_B3 head backward. Tile of 32 samples per iteration, 320 threads.
// ---------------------------------------------------------------------------
constexpr int HT = 32;
constexpr int HB_THREADS = 320;
constexpr int HB_A2S = 324;   // a2 tile row stride
constexpr int HB_DPS = 36;    // dpT row stride ([j][s])
constexpr int HB_HS = 53;     // h tile row stride
constexpr int HB_SMEM = (F1 * A2_PER + HT * HB_A2S + F1 * HB_DPS + HT * 12 + HT * HB_HS + NC * 52) * 4 + 16;
__global__ void __launch_bounds__(HB_THREADS, 1)
head_bwd_kernel(const float* __restrict__ params, const float* __restrict__ a2,
                const float* __restrict__ h, const float* __restrict__ logp,
                const float* __restrict__ dlogp, const int64_t* __restrict__ target,
                float nll_coef, float drop_scale, float* __restrict__ da2,
                float* __restrict__ partials, int B) {
  extern __shared__ __align__(128) float smem[];
  // fc1.weight (64 KB) and each a2 tile arrive by cp.async.bulk + mbarrier (no register staging)
  float* w1_s = smem;                     // [50][320] natural layout
  float* a2_s = w1_s + F1 * A2_PER;       // [32][324]
  float* dp_s = a2_s + HT * HB_A2S;       // [50][36]  dL/d(fc1 pre-activation), transposed
  float* dz_s = dp_s + F1 * HB_DPS;       // [32][12]
  float* h_s = dz_s + HT * 12;            // [32][53]
  float* w2_s = h_s + HT * HB_HS;         // [10][52]
  uint64_t* bar = reinterpret_cast<uint64_t*>(w2_s + NC * 52);
  const int tid = threadIdx.x;
  const int ntiles_all = (B + HT - 1) / HT;
  if (tid == 0) mbarrier_init(bar, 1);
  __syncthreads();
  auto issue_tile = [&](int tile, bool with_w1) {       // thread 0 only
    const int tb0 = tile * HT;
    const int nv = min(HT, B - tb0);
    mbarrier_expect_tx(bar, uint32_t(nv) * (A2_PER * 4) + (with_w1 ? F1 * A2_PER * 4 : 0));
    if (with_w1) {
      bulk_copy_g2s(w1_s, params + OFF_F1W, F1 * A2_PER * 4, bar);   // parameters: fetched while the previous kernel drains
      pdl_wait();
    }
    for (int s = 0; s < nv; ++s)
      bulk_copy_g2s(a2_s + s * HB_A2S, a2 + size_t(tb0 + s) * A2_PER, A2_PER * 4, bar);
  };
  pdl_trigger();
  if (tid == 0 && int(blockIdx.x) < ntiles_all) issue_tile(blockIdx.x, true);
  uint32_t phase = 0;
  for (int i = tid; i < NC * 52; i += HB_THREADS) {
    int c = i / 52, j = i % 52;
    w2_s[i] = j < F1 ? params[OFF_F2W + c * F1 + j] : 0.f;
  }
  // persistent accumulators
  unsigned long long accW1[5][5];          // fc1.weight rows jg*5.., cols ig*10.. as packed pairs (fma.rn.f32x2)
#pragma unroll
  for (int a = 0; a < 5; ++a)
#pragma unroll
    for (int c = 0; c < 5; ++c) accW1[a][c] = 0ull;
  float accW2[2] = {0.f, 0.f};             // fc2.weight element tid, tid+320
  float accB = 0.f;                        // tid<50: fc1.bias[tid]; 64<=tid<74: fc2.bias[tid-64]
  const int jgW = tid >> 5, igW = tid & 31;  // fc1 wgrad mapping
  const int sgD = tid / 40, igD = tid % 40;  // fc1 dgrad mapping
  const int ntiles = (B + HT - 1) / HT;
  pdl_wait();                               // h, logp (head_fwd) are read and da2 / partials written only after this
  for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const int b0 = tile * HT;
    const int nvalid = min(HT, B - b0);
    __syncthreads();                        // previous tile fully consumed (h_s, dz_s, dp_s)
    if (tile != int(blockIdx.x) && tid == 0) issue_tile(tile, false);   // first tile was issued before the loop
    for (int i = tid; i < (HT - nvalid) * A2_PER; i += HB_THREADS)       // rows past the batch: keep them zero
      a2_s[(nvalid + i / A2_PER) * HB_A2S + i % A2_PER] = 0.f;
    for (int i = tid; i < HT * F1; i += HB_THREADS) {
      int s = i / F1, j = i % F1;
      h_s[s * HB_HS + j] = s < nvalid ? h[size_t(b0) * F1 + i] : 0.f;
    }
    // dz = dlogp - exp(logp) * sum_c dlogp   (log_softmax backward)
    if (tid < HT) {
      const int s = tid;
      float g[NC], lp[NC];
      float sum = 0.f;
#pragma unroll
      for (int c = 0; c < NC; ++c) { g[c] = 0.f; lp[c] = 0.f; }
      if (s < nvalid) {
        const size_t o = size_t(b0 + s) * NC;
        if (dlogp) {
#pragma unroll
          for (int c = 0; c < NC; ++c) g[c] = dlogp[o + c];
        } else {
          const int t = int(target[b0 + s]);
#pragma unroll
          for (int c = 0; c < NC; ++c) g[c] = (c == t) ? nll_coef : 0.f;
        }
#pragma unroll
        for (int c = 0; c < NC; ++c) { lp[c] = logp[o + c]; sum += g[c]; }
      }
#pragma unroll
      for (int c = 0; c < NC; ++c)
        dz_s[s * 12 + c] = (s < nvalid) ? g[c] - expf(lp[c]) * sum : 0.f;
    }
    mbarrier_wait(bar, phase);              // a2 tile (and, first time, fc1.weight) have landed
    phase ^= 1u;
    __syncthreads();
    // dp[s][j] = relu'/dropout' * sum_c dz[s][c] * W2[c][j]
    for (int e = tid; e < HT * F1; e += HB_THREADS) {
      const int s = e / F1, j = e % F1;
      float dh = 0.f;
#pragma unroll
      for (int c = 0; c < NC; ++c) dh = fmaf(dz_s[s * 12 + c], w2_s[c * 52 + j], dh);
      dp_s[j * HB_DPS + s] = h_s[s * HB_HS + j] > 0.f ? dh * drop_scale : 0.f;
    }
    __syncthreads();
    // fc2 weight / bias gradients and fc1 bias gradient
#pragma unroll
    for (int q = 0; q < 2; ++q) {
      const int e = tid + q * HB_THREADS;
      if (e < NC * F1) {
        const int c = e / F1, j = e % F1;
        float a = accW2[q];
        for (int s = 0; s < HT; ++s) a = fmaf(dz_s[s * 12 + c], h_s[s * HB_HS + j], a);
        accW2[q] = a;
      }
    }
    if (tid < F1) {
      float a = accB;
      for (int s = 0; s < HT; ++s) a += dp_s[tid * HB_DPS + s];
      accB = a;
    } else if (tid >= 64 && tid < 64 + NC) {
      float a = accB;
      for (int s = 0; s < HT; ++s) a += dz_s[s * 12 + (tid - 64)];
      accB = a;
    }
    // fc1 weight gradient: dW1[j][i] += dp[s][j] * a2[s][i]
    {
      const float* ar = a2_s + igW * 10;
      const float* dr = dp_s + (jgW * 5) * HB_DPS;
#pragma unroll 2
      for (int s = 0; s < HT; ++s) {
        unsigned long long av[5];
#pragma unroll
        for (int q = 0; q < 5; ++q) av[q] = *reinterpret_cast<const unsigned long long*>(ar + s * HB_A2S + 2 * q);
#pragma unroll
        for (int a = 0; a < 5; ++a) {
          const float d = dr[a * HB_DPS + s];
          const unsigned long long d2 = pack2(d, d);
#pragma unroll
          for (int c = 0; c < 5; ++c) fma2(accW1[a][c], d2, av[c]);
        }
      }
    }
    // fc1 data gradient: da2[s][i] = sum_j dp[s][j] * W1[j][i]
    {
      unsigned long long acc[4][4];
#pragma unroll
      for (int a = 0; a < 4; ++a)
#pragma unroll
        for (int c = 0; c < 4; ++c) acc[a][c] = 0ull;
#pragma unroll 2
      for (int j = 0; j < F1; ++j) {
        const float4 d = *reinterpret_cast<const float4*>(dp_s + j * HB_DPS + sgD * 4);
        const ulonglong2 wa = *reinterpret_cast<const ulonglong2*>(w1_s + j * A2_PER + igD * 4);
        const ulonglong2 wb = *reinterpret_cast<const ulonglong2*>(w1_s + j * A2_PER + 160 + igD * 4);
        const unsigned long long dv[4] = {pack2(d.x, d.x), pack2(d.y, d.y), pack2(d.z, d.z), pack2(d.w, d.w)};
        const unsigned long long wv[4] = {wa.x, wa.y, wb.x, wb.y};
#pragma unroll
        for (int a = 0; a < 4; ++a)
#pragma unroll
          for (int c = 0; c < 4; ++c) fma2(acc[a][c], dv[a], wv[c]);
      }
#pragma unroll
      for (int a = 0; a < 4; ++a) {
        const int s = sgD * 4 + a;
        if (s < nvalid) {
          float* o = da2 + size_t(b0 + s) * A2_PER + igD * 4;
          *reinterpret_cast<float4*>(o) = make_float4(lo_f(acc[a][0]), hi_f(acc[a][0]), lo_f(acc[a][1]), hi_f(acc[a][1]));
          *reinterpret_cast<float4*>(o + 160) = make_float4(lo_f(acc[a][2]), hi_f(acc[a][2]), lo_f(acc[a][3]), hi_f(acc[a][3]));
        }
      }
    }
  }
  // one partial per CTA
  float* part = partials + size_t(blockIdx.x) * P_ARENA;
#pragma unroll
  for (int a = 0; a < 5; ++a)
#pragma unroll
    for (int c = 0; c < 5; ++c)         // OFF_F1W and igW*10 are even: 8-byte aligned pairs
      *reinterpret_cast<unsigned long long*>(part + OFF_F1W + (jgW * 5 + a) * A2_PER + igW * 10 + 2 * c) = accW1[a][c];
#pragma unroll
  for (int q = 0; q < 2; ++q) {
    const int e = tid + q * HB_THREADS;
    if (e < NC * F1) part[OFF_F2W + e] = accW2[q];
  }
  if (tid < F1) part[OFF_F1B + tid] = accB;
  else if (tid >= 64 && tid < 64 + NC) part[OFF_F2B + tid - 64] = accB;
}

// ---------------------------------------------------------------------------
// K_B2d conv2 data gradient.  One warp per sample; lane = g*16 + tile, tile =
// 3x3 block of the 12x12 input plane, g = half of the 10 input channels.
//   da1[ci][y][x] = sum_co sum_kh,kw dy2[co][y-kh][x-kw] * W[co][ci][kh][kw]
// dy2 (8x8 per co, one non-zero per 2x2 pool window) is rebuilt in smem from
// da2/a2/idx2; out-of-range taps are predicated to zero.
// Output dp1 = da1 * (a1 > 0): gradient w.r.t. the pooled conv1 map before ReLU.
// ---------------------------------------------------------------------------
constexpr int D2_WARPS = 8;
constexpr int D2_THREADS = D2_WARPS * 32;
constexpr int D2_WS = C2 * 25 * 2 * 8;              // [co][tap][g][8]
constexpr int D2_PER_WARP = A1_PER;                // dy2 [20][8][8] (1280 floats), later reused for the staged output [10][144]
constexpr int D2_SMEM = (D2_WS + D2_WARPS * D2_PER_WARP) * 4;
// S warps share one sample (S = 1, 2, 4): warp `part` of the S takes the live output channels co = part, part+S, ...
// and the S partial planes are added in part order.  S > 1 is the small-batch mode: with one warp per sample the
// kernel is a single ~40 us dependency chain however few samples there are; at B <= 1024 the SMs are mostly empty
// and splitting the chain four ways is nearly free.  S = 1 is the throughput mode (no redundant dy2 rebuild).
template <int S>
__global__ void __launch_bounds__(D2_THREADS, 2)
conv2_dgrad_kernel(const float* __restrict__ params, const float* __restrict__ da2,
                   const float* __restrict__ a2, const uint8_t* __restrict__ idx2,
                   const float* __restrict__ a1, float drop_scale, float* __restrict__ dp1, int B) {
  extern __shared__ __align__(128) float smem[];
  float* w_s = smem;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  constexpr int SPC = D2_WARPS / S;                 // samples per CTA iteration
  const int slot = warp / S, part = warp % S;
  float* dy_s = smem + D2_WS + warp * D2_PER_WARP;  // [20][8][8], private to the warp
  float* out_s = dy_s;                              // [10][144]: written only after the last read of dy2
  pdl_trigger();
  for (int i = tid; i < D2_WS; i += D2_THREADS) {
    int cc = i & 7, g = (i >> 3) & 1, ct = i >> 4;  // ct = co*25 + tap
    int co = ct / 25, tap = ct % 25;
    w_s[i] = cc < 5 ? params[OFF_C2W + co * 250 + (g * 5 + cc) * 25 + tap] : 0.f;
  }
  __syncthreads();
  pdl_wait();                                       // da2 (head_bwd) is read from here on
  const int g = lane >> 4, tile = lane & 15, ty = tile >> 2, tx = tile & 3;
  // validity of patch rows/cols: dy2 row = 3*ty + r - 4 must be in [0,8)
  bool rok[7], cok[7];
#pragma unroll
  for (int r = 0; r < 7; ++r) {
    int rr = 3 * ty + r - 4, cc = 3 * tx + r - 4;
    rok[r] = rr >= 0 && rr < O2;
    cok[r] = cc >= 0 && cc < O2;
  }
  const int pbase = (3 * ty - 4) * O2 + (3 * tx - 4);
  const int total_slots = gridDim.x * SPC;
  for (int b = blockIdx.x * SPC + slot; b < B; b += total_slots) {
    // rebuild dy2: 320 (co,pp) entries, 10 per lane, each a 2x2 block with one non-zero.
    // live_mask bit co: the channel has any gradient for this sample (Dropout2d zeroes ~half of the
    // channels, ReLU some more) -- dead channels are skipped, a warp-uniform branch.
    unsigned live_mask = 0u;
#pragma unroll
    for (int q = 0; q < 10; ++q) {
      const int e = lane + 32 * q;               // = co*16 + pp
      const size_t o = size_t(b) * A2_PER + e;
      const float gv = a2[o] > 0.f ? da2[o] * drop_scale : 0.f;
      const int id = idx2[o];
      const unsigned bal = __ballot_sync(0xffffffffu, gv != 0.f);
      live_mask |= ((bal & 0xffffu) ? 1u : 0u) << (2 * q);
      live_mask |= ((bal >> 16) ? 1u : 0u) << (2 * q + 1);
      const int co = e >> 4, pp = e & 15, py = pp >> 2, px = pp & 3;
      float* d = dy_s + co * 64 + (2 * py) * O2 + 2 * px;
      *reinterpret_cast<float2*>(d) = make_float2(id == 0 ? gv : 0.f, id == 1 ? gv : 0.f);
      *reinterpret_cast<float2*>(d + O2) = make_float2(id == 2 ? gv : 0.f, id == 3 ? gv : 0.f);
    }
    __syncwarp();
    float acc[3][3][5];
#pragma unroll
    for (int a = 0; a < 3; ++a)
#pragma unroll
      for (int c = 0; c < 3; ++c)
#pragma unroll
        for (int k = 0; k < 5; ++k) acc[a][c][k] = 0.f;
#pragma unroll 1
    for (int co = part; co < C2; co += S) {
      if (!((live_mask >> co) & 1u)) continue;
      float patch[7][7];
      const float* dp = dy_s + co * 64 + pbase;
#pragma unroll
      for (int r = 0; r < 7; ++r)
#pragma unroll
        for (int c = 0; c < 7; ++c) patch[r][c] = (rok[r] && cok[c]) ? dp[r * O2 + c] : 0.f;
      const float* wb = w_s + (co * 25) * 16 + g * 8;
#pragma unroll
      for (int kh = 0; kh < 5; ++kh)
#pragma unroll
        for (int kw = 0; kw < 5; ++kw) {
          float4 w0 = *reinterpret_cast<const float4*>(wb + (kh * 5 + kw) * 16);
          float w4 = wb[(kh * 5 + kw) * 16 + 4];
          float w[5] = {w0.x, w0.y, w0.z, w0.w, w4};
#pragma unroll
          for (int oy = 0; oy < 3; ++oy)
#pragma unroll
            for (int ox = 0; ox < 3; ++ox) {
              float v = patch[oy - kh + 4][ox - kw + 4];
#pragma unroll
              for (int k = 0; k < 5; ++k) acc[oy][ox][k] = fmaf(v, w[k], acc[oy][ox][k]);
            }
        }
    }
    // stage through smem so the global store (and the a1 mask read) is coalesced
    __syncwarp();                                     // every lane is done reading dy2 before it is overwritten
#pragma unroll
    for (int k = 0; k < 5; ++k)
#pragma unroll
      for (int oy = 0; oy < 3; ++oy)
#pragma unroll
        for (int ox = 0; ox < 3; ++ox)
          out_s[(g * 5 + k) * 144 + (3 * ty + oy) * P1 + 3 * tx + ox] = acc[oy][ox][k];
    if (S == 1) __syncwarp();
    else asm volatile("bar.sync %0, %1;" :: "r"(1 + slot), "n"(S * 32) : "memory");   // the S partial planes are complete
    const float4* a4 = reinterpret_cast<const float4*>(a1 + size_t(b) * A1_PER);
    float4* o4 = reinterpret_cast<float4*>(dp1 + size_t(b) * A1_PER);
    const float* first = smem + D2_WS + (slot * S) * D2_PER_WARP;                     // plane of part 0
    for (int i = part * 32 + lane; i < A1_PER / 4; i += S * 32) {
      float4 v = reinterpret_cast<const float4*>(first)[i];
#pragma unroll
      for (int p = 1; p < S; ++p) {                   // fixed order: part 0 + part 1 + ...
        const float4 u = reinterpret_cast<const float4*>(first + p * D2_PER_WARP)[i];
        v.x += u.x; v.y += u.y; v.z += u.z; v.w += u.w;
      }
      float4 a = a4[i];
      v.x = a.x > 0.f ? v.x : 0.f; v.y = a.y > 0.f ? v.y : 0.f;
      v.z = a.z > 0.f ? v.z : 0.f; v.w = a.w > 0.f ? v.w : 0.f;
      o4[i] = v;
    }
    if (S == 1) __syncwarp();
    else asm volatile("bar.sync %0, %1;" :: "r"(1 + slot), "n"(S * 32) : "memory");   // planes may be overwritten again
  }
}

// ---------------------------------------------------------------------------
// K_B2w conv2 weight/bias gradient.  thread <-> (ci, kh, cq): accumulators for
// co in {cq, cq+5, cq+10, cq+15} x kw 0..4; slides over the 8 output columns.
//   dW[co][ci][kh][kw] = sum_b sum_oy,ox dy2[b][co][oy][ox] * a1[b][ci][oy+kh][ox+kw]
// ---------------------------------------------------------------------------
constexpr int W2_THREADS = 256;
constexpr int W2_NS = 2;                              // samples per stage
constexpr int W2_DYS = O2 * C2 * O2;                  // [oy][co][ox] = 1280
constexpr int W2_CS = 156;                            // a1 channel stride in smem (144 used): bank = 12*(ci*5+kh) mod 32
constexpr int W2_A1S = C1 * W2_CS;                    // per-sample a1 footprint in smem
// shared-memory layout (dynamic): a1 double-buffered in its padded layout (bulk-copied per channel), raw a2 / da2 /
// idx2 double-buffered, dy2 + bias staging + live flags single
constexpr int W2_RAW = W2_NS * A2_PER;                          // floats of a2 (or da2) per stage
constexpr int W2_SMEM = (2 * W2_NS * W2_A1S + 2 * 2 * W2_RAW + W2_NS * W2_DYS + A2_PER) * 4 + 2 * W2_RAW + W2_NS * C2 * 4 + 32;
__global__ void __launch_bounds__(W2_THREADS, 3)
conv2_wgrad_kernel(const float* __restrict__ da2, const float* __restrict__ a2,
                   const uint8_t* __restrict__ idx2, const float* __restrict__ a1,
                   float drop_scale, float* __restrict__ partials, int B) {
  extern __shared__ __align__(128) float smem[];
  float* a1_s2 = smem;                                   // [2][W2_NS][10][156]
  float* ra2 = a1_s2 + 2 * W2_NS * W2_A1S;               // [2][W2_NS*320] raw a2
  float* rda2 = ra2 + 2 * W2_RAW;                        // [2][W2_NS*320] raw da2
  float* dy_s = rda2 + 2 * W2_RAW;                       // [W2_NS][8][20][8]
  float* bsum_s = dy_s + W2_NS * W2_DYS;                 // [320]
  int* live_s = reinterpret_cast<int*>(bsum_s + A2_PER); // [W2_NS*20]
  uint64_t* bar = reinterpret_cast<uint64_t*>(live_s + W2_NS * C2);   // [2]
  uint8_t* ridx = reinterpret_cast<uint8_t*>(bar + 2);   // [2][W2_NS*320] raw idx2
  const int tid = threadIdx.x;
  pdl_trigger();
  pdl_wait();                                            // no parameter prologue here: only the launch latency is hidden
  // cq (output-channel residue mod 5) is the SLOWEST index so that a warp is (almost) uniform in co
  // and the dead-channel skip below is (almost) divergence-free
  const int cq = tid / 50, kh = tid % 5, ci = (tid % 50) / 5;   // cq < 5 for tid < 250
  const bool active = tid < 250;
  const int nstages = (B + W2_NS - 1) / W2_NS;
  auto issue = [&](int st, int buf) {                    // thread 0: prefetch one stage's inputs (cp.async.bulk)
    const int b0 = st * W2_NS;
    const int nv = min(W2_NS, B - b0);
    mbarrier_expect_tx(&bar[buf], uint32_t(nv) * (A1_PER * 4 + 2 * A2_PER * 4 + A2_PER));
    for (int s = 0; s < nv; ++s)
      for (int c = 0; c < C1; ++c)                       // channel by channel into the conflict-free padded layout
        bulk_copy_g2s(a1_s2 + (buf * W2_NS + s) * W2_A1S + c * W2_CS, a1 + size_t(b0 + s) * A1_PER + c * 144, 144 * 4, &bar[buf]);
    bulk_copy_g2s(ra2 + buf * W2_RAW, a2 + size_t(b0) * A2_PER, uint32_t(nv) * A2_PER * 4, &bar[buf]);
    bulk_copy_g2s(rda2 + buf * W2_RAW, da2 + size_t(b0) * A2_PER, uint32_t(nv) * A2_PER * 4, &bar[buf]);
    bulk_copy_g2s(ridx + buf * W2_RAW, idx2 + size_t(b0) * A2_PER, uint32_t(nv) * A2_PER, &bar[buf]);
  };
  if (tid == 0) {
    mbarrier_init(&bar[0], 1); mbarrier_init(&bar[1], 1);
    if (int(blockIdx.x) < nstages) issue(blockIdx.x, 0);
    if (int(blockIdx.x + gridDim.x) < nstages) issue(blockIdx.x + gridDim.x, 1);
  }
  __syncthreads();
  float acc[4][5];
#pragma unroll
  for (int j = 0; j < 4; ++j)
#pragma unroll
    for (int k = 0; k < 5; ++k) acc[j][k] = 0.f;
  float bacc[2] = {0.f, 0.f};  // bias-gradient share of entries tid and tid+256
  uint32_t it = 0;
  for (int st = blockIdx.x; st < nstages; st += gridDim.x, ++it) {
    const int buf = it & 1;
    const int b0 = st * W2_NS;
    const int nvalid = min(W2_NS, B - b0);
    mbarrier_wait(&bar[buf], (it >> 1) & 1);
    const float* a1_s = a1_s2 + buf * W2_NS * W2_A1S;
    // dy2 (one non-zero per 2x2 pool window) from the raw tiles, all in shared memory
    for (int s = 0; s < nvalid; ++s) {
#pragma unroll
      for (int q = 0; q < 2; ++q) {
        const int e = tid + q * W2_THREADS;
        if (e < A2_PER) {
          const int o = buf * W2_RAW + s * A2_PER + e;
          const float gv = ra2[o] > 0.f ? rda2[o] * drop_scale : 0.f;
          const int id = ridx[o];
          const int co = e >> 4, pp = e & 15, py = pp >> 2, px = pp & 3;
          bacc[q] += gv;
          const unsigned bal = __ballot_sync(0xffffffffu, gv != 0.f);   // e < 320 is warp-uniform here
          if ((tid & 15) == 0) live_s[s * C2 + co] = ((bal >> (tid & 16)) & 0xffffu) != 0u;
          float* d = dy_s + s * W2_DYS + (2 * py) * (C2 * O2) + co * O2 + 2 * px;
          *reinterpret_cast<float2*>(d) = make_float2(id == 0 ? gv : 0.f, id == 1 ? gv : 0.f);
          *reinterpret_cast<float2*>(d + C2 * O2) = make_float2(id == 2 ? gv : 0.f, id == 3 ? gv : 0.f);
        }
      }
    }
    __syncthreads();
    if (active) {
      for (int s = 0; s < nvalid; ++s) {
        const float* ab = a1_s + s * W2_A1S + ci * W2_CS + kh * P1;
        const float* db = dy_s + s * W2_DYS + cq * O2;
        bool lv[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) lv[j] = live_s[s * C2 + cq + 5 * j] != 0;
        if (!(lv[0] || lv[1] || lv[2] || lv[3])) continue;
#pragma unroll 2
        for (int oy = 0; oy < O2; ++oy) {
          float a[12];
#pragma unroll
          for (int q = 0; q < 3; ++q) {
            float4 v = *reinterpret_cast<const float4*>(ab + oy * P1 + 4 * q);
            a[4 * q] = v.x; a[4 * q + 1] = v.y; a[4 * q + 2] = v.z; a[4 * q + 3] = v.w;
          }
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            if (!lv[j]) continue;
            const float* dr = db + oy * (C2 * O2) + j * 5 * O2;
            float4 d0 = *reinterpret_cast<const float4*>(dr);
            float4 d1 = *reinterpret_cast<const float4*>(dr + 4);
            float d[8] = {d0.x, d0.y, d0.z, d0.w, d1.x, d1.y, d1.z, d1.w};
#pragma unroll
            for (int ox = 0; ox < 8; ++ox)
#pragma unroll
              for (int k = 0; k < 5; ++k) acc[j][k] = fmaf(d[ox], a[ox + k], acc[j][k]);
          }
        }
      }
    }
    __syncthreads();                                     // dy_s and the buffers of `buf` are free
    if (tid == 0 && st + 2 * int(gridDim.x) < nstages) issue(st + 2 * gridDim.x, buf);
  }
  float* part = partials + size_t(blockIdx.x) * P_ARENA;
  if (active) {
#pragma unroll
    for (int j = 0; j < 4; ++j)
#pragma unroll
      for (int k = 0; k < 5; ++k)
        part[OFF_C2W + (cq + 5 * j) * 250 + ci * 25 + kh * 5 + k] = acc[j][k];
  }
  __syncthreads();
  bsum_s[tid] = bacc[0];
  if (tid + W2_THREADS < A2_PER) bsum_s[tid + W2_THREADS] = bacc[1];
  __syncthreads();
  if (tid < C2) {
    float s = 0.f;
#pragma unroll
    for (int p = 0; p < 16; ++p) s += bsum_s[tid * 16 + p];
    part[OFF_C2B + tid] = s;
  }
}

// ---------------------------------------------------------------------------
// K_B1 conv1 weight/bias gradient.  Two samples per stage; dy1 (24x24 per channel, one non-zero per pool
// window) is rebuilt dense in smem from dp1/idx1.  thread <-> (sample slot, output row oy, kh): it holds the
// input row oy+kh in registers and sweeps all 10 channels over it -- 50 accumulators (c, kw), 18 FFMA per LDS.128.
//   dW[c][kh][kw] = sum_b sum_oy,ox dy1[b][c][oy][ox] * x[b][oy+kh][ox+kw]
// Raw inputs (x, dp1, idx1) of the next stage are bulk-prefetched (cp.async.bulk + mbarrier) while this one computes.
// ---------------------------------------------------------------------------
constexpr int W1_THREADS = 256;
constexpr int W1_NS = 2;                                    // samples per stage
constexpr int W1_ACTIVE = W1_NS * 120;                      // (slot, oy 24, kh 5)
constexpr int W1_DYS = C1 * 576;                            // dense dy1 of one sample
constexpr int W1_RAWF = W1_NS * (IMG_PIX + A1_PER);         // floats per raw buffer: x + dp1
constexpr int W1_SMEM = (W1_NS * W1_DYS + 2 * W1_RAWF) * 4 + 2 * W1_NS * A1_PER + 64;
__global__ void __launch_bounds__(W1_THREADS, 2)
conv1_wgrad_kernel(const float* __restrict__ dp1, const uint8_t* __restrict__ idx1,
                   const float* __restrict__ x, float* __restrict__ partials, int B) {
  extern __shared__ __align__(128) float smem[];
  float* dy_s = smem;                                        // [W1_NS][10][24][24]; reused for the final reductions
  float* rawf = dy_s + W1_NS * W1_DYS;                       // [2 buffers][ x: NS*784 | dp1: NS*1440 ]
  uint64_t* bar = reinterpret_cast<uint64_t*>(rawf + 2 * W1_RAWF);   // [2]
  uint8_t* rawi = reinterpret_cast<uint8_t*>(bar + 2);       // [2 buffers][NS*1440] idx1
  const int tid = threadIdx.x;
  pdl_trigger();
  pdl_wait();
  const bool active = tid < W1_ACTIVE;
  const int sl = tid / 120, r = tid % 120, kh = r % 5, oy = r / 5;
  const int nstages = (B + W1_NS - 1) / W1_NS;
  auto issue = [&](int st, int buf) {                        // thread 0: prefetch one stage's x, dp1, idx1
    const int b0 = st * W1_NS;
    const uint32_t nv = uint32_t(min(W1_NS, B - b0));
    mbarrier_expect_tx(&bar[buf], nv * (IMG_PIX * 4 + A1_PER * 4 + A1_PER));
    bulk_copy_g2s(rawf + buf * W1_RAWF, x + size_t(b0) * IMG_PIX, nv * IMG_PIX * 4, &bar[buf]);
    bulk_copy_g2s(rawf + buf * W1_RAWF + W1_NS * IMG_PIX, dp1 + size_t(b0) * A1_PER, nv * A1_PER * 4, &bar[buf]);
    bulk_copy_g2s(rawi + buf * W1_NS * A1_PER, idx1 + size_t(b0) * A1_PER, nv * A1_PER, &bar[buf]);
  };
  if (tid == 0) {
    mbarrier_init(&bar[0], 1); mbarrier_init(&bar[1], 1);
    if (int(blockIdx.x) < nstages) issue(blockIdx.x, 0);
    if (int(blockIdx.x + gridDim.x) < nstages) issue(blockIdx.x + gridDim.x, 1);
  }
  __syncthreads();
  float acc[C1][5];
#pragma unroll
  for (int c = 0; c < C1; ++c)
#pragma unroll
    for (int k = 0; k < 5; ++k) acc[c][k] = 0.f;
  float bacc[2] = {0.f, 0.f};                                // bias share: channels 2*kh, 2*kh+1 of this thread's row
  int dst_off[6];                                            // where entry tid + 256*q (= cc*144 + p) lands in the dense dy1
#pragma unroll
  for (int q = 0; q < 6; ++q) {
    const int e = tid + q * W1_THREADS, cc = e / 144, p = e % 144;
    dst_off[q] = cc * 576 + (2 * (p / 12)) * 24 + 2 * (p % 12);
  }
  uint32_t it = 0;
  for (int st = blockIdx.x; st < nstages; st += gridDim.x, ++it) {
    const int buf = it & 1;
    const int nvalid = min(W1_NS, B - st * W1_NS);
    mbarrier_wait(&bar[buf], (it >> 1) & 1);
    const float* x_s = rawf + buf * W1_RAWF;
    const float* dp_s = x_s + W1_NS * IMG_PIX;
    const uint8_t* ix_s = rawi + buf * W1_NS * A1_PER;
#pragma unroll
    for (int q = 0; q < 6; ++q) {                                      // entry e = tid + 256*q = cc*144 + p of each sample
      const int e = tid + q * W1_THREADS;
      if (e < A1_PER) {
        for (int s2 = 0; s2 < nvalid; ++s2) {
          const float gv = dp_s[s2 * A1_PER + e];
          const int id = ix_s[s2 * A1_PER + e];
          float* d = dy_s + s2 * W1_DYS + dst_off[q];
          *reinterpret_cast<float2*>(d) = make_float2(id == 0 ? gv : 0.f, id == 1 ? gv : 0.f);
          *reinterpret_cast<float2*>(d + 24) = make_float2(id == 2 ? gv : 0.f, id == 3 ? gv : 0.f);
        }
      }
    }
    __syncthreads();
    if (active && sl < nvalid) {
      float xr[28];
      const float* xp = x_s + sl * IMG_PIX + (oy + kh) * IMG;
#pragma unroll
      for (int q = 0; q < 7; ++q) {
        float4 v = *reinterpret_cast<const float4*>(xp + 4 * q);
        xr[4 * q] = v.x; xr[4 * q + 1] = v.y; xr[4 * q + 2] = v.z; xr[4 * q + 3] = v.w;
      }
      const float* dbase = dy_s + sl * W1_DYS + oy * 24;
#pragma unroll
      for (int c = 0; c < C1; ++c) {
        float d[24];
#pragma unroll
        for (int q = 0; q < 6; ++q) {
          float4 v = *reinterpret_cast<const float4*>(dbase + c * 576 + 4 * q);
          d[4 * q] = v.x; d[4 * q + 1] = v.y; d[4 * q + 2] = v.z; d[4 * q + 3] = v.w;
        }
#pragma unroll
        for (int ox = 0; ox < 24; ++ox)
#pragma unroll
          for (int k = 0; k < 5; ++k) acc[c][k] = fmaf(d[ox], xr[ox + k], acc[c][k]);
        if ((c >> 1) == kh) {                                 // bias: every thread sums 2 of the 10 channels of its row
          float sum = 0.f;
#pragma unroll
          for (int ox = 0; ox < 24; ++ox) sum += d[ox];
          bacc[c & 1] += sum;
        }
      }
    }
    __syncthreads();                                         // dy_s and the raw buffers of `buf` are free
    if (tid == 0 && st + 2 * int(gridDim.x) < nstages) issue(st + 2 * gridDim.x, buf);
  }
  // fixed-order reduction over the 48 (slot, oy) threads of every (c, kh, kw): two passes of 5 channels
  float* part = partials + size_t(blockIdx.x) * P_ARENA;
  float* red = dy_s;                                         // [240][25]
#pragma unroll
  for (int half = 0; half < 2; ++half) {
    __syncthreads();
    if (active) {
#pragma unroll
      for (int c = 0; c < 5; ++c)
#pragma unroll
        for (int k = 0; k < 5; ++k) red[tid * 25 + c * 5 + k] = acc[half * 5 + c][k];
    }
    __syncthreads();
    if (tid < 125) {                                         // (c_local, kh, kw)
      const int cl = tid / 25, kh2 = (tid % 25) / 5, kw = tid % 5;
      float sum = 0.f;
      for (int t = 0; t < W1_NS * 24; ++t) sum += red[((t / 24) * 120 + (t % 24) * 5 + kh2) * 25 + cl * 5 + kw];
      part[OFF_C1W + (half * 5 + cl) * 25 + kh2 * 5 + kw] = sum;
    }
  }
  __syncthreads();
  if (active) { red[tid * 2] = bacc[0]; red[tid * 2 + 1] = bacc[1]; }
  __syncthreads();
  if (tid < C1) {                                            // channel c was summed by the threads with kh == c/2
    float sum = 0.f;
    for (int t = 0; t < W1_NS * 24; ++t) sum += red[((t / 24) * 120 + (t % 24) * 5 + (tid >> 1)) * 2 + (tid & 1)];
    part[OFF_C1B + tid] = sum;
  }
}

// ---------------------------------------------------------------------------
// K_R reduce per-CTA partials into the gradient bucket, fixed order.
// 1024 threads = 32 warps x 32 consecutive elements; warp w sums parts w, w+32, ... (all its loads in flight at
// once), then warp 0 adds the 32 warp sums in order.
// ---------------------------------------------------------------------------
__device__ __forceinline__ bool arena_is_pad(int e) {
  return (e >= 250 && e < OFF_C1B) || (e >= OFF_C1B + 10 && e < OFF_C2W) ||
         (e >= OFF_F1B + 50 && e < OFF_F2W) || (e >= OFF_F2B + 10);
}
constexpr int RP_WARPS = 32;
// 1024 threads = 32 warps x 32 lanes; lane <-> one float4 (4 consecutive arena elements), warp w sums the partial rows
// w, w+32, ... (all its 128-bit loads in flight at once), then warp 0 adds the 32 warp sums in order.
__global__ void __launch_bounds__(RP_WARPS * 32)
reduce_partials_kernel(const float* __restrict__ partials, int parts_c1, int parts_c2,
                       int parts_head, float* __restrict__ bucket) {
  __shared__ float4 red[RP_WARPS][33];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  pdl_trigger();
  pdl_wait();
  const int e = (blockIdx.x * 32 + lane) * 4;          // every tensor starts at a multiple of 4 floats: no float4 straddles two
  float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
  if (e < P_ARENA) {
    const int parts = e < OFF_C2W ? parts_c1 : (e < OFF_F1W ? parts_c2 : parts_head);
    const float* p = partials + e;
    constexpr int Q = (MAX_PARTS + RP_WARPS - 1) / RP_WARPS;
    float4 v[Q];
#pragma unroll
    for (int q = 0; q < Q; ++q) {
      const int i = warp + q * RP_WARPS;
      v[q] = i < parts ? __ldcg(reinterpret_cast<const float4*>(p + size_t(i) * P_ARENA)) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
#pragma unroll
    for (int q = 0; q < Q; ++q) { s.x += v[q].x; s.y += v[q].y; s.z += v[q].z; s.w += v[q].w; }
  }
  red[warp][lane] = s;
  __syncthreads();
  if (warp == 0 && e < P_ARENA) {
    float4 t = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
    for (int w = 0; w < RP_WARPS; ++w) { const float4 r = red[w][lane]; t.x += r.x; t.y += r.y; t.z += r.z; t.w += r.w; }
    // padding slots of the arena are never written by the producers: keep them zero
    if (arena_is_pad(e)) t.x = 0.f;
    if (arena_is_pad(e + 1)) t.y = 0.f;
    if (arena_is_pad(e + 2)) t.z = 0.f;
    if (arena_is_pad(e + 3)) t.w = 0.f;
    *reinterpret_cast<float4*>(bucket + e) = t;
  }
}

static int set_bwd_attrs() {   // function attributes are per device: a process may drive several GPUs (ddp_set_device per host thread)
  static bool attr_done[64] = {false};
  int dev = 0;
  DDP_CUDA(cudaGetDevice(&dev));
  if (dev >= 0 && dev < 64 && !attr_done[dev]) {
    DDP_CUDA(cudaFuncSetAttribute(head_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, HB_SMEM));
    DDP_CUDA(cudaFuncSetAttribute(conv2_dgrad_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, D2_SMEM));
    DDP_CUDA(cudaFuncSetAttribute(conv2_dgrad_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, D2_SMEM));
    DDP_CUDA(cudaFuncSetAttribute(conv2_dgrad_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, D2_SMEM));
    DDP_CUDA(cudaFuncSetAttribute(conv2_wgrad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, W2_SMEM));
    DDP_CUDA(cudaFuncSetAttribute(conv1_wgrad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, W1_SMEM));
    attr_done[dev] = true;
  }
  return 0;
}

// Everything behind the head: conv2 dgrad / wgrad, conv1 wgrad, fixed-order reduction of the partial rows into the
// bucket.  `parts_head` = number of partial rows the head kernel (K_B3 or the fused training head) wrote.
int launch_bwd_after_head(const float* params, const float* x, int training, int B, const Workspace& ws, const Scratch& sc,
                          int parts_head, float* grad_bucket, cudaStream_t stream, const PeerExchange* px) {
  if (int rc = set_bwd_attrs()) return rc;
  const int sms = sm_count();
  const float drop_scale = training ? 2.f : 1.f;
  int parts_c2, parts_c1;
  prof_mark(5, stream);
  // conv2_wgrad depends on head_bwd only, not on dgrad -> conv1_wgrad.  At large batches every kernel fills the SMs
  // by itself (forking was tried there: the CTAs do not co-reside, and shrinking the grids to make them fit lost 30 %),
  // but up to B = 1024 the machine is mostly empty and the step is a latency chain: there conv2_wgrad runs on the
  // library's side stream beside dgrad and joins before the reduction (captured as a fork/join in a CUDA graph).
  static int fork_max_b = -1;
  if (fork_max_b < 0) { const char* e = getenv("DDP_FORK_MAX_B"); fork_max_b = e ? atoi(e) : 1024; }
  bool fork = B <= fork_max_b && !prof_enabled();
  // With a peer exchange attached, large batches push the fc / conv2 gradients to the peers WHILE conv1_wgrad runs
  // (side stream, early kernel), so only conv1's 260 gradients and the flag round trip remain exposed afterwards.
  // (A/B switch DDP_EARLY_PUSH=1; measured at 2 GPUs, B = 4096: 277 us/step with the early push against 258 us without --
  //  the side-stream kernel and the smaller conv1_wgrad grid cost more than the hidden exchange saves -- so the default
  //  is the single final kernel, which reduces the partial rows itself instead of a separate reduce_partials launch.)
  static int early_env = -1;
  if (early_env < 0) { const char* e = getenv("DDP_EARLY_PUSH"); early_env = (e && e[0] == '1') ? 1 : 0; }
  const bool early = px != nullptr && !fork && !prof_enabled() && early_env == 1;
  cudaStream_t wstream = stream;
  cudaEvent_t join_ev = nullptr;
  if (fork) {
    if (int rc = side_fork(stream, &wstream, &join_ev)) return rc;
    if (!wstream) { fork = false; wstream = stream; }   // first call is inside a stream capture: no side stream yet
  }
  {
    int nstages = (B + W2_NS - 1) / W2_NS;
    parts_c2 = nstages < MAX_PARTS ? nstages : MAX_PARTS;
    if (parts_c2 > sms * 3) parts_c2 = sms * 3;      // 3 CTAs/SM: one stages while the others compute
  }
  const bool dense = bwd_dense();
  static int w2_dense_env = -1, d2_dense_env = -1;
  if (w2_dense_env < 0) { const char* e = getenv("DDP_W2_DENSE"); w2_dense_env = (e && e[0] == '1') ? 1 : 0; }
  if (d2_dense_env < 0) { const char* e = getenv("DDP_D2_DENSE"); d2_dense_env = (e && e[0] == '1') ? 1 : 0; }
  const bool w2_dense = dense || w2_dense_env == 1, d2_dense = dense || d2_dense_env == 1;
  auto launch_wgrad2 = [&]() -> int {
    if (!w2_dense && wgrad2_tc())
      return launch_conv2_wgrad_tc(sc.da2, ws.a2, ws.idx2, ws.a1, drop_scale, sc.partials, B, &parts_c2, wstream, !fork);
    if (!w2_dense)
      return launch_conv2_wgrad_sp(sc.da2, ws.a2, ws.idx2, ws.a1, drop_scale, sc.partials, B, &parts_c2, wstream, !fork);
    DDP_CUDA(launch_pdl(conv2_wgrad_kernel, dim3(parts_c2), dim3(W2_THREADS), W2_SMEM, wstream, !fork, sc.da2, ws.a2, ws.idx2, ws.a1,
                        drop_scale, sc.partials, B));
    return 0;
  };
  if (fork) {
    if (int rc = launch_wgrad2()) return rc;
    DDP_CUDA(cudaEventRecord(join_ev, wstream));
  }
  if (!d2_dense) {
    if (int rc = launch_conv2_dgrad_sp(params, sc.da2, ws.a2, ws.idx2, ws.a1, drop_scale, sc.dp1, B, stream)) return rc;
  } else {
    // warps per sample: 4 while the SMs are mostly empty (latency mode), 1 once one warp per sample fills them
    const int split = B <= 1024 ? 4 : (B <= 2048 ? 2 : 1);
    const int spc = D2_WARPS / split;
    int nblk = (B + spc - 1) / spc;
    int grid = nblk < sms * 2 ? nblk : sms * 2;
    if (split == 4) DDP_CUDA(launch_pdl(conv2_dgrad_kernel<4>, dim3(grid), dim3(D2_THREADS), D2_SMEM, stream, true, params, sc.da2, ws.a2, ws.idx2, ws.a1, drop_scale, sc.dp1, B));
    else if (split == 2) DDP_CUDA(launch_pdl(conv2_dgrad_kernel<2>, dim3(grid), dim3(D2_THREADS), D2_SMEM, stream, true, params, sc.da2, ws.a2, ws.idx2, ws.a1, drop_scale, sc.dp1, B));
    else DDP_CUDA(launch_pdl(conv2_dgrad_kernel<1>, dim3(grid), dim3(D2_THREADS), D2_SMEM, stream, true, params, sc.da2, ws.a2, ws.idx2, ws.a1, drop_scale, sc.dp1, B));
  }
  prof_mark(6, stream);
  if (!fork)
    if (int rc = launch_wgrad2()) return rc;
  prof_mark(7, stream);
  cudaStream_t estream = nullptr;
  cudaEvent_t ejoin = nullptr;
  bool early_on = early;
  if (early_on) {                                      // fc + conv2 partial rows are final: reduce + push them beside conv1_wgrad
    if (int rc = side_fork(stream, &estream, &ejoin)) return rc;
    if (!estream) early_on = false;                    // first call is inside a stream capture: no side stream yet
  }
  if (early_on) {
    if (int rc = launch_peer_early(*px, sc.partials, parts_c2, parts_head, estream)) return rc;
    DDP_CUDA(cudaEventRecord(ejoin, estream));
  }
  {
    // default: the column-walker kernel on packed pairs (two samples in flight per CTA); DDP_W1_SCALAR=1 keeps the scalar-FFMA kernel for A/B
    static int w1_scalar = -1;
    if (w1_scalar < 0) { const char* e = getenv("DDP_W1_SCALAR"); w1_scalar = (e && e[0] == '1') ? 1 : 0; }
    const int nst1 = w1_scalar ? (B + W1_NS - 1) / W1_NS : (B + 1) / 2;   // the grid policy counts pairs of samples either way
    // with the early push in flight leave room for its 32 small CTAs: 32 SMs keep one conv1_wgrad CTA instead of two
    static int w1_per_sm = -1;
    if (w1_per_sm < 0) { const char* e = getenv("DDP_W1_CTAS_PER_SM"); w1_per_sm = (e && e[0] == '1') ? 1 : 2; }
    int cap1 = early_on ? sms * 2 - 32 : sms * w1_per_sm;
    // small-batch mode: conv2_wgrad is still running on the side stream when this kernel starts; with two of these CTAs
    // per SM (64k registers) nothing co-resides and the second wave waits -- one CTA per SM runs beside it
    // (measured at B = 512 / 768 / 1024: 116.7 / 118.8 / 135.2 -> 108.5 / 108.6 / 129.1 us per step; B = 384 prefers two)
    if (fork && nst1 > sms + sms / 2) cap1 = sms;
    parts_c1 = nst1 < cap1 ? nst1 : cap1;
    if (parts_c1 > MAX_PARTS) parts_c1 = MAX_PARTS;
    if (w1_scalar) {
      DDP_CUDA(launch_pdl(conv1_wgrad_kernel, dim3(parts_c1), dim3(W1_THREADS), W1_SMEM, stream, !early_on, sc.dp1, ws.idx1, x, sc.partials, B));
    } else {
      if (int rc = launch_conv1_wgrad_col(sc.dp1, ws.idx1, x, sc.partials, B, parts_c1, stream, !early_on)) return rc;
    }
  }
  prof_mark(8, stream);
  if (fork) DDP_CUDA(cudaStreamWaitEvent(stream, join_ev, 0));
  if (early_on) {
    DDP_CUDA(cudaStreamWaitEvent(stream, ejoin, 0));
    if (int rc = launch_peer_final(*px, sc.partials, parts_c1, /*have_early=*/true, parts_c2, parts_head, grad_bucket, stream)) return rc;
  } else if (px) {
    if (int rc = launch_peer_final(*px, sc.partials, parts_c1, /*have_early=*/false, parts_c2, parts_head, grad_bucket, stream)) return rc;
  } else {
    DDP_CUDA(launch_pdl(reduce_partials_kernel, dim3((P_ARENA / 4 + 31) / 32), dim3(RP_WARPS * 32), 0, stream, true, sc.partials, parts_c1,
                        parts_c2, parts_head, grad_bucket));
  }
  prof_mark(9, stream);
  DDP_CUDA(cudaGetLastError());
  return 0;
}
// K_B3 (shared by ddp_net_backward and the whole-step entry point)
int launch_head_bwd(const float* params, const Workspace& ws, const float* logp, const float* dlogp, const int64_t* target,
                    float nll_coef, float drop_scale, const Scratch& sc, int B, int* parts_head, cudaStream_t stream) {
  if (int rc = set_bwd_attrs()) return rc;
  const int sms = sm_count();
  prof_mark(4, stream);
  int ntiles = (B + HT - 1) / HT;
  *parts_head = ntiles < sms ? ntiles : sms;
  DDP_CUDA(launch_pdl(head_bwd_kernel, dim3(*parts_head), dim3(HB_THREADS), HB_SMEM, stream, true, params, ws.a2, ws.h, logp,
                      dlogp, target, nll_coef, drop_scale, sc.da2, sc.partials, B));
  return 0;
}
}  // namespace ddp

using namespace ddp;

extern "C" int ddp_net_backward(const float* params, const float* x, const float* logp,
                                const float* dlogp, const int64_t* target, float loss_scale,
                                int training, int64_t batch, const void* workspace, void* scratch,
                                float* grad_bucket, void* stream_) {
  DDP_REQUIRE(params && x && logp && workspace && scratch && grad_bucket, "null required pointer");
  DDP_REQUIRE(dlogp || target, "either dlogp or target must be given");
  DDP_REQUIRE(batch >= 1 && batch <= (int64_t(1) << 24), "batch out of range");
  DDP_REQUIRE(aligned16(params) && aligned16(x) && aligned16(workspace) && aligned16(scratch),
              "params/x/workspace/scratch must be 16-byte aligned");
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  const int B = int(batch);
  const int sms = sm_count();
  Workspace ws = carve_workspace(const_cast<void*>(workspace), batch);
  Scratch sc = carve_scratch(scratch, batch);
  const float drop_scale = training ? 2.f : 1.f;
  const float nll_coef = -loss_scale / float(B);
  if (int rc = set_bwd_attrs()) return rc;
  int parts_head;
  if (int rc = launch_head_bwd(params, ws, logp, dlogp, target, nll_coef, drop_scale, sc, B, &parts_head, stream)) return rc;
  return launch_bwd_after_head(params, x, training, B, ws, sc, parts_head, grad_bucket, stream);
}
