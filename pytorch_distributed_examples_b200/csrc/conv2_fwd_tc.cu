// K_F2 on the 5th-generation tensor cores: conv2(10->20,5x5) + bias + Dropout2d + maxpool2 + ReLU as a
// shifted-window implicit GEMM with tcgen05.mma (kind::tf32, fp32 accumulators in TMEM), 3xTF32 split so the
// result stays within the fp32 parity bar (ref: horovod/mnist_horovod.py:20;  SURVEY §7.3 H1).
//
// GEMM view, per group of 4 samples:
//   A'[r][k]   r = y*32 + s*8 + ox   (y<12, s<4, ox<8)        k = kw*10 + ci  (50, zero-padded to 56)
//              = a1[s][ci][y][ox+kw]          -- im2col along kw only (5x duplication, built on the fly)
//   B[kh][k][n] = W[n][ci][kh][kw]            n = output channel (20, zero-padded to 32)
//   D[m][n] = sum_kh  A'[m + 32*kh][:] . B[kh][:, n]         m = oy*32 + s*8 + ox,  oy < 8
// The kh taps are NOT materialised: shifting the conv window down by one row is a shift of 32 rows of A',
// i.e. +512 bytes on the start address of the (un-swizzled, K-major) shared-memory descriptor.  With the image row
// as the SLOWEST row index the 8 valid output rows of the 4 samples are rows 0..255 of D: two full 128-row tiles,
// no discarded rows (a sample-major order needs three tiles per group, a third of them garbage).
// 3xTF32 with two MMAs per (kh, k-step).  B rows are packed [B_hi (20) | B_lo (20) | 8 zero rows]:
//   A_hi x B[0..47]  (N=48): hi*hi in columns 0-19, hi*lo in columns 20-39
//   A_lo x B[0..31]  (N=32): + lo*hi into columns 0-19; columns 20-31 also receive lo*lo of channels 0-11 (the fourth,
//                    2^-22-sized term of the exact product: harmless, deterministic, and it lets N be 48+32 instead of 64+32)
// and the epilogue adds column c and column 20+c.  tf32 MMAs retire about one accumulator column per clock per 128x8
// tile here, so the MMA time is proportional to the padded N: 80 columns per (kh, k-step) instead of 96.
//
// Warp roles (320 threads, 1 CTA/SM, persistent over groups):
//   warps 0-3  epilogue: tcgen05.ld accumulators -> bias, dropout, 2x2 max-pool (rows oy, oy+1 live in neighbouring warps:
//              exchanged through shared memory), ReLU -> a2, idx2
//   warps 4-7  producers: the group's raw a1 (23 KB) arrives by ONE cp.async.bulk (TMA engine, mbarrier complete_tx),
//              prefetched a group ahead; producers read it from smem, split into tf32 hi / lo, 16-byte STS into the ring
//   warps 8-9  one elected lane each issues the tcgen05.mma / tcgen05.commit of one M-tile; warp 8 owns the TMEM allocation
// Pipelines: full/empty mbarriers over a 4-stage ring of k-steps (producer <-> MMA) and tmem_full/tmem_empty
// over two accumulator buffers (MMA <-> epilogue), so the epilogue of group g overlaps the MMAs of group g+1.
#include "common.cuh"
#include "tc_common.cuh"

namespace ddp {
namespace tc2 {

// Developer aid: -DDDP_TC_TIMING accumulates clock64() per warp role into g_tc_dbg (read with ddp_debug_tc_timing,
// tools/tc_timing.py); it is how the producer / MMA-issue / epilogue balance of this kernel was tuned.
#ifdef DDP_TC_TIMING
__device__ unsigned long long g_tc_dbg[160 * 16];
#define TC_T0(var) const long long var = clock64()
#define TC_ACC(slot, t0) do { g_tc_dbg[blockIdx.x * 16 + (slot)] += (unsigned long long)(clock64() - (t0)); } while (0)
#else
#define TC_T0(var)
#define TC_ACC(slot, t0)
#endif

constexpr int G = 4;                        // samples per group
constexpr int ROWS = 12 * G * 8;            // rows of A': (y, s, ox) = 384
constexpr int ROWS_PAD = ROWS;              // the deepest kh shift (4*32 rows) of the last D row (255) is row 383
constexpr int KH_ROWS = G * 8;              // one conv row down = 32 rows of A'
constexpr int MT = 2;                       // M-tiles per group: D rows (oy < 8, s, ox) = 256
constexpr int KSTEPS = 7;                   // 56 / 8
constexpr int NPAD = 32;                    // output channels padded to the MMA N
constexpr int STAGES = 4;
constexpr int A_PLANE = ROWS_PAD * 16;      // bytes of one k-chunk plane (4 k values x all rows)
constexpr int A_HL = 2 * A_PLANE;           // one k-step of hi (or lo): 2 chunks
constexpr int A_STAGE = 2 * A_HL;           // hi + lo
constexpr int B_ROWS = 48;                  // [B_hi rows 0-19 | B_lo rows 20-39 | zero rows 40-47] of one (kh, kstep)
constexpr int B_PLANE = B_ROWS * 16;        // one k-chunk plane: 768 B
constexpr int B_TILE = 2 * B_PLANE;         // one (kh, kstep): 1.5 KB
constexpr int SMEM_A = STAGES * A_STAGE;    // 106,496
constexpr int SMEM_B = 5 * KSTEPS * B_TILE; // 53,760
constexpr int RAW_S = A1_PER + 8;             // per-sample stride of the raw tile (floats): == 8 mod 32, so the 32 lanes
                                             // (s, ox) of a producer warp read 32 distinct banks
constexpr int RAW_BYTES = G * RAW_S * 4;     // one group's a1: 23,168 B
constexpr int SMEM_RAW = 2 * RAW_BYTES;     // double-buffered
constexpr int SMEM_MISC = 1024;             // barriers, tmem pointer, bias, noise
constexpr int SMEM_XCH = 4 * C2 * 32 * 4;   // pooling exchange: [epilogue warp][channel][lane]
constexpr int SMEM_BYTES = SMEM_A + SMEM_B + SMEM_RAW + SMEM_MISC + SMEM_XCH;
constexpr int THREADS = 320;                // 4 epilogue + 4 producer + 2 MMA-issuer warps
constexpr int DCOLS = 2 * NPAD;              // accumulator column stride per tile (48 used: [hi*hi + lo*hi (20) | hi*lo (20) | pad])
constexpr int NHL = B_ROWS;                  // N of the A_hi MMA
constexpr int TMEM_COLS = 256;              // 2 buffers x 2 tiles x 64 columns

using namespace ::ddp::tc;

__global__ void __launch_bounds__(THREADS, 1)
conv2_fwd_tc_kernel(const float* __restrict__ params, const float* __restrict__ a1,
                    const float* __restrict__ mask2d, uint64_t seed, uint64_t offset,
                    const int64_t* __restrict__ offset_dev, int training,
                    float* __restrict__ a2, uint8_t* __restrict__ idx2, int B) {
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  unsigned char* a_ring = smem_raw;
  unsigned char* b_mat = smem_raw + SMEM_A;
  float* raw_s = reinterpret_cast<float*>(b_mat + SMEM_B);          // [2][G][RAW_S]
  unsigned char* misc = b_mat + SMEM_B + SMEM_RAW;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(misc);            // [STAGES]
  uint64_t* empty_bar = full_bar + STAGES;                           // [STAGES]
  uint64_t* tfull_bar = empty_bar + STAGES;                          // [2]
  uint64_t* tempty_bar = tfull_bar + 2;                              // [2]
  uint64_t* rfull_bar = tempty_bar + 2;                              // [2] raw a1 tile landed (tx bytes)
  uint64_t* rempty_bar = rfull_bar + 2;                              // [2] producers done reading it
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(rempty_bar + 2);
  float* bias_s = reinterpret_cast<float*>(misc + 192);              // [32]  (16 barriers + the TMEM pointer sit below)
  float* noise_s = bias_s + 32;                                      // [2 buffers][G][20]
  float* xch_s = reinterpret_cast<float*>(misc + SMEM_MISC);         // [4 warps][20][32]

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  pdl_trigger();
  if (offset_dev) offset += uint64_t(*offset_dev);   // the step counter is bumped by the update kernel, which precedes conv1_fwd
  TC_T0(t_kernel);

  // ---- one-time setup: weights -> B (hi, lo), barriers, TMEM ---------------------------------------
  // zero B (pad rows, pad k) and the whole stage ring once: pad rows / pad k stay finite forever
  for (int i = tid; i < SMEM_B / 16; i += THREADS) reinterpret_cast<float4*>(b_mat)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
  for (int i = tid; i < SMEM_A / 16; i += THREADS) reinterpret_cast<float4*>(a_ring)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
  if (tid < 32) bias_s[tid] = tid < C2 ? params[OFF_C2B + tid] : 0.f;
  __syncthreads();
  // ONE coalesced pass over the 5,000 weights (this prologue is exposed: conv1_fwd's CTAs fill the register file, so the
  // CTA cannot start before they drain): W[n][ci][kh][kw] -> k = kw*10 + ci, hi into row n, lo into row 20 + n
#pragma unroll 4
  for (int i = tid; i < C2 * 250; i += THREADS) {
    const float w = params[OFF_C2W + i];
    const int n = i / 250, r = i % 250, ci = r / 25, kh = (r % 25) / 5, kw = r % 5;
    const int k = kw * 10 + ci;
    float hi, lo;
    split_tf32(w, hi, lo);
    const int off = (kh * KSTEPS + (k >> 3)) * B_TILE + ((k >> 2) & 1) * B_PLANE + n * 16 + (k & 3) * 4;
    *reinterpret_cast<float*>(b_mat + off) = hi;
    *reinterpret_cast<float*>(b_mat + off + C2 * 16) = lo;
  }
  if (tid == 0) {
    for (int s = 0; s < STAGES; ++s) { mbar_init(&full_bar[s], 128); mbar_init(&empty_bar[s], MT); }
    for (int b = 0; b < 2; ++b) {
      mbar_init(&tfull_bar[b], MT); mbar_init(&tempty_bar[b], 128);
      mbar_init(&rfull_bar[b], 1); mbar_init(&rempty_bar[b], 128);
    }
    asm volatile("fence.mbarrier_init.release.cluster;");
  }
  if (warp == 8) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" :: "r"(smem_u32(tmem_ptr)), "n"(TMEM_COLS));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;");
  pdl_wait();                                  // a1 (conv1_fwd) is read from here on; everything above touched params + smem only
  const uint32_t tmem = *tmem_ptr;
  const int ngroups = (B + G - 1) / G;
  if (tid == 0) TC_ACC(0, t_kernel);   // slot 0: setup

  if (warp >= 4 && warp < 8) {
    // ======================= producers =======================
    const int pt = tid - 128;                       // 0..127
    // each thread owns rows pt, pt+128, pt+256 of A' (both k-chunks): decompose them once
    uint32_t it = 0, gi = 0;                        // running k-step / group counters -> ring slots / phases
    int row_src[3], row_s[3];
#pragma unroll
    for (int q = 0; q < 3; ++q) {
      const int r = pt + 128 * q;
      const int y = r / KH_ROWS, sidx = (r % KH_ROWS) >> 3, ox = r & 7;
      row_s[q] = sidx;
      row_src[q] = sidx * RAW_S + y * P1 + ox;                      // + ci*144 + kw per k
    }
    auto issue_raw = [&](int grp, int rb) {         // one elected thread: bulk-copy the group's a1 into raw_s[rb]
      const int b0 = grp * G;
      const int nv = min(G, B - b0);
      asm volatile("{\n\t.reg .b64 st;\n\tmbarrier.arrive.expect_tx.shared::cta.b64 st, [%0], %1;\n\t}"
                   :: "r"(smem_u32(&rfull_bar[rb])), "r"(uint32_t(nv) * (A1_PER * 4)) : "memory");
      for (int sm = 0; sm < nv; ++sm)               // one copy per sample into its padded slot
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                     :: "r"(smem_u32(raw_s + rb * (G * RAW_S) + sm * RAW_S)), "l"(a1 + size_t(b0 + sm) * A1_PER),
                        "r"(uint32_t(A1_PER * 4)), "r"(smem_u32(&rfull_bar[rb])) : "memory");
    };
    if (pt == 0 && blockIdx.x < ngroups) issue_raw(blockIdx.x, 0);
    for (int grp = blockIdx.x; grp < ngroups; grp += gridDim.x, ++gi) {
      const int rb = gi & 1;
      const int nvalid = min(G, B - grp * G);
      // prefetch the next group's tile into the other buffer once its previous readers are done
      if (pt == 0 && grp + int(gridDim.x) < ngroups) {
        if (gi >= 1) mbar_wait(&rempty_bar[rb ^ 1], ((gi - 1) >> 1) & 1);
        issue_raw(grp + gridDim.x, rb ^ 1);
      }
      { TC_T0(tw); mbar_wait(&rfull_bar[rb], (gi >> 1) & 1); if (pt == 0) TC_ACC(1, tw); }   // slot 1: producer waits raw tile
      const float* raw = raw_s + rb * (G * RAW_S);
      for (int j = 0; j < KSTEPS; ++j, ++it) {
        const int slot = it % STAGES;
        { TC_T0(tw); if (it >= STAGES) mbar_wait(&empty_bar[slot], ((it / STAGES) - 1) & 1); if (pt == 0) TC_ACC(2, tw); }  // slot 2: producer waits empty
        TC_T0(tb);
        unsigned char* st = a_ring + slot * A_STAGE;
        // this k-step's 8 k values (warp-uniform): source offset ci*144 + kw, or "padding" for k >= 50
        int koff[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) { const int k = j * 8 + e; koff[e] = (k % 10) * 144 + k / 10; }
        {
#pragma unroll
          for (int q = 0; q < 3; ++q) {
            const float* src = raw + row_src[q];
            const bool live = row_s[q] < nvalid;
            float hi[8], lo[8];
#pragma unroll
            for (int e = 0; e < 8; ++e) {
              float v = 0.f;
              if (live && (j * 8 + e) < 50) v = src[koff[e]];
              split_tf32(v, hi[e], lo[e]);
            }
            const int off = (pt + 128 * q) * 16;
            *reinterpret_cast<float4*>(st + off) = make_float4(hi[0], hi[1], hi[2], hi[3]);
            *reinterpret_cast<float4*>(st + A_PLANE + off) = make_float4(hi[4], hi[5], hi[6], hi[7]);
            *reinterpret_cast<float4*>(st + A_HL + off) = make_float4(lo[0], lo[1], lo[2], lo[3]);
            *reinterpret_cast<float4*>(st + A_HL + A_PLANE + off) = make_float4(lo[4], lo[5], lo[6], lo[7]);
          }
        }
        if (pt == 0) TC_ACC(3, tb);   // slot 3: producer build time
        { TC_T0(tf); asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        mbar_arrive(&full_bar[slot]); if (pt == 0) TC_ACC(4, tf); }   // slot 4: fence + arrive
      }
      mbar_arrive(&rempty_bar[rb]);
    }
  } else if (warp >= 8) {
    // ======================= MMA issuers: warp 8+t owns M-tile t (a single thread's issue rate is the limit) ====
    if (lane == 0) {
      const int t = warp - 8;
      const uint32_t idesc_base = (1u << 4) | (2u << 7) | (2u << 10) | (uint32_t(128 >> 4) << 24);   // D f32, A/B tf32 K-major, M=128
      const uint32_t idesc_hl = idesc_base | (uint32_t(NHL >> 3) << 17), idesc32 = idesc_base | (uint32_t(NPAD >> 3) << 17);
      const uint32_t a_base = smem_u32(a_ring), b_base = smem_u32(b_mat);
      uint32_t it = 0, gi = 0;
      for (int grp = blockIdx.x; grp < ngroups; grp += gridDim.x, ++gi) {
        const int buf = gi & 1;
        { TC_T0(tw); if (gi >= 2) mbar_wait(&tempty_bar[buf], ((gi >> 1) - 1) & 1); if (t == 0) TC_ACC(5, tw); }   // slot 5: MMA waits tmem empty
        asm volatile("tcgen05.fence::after_thread_sync;");
        for (int j = 0; j < KSTEPS; ++j, ++it) {
          const int slot = it % STAGES;
          { TC_T0(tw); mbar_wait(&full_bar[slot], (it / STAGES) & 1); if (t == 0) TC_ACC(6, tw); }   // slot 6: MMA waits full
          TC_T0(ti);
          asm volatile("tcgen05.fence::after_thread_sync;");
          // descriptor arithmetic: the start-address field is the low 14 bits in units of 16 B, so a byte offset
          // is added as (bytes >> 4); rows are 16 B apart, hence "+ rows" below
          const uint64_t a_hi0 = make_desc(a_base + slot * A_STAGE, A_PLANE, 128);
          const uint64_t a_lo0 = a_hi0 + (A_HL >> 4);
          const uint64_t b0 = make_desc(b_base + j * B_TILE, B_PLANE, 128);
          const uint32_t d = tmem + uint32_t((buf * MT + t) * DCOLS);
          {
#pragma unroll
            for (int kh = 0; kh < 5; ++kh) {
              const uint64_t rows = uint64_t(t * 128 + kh * KH_ROWS);
              const uint64_t bsel = uint64_t((kh * KSTEPS * B_TILE) >> 4);
              umma_tf32(d, a_hi0 + rows, b0 + bsel, idesc_hl, (j | kh) ? 1u : 0u);  // hi*hi (0-19) | hi*lo (20-39)
              umma_tf32(d, a_lo0 + rows, b0 + bsel, idesc32, 1u);                   // + lo*hi into columns 0-19 (and lo*lo into 20-31)
            }
          }
          umma_commit(&empty_bar[slot]);            // stage reusable once these MMAs have read it
          if (t == 0) TC_ACC(7, ti);   // slot 7: MMA issue time
        }
        umma_commit(&tfull_bar[buf]);               // accumulators of this group complete
      }
    }
  } else {
    // ======================= epilogue (warps 0-3 <-> TMEM lanes 32*warp ..) =======================
    uint32_t gi = 0;
    for (int grp = blockIdx.x; grp < ngroups; grp += gridDim.x, ++gi) {
      const int buf = gi & 1;
      const int b0 = grp * G;
      // dropout noise of this group's (sample, channel) pairs: x * noise, noise in {0, 2} (1 in eval mode)
      float* nz = noise_s + buf * (G * C2);
      if (tid < G * C2) {
        const int s = tid / C2, co = tid % C2, b = b0 + s;
        float noise = 1.f;
        if (training && b < B) {
          const float keep = mask2d ? mask2d[size_t(b) * C2 + co] : philox_keep(seed, offset, 1u, uint64_t(b) * C2 + co);
          noise = keep * 2.f;
        }
        nz[tid] = noise;
      }
      asm volatile("bar.sync 1, 128;" ::: "memory");
      { TC_T0(tw); mbar_wait(&tfull_bar[buf], (gi >> 1) & 1); if (tid == 0) TC_ACC(8, tw); }   // slot 8: epilogue waits accumulators
      TC_T0(te);
      asm volatile("tcgen05.fence::after_thread_sync;");
#pragma unroll 1
      for (int t = 0; t < MT; ++t) {
        uint32_t acc[32], acc_hl[32];
        const uint32_t taddr = tmem + (uint32_t(warp * 32) << 16) + uint32_t((buf * MT + t) * DCOLS);
        tmem_ld32(taddr, acc);
        tmem_ld32(taddr + NPAD, acc_hl);
        // this warp's 32 lanes are conv row oy = 4t + warp of the 4 samples: lane = s*8 + ox
        const int s = lane >> 3, ox = lane & 7;
        const int b = b0 + s;
        const float* nzs = nz + s * C2;
        float* mine = xch_s + warp * (C2 * 32);
#pragma unroll
        for (int co = 0; co < C2; ++co)
          mine[co * 32 + lane] = (__uint_as_float(acc[co]) + __uint_as_float(co < 12 ? acc[C2 + co] : acc_hl[co - 12]) + bias_s[co]) * nzs[co];
        // the pool partner row lives in the neighbouring warp (warps 0|1 and 2|3 pair up); each warp of a pair pools
        // half of the channels: first-max in window scan order (oy,ox), (oy,ox+1), (oy+1,ox), (oy+1,ox+1)
        asm volatile("bar.sync %0, 64;" :: "r"(2 + (warp >> 1)) : "memory");
        const float* top = xch_s + (warp & ~1) * (C2 * 32);
        const float* bot = top + C2 * 32;
        const bool writer = !(ox & 1) && (b < B);
        const int pp = (2 * t + (warp >> 1)) * 4 + (ox >> 1);
        const int c0 = (warp & 1) * (C2 / 2);
#pragma unroll
        for (int cc = 0; cc < C2 / 2; ++cc) {
          const int co = c0 + cc;
          const float v00 = top[co * 32 + lane], v01 = top[co * 32 + (lane ^ 1)];
          const float v10 = bot[co * 32 + lane], v11 = bot[co * 32 + (lane ^ 1)];
          float best = v00; int bi = 0;
          if (v01 > best) { best = v01; bi = 1; }
          if (v10 > best) { best = v10; bi = 2; }
          if (v11 > best) { best = v11; bi = 3; }
          if (writer) {
            const size_t o = size_t(b) * A2_PER + co * 16 + pp;
            a2[o] = fmaxf(best, 0.f);
            idx2[o] = uint8_t(bi);
          }
        }
        asm volatile("bar.sync %0, 64;" :: "r"(2 + (warp >> 1)) : "memory");   // exchange buffers are free again
      }
      asm volatile("tcgen05.fence::before_thread_sync;");
      mbar_arrive(&tempty_bar[buf]);
      if (tid == 0) TC_ACC(9, te);   // slot 9: epilogue work
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;");
  __syncthreads();
  if (tid == 0) TC_ACC(10, t_kernel);   // slot 10: whole kernel
  if (warp == 8) {
    asm volatile("tcgen05.fence::after_thread_sync;");
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" :: "r"(tmem), "n"(TMEM_COLS));
  }
}

}  // namespace tc2

#ifdef DDP_TC_TIMING
extern "C" int ddp_debug_tc_timing(unsigned long long* out, int reset) {
  if (reset) { static unsigned long long z[160 * 16] = {0}; return (int)cudaMemcpyToSymbol(tc2::g_tc_dbg, z, sizeof(z)); }
  return (int)cudaMemcpyFromSymbol(out, tc2::g_tc_dbg, sizeof(unsigned long long) * 160 * 16);
}
#endif

int launch_conv2_fwd_tc(const float* params, const float* a1, const float* mask2d, uint64_t seed, uint64_t offset,
                        const int64_t* offset_dev, int training, float* a2, uint8_t* idx2, int B, cudaStream_t stream) {
  static bool attr_done = false;
  if (!attr_done) {
    DDP_CUDA(cudaFuncSetAttribute(tc2::conv2_fwd_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, tc2::SMEM_BYTES));
    attr_done = true;
  }
  const int ngroups = (B + tc2::G - 1) / tc2::G;
  const int sms = sm_count();
  const int grid = ngroups < sms ? ngroups : sms;
  DDP_CUDA(launch_pdl(tc2::conv2_fwd_tc_kernel, dim3(grid), dim3(tc2::THREADS), tc2::SMEM_BYTES, stream, true,
                      params, a1, mask2d, seed, offset, offset_dev, training, a2, idx2, B));
  return 0;
}

}  // namespace ddp
