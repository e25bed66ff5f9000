// K_B2w on the 5th-generation tensor cores: conv2 weight gradient as a "phase-expanded" GEMM with tcgen05.mma
// (kind::tf32, 3xTF32 split, fp32 accumulators in TMEM).   ref: horovod/mnist_horovod.py:20,63 (autograd of conv2).
//
// The gradient of conv2's output is non-zero only at the arg-max of each 2x2 pool window: entry (co, cell (py,px)) with
// value g sits at (oy, ox) = (2py + idy, 2px + idx), phase ph = 2*idy + idx.  Writing kh' = kh + idy, kw' = kw + idx:
//   dW[co][ci][kh][kw] = sum_ph  D'[(ci, kh+idy, kw+idx)][(ph, co)]
//   D'[(ci,kh',kw')][(ph,co)] = sum over (sample, cell)  A'[(ci,kh',kw')][(s,cell)] * G[(s,cell)][(ph,co)]
//   A'[(ci,kh',kw')][(s,py,px)] = a1[s][ci][2py+kh'][2px+kw']          360 rows: a 6x6 window per pool cell, 10 channels
//   G[(s,cell)][(ph,co)]       = g(s,co,cell) if its phase is ph, else 0     80 columns
// i.e. ONE dense GEMM, M = 360 (3 tiles of 128), N = 80, K = 16 per sample, whose operands do not depend on the
// data-dependent positions any more -- the positions became the column index.  (Three of four G entries are structural
// zeros: the tensor pipe does 4x the useful multiply-adds, x3 for the tf32 split, and is still far faster than the FMA
// or shared-memory pipes on this shape.)  Accumulation runs in TMEM over all samples of a persistent CTA; the
// epilogue folds the four phases back into dW and writes one partial row per CTA (reduced in fixed order by K_R).
//
// Shared-memory bandwidth is what bounds a small-N tcgen05 kernel (an MMA that reads a 128x8 A tile from shared memory
// spends more time fetching it than multiplying), so A' never touches shared memory: each producer thread owns one ROW
// of A' (= one TMEM lane), gathers its 8 values of the k-step from the raw a1 tile, splits them into tf32 hi/lo in
// registers and writes them straight into TMEM with tcgen05.st; the MMAs take A from TMEM and only the small G tile
// (80 x 8) from shared memory.
//
// Warp roles (288 threads, 1 CTA/SM):
//   warps 0-7  producers: the sample's raw tiles (a1, a2, da2, idx2) arrive by cp.async.bulk (3-deep ring, mbarrier
//              complete_tx); per k-step (8 cells = half a sample) warp w fills the TMEM lanes 32*(w%4).. of its M-tiles
//              (warps 0-3: tiles 0 and 2, warps 4-7: tile 1) and 40 threads of warps 4-5 build G (hi/lo, un-swizzled
//              K-major 16-byte chunks) in a 3-stage ring
//   warp 8     one lane issues 9 MMAs per k-step (3 M-tiles x {hi*hi, hi*lo, lo*hi}) and the commits; owns TMEM
//   warps 0-3  after the last k-step: tcgen05.ld the accumulators, phase fold through shared memory
#include "common.cuh"
#include "tc_common.cuh"

namespace ddp {
namespace wtc {
using namespace ::ddp::tc;

constexpr int MROWS = 360;                  // (ci, kh', kw') = 10 x 6 x 6
constexpr int MPAD = 384;                   // 3 M-tiles of 128 (rows 360..383 are zero)
static_assert(MPAD == 3 * 128, "three M-tiles");
constexpr int MT = 3;
constexpr int NB = 80;                      // (phase, co) = 4 x 20
constexpr int TCOL = 96;                    // TMEM column stride between the tiles' accumulators
constexpr int TMEM_COLS = 512;              // 3 x 96 accumulator columns + 3 stages x 3 tiles x 16 operand columns = 432
constexpr int STAGES = 3;
constexpr int ACOL0 = MT * TCOL;            // first operand column: stage s, tile t -> ACOL0 + (s*3 + t)*16 (+8 for lo)
constexpr int B_PLANE = NB * 16;            // bytes of one k-chunk plane of G (4 k values x 80 rows)
constexpr int B_HL = 2 * B_PLANE;           // one k-step of hi (or lo)
constexpr int STAGE_BYTES = 2 * B_HL;       // B_hi | B_lo = 5,120
constexpr int RAWS = 3;
constexpr int RAW_BYTES = A1_PER * 4 + 2 * A2_PER * 4 + A2_PER;        // a1 | a2 | da2 | idx2 = 8,640
constexpr int PROD_THREADS = 256;
constexpr int THREADS = PROD_THREADS + 32;
constexpr int DS = 81;                      // row stride of the staged D' (floats)
constexpr int SMEM_MAIN = MROWS * DS * 4 > STAGES * STAGE_BYTES + RAWS * RAW_BYTES ? MROWS * DS * 4
                                                                                    : STAGES * STAGE_BYTES + RAWS * RAW_BYTES;
constexpr int SMEM_BYTES = ((SMEM_MAIN + 127) / 128) * 128 + 256;

// A from tensor memory (128 lanes x 8 columns of tf32), B from shared memory
__device__ __forceinline__ void umma_tf32_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t bdesc, uint32_t idesc, uint32_t acc) {
  asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
               "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n\t}"
               :: "r"(tmem_d), "r"(tmem_a), "l"(bdesc), "r"(idesc), "r"(acc) : "memory");
}

__global__ void __launch_bounds__(THREADS, 1)
conv2_wgrad_tc_kernel(const float* __restrict__ da2, const float* __restrict__ a2,
                      const uint8_t* __restrict__ idx2, const float* __restrict__ a1,
                      float drop_scale, float* __restrict__ partials, int B) {
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  unsigned char* ring = smem_raw;
  unsigned char* raws = smem_raw + STAGES * STAGE_BYTES;
  unsigned char* misc = smem_raw + ((SMEM_MAIN + 127) / 128) * 128;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(misc);            // [STAGES] producers -> MMA
  uint64_t* empty_bar = full_bar + STAGES;                           // [STAGES] MMA -> producers
  uint64_t* rfull_bar = empty_bar + STAGES;                          // [RAWS]   raw tiles landed (tx bytes)
  uint64_t* rempty_bar = rfull_bar + RAWS;                           // [RAWS]   producers done reading them
  uint64_t* tfull_bar = rempty_bar + RAWS;                           // accumulators complete
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(tfull_bar + 1);
  float* bias_s = reinterpret_cast<float*>(misc + 128);              // [20]

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  pdl_trigger();
  // ---- one-time setup: zero the operand ring (pad rows stay finite), barriers, TMEM ----
  for (int i = tid; i < STAGES * STAGE_BYTES / 16; i += THREADS) reinterpret_cast<float4*>(ring)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
  if (tid == 0) {
    for (int s = 0; s < STAGES; ++s) { mbar_init(&full_bar[s], PROD_THREADS); mbar_init(&empty_bar[s], 1); }
    for (int s = 0; s < RAWS; ++s) { mbar_init(&rfull_bar[s], 1); mbar_init(&rempty_bar[s], PROD_THREADS); }
    mbar_init(tfull_bar, 1);
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
  pdl_wait();                                  // da2 (head_bwd) and a1/a2/idx2 are read from here on
  const uint32_t tmem = *tmem_ptr;
  const int n = (B - int(blockIdx.x) + int(gridDim.x) - 1) / int(gridDim.x);   // samples of this CTA: b = blockIdx + k*grid

  if (warp < 8) {
    // ======================= producers =======================
    // this thread's rows of A' (= its TMEM lane in tiles t0 and, for warps 0-3, tile 2) are fixed: decode them once
    const int quad = warp & 3, lrow = quad * 32 + lane;             // TMEM lane this thread may write
    const int t0 = warp < 4 ? 0 : 1, nrows = warp < 4 ? 2 : 1;
    int src[2];
    bool live[2];
#pragma unroll
    for (int r = 0; r < 2; ++r) {
      const int m = (t0 + 2 * r) * 128 + lrow;
      live[r] = r < nrows && m < MROWS;
      const int mm = live[r] ? m : 0, ci = mm / 36, khp = (mm % 36) / 6, kwp = mm % 6;
      src[r] = ci * 144 + khp * P1 + kwp;                           // + (4j + 2*plane)*12 + 2*px
    }
    // G: thread 128 + 2*co + plane builds the four phase rows of (co, plane) from one look at the entries
    const bool doB = tid >= 128 && tid < 128 + 2 * C2;
    const int bco = (tid - 128) >> 1, bplane = (tid - 128) & 1;
    auto issue_raw = [&](int k) {                                   // thread 0: raw tiles of the CTA's k-th sample
      const size_t b = size_t(blockIdx.x) + size_t(k) * gridDim.x;
      unsigned char* r = raws + (k % RAWS) * RAW_BYTES;
      uint64_t* bar = &rfull_bar[k % RAWS];
      mbarrier_expect_tx(bar, RAW_BYTES);
      bulk_copy_g2s(r, a1 + b * A1_PER, A1_PER * 4, bar);
      bulk_copy_g2s(r + A1_PER * 4, a2 + b * A2_PER, A2_PER * 4, bar);
      bulk_copy_g2s(r + A1_PER * 4 + A2_PER * 4, da2 + b * A2_PER, A2_PER * 4, bar);
      bulk_copy_g2s(r + A1_PER * 4 + 2 * A2_PER * 4, idx2 + b * A2_PER, A2_PER, bar);
    };
    if (tid == 0) {
      if (0 < n) issue_raw(0);
      if (1 < n) issue_raw(1);
    }
    float bacc = 0.f;                                               // tid < 20: bias gradient of channel tid
    uint32_t it = 0;
    for (int k = 0; k < n; ++k) {
      const int rb = k % RAWS;
      if (tid == 0 && k + 2 < n) {                                  // refill the buffer sample k-1 used
        if (k >= 1) mbar_wait(&rempty_bar[(k + 2) % RAWS], ((k - 1) / RAWS) & 1);
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        issue_raw(k + 2);
      }
      mbar_wait(&rfull_bar[rb], (k / RAWS) & 1);
      const float* a1_s = reinterpret_cast<const float*>(raws + rb * RAW_BYTES);
      const float* a2_s = a1_s + A1_PER;
      const float* da2_s = a2_s + A2_PER;
      const uint8_t* id_s = reinterpret_cast<const uint8_t*>(da2_s + A2_PER);
#pragma unroll 1
      for (int j = 0; j < 2; ++j, ++it) {
        const int slot = it % STAGES;
        if (it >= STAGES) {
          mbar_wait(&empty_bar[slot], ((it / STAGES) - 1) & 1);     // the MMAs that read this stage have finished
          asm volatile("tcgen05.fence::after_thread_sync;");
        }
#pragma unroll
        for (int r = 0; r < 2; ++r) {
          if (r >= nrows) continue;                                 // warp-uniform
          uint32_t hi[8], lo[8];
          const float* s0 = a1_s + src[r] + 48 * j;
#pragma unroll
          for (int e = 0; e < 8; ++e) {                             // k = plane*4 + px
            float v = live[r] ? s0[(e >> 2) * 2 * P1 + 2 * (e & 3)] : 0.f;
            float h, l;
            split_tf32(v, h, l);
            hi[e] = __float_as_uint(h); lo[e] = __float_as_uint(l);
          }
          const uint32_t ta = tmem + (uint32_t(quad * 32) << 16) + uint32_t(ACOL0 + (slot * MT + t0 + 2 * r) * 16);
          asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};"
                       :: "r"(ta), "r"(hi[0]), "r"(hi[1]), "r"(hi[2]), "r"(hi[3]), "r"(hi[4]), "r"(hi[5]), "r"(hi[6]), "r"(hi[7]) : "memory");
          asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};"
                       :: "r"(ta + 8), "r"(lo[0]), "r"(lo[1]), "r"(lo[2]), "r"(lo[3]), "r"(lo[4]), "r"(lo[5]), "r"(lo[6]), "r"(lo[7]) : "memory");
        }
        if (doB) {
          unsigned char* st = ring + slot * STAGE_BYTES;
          const int e0 = bco * 16 + (2 * j + bplane) * 4;           // entries (co, py = 2j + plane, px 0..3)
          const float4 a = *reinterpret_cast<const float4*>(a2_s + e0);
          const float4 d = *reinterpret_cast<const float4*>(da2_s + e0);
          const unsigned ids = *reinterpret_cast<const unsigned*>(id_s + e0);
          float g[4];
          g[0] = a.x > 0.f ? d.x * drop_scale : 0.f;
          g[1] = a.y > 0.f ? d.y * drop_scale : 0.f;
          g[2] = a.z > 0.f ? d.z * drop_scale : 0.f;
          g[3] = a.w > 0.f ? d.w * drop_scale : 0.f;
          float gh[4], gl[4];
#pragma unroll
          for (int e = 0; e < 4; ++e) split_tf32(g[e], gh[e], gl[e]);
#pragma unroll
          for (int ph = 0; ph < 4; ++ph) {
            float4 h4, l4;
            h4.x = ((ids) & 3u) == unsigned(ph) ? gh[0] : 0.f;        l4.x = ((ids) & 3u) == unsigned(ph) ? gl[0] : 0.f;
            h4.y = ((ids >> 8) & 3u) == unsigned(ph) ? gh[1] : 0.f;   l4.y = ((ids >> 8) & 3u) == unsigned(ph) ? gl[1] : 0.f;
            h4.z = ((ids >> 16) & 3u) == unsigned(ph) ? gh[2] : 0.f;  l4.z = ((ids >> 16) & 3u) == unsigned(ph) ? gl[2] : 0.f;
            h4.w = ((ids >> 24) & 3u) == unsigned(ph) ? gh[3] : 0.f;  l4.w = ((ids >> 24) & 3u) == unsigned(ph) ? gl[3] : 0.f;
            const int off = bplane * B_PLANE + (ph * C2 + bco) * 16;
            *reinterpret_cast<float4*>(st + off) = h4;
            *reinterpret_cast<float4*>(st + B_HL + off) = l4;
          }
          asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        }
        asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
        asm volatile("tcgen05.fence::before_thread_sync;");
        mbar_arrive(&full_bar[slot]);
      }
      if (tid < C2) {
#pragma unroll
        for (int pp = 0; pp < 16; ++pp) {
          const int e = tid * 16 + pp;
          bacc += a2_s[e] > 0.f ? da2_s[e] * drop_scale : 0.f;
        }
      }
      mbar_arrive(&rempty_bar[rb]);
    }
    if (tid < C2) bias_s[tid] = bacc;
  } else if (lane == 0) {
    // ======================= MMA issuer =======================
    const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | (uint32_t(NB >> 3) << 17) | (uint32_t(128 >> 4) << 24);
    const uint32_t ring_base = smem_u32(ring);
    const int ksteps = 2 * n;
    for (int it = 0; it < ksteps; ++it) {
      const int slot = it % STAGES;
      mbar_wait(&full_bar[slot], (it / STAGES) & 1);
      asm volatile("tcgen05.fence::after_thread_sync;");
      const uint32_t sb = ring_base + slot * STAGE_BYTES;
      const uint64_t b_hi = make_desc(sb, B_PLANE, 128);
      const uint64_t b_lo = make_desc(sb + B_HL, B_PLANE, 128);
#pragma unroll
      for (int t = 0; t < MT; ++t) {
        const uint32_t a_hi = tmem + uint32_t(ACOL0 + (slot * MT + t) * 16), a_lo = a_hi + 8;
        const uint32_t d = tmem + uint32_t(t * TCOL);
        umma_tf32_ts(d, a_hi, b_hi, idesc, it ? 1u : 0u);
        umma_tf32_ts(d, a_hi, b_lo, idesc, 1u);
        umma_tf32_ts(d, a_lo, b_hi, idesc, 1u);
      }
      umma_commit(&empty_bar[slot]);             // stage reusable once these MMAs have read it
    }
    umma_commit(tfull_bar);                      // accumulators complete
  }
  __syncthreads();                               // producers are done with the ring / raw tiles, the issuer has issued everything
  float* Ds = reinterpret_cast<float*>(smem_raw);                    // [360][81] staged D' (aliases the operand ring)
  if (warp < 4) {
    mbar_wait(tfull_bar, 0);                     // ... and every MMA has finished reading the ring
    asm volatile("tcgen05.fence::after_thread_sync;");
#pragma unroll 1
    for (int t = 0; t < MT; ++t) {
      const int m = t * 128 + warp * 32 + lane;
#pragma unroll 1
      for (int c3 = 0; c3 < 3; ++c3) {
        uint32_t r[32];
        tmem_ld32(tmem + (uint32_t(warp * 32) << 16) + uint32_t(t * TCOL + 32 * c3), r);
        if (m < MROWS) {
#pragma unroll
          for (int c = 0; c < 32; ++c)
            if (32 * c3 + c < NB) Ds[m * DS + 32 * c3 + c] = __uint_as_float(r[c]);
        }
      }
    }
    asm volatile("tcgen05.fence::before_thread_sync;");
  }
  __syncthreads();
  // phase fold: dW[co][ci][kh][kw] = sum_ph D'[(ci, kh + idy, kw + idx)][ph*20 + co], fixed order
  float* part = partials + size_t(blockIdx.x) * P_ARENA;
  for (int o = tid; o < C2 * 250; o += THREADS) {
    const int co = o / 250, r = o % 250, ci = r / 25, kh = (r % 25) / 5, kw = r % 5;
    const float* d = Ds + (ci * 36 + kh * 6 + kw) * DS + co;
    part[OFF_C2W + o] = ((d[0] + d[DS + C2]) + d[6 * DS + 2 * C2]) + d[7 * DS + 3 * C2];
  }
  if (tid < C2) part[OFF_C2B + tid] = bias_s[tid];
  if (warp == 8) {
    asm volatile("tcgen05.fence::after_thread_sync;");
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" :: "r"(tmem), "n"(TMEM_COLS));
  }
}

}  // namespace wtc

int launch_conv2_wgrad_tc(const float* da2, const float* a2, const uint8_t* idx2, const float* a1, float drop_scale,
                          float* partials, int B, int* parts_out, cudaStream_t stream, bool pdl) {
  static bool done[64] = {false};
  int dev = 0;
  DDP_CUDA(cudaGetDevice(&dev));
  if (dev >= 0 && dev < 64 && !done[dev]) {
    DDP_CUDA(cudaFuncSetAttribute(wtc::conv2_wgrad_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, wtc::SMEM_BYTES));
    done[dev] = true;
  }
  int grid = sm_count();
  if (grid > B) grid = B;
  if (grid > MAX_PARTS) grid = MAX_PARTS;
  *parts_out = grid;
  DDP_CUDA(launch_pdl(wtc::conv2_wgrad_tc_kernel, dim3(grid), dim3(wtc::THREADS), wtc::SMEM_BYTES, stream, pdl, da2, a2, idx2, a1,
                      drop_scale, partials, B));
  return 0;
}

}  // namespace ddp
