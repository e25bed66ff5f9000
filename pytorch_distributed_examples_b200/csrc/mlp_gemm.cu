// 3xTF32 GEMM on the 5th-generation tensor cores for the MLP of the DDP script
// (ref: pytorch_elastic/mnist_ddp_elastic.py:133-159, Model(hidden_layers=5, features=1024)).
//
//   C[M][N] = A . B^T,  each operand K-major ([rows][K], K contiguous) or MN-major ([K][rows]) in global memory:
//     forward        x @ W^T        A = X   [B][in]  K-major     B = W [out][in] K-major
//     data gradient  dY @ W         A = dY  [B][out] K-major     B = W [out][in] MN-major (n = in, k = out)
//     weight grad.   dY^T @ X       A = dY  [B][out] MN-major    B = X [B][in]   MN-major (k = batch)
//   so ONE row-major copy of every activation, gradient and weight plane feeds all three (no transposes anywhere).
// fp32 results within the 1e-5 parity bar from tf32 tensor cores: every operand is kept as a pair of tf32-exact
// planes (hi = rna_tf32(v), lo = rna_tf32(v - hi)) and each k-step issues three tcgen05.mma.kind::tf32 into one
// fp32 TMEM accumulator:  lo*hi + hi*lo + hi*hi   (lo*lo, 2^-22 relative, is dropped).
// The tensor core adds into its fp32 accumulator with TRUNCATION, a bias that grows linearly with the number of
// accumulation steps (measured: 4e-3 absolute at K = 4096 on unit-variance data, 30x worse than an fp32 dot
// product).  So an accumulator only ever sees KCHUNK k-blocks (128 k values): the epilogue warps drain it into
// round-to-nearest fp32 running sums held in registers, alternating between the two TMEM buffers.
//
// One persistent CTA per SM, 192 threads, warp-specialised:
//   warp 0    producer: one lane issues four TMA tile loads per k-block (cp.async.bulk.tensor, 128-byte swizzle,
//             out-of-range k zero-filled by the TMA unit) into a STAGES-deep shared-memory ring.  K-major tile:
//             2-D box {32 k, 128 rows}.  MN-major tile: 3-D box {32 rows, 32 k, 4 row-blocks} over the view
//             (row % 32, k, row / 32) of the same matrix, which lands as [row-block][k][32 rows] = the canonical
//             MN-major layout for 32-bit operands (128-byte swizzle with 32-byte atoms, the only one tf32 allows)
//   warp 1    one lane issues the 12 MMAs of a k-block (4 k-steps of 8 x 3 products) and tcgen05.commit's the
//             stage back to the producer; owns the TMEM allocation (two accumulator buffers of BN columns)
//   warps 2-5 epilogue: tcgen05.ld every finished 128 x BN chunk accumulator (one row per thread) and add it to the
//             running sums; after the last chunk apply bias+ReLU (recording the ReLU mask as one bit per element)
//             or the mask, split into hi/lo, stage the tile in swizzled shared memory and TMA-store it --
//             overlapped with the following MMAs through the second accumulator buffer.
// Several problems (the six weight-gradient GEMMs of a step) can share one launch: tiles are numbered across
// problems and dealt round-robin to the CTAs.
#include "mlp_gemm.cuh"

namespace ddp {
namespace mlp {

// Developer aid: -DDDP_MLP_TIMING accumulates clock64() per warp role into g_mlp_dbg (ddp_debug_mlp_timing,
// tools/mlp_timing.py): how the producer / MMA / epilogue balance of this kernel was tuned.
#ifdef DDP_MLP_TIMING
__device__ unsigned long long g_mlp_dbg[160 * 16];
#define MT_T0(var) const long long var = clock64()
#define MT_LOCALS long long mt_loc[16] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0}
#define MT_ACC(slot, t0) do { mt_loc[slot] += clock64() - (t0); } while (0)          // registers: no memory traffic in the loops
#define MT_FLUSH(slot) do { g_mlp_dbg[blockIdx.x * 16 + (slot)] += (unsigned long long)mt_loc[slot]; } while (0)
#else
#define MT_T0(var)
#define MT_LOCALS
#define MT_ACC(slot, t0)
#define MT_FLUSH(slot)
#endif

constexpr int BM = 128;
constexpr int BK = 32;                       // 32 tf32 = 128 bytes = one swizzle row
constexpr int THREADS = 192;
constexpr int A_BYTES = BM * BK * 4;         // 16 KB per plane
constexpr int KCHUNK = 4;                    // k-blocks accumulated in TMEM before the epilogue drains them

template <int BN>
struct Cfg {
  static constexpr int STAGES = 3;
  static constexpr int B_BYTES = BN * BK * 4;
  static constexpr int STAGE_BYTES = 2 * A_BYTES + 2 * B_BYTES;
  static constexpr int STORE_BYTES = BM * 32 * 4;                        // one 128 x 32 fp32 staging tile for the TMA store
  static constexpr int BAR_BYTES = 256;
  static constexpr int SMEM = STAGES * STAGE_BYTES + 2 * STORE_BYTES + BAR_BYTES + 1024;   // + slack to align to 1024
  static constexpr int TMEM_COLS = 2 * BN;
};

__device__ __forceinline__ uint32_t s32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void bar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(bar), "r"(count));
}
__device__ __forceinline__ void bar_arrive(uint32_t bar) {
  asm volatile("{\n\t.reg .b64 st;\n\tmbarrier.arrive.shared::cta.b64 st, [%0];\n\t}" :: "r"(bar) : "memory");
}
__device__ __forceinline__ void bar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("{\n\t.reg .b64 st;\n\tmbarrier.arrive.expect_tx.shared::cta.b64 st, [%0], %1;\n\t}" :: "r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bar_wait(uint32_t bar, uint32_t parity) {
  uint32_t done = 0, spins = 0;
  while (!done) {
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.b32 %0, 1, 0, p;\n\t}"
                 : "=r"(done) : "r"(bar), "r"(parity) : "memory");
    if (++spins > (1u << 26)) __trap();      // a protocol bug must abort the launch, never hang the GPU
  }
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap* map, int c0, int c1, uint32_t bar) {
  asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
               :: "r"(dst), "l"(map), "r"(c0), "r"(c1), "r"(bar) : "memory");
}
__device__ __forceinline__ void tma_load_3d(uint32_t dst, const CUtensorMap* map, int c0, int c1, int c2, uint32_t bar) {
  asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
               :: "r"(dst), "l"(map), "r"(c0), "r"(c1), "r"(c2), "r"(bar) : "memory");
}
__device__ __forceinline__ void tma_store_2d(const CUtensorMap* map, uint32_t src, int c0, int c1) {
  asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.tile.bulk_group [%0, {%2, %3}], [%1];"
               :: "l"(map), "r"(src), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void sts128(uint32_t addr, float a, float b, float c, float d) {
  asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" :: "r"(addr), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}
// K-major operand tile written by TMA with 128-byte swizzle: rows of 128 B, 8-row groups 1024 B apart
__device__ __forceinline__ uint64_t sw128_desc(uint32_t saddr) {
  return (uint64_t)((saddr & 0x3FFFFu) >> 4) | ((uint64_t)1 << 16) | ((uint64_t)(1024 >> 4) << 32) |
         ((uint64_t)1 << 46) | ((uint64_t)2 << 61);
}
// MN-major tf32 operand tile: [row-block of 32][k][32 rows], written by TMA with CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B.
// 32-bit MN-major operands have ONE legal layout, "128-byte swizzle with 32-byte atoms" (layout type 1): 32-byte chunks
// of a 128-byte row XOR-ed with (row % 4), atoms of 4 k-rows.  LBO = bytes between row-blocks, SBO = bytes between
// 4-row k groups (an MMA of K = 8 spans two of them).
__device__ __forceinline__ uint64_t sw128_desc_mn(uint32_t saddr) {
  return (uint64_t)((saddr & 0x3FFFFu) >> 4) | ((uint64_t)((BK * 128) >> 4) << 16) | ((uint64_t)(512 >> 4) << 32) |
         ((uint64_t)1 << 46) | ((uint64_t)1 << 61);
}
__device__ __forceinline__ void umma(uint32_t d, uint64_t a, uint64_t b, uint32_t idesc, uint32_t acc) {
  asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
               "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
               :: "r"(d), "l"(a), "l"(b), "r"(idesc), "r"(acc) : "memory");
}
__device__ __forceinline__ void umma_commit_to(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" :: "r"(bar) : "memory");
}
__device__ __forceinline__ void tmem_ld_32cols(uint32_t taddr, uint32_t (&r)[32]) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 "
               "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
                 "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
                 "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
                 "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
               : "r"(taddr));
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ float to_tf32(float v) {
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(v));
  return __uint_as_float(r);
}

struct TileCoord { int prob, m0, n0; };
template <int BN>
__device__ __forceinline__ TileCoord locate(const GemmBatch& batch, int t) {
  int pi = 0;
  while (pi + 1 < batch.count && t >= batch.tile_start[pi + 1]) ++pi;
  const int local = t - batch.tile_start[pi];
  const int mt = local % batch.p[pi].mtiles, nt = local / batch.p[pi].mtiles;
  return TileCoord{pi, mt * BM, nt * BN};
}

template <int BN, int EPI, bool AMN, bool BMN>
__global__ void __launch_bounds__(THREADS, 1)
gemm3x_kernel(const __grid_constant__ GemmBatch batch) {
  using C = Cfg<BN>;
  extern __shared__ unsigned char smem_raw[];
  const uint32_t raw = s32(smem_raw);
  const uint32_t ring = (raw + 1023u) & ~1023u;                 // 128-byte swizzle needs 1024-byte aligned tiles
  unsigned char* gen_ring = smem_raw + (ring - raw);
  const uint32_t stage_out = ring + C::STAGES * C::STAGE_BYTES;  // two swizzled 128 x 32 tiles feeding the TMA stores
  const uint32_t bars = stage_out + 2 * C::STORE_BYTES;         // full[STAGES], empty[STAGES], tfull[2], tempty[2], tmem ptr
  auto full = [&](int s) { return bars + 8u * s; };
  auto empty = [&](int s) { return bars + 8u * (C::STAGES + s); };
  auto tfull = [&](int b) { return bars + 8u * (2 * C::STAGES + b); };
  auto tempty = [&](int b) { return bars + 8u * (2 * C::STAGES + 2 + b); };
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(gen_ring + C::STAGES * C::STAGE_BYTES + 2 * C::STORE_BYTES + 8 * (2 * C::STAGES + 4));

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  pdl_trigger();
  MT_T0(t_kernel);
  MT_LOCALS;
  if (tid == 0) {
    for (int s = 0; s < C::STAGES; ++s) { bar_init(full(s), 1); bar_init(empty(s), 1); }
    for (int b = 0; b < 2; ++b) { bar_init(tfull(b), 1); bar_init(tempty(b), 128); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" :: "r"(s32(tmem_slot)), "n"(C::TMEM_COLS));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  asm volatile("tcgen05.fence::before_thread_sync;");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;");
  pdl_wait();                                  // barriers, TMEM and the tensor maps are set up; operands come from earlier kernels
  const uint32_t tmem = *tmem_slot;
  const int total = batch.tile_start[batch.count];

  if (warp == 0) {
    // ======================= producer =======================
    if (lane == 0) {
      uint32_t it = 0;
      for (int t = blockIdx.x; t < total; t += gridDim.x) {
        const TileCoord tc = locate<BN>(batch, t);
        const GemmProblem& P = batch.p[tc.prob];
        const int kblocks = (P.K + BK - 1) / BK;
        for (int kb = 0; kb < kblocks; ++kb, ++it) {
          const int slot = it % C::STAGES;
          { MT_T0(tw); if (it >= C::STAGES) bar_wait(empty(slot), ((it / C::STAGES) - 1) & 1); MT_ACC(1, tw); }   // 1: producer waits empty
          const uint32_t st = ring + slot * C::STAGE_BYTES;
          bar_expect_tx(full(slot), C::STAGE_BYTES);
          if (AMN) {
            tma_load_3d(st, &P.a_hi, 0, kb * BK, tc.m0 >> 5, full(slot));
            tma_load_3d(st + A_BYTES, &P.a_lo, 0, kb * BK, tc.m0 >> 5, full(slot));
          } else {
            tma_load_2d(st, &P.a_hi, kb * BK, tc.m0, full(slot));
            tma_load_2d(st + A_BYTES, &P.a_lo, kb * BK, tc.m0, full(slot));
          }
          if (BMN) {
            tma_load_3d(st + 2 * A_BYTES, &P.b_hi, 0, kb * BK, tc.n0 >> 5, full(slot));
            tma_load_3d(st + 2 * A_BYTES + C::B_BYTES, &P.b_lo, 0, kb * BK, tc.n0 >> 5, full(slot));
          } else {
            tma_load_2d(st + 2 * A_BYTES, &P.b_hi, kb * BK, tc.n0, full(slot));
            tma_load_2d(st + 2 * A_BYTES + C::B_BYTES, &P.b_lo, kb * BK, tc.n0, full(slot));
          }
        }
      }
      MT_FLUSH(1);
    }
  } else if (warp == 1) {
    // ======================= MMA issuer =======================
    if (lane == 0) {
      const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | (AMN ? (1u << 15) : 0u) | (BMN ? (1u << 16) : 0u) |
                             (uint32_t(BN >> 3) << 17) | (uint32_t(BM >> 4) << 24);
      uint32_t it = 0, sub = 0;                                  // running k-block / chunk counters
      for (int t = blockIdx.x; t < total; t += gridDim.x) {
        const TileCoord tc = locate<BN>(batch, t);
        const int kblocks = (batch.p[tc.prob].K + BK - 1) / BK;
        for (int kb0 = 0; kb0 < kblocks; kb0 += KCHUNK, ++sub) {
          const int buf = sub & 1;
          { MT_T0(tw); if (sub >= 2) bar_wait(tempty(buf), ((sub >> 1) - 1) & 1); MT_ACC(2, tw); }               // 2: MMA waits tmem empty
          asm volatile("tcgen05.fence::after_thread_sync;");
          const uint32_t d = tmem + uint32_t(buf * BN);
          const int kend = kb0 + KCHUNK < kblocks ? kb0 + KCHUNK : kblocks;
          for (int kb = kb0; kb < kend; ++kb, ++it) {
            const int slot = it % C::STAGES;
            { MT_T0(tw); bar_wait(full(slot), (it / C::STAGES) & 1); MT_ACC(3, tw); }                               // 3: MMA waits full
            if (it == 0) MT_ACC(8, t_kernel);                                                                     // 8: kernel start -> first operands landed
            MT_T0(ti);
            asm volatile("tcgen05.fence::after_thread_sync;");
            const uint32_t st = ring + slot * C::STAGE_BYTES;
            const uint64_t a_hi = AMN ? sw128_desc_mn(st) : sw128_desc(st);
            const uint64_t a_lo = AMN ? sw128_desc_mn(st + A_BYTES) : sw128_desc(st + A_BYTES);
            const uint64_t b_hi = BMN ? sw128_desc_mn(st + 2 * A_BYTES) : sw128_desc(st + 2 * A_BYTES);
            const uint64_t b_lo = BMN ? sw128_desc_mn(st + 2 * A_BYTES + C::B_BYTES) : sw128_desc(st + 2 * A_BYTES + C::B_BYTES);
#pragma unroll
            for (int ks = 0; ks < BK / 8; ++ks) {
              // one k-step = 8 tf32: 32 bytes along a K-major row, or 8 rows = one 1024-byte atom of an MN-major tile
              const uint64_t oa = uint64_t(ks * (AMN ? 64 : 2)), ob = uint64_t(ks * (BMN ? 64 : 2));   // 16-byte units
              umma(d, a_lo + oa, b_hi + ob, idesc, (kb > kb0 || ks) ? 1u : 0u);
              umma(d, a_hi + oa, b_lo + ob, idesc, 1u);
              umma(d, a_hi + oa, b_hi + ob, idesc, 1u);
            }
            umma_commit_to(empty(slot));
            MT_ACC(4, ti);                                                                                        // 4: MMA issue
          }
          umma_commit_to(tfull(buf));
        }
      }
      MT_ACC(9, t_kernel);                                                                                        // 9: kernel start -> last MMA issued
      MT_FLUSH(2); MT_FLUSH(3); MT_FLUSH(4); MT_FLUSH(8); MT_FLUSH(9);
    }
  } else {
    // ======================= epilogue: warp w reads TMEM lanes 32*(w%4).. =======================
    const int q = warp & 3;
    uint32_t sub = 0;
    for (int t = blockIdx.x; t < total; t += gridDim.x) {
      const TileCoord tc = locate<BN>(batch, t);
      const GemmProblem& P = batch.p[tc.prob];
      const int kblocks = (P.K + BK - 1) / BK;
      float sum[BN];
#pragma unroll
      for (int j = 0; j < BN; ++j) sum[j] = 0.f;
      for (int kb0 = 0; kb0 < kblocks; kb0 += KCHUNK, ++sub) {
        const int buf = sub & 1;
        { MT_T0(tw); bar_wait(tfull(buf), (sub >> 1) & 1); MT_ACC(5, tw); }                          // 5: epilogue waits accumulators
        MT_T0(td);
        asm volatile("tcgen05.fence::after_thread_sync;");
#pragma unroll
        for (int c = 0; c < BN / 32; ++c) {
          uint32_t acc[32];
          tmem_ld_32cols(tmem + (uint32_t(q * 32) << 16) + uint32_t(buf * BN + c * 32), acc);
#pragma unroll
          for (int j = 0; j < 32; ++j) sum[c * 32 + j] += __uint_as_float(acc[j]);
        }
        asm volatile("tcgen05.fence::before_thread_sync;");
        bar_arrive(tempty(buf));
        MT_ACC(6, td);                                                                             // 6: drain
      }
      MT_T0(ts);
      const int row = q * 32 + lane;
      const int m = tc.m0 + row;
      // problem fields into registers once per tile: left in param space they are re-fetched (indexed LDC) per store
      const int pM = P.M, pN = P.N, bits_ld = P.bits_ld;
      const float* bias = P.bias;
      const uint32_t* bits_in = P.bits_in;
      uint32_t* bits_out = P.bits_out;
      const bool mok = m < pM;
      const bool elected = (tid == 64);
      const uint32_t st0 = stage_out + row * 128, st1 = st0 + C::STORE_BYTES;
#pragma unroll
      for (int c = 0; c < BN / 32; ++c) {
        const int nb = tc.n0 + c * 32;
        if (nb >= pN) continue;                                 // uniform over the four epilogue warps
        const bool full_chunk = nb + 32 <= pN;
        float* acc = sum + c * 32;
        float lo[32];
        if (EPI == EPI_BIAS_RELU_SPLIT) {
          uint32_t bits = 0u;
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            const float b = (full_chunk || nb + j < pN) ? __ldg(bias + nb + j) : 0.f;
            const float v = fmaxf(acc[j] + b, 0.f);
            bits |= (v > 0.f ? 1u : 0u) << j;
            acc[j] = to_tf32(v); lo[j] = to_tf32(v - acc[j]);
          }
          if (mok) bits_out[size_t(m) * bits_ld + (nb >> 5)] = bits;
        } else if (EPI == EPI_MASK_SPLIT) {
          const uint32_t bits = mok ? __ldg(bits_in + size_t(m) * bits_ld + (nb >> 5)) : 0u;
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            const float v = ((bits >> j) & 1u) ? acc[j] : 0.f;
            acc[j] = to_tf32(v); lo[j] = to_tf32(v - acc[j]);
          }
        }
        // row-major planes: swizzled staging tile -> one TMA store per plane (rows / columns past M, N are clipped)
        if (elected) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");   // previous store has read the staging tiles
        asm volatile("bar.sync 1, 128;" ::: "memory");
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          const uint32_t o = uint32_t((k ^ (row & 7)) * 16);
          sts128(st0 + o, acc[4 * k], acc[4 * k + 1], acc[4 * k + 2], acc[4 * k + 3]);
          if (EPI != EPI_PLAIN) sts128(st1 + o, lo[4 * k], lo[4 * k + 1], lo[4 * k + 2], lo[4 * k + 3]);
        }
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        asm volatile("bar.sync 1, 128;" ::: "memory");
        if (elected) {
          tma_store_2d(&P.o0, stage_out, nb, tc.m0);
          if (EPI != EPI_PLAIN) tma_store_2d(&P.o1, stage_out + C::STORE_BYTES, nb, tc.m0);
          asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        }
      }
      MT_ACC(7, ts);                                                                               // 7: store phase
    }
    if (tid == 64) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
    if (tid == 64) { MT_FLUSH(5); MT_FLUSH(6); MT_FLUSH(7); }
  }
  asm volatile("tcgen05.fence::before_thread_sync;");
  __syncthreads();
  if (tid == 0) { MT_ACC(0, t_kernel); MT_FLUSH(0); }                                                                             // 0: whole kernel
  if (warp == 1) {
    asm volatile("tcgen05.fence::after_thread_sync;");
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" :: "r"(tmem), "n"(C::TMEM_COLS));
  }
}

#ifdef DDP_MLP_TIMING
}  // namespace mlp
}  // namespace ddp
extern "C" int ddp_debug_mlp_timing(unsigned long long* out, int reset) {
  if (reset) { static unsigned long long z[160 * 16] = {0}; return (int)cudaMemcpyToSymbol(ddp::mlp::g_mlp_dbg, z, sizeof(z)); }
  return (int)cudaMemcpyFromSymbol(out, ddp::mlp::g_mlp_dbg, sizeof(unsigned long long) * 160 * 16);
}
namespace ddp {
namespace mlp {
#endif

// ---- host side -------------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  }
  return fn;
}
// K-major operand / row-major output: 2-D map over a rows x cols matrix (cols contiguous, row pitch ld), box {32, box_rows}
static int encode_map(CUtensorMap* map, const float* base, int rows, int cols, int box_rows, int ld) {
  EncodeTiledFn fn = encode_fn();
  if (!fn) { set_error("cuTensorMapEncodeTiled is not available from this driver"); return DDP_ENOSYS; }
  if (!aligned16(base) || (ld & 3) || ld < cols) { set_error("GEMM operands and outputs need 16-byte aligned rows (row pitch %% 4 == 0)"); return DDP_EINVAL; }
  cuuint64_t dims[2] = {cuuint64_t(cols), cuuint64_t(rows)};
  cuuint64_t strides[1] = {cuuint64_t(ld) * 4};
  cuuint32_t box[2] = {cuuint32_t(BK), cuuint32_t(box_rows)};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float*>(base), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { set_error("cuTensorMapEncodeTiled failed (CUresult %d) for a %d x %d matrix", int(r), rows, cols); return DDP_EINVAL; }
  return 0;
}
// MN-major operand: the matrix is K x rows (rows contiguous, pitch ld).  3-D view (row % 32, k, row / 32), box {32, 32, blocks}:
// the tile lands as [row-block][k][32 rows].  A ragged last row-block reads past the row end (into the next row / the
// 256-byte plane padding): those operand rows only feed output rows/columns >= M/N, which the TMA store clips.
static int encode_map_mn(CUtensorMap* map, const float* base, int rows, int K, int box_rows, int ld) {
  EncodeTiledFn fn = encode_fn();
  if (!fn) { set_error("cuTensorMapEncodeTiled is not available from this driver"); return DDP_ENOSYS; }
  if (!aligned16(base) || (ld & 3) || ld < rows) { set_error("GEMM operands need 16-byte aligned rows (row pitch %% 4 == 0)"); return DDP_EINVAL; }
  cuuint64_t dims[3] = {32, cuuint64_t(K), cuuint64_t((rows + 31) / 32)};
  cuuint64_t strides[2] = {cuuint64_t(ld) * 4, 128};
  cuuint32_t box[3] = {32, cuuint32_t(BK), cuuint32_t(box_rows / 32)};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, const_cast<float*>(base), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { set_error("cuTensorMapEncodeTiled (MN-major) failed (CUresult %d) for a %d x %d matrix", int(r), K, rows); return DDP_EINVAL; }
  return 0;
}

int make_operand_maps(GemmProblem* p, const float* a_hi, const float* a_lo, bool a_mn, int lda,
                      const float* b_hi, const float* b_lo, bool b_mn, int ldb, int M, int N, int K) {
  constexpr int BN = 128;
  if (a_mn) {
    if (int rc = encode_map_mn(&p->a_hi, a_hi, M, K, BM, lda)) return rc;
    if (int rc = encode_map_mn(&p->a_lo, a_lo, M, K, BM, lda)) return rc;
  } else {
    if (int rc = encode_map(&p->a_hi, a_hi, M, K, BM, lda)) return rc;
    if (int rc = encode_map(&p->a_lo, a_lo, M, K, BM, lda)) return rc;
  }
  if (b_mn) {
    if (int rc = encode_map_mn(&p->b_hi, b_hi, N, K, BN, ldb)) return rc;
    if (int rc = encode_map_mn(&p->b_lo, b_lo, N, K, BN, ldb)) return rc;
  } else {
    if (int rc = encode_map(&p->b_hi, b_hi, N, K, BN, ldb)) return rc;
    if (int rc = encode_map(&p->b_lo, b_lo, N, K, BN, ldb)) return rc;
  }
  p->M = M; p->N = N; p->K = K;
  p->mtiles = (M + BM - 1) / BM;
  p->ntiles = (N + BN - 1) / BN;
  return 0;
}

int make_output_maps(GemmProblem* p, float* out0, float* out1, int M, int N, int ld) {
  if (int rc = encode_map(&p->o0, out0, M, N, BM, ld)) return rc;
  if (out1) return encode_map(&p->o1, out1, M, N, BM, ld);
  return 0;
}

template <int EPI, bool AMN, bool BMN>
static int launch_one(const GemmBatch& batch, cudaStream_t stream) {
  constexpr int BN = 128;
  static bool attr_done = false;
  if (!attr_done) {
    DDP_CUDA(cudaFuncSetAttribute(gemm3x_kernel<BN, EPI, AMN, BMN>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg<BN>::SMEM));
    attr_done = true;
  }
  const int total = batch.tile_start[batch.count];
  const int sms = sm_count();
  const int grid = total < sms ? total : sms;
  DDP_CUDA(launch_pdl(gemm3x_kernel<BN, EPI, AMN, BMN>, dim3(grid), dim3(THREADS), Cfg<BN>::SMEM, stream, true, batch));
  DDP_CUDA(cudaGetLastError());
  return 0;
}

int launch_gemm(const GemmBatch& batch, int epilogue, bool a_mn, bool b_mn, cudaStream_t stream) {
  if (batch.count < 1 || batch.count > MAX_PROBLEMS || batch.tile_start[batch.count] < 1) {
    set_error("launch_gemm: empty or oversized batch");
    return DDP_EINVAL;
  }
  if (epilogue == EPI_PLAIN) {
    if (!a_mn && !b_mn) return launch_one<EPI_PLAIN, false, false>(batch, stream);
    if (!a_mn && b_mn) return launch_one<EPI_PLAIN, false, true>(batch, stream);
    if (a_mn && !b_mn) return launch_one<EPI_PLAIN, true, false>(batch, stream);
    return launch_one<EPI_PLAIN, true, true>(batch, stream);
  }
  if (epilogue == EPI_BIAS_RELU_SPLIT && !a_mn && !b_mn) return launch_one<EPI_BIAS_RELU_SPLIT, false, false>(batch, stream);
  if (epilogue == EPI_MASK_SPLIT && !a_mn && b_mn) return launch_one<EPI_MASK_SPLIT, false, true>(batch, stream);
  set_error("launch_gemm: unsupported epilogue %d / operand layout (%d, %d)", epilogue, int(a_mn), int(b_mn));
  return DDP_EINVAL;
}

}  // namespace mlp
}  // namespace ddp
