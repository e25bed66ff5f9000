// 3xTF32 GEMM on the 5th-generation tensor cores for the MLP of the DDP script
// (ref: pytorch_elastic/mnist_ddp_elastic.py:133-159, Model(hidden_layers=5, features=1024)).
//
//   C[M][N] = A[M][K] . B[N][K]^T          (both operands K-contiguous: nn.Linear's x @ W^T is this shape as stored)
// fp32 results within the 1e-5 parity bar from tf32 tensor cores: every operand is kept as a pair of tf32-exact
// planes (hi = rna_tf32(v), lo = rna_tf32(v - hi)) and each k-step issues three tcgen05.mma.kind::tf32 into one
// fp32 TMEM accumulator:  lo*hi + hi*lo + hi*hi   (lo*lo, 2^-22 relative, is dropped).
// The tensor core adds into its fp32 accumulator with TRUNCATION, a bias that grows linearly with the number of
// accumulation steps (measured: 4e-3 absolute at K = 4096 on unit-variance data, 30x worse than an fp32 dot
// product).  So an accumulator only ever sees KCHUNK k-blocks (128 k values): the epilogue warps drain it into
// round-to-nearest fp32 running sums held in registers, alternating between the two TMEM buffers.
//
// One persistent CTA per SM, 192 threads, warp-specialised:
//   warp 0    producer: one lane issues four TMA tile loads per k-block (cp.async.bulk.tensor.2d, 128-byte swizzle,
//             out-of-range rows / k zero-filled by the TMA unit) into a STAGES-deep shared-memory ring
//   warp 1    one lane issues the 12 MMAs of a k-block (4 k-steps of 8 x 3 products) and tcgen05.commit's the
//             stage back to the producer; owns the TMEM allocation (two accumulator buffers of BN columns)
//   warps 2-5 epilogue: tcgen05.ld every finished 128 x BN chunk accumulator (one row per thread) and add it to the
//             running sums; after the last chunk apply bias+ReLU or the ReLU mask, split into hi/lo and store
//             row-major AND transposed (the transposed copy is what the weight-gradient GEMM consumes as its
//             K-contiguous operand) -- overlapped with the following MMAs through the second accumulator buffer.
// Several problems (the six weight-gradient GEMMs of a step) can share one launch: tiles are numbered across
// problems and dealt round-robin to the CTAs.
#include "mlp_gemm.cuh"

namespace ddp {
namespace mlp {

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
  static constexpr int BAR_BYTES = 256;
  static constexpr int SMEM = STAGES * STAGE_BYTES + BAR_BYTES + 1024;   // + slack to align the ring to 1024
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
// K-major operand tile written by TMA with 128-byte swizzle: rows of 128 B, 8-row groups 1024 B apart
__device__ __forceinline__ uint64_t sw128_desc(uint32_t saddr) {
  return (uint64_t)((saddr & 0x3FFFFu) >> 4) | ((uint64_t)1 << 16) | ((uint64_t)(1024 >> 4) << 32) |
         ((uint64_t)1 << 46) | ((uint64_t)2 << 61);
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

template <int BN, int EPI>
__global__ void __launch_bounds__(THREADS, 1)
gemm3x_kernel(const __grid_constant__ GemmBatch batch) {
  using C = Cfg<BN>;
  extern __shared__ unsigned char smem_raw[];
  const uint32_t raw = s32(smem_raw);
  const uint32_t ring = (raw + 1023u) & ~1023u;                 // 128-byte swizzle needs 1024-byte aligned tiles
  unsigned char* gen_ring = smem_raw + (ring - raw);
  const uint32_t bars = ring + C::STAGES * C::STAGE_BYTES;      // full[STAGES], empty[STAGES], tfull[2], tempty[2], tmem ptr
  auto full = [&](int s) { return bars + 8u * s; };
  auto empty = [&](int s) { return bars + 8u * (C::STAGES + s); };
  auto tfull = [&](int b) { return bars + 8u * (2 * C::STAGES + b); };
  auto tempty = [&](int b) { return bars + 8u * (2 * C::STAGES + 2 + b); };
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(gen_ring + C::STAGES * C::STAGE_BYTES + 8 * (2 * C::STAGES + 4));

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
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
          if (it >= C::STAGES) bar_wait(empty(slot), ((it / C::STAGES) - 1) & 1);
          const uint32_t st = ring + slot * C::STAGE_BYTES;
          bar_expect_tx(full(slot), C::STAGE_BYTES);
          tma_load_2d(st, &P.a_hi, kb * BK, tc.m0, full(slot));
          tma_load_2d(st + A_BYTES, &P.a_lo, kb * BK, tc.m0, full(slot));
          tma_load_2d(st + 2 * A_BYTES, &P.b_hi, kb * BK, tc.n0, full(slot));
          tma_load_2d(st + 2 * A_BYTES + C::B_BYTES, &P.b_lo, kb * BK, tc.n0, full(slot));
        }
      }
    }
  } else if (warp == 1) {
    // ======================= MMA issuer =======================
    if (lane == 0) {
      const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | (uint32_t(BN >> 3) << 17) | (uint32_t(BM >> 4) << 24);
      uint32_t it = 0, sub = 0;                                  // running k-block / chunk counters
      for (int t = blockIdx.x; t < total; t += gridDim.x) {
        const TileCoord tc = locate<BN>(batch, t);
        const int kblocks = (batch.p[tc.prob].K + BK - 1) / BK;
        for (int kb0 = 0; kb0 < kblocks; kb0 += KCHUNK, ++sub) {
          const int buf = sub & 1;
          if (sub >= 2) bar_wait(tempty(buf), ((sub >> 1) - 1) & 1);
          asm volatile("tcgen05.fence::after_thread_sync;");
          const uint32_t d = tmem + uint32_t(buf * BN);
          const int kend = kb0 + KCHUNK < kblocks ? kb0 + KCHUNK : kblocks;
          for (int kb = kb0; kb < kend; ++kb, ++it) {
            const int slot = it % C::STAGES;
            bar_wait(full(slot), (it / C::STAGES) & 1);
            asm volatile("tcgen05.fence::after_thread_sync;");
            const uint32_t st = ring + slot * C::STAGE_BYTES;
            const uint64_t a_hi = sw128_desc(st), a_lo = sw128_desc(st + A_BYTES);
            const uint64_t b_hi = sw128_desc(st + 2 * A_BYTES), b_lo = sw128_desc(st + 2 * A_BYTES + C::B_BYTES);
#pragma unroll
            for (int ks = 0; ks < BK / 8; ++ks) {
              const uint64_t o = uint64_t(ks * 2);               // 8 tf32 = 32 bytes, in 16-byte units
              umma(d, a_lo + o, b_hi + o, idesc, (kb > kb0 || ks) ? 1u : 0u);
              umma(d, a_hi + o, b_lo + o, idesc, 1u);
              umma(d, a_hi + o, b_hi + o, idesc, 1u);
            }
            umma_commit_to(empty(slot));
          }
          umma_commit_to(tfull(buf));
        }
      }
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
        bar_wait(tfull(buf), (sub >> 1) & 1);
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
      }
      const int m = tc.m0 + q * 32 + lane;
      const bool mok = m < P.M;
#pragma unroll
      for (int c = 0; c < BN / 32; ++c) {
        const int nb = tc.n0 + c * 32;
        if (nb >= P.N) continue;                                // warp-uniform
        const bool full_chunk = (nb + 32 <= P.N) && ((P.ldo & 3) == 0);
        float* acc = sum + c * 32;
        if (EPI == EPI_PLAIN) {
          if (mok) {
            float* o = P.out + size_t(m) * P.ldo + nb;
            if (full_chunk) {
#pragma unroll
              for (int j = 0; j < 32; j += 4)
                *reinterpret_cast<float4*>(o + j) = make_float4(acc[j], acc[j + 1], acc[j + 2], acc[j + 3]);
            } else {
#pragma unroll
              for (int j = 0; j < 32; ++j)
                if (nb + j < P.N) o[j] = acc[j];
            }
          }
        } else {
          float hi[32], lo[32];
          if (EPI == EPI_BIAS_RELU_SPLIT) {
#pragma unroll
            for (int j = 0; j < 32; ++j) {
              const float b = (nb + j < P.N) ? __ldg(P.bias + nb + j) : 0.f;
              const float v = fmaxf(acc[j] + b, 0.f);
              hi[j] = to_tf32(v); lo[j] = to_tf32(v - hi[j]);
            }
          } else {
            const float* mk = P.mask + size_t(mok ? m : 0) * P.ldo + nb;
#pragma unroll
            for (int j = 0; j < 32; ++j) {
              const float g = (nb + j < P.N) ? __ldg(mk + j) : 0.f;
              const float v = g > 0.f ? acc[j] : 0.f;
              hi[j] = to_tf32(v); lo[j] = to_tf32(v - hi[j]);
            }
          }
          if (mok) {
            float* oh = P.out_hi + size_t(m) * P.ldo + nb;
            float* ol = P.out_lo + size_t(m) * P.ldo + nb;
            if (full_chunk) {
#pragma unroll
              for (int j = 0; j < 32; j += 4) {
                *reinterpret_cast<float4*>(oh + j) = make_float4(hi[j], hi[j + 1], hi[j + 2], hi[j + 3]);
                *reinterpret_cast<float4*>(ol + j) = make_float4(lo[j], lo[j + 1], lo[j + 2], lo[j + 3]);
              }
            } else {
#pragma unroll
              for (int j = 0; j < 32; ++j)
                if (nb + j < P.N) { oh[j] = hi[j]; ol[j] = lo[j]; }
            }
            // transposed copies: lanes are consecutive m -> each store instruction writes 128 contiguous bytes
#pragma unroll
            for (int j = 0; j < 32; ++j)
              if (nb + j < P.N) {
                P.outT_hi[size_t(nb + j) * P.ldt + m] = hi[j];
                P.outT_lo[size_t(nb + j) * P.ldt + m] = lo[j];
              }
          }
        }
      }
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;");
  __syncthreads();
  if (warp == 1) {
    asm volatile("tcgen05.fence::after_thread_sync;");
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" :: "r"(tmem), "n"(C::TMEM_COLS));
  }
}

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
static int encode_map(CUtensorMap* map, const float* base, int rows, int K, int box_rows) {
  EncodeTiledFn fn = encode_fn();
  if (!fn) { set_error("cuTensorMapEncodeTiled is not available from this driver"); return DDP_ENOSYS; }
  if (!aligned16(base) || (K & 3)) { set_error("GEMM operands need 16-byte aligned rows (K %% 4 == 0)"); return DDP_EINVAL; }
  cuuint64_t dims[2] = {cuuint64_t(K), cuuint64_t(rows)};
  cuuint64_t strides[1] = {cuuint64_t(K) * 4};
  cuuint32_t box[2] = {cuuint32_t(BK), cuuint32_t(box_rows)};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float*>(base), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { set_error("cuTensorMapEncodeTiled failed (CUresult %d) for a %d x %d operand", int(r), rows, K); return DDP_EINVAL; }
  return 0;
}

int make_operand_maps(GemmProblem* p, const float* a_hi, const float* a_lo, const float* b_hi, const float* b_lo,
                      int M, int N, int K, int bn) {
  if (int rc = encode_map(&p->a_hi, a_hi, M, K, BM)) return rc;
  if (int rc = encode_map(&p->a_lo, a_lo, M, K, BM)) return rc;
  if (int rc = encode_map(&p->b_hi, b_hi, N, K, bn)) return rc;
  if (int rc = encode_map(&p->b_lo, b_lo, N, K, bn)) return rc;
  p->M = M; p->N = N; p->K = K;
  p->mtiles = (M + BM - 1) / BM;
  p->ntiles = (N + bn - 1) / bn;
  return 0;
}

template <int BN, int EPI>
static int launch_one(const GemmBatch& batch, cudaStream_t stream) {
  static bool attr_done = false;
  if (!attr_done) {
    DDP_CUDA(cudaFuncSetAttribute(gemm3x_kernel<BN, EPI>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg<BN>::SMEM));
    attr_done = true;
  }
  const int total = batch.tile_start[batch.count];
  const int sms = sm_count();
  const int grid = total < sms ? total : sms;
  gemm3x_kernel<BN, EPI><<<grid, THREADS, Cfg<BN>::SMEM, stream>>>(batch);
  DDP_CUDA(cudaGetLastError());
  return 0;
}

int launch_gemm(const GemmBatch& batch, int epilogue, int bn, cudaStream_t stream) {
  if (batch.count < 1 || batch.count > MAX_PROBLEMS || batch.tile_start[batch.count] < 1) {
    set_error("launch_gemm: empty or oversized batch");
    return DDP_EINVAL;
  }
  if (bn == 128) {
    if (epilogue == EPI_PLAIN) return launch_one<128, EPI_PLAIN>(batch, stream);
    if (epilogue == EPI_BIAS_RELU_SPLIT) return launch_one<128, EPI_BIAS_RELU_SPLIT>(batch, stream);
    if (epilogue == EPI_MASK_SPLIT) return launch_one<128, EPI_MASK_SPLIT>(batch, stream);
  }
  set_error("launch_gemm: unsupported tile width %d / epilogue %d", bn, epilogue);
  return DDP_EINVAL;
}

}  // namespace mlp
}  // namespace ddp
