// Interface of the tcgen05 3xTF32 GEMM used by the MLP path (mlp_gemm.cu <-> mlp.cu).
#pragma once
#include <cuda.h>
#include "common.cuh"

namespace ddp {
namespace mlp {

enum Epilogue { EPI_PLAIN = 0, EPI_BIAS_RELU_SPLIT = 1, EPI_MASK_SPLIT = 2 };

// One C[M][N] = A[M][K] . B[N][K]^T problem; A and B are given as tf32 hi/lo pairs (K contiguous).
struct GemmProblem {
  CUtensorMap a_hi, a_lo, b_hi, b_lo;   // 2-D fp32 maps, box {32 (k), 128 | BN rows}, 128-byte swizzle
  int M, N, K;
  int mtiles, ntiles;
  // EPI_PLAIN: out[m*ldo + n] = acc
  float* out;
  // EPI_BIAS_RELU_SPLIT: v = relu(acc + bias[n]);  EPI_MASK_SPLIT: v = mask[m*ldo + n] > 0 ? acc : 0
  // both write v as tf32 hi/lo, row-major [M][ldo] and transposed [N][ldt]
  const float* bias;
  const float* mask;
  float *out_hi, *out_lo, *outT_hi, *outT_lo;
  int ldo, ldt;
};
constexpr int MAX_PROBLEMS = 6;
struct GemmBatch {
  GemmProblem p[MAX_PROBLEMS];
  int count;
  int tile_start[MAX_PROBLEMS + 1];
};

// Fill the four tensor maps of `p` (row-major operands, leading dimension = K elements, 16-byte aligned rows).
int make_operand_maps(GemmProblem* p, const float* a_hi, const float* a_lo, const float* b_hi, const float* b_lo,
                      int M, int N, int K, int bn);
// Launch all problems of `batch` in one persistent grid; bn = 128 or 256 (the N tile).
int launch_gemm(const GemmBatch& batch, int epilogue, int bn, cudaStream_t stream);

}  // namespace mlp
}  // namespace ddp
