// Interface of the tcgen05 3xTF32 GEMM used by the MLP path (mlp_gemm.cu <-> mlp.cu).
#pragma once
#include <cuda.h>
#include "common.cuh"

namespace ddp {
namespace mlp {

enum Epilogue { EPI_PLAIN = 0, EPI_BIAS_RELU_SPLIT = 1, EPI_MASK_SPLIT = 2 };

// One C[M][N] = A . B^T problem.  Each operand is a pair of tf32 hi/lo planes stored either "K-major"
// (A as [M][K], B as [N][K], K contiguous) or "MN-major" (A as [K][M], B as [K][N]): the tensor core reads both,
// so a row-major activation / gradient / weight plane serves every GEMM of the step without a transposed copy.
struct GemmProblem {
  CUtensorMap a_hi, a_lo, b_hi, b_lo;   // fp32 tiles, 128-byte swizzle (2-D maps for K-major, 3-D for MN-major operands)
  CUtensorMap o0, o1;                   // row-major outputs (TMA store, box {32, 128}): out | out_hi, out_lo
  int M, N, K;
  int mtiles, ntiles;
  // EPI_PLAIN: out[m][n] = acc (through o0)
  // EPI_BIAS_RELU_SPLIT: v = relu(acc + bias[n]), and bit (n % 32) of bits_out[m*bits_ld + n/32] = (v > 0)
  // EPI_MASK_SPLIT:      v = that bit of bits_in ? acc : 0        (the ReLU mask of the forward activation)
  // both write v as tf32 hi/lo planes through o0/o1
  const float* bias;
  const uint32_t* bits_in;
  uint32_t* bits_out;
  int bits_ld;
};
constexpr int MAX_PROBLEMS = 6;
struct GemmBatch {
  GemmProblem p[MAX_PROBLEMS];
  int count;
  int tile_start[MAX_PROBLEMS + 1];
};

// Fill the operand maps of `p`.  K-major operand: rows x K matrix, row pitch ld floats.  MN-major operand: K x rows matrix,
// row pitch ld floats (rows = M for A, N for B).  ld % 4 == 0, 16-byte aligned bases.
int make_operand_maps(GemmProblem* p, const float* a_hi, const float* a_lo, bool a_mn, int lda,
                      const float* b_hi, const float* b_lo, bool b_mn, int ldb, int M, int N, int K);
// Row-major output maps: one plane (EPI_PLAIN, out1 = nullptr) or the hi/lo pair; ld = row pitch in floats.
int make_output_maps(GemmProblem* p, float* out0, float* out1, int M, int N, int ld);
// Launch all problems of `batch` (same epilogue and operand majors) in one persistent grid.
int launch_gemm(const GemmBatch& batch, int epilogue, bool a_mn, bool b_mn, cudaStream_t stream);

}  // namespace mlp
}  // namespace ddp
