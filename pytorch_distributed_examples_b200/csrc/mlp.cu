// Native forward/backward of the DDP script's own model:
//   Model(hidden_layers=H, features=F): Linear(in,F) -> ReLU -> H x [Linear(F,F) -> ReLU] -> Linear(F,out)
//   (ref: pytorch_elastic/mnist_ddp_elastic.py:133-159; built with H=5, F=1024 at :172) + nn.CrossEntropyLoss (:174).
// The H+1 wide layers run on the tensor cores (mlp_gemm.cu, 3xTF32); the 10-wide classifier, the loss and the
// bias gradients are small FFMA / reduction kernels here.  Data layout in the workspace: every activation X[l], every
// gradient D[l] and every weight matrix is ONE pair of row-major tf32 hi/lo planes; the GEMM reads them K-major or
// MN-major as each product needs, so nothing is ever transposed.  The ReLU masks travel as one bit per element.
// Weights are split into their planes at the start of every forward.
//   forward : X[l+1] = relu(X[l] W_l^T + b_l)                      GEMM M=B N=F K=in_l   epilogue bias+ReLU+split
//   head    : z = X[G] W_G^T + b_G ; loss = mean_b(lse(z_b) - z_b[y_b])
//   backward: D[G] = (dz W_G) * (X[G] > 0)
//             D[l] = (D[l+1] W_l) * (X[l] > 0)      l = G-1..1      GEMM M=B N=F K=F      epilogue mask+split
//             dW_l = D[l+1]^T X[l]                  l = 0..G-1      ONE grouped GEMM launch, M=F N=in_l K=B
//             db_l = sum_b D[l+1]
// All reductions run in a fixed order (no floating-point atomics): reruns and replicas are bit-identical.
#include <math.h>
#include <stdlib.h>
#include "mlp_gemm.cuh"

namespace ddp {
namespace mlp {

constexpr int MAXC = 16;            // classifier width supported by the head kernels
constexpr int HEAD_CHUNK = 32;      // samples per partial sum of the classifier weight gradient

struct Shape { int in, F, H, out, G; };     // G = H + 1 tensor-core layers
static int check_shape(const ddp_mlp_shape_t* s, Shape* o) {
  if (!s) { set_error("mlp: shape is NULL"); return DDP_EINVAL; }
  o->in = s->in_features; o->F = s->features; o->H = s->hidden_layers; o->out = s->out_features; o->G = s->hidden_layers + 1;
  if (o->in < 4 || (o->in & 3) || o->F < 4 || (o->F & 3) || o->H < 0 || o->H + 1 > MAX_PROBLEMS || o->out < 1 || o->out > MAXC) {
    set_error("mlp: need in_features %% 4 == 0, features %% 4 == 0, 0 <= hidden_layers <= %d, 1 <= out_features <= %d",
              MAX_PROBLEMS - 1, MAXC);
    return DDP_EINVAL;
  }
  return 0;
}
static inline int64_t pad4(int64_t v) { return (v + 3) & ~int64_t(3); }
// parameter arena, reference state_dict order: input_layer.{weight,bias}, hidden_layers.i.{weight,bias}, final_layer.{weight,bias}
struct ParamLayout { int64_t w[MAX_PROBLEMS + 1], b[MAX_PROBLEMS + 1], total; };
static ParamLayout param_layout(const Shape& s) {
  ParamLayout L{};
  int64_t off = 0;
  for (int l = 0; l <= s.G; ++l) {
    const int64_t rows = l == s.G ? s.out : s.F, cols = l == 0 ? s.in : s.F;
    L.w[l] = off; off += pad4(rows * cols);
    L.b[l] = off; off += pad4(rows);
  }
  L.total = off;
  return L;
}

struct Planes { float *hi, *lo, *hiT, *loT; };
struct Work {
  Planes W[MAX_PROBLEMS];       // W[l].hi/lo: [F][in_l]
  Planes X[MAX_PROBLEMS + 1];   // X[0]: [B][in]; X[l>=1]: [B][F]
  Planes D[MAX_PROBLEMS + 1];   // D[1..G]: [B][F]
  uint32_t* bits[MAX_PROBLEMS + 1];   // bits[l], l = 1..G: ReLU mask of X[l], [B][ceil(F/32)]
  float* logits;                // [B][MAXC]
  float* dz;                    // [B][MAXC]
  float* loss_part;             // [B]
  float* head_part;             // [B/HEAD_CHUNK][out][F]
  float* bias_part;             // [BIAS_RC][G][F]
};
static int64_t carve(const Shape& s, int64_t B, char* base, Work* w) {
  int64_t off = 0;
  auto take = [&](int64_t floats) { float* p = base ? reinterpret_cast<float*>(base + off) : nullptr; off += align256(floats * 4); return p; };
  auto planes = [&](int64_t floats, bool transposed) {
    Planes p{};
    p.hi = take(floats); p.lo = take(floats);
    if (transposed) { p.hiT = take(floats); p.loT = take(floats); }
    return p;
  };
  Work tmp{};
  for (int l = 0; l < s.G; ++l) tmp.W[l] = planes(int64_t(s.F) * (l == 0 ? s.in : s.F), false);
  for (int l = 0; l <= s.G; ++l) tmp.X[l] = planes(B * (l == 0 ? s.in : s.F), false);
  for (int l = 1; l <= s.G; ++l) tmp.D[l] = planes(B * s.F, false);
  for (int l = 1; l <= s.G; ++l) tmp.bits[l] = reinterpret_cast<uint32_t*>(take(B * ((s.F + 31) / 32)));
  tmp.logits = take(B * MAXC);
  tmp.dz = take(B * MAXC);
  tmp.loss_part = take(B);
  tmp.head_part = take(((B + HEAD_CHUNK - 1) / HEAD_CHUNK) * int64_t(s.out) * s.F);
  tmp.bias_part = take(int64_t(16) * s.G * s.F);
  if (w) *w = tmp;
  return off;
}

__device__ __forceinline__ float tf32_rna(float v) {
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(v));
  return __uint_as_float(r);
}

// ---- src[rows][cols] -> tf32 hi/lo planes, row-major and (optionally) transposed [cols][rows] ----------------
__global__ void __launch_bounds__(256)
split_kernel(const float* __restrict__ src, int rows, int cols, float* __restrict__ hi, float* __restrict__ lo,
             float* __restrict__ hiT, float* __restrict__ loT) {
  pdl_trigger();
  pdl_wait();                                   // nothing of the step is touched before the previous kernel is complete
  __shared__ float th[32][33], tl[32][33];
  const int tx = threadIdx.x, ty = threadIdx.y;
  const int c = blockIdx.x * 32 + tx;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int rr = ty + 8 * i, r = blockIdx.y * 32 + rr;
    float h = 0.f, l = 0.f;
    if (r < rows && c < cols) {
      const float v = src[size_t(r) * cols + c];
      h = tf32_rna(v); l = tf32_rna(v - h);
      hi[size_t(r) * cols + c] = h; lo[size_t(r) * cols + c] = l;
    }
    th[rr][tx] = h; tl[rr][tx] = l;
  }
  if (!hiT) return;
  __syncthreads();
  const int r = blockIdx.y * 32 + tx;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int cc = ty + 8 * i, c2 = blockIdx.x * 32 + cc;
    if (c2 < cols && r < rows) {
      hiT[size_t(c2) * rows + r] = th[tx][cc];
      loT[size_t(c2) * rows + r] = tl[tx][cc];
    }
  }
}
static void launch_split(const float* src, int rows, int cols, const Planes& p, bool transposed, cudaStream_t st) {
  dim3 grid((cols + 31) / 32, (rows + 31) / 32), block(32, 8);
  launch_pdl(split_kernel, dim3(grid), dim3(block), 0, st, true, src, rows, cols, p.hi, p.lo, transposed ? p.hiT : nullptr, transposed ? p.loT : nullptr);
}

// ---- classifier forward: one warp per HEAD_SPW samples (the classifier rows are loaded once for all of them) -----
constexpr int HEAD_SPW = 1;
__global__ void __launch_bounds__(256)
head_fwd_kernel(const float* __restrict__ xh, const float* __restrict__ xl, const float* __restrict__ w, const float* __restrict__ bias,
                const int64_t* __restrict__ target, int B, int F, int out, float* __restrict__ logits_ws,
                float* __restrict__ out_logits, float* __restrict__ loss_part) {
  pdl_trigger();
  pdl_wait();                                   // nothing of the step is touched before the previous kernel is complete
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  const int b0 = warp * HEAD_SPW;
  if (b0 >= B) return;
  float acc[HEAD_SPW][MAXC];
#pragma unroll
  for (int s = 0; s < HEAD_SPW; ++s)
#pragma unroll
    for (int c = 0; c < MAXC; ++c) acc[s][c] = 0.f;
#pragma unroll 4
  for (int k = lane; k < F; k += 32) {
    float v[HEAD_SPW];
#pragma unroll
    for (int s = 0; s < HEAD_SPW; ++s) {
      const size_t o = size_t(min(b0 + s, B - 1)) * F + k;
      v[s] = xh[o] + xl[o];
    }
#pragma unroll
    for (int c = 0; c < MAXC; ++c)
      if (c < out) {
        const float wv = __ldg(w + size_t(c) * F + k);
#pragma unroll
        for (int s = 0; s < HEAD_SPW; ++s) acc[s][c] = fmaf(v[s], wv, acc[s][c]);
      }
  }
#pragma unroll
  for (int s = 0; s < HEAD_SPW; ++s) {
#pragma unroll
    for (int c = 0; c < MAXC; ++c)
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) acc[s][c] += __shfl_xor_sync(0xffffffffu, acc[s][c], o);
    const int b = b0 + s;
    if (b >= B) break;                                         // warp-uniform
    float z = -INFINITY;
#pragma unroll
    for (int c = 0; c < MAXC; ++c)
      if (lane == c && c < out) z = acc[s][c] + bias[c];
    if (lane < MAXC) logits_ws[size_t(b) * MAXC + lane] = lane < out ? z : 0.f;
    if (out_logits && lane < out) out_logits[size_t(b) * out + lane] = z;
    if (target && loss_part) {
      float mx = z;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
      float e = lane < out ? expf(z - mx) : 0.f;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) e += __shfl_xor_sync(0xffffffffu, e, o);
      const int y = int(target[b]);
      const float zy = __shfl_sync(0xffffffffu, z, y);
      if (lane == 0) loss_part[b] = (mx + logf(e)) - zy;
    }
  }
}
// mean of loss_part in a fixed order (one block)
__global__ void __launch_bounds__(1024)
loss_mean_kernel(const float* __restrict__ part, int B, float* __restrict__ out) {
  pdl_trigger();
  pdl_wait();                                   // nothing of the step is touched before the previous kernel is complete
  __shared__ float s[1024];
  float a = 0.f;
  for (int i = threadIdx.x; i < B; i += 1024) a += part[i];
  s[threadIdx.x] = a;
  __syncthreads();
  for (int w = 512; w > 0; w >>= 1) {
    if (threadIdx.x < w) s[threadIdx.x] += s[threadIdx.x + w];
    __syncthreads();
  }
  if (threadIdx.x == 0) out[0] = s[0] / float(B);
}

// ---- classifier backward ---------------------------------------------------------------------------------------
// dz = (softmax(z) - onehot(y)) * scale / B   when the loss is ours;  dz = dlogits * scale   when autograd hands it in
__global__ void __launch_bounds__(256)
head_dz_kernel(const float* __restrict__ logits_ws, const int64_t* __restrict__ target, const float* __restrict__ dlogits,
               float scale, int B, int out, float* __restrict__ dz) {
  pdl_trigger();
  pdl_wait();                                   // nothing of the step is touched before the previous kernel is complete
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= B) return;
  float v[MAXC];
  if (dlogits) {
#pragma unroll
    for (int c = 0; c < MAXC; ++c) v[c] = c < out ? dlogits[size_t(b) * out + c] * scale : 0.f;
  } else {
    float mx = -INFINITY;
#pragma unroll
    for (int c = 0; c < MAXC; ++c) { v[c] = c < out ? logits_ws[size_t(b) * MAXC + c] : -INFINITY; mx = fmaxf(mx, v[c]); }
    float sum = 0.f;
#pragma unroll
    for (int c = 0; c < MAXC; ++c) { v[c] = c < out ? expf(v[c] - mx) : 0.f; sum += v[c]; }
    const int y = int(target[b]);
    const float inv = 1.f / sum, k = scale / float(B);
#pragma unroll
    for (int c = 0; c < MAXC; ++c) v[c] = c < out ? (v[c] * inv - (c == y ? 1.f : 0.f)) * k : 0.f;
  }
#pragma unroll
  for (int c = 0; c < MAXC; c += 4)
    *reinterpret_cast<float4*>(dz + size_t(b) * MAXC + c) = make_float4(v[c], v[c + 1], v[c + 2], v[c + 3]);
}
// D[G][b][k] = (sum_c dz[b][c] W[c][k]) * (X[G][b][k] > 0), written as tf32 hi/lo planes
__global__ void __launch_bounds__(256)
head_dx_kernel(const float* __restrict__ dz, const float* __restrict__ w, const float* __restrict__ xh, int B, int F, int out,
               float* __restrict__ hi, float* __restrict__ lo) {
  pdl_trigger();
  pdl_wait();                                   // nothing of the step is touched before the previous kernel is complete
  __shared__ float dzs[32][MAXC];
  const int tx = threadIdx.x, ty = threadIdx.y, t = ty * 32 + tx;
  const int b0 = blockIdx.y * 32, k = blockIdx.x * 32 + tx;
  for (int i = t; i < 32 * MAXC; i += 256) {
    const int bb = b0 + i / MAXC;
    dzs[i / MAXC][i % MAXC] = bb < B ? dz[size_t(bb) * MAXC + i % MAXC] : 0.f;
  }
  __syncthreads();
  float wv[MAXC];
#pragma unroll
  for (int c = 0; c < MAXC; ++c) wv[c] = (c < out && k < F) ? __ldg(w + size_t(c) * F + k) : 0.f;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int rr = ty + 8 * i, b = b0 + rr;
    if (b < B && k < F) {
      float a = 0.f;
#pragma unroll
      for (int c = 0; c < MAXC; ++c) a = fmaf(dzs[rr][c], wv[c], a);
      const float v = xh[size_t(b) * F + k] > 0.f ? a : 0.f;
      const float h = tf32_rna(v);
      hi[size_t(b) * F + k] = h; lo[size_t(b) * F + k] = tf32_rna(v - h);
    }
  }
}
// partial[chunk][c][k] = sum_{b in chunk} dz[b][c] * X[G][b][k]
__global__ void __launch_bounds__(128)
head_dw_partial_kernel(const float* __restrict__ dz, const float* __restrict__ xh, const float* __restrict__ xl, int B, int F, int out,
                       float* __restrict__ part) {
  pdl_trigger();
  pdl_wait();                                   // nothing of the step is touched before the previous kernel is complete
  __shared__ float dzs[HEAD_CHUNK][MAXC];
  const int k = blockIdx.x * 128 + threadIdx.x, b0 = blockIdx.y * HEAD_CHUNK;
  for (int i = threadIdx.x; i < HEAD_CHUNK * MAXC; i += 128) {
    const int bb = b0 + i / MAXC;
    dzs[i / MAXC][i % MAXC] = bb < B ? dz[size_t(bb) * MAXC + i % MAXC] : 0.f;
  }
  __syncthreads();
  if (k >= F) return;
  float acc[MAXC];
#pragma unroll
  for (int c = 0; c < MAXC; ++c) acc[c] = 0.f;
  const int nb = min(HEAD_CHUNK, B - b0);
#pragma unroll 8
  for (int i = 0; i < nb; ++i) {                               // rows past B have dz = 0 in smem, but are not read
    const size_t o = size_t(b0 + i) * F + k;
    const float v = xh[o] + xl[o];
#pragma unroll
    for (int c = 0; c < MAXC; ++c) acc[c] = fmaf(dzs[i][c], v, acc[c]);
  }
#pragma unroll
  for (int c = 0; c < MAXC; ++c)
    if (c < out) part[(size_t(blockIdx.y) * out + c) * F + k] = acc[c];
}
__global__ void __launch_bounds__(256)
head_dw_reduce_kernel(const float* __restrict__ part, int chunks, int n /* out*F */, float* __restrict__ dw) {
  pdl_trigger();
  pdl_wait();                                   // nothing of the step is touched before the previous kernel is complete
  const int i = blockIdx.x * 256 + threadIdx.x;
  if (i >= n) return;
  float a = 0.f;
  for (int c = 0; c < chunks; ++c) a += part[size_t(c) * n + i];
  dw[i] = a;
}
// db[c] = sum_b dz[b][c]  (one block, fixed order)
__global__ void __launch_bounds__(1024)
head_db_kernel(const float* __restrict__ dz, int B, int out, float* __restrict__ db) {
  pdl_trigger();
  pdl_wait();                                   // nothing of the step is touched before the previous kernel is complete
  __shared__ float s[1024];
  const int c = threadIdx.x % MAXC, lane_b = threadIdx.x / MAXC;        // 64 sample lanes x 16 classes
  float a = 0.f;
  for (int b = lane_b; b < B; b += 1024 / MAXC) a += dz[size_t(b) * MAXC + c];
  s[threadIdx.x] = a;
  __syncthreads();
  for (int w = 512; w >= MAXC; w >>= 1) {
    if (threadIdx.x < w) s[threadIdx.x] += s[threadIdx.x + w];
    __syncthreads();
  }
  if (threadIdx.x < out) db[threadIdx.x] = s[threadIdx.x];
}
// db_l[n] = sum_b (D_hi[b][n] + D_lo[b][n]) in two fixed-order stages.  Stage 1: block (x = 128-column group, y = layer,
// z = row chunk), 8 warps; a lane owns 4 columns (float4 loads, a warp covers 512 contiguous bytes of a row), warp w
// takes rows w, w+8, ... of the chunk, the 8 warp sums are added in warp order -> part[z][layer][n].  Stage 2 adds
// the BIAS_RC chunk sums in order.
constexpr int BIAS_RC = 16;
struct BiasGradArgs { const float* dh[MAX_PROBLEMS]; const float* dl[MAX_PROBLEMS]; float* db[MAX_PROBLEMS]; };
__global__ void __launch_bounds__(256)
bias_grad_partial_kernel(BiasGradArgs args, int F, int B, float* __restrict__ part) {
  pdl_trigger();
  pdl_wait();                                   // nothing of the step is touched before the previous kernel is complete
  __shared__ float4 s[8][32];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n = blockIdx.x * 128 + lane * 4;
  const int rows_per = (B + BIAS_RC - 1) / BIAS_RC;
  const int r0 = blockIdx.z * rows_per, r1 = min(B, r0 + rows_per);
  const float* h = args.dh[blockIdx.y];
  const float* l = args.dl[blockIdx.y];
  float4 a = make_float4(0.f, 0.f, 0.f, 0.f);
  if (n < F) {                                                 // F % 4 == 0: a float4 is all in or all out
#pragma unroll 4
    for (int b = r0 + warp; b < r1; b += 8) {
      const float4 x = *reinterpret_cast<const float4*>(h + size_t(b) * F + n);
      const float4 y = *reinterpret_cast<const float4*>(l + size_t(b) * F + n);
      a.x += x.x + y.x; a.y += x.y + y.y; a.z += x.z + y.z; a.w += x.w + y.w;
    }
  }
  s[warp][lane] = a;
  __syncthreads();
  if (warp == 0 && n < F) {
    float4 t = s[0][lane];
#pragma unroll
    for (int i = 1; i < 8; ++i) { const float4 v = s[i][lane]; t.x += v.x; t.y += v.y; t.z += v.z; t.w += v.w; }
    *reinterpret_cast<float4*>(part + (size_t(blockIdx.z) * gridDim.y + blockIdx.y) * F + n) = t;
  }
}
__global__ void __launch_bounds__(256)
bias_grad_reduce_kernel(BiasGradArgs args, int F, int layers, const float* __restrict__ part) {
  pdl_trigger();
  pdl_wait();                                   // nothing of the step is touched before the previous kernel is complete
  const int n = blockIdx.x * 256 + threadIdx.x;
  if (n >= F) return;
  float t = 0.f;
#pragma unroll
  for (int z = 0; z < BIAS_RC; ++z) t += part[(size_t(z) * layers + blockIdx.y) * F + n];
  args.db[blockIdx.y][n] = t;
}

// ---- all weight matrices -> tf32 hi/lo planes in one launch (blockIdx.z = layer) ------------------------------------
struct SplitArgs { const float* src[MAX_PROBLEMS]; float* hi[MAX_PROBLEMS]; float* lo[MAX_PROBLEMS]; int64_t n[MAX_PROBLEMS]; };
__global__ void __launch_bounds__(256)
split_flat_kernel(SplitArgs args) {
  pdl_trigger();
  pdl_wait();                                   // nothing of the step is touched before the previous kernel is complete
  const int64_t n4 = args.n[blockIdx.z] >> 2;                  // every tensor is a multiple of 4 floats (F % 4 == in % 4 == 0)
  const float4* src = reinterpret_cast<const float4*>(args.src[blockIdx.z]);
  float4* hi = reinterpret_cast<float4*>(args.hi[blockIdx.z]);
  float4* lo = reinterpret_cast<float4*>(args.lo[blockIdx.z]);
  for (int64_t i = int64_t(blockIdx.x) * 256 + threadIdx.x; i < n4; i += int64_t(gridDim.x) * 256) {
    const float4 v = src[i];
    float4 h, l;
    h.x = tf32_rna(v.x); h.y = tf32_rna(v.y); h.z = tf32_rna(v.z); h.w = tf32_rna(v.w);
    l.x = tf32_rna(v.x - h.x); l.y = tf32_rna(v.y - h.y); l.z = tf32_rna(v.z - h.z); l.w = tf32_rna(v.w - h.w);
    hi[i] = h; lo[i] = l;
  }
}

}  // namespace mlp
}  // namespace ddp

using namespace ddp;
using namespace ddp::mlp;

extern "C" int64_t ddp_mlp_param_count(const ddp_mlp_shape_t* shape) {
  Shape s;
  if (check_shape(shape, &s)) return -1;
  return param_layout(s).total;
}

extern "C" int ddp_mlp_param_layout(const ddp_mlp_shape_t* shape, int64_t* offsets, int64_t* sizes, int max_tensors) {
  Shape s;
  if (int rc = check_shape(shape, &s)) return rc;
  const int n = 2 * (s.G + 1);
  DDP_REQUIRE(offsets && sizes && max_tensors >= n, "offsets/sizes must hold 2 * (hidden_layers + 2) entries");
  const ParamLayout L = param_layout(s);
  for (int l = 0; l <= s.G; ++l) {
    const int64_t rows = l == s.G ? s.out : s.F, cols = l == 0 ? s.in : s.F;
    offsets[2 * l] = L.w[l]; sizes[2 * l] = rows * cols;
    offsets[2 * l + 1] = L.b[l]; sizes[2 * l + 1] = rows;
  }
  return n;
}

extern "C" int64_t ddp_mlp_workspace_bytes(const ddp_mlp_shape_t* shape, int64_t batch) {
  Shape s;
  if (check_shape(shape, &s)) return -1;
  if (batch < 4 || (batch & 3)) { set_error("ddp_mlp_workspace_bytes: batch must be a positive multiple of 4"); return -1; }
  return carve(s, batch, nullptr, nullptr);
}

extern "C" int ddp_mlp_forward(const ddp_mlp_shape_t* shape, const float* params, const float* x, const int64_t* target,
                               int64_t batch, void* workspace, float* out_logits, float* out_loss, void* stream_) {
  Shape s;
  if (int rc = check_shape(shape, &s)) return rc;
  DDP_REQUIRE(params && x && workspace, "params, x and workspace are required");
  DDP_REQUIRE(batch >= 4 && (batch & 3) == 0 && batch <= (int64_t(1) << 20), "batch must be a multiple of 4 (<= 2^20)");
  DDP_REQUIRE(aligned16(params) && aligned16(x) && (reinterpret_cast<uintptr_t>(workspace) & 255) == 0,
              "params/x must be 16-byte aligned, workspace 256-byte aligned");
  DDP_REQUIRE(!out_loss || target, "out_loss needs target");
  cudaStream_t st = static_cast<cudaStream_t>(stream_);
  const int B = int(batch);
  Work w;
  carve(s, batch, static_cast<char*>(workspace), &w);
  const ParamLayout L = param_layout(s);
  // operands: weights and the input batch as tf32 hi/lo planes
  {
    SplitArgs sa{};
    for (int l = 0; l < s.G; ++l) {
      sa.src[l] = params + L.w[l]; sa.hi[l] = w.W[l].hi; sa.lo[l] = w.W[l].lo;
      sa.n[l] = int64_t(s.F) * (l == 0 ? s.in : s.F);
    }
    launch_pdl(split_flat_kernel, dim3(dim3(128, 1, s.G)), dim3(256), 0, st, true, sa);
    SplitArgs sx{};
    sx.src[0] = x; sx.hi[0] = w.X[0].hi; sx.lo[0] = w.X[0].lo; sx.n[0] = int64_t(B) * s.in;
    launch_pdl(split_flat_kernel, dim3(dim3(256, 1, 1)), dim3(256), 0, st, true, sx);
  }
  DDP_CUDA(cudaGetLastError());
  for (int l = 0; l < s.G; ++l) {
    GemmBatch gb{};
    GemmProblem& p = gb.p[0];
    const int K = l == 0 ? s.in : s.F;
    if (int rc = make_operand_maps(&p, w.X[l].hi, w.X[l].lo, false, K, w.W[l].hi, w.W[l].lo, false, K, B, s.F, K)) return rc;
    p.bias = params + L.b[l];
    if (int rc = make_output_maps(&p, w.X[l + 1].hi, w.X[l + 1].lo, B, s.F, s.F)) return rc;
    p.bits_out = w.bits[l + 1]; p.bits_ld = (s.F + 31) / 32;
    gb.count = 1; gb.tile_start[0] = 0; gb.tile_start[1] = p.mtiles * p.ntiles;
    if (int rc = launch_gemm(gb, EPI_BIAS_RELU_SPLIT, false, false, st)) return rc;
  }
  const int spb = 8 * HEAD_SPW;                               // samples per 256-thread block
  launch_pdl(head_fwd_kernel, dim3((B + spb - 1) / spb), dim3(256), 0, st, true, w.X[s.G].hi, w.X[s.G].lo, params + L.w[s.G], params + L.b[s.G], target, B, s.F,
                                                         s.out, w.logits, out_logits, out_loss ? w.loss_part : nullptr);
  if (out_loss) launch_pdl(loss_mean_kernel, dim3(1), dim3(1024), 0, st, true, w.loss_part, B, out_loss);
  DDP_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int ddp_mlp_backward(const ddp_mlp_shape_t* shape, const float* params, const float* dlogits, const int64_t* target,
                                float loss_scale, int64_t batch, void* workspace, float* grad_bucket, void* stream_) {
  Shape s;
  if (int rc = check_shape(shape, &s)) return rc;
  DDP_REQUIRE(params && workspace && grad_bucket, "params, workspace and grad_bucket are required");
  DDP_REQUIRE(dlogits || target, "either dlogits (from autograd) or target (fused cross-entropy) is required");
  DDP_REQUIRE(batch >= 4 && (batch & 3) == 0 && batch <= (int64_t(1) << 20), "batch must be a multiple of 4 (<= 2^20)");
  DDP_REQUIRE(aligned16(params) && aligned16(grad_bucket) && (reinterpret_cast<uintptr_t>(workspace) & 255) == 0,
              "params/grad_bucket must be 16-byte aligned, workspace 256-byte aligned");
  cudaStream_t st = static_cast<cudaStream_t>(stream_);
  const int B = int(batch);
  Work w;
  carve(s, batch, static_cast<char*>(workspace), &w);
  const ParamLayout L = param_layout(s);
  const int G = s.G;
  // classifier
  launch_pdl(head_dz_kernel, dim3((B + 255) / 256), dim3(256), 0, st, true, w.logits, target, dlogits, loss_scale, B, s.out, w.dz);
  launch_pdl(head_dx_kernel, dim3(dim3((s.F + 31) / 32, (B + 31) / 32)), dim3(dim3(32, 8)), 0, st, true, w.dz, params + L.w[G], w.X[G].hi, B, s.F, s.out,
                                                                              w.D[G].hi, w.D[G].lo);
  const int chunks = (B + HEAD_CHUNK - 1) / HEAD_CHUNK;
  launch_pdl(head_dw_partial_kernel, dim3(dim3((s.F + 127) / 128, chunks)), dim3(128), 0, st, true, w.dz, w.X[G].hi, w.X[G].lo, B, s.F, s.out, w.head_part);
  launch_pdl(head_dw_reduce_kernel, dim3((s.out * s.F + 255) / 256), dim3(256), 0, st, true, w.head_part, chunks, s.out * s.F, grad_bucket + L.w[G]);
  launch_pdl(head_db_kernel, dim3(1), dim3(1024), 0, st, true, w.dz, B, s.out, grad_bucket + L.b[G]);
  DDP_CUDA(cudaGetLastError());
  // data gradients through the wide layers
  for (int l = G - 1; l >= 1; --l) {
    GemmBatch gb{};
    GemmProblem& p = gb.p[0];
    // D[l] = D[l+1] @ W_l: B operand (n = in, k = out) is W_l's row-major plane read MN-major
    if (int rc = make_operand_maps(&p, w.D[l + 1].hi, w.D[l + 1].lo, false, s.F, w.W[l].hi, w.W[l].lo, true, s.F, B, s.F, s.F)) return rc;
    p.bits_in = w.bits[l]; p.bits_ld = (s.F + 31) / 32;
    if (int rc = make_output_maps(&p, w.D[l].hi, w.D[l].lo, B, s.F, s.F)) return rc;
    gb.count = 1; gb.tile_start[0] = 0; gb.tile_start[1] = p.mtiles * p.ntiles;
    if (int rc = launch_gemm(gb, EPI_MASK_SPLIT, false, true, st)) return rc;
  }
  // all weight gradients in one grouped launch; bias gradients beside it
  {
    GemmBatch gb{};
    gb.count = G;
    gb.tile_start[0] = 0;
    for (int l = 0; l < G; ++l) {
      GemmProblem& p = gb.p[l];
      const int K_l = l == 0 ? s.in : s.F;
      // dW_l = D[l+1]^T @ X[l]: both operands are the row-major planes read MN-major (k = batch)
      if (int rc = make_operand_maps(&p, w.D[l + 1].hi, w.D[l + 1].lo, true, s.F, w.X[l].hi, w.X[l].lo, true, K_l, s.F, K_l, B)) return rc;
      if (int rc = make_output_maps(&p, grad_bucket + L.w[l], nullptr, s.F, K_l, K_l)) return rc;
      gb.tile_start[l + 1] = gb.tile_start[l] + p.mtiles * p.ntiles;
    }
    if (int rc = launch_gemm(gb, EPI_PLAIN, true, true, st)) return rc;
  }
  {
    BiasGradArgs ba{};
    for (int l = 0; l < G; ++l) { ba.dh[l] = w.D[l + 1].hi; ba.dl[l] = w.D[l + 1].lo; ba.db[l] = grad_bucket + L.b[l]; }
    static_assert(BIAS_RC == 16, "carve() reserves 16 row chunks");
    launch_pdl(bias_grad_partial_kernel, dim3(dim3((s.F + 127) / 128, G, BIAS_RC)), dim3(256), 0, st, true, ba, s.F, B, w.bias_part);
    launch_pdl(bias_grad_reduce_kernel, dim3(dim3((s.F + 255) / 256, G)), dim3(256), 0, st, true, ba, s.F, G, w.bias_part);
  }
  DDP_CUDA(cudaGetLastError());
  return 0;
}

// ---- building blocks, exported for tests and reuse ----------------------------------------------------------------
extern "C" int ddp_split_tf32(const float* src, int64_t rows, int64_t cols, float* hi, float* lo, float* hiT, float* loT,
                              void* stream) {
  DDP_REQUIRE(src && hi && lo, "src, hi and lo are required");
  DDP_REQUIRE((hiT == nullptr) == (loT == nullptr), "give both transposed planes or neither");
  DDP_REQUIRE(rows >= 1 && cols >= 1 && rows <= (1 << 24) && cols <= (1 << 24), "rows/cols out of range");
  Planes p{hi, lo, hiT, loT};
  launch_split(src, int(rows), int(cols), p, hiT != nullptr, static_cast<cudaStream_t>(stream));
  DDP_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int ddp_gemm_3xtf32(const float* a_hi, const float* a_lo, int a_mn_major, const float* b_hi, const float* b_lo,
                               int b_mn_major, int M, int N, int K, float* c, int ldc, void* stream) {
  DDP_REQUIRE(a_hi && a_lo && b_hi && b_lo && c, "null operand");
  DDP_REQUIRE(M >= 1 && N >= 1 && K >= 1 && ldc >= N, "need positive sizes and ldc >= N");
  DDP_REQUIRE((ldc & 3) == 0 && aligned16(c), "c must be 16-byte aligned with ldc %% 4 == 0");
  const int lda = a_mn_major ? M : K, ldb = b_mn_major ? N : K;
  DDP_REQUIRE((lda & 3) == 0 && (ldb & 3) == 0, "the contiguous dimension of each operand must be a multiple of 4");
  GemmBatch gb{};
  GemmProblem& p = gb.p[0];
  if (int rc = make_operand_maps(&p, a_hi, a_lo, a_mn_major != 0, lda, b_hi, b_lo, b_mn_major != 0, ldb, M, N, K)) return rc;
  if (int rc = make_output_maps(&p, c, nullptr, M, N, ldc)) return rc;
  gb.count = 1; gb.tile_start[0] = 0; gb.tile_start[1] = p.mtiles * p.ntiles;
  return launch_gemm(gb, EPI_PLAIN, a_mn_major != 0, b_mn_major != 0, static_cast<cudaStream_t>(stream));
}
