// Fused (1/N)*grad + optimizer update, in place, one pass over HBM.
// Restates torch's single-tensor rules (torch: optim/sgd.py:343-380, optim/adam.py:414-547,
// optim/adadelta.py) for the optimizers the reference scripts build:
//   SGD lr=.01            ref: horovod/mnist_horovod.py:50
//   AdamW lr=.01/sqrt(N)  ref: horovod/horovod_mnist_elastic.py:41
//   Adam lr=1e-3          ref: pytorch_elastic/mnist_ddp_elastic.py:173
// plus SGD-momentum and Adadelta, which BASELINE.json's north_star names.
// HBM-bound: bytes/element = 12 (SGD), 20 (SGD-momentum), 28 (Adam/AdamW/Adadelta) with f32 grads.
#include <cuda_bf16.h>
#include <math.h>
#include "common.cuh"
#include "update_rules.cuh"

namespace ddp {

__device__ unsigned int g_update_ticket = 0;   // last-CTA detection for bump_step_dev

template <int OPT, bool HAS1, bool HAS2, typename GT>
__global__ void __launch_bounds__(256)
fused_update_kernel(const GT* __restrict__ grad, float* __restrict__ params, float* __restrict__ state1,
                    float* __restrict__ state2, int64_t n, UpdateConsts kc, int64_t step,
                    const int64_t* __restrict__ step_dev, float* __restrict__ grad_wb) {
  __shared__ UpdateConsts ks;
  pdl_wait();                               // launched as a programmatic dependent of the gradient reduction
  if (threadIdx.x == 0) {
    UpdateConsts k = kc;
    int64_t st = step_dev ? (*step_dev + 1) : step;
    finish_consts(k, OPT, st);
    ks = k;
  }
  __syncthreads();
  const UpdateConsts k = ks;
  const int64_t nvec = n >> 2;
  const int64_t stride = int64_t(gridDim.x) * blockDim.x;
  for (int64_t i = int64_t(blockIdx.x) * blockDim.x + threadIdx.x; i < nvec; i += stride) {
    float g[4];
    if (sizeof(GT) == 4) {
      const float* gp = reinterpret_cast<const float*>(grad) + i * 4;
      float4 gv = grad_wb ? ld4(gp) : ld4_stream(gp);  // .nc only when the bucket is read-only
      g[0] = gv.x; g[1] = gv.y; g[2] = gv.z; g[3] = gv.w;
    } else {
      uint2 raw = *reinterpret_cast<const uint2*>(reinterpret_cast<const __nv_bfloat16*>(grad) + i * 4);
      __nv_bfloat162 lo = *reinterpret_cast<__nv_bfloat162*>(&raw.x);
      __nv_bfloat162 hi = *reinterpret_cast<__nv_bfloat162*>(&raw.y);
      g[0] = __low2float(lo); g[1] = __high2float(lo); g[2] = __low2float(hi); g[3] = __high2float(hi);
    }
    float4 pv = ld4(params + i * 4);
    float4 s1v = HAS1 ? ld4(state1 + i * 4) : make_float4(0.f, 0.f, 0.f, 0.f);
    float4 s2v = HAS2 ? ld4(state2 + i * 4) : make_float4(0.f, 0.f, 0.f, 0.f);
    float p[4] = {pv.x, pv.y, pv.z, pv.w};
    float s1[4] = {s1v.x, s1v.y, s1v.z, s1v.w};
    float s2[4] = {s2v.x, s2v.y, s2v.z, s2v.w};
    float go[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) update_one<OPT>(g[j], p[j], s1[j], s2[j], go[j], k);
    st4(params + i * 4, make_float4(p[0], p[1], p[2], p[3]));
    if (HAS1) st4(state1 + i * 4, make_float4(s1[0], s1[1], s1[2], s1[3]));
    if (HAS2) st4(state2 + i * 4, make_float4(s2[0], s2[1], s2[2], s2[3]));
    if (grad_wb) st4(grad_wb + i * 4, make_float4(go[0], go[1], go[2], go[3]));
  }
  // scalar tail (n % 4 elements), handled by block 0
  if (blockIdx.x == 0 && threadIdx.x < (n & 3)) {
    const int64_t i = (nvec << 2) + threadIdx.x;
    float g = sizeof(GT) == 4 ? reinterpret_cast<const float*>(grad)[i]
                              : __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(grad)[i]);
    float p = params[i], s1 = HAS1 ? state1[i] : 0.f, s2 = HAS2 ? state2[i] : 0.f, go;
    update_one<OPT>(g, p, s1, s2, go, k);
    params[i] = p;
    if (HAS1) state1[i] = s1;
    if (HAS2) state2[i] = s2;
    if (grad_wb) grad_wb[i] = go;
  }
  if (k.bump_step && step_dev) {            // every CTA read *step_dev at its start, before taking a ticket
    __syncthreads();
    if (threadIdx.x == 0) {
      __threadfence();
      if (atomicAdd(&g_update_ticket, 1u) == gridDim.x - 1) {
        g_update_ticket = 0u;
        *const_cast<int64_t*>(step_dev) += 1;
      }
    }
  }
}

__global__ void step_increment_kernel(int64_t* step) { *step += 1; }

template <typename T>
__global__ void __launch_bounds__(256) scale_kernel(T* buf, int64_t n, float s) {
  const int64_t stride = int64_t(gridDim.x) * blockDim.x;
  for (int64_t i = int64_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride) {
    if (sizeof(T) == 4) reinterpret_cast<float*>(buf)[i] *= s;
    else {
      __nv_bfloat16* b = reinterpret_cast<__nv_bfloat16*>(buf);
      b[i] = __float2bfloat16(__bfloat162float(b[i]) * s);
    }
  }
}
__global__ void __launch_bounds__(256) scale4_kernel(float4* buf, int64_t nvec, float s) {
  const int64_t stride = int64_t(gridDim.x) * blockDim.x;
  for (int64_t i = int64_t(blockIdx.x) * blockDim.x + threadIdx.x; i < nvec; i += stride) {
    float4 v = buf[i];
    v.x *= s; v.y *= s; v.z *= s; v.w *= s;
    buf[i] = v;
  }
}

template <int OPT, bool HAS1, bool HAS2>
static void launch_update(int grad_dtype, int grid, cudaStream_t stream, const void* grad, float* params,
                          float* s1, float* s2, int64_t n, const UpdateConsts& k, int64_t step,
                          const int64_t* step_dev, float* wb) {
  if (grad_dtype == 0)
    launch_pdl(fused_update_kernel<OPT, HAS1, HAS2, float>, dim3(grid), dim3(256), 0, stream, true,
               static_cast<const float*>(grad), params, s1, s2, n, k, step, step_dev, wb);
  else
    launch_pdl(fused_update_kernel<OPT, HAS1, HAS2, __nv_bfloat16>, dim3(grid), dim3(256), 0, stream, true,
               static_cast<const __nv_bfloat16*>(grad), params, s1, s2, n, k, step, step_dev, wb);
}

}  // namespace ddp

using namespace ddp;

extern "C" int ddp_fused_update(int opt, const void* grad, int grad_dtype, float* params, float* state1,
                                float* state2, int64_t n, float inv_world, const ddp_hyper_t* h,
                                int64_t step, const int64_t* step_dev, void* stream_) {
  DDP_REQUIRE(grad && params && h, "grad, params and hyper are required");
  DDP_REQUIRE(n >= 0, "negative element count");
  DDP_REQUIRE(grad_dtype == 0 || grad_dtype == 1, "grad_dtype must be 0 (f32) or 1 (bf16)");
  DDP_REQUIRE(aligned16(params) && (grad_dtype == 0 ? aligned16(grad) : (reinterpret_cast<uintptr_t>(grad) & 7) == 0),
              "grad/params must be 16-byte aligned (8 for bf16 grads)");
  DDP_REQUIRE(step >= 1 || step_dev, "step must be >= 1 (1-based) unless step_dev is given");
  const bool need1 = (opt == DDP_OPT_SGD) ? (h->momentum != 0.0) : true;
  const bool need2 = (opt != DDP_OPT_SGD);
  DDP_REQUIRE(!need1 || (state1 && aligned16(state1)), "state1 required (16-byte aligned) for this optimizer");
  DDP_REQUIRE(!need2 || (state2 && aligned16(state2)), "state2 required (16-byte aligned) for this optimizer");
  DDP_REQUIRE(!(h->write_back_grad && grad_dtype != 0), "write_back_grad needs f32 gradients");
  if (n == 0) return 0;
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  UpdateConsts k = make_update_consts(h, inv_world);
  float* wb = h->write_back_grad ? const_cast<float*>(static_cast<const float*>(grad)) : nullptr;
  int64_t nvec = n >> 2;
  int64_t want = (nvec + 255) / 256;
  int cap = sm_count() * 8;
  int grid = int(want < 1 ? 1 : (want > cap ? cap : want));
  switch (opt) {
    case DDP_OPT_SGD:
      if (need1) launch_update<DDP_OPT_SGD, true, false>(grad_dtype, grid, stream, grad, params, state1, nullptr, n, k, step, step_dev, wb);
      else launch_update<DDP_OPT_SGD, false, false>(grad_dtype, grid, stream, grad, params, nullptr, nullptr, n, k, step, step_dev, wb);
      break;
    case DDP_OPT_ADAM:
      launch_update<DDP_OPT_ADAM, true, true>(grad_dtype, grid, stream, grad, params, state1, state2, n, k, step, step_dev, wb);
      break;
    case DDP_OPT_ADAMW:
      launch_update<DDP_OPT_ADAMW, true, true>(grad_dtype, grid, stream, grad, params, state1, state2, n, k, step, step_dev, wb);
      break;
    case DDP_OPT_ADADELTA:
      launch_update<DDP_OPT_ADADELTA, true, true>(grad_dtype, grid, stream, grad, params, state1, state2, n, k, step, step_dev, wb);
      break;
    default:
      set_error("ddp_fused_update: unknown optimizer kind %d", opt);
      return DDP_EINVAL;
  }
  DDP_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int ddp_step_increment(int64_t* step_dev, void* stream_) {
  DDP_REQUIRE(step_dev, "step_dev is NULL");
  step_increment_kernel<<<1, 1, 0, static_cast<cudaStream_t>(stream_)>>>(step_dev);
  DDP_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int ddp_scale(void* buf, int dtype, int64_t n, float s, void* stream_) {
  DDP_REQUIRE(buf || n == 0, "buf is NULL");
  DDP_REQUIRE(dtype == 0 || dtype == 1, "dtype must be 0 (f32) or 1 (bf16)");
  if (n <= 0) return 0;
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  int cap = sm_count() * 8;
  if (dtype == 0 && aligned16(buf) && (n & 3) == 0) {
    int64_t nvec = n >> 2;
    int64_t want = (nvec + 255) / 256;
    scale4_kernel<<<int(want > cap ? cap : want), 256, 0, stream>>>(static_cast<float4*>(buf), nvec, s);
  } else {
    int64_t want = (n + 255) / 256;
    int grid = int(want > cap ? cap : want);
    if (dtype == 0) scale_kernel<float><<<grid, 256, 0, stream>>>(static_cast<float*>(buf), n, s);
    else scale_kernel<__nv_bfloat16><<<grid, 256, 0, stream>>>(static_cast<__nv_bfloat16*>(buf), n, s);
  }
  DDP_CUDA(cudaGetLastError());
  return 0;
}
