// Optimizer update rules shared by the fused update kernel (update.cu) and the peer-memory
// allreduce+update kernel (peer.cu).  Restates torch's single-tensor rules
// (torch: optim/sgd.py:343-380, optim/adam.py:414-547, optim/adadelta.py).
#pragma once
#include <math.h>
#include "common.cuh"

namespace ddp {

struct UpdateConsts {
  float inv_world, lr, momentum, one_minus_damp, wd;
  float beta1, beta2, one_minus_beta1, one_minus_beta2, eps, rho, one_minus_rho;
  float step_size, bc2_sqrt;   // Adam: lr/(1-beta1^t), sqrt(1-beta2^t)
  float decay;                 // AdamW: 1 - lr*wd
  double lr_d, beta1_d, beta2_d;
  int nesterov, first_step, write_back, bump_step;
};

template <int OPT>
__device__ __forceinline__ void update_one(float g_in, float& p, float& s1, float& s2, float& g_out,
                                           const UpdateConsts& k) {
  float g = g_in * k.inv_world;
  g_out = g;
  if (OPT == DDP_OPT_SGD) {
    if (k.wd != 0.f) g = fmaf(p, k.wd, g);
    if (k.momentum != 0.f) {
      float buf = k.first_step ? g : fmaf(g, k.one_minus_damp, s1 * k.momentum);
      s1 = buf;
      g = k.nesterov ? fmaf(buf, k.momentum, g) : buf;
    }
    p = fmaf(g, -k.lr, p);
  } else if (OPT == DDP_OPT_ADAM || OPT == DDP_OPT_ADAMW) {
    if (OPT == DDP_OPT_ADAMW) p = p * k.decay;
    else if (k.wd != 0.f) g = fmaf(p, k.wd, g);
    float m = fmaf(g - s1, k.one_minus_beta1, s1);           // exp_avg.lerp_(g, 1-beta1)
    float v = fmaf(k.one_minus_beta2 * g, g, s2 * k.beta2);  // mul_(beta2).addcmul_(g,g,1-beta2)
    s1 = m; s2 = v;
    float denom = sqrtf(v) / k.bc2_sqrt + k.eps;
    p = fmaf(m / denom, -k.step_size, p);                    // addcdiv_(m, denom, -step_size)
  } else {  // Adadelta
    if (k.wd != 0.f) g = fmaf(p, k.wd, g);
    float sq = fmaf(k.one_minus_rho * g, g, s1 * k.rho);
    float stdv = sqrtf(sq + k.eps);
    float delta = sqrtf(s2 + k.eps) / stdv * g;
    s1 = sq;
    s2 = fmaf(k.one_minus_rho * delta, delta, s2 * k.rho);
    p = fmaf(delta, -k.lr, p);
  }
}

__device__ __forceinline__ float4 ld4(const float* p) { return *reinterpret_cast<const float4*>(p); }
__device__ __forceinline__ void st4(float* p, float4 v) { *reinterpret_cast<float4*>(p) = v; }
__device__ __forceinline__ float4 ld4_stream(const float* p) {
  float4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p));
  return r;
}

__device__ __forceinline__ void finish_consts(UpdateConsts& k, int opt, int64_t step) {
  k.first_step = (step <= 1);
  if (opt == DDP_OPT_ADAM || opt == DDP_OPT_ADAMW) {
    // torch computes these as Python doubles: 1 - beta**step, lr / bc1, bc2 ** 0.5
    double bc1 = 1.0 - pow(k.beta1_d, double(step));
    double bc2 = 1.0 - pow(k.beta2_d, double(step));
    k.step_size = float(k.lr_d / bc1);
    k.bc2_sqrt = float(sqrt(bc2));
  }
}


// host side: ddp_hyper_t (doubles) -> UpdateConsts (derived constants formed in double, rounded once)
inline UpdateConsts make_update_consts(const ddp_hyper_t* h, float inv_world) {
  UpdateConsts k{};
  k.inv_world = inv_world; k.lr = float(h->lr); k.momentum = float(h->momentum);
  k.one_minus_damp = float(1.0 - h->dampening);
  k.wd = float(h->weight_decay); k.beta1 = float(h->beta1); k.beta2 = float(h->beta2);
  k.one_minus_beta1 = float(1.0 - h->beta1); k.one_minus_beta2 = float(1.0 - h->beta2);
  k.eps = float(h->eps); k.rho = float(h->rho); k.one_minus_rho = float(1.0 - h->rho);
  k.decay = float(1.0 - h->lr * h->weight_decay);
  k.lr_d = h->lr; k.beta1_d = h->beta1; k.beta2_d = h->beta2;
  k.nesterov = h->nesterov; k.write_back = h->write_back_grad; k.bump_step = h->bump_step_dev;
  return k;
}

}  // namespace ddp
