// The training step's head in ONE kernel, and the entry point that runs forward + backward back to back:
//   fc1 + bias -> ReLU -> dropout -> fc2 + bias -> log_softmax -> NLL mean (+ argmax, correct count)      K_F3
//   ... and, while fc1.weight and the a2 tile are still in shared memory ...
//   dL/dz (log_softmax + NLL backward) -> fc2 grads -> dropout / ReLU backward -> fc1 grads -> da2          K_B3
// ref: horovod/mnist_horovod.py:21-25,62-63 (forward, F.nll_loss, loss.backward()).
// In the reference loop the loss is always F.nll_loss(output, target) and backward follows forward immediately, so the
// head's backward needs nothing from the host.  The separate kernels (net_fwd.cu / net_bwd.cu) each stage the 64 KB
// fc1.weight and the a2 tile, write and re-read h / logp, and pay a launch; here that happens once.  The separate
// kernels stay the path of the module API (autograd hands an arbitrary dlogp to backward).
// Arithmetic (operation order) is that of head_fwd_kernel / head_bwd_kernel, so both paths agree to the last bits
// except for the fc1 weight-gradient partial sums (different tile size => different partial rows).
#include "common.cuh"

namespace ddp {

constexpr int FT = 16;                 // samples per tile
constexpr int FH_WARPS = 13;           // one warp per fc1 output quad (52 >= 50)
constexpr int FH_THREADS = FH_WARPS * 32;
constexpr int FH_A2S = 324;            // a2 row stride (floats)
constexpr int FH_HS = 53;              // h row stride
constexpr int FH_DPS = 20;             // dp row stride ([j][s], 16 used; float4-aligned)
constexpr int FH_FLOATS = F1 * A2_PER + FT * FH_A2S + 52 + FT * FH_HS + 50 * 12 + 12 + FT * 12 + F1 * FH_DPS + FT * 12;
constexpr int FH_SMEM = FH_FLOATS * 4 + 16;
__global__ void __launch_bounds__(FH_THREADS, 2)
head_fused_kernel(const float* __restrict__ params, const float* __restrict__ a2,
                  const int64_t* __restrict__ target, const float* __restrict__ mask1d,
                  uint64_t seed, uint64_t offset, const int64_t* __restrict__ offset_dev,
                  float nll_coef, float* __restrict__ h_out, float* __restrict__ logp_out,
                  float* __restrict__ loss_out, int64_t* __restrict__ argmax_out,
                  unsigned long long* __restrict__ correct_out, float* __restrict__ loss_part,
                  unsigned int* __restrict__ counters, float* __restrict__ da2,
                  float* __restrict__ partials, int B) {
  extern __shared__ __align__(128) float smem[];
  float* w1_s = smem;                    // [50][320] natural layout (bulk copy)
  float* a2_s = w1_s + F1 * A2_PER;      // [16][324]
  float* b1_s = a2_s + FT * FH_A2S;      // [52]
  float* h_s = b1_s + 52;                // [16][53]
  float* w2_s = h_s + FT * FH_HS;        // [50][12]   fc2.weight transposed: [k][c]
  float* b2_s = w2_s + 50 * 12;          // [12]
  float* z_s = b2_s + 12;                // [16][12]   logits, then log-probabilities
  float* dp_s = z_s + FT * 12;           // [50][20]   dL/d(fc1 pre-activation), transposed
  float* dz_s = dp_s + F1 * FH_DPS;      // [16][12]   dL/d(logits)
  uint64_t* bar = reinterpret_cast<uint64_t*>(dz_s + FT * 12);
  if (offset_dev) offset += uint64_t(*offset_dev);
  const int tid = threadIdx.x;
  const int ntiles = (B + FT - 1) / FT;
  if (tid == 0) mbarrier_init(bar, 1);
  __syncthreads();
  auto issue_tile = [&](int tile, bool with_w1) {       // thread 0 only
    const int b0 = tile * FT;
    const int nvalid = min(FT, B - b0);
    mbarrier_expect_tx(bar, uint32_t(nvalid) * (A2_PER * 4) + (with_w1 ? F1 * A2_PER * 4 : 0));
    if (with_w1) {
      bulk_copy_g2s(w1_s, params + OFF_F1W, F1 * A2_PER * 4, bar);   // parameters: may be fetched while conv2_fwd still runs
      pdl_wait();                                                    // a2 is conv2_fwd's output
    }
    for (int s = 0; s < nvalid; ++s)
      bulk_copy_g2s(a2_s + s * FH_A2S, a2 + size_t(b0 + s) * A2_PER, A2_PER * 4, bar);
  };
  pdl_trigger();
  if (tid == 0 && int(blockIdx.x) < ntiles) issue_tile(blockIdx.x, true);
  if (tid < 52) b1_s[tid] = tid < F1 ? params[OFF_F1B + tid] : 0.f;
  for (int i = tid; i < 50 * 12; i += FH_THREADS) {
    int k = i / 12, c = i % 12;
    w2_s[i] = c < NC ? params[OFF_F2W + c * F1 + k] : 0.f;
  }
  if (tid < 12) b2_s[tid] = tid < NC ? params[OFF_F2B + tid] : 0.f;

  // forward mapping: warp = 4 fc1 outputs, lane = (k-quarter, sample pair)
  const int jg = tid >> 5, lane = tid & 31, kp = lane >> 3, sg = lane & 7;
  int jrow[4];
#pragma unroll
  for (int jj = 0; jj < 4; ++jj) jrow[jj] = min(jg * 4 + jj, F1 - 1) * A2_PER + kp * 80;   // clamp: cols 50,51 discarded
  // backward mappings
  const int jgW = tid / 40, igW = tid % 40;     // fc1 wgrad: rows jgW*5.., cols igW*8..   (tid < 400)
  const int sgD = tid / 80, iqD = tid % 80;     // fc1 dgrad: samples sgD*4.., cols iqD*4.. (tid < 320)
  float accW1[5][8];
#pragma unroll
  for (int a = 0; a < 5; ++a)
#pragma unroll
    for (int c = 0; c < 8; ++c) accW1[a][c] = 0.f;
  float accW2[2] = {0.f, 0.f};                  // fc2.weight elements tid, tid + 416
  float accB = 0.f;                             // tid < 50: fc1.bias[tid]; 64 <= tid < 74: fc2.bias[tid-64]
  float loss_acc = 0.f;                         // thread 0 only
  int corr_acc = 0;
  uint32_t phase = 0;
  pdl_wait();                                   // every thread: outputs may only be written after this
  for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const int b0 = tile * FT;
    const int nvalid = min(FT, B - b0);
    for (int i = tid; i < (FT - nvalid) * A2_PER; i += FH_THREADS)    // rows past the batch are not copied: keep them zero
      a2_s[(nvalid + i / A2_PER) * FH_A2S + i % A2_PER] = 0.f;
    mbarrier_wait(bar, phase);
    phase ^= 1u;
    __syncthreads();
    // ---------------- forward: fc1 (split-K over the lanes) + bias + ReLU + dropout ----------------
    {
      float acc[2][4];
#pragma unroll
      for (int i = 0; i < 2; ++i)
#pragma unroll
        for (int jj = 0; jj < 4; ++jj) acc[i][jj] = 0.f;
      const float* ar0 = a2_s + sg * FH_A2S + kp * 80;
      const float* ar1 = a2_s + (sg + 8) * FH_A2S + kp * 80;
#pragma unroll 5
      for (int k = 0; k < 80; k += 4) {
        float4 x0 = *reinterpret_cast<const float4*>(ar0 + k);
        float4 x1 = *reinterpret_cast<const float4*>(ar1 + k);
#pragma unroll
        for (int jj = 0; jj < 4; ++jj) {
          float4 w = *reinterpret_cast<const float4*>(w1_s + jrow[jj] + k);
          acc[0][jj] = fmaf(x0.x, w.x, acc[0][jj]); acc[0][jj] = fmaf(x0.y, w.y, acc[0][jj]);
          acc[0][jj] = fmaf(x0.z, w.z, acc[0][jj]); acc[0][jj] = fmaf(x0.w, w.w, acc[0][jj]);
          acc[1][jj] = fmaf(x1.x, w.x, acc[1][jj]); acc[1][jj] = fmaf(x1.y, w.y, acc[1][jj]);
          acc[1][jj] = fmaf(x1.z, w.z, acc[1][jj]); acc[1][jj] = fmaf(x1.w, w.w, acc[1][jj]);
        }
      }
#pragma unroll
      for (int i = 0; i < 2; ++i)
#pragma unroll
        for (int jj = 0; jj < 4; ++jj) {
          float v = acc[i][jj];
          v += __shfl_xor_sync(0xffffffffu, v, 8);
          v += __shfl_xor_sync(0xffffffffu, v, 16);
          acc[i][jj] = v;
        }
      if (kp == 0) {
#pragma unroll
        for (int i = 0; i < 2; ++i) {
          const int s = sg + 8 * i;
          const int b = b0 + s;
#pragma unroll
          for (int jj = 0; jj < 4; ++jj) {
            const int j = jg * 4 + jj;
            if (j < F1) {
              float v = fmaxf(acc[i][jj] + b1_s[j], 0.f);
              if (s < nvalid) {
                float keep = mask1d ? mask1d[size_t(b) * F1 + j]
                                    : philox_keep(seed, offset, 2u, uint64_t(b) * F1 + j);
                v = v * (keep * 2.f);
              }
              h_s[s * FH_HS + j] = v;
              if (s < nvalid) h_out[size_t(b) * F1 + j] = v;
            }
          }
        }
      }
    }
    __syncthreads();
    // ---------------- forward: fc2, thread (s, class pair) ----------------
    if (tid < FT * 5) {
      const int s = tid / 5, c2 = (tid % 5) * 2;
      float z0 = b2_s[c2], z1 = b2_s[c2 + 1];
      const float* hr = h_s + s * FH_HS;
#pragma unroll 10
      for (int k = 0; k < F1; ++k) {
        float hv = hr[k];
        float2 w = *reinterpret_cast<const float2*>(w2_s + k * 12 + c2);
        z0 = fmaf(hv, w.x, z0); z1 = fmaf(hv, w.y, z1);
      }
      z_s[s * 12 + c2] = z0; z_s[s * 12 + c2 + 1] = z1;
    }
    __syncthreads();
    // ---------------- log_softmax, NLL, argmax; and dL/dz ----------------
    if (tid < 32) {                        // warp 0: lanes 0..15 hold the tile's samples
      const int s = tid;
      float nll = 0.f; int ok = 0;
      if (s < FT) {
        float dz[NC];
#pragma unroll
        for (int c = 0; c < NC; ++c) dz[c] = 0.f;
        if (s < nvalid) {
          float z[NC];
#pragma unroll
          for (int c = 0; c < NC; ++c) z[c] = z_s[s * 12 + c];
          float m = z[0];
#pragma unroll
          for (int c = 1; c < NC; ++c) m = fmaxf(m, z[c]);
          float sum = 0.f;
#pragma unroll
          for (int c = 0; c < NC; ++c) sum += expf(z[c] - m);
          const float lse = logf(sum);
          float lp[NC];
          int am = 0;
#pragma unroll
          for (int c = 0; c < NC; ++c) {
            lp[c] = z[c] - m - lse;
            logp_out[size_t(b0 + s) * NC + c] = lp[c];
          }
          float best = lp[0];
#pragma unroll
          for (int c = 1; c < NC; ++c) if (lp[c] > best) { best = lp[c]; am = c; }
          if (argmax_out) argmax_out[b0 + s] = am;
          const int t = int(target[b0 + s]);
          float picked = 0.f;
#pragma unroll
          for (int c = 0; c < NC; ++c) if (c == t) picked = lp[c];
          nll = -picked;
          ok = (am == t);
          // log_softmax + NLL backward, as head_bwd_kernel computes it: dz = g - exp(logp) * sum(g), g = nll_coef * onehot
#pragma unroll
          for (int c = 0; c < NC; ++c) dz[c] = ((c == t) ? nll_coef : 0.f) - expf(lp[c]) * nll_coef;
        }
#pragma unroll
        for (int c = 0; c < NC; ++c) dz_s[s * 12 + c] = dz[c];
      }
      float v = nll;                       // fixed-order shuffle tree
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
      int okc = __popc(__ballot_sync(0xffffffffu, ok));
      if (tid == 0) { loss_acc += v; corr_acc += okc; }
    }
    __syncthreads();
    // ---------------- backward: dp[s][j] = relu'/dropout' * sum_c dz[s][c] * W2[c][j] ----------------
    for (int e = tid; e < FT * F1; e += FH_THREADS) {
      const int s = e / F1, j = e % F1;
      float dh = 0.f;
#pragma unroll
      for (int c = 0; c < NC; ++c) dh = fmaf(dz_s[s * 12 + c], w2_s[j * 12 + c], dh);
      dp_s[j * FH_DPS + s] = h_s[s * FH_HS + j] > 0.f ? dh * 2.f : 0.f;     // training mode: dropout scale 2
    }
    __syncthreads();
    // fc2 weight / bias gradients and fc1 bias gradient
#pragma unroll
    for (int q = 0; q < 2; ++q) {
      const int e = tid + q * FH_THREADS;
      if (e < NC * F1) {
        const int c = e / F1, j = e % F1;
        float a = accW2[q];
        for (int s = 0; s < FT; ++s) a = fmaf(dz_s[s * 12 + c], h_s[s * FH_HS + j], a);
        accW2[q] = a;
      }
    }
    if (tid < F1) {
      float a = accB;
      for (int s = 0; s < FT; ++s) a += dp_s[tid * FH_DPS + s];
      accB = a;
    } else if (tid >= 64 && tid < 64 + NC) {
      float a = accB;
      for (int s = 0; s < FT; ++s) a += dz_s[s * 12 + (tid - 64)];
      accB = a;
    }
    // fc1 weight gradient: dW1[j][i] += dp[s][j] * a2[s][i]
    if (tid < 400) {
      const float* ar = a2_s + igW * 8;
      const float* dr = dp_s + (jgW * 5) * FH_DPS;
#pragma unroll 4
      for (int s = 0; s < FT; ++s) {
        const float4 v0 = *reinterpret_cast<const float4*>(ar + s * FH_A2S);
        const float4 v1 = *reinterpret_cast<const float4*>(ar + s * FH_A2S + 4);
        const float av[8] = {v0.x, v0.y, v0.z, v0.w, v1.x, v1.y, v1.z, v1.w};
#pragma unroll
        for (int a = 0; a < 5; ++a) {
          const float d = dr[a * FH_DPS + s];
#pragma unroll
          for (int c = 0; c < 8; ++c) accW1[a][c] = fmaf(d, av[c], accW1[a][c]);
        }
      }
    }
    // fc1 data gradient: da2[s][i] = sum_j dp[s][j] * W1[j][i]
    if (tid < 320) {
      float acc[4][4];
#pragma unroll
      for (int a = 0; a < 4; ++a)
#pragma unroll
        for (int c = 0; c < 4; ++c) acc[a][c] = 0.f;
#pragma unroll 5
      for (int j = 0; j < F1; ++j) {
        const float4 d = *reinterpret_cast<const float4*>(dp_s + j * FH_DPS + sgD * 4);
        const float4 w = *reinterpret_cast<const float4*>(w1_s + j * A2_PER + iqD * 4);
        const float dv[4] = {d.x, d.y, d.z, d.w};
        const float wv[4] = {w.x, w.y, w.z, w.w};
#pragma unroll
        for (int a = 0; a < 4; ++a)
#pragma unroll
          for (int c = 0; c < 4; ++c) acc[a][c] = fmaf(dv[a], wv[c], acc[a][c]);
      }
#pragma unroll
      for (int a = 0; a < 4; ++a) {
        const int s = sgD * 4 + a;
        if (s < nvalid)
          *reinterpret_cast<float4*>(da2 + size_t(b0 + s) * A2_PER + iqD * 4) = make_float4(acc[a][0], acc[a][1], acc[a][2], acc[a][3]);
      }
    }
    __syncthreads();                        // the tile (a2_s, h_s, dp_s, dz_s) is consumed
    if (tid == 0 && tile + int(gridDim.x) < ntiles) issue_tile(tile + gridDim.x, false);
  }
  // one partial gradient row per CTA
  float* part = partials + size_t(blockIdx.x) * P_ARENA;
  if (tid < 400) {
#pragma unroll
    for (int a = 0; a < 5; ++a) {
      float* o = part + OFF_F1W + (jgW * 5 + a) * A2_PER + igW * 8;        // OFF_F1W and 8*igW are multiples of 4
      *reinterpret_cast<float4*>(o) = make_float4(accW1[a][0], accW1[a][1], accW1[a][2], accW1[a][3]);
      *reinterpret_cast<float4*>(o + 4) = make_float4(accW1[a][4], accW1[a][5], accW1[a][6], accW1[a][7]);
    }
  }
#pragma unroll
  for (int q = 0; q < 2; ++q) {
    const int e = tid + q * FH_THREADS;
    if (e < NC * F1) part[OFF_F2W + e] = accW2[q];
  }
  if (tid < F1) part[OFF_F1B + tid] = accB;
  else if (tid >= 64 && tid < 64 + NC) part[OFF_F2B + tid - 64] = accB;
  // loss: one partial per CTA, the last CTA to arrive adds them in a fixed order
  __shared__ unsigned int ticket_s;
  if (tid == 0) {
    if (correct_out && corr_acc) atomicAdd(correct_out, (unsigned long long)corr_acc);
    loss_part[blockIdx.x] = loss_acc;
    __threadfence();
    ticket_s = atomicAdd(&counters[0], 1u) + 1u;
  }
  __syncthreads();
  if (ticket_s == gridDim.x && tid < 32) {
    __threadfence();
    float v = 0.f;
    for (unsigned int i = tid; i < gridDim.x; i += 32) v += __ldcg(loss_part + i);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if (tid == 0) {
      loss_out[0] = v / float(B);
      counters[0] = 0u;  // self-reset for the next launch
    }
  }
}

}  // namespace ddp

using namespace ddp;

extern "C" int ddp_net_train_fwdbwd(const float* params, const float* x, const int64_t* target,
                                    const float* mask2d, const float* mask1d, uint64_t seed, uint64_t offset,
                                    const int64_t* offset_dev, float loss_scale, int64_t batch, void* workspace,
                                    float* out_logp, float* out_loss, int64_t* out_argmax, int64_t* out_correct,
                                    void* scratch, float* grad_bucket, void* stream_) {
  DDP_REQUIRE(params && x && target && workspace && out_logp && out_loss && scratch && grad_bucket,
              "params, x, target, workspace, out_logp, out_loss, scratch and grad_bucket are required");
  DDP_REQUIRE(batch >= 1 && batch <= (int64_t(1) << 24), "batch out of range");
  DDP_REQUIRE(aligned16(params) && aligned16(x) && aligned16(workspace) && aligned16(scratch),
              "params/x/workspace/scratch must be 16-byte aligned");
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  const int B = int(batch);
  Workspace ws = carve_workspace(workspace, batch);
  Scratch sc = carve_scratch(scratch, batch);
  if (int rc = launch_fwd_convs(params, x, mask2d, seed, offset, offset_dev, 1, ws, B, stream)) return rc;
  {
    static bool attr_done[64] = {false};
    int dev = 0;
    DDP_CUDA(cudaGetDevice(&dev));
    if (dev >= 0 && dev < 64 && !attr_done[dev]) {
      DDP_CUDA(cudaFuncSetAttribute(head_fused_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, FH_SMEM));
      attr_done[dev] = true;
    }
  }
  const int ntiles = (B + FT - 1) / FT;
  int grid = sm_count() * 2;
  if (grid > ntiles) grid = ntiles;
  if (grid > MAX_PARTS) grid = MAX_PARTS;
  DDP_CUDA(launch_pdl(head_fused_kernel, dim3(grid), dim3(FH_THREADS), FH_SMEM, stream, true, params, ws.a2, target, mask1d,
                      seed, offset, offset_dev, -loss_scale / float(B), ws.h, out_logp, out_loss, out_argmax,
                      reinterpret_cast<unsigned long long*>(out_correct), sc.loss_part, sc.counters, sc.da2, sc.partials, B));
  prof_mark(3, stream);
  prof_mark(4, stream);
  prof_mark(5, stream);
  return launch_bwd_after_head(params, x, 1, B, ws, sc, grid, grad_bucket, stream);
}

// The whole data-parallel training step of one rank in ONE call: forward, loss, backward, and -- instead of writing the
// gradients into a bucket for a separate allreduce + update -- the partial gradient sums go straight into the NVLink
// peer-memory exchange (peer.cu), the fc / conv2 part while conv1's weight gradient is still being computed.
// ref: horovod/mnist_horovod.py:59-64 (the five-line loop incl. hvd.DistributedOptimizer.step()).
extern "C" int ddp_net_train_step_peer(float* params, const float* x, const int64_t* target, const float* mask2d,
                                       const float* mask1d, uint64_t seed, uint64_t offset, const int64_t* offset_dev,
                                       int64_t batch, void* workspace, float* out_logp, float* out_loss, void* scratch,
                                       float* grad_bucket, void* peer, int opt, float* state1, float* state2,
                                       const ddp_hyper_t* hyper, int64_t step, const int64_t* step_dev, int fused_head,
                                       void* stream_) {
  DDP_REQUIRE(params && x && target && workspace && out_logp && out_loss && scratch && peer && hyper,
              "params, x, target, workspace, out_logp, out_loss, scratch, peer and hyper are required");
  DDP_REQUIRE(batch >= 1 && batch <= (int64_t(1) << 24), "batch out of range");
  DDP_REQUIRE(aligned16(params) && aligned16(x) && aligned16(workspace) && aligned16(scratch),
              "params/x/workspace/scratch must be 16-byte aligned");
  DDP_REQUIRE(!hyper->write_back_grad || grad_bucket, "grad_bucket is required when hyper.write_back_grad is set");
  DDP_REQUIRE(step >= 1 || step_dev, "step must be >= 1 unless step_dev is given");
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  const int B = int(batch);
  Workspace ws = carve_workspace(workspace, batch);
  Scratch sc = carve_scratch(scratch, batch);
  if (int rc = launch_fwd_convs(params, x, mask2d, seed, offset, offset_dev, 1, ws, B, stream)) return rc;
  int parts_head = 0;
  const float nll_coef = -1.0f / float(B);
  if (fused_head) {
    static bool attr_done[64] = {false};
    int dev = 0;
    DDP_CUDA(cudaGetDevice(&dev));
    if (dev >= 0 && dev < 64 && !attr_done[dev]) {
      DDP_CUDA(cudaFuncSetAttribute(head_fused_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, FH_SMEM));
      attr_done[dev] = true;
    }
    const int ntiles = (B + FT - 1) / FT;
    int grid = sm_count() * 2;
    if (grid > ntiles) grid = ntiles;
    if (grid > MAX_PARTS) grid = MAX_PARTS;
    DDP_CUDA(launch_pdl(head_fused_kernel, dim3(grid), dim3(FH_THREADS), FH_SMEM, stream, true, params, ws.a2, target, mask1d,
                        seed, offset, offset_dev, nll_coef, ws.h, out_logp, out_loss, (int64_t*)nullptr,
                        (unsigned long long*)nullptr, sc.loss_part, sc.counters, sc.da2, sc.partials, B));
    parts_head = grid;
  } else {
    if (int rc = launch_head_fwd(params, ws, target, mask1d, seed, offset, offset_dev, 1, out_logp, out_loss, nullptr, nullptr,
                                 sc, B, stream)) return rc;
    if (int rc = launch_head_bwd(params, ws, out_logp, nullptr, target, nll_coef, 2.f, sc, B, &parts_head, stream)) return rc;
  }
  const PeerExchange px{peer, opt, params, state1, state2, hyper, step, step_dev};
  return launch_bwd_after_head(params, x, 1, B, ws, sc, parts_head, grad_bucket, stream, &px);
}
