// Forward kernels of the MNIST CNN `Net` (ref: horovod/mnist_horovod.py:18-25):
//   K_F1  conv1(1->10,5x5)+bias -> maxpool2 -> ReLU                     x  -> a1,idx1
//   K_F2  conv2(10->20,5x5)+bias -> Dropout2d -> maxpool2 -> ReLU       a1 -> a2,idx2   (tensor cores: conv2_fwd_tc.cu)
//   K_F3  fc1+bias -> ReLU -> dropout -> fc2+bias -> log_softmax (-> NLL mean, argmax)
// K_F1 / K_F3 are fp32 FFMA with register tiling; conv outputs before pooling never leave the SM.
#include <stdlib.h>
#include "common.cuh"

namespace ddp {

// first-max over the 2x2 window in scan order (0,0),(0,1),(1,0),(1,1); matches
// torch max_pool2d's `val > maxval` update rule, so ties keep the earliest index.
__device__ __forceinline__ void pool4(float v0, float v1, float v2, float v3, float& best, int& bi) {
  best = v0; bi = 0;
  if (v1 > best) { best = v1; bi = 1; }
  if (v2 > best) { best = v2; bi = 2; }
  if (v3 > best) { best = v3; bi = 3; }
}

// ---------------------------------------------------------------------------
// K_F1: one thread per pooled output position (144 per sample), all 10 channels:
// 6x6 input patch in registers, 40 accumulators, weights broadcast from smem.
// 288 threads = 2 samples per pass.
// ---------------------------------------------------------------------------
constexpr int F1_THREADS = 288;
__global__ void __launch_bounds__(F1_THREADS, 2)
conv1_fwd_kernel(const float* __restrict__ params, const float* __restrict__ x,
                 float* __restrict__ a1, uint8_t* __restrict__ idx1, int B) {
  __shared__ __align__(16) float w_s[25 * 12];   // [tap][c padded to 12]
  __shared__ float b_s[12];
  __shared__ __align__(128) float img_s[2][2 * IMG_PIX];   // double-buffered pair of images (bulk-copied)
  __shared__ __align__(8) uint64_t bar[2];
  const int tid = threadIdx.x;
  pdl_wait();                                              // first kernel of the step: the optimizer update before it must be complete
  pdl_trigger();                                           // ... and only then may conv2_fwd start building its weight operand
  const int npairs = (B + 1) >> 1;
  auto issue = [&](int pair, int buf) {                    // thread 0: prefetch a pair of images
    const int b0 = pair * 2;
    const uint32_t bytes = uint32_t(min(2, B - b0)) * (IMG_PIX * 4);
    mbarrier_expect_tx(&bar[buf], bytes);
    bulk_copy_g2s(img_s[buf], x + size_t(b0) * IMG_PIX, bytes, &bar[buf]);
  };
  if (tid == 0) {
    mbarrier_init(&bar[0], 1); mbarrier_init(&bar[1], 1);
    if (int(blockIdx.x) < npairs) issue(blockIdx.x, 0);
    if (int(blockIdx.x + gridDim.x) < npairs) issue(blockIdx.x + gridDim.x, 1);
  }
  for (int i = tid; i < 25 * 12; i += F1_THREADS) {
    int tap = i / 12, c = i % 12;
    w_s[i] = c < C1 ? params[OFF_C1W + c * 25 + tap] : 0.f;
  }
  if (tid < 12) b_s[tid] = tid < C1 ? params[OFF_C1B + tid] : 0.f;
  __syncthreads();
  const int s = tid / 144, p = tid % 144, py = p / 12, px = p % 12;
  uint32_t it = 0;
  for (int pair = blockIdx.x; pair < npairs; pair += gridDim.x, ++it) {
    const int buf = it & 1;
    const int b0 = pair * 2;
    const int nvalid = min(2, B - b0);
    mbarrier_wait(&bar[buf], (it >> 1) & 1);
    const float* img = img_s[buf];
    if (s < nvalid) {
    float patch[6][6];
    const float* ip = img + s * IMG_PIX + (2 * py) * IMG + 2 * px;
#pragma unroll
    for (int r = 0; r < 6; ++r)
#pragma unroll
      for (int j = 0; j < 3; ++j) {
        float2 v = *reinterpret_cast<const float2*>(ip + r * IMG + 2 * j);
        patch[r][2 * j] = v.x; patch[r][2 * j + 1] = v.y;
      }
    // accumulators as packed channel pairs: one FFMA2 (fma.rn.f32x2) updates channels (2p, 2p+1) of a position
    unsigned long long acc[4][5];
#pragma unroll
    for (int q = 0; q < 4; ++q)
#pragma unroll
      for (int p2 = 0; p2 < 5; ++p2) acc[q][p2] = pack2(b_s[2 * p2], b_s[2 * p2 + 1]);
#pragma unroll
    for (int kh = 0; kh < 5; ++kh)
#pragma unroll
      for (int kw = 0; kw < 5; ++kw) {
        const ulonglong2* wp = reinterpret_cast<const ulonglong2*>(w_s + (kh * 5 + kw) * 12);
        const ulonglong2 w01 = wp[0], w23 = wp[1];
        const unsigned long long w4 = *reinterpret_cast<const unsigned long long*>(w_s + (kh * 5 + kw) * 12 + 8);
        const unsigned long long w[5] = {w01.x, w01.y, w23.x, w23.y, w4};
#pragma unroll
        for (int dy = 0; dy < 2; ++dy)
#pragma unroll
          for (int dx = 0; dx < 2; ++dx) {
            const float v = patch[dy + kh][dx + kw];
            const unsigned long long v2 = pack2(v, v);
#pragma unroll
            for (int p2 = 0; p2 < 5; ++p2) fma2(acc[dy * 2 + dx][p2], v2, w[p2]);
          }
      }
    const size_t ob = size_t(b0 + s) * A1_PER + p;
#pragma unroll
    for (int c = 0; c < C1; ++c) {
      float best; int bi;
      if (c & 1) pool4(hi_f(acc[0][c >> 1]), hi_f(acc[1][c >> 1]), hi_f(acc[2][c >> 1]), hi_f(acc[3][c >> 1]), best, bi);
      else pool4(lo_f(acc[0][c >> 1]), lo_f(acc[1][c >> 1]), lo_f(acc[2][c >> 1]), lo_f(acc[3][c >> 1]), best, bi);
      a1[ob + c * 144] = fmaxf(best, 0.f);
      idx1[ob + c * 144] = uint8_t(bi);
    }
    }
    __syncthreads();                                        // everyone is done with img_s[buf]
    if (tid == 0 && pair + 2 * int(gridDim.x) < npairs) issue(pair + 2 * gridDim.x, buf);
  }
}

// ---------------------------------------------------------------------------
// K_F3: head.  Tile of 16 samples per CTA iteration.  fc1.weight (64 KB) and the a2 tile arrive in shared
// memory by cp.async.bulk (one thread issues, an mbarrier counts the bytes) -- no register staging, no chain
// of exposed L2 latencies.  fc1 is split-K over the lanes of a warp: warp = 4 fc1 outputs, lane =
// (k-quarter, sample pair); the 4 partial sums meet with two shuffles.  13 warps per CTA, 2 CTAs per SM.
// ---------------------------------------------------------------------------
constexpr int HT = 16;              // samples per tile
constexpr int HF_WARPS = 13;        // one warp per fc1 output quad (52 >= 50)
constexpr int HF_THREADS = HF_WARPS * 32;
constexpr int HF_A2S = 324;         // a2 row stride (floats): conflict-free LDS.128 across samples
constexpr int HF_HS = 53;           // h row stride
constexpr int HF_SMEM = (F1 * A2_PER + HT * HF_A2S + 52 + HT * HF_HS + 50 * 12 + 12 + HT * 12) * 4 + 16;
__global__ void __launch_bounds__(HF_THREADS, 2)
head_fwd_kernel(const float* __restrict__ params, const float* __restrict__ a2,
                const int64_t* __restrict__ target, const float* __restrict__ mask1d,
                uint64_t seed, uint64_t offset, const int64_t* __restrict__ offset_dev, int training,
                float* __restrict__ h_out,
                float* __restrict__ logp_out, float* __restrict__ loss_out,
                int64_t* __restrict__ argmax_out, unsigned long long* __restrict__ correct_out,
                float* __restrict__ loss_part, unsigned int* __restrict__ counters, int B) {
  extern __shared__ __align__(128) float smem[];
  float* w1_s = smem;                    // [50][320] natural layout (bulk copy)
  float* a2_s = w1_s + F1 * A2_PER;      // [16][324]
  float* b1_s = a2_s + HT * HF_A2S;      // [52]
  float* h_s = b1_s + 52;                // [16][53]
  float* w2_s = h_s + HT * HF_HS;        // [50][12]
  float* b2_s = w2_s + 50 * 12;          // [12]
  float* z_s = b2_s + 12;                // [16][12]
  uint64_t* bar = reinterpret_cast<uint64_t*>(z_s + HT * 12);   // 8-byte aligned: all sizes above are even
  if (offset_dev) offset += uint64_t(*offset_dev);
  const int tid = threadIdx.x;
  const int ntiles = (B + HT - 1) / HT;
  if (tid == 0) mbarrier_init(bar, 1);
  __syncthreads();
  auto issue_tile = [&](int tile, bool with_w1) {       // thread 0 only
    const int b0 = tile * HT;
    const int nvalid = min(HT, B - b0);
    mbarrier_expect_tx(bar, uint32_t(nvalid) * (A2_PER * 4) + (with_w1 ? F1 * A2_PER * 4 : 0));
    if (with_w1) {
      bulk_copy_g2s(w1_s, params + OFF_F1W, F1 * A2_PER * 4, bar);   // parameters: may be fetched while conv2_fwd still runs
      pdl_wait();                                                    // a2 is conv2_fwd's output
    }
    for (int s = 0; s < nvalid; ++s)
      bulk_copy_g2s(a2_s + s * HF_A2S, a2 + size_t(b0 + s) * A2_PER, A2_PER * 4, bar);
  };
  pdl_trigger();
  if (tid == 0 && blockIdx.x < ntiles) issue_tile(blockIdx.x, true);
  if (tid < 52) b1_s[tid] = tid < F1 ? params[OFF_F1B + tid] : 0.f;
  for (int i = tid; i < 50 * 12; i += HF_THREADS) {
    int k = i / 12, c = i % 12;
    w2_s[i] = c < NC ? params[OFF_F2W + c * F1 + k] : 0.f;
  }
  if (tid < 12) b2_s[tid] = tid < NC ? params[OFF_F2B + tid] : 0.f;

  const int jg = tid >> 5, lane = tid & 31, kp = lane >> 3, sg = lane & 7;
  int jrow[4];
#pragma unroll
  for (int jj = 0; jj < 4; ++jj) jrow[jj] = min(jg * 4 + jj, F1 - 1) * A2_PER + kp * 80;   // clamp: cols 50,51 discarded
  float loss_acc = 0.f;                    // thread 0 only
  int corr_acc = 0;
  uint32_t phase = 0;
  pdl_wait();                              // every thread: outputs (h, logp, loss partials) may only be written after this
  for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const int b0 = tile * HT;
    const int nvalid = min(HT, B - b0);
    // rows past the batch are not copied: keep them finite
    for (int i = tid; i < (HT - nvalid) * A2_PER; i += HF_THREADS)
      a2_s[(nvalid + i / A2_PER) * HF_A2S + i % A2_PER] = 0.f;
    mbarrier_wait(bar, phase);
    phase ^= 1u;
    __syncthreads();
    {
      float acc[2][4];
#pragma unroll
      for (int i = 0; i < 2; ++i)
#pragma unroll
        for (int jj = 0; jj < 4; ++jj) acc[i][jj] = 0.f;
      const float* ar0 = a2_s + sg * HF_A2S + kp * 80;
      const float* ar1 = a2_s + (sg + 8) * HF_A2S + kp * 80;
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
      // combine the 4 k-quarters (fixed order: xor 8 then xor 16)
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
              if (training && s < nvalid) {
                float keep = mask1d ? mask1d[size_t(b) * F1 + j]
                                    : philox_keep(seed, offset, 2u, uint64_t(b) * F1 + j);
                v = v * (keep * 2.f);
              }
              h_s[s * HF_HS + j] = v;
              if (s < nvalid) h_out[size_t(b) * F1 + j] = v;
            }
          }
        }
      }
    }
    __syncthreads();
    // the a2 tile is consumed: prefetch the next one while fc2 / softmax run
    if (tid == 0 && tile + int(gridDim.x) < ntiles) issue_tile(tile + gridDim.x, false);
    // fc2: thread (s, class pair)
    if (tid < HT * 5) {
      const int s = tid / 5, c2 = (tid % 5) * 2;
      float z0 = b2_s[c2], z1 = b2_s[c2 + 1];
      const float* hr = h_s + s * HF_HS;
#pragma unroll 10
      for (int k = 0; k < F1; ++k) {
        float hv = hr[k];
        float2 w = *reinterpret_cast<const float2*>(w2_s + k * 12 + c2);
        z0 = fmaf(hv, w.x, z0); z1 = fmaf(hv, w.y, z1);
      }
      z_s[s * 12 + c2] = z0; z_s[s * 12 + c2 + 1] = z1;
    }
    __syncthreads();
    if (tid < 32) {                        // warp 0: lanes 0..15 hold the tile's samples
      const int s = tid;
      float nll = 0.f; int ok = 0;
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
        if (target) {
          const int t = int(target[b0 + s]);
          float picked = 0.f;
#pragma unroll
          for (int c = 0; c < NC; ++c) if (c == t) picked = lp[c];
          nll = -picked;
          ok = (am == t);
        }
      }
      float v = nll;                       // fixed-order shuffle tree
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
      int okc = __popc(__ballot_sync(0xffffffffu, ok));
      if (tid == 0) { loss_acc += v; corr_acc += okc; }
    }
  }
  // loss: one partial per CTA, the last CTA to arrive adds them in a fixed order (lane-strided loads that
  // are all in flight at once, then a shuffle tree) -- deterministic, and no serial chain of L2 round trips
  __shared__ unsigned int ticket_s;
  if (tid == 0) {
    if (correct_out && corr_acc) atomicAdd(correct_out, (unsigned long long)corr_acc);
    ticket_s = 0u;
    if (loss_out) {
      loss_part[blockIdx.x] = loss_acc;
      __threadfence();
      ticket_s = atomicAdd(&counters[0], 1u) + 1u;
    }
  }
  __syncthreads();
  if (loss_out && ticket_s == gridDim.x && tid < 32) {
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

namespace ddp {
// K_F1 + K_F2: x -> a1, idx1 -> a2, idx2 (shared by ddp_net_forward and the fused training entry point)
int launch_fwd_convs(const float* params, const float* x, const float* mask2d, uint64_t seed, uint64_t offset,
                     const int64_t* offset_dev, int training, const Workspace& ws, int B, cudaStream_t stream) {
  const int sms = sm_count();
  prof_mark(0, stream);
  {
    int npairs = (B + 1) / 2;
    static int f1_per_sm = -1;
    if (f1_per_sm < 0) { const char* e = getenv("DDP_F1_CTAS_PER_SM"); f1_per_sm = e ? atoi(e) : 2; if (f1_per_sm < 1 || f1_per_sm > 2) f1_per_sm = 2; }
    int grid = npairs < sms * f1_per_sm ? npairs : sms * f1_per_sm;
    DDP_CUDA(launch_pdl(conv1_fwd_kernel, dim3(grid), dim3(F1_THREADS), 0, stream, true, params, x, ws.a1, ws.idx1, B));
  }
  prof_mark(1, stream);
  if (int rc = launch_conv2_fwd_tc(params, ws.a1, mask2d, seed, offset, offset_dev, training, ws.a2, ws.idx2, B, stream))
    return rc;
  prof_mark(2, stream);
  return 0;
}
// K_F3 (shared by ddp_net_forward and the whole-step entry point)
int launch_head_fwd(const float* params, const Workspace& ws, const int64_t* target, const float* mask1d, uint64_t seed,
                    uint64_t offset, const int64_t* offset_dev, int training, float* out_logp, float* out_loss,
                    int64_t* out_argmax, int64_t* out_correct, const Scratch& sc, int B, cudaStream_t stream) {
  int ntiles = (B + HT - 1) / HT;
  int grid = ntiles < MAX_PARTS ? ntiles : MAX_PARTS;   // one loss partial per CTA
  static bool attr_done[64] = {false};       // function attributes are per device
  int dev = 0;
  DDP_CUDA(cudaGetDevice(&dev));
  if (dev >= 0 && dev < 64 && !attr_done[dev]) {
    DDP_CUDA(cudaFuncSetAttribute(head_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, HF_SMEM));
    attr_done[dev] = true;
  }
  DDP_CUDA(launch_pdl(head_fwd_kernel, dim3(grid), dim3(HF_THREADS), HF_SMEM, stream, true,
                      params, ws.a2, target, mask1d, seed, offset, offset_dev, training, ws.h, out_logp, out_loss, out_argmax,
                      reinterpret_cast<unsigned long long*>(out_correct), sc.loss_part, sc.counters, B));
  return 0;
}
}  // namespace ddp

using namespace ddp;

extern "C" int ddp_net_forward(const float* params, const float* x, const int64_t* target,
                               const float* mask2d, const float* mask1d, uint64_t seed,
                               uint64_t offset, const int64_t* offset_dev, int training, int64_t batch,
                               void* workspace,
                               float* out_logp, float* out_loss, int64_t* out_argmax,
                               int64_t* out_correct, void* scratch, void* stream_) {
  DDP_REQUIRE(params && x && workspace && out_logp, "params, x, workspace and out_logp are required");
  DDP_REQUIRE(batch >= 1 && batch <= (int64_t(1) << 24), "batch out of range");
  DDP_REQUIRE(aligned16(params) && aligned16(x) && aligned16(workspace), "params/x/workspace must be 16-byte aligned");
  DDP_REQUIRE(!(out_loss || out_correct) || target, "target is required for out_loss/out_correct");
  DDP_REQUIRE(!out_loss || scratch, "scratch is required for out_loss");
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  const int B = int(batch);
  const int sms = sm_count();
  Workspace ws = carve_workspace(workspace, batch);

  if (int rc = launch_fwd_convs(params, x, mask2d, seed, offset, offset_dev, training, ws, B, stream)) return rc;
  {
    Scratch sc{};
    if (scratch) sc = carve_scratch(scratch, batch);
    if (int rc = launch_head_fwd(params, ws, target, mask1d, seed, offset, offset_dev, training, out_logp, out_loss, out_argmax,
                                 out_correct, sc, B, stream)) return rc;
  }
  prof_mark(3, stream);
  DDP_CUDA(cudaGetLastError());
  return 0;
}
