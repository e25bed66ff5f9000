// Standalone probe for the tcgen05 (UMMA) building blocks used by the tensor-core conv kernels:
//   - no-swizzle K-major smem descriptors (LBO/SBO meaning), kind::tf32 instruction descriptor
//   - row-shifted A start addresses (the "shifted window" implicit-GEMM trick)
//   - TMEM alloc / tcgen05.ld lane mapping / commit -> mbarrier
// Inputs are small integers (exact in TF32) so the result must match the host bit for bit.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o tools/probe/umma_probe tools/probe/umma_probe.cu
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); exit(2); } } while (0)

constexpr int ROWS = 192;   // A rows resident in smem (128 used per MMA + room for shifts)
constexpr int KTOT = 16;    // 2 k-steps of 8
constexpr int N = 32;

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ uint64_t make_desc(uint32_t saddr, uint32_t lbo, uint32_t sbo) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr & 0x3FFFFu) >> 4);
  d |= (uint64_t)((lbo >> 4) & 0x3FFFu) << 16;
  d |= (uint64_t)((sbo >> 4) & 0x3FFFu) << 32;
  d |= (uint64_t)1 << 46;  // descriptor version 1 (sm_100)
  return d;                // layout_type (bits 61..63) = 0: no swizzle
}

__global__ void __launch_bounds__(128, 1)
probe_kernel(const float* __restrict__ A_g, const float* __restrict__ B_g, float* __restrict__ D_g,
             int shift_rows, int swap_lbo_sbo, int m_dim) {
  __shared__ __align__(128) float A_s[(KTOT / 4) * ROWS * 4];
  __shared__ __align__(128) float B_s[(KTOT / 4) * N * 4];
  __shared__ __align__(8) uint64_t mbar;
  __shared__ uint32_t tmem_base_s;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  for (int i = tid; i < (KTOT / 4) * ROWS * 4; i += 128) A_s[i] = A_g[i];
  for (int i = tid; i < (KTOT / 4) * N * 4; i += 128) B_s[i] = B_g[i];
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" :: "r"(smem_u32(&tmem_base_s)), "n"(32));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  if (tid == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(smem_u32(&mbar)), "r"(1));
    asm volatile("fence.mbarrier_init.release.cluster;");
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic smem writes -> visible to the MMA's async proxy
  asm volatile("tcgen05.fence::before_thread_sync;");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;");
  const uint32_t tmem = tmem_base_s;
  if (tid == 0) {
    // instruction descriptor: D=f32, A=B=tf32, both K-major, N, M
    const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | (uint32_t(N >> 3) << 17) | (uint32_t(m_dim >> 4) << 24);
    const uint32_t a_lbo = ROWS * 16, a_sbo = 128, b_lbo = N * 16, b_sbo = 128;
    for (int ks = 0; ks < KTOT / 8; ++ks) {
      const uint32_t a_addr = smem_u32(A_s) + ks * 2 * a_lbo + shift_rows * 16;
      const uint32_t b_addr = smem_u32(B_s) + ks * 2 * b_lbo;
      const uint64_t ad = swap_lbo_sbo ? make_desc(a_addr, a_sbo, a_lbo) : make_desc(a_addr, a_lbo, a_sbo);
      const uint64_t bd = swap_lbo_sbo ? make_desc(b_addr, b_sbo, b_lbo) : make_desc(b_addr, b_lbo, b_sbo);
      const uint32_t acc = ks > 0 ? 1u : 0u;
      asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                   "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
                   :: "r"(tmem), "l"(ad), "l"(bd), "r"(idesc), "r"(acc) : "memory");
    }
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" :: "r"(smem_u32(&mbar)) : "memory");
  }
  // everyone waits for the MMAs (phase 0)
  {
    uint32_t done = 0;
    int spins = 0;
    while (!done) {
      asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.b32 %0, 1, 0, p;\n\t}"
                   : "=r"(done) : "r"(smem_u32(&mbar)), "r"(0) : "memory");
      if (++spins > (1 << 22)) { if (tid == 0) printf("TIMEOUT waiting for MMA commit\n"); break; }
    }
  }
  asm volatile("tcgen05.fence::after_thread_sync;");
  if (warp * 32 < m_dim) {
    uint32_t r[32];
    const uint32_t taddr = tmem + (uint32_t(warp * 32) << 16);
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 "
                 "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
                   "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
                   "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
                   "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
                 : "r"(taddr));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
    for (int c = 0; c < 32; ++c) D_g[(warp * 32 + lane) * N + c] = __uint_as_float(r[c]);
  }
  asm volatile("tcgen05.fence::before_thread_sync;");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" :: "r"(tmem), "n"(32));
}

static float aval(int r, int k) { return float(((r * 7 + k * 3) % 9) - 4); }
static float bval(int n, int k) { return float(((n * 5 + k * 11) % 7) - 3); }

int main() {
  std::vector<float> A((KTOT / 4) * ROWS * 4), B((KTOT / 4) * N * 4);
  for (int kc = 0; kc < KTOT / 4; ++kc)
    for (int r = 0; r < ROWS; ++r)
      for (int e = 0; e < 4; ++e) A[(kc * ROWS + r) * 4 + e] = aval(r, kc * 4 + e);
  for (int kc = 0; kc < KTOT / 4; ++kc)
    for (int n = 0; n < N; ++n)
      for (int e = 0; e < 4; ++e) B[(kc * N + n) * 4 + e] = bval(n, kc * 4 + e);
  float *Ad, *Bd, *Dd;
  CK(cudaMalloc(&Ad, A.size() * 4)); CK(cudaMalloc(&Bd, B.size() * 4)); CK(cudaMalloc(&Dd, 128 * N * 4));
  CK(cudaMemcpy(Ad, A.data(), A.size() * 4, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(Bd, B.data(), B.size() * 4, cudaMemcpyHostToDevice));
  struct Case { int shift, swap, m; const char* name; } cases[] = {
      {0, 0, 128, "M128 LBO=K-chunk stride, SBO=8-row stride"},
      {8, 0, 128, "M128 rows shifted by 8"}, {40, 0, 128, "M128 rows shifted by 40"}, {3, 0, 128, "M128 rows shifted by 3"},
      {0, 0, 64, "M64"}, {16, 0, 64, "M64 rows shifted by 16"}};
  int all_ok = 1;
  for (auto& c : cases) {
    CK(cudaMemset(Dd, 0xff, 128 * N * 4));
    probe_kernel<<<1, 128>>>(Ad, Bd, Dd, c.shift, c.swap, c.m);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("[%s] CUDA error: %s\n", c.name, cudaGetErrorString(e)); return 3; }
    std::vector<float> D(128 * N);
    CK(cudaMemcpy(D.data(), Dd, D.size() * 4, cudaMemcpyDeviceToHost));
    int bad = 0; float first_got = 0, first_exp = 0; int first_idx = -1;
    for (int m = 0; m < c.m; ++m)
      for (int n = 0; n < N; ++n) {
        float exp = 0;
        for (int k = 0; k < KTOT; ++k) exp += aval(m + c.shift, k) * bval(n, k);
        if (D[m * N + n] != exp) { if (!bad) { first_got = D[m * N + n]; first_exp = exp; first_idx = m * N + n; } ++bad; }
      }
    printf("[%s] %s  mismatches=%d", c.name, bad ? "FAIL" : "PASS", bad);
    if (bad) printf("  first at (m=%d,n=%d): got %g expected %g", first_idx / N, first_idx % N, first_got, first_exp);
    printf("\n");
    if (bad && !c.swap) all_ok = 0;
  }
  printf(all_ok ? "UMMA_PROBE_OK\n" : "UMMA_PROBE_FAILED\n");
  return 0;
}
