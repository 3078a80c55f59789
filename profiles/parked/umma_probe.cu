// Probe: one tcgen05.mma chain (M=128, N=256 or 16, K=32) on SWIZZLE_NONE K-major operands written by generic stores;
// checks which of (LBO, SBO) assignments reproduces the CPU product.  Build: nvcc -gencode arch=compute_100a,code=sm_100a
#include <cstdint>
#include <cuda_bf16.h>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
#include <cmath>
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint64_t umma_desc(uint32_t saddr, uint32_t lbo, uint32_t sbo) {
  return (uint64_t)((saddr & 0x3ffffu) >> 4) | ((uint64_t)(lbo >> 4) << 16) | ((uint64_t)(sbo >> 4) << 32) | (1ull << 46);
}
__device__ __forceinline__ uint32_t umma_idesc(int M, int N) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}
__global__ void __launch_bounds__(128) probe(const uint16_t* A /*[128][K]*/, const uint16_t* B /*[N][K]*/, float* D /*[128][N]*/,
                                             int N, int K, int swap) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ __align__(8) uint64_t bar;
  __shared__ uint32_t s_tmem;
  const int tid = threadIdx.x, warp = tid >> 5;
  uint8_t* sa = smem;                    // K/8 chunks x 128 rows x 16 B
  uint8_t* sb = smem + (K / 8) * 2048;   // K/8 chunks x N rows x 16 B
  for (int idx = tid; idx < 128 * K; idx += 128) {
    int r = idx / K, k = idx % K;
    *(uint16_t*)(sa + (k / 8) * 2048 + r * 16 + (k % 8) * 2) = A[idx];
  }
  for (int idx = tid; idx < N * K; idx += 128) {
    int n = idx / K, k = idx % K;
    *(uint16_t*)(sb + (k / 8) * N * 16 + n * 16 + (k % 8) * 2) = B[idx];
  }
  if (tid == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar)) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&s_tmem)), "r"(256u) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = s_tmem;
  if (tid == 0) {
    uint32_t id = umma_idesc(128, N);
    for (int s = 0; s < K / 16; s++) {
      uint32_t a_lbo = 2048, a_sbo = 128, b_lbo = N * 16, b_sbo = 128;
      if (swap) { uint32_t t = a_lbo; a_lbo = a_sbo; a_sbo = t; t = b_lbo; b_lbo = b_sbo; b_sbo = t; }
      uint64_t ad = umma_desc(smem_u32(sa) + s * 4096, a_lbo, a_sbo);
      uint64_t bd = umma_desc(smem_u32(sb) + s * 2 * N * 16, b_lbo, b_sbo);
      uint32_t acc = s > 0;
      asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem),
                   "l"(ad), "l"(bd), "r"(id), "r"(acc) : "memory");
    }
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar)) : "memory");
  }
  uint32_t done = 0;
  long long t0 = clock64();
  while (!done) {
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(done) : "r"(smem_u32(&bar)), "r"(0u) : "memory");
    if (clock64() - t0 > 2000000000ll) __trap();
  }
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  for (int c = 0; c < N; c += 8) {
    uint32_t r[8];
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
                 : "r"(tmem + ((uint32_t)(warp * 32) << 16) + c));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
    for (int k = 0; k < 8; k++) D[tid * N + c + k] = __uint_as_float(r[k]);
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(256u) : "memory");
}
static uint16_t f2bf(float f) { uint32_t u; memcpy(&u, &f, 4); u += 0x7fffu + ((u >> 16) & 1u); return (uint16_t)(u >> 16); }
static float bf2f(uint16_t b) { uint32_t u = (uint32_t)b << 16; float f; memcpy(&f, &u, 4); return f; }
int main() {
  for (int N : {256, 16}) for (int swap : {0, 1}) {
    const int K = 32;
    std::vector<uint16_t> A(128 * K), B(N * K);
    srand(1);
    for (auto& v : A) v = f2bf((rand() % 17 - 8) / 4.0f);
    for (auto& v : B) v = f2bf((rand() % 13 - 6) / 8.0f);
    uint16_t *dA, *dB; float* dD;
    cudaMalloc(&dA, A.size() * 2); cudaMalloc(&dB, B.size() * 2); cudaMalloc(&dD, 128 * N * 4);
    cudaMemcpy(dA, A.data(), A.size() * 2, cudaMemcpyHostToDevice); cudaMemcpy(dB, B.data(), B.size() * 2, cudaMemcpyHostToDevice);
    cudaMemset(dD, 0xff, 128 * N * 4);
    size_t sm = (K / 8) * 2048 + (K / 8) * N * 16;
    cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm);
    probe<<<1, 128, sm>>>(dA, dB, dD, N, K, swap);
    cudaError_t e = cudaDeviceSynchronize();
    std::vector<float> D(128 * N);
    cudaMemcpy(D.data(), dD, D.size() * 4, cudaMemcpyDeviceToHost);
    double maxerr = 0;
    for (int r = 0; r < 128; r++) for (int n = 0; n < N; n++) {
      float s = 0; for (int k = 0; k < K; k++) s += bf2f(A[r * K + k]) * bf2f(B[n * K + k]);
      double d = fabs(s - D[r * N + n]); if (!(d <= maxerr)) maxerr = d;
    }
    printf("probe N=%d swap=%d: %s maxerr=%g\n", N, swap, cudaGetErrorString(e), maxerr);
    if (e != cudaSuccess) return 1;
  }
  return 0;
}
