// Probe of tcgen05 MN-major shared-memory descriptors (no swizzle) on sm_100a: which of LBO / SBO is the stride between
// core matrices along K and which along M/N, for operands laid out like P8 ("pixel-major, 8 channels contiguous").
//   A: M = 128 "channels", K = 16 "pixels":  a[kgroup(2)][mgroup(16)][8 k][8 m] bf16  (K-core-matrix stride 128 B * 16? see below)
// We build smem as: element (m, k) at byte  (m/8)*MSTR + (k/8)*KSTR + (k%8)*16 + (m%8)*2   (a core matrix = 8 k-rows of 16 B)
// and try descriptor (LBO, SBO) = (MSTR, KSTR) and (KSTR, MSTR).  D[128][N=64] fp32 is read back and compared on the host.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o mn_probe mn_major_probe.cu ; run: ./mn_probe
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <vector>
#include <cmath>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

constexpr int M = 128, N = 64, K = 16;
constexpr int MSTR = 256, KSTR = 128;      // A: m-groups 256 B apart (two k-groups interleaved), k-groups 128 B apart
constexpr int NSTR = 256;                  // B likewise

__global__ void probe(const __nv_bfloat16* a, const __nv_bfloat16* b, float* d, int variant) {
  __shared__ __align__(128) uint8_t sa[M / 8 * MSTR];
  __shared__ __align__(128) uint8_t sb[N / 8 * NSTR];
  __shared__ __align__(8) uint64_t bar;
  __shared__ uint32_t tmem_base_s;
  const int tid = threadIdx.x;
  for (int i = tid; i < M * K; i += blockDim.x) {
    const int m = i / K, k = i % K;
    *reinterpret_cast<__nv_bfloat16*>(sa + (m / 8) * MSTR + (k / 8) * KSTR + (k % 8) * 16 + (m % 8) * 2) = a[m * K + k];
  }
  for (int i = tid; i < N * K; i += blockDim.x) {
    const int n = i / K, k = i % K;
    *reinterpret_cast<__nv_bfloat16*>(sb + (n / 8) * NSTR + (k / 8) * KSTR + (k % 8) * 16 + (n % 8) * 2) = b[n * K + k];
  }
  if (tid == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar)));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  __syncthreads();
  if (tid < 32) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_s)), "r"(64));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::);
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = tmem_base_s;
  if (tid == 0) {
    const uint32_t lbo = variant == 0 ? MSTR : KSTR, sbo = variant == 0 ? KSTR : MSTR;
    auto desc = [&](uint32_t addr, uint32_t l, uint32_t s) {
      return (uint64_t)((addr >> 4) & 0x3FFFu) | ((uint64_t)((l >> 4) & 0x3FFFu) << 16) | ((uint64_t)((s >> 4) & 0x3FFFu) << 32) | ((uint64_t)1 << 46);
    };
    const uint64_t da = desc(smem_u32(sa), lbo, sbo);
    const uint64_t db = desc(smem_u32(sb), variant == 0 ? NSTR : KSTR, variant == 0 ? KSTR : NSTR);
    // idesc: c=f32 (1<<4), a=bf16 (1<<7), b=bf16 (1<<10), a_major=MN (1<<15), b_major=MN (1<<16), N>>3 <<17, M>>4 <<24
    const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | (1u << 15) | (1u << 16) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
                 ::"r"(tmem), "l"(da), "l"(db), "r"(idesc), "r"(0u));
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar)) : "memory");
  }
  // wait
  {
    uint32_t ok = 0; int spins = 0;
    while (!ok && ++spins < (1 << 22))
      asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.b32 %0, 1, 0, p;\n\t}" : "=r"(ok) : "r"(smem_u32(&bar)), "r"(0u) : "memory");
  }
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  if (tid < 128) {
    const int warp = tid >> 5;
    uint32_t r[32];
    for (int c0 = 0; c0 < N; c0 += 32) {
      const uint32_t taddr = tmem + ((uint32_t)(warp * 32) << 16) + c0;
      asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
                   : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]),
                     "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
                     "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31]) : "r"(taddr));
      asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
      for (int e = 0; e < 32; ++e) d[tid * N + c0 + e] = __uint_as_float(r[e]);
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (tid < 32) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(64));
}

int main() {
  std::vector<__nv_bfloat16> a(M * K), b(N * K);
  std::vector<float> af(M * K), bf(N * K);
  srand(1);
  for (int i = 0; i < M * K; ++i) { af[i] = (float)(rand() % 17 - 8); a[i] = __float2bfloat16(af[i]); }
  for (int i = 0; i < N * K; ++i) { bf[i] = (float)(rand() % 13 - 6); b[i] = __float2bfloat16(bf[i]); }
  __nv_bfloat16 *da, *db; float* dd;
  cudaMalloc(&da, a.size() * 2); cudaMalloc(&db, b.size() * 2); cudaMalloc(&dd, M * N * 4);
  cudaMemcpy(da, a.data(), a.size() * 2, cudaMemcpyHostToDevice);
  cudaMemcpy(db, b.data(), b.size() * 2, cudaMemcpyHostToDevice);
  for (int variant = 0; variant < 2; ++variant) {
    cudaMemset(dd, 0, M * N * 4);
    probe<<<1, 128>>>(da, db, dd, variant);
    cudaError_t e = cudaDeviceSynchronize();
    std::vector<float> d(M * N);
    cudaMemcpy(d.data(), dd, M * N * 4, cudaMemcpyDeviceToHost);
    double maxerr = 0; int bad = 0;
    for (int m = 0; m < M; ++m)
      for (int n = 0; n < N; ++n) {
        float ref = 0; for (int k = 0; k < K; ++k) ref += af[m * K + k] * bf[n * K + k];
        double er = fabs(ref - d[m * N + n]); if (er > maxerr) maxerr = er; if (er > 1e-3) ++bad;
      }
    printf("variant %d (LBO=%s stride): err=%s max|diff|=%g mismatches=%d/%d  d[0][0]=%g d[1][0]=%g d[8][0]=%g\n", variant,
           variant == 0 ? "M/N" : "K", cudaGetErrorString(e), maxerr, bad, M * N, d[0], d[N], d[8 * N]);
  }
  return 0;
}
