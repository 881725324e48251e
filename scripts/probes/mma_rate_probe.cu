// Probe: issue rate of small-N tcgen05.mma kind::tf32 (A from TMEM, B from shared memory) as a function of N and of how
// many independent accumulators the instructions rotate over.  One CTA per SM is enough: prints cycles per MMA.
#include "../../speakingstyle_b200/csrc/tc_common.cuh"
using namespace sstc;

__device__ __forceinline__ void umma_ts(uint32_t d, uint32_t a, uint64_t db, uint32_t idesc, uint32_t accum) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n\t}" ::"r"(d), "r"(a), "l"(db), "r"(idesc), "r"(accum)
      : "memory");
}
__device__ __forceinline__ void umma_ss(uint32_t d, uint64_t da, uint64_t db, uint32_t idesc, uint32_t accum) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}" ::"r"(d), "l"(da), "l"(db), "r"(idesc), "r"(accum)
      : "memory");
}

// mode 0: A in TMEM (TS), mode 1: A in shared memory (SS), mode 2: TS alternating N = 2*N / N instructions (two idesc),
// mode 3: TS while the other three warps keep storing to TMEM (tcgen05.st + wait::st) and reading shared memory
template <int N, int NACC, int MODE>
__global__ void __launch_bounds__(128, 1) mma_rate_kernel(long long* out, int iters) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem0 = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* sm = smem_raw + (smem0 - smem_u32(smem_raw));
  // B: 128 rows x 128 B (zeros), A (SS mode): 128 rows x 128 B at +16 KB
  for (int i = threadIdx.x; i < 32768 / 4; i += 128) reinterpret_cast<float*>(sm)[i] = 0.f;
  const uint32_t bar = smem0 + 32768, tslot = bar + 8;
  volatile uint32_t* tslot_ptr = reinterpret_cast<volatile uint32_t*>(sm + 32768 + 8);
  if (threadIdx.x == 0) {
    mbar_init(bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  fence_proxy_async();
  const int warp = threadIdx.x >> 5;
  if (warp == 0) tmem_alloc(tslot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tslot_ptr;
  const uint32_t idesc = make_idesc(2u, N);
  const uint32_t idesc_b = make_idesc(2u, 2 * N <= 256 ? 2 * N : N);
  volatile int* flag = reinterpret_cast<volatile int*>(sm + 32768 + 48);
  if (threadIdx.x == 0) *flag = 0;
  __syncthreads();
  if (MODE == 3 && warp > 0) {
    uint32_t v[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) v[i] = i + threadIdx.x;
    const uint32_t ta = tmem + ((uint32_t)(warp * 32) << 16) + 256;
    float acc = 0.f;
    while (*flag == 0) {
      const float4 x = *reinterpret_cast<const float4*>(sm + ((threadIdx.x * 16) & 16383));
      acc += x.x;
      asm volatile("tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};" ::"r"(ta),
                   "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]), "r"(v[8]), "r"(v[9]), "r"(v[10]),
                   "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15]) : "memory");
      asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
    }
    if (acc == 123.f) out[3] = 1;
  }
  if (warp == 0 && elect_one()) {
    const uint32_t a_tmem = tmem + 448;  // 64 columns of A (garbage is fine)
    long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
      for (int u = 0; u < 12; ++u) {
        const uint32_t d = tmem + (u % NACC) * N;
        if (MODE == 0 || MODE == 3)
          umma_ts(d, a_tmem + (u & 3) * 8, make_desc(smem0 + (u & 3) * 32), idesc, 1u);
        else if (MODE == 2)
          umma_ts((u & 1) ? tmem + N : tmem, a_tmem + (u & 3) * 8 + (u & 1) * 32, make_desc(smem0 + (u & 3) * 32),
                  (u & 1) ? idesc : idesc_b, 1u);
        else
          umma_ss(d, make_desc(smem0 + 16384 + (u & 3) * 32), make_desc(smem0 + (u & 3) * 32), idesc, 1u);
      }
    }
    long long t1 = clock64();
    umma_commit(bar);
    mbar_wait(bar, 0);
    long long t2 = clock64();
    *flag = 1;
    if (blockIdx.x == 0) {
      out[0] = t1 - t0;
      out[1] = t2 - t0;
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 512);
}

template <int N, int NACC, int MODE>
static void run(long long* dev, long long* host, int iters, const char* tag) {
  const int smem = 32768 + 128 + 1024;
  cudaFuncSetAttribute(mma_rate_kernel<N, NACC, MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  mma_rate_kernel<N, NACC, MODE><<<148, 128, smem>>>(dev, iters);
  cudaDeviceSynchronize();
  cudaMemcpy(host, dev, 16, cudaMemcpyDeviceToHost);
  printf("%s N=%3d acc=%d : issue %.1f cyc/MMA, complete %.1f cyc/MMA (math %d)\n", tag, N, NACC,
         (double)host[0] / (12.0 * iters), (double)host[1] / (12.0 * iters), N / 2);
}

extern "C" __attribute__((visibility("default"))) int mma_rate_probe() {
  long long *dev, host[4];
  cudaMalloc(&dev, 32);
  const int it = 2000;
  run<32, 1, 0>(dev, host, it, "TS"); run<32, 3, 0>(dev, host, it, "TS"); run<32, 6, 0>(dev, host, it, "TS");
  run<64, 1, 0>(dev, host, it, "TS"); run<64, 3, 0>(dev, host, it, "TS"); run<64, 6, 0>(dev, host, it, "TS");
  run<128, 1, 0>(dev, host, it, "TS"); run<128, 3, 0>(dev, host, it, "TS");
  run<32, 1, 1>(dev, host, it, "SS"); run<32, 3, 1>(dev, host, it, "SS");
  run<64, 3, 1>(dev, host, it, "SS"); run<128, 3, 1>(dev, host, it, "SS");
  run<256, 1, 0>(dev, host, it, "TS");
  run<32, 1, 2>(dev, host, it, "TS alternating N=64/32"); run<64, 1, 2>(dev, host, it, "TS alternating N=128/64");
  run<32, 1, 3>(dev, host, it, "TS + concurrent tcgen05.st"); run<64, 1, 3>(dev, host, it, "TS + concurrent tcgen05.st");
  cudaFree(dev);
  return (int)cudaGetLastError();
}
