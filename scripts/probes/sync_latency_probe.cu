// Probe: latency of the hand-over primitives the tcgen05 pipelines are built from.
//  (1) mbarrier ping-pong between two warps (arrive -> try_wait success), per hop
//  (2) tcgen05.commit with nothing outstanding -> mbarrier completes -> same thread's try_wait returns
//  (3) 8 MMAs (N = 32, TS form) + commit -> wait
//  (4) tcgen05.st x2 + wait::st
#include "../../speakingstyle_b200/csrc/tc_common.cuh"
using namespace sstc;

__device__ __forceinline__ void umma_ts(uint32_t d, uint32_t a, uint64_t db, uint32_t idesc, uint32_t accum) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n\t}" ::"r"(d), "r"(a), "l"(db), "r"(idesc), "r"(accum)
      : "memory");
}

__global__ void __launch_bounds__(128, 1) sync_latency_kernel(long long* out, int iters) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem0 = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* sm = smem_raw + (smem0 - smem_u32(smem_raw));
  for (int i = threadIdx.x; i < 16384 / 4; i += 128) reinterpret_cast<float*>(sm)[i] = 0.f;
  const uint32_t bar1 = smem0 + 16384, bar2 = bar1 + 8, bar3 = bar1 + 16, tslot = bar1 + 32;
  volatile uint32_t* tslot_ptr = reinterpret_cast<volatile uint32_t*>(sm + 16384 + 32);
  if (threadIdx.x == 0) {
    mbar_init(bar1, 1); mbar_init(bar2, 1); mbar_init(bar3, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  fence_proxy_async();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp == 0) tmem_alloc(tslot, 256);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tslot_ptr;
  // (1) ping-pong
  if (warp == 0 && lane == 0) {
    long long t0 = clock64();
    for (int i = 0; i < iters; ++i) {
      mbar_arrive(bar1);
      mbar_wait(bar2, i & 1);
    }
    out[0] = clock64() - t0;
  } else if (warp == 1 && lane == 0) {
    for (int i = 0; i < iters; ++i) {
      mbar_wait(bar1, i & 1);
      mbar_arrive(bar2);
    }
  }
  __syncthreads();
  if (warp == 0 && elect_one()) {
    // (2) empty commit -> wait
    long long t0 = clock64();
    for (int i = 0; i < iters; ++i) {
      umma_commit(bar3);
      mbar_wait(bar3, i & 1);
    }
    out[1] = clock64() - t0;
    // (3) 8 MMAs N = 32 + commit -> wait
    const uint32_t idesc = make_idesc(2u, 32);
    t0 = clock64();
    for (int i = 0; i < iters; ++i) {
#pragma unroll
      for (int u = 0; u < 8; ++u) umma_ts(tmem, tmem + 128 + (u & 3) * 8, make_desc(smem0 + (u & 3) * 32), idesc, 1u);
      umma_commit(bar3);
      mbar_wait(bar3, (iters + i) & 1);
    }
    out[2] = clock64() - t0;
  }
  __syncthreads();
  // (4) tcgen05.st x2 (16 columns each) + wait::st, all four warps
  {
    uint32_t v[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) v[i] = i;
    const uint32_t ta = tmem + ((uint32_t)(warp * 32) << 16) + 192;
    long long t0 = clock64();
    for (int i = 0; i < iters; ++i) {
      asm volatile("tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};" ::"r"(ta),
                   "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]), "r"(v[8]), "r"(v[9]), "r"(v[10]),
                   "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15]) : "memory");
      asm volatile("tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};" ::"r"(ta + 16),
                   "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]), "r"(v[8]), "r"(v[9]), "r"(v[10]),
                   "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15]) : "memory");
      asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
      tc_fence_before();
    }
    if (threadIdx.x == 0) out[3] = clock64() - t0;
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 256);
}

extern "C" __attribute__((visibility("default"))) int sync_latency_probe() {
  long long *dev, host[4];
  cudaMalloc(&dev, 32);
  const int smem = 16384 + 64 + 1024, it = 2000;
  cudaFuncSetAttribute(sync_latency_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  sync_latency_kernel<<<1, 128, smem>>>(dev, it);
  cudaDeviceSynchronize();
  cudaMemcpy(host, dev, 32, cudaMemcpyDeviceToHost);
  printf("mbarrier hop (arrive -> other warp's wait returns): %.0f cycles\n", (double)host[0] / it / 2);
  printf("empty tcgen05.commit -> own wait returns:           %.0f cycles\n", (double)host[1] / it);
  printf("8 MMAs (N=32, 128 cycles of math) + commit -> wait: %.0f cycles\n", (double)host[2] / it);
  printf("2 x tcgen05.st.x16 + wait::st + fence:              %.0f cycles\n", (double)host[3] / it);
  cudaFree(dev);
  return (int)cudaGetLastError();
}
