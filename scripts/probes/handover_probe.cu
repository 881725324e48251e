// Probe: the operand hand-over skeleton of the TS-form conv kernels in isolation.  NW "split" warps and one "MMA"
// thread pass SLOTS slots back and forth: split: wait empty[s] -> arrive full[s];  MMA: wait full[s] -> (optionally
// 16 small MMAs) -> tcgen05.commit empty[s].  Prints cycles per hand-over.
#include "../../speakingstyle_b200/csrc/tc_common.cuh"
using namespace sstc;

__device__ __forceinline__ void umma_ts(uint32_t d, uint32_t a, uint64_t db, uint32_t idesc, uint32_t accum) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n\t}" ::"r"(d), "r"(a), "l"(db), "r"(idesc), "r"(accum)
      : "memory");
}

// release: 0 = tcgen05.commit, 1 = plain mbarrier.arrive by the MMA thread (no tensor-core involvement)
template <int NW, int SLOTS, int MMAS, int RELEASE, int WORK>
__global__ void __launch_bounds__(32 * (2 + NW), 1) handover_kernel(long long* out, int iters) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem0 = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* sm = smem_raw + (smem0 - smem_u32(smem_raw));
  for (int i = threadIdx.x; i < 16384 / 4; i += blockDim.x) reinterpret_cast<float*>(sm)[i] = 0.f;
  const uint32_t full0 = smem0 + 16384, empty0 = full0 + 64, tslot = empty0 + 64;
  volatile uint32_t* tslot_ptr = reinterpret_cast<volatile uint32_t*>(sm + 16384 + 128);
  if (threadIdx.x == 0) {
    for (int s = 0; s < SLOTS; ++s) {
      mbar_init(full0 + 8 * s, NW);
      mbar_init(empty0 + 8 * s, 1);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  fence_proxy_async();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp == 1) tmem_alloc(tslot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tslot_ptr;
  if (warp == 1) {
    if (elect_one()) {
      const uint32_t idesc = make_idesc(2u, 32);
      int slot = 0;
      uint32_t ph = 0;
      long long t0 = clock64();
      for (int i = 0; i < iters; ++i) {
        mbar_wait(full0 + 8 * slot, ph);
        tc_fence_after();
#pragma unroll
        for (int u = 0; u < MMAS; ++u)
          umma_ts(tmem + (u & 1) * 32, tmem + 256 + slot * 64 + (u & 3) * 8, make_desc(smem0 + (u & 3) * 32), idesc, 1u);
        if (RELEASE == 0) umma_commit(empty0 + 8 * slot); else mbar_arrive(empty0 + 8 * slot);
        if (++slot == SLOTS) { slot = 0; ph ^= 1; }
      }
      out[0] = clock64() - t0;
    }
  } else if (warp >= 2) {
    int slot = 0;
    uint32_t ph = 0;
    float acc = 0.f;
    for (int i = 0; i < iters; ++i) {
      if (WORK) {  // stand-in for the split: a few hundred cycles of dependent ALU work
#pragma unroll 1
        for (int w = 0; w < WORK; ++w) acc = fmaf(acc, 1.0001f, 0.5f);
      }
      mbar_wait(empty0 + 8 * slot, ph ^ 1);
      tc_fence_after();
      asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(full0 + 8 * slot);
      if (++slot == SLOTS) { slot = 0; ph ^= 1; }
    }
    if (acc == 12345.f) out[1] = 1;
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem, 512);
}

template <int NW, int SLOTS, int MMAS, int RELEASE, int WORK>
static void run(long long* dev, const char* tag) {
  long long host[2];
  const int smem = 16384 + 256 + 1024, it = 4000;
  cudaFuncSetAttribute(handover_kernel<NW, SLOTS, MMAS, RELEASE, WORK>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  handover_kernel<NW, SLOTS, MMAS, RELEASE, WORK><<<1, 32 * (2 + NW), smem>>>(dev, it);
  cudaError_t e = cudaDeviceSynchronize();
  cudaMemcpy(host, dev, 16, cudaMemcpyDeviceToHost);
  printf("%-34s split warps %d, slots %d, MMAs %2d, work %3d: %.0f cycles per hand-over%s\n", tag, NW, SLOTS, MMAS, WORK,
         (double)host[0] / it, e == cudaSuccess ? "" : " (ERROR)");
}

extern "C" __attribute__((visibility("default"))) int handover_probe() {
  long long* dev;
  cudaMalloc(&dev, 16);
  run<8, 2, 0, 0, 0>(dev, "commit release");
  run<8, 2, 0, 1, 0>(dev, "plain arrive release");
  run<4, 2, 0, 0, 0>(dev, "commit release");
  run<1, 2, 0, 0, 0>(dev, "commit release");
  run<8, 3, 0, 0, 0>(dev, "commit release");
  run<8, 2, 16, 0, 0>(dev, "commit release + MMAs");
  run<8, 3, 16, 0, 0>(dev, "commit release + MMAs");
  run<8, 2, 16, 0, 100>(dev, "commit release + MMAs + work");
  run<8, 3, 16, 0, 100>(dev, "commit release + MMAs + work");
  run<4, 3, 16, 0, 100>(dev, "commit release + MMAs + work");
  cudaFree(dev);
  return (int)cudaGetLastError();
}
