// Probe: can a K-major SWIZZLE_128B shared-memory operand of tcgen05.mma start at an arbitrary ROW of a swizzled
// panel (start address = panel + s*128 bytes, s not a multiple of 8), and which value of the descriptor's
// base_offset field (bits 49-51) does that need?  A = 176-row fp32 panel written in the layout TMA produces
// (16-byte unit u of row r at unit u ^ (r & 7)), B = 32x32 identity, D[i][n] must come out as A[s + i][n].
// out[variant][s][128][32]; variant 0: base_offset = 0, variant 1: base_offset = s & 7.
#include "../../speakingstyle_b200/csrc/tc_common.cuh"
using namespace sstc;

constexpr int ROWS = 176, NSHIFT = 40;

__global__ void __launch_bounds__(128, 1) desc_shift_probe_kernel(float* out) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem0 = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* sm = smem_raw + (smem0 - smem_u32(smem_raw));
  float* A = reinterpret_cast<float*>(sm);                       // ROWS x 32 fp32, swizzled
  float* Bm = reinterpret_cast<float*>(sm + ROWS * 128);         // 32 x 32 identity, swizzled (ROWS*128 % 1024 == 0)
  const uint32_t bar = smem0 + ROWS * 128 + 32 * 128, tslot = bar + 8;
  volatile uint32_t* tslot_ptr = reinterpret_cast<volatile uint32_t*>(sm + ROWS * 128 + 32 * 128 + 8);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int e = threadIdx.x; e < ROWS * 32; e += 128) {
    const int r = e >> 5, c = e & 31, u = c >> 2;
    A[r * 32 + ((u ^ (r & 7)) << 2) + (c & 3)] = (float)(r * 8 + u);
  }
  for (int e = threadIdx.x; e < 32 * 32; e += 128) {
    const int n = e >> 5, c = e & 31, u = c >> 2;
    Bm[n * 32 + ((u ^ (n & 7)) << 2) + (c & 3)] = n == c ? 1.f : 0.f;
  }
  if (threadIdx.x == 0) {
    mbar_init(bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  fence_proxy_async();
  if (warp == 0) tmem_alloc(tslot, 32);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tslot_ptr;
  const uint32_t idesc = make_idesc(2u, 32);
  uint32_t phase = 0;
  for (int variant = 0; variant < 2; ++variant) {
    for (int s = 0; s < NSHIFT; ++s) {
      if (warp == 0 && elect_one()) {
        const uint64_t bo = variant ? ((uint64_t)(s & 7) << 49) : 0ull;
#pragma unroll
        for (int kk = 0; kk < 4; ++kk)
          umma_tf32(tmem, make_desc(smem0 + s * 128 + kk * 32) | bo, make_desc(smem0 + ROWS * 128 + kk * 32), idesc,
                    kk != 0);
        umma_commit(bar);
      }
      mbar_wait(bar, phase);
      phase ^= 1;
      tc_fence_after();
      float v[32];
      tmem_ld32(tmem + ((uint32_t)(warp * 32) << 16), v);
      float* o = out + ((size_t)(variant * NSHIFT + s) * 128 + warp * 32 + lane) * 32;
#pragma unroll
      for (int i = 0; i < 32; ++i) o[i] = v[i];
      tc_fence_before();
      __syncthreads();
      tc_fence_after();
    }
  }
  if (warp == 0) tmem_dealloc(tmem, 32);
}

extern "C" __attribute__((visibility("default"))) int desc_shift_probe(float* out_dev) {
  const int smem = ROWS * 128 + 32 * 128 + 64 + 1024;
  cudaFuncSetAttribute(desc_shift_probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  desc_shift_probe_kernel<<<1, 128, smem>>>(out_dev);
  return (int)cudaDeviceSynchronize();
}
