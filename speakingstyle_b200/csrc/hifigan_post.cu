// HiFi-GAN output head: leaky_relu(0.01) -> Conv1d(C -> 1, k) -> tanh (-> int16 PCM).
// Bandwidth-bound: the C-channel activation is read once (128 B per sample at C=32), one float
// (and/or one int16) is written per sample.  A CTA stages 256 + k - 1 rows in shared memory with
// the activation applied on the way in; rows are padded to C+1 words so that the per-thread dot
// products (thread = sample, stride C between neighbours) are bank-conflict free.
#include "common.cuh"

namespace {

constexpr int TS = 256;

__global__ void __launch_bounds__(TS) hifigan_post_kernel(const float* __restrict__ x, const float* __restrict__ w,
                                                          const float* __restrict__ bias, float* __restrict__ wave,
                                                          int16_t* __restrict__ pcm, float pcm_scale, float in_slope,
                                                          int T, int C, int k, const int32_t* __restrict__ n_samples,
                                                          const int64_t* __restrict__ out_offset) {
  extern __shared__ float sm[];
  const int pitch = C + 1;
  const int halo = (k - 1) / 2;
  const int rows = TS + k - 1;
  float* xs = sm;                 // [rows][pitch]
  float* ws = sm + rows * pitch;  // [k*C]
  const int b = blockIdx.y;
  const int t0 = blockIdx.x * TS;
  const int t_end = n_samples ? min(max(n_samples[b], 0), T) : T;  // ragged form: samples past the utterance's
  if (t0 >= t_end) return;                                          // end are neither computed nor written
  const float* xb = x + (int64_t)b * T * C;

  for (int i = threadIdx.x; i < k * C; i += TS) ws[i] = w[i];
  const int C4 = C >> 2;
  for (int f = threadIdx.x; f < rows * C4; f += TS) {
    const int r = f / C4, c = (f - r * C4) << 2;
    const int t = t0 + r - halo;
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (t >= 0 && t < T) v = ss::ld_stream(reinterpret_cast<const float4*>(xb + (int64_t)t * C + c));
    float* d = xs + r * pitch + c;
    d[0] = v.x > 0.f ? v.x : v.x * in_slope;
    d[1] = v.y > 0.f ? v.y : v.y * in_slope;
    d[2] = v.z > 0.f ? v.z : v.z * in_slope;
    d[3] = v.w > 0.f ? v.w : v.w * in_slope;
  }
  __syncthreads();

  const int t = t0 + threadIdx.x;
  if (t >= t_end) return;
  float acc = 0.f;
  for (int j = 0; j < k; ++j) {
    const float* row = xs + (threadIdx.x + j) * pitch;
    const float* wj = ws + j * C;
#pragma unroll 8
    for (int c = 0; c < C; ++c) acc = fmaf(row[c], wj[c], acc);
  }
  const float y = tanhf(acc + (bias ? __ldg(bias) : 0.f));
  const int64_t o = (int64_t)b * T + t;
  if (wave) wave[o] = y;
  // truncation toward zero, low 16 bits (numpy astype); ragged form: utterance b starts at out_offset[b]
  if (pcm) pcm[out_offset ? out_offset[b] + t : o] = (int16_t)(int32_t)(y * pcm_scale);
}

}  // namespace

static int post_launch(const float* x, const float* w, const float* bias, float* wave, int16_t* pcm, float pcm_scale,
                       float in_slope, int B, int T, int C, int k, const int32_t* n_samples, const int64_t* out_offset,
                       void* stream) {
  SS_REQUIRE(x && w && (wave || pcm), SS_ERR_INVALID, "ss_hifigan_post: null pointer");
  SS_REQUIRE(B > 0 && T > 0 && C > 0 && C % 4 == 0 && C <= 64 && k > 0 && k % 2 == 1 && k <= 15 && B <= 65535,
             SS_ERR_INVALID, "ss_hifigan_post: bad shape B=%d T=%d C=%d k=%d", B, T, C, k);
  SS_REQUIRE(ss::aligned16(x), SS_ERR_ALIGNMENT, "ss_hifigan_post: x must be 16-byte aligned");
  const size_t smem = ((size_t)(TS + k - 1) * (C + 1) + (size_t)k * C) * sizeof(float);
  if (smem > 48 * 1024) {
    cudaError_t e = cudaFuncSetAttribute(hifigan_post_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return ss::cuda_fail(e, "ss_hifigan_post: smem opt-in");
  }
  dim3 grid((T + TS - 1) / TS, B);
  hifigan_post_kernel<<<grid, TS, smem, (cudaStream_t)stream>>>(x, w, bias, wave, pcm, pcm_scale, in_slope, T, C, k,
                                                                n_samples, out_offset);
  SS_CHECK_LAUNCH("ss_hifigan_post");
  return SS_OK;
}

extern "C" int ss_hifigan_post(const float* x, const float* w, const float* bias, float* wave, int16_t* pcm,
                               float pcm_scale, float in_slope, int B, int T, int C, int k, void* stream) {
  return post_launch(x, w, bias, wave, pcm, pcm_scale, in_slope, B, T, C, k, nullptr, nullptr, stream);
}

extern "C" int ss_hifigan_post_trim(const float* x, const float* w, const float* bias, const int32_t* n_samples,
                                    const int64_t* out_offset, int16_t* pcm, float pcm_scale, float in_slope, int B,
                                    int T, int C, int k, void* stream) {
  SS_REQUIRE(n_samples && out_offset && pcm, SS_ERR_INVALID, "ss_hifigan_post_trim: null pointer");
  return post_launch(x, w, bias, nullptr, pcm, pcm_scale, in_slope, B, T, C, k, n_samples, out_offset, stream);
}
