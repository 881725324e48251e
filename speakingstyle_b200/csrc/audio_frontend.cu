// Reference-mel front-end (audio/stft.py:51-81, 160-178, audio/audio_processing.py:85-91):
// wave -> reflect pad -> STFT as a strided convolution with the windowed Fourier basis -> magnitude -> mel
// filterbank -> log(clamp(., 1e-5)).  The two contractions (frames x 1024 -> 1026 Fourier rows, 513 -> 80
// mel rows) run on the 3xTF32 tensor-core convolution kernel: a 1024-sample frame at hop 256 is four consecutive
// 256-sample rows, so the strided convolution is an ordinary k = 4 "valid" convolution over hop-sized rows
// and needs no overlapping tensor map.  This file holds the three bandwidth-bound steps around them.
#include "common.cuh"

namespace {

// ypad[b, i] = y[b, reflect(i - half)], i < n_out  (F.pad(mode="reflect"): ... y2 y1 | y0 y1 ... yT-1 | yT-2 ...)
__global__ void reflect_pad_kernel(const float* __restrict__ y, float* __restrict__ out, int T, int half, int n_out) {
  const int b = blockIdx.y;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_out) return;
  int t = i - half;
  if (t < 0) t = -t;
  if (t >= T) t = 2 * (T - 1) - t;
  out[(int64_t)b * n_out + i] = (t >= 0 && t < T) ? y[(int64_t)b * T + t] : 0.f;
}

// One warp per frame: spec row = [re(0..nb-1) | im(0..nb-1) | pad], mag row = sqrt(re^2 + im^2) padded with zeros
// to mag_stride columns, energy = ||mag||_2 over the nb bins (torch.norm(magnitudes, dim=1)).
__global__ void __launch_bounds__(256) stft_magnitude_kernel(const float* __restrict__ spec, float* __restrict__ mag,
                                                             float* __restrict__ energy, int64_t rows, int nb,
                                                             int spec_stride, int mag_stride) {
  const int64_t row = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= rows) return;
  const int lane = threadIdx.x & 31;
  const float* s = spec + row * spec_stride;
  float* m = mag + row * mag_stride;
  float e = 0.f;
  for (int i = lane; i < mag_stride; i += 32) {
    float v = 0.f;
    if (i < nb) {
      const float re = s[i], im = s[nb + i];
      const float p = re * re + im * im;
      v = sqrtf(p);
      e += p;
    }
    m[i] = v;
  }
  e = ss::warp_sum(e);
  if (lane == 0 && energy) energy[row] = sqrtf(e);
}

__global__ void log_clamp_kernel(float* __restrict__ x, int64_t n, float clip) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) x[i] = logf(fmaxf(x[i], clip));
}

}  // namespace

// *flag |= 1 if any y[i] lies outside [lo, hi] (NaN counts as outside); flag must be zeroed by the caller
__global__ void __launch_bounds__(256) range_flag_kernel(const float* __restrict__ y, int64_t n, float lo, float hi,
                                                         int32_t* __restrict__ flag) {
  const int64_t base = (int64_t)blockIdx.x * 1024 + threadIdx.x;
  bool bad = false;
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const int64_t i = base + k * 256;
    if (i < n) {
      const float v = y[i];
      bad |= !(v >= lo && v <= hi);
    }
  }
  if (__any_sync(0xffffffffu, bad) && (threadIdx.x & 31) == 0) atomicOr(flag, 1);
}

extern "C" int ss_reflect_pad(const float* y, float* out, int B, int T, int half, int n_out, void* stream) {
  SS_REQUIRE(y && out, SS_ERR_INVALID, "ss_reflect_pad: null pointer");
  SS_REQUIRE(B > 0 && B <= 65535 && T > half && half >= 0 && n_out > 0 && n_out <= T + 2 * half, SS_ERR_INVALID,
             "ss_reflect_pad: bad shape B=%d T=%d half=%d n_out=%d (reflect padding needs T > half)", B, T, half, n_out);
  dim3 grid((n_out + 255) / 256, B);
  reflect_pad_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(y, out, T, half, n_out);
  SS_CHECK_LAUNCH("ss_reflect_pad");
  return SS_OK;
}

extern "C" int ss_stft_magnitude(const float* spec, float* mag, float* energy, int64_t rows, int n_bins,
                                 int spec_stride, int mag_stride, void* stream) {
  SS_REQUIRE(spec && mag, SS_ERR_INVALID, "ss_stft_magnitude: null pointer");
  SS_REQUIRE(rows > 0 && n_bins > 0 && spec_stride >= 2 * n_bins && mag_stride >= n_bins, SS_ERR_INVALID,
             "ss_stft_magnitude: bad shape rows=%lld bins=%d strides=%d,%d", (long long)rows, n_bins, spec_stride,
             mag_stride);
  const int64_t blocks = (rows + 7) / 8;
  SS_REQUIRE(blocks < (1ll << 31), SS_ERR_UNSUPPORTED, "ss_stft_magnitude: too many rows");
  stft_magnitude_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(spec, mag, energy, rows, n_bins, spec_stride,
                                                                          mag_stride);
  SS_CHECK_LAUNCH("ss_stft_magnitude");
  return SS_OK;
}

extern "C" int ss_range_flag(const float* y, int64_t n, float lo, float hi, int32_t* flag, void* stream) {
  SS_REQUIRE(y && flag && n > 0, SS_ERR_INVALID, "ss_range_flag: bad arguments");
  const int64_t blocks = (n + 1023) / 1024;
  SS_REQUIRE(blocks < (1ll << 31), SS_ERR_UNSUPPORTED, "ss_range_flag: too many elements");
  range_flag_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(y, n, lo, hi, flag);
  SS_CHECK_LAUNCH("ss_range_flag");
  return SS_OK;
}

extern "C" int ss_log_clamp(float* x, int64_t n, float clip, void* stream) {
  SS_REQUIRE(x && n > 0 && clip > 0.f, SS_ERR_INVALID, "ss_log_clamp: bad arguments");
  const int64_t blocks = (n + 255) / 256;
  SS_REQUIRE(blocks < (1ll << 31), SS_ERR_UNSUPPORTED, "ss_log_clamp: too many elements");
  log_clamp_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(x, n, clip);
  SS_CHECK_LAUNCH("ss_log_clamp");
  return SS_OK;
}
