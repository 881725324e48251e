// Duration rounding, prefix scan and the LengthRegulator frame expansion.
//
// The reference does this with a Python double loop and one .item() host sync per phoneme
// (model/modules.py:174-197).  Here: one tiny scan kernel (a warp per utterance) produces the
// integer repeat counts' inclusive prefix sums, the host reads B int32 once to size the output,
// and one bandwidth-bound kernel writes every output row exactly once - gathered source row,
// optional decoder position row added on the fly (transformer/Models.py:158-160), zeros on the
// padding - with 128-bit streaming stores.  The source rows (B*L*C floats) are re-read ~8x but
// stay in L2; algorithmic HBM bytes = 4*C*(sum_b src_len_b + B*T) + 4*B*L + 4*B.
#include <atomic>

#include "common.cuh"

namespace {

__device__ __forceinline__ float round_duration(float log_d, float control) {
  // same op order as the reference: exp, minus one, round half-to-even, times control, clamp
  const float v = rintf(expf(log_d) - 1.0f) * control;
  return v < 0.0f ? 0.0f : v;  // keeps NaN and -0.0 like torch.clamp(min=0)
}

__global__ void __launch_bounds__(128) duration_scan_kernel(const float* __restrict__ log_d, float control,
                                                            const float* __restrict__ control_vec,
                                                            float* __restrict__ d_rounded,
                                                            int32_t* __restrict__ ends,
                                                            int32_t* __restrict__ mel_len_expand,
                                                            int32_t* __restrict__ mel_len_sum, int B, int L) {
  const int b = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (b >= B) return;
  const int lane = threadIdx.x & 31;
  const float* ld = log_d ? log_d + (int64_t)b * L : nullptr;
  const float* cv = control_vec ? control_vec + (int64_t)b * L : nullptr;  // per-token d_control (modules.py:137-140)
  float* dr = d_rounded + (int64_t)b * L;
  int32_t* eb = ends + (int64_t)b * L;
  int carry = 0;
  for (int base = 0; base < L; base += 32) {
    const int l = base + lane;
    int reps = 0;
    if (l < L) {
      float d;
      if (ld) {
        d = round_duration(ld[l], cv ? cv[l] : control);
        dr[l] = d;
      } else {
        d = dr[l];
      }
      reps = max((int)d, 0);  // int(): truncation toward zero
    }
    int incl = reps;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int n = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += n;
    }
    if (l < L) eb[l] = carry + incl;
    carry += __shfl_sync(0xffffffffu, incl, 31);
  }
  if (lane == 0) {
    if (mel_len_expand) mel_len_expand[b] = carry;
    if (mel_len_sum) {
      // float32 running sum in index order, then .int() (modules.py:143)
      float s = 0.0f;
      for (int l = 0; l < L; ++l) s += ld ? round_duration(ld[l], cv ? cv[l] : control) : dr[l];
      mel_len_sum[b] = (int32_t)s;
    }
  }
}

constexpr int LR_WARPS = 8;
constexpr int LR_FRAMES_PER_WARP = 8;
constexpr int LR_FRAMES_PER_CTA = LR_WARPS * LR_FRAMES_PER_WARP;

__global__ void __launch_bounds__(LR_WARPS * 32) length_regulate_kernel(const float* __restrict__ x,
                                                                        const int32_t* __restrict__ ends,
                                                                        const float* __restrict__ pe,
                                                                        float* __restrict__ out,
                                                                        int32_t* __restrict__ mel2phone, int L,
                                                                        int T, int C) {
  extern __shared__ int32_t s_ends[];
  const int b = blockIdx.y;
  const int32_t* eb = ends + (int64_t)b * L;
  for (int i = threadIdx.x; i < L; i += blockDim.x) s_ends[i] = eb[i];
  __syncthreads();
  const int total = L > 0 ? s_ends[L - 1] : 0;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int tbase = blockIdx.x * LR_FRAMES_PER_CTA + warp * LR_FRAMES_PER_WARP;

  // lane f (< 8) finds the source row of frame tbase+f; the eight searches run side by side instead of
  // one after the other, then every lane picks the results up by shuffle
  int my_idx = -1;
  {
    const int t = tbase + lane;
    if (lane < LR_FRAMES_PER_WARP && t < T && t < total) {
      int lo = 0, hi = L;  // number of prefix sums <= t  (upper bound)
      while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (s_ends[mid] <= t) lo = mid + 1; else hi = mid;
      }
      my_idx = lo;
    }
    if (mel2phone && lane < LR_FRAMES_PER_WARP && t < T) mel2phone[(int64_t)b * T + t] = my_idx;
  }
  int src[LR_FRAMES_PER_WARP];
#pragma unroll
  for (int f = 0; f < LR_FRAMES_PER_WARP; ++f) src[f] = __shfl_sync(0xffffffffu, my_idx, f);

  const int C4 = C >> 2;
  const float4* xb = reinterpret_cast<const float4*>(x + (int64_t)b * L * C);
  for (int c = lane; c < C4; c += 32) {
    float4 v[LR_FRAMES_PER_WARP];
#pragma unroll
    for (int f = 0; f < LR_FRAMES_PER_WARP; ++f) {
      v[f] = make_float4(0.f, 0.f, 0.f, 0.f);
      if (src[f] >= 0) v[f] = __ldg(xb + (int64_t)src[f] * C4 + c);
    }
    if (pe) {
#pragma unroll
      for (int f = 0; f < LR_FRAMES_PER_WARP; ++f) {
        const int t = tbase + f;
        if (t < T) {
          const float4 q = __ldg(reinterpret_cast<const float4*>(pe + (int64_t)t * C) + c);
          v[f].x += q.x; v[f].y += q.y; v[f].z += q.z; v[f].w += q.w;
        }
      }
    }
#pragma unroll
    for (int f = 0; f < LR_FRAMES_PER_WARP; ++f) {
      const int t = tbase + f;
      if (t < T) ss::st_stream(reinterpret_cast<float4*>(out + ((int64_t)b * T + t) * C) + c, v[f]);
    }
  }
}

}  // namespace

extern "C" {

int ss_duration_scan(const float* log_d, float d_control, float* d_rounded, int32_t* ends, int32_t* mel_len_expand,
                     int32_t* mel_len_sum, int B, int L, void* stream) {
  SS_REQUIRE(d_rounded && ends, SS_ERR_INVALID, "ss_duration_scan: null pointer");
  SS_REQUIRE(B > 0 && L > 0, SS_ERR_INVALID, "ss_duration_scan: bad shape B=%d L=%d", B, L);
  const int warps = 4;
  duration_scan_kernel<<<(B + warps - 1) / warps, warps * 32, 0, (cudaStream_t)stream>>>(
      log_d, d_control, nullptr, d_rounded, ends, mel_len_expand, mel_len_sum, B, L);
  SS_CHECK_LAUNCH("ss_duration_scan");
  return SS_OK;
}

int ss_duration_scan_v(const float* log_d, const float* d_control_vec, float* d_rounded, int32_t* ends,
                       int32_t* mel_len_expand, int32_t* mel_len_sum, int B, int L, void* stream) {
  SS_REQUIRE(log_d && d_control_vec && d_rounded && ends, SS_ERR_INVALID, "ss_duration_scan_v: null pointer");
  SS_REQUIRE(B > 0 && L > 0, SS_ERR_INVALID, "ss_duration_scan_v: bad shape B=%d L=%d", B, L);
  const int warps = 4;
  duration_scan_kernel<<<(B + warps - 1) / warps, warps * 32, 0, (cudaStream_t)stream>>>(
      log_d, 1.0f, d_control_vec, d_rounded, ends, mel_len_expand, mel_len_sum, B, L);
  SS_CHECK_LAUNCH("ss_duration_scan_v");
  return SS_OK;
}

int ss_length_regulate(const float* x, const int32_t* ends, const float* pe, float* out, int32_t* mel2phone, int B,
                       int L, int T, int C, void* stream) {
  SS_REQUIRE(x && ends && out, SS_ERR_INVALID, "ss_length_regulate: null pointer");
  SS_REQUIRE(B > 0 && L > 0 && T >= 0 && C > 0 && C % 4 == 0, SS_ERR_INVALID,
             "ss_length_regulate: bad shape B=%d L=%d T=%d C=%d", B, L, T, C);
  SS_REQUIRE(B <= 65535, SS_ERR_UNSUPPORTED, "ss_length_regulate: B > 65535");
  SS_REQUIRE((size_t)L * 4 <= 200 * 1024, SS_ERR_UNSUPPORTED, "ss_length_regulate: L=%d too long for smem", L);
  SS_REQUIRE(ss::aligned16(x) && ss::aligned16(out) && ss::aligned16(pe), SS_ERR_ALIGNMENT,
             "ss_length_regulate: pointers must be 16-byte aligned");
  if (T == 0) return SS_OK;
  const size_t smem = (size_t)L * sizeof(int32_t);
  if (smem > 48 * 1024) {  // very long token sequences only: opt in once per device, up to the 227 KB an SM offers
    SS_REQUIRE(smem <= 227 * 1024, SS_ERR_UNSUPPORTED, "ss_length_regulate: L = %d tokens exceeds the shared-memory table", L);
    static std::atomic<unsigned long long> attr_mask{0};
    int d = 0;
    cudaGetDevice(&d);
    if ((attr_mask.load(std::memory_order_acquire) & (1ull << (d & 63))) == 0) {
      cudaError_t e = cudaFuncSetAttribute(length_regulate_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
      if (e != cudaSuccess) return ss::cuda_fail(e, "ss_length_regulate: smem opt-in");
      attr_mask.fetch_or(1ull << (d & 63), std::memory_order_release);
    }
  }
  dim3 grid((T + LR_FRAMES_PER_CTA - 1) / LR_FRAMES_PER_CTA, B);
  length_regulate_kernel<<<grid, LR_WARPS * 32, smem, (cudaStream_t)stream>>>(x, ends, pe, out, mel2phone, L, T, C);
  SS_CHECK_LAUNCH("ss_length_regulate");
  return SS_OK;
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------
// LengthRegulator for callers that cannot allocate device memory.
//
// The two-step form above needs a [B, L] int32 buffer of prefix sums between its two kernels, and
// the caller has to provide it.  Some callers have nowhere to get it from: a synthesis worker that
// was handed a fixed set of device buffers at start-up and runs inside a latency budget where a
// cudaMalloc (a device-wide synchronising call that can also fail late, once memory is fragmented)
// is not acceptable, or a plug-in hosted by another runtime that owns the allocator and does not
// expose it.  For them the prefix sums live in the array below, which is part of the library image:
// every GPU that loads the library carries these 256 KiB until the process exits.
//
// g_lr_prefix is a DEVICE symbol.  In host code its name is only a handle - passing it to a kernel
// would hand the kernel a host address - so the device address is looked up once per device with
// cudaGetSymbolAddress (below) and that pointer is what the kernels receive.
// ---------------------------------------------------------------------------------------------
namespace {

constexpr int kNoAllocMaxTokens = SS_LR_NOALLOC_MAX_TOKENS;
constexpr int kNoAllocMaxDevices = 64;

static __device__ __align__(256) int32_t g_lr_prefix[kNoAllocMaxTokens];

// Device address of g_lr_prefix on the CURRENT device, resolved on the first call made with that
// device current (each device has its own copy of the array, hence its own address).
int noalloc_prefix_ptr(int32_t** out) {
  static int32_t* cache[kNoAllocMaxDevices] = {};
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return ss::cuda_fail(e, "ss_length_regulate_noalloc: cudaGetDevice");
  SS_REQUIRE(dev >= 0 && dev < kNoAllocMaxDevices, SS_ERR_UNSUPPORTED,
             "ss_length_regulate_noalloc: device ordinal %d is beyond the %d this entry point tracks", dev,
             kNoAllocMaxDevices);
  int32_t* p = __atomic_load_n(&cache[dev], __ATOMIC_ACQUIRE);
  if (p == nullptr) {
    void* sym = nullptr;
    e = cudaGetSymbolAddress(&sym, g_lr_prefix);
    if (e != cudaSuccess) return ss::cuda_fail(e, "ss_length_regulate_noalloc: cudaGetSymbolAddress");
    SS_REQUIRE(sym != nullptr, SS_ERR_CUDA, "ss_length_regulate_noalloc: null symbol address");
    p = static_cast<int32_t*>(sym);
    __atomic_store_n(&cache[dev], p, __ATOMIC_RELEASE);  // racing first calls store the same value
  }
  *out = p;
  return SS_OK;
}

}  // namespace

extern "C" int ss_length_regulate_noalloc(const float* x, const float* durations, const float* pe, float* out,
                                          int32_t* mel_len, int B, int L, int T, int C, void* stream) {
  SS_REQUIRE(x && durations && out && mel_len, SS_ERR_INVALID, "ss_length_regulate_noalloc: null pointer");
  SS_REQUIRE(B > 0 && L > 0 && T >= 0 && C > 0 && C % 4 == 0, SS_ERR_INVALID,
             "ss_length_regulate_noalloc: bad shape B=%d L=%d T=%d C=%d", B, L, T, C);
  SS_REQUIRE((int64_t)B * L <= kNoAllocMaxTokens, SS_ERR_UNSUPPORTED,
             "ss_length_regulate_noalloc: B*L = %lld exceeds the built-in scratch of %d prefix sums; use "
             "ss_duration_scan + ss_length_regulate with a caller-provided buffer",
             (long long)B * L, kNoAllocMaxTokens);
  SS_REQUIRE(B <= 65535, SS_ERR_UNSUPPORTED, "ss_length_regulate_noalloc: B > 65535");
  SS_REQUIRE((size_t)L * 4 <= 48 * 1024, SS_ERR_UNSUPPORTED, "ss_length_regulate_noalloc: L=%d too long", L);
  SS_REQUIRE(ss::aligned16(x) && ss::aligned16(out) && ss::aligned16(pe), SS_ERR_ALIGNMENT,
             "ss_length_regulate_noalloc: pointers must be 16-byte aligned");
  int32_t* prefix = nullptr;
  const int rc = noalloc_prefix_ptr(&prefix);
  if (rc != SS_OK) return rc;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  // Every prefix sum this call reads ([0, B*L)) is written by the scan first, so nothing in the
  // array has to be cleared; whatever an earlier call left beyond B*L is never looked at.
  const int warps = 4;
  duration_scan_kernel<<<(B + warps - 1) / warps, warps * 32, 0, st>>>(nullptr, 1.0f, nullptr, const_cast<float*>(durations),
                                                                      prefix, mel_len, nullptr, B, L);
  SS_CHECK_LAUNCH("ss_length_regulate_noalloc (scan)");
  if (T == 0) return SS_OK;
  dim3 grid((T + LR_FRAMES_PER_CTA - 1) / LR_FRAMES_PER_CTA, B);
  length_regulate_kernel<<<grid, LR_WARPS * 32, (size_t)L * sizeof(int32_t), st>>>(x, prefix, pe, out, nullptr, L, T, C);
  SS_CHECK_LAUNCH("ss_length_regulate_noalloc (expand)");
  return SS_OK;
}
