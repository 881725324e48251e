// Fused masked-softmax self-attention, fp32 FFMA, flash-style (scores and probabilities live in
// registers / shared memory only; the reference materialises [H*B, S, S] in HBM - 1.3 GB per layer
// at B=64, S=800, H=8 - plus three permuted copies of q, k, v).
//
// CTA = (64 queries) x (one head) x (one utterance); the key loop walks 64-key tiles up to the
// utterance's valid length, so padded keys are never loaded and padded query tiles exit at once.
// Thread (ty, tx) of a 16x16 grid owns a 4x4 block of the score tile and a 4 x DK/16 block of the
// output tile for the same four query rows, so the online-softmax rescale never leaves registers.
#include <math.h>

#include "common.cuh"

namespace {

constexpr int BQ = 64, BKV = 64, PQ = BQ + 4, NT = 256;

template <int DK>
struct Smem {
  static constexpr int q = 0;
  static constexpr int k = q + DK * PQ;
  static constexpr int v = k + DK * PQ;
  static constexpr int p = v + BKV * DK;
  static constexpr int floats = p + BKV * PQ;
};

// tile [64 rows][DK] from global (row stride ld) -> shared, transposed to [DK][PQ]
template <int DK>
__device__ __forceinline__ void load_transposed(float* dst, const float* __restrict__ src, int64_t ld, int row0,
                                                int row_end, int tid) {
  const int r = tid >> 2;
  const int dq = (tid & 3) << 2;
#pragma unroll
  for (int it = 0; it < DK / 16; ++it) {
    const int d = it * 16 + dq;
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (row0 + r < row_end) v = __ldg(reinterpret_cast<const float4*>(src + (int64_t)(row0 + r) * ld + d));
    dst[(d + 0) * PQ + r] = v.x;
    dst[(d + 1) * PQ + r] = v.y;
    dst[(d + 2) * PQ + r] = v.z;
    dst[(d + 3) * PQ + r] = v.w;
  }
}

template <int DK>
__global__ void __launch_bounds__(NT) attention_f32_kernel(const float* __restrict__ qkv,
                                                           const int32_t* __restrict__ lens,
                                                           float* __restrict__ ctx, int S, int H,
                                                           float temperature) {
  extern __shared__ __align__(16) float smem[];
  float* Qs = smem + Smem<DK>::q;
  float* Ks = smem + Smem<DK>::k;
  float* Vs = smem + Smem<DK>::v;
  float* Pt = smem + Smem<DK>::p;

  constexpr int CPT = DK / 16;  // output columns per thread
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  const int b = blockIdx.z, h = blockIdx.y, q0 = blockIdx.x * BQ;
  const int D = H * DK;
  const int64_t ld = 3 * (int64_t)D;
  int len = lens ? lens[b] : S;
  len = min(max(len, 0), S);
  const float* base = qkv + (int64_t)b * S * ld + h * DK;
  float* out = ctx + (int64_t)b * S * D + h * DK;
  const int q_end = min(q0 + BQ, S);

  if (q0 >= len) {  // whole tile is padding: the caller's mask zeroes these rows anyway
    for (int f = tid; f < (q_end - q0) * (DK / 4); f += NT) {
      const int r = f / (DK / 4), c = f - r * (DK / 4);
      reinterpret_cast<float4*>(out + (int64_t)(q0 + r) * D)[c] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
    return;
  }

  load_transposed<DK>(Qs, base, ld, q0, q_end, tid);

  float m[4], l[4], o[4][CPT];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    m[i] = -INFINITY;
    l[i] = 0.f;
#pragma unroll
    for (int c = 0; c < CPT; ++c) o[i][c] = 0.f;
  }

  for (int k0 = 0; k0 < len; k0 += BKV) {
    __syncthreads();  // previous tile's readers are done with Ks / Vs / Pt
    load_transposed<DK>(Ks, base + D, ld, k0, len, tid);
    for (int f = tid; f < BKV * (DK / 4); f += NT) {
      const int r = f / (DK / 4), c = f - r * (DK / 4);
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      if (k0 + r < len) v = __ldg(reinterpret_cast<const float4*>(base + 2 * D + (int64_t)(k0 + r) * ld) + c);
      reinterpret_cast<float4*>(Vs + r * DK)[c] = v;
    }
    __syncthreads();

    // ---- scores: 4 queries x 4 keys per thread ----
    float s[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) s[i][j] = 0.f;
#pragma unroll 8
    for (int d = 0; d < DK; ++d) {
      const float4 qv = *reinterpret_cast<const float4*>(Qs + d * PQ + ty * 4);
      const float4 kv = *reinterpret_cast<const float4*>(Ks + d * PQ + tx * 4);
      const float qa[4] = {qv.x, qv.y, qv.z, qv.w};
      const float ka[4] = {kv.x, kv.y, kv.z, kv.w};
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) s[i][j] = fmaf(qa[i], ka[j], s[i][j]);
    }

    // ---- online softmax (row statistics shared by the 16 lanes with the same ty) ----
    float alpha[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      float mx = -INFINITY;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        s[i][j] = (k0 + tx * 4 + j < len) ? s[i][j] / temperature : -INFINITY;
        mx = fmaxf(mx, s[i][j]);
      }
#pragma unroll
      for (int off = 8; off > 0; off >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, off));
      const float m_new = fmaxf(m[i], mx);  // finite: key 0 is always valid when len >= 1
      alpha[i] = expf(m[i] - m_new);
      float sum = 0.f;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        s[i][j] = expf(s[i][j] - m_new);
        sum += s[i][j];
      }
#pragma unroll
      for (int off = 8; off > 0; off >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, off);
      l[i] = l[i] * alpha[i] + sum;
      m[i] = m_new;
    }
#pragma unroll
    for (int j = 0; j < 4; ++j)
      *reinterpret_cast<float4*>(Pt + (tx * 4 + j) * PQ + ty * 4) = make_float4(s[0][j], s[1][j], s[2][j], s[3][j]);
    __syncthreads();

    // ---- output: o = o*alpha + P.V ----
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int c = 0; c < CPT; ++c) o[i][c] *= alpha[i];
    const int kmax = min(BKV, len - k0);
    for (int k = 0; k < kmax; ++k) {
      const float4 pv = *reinterpret_cast<const float4*>(Pt + k * PQ + ty * 4);
      const float pa[4] = {pv.x, pv.y, pv.z, pv.w};
      float va[CPT];
      if constexpr (CPT >= 4) {
#pragma unroll
        for (int c4 = 0; c4 < CPT / 4; ++c4) {
          const float4 vv = *reinterpret_cast<const float4*>(Vs + k * DK + (tx + 16 * c4) * 4);
          va[c4 * 4 + 0] = vv.x; va[c4 * 4 + 1] = vv.y; va[c4 * 4 + 2] = vv.z; va[c4 * 4 + 3] = vv.w;
        }
      } else {
        const float2 vv = *reinterpret_cast<const float2*>(Vs + k * DK + tx * 2);
        va[0] = vv.x; va[1] = vv.y;
      }
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int c = 0; c < CPT; ++c) o[i][c] = fmaf(pa[i], va[c], o[i][c]);
    }
  }

  // ---- normalise and store ----
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int q = q0 + ty * 4 + i;
    if (q >= S) continue;
    const bool pad = q >= len;
    const float inv = pad ? 0.f : 1.0f / l[i];
    float* orow = out + (int64_t)q * D;
    if constexpr (CPT >= 4) {
#pragma unroll
      for (int c4 = 0; c4 < CPT / 4; ++c4)
        reinterpret_cast<float4*>(orow)[tx + 16 * c4] =
            make_float4(o[i][c4 * 4 + 0] * inv, o[i][c4 * 4 + 1] * inv, o[i][c4 * 4 + 2] * inv, o[i][c4 * 4 + 3] * inv);
    } else {
      reinterpret_cast<float2*>(orow)[tx] = make_float2(o[i][0] * inv, o[i][1] * inv);
    }
  }
}

template <int DK>
int launch(const float* qkv, const int32_t* lens, float* ctx, int B, int S, int H, cudaStream_t st) {
  const size_t smem = Smem<DK>::floats * sizeof(float);
  cudaError_t e = cudaFuncSetAttribute(attention_f32_kernel<DK>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return ss::cuda_fail(e, "ss_attention_f32: smem opt-in");
  dim3 grid((S + BQ - 1) / BQ, H, B);
  attention_f32_kernel<DK><<<grid, NT, smem, st>>>(qkv, lens, ctx, S, H, sqrtf((float)DK));
  SS_CHECK_LAUNCH("ss_attention_f32");
  return SS_OK;
}

}  // namespace

extern "C" int ss_attention_f32(const float* qkv, const int32_t* lens, float* ctx, int B, int S, int H, int dk,
                                void* stream) {
  SS_REQUIRE(qkv && ctx, SS_ERR_INVALID, "ss_attention_f32: null pointer");
  SS_REQUIRE(B > 0 && S > 0 && H > 0 && B <= 65535 && H <= 65535, SS_ERR_INVALID, "ss_attention_f32: bad shape");
  SS_REQUIRE(ss::aligned16(qkv) && ss::aligned16(ctx), SS_ERR_ALIGNMENT, "ss_attention_f32: 16-byte alignment");
  cudaStream_t st = (cudaStream_t)stream;
  if (dk == 128) return launch<128>(qkv, lens, ctx, B, S, H, st);
  if (dk == 64) return launch<64>(qkv, lens, ctx, B, S, H, st);
  if (dk == 32) return launch<32>(qkv, lens, ctx, B, S, H, st);
  SS_REQUIRE(false, SS_ERR_UNSUPPORTED, "ss_attention_f32: dk=%d not in {32,64,128}", dk);
}
