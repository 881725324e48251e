// Fused masked-softmax self-attention on the tensor cores for head size 32 (the ReferenceEncoder's
// 8 x 32 heads), fp32-accurate: both GEMMs are 3xTF32 error-compensated tcgen05.mma with fp32
// accumulation in TMEM; softmax statistics, probabilities and the output accumulator are fp32.
//
// One CTA = 128 queries of one head of one utterance; keys/values stream through in 64-key tiles up to
// the utterance's valid length.  Roles (320 threads):
//   warp 0      TMA: Q once, then K_j / V_j tiles into a 2-stage ring (rows outside [0,S) zero-filled)
//   warp 1      one thread issues  S_j = Q.K_j^T  (12 MMAs, N = 64) and  O_j = P_j.V_j  (24 MMAs, N = 32);
//               S_{j+1} is issued before O_j so it overlaps the softmax of tile j
//   warps 2-5   softmax: thread = query row.  tcgen05.ld S_j -> scale by log2(e)/sqrt(dk) -> mask keys >= len
//               -> online max/sum in the base-2 domain -> P_j = 2^(t - m) (MUFU.EX2), split into TF32
//               hi/lo and written straight into the 128-byte-swizzled K-major layout the MMA reads;
//               then  o = o*alpha_j + O_j  in registers once O_j lands (fp32 round-to-nearest adds, so the
//               truncating TMEM accumulation never runs longer than one tile)
//   warps 6-9   split Q, K_j, V_j into TF32 hi/lo in shared memory (element-wise, layout-agnostic)
// V arrives [keys x 32] (head dims contiguous) but the MMA wants its B operand K-major, i.e. V^T with the
// keys contiguous: the split warps transpose the 8 KB tile while they split it (conflict-free: a lane
// reads one head dim of four consecutive keys and writes one 16-byte unit of the swizzled V^T row).
// The [S x S] score matrix never leaves the SM.
#include "tc_common.cuh"

namespace {

using namespace sstc;

constexpr int DK = 32;            // head size = one 128-byte row of fp32
constexpr int BQ = 128;           // queries per CTA = TMEM lanes
constexpr int BKV = 64;           // keys per tile
constexpr int NTHREADS = 320;
constexpr int Q_BYTES = BQ * 128;      // 16 KB
constexpr int KV_BYTES = BKV * 128;    // 8 KB
constexpr int STAGE_BYTES = 5 * KV_BYTES;  // K_hi, K_lo, V (as loaded, unswizzled), V^T_hi, V^T_lo
constexpr int P_BYTES = 2 * Q_BYTES;   // 2 chunks of 32 keys x 128 rows
constexpr int SMEM_DATA = 2 * Q_BYTES + 2 * STAGE_BYTES + 2 * P_BYTES;  // 32 + 80 + 64 KB
constexpr int N_BARS = 2 + 6 + 4 + 2 + 4;
constexpr int SMEM_BYTES = SMEM_DATA + N_BARS * 8 + 16 + 1024;
constexpr int TMEM_COLS = 256;  // S0,S1: 2 x 64; O0,O1: 2 x 32

__device__ __forceinline__ float tf32_rn(float x) {
  return __uint_as_float((__float_as_uint(x) + 0x1000u) & 0xFFFFE000u);
}

__device__ __forceinline__ void split_tile(float4* hi, float4* lo, int n_vec, int tt) {
  for (int idx = tt; idx < n_vec; idx += 128) {
    const float4 x = hi[idx];
    const float4 h = make_float4(tf32_rn(x.x), tf32_rn(x.y), tf32_rn(x.z), tf32_rn(x.w));
    hi[idx] = h;
    lo[idx] = make_float4(tf32_rn(x.x - h.x), tf32_rn(x.y - h.y), tf32_rn(x.z - h.z), tf32_rn(x.w - h.w));
  }
}

__global__ void __launch_bounds__(NTHREADS, 1)
attention_tc_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmKV,
                    const __grid_constant__ CUtensorMap tmV, const int32_t* __restrict__ lens, float* __restrict__ ctx, int S, int H, float temperature) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int b = blockIdx.z, h = blockIdx.y, q0 = blockIdx.x * BQ;
  const int D = H * DK;
  int len = lens ? lens[b] : S;
  len = min(max(len, 0), S);
  float* out = ctx + (int64_t)b * S * D + h * DK;

  if (q0 >= len) {  // whole tile is padding: the caller's mask zeroes these rows anyway
    const int rows = min(BQ, S - q0);
    for (int f = threadIdx.x; f < rows * (DK / 4); f += NTHREADS)
      reinterpret_cast<float4*>(out + (int64_t)(q0 + f / (DK / 4)) * D)[f % (DK / 4)] = make_float4(0.f, 0.f, 0.f, 0.f);
    return;
  }

  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem0 = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* smem_al = smem_raw + (smem0 - smem_u32(smem_raw));
  constexpr int OFF_QH = 0, OFF_QL = Q_BYTES, OFF_ST = 2 * Q_BYTES, OFF_PH = OFF_ST + 2 * STAGE_BYTES,
                OFF_PL = OFF_PH + P_BYTES;
  const uint32_t bars = smem0 + SMEM_DATA;
  const uint32_t q_full = bars, q_xf = bars + 8;
  const uint32_t kv_full0 = bars + 16, kv_xf0 = bars + 32, kv_empty0 = bars + 48;
  const uint32_t s_full0 = bars + 64, s_empty0 = bars + 80;
  const uint32_t p_full = bars + 96, p_empty = bars + 104;
  const uint32_t o_full0 = bars + 112, o_empty0 = bars + 128;
  const uint32_t tslot = bars + N_BARS * 8;
  volatile uint32_t* tslot_ptr = reinterpret_cast<volatile uint32_t*>(smem_al + SMEM_DATA + N_BARS * 8);

  const int NT = (len + BKV - 1) / BKV;

  if (threadIdx.x == 0) {
    prefetch_tmap(&tmQ);
    prefetch_tmap(&tmKV);
    prefetch_tmap(&tmV);
    mbar_init(q_full, 1);
    mbar_init(q_xf, 4);
    for (int s = 0; s < 2; ++s) {
      mbar_init(kv_full0 + 8 * s, 1);
      mbar_init(kv_xf0 + 8 * s, 4);
      mbar_init(kv_empty0 + 8 * s, 1);
      mbar_init(s_full0 + 8 * s, 1);
      mbar_init(s_empty0 + 8 * s, 4);
      mbar_init(o_full0 + 8 * s, 1);
      mbar_init(o_empty0 + 8 * s, 4);
    }
    mbar_init(p_full, 4);
    mbar_init(p_empty, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) tmem_alloc(tslot, TMEM_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tslot_ptr;

  if (warp == 0) {
    // ================= TMA producer =================
    if (lane == 0) {
      mbar_expect_tx(q_full, Q_BYTES);
      tma_load_3d(smem0 + OFF_QH, &tmQ, q_full, h * DK, q0, b);
      for (int j = 0; j < NT; ++j) {
        const int s = j & 1;
        mbar_wait(kv_empty0 + 8 * s, ((j >> 1) & 1) ^ 1);
        mbar_expect_tx(kv_full0 + 8 * s, 2 * KV_BYTES);
        const uint32_t st = smem0 + OFF_ST + s * STAGE_BYTES;
        tma_load_3d(st, &tmKV, kv_full0 + 8 * s, D + h * DK, j * BKV, b);                    // K_j -> K_hi slot
        tma_load_3d(st + 2 * KV_BYTES, &tmV, kv_full0 + 8 * s, 2 * D + h * DK, j * BKV, b);   // V_j, unswizzled
      }
    }
  } else if (warp == 1) {
    // ================= MMA issuer =================
    if (lane == 0) {
      const uint32_t idesc_s = make_idesc(2u, BKV);               // tf32, N = 64, A and B K-major
      const uint32_t idesc_o = make_idesc(2u, DK);                // tf32, N = 32, B = V^T K-major
      const uint32_t qh = smem0 + OFF_QH, ql = smem0 + OFF_QL;
      auto issue_s = [&](int j) {
        const int s = j & 1;
        mbar_wait(kv_xf0 + 8 * s, (j >> 1) & 1);
        mbar_wait(s_empty0 + 8 * s, ((j >> 1) & 1) ^ 1);
        tc_fence_after();
        const uint32_t kh = smem0 + OFF_ST + s * STAGE_BYTES, kl = kh + KV_BYTES;
        const uint32_t d = tmem_base + s * BKV;
#pragma unroll
        for (int kk = 0; kk < DK / 8; ++kk) {
          const uint32_t o = kk * 32;
          umma_tf32(d, make_desc(ql + o), make_desc(kh + o), idesc_s, kk != 0);
          umma_tf32(d, make_desc(qh + o), make_desc(kl + o), idesc_s, 1u);
          umma_tf32(d, make_desc(qh + o), make_desc(kh + o), idesc_s, 1u);
        }
        umma_commit(s_full0 + 8 * s);
      };
      mbar_wait(q_xf, 0);
      issue_s(0);
      for (int j = 0; j < NT; ++j) {
        if (j + 1 < NT) issue_s(j + 1);
        const int s = j & 1;
        mbar_wait(p_full, j & 1);
        mbar_wait(o_empty0 + 8 * s, ((j >> 1) & 1) ^ 1);
        tc_fence_after();
        const uint32_t vh = smem0 + OFF_ST + s * STAGE_BYTES + 3 * KV_BYTES, vl = vh + KV_BYTES;  // V^T hi / lo
        const uint32_t d = tmem_base + 2 * BKV + s * DK;
#pragma unroll
        for (int ks = 0; ks < BKV / 8; ++ks) {  // 8 keys per instruction
          const uint32_t pa = (ks >> 2) * Q_BYTES + (ks & 3) * 32;  // chunk of 32 keys, then 32 bytes per step
          const uint32_t vb = (ks >> 2) * (DK * 128) + (ks & 3) * 32;  // V^T: 32 rows x 128 B per 32-key chunk
          umma_tf32(d, make_desc(smem0 + OFF_PL + pa), make_desc(vh + vb), idesc_o, ks != 0);
          umma_tf32(d, make_desc(smem0 + OFF_PH + pa), make_desc(vl + vb), idesc_o, 1u);
          umma_tf32(d, make_desc(smem0 + OFF_PH + pa), make_desc(vh + vb), idesc_o, 1u);
        }
        umma_commit(o_full0 + 8 * s);
        umma_commit(p_empty);
        umma_commit(kv_empty0 + 8 * s);
      }
    }
  } else if (warp < 6) {
    // ================= softmax + output accumulation (thread = query row) =================
    const int lg = warp & 3;
    const int r = lg * 32 + lane;
    const uint32_t lane_base = tmem_base + ((uint32_t)(lg * 32) << 16);
    float m = -INFINITY, l = 0.f, alpha_prev = 1.f;
    const float scale_log2 = 1.4426950408889634f / temperature;
    float o[DK];
#pragma unroll
    for (int i = 0; i < DK; ++i) o[i] = 0.f;
    uint8_t* ph_row = smem_al + OFF_PH + r * 128;
    uint8_t* pl_row = smem_al + OFF_PL + r * 128;
    const int sw = r & 7;

    for (int j = 0; j < NT; ++j) {
      const int sb = j & 1;
      if (j > 0) {  // fold in O_{j-1} (its P.V ran while we were busy with S_j's predecessor)
        const int ob = (j - 1) & 1;
        mbar_wait(o_full0 + 8 * ob, ((j - 1) >> 1) & 1);
        tc_fence_after();
        float v[32];
        tmem_ld32(lane_base + 2 * BKV + ob * DK, v);
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(o_empty0 + 8 * ob);
#pragma unroll
        for (int i = 0; i < DK; ++i) o[i] = fmaf(o[i], alpha_prev, v[i]);
      }
      mbar_wait(s_full0 + 8 * sb, (j >> 1) & 1);
      tc_fence_after();
      float s0[32], s1[32];
      tmem_ld32(lane_base + sb * BKV, s0);
      tmem_ld32(lane_base + sb * BKV + 32, s1);
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(s_empty0 + 8 * sb);

      // softmax in the base-2 domain: t = s * log2(e)/sqrt(dk); p = 2^(t - m).  One FMUL and one MUFU.EX2 per
      // score (the four softmax warps are the busiest part of this kernel; libm expf and a true division
      // cost ~10x more issue slots).  Relative error of p <= ~3e-7, far inside the kernel's 1e-6 class.
      const int kbase = j * BKV;
      float mx = -INFINITY;
#pragma unroll
      for (int i = 0; i < 32; ++i) {
        s0[i] = (kbase + i < len) ? s0[i] * scale_log2 : -INFINITY;
        s1[i] = (kbase + 32 + i < len) ? s1[i] * scale_log2 : -INFINITY;
        mx = fmaxf(mx, fmaxf(s0[i], s1[i]));
      }
      const float m_new = fmaxf(m, mx);  // finite: every tile holds at least one valid key
      const float alpha = exp2f(m - m_new);
      float sum = 0.f;
#pragma unroll
      for (int i = 0; i < 32; ++i) {
        s0[i] = exp2f(s0[i] - m_new);
        s1[i] = exp2f(s1[i] - m_new);
        sum += s0[i] + s1[i];
      }
      l = l * alpha + sum;
      m = m_new;
      alpha_prev = alpha;

      mbar_wait(p_empty, (j & 1) ^ 1);  // P.V of the previous tile has finished reading P
#pragma unroll
      for (int u = 0; u < 8; ++u) {  // 16-byte unit u of this row lives at unit (u ^ (row & 7)) - SWIZZLE_128B
        const int pu = (u ^ sw) * 16;
        const float4 a = make_float4(s0[4 * u], s0[4 * u + 1], s0[4 * u + 2], s0[4 * u + 3]);
        const float4 c = make_float4(s1[4 * u], s1[4 * u + 1], s1[4 * u + 2], s1[4 * u + 3]);
        const float4 ah = make_float4(tf32_rn(a.x), tf32_rn(a.y), tf32_rn(a.z), tf32_rn(a.w));
        const float4 ch = make_float4(tf32_rn(c.x), tf32_rn(c.y), tf32_rn(c.z), tf32_rn(c.w));
        *reinterpret_cast<float4*>(ph_row + pu) = ah;
        *reinterpret_cast<float4*>(ph_row + Q_BYTES + pu) = ch;
        *reinterpret_cast<float4*>(pl_row + pu) =
            make_float4(tf32_rn(a.x - ah.x), tf32_rn(a.y - ah.y), tf32_rn(a.z - ah.z), tf32_rn(a.w - ah.w));
        *reinterpret_cast<float4*>(pl_row + Q_BYTES + pu) =
            make_float4(tf32_rn(c.x - ch.x), tf32_rn(c.y - ch.y), tf32_rn(c.z - ch.z), tf32_rn(c.w - ch.w));
      }
      fence_proxy_async();
      __syncwarp();
      if (lane == 0) mbar_arrive(p_full);
    }
    {  // last tile's O
      const int ob = (NT - 1) & 1;
      mbar_wait(o_full0 + 8 * ob, ((NT - 1) >> 1) & 1);
      tc_fence_after();
      float v[32];
      tmem_ld32(lane_base + 2 * BKV + ob * DK, v);
#pragma unroll
      for (int i = 0; i < DK; ++i) o[i] = fmaf(o[i], alpha_prev, v[i]);
    }
    const int q = q0 + r;
    if (q < S) {
      const float inv = q < len ? 1.0f / l : 0.f;  // padded query rows are written as zeros
      float4* orow = reinterpret_cast<float4*>(out + (int64_t)q * D);
#pragma unroll
      for (int c = 0; c < DK / 4; ++c)
        orow[c] = make_float4(o[4 * c] * inv, o[4 * c + 1] * inv, o[4 * c + 2] * inv, o[4 * c + 3] * inv);
#ifdef ATT_DEBUG
      orow[0] = make_float4(l, m, o[0], o[1]);
      orow[1] = make_float4(inv, alpha_prev, (float)NT, (float)len);
#endif
    }
  } else {
    // ================= operand split (warps 6..9) =================
    const int tt = (warp - 6) * 32 + lane;
    mbar_wait(q_full, 0);
    split_tile(reinterpret_cast<float4*>(smem_al + OFF_QH), reinterpret_cast<float4*>(smem_al + OFF_QL), Q_BYTES / 16, tt);
    fence_proxy_async();
    __syncwarp();
    if (lane == 0) mbar_arrive(q_xf);
    for (int j = 0; j < NT; ++j) {
      const int s = j & 1;
      mbar_wait(kv_full0 + 8 * s, (j >> 1) & 1);
      uint8_t* st = smem_al + OFF_ST + s * STAGE_BYTES;
      split_tile(reinterpret_cast<float4*>(st), reinterpret_cast<float4*>(st + KV_BYTES), KV_BYTES / 16, tt);
      {  // V [64 keys x 32 dims] -> V^T hi / lo [32 dims x 64 keys], K-major, 128-byte swizzle
        const float* vraw = reinterpret_cast<const float*>(st + 2 * KV_BYTES);
        const int d = lane, w = warp - 6;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const int kq = w + 4 * i;  // quad of keys 4*kq .. 4*kq+3
          const float x0 = vraw[(4 * kq + 0) * DK + d], x1 = vraw[(4 * kq + 1) * DK + d];
          const float x2 = vraw[(4 * kq + 2) * DK + d], x3 = vraw[(4 * kq + 3) * DK + d];
          const float4 hv = make_float4(tf32_rn(x0), tf32_rn(x1), tf32_rn(x2), tf32_rn(x3));
          const int off = (kq >> 3) * (DK * 128) + d * 128 + (((kq & 7) ^ (d & 7)) << 4);
          *reinterpret_cast<float4*>(st + 3 * KV_BYTES + off) = hv;
          *reinterpret_cast<float4*>(st + 4 * KV_BYTES + off) =
              make_float4(tf32_rn(x0 - hv.x), tf32_rn(x1 - hv.y), tf32_rn(x2 - hv.z), tf32_rn(x3 - hv.w));
        }
      }
      fence_proxy_async();
      __syncwarp();
      if (lane == 0) mbar_arrive(kv_xf0 + 8 * s);
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    __syncwarp();
    tc_fence_after();
    tmem_dealloc(tmem_base, TMEM_COLS);
  }
}

}  // namespace

int ss_attention_tc128_launch(const float* qkv, const int32_t* lens, float* ctx, int B, int S, int H, void* stream);

extern "C" int ss_attention_tc(const float* qkv, const int32_t* lens, float* ctx, int B, int S, int H, int dk,
                               void* stream) {
  SS_REQUIRE(qkv && ctx, SS_ERR_INVALID, "ss_attention_tc: null pointer");
  SS_REQUIRE(B > 0 && S > 0 && H > 0 && B <= 65535 && H <= 65535, SS_ERR_INVALID, "ss_attention_tc: bad shape");
  SS_REQUIRE(dk == DK || dk == 128, SS_ERR_UNSUPPORTED, "ss_attention_tc: dk=%d (kernels exist for dk=32 and dk=128)", dk);
  SS_REQUIRE(ss::aligned16(qkv) && ss::aligned16(ctx), SS_ERR_ALIGNMENT, "ss_attention_tc: 16-byte alignment");
  if (dk == 128) return ss_attention_tc128_launch(qkv, lens, ctx, B, S, H, stream);
  const int D = H * DK;
  CUtensorMap tmQ, tmKV, tmV;
  int rc = make_map_3d(&tmQ, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, qkv, (uint64_t)3 * D, (uint64_t)S, (uint64_t)B,
                       (uint64_t)3 * D * 4, (uint64_t)S * 3 * D * 4, DK, BQ);
  if (rc != SS_OK) return rc;
  rc = make_map_3d(&tmKV, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, qkv, (uint64_t)3 * D, (uint64_t)S, (uint64_t)B,
                   (uint64_t)3 * D * 4, (uint64_t)S * 3 * D * 4, DK, BKV);
  if (rc != SS_OK) return rc;
  rc = make_map_3d(&tmV, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, qkv, (uint64_t)3 * D, (uint64_t)S, (uint64_t)B,
                   (uint64_t)3 * D * 4, (uint64_t)S * 3 * D * 4, DK, BKV, CU_TENSOR_MAP_SWIZZLE_NONE);
  if (rc != SS_OK) return rc;
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(attention_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES);
    if (e != cudaSuccess) return ss::cuda_fail(e, "ss_attention_tc: smem opt-in");
    attr_set = true;
  }
  dim3 grid((S + BQ - 1) / BQ, H, B);
  attention_tc_kernel<<<grid, NTHREADS, SMEM_BYTES, static_cast<cudaStream_t>(stream)>>>(tmQ, tmKV, tmV, lens, ctx, S,
                                                                                           H, sqrtf((float)dk));
  SS_CHECK_LAUNCH("ss_attention_tc");
  return SS_OK;
}
