// Fused masked-softmax self-attention on the tensor cores for head size 128 (Encoder / Decoder: 2 heads
// of 128), fp32-accurate (3xTF32 for both GEMMs, fp32 softmax and output accumulation).
//
// Same scheme as attention_tc.cu (head size 32) with the layout changes the larger head forces:
//   * Q hi / lo (2 x 64 KB) cannot stay in shared memory next to K, V and P, so the split warps put them
//     in TENSOR MEMORY once per CTA (tcgen05.st, thread = query row) and S = Q.K^T uses the TS form of
//     tcgen05.mma (A from TMEM) - 256 of the 512 TMEM columns;
//   * keys come in 32-key tiles: K [32 x 128] is four K-major [32 x 32] chunk tiles, V [32 x 128] is
//     transposed on chip into V^T [128 x 32] (thread = head dim), P [128 x 32] is one swizzle chunk;
//   * O_j = P_j.V_j is a full N = 128 accumulator (single buffer) that the softmax warps fold into 128
//     fp32 registers per row each tile (o = o*alpha + O_j).
// TMEM: S0,S1 (2 x 32) | O (128) | Q_hi (128) | Q_lo (128) = 448 columns.  Shared memory: 2 stages x 80 KB
// (K hi/lo, V as loaded, V^T hi/lo) + P hi/lo 32 KB; the Q staging area aliases the stages.
#include "tc_common.cuh"

namespace {

using namespace sstc;

constexpr int DK = 128, NCH = DK / 32;   // head size, 32-wide chunks of it
constexpr int BQ = 128, BKV = 32;
constexpr int NTHREADS = 320;
constexpr int CH_BYTES = BKV * 128;            // one [32 keys x 32 dims] chunk tile: 4 KB
constexpr int KV_BYTES = NCH * CH_BYTES;       // 16 KB
constexpr int STAGE_BYTES = 5 * KV_BYTES;      // K_hi, K_lo, V_raw, V^T_hi, V^T_lo
constexpr int P_BYTES = BQ * 128;              // 16 KB
constexpr int SMEM_DATA = 2 * STAGE_BYTES + 2 * P_BYTES;  // 160 + 32 KB
constexpr int N_BARS = 2 + 6 + 4 + 2 + 2;
constexpr int SMEM_BYTES = SMEM_DATA + N_BARS * 8 + 16 + 1024;
constexpr int COL_S = 0, COL_O = 2 * BKV, COL_QH = COL_O + DK, COL_QL = COL_QH + DK;
constexpr int TMEM_COLS = 512;
static_assert(COL_QL + DK <= TMEM_COLS, "TMEM budget");
static_assert(NCH * BQ * 128 <= 2 * STAGE_BYTES, "Q staging must fit in the stage area");

__device__ __forceinline__ float tf32_rn(float x) {
  return __uint_as_float((__float_as_uint(x) + 0x1000u) & 0xFFFFE000u);
}

__device__ __forceinline__ void umma_tf32_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t db, uint32_t idesc, uint32_t accum) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n\t"
      "}" ::"r"(tmem_d),
      "r"(tmem_a), "l"(db), "r"(idesc), "r"(accum)
      : "memory");
}

__device__ __forceinline__ void tmem_st32(uint32_t taddr, const float (&v)[32]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
      "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, "
      "%17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};" ::"r"(taddr),
      "r"(__float_as_uint(v[0])), "r"(__float_as_uint(v[1])), "r"(__float_as_uint(v[2])), "r"(__float_as_uint(v[3])),
      "r"(__float_as_uint(v[4])), "r"(__float_as_uint(v[5])), "r"(__float_as_uint(v[6])), "r"(__float_as_uint(v[7])),
      "r"(__float_as_uint(v[8])), "r"(__float_as_uint(v[9])), "r"(__float_as_uint(v[10])), "r"(__float_as_uint(v[11])),
      "r"(__float_as_uint(v[12])), "r"(__float_as_uint(v[13])), "r"(__float_as_uint(v[14])), "r"(__float_as_uint(v[15])),
      "r"(__float_as_uint(v[16])), "r"(__float_as_uint(v[17])), "r"(__float_as_uint(v[18])), "r"(__float_as_uint(v[19])),
      "r"(__float_as_uint(v[20])), "r"(__float_as_uint(v[21])), "r"(__float_as_uint(v[22])), "r"(__float_as_uint(v[23])),
      "r"(__float_as_uint(v[24])), "r"(__float_as_uint(v[25])), "r"(__float_as_uint(v[26])), "r"(__float_as_uint(v[27])),
      "r"(__float_as_uint(v[28])), "r"(__float_as_uint(v[29])), "r"(__float_as_uint(v[30])), "r"(__float_as_uint(v[31]))
      : "memory");
}

__global__ void __launch_bounds__(NTHREADS, 1)
attention_tc128_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                       const __grid_constant__ CUtensorMap tmV, const int32_t* __restrict__ lens,
                       float* __restrict__ ctx, int S, int H, float temperature) {
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
  constexpr int OFF_PH = 2 * STAGE_BYTES, OFF_PL = OFF_PH + P_BYTES;
  const uint32_t bars = smem0 + SMEM_DATA;
  const uint32_t q_full = bars, q_tm = bars + 8;
  const uint32_t kv_full0 = bars + 16, kv_xf0 = bars + 32, kv_empty0 = bars + 48;
  const uint32_t s_full0 = bars + 64, s_empty0 = bars + 80;
  const uint32_t p_full = bars + 96, p_empty = bars + 104;
  const uint32_t o_full = bars + 112, o_empty = bars + 120;
  const uint32_t tslot = bars + N_BARS * 8;
  volatile uint32_t* tslot_ptr = reinterpret_cast<volatile uint32_t*>(smem_al + SMEM_DATA + N_BARS * 8);

  const int NT = (len + BKV - 1) / BKV;

  if (threadIdx.x == 0) {
    prefetch_tmap(&tmQ);
    prefetch_tmap(&tmK);
    prefetch_tmap(&tmV);
    mbar_init(q_full, 1);
    mbar_init(q_tm, 4);
    for (int s = 0; s < 2; ++s) {
      mbar_init(kv_full0 + 8 * s, 1);
      mbar_init(kv_xf0 + 8 * s, 4);
      mbar_init(kv_empty0 + 8 * s, 1);
      mbar_init(s_full0 + 8 * s, 1);
      mbar_init(s_empty0 + 8 * s, 4);
    }
    mbar_init(p_full, 4);
    mbar_init(p_empty, 1);
    mbar_init(o_full, 1);
    mbar_init(o_empty, 4);
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
      mbar_expect_tx(q_full, NCH * BQ * 128);
      for (int c = 0; c < NCH; ++c)  // Q staging [chunk][128 rows x 128 B] over the (still unused) stage area
        tma_load_3d(smem0 + c * (BQ * 128), &tmQ, q_full, h * DK + c * 32, q0, b);
      mbar_wait(q_tm, 0);  // Q now lives in TMEM: the staging area may be overwritten
      for (int j = 0; j < NT; ++j) {
        const int s = j & 1;
        mbar_wait(kv_empty0 + 8 * s, ((j >> 1) & 1) ^ 1);
        mbar_expect_tx(kv_full0 + 8 * s, 2 * KV_BYTES);
        const uint32_t st = smem0 + s * STAGE_BYTES;
        for (int c = 0; c < NCH; ++c) {
          tma_load_3d(st + c * CH_BYTES, &tmK, kv_full0 + 8 * s, D + h * DK + c * 32, j * BKV, b);                    // K chunk
          tma_load_3d(st + 2 * KV_BYTES + c * CH_BYTES, &tmV, kv_full0 + 8 * s, 2 * D + h * DK + c * 32, j * BKV, b);  // V chunk
        }
      }
    }
  } else if (warp == 1) {
    // ================= MMA issuer =================
    if (lane == 0) {
      const uint32_t idesc_s = make_idesc(2u, BKV);  // N = 32 keys
      const uint32_t idesc_o = make_idesc(2u, DK);   // N = 128 head dims
      const uint32_t qh = tmem_base + COL_QH, ql = tmem_base + COL_QL;
      auto issue_s = [&](int j) {
        const int s = j & 1;
        mbar_wait(kv_xf0 + 8 * s, (j >> 1) & 1);
        mbar_wait(s_empty0 + 8 * s, ((j >> 1) & 1) ^ 1);
        tc_fence_after();
        const uint32_t kh = smem0 + s * STAGE_BYTES, kl = kh + KV_BYTES;
        const uint32_t d = tmem_base + COL_S + s * BKV;
#pragma unroll
        for (int ks = 0; ks < DK / 8; ++ks) {  // 16 steps of 8 head dims
          const uint32_t ko = (ks >> 2) * CH_BYTES + (ks & 3) * 32;
          umma_tf32_ts(d, ql + ks * 8, make_desc(kh + ko), idesc_s, ks != 0);
          umma_tf32_ts(d, qh + ks * 8, make_desc(kl + ko), idesc_s, 1u);
          umma_tf32_ts(d, qh + ks * 8, make_desc(kh + ko), idesc_s, 1u);
        }
        umma_commit(s_full0 + 8 * s);
      };
      mbar_wait(q_tm, 0);
      tc_fence_after();
      issue_s(0);
      for (int j = 0; j < NT; ++j) {
        if (j + 1 < NT) issue_s(j + 1);
        const int s = j & 1;
        mbar_wait(p_full, j & 1);
        mbar_wait(o_empty, (j & 1) ^ 1);
        tc_fence_after();
        const uint32_t vh = smem0 + s * STAGE_BYTES + 3 * KV_BYTES, vl = vh + KV_BYTES;  // V^T hi / lo [128 x 32]
        const uint32_t d = tmem_base + COL_O;
#pragma unroll
        for (int kk = 0; kk < BKV / 8; ++kk) {  // 4 steps of 8 keys
          const uint32_t o = kk * 32;
          umma_tf32(d, make_desc(smem0 + OFF_PL + o), make_desc(vh + o), idesc_o, kk != 0);
          umma_tf32(d, make_desc(smem0 + OFF_PH + o), make_desc(vl + o), idesc_o, 1u);
          umma_tf32(d, make_desc(smem0 + OFF_PH + o), make_desc(vh + o), idesc_o, 1u);
        }
        umma_commit(o_full);
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

    auto fold_o = [&](int jj) {  // o = o*alpha + O_jj
      mbar_wait(o_full, jj & 1);
      tc_fence_after();
#pragma unroll
      for (int c = 0; c < NCH; ++c) {
        float v[32];
        tmem_ld32(lane_base + COL_O + c * 32, v);
#pragma unroll
        for (int i = 0; i < 32; ++i) o[c * 32 + i] = fmaf(o[c * 32 + i], alpha_prev, v[i]);
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(o_empty);
    };

    for (int j = 0; j < NT; ++j) {
      const int sb = j & 1;
      if (j > 0) fold_o(j - 1);
      mbar_wait(s_full0 + 8 * sb, (j >> 1) & 1);
      tc_fence_after();
      float s0[32];
      tmem_ld32(lane_base + COL_S + sb * BKV, s0);
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(s_empty0 + 8 * sb);

      const int kbase = j * BKV;
      float mx = -INFINITY;
#pragma unroll
      for (int i = 0; i < 32; ++i) {
        s0[i] = (kbase + i < len) ? s0[i] * scale_log2 : -INFINITY;
        mx = fmaxf(mx, s0[i]);
      }
      const float m_new = fmaxf(m, mx);  // finite: every tile holds at least one valid key
      const float alpha = exp2f(m - m_new);
      float sum = 0.f;
#pragma unroll
      for (int i = 0; i < 32; ++i) {
        s0[i] = exp2f(s0[i] - m_new);
        sum += s0[i];
      }
      l = l * alpha + sum;
      m = m_new;
      alpha_prev = alpha;

      mbar_wait(p_empty, (j & 1) ^ 1);  // P.V of the previous tile has finished reading P
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int pu = (u ^ sw) * 16;
        const float4 a = make_float4(s0[4 * u], s0[4 * u + 1], s0[4 * u + 2], s0[4 * u + 3]);
        const float4 ah = make_float4(tf32_rn(a.x), tf32_rn(a.y), tf32_rn(a.z), tf32_rn(a.w));
        *reinterpret_cast<float4*>(ph_row + pu) = ah;
        *reinterpret_cast<float4*>(pl_row + pu) =
            make_float4(tf32_rn(a.x - ah.x), tf32_rn(a.y - ah.y), tf32_rn(a.z - ah.z), tf32_rn(a.w - ah.w));
      }
      fence_proxy_async();
      __syncwarp();
      if (lane == 0) mbar_arrive(p_full);
    }
    fold_o(NT - 1);
    const int q = q0 + r;
    if (q < S) {
      const float inv = q < len ? 1.0f / l : 0.f;  // padded query rows are written as zeros
      float4* orow = reinterpret_cast<float4*>(out + (int64_t)q * D);
#pragma unroll
      for (int c = 0; c < DK / 4; ++c)
        orow[c] = make_float4(o[4 * c] * inv, o[4 * c + 1] * inv, o[4 * c + 2] * inv, o[4 * c + 3] * inv);
    }
  } else {
    // ================= operand split / transpose (warps 6..9) =================
    const int lg = warp & 3;
    const int r = lg * 32 + lane;         // Q: this thread's query row; V: this thread's head dim
    const int tt = (warp - 6) * 32 + lane;  // linear id for element-wise work
    const int sw = r & 7;
    {  // Q: shared staging (swizzled chunk tiles) -> registers -> TMEM hi / lo, thread = row
      mbar_wait(q_full, 0);
      const uint32_t tq = tmem_base + ((uint32_t)(lg * 32) << 16);
#pragma unroll 1
      for (int c = 0; c < NCH; ++c) {
        const uint8_t* row = smem_al + c * (BQ * 128) + r * 128;
        float hi[32], lo[32];
#pragma unroll
        for (int u = 0; u < 8; ++u) {
          const float4 x = *reinterpret_cast<const float4*>(row + ((u ^ sw) << 4));
          hi[4 * u + 0] = tf32_rn(x.x); lo[4 * u + 0] = tf32_rn(x.x - hi[4 * u + 0]);
          hi[4 * u + 1] = tf32_rn(x.y); lo[4 * u + 1] = tf32_rn(x.y - hi[4 * u + 1]);
          hi[4 * u + 2] = tf32_rn(x.z); lo[4 * u + 2] = tf32_rn(x.z - hi[4 * u + 2]);
          hi[4 * u + 3] = tf32_rn(x.w); lo[4 * u + 3] = tf32_rn(x.w - hi[4 * u + 3]);
        }
        tmem_st32(tq + COL_QH + c * 32, hi);
        tmem_st32(tq + COL_QL + c * 32, lo);
      }
      asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(q_tm);
    }
    for (int j = 0; j < NT; ++j) {
      const int s = j & 1;
      mbar_wait(kv_full0 + 8 * s, (j >> 1) & 1);
      uint8_t* st = smem_al + s * STAGE_BYTES;
      {  // K: element-wise split, hi in place, lo next to it
        float4* hi = reinterpret_cast<float4*>(st);
        float4* lo = reinterpret_cast<float4*>(st + KV_BYTES);
#pragma unroll
        for (int i = 0; i < KV_BYTES / 16 / 128; ++i) {
          const int idx = tt + i * 128;
          const float4 x = hi[idx];
          const float4 hv = make_float4(tf32_rn(x.x), tf32_rn(x.y), tf32_rn(x.z), tf32_rn(x.w));
          hi[idx] = hv;
          lo[idx] = make_float4(tf32_rn(x.x - hv.x), tf32_rn(x.y - hv.y), tf32_rn(x.z - hv.z), tf32_rn(x.w - hv.w));
        }
      }
      {  // V [32 keys x 128 dims] (chunked [c][key][32]) -> V^T hi / lo [128 dims x 32 keys], swizzled rows
        const float* vraw = reinterpret_cast<const float*>(st + 2 * KV_BYTES) + (r >> 5) * (CH_BYTES / 4) + (r & 31);
        uint8_t* th = st + 3 * KV_BYTES + r * 128;
        uint8_t* tl = st + 4 * KV_BYTES + r * 128;
#pragma unroll
        for (int u = 0; u < 8; ++u) {  // keys 4u .. 4u+3
          const float x0 = vraw[(4 * u + 0) * 32], x1 = vraw[(4 * u + 1) * 32];
          const float x2 = vraw[(4 * u + 2) * 32], x3 = vraw[(4 * u + 3) * 32];
          const float4 hv = make_float4(tf32_rn(x0), tf32_rn(x1), tf32_rn(x2), tf32_rn(x3));
          const int pu = (u ^ sw) << 4;
          *reinterpret_cast<float4*>(th + pu) = hv;
          *reinterpret_cast<float4*>(tl + pu) =
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

// dispatched from ss_attention_tc (attention_tc.cu) for dk == 128
int ss_attention_tc128_launch(const float* qkv, const int32_t* lens, float* ctx, int B, int S, int H, void* stream) {
  const int D = H * DK;
  CUtensorMap tmQ, tmK, tmV;
  int rc = make_map_3d(&tmQ, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, qkv, (uint64_t)3 * D, (uint64_t)S, (uint64_t)B,
                       (uint64_t)3 * D * 4, (uint64_t)S * 3 * D * 4, 32, BQ);
  if (rc != SS_OK) return rc;
  rc = make_map_3d(&tmK, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, qkv, (uint64_t)3 * D, (uint64_t)S, (uint64_t)B,
                   (uint64_t)3 * D * 4, (uint64_t)S * 3 * D * 4, 32, BKV);
  if (rc != SS_OK) return rc;
  rc = make_map_3d(&tmV, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, qkv, (uint64_t)3 * D, (uint64_t)S, (uint64_t)B,
                   (uint64_t)3 * D * 4, (uint64_t)S * 3 * D * 4, 32, BKV, CU_TENSOR_MAP_SWIZZLE_NONE);
  if (rc != SS_OK) return rc;
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(attention_tc128_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES);
    if (e != cudaSuccess) return ss::cuda_fail(e, "ss_attention_tc: smem opt-in (dk=128)");
    attr_set = true;
  }
  dim3 grid((S + BQ - 1) / BQ, H, B);
  attention_tc128_kernel<<<grid, NTHREADS, SMEM_BYTES, static_cast<cudaStream_t>(stream)>>>(tmQ, tmK, tmV, lens, ctx, S,
                                                                                            H, sqrtf((float)DK));
  SS_CHECK_LAUNCH("ss_attention_tc (dk=128)");
  return SS_OK;
}
