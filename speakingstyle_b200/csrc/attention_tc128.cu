// Fused masked-softmax self-attention on the tensor cores for head size 128 (Encoder / Decoder: 2 heads
// of 128), fp32-accurate (3xTF32 for both GEMMs, fp32 softmax and output accumulation).
//
// Same scheme as attention_tc.cu (head size 32) with the layout changes the larger head forces:
//   * Q hi / lo (2 x 64 KB) cannot stay in shared memory next to K, V and P, so they are put in TENSOR
//     MEMORY once per CTA (tcgen05.st, thread = query row) and S = Q.K^T uses the TS form of tcgen05.mma
//     (A from TMEM) - 256 of the 512 TMEM columns;
//   * keys come in 32-key tiles: K [32 x 128] is four K-major [32 x 32] chunk tiles, V [32 x 128] is
//     transposed on chip into V^T [128 x 32]; P [128 x 32] goes to TMEM too (tcgen05.st by the softmax warps, TS-form
//     P.V), which leaves shared memory for a three-deep K ring;
//   * O_j = P_j.V_j is a full N = 128 accumulator (single TMEM buffer) that the softmax warps fold into
//     fp32 registers each tile (o = o*alpha_j + O_j).  EIGHT softmax warps: warps w and w+4 own the same
//     32 query rows; both read the whole S row (same running max), but each exponentiates / writes half
//     of the P row and folds half of the O columns, keeping a partial row sum that is merged at the end.
//   * the exponentials of tile j are computed BEFORE fold(j-1), so they overlap P.V of the previous tile on the
//     tensor pipe; P_j itself is stored after the fold (the single P buffer is free once P.V_{j-1} has retired).
// Rings: K hi/lo three deep (freed by S_j: the loop "S_j retires -> K_{j+3} lands -> split -> S_{j+3}" was the
// critical one with two stages), V as loaded (freed by the transpose) and V^T hi/lo (freed by P.V_j) two deep.
// TMEM: S0,S1 (2 x 32) | O (128) | Q_hi (128) | Q_lo (128) | P hi, lo (64) = 512 columns.  Shared memory 192 KB:
// K 3x32, V 2x16, V^T 2x32 KB; the Q staging area (64 KB) aliases the K ring.
#include "tc_common.cuh"

namespace {

using namespace sstc;

constexpr int DK = 128, NCH = DK / 32;   // head size, 32-wide chunks of it
constexpr int BQ = 128, BKV = 32;
constexpr int N_SOFTMAX_WARPS = 8, N_XF_WARPS = 2;
constexpr int NTHREADS = 32 * (2 + N_SOFTMAX_WARPS + N_XF_WARPS);  // 384: three warps per scheduler -> 168 registers
constexpr int CH_BYTES = BKV * 128;            // one [32 keys x 32 dims] chunk tile: 4 KB
constexpr int KV_BYTES = NCH * CH_BYTES;       // 16 KB
constexpr int NK = 3;                          // K ring depth (the loop S_j retires -> K_{j+NK} lands -> split -> S_{j+NK})
constexpr int OFF_K = 0;                       // [NK][K_hi 16 | K_lo 16]
constexpr int OFF_VR = OFF_K + NK * 2 * KV_BYTES;  // [2][V as loaded 16]
constexpr int OFF_VT = OFF_VR + 2 * KV_BYTES;  // [2][V^T_hi 16 | V^T_lo 16]
constexpr int SMEM_DATA = OFF_VT + 4 * KV_BYTES;  // 192 KB
constexpr int N_BARS = 3 * NK + 4 * 2 + 2 * 2 + 5;
constexpr int SMEM_BYTES = SMEM_DATA + N_BARS * 8 + 16 + 1024;
constexpr int COL_S = 0, COL_O = 2 * BKV, COL_QH = COL_O + DK, COL_QL = COL_QH + DK, COL_P = COL_QL + DK;
constexpr int TMEM_COLS = 512;
static_assert(COL_P + 2 * BKV <= TMEM_COLS, "TMEM budget");
static_assert(NCH * BQ * 128 <= NK * 2 * KV_BYTES, "Q staging must fit in the K ring");
static_assert(SMEM_BYTES <= 232448, "shared memory budget");

// Debug build (-DSS_ATT_TRACE, scripts/attn_trace.py): CTA (0,0,0) stamps clock64() at the pipeline hand-overs.
#ifdef SS_ATT_TRACE
__device__ long long g_att_trace[16 * 64];
#define ATT_TRACE(ev, j)                                                                   \
  do {                                                                                     \
    if ((blockIdx.x | blockIdx.y | blockIdx.z) == 0 && (j) < 64) g_att_trace[(ev) * 64 + (j)] = clock64(); \
  } while (0)
#else
#define ATT_TRACE(ev, j)
#endif

// Only hi is rounded here; lo = x - hi goes to the MMA as is - the tensor core truncates an fp32 operand to TF32
// itself, and lo only carries bits 12-22 of x (same finding as in conv1d_tf32_ts.cu: error unchanged, fewer ALU ops
// on the split / softmax warps that pace these kernels).



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
  const uint32_t bars = smem0 + SMEM_DATA;
  // rings: index with + 8 * slot (K: NK deep; V, V^T, S: two deep)
  const uint32_t k_full0 = bars, k_xf0 = k_full0 + 8 * NK, k_empty0 = k_xf0 + 8 * NK;
  const uint32_t v_full0 = k_empty0 + 8 * NK, vr_empty0 = v_full0 + 16, vt_full0 = vr_empty0 + 16, vt_empty0 = vt_full0 + 16;
  const uint32_t s_full0 = vt_empty0 + 16, s_empty0 = s_full0 + 16;
  const uint32_t p_full = s_empty0 + 16, o_full = p_full + 8, o_empty = o_full + 8, q_full = o_empty + 8, q_tm = q_full + 8;
  const uint32_t tslot = bars + N_BARS * 8;
  volatile uint32_t* tslot_ptr = reinterpret_cast<volatile uint32_t*>(smem_al + SMEM_DATA + N_BARS * 8);

  const int NT = (len + BKV - 1) / BKV;

  if (threadIdx.x == 0) {
    prefetch_tmap(&tmQ);
    prefetch_tmap(&tmK);
    prefetch_tmap(&tmV);
    for (int s = 0; s < NK; ++s) {
      mbar_init(k_full0 + 8 * s, 1);
      mbar_init(k_xf0 + 8 * s, N_SOFTMAX_WARPS);
      mbar_init(k_empty0 + 8 * s, 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(v_full0 + 8 * s, 1);
      mbar_init(vr_empty0 + 8 * s, N_XF_WARPS);
      mbar_init(vt_full0 + 8 * s, N_XF_WARPS);
      mbar_init(vt_empty0 + 8 * s, 1);
      mbar_init(s_full0 + 8 * s, 1);
      mbar_init(s_empty0 + 8 * s, N_SOFTMAX_WARPS);
    }
    mbar_init(p_full, N_SOFTMAX_WARPS);
    mbar_init(o_full, 1);
    mbar_init(o_empty, N_SOFTMAX_WARPS);
    mbar_init(q_full, 1);
    mbar_init(q_tm, N_SOFTMAX_WARPS);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) tmem_alloc(tslot, TMEM_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tslot_ptr;

  if (warp == 0) {
    // ================= TMA producer =================
    if (elect_one()) {  // elect.sync, not lane == 0: lets ptxas issue TMA / MMA without a per-instruction elect loop
      mbar_expect_tx(q_full, NCH * BQ * 128);
      for (int c = 0; c < NCH; ++c)  // Q staging [chunk][128 rows x 128 B] over the (still unused) K ring
        tma_load_3d(smem0 + OFF_K + c * (BQ * 128), &tmQ, q_full, h * DK + c * 32, q0, b);
      for (int j = 0; j < NT; ++j) {
        const int s = j & 1;
        const uint32_t ph = ((j >> 1) & 1) ^ 1;
        mbar_wait(vr_empty0 + 8 * s, ph);
        ATT_TRACE(0, j);
        mbar_expect_tx(v_full0 + 8 * s, KV_BYTES);
        for (int c = 0; c < NCH; ++c)
          tma_load_3d(smem0 + OFF_VR + s * KV_BYTES + c * CH_BYTES, &tmV, v_full0 + 8 * s, 2 * D + h * DK + c * 32,
                      j * BKV, b);
        if (j == 0) mbar_wait(q_tm, 0);  // Q now lives in TMEM: its staging area may be overwritten
        const int ks = j % NK;
        mbar_wait(k_empty0 + 8 * ks, ((j / NK) & 1) ^ 1);
        ATT_TRACE(1, j);
        mbar_expect_tx(k_full0 + 8 * ks, KV_BYTES);
        for (int c = 0; c < NCH; ++c)
          tma_load_3d(smem0 + OFF_K + ks * 2 * KV_BYTES + c * CH_BYTES, &tmK, k_full0 + 8 * ks, D + h * DK + c * 32,
                      j * BKV, b);
      }
    }
  } else if (warp == 1) {
    // ================= MMA issuer =================
    if (elect_one()) {
      const uint32_t idesc_s = make_idesc(2u, BKV);  // N = 32 keys
      const uint32_t idesc_o = make_idesc(2u, DK);   // N = 128 head dims
      const uint32_t qh = tmem_base + COL_QH, ql = tmem_base + COL_QL;
      auto issue_s = [&](int j) {
        const int s = j & 1, ks_ = j % NK;
        mbar_wait(k_xf0 + 8 * ks_, (j / NK) & 1);
        mbar_wait(s_empty0 + 8 * s, ((j >> 1) & 1) ^ 1);
        tc_fence_after();
        ATT_TRACE(6, j);
        const uint32_t kh = desc_lo(smem0 + OFF_K + ks_ * 2 * KV_BYTES), kl = kh + (KV_BYTES >> 4);
        const uint32_t d = tmem_base + COL_S + s * BKV;
#pragma unroll
        for (int ks = 0; ks < DK / 8; ++ks) {  // 16 steps of 8 head dims
          const uint32_t ko = ((ks >> 2) * CH_BYTES + (ks & 3) * 32) >> 4;
          umma_tf32_ts_w(d, ql + ks * 8, kh + ko, idesc_s, ks != 0);
          umma_tf32_ts_w(d, qh + ks * 8, kl + ko, idesc_s, 1u);
          umma_tf32_ts_w(d, qh + ks * 8, kh + ko, idesc_s, 1u);
        }
        umma_commit(s_full0 + 8 * s);
        umma_commit(k_empty0 + 8 * ks_);
        ATT_TRACE(7, j);
      };
      mbar_wait(q_tm, 0);
      tc_fence_after();
      issue_s(0);
      for (int j = 0; j < NT; ++j) {
        if (j + 1 < NT) issue_s(j + 1);
        const int s = j & 1;
        mbar_wait(vt_full0 + 8 * s, (j >> 1) & 1);
        mbar_wait(p_full, j & 1);
        mbar_wait(o_empty, (j & 1) ^ 1);
        tc_fence_after();
        ATT_TRACE(8, j);
        const uint32_t vh = desc_lo(smem0 + OFF_VT + s * 2 * KV_BYTES), vl = vh + (KV_BYTES >> 4);  // V^T hi / lo
        const uint32_t ph = tmem_base + COL_P, pl = ph + BKV;                                       // P hi / lo in TMEM
        const uint32_t d = tmem_base + COL_O;
#pragma unroll
        for (int kk = 0; kk < BKV / 8; ++kk) {  // 4 steps of 8 keys: 8 TMEM columns of P, 32 bytes of each V^T row
          umma_tf32_ts_w(d, pl + kk * 8, vh + kk * 2, idesc_o, kk != 0);
          umma_tf32_ts_w(d, ph + kk * 8, vl + kk * 2, idesc_o, 1u);
          umma_tf32_ts_w(d, ph + kk * 8, vh + kk * 2, idesc_o, 1u);
        }
        umma_commit(o_full);
        umma_commit(vt_empty0 + 8 * s);
        ATT_TRACE(9, j);
      }
    }
  } else if (warp < 2 + N_SOFTMAX_WARPS) {
    // ================= softmax + output accumulation =================
    // thread = query row (lane group = warp % 4); half = which 16 keys of P / 64 columns of O it owns
    const int lg = warp & 3;
    const int half = (warp - 2) >> 2;
    const int r = lg * 32 + lane;
    const uint32_t lane_base = tmem_base + ((uint32_t)(lg * 32) << 16);
    const int sw = r & 7;
    {  // Q: shared staging (swizzled chunk tiles) -> registers -> TMEM hi / lo; this half does two of the chunks
      mbar_wait(q_full, 0);
#pragma unroll 1
      for (int c = 2 * half; c < 2 * half + 2; ++c) {
        const uint8_t* row = smem_al + OFF_K + c * (BQ * 128) + r * 128;
        float hi[32], lo[32];
        const float qs = 1.4426950408889634f / temperature;  // log2(e)/sqrt(dk) folded into Q
#pragma unroll
        for (int u = 0; u < 8; ++u) {
          float4 x = *reinterpret_cast<const float4*>(row + ((u ^ sw) << 4));
          x.x *= qs; x.y *= qs; x.z *= qs; x.w *= qs;
          hi[4 * u + 0] = tf32_rn(x.x); lo[4 * u + 0] = (x.x - hi[4 * u + 0]);
          hi[4 * u + 1] = tf32_rn(x.y); lo[4 * u + 1] = (x.y - hi[4 * u + 1]);
          hi[4 * u + 2] = tf32_rn(x.z); lo[4 * u + 2] = (x.z - hi[4 * u + 2]);
          hi[4 * u + 3] = tf32_rn(x.w); lo[4 * u + 3] = (x.w - hi[4 * u + 3]);
        }
        tmem_st32(lane_base + COL_QH + c * 32, hi);
        tmem_st32(lane_base + COL_QL + c * 32, lo);
      }
      asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(q_tm);
    }
    // K_j is split into TF32 hi / lo by these warps too (element-wise, 4 vectors per thread, two tiles ahead of its
    // S_j): with the split on the two transpose warps as well, those paced the kernel (~2 k cycles per tile)
    const int st_id = (warp - 2) * 32 + lane;  // 0..255
    auto split_k = [&](int jj) {
      if (jj >= NT) return;
      const int ks = jj % NK;
      mbar_wait(k_full0 + 8 * ks, (jj / NK) & 1);
      if (st_id == 0) ATT_TRACE(2, jj);
      float4* khi = reinterpret_cast<float4*>(smem_al + OFF_K + ks * 2 * KV_BYTES);
      float4* klo = khi + KV_BYTES / 16;
      float4 x[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) x[i] = khi[st_id + i * (N_SOFTMAX_WARPS * 32)];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const float4 hv = make_float4(tf32_rn(x[i].x), tf32_rn(x[i].y), tf32_rn(x[i].z), tf32_rn(x[i].w));
        khi[st_id + i * (N_SOFTMAX_WARPS * 32)] = hv;
        klo[st_id + i * (N_SOFTMAX_WARPS * 32)] = make_float4(x[i].x - hv.x, x[i].y - hv.y, x[i].z - hv.z, x[i].w - hv.w);
      }
      fence_proxy_async();
      __syncwarp();
      if (lane == 0) mbar_arrive(k_xf0 + 8 * ks);
      if (st_id == 0) ATT_TRACE(3, jj);
    };
    split_k(0);
    split_k(1);

    float m = -INFINITY, l = 0.f, alpha_fold = 1.f;
    float o[DK / 2];
#pragma unroll
    for (int i = 0; i < DK / 2; ++i) o[i] = 0.f;

    auto fold_o = [&](int jj, float alpha) {  // o = o*alpha + O_jj (this half's 64 columns)
      mbar_wait(o_full, jj & 1);
      tc_fence_after();
      if (warp == 2 && lane == 0) ATT_TRACE(12, jj);
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        float v[32];
        tmem_ld32(lane_base + COL_O + half * 64 + c * 32, v);
#pragma unroll
        for (int i = 0; i < 32; ++i) o[c * 32 + i] = fmaf(o[c * 32 + i], alpha, v[i]);
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(o_empty);
      if (warp == 2 && lane == 0) ATT_TRACE(13, jj);
    };

    for (int j = 0; j < NT; ++j) {
      const int sb = j & 1;
      mbar_wait(s_full0 + 8 * sb, (j >> 1) & 1);
      tc_fence_after();
      if (warp == 2 && lane == 0) ATT_TRACE(10, j);
      float s0[32];
      tmem_ld32(lane_base + COL_S + sb * BKV, s0);
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(s_empty0 + 8 * sb);

      const int kbase = j * BKV;
      float mx4[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};  // four chains: the warp has little else to overlap
      if (kbase + BKV <= len) {
#pragma unroll
        for (int i = 0; i < 32; ++i) mx4[i & 3] = fmaxf(mx4[i & 3], s0[i]);
      } else {
#pragma unroll
        for (int i = 0; i < 32; ++i) {
          s0[i] = (kbase + i < len) ? s0[i] : -INFINITY;
          mx4[i & 3] = fmaxf(mx4[i & 3], s0[i]);
        }
      }
      const float m_new = fmaxf(fmaxf(m, mx4[0]), fmaxf(fmaxf(mx4[1], mx4[2]), mx4[3]));  // finite: >= 1 valid key per tile
      const float alpha = ex2_approx(m - m_new);
      m = m_new;
      // this half's 16 keys: exponentiate, sum, split into TF32 hi / lo
      float pv_hi[16], pv_lo[16];
      float sum4[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
      for (int uu = 0; uu < 4; ++uu) {
        float4 a;  // the s0 index has to be a compile-time constant (registers): select on half
        a.x = ex2_approx((half ? s0[16 + 4 * uu + 0] : s0[4 * uu + 0]) - m_new);
        a.y = ex2_approx((half ? s0[16 + 4 * uu + 1] : s0[4 * uu + 1]) - m_new);
        a.z = ex2_approx((half ? s0[16 + 4 * uu + 2] : s0[4 * uu + 2]) - m_new);
        a.w = ex2_approx((half ? s0[16 + 4 * uu + 3] : s0[4 * uu + 3]) - m_new);
        sum4[uu] = (a.x + a.y) + (a.z + a.w);
        pv_hi[4 * uu + 0] = tf32_rn(a.x); pv_lo[4 * uu + 0] = (a.x - pv_hi[4 * uu + 0]);
        pv_hi[4 * uu + 1] = tf32_rn(a.y); pv_lo[4 * uu + 1] = (a.y - pv_hi[4 * uu + 1]);
        pv_hi[4 * uu + 2] = tf32_rn(a.z); pv_lo[4 * uu + 2] = (a.z - pv_hi[4 * uu + 2]);
        pv_hi[4 * uu + 3] = tf32_rn(a.w); pv_lo[4 * uu + 3] = (a.w - pv_hi[4 * uu + 3]);
      }
      l = l * alpha + ((sum4[0] + sum4[1]) + (sum4[2] + sum4[3]));
      // P lives in TMEM (single buffer): P.V_{j-1} must have retired before it is overwritten - that is the same
      // event fold(j-1) waits for, so the fold comes first; S_{j+1} keeps the tensor pipe busy meanwhile
      if (j > 0) fold_o(j - 1, alpha_fold);
      alpha_fold = alpha;
      tmem_st16(lane_base + COL_P + half * 16, pv_hi);
      tmem_st16(lane_base + COL_P + BKV + half * 16, pv_lo);
      asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
      tc_fence_before();
      __syncwarp();
      if (warp == 2 && lane == 0) ATT_TRACE(11, j);
      if (lane == 0) mbar_arrive(p_full);
      split_k(j + 2);
    }
    fold_o(NT - 1, alpha_fold);

    // merge the two halves' row sums through shared memory (all MMAs have retired: the K ring is free)
    float* lbuf = reinterpret_cast<float*>(smem_al + OFF_K);
    lbuf[half * BQ + r] = l;
    asm volatile("bar.sync 1, %0;" ::"n"(N_SOFTMAX_WARPS * 32) : "memory");
    l += lbuf[(half ^ 1) * BQ + r];
    const int q = q0 + r;
    if (q < S) {
      const float inv = q < len ? 1.0f / l : 0.f;  // padded query rows are written as zeros
      float4* orow = reinterpret_cast<float4*>(out + (int64_t)q * D + half * 64);
#pragma unroll
      for (int c = 0; c < DK / 8; ++c)
        orow[c] = make_float4(o[4 * c] * inv, o[4 * c + 1] * inv, o[4 * c + 2] * inv, o[4 * c + 3] * inv);
    }
  } else {
    // ================= V transpose (2 warps, 64 threads) =================
    const int tt = (warp - 2 - N_SOFTMAX_WARPS) * 32 + lane;  // 0..63
    for (int j = 0; j < NT; ++j) {
      const int s = j & 1;
      const uint32_t ph = (j >> 1) & 1;
      {  // V [32 keys x 128 dims] (chunked [c][key][32]) -> V^T hi / lo [128 dims x 32 keys], swizzled rows
        mbar_wait(v_full0 + 8 * s, ph);
        mbar_wait(vt_empty0 + 8 * s, ph ^ 1);  // P.V_{j-2} has retired
        if (tt == 0) ATT_TRACE(4, j);
        const float* vr = reinterpret_cast<const float*>(smem_al + OFF_VR + s * KV_BYTES);
#pragma unroll 1
        for (int dd = 0; dd < DK / (N_XF_WARPS * 32); ++dd) {
          const int d = tt + dd * (N_XF_WARPS * 32);  // head dim = V^T row
          const float* vraw = vr + (d >> 5) * (CH_BYTES / 4) + (d & 31);
          uint8_t* th = smem_al + OFF_VT + s * 2 * KV_BYTES + d * 128;
          uint8_t* tl = th + KV_BYTES;
          const int swd = d & 7;
          float xv[32];
#pragma unroll
          for (int k = 0; k < 32; ++k) xv[k] = vraw[k * 32];  // all 32 keys of this head dim first
#pragma unroll
          for (int u = 0; u < 8; ++u) {  // keys 4u .. 4u+3
            const float x0 = xv[4 * u + 0], x1 = xv[4 * u + 1], x2 = xv[4 * u + 2], x3 = xv[4 * u + 3];
            const float4 hv = make_float4(tf32_rn(x0), tf32_rn(x1), tf32_rn(x2), tf32_rn(x3));
            const int pu = (u ^ swd) << 4;
            *reinterpret_cast<float4*>(th + pu) = hv;
            *reinterpret_cast<float4*>(tl + pu) =
                make_float4((x0 - hv.x), (x1 - hv.y), (x2 - hv.z), (x3 - hv.w));
          }
        }
        fence_proxy_async();
        __syncwarp();
        if (lane == 0) {
          mbar_arrive(vr_empty0 + 8 * s);
          mbar_arrive(vt_full0 + 8 * s);
        }
        if (tt == 0) ATT_TRACE(5, j);
      }
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

#ifdef SS_ATT_TRACE
extern "C" __attribute__((visibility("default"))) int ss_debug_att_trace(long long* host) {
  return (int)cudaMemcpyFromSymbol(host, g_att_trace, sizeof(long long) * 16 * 64);
}
#endif

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
  static std::atomic<unsigned long long> attr_mask{0};  // per device
  if (needs_attr_on_device(attr_mask)) {
    cudaError_t e = cudaFuncSetAttribute(attention_tc128_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES);
    if (e != cudaSuccess) return ss::cuda_fail(e, "ss_attention_tc: smem opt-in (dk=128)");
    attr_done_on_device(attr_mask);
  }
  dim3 grid((S + BQ - 1) / BQ, H, B);
  attention_tc128_kernel<<<grid, NTHREADS, SMEM_BYTES, static_cast<cudaStream_t>(stream)>>>(tmQ, tmK, tmV, lens, ctx, S,
                                                                                            H, sqrtf((float)DK));
  SS_CHECK_LAUNCH("ss_attention_tc (dk=128)");
  return SS_OK;
}
