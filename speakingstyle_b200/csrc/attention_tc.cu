// Fused masked-softmax self-attention on the tensor cores for head size 32 (the ReferenceEncoder's
// 8 x 32 heads), fp32-accurate: both GEMMs are 3xTF32 error-compensated tcgen05.mma with fp32
// accumulation in TMEM; softmax statistics, probabilities and the output accumulator are fp32.
//
// One CTA = 128 queries of one head of one utterance; keys/values stream through in 64-key tiles up to
// the utterance's valid length.  The A operands of both GEMMs live in TENSOR MEMORY: with N = 64 / 32 the
// MMAs are small, and re-reading a 4 KB Q or P slice from shared memory for each of them (the first
// version) held them at ~45 cycles apiece where the tensor pipe needs 32 / 16.
// Roles (384 threads = three warps per scheduler, 168 registers each):
//   warp 0      TMA: Q once, then V_j / K_j tiles into four-deep rings (rows outside [0,S) zero-filled)
//   warp 1      one elected thread issues  S_j = Q.K_j^T  (12 MMAs, N = 64, A = Q hi/lo in TMEM) and
//               O_j = P_j.V_j  (24 MMAs, N = 32, A = P_j hi/lo in TMEM); S runs two tiles ahead of P.V
//   warps 2-9   softmax: thread = query row, warps w and w+4 share the rows.  First they split this CTA's
//               Q rows into TMEM.  Per tile both read the whole S_j row (same running max, base-2 domain);
//               each exponentiates its own 32 keys (MUFU.EX2), tcgen05.st's them as TF32 hi/lo into the
//               P buffer (double-buffered), keeps a partial row sum, and folds its 16 columns of O_{j-1}
//               in registers (o = o*alpha + O, fp32 round-to-nearest, so the truncating TMEM accumulation
//               never runs longer than one tile) - after P_j is handed over, off the critical path.
//   warps 10-11 split K_j into TF32 hi/lo in shared memory and transpose V_j while splitting it: V arrives
//               [keys x 32] but the MMA wants its B operand K-major, i.e. V^T with the keys contiguous
//               (conflict-free: a lane reads one head dim of four consecutive keys and writes one 16-byte
//               unit of the swizzled V^T row).
// Rings: K hi/lo freed by S_j, V as loaded freed by the transpose, V^T hi/lo freed by P.V_j.
// TMEM (512 columns): S x2 (128) | O x2 (64) | Q hi, lo (64) | P hi, lo x2 (256).
// The [S x S] score matrix never leaves the SM.
#include "tc_common.cuh"

namespace {

using namespace sstc;

constexpr int DK = 32;            // head size = one 128-byte row of fp32
constexpr int BQ = 128;           // queries per CTA = TMEM lanes
constexpr int BKV = 64;           // keys per tile
constexpr int NS = 4;             // ring depth
constexpr int N_SOFTMAX_WARPS = 8, N_XF_WARPS = 2, XF_THREADS = N_XF_WARPS * 32;
constexpr int NTHREADS = 32 * (2 + N_SOFTMAX_WARPS + N_XF_WARPS);
constexpr int Q_BYTES = BQ * 128;      // 16 KB
constexpr int KV_BYTES = BKV * 128;    // 8 KB
constexpr int OFF_Q = 0;                             // Q as loaded (read once by the softmax warps)
constexpr int OFF_K = Q_BYTES;                       // [NS][K_hi 8 | K_lo 8]
constexpr int OFF_VR = OFF_K + NS * 2 * KV_BYTES;    // [NS][V as loaded (unswizzled) 8]
constexpr int OFF_VT = OFF_VR + NS * KV_BYTES;       // [NS][V^T_hi 8 | V^T_lo 8]
constexpr int SMEM_DATA = OFF_VT + NS * 2 * KV_BYTES;  // 16 + 160 KB
constexpr int N_BARS = 7 * NS + 12;
constexpr int SMEM_BYTES = SMEM_DATA + N_BARS * 8 + 16 + 1024;
constexpr int COL_S = 0, COL_O = 2 * BKV, COL_QH = COL_O + 2 * DK, COL_QL = COL_QH + DK, COL_P = COL_QL + DK;
constexpr int TMEM_COLS = 512;
static_assert(COL_P + 2 * 2 * BKV <= TMEM_COLS, "TMEM budget");
static_assert(SMEM_BYTES <= 232448, "shared memory budget");

// Debug build (-DSS_ATT_TRACE, scripts/attn_trace.py): CTA (0,0,0) stamps clock64() at the pipeline hand-overs.
#ifdef SS_ATT_TRACE
__device__ long long g_att32_trace[16 * 64];
#define ATT_TRACE(ev, j)                                                                   \
  do {                                                                                     \
    if ((blockIdx.x | blockIdx.y | blockIdx.z) == 0 && (j) < 64) g_att32_trace[(ev) * 64 + (j)] = clock64(); \
  } while (0)
#else
#define ATT_TRACE(ev, j)
#endif

// Only hi is rounded here; lo = x - hi goes to the MMA as is - the tensor core truncates an fp32 operand to TF32
// itself, and lo only carries bits 12-22 of x (same finding as in conv1d_tf32_ts.cu: error unchanged, fewer ALU ops
// on the split / softmax warps that pace these kernels).

// All loads of a batch come first: with the in-place stores in between, every load would otherwise be serialised
// behind the previous store (aliasing) on a warp with no ILP to spare.
template <int N_VEC>
__device__ __forceinline__ void split_tile(float4* hi, float4* lo, int tt) {
  constexpr int PER = N_VEC / XF_THREADS;  // vectors per thread
  static_assert(N_VEC % XF_THREADS == 0 && PER >= 1 && PER <= 8, "split_tile batch");
  float4 x[PER];
#pragma unroll
  for (int i = 0; i < PER; ++i) x[i] = hi[tt + i * XF_THREADS];
#pragma unroll
  for (int i = 0; i < PER; ++i) {
    const float4 h = make_float4(tf32_rn(x[i].x), tf32_rn(x[i].y), tf32_rn(x[i].z), tf32_rn(x[i].w));
    hi[tt + i * XF_THREADS] = h;
    lo[tt + i * XF_THREADS] = make_float4((x[i].x - h.x), (x[i].y - h.y), (x[i].z - h.z), (x[i].w - h.w));
  }
}

__device__ __forceinline__ void tmem_ld16(uint32_t taddr, float (&v)[16]) {
  uint32_t r[16];
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr));
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
  for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(r[i]);
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
  const uint32_t bars = smem0 + SMEM_DATA;
  // NS-deep rings: index with + 8 * slot
  const uint32_t k_full0 = bars, k_xf0 = k_full0 + 8 * NS, k_empty0 = k_xf0 + 8 * NS;
  const uint32_t v_full0 = k_empty0 + 8 * NS, vr_empty0 = v_full0 + 8 * NS, vt_full0 = vr_empty0 + 8 * NS,
                 vt_empty0 = vt_full0 + 8 * NS;
  // two-deep: S, O, P buffers
  const uint32_t s_full0 = vt_empty0 + 8 * NS, s_empty0 = s_full0 + 16, o_full0 = s_empty0 + 16, o_empty0 = o_full0 + 16,
                 p_full0 = o_empty0 + 16, q_full = p_full0 + 16, q_tm = q_full + 8;
  const uint32_t tslot = bars + N_BARS * 8;
  volatile uint32_t* tslot_ptr = reinterpret_cast<volatile uint32_t*>(smem_al + SMEM_DATA + N_BARS * 8);

  const int NT = (len + BKV - 1) / BKV;

  if (threadIdx.x == 0) {
    prefetch_tmap(&tmQ);
    prefetch_tmap(&tmKV);
    prefetch_tmap(&tmV);
    for (int s = 0; s < NS; ++s) {
      mbar_init(k_full0 + 8 * s, 1);
      mbar_init(k_xf0 + 8 * s, N_XF_WARPS);
      mbar_init(k_empty0 + 8 * s, 1);
      mbar_init(v_full0 + 8 * s, 1);
      mbar_init(vr_empty0 + 8 * s, N_XF_WARPS);
      mbar_init(vt_full0 + 8 * s, N_XF_WARPS);
      mbar_init(vt_empty0 + 8 * s, 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(s_full0 + 8 * s, 1);
      mbar_init(s_empty0 + 8 * s, N_SOFTMAX_WARPS);
      mbar_init(o_full0 + 8 * s, 1);
      mbar_init(o_empty0 + 8 * s, N_SOFTMAX_WARPS);
      mbar_init(p_full0 + 8 * s, N_SOFTMAX_WARPS);
    }
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
      mbar_expect_tx(q_full, Q_BYTES);
      tma_load_3d(smem0 + OFF_Q, &tmQ, q_full, h * DK, q0, b);
      int s = 0;
      uint32_t ph = 1;  // "empty" parity of a fresh barrier
      for (int j = 0; j < NT; ++j) {
        mbar_wait(vr_empty0 + 8 * s, ph);
        ATT_TRACE(0, j);
        mbar_expect_tx(v_full0 + 8 * s, KV_BYTES);
        tma_load_3d(smem0 + OFF_VR + s * KV_BYTES, &tmV, v_full0 + 8 * s, 2 * D + h * DK, j * BKV, b);  // V_j, unswizzled
        mbar_wait(k_empty0 + 8 * s, ph);
        ATT_TRACE(1, j);
        mbar_expect_tx(k_full0 + 8 * s, KV_BYTES);
        tma_load_3d(smem0 + OFF_K + s * 2 * KV_BYTES, &tmKV, k_full0 + 8 * s, D + h * DK, j * BKV, b);  // K_j -> K_hi slot
        if (++s == NS) { s = 0; ph ^= 1; }
      }
    }
  } else if (warp == 1) {
    // ================= MMA issuer =================
    if (elect_one()) {  // elect.sync, not lane == 0: lets ptxas issue TMA / MMA without a per-instruction elect loop
      const uint32_t idesc_s = make_idesc(2u, BKV);               // tf32, N = 64, B = K K-major
      const uint32_t idesc_o = make_idesc(2u, DK);                // tf32, N = 32, B = V^T K-major
      const uint32_t qh = tmem_base + COL_QH, ql = tmem_base + COL_QL;
      auto issue_s = [&](int j) {
        const int s = j % NS, sb = j & 1;
        mbar_wait(k_xf0 + 8 * s, (j / NS) & 1);
        mbar_wait(s_empty0 + 8 * sb, ((j >> 1) & 1) ^ 1);
        tc_fence_after();
        ATT_TRACE(6, j);
        const uint32_t kh = desc_lo(smem0 + OFF_K + s * 2 * KV_BYTES), kl = kh + (KV_BYTES >> 4);
        const uint32_t d = tmem_base + COL_S + sb * BKV;
#pragma unroll
        for (int kk = 0; kk < DK / 8; ++kk) {  // 8 head dims per instruction: 8 TMEM columns of Q, 32 bytes of each K row
          umma_tf32_ts_w(d, ql + kk * 8, kh + kk * 2, idesc_s, kk != 0);
          umma_tf32_ts_w(d, qh + kk * 8, kl + kk * 2, idesc_s, 1u);
          umma_tf32_ts_w(d, qh + kk * 8, kh + kk * 2, idesc_s, 1u);
        }
        umma_commit(s_full0 + 8 * sb);
        umma_commit(k_empty0 + 8 * s);
        ATT_TRACE(7, j);
      };
      mbar_wait(q_tm, 0);
      tc_fence_after();
      // S runs TWO tiles ahead of P.V: S_{j+2} only needs the softmax warps to have READ S_j (its buffer), which
      // happens ~1 k cycles before they hand over P_j - issued after P.V_j instead, it sat on the critical loop
      // P_j -> P.V_j -> S_{j+2} -> softmax_{j+2} -> P_{j+2}
      issue_s(0);
      if (NT > 1) issue_s(1);
      for (int j = 0; j < NT; ++j) {
        if (j + 2 < NT) issue_s(j + 2);
        const int s = j % NS, pb = j & 1;
        mbar_wait(vt_full0 + 8 * s, (j / NS) & 1);
        mbar_wait(p_full0 + 8 * pb, (j >> 1) & 1);
        mbar_wait(o_empty0 + 8 * pb, ((j >> 1) & 1) ^ 1);
        tc_fence_after();
        ATT_TRACE(8, j);
        const uint32_t vh = desc_lo(smem0 + OFF_VT + s * 2 * KV_BYTES), vl = vh + (KV_BYTES >> 4);  // V^T hi / lo
        const uint32_t ph = tmem_base + COL_P + pb * 2 * BKV, pl = ph + BKV;
        const uint32_t d = tmem_base + COL_O + pb * DK;
#pragma unroll
        for (int ks = 0; ks < BKV / 8; ++ks) {  // 8 keys per instruction: 8 TMEM columns of P, 32 bytes of each V^T row
          const uint32_t vb = ((ks >> 2) * (DK * 128) + (ks & 3) * 32) >> 4;  // V^T: 32 rows x 128 B per 32-key chunk
          umma_tf32_ts_w(d, pl + ks * 8, vh + vb, idesc_o, ks != 0);
          umma_tf32_ts_w(d, ph + ks * 8, vl + vb, idesc_o, 1u);
          umma_tf32_ts_w(d, ph + ks * 8, vh + vb, idesc_o, 1u);
        }
        umma_commit(o_full0 + 8 * pb);
        umma_commit(vt_empty0 + 8 * s);
        ATT_TRACE(9, j);
      }
    }
  } else if (warp < 2 + N_SOFTMAX_WARPS) {
    // ================= softmax + output accumulation =================
    // thread = query row (lane group = warp % 4); half = which 32 keys of the tile / 16 columns of O it owns
    const int lg = warp & 3;
    const int half = (warp - 2) >> 2;
    const int r = lg * 32 + lane;
    const uint32_t lane_base = tmem_base + ((uint32_t)(lg * 32) << 16);
    {  // Q row: shared staging (one swizzled 128-byte row) -> TMEM; half 0 stores hi, half 1 stores lo
      mbar_wait(q_full, 0);
      const uint8_t* row = smem_al + OFF_Q + r * 128;
      const int sw = r & 7;
      float qv[32];
      const float qs = 1.4426950408889634f / temperature;  // log2(e)/sqrt(dk) folded into Q: S comes out in the base-2 domain
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        float4 x = *reinterpret_cast<const float4*>(row + ((u ^ sw) << 4));
        x.x *= qs; x.y *= qs; x.z *= qs; x.w *= qs;
        const float4 hq = make_float4(tf32_rn(x.x), tf32_rn(x.y), tf32_rn(x.z), tf32_rn(x.w));
        qv[4 * u + 0] = half ? (x.x - hq.x) : hq.x;
        qv[4 * u + 1] = half ? (x.y - hq.y) : hq.y;
        qv[4 * u + 2] = half ? (x.z - hq.z) : hq.z;
        qv[4 * u + 3] = half ? (x.w - hq.w) : hq.w;
      }
      tmem_st32(lane_base + (half ? COL_QL : COL_QH), qv);
      asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(q_tm);
    }
    float m = -INFINITY, l = 0.f, alpha_fold = 1.f;
    float o[DK / 2];
#pragma unroll
    for (int i = 0; i < DK / 2; ++i) o[i] = 0.f;

    auto fold_o = [&](int jj, float alpha) {  // o = o*alpha + O_jj (this half's 16 columns)
      const int ob = jj & 1;
      mbar_wait(o_full0 + 8 * ob, (jj >> 1) & 1);
      tc_fence_after();
      if (warp == 2 && lane == 0) ATT_TRACE(12, jj);
      float v[16];
      tmem_ld16(lane_base + COL_O + ob * DK + half * 16, v);
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(o_empty0 + 8 * ob);
#pragma unroll
      for (int i = 0; i < 16; ++i) o[i] = fmaf(o[i], alpha, v[i]);
      if (warp == 2 && lane == 0) ATT_TRACE(13, jj);
    };

    for (int j = 0; j < NT; ++j) {
      const int sb = j & 1;
      mbar_wait(s_full0 + 8 * sb, (j >> 1) & 1);
      tc_fence_after();
      if (warp == 2 && lane == 0) ATT_TRACE(10, j);
      float sa[32], sc[32];  // sa = this half's 32 keys, sc = the other half's (needed for the row max only)
      tmem_ld32x2(lane_base + COL_S + sb * BKV + half * 32, sa, lane_base + COL_S + sb * BKV + (half ^ 1) * 32, sc);
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(s_empty0 + 8 * sb);

      // softmax in the base-2 domain: S already carries log2(e)/sqrt(dk) (folded into Q); p = 2^(t - m), one
      // MUFU.EX2 per score; relative error of p <= ~3e-7, far inside the kernel's 1e-6 class.
      const int kbase = j * BKV;
      float mx4[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};  // four chains: the warp has little else to overlap
      if (kbase + BKV <= len) {
#pragma unroll
        for (int i = 0; i < 32; ++i) mx4[i & 3] = fmaxf(mx4[i & 3], fmaxf(sa[i], sc[i]));
      } else {
        const int ka = kbase + half * 32, kc = kbase + (half ^ 1) * 32;
#pragma unroll
        for (int i = 0; i < 32; ++i) {
          sa[i] = (ka + i < len) ? sa[i] : -INFINITY;
          const float c = (kc + i < len) ? sc[i] : -INFINITY;
          mx4[i & 3] = fmaxf(mx4[i & 3], fmaxf(sa[i], c));
        }
      }
      const float m_new = fmaxf(fmaxf(m, mx4[0]), fmaxf(fmaxf(mx4[1], mx4[2]), mx4[3]));  // finite: >= 1 valid key per tile
      const float alpha = ex2_approx(m - m_new);
      m = m_new;
      float sum4[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
      for (int i = 0; i < 32; ++i) {
        sa[i] = ex2_approx(sa[i] - m_new);
        sum4[i & 3] += sa[i];
      }
      l = l * alpha + ((sum4[0] + sum4[1]) + (sum4[2] + sum4[3]));
      if (warp == 2 && lane == 0) ATT_TRACE(14, j);

      // P_j (this half's 32 keys) -> TMEM as TF32 hi / lo.  Buffer j & 1 was last read by P.V_{j-2}, which had
      // retired before fold(j-2) returned in the previous iteration.
      const uint32_t pcol = lane_base + COL_P + sb * 2 * BKV + half * 32;
#pragma unroll
      for (int i = 0; i < 32; ++i) sc[i] = tf32_rn(sa[i]);
      tmem_st32(pcol, sc);
#pragma unroll
      for (int i = 0; i < 32; ++i) sa[i] = (sa[i] - sc[i]);
      tmem_st32(pcol + BKV, sa);
      asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(p_full0 + 8 * sb);
      if (warp == 2 && lane == 0) ATT_TRACE(11, j);

      if (j > 0) fold_o(j - 1, alpha_fold);
      alpha_fold = alpha;
    }
    fold_o(NT - 1, alpha_fold);

    // merge the two halves' row sums through shared memory (all loads have landed and been consumed: Q staging is free)
    float* lbuf = reinterpret_cast<float*>(smem_al + OFF_Q);
    lbuf[half * BQ + r] = l;
    asm volatile("bar.sync 1, %0;" ::"n"(N_SOFTMAX_WARPS * 32) : "memory");
    l += lbuf[(half ^ 1) * BQ + r];
    const int q = q0 + r;
    if (q < S) {
      const float inv = q < len ? 1.0f / l : 0.f;  // padded query rows are written as zeros
      float4* orow = reinterpret_cast<float4*>(out + (int64_t)q * D + half * 16);
#pragma unroll
      for (int c = 0; c < DK / 8; ++c)
        orow[c] = make_float4(o[4 * c] * inv, o[4 * c + 1] * inv, o[4 * c + 2] * inv, o[4 * c + 3] * inv);
    }
  } else {
    // ================= operand split / transpose (2 warps) =================
    const int w = warp - 2 - N_SOFTMAX_WARPS;
    const int tt = w * 32 + lane;
    int s = 0;
    uint32_t ph = 0;
    for (int j = 0; j < NT; ++j) {
      mbar_wait(k_full0 + 8 * s, ph);
      if (tt == 0) ATT_TRACE(2, j);
      split_tile<KV_BYTES / 16>(reinterpret_cast<float4*>(smem_al + OFF_K + s * 2 * KV_BYTES),
                                reinterpret_cast<float4*>(smem_al + OFF_K + s * 2 * KV_BYTES + KV_BYTES), tt);
      fence_proxy_async();
      __syncwarp();
      if (lane == 0) mbar_arrive(k_xf0 + 8 * s);
      if (tt == 0) ATT_TRACE(3, j);

      mbar_wait(v_full0 + 8 * s, ph);
      mbar_wait(vt_empty0 + 8 * s, ph ^ 1);  // P.V_{j-NS} has retired
      if (tt == 0) ATT_TRACE(4, j);
      {  // V [64 keys x 32 dims] -> V^T hi / lo [32 dims x 64 keys], K-major, 128-byte swizzle
        const float* vraw = reinterpret_cast<const float*>(smem_al + OFF_VR + s * KV_BYTES);
        uint8_t* vt = smem_al + OFF_VT + s * 2 * KV_BYTES;
        const int d = lane;
        float xv[16 / N_XF_WARPS][4];
#pragma unroll
        for (int i = 0; i < 16 / N_XF_WARPS; ++i)  // all loads first (the stores below could alias them)
#pragma unroll
          for (int e = 0; e < 4; ++e) xv[i][e] = vraw[(4 * (w + N_XF_WARPS * i) + e) * DK + d];
#pragma unroll
        for (int i = 0; i < 16 / N_XF_WARPS; ++i) {
          const int kq = w + N_XF_WARPS * i;  // quad of keys 4*kq .. 4*kq+3
          const float x0 = xv[i][0], x1 = xv[i][1], x2 = xv[i][2], x3 = xv[i][3];
          const float4 hv = make_float4(tf32_rn(x0), tf32_rn(x1), tf32_rn(x2), tf32_rn(x3));
          const int off = (kq >> 3) * (DK * 128) + d * 128 + (((kq & 7) ^ (d & 7)) << 4);
          *reinterpret_cast<float4*>(vt + off) = hv;
          *reinterpret_cast<float4*>(vt + KV_BYTES + off) =
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
      if (++s == NS) { s = 0; ph ^= 1; }
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
extern "C" __attribute__((visibility("default"))) int ss_debug_att32_trace(long long* host) {
  return (int)cudaMemcpyFromSymbol(host, g_att32_trace, sizeof(long long) * 16 * 64);
}
#endif

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
  static std::atomic<unsigned long long> attr_mask{0};  // per device
  if (needs_attr_on_device(attr_mask)) {
    cudaError_t e = cudaFuncSetAttribute(attention_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES);
    if (e != cudaSuccess) return ss::cuda_fail(e, "ss_attention_tc: smem opt-in");
    attr_done_on_device(attr_mask);
  }
  dim3 grid((S + BQ - 1) / BQ, H, B);
  attention_tc_kernel<<<grid, NTHREADS, SMEM_BYTES, static_cast<cudaStream_t>(stream)>>>(tmQ, tmKV, tmV, lens, ctx, S,
                                                                                           H, sqrtf((float)dk));
  SS_CHECK_LAUNCH("ss_attention_tc");
  return SS_OK;
}
