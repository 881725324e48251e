// One HiFi-GAN ResBlock unit, fused (hifigan/models.py:96-103):  y' = y + c2( leaky_relu( c1( leaky_relu(y) ) ) )
// for the 32-channel stage (every kernel size) and the 64-channel k = 3 units, on the split-bf16 arithmetic of conv1d_bf16x3.cu (same MMAs in the same order, same
// epilogue arithmetic: results are bit-identical to the two-launch form).
//
// Why: at 32 channels both convolutions are HBM-bound (8 bytes per element for c1, 16 for c2).  The intermediate
// t = leaky_relu(c1(.)) is written as pair rows by c1's epilogue and read back by c2 - 8 of the unit's 24 bytes per
// element.  Here it never leaves the SM: 16 bytes per element.
//
// Tile.  c2 (dilation 1) needs k-1 rows of halo of t, so a tile computes 128 rows of t and TM = 128 - (k-1) rows of
// output:  output rows r0 + [0, TM),  t rows r0 - p2 + [0, 128)  (p2 = (k-1)/2),  input panel rows
// r0 - p2 - p1 + [0, 128 + (k-1)*d)  (p1 = d*(k-1)/2).  Consecutive tiles recompute k-1 rows of t (8 % at k = 11).
//   1. TMA: the input panel (pair rows [hi(32) | lo(32)], one 128-byte swizzle row each), weights of both
//      convolutions resident.
//   2. MMA: c1 over the panel (taps = shifted views, 4 MMAs per tap) -> accumulator 1 (TMEM).
//   3. Middle warps: accumulator 1 -> registers, + bias, zero outside [0, S) (c2's zero padding), leaky-ReLU, split
//      hi / lo, written into a shared-memory panel in the SWIZZLE_128B layout TMA would have produced (16-byte chunk c
//      of row r at chunk c ^ (r & 7)); fence.proxy.async; barrier.
//   4. MMA: c2 over that panel -> accumulator 2.  Rows >= TM of its M = 128 read past the 128 rows of t (stale shared
//      memory): they only reach output rows >= TM, which are not stored.
//   5. Final warps (two sets, alternate tiles): accumulator 2 + bias + residual -> y', pair rows of leaky_relu(y') for
//      the next unit, or the running stage mean.
// The MMA thread issues c1 of tile i+1 before c2 of tile i, so the tensor pipe has work while the middle warps run.
// 64 channels (hi and lo panels are separate 128-byte-row panels; weights of both convolutions resident only for
// k = 3): shared memory holds ONE intermediate panel pair and one set of final warps, so c2 of tile i and the middle
// warps' conversion of tile i+1 alternate on it; the unit is HBM-bound either way.
#include "x3_common.cuh"

namespace {

using namespace sstc;

constexpr int MID_WARPS = 4;
constexpr int MAX_A = 4;
constexpr int N_BARS = 2 * MAX_A + 1 + 12;
constexpr int MAX_SMEM = 227 * 1024;
// per channel count: 32 = one 128-byte row [hi | lo] per time step, two taps per weight row, double-buffered t panel, two
// sets of final warps; 64 = hi panel + lo panel, one tap per weight tile, one t panel pair, one set of final warps
template <int C> struct Unit {
  static constexpr bool narrow = C == 32;
  static constexpr int np = narrow ? 1 : 2;                 // 128-byte-row panels per operand (hi | lo in one, or two)
  static constexpr int t_rows = narrow ? 128 + 16 : 128 + 8;  // rows of a t panel: 128 written + up to k-1 read past them
  static constexpr int T_BYTES = t_rows * ROW_BYTES;        // a multiple of 1024
  static constexpr int WT = 2 * C * ROW_BYTES;              // one weight tile: [W_hi rows ; W_lo rows]
  static constexpr int NT = narrow ? 2 : 1;                 // t panel buffers
  static constexpr int FIN_WARPS = narrow ? 8 : 4;
  static constexpr int NSET = FIN_WARPS / 4;
  static constexpr int nthreads = 32 * (2 + MID_WARPS + FIN_WARPS);
  static constexpr int tmem_cols = 8 * C;                   // accumulator 1 x 2, accumulator 2 x 2, [main | correction] each
  static constexpr int fixed_smem = 1024 + NT * np * T_BYTES + FIN_WARPS * STG_WARP_BYTES + N_BARS * 8 + 16;
  static constexpr int tiles_per_conv(int k) { return narrow ? (k + 1) / 2 : k; }
};

struct UnitArgs {
  const float* b1;
  int32_t R1;           // input panel rows = 128 + (k-1)*dilation
  int32_t panel_bytes;  // R1 * 128 rounded up to 1024
  int32_t n_a;          // input panels in the ring
  int32_t npair;        // weight tiles per convolution: ceil(k/2) at 32 channels (two taps per row), k at 64
  int32_t TM;           // output rows per tile = 128 - (k-1)
  int32_t dilation;     // of c1 (c2: 1)
  float slope;          // leaky-ReLU between c1 and c2
};

template <int C>
__global__ void __launch_bounds__(Unit<C>::nthreads, 1)
resunit_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmW1,
               const __grid_constant__ CUtensorMap tmW2, const TcArgs p, const UnitArgs q) {
  using U = Unit<C>;
  constexpr bool NARROW = U::narrow;
  constexpr int WT = U::WT, NT = U::NT, FIN_WARPS = U::FIN_WARPS, NSET = U::NSET;
  constexpr int T_BYTES = U::T_BYTES;
  constexpr int TP = U::np * T_BYTES;  // one t panel buffer (hi | lo in one panel, or hi panel + lo panel)
  pdl_launch_dependents();
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem0 = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* smem_al = smem_raw + (smem0 - smem_u32(smem_raw));
  const int a_stage = U::np * q.panel_bytes;
  const int t_off = q.n_a * a_stage;
  const int w_off = t_off + NT * TP;
  const int stg_off = w_off + 2 * q.npair * WT;
  const int bar_off = stg_off + FIN_WARPS * STG_WARP_BYTES;
  float* stg_all = reinterpret_cast<float*>(smem_al + stg_off);
  const uint32_t bars = smem0 + bar_off;
  const uint32_t afull0 = bars, aempty0 = afull0 + 8 * MAX_A, wres = aempty0 + 8 * MAX_A, t1full0 = wres + 8,
                 t1empty0 = t1full0 + 16, pfull0 = t1empty0 + 16, pempty0 = pfull0 + 16, t2full0 = pempty0 + 16,
                 t2empty0 = t2full0 + 16, tslot = t2empty0 + 16;
  volatile uint32_t* tslot_ptr = reinterpret_cast<volatile uint32_t*>(smem_al + bar_off + N_BARS * 8);
  int* const rag = reinterpret_cast<int*>(smem_al + bar_off + N_BARS * 8 + 16);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const bool ragged = p.rows != nullptr;
  if (ragged) ragged_build(p.rows, p.B, p.S_out, q.TM, rag);

  if (threadIdx.x == 0) {
    prefetch_tmap(&tmA);
    prefetch_tmap(&tmW1);
    prefetch_tmap(&tmW2);
    for (int s = 0; s < MAX_A; ++s) {
      mbar_init(afull0 + 8 * s, 1);
      mbar_init(aempty0 + 8 * s, 1);
    }
    mbar_init(wres, 1);
    for (int a = 0; a < 2; ++a) {
      mbar_init(t1full0 + 8 * a, 1);
      mbar_init(t1empty0 + 8 * a, MID_WARPS);
      mbar_init(pfull0 + 8 * a, MID_WARPS);
      mbar_init(pempty0 + 8 * a, 1);
      mbar_init(t2full0 + 8 * a, 1);
      mbar_init(t2empty0 + 8 * a, 4);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) tmem_alloc(tslot, (uint32_t)U::tmem_cols);  // accumulator 1 x 2 (2*C columns each), accumulator 2 x 2
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  pdl_wait();  // everything below may touch memory the previous launch wrote
  const uint32_t tmem_base = *tslot_ptr;
  const int num_tiles = ragged ? rag[p.B] : p.B * p.m_tiles;
  auto decode = [&](int tile, int& mt, int& b) {
    if (ragged) {
      ragged_decode(rag, p.B, tile, b, mt);
    } else {
      mt = tile % p.m_tiles;
      b = tile / p.m_tiles;
    }
  };
  const int p2 = (p.k - 1) / 2, p1 = q.dilation * p2;

  if (warp == 0) {
    // ================= TMA producer =================
    if (elect_one()) {
      mbar_expect_tx(wres, (uint32_t)(2 * q.npair * WT));
      for (int i = 0; i < q.npair; ++i) {
        tma_load_3d(smem0 + w_off + i * WT, &tmW1, wres, 0, 0, i);
        tma_load_3d(smem0 + w_off + (q.npair + i) * WT, &tmW2, wres, 0, 0, i);
      }
      int as = 0;
      uint32_t aph = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        int mt, b;
        decode(tile, mt, b);
        mbar_wait(aempty0 + 8 * as, aph ^ 1);
        mbar_expect_tx(afull0 + 8 * as, (uint32_t)(U::np * q.R1 * ROW_BYTES));
        tma_load_3d(smem0 + as * a_stage, &tmA, afull0 + 8 * as, 0, mt * q.TM - p2 - p1, b);
        if (!NARROW) tma_load_3d(smem0 + as * a_stage + q.panel_bytes, &tmA, afull0 + 8 * as, C, mt * q.TM - p2 - p1, b);
        if (++as == q.n_a) { as = 0; aph ^= 1; }
      }
    }
  } else if (warp == 1) {
    // ================= MMA issuer =================
    if (elect_one()) {
      const uint32_t idesc_m = make_idesc(1u, 2 * C);  // x_hi * [W_hi ; W_lo] -> [main | correction]
      const uint32_t idesc_h = make_idesc(1u, C);      // x_lo * W_hi -> correction
      const uint32_t w1_lo = desc_lo(smem0 + w_off), w2_lo = desc_lo(smem0 + w_off + q.npair * WT);
      const uint32_t step1 = (uint32_t)(q.dilation * ROW_BYTES) >> 4, step2 = ROW_BYTES >> 4;
      mbar_wait(wres, 0);
      tc_fence_after();
      int my_tiles = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) ++my_tiles;
      int as = 0;
      uint32_t aph = 0;
      // one convolution of one tile: taps are shifted views of the operand panel
      // lo_off: descriptor units from the hi panel to the lo panel of the operand (64 channels only)
      auto conv = [&](uint32_t d_main, uint32_t a0, uint32_t step, uint32_t w_lo, uint32_t lo_off) {
        const uint32_t d_corr = d_main + C;
        uint32_t a = a0;
        for (int j = 0; j < p.k; ++j, a += step) {
          if constexpr (NARROW) {
            const uint32_t b = w_lo + (uint32_t)(j >> 1) * (WT >> 4) + (uint32_t)(j & 1) * 4;
            umma_bf16_w(d_main, a, b, idesc_m, j != 0);
            umma_bf16_w(d_main, a + 2, b + 2, idesc_m, 1);
            umma_bf16_w(d_corr, a + 4, b, idesc_h, 1);
            umma_bf16_w(d_corr, a + 6, b + 2, idesc_h, 1);
          } else {
            const uint32_t b = w_lo + (uint32_t)j * (WT >> 4), al = a + lo_off;
            umma_bf16_w(d_main, a, b, idesc_m, j != 0);
            umma_bf16_w(d_main, a + 2, b + 2, idesc_m, 1);
            umma_bf16_w(d_main, a + 4, b + 4, idesc_m, 1);
            umma_bf16_w(d_main, a + 6, b + 6, idesc_m, 1);
            umma_bf16_w(d_corr, al, b, idesc_h, 1);
            umma_bf16_w(d_corr, al + 2, b + 2, idesc_h, 1);
            umma_bf16_w(d_corr, al + 4, b + 4, idesc_h, 1);
            umma_bf16_w(d_corr, al + 6, b + 6, idesc_h, 1);
          }
        }
      };
      auto issue1 = [&](int i) {
        const int ab = i & 1;
        mbar_wait(t1empty0 + 8 * ab, ((i >> 1) & 1) ^ 1);  // middle warps have read this accumulator
        mbar_wait(afull0 + 8 * as, aph);
        tc_fence_after();
        conv(tmem_base + ab * 2 * C, desc_lo(smem0 + as * a_stage), step1, w1_lo, (uint32_t)q.panel_bytes >> 4);
        umma_commit(aempty0 + 8 * as);
        umma_commit(t1full0 + 8 * ab);
        if (++as == q.n_a) { as = 0; aph ^= 1; }
      };
      auto issue2 = [&](int i) {
        const int ab = i & 1, pb = i % NT;
        mbar_wait(t2empty0 + 8 * ab, ((i >> 1) & 1) ^ 1);  // final warps have read this accumulator
        mbar_wait(pfull0 + 8 * pb, (i / NT) & 1);          // t panel written (generic proxy -> fence.proxy.async)
        tc_fence_after();
        conv(tmem_base + 4 * C + ab * 2 * C, desc_lo(smem0 + t_off + pb * TP), step2, w2_lo, (uint32_t)T_BYTES >> 4);
        umma_commit(pempty0 + 8 * pb);
        umma_commit(t2full0 + 8 * ab);
      };
      if (my_tiles > 0) issue1(0);
      for (int i = 0; i < my_tiles; ++i) {
        if (i + 1 < my_tiles) issue1(i + 1);
        issue2(i);
      }
    }
  } else if (warp < 2 + MID_WARPS) {
    // ================= middle warps: accumulator 1 -> t panel =================
    const int lg = warp & 3;
    const int r = lg * 32 + lane;  // row of t this thread owns
    const uint32_t lane_base = tmem_base + ((uint32_t)(lg * 32) << 16);
    const uint32_t sw = (uint32_t)(r & 7);
    int i = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++i) {
      int mt, b;
      decode(tile, mt, b);
      const int ab = i & 1, pb = i % NT;
      mbar_wait(t1full0 + 8 * ab, (i >> 1) & 1);
      tc_fence_after();
      const int g = mt * q.TM - p2 + r;  // global row of this t row: outside [0, S) it is c2's zero padding
      const bool inside = g >= 0 && g < p.S_out;
      uint8_t* const row = smem_al + t_off + pb * TP + r * ROW_BYTES;
#pragma unroll
      for (int h = 0; h < C / 32; ++h) {  // 32 channels at a time
        float v[32], cr[32];
        tmem_ld32x2(lane_base + ab * 2 * C + 32 * h, v, lane_base + ab * 2 * C + C + 32 * h, cr);
        if (h == C / 32 - 1) {  // accumulator 1 is in registers: hand it back
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(t1empty0 + 8 * ab);
        }
        uint32_t hi[16], lo[16];
#pragma unroll
        for (int c = 0; c < 32; c += 4) {
          const float4 b4 = __ldg(reinterpret_cast<const float4*>(q.b1 + 32 * h + c));
          const float bb[4] = {b4.x, b4.y, b4.z, b4.w};
#pragma unroll
          for (int u = 0; u < 4; u += 2) {
            float f0 = (v[c + u] + cr[c + u]) + bb[u], f1 = (v[c + u + 1] + cr[c + u + 1]) + bb[u + 1];
            f0 = f0 > 0.f ? f0 : f0 * q.slope;
            f1 = f1 > 0.f ? f1 : f1 * q.slope;
            if (!inside) f0 = f1 = 0.f;
            const uint32_t hh = pack_bf16(f0, f1);
            float a0, a1;
            unpack_bf16(hh, a0, a1);
            hi[(c + u) >> 1] = hh;
            lo[(c + u) >> 1] = pack_bf16(f0 - a0, f1 - a1);
          }
        }
        if (h == 0) mbar_wait(pempty0 + 8 * pb, ((i / NT) & 1) ^ 1);  // c2 of the tile before has read this panel
        // physical 16-byte chunk = logical ^ (r & 7).  32 channels: logical chunks 0..3 = hi, 4..7 = lo of ONE panel;
        // 64 channels: chunks 4h..4h+3 of the hi panel and of the lo panel
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          const uint32_t ch = NARROW ? (uint32_t)c : (uint32_t)(4 * h + c), cl = NARROW ? (uint32_t)(c + 4) : ch;
          uint8_t* const lo_row = NARROW ? row : row + T_BYTES;
          *reinterpret_cast<uint4*>(row + ((ch ^ sw) << 4)) = make_uint4(hi[4 * c], hi[4 * c + 1], hi[4 * c + 2], hi[4 * c + 3]);
          *reinterpret_cast<uint4*>(lo_row + ((cl ^ sw) << 4)) = make_uint4(lo[4 * c], lo[4 * c + 1], lo[4 * c + 2], lo[4 * c + 3]);
        }
      }
      fence_proxy_async();  // generic-proxy writes -> visible to the MMA's operand reads
      __syncwarp();
      if (lane == 0) mbar_arrive(pfull0 + 8 * pb);
    }
  } else {
    // ================= final warps: accumulator 2 -> global memory; set e = tiles with i % NSET == e =================
    const int w = warp - 2 - MID_WARPS;
    const int e = w >> 2, lg = warp & 3;
    float* stg = stg_all + w * 32 * STG_PITCH;
    int i = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++i) {
      if (i % NSET != e) continue;
      int mt, b;
      decode(tile, mt, b);
      const int ab = i & 1;
      const uint32_t acc = tmem_base + ((uint32_t)(lg * 32) << 16) + 4 * C + ab * 2 * C;
      const int row0 = mt * q.TM, nrows = min(q.TM, p.S_out - row0);
      prefetch_epilogue_rows(p, lg, lane, b, row0, nrows, 0, C);
      mbar_wait(t2full0 + 8 * ab, (i >> 1) & 1);
      tc_fence_after();
#pragma unroll 1
      for (int c0 = 0; c0 < C; c0 += 32) {
        float v[32], cr[32];
        tmem_ld32x2(acc + c0, v, acc + C + c0, cr);
#pragma unroll
        for (int c = 0; c < 32; ++c) v[c] += cr[c];
        if (c0 + 32 >= C) {
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(t2empty0 + 8 * ab);
        }
        drain_block_dense(p, v, stg, lg, lane, b, row0, nrows, c0);
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    __syncwarp();
    tc_fence_after();
    tmem_dealloc(tmem_base, (uint32_t)U::tmem_cols);
  }
}

// Shared-memory plan; false when the layer does not fit.
template <int C>
bool plan(int k, int dilation, int ragged_b, UnitArgs* q) {
  using U = Unit<C>;
  if (k < 1 || k - 1 > U::t_rows - 128 || (k & 1) == 0 || dilation < 1) return false;
  q->R1 = 128 + (k - 1) * dilation;
  if (q->R1 > 256) return false;
  q->panel_bytes = (q->R1 * ROW_BYTES + 1023) & ~1023;
  q->npair = U::tiles_per_conv(k);
  q->TM = 128 - (k - 1);
  q->dilation = dilation;
  const int avail = MAX_SMEM - U::fixed_smem - 2 * q->npair * U::WT - (ragged_b > 0 ? (ragged_b + 1) * 4 : 0);
  q->n_a = avail / (U::np * q->panel_bytes);
  if (q->n_a > MAX_A) q->n_a = MAX_A;
  return q->n_a >= 2;
}

template <int C>
int run(const ss_resunit_params* p, void* stream) {
  using U = Unit<C>;
  UnitArgs q;
  SS_REQUIRE(plan<C>(p->k, p->dilation, p->rows ? p->B : 0, &q), SS_ERR_UNSUPPORTED,
             "ss_resunit_bf16x3: C=%d k=%d dilation=%d is not handled", C, p->k, p->dilation);
  q.b1 = p->b1;
  q.slope = p->slope;

  TcArgs a;  // what the final epilogue (drain_block_dense) reads
  a.bias = p->b2; a.residual = p->residual; a.lens = nullptr; a.y = p->y;
  a.y_act = static_cast<__nv_bfloat16*>(p->y_pair); a.accum = p->accum;
  a.pair_c = p->y_pair ? C : 0;
  a.y_batch_stride = (int64_t)p->S * C; a.y_row_stride = C; a.y_offset = 0; a.y_batch_elems = (int64_t)p->S * C;
  a.B = p->B; a.S_out = p->S; a.Cin = C; a.Cout = C; a.k = p->k; a.dilation = 1; a.pad = 0;
  a.n_tiles = 1;
  a.m_tiles = (p->S + q.TM - 1) / q.TM;
  a.act = SS_ACT_NONE; a.act2 = p->act2; a.y_dtype = SS_F32; a.residual_dtype = SS_F32;
  a.accum_init = p->accum_init; a.slope = 0.f; a.slope2 = p->slope2; a.accum_scale = p->accum_scale;
  a.rows = p->rows;
  a.l2_prefetch = x3_prefetch_default();
  const int64_t tiles = (int64_t)a.B * a.m_tiles;
  SS_REQUIRE(tiles < (1ll << 31), SS_ERR_UNSUPPORTED, "ss_resunit_bf16x3: too many tiles");

  CUtensorMap tmA, tmW1, tmW2;
  int rc = make_map_3d(&tmA, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, p->x, (uint64_t)2 * C, (uint64_t)p->S, (uint64_t)p->B,
                       (uint64_t)2 * C * 2, (uint64_t)p->S * 2 * C * 2, 64, (uint32_t)q.R1);
  if (rc != SS_OK) return rc;
  rc = make_map_3d(&tmW1, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, p->w1, 64, (uint64_t)2 * C, (uint64_t)q.npair, 128,
                   (uint64_t)2 * C * 128, 64, (uint32_t)(2 * C));
  if (rc != SS_OK) return rc;
  rc = make_map_3d(&tmW2, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, p->w2, 64, (uint64_t)2 * C, (uint64_t)q.npair, 128,
                   (uint64_t)2 * C * 128, 64, (uint32_t)(2 * C));
  if (rc != SS_OK) return rc;

  static std::atomic<unsigned long long> attr_mask{0};  // per device
  if (needs_attr_on_device(attr_mask)) {
    cudaError_t e = cudaFuncSetAttribute(resunit_kernel<C>, cudaFuncAttributeMaxDynamicSharedMemorySize, MAX_SMEM);
    if (e != cudaSuccess) return ss::cuda_fail(e, "ss_resunit_bf16x3: smem opt-in");
    attr_done_on_device(attr_mask);
  }
  const int sms = sm_count();
  SS_REQUIRE(sms > 0, SS_ERR_CUDA, "ss_resunit_bf16x3: no CUDA device");
  const int grid = (int)(tiles < sms ? tiles : sms);
  const int smem = U::fixed_smem + q.n_a * U::np * q.panel_bytes + 2 * q.npair * U::WT + (a.rows ? (a.B + 1) * 4 : 0);
  cudaError_t le = launch_pdl(resunit_kernel<C>, grid, U::nthreads, smem, static_cast<cudaStream_t>(stream), tmA, tmW1,
                              tmW2, a, q);
  if (le != cudaSuccess) return ss::cuda_fail(le, "ss_resunit_bf16x3: launch");
  SS_CHECK_LAUNCH("ss_resunit_bf16x3");
  return SS_OK;
}

}  // namespace

extern "C" int ss_resunit_bf16x3_supported(int channels, int k, int dilation) {
  UnitArgs q;
  if (channels == 32) return plan<32>(k, dilation, RAGGED_MAX_B, &q) ? 1 : 0;
  if (channels == 64) return plan<64>(k, dilation, RAGGED_MAX_B, &q) ? 1 : 0;
  return 0;
}

extern "C" int ss_resunit_bf16x3(const ss_resunit_params* p, void* stream) {
  SS_REQUIRE(p && p->x && p->w1 && p->w2 && p->b1 && p->b2 && (p->y || p->y_pair || p->accum), SS_ERR_INVALID,
             "ss_resunit_bf16x3: null pointer");
  SS_REQUIRE(p->B > 0 && p->S > 0 && (p->C == 32 || p->C == 64), SS_ERR_UNSUPPORTED,
             "ss_resunit_bf16x3: B=%d S=%d C=%d (C must be 32 or 64)", p->B, p->S, p->C);
  SS_REQUIRE(p->rows == nullptr || p->B <= RAGGED_MAX_B, SS_ERR_UNSUPPORTED,
             "ss_resunit_bf16x3: the ragged schedule (rows) takes at most %d utterances per call", RAGGED_MAX_B);
  SS_REQUIRE(ss::aligned16(p->x) && ss::aligned16(p->w1) && ss::aligned16(p->w2) && ss::aligned16(p->y) &&
                 ss::aligned16(p->y_pair) && ss::aligned16(p->b1) && ss::aligned16(p->b2) &&
                 ss::aligned16(p->residual) && ss::aligned16(p->accum),
             SS_ERR_ALIGNMENT, "ss_resunit_bf16x3: pointers must be 16-byte aligned");
  SS_REQUIRE(p->y_pair == nullptr || p->y_pair != p->x, SS_ERR_INVALID,
             "ss_resunit_bf16x3: y_pair must not alias x (other tiles still read their halo rows of x)");
  return p->C == 32 ? run<32>(p, stream) : run<64>(p, stream);
}
