// 3xTF32 implicit-GEMM convolution as a CTA PAIR (tcgen05 cta_group::2): two CTAs of a 2-CTA cluster compute one
// M = 256 x N = 128 tile.  Same mathematics, accuracy measures and epilogue as conv1d_tf32_ts.cu; what changes is
// who holds the weights.  With one CTA per tile every SM streams the whole [128 x 32] W_hi / W_lo chunk through its
// own shared memory (48 KB of MMA reads + 32 KB of TMA fill per chunk, 112 KB with the activations = 875 cycles of
// the 128 B/clk shared-memory pipe against 768 cycles of tensor time).  In the pair form each SM holds HALF of the
// weight tile (64 of the 128 output channels): the M = 256 MMA reads the two halves from the two SMs, so per SM it
// is 24 KB of reads + 16 KB of fill (72 KB per chunk in all).  Measured: bit-identical results, ~4 % faster on long
// reductions (282 vs 271 TFLOP/s on 1024->1024 k3), slower on 8-chunk tiles where the cluster-wide hand-overs
// weigh more - ops.py picks it from 16 chunks up (DESIGN.md 3.1 has the whole story, including what it disproved).
//   * CTA r of the pair owns rows [128 r, 128 r + 128) of the tile: its own activation slab by TMA, its own split
//     warps writing hi / lo into its own TMEM, its own accumulators (same TMEM columns in both SMs), its own
//     epilogue.  It loads weight rows [64 r, 64 r + 64).
//   * Only the leader (rank 0) issues MMAs (tcgen05.mma.cta_group::2, A from TMEM).  It learns that the peer's
//     operands are in place from the peer's split warps arriving REMOTELY on the leader's barrier (they have
//     waited on their own TMA barrier first, so the peer's weight half has landed too), and that the peer's
//     epilogue has drained an accumulator the same way.  Completion goes back to both CTAs with a multicast
//     tcgen05.commit.
#include "tc_common.cuh"

namespace {

using namespace sstc;

constexpr int KCHUNK = 32;  // fp32 elements per 128-byte swizzle row
// Eight epilogue warps: two per TMEM lane group, each owning 64 of the tile's 128 columns (accumulator row of 64 floats
// per thread instead of 128).  With four, the epilogue of a 128 x 128 tile took ~18 k cycles - longer than the main loop
// of a 12-chunk tile (128 channels, k = 3), and because the correction accumulator is single-buffered the next tile's
// MMAs waited ~8 k cycles for the slower CTA's epilogue (scripts/conv_trace.py ts2 128 128 3).
constexpr int EPI_WARPS = 8;
constexpr int NTHREADS = 32 * (2 + EPI_WARPS + 4);
constexpr int EPI_STG_BYTES = EPI_WARPS * STG_WARP_BYTES;
constexpr int BN2 = 128;      // tile width; each CTA of the pair holds BN2 / 2 weight rows
constexpr int STAGES2 = 5;  // 5 x 32 KB + 8 staging tiles
constexpr int A_SLOTS = 2;
constexpr int A_SLOT_COLS = 2 * KCHUNK;  // hi | lo

template <int BN, int STAGES>
struct Cfg {
  static constexpr int b_bytes = (BN / 2) * ROW_BYTES;  // this CTA's half of the weight tile
  static constexpr int stage_bytes = A_BYTES + 2 * b_bytes;  // x (fp32), W_hi, W_lo
  static constexpr int a_col0 = 3 * BN;                      // after main0, main1, corr
  static constexpr int used_cols = a_col0 + A_SLOTS * A_SLOT_COLS;
  static constexpr int tmem_cols = used_cols <= 32 ? 32 : used_cols <= 64 ? 64 : used_cols <= 128 ? 128 : used_cols <= 256 ? 256 : 512;
  static_assert(used_cols <= 512, "TMEM budget");
  static constexpr int n_bars = 2 * STAGES + 2 * A_SLOTS + 4 + 2;
  static constexpr int ragged_off = STAGES * stage_bytes + EPI_STG_BYTES + n_bars * 8 + 16;  // tile prefix table (ragged schedule)
  static constexpr int smem_bytes = ragged_off + 1024;  // + (B + 1) * 4 at launch when rows are given
  static_assert(smem_bytes + (RAGGED_MAX_B + 1) * 4 <= 227 * 1024, "shared memory budget");
};

// Debug build (-DSS_CONV2_TRACE, scripts/conv_trace.py): CTA 0 stamps clock64() per K-chunk at the hand-overs.
#ifdef SS_CONV2_TRACE
__device__ long long g_conv_ts2_trace[12 * 256];
#define CONV_TRACE(ev, i)                                                              \
  do {                                                                                 \
    if (blockIdx.x == 0 && (i) < 256) g_conv_ts2_trace[(ev) * 256 + (i)] = clock64();           \
  } while (0)
#else
#define CONV_TRACE(ev, i)
#endif


__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {  // every thread of both CTAs
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// arrive on the barrier at the same shared-memory offset in CTA `cta` of the cluster
__device__ __forceinline__ void mbar_arrive_cta(uint32_t bar, uint32_t cta) {
  asm volatile(
      "{\n\t"
      ".reg .b32 ra;\n\t"
      "mapa.shared::cluster.u32 ra, %0, %1;\n\t"
      "mbarrier.arrive.shared::cluster.b64 _, [ra];\n\t"
      "}" ::"r"(bar),
      "r"(cta)
      : "memory");
}
__device__ __forceinline__ void mbar_wait_cluster(uint32_t bar, uint32_t parity) {  // arrivals come from both CTAs
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "WAIT_LOOP_C:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra WAIT_DONE_C;\n\t"
      "bra WAIT_LOOP_C;\n\t"
      "WAIT_DONE_C:\n\t"
      "}" ::"r"(bar),
      "r"(parity)
      : "memory");
}
__device__ __forceinline__ void tmem_alloc2(uint32_t slot_smem, uint32_t cols) {
  asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(slot_smem), "r"(cols) : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc2(uint32_t taddr, uint32_t cols) {
  asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(cols) : "memory");
}
// D[tmem, both SMs] (+)= A[tmem, both SMs] . B[smem halves of both SMs]: M = 256, issued by the leader only
__device__ __forceinline__ void umma_tf32_ts2(uint32_t tmem_d, uint32_t tmem_a, uint64_t db, uint32_t idesc, uint32_t accum) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::tf32 [%0], [%1], %2, %3, {%5, %5, %5, %5, %5, %5, %5, %5}, p;\n\t"
      "}" ::"r"(tmem_d),
      "r"(tmem_a), "l"(db), "r"(idesc), "r"(accum), "r"(0u)
      : "memory");
}
// completion of all MMAs issued so far -> the barrier at this offset in BOTH CTAs
__device__ __forceinline__ void umma_commit2(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(bar),
               "h"((uint16_t)3)
               : "memory");
}


template <int BN, int STAGES>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(NTHREADS, 1)
conv1d_tf32_ts2_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmBh,
                      const __grid_constant__ CUtensorMap tmBl, const TcArgs p, const int in_act, const float in_slope,
                      const int group) {
  using C = Cfg<BN, STAGES>;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem0 = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* smem_al = smem_raw + (smem0 - smem_u32(smem_raw));
  auto off_a = [](int s) { return s * C::stage_bytes; };
  auto off_bh = [](int s) { return s * C::stage_bytes + A_BYTES; };
  auto off_bl = [](int s) { return s * C::stage_bytes + A_BYTES + C::b_bytes; };
  float* stg_all = reinterpret_cast<float*>(smem_al + STAGES * C::stage_bytes);
  const uint32_t bars = smem0 + STAGES * C::stage_bytes + EPI_STG_BYTES;
  const uint32_t full0 = bars, empty0 = full0 + 8 * STAGES, tafull0 = empty0 + 8 * STAGES,
                 taempty0 = tafull0 + 8 * A_SLOTS, tfull0 = taempty0 + 8 * A_SLOTS, tempty0 = tfull0 + 16,
                 cfull = tempty0 + 16, cempty = cfull + 8, tslot = cempty + 8;
  volatile uint32_t* tslot_ptr =
      reinterpret_cast<volatile uint32_t*>(smem_al + STAGES * C::stage_bytes + EPI_STG_BYTES + C::n_bars * 8);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nchunk = (p.Cin + KCHUNK - 1) / KCHUNK;
  const int KT = p.k * nchunk;
  // p.m_tiles counts PAIR tiles (256 rows); this CTA's 128-row tile is 2 * pair_tile + rank
  int* const rag = reinterpret_cast<int*>(smem_al + C::ragged_off);
  const bool ragged = p.rows != nullptr;
  if (ragged) ragged_build(p.rows, p.B, p.S_out, 2 * BM, rag);  // both CTAs of the pair build the same table
  const uint32_t cta_rank = cluster_ctarank();
  const int pair0 = blockIdx.x >> 1, pair_step = gridDim.x >> 1;

  if (threadIdx.x == 0) {
    prefetch_tmap(&tmA);
    prefetch_tmap(&tmBh);
    prefetch_tmap(&tmBl);
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(full0 + 8 * s, 1);
      mbar_init(empty0 + 8 * s, 1);
    }
    for (int t = 0; t < A_SLOTS; ++t) {
      mbar_init(tafull0 + 8 * t, 8);  // one arrival per split warp of BOTH CTAs (used in the leader only)
      mbar_init(taempty0 + 8 * t, 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(tfull0 + 8 * a, 1);
      mbar_init(tempty0 + 8 * a, 2 * EPI_WARPS);  // epilogue warps of both CTAs
    }
    mbar_init(cfull, 1);
    mbar_init(cempty, 2 * EPI_WARPS);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) tmem_alloc2(tslot, (uint32_t)C::tmem_cols);
  tc_fence_before();
  cluster_sync_all();  // barriers of both CTAs initialised, TMEM of both SMs allocated
  tc_fence_after();
  const uint32_t tmem_base = *tslot_ptr;
  const int num_tiles = ragged ? rag[p.B] * p.n_tiles : (p.flat ? 1 : p.B) * p.m_tiles * p.n_tiles;
  // linear pair-tile index -> (column tile, this CTA's 128-row tile, utterance)
  auto decode = [&](int tile, int& nt, int& mt, int& b) {
    nt = tile % p.n_tiles;
    const int rest = tile / p.n_tiles;
    int pm;
    if (ragged) {
      ragged_decode(rag, p.B, rest, b, pm);
    } else {
      pm = rest % p.m_tiles;
      b = rest / p.m_tiles;
    }
    mt = pm * 2 + (int)cta_rank;
  };

  if (warp == 0) {
    // ================= TMA producer =================
    if (elect_one()) {  // elect.sync, not lane == 0: lets ptxas issue TMA / MMA without a per-instruction elect loop
      int stage = 0;
      uint32_t phase = 0;
      [[maybe_unused]] int gi = 0;
      for (int tile = pair0; tile < num_tiles; tile += pair_step) {
        int nt, mt, b;
        decode(tile, nt, mt, b);
        const int t0 = mt * BM, n0 = nt * BN;
        for (int it = 0; it < KT; ++it) {
          const int j = it / nchunk;
          const int c0 = (it - j * nchunk) * KCHUNK;
          mbar_wait(empty0 + 8 * stage, phase ^ 1);
          CONV_TRACE(0, gi); ++gi;
          mbar_expect_tx(full0 + 8 * stage, C::stage_bytes);
          tma_load_3d(smem0 + off_a(stage), &tmA, full0 + 8 * stage, c0, t0 + j * p.dilation - p.pad, b);
          tma_load_3d(smem0 + off_bh(stage), &tmBh, full0 + 8 * stage, c0, n0 + (int)cta_rank * (BN / 2), j);
          tma_load_3d(smem0 + off_bl(stage), &tmBl, full0 + 8 * stage, c0, n0 + (int)cta_rank * (BN / 2), j);
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ================= MMA issuer (leader CTA only) =================
    if (cta_rank == 0 && elect_one()) {
      const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(256 >> 4) << 24);  // tf32, M = 256
      int stage = 0, slot = 0;
      uint32_t phase = 0, sphase = 0;
      int local = 0, gcount = 0;
      [[maybe_unused]] int gi = 0;
      for (int tile = pair0; tile < num_tiles; tile += pair_step, ++local) {
        mbar_wait_cluster(cempty, (local & 1) ^ 1);  // both epilogues have drained the correction accumulator
        CONV_TRACE(8, gi);
        const uint32_t c_tmem = tmem_base + 2 * BN;
        for (int it0 = 0; it0 < KT; it0 += group, ++gcount) {
          const int as = gcount & 1;
          mbar_wait_cluster(tempty0 + 8 * as, ((gcount >> 1) & 1) ^ 1);
          tc_fence_after();
          CONV_TRACE(9, gi);
          const uint32_t d_tmem = tmem_base + as * BN;
          const int it1 = min(it0 + group, KT);
          for (int it = it0; it < it1; ++it) {
            mbar_wait_cluster(tafull0 + 8 * slot, sphase);  // hi / lo in the TMEM of both SMs (so both TMA fills landed)
            tc_fence_after();
            CONV_TRACE(3, gi);
            const uint32_t a_hi = tmem_base + C::a_col0 + slot * A_SLOT_COLS, a_lo = a_hi + KCHUNK;
            const uint32_t b_hi = smem0 + off_bh(stage), b_lo = smem0 + off_bl(stage);
#pragma unroll
            for (int kk = 0; kk < KCHUNK / 8; ++kk) {  // K = 8 tf32 per instruction: 8 TMEM columns / 32 bytes of B
              umma_tf32_ts2(c_tmem, a_lo + kk * 8, make_desc(b_hi + kk * 32), idesc, (it | kk) != 0);
              umma_tf32_ts2(c_tmem, a_hi + kk * 8, make_desc(b_lo + kk * 32), idesc, 1u);
              umma_tf32_ts2(d_tmem, a_hi + kk * 8, make_desc(b_hi + kk * 32), idesc, ((it - it0) | kk) != 0);
            }
            umma_commit2(empty0 + 8 * stage);    // smem stage free in both CTAs
            umma_commit2(taempty0 + 8 * slot);   // TMEM A slot free in both CTAs
            CONV_TRACE(4, gi); ++gi;
            if (++stage == STAGES) { stage = 0; phase ^= 1; }
            if (++slot == A_SLOTS) { slot = 0; sphase ^= 1; }
          }
          umma_commit2(tfull0 + 8 * as);
        }
        umma_commit2(cfull);
      }
    }
  } else if (warp < 2 + EPI_WARPS) {
    // ================= epilogue (warps 2..9): warp w and w + 4 share a lane group, 64 columns each =================
    const int chalf = (warp - 2) >> 2;          // which 64-column half of the tile this warp owns
    constexpr int CW = BN / 2;                  // columns per epilogue warp
    const int lg = warp & 3;
    float* stg = stg_all + (warp - 2) * 32 * STG_PITCH;
    const uint32_t lane_base = tmem_base + ((uint32_t)(lg * 32) << 16);
    int local = 0, gcount = 0;
    for (int tile = pair0; tile < num_tiles; tile += pair_step, ++local) {
      int nt, mt, b;
      decode(tile, nt, mt, b);
      int b_t = b, t_tile0 = mt * BM;
      bool straddle = false;
      if (p.flat) {  // un-flatten the tile's first row once; most tiles lie inside one utterance
        b_t = t_tile0 / p.S_out;
        t_tile0 -= b_t * p.S_out;
        straddle = t_tile0 + BM > p.S_out;
      }
      const int len = p.lens ? p.lens[b_t] : 0x7fffffff;
      const bool lean = drain_is_lean<true>(p);  // warp-uniform
      const bool pre_accum = lean && p.accum != nullptr && !p.accum_init;  // running mean: read-modify-write
      if (lean && (p.residual || pre_accum)) {
        // Pull this tile's residual rows into L2 now, while the MMAs run: the drain below keeps only four 16-byte
        // loads per lane in flight, which from HBM (~1 us) capped the whole kernel at ~1.2 TB/s of residual reads.
        int t = t_tile0 + lg * 32 + lane, bb = b_t;
        if (straddle)
          while (t >= p.S_out) { t -= p.S_out; ++bb; }
        const int64_t off = p.y_offset + (int64_t)t * p.y_row_stride + nt * BN + chalf * CW;
        if (bb < p.B && t < p.S_out && off >= 0 && off < p.y_batch_elems) {
          const int64_t g0 = (int64_t)bb * p.y_batch_stride + off;
          const int ncol = min(CW, p.Cout - nt * BN - chalf * CW);
          for (int c = 0; c < ncol; c += 32) {  // one 128-byte line per step
            if (p.residual) asm volatile("prefetch.global.L2 [%0];" ::"l"(static_cast<const float*>(p.residual) + g0 + c));
            if (pre_accum) asm volatile("prefetch.global.L2 [%0];" ::"l"(p.accum + g0 + c));
          }
        }
      }
      float acc[CW];
#pragma unroll
      for (int i = 0; i < CW; ++i) acc[i] = 0.f;
      for (int it0 = 0; it0 < KT; it0 += group, ++gcount) {
        if (it0 + group >= KT) {
          // Last group: its MMAs and the correction terms finish together.  Take the (single-buffered) correction
          // accumulator first and hand it back at once - the next tile's MMAs wait for it, and with a short
          // reduction (k = 1 layers: 8 chunks) anything slower would stall the tensor pipe between tiles.
          mbar_wait(cfull, local & 1);
          tc_fence_after();
          if (threadIdx.x == 64) CONV_TRACE(10, local * KT + KT - 1);
#pragma unroll
          for (int c0 = 0; c0 < CW; c0 += 32) {
            float v[32];
            tmem_ld32(lane_base + 2 * BN + chalf * CW + c0, v);
#pragma unroll
            for (int i = 0; i < 32; ++i) acc[c0 + i] += v[i];
          }
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive_cta(cempty, 0);
          if (threadIdx.x == 64) CONV_TRACE(11, local * KT + KT - 1);
        }
        const int as = gcount & 1;
        mbar_wait(tfull0 + 8 * as, (gcount >> 1) & 1);
        tc_fence_after();
#pragma unroll
        for (int c0 = 0; c0 < CW; c0 += 32) {
          float v[32];
          tmem_ld32(lane_base + as * BN + chalf * CW + c0, v);
#pragma unroll
          for (int i = 0; i < 32; ++i) acc[c0 + i] += v[i];  // round-to-nearest fp32 adds
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive_cta(tempty0 + 8 * as, 0);
        if (threadIdx.x == 64) CONV_TRACE(7, local * KT + min(it0 + group, KT) - 1);  // stamped at the group's last chunk
      }
#pragma unroll
      for (int c0 = 0; c0 < CW; c0 += 32) {  // TMEM is free again: the stores overlap the next tile's MMAs
        float v[32];
#pragma unroll
        for (int i = 0; i < 32; ++i) v[i] = acc[c0 + i];
        if (lean)
          drain_block_lean<true>(p, v, stg, lg, lane, b_t, t_tile0, straddle, nt * BN + chalf * CW + c0, len);
        else
          drain_block(p, v, stg, lg, lane, b, mt, nt * BN + chalf * CW + c0, len);
      }
    }
  } else {
    // ================= operand split (warps 10..13): thread = row, shared memory -> registers -> TMEM =========
    const int lg = warp & 3;  // TMEM lane group this warp may write
    const int r = lg * 32 + lane;
    const uint32_t lane_base = tmem_base + ((uint32_t)(lg * 32) << 16) + C::a_col0;
    const int sw = r & 7;
    int stage = 0, slot = 0;
    uint32_t phase = 0, sphase = 0;
    [[maybe_unused]] int gi = 0;
    for (int tile = pair0; tile < num_tiles; tile += pair_step) {
      // flat tiling: frame index of this row inside its own utterance, and whether any row of the tile (with its
      // halo) can reach across an utterance end - only then do taps have to be checked row by row
      int t_row = 0;
      bool edge_tile = false;
      if (p.flat) {
        const int R0 = (((tile / p.n_tiles) % p.m_tiles) * 2 + (int)cta_rank) * BM;
        const int t0 = R0 % p.S_in;
        t_row = t0 + r;
        if (t_row >= p.S_in) t_row %= p.S_in;
        edge_tile = t0 - p.pad < 0 || t0 + BM - 1 + (p.k - 1) * p.dilation - p.pad >= p.S_in;
      }
      int t_src = t_row - p.pad, cleft = nchunk;  // source frame of this row for the current tap
      for (int it = 0; it < KT; ++it) {
        mbar_wait(full0 + 8 * stage, phase);
        if (threadIdx.x == 320) CONV_TRACE(1, gi);
        const uint8_t* row = smem_al + off_a(stage) + r * ROW_BYTES;
        // flat tiling: a tap that leaves this row's utterance must read the zero padding, not the neighbour's frames
        const bool outside = edge_tile && (t_src < 0 || t_src >= p.S_in);
        if (--cleft == 0) { cleft = nchunk; t_src += p.dilation; }
        float hi[32], lo[32];
#pragma unroll
        for (int u = 0; u < 8; ++u) {  // 16-byte unit u of the row sits at unit (u ^ (row & 7)) - SWIZZLE_128B
          float4 x = *reinterpret_cast<const float4*>(row + ((u ^ sw) << 4));
          if (in_act == SS_ACT_LRELU) {
            x.x = fmaxf(x.x, x.x * in_slope);
            x.y = fmaxf(x.y, x.y * in_slope);
            x.z = fmaxf(x.z, x.z * in_slope);
            x.w = fmaxf(x.w, x.w * in_slope);
          }
          hi[4 * u + 0] = tf32_rn(x.x); lo[4 * u + 0] = x.x - hi[4 * u + 0];  // truncated to TF32 by the MMA (see conv1d_tf32_ts.cu)
          hi[4 * u + 1] = tf32_rn(x.y); lo[4 * u + 1] = x.y - hi[4 * u + 1];  // truncated to TF32 by the MMA (see conv1d_tf32_ts.cu)
          hi[4 * u + 2] = tf32_rn(x.z); lo[4 * u + 2] = x.z - hi[4 * u + 2];  // truncated to TF32 by the MMA (see conv1d_tf32_ts.cu)
          hi[4 * u + 3] = tf32_rn(x.w); lo[4 * u + 3] = x.w - hi[4 * u + 3];  // truncated to TF32 by the MMA (see conv1d_tf32_ts.cu)
        }
        if (edge_tile && __any_sync(0xffffffffu, outside)) {  // warp-uniform branch: skipped in interior tiles
#pragma unroll
          for (int i = 0; i < 32; ++i) {
            hi[i] = outside ? 0.f : hi[i];
            lo[i] = outside ? 0.f : lo[i];
          }
        }
        if (threadIdx.x == 320) CONV_TRACE(5, gi);
        mbar_wait(taempty0 + 8 * slot, sphase ^ 1);  // the MMAs that read this TMEM slot have completed
        tc_fence_after();
        if (threadIdx.x == 320) CONV_TRACE(6, gi);
        tmem_st32(lane_base + slot * A_SLOT_COLS, hi);
        tmem_st32(lane_base + slot * A_SLOT_COLS + KCHUNK, lo);
        asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive_cta(tafull0 + 8 * slot, 0);
        if (threadIdx.x == 320) CONV_TRACE(2, gi);
        ++gi;
        if (++stage == STAGES) { stage = 0; phase ^= 1; }
        if (++slot == A_SLOTS) { slot = 0; sphase ^= 1; }
      }
    }
  }

  tc_fence_before();
  cluster_sync_all();  // no MMA of the pair may still be reading either SM's TMEM / shared memory
  if (warp == 1) {
    __syncwarp();
    tc_fence_after();
    tmem_dealloc2(tmem_base, (uint32_t)C::tmem_cols);
  }
}

int launch2(const CUtensorMap& tmA, const CUtensorMap& tmBh, const CUtensorMap& tmBl, TcArgs a, int in_act, float in_slope,
            int group, cudaStream_t st) {
  using C = Cfg<BN2, STAGES2>;
  a.n_tiles = (a.Cout + BN2 - 1) / BN2;
  const int64_t rows = a.flat ? (int64_t)a.B * a.S_out : a.S_out;
  a.m_tiles = (int)((rows + 2 * BM - 1) / (2 * BM));  // pair tiles of 256 rows
  const int64_t tiles = (int64_t)(a.flat ? 1 : a.B) * a.m_tiles * a.n_tiles;
  SS_REQUIRE(tiles < (1ll << 31), SS_ERR_UNSUPPORTED, "ss_conv1d_tf32x3_ts2: too many tiles");
  static std::atomic<unsigned long long> attr_mask{0};  // per device
  if (needs_attr_on_device(attr_mask)) {
    cudaError_t e = cudaFuncSetAttribute(conv1d_tf32_ts2_kernel<BN2, STAGES2>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         C::smem_bytes + (RAGGED_MAX_B + 1) * 4);
    if (e != cudaSuccess) return ss::cuda_fail(e, "ss_conv1d_tf32x3_ts2: smem opt-in");
    attr_done_on_device(attr_mask);
  }
  const int sms = sm_count();
  SS_REQUIRE(sms >= 2, SS_ERR_CUDA, "ss_conv1d_tf32x3_ts2: no CUDA device");
  const int pairs = (int)(tiles < sms / 2 ? tiles : sms / 2);
  const int smem = C::smem_bytes + (a.rows ? (a.B + 1) * 4 : 0);
  conv1d_tf32_ts2_kernel<BN2, STAGES2><<<2 * pairs, NTHREADS, smem, st>>>(tmA, tmBh, tmBl, a, in_act, in_slope, group);
  SS_CHECK_LAUNCH("ss_conv1d_tf32x3_ts2");
  return SS_OK;
}

}  // namespace

#ifdef SS_CONV2_TRACE
extern "C" __attribute__((visibility("default"))) int ss_debug_conv_ts2_trace(long long* host) {
  return (int)cudaMemcpyFromSymbol(host, g_conv_ts2_trace, sizeof(long long) * 12 * 256);
}
#endif

extern "C" int ss_conv1d_tf32x3_ts(const ss_conv1d_tf32_params* p, void* stream);

extern "C" int ss_conv1d_tf32x3_ts2(const ss_conv1d_tf32_params* p, void* stream) {
  SS_REQUIRE(p && p->x && p->w_hi && p->w_lo && (p->y || p->y_act || p->accum), SS_ERR_INVALID,
             "ss_conv1d_tf32x3_ts2: null pointer");
  SS_REQUIRE(p->B > 0 && p->S_in > 0 && p->S_out > 0 && p->Cin > 0 && p->Cout > 0 && p->k > 0 && p->dilation > 0,
             SS_ERR_INVALID, "ss_conv1d_tf32x3_ts2: bad shape B=%d S_in=%d S_out=%d Cin=%d Cout=%d k=%d dil=%d", p->B,
             p->S_in, p->S_out, p->Cin, p->Cout, p->k, p->dilation);
  SS_REQUIRE(p->Cin % 4 == 0 && p->Cout % 4 == 0, SS_ERR_UNSUPPORTED,
             "ss_conv1d_tf32x3_ts2: Cin (%d) and Cout (%d) must be multiples of 4", p->Cin, p->Cout);
  SS_REQUIRE(p->x_row_stride % 4 == 0 && p->x_batch_stride % 4 == 0 && p->y_row_stride % 4 == 0 &&
                 p->y_batch_stride % 4 == 0 && p->y_offset % 4 == 0 && p->y_batch_elems % 4 == 0,
             SS_ERR_ALIGNMENT, "ss_conv1d_tf32x3_ts2: strides/offsets must be multiples of 4 elements");
  SS_REQUIRE(ss::aligned16(p->x) && ss::aligned16(p->w_hi) && ss::aligned16(p->w_lo) && ss::aligned16(p->y) &&
                 ss::aligned16(p->y_act) && ss::aligned16(p->bias) && ss::aligned16(p->residual) && ss::aligned16(p->accum),
             SS_ERR_ALIGNMENT, "ss_conv1d_tf32x3_ts2: pointers must be 16-byte aligned");
  SS_REQUIRE(p->in_act == SS_ACT_NONE || p->in_act == SS_ACT_LRELU, SS_ERR_UNSUPPORTED,
             "ss_conv1d_tf32x3_ts2: in_act must be none or lrelu");
  SS_REQUIRE(p->in_act != SS_ACT_LRELU || (p->in_slope >= 0.f && p->in_slope <= 1.f), SS_ERR_UNSUPPORTED,
             "ss_conv1d_tf32x3_ts2: leaky-ReLU on load is computed as max(x, slope*x) and needs 0 <= slope <= 1 (got %g)", p->in_slope);

  if (p->Cout <= 64) return ss_conv1d_tf32x3_ts(p, stream);  // narrow layers: the single-CTA kernel's BN = 64 / 32 forms
  const int bn = BN2 / 2;  // rows of the weight tile each CTA loads
  // Flat tiling (see TcArgs::flat): stride-1 same-length layer, batch-contiguous input, lean epilogue, and a
  // sequence length that would otherwise leave a partly empty last tile in every utterance.
  SS_REQUIRE(p->rows == nullptr || p->B <= RAGGED_MAX_B, SS_ERR_UNSUPPORTED,
             "ss_conv1d_tf32x3_ts2: the ragged schedule (rows) takes at most %d utterances per call", RAGGED_MAX_B);
  const bool flat = p->rows == nullptr && p->B > 1 && p->S_in == p->S_out && p->S_out % (2 * BM) != 0 &&
                    p->x_batch_stride == (int64_t)p->S_in * p->x_row_stride && (int64_t)p->B * p->S_in < (1ll << 30) &&
                    p->y != nullptr && p->accum == nullptr && p->y_act == nullptr && p->act != SS_ACT_TANH;
  CUtensorMap tmA, tmBh, tmBl;
  int rc = flat ? make_map_3d(&tmA, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, p->x, (uint64_t)p->Cin,
                              (uint64_t)p->B * (uint64_t)p->S_in, 1, (uint64_t)p->x_row_stride * 4,
                              (uint64_t)p->B * (uint64_t)p->x_batch_stride * 4, KCHUNK, BM)
                : make_map_3d(&tmA, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, p->x, (uint64_t)p->Cin, (uint64_t)p->S_in,
                              (uint64_t)p->B, (uint64_t)p->x_row_stride * 4, (uint64_t)p->x_batch_stride * 4, KCHUNK, BM);
  if (rc != SS_OK) return rc;
  rc = make_map_3d(&tmBh, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, p->w_hi, (uint64_t)p->Cin, (uint64_t)p->Cout, (uint64_t)p->k,
                   (uint64_t)p->Cin * 4, (uint64_t)p->Cout * p->Cin * 4, KCHUNK, (uint32_t)bn);
  if (rc != SS_OK) return rc;
  rc = make_map_3d(&tmBl, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, p->w_lo, (uint64_t)p->Cin, (uint64_t)p->Cout, (uint64_t)p->k,
                   (uint64_t)p->Cin * 4, (uint64_t)p->Cout * p->Cin * 4, KCHUNK, (uint32_t)bn);
  if (rc != SS_OK) return rc;

  TcArgs a;
  a.bias = p->bias; a.residual = p->residual; a.lens = p->lens; a.y = p->y;
  a.y_act = static_cast<__nv_bfloat16*>(p->y_act); a.accum = p->accum;
  a.y_batch_stride = p->y_batch_stride; a.y_row_stride = p->y_row_stride; a.y_offset = p->y_offset;
  a.y_batch_elems = p->y_batch_elems;
  a.B = p->B; a.S_out = p->S_out; a.Cin = p->Cin; a.Cout = p->Cout; a.k = p->k; a.dilation = p->dilation; a.pad = p->pad;
  a.m_tiles = a.n_tiles = 0;
  a.act = p->act; a.act2 = p->act2; a.y_dtype = SS_F32; a.residual_dtype = SS_F32;
  a.accum_init = p->accum_init; a.slope = p->slope; a.slope2 = p->slope2; a.accum_scale = p->accum_scale;
  a.flat = flat ? 1 : 0; a.S_in = p->S_in; a.rows = p->rows;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int group = p->k_group > 0 ? p->k_group : 8;
  return launch2(tmA, tmBh, tmBl, a, p->in_act, p->in_slope, group, st);
}
