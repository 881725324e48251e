// 3xTF32 implicit-GEMM convolution, A operand in TENSOR MEMORY (the "TS" form of tcgen05.mma).
//
// Same mathematics, interface, accuracy measures (separate correction accumulator, main accumulator
// restarted every k_group chunks, partial tiles summed in registers) and epilogue as conv1d_tf32.cu;
// what changes is where the split activations live.  In conv1d_tf32.cu the split warps write hi and lo
// back to shared memory and every one of the 12 MMAs of a chunk re-reads a 4 KB A slice from it: the
// shared-memory pipe, not the tensor pipe, sets the pace (ncu: 59 % tensor, 53 % LSU wavefronts).
// Here the split warps read the TMA-landed fp32 tile ONCE (thread = row, un-swizzling the eight 16-byte
// units of its row), split it in registers and store hi / lo with tcgen05.st into TMEM, from where the
// MMAs take A for free; shared memory then only carries the TMA fill and the B (weight) reads.
// Stride-1 same-length layers are tiled over the flattened B*S rows (TcArgs::flat); the common epilogue is the
// lean straight-line one of tc_common.cuh; long reductions at BN = 128 go to the CTA-pair form
// (conv1d_tf32_ts2.cu).  DESIGN.md 3.1 has the measurements behind each of these.
//
// TMEM columns (BN = 128): main accumulator x2 (256) | correction accumulator (128) | A hi/lo, 2 slots x
// (32 + 32) columns (128) = 512.  Shared memory per stage: x tile 16 KB + W_hi + W_lo.
// Warps: TMA, MMA, 4 epilogue, 4 split (8 split warps for the narrow tiles BN = 64 / 32).
#include "tc_common.cuh"

namespace {

using namespace sstc;

constexpr int KCHUNK = 32;  // fp32 elements per 128-byte swizzle row
// Split warps: four (one per TMEM lane group) for BN = 128; eight for the narrow tiles, where the split (~650 cycles
// per chunk whatever N is) - not the N = 64 / 32 MMAs (384 / 192 cycles) - sets the pace: warps w and w+4 share a lane
// group and each splits half of the row's 32 channels.  (At BN = 128 the epilogue's 128-float row caps the kernel at
// three warps per scheduler.)
template <int BN>
constexpr int xf_warps() { return BN <= 64 ? 8 : 4; }
template <int BN>
constexpr int n_threads() { return 32 * (6 + xf_warps<BN>()); }
constexpr int A_SLOTS = 2;
constexpr int A_SLOT_COLS = 2 * KCHUNK;  // hi | lo

template <int BN, int STAGES>
struct Cfg {
  static constexpr int b_bytes = BN * ROW_BYTES;
  static constexpr int stage_bytes = A_BYTES + 2 * b_bytes;  // x (fp32), W_hi, W_lo
  // BN = 96 is the short-reduction form: with a single accumulator restart group per tile, the one correction
  // accumulator of the BN = 128 layout stalls the tensor pipe between tiles until the epilogue has drained it
  // (~3.6 k cycles against a 7 k-cycle tile at K = 256); 4 x 96 + 128 columns fit a second one.
  // The narrow tiles (BN = 64 / 32, the HiFi-GAN 64- and 32-channel stages) have the room as well, and need it more:
  // their tiles are short (3 to 22 chunks) and with one correction accumulator the next tile's MMAs could not start
  // before the epilogue - still busy storing the previous tile - had read it (scripts/conv_trace.py ts 32 32 3:
  // ~3.7 k cycles per tile against 1.1 k of main loop).
  static constexpr int corr_bufs = BN <= 96 ? 2 : 1;
  static constexpr int a_col0 = (2 + corr_bufs) * BN;        // after main0, main1, corr (x corr_bufs)
  static constexpr int used_cols = a_col0 + A_SLOTS * A_SLOT_COLS;
  static constexpr int tmem_cols = used_cols <= 32 ? 32 : used_cols <= 64 ? 64 : used_cols <= 128 ? 128 : used_cols <= 256 ? 256 : 512;
  static_assert(used_cols <= 512, "TMEM budget");
  static constexpr int n_bars = 2 * STAGES + 2 * A_SLOTS + 4 + 4;
  // the 96-column form stages a whole output row per epilogue thread (drain_tile_lean_wide)
  static constexpr int stg_pitch = BN == 96 ? BN + 4 : STG_PITCH;
  static constexpr int stg_bytes = 4 * 32 * stg_pitch * 4;
  static constexpr int ragged_off = STAGES * stage_bytes + stg_bytes + n_bars * 8 + 16;  // tile prefix table (ragged schedule)
  static constexpr int smem_bytes = ragged_off + 1024;  // + (B + 1) * 4 at launch when rows are given
  static_assert(smem_bytes + (RAGGED_MAX_B + 1) * 4 <= 227 * 1024, "shared memory budget");
};

// Debug build (-DSS_CONV_TRACE, scripts/conv_trace.py): CTA 0 stamps clock64() per K-chunk at the hand-overs.
#ifdef SS_CONV_TRACE
__device__ long long g_conv_ts_trace[12 * 256];
#define CONV_TRACE(ev, i)                                                              \
  do {                                                                                 \
    if (blockIdx.x == 0 && (i) < 256) g_conv_ts_trace[(ev) * 256 + (i)] = clock64();           \
  } while (0)
#else
#define CONV_TRACE(ev, i)
#endif

// lo = x - hi is handed to the MMA as is: the tensor core reads the top 19 bits of an fp32 operand, i.e. truncates lo
// to TF32 itself.  Rounding it to nearest first costs two more ALU operations per element on the split warps (which sit
// on the A-slot round trip that paces this kernel) and changes nothing measurable: rms / max error vs fp64 are identical
// to three digits (scripts/tf32_group_error.py), since lo only carries bits 12-22 of x.
#define LO_RN(v) (v)

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



template <int BN, int STAGES>
__global__ void __launch_bounds__(n_threads<BN>(), 1)
conv1d_tf32_ts_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmBh,
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
  const uint32_t bars = smem0 + STAGES * C::stage_bytes + C::stg_bytes;
  const uint32_t full0 = bars, empty0 = full0 + 8 * STAGES, tafull0 = empty0 + 8 * STAGES,
                 taempty0 = tafull0 + 8 * A_SLOTS, tfull0 = taempty0 + 8 * A_SLOTS, tempty0 = tfull0 + 16,
                 cfull0 = tempty0 + 16, cempty0 = cfull0 + 16, tslot = cempty0 + 16;
  volatile uint32_t* tslot_ptr =
      reinterpret_cast<volatile uint32_t*>(smem_al + STAGES * C::stage_bytes + C::stg_bytes + C::n_bars * 8);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nchunk = (p.Cin + KCHUNK - 1) / KCHUNK;
  const int KT = p.k * nchunk;
  int* const rag = reinterpret_cast<int*>(smem_al + C::ragged_off);
  const bool ragged = p.rows != nullptr;
  if (ragged) ragged_build(p.rows, p.B, p.S_out, BM, rag);

  if (threadIdx.x == 0) {
    prefetch_tmap(&tmA);
    prefetch_tmap(&tmBh);
    prefetch_tmap(&tmBl);
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(full0 + 8 * s, 1);
      mbar_init(empty0 + 8 * s, 1);
    }
    for (int t = 0; t < A_SLOTS; ++t) {
      mbar_init(tafull0 + 8 * t, xf_warps<BN>());  // one arrival per split warp
      mbar_init(taempty0 + 8 * t, 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(tfull0 + 8 * a, 1);
      mbar_init(tempty0 + 8 * a, 4);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(cfull0 + 8 * a, 1);
      mbar_init(cempty0 + 8 * a, 4);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) tmem_alloc(tslot, (uint32_t)C::tmem_cols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tslot_ptr;
  // flat: m_tiles spans all utterances; ragged: the prefix table counts each utterance's own tiles
  const int num_tiles = ragged ? rag[p.B] * p.n_tiles : (p.flat ? 1 : p.B) * p.m_tiles * p.n_tiles;
  // linear tile index -> (column tile, row tile, utterance)
  auto decode = [&](int tile, int& nt, int& mt, int& b) {
    nt = tile % p.n_tiles;
    const int rest = tile / p.n_tiles;
    if (ragged) {
      ragged_decode(rag, p.B, rest, b, mt);
    } else {
      mt = rest % p.m_tiles;
      b = rest / p.m_tiles;
    }
  };

  if (warp == 0) {
    // ================= TMA producer =================
    if (elect_one()) {  // elect.sync, not lane == 0: lets ptxas issue TMA / MMA without a per-instruction elect loop
      int stage = 0;
      uint32_t phase = 0;
      [[maybe_unused]] int gi = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
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
          tma_load_3d(smem0 + off_bh(stage), &tmBh, full0 + 8 * stage, c0, n0, j);
          tma_load_3d(smem0 + off_bl(stage), &tmBl, full0 + 8 * stage, c0, n0, j);
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ================= MMA issuer =================
    if (elect_one()) {  // elect.sync, not lane == 0: lets ptxas issue TMA / MMA without a per-instruction elect loop
      const uint32_t idesc = make_idesc(2u, BN);  // tf32 x tf32 -> f32
      int stage = 0, slot = 0;
      uint32_t phase = 0, sphase = 0;
      int local = 0, gcount = 0;
      [[maybe_unused]] int gi = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++local) {
        const int cs = C::corr_bufs == 2 ? (local & 1) : 0;
        const uint32_t cph = C::corr_bufs == 2 ? ((local >> 1) & 1) : (local & 1);
        mbar_wait(cempty0 + 8 * cs, cph ^ 1);  // the epilogue has drained this correction accumulator
        CONV_TRACE(8, gi);
        const uint32_t c_tmem = tmem_base + (2 + cs) * BN;
        for (int it0 = 0; it0 < KT; it0 += group, ++gcount) {
          const int as = gcount & 1;
          mbar_wait(tempty0 + 8 * as, ((gcount >> 1) & 1) ^ 1);
          tc_fence_after();
          CONV_TRACE(9, gi);
          const uint32_t d_tmem = tmem_base + as * BN;
          const int it1 = min(it0 + group, KT);
          for (int it = it0; it < it1; ++it) {
            mbar_wait(tafull0 + 8 * slot, sphase);  // hi / lo of this chunk are in TMEM (so its TMA has landed too)
            tc_fence_after();
            CONV_TRACE(3, gi);
            const uint32_t a_hi = tmem_base + C::a_col0 + slot * A_SLOT_COLS, a_lo = a_hi + KCHUNK;
            const uint32_t b_hi = smem0 + off_bh(stage), b_lo = smem0 + off_bl(stage);
#pragma unroll
            for (int kk = 0; kk < KCHUNK / 8; ++kk) {  // K = 8 tf32 per instruction: 8 TMEM columns / 32 bytes of B
              umma_tf32_ts(c_tmem, a_lo + kk * 8, make_desc(b_hi + kk * 32), idesc, (it | kk) != 0);
              umma_tf32_ts(c_tmem, a_hi + kk * 8, make_desc(b_lo + kk * 32), idesc, 1u);
              umma_tf32_ts(d_tmem, a_hi + kk * 8, make_desc(b_hi + kk * 32), idesc, ((it - it0) | kk) != 0);
            }
            umma_commit(empty0 + 8 * stage);    // smem stage free (x was consumed by the split warps earlier)
            umma_commit(taempty0 + 8 * slot);   // TMEM A slot free
            CONV_TRACE(4, gi); ++gi;
            if (++stage == STAGES) { stage = 0; phase ^= 1; }
            if (++slot == A_SLOTS) { slot = 0; sphase ^= 1; }
          }
          umma_commit(tfull0 + 8 * as);
        }
        umma_commit(cfull0 + 8 * cs);
      }
    }
  } else if (warp < 6) {
    // ================= epilogue (warps 2..5) =================
    const int lg = warp & 3;
    float* stg = stg_all + lg * 32 * C::stg_pitch;
    const uint32_t lane_base = tmem_base + ((uint32_t)(lg * 32) << 16);
    int local = 0, gcount = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++local) {
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
      const bool lean = drain_is_lean<(BN != 96)>(p);  // warp-uniform
      const bool pre_accum = lean && p.accum != nullptr && !p.accum_init;  // running mean: read-modify-write
      if (lean && (p.residual || pre_accum)) {
        // Pull this tile's residual rows into L2 now, while the MMAs run: the drain below keeps only four 16-byte
        // loads per lane in flight, which from HBM (~1 us) capped the whole kernel at ~1.2 TB/s of residual reads.
        int t = t_tile0 + lg * 32 + lane, bb = b_t;
        if (straddle)
          while (t >= p.S_out) { t -= p.S_out; ++bb; }
        const int64_t off = p.y_offset + (int64_t)t * p.y_row_stride + nt * BN;
        if (bb < p.B && t < p.S_out && off >= 0 && off < p.y_batch_elems) {
          const int64_t g0 = (int64_t)bb * p.y_batch_stride + off;
          const int ncol = min(BN, p.Cout - nt * BN);
          for (int c = 0; c < ncol; c += 32) {  // one 128-byte line per step
            if (p.residual) asm volatile("prefetch.global.L2 [%0];" ::"l"(static_cast<const float*>(p.residual) + g0 + c));
            if (pre_accum) asm volatile("prefetch.global.L2 [%0];" ::"l"(p.accum + g0 + c));
          }
        }
      }
      float acc[BN];
#pragma unroll
      for (int i = 0; i < BN; ++i) acc[i] = 0.f;
      for (int it0 = 0; it0 < KT; it0 += group, ++gcount) {
        if (it0 + group >= KT) {
          // Last group: its MMAs and the correction terms finish together.  Take the (single-buffered) correction
          // accumulator first and hand it back at once - the next tile's MMAs wait for it, and with a short
          // reduction (k = 1 layers: 8 chunks) anything slower would stall the tensor pipe between tiles.
          const int cs = C::corr_bufs == 2 ? (local & 1) : 0;
          const uint32_t cph = C::corr_bufs == 2 ? ((local >> 1) & 1) : (local & 1);
          mbar_wait(cfull0 + 8 * cs, cph);
          tc_fence_after();
          if (threadIdx.x == 64) CONV_TRACE(10, local * KT + KT - 1);
#pragma unroll
          for (int c0 = 0; c0 < BN; c0 += 32) {
            float v[32];
            tmem_ld32(lane_base + (2 + cs) * BN + c0, v);
#pragma unroll
            for (int i = 0; i < 32; ++i) acc[c0 + i] += v[i];
          }
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(cempty0 + 8 * cs);
          if (threadIdx.x == 64) CONV_TRACE(11, local * KT + KT - 1);
        }
        const int as = gcount & 1;
        mbar_wait(tfull0 + 8 * as, (gcount >> 1) & 1);
        tc_fence_after();
#pragma unroll
        for (int c0 = 0; c0 < BN; c0 += 32) {
          float v[32];
          tmem_ld32(lane_base + as * BN + c0, v);
#pragma unroll
          for (int i = 0; i < 32; ++i) acc[c0 + i] += v[i];  // round-to-nearest fp32 adds
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(tempty0 + 8 * as);
        if (threadIdx.x == 64) CONV_TRACE(7, local * KT + min(it0 + group, KT) - 1);  // stamped at the group's last chunk
      }
      if constexpr (BN == 96) {
        if (lean) {
          drain_tile_lean_wide<BN / 32>(p, acc, stg, lg, lane, b_t, t_tile0, straddle, nt * BN, len);
          continue;
        }
      }
#pragma unroll
      for (int c0 = 0; c0 < BN; c0 += 32) {  // TMEM is free again: the stores overlap the next tile's MMAs
        float v[32];
#pragma unroll
        for (int i = 0; i < 32; ++i) v[i] = acc[c0 + i];
        if (lean)
          drain_block_lean<(BN != 96)>(p, v, stg, lg, lane, b_t, t_tile0, straddle, nt * BN + c0, len);
        else
          drain_block(p, v, stg, lg, lane, b, mt, nt * BN + c0, len);
      }
    }
  } else {
    // ================= operand split (warps 6..9): thread = row, shared memory -> registers -> TMEM =========
    const int lg = warp & 3;  // TMEM lane group this warp may write
    const int xhalf = (warp - 6) >> 2;  // which half of the row's channels (eight split warps only)
    const int r = lg * 32 + lane;
    const uint32_t lane_base = tmem_base + ((uint32_t)(lg * 32) << 16) + C::a_col0;
    const int sw = r & 7;
    int stage = 0, slot = 0;
    uint32_t phase = 0, sphase = 0;
    [[maybe_unused]] int gi = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
      // flat tiling: frame index of this row inside its own utterance, and whether any row of the tile (with its
      // halo) can reach across an utterance end - only then do taps have to be checked row by row
      int t_row = 0;
      bool edge_tile = false;
      if (p.flat) {
        const int R0 = ((tile / p.n_tiles) % p.m_tiles) * BM;
        const int t0 = R0 % p.S_in;
        t_row = t0 + r;
        if (t_row >= p.S_in) t_row %= p.S_in;
        edge_tile = t0 - p.pad < 0 || t0 + BM - 1 + (p.k - 1) * p.dilation - p.pad >= p.S_in;
      }
      int t_src = t_row - p.pad, cleft = nchunk;  // source frame of this row for the current tap
      for (int it = 0; it < KT; ++it) {
        mbar_wait(full0 + 8 * stage, phase);
        if (threadIdx.x == 192) CONV_TRACE(1, gi);
        const uint8_t* row = smem_al + off_a(stage) + r * ROW_BYTES;
        // flat tiling: a tap that leaves this row's utterance must read the zero padding, not the neighbour's frames
        const bool outside = edge_tile && (t_src < 0 || t_src >= p.S_in);
        if (--cleft == 0) { cleft = nchunk; t_src += p.dilation; }
        constexpr int NU = 32 / xf_warps<BN>();  // 16-byte units of the row this warp splits: 8 (whole row) or 4 (half)
        float hi[4 * NU], lo[4 * NU];
#pragma unroll
        for (int uu = 0; uu < NU; ++uu) {  // 16-byte unit u of the row sits at unit (u ^ (row & 7)) - SWIZZLE_128B
          const int u = xhalf * NU + uu;
          float4 x = *reinterpret_cast<const float4*>(row + ((u ^ sw) << 4));
          if (in_act == SS_ACT_LRELU) {
            x.x = fmaxf(x.x, x.x * in_slope);
            x.y = fmaxf(x.y, x.y * in_slope);
            x.z = fmaxf(x.z, x.z * in_slope);
            x.w = fmaxf(x.w, x.w * in_slope);
          }
          hi[4 * uu + 0] = tf32_rn(x.x); lo[4 * uu + 0] = LO_RN(x.x - hi[4 * uu + 0]);
          hi[4 * uu + 1] = tf32_rn(x.y); lo[4 * uu + 1] = LO_RN(x.y - hi[4 * uu + 1]);
          hi[4 * uu + 2] = tf32_rn(x.z); lo[4 * uu + 2] = LO_RN(x.z - hi[4 * uu + 2]);
          hi[4 * uu + 3] = tf32_rn(x.w); lo[4 * uu + 3] = LO_RN(x.w - hi[4 * uu + 3]);
        }
        if (edge_tile && __any_sync(0xffffffffu, outside)) {  // warp-uniform branch: skipped in interior tiles
#pragma unroll
          for (int i = 0; i < 4 * NU; ++i) {
            hi[i] = outside ? 0.f : hi[i];
            lo[i] = outside ? 0.f : lo[i];
          }
        }
        if (threadIdx.x == 192) CONV_TRACE(5, gi);
        mbar_wait(taempty0 + 8 * slot, sphase ^ 1);  // the MMAs that read this TMEM slot have completed
        tc_fence_after();
        if (threadIdx.x == 192) CONV_TRACE(6, gi);
        if constexpr (NU == 8) {
          tmem_st32(lane_base + slot * A_SLOT_COLS, hi);
          tmem_st32(lane_base + slot * A_SLOT_COLS + KCHUNK, lo);
        } else {
          tmem_st16(lane_base + slot * A_SLOT_COLS + xhalf * 16, hi);
          tmem_st16(lane_base + slot * A_SLOT_COLS + KCHUNK + xhalf * 16, lo);
        }
        asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(tafull0 + 8 * slot);
        if (threadIdx.x == 192) CONV_TRACE(2, gi);
        ++gi;
        if (++stage == STAGES) { stage = 0; phase ^= 1; }
        if (++slot == A_SLOTS) { slot = 0; sphase ^= 1; }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    __syncwarp();
    tc_fence_after();
    tmem_dealloc(tmem_base, (uint32_t)C::tmem_cols);
  }
}

template <int BN, int STAGES>
int launch(const CUtensorMap& tmA, const CUtensorMap& tmBh, const CUtensorMap& tmBl, TcArgs a, int in_act,
           float in_slope, int group, cudaStream_t st) {
  using C = Cfg<BN, STAGES>;
  a.n_tiles = (a.Cout + BN - 1) / BN;
  a.m_tiles = a.flat ? (int)(((int64_t)a.B * a.S_out + BM - 1) / BM) : (a.S_out + BM - 1) / BM;
  const int64_t tiles = (int64_t)(a.flat ? 1 : a.B) * a.m_tiles * a.n_tiles;
  SS_REQUIRE(tiles < (1ll << 31), SS_ERR_UNSUPPORTED, "ss_conv1d_tf32x3_ts: too many tiles");
  static std::atomic<unsigned long long> attr_mask{0};  // per device
  if (needs_attr_on_device(attr_mask)) {
    cudaError_t e = cudaFuncSetAttribute(conv1d_tf32_ts_kernel<BN, STAGES>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         C::smem_bytes + (RAGGED_MAX_B + 1) * 4);
    if (e != cudaSuccess) return ss::cuda_fail(e, "ss_conv1d_tf32x3_ts: smem opt-in");
    attr_done_on_device(attr_mask);
  }
  const int sms = sm_count();
  SS_REQUIRE(sms > 0, SS_ERR_CUDA, "ss_conv1d_tf32x3_ts: no CUDA device");
  // ragged schedule: `tiles` is the upper bound (every row wanted); the CTAs read the real count from the prefix table
  const int grid = (int)(tiles < sms ? tiles : sms);
  const int smem = C::smem_bytes + (a.rows ? (a.B + 1) * 4 : 0);
  conv1d_tf32_ts_kernel<BN, STAGES><<<grid, n_threads<BN>(), smem, st>>>(tmA, tmBh, tmBl, a, in_act, in_slope, group);
  SS_CHECK_LAUNCH("ss_conv1d_tf32x3_ts");
  return SS_OK;
}

}  // namespace

#ifdef SS_CONV_TRACE
extern "C" __attribute__((visibility("default"))) int ss_debug_conv_ts_trace(long long* host) {
  return (int)cudaMemcpyFromSymbol(host, g_conv_ts_trace, sizeof(long long) * 12 * 256);
}
#endif

extern "C" int ss_conv1d_tf32x3_ts(const ss_conv1d_tf32_params* p, void* stream) {
  SS_REQUIRE(p && p->x && p->w_hi && p->w_lo && (p->y || p->y_act || p->accum), SS_ERR_INVALID,
             "ss_conv1d_tf32x3_ts: null pointer");
  SS_REQUIRE(p->B > 0 && p->S_in > 0 && p->S_out > 0 && p->Cin > 0 && p->Cout > 0 && p->k > 0 && p->dilation > 0,
             SS_ERR_INVALID, "ss_conv1d_tf32x3_ts: bad shape B=%d S_in=%d S_out=%d Cin=%d Cout=%d k=%d dil=%d", p->B,
             p->S_in, p->S_out, p->Cin, p->Cout, p->k, p->dilation);
  SS_REQUIRE(p->Cin % 4 == 0 && p->Cout % 4 == 0, SS_ERR_UNSUPPORTED,
             "ss_conv1d_tf32x3_ts: Cin (%d) and Cout (%d) must be multiples of 4", p->Cin, p->Cout);
  SS_REQUIRE(p->x_row_stride % 4 == 0 && p->x_batch_stride % 4 == 0 && p->y_row_stride % 4 == 0 &&
                 p->y_batch_stride % 4 == 0 && p->y_offset % 4 == 0 && p->y_batch_elems % 4 == 0,
             SS_ERR_ALIGNMENT, "ss_conv1d_tf32x3_ts: strides/offsets must be multiples of 4 elements");
  SS_REQUIRE(ss::aligned16(p->x) && ss::aligned16(p->w_hi) && ss::aligned16(p->w_lo) && ss::aligned16(p->y) &&
                 ss::aligned16(p->y_act) && ss::aligned16(p->bias) && ss::aligned16(p->residual) && ss::aligned16(p->accum),
             SS_ERR_ALIGNMENT, "ss_conv1d_tf32x3_ts: pointers must be 16-byte aligned");
  SS_REQUIRE(p->in_act == SS_ACT_NONE || p->in_act == SS_ACT_LRELU, SS_ERR_UNSUPPORTED,
             "ss_conv1d_tf32x3_ts: in_act must be none or lrelu");
  SS_REQUIRE(p->in_act != SS_ACT_LRELU || (p->in_slope >= 0.f && p->in_slope <= 1.f), SS_ERR_UNSUPPORTED,
             "ss_conv1d_tf32x3_ts: leaky-ReLU on load is computed as max(x, slope*x) and needs 0 <= slope <= 1 (got %g)", p->in_slope);

  const int group = p->k_group > 0 ? p->k_group : 8;
  const int chunks = p->k * ((p->Cin + KCHUNK - 1) / KCHUNK);
  // short reductions (under 16 chunks: the k = 1 layers with K <= 256, the 80-channel input layers): the 96-column
  // form, whose correction accumulator is double-buffered, unless that would waste more than a quarter of the
  // last column tile.  Per output column its steady state costs the same as the 128-column form (it is bound by
  // the split, 650 cycles per chunk), and it has no stall between tiles.
  const bool short_k = chunks < 16 && p->Cout > 64 && ((p->Cout + 95) / 96) * 96 * 4 <= p->Cout * 5;
  const int bn = short_k ? 96 : p->Cout > 64 ? 128 : p->Cout > 32 ? 64 : 32;
  // Flat tiling (see TcArgs::flat): stride-1 same-length layer, batch-contiguous input, lean epilogue, and a
  // sequence length that would otherwise leave a partly empty last tile in every utterance.
  SS_REQUIRE(p->rows == nullptr || p->B <= RAGGED_MAX_B, SS_ERR_UNSUPPORTED,
             "ss_conv1d_tf32x3_ts: the ragged schedule (rows) takes at most %d utterances per call", RAGGED_MAX_B);
  const bool flat = p->rows == nullptr && p->B > 1 && p->S_in == p->S_out && p->S_out % BM != 0 &&
                    p->x_batch_stride == (int64_t)p->S_in * p->x_row_stride && (int64_t)p->B * p->S_in < (1ll << 30) &&
                    (p->accum == nullptr ? p->y != nullptr : p->Cout <= 64) && p->y_act == nullptr &&
                    p->act != SS_ACT_TANH;  // = the conditions of the lean epilogue, the only one that un-flattens rows
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
  switch (bn) {
    case 128: return launch<128, 4>(tmA, tmBh, tmBl, a, p->in_act, p->in_slope, group, st);
    case 96: return launch<96, 4>(tmA, tmBh, tmBl, a, p->in_act, p->in_slope, group, st);
    case 64: return launch<64, 6>(tmA, tmBh, tmBl, a, p->in_act, p->in_slope, group, st);
    default: return launch<32, 8>(tmA, tmBh, tmBl, a, p->in_act, p->in_slope, group, st);
  }
}
