// 3xTF32 implicit-GEMM convolution for NARROW layers (Cin, Cout <= 64: HiFi-GAN's 64- and 32-channel MRF stages),
// "panel" form.  Same mathematics, MMA order, accumulator handling and epilogue as conv1d_tf32_ts.cu - results are
// bit-identical - but a different way of feeding the tensor core.
//
// Why: in conv1d_tf32_ts.cu every (tap, 32-channel chunk) step fetches a fresh 128-row activation tile AND its weight
// tile by TMA, so a k = 11 layer pulls every activation row through L2 -> shared memory eleven times and the same
// weights once per 128-row tile: 264 KB (32 channels) / 704 KB (64 channels) per tile for 16 / 32 KB of new input.
// On the narrow layers that traffic, not the tensor pipe, sets the pace (scripts/conv_trace.py: TMA issue -> data
// landed takes 3.4-6.8 k cycles with eight stages in flight, i.e. ~30 B/clk per SM, ~2/3 of what L2 can deliver to
// 148 SMs; tensor pipe 13-28 % busy at 32 channels, profiles/r2a_ncu_hifigan_shapes.txt).
//
// Here
//   * the activations of a tile arrive ONCE, as a panel of R = 128 + (k-1)*dilation rows per 32-channel chunk (one TMA
//     box; rows outside the utterance are zero-filled by TMA as before).  Tap j of output row r is panel row
//     r + j*dilation: the split warps just read a shifted row.  (A SWIZZLE_128B panel can be read at any row offset
//     because the swizzle phase is a function of the absolute shared-memory address - scripts/desc_shift_probe.py.)
//   * the weights of ALL taps stay resident in shared memory, loaded once per CTA.  That is the condition for using this
//     kernel (32 channels: k <= 12; 64 channels: k = 3); layers whose weights do not fit stay with conv1d_tf32_ts.cu.
// Per 128-row tile that is 18-46 KB of TMA traffic instead of 264 / 192 KB.
//
//   * the MMA-issuing thread, not the tensor pipe, paced the narrow tiles (an N = 32 MMA is 16 cycles of tensor work; a
//     microbenchmark, scripts/mma_rate_probe.py, shows the pipe itself runs back-to-back N = 32 TS-form MMAs at full
//     rate, while the kernel's issue loop with its per-chunk barrier round trips needed ~50 cycles per MMA).  So per
//     K-step the products hi*W_hi and hi*W_lo are ONE instruction with N = 2*BN (W_hi and W_lo tiles are adjacent in
//     shared memory and form one [2*BN x 32] operand; the accumulator is [main | correction], 2*BN columns), followed by
//     lo*W_hi into the correction half: 8 instructions per chunk instead of 12, and one operand hand-over covers TWO
//     (tap, chunk) steps.  Main and correction accumulators are restarted together every k_group chunks and summed in
//     registers by the epilogue, so results differ from conv1d_tf32_ts.cu in the last bits (same error class).
//
// Warps: TMA, MMA, 4 epilogue, 8 split (two per TMEM lane group, half a row each).  TMEM: [main | correction] x2 |
// two A slots of 2 x (32 hi + 32 lo) columns.
#include <type_traits>

#include "tc_common.cuh"

namespace {

using namespace sstc;

constexpr int KCHUNK = 32;
constexpr int XF_WARPS = 8;
constexpr int NTHREADS = 32 * (6 + XF_WARPS);
constexpr int MAX_A_SLOTS = 3;
constexpr int A_STEPS = 2;                         // (tap, chunk) steps per operand hand-over
constexpr int A_SLOT_COLS = A_STEPS * 2 * KCHUNK;  // [hi | lo] per step
constexpr int MAX_PAN = 4;               // activation panel ring depth (upper bound)
constexpr int W_RING = 4;                // weight ring depth when the weights of all taps do not fit (64 channels, k > 3)
constexpr int N_BARS = 2 * MAX_PAN + 2 * W_RING + 1 + 2 * MAX_A_SLOTS + 4;
constexpr int MAX_SMEM = 227 * 1024;

struct PanelArgs {
  int32_t R;            // panel rows = 128 + (k-1)*dilation
  int32_t panel_bytes;  // per 32-channel chunk, multiple of 1024
  int32_t n_pan;        // panels in the ring (each holds all chunks of one tile)
};

// Debug build (-DSS_PANEL_TRACE, scripts/panel_trace.py): CTA 0 stamps clock64() per operand hand-over.
#ifdef SS_PANEL_TRACE
__device__ long long g_panel_trace[8 * 256];
#define PANEL_TRACE(ev, i)                                                        \
  do {                                                                            \
    if (blockIdx.x == 0 && (i) < 256) g_panel_trace[(ev) * 256 + (i)] = clock64(); \
  } while (0)
#else
#define PANEL_TRACE(ev, i)
#endif



// NCHUNK = Cin / 32 is a template parameter, and the weights are always resident: the MMA-issuing thread's loop is the
// critical path of the narrow tiles (one thread, ~10 % issue efficiency on branchy code - ncu source view), so it is
// kept free of run-time configuration branches.
// RESIDENT: the weights of all taps stay in shared memory (loaded once per CTA); otherwise one [W_hi ; W_lo] tile per
// (tap, chunk) step streams through a W_RING-deep ring, fetched once per 128-row tile - still half of the per-tap
// kernel's L2 -> shared-memory traffic, because the activations come once per tile.
template <int BN, int NCHUNK, bool RESIDENT>
__global__ void __launch_bounds__(NTHREADS, 1)
conv1d_tf32_panel_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmBh,
                         const __grid_constant__ CUtensorMap tmBl, const TcArgs p, const PanelArgs q, const int in_act,
                         const float in_slope, const int group) {
  constexpr int B_BYTES = BN * ROW_BYTES;  // one weight tile (hi or lo) of one (tap, chunk)
  constexpr int WT = 2 * B_BYTES;          // hi + lo
  constexpr int A_COL0 = 4 * BN;           // after [main | corr] x 2
  constexpr int A_SLOTS = (512 - A_COL0) / A_SLOT_COLS;  // 3 at BN = 32, 2 at BN = 64
  static_assert(A_SLOTS >= 2 && A_SLOTS <= MAX_A_SLOTS, "operand slots");
  static_assert(A_COL0 + A_SLOTS * A_SLOT_COLS <= 512, "TMEM budget");
  static_assert(2 * BN <= 256 && BN % 8 == 0, "[W_hi ; W_lo] must be one legal MMA operand");
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem0 = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* smem_al = smem_raw + (smem0 - smem_u32(smem_raw));
  constexpr int nchunk = NCHUNK;
  // Split ONCE per tile (32-channel layers): the split warps rewrite the landed panel in place as hi and write lo to a
  // second panel, then every tap only COPIES its shifted rows into TMEM.  The per-tap split did the same
  // arithmetic k times per row and made the split warps (issue-bound, ~5 ALU operations per element and tap) the pace
  // setter once the MMA loop was lean.  64-channel layers keep the per-tap split: two panels of [hi | lo] for two
  // chunks do not fit next to their resident weights.
  constexpr bool ONCE = NCHUNK == 1;
  constexpr int PAN_PARTS = ONCE ? 2 : 1;
  const int KT = p.k * nchunk;
  const int pan_stage = nchunk * PAN_PARTS * q.panel_bytes;           // one tile's panels (all chunks)
  const int w_off = q.n_pan * pan_stage;
  const int w_bytes = (RESIDENT ? KT : W_RING) * WT;
  const int stg_off = w_off + w_bytes;
  const int bar_off = stg_off + STG_BYTES;
  float* stg_all = reinterpret_cast<float*>(smem_al + stg_off);
  const uint32_t bars = smem0 + bar_off;
  const uint32_t pfull0 = bars, pempty0 = pfull0 + 8 * MAX_PAN, wfull0 = pempty0 + 8 * MAX_PAN, wempty0 = wfull0 + 8 * W_RING,
                 wres = wempty0 + 8 * W_RING, tafull0 = wres + 8, taempty0 = tafull0 + 8 * MAX_A_SLOTS,
                 tfull0 = taempty0 + 8 * MAX_A_SLOTS, tempty0 = tfull0 + 16, tslot = tempty0 + 16;
  volatile uint32_t* tslot_ptr = reinterpret_cast<volatile uint32_t*>(smem_al + bar_off + N_BARS * 8);
  int* const rag = reinterpret_cast<int*>(smem_al + bar_off + N_BARS * 8 + 16);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const bool ragged = p.rows != nullptr;
  if (ragged) ragged_build(p.rows, p.B, p.S_out, BM, rag);

  if (threadIdx.x == 0) {
    prefetch_tmap(&tmA);
    prefetch_tmap(&tmBh);
    prefetch_tmap(&tmBl);
    for (int s = 0; s < MAX_PAN; ++s) {
      mbar_init(pfull0 + 8 * s, 1);
      mbar_init(pempty0 + 8 * s, XF_WARPS);  // every split warp has read its last row of the panel
    }
    for (int w = 0; w < W_RING; ++w) {
      mbar_init(wfull0 + 8 * w, 1);
      mbar_init(wempty0 + 8 * w, 1);
    }
    mbar_init(wres, 1);
    for (int t = 0; t < MAX_A_SLOTS; ++t) {
      mbar_init(tafull0 + 8 * t, XF_WARPS);
      mbar_init(taempty0 + 8 * t, 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(tfull0 + 8 * a, 1);
      mbar_init(tempty0 + 8 * a, 4);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) tmem_alloc(tslot, 512u);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tslot_ptr;
  const int num_tiles = ragged ? rag[p.B] : p.B * p.m_tiles;  // one column tile (Cout <= BN)
  auto decode = [&](int tile, int& mt, int& b) {
    if (ragged) {
      ragged_decode(rag, p.B, tile, b, mt);
    } else {
      mt = tile % p.m_tiles;
      b = tile / p.m_tiles;
    }
  };

  if (warp == 0) {
    // ================= TMA producer =================
    if (elect_one()) {
      if constexpr (RESIDENT) {  // every (tap, chunk) weight tile, once per CTA
        mbar_expect_tx(wres, (uint32_t)(KT * WT));
        for (int it = 0; it < KT; ++it) {
          const int j = it / nchunk, c0 = (it - j * nchunk) * KCHUNK;
          tma_load_3d(smem0 + w_off + it * WT, &tmBh, wres, c0, 0, j);
          tma_load_3d(smem0 + w_off + it * WT + B_BYTES, &tmBl, wres, c0, 0, j);
        }
      }
      int ps = 0, ws = 0;
      uint32_t pphase = 0, wphase = 0;
      auto load_panel = [&](int tile) {
        int mt, b;
        decode(tile, mt, b);
        mbar_wait(pempty0 + 8 * ps, pphase ^ 1);
        mbar_expect_tx(pfull0 + 8 * ps, (uint32_t)(nchunk * q.R * ROW_BYTES));
#pragma unroll
        for (int c = 0; c < nchunk; ++c)
          tma_load_3d(smem0 + ps * pan_stage + c * PAN_PARTS * q.panel_bytes, &tmA, pfull0 + 8 * ps, c * KCHUNK,
                      mt * BM - p.pad, b);
        if (++ps == q.n_pan) { ps = 0; pphase ^= 1; }
      };
      if constexpr (RESIDENT) {
        for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) load_panel(tile);
      } else {
        // the panel of tile i+1 is requested BEFORE the weights of tile i: the weight ring only moves as fast as the
        // MMAs consume it, and a panel requested behind it would arrive late
        if ((int)blockIdx.x < num_tiles) load_panel(blockIdx.x);
        for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
          if (tile + (int)gridDim.x < num_tiles) load_panel(tile + gridDim.x);
          for (int it = 0; it < KT; ++it) {
            const int j = it / nchunk, c0 = (it - j * nchunk) * KCHUNK;
            mbar_wait(wempty0 + 8 * ws, wphase ^ 1);
            mbar_expect_tx(wfull0 + 8 * ws, (uint32_t)WT);
            tma_load_3d(smem0 + w_off + ws * WT, &tmBh, wfull0 + 8 * ws, c0, 0, j);
            tma_load_3d(smem0 + w_off + ws * WT + B_BYTES, &tmBl, wfull0 + 8 * ws, c0, 0, j);
            if (++ws == W_RING) { ws = 0; wphase ^= 1; }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ================= MMA issuer =================
    if (elect_one()) {
      const uint32_t idesc2 = make_idesc(2u, 2 * BN);  // hi x [W_hi ; W_lo] -> [main | corr]
      const uint32_t idesc1 = make_idesc(2u, BN);      // lo x W_hi -> corr
      const uint32_t w_lo0 = desc_lo(smem0 + w_off);   // descriptor low word of weight tile 0; tile t is + t * (WT >> 4)
      constexpr uint32_t WSTEP = WT >> 4;
      const uint32_t a_base = tmem_base + A_COL0;
      int slot = 0;
      uint32_t sphase = 0;
      [[maybe_unused]] int ws = 0;
      [[maybe_unused]] uint32_t wphase = 0;
      int gcount = 0;
      [[maybe_unused]] int hcount = 0;
      if constexpr (RESIDENT) mbar_wait(wres, 0);
      // one operand hand-over: STEPS (tap, chunk) steps, 8 MMAs each, straight-line
      auto hand_over = [&](auto steps_c, uint32_t d_tmem, uint32_t bw, uint32_t& accum) {
        constexpr int STEPS = decltype(steps_c)::value;
        PANEL_TRACE(0, hcount);
        mbar_wait(tafull0 + 8 * slot, sphase);  // hi / lo of these steps are in TMEM
        tc_fence_after();
        PANEL_TRACE(1, hcount);
        const uint32_t a0 = a_base + slot * A_SLOT_COLS;
#pragma unroll
        for (int sidx = 0; sidx < STEPS; ++sidx) {
          uint32_t bws = bw + sidx * WSTEP;  // resident: weight tile `it` of the tile
          if constexpr (!RESIDENT) {       // streamed: the ring slot this step's weights landed in
            mbar_wait(wfull0 + 8 * ws, wphase);
            tc_fence_after();
            bws = w_lo0 + ws * WSTEP;
          }
#pragma unroll
          for (int kk = 0; kk < KCHUNK / 8; ++kk) {  // K = 8 tf32 per instruction: 8 TMEM columns / 32 bytes of B
            umma_tf32_ts_w(d_tmem, a0 + sidx * 2 * KCHUNK + kk * 8, bws + kk * 2, idesc2, accum);
            accum = 1u;
            umma_tf32_ts_w(d_tmem + BN, a0 + sidx * 2 * KCHUNK + KCHUNK + kk * 8, bws + kk * 2, idesc1, 1u);
          }
          if constexpr (!RESIDENT) {
            umma_commit(wempty0 + 8 * ws);
            if (++ws == W_RING) { ws = 0; wphase ^= 1; }
          }
        }
        umma_commit(taempty0 + 8 * slot);
        PANEL_TRACE(2, hcount); ++hcount;
        if (++slot == A_SLOTS) { slot = 0; sphase ^= 1; }
      };
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        uint32_t bw = w_lo0;
        for (int it0 = 0; it0 < KT; it0 += group, ++gcount) {
          const int as = gcount & 1;
          mbar_wait(tempty0 + 8 * as, ((gcount >> 1) & 1) ^ 1);  // the epilogue has drained this accumulator pair
          tc_fence_after();
          const uint32_t d_tmem = tmem_base + as * 2 * BN;
          const int n = min(group, KT - it0);  // steps of this group; hand-overs never straddle a group (group is even)
          uint32_t accum = 0u;                 // the first MMA of a group restarts [main | corr]
          for (int h = 0; h < (n >> 1); ++h) {
            hand_over(std::integral_constant<int, 2>{}, d_tmem, bw, accum);
            bw += 2 * WSTEP;
          }
          if (n & 1) {
            hand_over(std::integral_constant<int, 1>{}, d_tmem, bw, accum);
            bw += WSTEP;
          }
          umma_commit(tfull0 + 8 * as);
        }
      }
    }
  } else if (warp < 6) {
    // ================= epilogue (warps 2..5) =================
    const int lg = warp & 3;
    float* stg = stg_all + lg * 32 * STG_PITCH;
    const uint32_t lane_base = tmem_base + ((uint32_t)(lg * 32) << 16);
    int gcount = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
      int mt, b;
      decode(tile, mt, b);
      const int t_tile0 = mt * BM;
      const int len = p.lens ? p.lens[b] : 0x7fffffff;
      const bool lean = drain_is_lean<true>(p);  // warp-uniform
      const bool pre_accum = lean && p.accum != nullptr && !p.accum_init;  // running mean: read-modify-write
      if (lean && (p.residual || pre_accum)) {  // pull this tile's residual / running-mean rows into L2 while its MMAs run
        const int t = t_tile0 + lg * 32 + lane;
        const int64_t off = p.y_offset + (int64_t)t * p.y_row_stride;
        if (t < p.S_out && off >= 0 && off < p.y_batch_elems) {
          const int64_t g0 = (int64_t)b * p.y_batch_stride + off;
          for (int c = 0; c < min(BN, p.Cout); c += 32) {
            if (p.residual) asm volatile("prefetch.global.L2 [%0];" ::"l"(static_cast<const float*>(p.residual) + g0 + c));
            if (pre_accum) asm volatile("prefetch.global.L2 [%0];" ::"l"(p.accum + g0 + c));
          }
        }
      }
      float acc[BN];
#pragma unroll
      for (int i = 0; i < BN; ++i) acc[i] = 0.f;
      for (int it0 = 0; it0 < KT; it0 += group, ++gcount) {
        const int as = gcount & 1;
        mbar_wait(tfull0 + 8 * as, (gcount >> 1) & 1);
        tc_fence_after();
#pragma unroll
        for (int c0 = 0; c0 < BN; c0 += 32) {
          float v[32], w[32];
          tmem_ld32x2(lane_base + as * 2 * BN + c0, v, lane_base + as * 2 * BN + BN + c0, w);
#pragma unroll
          for (int i = 0; i < 32; ++i) acc[c0 + i] += v[i] + w[i];  // round-to-nearest fp32 adds: main + correction
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(tempty0 + 8 * as);
      }
#pragma unroll
      for (int c0 = 0; c0 < BN; c0 += 32) {  // TMEM is free again: the stores overlap the next tile's MMAs
        float v[32];
#pragma unroll
        for (int i = 0; i < 32; ++i) v[i] = acc[c0 + i];
        if (lean)
          drain_block_lean<true>(p, v, stg, lg, lane, b, t_tile0, false, c0, len);
        else
          drain_block(p, v, stg, lg, lane, b, mt, c0, len);
      }
    }
  } else {
    // ================= operand split (warps 6..13): thread = output row, half of the 32 channels each =========
    const int lg = warp & 3;             // TMEM lane group this warp may write
    const int xhalf = (warp - 6) >> 2;   // which half of the chunk's channels
    const int r = lg * 32 + lane;
    const uint32_t lane_base = tmem_base + ((uint32_t)(lg * 32) << 16) + A_COL0;
    int ps = 0, slot = 0;
    uint32_t pphase = 0, sphase = 0;
    [[maybe_unused]] int hcount = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
      if (threadIdx.x == 192) PANEL_TRACE(6, hcount);
      mbar_wait(pfull0 + 8 * ps, pphase);
      if (threadIdx.x == 192) PANEL_TRACE(7, hcount);
      uint8_t* pan = smem_al + ps * pan_stage;
      if constexpr (ONCE) {
        // phase 1: split the whole panel (R rows, halo included) element-wise, 16-byte units dealt round-robin to the
        // 256 split threads - in place, so the swizzle never has to be decoded
        const int tid = threadIdx.x - 6 * 32;
        const int n_units = q.R * (ROW_BYTES / 16);
        for (int un = tid; un < n_units; un += XF_WARPS * 32) {
          float4* ph = reinterpret_cast<float4*>(pan + un * 16);
          float4 x = *ph;
          if (in_act == SS_ACT_LRELU) {
            x.x = fmaxf(x.x, x.x * in_slope);
            x.y = fmaxf(x.y, x.y * in_slope);
            x.z = fmaxf(x.z, x.z * in_slope);
            x.w = fmaxf(x.w, x.w * in_slope);
          }
          const float4 h = make_float4(tf32_rn(x.x), tf32_rn(x.y), tf32_rn(x.z), tf32_rn(x.w));
          *ph = h;
          // lo stays unrounded: the MMA truncates an fp32 operand to TF32 itself (see conv1d_tf32_ts.cu)
          *reinterpret_cast<float4*>(pan + q.panel_bytes + un * 16) = make_float4(x.x - h.x, x.y - h.y, x.z - h.z, x.w - h.w);
        }
        asm volatile("bar.sync 1, %0;" ::"n"(XF_WARPS * 32) : "memory");  // the split warps only
      }
      int prow = r, c = 0;  // panel row of this output row for the current tap (tap j reads row r + j*dilation)
      for (int it = 0; it < KT; it += A_STEPS) {
        const int steps = min(A_STEPS, KT - it);  // the MMA issuer cuts hand-overs the same way (group is even)
        float hi[A_STEPS][16], lo[A_STEPS][16];
#pragma unroll
        for (int sidx = 0; sidx < A_STEPS; ++sidx) {
          if (sidx < steps) {
            const uint8_t* row = pan + c * PAN_PARTS * q.panel_bytes + prow * ROW_BYTES;
            const int sw = prow & 7;  // SWIZZLE_128B: 16-byte unit u of a row sits at unit (u ^ (row & 7)); panels are 1024-aligned
            if (++c == nchunk) { c = 0; prow += p.dilation; }
#pragma unroll
            for (int uu = 0; uu < 4; ++uu) {
              const int u = xhalf * 4 + uu;
              if constexpr (ONCE) {  // phase 2: copy the shifted row of the hi and lo panels
                const float4 h = *reinterpret_cast<const float4*>(row + ((u ^ sw) << 4));
                const float4 l = *reinterpret_cast<const float4*>(row + q.panel_bytes + ((u ^ sw) << 4));
                hi[sidx][4 * uu + 0] = h.x; hi[sidx][4 * uu + 1] = h.y; hi[sidx][4 * uu + 2] = h.z; hi[sidx][4 * uu + 3] = h.w;
                lo[sidx][4 * uu + 0] = l.x; lo[sidx][4 * uu + 1] = l.y; lo[sidx][4 * uu + 2] = l.z; lo[sidx][4 * uu + 3] = l.w;
              } else {
                float4 x = *reinterpret_cast<const float4*>(row + ((u ^ sw) << 4));
                if (in_act == SS_ACT_LRELU) {
                  x.x = fmaxf(x.x, x.x * in_slope);
                  x.y = fmaxf(x.y, x.y * in_slope);
                  x.z = fmaxf(x.z, x.z * in_slope);
                  x.w = fmaxf(x.w, x.w * in_slope);
                }
                // lo is handed over unrounded: the MMA truncates an fp32 operand to TF32 itself (see conv1d_tf32_ts.cu)
                hi[sidx][4 * uu + 0] = tf32_rn(x.x); lo[sidx][4 * uu + 0] = x.x - hi[sidx][4 * uu + 0];
                hi[sidx][4 * uu + 1] = tf32_rn(x.y); lo[sidx][4 * uu + 1] = x.y - hi[sidx][4 * uu + 1];
                hi[sidx][4 * uu + 2] = tf32_rn(x.z); lo[sidx][4 * uu + 2] = x.z - hi[sidx][4 * uu + 2];
                hi[sidx][4 * uu + 3] = tf32_rn(x.w); lo[sidx][4 * uu + 3] = x.w - hi[sidx][4 * uu + 3];
              }
            }
          }
        }
        if (threadIdx.x == 192) PANEL_TRACE(3, hcount);
        if (it + steps >= KT) {  // this warp has read its last row of the panel: hand it back to the TMA producer
          __syncwarp();
          if (lane == 0) mbar_arrive(pempty0 + 8 * ps);
        }
        mbar_wait(taempty0 + 8 * slot, sphase ^ 1);  // the MMAs that read this TMEM slot have completed
        tc_fence_after();
        if (threadIdx.x == 192) PANEL_TRACE(4, hcount);
#pragma unroll
        for (int sidx = 0; sidx < A_STEPS; ++sidx) {
          if (sidx < steps) {
            tmem_st16(lane_base + slot * A_SLOT_COLS + sidx * 2 * KCHUNK + xhalf * 16, hi[sidx]);
            tmem_st16(lane_base + slot * A_SLOT_COLS + sidx * 2 * KCHUNK + KCHUNK + xhalf * 16, lo[sidx]);
          }
        }
        asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(tafull0 + 8 * slot);
        if (threadIdx.x == 192) PANEL_TRACE(5, hcount);
        ++hcount;
        if (++slot == A_SLOTS) { slot = 0; sphase ^= 1; }
      }
      if (++ps == q.n_pan) { ps = 0; pphase ^= 1; }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    __syncwarp();
    tc_fence_after();
    tmem_dealloc(tmem_base, 512u);
  }
}

// Shared-memory plan for one layer: weights resident (all taps) or streamed through a ring, plus 2 or 3 activation
// panels.  Returns 0 when nothing fits, 1 for resident weights, 2 for streamed.
inline int plan_smem(int bn, int k, int dilation, int nchunk, int B_ragged, PanelArgs& q, int& smem_bytes) {
  const int wt = 2 * bn * ROW_BYTES;
  q.R = BM + (k - 1) * dilation;
  q.panel_bytes = (q.R * ROW_BYTES + 1023) / 1024 * 1024;
  const int fixed = STG_BYTES + N_BARS * 8 + 16 + (B_ragged ? (B_ragged + 1) * 4 : 0) + 1024;
  const int parts = nchunk == 1 ? 2 : 1;  // 32-channel layers keep [hi | lo] panels (split once per tile)
  for (int resident = 1; resident >= 0; --resident)
    for (int n_pan = 3; n_pan >= 2; --n_pan) {
      const int total = n_pan * nchunk * parts * q.panel_bytes + (resident ? k * nchunk : W_RING) * wt + fixed;
      if (total <= MAX_SMEM) {
        q.n_pan = n_pan;
        smem_bytes = total;
        return resident ? 1 : 2;
      }
    }
  return 0;
}

template <int BN, int NCHUNK, bool RESIDENT>
int launch_k(const CUtensorMap& tmA, const CUtensorMap& tmBh, const CUtensorMap& tmBl, const TcArgs& a, const PanelArgs& q,
             int smem, int grid, int in_act, float in_slope, int group, cudaStream_t st) {
  static std::atomic<unsigned long long> attr_mask{0};  // per device
  if (needs_attr_on_device(attr_mask)) {
    cudaError_t e = cudaFuncSetAttribute(conv1d_tf32_panel_kernel<BN, NCHUNK, RESIDENT>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, MAX_SMEM);
    if (e != cudaSuccess) return ss::cuda_fail(e, "ss_conv1d_tf32x3_panel: smem opt-in");
    attr_done_on_device(attr_mask);
  }
  conv1d_tf32_panel_kernel<BN, NCHUNK, RESIDENT><<<grid, NTHREADS, smem, st>>>(tmA, tmBh, tmBl, a, q, in_act, in_slope, group);
  SS_CHECK_LAUNCH("ss_conv1d_tf32x3_panel");
  return SS_OK;
}

template <int BN, int NCHUNK>
int launch(const ss_conv1d_tf32_params* p, const TcArgs& a0, int group, cudaStream_t st) {
  TcArgs a = a0;
  PanelArgs q;
  int smem = 0;
  const int mode = plan_smem(BN, p->k, p->dilation, NCHUNK, p->rows ? p->B : 0, q, smem);
  SS_REQUIRE(mode != 0, SS_ERR_UNSUPPORTED, "ss_conv1d_tf32x3_panel: k=%d dilation=%d does not fit shared memory", p->k,
             p->dilation);
  CUtensorMap tmA, tmBh, tmBl;
  int rc = make_map_3d(&tmA, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, p->x, (uint64_t)p->Cin, (uint64_t)p->S_in, (uint64_t)p->B,
                       (uint64_t)p->x_row_stride * 4, (uint64_t)p->x_batch_stride * 4, KCHUNK, (uint32_t)q.R);
  if (rc != SS_OK) return rc;
  rc = make_map_3d(&tmBh, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, p->w_hi, (uint64_t)p->Cin, (uint64_t)p->Cout, (uint64_t)p->k,
                   (uint64_t)p->Cin * 4, (uint64_t)p->Cout * p->Cin * 4, KCHUNK, (uint32_t)BN);
  if (rc != SS_OK) return rc;
  rc = make_map_3d(&tmBl, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, p->w_lo, (uint64_t)p->Cin, (uint64_t)p->Cout, (uint64_t)p->k,
                   (uint64_t)p->Cin * 4, (uint64_t)p->Cout * p->Cin * 4, KCHUNK, (uint32_t)BN);
  if (rc != SS_OK) return rc;
  a.n_tiles = 1;
  a.m_tiles = (a.S_out + BM - 1) / BM;
  const int64_t tiles = (int64_t)a.B * a.m_tiles;
  SS_REQUIRE(tiles < (1ll << 31), SS_ERR_UNSUPPORTED, "ss_conv1d_tf32x3_panel: too many tiles");
  const int sms = sm_count();
  SS_REQUIRE(sms > 0, SS_ERR_CUDA, "ss_conv1d_tf32x3_panel: no CUDA device");
  const int grid = (int)(tiles < sms ? tiles : sms);
  return mode == 1 ? launch_k<BN, NCHUNK, true>(tmA, tmBh, tmBl, a, q, smem, grid, p->in_act, p->in_slope, group, st)
                   : launch_k<BN, NCHUNK, false>(tmA, tmBh, tmBl, a, q, smem, grid, p->in_act, p->in_slope, group, st);
}

}  // namespace

#ifdef SS_PANEL_TRACE
extern "C" __attribute__((visibility("default"))) int ss_debug_panel_trace(long long* host) {
  return (int)cudaMemcpyFromSymbol(host, g_panel_trace, sizeof(long long) * 8 * 256);
}
#endif

// Non-zero if ss_conv1d_tf32x3_panel takes this layer: stride-1 same-length convolution with at least two taps, 32 or 64
// input channels, at most 64 output channels, a halo that fits one TMA box.  1: the weights of all taps stay resident in
// shared memory; 2: they stream through a ring once per 128-row tile.
extern "C" int ss_conv1d_tf32x3_panel_supported(const ss_conv1d_tf32_params* p) {
  if (!p) return 0;
  if (!(p->S_in == p->S_out && p->k >= 2 && (p->Cin == 32 || p->Cin == 64) && p->Cout > 0 && p->Cout <= 64 &&
        p->Cout % 4 == 0 && p->dilation > 0 && BM + (p->k - 1) * p->dilation <= 256 && p->pad >= 0 &&
        p->pad <= (p->k - 1) * p->dilation))
    return 0;
  PanelArgs q;
  int smem = 0;
  return plan_smem(p->Cout > 32 ? 64 : 32, p->k, p->dilation, p->Cin / KCHUNK, p->rows ? p->B : 0, q, smem);
}

extern "C" int ss_conv1d_tf32x3_panel(const ss_conv1d_tf32_params* p, void* stream) {
  SS_REQUIRE(p && p->x && p->w_hi && p->w_lo && (p->y || p->y_act || p->accum), SS_ERR_INVALID,
             "ss_conv1d_tf32x3_panel: null pointer");
  SS_REQUIRE(p->B > 0 && p->S_in > 0 && p->S_out > 0 && p->Cin > 0 && p->Cout > 0 && p->k > 0 && p->dilation > 0,
             SS_ERR_INVALID, "ss_conv1d_tf32x3_panel: bad shape B=%d S_in=%d S_out=%d Cin=%d Cout=%d k=%d dil=%d", p->B,
             p->S_in, p->S_out, p->Cin, p->Cout, p->k, p->dilation);
  SS_REQUIRE(ss_conv1d_tf32x3_panel_supported(p), SS_ERR_UNSUPPORTED,
             "ss_conv1d_tf32x3_panel: needs S_in == S_out, k >= 2, Cin in {32, 64}, Cout <= 64, 128 + (k-1)*dilation <= 256 "
             "(got S_in=%d S_out=%d Cin=%d Cout=%d k=%d dil=%d pad=%d)", p->S_in, p->S_out, p->Cin, p->Cout, p->k,
             p->dilation, p->pad);
  SS_REQUIRE(p->x_row_stride % 4 == 0 && p->x_batch_stride % 4 == 0 && p->y_row_stride % 4 == 0 &&
                 p->y_batch_stride % 4 == 0 && p->y_offset % 4 == 0 && p->y_batch_elems % 4 == 0,
             SS_ERR_ALIGNMENT, "ss_conv1d_tf32x3_panel: strides/offsets must be multiples of 4 elements");
  SS_REQUIRE(ss::aligned16(p->x) && ss::aligned16(p->w_hi) && ss::aligned16(p->w_lo) && ss::aligned16(p->y) &&
                 ss::aligned16(p->y_act) && ss::aligned16(p->bias) && ss::aligned16(p->residual) && ss::aligned16(p->accum),
             SS_ERR_ALIGNMENT, "ss_conv1d_tf32x3_panel: pointers must be 16-byte aligned");
  SS_REQUIRE(p->in_act == SS_ACT_NONE || p->in_act == SS_ACT_LRELU, SS_ERR_UNSUPPORTED,
             "ss_conv1d_tf32x3_panel: in_act must be none or lrelu");
  SS_REQUIRE(p->in_act != SS_ACT_LRELU || (p->in_slope >= 0.f && p->in_slope <= 1.f), SS_ERR_UNSUPPORTED,
             "ss_conv1d_tf32x3_panel: leaky-ReLU on load is computed as max(x, slope*x) and needs 0 <= slope <= 1 (got %g)", p->in_slope);
  SS_REQUIRE(p->rows == nullptr || p->B <= RAGGED_MAX_B, SS_ERR_UNSUPPORTED,
             "ss_conv1d_tf32x3_panel: the ragged schedule (rows) takes at most %d utterances per call", RAGGED_MAX_B);
  TcArgs a;
  a.bias = p->bias; a.residual = p->residual; a.lens = p->lens; a.y = p->y;
  a.y_act = static_cast<__nv_bfloat16*>(p->y_act); a.accum = p->accum;
  a.y_batch_stride = p->y_batch_stride; a.y_row_stride = p->y_row_stride; a.y_offset = p->y_offset;
  a.y_batch_elems = p->y_batch_elems;
  a.B = p->B; a.S_out = p->S_out; a.Cin = p->Cin; a.Cout = p->Cout; a.k = p->k; a.dilation = p->dilation; a.pad = p->pad;
  a.m_tiles = a.n_tiles = 0;
  a.act = p->act; a.act2 = p->act2; a.y_dtype = SS_F32; a.residual_dtype = SS_F32;
  a.accum_init = p->accum_init; a.slope = p->slope; a.slope2 = p->slope2; a.accum_scale = p->accum_scale;
  a.flat = 0; a.S_in = p->S_in; a.rows = p->rows;
  const int group = ((p->k_group > 0 ? p->k_group : 8) + 1) & ~1;  // even: operand hand-overs cover two steps
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (p->Cin == 32) return p->Cout > 32 ? launch<64, 1>(p, a, group, st) : launch<32, 1>(p, a, group, st);
  return p->Cout > 32 ? launch<64, 2>(p, a, group, st) : launch<32, 2>(p, a, group, st);
}
