// Split-bf16 ("bf16x3") implicit-GEMM convolution on tcgen05: fp32-class accuracy from THREE bf16 tensor-core
// products per multiply instead of three TF32 ones - twice the tensor-pipe rate and half the operand bytes of the
// 3xTF32 kernels.  Used for HiFi-GAN's MRF ResBlock convolutions in the fp32 mode (hifigan.py), where no integer
// decision hangs on the last bits: the reference's fp32 wave is matched to ~1e-5 (bar: 1e-3).
//
// Numerics.  Every fp32 value v is carried as a PAIR of bf16 numbers, hi = bf16(v), lo = bf16(v - hi): 16 significant
// bits, |v - hi - lo| <= 2^-18 |v|.  x*w ~= x_hi*w_hi + x_hi*w_lo + x_lo*w_hi (the dropped terms are <= 2^-16 relative),
// accumulated in fp32 in TMEM: the leading product in a "main" accumulator, both small ones in a "correction"
// accumulator, added in registers by the epilogue.
//
// Data layout.  Activations live in HBM as pair rows  [B, S, 2*C] bf16 = [hi(C) | lo(C)]  - the same bytes per row
// as fp32 - written by the producing kernel's epilogue (TcArgs::pair_c), already leaky-ReLU'd, so the operands go from
// TMA straight into the MMAs: no operand-split warps, no TMEM A operand, no activation between TMA and MMA.
// Weights are packed once (ops.ConvW.bf16x3) into the tiles one MMA group reads: [2*BN rows = W_hi rows ; W_lo rows]
// x 64 columns (128-byte rows, SWIZZLE_128B), so that x_hi * [W_hi ; W_lo] is ONE instruction with N = 2*BN that feeds
// main and correction accumulator at once (A read once from shared memory instead of twice), followed by
// x_lo * W_hi (N = BN) into the correction columns.
//   wide  (Cin % 64 == 0): tile (nt, chunk c, tap j) = W[j, nt*BN + r, 64c .. 64c+64), index (nt*nchunk + c)*k + j
//   narrow (Cin == 32):    one row holds TWO taps, [tap 2q (32 ch) | tap 2q+1 (32 ch)], index nt*ceil(k/2) + q; and an
//                          activation row [hi(32) | lo(32)] is exactly one 128-byte swizzle row.
//
// Feeding.  As in conv1d_tf32_panel.cu the activations of a tile arrive ONCE per 64-channel chunk, as a panel of
// R = 128 + (k-1)*dilation rows (hi panel + lo panel); tap j is the same panel read R rows further down - a
// SWIZZLE_128B operand may start at any row (scripts/desc_shift_probe.py) - so a k = 11 layer moves each activation
// row L2 -> shared memory once, not eleven times.  Weights stay resident in shared memory when all taps fit
// (narrow layers, 64 channels with k = 3); otherwise they stream through a ring, one tile per (chunk, tap).
//
// Roles (persistent CTA): warp 0 TMA, warp 1 MMA (one elected thread each), then NBUF epilogue sets of four warps (one
// warp per TMEM lane group); set e drains accumulator buffer e, i.e. every NBUF-th tile, so each set has NBUF tile
// periods for its instruction stream and its global-memory round trips.  NBUF = 4 for BN <= 64 (the narrow layers'
// tiles are short - a 32-channel k = 3 tile is 12 MMAs - and the epilogue, ~500 warp instructions per 32 x 32 block,
// is what has to keep up: ncu showed two warps per scheduler issue-latency bound), 2 for BN = 128.
// TMEM: NBUF x [main | correction] = NBUF * 2*BN columns.
#include "x3_common.cuh"

namespace {

using namespace sstc;

constexpr int MAX_BUF = 4;  // accumulator buffers = epilogue sets (upper bound)
constexpr int MAX_A = 4;    // activation panel ring depth (upper bound)
constexpr int MAX_W = 6;    // weight ring depth (upper bound)
constexpr int N_BARS = 2 * MAX_A + 2 * MAX_W + 1 + 2 * MAX_BUF;
constexpr int MAX_SMEM = 227 * 1024;
constexpr int nbuf_of(int bn) { return bn <= 64 ? 4 : 2; }
constexpr int nthreads_of(int bn) { return 32 * (2 + 4 * nbuf_of(bn)); }
constexpr int fixed_smem_of(int bn) { return 1024 /* alignment */ + 4 * nbuf_of(bn) * STG_WARP_BYTES + N_BARS * 8 + 16; }

struct X3Args {
  int32_t R;            // panel rows = 128 + (k-1)*dilation
  int32_t panel_bytes;  // R * 128 rounded up to 1024
  int32_t n_a;          // activation stages in the ring (one stage = one 64-channel chunk: hi panel + lo panel)
  int32_t n_w;          // weight tiles in shared memory: all of them (resident) or the ring depth
};

// SPLIT = false is the plain bf16 convolution of the bf16 mode (ss_conv1d_bf16_panel): the same feeding and roles with
// one bf16 number per value - activation rows [B, S, C], weight tiles of BN rows, one accumulator, a third of the MMAs.
template <int BN, bool NARROW, bool RESIDENT, bool SPLIT>
__global__ void __launch_bounds__(nthreads_of(BN), 1)
conv1d_bf16x3_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmW, const TcArgs p,
                     const X3Args q) {
  pdl_launch_dependents();
  constexpr int NP = SPLIT ? 2 : 1;        // bf16 numbers per value
  constexpr int WT = NP * BN * ROW_BYTES;  // one weight tile: [W_hi rows ; W_lo rows] x 128 bytes
  constexpr int ACC = NP * BN;             // columns of one accumulator buffer: [main | correction]
  constexpr int NBUF = nbuf_of(BN);
  constexpr int EPI_WARPS = 4 * NBUF;
  constexpr int TMEM_COLS = NBUF * ACC;
  static_assert(TMEM_COLS <= 512 && TMEM_COLS >= 32 && (TMEM_COLS & (TMEM_COLS - 1)) == 0 && 2 * BN <= 256,
                "TMEM / MMA N budget");
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem0 = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* smem_al = smem_raw + (smem0 - smem_u32(smem_raw));
  const int nchunk = NARROW ? 1 : p.Cin / 64;
  const int a_stage = (NARROW ? 1 : NP) * q.panel_bytes;
  const uint32_t a_tx = (uint32_t)((NARROW ? 1 : NP) * q.R * ROW_BYTES);
  const int w_off = q.n_a * a_stage;
  const int stg_off = w_off + q.n_w * WT;
  const int bar_off = stg_off + EPI_WARPS * STG_WARP_BYTES;
  float* stg_all = reinterpret_cast<float*>(smem_al + stg_off);
  const uint32_t bars = smem0 + bar_off;
  const uint32_t afull0 = bars, aempty0 = afull0 + 8 * MAX_A, wfull0 = aempty0 + 8 * MAX_A, wempty0 = wfull0 + 8 * MAX_W,
                 wres = wempty0 + 8 * MAX_W, tfull0 = wres + 8, tempty0 = tfull0 + 8 * MAX_BUF,
                 tslot = tempty0 + 8 * MAX_BUF;
  volatile uint32_t* tslot_ptr = reinterpret_cast<volatile uint32_t*>(smem_al + bar_off + N_BARS * 8);
  int* const rag = reinterpret_cast<int*>(smem_al + bar_off + N_BARS * 8 + 16);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const bool ragged = p.rows != nullptr;
  if (ragged) ragged_build(p.rows, p.B, p.S_out, BM, rag);

  if (threadIdx.x == 0) {
    prefetch_tmap(&tmA);
    prefetch_tmap(&tmW);
    for (int s = 0; s < MAX_A; ++s) {
      mbar_init(afull0 + 8 * s, 1);
      mbar_init(aempty0 + 8 * s, 1);
    }
    for (int s = 0; s < MAX_W; ++s) {
      mbar_init(wfull0 + 8 * s, 1);
      mbar_init(wempty0 + 8 * s, 1);
    }
    mbar_init(wres, 1);
    for (int a = 0; a < NBUF; ++a) {
      mbar_init(tfull0 + 8 * a, 1);
      mbar_init(tempty0 + 8 * a, 4);  // the four warps of epilogue set a
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) tmem_alloc(tslot, (uint32_t)TMEM_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  pdl_wait();  // everything below may touch memory the previous launch wrote
  const uint32_t tmem_base = *tslot_ptr;
  const int num_tiles = ragged ? rag[p.B] * p.n_tiles : p.B * p.m_tiles * p.n_tiles;
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
  const int wt_per_nt = NARROW ? (p.k + 1) / 2 : nchunk * p.k;  // weight tiles of one column tile

  if (warp == 0) {
    // ================= TMA producer =================
    if (elect_one()) {
      if (RESIDENT) {  // n_tiles == 1: every CTA keeps the whole layer
        mbar_expect_tx(wres, (uint32_t)(wt_per_nt * WT));
        for (int i = 0; i < wt_per_nt; ++i) tma_load_3d(smem0 + w_off + i * WT, &tmW, wres, 0, 0, i);
      }
      int as = 0, ws = 0;
      uint32_t aph = 0, wph = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        int nt, mt, b;
        decode(tile, nt, mt, b);
        const int t0 = mt * BM - p.pad;
        for (int c = 0; c < nchunk; ++c) {
          mbar_wait(aempty0 + 8 * as, aph ^ 1);
          mbar_expect_tx(afull0 + 8 * as, a_tx);
          const uint32_t dst = smem0 + as * a_stage;
          if (NARROW) {
            // [hi(32) | lo(32)] is one 128-byte row; plain bf16: [x(32) | zero fill past the 32 channels]
            tma_load_3d(dst, &tmA, afull0 + 8 * as, 0, t0, b);
          } else {
            tma_load_3d(dst, &tmA, afull0 + 8 * as, 64 * c, t0, b);                                     // hi
            if (SPLIT) tma_load_3d(dst + q.panel_bytes, &tmA, afull0 + 8 * as, p.Cin + 64 * c, t0, b);  // lo
          }
          if (++as == q.n_a) { as = 0; aph ^= 1; }
          if (!RESIDENT) {
            for (int j = 0; j < p.k; ++j) {
              mbar_wait(wempty0 + 8 * ws, wph ^ 1);
              mbar_expect_tx(wfull0 + 8 * ws, (uint32_t)WT);
              tma_load_3d(smem0 + w_off + ws * WT, &tmW, wfull0 + 8 * ws, 0, 0, (nt * nchunk + c) * p.k + j);
              if (++ws == q.n_w) { ws = 0; wph ^= 1; }
            }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ================= MMA issuer =================
    if (elect_one()) {
      const uint32_t idesc_m = make_idesc(1u, ACC);     // x_hi * [W_hi ; W_lo] -> [main | correction]
      const uint32_t idesc_h = make_idesc(1u, BN);      // x_lo * W_hi -> correction
      const uint32_t w_lo0 = desc_lo(smem0 + w_off);
      const uint32_t tap_step = (uint32_t)(p.dilation * ROW_BYTES) >> 4;  // descriptor units (16 bytes) per tap
      if (RESIDENT) {
        mbar_wait(wres, 0);
        tc_fence_after();
      }
      int as = 0, ws = 0, local = 0;
      uint32_t aph = 0, wph = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++local) {
        const int ab = local % NBUF;
        mbar_wait(tempty0 + 8 * ab, ((local / NBUF) & 1) ^ 1);  // epilogue set ab has read this buffer
        tc_fence_after();
        const uint32_t d_main = tmem_base + ab * ACC, d_corr = d_main + BN;
        for (int c = 0; c < nchunk; ++c) {
          mbar_wait(afull0 + 8 * as, aph);
          tc_fence_after();
          const uint32_t a0 = desc_lo(smem0 + as * a_stage);
          if constexpr (NARROW) {
            // a row is [hi k-slices 0,1 | lo k-slices 2,3]; a weight row is [tap 2q k-slices 0,1 | tap 2q+1 k-slices 2,3]
            uint32_t a = a0;
            for (int j = 0; j < p.k; ++j, a += tap_step) {
              const uint32_t b = w_lo0 + (uint32_t)(j >> 1) * (WT >> 4) + (uint32_t)(j & 1) * 4;
              umma_bf16_w(d_main, a, b, idesc_m, j != 0);
              umma_bf16_w(d_main, a + 2, b + 2, idesc_m, 1);
              if (SPLIT) {
                umma_bf16_w(d_corr, a + 4, b, idesc_h, 1);
                umma_bf16_w(d_corr, a + 6, b + 2, idesc_h, 1);
              }
            }
          } else {
            uint32_t ah = a0;
            const uint32_t lo_off = (uint32_t)q.panel_bytes >> 4;
            for (int j = 0; j < p.k; ++j, ah += tap_step) {
              uint32_t b;
              if (RESIDENT) {
                b = w_lo0 + (uint32_t)(c * p.k + j) * (WT >> 4);
              } else {
                mbar_wait(wfull0 + 8 * ws, wph);
                tc_fence_after();
                b = w_lo0 + (uint32_t)ws * (WT >> 4);
              }
              const uint32_t al = ah + lo_off;
              umma_bf16_w(d_main, ah, b, idesc_m, (c | j) != 0);
              umma_bf16_w(d_main, ah + 2, b + 2, idesc_m, 1);
              umma_bf16_w(d_main, ah + 4, b + 4, idesc_m, 1);
              umma_bf16_w(d_main, ah + 6, b + 6, idesc_m, 1);
              if (SPLIT) {
                umma_bf16_w(d_corr, al, b, idesc_h, 1);
                umma_bf16_w(d_corr, al + 2, b + 2, idesc_h, 1);
                umma_bf16_w(d_corr, al + 4, b + 4, idesc_h, 1);
                umma_bf16_w(d_corr, al + 6, b + 6, idesc_h, 1);
              }
              if (!RESIDENT) {
                umma_commit(wempty0 + 8 * ws);
                if (++ws == q.n_w) { ws = 0; wph ^= 1; }
              }
            }
          }
          umma_commit(aempty0 + 8 * as);
          if (++as == q.n_a) { as = 0; aph ^= 1; }
        }
        umma_commit(tfull0 + 8 * ab);
      }
    }
  } else {
    // ================= epilogue: set e = tiles with local % NBUF == e =================
    const int e = (warp - 2) >> 2, lg = warp & 3;
    const bool dense = drain_is_dense(p);  // warp-uniform
    float* stg = stg_all + (warp - 2) * 32 * STG_PITCH;
    const uint32_t acc = tmem_base + ((uint32_t)(lg * 32) << 16) + e * ACC;
    int local = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++local) {
      if (local % NBUF != e) continue;
      int nt, mt, b;
      decode(tile, nt, mt, b);
      const int len = p.lens ? p.lens[b] : 0x7fffffff;
      if (dense) prefetch_epilogue_rows(p, lg, lane, b, mt * BM, p.S_out - mt * BM, nt * BN, min(BN, p.Cout - nt * BN));
      mbar_wait(tfull0 + 8 * e, (local / NBUF) & 1);
      tc_fence_after();
#pragma unroll 1
      for (int c0 = 0; c0 < BN; c0 += 32) {
        float v[32];
        if constexpr (SPLIT) {
          float cr[32];
          tmem_ld32x2(acc + c0, v, acc + BN + c0, cr);
#pragma unroll
          for (int i = 0; i < 32; ++i) v[i] += cr[i];
        } else {
          tmem_ld32(acc + c0, v);
        }
        if (c0 + 32 >= BN) {  // the accumulator is in registers: hand the buffer back before the global round trips
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(tempty0 + 8 * e);
        }
        if (dense)
          drain_block_dense(p, v, stg, lg, lane, b, mt * BM, p.S_out - mt * BM, nt * BN + c0);
        else
          drain_block(p, v, stg, lg, lane, b, mt, nt * BN + c0, len);
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    __syncwarp();
    tc_fence_after();
    tmem_dealloc(tmem_base, (uint32_t)TMEM_COLS);
  }
}

// ---------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------
// Tile width for a layer, 0 = not handled by this kernel.
int plan_bn(int Cin, int Cout, int k, int dilation) {
  if (k < 1 || dilation < 1 || 128 + (k - 1) * dilation > 256) return 0;
  if (Cin == 32) return (Cout == 32 && (k + 1) / 2 * (2 * 32 * ROW_BYTES) <= 64 * 1024) ? 32 : 0;
  if (Cin % 64 != 0 || Cin > 1024) return 0;
  if (Cout % 128 == 0) return 128;
  if (Cout == 64) return 64;
  return 0;
}

// Shared-memory plan: returns 1 (weights resident), 0 (weight ring) or -1 (does not fit).
int plan_smem(int bn, bool narrow, bool split, int Cin, int Cout, int k, int dilation, int ragged_b, X3Args* q) {
  const int FIXED_SMEM = bn <= 64 ? fixed_smem_of(32) : fixed_smem_of(128);
  q->R = 128 + (k - 1) * dilation;
  q->panel_bytes = (q->R * ROW_BYTES + 1023) & ~1023;
  const int np = split ? 2 : 1;
  const int a_stage = (narrow ? 1 : np) * q->panel_bytes;
  const int WT = np * bn * ROW_BYTES;
  const int avail = MAX_SMEM - FIXED_SMEM - (ragged_b > 0 ? (ragged_b + 1) * 4 : 0);
  const int n_tiles = (Cout + bn - 1) / bn;
  const int n_wt = narrow ? (k + 1) / 2 : (Cin / 64) * k;
  if (n_tiles == 1 && n_wt * WT + 2 * a_stage <= avail) {
    q->n_w = n_wt;
    q->n_a = (avail - n_wt * WT) / a_stage;
    if (q->n_a > MAX_A) q->n_a = MAX_A;
    return 1;
  }
  if (narrow) return -1;
  q->n_a = 2;
  q->n_w = (avail - 2 * a_stage) / WT;
  if (q->n_w > MAX_W) q->n_w = MAX_W;
  if (q->n_w < 2) return -1;
  if (q->n_w == MAX_W && avail - MAX_W * WT >= 3 * a_stage) q->n_a = 3;
  return 0;
}

template <int BN, bool NARROW, bool RESIDENT, bool SPLIT>
int launch_k(const CUtensorMap& tmA, const CUtensorMap& tmW, const TcArgs& a, const X3Args& q, int64_t tiles,
             cudaStream_t st) {
  static std::atomic<unsigned long long> attr_mask{0};  // per device
  if (needs_attr_on_device(attr_mask)) {
    cudaError_t e = cudaFuncSetAttribute(conv1d_bf16x3_kernel<BN, NARROW, RESIDENT, SPLIT>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, MAX_SMEM);
    if (e != cudaSuccess) return ss::cuda_fail(e, "ss_conv1d_bf16x3: smem opt-in");
    attr_done_on_device(attr_mask);
  }
  const int sms = sm_count();
  SS_REQUIRE(sms > 0, SS_ERR_CUDA, "ss_conv1d_bf16x3: no CUDA device");
  const int grid = (int)(tiles < sms ? tiles : sms);
  constexpr int NP = SPLIT ? 2 : 1;
  const int a_stage = (NARROW ? 1 : NP) * q.panel_bytes;
  const int smem = fixed_smem_of(BN) + q.n_a * a_stage + q.n_w * NP * BN * ROW_BYTES + (a.rows ? (a.B + 1) * 4 : 0);
  SS_REQUIRE(smem <= MAX_SMEM, SS_ERR_UNSUPPORTED, "ss_conv1d_bf16x3: shared-memory plan of %d bytes", smem);
  cudaError_t le = launch_pdl(conv1d_bf16x3_kernel<BN, NARROW, RESIDENT, SPLIT>, grid, nthreads_of(BN), smem, st, tmA, tmW, a, q);
  if (le != cudaSuccess) return ss::cuda_fail(le, "ss_conv1d_bf16x3: launch");
  SS_CHECK_LAUNCH("ss_conv1d_bf16x3");
  return SS_OK;
}

template <bool SPLIT>
int run(const ss_conv1d_tc_params* p, void* stream, const char* who) {
  constexpr int NP = SPLIT ? 2 : 1;
  SS_REQUIRE(p && p->x && p->w && (p->y || p->y_act || p->accum), SS_ERR_INVALID, "%s: null pointer", who);
  SS_REQUIRE(p->B > 0 && p->S_in > 0 && p->S_out > 0 && p->Cin > 0 && p->Cout > 0 && p->k > 0 && p->dilation > 0,
             SS_ERR_INVALID, "%s: bad shape B=%d S_in=%d S_out=%d Cin=%d Cout=%d k=%d dil=%d", who, p->B, p->S_in,
             p->S_out, p->Cin, p->Cout, p->k, p->dilation);
  const int bn = plan_bn(p->Cin, p->Cout, p->k, p->dilation);
  SS_REQUIRE(bn != 0, SS_ERR_UNSUPPORTED, "%s: layer Cin=%d Cout=%d k=%d dil=%d is not handled", who, p->Cin, p->Cout,
             p->k, p->dilation);
  SS_REQUIRE(p->x_row_stride >= NP * p->Cin && p->x_row_stride % 8 == 0 && p->x_batch_stride % 8 == 0 &&
                 p->y_row_stride % 8 == 0 && p->y_batch_stride % 8 == 0 && p->y_offset % 8 == 0 &&
                 p->y_batch_elems % 8 == 0,
             SS_ERR_ALIGNMENT, "%s: strides/offsets must be multiples of 8 elements (rows of %d*Cin)", who, NP);
  SS_REQUIRE(ss::aligned16(p->x) && ss::aligned16(p->w) && ss::aligned16(p->y) && ss::aligned16(p->y_act) &&
                 ss::aligned16(p->bias) && ss::aligned16(p->residual) && ss::aligned16(p->accum),
             SS_ERR_ALIGNMENT, "%s: pointers must be 16-byte aligned", who);
  SS_REQUIRE((p->y_dtype == SS_F32 || p->y_dtype == SS_BF16) && (p->residual_dtype == SS_F32 || p->residual_dtype == SS_BF16),
             SS_ERR_INVALID, "%s: bad dtype code", who);
  SS_REQUIRE(p->rows == nullptr || p->B <= RAGGED_MAX_B, SS_ERR_UNSUPPORTED,
             "%s: the ragged schedule (rows) takes at most %d utterances per call", who, RAGGED_MAX_B);
  const bool narrow = p->Cin == 32;
  X3Args q;
  const int resident = plan_smem(bn, narrow, SPLIT, p->Cin, p->Cout, p->k, p->dilation, p->rows ? p->B : 0, &q);
  SS_REQUIRE(resident >= 0, SS_ERR_UNSUPPORTED, "%s: no shared-memory plan for Cin=%d k=%d dil=%d", who, p->Cin, p->k,
             p->dilation);

  TcArgs a;
  a.bias = p->bias; a.residual = p->residual; a.lens = p->lens; a.y = p->y;
  a.y_act = static_cast<__nv_bfloat16*>(p->y_act); a.accum = p->accum;
  // split: the activated copy is a pair row [hi | lo] of pair_c channels; plain: a bf16 copy laid out like y
  a.pair_c = (SPLIT && p->y_act) ? (p->pair_c > 0 ? p->pair_c : p->Cout) : 0;
  SS_REQUIRE(a.pair_c % 8 == 0 && (a.pair_c == 0 || p->Cout % a.pair_c == 0), SS_ERR_INVALID,
             "%s: pair_c (%d) must be a multiple of 8 that divides Cout (%d)", who, a.pair_c, p->Cout);
  a.y_batch_stride = p->y_batch_stride; a.y_row_stride = p->y_row_stride; a.y_offset = p->y_offset;
  a.y_batch_elems = p->y_batch_elems;
  a.B = p->B; a.S_out = p->S_out; a.Cin = p->Cin; a.Cout = p->Cout; a.k = p->k; a.dilation = p->dilation; a.pad = p->pad;
  a.n_tiles = (p->Cout + bn - 1) / bn;
  a.m_tiles = (p->S_out + BM - 1) / BM;
  a.act = p->act; a.act2 = p->act2; a.y_dtype = p->y_dtype; a.residual_dtype = p->residual_dtype;
  a.accum_init = p->accum_init; a.slope = p->slope; a.slope2 = p->slope2; a.accum_scale = p->accum_scale;
  a.rows = p->rows;
  a.l2_prefetch = x3_prefetch_default();
  const int64_t tiles = (int64_t)a.B * a.m_tiles * a.n_tiles;
  SS_REQUIRE(tiles < (1ll << 31), SS_ERR_UNSUPPORTED, "%s: too many tiles", who);

  CUtensorMap tmA, tmW;
  // rows of NP*Cin bf16; a 64-element box past the row's end (plain, 32 channels) is zero-filled by TMA
  int rc = make_map_3d(&tmA, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, p->x, (uint64_t)NP * p->Cin, (uint64_t)p->S_in, (uint64_t)p->B,
                       (uint64_t)p->x_row_stride * 2, (uint64_t)p->x_batch_stride * 2, 64, (uint32_t)q.R);
  if (rc != SS_OK) return rc;
  const int wt_total = a.n_tiles * (narrow ? (p->k + 1) / 2 : (p->Cin / 64) * p->k);
  rc = make_map_3d(&tmW, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, p->w, 64, (uint64_t)NP * bn, (uint64_t)wt_total, 128,
                   (uint64_t)NP * bn * 128, 64, (uint32_t)(NP * bn));
  if (rc != SS_OK) return rc;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (narrow) return launch_k<32, true, true, SPLIT>(tmA, tmW, a, q, tiles, st);
  if (bn == 64)
    return resident ? launch_k<64, false, true, SPLIT>(tmA, tmW, a, q, tiles, st)
                    : launch_k<64, false, false, SPLIT>(tmA, tmW, a, q, tiles, st);
  return resident ? launch_k<128, false, true, SPLIT>(tmA, tmW, a, q, tiles, st)
                  : launch_k<128, false, false, SPLIT>(tmA, tmW, a, q, tiles, st);
}

}  // namespace

extern "C" int ss_conv1d_bf16x3_plan(int Cin, int Cout, int k, int dilation) {
  const int bn = plan_bn(Cin, Cout, k, dilation);
  if (bn == 0) return 0;
  X3Args q;
  return plan_smem(bn, Cin == 32, true, Cin, Cout, k, dilation, RAGGED_MAX_B, &q) < 0 ? 0 : bn;
}

extern "C" int ss_conv1d_bf16x3(const ss_conv1d_tc_params* p, void* stream) {
  return run<true>(p, stream, "ss_conv1d_bf16x3");
}

extern "C" int ss_conv1d_bf16_panel(const ss_conv1d_tc_params* p, void* stream) {
  return run<false>(p, stream, "ss_conv1d_bf16_panel");
}
