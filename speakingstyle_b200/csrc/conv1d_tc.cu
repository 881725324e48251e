// Implicit-GEMM 1-D convolution on the 5th-generation tensor cores (tcgen05), bf16 operands,
// fp32 accumulation in TMEM, operands fed by TMA.  sm_100a only.
//
//   D[t, n] = sum_{tap j} sum_{ci}  X[b, t + j*dil - pad, ci] * W[j, n, ci]
//
// GEMM view per output tile: M = 128 positions (one TMEM lane each), N = BN output channels (TMEM
// columns), K = taps x Cin walked in 64-channel chunks.  There is no im2col: tap j of the A operand
// is the SAME activation tensor fetched by TMA at row coordinate t0 + j*dil - pad; rows outside
// [0, S) - the convolution's zero padding and the halo across utterance boundaries - are zero-filled
// by the TMA unit itself (signed coordinates, OOB fill), so the kernel has no boundary code in its
// main loop.  Both operands are K-major with the 128-byte swizzle (64 bf16 per row), which is
// exactly what one TMA box {64 ch, 128 rows} lands in shared memory.
//
// Warp roles (192 threads, persistent CTA, static round-robin tile schedule):
//   warp 0   one lane issues TMA loads into a STAGES-deep ring (full/empty mbarriers)
//   warp 1   allocates TMEM; one lane issues tcgen05.mma (4 x K=16 per chunk) and commits
//   warps 2-5 epilogue: tcgen05.ld 32 columns at a time from one of TWO accumulator buffers, so the
//            epilogue of tile i overlaps the main loop of tile i+1; bias, activation, residual,
//            row mask, optional running mean, fp32 and/or bf16 stores.
#include "tc_common.cuh"

namespace {

using namespace sstc;

constexpr int KCHUNK = 64;    // bf16 elements per 128-byte swizzle row
// epilogue warps per TMEM lane group: the epilogue is global-memory-latency bound, and a lane group can
// only be read by warps with the same (warp id % 4), so more parallelism means more warps per group
template <int BN> struct EpiG { static constexpr int value = BN <= 128 ? 2 : 1; };
// M sub-tiles per scheduling unit.  With few output channels a 128 x BN tile is tiny (a 32-channel k=3
// convolution is 12 MMAs), and the TMA -> MMA -> epilogue -> MMA hand-offs, not the work, set the pace; so a
// unit is MT consecutive 128-row tiles that share one accumulator buffer (MT*BN columns) and one hand-off.
template <int BN> struct SubTiles { static constexpr int value = BN >= 256 ? 1 : 256 / BN; };

// ---------------------------------------------------------------------------------------------
template <int BN, int STAGES>
struct Cfg {
  static constexpr int b_bytes = BN * KCHUNK * 2;
  static constexpr int stage_bytes = A_BYTES + b_bytes;
  static constexpr int MT = SubTiles<BN>::value;
  static constexpr int acc_cols = MT * BN;  // columns of one accumulator buffer
  static constexpr int tmem_cols = (2 * acc_cols <= 32) ? 32 : (2 * acc_cols <= 64) ? 64 : (2 * acc_cols <= 128) ? 128 : (2 * acc_cols <= 256) ? 256 : 512;
  static constexpr int bar_bytes = (2 * STAGES + 4) * 8 + 16;
  static constexpr int G = EpiG<BN>::value;
  static constexpr int stg_bytes = 4 * G * STG_WARP_BYTES;
  static constexpr int nthreads = 64 + 128 * G;
  static constexpr int ragged_off = STAGES * stage_bytes + stg_bytes + bar_bytes;  // tile prefix table (ragged schedule)
  static constexpr int smem_bytes = ragged_off + 1024;  // +1024: alignment slack; + (B + 1) * 4 at launch when rows are given
  static_assert(smem_bytes + (RAGGED_MAX_B + 1) * 4 <= 227 * 1024, "shared memory budget");
};

template <int BN, int STAGES>
__global__ void __launch_bounds__(Cfg<BN, STAGES>::nthreads, 1)
conv1d_tc_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, const TcArgs p) {
  using C = Cfg<BN, STAGES>;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem0 = (smem_u32(smem_raw) + 1023u) & ~1023u;  // SWIZZLE_128B tiles need 1024-byte alignment
  const uint32_t smA = smem0;
  const uint32_t smB = smem0 + STAGES * A_BYTES;
  const uint32_t bars = smem0 + STAGES * C::stage_bytes + C::stg_bytes;
  const uint32_t full0 = bars, empty0 = bars + 8 * STAGES, tfull0 = bars + 16 * STAGES, tempty0 = tfull0 + 16;
  const uint32_t tslot = tempty0 + 16;
  uint8_t* smem_al = smem_raw + (smem0 - smem_u32(smem_raw));
  float* stg_all = reinterpret_cast<float*>(smem_al + STAGES * C::stage_bytes);
  volatile uint32_t* tslot_ptr =
      reinterpret_cast<volatile uint32_t*>(smem_al + STAGES * C::stage_bytes + C::stg_bytes + (2 * STAGES + 4) * 8);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nchunk = (p.Cin + KCHUNK - 1) / KCHUNK;
  const int KT = p.k * nchunk;
  constexpr int MT = C::MT;
  const int m_units = (p.m_tiles + MT - 1) / MT;
  int* const rag = reinterpret_cast<int*>(smem_al + C::ragged_off);
  const bool ragged = p.rows != nullptr;
  if (ragged) ragged_build(p.rows, p.B, p.S_out, MT * BM, rag);

  if (threadIdx.x == 0) {
    prefetch_tmap(&tmA);
    prefetch_tmap(&tmB);
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(full0 + 8 * s, 1);
      mbar_init(empty0 + 8 * s, 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(tfull0 + 8 * a, 1);
      mbar_init(tempty0 + 8 * a, 4 * C::G);  // one arrival per epilogue warp
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    tmem_alloc(tslot, (uint32_t)C::tmem_cols);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tslot_ptr;
  const int num_tiles = ragged ? rag[p.B] * p.n_tiles : p.B * m_units * p.n_tiles;  // scheduling units
  // linear unit index -> (column tile, unit of MT row tiles, utterance)
  auto decode = [&](int tile, int& nt, int& um, int& b) {
    nt = tile % p.n_tiles;
    const int rest = tile / p.n_tiles;
    if (ragged) {
      ragged_decode(rag, p.B, rest, b, um);
    } else {
      um = rest % m_units;
      b = rest / m_units;
    }
  };

  if (warp == 0) {
    // ================= TMA producer =================
    if (elect_one()) {  // elect.sync, not lane == 0: lets ptxas issue TMA / MMA without a per-instruction elect loop
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        int nt, um, b;
        decode(tile, nt, um, b);
        const int n0 = nt * BN;
        for (int sub = 0; sub < MT; ++sub) {
          const int mt = um * MT + sub;
          if (mt >= p.m_tiles) break;
          const int t0 = mt * BM;
          for (int it = 0; it < KT; ++it) {
            const int j = it / nchunk;
            const int c0 = (it - j * nchunk) * KCHUNK;
            mbar_wait(empty0 + 8 * stage, phase ^ 1);
            mbar_expect_tx(full0 + 8 * stage, C::stage_bytes);
            tma_load_3d(smA + stage * A_BYTES, &tmA, full0 + 8 * stage, c0, t0 + j * p.dilation - p.pad, b);
            tma_load_3d(smB + stage * C::b_bytes, &tmB, full0 + 8 * stage, c0, n0, j);
            if (++stage == STAGES) { stage = 0; phase ^= 1; }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ================= MMA issuer =================
    if (elect_one()) {  // elect.sync, not lane == 0: lets ptxas issue TMA / MMA without a per-instruction elect loop
      // instruction descriptor (cute::UMMA::InstrDescriptor): c=f32 [4,6)=1, a=bf16 [7,10)=1, b=bf16 [10,13)=1,
      // a/b K-major (bits 15,16 = 0), N>>3 [17,23), M>>4 [24,29)
      const uint32_t idesc = make_idesc(1u, BN);
      int stage = 0;
      uint32_t phase = 0;
      int local = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++local) {
        const int as = local & 1;
        const uint32_t aphase = (local >> 1) & 1;
        mbar_wait(tempty0 + 8 * as, aphase ^ 1);  // epilogue has drained this accumulator buffer
        tc_fence_after();
        int nt_, um, b_;
        decode(tile, nt_, um, b_);
        for (int sub = 0; sub < MT; ++sub) {
          if (um * MT + sub >= p.m_tiles) break;
          const uint32_t d_tmem = tmem_base + as * C::acc_cols + sub * BN;
          for (int it = 0; it < KT; ++it) {
            mbar_wait(full0 + 8 * stage, phase);
            tc_fence_after();
            const uint32_t a_addr = smA + stage * A_BYTES;
            const uint32_t b_addr = smB + stage * C::b_bytes;
#pragma unroll
            for (int kk = 0; kk < KCHUNK / 16; ++kk) {
              // advancing K by 16 bf16 = 32 bytes inside the 128-byte swizzle row: bump the start address
              umma_bf16(d_tmem, make_desc(a_addr + kk * 32), make_desc(b_addr + kk * 32), idesc, (it | kk) != 0);
            }
            umma_commit(empty0 + 8 * stage);  // frees the smem slot once these MMAs have read it
            if (++stage == STAGES) { stage = 0; phase ^= 1; }
          }
        }
        umma_commit(tfull0 + 8 * as);  // accumulators of the whole unit complete -> epilogue
      }
    }
  } else {
    // ================= epilogue (warps 2..5) =================
    const int lg = warp & 3;         // TMEM lane group this warp may access
    const int sub = (warp - 2) >> 2;  // which of the G warps of that group
    float* stg = stg_all + (warp - 2) * 32 * STG_PITCH;
    int local = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++local) {
      int nt, um, b;
      decode(tile, nt, um, b);
      const int as = local & 1;
      const uint32_t aphase = (local >> 1) & 1;
      mbar_wait(tfull0 + 8 * as, aphase);
      tc_fence_after();
      const uint32_t acc = tmem_base + ((uint32_t)(lg * 32) << 16) + as * C::acc_cols;
      if constexpr (MT == 1) {
        drain_tile<BN, C::G>(p, acc, stg, lg, lane, b, um, nt * BN, sub);
      } else {
        // the G warps of a lane group share the unit's 32-column blocks round-robin across its sub-tiles
        constexpr int BLK = BN / 32;
        const int len = p.lens ? p.lens[b] : 0x7fffffff;
#pragma unroll 1
        for (int q = sub; q < MT * BLK; q += C::G) {
          const int st = q / BLK, c0 = (q - st * BLK) * 32;
          const int mt = um * MT + st;
          if (mt >= p.m_tiles) break;
          float v[32];
          __syncwarp();
          tmem_ld32(acc + st * BN + c0, v);
          drain_block(p, v, stg, lg, lane, b, mt, nt * BN + c0, len);
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(tempty0 + 8 * as);
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

// ---------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------
template <int BN, int STAGES>
int launch(const CUtensorMap& tmA, const CUtensorMap& tmB, TcArgs a, cudaStream_t st) {
  using C = Cfg<BN, STAGES>;
  a.n_tiles = (a.Cout + BN - 1) / BN;
  a.m_tiles = (a.S_out + BM - 1) / BM;
  const int64_t tiles = (int64_t)a.B * ((a.m_tiles + C::MT - 1) / C::MT) * a.n_tiles;  // scheduling units
  SS_REQUIRE(tiles < (1ll << 31), SS_ERR_UNSUPPORTED, "ss_conv1d_tc: too many tiles");
  static std::atomic<unsigned long long> attr_mask{0};  // per device
  if (needs_attr_on_device(attr_mask)) {
    cudaError_t e = cudaFuncSetAttribute(conv1d_tc_kernel<BN, STAGES>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         C::smem_bytes + (RAGGED_MAX_B + 1) * 4);
    if (e != cudaSuccess) return ss::cuda_fail(e, "ss_conv1d_tc: smem opt-in");
    attr_done_on_device(attr_mask);
  }
  const int sms = sm_count();
  SS_REQUIRE(sms > 0, SS_ERR_CUDA, "ss_conv1d_tc: no CUDA device");
  const int grid = (int)(tiles < sms ? tiles : sms);
  const int smem = C::smem_bytes + (a.rows ? (a.B + 1) * 4 : 0);
  conv1d_tc_kernel<BN, STAGES><<<grid, C::nthreads, smem, st>>>(tmA, tmB, a);
  SS_CHECK_LAUNCH("ss_conv1d_tc");
  return SS_OK;
}

}  // namespace

extern "C" int ss_conv1d_tc(const ss_conv1d_tc_params* p, void* stream) {
  SS_REQUIRE(p && p->x && p->w && (p->y || p->y_act || p->accum), SS_ERR_INVALID, "ss_conv1d_tc: null pointer");
  SS_REQUIRE(p->B > 0 && p->S_in > 0 && p->S_out > 0 && p->Cin > 0 && p->Cout > 0 && p->k > 0 && p->dilation > 0,
             SS_ERR_INVALID, "ss_conv1d_tc: bad shape B=%d S_in=%d S_out=%d Cin=%d Cout=%d k=%d dil=%d", p->B, p->S_in,
             p->S_out, p->Cin, p->Cout, p->k, p->dilation);
  SS_REQUIRE(p->Cin % 8 == 0 && p->Cout % 8 == 0, SS_ERR_UNSUPPORTED,
             "ss_conv1d_tc: Cin (%d) and Cout (%d) must be multiples of 8", p->Cin, p->Cout);
  SS_REQUIRE(p->x_row_stride % 8 == 0 && p->x_batch_stride % 8 == 0 && p->y_row_stride % 8 == 0 &&
                 p->y_batch_stride % 8 == 0 && p->y_offset % 8 == 0 && p->y_batch_elems % 8 == 0,
             SS_ERR_ALIGNMENT, "ss_conv1d_tc: strides/offsets must be multiples of 8 elements");
  SS_REQUIRE(ss::aligned16(p->x) && ss::aligned16(p->w) && ss::aligned16(p->y) && ss::aligned16(p->y_act) &&
                 ss::aligned16(p->bias) && ss::aligned16(p->residual) && ss::aligned16(p->accum),
             SS_ERR_ALIGNMENT, "ss_conv1d_tc: pointers must be 16-byte aligned");
  SS_REQUIRE((p->y_dtype == SS_F32 || p->y_dtype == SS_BF16) && (p->residual_dtype == SS_F32 || p->residual_dtype == SS_BF16),
             SS_ERR_INVALID, "ss_conv1d_tc: bad dtype code");

  SS_REQUIRE(p->rows == nullptr || p->B <= RAGGED_MAX_B, SS_ERR_UNSUPPORTED,
             "ss_conv1d_tc: the ragged schedule (rows) takes at most %d utterances per call", RAGGED_MAX_B);
  CUtensorMap tmA, tmB;
  int rc = make_map_3d(&tmA, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, p->x, (uint64_t)p->Cin, (uint64_t)p->S_in, (uint64_t)p->B, (uint64_t)p->x_row_stride * 2,
                       (uint64_t)p->x_batch_stride * 2, KCHUNK, BM);
  if (rc != SS_OK) return rc;
  const int bn = p->Cout >= 256 ? 256 : p->Cout > 64 ? 128 : p->Cout > 32 ? 64 : 32;
  rc = make_map_3d(&tmB, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, p->w, (uint64_t)p->Cin, (uint64_t)p->Cout, (uint64_t)p->k, (uint64_t)p->Cin * 2,
                   (uint64_t)p->Cout * p->Cin * 2, KCHUNK, (uint32_t)bn);
  if (rc != SS_OK) return rc;

  TcArgs a;
  a.bias = p->bias; a.residual = p->residual; a.lens = p->lens; a.y = p->y;
  a.y_act = static_cast<__nv_bfloat16*>(p->y_act); a.accum = p->accum;
  a.y_batch_stride = p->y_batch_stride; a.y_row_stride = p->y_row_stride; a.y_offset = p->y_offset;
  a.y_batch_elems = p->y_batch_elems;
  a.B = p->B; a.S_out = p->S_out; a.Cin = p->Cin; a.Cout = p->Cout; a.k = p->k; a.dilation = p->dilation; a.pad = p->pad;
  a.m_tiles = a.n_tiles = 0;
  a.act = p->act; a.act2 = p->act2; a.y_dtype = p->y_dtype; a.residual_dtype = p->residual_dtype;
  a.accum_init = p->accum_init; a.slope = p->slope; a.slope2 = p->slope2; a.accum_scale = p->accum_scale;
  a.rows = p->rows;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  switch (bn) {
    case 256: return launch<256, 4>(tmA, tmB, a, st);
    case 128: return launch<128, 5>(tmA, tmB, a, st);
    case 64: return launch<64, 7>(tmA, tmB, a, st);
    default: return launch<32, 8>(tmA, tmB, a, st);
  }
}
