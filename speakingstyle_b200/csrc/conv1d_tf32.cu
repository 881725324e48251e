// fp32-accurate implicit-GEMM 1-D convolution on the tensor cores: 3xTF32 error-compensated
// products (tcgen05.mma kind::tf32, fp32 accumulation in TMEM), fp32 activations in and out.
//
// Why: every integer this path outputs (durations, pitch / energy bucket ids) is decided by the
// ReferenceEncoder -> Encoder -> predictor chain and flips at ~1e-5 relative error, so that chain
// cannot run in bf16 or single-pass TF32.  Splitting each fp32 operand as  x = hi + lo  with hi the
// nearest TF32 value and lo = tf32(x - hi), and summing  lo*Whi + hi*Wlo + hi*Whi  in fp32, leaves
// ~2^-22 relative error per product - measured below plain fp32 FMA accumulation.
//
// Data flow per K-chunk (32 channels = one 128-byte swizzle row):
//   TMA      x tile [128 x 32] fp32 -> smem A ; pre-split weights Whi, Wlo [BN x 32] -> smem
//   4 transform warps: read A, (optional leaky-ReLU), write hi back in place and lo to a second
//            buffer - purely element-wise, so the swizzled layout never has to be decoded -
//            then fence.proxy.async + arrive
//   1 thread: 3 x 4 tcgen05.mma (K = 8 each) into the TMEM accumulator; commit frees the stage
//   4 epilogue warps: the same coalesced drain as the bf16 kernel (tc_common.cuh)
// Same persistent, warp-specialised structure and the same TMA zero-fill for padding / halo.
//
// Accumulation.  Measured on B200 (scripts/tf32_probe.py): tcgen05.mma adds into the TMEM accumulator
// with truncation, losing ~0.5 ulp of the ACCUMULATOR per instruction - a systematic shrink that
// reaches 2e-5 relative after the 864 instructions of a K = 2304 reduction, 20x the error of fp32 FMA.
// Two measures bring it back under the FMA kernel's error:
//   * the two correction products (lo*Whi, hi*Wlo) go to their OWN accumulator: they are 2^-11 of the
//     main term, so neither their truncation nor their being added last matters, and the main
//     accumulator sees one instruction per K=8 step instead of three;
//   * the main accumulator is restarted every GROUP chunks (GROUP*4 instructions) and the epilogue
//     warps add the partial tiles in registers with round-to-nearest fp32 adds, overlapped with the
//     MMAs of the next group (two main buffers + two correction buffers = 4*BN TMEM columns).
#include "tc_common.cuh"

namespace {

using namespace sstc;

constexpr int KCHUNK = 32;  // fp32 elements per 128-byte swizzle row
constexpr int NTHREADS = 320;

template <int BN, int STAGES>
struct Cfg {
  static constexpr int b_bytes = BN * ROW_BYTES;
  static constexpr int stage_bytes = 2 * A_BYTES + 2 * b_bytes;  // A (-> hi), A_lo, W_hi, W_lo
  static constexpr int tx_bytes = A_BYTES + 2 * b_bytes;         // what TMA writes per stage
  static constexpr int tmem_cols = (4 * BN <= 32) ? 32 : (4 * BN <= 64) ? 64 : (4 * BN <= 128) ? 128 : (4 * BN <= 256) ? 256 : 512;
  static constexpr int n_bars = 3 * STAGES + 8;
  static constexpr int smem_bytes = STAGES * stage_bytes + STG_BYTES + n_bars * 8 + 16 + 1024;
};

// Debug build (-DSS_CONV_TRACE, scripts/conv_trace.py): CTA 0 stamps clock64() per K-chunk at the hand-overs.
#ifdef SS_CONV_TRACE
__device__ long long g_conv_ss_trace[8 * 256];
#define CONV_TRACE(ev, i)                                                              \
  do {                                                                                 \
    if (blockIdx.x == 0 && (i) < 256) g_conv_ss_trace[(ev) * 256 + (i)] = clock64();           \
  } while (0)
#else
#define CONV_TRACE(ev, i)
#endif


template <int BN, int STAGES>
__global__ void __launch_bounds__(NTHREADS, 1)
conv1d_tf32_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmBh,
                   const __grid_constant__ CUtensorMap tmBl, const TcArgs p, const int in_act, const float in_slope,
                   const int group) {
  using C = Cfg<BN, STAGES>;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem0 = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* smem_al = smem_raw + (smem0 - smem_u32(smem_raw));
  // stage s: [A | A_lo | W_hi | W_lo]
  auto off_a = [](int s) { return s * C::stage_bytes; };
  auto off_alo = [](int s) { return s * C::stage_bytes + A_BYTES; };
  auto off_bh = [](int s) { return s * C::stage_bytes + 2 * A_BYTES; };
  auto off_bl = [](int s) { return s * C::stage_bytes + 2 * A_BYTES + C::b_bytes; };
  float* stg_all = reinterpret_cast<float*>(smem_al + STAGES * C::stage_bytes);
  const uint32_t bars = smem0 + STAGES * C::stage_bytes + STG_BYTES;
  const uint32_t full0 = bars, xf0 = bars + 8 * STAGES, empty0 = bars + 16 * STAGES, tfull0 = bars + 24 * STAGES,
                 tempty0 = tfull0 + 16, cfull0 = tempty0 + 16, cempty0 = cfull0 + 16, tslot = cempty0 + 16;
  volatile uint32_t* tslot_ptr =
      reinterpret_cast<volatile uint32_t*>(smem_al + STAGES * C::stage_bytes + STG_BYTES + C::n_bars * 8);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nchunk = (p.Cin + KCHUNK - 1) / KCHUNK;
  const int KT = p.k * nchunk;
  const int num_tiles = p.B * p.m_tiles * p.n_tiles;

  if (threadIdx.x == 0) {
    prefetch_tmap(&tmA);
    prefetch_tmap(&tmBh);
    prefetch_tmap(&tmBl);
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(full0 + 8 * s, 1);
      mbar_init(xf0 + 8 * s, 4);  // one arrival per transform warp
      mbar_init(empty0 + 8 * s, 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(tfull0 + 8 * a, 1);
      mbar_init(tempty0 + 8 * a, 4);
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

  if (warp == 0) {
    // ================= TMA producer =================
    if (elect_one()) {  // elect.sync, not lane == 0: lets ptxas issue TMA / MMA without a per-instruction elect loop
      int stage = 0;
      uint32_t phase = 0;
      [[maybe_unused]] int gi = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        const int nt = tile % p.n_tiles;
        const int rest = tile / p.n_tiles;
        const int mt = rest % p.m_tiles;
        const int b = rest / p.m_tiles;
        const int t0 = mt * BM, n0 = nt * BN;
        for (int it = 0; it < KT; ++it) {
          const int j = it / nchunk;
          const int c0 = (it - j * nchunk) * KCHUNK;
          mbar_wait(empty0 + 8 * stage, phase ^ 1);
          CONV_TRACE(0, gi); ++gi;
          mbar_expect_tx(full0 + 8 * stage, C::tx_bytes);
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
      int stage = 0;
      uint32_t phase = 0;
      int local = 0;   // tiles done by this CTA -> correction buffer / phase
      int gcount = 0;  // K-groups done by this CTA -> main buffer / phase
      [[maybe_unused]] int gi = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++local) {
        const int cs = local & 1;
        mbar_wait(cempty0 + 8 * cs, ((local >> 1) & 1) ^ 1);
        const uint32_t c_tmem = tmem_base + (2 + cs) * BN;
        for (int it0 = 0; it0 < KT; it0 += group, ++gcount) {
          const int as = gcount & 1;
          mbar_wait(tempty0 + 8 * as, ((gcount >> 1) & 1) ^ 1);
          tc_fence_after();
          const uint32_t d_tmem = tmem_base + as * BN;
          const int it1 = min(it0 + group, KT);
          for (int it = it0; it < it1; ++it) {
            mbar_wait(xf0 + 8 * stage, phase);  // hi / lo written (implies the TMA data has landed)
            tc_fence_after();
            CONV_TRACE(3, gi);
            const uint32_t a_hi = smem0 + off_a(stage), a_lo = smem0 + off_alo(stage);
            const uint32_t b_hi = smem0 + off_bh(stage), b_lo = smem0 + off_bl(stage);
#pragma unroll
            for (int kk = 0; kk < KCHUNK / 8; ++kk) {  // K = 8 tf32 = 32 bytes per instruction
              const uint32_t o = kk * 32;
              umma_tf32(c_tmem, make_desc(a_lo + o), make_desc(b_hi + o), idesc, (it | kk) != 0);
              umma_tf32(c_tmem, make_desc(a_hi + o), make_desc(b_lo + o), idesc, 1u);
              umma_tf32(d_tmem, make_desc(a_hi + o), make_desc(b_hi + o), idesc, ((it - it0) | kk) != 0);
            }
            umma_commit(empty0 + 8 * stage);
            CONV_TRACE(4, gi); ++gi;
            if (++stage == STAGES) { stage = 0; phase ^= 1; }
          }
          umma_commit(tfull0 + 8 * as);  // this group's partial sum is complete
        }
        umma_commit(cfull0 + 8 * cs);  // correction terms of the whole tile are complete
      }
    }
  } else if (warp < 6) {
    // ================= epilogue (warps 2..5) =================
    const int lg = warp & 3;
    float* stg = stg_all + lg * 32 * STG_PITCH;
    const uint32_t lane_base = tmem_base + ((uint32_t)(lg * 32) << 16);
    int local = 0, gcount = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++local) {
      const int nt = tile % p.n_tiles;
      const int rest = tile / p.n_tiles;
      const int mt = rest % p.m_tiles;
      const int b = rest / p.m_tiles;
      int b_t = b, t_tile0 = mt * BM;
      bool straddle = false;
      if (p.flat) {  // un-flatten the tile's first row once; most tiles lie inside one utterance
        b_t = t_tile0 / p.S_out;
        t_tile0 -= b_t * p.S_out;
        straddle = t_tile0 + BM > p.S_out;
      }
      const int len = p.lens ? p.lens[b_t] : 0x7fffffff;
      const bool lean = drain_is_lean<(BN <= 64)>(p);  // warp-uniform
      if (lean && p.residual) {
        // Pull this tile's residual rows into L2 now, while the MMAs run: the drain below keeps only four 16-byte
        // loads per lane in flight, which from HBM (~1 us) capped the whole kernel at ~1.2 TB/s of residual reads.
        int t = t_tile0 + lg * 32 + lane, bb = b_t;
        if (straddle)
          while (t >= p.S_out) { t -= p.S_out; ++bb; }
        const int64_t off = p.y_offset + (int64_t)t * p.y_row_stride + nt * BN;
        if (bb < p.B && t < p.S_out && off >= 0 && off < p.y_batch_elems) {
          const float* rp = static_cast<const float*>(p.residual) + (int64_t)bb * p.y_batch_stride + off;
          const int ncol = min(BN, p.Cout - nt * BN);
          for (int c = 0; c < ncol; c += 32)  // one 128-byte line per step
            asm volatile("prefetch.global.L2 [%0];" ::"l"(rp + c));
        }
      }
      float acc[BN];
#pragma unroll
      for (int i = 0; i < BN; ++i) acc[i] = 0.f;
      const int cs = local & 1;
      for (int it0 = 0; it0 < KT; it0 += group, ++gcount) {
        if (it0 + group >= KT) {  // last group: correction terms first (same summation order as the TS kernel)
          mbar_wait(cfull0 + 8 * cs, (local >> 1) & 1);
          tc_fence_after();
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
      }
#pragma unroll
      for (int c0 = 0; c0 < BN; c0 += 32) {  // TMEM is free again: the stores overlap the next tile's MMAs
        float v[32];
#pragma unroll
        for (int i = 0; i < 32; ++i) v[i] = acc[c0 + i];
        if (lean)
          drain_block_lean<(BN <= 64)>(p, v, stg, lg, lane, b_t, t_tile0, straddle, nt * BN + c0, len);
        else
          drain_block(p, v, stg, lg, lane, b, mt, nt * BN + c0, len);
      }
    }
  } else {
    // ================= operand split (warps 6..9) =================
    const int tt = (warp - 6) * 32 + lane;  // 0..127
    int stage = 0;
    uint32_t phase = 0;
    [[maybe_unused]] int gi = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
      for (int it = 0; it < KT; ++it) {
        mbar_wait(full0 + 8 * stage, phase);
        if (threadIdx.x == 192) CONV_TRACE(1, gi);
        float4* a = reinterpret_cast<float4*>(smem_al + off_a(stage));
        float4* alo = reinterpret_cast<float4*>(smem_al + off_alo(stage));
#pragma unroll
        for (int i = 0; i < A_BYTES / 16 / 128; ++i) {
          const int idx = tt + i * 128;
          float4 x = a[idx];
          if (in_act == SS_ACT_LRELU) {
            x.x = x.x > 0.f ? x.x : x.x * in_slope;
            x.y = x.y > 0.f ? x.y : x.y * in_slope;
            x.z = x.z > 0.f ? x.z : x.z * in_slope;
            x.w = x.w > 0.f ? x.w : x.w * in_slope;
          }
          const float4 h = make_float4(tf32_rn(x.x), tf32_rn(x.y), tf32_rn(x.z), tf32_rn(x.w));
          a[idx] = h;
          alo[idx] = make_float4(x.x - h.x, x.y - h.y, x.z - h.z, x.w - h.w);  // truncated to TF32 by the MMA
        }
        fence_proxy_async();  // make these generic-proxy writes visible to tcgen05.mma
        __syncwarp();
        if (lane == 0) mbar_arrive(xf0 + 8 * stage);
        if (threadIdx.x == 192) CONV_TRACE(2, gi);
        ++gi;
        if (++stage == STAGES) { stage = 0; phase ^= 1; }
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
  a.m_tiles = (a.S_out + BM - 1) / BM;
  const int64_t tiles = (int64_t)a.B * a.m_tiles * a.n_tiles;
  SS_REQUIRE(tiles < (1ll << 31), SS_ERR_UNSUPPORTED, "ss_conv1d_tf32x3: too many tiles");
  static std::atomic<unsigned long long> attr_mask{0};  // per device
  if (needs_attr_on_device(attr_mask)) {
    cudaError_t e = cudaFuncSetAttribute(conv1d_tf32_kernel<BN, STAGES>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         C::smem_bytes);
    if (e != cudaSuccess) return ss::cuda_fail(e, "ss_conv1d_tf32x3: smem opt-in");
    attr_done_on_device(attr_mask);
  }
  const int sms = sm_count();
  SS_REQUIRE(sms > 0, SS_ERR_CUDA, "ss_conv1d_tf32x3: no CUDA device");
  const int grid = (int)(tiles < sms ? tiles : sms);
  conv1d_tf32_kernel<BN, STAGES><<<grid, NTHREADS, C::smem_bytes, st>>>(tmA, tmBh, tmBl, a, in_act, in_slope, group);
  SS_CHECK_LAUNCH("ss_conv1d_tf32x3");
  return SS_OK;
}

}  // namespace

#ifdef SS_CONV_TRACE
extern "C" __attribute__((visibility("default"))) int ss_debug_conv_ss_trace(long long* host) {
  return (int)cudaMemcpyFromSymbol(host, g_conv_ss_trace, sizeof(long long) * 8 * 256);
}
#endif

extern "C" int ss_conv1d_tf32x3(const ss_conv1d_tf32_params* p, void* stream) {
  SS_REQUIRE(p && p->x && p->w_hi && p->w_lo && (p->y || p->y_act || p->accum), SS_ERR_INVALID,
             "ss_conv1d_tf32x3: null pointer");
  SS_REQUIRE(p->B > 0 && p->S_in > 0 && p->S_out > 0 && p->Cin > 0 && p->Cout > 0 && p->k > 0 && p->dilation > 0,
             SS_ERR_INVALID, "ss_conv1d_tf32x3: bad shape B=%d S_in=%d S_out=%d Cin=%d Cout=%d k=%d dil=%d", p->B,
             p->S_in, p->S_out, p->Cin, p->Cout, p->k, p->dilation);
  SS_REQUIRE(p->Cin % 4 == 0 && p->Cout % 4 == 0, SS_ERR_UNSUPPORTED,
             "ss_conv1d_tf32x3: Cin (%d) and Cout (%d) must be multiples of 4", p->Cin, p->Cout);
  SS_REQUIRE(p->x_row_stride % 4 == 0 && p->x_batch_stride % 4 == 0 && p->y_row_stride % 4 == 0 &&
                 p->y_batch_stride % 4 == 0 && p->y_offset % 4 == 0 && p->y_batch_elems % 4 == 0,
             SS_ERR_ALIGNMENT, "ss_conv1d_tf32x3: strides/offsets must be multiples of 4 elements");
  SS_REQUIRE(ss::aligned16(p->x) && ss::aligned16(p->w_hi) && ss::aligned16(p->w_lo) && ss::aligned16(p->y) &&
                 ss::aligned16(p->y_act) && ss::aligned16(p->bias) && ss::aligned16(p->residual) && ss::aligned16(p->accum),
             SS_ERR_ALIGNMENT, "ss_conv1d_tf32x3: pointers must be 16-byte aligned");
  SS_REQUIRE(p->in_act == SS_ACT_NONE || p->in_act == SS_ACT_LRELU, SS_ERR_UNSUPPORTED,
             "ss_conv1d_tf32x3: in_act must be none or lrelu");

  const int bn = p->Cout > 64 ? 128 : p->Cout > 32 ? 64 : 32;
  CUtensorMap tmA, tmBh, tmBl;
  int rc = make_map_3d(&tmA, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, p->x, (uint64_t)p->Cin, (uint64_t)p->S_in, (uint64_t)p->B,
                       (uint64_t)p->x_row_stride * 4, (uint64_t)p->x_batch_stride * 4, KCHUNK, BM);
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
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int group = p->k_group > 0 ? p->k_group : 8;  // chunks (of 32 channels) per main-accumulator restart
  switch (bn) {
    case 128: return launch<128, 3>(tmA, tmBh, tmBl, a, p->in_act, p->in_slope, group, st);
    case 64: return launch<64, 4>(tmA, tmBh, tmBl, a, p->in_act, p->in_slope, group, st);
    default: return launch<32, 4>(tmA, tmBh, tmBl, a, p->in_act, p->in_slope, group, st);
  }
}
