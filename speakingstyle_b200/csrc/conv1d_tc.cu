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
#include <cuda.h>
#include <cuda_bf16.h>

#include "common.cuh"

namespace {

constexpr int BM = 128;       // rows per tile = TMEM lanes
constexpr int KCHUNK = 64;    // bf16 elements per 128-byte swizzle row
constexpr int A_BYTES = BM * KCHUNK * 2;
constexpr int NTHREADS = 192;

struct TcArgs {
  const float* bias;
  const void* residual;
  const int32_t* lens;
  void* y;
  __nv_bfloat16* y_act;
  float* accum;
  int64_t y_batch_stride, y_row_stride, y_offset, y_batch_elems;
  int32_t B, S_out, Cin, Cout, k, dilation, pad;
  int32_t m_tiles, n_tiles;
  int32_t act, act2, y_dtype, residual_dtype, accum_init;
  float slope, slope2, accum_scale;
};

// ---------------------------------------------------------------------------------------------
// PTX wrappers
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "WAIT_LOOP:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra WAIT_DONE;\n\t"
      "bra WAIT_LOOP;\n\t"
      "WAIT_DONE:\n\t"
      "}" ::"r"(bar),
      "r"(parity)
      : "memory");
}
__device__ __forceinline__ void tma_load_3d(uint32_t dst, const CUtensorMap* map, uint32_t bar, int c0, int c1,
                                            int c2) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];" ::
          "r"(dst),
      "l"(reinterpret_cast<uint64_t>(map)), "r"(bar), "r"(c0), "r"(c1), "r"(c2)
      : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// K-major, SWIZZLE_128B shared-memory matrix descriptor (cute::UMMA::SmemDescriptor bit layout):
// start>>4 [0,14) | LBO>>4 [16,30) (unused for swizzled K-major, 1) | SBO>>4 [32,46) = 1024 B between
// 8-row groups | version=1 [46,48) | layout_type=2 (SWIZZLE_128B) [61,64)
__device__ __forceinline__ uint64_t make_desc(uint32_t smem_addr) {
  return (uint64_t)((smem_addr & 0x3FFFFu) >> 4) | (1ull << 16) | (64ull << 32) | (1ull << 46) | (2ull << 61);
}
__device__ __forceinline__ void umma_bf16(uint32_t tmem_d, uint64_t da, uint64_t db, uint32_t idesc, uint32_t accum) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t"
      "}" ::"r"(tmem_d),
      "l"(da), "l"(db), "r"(idesc), "r"(accum)
      : "memory");
}
__device__ __forceinline__ void umma_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, float (&v)[32]) {
  uint32_t r[32];
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
        "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
        "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr));
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
  for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(r[i]);
}

__device__ __forceinline__ uint32_t pack_bf16(float a, float b) {
  __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&h);
}
__device__ __forceinline__ void unpack_bf16(uint32_t u, float& a, float& b) {
  a = __uint_as_float(u << 16);
  b = __uint_as_float(u & 0xFFFF0000u);
}

// ---------------------------------------------------------------------------------------------
template <int BN, int STAGES>
struct Cfg {
  static constexpr int b_bytes = BN * KCHUNK * 2;
  static constexpr int stage_bytes = A_BYTES + b_bytes;
  static constexpr int tmem_cols = (2 * BN <= 32) ? 32 : (2 * BN <= 64) ? 64 : (2 * BN <= 128) ? 128 : (2 * BN <= 256) ? 256 : 512;
  static constexpr int bar_bytes = (2 * STAGES + 4) * 8 + 16;
  static constexpr int smem_bytes = STAGES * stage_bytes + bar_bytes + 1024;  // +1024: manual alignment slack
};

template <int BN, int STAGES>
__global__ void __launch_bounds__(NTHREADS, 1)
conv1d_tc_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, const TcArgs p) {
  using C = Cfg<BN, STAGES>;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem0 = (smem_u32(smem_raw) + 1023u) & ~1023u;  // SWIZZLE_128B tiles need 1024-byte alignment
  const uint32_t smA = smem0;
  const uint32_t smB = smem0 + STAGES * A_BYTES;
  const uint32_t bars = smem0 + STAGES * C::stage_bytes;
  const uint32_t full0 = bars, empty0 = bars + 8 * STAGES, tfull0 = bars + 16 * STAGES, tempty0 = tfull0 + 16;
  const uint32_t tslot = tempty0 + 16;
  uint8_t* smem_al = smem_raw + (smem0 - smem_u32(smem_raw));
  volatile uint32_t* tslot_ptr = reinterpret_cast<volatile uint32_t*>(smem_al + STAGES * C::stage_bytes + (2 * STAGES + 4) * 8);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nchunk = (p.Cin + KCHUNK - 1) / KCHUNK;
  const int KT = p.k * nchunk;
  const int num_tiles = p.B * p.m_tiles * p.n_tiles;

  if (threadIdx.x == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmA)) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmB)) : "memory");
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(full0 + 8 * s, 1);
      mbar_init(empty0 + 8 * s, 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(tfull0 + 8 * a, 1);
      mbar_init(tempty0 + 8 * a, 4);  // one arrival per epilogue warp
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tslot), "r"((uint32_t)C::tmem_cols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tslot_ptr;

  if (warp == 0) {
    // ================= TMA producer =================
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
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
          mbar_expect_tx(full0 + 8 * stage, C::stage_bytes);
          tma_load_3d(smA + stage * A_BYTES, &tmA, full0 + 8 * stage, c0, t0 + j * p.dilation - p.pad, b);
          tma_load_3d(smB + stage * C::b_bytes, &tmB, full0 + 8 * stage, c0, n0, j);
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ================= MMA issuer =================
    if (lane == 0) {
      // instruction descriptor (cute::UMMA::InstrDescriptor): c=f32 [4,6)=1, a=bf16 [7,10)=1, b=bf16 [10,13)=1,
      // a/b K-major (bits 15,16 = 0), N>>3 [17,23), M>>4 [24,29)
      const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);
      int stage = 0;
      uint32_t phase = 0;
      int local = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++local) {
        const int as = local & 1;
        const uint32_t aphase = (local >> 1) & 1;
        mbar_wait(tempty0 + 8 * as, aphase ^ 1);  // epilogue has drained this accumulator buffer
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + as * BN;
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
        umma_commit(tfull0 + 8 * as);  // accumulator complete -> epilogue
      }
    }
  } else {
    // ================= epilogue (warps 2..5) =================
    const int lg = warp & 3;  // TMEM lane group this warp may access
    const int row = lg * 32 + lane;
    int local = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++local) {
      const int nt = tile % p.n_tiles;
      const int rest = tile / p.n_tiles;
      const int mt = rest % p.m_tiles;
      const int b = rest / p.m_tiles;
      const int t = mt * BM + row, n0 = nt * BN;
      const int as = local & 1;
      const uint32_t aphase = (local >> 1) & 1;
      mbar_wait(tfull0 + 8 * as, aphase);
      tc_fence_after();
      const uint32_t taddr = tmem_base + ((uint32_t)(lg * 32) << 16) + as * BN;
      const bool row_ok = t < p.S_out;
      const bool masked = p.lens && t >= p.lens[b];
      const int64_t ybase = (int64_t)b * p.y_batch_stride;
      const int64_t roff = p.y_offset + (int64_t)t * p.y_row_stride;
#pragma unroll 1
      for (int c0 = 0; c0 < BN; c0 += 32) {
        float v[32];
        __syncwarp();              // tcgen05.ld is .sync.aligned: reconverge after the divergent stores below
        tmem_ld32(taddr + c0, v);  // warp-collective: always executed
        if (!row_ok || n0 + c0 >= p.Cout) continue;
#pragma unroll
        for (int g = 0; g < 4; ++g) {
          const int n = n0 + c0 + g * 8;
          if (n >= p.Cout) break;
          const int64_t off = roff + n;
          if (off < 0 || off >= p.y_batch_elems) continue;
          float* f = &v[g * 8];
          if (p.bias) {
            const float4 b0 = __ldg(reinterpret_cast<const float4*>(p.bias + n));
            const float4 b1 = __ldg(reinterpret_cast<const float4*>(p.bias + n) + 1);
            f[0] += b0.x; f[1] += b0.y; f[2] += b0.z; f[3] += b0.w;
            f[4] += b1.x; f[5] += b1.y; f[6] += b1.z; f[7] += b1.w;
          }
#pragma unroll
          for (int i = 0; i < 8; ++i) f[i] = ss::apply_act(f[i], p.act, p.slope);
          if (p.residual) {
            if (p.residual_dtype == SS_F32) {
              const float4* rp = reinterpret_cast<const float4*>(static_cast<const float*>(p.residual) + ybase + off);
              const float4 r0 = __ldg(rp), r1 = __ldg(rp + 1);
              f[0] += r0.x; f[1] += r0.y; f[2] += r0.z; f[3] += r0.w;
              f[4] += r1.x; f[5] += r1.y; f[6] += r1.z; f[7] += r1.w;
            } else {
              const uint4 r = __ldg(reinterpret_cast<const uint4*>(static_cast<const __nv_bfloat16*>(p.residual) + ybase + off));
              float a, c;
              unpack_bf16(r.x, a, c); f[0] += a; f[1] += c;
              unpack_bf16(r.y, a, c); f[2] += a; f[3] += c;
              unpack_bf16(r.z, a, c); f[4] += a; f[5] += c;
              unpack_bf16(r.w, a, c); f[6] += a; f[7] += c;
            }
          }
          if (masked) {
#pragma unroll
            for (int i = 0; i < 8; ++i) f[i] = 0.f;
          }
          if (p.y) {
            if (p.y_dtype == SS_F32) {
              float4* yp = reinterpret_cast<float4*>(static_cast<float*>(p.y) + ybase + off);
              yp[0] = make_float4(f[0], f[1], f[2], f[3]);
              yp[1] = make_float4(f[4], f[5], f[6], f[7]);
            } else {
              *reinterpret_cast<uint4*>(static_cast<__nv_bfloat16*>(p.y) + ybase + off) =
                  make_uint4(pack_bf16(f[0], f[1]), pack_bf16(f[2], f[3]), pack_bf16(f[4], f[5]), pack_bf16(f[6], f[7]));
            }
          }
          if (p.accum) {  // running mean; the activated copy below then follows the accumulated value
            float4* ap = reinterpret_cast<float4*>(p.accum + ybase + off);
            float4 a0 = make_float4(0.f, 0.f, 0.f, 0.f), a1 = a0;
            if (!p.accum_init) { a0 = ap[0]; a1 = ap[1]; }
            f[0] = fmaf(p.accum_scale, f[0], a0.x); f[1] = fmaf(p.accum_scale, f[1], a0.y);
            f[2] = fmaf(p.accum_scale, f[2], a0.z); f[3] = fmaf(p.accum_scale, f[3], a0.w);
            f[4] = fmaf(p.accum_scale, f[4], a1.x); f[5] = fmaf(p.accum_scale, f[5], a1.y);
            f[6] = fmaf(p.accum_scale, f[6], a1.z); f[7] = fmaf(p.accum_scale, f[7], a1.w);
            ap[0] = make_float4(f[0], f[1], f[2], f[3]);
            ap[1] = make_float4(f[4], f[5], f[6], f[7]);
          }
          if (p.y_act) {
#pragma unroll
            for (int i = 0; i < 8; ++i) f[i] = ss::apply_act(f[i], p.act2, p.slope2);
            *reinterpret_cast<uint4*>(p.y_act + ybase + off) =
                make_uint4(pack_bf16(f[0], f[1]), pack_bf16(f[2], f[3]), pack_bf16(f[4], f[5]), pack_bf16(f[6], f[7]));
          }
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
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"((uint32_t)C::tmem_cols) : "memory");
  }
}

// ---------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = []() -> EncodeTiledFn {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess ||
        q != cudaDriverEntryPointSuccess)
      return nullptr;
    return reinterpret_cast<EncodeTiledFn>(p);
  }();
  return fn;
}

int make_map_3d(CUtensorMap* m, const void* base, uint64_t d0, uint64_t d1, uint64_t d2, uint64_t s1_bytes,
                uint64_t s2_bytes, uint32_t box0, uint32_t box1) {
  EncodeTiledFn fn = encode_fn();
  SS_REQUIRE(fn != nullptr, SS_ERR_CUDA, "ss_conv1d_tc: cuTensorMapEncodeTiled is not available from this driver");
  cuuint64_t dims[3] = {d0, d1, d2};
  cuuint64_t strides[2] = {s1_bytes, s2_bytes};
  cuuint32_t box[3] = {box0, box1, 1};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = fn(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(base), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  SS_REQUIRE(r == CUDA_SUCCESS, SS_ERR_CUDA, "ss_conv1d_tc: cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
  return SS_OK;
}

int sm_count() {
  static int n = []() {
    int dev = 0, v = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return 0;
    if (cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) return 0;
    return v;
  }();
  return n;
}

template <int BN, int STAGES>
int launch(const CUtensorMap& tmA, const CUtensorMap& tmB, TcArgs a, cudaStream_t st) {
  using C = Cfg<BN, STAGES>;
  a.n_tiles = (a.Cout + BN - 1) / BN;
  a.m_tiles = (a.S_out + BM - 1) / BM;
  const int64_t tiles = (int64_t)a.B * a.m_tiles * a.n_tiles;
  SS_REQUIRE(tiles < (1ll << 31), SS_ERR_UNSUPPORTED, "ss_conv1d_tc: too many tiles");
  static bool attr_set = false;  // per template instantiation
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(conv1d_tc_kernel<BN, STAGES>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         C::smem_bytes);
    if (e != cudaSuccess) return ss::cuda_fail(e, "ss_conv1d_tc: smem opt-in");
    attr_set = true;
  }
  const int sms = sm_count();
  SS_REQUIRE(sms > 0, SS_ERR_CUDA, "ss_conv1d_tc: no CUDA device");
  const int grid = (int)(tiles < sms ? tiles : sms);
  conv1d_tc_kernel<BN, STAGES><<<grid, NTHREADS, C::smem_bytes, st>>>(tmA, tmB, a);
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

  CUtensorMap tmA, tmB;
  int rc = make_map_3d(&tmA, p->x, (uint64_t)p->Cin, (uint64_t)p->S_in, (uint64_t)p->B, (uint64_t)p->x_row_stride * 2,
                       (uint64_t)p->x_batch_stride * 2, KCHUNK, BM);
  if (rc != SS_OK) return rc;
  const int bn = p->Cout >= 256 ? 256 : p->Cout > 64 ? 128 : p->Cout > 32 ? 64 : 32;
  rc = make_map_3d(&tmB, p->w, (uint64_t)p->Cin, (uint64_t)p->Cout, (uint64_t)p->k, (uint64_t)p->Cin * 2,
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
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  switch (bn) {
    case 256: return launch<256, 4>(tmA, tmB, a, st);
    case 128: return launch<128, 6>(tmA, tmB, a, st);
    case 64: return launch<64, 8>(tmA, tmB, a, st);
    default: return launch<32, 8>(tmA, tmB, a, st);
  }
}
