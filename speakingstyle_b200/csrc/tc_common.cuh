// Shared pieces of the tcgen05 convolution kernels (bf16 and 3xTF32): PTX wrappers for mbarrier,
// TMA, tcgen05 / TMEM, the shared-memory matrix descriptor, and the epilogue that drains one
// accumulator tile from TMEM to global memory with coalesced 128-bit accesses.
#pragma once
#include <cuda.h>
#include <cuda_bf16.h>

#include <atomic>

#include "common.cuh"

namespace sstc {

constexpr int BM = 128;          // rows per tile = TMEM lanes
constexpr int ROW_BYTES = 128;   // one K-chunk row = one 128-byte swizzle row (64 bf16 or 32 tf32)
constexpr int A_BYTES = BM * ROW_BYTES;
constexpr int STG_PITCH = 36;    // floats; 32 + 4 keeps both the float4 writes and reads conflict-free
constexpr int STG_WARP_BYTES = 32 * STG_PITCH * 4;  // one 32x32 fp32 staging tile per epilogue warp
constexpr int STG_BYTES = 4 * STG_WARP_BYTES;       // four epilogue warps (one per TMEM lane group)

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
  // "flat" tiling (3xTF32 TS kernel, stride-1 same-length layers with batch-contiguous x): the B*S rows of all
  // utterances are tiled as ONE sequence, so no 128-row tile is padded at every utterance end (S = 696 wastes
  // 9 % of each M = 128 MMA otherwise).  m_tiles then counts tiles of the flattened sequence; rows whose tap
  // falls outside their own utterance are zeroed by the operand-split warps, and the epilogue maps each
  // flattened row back to (utterance, frame).
  int32_t flat = 0, S_in = 0;
  // Ragged schedule: rows[b] = number of output rows of utterance b that are wanted (clamped to [0, S_out]); tiles past
  // that are neither computed nor written (the tile holding row rows[b]-1 is computed whole).  NULL: every row.
  // The CTA builds the per-utterance tile prefix in shared memory once (ragged_build) and every role maps its linear
  // tile index to (utterance, tile) with a binary search (ragged_decode); never combined with `flat`.
  const int32_t* rows = nullptr;
  // Split-bf16 pair output (conv1d_bf16x3.cu): when pair_c > 0, y_act is not a plain bf16 copy laid out like y but a
  // tensor of PAIR rows [.., 2*pair_c] = [hi(pair_c) | lo(pair_c)], hi = bf16(v), lo = bf16(v - hi), for the element
  // that y holds at flat index g = (row)*pair_c + channel (pair_c = channels per output row).
  int32_t pair_c = 0;
  int32_t l2_prefetch = 0;  // split-bf16 kernels: prefetch a tile's residual / running-mean rows into L2 (x3_common.cuh)
};
constexpr int RAGGED_MAX_B = 1024;  // utterances per launch in the ragged schedule (prefix table in shared memory)

// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// true in exactly one lane of a fully converged warp (elect.sync); ptxas knows a region guarded by it runs
// single-threaded, which `lane == 0` does not tell it
__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "elect.sync _|p, 0xffffffff;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t"
      "}"
      : "=r"(pred));
  return pred != 0;
}
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
__device__ __forceinline__ void prefetch_tmap(const CUtensorMap* m) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(m)) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
// generic-proxy shared-memory writes -> visible to the async proxy (tcgen05.mma operand reads)
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

__device__ __forceinline__ void tmem_alloc(uint32_t slot_smem, uint32_t cols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(slot_smem), "r"(cols) : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t cols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(cols) : "memory");
}

// K-major, SWIZZLE_128B shared-memory matrix descriptor (cute::UMMA::SmemDescriptor bit layout):
// start>>4 [0,14) | LBO>>4 [16,30) (unused for swizzled K-major, 1) | SBO>>4 [32,46) = 1024 B between
// 8-row groups | version=1 [46,48) | layout_type=2 (SWIZZLE_128B) [61,64)
__device__ __forceinline__ uint64_t make_desc(uint32_t smem_addr) {
  return (uint64_t)((smem_addr & 0x3FFFFu) >> 4) | (1ull << 16) | (64ull << 32) | (1ull << 46) | (2ull << 61);
}
// instruction descriptor (cute::UMMA::InstrDescriptor): c=f32 [4,6)=1, a_format [7,10), b_format [10,13)
// (1 = bf16, 2 = tf32), a/b K-major (bits 15,16 = 0), N>>3 [17,23), M>>4 [24,29)
__device__ __forceinline__ uint32_t make_idesc(uint32_t fmt, int n) {
  return (1u << 4) | (fmt << 7) | (fmt << 10) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);
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
__device__ __forceinline__ void umma_tf32(uint32_t tmem_d, uint64_t da, uint64_t db, uint32_t idesc, uint32_t accum) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t"
      "}" ::"r"(tmem_d),
      "l"(da), "l"(db), "r"(idesc), "r"(accum)
      : "memory");
}
// 2^x on the special-function unit (one MUFU.EX2; x <= 0 here, -inf -> 0, denormal results flush to 0)
__device__ __forceinline__ float ex2_approx(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
// Low-word forms for issue-bound kernels (many small MMAs from one thread): the descriptor's high word is
// the constant 0x40004040 (SBO = 64, version 1, SWIZZLE_128B) and the low word is start>>4 | 1<<16, so an
// operand that sits `off` bytes further on is just  lo + (off >> 4)  - one add instead of the shift / mask /
// or chain per MMA that make_desc() costs on the (serial, high-latency) uniform datapath.
__device__ __forceinline__ uint32_t desc_lo(uint32_t smem_addr) { return ((smem_addr & 0x3FFFFu) >> 4) | (1u << 16); }
__device__ __forceinline__ void umma_tf32_w(uint32_t tmem_d, uint32_t a_lo, uint32_t b_lo, uint32_t idesc, uint32_t accum) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      ".reg .b32 hi;\n\t"
      ".reg .b64 da, db;\n\t"
      "mov.u32 hi, 0x40004040;\n\t"
      "mov.b64 da, {%1, hi};\n\t"
      "mov.b64 db, {%2, hi};\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], da, db, %3, p;\n\t"
      "}" ::"r"(tmem_d),
      "r"(a_lo), "r"(b_lo), "r"(idesc), "r"(accum)
      : "memory");
}
__device__ __forceinline__ void umma_tf32_ts_w(uint32_t tmem_d, uint32_t tmem_a, uint32_t b_lo, uint32_t idesc, uint32_t accum) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      ".reg .b32 hi;\n\t"
      ".reg .b64 db;\n\t"
      "mov.u32 hi, 0x40004040;\n\t"
      "mov.b64 db, {%2, hi};\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], db, %3, p;\n\t"
      "}" ::"r"(tmem_d),
      "r"(tmem_a), "r"(b_lo), "r"(idesc), "r"(accum)
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

// two 32-column loads in flight, one wait
__device__ __forceinline__ void tmem_ld32x2(uint32_t ta, float (&a)[32], uint32_t tb, float (&b)[32]) {
  uint32_t r[32], q[32];
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
        "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
        "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(ta));
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(q[0]), "=r"(q[1]), "=r"(q[2]), "=r"(q[3]), "=r"(q[4]), "=r"(q[5]), "=r"(q[6]), "=r"(q[7]), "=r"(q[8]),
        "=r"(q[9]), "=r"(q[10]), "=r"(q[11]), "=r"(q[12]), "=r"(q[13]), "=r"(q[14]), "=r"(q[15]), "=r"(q[16]),
        "=r"(q[17]), "=r"(q[18]), "=r"(q[19]), "=r"(q[20]), "=r"(q[21]), "=r"(q[22]), "=r"(q[23]), "=r"(q[24]),
        "=r"(q[25]), "=r"(q[26]), "=r"(q[27]), "=r"(q[28]), "=r"(q[29]), "=r"(q[30]), "=r"(q[31])
      : "r"(tb));
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
  for (int i = 0; i < 32; ++i) {
    a[i] = __uint_as_float(r[i]);
    b[i] = __uint_as_float(q[i]);
  }
}

// ---------------------------------------------------------------------------------------------
// Ragged schedule helpers.  pre[0..B] (shared memory) = exclusive prefix of the utterances' scheduling units
// (unit_rows output rows each).  ragged_build is called by every thread of the CTA before the role split (warp 0
// does the work; the caller's __syncthreads / cluster sync publishes it).
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ void ragged_build(const int32_t* rows, int B, int S_out, int unit_rows, int* pre) {
  if (threadIdx.x >= 32) return;
  const int lane = threadIdx.x;
  int carry = 0;
  for (int base = 0; base < B; base += 32) {
    const int b = base + lane;
    int v = 0;
    if (b < B) v = (min(max(__ldg(rows + b), 0), S_out) + unit_rows - 1) / unit_rows;
    int incl = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int n = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += n;
    }
    if (b < B) pre[b] = carry + incl - v;
    carry += __shfl_sync(0xffffffffu, incl, 31);
  }
  if (lane == 0) pre[B] = carry;
}
// u < pre[B]: the utterance holding unit u (the LAST b with pre[b] <= u, which skips empty utterances) and the
// unit's index inside it
__device__ __forceinline__ void ragged_decode(const int* pre, int B, int u, int& b, int& mt) {
  int lo = 0, hi = B;
  while (hi - lo > 1) {
    const int mid = (lo + hi) >> 1;
    if (pre[mid] <= u) lo = mid; else hi = mid;
  }
  b = lo;
  mt = u - pre[lo];
}

// ---------------------------------------------------------------------------------------------
// Pieces every 3xTF32 kernel (convolutions, attention) uses for its operand split
// ---------------------------------------------------------------------------------------------
// nearest TF32 value (10 explicit mantissa bits), ties away from zero in magnitude
__device__ __forceinline__ float tf32_rn(float x) {
  return __uint_as_float((__float_as_uint(x) + 0x1000u) & 0xFFFFE000u);
}

__device__ __forceinline__ void tmem_st32(uint32_t taddr, const float (&v)[32]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
      "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, "
      "%17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};" ::"r"(taddr),
      "r"(__float_as_uint(v[0])), "r"(__float_as_uint(v[1])), "r"(__float_as_uint(v[2])), "r"(__float_as_uint(v[3])),
      "r"(__float_as_uint(v[4])), "r"(__float_as_uint(v[5])), "r"(__float_as_uint(v[6])), "r"(__float_as_uint(v[7])),
      "r"(__float_as_uint(v[8])), "r"(__float_as_uint(v[9])), "r"(__float_as_uint(v[10])), "r"(__float_as_uint(v[11])),
      "r"(__float_as_uint(v[12])), "r"(__float_as_uint(v[13])), "r"(__float_as_uint(v[14])), "r"(__float_as_uint(v[15])),
      "r"(__float_as_uint(v[16])), "r"(__float_as_uint(v[17])), "r"(__float_as_uint(v[18])), "r"(__float_as_uint(v[19])),
      "r"(__float_as_uint(v[20])), "r"(__float_as_uint(v[21])), "r"(__float_as_uint(v[22])), "r"(__float_as_uint(v[23])),
      "r"(__float_as_uint(v[24])), "r"(__float_as_uint(v[25])), "r"(__float_as_uint(v[26])), "r"(__float_as_uint(v[27])),
      "r"(__float_as_uint(v[28])), "r"(__float_as_uint(v[29])), "r"(__float_as_uint(v[30])), "r"(__float_as_uint(v[31]))
      : "memory");
}

__device__ __forceinline__ void tmem_st16(uint32_t taddr, const float (&v)[16]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], "
      "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};" ::"r"(taddr),
      "r"(__float_as_uint(v[0])), "r"(__float_as_uint(v[1])), "r"(__float_as_uint(v[2])), "r"(__float_as_uint(v[3])),
      "r"(__float_as_uint(v[4])), "r"(__float_as_uint(v[5])), "r"(__float_as_uint(v[6])), "r"(__float_as_uint(v[7])),
      "r"(__float_as_uint(v[8])), "r"(__float_as_uint(v[9])), "r"(__float_as_uint(v[10])), "r"(__float_as_uint(v[11])),
      "r"(__float_as_uint(v[12])), "r"(__float_as_uint(v[13])), "r"(__float_as_uint(v[14])), "r"(__float_as_uint(v[15]))
      : "memory");
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
// Drain one 128 x BN fp32 accumulator tile.  Called by the four epilogue warps; warp `lg` owns TMEM
// lanes [32*lg, 32*lg+32).  tcgen05.ld hands each thread one ROW (32 consecutive columns); a global
// store from that layout would touch 32 different cache lines per instruction, so each 32x32 block
// goes through a per-warp shared-memory staging tile and is re-read with lanes running along the
// row: every global access below covers 4 rows x 128 contiguous bytes.
// All fused element-wise work (bias, activation, residual, row mask, running mean, activated bf16
// copy) happens in that coalesced domain, where a lane's four columns are the same for all rows.
// ---------------------------------------------------------------------------------------------
// Activation on four values with ONE dispatch (the epilogue is instruction-cache sensitive: it runs on
// four warps, one per scheduler, so its body must stay small - tanhf is kept out of line).
static __device__ __noinline__ float4 tanh4(float4 f) { return make_float4(tanhf(f.x), tanhf(f.y), tanhf(f.z), tanhf(f.w)); }
__device__ __forceinline__ float4 act4(float4 f, int act, float slope) {
  if (act == SS_ACT_NONE) return f;
  if (act == SS_ACT_TANH) return tanh4(f);
  const float neg = act == SS_ACT_RELU ? 0.f : slope;  // relu is leaky-relu with slope 0
  f.x = f.x > 0.f ? f.x : f.x * neg;
  f.y = f.y > 0.f ? f.y : f.y * neg;
  f.z = f.z > 0.f ? f.z : f.z * neg;
  f.w = f.w > 0.f ? f.w : f.w * neg;
  return f;
}

// The activated copy of four consecutive channels whose y index is g: a plain bf16 copy, or (pair_c > 0) the hi / lo
// halves of a split-bf16 pair row.  co = channel of the first of the four inside its row (only used for pairs).
__device__ __forceinline__ void store_act4(const TcArgs& p, int64_t g, int co, float4 f) {
  if (p.pair_c == 0) {
    *reinterpret_cast<uint2*>(p.y_act + g) = make_uint2(pack_bf16(f.x, f.y), pack_bf16(f.z, f.w));
    return;
  }
  const uint2 hi = make_uint2(pack_bf16(f.x, f.y), pack_bf16(f.z, f.w));
  float a, b, c, d;
  unpack_bf16(hi.x, a, b);
  unpack_bf16(hi.y, c, d);
  __nv_bfloat16* dst = p.y_act + 2 * (g - co) + co;
  *reinterpret_cast<uint2*>(dst) = hi;
  *reinterpret_cast<uint2*>(dst + p.pair_c) = make_uint2(pack_bf16(f.x - a, f.y - b), pack_bf16(f.z - c, f.w - d));
}

// One 32-column block: v[] holds this thread's ROW (TMEM lane) of the block; n_blk = first column.
__device__ __forceinline__ void drain_block(const TcArgs& p, const float (&v)[32], float* stg, int lg, int lane, int b,
                                            int mt, int n_blk, int len) {
  const int col = (lane & 7) * 4;
  const int rsub = lane >> 3;
  float* my = stg + lane * STG_PITCH;
  __syncwarp();  // the previous block has been fully read back
#pragma unroll
  for (int q = 0; q < 8; ++q)
    *reinterpret_cast<float4*>(my + q * 4) = make_float4(v[q * 4], v[q * 4 + 1], v[q * 4 + 2], v[q * 4 + 3]);
  __syncwarp();
  const int n = n_blk + col;
  if (n >= p.Cout) return;
  float4 bias4 = make_float4(0.f, 0.f, 0.f, 0.f);
  if (p.bias) bias4 = __ldg(reinterpret_cast<const float4*>(p.bias + n));
  const int pair_co = p.pair_c > 0 ? n % p.pair_c : 0;
  const int t_first = mt * BM + lg * 32 + rsub;
  const int64_t ybase = (int64_t)b * p.y_batch_stride;
  const float* src = stg + rsub * STG_PITCH + col;
  const bool pre_res = p.residual != nullptr && p.residual_dtype == SS_F32;
  const bool pre_acc = p.accum != nullptr && !p.accum_init;
  if (!pre_res && !pre_acc) {
    // No global reads to wait for: a compact rolled loop (the epilogue runs on four warps, one per
    // scheduler, and is instruction-cache sensitive - keep the common case small).
#pragma unroll 1
    for (int it = 0; it < 8; ++it) {
      const int t = t_first + it * 4;
      const int64_t off = p.y_offset + (int64_t)t * p.y_row_stride + n;
      if (t >= p.S_out || off < 0 || off >= p.y_batch_elems) continue;
      float4 f = *reinterpret_cast<const float4*>(src + it * 4 * STG_PITCH);
      f.x += bias4.x; f.y += bias4.y; f.z += bias4.z; f.w += bias4.w;
      f = act4(f, p.act, p.slope);
      const int64_t g = ybase + off;
      if (p.residual) {  // bf16 residual
        const uint2 r = *reinterpret_cast<const uint2*>(static_cast<const __nv_bfloat16*>(p.residual) + g);
        float a, c;
        unpack_bf16(r.x, a, c); f.x += a; f.y += c;
        unpack_bf16(r.y, a, c); f.z += a; f.w += c;
      }
      if (t >= len) f = make_float4(0.f, 0.f, 0.f, 0.f);
      if (p.y) {
        if (p.y_dtype == SS_F32)
          *reinterpret_cast<float4*>(static_cast<float*>(p.y) + g) = f;
        else
          *reinterpret_cast<uint2*>(static_cast<__nv_bfloat16*>(p.y) + g) = make_uint2(pack_bf16(f.x, f.y), pack_bf16(f.z, f.w));
      }
      if (p.accum) {  // accum_init: plain store of the scaled value
        f.x *= p.accum_scale; f.y *= p.accum_scale; f.z *= p.accum_scale; f.w *= p.accum_scale;
        *reinterpret_cast<float4*>(p.accum + g) = f;
      }
      if (p.y_act) {
        f = act4(f, p.act2, p.slope2);
        store_act4(p, g, pair_co, f);
      }
    }
    return;
  }
  // Two half-loops of four rows: the global reads of a half (residual, running mean) are all issued
  // before any of them is consumed, so four loads are in flight per lane instead of one.
#pragma unroll 1
  for (int half = 0; half < 2; ++half) {
    float4 rpre[4], apre[4];
    int64_t g[4];
    bool ok[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int t = t_first + (half * 4 + u) * 4;
      const int64_t off = p.y_offset + (int64_t)t * p.y_row_stride + n;
      ok[u] = t < p.S_out && off >= 0 && off < p.y_batch_elems;
      g[u] = ybase + off;
      rpre[u] = make_float4(0.f, 0.f, 0.f, 0.f);
      apre[u] = rpre[u];
      if (ok[u] && pre_res) rpre[u] = *reinterpret_cast<const float4*>(static_cast<const float*>(p.residual) + g[u]);  // plain load: may alias y
      if (ok[u] && pre_acc) apre[u] = *reinterpret_cast<const float4*>(p.accum + g[u]);
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      if (!ok[u]) continue;
      const int t = t_first + (half * 4 + u) * 4;
      float4 f = *reinterpret_cast<const float4*>(src + (half * 4 + u) * 4 * STG_PITCH);
      f.x += bias4.x; f.y += bias4.y; f.z += bias4.z; f.w += bias4.w;
      f = act4(f, p.act, p.slope);
      if (pre_res) {
        f.x += rpre[u].x; f.y += rpre[u].y; f.z += rpre[u].z; f.w += rpre[u].w;
      } else if (p.residual) {
        const uint2 r = *reinterpret_cast<const uint2*>(static_cast<const __nv_bfloat16*>(p.residual) + g[u]);
        float a, c;
        unpack_bf16(r.x, a, c); f.x += a; f.y += c;
        unpack_bf16(r.y, a, c); f.z += a; f.w += c;
      }
      if (t >= len) f = make_float4(0.f, 0.f, 0.f, 0.f);
      if (p.y) {
        if (p.y_dtype == SS_F32)
          *reinterpret_cast<float4*>(static_cast<float*>(p.y) + g[u]) = f;
        else
          *reinterpret_cast<uint2*>(static_cast<__nv_bfloat16*>(p.y) + g[u]) =
              make_uint2(pack_bf16(f.x, f.y), pack_bf16(f.z, f.w));
      }
      if (p.accum) {  // running mean; the activated copy below then follows the accumulated value
        f.x = fmaf(p.accum_scale, f.x, apre[u].x);
        f.y = fmaf(p.accum_scale, f.y, apre[u].y);
        f.z = fmaf(p.accum_scale, f.z, apre[u].z);
        f.w = fmaf(p.accum_scale, f.w, apre[u].w);
        *reinterpret_cast<float4*>(p.accum + g[u]) = f;
      }
      if (p.y_act) {
        f = act4(f, p.act2, p.slope2);
        store_act4(p, g[u], pair_co, f);
      }
    }
  }
}

// Lean variant for the fp32 kernels' common case (fp32 y and / or running mean, optional bias / fp32 residual / row
// mask, activation none / relu / leaky-relu; no bf16 copy, no tanh).  The 3xTF32 kernels keep the whole output
// row in registers, so their column-block loop must be unrolled and this body is instantiated BN/32 times:
// it has to be small (one warp per scheduler runs it - the general drain_block above, four times over, spills
// out of the instruction cache and took ~4.5k cycles per block) and its eight row steps are straight-line
// with all residual loads in flight before the first is used.
// WITH_ACCUM: the running-mean form is compiled in only for the narrow tiles (BN <= 64, the HiFi-GAN stages that use
// it); at BN = 128 the epilogue already holds a 128-float row and the extra live values would spill.
template <bool WITH_ACCUM>
__device__ __forceinline__ bool drain_is_lean(const TcArgs& p) {
  return (p.y != nullptr || (WITH_ACCUM && p.accum != nullptr)) && (WITH_ACCUM || p.accum == nullptr) &&
         p.y_dtype == SS_F32 && p.y_act == nullptr && p.act != SS_ACT_TANH &&
         (p.residual == nullptr || p.residual_dtype == SS_F32);
}
// (b, t_tile0) = utterance and frame of the tile's first row; straddle = the tile runs past the end of utterance b
// (flat tiling only), in which case rows are un-flattened one by one and the row mask is looked up per row.
template <bool WITH_ACCUM>
__device__ __forceinline__ void drain_block_lean(const TcArgs& p, const float (&v)[32], float* stg, int lg, int lane,
                                                 int b, int t_tile0, bool straddle, int n_blk, int len) {
  const int col = (lane & 7) * 4;
  const int rsub = lane >> 3;
  float* my = stg + lane * STG_PITCH;
  __syncwarp();  // the previous block has been fully read back
#pragma unroll
  for (int q = 0; q < 8; ++q)
    *reinterpret_cast<float4*>(my + q * 4) = make_float4(v[q * 4], v[q * 4 + 1], v[q * 4 + 2], v[q * 4 + 3]);
  __syncwarp();
  const int n = n_blk + col;
  if (n >= p.Cout) return;
  float4 bias4 = make_float4(0.f, 0.f, 0.f, 0.f);
  if (p.bias) bias4 = __ldg(reinterpret_cast<const float4*>(p.bias + n));
  const float neg = p.act == SS_ACT_NONE ? 1.f : p.act == SS_ACT_RELU ? 0.f : p.slope;  // x > 0 ? x : x * neg
  const float* src = stg + rsub * STG_PITCH + col;
  const float* res = static_cast<const float*>(p.residual);
  float* y = static_cast<float*>(p.y);
  int t = t_tile0 + lg * 32 + rsub, bb = b;  // first row of this lane
  // offsets are carried along incrementally (64-bit multiplies per row showed up in the profile of this loop)
  const int64_t row4 = 4 * p.y_row_stride;
  int64_t off_rel = p.y_offset + (int64_t)t * p.y_row_stride + n;  // inside the utterance's output window
  int64_t base = (int64_t)bb * p.y_batch_stride;
  // two half-loops of four rows (register budget: the caller holds a whole output row); the residual loads of a
  // half are all in flight before the first is used
#pragma unroll 1
  for (int half = 0; half < 2; ++half) {
    int64_t off[4];
    bool ok[4], keep[4];
    float4 r[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      if (straddle)
        while (t >= p.S_out) {
          t -= p.S_out;
          ++bb;
          off_rel -= (int64_t)p.S_out * p.y_row_stride;
          base += p.y_batch_stride;
        }
      ok[u] = bb < p.B && t < p.S_out && off_rel >= 0 && off_rel < p.y_batch_elems;
      off[u] = base + off_rel;
      keep[u] = t < (straddle ? (p.lens && ok[u] ? __ldg(p.lens + bb) : 0x7fffffff) : len);
      r[u] = make_float4(0.f, 0.f, 0.f, 0.f);
      if (res && ok[u]) r[u] = *reinterpret_cast<const float4*>(res + off[u]);  // plain load: the residual may be the buffer this launch writes (out == residual)
      t += 4;
      off_rel += row4;
    }
    if (WITH_ACCUM && p.accum) {  // running mean (HiFi-GAN MRF): accum = scale*v (first branch) or accum += scale*v
      const bool rd = !p.accum_init;
      float4 a[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        a[u] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (rd && ok[u]) a[u] = *reinterpret_cast<const float4*>(p.accum + off[u]);
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        float4 f = *reinterpret_cast<const float4*>(src + (half * 4 + u) * 4 * STG_PITCH);
        f.x += bias4.x; f.y += bias4.y; f.z += bias4.z; f.w += bias4.w;
        f.x = f.x > 0.f ? f.x : f.x * neg;
        f.y = f.y > 0.f ? f.y : f.y * neg;
        f.z = f.z > 0.f ? f.z : f.z * neg;
        f.w = f.w > 0.f ? f.w : f.w * neg;
        f.x += r[u].x; f.y += r[u].y; f.z += r[u].z; f.w += r[u].w;
        if (!keep[u]) f = make_float4(0.f, 0.f, 0.f, 0.f);
        if (!ok[u]) continue;
        if (y) *reinterpret_cast<float4*>(y + off[u]) = f;
        f.x = fmaf(p.accum_scale, f.x, a[u].x);
        f.y = fmaf(p.accum_scale, f.y, a[u].y);
        f.z = fmaf(p.accum_scale, f.z, a[u].z);
        f.w = fmaf(p.accum_scale, f.w, a[u].w);
        *reinterpret_cast<float4*>(p.accum + off[u]) = f;
      }
      continue;
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      float4 f = *reinterpret_cast<const float4*>(src + (half * 4 + u) * 4 * STG_PITCH);
      f.x += bias4.x; f.y += bias4.y; f.z += bias4.z; f.w += bias4.w;
      f.x = f.x > 0.f ? f.x : f.x * neg;
      f.y = f.y > 0.f ? f.y : f.y * neg;
      f.z = f.z > 0.f ? f.z : f.z * neg;
      f.w = f.w > 0.f ? f.w : f.w * neg;
      f.x += r[u].x; f.y += r[u].y; f.z += r[u].z; f.w += r[u].w;
      if (!keep[u]) f = make_float4(0.f, 0.f, 0.f, 0.f);
      if (ok[u]) *reinterpret_cast<float4*>(y + off[u]) = f;
    }
  }
}

// Whole-row form of the lean drain for tiles whose row fits a per-warp staging tile (BN = 96): the thread's
// BN-float row goes to shared memory in one go - the accumulator registers are dead afterwards - and the row
// bookkeeping (utterance, frame, mask, offset) is done once per row quad for all column blocks instead of once per
// block; two warp syncs per tile instead of two per block.  With cold operands this epilogue, not the main loop,
// paces the short-reduction layers.
template <int NB>  // column blocks of 32
__device__ __forceinline__ void drain_tile_lean_wide(const TcArgs& p, const float (&acc)[NB * 32], float* stg, int lg,
                                                     int lane, int b, int t_tile0, bool straddle, int n0, int len) {
  constexpr int PW = NB * 32 + 4;  // pitch in floats: +4 keeps the float4 writes and reads conflict-free
  const int col = (lane & 7) * 4;
  const int rsub = lane >> 3;
  float* my = stg + lane * PW;
  __syncwarp();  // the previous tile has been fully read back
#pragma unroll
  for (int q = 0; q < NB * 8; ++q)
    *reinterpret_cast<float4*>(my + q * 4) = make_float4(acc[q * 4], acc[q * 4 + 1], acc[q * 4 + 2], acc[q * 4 + 3]);
  __syncwarp();
  float4 bias4[NB];
  bool colok[NB];
#pragma unroll
  for (int k = 0; k < NB; ++k) {
    const int n = n0 + k * 32 + col;
    colok[k] = n < p.Cout;
    bias4[k] = make_float4(0.f, 0.f, 0.f, 0.f);
    if (p.bias && colok[k]) bias4[k] = __ldg(reinterpret_cast<const float4*>(p.bias + n));
  }
  const float neg = p.act == SS_ACT_NONE ? 1.f : p.act == SS_ACT_RELU ? 0.f : p.slope;  // x > 0 ? x : x * neg
  const float* src = stg + rsub * PW + col;
  const float* res = static_cast<const float*>(p.residual);
  float* y = static_cast<float*>(p.y);
  int t = t_tile0 + lg * 32 + rsub, bb = b;  // first row of this lane
  const int64_t row4 = 4 * p.y_row_stride;
  int64_t off_rel = p.y_offset + (int64_t)t * p.y_row_stride + n0 + col;
  int64_t base = (int64_t)bb * p.y_batch_stride;
#pragma unroll 1
  for (int half = 0; half < 2; ++half) {
    int64_t off[4];
    bool ok[4], keep[4];
    float4 r[4][NB];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      if (straddle)
        while (t >= p.S_out) {
          t -= p.S_out;
          ++bb;
          off_rel -= (int64_t)p.S_out * p.y_row_stride;
          base += p.y_batch_stride;
        }
      ok[u] = bb < p.B && t < p.S_out && off_rel >= 0 && off_rel < p.y_batch_elems;
      off[u] = base + off_rel;
      keep[u] = t < (straddle ? (p.lens && ok[u] ? __ldg(p.lens + bb) : 0x7fffffff) : len);
#pragma unroll
      for (int k = 0; k < NB; ++k) {
        r[u][k] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (res && ok[u] && colok[k]) r[u][k] = *reinterpret_cast<const float4*>(res + off[u] + k * 32);
      }
      t += 4;
      off_rel += row4;
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
#pragma unroll
      for (int k = 0; k < NB; ++k) {
        float4 f = *reinterpret_cast<const float4*>(src + (half * 4 + u) * 4 * PW + k * 32);
        f.x += bias4[k].x; f.y += bias4[k].y; f.z += bias4[k].z; f.w += bias4[k].w;
        f.x = f.x > 0.f ? f.x : f.x * neg;
        f.y = f.y > 0.f ? f.y : f.y * neg;
        f.z = f.z > 0.f ? f.z : f.z * neg;
        f.w = f.w > 0.f ? f.w : f.w * neg;
        f.x += r[u][k].x; f.y += r[u][k].y; f.z += r[u][k].z; f.w += r[u][k].w;
        if (!keep[u]) f = make_float4(0.f, 0.f, 0.f, 0.f);
        if (ok[u] && colok[k]) *reinterpret_cast<float4*>(y + off[u] + k * 32) = f;
      }
    }
  }
}

// G epilogue warps share one TMEM lane group; warp `sub` of the group takes every G-th 32-column block.
template <int BN, int G = 1>
__device__ __forceinline__ void drain_tile(const TcArgs& p, uint32_t taddr, float* stg, int lg, int lane, int b, int mt,
                                           int n0, int sub = 0) {
  const int len = p.lens ? p.lens[b] : 0x7fffffff;
#pragma unroll 1
  for (int c0 = sub * 32; c0 < BN; c0 += 32 * G) {
    float v[32];
    __syncwarp();  // reconverge: tcgen05.ld is .sync.aligned
    tmem_ld32(taddr + c0, v);
    drain_block(p, v, stg, lg, lane, b, mt, n0 + c0, len);
  }
}

// ---------------------------------------------------------------------------------------------
// host helpers
// ---------------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

inline EncodeTiledFn encode_fn() {
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

// 3-D tensor map {d0 (contiguous), d1, d2} with a {box0, box1, 1} box and the 128-byte swizzle.
inline int make_map_3d(CUtensorMap* m, CUtensorMapDataType dt, const void* base, uint64_t d0, uint64_t d1, uint64_t d2,
                       uint64_t s1_bytes, uint64_t s2_bytes, uint32_t box0, uint32_t box1,
                       CUtensorMapSwizzle swizzle = CU_TENSOR_MAP_SWIZZLE_128B) {
  EncodeTiledFn fn = encode_fn();
  SS_REQUIRE(fn != nullptr, SS_ERR_CUDA, "tensor-core conv: cuTensorMapEncodeTiled is not available from this driver");
  cuuint64_t dims[3] = {d0, d1, d2};
  cuuint64_t strides[2] = {s1_bytes, s2_bytes};
  cuuint32_t box[3] = {box0, box1, 1};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = fn(m, dt, 3, const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                  swizzle, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  SS_REQUIRE(r == CUDA_SUCCESS, SS_ERR_CUDA, "tensor-core conv: cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
  return SS_OK;
}

// Per-device, thread-safe bookkeeping of kernel attributes (the dynamic shared-memory opt-in is per device and per
// kernel).  needs_attr_on_device() is true until attr_done_on_device() has been called for the current device; the bit
// is published only AFTER cudaFuncSetAttribute succeeded, so a second host thread can at worst repeat the (idempotent)
// call - it can never launch before the opt-in has landed.
inline bool needs_attr_on_device(const std::atomic<unsigned long long>& mask) {
  int d = 0;
  if (cudaGetDevice(&d) != cudaSuccess) return true;
  return (mask.load(std::memory_order_acquire) & (1ull << (d & 63))) == 0;
}
inline void attr_done_on_device(std::atomic<unsigned long long>& mask) {
  int d = 0;
  if (cudaGetDevice(&d) != cudaSuccess) return;
  mask.fetch_or(1ull << (d & 63), std::memory_order_release);
}

// SM count of the CURRENT device (cached per device ordinal)
inline int sm_count() {
  static std::atomic<int> cache[64];
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 0;
  const int slot = dev & 63;
  int v = cache[slot].load(std::memory_order_relaxed);
  if (v > 0) return v;
  if (cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) return 0;
  cache[slot].store(v, std::memory_order_relaxed);
  return v;
}

}  // namespace sstc
