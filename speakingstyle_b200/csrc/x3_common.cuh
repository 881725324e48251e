// Pieces shared by the split-bf16 kernels (conv1d_bf16x3.cu, resunit_bf16x3.cu): the low-word-descriptor bf16 MMA and
// the dense epilogue.
#pragma once
#include <cstdlib>

#include "tc_common.cuh"

namespace sstc {

// Programmatic dependent launch: the split-bf16 kernels are launched with programmaticStreamSerialization, announce
// `launch_dependents` at their first instruction and do their prologue (barrier init, TMEM allocation, tensor-map
// prefetch) before `griddepcontrol.wait`; nothing that reads or writes activation memory precedes the wait.  A CTA of the
// next launch therefore takes over an SM the moment this launch's CTA leaves it, instead of after the whole grid has
// drained and the next one has been scheduled (~300 convolution launches per cfg3 step).  SS_PDL=0 turns it off.
inline int x3_pdl_enabled() {
  static const int v = [] {
    const char* e = getenv("SS_PDL");
    return (e && e[0] == '0') ? 0 : 1;
  }();
  return v;
}
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kern)(KArgs...), int grid, int threads, int smem, cudaStream_t st, Args&&... args) {
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3((unsigned)grid);
  cfg.blockDim = dim3((unsigned)threads);
  cfg.dynamicSmemBytes = (size_t)smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = x3_pdl_enabled();
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, kern, static_cast<KArgs>(args)...);
}

// debug switch for A/B runs: SS_X3_PREFETCH=0 turns the epilogue's L2 prefetch off
inline int x3_prefetch_default() {
  static const int v = [] {
    const char* e = getenv("SS_X3_PREFETCH");
    return (e && e[0] == '0') ? 0 : 1;
  }();
  return v;
}

__device__ __forceinline__ void umma_bf16_w(uint32_t tmem_d, uint32_t a_lo, uint32_t b_lo, uint32_t idesc, uint32_t accum) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      ".reg .b32 hi;\n\t"
      ".reg .b64 da, db;\n\t"
      "mov.u32 hi, 0x40004040;\n\t"
      "mov.b64 da, {%1, hi};\n\t"
      "mov.b64 db, {%2, hi};\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], da, db, %3, p;\n\t"
      "}" ::"r"(tmem_d),
      "r"(a_lo), "r"(b_lo), "r"(idesc), "r"(accum)
      : "memory");
}

// Lean drain of one 32 x 32 block for the dense case HiFi-GAN's ResBlocks are (no conv activation, fp32 y / residual,
// rows of Cout contiguous channels, no row mask): v = acc + bias (+ residual); y = v; running mean; pair row of
// act2(v).  Same staging transposition as drain_block (tc_common.cuh), but one flat index per lane, advanced by a
// constant, and the residual / running-mean reads of four rows in flight at a time.
__device__ __forceinline__ bool drain_is_dense(const TcArgs& p) {
  return p.act == SS_ACT_NONE && p.y_dtype == SS_F32 && p.residual_dtype == SS_F32 && p.lens == nullptr &&
         p.y_offset == 0 && p.y_row_stride == p.Cout && p.y_batch_stride == (int64_t)p.S_out * p.Cout &&
         p.y_batch_elems == (int64_t)p.S_out * p.Cout && (p.y_act == nullptr || p.pair_c == p.Cout || p.pair_c == 0);
}
// Pull the residual / running-mean rows a tile's epilogue will read into L2 ahead of time (called before the warp waits
// for the tile's accumulator): drain_block_dense keeps only four 16-byte loads per lane in flight, which from HBM
// (~1-2 us under load) is too few to keep a bandwidth-bound launch fed.  Dense layout only.
__device__ __forceinline__ void prefetch_epilogue_rows(const TcArgs& p, int lg, int lane, int b, int row0, int nrows,
                                                       int n0, int ncols) {
  const bool pre_acc = p.accum != nullptr && !p.accum_init;
  const int r = lg * 32 + lane;
  if (!p.l2_prefetch || (p.residual == nullptr && !pre_acc) || r >= nrows) return;
  const int64_t g = ((int64_t)b * p.S_out + row0 + r) * p.Cout + n0;
  for (int c = 0; c < ncols; c += 32) {  // one 128-byte line per step
    if (p.residual) asm volatile("prefetch.global.L2 [%0];" ::"l"(static_cast<const float*>(p.residual) + g + c));
    if (pre_acc) asm volatile("prefetch.global.L2 [%0];" ::"l"(p.accum + g + c));
  }
}

// The block holds rows row0 + [32*lg, 32*lg + 32) of utterance b; only the tile's first `nrows` rows are stored.
__device__ __forceinline__ void drain_block_dense(const TcArgs& p, const float (&v)[32], float* stg, int lg, int lane, int b,
                                                  int row0, int nrows, int n_blk) {
  const int col = (lane & 7) * 4;
  const int rsub = lane >> 3;
  float* my = stg + lane * STG_PITCH;
  __syncwarp();  // the previous block has been fully read back
#pragma unroll
  for (int q = 0; q < 8; ++q)
    *reinterpret_cast<float4*>(my + q * 4) = make_float4(v[q * 4], v[q * 4 + 1], v[q * 4 + 2], v[q * 4 + 3]);
  __syncwarp();
  const int n = n_blk + col;
  const float4 bias4 = p.bias ? __ldg(reinterpret_cast<const float4*>(p.bias + n)) : make_float4(0.f, 0.f, 0.f, 0.f);
  const int r_first = lg * 32 + rsub;  // row inside the tile
  const int64_t g0 = ((int64_t)b * p.S_out + row0 + r_first) * p.Cout + n;
  const int64_t gstep = 4 * (int64_t)p.Cout;  // four rows per step
  const float* src = stg + rsub * STG_PITCH + col;
  const float* res = p.residual ? static_cast<const float*>(p.residual) + g0 : nullptr;
  const bool pre_acc = p.accum != nullptr && !p.accum_init;
  float* const y = p.y ? static_cast<float*>(p.y) + g0 : nullptr;
  float* const acc = p.accum ? p.accum + g0 : nullptr;
  const bool plain = p.pair_c == 0;  // y_act is a plain bf16 copy laid out like y, not a pair row
  __nv_bfloat16* const pair = p.y_act ? (plain ? p.y_act + g0 : p.y_act + 2 * g0 - n) : nullptr;
#pragma unroll
  for (int half = 0; half < 2; ++half) {
    float4 rpre[4], apre[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int i = half * 4 + u;
      const bool ok = r_first + 4 * i < nrows;
      rpre[u] = make_float4(0.f, 0.f, 0.f, 0.f);
      apre[u] = rpre[u];
      if (res && ok) rpre[u] = *reinterpret_cast<const float4*>(res + i * gstep);  // plain load: may alias y
      if (pre_acc && ok) apre[u] = *reinterpret_cast<const float4*>(acc + i * gstep);
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int i = half * 4 + u;
      if (r_first + 4 * i >= nrows) continue;
      float4 f = *reinterpret_cast<const float4*>(src + i * 4 * STG_PITCH);
      f.x = (f.x + bias4.x) + rpre[u].x; f.y = (f.y + bias4.y) + rpre[u].y;
      f.z = (f.z + bias4.z) + rpre[u].z; f.w = (f.w + bias4.w) + rpre[u].w;
      if (y) *reinterpret_cast<float4*>(y + i * gstep) = f;
      if (acc) {  // running mean; the pair copy below then follows the accumulated value
        f.x = fmaf(p.accum_scale, f.x, apre[u].x);
        f.y = fmaf(p.accum_scale, f.y, apre[u].y);
        f.z = fmaf(p.accum_scale, f.z, apre[u].z);
        f.w = fmaf(p.accum_scale, f.w, apre[u].w);
        *reinterpret_cast<float4*>(acc + i * gstep) = f;
      }
      if (pair) {
        f = act4(f, p.act2, p.slope2);
        const uint2 hi = make_uint2(pack_bf16(f.x, f.y), pack_bf16(f.z, f.w));
        float a0, a1, a2, a3;
        unpack_bf16(hi.x, a0, a1);
        unpack_bf16(hi.y, a2, a3);
        if (plain) {
          *reinterpret_cast<uint2*>(pair + i * gstep) = hi;
        } else {
          __nv_bfloat16* dst = pair + 2 * i * gstep;
          *reinterpret_cast<uint2*>(dst) = hi;
          *reinterpret_cast<uint2*>(dst + p.Cout) = make_uint2(pack_bf16(f.x - a0, f.y - a1), pack_bf16(f.z - a2, f.w - a3));
        }
      }
    }
  }
}


}  // namespace sstc
