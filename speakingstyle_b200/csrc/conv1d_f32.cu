// fp32 implicit-GEMM 1-D convolution over channels-last activations (FFMA pipe).
//
// This is the accuracy-class-of-the-reference kernel: everything that feeds an integer decision
// (durations, pitch/energy buckets) runs through it, as does the fp32 precision mode of the rest
// of the model.  One CTA computes a 128 (positions) x 128 (output channels) tile; the K loop walks
// taps x input-channel chunks of 16, double-buffered through shared memory with the next tile's
// global loads in flight during the FMAs.  No im2col buffer exists anywhere: a tap is just a row
// offset into the same activation tensor, and rows outside [0, S_in) read as zero.
//
// Roofline: compute-bound on the FP32 pipe (arithmetic intensity ~ 2*128*128*16 / (2*128*16*4 B)
// = 32 FLOP/B from L2 per tile step; weights and the k-fold re-read of activations stay in L2).
#include "common.cuh"

namespace {

constexpr int BM = 128, BN = 128, BK = 16, NT = 256;

struct Frag {
  float4 a[2];
  float4 b[2];
};

__device__ __forceinline__ float lrelu(float v, float s) { return v > 0.0f ? v : v * s; }

__global__ void __launch_bounds__(NT, 2) conv1d_f32_kernel(const ss_conv1d_f32_params p) {
  __shared__ __align__(16) float As[2][BK][BM + 4];  // +4: transposed stores become 2-way instead of 4-way conflicts
  __shared__ __align__(16) float Bs[2][BK][BN];

  const int tid = threadIdx.x;
  const int tx = tid & 15, ty = tid >> 4;
  const int t0 = blockIdx.x * BM;
  const int n0 = blockIdx.y * BN;
  const int b = blockIdx.z;

  const float* __restrict__ xb = p.x + (int64_t)b * p.x_batch_stride;
  const float* __restrict__ w = p.w;
  const int nchunk = (p.Cin + BK - 1) / BK;
  const int KT = p.k * nchunk;

  // loader coordinates
  const int a_row = tid >> 2;        // 0..63 (+64)
  const int a_q = (tid & 3) << 2;    // 0,4,8,12
  const int b_k = tid >> 5;          // 0..7 (+8)
  const int b_n = (tid & 31) << 2;   // 0..124

  float acc[8][8];
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[i][j] = 0.0f;

  Frag st;
  auto load_global = [&](int kt) {
    const int j = kt / nchunk;
    const int ci0 = (kt - j * nchunk) * BK;
    const int shift = j * p.dilation - p.pad;
#pragma unroll
    for (int i = 0; i < 2; ++i) {
      const int row = t0 + a_row + i * 64 + shift;
      const int ci = ci0 + a_q;
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      if (row >= 0 && row < p.S_in && ci < p.Cin)
        v = __ldg(reinterpret_cast<const float4*>(xb + (int64_t)row * p.x_row_stride + ci));
      if (p.in_act == SS_ACT_LRELU) {
        v.x = lrelu(v.x, p.in_slope);
        v.y = lrelu(v.y, p.in_slope);
        v.z = lrelu(v.z, p.in_slope);
        v.w = lrelu(v.w, p.in_slope);
      }
      st.a[i] = v;
    }
#pragma unroll
    for (int i = 0; i < 2; ++i) {
      const int ci = ci0 + b_k + i * 8;
      const int n = n0 + b_n;
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      if (ci < p.Cin && n < p.Cout)
        v = __ldg(reinterpret_cast<const float4*>(w + ((int64_t)j * p.Cin + ci) * p.Cout + n));
      st.b[i] = v;
    }
  };
  auto store_smem = [&](int buf) {
#pragma unroll
    for (int i = 0; i < 2; ++i) {
      const int r = a_row + i * 64;
      As[buf][a_q + 0][r] = st.a[i].x;
      As[buf][a_q + 1][r] = st.a[i].y;
      As[buf][a_q + 2][r] = st.a[i].z;
      As[buf][a_q + 3][r] = st.a[i].w;
      *reinterpret_cast<float4*>(&Bs[buf][b_k + i * 8][b_n]) = st.b[i];
    }
  };

  load_global(0);
  store_smem(0);
  __syncthreads();

  int buf = 0;
  for (int kt = 0; kt < KT; ++kt) {
    if (kt + 1 < KT) load_global(kt + 1);
#pragma unroll
    for (int kk = 0; kk < BK; ++kk) {
      const float4 a0 = *reinterpret_cast<const float4*>(&As[buf][kk][ty * 4]);
      const float4 a1 = *reinterpret_cast<const float4*>(&As[buf][kk][64 + ty * 4]);
      const float4 b0 = *reinterpret_cast<const float4*>(&Bs[buf][kk][tx * 4]);
      const float4 b1 = *reinterpret_cast<const float4*>(&Bs[buf][kk][64 + tx * 4]);
      const float av[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
      const float bv[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
#pragma unroll
      for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
    }
    if (kt + 1 < KT) store_smem(buf ^ 1);
    __syncthreads();
    buf ^= 1;
  }

  // ---- epilogue ----
  const int len = p.lens ? p.lens[b] : 0x7fffffff;
  const int64_t ybase = (int64_t)b * p.y_batch_stride;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int t = t0 + (i < 4 ? ty * 4 + i : 64 + ty * 4 + (i - 4));
    if (t >= p.S_out) continue;
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const int n = n0 + tx * 4 + h * 64;
      if (n >= p.Cout) continue;
      const int64_t off = p.y_offset + (int64_t)t * p.y_row_stride + n;
      if (off < 0 || off >= p.y_batch_elems) continue;
      float4 v = make_float4(acc[i][h * 4 + 0], acc[i][h * 4 + 1], acc[i][h * 4 + 2], acc[i][h * 4 + 3]);
      if (p.bias) {
        const float4 bb = __ldg(reinterpret_cast<const float4*>(p.bias + n));
        v.x += bb.x; v.y += bb.y; v.z += bb.z; v.w += bb.w;
      }
      v.x = ss::apply_act(v.x, p.act, p.slope);
      v.y = ss::apply_act(v.y, p.act, p.slope);
      v.z = ss::apply_act(v.z, p.act, p.slope);
      v.w = ss::apply_act(v.w, p.act, p.slope);
      if (p.residual) {
        const float4 r = *reinterpret_cast<const float4*>(p.residual + ybase + off);  // plain load: may alias y
        v.x += r.x; v.y += r.y; v.z += r.z; v.w += r.w;
      }
      if (t >= len) v = make_float4(0.f, 0.f, 0.f, 0.f);
      if (p.y) *reinterpret_cast<float4*>(p.y + ybase + off) = v;
      if (p.accum) {
        float4* ap = reinterpret_cast<float4*>(p.accum + ybase + off);
        float4 a = make_float4(0.f, 0.f, 0.f, 0.f);
        if (!p.accum_init) a = *ap;
        a.x = fmaf(p.accum_scale, v.x, a.x);
        a.y = fmaf(p.accum_scale, v.y, a.y);
        a.z = fmaf(p.accum_scale, v.z, a.z);
        a.w = fmaf(p.accum_scale, v.w, a.w);
        *ap = a;
      }
    }
  }
}

}  // namespace

extern "C" int ss_conv1d_f32(const ss_conv1d_f32_params* p, void* stream) {
  SS_REQUIRE(p && p->x && p->w && (p->y || p->accum), SS_ERR_INVALID, "ss_conv1d_f32: null pointer");
  SS_REQUIRE(p->B > 0 && p->S_in > 0 && p->S_out > 0 && p->Cin > 0 && p->Cout > 0 && p->k > 0 && p->dilation > 0,
             SS_ERR_INVALID, "ss_conv1d_f32: bad shape B=%d S_in=%d S_out=%d Cin=%d Cout=%d k=%d dil=%d", p->B,
             p->S_in, p->S_out, p->Cin, p->Cout, p->k, p->dilation);
  SS_REQUIRE(p->Cin % 4 == 0 && p->Cout % 4 == 0, SS_ERR_UNSUPPORTED,
             "ss_conv1d_f32: Cin (%d) and Cout (%d) must be multiples of 4", p->Cin, p->Cout);
  SS_REQUIRE(p->x_row_stride % 4 == 0 && p->x_batch_stride % 4 == 0 && p->y_row_stride % 4 == 0 &&
                 p->y_batch_stride % 4 == 0 && p->y_offset % 4 == 0 && p->y_batch_elems % 4 == 0,
             SS_ERR_ALIGNMENT, "ss_conv1d_f32: strides/offsets must be multiples of 4 elements");
  SS_REQUIRE(ss::aligned16(p->x) && ss::aligned16(p->w) && ss::aligned16(p->y) && ss::aligned16(p->bias) &&
                 ss::aligned16(p->residual) && ss::aligned16(p->accum),
             SS_ERR_ALIGNMENT, "ss_conv1d_f32: pointers must be 16-byte aligned");
  SS_REQUIRE(p->in_act == SS_ACT_NONE || p->in_act == SS_ACT_LRELU, SS_ERR_UNSUPPORTED,
             "ss_conv1d_f32: in_act must be none or lrelu");
  SS_REQUIRE(p->B <= 65535, SS_ERR_UNSUPPORTED, "ss_conv1d_f32: B > 65535");
  dim3 grid((p->S_out + BM - 1) / BM, (p->Cout + BN - 1) / BN, p->B);
  conv1d_f32_kernel<<<grid, NT, 0, static_cast<cudaStream_t>(stream)>>>(*p);
  SS_CHECK_LAUNCH("ss_conv1d_f32");
  return SS_OK;
}
