// Bandwidth-bound row-wise kernels: embedding + position table, LayerNorm with its surrounding
// element-wise steps, variance-predictor tail, bucketize + embedding add, masks, mean pool.
// All are one-warp-per-row, 128-bit vectorised, each activation row read once and written once.
#include <cuda_bf16.h>

#include "common.cuh"

namespace {

constexpr int ROWS_PER_CTA = 8;  // 8 warps of 32 lanes

// ------------------------------------------------------------------------------------------
__global__ void embed_posenc_kernel(const int64_t* __restrict__ texts, const float* __restrict__ emb,
                                    const float* __restrict__ pe, float* __restrict__ out, int B, int L, int D,
                                    int V) {
  const int row = blockIdx.x * ROWS_PER_CTA + (threadIdx.x >> 5);
  if (row >= B * L) return;
  const int lane = threadIdx.x & 31;
  const int l = row % L;
  int64_t id = texts[row];
  if (id < 0 || id >= V) id = 0;  // torch would raise; keep memory safe
  const float4* e = reinterpret_cast<const float4*>(emb + id * D);
  const float4* q = reinterpret_cast<const float4*>(pe + (int64_t)l * D);
  float4* o = reinterpret_cast<float4*>(out + (int64_t)row * D);
  for (int c = lane; c < D / 4; c += 32) {
    const float4 a = __ldg(e + c), b = __ldg(q + c);
    o[c] = make_float4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w);
  }
}

// ------------------------------------------------------------------------------------------
__device__ __forceinline__ uint2 pack4_bf16(const float4& o) {
  __nv_bfloat162 a = __floats2bfloat162_rn(o.x, o.y), b = __floats2bfloat162_rn(o.z, o.w);
  return make_uint2(*reinterpret_cast<uint32_t*>(&a), *reinterpret_cast<uint32_t*>(&b));
}

__device__ __forceinline__ void store_bf16_or_pair(uint2* yb, int c, int lo_off, int pair, const float4& o) {
  const uint2 hi = pack4_bf16(o);
  yb[c] = hi;
  if (pair) {
    const float4 r = make_float4(o.x - __uint_as_float(hi.x << 16), o.y - __uint_as_float(hi.x & 0xFFFF0000u),
                                 o.z - __uint_as_float(hi.y << 16), o.w - __uint_as_float(hi.y & 0xFFFF0000u));
    yb[c + lo_off] = pack4_bf16(r);
  }
}

// LayerNorm over C = 128*V4 floats held in registers (V4 float4 per lane).
template <int V4>
__device__ __forceinline__ void ln_normalize(float4 (&v)[V4], int C) {
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < V4; ++i) s += (v[i].x + v[i].y) + (v[i].z + v[i].w);
  const float mean = ss::warp_sum(s) / (float)C;
  float q = 0.f;
#pragma unroll
  for (int i = 0; i < V4; ++i) {
    v[i].x -= mean; v[i].y -= mean; v[i].z -= mean; v[i].w -= mean;
    q += (v[i].x * v[i].x + v[i].y * v[i].y) + (v[i].z * v[i].z + v[i].w * v[i].w);
  }
  const float rstd = rsqrtf(ss::warp_sum(q) / (float)C + 1e-5f);
#pragma unroll
  for (int i = 0; i < V4; ++i) {
    v[i].x *= rstd; v[i].y *= rstd; v[i].z *= rstd; v[i].w *= rstd;
  }
}

template <int V4>
__global__ void __launch_bounds__(256) layernorm_kernel(const ss_layernorm_params p) {
  const int row = blockIdx.x * ROWS_PER_CTA + (threadIdx.x >> 5);
  if (row >= p.B * p.S) return;
  const int lane = threadIdx.x & 31;
  const int b = row / p.S, t = row - b * p.S;
  const int C = p.C;
  const bool masked = p.lens && t >= p.lens[b];
  float4* y = reinterpret_cast<float4*>(p.y + (int64_t)row * C);
  // bf16 copy of y, or (bf16_pair) the split-bf16 pair row [hi(C) | lo(C)] that ss_conv1d_bf16x3 reads
  uint2* yb = p.y_bf16 ? reinterpret_cast<uint2*>(static_cast<__nv_bfloat16*>(p.y_bf16) + (int64_t)row * C * (p.bf16_pair ? 2 : 1))
                       : nullptr;
  const int lo_off = C / 4;  // uint2 units from hi to lo inside a pair row
  if (masked) {  // result is known without touching x
#pragma unroll
    for (int i = 0; i < V4; ++i) {
      const int c = lane + i * 32;
      float4 o = make_float4(0.f, 0.f, 0.f, 0.f);
      if (p.pe) o = __ldg(reinterpret_cast<const float4*>(p.pe + (int64_t)t * C) + c);
      y[c] = o;
      if (yb) store_bf16_or_pair(yb, c, lo_off, p.bf16_pair, o);
    }
    return;
  }
  const float4* x = reinterpret_cast<const float4*>(p.x + (int64_t)row * C);
  float4 v[V4];
#pragma unroll
  for (int i = 0; i < V4; ++i) v[i] = x[lane + i * 32];  // plain loads: callers run this in place (y == x), which rules out ld.global.nc
  if (p.residual) {
    const float4* r = reinterpret_cast<const float4*>(p.residual + (int64_t)row * C);
#pragma unroll
    for (int i = 0; i < V4; ++i) {
      const float4 a = r[lane + i * 32];  // may alias y as well
      v[i].x += a.x; v[i].y += a.y; v[i].z += a.z; v[i].w += a.w;
    }
  }
  if (p.pre_relu) {
#pragma unroll
    for (int i = 0; i < V4; ++i) {
      v[i].x = fmaxf(v[i].x, 0.f); v[i].y = fmaxf(v[i].y, 0.f);
      v[i].z = fmaxf(v[i].z, 0.f); v[i].w = fmaxf(v[i].w, 0.f);
    }
  }
  ln_normalize<V4>(v, C);
  const float sg = p.gamma ? __ldg(p.s_gamma) : 0.f;
  const float sb = p.gamma ? __ldg(p.s_beta) : 0.f;
#pragma unroll
  for (int i = 0; i < V4; ++i) {
    const int c = lane + i * 32;
    const float4 w = __ldg(reinterpret_cast<const float4*>(p.ln_w) + c);
    const float4 bb = __ldg(reinterpret_cast<const float4*>(p.ln_b) + c);
    float4 o = make_float4(v[i].x * w.x + bb.x, v[i].y * w.y + bb.y, v[i].z * w.z + bb.z, v[i].w * w.w + bb.w);
    if (p.gamma) {
      const float4 g = __ldg(reinterpret_cast<const float4*>(p.gamma + (int64_t)b * C) + c);
      const float4 be = __ldg(reinterpret_cast<const float4*>(p.beta + (int64_t)b * C) + c);
      o.x = (sg * g.x + 1.0f) * o.x + sb * be.x;
      o.y = (sg * g.y + 1.0f) * o.y + sb * be.y;
      o.z = (sg * g.z + 1.0f) * o.z + sb * be.z;
      o.w = (sg * g.w + 1.0f) * o.w + sb * be.w;
    }
    if (p.pe) {
      const float4 q = __ldg(reinterpret_cast<const float4*>(p.pe + (int64_t)t * C) + c);
      o.x += q.x; o.y += q.y; o.z += q.z; o.w += q.w;
    }
    y[c] = o;
    if (yb) store_bf16_or_pair(yb, c, lo_off, p.bf16_pair, o);
  }
}

// ------------------------------------------------------------------------------------------
template <int V4>
__global__ void __launch_bounds__(256) predictor_head_kernel(const ss_predictor_head_params p) {
  const int row = blockIdx.x * ROWS_PER_CTA + (threadIdx.x >> 5);
  if (row >= p.B * p.L) return;
  const int lane = threadIdx.x & 31;
  const int b = row / p.L, l = row - b * p.L;
  const int C = p.C;
  if (p.lens && l >= p.lens[b]) {
    if (lane == 0) p.pred[row] = 0.0f;
    return;
  }
  const float4* x = reinterpret_cast<const float4*>(p.h + (int64_t)row * C);
  float4 v[V4];
#pragma unroll
  for (int i = 0; i < V4; ++i) v[i] = __ldg(x + lane + i * 32);
  ln_normalize<V4>(v, C);
  const float sg = p.gamma ? __ldg(p.s_gamma) : 0.f;
  const float sb = p.gamma ? __ldg(p.s_beta) : 0.f;
  float dot = 0.f;
#pragma unroll
  for (int i = 0; i < V4; ++i) {
    const int c = lane + i * 32;
    const float4 w = __ldg(reinterpret_cast<const float4*>(p.ln_w) + c);
    const float4 bb = __ldg(reinterpret_cast<const float4*>(p.ln_b) + c);
    float4 o = make_float4(v[i].x * w.x + bb.x, v[i].y * w.y + bb.y, v[i].z * w.z + bb.z, v[i].w * w.w + bb.w);
    if (p.gamma) {
      const float4 g = __ldg(reinterpret_cast<const float4*>(p.gamma + (int64_t)b * C) + c);
      const float4 be = __ldg(reinterpret_cast<const float4*>(p.beta + (int64_t)b * C) + c);
      o.x = (sg * g.x + 1.0f) * o.x + sb * be.x;
      o.y = (sg * g.y + 1.0f) * o.y + sb * be.y;
      o.z = (sg * g.z + 1.0f) * o.z + sb * be.z;
      o.w = (sg * g.w + 1.0f) * o.w + sb * be.w;
    }
    const float4 lw = __ldg(reinterpret_cast<const float4*>(p.lin_w) + c);
    dot += (o.x * lw.x + o.y * lw.y) + (o.z * lw.z + o.w * lw.w);
  }
  dot = ss::warp_sum(dot);
  if (lane == 0) p.pred[row] = dot + __ldg(p.lin_b);
}

// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) bucket_embed_add_kernel(float* __restrict__ pred, float control,
                                                               const float* __restrict__ control_vec,
                                                               const float* __restrict__ edges, int n_edges,
                                                               const float* __restrict__ table,
                                                               float* __restrict__ x, int32_t* __restrict__ idx_out,
                                                               int rows, int C) {
  extern __shared__ float s_edges[];
  for (int i = threadIdx.x; i < n_edges; i += blockDim.x) s_edges[i] = edges[i];
  __syncthreads();
  const int row = blockIdx.x * ROWS_PER_CTA + (threadIdx.x >> 5);
  if (row >= rows) return;
  const int lane = threadIdx.x & 31;
  const float v = pred[row] * (control_vec ? control_vec[row] : control);  // modules.py:88,99: prediction * control
  // lower bound: first edge that is >= v; NaN fails every comparison and ends at n_edges
  int lo = 0, hi = n_edges;
  while (lo < hi) {
    const int mid = (lo + hi) >> 1;
    if (!(s_edges[mid] >= v)) lo = mid + 1; else hi = mid;
  }
  __syncwarp();
  if (lane == 0) {
    pred[row] = v;
    if (idx_out) idx_out[row] = lo;
  }
  const float4* e = reinterpret_cast<const float4*>(table + (int64_t)lo * C);
  float4* xr = reinterpret_cast<float4*>(x + (int64_t)row * C);
  for (int c = lane; c < C / 4; c += 32) {
    float4 a = xr[c];
    const float4 t = __ldg(e + c);
    a.x += t.x; a.y += t.y; a.z += t.z; a.w += t.w;
    xr[c] = a;
  }
}

// ------------------------------------------------------------------------------------------
__global__ void length_mask_kernel(const int32_t* __restrict__ lens, uint8_t* __restrict__ mask, int B, int T) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (int64_t)B * T) return;
  const int b = (int)(i / T), t = (int)(i - (int64_t)b * T);
  mask[i] = t >= lens[b] ? 1 : 0;
}

// one CTA per (batch, 128-channel slab); threads = 32 channel-quads x 8 row groups
__global__ void __launch_bounds__(256) mean_pool_kernel(const float* __restrict__ x, float* __restrict__ out, int S_all,
                                                        int C, const int32_t* __restrict__ lens) {
  __shared__ float4 part[8][32];
  const int b = blockIdx.y;
  // lens given (batch-invariant mode): mean over the utterance's own frames; otherwise over the padded length
  const int S = lens ? min(max(lens[b], 1), S_all) : S_all;
  const int cq = blockIdx.x * 32 + (threadIdx.x & 31);  // float4 column
  const int g = threadIdx.x >> 5;
  float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
  if (cq < C / 4) {
    const float4* xp = reinterpret_cast<const float4*>(x + (int64_t)b * S_all * C) + cq;
    for (int t = g; t < S; t += 8) {
      const float4 a = __ldg(xp + (int64_t)t * (C / 4));
      s.x += a.x; s.y += a.y; s.z += a.z; s.w += a.w;
    }
  }
  part[g][threadIdx.x & 31] = s;
  __syncthreads();
  if (g == 0 && cq < C / 4) {
    for (int i = 1; i < 8; ++i) {
      const float4 a = part[i][threadIdx.x & 31];
      s.x += a.x; s.y += a.y; s.z += a.z; s.w += a.w;
    }
    const float inv = 1.0f / (float)S;
    reinterpret_cast<float4*>(out + (int64_t)b * C)[cq] = make_float4(s.x * inv, s.y * inv, s.z * inv, s.w * inv);
  }
}

// x[b, t, :] = 0 for t >= lens[b]: one warp per row, only padded rows are touched
__global__ void mask_rows_kernel(float* __restrict__ x, const int32_t* __restrict__ lens, int B, int S, int C) {
  const int64_t row = (int64_t)blockIdx.x * ROWS_PER_CTA + (threadIdx.x >> 5);
  if (row >= (int64_t)B * S) return;
  const int b = (int)(row / S), t = (int)(row - (int64_t)b * S);
  if (t < lens[b]) return;
  float4* xr = reinterpret_cast<float4*>(x + row * C);
  for (int c = threadIdx.x & 31; c < C / 4; c += 32) xr[c] = make_float4(0.f, 0.f, 0.f, 0.f);
}

__global__ void add_speaker_kernel(float* __restrict__ x, const int64_t* __restrict__ ids,
                                   const float* __restrict__ table, int B, int L, int C, int n_speakers) {
  const int row = blockIdx.x * ROWS_PER_CTA + (threadIdx.x >> 5);
  if (row >= B * L) return;
  const int lane = threadIdx.x & 31;
  int64_t id = ids[row / L];
  if (id < 0 || id >= n_speakers) id = 0;
  const float4* e = reinterpret_cast<const float4*>(table + id * C);
  float4* xr = reinterpret_cast<float4*>(x + (int64_t)row * C);
  for (int c = lane; c < C / 4; c += 32) {
    float4 a = xr[c];
    const float4 t = __ldg(e + c);
    a.x += t.x; a.y += t.y; a.z += t.z; a.w += t.w;
    xr[c] = a;
  }
}

__global__ void add_posenc_kernel(const float* __restrict__ x, const float* __restrict__ pe, float* __restrict__ out,
                                  int B, int S, int C) {
  const int row = blockIdx.x * ROWS_PER_CTA + (threadIdx.x >> 5);
  if (row >= B * S) return;
  const int lane = threadIdx.x & 31;
  const int t = row % S;
  const float4* xr = reinterpret_cast<const float4*>(x + (int64_t)row * C);
  const float4* q = reinterpret_cast<const float4*>(pe + (int64_t)t * C);
  float4* o = reinterpret_cast<float4*>(out + (int64_t)row * C);
  for (int c = lane; c < C / 4; c += 32) {
    const float4 a = __ldg(xr + c), b = __ldg(q + c);
    o[c] = make_float4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w);
  }
}

__global__ void cast_bf16_kernel(const float* __restrict__ x, uint4* __restrict__ out, int64_t n8, int act,
                                 float slope) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n8; i += (int64_t)gridDim.x * blockDim.x) {
    const float4 a = ss::ld_stream(reinterpret_cast<const float4*>(x) + 2 * i);
    const float4 b = ss::ld_stream(reinterpret_cast<const float4*>(x) + 2 * i + 1);
    float f[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
    uint32_t u[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      __nv_bfloat162 h = __floats2bfloat162_rn(ss::apply_act(f[2 * j], act, slope), ss::apply_act(f[2 * j + 1], act, slope));
      u[j] = *reinterpret_cast<uint32_t*>(&h);
    }
    out[i] = make_uint4(u[0], u[1], u[2], u[3]);
  }
}

// act(x) as split-bf16 pair rows: out[r, 0:C] = hi = bf16(v), out[r, C:2C] = lo = bf16(v - hi); 8 channels per thread
__global__ void split_bf16_kernel(const float* __restrict__ x, __nv_bfloat16* __restrict__ out, int64_t n8, int C8, int C,
                                  int act, float slope) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n8; i += (int64_t)gridDim.x * blockDim.x) {
    const float4 a = ss::ld_stream(reinterpret_cast<const float4*>(x) + 2 * i);
    const float4 b = ss::ld_stream(reinterpret_cast<const float4*>(x) + 2 * i + 1);
    float f[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
    uint32_t h[4], l[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float v0 = ss::apply_act(f[2 * j], act, slope), v1 = ss::apply_act(f[2 * j + 1], act, slope);
      __nv_bfloat162 hh = __floats2bfloat162_rn(v0, v1);
      __nv_bfloat162 ll = __floats2bfloat162_rn(v0 - __low2float(hh), v1 - __high2float(hh));
      h[j] = *reinterpret_cast<uint32_t*>(&hh);
      l[j] = *reinterpret_cast<uint32_t*>(&ll);
    }
    const int64_t row = i / C8;
    const int c = (int)(i - row * C8) * 8;
    __nv_bfloat16* dst = out + row * 2 * C + c;
    *reinterpret_cast<uint4*>(dst) = make_uint4(h[0], h[1], h[2], h[3]);
    *reinterpret_cast<uint4*>(dst + C) = make_uint4(l[0], l[1], l[2], l[3]);
  }
}

inline int row_grid(int64_t rows) { return (int)((rows + ROWS_PER_CTA - 1) / ROWS_PER_CTA); }

}  // namespace

extern "C" {

int ss_embed_posenc(const int64_t* texts, const float* emb, const float* pe, float* out, int B, int L, int D, int V,
                    void* stream) {
  SS_REQUIRE(texts && emb && pe && out, SS_ERR_INVALID, "ss_embed_posenc: null pointer");
  SS_REQUIRE(B > 0 && L > 0 && D > 0 && D % 4 == 0 && V > 0, SS_ERR_INVALID, "ss_embed_posenc: bad shape");
  embed_posenc_kernel<<<row_grid((int64_t)B * L), 256, 0, (cudaStream_t)stream>>>(texts, emb, pe, out, B, L, D, V);
  SS_CHECK_LAUNCH("ss_embed_posenc");
  return SS_OK;
}

int ss_add_posenc(const float* x, const float* pe, float* out, int B, int S, int C, void* stream) {
  SS_REQUIRE(x && pe && out && B > 0 && S > 0 && C > 0 && C % 4 == 0, SS_ERR_INVALID, "ss_add_posenc: bad argument");
  add_posenc_kernel<<<row_grid((int64_t)B * S), 256, 0, (cudaStream_t)stream>>>(x, pe, out, B, S, C);
  SS_CHECK_LAUNCH("ss_add_posenc");
  return SS_OK;
}

int ss_cast_bf16(const float* x, void* out, int64_t n, int act, float slope, void* stream) {
  SS_REQUIRE(x && out && n > 0 && n % 8 == 0, SS_ERR_INVALID, "ss_cast_bf16: bad argument (n must be a multiple of 8)");
  SS_REQUIRE(ss::aligned16(x) && ss::aligned16(out), SS_ERR_ALIGNMENT, "ss_cast_bf16: 16-byte alignment");
  const int64_t n8 = n / 8;
  const int64_t want = (n8 + 255) / 256;
  const int grid = (int)(want < 148 * 16 ? want : 148 * 16);
  cast_bf16_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(x, reinterpret_cast<uint4*>(out), n8, act, slope);
  SS_CHECK_LAUNCH("ss_cast_bf16");
  return SS_OK;
}

int ss_split_bf16(const float* x, void* out, int64_t rows, int C, int act, float slope, void* stream) {
  SS_REQUIRE(x && out && rows > 0 && C > 0 && C % 8 == 0, SS_ERR_INVALID,
             "ss_split_bf16: bad argument (C must be a multiple of 8)");
  SS_REQUIRE(ss::aligned16(x) && ss::aligned16(out), SS_ERR_ALIGNMENT, "ss_split_bf16: 16-byte alignment");
  const int64_t n8 = rows * (C / 8);
  const int64_t want = (n8 + 255) / 256;
  const int grid = (int)(want < 148 * 16 ? want : 148 * 16);
  split_bf16_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(x, static_cast<__nv_bfloat16*>(out), n8, C / 8, C, act, slope);
  SS_CHECK_LAUNCH("ss_split_bf16");
  return SS_OK;
}

int ss_layernorm(const ss_layernorm_params* p, void* stream) {
  SS_REQUIRE(p && p->x && p->ln_w && p->ln_b && p->y, SS_ERR_INVALID, "ss_layernorm: null pointer");
  SS_REQUIRE(p->B > 0 && p->S > 0, SS_ERR_INVALID, "ss_layernorm: bad shape");
  SS_REQUIRE(!p->gamma || (p->beta && p->s_gamma && p->s_beta), SS_ERR_INVALID, "ss_layernorm: FiLM needs all four");
  const int grid = row_grid((int64_t)p->B * p->S);
  cudaStream_t st = (cudaStream_t)stream;
  switch (p->C) {
    case 128: layernorm_kernel<1><<<grid, 256, 0, st>>>(*p); break;
    case 256: layernorm_kernel<2><<<grid, 256, 0, st>>>(*p); break;
    case 512: layernorm_kernel<4><<<grid, 256, 0, st>>>(*p); break;
    case 1024: layernorm_kernel<8><<<grid, 256, 0, st>>>(*p); break;
    default: SS_REQUIRE(false, SS_ERR_UNSUPPORTED, "ss_layernorm: C=%d not in {128,256,512,1024}", p->C);
  }
  SS_CHECK_LAUNCH("ss_layernorm");
  return SS_OK;
}

int ss_predictor_head(const ss_predictor_head_params* p, void* stream) {
  SS_REQUIRE(p && p->h && p->ln_w && p->ln_b && p->lin_w && p->lin_b && p->pred, SS_ERR_INVALID,
             "ss_predictor_head: null pointer");
  SS_REQUIRE(p->B > 0 && p->L > 0, SS_ERR_INVALID, "ss_predictor_head: bad shape");
  SS_REQUIRE(!p->gamma || (p->beta && p->s_gamma && p->s_beta), SS_ERR_INVALID,
             "ss_predictor_head: FiLM needs all four");
  const int grid = row_grid((int64_t)p->B * p->L);
  cudaStream_t st = (cudaStream_t)stream;
  switch (p->C) {
    case 128: predictor_head_kernel<1><<<grid, 256, 0, st>>>(*p); break;
    case 256: predictor_head_kernel<2><<<grid, 256, 0, st>>>(*p); break;
    case 512: predictor_head_kernel<4><<<grid, 256, 0, st>>>(*p); break;
    default: SS_REQUIRE(false, SS_ERR_UNSUPPORTED, "ss_predictor_head: C=%d not in {128,256,512}", p->C);
  }
  SS_CHECK_LAUNCH("ss_predictor_head");
  return SS_OK;
}

int ss_bucket_embed_add_v(float* pred, const float* control_vec, const float* edges, int n_edges, const float* table,
                          float* x, int32_t* idx_out, int B, int L, int C, void* stream) {
  SS_REQUIRE(pred && control_vec && edges && table && x, SS_ERR_INVALID, "ss_bucket_embed_add_v: null pointer");
  SS_REQUIRE(B > 0 && L > 0 && C > 0 && C % 4 == 0 && n_edges > 0 && n_edges <= 1024, SS_ERR_INVALID,
             "ss_bucket_embed_add_v: bad shape");
  bucket_embed_add_kernel<<<row_grid((int64_t)B * L), 256, n_edges * sizeof(float), (cudaStream_t)stream>>>(
      pred, 1.0f, control_vec, edges, n_edges, table, x, idx_out, B * L, C);
  SS_CHECK_LAUNCH("ss_bucket_embed_add_v");
  return SS_OK;
}

int ss_bucket_embed_add(float* pred, float control, const float* edges, int n_edges, const float* table, float* x,
                        int32_t* idx_out, int B, int L, int C, void* stream) {
  SS_REQUIRE(pred && edges && table && x, SS_ERR_INVALID, "ss_bucket_embed_add: null pointer");
  SS_REQUIRE(B > 0 && L > 0 && C > 0 && C % 4 == 0 && n_edges > 0 && n_edges <= 1024, SS_ERR_INVALID,
             "ss_bucket_embed_add: bad shape");
  bucket_embed_add_kernel<<<row_grid((int64_t)B * L), 256, n_edges * sizeof(float), (cudaStream_t)stream>>>(
      pred, control, nullptr, edges, n_edges, table, x, idx_out, B * L, C);
  SS_CHECK_LAUNCH("ss_bucket_embed_add");
  return SS_OK;
}

int ss_length_mask(const int32_t* lens, uint8_t* mask, int B, int T, void* stream) {
  SS_REQUIRE(lens && mask && B > 0 && T >= 0, SS_ERR_INVALID, "ss_length_mask: bad argument");
  if (T == 0) return SS_OK;
  const int64_t n = (int64_t)B * T;
  length_mask_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(lens, mask, B, T);
  SS_CHECK_LAUNCH("ss_length_mask");
  return SS_OK;
}

int ss_mean_pool(const float* x, float* out, int B, int S, int C, void* stream) {
  SS_REQUIRE(x && out && B > 0 && S > 0 && C > 0 && C % 4 == 0, SS_ERR_INVALID, "ss_mean_pool: bad argument");
  dim3 grid((C / 4 + 31) / 32, B);
  mean_pool_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(x, out, S, C, nullptr);
  SS_CHECK_LAUNCH("ss_mean_pool");
  return SS_OK;
}

int ss_mean_pool_lens(const float* x, const int32_t* lens, float* out, int B, int S, int C, void* stream) {
  SS_REQUIRE(x && lens && out && B > 0 && S > 0 && C > 0 && C % 4 == 0, SS_ERR_INVALID, "ss_mean_pool_lens: bad argument");
  dim3 grid((C / 4 + 31) / 32, B);
  mean_pool_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(x, out, S, C, lens);
  SS_CHECK_LAUNCH("ss_mean_pool_lens");
  return SS_OK;
}

int ss_mask_rows(float* x, const int32_t* lens, int B, int S, int C, void* stream) {
  SS_REQUIRE(x && lens && B > 0 && S > 0 && C > 0 && C % 4 == 0, SS_ERR_INVALID, "ss_mask_rows: bad argument");
  mask_rows_kernel<<<row_grid((int64_t)B * S), 256, 0, (cudaStream_t)stream>>>(x, lens, B, S, C);
  SS_CHECK_LAUNCH("ss_mask_rows");
  return SS_OK;
}

int ss_add_speaker(float* x, const int64_t* ids, const float* table, int B, int L, int C, int n_speakers,
                   void* stream) {
  SS_REQUIRE(x && ids && table && B > 0 && L > 0 && C > 0 && C % 4 == 0 && n_speakers > 0, SS_ERR_INVALID,
             "ss_add_speaker: bad argument");
  add_speaker_kernel<<<row_grid((int64_t)B * L), 256, 0, (cudaStream_t)stream>>>(x, ids, table, B, L, C, n_speakers);
  SS_CHECK_LAUNCH("ss_add_speaker");
  return SS_OK;
}

}  // extern "C"
