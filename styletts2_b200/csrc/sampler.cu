// sampler.cu — kernels of the style sampler: the StyleTransformer1d / Transformer1d
// denoiser of KDiffusion and the ADPM2 ancestral sampler around it.
//
// Reference: Modules/diffusion/modules.py:18-38 (AdaLayerNorm), :121-186 (StyleTransformer1d.run /
// forward), :232-278 (StyleAttention), :489-526 (AttentionBase), :629-652 (time embedding);
// Modules/diffusion/sampler.py:184-208 (KDiffusion.denoise_fn), :480-519 (ADPM2Sampler).
//
// The dense token-level contractions (to_q, to_kv, to_out, feed_forward) run on the conv kernels
// (tcgen05 on the bf16 path) as 1x1 convs over the flattened [items * tokens] rows; this file holds
// what sits between them.  The residual stream X is kept in fp32 on both paths; AT (fp32 or bf16)
// is the storage type of the GEMM operands.
#include "common.cuh"

namespace st2 {

__device__ __forceinline__ float gelu_erf(float v) { return 0.5f * v * (1.f + erff(v * 0.70710678118654752f)); }

// four consecutive elements (16-byte fp32 / 8-byte bf16 accesses; the address must be aligned accordingly)
template <typename T>
__device__ __forceinline__ float4 ld4_f(const T* p);
template <>
__device__ __forceinline__ float4 ld4_f<float>(const float* p) { return *reinterpret_cast<const float4*>(p); }
template <>
__device__ __forceinline__ float4 ld4_f<__nv_bfloat16>(const __nv_bfloat16* p) {
  const uint2 r = *reinterpret_cast<const uint2*>(p);
  return make_float4(__uint_as_float(r.x << 16), __uint_as_float(r.x & 0xFFFF0000u), __uint_as_float(r.y << 16),
                     __uint_as_float(r.y & 0xFFFF0000u));
}
template <typename T>
__device__ __forceinline__ void st4_f(T* p, float4 v);
template <>
__device__ __forceinline__ void st4_f<float>(float* p, float4 v) { *reinterpret_cast<float4*>(p) = v; }
template <>
__device__ __forceinline__ void st4_f<__nv_bfloat16>(__nv_bfloat16* p, float4 v) {
  __nv_bfloat162 a = __floats2bfloat162_rn(v.x, v.y), b = __floats2bfloat162_rn(v.z, v.w);
  uint2 r;
  r.x = *reinterpret_cast<uint32_t*>(&a);
  r.y = *reinterpret_cast<uint32_t*>(&b);
  *reinterpret_cast<uint2*>(p) = r;
}

// ------------------------------------------------------------------ small dense layers (fp32)
// y[r][o] = act(bias[o] + sum_i W[o][i] * (x[r][i] + xb[i])).  A warp owns OUTS output features x ROWS rows
// (register block: OUTS + ROWS 16-byte loads per 4 * OUTS * ROWS FMAs), lanes split the reduction.
// VEC: 16-byte loads (I, ldx multiples of 4, 16-byte aligned bases)
template <int ROWS, int OUTS, bool VEC>
__global__ void __launch_bounds__(256)
small_linear_kernel(const float* __restrict__ x, int ldx, const float* __restrict__ xb, const float* __restrict__ W,
                    const float* __restrict__ bias, float* __restrict__ y, int ldy, int R, int I, int O, int act) {
  const int o0 = (blockIdx.x * 8 + (threadIdx.x >> 5)) * OUTS, lane = threadIdx.x & 31;
  const int r0 = blockIdx.y * ROWS;
  if (o0 >= O) return;
  float acc[OUTS][ROWS];
#pragma unroll
  for (int u = 0; u < OUTS; ++u)
#pragma unroll
    for (int j = 0; j < ROWS; ++j) acc[u][j] = 0.f;
  const float* wr[OUTS];
#pragma unroll
  for (int u = 0; u < OUTS; ++u) wr[u] = W + (int64_t)min(o0 + u, O - 1) * I;
  if (VEC) {
    for (int i = lane * 4; i < I; i += 128) {
      float4 add = make_float4(0.f, 0.f, 0.f, 0.f);
      if (xb) add = __ldg(reinterpret_cast<const float4*>(xb + i));
      float4 w[OUTS];
#pragma unroll
      for (int u = 0; u < OUTS; ++u) w[u] = __ldg(reinterpret_cast<const float4*>(wr[u] + i));
#pragma unroll
      for (int j = 0; j < ROWS; ++j) {
        const int r = min(r0 + j, R - 1);
        float4 v = __ldg(reinterpret_cast<const float4*>(x + (int64_t)r * ldx + i));
        v.x += add.x; v.y += add.y; v.z += add.z; v.w += add.w;
#pragma unroll
        for (int u = 0; u < OUTS; ++u) {
          acc[u][j] = fmaf(w[u].x, v.x, acc[u][j]);
          acc[u][j] = fmaf(w[u].y, v.y, acc[u][j]);
          acc[u][j] = fmaf(w[u].z, v.z, acc[u][j]);
          acc[u][j] = fmaf(w[u].w, v.w, acc[u][j]);
        }
      }
    }
  } else {
    for (int i = lane; i < I; i += 32) {
      const float add = xb ? __ldg(xb + i) : 0.f;
      float w[OUTS];
#pragma unroll
      for (int u = 0; u < OUTS; ++u) w[u] = __ldg(wr[u] + i);
#pragma unroll
      for (int j = 0; j < ROWS; ++j) {
        const int r = min(r0 + j, R - 1);
        const float v = __ldg(x + (int64_t)r * ldx + i) + add;
#pragma unroll
        for (int u = 0; u < OUTS; ++u) acc[u][j] = fmaf(w[u], v, acc[u][j]);
      }
    }
  }
#pragma unroll
  for (int u = 0; u < OUTS; ++u)
#pragma unroll
    for (int j = 0; j < ROWS; ++j) acc[u][j] = warp_sum(acc[u][j]);
  if (lane == 0) {
#pragma unroll
    for (int u = 0; u < OUTS; ++u) {
      if (o0 + u >= O) continue;
      const float b = bias ? bias[o0 + u] : 0.f;
#pragma unroll
      for (int j = 0; j < ROWS; ++j)
        if (r0 + j < R) {
          const float v = acc[u][j] + b;
          y[(int64_t)(r0 + j) * ldy + o0 + u] = act ? gelu_erf(v) : v;
        }
    }
  }
}
int launch_small_linear(const float* x, int ldx, const float* xb, const float* W, const float* bias, float* y, int ldy,
                        int R, int I, int O, int gelu, cudaStream_t st) {
  const bool vec = I % 4 == 0 && ldx % 4 == 0 && (((uintptr_t)x | (uintptr_t)W | (uintptr_t)xb) & 15) == 0;
  if (R == 1) {
    dim3 grid(ceil_div(O, 8), 1);
    if (vec) small_linear_kernel<1, 1, true><<<grid, 256, 0, st>>>(x, ldx, xb, W, bias, y, ldy, R, I, O, gelu);
    else small_linear_kernel<1, 1, false><<<grid, 256, 0, st>>>(x, ldx, xb, W, bias, y, ldy, R, I, O, gelu);
  } else {
    dim3 grid(ceil_div(O, 8 * 4), ceil_div(R, 8));
    if (vec) small_linear_kernel<8, 4, true><<<grid, 256, 0, st>>>(x, ldx, xb, W, bias, y, ldy, R, I, O, gelu);
    else small_linear_kernel<8, 4, false><<<grid, 256, 0, st>>>(x, ldx, xb, W, bias, y, ldy, R, I, O, gelu);
  }
  ST2_LAUNCH_OK("small_linear");
  return ST2_OK;
}

// LearnedPositionalEmbedding (modules.py:629-641): out = [t, sin(2 pi t w), cos(2 pi t w)]
__global__ void time_fourier_kernel(float t, const float* __restrict__ w, int half, float* __restrict__ out) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j == 0) out[0] = t;
  if (j < half) {
    const float f = t * w[j] * 2.f * 3.14159265358979323846f;
    out[1 + j] = sinf(f);
    out[1 + half + j] = cosf(f);
  }
}
int launch_time_fourier(float t, const float* w, int half, float* out, cudaStream_t st) {
  time_fourier_kernel<<<ceil_div(half, 128), 128, 0, st>>>(t, w, half, out);
  ST2_LAUNCH_OK("time_fourier");
  return ST2_OK;
}

// ------------------------------------------------------------------ token rows
// X[(bx, t)][:] = cat(c_in * x[b][:C], embedding[b][t][:] or fixed[t][:])   (modules.py:149; bx >= B is
// the classifier-free-guidance run on the fixed embedding, :180-181)
__global__ void __launch_bounds__(256)
build_rows_kernel(const float* __restrict__ x, float c_in, const float* __restrict__ emb, const float* __restrict__ fixed,
                  float* __restrict__ X, int B, int n, int C, int E) {
  const int F = C + E;
  const int64_t row = blockIdx.x;
  const int bx = (int)(row / n), t = (int)(row % n), b = bx % B;
  const float* e = bx < B ? emb + ((int64_t)b * n + t) * E : fixed + (int64_t)t * E;
  float* dst = X + row * F;
  for (int c = threadIdx.x; c < F; c += blockDim.x) dst[c] = c < C ? c_in * x[(int64_t)b * C + c] : e[c - C];
}
int launch_build_rows(const float* x, float c_in, const float* emb, const float* fixed, float* X, int B, int Bx, int n,
                      int C, int E, cudaStream_t st) {
  build_rows_kernel<<<(unsigned)((int64_t)Bx * n), 256, 0, st>>>(x, c_in, emb, fixed, X, B, n, C, E);
  ST2_LAUNCH_OK("build_rows");
  return ST2_OK;
}

// x = x + delta + mapping[b]  (written back, fp32), then the two (adaptive) layer norms of StyleAttention
// (modules.py:268-275): nq = (1 + g1) * LN(x) + b1, nk = (1 + g2) * LN(x) + b2, LN biased variance, eps 1e-5.
// One warp per row.  gb = [B or 1][gb_stride]: this block's (g1, b1, g2, b2), F values each.
template <typename AT, int NV>
__global__ void __launch_bounds__(256)
adaln2_kernel(float* __restrict__ X, const AT* __restrict__ delta, const float* __restrict__ mapping, int64_t map_bstride,
              const float* __restrict__ gb, int64_t gb_bstride, AT* __restrict__ nq, AT* __restrict__ nk, int64_t rows,
              int n, int B, int F) {
  // lane owns the 4-channel groups c = 4 * (lane + 32 j), j < NV (F <= 128 NV): the row lives in registers
  const int64_t row = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (row >= rows) return;
  const int b = (int)((row / n) % B);
  float* xr = X + row * F;
  const AT* dr = delta ? delta + row * F : nullptr;
  const float* mp = mapping + (int64_t)b * map_bstride;
  float4 v[NV];
  float s = 0.f;
#pragma unroll
  for (int j = 0; j < NV; ++j) {
    const int c = 4 * (lane + 32 * j);
    v[j] = make_float4(0.f, 0.f, 0.f, 0.f);
    if (c < F) {
      const float4 a = *reinterpret_cast<const float4*>(xr + c), m = __ldg(reinterpret_cast<const float4*>(mp + c));
      v[j] = make_float4(a.x + m.x, a.y + m.y, a.z + m.z, a.w + m.w);
      if (dr) {
        const float4 d = ld4_f(dr + c);
        v[j].x += d.x; v[j].y += d.y; v[j].z += d.z; v[j].w += d.w;
      }
      *reinterpret_cast<float4*>(xr + c) = v[j];
      s += (v[j].x + v[j].y) + (v[j].z + v[j].w);
    }
  }
  const float mean = warp_sum(s) / (float)F;
  float q = 0.f;
#pragma unroll
  for (int j = 0; j < NV; ++j) {
    if (4 * (lane + 32 * j) < F) {
      const float d0 = v[j].x - mean, d1 = v[j].y - mean, d2 = v[j].z - mean, d3 = v[j].w - mean;
      q += (d0 * d0 + d1 * d1) + (d2 * d2 + d3 * d3);
    }
  }
  const float rstd = rsqrtf(warp_sum(q) / (float)F + 1e-5f);
  const float* g = gb + (int64_t)b * gb_bstride;
  AT* q_o = nq + row * F;
  AT* k_o = nk + row * F;
#pragma unroll
  for (int j = 0; j < NV; ++j) {
    const int c = 4 * (lane + 32 * j);
    if (c < F) {
      const float xh[4] = {(v[j].x - mean) * rstd, (v[j].y - mean) * rstd, (v[j].z - mean) * rstd, (v[j].w - mean) * rstd};
      const float4 g1 = __ldg(reinterpret_cast<const float4*>(g + c)), b1 = __ldg(reinterpret_cast<const float4*>(g + F + c));
      const float4 g2 = __ldg(reinterpret_cast<const float4*>(g + 2 * F + c)), b2 = __ldg(reinterpret_cast<const float4*>(g + 3 * F + c));
      const float ga[4] = {g1.x, g1.y, g1.z, g1.w}, ba[4] = {b1.x, b1.y, b1.z, b1.w};
      const float gc[4] = {g2.x, g2.y, g2.z, g2.w}, bc[4] = {b2.x, b2.y, b2.z, b2.w};
      st4_f(q_o + c, make_float4(fmaf(1.f + ga[0], xh[0], ba[0]), fmaf(1.f + ga[1], xh[1], ba[1]),
                                 fmaf(1.f + ga[2], xh[2], ba[2]), fmaf(1.f + ga[3], xh[3], ba[3])));
      st4_f(k_o + c, make_float4(fmaf(1.f + gc[0], xh[0], bc[0]), fmaf(1.f + gc[1], xh[1], bc[1]),
                                 fmaf(1.f + gc[2], xh[2], bc[2]), fmaf(1.f + gc[3], xh[3], bc[3])));
    }
  }
}
template <typename AT>
int launch_adaln2(float* X, const AT* delta, const float* mapping, int64_t map_bstride, const float* gb, int64_t gb_bstride,
                  AT* nq, AT* nk, int64_t rows, int n, int B, int F, cudaStream_t st) {
  ST2_REQUIRE(F % 4 == 0 && F <= 2048, "adaln2: features must be a multiple of 4, at most 2048");
  const unsigned grid = (unsigned)((rows + 7) / 8);
  if (F <= 1024) adaln2_kernel<AT, 8><<<grid, 256, 0, st>>>(X, delta, mapping, map_bstride, gb, gb_bstride, nq, nk, rows, n, B, F);
  else adaln2_kernel<AT, 16><<<grid, 256, 0, st>>>(X, delta, mapping, map_bstride, gb, gb_bstride, nq, nk, rows, n, B, F);
  ST2_LAUNCH_OK("adaln2");
  return ST2_OK;
}
template int launch_adaln2<float>(float*, const float*, const float*, int64_t, const float*, int64_t, float*, float*, int64_t,
                                  int, int, int, cudaStream_t);
template int launch_adaln2<__nv_bfloat16>(float*, const __nv_bfloat16*, const float*, int64_t, const float*, int64_t,
                                          __nv_bfloat16*, __nv_bfloat16*, int64_t, int, int, int, cudaStream_t);

// X += delta; Xc = AT(X)   (the attention residual, modules.py:229; Xc feeds feed_forward.0)
template <typename AT>
__global__ void __launch_bounds__(256)
add_cast_kernel(float* __restrict__ X, const AT* __restrict__ delta, AT* __restrict__ Xc, int64_t n) {
  const int64_t i = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * 4;
  if (i >= n) return;
  float4 v = *reinterpret_cast<float4*>(X + i);
  const float4 d = ld4_f(delta + i);
  v.x += d.x; v.y += d.y; v.z += d.z; v.w += d.w;
  *reinterpret_cast<float4*>(X + i) = v;
  st4_f(Xc + i, v);
}
template <typename AT>
int launch_add_cast(float* X, const AT* delta, AT* Xc, int64_t n, cudaStream_t st) {
  add_cast_kernel<AT><<<(unsigned)((n / 4 + 255) / 256), 256, 0, st>>>(X, delta, Xc, n);
  ST2_LAUNCH_OK("add_cast");
  return ST2_OK;
}
template int launch_add_cast<float>(float*, const float*, float*, int64_t, cudaStream_t);
template int launch_add_cast<__nv_bfloat16>(float*, const __nv_bfloat16*, __nv_bfloat16*, int64_t, cudaStream_t);

// exact (erf) GELU in place: nn.GELU() of FeedForward (modules.py:476-482)
template <typename AT>
__global__ void __launch_bounds__(256) gelu_kernel(AT* __restrict__ x, int64_t n) {
  const int64_t i = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * 4;
  if (i >= n) return;
#pragma unroll
  for (int j = 0; j < 4; ++j) st_f(x + i + j, gelu_erf(ld_f(x + i + j)));
}
template <>
__global__ void __launch_bounds__(256) gelu_kernel<__nv_bfloat16>(__nv_bfloat16* __restrict__ x, int64_t n) {
  const int64_t i = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * 8;
  if (i >= n) return;
  uint4 raw = *reinterpret_cast<uint4*>(x + i);
  __nv_bfloat162* p = reinterpret_cast<__nv_bfloat162*>(&raw);
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    float2 v = __bfloat1622float2(p[j]);
    v.x = gelu_erf(v.x);
    v.y = gelu_erf(v.y);
    p[j] = __floats2bfloat162_rn(v.x, v.y);
  }
  *reinterpret_cast<uint4*>(x + i) = raw;
}
template <typename AT>
int launch_gelu(AT* x, int64_t n, cudaStream_t st) {
  const int per = sizeof(AT) == 2 ? 8 : 4;
  gelu_kernel<AT><<<(unsigned)((n / per + 255) / 256), 256, 0, st>>>(x, n);
  ST2_LAUNCH_OK("gelu");
  return ST2_OK;
}
template int launch_gelu<float>(float*, int64_t, cudaStream_t);
template int launch_gelu<__nv_bfloat16>(__nv_bfloat16*, int64_t, cudaStream_t);

// ------------------------------------------------------------------ attention
// softmax(q k^T / sqrt(D)) v per (item, head) (modules.py:512-526), D = 64.  fp32 parity path: one thread per query row with
// q and the output accumulator in registers; keys / values of the head are staged through shared memory in
// tiles of 64 rows (fp32) and read as warp broadcasts; online softmax in fp32.
template <typename AT>
__global__ void __launch_bounds__(128)
attention_kernel(const AT* __restrict__ Q, int ldq, const AT* __restrict__ KV, int ldkv, AT* __restrict__ O, int ldo,
                 int n, int mid, float scale) {
  constexpr int D = 64, KT = 64;
  __shared__ __align__(16) float ks[KT][D];
  __shared__ __align__(16) float vs[KT][D];
  const int h = blockIdx.y, bx = blockIdx.z;
  const int qi = blockIdx.x * 128 + threadIdx.x;
  const bool live = qi < n;
  const int64_t base = (int64_t)bx * n;
  float q[D], o[D];
  {
    const AT* qp = Q + (base + (live ? qi : 0)) * ldq + h * D;
#pragma unroll
    for (int d = 0; d < D; ++d) {
      q[d] = ld_f(qp + d) * scale;
      o[d] = 0.f;
    }
  }
  float m = -INFINITY, l = 0.f;
  for (int k0 = 0; k0 < n; k0 += KT) {
    const int kn = min(KT, n - k0);
    __syncthreads();
    for (int e = threadIdx.x; e < KT * D; e += 128) {
      const int r = e / D, d = e % D;
      float kv = 0.f, vv = 0.f;
      if (r < kn) {
        const AT* p = KV + (base + k0 + r) * ldkv + h * D + d;
        kv = ld_f(p);
        vv = ld_f(p + mid);
      }
      ks[r][d] = kv;
      vs[r][d] = vv;
    }
    __syncthreads();
    for (int j0 = 0; j0 < kn; j0 += 8) {
      float s[8];
      float mx = m;
#pragma unroll
      for (int jj = 0; jj < 8; ++jj) {
        const int j = min(j0 + jj, KT - 1);
        float a0 = 0.f, a1 = 0.f;
#pragma unroll
        for (int d = 0; d < D; d += 4) {
          const float4 kk = *reinterpret_cast<const float4*>(&ks[j][d]);
          a0 = fmaf(q[d], kk.x, a0);
          a1 = fmaf(q[d + 1], kk.y, a1);
          a0 = fmaf(q[d + 2], kk.z, a0);
          a1 = fmaf(q[d + 3], kk.w, a1);
        }
        s[jj] = j0 + jj < kn ? a0 + a1 : -INFINITY;
        mx = fmaxf(mx, s[jj]);
      }
      const float corr = __expf(m - mx);  // m = -inf on the first pass: exp(-inf) = 0
      m = mx;
      l *= corr;
#pragma unroll
      for (int d = 0; d < D; ++d) o[d] *= corr;
#pragma unroll
      for (int jj = 0; jj < 8; ++jj) {
        const int j = min(j0 + jj, KT - 1);
        const float p = __expf(s[jj] - mx);
        l += p;
#pragma unroll
        for (int d = 0; d < D; d += 4) {
          const float4 vv = *reinterpret_cast<const float4*>(&vs[j][d]);
          o[d] = fmaf(p, vv.x, o[d]);
          o[d + 1] = fmaf(p, vv.y, o[d + 1]);
          o[d + 2] = fmaf(p, vv.z, o[d + 2]);
          o[d + 3] = fmaf(p, vv.w, o[d + 3]);
        }
      }
    }
  }
  if (live) {
    const float inv = 1.f / l;
    AT* op = O + (base + qi) * ldo + h * D;
#pragma unroll
    for (int d = 0; d < D; ++d) st_f(op + d, o[d] * inv);
  }
}
// bf16 path: flash-attention form on the warp-level tensor-core MMA (mma.sync m16n8k16, bf16 x bf16 -> fp32).
// The attention core is 3 % of the sampler's FLOPs (n = 128 keys of 64 features per head); it is latency- and
// shared-memory-bound, so each warp keeps its 16 query rows, the 16 x 64 score tile and the 16 x 64 output
// accumulator in registers (scores never reach shared memory) and K / V tiles of 64 keys are staged with
// cp.async (rows padded to 144 bytes: conflict-free ldmatrix).  CTA = 4 warps = 64 queries of one (item, head).
__device__ __forceinline__ void mma_bf16_16816(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ void ldsm_x4(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];" : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}
__device__ __forceinline__ void ldsm_x4_t(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}
__device__ __forceinline__ uint32_t pack2_bf16(float lo, float hi) {
  __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&v);
}
__global__ void __launch_bounds__(128)
attention_mma_kernel(const __nv_bfloat16* __restrict__ Q, int ldq, const __nv_bfloat16* __restrict__ KV, int ldkv,
                     __nv_bfloat16* __restrict__ O, int ldo, int n, int mid, float scale_log2e) {
  constexpr int D = 64, KT = 64, LDS = 72;  // padded row: 144 bytes
  __shared__ __align__(16) __nv_bfloat16 ks[KT * LDS];
  __shared__ __align__(16) __nv_bfloat16 vs[KT * LDS];
  const int h = blockIdx.y, bx = blockIdx.z, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int g = lane >> 2, t = lane & 3;
  const int64_t base = (int64_t)bx * n;
  const int q0 = blockIdx.x * 64 + warp * 16;
  // A fragments of this warp's 16 query rows (rows g and g + 8), 4 k-steps of 16 features
  uint32_t qa[4][4];
  {
    const int r0 = min(q0 + g, n - 1), r1 = min(q0 + g + 8, n - 1);
    const __nv_bfloat16* p0 = Q + (base + r0) * ldq + h * D + 2 * t;
    const __nv_bfloat16* p1 = Q + (base + r1) * ldq + h * D + 2 * t;
#pragma unroll
    for (int kk = 0; kk < 4; ++kk) {
      qa[kk][0] = *reinterpret_cast<const uint32_t*>(p0 + kk * 16);
      qa[kk][1] = *reinterpret_cast<const uint32_t*>(p1 + kk * 16);
      qa[kk][2] = *reinterpret_cast<const uint32_t*>(p0 + kk * 16 + 8);
      qa[kk][3] = *reinterpret_cast<const uint32_t*>(p1 + kk * 16 + 8);
    }
  }
  float o[8][4];
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) o[i][j] = 0.f;
  float m0 = -INFINITY, m1 = -INFINITY, l0 = 0.f, l1 = 0.f;
  const uint32_t ks_u = (uint32_t)__cvta_generic_to_shared(ks), vs_u = (uint32_t)__cvta_generic_to_shared(vs);
  for (int k0 = 0; k0 < n; k0 += KT) {
    const int kn = min(KT, n - k0);
    __syncthreads();
    // stage K and V rows [k0, k0 + 64) of this head (16-byte chunks; rows past n are zero)
    for (int e = threadIdx.x; e < KT * 8; e += 128) {
      const int r = e >> 3, ch = e & 7;
      const uint32_t off = (uint32_t)(r * LDS + ch * 8) * 2u;
      if (r < kn) {
        const __nv_bfloat16* src = KV + (base + k0 + r) * ldkv + h * D + ch * 8;
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(ks_u + off), "l"(src));
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(vs_u + off), "l"(src + mid));
      } else {
        *reinterpret_cast<uint4*>(reinterpret_cast<uint8_t*>(ks) + off) = make_uint4(0u, 0u, 0u, 0u);
        *reinterpret_cast<uint4*>(reinterpret_cast<uint8_t*>(vs) + off) = make_uint4(0u, 0u, 0u, 0u);
      }
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    __syncthreads();
    // S = Q K^T: 8 blocks of 8 keys
    float sc[8][4];
#pragma unroll
    for (int nb = 0; nb < 8; ++nb)
#pragma unroll
      for (int j = 0; j < 4; ++j) sc[nb][j] = 0.f;
#pragma unroll
    for (int nb2 = 0; nb2 < 4; ++nb2) {  // 16 keys per ldmatrix.x4 pair
#pragma unroll
      for (int kk2 = 0; kk2 < 2; ++kk2) {  // 32 features per iteration
        // matrices: (keys nb2*16 + [0,8), d kk2*32 + [0,8)), (.., d +8), (.., d +16), (.., d +24)
        uint32_t b[4], c2[4];
        const int row = nb2 * 16 + (lane & 7), col = kk2 * 32 + (lane >> 3) * 8;
        ldsm_x4(ks_u + (uint32_t)(row * LDS + col) * 2u, b[0], b[1], b[2], b[3]);
        ldsm_x4(ks_u + (uint32_t)((row + 8) * LDS + col) * 2u, c2[0], c2[1], c2[2], c2[3]);
        mma_bf16_16816(sc[nb2 * 2], qa[kk2 * 2], b[0], b[1]);
        mma_bf16_16816(sc[nb2 * 2], qa[kk2 * 2 + 1], b[2], b[3]);
        mma_bf16_16816(sc[nb2 * 2 + 1], qa[kk2 * 2], c2[0], c2[1]);
        mma_bf16_16816(sc[nb2 * 2 + 1], qa[kk2 * 2 + 1], c2[2], c2[3]);
      }
    }
    // online softmax on rows g (c0, c1) and g + 8 (c2, c3); keys past n are masked
    float mx0 = m0, mx1 = m1;
#pragma unroll
    for (int nb = 0; nb < 8; ++nb) {
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int key = nb * 8 + 2 * t + (j & 1);
        const float v = key < kn ? sc[nb][j] * scale_log2e : -INFINITY;
        sc[nb][j] = v;
        if (j < 2) mx0 = fmaxf(mx0, v);
        else mx1 = fmaxf(mx1, v);
      }
    }
    mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 1));
    mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 2));
    mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 1));
    mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 2));
    const float cr0 = exp2f(m0 - mx0), cr1 = exp2f(m1 - mx1);
    m0 = mx0; m1 = mx1;
    l0 *= cr0; l1 *= cr1;
#pragma unroll
    for (int nb = 0; nb < 8; ++nb) {
      o[nb][0] *= cr0; o[nb][1] *= cr0; o[nb][2] *= cr1; o[nb][3] *= cr1;
    }
    uint32_t pa[4][4];  // P as A fragments: k-step j covers keys [16 j, 16 j + 16) = score blocks 2j, 2j + 1
#pragma unroll
    for (int nb = 0; nb < 8; ++nb) {
      const float p0 = exp2f(sc[nb][0] - mx0), p1 = exp2f(sc[nb][1] - mx0);
      const float p2 = exp2f(sc[nb][2] - mx1), p3 = exp2f(sc[nb][3] - mx1);
      l0 += p0 + p1;
      l1 += p2 + p3;
      pa[nb >> 1][(nb & 1) * 2] = pack2_bf16(p0, p1);
      pa[nb >> 1][(nb & 1) * 2 + 1] = pack2_bf16(p2, p3);
    }
    // O += P V: B[k = key][n = d] from V[key][d] through ldmatrix.trans
#pragma unroll
    for (int j = 0; j < 4; ++j) {      // 16 keys
#pragma unroll
      for (int db = 0; db < 4; ++db) {  // 16 features
        // matrices: (keys 16j + [0,8), d 16db + [0,8)), (keys +8, d), (keys, d + 8), (keys + 8, d + 8)
        uint32_t b0, b1, b2, b3;
        const int row = j * 16 + (lane & 7) + ((lane >> 3) & 1) * 8, col = db * 16 + (lane >> 4) * 8;
        ldsm_x4_t(vs_u + (uint32_t)(row * LDS + col) * 2u, b0, b1, b2, b3);
        mma_bf16_16816(o[db * 2], pa[j], b0, b1);
        mma_bf16_16816(o[db * 2 + 1], pa[j], b2, b3);
      }
    }
  }
  l0 += __shfl_xor_sync(0xffffffffu, l0, 1);
  l0 += __shfl_xor_sync(0xffffffffu, l0, 2);
  l1 += __shfl_xor_sync(0xffffffffu, l1, 1);
  l1 += __shfl_xor_sync(0xffffffffu, l1, 2);
  const float i0 = 1.f / l0, i1 = 1.f / l1;
  const int r0 = q0 + g, r1 = q0 + g + 8;
#pragma unroll
  for (int nb = 0; nb < 8; ++nb) {
    if (r0 < n) *reinterpret_cast<uint32_t*>(O + (base + r0) * ldo + h * D + nb * 8 + 2 * t) = pack2_bf16(o[nb][0] * i0, o[nb][1] * i0);
    if (r1 < n) *reinterpret_cast<uint32_t*>(O + (base + r1) * ldo + h * D + nb * 8 + 2 * t) = pack2_bf16(o[nb][2] * i1, o[nb][3] * i1);
  }
}

template <typename AT>
int launch_attention(const AT* Q, int ldq, const AT* KV, int ldkv, AT* O, int ldo, int Bx, int n, int heads, int head_dim,
                     cudaStream_t st) {
  ST2_REQUIRE(head_dim == 64, "attention kernel is built for head_features = 64");
  ST2_REQUIRE(Bx <= 65535 && heads <= 65535, "attention grid");
  const float scale = 1.f / sqrtf((float)head_dim);
  if constexpr (sizeof(AT) == 2) {
    ST2_REQUIRE(ldq % 8 == 0 && ldkv % 8 == 0 && ldo % 2 == 0, "attention: 16-byte aligned rows");
    dim3 grid(ceil_div(n, 64), heads, Bx);
    attention_mma_kernel<<<grid, 128, 0, st>>>(Q, ldq, KV, ldkv, O, ldo, n, heads * head_dim, scale * 1.4426950408889634f);
    ST2_LAUNCH_OK("attention_mma");
    return ST2_OK;
  } else {
    dim3 grid(ceil_div(n, 128), heads, Bx);
    attention_kernel<AT><<<grid, 128, 0, st>>>(Q, ldq, KV, ldkv, O, ldo, n, heads * head_dim, scale);
    ST2_LAUNCH_OK("attention");
    return ST2_OK;
  }
}
template int launch_attention<float>(const float*, int, const float*, int, float*, int, int, int, int, int, cudaStream_t);
template int launch_attention<__nv_bfloat16>(const __nv_bfloat16*, int, const __nv_bfloat16*, int, __nv_bfloat16*, int, int,
                                             int, int, int, cudaStream_t);

// out[bx][c] = mean_t (X[bx, t, c] + delta[bx, t, c])   (modules.py:157)
template <typename AT>
__global__ void __launch_bounds__(256)
token_mean_kernel(const float* __restrict__ X, const AT* __restrict__ delta, float* __restrict__ out, int n, int F) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x, bx = blockIdx.y;
  if (c >= F) return;
  const float* xp = X + (int64_t)bx * n * F + c;
  const AT* dp = delta + (int64_t)bx * n * F + c;
  float s = 0.f;
  for (int t = 0; t < n; ++t) s += xp[(int64_t)t * F] + ld_f(dp + (int64_t)t * F);
  out[(int64_t)bx * F + c] = s / (float)n;
}
template <typename AT>
int launch_token_mean(const float* X, const AT* delta, float* out, int Bx, int n, int F, cudaStream_t st) {
  token_mean_kernel<AT><<<dim3(ceil_div(F, 256), Bx), 256, 0, st>>>(X, delta, out, n, F);
  ST2_LAUNCH_OK("token_mean");
  return ST2_OK;
}
template int launch_token_mean<float>(const float*, const float*, float*, int, int, int, cudaStream_t);
template int launch_token_mean<__nv_bfloat16>(const float*, const __nv_bfloat16*, float*, int, int, int, cudaStream_t);

// ------------------------------------------------------------------ sampler arithmetic on [B][C]
// classifier-free guidance (modules.py:178-184), KDiffusion skip connection (sampler.py:204-206) and one
// ADPM2 half step (sampler.py:500-510):
//   net = pm + (p - pm) * cfg_scale            (p = pred[b], pm = pred[B + b]; net = p without guidance)
//   den = c_skip * x_eval + c_out * net
//   x_new = x_base + ((x_eval - den) / sigma_eval) * dt + noise * sigma_up        (optionally clamped)
__global__ void __launch_bounds__(256)
sampler_step_kernel(const float* __restrict__ pred, const float* __restrict__ x_eval, const float* __restrict__ x_base,
                    const float* __restrict__ noise, uint64_t seed, uint32_t step, int64_t item0, float* __restrict__ x_new,
                    float* __restrict__ den_out, int B, int C, int cfg, float cfg_scale, float c_skip, float c_out,
                    float sigma_eval, float dt, float sigma_up, int clamp) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= B * C) return;
  float net = pred[i];
  if (cfg) {
    const float pm = pred[(int64_t)B * C + i];
    net = pm + (net - pm) * cfg_scale;
  }
  const float xe = x_eval[i];
  const float den = c_skip * xe + c_out * net;
  if (den_out) den_out[i] = den;
  if (!x_new) return;
  const float d = (xe - den) / sigma_eval;
  float v = x_base[i] + d * dt;
  if (sigma_up != 0.f) {
    float z;
    if (noise) {
      z = noise[i];
    } else {  // Philox4x32-10 keyed by (seed), counter (global item, channel pair, step)
      const int b = i / C, c = i % C;
      uint32_t r[4];
      philox4x32((uint32_t)(item0 + b), (uint32_t)((item0 + b) >> 32), (uint32_t)(c >> 1), 0x53414d50u ^ step,
                 (uint32_t)seed, (uint32_t)(seed >> 32), r);
      float z0, z1;
      box_muller(r[0], r[1], z0, z1);
      z = (c & 1) ? z1 : z0;
    }
    v += z * sigma_up;
  }
  if (clamp) v = fminf(fmaxf(v, -1.f), 1.f);
  x_new[i] = v;
}
int launch_sampler_step(const float* pred, const float* x_eval, const float* x_base, const float* noise, uint64_t seed,
                        uint32_t step, int64_t item0, float* x_new, float* den_out, int B, int C, int cfg, float cfg_scale,
                        float c_skip, float c_out, float sigma_eval, float dt, float sigma_up, int clamp, cudaStream_t st) {
  sampler_step_kernel<<<ceil_div(B * C, 256), 256, 0, st>>>(pred, x_eval, x_base, noise, seed, step, item0, x_new, den_out,
                                                            B, C, cfg, cfg_scale, c_skip, c_out, sigma_eval, dt, sigma_up,
                                                            clamp);
  ST2_LAUNCH_OK("sampler_step");
  return ST2_OK;
}

__global__ void scale_kernel(const float* __restrict__ x, float a, float* __restrict__ y, int n) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) y[i] = a * x[i];
}
int launch_scale(const float* x, float a, float* y, int n, cudaStream_t st) {
  scale_kernel<<<ceil_div(n, 256), 256, 0, st>>>(x, a, y, n);
  ST2_LAUNCH_OK("scale");
  return ST2_OK;
}

}  // namespace st2
