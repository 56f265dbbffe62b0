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

// ------------------------------------------------------------------ small dense layers (fp32)
// y[r][o] = act(bias[o] + sum_i W[o][i] * (x[r][i] + xb[i])), one warp per output feature and 8 rows
template <int ROWS>
__global__ void __launch_bounds__(256)
small_linear_kernel(const float* __restrict__ x, int ldx, const float* __restrict__ xb, const float* __restrict__ W,
                    const float* __restrict__ bias, float* __restrict__ y, int ldy, int R, int I, int O, int act) {
  const int o = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  const int r0 = blockIdx.y * ROWS;
  if (o >= O) return;
  float acc[ROWS];
#pragma unroll
  for (int j = 0; j < ROWS; ++j) acc[j] = 0.f;
  const float* wr = W + (int64_t)o * I;
  for (int i = lane; i < I; i += 32) {
    const float w = __ldg(wr + i);
    const float add = xb ? __ldg(xb + i) : 0.f;
#pragma unroll
    for (int j = 0; j < ROWS; ++j) {
      const int r = min(r0 + j, R - 1);
      acc[j] = fmaf(w, __ldg(x + (int64_t)r * ldx + i) + add, acc[j]);
    }
  }
#pragma unroll
  for (int j = 0; j < ROWS; ++j) acc[j] = warp_sum(acc[j]);
  if (lane == 0) {
    const float b = bias ? bias[o] : 0.f;
#pragma unroll
    for (int j = 0; j < ROWS; ++j)
      if (r0 + j < R) {
        const float v = acc[j] + b;
        y[(int64_t)(r0 + j) * ldy + o] = act ? gelu_erf(v) : v;
      }
  }
}
int launch_small_linear(const float* x, int ldx, const float* xb, const float* W, const float* bias, float* y, int ldy,
                        int R, int I, int O, int gelu, cudaStream_t st) {
  dim3 grid(ceil_div(O, 8), ceil_div(R, 8));
  small_linear_kernel<8><<<grid, 256, 0, st>>>(x, ldx, xb, W, bias, y, ldy, R, I, O, gelu);
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
template <typename AT>
__global__ void __launch_bounds__(256)
adaln2_kernel(float* __restrict__ X, const AT* __restrict__ delta, const float* __restrict__ mapping, int64_t map_bstride,
              const float* __restrict__ gb, int64_t gb_bstride, AT* __restrict__ nq, AT* __restrict__ nk, int64_t rows,
              int n, int B, int F) {
  const int64_t row = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (row >= rows) return;
  const int b = (int)((row / n) % B);
  float* xr = X + row * F;
  const AT* dr = delta ? delta + row * F : nullptr;
  const float* mp = mapping + (int64_t)b * map_bstride;
  float s = 0.f;
  for (int c = lane; c < F; c += 32) {
    float v = xr[c] + mp[c];
    if (dr) v += ld_f(dr + c);
    xr[c] = v;
    s += v;
  }
  const float mean = warp_sum(s) / (float)F;
  float q = 0.f;
  for (int c = lane; c < F; c += 32) {
    const float d = xr[c] - mean;
    q = fmaf(d, d, q);
  }
  const float rstd = rsqrtf(warp_sum(q) / (float)F + 1e-5f);
  const float* g = gb + (int64_t)b * gb_bstride;
  AT* q_o = nq + row * F;
  AT* k_o = nk + row * F;
  for (int c = lane; c < F; c += 32) {
    const float xh = (xr[c] - mean) * rstd;
    st_f(q_o + c, fmaf(1.f + g[c], xh, g[F + c]));
    st_f(k_o + c, fmaf(1.f + g[2 * F + c], xh, g[3 * F + c]));
  }
}
template <typename AT>
int launch_adaln2(float* X, const AT* delta, const float* mapping, int64_t map_bstride, const float* gb, int64_t gb_bstride,
                  AT* nq, AT* nk, int64_t rows, int n, int B, int F, cudaStream_t st) {
  adaln2_kernel<AT><<<(unsigned)((rows + 7) / 8), 256, 0, st>>>(X, delta, mapping, map_bstride, gb, gb_bstride, nq, nk, rows, n, B, F);
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
  v.x += ld_f(delta + i); v.y += ld_f(delta + i + 1); v.z += ld_f(delta + i + 2); v.w += ld_f(delta + i + 3);
  *reinterpret_cast<float4*>(X + i) = v;
  st_f(Xc + i, v.x); st_f(Xc + i + 1, v.y); st_f(Xc + i + 2, v.z); st_f(Xc + i + 3, v.w);
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
// softmax(q k^T / sqrt(D)) v per (item, head) (modules.py:512-526), D = 64.  One thread per query row with
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
template <typename AT>
int launch_attention(const AT* Q, int ldq, const AT* KV, int ldkv, AT* O, int ldo, int Bx, int n, int heads, int head_dim,
                     cudaStream_t st) {
  ST2_REQUIRE(head_dim == 64, "attention kernel is built for head_features = 64");
  ST2_REQUIRE(Bx <= 65535 && heads <= 65535, "attention grid");
  dim3 grid(ceil_div(n, 128), heads, Bx);
  attention_kernel<AT><<<grid, 128, 0, st>>>(Q, ldq, KV, ldkv, O, ldo, n, heads * head_dim, 1.f / sqrtf((float)head_dim));
  ST2_LAUNCH_OK("attention");
  return ST2_OK;
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
