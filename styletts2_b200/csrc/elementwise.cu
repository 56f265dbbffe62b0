// elementwise.cu — the small memory-bound kernels around the convolutions:
//   style_fc        all AdaIN1d.fc layers as one [sum 2C, style] x [style, B] product (istftnet.py:19,22)
//   stats_finalize  InstanceNorm1d statistics (biased var, eps 1e-5, istftnet.py:18) from per-tile
//                   partial sums -> per-(b,c) (scale, shift) with the style affine folded in (:25)
//   colstats        partial sums of a channels-last tensor that no conv epilogue produced
//   prep_inputs     NCL -> channels-last transpose of asr, F0_conv / N_conv (istftnet.py:510-513)
//   pool_up2        AdaIN + LeakyReLU(0.2) + depthwise ConvTranspose1d(k3,s2,p1,op1) (:441-443, :423)
//   reflect_row0    row 0 of ReflectionPad1d((1,0)) after the last upsampling (:364-367)
//   layout helpers  NCL <-> channels-last for the injected / exported `har`
#include "common.cuh"

namespace st2 {

// ------------------------------------------------------------------------- style fc
// lane = output row j (a warp owns 32 consecutive rows and streams its 32 weight rows with 16-byte loads: every
// 128-byte line is used completely over 8 iterations), the accumulators of up to SFC_NB batch items live in
// registers and their style vectors are read from shared memory as warp-wide broadcasts: 1 LDG + SFC_NB LDS
// per 4 * SFC_NB FMAs, and every weight is read from HBM exactly once per group of SFC_NB items.
constexpr int SFC_NB = 32;
__global__ void __launch_bounds__(256)
style_fc_kernel(const float* __restrict__ s, const float* __restrict__ W,
                const float* __restrict__ bvec, float* __restrict__ h, int B, int style, int rows) {
  extern __shared__ __align__(16) float ssm[];  // [SFC_NB][style]
  const int b0 = blockIdx.y * SFC_NB;
  const int nb = min(SFC_NB, B - b0);
  for (int i = threadIdx.x; i < SFC_NB * style; i += blockDim.x)
    ssm[i] = i < nb * style ? s[(int64_t)b0 * style + i] : 0.f;
  __syncthreads();
  const int j = blockIdx.x * 256 + threadIdx.x;
  if (j >= rows) return;
  const float4* wr = reinterpret_cast<const float4*>(W + (int64_t)j * style);
  float acc[SFC_NB];
#pragma unroll
  for (int b = 0; b < SFC_NB; ++b) acc[b] = 0.f;
  for (int i4 = 0; i4 < style / 4; ++i4) {
    const float4 w = __ldg(wr + i4);
#pragma unroll
    for (int b = 0; b < SFC_NB; ++b) {
      const float4 v = *reinterpret_cast<const float4*>(ssm + b * style + 4 * i4);
      acc[b] = fmaf(w.x, v.x, acc[b]);
      acc[b] = fmaf(w.y, v.y, acc[b]);
      acc[b] = fmaf(w.z, v.z, acc[b]);
      acc[b] = fmaf(w.w, v.w, acc[b]);
    }
  }
  const float bj = bvec[j];
#pragma unroll
  for (int b = 0; b < SFC_NB; ++b)
    if (b < nb) h[(int64_t)(b0 + b) * rows + j] = acc[b] + bj;
}

int launch_style_fc(const float* s, const float* W, const float* bvec, float* h, int B, int style,
                    int rows, cudaStream_t st) {
  dim3 grid(ceil_div(rows, 256), ceil_div(B, SFC_NB));
  size_t smem = (size_t)SFC_NB * style * sizeof(float);
  ST2_REQUIRE(smem <= 48 * 1024 && style % 4 == 0, "style_dim must be a multiple of 4, at most 384");
  style_fc_kernel<<<grid, 256, smem, st>>>(s, W, bvec, h, B, style, rows);
  ST2_LAUNCH_OK("style_fc");
  return ST2_OK;
}

// ------------------------------------------------------------------- stats finalize
// block = 32 channels of one batch item, 8 warps: warp w reduces tiles w, w+8, ... (lane = channel,
// so every load is a 256-byte coalesced row of (sum, sumsq) pairs), then the 8 warp partials are
// combined in a fixed order.  Partials are [B][n_tiles][n_cols][2]; channel c owns columns c, c+C,
// ... (n_cols = phases*C for a polyphase transposed conv).  Sums run in fp64 in a fixed order, so
// the result is deterministic and independent of how the producing grid was scheduled.
constexpr int SF_WARPS = 32;
// up to 4 norm layers of the same shape per launch (blockIdx.z): the three resblocks of a generator stage in lock step,
// the four decode blocks' norm1 over the constant concat columns
struct SfJobs {
  const float* partials[4];
  const float* hg[4];
  float* ss[4];
};
__global__ void __launch_bounds__(SF_WARPS * 32)
stats_finalize_kernel(const SfJobs jobs, int n_tiles, int n_cols, int C,
                      double inv_count, int h_stride, int Cn,
                      int ss_stride, int c_off, int reps) {
  __shared__ double red[SF_WARPS][32][2];
  pdl_wait_then_trigger();
  const float* __restrict__ partials = jobs.partials[blockIdx.z];
  const float* __restrict__ hg = jobs.hg[blockIdx.z];
  float* __restrict__ ss = jobs.ss[blockIdx.z];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int c = blockIdx.x * 32 + lane;
  const int b = blockIdx.y;
  const float2* pp = reinterpret_cast<const float2*>(partials) + (int64_t)b * n_tiles * n_cols;
  double s = 0.0, q = 0.0;
  if (c < C) {
    for (int col = c; col < n_cols; col += C) {
      int t = warp;
      for (; t + 7 * SF_WARPS < n_tiles; t += 8 * SF_WARPS) {  // 8 independent loads in flight
        float2 v[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) v[u] = pp[(int64_t)(t + u * SF_WARPS) * n_cols + col];
#pragma unroll
        for (int u = 0; u < 8; ++u) {
          s += (double)v[u].x;
          q += (double)v[u].y;
        }
      }
      for (; t + 3 * SF_WARPS < n_tiles; t += 4 * SF_WARPS) {  // 4 independent loads in flight
        const float2 v0 = pp[(int64_t)t * n_cols + col];
        const float2 v1 = pp[(int64_t)(t + SF_WARPS) * n_cols + col];
        const float2 v2 = pp[(int64_t)(t + 2 * SF_WARPS) * n_cols + col];
        const float2 v3 = pp[(int64_t)(t + 3 * SF_WARPS) * n_cols + col];
        s += (double)v0.x; q += (double)v0.y;
        s += (double)v1.x; q += (double)v1.y;
        s += (double)v2.x; q += (double)v2.y;
        s += (double)v3.x; q += (double)v3.y;
      }
      for (; t < n_tiles; t += SF_WARPS) {
        const float2 v = pp[(int64_t)t * n_cols + col];
        s += (double)v.x; q += (double)v.y;
      }
    }
  }
  red[warp][lane][0] = s;
  red[warp][lane][1] = q;
  __syncthreads();
  if (warp != 0 || c >= C) return;
  s = 0.0; q = 0.0;
#pragma unroll
  for (int w = 0; w < SF_WARPS; ++w) {
    s += red[w][lane][0];
    q += red[w][lane][1];
  }
  const double mean = s * inv_count;
  double var = q * inv_count - mean * mean;
  if (var < 0.0) var = 0.0;
  const double rstd = 1.0 / sqrt(var + 1e-5);
  const double gamma = hg[(int64_t)b * h_stride + c_off + c];
  const double beta = hg[(int64_t)b * h_stride + Cn + c_off + c];
  const double scale = (1.0 + gamma) * rstd;
  // reps > 1: the consumer reads the tensor as rows of reps*C columns (row-pair view of a
  // 32-channel stage), so the pair is repeated at c, c + C, ...
  for (int r = 0; r < reps; ++r) {
    float* o = ss + ((int64_t)b * ss_stride + c_off + c + r * C) * 2;
    o[0] = (float)scale;
    o[1] = (float)(beta - mean * scale);
  }
}

// hg points at this norm layer's gamma block inside the concatenated fc output; the layer
// normalises Cn channels, of which [c_off, c_off + C) are finalised by this launch.
int launch_stats_finalize(const float* partials, int n_tiles, int n_cols, int C, int64_t count,
                          const float* hg, int h_stride, int Cn, float* ss, int ss_stride,
                          int c_off, int B, cudaStream_t st, int reps) {
  SfJobs jobs = {};
  jobs.partials[0] = partials; jobs.hg[0] = hg; jobs.ss[0] = ss;
  dim3 grid(ceil_div(C, 32), B, 1);
  ST2_CUDA_OK(launch_pdl(stats_finalize_kernel, grid, dim3(SF_WARPS * 32), 0, st, jobs, n_tiles, n_cols, C,
                         1.0 / (double)count, h_stride, Cn, ss_stride, c_off, reps));
  ST2_LAUNCH_OK("stats_finalize");
  return ST2_OK;
}
// n (<= 4) norm layers with the same shapes (tiles, columns, channels, count) in one launch
int launch_stats_finalize_multi(const float* const* partials, int n, int n_tiles, int n_cols, int C, int64_t count,
                                const float* const* hg, int h_stride, int Cn, float* const* ss, int ss_stride,
                                int c_off, int B, cudaStream_t st, int reps) {
  ST2_REQUIRE(n >= 1 && n <= 4, "stats_finalize_multi: 1..4 layers");
  SfJobs jobs = {};
  for (int j = 0; j < n; ++j) {
    jobs.partials[j] = partials[j];
    jobs.hg[j] = hg[j];
    jobs.ss[j] = ss[j];
  }
  dim3 grid(ceil_div(C, 32), B, n);
  ST2_CUDA_OK(launch_pdl(stats_finalize_kernel, grid, dim3(SF_WARPS * 32), 0, st, jobs, n_tiles, n_cols, C,
                         1.0 / (double)count, h_stride, Cn, ss_stride, c_off, reps));
  ST2_LAUNCH_OK("stats_finalize");
  return ST2_OK;
}

// ------------------------------------------------------------------------- colstats
constexpr int CS_ROWS = 256;
int colstats_tiles(int L) { return ceil_div(L, CS_ROWS); }

// block = 64 columns x CS_GROUPS row groups of one 256-row tile: 16 rows per thread, 8 loads in flight (the kernel
// is latency-bound: with 4 row groups and a rolled loop a 200-row tile took 11 us)
constexpr int CS_GROUPS = 16;
template <typename T>
__global__ void __launch_bounds__(64 * CS_GROUPS)
colstats_kernel(const T* __restrict__ x, int64_t bstride, int ld, int L, int c0, int ncols,
                float* __restrict__ partials, int n_tiles) {
  __shared__ float red[2][CS_GROUPS][64];
  const int tx = threadIdx.x & 63, ty = threadIdx.x >> 6;
  const int tile = blockIdx.x, b = blockIdx.z;
  const int col = blockIdx.y * 64 + tx;
  float s = 0.f, q = 0.f;
  if (col < ncols) {
    const T* xb = x + (int64_t)b * bstride + c0 + col;
    const int r1 = min(L, (tile + 1) * CS_ROWS);
    int r = tile * CS_ROWS + ty;
    for (; r + 7 * CS_GROUPS < r1; r += 8 * CS_GROUPS) {
      float v[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) v[u] = ld_f<T>(xb + (int64_t)(r + u * CS_GROUPS) * ld);
#pragma unroll
      for (int u = 0; u < 8; ++u) { s += v[u]; q = fmaf(v[u], v[u], q); }
    }
    for (; r < r1; r += CS_GROUPS) {
      const float v = ld_f<T>(xb + (int64_t)r * ld);
      s += v;
      q = fmaf(v, v, q);
    }
  }
  red[0][ty][tx] = s;
  red[1][ty][tx] = q;
  __syncthreads();
  if (ty == 0 && col < ncols) {
    float a = 0.f, a2 = 0.f;
#pragma unroll
    for (int g = 0; g < CS_GROUPS; ++g) {  // fixed order: deterministic
      a += red[0][g][tx];
      a2 += red[1][g][tx];
    }
    float* o = partials + (((int64_t)b * n_tiles + tile) * ncols + col) * 2;
    o[0] = a;
    o[1] = a2;
  }
}

template <typename T>
int launch_colstats(const T* x, int64_t bstride, int ld, int L, int c0, int ncols, float* partials,
                    int B, cudaStream_t st, int n_tiles_total) {
  const int nt = colstats_tiles(L);  // n_tiles_total > nt: the partial buffer has extra tiles other kernels fill
  dim3 grid(nt, ceil_div(ncols, 64), B);
  colstats_kernel<T><<<grid, 64 * CS_GROUPS, 0, st>>>(x, bstride, ld, L, c0, ncols, partials, n_tiles_total > nt ? n_tiles_total : nt);
  ST2_LAUNCH_OK("colstats");
  return ST2_OK;
}
template int launch_colstats<float>(const float*, int64_t, int, int, int, int, float*, int, cudaStream_t, int);
template int launch_colstats<__nv_bfloat16>(const __nv_bfloat16*, int64_t, int, int, int, int, float*,
                                            int, cudaStream_t, int);

// ---------------------------------------------------------------------- prep inputs
// grid (ceil(T2/32), ceil(ld0/32), B), block (32, 8).  f0w / nw = {w0, w1, w2, bias} (folded).
template <typename T>
__global__ void __launch_bounds__(256)
prep_inputs_kernel(const float* __restrict__ asr, const float* __restrict__ f0,
                   const float* __restrict__ nn, const float* __restrict__ f0w,
                   const float* __restrict__ nw, T* __restrict__ x0, int ld0, T* __restrict__ ca,
                   T* __restrict__ cb, int ldc, int c_f0, int dim_in, int T2,
                   const int* __restrict__ align_idx, int align_len) {
  __shared__ float tile[32][33];
  const int b = blockIdx.z;
  const int t0 = blockIdx.x * 32, ch0 = blockIdx.y * 32;
  const int tx = threadIdx.x, ty = threadIdx.y;
  // read: channel rows, time contiguous
  for (int i = ty; i < 32; i += 8) {
    const int ch = ch0 + i, t = t0 + tx;
    float v = 0.f;
    if (t < T2) {
      if (ch < dim_in) {
        if (align_idx) {  // asr = t_en @ one-hot alignment, fused: frame t reads token align_idx[b][t]
          const int tok = align_idx[(int64_t)b * T2 + t];
          v = (tok >= 0 && tok < align_len) ? asr[((int64_t)b * dim_in + ch) * align_len + tok] : 0.f;
        } else {
          v = asr[((int64_t)b * dim_in + ch) * T2 + t];
        }
      } else if (ch == dim_in || ch == dim_in + 1) {
        const float* src = (ch == dim_in ? f0 : nn) + (int64_t)b * 2 * T2;
        const float* w = ch == dim_in ? f0w : nw;
        float a = w[3];
#pragma unroll
        for (int j = 0; j < 3; ++j) {
          const int ti = 2 * t + j - 1;
          if (ti >= 0 && ti < 2 * T2) a = fmaf(w[j], src[ti], a);
        }
        v = a;
      }
    }
    tile[i][tx] = v;
  }
  __syncthreads();
  // write: time rows, channels contiguous
  for (int i = ty; i < 32; i += 8) {
    const int t = t0 + i, ch = ch0 + tx;
    if (t >= T2 || ch >= ld0) continue;
    const float v = tile[tx][i];
    st_f<T>(x0 + ((int64_t)b * T2 + t) * ld0 + ch, v);
    if (ch >= dim_in) {  // F0 / N / zero padding of the concat buffers
      const int cc = c_f0 + (ch - dim_in);
      if (cc < ldc) {
        st_f<T>(ca + ((int64_t)b * T2 + t) * ldc + cc, v);
        st_f<T>(cb + ((int64_t)b * T2 + t) * ldc + cc, v);
      }
    }
  }
}

template <typename T>
int launch_prep_inputs(const float* asr, const float* f0, const float* n, const float* f0w,
                       const float* nw, T* x0, int ld0, T* ca, T* cb, int ldc, int c_f0, int B,
                       int dim_in, int T2, cudaStream_t st, const int* align_idx, int align_len) {
  ST2_REQUIRE(ld0 - dim_in >= ldc - c_f0, "prep_inputs: x0 padding must cover the concat padding");
  dim3 grid(ceil_div(T2, 32), ceil_div(ld0, 32), B), block(32, 8);
  prep_inputs_kernel<T><<<grid, block, 0, st>>>(asr, f0, n, f0w, nw, x0, ld0, ca, cb, ldc, c_f0, dim_in, T2,
                                                align_idx, align_len);
  ST2_LAUNCH_OK("prep_inputs");
  return ST2_OK;
}
template int launch_prep_inputs<float>(const float*, const float*, const float*, const float*,
                                       const float*, float*, int, float*, float*, int, int, int, int,
                                       int, cudaStream_t, const int*, int);
template int launch_prep_inputs<__nv_bfloat16>(const float*, const float*, const float*, const float*,
                                               const float*, __nv_bfloat16*, int, __nv_bfloat16*,
                                               __nv_bfloat16*, int, int, int, int, int, cudaStream_t,
                                               const int*, int);

// ------------------------------------------------------------------------ alignment
// The reference expands token-rate features to frame rate with a one-hot matrix built on the host
// (Demo/Inference_LJSpeech.ipynb cell 17: pred_aln_trg[i, c : c + dur[i]] = 1; asr = t_en @ pred_aln_trg).
// Here: idx[b][f] = the token whose duration interval contains frame f (-1 past the item's total: the
// matmul's zero columns); shift = HiFi-GAN's one-frame right shift (Inference_LibriTTS.ipynb cell 16:
// asr_new[..., 0] = asr[..., 0]; asr_new[..., 1:] = asr[..., :-1]).  One block per item.
__global__ void __launch_bounds__(256)
alignment_idx_kernel(const int* __restrict__ dur, int Ltext, int frames, int shift, int* __restrict__ idx,
                     int* __restrict__ totals) {
  __shared__ int scan[256];
  __shared__ int carry_s;
  const int b = blockIdx.x, tid = threadIdx.x;
  const int* d = dur + (int64_t)b * Ltext;
  int* out = idx + (int64_t)b * frames;
  for (int f = tid; f < frames; f += 256) out[f] = -1;
  if (tid == 0) carry_s = 0;
  __syncthreads();
  for (int base = 0; base < Ltext; base += 256) {
    const int i = base + tid;
    const int v = i < Ltext ? max(d[i], 0) : 0;
    scan[tid] = v;
    __syncthreads();
    for (int o = 1; o < 256; o <<= 1) {  // inclusive Hillis-Steele scan of one 256-token chunk
      const int t = tid >= o ? scan[tid - o] : 0;
      __syncthreads();
      scan[tid] += t;
      __syncthreads();
    }
    const int carry = carry_s;
    const int end = carry + scan[tid], start = end - v;
    for (int f = start; f < end; ++f) {
      const int g = f + (shift ? 1 : 0);  // frame g shows what frame f held before the shift
      if (g < frames) out[g] = i;
      if (shift && f == 0) out[0] = i;
    }
    __syncthreads();
    if (tid == 255) carry_s = carry + scan[255];
    __syncthreads();
  }
  if (tid == 0 && totals) totals[b] = carry_s;
}
int launch_alignment_idx(const int* dur, int B, int Ltext, int frames, int shift, int* idx, int* totals,
                         cudaStream_t st) {
  alignment_idx_kernel<<<B, 256, 0, st>>>(dur, Ltext, frames, shift, idx, totals);
  ST2_LAUNCH_OK("alignment_idx");
  return ST2_OK;
}
// out[b][c][f] = idx[b][f] >= 0 ? x[b][c][idx[b][f]] : 0      (x NCL [B][C][Ltext], out NCL [B][C][frames])
__global__ void expand_alignment_kernel(const float* __restrict__ x, const int* __restrict__ idx, int C, int Ltext,
                                        int frames, float* __restrict__ out) {
  const int f = blockIdx.x * blockDim.x + threadIdx.x, c = blockIdx.y, b = blockIdx.z;
  if (f >= frames) return;
  const int tok = idx[(int64_t)b * frames + f];
  out[((int64_t)b * C + c) * frames + f] = (tok >= 0 && tok < Ltext) ? x[((int64_t)b * C + c) * Ltext + tok] : 0.f;
}
int launch_expand_alignment(const float* x, const int* idx, int B, int C, int Ltext, int frames, float* out,
                            cudaStream_t st) {
  dim3 grid(ceil_div(frames, 256), C, B);
  expand_alignment_kernel<<<grid, 256, 0, st>>>(x, idx, C, Ltext, frames, out);
  ST2_LAUNCH_OK("expand_alignment");
  return ST2_OK;
}

// ------------------------------------------------------------------------- norm + act
// y = lrelu(x * scale[b,c] + shift[b,c]) on a bf16 channels-last tensor, 8 channels per thread.
// Used where one activation tile feeds many N tiles (N >= 512: the decode / encode blocks and the
// F0/N predictor): applying AdaIN + LeakyReLU once here is cheaper than once per N tile in the conv
// prologue, which then degenerates to plain cp.async staging.
constexpr int NA_ROWS = 8;
__global__ void __launch_bounds__(256)
norm_act_bf16_kernel(const __nv_bfloat16* __restrict__ x, int64_t x_bstride, int ldx, const float* __restrict__ ss,
                     int ss_stride, float slope, __nv_bfloat16* __restrict__ y, int64_t y_bstride, int ldy,
                     int L, int C8) {
  const int g = blockIdx.x * blockDim.x + threadIdx.x;  // group of 8 channels
  const int b = blockIdx.z;
  if (g >= C8) return;
  float sc[8], sh[8];
  const float4* sp = reinterpret_cast<const float4*>(ss + ((int64_t)b * ss_stride + g * 8) * 2);
#pragma unroll
  for (int e = 0; e < 4; ++e) {
    const float4 t = __ldg(sp + e);
    sc[2 * e] = t.x; sh[2 * e] = t.y; sc[2 * e + 1] = t.z; sh[2 * e + 1] = t.w;
  }
  const int r0 = blockIdx.y * NA_ROWS;
#pragma unroll
  for (int i = 0; i < NA_ROWS; ++i) {
    const int r = r0 + i;
    if (r >= L) break;
    const uint4 raw = *reinterpret_cast<const uint4*>(x + (int64_t)b * x_bstride + (int64_t)r * ldx + g * 8);
    const uint32_t w[4] = {raw.x, raw.y, raw.z, raw.w};
    uint32_t o[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      float lo = fmaf(__uint_as_float(w[e] << 16), sc[2 * e], sh[2 * e]);
      float hi = fmaf(__uint_as_float(w[e] & 0xFFFF0000u), sc[2 * e + 1], sh[2 * e + 1]);
      lo = fmaxf(lo, lo * slope);
      hi = fmaxf(hi, hi * slope);
      const __nv_bfloat162 pk = __floats2bfloat162_rn(lo, hi);
      o[e] = *reinterpret_cast<const uint32_t*>(&pk);
    }
    *reinterpret_cast<uint4*>(y + (int64_t)b * y_bstride + (int64_t)r * ldy + g * 8) = make_uint4(o[0], o[1], o[2], o[3]);
  }
}
int launch_norm_act_bf16(const __nv_bfloat16* x, int64_t x_bstride, int ldx, const float* ss, int ss_stride,
                         float slope, __nv_bfloat16* y, int64_t y_bstride, int ldy, int B, int L, int C,
                         cudaStream_t st) {
  ST2_REQUIRE(C % 8 == 0 && ldx % 8 == 0 && ldy % 8 == 0 && slope >= 0.f && slope <= 1.f, "norm_act: shape");
  const int C8 = C / 8, threads = C8 < 128 ? 64 : 128;
  dim3 grid(ceil_div(C8, threads), ceil_div(L, NA_ROWS), B);
  norm_act_bf16_kernel<<<grid, threads, 0, st>>>(x, x_bstride, ldx, ss, ss_stride, slope, y, y_bstride, ldy, L, C8);
  ST2_LAUNCH_OK("norm_act");
  return ST2_OK;
}

// -------------------------------------------------------------------------- pool x2
// y[2q]   = w[c][1] * a[q] + bias[c]
// y[2q+1] = w[c][2] * a[q] + w[c][0] * a[q+1] + bias[c],   a = lrelu(x*scale+shift), a[Lin] = 0
// (ConvTranspose1d: y[t] = sum_i a[i] w[t + 1 - 2i], 0 <= t+1-2i < 3.)
template <typename T>
__global__ void __launch_bounds__(256)
pool_up2_kernel(const T* __restrict__ x, int64_t x_bstride, int ldx, const float* __restrict__ ss,
                int ss_stride, float slope, const float* __restrict__ w3,
                const float* __restrict__ bias, T* __restrict__ y, int64_t y_bstride, int ldy,
                int Lin, int C, int Cpad) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  const int q = blockIdx.y, b = blockIdx.z;
  if (c >= Cpad) return;
  T* y0 = y + (int64_t)b * y_bstride + (int64_t)(2 * q) * ldy + c;
  if (c >= C) {
    st_f<T>(y0, 0.f);
    st_f<T>(y0 + ldy, 0.f);
    return;
  }
  const float sc = ss[((int64_t)b * ss_stride + c) * 2], sh = ss[((int64_t)b * ss_stride + c) * 2 + 1];
  const T* xp = x + (int64_t)b * x_bstride + (int64_t)q * ldx + c;
  float a0 = fmaf(ld_f<T>(xp), sc, sh);
  a0 = a0 > 0.f ? a0 : a0 * slope;
  float a1 = 0.f;
  if (q + 1 < Lin) {
    a1 = fmaf(ld_f<T>(xp + ldx), sc, sh);
    a1 = a1 > 0.f ? a1 : a1 * slope;
  }
  const float w0 = w3[c * 3], w1 = w3[c * 3 + 1], w2 = w3[c * 3 + 2], bb = bias[c];
  st_f<T>(y0, fmaf(w1, a0, bb));
  st_f<T>(y0 + ldy, fmaf(w2, a0, fmaf(w0, a1, bb)));
}

template <typename T>
int launch_pool_up2(const T* x, int64_t x_bstride, int ldx, const float* ss, int ss_stride,
                    float slope, const float* w3, const float* bias, T* y, int64_t y_bstride, int ldy,
                    int B, int Lin, int C, int Cpad, cudaStream_t st) {
  ST2_REQUIRE(Lin <= 65535, "pool_up2: sequence too long for grid.y");
  dim3 grid(ceil_div(Cpad, 256), Lin, B);
  pool_up2_kernel<T><<<grid, 256, 0, st>>>(x, x_bstride, ldx, ss, ss_stride, slope, w3, bias, y,
                                           y_bstride, ldy, Lin, C, Cpad);
  ST2_LAUNCH_OK("pool_up2");
  return ST2_OK;
}
template int launch_pool_up2<float>(const float*, int64_t, int, const float*, int, float, const float*,
                                    const float*, float*, int64_t, int, int, int, int, int, cudaStream_t);
template int launch_pool_up2<__nv_bfloat16>(const __nv_bfloat16*, int64_t, int, const float*, int, float,
                                            const float*, const float*, __nv_bfloat16*, int64_t, int,
                                            int, int, int, int, cudaStream_t);

// --------------------------------------------------------------------- reflect row 0
// ReflectionPad1d((1,0)) after ups[-1]: padded row 0 = ups output at t = 1 = phase 1 of q = 0,
// which only sees input row 0 through tap (1 + padding):  y0[co] = b[co] + sum_ci a[0,ci] w[ci,co],
// then + x_source[0].  One block per batch item, one thread per output channel.
template <typename T>
__global__ void reflect_row0_kernel(const T* __restrict__ xin, int64_t xin_bstride, int cin,
                                    float slope, const float* __restrict__ w_tap,
                                    const float* __restrict__ bias, const T* __restrict__ xs,
                                    int64_t xs_bstride, T* __restrict__ y, int64_t y_bstride, int cout,
                                    float* __restrict__ stats_slot, int64_t stats_bstride) {
  const int b = blockIdx.x, co = threadIdx.x;
  if (co >= cout) return;
  const T* xr = xin + (int64_t)b * xin_bstride;
  float a = bias[co];
  for (int ci = 0; ci < cin; ++ci) {
    float v = ld_f<T>(xr + ci);
    v = v > 0.f ? v : v * slope;
    a = fmaf(v, w_tap[(int64_t)ci * cout + co], a);
  }
  a += ld_f<T>(xs + (int64_t)b * xs_bstride + co);
  st_f<T>(y + (int64_t)b * y_bstride + co, a);
  if (stats_slot) {
    const float r = round_to<T>(a);
    float* o = stats_slot + (int64_t)b * stats_bstride + co * 2;
    o[0] = r;
    o[1] = r * r;
  }
}

template <typename T>
int launch_reflect_row0(const T* xin, int64_t xin_bstride, int cin, float slope, const float* w_tap,
                        const float* bias, const T* xs, int64_t xs_bstride, T* y, int64_t y_bstride,
                        int cout, float* stats_slot, int64_t stats_bstride, int B, cudaStream_t st) {
  ST2_REQUIRE(cout <= 1024, "reflect_row0: cout");
  reflect_row0_kernel<T><<<B, round_up(cout, 32), 0, st>>>(xin, xin_bstride, cin, slope, w_tap, bias, xs,
                                                           xs_bstride, y, y_bstride, cout, stats_slot,
                                                           stats_bstride);
  ST2_LAUNCH_OK("reflect_row0");
  return ST2_OK;
}
template int launch_reflect_row0<float>(const float*, int64_t, int, float, const float*, const float*,
                                        const float*, int64_t, float*, int64_t, int, float*, int64_t,
                                        int, cudaStream_t);
template int launch_reflect_row0<__nv_bfloat16>(const __nv_bfloat16*, int64_t, int, float, const float*,
                                                const float*, const __nv_bfloat16*, int64_t,
                                                __nv_bfloat16*, int64_t, int, float*, int64_t, int,
                                                cudaStream_t);

// ------------------------------------------------------------------- layout helpers
__global__ void cl_to_ncl_kernel(const float* __restrict__ cl, int ld, int C, int L,
                                 float* __restrict__ ncl) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;  // over C*L of item b
  const int b = blockIdx.y;
  if (i >= (int64_t)C * L) return;
  const int c = (int)(i / L), t = (int)(i - (int64_t)c * L);
  ncl[(int64_t)b * C * L + i] = cl[((int64_t)b * L + t) * ld + c];
}
__global__ void ncl_to_cl_kernel(const float* __restrict__ ncl, int C, int L, float* __restrict__ cl,
                                 int ld) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;  // over L*ld of item b
  const int b = blockIdx.y;
  if (i >= (int64_t)L * ld) return;
  const int t = (int)(i / ld), c = (int)(i - (int64_t)t * ld);
  cl[(int64_t)b * L * ld + i] = c < C ? ncl[((int64_t)b * C + c) * L + t] : 0.f;
}
int launch_cl_to_ncl(const float* cl, int ld, int C, int L, float* ncl, int B, cudaStream_t st) {
  dim3 grid((unsigned)(((int64_t)C * L + 255) / 256), B);
  cl_to_ncl_kernel<<<grid, 256, 0, st>>>(cl, ld, C, L, ncl);
  ST2_LAUNCH_OK("cl_to_ncl");
  return ST2_OK;
}
int launch_ncl_to_cl(const float* ncl, int C, int L, float* cl, int ld, int B, cudaStream_t st) {
  dim3 grid((unsigned)(((int64_t)L * ld + 255) / 256), B);
  ncl_to_cl_kernel<<<grid, 256, 0, st>>>(ncl, C, L, cl, ld);
  ST2_LAUNCH_OK("ncl_to_cl");
  return ST2_OK;
}

// injected / exported `har` <-> the bf16 (hi, lo) row layout of stft_hilo_kernel
__global__ void ncl_to_hilo_kernel(const float* __restrict__ ncl, int C, int L, int rows_pad,
                                   __nv_bfloat16* __restrict__ har) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;  // over rows_pad * 32 of item b
  const int b = blockIdx.y;
  if (i >= (int64_t)rows_pad * 32) return;
  const int t = (int)(i >> 5), c = (int)(i & 31);
  const float v = (c < C && t < L) ? ncl[((int64_t)b * C + c) * L + t] : 0.f;
  const __nv_bfloat16 h = __float2bfloat16_rn(v);
  __nv_bfloat16* row = har + ((int64_t)b * rows_pad + t) * 64;
  row[c] = h;
  row[32 + c] = __float2bfloat16_rn(v - __bfloat162float(h));
}
__global__ void hilo_to_ncl_kernel(const __nv_bfloat16* __restrict__ har, int rows_pad, int C, int L,
                                   float* __restrict__ ncl) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;  // over C*L of item b
  const int b = blockIdx.y;
  if (i >= (int64_t)C * L) return;
  const int c = (int)(i / L), t = (int)(i - (int64_t)c * L);
  const __nv_bfloat16* row = har + ((int64_t)b * rows_pad + t) * 64;
  ncl[(int64_t)b * C * L + i] = __bfloat162float(row[c]) + __bfloat162float(row[32 + c]);
}
int launch_ncl_to_hilo(const float* ncl, int C, int L, int rows_pad, __nv_bfloat16* har, int B, cudaStream_t st) {
  ST2_REQUIRE(C <= 32 && rows_pad >= L, "ncl_to_hilo: shape");
  dim3 grid((unsigned)(((int64_t)rows_pad * 32 + 255) / 256), B);
  ncl_to_hilo_kernel<<<grid, 256, 0, st>>>(ncl, C, L, rows_pad, har);
  ST2_LAUNCH_OK("ncl_to_hilo");
  return ST2_OK;
}
int launch_hilo_to_ncl(const __nv_bfloat16* har, int rows_pad, int C, int L, float* ncl, int B, cudaStream_t st) {
  dim3 grid((unsigned)(((int64_t)C * L + 255) / 256), B);
  hilo_to_ncl_kernel<<<grid, 256, 0, st>>>(har, rows_pad, C, L, ncl);
  ST2_LAUNCH_OK("hilo_to_ncl");
  return ST2_OK;
}

// ------------------------------------------------------- mean-free storage of x_source (hifigan)
// The time-domain noise convs of the HiFi-GAN generator (hifigan.py:295-302) see a one-channel
// source whose mean is far from zero, so each output channel is a small signal riding on a large
// constant kappa[b,c] = bias[c] + mean_t(har_source[b]) * sum_taps w[c]; rounding that sum to bf16
// and then instance-normalising it amplifies the rounding noise by |mean|/std (20-5000 here).
// InstanceNorm ignores a per-(b,c) constant, so the bf16 path stores x_source - kappa and adds
// kappa back where the un-normalised tensor is consumed (the residual of noise_res iteration 0).
__global__ void __launch_bounds__(256) row_mean_kernel(const float* __restrict__ x, int64_t L, float* __restrict__ mean) {
  __shared__ double red[256];
  const float* xb = x + (int64_t)blockIdx.x * L;
  double s = 0.0;
  for (int64_t i = threadIdx.x; i < L; i += 256) s += (double)xb[i];
  red[threadIdx.x] = s;
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) {
    if (threadIdx.x < o) red[threadIdx.x] += red[threadIdx.x + o];
    __syncthreads();
  }
  if (threadIdx.x == 0) mean[blockIdx.x] = (float)(red[0] / (double)L);
}
// bias_nc_shift[b,c] = -mean[b]*wsum[c]  (noise conv bias with kappa removed)
// bias_c2_add[b,c]   = bias_c2[c] + bias_nc[c] + mean[b]*wsum[c]  (conv2 bias with kappa restored)
__global__ void noise_bias_kernel(const float* __restrict__ mean, const float* __restrict__ wsum,
                                  const float* __restrict__ bias_nc, const float* __restrict__ bias_c2,
                                  float* __restrict__ bias_nc_shift, float* __restrict__ bias_c2_add, int C,
                                  int reps) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x, b = blockIdx.y;
  if (c >= C) return;
  const float k = mean[b] * wsum[c];
  bias_nc_shift[(int64_t)b * C + c] = -k;
  for (int r = 0; r < reps; ++r)  // repeated for the row-pair view (rows of reps*C columns)
    bias_c2_add[((int64_t)b * reps + r) * C + c] = bias_c2[c] + bias_nc[c] + k;
}
int launch_row_mean(const float* x, int64_t L, float* mean, int B, cudaStream_t st) {
  row_mean_kernel<<<B, 256, 0, st>>>(x, L, mean);
  ST2_LAUNCH_OK("row_mean");
  return ST2_OK;
}
int launch_noise_bias(const float* mean, const float* wsum, const float* bias_nc, const float* bias_c2,
                      float* bias_nc_shift, float* bias_c2_add, int C, int B, cudaStream_t st, int reps) {
  dim3 grid(ceil_div(C, 128), B);
  noise_bias_kernel<<<grid, 128, 0, st>>>(mean, wsum, bias_nc, bias_c2, bias_nc_shift, bias_c2_add, C, reps);
  ST2_LAUNCH_OK("noise_bias");
  return ST2_OK;
}

// debug / parity taps: channels-last activation (fp32 or bf16) -> fp32 NCL
template <typename T>
__global__ void act_to_ncl_kernel(const T* __restrict__ cl, int64_t bstride, int ld, int C, int L,
                                  float* __restrict__ ncl) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int b = blockIdx.y;
  if (i >= (int64_t)C * L) return;
  const int c = (int)(i / L), t = (int)(i - (int64_t)c * L);
  ncl[(int64_t)b * C * L + i] = ld_f<T>(cl + (int64_t)b * bstride + (int64_t)t * ld + c);
}
template <typename T>
int launch_act_to_ncl(const T* cl, int64_t bstride, int ld, int C, int L, float* ncl, int B, cudaStream_t st) {
  dim3 grid((unsigned)(((int64_t)C * L + 255) / 256), B);
  act_to_ncl_kernel<T><<<grid, 256, 0, st>>>(cl, bstride, ld, C, L, ncl);
  ST2_LAUNCH_OK("act_to_ncl");
  return ST2_OK;
}
template int launch_act_to_ncl<float>(const float*, int64_t, int, int, int, float*, int, cudaStream_t);
template int launch_act_to_ncl<__nv_bfloat16>(const __nv_bfloat16*, int64_t, int, int, int, float*, int,
                                              cudaStream_t);

// tanh head of the HiFi-GAN generator (hifigan.py:344), in place on the fp32 waveform
__global__ void tanh_kernel(float* __restrict__ x, int64_t n) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) x[i] = tanhf(x[i]);
}
int launch_tanh_inplace(float* x, int64_t n, cudaStream_t st) {
  tanh_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(x, n);
  ST2_LAUNCH_OK("tanh");
  return ST2_OK;
}

// Output stage (next row N4): drop the last trim samples of every item (Inference_LibriTTS.ipynb cell 16:
// out[..., :-50]) and convert to 16-bit PCM: clip(rint(x * 32767), -32768, 32767), round half to even.
__global__ void wave_to_pcm16_kernel(const float* __restrict__ wav, int64_t L, int64_t Lo, int16_t* __restrict__ out) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int b = blockIdx.y;
  if (i >= Lo) return;
  const float v = fminf(fmaxf(wav[(int64_t)b * L + i] * 32767.f, -32768.f), 32767.f);
  out[(int64_t)b * Lo + i] = (int16_t)__float2int_rn(v);
}
int launch_wave_to_pcm16(const float* wav, int B, int64_t L, int64_t trim, int16_t* out, cudaStream_t st) {
  const int64_t Lo = L - trim;
  dim3 grid((unsigned)((Lo + 255) / 256), B);
  wave_to_pcm16_kernel<<<grid, 256, 0, st>>>(wav, L, Lo, out);
  ST2_LAUNCH_OK("wave_to_pcm16");
  return ST2_OK;
}

// HiFi-GAN head on the tensor-core path: conv_post ran on row pairs and left fp32 rows of ld columns
// whose first ncol entries are consecutive samples; wav[i] = tanh(cp[i / ncol][i % ncol]).
__global__ void cols_tanh_kernel(const float* __restrict__ cp, int ld, int ncol, float* __restrict__ wav, int64_t n) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) wav[i] = tanhf(cp[(i / ncol) * ld + (i % ncol)]);
}
int launch_cols_tanh(const float* cp, int ld, int ncol, float* wav, int64_t n, cudaStream_t st) {
  cols_tanh_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(cp, ld, ncol, wav, n);
  ST2_LAUNCH_OK("cols_tanh");
  return ST2_OK;
}

}  // namespace st2

// ------------------------------------------------- fp32 parity path on the tensor cores (ST2_PREC_FP32_TC)
// A bf16 x bf16 -> fp32 MMA reproduces an fp32 product to ~2^-17 when both operands are split into bf16 (hi, lo) pairs
// and the three significant products are kept:  x w ~= x_hi w_hi + x_lo w_hi + x_hi w_lo.  As ONE conv on the existing
// tcgen05 kernel: input channels [x_hi | x_lo | x_hi] (3 * cin_pad bf16 per row, written by split3_kernel below
// together with the layer's fused prologue AdaIN scale / shift + LeakyReLU / Snake, in fp32, exactly as
// conv_simt_kernel applies it) against weight rows [w_hi | w_hi | w_lo] (packed at load time), fp32 accumulation in
// TMEM, fp32 output; residual / scale / running mean are applied by post32_kernel, statistics by colstats_kernel.
namespace st2 {

// grid (ceil(L / 8), B), block 256: a warp converts rows, lanes stride over channels (coalesced 4-byte reads,
// 2-byte writes into each of the three blocks)
__global__ void __launch_bounds__(256)
split3_kernel(const float* __restrict__ x, int64_t x_bstride, int ldx, int L, int cin, int cin_pad,
              const float* __restrict__ ss, int ss_stride, int act, float slope, const float* __restrict__ alpha,
              __nv_bfloat16* __restrict__ y) {
  pdl_wait_then_trigger();
  const int b = blockIdx.y, lane = threadIdx.x & 31, wrp = threadIdx.x >> 5;
  const int t = blockIdx.x * 8 + wrp;
  if (t >= L) return;
  const float* xr = x + (int64_t)b * x_bstride + (int64_t)t * ldx;
  const float* ssb = ss ? ss + (int64_t)b * ss_stride * 2 : nullptr;
  __nv_bfloat16* yr = y + ((int64_t)b * L + t) * (3 * (int64_t)cin_pad);
  for (int c = lane; c < cin_pad; c += 32) {
    float v = 0.f;
    if (c < cin) {
      v = xr[c];
      if (ssb) v = fmaf(v, ssb[2 * c], ssb[2 * c + 1]);
      float a = 1.f;
      if (act == ST2_ACT_SNAKE) a = alpha[c];
      v = apply_act<true>(v, act, slope, a, 1.f / a);
    }
    const __nv_bfloat16 hi = __float2bfloat16_rn(v);
    const __nv_bfloat16 lo = __float2bfloat16_rn(v - __bfloat162float(hi));
    yr[c] = hi;
    yr[cin_pad + c] = lo;
    yr[2 * cin_pad + c] = hi;
  }
}

int launch_split3(const float* x, int64_t x_bstride, int ldx, int L, int cin, int cin_pad, const float* ss,
                  int ss_stride, int act, float slope, const float* alpha, __nv_bfloat16* y, int B, cudaStream_t st) {
  ST2_CUDA_OK(launch_pdl(split3_kernel, dim3(ceil_div(L, 8), B), dim3(256), 0, st, x, x_bstride, ldx, L, cin, cin_pad, ss,
                         ss_stride, act, slope, alpha, y));
  ST2_LAUNCH_OK("split3");
  return ST2_OK;
}

// y[b,t,n] = (acc[b,t,n] + res[b, t >> res_shift, n]) * out_scale (+ y_old[b,t,n]); acc [B][L][ldacc] fp32 (the conv
// with its bias), n < n_cols.  grid (ceil(L * n_cols / 256), B)
__global__ void __launch_bounds__(256)
post32_kernel(const float* __restrict__ acc, int ldacc, int L, int n_cols, const float* __restrict__ res,
              int64_t res_bstride, int ldres, int res_shift, float out_scale, int accumulate, float* __restrict__ y,
              int64_t y_bstride, int ldy) {
  pdl_wait_then_trigger();
  const int b = blockIdx.y;
  const int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x;
  if (i >= (int64_t)L * n_cols) return;
  const int t = (int)(i / n_cols), n = (int)(i - (int64_t)t * n_cols);
  float v = acc[((int64_t)b * L + t) * ldacc + n];
  if (res) v += res[(int64_t)b * res_bstride + (int64_t)(t >> res_shift) * ldres + n];
  v *= out_scale;
  float* yp = y + (int64_t)b * y_bstride + (int64_t)t * ldy + n;
  if (accumulate) v += *yp;
  *yp = v;
}

int launch_post32(const float* acc, int ldacc, int L, int n_cols, const float* res, int64_t res_bstride, int ldres,
                  int res_shift, float out_scale, int accumulate, float* y, int64_t y_bstride, int ldy, int B,
                  cudaStream_t st) {
  const int64_t n = (int64_t)L * n_cols;
  ST2_CUDA_OK(launch_pdl(post32_kernel, dim3((unsigned)((n + 255) / 256), B), dim3(256), 0, st, acc, ldacc, L, n_cols, res,
                         res_bstride, ldres, res_shift, out_scale, accumulate, y, y_bstride, ldy));
  ST2_LAUNCH_OK("post32");
  return ST2_OK;
}

}  // namespace st2
