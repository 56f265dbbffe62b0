// conv_simt.cu — generic fused 1-D convolution on CUDA cores (fp32 FFMA).
//
// This is the parity path (ST2_PREC_FP32) for every convolution of the decoder and the
// production path for the thin layers that do not map onto 128-wide UMMA tiles
// (F0/N 3-tap convs, asr_res 512->64, noise_convs with Cin = 22 / 1, conv_post Cout = 22 / 1).
// Semantics: conv_params.h.  One block = 64 time rows x 64 output columns, 256 threads,
// 4x4 register micro-tile per thread; the transformed input tile (AdaIN affine + Snake /
// LeakyReLU applied once per element) and a [taps][channels][64] weight slab are staged in
// shared memory.
#include "common.cuh"

namespace st2 {

constexpr int SIMT_TT = 64;   // time rows per block
constexpr int SIMT_TC = 64;   // output columns per block
constexpr int SIMT_THREADS = 256;

template <typename InT, typename OutT, bool PRECISE>
__global__ void __launch_bounds__(SIMT_THREADS)
conv_simt_kernel(const St2ConvParams p, const int CK, const int KJ, const int rows_in) {
  extern __shared__ float smem[];
  float* Xs = smem;                           // [rows_in][CK]
  float* Ws = smem + (((size_t)rows_in * CK + 3) & ~(size_t)3);  // [KJ][CK][64], 16-byte aligned
  __shared__ float red[2][16][SIMT_TC];

  const int tid = threadIdx.x;
  const int tx = tid & 15, ty = tid >> 4;
  const int tile = blockIdx.x, b = blockIdx.z;
  const int t0 = tile * SIMT_TT, n0 = blockIdx.y * SIMT_TC;

  const InT* xb = reinterpret_cast<const InT*>(p.x) + (int64_t)b * p.x_bstride;
  const float* ssb = p.ss ? p.ss + (int64_t)b * p.ss_stride * 2 : nullptr;
  const float* wg = reinterpret_cast<const float*>(p.w);

  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int q = 0; q < 4; ++q) acc[i][q] = 0.f;

  const int row_base = t0 * p.stride - p.pad;
  for (int c0 = 0; c0 < p.cin; c0 += CK) {
    const int ck = min(CK, p.cin - c0);
    __syncthreads();
    for (int idx = tid; idx < rows_in * CK; idx += SIMT_THREADS) {
      const int r = idx / CK, c = idx - r * CK;
      const int tin = row_base + r, ch = c0 + c;
      float v = 0.f;
      if (c < ck && tin >= 0 && tin < p.Lin) {
        v = ld_f<InT>(xb + (int64_t)tin * p.ldx + ch);
        if (ssb) v = fmaf(v, ssb[2 * ch], ssb[2 * ch + 1]);
        float a = 1.f;
        if (p.act == ST2_ACT_SNAKE) a = p.alpha[ch];
        v = apply_act<PRECISE>(v, p.act, p.slope, a, 1.f / a);
      }
      Xs[idx] = v;
    }
    for (int j0 = 0; j0 < p.k; j0 += KJ) {
      const int kj = min(KJ, p.k - j0);
      if (j0 > 0) __syncthreads();
      for (int idx = tid; idx < kj * CK * SIMT_TC; idx += SIMT_THREADS) {
        const int j = idx / (CK * SIMT_TC);
        const int rem = idx - j * CK * SIMT_TC;
        const int c = rem / SIMT_TC, n = rem - c * SIMT_TC;
        const int col = n0 + n;
        float wv = 0.f;
        if (c < ck && col < p.n_cols)
          wv = wg[((int64_t)(j0 + j) * p.cin + (c0 + c)) * p.n_cols + col];
        Ws[idx] = wv;
      }
      __syncthreads();
      for (int j = 0; j < kj; ++j) {
        const float* xrow = Xs + ((size_t)(ty * 4) * p.stride + (size_t)(j0 + j) * p.dil) * CK;
        const float* wrow = Ws + (size_t)j * CK * SIMT_TC + tx * 4;
        for (int c = 0; c < ck; ++c) {
          const float4 w4 = *reinterpret_cast<const float4*>(wrow + c * SIMT_TC);
          float xv[4];
#pragma unroll
          for (int i = 0; i < 4; ++i) xv[i] = xrow[(size_t)i * p.stride * CK + c];
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            acc[i][0] = fmaf(xv[i], w4.x, acc[i][0]);
            acc[i][1] = fmaf(xv[i], w4.y, acc[i][1]);
            acc[i][2] = fmaf(xv[i], w4.z, acc[i][2]);
            acc[i][3] = fmaf(xv[i], w4.w, acc[i][3]);
          }
        }
      }
    }
  }

  // ---- epilogue
  OutT* yb = reinterpret_cast<OutT*>(p.y) + (int64_t)b * p.y_bstride;
  const OutT* rb = p.res ? reinterpret_cast<const OutT*>(p.res) + (int64_t)b * p.res_bstride : nullptr;
  float s1[4] = {0.f, 0.f, 0.f, 0.f}, s2[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int t = t0 + ty * 4 + i;
    if (t >= p.Lout) continue;
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int col = n0 + tx * 4 + q;
      if (col >= p.n_cols) continue;
      float v = acc[i][q];
      if (p.bias) v += p.bias[(int64_t)b * p.bias_bstride + col];
      if (rb) v += ld_f<OutT>(rb + (int64_t)(t >> p.res_shift) * p.ldres + col);
      v *= p.out_scale;
      OutT* yp = yb + (int64_t)t * p.ldy + col;
      if (p.accumulate) v += ld_f<OutT>(yp);
      st_f<OutT>(yp, v);
      const float vr = round_to<OutT>(v);
      s1[q] += vr;
      s2[q] += vr * vr;
    }
  }
  if (p.stats) {
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      red[0][ty][tx * 4 + q] = s1[q];
      red[1][ty][tx * 4 + q] = s2[q];
    }
    __syncthreads();
    if (tid < SIMT_TC) {
      const int col = n0 + tid;
      if (col < p.n_cols) {
        float a = 0.f, q2 = 0.f;
#pragma unroll
        for (int r = 0; r < 16; ++r) {
          a += red[0][r][tid];
          q2 += red[1][r][tid];
        }
        float* sp = p.stats + (((int64_t)b * p.n_tiles + tile) * p.n_cols + col) * 2;
        sp[0] = a;
        sp[1] = q2;
      }
    }
  }
}

template <typename InT, typename OutT, bool PRECISE>
static int launch_t(const St2ConvParams& p, cudaStream_t st) {
  const int CK = p.cin < 16 ? p.cin : 16;
  const int KJ = p.k < 8 ? p.k : 8;
  const int rows_in = (SIMT_TT - 1) * p.stride + (p.k - 1) * p.dil + 1;
  const size_t smem = ((((size_t)rows_in * CK + 3) & ~(size_t)3) + (size_t)KJ * CK * SIMT_TC) * sizeof(float);
  ST2_REQUIRE(smem <= 200 * 1024, "conv_simt: tile does not fit shared memory");
  auto kern = conv_simt_kernel<InT, OutT, PRECISE>;
  if (smem > 48 * 1024) {
    ST2_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  }
  dim3 grid(ceil_div(p.Lout, SIMT_TT), ceil_div(p.n_cols, SIMT_TC), p.B);
  kern<<<grid, SIMT_THREADS, smem, st>>>(p, CK, KJ, rows_in);
  ST2_LAUNCH_OK("conv_simt");
  return ST2_OK;
}

// ---------------------------------------------------------------------------- thin conv
// One input channel, many output channels (HiFi-GAN's strided noise convs on the raw excitation,
// hifigan.py:301-311: Conv1d(1, C, k = 2 s, stride s)): HBM-write bound (k <= 4) or FMA-bound (k = 12, 60), so no GEMM
// tiling; a block stages its input window and the [k][C] weights in shared memory, a thread owns 8 channels of a
// row and writes one 16-byte bf16 vector.  Same argument block, tile size (64 rows) and statistics
// layout as conv_simt_kernel.
constexpr int THIN_TT = SIMT_TT;
// A block owns `tiles_per_block` consecutive 64-row tiles of one item: it stages their whole input window and the [k][n]
// weights once (one __syncthreads per block, not per tile), a thread computes RT rows x 8 channels at a time (one pair
// of 16-byte weight loads feeds 8 RT FMAs) and keeps its (sum, sumsq) in registers over all its rows; one reduction at
// the end.  The block's statistics go to the partial slot of its FIRST tile and zeros to the slots of its other tiles
// (the consumer sums the slots of an item in a fixed order: same result, deterministic).
template <int RT>
__global__ void __launch_bounds__(256)
conv_thin_kernel(const St2ConvParams p, int tiles_per_block) {
  extern __shared__ float thin_smem[];
  const int n = p.n_cols, G = n >> 3, R = 256 / G;       // channel groups, row threads
  const int win_max = (tiles_per_block * THIN_TT - 1) * p.stride + p.k;
  float* xs = thin_smem;                                  // [win_max]
  float* ws = thin_smem + ((win_max + 3) & ~3);           // [k][n]
  float* red = ws + p.k * n;                              // [2][8 warps][n]
  const int b = blockIdx.y, tid = threadIdx.x;
  const float* xb = reinterpret_cast<const float*>(p.x) + (int64_t)b * p.x_bstride;
  __nv_bfloat16* yb = reinterpret_cast<__nv_bfloat16*>(p.y) + (int64_t)b * p.y_bstride;
  const float* wg = reinterpret_cast<const float*>(p.w);
  const int n_tiles_item = (p.Lout + THIN_TT - 1) / THIN_TT;
  const int tile0 = blockIdx.x * tiles_per_block;
  const int tile_end = min(n_tiles_item, tile0 + tiles_per_block);
  const int t_begin = tile0 * THIN_TT, t_end = min(p.Lout, tile_end * THIN_TT);
  const int rows = t_end - t_begin;
  for (int i = tid; i < p.k * n; i += 256) ws[i] = wg[i];
  {
    const int i0 = t_begin * p.stride - p.pad;
    const int win = (rows - 1) * p.stride + p.k;
    for (int i = tid; i < win; i += 256) {
      const int r = i0 + i;
      xs[i] = (r >= 0 && r < p.Lin) ? xb[(int64_t)r * p.ldx] : 0.f;
    }
  }
  const int cg = tid % G, rr = tid / G;
  float bias[8];
#pragma unroll
  for (int e = 0; e < 8; ++e) bias[e] = p.bias ? p.bias[(int64_t)b * p.bias_bstride + cg * 8 + e] : 0.f;
  __syncthreads();
  float s1[8], s2[8];
#pragma unroll
  for (int e = 0; e < 8; ++e) s1[e] = s2[e] = 0.f;
  const float* wr = ws + cg * 8;
  for (int tl = rr * RT; tl < rows; tl += R * RT) {
    float acc[RT][8];
#pragma unroll
    for (int i = 0; i < RT; ++i)
#pragma unroll
      for (int e = 0; e < 8; ++e) acc[i][e] = bias[e];
    // rows past the block's last one read inside the staged window or just behind it (never stored): clamp the offset
    int xo[RT];
#pragma unroll
    for (int i = 0; i < RT; ++i) xo[i] = min(tl + i, rows - 1) * p.stride;
    for (int j = 0; j < p.k; ++j) {
      const float4 w0 = *reinterpret_cast<const float4*>(wr + j * n);
      const float4 w1 = *reinterpret_cast<const float4*>(wr + j * n + 4);
#pragma unroll
      for (int i = 0; i < RT; ++i) {
        const float xv = xs[xo[i] + j];
        acc[i][0] = fmaf(w0.x, xv, acc[i][0]); acc[i][1] = fmaf(w0.y, xv, acc[i][1]);
        acc[i][2] = fmaf(w0.z, xv, acc[i][2]); acc[i][3] = fmaf(w0.w, xv, acc[i][3]);
        acc[i][4] = fmaf(w1.x, xv, acc[i][4]); acc[i][5] = fmaf(w1.y, xv, acc[i][5]);
        acc[i][6] = fmaf(w1.z, xv, acc[i][6]); acc[i][7] = fmaf(w1.w, xv, acc[i][7]);
      }
    }
#pragma unroll
    for (int i = 0; i < RT; ++i) {
      if (tl + i < rows) {
        uint32_t o[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const __nv_bfloat162 pk =
              __floats2bfloat162_rn(acc[i][2 * e] * p.out_scale, acc[i][2 * e + 1] * p.out_scale);
          o[e] = *reinterpret_cast<const uint32_t*>(&pk);
          const float lo = __bfloat162float(pk.x), hi = __bfloat162float(pk.y);
          s1[2 * e] += lo; s2[2 * e] = fmaf(lo, lo, s2[2 * e]);
          s1[2 * e + 1] += hi; s2[2 * e + 1] = fmaf(hi, hi, s2[2 * e + 1]);
        }
        *reinterpret_cast<uint4*>(yb + (int64_t)(t_begin + tl + i) * p.ldy + cg * 8) = make_uint4(o[0], o[1], o[2], o[3]);
      }
    }
  }
  if (!p.stats) return;
  // rows that share a warp first (lanes cg, cg + G, ...), then one partial per warp
  for (int off = G; off < 32; off <<= 1) {
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      s1[e] += __shfl_xor_sync(0xffffffffu, s1[e], off);
      s2[e] += __shfl_xor_sync(0xffffffffu, s2[e], off);
    }
  }
  if ((tid & 31) < G) {
    float4* r1 = reinterpret_cast<float4*>(red + (0 * 8 + (tid >> 5)) * n + cg * 8);
    float4* r2 = reinterpret_cast<float4*>(red + (1 * 8 + (tid >> 5)) * n + cg * 8);
    r1[0] = make_float4(s1[0], s1[1], s1[2], s1[3]); r1[1] = make_float4(s1[4], s1[5], s1[6], s1[7]);
    r2[0] = make_float4(s2[0], s2[1], s2[2], s2[3]); r2[1] = make_float4(s2[4], s2[5], s2[6], s2[7]);
  }
  __syncthreads();
  // G < 32: every warp holds all channel groups; G = 32: warp w holds rows of all 32 groups too (cg = lane)
  for (int col = tid; col < n; col += 256) {
    float a = 0.f, q2 = 0.f;
#pragma unroll
    for (int r = 0; r < 8; ++r) {  // fixed order: deterministic
      a += red[(0 * 8 + r) * n + col];
      q2 += red[(1 * 8 + r) * n + col];
    }
    float2* sp = reinterpret_cast<float2*>(p.stats + (((int64_t)b * p.n_tiles + tile0) * p.n_cols + col) * 2);
    sp[0] = make_float2(a, q2);
    for (int t = 1; t < tile_end - tile0; ++t) sp[(int64_t)t * p.n_cols] = make_float2(0.f, 0.f);
  }
}
static bool thin_ok(const St2ConvParams& p) {
  const int n = p.n_cols;
  return p.cin == 1 && p.x_dtype == ST2_DT_F32 && p.y_dtype == ST2_DT_BF16 && !p.ss && p.act == ST2_ACT_NONE && !p.res &&
         !p.accumulate && p.dil == 1 && (n == 32 || n == 64 || n == 128 || n == 256) && p.k <= 64 &&
         p.ldy % 8 == 0 && p.y_bstride % 8 == 0 && (((uintptr_t)p.y) & 15) == 0;
}
template <int RT>
static int launch_thin_t(const St2ConvParams& p, cudaStream_t st) {
  const int n = p.n_cols;
  // about 128 KB of output per block, so the [k][n] weights are read once per many tiles, but at least 16 blocks per
  // item.  A function of the item's shape only, never of B: the grouping of rows into partial sums must not depend on
  // the batch an item is decoded in (a shard gives bit-identical rows).
  int tpb = max(1, (128 * 1024) / (THIN_TT * n * 2));
  const int nt = ceil_div(p.Lout, THIN_TT);
  while (tpb > 1 && ceil_div(nt, tpb) < 16) tpb >>= 1;
  const int win = (tpb * THIN_TT - 1) * p.stride + p.k;
  const size_t smem = ((size_t)((win + 3) & ~3) + (size_t)p.k * n + (size_t)2 * 8 * n) * sizeof(float);
  ST2_REQUIRE(smem <= 160 * 1024, "conv_thin: tile does not fit shared memory");
  if (smem > 48 * 1024)
    ST2_CUDA_OK(cudaFuncSetAttribute(conv_thin_kernel<RT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  dim3 grid(ceil_div(nt, tpb), p.B);
  conv_thin_kernel<RT><<<grid, 256, smem, st>>>(p, tpb);
  ST2_LAUNCH_OK("conv_thin");
  return ST2_OK;
}
static int launch_thin(const St2ConvParams& p, cudaStream_t st) {
  // row threads R = 2048 / n, RT rows each per pass.  The long-tap layers (k = 12, 60 at n >= 128) are FMA- and
  // shared-memory-bound: 8 rows per weight load; the short-tap ones (k <= 4 at n <= 64) are HBM-write-bound and want
  // threads rather than registers: 2 rows (77 registers, 3 blocks per SM; measured 127 us = 3.9 TB/s at n = 32)
  switch (p.n_cols) {
    case 32: return launch_thin_t<2>(p, st);
    case 64: return launch_thin_t<2>(p, st);
    default: return launch_thin_t<8>(p, st);
  }
}

// launches of this thread with a GEMM-sized contraction (cin >= 64 at stride 1: the layers the tcgen05 kernel can take);
// tests use it to show that ST2_PREC_FP32_TC leaves no dense conv on the FFMA kernel
static thread_local int64_t g_simt_dense = 0;
int64_t conv_simt_dense_launches(bool reset) {
  const int64_t v = g_simt_dense;
  if (reset) g_simt_dense = 0;
  return v;
}

int launch_conv_simt(const St2ConvParams& p, cudaStream_t st) {
  ST2_REQUIRE(p.B > 0 && p.Lout > 0 && p.cin > 0 && p.n_cols > 0 && p.k > 0, "conv_simt: empty shape");
  if (p.cin >= 64 && p.stride == 1) ++g_simt_dense;
  ST2_REQUIRE(p.stride >= 1 && p.dil >= 1, "conv_simt: stride/dilation");
  ST2_REQUIRE(!(p.stats && p.accumulate), "conv_simt: stats with accumulate");
  ST2_REQUIRE(!p.stats || p.n_tiles >= ceil_div(p.Lout, SIMT_TT), "conv_simt: n_tiles");
  ST2_REQUIRE(p.act != ST2_ACT_SNAKE || p.alpha, "conv_simt: snake needs alpha");
  if (thin_ok(p)) return launch_thin(p, st);
  const bool xi = p.x_dtype == ST2_DT_BF16, yo = p.y_dtype == ST2_DT_BF16;
  if (!xi && !yo) return launch_t<float, float, true>(p, st);
  if (xi && yo) return launch_t<__nv_bfloat16, __nv_bfloat16, false>(p, st);
  if (!xi && yo) return launch_t<float, __nv_bfloat16, false>(p, st);
  return launch_t<__nv_bfloat16, float, false>(p, st);
}

}  // namespace st2
