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

int launch_conv_simt(const St2ConvParams& p, cudaStream_t st) {
  ST2_REQUIRE(p.B > 0 && p.Lout > 0 && p.cin > 0 && p.n_cols > 0 && p.k > 0, "conv_simt: empty shape");
  ST2_REQUIRE(p.stride >= 1 && p.dil >= 1, "conv_simt: stride/dilation");
  ST2_REQUIRE(!(p.stats && p.accumulate), "conv_simt: stats with accumulate");
  ST2_REQUIRE(!p.stats || p.n_tiles >= ceil_div(p.Lout, SIMT_TT), "conv_simt: n_tiles");
  ST2_REQUIRE(p.act != ST2_ACT_SNAKE || p.alpha, "conv_simt: snake needs alpha");
  const bool xi = p.x_dtype == ST2_DT_BF16, yo = p.y_dtype == ST2_DT_BF16;
  if (!xi && !yo) return launch_t<float, float, true>(p, st);
  if (xi && yo) return launch_t<__nv_bfloat16, __nv_bfloat16, false>(p, st);
  if (!xi && yo) return launch_t<float, __nv_bfloat16, false>(p, st);
  return launch_t<__nv_bfloat16, float, false>(p, st);
}

}  // namespace st2
