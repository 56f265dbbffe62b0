// source.cu — harmonic excitation, forward STFT and the inverse-STFT output head.
//
//   phase_scan   SineGen._f02sine up to the frame-rate phase (istftnet.py:152,174,183-184):
//                rad = (f0*h/24000) % 1 per frame, block-level inclusive scan in fp64 (ATen's CPU
//                cumsum accumulates float in double), then *2*pi*spf with fp32 rounding points
//                identical to the reference expression order.
//   har_source   linear x spf resampling of the phase with ATen's exact fp32 arithmetic
//                (src = fma(1/spf, n+0.5, -0.5); out = fma(1-l, p[i0], l*p[i1])), sin, *0.1, uv mask,
//                additive noise (injected or Philox), Linear(9->1) + tanh (istftnet.py:186,230-246,292).
//   stft         torch.stft(n_fft 20, hop 5, hann periodic, center/reflect) -> |X|, angle(X) as 22
//                channels-last columns (istftnet.py:91-97,355-356): a register-resident 20-point DFT.
//   istft        exp / sin heads, inverse 20-point DFT, windowed overlap-add, envelope division and
//                trimming (istftnet.py:377-379, 99-104); writes the final waveform.
// All four are memory-bound; bytes per unit are listed in DESIGN.md.
#include <mutex>

#include "common.cuh"

namespace st2 {

constexpr int NH = 9;  // fundamental + 8 overtones (istftnet.py:312)
constexpr float F_PI = 3.14159265358979323846f;

__constant__ float c_cos20[20];
__constant__ float c_sin20[20];
__constant__ float c_hann20[20];
static bool g_tables_ready[64] = {};
static std::mutex g_tables_mu;

// constant tables of the 20-point DFT, once per device (st2_decoder_load calls it through source_prepare, so a forward
// never runs the synchronous cudaMemcpyToSymbol)
static int ensure_tables() {
  int dev = 0;
  ST2_CUDA_OK(cudaGetDevice(&dev));
  std::lock_guard<std::mutex> lk(g_tables_mu);
  if (dev < 64 && g_tables_ready[dev]) return ST2_OK;
  float c[20], s[20], w[20];
  const double pi = 3.14159265358979323846;
  for (int m = 0; m < 20; ++m) {
    double cv = cos(2.0 * pi * m / 20.0), sv = sin(2.0 * pi * m / 20.0);
    if (fabs(cv) < 1e-12) cv = 0.0;
    if (fabs(sv) < 1e-12) sv = 0.0;
    c[m] = (float)cv;
    s[m] = (float)sv;
    w[m] = (float)(0.5 - 0.5 * cos(2.0 * pi * m / 20.0));  // scipy get_window('hann', 20, fftbins=True)
  }
  ST2_CUDA_OK(cudaMemcpyToSymbol(c_cos20, c, sizeof(c)));
  ST2_CUDA_OK(cudaMemcpyToSymbol(c_sin20, s, sizeof(s)));
  ST2_CUDA_OK(cudaMemcpyToSymbol(c_hann20, w, sizeof(w)));
  if (dev < 64) g_tables_ready[dev] = true;
  return ST2_OK;
}

int source_prepare() { return ensure_tables(); }

// ----------------------------------------------------------------------- phase scan
// grid B, block 256.  phase layout [B][9][T] (harmonic-major).
__global__ void __launch_bounds__(256)
phase_scan_kernel(const float* __restrict__ f0, int T, float spf_f, float* __restrict__ phase) {
  __shared__ double part[NH][256];
  const int b = blockIdx.x, tid = threadIdx.x;
  const int per = (T + 255) / 256;
  const int q0 = min(T, tid * per), q1 = min(T, q0 + per);
  const float* f = f0 + (int64_t)b * T;
  double loc[NH];
#pragma unroll
  for (int h = 0; h < NH; ++h) loc[h] = 0.0;
  for (int q = q0; q < q1; ++q) {
    const float fv = f[q];
#pragma unroll
    for (int h = 0; h < NH; ++h) {
      float r = __fdiv_rn(__fmul_rn(fv, (float)(h + 1)), 24000.f);
      r = r - floorf(r);  // torch `% 1` (python modulo)
      loc[h] += (double)r;
    }
  }
#pragma unroll
  for (int h = 0; h < NH; ++h) part[h][tid] = loc[h];
  __syncthreads();
  if (tid < NH) {  // exclusive scan over the 256 chunk sums
    double run = 0.0;
    for (int i = 0; i < 256; ++i) {
      const double v = part[tid][i];
      part[tid][i] = run;
      run += v;
    }
  }
  __syncthreads();
#pragma unroll
  for (int h = 0; h < NH; ++h) loc[h] = part[h][tid];
  for (int q = q0; q < q1; ++q) {
    const float fv = f[q];
#pragma unroll
    for (int h = 0; h < NH; ++h) {
      float r = __fdiv_rn(__fmul_rn(fv, (float)(h + 1)), 24000.f);
      r = r - floorf(r);
      loc[h] += (double)r;
      float p = (float)loc[h];               // cumsum result, fp32
      p = __fmul_rn(__fmul_rn(p, 2.f), F_PI);  // * 2 * np.pi
      p = __fmul_rn(p, spf_f);                 // * upsample_scale (istftnet.py:184)
      phase[((int64_t)b * NH + h) * T + q] = p;
    }
  }
}

int launch_phase_scan(const float* f0, int B, int T, int spf, float* phase, cudaStream_t st) {
  phase_scan_kernel<<<B, 256, 0, st>>>(f0, T, (float)spf, phase);
  ST2_LAUNCH_OK("phase_scan");
  return ST2_OK;
}

// ----------------------------------------------------------------------- har source
__global__ void __launch_bounds__(256)
har_source_kernel(const float* __restrict__ f0, const float* __restrict__ phase,
                  const float* __restrict__ noise, uint64_t seed, const float* __restrict__ lin_w,
                  float lin_b, int T, int spf, float ratio, float* __restrict__ out, int64_t item0,
                  const uint64_t* __restrict__ seed_dev) {
  if (seed_dev) seed = *seed_dev;
  const int b = blockIdx.y;
  const int Lw = T * spf;  // < 2^31 (checked at launch): 32-bit index arithmetic, no 64-bit division
  const int n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= Lw) return;
  float src = fmaf(ratio, (float)n + 0.5f, -0.5f);
  src = src < 0.f ? 0.f : src;
  int i0 = (int)src;
  i0 = min(i0, T - 1);
  const int i1 = min(i0 + 1, T - 1);
  const float l1 = __fsub_rn(src, (float)i0);
  const float l0 = __fsub_rn(1.f, l1);
  const float fv = f0[(int64_t)b * T + n / spf];
  const float uv = fv > 10.f ? 1.f : 0.f;                         // voiced_threshold 10 (istftnet.py:312)
  const float namp = uv > 0.f ? 0.003f : __fdiv_rn(0.1f, 3.f);   // istftnet.py:240
  float z[12];
  if (noise) {
    const float* np_ = noise + ((int64_t)b * Lw + n) * NH;
#pragma unroll
    for (int h = 0; h < NH; ++h) z[h] = np_[h];
  } else {
    // counter = (global item, sample): a shard / micro-batch starting at item0 draws the same noise
    const uint64_t ctr = (uint64_t)(b + item0) * (uint64_t)Lw + (uint64_t)n;
#pragma unroll
    for (int g = 0; g < 3; ++g) {
      uint32_t r[4];
      philox4x32((uint32_t)ctr, (uint32_t)(ctr >> 32), (uint32_t)g, 0x5742u, (uint32_t)seed,
                 (uint32_t)(seed >> 32), r);
      box_muller(r[0], r[1], z[4 * g], z[4 * g + 1]);
      box_muller(r[2], r[3], z[4 * g + 2], z[4 * g + 3]);
    }
  }
  const float* pb = phase + (int64_t)b * NH * T;
  float acc = 0.f;
#pragma unroll
  for (int h = 0; h < NH; ++h) {
    const float p0 = pb[(int64_t)h * T + i0], p1 = pb[(int64_t)h * T + i1];
    const float ph = fmaf(l0, p0, __fmul_rn(l1, p1));
    // sin of an fp32 phase that reaches 1e5 .. 1e6 rad (libdevice sinf would take its slow Payne-Hanek path):
    // three-term Cody-Waite reduction, 2*pi = 6.28125 + 1.9350051879882812e-3 + 3.019916e-7.  The first fma is
    // exact (ph and k * 6.28125 are multiples of 2^-5 or coarser and the difference is < 8), the other two round
    // at 4e-7; then MUFU.SIN on [-pi, pi] (abs error < 4e-7).
    const float kq = rintf(ph * 0.15915494309189535f);
    float red = fmaf(-kq, 6.28125f, ph);
    red = fmaf(-kq, 1.9350051879882812e-3f, red);
    red = fmaf(-kq, 3.019916050561733e-7f, red);
    const float sv = __fmul_rn(__sinf(red), 0.1f);  // sine_amp
    const float sw = __fadd_rn(__fmul_rn(sv, uv), __fmul_rn(namp, z[h]));
    acc = fmaf(lin_w[h], sw, acc);
  }
  out[(int64_t)b * Lw + n] = tanhf(acc + lin_b);
}

int launch_har_source(const float* f0, const float* phase, const float* noise, uint64_t seed,
                      const float* lin_w9, float lin_b, int B, int T, int spf, float* out,
                      cudaStream_t st, int64_t item0, const uint64_t* seed_dev) {
  const int64_t Lw = (int64_t)T * spf;
  ST2_REQUIRE(Lw < (1ll << 31) - 256, "har_source: T * samples_per_frame must fit 31 bits");
  dim3 grid((unsigned)((Lw + 255) / 256), B);
  const float ratio = (float)(1.0 / (double)spf);
  har_source_kernel<<<grid, 256, 0, st>>>(f0, phase, noise, seed, lin_w9, lin_b, T, spf, ratio, out, item0, seed_dev);
  ST2_LAUNCH_OK("har_source");
  return ST2_OK;
}

// ----------------------------------------------------------------------------- STFT
// thread per (b, frame).  har_cl row = [|X_0..10|, angle(X_0..10), zero padding to ld].
__global__ void __launch_bounds__(128)
stft_kernel(const float* __restrict__ hs, int Lw, int frames, float* __restrict__ har, int ld) {
  const int b = blockIdx.y;
  const int f = blockIdx.x * blockDim.x + threadIdx.x;
  if (f >= frames) return;
  const float* x = hs + (int64_t)b * Lw;
  float xw[20];
#pragma unroll
  for (int n = 0; n < 20; ++n) {
    int idx = 5 * f + n - 10;
    if (idx < 0) idx = -idx;                      // reflect, center=True
    if (idx >= Lw) idx = 2 * (Lw - 1) - idx;
    xw[n] = x[idx] * c_hann20[n];
  }
  float* o = har + ((int64_t)b * frames + f) * ld;
  float val[24];
  val[22] = val[23] = 0.f;
#pragma unroll
  for (int k = 0; k <= 10; ++k) {
    float re = 0.f, im = 0.f;
#pragma unroll
    for (int n = 0; n < 20; ++n) {
      const int m = (k * n) % 20;
      re = fmaf(xw[n], c_cos20[m], re);
      im = fmaf(xw[n], -c_sin20[m], im);
    }
    if (k == 0 || k == 10) im = 0.f;  // r2c: DC / Nyquist have Im == +0 -> angle in {0, +pi}
    val[k] = sqrtf(re * re + im * im);
    val[11 + k] = atan2f(im, re);
  }
  if ((ld & 3) == 0 && ld >= 24 && (((uintptr_t)har) & 15) == 0) {  // 16-byte stores (the decoder's rows: ld 32)
    float4* o4 = reinterpret_cast<float4*>(o);
#pragma unroll
    for (int i = 0; i < 6; ++i) o4[i] = make_float4(val[4 * i], val[4 * i + 1], val[4 * i + 2], val[4 * i + 3]);
    for (int c = 6; c < ld / 4; ++c) o4[c] = make_float4(0.f, 0.f, 0.f, 0.f);
  } else {
#pragma unroll
    for (int c = 0; c < 22; ++c) o[c] = val[c];
    for (int c = 22; c < ld; ++c) o[c] = 0.f;
  }
}

// Same transform written as bf16 (hi, lo) pairs for the tcgen05 noise convs: row = 64 bf16 =
// [hi(22) | 0(10) | lo(22) | 0(10)] with hi = bf16(v), lo = bf16(v - hi), so hi + lo carries 16
// mantissa bits of v.  Rows [frames, rows_pad) are zero (the strided noise conv reads whole groups
// of rows as one wide row).
__global__ void __launch_bounds__(128)
stft_hilo_kernel(const float* __restrict__ hs, int Lw, int frames, int rows_pad,
                 __nv_bfloat16* __restrict__ har) {
  const int b = blockIdx.y;
  const int f = blockIdx.x * blockDim.x + threadIdx.x;
  if (f >= rows_pad) return;
  uint4* o4 = reinterpret_cast<uint4*>(har + ((int64_t)b * rows_pad + f) * 64);
  if (f >= frames) {
#pragma unroll
    for (int i = 0; i < 8; ++i) o4[i] = make_uint4(0u, 0u, 0u, 0u);
    return;
  }
  const float* x = hs + (int64_t)b * Lw;
  float xw[20];
#pragma unroll
  for (int n = 0; n < 20; ++n) {
    int idx = 5 * f + n - 10;
    if (idx < 0) idx = -idx;
    if (idx >= Lw) idx = 2 * (Lw - 1) - idx;
    xw[n] = x[idx] * c_hann20[n];
  }
  float val[32];
#pragma unroll
  for (int c = 22; c < 32; ++c) val[c] = 0.f;
#pragma unroll
  for (int k = 0; k <= 10; ++k) {
    float re = 0.f, im = 0.f;
#pragma unroll
    for (int n = 0; n < 20; ++n) {
      const int m = (k * n) % 20;
      re = fmaf(xw[n], c_cos20[m], re);
      im = fmaf(xw[n], -c_sin20[m], im);
    }
    if (k == 0 || k == 10) im = 0.f;
    val[k] = sqrtf(re * re + im * im);
    val[11 + k] = atan2f(im, re);
  }
  uint32_t hi[16], lo[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) {
    const __nv_bfloat16 h0 = __float2bfloat16_rn(val[2 * i]), h1 = __float2bfloat16_rn(val[2 * i + 1]);
    const __nv_bfloat16 l0 = __float2bfloat16_rn(val[2 * i] - __bfloat162float(h0));
    const __nv_bfloat16 l1 = __float2bfloat16_rn(val[2 * i + 1] - __bfloat162float(h1));
    hi[i] = (uint32_t)__bfloat16_as_ushort(h0) | ((uint32_t)__bfloat16_as_ushort(h1) << 16);
    lo[i] = (uint32_t)__bfloat16_as_ushort(l0) | ((uint32_t)__bfloat16_as_ushort(l1) << 16);
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    o4[i] = make_uint4(hi[4 * i], hi[4 * i + 1], hi[4 * i + 2], hi[4 * i + 3]);
    o4[4 + i] = make_uint4(lo[4 * i], lo[4 * i + 1], lo[4 * i + 2], lo[4 * i + 3]);
  }
}

int launch_stft_hilo(const float* hs, int B, int Lw, int rows_pad, __nv_bfloat16* har, cudaStream_t st) {
  int rc = ensure_tables();
  if (rc) return rc;
  const int frames = Lw / 5 + 1;
  ST2_REQUIRE(Lw % 5 == 0 && Lw >= 20 && rows_pad >= frames, "stft_hilo: shape");
  dim3 grid(ceil_div(rows_pad, 128), B);
  stft_hilo_kernel<<<grid, 128, 0, st>>>(hs, Lw, frames, rows_pad, har);
  ST2_LAUNCH_OK("stft_hilo");
  return ST2_OK;
}

int launch_stft(const float* hs, int B, int Lw, float* har_cl, int ld, cudaStream_t st) {
  int rc = ensure_tables();
  if (rc) return rc;
  ST2_REQUIRE(Lw % 5 == 0 && Lw >= 20 && ld >= 22, "stft: shape");
  const int frames = Lw / 5 + 1;
  dim3 grid(ceil_div(frames, 128), B);
  stft_kernel<<<grid, 128, 0, st>>>(hs, Lw, frames, har_cl, ld);
  ST2_LAUNCH_OK("stft");
  return ST2_OK;
}

// ---------------------------------------------------------------------------- iSTFT
constexpr int IS_NS = 1024;            // output samples per block
constexpr int IS_MAXF = IS_NS / 5 + 8; // frames touching them

// Block = IS_NS output samples of one item.  Phase 0: the block's rows ([frames][ld] fp32, 22 of them used) come in as
// 128-bit loads, four rows per warp instruction and every load of the block in flight at once (the kernel is
// HBM-latency-bound: one 88-byte load per warp and iteration kept too few bytes in flight, ncu: 7.3 warps per issue
// stalled on long scoreboard).  Phase 1, one thread per frame: exp / sin heads (istftnet.py:377-378), then the inverse
// 20-point DFT of the frame entirely in registers -- the twiddles are compile-time indices into constant memory (FFMA
// constant operands, no table loads), and the symmetry x[20 - n] = A_n + B_n, x[n] = A_n - B_n halves the arithmetic --
// and the 20 windowed samples go back to the frame's row.  Phase 2, one thread per output sample: overlap-add of the 4
// frames that cover it, division by the window envelope, trimmed to [10, 10 + Lw).  Row stride IS_LD = 37 floats
// (= 5 mod 32): conflict-free for the row-per-thread accesses of phase 1 AND for phase 2, whose 32 consecutive samples
// read rows i / 5 at column i % 5 (+ 5 j): bank 5 (i / 5) + i % 5 = i.  FAST = SFU exp / sin (bf16 path).
constexpr int IS_LD = 37;
template <bool FAST>
__global__ void __launch_bounds__(256)
istft_kernel(const float* __restrict__ x, int ld, int frames, float* __restrict__ wav, bool vec) {
  __shared__ float fo[IS_MAXF][IS_LD];  // row fr: first the frame's head inputs, then (same thread) its 20 samples
  __shared__ float s_w2[20];
  const int b = blockIdx.y;
  const int Lw = 5 * (frames - 1);
  const int i0 = blockIdx.x * IS_NS;
  const int i1 = min(Lw, i0 + IS_NS);
  if (threadIdx.x < 20) s_w2[threadIdx.x] = c_hann20[threadIdx.x] * c_hann20[threadIdx.x];
  // frames f with 0 <= (i+10) - 5f <= 19 for some i in [i0, i1): f in [(i+10)/5 - 3, (i+10)/5]
  const int fmin = max(0, (i0 + 10) / 5 - 3);
  const int fmax = min(frames - 1, (i1 - 1 + 10) / 5);
  const int nf = fmax - fmin + 1;
  const float* base = x + ((int64_t)b * frames + fmin) * ld;
  if (vec) {  // ld == 32, 16-byte aligned: rows of 128 bytes: lane -> (row l / 8 of 4, columns 4 (l % 8) ..): all loads first, then the stores
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int rq = lane >> 3, c4 = (lane & 7) * 4;
    constexpr int NIT = (IS_MAXF + 31) / 32;
    float4 v[NIT];
#pragma unroll
    for (int it = 0; it < NIT; ++it) {
      const int r = (it * 8 + warp) * 4 + rq;
      v[it] = r < nf ? __ldg(reinterpret_cast<const float4*>(base + (int64_t)r * 32 + c4)) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
#pragma unroll
    for (int it = 0; it < NIT; ++it) {
      const int r = (it * 8 + warp) * 4 + rq;
      if (r < nf) { fo[r][c4] = v[it].x; fo[r][c4 + 1] = v[it].y; fo[r][c4 + 2] = v[it].z; fo[r][c4 + 3] = v[it].w; }
    }
  } else {  // any other row stride: one coalesced load per row (lane = column)
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (lane < 22)
      for (int r = warp; r < nf; r += 8) fo[r][lane] = __ldg(base + (int64_t)r * ld + lane);
  }
  __syncthreads();
  for (int fr = threadIdx.x; fr < nf; fr += blockDim.x) {
    float re[11], im[11];
#pragma unroll
    for (int k = 0; k <= 10; ++k) {
      const float lx = fo[fr][k], px = fo[fr][11 + k];
      const float spec = FAST ? __expf(lx) : expf(lx);   // istftnet.py:377
      const float ph = FAST ? __sinf(px) : sinf(px);     // istftnet.py:378
      float sn, cs;
      if (FAST) __sincosf(ph, &sn, &cs);
      else sincosf(ph, &sn, &cs);
      const float ck = (k == 0 || k == 10) ? 0.05f : 0.1f;  // c_k / n_fft
      re[k] = spec * cs * ck;
      im[k] = spec * sn * ck;
    }
#pragma unroll
    for (int n = 0; n <= 10; ++n) {
      float A = 0.f, Bn = 0.f;
#pragma unroll
      for (int k = 0; k <= 10; ++k) {
        const int mm = (k * n) % 20;
        A = fmaf(re[k], c_cos20[mm], A);
        Bn = fmaf(im[k], c_sin20[mm], Bn);
      }
      fo[fr][n] = (A - Bn) * c_hann20[n];
      if (n >= 1 && n <= 9) fo[fr][20 - n] = (A + Bn) * c_hann20[20 - n];
    }
  }
  __syncthreads();
  for (int i = i0 + threadIdx.x; i < i1; i += blockDim.x) {
    const int m = i + 10;
    const int fq = m / 5, r = m - 5 * fq;
    float val = 0.f, env = 0.f;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int f = fq - j, n = r + 5 * j;
      if (f < 0 || f >= frames) continue;
      val += fo[f - fmin][n];
      env += s_w2[n];
    }
    wav[(int64_t)b * Lw + i] = val / env;
  }
}

int launch_istft(const float* x_cl, int ld, int B, int frames, float* wav, cudaStream_t st, bool fast) {
  int rc = ensure_tables();
  if (rc) return rc;
  ST2_REQUIRE(frames >= 5 && ld >= 22, "istft: shape");
  const int Lw = 5 * (frames - 1);
  dim3 grid(ceil_div(Lw, IS_NS), B);
  const bool vec = ld == 32 && (reinterpret_cast<uintptr_t>(x_cl) & 15u) == 0;
  if (fast) istft_kernel<true><<<grid, 256, 0, st>>>(x_cl, ld, frames, wav, vec);
  else istft_kernel<false><<<grid, 256, 0, st>>>(x_cl, ld, frames, wav, vec);
  ST2_LAUNCH_OK("istft");
  return ST2_OK;
}

}  // namespace st2

// ------------------------------------------------------------------------------------------------------------------
// st2_source_har_static: st2_source_har without a caller-provided phase scratch and without any allocation.
// No counterpart in the reference; added for callers that cannot call
// cudaMalloc (CUDA-graph capture, tight serving loops, harnesses that forbid it).
// The frame-rate phase table [B][9][T] of phase_scan lives in this file-scope device array (4 MiB per GPU that
// loads the library, held until exit).  In host code the array's name is only a stand-in for the symbol: its device
// address is looked up once per device with cudaGetSymbolAddress (that device current) and cached.  phase_scan
// overwrites every element har_source reads, so nothing has to be cleared between calls; all calls share the one
// array, so two calls that overlap on the GPU (different streams) race.
#define ST2_STATIC_PHASE_FLOATS (1 << 20)
static __device__ __align__(256) float g_static_phase[ST2_STATIC_PHASE_FLOATS];
namespace st2 {
static float* g_static_phase_ptr[64] = {};
static int static_phase_ptr(float** out) {
  int dev = 0;
  ST2_CUDA_OK(cudaGetDevice(&dev));
  ST2_REQUIRE(dev >= 0 && dev < 64, "st2_source_har_static: device index out of range");
  if (!g_static_phase_ptr[dev]) {
    float* p = nullptr;
    ST2_CUDA_OK(cudaGetSymbolAddress((void**)&p, g_static_phase));
    ST2_REQUIRE(p != nullptr, "st2_source_har_static: no device address for the scratch array");
    g_static_phase_ptr[dev] = p;
  }
  *out = g_static_phase_ptr[dev];
  return ST2_OK;
}
}  // namespace st2

extern "C" int st2_source_har_static(const float* f0, const float* noise, uint64_t seed, const float* lin_w9,
                                     float lin_b, int B, int T, int spf, float* har_source, void* stream) {
  using namespace st2;
  ST2_REQUIRE(f0 && lin_w9 && har_source, "st2_source_har_static: null pointer");
  ST2_REQUIRE(B >= 1 && T >= 1 && spf >= 1, "st2_source_har_static: empty shape");
  if ((int64_t)B * NH * T > (int64_t)ST2_STATIC_PHASE_FLOATS) {
    set_error("st2_source_har_static: B * 9 * T = " + std::to_string((int64_t)B * NH * T) +
              " floats exceeds the static scratch of " + std::to_string((int64_t)ST2_STATIC_PHASE_FLOATS) +
              " floats (B * T <= 116508 frames); use st2_source_har with a caller-owned scratch");
    return ST2_EINVAL;
  }
  float* phase = nullptr;
  int rc = static_phase_ptr(&phase);
  if (rc) return rc;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  rc = launch_phase_scan(f0, B, T, spf, phase, st);
  if (rc) return rc;
  return launch_har_source(f0, phase, noise, seed, lin_w9, lin_b, B, T, spf, har_source, st, 0, nullptr);
}
