// common.cuh — shared device/host helpers for libst2dec (sm_100a only).
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include <string>
#include <vector>

#include "../../include/st2_decoder.h"
#include "conv_params.h"

namespace st2 {

void set_error(const std::string& msg);  // decoder.cu (thread-local)
int64_t& launch_counter();               // decoder.cu (thread-local running count)

#define ST2_CUDA_OK(expr)                                                              \
  do {                                                                                 \
    cudaError_t _e = (expr);                                                           \
    if (_e != cudaSuccess) {                                                           \
      st2::set_error(std::string(#expr) + ": " + cudaGetErrorString(_e));              \
      return ST2_ECUDA;                                                                \
    }                                                                                  \
  } while (0)

// Programmatic dependent launch: a kernel that starts with pdl_wait_then_trigger() may be launched with
// the programmatic-stream-serialization attribute (launch_pdl below).  It is then scheduled while its
// predecessor is still running (its launch latency hides behind the predecessor's tail), blocks in
// griddepcontrol.wait until the predecessor has completed and flushed, and only then lets its own
// successor be scheduled -- so nothing it reads can be stale, whatever the predecessor is.
#ifdef __CUDACC__
__device__ __forceinline__ void pdl_wait_then_trigger() {
  asm volatile("griddepcontrol.wait;" ::: "memory");
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
}
#endif
bool pdl_enabled();  // decoder.cu: ST2_PDL=0 turns the attribute off (A/B switch); consumes pdl_plain_next()
void pdl_plain_next();
template <typename... KArgs, typename... Args>
static inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st,
                                     Args... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = pdl_enabled() ? 1 : 0;
  cfg.attrs = attr; cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...);
}

#define ST2_LAUNCH_OK(name)                                                            \
  do {                                                                                 \
    st2::launch_counter()++;                                                           \
    cudaError_t _e = cudaGetLastError();                                               \
    if (_e != cudaSuccess) {                                                           \
      st2::set_error(std::string("launch ") + name + ": " + cudaGetErrorString(_e));   \
      return ST2_ECUDA;                                                                \
    }                                                                                  \
  } while (0)

#define ST2_REQUIRE(cond, msg)                                                         \
  do {                                                                                 \
    if (!(cond)) {                                                                     \
      st2::set_error(std::string("invalid argument: ") + msg + " [" #cond "]");        \
      return ST2_EINVAL;                                                               \
    }                                                                                  \
  } while (0)

static inline int ceil_div(int a, int b) { return (a + b - 1) / b; }
static inline int64_t round_up(int64_t a, int64_t b) { return (a + b - 1) / b * b; }

// ---- dtype-generic element access (fp32 or bf16 storage, fp32 math) -------------------
template <typename T>
__device__ __forceinline__ float ld_f(const T* p);
template <>
__device__ __forceinline__ float ld_f<float>(const float* p) { return *p; }
template <>
__device__ __forceinline__ float ld_f<__nv_bfloat16>(const __nv_bfloat16* p) {
  return __bfloat162float(*p);
}
template <typename T>
__device__ __forceinline__ void st_f(T* p, float v);
template <>
__device__ __forceinline__ void st_f<float>(float* p, float v) { *p = v; }
template <>
__device__ __forceinline__ void st_f<__nv_bfloat16>(__nv_bfloat16* p, float v) {
  *p = __float2bfloat16_rn(v);
}
// value as it will be read back (statistics are taken on the stored value)
template <typename T>
__device__ __forceinline__ float round_to(float v);
template <>
__device__ __forceinline__ float round_to<float>(float v) { return v; }
template <>
__device__ __forceinline__ float round_to<__nv_bfloat16>(float v) {
  return __bfloat162float(__float2bfloat16_rn(v));
}

// pre-activation of the conv prologue (conv_params.h).  PRECISE selects sinf (fp32 parity
// path) vs the SFU approximation (bf16 throughput path).
template <bool PRECISE>
__device__ __forceinline__ float apply_act(float v, int act, float slope, float a, float inv_a) {
  if (act == ST2_ACT_LRELU) return v > 0.f ? v : v * slope;
  if (act == ST2_ACT_SNAKE) {
    float s = PRECISE ? sinf(a * v) : __sinf(a * v);
    return v + inv_a * (s * s);  // x + (1/a) sin^2(a x), istftnet.py:69
  }
  return v;
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// ---- Philox4x32-10 + Box-Muller normal (source.cu: SineGen noise; sampler.cu: ADPM2 step noise) ----
#ifdef __CUDACC__
__device__ __forceinline__ void philox4x32(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3,
                                           uint32_t k0, uint32_t k1, uint32_t out[4]) {
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
    const uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
    c0 = hi1 ^ c1 ^ k0;
    c1 = lo1;
    c2 = hi0 ^ c3 ^ k1;
    c3 = lo0;
    k0 += 0x9E3779B9u;
    k1 += 0xBB67AE85u;
  }
  out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}
__device__ __forceinline__ void box_muller(uint32_t a, uint32_t b, float& z0, float& z1) {
  const float u1 = ((float)a + 1.f) * 2.3283064365386963e-10f;  // (0,1]
  const float u2 = (float)b * 2.3283064365386963e-10f;
  const float r = sqrtf(-2.f * __logf(u1));
  float s, c;
  __sincosf(2.f * 3.14159265358979323846f * u2, &s, &c);
  z0 = r * c;
  z1 = r * s;
}
#endif

// ---- kernel launchers implemented across the .cu files --------------------------------
int launch_conv_simt(const St2ConvParams& p, cudaStream_t st);
int launch_conv_umma(const St2ConvParams& p, cudaStream_t st);
int launch_conv_umma_multi(const St2ConvParams* ps, int n, cudaStream_t st);  // up to 3 same-shaped convs, one kernel
                                                                               // (conv_umma_multi.cu)
void umma_multi_reload_env();
int umma_multi_wait_read(unsigned long long* h16, bool reset);
// per-launch event timing of the tcgen05 conv launches of the calling thread (st2_prof_enable / st2_prof_read)
struct UmProf {
  bool on = false;
  std::vector<cudaEvent_t> ev;  // pairs
  size_t used = 0;
  double flops = 0.0;
};
UmProf& umma_prof();
int conv_tile_rows(bool umma);
int umma_n_tile(int n_cols);  // GEMM columns per tcgen05 tile for a layer of n_cols (0: none)
int umma_prepare(int acc_mult);  // conv_umma's one-time per-device setup (called by st2_decoder_load, never by forward)
int source_prepare();            // the DFT constant tables of source.cu, likewise

int launch_style_fc(const float* s, const float* W, const float* bvec, float* h, int B, int style,
                    int rows, cudaStream_t st);
int launch_stats_finalize(const float* partials, int n_tiles, int n_cols, int C, int64_t count,
                          const float* h, int h_stride, int gamma_off, float* ss, int ss_stride,
                          int c_off, int B, cudaStream_t st, int reps = 1);
int launch_stats_finalize_multi(const float* const* partials, int n, int n_tiles, int n_cols, int C, int64_t count,
                                const float* const* hg, int h_stride, int Cn, float* const* ss, int ss_stride,
                                int c_off, int B, cudaStream_t st, int reps);
template <typename T>
int launch_colstats(const T* x, int64_t bstride, int ld, int L, int c0, int ncols, float* partials,
                    int B, cudaStream_t st, int n_tiles_total = 0);
// fp32 parity path on the tensor cores (elementwise.cu): prologue + (hi, lo, hi) bf16 split, and the fp32 epilogue
int launch_split3(const float* x, int64_t x_bstride, int ldx, int L, int cin, int cin_pad, const float* ss,
                  int ss_stride, int act, float slope, const float* alpha, __nv_bfloat16* y, int B, cudaStream_t st);
int launch_post32(const float* acc, int ldacc, int L, int n_cols, const float* res, int64_t res_bstride, int ldres,
                  int res_shift, float out_scale, int accumulate, float* y, int64_t y_bstride, int ldy, int B,
                  cudaStream_t st);
int colstats_tiles(int L);
template <typename T>
int launch_prep_inputs(const float* asr, const float* f0, const float* n, const float* f0w,
                       const float* nw, T* x0, int ld0, T* ca, T* cb, int ldc, int c_f0, int B,
                       int dim_in, int T2, cudaStream_t st, const int* align_idx = nullptr, int align_len = 0);
int launch_norm_act_bf16(const __nv_bfloat16* x, int64_t x_bstride, int ldx, const float* ss, int ss_stride,
                         float slope, __nv_bfloat16* y, int64_t y_bstride, int ldy, int B, int L, int C,
                         cudaStream_t st);
int launch_wave_to_pcm16(const float* wav, int B, int64_t L, int64_t trim, int16_t* out, cudaStream_t st);
int launch_alignment_idx(const int* dur, int B, int Ltext, int frames, int shift, int* idx, int* totals,
                         cudaStream_t st);
int launch_expand_alignment(const float* x, const int* idx, int B, int C, int Ltext, int frames, float* out,
                            cudaStream_t st);
template <typename T>
int launch_pool_up2(const T* x, int64_t x_bstride, int ldx, const float* ss, int ss_stride,
                    float slope, const float* w3, const float* bias, T* y, int64_t y_bstride, int ldy,
                    int B, int Lin, int C, int Cpad, cudaStream_t st);
template <typename T>
int launch_reflect_row0(const T* xin, int64_t xin_bstride, int cin, float slope, const float* w_tap,
                        const float* bias, const T* xs, int64_t xs_bstride, T* y, int64_t y_bstride,
                        int cout, float* stats_slot, int64_t stats_bstride, int B, cudaStream_t st);
template <typename T>
int launch_act_to_ncl(const T* cl, int64_t bstride, int ld, int C, int L, float* ncl, int B, cudaStream_t st);
int launch_cl_to_ncl(const float* cl, int ld, int C, int L, float* ncl, int B, cudaStream_t st);
int launch_ncl_to_cl(const float* ncl, int C, int L, float* cl, int ld, int B, cudaStream_t st);
// lstm.cu: NCL fp32 -> channels-last activations; recurrence of the bidirectional LSTM (hidden 256)
template <typename AT>
int launch_ncl_to_act(const float* ncl, int C, int L, AT* cl, int ld, int B, cudaStream_t st);
template <typename AT>
int launch_lstm_scan(const AT* pre, const float* whh, AT* out, int B, int T, cudaStream_t st);

int launch_row_mean(const float* x, int64_t L, float* mean, int B, cudaStream_t st);
int launch_noise_bias(const float* mean, const float* wsum, const float* bias_nc, const float* bias_c2,
                      float* bias_nc_shift, float* bias_c2_add, int C, int B, cudaStream_t st, int reps = 1);
int launch_cols_tanh(const float* cp, int ld, int ncol, float* wav, int64_t n, cudaStream_t st);
int launch_phase_scan(const float* f0, int B, int T, int spf, float* phase, cudaStream_t st);
int launch_har_source(const float* f0, const float* phase, const float* noise, uint64_t seed,
                      const float* lin_w9, float lin_b, int B, int T, int spf, float* out,
                      cudaStream_t st, int64_t item0 = 0, const uint64_t* seed_dev = nullptr);
int launch_stft(const float* hs, int B, int Lw, float* har_cl, int ld, cudaStream_t st);
int launch_stft_hilo(const float* hs, int B, int Lw, int rows_pad, __nv_bfloat16* har, cudaStream_t st);
int launch_ncl_to_hilo(const float* ncl, int C, int L, int rows_pad, __nv_bfloat16* har, int B, cudaStream_t st);
int launch_hilo_to_ncl(const __nv_bfloat16* har, int rows_pad, int C, int L, float* ncl, int B, cudaStream_t st);
int launch_istft(const float* x_cl, int ld, int B, int frames, float* wav, cudaStream_t st, bool fast = false);
int launch_tanh_inplace(float* x, int64_t n, cudaStream_t st);

// sampler.cu: style sampler (StyleTransformer1d / Transformer1d denoiser + ADPM2 steps)
int launch_small_linear(const float* x, int ldx, const float* xb, const float* W, const float* bias, float* y, int ldy,
                        int R, int I, int O, int gelu, cudaStream_t st);
int launch_time_fourier(float t, const float* w, int half, float* out, cudaStream_t st);
int launch_build_rows(const float* x, float c_in, const float* emb, const float* fixed, float* X, int B, int Bx, int n,
                      int C, int E, cudaStream_t st);
template <typename AT>
int launch_adaln2(float* X, const AT* delta, const float* mapping, int64_t map_bstride, const float* gb, int64_t gb_bstride,
                  AT* nq, AT* nk, int64_t rows, int n, int B, int F, cudaStream_t st);
template <typename AT>
int launch_add_cast(float* X, const AT* delta, AT* Xc, int64_t n, cudaStream_t st);
template <typename AT>
int launch_gelu(AT* x, int64_t n, cudaStream_t st);
template <typename AT>
int launch_attention(const AT* Q, int ldq, const AT* KV, int ldkv, AT* O, int ldo, int Bx, int n, int heads, int head_dim,
                     cudaStream_t st);
template <typename AT>
int launch_token_mean(const float* X, const AT* delta, float* out, int Bx, int n, int F, cudaStream_t st);
int launch_sampler_step(const float* pred, const float* x_eval, const float* x_base, const float* noise, uint64_t seed,
                        uint32_t step, int64_t item0, float* x_new, float* den_out, int B, int C, int cfg, float cfg_scale,
                        float c_skip, float c_out, float sigma_eval, float dt, float sigma_up, int clamp, cudaStream_t st);
int launch_scale(const float* x, float a, float* y, int n, cudaStream_t st);

}  // namespace st2
