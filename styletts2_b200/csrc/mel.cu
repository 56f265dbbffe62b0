// mel.cu — the reference's log-mel front end on the device (the quality metric the
// reference itself trains and evaluates with).
//
// Reference: /root/reference/meldataset.py:35-43
//   to_mel = torchaudio.transforms.MelSpectrogram(n_mels=80, n_fft=2048, win_length=1200, hop_length=300)
//   mel = (log(1e-5 + to_mel(wave)) - mean) / std,  mean = -4, std = 4
// torchaudio's defaults for everything the call leaves out: sample_rate 16000 (so the filterbank spans 0..8000 Hz over
// the 1025 bins), f_min 0, power 2, centre = True with reflect padding, periodic Hann window of win_length centred in
// the n_fft frame, HTK mel scale, no filter normalisation.
//
// One CTA per (frame, item): 2048 windowed samples -> radix-2 FFT in shared memory -> |X|^2 of bins 0..1024 -> 80
// triangular filters (sparse table built on the host in double) -> log / normalise.
#include <math.h>

#include <mutex>
#include <vector>

#include "common.cuh"

namespace st2 {

constexpr int MEL_NFFT = 2048, MEL_WIN = 1200, MEL_HOP = 300, MEL_N = 80, MEL_BINS = MEL_NFFT / 2 + 1;
constexpr int MEL_MAX_W = 4096;  // non-zero filter weights (about 2 per bin)
constexpr int MEL_THREADS = 256;

__constant__ int c_mel_start[MEL_N];  // first bin of filter m
__constant__ int c_mel_cnt[MEL_N];    // bins it covers
__constant__ int c_mel_off[MEL_N];    // offset of its weights in g_mel_w
__device__ float g_mel_w[MEL_MAX_W];
static bool g_mel_ready[64] = {};
static std::mutex g_mel_mu;

// torchaudio.functional.melscale_fbanks(n_freqs 1025, f_min 0, f_max sample_rate // 2, n_mels 80, sample_rate 16000,
// norm None, mel_scale "htk"): fb[k][m] = max(0, min(down_slope, up_slope)), in double, stored per filter
static int mel_tables() {
  int dev = 0;
  ST2_CUDA_OK(cudaGetDevice(&dev));
  std::lock_guard<std::mutex> lk(g_mel_mu);
  if (dev < 64 && g_mel_ready[dev]) return ST2_OK;
  const double f_max = 8000.0;  // sample_rate // 2 of torchaudio's default sample_rate
  auto hz2mel = [](double f) { return 2595.0 * log10(1.0 + f / 700.0); };
  std::vector<double> f_pts(MEL_N + 2);
  const double m_lo = hz2mel(0.0), m_hi = hz2mel(f_max);
  for (int i = 0; i < MEL_N + 2; ++i) {
    const double m = m_lo + (m_hi - m_lo) * i / (MEL_N + 1);
    f_pts[i] = 700.0 * (pow(10.0, m / 2595.0) - 1.0);
  }
  std::vector<int> start(MEL_N), cnt(MEL_N), off(MEL_N);
  std::vector<float> w;
  for (int m = 0; m < MEL_N; ++m) {
    int first = -1, last = -1;
    std::vector<float> wm;
    for (int k = 0; k < MEL_BINS; ++k) {
      const double f = f_max * k / (MEL_BINS - 1);
      const double down = (f - f_pts[m]) / (f_pts[m + 1] - f_pts[m]);
      const double up = (f_pts[m + 2] - f) / (f_pts[m + 2] - f_pts[m + 1]);
      const double v = fmax(0.0, fmin(down, up));
      if (v > 0.0) {
        if (first < 0) first = k;
        last = k;
      }
      if (first >= 0) wm.push_back((float)v);
    }
    start[m] = first < 0 ? 0 : first;
    cnt[m] = first < 0 ? 0 : last - first + 1;
    off[m] = (int)w.size();
    for (int i = 0; i < cnt[m]; ++i) w.push_back(wm[i]);
  }
  ST2_REQUIRE((int)w.size() <= MEL_MAX_W, "mel filter table too large");
  ST2_CUDA_OK(cudaMemcpyToSymbol(c_mel_start, start.data(), sizeof(int) * MEL_N));
  ST2_CUDA_OK(cudaMemcpyToSymbol(c_mel_cnt, cnt.data(), sizeof(int) * MEL_N));
  ST2_CUDA_OK(cudaMemcpyToSymbol(c_mel_off, off.data(), sizeof(int) * MEL_N));
  ST2_CUDA_OK(cudaMemcpyToSymbol(g_mel_w, w.data(), sizeof(float) * w.size()));
  if (dev < 64) g_mel_ready[dev] = true;
  return ST2_OK;
}

// grid (frames, B), block 256.  out [B][80][frames]
__global__ void __launch_bounds__(MEL_THREADS) log_mel_kernel(const float* __restrict__ wav, int64_t L, int frames,
                                                               float* __restrict__ out) {
  __shared__ float re[MEL_NFFT], im[MEL_NFFT];
  __shared__ float2 tw[MEL_NFFT / 2];  // exp(-2 pi i k / 2048)
  pdl_wait_then_trigger();
  const int f = blockIdx.x, b = blockIdx.y, tid = threadIdx.x;
  const float* w = wav + (int64_t)b * L;
  for (int k = tid; k < MEL_NFFT / 2; k += MEL_THREADS) {
    float s, c;
    sincospif(-(float)k / (float)(MEL_NFFT / 2), &s, &c);
    tw[k] = make_float2(c, s);
  }
  // frame sample n = wave sample f*hop - n_fft/2 + n (reflect padding), window = periodic Hann(1200) centred in 2048;
  // stored in bit-reversed order for the in-place decimation-in-time FFT
  constexpr int W0 = (MEL_NFFT - MEL_WIN) / 2;
  for (int n = tid; n < MEL_NFFT; n += MEL_THREADS) {
    float v = 0.f;
    if (n >= W0 && n < W0 + MEL_WIN) {
      int64_t t = (int64_t)f * MEL_HOP - MEL_NFFT / 2 + n;
      if (t < 0) t = -t;
      if (t >= L) t = 2 * (L - 1) - t;
      const float win = 0.5f - 0.5f * cospif(2.f * (float)(n - W0) / (float)MEL_WIN);
      v = w[t] * win;
    }
    const int r = (int)(__brev((unsigned)n) >> (32 - 11));
    re[r] = v;
    im[r] = 0.f;
  }
  __syncthreads();
#pragma unroll 1
  for (int half = 1; half < MEL_NFFT; half <<= 1) {
    const int tstep = (MEL_NFFT / 2) / half;
    for (int j = tid; j < MEL_NFFT / 2; j += MEL_THREADS) {
      const int pos = j & (half - 1), i0 = ((j - pos) << 1) + pos, i1 = i0 + half;
      const float2 t = tw[pos * tstep];
      const float xr = re[i1] * t.x - im[i1] * t.y, xi = re[i1] * t.y + im[i1] * t.x;
      const float ar = re[i0], ai = im[i0];
      re[i0] = ar + xr; im[i0] = ai + xi;
      re[i1] = ar - xr; im[i1] = ai - xi;
    }
    __syncthreads();
  }
  for (int k = tid; k < MEL_BINS; k += MEL_THREADS) re[k] = re[k] * re[k] + im[k] * im[k];  // power spectrum
  __syncthreads();
  if (tid < MEL_N) {
    const int s0 = c_mel_start[tid], n = c_mel_cnt[tid];
    const float* fw = g_mel_w + c_mel_off[tid];
    float acc = 0.f;
    for (int i = 0; i < n; ++i) acc = fmaf(fw[i], re[s0 + i], acc);
    out[((int64_t)b * MEL_N + tid) * frames + f] = (logf(1e-5f + acc) + 4.f) * 0.25f;
  }
}

int launch_log_mel(const float* wav, int B, int64_t L, float* out, cudaStream_t st) {
  int rc = mel_tables();
  if (rc) return rc;
  const int frames = (int)(L / MEL_HOP) + 1;
  ST2_CUDA_OK(launch_pdl(log_mel_kernel, dim3(frames, B), dim3(MEL_THREADS), 0, st, wav, L, frames, out));
  ST2_LAUNCH_OK("log_mel");
  return ST2_OK;
}

}  // namespace st2

extern "C" int st2_log_mel(const float* wav, int B, int64_t L, float* mel_out, void* stream) {
  ST2_REQUIRE(wav && mel_out && B >= 1 && B <= 65535, "st2_log_mel: bad argument");
  ST2_REQUIRE(L > st2::MEL_NFFT / 2, "st2_log_mel: reflect padding needs more than n_fft / 2 = 1024 samples per item");
  ST2_REQUIRE(L / st2::MEL_HOP + 1 <= 2147483647ll / 4, "st2_log_mel: too many frames");
  return st2::launch_log_mel(wav, B, L, mel_out, reinterpret_cast<cudaStream_t>(stream));
}
extern "C" int st2_log_mel_frames(int64_t L) { return (int)(L / st2::MEL_HOP) + 1; }
