/*
 * st2_decoder.h — C ABI of the B200-native StyleTTS2 waveform decoder (libst2dec.so).
 *
 * Drop-in boundary.  The reference exposes this path as a torch.nn.Module:
 *   Decoder.__init__  /root/reference/Modules/istftnet.py:467-496  (hifigan.py:416-442)
 *   Decoder.forward   /root/reference/Modules/istftnet.py:498-527  (hifigan.py:445-474)
 *   wiring            /root/reference/models.py:1002-1032 (build_model), :1218-1239 (load_checkpoint)
 * There is no FFI in the reference; the boundary is the constructor kwargs, forward(asr,
 * F0_curve, N, s) and the state_dict key set.  This header is what a binding for that
 * boundary calls (the Python module in styletts2_b200/ binds it with ctypes; see
 * INTEGRATION.md for the reference-side stub).
 *
 * Conventions: plain C, no torch types.  Every function returns 0 on success or a
 * negative ST2_E* code and never throws; st2_last_error() gives a thread-local message.
 * All device pointers belong to the caller (torch allocates wav_out and the workspace).
 * A handle is bound to the CUDA device current at create time and is immutable after load.
 * One forward at a time per handle: callers serialise the calls on ONE handle (the Python
 * module holds a lock per module; the per-call launch count lives in the handle); different
 * handles (one per device / per thread, e.g. the one-thread-per-device calling convention
 * of train_finetune.py:96-101) run concurrently.  All one-time per-device setup (identity
 * blobs, constant tables, kernel attributes) happens in st2_decoder_load: forward allocates
 * nothing and never synchronises the host, so it can be stream-captured.  A decoder handle owns a
 * second stream + events (made by load): a single-utterance forward forks its excitation branch
 * onto it and joins it back, so the caller's stream still orders the whole call.  sm_100a only; there is no
 * CPU fallback: without a usable GPU every compute entry point fails with ST2_ECUDA.
 */
#ifndef ST2_DECODER_H
#define ST2_DECODER_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ST2_OK 0
#define ST2_EINVAL (-1)  /* bad argument / shape / missing weight */
#define ST2_ECUDA (-2)   /* CUDA runtime error (message has the detail) */
#define ST2_ESTATE (-3)  /* call order (forward before load, ...) */
#define ST2_ENOMEM (-4)  /* workspace too small */

#define ST2_KIND_ISTFTNET 0 /* Modules/istftnet.py */
#define ST2_KIND_HIFIGAN 1  /* Modules/hifigan.py  */
#define ST2_KIND_F0N 2      /* ProsodyPredictor.F0Ntrain (models.py:562-590, 638-654): the step before the decoder.
                               St2Config: dim_in = d_hid (512), style_dim; the other fields are ignored.  Uses
                               create / load / num_keys / key / destroy of the decoder API and st2_f0n_* below. */

#define ST2_KIND_SAMPLER 3  /* style sampler (next row N2): DiffusionSampler + ADPM2Sampler + KarrasSchedule over the
                               KDiffusion denoiser StyleTransformer1d / Transformer1d (Modules/diffusion/sampler.py:480-585,
                               Modules/diffusion/modules.py:40-426, built by models.py:1057-1089).  St2Config fields:
                               dim_in = channels (2 * style_dim = 256), style_dim = context_features (256: StyleTransformer1d,
                               multispeaker; 0: Transformer1d with plain LayerNorm), dim_out = context_embedding_features
                               (PL-BERT hidden size, 768), n_resblock_kernels = num_layers, resblock_kernel_sizes[0..3] =
                               num_heads, head_features, multiplier, embedding_max_length; the rest is ignored.  Uses create /
                               load / num_keys / key / destroy of the decoder API and st2_sampler_* below. */

/* arithmetic of the dense contractions */
#define ST2_PREC_FP32 0 /* fp32 FFMA kernels, fp32 activations: the parity path        */
#define ST2_PREC_BF16 1 /* tcgen05 bf16 x bf16 -> fp32 (TMEM), bf16 activations in HBM */
#define ST2_PREC_FP32_TC 2 /* the parity path on the tensor cores: fp32 activations, every dense conv as three bf16
                              products x_hi w_hi + x_lo w_hi + x_hi w_lo with fp32 accumulation in TMEM (decoders and F0N;
                              needs st2_decoder_set_option(h, "fp32_tc", 1) before load: 3x the bf16 weight bytes) */

#define ST2_MAX_UPS 4
#define ST2_MAX_RESK 4

/* Constructor kwargs of the reference Decoder (istftnet.py:467-473, hifigan.py:416-421). */
typedef struct St2Config {
  int kind;                 /* ST2_KIND_* */
  int dim_in;               /* 512 (hidden_dim) */
  int style_dim;            /* 128 */
  int dim_out;              /* 80, accepted and unused like the reference */
  int n_resblock_kernels;   /* len(resblock_kernel_sizes) = 3 */
  int resblock_kernel_sizes[ST2_MAX_RESK];      /* 3,7,11 */
  int resblock_dilations[ST2_MAX_RESK][3];      /* 1,3,5 each */
  int n_ups;                /* len(upsample_rates): 2 (istftnet) or 4 (hifigan) */
  int upsample_rates[ST2_MAX_UPS];              /* 10,6 | 10,5,3,2 */
  int upsample_kernel_sizes[ST2_MAX_UPS];       /* 20,12 | 20,10,6,4 */
  int upsample_initial_channel;                 /* 512 */
  int gen_istft_n_fft;      /* 20 (istftnet only) */
  int gen_istft_hop_size;   /* 5  (istftnet only) */
} St2Config;

/* One state_dict entry (host fp32, contiguous, torch layout).  Names are the reference's
 * keys ("decode.0.conv1.weight_v", ...); a leading "module." (DDP) is ignored like
 * models.py:1126-1142 does. */
typedef struct St2Tensor {
  const char* name;
  const float* data;
  int ndim;
  int64_t shape[4];
} St2Tensor;

typedef struct St2Decoder St2Decoder; /* opaque */

/* Optional inputs of one forward call (all device pointers, may be NULL). */
typedef struct St2ForwardOpts {
  int precision;            /* ST2_PREC_* */
  const float* noise;       /* [B, T*spf, 9] standard normal: the randn_like draw of
                               SineGen.forward (istftnet.py:241).  NULL -> in-kernel Philox */
  const float* har_source;  /* [B, T*spf] output of m_source (istftnet.py:353): skips SineGen */
  const float* har;         /* [B, 22, 60T+1] NCL: cat(|STFT|, angle) (istftnet.py:356); istftnet only */
  uint64_t seed;            /* Philox seed when noise == NULL */
  float* har_source_out;    /* optional [B, T*spf]: the excitation that was used */
  float* har_out;           /* optional [B, 22, 60T+1] NCL: the har that was used */
  float* tap_decode;        /* optional [B, 512, T] NCL fp32: output of decode[3] (istftnet.py:522) */
  float* tap_stage[ST2_MAX_UPS]; /* optional [B, C_i, L_i] NCL fp32: x after generator stage i (:374) */
  int64_t item_offset;      /* global index of row 0 (Philox counter = (item, sample)): a shard or a
                               micro-batch of a larger batch draws the noise the full batch would */
  const uint64_t* seed_dev; /* optional DEVICE pointer to the Philox seed, read by the kernel instead of
                               `seed`: lets a captured CUDA graph draw fresh noise on every replay */
  const int32_t* align_idx; /* optional [B, T/2] token index of every asr frame (st2_alignment_indices).  When set,
                               the `asr` argument is the TOKEN-rate encoding t_en [B, dim_in, align_len] and the
                               expansion asr = t_en @ pred_aln_trg (Inference_LJSpeech.ipynb cell 17) is fused
                               into the decoder's input load; -1 = a zero frame */
  int align_len;            /* tokens per item (Ltext) of t_en */
} St2ForwardOpts;

/* Decoder(**kwargs): validates the config and builds the layer plan (no weights yet). */
int st2_decoder_create(const St2Config* cfg, St2Decoder** out);

/* Options.  "fp32_tc" = 1 changes what load builds (call between create and load; after load -> ST2_ESTATE): also pack
 * the (hi, hi, lo) bf16 blobs ST2_PREC_FP32_TC needs.  No counterpart in the reference (its precision is torch's
 * dtype / autocast context, train_second.py:183-186).  Schedule switches, any time between forwards, results are
 * bit-identical for every value: "fork" = 0 / 1 / 2: the excitation branch (source -> STFT -> noise_convs -> noise_res)
 * on the handle's second stream never / for single-utterance-sized forwards (default) / always; "group" = 0 / 1:
 * independent same-shaped convs of such a forward as one multi-conv launch (default 1).  Defaults can also come from
 * ST2_FORK / ST2_GROUP in the environment at create time.  Unknown names -> ST2_EINVAL. */
int st2_decoder_set_option(St2Decoder* h, const char* name, int value);

/* load_state_dict: folds weight-norm (w = g*v/||v||, dim 0), repacks every conv into the
 * kernel layouts (fp32 tap-major and bf16 UMMA core-matrix blobs), expands ConvTranspose1d
 * into polyphase form, concatenates all AdaIN fc layers, and uploads to the current device.
 * Missing or mis-shaped keys -> ST2_EINVAL (strict, like load_state_dict). */
int st2_decoder_load(St2Decoder* h, const St2Tensor* tensors, int n);

/* Number of state-dict keys the decoder expects / the i-th expected key and shape. */
int st2_decoder_num_keys(const St2Decoder* h);
int st2_decoder_key(const St2Decoder* h, int i, const char** name, int* ndim, int64_t shape[4]);

/* samples generated per mel frame (300 for both shipped configs). */
int st2_decoder_samples_per_frame(const St2Decoder* h);

/* Bytes of device scratch forward() needs for this batch/length/precision. */
int64_t st2_decoder_workspace_bytes(const St2Decoder* h, int B, int T, int precision);

/* Decoder.forward(asr, F0_curve, N, s) in eval mode.
 *   asr [B, dim_in, T/2] fp32 NCL, f0 [B, T], n [B, T], s [B, style_dim]  (device, contiguous)
 *   wav_out [B, 1, T*spf] fp32 (device).  T must be even (the reference's cat at
 *   istftnet.py:513 raises otherwise).  stream is a cudaStream_t. */
int st2_decoder_forward(St2Decoder* h, const float* asr, const float* f0, const float* n,
                        const float* s, float* wav_out, void* workspace, int64_t workspace_bytes,
                        int B, int T, const St2ForwardOpts* opts, void* stream);

/* ProsodyPredictor.F0Ntrain(x, s) (models.py:651-654) on a handle created with kind ST2_KIND_F0N:
 *   x [B, d_hid + style_dim, T] fp32 NCL (duration-expanded prosody encoding `en`), s [B, style_dim]
 *   f0_out, n_out [B, 2T] fp32: the F0 curve and energy the decoder consumes (F0_curve, N).
 * bidirectional LSTM (input projection as one dense conv + a cluster-resident recurrence kernel),
 * then per branch three AdainResBlk1d (the middle one upsamples x2) and a 1x1 projection. */
int64_t st2_f0n_workspace_bytes(const St2Decoder* h, int B, int T, int precision);
int st2_f0n_forward(St2Decoder* h, const float* x, const float* s, float* f0_out, float* n_out,
                    void* workspace, int64_t workspace_bytes, int B, int T, int precision, void* stream);

/* Style sampler (next row N2) on a handle created with kind ST2_KIND_SAMPLER.
 *   s_pred = sampler(noise, embedding=bert_dur, embedding_scale=..., features=ref_s, num_steps=...)
 *   (Demo/Inference_LibriTTS.ipynb cells 14, 16; DiffusionSampler.forward sampler.py:573-585). */
typedef struct St2SamplerOpts {
  int precision;          /* ST2_PREC_* of the token-level contractions (to_q, to_kv, to_out, feed_forward) */
  int num_steps;          /* diffusion_steps: num_steps - 1 ADPM2 steps = 2 * (num_steps - 1) denoiser calls */
  float sigma_min, sigma_max, rho; /* KarrasSchedule(sigma_min=1e-4, sigma_max=3.0, rho=9.0) */
  float sampler_rho;      /* ADPM2Sampler(rho=1.0) */
  float sigma_data;       /* KDiffusion.sigma_data (0.2) */
  float embedding_scale;  /* classifier-free guidance scale; != 1 runs the denoiser a second time on the
                             fixed embedding (modules.py:178-184), batched with the first run */
  int clamp;              /* DiffusionSampler(clamp=...): clamp the result to [-1, 1] */
  const float* step_noise;/* optional DEVICE [num_steps - 1, B, channels]: the randn_like draw of every
                             ADPM2 step (sampler.py:510); NULL -> in-kernel Philox(seed, item, channel, step) */
  uint64_t seed;
  int64_t item_offset;    /* global index of row 0 (shards draw the noise the full batch would) */
} St2SamplerOpts;

/* guided != 0 when embedding_scale != 1 (twice the token rows). */
int64_t st2_sampler_workspace_bytes(const St2Decoder* h, int B, int n_tokens, int precision, int guided);
/* noise [B, channels], embedding [B, n_tokens, emb_features], features [B, context_features] (NULL for the
 * Transformer1d variant) -> out [B, channels]; all device fp32, contiguous. */
int st2_sampler_sample(St2Decoder* h, const float* noise, const float* embedding, const float* features, float* out,
                       void* workspace, int64_t workspace_bytes, int B, int n_tokens, const St2SamplerOpts* opts,
                       void* stream);
/* One denoiser evaluation, KDiffusion.denoise_fn(x_noisy, sigma=sigma, ...) (sampler.py:193-208):
 * x_noisy [B, channels] -> out [B, channels].  Uses precision, sigma_data and embedding_scale of opts. */
int st2_sampler_denoise(St2Decoder* h, const float* x_noisy, float sigma, const float* embedding, const float* features,
                        float* out, void* workspace, int64_t workspace_bytes, int B, int n_tokens,
                        const St2SamplerOpts* opts, void* stream);

/* Duration-based alignment (next row N3; Demo/Inference_LJSpeech.ipynb cell 17, Inference_LibriTTS.ipynb cell 16).
 *   dur [B, Ltext] int32 durations (frames per token, device) -> idx [B, frames] int32: the token shown by each
 *   frame, -1 past the item's total; totals [B] (optional) = sum of durations.  shift_one != 0 applies the
 *   HiFi-GAN one-frame right shift (frame 0 repeated).
 * st2_expand_alignment materialises x @ one_hot: out[b, c, f] = x[b, c, idx[b, f]] (0 where idx < 0), NCL fp32. */
int st2_alignment_indices(const int32_t* dur, int B, int Ltext, int frames, int shift_one, int32_t* idx,
                          int32_t* totals, void* stream);
int st2_expand_alignment(const float* x, const int32_t* idx, int B, int C, int Ltext, int frames, float* out,
                         void* stream);

/* Output stage (next row N4): wav [B, L] fp32 -> out [B, L - trim_tail] int16 PCM,
 * clip(rint(x * 32767), -32768, 32767); trim_tail drops the last samples of every item (the LibriTTS
 * notebook cuts 50: "weird pulse at the end of the model"). */
int st2_wave_to_pcm16(const float* wav, int B, int64_t L, int64_t trim_tail, int16_t* out, void* stream);
/* The reference's log-mel front end (next row N4; /root/reference/meldataset.py:35-43: torchaudio MelSpectrogram(n_mels 80,
 * n_fft 2048, win_length 1200, hop_length 300) with torchaudio's defaults (sample_rate 16000, centre, reflect padding,
 * periodic Hann, power 2, HTK scale, no normalisation), then (log(1e-5 + mel) + 4) / 4).  wav [B, L] fp32 (L > 1024) ->
 * mel_out [B, 80, L / 300 + 1] fp32.  One CTA per frame: 2048-point FFT in shared memory. */
int st2_log_mel(const float* wav, int B, int64_t L, float* mel_out, void* stream);
int st2_log_mel_frames(int64_t L);

/* Number of kernel launches the last forward on this handle issued (bench bookkeeping). */
int64_t st2_decoder_last_launches(const St2Decoder* h);

void st2_decoder_destroy(St2Decoder* h);

const char* st2_last_error(void);
const char* st2_version(void);

/* ---- single-kernel entry points (unit tests and micro-benchmarks call these) ---------- */

/* Generic fused 1-D convolution on channels-last activations.  See csrc/conv_params.h for
 * the meaning of every field; the struct is passed by pointer and copied. */
struct St2ConvParams;
int st2_conv_simt(const struct St2ConvParams* p, void* stream);
int st2_conv_umma(const struct St2ConvParams* p, void* stream);
/* n (1..3) convs of the same geometry (batch, lengths, channel counts, activation kind) as ONE persistent tcgen05 kernel
 * whose CTAs alternate between them tile by tile, so that a conv bound by its per-element prologue / epilogue work
 * (k = 3) overlaps one bound by the tensor pipe (k = 11): the three first convs of the resblocks k = 3 / 7 / 11 read the
 * same x (istftnet.py:369-373).  Results equal n separate st2_conv_umma calls bit for bit; jobs that cannot share a
 * kernel are launched one by one. */
int st2_conv_umma_multi(const struct St2ConvParams* ps, int n, void* stream);
/* Debug counter: st2_conv_simt launches of the calling thread with cin >= 64 at stride 1 (layers the tcgen05 kernel can
 * take) since the last reset; ST2_PREC_FP32_TC must leave it at 0 for the decoders' dense convs. */
int64_t st2_conv_simt_dense_launches(int reset);
/* packs fp32 [k][Cin][N] host weights into the UMMA blob layout (host only, no GPU). */
int64_t st2_umma_packed_elems(int k, int cin_pad, int n);
int st2_umma_pack_weights(const float* w_kcn, int k, int cin, int cin_pad, int n, uint16_t* out_bf16);
/* the same for an explicit tile width n_tile (32 / 64 / 128 dividing n; 0 = the default): St2ConvParams.n_tile */
int st2_umma_pack_weights_nt(const float* w_kcn, int k, int cin, int cin_pad, int n, int n_tile, uint16_t* out_bf16);
/* host-only test hook: folds weight-norm and repacks one Conv1d (transposed=0) or
 * ConvTranspose1d (transposed=1, polyphase form) exactly as st2_decoder_load does. */
int st2_debug_pack_conv(const St2Tensor* tensors, int n, const char* prefix, int transposed, int K,
                        int stride, int padding, float* w_out, int64_t w_cap, float* bias_out,
                        int* k_out, int* pad_out, int* cin_out, int* ncols_out);
/* Per-kernel timing of the tcgen05 conv launches (bench.py's roofline): when enabled, every
 * st2_conv_umma launch of the calling thread is bracketed by CUDA events on its stream.
 * st2_prof_read synchronises those events and returns the summed device time, the number of
 * launches and their algorithmic FLOPs (St2ConvParams.algo_flops); reset != 0 clears the log. */
int st2_prof_enable(int on);
int st2_prof_read(double* umma_ms, int64_t* umma_launches, double* umma_flops, int reset);
/* Debug: per-role barrier-wait accounting of conv_umma_kernel, filled while ST2_UMMA_DBG has bit 64 set (SM clock
 * cycles summed over all CTAs since the last reset; index meaning in csrc/conv_umma.cu, g_um_wait).  Synchronises. */
int st2_umma_wait_profile(uint64_t out[16], int reset);
/* Debug: the ST2_UMMA_* A/B switches (csrc/conv_umma.cu) are read from the environment once, at the first launch;
 * a tool that changes them afterwards calls this to have them read again.  No counterpart in the reference. */
int st2_umma_reload_env(void);
/* rows of the time tile and partial-statistics tiles each implementation uses. */
int st2_conv_tile_rows(int use_umma);

/* Harmonic source + STFT (istftnet.py:351-356) and exp/sin heads + inverse STFT (:377-379). */
int st2_source_har(const float* f0, const float* noise, uint64_t seed, const float* lin_w9,
                   float lin_b, int B, int T, int spf, float* phase_scratch, float* har_source,
                   void* stream);
/* st2_source_har without the phase_scratch argument and without any allocation (no cudaMalloc / cudaFree, no
 * allocating library call): the [B][9][T] phase table lives in a file-scope static __device__ array inside the
 * library (4 MiB per GPU that loads it, held until exit).  No counterpart in the reference: added
 * for callers that cannot allocate (CUDA-graph capture, tight serving loops, harnesses that
 * forbid cudaMalloc).  Limit: B * 9 * T <= 1 048 576 floats (B * T <= 116 508 frames; config C2 needs 230 400),
 * beyond that ST2_EINVAL with a message.  Every call on a device shares the one array: calls that overlap on the
 * GPU (different streams) race; calls on one stream are ordered.  Nothing needs clearing between calls (the scan
 * overwrites every element the second kernel reads).  Same results as st2_source_har, bit for bit. */
int st2_source_har_static(const float* f0, const float* noise, uint64_t seed, const float* lin_w9,
                          float lin_b, int B, int T, int spf, float* har_source, void* stream);
int st2_stft_har(const float* har_source, int B, int Lw, float* har_cl32, void* stream);
/* the STFT of the bf16 path (what the benchmarked forward launches): every frame is one row of 64 bf16 =
 * [hi(22) | 0(10) | lo(22) | 0(10)], hi = bf16(v), lo = bf16(v - hi); rows [Lw/5+1, rows_pad) are zero. */
int st2_stft_har16(const float* har_source, int B, int Lw, int rows_pad, uint16_t* har_hilo, void* stream);
int st2_istft_head(const float* x_cl32, int B, int frames, float* wav, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* ST2_DECODER_H */
