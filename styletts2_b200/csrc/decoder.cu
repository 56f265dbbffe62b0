// decoder.cu — host runtime of libst2dec: handle, load-time weight folding / packing, workspace
// planning and the layer schedule of Decoder.forward (Modules/istftnet.py:498-527,
// Modules/hifigan.py:445-474), plus the extern "C" entry points declared in include/st2_decoder.h.
#include <math.h>
#include <string.h>

#include <map>
#include <string>
#include <vector>

#include <nvtx3/nvToolsExt.h>  // header-only: the ranges are no-ops unless a profiler injects the NVTX library

#include "common.cuh"

namespace st2 {

// NVTX ranges around the stages of a forward (ST2_NVTX=1; read once).  Host-side markers: they bracket the LAUNCHES
// of a stage on the calling thread, which is what a timeline view (nsys / ncu --nvtx) groups kernels by.
static bool nvtx_enabled() {
  static const bool on = [] {
    const char* e = getenv("ST2_NVTX");
    return e && e[0] == '1';
  }();
  return on;
}
struct NvtxRange {
  bool on;
  explicit NvtxRange(const char* name, bool enable = true) : on(enable && nvtx_enabled()) {
    if (on) nvtxRangePushA(name);
  }
  ~NvtxRange() {
    if (on) nvtxRangePop();
  }
  void next(const char* name) {  // close the current range, open the next one
    if (on) {
      nvtxRangePop();
      nvtxRangePushA(name);
    }
  }
};

static thread_local std::string g_err;
static thread_local int64_t g_launches = 0;
void set_error(const std::string& msg) { g_err = msg; }
// pdl_plain_next(): the next launch of this thread goes out without the programmatic attribute (the first kernel after a
// cross-stream event wait: its dependencies are the event's, plain ones)
static thread_local bool g_pdl_plain_next = false;
void pdl_plain_next() { g_pdl_plain_next = true; }
bool pdl_enabled() {
  static const bool on = [] {
    const char* e = getenv("ST2_PDL");
    return !(e && e[0] == '0');
  }();
  if (g_pdl_plain_next) {
    g_pdl_plain_next = false;
    return false;
  }
  return on;
}
int64_t& launch_counter() { return g_launches; }
int launch_tanh_inplace(float* x, int64_t n, cudaStream_t st);  // elementwise.cu

// ------------------------------------------------------------------ layer records
struct ConvW {
  int k = 1, dil = 1, pad = 0, stride = 1;
  int cin = 0;      // contraction channels of the fp32 layout
  int cin_pad = 0;  // channels of the UMMA blob (multiple of 64), 0 = no blob
  int n = 0;        // GEMM columns (Cout, or phases*Cout for a transposed conv)
  int n16 = 0;      // columns of the UMMA blob (n rounded up to a multiple of 32)
  int cstat = 0;    // real channels the columns fold onto for statistics
  double flops_per_row = 0.0;  // algorithmic 2*MAC per output row (per INPUT row for a transposed conv)
  int64_t w32 = -1, w16 = -1, bias = -1, bias16 = -1;  // byte offsets into the device weight arena
  // fp32 parity path on the tensor cores (ST2_PREC_FP32_TC): bf16 blob over 3 * cin_pad input channels, weight rows
  // [w_hi | w_hi | w_lo] against activations [x_hi | x_lo | x_hi]; packed for tiles of nt48 columns (fp32 output)
  int64_t w48 = -1;
  int nt48 = 0;
};
struct NormW {
  int C = 0;
  int h_off = 0;   // row offset of gamma in the concatenated fc output (beta at h_off + C)
  int ss_off = 0;  // pair offset inside one batch row of the (scale, shift) arena
};
struct ResBlock1W {  // AdaINResBlock1, istftnet.py:27-75
  int C = 0, k = 0;
  ConvW c1[3], c2[3];
  NormW n1[3], n2[3];
  int64_t a1[3], a2[3];
  // row-pair form of a 32-channel block (Loader::pair_conv): the [L][32] tensor is read as
  // [L/2][64], so the tcgen05 kernel (64-channel chunks, 128-byte output rows) can run it
  bool paired = false;
  ConvW c1p[3], c2p[3];
  int64_t a1p[3], a2p[3];
};
struct ResBlk1dW {  // AdainResBlk1d, istftnet.py:409-453
  int din = 0, dout = 0, din_pad = 0;
  bool up = false, has_sc = false;
  ConvW conv1, conv2, conv1x1;
  NormW n1, n2;
  int64_t pool_w3 = -1, pool_b = -1;
};
struct UpW {
  ConvW conv;         // polyphase form
  int rate = 0, cout = 0;
  int64_t row0_tap = -1;  // [Cin][Cout] tap (1 + padding) for the reflection-padded row (istftnet only)
};

}  // namespace st2

struct St2Decoder {
  St2Config cfg;
  int device = 0;
  bool loaded = false;
  bool opt_fp32_tc = false;  // st2_decoder_set_option("fp32_tc", 1) before load: pack the 3 x bf16 blobs
  // schedule switches (st2_decoder_set_option any time; defaults from ST2_FORK / ST2_GROUP, else 1):
  int opt_fork = 1;   // excitation branch on the second stream: 0 never, 1 small forwards, 2 every size
  int opt_group = 1;  // independent same-shaped convs of a small forward as one multi-conv launch: 0 never
  int spf = 0;
  std::vector<std::string> key_names;
  std::vector<std::vector<int64_t>> key_shapes;
  // device weights
  uint8_t* dev = nullptr;
  int64_t dev_bytes = 0;
  // plan
  st2::ResBlk1dW encode, decode[4];
  st2::ConvW asr_res;
  int64_t f0w = -1, nw = -1;  // {w0,w1,w2,bias}
  int64_t fcW = -1, fcB = -1;
  int fc_rows = 0;
  int ss_pairs = 0;
  int64_t lin_w = -1;
  float lin_b = 0.f;
  std::vector<st2::UpW> ups;
  std::vector<st2::ConvW> noise_convs;
  std::vector<st2::ConvW> noise_convs16;  // istftnet, bf16 path: grouped (hi, lo) form for tcgen05
  std::vector<int64_t> noise_wsum;        // hifigan: [C] sum over taps of noise_convs[i].weight (fp32)
  std::vector<int> noise_group;           // har rows per wide row of noise_convs16[i] (stride of noise_convs[i])
  std::vector<st2::ResBlock1W> noise_res, resblocks;
  std::vector<int64_t> alphas;  // hifigan generator.alphas
  st2::ConvW conv_post;
  // ST2_KIND_F0N: ProsodyPredictor.F0Ntrain
  st2::ConvW lstm_ih;         // both directions' W_ih as one 1x1 conv (n = 2 * 4H), bias = b_ih + b_hh
  int64_t lstm_whh = -1;      // [2][4H][H] fp32
  st2::ResBlk1dW f0n_blk[2][3];
  st2::ConvW f0n_proj[2];
  st2::ConvW conv_post_p;     // row-pair form of a 32-channel conv_post (hifigan)
  int64_t alpha_post_p = -1;  // alphas[n_ups] repeated twice
  // ST2_KIND_SAMPLER: StyleTransformer1d / Transformer1d denoiser (Modules/diffusion/modules.py:40-426)
  struct SamplerBlk {
    st2::ConvW to_q, to_kv, to_out, ff1, ff2;
  };
  std::vector<SamplerBlk> sblk;
  int64_t s_fcW = -1, s_fcB = -1;  // AdaLayerNorm fc of every block, rows (l, norm|norm_context, gamma|beta, F)
  int64_t s_ln = -1;               // Transformer1d: [layers][4F] (w - 1, b, w_ctx - 1, b_ctx)
  int64_t s_four = -1, s_timeW = -1, s_timeB = -1, s_featW = -1, s_featB = -1;
  int64_t s_mapW[2] = {-1, -1}, s_mapB[2] = {-1, -1}, s_outW = -1, s_outB = -1, s_fixed = -1;
  int64_t last_launches = 0;
  // excitation branch of a small forward (Fwd::run): second stream + fork / join events, created by st2_decoder_load
  cudaStream_t side_st = nullptr;
  cudaEvent_t ev_fork = nullptr, ev_join[ST2_MAX_UPS] = {};
};

namespace st2 {

// ------------------------------------------------------------------- key list
static void add_key(St2Decoder* h, const std::string& n, std::vector<int64_t> s) {
  h->key_names.push_back(n);
  h->key_shapes.push_back(std::move(s));
}
static void keys_wn_conv(St2Decoder* h, const std::string& p, int64_t d0, int64_t d1, int64_t k, bool bias,
                         int64_t bias_n) {
  if (bias) add_key(h, p + ".bias", {bias_n});
  add_key(h, p + ".weight_g", {d0, 1, 1});
  add_key(h, p + ".weight_v", {d0, d1, k});
}
static void keys_adain(St2Decoder* h, const std::string& p, int style, int c) {
  add_key(h, p + ".fc.weight", {2 * c, style});
  add_key(h, p + ".fc.bias", {2 * c});
}
static void keys_resblk1d(St2Decoder* h, const std::string& p, int din, int dout, int style, bool up) {
  keys_wn_conv(h, p + ".conv1", dout, din, 3, true, dout);
  keys_wn_conv(h, p + ".conv2", dout, dout, 3, true, dout);
  keys_adain(h, p + ".norm1", style, din);
  keys_adain(h, p + ".norm2", style, dout);
  if (din != dout) keys_wn_conv(h, p + ".conv1x1", dout, din, 1, false, 0);
  if (up) keys_wn_conv(h, p + ".pool", din, 1, 3, true, din);
}
static void keys_resblock1(St2Decoder* h, const std::string& p, int c, int k, int style) {
  const char* grp[2] = {"convs1", "convs2"};
  for (auto g : grp)
    for (int i = 0; i < 3; ++i) keys_wn_conv(h, p + "." + g + "." + std::to_string(i), c, c, k, true, c);
  const char* ng[2] = {"adain1", "adain2"};
  for (auto g : ng)
    for (int i = 0; i < 3; ++i) keys_adain(h, p + "." + g + "." + std::to_string(i), style, c);
  const char* ag[2] = {"alpha1", "alpha2"};
  for (auto g : ag)
    for (int i = 0; i < 3; ++i) add_key(h, p + "." + g + "." + std::to_string(i), {1, c, 1});
}

static int prod_from(const int* r, int from, int n) {
  int p = 1;
  for (int i = from; i < n; ++i) p *= r[i];
  return p;
}

static void build_keys(St2Decoder* h) {
  const St2Config& c = h->cfg;
  const int style = c.style_dim;
  if (c.kind == ST2_KIND_SAMPLER) {  // state_dict of StyleTransformer1d / Transformer1d (modules.py:40-119, 280-355)
    const int C = c.dim_in, S = c.style_dim, E = c.dim_out, F = C + E;
    const int mid = c.resblock_kernel_sizes[0] * c.resblock_kernel_sizes[1], mult = c.resblock_kernel_sizes[2];
    for (int i = 0; i < c.n_resblock_kernels; ++i) {
      const std::string p = "blocks." + std::to_string(i);
      for (const char* nm : {".attention.norm", ".attention.norm_context"}) {
        if (S) {
          add_key(h, p + nm + ".fc.weight", {2 * F, S});
          add_key(h, p + nm + ".fc.bias", {2 * F});
        } else {
          add_key(h, p + nm + ".weight", {F});
          add_key(h, p + nm + ".bias", {F});
        }
      }
      add_key(h, p + ".attention.to_q.weight", {mid, F});
      add_key(h, p + ".attention.to_kv.weight", {2 * mid, F});
      add_key(h, p + ".attention.attention.to_out.weight", {F, mid});
      add_key(h, p + ".attention.attention.to_out.bias", {F});
      add_key(h, p + ".feed_forward.0.weight", {F * mult, F});
      add_key(h, p + ".feed_forward.0.bias", {F * mult});
      add_key(h, p + ".feed_forward.2.weight", {F, F * mult});
      add_key(h, p + ".feed_forward.2.bias", {F});
    }
    add_key(h, "to_out.1.weight", {C, F, 1});
    add_key(h, "to_out.1.bias", {C});
    for (int j : {0, 2}) {
      add_key(h, "to_mapping." + std::to_string(j) + ".weight", {F, F});
      add_key(h, "to_mapping." + std::to_string(j) + ".bias", {F});
    }
    add_key(h, "to_time.0.0.weights", {C / 2});
    add_key(h, "to_time.0.1.weight", {F, C + 1});
    add_key(h, "to_time.0.1.bias", {F});
    if (S) {
      add_key(h, "to_features.0.weight", {F, S});
      add_key(h, "to_features.0.bias", {F});
    }
    add_key(h, "fixed_embedding.embedding.weight", {c.resblock_kernel_sizes[3], E});
    return;
  }
  if (c.kind == ST2_KIND_F0N) {  // models.py:562-590
    const int d = c.dim_in, hid = d / 2;
    for (int r = 0; r < 2; ++r) {
      const std::string sfx = r ? "_reverse" : "";
      add_key(h, "shared.weight_ih_l0" + sfx, {4 * hid, d + style});
      add_key(h, "shared.weight_hh_l0" + sfx, {4 * hid, hid});
      add_key(h, "shared.bias_ih_l0" + sfx, {4 * hid});
      add_key(h, "shared.bias_hh_l0" + sfx, {4 * hid});
    }
    for (int br = 0; br < 2; ++br) {
      const std::string b = br ? "N" : "F0";
      keys_resblk1d(h, b + ".0", d, d, style, false);
      keys_resblk1d(h, b + ".1", d, hid, style, true);
      keys_resblk1d(h, b + ".2", hid, hid, style, false);
      add_key(h, b + "_proj.weight", {1, hid, 1});
      add_key(h, b + "_proj.bias", {1});
    }
    return;
  }
  for (int i = 0; i < 3; ++i) keys_resblk1d(h, "decode." + std::to_string(i), 1024 + 2 + 64, 1024, style, false);
  keys_resblk1d(h, "decode.3", 1024 + 2 + 64, 512, style, true);
  keys_resblk1d(h, "encode", c.dim_in + 2, 1024, style, false);
  keys_wn_conv(h, "F0_conv", 1, 1, 3, true, 1);
  keys_wn_conv(h, "N_conv", 1, 1, 3, true, 1);
  keys_wn_conv(h, "asr_res.0", 64, 512, 1, true, 64);
  const std::string g = "generator";
  add_key(h, g + ".m_source.l_linear.weight", {1, 9});
  add_key(h, g + ".m_source.l_linear.bias", {1});
  const bool istft = c.kind == ST2_KIND_ISTFTNET;
  const int src_ch = istft ? c.gen_istft_n_fft + 2 : 1;
  const int c0 = c.upsample_initial_channel;
  for (int i = 0; i < c.n_ups; ++i) {
    const int cin = c0 >> i, cout = c0 >> (i + 1);
    keys_wn_conv(h, g + ".ups." + std::to_string(i), cin, cout, c.upsample_kernel_sizes[i], true, cout);
    int kk = 1, rk = 11;
    if (i + 1 < c.n_ups) {
      kk = 2 * prod_from(c.upsample_rates, i + 1, c.n_ups);
      rk = 7;
    }
    add_key(h, g + ".noise_convs." + std::to_string(i) + ".weight", {cout, src_ch, kk});
    add_key(h, g + ".noise_convs." + std::to_string(i) + ".bias", {cout});
    keys_resblock1(h, g + ".noise_res." + std::to_string(i), cout, rk, style);
    for (int j = 0; j < c.n_resblock_kernels; ++j)
      keys_resblock1(h, g + ".resblocks." + std::to_string(i * c.n_resblock_kernels + j), cout,
                     c.resblock_kernel_sizes[j], style);
  }
  const int ch_last = c0 >> c.n_ups;
  if (istft) {
    keys_wn_conv(h, g + ".conv_post", c.gen_istft_n_fft + 2, ch_last, 7, true, c.gen_istft_n_fft + 2);
  } else {
    keys_wn_conv(h, g + ".conv_post", 1, ch_last, 7, true, 1);
    add_key(h, g + ".alphas.0", {1, c0, 1});
    for (int i = 0; i < c.n_ups; ++i) add_key(h, g + ".alphas." + std::to_string(i + 1), {1, c0 >> (i + 1), 1});
  }
}

// ------------------------------------------------------------------- host blob
struct HostBlob {
  std::vector<uint8_t> data;
  int64_t add(const void* src, int64_t bytes) {
    const int64_t off = round_up((int64_t)data.size(), 256);
    data.resize(off + bytes);
    if (src) memcpy(data.data() + off, src, bytes);
    return off;
  }
};

struct Loader {
  std::map<std::string, const St2Tensor*> by_name;
  HostBlob blob;
  std::vector<float> fcW, fcB;  // concatenated AdaIN fc layers
  int style = 0;
  int ss_pairs = 0;
  bool fp32_tc = false;  // also pack the 3 x bf16 blobs of the ST2_PREC_FP32_TC path
  std::string err;

  const St2Tensor* get(const std::string& n) {
    auto it = by_name.find(n);
    if (it == by_name.end()) {
      if (err.empty()) err = "missing key " + n;
      return nullptr;
    }
    return it->second;
  }
  // old-style weight_norm, dim 0: w = g * v / ||v|| over all dims but 0 (istftnet.py:5)
  bool folded(const std::string& p, std::vector<float>& w, int64_t& d0, int64_t& d1, int64_t& d2) {
    const St2Tensor* v = get(p + ".weight_v");
    const St2Tensor* g = get(p + ".weight_g");
    if (!v || !g) return false;
    d0 = v->shape[0]; d1 = v->shape[1]; d2 = v->shape[2];
    const int64_t inner = d1 * d2;
    w.resize(d0 * inner);
    for (int64_t o = 0; o < d0; ++o) {
      double nrm = 0.0;
      for (int64_t i = 0; i < inner; ++i) nrm += (double)v->data[o * inner + i] * v->data[o * inner + i];
      const float scale = (float)((double)g->data[o] / sqrt(nrm));
      for (int64_t i = 0; i < inner; ++i) w[o * inner + i] = v->data[o * inner + i] * scale;
    }
    return true;
  }
  void finish_conv(ConvW& c, const std::vector<float>& w_kcn, const std::vector<float>* bias, bool want_umma) {
    c.w32 = blob.add(w_kcn.data(), (int64_t)w_kcn.size() * 4);
    if (bias) c.bias = blob.add(bias->data(), (int64_t)bias->size() * 4);
    if (want_umma && c.stride == 1 && (c.k - 1) * c.dil <= 50) {
      // GEMM columns are padded with zero weights / zero bias up to a multiple of 32 (conv_post: 22 -> 32)
      c.cin_pad = (int)round_up(c.cin, 64);
      c.n16 = (int)round_up(c.n, 32);
      const std::vector<float>* wsrc = &w_kcn;
      std::vector<float> wpad, bpad;
      if (c.n16 != c.n) {
        wpad.assign((size_t)c.k * c.cin * c.n16, 0.f);
        for (int64_t r = 0; r < (int64_t)c.k * c.cin; ++r)
          memcpy(&wpad[(size_t)r * c.n16], &w_kcn[(size_t)r * c.n], (size_t)c.n * 4);
        wsrc = &wpad;
        if (bias) {
          bpad.assign((size_t)c.n16, 0.f);
          memcpy(bpad.data(), bias->data(), (size_t)c.n * 4);
          c.bias16 = blob.add(bpad.data(), (int64_t)bpad.size() * 4);
        }
      } else {
        c.bias16 = c.bias;
      }
      const int64_t elems = st2_umma_packed_elems(c.k, c.cin_pad, c.n16);
      c.w16 = blob.add(nullptr, elems * 2);
      st2_umma_pack_weights(wsrc->data(), c.k, c.cin, c.cin_pad, c.n16,
                            reinterpret_cast<uint16_t*>(blob.data.data() + c.w16));
      if (fp32_tc) {
        // [k][3 * cin_pad][n16] fp32 with rows [w_hi | w_hi | w_lo] (all exactly representable in bf16)
        const int cp = c.cin_pad, n16 = c.n16;
        std::vector<float> w3((size_t)c.k * 3 * cp * n16, 0.f);
        for (int j = 0; j < c.k; ++j)
          for (int ch = 0; ch < c.cin; ++ch)
            for (int nn = 0; nn < n16; ++nn) {
              const float w = (*wsrc)[((size_t)j * c.cin + ch) * n16 + nn];
              uint32_t u;
              memcpy(&u, &w, 4);
              uint32_t r = u + 0x7fffu + ((u >> 16) & 1u);  // round to nearest even bf16 (finite weights)
              r &= 0xffff0000u;
              float hi;
              memcpy(&hi, &r, 4);
              const float lo = w - hi;
              const size_t base = (size_t)j * 3 * cp;
              w3[(base + ch) * n16 + nn] = hi;
              w3[(base + cp + ch) * n16 + nn] = hi;
              w3[(base + 2 * cp + ch) * n16 + nn] = lo;
            }
        c.nt48 = n16 % 64 == 0 ? 64 : 32;
        c.w48 = blob.add(nullptr, st2_umma_packed_elems(c.k, 3 * cp, n16) * 2);
        st2_umma_pack_weights_nt(w3.data(), c.k, 3 * cp, 3 * cp, n16, c.nt48,
                                 reinterpret_cast<uint16_t*>(blob.data.data() + c.w48));
      }
    }
  }
  // Conv1d: torch [Cout][Cin][k] -> [k][Cin][Cout]
  // Row-pair form of a stride-1 conv: with t = 2q + e, y[2q+e][o] = sum_j W[j][c][o] x[2q + e + j*dil - pad][c]
  // and 2q + e + j*dil - pad = 2(q + m) + h, so on rows of (h, c) / (e, o) pairs it is the dense conv
  //   Y[q][e*n + o] = sum_m W2[m][h*cin + c][e*n + o] X[q + m][h*cin + c],  m = floor((e + j*dil - pad) / 2)
  // with dilation 1 (most 32x32 blocks of W2 are zero: 2k of them are not).
  void pair_conv(ConvW& c2, const ConvW& c, const std::vector<float>& kcn, const std::vector<float>* bias) {
    auto fl2 = [](int v) { return v >= 0 ? v / 2 : -((-v + 1) / 2); };
    const int m_min = fl2(-c.pad), m_max = fl2(1 + (c.k - 1) * c.dil - c.pad);
    c2 = ConvW();
    c2.k = m_max - m_min + 1; c2.dil = 1; c2.pad = -m_min; c2.stride = 1;
    c2.cin = 2 * c.cin; c2.n = 2 * c.n; c2.cstat = c.n;
    c2.flops_per_row = 2.0 * c.flops_per_row;  // two original rows per pair row
    std::vector<float> w2((size_t)c2.k * c2.cin * c2.n, 0.f);
    for (int e = 0; e < 2; ++e)
      for (int j = 0; j < c.k; ++j) {
        const int s = e + j * c.dil - c.pad, m = fl2(s), hh = s - 2 * m;
        for (int ci = 0; ci < c.cin; ++ci)
          for (int o = 0; o < c.n; ++o)
            w2[((size_t)(m - m_min) * c2.cin + hh * c.cin + ci) * c2.n + e * c.n + o] =
                kcn[((size_t)j * c.cin + ci) * c.n + o];
      }
    std::vector<float> b2;
    if (bias) {
      b2.resize((size_t)c2.n);
      for (int e = 0; e < 2; ++e)
        for (int o = 0; o < c.n; ++o) b2[e * c.n + o] = (*bias)[o];
    }
    finish_conv(c2, w2, bias ? &b2 : nullptr, true);
  }
  bool vec2(int64_t& off, const std::string& name, int64_t n) {  // parameter vector repeated twice
    const St2Tensor* t = get(name);
    if (!t) return false;
    std::vector<float> v((size_t)2 * n);
    for (int64_t i = 0; i < n; ++i) v[i] = v[n + i] = t->data[i];
    off = blob.add(v.data(), 2 * n * 4);
    return true;
  }
  bool conv1d(ConvW& c, const std::string& p, bool weight_norm, bool has_bias, int k, int dil, int pad,
              int stride, bool want_umma, ConvW* pair = nullptr) {
    std::vector<float> w;
    int64_t d0, d1, d2;
    if (weight_norm) {
      if (!folded(p, w, d0, d1, d2)) return false;
    } else {
      const St2Tensor* t = get(p + ".weight");
      if (!t) return false;
      d0 = t->shape[0]; d1 = t->shape[1]; d2 = t->shape[2];
      w.assign(t->data, t->data + d0 * d1 * d2);
    }
    c.k = k; c.dil = dil; c.pad = pad; c.stride = stride;
    c.cin = (int)d1; c.n = (int)d0; c.cstat = (int)d0;
    c.flops_per_row = 2.0 * (double)d0 * (double)d1 * (double)d2;
    std::vector<float> kcn((size_t)d2 * d1 * d0);
    for (int64_t co = 0; co < d0; ++co)
      for (int64_t ci = 0; ci < d1; ++ci)
        for (int64_t j = 0; j < d2; ++j) kcn[(j * d1 + ci) * d0 + co] = w[(co * d1 + ci) * d2 + j];
    std::vector<float> b;
    if (has_bias) {
      const St2Tensor* bt = get(p + ".bias");
      if (!bt) return false;
      b.assign(bt->data, bt->data + d0);
    }
    finish_conv(c, kcn, has_bias ? &b : nullptr, want_umma);
    if (pair) pair_conv(*pair, c, kcn, has_bias ? &b : nullptr);
    return true;
  }
  // ConvTranspose1d [Cin][Cout][K] (stride s, padding p) -> polyphase conv over input offsets
  // o in [omin, omax]: y[s*q + r] += x[q + o] * w[.., r + p - s*o]  (0 <= r + p - s*o < K)
  bool conv_transpose(UpW& u, const std::string& p, int K, int s, int pd) {
    std::vector<float> w;
    int64_t d0, d1, d2;
    if (!folded(p, w, d0, d1, d2)) return false;
    const St2Tensor* bt = get(p + ".bias");
    if (!bt) return false;
    const int cin = (int)d0, cout = (int)d1;
    const int omax = (s - 1 + pd) / s, omin = -((K - 1 - pd) / s);
    ConvW& c = u.conv;
    c.k = omax - omin + 1; c.dil = 1; c.pad = -omin; c.stride = 1;
    c.cin = cin; c.n = s * cout; c.cstat = cout;
    c.flops_per_row = 2.0 * (double)cin * (double)cout * (double)K;  // 2*Cin*Lin*Cout*K in total
    std::vector<float> kcn((size_t)c.k * cin * c.n, 0.f);
    for (int j = 0; j < c.k; ++j) {
      const int o = j + omin;
      for (int r = 0; r < s; ++r) {
        const int kk = r + pd - s * o;
        if (kk < 0 || kk >= K) continue;
        for (int ci = 0; ci < cin; ++ci)
          for (int co = 0; co < cout; ++co)
            kcn[((size_t)j * cin + ci) * c.n + r * cout + co] = w[((size_t)ci * cout + co) * K + kk];
      }
    }
    std::vector<float> b((size_t)c.n);
    for (int r = 0; r < s; ++r)
      for (int co = 0; co < cout; ++co) b[r * cout + co] = bt->data[co];
    finish_conv(c, kcn, &b, true);
    u.rate = s; u.cout = cout;
    // tap seen by output t = 1 from input row 0 (reflection-padded row of the last iSTFTNet stage)
    std::vector<float> tap((size_t)cin * cout);
    const int kk1 = 1 + pd;
    for (int ci = 0; ci < cin; ++ci)
      for (int co = 0; co < cout; ++co) tap[(size_t)ci * cout + co] = kk1 < K ? w[((size_t)ci * cout + co) * K + kk1] : 0.f;
    u.row0_tap = blob.add(tap.data(), (int64_t)tap.size() * 4);
    return true;
  }
  // Strided noise conv Conv1d(src_ch, C, k = 2*sf, stride sf, padding (sf+1)/2) over the bf16 (hi, lo)
  // rows of `har` (64 columns per row, source.cu stft_hilo_kernel), regrouped so that the tensor
  // core can run it with stride 1: sf consecutive har rows form one wide row of sf*64 channels and
  //   y[q] = sum_{j<3} sum_{r<sf} W'[j][r*64 + c] har[sf*(q+j-1) + r][c],  tap = sf*(j-1) + r + pad
  // (zero weight when the tap falls outside [0, k)); the lo half repeats the hi weights.
  // sf == 1 is the plain 1x1 conv of the last stage.
  bool noise_conv16(ConvW& c, const std::string& p, int sf, int src_ch) {
    const St2Tensor* t = get(p + ".weight");
    const St2Tensor* bt = get(p + ".bias");
    if (!t || !bt) return false;
    const int cout = (int)t->shape[0], K = (int)t->shape[2];
    if ((int)t->shape[1] != src_ch || src_ch > 32) return false;
    const int pd = sf > 1 ? (sf + 1) / 2 : 0;
    c.k = sf > 1 ? 3 : 1; c.dil = 1; c.pad = sf > 1 ? 1 : 0; c.stride = 1;
    c.cin = sf * 64; c.n = cout; c.cstat = cout;
    c.flops_per_row = 2.0 * (double)cout * (double)src_ch * (double)K;
    std::vector<float> kcn((size_t)c.k * c.cin * cout, 0.f);
    for (int j = 0; j < c.k; ++j)
      for (int r = 0; r < sf; ++r) {
        const int tap = sf > 1 ? sf * (j - 1) + r + pd : 0;
        if (tap < 0 || tap >= K) continue;
        for (int ch = 0; ch < src_ch; ++ch)
          for (int co = 0; co < cout; ++co) {
            const float w = t->data[((size_t)co * src_ch + ch) * K + tap];
            kcn[((size_t)j * c.cin + r * 64 + ch) * cout + co] = w;
            kcn[((size_t)j * c.cin + r * 64 + 32 + ch) * cout + co] = w;
          }
      }
    std::vector<float> b(bt->data, bt->data + cout);
    // only the UMMA blob is kept (w32 stays -1: this record is never run on the fp32 path)
    c.bias = blob.add(b.data(), (int64_t)b.size() * 4);
    c.bias16 = c.bias;
    c.cin_pad = c.cin; c.n16 = cout;
    if (umma_n_tile(cout) == 0) return false;
    const int64_t elems = st2_umma_packed_elems(c.k, c.cin_pad, c.n16);
    c.w16 = blob.add(nullptr, elems * 2);
    st2_umma_pack_weights(kcn.data(), c.k, c.cin, c.cin_pad, c.n16,
                          reinterpret_cast<uint16_t*>(blob.data.data() + c.w16));
    return true;
  }
  // nn.Linear weight [out][in] as a 1x1 conv (GEMM columns = out)
  bool linear(ConvW& c, const std::string& p, bool has_bias) {
    const St2Tensor* t = get(p + ".weight");
    const St2Tensor* bt = has_bias ? get(p + ".bias") : nullptr;
    if (!t || (has_bias && !bt)) return false;
    const int64_t out = t->shape[0], in = t->shape[1];
    c.k = 1; c.dil = 1; c.pad = 0; c.stride = 1;
    c.cin = (int)in; c.n = (int)out; c.cstat = (int)out;
    c.flops_per_row = 2.0 * (double)in * (double)out;
    std::vector<float> kcn((size_t)in * out);
    for (int64_t o = 0; o < out; ++o)
      for (int64_t i = 0; i < in; ++i) kcn[(size_t)i * out + o] = t->data[(size_t)o * in + i];
    std::vector<float> b;
    if (has_bias) b.assign(bt->data, bt->data + out);
    finish_conv(c, kcn, has_bias ? &b : nullptr, true);
    return true;
  }
  bool norm(NormW& n, const std::string& p, int C) {
    const St2Tensor* w = get(p + ".fc.weight");
    const St2Tensor* b = get(p + ".fc.bias");
    if (!w || !b) return false;
    n.C = C;
    n.h_off = (int)fcB.size();
    fcW.insert(fcW.end(), w->data, w->data + (int64_t)2 * C * style);
    fcB.insert(fcB.end(), b->data, b->data + 2 * C);
    n.ss_off = ss_pairs;
    ss_pairs += (int)round_up(C, 64);
    return true;
  }
  bool vec(int64_t& off, const std::string& name, int64_t n) {
    const St2Tensor* t = get(name);
    if (!t) return false;
    off = blob.add(t->data, n * 4);
    return true;
  }
  bool resblk1d(ResBlk1dW& r, const std::string& p, int din, int dout, bool up) {
    r.din = din; r.dout = dout; r.up = up; r.has_sc = din != dout;
    r.din_pad = (int)round_up(din, 64);
    bool ok = conv1d(r.conv1, p + ".conv1", true, true, 3, 1, 1, 1, true) &&
              conv1d(r.conv2, p + ".conv2", true, true, 3, 1, 1, 1, true) && norm(r.n1, p + ".norm1", din) &&
              norm(r.n2, p + ".norm2", dout);
    if (ok && r.has_sc) ok = conv1d(r.conv1x1, p + ".conv1x1", true, false, 1, 1, 0, 1, true);
    if (ok && up) {
      std::vector<float> w;
      int64_t d0, d1, d2;
      ok = folded(p + ".pool", w, d0, d1, d2);  // [din][1][3], g per channel
      if (ok) {
        r.pool_w3 = blob.add(w.data(), (int64_t)w.size() * 4);
        ok = vec(r.pool_b, p + ".pool.bias", din);
      }
    }
    return ok;
  }
  bool resblock1(ResBlock1W& r, const std::string& p, int C, int k, const int* dils) {
    r.C = C; r.k = k;
    r.paired = C == 32;
    bool ok = true;
    for (int i = 0; i < 3 && ok; ++i) {
      const std::string si = std::to_string(i);
      ok = conv1d(r.c1[i], p + ".convs1." + si, true, true, k, dils[i], dils[i] * (k - 1) / 2, 1, true,
                  r.paired ? &r.c1p[i] : nullptr) &&
           conv1d(r.c2[i], p + ".convs2." + si, true, true, k, 1, (k - 1) / 2, 1, true,
                  r.paired ? &r.c2p[i] : nullptr) &&
           norm(r.n1[i], p + ".adain1." + si, C) && norm(r.n2[i], p + ".adain2." + si, C) &&
           vec(r.a1[i], p + ".alpha1." + si, C) && vec(r.a2[i], p + ".alpha2." + si, C);
      if (ok && r.paired) {
        ok = vec2(r.a1p[i], p + ".alpha1." + si, C) && vec2(r.a2p[i], p + ".alpha2." + si, C);
        if (r.c1p[i].w16 < 0 || r.c2p[i].w16 < 0) r.paired = false;  // halo too long for the tile
      }
    }
    return ok;
  }
};

static int finish_load(St2Decoder* h, Loader& L) {
  h->fc_rows = (int)L.fcB.size();
  h->ss_pairs = L.ss_pairs;
  h->fcW = L.blob.add(L.fcW.data(), (int64_t)L.fcW.size() * 4);
  h->fcB = L.blob.add(L.fcB.data(), (int64_t)L.fcB.size() * 4);

  ST2_CUDA_OK(cudaSetDevice(h->device));
  if (h->dev) ST2_CUDA_OK(cudaFree(h->dev));
  h->dev = nullptr;
  h->dev_bytes = (int64_t)L.blob.data.size();
  ST2_CUDA_OK(cudaMalloc(&h->dev, h->dev_bytes));
  ST2_CUDA_OK(cudaMemcpy(h->dev, L.blob.data.data(), h->dev_bytes, cudaMemcpyHostToDevice));
  h->loaded = true;
  return ST2_OK;
}

static int do_load(St2Decoder* h, const St2Tensor* tensors, int n) {
  const St2Config& c = h->cfg;
  Loader L;
  L.style = c.style_dim;
  L.fp32_tc = h->opt_fp32_tc && (c.kind == ST2_KIND_ISTFTNET || c.kind == ST2_KIND_HIFIGAN || c.kind == ST2_KIND_F0N);
  std::vector<std::string> names(n);
  for (int i = 0; i < n; ++i) {
    ST2_REQUIRE(tensors[i].name && tensors[i].data, "tensor without name/data");
    std::string nm = tensors[i].name;
    if (nm.rfind("module.", 0) == 0) nm = nm.substr(7);  // DDP prefix, models.py:1126-1142
    L.by_name[nm] = &tensors[i];
  }
  // strict key / shape check (load_state_dict semantics)
  for (size_t i = 0; i < h->key_names.size(); ++i) {
    auto it = L.by_name.find(h->key_names[i]);
    ST2_REQUIRE(it != L.by_name.end(), "missing key " + h->key_names[i]);
    const St2Tensor* t = it->second;
    bool same = t->ndim == (int)h->key_shapes[i].size();
    for (int d = 0; same && d < t->ndim; ++d) same = t->shape[d] == h->key_shapes[i][d];
    ST2_REQUIRE(same, "shape mismatch for " + h->key_names[i]);
  }
  ST2_REQUIRE(L.by_name.size() == h->key_names.size(), "unexpected keys in state dict");

  bool ok = true;
  if (c.kind == ST2_KIND_SAMPLER) {
    const int C = c.dim_in, S = c.style_dim, E = c.dim_out, F = C + E, layers = c.n_resblock_kernels;
    std::vector<float> fcW, fcB, ln;
    h->sblk.assign(layers, St2Decoder::SamplerBlk());
    for (int i = 0; i < layers && ok; ++i) {
      const std::string p = "blocks." + std::to_string(i);
      for (const char* nm : {".attention.norm", ".attention.norm_context"}) {
        if (S) {
          const St2Tensor* w = L.get(p + nm + ".fc.weight");
          const St2Tensor* b = L.get(p + nm + ".fc.bias");
          ok = ok && w && b;
          if (!ok) break;
          fcW.insert(fcW.end(), w->data, w->data + (int64_t)2 * F * S);
          fcB.insert(fcB.end(), b->data, b->data + 2 * F);
        } else {
          const St2Tensor* w = L.get(p + nm + ".weight");
          const St2Tensor* b = L.get(p + nm + ".bias");
          ok = ok && w && b;
          if (!ok) break;
          for (int j = 0; j < F; ++j) ln.push_back(w->data[j] - 1.f);  // adaln2 applies (1 + g) * xhat + b
          ln.insert(ln.end(), b->data, b->data + F);
        }
      }
      St2Decoder::SamplerBlk& k = h->sblk[i];
      ok = ok && L.linear(k.to_q, p + ".attention.to_q", false) && L.linear(k.to_kv, p + ".attention.to_kv", false) &&
           L.linear(k.to_out, p + ".attention.attention.to_out", true) && L.linear(k.ff1, p + ".feed_forward.0", true) &&
           L.linear(k.ff2, p + ".feed_forward.2", true);
    }
    if (ok && S) {
      h->s_fcW = L.blob.add(fcW.data(), (int64_t)fcW.size() * 4);
      h->s_fcB = L.blob.add(fcB.data(), (int64_t)fcB.size() * 4);
      ok = L.vec(h->s_featW, "to_features.0.weight", (int64_t)F * S) && L.vec(h->s_featB, "to_features.0.bias", F);
    } else if (ok) {
      h->s_ln = L.blob.add(ln.data(), (int64_t)ln.size() * 4);
    }
    ok = ok && L.vec(h->s_four, "to_time.0.0.weights", C / 2) && L.vec(h->s_timeW, "to_time.0.1.weight", (int64_t)F * (C + 1)) &&
         L.vec(h->s_timeB, "to_time.0.1.bias", F) && L.vec(h->s_mapW[0], "to_mapping.0.weight", (int64_t)F * F) &&
         L.vec(h->s_mapB[0], "to_mapping.0.bias", F) && L.vec(h->s_mapW[1], "to_mapping.2.weight", (int64_t)F * F) &&
         L.vec(h->s_mapB[1], "to_mapping.2.bias", F) && L.vec(h->s_outW, "to_out.1.weight", (int64_t)C * F) &&
         L.vec(h->s_outB, "to_out.1.bias", C) &&
         L.vec(h->s_fixed, "fixed_embedding.embedding.weight", (int64_t)c.resblock_kernel_sizes[3] * E);
    ST2_REQUIRE(ok, L.err);
    return finish_load(h, L);
  }
  if (c.kind == ST2_KIND_F0N) {
    const int d = c.dim_in, hid = d / 2, in = d + c.style_dim, g4 = 4 * hid;
    // both directions' input projections as one GEMM: column dir * 4H + gate * H + unit
    std::vector<float> kcn((size_t)in * 2 * g4), bsum((size_t)2 * g4), whh((size_t)2 * g4 * hid);
    for (int r = 0; r < 2 && ok; ++r) {
      const std::string sfx = r ? "_reverse" : "";
      const St2Tensor* wih = L.get("shared.weight_ih_l0" + sfx);
      const St2Tensor* whp = L.get("shared.weight_hh_l0" + sfx);
      const St2Tensor* bi = L.get("shared.bias_ih_l0" + sfx);
      const St2Tensor* bh = L.get("shared.bias_hh_l0" + sfx);
      ok = wih && whp && bi && bh;
      if (!ok) break;
      for (int row = 0; row < g4; ++row) {
        for (int ci = 0; ci < in; ++ci) kcn[(size_t)ci * 2 * g4 + r * g4 + row] = wih->data[(size_t)row * in + ci];
        bsum[(size_t)r * g4 + row] = bi->data[row] + bh->data[row];
      }
      memcpy(&whh[(size_t)r * g4 * hid], whp->data, (size_t)g4 * hid * 4);
    }
    if (ok) {
      ConvW& cw = h->lstm_ih;
      cw.k = 1; cw.dil = 1; cw.pad = 0; cw.stride = 1;
      cw.cin = in; cw.n = 2 * g4; cw.cstat = 2 * g4;
      cw.flops_per_row = 2.0 * (double)in * (double)(2 * g4);
      L.finish_conv(cw, kcn, &bsum, true);
      h->lstm_whh = L.blob.add(whh.data(), (int64_t)whh.size() * 4);
    }
    for (int br = 0; br < 2 && ok; ++br) {
      const std::string b = br ? "N" : "F0";
      ok = L.resblk1d(h->f0n_blk[br][0], b + ".0", d, d, false) && L.resblk1d(h->f0n_blk[br][1], b + ".1", d, hid, true) &&
           L.resblk1d(h->f0n_blk[br][2], b + ".2", hid, hid, false) &&
           L.conv1d(h->f0n_proj[br], b + "_proj", false, true, 1, 1, 0, 1, false);
    }
    ST2_REQUIRE(ok, L.err);
    return finish_load(h, L);
  }
  for (int i = 0; i < 3 && ok; ++i) ok = L.resblk1d(h->decode[i], "decode." + std::to_string(i), 1090, 1024, false);
  ok = ok && L.resblk1d(h->decode[3], "decode.3", 1090, 512, true);
  ok = ok && L.resblk1d(h->encode, "encode", c.dim_in + 2, 1024, false);
  ok = ok && L.conv1d(h->asr_res, "asr_res.0", true, true, 1, 1, 0, 1, true);
  for (int which = 0; which < 2 && ok; ++which) {
    const std::string p = which == 0 ? "F0_conv" : "N_conv";
    std::vector<float> w;
    int64_t d0, d1, d2;
    ok = L.folded(p, w, d0, d1, d2);
    const St2Tensor* b = ok ? L.get(p + ".bias") : nullptr;
    ok = ok && b;
    if (ok) {
      float four[4] = {w[0], w[1], w[2], b->data[0]};
      (which == 0 ? h->f0w : h->nw) = L.blob.add(four, sizeof(four));
    }
  }
  const std::string g = "generator";
  if (ok) {
    ok = L.vec(h->lin_w, g + ".m_source.l_linear.weight", 9);
    const St2Tensor* lb = L.get(g + ".m_source.l_linear.bias");
    ok = ok && lb;
    if (ok) h->lin_b = lb->data[0];
  }
  const bool istft = c.kind == ST2_KIND_ISTFTNET;
  const int c0 = c.upsample_initial_channel;
  h->ups.assign(c.n_ups, UpW());
  h->noise_convs.assign(c.n_ups, ConvW());
  h->noise_convs16.assign(c.n_ups, ConvW());
  h->noise_group.assign(c.n_ups, 1);
  h->noise_wsum.assign(c.n_ups, -1);
  h->noise_res.assign(c.n_ups, ResBlock1W());
  h->resblocks.assign(c.n_ups * c.n_resblock_kernels, ResBlock1W());
  const int d135[3] = {1, 3, 5};
  for (int i = 0; i < c.n_ups && ok; ++i) {
    const std::string si = std::to_string(i);
    const int cout = c0 >> (i + 1);
    const int u = c.upsample_rates[i], K = c.upsample_kernel_sizes[i];
    const int pd = istft ? (K - u) / 2 : (u / 2 + u % 2);  // istftnet.py:321, hifigan.py:293
    ok = L.conv_transpose(h->ups[i], g + ".ups." + si, K, u, pd);
    int rk = 11;
    if (ok && i + 1 < c.n_ups) {
      const int sf = prod_from(c.upsample_rates, i + 1, c.n_ups);
      ok = L.conv1d(h->noise_convs[i], g + ".noise_convs." + si, false, true, 2 * sf, 1, (sf + 1) / 2, sf, false);
      rk = 7;
    } else if (ok) {
      ok = L.conv1d(h->noise_convs[i], g + ".noise_convs." + si, false, true, 1, 1, 0, 1, false);
    }
    if (ok && !istft) {  // sum of each output channel's taps (mean-free storage of x_source)
      const St2Tensor* t = L.get(g + ".noise_convs." + si + ".weight");
      ok = t != nullptr;
      if (ok) {
        const int64_t inner = t->shape[1] * t->shape[2];
        std::vector<float> ws((size_t)cout);
        for (int co = 0; co < cout; ++co) {
          double a = 0.0;
          for (int64_t q = 0; q < inner; ++q) a += (double)t->data[co * inner + q];
          ws[co] = (float)a;
        }
        h->noise_wsum[i] = L.blob.add(ws.data(), (int64_t)ws.size() * 4);
      }
    }
    if (ok && istft) {
      const int sf = i + 1 < c.n_ups ? prod_from(c.upsample_rates, i + 1, c.n_ups) : 1;
      h->noise_group[i] = sf;
      // best effort: a layer the grouped form cannot express keeps the fp32-input FFMA kernel
      if (!L.noise_conv16(h->noise_convs16[i], g + ".noise_convs." + si, sf, c.gen_istft_n_fft + 2))
        h->noise_convs16[i] = ConvW();
    }
    ok = ok && L.resblock1(h->noise_res[i], g + ".noise_res." + si, cout, rk, d135);
    for (int j = 0; j < c.n_resblock_kernels && ok; ++j)
      ok = L.resblock1(h->resblocks[i * c.n_resblock_kernels + j],
                       g + ".resblocks." + std::to_string(i * c.n_resblock_kernels + j), cout,
                       c.resblock_kernel_sizes[j], c.resblock_dilations[j]);
  }
  const bool post_pair = !istft && (c0 >> c.n_ups) == 32;
  ok = ok && L.conv1d(h->conv_post, g + ".conv_post", true, true, 7, 1, 3, 1, istft,
                      post_pair ? &h->conv_post_p : nullptr);
  if (ok && post_pair) ok = L.vec2(h->alpha_post_p, g + ".alphas." + std::to_string(c.n_ups), 32);
  if (ok && !istft) {
    h->alphas.assign(c.n_ups + 1, -1);
    for (int i = 0; i <= c.n_ups && ok; ++i)
      ok = L.vec(h->alphas[i], g + ".alphas." + std::to_string(i), i == 0 ? c0 : c0 >> i);
  }
  ST2_REQUIRE(ok, L.err);
  return finish_load(h, L);
}

}  // namespace st2

// =============================================================================== forward
namespace st2 {

struct Arena {
  uint8_t* base = nullptr;
  int64_t cap = 0, off = 0, peak = 0;
  bool dry = false;
  template <typename T>
  T* alloc(int64_t count) {
    off = round_up(off, 256);
    T* p = reinterpret_cast<T*>((dry ? reinterpret_cast<uint8_t*>((uintptr_t)4096) : base) + off);
    off += count * (int64_t)sizeof(T);
    if (off > peak) peak = off;
    return p;
  }
  int64_t mark() const { return off; }
  void release(int64_t m) { off = m; }
};

template <typename AT>
struct DT;
template <>
struct DT<float> { static constexpr int id = ST2_DT_F32; };
template <>
struct DT<__nv_bfloat16> { static constexpr int id = ST2_DT_BF16; };

// channels-last activation view
template <typename AT>
struct Act {
  AT* p = nullptr;
  int64_t bstride = 0;
  int ld = 0;
  int L = 0;
};

template <typename AT>
struct Fwd {
  St2Decoder* h;
  Arena ar;
  cudaStream_t st;
  int B, T;
  bool dry;
  bool tc32 = false;            // ST2_PREC_FP32_TC: fp32 activations, dense convs as 3 x bf16 on tcgen05
  // small forwards run the excitation branch (source -> STFT -> noise_convs -> noise_res, istftnet.py:350-356 and
  // :366-368) on the handle's second stream, beside encode / decode; it allocates from its own arena (`side`, the
  // first side_bytes of the workspace) so that the two branches never alias.  to_side() / to_main() swap (ar, st).
  Arena side;
  cudaStream_t st_side = nullptr;
  bool on_side = false;
  void to_side() { if (!on_side) { std::swap(ar, side); std::swap(st, st_side); on_side = true; } }
  void to_main() { if (on_side) { std::swap(ar, side); std::swap(st, st_side); on_side = false; } }
  const float* hall = nullptr;  // [B][fc_rows]
  float* ss = nullptr;          // [B][ss_pairs][2]

  template <typename U>
  U* W(int64_t off) const { return off < 0 ? nullptr : reinterpret_cast<U*>(h->dev + off); }

#define RUN(x)                 \
  do {                         \
    if (!dry) {                \
      int _rc = (x);           \
      if (_rc) return _rc;     \
    }                          \
  } while (0)

  const float* ss_of(const NormW& n) const { return ss + (int64_t)n.ss_off * 2; }

  // finalize InstanceNorm statistics of channels [c_off, c_off+C) of norm layer n
  int finalize(const float* partials, int n_tiles, int n_cols, int C, int64_t count, const NormW& n,
               int c_off, int reps = 1) {
    RUN(launch_stats_finalize(partials, n_tiles, n_cols, C, count, hall + n.h_off, h->fc_rows, n.C,
                              ss + (int64_t)n.ss_off * 2, h->ss_pairs, c_off, B, st, reps));
    return ST2_OK;
  }

  // the same for n (<= 4) norm layers with equal shapes, one launch
  int finalize_multi(float* const* partials, int n, int n_tiles, int n_cols, int C, int64_t count,
                     const NormW* const* norms, int reps, int c_off = 0) {
    const float* pp[4]; const float* hg[4]; float* so[4];
    for (int j = 0; j < n; ++j) {
      if (norms[j]->C != norms[0]->C) {  // different layers after all: one by one
        for (int q = 0; q < n; ++q) {
          int rc = finalize(partials[q], n_tiles, n_cols, C, count, *norms[q], c_off, reps);
          if (rc) return rc;
        }
        return ST2_OK;
      }
      pp[j] = partials[j]; hg[j] = hall + norms[j]->h_off; so[j] = ss + (int64_t)norms[j]->ss_off * 2;
    }
    RUN(launch_stats_finalize_multi(pp, n, n_tiles, n_cols, C, count, hg, h->fc_rows, norms[0]->C, so, h->ss_pairs, c_off,
                                    B, st, reps));
    return ST2_OK;
  }

  struct ConvIO {
    const void* x; int x_dtype; int64_t x_bstride; int ldx; int Lin;
    void* y; int y_dtype; int64_t y_bstride; int ldy; int Lout;
    const void* res = nullptr; int64_t res_bstride = 0; int ldres = 0; int res_shift = 0;
    const NormW* pre_norm = nullptr; int act = ST2_ACT_NONE; float slope = 0.f; const float* alpha = nullptr;
    float out_scale = 1.f; int accumulate = 0;
    bool want_stats = false;
    int extra_stat_tiles = 0;
    const float* bias_b = nullptr;  // per-batch bias [B][n] replacing the layer's own
  };
  // stats result of the last conv() call
  float* last_partials = nullptr;
  int last_tiles = 0, last_cols = 0;

  // ST2_PREC_FP32_TC: the same fused conv with fp32 activations on the tensor cores.  (1) split3: the layer's prologue
  // in fp32 + (hi, lo, hi) bf16 split of every input value; (2) conv_umma without prologue over 3 * cin_pad channels
  // against the [w_hi | w_hi | w_lo] blob, fp32 accumulation, fp32 output (bias included); (3) post32 when a residual,
  // scale or running mean follows; (4) colstats for the next InstanceNorm.  Same St2ConvParams semantics as conv().
  int conv_tc32(const ConvW& c, const ConvIO& io) {
    float* stats = nullptr;
    int n_tiles = 0;
    if (io.want_stats) {
      n_tiles = colstats_tiles(io.Lout) + io.extra_stat_tiles;
      stats = ar.alloc<float>((int64_t)B * n_tiles * c.n * 2);
      last_partials = stats; last_tiles = n_tiles; last_cols = c.n;
      if (io.extra_stat_tiles) RUN((cudaMemsetAsync(stats, 0, (size_t)B * n_tiles * c.n * 8, st) == cudaSuccess) ? ST2_OK : ST2_ECUDA);
    }
    const int64_t m0 = ar.mark();
    const int cp3 = 3 * c.cin_pad;
    __nv_bfloat16* xs = ar.alloc<__nv_bfloat16>((int64_t)B * io.Lin * cp3);
    RUN(launch_split3(reinterpret_cast<const float*>(io.x), io.x_bstride, io.ldx, io.Lin, c.cin, c.cin_pad,
                      io.pre_norm ? ss_of(*io.pre_norm) : nullptr, h->ss_pairs, io.act, io.slope, io.alpha, xs, B, st));
    const bool post = io.res || io.accumulate || io.out_scale != 1.f || c.n16 > io.ldy || (io.ldy & 3) ||
                      (io.y_bstride & 3) || ((uintptr_t)io.y & 15);
    float* acc = post ? ar.alloc<float>((int64_t)B * io.Lout * c.n16) : reinterpret_cast<float*>(io.y);
    St2ConvParams p;
    memset(&p, 0, sizeof(p));
    p.B = B;
    p.x = xs; p.x_dtype = ST2_DT_BF16; p.x_bstride = (int64_t)io.Lin * cp3; p.ldx = cp3; p.Lin = io.Lin; p.cin = cp3;
    p.act = ST2_ACT_NONE;
    p.k = c.k; p.dil = c.dil; p.pad = c.pad; p.stride = 1;
    p.w = W<uint16_t>(c.w48); p.bias = W<float>(c.bias16); p.n_cols = c.n16; p.n_tile = c.nt48;
    p.y = acc; p.y_dtype = ST2_DT_F32;
    p.y_bstride = post ? (int64_t)io.Lout * c.n16 : io.y_bstride;
    p.ldy = post ? c.n16 : io.ldy;
    p.Lout = io.Lout;
    p.out_scale = 1.f;
    p.algo_flops = c.flops_per_row * (double)io.Lout * (double)B;
    RUN(launch_conv_umma(p, st));
    if (post)
      RUN(launch_post32(acc, c.n16, io.Lout, c.n, reinterpret_cast<const float*>(io.res), io.res_bstride, io.ldres,
                        io.res_shift, io.out_scale, io.accumulate, reinterpret_cast<float*>(io.y), io.y_bstride, io.ldy,
                        B, st));
    if (io.want_stats)
      RUN(launch_colstats<float>(reinterpret_cast<const float*>(io.y), io.y_bstride, io.ldy, io.Lout, 0, c.n, stats, B, st,
                                 n_tiles));
    ar.release(m0);
    return ST2_OK;
  }

  // runs one fused conv; partial statistics (if requested) are left in arena scratch that the
  // caller must consume (finalize) before the next arena allocation that follows a release.
  // St2ConvParams of one fused conv (statistics buffer allocated from the arena when requested); *umma = it runs on
  // the tcgen05 kernel
  int conv_params(const ConvW& c, const ConvIO& io, St2ConvParams& p, bool* umma_o) {
    // tcgen05 path: bf16 activations in; bf16 out, or fp32 out when the layer's padded width is the
    // row stride of the destination (conv_post -> [rows][32] fp32)
    const bool umma = DT<AT>::id == ST2_DT_BF16 && c.w16 >= 0 && io.x_dtype == ST2_DT_BF16 &&
                      io.ldx >= c.cin_pad &&
                      ((io.y_dtype == ST2_DT_BF16 && c.n16 == c.n) ||
                       (io.y_dtype == ST2_DT_F32 && c.n16 <= io.ldy && c.n16 <= 64 && !io.want_stats));
    memset(&p, 0, sizeof(p));
    p.B = B;
    p.x = io.x; p.x_dtype = io.x_dtype; p.x_bstride = io.x_bstride; p.ldx = io.ldx; p.Lin = io.Lin;
    p.cin = umma ? c.cin_pad : c.cin;
    if (io.pre_norm) {
      p.ss = ss_of(*io.pre_norm);
      p.ss_stride = h->ss_pairs;
    }
    p.act = io.act; p.slope = io.slope; p.alpha = io.alpha;
    p.k = c.k; p.dil = c.dil; p.pad = c.pad; p.stride = c.stride;
    p.w = umma ? (const void*)W<uint16_t>(c.w16) : (const void*)W<float>(c.w32);
    p.bias = W<float>(umma ? c.bias16 : c.bias);
    if (io.bias_b) {
      p.bias = io.bias_b;
      p.bias_bstride = c.n;
    }
    p.n_cols = umma ? c.n16 : c.n;
    p.y = io.y; p.y_dtype = io.y_dtype; p.y_bstride = io.y_bstride; p.ldy = io.ldy; p.Lout = io.Lout;
    p.res = io.res; p.res_bstride = io.res_bstride; p.ldres = io.ldres; p.res_shift = io.res_shift;
    p.out_scale = io.out_scale; p.accumulate = io.accumulate;
    p.algo_flops = c.flops_per_row * (double)io.Lout * (double)B;
    if (io.want_stats) {
      const int tiles = ceil_div(io.Lout, conv_tile_rows(umma));
      p.n_tiles = tiles + io.extra_stat_tiles;
      p.stats = ar.alloc<float>((int64_t)B * p.n_tiles * c.n * 2);
      last_partials = p.stats; last_tiles = p.n_tiles; last_cols = c.n;
      if (io.extra_stat_tiles) RUN((cudaMemsetAsync(p.stats, 0, (size_t)B * p.n_tiles * c.n * 8, st) == cudaSuccess) ? ST2_OK : ST2_ECUDA);
    }
    *umma_o = umma;
    return ST2_OK;
  }

  int conv(const ConvW& c, const ConvIO& io) {
    if (tc32 && DT<AT>::id == ST2_DT_F32 && c.w48 >= 0 && io.x_dtype == ST2_DT_F32 && io.y_dtype == ST2_DT_F32 &&
        !io.bias_b)
      return conv_tc32(c, io);
    St2ConvParams p;
    bool umma = false;
    int rc = conv_params(c, io, p, &umma);
    if (rc) return rc;
    RUN(umma ? launch_conv_umma(p, st) : launch_conv_simt(p, st));
    return ST2_OK;
  }

  // n (<= 3) convs of the same shape as ONE tcgen05 kernel whose CTAs alternate between them tile by tile
  // (launch_conv_umma_multi): the convs of the three resblocks k = 3 / 7 / 11 at the same position of their chains.
  // parts[j] / tiles[j] receive the statistics partials of conv j.  Other precisions run them one by one.
  int conv_multi(const ConvW* const* cs, const ConvIO* ios, int n, float** parts, int* tiles, int* cols) {
    St2ConvParams ps[3];
    bool all_umma = n <= 3;
    if (DT<AT>::id != ST2_DT_BF16) all_umma = false;
    if (all_umma) {
      for (int j = 0; j < n; ++j) {
        bool umma = false;
        int rc = conv_params(*cs[j], ios[j], ps[j], &umma);
        if (rc) return rc;
        parts[j] = last_partials; tiles[j] = last_tiles; cols[j] = last_cols;
        all_umma = all_umma && umma;
      }
      if (all_umma) {
        RUN(launch_conv_umma_multi(ps, n, st));
        // bookkeeping: one launch
        return ST2_OK;
      }
      for (int j = 0; j < n; ++j) RUN(launch_conv_simt(ps[j], st));  // (not reached for the resblock convs)
      return ST2_OK;
    }
    for (int j = 0; j < n; ++j) {
      int rc = conv(*cs[j], ios[j]);
      if (rc) return rc;
      parts[j] = last_partials; tiles[j] = last_tiles; cols[j] = last_cols;
    }
    return ST2_OK;
  }

  // independent convs of the same shape (conv1 + shortcut of an AdainResBlk1d, the two asr_res copies): one multi-conv
  // launch when all their tiles fit the SMs at once (a single utterance), one launch each otherwise.  Option "group" /
  // ST2_GROUP = 0: never.
  int conv_group(const ConvW* const* cs, const ConvIO* ios, int n) {
    const bool group_env = h->opt_group != 0;
    const int64_t tiles = (int64_t)B * ceil_div(ios[0].Lout, 256) * ceil_div(cs[0]->n16 > 0 ? cs[0]->n16 : cs[0]->n, 128);
    if (group_env && DT<AT>::id == ST2_DT_BF16 && !tc32 && n <= 3 && tiles * n <= 148) {
      float* parts[3]; int tl[3], cl[3];
      return conv_multi(cs, ios, n, parts, tl, cl);  // last_* = the last conv that asked for statistics
    }
    for (int j = 0; j < n; ++j) {
      int rc = conv(*cs[j], ios[j]);
      if (rc) return rc;
    }
    return ST2_OK;
  }

  ConvIO io_of(const Act<AT>& x, const Act<AT>& y) const {
    ConvIO io;
    io.x = x.p; io.x_dtype = DT<AT>::id; io.x_bstride = x.bstride; io.ldx = x.ld; io.Lin = x.L;
    io.y = y.p; io.y_dtype = DT<AT>::id; io.y_bstride = y.bstride; io.ldy = y.ld; io.Lout = y.L;
    return io;
  }
  Act<AT> new_act(int L, int C) {
    Act<AT> a;
    a.p = ar.alloc<AT>((int64_t)B * L * C);
    a.bstride = (int64_t)L * C; a.ld = C; a.L = L;
    return a;
  }

  // y = lrelu(AdaIN(x)) as its own pass (bf16 activations only)
  int norm_act(const Act<AT>& x, const NormW& n, float slope, const Act<AT>& y, int C) {
    return launch_norm_act_bf16(reinterpret_cast<const __nv_bfloat16*>(x.p), x.bstride, x.ld, ss_of(n), h->ss_pairs,
                                slope, reinterpret_cast<__nv_bfloat16*>(y.p), y.bstride, y.ld, B, x.L, C, st);
  }

  // AdainResBlk1d (istftnet.py:440-453).  `in` has r.din_pad columns; norm1 of `in` must already be
  // finalized.  Output goes to out (any ld).  want_out_stats leaves partials in last_*.
  int resblk1d(const ResBlk1dW& r, const Act<AT>& in, const Act<AT>& out, bool want_out_stats) {
    const int64_t m0 = ar.mark();
    const int L = in.L, Lo = r.up ? 2 * L : L;
    Act<AT> cin1 = in;
    if (r.up) {  // AdaIN + LeakyReLU + depthwise transposed conv, materialised at 2L
      cin1 = new_act(Lo, r.din_pad);
      RUN(launch_pool_up2<AT>(in.p, in.bstride, in.ld, ss_of(r.n1), h->ss_pairs, 0.2f, W<float>(r.pool_w3),
                              W<float>(r.pool_b), cin1.p, cin1.bstride, cin1.ld, B, L, r.din, r.din_pad, st));
    }
    Act<AT> hbuf = new_act(Lo, r.dout);
    Act<AT> sc = in;  // identity shortcut when dims match (never the case in this decoder)
    bool sc_done = false;
    // wide layers (N >= 512: every activation tile feeds >= 4 N tiles of the tcgen05 kernel): AdaIN +
    // LeakyReLU are applied once by an elementwise pass instead of once per N tile in the conv prologue
    const bool pre_act = DT<AT>::id == ST2_DT_BF16 && r.conv1.w16 >= 0 && r.dout >= 512;
    {
      if (pre_act && !r.up) {
        Act<AT> a1 = new_act(Lo, r.din_pad);
        RUN(norm_act(cin1, r.n1, 0.2f, a1, r.din_pad));
        cin1 = a1;
      }
      ConvIO io = io_of(cin1, hbuf);
      if (!r.up && !pre_act) { io.pre_norm = &r.n1; io.act = ST2_ACT_LRELU; io.slope = 0.2f; }
      io.want_stats = true;
      // conv1 and the 1x1 shortcut read the same rows and are independent: one launch for both when the pair still
      // fits the SMs (a single utterance: 2 x 8 CTAs)
      if (r.has_sc && !r.up) {
        sc = new_act(L, r.dout);
        sc_done = true;
        const int64_t m1 = ar.mark();
        const ConvW* cs2[2] = {&r.conv1, &r.conv1x1};
        ConvIO ios2[2] = {io, io_of(in, sc)};
        int rc = conv_group(cs2, ios2, 2);  // statistics of conv1 = the only ones asked for
        if (rc) return rc;
        rc = finalize(last_partials, last_tiles, last_cols, r.dout, Lo, r.n2, 0);
        if (rc) return rc;
        ar.release(m1);
      } else {
        const int64_t m1 = ar.mark();
        int rc = conv(r.conv1, io);
        if (rc) return rc;
        rc = finalize(last_partials, last_tiles, last_cols, r.dout, Lo, r.n2, 0);
        if (rc) return rc;
        ar.release(m1);
      }
    }
    if (r.has_sc && !sc_done) {
      sc = new_act(L, r.dout);
      ConvIO io = io_of(in, sc);
      int rc = conv(r.conv1x1, io);
      if (rc) return rc;
    }
    {
      Act<AT> cin2 = hbuf;
      if (pre_act) {
        cin2 = new_act(Lo, r.dout);
        RUN(norm_act(hbuf, r.n2, 0.2f, cin2, r.dout));
      }
      ConvIO io = io_of(cin2, out);
      if (!pre_act) { io.pre_norm = &r.n2; io.act = ST2_ACT_LRELU; io.slope = 0.2f; }
      io.res = sc.p; io.res_bstride = sc.bstride; io.ldres = sc.ld; io.res_shift = r.up ? 1 : 0;
      io.out_scale = 0.70710678118654752440f;  // 1/sqrt(2), istftnet.py:452
      io.want_stats = want_out_stats;
      // partials of this conv stay allocated for the caller (it finalizes, then releases)
      int rc = conv(r.conv2, io);
      if (rc) return rc;
    }
    if (!want_out_stats) ar.release(m0);
    return ST2_OK;
  }

  // AdaINResBlock1 (istftnet.py:66-75).  x: block input whose adain1[0] scale/shift are finalized.
  // run: private running buffer (may alias x for the in-place noise_res), tmp: conv1 output.
  // final: destination of the last iteration; final_scale / final_acc implement mean over resblocks.
  // pr: the views are row pairs ([L/2][2C]) and the block runs in its row-pair form (Loader::pair_conv).
  int resblock1(const ResBlock1W& r, const Act<AT>& x, const Act<AT>& run, const Act<AT>& tmp,
                const Act<AT>& fin, float final_scale, int final_acc, const float* c2_bias0 = nullptr,
                bool pr = false) {
    const int reps = pr ? 2 : 1;
    for (int it = 0; it < 3; ++it) {
      const Act<AT>& src = it == 0 ? x : run;
      const bool last = it == 2;
      const int64_t m = ar.mark();
      {
        ConvIO io = io_of(src, tmp);
        io.pre_norm = &r.n1[it]; io.act = ST2_ACT_SNAKE; io.alpha = W<float>(pr ? r.a1p[it] : r.a1[it]);
        io.want_stats = true;
        int rc = conv(pr ? r.c1p[it] : r.c1[it], io);
        if (rc) return rc;
        rc = finalize(last_partials, last_tiles, last_cols, r.C, (int64_t)tmp.L * reps, r.n2[it], 0, reps);
        if (rc) return rc;
        ar.release(m);
      }
      {
        const Act<AT>& dst = last ? fin : run;
        ConvIO io = io_of(tmp, dst);
        io.pre_norm = &r.n2[it]; io.act = ST2_ACT_SNAKE; io.alpha = W<float>(pr ? r.a2p[it] : r.a2[it]);
        io.res = src.p; io.res_bstride = src.bstride; io.ldres = src.ld;
        if (last) { io.out_scale = final_scale; io.accumulate = final_acc; }
        if (it == 0) io.bias_b = c2_bias0;
        io.want_stats = !last;
        int rc = conv(pr ? r.c2p[it] : r.c2[it], io);
        if (rc) return rc;
        if (!last) {
          rc = finalize(last_partials, last_tiles, last_cols, r.C, (int64_t)dst.L * reps, r.n1[it + 1], 0, reps);
          if (rc) return rc;
        }
        ar.release(m);
      }
    }
    return ST2_OK;
  }

  // one chain position of the lock-step schedule: ONE launch of all nk convs.  (ST2_LOCKSTEP=2, an A/B switch: large
  // problems run the k-smallest + k-largest conv fused and the middle one alone.  The microbenchmark has that 2 % ahead
  // of three in one -- the SIMT-bound k = 3 tiles pair best with the tensor-bound k = 11 tiles -- but whole C2 / C4
  // steps on one box do not: 20.05 vs 20.00 ms and 41.9 vs 41.7 ms, profiles/r2_lockstep_ab.txt.)
  int conv_position(const ConvW* const* cs, const ConvIO* ios, int nk, float** parts, int* tiles, int* cols) {
    const int64_t tiles_per_conv = (int64_t)B * ceil_div(ios[0].Lout, 256) * ceil_div(cs[0]->n, 128);
    if (nk != 3 || tiles_per_conv < 2 * 148 || lockstep_mode != 2) return conv_multi(cs, ios, nk, parts, tiles, cols);
    // (mode 2 is only reached with ST2_LOCKSTEP=2, which also forces lock step for every size)
    const ConvW* c2[2] = {cs[0], cs[2]};
    ConvIO io2[2] = {ios[0], ios[2]};
    float* p2[2]; int t2[2], n2[2];
    int rc = conv_multi(c2, io2, 2, p2, t2, n2);
    if (rc) return rc;
    parts[0] = p2[0]; tiles[0] = t2[0]; cols[0] = n2[0];
    parts[2] = p2[1]; tiles[2] = t2[1]; cols[2] = n2[1];
    rc = conv_multi(cs + 1, ios + 1, 1, parts + 1, tiles + 1, cols + 1);
    return rc;
  }
  int lockstep_mode = 0;  // ST2_LOCKSTEP: 0 off, 1 (default) small problems only, 2 see above, 3 every size

  // The nk resblocks of a generator stage (istftnet.py:369-374: xs = sum_j resblocks[i*nk+j](x); x = xs / nk) run in
  // lock step: their convs at the same position of the chain have the same shape and no dependence on each other, so
  // each position is ONE kernel launch whose CTAs alternate between the resblocks tile by tile (conv_multi) -- the
  // k = 3 conv's per-element work overlaps the k = 11 conv's MMAs, and a single utterance fills 3x the SMs.  Only the
  // last conv of every chain is launched on its own: it accumulates the running mean into `fin` (resblock j reads what
  // resblock j-1 wrote).  Each resblock needs its own run / tmp buffers.
  int resblocks_lockstep(const ResBlock1W* rs, int nk, const Act<AT>& x, const Act<AT>* run, const Act<AT>* tmp,
                         const Act<AT>& fin, bool pr) {
    const int reps = pr ? 2 : 1;
    const ConvW* cs[3];
    ConvIO ios[3];
    float* parts[3];
    int tiles[3], cols[3];
    for (int it = 0; it < 3; ++it) {
      const bool last = it == 2;
      const int64_t m = ar.mark();
      for (int j = 0; j < nk; ++j) {
        const ResBlock1W& r = rs[j];
        const Act<AT>& src = it == 0 ? x : run[j];
        ios[j] = io_of(src, tmp[j]);
        ios[j].pre_norm = &r.n1[it]; ios[j].act = ST2_ACT_SNAKE; ios[j].alpha = W<float>(pr ? r.a1p[it] : r.a1[it]);
        ios[j].want_stats = true;
        cs[j] = pr ? &r.c1p[it] : &r.c1[it];
      }
      int rc = conv_position(cs, ios, nk, parts, tiles, cols);
      if (rc) return rc;
      const NormW* norms[3];
      for (int j = 0; j < nk; ++j) norms[j] = &rs[j].n2[it];
      rc = finalize_multi(parts, nk, tiles[0], cols[0], rs[0].C, (int64_t)tmp[0].L * reps, norms, reps);
      if (rc) return rc;
      ar.release(m);
      for (int j = 0; j < nk; ++j) {
        const ResBlock1W& r = rs[j];
        const Act<AT>& src = it == 0 ? x : run[j];
        const Act<AT>& dst = last ? fin : run[j];
        ios[j] = io_of(tmp[j], dst);
        ios[j].pre_norm = &r.n2[it]; ios[j].act = ST2_ACT_SNAKE; ios[j].alpha = W<float>(pr ? r.a2p[it] : r.a2[it]);
        ios[j].res = src.p; ios[j].res_bstride = src.bstride; ios[j].ldres = src.ld;
        if (last) { ios[j].out_scale = 1.f / (float)nk; ios[j].accumulate = j > 0 ? 1 : 0; }
        ios[j].want_stats = !last;
        cs[j] = pr ? &r.c2p[it] : &r.c2[it];
      }
      if (last) {
        for (int j = 0; j < nk; ++j) {
          rc = conv(*cs[j], ios[j]);
          if (rc) return rc;
        }
      } else {
        rc = conv_position(cs, ios, nk, parts, tiles, cols);
        if (rc) return rc;
        const NormW* norms[3];
        for (int j = 0; j < nk; ++j) norms[j] = &rs[j].n1[it + 1];
        rc = finalize_multi(parts, nk, tiles[0], cols[0], rs[0].C, (int64_t)run[0].L * reps, norms, reps);
        if (rc) return rc;
      }
      ar.release(m);
    }
    return ST2_OK;
  }

  int run(const float* asr, const float* f0, const float* nn, const float* s, float* wav,
          const St2ForwardOpts& o);
  int run_f0n(const float* x, const float* s, float* f0_out, float* n_out);
};

// ProsodyPredictor.F0Ntrain (models.py:638-654): x [B, d_hid + style, T] NCL -> F0, N [B, 2T]
template <typename AT>
int Fwd<AT>::run_f0n(const float* x, const float* s, float* f0_out, float* n_out) {
  const St2Config& c = h->cfg;
  const int d = c.dim_in, hid = d / 2, in = d + c.style_dim;
  float* hall_w = ar.alloc<float>((int64_t)B * h->fc_rows);
  hall = hall_w;
  ss = ar.alloc<float>((int64_t)B * h->ss_pairs * 2);
  RUN(launch_style_fc(s, W<float>(h->fcW), W<float>(h->fcB), hall_w, B, c.style_dim, h->fc_rows, st));
  RUN(cudaMemsetAsync(ss, 0, (size_t)B * h->ss_pairs * 8, st) == cudaSuccess ? ST2_OK : ST2_ECUDA);

  // ---- x, _ = self.shared(x.transpose(-1, -2)) (models.py:639)
  Act<AT> hx = new_act(T, d);  // LSTM output, channels-last [B][T][2H]
  {
    const int64_t m = ar.mark();
    Act<AT> xc = new_act(T, in), pre = new_act(T, h->lstm_ih.n);
    RUN(launch_ncl_to_act<AT>(x, in, T, xc.p, in, B, st));
    int rc = conv(h->lstm_ih, io_of(xc, pre));  // W_ih x + b_ih + b_hh for every step and both directions
    if (rc) return rc;
    RUN(launch_lstm_scan<AT>(pre.p, W<float>(h->lstm_whh), hx.p, B, T, st));
    ar.release(m);
  }
  // statistics of the LSTM output for norm1 of the first block of both branches
  {
    const int64_t m = ar.mark();
    const int tiles = colstats_tiles(T);
    float* part = ar.alloc<float>((int64_t)B * tiles * d * 2);
    RUN(launch_colstats<AT>(hx.p, hx.bstride, hx.ld, T, 0, d, part, B, st));
    for (int br = 0; br < 2; ++br) {
      int rc = finalize(part, tiles, d, d, T, h->f0n_blk[br][0].n1, 0);
      if (rc) return rc;
    }
    ar.release(m);
  }
  for (int br = 0; br < 2; ++br) {
    const int64_t m = ar.mark();
    Act<AT> a1 = new_act(T, d), a2 = new_act(2 * T, hid), a3 = new_act(2 * T, hid);
    const Act<AT> ins[3] = {hx, a1, a2}, outs[3] = {a1, a2, a3};
    for (int i = 0; i < 3; ++i) {
      const int64_t m2 = ar.mark();
      const bool more = i < 2;
      int rc = resblk1d(h->f0n_blk[br][i], ins[i], outs[i], more);
      if (rc) return rc;
      if (more) {
        rc = finalize(last_partials, last_tiles, last_cols, h->f0n_blk[br][i].dout, outs[i].L,
                      h->f0n_blk[br][i + 1].n1, 0);
        if (rc) return rc;
      }
      ar.release(m2);
    }
    // F0_proj / N_proj: Conv1d(hid, 1, 1) -> [B, 2T] fp32
    ConvIO io;
    io.x = a3.p; io.x_dtype = DT<AT>::id; io.x_bstride = a3.bstride; io.ldx = a3.ld; io.Lin = a3.L;
    float* dst = br == 0 ? f0_out : n_out;
    io.y = dst; io.y_dtype = ST2_DT_F32; io.y_bstride = 2 * T; io.ldy = 1; io.Lout = 2 * T;
    int rc = conv(h->f0n_proj[br], io);
    if (rc) return rc;
    ar.release(m);
  }
  return ST2_OK;
}

template <typename AT>
int Fwd<AT>::run(const float* asr, const float* f0, const float* nn, const float* s, float* wav,
                 const St2ForwardOpts& o) {
  const St2Config& c = h->cfg;
  const bool istft = c.kind == ST2_KIND_ISTFTNET;
  const int T2 = T / 2;
  const int spf = h->spf;
  const int64_t Lw = (int64_t)T * spf;

  NvtxRange nvtx("st2.decoder.style_fc", !dry);
  // ---- all AdaIN fc layers at once; zeroed (scale, shift) arena (padding channels stay 0)
  float* hall_w = ar.alloc<float>((int64_t)B * h->fc_rows);
  hall = hall_w;
  ss = ar.alloc<float>((int64_t)B * h->ss_pairs * 2);
  RUN(launch_style_fc(s, W<float>(h->fcW), W<float>(h->fcB), hall_w, B, c.style_dim, h->fc_rows, st));
  RUN(cudaMemsetAsync(ss, 0, (size_t)B * h->ss_pairs * 8, st) == cudaSuccess ? ST2_OK : ST2_ECUDA);

  // ---- stage geometry of the generator (istftnet.py:358-374 / hifigan.py:327-341)
  const int nk = c.n_resblock_kernels;
  struct StageGeo { int Lin, Ls, C, pad_rows, reps; bool pr; };
  StageGeo geo[ST2_MAX_UPS];
  for (int i = 0, Lin = T; i < c.n_ups; ++i) {
    StageGeo& g = geo[i];
    const UpW& up = h->ups[i];
    g.Lin = Lin; g.C = up.cout;
    g.pad_rows = (istft && i == c.n_ups - 1) ? 1 : 0;  // ReflectionPad1d((1,0)), istftnet.py:364-365
    g.Ls = Lin * up.rate + g.pad_rows;
    // a 32-channel stage runs on row pairs ([Ls/2][64] views of the same memory) so that its convs
    // fit the tcgen05 kernel's 64-channel chunks (Loader::pair_conv)
    g.pr = DT<AT>::id == ST2_DT_BF16 && g.C == 32 && g.pad_rows == 0 && g.Ls % 2 == 0 && up.rate % 2 == 0 &&
           h->noise_res[i].paired;
    for (int j = 0; j < nk; ++j) g.pr = g.pr && h->resblocks[i * nk + j].paired;
    g.reps = g.pr ? 2 : 1;
    Lin = g.Ls;
  }
  // fork the excitation branch when the last stage has fewer 256-row tiles than 2 x SMs (a single utterance: the step
  // is a chain of ~100 short dependent launches, a third of them in this branch); option "fork" / ST2_FORK: 0 never, 2 always
  const bool forked = h->side_st && h->opt_fork != 0 &&
                      ((int64_t)B * ceil_div(geo[c.n_ups - 1].Ls, 256) < 2 * 148 || h->opt_fork >= 2);
  st_side = h->side_st;

  const int n_fft = c.gen_istft_n_fft, hop = c.gen_istft_hop_size;
  const int L2 = istft ? (int)(Lw / hop) + 1 : 0;
  const int ldh = 32;
  const float* hs = o.har_source;
  bool hilo = istft && DT<AT>::id == ST2_DT_BF16;
  int rows_pad = L2;
  float* har = nullptr;
  __nv_bfloat16* har16 = nullptr;
  const bool meanfree = !istft && DT<AT>::id == ST2_DT_BF16;
  float* hs_mean = nullptr;
  Act<AT> xs_of[ST2_MAX_UPS];
  float* bias_c2_of[ST2_MAX_UPS] = {};

  // ---- excitation (istftnet.py:350-356 / hifigan.py:320-323)
  auto excitation = [&]() -> int {
  // bf16 path: `har` is kept as bf16 (hi, lo) rows and the noise convs run on the tensor cores
  if (hilo) {
    int len = T;
    for (int i = 0; i < c.n_ups; ++i) {
      len = len * c.upsample_rates[i] + (i == c.n_ups - 1 ? 1 : 0);
      hilo = hilo && h->noise_convs16[i].w16 >= 0;
      const int need = h->noise_group[i] > 1 ? h->noise_group[i] * (len + 1) : len;
      if (need > rows_pad) rows_pad = need;
    }
  }
  if (istft && hilo) har16 = ar.alloc<__nv_bfloat16>((int64_t)B * rows_pad * 64);
  else if (istft) har = ar.alloc<float>((int64_t)B * L2 * ldh);
  if (istft && o.har) {
    if (hilo) RUN(launch_ncl_to_hilo(o.har, n_fft + 2, L2, rows_pad, har16, B, st));
    else RUN(launch_ncl_to_cl(o.har, n_fft + 2, L2, har, ldh, B, st));
  } else {
    if (!hs) {
      float* ph = ar.alloc<float>((int64_t)B * 9 * T);
      float* hsw = ar.alloc<float>((int64_t)B * Lw);
      RUN(launch_phase_scan(f0, B, T, spf, ph, st));
      RUN(launch_har_source(f0, ph, o.noise, o.seed, W<float>(h->lin_w), h->lin_b, B, T, spf, hsw, st, o.item_offset, o.seed_dev));
      hs = hsw;
    }
    if (istft && hilo) RUN(launch_stft_hilo(hs, B, (int)Lw, rows_pad, har16, st));
    else if (istft) RUN(launch_stft(hs, B, (int)Lw, har, ldh, st));
  }
  if (o.har_source_out && hs)
    RUN(cudaMemcpyAsync(o.har_source_out, hs, (size_t)B * Lw * 4, cudaMemcpyDeviceToDevice, st) == cudaSuccess ? ST2_OK : ST2_ECUDA);
  if (o.har_out && istft) {
    if (hilo) RUN(launch_hilo_to_ncl(har16, rows_pad, n_fft + 2, L2, o.har_out, B, st));
    else RUN(launch_cl_to_ncl(har, ldh, n_fft + 2, L2, o.har_out, B, st));
  }
  // hifigan, bf16 path: x_source is stored without its per-(b,c) constant (elementwise.cu, noise_bias)
  if (meanfree) {
    hs_mean = ar.alloc<float>(B);
    RUN(launch_row_mean(hs, Lw, hs_mean, B, st));
  }
  return ST2_OK;
  };

  // ---- x_source of stage i = noise_res[i](noise_convs[i](har | har_source)) into xs, tmp = scratch of the resblock
  auto x_source = [&](int i, const Act<AT>& xs, const Act<AT>& tmp) -> int {
    const StageGeo& g = geo[i];
    const int C = g.C, Ls = g.Ls, reps = g.reps;
    const bool pr = g.pr;
    auto vw = [pr](const Act<AT>& a) {
      Act<AT> v = a;
      if (pr) { v.L = a.L / 2; v.ld = a.ld * 2; }
      return v;
    };
    const ConvW& nc = hilo ? h->noise_convs16[i] : h->noise_convs[i];
    ConvIO io;
    if (hilo) {  // sf har rows = one wide row of sf * 64 channels
      const int sf = h->noise_group[i];
      io.x = har16; io.x_dtype = ST2_DT_BF16; io.x_bstride = (int64_t)rows_pad * 64; io.ldx = sf * 64;
      io.Lin = sf > 1 ? Ls + 1 : L2;
    } else if (istft) {
      io.x = har; io.x_dtype = ST2_DT_F32; io.x_bstride = (int64_t)L2 * ldh; io.ldx = ldh; io.Lin = L2;
    } else {
      io.x = hs; io.x_dtype = ST2_DT_F32; io.x_bstride = Lw; io.ldx = 1; io.Lin = (int)Lw;
    }
    io.y = xs.p; io.y_dtype = DT<AT>::id; io.y_bstride = xs.bstride; io.ldy = xs.ld; io.Lout = Ls;
    io.want_stats = true;
    float* bias_c2 = nullptr;
    if (meanfree) {
      float* bias_nc = ar.alloc<float>((int64_t)B * C);
      bias_c2 = ar.alloc<float>((int64_t)B * C * reps);
      RUN(launch_noise_bias(hs_mean, W<float>(h->noise_wsum[i]), W<float>(nc.bias),
                            W<float>(h->noise_res[i].c2[0].bias), bias_nc, bias_c2, C, B, st, reps));
      io.bias_b = bias_nc;
    }
    const int64_t m = ar.mark();
    int rc = conv(nc, io);
    if (rc) return rc;
    rc = finalize(last_partials, last_tiles, last_cols, C, Ls, h->noise_res[i].n1[0], 0, reps);
    if (rc) return rc;
    ar.release(m);
    return resblock1(h->noise_res[i], vw(xs), vw(xs), vw(tmp), vw(xs), 1.f, 0, bias_c2, pr);
  };

  if (forked) {
    // fork: everything below up to to_main() is issued on the second stream, from the side arena
    RUN(cudaEventRecord(h->ev_fork, st) == cudaSuccess ? ST2_OK : ST2_ECUDA);
    to_side();
    RUN(cudaStreamWaitEvent(st, h->ev_fork, 0) == cudaSuccess ? ST2_OK : ST2_ECUDA);
    if (!dry) pdl_plain_next();
    int rc = excitation();
    for (int i = 0; i < c.n_ups && !rc; ++i) {
      xs_of[i] = new_act(geo[i].Ls, geo[i].C);
      const int64_t m = ar.mark();
      Act<AT> tmp = new_act(geo[i].Ls, geo[i].C);
      rc = x_source(i, xs_of[i], tmp);
      ar.release(m);
      if (!rc) RUN(cudaEventRecord(h->ev_join[i], st) == cudaSuccess ? ST2_OK : ST2_ECUDA);
    }
    to_main();
    if (rc) return rc;
  }

  nvtx.next("st2.decoder.encode_decode");
  // ---- encode / decode (istftnet.py:510-524)
  const int din0 = c.dim_in + 2, ld0 = (int)round_up(din0, 64);
  const int ldc = h->decode[0].din_pad, c_f0 = 1024 + 64;  // [x(1024) | asr_res(64) | F0 | N | 0...]
  Act<AT> xg = new_act(T, 512);  // decode.3 output = generator input; lives past the decode scratch
  const int64_t m_dec = ar.mark();
  {
    Act<AT> x0 = new_act(T2, ld0), ca = new_act(T2, ldc), cb = new_act(T2, ldc);
    RUN(launch_prep_inputs<AT>(asr, f0, nn, W<float>(h->f0w), W<float>(h->nw), x0.p, ld0, ca.p, cb.p, ldc,
                               c_f0, B, c.dim_in, T2, st, o.align_idx, o.align_len));
    // asr_res = Conv1d(512 -> 64, k1)(asr) written into columns [1024, 1088) of both concat buffers
    {
      const ConvW* cs2[2] = {&h->asr_res, &h->asr_res};
      ConvIO ios2[2] = {io_of(x0, ca), io_of(x0, cb)};
      ios2[0].y = ca.p + 1024;
      ios2[1].y = cb.p + 1024;
      int rc = conv_group(cs2, ios2, 2);
      if (rc) return rc;
    }
    // statistics of tensors no conv epilogue produced: x0 (encode.norm1) and the constant concat columns
    const int cs_tiles = colstats_tiles(T2);
    float* p_x0 = ar.alloc<float>((int64_t)B * cs_tiles * din0 * 2);
    const int n_const = h->decode[0].din - 1024;  // 66
    float* p_const = ar.alloc<float>((int64_t)B * cs_tiles * n_const * 2);
    RUN(launch_colstats<AT>(x0.p, x0.bstride, x0.ld, T2, 0, din0, p_x0, B, st));
    RUN(launch_colstats<AT>(ca.p, ca.bstride, ca.ld, T2, 1024, n_const, p_const, B, st));
    int rc = finalize(p_x0, cs_tiles, din0, din0, T2, h->encode.n1, 0);
    if (rc) return rc;
    {
      float* pp4[4]; const NormW* norms4[4];
      for (int i = 0; i < 4; ++i) { pp4[i] = p_const; norms4[i] = &h->decode[i].n1; }
      rc = finalize_multi(pp4, 4, cs_tiles, n_const, n_const, T2, norms4, 1, 1024);
      if (rc) return rc;
    }
    // encode: x0 -> ca[:, :1024]
    Act<AT> in = x0;
    Act<AT> bufs[2] = {ca, cb};
    for (int blk = 0; blk < 5; ++blk) {
      const ResBlk1dW& r = blk == 0 ? h->encode : h->decode[blk - 1];
      const bool lastb = blk == 4;
      Act<AT> out = lastb ? xg : bufs[blk & 1];
      if (!lastb) out.L = T2;
      const int64_t m = ar.mark();
      rc = resblk1d(r, in, out, !lastb);
      if (rc) return rc;
      if (!lastb) {
        rc = finalize(last_partials, last_tiles, last_cols, 1024, T2, h->decode[blk].n1, 0);
        if (rc) return rc;
        in = bufs[blk & 1];
      }
      ar.release(m);
    }
  }
  ar.release(m_dec);

  if (o.tap_decode) RUN(launch_act_to_ncl<AT>(xg.p, xg.bstride, xg.ld, 512, T, o.tap_decode, B, st));

  if (!forked) {
    nvtx.next("st2.generator.excitation");
    int rc = excitation();
    if (rc) return rc;
  }

  nvtx.next("st2.generator.stages");
  // ---- generator stages (istftnet.py:358-374 / hifigan.py:327-341)
  Act<AT> x = xg;
  // bf16 path: the resblocks of a stage run in lock step, one fused launch per position (ST2_LOCKSTEP=0: one by one)
  static const int lockstep_env = [] { const char* e = getenv("ST2_LOCKSTEP"); return e ? atoi(e) : 1; }();
  const bool lockstep = lockstep_env != 0 && DT<AT>::id == ST2_DT_BF16;
  lockstep_mode = lockstep_env;
  for (int i = 0; i < c.n_ups; ++i) {
    const UpW& up = h->ups[i];
    const StageGeo& g = geo[i];
    const int Lin = g.Lin, C = g.C, pad_rows = g.pad_rows, Ls = g.Ls, reps = g.reps;
    const bool pr = g.pr;
    Act<AT> xs = forked ? xs_of[i] : new_act(Ls, C);
    Act<AT> xm = new_act(Ls, C), tmp = new_act(Ls, C), runb = new_act(Ls, C), outb = new_act(Ls, C);
    auto vw = [pr](const Act<AT>& a) {
      Act<AT> v = a;
      if (pr) { v.L = a.L / 2; v.ld = a.ld * 2; }
      return v;
    };
    if (forked) {  // join: x_source of this stage comes from the second stream
      RUN(cudaStreamWaitEvent(st, h->ev_join[i], 0) == cudaSuccess ? ST2_OK : ST2_ECUDA);
      if (!dry) pdl_plain_next();
    } else {
      int rc = x_source(i, xs, tmp);
      if (rc) return rc;
    }
    // x = ups[i](act(x)) (+ reflection pad) + x_source, with statistics for the three resblocks
    {
      ConvIO io;
      io.x = x.p; io.x_dtype = DT<AT>::id; io.x_bstride = x.bstride; io.ldx = x.ld; io.Lin = Lin;
      if (istft) { io.act = ST2_ACT_LRELU; io.slope = 0.1f; }
      else { io.act = ST2_ACT_SNAKE; io.alpha = W<float>(h->alphas[i]); }
      io.y = xm.p + (int64_t)pad_rows * C; io.y_dtype = DT<AT>::id; io.y_bstride = xm.bstride;
      io.ldy = up.rate * C; io.Lout = Lin;
      io.res = xs.p + (int64_t)pad_rows * C; io.res_bstride = xs.bstride; io.ldres = up.rate * C;
      io.want_stats = true; io.extra_stat_tiles = pad_rows;
      const int64_t m = ar.mark();
      int rc = conv(up.conv, io);
      if (rc) return rc;
      if (pad_rows) {
        float* slot = last_partials + (int64_t)(last_tiles - 1) * last_cols * 2;
        RUN(launch_reflect_row0<AT>(x.p, x.bstride, x.ld, 0.1f, W<float>(up.row0_tap), W<float>(up.conv.bias),
                                    xs.p, xs.bstride, xm.p, xm.bstride, C, slot,
                                    (int64_t)last_tiles * last_cols * 2, B, st));
      }
      if (nk <= 3) {
        float* pp3[3]; const NormW* norms[3];
        for (int j = 0; j < nk; ++j) { pp3[j] = last_partials; norms[j] = &h->resblocks[i * nk + j].n1[0]; }
        rc = finalize_multi(pp3, nk, last_tiles, last_cols, C, Ls, norms, reps);
        if (rc) return rc;
      } else {
        for (int j = 0; j < nk; ++j) {
          rc = finalize(last_partials, last_tiles, last_cols, C, Ls, h->resblocks[i * nk + j].n1[0], 0, reps);
          if (rc) return rc;
        }
      }
      ar.release(m);
    }
    // lock step only where it pays: a conv with fewer tiles than 2 x SMs (a single utterance) gains from one launch of
    // three convs (3 x the SMs busy, two launches less per position: C1 1.68 -> 1.49 ms); large batches run one conv per
    // launch on conv_umma.cu's kernel, whose MMA issue loop is 5 % faster (conv_umma_multi.cu header).  ST2_LOCKSTEP=3
    // forces it for every size, 0 turns it off.
    const int64_t tiles_per_conv = (int64_t)B * ceil_div(pr ? Ls / 2 : Ls, 256) * ceil_div(pr ? 2 * C : C, 128);
    if (lockstep && nk <= 3 && (tiles_per_conv < 2 * 148 || lockstep_mode >= 2)) {
      // the nk resblocks in lock step: each needs its own run / tmp buffers (resblock 0 reuses the stage's)
      Act<AT> runs[3], tmps[3];
      runs[0] = vw(runb); tmps[0] = vw(tmp);
      for (int j = 1; j < nk; ++j) { runs[j] = vw(new_act(Ls, C)); tmps[j] = vw(new_act(Ls, C)); }
      int rc = resblocks_lockstep(&h->resblocks[i * nk], nk, vw(xm), runs, tmps, vw(outb), pr);
      if (rc) return rc;
    } else {
      for (int j = 0; j < nk; ++j) {
        int rc = resblock1(h->resblocks[i * nk + j], vw(xm), vw(runb), vw(tmp), vw(outb), 1.f / (float)nk,
                           j > 0 ? 1 : 0, nullptr, pr);
        if (rc) return rc;
      }
    }
    x = outb;
    if (o.tap_stage[i]) RUN(launch_act_to_ncl<AT>(x.p, x.bstride, x.ld, C, Ls, o.tap_stage[i], B, st));
  }

  nvtx.next("st2.generator.head");
  // ---- output head
  if (istft) {
    float* cp = ar.alloc<float>((int64_t)B * L2 * ldh);
    ConvIO io;
    io.x = x.p; io.x_dtype = DT<AT>::id; io.x_bstride = x.bstride; io.ldx = x.ld; io.Lin = x.L;
    io.act = ST2_ACT_LRELU; io.slope = 0.01f;  // F.leaky_relu default slope, istftnet.py:375
    io.y = cp; io.y_dtype = ST2_DT_F32; io.y_bstride = (int64_t)L2 * ldh; io.ldy = ldh; io.Lout = L2;
    int rc = conv(h->conv_post, io);
    if (rc) return rc;
    RUN(launch_istft(cp, ldh, B, L2, wav, st, DT<AT>::id == ST2_DT_BF16));  // bf16 path: SFU exp / sin heads
  } else if (DT<AT>::id == ST2_DT_BF16 && h->conv_post_p.w16 >= 0 && x.ld == 32 && x.L % 2 == 0) {
    // row-pair form on the tensor cores: fp32 rows of 32 columns, columns 0/1 = samples 2q / 2q+1
    const int Lp = x.L / 2, ldp = h->conv_post_p.n16;
    float* cp = ar.alloc<float>((int64_t)B * Lp * ldp);
    ConvIO io;
    io.x = x.p; io.x_dtype = DT<AT>::id; io.x_bstride = x.bstride; io.ldx = 2 * x.ld; io.Lin = Lp;
    io.act = ST2_ACT_SNAKE; io.alpha = W<float>(h->alpha_post_p);  // hifigan.py:342
    io.y = cp; io.y_dtype = ST2_DT_F32; io.y_bstride = (int64_t)Lp * ldp; io.ldy = ldp; io.Lout = Lp;
    int rc = conv(h->conv_post_p, io);
    if (rc) return rc;
    RUN(launch_cols_tanh(cp, ldp, 2, wav, (int64_t)B * Lw, st));
  } else {
    ConvIO io;
    io.x = x.p; io.x_dtype = DT<AT>::id; io.x_bstride = x.bstride; io.ldx = x.ld; io.Lin = x.L;
    io.act = ST2_ACT_SNAKE; io.alpha = W<float>(h->alphas[c.n_ups]);  // hifigan.py:342
    io.y = wav; io.y_dtype = ST2_DT_F32; io.y_bstride = Lw; io.ldy = 1; io.Lout = (int)Lw;
    int rc = conv(h->conv_post, io);
    if (rc) return rc;
    RUN(launch_tanh_inplace(wav, (int64_t)B * Lw, st));
  }
  return ST2_OK;
#undef RUN
}

}  // namespace st2

// ============================================================================ C ABI
using namespace st2;

static int validate_cfg(const St2Config& c) {
  if (c.kind == ST2_KIND_SAMPLER) {
    ST2_REQUIRE(c.dim_in >= 2 && c.dim_in % 2 == 0 && c.dim_out >= 1, "channels / context_embedding_features");
    ST2_REQUIRE((c.dim_in + c.dim_out) % 64 == 0, "channels + context_embedding_features must be a multiple of 64");
    ST2_REQUIRE(c.style_dim >= 0, "context_features");
    ST2_REQUIRE(c.n_resblock_kernels >= 1 && c.n_resblock_kernels <= 64, "num_layers");
    ST2_REQUIRE(c.resblock_kernel_sizes[0] >= 1 && c.resblock_kernel_sizes[1] == 64,
                "num_heads >= 1, head_features must be 64 (attention kernel)");
    ST2_REQUIRE(c.resblock_kernel_sizes[2] >= 1 && c.resblock_kernel_sizes[3] >= 1, "multiplier / embedding_max_length");
    return ST2_OK;
  }
  if (c.kind == ST2_KIND_F0N) {
    ST2_REQUIRE(c.dim_in == 512, "d_hid must be 512 (the LSTM kernel keeps 256 hidden units per direction)");
    ST2_REQUIRE(c.style_dim > 0 && c.style_dim <= 192 && (c.dim_in + c.style_dim) % 64 == 0, "style_dim");
    return ST2_OK;
  }
  ST2_REQUIRE(c.kind == ST2_KIND_ISTFTNET || c.kind == ST2_KIND_HIFIGAN, "kind");
  ST2_REQUIRE(c.dim_in == 512, "dim_in must be 512 (asr_res is Conv1d(512, 64), istftnet.py:490)");
  ST2_REQUIRE(c.style_dim > 0 && c.style_dim <= 192, "style_dim");
  ST2_REQUIRE(c.upsample_initial_channel == 512, "upsample_initial_channel must equal decode.3's 512 outputs");
  ST2_REQUIRE(c.n_ups >= 1 && c.n_ups <= ST2_MAX_UPS, "n_ups");
  ST2_REQUIRE(c.n_resblock_kernels >= 1 && c.n_resblock_kernels <= ST2_MAX_RESK, "n_resblock_kernels");
  for (int i = 0; i < c.n_ups; ++i) {
    const int u = c.upsample_rates[i], k = c.upsample_kernel_sizes[i];
    ST2_REQUIRE(u >= 1 && k >= u, "upsample rate / kernel");
    if (c.kind == ST2_KIND_ISTFTNET) ST2_REQUIRE((k - u) % 2 == 0, "istftnet needs (kernel - rate) even so that Lout = rate * Lin");
    else ST2_REQUIRE(k == 2 * u, "hifigan upsample kernels must be 2 * rate");
  }
  for (int j = 0; j < c.n_resblock_kernels; ++j) {
    ST2_REQUIRE(c.resblock_kernel_sizes[j] % 2 == 1, "resblock kernel sizes must be odd");
    for (int d = 0; d < 3; ++d) ST2_REQUIRE(c.resblock_dilations[j][d] >= 1, "dilation");
  }
  if (c.kind == ST2_KIND_ISTFTNET)
    ST2_REQUIRE(c.gen_istft_n_fft == 20 && c.gen_istft_hop_size == 5,
                "only the shipped n_fft 20 / hop 5 inverse STFT is implemented (Configs/config.yml:59-60)");
  return ST2_OK;
}

extern "C" int st2_decoder_create(const St2Config* cfg, St2Decoder** out) {
  if (!cfg || !out) {
    set_error("st2_decoder_create: null argument");
    return ST2_EINVAL;
  }
  int rc = validate_cfg(*cfg);
  if (rc) return rc;
  St2Decoder* h = new St2Decoder();
  h->cfg = *cfg;
  if (const char* e = getenv("ST2_FORK")) h->opt_fork = std::min(2, std::max(0, atoi(e)));
  if (const char* e = getenv("ST2_GROUP")) h->opt_group = e[0] == '0' ? 0 : 1;
  h->spf = 1;
  for (int i = 0; i < cfg->n_ups; ++i) h->spf *= cfg->upsample_rates[i];
  if (cfg->kind == ST2_KIND_ISTFTNET) h->spf *= cfg->gen_istft_hop_size;
  build_keys(h);
  if (cudaGetDevice(&h->device) != cudaSuccess) h->device = -1;  // reported at load()
  (void)cudaGetLastError();
  *out = h;
  return ST2_OK;
}

extern "C" int st2_decoder_num_keys(const St2Decoder* h) { return h ? (int)h->key_names.size() : 0; }

extern "C" int st2_decoder_key(const St2Decoder* h, int i, const char** name, int* ndim, int64_t shape[4]) {
  if (!h || i < 0 || i >= (int)h->key_names.size()) {
    set_error("st2_decoder_key: index");
    return ST2_EINVAL;
  }
  if (name) *name = h->key_names[i].c_str();
  if (ndim) *ndim = (int)h->key_shapes[i].size();
  if (shape)
    for (size_t d = 0; d < 4; ++d) shape[d] = d < h->key_shapes[i].size() ? h->key_shapes[i][d] : 1;
  return ST2_OK;
}

extern "C" int st2_decoder_samples_per_frame(const St2Decoder* h) { return h ? h->spf : 0; }

extern "C" int st2_decoder_set_option(St2Decoder* h, const char* name, int value) {
  if (!h || !name) {
    set_error("st2_decoder_set_option: null argument");
    return ST2_EINVAL;
  }
  if (strcmp(name, "fp32_tc") == 0) {
    if (h->loaded && (value != 0) != h->opt_fp32_tc) {
      set_error("st2_decoder_set_option: fp32_tc must be set before st2_decoder_load (it adds weight blobs)");
      return ST2_ESTATE;
    }
    h->opt_fp32_tc = value != 0;
    return ST2_OK;
  }
  if (strcmp(name, "fork") == 0 || strcmp(name, "group") == 0) {
    if (value < 0 || value > 2) {
      set_error("st2_decoder_set_option: fork / group take 0, 1 or 2");
      return ST2_EINVAL;
    }
    (name[0] == 'f' ? h->opt_fork : h->opt_group) = value;
    return ST2_OK;
  }
  set_error(std::string("st2_decoder_set_option: unknown option ") + name);
  return ST2_EINVAL;
}

extern "C" int st2_decoder_load(St2Decoder* h, const St2Tensor* tensors, int n) {
  if (!h || !tensors || n <= 0) {
    set_error("st2_decoder_load: null argument");
    return ST2_EINVAL;
  }
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
    (void)cudaGetLastError();
    set_error("st2_decoder_load: no CUDA device (this library has no CPU fallback)");
    return ST2_ECUDA;
  }
  if (h->device < 0) ST2_CUDA_OK(cudaGetDevice(&h->device));
  cudaDeviceProp prop;
  ST2_CUDA_OK(cudaGetDeviceProperties(&prop, h->device));
  if (prop.major != 10) {
    set_error("st2_decoder_load: device is sm_" + std::to_string(prop.major) + std::to_string(prop.minor) +
              ", this library is built for sm_100a only");
    return ST2_ECUDA;
  }
  // one-time per-device setup of the kernels (identity blobs, constant tables, shared-memory attributes, the
  // tensor-map encoder) happens here, so that forward neither allocates nor synchronises the host
  {
    int cur = 0;
    ST2_CUDA_OK(cudaGetDevice(&cur));
    if (cur != h->device) ST2_CUDA_OK(cudaSetDevice(h->device));
    const bool dec = h->cfg.kind == ST2_KIND_ISTFTNET || h->cfg.kind == ST2_KIND_HIFIGAN;
    int rc = umma_prepare(dec ? h->cfg.n_resblock_kernels : 0);
    if (!rc) rc = source_prepare();
    if (!rc && dec && !h->side_st) {
      bool ok = cudaStreamCreateWithFlags(&h->side_st, cudaStreamNonBlocking) == cudaSuccess &&
                cudaEventCreateWithFlags(&h->ev_fork, cudaEventDisableTiming) == cudaSuccess;
      for (int i = 0; i < ST2_MAX_UPS && ok; ++i)
        ok = cudaEventCreateWithFlags(&h->ev_join[i], cudaEventDisableTiming) == cudaSuccess;
      if (!ok) {
        set_error(std::string("st2_decoder_load: side stream / events: ") + cudaGetErrorString(cudaGetLastError()));
        rc = ST2_ECUDA;
      }
    }
    if (cur != h->device) ST2_CUDA_OK(cudaSetDevice(cur));
    if (rc) return rc;
  }
  return do_load(h, tensors, n);
}

template <typename AT>
static int forward_t(St2Decoder* h, const float* asr, const float* f0, const float* n, const float* s,
                     float* wav, void* ws, int64_t ws_bytes, int B, int T, const St2ForwardOpts& o,
                     cudaStream_t st, bool dry, int64_t* peak, int64_t* side_bytes) {
  // workspace = [side arena (*side_bytes, from the dry run) | main arena]
  Fwd<AT> f;
  f.h = h; f.st = st; f.B = B; f.T = T; f.dry = dry;
  f.tc32 = o.precision == ST2_PREC_FP32_TC;
  const int64_t sb = dry ? 0 : *side_bytes;
  f.side.base = reinterpret_cast<uint8_t*>(ws); f.side.cap = sb; f.side.dry = dry;
  f.ar.base = reinterpret_cast<uint8_t*>(ws) + sb; f.ar.cap = ws_bytes - sb; f.ar.dry = dry;
  int rc = f.run(asr, f0, n, s, wav, o);
  if (peak) *peak = f.ar.peak;
  if (dry) *side_bytes = round_up(f.side.peak, 256);
  return rc;
}

static int plan_bytes(St2Decoder* h, int B, int T, const St2ForwardOpts& o, int64_t* bytes, int64_t* side_bytes = nullptr) {
  int64_t peak = 0, side = 0;
  int rc;
  if (o.precision == ST2_PREC_BF16)
    rc = forward_t<__nv_bfloat16>(h, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, 0, B, T, o, 0, true, &peak, &side);
  else
    rc = forward_t<float>(h, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, 0, B, T, o, 0, true, &peak, &side);
  *bytes = side + round_up(peak, 256) + 256;
  if (side_bytes) *side_bytes = side;
  return rc;
}

extern "C" int64_t st2_decoder_workspace_bytes(const St2Decoder* hc, int B, int T, int precision) {
  St2Decoder* h = const_cast<St2Decoder*>(hc);
  if (!h || !h->loaded || (h->cfg.kind != ST2_KIND_ISTFTNET && h->cfg.kind != ST2_KIND_HIFIGAN) || B < 1 || T < 2 ||
      (T & 1) || (precision != ST2_PREC_FP32 && precision != ST2_PREC_BF16 && precision != ST2_PREC_FP32_TC) ||
      (precision == ST2_PREC_FP32_TC && !h->opt_fp32_tc)) {
    set_error("st2_decoder_workspace_bytes: bad argument, weights not loaded, not a decoder handle, or ST2_PREC_FP32_TC "
              "without st2_decoder_set_option(\"fp32_tc\", 1) before load");
    return ST2_EINVAL;
  }
  St2ForwardOpts o;
  memset(&o, 0, sizeof(o));
  o.precision = precision;  // no injected excitation = the largest plan
  int64_t bytes = 0;
  int rc = plan_bytes(h, B, T, o, &bytes);
  return rc ? rc : bytes;
}

extern "C" int st2_decoder_forward(St2Decoder* h, const float* asr, const float* f0, const float* n,
                                   const float* s, float* wav_out, void* workspace, int64_t workspace_bytes,
                                   int B, int T, const St2ForwardOpts* opts, void* stream) {
  if (!h || !asr || !f0 || !n || !s || !wav_out || !workspace) {
    set_error("st2_decoder_forward: null argument");
    return ST2_EINVAL;
  }
  if (!h->loaded) {
    set_error("st2_decoder_forward: weights not loaded");
    return ST2_ESTATE;
  }
  ST2_REQUIRE(h->cfg.kind == ST2_KIND_ISTFTNET || h->cfg.kind == ST2_KIND_HIFIGAN,
              "st2_decoder_forward: this handle is not a decoder (F0/N predictor: st2_f0n_forward, sampler: st2_sampler_sample)");
  ST2_REQUIRE(B >= 1 && T >= 2, "batch / length");
  ST2_REQUIRE((T & 1) == 0, "len(F0_curve) must equal 2 * asr.shape[-1] (istftnet.py:513)");
  ST2_REQUIRE(B <= 65535, "batch too large for one call");
  St2ForwardOpts o;
  memset(&o, 0, sizeof(o));
  if (opts) o = *opts;
  ST2_REQUIRE(o.precision == ST2_PREC_FP32 || o.precision == ST2_PREC_BF16 || o.precision == ST2_PREC_FP32_TC, "precision");
  ST2_REQUIRE(o.precision != ST2_PREC_FP32_TC || h->opt_fp32_tc,
              "ST2_PREC_FP32_TC needs st2_decoder_set_option(h, \"fp32_tc\", 1) before st2_decoder_load");
  ST2_REQUIRE(!(o.har && h->cfg.kind != ST2_KIND_ISTFTNET), "har injection is istftnet-only");
  ST2_REQUIRE(!o.align_idx || o.align_len >= 1, "align_len must be the token count of t_en");
  ST2_REQUIRE(((uintptr_t)workspace & 255) == 0, "workspace must be 256-byte aligned");
  int dev = -1;
  ST2_CUDA_OK(cudaGetDevice(&dev));
  ST2_REQUIRE(dev == h->device, "handle belongs to another device");
  int64_t need = 0, side_bytes = 0;
  int rc = plan_bytes(h, B, T, o, &need, &side_bytes);
  if (rc) return rc;
  if (need > workspace_bytes) {
    set_error("st2_decoder_forward: workspace too small (" + std::to_string(workspace_bytes) + " < " +
              std::to_string(need) + ")");
    return ST2_ENOMEM;
  }
  const int64_t before = launch_counter();
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (o.precision == ST2_PREC_BF16)
    rc = forward_t<__nv_bfloat16>(h, asr, f0, n, s, wav_out, workspace, workspace_bytes, B, T, o, st, false, nullptr,
                                  &side_bytes);
  else
    rc = forward_t<float>(h, asr, f0, n, s, wav_out, workspace, workspace_bytes, B, T, o, st, false, nullptr, &side_bytes);
  h->last_launches = launch_counter() - before;
  return rc;
}

template <typename AT>
static int f0n_t(St2Decoder* h, const float* x, const float* s, float* f0, float* n, void* ws, int64_t ws_bytes,
                 int B, int T, cudaStream_t st, bool dry, int64_t* peak, bool tc32 = false) {
  Fwd<AT> f;
  f.h = h; f.st = st; f.B = B; f.T = T; f.dry = dry;
  f.tc32 = tc32;
  f.ar.base = reinterpret_cast<uint8_t*>(ws); f.ar.cap = ws_bytes; f.ar.dry = dry;
  int rc = f.run_f0n(x, s, f0, n);
  if (peak) *peak = f.ar.peak;
  return rc;
}
static int f0n_plan(St2Decoder* h, int B, int T, int precision, int64_t* bytes) {
  int64_t peak = 0;
  int rc = precision == ST2_PREC_BF16
               ? f0n_t<__nv_bfloat16>(h, nullptr, nullptr, nullptr, nullptr, nullptr, 0, B, T, 0, true, &peak)
               : f0n_t<float>(h, nullptr, nullptr, nullptr, nullptr, nullptr, 0, B, T, 0, true, &peak,
                              precision == ST2_PREC_FP32_TC);
  *bytes = round_up(peak, 256) + 256;
  return rc;
}
extern "C" int64_t st2_f0n_workspace_bytes(const St2Decoder* hc, int B, int T, int precision) {
  St2Decoder* h = const_cast<St2Decoder*>(hc);
  if (!h || !h->loaded || h->cfg.kind != ST2_KIND_F0N || B < 1 || T < 1 ||
      (precision != ST2_PREC_FP32 && precision != ST2_PREC_BF16 && precision != ST2_PREC_FP32_TC) ||
      (precision == ST2_PREC_FP32_TC && !h->opt_fp32_tc)) {
    set_error("st2_f0n_workspace_bytes: bad argument, wrong handle kind or weights not loaded");
    return ST2_EINVAL;
  }
  int64_t bytes = 0;
  int rc = f0n_plan(h, B, T, precision, &bytes);
  return rc ? rc : bytes;
}
extern "C" int st2_f0n_forward(St2Decoder* h, const float* x, const float* s, float* f0_out, float* n_out,
                               void* workspace, int64_t workspace_bytes, int B, int T, int precision,
                               void* stream) {
  if (!h || !x || !s || !f0_out || !n_out || !workspace) {
    set_error("st2_f0n_forward: null argument");
    return ST2_EINVAL;
  }
  ST2_REQUIRE(h->cfg.kind == ST2_KIND_F0N, "st2_f0n_forward: handle was not created with ST2_KIND_F0N");
  if (!h->loaded) {
    set_error("st2_f0n_forward: weights not loaded");
    return ST2_ESTATE;
  }
  ST2_REQUIRE(B >= 1 && B <= 65535 && T >= 1, "batch / length");
  ST2_REQUIRE(precision == ST2_PREC_FP32 || precision == ST2_PREC_BF16 || precision == ST2_PREC_FP32_TC, "precision");
  ST2_REQUIRE(precision != ST2_PREC_FP32_TC || h->opt_fp32_tc,
              "ST2_PREC_FP32_TC needs st2_decoder_set_option(h, \"fp32_tc\", 1) before st2_decoder_load");
  ST2_REQUIRE(((uintptr_t)workspace & 255) == 0, "workspace must be 256-byte aligned");
  int dev = -1;
  ST2_CUDA_OK(cudaGetDevice(&dev));
  ST2_REQUIRE(dev == h->device, "handle belongs to another device");
  int64_t need = 0;
  int rc = f0n_plan(h, B, T, precision, &need);
  if (rc) return rc;
  if (need > workspace_bytes) {
    set_error("st2_f0n_forward: workspace too small (" + std::to_string(workspace_bytes) + " < " + std::to_string(need) + ")");
    return ST2_ENOMEM;
  }
  const int64_t before = launch_counter();
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  rc = precision == ST2_PREC_BF16
           ? f0n_t<__nv_bfloat16>(h, x, s, f0_out, n_out, workspace, workspace_bytes, B, T, st, false, nullptr)
           : f0n_t<float>(h, x, s, f0_out, n_out, workspace, workspace_bytes, B, T, st, false, nullptr,
                          precision == ST2_PREC_FP32_TC);
  h->last_launches = launch_counter() - before;
  return rc;
}

// ---------------------------------------------------------------------------------------------
// Style sampler.  DiffusionSampler.forward (Modules/diffusion/sampler.py:573-585) =
// KarrasSchedule (:318-336) + ADPM2Sampler.forward / .step (:496-519) around KDiffusion.denoise_fn
// (:193-208) around StyleTransformer1d.forward / .run (modules.py:146-186; Transformer1d :382-426).
// Token rows of all items (and of the guidance run on the fixed embedding) are flattened into one
// [Bx * n][F] matrix, so every nn.Linear over tokens is one 1x1 conv (tcgen05 GEMM on the bf16 path).
// ---------------------------------------------------------------------------------------------
struct SamplerSigmas {
  std::vector<float> sig;
};
// KarrasSchedule.forward in fp32 like the reference's tensor arithmetic
static void karras_sigmas(int num_steps, float smin, float smax, float rho, std::vector<float>& out) {
  const double rho_inv = 1.0 / (double)rho;
  const float a = (float)pow((double)smax, rho_inv), d = (float)(pow((double)smin, rho_inv) - pow((double)smax, rho_inv));
  out.resize(num_steps + 1);
  for (int i = 0; i < num_steps; ++i) {
    const float t = (float)i / (float)(num_steps - 1);
    out[i] = powf(a + t * d, rho);
  }
  out[num_steps] = 0.f;
}

template <typename AT>
static int sampler_t(St2Decoder* h, const float* x_in, const float* emb, const float* feat, float* out, void* ws,
                     int64_t ws_bytes, int B, int n, const St2SamplerOpts& o, bool single, float single_sigma,
                     cudaStream_t st, bool dry, int64_t* peak) {
  Fwd<AT> f;
  f.h = h; f.st = st; f.B = 1; f.T = n; f.dry = dry;
  f.ar.base = reinterpret_cast<uint8_t*>(ws); f.ar.cap = ws_bytes; f.ar.dry = dry;
  Arena& ar = f.ar;
  const St2Config& c = h->cfg;
  const int C = c.dim_in, S = c.style_dim, E = c.dim_out, F = C + E, layers = c.n_resblock_kernels;
  const int H = c.resblock_kernel_sizes[0], D = c.resblock_kernel_sizes[1], mid = H * D, mult = c.resblock_kernel_sizes[2];
  const bool guided = o.embedding_scale != 1.f;
  const int Bx = guided ? 2 * B : B;
  const int64_t rows = (int64_t)Bx * n;
  auto Wf = [&](int64_t off) { return f.template W<float>(off); };

  float* hall = S ? ar.alloc<float>((int64_t)B * layers * 4 * F) : nullptr;
  float* feat_act = S ? ar.alloc<float>((int64_t)B * F) : nullptr;
  float* four = ar.alloc<float>(C + 1);
  float* t_act = ar.alloc<float>(F);
  float* m1 = ar.alloc<float>((int64_t)B * F);
  float* mapping = ar.alloc<float>((int64_t)B * F);
  float* X = ar.alloc<float>(rows * F);
  AT* nq = ar.alloc<AT>(rows * F);
  AT* nk = ar.alloc<AT>(rows * F);
  AT* q = ar.alloc<AT>(rows * mid);
  AT* kv = ar.alloc<AT>(rows * 2 * mid);
  AT* att = ar.alloc<AT>(rows * mid);
  AT* delta = ar.alloc<AT>(rows * F);
  AT* xc = ar.alloc<AT>(rows * F);
  AT* hf = ar.alloc<AT>(rows * F * mult);
  float* xm = ar.alloc<float>((int64_t)Bx * F);
  float* pred = ar.alloc<float>((int64_t)Bx * C);
  float* x = ar.alloc<float>((int64_t)B * C);
  float* x_mid = ar.alloc<float>((int64_t)B * C);
  if (peak) *peak = ar.peak;
  if (dry) return ST2_OK;
#define SRUN(x)          \
  do {                   \
    int _rc = (x);       \
    if (_rc) return _rc; \
  } while (0)

  auto gemm = [&](const ConvW& w, const AT* src, int ldx, AT* dst, int ldy) -> int {
    typename Fwd<AT>::ConvIO io;
    io.x = src; io.x_dtype = DT<AT>::id; io.x_bstride = 0; io.ldx = ldx; io.Lin = (int)rows;
    io.y = dst; io.y_dtype = DT<AT>::id; io.y_bstride = 0; io.ldy = ldy; io.Lout = (int)rows;
    return f.conv(w, io);
  };
  // features are constant over the sampling run: AdaLayerNorm fc of all blocks and to_features once
  if (S) {
    SRUN(launch_small_linear(feat, S, nullptr, Wf(h->s_fcW), Wf(h->s_fcB), hall, layers * 4 * F, B, S, layers * 4 * F, 0, st));
    SRUN(launch_small_linear(feat, S, nullptr, Wf(h->s_featW), Wf(h->s_featB), feat_act, F, B, S, F, 1, st));
  }
  const float sd = o.sigma_data;
  float c_skip = 0.f, c_out = 0.f;
  // net(c_in * xin, c_noise, ...) -> pred [Bx][C]; sets c_skip / c_out of this sigma (sampler.py:184-191)
  auto eval = [&](const float* xin, float sigma) -> int {
    const float s2 = sigma * sigma + sd * sd;
    const float c_in = powf(s2, -0.5f), c_noise = logf(sigma) * 0.25f;
    c_skip = (sd * sd) / s2;
    c_out = sigma * sd * powf(s2, -0.5f);
    // mapping = to_mapping(to_time(c_noise) + to_features(features))  (modules.py:121-144)
    SRUN(launch_time_fourier(c_noise, Wf(h->s_four), C / 2, four, st));
    SRUN(launch_small_linear(four, C + 1, nullptr, Wf(h->s_timeW), Wf(h->s_timeB), t_act, F, 1, C + 1, F, 1, st));
    const int Rm = S ? B : 1;
    if (S) SRUN(launch_small_linear(feat_act, F, t_act, Wf(h->s_mapW[0]), Wf(h->s_mapB[0]), m1, F, B, F, F, 1, st));
    else SRUN(launch_small_linear(t_act, F, nullptr, Wf(h->s_mapW[0]), Wf(h->s_mapB[0]), m1, F, 1, F, F, 1, st));
    SRUN(launch_small_linear(m1, F, nullptr, Wf(h->s_mapW[1]), Wf(h->s_mapB[1]), mapping, F, Rm, F, F, 1, st));
    SRUN(launch_build_rows(xin, c_in, emb, Wf(h->s_fixed), X, B, Bx, n, C, E, st));
    for (int l = 0; l < layers; ++l) {
      const St2Decoder::SamplerBlk& k = h->sblk[l];
      const float* gb = S ? hall + (int64_t)l * 4 * F : Wf(h->s_ln) + (int64_t)l * 4 * F;
      SRUN(launch_adaln2<AT>(X, l ? delta : nullptr, mapping, S ? F : 0, gb, S ? (int64_t)layers * 4 * F : 0, nq, nk, rows,
                             n, B, F, st));
      SRUN(gemm(k.to_q, nq, F, q, mid));
      SRUN(gemm(k.to_kv, nk, F, kv, 2 * mid));
      SRUN(launch_attention<AT>(q, mid, kv, 2 * mid, att, mid, Bx, n, H, D, st));
      SRUN(gemm(k.to_out, att, mid, delta, F));
      SRUN(launch_add_cast<AT>(X, delta, xc, rows * F, st));
      SRUN(gemm(k.ff1, xc, F, hf, F * mult));
      SRUN(launch_gelu<AT>(hf, rows * F * mult, st));
      SRUN(gemm(k.ff2, hf, F * mult, delta, F));
    }
    SRUN(launch_token_mean<AT>(X, delta, xm, Bx, n, F, st));
    SRUN(launch_small_linear(xm, F, nullptr, Wf(h->s_outW), Wf(h->s_outB), pred, C, Bx, F, C, 0, st));
    return ST2_OK;
  };
  if (single) {
    SRUN(eval(x_in, single_sigma));
    return launch_sampler_step(pred, x_in, nullptr, nullptr, 0, 0, 0, nullptr, out, B, C, guided, o.embedding_scale, c_skip,
                               c_out, single_sigma, 0.f, 0.f, 0, st);
  }
  std::vector<float> sig;
  karras_sigmas(o.num_steps, o.sigma_min, o.sigma_max, o.rho, sig);
  SRUN(launch_scale(x_in, sig[0], x, B * C, st));  // x = sigmas[0] * noise (sampler.py:499)
  const int nst = o.num_steps - 1;
  for (int i = 0; i < nst; ++i) {
    // ADPM2Sampler.get_sigmas (sampler.py:489-494): fp32 tensor arithmetic, math.sqrt in double
    const float s0 = sig[i], s1 = sig[i + 1];
    const double up = sqrt((double)(s1 * s1 * (s0 * s0 - s1 * s1) / (s0 * s0)));
    const double down = sqrt((double)(float)((double)(s1 * s1) - up * up));
    const float r = o.sampler_rho;
    const float smid = r == 1.f ? (s0 + (float)down) / 2.f : powf((powf(s0, 1.f / r) + (float)pow(down, 1.0 / r)) / 2.f, r);
    const bool last = i == nst - 1;
    SRUN(eval(x, s0));
    SRUN(launch_sampler_step(pred, x, x, nullptr, 0, 0, 0, x_mid, nullptr, B, C, guided, o.embedding_scale, c_skip, c_out, s0,
                             smid - s0, 0.f, 0, st));
    SRUN(eval(x_mid, smid));
    SRUN(launch_sampler_step(pred, x_mid, x, o.step_noise ? o.step_noise + (int64_t)i * B * C : nullptr, o.seed, (uint32_t)i,
                             o.item_offset, last ? out : x, nullptr, B, C, guided, o.embedding_scale, c_skip, c_out, smid,
                             (float)down - s0, (float)up, last && o.clamp, st));
  }
#undef SRUN
  return ST2_OK;
}

static int sampler_check(St2Decoder* h, int B, int n, int precision, const char* who) {
  if (!h) {
    set_error(std::string(who) + ": null handle");
    return ST2_EINVAL;
  }
  ST2_REQUIRE(h->cfg.kind == ST2_KIND_SAMPLER, std::string(who) + ": handle was not created with ST2_KIND_SAMPLER");
  if (!h->loaded) {
    set_error(std::string(who) + ": weights not loaded");
    return ST2_ESTATE;
  }
  ST2_REQUIRE(B >= 1 && B <= 32767, "batch");
  ST2_REQUIRE(n >= 1 && n <= h->cfg.resblock_kernel_sizes[3],
              "n_tokens must be in [1, embedding_max_length] (FixedEmbedding, modules.py:654-690)");
  ST2_REQUIRE((int64_t)2 * B * n < (int64_t)1 << 30, "too many token rows");
  ST2_REQUIRE(precision == ST2_PREC_FP32 || precision == ST2_PREC_BF16, "precision");
  return ST2_OK;
}
static int sampler_plan(St2Decoder* h, int B, int n, int precision, bool guided, int64_t* bytes) {
  St2SamplerOpts o;
  memset(&o, 0, sizeof(o));
  o.embedding_scale = guided ? 2.f : 1.f;
  int64_t peak = 0;
  int rc = precision == ST2_PREC_BF16
               ? sampler_t<__nv_bfloat16>(h, nullptr, nullptr, nullptr, nullptr, nullptr, 0, B, n, o, true, 1.f, 0, true, &peak)
               : sampler_t<float>(h, nullptr, nullptr, nullptr, nullptr, nullptr, 0, B, n, o, true, 1.f, 0, true, &peak);
  *bytes = round_up(peak, 256) + 256;
  return rc;
}
extern "C" int64_t st2_sampler_workspace_bytes(const St2Decoder* hc, int B, int n_tokens, int precision, int guided) {
  St2Decoder* h = const_cast<St2Decoder*>(hc);
  int rc = sampler_check(h, B, n_tokens, precision, "st2_sampler_workspace_bytes");
  if (rc) return rc;
  int64_t bytes = 0;
  rc = sampler_plan(h, B, n_tokens, precision, guided != 0, &bytes);
  return rc ? rc : bytes;
}
static int sampler_call(St2Decoder* h, const float* x_in, const float* emb, const float* feat, float* out, void* ws,
                        int64_t ws_bytes, int B, int n, const St2SamplerOpts* opts, bool single, float sigma, void* stream,
                        const char* who) {
  if (!opts || !x_in || !emb || !out || !ws) {
    set_error(std::string(who) + ": null argument");
    return ST2_EINVAL;
  }
  int rc = sampler_check(h, B, n, opts->precision, who);
  if (rc) return rc;
  ST2_REQUIRE((h->cfg.style_dim > 0) == (feat != nullptr),
              "features are required by StyleTransformer1d and not accepted by Transformer1d");
  ST2_REQUIRE(opts->sigma_data > 0.f, "sigma_data");
  if (single) ST2_REQUIRE(sigma > 0.f, "sigma");
  else ST2_REQUIRE(opts->num_steps >= 2 && opts->sigma_min > 0.f && opts->sigma_max > opts->sigma_min && opts->rho > 0.f &&
                   opts->sampler_rho > 0.f, "num_steps >= 2, 0 < sigma_min < sigma_max, rho > 0");
  ST2_REQUIRE(((uintptr_t)ws & 255) == 0, "workspace must be 256-byte aligned");
  int dev = -1;
  ST2_CUDA_OK(cudaGetDevice(&dev));
  ST2_REQUIRE(dev == h->device, "handle belongs to another device");
  int64_t need = 0;
  rc = sampler_plan(h, B, n, opts->precision, opts->embedding_scale != 1.f, &need);
  if (rc) return rc;
  if (need > ws_bytes) {
    set_error(std::string(who) + ": workspace too small (" + std::to_string(ws_bytes) + " < " + std::to_string(need) + ")");
    return ST2_ENOMEM;
  }
  const int64_t before = launch_counter();
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  rc = opts->precision == ST2_PREC_BF16
           ? sampler_t<__nv_bfloat16>(h, x_in, emb, feat, out, ws, ws_bytes, B, n, *opts, single, sigma, st, false, nullptr)
           : sampler_t<float>(h, x_in, emb, feat, out, ws, ws_bytes, B, n, *opts, single, sigma, st, false, nullptr);
  h->last_launches = launch_counter() - before;
  return rc;
}
extern "C" int st2_sampler_sample(St2Decoder* h, const float* noise, const float* embedding, const float* features,
                                  float* out, void* workspace, int64_t workspace_bytes, int B, int n_tokens,
                                  const St2SamplerOpts* opts, void* stream) {
  return sampler_call(h, noise, embedding, features, out, workspace, workspace_bytes, B, n_tokens, opts, false, 0.f, stream,
                      "st2_sampler_sample");
}
extern "C" int st2_sampler_denoise(St2Decoder* h, const float* x_noisy, float sigma, const float* embedding,
                                   const float* features, float* out, void* workspace, int64_t workspace_bytes, int B,
                                   int n_tokens, const St2SamplerOpts* opts, void* stream) {
  return sampler_call(h, x_noisy, embedding, features, out, workspace, workspace_bytes, B, n_tokens, opts, true, sigma, stream,
                      "st2_sampler_denoise");
}

extern "C" int st2_wave_to_pcm16(const float* wav, int B, int64_t L, int64_t trim_tail, int16_t* out, void* stream) {
  ST2_REQUIRE(wav && out && B >= 1 && B <= 65535 && trim_tail >= 0 && L > trim_tail, "st2_wave_to_pcm16: bad argument");
  return launch_wave_to_pcm16(wav, B, L, trim_tail, out, reinterpret_cast<cudaStream_t>(stream));
}
extern "C" int st2_alignment_indices(const int32_t* dur, int B, int Ltext, int frames, int shift_one, int32_t* idx,
                                     int32_t* totals, void* stream) {
  ST2_REQUIRE(dur && idx && B >= 1 && B <= 65535 && Ltext >= 1 && frames >= 1, "st2_alignment_indices: bad argument");
  return launch_alignment_idx(dur, B, Ltext, frames, shift_one, idx, totals, reinterpret_cast<cudaStream_t>(stream));
}
extern "C" int st2_expand_alignment(const float* x, const int32_t* idx, int B, int C, int Ltext, int frames, float* out,
                                    void* stream) {
  ST2_REQUIRE(x && idx && out && B >= 1 && B <= 65535 && C >= 1 && C <= 65535 && Ltext >= 1 && frames >= 1,
              "st2_expand_alignment: bad argument");
  return launch_expand_alignment(x, idx, B, C, Ltext, frames, out, reinterpret_cast<cudaStream_t>(stream));
}

extern "C" int64_t st2_decoder_last_launches(const St2Decoder* h) { return h ? h->last_launches : 0; }

extern "C" void st2_decoder_destroy(St2Decoder* h) {
  if (!h) return;
  if (h->dev) {
    cudaFree(h->dev);
  }
  if (h->side_st) cudaStreamDestroy(h->side_st);
  if (h->ev_fork) cudaEventDestroy(h->ev_fork);
  for (cudaEvent_t e : h->ev_join)
    if (e) cudaEventDestroy(e);
  delete h;
}

extern "C" const char* st2_last_error(void) { return g_err.c_str(); }
extern "C" const char* st2_version(void) { return "st2dec 0.1 (sm_100a)"; }

// ---- single-kernel entry points
extern "C" int st2_conv_simt(const St2ConvParams* p, void* stream) {
  if (!p) return ST2_EINVAL;
  return launch_conv_simt(*p, reinterpret_cast<cudaStream_t>(stream));
}
extern "C" int st2_conv_umma_multi(const St2ConvParams* ps, int n, void* stream) {
  if (!ps) return ST2_EINVAL;
  return launch_conv_umma_multi(ps, n, reinterpret_cast<cudaStream_t>(stream));
}
namespace st2 { int64_t conv_simt_dense_launches(bool reset); }
extern "C" int64_t st2_conv_simt_dense_launches(int reset) { return st2::conv_simt_dense_launches(reset != 0); }
extern "C" int st2_conv_umma(const St2ConvParams* p, void* stream) {
  if (!p) return ST2_EINVAL;
  return launch_conv_umma(*p, reinterpret_cast<cudaStream_t>(stream));
}
extern "C" int st2_conv_tile_rows(int use_umma) { return conv_tile_rows(use_umma != 0); }

extern "C" int st2_source_har(const float* f0, const float* noise, uint64_t seed, const float* lin_w9,
                              float lin_b, int B, int T, int spf, float* phase_scratch, float* har_source,
                              void* stream) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  int rc = launch_phase_scan(f0, B, T, spf, phase_scratch, st);
  if (rc) return rc;
  return launch_har_source(f0, phase_scratch, noise, seed, lin_w9, lin_b, B, T, spf, har_source, st);
}
extern "C" int st2_stft_har(const float* har_source, int B, int Lw, float* har_cl32, void* stream) {
  return launch_stft(har_source, B, Lw, har_cl32, 32, reinterpret_cast<cudaStream_t>(stream));
}
extern "C" int st2_stft_har16(const float* har_source, int B, int Lw, int rows_pad, uint16_t* har_hilo, void* stream) {
  return launch_stft_hilo(har_source, B, Lw, rows_pad, reinterpret_cast<__nv_bfloat16*>(har_hilo),
                          reinterpret_cast<cudaStream_t>(stream));
}
extern "C" int st2_istft_head(const float* x_cl32, int B, int frames, float* wav, void* stream) {
  return launch_istft(x_cl32, 32, B, frames, wav, reinterpret_cast<cudaStream_t>(stream));
}

// ---- host-only test hook: fold weight-norm and repack one conv exactly as load() does.
// transposed = 0: Conv1d prefix (weight_g/weight_v/bias);  1: ConvTranspose1d (K, stride, padding).
// Writes the fp32 [k][cin][n_cols] weights and [n_cols] bias; returns k / pad / cin / n_cols.
extern "C" int st2_debug_pack_conv(const St2Tensor* tensors, int n, const char* prefix, int transposed,
                                   int K, int stride, int padding, float* w_out, int64_t w_cap,
                                   float* bias_out, int* k_out, int* pad_out, int* cin_out, int* ncols_out) {
  Loader L;
  for (int i = 0; i < n; ++i) L.by_name[tensors[i].name] = &tensors[i];
  ConvW c;
  UpW u;
  bool ok;
  if (transposed) {
    ok = L.conv_transpose(u, prefix, K, stride, padding);
    c = u.conv;
  } else {
    ok = L.conv1d(c, prefix, true, true, K, 1, padding, stride, false);
  }
  ST2_REQUIRE(ok, L.err);
  const int64_t elems = (int64_t)c.k * c.cin * c.n;
  ST2_REQUIRE(elems <= w_cap, "output buffer too small");
  memcpy(w_out, L.blob.data.data() + c.w32, elems * 4);
  if (bias_out && c.bias >= 0) memcpy(bias_out, L.blob.data.data() + c.bias, (size_t)c.n * 4);
  *k_out = c.k; *pad_out = c.pad; *cin_out = c.cin; *ncols_out = c.n;
  return ST2_OK;
}
