// conv_params.h — argument block of the fused 1-D convolution kernels (SIMT and tcgen05).
//
// Activations are channels-last: element (b, t, c) of a tensor lives at
//   base + b * bstride + t * ld + c        (all in elements of the tensor's dtype)
// so one conv tap is a row offset and the GEMM K dimension (channels) is contiguous.
//
//   y[b,t,n] = post( bias[b*bias_bstride + n] + sum_{j<k} sum_{c<cin} W[j][c][n] * pre(x[b, t*stride + j*dil - pad, c]) )
//   pre(v)   = act( v * scale[b,c] + shift[b,c] ),  0 for rows outside [0, Lin)   (conv zero padding
//              applies AFTER the activation, exactly like AdaIN -> Snake -> Conv1d(padding=..))
//   post(a)  = (a + res[b, t >> res_shift, n]) * out_scale (+ y_old[b,t,n] if accumulate)
//   stats    : per-(b, time-tile, n) partial sum and sum of squares of the stored value, fp32
//
// Reference semantics covered by one call:
//   AdaIN1d + Snake + Conv1d(dilated)   Modules/istftnet.py:68-73
//   AdaIN1d + LeakyReLU(0.2) + Conv1d   Modules/istftnet.py:441-447
//   residual add / (.)/sqrt(2) / mean over resblocks      :74, :452, :371-374
//   LeakyReLU(0.1) + ConvTranspose1d (polyphase, as a k=3 conv with N = stride*Cout) :359-363
#ifndef ST2_CONV_PARAMS_H
#define ST2_CONV_PARAMS_H
#include <stdint.h>

#define ST2_DT_F32 0
#define ST2_DT_BF16 1

#define ST2_ACT_NONE 0
#define ST2_ACT_LRELU 1
#define ST2_ACT_SNAKE 2

typedef struct St2ConvParams {
  int B;
  // ---- input
  const void* x;
  int x_dtype;        // ST2_DT_*
  int64_t x_bstride;  // elements between batch items
  int ldx;            // elements between time rows
  int Lin;            // rows per item
  int cin;            // channels the contraction runs over (SIMT: real count; UMMA: multiple of 64)
  // ---- prologue
  const float* ss;    // [B][ss_stride] interleaved (scale, shift) pairs, or NULL for identity
  int ss_stride;      // pairs per batch row
  int act;            // ST2_ACT_*
  float slope;        // LeakyReLU slope
  const float* alpha; // [cin] Snake alpha (ST2_ACT_SNAKE)
  // ---- contraction
  int k, dil, pad, stride;
  const void* w;      // SIMT: fp32 [k][cin][n_cols]; UMMA: bf16 blob from st2_umma_pack_weights
  const float* bias;  // [n_cols] or NULL
  int n_cols;         // GEMM N (= Cout, or stride*Cout for a polyphase transposed conv)
  // ---- epilogue
  void* y;
  int y_dtype;
  int64_t y_bstride;
  int ldy;
  int Lout;           // rows written per item
  const void* res;    // same dtype as y, or NULL
  int64_t res_bstride;
  int ldres;
  int res_shift;      // residual row = t >> res_shift (nearest x2 upsampled shortcut)
  float out_scale;
  int accumulate;     // y = y_old + value
  float* stats;       // [B][n_tiles][n_cols][2] partial (sum, sumsq), or NULL
  int n_tiles;        // ceil(Lout / tile_rows) of the implementation that is called
  double algo_flops;  // algorithmic FLOPs of this call (2*MAC, no padding), bookkeeping only
  int64_t bias_bstride;  // elements between the bias rows of two batch items (0: one shared bias)
  int n_tile;            // UMMA only: GEMM columns per tile the blob was packed for (0 = the default for n_cols)
} St2ConvParams;

#endif
