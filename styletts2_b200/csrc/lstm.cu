// lstm.cu — the recurrent half of ProsodyPredictor.shared (bidirectional nn.LSTM, hidden 256,
// /root/reference/models.py:562-564, 639).  The input projection W_ih x + b_ih + b_hh of every time
// step is one dense conv (run by the conv kernels); what is left is T dependent steps of
//      g = pre[t] + W_hh h;  c = sig(f) c + sig(i) tanh(g);  h = sig(o) tanh(c)      (gate order i, f, g, o)
// per (item, direction).  One thread-block CLUSTER of 8 CTAs owns one direction of a group of NB
// items: CTA q keeps the 4 x 32 gate rows of hidden units [32 q, 32 q + 32) of W_hh in REGISTERS
// (128 rows x 256 = 128 fp32 per thread), computes those gates for all NB items from the full
// hidden vector in its shared memory, updates its 32 units and writes the new values into the
// shared memory of all 8 CTAs (distributed shared memory), one cluster barrier per step.
#include <cooperative_groups.h>

#include "common.cuh"

namespace st2 {

constexpr int LS_H = 256;       // hidden units per direction
constexpr int LS_CTAS = 8;      // cluster size
constexpr int LS_UNITS = LS_H / LS_CTAS;  // 32 units per CTA
constexpr int LS_ROWS = 4 * LS_UNITS;     // 128 gate rows per CTA
constexpr int LS_NB = 4;        // items per cluster
constexpr int LS_THREADS = 256;

template <typename T>
__device__ __forceinline__ float ls_ld(const T* p);
template <>
__device__ __forceinline__ float ls_ld<float>(const float* p) { return __ldg(p); }
template <>
__device__ __forceinline__ float ls_ld<__nv_bfloat16>(const __nv_bfloat16* p) { return __bfloat162float(*p); }
template <typename T>
__device__ __forceinline__ void ls_st(T* p, float v);
template <>
__device__ __forceinline__ void ls_st<float>(float* p, float v) { *p = v; }
template <>
__device__ __forceinline__ void ls_st<__nv_bfloat16>(__nv_bfloat16* p, float v) { *p = __float2bfloat16_rn(v); }

__device__ __forceinline__ float ls_sigmoid(float x) { return 1.f / (1.f + expf(-x)); }

// pre [B][T][2][4H] (direction-major columns: dir * 4H + gate * H + unit), whh [2][4H][H] fp32,
// out [B][T][2H] (forward units then backward units, like torch's bidirectional output)
template <typename AT>
__global__ void __cluster_dims__(LS_CTAS, 1, 1) __launch_bounds__(LS_THREADS, 1)
lstm_scan_kernel(const AT* __restrict__ pre, const float* __restrict__ whh, AT* __restrict__ out, int B, int T) {
  namespace cg = cooperative_groups;
  cg::cluster_group cluster = cg::this_cluster();
  const int q = (int)cluster.block_rank();
  const int cl = blockIdx.x / LS_CTAS;
  const int n_groups = (B + LS_NB - 1) / LS_NB;
  const int dir = cl / n_groups, b0 = (cl % n_groups) * LS_NB;
  const int tid = threadIdx.x;

  __shared__ __align__(16) float h_s[2][LS_NB][LS_H];       // full hidden vectors, double-buffered by step
  __shared__ __align__(16) float part_s[2][LS_ROWS][LS_NB]; // partial dot products of the two K halves

  // this thread's share of W_hh: local row lr (gate g = lr / 32, unit 32 q + lr % 32), K half kh
  const int lr = tid & (LS_ROWS - 1), kh = tid >> 7;
  const int grow = (lr >> 5) * LS_H + q * LS_UNITS + (lr & 31);
  float w[LS_H / 2];
  {
    const float4* wp = reinterpret_cast<const float4*>(whh + ((int64_t)dir * 4 * LS_H + grow) * LS_H + kh * (LS_H / 2));
#pragma unroll
    for (int k = 0; k < LS_H / 8; ++k) {
      const float4 t = __ldg(wp + k);
      w[4 * k] = t.x; w[4 * k + 1] = t.y; w[4 * k + 2] = t.z; w[4 * k + 3] = t.w;
    }
  }
  for (int i = tid; i < 2 * LS_NB * LS_H; i += LS_THREADS) (&h_s[0][0][0])[i] = 0.f;

  // cell owner: threads 0..127 = (unit jj, item n)
  const int jj = tid & 31, n = tid >> 5;
  const bool owner = tid < LS_UNITS * LS_NB && b0 + n < B;
  float c = 0.f;
  float pn[4] = {0.f, 0.f, 0.f, 0.f};
  auto load_pre = [&](int step, float (&p4)[4]) {
    const int t = dir ? T - 1 - step : step;
    const AT* pp = pre + (((int64_t)(b0 + n) * T + t) * 2 + dir) * (4 * LS_H) + q * LS_UNITS + jj;
#pragma unroll
    for (int g = 0; g < 4; ++g) p4[g] = ls_ld<AT>(pp + g * LS_H);
  };
  if (owner && T > 0) load_pre(0, pn);
  cluster.sync();  // every CTA's h_s is zeroed before anyone writes into it remotely

  for (int step = 0; step < T; ++step) {
    const int cur = step & 1, nxt = cur ^ 1;
    float pc[4] = {pn[0], pn[1], pn[2], pn[3]};
    if (owner && step + 1 < T) load_pre(step + 1, pn);  // in flight during this step's arithmetic
    float acc[LS_NB];
#pragma unroll
    for (int i = 0; i < LS_NB; ++i) acc[i] = 0.f;
    const float4* hp = reinterpret_cast<const float4*>(&h_s[cur][0][kh * (LS_H / 2)]);
#pragma unroll
    for (int k = 0; k < LS_H / 8; ++k) {
#pragma unroll
      for (int i = 0; i < LS_NB; ++i) {
        const float4 hv = hp[i * (LS_H / 4) + k];  // same address across the warp: broadcast
        acc[i] = fmaf(w[4 * k], hv.x, acc[i]);
        acc[i] = fmaf(w[4 * k + 1], hv.y, acc[i]);
        acc[i] = fmaf(w[4 * k + 2], hv.z, acc[i]);
        acc[i] = fmaf(w[4 * k + 3], hv.w, acc[i]);
      }
    }
    static_assert(LS_NB == 4, "one float4 of partial sums per gate row");
    *reinterpret_cast<float4*>(&part_s[kh][lr][0]) = make_float4(acc[0], acc[1], acc[2], acc[3]);
    __syncthreads();
    if (owner) {
      float g4[4];
#pragma unroll
      for (int g = 0; g < 4; ++g) g4[g] = pc[g] + (part_s[0][g * LS_UNITS + jj][n] + part_s[1][g * LS_UNITS + jj][n]);
      c = ls_sigmoid(g4[1]) * c + ls_sigmoid(g4[0]) * tanhf(g4[2]);
      const float hv = ls_sigmoid(g4[3]) * tanhf(c);
      float* slot = &h_s[nxt][n][q * LS_UNITS + jj];
#pragma unroll
      for (int r = 0; r < LS_CTAS; ++r) *cluster.map_shared_rank(slot, r) = hv;
      const int t = dir ? T - 1 - step : step;
      ls_st<AT>(out + ((int64_t)(b0 + n) * T + t) * (2 * LS_H) + dir * LS_H + q * LS_UNITS + jj, hv);
    }
    cluster.sync();  // new hidden vector visible everywhere; part_s may be overwritten
  }
}

template <typename AT>
int launch_lstm_scan(const AT* pre, const float* whh, AT* out, int B, int T, cudaStream_t st) {
  const int n_groups = (B + LS_NB - 1) / LS_NB;
  lstm_scan_kernel<AT><<<2 * n_groups * LS_CTAS, LS_THREADS, 0, st>>>(pre, whh, out, B, T);
  ST2_LAUNCH_OK("lstm_scan");
  return ST2_OK;
}
template int launch_lstm_scan<float>(const float*, const float*, float*, int, int, cudaStream_t);
template int launch_lstm_scan<__nv_bfloat16>(const __nv_bfloat16*, const float*, __nv_bfloat16*, int, int, cudaStream_t);

// NCL fp32 -> channels-last (fp32 or bf16) with row stride ld (columns >= C are left untouched)
template <typename AT>
__global__ void ncl_to_act_kernel(const float* __restrict__ src, int C, int L, AT* __restrict__ dst, int ld) {
  __shared__ float tile[32][33];
  const int b = blockIdx.z, c0 = blockIdx.y * 32, t0 = blockIdx.x * 32;
  const int tx = threadIdx.x, ty = threadIdx.y;  // 32 x 8
  for (int i = ty; i < 32; i += 8) {
    const int c = c0 + i, t = t0 + tx;
    tile[i][tx] = (c < C && t < L) ? src[((int64_t)b * C + c) * L + t] : 0.f;
  }
  __syncthreads();
  for (int i = ty; i < 32; i += 8) {
    const int t = t0 + i, c = c0 + tx;
    if (t < L && c < C) ls_st<AT>(dst + ((int64_t)b * L + t) * ld + c, tile[tx][i]);
  }
}
template <typename AT>
int launch_ncl_to_act(const float* src, int C, int L, AT* dst, int ld, int B, cudaStream_t st) {
  dim3 grid((L + 31) / 32, (C + 31) / 32, B), block(32, 8);
  ncl_to_act_kernel<AT><<<grid, block, 0, st>>>(src, C, L, dst, ld);
  ST2_LAUNCH_OK("ncl_to_act");
  return ST2_OK;
}
template int launch_ncl_to_act<float>(const float*, int, int, float*, int, int, cudaStream_t);
template int launch_ncl_to_act<__nv_bfloat16>(const float*, int, int, __nv_bfloat16*, int, int, cudaStream_t);

}  // namespace st2
