// lstm.cu — the recurrent half of ProsodyPredictor.shared (bidirectional nn.LSTM, hidden 256,
// /root/reference/models.py:562-564, 639).  The input projection W_ih x + b_ih + b_hh of every time
// step is one dense conv (run by the conv kernels); what is left is T dependent steps of
//      g = pre[t] + W_hh h;  c = sig(f) c + sig(i) tanh(g);  h = sig(o) tanh(c)      (gate order i, f, g, o)
// per (item, direction).  One thread-block CLUSTER of 8 CTAs owns one direction of a group of NB
// items: CTA q keeps the 4 x 32 gate rows of hidden units [32 q, 32 q + 32) of W_hh in REGISTERS
// (128 rows x 256 = 128 fp32 per thread: lane = 4 rows, warp = a 32-wide K slice, so one broadcast
// LDS.128 of the hidden vector feeds 16 FMAs), computes those gates for all NB items from the full
// hidden vector in its shared memory, updates its 32 units and writes the new values into the
// shared memory of all 8 CTAs (distributed shared memory) with st.async: every remote store reports its
// 4 bytes on an mbarrier of the DESTINATION CTA, so a CTA starts step t+1 as soon as its own copy of h(t)
// is complete (dataflow synchronisation; no cluster-wide barrier inside the time loop).
#include <cooperative_groups.h>

#include "common.cuh"

namespace st2 {

constexpr int LS_H = 256;       // hidden units per direction
constexpr int LS_CTAS = 8;      // cluster size
constexpr int LS_UNITS = LS_H / LS_CTAS;  // 32 units per CTA
constexpr int LS_ROWS = 4 * LS_UNITS;     // 128 gate rows per CTA
constexpr int LS_NB = 4;        // items per cluster
constexpr int LS_THREADS = 256;
constexpr int LS_KS = 8;                  // K slices (one per warp)
constexpr int LS_KW = LS_H / LS_KS;       // 32 hidden values per slice

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

// FAST (bf16 path): SFU exp; the fp32 parity path keeps libdevice expf / tanhf
template <bool FAST>
__device__ __forceinline__ float ls_sigmoid(float x) {
  return FAST ? __fdividef(1.f, 1.f + __expf(-x)) : 1.f / (1.f + expf(-x));
}
template <bool FAST>
__device__ __forceinline__ float ls_tanh(float x) {
  if (!FAST) return tanhf(x);
  const float e = __expf(-2.f * fabsf(x));  // in (0, 1]: no overflow
  return copysignf(__fdividef(1.f - e, 1.f + e), x);
}
__device__ __forceinline__ uint32_t ls_smem(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint32_t ls_mapa(uint32_t addr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(addr), "r"(rank));
  return r;
}
// 4-byte store into the shared memory of a CTA of the cluster; completion is reported as 4 transaction bytes
// on that CTA's mbarrier
__device__ __forceinline__ void ls_st_async(uint32_t remote_addr, float v, uint32_t remote_bar) {
  asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.b32 [%0], %1, [%2];" ::"r"(remote_addr),
               "r"(__float_as_uint(v)), "r"(remote_bar)
               : "memory");
}
// bounded wait (a protocol bug traps instead of hanging the GPU): ~2 s of 1 us suspended probes
__device__ __forceinline__ void ls_bar_wait(uint32_t bar, uint32_t parity) {
  uint32_t done = 0;
  for (uint32_t spin = 0; spin < (1u << 21); ++spin) {
    asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3; selp.u32 %0, 1, 0, p; }"
                 : "=r"(done)
                 : "r"(bar), "r"(parity), "r"(1000u)
                 : "memory");
    if (done) return;
  }
  printf("lstm_scan: hidden-state barrier timed out (block %d)\n", (int)blockIdx.x);
  __trap();
}

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
  __shared__ __align__(16) float part_s[LS_KS][LS_NB][LS_ROWS]; // partial dot products of the 8 K slices
  __shared__ __align__(8) uint64_t h_bar[2];                // h_s[i] of the coming step is complete (transaction bytes)
  constexpr bool FAST = sizeof(AT) == 2;
  if (tid == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(ls_smem(&h_bar[0])));
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(ls_smem(&h_bar[1])));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }

  // this thread's share of W_hh: local rows 4 rg .. 4 rg + 3 (gate = row / 32, unit 32 q + row % 32), K slice ks.
  // All lanes of a warp share ks, so the hidden-vector loads below are warp-wide broadcasts, and each of them
  // feeds 4 rows x 4 values: 32 LDS.128 per thread and step instead of one per 4 FMAs.
  const int rg = tid & 31, ks = tid >> 5;
  float w[4][LS_KW];
#pragma unroll
  for (int r = 0; r < 4; ++r) {
    const int lr = 4 * rg + r;
    const int grow = (lr >> 5) * LS_H + q * LS_UNITS + (lr & 31);
    const float4* wp = reinterpret_cast<const float4*>(whh + ((int64_t)dir * 4 * LS_H + grow) * LS_H + ks * LS_KW);
#pragma unroll
    for (int k = 0; k < LS_KW / 4; ++k) {
      const float4 t = __ldg(wp + k);
      w[r][4 * k] = t.x; w[r][4 * k + 1] = t.y; w[r][4 * k + 2] = t.z; w[r][4 * k + 3] = t.w;
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
  // bytes every step delivers into this CTA's h_s: 8 CTAs x 32 units x valid items x 4
  const uint32_t step_bytes = (uint32_t)(LS_CTAS * LS_UNITS * 4 * min(LS_NB, B - b0));
  // remote addresses of this owner's slot (both buffers) and of the barriers in every CTA of the cluster
  const uint32_t bar_u[2] = {ls_smem(&h_bar[0]), ls_smem(&h_bar[1])};
  cluster.sync();  // every CTA's h_s is zeroed and its barriers are initialised before anyone writes remotely

  for (int step = 0; step < T; ++step) {
    const int cur = step & 1, nxt = cur ^ 1;
    // h(step - 1) has arrived from all 8 CTAs (phase (step - 1) / 2 of the barrier of buffer `cur`)
    if (step > 0) ls_bar_wait(bar_u[cur], (uint32_t)(((step - 1) >> 1) & 1));
    if (tid == 0 && step + 1 < T)  // arm the barrier of the buffer this step fills (its previous phase was consumed)
      asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar_u[nxt]), "r"(step_bytes) : "memory");
    float pc[4] = {pn[0], pn[1], pn[2], pn[3]};
    if (owner && step + 1 < T) load_pre(step + 1, pn);  // in flight during this step's arithmetic
    float acc[LS_NB][4];
#pragma unroll
    for (int i = 0; i < LS_NB; ++i)
#pragma unroll
      for (int r = 0; r < 4; ++r) acc[i][r] = 0.f;
    const float4* hp = reinterpret_cast<const float4*>(&h_s[cur][0][ks * LS_KW]);
#pragma unroll
    for (int k = 0; k < LS_KW / 4; ++k) {
#pragma unroll
      for (int i = 0; i < LS_NB; ++i) {
        const float4 hv = hp[i * (LS_H / 4) + k];  // same address across the warp: broadcast
#pragma unroll
        for (int r = 0; r < 4; ++r) {
          acc[i][r] = fmaf(w[r][4 * k], hv.x, acc[i][r]);
          acc[i][r] = fmaf(w[r][4 * k + 1], hv.y, acc[i][r]);
          acc[i][r] = fmaf(w[r][4 * k + 2], hv.z, acc[i][r]);
          acc[i][r] = fmaf(w[r][4 * k + 3], hv.w, acc[i][r]);
        }
      }
    }
#pragma unroll
    for (int i = 0; i < LS_NB; ++i)
      *reinterpret_cast<float4*>(&part_s[ks][i][4 * rg]) = make_float4(acc[i][0], acc[i][1], acc[i][2], acc[i][3]);
    __syncthreads();
    if (owner) {
      float g4[4];
#pragma unroll
      for (int g = 0; g < 4; ++g) {
        float a = 0.f;
#pragma unroll
        for (int k = 0; k < LS_KS; ++k) a += part_s[k][n][g * LS_UNITS + jj];  // fixed order: deterministic
        g4[g] = pc[g] + a;
      }
      c = ls_sigmoid<FAST>(g4[1]) * c + ls_sigmoid<FAST>(g4[0]) * ls_tanh<FAST>(g4[2]);
      const float hv = ls_sigmoid<FAST>(g4[3]) * ls_tanh<FAST>(c);
      if (step + 1 < T) {
        const uint32_t slot = ls_smem(&h_s[nxt][n][q * LS_UNITS + jj]);
#pragma unroll
        for (int r = 0; r < LS_CTAS; ++r) ls_st_async(ls_mapa(slot, (uint32_t)r), hv, ls_mapa(bar_u[nxt], (uint32_t)r));
      }
      const int t = dir ? T - 1 - step : step;
      ls_st<AT>(out + ((int64_t)(b0 + n) * T + t) * (2 * LS_H) + dir * LS_H + q * LS_UNITS + jj, hv);
    }
    // No barrier here: part_s is rewritten only after the next wait, which cannot pass before this CTA's own
    // owners have sent h(step), i.e. after they have read part_s; h_s[cur] is rewritten (by step + 1 of a faster
    // CTA) only after that CTA has received h(step) from every CTA, i.e. after everyone has finished reading it.
  }
  cluster.sync();  // no CTA exits while stores into its shared memory may still be in flight
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
