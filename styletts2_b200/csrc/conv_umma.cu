// conv_umma.cu — fused 1-D convolution as an implicit GEMM on tcgen05 / TMEM (sm_100a).
//
//   D[time 128 x Cout 128] (fp32, TMEM) += A[time x 16 ch] (bf16, smem) * B[Cout x 16 ch] (bf16, smem)
//
// for every (64-channel chunk, tap, 16-channel K step).  Activations are channels-last, so a conv
// tap is a ROW OFFSET into one staged activation tile: the tile (256 output rows + (k-1)*dil halo)
// is loaded, normalised (AdaIN scale/shift), activated (Snake / LeakyReLU) and converted to bf16
// ONCE by the producer warps, written in the UMMA no-swizzle K-major core-matrix layout
//      byte(row r, 16-byte channel group v) = v * (RPAD*16) + r * 16
// in which every row shift keeps the layout canonical (core matrix = 8 consecutive rows x 16 B,
// SBO = 128 B, LBO = RPAD*16 B), so all k taps read the same smem tile through descriptors that
// differ only in their start address.  RPAD is odd, which makes the producers' 16-byte stores
// bank-conflict free.  Weights are pre-packed at load time into the same core-matrix layout, one
// 16 KB blob per (N tile, channel chunk, tap), and streamed with cp.async.bulk (TMA engine,
// mbarrier complete_tx) through a 4-deep ring.
//
// Persistent, warp-specialised CTA (1 per SM, 448 threads):
//   warp 0      MMA issuer (one lane) + TMEM allocation (512 columns = 2 tiles x 2 accumulators)
//   warp 1      weight producer (one lane, bulk copies)
//   warps 2-5   epilogue: tcgen05.ld -> bias / residual / scale / accumulate -> bf16 stores, plus
//               per-column sum / sum-of-squares for the next InstanceNorm via a shuffle transpose-reduce
//   warps 6-13  activation producers (fused conv prologue)
// Pipelines: A ring (3 stages, producers <-> MMA), W ring (4 stages, bulk copy <-> MMA),
// TMEM double buffer (MMA <-> epilogue), so the prologue of tile i+1 and the epilogue of tile i-1
// overlap the MMAs of tile i.
//
// Reference semantics: conv_params.h (Modules/istftnet.py:66-75, 440-453, 358-374).
#include <string.h>

#include <vector>

#include "common.cuh"

namespace st2 {

constexpr int UM_TM = 256;      // output rows per tile
constexpr int UM_NACC = 2;      // 128-row accumulators per tile
constexpr int UM_N = 128;       // output columns per tile
constexpr int UM_CK = 64;       // channels per stage
constexpr int UM_A_STAGES = 3;
constexpr int UM_W_STAGES = 4;
constexpr int UM_MAX_HALO = 50;               // (k-1)*dil of the widest resblock conv (k 11, d 5)
constexpr int UM_RPAD = UM_TM + UM_MAX_HALO + 1;  // 307, odd
constexpr int UM_A_BYTES = 8 * UM_RPAD * 16;      // 39296
constexpr int UM_W_BYTES = 8 * UM_N * 16;         // 16384
constexpr int UM_W_ELEMS = UM_W_BYTES / 2;
constexpr int UM_THREADS = 448;
constexpr int UM_PROD_WARP0 = 6, UM_PROD_THREADS = 256;
constexpr int UM_EPI_WARP0 = 2, UM_EPI_THREADS = 128;
constexpr int UM_NI = (UM_RPAD + 31) / 32;  // row iterations per producer thread and chunk

struct __align__(16) UmSmem {
  uint8_t a[UM_A_STAGES][UM_A_BYTES];
  uint8_t w[UM_W_STAGES][UM_W_BYTES];
  uint64_t a_full[UM_A_STAGES], a_empty[UM_A_STAGES];
  uint64_t w_full[UM_W_STAGES], w_empty[UM_W_STAGES];
  uint64_t t_full[2], t_empty[2];
  uint32_t tmem_base;
  float bias[UM_N];
  float sred[4][UM_N][2];
};

// ------------------------------------------------------------------ PTX wrappers
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("{ .reg .b64 st; mbarrier.arrive.shared::cta.b64 st, [%0]; }" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("{ .reg .b64 st; mbarrier.arrive.expect_tx.shared::cta.b64 st, [%0], %1; }" ::"r"(smem_u32(bar)),
               "r"(bytes)
               : "memory");
}
// bounded wait: a protocol bug traps (launch error) instead of hanging the GPU
__device__ __noinline__ void mbar_timeout(int which, uint32_t parity) {
  printf("st2 conv_umma: mbarrier wait timed out (block %d thread %d barrier-id %d parity %u)\n",
         (int)blockIdx.x, (int)threadIdx.x, which, parity);
  __trap();
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity, int which) {
  const uint32_t addr = smem_u32(bar);
  uint32_t done = 0;
  for (uint32_t spin = 0; !done; ++spin) {
    asm volatile(
        "{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
        : "=r"(done)
        : "r"(addr), "r"(parity)
        : "memory");
    if (!done && spin > (1u << 24)) mbar_timeout(which, parity);
  }
}
__device__ __forceinline__ void fence_barrier_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void fence_proxy_async_smem() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void tc_fence_before() {
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
}
__device__ __forceinline__ void tc_fence_after() {
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
}
__device__ __forceinline__ void bulk_g2s(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   smem_u32(smem_dst)),
               "l"(gsrc), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void umma_bf16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                          uint32_t accumulate) {
  asm volatile(
      "{ .reg .pred p; setp.ne.b32 p, %4, 0; tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p; }" ::"r"(
          tmem_d),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// K-major, SWIZZLE_NONE shared-memory matrix descriptor (cute::UMMA::SmemDescriptor, version 1):
// [0,14) start>>4, [16,30) leading (K-direction core-matrix) byte offset>>4,
// [32,46) stride (8-row group) byte offset>>4, [46,48) version=1, [61,64) layout type 0.
__device__ __forceinline__ uint64_t umma_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  uint64_t d = (uint64_t)((saddr >> 4) & 0x3FFF);
  d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16;
  d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32;
  d |= (uint64_t)1 << 46;
  return d;
}
// instruction descriptor (cute::UMMA::InstrDescriptor): D fp32, A/B bf16, both K-major, M 128, N 128
constexpr uint32_t UM_IDESC = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(UM_N >> 3) << 17) |
                              ((uint32_t)(128 >> 4) << 24);

__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&v)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
        "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]),
        "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]),
        "=r"(v[23]), "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]),
        "=r"(v[30]), "=r"(v[31])
      : "r"(taddr)
      : "memory");
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void named_bar_sync(int id, int nthreads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}
__device__ __forceinline__ uint4 ldg_nc16(const void* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0, %1, %2, %3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}
__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
  __nv_bfloat162 t = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&t);
}
__device__ __forceinline__ float bf16lo(uint32_t u) { return __uint_as_float(u << 16); }
__device__ __forceinline__ float bf16hi(uint32_t u) { return __uint_as_float(u & 0xFFFF0000u); }

// sum over the 32 lanes of v[c] for every c, result for column `lane` returned (31 shuffles)
__device__ __forceinline__ float transpose_reduce32(float (&v)[32], int lane) {
#pragma unroll
  for (int off = 16; off >= 1; off >>= 1) {
    const bool up = (lane & off) != 0;
#pragma unroll
    for (int i = 0; i < off; ++i) {
      const float keep = up ? v[i + off] : v[i];
      const float send = up ? v[i] : v[i + off];
      v[i] = keep + __shfl_xor_sync(0xffffffffu, send, off);
    }
  }
  return v[0];
}

// ------------------------------------------------------------------------- kernel
struct UmTile {
  int b, t0, nt;
};
__device__ __forceinline__ UmTile um_tile(int tile, int n_ttiles, int n_ntiles) {
  UmTile t;
  t.nt = tile % n_ntiles;
  const int bt = tile / n_ntiles;
  t.t0 = (bt % n_ttiles) * UM_TM;
  t.b = bt / n_ttiles;
  return t;
}

template <int ACT, bool HAS_SS>
__device__ __forceinline__ void um_produce(const St2ConvParams& p, UmSmem* sm, int ptid, int total_tiles,
                                           int n_ttiles, int n_ntiles, int nchunks) {
  const int v = ptid & 7, rbase = ptid >> 3;
  const int halo = (p.k - 1) * p.dil;
  const int r_tot = UM_TM + halo;
  const __nv_bfloat16* xg = reinterpret_cast<const __nv_bfloat16*>(p.x);
  uint32_t a_it = 0;
  for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
    const UmTile tl = um_tile(tile, n_ttiles, n_ntiles);
    const int row0 = tl.t0 - p.pad;
    const __nv_bfloat16* xb = xg + (int64_t)tl.b * p.x_bstride;
    for (int c = 0; c < nchunks; ++c, ++a_it) {
      const int stage = a_it % UM_A_STAGES;
      const uint32_t ph = (a_it / UM_A_STAGES) & 1;
      const int ch0 = c * UM_CK + v * 8;
      // issue all global loads of this chunk first (up to 10 x 16 B in flight per thread)
      uint4 raw[UM_NI];
#pragma unroll
      for (int i = 0; i < UM_NI; ++i) {
        const int r = rbase + 32 * i;
        const int tin = row0 + r;
        raw[i] = make_uint4(0u, 0u, 0u, 0u);
        if (r < r_tot && tin >= 0 && tin < p.Lin) raw[i] = ldg_nc16(xb + (int64_t)tin * p.ldx + ch0);
      }
      float sc[8], sh[8], al[8], ial[8];
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        sc[e] = 1.f; sh[e] = 0.f; al[e] = 1.f; ial[e] = 1.f;
        if (HAS_SS) {
          const float2 t = *reinterpret_cast<const float2*>(p.ss + ((int64_t)tl.b * p.ss_stride + ch0 + e) * 2);
          sc[e] = t.x; sh[e] = t.y;
        }
        if (ACT == ST2_ACT_SNAKE) {
          al[e] = p.alpha[ch0 + e];
          ial[e] = 1.f / al[e];
        }
      }
      mbar_wait(&sm->a_empty[stage], ph ^ 1, 0);
      uint8_t* abase = sm->a[stage] + (size_t)v * (UM_RPAD * 16);
#pragma unroll
      for (int i = 0; i < UM_NI; ++i) {
        const int r = rbase + 32 * i;
        if (r >= r_tot) continue;
        const int tin = row0 + r;
        uint4 o = make_uint4(0u, 0u, 0u, 0u);
        if (tin >= 0 && tin < p.Lin) {
          const uint32_t w[4] = {raw[i].x, raw[i].y, raw[i].z, raw[i].w};
          uint32_t q[4];
#pragma unroll
          for (int e2 = 0; e2 < 4; ++e2) {
            float lo = bf16lo(w[e2]), hi = bf16hi(w[e2]);
            if (HAS_SS) {
              lo = fmaf(lo, sc[2 * e2], sh[2 * e2]);
              hi = fmaf(hi, sc[2 * e2 + 1], sh[2 * e2 + 1]);
            }
            lo = apply_act<false>(lo, ACT, p.slope, al[2 * e2], ial[2 * e2]);
            hi = apply_act<false>(hi, ACT, p.slope, al[2 * e2 + 1], ial[2 * e2 + 1]);
            q[e2] = pack_bf16x2(lo, hi);
          }
          o = make_uint4(q[0], q[1], q[2], q[3]);
        }
        *reinterpret_cast<uint4*>(abase + (size_t)r * 16) = o;
      }
      fence_proxy_async_smem();
      mbar_arrive(&sm->a_full[stage]);
    }
  }
}

__global__ void __launch_bounds__(UM_THREADS, 1) conv_umma_kernel(const St2ConvParams p, int total_tiles,
                                                                   int n_ttiles, int n_ntiles) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  UmSmem* sm = reinterpret_cast<UmSmem*>(smem_raw);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nchunks = p.cin / UM_CK;

  if (threadIdx.x == 0) {
    for (int i = 0; i < UM_A_STAGES; ++i) {
      mbar_init(&sm->a_full[i], UM_PROD_THREADS);
      mbar_init(&sm->a_empty[i], 1);
    }
    for (int i = 0; i < UM_W_STAGES; ++i) {
      mbar_init(&sm->w_full[i], 1);
      mbar_init(&sm->w_empty[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&sm->t_full[i], 1);
      mbar_init(&sm->t_empty[i], UM_EPI_THREADS);
    }
    fence_barrier_init();
  }
  if (warp == 0) {  // TMEM: all 512 columns (this CTA owns the SM: ~184 KB smem, 1 block/SM)
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                     smem_u32(&sm->tmem_base)),
                 "r"(512)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = sm->tmem_base;

  if (warp == 0) {
    // ===================================================================== MMA issuer
    if (lane == 0) {
      uint32_t a_it = 0, w_it = 0, t_it = 0;
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++t_it) {
        const uint32_t tb = t_it & 1, tph = (t_it >> 1) & 1;
        mbar_wait(&sm->t_empty[tb], tph ^ 1, 1);
        tc_fence_after();
        for (int c = 0; c < nchunks; ++c, ++a_it) {
          const int as = a_it % UM_A_STAGES;
          mbar_wait(&sm->a_full[as], (a_it / UM_A_STAGES) & 1, 2);
          tc_fence_after();
          const uint32_t a_addr = smem_u32(sm->a[as]);
          for (int j = 0; j < p.k; ++j, ++w_it) {
            const int ws = w_it % UM_W_STAGES;
            mbar_wait(&sm->w_full[ws], (w_it / UM_W_STAGES) & 1, 3);
            tc_fence_after();
            const uint32_t w_addr = smem_u32(sm->w[ws]);
#pragma unroll
            for (int acc = 0; acc < UM_NACC; ++acc) {
              const uint32_t d_tmem = tmem_base + (tb * UM_NACC + acc) * UM_N;
#pragma unroll
              for (int ks = 0; ks < UM_CK / 16; ++ks) {
                const uint64_t ad = umma_desc(
                    a_addr + (uint32_t)(2 * ks) * (UM_RPAD * 16) + (uint32_t)(acc * 128 + j * p.dil) * 16,
                    UM_RPAD * 16, 128);
                const uint64_t bd = umma_desc(w_addr + (uint32_t)(2 * ks) * (UM_N * 16), UM_N * 16, 128);
                umma_bf16(d_tmem, ad, bd, UM_IDESC, (c | j | ks) != 0 ? 1u : 0u);
              }
            }
            umma_commit(&sm->w_empty[ws]);
          }
          umma_commit(&sm->a_empty[as]);
        }
        umma_commit(&sm->t_full[tb]);
      }
    }
  } else if (warp == 1) {
    // ================================================================ weight producer
    if (lane == 0) {
      const __nv_bfloat16* wg = reinterpret_cast<const __nv_bfloat16*>(p.w);
      const int per_nt = nchunks * p.k;
      uint32_t w_it = 0;
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
        const UmTile tl = um_tile(tile, n_ttiles, n_ntiles);
        const __nv_bfloat16* src = wg + (int64_t)tl.nt * per_nt * UM_W_ELEMS;
        for (int s = 0; s < per_nt; ++s, ++w_it) {
          const int ws = w_it % UM_W_STAGES;
          mbar_wait(&sm->w_empty[ws], ((w_it / UM_W_STAGES) & 1) ^ 1, 4);
          mbar_arrive_expect_tx(&sm->w_full[ws], UM_W_BYTES);
          bulk_g2s(sm->w[ws], src + (int64_t)s * UM_W_ELEMS, UM_W_BYTES, &sm->w_full[ws]);
        }
      }
    }
  } else if (warp < UM_PROD_WARP0) {
    // ======================================================================= epilogue
    const int q = warp & 3;                       // TMEM lane group this warp may access
    const int et = threadIdx.x - UM_EPI_WARP0 * 32;  // 0..127
    const int ew = warp - UM_EPI_WARP0;
    __nv_bfloat16* yg = reinterpret_cast<__nv_bfloat16*>(p.y);
    const __nv_bfloat16* rg = reinterpret_cast<const __nv_bfloat16*>(p.res);
    uint32_t t_it = 0;
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++t_it) {
      const UmTile tl = um_tile(tile, n_ttiles, n_ntiles);
      const uint32_t tb = t_it & 1, tph = (t_it >> 1) & 1;
      sm->bias[et] = p.bias ? p.bias[tl.nt * UM_N + et] : 0.f;
      named_bar_sync(1, UM_EPI_THREADS);
      mbar_wait(&sm->t_full[tb], tph, 5);
      tc_fence_after();
      float st1[4] = {0.f, 0.f, 0.f, 0.f}, st2[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
      for (int acc = 0; acc < UM_NACC; ++acc) {
        const int row = tl.t0 + acc * 128 + q * 32 + lane;
        const bool valid = row < p.Lout;
        __nv_bfloat16* yrow = yg + (int64_t)tl.b * p.y_bstride + (int64_t)row * p.ldy + tl.nt * UM_N;
        const __nv_bfloat16* rrow =
            rg ? rg + (int64_t)tl.b * p.res_bstride + (int64_t)(row >> p.res_shift) * p.ldres + tl.nt * UM_N
               : nullptr;
#pragma unroll
        for (int cb = 0; cb < 4; ++cb) {
          uint32_t v[32];
          tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + (tb * UM_NACC + acc) * UM_N + cb * 32, v);
          float o[32];
#pragma unroll
          for (int i = 0; i < 32; ++i) o[i] = __uint_as_float(v[i]) + sm->bias[cb * 32 + i];
          if (valid) {
            if (rrow) {
#pragma unroll
              for (int g = 0; g < 4; ++g) {
                const uint4 r4 = *reinterpret_cast<const uint4*>(rrow + cb * 32 + g * 8);
                const uint32_t rw[4] = {r4.x, r4.y, r4.z, r4.w};
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                  o[g * 8 + 2 * e] += bf16lo(rw[e]);
                  o[g * 8 + 2 * e + 1] += bf16hi(rw[e]);
                }
              }
            }
#pragma unroll
            for (int i = 0; i < 32; ++i) o[i] *= p.out_scale;
            if (p.accumulate) {
#pragma unroll
              for (int g = 0; g < 4; ++g) {
                const uint4 r4 = *reinterpret_cast<const uint4*>(yrow + cb * 32 + g * 8);
                const uint32_t rw[4] = {r4.x, r4.y, r4.z, r4.w};
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                  o[g * 8 + 2 * e] += bf16lo(rw[e]);
                  o[g * 8 + 2 * e + 1] += bf16hi(rw[e]);
                }
              }
            }
#pragma unroll
            for (int g = 0; g < 4; ++g) {
              uint4 w4;
              w4.x = pack_bf16x2(o[g * 8 + 0], o[g * 8 + 1]);
              w4.y = pack_bf16x2(o[g * 8 + 2], o[g * 8 + 3]);
              w4.z = pack_bf16x2(o[g * 8 + 4], o[g * 8 + 5]);
              w4.w = pack_bf16x2(o[g * 8 + 6], o[g * 8 + 7]);
              *reinterpret_cast<uint4*>(yrow + cb * 32 + g * 8) = w4;
            }
          }
          if (p.stats) {
            float a1[32], a2[32];
#pragma unroll
            for (int i = 0; i < 32; ++i) {
              const float r = valid ? round_to<__nv_bfloat16>(o[i]) : 0.f;
              a1[i] = r;
              a2[i] = r * r;
            }
            st1[cb] += transpose_reduce32(a1, lane);
            st2[cb] += transpose_reduce32(a2, lane);
          }
        }
      }
      // all TMEM reads of this buffer are complete (tcgen05.wait::ld inside tmem_ld32)
      tc_fence_before();
      mbar_arrive(&sm->t_empty[tb]);
      if (p.stats) {
#pragma unroll
        for (int cb = 0; cb < 4; ++cb) {
          sm->sred[ew][cb * 32 + lane][0] = st1[cb];
          sm->sred[ew][cb * 32 + lane][1] = st2[cb];
        }
        named_bar_sync(1, UM_EPI_THREADS);
        const float s = sm->sred[0][et][0] + sm->sred[1][et][0] + sm->sred[2][et][0] + sm->sred[3][et][0];
        const float s2 = sm->sred[0][et][1] + sm->sred[1][et][1] + sm->sred[2][et][1] + sm->sred[3][et][1];
        float* sp = p.stats +
                    (((int64_t)tl.b * p.n_tiles + tl.t0 / UM_TM) * p.n_cols + tl.nt * UM_N + et) * 2;
        sp[0] = s;
        sp[1] = s2;
      }
    }
  } else {
    // ============================================================ activation producers
    const int ptid = threadIdx.x - UM_PROD_WARP0 * 32;
    if (p.ss) {
      if (p.act == ST2_ACT_SNAKE) um_produce<ST2_ACT_SNAKE, true>(p, sm, ptid, total_tiles, n_ttiles, n_ntiles, nchunks);
      else if (p.act == ST2_ACT_LRELU) um_produce<ST2_ACT_LRELU, true>(p, sm, ptid, total_tiles, n_ttiles, n_ntiles, nchunks);
      else um_produce<ST2_ACT_NONE, true>(p, sm, ptid, total_tiles, n_ttiles, n_ntiles, nchunks);
    } else {
      if (p.act == ST2_ACT_SNAKE) um_produce<ST2_ACT_SNAKE, false>(p, sm, ptid, total_tiles, n_ttiles, n_ntiles, nchunks);
      else if (p.act == ST2_ACT_LRELU) um_produce<ST2_ACT_LRELU, false>(p, sm, ptid, total_tiles, n_ttiles, n_ntiles, nchunks);
      else um_produce<ST2_ACT_NONE, false>(p, sm, ptid, total_tiles, n_ttiles, n_ntiles, nchunks);
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 0) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512) : "memory");
  }
}

// ------------------------------------------------------------------------ host side
int conv_tile_rows(bool umma) { return umma ? UM_TM : 64; }

static int g_num_sms[64] = {};

// optional per-launch event timing (st2_prof_enable / st2_prof_read)
struct UmProf {
  bool on = false;
  std::vector<cudaEvent_t> ev;  // pairs
  size_t used = 0;
  double flops = 0.0;
};
static thread_local UmProf g_prof;

int launch_conv_umma(const St2ConvParams& p, cudaStream_t st) {
  ST2_REQUIRE(p.B > 0 && p.Lout > 0, "conv_umma: empty shape");
  ST2_REQUIRE(p.x_dtype == ST2_DT_BF16 && p.y_dtype == ST2_DT_BF16, "conv_umma: bf16 activations only");
  ST2_REQUIRE(p.cin % UM_CK == 0 && p.cin > 0, "conv_umma: cin must be a multiple of 64");
  ST2_REQUIRE(p.n_cols % UM_N == 0, "conv_umma: n_cols must be a multiple of 128");
  ST2_REQUIRE(p.stride == 1, "conv_umma: stride 1 only");
  ST2_REQUIRE(p.k >= 1 && p.dil >= 1 && (p.k - 1) * p.dil <= UM_MAX_HALO, "conv_umma: halo too wide");
  ST2_REQUIRE(p.ldx % 8 == 0 && p.ldy % 8 == 0 && p.x_bstride % 8 == 0 && p.y_bstride % 8 == 0,
              "conv_umma: 16-byte alignment of rows");
  ST2_REQUIRE(((uintptr_t)p.x & 15) == 0 && ((uintptr_t)p.y & 15) == 0 && ((uintptr_t)p.w & 15) == 0,
              "conv_umma: 16-byte alignment of pointers");
  ST2_REQUIRE(!p.res || (p.ldres % 8 == 0 && p.res_bstride % 8 == 0 && ((uintptr_t)p.res & 15) == 0),
              "conv_umma: residual alignment");
  ST2_REQUIRE(!(p.stats && p.accumulate), "conv_umma: stats with accumulate");
  ST2_REQUIRE(p.act != ST2_ACT_SNAKE || p.alpha, "conv_umma: snake needs alpha");
  const int n_ttiles = ceil_div(p.Lout, UM_TM), n_ntiles = p.n_cols / UM_N;
  ST2_REQUIRE(!p.stats || p.n_tiles >= n_ttiles, "conv_umma: n_tiles");
  const int64_t total = (int64_t)p.B * n_ttiles * n_ntiles;
  ST2_REQUIRE(total < (1ll << 31), "conv_umma: too many tiles");
  int dev = 0;
  ST2_CUDA_OK(cudaGetDevice(&dev));
  if (dev >= 64 || g_num_sms[dev] == 0) {
    int n = 0;
    ST2_CUDA_OK(cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev));
    ST2_CUDA_OK(cudaFuncSetAttribute(conv_umma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     (int)sizeof(UmSmem)));
    if (dev < 64) g_num_sms[dev] = n;
  }
  int sms = dev < 64 ? g_num_sms[dev] : 148;
  const int grid = (int)(total < sms ? total : sms);
  cudaEvent_t e0 = nullptr, e1 = nullptr;
  if (g_prof.on) {
    if (g_prof.used + 2 > g_prof.ev.size()) {
      for (int i = 0; i < 2; ++i) {
        cudaEvent_t e;
        ST2_CUDA_OK(cudaEventCreate(&e));
        g_prof.ev.push_back(e);
      }
    }
    e0 = g_prof.ev[g_prof.used];
    e1 = g_prof.ev[g_prof.used + 1];
    g_prof.used += 2;
    g_prof.flops += p.algo_flops;
    ST2_CUDA_OK(cudaEventRecord(e0, st));
  }
  conv_umma_kernel<<<grid, UM_THREADS, sizeof(UmSmem), st>>>(p, (int)total, n_ttiles, n_ntiles);
  ST2_LAUNCH_OK("conv_umma");
  if (e1) ST2_CUDA_OK(cudaEventRecord(e1, st));
  return ST2_OK;
}

}  // namespace st2

extern "C" int st2_prof_enable(int on) {
  st2::g_prof.on = on != 0;
  return ST2_OK;
}
extern "C" int st2_prof_read(double* umma_ms, int64_t* umma_launches, double* umma_flops, int reset) {
  double ms = 0.0;
  for (size_t i = 0; i + 1 < st2::g_prof.used; i += 2) {
    float t = 0.f;
    ST2_CUDA_OK(cudaEventSynchronize(st2::g_prof.ev[i + 1]));
    ST2_CUDA_OK(cudaEventElapsedTime(&t, st2::g_prof.ev[i], st2::g_prof.ev[i + 1]));
    ms += t;
  }
  if (umma_ms) *umma_ms = ms;
  if (umma_launches) *umma_launches = (int64_t)(st2::g_prof.used / 2);
  if (umma_flops) *umma_flops = st2::g_prof.flops;
  if (reset) {
    st2::g_prof.used = 0;
    st2::g_prof.flops = 0.0;
  }
  return ST2_OK;
}

// blob order: [n tile][64-channel chunk][tap][16-byte channel group v (8)][column n (128)][8 channels]
extern "C" int64_t st2_umma_packed_elems(int k, int cin_pad, int n) {
  return (int64_t)(n / st2::UM_N) * (cin_pad / st2::UM_CK) * k * st2::UM_W_ELEMS;
}

static inline uint16_t f32_to_bf16_rn(float f) {
  uint32_t u;
  memcpy(&u, &f, 4);
  if ((u & 0x7fffffffu) > 0x7f800000u) return (uint16_t)((u >> 16) | 0x40);
  u += 0x7fffu + ((u >> 16) & 1u);
  return (uint16_t)(u >> 16);
}

extern "C" int st2_umma_pack_weights(const float* w_kcn, int k, int cin, int cin_pad, int n,
                                     uint16_t* out) {
  if (!w_kcn || !out || k < 1 || cin < 1 || cin_pad < cin || cin_pad % st2::UM_CK || n % st2::UM_N) {
    st2::set_error("st2_umma_pack_weights: bad shape");
    return ST2_EINVAL;
  }
  const int nch = cin_pad / st2::UM_CK, nnt = n / st2::UM_N;
  for (int nt = 0; nt < nnt; ++nt)
    for (int c = 0; c < nch; ++c)
      for (int j = 0; j < k; ++j) {
        uint16_t* blob = out + (((int64_t)nt * nch + c) * k + j) * st2::UM_W_ELEMS;
        for (int v = 0; v < 8; ++v)
          for (int nn = 0; nn < st2::UM_N; ++nn)
            for (int e = 0; e < 8; ++e) {
              const int ch = c * st2::UM_CK + v * 8 + e;
              const float val = ch < cin ? w_kcn[((int64_t)j * cin + ch) * n + nt * st2::UM_N + nn] : 0.f;
              blob[(v * st2::UM_N + nn) * 8 + e] = f32_to_bf16_rn(val);
            }
      }
  return ST2_OK;
}
