// conv_umma.cu — fused 1-D convolution as an implicit GEMM on tcgen05 / TMEM (sm_100a).
//
//   D[time 128 x Cout NT] (fp32, TMEM) += A[time x 16 ch] (bf16, smem) * B[Cout x 16 ch] (bf16, smem)
//
// for every (64-channel chunk, tap, 16-channel K step); NT (columns per tile) is 128, 64 or 32.
// Activations are channels-last, so a conv tap is a ROW OFFSET into one staged activation tile: the
// tile (256 output rows + (k-1)*dil halo) lives in the UMMA no-swizzle K-major core-matrix layout
//      byte(row r, 16-byte channel group v) = v * (RPAD*16) + r * 16
// in which every row shift keeps the layout canonical (core matrix = 8 consecutive rows x 16 B,
// SBO = 128 B, LBO = RPAD*16 B), so all k taps read the same smem tile through descriptors that
// differ only in their start address.  RPAD is odd, which makes 16-byte accesses of 8 planes
// bank-conflict free.  Weights are pre-packed at load time into the same core-matrix layout, one
// blob per (N tile, channel chunk, tap), streamed with cp.async.bulk (mbarrier complete_tx).
//
// Persistent, warp-specialised CTA (1 per SM, 448 threads):
//   warp 0      MMA issuer (one lane) + TMEM allocation (512 columns: 2 tiles x 2 accumulators)
//   warp 1      weight producer (one lane, bulk copies through a 3-deep ring)
//   warps 2-5   epilogue: tcgen05.ld -> (+residual) * scale + bias (+ y_old) -> staging tile in smem
//               -> one bulk store per output row (coalesced, asynchronous), and per-column
//               sum / sum-of-squares of the stored values for the next InstanceNorm, read back
//               from the staging tile (2 columns per thread, conflict free)
//   warps 6-13  activation producers (the fused conv prologue): cp.async (LDGSTS) the raw bf16 tile
//               straight into its final place in the A ring D chunks ahead, then normalise (AdaIN
//               scale/shift), activate (Snake / LeakyReLU) and convert IN PLACE
// Pipelines: A ring (3-4 stages: loading / transforming / feeding the MMAs), W ring, TMEM double
// buffer (MMA <-> epilogue): the prologue of tile i+1 and the epilogue of tile i-1 overlap the
// MMAs of tile i.  No thread of this kernel waits on a global load it has just issued.
//
// Reference semantics: conv_params.h (Modules/istftnet.py:66-75, 440-453, 358-374).
#include <string.h>

#include <vector>

#include "common.cuh"

namespace st2 {

constexpr int UM_TM = 256;      // output rows per tile
constexpr int UM_NACC = 2;      // 128-row accumulators per tile
constexpr int UM_CK = 64;       // channels per A stage
constexpr int UM_MAX_A = 4;     // A ring stages (3 when the halo is wide)
constexpr int UM_W_STAGES = 3;
constexpr int UM_MAX_HALO = 50; // (k-1)*dil of the widest resblock conv (k 11, d 5)
constexpr int UM_THREADS = 448;
constexpr int UM_PROD_WARP0 = 6, UM_PROD_THREADS = 256;
constexpr int UM_EPI_WARP0 = 2, UM_EPI_THREADS = 128;
constexpr int UM_NI = (UM_TM + UM_MAX_HALO + 31) / 32;  // row iterations per producer thread and chunk
constexpr int UM_SMEM_LIMIT = 227 * 1024;

struct UmLayout {
  int n_tile;         // GEMM columns per tile (128 / 64 / 32)
  int rpad;           // rows per plane of an A stage (odd)
  int a_bytes;        // bytes per A stage (8 planes)
  int na;             // A stages in use
  int dist;           // prefetch distance of the producers, in chunks (na - 2)
  int w_bytes;        // bytes per weight stage (= 8 * n_tile * 16)
  int stg_row;        // bytes per row of the epilogue staging tile (row + 16 B pad)
  int out_row;        // bytes per output row segment (n_tile * element size)
  uint32_t off_w, off_stg, off_misc, total;
  uint32_t idesc;
};

struct UmMisc {
  uint64_t a_full[UM_MAX_A], a_empty[UM_MAX_A];
  uint64_t w_full[UM_W_STAGES], w_empty[UM_W_STAGES];
  uint64_t t_full[2], t_empty[2];
  uint32_t tmem_base, pad_;
  alignas(16) float bias[2][128];
  alignas(16) float sred[256][2];
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
// smem -> global bulk store (TMA engine), tracked by the issuing thread's bulk async-group
__device__ __forceinline__ void bulk_s2g(void* gdst, const void* smem_src, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gdst),
               "r"(smem_u32(smem_src)), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read0() {
  asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
}
__device__ __forceinline__ void bulk_wait_all0() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
// 16-byte LDGSTS, L2 only
__device__ __forceinline__ void cp_async16(uint32_t smem_dst, const void* gsrc) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_dst), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}
__device__ __forceinline__ uint4 lds128(uint32_t addr) {
  uint4 r;
  asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "r"(addr) : "memory");
  return r;
}
__device__ __forceinline__ void sts128(uint32_t addr, uint4 v) {
  asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ uint32_t lds32(uint32_t addr) {
  uint32_t r;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(r) : "r"(addr) : "memory");
  return r;
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
// instruction descriptor (cute::UMMA::InstrDescriptor): D fp32, A/B bf16, both K-major, M 128, N n_tile
static inline uint32_t um_idesc(int n_tile) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(n_tile >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
}

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

// one producer cursor: which (tile, chunk) an operation is at, and the A stage / phase it maps to
struct UmCur {
  int tile, c, stage;
  uint32_t ph;
};
__device__ __forceinline__ void um_advance(UmCur& q, int nchunks, int na) {
  if (++q.c == nchunks) {
    q.c = 0;
    q.tile += gridDim.x;
  }
  if (++q.stage == na) {
    q.stage = 0;
    q.ph ^= 1u;
  }
}

// Snake through one MUFU: x + (1/a) sin^2(a x) = x + 1/(2a) - (1/(2a)) cos(2 a x)   (istftnet.py:69)
template <int ACT>
__device__ __forceinline__ float um_act(float v, float sc, float sh, float wsc, float wsh, float hia,
                                        float slope) {
  if (ACT == ST2_ACT_SNAKE) {
    const float u2 = fmaf(v, sc, sh);     // u + 1/(2a)
    const float w2 = fmaf(v, wsc, wsh);   // 2 a u
    return fmaf(-hia, __cosf(w2), u2);
  }
  const float u = fmaf(v, sc, sh);
  if (ACT == ST2_ACT_LRELU) return fmaxf(u, u * slope);  // 0 <= slope <= 1
  return u;
}

template <int ACT, bool HAS_SS>
__device__ __forceinline__ void um_produce(const St2ConvParams& p, const UmLayout& lay, uint8_t* smem,
                                           UmMisc* misc, int ptid, int total_tiles, int n_ttiles,
                                           int n_ntiles, int nchunks) {
  const int v = ptid & 7, rbase = ptid >> 3;
  const int r_tot = UM_TM + (p.k - 1) * p.dil;
  const __nv_bfloat16* xg = reinterpret_cast<const __nv_bfloat16*>(p.x);
  const uint32_t a0 = smem_u32(smem) + (uint32_t)v * (uint32_t)(lay.rpad * 16);
  constexpr bool IDENT = (ACT == ST2_ACT_NONE) && !HAS_SS;

  auto issue_loads = [&](const UmCur& q) {
    mbar_wait(&misc->a_empty[q.stage], q.ph ^ 1u, 0);
    const UmTile tl = um_tile(q.tile, n_ttiles, n_ntiles);
    const int row0 = tl.t0 - p.pad;
    const __nv_bfloat16* xb = xg + (int64_t)tl.b * p.x_bstride + (q.c * UM_CK + v * 8);
    const uint32_t dst = a0 + (uint32_t)q.stage * (uint32_t)lay.a_bytes;
#pragma unroll
    for (int i = 0; i < UM_NI; ++i) {
      const int r = rbase + 32 * i;
      const int tin = row0 + r;
      if (r < r_tot && tin >= 0 && tin < p.Lin) cp_async16(dst + (uint32_t)r * 16u, xb + (int64_t)tin * p.ldx);
    }
  };

  auto transform = [&](const UmCur& q) {
    const UmTile tl = um_tile(q.tile, n_ttiles, n_ntiles);
    const int row0 = tl.t0 - p.pad;
    const int ch0 = q.c * UM_CK + v * 8;
    float sc[8], sh[8], wsc[8], wsh[8], hia[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      sc[e] = 1.f; sh[e] = 0.f; wsc[e] = 0.f; wsh[e] = 0.f; hia[e] = 0.f;
      if (HAS_SS) {
        const float2 t = __ldg(reinterpret_cast<const float2*>(p.ss + ((int64_t)tl.b * p.ss_stride + ch0 + e) * 2));
        sc[e] = t.x; sh[e] = t.y;
      }
      if (ACT == ST2_ACT_SNAKE) {
        const float a = __ldg(p.alpha + ch0 + e);
        hia[e] = __fdividef(0.5f, a);
        wsc[e] = 2.f * a * sc[e];
        wsh[e] = 2.f * a * sh[e];
        sh[e] += hia[e];
      }
    }
    // chunk q's LDGSTS group is complete once at most `dist` newer groups are pending
    if (lay.dist == 2) cp_async_wait<2>();
    else cp_async_wait<1>();
    const uint32_t dst = a0 + (uint32_t)q.stage * (uint32_t)lay.a_bytes;
#pragma unroll
    for (int i = 0; i < UM_NI; ++i) {
      const int r = rbase + 32 * i;
      if (r >= r_tot) continue;
      const int tin = row0 + r;
      const uint32_t addr = dst + (uint32_t)r * 16u;
      if (tin >= 0 && tin < p.Lin) {
        if (!IDENT) {
          const uint4 raw = lds128(addr);
          const uint32_t w[4] = {raw.x, raw.y, raw.z, raw.w};
          uint32_t o[4];
#pragma unroll
          for (int e2 = 0; e2 < 4; ++e2) {
            const float lo = um_act<ACT>(bf16lo(w[e2]), sc[2 * e2], sh[2 * e2], wsc[2 * e2], wsh[2 * e2],
                                         hia[2 * e2], p.slope);
            const float hi = um_act<ACT>(bf16hi(w[e2]), sc[2 * e2 + 1], sh[2 * e2 + 1], wsc[2 * e2 + 1],
                                         wsh[2 * e2 + 1], hia[2 * e2 + 1], p.slope);
            o[e2] = pack_bf16x2(lo, hi);
          }
          sts128(addr, make_uint4(o[0], o[1], o[2], o[3]));
        }
      } else {
        sts128(addr, make_uint4(0u, 0u, 0u, 0u));  // conv zero padding applies after the activation
      }
    }
    fence_proxy_async_smem();
    mbar_arrive(&misc->a_full[q.stage]);
  };

  UmCur lc{(int)blockIdx.x, 0, 0, 0u}, tc{(int)blockIdx.x, 0, 0, 0u};
  for (int d = 0; d < lay.dist; ++d) {
    if (lc.tile < total_tiles) {
      issue_loads(lc);
      um_advance(lc, nchunks, lay.na);
    }
    cp_async_commit();
  }
  while (tc.tile < total_tiles) {
    if (lc.tile < total_tiles) {
      issue_loads(lc);
      um_advance(lc, nchunks, lay.na);
    }
    cp_async_commit();
    transform(tc);
    um_advance(tc, nchunks, lay.na);
  }
  cp_async_wait<0>();
}

// epilogue output conversion + staging store of 32 consecutive columns of one row
template <bool OUT_F32>
__device__ __forceinline__ void um_stage32(uint32_t addr, const float (&o)[32]) {
  if (OUT_F32) {
#pragma unroll
    for (int g = 0; g < 8; ++g)
      sts128(addr + g * 16, make_uint4(__float_as_uint(o[4 * g]), __float_as_uint(o[4 * g + 1]),
                                       __float_as_uint(o[4 * g + 2]), __float_as_uint(o[4 * g + 3])));
  } else {
#pragma unroll
    for (int g = 0; g < 4; ++g)
      sts128(addr + g * 16, make_uint4(pack_bf16x2(o[8 * g], o[8 * g + 1]), pack_bf16x2(o[8 * g + 2], o[8 * g + 3]),
                                       pack_bf16x2(o[8 * g + 4], o[8 * g + 5]), pack_bf16x2(o[8 * g + 6], o[8 * g + 7])));
  }
}

template <bool OUT_F32>
__device__ __forceinline__ void um_epilogue(const St2ConvParams& p, const UmLayout& lay, uint8_t* smem,
                                            UmMisc* misc, uint32_t tmem_base, int warp, int lane,
                                            int total_tiles, int n_ttiles, int n_ntiles) {
  const int q = warp & 3;                          // TMEM lane group this warp may access
  const int et = threadIdx.x - UM_EPI_WARP0 * 32;  // 0..127
  const int rl = q * 32 + lane;                    // row inside an accumulator = TMEM lane
  const int n_tile = lay.n_tile, ncb = n_tile >> 5;
  const int esz = OUT_F32 ? 4 : 2;
  uint8_t* stg = smem + lay.off_stg;
  const uint32_t my_stg = smem_u32(stg) + (uint32_t)rl * (uint32_t)lay.stg_row;
  // statistics pass: ncp column pairs x (128 / ncp) row groups of ncp rows
  const int ncp = n_tile >> 1;
  const int cp = et % ncp, grp = et / ncp;
  const bool has_res = p.res != nullptr, has_acc = p.accumulate != 0;
  uint8_t* yg = reinterpret_cast<uint8_t*>(p.y);
  const uint8_t* rg = reinterpret_cast<const uint8_t*>(p.res);
  uint32_t t_it = 0;
  for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++t_it) {
    const UmTile tl = um_tile(tile, n_ttiles, n_ntiles);
    const uint32_t tb = t_it & 1, tph = (t_it >> 1) & 1;
    float* bias_s = misc->bias[tb];
    if (et < n_tile) bias_s[et] = (p.bias ? p.bias[tl.nt * n_tile + et] : 0.f) * p.out_scale;
    named_bar_sync(1, UM_EPI_THREADS);
    float s0 = 0.f, s1 = 0.f, q0 = 0.f, q1 = 0.f;

    // residual / y_old of block g = acc * ncb + cb of this thread's row (16-byte pieces of 32 bf16)
    auto row_of = [&](int acc) { return tl.t0 + acc * 128 + rl; };
    auto load_blk = [&](int g, uint4 (&rr)[4], uint4 (&yy)[4]) {
      const int acc = g / ncb, cb = g - acc * ncb;
      const int row = row_of(acc);
      if (row >= p.Lout) return;
      if (has_res) {
        const uint8_t* rp = rg + ((int64_t)tl.b * p.res_bstride + (int64_t)(row >> p.res_shift) * p.ldres +
                                  tl.nt * n_tile + cb * 32) * 2;
#pragma unroll
        for (int i = 0; i < 4; ++i) rr[i] = ldg_nc16(rp + i * 16);
      }
      if (has_acc) {
        const uint8_t* yp = yg + ((int64_t)tl.b * p.y_bstride + (int64_t)row * p.ldy + tl.nt * n_tile + cb * 32) * 2;
#pragma unroll
        for (int i = 0; i < 4; ++i) yy[i] = ldg_nc16(yp + i * 16);
      }
    };
    uint4 rr[4], yy[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) rr[i] = yy[i] = make_uint4(0u, 0u, 0u, 0u);
    if (!OUT_F32) load_blk(0, rr, yy);

    mbar_wait(&misc->t_full[tb], tph, 5);
    tc_fence_after();
#pragma unroll 1
    for (int acc = 0; acc < UM_NACC; ++acc) {
      const int row = row_of(acc);
      const bool valid = row < p.Lout;
      // the staging tile is free once this thread's bulk store has read its row and every thread has
      // finished the statistics pass over the previous contents
      bulk_wait_read0();
      if (p.stats) named_bar_sync(2, UM_EPI_THREADS);
#pragma unroll 1
      for (int cb = 0; cb < ncb; ++cb) {
        uint32_t v[32];
        tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)((tb * UM_NACC + acc) * n_tile + cb * 32), v);
        uint4 rn[4], yn[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) rn[i] = yn[i] = make_uint4(0u, 0u, 0u, 0u);
        const int gnext = acc * ncb + cb + 1;
        if (!OUT_F32 && gnext < UM_NACC * ncb) load_blk(gnext, rn, yn);
        float o[32];
#pragma unroll
        for (int i = 0; i < 32; ++i) o[i] = __uint_as_float(v[i]);
        if (!OUT_F32 && has_res) {
#pragma unroll
          for (int g = 0; g < 4; ++g) {
            const uint32_t rw[4] = {rr[g].x, rr[g].y, rr[g].z, rr[g].w};
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              o[g * 8 + 2 * e] += bf16lo(rw[e]);
              o[g * 8 + 2 * e + 1] += bf16hi(rw[e]);
            }
          }
        }
#pragma unroll
        for (int g = 0; g < 8; ++g) {
          const float4 b4 = *reinterpret_cast<const float4*>(bias_s + cb * 32 + g * 4);
          o[4 * g] = fmaf(o[4 * g], p.out_scale, b4.x);
          o[4 * g + 1] = fmaf(o[4 * g + 1], p.out_scale, b4.y);
          o[4 * g + 2] = fmaf(o[4 * g + 2], p.out_scale, b4.z);
          o[4 * g + 3] = fmaf(o[4 * g + 3], p.out_scale, b4.w);
        }
        if (!OUT_F32 && has_acc) {
#pragma unroll
          for (int g = 0; g < 4; ++g) {
            const uint32_t yw[4] = {yy[g].x, yy[g].y, yy[g].z, yy[g].w};
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              o[g * 8 + 2 * e] += bf16lo(yw[e]);
              o[g * 8 + 2 * e + 1] += bf16hi(yw[e]);
            }
          }
        }
        if (!valid) {
#pragma unroll
          for (int i = 0; i < 32; ++i) o[i] = 0.f;
        }
        um_stage32<OUT_F32>(my_stg + (uint32_t)(cb * 32 * esz), o);
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          rr[i] = rn[i];
          yy[i] = yn[i];
        }
      }
      if (acc == UM_NACC - 1) {
        // all TMEM reads of this buffer are complete (tcgen05.wait::ld inside tmem_ld32)
        tc_fence_before();
        mbar_arrive(&misc->t_empty[tb]);
      }
      // this thread's row: staging (generic proxy) -> bulk store (async proxy)
      fence_proxy_async_smem();
      if (valid) {
        uint8_t* yrow = yg + ((int64_t)tl.b * p.y_bstride + (int64_t)row * p.ldy + tl.nt * n_tile) * esz;
        bulk_s2g(yrow, stg + (size_t)rl * lay.stg_row, (uint32_t)lay.out_row);
      }
      bulk_commit();
      if (p.stats) {
        named_bar_sync(1, UM_EPI_THREADS);  // every row of the staging tile is written
        const uint32_t base = smem_u32(stg) + (uint32_t)(grp * ncp) * (uint32_t)lay.stg_row + (uint32_t)cp * 4u;
#pragma unroll 8
        for (int r = 0; r < ncp; ++r) {
          const uint32_t w = lds32(base + (uint32_t)r * (uint32_t)lay.stg_row);
          const float a = bf16lo(w), b = bf16hi(w);
          s0 += a; q0 = fmaf(a, a, q0);
          s1 += b; q1 = fmaf(b, b, q1);
        }
      }
    }
    if (p.stats) {
      // combine the row groups: sred[grp * n_tile + col]
      misc->sred[grp * n_tile + 2 * cp][0] = s0;
      misc->sred[grp * n_tile + 2 * cp][1] = q0;
      misc->sred[grp * n_tile + 2 * cp + 1][0] = s1;
      misc->sred[grp * n_tile + 2 * cp + 1][1] = q1;
      named_bar_sync(1, UM_EPI_THREADS);
      if (et < n_tile) {
        const int ngrp = 128 / ncp;
        float s = 0.f, s2 = 0.f;
        for (int gidx = 0; gidx < ngrp; ++gidx) {
          s += misc->sred[gidx * n_tile + et][0];
          s2 += misc->sred[gidx * n_tile + et][1];
        }
        float* sp = p.stats + (((int64_t)tl.b * p.n_tiles + tl.t0 / UM_TM) * p.n_cols + tl.nt * n_tile + et) * 2;
        sp[0] = s;
        sp[1] = s2;
      }
    }
  }
  bulk_wait_all0();
}

__global__ void __launch_bounds__(UM_THREADS, 1) conv_umma_kernel(const St2ConvParams p, const UmLayout lay,
                                                                   int total_tiles, int n_ttiles, int n_ntiles) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  UmMisc* misc = reinterpret_cast<UmMisc*>(smem_raw + lay.off_misc);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nchunks = p.cin / UM_CK;

  if (threadIdx.x == 0) {
    for (int i = 0; i < UM_MAX_A; ++i) {
      mbar_init(&misc->a_full[i], UM_PROD_THREADS);
      mbar_init(&misc->a_empty[i], 1);
    }
    for (int i = 0; i < UM_W_STAGES; ++i) {
      mbar_init(&misc->w_full[i], 1);
      mbar_init(&misc->w_empty[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&misc->t_full[i], 1);
      mbar_init(&misc->t_empty[i], UM_EPI_THREADS);
    }
    fence_barrier_init();
  }
  if (warp == 0) {  // TMEM: all 512 columns (this CTA owns the SM: > 114 KB smem, 1 block/SM)
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                     smem_u32(&misc->tmem_base)),
                 "r"(512)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = misc->tmem_base;

  if (warp == 0) {
    // ===================================================================== MMA issuer
    if (lane == 0) {
      const uint32_t a_base = smem_u32(smem_raw), w_base = smem_u32(smem_raw + lay.off_w);
      const uint32_t a_lbo = (uint32_t)lay.rpad * 16u, w_lbo = (uint32_t)lay.n_tile * 16u;
      int as = 0, ws = 0;
      uint32_t aph = 0, wph = 0, t_it = 0;
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++t_it) {
        const uint32_t tb = t_it & 1, tph = (t_it >> 1) & 1;
        mbar_wait(&misc->t_empty[tb], tph ^ 1, 1);
        tc_fence_after();
        for (int c = 0; c < nchunks; ++c) {
          mbar_wait(&misc->a_full[as], aph, 2);
          tc_fence_after();
          const uint32_t a_addr = a_base + (uint32_t)as * (uint32_t)lay.a_bytes;
          for (int j = 0; j < p.k; ++j) {
            mbar_wait(&misc->w_full[ws], wph, 3);
            tc_fence_after();
            const uint32_t w_addr = w_base + (uint32_t)ws * (uint32_t)lay.w_bytes;
#pragma unroll
            for (int acc = 0; acc < UM_NACC; ++acc) {
              const uint32_t d_tmem = tmem_base + (uint32_t)((tb * UM_NACC + acc) * lay.n_tile);
#pragma unroll
              for (int ks = 0; ks < UM_CK / 16; ++ks) {
                const uint64_t ad =
                    umma_desc(a_addr + (uint32_t)(2 * ks) * a_lbo + (uint32_t)(acc * 128 + j * p.dil) * 16u, a_lbo, 128);
                const uint64_t bd = umma_desc(w_addr + (uint32_t)(2 * ks) * w_lbo, w_lbo, 128);
                umma_bf16(d_tmem, ad, bd, lay.idesc, (c | j | ks) != 0 ? 1u : 0u);
              }
            }
            umma_commit(&misc->w_empty[ws]);
            if (++ws == UM_W_STAGES) { ws = 0; wph ^= 1u; }
          }
          umma_commit(&misc->a_empty[as]);
          if (++as == lay.na) { as = 0; aph ^= 1u; }
        }
        umma_commit(&misc->t_full[tb]);
      }
    }
  } else if (warp == 1) {
    // ================================================================ weight producer
    if (lane == 0) {
      const uint8_t* wg = reinterpret_cast<const uint8_t*>(p.w);
      uint8_t* wsm = smem_raw + lay.off_w;
      const int per_nt = nchunks * p.k;
      int ws = 0;
      uint32_t wph = 0;
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
        const UmTile tl = um_tile(tile, n_ttiles, n_ntiles);
        const uint8_t* src = wg + (int64_t)tl.nt * per_nt * lay.w_bytes;
        for (int s = 0; s < per_nt; ++s) {
          mbar_wait(&misc->w_empty[ws], wph ^ 1u, 4);
          mbar_arrive_expect_tx(&misc->w_full[ws], (uint32_t)lay.w_bytes);
          bulk_g2s(wsm + (size_t)ws * lay.w_bytes, src + (int64_t)s * lay.w_bytes, (uint32_t)lay.w_bytes,
                   &misc->w_full[ws]);
          if (++ws == UM_W_STAGES) { ws = 0; wph ^= 1u; }
        }
      }
    }
  } else if (warp < UM_PROD_WARP0) {
    // ======================================================================= epilogue
    if (p.y_dtype == ST2_DT_F32) um_epilogue<true>(p, lay, smem_raw, misc, tmem_base, warp, lane, total_tiles, n_ttiles, n_ntiles);
    else um_epilogue<false>(p, lay, smem_raw, misc, tmem_base, warp, lane, total_tiles, n_ttiles, n_ntiles);
  } else {
    // ============================================================ activation producers
    const int ptid = threadIdx.x - UM_PROD_WARP0 * 32;
    if (p.ss) {
      if (p.act == ST2_ACT_SNAKE) um_produce<ST2_ACT_SNAKE, true>(p, lay, smem_raw, misc, ptid, total_tiles, n_ttiles, n_ntiles, nchunks);
      else if (p.act == ST2_ACT_LRELU) um_produce<ST2_ACT_LRELU, true>(p, lay, smem_raw, misc, ptid, total_tiles, n_ttiles, n_ntiles, nchunks);
      else um_produce<ST2_ACT_NONE, true>(p, lay, smem_raw, misc, ptid, total_tiles, n_ttiles, n_ntiles, nchunks);
    } else {
      if (p.act == ST2_ACT_SNAKE) um_produce<ST2_ACT_SNAKE, false>(p, lay, smem_raw, misc, ptid, total_tiles, n_ttiles, n_ntiles, nchunks);
      else if (p.act == ST2_ACT_LRELU) um_produce<ST2_ACT_LRELU, false>(p, lay, smem_raw, misc, ptid, total_tiles, n_ttiles, n_ntiles, nchunks);
      else um_produce<ST2_ACT_NONE, false>(p, lay, smem_raw, misc, ptid, total_tiles, n_ttiles, n_ntiles, nchunks);
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

// columns per tile for a GEMM N of n_cols (0 = not representable)
int umma_n_tile(int n_cols) {
  if (n_cols <= 0) return 0;
  if (n_cols % 128 == 0) return 128;
  if (n_cols % 64 == 0) return 64;
  if (n_cols % 32 == 0) return 32;
  return 0;
}

static int um_layout(const St2ConvParams& p, UmLayout& lay) {
  lay.n_tile = umma_n_tile(p.n_cols);
  ST2_REQUIRE(lay.n_tile != 0, "conv_umma: n_cols must be a multiple of 32");
  const int halo = (p.k - 1) * p.dil;
  lay.rpad = (UM_TM + halo) | 1;  // >= rows staged, odd
  lay.a_bytes = (int)round_up((int64_t)8 * lay.rpad * 16, 128);
  lay.w_bytes = 8 * lay.n_tile * 16;
  const int esz = p.y_dtype == ST2_DT_F32 ? 4 : 2;
  lay.out_row = lay.n_tile * esz;
  lay.stg_row = lay.out_row + 16;
  const int stg_bytes = (int)round_up((int64_t)128 * lay.stg_row, 128);
  const int misc_bytes = (int)round_up((int64_t)sizeof(UmMisc), 128);
  const int fixed = UM_W_STAGES * lay.w_bytes + stg_bytes + misc_bytes;
  int na = (UM_SMEM_LIMIT - fixed) / lay.a_bytes;
  if (na > UM_MAX_A) na = UM_MAX_A;
  ST2_REQUIRE(na >= 3, "conv_umma: tile does not fit shared memory");
  lay.na = na;
  lay.dist = na - 2;
  lay.off_w = (uint32_t)(na * lay.a_bytes);
  lay.off_stg = lay.off_w + (uint32_t)(UM_W_STAGES * lay.w_bytes);
  lay.off_misc = lay.off_stg + (uint32_t)stg_bytes;
  lay.total = lay.off_misc + (uint32_t)misc_bytes;
  lay.idesc = um_idesc(lay.n_tile);
  return ST2_OK;
}

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
  ST2_REQUIRE(p.x_dtype == ST2_DT_BF16, "conv_umma: bf16 input activations only");
  ST2_REQUIRE(p.cin % UM_CK == 0 && p.cin > 0, "conv_umma: cin must be a multiple of 64");
  ST2_REQUIRE(p.stride == 1, "conv_umma: stride 1 only");
  ST2_REQUIRE(p.k >= 1 && p.dil >= 1 && (p.k - 1) * p.dil <= UM_MAX_HALO, "conv_umma: halo too wide");
  ST2_REQUIRE(p.ldx % 8 == 0 && p.x_bstride % 8 == 0, "conv_umma: 16-byte alignment of input rows");
  ST2_REQUIRE(((uintptr_t)p.x & 15) == 0 && ((uintptr_t)p.y & 15) == 0 && ((uintptr_t)p.w & 15) == 0,
              "conv_umma: 16-byte alignment of pointers");
  if (p.y_dtype == ST2_DT_F32) {
    ST2_REQUIRE(!p.res && !p.accumulate && !p.stats, "conv_umma: fp32 output takes no residual / accumulate / stats");
    ST2_REQUIRE(p.ldy % 4 == 0 && p.y_bstride % 4 == 0, "conv_umma: 16-byte alignment of fp32 output rows");
    ST2_REQUIRE(umma_n_tile(p.n_cols) != 128, "conv_umma: fp32 output needs a 32- or 64-column tile");
  } else {
    ST2_REQUIRE(p.y_dtype == ST2_DT_BF16, "conv_umma: output dtype");
    ST2_REQUIRE(p.ldy % 8 == 0 && p.y_bstride % 8 == 0, "conv_umma: 16-byte alignment of output rows");
  }
  ST2_REQUIRE(!p.res || (p.ldres % 8 == 0 && p.res_bstride % 8 == 0 && ((uintptr_t)p.res & 15) == 0),
              "conv_umma: residual alignment");
  ST2_REQUIRE(!(p.stats && p.accumulate), "conv_umma: stats with accumulate");
  ST2_REQUIRE(p.act != ST2_ACT_SNAKE || p.alpha, "conv_umma: snake needs alpha");
  ST2_REQUIRE(p.act != ST2_ACT_LRELU || (p.slope >= 0.f && p.slope <= 1.f), "conv_umma: LeakyReLU slope in [0, 1]");
  UmLayout lay;
  memset(&lay, 0, sizeof(lay));
  int rc = um_layout(p, lay);
  if (rc) return rc;
  const int n_ttiles = ceil_div(p.Lout, UM_TM), n_ntiles = p.n_cols / lay.n_tile;
  ST2_REQUIRE(!p.stats || p.n_tiles >= n_ttiles, "conv_umma: n_tiles");
  const int64_t total = (int64_t)p.B * n_ttiles * n_ntiles;
  ST2_REQUIRE(total < (1ll << 31), "conv_umma: too many tiles");
  int dev = 0;
  ST2_CUDA_OK(cudaGetDevice(&dev));
  if (dev >= 64 || g_num_sms[dev] == 0) {
    int n = 0;
    ST2_CUDA_OK(cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev));
    ST2_CUDA_OK(cudaFuncSetAttribute(conv_umma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     UM_SMEM_LIMIT));
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
  conv_umma_kernel<<<grid, UM_THREADS, lay.total, st>>>(p, lay, (int)total, n_ttiles, n_ntiles);
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

// blob order: [n tile][64-channel chunk][tap][16-byte channel group v (8)][column n (n_tile)][8 channels]
extern "C" int64_t st2_umma_packed_elems(int k, int cin_pad, int n) {
  const int nt = st2::umma_n_tile(n);
  if (nt == 0) return 0;
  return (int64_t)(n / nt) * (cin_pad / st2::UM_CK) * k * (8 * nt * 8);
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
  const int ntile = st2::umma_n_tile(n);
  if (!w_kcn || !out || k < 1 || cin < 1 || cin_pad < cin || cin_pad % st2::UM_CK || ntile == 0) {
    st2::set_error("st2_umma_pack_weights: bad shape");
    return ST2_EINVAL;
  }
  const int nch = cin_pad / st2::UM_CK, nnt = n / ntile;
  const int64_t blob_elems = (int64_t)8 * ntile * 8;
  for (int nt = 0; nt < nnt; ++nt)
    for (int c = 0; c < nch; ++c)
      for (int j = 0; j < k; ++j) {
        uint16_t* blob = out + (((int64_t)nt * nch + c) * k + j) * blob_elems;
        for (int v = 0; v < 8; ++v)
          for (int nn = 0; nn < ntile; ++nn)
            for (int e = 0; e < 8; ++e) {
              const int ch = c * st2::UM_CK + v * 8 + e;
              const float val = ch < cin ? w_kcn[((int64_t)j * cin + ch) * n + nt * ntile + nn] : 0.f;
              blob[(v * ntile + nn) * 8 + e] = f32_to_bf16_rn(val);
            }
      }
  return ST2_OK;
}
