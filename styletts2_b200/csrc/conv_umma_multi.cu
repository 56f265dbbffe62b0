// conv_umma_multi.cu -- up to three same-shaped convs as ONE persistent tcgen05 kernel (st2_conv_umma_multi).
//
// The same kernel design as conv_umma.cu (read its header first: roles, rings, layouts), with one difference: a launch
// carries an array of conv descriptions (UmJobs) and the CTAs alternate between them tile by tile (work item wi -> conv
// wi % n, tile slot wi / n); every role looks the conv's parameters (taps, dilation, weights, AdaIN parameters, residual,
// output, statistics, tensor maps) up per tile, the rings, barriers and the TMEM double buffer run straight through.
// Used by the lock-step resblock schedule of decoder.cu for SMALL problems (fewer tiles per conv than 2 x SMs: a single
// utterance), where one launch of three convs fills three times the SMs and saves two launches per chain position.
//
// Why a second translation unit instead of n = 1 of this kernel everywhere: the per-tile lookups change what ptxas
// keeps in uniform registers around the MMA issue loop, and the eight UTCHMMAs of a tap come out with 4-8 more UMOVs
// between them (compare `cuobjdump -sass` of the two objects): every k >= 7 conv ran 5-6 % slower as n = 1 of this
// kernel than on conv_umma.cu's (same box: k = 11 431 vs 402 us, C2 step 19.98 vs 18.90 ms,
// profiles/r2_conv_umma_multi_vs_single.txt), which is more than large batches gain from the overlap (k3 + k11 -6 %).
// Everything lives in namespace st2::mj; the only exported function is st2::launch_conv_umma_multi.
#include <cuda.h>
#include <stdlib.h>
#include <string.h>

#include <map>
#include <mutex>
#include <vector>

#include "common.cuh"

// compile-time A/B switches (tools build variants with -D...)
#ifndef UM_OPT_NOMASK
#define UM_OPT_NOMASK 1     // producers: no row mask in the arithmetic, out-of-range rows zeroed afterwards (edge tiles)
#endif
#ifndef UM_OPT_SNAKE4
#define UM_OPT_SNAKE4 1     // Snake with 4 parameters per channel: w = 2a (x + h) - 1 instead of its own scale / shift pair
#endif
#ifndef UM_OPT_STATS16
#define UM_OPT_STATS16 0    // epilogue statistics pass: 16 instead of 8 shared-memory loads in flight
#endif

namespace st2 {
namespace mj {

constexpr int UM_TM = 256;      // output rows per tile
constexpr int UM_NACC = 2;      // 128-row accumulators per tile
constexpr int UM_CK = 64;       // channels per A stage
constexpr int UM_MAX_A = 4;     // A ring stages (3 when the halo is wide)
constexpr int UM_W_STAGES = 3;  // default depth of the weight ring (UmLayout::nw)
constexpr int UM_MAX_W = 8;
constexpr int UM_MAX_HALO = 50; // (k-1)*dil of the widest resblock conv (k 11, d 5)
constexpr int UM_THREADS = 480;
// Warp roles, by warp id.  The SMSP arbiter prefers the HIGHEST warp id among eligible warps (B300_MICROARCH.md), so the
// roles on the critical path of the tensor pipe sit at the top: MMA issuer, weight / activation loaders, then the
// epilogue (it frees TMEM buffers); the producers, which have the most work and the most slack, at the bottom.
constexpr int UM_PROD_WARP0 = 0, UM_PROD_THREADS = 256;   // warps 0-7
constexpr int UM_EPI_WARP0 = 8, UM_EPI_THREADS = 128;     // warps 8-11 (TMEM lane quarter = warp & 3)
constexpr int UM_LOAD_WARP = 12;  // activation-tile loader (one lane), UmLayout::loader
constexpr int UM_W_WARP = 13;     // weight producer (one lane)
constexpr int UM_MMA_WARP = 14;   // MMA issuer (one lane) + TMEM allocation
constexpr int UM_NI = (UM_TM + UM_MAX_HALO + 31) / 32;  // row iterations per producer thread and chunk
constexpr int UM_SMEM_LIMIT = 227 * 1024;

struct UmLayout {
  int n_tile;         // GEMM columns per tile (128 / 64 / 32)
  int a_bytes;        // bytes per A stage (rows x 128 B, multiple of 1024)
  int na;             // A stages in use
  int nw;             // weight stages in use
  int pair;           // 1 = two CTAs of a cluster run one M = 256 cta_group::2 MMA per step: each CTA feeds its own
                      // 256-row tile (two 128-row accumulators) and HALF of the weight columns, so the weight stream
                      // per SM and the B-operand reads of shared memory halve.  Same N tile in both, any (item, t0).
  int n_bt;           // pair mode: B * n_ttiles
  int total_ptiles;   // pair mode: n_ntiles * ceil(n_bt / 2)
  uint32_t idesc_pair;  // M = 256
  int dist;           // prefetch distance of the producers, in chunks (na - 2)
  uint32_t spin;      // ST2_UMMA_SPIN: barrier classes that are polled instead of waited on with a suspend hint:
                      // 1 t_empty (MMA warp), 2 a_full (MMA warp / relay), 4 w_full (MMA warp), 8 w_empty (weight
                      // producer), 16 t_full (epilogue); 32 = the MMA warp's loop runs on its elected lane only
  int dbg;            // ST2_UMMA_DBG timing experiments, all with wrong results: 1 reuse the weight stages without
                      // reloading, 2 producers skip the prologue arithmetic, 4 epilogue skips the statistics pass,
                      // 8 epilogue skips TMEM read / staging / store, 16 no conv MMAs are issued (barriers only)
  int w_bytes;        // bytes per weight blob (= 8 * n_tile * 16)
  int w_slot;         // bytes per weight stage in shared memory (= w_bytes; half of it in pair mode)
  int out_row;        // bytes per output row segment (n_tile * element size), multiple of 128
  int n_res, n_acc;   // extra 64-channel chunks per tile: residual rows, y_old rows
  int res_tma, acc_tma;  // those rows come by TMA tensor loads (one thread) instead of 256 threads' cp.async
  int x_tma;          // the raw activation tile too: one 256-row box + ceil(halo / 8) boxes of 8 rows
  int loader;         // every chunk type comes by TMA: one lane of warp UM_LOAD_WARP issues all activation / residual /
                      // y_old loads as soon as the MMAs release a stage (up to `na` chunks ahead) and the 256 producer
                      // threads only wait for the bytes (raw_full), transform and arrive; 0 = the producers also load
                      // (cp.async fallback for unaligned views and the x2-repeated shortcut rows)
  uint32_t nt_m, nt_l, tt_m, tt_l;  // multiply-shift division by n_ntiles / n_ttiles (um_tile)
  const uint8_t* ident_res;  // n_tile/64 identity weight blobs (x 1)
  const uint8_t* ident_acc;  // n_tile/64 identity weight blobs (x 1/out_scale)
  uint32_t off_a, off_w, off_misc, total;  // the staging tile sits at offset 0 (1024-byte aligned)
  uint32_t idesc;
  uint32_t idesc64;   // the same with N = 64 (identity chunks of a 128-column tile)
};

struct UmMisc {
  uint64_t a_full[UM_MAX_A], a_empty[UM_MAX_A];
  uint64_t raw_full[UM_MAX_A];  // raw activation tile of the stage has landed (TMA transaction bytes)
  uint64_t w_full[UM_MAX_W], w_empty[UM_MAX_W];
  uint64_t t_full[2], t_empty[2];
  uint64_t peer_a_full[UM_MAX_A];  // pair mode, leader CTA: the peer's A stage is ready
  uint32_t tmem_base, pad_;
  // bias through the tensor core: one K=16 step with A = ones_a (every row (1, 1, 0, ...)) and
  // B = bias_b[tb] (per column (hi, lo, 0, ...) bf16 split of the fp32 bias), issued first for each tile
  alignas(128) uint4 bias_b[2][128];
  alignas(128) uint4 ones_a[16];   // core matrix k 0..7 = rows of (1, 1, 0, 0, 0, 0, 0, 0); k 8..15 = zeros
  alignas(16) float sred[256][2];
};

// Up to UM_MAX_JOBS convs of the same geometry (batch, length, input / output channels, tile) run as ONE persistent
// kernel whose CTAs alternate between them tile by tile (work item wi -> job wi % n, tile slot wi / n): a conv whose
// tiles are bound by the SIMT roles (k = 3: prologue + epilogue per element, 3 taps of MMAs) then overlaps one whose
// tiles are bound by the tensor pipe (k = 11), through the same A / W rings and TMEM double buffer that already
// overlap consecutive tiles of one conv.  Everything that differs between the jobs (taps, dilation, weights, AdaIN
// parameters, residual, output, statistics) is looked up per tile; the shared-memory geometry is common.
constexpr int UM_MAX_JOBS = 3;
struct alignas(64) UmJob {
  CUtensorMap ymap, rmap, xmap, xmap8;
  St2ConvParams p;
  UmLayout lay;
};
struct alignas(64) UmJobs {
  UmJob j[UM_MAX_JOBS];
  CUtensorMap wmap, wmap_res, wmap_acc;  // pair mode (single job only)
  int n;
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
__device__ __forceinline__ uint64_t global_timer_ns() {
  uint64_t t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
// try_wait carries a suspend-time hint, so a waiting thread sleeps in hardware instead of spinning
// through issue slots the working warps need; the 4 s deadline is checked every 64 wake-ups.  poll != 0 polls
// test_wait instead: no suspension, so no wake-up latency (UmLayout::spin selects it per barrier class).
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity, int which, uint32_t poll = 0u) {
  const uint32_t addr = smem_u32(bar);
  uint32_t done = 0, spin = 0;
  uint64_t t0 = 0;
  while (true) {
    if (poll)
      asm volatile("{ .reg .pred p; mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                   : "=r"(done)
                   : "r"(addr), "r"(parity)
                   : "memory");
    else
      asm volatile(
          "{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3; selp.u32 %0, 1, 0, p; }"
          : "=r"(done)
          : "r"(addr), "r"(parity), "r"(0x989680u)
          : "memory");
    if (done) break;
    if ((++spin & (poll ? 0xFFFFu : 63u)) == 0u) {
      const uint64_t now = global_timer_ns();
      if (t0 == 0) t0 = now;
      else if (now - t0 > 4000000000ull) mbar_timeout(which, parity);
    }
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
// smem -> global TMA tensor store (3-D map: column, row, batch), tracked by the issuing thread's
// bulk async-group; rows / columns outside the tensor are clipped by the hardware
__device__ __forceinline__ void tma_store_3d(const CUtensorMap* map, uint32_t smem_src, int c0, int c1, int c2) {
  asm volatile("cp.async.bulk.tensor.3d.global.shared::cta.bulk_group [%0, {%2, %3, %4}], [%1];" ::"l"(map),
               "r"(smem_src), "r"(c0), "r"(c1), "r"(c2)
               : "memory");
}
// global -> smem TMA tensor load (3-D map), completion reported as transaction bytes on an mbarrier;
// elements outside the tensor are zero-filled
__device__ __forceinline__ void tma_load_3d(const CUtensorMap* map, uint32_t smem_dst, int c0, int c1, int c2,
                                            uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];" ::"r"(
          smem_dst),
      "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2)
      : "memory");
}
// the same from a CTA of a pair: the bytes are counted on a barrier of either CTA (shared::cluster address)
__device__ __forceinline__ void tma_load_3d_pair(const CUtensorMap* map, uint32_t smem_dst, int c0, int c1, int c2,
                                                 uint32_t bar_cluster_addr) {
  asm volatile(
      "cp.async.bulk.tensor.3d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];" ::"r"(
          smem_dst),
      "l"(map), "r"(bar_cluster_addr), "r"(c0), "r"(c1), "r"(c2)
      : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.expect_tx.relaxed.cta.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
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
// branch-free variants for the issuing warp: executed by all lanes, performed by the one with on != 0
__device__ __forceinline__ void umma_commit_if(uint64_t* bar, uint32_t on) {
  asm volatile(
      "{ .reg .pred q; setp.ne.b32 q, %1, 0;\n"
      "@q tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0]; }" ::"r"(smem_u32(bar)),
      "r"(on)
      : "memory");
}
__device__ __forceinline__ void umma_bf16_if(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                             uint32_t accumulate, uint32_t on) {
  asm volatile(
      "{ .reg .pred p, q; setp.ne.b32 p, %4, 0; setp.ne.b32 q, %5, 0;\n"
      "@q tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p; }" ::"r"(tmem_d),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate), "r"(on)
      : "memory");
}
// One tap of a 64-channel chunk in a single asm block: 2 accumulators x 4 K steps.  Issued one by one, every MMA
// costs the issuing thread a fresh transfer of all its operands into uniform registers (ELECT + 7 R2UR + vote loop in
// the SASS), which is longer than the 64 cycles the MMA itself runs; inside one block the descriptors advance with
// plain adds on values that are already there.
#define ST2_UMMA_TAP8(GROUP)                                                                                 \
  asm volatile(                                                                                              \
      "{ .reg .pred q, t; .reg .b64 a, b, a1; .reg .b32 d1;\n"                                               \
      "setp.ne.b32 q, %5, 0; setp.eq.b32 t, 0, 0;\n"                                                         \
      "add.u32 d1, %0, %6; add.u64 a1, %1, 1024;\n"                                                          \
      "@q tcgen05.mma.cta_group::" GROUP ".kind::f16 [%0], %1, %2, %3, t;\n"                                 \
      "@q tcgen05.mma.cta_group::" GROUP ".kind::f16 [d1], a1, %2, %3, t;\n"                                 \
      "add.u64 a, %1, 2; add.u64 a1, a1, 2; add.u64 b, %2, %4;\n"                                            \
      "@q tcgen05.mma.cta_group::" GROUP ".kind::f16 [%0], a, b, %3, t;\n"                                   \
      "@q tcgen05.mma.cta_group::" GROUP ".kind::f16 [d1], a1, b, %3, t;\n"                                  \
      "add.u64 a, a, 2; add.u64 a1, a1, 2; add.u64 b, b, %4;\n"                                              \
      "@q tcgen05.mma.cta_group::" GROUP ".kind::f16 [%0], a, b, %3, t;\n"                                   \
      "@q tcgen05.mma.cta_group::" GROUP ".kind::f16 [d1], a1, b, %3, t;\n"                                  \
      "add.u64 a, a, 2; add.u64 a1, a1, 2; add.u64 b, b, %4;\n"                                              \
      "@q tcgen05.mma.cta_group::" GROUP ".kind::f16 [%0], a, b, %3, t;\n"                                   \
      "@q tcgen05.mma.cta_group::" GROUP ".kind::f16 [d1], a1, b, %3, t; }" ::"r"(d),                         \
      "l"(ad), "l"(bd), "r"(idesc), "l"((uint64_t)w_kstep), "r"(on), "r"(acc_cols)                           \
      : "memory")
__device__ __forceinline__ void umma_tap8_if(uint32_t d, uint32_t acc_cols, uint64_t ad, uint64_t bd, uint32_t w_kstep,
                                             uint32_t idesc, uint32_t on) {
  ST2_UMMA_TAP8("1");
}
__device__ __forceinline__ void umma2_tap8_if(uint32_t d, uint32_t acc_cols, uint64_t ad, uint64_t bd, uint32_t w_kstep,
                                              uint32_t idesc, uint32_t on) {
  ST2_UMMA_TAP8("2");
}
// ---- CTA-pair variants (cta_group::2): issued by the leader CTA of a 2-CTA cluster; the commit arrives on the
// barrier at the same shared-memory offset in both CTAs
__device__ __forceinline__ void umma2_commit_if(uint64_t* bar, uint32_t on) {
  asm volatile(
      "{ .reg .pred q; .reg .b16 m; setp.ne.b32 q, %1, 0; mov.b16 m, 3;\n"
      "@q tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], m; }" ::"r"(
          smem_u32(bar)),
      "r"(on)
      : "memory");
}
__device__ __forceinline__ void umma2_bf16_if(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                              uint32_t accumulate, uint32_t on) {
  asm volatile(
      "{ .reg .pred p, q; setp.ne.b32 p, %4, 0; setp.ne.b32 q, %5, 0;\n"
      "@q tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p; }" ::"r"(tmem_d),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate), "r"(on)
      : "memory");
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// shared::cluster address of `addr` (a shared::cta address of this CTA) in CTA `cta` of the cluster
__device__ __forceinline__ uint32_t mapa_cta(uint32_t addr, uint32_t cta) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(addr), "r"(cta));
  return r;
}
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_addr) {
  asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
// one non-blocking probe of an mbarrier phase
__device__ __forceinline__ uint32_t mbar_try(uint64_t* bar, uint32_t parity) {
  uint32_t done;
  asm volatile(
      "{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
      : "=r"(done)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return done;
}
// K-major shared-memory matrix descriptor (cute::UMMA::SmemDescriptor, version 1):
// [0,14) start>>4, [16,30) leading (K-direction core-matrix) byte offset>>4,
// [32,46) stride (8-row group) byte offset>>4, [46,48) version=1, [49,52) base offset,
// [61,64) layout type (0 = no swizzle, 2 = SWIZZLE_128B).
__device__ __forceinline__ uint64_t umma_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes,
                                              uint32_t layout_type = 0) {
  uint64_t d = (uint64_t)((saddr >> 4) & 0x3FFF);
  d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16;
  d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)layout_type << 61;
  return d;
}
// instruction descriptor (cute::UMMA::InstrDescriptor): D fp32, A/B bf16, both K-major, M 128, N n_tile
static inline uint32_t um_idesc(int n_tile, int m = 128) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(n_tile >> 3) << 17) | ((uint32_t)(m >> 4) << 24);
}

__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
// empty asm that "modifies" the loaded registers: their uses cannot be scheduled above the wait
__device__ __forceinline__ void tmem_ld_fence(uint32_t (&v)[32]) {
  asm volatile(""
               : "+r"(v[0]), "+r"(v[1]), "+r"(v[2]), "+r"(v[3]), "+r"(v[4]), "+r"(v[5]), "+r"(v[6]), "+r"(v[7]),
                 "+r"(v[8]), "+r"(v[9]), "+r"(v[10]), "+r"(v[11]), "+r"(v[12]), "+r"(v[13]), "+r"(v[14]),
                 "+r"(v[15]), "+r"(v[16]), "+r"(v[17]), "+r"(v[18]), "+r"(v[19]), "+r"(v[20]), "+r"(v[21]),
                 "+r"(v[22]), "+r"(v[23]), "+r"(v[24]), "+r"(v[25]), "+r"(v[26]), "+r"(v[27]), "+r"(v[28]),
                 "+r"(v[29]), "+r"(v[30]), "+r"(v[31])
               :
               : "memory");
}
__device__ __forceinline__ void tmem_ld32_nowait(uint32_t taddr, uint32_t (&v)[32]) {
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
}
// one lane of a converged warp (elect.sync)
__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile("{ .reg .pred p; elect.sync _|p, 0xffffffff; selp.u32 %0, 1, 0, p; }" : "=r"(pred));
  return pred != 0;
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
// packed fp32x2 arithmetic (sm_100: FADD2 / FFMA2, one issue slot for two lanes of math)
__device__ __forceinline__ uint64_t f2_pack(uint32_t lo, uint32_t hi) {
  uint64_t r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "r"(lo), "r"(hi));
  return r;
}
__device__ __forceinline__ void f2_unpack(uint64_t v, float& lo, float& hi) {
  asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ uint64_t f2_add(uint64_t a, uint64_t b) {
  uint64_t d;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
  return d;
}
__device__ __forceinline__ uint64_t f2_fma(uint64_t a, uint64_t b, uint64_t c) {
  uint64_t d;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
  return d;
}
// bf16x2 word -> its two values as a packed fp32x2 (low half first)
__device__ __forceinline__ uint64_t f2_from_bf16x2(uint32_t w) { return f2_pack(w << 16, w & 0xFFFF0000u); }
__device__ __forceinline__ float bf16lo(uint32_t u) { return __uint_as_float(u << 16); }
__device__ __forceinline__ float bf16hi(uint32_t u) { return __uint_as_float(u & 0xFFFF0000u); }

// ------------------------------------------------------------------------- kernel
// ST2_UMMA_DBG bit 64: per-role wait accounting (SM clock cycles summed over all CTAs; st2_umma_wait_profile):
//   [0] MMA thread: whole loop   [1] waiting on t_empty (TMEM buffer held by the epilogue)   [2] on a_full (activation
//   stage not transformed yet)   [3] on w_full (weight blob not landed)   [4] tiles
//   [5] epilogue thread 0: whole loop   [6] waiting on t_full (accumulator not complete)   [7] staging tile not free
//   (previous TMA store still reading + barrier)   [8] statistics pass
//   [9] producer thread 0: whole loop   [10] waiting on a_empty (stage still feeding MMAs)   [11] on raw_full (TMA
//   bytes not landed)   [12] prologue arithmetic
__device__ unsigned long long g_um_wait[16];
__device__ __forceinline__ long long um_clk(bool on) { return on ? clock64() : 0ll; }

struct UmTile {
  int b, t0, nt;
};
// n / d for n < 2^31 with the host-computed (m, l) of um_fastdiv: no integer-division sequence in the
// per-tile bookkeeping of 448 threads
__device__ __forceinline__ uint32_t um_div(uint32_t n, uint32_t m, uint32_t l) {
  if (l == 0) return n;
  const uint32_t t = __umulhi(n, m);
  return (t + ((n - t) >> 1)) >> (l - 1);
}
#define um_tile(tile, n_ttiles, n_ntiles) um_tile_(tile, n_ttiles, n_ntiles, lay)
__device__ __forceinline__ UmTile um_tile_(int tile, int n_ttiles, int n_ntiles, const UmLayout& lay) {
  UmTile t;
  const uint32_t bt = um_div((uint32_t)tile, lay.nt_m, lay.nt_l);
  t.nt = tile - (int)bt * n_ntiles;
  const uint32_t b = um_div(bt, lay.tt_m, lay.tt_l);
  t.t0 = (int)(bt - b * (uint32_t)n_ttiles) * UM_TM;
  t.b = (int)b;
  return t;
}

// the it-th tile of this CTA (total_tiles = none left).  Pair mode: the cluster takes pair-tile q = (nt, btp) and its
// CTA of rank r the (item, time) tile 2 btp + r, clamped to the last one (an odd count makes both CTAs compute the
// same tile: identical bytes stored twice).
// Work items: the (conv, tile) pairs of a launch are numbered g = tile * n + conv and CTA b takes g = b, b + grid,
// b + 2 grid, ... -- with a grid of min(n * tiles, SMs) a small launch (a single utterance: 32 tiles x 3 convs) spreads
// over n times the SMs, a large one still alternates between the convs (grid % n != 0 rotates them).
__device__ __forceinline__ int64_t um_gid(int wi) { return (int64_t)blockIdx.x + (int64_t)wi * gridDim.x; }
__device__ __forceinline__ int64_t um_slot_of(int n, int64_t g) { return n == 1 ? g : (n == 2 ? g >> 1 : g / 3); }
__device__ __forceinline__ int um_job_of(int n, int wi) {
  if (n == 1) return 0;
  const int64_t g = um_gid(wi);
  return (int)(g - um_slot_of(n, g) * n);
}
#define um_tile_at(it, total_tiles, n_ntiles) um_tile_at_<PAIR>(J.n, it, total_tiles, n_ntiles, lay)
template <bool PAIR>
__device__ __forceinline__ int um_tile_at_(int n, int it, int total_tiles, int n_ntiles, const UmLayout& lay) {
  if (!PAIR) {
    const int64_t t = um_slot_of(n, um_gid(it));
    return t < total_tiles ? (int)t : total_tiles;
  }
  const int64_t q = (int64_t)(blockIdx.x >> 1) + (int64_t)it * (gridDim.x >> 1);  // single conv only
  if (q >= lay.total_ptiles) return total_tiles;
  const uint32_t btp = um_div((uint32_t)q, lay.nt_m, lay.nt_l);
  const int nt = (int)q - (int)btp * n_ntiles;
  int bt = 2 * (int)btp + (int)(blockIdx.x & 1u);
  if (bt >= lay.n_bt) bt = lay.n_bt - 1;
  return bt * n_ntiles + nt;
}

struct UmCur {
  int tile, c, stage;
  uint32_t ph;
  int b, t0, nt;
  int it;   // work item index
  int job;
};
// Order of a tile's chunks in the A ring.  nh "heavy" chunks (64 input channels: k taps of MMAs, prologue arithmetic)
// and nl "light" ones (64 residual / y_old columns: one identity tap, no arithmetic) are interleaved H L H L ...,
// the rest of the longer list last.  A chunk reuses the stage of the chunk `na` positions earlier, so it can only be
// loaded once that one's MMAs are done; with the light chunks after all the heavy ones (3 stages: H H L L | H ...) the
// first heavy chunk of the next tile had to wait for the LAST heavy chunk of this tile, and its load + prologue
// (~3-4 k cycles) were exposed with only two identity taps of MMA work to hide behind (measured: +3.7 k cycles per tile).
struct UmSlot {
  bool heavy;
  int idx;
};
__device__ __forceinline__ UmSlot um_slot(int c, int nh, int nl) {
  const int m = nh < nl ? nh : nl;
  UmSlot s;
  if (c < 2 * m) {
    s.heavy = !(c & 1);
    s.idx = c >> 1;
  } else {
    s.heavy = nh > nl;
    s.idx = c - m;
  }
  return s;
}
#define um_cur_tile(q, total_tiles, n_ttiles, n_ntiles) um_cur_tile_(q, total_tiles, n_ttiles, n_ntiles, lay)
#define um_advance(q, nchunks, na, total_tiles, n_ttiles, n_ntiles) um_advance_<PAIR>(q, nchunks, na, total_tiles, n_ttiles, n_ntiles, lay, J.n)
__device__ __forceinline__ void um_cur_tile_(UmCur& q, int total_tiles, int n_ttiles, int n_ntiles, const UmLayout& lay) {
  if (q.tile < total_tiles) {
    const UmTile t = um_tile(q.tile, n_ttiles, n_ntiles);
    q.b = t.b; q.t0 = t.t0; q.nt = t.nt;
  }
}
template <bool PAIR>
__device__ __forceinline__ void um_advance_(UmCur& q, int nchunks, int na, int total_tiles, int n_ttiles,
                                            int n_ntiles, const UmLayout& lay, int njobs) {
  if (++q.c == nchunks) {
    q.c = 0;
    ++q.it;
    q.job = um_job_of(njobs, q.it);
    q.tile = um_tile_at_<PAIR>(njobs, q.it, total_tiles, n_ntiles, lay);
    um_cur_tile(q, total_tiles, n_ttiles, n_ntiles);
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
#if UM_OPT_SNAKE4
    const float w2 = fmaf(wsc, u2, -1.f); // wsc = 2a: 2a (u + 1/(2a)) - 1 = 2 a u
#else
    const float w2 = fmaf(v, wsc, wsh);   // 2 a u
#endif
    return fmaf(-hia, __cosf(w2), u2);
  }
  const float u = fmaf(v, sc, sh);
  if (ACT == ST2_ACT_LRELU) return fmaxf(u, u * slope);  // 0 <= slope <= 1
  return u;
}

template <int ACT, bool HAS_SS, bool PAIR>
__device__ __forceinline__ void um_produce(const UmJobs& J, uint8_t* smem,
                                           UmMisc* misc, int ptid, int total_tiles, int n_ttiles,
                                           int n_ntiles, int nchunks) {
  const int v = ptid & 7, rbase = ptid >> 3;
  const UmLayout& lay = J.j[0].lay;  // the shared-memory geometry and the switches are common to all jobs
  // chunks per tile of a job, incl. its residual / y_old rows
  auto nct_of = [&](int job) { return nchunks + J.j[job].lay.n_res + J.j[job].lay.n_acc; };
  // this thread's 16-byte slot of row rbase in stage 0; row rbase + 32 i is i * 4096 bytes further
  // (same row & 7, hence the same swizzled slot)
  const uint32_t a0 = smem_u32(smem + lay.off_a) + (uint32_t)rbase * 128u + (uint32_t)((v ^ (rbase & 7)) << 4);
  constexpr bool IDENT = (ACT == ST2_ACT_NONE) && !HAS_SS;
  // per-job view inside the lambdas below (the names shadow the common ones on purpose)
#define UM_JOB_VIEW(q)                                                                                   \
  const UmJob& jb = J.j[(q).job];                                                                        \
  const St2ConvParams& p = jb.p;                                                                         \
  const UmLayout& lay = jb.lay;                                                                          \
  const CUtensorMap *ymap = &jb.ymap, *rmap = &jb.rmap, *xmap = &jb.xmap, *xmap8 = &jb.xmap8;            \
  (void)ymap; (void)rmap; (void)xmap; (void)xmap8;                                                       \
  const int r_tot = UM_TM + (p.k - 1) * p.dil, n_light = lay.n_res + lay.n_acc;                          \
  (void)r_tot; (void)n_light
  const bool prof = (lay.dbg & 64) && ptid == 0;
  long long w_raw = 0, w_math = 0, w_empty_c = 0;
  const long long c_start = um_clk(prof);

  // the caller has seen a_empty[q.stage] complete for this round of the ring
  auto issue_loads = [&](const UmCur& q) {
    UM_JOB_VIEW(q);
    const __nv_bfloat16* xg = reinterpret_cast<const __nv_bfloat16*>(p.x);
    const int64_t x_step = (int64_t)32 * p.ldx;
    const uint32_t dst = a0 + (uint32_t)q.stage * (uint32_t)lay.a_bytes;
    const UmSlot sl = um_slot(q.c, nchunks, n_light);
    if (sl.heavy && lay.x_tma) {
      // tile row r holds input row t0 - pad + r: one 256-row box and the halo in 8-row boxes, all
      // zero-filled outside [0, Lin).  A conv without prologue arithmetic (IDENT) hands the bytes
      // straight to the MMA barrier; otherwise the producers wait for them on raw_full.
      if (ptid == 0) {
        uint64_t* bar = IDENT ? &misc->a_full[q.stage] : &misc->raw_full[q.stage];
        const uint32_t base = smem_u32(smem + lay.off_a) + (uint32_t)q.stage * (uint32_t)lay.a_bytes;
        const int nh = (r_tot - UM_TM + 7) >> 3;
        const int row0 = q.t0 - p.pad, col = sl.idx * UM_CK;
        const uint32_t bytes = (uint32_t)(UM_TM * 128 + nh * 1024);
        if (IDENT) mbar_expect_tx(bar, bytes);      // next to the 256 producer arrivals
        else mbar_arrive_expect_tx(bar, bytes);     // raw_full: this is its only arrival
        tma_load_3d(xmap, base, col, row0, q.b, bar);
        for (int i = 0; i < nh; ++i)
          tma_load_3d(xmap8, base + (uint32_t)(UM_TM * 128 + i * 1024), col, row0 + UM_TM + 8 * i, q.b, bar);
      }
    } else if (sl.heavy) {
      // tile row r holds input row t0 - pad + r; rows outside [0, Lin) stay untouched (zeroed later)
      const int row0 = q.t0 - p.pad + rbase;
      const int lo = -row0, hi = min(r_tot - rbase, p.Lin - row0);  // valid: lo <= 32 i < hi
      const __nv_bfloat16* xb = xg + (int64_t)q.b * p.x_bstride + (int64_t)row0 * p.ldx + (sl.idx * UM_CK + v * 8);
#pragma unroll
      for (int i = 0; i < UM_NI; ++i)
        if (32 * i >= lo && 32 * i < hi) cp_async16(dst + (uint32_t)i * 4096u, xb + i * x_step);
    } else {
      // residual rows (row t >> res_shift of res) or y_old rows (row t of y): 64 columns of this N tile
      const int e = sl.idx;
      const bool is_res = e < lay.n_res;
      const int hh = is_res ? e : e - lay.n_res;
      // keep raw_full[stage] in step with the ring: this use of the stage has no raw tile
      if (lay.x_tma && !IDENT && ptid == 0) mbar_arrive(&misc->raw_full[q.stage]);
      if (is_res ? lay.res_tma : lay.acc_tma) {
        // two 128-row boxes in the SWIZZLE_128B pattern of the A tile, zero-filled past Lout; the bytes
        // are expected on a_full[stage] next to the producers' 256 arrivals
        if (ptid == 0) {
          const uint32_t base = smem_u32(smem + lay.off_a) + (uint32_t)q.stage * (uint32_t)lay.a_bytes;
          const int col = q.nt * lay.n_tile + hh * UM_CK;
          const bool two = q.t0 + 128 < p.Lout;
          mbar_expect_tx(&misc->a_full[q.stage], two ? 32768u : 16384u);
          const CUtensorMap* mp = is_res ? rmap : ymap;
          tma_load_3d(mp, base, col, q.t0, q.b, &misc->a_full[q.stage]);
          if (two) tma_load_3d(mp, base + 16384u, col, q.t0 + 128, q.b, &misc->a_full[q.stage]);
        }
        return;
      }
      const __nv_bfloat16* sb;
      int ld, shift;
      if (is_res) {
        sb = reinterpret_cast<const __nv_bfloat16*>(p.res) + (int64_t)q.b * p.res_bstride;
        ld = p.ldres; shift = p.res_shift;
      } else {
        sb = reinterpret_cast<const __nv_bfloat16*>(p.y) + (int64_t)q.b * p.y_bstride;
        ld = p.ldy; shift = 0;
      }
      sb += q.nt * lay.n_tile + hh * UM_CK + v * 8;
      const int t = q.t0 + rbase;
#pragma unroll
      for (int i = 0; i < UM_TM / 32; ++i)
        if (t + 32 * i < p.Lout) cp_async16(dst + (uint32_t)i * 4096u, sb + (int64_t)((t + 32 * i) >> shift) * ld);
    }
  };

  // raw per-channel parameters of the chunk (scale, shift pairs and alpha), loaded one iteration early
  auto load_params = [&](const UmCur& q, float2 (&ssr)[8], float (&alr)[8]) {
    UM_JOB_VIEW(q);
    const UmSlot sl = um_slot(q.c, nchunks, n_light);
    if (!sl.heavy || q.tile >= total_tiles) return;
    const int ch0 = sl.idx * UM_CK + v * 8;
    if (HAS_SS) {
      const float4* sp = reinterpret_cast<const float4*>(p.ss + ((int64_t)q.b * p.ss_stride + ch0) * 2);
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const float4 t = __ldg(sp + e);
        ssr[2 * e] = make_float2(t.x, t.y);
        ssr[2 * e + 1] = make_float2(t.z, t.w);
      }
    }
    if (ACT == ST2_ACT_SNAKE) {
      const float4* ap = reinterpret_cast<const float4*>(p.alpha + ch0);
      const float4 t0 = __ldg(ap), t1 = __ldg(ap + 1);
      alr[0] = t0.x; alr[1] = t0.y; alr[2] = t0.z; alr[3] = t0.w;
      alr[4] = t1.x; alr[5] = t1.y; alr[6] = t1.z; alr[7] = t1.w;
    }
  };

  // `newer` = LDGSTS groups committed after chunk q's: q's group is complete once at most that many
  // are pending
  auto transform = [&](const UmCur& q, const float2 (&ssr)[8], const float (&alr)[8], int newer) {
    UM_JOB_VIEW(q);
    const uint32_t dst = a0 + (uint32_t)q.stage * (uint32_t)lay.a_bytes;
    const UmSlot sl = um_slot(q.c, nchunks, n_light);
    if (lay.loader) {
      // the loader lane has armed this use of the stage (expect_tx on raw_full or a_full) before raw_full completes
      const long long c0 = um_clk(prof);
      mbar_wait(&misc->raw_full[q.stage], q.ph, 6);
      w_raw += um_clk(prof) - c0;
      if (!sl.heavy || IDENT) {  // bytes go straight to a_full; the TMA zero-fill is the conv's zero padding
        mbar_arrive(&misc->a_full[q.stage]);
        return;
      }
    } else {
      if (newer >= 3) cp_async_wait<3>();
      else if (newer == 2) cp_async_wait<2>();
      else if (newer == 1) cp_async_wait<1>();
      else cp_async_wait<0>();
    }
    if (!sl.heavy) {
      const bool tma = sl.idx < lay.n_res ? lay.res_tma : lay.acc_tma;
      if (!tma) {
        const int t = q.t0 + rbase;
#pragma unroll
        for (int i = 0; i < UM_TM / 32; ++i)
          if (t + 32 * i >= p.Lout) sts128(dst + (uint32_t)i * 4096u, make_uint4(0u, 0u, 0u, 0u));
        fence_proxy_async_smem();
      }
      // TMA rows: nothing to do (rows past Lout of a half-loaded tile are never stored or counted)
      mbar_arrive(&misc->a_full[q.stage]);
      return;
    }
    if (lay.x_tma && !lay.loader) {
      if (IDENT) {  // the TMA zero-fill is the conv's zero padding: nothing to touch
        mbar_arrive(&misc->a_full[q.stage]);
        return;
      }
      const long long c0 = um_clk(prof);
      mbar_wait(&misc->raw_full[q.stage], q.ph, 6);
      w_raw += um_clk(prof) - c0;
    }
    const long long c_ar = um_clk(prof);
    if (lay.dbg & 2) {
      mbar_arrive(&misc->a_full[q.stage]);
      return;
    }
    const int row0 = q.t0 - p.pad + rbase;
    const int lo = -row0, hi = min(r_tot - rbase, p.Lin - row0), tot = r_tot - rbase;
    float sc[8], sh[8], wsc[8], wsh[8], hia[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      sc[e] = 1.f; sh[e] = 0.f; wsc[e] = 0.f; wsh[e] = 0.f; hia[e] = 0.f;
      if (HAS_SS) {
        sc[e] = ssr[e].x; sh[e] = ssr[e].y;
      }
      if (ACT == ST2_ACT_SNAKE) {
        const float a = alr[e];
        hia[e] = __fdividef(0.5f, a);
#if UM_OPT_SNAKE4
        wsc[e] = a + a;
#else
        wsc[e] = 2.f * a * sc[e];
        wsh[e] = 2.f * a * sh[e];
#endif
        sh[e] += hia[e];
      }
    }
    // Rows in batches of 4 (+ a tail of 2 halo rows): all shared-memory reads of a batch are issued
    // before its arithmetic, and the arithmetic is branch-free (one basic block per batch: 32
    // independent chains for the scheduler to interleave behind the MUFU latency).  Rows outside
    // [0, Lin) are computed on whatever the slot holds and masked to zero (the conv's zero padding
    // applies after the activation).  Only the tail rows i >= UM_TM/32 can lie outside the tile:
    // the tail is skipped by a warp-uniform test and its rows are predicated.
    constexpr int NB = UM_TM / 32 / 4;  // full batches
#pragma unroll
    for (int h = 0; h <= NB; ++h) {
      const int i0 = h * 4, cnt = h < NB ? 4 : UM_NI - 4 * NB;
      if (h == NB && 32 * i0 + 4 * (ptid >> 5) >= r_tot) break;  // no row of this warp reaches the tail
      uint4 raw[4];
      if (!IDENT) {
#pragma unroll
        for (int ii = 0; ii < 4; ++ii) {
          const int i = i0 + ii;
          raw[ii] = make_uint4(0u, 0u, 0u, 0u);
          if (ii < cnt && (h < NB || 32 * i < tot)) raw[ii] = lds128(dst + (uint32_t)i * 4096u);
        }
      }
#pragma unroll
      for (int ii = 0; ii < 4; ++ii) {
        if (ii >= cnt) continue;
        const int i = i0 + ii;
        const bool in_tile = h < NB || 32 * i < tot;
        const uint32_t addr = dst + (uint32_t)i * 4096u;
#if UM_OPT_NOMASK
        if (IDENT) continue;  // rows outside [0, Lin) are zeroed below
#else
        const uint32_t mask = (32 * i >= lo && 32 * i < hi) ? 0xFFFFFFFFu : 0u;
        if (IDENT) {
          if (in_tile && mask == 0u) sts128(addr, make_uint4(0u, 0u, 0u, 0u));
          continue;
        }
#endif
        const uint32_t w[4] = {raw[ii].x, raw[ii].y, raw[ii].z, raw[ii].w};
        uint32_t o[4];
#pragma unroll
        for (int e2 = 0; e2 < 4; ++e2) {
          const float lo_ = um_act<ACT>(bf16lo(w[e2]), sc[2 * e2], sh[2 * e2], wsc[2 * e2], wsh[2 * e2],
                                        hia[2 * e2], p.slope);
          const float hi_ = um_act<ACT>(bf16hi(w[e2]), sc[2 * e2 + 1], sh[2 * e2 + 1], wsc[2 * e2 + 1],
                                        wsh[2 * e2 + 1], hia[2 * e2 + 1], p.slope);
#if UM_OPT_NOMASK
          o[e2] = pack_bf16x2(lo_, hi_);
#else
          o[e2] = pack_bf16x2(lo_, hi_) & mask;
#endif
        }
        if (in_tile) sts128(addr, make_uint4(o[0], o[1], o[2], o[3]));
      }
    }
#if UM_OPT_NOMASK
    // the conv's zero padding applies after the activation: rows outside [0, Lin) become zeros (first / last tiles
    // of an item only; a branch no thread of an interior tile takes).  Measured: -2..8 % on every conv shape.
    if (lo > 0 || hi < tot) {
#pragma unroll 1
      for (int i = 0; 32 * i < tot; ++i)
        if (32 * i < lo || 32 * i >= hi) sts128(dst + (uint32_t)i * 4096u, make_uint4(0u, 0u, 0u, 0u));
    }
#endif
    fence_proxy_async_smem();
    mbar_arrive(&misc->a_full[q.stage]);
    w_math += um_clk(prof) - c_ar;
  };

  // Two cursors over the same chunk sequence: lc = next chunk to load, tc = next chunk to transform.
  // Loads are issued opportunistically: before and after every transform each thread probes the
  // a_empty barrier of the next slot and starts that chunk's loads as soon as the MMAs have released
  // it, up to na-1 chunks ahead.  (Residual / y_old chunks cost no arithmetic, so a fixed distance
  // counted in chunks would leave the following activation chunk no time to arrive.)
  UmCur lc{um_tile_at(0, total_tiles, n_ntiles), 0, 0, 0u, 0, 0, 0, 0, um_job_of(J.n, 0)};
  um_cur_tile(lc, total_tiles, n_ttiles, n_ntiles);
  UmCur tc = lc;
  int n_l = 0, n_t = 0;  // chunks whose loads are issued / that are transformed
  auto try_issue = [&](bool block) {
    if (lay.loader) return;  // loads are the loader lane's job
    if (lay.dist > 0) {  // fixed prefetch distance (A/B switch ST2_UMMA_DIST): blocking, dist chunks ahead
      while (lc.tile < total_tiles && n_l - n_t < lay.dist + (block ? 1 : 0)) {
        const long long c0 = um_clk(prof);
        mbar_wait(&misc->a_empty[lc.stage], lc.ph ^ 1u, 0);
        w_empty_c += um_clk(prof) - c0;
        issue_loads(lc);
        cp_async_commit();
        um_advance(lc, nct_of(lc.job), lay.na, total_tiles, n_ttiles, n_ntiles);
        ++n_l;
      }
      return;
    }
    while (lc.tile < total_tiles && n_l - n_t < lay.na - 1 + (block ? 1 : 0)) {
      if (block && n_l == n_t) mbar_wait(&misc->a_empty[lc.stage], lc.ph ^ 1u, 0);
      else if (!mbar_try(&misc->a_empty[lc.stage], lc.ph ^ 1u)) break;
      issue_loads(lc);
      cp_async_commit();
      um_advance(lc, nct_of(lc.job), lay.na, total_tiles, n_ttiles, n_ntiles);
      ++n_l;
      block = false;
    }
  };
  float2 ssr[8];
  float alr[8];
#pragma unroll
  for (int e = 0; e < 8; ++e) {
    ssr[e] = make_float2(1.f, 0.f);
    alr[e] = 1.f;
  }
  load_params(tc, ssr, alr);
  while (tc.tile < total_tiles) {
    try_issue(n_l == n_t);  // blocks only when the chunk to transform has not been requested yet
    // next chunk's parameters: their global loads are in flight during this chunk's arithmetic
    UmCur nx = tc;
    um_advance(nx, nct_of(nx.job), lay.na, total_tiles, n_ttiles, n_ntiles);
    float2 ssn[8];
    float aln[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      ssn[e] = ssr[e];
      aln[e] = alr[e];
    }
    load_params(nx, ssn, aln);
    transform(tc, ssr, alr, n_l - n_t - 1);
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      ssr[e] = ssn[e];
      alr[e] = aln[e];
    }
    tc = nx;
    ++n_t;
    try_issue(false);
  }
  cp_async_wait<0>();
  if (prof) {
    atomicAdd(&g_um_wait[9], (unsigned long long)(clock64() - c_start));
    atomicAdd(&g_um_wait[10], (unsigned long long)w_empty_c);
    atomicAdd(&g_um_wait[11], (unsigned long long)w_raw);
    atomicAdd(&g_um_wait[12], (unsigned long long)w_math);
  }
#undef UM_JOB_VIEW
}

// fp32 bias -> (hi, lo, 0, ...) bf16: one 16-byte K group of the bias operand
__device__ __forceinline__ uint4 um_bias_split(float b) {
  const __nv_bfloat16 hi = __float2bfloat16_rn(b);
  const __nv_bfloat16 lo = __float2bfloat16_rn(b - __bfloat162float(hi));
  return make_uint4((uint32_t)__bfloat16_as_ushort(hi) | ((uint32_t)__bfloat16_as_ushort(lo) << 16), 0u, 0u, 0u);
}

// Staging tile: 128 rows x out_row bytes as out_row/128 column segments of [128 rows][128 B], each in
// the TMA SWIZZLE_128B pattern (16-byte chunk index ^ (row & 7)), so that row-per-thread 16-byte
// stores and the column reads of the statistics pass are both bank-conflict free.
__device__ __forceinline__ uint32_t um_stg_addr(uint32_t stg, int row, int byte_in_row) {
  const int seg = byte_in_row >> 7, chunk = (byte_in_row >> 4) & 7;
  return stg + (uint32_t)seg * 16384u + (uint32_t)row * 128u + (uint32_t)((chunk ^ (row & 7)) << 4) +
         (uint32_t)(byte_in_row & 15);
}

// scale one 32-column block of this thread's row and write it to the staging tile.
// chunk_addr[j] = address of 16-byte chunk j of this row in column segment 0 (swizzle applied).
template <bool OUT_F32>
__device__ __forceinline__ void um_stage32(const uint32_t (&chunk_addr)[8], int cb, const uint32_t (&v)[32],
                                           float scale) {
  float o[32];
  if (scale == 1.f) {  // (warp-uniform) most convs: no multiply at all
#pragma unroll
    for (int g = 0; g < 32; ++g) o[g] = __uint_as_float(v[g]);
  } else {
#pragma unroll
    for (int g = 0; g < 32; ++g) o[g] = __uint_as_float(v[g]) * scale;  // the bias is already in the accumulator
  }
  if (OUT_F32) {  // 32 columns = one 128-byte segment
    const uint32_t seg = (uint32_t)cb * 16384u;
#pragma unroll
    for (int g = 0; g < 8; ++g)
      sts128(chunk_addr[g] + seg,
             make_uint4(__float_as_uint(o[4 * g]), __float_as_uint(o[4 * g + 1]), __float_as_uint(o[4 * g + 2]),
                        __float_as_uint(o[4 * g + 3])));
  } else {        // 32 columns = half a segment
    const uint32_t seg = (uint32_t)(cb >> 1) * 16384u;
    const int half = (cb & 1) * 4;
#pragma unroll
    for (int g = 0; g < 4; ++g)
      sts128((half ? chunk_addr[4 + g] : chunk_addr[g]) + seg,
             make_uint4(pack_bf16x2(o[8 * g], o[8 * g + 1]), pack_bf16x2(o[8 * g + 2], o[8 * g + 3]),
                        pack_bf16x2(o[8 * g + 4], o[8 * g + 5]), pack_bf16x2(o[8 * g + 6], o[8 * g + 7])));
  }
}

template <bool OUT_F32, bool PAIR>
__device__ __forceinline__ void um_epilogue(const UmJobs& J, uint8_t* smem, UmMisc* misc, uint32_t tmem_base, int warp,
                                            int lane, int total_tiles, int n_ttiles, int n_ntiles) {
  const UmLayout& lay = J.j[0].lay;  // tile geometry and switches are common to all jobs
  const int q = warp & 3;                          // TMEM lane group this warp may access
  const int et = threadIdx.x - UM_EPI_WARP0 * 32;  // 0..127
  const int rl = q * 32 + lane;                    // row inside an accumulator = TMEM lane
  const int n_tile = lay.n_tile, ncb = n_tile >> 5;
  const uint32_t stg = smem_u32(smem);
  const int nseg = lay.out_row >> 7;               // 128-byte column segments per row
  const int seg_cols = OUT_F32 ? 32 : 64;
  // statistics pass: ncp column pairs x (128 / ncp) row groups of ncp rows
  const int ncp = n_tile >> 1;
  const int cp = et % ncp, grp = et / ncp;
  uint32_t chunk_addr[8];   // this row's 16-byte chunks in segment 0
  uint32_t col_addr[8];     // statistics pass: (row grp*ncp + i, column pair cp), i = 0..7
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    chunk_addr[j] = um_stg_addr(stg, rl, j * 16);
    col_addr[j] = um_stg_addr(stg, grp * ncp + j, cp * 4);
  }
  // pair mode: this CTA owns half of the bias operand's columns, and the TMEM buffer goes back to the leader CTA
  const int nb_cols = PAIR ? n_tile >> 1 : n_tile;
  const int b_col0 = PAIR ? (int)(blockIdx.x & 1u) * nb_cols : 0;
  const uint32_t t_empty0 = PAIR ? mapa_cta(smem_u32(&misc->t_empty[0]), 0u) : smem_u32(&misc->t_empty[0]);
  auto release_tmem = [&](uint32_t tb) {
    if (PAIR) mbar_arrive_cluster(t_empty0 + tb * 8u);
    else mbar_arrive(&misc->t_empty[tb]);
  };
  uint32_t t_it = 0;
  const bool prof = (lay.dbg & 64) && et == 0;
  long long w_tfull = 0, w_stg = 0, w_stats = 0;
  const long long c_start = um_clk(prof);
  for (int it = 0;; ++it, ++t_it) {
    const int tile = um_tile_at(it, total_tiles, n_ntiles);
    if (tile >= total_tiles) break;
    const UmTile tl = um_tile(tile, n_ttiles, n_ntiles);
    const UmJob& jb = J.j[um_job_of(J.n, it)];  // this tile's conv
    const St2ConvParams& p = jb.p;
    const CUtensorMap* ymap = &jb.ymap;
    const float scale = p.out_scale;
    const uint32_t tb = t_it & 1, tph = (t_it >> 1) & 1;
    // bias of the tile that will use this TMEM buffer next (two work items ahead), loaded early
    const int tile2 = um_tile_at(it + 2, total_tiles, n_ntiles);
    const St2ConvParams& p2 = J.j[um_job_of(J.n, it + 2)].p;
    float bias2 = 0.f;
    if (et < nb_cols && tile2 < total_tiles && p2.bias) {
      const UmTile t2 = um_tile(tile2, n_ttiles, n_ntiles);
      bias2 = __ldg(p2.bias + (int64_t)t2.b * p2.bias_bstride + t2.nt * n_tile + b_col0 + et);
    }
    // two independent accumulator sets (even / odd rows) halve the dependent FADD/FFMA chains
    // (sum, sum of squares) of this thread's two columns as packed fp32x2
    uint64_t sa = 0ull, qa = 0ull, sb = 0ull, qb = 0ull;
    {
      const long long c0 = um_clk(prof);
      mbar_wait(&misc->t_full[tb], tph, 5, lay.spin & 16u);
      w_tfull += um_clk(prof) - c0;
    }
    tc_fence_after();
    // this tile's bias MMA has completed: its operand slot can take the next bias; the t_empty arrive
    // below (after the fence) hands it to the MMA warp together with the TMEM buffer
    if (et < nb_cols) misc->bias_b[tb][et] = um_bias_split(bias2);
    if (lay.dbg & 8) {
      fence_proxy_async_smem();
      tc_fence_before();
      release_tmem(tb);
      continue;
    }
#pragma unroll 1
    for (int acc = 0; acc < UM_NACC; ++acc) {
      // the staging tile is free once the previous TMA store has read it (issued by thread 0) and
      // every thread has finished its statistics pass
      const long long c_s0 = um_clk(prof);
      if (et == 0) bulk_wait_read0();
      named_bar_sync(2, UM_EPI_THREADS);
      w_stg += um_clk(prof) - c_s0;
      const uint32_t tcol = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)((tb * UM_NACC + acc) * n_tile);
      if (ncb == 1) {
        uint32_t v[32];
        tmem_ld32_nowait(tcol, v);
        tmem_ld_wait();
        tmem_ld_fence(v);
        um_stage32<OUT_F32>(chunk_addr, 0, v, scale);
      } else {
#pragma unroll 1
        for (int cb = 0; cb < ncb; cb += 2) {
          uint32_t v0[32], v1[32];
          tmem_ld32_nowait(tcol + (uint32_t)(cb * 32), v0);
          tmem_ld32_nowait(tcol + (uint32_t)(cb * 32 + 32), v1);
          tmem_ld_wait();
          tmem_ld_fence(v0);
          tmem_ld_fence(v1);
          um_stage32<OUT_F32>(chunk_addr, cb, v0, scale);
          um_stage32<OUT_F32>(chunk_addr, cb + 1, v1, scale);
        }
      }
      // staging and bias_b (generic proxy) -> TMA store / next bias MMA (async proxy)
      fence_proxy_async_smem();
      if (acc == UM_NACC - 1) {
        // all TMEM reads of this buffer are complete (tcgen05.wait::ld above)
        tc_fence_before();
        release_tmem(tb);
      }
      named_bar_sync(1, UM_EPI_THREADS);
      const int row0 = tl.t0 + acc * 128;
      if (et == 0 && row0 < p.Lout) {
        for (int sgi = 0; sgi < nseg; ++sgi)
          tma_store_3d(ymap, stg + (uint32_t)sgi * 16384u, tl.nt * n_tile + sgi * seg_cols, row0, tl.b);
        bulk_commit();
      }
      const long long c_st0 = um_clk(prof);
      if (!OUT_F32 && p.stats && !(lay.dbg & 4)) {
        const int r1 = p.Lout - row0 - grp * ncp;  // valid rows of this thread's row group
        if (r1 >= ncp) {
          // rows grp*ncp + 8 m + i sit 1024 m bytes after col_addr[i] (the swizzle repeats every 8 rows)
#if UM_OPT_STATS16
#pragma unroll 2
#else
#pragma unroll 1
#endif
          for (int m = 0; m < ncp; m += 8) {
#pragma unroll
            for (int i = 0; i < 8; i += 2) {
              const uint32_t w = lds32(col_addr[i] + (uint32_t)m * 128u);
              const uint32_t w2 = lds32(col_addr[i + 1] + (uint32_t)m * 128u);
              const uint64_t x = f2_from_bf16x2(w), x2 = f2_from_bf16x2(w2);
              sa = f2_add(sa, x); qa = f2_fma(x, x, qa);
              sb = f2_add(sb, x2); qb = f2_fma(x2, x2, qb);
            }
          }
        } else {
          for (int r = 0; r < r1; ++r) {
            const uint32_t w = lds32(col_addr[r & 7] + (uint32_t)(r >> 3) * 1024u);
            const uint64_t x = f2_from_bf16x2(w);
            sa = f2_add(sa, x); qa = f2_fma(x, x, qa);
          }
        }
      }
      w_stats += um_clk(prof) - c_st0;
    }
    if (!OUT_F32 && p.stats) {
      // combine the row groups: sred[grp * n_tile + col] (reads of the previous tile's sums finished
      // before that tile's threads passed the barriers above)
      float s0, s1, q0, q1;
      f2_unpack(f2_add(sa, sb), s0, s1);
      f2_unpack(f2_add(qa, qb), q0, q1);
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
  if (et == 0) bulk_wait_all0();
  if (prof) {
    atomicAdd(&g_um_wait[5], (unsigned long long)(clock64() - c_start));
    atomicAdd(&g_um_wait[6], (unsigned long long)w_tfull);
    atomicAdd(&g_um_wait[7], (unsigned long long)w_stg);
    atomicAdd(&g_um_wait[8], (unsigned long long)w_stats);
  }
}

template <bool PAIR>
__global__ void __launch_bounds__(UM_THREADS, 1) conv_umma_kernel(const __grid_constant__ UmJobs J, int total_tiles,
                                                                   int n_ttiles, int n_ntiles) {
  const St2ConvParams& p = J.j[0].p;   // what all jobs share: shapes, channel counts, activation kind, dtypes
  const UmLayout& lay = J.j[0].lay;    // shared-memory geometry, ring depths, switches
  const CUtensorMap& wmap = J.wmap;
  const CUtensorMap& wmap_res = J.wmap_res;
  const CUtensorMap& wmap_acc = J.wmap_acc;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  // launched with the programmatic-serialization attribute (common.cuh).  Waiting after the CTA-local setup
  // (barrier init, TMEM allocation) instead of here made no measurable difference (same-box A/B, C2 step).
  pdl_wait_then_trigger();
  UmMisc* misc = reinterpret_cast<UmMisc*>(smem_raw + lay.off_misc);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nchunks = p.cin / UM_CK;

  if (threadIdx.x == 0) {
    for (int i = 0; i < J.n; ++i) asm volatile("prefetch.tensormap [%0];" ::"l"(&J.j[i].ymap) : "memory");
    for (int i = 0; i < UM_MAX_A; ++i) {
      mbar_init(&misc->a_full[i], UM_PROD_THREADS);
      mbar_init(&misc->a_empty[i], 1);
      mbar_init(&misc->raw_full[i], 1);
    }
    for (int i = 0; i < lay.nw; ++i) {
      mbar_init(&misc->w_full[i], 1);
      mbar_init(&misc->w_empty[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&misc->t_full[i], 1);
      mbar_init(&misc->t_empty[i], PAIR ? 2 * UM_EPI_THREADS : UM_EPI_THREADS);  // pair: both CTAs' epilogues
    }
    if (PAIR) {
      for (int i = 0; i < UM_MAX_A; ++i) mbar_init(&misc->peer_a_full[i], 1);
    }
    fence_barrier_init();
  }
  {  // operands of the bias MMA: ones_a, and bias_b of this CTA's first two tiles
    const int et0 = (int)threadIdx.x - UM_EPI_WARP0 * 32;
    if (et0 >= 0 && et0 < UM_EPI_THREADS) {
      if (et0 < 16) misc->ones_a[et0] = make_uint4(et0 < 8 ? 0x3F803F80u : 0u, 0u, 0u, 0u);
      const int nb_cols = PAIR ? lay.n_tile >> 1 : lay.n_tile;
      const int b_col0 = PAIR ? (int)(blockIdx.x & 1u) * nb_cols : 0;
      if (et0 < nb_cols) {
        for (int i = 0; i < 2; ++i) {
          const int tile = um_tile_at(i, total_tiles, n_ntiles);
          const St2ConvParams& pi = J.j[um_job_of(J.n, i)].p;
          float b = 0.f;
          if (tile < total_tiles && pi.bias) {
            const UmTile t = um_tile(tile, n_ttiles, n_ntiles);
            b = __ldg(pi.bias + (int64_t)t.b * pi.bias_bstride + t.nt * lay.n_tile + b_col0 + et0);
          }
          misc->bias_b[i][et0] = um_bias_split(b);
        }
      }
      fence_proxy_async_smem();
    }
  }
  if (warp == UM_MMA_WARP) {  // TMEM: all 512 columns (this CTA owns the SM: > 114 KB smem, 1 block/SM)
    if (PAIR) {  // one warp of each CTA of the pair
      asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                       smem_u32(&misc->tmem_base)),
                   "r"(512)
                   : "memory");
      asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
    } else {
      asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                       smem_u32(&misc->tmem_base)),
                   "r"(512)
                   : "memory");
      asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
  }
  tc_fence_before();
  if (PAIR) cluster_sync_all();  // the peer's barriers are initialised before any remote arrive / multicast commit
  else __syncthreads();
  tc_fence_after();
  const uint32_t rank = PAIR ? (blockIdx.x & 1u) : 0u;  // %cluster_ctarank of a (2, 1, 1) cluster
  const uint32_t tmem_base = misc->tmem_base;

  if (warp == UM_MMA_WARP && rank != 0) {
    // ============================================================ pair mode, peer CTA: relay
    // The leader's MMA warp cannot wait on this CTA's barriers; this warp does, and forwards every completed
    // a_full phase to the leader's peer_a_full (same ring positions in both CTAs).  Weight stages need no relay: the
    // peer's TMA loads are counted on the leader's w_full directly.
    const int nct = nchunks + lay.n_res + lay.n_acc;
    const uint32_t pa = mapa_cta(smem_u32(&misc->peer_a_full[0]), 0u);
    int as = 0;
    uint32_t aph = 0;
    if (!(lay.spin & 32u) || lane == 0)
    for (int it = 0; um_tile_at(it, total_tiles, n_ntiles) < total_tiles; ++it) {
      for (int c = 0; c < nct; ++c) {
        mbar_wait(&misc->a_full[as], aph, 2, lay.spin & 2u);
        fence_proxy_async_smem();
        if (lane == 0) mbar_arrive_cluster(pa + (uint32_t)as * 8u);
        if (++as == lay.na) { as = 0; aph ^= 1u; }
      }
    }
  } else if (warp == UM_MMA_WARP) {
    // ===================================================================== MMA issuer
    // The whole warp runs the (warp-uniform) loop and waits on the barriers; one elected lane issues
    // the MMAs and commits.  Descriptors are built once; per MMA only the 14-bit start-address field
    // moves (16-byte units): A (SWIZZLE_128B rows of 128 B) + 2 per 16-channel K step, + 8 per row
    // (128 rows per accumulator, dil rows per tap); W (core matrices) + 2*n_tile per K step.
    const uint32_t a_base = smem_u32(smem_raw + lay.off_a), w_base = smem_u32(smem_raw + lay.off_w);
    // pair mode: this CTA's weight stage holds 64 of the 128 columns
    const uint32_t w_lbo = PAIR ? 1024u : (uint32_t)lay.n_tile * 16u;
    const uint64_t a_desc0 = umma_desc(0, 16, 1024, 2), w_desc0 = umma_desc(0, w_lbo, 128);
    const uint32_t a_kstep = 2u, w_kstep = w_lbo >> 3;  // in 16-byte units (two K groups)
    const uint32_t idesc_main = PAIR ? lay.idesc_pair : lay.idesc;
    auto mma = [&](uint32_t d, uint64_t ad, uint64_t bd, uint32_t id, uint32_t accum, uint32_t on) {
      if (PAIR) umma2_bf16_if(d, ad, bd, id, accum, on);
      else umma_bf16_if(d, ad, bd, id, accum, on);
    };
    auto commit = [&](uint64_t* bar, uint32_t on) {
      if (!PAIR && (lay.dbg & 32)) {  // timing experiment (with bit 16): a plain arrive instead of tcgen05.commit
        if (on) mbar_arrive(bar);
        return;
      }
      if (PAIR) umma2_commit_if(bar, on);
      else umma_commit_if(bar, on);
    };
    // the loop runs on ONE elected lane: the other lanes wait at the end of the kernel.  (With all lanes in the
    // loop every tcgen05.mma becomes a vote / ELECT / R2UR loop over the active lanes in the SASS.)
    const uint32_t leader = 1u;
    const uint32_t mma_on = (lay.dbg & 16) ? 0u : 1u;
    int as = 0, ws = 0;
    uint32_t aph = 0, wph = 0, t_it = 0;
    // The tensor-core queue is shallow, so everything between two groups of MMAs is on the critical
    // path: the next weight stage's barrier is probed BEFORE the current group is issued (its
    // latency hides behind the issue), and issue / commit are predicated instead of branched.
    uint32_t w_ready = 0;
    const bool prof = (lay.dbg & 64) != 0;
    long long w_te = 0, w_af = 0, w_wf = 0, n_tl = 0;
    const long long c_start = um_clk(prof);
    // ST2_UMMA_SPIN bit 32: only the elected lane runs the loop (an mbarrier probe by 32 lanes costs the warp
    // hundreds of cycles per tap)
    if (elect_one()) {
    for (int it = 0; um_tile_at(it, total_tiles, n_ntiles) < total_tiles; ++it, ++t_it) {
      const uint32_t tb = t_it & 1, tph = (t_it >> 1) & 1;
      {
        const long long c0 = um_clk(prof);
        mbar_wait(&misc->t_empty[tb], tph ^ 1, 1, lay.spin & 1u);
        w_te += um_clk(prof) - c0;
        ++n_tl;
      }
      const uint32_t d0 = tmem_base + (uint32_t)(tb * UM_NACC * lay.n_tile);
      tc_fence_after();
      {  // D = ones x bias: A rows all alias one core matrix (SBO 0), B's second K group aliases the first (LBO 0, x 0)
        const uint64_t od = umma_desc(smem_u32(misc->ones_a), 128, 0);
        const uint64_t bd = umma_desc(smem_u32(misc->bias_b[tb]), 0, 128);
#pragma unroll
        for (int acc = 0; acc < UM_NACC; ++acc)
          mma(d0 + (uint32_t)(acc * lay.n_tile), od, bd, idesc_main, 0u, leader);
      }
      // this tile's conv: taps, dilation and residual / running-mean chunks
      const St2ConvParams& pj = J.j[um_job_of(J.n, it)].p;
      const UmLayout& lj = J.j[um_job_of(J.n, it)].lay;
      const int nct = nchunks + lj.n_res + lj.n_acc;
      for (int c = 0; c < nct; ++c) {
        {
          const long long c0 = um_clk(prof);
          mbar_wait(&misc->a_full[as], aph, 2, lay.spin & 2u);
          if (PAIR) mbar_wait(&misc->peer_a_full[as], aph, 7, lay.spin & 2u);
          w_af += um_clk(prof) - c0;
        }
        const uint32_t a_u = (a_base + (uint32_t)as * (uint32_t)lay.a_bytes) >> 4;
        // residual / y_old chunks: one identity "tap" at row offset 0
        const UmSlot sl = um_slot(c, nchunks, lj.n_res + lj.n_acc);
        const int taps = sl.heavy ? pj.k : 1;
        const uint32_t dil = sl.heavy ? (uint32_t)pj.dil : 0u;
        for (int j = 0; j < taps; ++j) {
          if (!w_ready) {
            const long long c0 = um_clk(prof);
            mbar_wait(&misc->w_full[ws], wph, 3, lay.spin & 4u);
            w_wf += um_clk(prof) - c0;
          }
          if (!(lay.spin & 64u)) tc_fence_after();
          int wn = ws + 1;
          uint32_t wnph = wph;
          if (wn == lay.nw) { wn = 0; wnph ^= 1u; }
          w_ready = mbar_try(&misc->w_full[wn], wnph);  // consumed after the MMAs below
          const uint32_t shift = (uint32_t)j * dil;  // rows
          // the swizzle is a function of the absolute smem address: a start that is not a multiple of
          // 8 rows needs no base offset (measured: base-offset variants fail the parity tests)
          const uint64_t ad = a_desc0 + (uint64_t)(a_u + shift * 8u);
          uint64_t bd = w_desc0 + (uint64_t)((w_base + (uint32_t)ws * (uint32_t)lay.w_slot) >> 4);
          // identity chunk e of a 128-column tile only touches columns [64 e', 64 e' + 64): a 64-column MMA
          // (pair mode: the identity blob over all 128 columns, each CTA holding its half)
          uint32_t idesc = idesc_main, dcol = 0;
          if (!sl.heavy && lay.n_tile == 128 && !PAIR) {
            const int e = sl.idx;
            const uint32_t hh = (uint32_t)(e < lj.n_res ? e : e - lj.n_res);
            idesc = lay.idesc64;
            dcol = hh * 64u;
            bd += (uint64_t)(hh * 64u);  // 64 columns x 16 bytes, in 16-byte units
          }
          static_assert(UM_NACC == 2 && UM_CK == 64, "umma_tap8_if issues 2 accumulators x 4 K steps");
          if (PAIR) umma2_tap8_if(d0 + dcol, (uint32_t)lay.n_tile, ad, bd, w_kstep, idesc, mma_on);
          else umma_tap8_if(d0 + dcol, (uint32_t)lay.n_tile, ad, bd, w_kstep, idesc, mma_on);
          commit(&misc->w_empty[ws], leader);
          ws = wn;
          wph = wnph;
        }
        commit(&misc->a_empty[as], leader);
        if (++as == lay.na) { as = 0; aph ^= 1u; }
      }
      commit(&misc->t_full[tb], leader);
    }
    if (prof) {
      atomicAdd(&g_um_wait[0], (unsigned long long)(clock64() - c_start));
      atomicAdd(&g_um_wait[1], (unsigned long long)w_te);
      atomicAdd(&g_um_wait[2], (unsigned long long)w_af);
      atomicAdd(&g_um_wait[3], (unsigned long long)w_wf);
      atomicAdd(&g_um_wait[4], (unsigned long long)n_tl);
    }
    }
  } else if (warp == UM_W_WARP) {
    // ================================================================ weight producer
    if (elect_one()) {
      uint8_t* wsm = smem_raw + lay.off_w;
      const uint32_t wf0 = PAIR ? mapa_cta(smem_u32(&misc->w_full[0]), 0u) : 0u;
      int ws = 0;
      uint32_t wph = 0;
      for (int it = 0;; ++it) {
        const int tile = um_tile_at(it, total_tiles, n_ntiles);
        if (tile >= total_tiles) break;
        const UmTile tl = um_tile(tile, n_ttiles, n_ntiles);
        // this tile's conv: its blobs, tap count and identity blocks
        const St2ConvParams& p = J.j[um_job_of(J.n, it)].p;
        const UmLayout& lay = J.j[um_job_of(J.n, it)].lay;
        const int per_nt = nchunks * p.k;
        const uint8_t* src = reinterpret_cast<const uint8_t*>(p.w) + (int64_t)tl.nt * per_nt * lay.w_bytes;
        // blobs in the order the MMA thread consumes them: chunk slots (um_slot), k taps per heavy chunk, one
        // identity blob per light chunk
        const int n_light = lay.n_res + lay.n_acc;
        int c = 0, j = 0;
        for (int sq = 0; sq < per_nt + n_light; ++sq) {
          const UmSlot sl = um_slot(c, nchunks, n_light);
          const int s = sl.heavy ? sl.idx * p.k + j : per_nt + sl.idx;
          if (!sl.heavy || ++j == p.k) { ++c; j = 0; }
          const uint8_t* blob = src + (int64_t)s * lay.w_bytes;
          if (s >= per_nt) {  // identity blocks that add the residual / y_old rows
            const int e = s - per_nt;
            blob = e < lay.n_res ? lay.ident_res + (int64_t)e * lay.w_bytes
                                 : lay.ident_acc + (int64_t)(e - lay.n_res) * lay.w_bytes;
          }
          mbar_wait(&misc->w_empty[ws], wph ^ 1u, 4, lay.spin & 8u);
          if ((lay.dbg & 1) && (it != 0 || sq >= lay.nw)) {
            mbar_arrive(&misc->w_full[ws]);  // timing experiment only
          } else if (PAIR) {
            // this CTA's 64 columns of each of the 8 K groups of the [K group][128 columns][16 B] blob: one TMA box,
            // counted on the LEADER's barrier next to the leader's own half (no hand-over through this CTA)
            const CUtensorMap* mp = &wmap;
            int row = (tl.nt * per_nt + s) * 8;
            if (s >= per_nt) {
              const int e = s - per_nt;
              mp = e < lay.n_res ? &wmap_res : &wmap_acc;
              row = (e < lay.n_res ? e : e - lay.n_res) * 8;
            }
            if (rank == 0) mbar_arrive_expect_tx(&misc->w_full[ws], (uint32_t)lay.w_bytes);
            tma_load_3d_pair(mp, smem_u32(wsm + (size_t)ws * lay.w_slot), 0, 64 * (int)rank, row, wf0 + (uint32_t)ws * 8u);
          } else {
            mbar_arrive_expect_tx(&misc->w_full[ws], (uint32_t)lay.w_bytes);
            bulk_g2s(wsm + (size_t)ws * lay.w_slot, blob, (uint32_t)lay.w_bytes, &misc->w_full[ws]);
          }
          if (++ws == lay.nw) { ws = 0; wph ^= 1u; }
        }
      }
    }
  } else if (warp == UM_LOAD_WARP) {
    // ======================================================= activation / residual loader (UmLayout::loader)
    // One lane walks the chunk sequence of this CTA's tiles and issues every chunk's TMA loads the moment the MMAs
    // have released its stage (a_empty), up to `na` chunks ahead of them; no producer thread ever waits to load.
    //   heavy chunk with prologue arithmetic: bytes counted on raw_full[stage] (its only arrival); the producers wait
    //     for it, transform in place and arrive on a_full
    //   light chunk / heavy chunk without arithmetic: bytes counted on a_full[stage] next to the producers' 256
    //     arrivals; raw_full gets a plain arrive AFTER the expect_tx, and the producers arrive on a_full only after
    //     they have seen raw_full, so the phase cannot complete before its bytes are expected
    if (lay.loader && elect_one()) {
      const bool ident = p.act == ST2_ACT_NONE && p.ss == nullptr;  // the same for every job (host-checked)
      UmCur q{um_tile_at(0, total_tiles, n_ntiles), 0, 0, 0u, 0, 0, 0, 0, um_job_of(J.n, 0)};
      um_cur_tile(q, total_tiles, n_ttiles, n_ntiles);
      while (q.tile < total_tiles) {
        mbar_wait(&misc->a_empty[q.stage], q.ph ^ 1u, 0);
        // this chunk's conv: halo, padding, tensor maps, residual / running-mean chunks
        const UmJob& jb = J.j[q.job];
        const St2ConvParams& p = jb.p;
        const UmLayout& lay = jb.lay;
        const int n_light = lay.n_res + lay.n_acc, nct = nchunks + n_light;
        const int r_tot = UM_TM + (p.k - 1) * p.dil;
        const int nhb = (r_tot - UM_TM + 7) >> 3;
        const uint32_t hbytes = (uint32_t)(UM_TM * 128 + nhb * 1024);
        const UmSlot sl = um_slot(q.c, nchunks, n_light);
        const uint32_t base = smem_u32(smem_raw + lay.off_a) + (uint32_t)q.stage * (uint32_t)lay.a_bytes;
        if (sl.heavy) {
          // tile row r holds input row t0 - pad + r: one 256-row box and the halo in 8-row boxes, zero-filled
          // outside [0, Lin)
          uint64_t* bar = ident ? &misc->a_full[q.stage] : &misc->raw_full[q.stage];
          const int row0 = q.t0 - p.pad, col = sl.idx * UM_CK;
          if (ident) mbar_expect_tx(bar, hbytes);
          else mbar_arrive_expect_tx(bar, hbytes);
          tma_load_3d(&jb.xmap, base, col, row0, q.b, bar);
          for (int i = 0; i < nhb; ++i)
            tma_load_3d(&jb.xmap8, base + (uint32_t)(UM_TM * 128 + i * 1024), col, row0 + UM_TM + 8 * i, q.b, bar);
          if (ident) mbar_arrive(&misc->raw_full[q.stage]);
        } else {
          // two 128-row boxes of residual / y_old rows in the SWIZZLE_128B pattern of the A tile, zero-filled past Lout
          const bool is_res = sl.idx < lay.n_res;
          const int hh = is_res ? sl.idx : sl.idx - lay.n_res;
          const int col = q.nt * lay.n_tile + hh * UM_CK;
          const bool two = q.t0 + 128 < p.Lout;
          uint64_t* bar = &misc->a_full[q.stage];
          mbar_expect_tx(bar, two ? 32768u : 16384u);
          const CUtensorMap* mp = is_res ? &jb.rmap : &jb.ymap;
          tma_load_3d(mp, base, col, q.t0, q.b, bar);
          if (two) tma_load_3d(mp, base + 16384u, col, q.t0 + 128, q.b, bar);
          mbar_arrive(&misc->raw_full[q.stage]);
        }
        um_advance(q, nct, lay.na, total_tiles, n_ttiles, n_ntiles);
      }
    }
  } else if (warp >= UM_EPI_WARP0) {
    // ======================================================================= epilogue
    if (p.y_dtype == ST2_DT_F32) um_epilogue<true, PAIR>(J, smem_raw, misc, tmem_base, warp, lane, total_tiles, n_ttiles, n_ntiles);
    else um_epilogue<false, PAIR>(J, smem_raw, misc, tmem_base, warp, lane, total_tiles, n_ttiles, n_ntiles);
  } else {
    // ============================================================ activation producers (warps 0-7)
    const int ptid = threadIdx.x - UM_PROD_WARP0 * 32;
    if (p.ss) {
      if (p.act == ST2_ACT_SNAKE) um_produce<ST2_ACT_SNAKE, true, PAIR>(J, smem_raw, misc, ptid, total_tiles, n_ttiles, n_ntiles, nchunks);
      else if (p.act == ST2_ACT_LRELU) um_produce<ST2_ACT_LRELU, true, PAIR>(J, smem_raw, misc, ptid, total_tiles, n_ttiles, n_ntiles, nchunks);
      else um_produce<ST2_ACT_NONE, true, PAIR>(J, smem_raw, misc, ptid, total_tiles, n_ttiles, n_ntiles, nchunks);
    } else {
      if (p.act == ST2_ACT_SNAKE) um_produce<ST2_ACT_SNAKE, false, PAIR>(J, smem_raw, misc, ptid, total_tiles, n_ttiles, n_ntiles, nchunks);
      else if (p.act == ST2_ACT_LRELU) um_produce<ST2_ACT_LRELU, false, PAIR>(J, smem_raw, misc, ptid, total_tiles, n_ttiles, n_ntiles, nchunks);
      else um_produce<ST2_ACT_NONE, false, PAIR>(J, smem_raw, misc, ptid, total_tiles, n_ttiles, n_ntiles, nchunks);
    }
  }

  tc_fence_before();
  if (PAIR) cluster_sync_all();  // the leader's MMAs read this CTA's shared memory until its last commit
  else __syncthreads();
  if (warp == UM_MMA_WARP) {
    tc_fence_after();
    if (PAIR) asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512) : "memory");
    else asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512) : "memory");
  }
}

// ------------------------------------------------------------------------ host side

// columns per tile for a GEMM N of n_cols (0 = not representable)
int umma_n_tile(int n_cols) {
  if (n_cols <= 0) return 0;
  if (n_cols % 128 == 0) return 128;
  if (n_cols % 64 == 0) return 64;
  if (n_cols % 32 == 0) return 32;
  return 0;
}

static inline uint16_t f32_to_bf16_rn(float f) {
  uint32_t u;
  memcpy(&u, &f, 4);
  if ((u & 0x7fffffffu) > 0x7f800000u) return (uint16_t)((u >> 16) | 0x40);
  u += 0x7fffu + ((u >> 16) & 1u);
  return (uint16_t)(u >> 16);
}
static inline float bf16_to_f32(uint16_t h) {
  uint32_t u = (uint32_t)h << 16;
  float f;
  memcpy(&f, &u, 4);
  return f;
}

// Identity weight blobs (n_tile/64 of them, one per 64-channel chunk of the residual row):
// blob h, element (v, n, e) = mult if n == h*64 + v*8 + e else 0.  Built once per (device, n_tile,
// mult) on first use (never inside a stream capture: the first forward of a handle is eager).
static std::mutex g_ident_mu;
static std::map<uint64_t, uint8_t*> g_ident;
static int um_ident(int dev, int n_tile, float mult, const uint8_t** out) {
  const uint16_t mb = f32_to_bf16_rn(mult);
  const uint64_t key = ((uint64_t)dev << 40) | ((uint64_t)n_tile << 16) | mb;
  std::lock_guard<std::mutex> lk(g_ident_mu);
  auto it = g_ident.find(key);
  if (it != g_ident.end()) {
    *out = it->second;
    return ST2_OK;
  }
  const int nb = n_tile / 64;
  const size_t blob_elems = (size_t)8 * n_tile * 8;
  std::vector<uint16_t> h(nb * blob_elems, 0);
  for (int hh = 0; hh < nb; ++hh)
    for (int v = 0; v < 8; ++v)
      for (int e = 0; e < 8; ++e) {
        const int n = hh * 64 + v * 8 + e;
        h[hh * blob_elems + ((size_t)v * n_tile + n) * 8 + e] = mb;
      }
  uint8_t* d = nullptr;
  ST2_CUDA_OK(cudaMalloc(&d, h.size() * 2));
  ST2_CUDA_OK(cudaMemcpy(d, h.data(), h.size() * 2, cudaMemcpyHostToDevice));
  g_ident[key] = d;
  *out = d;
  return ST2_OK;
}

// A/B and debug switches (ST2_UMMA_*), read from the environment ONCE (first launch) and again only when a tool asks
// for it with st2_umma_reload_env(): no getenv on the launch path.
struct UmEnv {
  int dbg = 0;
  uint32_t spin = 0;
  bool pair = false;
  int wst = 0;    // 0 = default depth
  int dist = -1;  // -1 = default
  bool res_tma = true, x_tma = true, loader = true;
  bool no_multi = false;  // ST2_UMMA_MULTI=0: launch the jobs of launch_conv_umma_multi one by one (A/B switch)
};
static std::mutex g_env_mu;
static UmEnv g_env;
static bool g_env_ready = false;
static void um_env_read_locked() {
  UmEnv e;
  if (const char* v = getenv("ST2_UMMA_DBG")) e.dbg = atoi(v);
  if (const char* v = getenv("ST2_UMMA_SPIN")) e.spin = (uint32_t)atoi(v);
  if (const char* v = getenv("ST2_UMMA_PAIR")) e.pair = v[0] == '1';
  if (const char* v = getenv("ST2_UMMA_WST")) e.wst = atoi(v);
  if (const char* v = getenv("ST2_UMMA_DIST")) e.dist = atoi(v);
  if (const char* v = getenv("ST2_UMMA_RES_TMA")) e.res_tma = v[0] != '0';
  if (const char* v = getenv("ST2_UMMA_X_TMA")) e.x_tma = v[0] != '0';
  if (const char* v = getenv("ST2_UMMA_LOADER")) e.loader = v[0] != '0';
  if (const char* v = getenv("ST2_UMMA_MULTI")) e.no_multi = v[0] == '0';
  g_env = e;
  g_env_ready = true;
}
static UmEnv um_env() {
  std::lock_guard<std::mutex> lk(g_env_mu);
  if (!g_env_ready) um_env_read_locked();
  return g_env;
}

static int um_layout(const St2ConvParams& p, int dev, const UmEnv& env, UmLayout& lay) {
  lay.n_tile = p.n_tile ? p.n_tile : umma_n_tile(p.n_cols);
  ST2_REQUIRE(lay.n_tile != 0, "conv_umma: n_cols must be a multiple of 32");
  ST2_REQUIRE((lay.n_tile == 32 || lay.n_tile == 64 || lay.n_tile == 128) && p.n_cols % lay.n_tile == 0,
              "conv_umma: n_tile must be 32, 64 or 128 and divide n_cols");
  const int halo = (p.k - 1) * p.dil;
  lay.a_bytes = (int)round_up((int64_t)(UM_TM + halo) * 128, 1024);
  lay.w_bytes = 8 * lay.n_tile * 16;
  const int esz = p.y_dtype == ST2_DT_F32 ? 4 : 2;
  lay.out_row = lay.n_tile * esz;
  ST2_REQUIRE(lay.out_row % 128 == 0, "conv_umma: an output row segment must be a multiple of 128 bytes");
  const int stg_bytes = (lay.out_row / 128) * 16384;
  const int misc_bytes = (int)round_up((int64_t)sizeof(UmMisc), 128);
  // CTA-pair mode (UmLayout::pair): 128-column bf16 tiles whose bias does not depend on the item.  Off unless
  // ST2_UMMA_PAIR=1 (A/B switch).  A stage then holds half a blob, and the ring is twice as deep in the same bytes:
  // a stage's round trip includes the relay through the peer CTA.
  lay.dbg = env.dbg;
  lay.spin = env.spin;
  lay.pair = 0;
  if (env.pair && lay.n_tile == 128 && p.y_dtype == ST2_DT_BF16 && (!p.bias || p.bias_bstride == 0) && !(lay.dbg & 1))
    lay.pair = 1;
  lay.w_slot = lay.pair ? lay.w_bytes / 2 : lay.w_bytes;
  lay.nw = lay.pair ? 2 * UM_W_STAGES : UM_W_STAGES;
  // With 5 or more taps per chunk the weight stream, not the activation ring, is what the MMAs wait for: up to 5
  // stages as long as 3 A stages still fit (measured after the issue-loop change: k = 11 460 -> 424 us, k = 7
  // 332 -> 310 us at 5 stages; k = 3 loses 3 % when its A ring shrinks from 4 to 3, so it keeps 3 weight stages).
  if (!lay.pair && p.k >= 5) {
    int nw = 5;
    while (nw > UM_W_STAGES && (UM_SMEM_LIMIT - (nw * lay.w_slot + stg_bytes + misc_bytes)) / lay.a_bytes < 3) --nw;
    lay.nw = nw;
  }
  if (env.wst > 0) {  // A/B switch
    lay.nw = env.wst;
    if (lay.nw < 2) lay.nw = 2;
    if (lay.nw > UM_MAX_W) lay.nw = UM_MAX_W;
  }
  const int fixed = lay.nw * lay.w_slot + stg_bytes + misc_bytes;
  int na = (UM_SMEM_LIMIT - fixed) / lay.a_bytes;
  if (na > UM_MAX_A) na = UM_MAX_A;
  ST2_REQUIRE(na >= 3, "conv_umma: tile does not fit shared memory");
  lay.na = na;
  // producers' prefetch policy: 0 = opportunistic (issue as soon as a slot is released), > 0 = fixed
  // distance in chunks.  ST2_UMMA_DIST overrides for A/B measurements.
  lay.dist = na - 2;  // measured: fixed distance beats opportunistic issue by 5-10 % on residual convs
  if (env.dist >= 0) {
    lay.dist = env.dist;
    if (lay.dist > na - 1) lay.dist = na - 1;
    if (lay.dist < 0) lay.dist = 0;
  }
  lay.off_a = (uint32_t)stg_bytes;
  lay.off_w = lay.off_a + (uint32_t)(na * lay.a_bytes);
  lay.off_misc = lay.off_w + (uint32_t)(lay.nw * lay.w_slot);
  lay.total = lay.off_misc + (uint32_t)misc_bytes;
  lay.idesc = um_idesc(lay.n_tile);
  lay.idesc64 = um_idesc(64);
  lay.n_res = lay.n_acc = 0;
  if (p.res) {
    ST2_REQUIRE(lay.n_tile >= 64, "conv_umma: residual needs a 64- or 128-column tile");
    lay.n_res = lay.n_tile / 64;
    int rc = um_ident(dev, lay.n_tile, 1.f, &lay.ident_res);
    if (rc) return rc;
  }
  if (p.accumulate) {
    ST2_REQUIRE(lay.n_tile >= 64, "conv_umma: accumulate needs a 64- or 128-column tile");
    const float mult = 1.f / p.out_scale;
    const float mr = bf16_to_f32(f32_to_bf16_rn(mult));
    ST2_REQUIRE(fabsf(mr - mult) <= 1e-6f * fabsf(mult),
                "conv_umma: accumulate needs 1/out_scale representable in bf16 (mean over 1, 2, 3, 4 ... resblocks)");
    lay.n_acc = lay.n_tile / 64;
    int rc = um_ident(dev, lay.n_tile, mr, &lay.ident_acc);
    if (rc) return rc;
  }
  return ST2_OK;
}

// cuTensorMapEncodeTiled through the runtime (no link-time dependency on libcuda)
typedef CUresult (*UmEncodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static UmEncodeTiled g_encode = nullptr;
static std::mutex g_init_mu;  // g_encode, g_num_sms
static int um_encode_fn() {
  std::lock_guard<std::mutex> lk(g_init_mu);
  if (g_encode) return ST2_OK;
  void* fn = nullptr;
  cudaDriverEntryPointQueryResult qres;
  ST2_CUDA_OK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres));
  ST2_REQUIRE(fn != nullptr && qres == cudaDriverEntryPointSuccess, "cuTensorMapEncodeTiled is not available");
  g_encode = reinterpret_cast<UmEncodeTiled>(fn);
  return ST2_OK;
}
// output view [B][Lout][n_cols] (row stride ldy, batch stride y_bstride); box = 128 bytes x 128 rows
static int um_ymap(const St2ConvParams& p, CUtensorMap* map) {
  int rc = um_encode_fn();
  if (rc) return rc;
  const bool f32 = p.y_dtype == ST2_DT_F32;
  const uint64_t esz = f32 ? 4 : 2;
  const cuuint64_t gdim[3] = {(cuuint64_t)p.n_cols, (cuuint64_t)p.Lout, (cuuint64_t)p.B};
  const cuuint64_t gstr[2] = {(cuuint64_t)p.ldy * esz, (cuuint64_t)(p.B > 1 ? p.y_bstride : (int64_t)p.ldy * p.Lout) * esz};
  const cuuint32_t box[3] = {(cuuint32_t)(128 / esz), 128u, 1u};
  const cuuint32_t estr[3] = {1u, 1u, 1u};
  const CUresult r = g_encode(map, f32 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, p.y,
                              gdim, gstr, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                              CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("conv_umma: cuTensorMapEncodeTiled failed with code " + std::to_string((int)r));
    return ST2_ECUDA;
  }
  return ST2_OK;
}

// residual view [B][Lout][n_cols] of p.res (row stride ldres), same box / swizzle as the output map
static int um_rmap(const St2ConvParams& p, CUtensorMap* map) {
  int rc = um_encode_fn();
  if (rc) return rc;
  const cuuint64_t gdim[3] = {(cuuint64_t)p.n_cols, (cuuint64_t)p.Lout, (cuuint64_t)p.B};
  const cuuint64_t gstr[2] = {(cuuint64_t)p.ldres * 2, (cuuint64_t)(p.B > 1 ? p.res_bstride : (int64_t)p.ldres * p.Lout) * 2};
  const cuuint32_t box[3] = {64u, 128u, 1u};
  const cuuint32_t estr[3] = {1u, 1u, 1u};
  const CUresult r = g_encode(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(p.res), gdim, gstr, box, estr,
                              CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                              CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("conv_umma: cuTensorMapEncodeTiled (residual) failed with code " + std::to_string((int)r));
    return ST2_ECUDA;
  }
  return ST2_OK;
}

// input view [B][Lin][cin] of p.x (row stride ldx); rows = 256 for the body of a tile, 8 for its halo
static int um_xmap(const St2ConvParams& p, int rows, CUtensorMap* map) {
  int rc = um_encode_fn();
  if (rc) return rc;
  const cuuint64_t gdim[3] = {(cuuint64_t)p.cin, (cuuint64_t)p.Lin, (cuuint64_t)p.B};
  const cuuint64_t gstr[2] = {(cuuint64_t)p.ldx * 2, (cuuint64_t)(p.B > 1 ? p.x_bstride : (int64_t)p.ldx * p.Lin) * 2};
  const cuuint32_t box[3] = {64u, (cuuint32_t)rows, 1u};
  const cuuint32_t estr[3] = {1u, 1u, 1u};
  const CUresult r = g_encode(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(p.x), gdim, gstr, box, estr,
                              CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                              CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("conv_umma: cuTensorMapEncodeTiled (input) failed with code " + std::to_string((int)r));
    return ST2_ECUDA;
  }
  return ST2_OK;
}

// weight blobs [blob][K group (8)][column (128)][8 channels] as a 3-D tensor: a box of 8 K groups x 64 columns is one
// CTA's half of a blob in pair mode
static int um_wmap(const void* ptr, int n_blobs, CUtensorMap* map) {
  int rc = um_encode_fn();
  if (rc) return rc;
  const cuuint64_t gdim[3] = {8u, 128u, (cuuint64_t)n_blobs * 8u};
  const cuuint64_t gstr[2] = {16u, 2048u};
  const cuuint32_t box[3] = {8u, 64u, 8u};
  const cuuint32_t estr[3] = {1u, 1u, 1u};
  const CUresult r = g_encode(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(ptr), gdim, gstr, box, estr,
                              CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                              CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("conv_umma: cuTensorMapEncodeTiled (weights) failed with code " + std::to_string((int)r));
    return ST2_ECUDA;
  }
  return ST2_OK;
}

static void um_fastdiv(uint32_t d, uint32_t& m, uint32_t& l) {
  l = 0;
  while ((1u << l) < d) ++l;
  m = (uint32_t)((((uint64_t)1 << 32) * (((uint64_t)1 << l) - d)) / d + 1);
}

static int g_num_sms[64] = {};

// SM count + the kernels' dynamic shared-memory limit, once per device
static int um_device_setup(int dev, int* sms) {
  std::lock_guard<std::mutex> lk(g_init_mu);
  if (dev >= 64 || g_num_sms[dev] == 0) {
    int n = 0;
    ST2_CUDA_OK(cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev));
    ST2_CUDA_OK(cudaFuncSetAttribute(conv_umma_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     UM_SMEM_LIMIT));
    ST2_CUDA_OK(cudaFuncSetAttribute(conv_umma_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     UM_SMEM_LIMIT));
    if (dev < 64) g_num_sms[dev] = n;
    else { *sms = n; return ST2_OK; }
  }
  *sms = g_num_sms[dev];
  return ST2_OK;
}

// optional per-launch event timing (st2_prof_enable / st2_prof_read)
// per-launch event timing: the state of conv_umma.cu (st2_prof_enable / st2_prof_read)
#define g_prof (::st2::umma_prof())

// argument checks + layout + tensor maps of one conv -> jb; n_ttiles / n_ntiles / total of its tile grid
static int um_build_job(const St2ConvParams& p, int dev, const UmEnv& env, UmJob& jb, int* n_ttiles_o, int* n_ntiles_o,
                        int64_t* total_o) {
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
    ST2_REQUIRE((p.n_tile ? p.n_tile : umma_n_tile(p.n_cols)) != 128, "conv_umma: fp32 output needs a 32- or 64-column tile");
  } else {
    ST2_REQUIRE(p.y_dtype == ST2_DT_BF16, "conv_umma: output dtype");
    ST2_REQUIRE(p.ldy % 8 == 0 && p.y_bstride % 8 == 0, "conv_umma: 16-byte alignment of output rows");
  }
  ST2_REQUIRE(!p.res || (p.ldres % 8 == 0 && p.res_bstride % 8 == 0 && ((uintptr_t)p.res & 15) == 0),
              "conv_umma: residual alignment");
  ST2_REQUIRE(!(p.stats && p.accumulate), "conv_umma: stats with accumulate");
  ST2_REQUIRE(p.act != ST2_ACT_SNAKE || p.alpha, "conv_umma: snake needs alpha");
  ST2_REQUIRE(p.act != ST2_ACT_LRELU || (p.slope >= 0.f && p.slope <= 1.f), "conv_umma: LeakyReLU slope in [0, 1]");
  memset(&jb, 0, sizeof(jb));
  jb.p = p;
  UmLayout& lay = jb.lay;
  int rc = um_layout(p, dev, env, lay);
  if (rc) return rc;
  rc = um_ymap(p, &jb.ymap);
  if (rc) return rc;
  // residual / y_old rows by TMA when the source is a plain [B][Lout][>= n_cols] bf16 view (not the x2-repeated
  // shortcut of the upsampling block); ST2_UMMA_RES_TMA=0 keeps the cp.async path (A/B switch)
  jb.rmap = jb.ymap;
  lay.res_tma = lay.acc_tma = 0;
  if (env.res_tma && p.y_dtype == ST2_DT_BF16) {
    if (lay.n_acc) lay.acc_tma = 1;
    if (lay.n_res && p.res_shift == 0 && (((uintptr_t)p.res) & 15) == 0 && (p.ldres % 8) == 0 &&
        (p.res_bstride % 8) == 0) {
      rc = um_rmap(p, &jb.rmap);
      if (rc) return rc;
      lay.res_tma = 1;
    }
  }
  jb.xmap = jb.ymap;
  jb.xmap8 = jb.ymap;
  lay.x_tma = 0;
  if (env.x_tma && (((uintptr_t)p.x) & 15) == 0 && (p.ldx % 8) == 0 && (p.x_bstride % 8) == 0) {
    rc = um_xmap(p, UM_TM, &jb.xmap);
    if (rc) return rc;
    rc = um_xmap(p, 8, &jb.xmap8);
    if (rc) return rc;
    lay.x_tma = 1;
  }
  lay.loader = env.loader && lay.x_tma && (!lay.n_res || lay.res_tma) && (!lay.n_acc || lay.acc_tma);
  const int n_ttiles = ceil_div(p.Lout, UM_TM), n_ntiles = p.n_cols / lay.n_tile;
  ST2_REQUIRE(!p.stats || p.n_tiles >= n_ttiles, "conv_umma: n_tiles");
  um_fastdiv((uint32_t)n_ntiles, lay.nt_m, lay.nt_l);
  um_fastdiv((uint32_t)n_ttiles, lay.tt_m, lay.tt_l);
  const int64_t total = (int64_t)p.B * n_ttiles * n_ntiles;
  ST2_REQUIRE(total < (1ll << 31), "conv_umma: too many tiles");
  *n_ttiles_o = n_ttiles; *n_ntiles_o = n_ntiles; *total_o = total;
  return ST2_OK;
}

// can jobs a and b alternate in one kernel?  Same tile grid and channel counts, same prologue kind (the producers are
// compiled per activation), every chunk by TMA through the loader lane, no CTA-pair mode.
static bool um_jobs_compatible(const UmJob& a, const UmJob& b) {
  const St2ConvParams &p = a.p, &q = b.p;
  return p.B == q.B && p.Lin == q.Lin && p.Lout == q.Lout && p.cin == q.cin && p.n_cols == q.n_cols &&
         p.y_dtype == q.y_dtype && p.act == q.act && (p.ss != nullptr) == (q.ss != nullptr) && p.slope == q.slope &&
         a.lay.n_tile == b.lay.n_tile && a.lay.out_row == b.lay.out_row && a.lay.loader && b.lay.loader &&
         !a.lay.pair && !b.lay.pair && a.lay.dbg == b.lay.dbg;
}

static int um_launch_jobs(UmJobs& J, int n_ttiles, int n_ntiles, int64_t total, int dev, cudaStream_t st) {
  UmLayout& lay = J.j[0].lay;
  int sms = 148;
  int rc = um_device_setup(dev, &sms);
  if (rc) return rc;
  const int64_t items = total * J.n;  // (conv, tile) pairs
  ST2_REQUIRE(items < (1ll << 31), "conv_umma_multi: too many tiles");
  int grid = (int)(items < sms ? items : sms);
  J.wmap = J.j[0].ymap; J.wmap_res = J.j[0].ymap; J.wmap_acc = J.j[0].ymap;
  if (lay.pair) {  // decided in um_layout; single job only
    const St2ConvParams& p = J.j[0].p;
    rc = um_wmap(p.w, n_ntiles * (p.cin / UM_CK) * p.k, &J.wmap);
    if (rc) return rc;
    if (lay.n_res) rc = um_wmap(lay.ident_res, lay.n_res, &J.wmap_res);
    if (rc) return rc;
    if (lay.n_acc) rc = um_wmap(lay.ident_acc, lay.n_acc, &J.wmap_acc);
    if (rc) return rc;
    lay.n_bt = p.B * n_ttiles;
    lay.total_ptiles = n_ntiles * ceil_div(lay.n_bt, 2);
    lay.idesc_pair = um_idesc(128, 256);
    const int64_t want_grid = 2 * (int64_t)lay.total_ptiles;
    grid = (int)(want_grid < (sms & ~1) ? want_grid : (sms & ~1));
  }
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
    for (int i = 0; i < J.n; ++i) g_prof.flops += J.j[i].p.algo_flops;
    ST2_CUDA_OK(cudaEventRecord(e0, st));
  }
  if (lay.pair) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid); cfg.blockDim = dim3(UM_THREADS); cfg.dynamicSmemBytes = (size_t)lay.total; cfg.stream = st;
    cudaLaunchAttribute attr[2];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = pdl_enabled() ? 1 : 0;
    attr[1].id = cudaLaunchAttributeClusterDimension;
    attr[1].val.clusterDim.x = 2; attr[1].val.clusterDim.y = 1; attr[1].val.clusterDim.z = 1;
    cfg.attrs = attr; cfg.numAttrs = 2;
    ST2_CUDA_OK(cudaLaunchKernelEx(&cfg, conv_umma_kernel<true>, J, (int)total, n_ttiles, n_ntiles));
  } else {
    ST2_CUDA_OK(launch_pdl(conv_umma_kernel<false>, dim3(grid), dim3(UM_THREADS), (size_t)lay.total, st, J, (int)total,
                           n_ttiles, n_ntiles));
  }
  ST2_LAUNCH_OK("conv_umma");
  if (e1) ST2_CUDA_OK(cudaEventRecord(e1, st));
  return ST2_OK;
}

// n (<= UM_MAX_JOBS) convs of the same geometry as ONE kernel whose CTAs alternate between them tile by tile (UmJobs).
// Jobs that cannot share a kernel (different shapes, a fallback load path, pair mode) are launched one by one: the
// results are the same either way, bit for bit (every tile is computed by the same code on the same operands).
int launch_conv_umma_multi(const St2ConvParams* ps, int n, cudaStream_t st) {
  ST2_REQUIRE(ps && n >= 1 && n <= UM_MAX_JOBS, "conv_umma_multi: 1..3 jobs");
  if (n == 1) return ::st2::launch_conv_umma(ps[0], st);
  int dev = 0;
  ST2_CUDA_OK(cudaGetDevice(&dev));
  const UmEnv env = um_env();
  static thread_local UmJobs J;
  int n_ttiles = 0, n_ntiles = 0;
  int64_t total = 0;
  bool ok = !env.no_multi;
  for (int i = 0; i < n; ++i) {
    int tt = 0, nt = 0;
    int64_t tot = 0;
    int rc = um_build_job(ps[i], dev, env, J.j[i], &tt, &nt, &tot);
    if (rc) return rc;
    if (i == 0) { n_ttiles = tt; n_ntiles = nt; total = tot; }
    else ok = ok && um_jobs_compatible(J.j[0], J.j[i]) && tt == n_ttiles && nt == n_ntiles;
  }
  if (!ok) {
    for (int i = 0; i < n; ++i) {
      int rc = ::st2::launch_conv_umma(ps[i], st);
      if (rc) return rc;
    }
    return ST2_OK;
  }
  // common shared-memory geometry: the job with the widest halo decides stage size and ring depths
  int ref = 0;
  for (int i = 1; i < n; ++i)
    if (J.j[i].lay.a_bytes > J.j[ref].lay.a_bytes ||
        (J.j[i].lay.a_bytes == J.j[ref].lay.a_bytes && J.j[i].p.k > J.j[ref].p.k)) ref = i;
  const UmLayout g = J.j[ref].lay;
  for (int i = 0; i < n; ++i) {
    UmLayout& l = J.j[i].lay;
    l.a_bytes = g.a_bytes; l.na = g.na; l.nw = g.nw; l.w_slot = g.w_slot; l.dist = g.dist;
    l.off_a = g.off_a; l.off_w = g.off_w; l.off_misc = g.off_misc; l.total = g.total;
  }
  J.n = n;
  return um_launch_jobs(J, n_ttiles, n_ntiles, total, dev, st);
}

}  // namespace mj

int launch_conv_umma_multi(const St2ConvParams* ps, int n, cudaStream_t st) { return mj::launch_conv_umma_multi(ps, n, st); }
int umma_multi_wait_read(unsigned long long* h16, bool reset) {  // + the counters of conv_umma.cu's kernel
  unsigned long long h[16];
  ST2_CUDA_OK(cudaMemcpyFromSymbol(h, mj::g_um_wait, sizeof(h)));
  for (int i = 0; i < 16; ++i) h16[i] += h[i];
  if (reset) {
    memset(h, 0, sizeof(h));
    ST2_CUDA_OK(cudaMemcpyToSymbol(mj::g_um_wait, h, sizeof(h)));
  }
  return ST2_OK;
}
void umma_multi_reload_env() {
  std::lock_guard<std::mutex> lk(mj::g_env_mu);
  mj::um_env_read_locked();
}

}  // namespace st2
