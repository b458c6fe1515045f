// swarm_common.cuh — device-side parameter block and small helpers shared by every kernel.
//
// Resident layout in HBM (struct-of-arrays, one handle per GPU):
//   cov   uint32 [N][env_words]   COVERAGE MAP, 1 bit per cell, 32 cells per word, flat cell index c = y*R + x (the
//                                 reference's obs order, A2C/noVis_sim.py:168-170): bit = 1 <=> the cell is still unmapped
//                                 (reference grid value > 0).  Bits past R*C are 0.  env_words is a multiple of 4 (16 B).
//   alist uint32 [N][8]           LIST mode (n_anamolous <= 8 and R, C <= 256 — every reference call site): the env's
//                                 anomalous cells as x | y << 16, entries >= n_anamolous unused.  Static within an episode:
//                                 a cell's value is  bit ? (listed ? U : 1) : 0, so "anomaly found" is the cleared bit.
//   anomp uint32 [N][env_words]   PLANE mode (anything else): the anomalous cells as a second bit plane (static per episode).
//   pos   float2 [N][D]           (x, y) = reference n_drones_pos[i]
//   meta  int4   [N]              {step_func_count, remaining (#cells > 0), episode_return, episode}
//   wait  int16  [N][R*C]         dwell counters, only allocated when dwell_normal|dwell_anom != 0
// Round 1 kept 2 bits per cell; the 1-bit map halves the largest input stream of the step (cfg3: 64 -> 32 MiB per launch)
// and the shared-memory footprint of a staged env, which is what buys the deeper TMA ring of swarm_warp.cuh.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace swarm {

constexpr unsigned FULL = 0xffffffffu;

enum : int { F_MOVE = 1, F_SCATTER = 2, F_REWARD = 4, F_OBS = 8, F_ALL = 15 };

// Peer-memory statistics exchange (SURVEY §8e "fused one-shot variant"): every rank owns a mailbox of SWARM_MAX_RANKS slots,
// slot r = eight {value, version} pairs written by rank r with 16-byte stores over NVLink (peer-mapped through CUDA IPC).
// The push rides at the START of a step launch (CTA 0, thread 0): it publishes the running totals as they stand after the
// PREVIOUS launch — final by stream order — so it overlaps the whole kernel and costs the step nothing; the stores are
// posted (no fence, no read over NVLink).  swarm_stats_push publishes the current totals explicitly when exactness matters.
constexpr int SWARM_MAX_RANKS = 8;
constexpr int SWARM_MBOX_SLOT = 16;   // int64 words per slot: 8 x {value, version}
struct PeerStats {
  unsigned long long* mbox[SWARM_MAX_RANKS];   // mailbox base of every rank (own one included), null beyond n_ranks
  unsigned long long* pushes;                  // number of pushes this rank has made (device counter, = the version)
  int n_ranks, rank;                           // n_ranks == 0: exchange disabled
};

struct DevParams {
  // configuration (mirrors DroneEnv.__init__, A2C/noVis_sim.py:8-17)
  int N, D, R, C, RC, U, n_anom;
  int reward_mode, obs_mode, W, obs_dtype, auto_reset;
  int dwell_normal, dwell_anom, has_wait;
  int punish_thr;     // floor(R*C / step_size): punish <=> step_func_count > punish_thr   (noVis_sim.py:141-143)
  int cd_fast;        // collision_dist <= 0.5: at most one lattice cell can be hit
  int scan_m;         // ceil(cd) + 1: half-width of the bounded scan window for cd > 0.5
  int env_words;      // uint32 words per env of ONE bit plane (multiple of 4)
  int list_mode;      // 1: anomalies in `alist` (x | y << 16, n_anom <= 8, R, C <= 256); 0: second bit plane `anomp`
  int f32_exact;      // positions / squared distances are exact in float32 (step_size = m 2^-k, k <= 12, coordinates small):
                      // the FAST kernels compare in float; otherwise the generic double path runs
  float cd2_thr_f;    // smallest float >= cd2_thr:  d2 < cd2_thr <=> d2 < cd2_thr_f   for every d2 that is exactly a float
  float spread2_thr_f;// largest float <= spread2_thr: d2 > spread2_thr <=> d2 > spread2_thr_f   likewise
  int n_stages;       // depth of the CTA input ring of the staged kernel (TMA stages in flight)
  int obs_per_env;    // R*C or D*W*W
  float ss, xmax, ymax;
  double cd;          // collision_dist
  double spread_thr;  // row_count // n_drones                                            (noVis_sim.py:100)
  double cd2_thr;     // smallest double z with sqrt(z) >= cd:   sqrt(d2) < cd  <=>  d2 < cd2_thr   (exact, host-computed)
  double spread2_thr; // largest double z with sqrt(z) <= spread_thr: sqrt(d2) > thr <=> d2 > spread2_thr
  // warp-autonomous staged kernel (swarm_warp.cuh): per-warp shared-memory carve-up, bytes
  int stage_bytes;    // obs staging (WINDOW mode), multiple of 16, >= 16
  int in_bytes;       // one CTA input stage for EPC = wpb*EPW envs: [cov EPC*env_words*4][alist EPC*32][pos EPC*D*8][meta EPC*16]
                      // [actions EPC*D], each part 16 B aligned
  int off_alist, off_pos, off_meta, off_act;   // byte offsets inside an input stage
  int warp_smem;      // per-warp footprint used to size the CTA (stage_bytes + 2 * per-warp inputs + slack)
  int cta_smem;       // dynamic shared memory of a CTA: wpb * stage_bytes + n_stages * in_bytes + slack
  uint32_t nq_magic;  // ceil(2^32 / (RC/4)) for the FULL-mode vector -> (env, quad) split
  uint32_t seed_lo, seed_hi;
  long long env_id_base;
  // resident state
  PeerStats peers;
  uint32_t* cov;      // coverage bit plane
  uint32_t* anomp;    // anomaly bit plane (PLANE mode) or null
  uint32_t* alist;    // anomaly list (LIST mode) or null
  int* sched;         // int[4] dynamic CTA-group scheduler of the staged kernel: {next group, CTAs done, -, -}
  const uint32_t* tables;   // W == 5 window tables built once per handle (build_tables_kernel): [lut 1024 words][padx 256][pady 256]
  float2* pos;
  int4* meta;
  int16_t* wait;
  int* err;
};

__device__ __forceinline__ uint32_t ld_grid(const uint32_t* p) { return __ldcg(p); }
__device__ __forceinline__ void st_grid(uint32_t* p, uint32_t v) { __stcg(p, v); }
// one thread's ATOMS.ADD (the C++ atomicAdd on shared memory comes with warp-aggregation code this use does not need)
__device__ __forceinline__ uint32_t smem_atomic_inc(uint32_t addr_s) {
  uint32_t old;
  asm volatile("atom.shared.add.u32 %0, [%1], 1;" : "=r"(old) : "r"(addr_s) : "memory");
  return old;
}

// ---- sm_90+/sm_100a async-copy (TMA bulk) and mbarrier wrappers -----------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* ptr) { return (uint32_t)__cvta_generic_to_shared(ptr); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void fence_mbar_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
// Programmatic dependent launch (sm_90+): a step launch carries cudaLaunchAttributeProgrammaticStreamSerialization.  Once every
// CTA of the launch before it has called `pdl_launch_dependents` (after its main loop), its own CTAs may become resident and run
// their prologue (barrier init, table build); `pdl_wait` then returns once every prerequisite grid has completed and its memory
// is visible: it must precede the first access to anything an earlier launch wrote or read.  Both are no-ops in a launch
// without the attribute.
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
// orders this thread's earlier generic-proxy shared-memory accesses before later async-proxy (TMA) ones
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "SWARM_WAIT_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra SWARM_DONE_%=;\n"
      "bra SWARM_WAIT_%=;\n"
      "SWARM_DONE_%=:\n"
      "}\n" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
// global -> shared bulk copy (bytes % 16 == 0, both addresses 16 B aligned), completion on `bar`   [SASS: UBLKCP]
__device__ __forceinline__ void bulk_g2s(void* dst_smem, const void* src_gmem, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst_smem)),
               "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
// shared -> global bulk store (same alignment rules), tracked by the issuing thread's bulk async-group
__device__ __forceinline__ void bulk_s2g(void* dst_gmem, const void* src_smem, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst_gmem), "r"(smem_u32(src_smem)), "r"(bytes)
               : "memory");
}
// variants taking 32-bit shared-memory addresses computed once (saves the generic->shared conversion per use)
__device__ __forceinline__ void mbar_expect_tx_s(uint32_t bar_s, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar_s), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s_s(uint32_t dst_s, const void* src_gmem, uint32_t bytes, uint32_t bar_s) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst_s), "l"(src_gmem),
               "r"(bytes), "r"(bar_s)
               : "memory");
}
__device__ __forceinline__ void mbar_wait_s(uint32_t bar_s, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "SWARM_WAITS_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra SWARM_DONES_%=;\n"
      "bra SWARM_WAITS_%=;\n"
      "SWARM_DONES_%=:\n"
      "}\n" ::"r"(bar_s), "r"(parity) : "memory");
}
__device__ __forceinline__ void mbar_arrive_s(uint32_t bar_s) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar_s) : "memory");
}
__device__ __forceinline__ void bulk_s2g_s(void* dst_gmem, uint32_t src_s, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst_gmem), "r"(src_s), "r"(bytes) : "memory");
}
// the same copies with an L2 eviction-priority policy (createpolicy)
__device__ __forceinline__ uint64_t l2_policy_evict_first() {
  uint64_t pol;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
  return pol;
}
__device__ __forceinline__ void bulk_s2g_s_hint(void* dst_gmem, uint32_t src_s, uint32_t bytes, uint64_t pol) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group.L2::cache_hint [%0], [%1], %2, %3;" ::"l"(dst_gmem), "r"(src_s), "r"(bytes), "l"(pol) : "memory");
}
__device__ __forceinline__ void bulk_g2s_s_hint(uint32_t dst_s, const void* src_gmem, uint32_t bytes, uint32_t bar_s, uint64_t pol) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;" ::"r"(dst_s),
               "l"(src_gmem), "r"(bytes), "r"(bar_s), "l"(pol)
               : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read0() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }

// ---- cell codes ---------------------------------------------------------------------------------------------------
// Everything that turns state into values goes through a 2-bit CODE: 0 mapped (value 0), 1 normal (value 1),
// 2 pad (window cell outside the grid: -1 / 255), 3 anomalous (value U).  code = unmapped | (anomalous_or_pad << 1).
constexpr uint32_t CODE_MAPPED = 0u, CODE_NORMAL = 1u, CODE_PAD = 2u, CODE_ANOM = 3u;
// an anomaly-list entry that names no cell: (0x7000, 0x7000) is outside every LIST-mode grid (sides <= 256) and far enough
// from the 16-bit limits for the packed window test of swarm_warp.cuh
constexpr uint32_t LIST_EMPTY = 0x70007000u;
// obs element type of the packed 2-bit window format (SWARM_DTYPE_P2): one word per drone, bit k = LO plane, bit 32 + k = HI plane
using Packed2 = unsigned long long;

// Is `cell` (flat index) / (x, y) one of the env's anomalous cells?
__device__ __forceinline__ bool is_anomalous(const DevParams& p, int env, int cell, int x, int y) {
  if (p.list_mode) {
    const uint32_t* l = p.alist + (size_t)env * 8;
    const uint32_t key = (uint32_t)x | ((uint32_t)y << 16);
    bool hit = false;
    for (int k = 0; k < p.n_anom; ++k) hit = hit || (l[k] == key);
    return hit;
  }
  return (__ldcg(p.anomp + (size_t)env * p.env_words + (cell >> 5)) >> (cell & 31)) & 1u;
}
// Code of a cell of an env, read from HBM / L2 (state kernels, CTA kernel fallbacks; not the hot path).
__device__ __forceinline__ uint32_t cell_code_global(const DevParams& p, int env, int cell) {
  const uint32_t bit = (__ldcg(p.cov + (size_t)env * p.env_words + (cell >> 5)) >> (cell & 31)) & 1u;
  if (!bit) return CODE_MAPPED;
  const int y = cell / p.R, x = cell - y * p.R;
  return is_anomalous(p, env, cell, x, y) ? CODE_ANOM : CODE_NORMAL;
}

// ---- Philox4x32-10 (Salmon et al., SC'11); same constants/round structure as oracle_philox4x32_10 ----
__device__ __forceinline__ uint4 philox4x32_10(uint4 c, uint32_t k0, uint32_t k1) {
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    uint32_t hi0 = __umulhi(0xD2511F53u, c.x), lo0 = 0xD2511F53u * c.x;
    uint32_t hi1 = __umulhi(0xCD9E8D57u, c.z), lo1 = 0xCD9E8D57u * c.z;
    c = make_uint4(hi1 ^ c.y ^ k0, lo1, hi0 ^ c.w ^ k1, lo0);
    k0 += 0x9E3779B9u;
    k1 += 0xBB67AE85u;
  }
  return c;
}
__device__ __forceinline__ uint4 draw_block(const DevParams& p, long long env_id, uint32_t episode, uint32_t block) {
  return philox4x32_10(make_uint4((uint32_t)(unsigned long long)env_id, (uint32_t)((unsigned long long)env_id >> 32), episode, block),
                       p.seed_lo, p.seed_hi);
}

// One thread: publish stats[0..8) (read from L2, where the previous launch's atomics left them) to every rank's mailbox.
__device__ __forceinline__ void peer_stats_push(const PeerStats& ps, const unsigned long long* stats) {
  const unsigned long long ver = atomicAdd(ps.pushes, 1ull) + 1ull;   // atomic: two launches of one handle may overlap on different streams
  unsigned long long v[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) v[k] = __ldcg(stats + k);
  for (int r = 0; r < ps.n_ranks; ++r) {
    ulonglong2* slot = reinterpret_cast<ulonglong2*>(ps.mbox[r] + (size_t)ps.rank * SWARM_MBOX_SLOT);
#pragma unroll
    for (int k = 0; k < 8; ++k) __stcg(slot + k, make_ulonglong2(v[k], ver));   // 16-byte posted stores
  }
}
__global__ void peer_stats_push_kernel(PeerStats ps, const unsigned long long* stats) { peer_stats_push(ps, stats); }

// Sums the latest complete totals of all ranks' slots of THIS rank's mailbox into out[8]; waits (bounded) until every slot has
// been pushed at least min_push times.  One thread per rank slot.  A slot is read as eight 16-byte loads (the writer's unit:
// one {value, version} pair per posted 16-byte store), twice: the pass is accepted only when all sixteen versions agree, i.e.
// no push landed between or during the two passes.  A slot that never becomes consistent within the bound (~3 s) is a
// failure: bit 2 of the device error word is set and out[] is filled with -1 instead of a torn sum.
__device__ __forceinline__ ulonglong2 ld_pair_sys(const unsigned long long* p) {
  ulonglong2 v;
  asm volatile("ld.relaxed.sys.global.v2.u64 {%0, %1}, [%2];" : "=l"(v.x), "=l"(v.y) : "l"(p) : "memory");
  return v;
}
__global__ void peer_stats_reduce_kernel(const unsigned long long* mbox, int n_ranks, unsigned long long min_push, unsigned long long* out,
                                         int* err) {
  __shared__ unsigned long long acc[8];
  __shared__ int failed;
  if (threadIdx.x < 8) acc[threadIdx.x] = 0ull;
  if (threadIdx.x == 0) failed = 0;
  __syncthreads();
  if ((int)threadIdx.x < n_ranks) {
    const unsigned long long* slot = mbox + (size_t)threadIdx.x * SWARM_MBOX_SLOT;
    unsigned long long v[8];
    bool ok = false;
    for (long long spin = 0; spin < (1ll << 22) && !ok; ++spin) {
      const unsigned long long ver = ld_pair_sys(slot).y;
      ok = ver >= min_push;
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const ulonglong2 pr = ld_pair_sys(slot + 2 * k);
        v[k] = pr.x;
        ok = ok && pr.y == ver;
      }
#pragma unroll
      for (int k = 0; k < 8; ++k) {   // second pass: nothing changed while the first one ran
        const ulonglong2 pr = ld_pair_sys(slot + 2 * k);
        ok = ok && pr.y == ver && pr.x == v[k];
      }
      if (!ok) __nanosleep(500);
    }
    if (ok) {
#pragma unroll
      for (int k = 0; k < 8; ++k) atomicAdd(&acc[k], v[k]);
    } else {
      atomicOr(&failed, 1);
    }
  }
  __syncthreads();
  if (threadIdx.x < 8) out[threadIdx.x] = failed ? ~0ull : acc[threadIdx.x];
  if (threadIdx.x == 0 && failed && err) atomicOr(err, 4);
}

// per-thread statistics accumulator; reduced once per CTA, 6 atomics per CTA
struct StatAcc {
  long long rew = 0, ret = 0, len = 0;
  int done = 0, cells = 0, anom = 0;
};

}  // namespace swarm
