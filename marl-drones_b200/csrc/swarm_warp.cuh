// swarm_warp.cuh — the fused env step as a warp-autonomous, shared-memory-staged kernel (the hot path for D <= 32 and
// coverage maps of a few KiB: every BASELINE config except the 256x256 stress case).  LIST-mode anomalies only.
//
// One warp owns EPW = 32/TS consecutive envs (a "group"); a tile of TS lanes serves one env, one lane per drone.
// A CTA of wpb warps works on CTA-groups of EPC = wpb * EPW consecutive envs, whose inputs are contiguous in every array:
//   (a) INPUT RING.  The CTA-group's coverage planes, anomaly lists, positions, meta and actions arrive in one of
//       n_stages shared-memory stages through five TMA bulk copies (cp.async.bulk, SASS UBLKCP.S.G) on the stage's "full"
//       mbarrier.  There is no producer thread and no "empty" barrier: the LAST WARP TO LEAVE a stage (a shared-memory
//       counter) immediately issues the copies of the CTA-group n_stages iterations ahead into it — the refill goes out at the
//       earliest moment the stage is free, whichever warp that is, and up to n_stages - 1 CTA-groups (cfg3: 11 KB each)
//       are in flight per CTA while one is being computed.  (Round 1: a 2-deep ring refilled by thread 0, which had to wait
//       for the slowest warp first; 22 % of the kernel's stall samples sat on the full barrier.)
//       CTA-groups are handed out by an atomic counter (io.dyn) or statically (blockIdx.x + k * gridDim.x).
//   (b) stages (1)-(3) run on the shared-memory copy (cleared bits are written through to HBM with RED.AND);
//   (c) WINDOW obs: every lane builds its drone's W x W window as two 25-bit masks (unmapped, anomalous-or-pad), expands
//       them through a 256-entry float4 LUT (4 cells per LDS.128) into the warp's staging slice (lane stride W*W words:
//       bank-conflict free), and ONE TMA bulk store (UBLKCP.G.S) streams the group's [EPW][D][W*W] block to the obs tensor;
//       FULL obs: the warp expands the bit plane straight from shared memory with 128-bit streaming stores.
// There is no __syncthreads in the group loop: warps run wait / compute / store independently.  T > 1 (swarm_rollout)
// keeps coverage, positions and meta on chip across steps.
//
// Reference lines per stage: see swarm_step.cuh.
#pragma once
#include "swarm_common.cuh"
#include "swarm_emit.cuh"
#include "swarm_step.cuh"
#include "swarm_window.cuh"

// Launch bounds of the staged kernel: CTAs of up to SWARM_MAXT threads, SWARM_MINB of them per SM — 2 x 448 threads = 28
// resident warps at <= 72 registers.  The host picks the CTA shape (warps per CTA, ring depth) per env shape so that the
// CTA-groups divide into whole rounds over the resident CTAs (swarm_api.cu, swarm_create): cfg3 = 14 warps x 2 CTAs per SM,
// 2341 CTA-groups over 296 CTAs = 7.91 -> 8 rounds (round 1: 8 warps x 3 CTAs, 9.22 -> 10 rounds).
#ifndef SWARM_MAXT
#define SWARM_MAXT 448
#endif
#ifndef SWARM_MINB
#define SWARM_MINB 2
#endif

namespace swarm {

constexpr int MAX_STAGES = 6;

// FULL obs of the group's `nenv` envs from their shared-memory planes through the 16-entry nibble LUT (`lutq`: 4 cells ->
// one float4 / four bytes), then — after a __syncwarp — the listed anomalous cells that are still unmapped are patched to U.
template <typename OT>
__device__ __forceinline__ void emit_full_lut(const DevParams& p, const ObsCvt<OT>& cvt, const uint32_t* lutq, const uint32_t* sgrid, const uint32_t* slist,
                                              int nenv, int lane, OT* __restrict__ o) {
  const int RC = p.RC;
  if ((RC & 3) == 0 && ((uintptr_t)o & 15) == 0) {
    const int nq = RC >> 2, total = nenv * nq;
#pragma unroll 2
    for (int v = lane; v < total; v += 32) {
      const int e = p.nq_magic ? (int)__umulhi((uint32_t)v, p.nq_magic) : v;   // nq == 1: one quad per env
      const int qd = v - e * nq;
      const uint32_t lo = (sgrid[e * p.env_words + (qd >> 3)] >> ((qd & 7) * 4)) & 15u;
      if constexpr (sizeof(OT) == 4) __stcs(reinterpret_cast<float4*>(o + (size_t)e * RC + 4 * qd), reinterpret_cast<const float4*>(lutq)[lo]);
      else __stcs(reinterpret_cast<uint32_t*>(o + (size_t)e * RC + 4 * qd), lutq[lo]);
    }
  } else {
    for (int e = 0; e < nenv; ++e)
      for (int c = lane; c < RC; c += 32) o[(size_t)e * RC + c] = cvt.one((sgrid[e * p.env_words + (c >> 5)] >> (c & 31)) & 1u);
  }
  __syncwarp();   // the patch below overwrites cells another lane has just written
  emit_full_patch<OT>(p, cvt, sgrid, slist, nenv, lane, 32, o);
}

// (the packed 2-bit format exists for WINDOW obs only; the FULL-mode branch of the kernel is dead code in its instantiations)
template <>
__device__ __forceinline__ void emit_full_lut<Packed2>(const DevParams&, const ObsCvt<Packed2>&, const uint32_t*, const uint32_t*, const uint32_t*, int, int,
                                                       Packed2* __restrict__) {}

// init_env (A2C/noVis_sim.py:36-63) for the env of this tile, built in its shared-memory stage and copied out to HBM.
// Same RNG spec as reset_tile / oracle_reset_env.  Called by all 32 lanes; `doit` is tile-uniform; q = the lane's new position.
template <int TS>
__device__ __forceinline__ void reset_tile_smem(const DevParams& p, int env, int episode, int tl, bool doit, uint32_t* sc, uint32_t* sl, float2& q) {
  reset_planes<TS>(p, env, episode, tl, doit, sc, nullptr, sl);
  if (doit) {
    const long long gid = p.env_id_base + env;
    uint4* g4 = reinterpret_cast<uint4*>(p.cov + (size_t)env * p.env_words);
    const uint4* s4 = reinterpret_cast<const uint4*>(sc);
    for (int w = tl; w < (p.env_words >> 2); w += TS) __stcg(g4 + w, s4[w]);
    if (tl == 0) {
      __stcg(reinterpret_cast<uint4*>(p.alist + (size_t)env * 8), *reinterpret_cast<const uint4*>(sl));
      __stcg(reinterpret_cast<uint4*>(p.alist + (size_t)env * 8) + 1, *(reinterpret_cast<const uint4*>(sl) + 1));
      p.meta[env] = make_int4(0, p.RC, 0, episode);
    }
    if (tl < p.D) {
      const uint4 r = draw_block(p, gid, (uint32_t)episode, 0x80000000u | ((uint32_t)tl >> 1));
      q.x = (float)__umulhi((tl & 1) ? r.z : r.x, (uint32_t)p.R);
      q.y = (float)__umulhi((tl & 1) ? r.w : r.y, (uint32_t)p.C);
      __stcg(p.pos + (size_t)env * p.D + tl, q);
    }
  }
  __syncwarp();
}

// ------------------------------------------------------------------------------------------------ the kernel
// MODE: 0 FULL obs, 1 WINDOW obs.  WC: compile-time window size (5) or 0 = runtime p.W.
// FAST: the production configuration folded at compile time — all four stages in one pass (swarm_step / swarm_rollout),
//       collision_dist <= 0.5, no dwell counters, float32-exact coordinates; actions of any dtype (uint8 ones ride the stage's
//       bulk copy, int32 / int64 — what torch's Categorical.sample() hands over — are read directly).  FAST = false keeps every
//       switch at run time (stage entry points, cd > 0.5, dwell, exotic step sizes).
// LEAN (implies FAST): additionally one step per launch (io.T == 1), N a multiple of the warp's env count, D == TS, obs and
//       statistics requested, every per-group array on the TMA path (16-byte multiples for any whole number of groups) —
//       checked by the host (launch_warp_ts).  Folds the rollout / alignment bookkeeping of the group loop away; the
//       arithmetic of the four stages is the same code.
// The grid is persistent (resident CTAs only): prologue, LUT build and the statistics epilogue are paid once per warp, and
// the int64[8] statistics cost six RED.ADD per CTA.
template <int TS, typename OT, int MODE, int WC, bool FAST, bool LEAN = false>
__global__ void __launch_bounds__(SWARM_MAXT, SWARM_MINB) step_warp_kernel(const __grid_constant__ DevParams p, const __grid_constant__ StepIO io) {
  constexpr int EPW = 32 / TS;
  constexpr bool WIN5 = (MODE == 1 && WC == 5);
  constexpr int LUTN = (MODE == 1 ? 256 : 16) * (sizeof(OT) == 4 ? 4 : 1);
  extern __shared__ __align__(128) unsigned char smem_raw[];
  __shared__ uint64_t full_bar[MAX_STAGES];       // per input stage: "inputs landed" (TMA transaction count)
  __shared__ int s_cnt[MAX_STAGES];               // warps that have left the stage in its current use
  __shared__ int s_cg[MAX_STAGES];                // io.dyn: the CTA-group the stage holds (-1: no work left)
  __shared__ unsigned long long s_stats[6];
  __shared__ __align__(16) uint32_t lut[LUTN];    // nibble pair (lo4 | hi4 << 4) -> four cells (float4 / 4 bytes); FULL: lo4 only
  __shared__ uint32_t padx[WIN5 ? 256 : 1], pady[WIN5 ? 256 : 1];   // 5x5 pad masks by centre column / row
  __shared__ float2 s_dq[8];                      // position delta by action code (noVis_sim.py:121-128)
  const int lane = threadIdx.x & 31, wpb = blockDim.x >> 5;
  const int wib = __shfl_sync(FULL, (int)(threadIdx.x >> 5), 0);   // provably warp-uniform: group bookkeeping runs on the uniform datapath
  const int flags = FAST ? (F_ALL | (io.flags & F_RESET)) : io.flags;
  const bool cd_fast = FAST ? true : (p.cd_fast != 0);
  const bool has_wait = FAST ? false : (p.has_wait != 0);
  const bool want_stats = LEAN ? true : (io.stats != nullptr && (flags & F_REWARD));
  const ObsCvt<OT> cvt(p);
  const int S = p.n_stages;
  if (threadIdx.x == 0) {   // the thread that initialises the barriers also issues the first S stage loads (below), BEFORE the
    for (int s = 0; s < S; ++s) mbar_init(&full_bar[s], 1);   // CTA builds its tables: their HBM latency overlaps the prologue
    fence_mbar_init();
  }
  // CTA shared memory: [obs staging: wpb x stage_bytes][input stage 0] ... [input stage S-1][slack]
  OT* const stage = reinterpret_cast<OT*>(smem_raw + (size_t)wib * p.stage_bytes);   // this warp's [EPW][D][W*W] staging (WINDOW)
  unsigned char* const inbuf0 = smem_raw + (size_t)wpb * p.stage_bytes;
  const int tile = lane / TS, tl = lane % TS, lead = tile * TS;
  const int D = LEAN ? TS : p.D, N = p.N, env_words = p.env_words;
  const int T = LEAN ? 1 : io.T;
  const int ngroups = (N + EPW - 1) / EPW;
  const int EPC = wpb * EPW;                      // envs per CTA-iteration (wpb consecutive groups)
  const int ncg = (N + EPC - 1) / EPC;            // CTA-groups
  const bool dyn = io.dyn != 0;
  const bool want_obs = LEAN ? true : ((flags & F_OBS) && io.obs_out != nullptr);
  const int WW = MODE == 1 ? (sizeof(OT) == 8 ? 1 : (WC ? WC : p.W) * (WC ? WC : p.W)) : 1;   // elements per drone (packed: one word)
  OT* const my_stage = stage + (tile * D + tl) * WW;
  // which per-group arrays may ride the bulk copy (size and address multiples of 16 B; planes, lists and meta always do)
  const bool act_tma_ok = LEAN ? true : (flags & F_MOVE) && io.act_dtype == 0 && (((uintptr_t)io.actions) & 15) == 0;
  const bool meta_tma = LEAN ? true : (flags & F_REWARD) != 0;
  // statistics accumulators are warp-uniform (REDUX results), so they live in uniform registers
  long long acc_rew = 0, acc_len = 0, acc_ret = 0;
  int acc_done = 0, acc_cells = 0, acc_anom = 0;

  // The two big streams — coverage planes in, observations out — are touched once per step: evict-first keeps them from
  // pushing the small read-modify-write state (positions, meta) out of L2 (measured in round 1: only the pair pays).
  const uint64_t pol_ef = l2_policy_evict_first();
  const uint32_t in_s = smem_u32(inbuf0), full_s = smem_u32(&full_bar[0]), stage_s = smem_u32(stage), lut_s = smem_u32(lut), cnt_s = smem_u32(&s_cnt[0]);
  const bool pos_tma_c = LEAN ? true : ((EPC * D * 8) & 15) == 0;
  const bool act_tma_c = LEAN ? true : (act_tma_ok && ((EPC * D) & 15) == 0 && io.T == 1);
  const uint32_t cbytes_c = (uint32_t)EPC * (uint32_t)env_words * 4u, posb_c = (uint32_t)EPC * (uint32_t)D * 8u;
  const uint32_t tx_c = cbytes_c + (uint32_t)EPC * 32u + (pos_tma_c ? posb_c : 0u) + (meta_tma ? (uint32_t)EPC * 16u : 0u) + (act_tma_c ? (uint32_t)(EPC * D) : 0u);
  // Request the inputs of CTA-group `cg` into stage `s`: up to five bulk copies on full_bar[s] (one thread).  The partial
  // tail CTA-group moves only planes, lists and meta by TMA (their sizes are always multiples of 16 B) and its warps load
  // positions / actions directly.
  auto issue = [&](int cg, int s) {
    const int e0 = cg * EPC;
    const uint32_t dst = in_s + (uint32_t)s * (uint32_t)p.in_bytes, bs = full_s + 8u * (uint32_t)s;
    if (LEAN && e0 + EPC > N) {   // the last CTA-group holds fewer (whole) groups: every size stays a multiple of 16 B
      const uint32_t ne = (uint32_t)(N - e0);
      mbar_expect_tx_s(bs, ne * ((uint32_t)env_words * 4u + 32u + (uint32_t)D * 8u + 16u + (uint32_t)D));
      bulk_g2s_s_hint(dst, p.cov + (size_t)e0 * env_words, ne * (uint32_t)env_words * 4u, bs, pol_ef);
      bulk_g2s_s(dst + p.off_alist, p.alist + (size_t)e0 * 8, ne * 32u, bs);
      bulk_g2s_s(dst + p.off_pos, p.pos + (size_t)e0 * D, ne * (uint32_t)D * 8u, bs);
      bulk_g2s_s(dst + p.off_meta, p.meta + e0, ne * 16u, bs);
      bulk_g2s_s(dst + p.off_act, (const uint8_t*)io.actions + (size_t)e0 * D, ne * (uint32_t)D, bs);
    } else if (LEAN || e0 + EPC <= N) {
      mbar_expect_tx_s(bs, tx_c);
#ifdef SWARM_EXP_NOHINT_LD
      bulk_g2s_s(dst, p.cov + (size_t)e0 * env_words, cbytes_c, bs);
#else
      bulk_g2s_s_hint(dst, p.cov + (size_t)e0 * env_words, cbytes_c, bs, pol_ef);
#endif
      bulk_g2s_s(dst + p.off_alist, p.alist + (size_t)e0 * 8, (uint32_t)EPC * 32u, bs);
      if (pos_tma_c) bulk_g2s_s(dst + p.off_pos, p.pos + (size_t)e0 * D, posb_c, bs);
      if (meta_tma) bulk_g2s_s(dst + p.off_meta, p.meta + e0, (uint32_t)EPC * 16u, bs);
      if (act_tma_c) bulk_g2s_s(dst + p.off_act, (const uint8_t*)io.actions + (size_t)e0 * D, (uint32_t)(EPC * D), bs);
    } else {
      const uint32_t ne = (uint32_t)(N - e0);
      mbar_expect_tx_s(bs, ne * (uint32_t)env_words * 4u + ne * 32u + (meta_tma ? ne * 16u : 0u));
      bulk_g2s_s(dst, p.cov + (size_t)e0 * env_words, ne * (uint32_t)env_words * 4u, bs);
      bulk_g2s_s(dst + p.off_alist, p.alist + (size_t)e0 * 8, ne * 32u, bs);
      if (meta_tma) bulk_g2s_s(dst + p.off_meta, p.meta + e0, ne * 16u, bs);
    }
  };
  // The CTA-group of this CTA's k-th iteration goes into stage k % S: claim it (dynamic) or compute it (static), then either
  // request its inputs or — dynamic schedule only — tell the waiting warps that nothing is left.
  auto refill = [&](int k, int s) {
    int cg;
    if (dyn) {
      cg = atomicAdd(p.sched, 1);
      if (cg >= ncg) cg = -1;
      s_cg[s] = cg;
      if (cg < 0) { mbar_arrive_s(full_s + 8u * (uint32_t)s); return; }
    } else {
#ifdef SWARM_EXP_BLOCKED
      { const int per = (ncg + (int)gridDim.x - 1) / (int)gridDim.x; cg = (k < per) ? (int)blockIdx.x * per + k : ncg; }
#else
      cg = (int)blockIdx.x + k * (int)gridDim.x;
#endif
      if (cg >= ncg) return;
    }
    issue(cg, s);
  };
  if (threadIdx.x == 0) {
    pdl_wait();   // everything above touched only shared memory and kernel parameters; from here on we read what earlier launches wrote
    if (p.peers.n_ranks > 0 && (io.flags & F_PUSH) && io.stats && blockIdx.x == 0)
      peer_stats_push(p.peers, io.stats);   // multi-GPU: publish the totals of the launches before this one (see PeerStats)
    for (int k = 0; k < S; ++k) refill(k, k);
  }
  if (threadIdx.x < 6) s_stats[threadIdx.x] = 0ull;
  if (threadIdx.x < MAX_STAGES) s_cnt[threadIdx.x] = 0;
  if (threadIdx.x < 8)
    s_dq[threadIdx.x] = make_float2(threadIdx.x == 1 ? p.ss : (threadIdx.x == 3 ? -p.ss : 0.f), threadIdx.x == 2 ? p.ss : (threadIdx.x == 0 ? -p.ss : 0.f));
  build_window_tables<OT>(p, cvt, lut, MODE == 1 ? 256 : 16, WIN5 ? padx : nullptr, WIN5 ? pady : nullptr, threadIdx.x, blockDim.x);
  pdl_wait();   // every thread, before its own global accesses (thread 0 has passed it already)
  __syncthreads();

  uint32_t par = 0u;   // parity of the current use of the stages (flips every S iterations)
  int s = 0;
  for (int it = 0;; ++it) {
#ifdef SWARM_EXP_BLOCKED
    const int per_b = (ncg + (int)gridDim.x - 1) / (int)gridDim.x;
    int cg = (it < per_b) ? (int)blockIdx.x * per_b + it : ncg;
#else
    int cg = (int)blockIdx.x + it * (int)gridDim.x;
#endif
    if (!dyn && cg >= ncg) break;
    mbar_wait_s(full_s + 8u * (uint32_t)s, par);
    if (dyn) {
      cg = s_cg[s];
      if (cg < 0) break;
    }
    const int grp = cg * wpb + wib;
    if (grp < ngroups) {
    unsigned char* const in = inbuf0 + s * p.in_bytes;
    uint32_t* const sgrid = reinterpret_cast<uint32_t*>(in) + wib * EPW * env_words;          // the group's planes
    uint32_t* const slist = reinterpret_cast<uint32_t*>(in + p.off_alist) + wib * EPW * 8;     // the group's lists
    const bool cta_full = LEAN ? true : (cg * EPC + EPC <= N);
    const int env0 = grp * EPW, nenv = LEAN ? EPW : min(EPW, N - env0);
    const bool valid = LEAN ? true : (tile < nenv);
    const int env = valid ? env0 + tile : env0;
    const bool active = LEAN ? true : (valid && tl < D);
    uint32_t* const sc = sgrid + tile * env_words;
    uint32_t* const sl = slist + tile * 8;
    const uint32_t sl_s = in_s + (uint32_t)s * (uint32_t)p.in_bytes + (uint32_t)p.off_alist + (uint32_t)((wib * EPW + tile) * 32);
    const uint32_t sc_s = in_s + (uint32_t)s * (uint32_t)p.in_bytes + (uint32_t)((wib * EPW + tile) * env_words) * 4u;
    uint32_t* const gc = p.cov + (size_t)env * env_words;
    const bool pos_tma = pos_tma_c && cta_full, act_tma = act_tma_c && cta_full;
    float2* const pp = p.pos + (size_t)env * D + (active ? tl : 0);
    size_t aidx = (size_t)env * D + tl;                 // action index, advanced by N*D per step
    size_t ridx = (size_t)env;                          // reward / done index, advanced by N per step

    float2 q = make_float2(0.f, 0.f);
    int a = 0;
    int4 m = make_int4(0, 0, 0, 0);
    if (active) q = pos_tma ? reinterpret_cast<const float2*>(in + p.off_pos)[(wib * EPW + tile) * D + tl] : __ldcg(pp);
    if (active && (flags & F_MOVE)) {
      if (act_tma) a = (int)in[p.off_act + (wib * EPW + tile) * D + tl];
      else a = load_action(io.actions, io.act_dtype, (long long)aidx);
    }
    if (valid && tl == 0) m = meta_tma ? reinterpret_cast<const int4*>(in + p.off_meta)[wib * EPW + tile] : make_int4(0, 0, 0, 0);

    int a_next = 0;   // rollout: the NEXT step's action is requested a whole step ahead, so its HBM latency is off the chain
    for (int t = 0; t < T; ++t, aidx += (size_t)N * D, ridx += N) {
      if (t > 0) a = a_next;   // rollout: positions / meta are carried in registers, actions are re-read
      if (!LEAN && t + 1 < T && active && (flags & F_MOVE)) {
        const size_t an = aidx + (size_t)N * D;
        a_next = FAST ? (int)__ldg((const uint8_t*)io.actions + an) : load_action(io.actions, io.act_dtype, (long long)an);   // FAST rollouts: uint8 (host-checked)
      }

      // ---- stage (1): action decode + move + clip (noVis_sim.py:117-135)
      if (active && (flags & F_MOVE)) {
        if ((unsigned)a > 3u) atomicOr(p.err, 1);
        const float2 dq = s_dq[(unsigned)a > 3u ? 4 : a];   // a=0: y-=ss, 1: x+=ss, 2: y+=ss, 3: x-=ss; anything else: no move
        q.x = fminf(fmaxf(q.x + dq.x, 0.f), p.xmax);
        q.y = fminf(fmaxf(q.y + dq.y, 0.f), p.ymax);
        __stcg(pp, q);
      }
      int cx = (int)floorf(q.x + 0.5f), cy = (int)floorf(q.y + 0.5f);   // nearest lattice cell = window centre
      uint32_t A = 0u;                                                    // WIN5: anomalous cells inside the drone's window
      if constexpr (WIN5) A = window_anomalies5(sl_s, p.n_anom, cx, cy);

      // ---- stage (2): coverage scatter on the shared-memory plane
      int rew = 0, ncl = 0, an_w = 0;
      if (flags & F_SCATTER) {
        int pkv = 0;   // this lane's (reward | cells << 10)
        if (cd_fast) {   // cd <= 0.5: the only candidate is the nearest lattice cell
          bool cand = false;
          if (active) {
            if (FAST) {   // float32-exact coordinates (DevParams::f32_exact)
              const float dx = (float)cx - q.x, dy = (float)cy - q.y;
              cand = __fmaf_rn(dx, dx, dy * dy) < p.cd2_thr_f;          // sqrt(d2) < collision_dist (noVis_sim.py:81)
            } else {
              const double dx = (double)((float)cx - q.x), dy = (double)((float)cy - q.y);   // both differences are exact
              cand = dx * dx + dy * dy < p.cd2_thr;
            }
          }
          const int cell = cy * p.R + cx;
          const uint32_t bit = 1u << (cell & 31);
          bool won = false;
          if (FAST) {
            // Several drones of the env on one unmapped cell: the shared-memory atomic elects exactly one of them (the one
            // that still saw the bit), so the cell is cleared and rewarded once — the env's reward is a sum, the order does
            // not matter.
            if (cand && (sc[cell >> 5] & bit)) won = (atomicAnd(sc + (cell >> 5), ~bit) & bit) != 0u;
          } else {
            // generic: drones of one env on one cell form a match group; its lowest lane resolves the conflict for all of
            // them (the dwell counters need the group's size)
            const unsigned key = cand ? (((unsigned)tile << 24) | (unsigned)cell) : (0x80000000u | (unsigned)lane);
            const unsigned mm = __match_any_sync(FULL, key);
            if (cand && lane == __ffs(mm) - 1 && (sc[cell >> 5] & bit)) {
              won = true;
              if (has_wait) {   // dwell (App. B.2): k drones in index order, each decrements or (at 0) maps
                const int k = __popc(mm);
                int16_t* w = p.wait + (size_t)env * p.RC + cell;
                const int wv = *w;
                if (wv >= k) { *w = (int16_t)(wv - k); won = false; }
                else if (wv > 0) *w = 0;
              }
              if (won) atomicAnd(sc + (cell >> 5), ~bit);   // another leader of this tile may clear a cell of the same word
            }
          }
          if (won) {
            atomicAnd(gc + (cell >> 5), ~bit);   // write-through to HBM (RED.AND, fire and forget)
            int an;
            if constexpr (WIN5) an = (int)((A >> 12) & 1u);   // the window's centre cell
            else {
              const uint32_t key = (uint32_t)cx | ((uint32_t)cy << 16);
              an = 0;
              for (int k = 0; k < p.n_anom; ++k) an |= (sl[k] == key) ? 1 : 0;
            }
            pkv = (10 + 20 * an) | (1 << 10);    // 10 + _check_uncertain_mapped (noVis_sim.py:83-84,104-110)
            an_w = an;
          }
        } else {
          __syncwarp();
          if (valid && tl == 0) {
            int r2 = 0, c2 = 0, n2 = 0;
            scatter_sequential(p, env, sc, r2, c2, n2);
            pkv = r2 | (c2 << 10);
            an_w = n2;
          }
        }
        // tile totals: reward <= 30 * 32 < 2^10, cells <= 32 < 2^6 -> 16 bits per tile
        if constexpr (TS >= 16) {   // both tiles in ONE warp-wide REDUX
          const unsigned tot = __reduce_add_sync(FULL, (unsigned)pkv << (16 * tile));
          pkv = (int)((tot >> (16 * tile)) & 0xffffu);
        } else {
#pragma unroll
          for (int o = TS / 2; o > 0; o >>= 1) pkv += __shfl_xor_sync(FULL, pkv, o);
        }
        rew = pkv & 1023; ncl = pkv >> 10;
      }
      __syncwarp();   // the tile's clears are visible in shared memory

      // ---- stage (3): reward tail, spread, done; auto-reset
      if (flags & F_REWARD) {
        const int tcount = __shfl_sync(FULL, m.x, lead);
        const bool punish = tcount > p.punish_thr;                      // noVis_sim.py:141-143 (count before increment)
        const bool need = valid && !punish;
        bool spread = false;
        if (__any_sync(FULL, need)) spread = tile_spread<TS, FAST>(p, env, tile, tl, need, q.x, q.y);
        bool do_reset = false;
        int s_r = 0, s_len = 0, s_ret = 0;   // this lane's contribution to the statistics
        bool s_done = false;
        if (valid && tl == 0) {
          int r = (flags & F_SCATTER) ? rew : io.cover_in[env];
          if (p.reward_mode == 0) r += ((!punish && spread) ? 1 : 0) - 1;   // A2C noVis_sim.py:150-152
          else r += punish ? -1 : (spread ? 1 : 0);                         // DQL noVis_sim.py:141-148
          const int rem = m.y - ncl;
          const bool done = rem == 0;                                       // noVis_sim.py:157-166
          if (p.reward_mode == 0 && done) r += 100;                         // noVis_sim.py:172-173
          const int t1 = tcount + 1, epret = m.z + r;
          io.reward_out[ridx] = (float)r;
          io.done_out[ridx] = done ? 1 : 0;
          s_r = r;
          if (done && m.y > 0) { s_done = true; s_len = t1; s_ret = epret; }
          do_reset = done && (flags & F_RESET);
          if (!do_reset) {
            m = make_int4(t1, rem, epret, m.w);
            __stcg(p.meta + env, m);
          } else {
            m = make_int4(0, p.RC, 0, m.w + 1);   // what reset_tile_smem writes
          }
        }
        if (want_stats) {
          acc_rew += __reduce_add_sync(FULL, s_r);
          acc_cells += __reduce_add_sync(FULL, (valid && tl == 0) ? ncl : 0);
          acc_anom += __reduce_add_sync(FULL, an_w);
          if (__any_sync(FULL, s_done)) {   // rare: an episode finished
            acc_done += __popc(__ballot_sync(FULL, s_done));
            acc_len += __reduce_add_sync(FULL, s_len);
            acc_ret += __reduce_add_sync(FULL, s_ret);
          }
        }
        if (flags & F_RESET) {
          do_reset = __shfl_sync(FULL, (int)do_reset, lead) != 0;
          if (__any_sync(FULL, do_reset)) {   // rare: an episode finished
            const int episode = __shfl_sync(FULL, m.w, lead);   // already the new episode index where do_reset
            reset_tile_smem<TS>(p, env, episode, tl, do_reset, sc, sl, q);
            cx = (int)floorf(q.x + 0.5f); cy = (int)floorf(q.y + 0.5f);
            if constexpr (WIN5) A = window_anomalies5(sl_s, p.n_anom, cx, cy);
          }
        }
      } else if (flags & F_SCATTER) {   // staged scatter (swarm_stage_scatter)
        if (valid && tl == 0) {
          io.cover_out[env] = rew;
          if (ncl) p.meta[env].y -= ncl;
        }
      }

      // ---- stage (4): observation emission
      if (want_obs && (LEAN || !io.obs_last_only || t == T - 1)) {
        OT* obase = reinterpret_cast<OT*>(io.obs_out) + ((LEAN || io.obs_last_only) ? (size_t)0 : (size_t)t * N * p.obs_per_env);
        if (MODE == 0) {
          emit_full_lut<OT>(p, cvt, lut, sgrid, slist, nenv, lane, obase + (size_t)env0 * p.RC);
        } else {
          if (lane == 0) bulk_wait_read0();   // the previous bulk store has drained `stage` (long ago)
          __syncwarp();
#ifndef SWARM_EXP_NOWINDOW   // experiment build: the obs stream without the window arithmetic
          if (active) {
            if constexpr (WIN5) build_window5<OT>(p.R, sc_s, lut_s, padx[cx] | pady[cy], A, cx - 2, cy - 2, my_stage);
            else build_window_generic<OT>(p, cvt, sc, sl, cx, cy, my_stage);
          }
#endif
          fence_proxy_async();   // generic-proxy writes to `stage` ordered before the async-proxy read below
          __syncwarp();
          const int n = nenv * D * WW;
          OT* dst = obase + (size_t)env0 * D * WW;
          const uint32_t bytes = (uint32_t)n * (uint32_t)sizeof(OT);
          if (LEAN || ((((uintptr_t)dst) & 15) == 0 && (bytes & 15u) == 0)) {
#ifndef SWARM_EXP_NOSTORE   // experiment build: everything but the obs stream (what the compute side alone costs)
            if (lane == 0) {   // one TMA bulk store for the whole group
#ifdef SWARM_EXP_NOHINT_ST
              bulk_s2g_s(dst, stage_s, bytes);
#else
              bulk_s2g_s_hint(dst, stage_s, bytes, pol_ef);
#endif
              bulk_commit();
            }
#endif
          } else {
            for (int k = lane; k < n; k += 32) __stcs(dst + k, stage[k]);
          }
        }
      }
      __syncwarp();
    }
    }   // grp < ngroups
    // leave the stage; the last warp out refills it with the CTA-group S iterations ahead.  This group's generic accesses
    // to the stage must precede the bulk copies that refill it: in the LEAN WINDOW kernel the proxy fence + __syncwarp issued
    // before the obs bulk store already follows the last of them.
    if (!(LEAN && MODE == 1)) {
      fence_proxy_async();
      __syncwarp();
    }
    if (lane == 0) {
      if ((int)smem_atomic_inc(cnt_s + 4u * (uint32_t)s) == wpb - 1) {
        s_cnt[s] = 0;
        refill(it + S, s);
      }
    }
    if (++s == S) { s = 0; par ^= 1u; }
  }
  pdl_launch_dependents();   // main loop done: the next launch on the stream may bring its CTAs in (they run their prologue, then pdl_wait)
  if (lane == 0) bulk_wait_read0();   // shared memory must outlive the last bulk store's reads

  if (want_stats) {   // one lane per warp adds the warp's (uniform) totals; zero entries are skipped
    if (lane == 0) {
      const long long v[6] = {acc_rew, (long long)acc_done, acc_len, acc_ret, (long long)acc_cells, (long long)acc_anom};
#pragma unroll
      for (int k = 0; k < 6; ++k)
        if (v[k] != 0) atomicAdd(&s_stats[k], (unsigned long long)v[k]);
    }
    __syncthreads();
    if (threadIdx.x < 6 && s_stats[threadIdx.x] != 0ull) atomicAdd(io.stats + 2 + threadIdx.x, s_stats[threadIdx.x]);
    if (blockIdx.x == 0 && threadIdx.x == 0) {
      atomicAdd(io.stats + 0, (unsigned long long)((long long)N * T));
      atomicAdd(io.stats + 1, (unsigned long long)((long long)N * D * T));
    }
  }
  if (dyn) {   // the last CTA out rearms the scheduler for the next launch (every CTA has made its last claim by now)
    __syncthreads();
    if (threadIdx.x == 0 && atomicAdd(p.sched + 1, 1) == (int)gridDim.x - 1) {
      p.sched[0] = 0;
      p.sched[1] = 0;
      __threadfence();
    }
  }
}

}  // namespace swarm
