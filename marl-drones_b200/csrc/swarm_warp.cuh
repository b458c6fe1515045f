// swarm_warp.cuh — the fused env step as a warp-autonomous, shared-memory-staged kernel (the hot path for D <= 32 and
// grids whose packed form fits a few KiB: every BASELINE config except the 256x256 stress case).
//
// One warp owns EPW = 32/TS consecutive envs (a "group"); a tile of TS lanes serves one env, one lane per drone.
// Per group the warp
//   (a) receives the group's inputs — packed grids (EPW * env_words * 4 bytes, contiguous in HBM), positions, meta and
//       actions — in ITS OWN slice of shared memory through TMA bulk copies (cp.async.bulk, one mbarrier per buffer),
//       requested one group ahead into a double buffer, so HBM latency overlaps the previous group's arithmetic;
//   (b) runs stages (1)-(3) on the shared-memory copy (clears are written through to HBM with RED.AND);
//   (c) WINDOW obs: every lane gathers its drone's W x W window from shared memory into a staging slice (bank-conflict
//       free: lane stride W*W words), and one TMA bulk store streams the group's [EPW][D][W*W] block to the obs tensor;
//       FULL obs: the warp expands the 2-bit grid straight from shared memory with 128-bit streaming stores.
// There is no __syncthreads in the group loop: warps run load / compute / store independently.  T > 1 (swarm_rollout)
// keeps grid, positions and meta on chip across steps.
//
// Reference lines per stage: see swarm_step.cuh.
#pragma once
#include "swarm_common.cuh"
#include "swarm_emit.cuh"
#include "swarm_step.cuh"

#ifndef SWARM_TRACE
#define SWARM_TRACE 0   // 1: instrumented build for tools/trace.py (per-CTA start / end / SM id / input-wait cycles, CTA 0 timeline)
#endif
#ifndef SWARM_MINB
#define SWARM_MINB 3   // resident CTAs per SM the staged kernel is compiled for (register cap 65536 / (256 * SWARM_MINB))
#endif

namespace swarm {

#if SWARM_TRACE
__device__ unsigned long long g_trace[16 * 8 * 8];   // CTA 0: clock64 at events 0..6 of its first 16 iterations, per warp
__device__ unsigned long long g_cta_t[2048 * 2];     // [2b], [2b+1]: globaltimer at CTA b's start / end; [2048+b] wait cycles; [3072+b] SM id
__device__ __forceinline__ unsigned long long gtimer() { unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); return t; }
#define TR(ev) do { if (blockIdx.x == 0 && lane == 0 && it < 16) g_trace[(it * 8 + wib) * 8 + (ev)] = clock64(); } while (0)
#else
#define TR(ev) do { } while (0)
#endif

// init_env (A2C/noVis_sim.py:36-63) for the env of this tile, built in shared memory and copied out to HBM.
// Same RNG spec as reset_tile / oracle_reset_env.  Called by all 32 lanes; `doit` is tile-uniform; q = the lane's new position.
template <int TS>
__device__ __forceinline__ void reset_tile_smem(const DevParams& p, int env, int episode, int tl, bool doit, uint32_t* sg, float2& q) {
  const long long gid = p.env_id_base + env;
  int16_t* wt = p.has_wait ? p.wait + (size_t)(doit ? env : 0) * p.RC : nullptr;
  if (doit) {
    for (int w = tl; w < p.env_words; w += TS) {
      const int n = p.RC - w * 16;
      sg[w] = n >= 16 ? 0x55555555u : (n <= 0 ? 0u : (0x55555555u & ((1u << (2 * n)) - 1u)));
    }
    if (wt)
      for (int k = tl; k < p.RC; k += TS) wt[k] = (int16_t)p.dwell_normal;
  }
  __syncwarp();
  if (doit && tl == 0) {
    int placed = 0;
    uint4 r = make_uint4(0, 0, 0, 0);
    for (uint32_t j = 0; placed < p.n_anom; ++j) {
      if ((j & 1u) == 0) r = draw_block(p, gid, (uint32_t)episode, j >> 1);
      const uint32_t a = __umulhi((j & 1u) ? r.z : r.x, (uint32_t)p.R);
      const uint32_t b = __umulhi((j & 1u) ? r.w : r.y, (uint32_t)p.C);
      const int cell = (int)(b * (uint32_t)p.R + a);
      const int sh = (cell & 15) * 2;
      const uint32_t word = sg[cell >> 4];
      if (((word >> sh) & 3u) == 1u) {
        sg[cell >> 4] = word ^ (3u << sh);   // code 1 -> 2
        if (wt) wt[cell] = (int16_t)p.dwell_anom;
        ++placed;
      }
    }
    p.meta[env] = make_int4(0, p.RC, 0, episode);
  }
  __syncwarp();
  if (doit) {
    uint4* g4 = reinterpret_cast<uint4*>(p.grid + (size_t)env * p.env_words);
    const uint4* s4 = reinterpret_cast<const uint4*>(sg);
    for (int w = tl; w < (p.env_words >> 2); w += TS) __stcg(g4 + w, s4[w]);
    if (tl < p.D) {
      const uint4 r = draw_block(p, gid, (uint32_t)episode, 0x80000000u | ((uint32_t)tl >> 1));
      q.x = (float)__umulhi((tl & 1) ? r.z : r.x, (uint32_t)p.R);
      q.y = (float)__umulhi((tl & 1) ? r.w : r.y, (uint32_t)p.C);
      __stcg(p.pos + (size_t)env * p.D + tl, q);
    }
  }
  __syncwarp();
}

// Stage (4), WINDOW: the lane's drone gathers its W x W window from the shared-memory grid `sg` into
// st[0 .. W*W).  For W <= 15 a window row is one funnel-shifted 32-bit field (16 cells) of the 2-bit stream; cells that
// belong to out-of-range columns are forced to the pad code, which also covers the neighbouring-row / out-of-buffer bits
// such a field can contain (the buffer has readable slack on both sides).
// float obs: `lut4` (shared, 256 x float4) maps a byte of four 2-bit codes to its four float values, so a row costs one
// LDS.128 per four cells instead of per-cell bit twiddling.
template <int WC, typename OT>
__device__ __forceinline__ void build_window(const DevParams& p, const ObsCvt<OT>& cvt, const float4* lut4, const uint32_t* sg,
                                             float2 q, OT* st) {
  const int W = WC ? WC : p.W;
  const int h = (W - 1) >> 1;
  const int cx = (int)floorf(q.x + 0.5f), cy = (int)floorf(q.y + 0.5f);
  if (W <= 15) {
    uint32_t colpad = 0u;
#pragma unroll
    for (int a = 0; a < W; ++a)
      if ((unsigned)(cx - h + a) >= (unsigned)p.R) colpad |= 3u << (2 * a);
#pragma unroll
    for (int b = 0; b < W; ++b) {
      const int gy = cy - h + b;
      uint32_t field = 0xffffffffu;
      if ((unsigned)gy < (unsigned)p.C) {
        const int s0 = gy * p.R + cx - h;   // may be < 0 (gy == 0, cx < h): arithmetic shift / mask still address the stream
        const int wi = s0 >> 4;
        field = __funnelshift_r(sg[wi], sg[wi + 1], (s0 & 15) * 2) | colpad;
      }
      if constexpr (sizeof(OT) == 4) {
        float* o = reinterpret_cast<float*>(st) + b * W;
#pragma unroll
        for (int a = 0; a < W; a += 4) {
          const char* e = reinterpret_cast<const char*>(lut4) + (((a ? (field >> (2 * a)) : field) & 0xffu) << 4);
          if (a + 4 <= W) {
            const float4 v = *reinterpret_cast<const float4*>(e);
            o[a] = v.x; o[a + 1] = v.y; o[a + 2] = v.z; o[a + 3] = v.w;
          } else if (a + 3 == W) {
            const float2 v = *reinterpret_cast<const float2*>(e);
            o[a] = v.x; o[a + 1] = v.y; o[a + 2] = reinterpret_cast<const float*>(e)[2];
          } else if (a + 2 == W) {
            const float2 v = *reinterpret_cast<const float2*>(e);
            o[a] = v.x; o[a + 1] = v.y;
          } else {
            o[a] = *reinterpret_cast<const float*>(e);
          }
        }
      } else {
#pragma unroll
        for (int a = 0; a < W; ++a) st[b * W + a] = cvt.sel(field, a);
      }
    }
  } else {
    for (int b = 0; b < W; ++b)
      for (int a = 0; a < W; ++a) {
        const int gx = cx - h + a, gy = cy - h + b;
        uint32_t code = 3u;
        if ((unsigned)gx < (unsigned)p.R && (unsigned)gy < (unsigned)p.C) {
          const int cell = gy * p.R + gx;
          code = (sg[cell >> 4] >> ((cell & 15) * 2)) & 3u;
        }
        st[b * W + a] = cvt.one(code);
      }
  }
}

// Stage (4), FULL: expand the group's packed grids from shared memory.  obs[n][y*R+x], envs of a group are contiguous.
template <typename OT>
__device__ __forceinline__ void emit_full_smem(const DevParams& p, const ObsCvt<OT>& cvt, const uint32_t* sgrid, int nenv, int lane,
                                               OT* __restrict__ o) {
  const int RC = p.RC;
  if ((RC & 3) == 0 && ((uintptr_t)o & 15) == 0) {
    const int nq = RC >> 2, total = nenv * nq;
#pragma unroll 2
    for (int v = lane; v < total; v += 32) {
      const int e = p.nq_magic ? (int)__umulhi((uint32_t)v, p.nq_magic) : v;   // nq == 1: one quad per env
      const int qd = v - e * nq;
      const uint32_t w = sgrid[e * p.env_words + (qd >> 2)];
      cvt.store4(o + (size_t)e * RC + 4 * qd, (w >> ((qd & 3) * 8)) & 0xffu);
    }
  } else {
    for (int e = 0; e < nenv; ++e)
      for (int c = lane; c < RC; c += 32) o[(size_t)e * RC + c] = cvt.one((sgrid[e * p.env_words + (c >> 4)] >> ((c & 15) * 2)) & 3u);
  }
}

// ------------------------------------------------------------------------------------------------ the kernel
// MODE: 0 FULL obs, 1 WINDOW obs.  WC: compile-time window size (5) or 0 = runtime p.W.
// FAST: the production configuration folded at compile time — all four stages in one pass (swarm_step / swarm_rollout),
//       uint8 actions, collision_dist <= 0.5, no dwell counters.  FAST = false keeps every switch at run time (stage entry
//       points, int32/int64 actions, cd > 0.5, dwell).
//
// Software pipeline: a warp's inputs for group g+1 (packed grids, positions, meta, actions — four TMA bulk copies on one
// mbarrier, into the other half of a double buffer) are requested BEFORE it computes group g, so HBM latency overlaps the
// arithmetic even when every warp of the GPU runs in lock step.  The copies need no registers.  The grid is persistent
// (resident CTAs only): prologue, LUT build and the statistics epilogue are paid once per warp, and the int64[8]
// statistics cost six RED.ADD per CTA.
// LEAN (implies FAST): additionally one step per launch (io.T == 1), N a multiple of the CTA's env count, D == TS, obs and
// statistics requested, every per-group array on the TMA path — checked by the host (launch_warp_ts).  Folds the tail /
// rollout / alignment bookkeeping of the group loop away; the arithmetic of the four stages is the same code.
template <int TS, typename OT, int MODE, int WC, bool FAST, bool LEAN = false>
__global__ void __launch_bounds__(256, SWARM_MINB) step_warp_kernel(const __grid_constant__ DevParams p, const __grid_constant__ StepIO io) {
  constexpr int EPW = 32 / TS;
  extern __shared__ __align__(128) unsigned char smem_raw[];
  __shared__ uint64_t full_bar[2], empty_bar[2];   // per input buffer: "inputs landed" (TMA tx count) / "all warps done with it"
  __shared__ unsigned long long s_stats[6];
  __shared__ float4 lut4[(MODE == 1 && sizeof(OT) == 4) ? 256 : 1];   // byte of four codes -> four floats (WINDOW, float obs)
  const int lane = threadIdx.x & 31, wpb = blockDim.x >> 5;
  const int wib = __shfl_sync(FULL, (int)(threadIdx.x >> 5), 0);   // provably warp-uniform: group bookkeeping runs on the uniform datapath
  const int flags = FAST ? (F_ALL | (io.flags & F_RESET)) : io.flags;
  const bool cd_fast = FAST ? true : (p.cd_fast != 0);
  const bool has_wait = FAST ? false : (p.has_wait != 0);
  const bool want_stats = LEAN ? true : (io.stats != nullptr && (flags & F_REWARD));
  const ObsCvt<OT> cvt(p);
  if (p.peers.n_ranks > 0 && (io.flags & F_PUSH) && io.stats && blockIdx.x == 0 && threadIdx.x == 0)
    peer_stats_push(p.peers, io.stats);   // multi-GPU: publish the totals of the launches before this one (see PeerStats)
  if (threadIdx.x < 6) s_stats[threadIdx.x] = 0ull;
  if constexpr (MODE == 1 && sizeof(OT) == 4) {
    for (int b = threadIdx.x; b < 256; b += blockDim.x) lut4[b] = make_float4(cvt.sel(b, 0), cvt.sel(b, 1), cvt.sel(b, 2), cvt.sel(b, 3));
  }
  if (threadIdx.x == 0) {
    mbar_init(&full_bar[0], 1);
    mbar_init(&full_bar[1], 1);
    mbar_init(&empty_bar[0], (uint32_t)wpb);
    mbar_init(&empty_bar[1], (uint32_t)wpb);
    fence_mbar_init();
  }
  __syncthreads();
  // CTA shared memory: [obs staging: wpb x stage_bytes][input buffer 0][input buffer 1][slack]
  OT* const stage = reinterpret_cast<OT*>(smem_raw + (size_t)wib * p.stage_bytes);   // this warp's [EPW][D][W*W] staging (WINDOW)
  unsigned char* const inbuf0 = smem_raw + (size_t)wpb * p.stage_bytes;              // two CTA input buffers of p.in_bytes each
  const int tile = lane / TS, tl = lane % TS, lead = tile * TS;
  const int D = LEAN ? TS : p.D, N = p.N, env_words = p.env_words;
  const int T = LEAN ? 1 : io.T;
  const int ngroups = (N + EPW - 1) / EPW;
  const int EPC = wpb * EPW;                      // envs per CTA-iteration (wpb consecutive groups)
  const int ncg = (N + EPC - 1) / EPC;            // CTA-groups
  const bool want_obs = LEAN ? true : ((flags & F_OBS) && io.obs_out != nullptr);
  const int WW = MODE == 1 ? (WC ? WC : p.W) * (WC ? WC : p.W) : 1;
  OT* const my_stage = stage + (tile * D + tl) * WW;
  // which per-group arrays may ride the bulk copy (size and address multiples of 16 B; the grid and meta always do)
  const bool act_tma_ok = LEAN ? true : (flags & F_MOVE) && (FAST || io.act_dtype == 0) && (((uintptr_t)io.actions) & 15) == 0;
  const bool meta_tma = LEAN ? true : (flags & F_REWARD) != 0;
  // statistics accumulators are warp-uniform (REDUX results), so they live in uniform registers
  long long acc_rew = 0, acc_len = 0, acc_ret = 0;
  int acc_done = 0, acc_cells = 0, acc_anom = 0;

  // Producer (thread 0 of the CTA): request the inputs of CTA-group `cg` — wpb consecutive groups, contiguous in every
  // array — into input buffer `b`: four bulk copies on full_bar[b].  The partial tail CTA-group moves only its grid and
  // meta by TMA (their sizes are always multiples of 16 B) and its warps load positions / actions directly.
  // The two big streams — packed grids in, observations out — are touched once per step: evict-first keeps them from pushing
  // the small read-modify-write state (positions, meta) out of L2.  Measured 34.3 -> 33.5 us on cfg3; either hint alone,
  // or hints on the small copies too, are slower than none.
  const uint64_t pol_ef = l2_policy_evict_first();
  const uint32_t in_s = smem_u32(inbuf0), full_s = smem_u32(&full_bar[0]), empty_s = smem_u32(&empty_bar[0]), stage_s = smem_u32(stage);
  const bool pos_tma_c = LEAN ? true : ((EPC * D * 8) & 15) == 0;
  const bool act_tma_c = LEAN ? true : (act_tma_ok && ((EPC * D) & 15) == 0 && io.T == 1);
  const uint32_t gbytes_c = (uint32_t)EPC * (uint32_t)env_words * 4u, posb_c = (uint32_t)EPC * (uint32_t)D * 8u;
  const uint32_t tx_c = gbytes_c + (pos_tma_c ? posb_c : 0u) + (meta_tma ? (uint32_t)EPC * 16u : 0u) + (act_tma_c ? (uint32_t)(EPC * D) : 0u);
  auto produce = [&](int cg, int b) {
    const int e0 = cg * EPC;
    const uint32_t dst = in_s + (uint32_t)b * (uint32_t)p.in_bytes, bs = full_s + 8u * (uint32_t)b;
    if (LEAN || e0 + EPC <= N) {
      mbar_expect_tx_s(bs, tx_c);
      bulk_g2s_s_hint(dst, p.grid + (size_t)e0 * env_words, gbytes_c, bs, pol_ef);
      if (pos_tma_c) bulk_g2s_s(dst + p.off_pos, p.pos + (size_t)e0 * D, posb_c, bs);
      if (meta_tma) bulk_g2s_s(dst + p.off_meta, p.meta + e0, (uint32_t)EPC * 16u, bs);
      if (act_tma_c) bulk_g2s_s(dst + p.off_act, (const uint8_t*)io.actions + (size_t)e0 * D, (uint32_t)(EPC * D), bs);
    } else {
      const uint32_t ne = (uint32_t)(N - e0);
      mbar_expect_tx_s(bs, ne * (uint32_t)env_words * 4u + (meta_tma ? ne * 16u : 0u));
      bulk_g2s_s(dst, p.grid + (size_t)e0 * env_words, ne * (uint32_t)env_words * 4u, bs);
      if (meta_tma) bulk_g2s_s(dst + p.off_meta, p.meta + e0, ne * 16u, bs);
    }
  };

#if SWARM_TRACE
  if (threadIdx.x == 0) { unsigned sm; asm volatile("mov.u32 %0, %%smid;" : "=r"(sm)); g_cta_t[blockIdx.x * 2] = gtimer(); g_cta_t[2048 + blockIdx.x] = 0; g_cta_t[1024 * 3 + blockIdx.x] = sm; }
#endif
  uint32_t phases = 0;
  const bool producer = threadIdx.x == 0;
  if (producer && (int)blockIdx.x < ncg) produce(blockIdx.x, 0);
  int it = 0;
  for (int cg = blockIdx.x; cg < ncg; cg += gridDim.x, ++it) {
    const int b = it & 1;
    // refill the other buffer for the CTA's next iteration as soon as every warp has released it (it was read in iteration
    // it-1; its (it-1)/2-th release completes phase ((it-1)/2) & 1 of empty_bar[b^1])
    TR(0);
    if (producer && cg + (int)gridDim.x < ncg) {
      if (it >= 1) mbar_wait_s(empty_s + 8u * (uint32_t)(b ^ 1), (uint32_t)((it - 1) >> 1) & 1u);
      produce(cg + gridDim.x, b ^ 1);
    }
    TR(1);
    const int grp = cg * wpb + wib;
    if (!LEAN && grp >= ngroups) continue;   // idle warp of the tail CTA-group (last iteration only)
    unsigned char* const in = inbuf0 + b * p.in_bytes;
    uint32_t* const sgrid = reinterpret_cast<uint32_t*>(in) + wib * EPW * env_words;
    const bool cta_full = LEAN ? true : (cg * EPC + EPC <= N);
    const int env0 = grp * EPW, nenv = LEAN ? EPW : min(EPW, N - env0);
    const bool valid = LEAN ? true : (tile < nenv);
    const int env = valid ? env0 + tile : env0;
    const bool active = LEAN ? true : (valid && tl < D);
    uint32_t* const sg = sgrid + tile * env_words;
    uint32_t* const gg = p.grid + (size_t)env * env_words;
    const bool pos_tma = pos_tma_c && cta_full, act_tma = act_tma_c && cta_full;
    float2* const pp = p.pos + (size_t)env * D + (active ? tl : 0);
    size_t aidx = (size_t)env * D + tl;                 // action index, advanced by N*D per step
    size_t ridx = (size_t)env;                          // reward / done index, advanced by N per step

    // inputs that did not ride the bulk copy (odd shapes): plain loads, issued before the wait
    float2 q = make_float2(0.f, 0.f);
    int a = 0;
    int4 m = make_int4(0, 0, 0, 0);
    if (active && !pos_tma) q = __ldcg(pp);
    if (active && !act_tma && (flags & F_MOVE)) a = FAST ? (int)__ldg((const uint8_t*)io.actions + aidx) : load_action(io.actions, io.act_dtype, (long long)aidx);

    TR(2);
#if SWARM_TRACE
    const long long tw0 = clock64();
#endif
    mbar_wait_s(full_s + 8u * (uint32_t)b, (phases >> b) & 1u);
    phases ^= 1u << b;
#if SWARM_TRACE
    if (threadIdx.x == 32) g_cta_t[2048 + blockIdx.x] += (unsigned long long)(clock64() - tw0);   // warp 1: cycles waiting for input
#endif
    TR(3);
    if (active && pos_tma) q = reinterpret_cast<const float2*>(in + p.off_pos)[(wib * EPW + tile) * D + tl];
    if (active && act_tma) a = (int)in[p.off_act + (wib * EPW + tile) * D + tl];
    if (valid && tl == 0 && meta_tma) m = reinterpret_cast<const int4*>(in + p.off_meta)[wib * EPW + tile];

    for (int t = 0; t < T; ++t, aidx += (size_t)N * D, ridx += N) {
      if (t > 0 && active && (flags & F_MOVE))   // rollout: positions / meta are carried in registers, actions are re-read
        a = FAST ? (int)__ldg((const uint8_t*)io.actions + aidx) : load_action(io.actions, io.act_dtype, (long long)aidx);

      // ---- stage (1): action decode + move + clip (noVis_sim.py:117-135)
      if (active && (flags & F_MOVE)) {
        float sx = (a == 1) ? p.ss : 0.f, sy = (a == 2) ? p.ss : 0.f;   // a=0: y-=ss, 1: x+=ss, 2: y+=ss, 3: x-=ss
        sx = (a == 3) ? -p.ss : sx;
        sy = (a == 0) ? -p.ss : sy;
        if ((unsigned)a > 3u) atomicOr(p.err, 1);
        q.x = fminf(fmaxf(q.x + sx, 0.f), p.xmax);
        q.y = fminf(fmaxf(q.y + sy, 0.f), p.ymax);
        __stcg(pp, q);
      }

      // ---- stage (2): coverage scatter on the shared-memory grid, conflicts resolved with a warp match
      int rew = 0, ncl = 0, nan = 0;
      if (flags & F_SCATTER) {
        if (cd_fast) {   // cd <= 0.5: the only candidate is the nearest lattice cell
          bool cand = false;
          int cell = 0;
          if (active) {
            const int cx = (int)floorf(q.x + 0.5f), cy = (int)floorf(q.y + 0.5f);
            const double dx = (double)((float)cx - q.x), dy = (double)((float)cy - q.y);   // both differences are exact
            cand = dx * dx + dy * dy < p.cd2_thr;     // sqrt(d2) < collision_dist (noVis_sim.py:81)
            cell = cy * p.R + cx;
          }
          // drones of one env on one cell form a match group; its lowest lane resolves the conflict for all of them.
          // (Measured: guarding the match with "does any lane stand on an unmapped cell", or electing the winner by the return
          // value of a shared-memory atomicAnd instead, both make the cfg3 step SLOWER — the warps then run in closer lock
          // step and a random ~10 % of the CTAs is starved of TMA input, DESIGN.md §5 — so the match is unconditional.)
          const unsigned key = cand ? (((unsigned)tile << 24) | (unsigned)cell) : (0x80000000u | (unsigned)lane);
          const unsigned mm = __match_any_sync(FULL, key);
          if (cand && lane == __ffs(mm) - 1) {
            const int sh = (cell & 15) * 2;
            const uint32_t code = (sg[cell >> 4] >> sh) & 3u;
            if (code) {
              bool clear = true;
              if (has_wait) {   // dwell (App. B.2): k drones in index order, each decrements or (at 0) maps
                const int k = __popc(mm);
                int16_t* w = p.wait + (size_t)env * p.RC + cell;
                const int wv = *w;
                if (wv >= k) { *w = (int16_t)(wv - k); clear = false; }
                else if (wv > 0) *w = 0;
              }
              if (clear) {
                atomicAnd(sg + (cell >> 4), ~(3u << sh));   // another leader of this tile may clear a cell of the same word
                atomicAnd(gg + (cell >> 4), ~(3u << sh));   // write-through to HBM (RED.AND, fire and forget)
                rew = 10 + (code == 2u ? 20 : 0);           // 10 + _check_uncertain_mapped (noVis_sim.py:83-84,104-110)
                ncl = 1;
                nan = (code == 2u);
              }
            }
          }
        } else {
          __syncwarp();
          if (valid && tl == 0) scatter_sequential(p, env, gg, sg, rew, ncl, nan);
        }
        int pk = rew | (ncl << 16) | (nan << 24);   // rew <= 30*32 < 2^16, ncl, nan <= 32
#pragma unroll
        for (int o = TS / 2; o > 0; o >>= 1) pk += __shfl_xor_sync(FULL, pk, o);
        rew = pk & 0xffff; ncl = (pk >> 16) & 0xff; nan = pk >> 24;
      }
      __syncwarp();   // the tile's clears are visible in shared memory

      // ---- stage (3): reward tail, spread, done; auto-reset
      if (flags & F_REWARD) {
        const int tcount = __shfl_sync(FULL, m.x, lead);
        const bool punish = tcount > p.punish_thr;                      // noVis_sim.py:141-143 (count before increment)
        const bool need = valid && !punish;
        bool spread = false;
        if (__any_sync(FULL, need)) spread = tile_spread<TS>(p, env, tile, tl, need, q.x, q.y);
        bool do_reset = false;
        int s_r = 0, s_cl = 0, s_an = 0, s_len = 0, s_ret = 0;   // this lane's contribution to the statistics
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
          s_r = r; s_cl = ncl; s_an = nan;
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
          acc_cells += __reduce_add_sync(FULL, s_cl);
          acc_anom += __reduce_add_sync(FULL, s_an);
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
            reset_tile_smem<TS>(p, env, episode, tl, do_reset, sg, q);
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
          emit_full_smem<OT>(p, cvt, sgrid, nenv, lane, obase + (size_t)env0 * p.RC);
        } else {
          TR(4);
          if (lane == 0) bulk_wait_read0();   // the previous bulk store has drained `stage` (long ago)
          __syncwarp();
          TR(7);
          if (active) build_window<WC, OT>(p, cvt, lut4, sg, q, my_stage);
          fence_proxy_async();   // generic-proxy writes to `stage` ordered before the async-proxy read below
          __syncwarp();
          const int n = nenv * D * WW;
          OT* dst = obase + (size_t)env0 * D * WW;
          const uint32_t bytes = (uint32_t)n * (uint32_t)sizeof(OT);
          if (LEAN || ((((uintptr_t)dst) & 15) == 0 && (bytes & 15u) == 0)) {
            if (lane == 0) {   // one TMA bulk store for the whole group
              bulk_s2g_s_hint(dst, stage_s, bytes, pol_ef);
              bulk_commit();
            }
          } else {
            for (int k = lane; k < n; k += 32) __stcs(dst + k, stage[k]);
          }
        }
      }
      __syncwarp();
    }
    fence_proxy_async();   // this group's generic accesses to the input buffer precede the bulk copy that refills it
    __syncwarp();
    TR(5);
    if (lane == 0) mbar_arrive_s(empty_s + 8u * (uint32_t)b);   // release the buffer to the producer
    TR(6);
  }
  if (lane == 0) bulk_wait_read0();   // shared memory must outlive the last bulk store's reads
#if SWARM_TRACE
  if (threadIdx.x == 0) g_cta_t[blockIdx.x * 2 + 1] = gtimer();
#endif

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
}

}  // namespace swarm
