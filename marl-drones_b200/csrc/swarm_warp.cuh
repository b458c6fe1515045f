// swarm_warp.cuh — the fused env step as a warp-autonomous, shared-memory-staged kernel (the hot path for D <= 32 and
// grids whose packed form fits a few KiB: every BASELINE config except the 256x256 stress case).
//
// One warp owns EPW = 32/TS consecutive envs (a "group"); a tile of TS lanes serves one env, one lane per drone.
// Per group the warp
//   (a) pulls the group's packed grids (EPW * env_words * 4 bytes, contiguous in HBM) into ITS OWN slice of shared memory
//       with one TMA bulk copy (cp.async.bulk, mbarrier completion) while positions / actions / meta loads are in flight,
//       so the whole step costs ONE DRAM round trip;
//   (b) runs stages (1)-(3) on the shared-memory copy (clears are written through to HBM with RED.AND);
//   (c) WINDOW obs: every lane gathers its drone's W x W window from shared memory into a staging slice (bank-conflict
//       free: lane stride W*W words), and one TMA bulk store streams the group's [EPW][D][W*W] block to the obs tensor;
//       FULL obs: the warp expands the 2-bit grid straight from shared memory with 128-bit streaming stores.
// There is no __syncthreads anywhere: warps drift through load / compute / store phases independently, which is what
// keeps HBM busy.  T > 1 (swarm_rollout) keeps the grid in shared memory across steps.
//
// Reference lines per stage: see swarm_step.cuh.
#pragma once
#include "swarm_common.cuh"
#include "swarm_emit.cuh"
#include "swarm_step.cuh"

#ifndef SWARM_MINB
#define SWARM_MINB 4   // resident CTAs per SM the staged kernel is compiled for (register cap 65536 / (256 * SWARM_MINB))
#endif

namespace swarm {

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
template <int WC, typename OT>
__device__ __forceinline__ void build_window(const DevParams& p, const ObsCvt<OT>& cvt, const uint32_t* sg, float2 q, OT* st) {
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
#pragma unroll
      for (int a = 0; a < W; ++a) st[b * W + a] = cvt.sel(field, a);
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
      const int e = (int)__umulhi((uint32_t)v, p.nq_magic);
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
// The grid is persistent (resident CTAs only); a warp strides over its groups, so the prologue and the statistics epilogue
// are paid once per warp, and the int64[8] statistics cost six RED.ADD per CTA.
template <int TS, typename OT, int MODE, int WC, bool FAST>
__global__ void __launch_bounds__(256, SWARM_MINB) step_warp_kernel(const __grid_constant__ DevParams p, const __grid_constant__ StepIO io) {
  constexpr int EPW = 32 / TS;
  extern __shared__ __align__(128) unsigned char smem_raw[];
  __shared__ uint64_t bars[8];
  __shared__ unsigned long long s_stats[6];
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5, wpb = blockDim.x >> 5;
  const int flags = FAST ? (F_ALL | (io.flags & F_RESET)) : io.flags;
  const bool cd_fast = FAST ? true : (p.cd_fast != 0);
  const bool has_wait = FAST ? false : (p.has_wait != 0);
  const bool want_stats = io.stats != nullptr && (flags & F_REWARD);
  if (want_stats) {
    if (threadIdx.x < 6) s_stats[threadIdx.x] = 0ull;
    __syncthreads();
  }
  unsigned char* wbase = smem_raw + (size_t)wib * p.warp_smem;
  OT* stage = reinterpret_cast<OT*>(wbase);                                   // [EPW][D][W*W] obs staging (WINDOW)
  uint32_t* sgrid = reinterpret_cast<uint32_t*>(wbase + p.stage_bytes);       // [EPW][env_words] packed grids
  uint64_t* bar = &bars[wib];
  if (lane == 0) {
    mbar_init(bar, 1);
    fence_mbar_init();
  }
  __syncwarp();
  uint32_t phase = 0;
  const int tile = lane / TS, tl = lane % TS, lead = tile * TS;
  const int D = p.D, N = p.N, env_words = p.env_words;
  const int ngroups = (N + EPW - 1) / EPW;
  const int gw = blockIdx.x * wpb + wib, nw = gridDim.x * wpb;
  const ObsCvt<OT> cvt(p);
  const bool want_obs = (flags & F_OBS) && io.obs_out != nullptr;
  const int WW = MODE == 1 ? (WC ? WC : p.W) * (WC ? WC : p.W) : 1;
  uint32_t* const sg = sgrid + tile * env_words;
  OT* const my_stage = stage + (tile * D + tl) * WW;
  StatAcc acc;

  for (int grp = gw; grp < ngroups; grp += nw) {
    const int env0 = grp * EPW, nenv = min(EPW, N - env0);
    const bool valid = tile < nenv;
    const int env = valid ? env0 + tile : env0;
    const bool active = valid && tl < D;
    uint32_t* const gg = p.grid + (size_t)env * env_words;

    // (a) stage the group's packed grids: one TMA bulk copy, completion on the warp's mbarrier
    if (lane == 0) {
      bulk_wait_read0();   // the previous group's obs bulk store no longer reads `stage`
      const uint32_t bytes = (uint32_t)nenv * (uint32_t)env_words * 4u;
      mbar_expect_tx(bar, bytes);
      bulk_g2s(sgrid, p.grid + (size_t)env0 * env_words, bytes, bar);
    }
    float2* const pp = p.pos + (size_t)env * D + (active ? tl : 0);
    size_t aidx = (size_t)env * D + tl;                 // action index, advanced by N*D per step
    size_t ridx = (size_t)env;                          // reward / done index, advanced by N per step

    for (int t = 0; t < io.T; ++t, aidx += (size_t)N * D, ridx += N) {
      // ---- loads for this step, all in flight together with the bulk copy
      float2 q = make_float2(0.f, 0.f);
      int a = 0;
      int4 m = make_int4(0, 0, 0, 0);
      if (active) {
        q = __ldcg(pp);
        if (flags & F_MOVE) a = FAST ? (int)__ldg((const uint8_t*)io.actions + aidx) : load_action(io.actions, io.act_dtype, (long long)aidx);
      }
      if (valid && tl == 0 && (flags & F_REWARD)) m = __ldcg(p.meta + env);
      if (t == 0) {
        mbar_wait(bar, phase);
        phase ^= 1u;
      }

      // ---- stage (1): action decode + move + clip (noVis_sim.py:117-135)
      if (active && (flags & F_MOVE)) {
        const float sx = (a == 1) ? p.ss : ((a == 3) ? -p.ss : 0.f);
        const float sy = (a == 2) ? p.ss : ((a == 0) ? -p.ss : 0.f);
        if ((unsigned)a > 3u) atomicOr(p.err, 1);
        q.x = fminf(fmaxf(q.x + sx, 0.f), p.xmax);
        q.y = fminf(fmaxf(q.y + sy, 0.f), p.ymax);
        __stcg(pp, q);
      }
      const int cx = (int)floorf(q.x + 0.5f), cy = (int)floorf(q.y + 0.5f);   // nearest lattice cell = window centre

      // ---- stage (2): coverage scatter on the shared-memory grid, conflicts resolved with a warp match
      int rew = 0, ncl = 0, nan = 0;
      if (flags & F_SCATTER) {
        if (cd_fast) {   // cd <= 0.5: the only candidate is the nearest lattice cell
          bool cand = false;
          const int cell = cy * p.R + cx;
          if (active) {
            const double dx = (double)((float)cx - q.x), dy = (double)((float)cy - q.y);   // exact differences
            cand = dx * dx + dy * dy < p.cd2_thr;     // sqrt(d2) < collision_dist (noVis_sim.py:81)
          }
          // drones of one env on one cell form a match group; its lowest lane resolves the conflict for all of them
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
          acc.rew += r; acc.cells += ncl; acc.anom += nan;
          if (done && m.y > 0) { acc.done += 1; acc.len += t1; acc.ret += epret; }
          do_reset = done && (flags & F_RESET);
          if (!do_reset) __stcg(p.meta + env, make_int4(t1, rem, epret, m.w));
        }
        if (flags & F_RESET) {
          do_reset = __shfl_sync(FULL, (int)do_reset, lead) != 0;
          if (__any_sync(FULL, do_reset)) {   // rare: an episode finished
            const int episode = __shfl_sync(FULL, m.w, lead);
            reset_tile_smem<TS>(p, env, episode + 1, tl, do_reset, sg, q);
          }
        }
      } else if (flags & F_SCATTER) {   // staged scatter (swarm_stage_scatter)
        if (valid && tl == 0) {
          io.cover_out[env] = rew;
          if (ncl) p.meta[env].y -= ncl;
        }
      }

      // ---- stage (4): observation emission
      if (want_obs && (!io.obs_last_only || t == io.T - 1)) {
        OT* obase = reinterpret_cast<OT*>(io.obs_out) + (io.obs_last_only ? (size_t)0 : (size_t)t * N * p.obs_per_env);
        if (MODE == 0) {
          emit_full_smem<OT>(p, cvt, sgrid, nenv, lane, obase + (size_t)env0 * p.RC);
        } else {
          if (t > 0) {   // the previous step's bulk store must have drained `stage`
            if (lane == 0) bulk_wait_read0();
            __syncwarp();
          }
          if (active) build_window<WC, OT>(p, cvt, sg, q, my_stage);
          fence_proxy_async();   // generic-proxy writes to `stage` ordered before the async-proxy read below
          __syncwarp();
          const int n = nenv * D * WW;
          OT* dst = obase + (size_t)env0 * D * WW;
          const uint32_t bytes = (uint32_t)n * (uint32_t)sizeof(OT);
          if ((((uintptr_t)dst) & 15) == 0 && (bytes & 15u) == 0) {
            if (lane == 0) {   // one TMA bulk store for the whole group
              bulk_s2g(dst, stage, bytes);
              bulk_commit();
            }
          } else {
            for (int k = lane; k < n; k += 32) __stcs(dst + k, stage[k]);
          }
        }
      }
      __syncwarp();
    }
    fence_proxy_async();   // this group's generic accesses to sgrid/stage precede the next group's bulk copy
    __syncwarp();
  }
  if (lane == 0) bulk_wait_read0();   // shared memory must outlive the last bulk store's reads

  if (want_stats) {   // only tile leaders accumulated; zero entries are skipped
    if (tl == 0) {
      const long long v[6] = {acc.rew, (long long)acc.done, acc.len, acc.ret, (long long)acc.cells, (long long)acc.anom};
#pragma unroll
      for (int k = 0; k < 6; ++k)
        if (v[k] != 0) atomicAdd(&s_stats[k], (unsigned long long)v[k]);
    }
    __syncthreads();
    if (threadIdx.x < 6 && s_stats[threadIdx.x] != 0ull) atomicAdd(io.stats + 2 + threadIdx.x, s_stats[threadIdx.x]);
    if (blockIdx.x == 0 && threadIdx.x == 0) {
      atomicAdd(io.stats + 0, (unsigned long long)((long long)N * io.T));
      atomicAdd(io.stats + 1, (unsigned long long)((long long)N * D * io.T));
    }
  }
}

}  // namespace swarm
