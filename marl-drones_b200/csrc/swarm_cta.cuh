// swarm_cta.cuh — the fused env step for shapes the shared-memory-staged warp kernel does not take: more than 32 drones per
// env, a coverage map too large to stage per warp (the 256x256 stress config), or PLANE-mode anomalies (more than 8
// anomalous cells, or a side longer than 256).  One CTA per env:
//   TMA bulk copies bring the env's coverage plane (8 KiB at 256x256; plus the anomaly plane in PLANE mode, or the 32-byte
//   list) into the CTA's shared memory,
//   stages (1)-(2) run on all threads, one drone per thread (production path), or on warp 0 in index-ordered chunks of 32
//   drones (generic path); cleared bits go to shared memory and, write-through, to HBM,
//   warp 0 runs stage (3), then the WHOLE CTA streams stage (4) from shared memory (a 256x256 float obs is 256 KiB per env:
//   the emission is the kernel).  Planes beyond the shared-memory budget fall back to the HBM-resident state.
// Reference lines per stage: see swarm_step.cuh.
#pragma once
#include "swarm_common.cuh"
#include "swarm_emit.cuh"
#include "swarm_step.cuh"
#include "swarm_window.cuh"

namespace swarm {

// The env's state as the CTA sees it: planes / list in shared memory when staged, else null (HBM accessors are used).
struct CtaView {
  uint32_t* sc;          // coverage plane (shared) or null
  const uint32_t* sa;    // anomaly plane (shared, PLANE mode) or null
  const uint32_t* sl;    // anomaly list (shared, LIST mode) or null
};

__device__ __forceinline__ bool cta_is_anomalous(const DevParams& p, const CtaView& v, int env, int cell, int x, int y) {
  if (v.sl) {
    const uint32_t key = (uint32_t)x | ((uint32_t)y << 16);
    bool hit = false;
    for (int k = 0; k < p.n_anom; ++k) hit = hit || (v.sl[k] == key);
    return hit;
  }
  if (v.sa) return (v.sa[cell >> 5] >> (cell & 31)) & 1u;
  return is_anomalous(p, env, cell, x, y);
}

// Stages (1)-(3) (+ auto-reset) of one env by one warp.
// `pre_pk` (may be null): stages (1)-(2) were already run by the whole CTA (move_drone / scatter_drone); *pre_pk holds the
// env's packed coverage totals (reward | cells << 16 | anomalous << 24) and this warp runs stage (3) only.
// Returns true when the env was reset (the caller refreshes its shared-memory view).
__device__ __forceinline__ bool step_env_warp(const DevParams& p, const StepIO& io, int t, int env, int lane, StatAcc& acc, const CtaView& v,
                                              const int* pre_pk = nullptr, float2* sp = nullptr, const int4* m_pre = nullptr) {
  constexpr int TS = 32;
  const int flags = pre_pk ? (io.flags & ~(F_MOVE | F_SCATTER)) : io.flags;
  const int D = p.D;
  const int nch = (D + TS - 1) / TS;
  uint32_t* g = p.cov + (size_t)env * p.env_words;
  int rew = 0, ncl = 0, nan = 0;
  float x = 0.f, y = 0.f;   // chunk-0 drone of this lane (all drones when D <= 32)

  if (pre_pk) {
    const int pk = *pre_pk;
    rew = pk & 0xffff; ncl = (pk >> 16) & 0xff; nan = (pk >> 24) & 0xff;
    if (lane < D) { const float2 q = sp ? sp[lane] : __ldcg(p.pos + (size_t)env * D + lane); x = q.x; y = q.y; }
  }
  for (int c = 0; c < (pre_pk ? 0 : nch); ++c) {
    const int d = c * TS + lane;
    const bool active = d < D;
    float2 q = make_float2(0.f, 0.f);
    if (active) {
      float2* pp = p.pos + (size_t)env * D + d;
      q = __ldcg(pp);
      if (flags & F_MOVE) {   // stage (1), noVis_sim.py:117-135
        const int a = load_action(io.actions, io.act_dtype, ((long long)t * p.N + env) * D + d);
        if (a == 0) q.y -= p.ss;
        else if (a == 1) q.x += p.ss;
        else if (a == 2) q.y += p.ss;
        else if (a == 3) q.x -= p.ss;
        else atomicOr(p.err, 1);
        q.x = fminf(fmaxf(q.x, 0.f), p.xmax);
        q.y = fminf(fmaxf(q.y, 0.f), p.ymax);
        __stcg(pp, q);
      }
    }
    if (c == 0) { x = q.x; y = q.y; }
    if ((flags & F_SCATTER) && p.cd_fast) {   // stage (2), cd <= 0.5: the only candidate is the nearest lattice cell
      bool cand = false;
      int cell = 0, cx = 0, cy = 0;
      if (active) {
        cx = (int)floorf(q.x + 0.5f); cy = (int)floorf(q.y + 0.5f);
        const double dx = (double)cx - (double)q.x, dy = (double)cy - (double)q.y;
        cand = dx * dx + dy * dy < p.cd2_thr;
        cell = cy * p.R + cx;
      }
      const unsigned key = cand ? (unsigned)cell : (0x80000000u | (unsigned)lane);
      const unsigned m = __match_any_sync(FULL, key);
      if (cand && lane == __ffs(m) - 1) {
        const int k = __popc(m);
        uint32_t* wp = g + (cell >> 5);
        const uint32_t bit = 1u << (cell & 31);
        if ((v.sc ? v.sc[cell >> 5] : ld_grid(wp)) & bit) {
          bool clear = true;
          if (p.has_wait) {   // dwell (App. B.2)
            int16_t* w = p.wait + (size_t)env * p.RC + cell;
            const int wv = *w;
            if (wv >= k) { *w = (int16_t)(wv - k); clear = false; }
            else if (wv > 0) *w = 0;
          }
          if (clear) {
            atomicAnd(wp, ~bit);
            if (v.sc) atomicAnd(v.sc + (cell >> 5), ~bit);
            const int an = cta_is_anomalous(p, v, env, cell, cx, cy) ? 1 : 0;
            rew += 10 + 20 * an;   // 10 + _check_uncertain_mapped (noVis_sim.py:83-84,104-110)
            ncl += 1;
            nan += an;
          }
        }
      }
      if (nch > 1) __syncwarp();   // later chunks see this chunk's clears (index order)
    }
  }
  if ((flags & F_SCATTER) && !p.cd_fast) {
    __syncwarp();
    if (lane == 0) scatter_sequential(p, env, v.sc, rew, ncl, nan);
  }
#pragma unroll
  for (int o = 16; o > 0 && !pre_pk; o >>= 1) {
    rew += __shfl_xor_sync(FULL, rew, o);
    ncl += __shfl_xor_sync(FULL, ncl, o);
    nan += __shfl_xor_sync(FULL, nan, o);
  }

  bool did_reset = false;
  if (flags & F_REWARD) {   // stage (3)
    int4 m = make_int4(0, 0, 0, 0);
    if (lane == 0) m = m_pre ? *m_pre : __ldcg(p.meta + env);
    const int tcount = __shfl_sync(FULL, m.x, 0);
    const bool punish = tcount > p.punish_thr;                    // noVis_sim.py:141-143 (count before increment)
    const bool need = !punish;
    bool spread = false;
    __syncwarp();   // every chunk's position stores precede the D > 32 re-reads in tile_spread
    if (need) spread = tile_spread<TS>(p, env, 0, lane, need, x, y, sp);
    bool do_reset = false;
    if (lane == 0) {
      int r = ((flags & F_SCATTER) || pre_pk) ? rew : io.cover_in[env];
      if (p.reward_mode == 0) r += ((!punish && spread) ? 1 : 0) - 1;   // A2C noVis_sim.py:150-152
      else r += punish ? -1 : (spread ? 1 : 0);                         // DQL noVis_sim.py:141-148
      const int rem = m.y - ncl;
      const bool done = rem == 0;                                       // noVis_sim.py:157-166
      if (p.reward_mode == 0 && done) r += 100;                         // noVis_sim.py:172-173
      const int t1 = tcount + 1, epret = m.z + r;
      io.reward_out[(size_t)t * p.N + env] = (float)r;
      io.done_out[(size_t)t * p.N + env] = done ? 1 : 0;
      acc.rew += r; acc.cells += ncl; acc.anom += nan;
      if (done && m.y > 0) { acc.done += 1; acc.len += t1; acc.ret += epret; }
      do_reset = done && (flags & F_RESET);
      if (!do_reset) __stcg(p.meta + env, make_int4(t1, rem, epret, m.w));
    }
    do_reset = __shfl_sync(FULL, (int)do_reset, 0) != 0;
    const int episode = __shfl_sync(FULL, m.w, 0);
    if (do_reset) {
      reset_tile<TS>(p, env, episode + 1, lane, true);
      did_reset = true;
    }
  } else if (flags & F_SCATTER) {   // staged scatter (swarm_stage_scatter)
    if (lane == 0) {
      io.cover_out[env] = rew;
      if (ncl) p.meta[env].y -= ncl;
    }
  }
  __syncwarp();
  return did_reset;
}

// Stages (1)-(2) by the WHOLE CTA for collision_dist <= 0.5 without dwell (the production path): one thread per drone moves,
// clips and tests its nearest lattice cell on the shared-memory plane.  Several drones on one cell: the return value of the
// shared-memory atomicAnd elects exactly one of them (the one that still saw the bit), so the cell is cleared and rewarded
// once — the per-env reward is a sum, the order does not matter.  Totals are packed into *s_pk (at most 255
// drones fit the 8-bit fields — larger D fall back to the warp path).
__device__ __forceinline__ float2 move_drone(const DevParams& p, const StepIO& io, int t, int env, int d) {   // stage (1)
  float2* pp = p.pos + (size_t)env * p.D + d;
  float2 q = __ldcg(pp);
  const int a = load_action(io.actions, io.act_dtype, ((long long)t * p.N + env) * p.D + d);
  if (a == 0) q.y -= p.ss;                      // noVis_sim.py:121-128
  else if (a == 1) q.x += p.ss;
  else if (a == 2) q.y += p.ss;
  else if (a == 3) q.x -= p.ss;
  else atomicOr(p.err, 1);
  q.x = fminf(fmaxf(q.x, 0.f), p.xmax);         // noVis_sim.py:130-135
  q.y = fminf(fmaxf(q.y, 0.f), p.ymax);
  __stcg(pp, q);
  return q;
}
__device__ __forceinline__ void scatter_drone(const DevParams& p, const CtaView& v, int env, float2 q, int* s_pk) {   // stage (2)
  const int cx = (int)floorf(q.x + 0.5f), cy = (int)floorf(q.y + 0.5f);
  const double dx = (double)cx - (double)q.x, dy = (double)cy - (double)q.y;
  if (dx * dx + dy * dy < p.cd2_thr) {          // sqrt(d2) < collision_dist (noVis_sim.py:81)
    const int cell = cy * p.R + cx;
    const uint32_t bit = 1u << (cell & 31);
    if (v.sc[cell >> 5] & bit) {
      if (atomicAnd(v.sc + (cell >> 5), ~bit) & bit) {
        atomicAnd(p.cov + (size_t)env * p.env_words + (cell >> 5), ~bit);   // write-through to HBM
        const int an = cta_is_anomalous(p, v, env, cell, cx, cy) ? 1 : 0;
        atomicAdd(s_pk, (10 + 20 * an) | (1 << 16) | (an << 24));
      }
    }
  }
}

// Stage (2) with dwell goals (App. B.2) on the all-thread path: `s_cell[j]` holds drone j's candidate cell of this tick (-1: none).
// The drones on one cell form a group; its lowest-indexed drone acts for all k of them, which is what processing them in index
// order amounts to: the counter absorbs k ticks, or runs out — then the next drone in line maps the cell and the rest find it mapped.
__device__ __forceinline__ int drone_candidate_cell(const DevParams& p, float2 q) {
  const int cx = (int)floorf(q.x + 0.5f), cy = (int)floorf(q.y + 0.5f);
  const double dx = (double)cx - (double)q.x, dy = (double)cy - (double)q.y;
  return (dx * dx + dy * dy < p.cd2_thr) ? cy * p.R + cx : -1;
}
__device__ __forceinline__ void scatter_drone_dwell(const DevParams& p, const CtaView& v, int env, int d, const int* s_cell, int* s_pk) {
  const int cell = s_cell[d];
  if (cell < 0) return;
  const uint32_t bit = 1u << (cell & 31);
  if (!(v.sc[cell >> 5] & bit)) return;   // mapped before this tick (bits of OTHER cells of the word may be cleared meanwhile)
  int k = 0;
  bool leader = true;
  for (int j = 0; j < p.D; ++j) {
    const bool same = s_cell[j] == cell;
    k += same ? 1 : 0;
    leader = leader && !(same && j < d);
  }
  if (!leader) return;
  int16_t* w = p.wait + (size_t)env * p.RC + cell;
  const int wv = *w;
  if (wv >= k) { *w = (int16_t)(wv - k); return; }
  if (wv > 0) *w = 0;
  atomicAnd(v.sc + (cell >> 5), ~bit);
  atomicAnd(p.cov + (size_t)env * p.env_words + (cell >> 5), ~bit);   // write-through to HBM
  const int cy = cell / p.R, cx = cell - cy * p.R;
  const int an = cta_is_anomalous(p, v, env, cell, cx, cy) ? 1 : 0;
  atomicAdd(s_pk, (10 + 20 * an) | (1 << 16) | (an << 24));
}

// CTA-wide stage (4) from the shared-memory view.  FULL: bits -> elements with 128-bit stores (+ the LIST patch after a
// CTA barrier); WINDOW: one item = one window row of one drone.
template <typename OT>
__device__ __forceinline__ void emit_obs_cta_smem(const DevParams& p, const CtaView& v, int env, int tid, int nthr, OT* __restrict__ obs,
                                                  const float2* sp = nullptr) {
  const ObsCvt<OT> cvt(p);
  if (p.obs_mode == 0) {
    OT* o = obs + (size_t)env * p.RC;
    emit_full_bits<OT>(p, cvt, v.sc, v.sa, 1, tid, nthr, o);
    if (v.sl) {
      __syncthreads();   // the patch overwrites cells other threads have just written
      emit_full_patch<OT>(p, cvt, v.sc, v.sl, 1, tid, nthr, o);
    }
  } else {
    const int W = p.W, WW = W * W, h = (W - 1) >> 1, nitem = p.D * W;
    const float2* ps = p.pos + (size_t)env * p.D;
    OT* o = obs + (size_t)env * p.D * WW;
    for (int it = tid; it < nitem; it += nthr) {
      const int i = it / W, b = it - i * W;
      const float2 q = sp ? sp[i] : __ldcg(ps + i);
      const int cx = (int)floorf(q.x + 0.5f), gy = (int)floorf(q.y + 0.5f) - h + b;
      OT* dst = o + (size_t)i * WW + b * W;
      const bool row_ok = (unsigned)gy < (unsigned)p.C;
      for (int a = 0; a < W; ++a) {
        const int gx = cx - h + a;
        uint32_t code = CODE_PAD;
        if (row_ok && (unsigned)gx < (unsigned)p.R) {
          const int cell = gy * p.R + gx;
          code = (v.sc[cell >> 5] >> (cell & 31)) & 1u;
          if (code && cta_is_anomalous(p, v, env, cell, gx, gy)) code = CODE_ANOM;
        }
        dst[a] = cvt.one(code);
      }
    }
  }
}

// `staged` != 0: dynamic shared memory holds the env's plane(s) (+ slack) and the state is staged through it.
template <typename OT>
__global__ void __launch_bounds__(256) step_cta_kernel(const __grid_constant__ DevParams p, const __grid_constant__ StepIO io, int staged) {
  extern __shared__ __align__(128) unsigned char cta_smem[];
  __shared__ uint64_t bar;
  __shared__ int s_pk, s_reset;
  __shared__ __align__(16) uint32_t s_list[8];
  __shared__ float2 s_pos[256];   // this step's positions of the env (wide path): spread test and window emission read them here
  __shared__ int s_cell[256];     // dwell goals on the wide path: each drone's candidate cell of this tick (-1: none)
  // W == 5 window fast path (LIST mode, wide path): the tables of swarm_window.cuh, and a staging image of the env's
  // [D][25] obs block behind the plane in dynamic shared memory, streamed out with ONE TMA bulk store per env
  __shared__ __align__(16) uint32_t lut[256 * (sizeof(OT) == 4 ? 4 : 1)];
  __shared__ __align__(16) uint32_t padx[256], pady[256];
  CtaView v;
  v.sc = staged ? reinterpret_cast<uint32_t*>(cta_smem) : nullptr;
  v.sa = (staged && !p.list_mode) ? reinterpret_cast<const uint32_t*>(cta_smem) + p.env_words : nullptr;
  v.sl = (staged && p.list_mode) ? s_list : nullptr;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int flags = io.flags;
  const bool want_obs = (flags & F_OBS) && io.obs_out != nullptr;
  const bool stepping = (flags & (F_MOVE | F_SCATTER | F_REWARD)) != 0;
  if (staged && threadIdx.x == 0) {
    mbar_init(&bar, 1);
    fence_mbar_init();
  }
  pdl_wait();
  if (p.peers.n_ranks > 0 && (io.flags & F_PUSH) && io.stats && blockIdx.x == 0 && threadIdx.x == 0)
    peer_stats_push(p.peers, io.stats);   // multi-GPU: publish the totals of the launches before this one (see PeerStats)
  __syncthreads();
  uint32_t phase = 0;
  StatAcc acc;
  const uint32_t plane_bytes = (uint32_t)p.env_words * 4u;
  constexpr uint32_t lut_bytes = 256u * (sizeof(OT) == 4 ? 16u : 4u);
  auto load_view = [&](int env, bool with_tables = false) {   // one thread: the env's plane(s) / list (+ the window tables) by TMA
    mbar_expect_tx(&bar, (p.list_mode ? plane_bytes + 32u : 2u * plane_bytes) + (with_tables ? lut_bytes + 2048u : 0u));
    if (with_tables) {
      bulk_g2s(lut, p.tables, lut_bytes, &bar);
      bulk_g2s(padx, p.tables + 1024, 1024u, &bar);
      bulk_g2s(pady, p.tables + 1280, 1024u, &bar);
    }
    bulk_g2s(v.sc, p.cov + (size_t)env * p.env_words, plane_bytes, &bar);
    if (p.list_mode) bulk_g2s(s_list, p.alist + (size_t)env * 8, 32u, &bar);
    else bulk_g2s(const_cast<uint32_t*>(v.sa), p.anomp + (size_t)env * p.env_words, plane_bytes, &bar);
  };
  // production path: stages (1)-(2) on all threads, one drone per thread (30 * 255 < 2^16 keeps the packed totals exact)
  const bool wide = staged && (flags & (F_MOVE | F_SCATTER | F_REWARD)) == (F_MOVE | F_SCATTER | F_REWARD) && p.cd_fast &&
                    p.D <= 255 && p.D <= (int)blockDim.x;
  const uint32_t win_bytes = (uint32_t)p.D * 25u * (uint32_t)sizeof(OT);
  const bool win5 = wide && want_obs && p.list_mode && p.obs_mode == 1 && p.W == 5 && (win_bytes & 15u) == 0u && staged == 2;
  OT* const s_win = reinterpret_cast<OT*>(cta_smem + (((size_t)p.env_words * 4 + 15) & ~(size_t)15));   // staged == 2: room behind the plane

  for (int env = blockIdx.x; env < p.N; env += gridDim.x) {
    int4 m_pre = make_int4(0, 0, 0, 0);
    if (staged && threadIdx.x == 0) {
      load_view(env, win5 && env == (int)blockIdx.x);   // the CTA's first env brings the window tables along
      if (wide) m_pre = __ldcg(p.meta + env);   // in flight together with the planes and the positions
    }
    bool landed = !staged;
    for (int t = 0; t < io.T; ++t) {
      if (threadIdx.x == 0) s_reset = 0;
      if (stepping && wide) {
        // stage (1) needs no grid: its loads overlap the bulk copy
        const bool mine = (int)threadIdx.x < p.D;
        float2 q = make_float2(0.f, 0.f);
        if (threadIdx.x == 0) s_pk = 0;
        if (mine) {
          q = move_drone(p, io, t, env, threadIdx.x);
          s_pos[threadIdx.x] = q;
          if (p.has_wait) s_cell[threadIdx.x] = drone_candidate_cell(p, q);
        }
        if (!landed) { mbar_wait(&bar, phase); phase ^= 1u; landed = true; }
        __syncthreads();   // s_pk is zero, the planes have landed
        if (mine) {
          if (p.has_wait) scatter_drone_dwell(p, v, env, threadIdx.x, s_cell, &s_pk);
          else scatter_drone(p, v, env, q, &s_pk);
        }
        __syncthreads();   // clears (shared memory), positions (s_pos) and totals are visible to warp 0
        if (warp == 0 && step_env_warp(p, io, t, env, lane, acc, v, &s_pk, s_pos, t == 0 ? &m_pre : nullptr) && lane == 0) s_reset = 1;
        __syncthreads();   // warp 0's updates (auto-reset) are visible to the CTA
      } else {
        if (!landed) { mbar_wait(&bar, phase); phase ^= 1u; landed = true; }
        __syncthreads();
        if (stepping) {
          if (warp == 0 && step_env_warp(p, io, t, env, lane, acc, v) && lane == 0) s_reset = 1;
          __syncthreads();   // warp 0's updates are visible to the CTA
        }
      }
      if (staged && s_reset) {   // rare: the env was reset in HBM — bring the fresh planes / list / positions on chip
        fence_proxy_async();
        __syncthreads();
        if (threadIdx.x == 0) load_view(env);
        if (stepping && wide && (int)threadIdx.x < p.D) s_pos[threadIdx.x] = __ldcg(p.pos + (size_t)env * p.D + threadIdx.x);
        mbar_wait(&bar, phase); phase ^= 1u;
        __syncthreads();
      }
      if (want_obs && (!io.obs_last_only || t == io.T - 1)) {
        OT* o = reinterpret_cast<OT*>(io.obs_out) + (io.obs_last_only ? (size_t)0 : (size_t)t * p.N * p.obs_per_env);
        if (win5) {   // one thread per drone builds its window into the staging image; one bulk store per env
          if (threadIdx.x == 0) bulk_wait_read0();   // the previous env's / step's bulk store has drained the image
          __syncthreads();
          if ((int)threadIdx.x < p.D) {
            const float2 q = s_pos[threadIdx.x];
            const int cx = (int)floorf(q.x + 0.5f), cy = (int)floorf(q.y + 0.5f);
            const uint32_t A = window_anomalies5(smem_u32(s_list), p.n_anom, cx, cy);
            build_window5<OT>(p.R, smem_u32(v.sc), smem_u32(lut), padx[cx] | pady[cy], A, cx - 2, cy - 2, s_win + (size_t)threadIdx.x * 25);
          }
          fence_proxy_async();   // generic-proxy writes to the image ordered before the async-proxy read below
          __syncthreads();
          OT* dst = o + (size_t)env * p.D * 25;
          if (threadIdx.x == 0) {
            if ((((uintptr_t)dst) & 15) == 0) {
              bulk_s2g(dst, s_win, win_bytes);
              bulk_commit();
            } else {
              for (uint32_t k = 0; k < (uint32_t)p.D * 25u; ++k) dst[k] = s_win[k];
            }
          }
        } else if (staged) emit_obs_cta_smem<OT>(p, v, env, threadIdx.x, blockDim.x, o, (stepping && wide) ? s_pos : nullptr);
        else emit_obs_global<OT>(p, env, threadIdx.x, blockDim.x, o);
      }
      if (t + 1 < io.T) __syncthreads();   // emission reads precede the next step's clears; s_reset is re-armed
    }
    if (staged) {
      fence_proxy_async();   // generic accesses to the staged view precede the next env's bulk copies
      __syncthreads();
    }
  }

  pdl_launch_dependents();   // main loop done: the next launch on the stream may bring its CTAs in (they run their prologue, then pdl_wait)
  if (win5 && threadIdx.x == 0) bulk_wait_read0();   // shared memory must outlive the last bulk store's reads
  if (io.stats && (flags & F_REWARD)) {   // only lane 0 of warp 0 accumulated (step_env_warp): no reduction needed
    if (threadIdx.x == 0) {
      const long long vv[6] = {acc.rew, (long long)acc.done, acc.len, acc.ret, (long long)acc.cells, (long long)acc.anom};
#pragma unroll
      for (int k = 0; k < 6; ++k)
        if (vv[k] != 0) atomicAdd(io.stats + 2 + k, (unsigned long long)vv[k]);
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) {
      atomicAdd(io.stats + 0, (unsigned long long)((long long)p.N * io.T));
      atomicAdd(io.stats + 1, (unsigned long long)((long long)p.N * p.D * io.T));
    }
  }
}

}  // namespace swarm
