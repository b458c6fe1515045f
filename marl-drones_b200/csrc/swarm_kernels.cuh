// swarm_kernels.cuh — the fused step / rollout kernel and the state-management kernels.
#pragma once
#include "swarm_common.cuh"
#include "swarm_emit.cuh"
#include "swarm_step.cuh"

namespace swarm {

constexpr int F_RESET = 16;   // auto-reset on done inside the step (A2C/envs.py:22-26)

struct StepIO {
  const void* actions;     // [T][N][D] of act_dtype
  int act_dtype;
  void* obs_out;           // [T][N][obs_per_env] (or one step when obs_last_only), may be null
  int obs_last_only;
  float* reward_out;       // [T][N]
  uint8_t* done_out;       // [T][N]
  const int* cover_in;     // staged reward: coverage reward per env
  int* cover_out;          // staged scatter
  unsigned long long* stats;   // int64[8] accumulated, may be null
  int T;
  int flags;
};

// One env step for the EPW envs of this warp.
template <int TS, typename OT>
__device__ __forceinline__ void step_group(const DevParams& p, const StepIO& io, int t, int env0, int lane, StatAcc& acc) {
  constexpr int EPW = 32 / TS;
  const int flags = io.flags;
  const int tile = lane / TS, tl = lane % TS;
  const int env = env0 + tile;
  const bool valid = env < p.N;
  const int lead = tile * TS;
  const int D = p.D;
  const int nch = (D + TS - 1) / TS;
  uint32_t* g = p.grid + (size_t)(valid ? env : 0) * p.env_words;
  int rew = 0, ncl = 0, nan = 0;
  float x = 0.f, y = 0.f;   // chunk-0 drone of this lane (all drones when D <= TS)

  if (flags & (F_MOVE | F_SCATTER | F_REWARD)) {
    for (int c = 0; c < nch; ++c) {
      const int d = c * TS + tl;
      const bool active = valid && d < D;
      float2 q = make_float2(0.f, 0.f);
      if (active) {
        float2* pp = p.pos + (size_t)env * D + d;
        q = *pp;
        if (flags & F_MOVE) {   // stage (1), noVis_sim.py:117-135
          const int a = load_action(io.actions, io.act_dtype, ((long long)t * p.N + env) * D + d);
          if (a == 0) q.y -= p.ss;
          else if (a == 1) q.x += p.ss;
          else if (a == 2) q.y += p.ss;
          else if (a == 3) q.x -= p.ss;
          else atomicOr(p.err, 1);
          q.x = fminf(fmaxf(q.x, 0.f), p.xmax);
          q.y = fminf(fmaxf(q.y, 0.f), p.ymax);
          *pp = q;
        }
      }
      if (c == 0) { x = q.x; y = q.y; }
      if ((flags & F_SCATTER) && p.cd_fast) {   // stage (2), cd <= 0.5: the only candidate is the nearest lattice cell
        bool cand = false;
        int cell = 0;
        if (active) {
          const int cx = (int)floorf(q.x + 0.5f), cy = (int)floorf(q.y + 0.5f);
          const double dx = (double)cx - (double)q.x, dy = (double)cy - (double)q.y;
          cand = sqrt(dx * dx + dy * dy) < p.cd;
          cell = cy * p.R + cx;
        }
        // drones of one env on one cell form a match group; its lowest lane resolves the conflict for all of them
        const unsigned long long key = cand ? (((unsigned long long)tile << 32) | (unsigned)cell) : ((1ull << 63) | (unsigned)lane);
        const unsigned m = __match_any_sync(FULL, key);
        if (cand && lane == __ffs(m) - 1) {
          const int k = __popc(m);
          uint32_t* wp = g + (cell >> 4);
          const int sh = (cell & 15) * 2;
          const uint32_t code = (ld_grid(wp) >> sh) & 3u;
          if (code) {
            bool clear = true;
            if (p.has_wait) {   // dwell (App. B.2): k drones in index order, each decrements or (at 0) maps
              int16_t* w = p.wait + (size_t)env * p.RC + cell;
              const int wv = *w;
              if (wv >= k) { *w = (int16_t)(wv - k); clear = false; }
              else if (wv > 0) *w = 0;
            }
            if (clear) {
              atomicAnd(wp, ~(3u << sh));
              rew += 10 + (code == 2u ? 20 : 0);   // 10 + _check_uncertain_mapped (noVis_sim.py:83-84,104-110)
              ncl += 1;
              nan += (code == 2u);
            }
          }
        }
        if (nch > 1) __syncwarp();
      }
    }
    if ((flags & F_SCATTER) && !p.cd_fast) {
      __syncwarp();
      if (valid && tl == 0) scatter_sequential(p, env, g, rew, ncl, nan);
    }
  }

  // tile reduction of the coverage reward / newly mapped counts
#pragma unroll
  for (int o = TS / 2; o > 0; o >>= 1) {
    rew += __shfl_xor_sync(FULL, rew, o);
    ncl += __shfl_xor_sync(FULL, ncl, o);
    nan += __shfl_xor_sync(FULL, nan, o);
  }

  if (flags & F_REWARD) {   // stage (3)
    int4 m = make_int4(0, 0, 0, 0);
    if (valid && tl == 0) m = p.meta[env];
    const int tcount = __shfl_sync(FULL, m.x, lead);
    const bool punish = tcount > p.punish_thr;                    // noVis_sim.py:141-143 (count before increment)
    const bool need = valid && !punish;
    bool spread = false;
    if (__any_sync(FULL, need)) spread = tile_spread<TS>(p, valid ? env : 0, tile, tl, need, x, y);
    bool do_reset = false;
    if (valid && tl == 0) {
      int r = (flags & F_SCATTER) ? rew : io.cover_in[env];
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
      if (!do_reset) p.meta[env] = make_int4(t1, rem, epret, m.w);
    }
    do_reset = __shfl_sync(FULL, (int)do_reset, lead) != 0;
    const int episode = __shfl_sync(FULL, m.w, lead);
    if (__any_sync(FULL, do_reset)) reset_tile<TS>(p, env, episode + 1, tl, do_reset);
  } else if (flags & F_SCATTER) {   // staged scatter (swarm_stage_scatter)
    if (valid && tl == 0) {
      io.cover_out[env] = rew;
      if (ncl) p.meta[env].y -= ncl;
    }
  }

  if ((flags & F_OBS) && io.obs_out && (!io.obs_last_only || t == io.T - 1)) {   // stage (4)
    __syncwarp();
    const int nenv = min(EPW, p.N - env0);
    OT* o = reinterpret_cast<OT*>(io.obs_out) + (io.obs_last_only ? 0 : (size_t)t * p.N * p.obs_per_env);
    emit_obs<OT>(p, env0, nenv, lane, o);
  }
  __syncwarp();
}

// Persistent warp-per-env-group kernel: T steps of every env group in one launch (T = 1 is swarm_step).
// Envs are independent, so a warp runs its group through all T steps without any grid-wide synchronisation.
template <int TS, typename OT>
__global__ void __launch_bounds__(256) step_kernel(const __grid_constant__ DevParams p, const __grid_constant__ StepIO io) {
  constexpr int EPW = 32 / TS;
  const int lane = threadIdx.x & 31;
  const int wpb = blockDim.x >> 5;
  const int gw = blockIdx.x * wpb + (threadIdx.x >> 5);
  const int nw = gridDim.x * wpb;
  const int ngroups = (p.N + EPW - 1) / EPW;
  StatAcc acc;
  for (int grp = gw; grp < ngroups; grp += nw)
    for (int t = 0; t < io.T; ++t) step_group<TS, OT>(p, io, t, grp * EPW, lane, acc);

  if (io.stats && (io.flags & F_REWARD)) {
    __shared__ unsigned long long s[6];
    if (threadIdx.x < 6) s[threadIdx.x] = 0ull;
    __syncthreads();
    long long v[6] = {acc.rew, (long long)acc.done, acc.len, acc.ret, (long long)acc.cells, (long long)acc.anom};
#pragma unroll
    for (int k = 0; k < 6; ++k) {
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) v[k] += __shfl_xor_sync(FULL, v[k], o);
      if (lane == 0 && v[k] != 0) atomicAdd(&s[k], (unsigned long long)v[k]);
    }
    __syncthreads();
    if (threadIdx.x < 6 && s[threadIdx.x] != 0ull) atomicAdd(io.stats + 2 + threadIdx.x, s[threadIdx.x]);
    if (blockIdx.x == 0 && threadIdx.x == 0) {
      atomicAdd(io.stats + 0, (unsigned long long)((long long)p.N * io.T));
      atomicAdd(io.stats + 1, (unsigned long long)((long long)p.N * p.D * io.T));
    }
  }
}

// ------------------------------------------------------------------------------------------------ state kernels
// mode 0: create (episode 0, every env); mode 1: reset masked envs to their next episode.
__global__ void __launch_bounds__(256) reset_kernel(const __grid_constant__ DevParams p, const uint8_t* mask, int mode) {
  const int lane = threadIdx.x & 31;
  const int gw = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int nw = gridDim.x * (blockDim.x >> 5);
  for (int env = gw; env < p.N; env += nw) {
    const bool doit = mode == 0 || mask == nullptr || mask[env] != 0;
    int episode = 0;
    if (mode == 1) episode = p.meta[env].w + 1;
    __syncwarp();
    reset_tile<32>(p, env, episode, lane, doit);
  }
}

// set_state: boundary int8 [count][RC] (values 0 / 1 / U, obs order) -> packed grid; recomputes `remaining`.
__global__ void __launch_bounds__(256) set_state_kernel(const __grid_constant__ DevParams p, const int8_t* grid, const float* pos,
                                                        const int* t, const int16_t* wait, int first, int count) {
  const int lane = threadIdx.x & 31;
  const int gw = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int nw = gridDim.x * (blockDim.x >> 5);
  for (int e = gw; e < count; e += nw) {
    const int env = first + e;
    int rem = 0;
    if (grid) {
      const int8_t* src = grid + (size_t)e * p.RC;
      for (int w = lane; w < p.env_words; w += 32) {
        uint32_t word = 0;
        for (int k = 0; k < 16; ++k) {
          const int c = w * 16 + k;
          if (c < p.RC) {
            const int v = src[c];
            const uint32_t code = v <= 0 ? 0u : (v == p.U ? 2u : 1u);
            word |= code << (2 * k);
            rem += code != 0u;
          }
        }
        st_grid(p.grid + (size_t)env * p.env_words + w, word);
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) rem += __shfl_xor_sync(FULL, rem, o);
    }
    if (pos)
      for (int k = lane; k < p.D; k += 32)
        p.pos[(size_t)env * p.D + k] = make_float2(pos[((size_t)e * p.D + k) * 2], pos[((size_t)e * p.D + k) * 2 + 1]);
    if (wait && p.has_wait)
      for (int k = lane; k < p.RC; k += 32) p.wait[(size_t)env * p.RC + k] = wait[(size_t)e * p.RC + k];
    if (lane == 0) {
      int4 m = p.meta[env];
      if (t) m.x = t[e];
      if (grid) m.y = rem;
      m.z = 0;
      p.meta[env] = m;
    }
  }
}

__global__ void __launch_bounds__(256) get_state_kernel(const __grid_constant__ DevParams p, int8_t* grid, float* pos, int* t,
                                                        int16_t* wait, int first, int count) {
  const int lane = threadIdx.x & 31;
  const int gw = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int nw = gridDim.x * (blockDim.x >> 5);
  for (int e = gw; e < count; e += nw) {
    const int env = first + e;
    if (grid)
      for (int c = lane; c < p.RC; c += 32) {
        const uint32_t code = (ld_grid(p.grid + (size_t)env * p.env_words + (c >> 4)) >> ((c & 15) * 2)) & 3u;
        grid[(size_t)e * p.RC + c] = (int8_t)(code == 2u ? p.U : (int)code);
      }
    if (pos)
      for (int k = lane; k < p.D; k += 32) {
        const float2 q = p.pos[(size_t)env * p.D + k];
        pos[((size_t)e * p.D + k) * 2] = q.x;
        pos[((size_t)e * p.D + k) * 2 + 1] = q.y;
      }
    if (wait)
      for (int k = lane; k < p.RC; k += 32) wait[(size_t)e * p.RC + k] = p.has_wait ? p.wait[(size_t)env * p.RC + k] : (int16_t)0;
    if (t && lane == 0) t[e] = p.meta[env].x;
  }
}

// K1 as a standalone stage: one thread per (env, drone).
__global__ void __launch_bounds__(256) move_kernel(const __grid_constant__ DevParams p, const void* actions, int act_dtype) {
  const long long n = (long long)p.N * p.D;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    float2 q = p.pos[i];
    const int a = load_action(actions, act_dtype, i);
    if (a == 0) q.y -= p.ss;
    else if (a == 1) q.x += p.ss;
    else if (a == 2) q.y += p.ss;
    else if (a == 3) q.x -= p.ss;
    else atomicOr(p.err, 1);
    q.x = fminf(fmaxf(q.x, 0.f), p.xmax);
    q.y = fminf(fmaxf(q.y, 0.f), p.ymax);
    p.pos[i] = q;
  }
}

// Scripted greedy controller on the noVis lattice (restates swarmsim-2d/drone.py:29-58; see oracle_greedy_actions).
__global__ void __launch_bounds__(256) greedy_kernel(const __grid_constant__ DevParams p, uint8_t* actions) {
  const long long n = (long long)p.N * p.D;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const int env = (int)(i / p.D);
    const uint32_t* g = p.grid + (size_t)env * p.env_words;
    const float2 q = p.pos[i];
    const double x = q.x, y = q.y;
    double best = -1.0;
    int bx = -1, by = -1;
    for (int gx = 0; gx < p.R; ++gx)
      for (int gy = 0; gy < p.C; ++gy) {
        const int cell = gy * p.R + gx;
        if ((ld_grid(g + (cell >> 4)) >> ((cell & 15) * 2)) & 3u) {
          const double d2 = (gx - x) * (gx - x) + (gy - y) * (gy - y);
          if (best < 0.0 || d2 < best) { best = d2; bx = gx; by = gy; }
        }
      }
    uint8_t a = 0;
    if (bx >= 0) {
      const double dx = bx - x, dy = by - y;
      if (dx == 0.0 && dy == 0.0) a = 0;
      else if (fabs(dx) >= fabs(dy)) a = dx > 0 ? 1 : 3;
      else a = dy > 0 ? 2 : 0;
    }
    actions[i] = a;
  }
}

}  // namespace swarm
