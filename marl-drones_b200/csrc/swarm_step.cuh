// swarm_step.cuh — the fused env step, written warp-cooperatively.
//
// A warp owns EPW = 32/TS consecutive envs per iteration; each env gets a tile of TS lanes, one lane per drone
// (TS = smallest power of two >= D, capped at 32; D > 32 runs TS = 32 in index-ordered chunks).
//
// Stages (north-star numbering) and the reference lines they replace:
//   (1) action decode + move + clip                A2C/noVis_sim.py:117-135
//   (2) coverage scatter + conflict resolution     A2C/noVis_sim.py:75-90, 104-110   (__match_any_sync dedupe)
//   (3) reward tail, spread, done (warp shuffles)  A2C/noVis_sim.py:92-102, 141-166, 172-173
//                                                  DQL_model/noVis_sim.py:138-148
//       auto-reset on done                         A2C/envs.py:22-26 (+ init_env, noVis_sim.py:36-63)
//   (4) observation emission                       A2C/noVis_sim.py:168-170 / SURVEY App. B.1
#pragma once
#include "swarm_common.cuh"

namespace swarm {

constexpr int F_RESET = 16;   // auto-reset on done inside the step (A2C/envs.py:22-26)
constexpr int F_PUSH = 32;    // publish the running statistics to the ranks' mailboxes (swarm_step / swarm_rollout only: their
                              // statistics buffer holds running totals; swarm_step_host's is zeroed per call)

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
  int dyn;                 // staged kernel: CTA-groups handed out by the atomic counter DevParams::sched (else static striding)
};

__device__ __forceinline__ int load_action(const void* a, int dt, long long idx) {
  if (dt == 0) return (int)__ldg((const uint8_t*)a + idx);   // actions are read-only for the whole launch
  if (dt == 1) return __ldg((const int*)a + idx);
  long long v = __ldg((const long long*)a + idx);
  return (v < 0 || v > 3) ? -1 : (int)v;
}

// ------------------------------------------------------------------------------------------------ reset
// init_env (A2C/noVis_sim.py:36-63) for one env by its tile.  Called by all 32 lanes; `doit` is tile-uniform.
// RNG spec (identical in oracle_reset_env): anomaly attempt j -> Philox block j>>1, words (2(j&1), 2(j&1)+1);
// drone i -> block 0x80000000|(i>>1); bounded ints by mulhi.
// `cov` / `anomp` / `alist` point at the env's planes / list — in HBM (reset_kernel, CTA kernel) or in the shared-memory
// stage of the warp kernel (which then copies them out); generic addressing serves both.
template <int TS>
__device__ __forceinline__ void reset_planes(const DevParams& p, int env, int episode, int tl, bool doit, uint32_t* cov, uint32_t* anomp,
                                             uint32_t* alist) {
  int16_t* wt = p.has_wait ? p.wait + (size_t)(doit ? env : 0) * p.RC : nullptr;
  const long long gid = p.env_id_base + env;
  if (doit) {
    for (int w = tl; w < p.env_words; w += TS) {
      const int n = p.RC - w * 32;
      cov[w] = n >= 32 ? 0xffffffffu : (n <= 0 ? 0u : ((1u << n) - 1u));
      if (anomp) anomp[w] = 0u;
    }
    if (alist && tl < 8) alist[tl] = LIST_EMPTY;
    if (wt)
      for (int k = tl; k < p.RC; k += TS) wt[k] = (int16_t)p.dwell_normal;
  }
  __syncwarp();
  if (doit && tl == 0) {
    int placed = 0;
    uint4 r = make_uint4(0, 0, 0, 0);
    for (uint32_t j = 0; placed < p.n_anom; ++j) {   // noVis_sim.py:47-55: redraw until n_anamolous DISTINCT cells are marked
      if ((j & 1u) == 0) r = draw_block(p, gid, (uint32_t)episode, j >> 1);
      const uint32_t a = __umulhi((j & 1u) ? r.z : r.x, (uint32_t)p.R);
      const uint32_t b = __umulhi((j & 1u) ? r.w : r.y, (uint32_t)p.C);
      const int cell = (int)(b * (uint32_t)p.R + a);
      bool fresh = true;
      if (alist) {
        const uint32_t key = a | (b << 16);
        for (int k = 0; k < placed; ++k) fresh = fresh && alist[k] != key;
        if (fresh) alist[placed] = key;
      } else {
        const uint32_t word = anomp[cell >> 5];
        fresh = ((word >> (cell & 31)) & 1u) == 0u;
        if (fresh) anomp[cell >> 5] = word | (1u << (cell & 31));
      }
      if (fresh) {
        if (wt) wt[cell] = (int16_t)p.dwell_anom;
        ++placed;
      }
    }
  }
  __syncwarp();
}

// reset of an env whose state lives in HBM only (reset_kernel, the CTA kernel)
template <int TS>
__device__ __forceinline__ void reset_tile(const DevParams& p, int env, int episode, int tl, bool doit) {
  const size_t e = (size_t)(doit ? env : 0);
  reset_planes<TS>(p, env, episode, tl, doit, p.cov + e * p.env_words, p.anomp ? p.anomp + e * p.env_words : nullptr,
                   p.alist ? p.alist + e * 8 : nullptr);
  const long long gid = p.env_id_base + env;
  if (doit) {
    if (tl == 0) p.meta[env] = make_int4(0, p.RC, 0, episode);
    for (int d = tl; d < p.D; d += TS) {
      const uint4 r = draw_block(p, gid, (uint32_t)episode, 0x80000000u | ((uint32_t)d >> 1));
      const float x = (float)__umulhi((d & 1) ? r.z : r.x, (uint32_t)p.R);
      const float y = (float)__umulhi((d & 1) ? r.w : r.y, (uint32_t)p.C);
      p.pos[(size_t)env * p.D + d] = make_float2(x, y);
    }
  }
  __syncwarp();
}

// ------------------------------------------------------------------------------------------------ spread
// _drone_dist() > 0 (A2C/noVis_sim.py:92-102): any unordered pair farther apart than row_count // n_drones.
// Rotation s pairs lane i with lane (i+s) mod D; s = 1..floor(D/2) covers every pair; exits as soon as every
// tile that needs an answer has one (the first rotation almost always settles it).
// F32: squared distances are exact in float (DevParams::f32_exact) — same decisions, no FP64 pipe.
template <int TS, bool F32 = false>
__device__ __forceinline__ bool tile_spread(const DevParams& p, int env, int tile, int tl, bool need, float x, float y,
                                            const float2* sp = nullptr) {   // sp: the env's positions in shared memory (D > 32 path), or null
  const int D = p.D;
  const double thr2 = p.spread2_thr;   // sqrt(d2) > row_count // n_drones  <=>  d2 > spread2_thr
  const float thr2f = p.spread2_thr_f;
  const unsigned tm = (TS == 32) ? FULL : (((1u << TS) - 1u) << (tile * TS));
  bool found = false;
  if (D <= TS) {
    for (int s = 1; s <= D / 2; ++s) {
      int pl = tl + s;
      if (pl >= D) pl -= D;
      const float xs = __shfl_sync(FULL, x, tile * TS + (pl & (TS - 1)));
      const float ys = __shfl_sync(FULL, y, tile * TS + (pl & (TS - 1)));
      bool far;
      if (F32) {
        const float dx = xs - x, dy = ys - y;
        far = need && tl < D && (__fmaf_rn(dx, dx, dy * dy) > thr2f);
      } else {
        const double dx = (double)xs - (double)x, dy = (double)ys - (double)y;
        far = need && tl < D && (dx * dx + dy * dy > thr2);
      }
      const unsigned bal = __ballot_sync(FULL, far);
      found = found || ((bal & tm) != 0u);
      if (!__any_sync(FULL, need && !found)) break;
    }
  } else {   // D > 32: one env per warp, positions re-read from L2
    const float2* ps = p.pos + (size_t)env * D;
    for (int s = 1; s <= D / 2; ++s) {
      bool far = false;
      if (need)
        for (int i = tl; i < D; i += TS) {
          int j = i + s;
          if (j >= D) j -= D;
          const float2 a = sp ? sp[i] : __ldcg(ps + i), b = sp ? sp[j] : __ldcg(ps + j);
          const double dx = (double)b.x - (double)a.x, dy = (double)b.y - (double)a.y;
          far = far || (dx * dx + dy * dy > thr2);
        }
      found = found || (__ballot_sync(FULL, far) != 0u);
      if (!__any_sync(FULL, need && !found)) break;
    }
  }
  return found;
}

// ------------------------------------------------------------------------------------------------ scatter (generic cd)
// Sequential _det_collision for collision_dist > 0.5 (later drones see earlier clears, noVis_sim.py:82): the tile
// leader walks the drones in index order over the bounded window |g - pos| <= ceil(cd)+1.
// `sc` (may be null) is the shared-memory copy of the env's coverage plane kept by the staged kernels: reads come from it and
// every clear is written to both copies (only this thread touches this env's words during the launch).
__device__ __forceinline__ void scatter_sequential(const DevParams& p, int env, uint32_t* sc, int& rew, int& ncl, int& nan) {
  int16_t* wt = p.has_wait ? p.wait + (size_t)env * p.RC : nullptr;
  uint32_t* g = p.cov + (size_t)env * p.env_words;
  for (int d = 0; d < p.D; ++d) {
    const float2 q = __ldcg(p.pos + (size_t)env * p.D + d);
    const double x = q.x, y = q.y;
    const int x0 = max((int)floor(x) - p.scan_m, 0), x1 = min((int)ceil(x) + p.scan_m, p.R - 1);
    const int y0 = max((int)floor(y) - p.scan_m, 0), y1 = min((int)ceil(y) + p.scan_m, p.C - 1);
    bool hit = false;
    for (int gx = x0; gx <= x1 && !hit; ++gx)
      for (int gy = y0; gy <= y1; ++gy) {
        const double dx = (double)gx - x, dy = (double)gy - y;
        if (!(dx * dx + dy * dy < p.cd2_thr)) continue;   // sqrt(d2) < cd (noVis_sim.py:81)
        const int cell = gy * p.R + gx;
        const uint32_t word = sc ? sc[cell >> 5] : ld_grid(g + (cell >> 5));
        if (!((word >> (cell & 31)) & 1u)) continue;
        if (wt && wt[cell] > 0) {
          wt[cell] -= 1;
        } else {
          const uint32_t nw = word & ~(1u << (cell & 31));
          st_grid(g + (cell >> 5), nw);
          if (sc) sc[cell >> 5] = nw;
          const int an = is_anomalous(p, env, cell, gx, gy) ? 1 : 0;
          rew += 10 + 20 * an;                             // 10 + _check_uncertain_mapped (noVis_sim.py:83-84,104-110)
          ncl += 1;
          nan += an;
        }
        hit = true;
        break;
      }
  }
}

}  // namespace swarm
