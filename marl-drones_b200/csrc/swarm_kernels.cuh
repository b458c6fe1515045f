// swarm_kernels.cuh — the state-management kernels (create / reset, state injection / extraction, the standalone K1 move
// stage, the scripted greedy action source).  The fused step kernels live in swarm_warp.cuh and swarm_cta.cuh.
#pragma once
#include "swarm_common.cuh"
#include "swarm_cta.cuh"
#include "swarm_emit.cuh"
#include "swarm_step.cuh"
#include "swarm_warp.cuh"

namespace swarm {

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

// set_state: boundary int8 [count][RC] (values 0 / 1 / U, obs order) -> coverage plane + anomalies; recomputes `remaining`.
// LIST mode: the cells of value U become the env's list (at most n_anamolous of them — more is flagged in the device error
// word, SWARM_DEVERR_BAD_STATE); when the injected grid holds fewer (a state from the middle of an episode: found anomalies
// read 0), the spare entries repeat a listed cell, or hold an out-of-grid key when there is none — entries that can never
// turn a normal cell into an anomalous one.
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
        uint32_t cw = 0u, aw = 0u;
        for (int k = 0; k < 32; ++k) {
          const int c = w * 32 + k;
          if (c < p.RC) {
            const int v = src[c];
            if (v > 0) { cw |= 1u << k; ++rem; }
            if (v == p.U) aw |= 1u << k;
          }
        }
        st_grid(p.cov + (size_t)env * p.env_words + w, cw);
        if (p.anomp) st_grid(p.anomp + (size_t)env * p.env_words + w, aw);
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) rem += __shfl_xor_sync(FULL, rem, o);
      if (p.alist) {   // one lane walks the grid in flat order (state injection is not a hot path)
        __syncwarp();
        if (lane == 0) {
          uint32_t* l = p.alist + (size_t)env * 8;
          int n = 0, mapped = -1;
          for (int c = 0; c < p.RC; ++c) {
            const int v = src[c];
            if (v == p.U) {
              if (n < p.n_anom) l[n] = (uint32_t)(c % p.R) | ((uint32_t)(c / p.R) << 16);
              ++n;
            } else if (v <= 0 && mapped < 0) mapped = c;
          }
          if (n > p.n_anom) atomicOr(p.err, 8);
          uint32_t fill = LIST_EMPTY;                                // outside every LIST-mode grid
          if (n > 0) fill = l[0];
          else if (mapped >= 0) fill = (uint32_t)(mapped % p.R) | ((uint32_t)(mapped / p.R) << 16);
          for (int k = min(n, p.n_anom); k < 8; ++k) l[k] = fill;
        }
      }
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
        const uint32_t code = cell_code_global(p, env, c);
        grid[(size_t)e * p.RC + c] = (int8_t)(code == CODE_ANOM ? p.U : (int)code);
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

// The W == 5 window tables of swarm_window.cuh, built once per handle into DevParams::tables (the CTA kernel brings them on
// chip with three bulk copies instead of recomputing them in each of its one-env CTAs).
template <typename OT>
__global__ void __launch_bounds__(256) build_tables_kernel(const __grid_constant__ DevParams p, uint32_t* tab) {
  const ObsCvt<OT> cvt(p);
  build_window_tables<OT>(p, cvt, tab, 256, tab + 1024, tab + 1280, threadIdx.x, blockDim.x);
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

// Scripted greedy action source on the noVis lattice.  The TARGET is the 2D simulator's (Drone.update,
// swarmsim-2d/drone.py:29-47; oracle: drone_update_target, pinned by tests/golden/drone_update_targets.npz): row-major scan,
// mapped cells skipped, strict `min_dist > dist`, and the `if not min_dist` quirk — a running minimum of exactly 0.0 counts as
// "no minimum yet", so the next unmapped cell in scan order replaces it unconditionally.  The MOVE is a 4-action variant (the
// reference steps +-1 on both axes at once, which the noVis action set cannot express): the axis with the larger |delta|.
// Distances: the reference compares ((dx)**2 + (dy)**2) ** 0.5; on the lattice (multiples of 2^-k) the squares are exact and
// sqrt is strictly monotone on them at these magnitudes, so comparing the squared distances picks the same cell.
__global__ void __launch_bounds__(256) greedy_kernel(const __grid_constant__ DevParams p, uint8_t* actions) {
  const long long n = (long long)p.N * p.D;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const int env = (int)(i / p.D);
    const uint32_t* g = p.cov + (size_t)env * p.env_words;
    const float2 q = p.pos[i];
    const double x = q.x, y = q.y;
    double best = 0.0;
    bool have = false;
    int bx = -1, by = -1;
    for (int gx = 0; gx < p.R; ++gx)
      for (int gy = 0; gy < p.C; ++gy) {
        const int cell = gy * p.R + gx;
        if ((ld_grid(g + (cell >> 5)) >> (cell & 31)) & 1u) {
          const double d2 = (gx - x) * (gx - x) + (gy - y) * (gy - y);
          if (!have || best == 0.0 || d2 < best) { have = true; best = d2; bx = gx; by = gy; }
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
