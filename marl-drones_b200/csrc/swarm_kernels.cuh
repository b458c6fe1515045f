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

// ---- byte <-> bit helpers of the state kernels: four cells at a time, no per-byte memory traffic --------------------------
// bit 7 of every byte of q that holds a value > 0 (int8) / a value == v
__device__ __forceinline__ uint32_t bytes_positive(uint32_t q) { return (((q & 0x7f7f7f7fu) + 0x7f7f7f7fu) | q) & ~q & 0x80808080u; }
__device__ __forceinline__ uint32_t bytes_equal(uint32_t q, uint32_t v4) {
  const uint32_t z = q ^ v4;
  return ~(((z & 0x7f7f7f7fu) + 0x7f7f7f7fu) | z) & 0x80808080u;
}
// bits 7 / 15 / 23 / 31 -> a nibble (the partial products land on distinct bits: no carries)
__device__ __forceinline__ uint32_t gather_nibble(uint32_t m) { return (((m >> 7) * 0x01020408u) >> 24) & 0xfu; }
// a nibble -> four bytes of 0 / 1
__device__ __forceinline__ uint32_t spread_nibble(uint32_t n) { return (n * 0x00204081u) & 0x01010101u; }

// set_state: boundary int8 [count][RC] (values 0 / 1 / U, obs order) -> coverage plane + anomalies; recomputes `remaining`.
// LIST mode: the cells of value U become the env's list in flat-cell order (at most n_anamolous of them — more is flagged in the
// device error word, SWARM_DEVERR_BAD_STATE); when the injected grid holds fewer (a state from the middle of an episode: found
// anomalies read 0), the spare entries repeat a listed cell, or hold an out-of-grid key when there is none — entries that can
// never turn a normal cell into an anomalous one.
// One warp per env.  A lane turns 32 cells into one plane word from two 16-byte loads (grids whose rows start 16-byte aligned;
// byte loads otherwise) and the list is a warp-wide ordered compaction of the U bits: 268 MB of cfg3 grids in ~70 us, where the
// per-byte version with its one-lane list walk took 4.1 ms.
__global__ void __launch_bounds__(256) set_state_kernel(const __grid_constant__ DevParams p, const int8_t* grid, const float* pos,
                                                        const int* t, const int16_t* wait, int first, int count) {
  const int lane = threadIdx.x & 31;
  const int gw = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int nw = gridDim.x * (blockDim.x >> 5);
  const bool vec = grid && (p.RC & 15) == 0 && (((uintptr_t)grid) & 15) == 0;
  const uint32_t u4 = (uint32_t)p.U * 0x01010101u;
  for (int e = gw; e < count; e += nw) {
    const int env = first + e;
    int rem = 0;
    if (grid) {
      const int8_t* src = grid + (size_t)e * p.RC;
      uint32_t* l = p.alist ? p.alist + (size_t)env * 8 : nullptr;
      int n = 0, mapped = 0x7fffffff;   // n: anomalies seen so far (warp-uniform); mapped: first cell of value <= 0 (per lane)
      for (int wb = 0; wb < p.env_words; wb += 32) {
        const int w = wb + lane;
        const int nvalid = min(max(p.RC - w * 32, 0), 32);
        uint32_t cw = 0u, aw = 0u;
        if (w < p.env_words && nvalid > 0) {
          if (vec && nvalid == 32) {
            const uint4* s4 = reinterpret_cast<const uint4*>(src + (size_t)w * 32);
            const uint4 a = __ldg(s4), b = __ldg(s4 + 1);
            const uint32_t q[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
#pragma unroll
            for (int k = 0; k < 8; ++k) {
              cw |= gather_nibble(bytes_positive(q[k])) << (4 * k);
              aw |= gather_nibble(bytes_equal(q[k], u4)) << (4 * k);
            }
          } else {
            for (int k = 0; k < nvalid; ++k) {
              const int v = src[w * 32 + k];
              if (v > 0) cw |= 1u << k;
              if (v == p.U) aw |= 1u << k;
            }
          }
          rem += __popc(cw);
          const uint32_t inv = ~cw & (nvalid == 32 ? 0xffffffffu : ((1u << nvalid) - 1u));
          if (inv) mapped = min(mapped, w * 32 + __ffs(inv) - 1);
        }
        if (w < p.env_words) {
          st_grid(p.cov + (size_t)env * p.env_words + w, cw);
          if (p.anomp) st_grid(p.anomp + (size_t)env * p.env_words + w, aw);
        }
        if (l) {   // ordered compaction: lanes hold ascending words, bits ascend inside a word
          const int cnt = __popc(aw);
          int incl = cnt;
#pragma unroll
          for (int o = 1; o < 32; o <<= 1) {
            const int v = __shfl_up_sync(FULL, incl, o);
            if (lane >= o) incl += v;
          }
          int at = n + incl - cnt;
          while (aw && at < p.n_anom) {
            const int c = w * 32 + __ffs(aw) - 1;
            l[at++] = (uint32_t)(c % p.R) | ((uint32_t)(c / p.R) << 16);
            aw &= aw - 1u;
          }
          n += __shfl_sync(FULL, incl, 31);
        }
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        rem += __shfl_xor_sync(FULL, rem, o);
        mapped = min(mapped, __shfl_xor_sync(FULL, mapped, o));
      }
      if (l) {
        __syncwarp();   // the entries other lanes wrote are visible to lane 0
        if (lane == 0) {
          if (n > p.n_anom) atomicOr(p.err, 8);
          uint32_t fill = LIST_EMPTY;                                // outside every LIST-mode grid
          if (n > 0) fill = l[0];
          else if (mapped != 0x7fffffff) fill = (uint32_t)(mapped % p.R) | ((uint32_t)(mapped / p.R) << 16);
          for (int k = min(n, p.n_anom); k < 8; ++k) l[k] = fill;
        }
      }
    }
    if (pos)
      for (int k = lane; k < p.D; k += 32)
        p.pos[(size_t)env * p.D + k] = make_float2(pos[((size_t)e * p.D + k) * 2], pos[((size_t)e * p.D + k) * 2 + 1]);
    if (wait && p.has_wait) {
      if ((p.RC & 7) == 0 && (((uintptr_t)wait) & 15) == 0) {   // eight counters per access
        const uint4* s4 = reinterpret_cast<const uint4*>(wait + (size_t)e * p.RC);
        uint4* d4 = reinterpret_cast<uint4*>(p.wait + (size_t)env * p.RC);
        for (int k = lane; k < (p.RC >> 3); k += 32) d4[k] = __ldg(s4 + k);
      } else {
        for (int k = lane; k < p.RC; k += 32) p.wait[(size_t)env * p.RC + k] = wait[(size_t)e * p.RC + k];
      }
    }
    if (lane == 0) {
      int4 m = p.meta[env];
      if (t) m.x = t[e];
      if (grid) m.y = rem;
      m.z = 0;
      p.meta[env] = m;
    }
  }
}

// get_state: the inverse.  16 cells per lane and store (plane bits spread to bytes; PLANE mode adds (U - 1) where the anomaly
// bit is set too; LIST mode patches the listed cells that are still unmapped afterwards) when the env's grid starts 16-byte
// aligned, else one cell at a time.
__global__ void __launch_bounds__(256) get_state_kernel(const __grid_constant__ DevParams p, int8_t* grid, float* pos, int* t,
                                                        int16_t* wait, int first, int count) {
  const int lane = threadIdx.x & 31;
  const int gw = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int nw = gridDim.x * (blockDim.x >> 5);
  const bool vec = grid && (p.RC & 15) == 0 && (((uintptr_t)grid) & 15) == 0;
  for (int e = gw; e < count; e += nw) {
    const int env = first + e;
    if (grid && vec) {
      const uint32_t* cv = p.cov + (size_t)env * p.env_words;
      const uint32_t* av = p.anomp ? p.anomp + (size_t)env * p.env_words : nullptr;
      int8_t* dst = grid + (size_t)e * p.RC;
      const uint32_t um1 = (uint32_t)(p.U - 1);
      for (int ch = lane; ch < (p.RC >> 4); ch += 32) {
        const int c0 = ch * 16;
        const uint32_t bits = (__ldcg(cv + (c0 >> 5)) >> (c0 & 31)) & 0xffffu;
        const uint32_t ab = av ? ((__ldcg(av + (c0 >> 5)) >> (c0 & 31)) & bits) : 0u;
        uint4 o;
        o.x = spread_nibble(bits & 15u) + spread_nibble(ab & 15u) * um1;
        o.y = spread_nibble((bits >> 4) & 15u) + spread_nibble((ab >> 4) & 15u) * um1;
        o.z = spread_nibble((bits >> 8) & 15u) + spread_nibble((ab >> 8) & 15u) * um1;
        o.w = spread_nibble((bits >> 12) & 15u) + spread_nibble((ab >> 12) & 15u) * um1;
        reinterpret_cast<uint4*>(dst)[ch] = o;
      }
      if (p.alist) {
        __syncwarp();   // the patch overwrites bytes other lanes have just written
        if (lane < p.n_anom) {
          const uint32_t key = p.alist[(size_t)env * 8 + lane];
          const int x = (int)(key & 0xffffu), y = (int)(key >> 16);
          if (x < p.R && y < p.C) {
            const int c = y * p.R + x;
            if ((__ldcg(cv + (c >> 5)) >> (c & 31)) & 1u) dst[c] = (int8_t)p.U;
          }
        }
      }
    } else if (grid) {
      for (int c = lane; c < p.RC; c += 32) {
        const uint32_t code = cell_code_global(p, env, c);
        grid[(size_t)e * p.RC + c] = (int8_t)(code == CODE_ANOM ? p.U : (int)code);
      }
    }
    if (pos)
      for (int k = lane; k < p.D; k += 32) {
        const float2 q = p.pos[(size_t)env * p.D + k];
        pos[((size_t)e * p.D + k) * 2] = q.x;
        pos[((size_t)e * p.D + k) * 2 + 1] = q.y;
      }
    if (wait) {
      if ((p.RC & 7) == 0 && (((uintptr_t)wait) & 15) == 0) {   // eight counters per access
        const uint4* s4 = p.has_wait ? reinterpret_cast<const uint4*>(p.wait + (size_t)env * p.RC) : nullptr;
        uint4* d4 = reinterpret_cast<uint4*>(wait + (size_t)e * p.RC);
        for (int k = lane; k < (p.RC >> 3); k += 32) d4[k] = s4 ? s4[k] : make_uint4(0u, 0u, 0u, 0u);
      } else {
        for (int k = lane; k < p.RC; k += 32) wait[(size_t)e * p.RC + k] = p.has_wait ? p.wait[(size_t)env * p.RC + k] : (int16_t)0;
      }
    }
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
//
// The scan itself is NOT the reference's R*C loop (4.3 ms per call at cfg3): the winner of that loop is the cell with the smallest
// key (d2, gx, gy) — strict `>` keeps the first of equal distances, and the scan order is gx-major — so any order of evaluation
// finds it.  Rows are visited outward from the drone's row (offset j: cy - j, cy + j); in a row only the nearest unmapped cell on
// each side of x can win, found per 32-cell word with CLZ / FFS; the search stops once (j - 0.5)^2 exceeds the best d2 (no row
// further out can even tie).  The quirk: a drone standing exactly on an unmapped cell c0 ends up with the nearest cell AFTER c0
// in scan order (gx > x0, or gx == x0 and gy > y0), or with c0 itself when there is none — the same search with a lower bound
// on gx per row.
struct GreedyBest {
  double d2;
  int gx, gy;
  __device__ __forceinline__ void consider(int x_, int y_, double x, double y) {
    const double dx = x_ - x, dy = y_ - y, d = dx * dx + dy * dy;
    if (gx < 0 || d < d2 || (d == d2 && (x_ < gx || (x_ == gx && y_ < gy)))) { d2 = d; gx = x_; gy = y_; }
  }
};

// One 32-cell word `w` (word index k of the plane) of row gy = bits [s, s + R): feeds the unmapped cells nearest to `t`
// (= floor(x)) on both sides into `best`, among cells with gx >= gmin.
__device__ __forceinline__ void greedy_word(uint32_t w, int k, int s, int R, int gy, int t, int gmin, double x, double y, GreedyBest& best) {
  const int base = k * 32 - s;                         // gx of bit 0 of this word
  const int lo = max(max(s - k * 32, 0), gmin - base), hi = min(s + R - k * 32, 32);
  if (lo >= hi) return;
  w &= (hi == 32 ? 0xffffffffu : ((1u << hi) - 1u)) & ~((1u << lo) - 1u);   // lo < 32 here
  if (!w) return;
  const int tw = t - base;                             // t in word coordinates
  const uint32_t left = tw >= 31 ? w : (tw < 0 ? 0u : (w & ((2u << tw) - 1u)));
  const uint32_t right = w & ~left;
  if (left) best.consider(base + 31 - __clz(left), gy, x, y);
  if (right) best.consider(base + __ffs(right) - 1, gy, x, y);
}

// Row gy of the env's plane `g` (any alignment; rows of whole 128-cell chunks are read 16 bytes at a time and empty chunks —
// most of a late-episode grid — cost one test).
__device__ __forceinline__ void greedy_row(const uint32_t* __restrict__ g, int R, int gy, int t, int gmin, double x, double y, GreedyBest& best) {
  const int s = gy * R;
  if ((R & 127) == 0) {
    const uint4* g4 = reinterpret_cast<const uint4*>(g) + (s >> 7);
    for (int c = 0; c < (R >> 7); ++c) {
      const uint4 v = __ldg(g4 + c);
      if (!(v.x | v.y | v.z | v.w)) continue;
      const int k = (s >> 5) + 4 * c;
      if (v.x) greedy_word(v.x, k, s, R, gy, t, gmin, x, y, best);
      if (v.y) greedy_word(v.y, k + 1, s, R, gy, t, gmin, x, y, best);
      if (v.z) greedy_word(v.z, k + 2, s, R, gy, t, gmin, x, y, best);
      if (v.w) greedy_word(v.w, k + 3, s, R, gy, t, gmin, x, y, best);
    }
    return;
  }
  for (int k = s >> 5; k <= (s + R - 1) >> 5; ++k) {
    const uint32_t w = __ldg(g + k);
    if (w) greedy_word(w, k, s, R, gy, t, gmin, x, y, best);
  }
}

__global__ void __launch_bounds__(256) greedy_kernel(const __grid_constant__ DevParams p, uint8_t* actions) {
  const long long n = (long long)p.N * p.D;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const int env = (int)(i / p.D);
    const uint32_t* g = p.cov + (size_t)env * p.env_words;
    const float2 q = p.pos[i];
    const double x = q.x, y = q.y;
    const int cx = (int)floorf(q.x + 0.5f), cy = (int)floorf(q.y + 0.5f), t = (int)floorf(q.x);
    // does the drone stand exactly on an unmapped cell?  (then only cells after it in scan order compete)
    bool on_cell = false;
    if ((float)cx == q.x && (float)cy == q.y) {
      const int c0 = cy * p.R + cx;
      on_cell = (__ldg(g + (c0 >> 5)) >> (c0 & 31)) & 1u;
    }
    GreedyBest best;
    best.d2 = 0.0; best.gx = -1; best.gy = -1;
    const int jmax = p.meta[env].y == 0 ? -1 : max(cy, p.C - 1 - cy);   // nothing left to map: no search (action 0)
    for (int j = 0; j <= jmax; ++j) {
      if (j > 0 && best.gx >= 0 && ((double)j - 0.5) * ((double)j - 0.5) > best.d2) break;   // rows at offset >= j: d2 >= (j - 0.5)^2
      const int ya = cy - j, yb = cy + j;
      if (ya >= 0) greedy_row(g, p.R, ya, t, on_cell ? cx + 1 : 0, x, y, best);                       // gy <= y0: gx > x0
      if (j > 0 && yb < p.C) greedy_row(g, p.R, yb, t, on_cell ? cx : 0, x, y, best);                 // gy > y0: gx >= x0
    }
    int bx = best.gx, by = best.gy;
    if (bx < 0 && on_cell) { bx = cx; by = cy; }       // nothing after c0: the running minimum stays c0
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
