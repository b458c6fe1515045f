/*
 * swarm_oracle.c — CPU restatement of the MARL-Drones headless 2D env (TEST INFRASTRUCTURE).
 *
 * This file is the ORACLE: only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * `--impl reference` legs may load it.  It is never on the product path (the product is the
 * CUDA library in marl-drones_b200/csrc; it has no CPU fallback).
 *
 * What it restates (reference paths relative to /root/reference):
 *   init_env            A2C/noVis_sim.py:36-63   (RNG replaced by Philox, see oracle_reset_env;
 *                                                 a numpy-MT19937-exact variant is oracle_reset_numpy)
 *   step                A2C/noVis_sim.py:112-175 (DQL reward tail: DQL_model/noVis_sim.py:138-148)
 *   _det_collision      A2C/noVis_sim.py:75-90   (literal full scan, or the bounded window that
 *                                                 SURVEY App. A proves equivalent)
 *   _check_uncertain_mapped A2C/noVis_sim.py:104-110 (== +20 iff the cleared cell held U, U>1)
 *   _drone_dist         A2C/noVis_sim.py:92-102
 *   obs                 A2C/noVis_sim.py:168-170 (grid.T.flatten)
 *   auto-reset on done  A2C/envs.py:22-26
 *   dwell counters      swarmsim-2d/goals.py:4-41 + run_env.py:183-201 (the counter semantics are
 *                                                 pinned by tests/golden/goals_dwell.npz, frozen from the unmodified goals.py)
 *   greedy target       swarmsim-2d/drone.py:29-47 (pinned by tests/golden/drone_update_targets.npz)
 *   window obs          SURVEY App. B.1 (ours; parity unpinned)
 *
 * Parity pin: tests/test_oracle_vs_reference.py drives this file and the live reference with the
 * same states/actions (when /root/reference is present) and tests/golden/ holds traces frozen
 * from the live reference (tests/golden/make_golden.py).
 *
 * Arithmetic is the reference's: positions and distances in IEEE double, sqrt() as math.sqrt.
 * Layout: grid_t[y*R + x] == reference grid[x][y] (the obs order), one int8 per cell.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

typedef struct OracleConfig {
  int32_t row_count, col_count, n_drones, n_anamolous, uncertainity;
  int32_t reward_mode;   /* 0 A2C, 1 DQL */
  int32_t obs_mode;      /* 0 full, 1 window */
  int32_t window;
  int32_t auto_reset;
  int32_t dwell_normal, dwell_anom;
  int32_t scan_mode;     /* 0 literal R*C scan (the reference), 1 bounded window (App. A) */
  double collision_dist, step_size;
  uint64_t seed;
  int64_t env_id_base;
} OracleConfig;

enum { META_T = 0, META_REMAINING = 1, META_EPRET = 2, META_EPISODE = 3 };

/* ---------------------------------------------------------------- Philox4x32-10 (Salmon et al. 2011) */
static inline uint32_t mulhi32(uint32_t a, uint32_t b) { return (uint32_t)(((uint64_t)a * b) >> 32); }

void oracle_philox4x32_10(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]) {
  uint32_t c0 = ctr[0], c1 = ctr[1], c2 = ctr[2], c3 = ctr[3], k0 = key[0], k1 = key[1];
  for (int r = 0; r < 10; ++r) {
    uint32_t hi0 = mulhi32(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
    uint32_t hi1 = mulhi32(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
    uint32_t n0 = hi1 ^ c1 ^ k0, n1 = lo1, n2 = hi0 ^ c3 ^ k1, n3 = lo0;
    c0 = n0; c1 = n1; c2 = n2; c3 = n3;
    k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
  }
  out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

static void draw_block(const OracleConfig* cfg, int64_t env_id, uint32_t episode, uint32_t block, uint32_t out[4]) {
  uint32_t ctr[4] = {(uint32_t)(uint64_t)env_id, (uint32_t)((uint64_t)env_id >> 32), episode, block};
  uint32_t key[2] = {(uint32_t)cfg->seed, (uint32_t)(cfg->seed >> 32)};
  oracle_philox4x32_10(ctr, key, out);
}

/* init_env (A2C/noVis_sim.py:36-63) with the device RNG spec:
 *   anomaly attempt j  -> block j>>1, words (2*(j&1), 2*(j&1)+1) -> (a, b) = (mulhi(r, R), mulhi(r', C))
 *   drone i            -> block 0x80000000 | (i>>1), same word choice by i&1
 * grid all 1, n_anamolous distinct cells set to U (only where the cell is still 1), drones anywhere. */
void oracle_reset_env(const OracleConfig* cfg, int64_t env_id, int32_t episode, int8_t* grid_t, double* pos,
                      int32_t* meta, int16_t* wait) {
  const int R = cfg->row_count, C = cfg->col_count, D = cfg->n_drones;
  memset(grid_t, 1, (size_t)R * C);
  if (wait) for (int k = 0; k < R * C; ++k) wait[k] = (int16_t)cfg->dwell_normal;
  int placed = 0;
  uint32_t r[4];
  for (uint32_t j = 0; placed < cfg->n_anamolous; ++j) {
    if ((j & 1u) == 0) draw_block(cfg, env_id, (uint32_t)episode, j >> 1, r);
    uint32_t a = mulhi32(r[2 * (j & 1u)], (uint32_t)R), b = mulhi32(r[2 * (j & 1u) + 1], (uint32_t)C);
    if (grid_t[b * R + a] == 1) {
      grid_t[b * R + a] = (int8_t)cfg->uncertainity;
      if (wait) wait[b * R + a] = (int16_t)cfg->dwell_anom;
      ++placed;
    }
  }
  for (int i = 0; i < D; ++i) {
    draw_block(cfg, env_id, (uint32_t)episode, 0x80000000u | ((uint32_t)i >> 1), r);
    pos[2 * i + 0] = (double)mulhi32(r[2 * (i & 1)], (uint32_t)R);
    pos[2 * i + 1] = (double)mulhi32(r[2 * (i & 1) + 1], (uint32_t)C);
  }
  meta[META_T] = 0; meta[META_REMAINING] = R * C; meta[META_EPRET] = 0; meta[META_EPISODE] = episode;
}

/* ---------------------------------------------------------------- numpy legacy MT19937 (for reference-exact resets) */
typedef struct { uint32_t mt[624]; int idx; } MT;
static void mt_seed(MT* s, uint32_t seed) {   /* numpy _legacy_seeding with an int seed == init_genrand */
  s->mt[0] = seed;
  for (int i = 1; i < 624; ++i) s->mt[i] = 1812433253u * (s->mt[i - 1] ^ (s->mt[i - 1] >> 30)) + (uint32_t)i;
  s->idx = 624;
}
static uint32_t mt_next(MT* s) {
  if (s->idx >= 624) {
    uint32_t* mt = s->mt;
    for (int i = 0; i < 624; ++i) {
      uint32_t y = (mt[i] & 0x80000000u) | (mt[(i + 1) % 624] & 0x7fffffffu);
      mt[i] = mt[(i + 397) % 624] ^ (y >> 1) ^ ((y & 1u) ? 0x9908b0dfu : 0u);
    }
    s->idx = 0;
  }
  uint32_t y = s->mt[s->idx++];
  y ^= y >> 11; y ^= (y << 7) & 0x9d2c5680u; y ^= (y << 15) & 0xefc60000u; y ^= y >> 18;
  return y;
}
/* np.random.randint(n) (legacy RandomState, default dtype): masked rejection on 32-bit draws; no draw when n == 1 */
static uint32_t mt_randint(MT* s, uint32_t n) {
  uint32_t rng = n - 1;
  if (rng == 0) return 0;
  uint32_t mask = rng;
  mask |= mask >> 1; mask |= mask >> 2; mask |= mask >> 4; mask |= mask >> 8; mask |= mask >> 16;
  uint32_t v;
  do { v = mt_next(s) & mask; } while (v > rng);
  return v;
}
/* init_env exactly as `np.random.seed(seed); DroneEnv.init_env()` (A2C/noVis_sim.py:36-63), continuing the
 * stream for `n_resets` consecutive resets and returning the LAST one (so seed + reset count pins a state). */
void oracle_reset_numpy(const OracleConfig* cfg, uint32_t seed, int32_t n_resets, int8_t* grid_t, double* pos,
                        int32_t* meta, int16_t* wait) {
  const int R = cfg->row_count, C = cfg->col_count, D = cfg->n_drones;
  MT s; mt_seed(&s, seed);
  for (int k = 0; k < n_resets; ++k) {
    memset(grid_t, 1, (size_t)R * C);
    if (wait) for (int q = 0; q < R * C; ++q) wait[q] = (int16_t)cfg->dwell_normal;
    int placed = 0;
    while (placed < cfg->n_anamolous) {
      uint32_t a = mt_randint(&s, (uint32_t)R), b = mt_randint(&s, (uint32_t)C);
      if (grid_t[b * R + a] == 1) {
        grid_t[b * R + a] = (int8_t)cfg->uncertainity;
        if (wait) wait[b * R + a] = (int16_t)cfg->dwell_anom;
        ++placed;
      }
    }
    for (int i = 0; i < D; ++i) {
      uint32_t x = mt_randint(&s, (uint32_t)R), y = mt_randint(&s, (uint32_t)C);
      pos[2 * i] = (double)x; pos[2 * i + 1] = (double)y;
    }
  }
  meta[META_T] = 0; meta[META_REMAINING] = R * C; meta[META_EPRET] = 0;
}

/* ---------------------------------------------------------------- observation */
void oracle_emit_obs(const OracleConfig* cfg, const int8_t* grid_t, const double* pos, float* obs) {
  const int R = cfg->row_count, C = cfg->col_count, D = cfg->n_drones;
  if (cfg->obs_mode == 0) {                               /* grid.T.flatten(), A2C/noVis_sim.py:168-170 */
    for (int k = 0; k < R * C; ++k) obs[k] = (float)grid_t[k];
    return;
  }
  const int W = cfg->window, h = (W - 1) / 2;             /* SURVEY App. B.1 */
  for (int i = 0; i < D; ++i) {
    int cx = (int)floor(pos[2 * i] + 0.5), cy = (int)floor(pos[2 * i + 1] + 0.5);
    for (int b = 0; b < W; ++b)
      for (int a = 0; a < W; ++a) {
        int gx = cx - h + a, gy = cy - h + b;
        obs[(size_t)i * W * W + b * W + a] = (gx >= 0 && gx < R && gy >= 0 && gy < C) ? (float)grid_t[gy * R + gx] : -1.0f;
      }
  }
}

/* ---------------------------------------------------------------- one env, one step */
static inline double clipd(double v, double lo, double hi) { return v < lo ? lo : (v > hi ? hi : v); }

/* returns the step reward; *done_out as the reference's done; counts of newly mapped cells go to cleared[2] */
static double step_env(const OracleConfig* cfg, int8_t* g, double* pos, int32_t* meta, int16_t* wait,
                       const uint8_t* act, uint8_t* done_out, int32_t cleared[2], int32_t* bad_action) {
  const int R = cfg->row_count, C = cfg->col_count, D = cfg->n_drones, U = cfg->uncertainity;
  const double cd = cfg->collision_dist, ss = cfg->step_size;
  double total = 0.0;
  for (int i = 0; i < D; ++i) {                            /* A2C/noVis_sim.py:117-139, drones in index order */
    double x = pos[2 * i], y = pos[2 * i + 1];
    switch (act[i]) {
      case 0: y -= ss; break;
      case 1: x += ss; break;
      case 2: y += ss; break;
      case 3: x -= ss; break;
      default: *bad_action = 1; break;                      /* the reference asserts (:118-120) */
    }
    x = clipd(x, 0.0, (double)(R - 1));
    y = clipd(y, 0.0, (double)(C - 1));
    pos[2 * i] = x; pos[2 * i + 1] = y;
    /* _det_collision (:75-90): row-major scan (x outer, y inner), first hit only */
    int x0 = 0, x1 = R - 1, y0 = 0, y1 = C - 1;
    if (cfg->scan_mode == 1) {
      int m = (int)ceil(cd) + 1;
      x0 = (int)floor(x) - m; if (x0 < 0) x0 = 0;
      x1 = (int)ceil(x) + m;  if (x1 > R - 1) x1 = R - 1;
      y0 = (int)floor(y) - m; if (y0 < 0) y0 = 0;
      y1 = (int)ceil(y) + m;  if (y1 > C - 1) y1 = C - 1;
    }
    int hit = 0;
    for (int gx = x0; gx <= x1 && !hit; ++gx)
      for (int gy = y0; gy <= y1; ++gy) {
        double dx = (double)gx - x, dy = (double)gy - y;
        double dist = sqrt(dx * dx + dy * dy);
        int8_t* cell = &g[gy * R + gx];
        if (dist < cd && *cell > 0) {
          if (wait && wait[gy * R + gx] > 0) {              /* dwell: App. B.2 */
            wait[gy * R + gx] -= 1;
          } else {
            total += 10.0 + ((*cell == U) ? 20.0 : 0.0);    /* :83-84 + :104-110 */
            cleared[0] += 1; cleared[1] += (*cell == U);
            *cell = 0;
            meta[META_REMAINING] -= 1;
          }
          hit = 1;
          break;
        }
      }
  }
  /* reward tail */
  const int punish = (double)meta[META_T] > ((double)(R * C)) / ss;     /* :141-143 */
  int spread = 0;
  if (!punish) {                                                         /* _drone_dist :92-102 */
    const int thr = R / D;
    for (int i = 0; i < D - 1 && !spread; ++i)
      for (int j = i + 1; j < D; ++j) {
        double dx = pos[2 * j] - pos[2 * i], dy = pos[2 * j + 1] - pos[2 * i + 1];
        if (sqrt(dx * dx + dy * dy) > (double)thr) { spread = 1; break; }
      }
  }
  if (cfg->reward_mode == 0) {           /* A2C :150-152 */
    if (!punish) total += spread ? 1.0 : 0.0;
    total -= 1.0;
  } else {                               /* DQL DQL_model/noVis_sim.py:141-148 */
    if (punish) total -= 1.0; else total += spread ? 1.0 : 0.0;
  }
  meta[META_T] += 1;                                                     /* :156 */
  int done = 1;                                                          /* :157-166 */
  for (int k = 0; k < R * C; ++k) if (g[k] > 0) { done = 0; break; }
  if (cfg->reward_mode == 0 && done) total += 100.0;                     /* :172-173 */
  *done_out = (uint8_t)done;
  return total;
}

/* ---------------------------------------------------------------- batched entry points */
/* stats: int64[8] accumulated. obs may be NULL. */
void oracle_step_batch(const OracleConfig* cfg, int32_t N, int8_t* grid_t, double* pos, int32_t* meta, int16_t* wait,
                       const uint8_t* actions, float* obs, float* reward, uint8_t* done, int64_t* stats,
                       int32_t* bad_action, int32_t nthreads) {
  const int R = cfg->row_count, C = cfg->col_count, D = cfg->n_drones;
  const size_t RC = (size_t)R * C;
  const size_t obs_per_env = cfg->obs_mode == 0 ? RC : (size_t)D * cfg->window * cfg->window;
  int64_t s_rew = 0, s_done = 0, s_len = 0, s_ret = 0, s_cells = 0, s_anom = 0;
  int32_t bad = 0;
#ifdef _OPENMP
  if (nthreads <= 0) nthreads = omp_get_max_threads();
#pragma omp parallel for num_threads(nthreads) schedule(static) reduction(+ : s_rew, s_done, s_len, s_ret, s_cells, s_anom) reduction(| : bad)
#endif
  for (int32_t n = 0; n < N; ++n) {
    int8_t* g = grid_t + n * RC;
    double* p = pos + (size_t)n * D * 2;
    int32_t* m = meta + (size_t)n * 4;
    int16_t* w = wait ? wait + n * RC : NULL;
    int32_t cleared[2] = {0, 0};
    int32_t b = 0;
    const int was_remaining = m[META_REMAINING];
    uint8_t dn = 0;
    double r = step_env(cfg, g, p, m, w, actions + (size_t)n * D, &dn, cleared, &b);
    bad |= b;
    m[META_EPRET] += (int32_t)r;
    if (reward) reward[n] = (float)r;
    if (done) done[n] = dn;
    s_rew += (int64_t)r; s_cells += cleared[0]; s_anom += cleared[1];
    if (dn && was_remaining > 0) { s_done += 1; s_len += m[META_T]; s_ret += m[META_EPRET]; }
    if (dn && cfg->auto_reset)                         /* A2C/envs.py:22-26: obs below is the RESET obs */
      oracle_reset_env(cfg, cfg->env_id_base + n, m[META_EPISODE] + 1, g, p, m, w);
    if (obs) oracle_emit_obs(cfg, g, p, obs + (size_t)n * obs_per_env);
  }
  if (stats) {
    stats[0] += N; stats[1] += (int64_t)N * D; stats[2] += s_rew; stats[3] += s_done;
    stats[4] += s_len; stats[5] += s_ret; stats[6] += s_cells; stats[7] += s_anom;
  }
  if (bad_action) *bad_action |= bad;
}

/* mask NULL = all.  Every reset env moves to episode+1 (episode 0 is written by oracle_init_batch). */
void oracle_reset_batch(const OracleConfig* cfg, int32_t N, const uint8_t* mask, int8_t* grid_t, double* pos,
                        int32_t* meta, int16_t* wait, float* obs) {
  const size_t RC = (size_t)cfg->row_count * cfg->col_count;
  const int D = cfg->n_drones;
  const size_t obs_per_env = cfg->obs_mode == 0 ? RC : (size_t)D * cfg->window * cfg->window;
  for (int32_t n = 0; n < N; ++n) {
    if (!mask || mask[n])
      oracle_reset_env(cfg, cfg->env_id_base + n, meta[(size_t)n * 4 + META_EPISODE] + 1, grid_t + n * RC,
                       pos + (size_t)n * D * 2, meta + (size_t)n * 4, wait ? wait + n * RC : NULL);
    if (obs) oracle_emit_obs(cfg, grid_t + n * RC, pos + (size_t)n * D * 2, obs + (size_t)n * obs_per_env);
  }
}

void oracle_init_batch(const OracleConfig* cfg, int32_t N, int8_t* grid_t, double* pos, int32_t* meta, int16_t* wait) {
  const size_t RC = (size_t)cfg->row_count * cfg->col_count;
  const int D = cfg->n_drones;
  for (int32_t n = 0; n < N; ++n)
    oracle_reset_env(cfg, cfg->env_id_base + n, 0, grid_t + n * RC, pos + (size_t)n * D * 2, meta + (size_t)n * 4,
                     wait ? wait + n * RC : NULL);
}

void oracle_emit_obs_batch(const OracleConfig* cfg, int32_t N, const int8_t* grid_t, const double* pos, float* obs) {
  const size_t RC = (size_t)cfg->row_count * cfg->col_count;
  const int D = cfg->n_drones;
  const size_t obs_per_env = cfg->obs_mode == 0 ? RC : (size_t)D * cfg->window * cfg->window;
  for (int32_t n = 0; n < N; ++n) oracle_emit_obs(cfg, grid_t + n * RC, pos + (size_t)n * D * 2, obs + (size_t)n * obs_per_env);
}

/* Target choice of the 2D simulator's scripted controller, Drone.update (swarmsim-2d/drone.py:29-47), on the noVis
 * lattice: row-major scan (row == x outer, col == y inner), cells that are mapped are skipped (`grid == -1` there,
 * `grid == 0` here), the distance is `((dx)**2 + (dy)**2) ** 0.5` compared with a strict `min_dist > dist`, and the
 * QUIRK of `if not min_dist:` — Python truthiness: it fires when min_dist is None AND when the running minimum is
 * exactly 0.0, so after a cell at distance 0 the NEXT unmapped cell in scan order replaces it unconditionally.
 * Returns the flat target cell gy*R + gx, or -1 when nothing is left (the reference returns 1, drone.py:49-50).
 * Pinned by tests/golden/drone_update_targets.npz (frozen from the unmodified drone.py, tests/golden/make_donor_golden.py). */
static int drone_update_target(const OracleConfig* cfg, const int8_t* g, double x, double y) {
  const int R = cfg->row_count, C = cfg->col_count;
  int have = 0, target = -1;
  double min_dist = 0.0;
  for (int gx = 0; gx < R; ++gx)
    for (int gy = 0; gy < C; ++gy) {
      if (g[gy * R + gx] <= 0) continue;                                  /* drone.py:34-35 */
      const double dist = pow((x - gx) * (x - gx) + (y - gy) * (y - gy), 0.5);   /* drone.py:26-27 */
      if (!have || min_dist == 0.0) {                                     /* `if not min_dist` drone.py:39-42 */
        have = 1; min_dist = dist; target = gy * R + gx;
        continue;
      }
      if (min_dist > dist) { min_dist = dist; target = gy * R + gx; }     /* drone.py:43-45 */
    }
  return target;
}

/* Drone.update for every drone: target[n][i] (flat cell or -1) and the reference's 8-connected move direction
 * dir[n][i] = the net (dx, dy) of drone.py:52-59 (+-1 on EACH axis toward the target; (0, 0) when there is none). */
void oracle_drone_update(const OracleConfig* cfg, int32_t N, const int8_t* grid_t, const double* pos, int32_t* target, int8_t* dir) {
  const int R = cfg->row_count, C = cfg->col_count, D = cfg->n_drones;
  for (int32_t n = 0; n < N; ++n)
    for (int i = 0; i < D; ++i) {
      const double x = pos[((size_t)n * D + i) * 2], y = pos[((size_t)n * D + i) * 2 + 1];
      const int t = drone_update_target(cfg, grid_t + (size_t)n * R * C, x, y);
      int sx = 0, sy = 0;
      if (t >= 0) {
        const double tx = t % R, ty = t / R;
        /* drone.py:52-59, literally: four sequential, NON-exclusive ifs — a drone less than one unit above its target is
         * decremented and then incremented again (net 0), which a sign() restatement gets wrong */
        double nx = x, ny = y;
        if (ny > ty) ny -= 1;      /* center_x <-> column <-> noVis y */
        if (ny < ty) ny += 1;
        if (nx > tx) nx -= 1;      /* center_y <-> row <-> noVis x */
        if (nx < tx) nx += 1;
        sx = (int)(nx - x); sy = (int)(ny - y);
      }
      if (target) target[(size_t)n * D + i] = t;
      if (dir) { dir[((size_t)n * D + i) * 2] = (int8_t)sx; dir[((size_t)n * D + i) * 2 + 1] = (int8_t)sy; }
    }
}

/* Scripted greedy ACTION source on the noVis lattice — a 4-action VARIANT of Drone.update, not a restatement: the target
 * is the reference's (drone_update_target above, pinned), but the reference then moves +-1 on both axes at once, which
 * the noVis action set cannot express; here the drone takes the axis with the larger |delta| (x on ties):
 * x+ -> 1, x- -> 3, y+ -> 2, y- -> 0; action 0 when nothing is left or when it already sits on the target. */
void oracle_greedy_actions(const OracleConfig* cfg, int32_t N, const int8_t* grid_t, const double* pos, uint8_t* actions) {
  const int R = cfg->row_count, C = cfg->col_count, D = cfg->n_drones;
  for (int32_t n = 0; n < N; ++n)
    for (int i = 0; i < D; ++i) {
      const int8_t* g = grid_t + (size_t)n * R * C;
      double x = pos[((size_t)n * D + i) * 2], y = pos[((size_t)n * D + i) * 2 + 1];
      const int t = drone_update_target(cfg, g, x, y);
      uint8_t a = 0;
      if (t >= 0) {
        double dx = (t % R) - x, dy = (t / R) - y;
        if (dx == 0.0 && dy == 0.0) a = 0;
        else if (fabs(dx) >= fabs(dy)) a = dx > 0 ? 1 : 3;
        else a = dy > 0 ? 2 : 0;
      }
      actions[(size_t)n * D + i] = a;
    }
}

int oracle_max_threads(void) {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}
