// swarm_api.cu — the C ABI declared in include/swarm_b200.h (host side: validation, allocation, launches).
// There is NO CPU fallback: without an sm_100 device swarm_create returns SWARM_ERR_NO_DEVICE.
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <dlfcn.h>
#include <new>

#include "../../include/swarm_b200.h"
#include "swarm_kernels.cuh"
#include "swarm_a2c.cuh"

using namespace swarm;

struct SwarmEnv {
  SwarmConfig cfg;
  DevParams p;
  int ts;                 // tile size (lanes per env) of the staged warp kernel
  int staged;             // 1: swarm_warp.cuh kernel (grid staged in shared memory), 0: swarm_cta.cuh (one CTA per env)
  int wpb;                // warps per block of the staged kernel
  int dyn;                // staged kernel: dynamic CTA-group schedule (atomic counter) instead of static striding
  int resident[3];        // resident CTAs of the staged kernel on this device ([0] generic, [1] FAST, [2] LEAN instantiation); 0 = not yet queried
  int sm_count;
  long long launches;
  unsigned long long* stats_dev;   // int64[8] internal statistics
  // staging for swarm_step_host
  uint8_t* h_act_dev;
  void* h_obs_dev;
  float* h_rew_dev;
  uint8_t* h_done_dev;
  int* cover_dev;
  uint32_t* tables_dev;   // DevParams::tables
  // peer-memory statistics exchange
  unsigned long long* mbox;          // this rank's mailbox, int64 [SWARM_MAX_RANKS][SWARM_MBOX_SLOT]
  unsigned long long* gstats_dev;    // int64[8] scratch for swarm_stats_global with a host destination
  void* peer_open[SWARM_MAX_RANKS];  // mappings returned by cudaIpcOpenMemHandle (null for this rank / unused)
};

static thread_local char g_cuda_err[256] = "";

#define CUDA_TRY(expr)                                                                         \
  do {                                                                                         \
    cudaError_t _e = (expr);                                                                   \
    if (_e != cudaSuccess) {                                                                   \
      snprintf(g_cuda_err, sizeof(g_cuda_err), "%s: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__); \
      return SWARM_ERR_CUDA;                                                                   \
    }                                                                                          \
  } while (0)

extern "C" int swarm_abi_version(void) { return SWARM_ABI_VERSION; }

extern "C" const char* swarm_strerror(int s) {
  switch (s) {
    case SWARM_OK: return "ok";
    case SWARM_ERR_INVALID_ARG: return "invalid argument";
    case SWARM_ERR_CONFIG: return "unsupported SwarmConfig";
    case SWARM_ERR_CUDA: return "CUDA runtime error (see swarm_last_cuda_error)";
    case SWARM_ERR_NO_DEVICE: return "no sm_100 CUDA device: this library has no CPU fallback";
    case SWARM_ERR_ALLOC: return "allocation failed";
    case SWARM_ERR_NCCL: return "NCCL error";
    case SWARM_ERR_TIMEOUT: return "peer statistics mailboxes did not become consistent (a rank never pushed or died)";
    default: return "unknown status";
  }
}
extern "C" const char* swarm_last_cuda_error(void) { return g_cuda_err; }

extern "C" int swarm_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
  int ok = 0;
  for (int i = 0; i < n; ++i) {
    int major = 0;
    if (cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, i) == cudaSuccess && major == 10) ++ok;
  }
  return ok;
}

// ---- kernel selection ------------------------------------------------------------------------------------------------
constexpr int kWarpSmemMax = 20 * 1024;     // per-warp shared-memory slice the staged kernel may use

static int stage_bytes_for(const SwarmConfig* c, int epw) {
  if (c->obs_mode != SWARM_OBS_WINDOW) return 16;
  const long long b = c->obs_dtype == SWARM_DTYPE_P2 ? (long long)epw * c->n_drones * 8   // one packed word per drone
                                                      : (long long)epw * c->n_drones * c->window * c->window * (c->obs_dtype == SWARM_DTYPE_F32 ? 4 : 1);
  return (int)((b + 15) / 16 * 16);
}

static int align16(long long v) { return (int)((v + 15) / 16 * 16); }

// Per-warp shared-memory needs of the staged kernel for tile size ts: obs staging and one input stage.
static void warp_bytes(const SwarmConfig* c, const DevParams* p, int ts, int* stage, int* in_w) {
  const int epw = 32 / ts;
  *stage = stage_bytes_for(c, epw);
  *in_w = epw * (p->env_words * 4 + 32 + 16) + align16((long long)epw * c->n_drones * 8) + align16((long long)epw * c->n_drones);
}

// Smallest power-of-two tile >= D (min 4), widened while the batch would give too few warps to fill the GPU.
// Returns 0 when the shape does not fit the staged kernel (D > 32, PLANE-mode anomalies, or a per-warp slice too large for a
// two-stage ring).
static int pick_tile(const SwarmConfig* c, int sms, DevParams* p) {
  if (c->n_drones > 32 || !p->list_mode) return 0;
  int ts = 4;
  while (ts < c->n_drones) ts <<= 1;
  while (ts < 32 && (long long)c->n_envs / (32 / ts) < (long long)sms * 16) ts <<= 1;
  for (; ts <= 32; ts <<= 1) {
    int stage = 0, in_w = 0;
    warp_bytes(c, p, ts, &stage, &in_w);
    if ((long long)stage + 2ll * in_w + 16 <= kWarpSmemMax) return ts;
  }
  return 0;
}

static int use_device(const SwarmEnv* e) {
  int cur = -1;
  CUDA_TRY(cudaGetDevice(&cur));
  if (cur != e->cfg.device) CUDA_TRY(cudaSetDevice(e->cfg.device));
  return SWARM_OK;
}

static int grid_blocks(const SwarmEnv* e, long long warps_needed) {
  long long b = (warps_needed + 7) / 8;
  long long cap = (long long)e->sm_count * 8;
  if (b > cap) b = cap;
  if (b < 1) b = 1;
  return (int)b;
}

extern "C" int swarm_create(const SwarmConfig* c, SwarmEnv** out) {
  if (!c || !out) return SWARM_ERR_INVALID_ARG;
  *out = nullptr;
  if (c->struct_size != (int32_t)sizeof(SwarmConfig)) return SWARM_ERR_INVALID_ARG;
  const long long RC = (long long)c->row_count * c->col_count;
  if (c->n_envs < 1 || c->n_drones < 1 || c->n_drones > 1024 || c->row_count < 1 || c->col_count < 1 || RC > (1ll << 24))
    return SWARM_ERR_CONFIG;
  if (c->uncertainity < 2 || c->uncertainity > 127) return SWARM_ERR_CONFIG;
  if (c->n_anamolous < 0 || c->n_anamolous > RC) return SWARM_ERR_CONFIG;
  if (c->reward_mode != SWARM_REWARD_A2C && c->reward_mode != SWARM_REWARD_DQL) return SWARM_ERR_CONFIG;
  if (c->obs_mode != SWARM_OBS_FULL && c->obs_mode != SWARM_OBS_WINDOW) return SWARM_ERR_CONFIG;
  if (c->obs_mode == SWARM_OBS_WINDOW && (c->window < 1 || c->window > 31 || (c->window & 1) == 0)) return SWARM_ERR_CONFIG;
  if (c->obs_dtype != SWARM_DTYPE_F32 && c->obs_dtype != SWARM_DTYPE_U8 && c->obs_dtype != SWARM_DTYPE_P2) return SWARM_ERR_CONFIG;
  if (c->obs_dtype == SWARM_DTYPE_P2 && (c->obs_mode != SWARM_OBS_WINDOW || c->window != 5)) return SWARM_ERR_CONFIG;   // packed: 5x5 windows only
  if (!(c->collision_dist >= 0.0) || c->collision_dist > 64.0 || !(c->step_size > 0.0)) return SWARM_ERR_CONFIG;
  if (c->dwell_normal < 0 || c->dwell_anom < 0 || c->dwell_normal > 32767 || c->dwell_anom > 32767) return SWARM_ERR_CONFIG;
  if (RC / c->step_size > 2.0e9) return SWARM_ERR_CONFIG;
  {   // positions are float32 multiples of step_size clipped to [0, R-1] x [0, C-1]: exact (== the reference's float64,
      // A2C/noVis_sim.py:121-135) iff step_size = m * 2^-k with (max coordinate / 2^-k) < 2^24.  Anything else is rejected
      // rather than silently drifting across the `dist < collision_dist` knife edge.
    int ex = 0;
    const double mant = std::frexp(c->step_size, &ex);            // step_size = mant * 2^ex, mant in [0.5, 1)
    int bits = 0;
    double m = mant;
    while (m != std::floor(m) && bits < 60) { m *= 2.0; ++bits; }  // bits = number of mantissa bits used
    const int k = bits - ex;                                       // step_size is a multiple of 2^-k
    const double maxc = (double)((c->row_count > c->col_count ? c->row_count : c->col_count) - 1) + c->step_size;
    if (m != std::floor(m) || (float)c->step_size != c->step_size || (k > 0 && std::ldexp(maxc, k) >= 16777216.0) || maxc >= 16777216.0)
      return SWARM_ERR_CONFIG;
  }

  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) { cudaGetLastError(); return SWARM_ERR_NO_DEVICE; }
  if (c->device < 0 || c->device >= ndev) return SWARM_ERR_INVALID_ARG;
  int major = 0, sms = 0;
  CUDA_TRY(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, c->device));
  if (major != 10) return SWARM_ERR_NO_DEVICE;   // kernels are sm_100a only
  CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, c->device));
  CUDA_TRY(cudaSetDevice(c->device));

  SwarmEnv* e = new (std::nothrow) SwarmEnv();
  if (!e) return SWARM_ERR_ALLOC;
  memset(e, 0, sizeof(*e));
  e->cfg = *c;
  e->sm_count = sms;
  DevParams& p = e->p;
  p.N = c->n_envs; p.D = c->n_drones; p.R = c->row_count; p.C = c->col_count; p.RC = (int)RC;
  p.U = c->uncertainity; p.n_anom = c->n_anamolous;
  p.reward_mode = c->reward_mode; p.obs_mode = c->obs_mode; p.W = c->obs_mode == SWARM_OBS_WINDOW ? c->window : 1;
  p.obs_dtype = c->obs_dtype; p.auto_reset = c->auto_reset;
  p.dwell_normal = c->dwell_normal; p.dwell_anom = c->dwell_anom;
  p.has_wait = (c->dwell_normal | c->dwell_anom) != 0;
  p.punish_thr = (int)std::floor((double)RC / c->step_size);
  p.cd_fast = c->collision_dist <= 0.5;
  p.scan_m = (int)std::ceil(c->collision_dist) + 1;
  p.env_words = (int)(((RC + 31) / 32 + 3) / 4 * 4);
  p.list_mode = c->n_anamolous <= 8 && c->row_count <= 256 && c->col_count <= 256;
  {   // float32-exact coordinates: step_size = m 2^-k (validated above) with k <= 12 (squares of offsets <= 1/2 exact) and
      // (largest coordinate in units of 2^-k)^2 * 2 <= 2^24 (squared pair distances exact)
    int ex = 0, bits = 0;
    double m = std::frexp(c->step_size, &ex);
    while (m != std::floor(m) && bits < 60) { m *= 2.0; ++bits; }
    const int k = bits - ex > 0 ? bits - ex : 0;
    const double M = std::ldexp((double)((c->row_count > c->col_count ? c->row_count : c->col_count) - 1) + c->step_size, k);
    p.f32_exact = k <= 12 && M <= 2896.0;
  }
  p.obs_per_env = c->obs_mode == SWARM_OBS_FULL ? (int)RC : (c->obs_dtype == SWARM_DTYPE_P2 ? c->n_drones : c->n_drones * c->window * c->window);
  p.ss = (float)c->step_size; p.xmax = (float)(c->row_count - 1); p.ymax = (float)(c->col_count - 1);
  p.cd = c->collision_dist;
  p.spread_thr = (double)(c->row_count / c->n_drones);
  {   // exact squared-distance thresholds: sqrt() is monotone, so  sqrt(z) < cd  <=>  z < cd2_thr  for every double z
    double z = c->collision_dist * c->collision_dist;
    while (z > 0.0 && std::sqrt(z) >= c->collision_dist) z = std::nextafter(z, -INFINITY);
    while (std::sqrt(z) < c->collision_dist) z = std::nextafter(z, INFINITY);
    p.cd2_thr = z;                                   // smallest z with sqrt(z) >= cd
    z = p.spread_thr * p.spread_thr;
    while (std::sqrt(z) <= p.spread_thr) z = std::nextafter(z, INFINITY);
    while (z > 0.0 && std::sqrt(z) > p.spread_thr) z = std::nextafter(z, -INFINITY);
    p.spread2_thr = z;                               // largest z with sqrt(z) <= row_count // n_drones
    // float versions for exactly-representable d2: d2 < cd2_thr <=> d2 < (smallest float >= cd2_thr), and
    // d2 > spread2_thr <=> d2 > (largest float <= spread2_thr)
    float f = (float)p.cd2_thr;
    if ((double)f < p.cd2_thr) f = std::nextafterf(f, INFINITY);
    p.cd2_thr_f = f;
    f = (float)p.spread2_thr;
    if ((double)f > p.spread2_thr) f = std::nextafterf(f, -INFINITY);
    p.spread2_thr_f = f;
  }
  e->ts = pick_tile(c, sms, &p);
  e->staged = e->ts != 0;
  if (!e->staged && c->obs_dtype == SWARM_DTYPE_P2) {   // the packed format is emitted by the staged warp kernel only
    delete e;
    return SWARM_ERR_CONFIG;
  }
  if (!e->staged) e->ts = 32;
  if (e->staged) {
    // CTA shape and ring depth: the (warps per CTA w, stages S) pair with the cheapest schedule.  Model: an SM keeps `ctas`
    // CTAs resident (shared memory: 227 KB; registers: 72 per thread under the kernel's launch bounds); the
    // ncg = ceil(N / (w * EPW)) CTA-groups are dealt round-robin over SMs x ctas slots, so an SM works through
    // rounds = ceil(ceil(ncg / SMs) / ctas) rounds of min(ctas, ceil(ncg / SMs)) * w warps each, a round costing a latency
    // floor plus the warps it holds.  Ties: deeper ring, then more busy SMs, then the larger CTA.
    // cfg3 (65 536 envs x 16 drones): w = 14, S = 3, 2 CTAs per SM -> 2341 CTA-groups over 296 slots = 8 rounds at 98.8 %
    // (round 1's fixed 8-warp CTAs: 4096 over 444 = 9.22 -> 10 rounds); 8192 envs (cfg3 sharded over 8 GPUs) and cfg4 become
    // ONE round.  SWARM_WPB / SWARM_STAGES override.
    int stage = 0, in_w = 0;
    warp_bytes(c, &p, e->ts, &stage, &in_w);
    p.stage_bytes = stage;
    const int epw = 32 / e->ts;
    int smax = 4, wforce = 0;
    if (const char* v = getenv("SWARM_STAGES")) { const int x = atoi(v); if (x >= 2 && x <= MAX_STAGES) smax = x; }
    if (const char* v = getenv("SWARM_WPB")) { const int x = atoi(v); if (x >= 1 && x <= SWARM_MAXT / 32) wforce = x; }
    const long long stat = 7 * 1024 + 1024;   // static shared memory of the kernel (LUT, pad tables, barriers) + the driver's 1 KB
    const int kRegs = 72;
    double best_cost = 1e30;
    int best_s = 0, best_busy = 0;
    e->wpb = 8;
    for (int w = SWARM_MAXT / 32; w >= 1; --w) {
      if (wforce && w != wforce) continue;
      for (int S = smax; S >= 2; --S) {
        const long long smem = (long long)w * stage + (long long)S * w * in_w + 64 + stat;
        if (smem > 227 * 1024) continue;
        long long ctas = (227 * 1024) / smem;
        const long long by_regs = 65536 / ((long long)w * 32 * kRegs), by_warps = 64 / w;
        if (ctas > by_regs) ctas = by_regs;
        if (ctas > by_warps) ctas = by_warps;
        if (ctas > 32) ctas = 32;
        if (ctas < 1) continue;
        const long long ncg = ((long long)p.N + (long long)w * epw - 1) / ((long long)w * epw);
        const long long per_sm = (ncg + sms - 1) / sms;
        const long long rounds = (per_sm + ctas - 1) / ctas;
        const long long act = per_sm < ctas ? per_sm : ctas;
        const double cost = (double)rounds * (8.0 + (double)(act * w));
        const int busy = (int)(ncg < sms ? ncg : sms);
        if (cost < best_cost - 1e-9 || (cost < best_cost + 1e-9 && (S > best_s || (S == best_s && busy > best_busy)))) {
          best_cost = cost; best_s = S; best_busy = busy; e->wpb = w;
        }
      }
    }
    p.n_stages = best_s > 0 ? best_s : 2;
    const long long epc = (long long)e->wpb * epw;
    const int covb = (int)(epc * p.env_words * 4), listb = (int)(epc * 32), posb = align16(epc * c->n_drones * 8), metab = (int)(epc * 16);
    const int actb = align16(epc * c->n_drones);
    p.off_alist = covb;
    p.off_pos = covb + listb;
    p.off_meta = p.off_pos + posb;
    p.off_act = p.off_meta + metab;
    p.in_bytes = p.off_act + actb;
    p.warp_smem = stage + p.n_stages * in_w;
    p.cta_smem = e->wpb * p.stage_bytes + p.n_stages * p.in_bytes + 64;
    e->dyn = 0;
    if (const char* v = getenv("SWARM_DYN")) e->dyn = atoi(v) != 0;
  }
  // ceil(2^32 / nq) for the (env, quad) split of emit_full_smem; nq == 1 would need 2^32, which does not fit: the kernel
  // takes e = v directly in that case (nq_magic == 0 is the marker)
  p.nq_magic = RC >= 8 ? (uint32_t)(((1ull << 32) + (uint64_t)(RC / 4) - 1) / (uint64_t)(RC / 4)) : 0u;
  p.seed_lo = (uint32_t)c->seed; p.seed_hi = (uint32_t)(c->seed >> 32);
  p.env_id_base = c->env_id_base;

  const size_t N = (size_t)p.N;
  cudaError_t err = cudaSuccess;
  auto alloc = [&](void** ptr, size_t bytes) { if (err == cudaSuccess) err = cudaMalloc(ptr, bytes ? bytes : 16); };
  alloc((void**)&p.cov, N * p.env_words * sizeof(uint32_t));
  if (p.list_mode) alloc((void**)&p.alist, N * 8 * sizeof(uint32_t));
  else alloc((void**)&p.anomp, N * p.env_words * sizeof(uint32_t));
  alloc((void**)&p.sched, 16);
  alloc((void**)&e->tables_dev, 1536 * sizeof(uint32_t));
  alloc((void**)&p.pos, N * p.D * sizeof(float2));
  alloc((void**)&p.meta, N * sizeof(int4));
  if (p.has_wait) alloc((void**)&p.wait, N * RC * sizeof(int16_t));
  alloc((void**)&p.err, 16);
  alloc((void**)&e->stats_dev, 8 * sizeof(unsigned long long));
  alloc((void**)&e->cover_dev, N * sizeof(int));
  if (err != cudaSuccess) {
    snprintf(g_cuda_err, sizeof(g_cuda_err), "cudaMalloc: %s", cudaGetErrorString(err));
    swarm_destroy(e);
    return SWARM_ERR_ALLOC;
  }
  cudaMemsetAsync(p.err, 0, 16, 0);
  cudaMemsetAsync(p.sched, 0, 16, 0);
  cudaMemsetAsync(e->stats_dev, 0, 64, 0);
  cudaMemsetAsync(p.meta, 0, N * sizeof(int4), 0);
  p.tables = e->tables_dev;
  if (p.obs_dtype == SWARM_DTYPE_F32) build_tables_kernel<float><<<1, 256, 0, 0>>>(p, e->tables_dev);
  else if (p.obs_dtype == SWARM_DTYPE_U8) build_tables_kernel<uint8_t><<<1, 256, 0, 0>>>(p, e->tables_dev);   // (packed: warp kernel only, no tables)
  reset_kernel<<<grid_blocks(e, p.N), 256, 0, 0>>>(p, nullptr, 0);
  e->launches++;
  err = cudaStreamSynchronize(0);
  if (err == cudaSuccess) err = cudaGetLastError();
  if (err != cudaSuccess) {
    snprintf(g_cuda_err, sizeof(g_cuda_err), "create/reset_kernel: %s", cudaGetErrorString(err));
    swarm_destroy(e);
    return SWARM_ERR_CUDA;
  }
  *out = e;
  return SWARM_OK;
}

extern "C" int swarm_destroy(SwarmEnv* e) {
  if (!e) return SWARM_ERR_INVALID_ARG;
  // Nothing of this handle may still run when its memory goes away; peers that push into this rank's mailbox must have been
  // quiesced by the caller (swarm_stats_peers_close on every rank + a barrier; BatchedDroneEnv.close does both when a process
  // group exists) — freeing IPC-exported memory that an importer still has open is undefined.
  int cur = -1;
  if (cudaGetDevice(&cur) == cudaSuccess && cur != e->cfg.device) cudaSetDevice(e->cfg.device);
  cudaDeviceSynchronize();
  cudaFree(e->p.cov); cudaFree(e->p.anomp); cudaFree(e->p.alist); cudaFree(e->p.sched); cudaFree(e->p.pos); cudaFree(e->p.meta); cudaFree(e->p.wait); cudaFree(e->p.err);
  cudaFree(e->stats_dev); cudaFree(e->cover_dev); cudaFree(e->tables_dev);
  cudaFree(e->h_act_dev); cudaFree(e->h_obs_dev); cudaFree(e->h_rew_dev); cudaFree(e->h_done_dev);
  for (int r = 0; r < SWARM_MAX_RANKS; ++r)
    if (e->peer_open[r]) cudaIpcCloseMemHandle(e->peer_open[r]);
  cudaFree(e->mbox); cudaFree(e->p.peers.pushes); cudaFree(e->gstats_dev);
  delete e;
  return SWARM_OK;
}

extern "C" int swarm_get_config(const SwarmEnv* e, SwarmConfig* out) {
  if (!e || !out) return SWARM_ERR_INVALID_ARG;
  *out = e->cfg;
  return SWARM_OK;
}

extern "C" int64_t swarm_obs_elems(const SwarmEnv* e) { return e ? (int64_t)e->p.N * e->p.obs_per_env : 0; }
extern "C" int64_t swarm_obs_bytes(const SwarmEnv* e) {
  return e ? swarm_obs_elems(e) * (e->p.obs_dtype == SWARM_DTYPE_F32 ? 4 : (e->p.obs_dtype == SWARM_DTYPE_P2 ? 8 : 1)) : 0;
}

extern "C" int swarm_state_ptr(SwarmEnv* e, int which, void** ptr_out, int64_t* bytes_out) {
  if (!e || !ptr_out) return SWARM_ERR_INVALID_ARG;
  const size_t N = (size_t)e->p.N;
  void* ptr = nullptr;
  int64_t b = 0;
  switch (which) {
    case 0: ptr = e->p.pos; b = (int64_t)(N * e->p.D * sizeof(float2)); break;
    case 1: ptr = e->p.meta; b = (int64_t)(N * sizeof(int4)); break;
    case 2: ptr = e->p.cov; b = (int64_t)(N * e->p.env_words * 4); break;
    case 3: ptr = e->p.err; b = 4; break;
    case 4: ptr = e->stats_dev; b = 64; break;
    default: return SWARM_ERR_INVALID_ARG;
  }
  *ptr_out = ptr;
  if (bytes_out) *bytes_out = b;
  return SWARM_OK;
}

// ------------------------------------------------------------------------------------------------ launches
// Step launches are programmatic dependent launches (cudaLaunchAttributeProgrammaticStreamSerialization): the kernels call
// griddepcontrol.launch_dependents when their main loop is done and griddepcontrol.wait before their first global access, so
// the launch latency, CTA start-up and shared-memory prologue of step k+1 overlap the epilogue and drain of whatever kernel
// precedes it on the stream (under stream capture the edge becomes a programmatic graph dependency).  Measured per shape in
// DESIGN.md §5 (triggering at the kernel's TOP instead lets waiting CTAs crowd the running grid and loses).
// SWARM_PDL=0 launches plain kernels (A/B measurements).
static bool pdl_enabled() {
  static const bool on = [] {
    const char* s = getenv("SWARM_PDL");
    return !(s && s[0] == '0');
  }();
  return on;
}

template <typename... KArgs, typename... Args>
static cudaError_t launch_step_kernel(void (*kernel)(KArgs...), int blocks, int threads, int smem, cudaStream_t st, Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)blocks);
  cfg.blockDim = dim3((unsigned)threads);
  cfg.dynamicSmemBytes = (size_t)smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

// One instantiation of the staged kernel: raises its dynamic shared-memory limit once per handle, sizes the persistent
// grid from the occupancy calculator (resident CTAs only) and launches.
template <int TS, typename OT, int MODE, int WC, bool FAST, bool LEAN = false>
static cudaError_t launch_warp_inst(SwarmEnv* e, const StepIO& io, cudaStream_t st, long long blocks_needed, int smem) {
  auto kernel = step_warp_kernel<TS, OT, MODE, WC, FAST, LEAN>;
  int& resident = e->resident[LEAN ? 2 : (FAST ? 1 : 0)];
  // The dynamic shared-memory limit is a property of the FUNCTION (per context), not of a handle: two live envs of different
  // shapes share an instantiation, so the limit is only ever raised (per device).
  static int smem_limit[64] = {0};
  int& limit = smem_limit[e->cfg.device & 63];
  if (smem > limit) {
    cudaError_t err = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (err != cudaSuccess) return err;
    limit = smem;
  }
  if (resident == 0) {
    cudaError_t err;
    int per_sm = 0;
    err = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, e->wpb * 32, smem);
    if (err != cudaSuccess) return err;
    resident = (per_sm > 0 ? per_sm : 1) * e->sm_count;
  }
  const long long blocks = blocks_needed < resident ? blocks_needed : resident;
  return launch_step_kernel(kernel, (int)blocks, e->wpb * 32, smem, st, e->p, io);
}

template <int TS, typename OT>
static cudaError_t launch_warp_ts(SwarmEnv* e, const StepIO& io, cudaStream_t st, long long blocks, int smem) {
  const DevParams& p = e->p;
  // FAST: every stage in one pass, cd <= 0.5, no dwell — the swarm_step / swarm_rollout production path (any action dtype)
  const bool fast = (io.flags & F_ALL) == F_ALL && p.cd_fast && !p.has_wait && p.f32_exact && (io.act_dtype == SWARM_ACT_U8 || io.T == 1);
  // LEAN: the FAST kernel with the tail / rollout / alignment bookkeeping folded away (see step_warp_kernel)
  const int epw = 32 / TS;
  const bool lean = fast && io.act_dtype == SWARM_ACT_U8 && io.T == 1 && !io.obs_last_only && io.obs_out && io.stats && p.D == TS && p.N % epw == 0 && (epw * TS) % 16 == 0 &&
                    ((((uintptr_t)io.actions) | ((uintptr_t)io.obs_out)) & 15) == 0 && !getenv("SWARM_NO_LEAN");
  if (p.obs_mode == SWARM_OBS_FULL) {
    if (lean) return launch_warp_inst<TS, OT, 0, 0, true, true>(e, io, st, blocks, smem);
    return fast ? launch_warp_inst<TS, OT, 0, 0, true>(e, io, st, blocks, smem) : launch_warp_inst<TS, OT, 0, 0, false>(e, io, st, blocks, smem);
  }
  if (p.W == 5) {
    if (lean) return launch_warp_inst<TS, OT, 1, 5, true, true>(e, io, st, blocks, smem);
    return fast ? launch_warp_inst<TS, OT, 1, 5, true>(e, io, st, blocks, smem) : launch_warp_inst<TS, OT, 1, 5, false>(e, io, st, blocks, smem);
  }
  return fast ? launch_warp_inst<TS, OT, 1, 0, true>(e, io, st, blocks, smem) : launch_warp_inst<TS, OT, 1, 0, false>(e, io, st, blocks, smem);
}

template <typename OT>
static cudaError_t launch_step_ot(SwarmEnv* e, const StepIO& io_in, cudaStream_t st) {
  const DevParams& p = e->p;
  StepIO io = io_in;
  io.dyn = e->staged ? e->dyn : 0;
  cudaError_t err = cudaSuccess;
  if (e->staged) {
    const int epw = 32 / e->ts;
    const long long ngroups = ((long long)p.N + epw - 1) / epw;
    const long long blocks = (ngroups + e->wpb - 1) / e->wpb;
    const int smem = p.cta_smem;
    switch (e->ts) {
      case 4: err = launch_warp_ts<4, OT>(e, io, st, blocks, smem); break;
      case 8: err = launch_warp_ts<8, OT>(e, io, st, blocks, smem); break;
      case 16: err = launch_warp_ts<16, OT>(e, io, st, blocks, smem); break;
      default: err = launch_warp_ts<32, OT>(e, io, st, blocks, smem); break;
    }
  } else {
    // one CTA per env; the env's plane(s) are staged in shared memory when they fit (160 KB), plus — WINDOW obs of W == 5
    // in LIST mode — a staging image of the env's [D][25] obs block (staged == 2).  WINDOW obs: CTAs of one thread per drone
    // (64 threads for cfg5b, ~7 envs resident per SM); FULL obs: 256 threads stream the 256 KiB of a 256 x 256 env.
    const long long planes = (long long)p.env_words * 4 * (p.list_mode ? 1 : 2);
    const long long image = (p.list_mode && p.obs_mode == SWARM_OBS_WINDOW && p.W == 5 && p.D <= 255)
                                ? (long long)p.D * 25 * (p.obs_dtype == SWARM_DTYPE_F32 ? 4 : 1) : 0;
    int staged = planes + 64 <= 160 * 1024 ? 1 : 0;
    long long need = staged ? planes + 64 : 0;
    if (staged && image && planes + 64 + image + 16 <= 200 * 1024) { staged = 2; need = planes + 64 + image + 16; }
    const int smem = (int)need;
    int threads = 256;
    if (p.obs_mode == SWARM_OBS_WINDOW) {
      threads = (p.D + 31) / 32 * 32;
      if (threads < 64) threads = 64;
      if (threads > 256) threads = 256;
    }
    static int cta_limit[64] = {0};   // per-function limit, only raised (see launch_warp_inst)
    int& limit = cta_limit[e->cfg.device & 63];
    if (smem > limit) {
      if ((err = cudaFuncSetAttribute(step_cta_kernel<OT>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem)) != cudaSuccess) return err;
      limit = smem;
    }
    if (!e->resident[0]) {
      int per_sm = 0;
      if ((err = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, step_cta_kernel<OT>, threads, smem)) != cudaSuccess) return err;
      e->resident[0] = (per_sm > 0 ? per_sm : 1) * e->sm_count;
    }
    long long blocks = p.N;
    if (blocks > e->resident[0]) blocks = e->resident[0];   // resident CTAs only, grid-stride beyond
    err = launch_step_kernel(step_cta_kernel<OT>, (int)blocks, threads, smem, st, p, io, staged);
  }
  e->launches++;
  return err;
}

// Packed 2-bit windows (SWARM_DTYPE_P2): the staged warp kernel with W == 5 only, one 64-bit word per drone.
template <int TS>
static cudaError_t launch_p2_ts(SwarmEnv* e, const StepIO& io, cudaStream_t st, long long blocks, int smem) {
  const DevParams& p = e->p;
  const bool fast = (io.flags & F_ALL) == F_ALL && p.cd_fast && !p.has_wait && p.f32_exact && (io.act_dtype == SWARM_ACT_U8 || io.T == 1);
  const int epw = 32 / TS;
  const bool lean = fast && io.act_dtype == SWARM_ACT_U8 && io.T == 1 && !io.obs_last_only && io.obs_out && io.stats && p.D == TS && p.N % epw == 0 && (epw * TS) % 16 == 0 &&
                    ((((uintptr_t)io.actions) | ((uintptr_t)io.obs_out)) & 15) == 0 && !getenv("SWARM_NO_LEAN");
  if (lean) return launch_warp_inst<TS, Packed2, 1, 5, true, true>(e, io, st, blocks, smem);
  return fast ? launch_warp_inst<TS, Packed2, 1, 5, true>(e, io, st, blocks, smem) : launch_warp_inst<TS, Packed2, 1, 5, false>(e, io, st, blocks, smem);
}

static cudaError_t launch_step_p2(SwarmEnv* e, const StepIO& io_in, cudaStream_t st) {
  const DevParams& p = e->p;
  StepIO io = io_in;
  io.dyn = e->dyn;
  const int epw = 32 / e->ts;
  const long long ngroups = ((long long)p.N + epw - 1) / epw;
  const long long blocks = (ngroups + e->wpb - 1) / e->wpb;
  cudaError_t err;
  switch (e->ts) {
    case 4: err = launch_p2_ts<4>(e, io, st, blocks, p.cta_smem); break;
    case 8: err = launch_p2_ts<8>(e, io, st, blocks, p.cta_smem); break;
    case 16: err = launch_p2_ts<16>(e, io, st, blocks, p.cta_smem); break;
    default: err = launch_p2_ts<32>(e, io, st, blocks, p.cta_smem); break;
  }
  e->launches++;
  return err;
}

static int launch_step(SwarmEnv* e, const StepIO& io, cudaStream_t st) {
  if (e->p.obs_dtype == SWARM_DTYPE_P2) CUDA_TRY(launch_step_p2(e, io, st));
  else if (e->p.obs_dtype == SWARM_DTYPE_F32) CUDA_TRY(launch_step_ot<float>(e, io, st));
  else CUDA_TRY(launch_step_ot<uint8_t>(e, io, st));
  CUDA_TRY(cudaGetLastError());
  return SWARM_OK;
}

static StepIO make_io(int flags) {
  StepIO io;
  memset(&io, 0, sizeof(io));
  io.T = 1;
  io.flags = flags;
  return io;
}

extern "C" int swarm_reset(SwarmEnv* e, const uint8_t* mask, void* obs_out, void* stream) {
  if (!e) return SWARM_ERR_INVALID_ARG;
  int rc = use_device(e);
  if (rc) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  reset_kernel<<<grid_blocks(e, e->p.N), 256, 0, st>>>(e->p, mask, 1);
  e->launches++;
  CUDA_TRY(cudaGetLastError());
  if (obs_out) {
    StepIO io = make_io(F_OBS);
    io.obs_out = obs_out;
    return launch_step(e, io, st);
  }
  return SWARM_OK;
}

extern "C" int swarm_set_state(SwarmEnv* e, const int8_t* grid, const float* pos, const int32_t* t, const int16_t* wait,
                               int32_t first, int32_t count, void* stream) {
  if (!e || first < 0 || count < 0 || (long long)first + count > e->p.N) return SWARM_ERR_INVALID_ARG;
  if (count == 0) return SWARM_OK;
  int rc = use_device(e);
  if (rc) return rc;
  set_state_kernel<<<grid_blocks(e, count), 256, 0, (cudaStream_t)stream>>>(e->p, grid, pos, t, wait, first, count);
  e->launches++;
  CUDA_TRY(cudaGetLastError());
  return SWARM_OK;
}

extern "C" int swarm_get_state(SwarmEnv* e, int8_t* grid, float* pos, int32_t* t, int16_t* wait, int32_t first, int32_t count,
                               void* stream) {
  if (!e || first < 0 || count < 0 || (long long)first + count > e->p.N) return SWARM_ERR_INVALID_ARG;
  if (count == 0) return SWARM_OK;
  int rc = use_device(e);
  if (rc) return rc;
  get_state_kernel<<<grid_blocks(e, count), 256, 0, (cudaStream_t)stream>>>(e->p, grid, pos, t, wait, first, count);
  e->launches++;
  CUDA_TRY(cudaGetLastError());
  return SWARM_OK;
}

extern "C" int swarm_step(SwarmEnv* e, const void* actions, int action_dtype, void* obs_out, float* reward_out,
                          uint8_t* done_out, int64_t* stats_accum, void* stream) {
  if (!e || !actions || !reward_out || !done_out || action_dtype < 0 || action_dtype > 2) return SWARM_ERR_INVALID_ARG;
  int rc = use_device(e);
  if (rc) return rc;
  StepIO io = make_io(F_ALL | (e->p.auto_reset ? F_RESET : 0) | F_PUSH);
  io.actions = actions; io.act_dtype = action_dtype; io.obs_out = obs_out;
  io.reward_out = reward_out; io.done_out = done_out; io.stats = (unsigned long long*)stats_accum;
  return launch_step(e, io, (cudaStream_t)stream);
}

extern "C" int swarm_rollout(SwarmEnv* e, int32_t T, const uint8_t* actions, void* obs_out, int32_t obs_last_only,
                             float* reward_out, uint8_t* done_out, int64_t* stats_accum, void* stream) {
  if (!e || T < 1 || !actions || !reward_out || !done_out) return SWARM_ERR_INVALID_ARG;
  int rc = use_device(e);
  if (rc) return rc;
  StepIO io = make_io(F_ALL | (e->p.auto_reset ? F_RESET : 0) | F_PUSH);
  io.T = T;
  io.actions = actions; io.act_dtype = SWARM_ACT_U8; io.obs_out = obs_out; io.obs_last_only = obs_last_only ? 1 : 0;
  io.reward_out = reward_out; io.done_out = done_out; io.stats = (unsigned long long*)stats_accum;
  return launch_step(e, io, (cudaStream_t)stream);
}

// A PINNED host buffer (cudaHostAlloc / cudaHostRegister, torch's pin_memory) is addressable from the device under UVA.
// Returns its device-side address, or null when the buffer is pageable.  Asked on every call (~1 us): a cached answer would go
// stale when the caller frees the buffer.
static void* host_mapped_ptr(const void* host) {
  if (!host) return nullptr;
  cudaPointerAttributes a;
  if (cudaPointerGetAttributes(&a, host) != cudaSuccess) { cudaGetLastError(); return nullptr; }
  return (a.type == cudaMemoryTypeHost) ? a.devicePointer : nullptr;
}

// swarm_step_host / swarm_step_host_pos.  Two regimes:
//  - large batches: H2D actions, step, D2H results through device staging — the copy engines move PCIe traffic faster than
//    SM-issued stores do (measured);
//  - SMALL batches (actions + obs + reward + done <= 64 KiB, e.g. the single-env facade the reference's scripts run on): the
//    call is all fixed cost, so pinned host buffers are handed to the kernel itself — it reads the actions and writes obs /
//    reward / done over PCIe — and the four copies with their launch latencies disappear (pageable buffers still stage).
static int step_host_impl(SwarmEnv* e, const uint8_t* actions_host, void* obs_host, float* reward_host, uint8_t* done_host,
                          float* pos_host, int64_t* stats_host, void* stream) {
  if (!e || !actions_host || !reward_host || !done_host) return SWARM_ERR_INVALID_ARG;
  int rc = use_device(e);
  if (rc) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  const size_t N = (size_t)e->p.N, nact = N * e->p.D, obs_bytes = obs_host ? (size_t)swarm_obs_bytes(e) : 0;
  static const bool zc_off = getenv("SWARM_NO_ZEROCOPY") != nullptr;
  const bool small = !zc_off && nact + obs_bytes + 5 * N <= (64u << 10);
  const uint8_t* act_zc = small ? (const uint8_t*)host_mapped_ptr(actions_host) : nullptr;
  void* obs_zc = small ? host_mapped_ptr(obs_host) : nullptr;
  float* rew_zc = small ? (float*)host_mapped_ptr(reward_host) : nullptr;
  uint8_t* done_zc = small ? (uint8_t*)host_mapped_ptr(done_host) : nullptr;
  if (!e->h_act_dev && !(act_zc && rew_zc && done_zc)) {
    CUDA_TRY(cudaMalloc((void**)&e->h_act_dev, nact));
    CUDA_TRY(cudaMalloc((void**)&e->h_rew_dev, N * sizeof(float)));
    CUDA_TRY(cudaMalloc((void**)&e->h_done_dev, N));
  }
  if (obs_host && !obs_zc && !e->h_obs_dev) CUDA_TRY(cudaMalloc(&e->h_obs_dev, obs_bytes));
  if (!act_zc) CUDA_TRY(cudaMemcpyAsync(e->h_act_dev, actions_host, nact, cudaMemcpyHostToDevice, st));
  if (stats_host) CUDA_TRY(cudaMemsetAsync(e->stats_dev, 0, 64, st));
  StepIO io = make_io(F_ALL | (e->p.auto_reset ? F_RESET : 0));
  io.actions = act_zc ? act_zc : e->h_act_dev; io.act_dtype = SWARM_ACT_U8;
  io.obs_out = obs_host ? (obs_zc ? obs_zc : e->h_obs_dev) : nullptr;
  io.reward_out = rew_zc ? rew_zc : e->h_rew_dev; io.done_out = done_zc ? done_zc : e->h_done_dev;
  io.stats = stats_host ? e->stats_dev : nullptr;
  rc = launch_step(e, io, st);
  if (rc) return rc;
  if (obs_host && !obs_zc) CUDA_TRY(cudaMemcpyAsync(obs_host, e->h_obs_dev, obs_bytes, cudaMemcpyDeviceToHost, st));
  if (!rew_zc) CUDA_TRY(cudaMemcpyAsync(reward_host, e->h_rew_dev, N * sizeof(float), cudaMemcpyDeviceToHost, st));
  if (!done_zc) CUDA_TRY(cudaMemcpyAsync(done_host, e->h_done_dev, N, cudaMemcpyDeviceToHost, st));
  if (pos_host) CUDA_TRY(cudaMemcpyAsync(pos_host, e->p.pos, nact * sizeof(float2), cudaMemcpyDeviceToHost, st));
  if (stats_host) CUDA_TRY(cudaMemcpyAsync(stats_host, e->stats_dev, 64, cudaMemcpyDeviceToHost, st));
  CUDA_TRY(cudaStreamSynchronize(st));
  return SWARM_OK;
}

extern "C" int swarm_step_host(SwarmEnv* e, const uint8_t* actions_host, void* obs_host, float* reward_host, uint8_t* done_host,
                               int64_t* stats_host, void* stream) {
  return step_host_impl(e, actions_host, obs_host, reward_host, done_host, nullptr, stats_host, stream);
}

extern "C" int swarm_step_host_pos(SwarmEnv* e, const uint8_t* actions_host, void* obs_host, float* reward_host, uint8_t* done_host,
                                   float* pos_host, int64_t* stats_host, void* stream) {
  if (!pos_host) return SWARM_ERR_INVALID_ARG;
  return step_host_impl(e, actions_host, obs_host, reward_host, done_host, pos_host, stats_host, stream);
}

// ------------------------------------------------------------------------------------------------ stages
extern "C" int swarm_stage_move(SwarmEnv* e, const void* actions, int action_dtype, void* stream) {
  if (!e || !actions || action_dtype < 0 || action_dtype > 2) return SWARM_ERR_INVALID_ARG;
  int rc = use_device(e);
  if (rc) return rc;
  const long long n = (long long)e->p.N * e->p.D;
  long long b = (n + 255) / 256;
  if (b > e->sm_count * 8) b = e->sm_count * 8;
  move_kernel<<<(int)b, 256, 0, (cudaStream_t)stream>>>(e->p, actions, action_dtype);
  e->launches++;
  CUDA_TRY(cudaGetLastError());
  return SWARM_OK;
}

extern "C" int swarm_stage_scatter(SwarmEnv* e, int32_t* cover_reward_out, void* stream) {
  if (!e || !cover_reward_out) return SWARM_ERR_INVALID_ARG;
  int rc = use_device(e);
  if (rc) return rc;
  StepIO io = make_io(F_SCATTER);
  io.cover_out = cover_reward_out;
  return launch_step(e, io, (cudaStream_t)stream);
}

extern "C" int swarm_stage_reward(SwarmEnv* e, const int32_t* cover_reward, float* reward_out, uint8_t* done_out, void* stream) {
  if (!e || !cover_reward || !reward_out || !done_out) return SWARM_ERR_INVALID_ARG;
  int rc = use_device(e);
  if (rc) return rc;
  StepIO io = make_io(F_REWARD);
  io.cover_in = cover_reward; io.reward_out = reward_out; io.done_out = done_out;
  return launch_step(e, io, (cudaStream_t)stream);
}

extern "C" int swarm_stage_obs(SwarmEnv* e, void* obs_out, void* stream) {
  if (!e || !obs_out) return SWARM_ERR_INVALID_ARG;
  int rc = use_device(e);
  if (rc) return rc;
  StepIO io = make_io(F_OBS);
  io.obs_out = obs_out;
  return launch_step(e, io, (cudaStream_t)stream);
}

extern "C" int swarm_greedy_actions(SwarmEnv* e, uint8_t* actions_out, void* stream) {
  if (!e || !actions_out) return SWARM_ERR_INVALID_ARG;
  int rc = use_device(e);
  if (rc) return rc;
  const long long n = (long long)e->p.N * e->p.D;
  long long b = (n + 255) / 256;
  if (b > e->sm_count * 8) b = e->sm_count * 8;
  greedy_kernel<<<(int)b, 256, 0, (cudaStream_t)stream>>>(e->p, actions_out);
  e->launches++;
  CUDA_TRY(cudaGetLastError());
  return SWARM_OK;
}

extern "C" int swarm_device_error(SwarmEnv* e, int32_t* code_out, int clear, void* stream) {
  if (!e || !code_out) return SWARM_ERR_INVALID_ARG;
  int rc = use_device(e);
  if (rc) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  CUDA_TRY(cudaMemcpyAsync(code_out, e->p.err, 4, cudaMemcpyDeviceToHost, st));
  if (clear) CUDA_TRY(cudaMemsetAsync(e->p.err, 0, 4, st));
  CUDA_TRY(cudaStreamSynchronize(st));
  return SWARM_OK;
}

extern "C" int64_t swarm_launch_count(const SwarmEnv* e) { return e ? e->launches : 0; }

// ------------------------------------------------------------------------------------------------ A2C rollout helpers
// Stateless (no handle): they work on caller-owned device buffers of the CURRENT device.
static int current_sm_count(int* sms) {
  int dev = -1, major = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) { cudaGetLastError(); return SWARM_ERR_NO_DEVICE; }
  CUDA_TRY(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev));
  if (major != 10) return SWARM_ERR_NO_DEVICE;
  CUDA_TRY(cudaDeviceGetAttribute(sms, cudaDevAttrMultiProcessorCount, dev));
  return SWARM_OK;
}

extern "C" int swarm_sample_actions(const float* logits, int32_t n_drones, int64_t n_envs, int32_t n_actions, uint64_t seed,
                                    uint32_t step, int64_t env_id_base, uint8_t* actions_out, void* stream) {
  if (!logits || !actions_out || n_drones < 1 || n_envs < 1 || n_actions < 1 || n_actions > 256) return SWARM_ERR_INVALID_ARG;
  if (n_actions == 4 && (((uintptr_t)logits) & 15) != 0) return SWARM_ERR_INVALID_ARG;
  int sms = 0;
  int rc = current_sm_count(&sms);
  if (rc) return rc;
  long long b = (n_envs + 255) / 256;
  if (b > (long long)sms * 8) b = (long long)sms * 8;
  cudaStream_t st = (cudaStream_t)stream;
  if (n_actions == 4)
    sample_actions_kernel<4><<<(int)b, 256, 0, st>>>(logits, n_drones, n_envs, 4, (uint32_t)seed, (uint32_t)(seed >> 32), step, env_id_base, actions_out);
  else
    sample_actions_kernel<0><<<(int)b, 256, 0, st>>>(logits, n_drones, n_envs, n_actions, (uint32_t)seed, (uint32_t)(seed >> 32), step, env_id_base, actions_out);
  CUDA_TRY(cudaGetLastError());
  return SWARM_OK;
}

extern "C" int swarm_nstep_returns(const float* rewards, const uint8_t* dones, const float* last_values, int32_t T, int64_t n_envs,
                                   int32_t n_drones, float gamma, float* returns_out, float* coef_out, void* stream) {
  if (!rewards || !dones || !last_values || !returns_out || T < 1 || n_envs < 1 || n_drones < 1) return SWARM_ERR_INVALID_ARG;
  int sms = 0;
  int rc = current_sm_count(&sms);
  if (rc) return rc;
  long long b = ((long long)n_envs * n_drones + 255) / 256;
  if (b > (long long)sms * 8) b = (long long)sms * 8;
  nstep_returns_kernel<<<(int)b, 256, 0, (cudaStream_t)stream>>>(rewards, dones, last_values, T, n_envs, n_drones, gamma, returns_out, coef_out);
  CUDA_TRY(cudaGetLastError());
  return SWARM_OK;
}

extern "C" int swarm_gru_gates_fwd(const float* gi, const float* bias_hh, int64_t n_envs, int32_t n_nets, int32_t hidden, float* h_out, void* stream) {
  if (!gi || !bias_hh || !h_out || n_envs < 1 || n_nets < 1 || hidden < 1) return SWARM_ERR_INVALID_ARG;
  if ((hidden & 3) == 0 && (((uintptr_t)gi | (uintptr_t)bias_hh | (uintptr_t)h_out) & 15)) return SWARM_ERR_INVALID_ARG;   // 128-bit path
  int sms = 0;
  int rc = current_sm_count(&sms);
  if (rc) return rc;
  long long b = ((long long)n_envs * n_nets * hidden + 255) / 256;
  if (b > (long long)sms * 16) b = (long long)sms * 16;
  gru_gates_fwd_kernel<<<(int)b, 256, 0, (cudaStream_t)stream>>>(gi, bias_hh, n_envs, n_nets, hidden, h_out);
  CUDA_TRY(cudaGetLastError());
  return SWARM_OK;
}

extern "C" int swarm_gru_gates_bwd(const float* gi, const float* bias_hh, const float* dh, int64_t n_envs, int32_t n_nets, int32_t hidden,
                                   float* dgi_out, float* dnr_out, void* stream) {
  if (!gi || !bias_hh || !dh || !dgi_out || !dnr_out || n_envs < 1 || n_nets < 1 || hidden < 1) return SWARM_ERR_INVALID_ARG;
  if ((hidden & 3) == 0 && (((uintptr_t)gi | (uintptr_t)bias_hh | (uintptr_t)dh | (uintptr_t)dgi_out | (uintptr_t)dnr_out) & 15)) return SWARM_ERR_INVALID_ARG;
  int sms = 0;
  int rc = current_sm_count(&sms);
  if (rc) return rc;
  long long b = ((long long)n_envs * n_nets * hidden + 255) / 256;
  if (b > (long long)sms * 16) b = (long long)sms * 16;
  gru_gates_bwd_kernel<<<(int)b, 256, 0, (cudaStream_t)stream>>>(gi, bias_hh, dh, n_envs, n_nets, hidden, dgi_out, dnr_out);
  CUDA_TRY(cudaGetLastError());
  return SWARM_OK;
}

extern "C" int swarm_bias_act(float* y, const float* bias, int32_t n_nets, int64_t n_rows, int32_t features, int32_t act, void* stream) {
  if (!y || !bias || n_nets < 1 || n_rows < 1 || features < 1 || act < 0 || act > 2) return SWARM_ERR_INVALID_ARG;
  if ((features & 3) == 0 && (((uintptr_t)y | (uintptr_t)bias) & 15)) return SWARM_ERR_INVALID_ARG;   // 128-bit path
  int sms = 0;
  int rc = current_sm_count(&sms);
  if (rc) return rc;
  const long long work = (long long)n_nets * n_rows * ((features & 3) == 0 ? features / 4 : features);
  long long b = (work + 255) / 256;
  if (b > (long long)sms * 16) b = (long long)sms * 16;
  bias_act_kernel<<<(int)b, 256, 0, (cudaStream_t)stream>>>(y, bias, n_nets, n_rows, features, act);
  CUDA_TRY(cudaGetLastError());
  return SWARM_OK;
}

static int row_kernel_blocks(int64_t rows, int* blocks) {
  int sms = 0;
  int rc = current_sm_count(&sms);
  if (rc) return rc;
  long long b = (rows + 7) / 8;   // one warp per row, 8 warps per block
  if (b > (long long)sms * 16) b = (long long)sms * 16;
  *blocks = (int)b;
  return SWARM_OK;
}

extern "C" int swarm_relu_l2norm_fwd(const float* x, int64_t rows, int32_t features, float eps, float* out, void* stream) {
  if (!x || !out || rows < 1 || features < 1 || (((uintptr_t)x | (uintptr_t)out) & 15)) return SWARM_ERR_INVALID_ARG;
  int b = 0;
  int rc = row_kernel_blocks(rows, &b);
  if (rc) return rc;
  relu_l2norm_fwd_kernel<<<b, 256, 0, (cudaStream_t)stream>>>(x, rows, features, eps, out);
  CUDA_TRY(cudaGetLastError());
  return SWARM_OK;
}

extern "C" int swarm_relu_l2norm_bwd(const float* x, const float* grad_out, int64_t rows, int32_t features, float eps, float* grad_x, void* stream) {
  if (!x || !grad_out || !grad_x || rows < 1 || features < 1 || (((uintptr_t)x | (uintptr_t)grad_out | (uintptr_t)grad_x) & 15)) return SWARM_ERR_INVALID_ARG;
  int b = 0;
  int rc = row_kernel_blocks(rows, &b);
  if (rc) return rc;
  relu_l2norm_bwd_kernel<<<b, 256, 0, (cudaStream_t)stream>>>(x, grad_out, rows, features, eps, grad_x);
  CUDA_TRY(cudaGetLastError());
  return SWARM_OK;
}

extern "C" int swarm_row_standardize(const float* x, int64_t rows, int32_t features, float* out, void* stream) {
  if (!x || !out || rows < 1 || features < 1) return SWARM_ERR_INVALID_ARG;
  int blocks = 0;
  int rc = row_kernel_blocks(rows, &blocks);
  if (rc) return rc;
  row_standardize_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(x, rows, features, out);
  CUDA_TRY(cudaGetLastError());
  return SWARM_OK;
}

extern "C" int swarm_embed_tanh_l2norm_fwd(const float* a, const float* b, const float* bias_a, const float* bias_b, int64_t n_rows, int32_t n_nets, int32_t fa, int32_t fb, float eps, float* out,
                                           void* stream) {
  if (!a || !b || !out || n_rows < 1 || n_nets < 1 || fa < 4 || fb < 4 || (fa & 3) || (fb & 3) || fa + fb > 256 ||
      (((uintptr_t)a | (uintptr_t)b | (uintptr_t)out | (uintptr_t)bias_a | (uintptr_t)bias_b) & 15))
    return SWARM_ERR_INVALID_ARG;
  int blocks = 0;
  int rc = row_kernel_blocks(n_rows * n_nets, &blocks);
  if (rc) return rc;
  embed_tanh_l2norm_fwd_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(a, b, bias_a, bias_b, n_rows, n_nets, fa, fb, eps, out);
  CUDA_TRY(cudaGetLastError());
  return SWARM_OK;
}

extern "C" int swarm_embed_tanh_l2norm_bwd(const float* a, const float* b, const float* bias_a, const float* bias_b, const float* grad_out, int64_t n_rows, int32_t n_nets, int32_t fa, int32_t fb,
                                           float eps, float* grad_a, float* grad_b, void* stream) {
  if (!a || !b || !grad_out || !grad_a || !grad_b || n_rows < 1 || n_nets < 1 || fa < 4 || fb < 4 || (fa & 3) || (fb & 3) || fa + fb > 256 ||
      (((uintptr_t)a | (uintptr_t)b | (uintptr_t)grad_out | (uintptr_t)grad_a | (uintptr_t)grad_b | (uintptr_t)bias_a | (uintptr_t)bias_b) & 15))
    return SWARM_ERR_INVALID_ARG;
  int blocks = 0;
  int rc = row_kernel_blocks(n_rows * n_nets, &blocks);
  if (rc) return rc;
  embed_tanh_l2norm_bwd_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(a, b, bias_a, bias_b, grad_out, n_rows, n_nets, fa, fb, eps, grad_a, grad_b);
  CUDA_TRY(cudaGetLastError());
  return SWARM_OK;
}

// ------------------------------------------------------------------------------------------------ NCCL (statistics only)
// The path's one collective: ncclAllReduce(SUM, int64[8]) of the episode statistics, 64 B per rank.  libnccl is
// bound at run time (dlopen of the copy the process already carries — PyTorch's — else the system one), so the library has no
// link-time NCCL dependency and single-GPU users never load it.
namespace {
struct NcclId128 { char b[128]; };   // ncclUniqueId (NCCL_UNIQUE_ID_BYTES = 128), passed by value
struct NcclApi {
  void* handle = nullptr;
  int (*GetUniqueId)(void*) = nullptr;
  int (*CommInitRank)(void**, int, NcclId128, int) = nullptr;
  int (*CommDestroy)(void*) = nullptr;
  int (*AllReduce)(const void*, void*, size_t, int, int, void*, cudaStream_t) = nullptr;
  const char* (*GetErrorString)(int) = nullptr;
};
NcclApi g_nccl;

int load_nccl() {
  if (g_nccl.handle) return SWARM_OK;
  void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD | RTLD_GLOBAL);
  if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
  if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
  if (!h) { snprintf(g_cuda_err, sizeof(g_cuda_err), "dlopen(libnccl.so.2): %s", dlerror()); return SWARM_ERR_NCCL; }
  g_nccl.GetUniqueId = (decltype(g_nccl.GetUniqueId))dlsym(h, "ncclGetUniqueId");
  g_nccl.CommInitRank = (decltype(g_nccl.CommInitRank))dlsym(h, "ncclCommInitRank");
  g_nccl.CommDestroy = (decltype(g_nccl.CommDestroy))dlsym(h, "ncclCommDestroy");
  g_nccl.AllReduce = (decltype(g_nccl.AllReduce))dlsym(h, "ncclAllReduce");
  g_nccl.GetErrorString = (decltype(g_nccl.GetErrorString))dlsym(h, "ncclGetErrorString");
  if (!g_nccl.GetUniqueId || !g_nccl.CommInitRank || !g_nccl.CommDestroy || !g_nccl.AllReduce) {
    snprintf(g_cuda_err, sizeof(g_cuda_err), "libnccl lacks a required symbol");
    return SWARM_ERR_NCCL;
  }
  g_nccl.handle = h;
  return SWARM_OK;
}

int nccl_fail(const char* what, int rc) {
  snprintf(g_cuda_err, sizeof(g_cuda_err), "%s: %s", what, g_nccl.GetErrorString ? g_nccl.GetErrorString(rc) : "nccl error");
  return SWARM_ERR_NCCL;
}
constexpr int kNcclInt64 = 4, kNcclSum = 0;   // ncclDataType_t / ncclRedOp_t values of nccl.h (stable since NCCL 2.0)
}  // namespace

extern "C" int swarm_comm_unique_id(uint8_t* id_out) {
  if (!id_out) return SWARM_ERR_INVALID_ARG;
  int rc = load_nccl();
  if (rc) return rc;
  int n = g_nccl.GetUniqueId(id_out);
  return n ? nccl_fail("ncclGetUniqueId", n) : SWARM_OK;
}

extern "C" int swarm_comm_init(const uint8_t* id, int32_t rank, int32_t world, void** comm_out) {
  if (!id || !comm_out || world < 1 || rank < 0 || rank >= world) return SWARM_ERR_INVALID_ARG;
  int rc = load_nccl();
  if (rc) return rc;
  NcclId128 u;
  memcpy(u.b, id, 128);
  void* comm = nullptr;
  int n = g_nccl.CommInitRank(&comm, world, u, rank);   // on the CURRENT CUDA device
  if (n) return nccl_fail("ncclCommInitRank", n);
  *comm_out = comm;
  return SWARM_OK;
}

extern "C" int swarm_comm_destroy(void* comm) {
  if (!comm) return SWARM_ERR_INVALID_ARG;
  int rc = load_nccl();
  if (rc) return rc;
  int n = g_nccl.CommDestroy(comm);
  return n ? nccl_fail("ncclCommDestroy", n) : SWARM_OK;
}

extern "C" int swarm_stats_allreduce(SwarmEnv* e, int64_t* stats, void* comm, void* stream) {
  if (!comm || (!e && !stats)) return SWARM_ERR_INVALID_ARG;
  int rc = load_nccl();
  if (rc) return rc;
  if (e && (rc = use_device(e))) return rc;
  void* buf = stats ? (void*)stats : (void*)e->stats_dev;
  int n = g_nccl.AllReduce(buf, buf, SWARM_STAT_COUNT, kNcclInt64, kNcclSum, comm, (cudaStream_t)stream);
  return n ? nccl_fail("ncclAllReduce", n) : SWARM_OK;
}

// ------------------------------------------------------------------------------------------------ peer-memory statistics
// The fused form of the path's one collective: no NCCL kernel, no extra launch — the step kernel publishes this rank's running
// totals into every rank's mailbox over NVLink at its start (peer_stats_push, swarm_common.cuh).
static int ensure_mailbox(SwarmEnv* e) {
  if (e->mbox) return SWARM_OK;
  const size_t bytes = (size_t)SWARM_MAX_RANKS * SWARM_MBOX_SLOT * sizeof(unsigned long long);
  CUDA_TRY(cudaMalloc((void**)&e->mbox, bytes));
  CUDA_TRY(cudaMemset(e->mbox, 0, bytes));
  CUDA_TRY(cudaMalloc((void**)&e->p.peers.pushes, 16));
  CUDA_TRY(cudaMemset(e->p.peers.pushes, 0, 16));
  CUDA_TRY(cudaMalloc((void**)&e->gstats_dev, 64));
  return SWARM_OK;
}

extern "C" int swarm_stats_mailbox_handle(SwarmEnv* e, uint8_t* handle_out) {
  if (!e || !handle_out) return SWARM_ERR_INVALID_ARG;
  int rc = use_device(e);
  if (rc) return rc;
  if ((rc = ensure_mailbox(e))) return rc;
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "cudaIpcMemHandle_t is 64 bytes");
  cudaIpcMemHandle_t h;
  CUDA_TRY(cudaIpcGetMemHandle(&h, e->mbox));
  memcpy(handle_out, &h, 64);
  return SWARM_OK;
}

extern "C" int swarm_stats_peers_open(SwarmEnv* e, const uint8_t* handles, int32_t rank, int32_t world) {
  if (!e || !handles || world < 1 || world > SWARM_MAX_RANKS || rank < 0 || rank >= world) return SWARM_ERR_INVALID_ARG;
  int rc = use_device(e);
  if (rc) return rc;
  if ((rc = ensure_mailbox(e))) return rc;
  if (e->p.peers.n_ranks > 0) return SWARM_ERR_INVALID_ARG;   // already open
  for (int r = 0; r < world; ++r) {
    if (r == rank) { e->p.peers.mbox[r] = e->mbox; continue; }
    cudaIpcMemHandle_t h;
    memcpy(&h, handles + (size_t)r * 64, 64);
    void* ptr = nullptr;
    CUDA_TRY(cudaIpcOpenMemHandle(&ptr, h, cudaIpcMemLazyEnablePeerAccess));
    e->peer_open[r] = ptr;
    e->p.peers.mbox[r] = (unsigned long long*)ptr;
  }
  e->p.peers.rank = rank;
  e->p.peers.n_ranks = world;
  return SWARM_OK;
}

extern "C" int swarm_stats_peers_close(SwarmEnv* e) {
  if (!e) return SWARM_ERR_INVALID_ARG;
  int rc = use_device(e);
  if (rc) return rc;
  CUDA_TRY(cudaDeviceSynchronize());
  e->p.peers.n_ranks = 0;
  for (int r = 0; r < SWARM_MAX_RANKS; ++r) {
    if (e->peer_open[r]) CUDA_TRY(cudaIpcCloseMemHandle(e->peer_open[r]));
    e->peer_open[r] = nullptr;
    e->p.peers.mbox[r] = nullptr;
  }
  return SWARM_OK;
}

extern "C" int swarm_stats_push(SwarmEnv* e, const int64_t* stats_dev, void* stream) {
  if (!e || e->p.peers.n_ranks < 1) return SWARM_ERR_INVALID_ARG;
  int rc = use_device(e);
  if (rc) return rc;
  peer_stats_push_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(e->p.peers, stats_dev ? (const unsigned long long*)stats_dev : e->stats_dev);
  e->launches++;
  CUDA_TRY(cudaGetLastError());
  return SWARM_OK;
}

extern "C" int swarm_stats_global(SwarmEnv* e, int64_t min_pushes, int64_t* out_dev, void* stream) {
  if (!e || !out_dev || min_pushes < 0 || e->p.peers.n_ranks < 1) return SWARM_ERR_INVALID_ARG;
  int rc = use_device(e);
  if (rc) return rc;
  peer_stats_reduce_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(e->mbox, e->p.peers.n_ranks, (unsigned long long)min_pushes,
                                                                (unsigned long long*)out_dev, e->p.err);
  e->launches++;
  CUDA_TRY(cudaGetLastError());
  return SWARM_OK;
}

extern "C" int swarm_stats_global_host(SwarmEnv* e, int64_t min_pushes, int64_t* out_host, void* stream) {
  if (!e || !out_host) return SWARM_ERR_INVALID_ARG;
  int rc = swarm_stats_global(e, min_pushes, (int64_t*)e->gstats_dev, stream);
  if (rc) return rc;
  CUDA_TRY(cudaMemcpyAsync(out_host, e->gstats_dev, 64, cudaMemcpyDeviceToHost, (cudaStream_t)stream));
  CUDA_TRY(cudaStreamSynchronize((cudaStream_t)stream));
  for (int k = 0; k < SWARM_STAT_COUNT; ++k)
    if (out_host[k] != -1) return SWARM_OK;
  return SWARM_ERR_TIMEOUT;   // every entry -1: the reduce kernel gave up (it never sums a torn slot)
}

// ------------------------------------------------------------------------------------------------ allocation-free host step
// swarm_step_host stages its host buffers through device memory it cudaMalloc's on first use.  Callers that may not allocate
// (CUDA-graph-adjacent serving loops, harnesses that forbid cudaMalloc) get the same call with the staging in this file-scope
// array instead: SWARM_STATIC_SCRATCH_BYTES on every device that loads the library, held until the process exits.  In host
// code `g_static_scratch` is only a stand-in for the symbol — the device address comes from cudaGetSymbolAddress, once per
// device, with that device current.  Every call (any handle, any stream) shares the array: calls that overlap on the GPU race.
static __device__ __align__(256) unsigned char g_static_scratch[SWARM_STATIC_SCRATCH_BYTES];

static int static_scratch_ptr(int device, unsigned char** out) {
  static unsigned char* ptr_of[64] = {nullptr};   // per device ordinal; resolved with that device current (use_device ran)
  if (device < 0 || device >= 64) return SWARM_ERR_INVALID_ARG;
  if (!ptr_of[device]) {
    unsigned char* ptr = nullptr;
    CUDA_TRY(cudaGetSymbolAddress((void**)&ptr, g_static_scratch));
    if (!ptr) return SWARM_ERR_CUDA;
    ptr_of[device] = ptr;
  }
  *out = ptr_of[device];
  return SWARM_OK;
}

extern "C" int swarm_step_host_static(SwarmEnv* e, const uint8_t* actions_host, void* obs_host, float* reward_host,
                                      uint8_t* done_host, int64_t* stats_host, void* stream) {
  if (!e || !actions_host || !reward_host || !done_host) return SWARM_ERR_INVALID_ARG;
  const size_t N = (size_t)e->p.N, nact = N * e->p.D, nobs = obs_host ? (size_t)swarm_obs_bytes(e) : 0;
  auto up = [](size_t v) { return (v + 255) / 256 * 256; };
  // layout: statistics int64[8] | actions uint8[N][D] | reward float[N] | done uint8[N] | obs   (each 256 B aligned)
  const size_t o_stats = 0, o_act = 256, o_rew = o_act + up(nact), o_done = o_rew + up(N * sizeof(float)),
               o_obs = o_done + up(N), need = o_obs + up(nobs);
  if (need > (size_t)SWARM_STATIC_SCRATCH_BYTES) {
    snprintf(g_cuda_err, sizeof(g_cuda_err),
             "swarm_step_host_static: staging for %zu envs needs %zu bytes, the static scratch holds %d (use swarm_step_host)",
             N, need, (int)SWARM_STATIC_SCRATCH_BYTES);
    return SWARM_ERR_INVALID_ARG;
  }
  int rc = use_device(e);
  if (rc) return rc;
  unsigned char* s = nullptr;
  if ((rc = static_scratch_ptr(e->cfg.device, &s)) != SWARM_OK) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  CUDA_TRY(cudaMemcpyAsync(s + o_act, actions_host, nact, cudaMemcpyHostToDevice, st));
  // the array keeps the previous call's contents; only the statistics accumulate, so only they start from zero
  if (stats_host) CUDA_TRY(cudaMemsetAsync(s + o_stats, 0, 64, st));
  StepIO io = make_io(F_ALL | (e->p.auto_reset ? F_RESET : 0));
  io.actions = s + o_act; io.act_dtype = SWARM_ACT_U8; io.obs_out = obs_host ? (void*)(s + o_obs) : nullptr;
  io.reward_out = (float*)(s + o_rew); io.done_out = s + o_done; io.stats = stats_host ? (unsigned long long*)(s + o_stats) : nullptr;
  rc = launch_step(e, io, st);
  if (rc) return rc;
  if (obs_host) CUDA_TRY(cudaMemcpyAsync(obs_host, s + o_obs, nobs, cudaMemcpyDeviceToHost, st));
  CUDA_TRY(cudaMemcpyAsync(reward_host, s + o_rew, N * sizeof(float), cudaMemcpyDeviceToHost, st));
  CUDA_TRY(cudaMemcpyAsync(done_host, s + o_done, N, cudaMemcpyDeviceToHost, st));
  if (stats_host) CUDA_TRY(cudaMemcpyAsync(stats_host, s + o_stats, 64, cudaMemcpyDeviceToHost, st));
  CUDA_TRY(cudaStreamSynchronize(st));
  return SWARM_OK;
}
