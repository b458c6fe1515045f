/*
 * swarm_b200.h — C ABI of the B200-native batched drone-swarm coverage environment.
 *
 * This is the drop-in boundary for ONE hot path of grohith327/MARL-Drones: the headless 2D
 * environment `DroneEnv.reset()/step()` (reference: A2C/noVis_sim.py:7-175 and
 * DQL_model/noVis_sim.py:7-168) and the stacked, auto-resetting vec-env surface around it
 * (reference: A2C/envs.py:17-41,58-117).  The reference has no FFI on this path — the
 * boundary there is the Python class — so these entry points are what a `ctypes` binding
 * on the reference side calls (see INTEGRATION.md for the stub).
 *
 * Conventions
 *  - every function returns int: 0 = SWARM_OK, negative = SwarmStatus error; no C++
 *    exception crosses the ABI; `swarm_strerror` turns a status into text.
 *  - the library owns the opaque handle and the resident struct-of-arrays state; the
 *    caller owns every input / output buffer.  Unless a name ends in `_host`, buffers are
 *    DEVICE pointers on the handle's device and all work is enqueued asynchronously on
 *    the `stream` argument (a `cudaStream_t` passed as `void*`; NULL = legacy default stream).
 *    No hidden synchronisation: every device entry point is CUDA-graph capturable.
 *    Step launches (swarm_step / swarm_rollout / swarm_stage_* / swarm_step_host*) are programmatic dependent launches: their
 *    kernels start their shared-memory prologue while the kernel before them on the stream drains and wait (griddepcontrol.wait)
 *    before touching global memory.  Nothing changes for the caller: work enqueued after them keeps its full stream-order
 *    dependency (unless it opts into programmatic launch itself).  SWARM_PDL=0 in the environment launches them as plain kernels.
 *  - invalid actions do not trap; they set a sticky device error word that
 *    `swarm_device_error` reads (the Python facade turns it into the reference's
 *    AssertionError, A2C/noVis_sim.py:118-120).
 *  - a handle is not thread-safe; one handle per device; multi-GPU = one handle per rank
 *    (`env_id_base` keys the RNG by GLOBAL env id, so any sharding yields identical envs).
 *
 * Layouts (N envs, D drones, R=row_count, C=col_count)
 *  - boundary grid      int8  [N][C][R]   grid[n][y][x] = reference `grid[x][y]`
 *                                         (== the reference obs order `np.array(grid).T.flatten()`,
 *                                         A2C/noVis_sim.py:168-170), values {0, 1, uncertainity}
 *  - positions          float [N][D][2]   (x, y) = reference `n_drones_pos[i]`
 *  - meta               int32 [N][4]      {step_func_count, remaining(#cells>0), episode_return, episode}
 *  - dwell counters     int16 [N][C][R]   only when dwell_normal|dwell_anom != 0
 *  - obs FULL           T     [N][R*C]    obs[n][y*R+x]                (T = float32 or uint8)
 *  - obs WINDOW         T     [N][D][W*W] win[n][i][b*W+a] = grid[cx-h+a][cy-h+b], pad -1 (255 for uint8)
 *  - obs WINDOW packed  u64   [N][D]      SWARM_DTYPE_P2: the 25 cells of a 5x5 window as two bit planes in one word (see the enum)
 *  - resident coverage map (private): 1 bit / cell (1 = still unmapped), 32 cells per uint32, flat index y*R+x, plus the
 *    env's anomalous cells as a list of 8 x uint32 (x | y << 16) or, for more than 8 anomalies / sides > 256, a second plane.
 */
#ifndef SWARM_B200_H_
#define SWARM_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SWARM_ABI_VERSION 1

typedef enum SwarmStatus {
  SWARM_OK = 0,
  SWARM_ERR_INVALID_ARG = -1,   /* NULL pointer, bad range, bad enum            */
  SWARM_ERR_CONFIG = -2,        /* unsupported SwarmConfig (see swarm_create)    */
  SWARM_ERR_CUDA = -3,          /* a CUDA runtime call failed (swarm_last_cuda_error) */
  SWARM_ERR_NO_DEVICE = -4,     /* no CUDA device / wrong architecture: there is NO CPU fallback */
  SWARM_ERR_ALLOC = -5,
  SWARM_ERR_NCCL = -6,
  SWARM_ERR_TIMEOUT = -7        /* the peer-statistics mailboxes never became consistent (swarm_stats_global_host) */
} SwarmStatus;

enum { SWARM_REWARD_A2C = 0, SWARM_REWARD_DQL = 1 };   /* A2C/noVis_sim.py:141-152,172-173 vs DQL_model/noVis_sim.py:138-148 */
enum { SWARM_OBS_FULL = 0, SWARM_OBS_WINDOW = 1 };     /* A2C/noVis_sim.py:168-170 vs SURVEY App. B.1 */
enum { SWARM_DTYPE_F32 = 0, SWARM_DTYPE_U8 = 1,        /* obs element type */
       SWARM_DTYPE_P2 = 2 };  /* packed 2-bit windows (WINDOW mode, W = 5, shapes the staged warp kernel serves: D <= 32, <= 8 anomalies,
                               * sides <= 256; anything else is SWARM_ERR_CONFIG): ONE uint64 per drone, obs uint64 [N][D].  For window
                               * cell k = b*5 + a (the layout of the WINDOW obs above): bit k = LO, bit 32 + k = HI, code = LO + 2 HI:
                               * 0 mapped -> 0, 1 unmapped -> 1, 3 unmapped and anomalous -> uncertainity, 2 outside the grid -> -1 (255).
                               * Lossless, 12.5x smaller than the float32 windows (8 B instead of 100 B per drone): the form for
                               * device-side consumers and for host links that are PCIe-bound; `swarm_b200.unpack_windows` expands it. */
enum { SWARM_ACT_U8 = 0, SWARM_ACT_I32 = 1, SWARM_ACT_I64 = 2 };

/* device error word bits (sticky until swarm_device_error(clear=1)) */
enum { SWARM_DEVERR_BAD_ACTION = 1, SWARM_DEVERR_STATS_TIMEOUT = 4,
       SWARM_DEVERR_BAD_STATE = 8 /* swarm_set_state: more cells of value U than n_anamolous */ };

/* statistics vector, int64[8], accumulated with atomics */
enum {
  SWARM_STAT_ENV_STEPS = 0, SWARM_STAT_AGENT_STEPS = 1, SWARM_STAT_REWARD_SUM = 2,
  SWARM_STAT_EPISODES_DONE = 3, SWARM_STAT_EPISODE_LEN_SUM = 4, SWARM_STAT_EPISODE_RETURN_SUM = 5,
  SWARM_STAT_CELLS_MAPPED = 6, SWARM_STAT_ANOMALOUS_MAPPED = 7, SWARM_STAT_COUNT = 8
};

/* Mirrors the kwargs of DroneEnv.__init__ (A2C/noVis_sim.py:8-17) plus the batch fields. */
typedef struct SwarmConfig {
  int32_t struct_size;     /* = sizeof(SwarmConfig); ABI guard                                   */
  int32_t n_envs;          /* N  (envs resident on this device)                                  */
  int32_t n_drones;        /* D  n_drones        (reference default 3)                           */
  int32_t row_count;       /* R  row_count       (10)                                            */
  int32_t col_count;       /* C  col_count       (10)                                            */
  int32_t n_anamolous;     /*    n_anamolous     (5)  [sic, reference spelling]                  */
  int32_t uncertainity;    /* U  uncertainity    (5)  must be in [2,127]                         */
  int32_t reward_mode;     /* SWARM_REWARD_*                                                     */
  int32_t obs_mode;        /* SWARM_OBS_*                                                        */
  int32_t window;          /* W, odd, WINDOW mode only (5)                                       */
  int32_t obs_dtype;       /* SWARM_DTYPE_*                                                      */
  int32_t auto_reset;      /* 1: reset on done and return the reset obs (A2C/envs.py:22-26)      */
  int32_t dwell_normal;    /* swarmsim-2d/run_env.py:46  (0 = noVis behaviour)                   */
  int32_t dwell_anom;      /* swarmsim-2d/run_env.py:47                                          */
  int32_t device;          /* CUDA device ordinal                                                */
  int32_t reserved0;
  double  collision_dist;  /* cd collision_dist  (0.5)                                           */
  double  step_size;       /* ss step_size       (0.5; every trainer passes 1.0).  Must be a dyadic rational
                            * m * 2^-k small enough that every reachable coordinate is exact in float32
                            * (the reference accumulates float64, A2C/noVis_sim.py:121-135); 0.1, 0.3 ... are
                            * SWARM_ERR_CONFIG rather than a silent drift across `dist < collision_dist`.   */
  uint64_t seed;           /* Philox key for device resets                                       */
  int64_t env_id_base;     /* global id of local env 0 (sharding: rank g owns [g*N, (g+1)*N))    */
} SwarmConfig;

typedef struct SwarmEnv SwarmEnv;   /* opaque */

/* -- lifetime ------------------------------------------------------------------------------ */
int swarm_abi_version(void);
const char* swarm_strerror(int status);
const char* swarm_last_cuda_error(void);
/* Number of CUDA devices with compute capability 10.x visible to this process (0 = none). */
int swarm_device_count(void);

/* Allocates the resident state for cfg->n_envs envs on cfg->device and initialises every env
 * (episode 0) exactly like DroneEnv.__init__ -> init_env (A2C/noVis_sim.py:34,36-63), with
 * Philox4x32-10 keyed by (seed, env_id_base+n, episode) in place of the reference's global
 * MT19937.  Rejects (SWARM_ERR_CONFIG): uncertainity outside [2,127], n_anamolous > R*C
 * (the reference loops forever, noVis_sim.py:47-55), even/oversized window, R*C > 2^24,
 * D > 1024. */
int swarm_create(const SwarmConfig* cfg, SwarmEnv** out);
int swarm_destroy(SwarmEnv* env);
int swarm_get_config(const SwarmEnv* env, SwarmConfig* out);

/* Elements of one obs tensor ([N][R*C] or [N][D][W*W]) and its byte size. */
int64_t swarm_obs_elems(const SwarmEnv* env);
int64_t swarm_obs_bytes(const SwarmEnv* env);

/* Device pointers into the resident state (views stay valid until swarm_destroy):
 * which = 0: positions float[N][D][2]; 1: meta int32[N][4]; 2: coverage bit plane uint32[N][env_words];
 * 3: device error word int32[1]; 4: internal statistics int64[8]. */
int swarm_state_ptr(SwarmEnv* env, int which, void** ptr_out, int64_t* bytes_out);

/* -- reset (DroneEnv.reset, A2C/noVis_sim.py:65-70; VecEnv.reset, A2C/envs.py:95-98) -------- */
/* mask: NULL = every env, else uint8[N] (non-zero = reset).  Each reset env advances to its
 * next episode.  obs_out may be NULL; otherwise the obs of ALL N envs is written. */
int swarm_reset(SwarmEnv* env, const uint8_t* mask, void* obs_out, void* stream);

/* -- state injection / extraction (parity harness, checkpoint) ------------------------------ */
/* Buffers describe envs [first, first+count).  grid int8[count][C][R]; pos float[count][D][2];
 * t int32[count] (step_func_count); wait int16[count][C][R] or NULL.  set_state recomputes
 * `remaining`, zeroes the episode return and keeps the episode counter. */
int swarm_set_state(SwarmEnv* env, const int8_t* grid, const float* pos, const int32_t* t,
                    const int16_t* wait, int32_t first, int32_t count, void* stream);
int swarm_get_state(SwarmEnv* env, int8_t* grid, float* pos, int32_t* t, int16_t* wait,
                    int32_t first, int32_t count, void* stream);

/* -- the hot path (DroneEnv.step, A2C/noVis_sim.py:112-175) --------------------------------- */
/* actions[N][D] of action_dtype; obs_out as described above (may be NULL to skip emission);
 * reward_out float[N]; done_out uint8[N]; stats_accum int64[8] or NULL (atomically accumulated;
 * the caller zeroes it).  One fused kernel launch: move+clip -> coverage scatter with
 * intra-warp conflict resolution -> reward/done reduction -> (auto-reset) -> obs emission. */
int swarm_step(SwarmEnv* env, const void* actions, int action_dtype, void* obs_out,
               float* reward_out, uint8_t* done_out, int64_t* stats_accum, void* stream);

/* T consecutive steps in ONE launch: actions uint8[T][N][D]; obs_out [T][...] (or NULL, or
 * only the last step's obs when obs_last_only != 0); reward_out float[T][N]; done_out uint8[T][N]. */
int swarm_rollout(SwarmEnv* env, int32_t T, const uint8_t* actions, void* obs_out,
                  int32_t obs_last_only, float* reward_out, uint8_t* done_out,
                  int64_t* stats_accum, void* stream);

/* Host-buffer form of swarm_step (what a CPU-side caller of the reference's VecEnv.step sees):
 * copies actions H2D, steps, copies obs/reward/done D2H through internal staging buffers and
 * synchronises `stream` before returning.  Host buffers should be pinned for full PCIe rate.
 * obs_host may be NULL (observation stays on the device for an on-device policy). */
int swarm_step_host(SwarmEnv* env, const uint8_t* actions_host, void* obs_host,
                    float* reward_host, uint8_t* done_host, int64_t* stats_host, void* stream);

/* swarm_step_host that also returns the positions after the step, float[N][D][2] — the reference's `env.n_drones_pos`, which
 * its callers read after every step (A2C/train.py:54,81,229,266; DQL_model/DQLearning.py:166,173; A2C/a2c_eval.py:77,97).  One call
 * is a whole `DroneEnv.step` as the single-env facade needs it.  For small batches (actions + obs + reward + done <= 64 KiB) both
 * entry points hand PINNED host buffers to the kernel itself (it reads the actions and writes the results over PCIe), so the
 * call costs one launch, the position copy and one synchronisation. */
int swarm_step_host_pos(SwarmEnv* env, const uint8_t* actions_host, void* obs_host, float* reward_host,
                        uint8_t* done_host, float* pos_host, int64_t* stats_host, void* stream);

/* swarm_step_host without any allocation.  NO COUNTERPART IN THE REFERENCE: added for
 * callers that cannot call cudaMalloc (code next to CUDA graph capture, tight serving loops, harnesses that forbid it).
 * Same arguments, results and synchronisation as swarm_step_host, but the device-side staging of actions / obs / reward /
 * done / statistics lives in ONE file-scope `static __device__` array of SWARM_STATIC_SCRATCH_BYTES that every GPU which
 * loads the library holds until the process exits; the call issues no cudaMalloc / cudaFree (swarm_create did the handle's).
 * Limit: 256 + up256(N*D) + up256(4*N) + up256(N) + up256(obs bytes, 0 when obs_host == NULL) <= SWARM_STATIC_SCRATCH_BYTES,
 * else SWARM_ERR_INVALID_ARG with the sizes spelled out in swarm_last_cuda_error().  The array is shared by every handle and
 * stream of the process on that device and keeps the last call's contents (the statistics slot is cleared per call): two
 * calls that overlap on the GPU race — serialise them. */
#define SWARM_STATIC_SCRATCH_BYTES (4 * 1024 * 1024)
int swarm_step_host_static(SwarmEnv* env, const uint8_t* actions_host, void* obs_host,
                           float* reward_host, uint8_t* done_host, int64_t* stats_host, void* stream);

/* -- per-stage entry points (unit parity of the four stages named in the north star) --------- */
/* K1 action decode + move + clip (A2C/noVis_sim.py:117-135): updates resident positions. */
int swarm_stage_move(SwarmEnv* env, const void* actions, int action_dtype, void* stream);
/* K2 coverage scatter (A2C/noVis_sim.py:75-90,104-110) for the CURRENT positions; writes the
 * per-env coverage reward (10 / 30 per newly mapped cell) as int32[N]. */
int swarm_stage_scatter(SwarmEnv* env, int32_t* cover_reward_out, void* stream);
/* K3 reward tail + spread + done (A2C/noVis_sim.py:92-102,141-166,172-173): consumes the
 * int32[N] coverage reward, advances step_func_count, never auto-resets. */
int swarm_stage_reward(SwarmEnv* env, const int32_t* cover_reward, float* reward_out,
                       uint8_t* done_out, void* stream);
/* K4 observation emission for the current state. */
int swarm_stage_obs(SwarmEnv* env, void* obs_out, void* stream);

/* -- scripted action source ------------------------------------------------------------------ */
/* Greedy nearest-unmapped-cell pursuit on the noVis lattice, writes uint8[N][D].  A 4-ACTION VARIANT of the 2D
 * simulator's controller (swarmsim-2d/drone.py:29-58), NOT a restatement of it: the target choice is the reference's
 * (nearest cell that is not mapped, row-major scan, strict <, including its `if not min_dist` quirk — pinned by
 * tests/golden/drone_update_targets.npz), but the reference then moves +-1 on BOTH axes at once (8-connected),
 * which the noVis action set {0: y-, 1: x+, 2: y+, 3: x-} cannot express; this source moves along the axis with the
 * larger |delta| (ties: x). */
int swarm_greedy_actions(SwarmEnv* env, uint8_t* actions_out, void* stream);

/* -- batched A2C rollout helpers (stateless, buffers on the current device) ---- */
/* softmax + Categorical(probs).sample() of A2C/train.py:66-73 for every (env, drone):
 * logits float32 [D][N][A] (16 B aligned when A == 4) -> actions uint8 [N][D], the env's action layout.
 * u ~ Philox4x32-10 keyed by (seed; env_id_base + n, step, d / 4); a = #{k : cum_k <= u * sum}. */
int swarm_sample_actions(const float* logits, int32_t n_drones, int64_t n_envs, int32_t n_actions,
                         uint64_t seed, uint32_t step, int64_t env_id_base, uint8_t* actions_out,
                         void* stream);
/* process_rollout's reverse scan (A2C/train.py:160-183): returns[t] = rewards[t] + returns[t+1] * gamma * (1 - dones[t])
 * in float32 with the reference's operation order, bootstrapped from last_values.
 * rewards float32 [T][N], dones uint8 [T][N], last_values float32 [D][N] -> returns float32 [T][D][N];
 * coef_out (may be NULL) float32 [T][N] = prod_{s>=t} gamma * (1 - dones[s]) = d returns[t] / d last_values. */
int swarm_nstep_returns(const float* rewards, const uint8_t* dones, const float* last_values, int32_t T,
                        int64_t n_envs, int32_t n_drones, float gamma, float* returns_out,
                        float* coef_out, void* stream);

/* The GRU cell of the stacked policy nets (A2C/models.py:104, `_, h = self.gru(x.unsqueeze(0))`: ONE step from h0 = 0, so
 * only bias_hh survives of the hidden side), after the input-side GEMM, fused into one pass:
 *   r = sigmoid(gi_r + bh_r), z = sigmoid(gi_z + bh_z), n = tanh(gi_n + r * bh_n), h = (1 - z) * n   (gate order r, z, n)
 * gi float32 [N][G][3H] (= x W_ih^T + b_ih), bias_hh float32 [G][3H] -> h_out float32 [G][N][H].  G = 2D stacked nets. */
int swarm_gru_gates_fwd(const float* gi, const float* bias_hh, int64_t n_envs, int32_t n_nets,
                        int32_t hidden, float* h_out, void* stream);
/* Its backward: dh float32 [G][N][H] -> dgi_out float32 [N][G][3H] (gradient of the pre-activations = of gi; summed over N
 * it is also the gradient of b_ih and of the r, z thirds of bias_hh) and dnr_out float32 [N][G][H] = d(pre_n) * r, whose
 * sum over N is the gradient of bias_hh's n third.  The gates are recomputed from gi. */
int swarm_gru_gates_bwd(const float* gi, const float* bias_hh, const float* dh, int64_t n_envs,
                        int32_t n_nets, int32_t hidden, float* dgi_out, float* dnr_out, void* stream);

/* Epilogue of the stacked linears (nn.Linear bias + F.relu, A2C/models.py:106-116; bias + torch.tanh of the deep-Q nets,
 * DQL_model/DQLearning.py:43-48) after a plain batched GEMM, in place:
 *   y[g][n][:] = act(y[g][n][:] + bias[g][:])      y float32 [G][N][F], bias float32 [G][F]
 * act: 0 identity, 1 max(., 0), 2 tanhf; anything else is SWARM_ERR_INVALID_ARG. */
int swarm_bias_act(float* y, const float* bias, int32_t n_nets, int64_t n_rows, int32_t features,
                   int32_t act, void* stream);
/* relu + F.normalize(dim=-1) of the policy's state embedding (A2C/models.py:107-108) in one pass:
 *   y = max(x, 0), out = y / max(||y||_2, eps)      x, out float32 [rows][features], 16 B aligned. */
int swarm_relu_l2norm_fwd(const float* x, int64_t rows, int32_t features, float eps, float* out,
                          void* stream);
/* Its backward (y and the norm are recomputed from x): grad_x = (g / d - y (g.y) / d^3) where x > 0, d = max(||y||, eps);
 * rows whose norm is below eps get g / eps, as torch's clamp_min does. */
int swarm_relu_l2norm_bwd(const float* x, const float* grad_out, int64_t rows, int32_t features,
                          float eps, float* grad_x, void* stream);

/* Input normalisation of both policy families at batch 1 (A2C/models.py:101-103, DQL_model/DQLearning.py:40-42), row-wise:
 *   out[r][:] = (x[r][:] - mean(x[r])) / std(x[r])      unbiased std (torch.Tensor.std()); x, out float32 [rows][features]. */
int swarm_row_standardize(const float* x, int64_t rows, int32_t features, float* out, void* stream);

/* Embedding of the stacked deep-Q nets (DQL_model/DQLearning.py:43-46) in one pass:
 *   out[d][n][:] = F.normalize(cat(tanh(a[n][d][:] + bias_a[d][:]), tanh(b[n][d][:] + bias_b[d][:])))   (L2 over fa + fb, clamp eps)
 * a float32 [n_rows][n_nets][fa] = state_input's product with the state rows and b float32 [n_rows][n_nets][fb] = drone_input's
 * with the drone positions, for all nets; bias_a [n_nets][fa], bias_b [n_nets][fb] may be NULL (bias already added);
 * out float32 [n_nets][n_rows][fa + fb].  fa, fb multiples of 4, fa + fb <= 256, 16 B aligned pointers. */
int swarm_embed_tanh_l2norm_fwd(const float* a, const float* b, const float* bias_a, const float* bias_b, int64_t n_rows, int32_t n_nets, int32_t fa, int32_t fb,
                                float eps, float* out, void* stream);
/* Its backward: grad_out float32 [n_nets][n_rows][fa + fb] -> grad_a [n_rows][n_nets][fa], grad_b [n_rows][n_nets][fb]
 * (tanh recomputed from a, b and the biases; the bias gradients are the column sums of grad_a / grad_b over n_rows). */
int swarm_embed_tanh_l2norm_bwd(const float* a, const float* b, const float* bias_a, const float* bias_b,
                                const float* grad_out, int64_t n_rows, int32_t n_nets,
                                int32_t fa, int32_t fb, float eps, float* grad_a, float* grad_b, void* stream);

/* -- the path's one collective: episode statistics over NCCL --------------------- */
/* libnccl.so.2 is bound with dlopen at first use (the process's own copy, e.g. PyTorch's); no link-time dependency.
 * `comm` is an ncclComm_t passed as void*: either one the caller already owns, or one made here:
 * rank 0 calls swarm_comm_unique_id (ncclGetUniqueId, 128 bytes), ships the bytes to every rank out of band
 * (e.g. torch.distributed.broadcast_object_list), and every rank calls swarm_comm_init on its own device. */
int swarm_comm_unique_id(uint8_t* id_out_128);
int swarm_comm_init(const uint8_t* id_128, int32_t rank, int32_t world, void** comm_out);
int swarm_comm_destroy(void* comm);
/* ncclAllReduce(SUM, int64[8]) in place on `stats` (device; NULL = the handle's internal statistics used by
 * swarm_step_host), enqueued on `stream`; integer sums, so the result does not depend on the reduction order. */
int swarm_stats_allreduce(SwarmEnv* env, int64_t* stats, void* comm, void* stream);

/* Fused form of the same exchange over NVLink peer memory (no collective kernel, no extra launch): every rank owns a
 * small mailbox; once the mailboxes are exchanged, every swarm_step / swarm_rollout launch that is given a statistics
 * buffer publishes, as its FIRST action (one thread, 16-byte posted stores, nothing waits on them), the buffer's running
 * totals as the previous launches left them into this rank's slot of every rank's mailbox — the feed lags the step by one
 * launch and costs it nothing.  swarm_stats_push publishes the current totals explicitly (stream-ordered, exact).
 * Ranks = processes of ONE node, world <= 8.
 *   swarm_stats_mailbox_handle  allocates the mailbox on first use, returns its cudaIpcMemHandle_t (64 bytes)
 *   swarm_stats_peers_open      handles = [world][64] gathered from all ranks (own entry ignored); enables the pushes
 *   swarm_stats_push            one-thread kernel: publish stats_dev (NULL = the handle's internal statistics) now
 *   swarm_stats_global          out_dev int64[8] (device) = sum over ranks of their latest complete totals; waits (bounded)
 *                               until every rank has pushed at least min_pushes times (0 = whatever has arrived).
 *                               Asynchronous on `stream`.  A slot that does not become consistent within the bound (a peer
 *                               died or never pushed) is NOT summed: out_dev is filled with -1 and
 *                               SWARM_DEVERR_STATS_TIMEOUT is set in the device error word (swarm_device_error).
 *   swarm_stats_global_host     the same into HOST memory int64[8]; synchronises `stream` and returns SWARM_ERR_TIMEOUT
 *                               in that case                                                                            */
int swarm_stats_mailbox_handle(SwarmEnv* env, uint8_t* handle_out_64);
int swarm_stats_peers_open(SwarmEnv* env, const uint8_t* handles, int32_t rank, int32_t world);
int swarm_stats_peers_close(SwarmEnv* env);
int swarm_stats_push(SwarmEnv* env, const int64_t* stats_dev, void* stream);
int swarm_stats_global(SwarmEnv* env, int64_t min_pushes, int64_t* out_dev, void* stream);
int swarm_stats_global_host(SwarmEnv* env, int64_t min_pushes, int64_t* out_host, void* stream);

/* -- diagnostics ---------------------------------------------------------------------------- */
/* Synchronises `stream`, reads the sticky device error word; clears it when clear != 0. */
int swarm_device_error(SwarmEnv* env, int32_t* code_out, int clear, void* stream);
/* Number of kernel launches issued through this handle since creation. */
int64_t swarm_launch_count(const SwarmEnv* env);

#ifdef __cplusplus
}
#endif
#endif  /* SWARM_B200_H_ */
