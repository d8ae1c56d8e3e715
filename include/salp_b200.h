/*
 * salp_b200.h -- C ABI of the B200-native batched SALP swimmer environment step.
 *
 * This is the drop-in boundary for ONE hot path of SungRoboticsGroup/UNDERWATER-SWIMMER_RL:
 * SalpSnakeEnv.reset / SalpSnakeEnv.step over the legacy SalpRobotEnv integrator
 *   R = scripts/utilities/salp_robot.py                 (reference tree)
 *   S = src/salp/environments/salp_snake_env.py
 * The reference has no FFI (it is pure Python); each entry point below names the Python
 * method(s) it replaces.  A Python host binds this with ctypes (INTEGRATION.md shows the
 * stub); the in-repo host is underwater_swimmer_rl_b200/vector_env.py.
 *
 * Conventions
 *   - Every function returns 0 on success, a negative SALP_E* code on failure; the
 *     message is available from salp_last_error() (thread-local).  No exceptions cross
 *     the ABI; CUDA errors are returned, never aborted on.
 *   - All DEVICE buffers are allocated and owned by the caller (PyTorch's caching
 *     allocator in the in-repo host).  The library never allocates, frees or copies
 *     tensor memory in the async API; a handle owns only its parameter block (and, for
 *     the optional *_host entry points, private staging buffers and streams).
 *   - salp_reset / salp_step are asynchronous on the caller-supplied CUDA stream and
 *     never synchronise.  One handle per GPU per process; a handle is not thread-safe,
 *     different handles are independent.
 *   - There is no CPU fallback: without a CUDA device every compute entry point fails
 *     with SALP_ECUDA.
 */
#ifndef SALP_B200_H_
#define SALP_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SALP_ABI_VERSION 5

enum { SALP_OK = 0, SALP_EINVAL = -1, SALP_ECUDA = -2, SALP_EUNBOUND = -3, SALP_EUNSUPPORTED = -4 };

/* arithmetic type of the integrator ("dtype" of the build, not a precision claim) */
enum { SALP_F32 = 0, /* throughput build: float state/actions/reward                      */
       SALP_F64 = 1  /* validation build: double, reference operation order, no FMA fusion (the float32 observation
                        columns are scaled with reciprocals: <= 1 ulp(double) before the cast)              */ };

enum { SALP_MAX_FOOD = 32, SALP_MAX_OBSERVED_FOOD = 8 };

/* flags for SalpParams.flags */
enum { SALP_AUTORESET = 1,      /* done lanes are reset inside the step (gymnasium 0.29 vector semantics)           */
       SALP_RNG_PHILOX = 2,     /* the words of both random streams are computed in the kernel (Philox4x32-10 keyed by
                                   salp_set_rng_seed, counter = (word index / 4, stream, GLOBAL env id)) instead of
                                   being read from the host-seeded MT19937 tables: unbounded, aperiodic, no memory;
                                   food_words / jit_words may be NULL.  Same consumption discipline as the tables
                                   (cursors in `ctrl` count words), no parity with the reference's generators.    */
       SALP_GENERIC_KERNEL = 4  /* diagnostics: never select the specialised forced-breathing step kernel          */ };

/* event bits written to SalpBuffers.events */
enum { SALP_EV_TERMINATED = 1, SALP_EV_TRUNCATED = 2, SALP_EV_COLLISION = 4, SALP_EV_FOOD = 8 };

/*
 * Task parameters == the keyword arguments of SalpSnakeEnv.__init__ (S:29-33) plus the
 * tank size of SalpRobotEnv.__init__ (R:28).  The robot constants of R:34-53 and the
 * food constants of S:53-54 are fixed in the kernel, as they are in the reference.
 */
typedef struct SalpParams {
    double width, height;             /* R:28            default 800 x 600           */
    int32_t num_food_items;           /* S:30,36         base_num_food_items, 0..32  */
    int32_t max_observed_food;        /* S:32,44         1..8, default 3             */
    int32_t forced_breathing;         /* S:31,43,66      1: action = [nozzle]; 0: [inhale, nozzle] */
    int32_t random_food_count;        /* S:32,37,144-146                             */
    int32_t respawn_food;             /* S:32,38,179                                 */
    int32_t max_steps_without_food;   /* S:33,46,185                                 */
    double food_reward;               /* S:30,39                                     */
    double collision_penalty;         /* S:30,40                                     */
    double time_penalty;              /* S:31,41                                     */
    double efficiency_bonus;          /* S:31,42,287-290                             */
    double proximity_reward_weight;   /* S:33,45,297-322                             */
    int32_t flags;                    /* SALP_AUTORESET                              */
    int32_t reserved_;
} SalpParams;

/* Sizes the caller needs to allocate SalpBuffers; filled by salp_query_layout(). */
typedef struct SalpLayout {
    int32_t real_bytes;      /* 4 (SALP_F32) or 8 (SALP_F64)                                   */
    int32_t obs_dim;         /* 10 + 4*max_observed_food + 2   (S:79-80)                       */
    int32_t act_dim;         /* 1 forced breathing, 2 manual   (S:69-74)                       */
    int32_t food_slots;      /* F: food records per env kept in HBM (>= num_food_items)        */
    int32_t body_reals;      /* 8  reals per env: [2][N] records of 4                          */
    int32_t ctrl_ints;       /* 8  int32 per env: [2][N] records of 4                          */
    int32_t stats_len;       /* 8  doubles                                                     */
    int32_t reserved_;
} SalpLayout;

/*
 * Device buffers.  N = num_envs of the handle.  "real" = float or double per dtype.
 * State lives in HBM as structure-of-arrays of 128-bit (f32) / 256-bit (f64) records so
 * that every per-env load/store is one vector access and a warp touches 512 contiguous bytes:
 *   body      real [2][N][4]   rec0 = (pos.x, pos.y, vel.x, vel.y)          R:61-62
 *                              rec1 = (angle, angular_vel, nozzle, water)    R:63-67,77
 *   ctrl      i32  [2][N][4]   rec0 = (fsm, steps_since_food, food_collected, food_mask)
 *                              rec1 = (food_cursor, jitter_cursor, episode_len | n_food << 26,
 *                                      jitter_cache)
 *             episode_len saturates at 2^26-1.  jitter_cache is the library-maintained copy of
 *             jit_words[row][jitter_cursor] (first word of the next np.random draw): the f32 step
 *             kernel reads it with the rest of the state instead of chasing cursor -> table (a
 *             dependent DRAM round trip) and refills it right after a draw.  salp_reset sets it for
 *             the lanes it resets; call salp_refresh_rng after binding new tables or after
 *             overwriting the cursors in `ctrl` from outside.
 *             fsm bits: [0:2) breathing phase 0 rest/1 inhaling/2 exhaling (R:71)
 *                       [2:10) breathing_timer (R:72)   [10:18) current_exhale_duration (R:223)
 *                       [18:20) body-shape source 0 reset(30,30)/1 rest/2 inhale/3 exhale
 *                       [20:28) timer value the ellipse axes were last computed from
 *             (ellipse_a/b of R:74-75 are recomputed from the shape fields, never stored)
 *   ep_return real [N]         running episode return
 *   food      real [F][N][2]   food_positions (S:52); liveness is ctrl.food_mask
 *   obs       f32  [K][N][obs_dim]      written by salp_step (and [N][obs_dim] by salp_reset)
 *   reward    real [K][N]
 *   terminated, truncated  u8 [K][N]    (torch.bool compatible)
 *   events    u8   [K][N]  SALP_EV_* bits, may be NULL
 *   final_obs f32  [K][N][obs_dim]  terminal observation, written ONLY for lanes that
 *             finished at that substep in SALP_AUTORESET mode; may be NULL
 *   final_info real [K][N][4]  (episode return, episode length, food_collected,
 *             steps_since_food) of the episode that just ended (the reference's terminal
 *             `info`, S:195-200), written ONLY for lanes that finished; may be NULL
 *   stats     f64  [8]  accumulated: episodes, sum return, sum length, sum food, collisions,
 *             truncations, non-finite lanes, env-steps; may be NULL
 *   food_words / jit_words  u32 [rows][len]  host-seeded MT19937 output streams the
 *             reference would draw from `random` (S:101-104,239-242) and `np.random`
 *             (R:311); env e (local index) reads row e % table_rows at its cursor (mod len).
 *             The host fills row r from seed (seed0 + env_id0 + r), so with
 *             table_rows == N results do not depend on how envs are sharded over GPUs.
 */
typedef struct SalpBuffers {
    void *body;
    int32_t *ctrl;
    void *ep_return;
    void *food;
    float *obs;
    void *reward;
    uint8_t *terminated;
    uint8_t *truncated;
    uint8_t *events;
    float *final_obs;
    void *final_info;
    double *stats;
    const uint32_t *food_words;
    const uint32_t *jit_words;
    int64_t table_rows;
    int64_t food_len; /* words per row, power of two */
    int64_t jit_len;  /* words per row, power of two, >= 2 */
    int32_t max_substeps; /* K the obs/reward/... buffers were sized for */
    int32_t reserved_;
} SalpBuffers;

typedef struct SalpHandle SalpHandle;

int salp_abi_version(void);

/* Thread-local message of the last failing call ("" if none). */
const char *salp_last_error(void);

/* Validate params and report buffer geometry.  Pure host code. */
int salp_query_layout(const SalpParams *params, int dtype, SalpLayout *out);

/*
 * Replaces SalpSnakeEnv.__init__ (S:29-90) for a slab of `num_envs` independent
 * environments with global ids env_id0 .. env_id0+num_envs-1 on CUDA device `device`.
 */
int salp_create(const SalpParams *params, int64_t num_envs, int64_t env_id0, int device, int dtype,
                SalpHandle **out);
int salp_destroy(SalpHandle *h);

/* Seed of the in-kernel Philox streams (SALP_RNG_PHILOX); takes effect for launches issued after the call. */
int salp_set_rng_seed(SalpHandle *h, uint64_t seed);

/*
 * Forced breathing (S:31,66 `forced_breathing=True`, R:158-167): phase, timer, ellipse axes, water volume and the thrust
 * window are functions of the position in the 273-step cycle.  The f32 forced-breathing kernel reads them from a table
 * built on the host in double with the reference's expressions (R:184-259); this call returns that table's sources so a
 * test can check them bit-for-bit against the fp64 validation kernel, which runs the recurrences:
 *   index 0 = fresh reset, index i = 1..273 = state after the i-th step of a cycle;  arrays of 274, any may be NULL.
 *   touch[i] bits: a clamp onto the low / right / bottom wall registers as a wall contact (R:335-352 + S:225-228).
 */
int salp_forced_cycle(const SalpParams *params, double *water_volume, double *ellipse_a, double *ellipse_b,
                      int32_t *fsm, int32_t *touch);

/* Attach caller-owned device buffers (raw pointers, e.g. torch.Tensor.data_ptr()). */
int salp_bind(SalpHandle *h, const SalpBuffers *buffers);

/*
 * Replaces SalpSnakeEnv.reset (S:133-155 over R:95-117) for every env whose byte in the
 * DEVICE array `mask` is non-zero (all envs when mask is NULL).  Writes the reset
 * observation to obs[0][e][:].  RNG cursors are NOT rewound (the reference keeps
 * consuming its global streams across episodes); pass rewind_rng != 0 to restart them.
 */
int salp_reset(SalpHandle *h, const uint8_t *mask, int rewind_rng, void *cuda_stream);

/*
 * Re-read ctrl.jitter_cache of every env from jit_words at the env's current jitter cursor
 * (see SalpBuffers).  Asynchronous on `cuda_stream`.
 */
int salp_refresh_rng(SalpHandle *h, void *cuda_stream);

/*
 * Replaces SalpSnakeEnv._get_extended_observation (S:366-428 over R:371-388) as the reference's callers use it:
 * they overwrite robot_pos / robot_velocity / robot_angle / robot_angular_velocity / food_positions /
 * steps_since_food and then ask for the observation of that state WITHOUT stepping
 * (eval/collect_navigation_data.py:76-89, eval/watch_navigation.py:63).  Writes obs[slot][e][:] for every env
 * from the current contents of body / ctrl / food; changes no state.  Asynchronous on `cuda_stream`.
 */
int salp_observe(SalpHandle *h, int slot, void *cuda_stream);

/*
 * Replaces the run-time assignment `env.base_num_food_items = n` (src/salp/training/continuous_trainer.py:408-411,
 * adaptive difficulty): episodes that start after this call (salp_reset, or the in-kernel auto-reset of later
 * steps) are laid out with n foods (or randint(1, n) with random_food_count).  Running episodes keep their
 * food.  n must not exceed SalpLayout.food_slots of the handle (SALP_EUNSUPPORTED otherwise: the kernel
 * instantiation and the `food` buffer were sized at salp_create).  Takes effect for launches issued after it.
 */
int salp_set_num_food_items(SalpHandle *h, int32_t num_food_items);

/*
 * Replaces K consecutive SalpSnakeEnv.step calls (S:157-202 over R:119-156) per env in ONE
 * kernel launch: state is read once, kept in registers for the K substeps, written once.
 * `actions` is a DEVICE pointer to real [K][N][act_dim] (any caller-owned memory, e.g. the
 * policy's output tensor -- no staging copy); it must stay valid until the launch has run.
 * 1 <= K <= buffers.max_substeps.
 */
int salp_step(SalpHandle *h, int K, const void *actions, void *cuda_stream);

/*
 * One env step whose outputs go to row `slot` (0 <= slot < buffers.max_substeps) of the [K][N]...
 * output buffers instead of row 0, so that a caller can keep the previous step's observation
 * alive (ping-pong) -- e.g. to store (obs, action, reward, next_obs) transitions without a copy.
 */
int salp_step_slot(SalpHandle *h, int slot, const void *actions, void *cuda_stream);

/*
 * Host-buffer convenience path (what a caller holding numpy arrays uses): copies
 * actions host->device (into the caller-owned DEVICE staging buffer `actions_staging`,
 * real [N][act_dim]), runs one step (K=1), copies obs/reward/terminated/truncated
 * device->host -- the observation rows in 8 chunks on internal non-blocking streams so that the copies overlap the
 * kernels of the later chunks, rewards and flags as one copy per array once every chunk has run -- and
 * returns after the results are in host memory.  The chunks start after the work already queued
 * on `caller_stream` (an event is recorded there); work queued on it later runs after the call
 * has returned.  Host pointers should be page-locked.  Any output pointer may be NULL.
 */
int salp_step_host(SalpHandle *h, void *actions_staging, const void *actions_host, float *obs_host,
                   void *reward_host, uint8_t *terminated_host, uint8_t *truncated_host, void *caller_stream);

/*
 * salp_step_host without any allocation.  This entry point has NO counterpart in the reference; it was added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight
 * serving loops, harnesses that forbid it).  Same inputs, outputs and result as salp_step_host, but
 *   - the device staging buffer the step kernel reads its actions from is not passed in: it is a file-scope
 *     `static __device__ __align__(256)` array of SALP_NOALLOC_STAGING_BYTES (1 MiB) inside the library, which every
 *     GPU that loads the library holds until the process exits.  num_envs * act_dim * real_bytes must fit:
 *     262144 envs (f32, forced breathing) ... 65536 envs (f64, manual breathing); more is rejected with
 *     SALP_EUNSUPPORTED and a message naming the limit -- use salp_step_host then;
 *   - everything runs on `caller_stream` alone (one H2D copy, one step kernel, up to four D2H copies, one
 *     cudaStreamSynchronize): no cudaMalloc / cudaFree, no stream or event creation, no allocating library call.
 * The first call on a device resolves the array's device address (cudaGetSymbolAddress, once per device).
 * Every call -- on any handle of the process -- shares the one array per GPU, and it keeps the last call's actions:
 * nothing in it has to start at zero (the kernel reads exactly the records the call has just copied in), but two
 * calls whose GPU work overlaps (different streams or threads on the same GPU) race on it.  Not for them.
 */
#define SALP_NOALLOC_STAGING_BYTES (1 << 20)
int salp_step_host_noalloc(SalpHandle *h, const void *actions_host, float *obs_host, void *reward_host,
                           uint8_t *terminated_host, uint8_t *truncated_host, void *caller_stream);

/*
 * Fill `rows` rows of `len` raw MT19937 outputs into HOST memory dst[rows][len]:
 *   kind 0: the stream of Python's random.seed(seed0 + first_row + r)     (init_by_array)
 *   kind 1: the stream of numpy.random.seed(seed0 + first_row + r)        (init_genrand)
 * Pure host code (multi-threaded); seeds must be < 2^32.
 */
int salp_fill_rng_words(uint32_t *dst, int64_t rows, int64_t len, uint64_t seed0, int64_t first_row,
                        int kind, int num_threads);

/*
 * Rollout -> replay, on the device.  Replaces the per-step
 *     replay_buffer.add(obs, action, reward, next_obs, done or truncated)
 * of the reference's episode loops (src/salp/training/continuous_trainer.py:296-303,
 * trainer.py:203-216) for all N envs of one step: appends N transitions to a caller-owned f32
 * ring buffer at rows (*head + e) mod capacity and advances *head (and *size, saturating).
 * next_obs of a lane whose episode ended is taken from `final_obs` (the terminal observation,
 * as the reference -- which never auto-resets -- would store); pass NULL without auto-reset.
 * With reward_aux != NULL the stored reward is w_env * reward + w_aux * reward_aux (the GAIL
 * mixing of src/salp/config/base_config.py:77-78).  `head`/`size` are DEVICE int64 scalars so the
 * call is CUDA-graph capturable.  Independent of any SalpHandle; asynchronous on `cuda_stream`.
 */
typedef struct SalpReplayAppend {
    const float *obs_prev;      /* [n][obs_dim]  observation the actions were computed from */
    const void *actions;        /* [n][act_dim]  real (dtype)                               */
    const void *reward;         /* [n]           real (dtype)                               */
    const float *reward_aux;    /* [n] f32 or NULL                                          */
    const uint8_t *terminated;  /* [n]                                                      */
    const uint8_t *truncated;   /* [n]                                                      */
    const float *obs_next;      /* [n][obs_dim]  obs returned by the step                   */
    const float *final_obs;     /* [n][obs_dim]  or NULL                                    */
    float *rb_obs, *rb_act, *rb_rew, *rb_next; /* ring: [C][obs_dim], [C][act_dim], [C], [C][obs_dim] */
    uint8_t *rb_done;           /* ring: [C]                                                */
    int64_t *head;              /* device scalar                                            */
    int64_t *size;              /* device scalar or NULL                                    */
    uint64_t *ticket;           /* device scalar, 0 between calls, or NULL (then a second, one-thread
                                   launch advances head/size)                               */
    int64_t n, capacity;        /* 0 < n <= capacity                                        */
    int32_t obs_dim, act_dim;   /* obs_dim % 4 == 0                                         */
    int32_t dtype;              /* SALP_F32 / SALP_F64: type of actions and reward          */
    float w_env, w_aux;
    int32_t reserved_;
} SalpReplayAppend;

int salp_replay_append(const SalpReplayAppend *p, void *cuda_stream);

/*
 * The rollout loop's policy step, fused: actions = tanh(mean + exp(clamp(log_std,-20,2)) * noise) with
 * (mean, log_std) = MLP(obs), for the reference's SAC actor shape obs -> 256 -> 256 -> 2*act with ReLU
 * (record_model.py:27-53; `agent.select_action(obs)` at src/salp/training/continuous_trainer.py:293).
 * One sm_100a kernel: bf16 tcgen05 MMAs with the accumulator in TMEM, weights resident in shared
 * memory, f32 epilogues; `obs` is read in place, `actions` written in the env build's real type.
 *   w1_packed / w2_packed: bf16 weights [256][K] (K = obs_dim zero-padded to 32, resp. 256) re-laid as
 *   [K/8][256][8] (rollout.pack_weight) -- the tcgen05 K-major operand layout.  w3: f32 [2*act][256].
 *   noise: f32 [n][act] standard normal draws, or NULL for the deterministic action tanh(mean).
 * Restrictions: hidden == 256, obs_dim (+ extra_dim) <= 32, act_dim in 1..4; otherwise SALP_EUNSUPPORTED.
 * Model (A) is obs 24 / act 1; model (B) is obs 6 / act 3 with the Box [0,1] x [0,1] x [-1,1]
 * (salp_robot_env.py:51-56), for which SB3's SAC maps the squashed action to the Box
 * (`unscale_action`: low + 0.5 * (a + 1) * (high - low)) -- act_rescale / act_low / act_high below.
 */
typedef struct SalpActorForward {
    const float *obs;        /* [n][obs_dim] */
    const void *w1_packed;   /* bf16 [4][256][8]  */
    const float *b1;         /* [256] */
    const void *w2_packed;   /* bf16 [32][256][8] */
    const float *b2;         /* [256] */
    const float *w3;         /* [2*act_dim][256]  (mode 1: [1][256]) */
    const float *b3;         /* [2*act_dim]       (mode 1: [1]) */
    const float *noise;      /* [n][act_dim] or NULL */
    void *actions;           /* mode 0: [n][act_dim] real (dtype);  mode 1: f32 [n] GAIL reward */
    const void *extra;       /* [n][extra_dim] real (dtype) appended to obs as input columns, or NULL */
    uint64_t *rng_state;     /* device u64[4] {seed, draw counter, 0, 0} or NULL: with noise == NULL the kernel
                                draws the normals itself (Philox4x32-10 keyed by seed, counter (row, draw)) and
                                advances the draw counter -- graph-replay safe; both NULL = deterministic */
    int64_t n;
    int32_t obs_dim, act_dim, hidden, dtype;
    int32_t extra_dim;       /* obs_dim + extra_dim <= 32 */
    int32_t mode;            /* 0: actor (tanh-Gaussian action)
                                1: GAIL reward -log(1 - sigmoid(MLP(cat(obs, extra))) + 1e-8) for a discriminator
                                   obs+act -> 256 -> 256 -> 1 (src/salp/agents/discriminator.py:43-85) */
    int32_t act_rescale;     /* mode 0: nonzero = write low + 0.5 * (tanh(.) + 1) * (high - low) instead of tanh(.) */
    int32_t reserved_;
    float act_low[4], act_high[4];
} SalpActorForward;

int salp_actor_forward(const SalpActorForward *p, void *cuda_stream);

/* =====================================================================================================
 * Model (B): one env step = one breathing cycle of the 3-D Newton-Euler robot.
 *
 * Replaces, for a slab of n environments (reference tree, Python):
 *   SalpRobotEnv.reset           src/salp/environments/salp_robot_env.py:100-127  (+ generate_target_point :266-328)
 *   SalpRobotEnv.step            src/salp/environments/salp_robot_env.py:139-201  (+ _rescale_action :129-137,
 *                                _calculate_reward :203-242, _get_observation :400-420)
 *   Nozzle.solve_angles / get_nozzle_direction   src/salp/environments/robot.py:55-83, :115-130, :153-178
 *   Robot.set_control / step_through_cycle / step and everything they call   robot.py:335-777
 * The calls are stateless: every buffer is caller-owned device memory passed in SalpCycleBuffers.
 *
 * Layout (N = n, `real` = float or double by dtype):
 *   state  real [6][N][4]  rec0 (pos.x, pos.y, pos.z, prev_dist)          rec1 (vel.x, vel.y, vel.z, water volume)
 *                          rec2 (omega.x, omega.y, omega.z, episode return) rec3 (roll, pitch, yaw, body length)
 *                          rec4 (prev_I.xx, prev_I.yy, prev_I.zz, body width) rec5 (vel_world.x, .y, .z, 0)
 *                          velocities are body-frame (robot.py:247-255); vel_world is kept for the reward heading term
 *   ctrl   i32  [N][4]     (cycle | phase << 16, rng cursor, target.x f32 bits, target.y f32 bits)
 *   actions real [N][3]    (contraction, coast time, nozzle yaw) in the env's Box [0,1] x [0,1] x [-1,1]
 *   obs    f32  [N][6]     (pos - target).xy, vel.xy (body), yaw, omega.z        (salp_robot_env.py:413-420)
 *   reward real [N]; terminated, truncated u8 [N]; n_sub i32 [N] integrator substeps of this cycle (0..1450)
 *   final_obs f32 [N][6]   terminal observation of lanes that finished (valid where terminated|truncated)
 *   stats  f64  [8]        += episodes, sum return, sum cycles, targets reached, truncations, non-finite lanes,
 *                             env steps, integrator substeps
 *   words  u32  [table_rows][words_len]  np.random MT19937 outputs per env (salp_fill_rng_words kind 1); env e
 *                          reads row (env_id0 + e) % table_rows; two 53-bit draws per reset (target point)
 * Finished environments auto-reset inside the step when SALP_CYCLE_AUTORESET is set (the vector-env convention;
 * the reference env itself never auto-resets): `obs` then holds the first observation of the next episode.
 */
#define SALP_CYCLE_AUTORESET 1
#define SALP_CYCLE_OBS_DIM 6
#define SALP_CYCLE_ACT_DIM 3
#define SALP_CYCLE_MAX_SUBSTEPS 4096

typedef struct SalpCycleParams {
    double dry_mass, init_length, init_width, max_contraction, density; /* Robot.__init__ / set_environment, robot.py:196-222,279 */
    double nozzle_length1, nozzle_length2, nozzle_length3, nozzle_area, nozzle_mass; /* Nozzle.__init__ robot.py:19-25 */
    double width, height;     /* SalpRobotEnv.__init__ (pixels; tank_margin 50, 200 px per metre) :32-39,266-277 */
    int32_t max_cycles;       /* 500 (salp_robot_env.py:187) */
    int32_t max_substeps;     /* safety cap on the cycle loop for actions outside the Box; 0 = SALP_CYCLE_MAX_SUBSTEPS */
    int32_t flags;            /* SALP_CYCLE_AUTORESET */
    int32_t reserved_;
} SalpCycleParams;

typedef struct SalpCycleBuffers {
    void *state;          /* real [6][n][4] */
    void *ctrl;           /* i32  [n][4]    */
    float *obs;           /* f32  [n][6]    */
    void *reward;         /* real [n]       */
    uint8_t *terminated;  /* [n] */
    uint8_t *truncated;   /* [n] */
    int32_t *n_sub;       /* [n] or NULL */
    float *final_obs;     /* [n][6] or NULL */
    double *stats;        /* [8] or NULL */
    const uint32_t *words; /* [table_rows][words_len] */
    int64_t table_rows, words_len;  /* words_len a power of two */
    int64_t n;
    int64_t env_id0;      /* global index of env 0 of this slab (RNG row selection; multi-GPU slabs) */
    int32_t dtype;        /* SALP_F32 | SALP_F64 */
    int32_t reserved_;
} SalpCycleBuffers;

/* reset: envs with mask[e] != 0 (mask == NULL: all) draw a new target and return to rest; rewind_rng != 0 puts
 * their cursor back to 0 first (reset(seed=...) semantics).  Writes `obs`. */
int salp_cycle_reset(const SalpCycleParams *p, const SalpCycleBuffers *b, const uint8_t *mask, int rewind_rng, void *cuda_stream);
/* step: one breathing cycle per env for actions real [n][3]. */
int salp_cycle_step(const SalpCycleParams *p, const SalpCycleBuffers *b, const void *actions, void *cuda_stream);

/*
 * Replaces the per-episode bookkeeping of the reference's loops -- score, steps and food of every finished episode
 * (src/salp/training/continuous_trainer.py:287-327, evaluation :343-372) -- in aggregate: copies the eight doubles the
 * step kernels accumulate in SalpBuffers.stats (episodes, sum return, sum length, sum food, collisions, truncations,
 * non-finite lanes, env-steps) to HOST memory `out8_host`, after the work already queued on `cuda_stream`, and returns
 * when they are there (it synchronises that stream; the only entry point besides the *_host steps that does).
 * reset != 0 zeroes the device vector afterwards.  SALP_EUNBOUND when no stats buffer is bound.  
 */
int salp_stats(SalpHandle *h, double *out8_host, int reset, void *cuda_stream);

/* Number of kernels this library has launched on behalf of `h` (bench "gpu_launches"). */
int64_t salp_launch_count(const SalpHandle *h);

#ifdef __cplusplus
}
#endif
#endif /* SALP_B200_H_ */
