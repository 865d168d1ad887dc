/*
 * urgym_b200 — C ABI of the B200-native batched physics backend for the URGym step path.
 *
 * The reference has no FFI: its hot path sits behind the Gymnasium Env interface and, below that, behind the
 * module-level pybullet calls listed in SURVEY.md §2a. Each entry point below names the reference interface it
 * replaces (paths under /root/reference). Conventions: every function returns 0 on success or a negative URGYM_E_*
 * code (text via urgym_last_error); no exception crosses the boundary; all `*_dev` pointers are device memory owned
 * by the caller (e.g. torch tensors) on the handle's device; calls are asynchronous on `stream` (a cudaStream_t
 * passed as void*); one host thread per handle at a time; the library never frees caller buffers.
 *
 * There is no CPU fallback: every call fails with URGYM_E_CUDA when no sm_100 device is usable.
 */
#ifndef URGYM_B200_H
#define URGYM_B200_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct urgym_env_s* urgym_handle;

/* Scalar type of every floating-point buffer that crosses this boundary. The product library liburgym_b200.so computes
 * and exchanges fp32. The parity build liburgym_b200_f64.so (same sources compiled with -DURGYM_F64, same entry points)
 * computes and exchanges fp64; it exists so that integer outputs can be compared with the fp64 CPU oracle without fp32
 * threshold flips. urgym_real_bytes() tells a binding which of the two it has loaded. */
#ifdef URGYM_F64
typedef double urgym_real;
#else
typedef float urgym_real;
#endif
int urgym_real_bytes(void);

/* 5 and 6 are the reference's variants of CubesPush / Golf: same world, other observation / action mapping
 * (urgym/envs/env_cubes_push_diff_v0.py:209-211; urgym/envs/env_golf_joints_v0.py:60-63,85,123-124,229-234). */
enum { URGYM_BALL_BALANCE = 0, URGYM_TWO_BALLS_BALANCE = 1, URGYM_CUBES_PUSH = 2, URGYM_CUBES_GRASP = 3, URGYM_GOLF = 4,
       URGYM_CUBES_PUSH_DIFF = 5, URGYM_GOLF_JOINTS = 6,
       URGYM_BOX = 7, /* urgym/envs/env_box_v0.py + urgym/urdf/skew-box-button.urdf: hinged door + button (cylinders as square prisms) */
       URGYM_ENV_COUNT = 8 };
enum { URGYM_OK = 0, URGYM_E_ARG = -1, URGYM_E_CUDA = -2, URGYM_E_STATE = -3 };
enum { URGYM_STATE_FLOATS = 96, URGYM_MAX_CONTACTS = 32, URGYM_MAX_PAIRS = 160, URGYM_PAIR_WORDS = 5, URGYM_NUM_LINKS = 19, URGYM_NUM_DOF = 12 };

/* Static facts per env type: urgym/envs/__init__.py:10-44 (max_episode_steps) and each env's spaces, e.g.
 * urgym/envs/env_ball_balance_v0.py:52-54. */
int urgym_obs_dim(int env_type);
int urgym_act_dim(int env_type);
int urgym_max_episode_steps(int env_type);
int urgym_num_shapes(int env_type);
/* body: 0 robot, 1..3 free body, -1 static; link: Bullet link index on the robot or sub-shape id otherwise */
int urgym_shape_info(int env_type, int shape, int* body, int* link);

/* Replaces Env.__init__ + UR5Robotiq85.load (env_ball_balance_v0.py:34-80, urgym/base/robot.py:39-77,178-211):
 * builds `num_envs` independent worlds resident in HBM on `device`. `env_index_offset` is the global index of this
 * shard's first env (the RNG is keyed on seed, global env index, episode and draw, so results do not depend on the
 * shard layout). reward_sparse mirrors reward_type='sparse' (env_cubes_push_v0.py:159, env_golf_v0.py:151,163). */
int urgym_create(int env_type, int num_envs, int device, uint64_t seed, int env_index_offset, int reward_sparse,
                 urgym_handle* out);
int urgym_destroy(urgym_handle h);

/* Replaces Env.reset (env_ball_balance_v0.py:145-170 and the four siblings; robot.py:85-101).
 * mask_dev: uint8[N], nullable = all. draws_dev: float[N,4] uniforms in [0,1) replacing the reference's global-random
 * draws (env_ball_balance_v0.py:193-194 etc.), nullable = counter-based RNG. obs_dev: float[N,obs_dim].
 * nticks_dev: int32[N] physics ticks spent, nullable. */
int urgym_reset(urgym_handle h, const uint8_t* mask_dev, const urgym_real* draws_dev, urgym_real* obs_dev, int32_t* nticks_dev,
                void* stream);

/* Replaces Env.step (env_ball_balance_v0.py:105-134 and siblings): IK -> joint motors -> wait_until_stable (1..480
 * stepSimulation ticks) -> contact predicates -> reward / terminated / observation, plus the TimeLimit wrapper's
 * `truncated`, and, when autoreset != 0, the reset of finished envs (obs_dev then holds the new episode's first
 * observation and final_obs_dev the last one of the old episode).
 * action_dev float[N,act_dim]; u_dev float[N] selector uniform for CubesGrasp (env_cubes_grasp_v0.py:127-130),
 * nullable = RNG; draws_dev as in urgym_reset; outputs: obs float[N,obs_dim], reward float[N], terminated /
 * truncated / success uint8[N], final_obs float[N,obs_dim] (nullable), nticks int32[N,2] = {settle ticks of the
 * action, ticks of the auto-reset} (nullable). */
int urgym_step(urgym_handle h, const urgym_real* action_dev, const urgym_real* u_dev, const urgym_real* draws_dev, urgym_real* obs_dev,
               urgym_real* reward_dev, uint8_t* terminated_dev, uint8_t* truncated_dev, uint8_t* success_dev,
               urgym_real* final_obs_dev, int32_t* nticks_dev, int autoreset, void* stream);

/* Allocation-free step. No counterpart in the reference: added for callers that
 * cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it). One gym step of
 * `num_envs` envs as urgym_step does it, but without a handle: the caller passes the state records in and out
 * (state_in_dev / state_out_dev float[N,URGYM_STATE_FLOATS], as urgym_get_state returns them; the two may be the same
 * buffer) together with what urgym_create would have fixed (env type, seed, env_index_offset, reward_sparse), and the
 * working memory urgym_create would have allocated (scene model, finished-env flags, episode counters) lives in a static
 * device array inside the library. The call itself performs no cudaMalloc / cudaFree and calls no allocating library
 * routine. Limit: num_envs <= URGYM_STATIC_MAX_ENVS (16384); larger batches are rejected with URGYM_E_ARG. The first call
 * per device writes the scene models into the array with blocking copies (make it outside a capture). Every call on a device
 * shares the array: calls that overlap on the GPU (different streams) race on it; calls on one stream are safe.
 * stats_dev double[8] (nullable): the counters of urgym_stats for THIS call only (they are cleared at the start of each call).
 * Other arguments as in urgym_step. Results are bit-identical to urgym_step from the same state. */
enum { URGYM_STATIC_MAX_ENVS = 16384 };
int urgym_step_static(int env_type, int device, int num_envs, uint64_t seed, int env_index_offset, int reward_sparse,
                      const urgym_real* state_in_dev, const urgym_real* action_dev, const urgym_real* u_dev, const urgym_real* draws_dev,
                      urgym_real* state_out_dev, urgym_real* obs_dev, urgym_real* reward_dev, uint8_t* terminated_dev,
                      uint8_t* truncated_dev, uint8_t* success_dev, urgym_real* final_obs_dev, int32_t* nticks_dev,
                      double* stats_dev, int autoreset, void* stream);

/* Same call with HOST buffers: copies inputs host->device and results device->host on the handle's stream and
 * synchronises. This is the entry the single-env Gymnasium shim and bench.py's e2e leg use. */
int urgym_step_host(urgym_handle h, const urgym_real* action, const urgym_real* u, const urgym_real* draws, urgym_real* obs, urgym_real* reward,
                    uint8_t* terminated, uint8_t* truncated, uint8_t* success, urgym_real* final_obs, int32_t* nticks,
                    int autoreset);
int urgym_reset_host(urgym_handle h, const uint8_t* mask, const urgym_real* draws, urgym_real* obs, int32_t* nticks);

/* Snapshot / restore of the per-env state records, float[N,URGYM_STATE_FLOATS] (integers stored as exact floats).
 * Stands in for pybullet's saveState/restoreState, unused by the reference, needed for identical-initial-state runs. */
int urgym_get_state(urgym_handle h, urgym_real* state_dev, void* stream);
int urgym_set_state(urgym_handle h, const urgym_real* state_dev, void* stream);

/* Replaces getContactPoints as the envs use it (env_cubes_push_v0.py:177,186; env_cubes_grasp_v0.py:210;
 * env_golf_v0.py:187-206): the manifold points of the last tick of the last step, as shape-index pairs.
 * pairs_dev int32[N,URGYM_MAX_CONTACTS,2] (-1 padded), count_dev int32[N]. */
int urgym_contacts(urgym_handle h, int32_t* pairs_dev, int32_t* count_dev, void* stream);
/* The same query at the granularity the reference's predicates use ("is there any contact between A and B"): one bit per
 * candidate pair of the scene's static pair list, set when the pair carried at least one manifold point in the last tick
 * of the last step. mask_dev uint32[N,URGYM_PAIR_WORDS], bit p of word p/32 = pair p. Unlike the point list above it does
 * not depend on the kernel's manifold-point capacity (16 / 24 / 32 points per tick by scene; ticks that had more are
 * counted in urgym_stats out[6]). urgym_num_pairs / urgym_pair_info give the pair list (shape indices as in
 * urgym_shape_info). */
int urgym_contact_pairs(urgym_handle h, uint32_t* mask_dev, void* stream);
int urgym_num_pairs(int env_type);
int urgym_pair_info(int env_type, int pair, int* shape_a, int* shape_b);

/* Env-logic check entry: step() with the physics skipped. Runs the part of step() before wait_until_stable (action ->
 * IK / motor targets, written to the state record) and the part after it (contact predicates, reward, termination,
 * success, observation, step counter) on the CURRENT state, with the contact set the predicates see given by the caller
 * as shape-index pairs (pairs_dev int32[N,URGYM_MAX_CONTACTS,2], count_dev int32[N]) instead of produced by the
 * simulation. Exists so that the device env logic can be compared case by case with fixtures recorded from the
 * reference's own step() (tests/golden/reference_env_logic.json: env_*_v0.py step() and its predicates). */
int urgym_step_logic(urgym_handle h, const urgym_real* action_dev, const urgym_real* u_dev, const int32_t* pairs_dev, const int32_t* count_dev,
                     urgym_real* obs_dev, urgym_real* reward_dev, uint8_t* terminated_dev, uint8_t* truncated_dev, uint8_t* success_dev,
                     void* stream);

/* Raw stepSimulation: `nticks` physics ticks on every env with the current motor targets, no env logic. */
int urgym_tick(urgym_handle h, int nticks, void* stream);

/* Stand-alone kinematics on `n` joint vectors (no handle state is touched):
 * urgym_fk replaces getLinkState(robot, i)[4:6] for all 19 links: q float[n,12] -> pose float[n,19,7] (xyz, quat xyzw).
 * urgym_ik replaces calculateInverseKinematics as robot.py:119-121 calls it: q float[n,12], target float[n,7]
 * -> arm joint targets float[n,6].
 * urgym_gripper_center replaces get_gripper_geometrical_center (env_ball_balance_v0.py:249-275): -> float[n,3]. */
int urgym_fk(int device, int n, const urgym_real* q_dev, urgym_real* pose_dev, void* stream);
int urgym_ik(int device, int n, const urgym_real* q_dev, const urgym_real* target_dev, urgym_real* out_dev, void* stream);
int urgym_gripper_center(int device, int n, const urgym_real* q_dev, urgym_real* out_dev, void* stream);

/* Episode statistics accumulated on device since creation: {episodes, return sum, length sum, successes, env steps,
 * physics ticks, ticks whose manifold points or rolling-friction manifolds exceeded the kernel's capacity, reserved}.
 * Synchronises the handle's last stream. (Optional NCCL reduction happens in the Python host.) */
int urgym_stats(urgym_handle h, double out[8]);
/* Launch geometry of the fused env kernel: lanes of a warp that run one env, and shared-memory bytes per env. */
int urgym_kernel_info(int env_type, int* group_lanes, int* smem_bytes_per_env);
/* Development aid: per-phase warp cycles of the tick (fk, collide, dynamics, row setup, T build, sweeps, integrate)
 * accumulated since the last call; returns URGYM_E_STATE unless the library was built with -DUG_PHASE_TIMING. */
int urgym_phase_cycles(unsigned long long out[8]);
/* Kernel launches issued by this library in this process (bench.py's gpu_launches claim). */
long long urgym_launch_count(void);
const char* urgym_last_error(void);

#ifdef __cplusplus
}
#endif
#endif
