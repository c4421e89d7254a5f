/* jaco_b200.h -- C ABI of the B200 batched Jaco simulator (libjaco_b200.so).
 *
 * The reference has no FFI: its seam is the Python class JacoMujocoEnv
 * (/root/reference/env_script/env_mujoco.py:9-167).  Each entry point below replaces the part of that class
 * (and of the objects beneath it) named in its comment; mujoco_jaco_b200/env.py binds them with ctypes and
 * rebuilds the reference's Python surface on top.  See INTEGRATION.md for the reference-side stub.
 *
 * Conventions: return 0 on success, a negative JACO_E* code otherwise (never throws); every device pointer is
 * owned by the caller, contiguous fp32 / uint8 row-major with the leading dimension n_envs; the library owns
 * the SoA simulation state; all calls are asynchronous and ordered on `stream` (a cudaStream_t passed as
 * void*, NULL = legacy default stream); one host thread per handle.  There is no CPU fallback.
 */
#ifndef JACO_B200_H
#define JACO_B200_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define JACO_NQ 23
#define JACO_NV 21
#define JACO_NU 9
#define JACO_NOBS 26
#define JACO_NACT 7
#define JACO_NAUX 72
#define JACO_NSTATS 16

enum { JACO_OK = 0, JACO_EINVAL = -1, JACO_ECUDA = -2, JACO_ENOMEM = -3 };
/* task ids: argparser.py --task */
enum { JACO_REACHING = 0, JACO_GRASPING = 1, JACO_PICKING = 2, JACO_PLACING = 3, JACO_PICKANDPLACE = 4, JACO_RELEASING = 5, JACO_CARRYING = 6 };
/* variant = observation variant (0 env_mujoco_util.py, 1 _cis, 2 _trans) | obs_mode << 4 | binary << 8
 * obs_mode: 0 reaching goal (util:255-270), 1 rule-based subgoal (util:240-254, what main.py uses), 2 subgoal_obs
 * binary: the _cis / _trans sparse picking reward (env_mujoco_util_cis.py:56,459,637-660: reward 0, +10 on success, -1 on a drop);
 *         the reference never sets it (self.binary = False), it is reachable here for completeness */

typedef struct jaco_sim_handle* jaco_handle;

/* JacoMujocoEnvUtil.__init__ (util:15-70): model load + sim + controller, for n_envs envs on `device`.
 * model_blob: float64 array produced by mujoco_jaco_b200/model/compiler.py (include/jaco_model_layout.h).
 * env_offset: index of this handle's first env in the global env numbering (RNG streams; multi-GPU sharding). */
int jaco_create(const void* model_blob, size_t nbytes, int n_envs, int task_id, int variant, int device,
                uint64_t seed, int env_offset, jaco_handle* out);
void jaco_destroy(jaco_handle h);
const char* jaco_last_error(jaco_handle h);

/* JacoMujocoEnv.reset (env_mujoco.py:99-102 -> util:92-174).  mask: device uint8[n] or NULL (= all envs).
 * obs: device float[n,26] or NULL. */
int jaco_reset(jaco_handle h, const uint8_t* mask, float* obs, void* stream);

/* JacoMujocoEnv.step (env_mujoco.py:116-139): clip, take_action, 50 x (OSC + mj_step), obs, reward, terminal.
 * action float[n,7]; obs float[n,26]; reward float[n]; done/succ uint8[n]; wb float[n] (get_wb()).
 * auto_reset != 0: finished envs are re-initialised in the same launch and `obs` holds the new episode's
 * first observation (VecEnv convention); the terminal observation is written to term_obs if non-NULL. */
int jaco_step(jaco_handle h, const float* action, float* obs, float* reward, uint8_t* done, uint8_t* succ,
              float* wb, float* term_obs, int auto_reset, void* stream);

/* jaco_step without any allocation.  NO COUNTERPART IN THE REFERENCE: added for callers
 * that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  Same inputs, outputs
 * and kernels as jaco_step (bit-identical results); the step kernels' scratch (work lists + counters) lives in a file-scope
 * `static __device__` array of the library instead of the handle's cudaMalloc'ed buffers, resolved once per device at set-up
 * (jaco_create) with cudaGetSymbolAddress.  Limits: n_envs <= JACO_NOALLOC_MAX_ENVS = 16384 (larger handles are rejected with
 * JACO_EINVAL and a message in jaco_last_error); every GPU that loads the library holds the 128.25 KB array until exit; all calls on
 * a device share it, so calls that overlap on the GPU (different streams / handles) race -- serialise them.  The entry point calls
 * no cudaMalloc / cudaFree or allocating library routine; the handle itself (state, statistics) is allocated by jaco_create. */
#define JACO_NOALLOC_MAX_ENVS 16384
int jaco_step_noalloc(jaco_handle h, const float* action, float* obs, float* reward, uint8_t* done, uint8_t* succ,
                      float* wb, float* term_obs, int auto_reset, void* stream);

/* Same call with HOST buffers (the reference-facing path: numpy in, numpy out).  Copies the action to the
 * device, steps, copies the results back and synchronises the stream before returning. */
int jaco_step_host(jaco_handle h, const float* action, float* obs, float* reward, uint8_t* done, uint8_t* succ,
                   float* wb, int auto_reset, void* stream);

/* mjData state exchange (mujoco.py:217-246,332-359 teleports / MjSimState): qpos float[n,23], qvel float[n,21],
 * aux float[n,72] (env-logic state: target, gripper, goals, counters, warm start, lagged controller state).
 * Any pointer may be NULL. */
int jaco_get_state(jaco_handle h, float* qpos, float* qvel, float* aux, void* stream);
int jaco_set_state(jaco_handle h, const float* qpos, const float* qvel, const float* aux, void* stream);

/* One raw mj_step with explicit ctrl float[n,9] (mujoco.py:258-278 send_forces); debug outputs, nullable:
 * ncon int32[n], sensordata float[n,20], qacc float[n,21]. */
int jaco_debug_substep(jaco_handle h, const float* ctrl, int32_t* ncon, float* sensordata, float* qacc,
                       void* stream);

/* Episode statistics accumulated on the device since the last call with clear != 0: out float[16] =
 * {episodes, successes, return_sum, timeouts, drops, singular_exits, nan_resets, env_steps, length_sum, 0...}.
 * This is the vector the multi-GPU harness all-reduces (env_mujoco.py:129-135 success-rate log). */
int jaco_stats(jaco_handle h, float* out16, int clear, void* stream);

/* JacoMujocoEnvUtil._get_camera (util:302-307: offscreen.render(640, 480, camera_id=0) + read_pixels(depth=True), flipped top-down)
 * for envs [env_begin, env_begin + env_count): a headless CUDA ray caster over the model's analytic primitives (the reference's
 * OpenGL path is unavailable on this box; mujoco_jaco_b200/csrc/jaco_render.cuh says what is drawn).
 * camera19 (host): pos[3], R[9] (row-major; columns = camera right / up / backward axes in the world, MuJoCo convention: the camera
 * looks along its -z), fovy in degrees, znear, zfar (metres), ambient, 3 unused.  prim_rgb (host): float[JACO_RENDER_PRIMS][3].
 * Outputs on the device: rgb uint8[env_count][height][width][3] (row 0 = top), depth float[env_count][height][width]
 * (OpenGL-normalised like read_pixels: (1/znear - 1/z) / (1/znear - 1/zfar), 1.0 where nothing is hit). */
#define JACO_RENDER_PRIMS 35
int jaco_render(jaco_handle h, const float* camera19, const float* prim_rgb, int width, int height, int env_begin, int env_count,
                uint8_t* rgb, float* depth, void* stream);

/* diagnostics: host copy of the step work-list counters, out int32[16] = {n_bail, n_reset of the last step,
 * cumulative bail-reason histogram (limit, hand, forearm, object, dest, osc), 8..11: env placement of the step kernel:
 * envs that were quiet / busy without a cooperative solve / busy with one / at the top workload key in the last step};
 * synchronises the device. */
int jaco_debug_counters(jaco_handle h, int32_t* out16);
int jaco_n_envs(jaco_handle h);
/* number of kernels this handle has launched (bench.py "gpu_launches") */
long long jaco_launch_count(jaco_handle h);

/* FP32 FMA peak of `device` in TFLOP/s, measured with a register-resident FMA loop (roofline denominator). */
double jaco_fp32_peak_tflops(int device, int reps);

#ifdef __cplusplus
}
#endif
#endif
