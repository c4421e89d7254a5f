/* jaco_oracle.h -- CPU ORACLE (test infrastructure, NOT the product).
 *
 * A plain-C restatement of the reference's env-step path:
 *   env_script/env_mujoco.py:116-150  (step / terminal_inspection)
 *   env_script/env_mujoco_util.py     (reset, take_action, _step_simulation, obs, reward, touch, terminal)
 *   + the un-vendored dependencies it calls, restated from their published algorithms:
 *     MuJoCo 2.0 mj_step and abr_control OSC / transformations (App. A.1, A.2).
 *
 * PARITY UNPINNED: mujoco-py / MuJoCo / abr_control cannot be installed in this environment and the reference
 * ships no tests or golden vectors, so this oracle is checked only against (a) analytic
 * invariants, (b) a second independent numpy restatement of the smooth dynamics and of Gohlke's Euler
 * conventions (tests/), and (c) the weak real-MuJoCo recordings in models_baseline/trajectories.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may call this.
 */
#ifndef JACO_ORACLE_H
#define JACO_ORACLE_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

enum { JO_TASK_REACHING = 0, JO_TASK_GRASPING = 1, JO_TASK_PICKING = 2, JO_TASK_PLACING = 3, JO_TASK_PICKANDPLACE = 4, JO_TASK_RELEASING = 5, JO_TASK_CARRYING = 6 };
enum { JO_VAR_BASE = 0, JO_VAR_CIS = 1, JO_VAR_TRANS = 2 };

#define JO_NQ 23
#define JO_NV 21
#define JO_NU 9
#define JO_OBS 26
#define JO_AUX 64
#define JO_MAXCON 96
#define JO_MAXEFC 1024
#define JO_GRASP_RESET_MAX 4000 /* bound (substeps) of each of the reference's two unbounded grasping-reset servo loops (util:135-167) */

typedef struct jo_env jo_env;

jo_env *jo_create(const double *blob, int nblob, int task, int variant, uint64_t seed, int env_id);
void jo_destroy(jo_env *e);
/* env surface (env_mujoco.py:99,116) */
void jo_reset(jo_env *e, float *obs);
void jo_step(jo_env *e, const double *action, float *obs, double *reward, int *done, int *succ, double *wb);
/* state injection for parity tests */
void jo_get_state(const jo_env *e, double *qpos, double *qvel, double *aux);
void jo_set_state(jo_env *e, const double *qpos, const double *qvel, const double *aux);
/* one raw physics substep with explicit ctrl (mj_step); fills debug outputs (nullable) */
void jo_mj_step(jo_env *e, const double *ctrl, int *ncon, double *sensordata);
/* mj_forward at the current (qpos,qvel,ctrl) without integrating; exposes derived quantities (nullable ptrs) */
void jo_forward(jo_env *e, const double *ctrl, double *M21x21, double *bias21, double *J6x6, double *ee_pos,
                double *ee_quat, double *qacc21, int *ncon, double *sensordata);
/* contacts of the last forward: rows of [pos3 normal3 dist geom1 geom2 normalforce] */
int jo_get_contacts(const jo_env *e, double *out10, int maxcon);
/* abr_control OSC.generate on explicit inputs (App. A.2) */
void jo_osc(const jo_env *e, const double *J6x6, const double *M6x6, const double *bias6, const double *ee_pos,
            const double *ee_quat, const double *dq6, const double *target6, double *u6);
/* parity-test probe: 1 when a touch-sensor attribution of the last forward flips if the site is resized by +-2 um */
int jo_touch_knife_edge(const jo_env *e);
/* parity-test probe: an OSC call since the last clear sat within 2 % of the |det| < 1e-3 branch threshold or had an eigenvalue within
   30 % of the rcond = 1e-4 cut-off of numpy.linalg.pinv (process-wide flag) */
int jo_osc_knife_edge(void);
void jo_osc_knife_edge_clear(void);
/* also report (through jo_touch_knife_edge) whether a geom pair is within 2 um of making / breaking contact in the last forward */
void jo_set_probe(jo_env *e, int on);
/* solver statistics of the last forward */
void jo_solver_stats(const jo_env *e, int *iters, double *gradnorm, int *nefc);
/* Philox4x32-10 uniforms, exposed so the tests can pin the generator */
void jo_philox(uint64_t seed, uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t out[4]);
int jo_sizeof_real(void);

#ifdef __cplusplus
}
#endif
#endif
