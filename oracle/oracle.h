/* CPU ORACLE — TEST INFRASTRUCTURE ONLY.  Never imported by the product path.
 *
 * Plain-C fp64 restatement of the physics the reference delegates to the closed
 * MuJoCo 2.00 binary through Physics.step()/forward()/reset()
 * (/root/reference/dm_control/mujoco/engine.py:149-166,268-305).  The engine's
 * source is NOT part of the reference: the algorithms below are
 * restated from MuJoCo's published documentation ("Computation" chapter) and
 * are pinned by the reference's own known answers (tests/test_oracle_pins.py:
 * README free-fall value 0.19996362, touch = 9.81 N, contact-force = weight,
 * object velocities, accelerometer = +9.81, LQR-style checks).  Digit-level
 * parity with real MuJoCo trajectories is otherwise UNPINNED (no stored
 * trajectories in the reference, no engine binary here).
 *
 * One environment per OData; all intermediates are exposed by name through
 * od_field() so tests can compare every stage with the CUDA path.
 */
#ifndef BP_ORACLE_H_
#define BP_ORACLE_H_
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct OModel OModel;
typedef struct OData OData;

OModel* om_create(const void* blob, int64_t nbytes, char* err, int errsz);
void om_free(OModel* m);
/* model field lookup (by mjModel-style name): returns pointer into the model's
 * own copy of the blob, count in *count, dtype 0=f64 1=i32 in *dtype */
void* om_field(OModel* m, const char* name, int64_t* count, int* dtype);

OData* od_create(const OModel* m);
void od_free(OData* d);
void* od_field(OData* d, const char* name, int64_t* count, int* dtype);

void o_reset(const OModel* m, OData* d);              /* mj_resetData      */
void o_forward(const OModel* m, OData* d);            /* mj_forward        */
void o_step(const OModel* m, OData* d);               /* mj_step           */
void o_step1(const OModel* m, OData* d);              /* mj_step1          */
void o_step2(const OModel* m, OData* d);              /* mj_step2          */
/* Physics.step() of engine.py:149-166: Euler -> step2 ; RK4 -> step, then step1 */
void o_physics_step(const OModel* m, OData* d);

/* mj_contactForce / mj_objectVelocity (wrapper/core.py:713-769) */
void o_contact_force(const OModel* m, const OData* d, int contact_id, double* f6);
void o_object_velocity(const OModel* m, const OData* d, int objtype, int objid, double* res6,
                       int flg_local);

/* CPU-baseline driver: nenv independent environments, `nsteps` calls of
 * o_physics_step x nsub each, ctrl[nsteps or 1][nenv][nu]; state arrays are
 * [nenv][n] row-major and updated in place; nthreads pthreads. */
void o_run_batch(const OModel* m, int nenv, int nsteps, int nsub,
                 double* qpos, double* qvel, double* act, double* warm,
                 double* time, const double* ctrl, int ctrl_per_step,
                 int nthreads);

/* the same with per-env height-field elevation data [nenv][nhfielddata] (NULL: the
 * model's), the batched form of quadruped.py:381 */
void o_run_batch_hf(const OModel* m, int nenv, int nsteps, int nsub,
                    double* qpos, double* qvel, double* act, double* warm,
                    double* time, const double* ctrl, int ctrl_per_step,
                    int nthreads, const double* hfield);

#ifdef __cplusplus
}
#endif
#endif
