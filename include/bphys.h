/* bphys.h — C ABI of the B200 batched physics backend (libbphys.so).
 *
 * Drop-in boundary for the one hot path of mhauskn/dm_control: the MuJoCo C
 * functions that `dm_control.mujoco.Physics` binds with ctypes
 * (/root/reference/dm_control/mujoco/wrapper/util.py:107-120 loads the library;
 * the call sites are cited per entry point below).  Plain C symbols, plain
 * pointers and sizes, no C++ or torch types, no exceptions across the boundary.
 * Every function returning `int` returns 0 on success, a negative BP_E* code
 * otherwise; the message is available from bp_last_error() (thread local).
 *
 * Where the reference has one mjData, this library has a *batch*: `nenv`
 * independent environments whose mjData-like "images" live in device memory as a
 * row-major [nenv][stride] array of doubles (ints packed at the tail).  Field
 * views (qpos, xpos, sensordata, ...) are strided windows of that one array —
 * the analogue of the reference's zero-copy numpy views of mjData
 * (wrapper/util.py:171-221).
 */
#ifndef BPHYS_H_
#define BPHYS_H_
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct bp_model bp_model;
typedef struct bp_batch bp_batch;

enum { BP_OK = 0, BP_EINVAL = -1, BP_ECUDA = -2, BP_ENOMEM = -3, BP_ENOFIELD = -4,
       BP_ESMEM = -5 };

/* mj_version (wrapper/core.py:65-67). */
int bp_version(void);
const char* bp_last_error(void);
/* number of CUDA devices visible; <= 0 means this library cannot run */
int bp_device_count(void);

/* mj_loadXML / mj_loadModel (wrapper/core.py:312-328, 460-468): the model arrives
 * already compiled, as the self-describing blob of dm_control_b200/model.py.
 * NULL + message in `err` on failure, like the reference's loader. */
bp_model* bp_model_create(const void* blob, int64_t nbytes, int device, char* err,
                          int errsz);
/* mj_deleteModel (wrapper/core.py:262-309). */
void bp_model_free(bp_model* m);
/* Run-time model edits (wrapper/core.py:581-620 `disable`, model.opt.*, and the
 * per-episode writes of suite/finger.py:140-142).  `name` is an mjModel field
 * name; `data` holds the whole field (int32 or float64 as in the blob). */
int bp_model_set_field(bp_model* m, const char* name, const void* data, int64_t nbytes);
int bp_model_get_field(bp_model* m, const char* name, void* data, int64_t nbytes);

/* mj_makeData (wrapper/core.py:637-651) for nenv environments.  njmax/nconmax <= 0
 * take the model's values.  `lanes` is the execution shape: GPU lanes cooperating
 * on one environment (1, 2, 4, 8, 16 or 32; <= 0 = default).  32 keeps the
 * environment's image in shared memory (one warp per environment) and fails when
 * the image does not fit an SM; smaller values address the image in global memory
 * with 32/lanes environments per warp (lanes == 1 interleaves the images of the 32
 * environments of a warp, see bp_batch_interleave). */
bp_batch* bp_batch_create(bp_model* m, int64_t nenv, int njmax, int nconmax, int lanes,
                          char* err, int errsz);
/* mj_deleteData. */
void bp_batch_free(bp_batch* b);

/* Device pointer to the [nenv][stride] image and its geometry. */
void* bp_batch_image(bp_batch* b);
int64_t bp_batch_stride(bp_batch* b);          /* doubles per environment        */
int64_t bp_batch_nenv(bp_batch* b);
int bp_batch_lanes(bp_batch* b);
/* 1: env e's element i is image[e*stride + i]; 32: image[(e/32)*stride*32 + i*32 + e%32]
 * (int32 fields: same rule in 4-byte units from the start of the int region). */
int bp_batch_interleave(bp_batch* b);
int64_t bp_batch_smem_bytes(bp_batch* b);      /* shared memory per environment  */
/* Enumerate / look up fields of the image: byte offset inside one env's image,
 * element count, dtype (0 = float64, 1 = int32).  Names follow mjData. */
int bp_field_count(bp_batch* b);
const char* bp_field_name(bp_batch* b, int index);
int bp_field_info(bp_batch* b, const char* name, int64_t* offset_bytes, int64_t* count,
                  int* dtype);

/* Host <-> device copies of one field for envs [env0, env0+n): `host` is a dense
 * row-major [n][count] array of the field's dtype. */
int bp_field_write(bp_batch* b, const char* name, const void* host, int64_t env0, int64_t n);
int bp_field_read(bp_batch* b, const char* name, void* host, int64_t env0, int64_t n);

/* Device <-> image copy of one field: `dev` is a dense [nenv][count] device array
 * (e.g. a torch tensor); asynchronous on the batch's stream. */
int bp_field_copy_device(bp_batch* b, const char* name, void* dev, int to_image);

/* Per-env height-field elevation data, the batched form of the run-time edit of
 * `model.hfield_data` in suite/quadruped.py:371-382 (Escape.initialize_episode writes a
 * new terrain every episode).  `host` is a dense [n][nhfielddata] float32 array
 * (mjModel.hfield_data is float32, normalised to [0, 1]) for envs [env0, env0+n). */
int bp_hfield_write(bp_batch* b, const float* host, int64_t env0, int64_t n);
int bp_hfield_read(bp_batch* b, float* host, int64_t env0, int64_t n);

/* Per-env copies of model arrays that tasks edit per episode
 * (`physics.named.model.geom_pos['target', 'x'] = ...`, suite/reacher.py:96-101;
 * `site_pos['target']`, suite/finger.py:180-181; `body_pos[target]` / `body_quat[target]`,
 * suite/manipulator.py:218-224; `model.wrap_prm[[0, 1]] = dir1`, suite/point_mass.py:85-101):
 * name is "geom_pos", "site_pos", "body_pos", "body_quat" or "wrap_prm", `host` a dense
 * [n][3 * ngeom | 3 * nsite | 3 * nbody | 4 * nbody | nwrap] float64 array for envs
 * [env0, env0+n).  Every env starts from the model's values; the copy is made on the first
 * write. */
int bp_env_model_write(bp_batch* b, const char* name, const double* host, int64_t env0,
                       int64_t n);
int bp_env_model_read(bp_batch* b, const char* name, double* host, int64_t env0, int64_t n);

/* Optional per-env mask (device or host bytes, nenv long; NULL = all envs): envs
 * whose byte is 0 are left untouched by the calls below. */
int bp_set_mask(bp_batch* b, const uint8_t* mask_host);

/* Physics.step() repeated nsub times (engine.py:149-166 inside the sub-step loop of
 * rl/control.py:100-106): Euler models mj_step2 + mj_step1, RK4 models mj_step +
 * mj_step1.  Asynchronous on the batch's stream. */
int bp_step(bp_batch* b, int nsub);
/* mj_step (forward + integrate), nsub times. */
int bp_mj_step(bp_batch* b, int nsub);
/* mj_forward (engine.py:297-305); disable_actuation mirrors the
 * `model.disable('actuation')` context of engine.py:287-295. */
int bp_forward(bp_batch* b, int disable_actuation);
int bp_step1(bp_batch* b);                      /* mj_step1 (engine.py:166) */
int bp_step2(bp_batch* b);                      /* mj_step2 (engine.py:162) */
/* mj_resetData / mj_resetDataKeyframe (engine.py:268-285) for the masked envs:
 * zero the image, qpos = qpos0 (or key `key` when >= 0).  Does not call forward. */
int bp_reset(bp_batch* b, int key);

/* Suite observation / reward kernels (suite/<task>.py get_observation/get_reward).
 * bp_set_task binds a task kind (ids in dm_control_b200/suite/tasks.py) with the
 * object ids (`ip`, <= 24) and constants (`fp`, <= 8) it needs, resolved by name on
 * the host.  obs is [nenv][nobs] float64 and reward [nenv] float64, device memory. */
int bp_set_task(bp_batch* b, int kind, const int32_t* ip, int nip, const double* fp, int nfp);
int bp_task_nobs(bp_batch* b);
int bp_obs_reward(bp_batch* b, double* obs_dev, double* reward_dev);
/* same, results copied to host buffers (synchronises) */
int bp_obs_reward_host(bp_batch* b, double* obs_host, double* reward_host);

/* rewards.tolerance (utils/rewards.py:93-135) evaluated on the device for host arrays;
 * sigmoid: 0 gaussian, 1 linear, 2 quadratic (the kinds the served tasks use). */
int bp_rewards_tolerance(int device, const double* x_host, double* out_host, int64_t n,
                         double lower, double upper, double margin, int sigmoid,
                         double value_at_margin);

/* bp_rewards_tolerance for callers that may not allocate device memory (code under CUDA
 * graph capture, tight serving loops, harnesses that forbid cudaMalloc).  NO counterpart in
 * the reference.  Same arguments and results
 * as bp_rewards_tolerance, but the device staging of x / out lives in one file-scope device
 * array (csrc/bp_noalloc.cu) instead of a cudaMalloc per call: it calls no cudaMalloc /
 * cudaFree / allocating library routine.  Limit: n <= BP_REWARDS_TOLERANCE_NOALLOC_MAX
 * (65536 elements; larger n returns BP_EINVAL with a message); every GPU that loads the
 * library holds the 1 MiB array until exit.  All calls share the array and it keeps the
 * previous call's contents (every element a call reads is written by that call first, so
 * nothing needs clearing); calls are serialised inside the library, and anything else that
 * would overlap with one on the GPU races. */
#define BP_REWARDS_TOLERANCE_NOALLOC_MAX 65536
int bp_rewards_tolerance_noalloc(int device, const double* x_host, double* out_host, int64_t n,
                                 double lower, double upper, double margin, int sigmoid,
                                 double value_at_margin);
int bp_rewards_tolerance_noalloc_max(void);   /* = BP_REWARDS_TOLERANCE_NOALLOC_MAX */

/* Fused Environment.step on DEVICE buffers, ONE kernel launch: scatter ctrl_dev
 * [nenv][nu] into the images, nsub physics steps, then the bound task's observation
 * ([nenv][nobs], may be NULL) and reward ([nenv]).  Asynchronous on the batch's stream. */
int bp_env_step(bp_batch* b, int nsub, const double* ctrl_dev, double* obs_dev,
                double* reward_dev);

/* End-to-end Environment.step for HOST buffers (rl/control.py:94-123): copy ctrl
 * [nenv][nu] to the device, nsub physics steps, observation/reward kernel, copy
 * obs/reward back and synchronise.  This is the call the e2e benchmark times. */
int bp_env_step_host(bp_batch* b, int nsub, const double* ctrl_host, double* obs_host,
                     double* reward_host);

/* 1 when the batch's last step / forward launch was served by a static (ahead-of-time
 * specialised) instance of the kernel, 0 when by a generic one.  Diagnostic: no reference
 * counterpart (the reference has one engine build). */
int bp_batch_last_static(bp_batch* b);

/* mj_copyData (wrapper/core.py:653-685 MjData.copy; engine.py:332-339 Physics.copy): the whole
 * state of `src` -- every env image, per-env height fields and model frames, the bound task --
 * into `dst`, created with the same sizes.  Device-to-device; synchronises. */
int bp_batch_copy(bp_batch* dst, bp_batch* src);

/* mj_contactForce as MjData.contact_force wraps it (wrapper/core.py:747-769): wrench of
 * contact `contact_id` of every env in the contact's frame, out_host [nenv][6] =
 * (normal, tangent, tangent, torsional, roll, roll).  Rows of envs that hold fewer contacts
 * than contact_id + 1 are NaN (the reference raises for such an id).  Synchronises. */
int bp_contact_force(bp_batch* b, int contact_id, double* out_host);
/* mj_objectVelocity as MjData.object_velocity wraps it (wrapper/core.py:713-745): 6D velocity
 * (angular, linear -- MuJoCo's order) of an object in a frame centred on it, world-aligned or
 * object-aligned (flg_local); objtype is an mjtObj value: 1 body (inertial frame), 2 xbody,
 * 5 geom, 6 site.  out_host [nenv][6].  Synchronises. */
int bp_object_velocity(bp_batch* b, int objtype, int objid, int flg_local, double* out_host);

/* Physics.check_invalid_state (engine.py:307-330): per-env warning counters
 * [nenv][8] int32 copied to the host (synchronises). */
int bp_warnings(bp_batch* b, int32_t* out_host);
int bp_sync(bp_batch* b);
/* DFMA saturation microbenchmark: measured FP64 vector peak of `device` in TFLOP/s
 * (the roofline denominator the reference's FP64 pipeline is judged against). */
int bp_fp64_peak(int device, double* tflops);
void* bp_batch_stream(bp_batch* b);             /* cudaStream_t of the batch */
/* milliseconds spent by the device in the last `n` bp_step-family launches is not
 * tracked here; time with CUDA events on bp_batch_stream(). */

#ifdef __cplusplus
}
#endif
#endif /* BPHYS_H_ */
