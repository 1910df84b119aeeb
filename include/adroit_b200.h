/* adroit_b200.h — C ABI of the batched Adroit step path.
 *
 * This is the drop-in boundary: the "operator" the reference's task layer sits on.
 * The reference reaches its physics through the mujoco-py object API in-process
 * (MjSim.step/forward/reset/get_state/set_state, data.qpos/qvel/ctrl/site_xpos/body_xpos/sensordata,
 * model.body_pos/site_pos/body_quat); every entry point below cites the reference call site it
 * replaces (paths relative to /root/reference/mj_envs_vision/hand_manipulation_suite/).
 *
 * Conventions
 *   - plain C, no exceptions: every call returns 0 on success, non-zero on error;
 *     adroit_last_error() returns a thread-local message.
 *   - the caller owns every I/O buffer.  For the CUDA library (libadroit_b200.so) `*_dev`
 *     arguments are DEVICE pointers, row-major [n_envs, dim]; `*_host` arguments are host pointers.
 *     The CPU oracle (oracle/liboracle.so) exports the same names with an `oracle_` prefix and takes host
 *     pointers everywhere.
 *   - all work is enqueued on the caller's CUDA stream (`stream` is a cudaStream_t passed as void*;
 *     NULL = default stream); no hidden synchronisation except where stated.
 *   - one host thread per handle; handles on different GPUs are independent.
 */
#ifndef ADROIT_B200_H
#define ADROIT_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct adroit_sim adroit_sim;

/* flags for adroit_create */
#define ADROIT_F64            1u   /* compute in fp64 instead of fp32 (the "fp64 build" of north_star) */
#define ADROIT_AUTO_RESET     2u   /* reset an env inside step() when its episode reaches the horizon */
#define ADROIT_RESET_ON_DONE  4u   /* with AUTO_RESET: also reset when the task reports done (pen drop, pen_v0.py:94-98) */

/* Build a simulator for `n_envs` independent environments from a compiled model blob
 * (mj_envs_b200/mjcf.py::Model.to_blob).  Replaces MujocoEnv.__init__ -> load_model_from_path -> MjSim
 * (hammer_v0.py:20, door_v0.py:22, pen_v0.py:27, relocate_v0.py:17).
 * `env_id_offset` is the global id of env 0 of this handle (multi-GPU sharding: the reset RNG stream is a
 * function of (seed, global env id, episode index) only, so results do not depend on the GPU count). */
int adroit_create(const void* model_blob, size_t blob_len, int n_envs, int device_id, uint64_t seed,
                  int64_t env_id_offset, uint32_t flags, adroit_sim** out);
void adroit_destroy(adroit_sim* sim);
const char* adroit_last_error(void);

/* sizes: nq(=nv), nu, obs_dim, npar, horizon, frame_skip, nbody, nsite (8 ints). */
int adroit_sizes(const adroit_sim* sim, int* out8);

/* env.reset(): mjrl MujocoEnv.reset -> sim.reset(); sim.forward(); reset_model()
 * (hammer_v0.py:106-132, door_v0.py:103-119, pen_v0.py:115-132, relocate_v0.py:85-103).
 * qpos=qvel=warmstart=0, per-env model parameters drawn from the counter-based RNG, mj_forward,
 * observation written.  `mask_dev` (u8[n], NULL = all) selects the envs to reset; `obs_dev` may be NULL. */
int adroit_reset(adroit_sim* sim, const uint8_t* mask_dev, float* obs_dev, void* stream);

/* env.step(a): clip + scale action, do_simulation(ctrl, frame_skip), get_obs, reward, goal_achieved, done
 * (hammer_v0.py:54-90, door_v0.py:55-85, pen_v0.py:65-102, relocate_v0.py:46-72).
 * action [n,nu] f32 in [-1,1]; obs [n,obs_dim] f32; reward [n] f32; done/goal [n] u8. */
int adroit_step(adroit_sim* sim, const float* action_dev, float* obs_dev, float* reward_dev,
                uint8_t* done_dev, uint8_t* goal_dev, void* stream);

/* Host-buffer convenience used by the reference-facing plugin: copies the action H2D, steps, copies the
 * four results D2H and synchronises `stream` (the copies are issued per sub-batch of envs, each on the stream
 * of its kernels, so they overlap the other sub-batches' compute).  Host pointers should be pinned for full speed. */
int adroit_step_host(adroit_sim* sim, const float* action_host, float* obs_host, float* reward_host,
                     uint8_t* done_host, uint8_t* goal_host, void* stream);

/* get_obs() (hammer_v0.py:92-104, door_v0.py:87-101, pen_v0.py:104-113, relocate_v0.py:74-83): packs the observation
 * from qpos / qvel and the task frames cached by the LAST forward.  Runs no forward itself -- like the reference, whose
 * get_obs reads data.site_xpos as mj_step / sim.forward() left it; call adroit_forward first where the reference does
 * (hammer_v0.py:131). */
int adroit_get_obs(adroit_sim* sim, float* obs_dev, void* stream);

/* sim.forward() = mj_forward on the current state of every env (hammer_v0.py:131,152; door_v0.py:110,138;
 * pen_v0.py:123,152; relocate_v0.py:94,129): refreshes frames, contacts, constraint forces and the cached task frames.
 * Enqueued on `stream`, no synchronisation. */
int adroit_forward(adroit_sim* sim, void* stream);

/* get_env_state / set_env_state payload (hammer_v0.py:134-153, door_v0.py:121-138, pen_v0.py:134-152,
 * relocate_v0.py:105-129): qpos [n,nq], qvel [n,nv], par [n,npar] (the per-env model fields reset_model
 * writes: board_pos | door_body_pos | desired_orien quat | obj body_pos + target site_pos), fp64 host buffers.
 * set_state also zeroes nothing else (qacc_warmstart is kept, as MjSim.set_state does) and runs mj_forward.
 * Both synchronise the stream. Any pointer may be NULL to skip that field. */
int adroit_get_state(adroit_sim* sim, double* qpos_host, double* qvel_host, double* par_host, void* stream);
int adroit_set_state(adroit_sim* sim, const double* qpos_host, const double* qvel_host, const double* par_host,
                     void* stream);
/* Episode step counter of every env (gym's TimeLimit._elapsed_steps, max_episode_steps of mj_envs_vision/__init__.py:4-29):
 * t [n] i32, host.  With ADROIT_AUTO_RESET an env resets when its counter reaches the horizon, so staggered counters
 * spread the resets of a batch over the episode (bench.py de-phases its batches with it).  Synchronises. */
int adroit_set_elapsed_steps(adroit_sim* sim, const int32_t* t_host, void* stream);
/* qacc_warmstart import/export [n,nv] (parity tests start both implementations from identical solver state). */
int adroit_get_warmstart(adroit_sim* sim, double* w_host, void* stream);
int adroit_set_warmstart(adroit_sim* sim, const double* w_host, void* stream);

/* data.body_xpos / body_xquat / site_xpos as the last forward left them (hammer_v0.py:141, relocate_v0.py:112-114 read
 * them in get_env_state WITHOUT a forward: after a step they lag qpos by one substep); fp64 host buffers
 * [n,nbody,3], [n,nbody,4], [n,nsite,3]; NULL to skip.  A pure read: no forward, no side effect.  Synchronises. */
int adroit_get_frames(adroit_sim* sim, double* body_xpos_host, double* body_xquat_host, double* site_xpos_host,
                      void* stream);

/* Solver introspection, fp64 host buffers (NULL to skip): data.qacc [n,nv], ncon [n] (i32), nefc [n] (i32) and
 * sensordata[S_nail] [n] (hammer_v0.py:102) as the last forward left them.  A pure read.  Synchronises. */
int adroit_get_solver(adroit_sim* sim, double* qacc_host, int32_t* ncon_host, int32_t* nefc_host,
                      double* touch_host, void* stream);

/* End-of-rollout statistics (eval.py:69-97 semantics and evaluate_success, hammer_v0.py:167-175):
 * out4 = { sum of finished-episode returns, #episodes with any goal_achieved,
 *          #episodes with sum(goal_achieved) > thr (25; pen 20), #finished episodes } as fp64 (host).
 * An episode finishes when an env is reset (explicitly or by AUTO_RESET).  `flush_running` != 0 also
 * counts the episodes in flight.  Synchronises; the caller all-reduces the 4 numbers across ranks. */
int adroit_reduce_episode(adroit_sim* sim, int flush_running, double* out4_host, void* stream);

/* adroit_reduce_episode for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid allocation).  NO COUNTERPART IN THE REFERENCE.
 * Same four statistics as adroit_reduce_episode, but the call allocates nothing, copies nothing to the host and does
 * not synchronise: it enqueues two kernels on `stream` that leave the result in the CALLER's device buffer
 * out4_dev (4 x fp64).  The partial sums live in a file-scope `static __device__` array of the library
 * (ADROIT_NOALLOC_MAX_ENVS / 256 rows of 4 doubles = 8 KB, held by every GPU that loads the library until exit);
 * a handle with more than ADROIT_NOALLOC_MAX_ENVS envs is rejected with an error, nothing is launched.
 * The array is shared by EVERY call on a device (all handles, all streams) and keeps the last call's contents; each
 * call overwrites every row it reads, so nothing has to be cleared, but two calls that overlap on the GPU (different
 * streams, or graphs replayed concurrently) race on it -- order them.  The sum is taken in a fixed order, so the
 * result is bit-reproducible; it agrees with adroit_reduce_episode to fp64 rounding of the return sum (counts exact). */
#define ADROIT_NOALLOC_MAX_ENVS 65536
int adroit_reduce_episode_noalloc(adroit_sim* sim, int flush_running, double* out4_dev, void* stream);

/* env.render() (hammer_v0.py:177-181 -> HeadlessObserver.render, headless_observer.py:34-52, + PlaNet's 128->64 resize):
 * 64x64 RGB of every env from the reference's free camera, ray-cast against the scene's analytic primitives (no GL on
 * this box; see csrc/render.cuh for what is and is not reproduced).  rgb_dev: u8 [n, 64, 64, 3].  Draws the frames of the
 * last forward (as sim.render does).  Enqueued on `stream`, no synchronisation. */
int adroit_render(adroit_sim* sim, uint8_t* rgb_dev, void* stream);

/* Number of kernel launches this handle has enqueued so far (bench.py's gpu_launches). */
int64_t adroit_launch_count(const adroit_sim* sim);

/* Per-phase device timing (bench.py's roofline: duration of the dominant kernel measured with CUDA events on the
 * launching stream).  adroit_profile(sim, 1) starts recording an event pair around every phase-kernel launch,
 * adroit_profile(sim, 0) stops.  adroit_phase_times synchronises the stream and returns, accumulated since the last
 * call, ms4 = device time of {smooth, collide, solve, finish} kernels and n4 = their launch counts, then clears them. */
int adroit_profile(adroit_sim* sim, int enable);
int adroit_phase_times(adroit_sim* sim, double* ms4_host, int64_t* n4_host, void* stream);

/* Counters of capacity overflows since creation: out[0] = contacts dropped (per-env contact buffer full),
 * out[1] = constraint rows dropped.  Both must stay 0 for results to be valid. */
int adroit_overflow(adroit_sim* sim, int64_t* out2_host, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* ADROIT_B200_H */
