/*
 * oige_b200.h -- C ABI of the B200-native locomotion environment step.
 *
 * The reference (tzmhuang/OmniIsaacGymEnvs, OIGE v1.1.0) is pure Python: it has no FFI of its
 * own.  Each entry point below therefore cites the reference *Python* interface it replaces
 * (paths relative to /root/reference/omniisaacgymenvs).  The Python host side
 * (omniisaacgymenvs_b200/) binds these with ctypes -- see INTEGRATION.md for the stub a
 * maintainer of the reference would add.
 *
 * Conventions
 *   - every pointer in OigeState / OigeTaskBuffers / OigeResetInit is a DEVICE pointer owned by
 *     the caller (PyTorch allocates everything); contiguous row-major; fp32 unless noted;
 *     16-byte aligned base addresses (torch allocations are).
 *   - quaternions are (w, x, y, z); root_vel is (lin xyz, ang xyz); joints in breadth-first order
 *     (docs/transfering_policies_from_isaac_gym.md:26-54).
 *   - every call is asynchronous on `stream` (a cudaStream_t passed as void*); the library never
 *     synchronises the device and never allocates per step.
 *   - return value: 0 = ok, <0 = argument error (OIGE_E_*), >0 = cudaError_t of a failed launch.
 *     oige_last_error() returns a thread-local message for the last non-zero return.
 *   - there is NO CPU fallback: without a CUDA device every compute entry point fails.
 */
#ifndef OIGE_B200_H
#define OIGE_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define OIGE_ABI_VERSION 4
#define OIGE_MAX_DOF 32      /* one lane per link in the articulation kernel */
#define OIGE_MAX_LINKS 32
#define OIGE_MAX_SPHERES 48  /* collision spheres per robot */
#define OIGE_MAX_SENSORS 4
#define OIGE_MAX_PAIRS 96    /* self-collision link pairs per robot */
#define OIGE_STATS_SLOTS 2048  /* episode_stats is striped over slots: same-address fp64 atomics serialise at ~50 ns each in L2 */

#define OIGE_OK 0
#define OIGE_E_NULL (-1)        /* null pointer argument */
#define OIGE_E_SHAPE (-2)       /* unsupported (num_dof, num_sensors, num_obs) combination */
#define OIGE_E_ALIGN (-3)       /* device pointer not 16-byte aligned */
#define OIGE_E_RANGE (-4)       /* n_envs / substeps / k out of range */
#define OIGE_E_MODEL (-5)       /* malformed articulation model */
#define OIGE_E_NODEVICE (-6)    /* no CUDA device / wrong architecture */

enum { OIGE_LIMIT_COST_HUMANOID = 0, OIGE_LIMIT_COST_ANT = 1 };
enum { OIGE_JOINT_HINGE = 0, OIGE_JOINT_PRISMATIC = 1 };

/* Scalar constants of one locomotion task.
 * Replaces: LocomotionTask.__init__ scalars (tasks/shared/locomotion.py:53-67), the per-env
 * replicated constants of post_reset (locomotion.py:152-163: targets, inv_start_rot, basis
 * vectors, dt = 1/60), joint_gears / motor_effort_ratio / dof limits (tasks/humanoid.py:82-112,
 * tasks/ant.py:80-84), clip_obs (tasks/base/rl_task.py:69). */
typedef struct OigeTaskParams {
    int32_t num_dof;          /* D: 21 Humanoid, 8 Ant */
    int32_t num_sensors;      /* S: 2 Humanoid, 4 Ant */
    int32_t num_obs;          /* 12 + 3D + 6S */
    int32_t limit_cost_kind;  /* OIGE_LIMIT_COST_* (humanoid.py:120-127 vs ant.py:92-95) */
    int32_t walk_target_uses_z; /* 1 = upstream compute_rot quirk atan2(dz, dx) (default), 0 = atan2(dy, dx) */
    int32_t progress_increment; /* 1: fuse `progress_buf += 1` (rl_task.py:242) into the pass; 0: leave it */
    int32_t closed_form_angles; /* 1: normalize_angle(x) evaluated as the closed-form wrap of x into (-pi,pi]
                                   (1-2 ulp from atan2(sin x, cos x)); 0: the literal chain (locomotion.py:190-192) */
    int32_t reserved_i[1];
    float dt;                 /* task dt, 1/60 (locomotion.py:163) */
    float dof_vel_scale;
    float angular_velocity_scale;
    float contact_force_scale;
    float heading_weight;
    float up_weight;
    float actions_cost_scale;
    float energy_cost_scale;
    float joints_at_limit_cost_scale;
    float death_cost;
    float termination_height;
    float alive_reward_scale;
    float max_episode_length; /* float, as in the jit signature (locomotion.py:265) */
    float clip_obs;           /* +inf = no clamp (envs/vec_env_rlgames.py:42) */
    float power_scale;
    float reserved_f;
    float targets[4];         /* xyz (+pad) */
    float inv_start_rot[4];   /* wxyz */
    float basis_vec0[4];      /* heading basis xyz (+pad) */
    float basis_vec1[4];      /* up basis xyz (+pad) */
    float dof_limits_lower[OIGE_MAX_DOF];
    float dof_limits_upper[OIGE_MAX_DOF];
    float motor_effort_ratio[OIGE_MAX_DOF];
    float joint_gears[OIGE_MAX_DOF];
} OigeTaskParams;

/* ArticulationView-style state tensors.
 * Replaces the buffers behind get_world_poses / get_velocities / get_joint_positions /
 * get_joint_velocities / _physics_view.get_force_sensor_forces (locomotion.py:81-89). */
typedef struct OigeState {
    float* root_pos;   /* [N,3] world frame */
    float* root_quat;  /* [N,4] wxyz */
    float* root_vel;   /* [N,6] lin, ang (world frame) */
    float* dof_pos;    /* [N,D] */
    float* dof_vel;    /* [N,D] */
    float* sensors;    /* [N,S,6] force, torque at each foot sensor (sensor frame) */
    float* dof_pos_targets; /* [N,D] PD position targets (set_joint_position_targets) or NULL = effort control only */
} OigeState;

/* RL buffers of the task.  Replaces RLTask.cleanup buffers (tasks/base/rl_task.py:98-107) and
 * the LocomotionTask extras of post_reset (locomotion.py:152-168). */
typedef struct OigeTaskBuffers {
    float* obs;              /* [N,OBS] */
    float* rew;              /* [N] */
    int64_t* reset_buf;      /* [N] in/out */
    int64_t* progress_buf;   /* [N] in/out */
    float* potentials;       /* [N] in/out */
    float* prev_potentials;  /* [N] out */
    float* up_vec;           /* [N,3] out */
    float* heading_vec;      /* [N,3] out */
    float* actions;          /* [N,D] the task's stored (clamped) actions */
    double* episode_stats;   /* [OIGE_STATS_SLOTS,4] accumulators or NULL: sum(rew) (see episode_return), #done,
                                sum(len at done), #env-steps */
    uint64_t* reset_counter; /* NULL, or TWO device words {counter, 0}: oige_env_step_f32 then takes the Philox reset
                                counter from word 0 (instead of its `counter` argument) and increments it at the end of
                                the step, so that the step can be captured once in a CUDA graph and replayed; word 1 is
                                scratch of the single-launch step (its CTAs take a ticket there: the last one to finish
                                advances the counter) and must be zero between steps */
    /* ---- ABI 4 ---- */
    float* episode_return;   /* NULL, or [N] in/out running return of each env's CURRENT episode: every pass adds the
                                step's reward, and an env whose episode ends in this pass (reset_buf set) contributes its
                                return to episode_stats[.][0] and starts again from 0.  episode_stats[.][0] is then the
                                sum of the RETURNS OF FINISHED EPISODES (what the reference's observer averages,
                                utils/rlgames/rlgames_utils.py:51-62) instead of the sum of every env-step's reward.
                                The caller zeroes the rows of envs it resets by hand (reset_idx). */
} OigeTaskBuffers;

/* Per-env initial state captured at post_reset (locomotion.py:148-150). */
typedef struct OigeResetInit {
    const float* initial_root_pos;  /* [N,3] */
    const float* initial_root_rot;  /* [N,4] */
    const float* initial_dof_pos;   /* [N,D] */
} OigeResetInit;

/* Host-side description of one robot for the articulation kernel (one link per DoF plus the
 * floating root; multi-DoF joints are chains of massless links).  No reference equivalent: the
 * reference loads USD assets from a Nucleus server (robots/articulations/humanoid.py:57, ant.py:57)
 * into PhysX.  Link 0 is the root. Links are in an order where parent[i] < i. */
typedef struct OigeModelDesc {
    int32_t num_links;                       /* 1 + D */
    int32_t num_spheres;
    int32_t num_sensors;
    int32_t fixed_base;                      /* 1: link 0 is welded to the world (Cartpole rail), 0: floating root */
    int32_t parent[OIGE_MAX_LINKS];          /* parent[0] = -1 */
    int32_t dof_index[OIGE_MAX_LINKS];       /* link -> column of dof_pos (BFS dof order); -1 for root */
    int32_t joint_type[OIGE_MAX_LINKS];      /* OIGE_JOINT_HINGE (rotation about joint_axis) or OIGE_JOINT_PRISMATIC (translation along it) */
    float joint_pos[OIGE_MAX_LINKS][3];      /* joint anchor in the parent link frame */
    float joint_quat[OIGE_MAX_LINKS][4];     /* fixed rotation parent link frame -> this link frame at q=0 (wxyz) */
    float joint_axis[OIGE_MAX_LINKS][3];     /* unit hinge axis in this link's frame */
    float mass[OIGE_MAX_LINKS];
    float com[OIGE_MAX_LINKS][3];            /* link frame */
    float inertia[OIGE_MAX_LINKS][6];        /* about COM, link frame: xx yy zz xy xz yz */
    float armature[OIGE_MAX_LINKS];
    float damping[OIGE_MAX_LINKS];
    float stiffness[OIGE_MAX_LINKS];         /* passive spring towards spring_ref */
    float spring_ref[OIGE_MAX_LINKS];
    float limit_lower[OIGE_MAX_LINKS];
    float limit_upper[OIGE_MAX_LINKS];
    float drive_stiffness[OIGE_MAX_LINKS];   /* PD position drive: tau = kp (target - q) - kd qd, |tau| <= max force */
    float drive_damping[OIGE_MAX_LINKS];
    float drive_max_force[OIGE_MAX_LINKS];   /* <= 0: unlimited */
    int32_t sphere_link[OIGE_MAX_SPHERES];
    float sphere_pos[OIGE_MAX_SPHERES][3];   /* link frame */
    float sphere_radius[OIGE_MAX_SPHERES];
    int32_t sensor_link[OIGE_MAX_SENSORS];   /* link whose net ground-contact wrench is reported */
    float gravity[3];
    float contact_stiffness;                 /* N/m */
    float contact_damping;                   /* N s/m, only while approaching */
    float friction_mu;                       /* cfg/task/Humanoid.yaml:46-48 */
    float friction_viscous;                  /* N s/m regularisation slope of the Coulomb cone */
    float limit_stiffness;                   /* N m/rad beyond a joint limit */
    float limit_damping;
    float max_joint_velocity;                /* JointAPI "Maximum Joint Velocity", rad/s (m/s for a prismatic joint); Isaac Sim
                                                default 1e6 deg/s (docs/transfering_policies_from_isaac_gym.md:85-87) */
    float angular_damping;                   /* RigidBodyAPI "Angular Damping", 1/s, on EVERY body: the root's angular
                                                velocity decays by (1 - dt c) per substep, every other link receives the
                                                torque -c I_link w_link; Isaac Sim default 0.05 (same doc :71-77) */
    /* ---- ABI 3 ---- */
    float max_angular_velocity;              /* RigidBodyAPI "Max Angular Velocity", rad/s (Isaac Sim default 5729.58 deg/s
                                                = 100 rad/s): clamps |w| of the root body and, in reduced coordinates, each
                                                hinge's relative angular velocity |qd|; <= 0: unlimited */
    float max_depenetration_velocity;        /* m/s (cfg/task/Humanoid.yaml:62,90): the penalty springs see at most
                                                max_depenetration_velocity * dt of penetration per substep; <= 0: unlimited */
    float self_contact_stiffness;            /* N/m, capsule-capsule penalty between the links of a collision pair */
    float self_contact_damping;              /* N s/m, only while approaching */
    int32_t num_collision_pairs;             /* enable_self_collisions (cfg/task/Humanoid.yaml:80): unordered LINK pairs whose
                                                collision spheres repel each other; 0 = no self-collision (Ant.yaml:82) */
    int32_t pair_link_a[OIGE_MAX_PAIRS];
    int32_t pair_link_b[OIGE_MAX_PAIRS];
    /* self-collision shape of a link: ONE capsule (segment a-b in the link frame, radius); radius <= 0 = none.  Every
     * link named by a collision pair must have one. */
    float capsule_a[OIGE_MAX_LINKS][3];
    float capsule_b[OIGE_MAX_LINKS][3];
    float capsule_radius[OIGE_MAX_LINKS];
} OigeModelDesc;

typedef struct OigeModel OigeModel; /* opaque; holds the immutable device-side model table */

/* ---- library ---- */
int oige_abi_version(void);
const char* oige_last_error(void);
/* number of kernels this library has launched since load (monotonic; for bench.py's gpu_launches) */
uint64_t oige_kernel_launch_count(void);
/* The OIGE_* environment knobs are read ONCE, at the first call that needs them; this re-reads
 * them (the test-suite switches the articulation kernel between tests).  No reference counterpart. */
void oige_reload_env(void);
/* Environments the articulation kernels parked below the termination height because their state became non-finite
 * (solver divergence), summed over the current device since the library was loaded.  A healthy run reports 0; a
 * parked env is indistinguishable from a fall in reset_buf / rew_buf, so episode statistics should be read next to
 * this counter.  Synchronises the device (diagnostic).  No reference counterpart. */
uint64_t oige_nonfinite_env_count(void);
/* Diagnostic: with OIGE_HOST_TRACE=1 in the environment, prints the per-chunk timeline (H2D / kernels / D2H done, us)
 * of the last oige_env_step_host_f32 call -- or CUDA-graph replay of it -- to stderr.  No reference counterpart. */
int oige_host_trace_dump(void);

/* ---- K1: fused observation / reward / termination pass ----
 * Replaces, in one launch: RLTask.post_physics_step's `progress_buf += 1` (rl_task.py:242),
 * LocomotionTask.get_observations + jit get_observations (locomotion.py:80-101, 194-254),
 * get_dof_at_limit_cost (humanoid.py:116-127 / ant.py:88-95), calculate_metrics + jit
 * (locomotion.py:173-178, 271-321), is_done + jit (locomotion.py:180-183, 257-268) and the obs
 * clamp of VecEnvRLGames._process_data (envs/vec_env_rlgames.py:42). */
int oige_obs_reward_reset_f32(const OigeTaskParams* params, const OigeState* state,
                              const OigeTaskBuffers* buf, int64_t n_envs, void* stream);
/* same pass with an explicit tile size (envs per CTA: 32 or 64; 0 = automatic) -- tuning hook */
int oige_obs_reward_reset_tile_f32(const OigeTaskParams* params, const OigeState* state,
                                   const OigeTaskBuffers* buf, int64_t n_envs, int32_t tile, void* stream);

/* ---- K3: reset ----
 * Replaces LocomotionTask.reset_idx (locomotion.py:116-145) and the nonzero()/len() host sync of
 * pre_physics_step (locomotion.py:104-106).
 * masked: resets every env whose reset_buf != 0 (device-side mask, no host sync).
 * idx:    resets the k envs listed in env_ids (device int64), the reference's calling form.
 * rand_pos / rand_vel: raw U[0,1) draws ([N,D] for masked, [k,D] for idx) -- the values
 * torch.rand would have produced (parity mode).  If NULL the kernel draws Philox4x32-10
 * uniforms keyed by (seed, env_id_offset + env, counter) so results do not depend on sharding. */
int oige_reset_masked_f32(const OigeTaskParams* params, const OigeState* state, const OigeTaskBuffers* buf,
                          const OigeResetInit* init, const float* rand_pos, const float* rand_vel,
                          uint64_t seed, uint64_t counter, int64_t env_id_offset,
                          int64_t n_envs, void* stream);
int oige_reset_idx_f32(const OigeTaskParams* params, const OigeState* state, const OigeTaskBuffers* buf,
                       const OigeResetInit* init, const int64_t* env_ids, int64_t k,
                       const float* rand_pos, const float* rand_vel,
                       uint64_t seed, uint64_t counter, int64_t env_id_offset, void* stream);

/* ---- K2: batched reduced-coordinate articulation step ----
 * Replaces set_joint_efforts (locomotion.py:108-114) + control_frequency_inv x World.step
 * (envs/vec_env_rlgames.py:64-66; PhysX, unavailable) with a warp-per-env articulated-body
 * forward-dynamics step.  efforts = actions * joint_gears * power_scale. */
OigeModel* oige_model_create(const OigeModelDesc* desc);
void oige_model_destroy(OigeModel* model);
int oige_articulation_step_f32(const OigeModel* model, const OigeTaskParams* params, const OigeState* state,
                               const float* actions, int64_t n_envs, int32_t substeps, float dt, void* stream);

/* ---- whole VecEnvRLGames.step on device buffers ----
 * Replaces VecEnvRLGames.step (envs/vec_env_rlgames.py:56-78) minus the rl_device copies: clamps `actions_in` to
 * +-clip_actions into buf->actions, masked reset_idx of the flagged envs (tasks/shared/locomotion.py:103-145),
 * `substeps` articulation substeps, progress += 1 / observations / reward / done (tasks/base/rl_task.py:242,
 * locomotion.py:190-321).  No host synchronisation.  ONE launch (K2-F, env_step_role_kernel: prologue + substeps +
 * epilogue in the role-split articulation kernel) whenever that kernel serves the model and env count; three launches
 * (K3 -> K2-W / K2-T -> K1) otherwise.  The rule (k2_uses_role_kernel, csrc/articulation_step.cu):
 *   - a model WITH collision pairs (the stock Humanoid) takes K2-F at EVERY env count -- only K2-F integrates
 *     self-collision -- and a forced OIGE_K2_MODE=warp|thread then fails with OIGE_E_MODEL instead of dropping the pairs;
 *   - without pairs, a three-role tree (Humanoid shape) takes K2-F from 4096 envs up; a four-role tree (Ant, Anymal) from
 *     1024 envs up to two waves (2 * 4 * 32 * SM count = 37,888 envs on a B200), beyond which the three launches with the
 *     thread-per-env kernel are faster for such small trees (OIGE_K2_MODE=role keeps K2-F);
 *   - a tree the role split cannot cut (trunk not a path, more than 4 chains ...) always takes the three launches.
 * Both paths produce the same bits as calling the three entry points above one after the other.
 * buf->reset_counter, when not NULL, MUST be the two-word form {counter, ticket = 0} described at OigeTaskBuffers: K2-F
 * does an atomic on word 1, so a one-word scalar from an older caller is written out of bounds. */
int oige_env_step_f32(const OigeModel* model, const OigeTaskParams* params, const OigeState* state,
                      const OigeTaskBuffers* buf, const OigeResetInit* init, const float* actions_in,
                      float clip_actions, uint64_t seed, uint64_t counter, int64_t env_id_offset,
                      int64_t n_envs, int32_t substeps, float dt, void* stream);

/* ---- the same step with HOST buffers (the call rl_games would make with CPU tensors) ----
 * actions_host [N,D] -> H2D into actions_dev_staging; after the step obs/rew/reset_buf are copied
 * D2H into the host buffers.  Internally the env range is cut into 4 tile-aligned chunks, each on
 * a library-owned stream (H2D -> K3 -> K2 -> K1 -> D2H), so copies overlap kernels; `stream` waits for
 * all of them.  Pinned host memory gives true asynchrony; the caller synchronises `stream` before
 * reading the outputs.  Results are identical to oige_env_step_f32 (envs are independent). */
int oige_env_step_host_f32(const OigeModel* model, const OigeTaskParams* params, const OigeState* state,
                           const OigeTaskBuffers* buf, const OigeResetInit* init,
                           const float* actions_host, float* actions_dev_staging,
                           float* obs_host, float* rew_host, int64_t* reset_host,
                           float clip_actions, uint64_t seed, uint64_t counter, int64_t env_id_offset,
                           int64_t n_envs, int32_t substeps, float dt, void* stream);

/* ---- the host-buffer step for callers that cannot provide device scratch ----
 * No counterpart in the reference (its .to(device) copies allocate through torch every step,
 * envs/vec_env_rlgames.py:57): this exists for the users of THIS library's host-buffer step.
 * oige_env_step_host_f32 wants `actions_dev_staging`, an [N,D] device buffer that is pure scratch.  The callers of the
 * host-buffer step are the ones least able to supply it: a policy loop that lives in host memory behind a ctypes /
 * cgo binding has no CUDA runtime of its own to allocate with, and inside a torch process a raw cudaMalloc from the
 * step loop bypasses the caching allocator (it can fail once that pool holds the HBM, and it stalls the device).
 * This variant keeps the staging rows in a fixed array inside the library image and allocates nothing: no cudaMalloc /
 * cudaFree, no stream or event creation, no torch.  Same arguments as oige_env_step_host_f32 minus the staging
 * pointer; same results, bit for bit, as oige_env_step_host_f32 / oige_env_step_f32.
 *   limit      n_envs * params->num_dof <= OIGE_STATIC_STAGING_FLOATS (2 Mi floats = 8 MiB: 99,864 Humanoids, 262,144
 *              Ants, 65,536 envs of a 32-DoF robot); beyond it the call returns OIGE_E_RANGE and does nothing.
 *   footprint  every GPU that loads the library holds those 8 MiB until the process exits.
 *   sharing    one array per GPU, used by every call: calls whose GPU work may overlap (other streams, other host
 *              threads, several envs on one GPU) race on it -- keep them ordered on the device.  The array keeps the
 *              previous call's rows; none of it needs clearing, the rows a call reads are the rows its own H2D wrote.
 *   first call on each GPU it looks up the array's device address (with that GPU current) and caches it; make that
 *              call -- n_envs = 0 suffices -- before beginning any stream capture.
 * Copies and kernel run back to back on `stream` (no library-owned chunk streams); the caller synchronises `stream`
 * before reading the host outputs. */
#define OIGE_STATIC_STAGING_FLOATS (2 * 1024 * 1024)
#define OIGE_STATIC_STAGING_BYTES (4 * OIGE_STATIC_STAGING_FLOATS)
int oige_env_step_host_static_f32(const OigeModel* model, const OigeTaskParams* params, const OigeState* state,
                                  const OigeTaskBuffers* buf, const OigeResetInit* init,
                                  const float* actions_host, float* obs_host, float* rew_host, int64_t* reset_host,
                                  float clip_actions, uint64_t seed, uint64_t counter, int64_t env_id_offset,
                                  int64_t n_envs, int32_t substeps, float dt, void* stream);

/* ================= Cartpole (tasks/cartpole.py) =================
 * Same three-kernel step as the locomotion tasks with the task-specific passes swapped:
 *   reset + efforts (pre_physics_step, cartpole.py:101-130)  ->  K2 on the 3-link welded-base tree
 *   ->  observations + reward + done (cartpole.py:80-99,143-162; rl_task.py:242).
 * dof_pos / dof_vel / efforts are [N,2] in DoF order (cartJoint, poleJoint by default). */
typedef struct OigeCartpoleParams {
    int32_t cart_dof;           /* get_dof_index("cartJoint") (cartpole.py:133) */
    int32_t pole_dof;           /* get_dof_index("poleJoint") (cartpole.py:134) */
    int32_t progress_increment; /* 1: fuse `progress_buf += 1` (rl_task.py:242) into the pass; 0: leave it */
    int32_t reserved_i;
    float reset_dist;           /* cfg resetDist, 3.0 (cartpole.py:59) */
    float max_push_effort;      /* cfg maxEffort, 400 (cartpole.py:60) */
    float max_episode_length;   /* 500 (cartpole.py:61); done when progress >= it (cartpole.py:161) */
    float clip_obs;             /* 5.0 (Cartpole.yaml:13); the STORED observations are clamped, reward and
                                   done are evaluated on the unclamped values like the reference */
} OigeCartpoleParams;

typedef struct OigeCartpoleBuffers {
    float* obs;              /* [N,4] cart_pos, cart_vel, pole_pos, pole_vel (cartpole.py:89-92) */
    float* rew;              /* [N] */
    int64_t* reset_buf;      /* [N] */
    int64_t* progress_buf;   /* [N] */
    float* efforts;          /* [N,2] what set_joint_efforts receives (cartpole.py:104-108) */
    double* episode_stats;   /* [OIGE_STATS_SLOTS,4] or NULL, as in OigeTaskBuffers */
    uint64_t* reset_counter; /* NULL or device scalar, as in OigeTaskBuffers */
} OigeCartpoleBuffers;

/* get_observations + calculate_metrics + is_done (+ progress += 1) in one pass (cartpole.py:80-99,143-162).
 * Only state->dof_pos / dof_vel are read. */
int oige_cartpole_obs_reward_reset_f32(const OigeCartpoleParams* params, const OigeState* state,
                                       const OigeCartpoleBuffers* buf, int64_t n_envs, void* stream);
/* pre_physics_step (cartpole.py:101-108): resets every env whose reset_buf != 0 on the device (no nonzero()
 * host sync) and, if actions_in ([N,1]) is not NULL, writes efforts = maxEffort * clamp(actions, +-clip_actions)
 * on the cart DoF and 0 on the pole.  rand: raw U[0,1) draws [N,4] in the reference's draw order (cart position,
 * pole angle, cart velocity, pole velocity; cartpole.py:115-121) or NULL for Philox keyed by
 * (seed, env_id_offset + env, counter). */
int oige_cartpole_reset_masked_f32(const OigeCartpoleParams* params, const OigeState* state,
                                   const OigeCartpoleBuffers* buf, const float* actions_in, float clip_actions,
                                   const float* rand, uint64_t seed, uint64_t counter, int64_t env_id_offset,
                                   int64_t n_envs, void* stream);
/* reset_idx(env_ids) (cartpole.py:110-130), the reference's calling form; rand is [k,4] or NULL. */
int oige_cartpole_reset_idx_f32(const OigeCartpoleParams* params, const OigeState* state,
                                const OigeCartpoleBuffers* buf, const int64_t* env_ids, int64_t k, const float* rand,
                                uint64_t seed, uint64_t counter, int64_t env_id_offset, void* stream);
/* Whole VecEnvRLGames.step (envs/vec_env_rlgames.py:56-78) for Cartpole on device buffers: three launches. */
int oige_cartpole_env_step_f32(const OigeModel* model, const OigeCartpoleParams* params, const OigeState* state,
                               const OigeCartpoleBuffers* buf, const float* actions_in, float clip_actions,
                               uint64_t seed, uint64_t counter, int64_t env_id_offset, int64_t n_envs,
                               int32_t substeps, float dt, void* stream);

/* ================= Anymal, flat terrain (tasks/anymal.py) =================
 * PD position-target actuation: pre_physics_step integrates and clamps the joint targets, K2 drives the joints
 * toward them with the model's drive gains (set_drive, anymal.py:128), post_physics_step forms 48 observations,
 * the velocity-tracking reward and the fall / time-out flags.  D = 12; all [N,12] tensors in the reference's
 * breadth-first DoF order (robots/articulations/anymal.py:72-83). */
#define OIGE_ANYMAL_DOF 12
#define OIGE_ANYMAL_OBS 48

typedef struct OigeAnymalParams {
    int32_t progress_increment;  /* 1: fuse `progress_buf += 1` (rl_task.py:242) into the pass */
    int32_t reserved_i[3];
    float lin_vel_scale, ang_vel_scale, dof_pos_scale, dof_vel_scale; /* anymal.py:57-60 */
    float action_scale;          /* 13.5 (anymal.py:61) */
    float dt;                    /* task dt, 1/60 (anymal.py:88) */
    /* reward scales ALREADY multiplied by dt in double precision (anymal.py:94-95) */
    float rew_lin_vel_xy, rew_ang_vel_z, rew_lin_vel_z, rew_joint_acc, rew_action_rate, rew_cosmetic;
    float command_x[2], command_y[2], command_yaw[2]; /* (min, max) of the resampled commands (anymal.py:72-74) */
    float fallen_threshold;      /* base height below which the env is reset with reward -1, 0.51 (anymal.py:284) */
    float max_episode_length;    /* int(episodeLength_s / dt + 0.5) = 3000; done when progress >= it - 1 */
    float clip_obs;              /* 5.0; the STORED observations are clamped */
    float reserved_f;
    float default_dof_pos[OIGE_ANYMAL_DOF];   /* defaultJointAngles in DoF order (anymal.py:130-135) */
    float dof_limits_lower[OIGE_ANYMAL_DOF];  /* anymal.py:234-235 */
    float dof_limits_upper[OIGE_ANYMAL_DOF];
} OigeAnymalParams;

typedef struct OigeAnymalBuffers {
    float* obs;              /* [N,48] lin vel 3, ang vel 3, projected gravity 3, commands 3, dof pos 12, dof vel 12,
                                actions 12 (anymal.py:158-169) */
    float* rew;              /* [N] */
    int64_t* reset_buf;      /* [N] */
    int64_t* progress_buf;   /* [N] */
    float* actions;          /* [N,12] the task's stored (clamped) actions */
    float* last_actions;     /* [N,12] in/out (anymal.py:280) */
    float* last_dof_vel;     /* [N,12] in/out (anymal.py:281) */
    float* commands;         /* [N,3] x, y, yaw velocity commands (anymal.py:237-240) */
    float* current_targets;  /* [N,12] joint position targets; the caller passes the same tensor as
                                OigeState.dof_pos_targets (what set_joint_position_targets receives) */
    double* episode_stats;   /* [OIGE_STATS_SLOTS,4] or NULL, as in OigeTaskBuffers */
    uint64_t* reset_counter; /* NULL or device scalar, as in OigeTaskBuffers */
} OigeAnymalBuffers;

/* get_observations + calculate_metrics + is_done (+ progress += 1) in one pass (anymal.py:137-179,252-293). */
int oige_anymal_obs_reward_reset_f32(const OigeAnymalParams* params, const OigeState* state,
                                     const OigeAnymalBuffers* buf, int64_t n_envs, void* stream);
/* pre_physics_step (anymal.py:181-190): device-side masked reset_idx (no nonzero() host sync), then -- if
 * actions_in ([N,12]) is not NULL -- actions = clamp(actions_in, +-clip_actions) and
 * current_targets = clamp(current_targets + action_scale * actions * dt, lower, upper).
 * rand_vel [N,12] / rand_cmd [N,3]: raw U[0,1) draws in the reference's draw order (joint velocities; command
 * x, y, yaw; anymal.py:195,211-219) or NULL for Philox keyed by (seed, env_id_offset + env, counter).
 * init: initial_root_pos / initial_root_rot (anymal.py:230); initial_dof_pos is not read. */
int oige_anymal_reset_masked_f32(const OigeAnymalParams* params, const OigeState* state, const OigeAnymalBuffers* buf,
                                 const OigeResetInit* init, const float* actions_in, float clip_actions,
                                 const float* rand_vel, const float* rand_cmd, uint64_t seed, uint64_t counter,
                                 int64_t env_id_offset, int64_t n_envs, void* stream);
/* reset_idx(env_ids) (anymal.py:192-227); rand_vel [k,12], rand_cmd [k,3] or NULL. */
int oige_anymal_reset_idx_f32(const OigeAnymalParams* params, const OigeState* state, const OigeAnymalBuffers* buf,
                              const OigeResetInit* init, const int64_t* env_ids, int64_t k, const float* rand_vel,
                              const float* rand_cmd, uint64_t seed, uint64_t counter, int64_t env_id_offset,
                              void* stream);
/* Whole VecEnvRLGames.step (envs/vec_env_rlgames.py:56-78) for Anymal on device buffers: three launches. */
int oige_anymal_env_step_f32(const OigeModel* model, const OigeAnymalParams* params, const OigeState* state,
                             const OigeAnymalBuffers* buf, const OigeResetInit* init, const float* actions_in,
                             float clip_actions, uint64_t seed, uint64_t counter, int64_t env_id_offset,
                             int64_t n_envs, int32_t substeps, float dt, void* stream);

/* ---- observation / action noise (utils/domain_randomization/randomize.py:212-305) -------------------------------
 * Replaces Randomizer.apply_actions_randomization / apply_observations_randomization, the two hooks of
 * VecEnvRLGames.step (envs/vec_env_rlgames.py:59-60,70-71), with one launch per buffer and no nonzero() host sync. */
enum { OIGE_DR_GAUSSIAN = 0, OIGE_DR_UNIFORM = 1, OIGE_DR_LOGUNIFORM = 2 };
enum { OIGE_DR_ADDITIVE = 0, OIGE_DR_SCALING = 1 };
enum { OIGE_DR_STREAM_ACTIONS = 0, OIGE_DR_STREAM_OBSERVATIONS = 2 }; /* + 0 on_reset, + 1 on_interval */

typedef struct OigeNoiseSpec {   /* one on_reset / on_interval block of the task YAML */
    int32_t enabled;             /* 0: block absent */
    int32_t operation;           /* OIGE_DR_ADDITIVE | OIGE_DR_SCALING */
    int32_t distribution;        /* OIGE_DR_* */
    int32_t frequency_interval;  /* on_interval only */
    float p0, p1;                /* (mean, std) or (low, high) */
} OigeNoiseSpec;

typedef struct OigeNoiseState {  /* randomize.py:192-193 / :209-210 */
    int32_t* counter;            /* [N] steps since the last interval noise (int32, as the reference's torch.int) */
    float* correlated;           /* [N,dim] on-reset noise kept until the env's next reset; may be NULL if !on_reset.enabled */
} OigeNoiseState;

/* out[N,dim] = clamp(noise(clamp(in[N,dim], +-clip_in)), +-clip_out); `out` may alias `in`; INFINITY skips a clamp.
 * clip_in folds the wrapper's action clamp (vec_env_rlgames.py:57, before the noise) and clip_out its observation
 * clamp (vec_env_rlgames.py:42, after the noise) into the launch.  reset_buf [N] int64 is only read.
 * noise_reset / noise_interval: [N,dim] ready-made noise values, read at the flagged envs only (parity mode: the
 * caller made the reference's torch.normal / torch.rand draws), or NULL for in-kernel Philox4x32-10 keyed by
 * (seed, env_id_offset + env, counter, stream + {0,1}) -- sharding-invariant.  If counter_dev is not NULL the
 * counter is read from that device scalar instead (CUDA-graph replays; the task's reset counter serves). */
int oige_dr_apply_f32(const OigeNoiseSpec* on_reset, const OigeNoiseSpec* on_interval, const OigeNoiseState* state,
                      const float* in, float* out, const int64_t* reset_buf, int64_t n_envs, int32_t dim, float clip_in,
                      float clip_out, const float* noise_reset, const float* noise_interval, uint64_t seed, uint64_t counter,
                      const uint64_t* counter_dev, int32_t stream_id, int64_t env_id_offset, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* OIGE_B200_H */
