// Anymal task passes (reference: tasks/anymal.py:137-293), D = 12, OBS = 48.
//
//   anymal_post_kernel   progress += 1, get_observations, calculate_metrics, is_done (anymal.py:137-179,252-293;
//                        rl_task.py:242).  HBM-bound: 77 f32 + 8 B read, 73 f32 + 16 B written = 624 B per env-step.
//   anymal_pre_kernel    pre_physics_step (anymal.py:181-190): device-side masked reset_idx + action clamp +
//                        target integration / clamp, one quartet of lanes per env.
//   anymal_reset_idx_kernel  reset_idx(env_ids) (anymal.py:192-227), the reference's calling form.
//
// Layout of the post pass: a CTA of 128 threads owns a tile of E = 32 consecutive envs.  Every [N,k] field of the
// tile is ONE contiguous chunk in global memory (row-major [N,k]), so the tile is staged by nine 1-D bulk copies
// (cp.async.bulk + mbarrier, issued by one thread) and drained by eight; ragged / misaligned tiles take a plain
// coalesced cooperative path.  Env e of the tile is served by the four adjacent lanes 4e..4e+3 of one warp
// ("quartet"); lane r owns dofs r, r+4, r+8 -- with the row stride of 12 words the 32 lanes of a warp then hit
// 32 distinct shared-memory banks -- and obs block r (0 linear velocity, 1 angular velocity, 2 projected gravity,
// 3 commands); the three per-dof reductions are finished with two xor-shuffles inside the quartet.
//
// Compiled with -fmad=false so every product/sum rounds once, as in torch eager; sums over dofs are evaluated in a
// different order than torch's vectorised reduction, hence floats are compared at 1e-6 and flags bit-exactly.
#include <math_constants.h>

#include "oige_common.cuh"

namespace oige {
namespace {

constexpr int D = OIGE_ANYMAL_DOF;
constexpr int OBS = OIGE_ANYMAL_OBS;
constexpr int E = 32;        // envs per tile
constexpr int NT = 4 * E;    // threads per CTA

struct AnyArgs {
    const float* pos; const float* quat; const float* vel;
    float* q; float* qd;
    float* obs; float* rew;
    long long* reset; long long* progress;
    float* actions; float* last_actions; float* last_qd; float* commands; float* targets;
    double* stats;
    unsigned long long* step_counter;
    const unsigned long long* counter_dev;
    float* wpos; float* wquat; float* wvel;                   // reset writes the root state
    const float* init_pos; const float* init_quat;
    const float* actions_in; const float* rand_vel; const float* rand_cmd;
    const long long* env_ids;
    unsigned long long seed, counter;
    long long env_id_offset, n;
    float clip_actions;
    int use_bulk;
};

template <typename T>
__device__ __forceinline__ void coop_copy(T* dst, const T* src, int count, int tid) {
    for (int i = tid; i < count; i += NT) dst[i] = src[i];
}
__device__ __forceinline__ float clampf(float x, float c) { return fminf(fmaxf(x, -c), c); }
__device__ __forceinline__ float sq(float x) { return x * x; }

// quat_rotate / quat_rotate_inverse of omni.isaac.core.utils.torch.rotations (wxyz):
//   a = v (2 w^2 - 1);  b = 2 w (q_vec x v);  c = 2 q_vec (q_vec . v);  rotate = a + b + c, inverse = a - b + c
template <bool INVERSE>
__device__ __forceinline__ void qrot(const float q[4], const float v[3], float out[3]) {
    const float w = q[0], x = q[1], y = q[2], z = q[3];
    const float s = 2.0f * (w * w) - 1.0f;
    const float cx = y * v[2] - z * v[1], cy = z * v[0] - x * v[2], cz = x * v[1] - y * v[0];
    const float dot = (x * v[0] + y * v[1]) + z * v[2];
    const float a[3] = {v[0] * s, v[1] * s, v[2] * s};
    const float b[3] = {cx * w * 2.0f, cy * w * 2.0f, cz * w * 2.0f};
    const float c[3] = {x * dot * 2.0f, y * dot * 2.0f, z * dot * 2.0f};
#pragma unroll
    for (int k = 0; k < 3; ++k) out[k] = INVERSE ? (a[k] - b[k]) + c[k] : (a[k] + b[k]) + c[k];
}

__global__ void __launch_bounds__(NT)
anymal_post_kernel(const __grid_constant__ OigeAnymalParams P, const __grid_constant__ AnyArgs A) {
    __shared__ __align__(128) long long s_prog[E];
    __shared__ __align__(16) long long s_reset[E];
    __shared__ __align__(16) float s_pos[3 * E];
    __shared__ __align__(16) float s_quat[4 * E];
    __shared__ __align__(16) float s_vel[6 * E];
    __shared__ __align__(16) float s_cmd[3 * E];
    __shared__ __align__(16) float s_q[D * E];
    __shared__ __align__(16) float s_qd[D * E];
    __shared__ __align__(16) float s_act[D * E];
    __shared__ __align__(16) float s_lact[D * E];
    __shared__ __align__(16) float s_lqd[D * E];
    __shared__ __align__(16) float o_obs[OBS * E];
    __shared__ __align__(16) float o_rew[E];
    __shared__ __align__(8) uint64_t s_bar;

    const int tid = threadIdx.x;
    const long long e0 = (long long)blockIdx.x * E;
    const int cnt = (int)((A.n - e0) < (long long)E ? (A.n - e0) : (long long)E);
    const bool bulk = A.use_bulk && cnt == E;

    if (bulk) {
        if (tid == 0) {
            mbar_init(&s_bar, 1);
            mbar_fence_init();
            mbar_expect_tx(&s_bar, (uint32_t)E * (4u * (3 + 4 + 6 + 3 + 5 * D) + 8u));
            bulk_g2s(s_q, A.q + e0 * D, E * D * 4, &s_bar);
            bulk_g2s(s_qd, A.qd + e0 * D, E * D * 4, &s_bar);
            bulk_g2s(s_act, A.actions + e0 * D, E * D * 4, &s_bar);
            bulk_g2s(s_lact, A.last_actions + e0 * D, E * D * 4, &s_bar);
            bulk_g2s(s_lqd, A.last_qd + e0 * D, E * D * 4, &s_bar);
            bulk_g2s(s_quat, A.quat + e0 * 4, E * 16, &s_bar);
            bulk_g2s(s_pos, A.pos + e0 * 3, E * 12, &s_bar);
            bulk_g2s(s_vel, A.vel + e0 * 6, E * 24, &s_bar);
            bulk_g2s(s_cmd, A.commands + e0 * 3, E * 12, &s_bar);
            bulk_g2s(s_prog, A.progress + e0, E * 8, &s_bar);
        }
        __syncthreads();
        mbar_wait(&s_bar, 0);
    } else {
        coop_copy(s_prog, (const long long*)A.progress + e0, cnt, tid);
        coop_copy(s_pos, A.pos + e0 * 3, cnt * 3, tid);
        coop_copy(s_quat, A.quat + e0 * 4, cnt * 4, tid);
        coop_copy(s_vel, A.vel + e0 * 6, cnt * 6, tid);
        coop_copy(s_cmd, (const float*)A.commands + e0 * 3, cnt * 3, tid);
        coop_copy(s_q, (const float*)A.q + e0 * D, cnt * D, tid);
        coop_copy(s_qd, (const float*)A.qd + e0 * D, cnt * D, tid);
        coop_copy(s_act, (const float*)A.actions + e0 * D, cnt * D, tid);
        coop_copy(s_lact, (const float*)A.last_actions + e0 * D, cnt * D, tid);
        coop_copy(s_lqd, (const float*)A.last_qd + e0 * D, cnt * D, tid);
        __syncthreads();
    }

    // ---- quartet (e, r): adjacent lanes, so the reductions are two xor-shuffles ---------------------------
    const int e = tid >> 2, r = tid & 3;
    const bool live = e < cnt;
    const int el = live ? e : 0;  // dead quartets (ragged tile) compute on env 0 of the tile and store nothing
    float* row = o_obs + el * OBS;
    float p_acc = 0.f, p_rate = 0.f, p_cos = 0.f;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const int d = r + 4 * k;
        const float q = s_q[el * D + d], qd = s_qd[el * D + d], a = s_act[el * D + d];
        const float dflt = P.default_dof_pos[d];
        if (live) {
            row[12 + d] = clampf((q - dflt) * P.dof_pos_scale, P.clip_obs);
            row[24 + d] = clampf(qd * P.dof_vel_scale, P.clip_obs);
            row[36 + d] = clampf(a, P.clip_obs);
        }
        p_acc = p_acc + sq(s_lqd[el * D + d] - qd);       // anymal.py:272
        p_rate = p_rate + sq(s_lact[el * D + d] - a);     // anymal.py:273
        if (k == 0) p_cos = fabsf(q - dflt);              // anymal.py:274: dofs 0..3 = the four HAA joints
        // last_actions / last_dof_vel (anymal.py:280-281) are the staged actions / dof_vel tiles: drained below
    }
#pragma unroll
    for (int o = 1; o <= 2; o <<= 1) {
        p_acc += __shfl_xor_sync(0xffffffffu, p_acc, o);
        p_rate += __shfl_xor_sync(0xffffffffu, p_rate, o);
        p_cos += __shfl_xor_sync(0xffffffffu, p_cos, o);
    }
    const float quat[4] = {s_quat[4 * el], s_quat[4 * el + 1], s_quat[4 * el + 2], s_quat[4 * el + 3]};
    const float vlin[3] = {s_vel[6 * el], s_vel[6 * el + 1], s_vel[6 * el + 2]};
    const float vang[3] = {s_vel[6 * el + 3], s_vel[6 * el + 4], s_vel[6 * el + 5]};
    const float cmd[3] = {s_cmd[3 * el], s_cmd[3 * el + 1], s_cmd[3 * el + 2]};
    float bl[3], ba[3];
    qrot<true>(quat, vlin, bl);
    qrot<true>(quat, vang, ba);
    if (live) {
        if (r == 0) {
#pragma unroll
            for (int k = 0; k < 3; ++k) row[k] = clampf(bl[k] * P.lin_vel_scale, P.clip_obs);
        } else if (r == 1) {
#pragma unroll
            for (int k = 0; k < 3; ++k) row[3 + k] = clampf(ba[k] * P.ang_vel_scale, P.clip_obs);
        } else if (r == 2) {
            const float g[3] = {0.f, 0.f, -1.f};
            float pg[3];
            qrot<false>(quat, g, pg);
#pragma unroll
            for (int k = 0; k < 3; ++k) row[6 + k] = clampf(pg[k], P.clip_obs);
        } else {
            row[9] = clampf(cmd[0] * P.lin_vel_scale, P.clip_obs);
            row[10] = clampf(cmd[1] * P.lin_vel_scale, P.clip_obs);
            row[11] = clampf(cmd[2] * P.ang_vel_scale, P.clip_obs);
        }
    }
    float st_rew = 0.f, st_done = 0.f, st_len = 0.f, st_cnt = 0.f;
    if (live && r == 0) {
        const float lin_err = sq(cmd[0] - bl[0]) + sq(cmd[1] - bl[1]);
        const float ang_err = sq(cmd[2] - ba[2]);
        const float rew_lin = expf(-lin_err / 0.25f) * P.rew_lin_vel_xy;
        const float rew_ang = expf(-ang_err / 0.25f) * P.rew_ang_vel_z;
        const float rew_z = sq(bl[2]) * P.rew_lin_vel_z;
        float total = ((((rew_lin + rew_ang) + p_acc * P.rew_joint_acc) + p_rate * P.rew_action_rate) +
                       p_cos * P.rew_cosmetic) + rew_z;  // anymal.py:276, left to right
        total = fmaxf(total, 0.0f);                      // torch.clip(total, 0.0, None)
        const bool fallen = s_pos[3 * el + 2] < P.fallen_threshold;  // AnymalView.is_base_below_threshold
        if (fallen) total = -1.0f;
        const long long prog = s_prog[el] + (P.progress_increment ? 1 : 0);
        const bool time_out = (float)prog >= P.max_episode_length - 1.0f;
        const long long rst = (time_out || fallen) ? 1 : 0;
        s_prog[el] = prog;
        s_reset[el] = rst;
        o_rew[el] = total;
        st_rew = total; st_cnt = 1.f;
        st_done = rst ? 1.f : 0.f;
        st_len = rst ? (float)prog : 0.f;
    }
    if (A.stats != nullptr) {  // uniform
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            st_rew += __shfl_xor_sync(0xffffffffu, st_rew, o);
            st_done += __shfl_xor_sync(0xffffffffu, st_done, o);
            st_len += __shfl_xor_sync(0xffffffffu, st_len, o);
            st_cnt += __shfl_xor_sync(0xffffffffu, st_cnt, o);
        }
        if ((tid & 31) == 0 && st_cnt > 0.f) {
            double* slot = A.stats + ((blockIdx.x * (NT / 32) + (tid >> 5)) % OIGE_STATS_SLOTS) * 4;
            atomicAdd(slot + 0, (double)st_rew);
            atomicAdd(slot + 1, (double)st_done);
            atomicAdd(slot + 2, (double)st_len);
            atomicAdd(slot + 3, (double)st_cnt);
        }
    }

    // ---- drain the tile: last_actions <- actions, last_dof_vel <- dof_vel straight from the staged inputs ----
    if (bulk) {
        fence_async_smem();
        __syncthreads();
        if (tid == 0) {
            bulk_s2g(A.obs + e0 * OBS, o_obs, E * OBS * 4);
            bulk_s2g(A.last_actions + e0 * D, s_act, E * D * 4);
            bulk_s2g(A.last_qd + e0 * D, s_qd, E * D * 4);
            bulk_s2g(A.rew + e0, o_rew, E * 4);
            bulk_s2g(A.reset + e0, s_reset, E * 8);
            bulk_s2g(A.progress + e0, s_prog, E * 8);
            bulk_commit();
        }
    } else {
        __syncthreads();
        coop_copy(A.obs + e0 * OBS, (const float*)o_obs, cnt * OBS, tid);
        coop_copy(A.last_actions + e0 * D, (const float*)s_act, cnt * D, tid);
        coop_copy(A.last_qd + e0 * D, (const float*)s_qd, cnt * D, tid);
        coop_copy(A.rew + e0, (const float*)o_rew, cnt, tid);
        coop_copy(A.reset + e0, (const long long*)s_reset, cnt, tid);
        coop_copy(A.progress + e0, (const long long*)s_prog, cnt, tid);
    }
    if (tid == 0) {
        if (A.step_counter != nullptr && blockIdx.x == 0) *A.step_counter += 1;
        if (bulk) bulk_wait_read0();  // smem must stay alive until the TMA engine has read it
    }
}

// ---- reset_idx arithmetic for one quartet lane (anymal.py:192-227) -------------------------------------------
// `row` indexes the caller-supplied draws (env index for the masked form, list position for the idx form).
__device__ __forceinline__ void reset_env_lane(const OigeAnymalParams& P, const AnyArgs& A, unsigned long long counter,
                                               long long env, long long row, int r) {
    const uint64_t gid = (uint64_t)(A.env_id_offset + env);
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const int d = r + 4 * k;
        const float u = A.rand_vel != nullptr ? A.rand_vel[row * D + d] : reset_uniform(A.seed, counter, gid, d, 1);
        const float dflt = P.default_dof_pos[d];
        A.q[env * D + d] = dflt;                                   // dof_pos = default_dof_pos[env_ids]
        A.qd[env * D + d] = (0.1f - (-0.1f)) * u + (-0.1f);        // torch_rand_float(-0.1, 0.1, ...)
        A.targets[env * D + d] = dflt;                             // current_targets[env_ids] = dof_pos
        A.last_actions[env * D + d] = 0.f;
        A.last_qd[env * D + d] = 0.f;
    }
    if (r < 3) {  // commands x, y, yaw: torch_rand_float(lo, hi, (k, 1)) each
        const float u = A.rand_cmd != nullptr ? A.rand_cmd[row * 3 + r] : reset_uniform(A.seed, counter, gid, r, 0);
        const float lo = r == 0 ? P.command_x[0] : (r == 1 ? P.command_y[0] : P.command_yaw[0]);
        const float hi = r == 0 ? P.command_x[1] : (r == 1 ? P.command_y[1] : P.command_yaw[1]);
        A.commands[env * 3 + r] = (hi - lo) * u + lo;
        A.wpos[env * 3 + r] = A.init_pos[env * 3 + r];
        A.wvel[env * 6 + r] = 0.f;
        A.wvel[env * 6 + 3 + r] = 0.f;
    }
    A.wquat[env * 4 + r] = A.init_quat[env * 4 + r];
    if (r == 3) {
        A.reset[env] = 0;
        A.progress[env] = 0;
    }
}

__global__ void __launch_bounds__(NT)
anymal_pre_kernel(const __grid_constant__ OigeAnymalParams P, const __grid_constant__ AnyArgs A) {
    const long long env = (long long)blockIdx.x * E + (threadIdx.x >> 2);
    const int r = threadIdx.x & 3;
    const bool live = env < A.n;
    const bool flagged = live && A.reset[env] != 0;
    __syncwarp();  // every lane of the quartet has read the flag before lane 3 clears it
    if (flagged) {
        const unsigned long long counter = A.counter_dev != nullptr ? *A.counter_dev : A.counter;
        reset_env_lane(P, A, counter, env, env, r);
    }
    if (live && A.actions_in != nullptr) {
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            const int d = r + 4 * k;
            const float a = clampf(A.actions_in[env * D + d], A.clip_actions);  // vec_env_rlgames.py:57
            A.actions[env * D + d] = a;                                         // anymal.py:187
            const float base = flagged ? P.default_dof_pos[d] : A.targets[env * D + d];
            const float t = base + (P.action_scale * a) * P.dt;                 // anymal.py:188
            A.targets[env * D + d] = fminf(fmaxf(t, P.dof_limits_lower[d]), P.dof_limits_upper[d]);
        }
    }
}

__global__ void __launch_bounds__(NT)
anymal_reset_idx_kernel(const __grid_constant__ OigeAnymalParams P, const __grid_constant__ AnyArgs A) {
    const long long i = (long long)blockIdx.x * E + (threadIdx.x >> 2);
    if (i >= A.n) return;
    reset_env_lane(P, A, A.counter, A.env_ids[i], i, threadIdx.x & 3);
}

bool aligned16(const void* p) { return ((uintptr_t)p & 15) == 0; }

int fill(AnyArgs& A, const char* who, const OigeAnymalParams* P, const OigeState* st, const OigeAnymalBuffers* b) {
    if (!P || !st || !b) return fail(OIGE_E_NULL, "%s: null struct pointer", who);
    if (!st->root_pos || !st->root_quat || !st->root_vel || !st->dof_pos || !st->dof_vel)
        return fail(OIGE_E_NULL, "%s: null state tensor", who);
    if (!b->reset_buf || !b->progress_buf || !b->actions || !b->last_actions || !b->last_dof_vel || !b->commands)
        return fail(OIGE_E_NULL, "%s: null task buffer", who);
    A.pos = st->root_pos; A.quat = st->root_quat; A.vel = st->root_vel;
    A.wpos = st->root_pos; A.wquat = st->root_quat; A.wvel = st->root_vel;
    A.q = st->dof_pos; A.qd = st->dof_vel;
    A.obs = b->obs; A.rew = b->rew;
    A.reset = reinterpret_cast<long long*>(b->reset_buf);
    A.progress = reinterpret_cast<long long*>(b->progress_buf);
    A.actions = b->actions; A.last_actions = b->last_actions; A.last_qd = b->last_dof_vel;
    A.commands = b->commands; A.targets = b->current_targets;
    A.stats = b->episode_stats;
    return OIGE_OK;
}

int fill_reset(AnyArgs& A, const char* who, const OigeAnymalBuffers* b, const OigeResetInit* init) {
    if (!init || !init->initial_root_pos || !init->initial_root_rot)
        return fail(OIGE_E_NULL, "%s: null initial root pose", who);
    if (!b->current_targets) return fail(OIGE_E_NULL, "%s: null current_targets", who);
    A.init_pos = init->initial_root_pos; A.init_quat = init->initial_root_rot;
    return OIGE_OK;
}

}  // namespace

int anymal_post(const OigeAnymalParams* P, const OigeState* st, const OigeAnymalBuffers* b, int64_t n,
                cudaStream_t stream, bool bump_counter) {
    AnyArgs A{};
    if (int rc = fill(A, "anymal_obs_reward_reset", P, st, b)) return rc;
    if (!b->obs || !b->rew) return fail(OIGE_E_NULL, "anymal_obs_reward_reset: null obs / rew buffer");
    A.n = n;
    A.step_counter = bump_counter ? reinterpret_cast<unsigned long long*>(b->reset_counter) : nullptr;
    A.use_bulk = 1;
    const void* ptrs[] = {A.pos, A.quat, A.vel, A.q, A.qd, A.obs, A.rew, A.reset, A.progress, A.actions,
                          A.last_actions, A.last_qd, A.commands};
    for (const void* p : ptrs)
        if (!aligned16(p)) A.use_bulk = 0;  // plain coalesced path for oddly offset views
    anymal_post_kernel<<<(unsigned)((n + E - 1) / E), NT, 0, stream>>>(*P, A);
    count_launch();
    return check_launch("anymal_post_kernel");
}

int anymal_pre(const OigeAnymalParams* P, const OigeState* st, const OigeAnymalBuffers* b, const OigeResetInit* init,
               const float* actions_in, float clip_actions, const float* rand_vel, const float* rand_cmd, uint64_t seed,
               uint64_t counter, int64_t env_id_offset, int64_t n, cudaStream_t stream, bool counter_from_device) {
    AnyArgs A{};
    if (int rc = fill(A, "anymal_reset_masked", P, st, b)) return rc;
    if (int rc = fill_reset(A, "anymal_reset_masked", b, init)) return rc;
    A.actions_in = actions_in; A.clip_actions = clip_actions; A.rand_vel = rand_vel; A.rand_cmd = rand_cmd;
    A.seed = seed; A.counter = counter; A.env_id_offset = env_id_offset; A.n = n;
    A.counter_dev = counter_from_device ? reinterpret_cast<const unsigned long long*>(b->reset_counter) : nullptr;
    anymal_pre_kernel<<<(unsigned)((n + E - 1) / E), NT, 0, stream>>>(*P, A);
    count_launch();
    return check_launch("anymal_pre_kernel");
}

int anymal_reset_idx(const OigeAnymalParams* P, const OigeState* st, const OigeAnymalBuffers* b,
                     const OigeResetInit* init, const int64_t* env_ids, int64_t k, const float* rand_vel,
                     const float* rand_cmd, uint64_t seed, uint64_t counter, int64_t env_id_offset,
                     cudaStream_t stream) {
    AnyArgs A{};
    if (int rc = fill(A, "anymal_reset_idx", P, st, b)) return rc;
    if (int rc = fill_reset(A, "anymal_reset_idx", b, init)) return rc;
    if (!env_ids) return fail(OIGE_E_NULL, "anymal_reset_idx: null env_ids");
    A.env_ids = reinterpret_cast<const long long*>(env_ids); A.rand_vel = rand_vel; A.rand_cmd = rand_cmd;
    A.seed = seed; A.counter = counter; A.env_id_offset = env_id_offset; A.n = k;
    anymal_reset_idx_kernel<<<(unsigned)((k + E - 1) / E), NT, 0, stream>>>(*P, A);
    count_launch();
    return check_launch("anymal_reset_idx_kernel");
}

}  // namespace oige
