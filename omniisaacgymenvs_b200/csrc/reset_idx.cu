// K3 -- device-side reset (no host sync).
//
// Replaces LocomotionTask.reset_idx (tasks/shared/locomotion.py:116-145) and removes the
// `reset_buf.nonzero()` + `len()` device->host sync of pre_physics_step (locomotion.py:104-106).
//
// masked form: each warp scans 32 consecutive reset flags with one coalesced load, ballots the
// non-zero ones, and then the whole warp rewrites the rows of each flagged env cooperatively
// (lane d writes dof d), so the row writes stay coalesced and no thread idles on a rare branch.
// idx form: one warp per listed env id (the reference's calling convention).
#include "oige_common.cuh"

namespace oige {
namespace {

struct K3Args {
    float* pos; float* quat; float* vel; float* q; float* qd;
    float* pot; float* prev_pot;
    long long* reset; long long* progress;
    const float* init_pos; const float* init_rot; const float* init_q;
    const float* rand_pos; const float* rand_vel;   // [rows, D] raw uniforms or null
    const long long* env_ids;                        // idx form only
    unsigned long long seed, counter;
    const unsigned long long* counter_dev;           // if set, overrides `counter` (CUDA-graph replays)
    const float* actions_in; float* actions_out; float clip;  // masked form: fused action clamp
    int num_dof;
    long long env_id_offset;
    long long n;                                     // envs (masked) or ids (idx)
};

// one env, executed by a full warp.  `rrow` = row of the rand tensors to consume.
__device__ __forceinline__ void reset_one_env(const OigeTaskParams& P, const K3Args& A, long long env,
                                              long long rrow, int lane) {
    const int D = P.num_dof;
    if (lane < D) {
        // torch_rand_float(-0.2, 0.2): (upper - lower) * rand + lower   (locomotion.py:120)
        const float up = A.rand_pos ? A.rand_pos[rrow * D + lane]
                                    : reset_uniform(A.seed, A.counter, (uint64_t)(A.env_id_offset + env), lane, 0);
        const float uv = A.rand_vel ? A.rand_vel[rrow * D + lane]
                                    : reset_uniform(A.seed, A.counter, (uint64_t)(A.env_id_offset + env), lane, 1);
        const float noise = __fadd_rn(__fmul_rn(0.2f - (-0.2f), up), -0.2f);
        float qv = __fadd_rn(A.init_q[env * D + lane], noise);
        // tensor_clamp = max(min(t, hi), lo)                              (locomotion.py:121-123)
        qv = fmaxf(fminf(qv, P.dof_limits_upper[lane]), P.dof_limits_lower[lane]);
        A.q[env * D + lane] = qv;
        A.qd[env * D + lane] = __fadd_rn(__fmul_rn(0.1f - (-0.1f), uv), -0.1f);  // locomotion.py:124
    }
    if (lane < 3) A.pos[env * 3 + lane] = A.init_pos[env * 3 + lane];
    if (lane < 4) A.quat[env * 4 + lane] = A.init_rot[env * 4 + lane];
    if (lane < 6) A.vel[env * 6 + lane] = 0.0f;
    if (lane == 0) {
        // potentials = -|| (targets - initial_root_pos)_xy || / dt         (locomotion.py:136-139)
        const float tx = __fsub_rn(P.targets[0], A.init_pos[env * 3]);
        const float ty = __fsub_rn(P.targets[1], A.init_pos[env * 3 + 1]);
        const float nrm = sqrtf(fmaf(0.0f, 0.0f, fmaf(ty, ty, __fmul_rn(tx, tx))));
        const float p = __fdiv_rn(-nrm, P.dt);
        A.prev_pot[env] = p;
        A.pot[env] = p;
        A.reset[env] = 0;     // locomotion.py:142-143
        A.progress[env] = 0;
    }
}

// One CTA (4 warps) per 32 consecutive envs: the action rows are clamped by all 128 threads; every warp ballots the
// same 32 reset flags and takes every 4th flagged env, so a group with several resets (random policies reset ~4 % of
// the envs per step) does not serialise them on one warp.
__global__ void __launch_bounds__(128) reset_masked_kernel(const __grid_constant__ OigeTaskParams P,
                                                           const __grid_constant__ K3Args A) {
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    const long long base = (long long)blockIdx.x * 32;
    pdl_trigger();  // the articulation kernel may stage its model table while this one runs
    if (base >= A.n) return;
    if (A.actions_in != nullptr) {
        // VecEnvRLGames.step's torch.clamp(actions, -clip, clip) (envs/vec_env_rlgames.py:57) for the 32 envs
        // of this group: one contiguous, coalesced chunk of the [N,D] action tensor
        const long long cnt = ((A.n - base) < 32 ? (A.n - base) : 32) * A.num_dof;
        const float* in = A.actions_in + base * A.num_dof;
        float* out = A.actions_out + base * A.num_dof;
        for (long long i = threadIdx.x; i < cnt; i += 128) {
            const float v = in[i];
            out[i] = v < -A.clip ? -A.clip : (v > A.clip ? A.clip : v);
        }
    }
    const long long e = base + lane;
    const bool flag = (e < A.n) && (A.reset[e] != 0);
    unsigned m = __ballot_sync(0xffffffffu, flag);
    if (m == 0) return;
    K3Args B = A;
    if (A.counter_dev != nullptr) B.counter = *A.counter_dev;
    // every warp must have read the flags before any warp clears one (reset_one_env zeroes reset[env])
    __syncthreads();
    for (int k = 0; m; ++k) {
        const int b = __ffs(m) - 1;
        m &= m - 1;
        if ((k & 3) == wib) reset_one_env(P, B, base + b, base + b, lane);
    }
}

__global__ void __launch_bounds__(128) reset_idx_kernel(const __grid_constant__ OigeTaskParams P,
                                                        const __grid_constant__ K3Args A) {
    const int lane = threadIdx.x & 31;
    const long long warp = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (warp >= A.n) return;
    reset_one_env(P, A, A.env_ids[warp], warp, lane);
}

int fill(K3Args& A, const OigeState* st, const OigeTaskBuffers* b, const OigeResetInit* init) {
    if (!st || !b || !init) return fail(OIGE_E_NULL, "reset: null struct pointer");
    A.pos = st->root_pos; A.quat = st->root_quat; A.vel = st->root_vel; A.q = st->dof_pos; A.qd = st->dof_vel;
    A.pot = b->potentials; A.prev_pot = b->prev_potentials;
    A.reset = reinterpret_cast<long long*>(b->reset_buf);
    A.progress = reinterpret_cast<long long*>(b->progress_buf);
    A.init_pos = init->initial_root_pos; A.init_rot = init->initial_root_rot; A.init_q = init->initial_dof_pos;
    if (!A.pos || !A.quat || !A.vel || !A.q || !A.qd || !A.pot || !A.prev_pot || !A.reset || !A.progress ||
        !A.init_pos || !A.init_rot || !A.init_q)
        return fail(OIGE_E_NULL, "reset: null device pointer");
    return OIGE_OK;
}

}  // namespace

int k3_masked(const OigeTaskParams* P, const OigeState* st, const OigeTaskBuffers* b, const OigeResetInit* init,
              const float* rand_pos, const float* rand_vel, uint64_t seed, uint64_t counter,
              int64_t env_id_offset, int64_t n, cudaStream_t stream, const float* actions_in, float clip) {
    K3Args A{};
    if (int rc = fill(A, st, b, init)) return rc;
    if (P->num_dof < 1 || P->num_dof > OIGE_MAX_DOF) return fail(OIGE_E_SHAPE, "reset: num_dof %d", P->num_dof);
    A.rand_pos = rand_pos; A.rand_vel = rand_vel; A.env_ids = nullptr;
    A.actions_in = actions_in; A.actions_out = b->actions; A.clip = clip; A.num_dof = P->num_dof;
    A.counter_dev = actions_in ? reinterpret_cast<const unsigned long long*>(b->reset_counter) : nullptr;
    if (actions_in && !b->actions) return fail(OIGE_E_NULL, "reset: null actions buffer");
    A.seed = seed; A.counter = counter; A.env_id_offset = env_id_offset; A.n = n;
    const unsigned blocks = (unsigned)((n + 31) / 32);
    reset_masked_kernel<<<blocks, 128, 0, stream>>>(*P, A);
    count_launch();
    return check_launch("reset_masked_kernel");
}

int k3_idx(const OigeTaskParams* P, const OigeState* st, const OigeTaskBuffers* b, const OigeResetInit* init,
           const int64_t* env_ids, int64_t k, const float* rand_pos, const float* rand_vel, uint64_t seed,
           uint64_t counter, int64_t env_id_offset, cudaStream_t stream) {
    K3Args A{};
    if (int rc = fill(A, st, b, init)) return rc;
    if (!env_ids) return fail(OIGE_E_NULL, "reset_idx: null env_ids");
    if (P->num_dof < 1 || P->num_dof > OIGE_MAX_DOF) return fail(OIGE_E_SHAPE, "reset: num_dof %d", P->num_dof);
    A.rand_pos = rand_pos; A.rand_vel = rand_vel;
    A.env_ids = reinterpret_cast<const long long*>(env_ids);
    A.seed = seed; A.counter = counter; A.env_id_offset = env_id_offset; A.n = k;
    const unsigned blocks = (unsigned)((k + 3) / 4);
    reset_idx_kernel<<<blocks, 128, 0, stream>>>(*P, A);
    count_launch();
    return check_launch("reset_idx_kernel");
}

}  // namespace oige
