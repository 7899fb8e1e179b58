// Cartpole task passes (reference: tasks/cartpole.py:80-162), one thread per environment.
//
//   cartpole_post_kernel   progress += 1, get_observations, calculate_metrics, is_done (cartpole.py:80-99,143-162;
//                          rl_task.py:242) -- 24 B read + 36 B written per env, pure streaming.
//   cartpole_pre_kernel    pre_physics_step (cartpole.py:101-108): device-side masked reset_idx (no nonzero()
//                          host sync) + action clamp (vec_env_rlgames.py:57) + efforts on the cart DoF.
//   cartpole_reset_idx_kernel  reset_idx(env_ids) (cartpole.py:110-130), the reference's calling form.
//
// Compiled with -fmad=false: every expression is evaluated in the reference's (torch eager) operation order with
// one rounding per op, so obs / rew / reset values / efforts are BIT-exact against oracle/cartpole_oracle.py.
// D = 2 makes each [N,2] row one 8-byte word and each obs row one 16-byte word: plain vector loads/stores are
// already fully coalesced, no shared-memory staging is needed.
#include <math_constants.h>

#include "oige_common.cuh"

namespace oige {
namespace {

struct CartArgs {
    float2* q;
    float2* qd;
    float4* obs;
    float* rew;
    long long* reset;
    long long* progress;
    float2* efforts;
    double* stats;
    unsigned long long* step_counter;       // bumped at the end of a whole step (CUDA-graph replays) or null
    const unsigned long long* counter_dev;  // if set, overrides `counter`
    const float* actions_in;                // [N,1] or null
    const float* rand;                      // [N,4] / [k,4] raw U[0,1) draws or null (Philox)
    const long long* env_ids;               // idx form only
    unsigned long long seed, counter;
    long long env_id_offset;
    long long n;
    float clip_actions;
};

__device__ __forceinline__ float pick(float2 v, int i) { return i == 0 ? v.x : v.y; }
__device__ __forceinline__ float2 place(float cart, float pole, int cart_dof) {
    return cart_dof == 0 ? make_float2(cart, pole) : make_float2(pole, cart);
}
__device__ __forceinline__ float clampf(float x, float c) { return fminf(fmaxf(x, -c), c); }

// reset_idx arithmetic (cartpole.py:115-121): python scalars are rounded to fp32 before they meet the tensor
__device__ __forceinline__ void reset_values(const float u[4], float& cart_q, float& pole_q, float& cart_v, float& pole_v) {
    cart_q = 1.0f * (1.0f - 2.0f * u[0]);
    pole_q = (float)(0.125 * 3.141592653589793) * (1.0f - 2.0f * u[1]);
    cart_v = 0.5f * (1.0f - 2.0f * u[2]);
    pole_v = (float)(0.25 * 3.141592653589793) * (1.0f - 2.0f * u[3]);
}

__device__ __forceinline__ void draw4(const CartArgs& A, unsigned long long counter, long long env, long long row, float u[4]) {
    if (A.rand != nullptr) {
        const float4 r = *reinterpret_cast<const float4*>(A.rand + row * 4);
        u[0] = r.x; u[1] = r.y; u[2] = r.z; u[3] = r.w;
    } else {  // one Philox block per (env, counter): the same keying as reset_uniform(dof 0..3, which 0)
        const uint64_t gid = (uint64_t)(A.env_id_offset + env);
        uint32_t o[4];
        philox4x32_10((uint32_t)gid, (uint32_t)(gid >> 32), (uint32_t)counter, ((uint32_t)(counter >> 32) << 8),
                      (uint32_t)A.seed, (uint32_t)(A.seed >> 32), o);
        u[0] = u01(o[0]); u[1] = u01(o[1]); u[2] = u01(o[2]); u[3] = u01(o[3]);
    }
}

__global__ void __launch_bounds__(256)
cartpole_post_kernel(const __grid_constant__ OigeCartpoleParams P, const __grid_constant__ CartArgs A) {
    const long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    float st_rew = 0.f, st_done = 0.f, st_len = 0.f, st_cnt = 0.f;
    if (e < A.n) {
        const float2 q = A.q[e], qd = A.qd[e];
        const float cart_pos = pick(q, P.cart_dof), cart_vel = pick(qd, P.cart_dof);
        const float pole_pos = pick(q, P.pole_dof), pole_vel = pick(qd, P.pole_dof);
        const long long prog = A.progress[e] + (P.progress_increment ? 1 : 0);
        // calculate_metrics (cartpole.py:149-151), evaluated left to right
        float rew = 1.0f - pole_pos * pole_pos;
        rew = rew - 0.01f * fabsf(cart_vel);
        rew = rew - 0.005f * fabsf(pole_vel);
        const bool out_of_rail = fabsf(cart_pos) > P.reset_dist;
        const bool fallen = fabsf(pole_pos) > (float)(3.141592653589793 / 2.0);
        if (out_of_rail) rew = -2.0f;
        if (fallen) rew = -2.0f;
        // is_done (cartpole.py:159-162): the incoming reset flag is not carried over
        long long rst = out_of_rail ? 1 : 0;
        if (fallen) rst = 1;
        if ((float)prog >= P.max_episode_length) rst = 1;
        const float c = P.clip_obs;
        A.obs[e] = make_float4(clampf(cart_pos, c), clampf(cart_vel, c), clampf(pole_pos, c), clampf(pole_vel, c));
        A.rew[e] = rew;
        A.reset[e] = rst;
        A.progress[e] = prog;
        st_rew = rew; st_cnt = 1.f;
        st_done = rst != 0 ? 1.f : 0.f;
        st_len = rst != 0 ? (float)prog : 0.f;
    }
    if (A.stats != nullptr) {  // uniform
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            st_rew += __shfl_xor_sync(0xffffffffu, st_rew, o);
            st_done += __shfl_xor_sync(0xffffffffu, st_done, o);
            st_len += __shfl_xor_sync(0xffffffffu, st_len, o);
            st_cnt += __shfl_xor_sync(0xffffffffu, st_cnt, o);
        }
        if ((threadIdx.x & 31) == 0 && st_cnt > 0.f) {
            const unsigned warp = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
            double* slot = A.stats + (warp % OIGE_STATS_SLOTS) * 4;
            atomicAdd(slot + 0, (double)st_rew);
            atomicAdd(slot + 1, (double)st_done);
            atomicAdd(slot + 2, (double)st_len);
            atomicAdd(slot + 3, (double)st_cnt);
        }
    }
    if (A.step_counter != nullptr && blockIdx.x == 0 && threadIdx.x == 0) *A.step_counter += 1;
}

__global__ void __launch_bounds__(256)
cartpole_pre_kernel(const __grid_constant__ OigeCartpoleParams P, const __grid_constant__ CartArgs A) {
    const long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= A.n) return;
    if (A.reset[e] != 0) {
        const unsigned long long counter = A.counter_dev != nullptr ? *A.counter_dev : A.counter;
        float u[4], cq, pq, cv, pv;
        draw4(A, counter, e, e, u);
        reset_values(u, cq, pq, cv, pv);
        A.q[e] = place(cq, pq, P.cart_dof);
        A.qd[e] = place(cv, pv, P.cart_dof);
        A.reset[e] = 0;
        A.progress[e] = 0;
    }
    if (A.actions_in != nullptr) {
        const float a = clampf(A.actions_in[e], A.clip_actions);  // vec_env_rlgames.py:57
        A.efforts[e] = place(P.max_push_effort * a, 0.f, P.cart_dof);  // cartpole.py:104-105
    }
}

__global__ void __launch_bounds__(256)
cartpole_reset_idx_kernel(const __grid_constant__ OigeCartpoleParams P, const __grid_constant__ CartArgs A) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= A.n) return;
    const long long e = A.env_ids[i];
    float u[4], cq, pq, cv, pv;
    draw4(A, A.counter, e, i, u);
    reset_values(u, cq, pq, cv, pv);
    A.q[e] = place(cq, pq, P.cart_dof);
    A.qd[e] = place(cv, pv, P.cart_dof);
    A.reset[e] = 0;
    A.progress[e] = 0;
}

int fill(CartArgs& A, const char* who, const OigeCartpoleParams* P, const OigeState* st, const OigeCartpoleBuffers* b) {
    if (!P || !st || !b) return fail(OIGE_E_NULL, "%s: null struct pointer", who);
    if (!st->dof_pos || !st->dof_vel || !b->reset_buf || !b->progress_buf)
        return fail(OIGE_E_NULL, "%s: null state / flag buffer", who);
    if (!((P->cart_dof == 0 && P->pole_dof == 1) || (P->cart_dof == 1 && P->pole_dof == 0)))
        return fail(OIGE_E_SHAPE, "%s: (cart_dof, pole_dof) = (%d, %d) must be a permutation of (0, 1)", who,
                    P->cart_dof, P->pole_dof);
    if (((uintptr_t)st->dof_pos | (uintptr_t)st->dof_vel) & 7)
        return fail(OIGE_E_ALIGN, "%s: dof_pos / dof_vel must be 8-byte aligned", who);
    A.q = reinterpret_cast<float2*>(st->dof_pos);
    A.qd = reinterpret_cast<float2*>(st->dof_vel);
    A.obs = reinterpret_cast<float4*>(b->obs);
    A.rew = b->rew;
    A.reset = reinterpret_cast<long long*>(b->reset_buf);
    A.progress = reinterpret_cast<long long*>(b->progress_buf);
    A.efforts = reinterpret_cast<float2*>(b->efforts);
    A.stats = b->episode_stats;
    return OIGE_OK;
}

}  // namespace

int cartpole_post(const OigeCartpoleParams* P, const OigeState* st, const OigeCartpoleBuffers* b, int64_t n,
                  cudaStream_t stream, bool bump_counter) {
    CartArgs A{};
    if (int rc = fill(A, "cartpole_obs_reward_reset", P, st, b)) return rc;
    if (!b->obs || !b->rew) return fail(OIGE_E_NULL, "cartpole_obs_reward_reset: null obs / rew buffer");
    if ((uintptr_t)b->obs & 15) return fail(OIGE_E_ALIGN, "cartpole_obs_reward_reset: obs must be 16-byte aligned");
    A.n = n;
    A.step_counter = bump_counter ? reinterpret_cast<unsigned long long*>(b->reset_counter) : nullptr;
    cartpole_post_kernel<<<(unsigned)((n + 255) / 256), 256, 0, stream>>>(*P, A);
    count_launch();
    return check_launch("cartpole_post_kernel");
}

int cartpole_pre(const OigeCartpoleParams* P, const OigeState* st, const OigeCartpoleBuffers* b, const float* actions_in,
                 float clip_actions, const float* rand, uint64_t seed, uint64_t counter, int64_t env_id_offset, int64_t n,
                 cudaStream_t stream, bool counter_from_device) {
    CartArgs A{};
    if (int rc = fill(A, "cartpole_reset_masked", P, st, b)) return rc;
    if (actions_in && (!b->efforts || ((uintptr_t)b->efforts & 7)))
        return fail(OIGE_E_NULL, "cartpole_reset_masked: efforts buffer missing or not 8-byte aligned");
    if (rand && ((uintptr_t)rand & 15)) return fail(OIGE_E_ALIGN, "cartpole_reset_masked: rand must be 16-byte aligned");
    A.actions_in = actions_in; A.clip_actions = clip_actions; A.rand = rand;
    A.seed = seed; A.counter = counter; A.env_id_offset = env_id_offset; A.n = n;
    A.counter_dev = counter_from_device ? reinterpret_cast<const unsigned long long*>(b->reset_counter) : nullptr;
    cartpole_pre_kernel<<<(unsigned)((n + 255) / 256), 256, 0, stream>>>(*P, A);
    count_launch();
    return check_launch("cartpole_pre_kernel");
}

int cartpole_reset_idx(const OigeCartpoleParams* P, const OigeState* st, const OigeCartpoleBuffers* b,
                       const int64_t* env_ids, int64_t k, const float* rand, uint64_t seed, uint64_t counter,
                       int64_t env_id_offset, cudaStream_t stream) {
    CartArgs A{};
    if (int rc = fill(A, "cartpole_reset_idx", P, st, b)) return rc;
    if (!env_ids) return fail(OIGE_E_NULL, "cartpole_reset_idx: null env_ids");
    if (rand && ((uintptr_t)rand & 15)) return fail(OIGE_E_ALIGN, "cartpole_reset_idx: rand must be 16-byte aligned");
    A.env_ids = reinterpret_cast<const long long*>(env_ids); A.rand = rand;
    A.seed = seed; A.counter = counter; A.env_id_offset = env_id_offset; A.n = k;
    cartpole_reset_idx_kernel<<<(unsigned)((k + 255) / 256), 256, 0, stream>>>(*P, A);
    count_launch();
    return check_launch("cartpole_reset_idx_kernel");
}

}  // namespace oige
