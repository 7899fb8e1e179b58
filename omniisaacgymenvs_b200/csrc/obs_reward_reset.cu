// K1 -- fused observation / reward / termination pass for the locomotion tasks (sm_100a).
//
// Replaces ~390 aten launches of the reference's TorchScript path with ONE streaming kernel:
//   progress_buf += 1                                   tasks/base/rl_task.py:242
//   jit get_observations (+ 3x normalize_angle)         tasks/shared/locomotion.py:190-254
//   get_dof_at_limit_cost                               tasks/humanoid.py:120-127, tasks/ant.py:92-95
//   jit calculate_metrics                               tasks/shared/locomotion.py:271-321
//   jit is_done                                         tasks/shared/locomotion.py:257-268
//   obs clamp of VecEnvRLGames._process_data            envs/vec_env_rlgames.py:42
//
// Data movement: a CTA owns a tile of E consecutive envs.  Every state tensor is [N,k] row-major,
// so a tile of each tensor is ONE contiguous 16-byte-aligned chunk: it is fetched with a 1-D bulk
// async copy (cp.async.bulk, the TMA engine; SASS UBLKCP) that completes on an mbarrier, and every
// output tile -- including the [E,OBS] observation block whose 348-byte rows are not 16-byte
// aligned individually -- leaves through one bulk store.  No thread ever issues a strided or
// misaligned global access.  The kernel is HBM-bound: 772 B (Humanoid) / 556 B (Ant) per env.
//
// Arithmetic is evaluated in the reference's operation order with FMA contraction disabled for
// this translation unit (-fmad=false); the two places where torch's CPU kernels themselves use
// FMA (vector norm, cross product) use explicit fmaf.
#include <math.h>

#include "oige_common.cuh"

namespace oige {

namespace {

constexpr float kTwoPi = 6.283185307179586f;  // (float)(2*pi), as torch casts the python scalar

__device__ __forceinline__ float normalize_angle(float x) {  // locomotion.py:190-192
    return atan2f(sinf(x), cosf(x));
}

// torch.cross CPU kernel contracts a1*b2 - a2*b1 into fma(a1, b2, -(a2*b1))
__device__ __forceinline__ void cross3(const float a[3], const float b[3], float c[3]) {
    c[0] = fmaf(a[1], b[2], -(a[2] * b[1]));
    c[1] = fmaf(a[2], b[0], -(a[0] * b[2]));
    c[2] = fmaf(a[0], b[1], -(a[1] * b[0]));
}

// v(2w^2-1) +- 2w (q x v) + 2 q (q.v)      (quat_rotate / quat_rotate_inverse)
template <bool INVERSE>
__device__ __forceinline__ void quat_rotate(const float q[4], const float v[3], float out[3]) {
    const float w = q[0];
    const float qv[3] = {q[1], q[2], q[3]};
    const float s = 2.0f * (w * w) - 1.0f;
    float cr[3];
    cross3(qv, v, cr);
    const float dot = (qv[0] * v[0] + qv[1] * v[1]) + qv[2] * v[2];
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        const float a = v[i] * s;
        const float b = (cr[i] * w) * 2.0f;
        const float c = (qv[i] * dot) * 2.0f;
        out[i] = INVERSE ? ((a - b) + c) : ((a + b) + c);
    }
}

// factored 9-multiply Hamilton product (quat_mul)
__device__ __forceinline__ void quat_mul(const float a[4], const float b[4], float o[4]) {
    const float w1 = a[0], x1 = a[1], y1 = a[2], z1 = a[3];
    const float w2 = b[0], x2 = b[1], y2 = b[2], z2 = b[3];
    const float ww = (z1 + x1) * (x2 + y2);
    const float yy = (w1 - y1) * (w2 + z2);
    const float zz = (w1 + y1) * (w2 - z2);
    const float xx = (ww + yy) + zz;
    const float qq = 0.5f * (xx + (z1 - x1) * (x2 - y2));
    o[0] = (qq - ww) + (z1 - y1) * (y2 - z2);
    o[1] = (qq - xx) + (x1 + w1) * (x2 + w2);
    o[2] = (qq - yy) + (w1 - x1) * (y2 + z2);
    o[3] = (qq - zz) + (z1 + y1) * (w2 - x2);
}

__device__ __forceinline__ float clampf(float v, float c) {  // torch.clamp semantics (NaN propagates)
    return v < -c ? -c : (v > c ? c : v);
}

template <int D, int S>
struct Tile {
    static constexpr int OBS = 12 + 3 * D + 6 * S;
    // per-env float counts of each staged tensor
    static constexpr int IN_F32 = 3 + 4 + 6 + D + D + 6 * S + D + 1;  // pos quat vel q qd sens act pot
    static constexpr int OUT_F32 = OBS + 1 + 1 + 3 + 3;                // obs rew pot up head (prev_pot = staged pot)
    static constexpr int BYTES_PER_ENV = 4 * (IN_F32 + OUT_F32) + 16;  // + progress, reset (int64 each)
};

struct K1Args {
    const float* pos;
    const float* quat;
    const float* vel;
    const float* q;
    const float* qd;
    const float* sens;
    const float* act;
    float* pot;
    float* prev_pot;
    float* obs;
    float* rew;
    float* up;
    float* head;
    long long* reset;
    long long* progress;
    double* stats;  // [OIGE_STATS_SLOTS][4] or null
    unsigned long long* step_counter;  // device-side reset counter to bump (whole-step calls only) or null
    long long n;
    int use_bulk;
};

template <typename T>
__device__ __forceinline__ void coop_load(T* dst, const T* src, int count, int tid, int nthr) {
    for (int i = tid; i < count; i += nthr) dst[i] = src[i];
}
template <typename T>
__device__ __forceinline__ void coop_store(T* dst, const T* src, int count, int tid, int nthr) {
    for (int i = tid; i < count; i += nthr) dst[i] = src[i];
}

// Correctly rounded a / b from y = RN(1/b): q0 = RN(a*y), r = a - q0*b (exact, one FMA), q1 = RN(q0 + r*y)
// (Markstein 1990: q1 == RN(a/b) whenever y is the correctly rounded reciprocal and nothing over/underflows).
// Three instructions instead of the ~15 of the IEEE division subroutine.  Inputs outside the normal
// range (|a/b| near the fp32 limits, non-finite values) take the IEEE path so NaN/inf propagate as in torch.
__device__ __forceinline__ float div_rn(float a, float b, float y) {
    const float q0 = a * y;
    const float aq = fabsf(q0);
    if (!(aq > 1e-30f && aq < 1e30f)) return a / b;
    const float r = fmaf(-q0, b, a);
    return fmaf(r, y, q0);
}

// wrap an angle x (|x| < 4 pi) into (-pi, pi] the way atan2(sin x, cos x) does, without the three
// transcendentals: sin/cos see the FLOAT x, so the exact result is x -/+ 2*pi (true pi), i.e.
// (x -/+ 2pi_f) +/- (2pi_f - 2pi).  Agrees with the literal chain to 1-2 ulp (tests run both).
__device__ __forceinline__ float wrap_pi(float x) {
    constexpr float kPi = 3.14159265358979f;  // rounds to pi_f, the smallest float above pi
    constexpr float kDelta = 1.7484555e-7f;   // 2pi_f - 2pi
    if (x >= kPi) return (x - kTwoPi) + kDelta;
    if (x < -kPi) return (x + kTwoPi) - kDelta;
    return x;
}

__device__ __forceinline__ float norm_angle(float x, bool closed_form) {
    return closed_form ? wrap_pi(x) : normalize_angle(x);
}

// atan2 result r in [-pi, pi]; torch's `r % (2*pi)` (python remainder) is r or r + 2pi_f
__device__ __forceinline__ float mod_2pi_of_atan2(float r) { return r < 0.0f ? r + kTwoPi : r; }

// Which of the LEN per-env columns thread r of env e handles in iteration k.  The lanes of a warp
// walk consecutive envs, i.e. shared-memory stride LEN words:
//   LEN odd        -> r + 4k: the stride is coprime with the 32 banks, conflict-free as is;
//   LEN % 8 == 0   -> the stride only reaches 4 banks, so ownership is skewed by the env index
//                     (column = 8m + ((2r + j + (e >> 2)) & 7)): 8 consecutive env groups hit the 8
//                     residues and a warp covers all 32 banks.  Each thread still owns exactly
//                     LEN/4 columns of its env; for the Ant (2 dofs per thread) the partial sums
//                     are order-independent.
template <int LEN>
__device__ __forceinline__ int owned_index(int r, int k, int e) {
    if constexpr (LEN % 8 == 0) {
        return 8 * (k >> 1) + ((2 * r + (k & 1) + (e >> 2)) & 7);
    } else {
        return r + 4 * k;
    }
}

// Block = 4*E threads.  After the flat phase every env of the tile is served by FOUR threads that
// sit in four different warps ("roles"), so the long transcendental chains of one env run
// concurrently and each warp executes one uniform instruction stream:
//   role 0  roll angle, up vector / up projection            (+ reward and termination at the end)
//   role 1  yaw angle, walk-target angle, angle to target
//   role 2  heading vector / heading projection, potentials
//   role 3  local linear / angular velocity, height
// and every thread also owns a quarter of the env's dofs for the element-wise part and the
// partial sums of the per-dof reductions.
template <int D, int S, int E>
__global__ void __launch_bounds__(4 * E) obs_reward_reset_kernel(const __grid_constant__ OigeTaskParams P,
                                                                 const __grid_constant__ K1Args A) {
    using T = Tile<D, S>;
    constexpr int OBS = T::OBS;
    constexpr int NT = 4 * E;
    constexpr int NW = E / 32;  // warps per role (r == 0 warps are warps 0 .. NW-1)
    extern __shared__ __align__(128) unsigned char smem_raw[];
    // ---- shared-memory carve-up (every region a multiple of 16 bytes because E % 4 == 0) ----
    long long* s_prog = reinterpret_cast<long long*>(smem_raw);
    long long* s_reset = s_prog + E;
    float* s_pos = reinterpret_cast<float*>(s_reset + E);
    float* s_quat = s_pos + 3 * E;
    float* s_vel = s_quat + 4 * E;
    float* s_q = s_vel + 6 * E;
    float* s_qd = s_q + D * E;
    float* s_sens = s_qd + D * E;
    float* s_act = s_sens + 6 * S * E;
    float* s_pot = s_act + D * E;  // staged potentials: stored back verbatim as prev_potentials
    float* o_obs = s_pot + E;
    float* o_rew = o_obs + OBS * E;
    float* o_pot = o_rew + E;
    float* o_up = o_pot + E;
    float* o_head = o_up + 3 * E;
    float* s_lo = o_head + 3 * E;  // per-dof constants (avoid divergent constant-bank indexing)
    float* s_hi = s_lo + 32;
    float* s_ratio = s_hi + 32;
    float* s_rcp = s_ratio + 32;   // RN(1 / (hi - lo)) per dof, for the correctly rounded fast division
    float* s_part = s_rcp + 32;    // [3 reductions][4 threads of an env][E] partial sums
    uint64_t* s_bar = reinterpret_cast<uint64_t*>(s_part + 12 * E);
    float* s_stat = reinterpret_cast<float*>(s_bar + 2);  // [NW][3] per-warp partial episode statistics

    const int tid = threadIdx.x;
    const long long e0 = (long long)blockIdx.x * E;
    const int cnt = (int)((A.n - e0) < (long long)E ? (A.n - e0) : (long long)E);
    const bool bulk = A.use_bulk && cnt == E;

    if (tid < 32) {
        s_lo[tid] = tid < D ? P.dof_limits_lower[tid] : 0.0f;
        s_hi[tid] = tid < D ? P.dof_limits_upper[tid] : 0.0f;
        s_ratio[tid] = tid < D ? P.motor_effort_ratio[tid] : 0.0f;
        s_rcp[tid] = tid < D ? 1.0f / (P.dof_limits_upper[tid] - P.dof_limits_lower[tid]) : 0.0f;
    }

    // ---- stage the tile -------------------------------------------------------------------
    pdl_wait();  // everything below reads what the articulation kernel wrote
    if (bulk) {
        if (tid == 0) {
            mbar_init(s_bar, 1);
            mbar_fence_init();
            constexpr uint32_t total = (uint32_t)E * (4u * T::IN_F32 + 16u);
            mbar_expect_tx(s_bar, total);
            bulk_g2s(s_q, A.q + e0 * D, E * D * 4, s_bar);
            bulk_g2s(s_qd, A.qd + e0 * D, E * D * 4, s_bar);
            bulk_g2s(s_act, A.act + e0 * D, E * D * 4, s_bar);
            bulk_g2s(s_sens, A.sens + e0 * (6 * S), E * 6 * S * 4, s_bar);
            bulk_g2s(s_quat, A.quat + e0 * 4, E * 16, s_bar);
            bulk_g2s(s_pos, A.pos + e0 * 3, E * 12, s_bar);
            bulk_g2s(s_vel, A.vel + e0 * 6, E * 24, s_bar);
            bulk_g2s(s_pot, A.pot + e0, E * 4, s_bar);
            bulk_g2s(s_prog, A.progress + e0, E * 8, s_bar);
            bulk_g2s(s_reset, A.reset + e0, E * 8, s_bar);
        }
        __syncthreads();  // barrier initialised (and per-dof constants written) before anyone polls
        mbar_wait(s_bar, 0);
    } else {
        coop_load(s_prog, A.progress + e0, cnt, tid, NT);
        coop_load(s_reset, A.reset + e0, cnt, tid, NT);
        coop_load(s_pos, A.pos + e0 * 3, cnt * 3, tid, NT);
        coop_load(s_quat, A.quat + e0 * 4, cnt * 4, tid, NT);
        coop_load(s_vel, A.vel + e0 * 6, cnt * 6, tid, NT);
        coop_load(s_q, A.q + e0 * D, cnt * D, tid, NT);
        coop_load(s_qd, A.qd + e0 * D, cnt * D, tid, NT);
        coop_load(s_sens, A.sens + e0 * (6 * S), cnt * 6 * S, tid, NT);
        coop_load(s_act, A.act + e0 * D, cnt * D, tid, NT);
        coop_load(s_pot, A.pot + e0, cnt, tid, NT);
        __syncthreads();
    }

    // ---- per-env work: thread (e, r) = (tid % E, tid / E); r is warp-uniform (E >= 32) ---------------
    // (1) flat part: thread r of env e owns dofs d = r, r+4, ... : obs[12+d] = unscale(dof_pos),
    //     obs[12+D+d] = dof_vel*scale, obs[12+2D+6S+d] = action, and the partial sums of the three
    //     per-dof reductions (limit cost / action cost / electricity) over ITS dofs.  Values are
    //     written UNCLAMPED: the reference computes its metrics on obs_buf before _process_data
    //     clamps a copy (envs/vec_env_rlgames.py:42).  Lanes walk consecutive envs: smem strides
    //     D and OBS are odd, so every access below is bank-conflict free.
    // (2) role part: the four threads of an env split the root-body terms (see kernel header).
    const int lane = tid & 31;
    const int r = tid / E;
    const int e = tid - r * E;
    const bool live = e < cnt;
    const bool closed = P.closed_form_angles != 0;
    if (live) {
        float* row = o_obs + e * OBS;
        float p_limit = 0.0f, p_act = 0.0f, p_elec = 0.0f;
        int p_count = 0;
#pragma unroll
        for (int k = 0; k < (D + 3) / 4; ++k) {
            const int d = owned_index<D>(r, k, e);
            if (d < D) {
                const float lo = s_lo[d], hi = s_hi[d];
                const float num = (2.0f * s_q[e * D + d] - hi) - lo;
                const float qs = div_rn(num, hi - lo, s_rcp[d]);  // maths.unscale
                const float dv = s_qd[e * D + d] * P.dof_vel_scale;
                const float a = s_act[e * D + d];
                row[12 + d] = qs;
                row[12 + D + d] = dv;
                row[12 + 2 * D + 6 * S + d] = a;
                const float ratio = s_ratio[d];
                if (P.limit_cost_kind == OIGE_LIMIT_COST_ANT) {
                    p_count += (qs > 0.99f) ? 1 : 0;  // ant.py:95
                } else {                               // humanoid.py:123-126 (masked-out terms are exact zeros)
                    const float aq = fabsf(qs);
                    const float sc = div_rn(P.joints_at_limit_cost_scale * (aq - 0.98f), 0.02f, 50.0f) * ratio;
                    p_limit = p_limit + (aq > 0.98f ? sc : 0.0f);
                }
                p_act = p_act + a * a;
                p_elec = p_elec + fabsf(a * dv) * ratio;
            }
        }
        if (P.limit_cost_kind == OIGE_LIMIT_COST_ANT) p_limit = (float)p_count;
        s_part[(0 * 4 + r) * E + e] = p_limit;
        s_part[(1 * 4 + r) * E + e] = p_act;
        s_part[(2 * 4 + r) * E + e] = p_elec;
#pragma unroll
        for (int k = 0; k < (6 * S + 3) / 4; ++k) {
            const int j = owned_index<6 * S>(r, k, e);
            if (j < 6 * S) row[12 + 2 * D + j] = s_sens[e * 6 * S + j] * P.contact_force_scale;
        }

        const float4 q4 = *reinterpret_cast<const float4*>(s_quat + 4 * e);
        const float rot[4] = {q4.x, q4.y, q4.z, q4.w};
        float tq[4];
        quat_mul(rot, P.inv_start_rot, tq);  // compute_heading_and_up: torso_quat
        const float w = tq[0], x = tq[1], y = tq[2], z = tq[3];
        const float px = s_pos[3 * e], py = s_pos[3 * e + 1], pz = s_pos[3 * e + 2];
        if (r == 0) {
            const float sinr_cosp = 2.0f * (w * x + y * z);
            const float cosr_cosp = ((w * w - x * x) - y * y) + z * z;
            const float roll = mod_2pi_of_atan2(atan2f(sinr_cosp, cosr_cosp));
            row[8] = norm_angle(roll, closed);
            float upv[3];
            quat_rotate<false>(tq, P.basis_vec1, upv);
            row[10] = upv[2];  // up_proj
            o_up[3 * e] = upv[0];
            o_up[3 * e + 1] = upv[1];
            o_up[3 * e + 2] = upv[2];
        } else if (r == 1) {
            const float siny_cosp = 2.0f * (w * z + x * y);
            const float cosy_cosp = ((w * w + x * x) - y * y) - z * z;
            const float yaw = mod_2pi_of_atan2(atan2f(siny_cosp, cosy_cosp));
            const float walk = P.walk_target_uses_z ? atan2f(P.targets[2] - pz, P.targets[0] - px)
                                                    : atan2f(P.targets[1] - py, P.targets[0] - px);
            row[7] = norm_angle(yaw, closed);
            row[9] = norm_angle(walk - yaw, closed);
        } else if (r == 2) {
            // heading: normalize(to_target) . quat_rotate(torso_quat, basis_vec0)
            const float tx = P.targets[0] - px, ty = P.targets[1] - py, tz = 0.0f;
            const float nrm = sqrtf(fmaf(tz, tz, fmaf(ty, ty, tx * tx)));  // torch's vector norm is an fma chain
            const float nc = fmaxf(nrm, 1e-9f);
            float hdv[3];
            quat_rotate<false>(tq, P.basis_vec0, hdv);
            row[11] = (hdv[0] * (tx / nc) + hdv[1] * (ty / nc)) + hdv[2] * (tz / nc);
            o_head[3 * e] = hdv[0];
            o_head[3 * e + 1] = hdv[1];
            o_head[3 * e + 2] = hdv[2];
            o_pot[e] = (-nrm) / P.dt;  // potentials (locomotion.py:219-223)
        } else {
            const float lin[3] = {s_vel[6 * e], s_vel[6 * e + 1], s_vel[6 * e + 2]};
            const float ang[3] = {s_vel[6 * e + 3], s_vel[6 * e + 4], s_vel[6 * e + 5]};
            float vloc[3], wloc[3];
            quat_rotate<true>(tq, lin, vloc);
            quat_rotate<true>(tq, ang, wloc);
            row[0] = pz;
            row[1] = vloc[0];
            row[2] = vloc[1];
            row[3] = vloc[2];
            row[4] = wloc[0] * P.angular_velocity_scale;
            row[5] = wloc[1] * P.angular_velocity_scale;
            row[6] = wloc[2] * P.angular_velocity_scale;
        }
    }
    __syncthreads();

    // ---- reward + termination: thread r == 0 of each env ------------------------------------------
    float st_rew = 0.0f, st_done = 0.0f, st_len = 0.0f;
    if (live && r == 0) {
        const float* row = o_obs + e * OBS;
        const float pz = row[0], up_proj = row[10], heading_proj = row[11];
        const float pot = o_pot[e], prev_pot = s_pot[e];
        // the four partial sums of each reduction are combined in a fixed order
        float red[3];
#pragma unroll
        for (int c = 0; c < 3; ++c)
            red[c] = ((s_part[(c * 4 + 0) * E + e] + s_part[(c * 4 + 1) * E + e]) + s_part[(c * 4 + 2) * E + e]) +
                     s_part[(c * 4 + 3) * E + e];
        const float limit_cost = red[0], actions_cost = red[1], electricity = red[2];
        // calculate_metrics (locomotion.py:288-321)
        const float heading_reward = heading_proj > 0.8f ? P.heading_weight : (P.heading_weight * heading_proj) / 0.8f;
        const float up_reward = up_proj > 0.93f ? (0.0f + P.up_weight) : 0.0f;
        const float progress_reward = pot - prev_pot;
        float total = progress_reward + P.alive_reward_scale;
        total = total + up_reward;
        total = total + heading_reward;
        total = total - P.actions_cost_scale * actions_cost;
        total = total - P.energy_cost_scale * electricity;
        total = total - limit_cost;
        const bool fallen = pz < P.termination_height;
        if (fallen) total = P.death_cost;
        // rl_task.py:242 then is_done (locomotion.py:266-267; compare in fp32 as torch promotes)
        const long long prog = s_prog[e] + (P.progress_increment ? 1 : 0);
        long long rst = s_reset[e];
        if (fallen) rst = 1;
        if ((float)prog >= P.max_episode_length - 1.0f) rst = 1;
        s_prog[e] = prog;
        s_reset[e] = rst;
        o_rew[e] = total;
        st_rew = total;
        st_done = rst != 0 ? 1.0f : 0.0f;
        st_len = rst != 0 ? (float)prog : 0.0f;
    }
    if (A.stats != nullptr && r == 0) {  // warp-uniform condition
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            st_rew += __shfl_xor_sync(0xffffffffu, st_rew, o);
            st_done += __shfl_xor_sync(0xffffffffu, st_done, o);
            st_len += __shfl_xor_sync(0xffffffffu, st_len, o);
        }
        if (lane == 0) {
            s_stat[(tid >> 5) * 3 + 0] = st_rew;
            s_stat[(tid >> 5) * 3 + 1] = st_done;
            s_stat[(tid >> 5) * 3 + 2] = st_len;
        }
    }
    if (P.clip_obs < INFINITY) {  // uniform; finite clip_obs only: clamp the rows in a second sweep
        __syncthreads();
        const float c = P.clip_obs;
        for (int i = tid; i < cnt * OBS; i += NT) o_obs[i] = clampf(o_obs[i], c);
    }

    // ---- drain the tile ----------------------------------------------------------------------
    if (bulk) {
        fence_async_smem();
        __syncthreads();
        if (tid == 0) {
            bulk_s2g(A.obs + e0 * OBS, o_obs, E * OBS * 4);
            bulk_s2g(A.rew + e0, o_rew, E * 4);
            bulk_s2g(A.pot + e0, o_pot, E * 4);
            bulk_s2g(A.prev_pot + e0, s_pot, E * 4);
            bulk_s2g(A.up + e0 * 3, o_up, E * 12);
            bulk_s2g(A.head + e0 * 3, o_head, E * 12);
            bulk_s2g(A.reset + e0, s_reset, E * 8);
            bulk_s2g(A.progress + e0, s_prog, E * 8);
            bulk_commit();
        }
    } else {
        __syncthreads();
        coop_store(A.obs + e0 * OBS, o_obs, cnt * OBS, tid, NT);
        coop_store(A.rew + e0, o_rew, cnt, tid, NT);
        coop_store(A.pot + e0, o_pot, cnt, tid, NT);
        coop_store(A.prev_pot + e0, s_pot, cnt, tid, NT);
        coop_store(A.up + e0 * 3, o_up, cnt * 3, tid, NT);
        coop_store(A.head + e0 * 3, o_head, cnt * 3, tid, NT);
        coop_store(A.reset + e0, s_reset, cnt, tid, NT);
        coop_store(A.progress + e0, s_prog, cnt, tid, NT);
    }
    if (tid == 0) {
        if (A.stats != nullptr) {
            double r = 0.0, d = 0.0, l = 0.0;
            for (int w = 0; w < NW; ++w) {
                r += (double)s_stat[w * 3 + 0];
                d += (double)s_stat[w * 3 + 1];
                l += (double)s_stat[w * 3 + 2];
            }
            double* slot = A.stats + (blockIdx.x % OIGE_STATS_SLOTS) * 4;
            atomicAdd(slot + 0, r);
            atomicAdd(slot + 1, d);
            atomicAdd(slot + 2, l);
            atomicAdd(slot + 3, (double)cnt);
        }
        if (A.step_counter != nullptr && blockIdx.x == 0) *A.step_counter += 1;  // K3 of this step already ran
        if (bulk) bulk_wait_read0();  // smem must stay alive until the TMA engine has read it
    }
}

template <int D, int S, int E>
constexpr size_t k1_smem_bytes() {
    return (size_t)E * Tile<D, S>::BYTES_PER_ENV + 4 * 32 * 4 + 12 * E * 4 + 16 + (E / 32) * 3 * 4 + 16;
}

template <int D, int S, int E>
int launch_k1(const OigeTaskParams& P, const K1Args& A, cudaStream_t stream) {
    static bool configured = false;
    constexpr size_t smem = k1_smem_bytes<D, S, E>();
    auto kern = obs_reward_reset_kernel<D, S, E>;
    if (!configured) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return fail((int)e, "cudaFuncSetAttribute(K1): %s", cudaGetErrorString(e));
        configured = true;
    }
    const long long tiles = (A.n + E - 1) / E;
    launch_pdl(kern, (unsigned)tiles, 4 * E, smem, stream, P, A);
    count_launch();
    return check_launch("obs_reward_reset_kernel");
}

inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

}  // namespace

// Tile size: 32 envs (128 threads, ~26 KB smem -> 8 CTAs/SM); 64-env tiles (256 threads) stay available
// through oige_obs_reward_reset_tile_f32 for sweeps.
int k1_dispatch(const OigeTaskParams* P, const OigeState* st, const OigeTaskBuffers* b, int64_t n,
                cudaStream_t stream, int force_tile, bool bump_counter) {
    K1Args A;
    A.pos = st->root_pos; A.quat = st->root_quat; A.vel = st->root_vel;
    A.q = st->dof_pos; A.qd = st->dof_vel; A.sens = st->sensors; A.act = b->actions;
    A.pot = b->potentials; A.prev_pot = b->prev_potentials; A.obs = b->obs; A.rew = b->rew;
    A.up = b->up_vec; A.head = b->heading_vec;
    A.reset = reinterpret_cast<long long*>(b->reset_buf);
    A.progress = reinterpret_cast<long long*>(b->progress_buf);
    A.stats = b->episode_stats;
    A.step_counter = bump_counter ? reinterpret_cast<unsigned long long*>(b->reset_counter) : nullptr;
    A.n = n;
    const void* ptrs[] = {A.pos, A.quat, A.vel, A.q, A.qd, A.sens, A.act, A.pot, A.prev_pot,
                          A.obs, A.rew, A.up, A.head, A.reset, A.progress};
    A.use_bulk = 1;
    for (const void* p : ptrs) {
        if (p == nullptr) return fail(OIGE_E_NULL, "oige_obs_reward_reset_f32: null device pointer");
        if (!aligned16(p)) A.use_bulk = 0;  // plain coalesced path for oddly offset views
    }
    int tile = force_tile;
    if (tile == 0) tile = 32;  // measured on B200: 32-env tiles beat 64 at every env count (profiles/)
    const int D = P->num_dof, S = P->num_sensors;
    if (P->num_obs != 12 + 3 * D + 6 * S)
        return fail(OIGE_E_SHAPE, "num_obs %d != 12 + 3*%d + 6*%d", P->num_obs, D, S);
#define OIGE_K1_CASE(DD, SS)                                                   \
    if (D == DD && S == SS) {                                                  \
        if (tile == 32) return launch_k1<DD, SS, 32>(*P, A, stream);           \
        if (tile == 64) return launch_k1<DD, SS, 64>(*P, A, stream);           \
        return fail(OIGE_E_RANGE, "unsupported K1 tile %d", tile);             \
    }
    OIGE_K1_CASE(21, 2)  // Humanoid
    OIGE_K1_CASE(8, 4)   // Ant
#undef OIGE_K1_CASE
    return fail(OIGE_E_SHAPE, "unsupported (num_dof, num_sensors) = (%d, %d); built: (21,2) (8,4)", D, S);
}

}  // namespace oige
