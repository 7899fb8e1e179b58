// Arithmetic of the locomotion observation / reward / termination pass, written ONCE for the two kernels that
// evaluate it: the standalone streaming pass (obs_reward_reset.cu, K1) and the epilogue of the whole-step kernel
// (env_step_role.cuh, K2-F).  Reference: tasks/shared/locomotion.py:190-321, tasks/humanoid.py:120-127,
// tasks/ant.py:92-95 and the omni.isaac.core helpers restated in SURVEY.md Appendix A.
//
// Every operation is spelled with a round-to-nearest intrinsic (__fmul_rn / __fadd_rn / __fsub_rn; never fused by
// nvcc or ptxas) in the reference's operation order, with explicit fmaf exactly where torch's CPU kernels use an FMA
// (vector norm, cross product).  The results therefore do not depend on the translation unit's -fmad setting: the
// articulation TU keeps FMA contraction for the dynamics and still produces the bits K1 produces.  (sinf / cosf /
// atan2f / sqrtf / IEEE division compile to identical SASS under both settings; checked with cuobjdump.)
#pragma once
#include <math.h>

#include "oige_b200.h"

namespace oige {
namespace lm {

__device__ __forceinline__ float mul(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ float add(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ float sub(float a, float b) { return __fsub_rn(a, b); }

constexpr float kTwoPi = 6.283185307179586f;  // (float)(2*pi), as torch casts the python scalar

__device__ __forceinline__ float normalize_angle(float x) {  // locomotion.py:190-192
    return atan2f(sinf(x), cosf(x));
}

// torch.cross CPU kernel contracts a1*b2 - a2*b1 into fma(a1, b2, -(a2*b1))
__device__ __forceinline__ void cross3(const float a[3], const float b[3], float c[3]) {
    c[0] = fmaf(a[1], b[2], -mul(a[2], b[1]));
    c[1] = fmaf(a[2], b[0], -mul(a[0], b[2]));
    c[2] = fmaf(a[0], b[1], -mul(a[1], b[0]));
}

// v(2w^2-1) +- 2w (q x v) + 2 q (q.v)      (quat_rotate / quat_rotate_inverse)
template <bool INVERSE>
__device__ __forceinline__ void quat_rotate(const float q[4], const float v[3], float out[3]) {
    const float w = q[0];
    const float qv[3] = {q[1], q[2], q[3]};
    const float s = sub(mul(2.0f, mul(w, w)), 1.0f);
    float cr[3];
    cross3(qv, v, cr);
    const float dot = add(add(mul(qv[0], v[0]), mul(qv[1], v[1])), mul(qv[2], v[2]));
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        const float a = mul(v[i], s);
        const float b = mul(mul(cr[i], w), 2.0f);
        const float c = mul(mul(qv[i], dot), 2.0f);
        out[i] = INVERSE ? add(sub(a, b), c) : add(add(a, b), c);
    }
}

// factored 9-multiply Hamilton product (quat_mul)
__device__ __forceinline__ void quat_mul(const float a[4], const float b[4], float o[4]) {
    const float w1 = a[0], x1 = a[1], y1 = a[2], z1 = a[3];
    const float w2 = b[0], x2 = b[1], y2 = b[2], z2 = b[3];
    const float ww = mul(add(z1, x1), add(x2, y2));
    const float yy = mul(sub(w1, y1), add(w2, z2));
    const float zz = mul(add(w1, y1), sub(w2, z2));
    const float xx = add(add(ww, yy), zz);
    const float qq = mul(0.5f, add(xx, mul(sub(z1, x1), sub(x2, y2))));
    o[0] = add(sub(qq, ww), mul(sub(z1, y1), sub(y2, z2)));
    o[1] = add(sub(qq, xx), mul(add(x1, w1), add(x2, w2)));
    o[2] = add(sub(qq, yy), mul(sub(w1, x1), add(y2, z2)));
    o[3] = add(sub(qq, zz), mul(add(z1, y1), sub(w2, x2)));
}

__device__ __forceinline__ float clampf(float v, float c) {  // torch.clamp semantics (NaN propagates)
    return v < -c ? -c : (v > c ? c : v);
}

// Correctly rounded a / b from y = RN(1/b): q0 = RN(a*y), r = a - q0*b (exact, one FMA), q1 = RN(q0 + r*y)
// (Markstein 1990: q1 == RN(a/b) whenever y is the correctly rounded reciprocal and nothing over/underflows).
// Three instructions instead of the ~15 of the IEEE division subroutine.  Inputs outside the normal
// range (|a/b| near the fp32 limits, non-finite values) take the IEEE path so NaN/inf propagate as in torch.
__device__ __forceinline__ float div_rn(float a, float b, float y) {
    const float q0 = mul(a, y);
    const float aq = fabsf(q0);
    if (!(aq > 1e-30f && aq < 1e30f)) return __fdiv_rn(a, b);
    const float r = fmaf(-q0, b, a);
    return fmaf(r, y, q0);
}

// wrap an angle x (|x| < 4 pi) into (-pi, pi] the way atan2(sin x, cos x) does, without the three
// transcendentals: sin/cos see the FLOAT x, so the exact result is x -/+ 2*pi (true pi), i.e.
// (x -/+ 2pi_f) +/- (2pi_f - 2pi).  Agrees with the literal chain to 1-2 ulp (tests run both).
__device__ __forceinline__ float wrap_pi(float x) {
    constexpr float kPi = 3.14159265358979f;  // rounds to pi_f, the smallest float above pi
    constexpr float kDelta = 1.7484555e-7f;   // 2pi_f - 2pi
    if (x >= kPi) return add(sub(x, kTwoPi), kDelta);
    if (x < -kPi) return sub(add(x, kTwoPi), kDelta);
    return x;
}

__device__ __forceinline__ float norm_angle(float x, bool closed_form) {
    return closed_form ? wrap_pi(x) : normalize_angle(x);
}

// atan2 result r in [-pi, pi]; torch's `r % (2*pi)` (python remainder) is r or r + 2pi_f
__device__ __forceinline__ float mod_2pi_of_atan2(float r) { return r < 0.0f ? add(r, kTwoPi) : r; }

// Which of the LEN per-env columns "quarter" r (0..3) of env e handles in iteration k.  In K1 the lanes of a warp
// walk consecutive envs, i.e. shared-memory stride LEN words:
//   LEN odd        -> r + 4k: the stride is coprime with the 32 banks, conflict-free as is;
//   LEN % 8 == 0   -> the stride only reaches 4 banks, so ownership is skewed by the env index
//                     (column = 8m + ((2r + j + (e >> 2)) & 7)): 8 consecutive env groups hit the 8
//                     residues and a warp covers all 32 banks.  Each quarter still owns exactly
//                     LEN/4 columns of its env; for the Ant (2 dofs per quarter) the partial sums
//                     are order-independent.
// The whole-step kernel uses the same ownership so that its partial sums are bit-identical to K1's.
template <int LEN>
__device__ __forceinline__ int owned_index(int r, int k, int e) {
    if constexpr (LEN % 8 == 0) {
        return 8 * (k >> 1) + ((2 * r + (k & 1) + (e >> 2)) & 7);
    } else {
        return r + 4 * k;
    }
}
__device__ __forceinline__ int owned_index_rt(int len, int r, int k, int e) {
    return (len % 8 == 0) ? 8 * (k >> 1) + ((2 * r + (k & 1) + (e >> 2)) & 7) : r + 4 * k;
}
// Same for columns that are only copied / scaled (the sensor block: no reduction, any ownership gives the same bits).
// LEN % 4 == 0 but not % 8 (the Humanoid's 12 sensor floats): stride 12 reaches 8 banks, lanes e and e + 8 collide
// (4-way); rotating the column by e >> 3 moves the four lane groups onto the four residues mod 4.
template <int LEN>
__device__ __forceinline__ int owned_index_elem(int r, int k, int e) {
    if constexpr (LEN % 8 != 0 && LEN % 4 == 0) {
        const int j = r + 4 * k + ((e >> 3) & 3);
        return j >= LEN ? j - LEN : j;
    } else {
        return owned_index<LEN>(r, k, e);
    }
}
__device__ __forceinline__ int owned_index_elem_rt(int len, int r, int k, int e) {
    if (len % 8 != 0 && len % 4 == 0) {
        const int j = r + 4 * k + ((e >> 3) & 3);
        return j >= len ? j - len : j;
    }
    return owned_index_rt(len, r, k, e);
}

// maths.unscale(dof_pos, lo, hi) = (2x - hi - lo) / (hi - lo)  (locomotion.py:233); rcp = RN(1 / (hi - lo))
__device__ __forceinline__ float unscale(float q, float lo, float hi, float rcp) {
    const float num = sub(sub(mul(2.0f, q), hi), lo);
    return div_rn(num, sub(hi, lo), rcp);
}

// Per-dof contributions to the three reductions of one quarter (limit cost / action cost / electricity).
struct Partial {
    float limit, act, elec;
    int count;
};
__device__ __forceinline__ void partial_init(Partial& p) { p.limit = 0.0f; p.act = 0.0f; p.elec = 0.0f; p.count = 0; }
__device__ __forceinline__ void partial_add(const OigeTaskParams& P, Partial& p, float qs, float dv, float a,
                                            float ratio) {
    if (P.limit_cost_kind == OIGE_LIMIT_COST_ANT) {
        p.count += (qs > 0.99f) ? 1 : 0;  // ant.py:95
    } else {                               // humanoid.py:123-126 (masked-out terms are exact zeros)
        const float aq = fabsf(qs);
        const float sc = mul(div_rn(mul(P.joints_at_limit_cost_scale, sub(aq, 0.98f)), 0.02f, 50.0f), ratio);
        p.limit = add(p.limit, (aq > 0.98f ? sc : 0.0f));
    }
    p.act = add(p.act, mul(a, a));
    p.elec = add(p.elec, mul(fabsf(mul(a, dv)), ratio));
}
__device__ __forceinline__ void partial_finish(const OigeTaskParams& P, Partial& p) {
    if (P.limit_cost_kind == OIGE_LIMIT_COST_ANT) p.limit = (float)p.count;
}

// The root-body terms of the observation row, split into four independent parts (K1 runs them on four threads of an
// env, the whole-step kernel on its role warps).  tq = quat_mul(torso_rotation, inv_start_rot).
//   part 0  roll angle, up vector / up projection
//   part 1  yaw angle, walk-target angle, angle to target
//   part 2  heading vector / heading projection, potentials
//   part 3  height, local linear / angular velocity
__device__ __forceinline__ void torso_quat(const OigeTaskParams& P, const float rot[4], float tq[4]) {
    quat_mul(rot, P.inv_start_rot, tq);  // compute_heading_and_up: torso_quat
}
__device__ __forceinline__ void root_part0(const OigeTaskParams& P, const float tq[4], bool closed, float* row,
                                           float* up3) {
    const float w = tq[0], x = tq[1], y = tq[2], z = tq[3];
    const float sinr_cosp = mul(2.0f, add(mul(w, x), mul(y, z)));
    const float cosr_cosp = add(sub(sub(mul(w, w), mul(x, x)), mul(y, y)), mul(z, z));
    const float roll = mod_2pi_of_atan2(atan2f(sinr_cosp, cosr_cosp));
    row[8] = norm_angle(roll, closed);
    float upv[3];
    quat_rotate<false>(tq, P.basis_vec1, upv);
    row[10] = upv[2];  // up_proj
    up3[0] = upv[0];
    up3[1] = upv[1];
    up3[2] = upv[2];
}
__device__ __forceinline__ void root_part1(const OigeTaskParams& P, const float tq[4], float px, float py, float pz,
                                           bool closed, float* row) {
    const float w = tq[0], x = tq[1], y = tq[2], z = tq[3];
    const float siny_cosp = mul(2.0f, add(mul(w, z), mul(x, y)));
    const float cosy_cosp = sub(sub(add(mul(w, w), mul(x, x)), mul(y, y)), mul(z, z));
    const float yaw = mod_2pi_of_atan2(atan2f(siny_cosp, cosy_cosp));
    const float walk = P.walk_target_uses_z ? atan2f(sub(P.targets[2], pz), sub(P.targets[0], px))
                                            : atan2f(sub(P.targets[1], py), sub(P.targets[0], px));
    row[7] = norm_angle(yaw, closed);
    row[9] = norm_angle(sub(walk, yaw), closed);
}
__device__ __forceinline__ void root_part2(const OigeTaskParams& P, const float tq[4], float px, float py, float* row,
                                           float* head3, float* pot) {
    // heading: normalize(to_target) . quat_rotate(torso_quat, basis_vec0)
    const float tx = sub(P.targets[0], px), ty = sub(P.targets[1], py), tz = 0.0f;
    const float nrm = sqrtf(fmaf(tz, tz, fmaf(ty, ty, mul(tx, tx))));  // torch's vector norm is an fma chain
    const float nc = fmaxf(nrm, 1e-9f);
    float hdv[3];
    quat_rotate<false>(tq, P.basis_vec0, hdv);
    row[11] = add(add(mul(hdv[0], __fdiv_rn(tx, nc)), mul(hdv[1], __fdiv_rn(ty, nc))), mul(hdv[2], __fdiv_rn(tz, nc)));
    head3[0] = hdv[0];
    head3[1] = hdv[1];
    head3[2] = hdv[2];
    *pot = __fdiv_rn(-nrm, P.dt);  // potentials (locomotion.py:219-223)
}
__device__ __forceinline__ void root_part3(const OigeTaskParams& P, const float tq[4], float pz, const float lin[3],
                                           const float ang[3], float* row) {
    float vloc[3], wloc[3];
    quat_rotate<true>(tq, lin, vloc);
    quat_rotate<true>(tq, ang, wloc);
    row[0] = pz;
    row[1] = vloc[0];
    row[2] = vloc[1];
    row[3] = vloc[2];
    row[4] = mul(wloc[0], P.angular_velocity_scale);
    row[5] = mul(wloc[1], P.angular_velocity_scale);
    row[6] = mul(wloc[2], P.angular_velocity_scale);
}

// calculate_metrics (locomotion.py:288-321) + rl_task.py:242 + is_done (locomotion.py:266-267; compared in fp32 as
// torch promotes).  red = {limit cost, actions cost, electricity}, the quarters combined in a fixed order.
__device__ __forceinline__ float combine4(float a, float b, float c, float d) { return add(add(add(a, b), c), d); }
__device__ __forceinline__ void reward_done(const OigeTaskParams& P, float pz, float up_proj, float heading_proj,
                                            float pot, float prev_pot, const float red[3], long long prog_in,
                                            long long reset_in, float& total_out, long long& prog_out,
                                            long long& reset_out) {
    const float limit_cost = red[0], actions_cost = red[1], electricity = red[2];
    const float heading_reward =
        heading_proj > 0.8f ? P.heading_weight : __fdiv_rn(mul(P.heading_weight, heading_proj), 0.8f);
    const float up_reward = up_proj > 0.93f ? add(0.0f, P.up_weight) : 0.0f;
    const float progress_reward = sub(pot, prev_pot);
    float total = add(progress_reward, P.alive_reward_scale);
    total = add(total, up_reward);
    total = add(total, heading_reward);
    total = sub(total, mul(P.actions_cost_scale, actions_cost));
    total = sub(total, mul(P.energy_cost_scale, electricity));
    total = sub(total, limit_cost);
    const bool fallen = pz < P.termination_height;
    if (fallen) total = P.death_cost;
    const long long prog = prog_in + (P.progress_increment ? 1 : 0);
    long long rst = reset_in;
    if (fallen) rst = 1;
    if ((float)prog >= sub(P.max_episode_length, 1.0f)) rst = 1;
    total_out = total;
    prog_out = prog;
    reset_out = rst;
}

}  // namespace lm
}  // namespace oige
