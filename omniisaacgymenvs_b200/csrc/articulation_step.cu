// K2 -- batched reduced-coordinate articulation step (sm_100a).
//
// Stands in for `set_joint_efforts` + `control_frequency_inv x World.step` of the reference
// (tasks/shared/locomotion.py:108-114, envs/vec_env_rlgames.py:64-66), i.e. for the PhysX 5 GPU
// articulation pipeline that cannot be installed here.  See DESIGN.md "K2" for the model.
//
// Three kernels integrate the same equations with the same device functions:
//   articulation_step_kernel      one WARP per env, lane = link (lane 0 = floating root, one link per
//                                 DoF); parent<->child exchange by warp shuffles; forward kinematics by
//                                 pointer jumping; small robots packed several envs per warp.  Lowest
//                                 latency for few environments.
//   articulation_step_tpe_kernel  one THREAD per env; per-link intermediates in per-thread local memory,
//                                 articulated inertias carried up heavy paths in registers.  ~7x fewer
//                                 issue slots per env: used when the launch spans several waves.
//   env_step_role_kernel          K2-F: lane = env, warp = ROLE (trunk + limb chains of the tree on separate warps),
//                                 every per-link intermediate in shared memory, and the WHOLE env step in the launch:
//                                 action clamp + masked reset before, observations / reward / termination after the
//                                 substeps (env_step_role.cuh).  The default from a few thousand envs on.
//
//   spatial vectors   = (angular xyz, linear xyz) in WORLD-aligned axes about the point O that
//                       coincides with the root origin at the start of the substep, so that
//                       parent->child propagation needs no frame transforms.
//   forward dynamics  = articulated-body algorithm: top-down kinematics, bottom-up articulated
//                       inertia (symmetric 6x6, 21 floats), 6x6 Cholesky at the root, top-down
//                       accelerations.
//   time stepping     = linearly-implicit Euler: joint springs/dampers/limit penalties and the
//                       ground-contact spring-damper-friction forces are evaluated at the
//                       end-of-step velocity, which adds a PSD term to the joint-space inertia
//                       (diagonal for joints; P^T C P on the link's 6x6 inertia for contacts)
//                       and keeps the 120 Hz step unconditionally stable for stiff contacts.
//   tree constants    = SoA table staged in shared memory once per CTA.
//
// Tensor cores are not used: the per-env matrices are 6x6 and the work is a latency-bound tree
// recursion, not a dense contraction.
#include <math.h>
#include <stdio.h>
#include <stdlib.h>

#include <atomic>
#include <mutex>

#include "locomotion_math.cuh"
#include "oige_common.cuh"

struct OigeModel {
    void* dev_table;      // DevModel in device memory
    void* dev_role;       // RoleModel in device memory (K2-F) or null
    int num_links;
    int num_dof;
    int num_sensors;
    int tpe_ok;
    int role_ok;          // K2-F applicable: trunk is a path, (roles, chain links per role) is an instantiated combination
    int nroles, maxc;     // roles per 32-env group; link slots per role the kernel is instantiated for (3 or 6)
    int group_floats;     // shared-memory floats (x32 lanes) of one group's scratch
    int has_sc;           // the model has self-collision pairs: only K2-F integrates it
};

namespace oige {
namespace {

// envs parked by the non-finite guard of ANY articulation kernel (K2-W, K2-T, K2-F) since load, on this device
// (oige_nonfinite_env_count; diagnostic)
__device__ unsigned long long g_nonfinite_envs = 0;

#ifndef TPE_MINBLOCKS
#define TPE_MINBLOCKS 4
#endif
constexpr int kWarpsPerBlock = 4;
constexpr int kMaxLevels = 16;
constexpr int kMaxChildren = 4;
constexpr int kMaxJump = 4;  // 2^4 = 16 >= kMaxLevels
constexpr int kMaxSlots = 8;
constexpr long long kRoleMinEnvs = 4096, kRoleMinEnvsSmall = 1024;  // measured, profiles/r1_k2_modes.txt
constexpr long long kTpeMinEnvs = 8192;  // measured crossover, see profiles/
constexpr int kHasHeavy = 1, kIsHead = 2, kFirstLight = 4;
constexpr int kMaxRoles = 4;

struct DevModel {
    int num_links, num_spheres, num_sensors, max_depth;
    int max_sph_per_link, jump_rounds, pack, tpe_ok;
    int fixed_base, pad_i[3];
    int parent[32];
    int dof[32];
    int jtype[32];                   // OIGE_JOINT_HINGE / OIGE_JOINT_PRISMATIC
    int level[32];
    int child[kMaxChildren][32];
    int maxch_at_level[kMaxLevels];  // max #children over parents at this level
    int anc[kMaxJump][32];           // 2^k-th ancestor (clamped at the root) for pointer jumping
    // thread-per-env kernel: post-order (heavy child last) and hand-over slots of the light children
    int order2[32];
    int tflags[32];                  // kHasHeavy | kIsHead | kFirstLight
    int light_slot[32];              // slot of a link that has light children, else -1
    int sph_start[32];
    int sph_count[32];
    int sensor_of_link[32];
    float jpos[3][32];
    float jquat[4][32];
    float axis[3][32];
    float mass[32];
    float com[3][32];
    float inertia[6][32];  // xx yy zz xy xz yz
    float armature[32], damping[32], stiffness[32], sref[32], lo[32], hi[32];
    float drive_kp[32], drive_kd[32], drive_fmax[32];
    float sph_pos[3][OIGE_MAX_SPHERES];
    float sph_r[OIGE_MAX_SPHERES];
    float gravity[4];
    float k_c, d_c, mu, c_visc, k_lim, d_lim, max_qd, ang_damp;
    float max_qd_lin, max_w, pen_cap_v, pad_f;  // prismatic clamp, root |w| clamp, depenetration cap (m/s); 1e30 = off
};
static_assert(sizeof(DevModel) % 16 == 0, "DevModel must be int4-copyable");

// ---- small vector helpers ---------------------------------------------------------------
struct V3 { float x, y, z; };
__device__ __forceinline__ V3 operator+(V3 a, V3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
__device__ __forceinline__ V3 operator-(V3 a, V3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
__device__ __forceinline__ V3 operator*(float s, V3 a) { return {s * a.x, s * a.y, s * a.z}; }
__device__ __forceinline__ V3 cross(V3 a, V3 b) {
    return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x};
}
__device__ __forceinline__ float dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }

struct Q4 { float w, x, y, z; };
// sin / cos of a joint HALF angle.  Joint limits keep |x| below ~1.4 rad (the knee: 160 deg / 2), where two short Taylor
// chains (x^13, x^12) replace sincosf's range reduction -- it heads the dependency chain of every link's kinematics.
// Accuracy is ABSOLUTE (what the components of a unit quaternion need), not relative to the result: up to the 1.45 rad
// cut-over |error| <= 1.03e-7 (sine) / 8.7e-8 (cosine), i.e. < 2 ulp of 1.0 against 0.5 ulp for a correctly rounded
// value, nearly all of it Horner rounding (the first dropped terms are x^15/15! = 2e-10 and x^14/14! = 2.1e-9 there).
// Near pi/2 the cosine is small and the same absolute error is many ulps OF THE RESULT; past 1.45 the dropped cosine
// term grows to 2.9e-8 at 1.75 and the total passes 1.3e-7, so anything beyond 1.45 (unlimited test joints) takes
// sincosf.  tests/test_oracle_math.py sweeps the restated chains against float64 sin / cos.
__device__ __forceinline__ void sincos_half_angle(float x, float* s, float* c) {
    if (fabsf(x) > 1.45f) {
        sincosf(x, s, c);
        return;
    }
    const float x2 = x * x;
    float ps = fmaf(x2, 1.6059044e-10f, -2.5052108e-8f);
    ps = fmaf(ps, x2, 2.7557319e-6f);
    ps = fmaf(ps, x2, -1.9841270e-4f);
    ps = fmaf(ps, x2, 8.3333333e-3f);
    ps = fmaf(ps, x2, -1.6666667e-1f);
    *s = fmaf(ps * x2, x, x);
    float pc = fmaf(x2, 2.0876757e-9f, -2.7557319e-7f);
    pc = fmaf(pc, x2, 2.4801587e-5f);
    pc = fmaf(pc, x2, -1.3888889e-3f);
    pc = fmaf(pc, x2, 4.1666667e-2f);
    pc = fmaf(pc, x2, -0.5f);
    *c = fmaf(pc, x2, 1.f);
}

__device__ __forceinline__ Q4 qmul(Q4 a, Q4 b) {
    return {a.w * b.w - a.x * b.x - a.y * b.y - a.z * b.z, a.w * b.x + a.x * b.w + a.y * b.z - a.z * b.y,
            a.w * b.y - a.x * b.z + a.y * b.w + a.z * b.x, a.w * b.z + a.x * b.y - a.y * b.x + a.z * b.w};
}
struct M3 { float m[9]; };  // row-major
__device__ __forceinline__ M3 qmat(Q4 q) {
    const float xx = q.x * q.x, yy = q.y * q.y, zz = q.z * q.z;
    const float xy = q.x * q.y, xz = q.x * q.z, yz = q.y * q.z;
    const float wx = q.w * q.x, wy = q.w * q.y, wz = q.w * q.z;
    M3 R;
    R.m[0] = 1.f - 2.f * (yy + zz); R.m[1] = 2.f * (xy - wz);       R.m[2] = 2.f * (xz + wy);
    R.m[3] = 2.f * (xy + wz);       R.m[4] = 1.f - 2.f * (xx + zz); R.m[5] = 2.f * (yz - wx);
    R.m[6] = 2.f * (xz - wy);       R.m[7] = 2.f * (yz + wx);       R.m[8] = 1.f - 2.f * (xx + yy);
    return R;
}
__device__ __forceinline__ V3 mulv(const M3& R, V3 v) {
    return {R.m[0] * v.x + R.m[1] * v.y + R.m[2] * v.z, R.m[3] * v.x + R.m[4] * v.y + R.m[5] * v.z,
            R.m[6] * v.x + R.m[7] * v.y + R.m[8] * v.z};
}
__device__ __forceinline__ V3 mulTv(const M3& R, V3 v) {
    return {R.m[0] * v.x + R.m[3] * v.y + R.m[6] * v.z, R.m[1] * v.x + R.m[4] * v.y + R.m[7] * v.z,
            R.m[2] * v.x + R.m[5] * v.y + R.m[8] * v.z};
}

// symmetric 6x6 in 21 registers; index of (i,j), i<=j
__host__ __device__ constexpr int sidx(int i, int j) {
    return i <= j ? (i * 6 - (i * (i - 1)) / 2 + (j - i)) : (j * 6 - (j * (j - 1)) / 2 + (i - j));
}
struct S6 { float a[21]; };
struct V6 { float a[6]; };  // (ang xyz, lin xyz)

__device__ __forceinline__ V6 s6mul(const S6& I, const V6& x) {
    V6 y;
#pragma unroll
    for (int i = 0; i < 6; ++i) {
        float acc = 0.f;
#pragma unroll
        for (int j = 0; j < 6; ++j) acc = fmaf(I.a[sidx(i, j)], x.a[j], acc);
        y.a[i] = acc;
    }
    return y;
}
__device__ __forceinline__ float dot6(const V6& a, const V6& b) {
    float acc = 0.f;
#pragma unroll
    for (int i = 0; i < 6; ++i) acc = fmaf(a.a[i], b.a[i], acc);
    return acc;
}
__device__ __forceinline__ V3 ang(const V6& v) { return {v.a[0], v.a[1], v.a[2]}; }
__device__ __forceinline__ V3 lin(const V6& v) { return {v.a[3], v.a[4], v.a[5]}; }
__device__ __forceinline__ V6 mk6(V3 a, V3 l) { return {{a.x, a.y, a.z, l.x, l.y, l.z}}; }

// Motion subspace of a joint whose axis is `aw` (world-aligned) and whose anchor is at p (rel. O):
// hinge (aw, p x aw), prismatic (0, aw).
__device__ __forceinline__ V6 joint_subspace(bool prismatic, V3 aw, V3 p) {
    const V3 zero = {0.f, 0.f, 0.f};
    return prismatic ? mk6(zero, aw) : mk6(aw, cross(p, aw));
}
// Velocity-product acceleration of a joint, c = qd * (v x s) (spatial motion cross product).
__device__ __forceinline__ V6 velocity_product(const V6& v, const V6& s, float qd) {
    const V3 w = ang(v), vl = lin(v), sa = ang(s);
    const V3 ca = cross(w, sa);
    const V3 cl = cross(w, lin(s)) + cross(vl, sa);
    return mk6(qd * ca, qd * cl);
}

__device__ __forceinline__ float shfl(float v, int src) { return __shfl_sync(0xffffffffu, v, src); }

// Ground contact of one sphere attached to a link (see DESIGN.md "contact model").
// Inputs in the substep's O-frame; root_z = absolute height of O.  Returns false if inactive.
struct Contact {
    V3 r;    // application point (bottom of the sphere), rel. O
    V3 F0;   // force at the current state incl. the velocity-product correction C*(P b)
    V3 C;    // dt * G (diagonal: tangential, tangential, normal)
};
template <class MT>
__device__ __forceinline__ bool contact_eval(const MT& M, int s, const M3& R, V3 p, const V6& v, const V6& b,
                                             float root_z, float dt, Contact& c) {
    const V3 local = {M.sph_pos[0][s], M.sph_pos[1][s], M.sph_pos[2][s]};
    const float rad = M.sph_r[s];
    const V3 x = p + mulv(R, local);
    const float pen_raw = rad - (root_z + x.z);
    if (!(pen_raw > 0.f)) return false;
    // max_depenetration_velocity (cfg/task/Humanoid.yaml:62,90): the spring sees at most v_dep * dt of penetration
    const float pen = fminf(pen_raw, M.pen_cap_v * dt);
    c.r = {x.x, x.y, x.z - rad};
    const V3 vp = lin(v) + cross(ang(v), c.r);
    const float vn = vp.z;
    const float Gn = dt * M.k_c + (vn < 0.f ? M.d_c : 0.f);
    const float fn0 = M.k_c * pen - Gn * vn;
    if (!(fn0 > 0.f)) return false;
    const float vt = sqrtf(vp.x * vp.x + vp.y * vp.y);
    const float ct = fminf(M.c_visc, M.mu * fn0 / (vt + 1e-6f));
    c.C = {dt * ct, dt * ct, dt * Gn};
    const V3 ab = lin(b) + cross(ang(b), c.r);
    c.F0 = {-ct * vp.x + c.C.x * ab.x, -ct * vp.y + c.C.y * ab.y, fn0 + c.C.z * ab.z};
    return true;
}

struct K2Args {
    float* pos; float* quat; float* vel; float* q; float* qd; float* sens;
    const float* actions;
    const float* targets;  // [N,D] PD position targets or null
    long long n;
    int substeps;
    float dt;
};

// ---- building blocks shared by the warp-per-env and the thread-per-env kernels ---------------------

// World-frame spatial inertia of link i about O (symmetric 6x6) and its bias force
// p = v x* (I v) - gravity wrench.  R, p: link frame in the O-frame; v: spatial velocity.
template <class MT>
__device__ __forceinline__ void link_inertia_bias(const MT& M, int li, const M3& R, V3 p, const V6& v, V3 grav,
                                                  S6& IA, V6& pA) {
    const float m = M.mass[li];
    const V3 cl = {M.com[0][li], M.com[1][li], M.com[2][li]};
    const V3 cw = p + mulv(R, cl);
    // Ic_world = R Iloc R^T
    const float ixx = M.inertia[0][li], iyy = M.inertia[1][li], izz = M.inertia[2][li];
    const float ixy = M.inertia[3][li], ixz = M.inertia[4][li], iyz = M.inertia[5][li];
    float T[9];  // T = R * Iloc
#pragma unroll
    for (int r = 0; r < 3; ++r) {
        const float r0 = R.m[3 * r], r1 = R.m[3 * r + 1], r2 = R.m[3 * r + 2];
        T[3 * r + 0] = r0 * ixx + r1 * ixy + r2 * ixz;
        T[3 * r + 1] = r0 * ixy + r1 * iyy + r2 * iyz;
        T[3 * r + 2] = r0 * ixz + r1 * iyz + r2 * izz;
    }
    const float cc = dot(cw, cw);
    const float cv[3] = {cw.x, cw.y, cw.z};
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = i; j < 3; ++j) {
            float icw = T[3 * i] * R.m[3 * j] + T[3 * i + 1] * R.m[3 * j + 1] + T[3 * i + 2] * R.m[3 * j + 2];
            icw += m * ((i == j ? cc : 0.f) - cv[i] * cv[j]);
            IA.a[sidx(i, j)] = icw;
        }
    // ang-lin block m*[c]x, lin-lin block m*1
    IA.a[sidx(0, 3)] = 0.f;         IA.a[sidx(0, 4)] = -m * cw.z;  IA.a[sidx(0, 5)] = m * cw.y;
    IA.a[sidx(1, 3)] = m * cw.z;    IA.a[sidx(1, 4)] = 0.f;        IA.a[sidx(1, 5)] = -m * cw.x;
    IA.a[sidx(2, 3)] = -m * cw.y;   IA.a[sidx(2, 4)] = m * cw.x;   IA.a[sidx(2, 5)] = 0.f;
    IA.a[sidx(3, 3)] = m; IA.a[sidx(3, 4)] = 0.f; IA.a[sidx(3, 5)] = 0.f;
    IA.a[sidx(4, 4)] = m; IA.a[sidx(4, 5)] = 0.f;
    IA.a[sidx(5, 5)] = m;
    const V6 h = s6mul(IA, v);
    const V3 w = ang(v), vl = lin(v);
    const V3 ba = cross(w, ang(h)) + cross(vl, lin(h));
    const V3 bl = cross(w, lin(h));
    const V3 fg = m * grav;
    const V3 tg = cross(cw, fg);
    pA = mk6(ba - tg, bl - fg);
    if (li > 0) {
        // RigidBodyAPI angular damping on every body: the torque -c Ic w about the link's centre of mass (a pure
        // couple, Ic = R Iloc R^T); the root's own damping is the velocity decay in integrate_root
        const V3 wl = mulTv(R, w);
        const float c = M.ang_damp;
        pA.a[0] = fmaf(c, T[0] * wl.x + T[1] * wl.y + T[2] * wl.z, pA.a[0]);
        pA.a[1] = fmaf(c, T[3] * wl.x + T[4] * wl.y + T[5] * wl.z, pA.a[1]);
        pA.a[2] = fmaf(c, T[6] * wl.x + T[7] * wl.y + T[8] * wl.z, pA.a[2]);
    }
}

// joint velocity limit: JointAPI max joint velocity; a hinge's relative angular velocity also obeys the body limit
template <class MT>
__device__ __forceinline__ float clamp_qd(const MT& M, int li, float qd) {
    const float m = M.jtype[li] == OIGE_JOINT_PRISMATIC ? M.max_qd_lin : M.max_qd;
    return fminf(fmaxf(qd, -m), m);
}

// Adds the linearly-implicit ground contacts of link li: P^T C P to the inertia, -wrench(F0) to the bias.
// `nloop` = loop bound (uniform across a warp-per-env warp: the model's max spheres per link).
template <class MT>
__device__ __forceinline__ void link_contacts(const MT& M, int li, const M3& R, V3 p, const V6& v, const V6& b,
                                              float root_z, float dt, S6& IA, V6& pA, int nloop = -1) {
    const int s0 = M.sph_start[li];
    const int sc = M.sph_count[li];
    const int n = nloop < 0 ? sc : nloop;
    for (int k = 0; k < n; ++k) {
        Contact ct;
        if (k < sc && contact_eval(M, s0 + k, R, p, v, b, root_z, dt, ct)) {
            const V3 tq = cross(ct.r, ct.F0);
            pA.a[0] -= tq.x; pA.a[1] -= tq.y; pA.a[2] -= tq.z;
            pA.a[3] -= ct.F0.x; pA.a[4] -= ct.F0.y; pA.a[5] -= ct.F0.z;
            // I += sum_k C_k u_k u_k^T with u_k = (r x e_k, e_k)
            const float Cd[3] = {ct.C.x, ct.C.y, ct.C.z};
            const float u[3][3] = {{0.f, ct.r.z, -ct.r.y}, {-ct.r.z, 0.f, ct.r.x}, {ct.r.y, -ct.r.x, 0.f}};
#pragma unroll
            for (int e = 0; e < 3; ++e) {
#pragma unroll
                for (int i = 0; i < 3; ++i) {
#pragma unroll
                    for (int j = i; j < 3; ++j) IA.a[sidx(i, j)] = fmaf(Cd[e] * u[e][i], u[e][j], IA.a[sidx(i, j)]);
                    IA.a[sidx(i, 3 + e)] = fmaf(Cd[e], u[e][i], IA.a[sidx(i, 3 + e)]);
                }
                IA.a[sidx(3 + e, 3 + e)] += Cd[e];
            }
        }
    }
}

// Joint-space force at the end-of-step position (actuation + passive spring/damper + limit penalty)
// and the implicit diagonal term armature + dt D + dt^2 K.
template <class MT>
__device__ __forceinline__ void joint_force(const MT& M, int li, float q, float qd, float effort, float dt,
                                            bool has_target, float target, float& tau0, float& dextra) {
    const float Ks = M.stiffness[li], Dj = M.damping[li];
    const float qn = fmaf(dt, qd, q);  // position the springs see at the end of the step
    float Kimp = Ks, Dimp = Dj;
    tau0 = effort - Ks * (qn - M.sref[li]) - Dj * qd;
    const float lo = M.lo[li], hi = M.hi[li];
    if (q < lo) {
        tau0 += -M.k_lim * (qn - lo) - M.d_lim * qd;
        Kimp += M.k_lim; Dimp += M.d_lim;
    } else if (q > hi) {
        tau0 += -M.k_lim * (qn - hi) - M.d_lim * qd;
        Kimp += M.k_lim; Dimp += M.d_lim;
    }
    if (has_target) {
        // PD position drive (ArticulationView.set_joint_position_targets): implicit like the passive spring.
        // Force limit: when the drive torque at the current state exceeds it, both gains are scaled by
        // fmax / |torque| -- the torque is then exactly +-fmax and the drive stays implicit (and passive), so a
        // saturated stiff drive cannot chatter the way a constant explicit torque would.
        const float kp = M.drive_kp[li], kd = M.drive_kd[li], fmax = M.drive_fmax[li];
        const float tpd = kp * (target - qn) - kd * qd;
        const float alpha = (fmax > 0.f && fabsf(tpd) > fmax) ? fmax / fabsf(tpd) : 1.f;
        tau0 += alpha * tpd;
        Kimp += alpha * kp; Dimp += alpha * kd;
    }
    dextra = M.armature[li] + dt * (Dimp + dt * Kimp);
}

// a = -(I^A)^-1 p^A for the floating root: 6x6 Cholesky in registers.
__device__ __forceinline__ void chol6_solve(const S6& IA, const V6& pA, V6& a) {
    float Lm[6][6];
#pragma unroll
    for (int i = 0; i < 6; ++i)
#pragma unroll
        for (int j = 0; j <= i; ++j) Lm[i][j] = IA.a[sidx(j, i)];
    float invd[6];
#pragma unroll
    for (int j = 0; j < 6; ++j) {
        float dj = Lm[j][j];
#pragma unroll
        for (int k = 0; k < j; ++k) dj = fmaf(-Lm[j][k], Lm[j][k], dj);
        const float ljj = sqrtf(dj);
        invd[j] = 1.f / ljj;
        Lm[j][j] = ljj;
#pragma unroll
        for (int i = j + 1; i < 6; ++i) {
            float t = Lm[i][j];
#pragma unroll
            for (int k = 0; k < j; ++k) t = fmaf(-Lm[i][k], Lm[j][k], t);
            Lm[i][j] = t * invd[j];
        }
    }
    float y[6];
#pragma unroll
    for (int i = 0; i < 6; ++i) {
        float t = -pA.a[i];
#pragma unroll
        for (int k = 0; k < i; ++k) t = fmaf(-Lm[i][k], y[k], t);
        y[i] = t * invd[i];
    }
#pragma unroll
    for (int i = 5; i >= 0; --i) {
        float t = y[i];
#pragma unroll
        for (int k = i + 1; k < 6; ++k) t = fmaf(-Lm[k][i], a.a[k], t);
        a.a[i] = t * invd[i];
    }
}

// Net ground wrench on link li (force, torque about the link origin) in the link frame -> out[6].
template <class MT>
__device__ __forceinline__ void sensor_wrench(const MT& M, int li, const M3& R, V3 p, const V6& v, const V6& b,
                                              const V6& a, float root_z, float dt, float* out) {
    V3 F = {0.f, 0.f, 0.f}, Tq = {0.f, 0.f, 0.f};
    const int s0 = M.sph_start[li], sc = M.sph_count[li];
    for (int k = 0; k < sc; ++k) {
        Contact ct;
        if (contact_eval(M, s0 + k, R, p, v, b, root_z, dt, ct)) {
            const V3 ap = lin(a) + cross(ang(a), ct.r);
            const V3 f = {ct.F0.x - ct.C.x * ap.x, ct.F0.y - ct.C.y * ap.y, ct.F0.z - ct.C.z * ap.z};
            F = F + f;
            Tq = Tq + cross(ct.r - p, f);
        }
    }
    const V3 Fl = mulTv(R, F), Tl = mulTv(R, Tq);
    out[0] = Fl.x; out[1] = Fl.y; out[2] = Fl.z; out[3] = Tl.x; out[4] = Tl.y; out[5] = Tl.z;
}

// Semi-implicit Euler for the floating root (exponential map for the quaternion).
template <class MT>
__device__ __forceinline__ void integrate_root(const MT& M, const V6& a, float dt, V3& rpos, Q4& rquat, V6& rvel) {
    const V3 w0 = ang(rvel), v0 = lin(rvel);
    const V3 vdot = lin(a) + cross(w0, v0);  // classical acceleration of the root origin
    float damp = 1.f - dt * M.ang_damp;
    damp = damp < 0.f ? 0.f : damp;
    V3 w1 = damp * (w0 + dt * ang(a));
    const float w2 = dot(w1, w1);
    if (w2 > M.max_w * M.max_w) w1 = (M.max_w * rsqrtf(w2)) * w1;  // RigidBodyAPI max angular velocity of the root body
    const V3 v1 = v0 + dt * vdot;
    rvel = mk6(w1, v1);
    rpos = rpos + dt * v1;
    const float wn = sqrtf(dot(w1, w1));
    const float th = 0.5f * dt * wn;
    float sh, ch;
    sincosf(th, &sh, &ch);
    const float k = wn > 1e-9f ? sh / wn : 0.5f * dt;
    const Q4 dq = {ch, k * w1.x, k * w1.y, k * w1.z};
    const Q4 nq = qmul(dq, rquat);
    const float inv = rsqrtf(nq.w * nq.w + nq.x * nq.x + nq.y * nq.y + nq.z * nq.z);
    rquat = {nq.w * inv, nq.x * inv, nq.y * inv, nq.z * inv};
}

template <bool PACKED>
__global__ void __launch_bounds__(kWarpsPerBlock * 32)
articulation_step_kernel(const DevModel* __restrict__ gmodel, const __grid_constant__ OigeTaskParams P,
                         const __grid_constant__ K2Args A) {
    __shared__ __align__(16) DevModel M;
    {
        const int4* src = reinterpret_cast<const int4*>(gmodel);
        int4* dst = reinterpret_cast<int4*>(&M);
        for (int i = threadIdx.x; i < (int)(sizeof(DevModel) / 16); i += blockDim.x) dst[i] = src[i];
    }
    __syncthreads();

    // Small robots are PACKED: a warp carries `pack` = floor(32 / num_links) environments side by side
    // (Ant: 3 x 9 lanes, Humanoid: 1 x 22).  Lane = (slot, link); every shuffle source is offset by the
    // slot's first lane, so the packed envs run the same instruction stream without interacting.
    const int lane = threadIdx.x & 31;
    const int L = M.num_links;
    const int pack = PACKED ? M.pack : 1;
    const long long warp_id = (long long)blockIdx.x * kWarpsPerBlock + (threadIdx.x >> 5);
    if (warp_id * pack >= A.n) return;  // warp-uniform
    const int slot = PACKED ? lane / L : (lane < L ? 0 : 1);
    const long long env = warp_id * pack + slot;
    const bool active = slot < pack && env < A.n;
    const int li = active ? lane - slot * L : 0;  // link index (clamped so table reads stay in range)
    const int ebase = active ? slot * L : lane;   // first lane of this env (inactive lanes talk to themselves)
    const bool is_root = active && li == 0;
    const int level = active ? M.level[li] : -1;
    const int parent = active && li > 0 ? ebase + M.parent[li] : lane;
    const int dofi = active ? M.dof[li] : -1;
    const bool is_joint = active && li > 0;
    const bool prismatic = is_joint && M.jtype[li] == OIGE_JOINT_PRISMATIC;
    const bool fixed_base = M.fixed_base != 0;
    const float dt = A.dt;
    const int D = P.num_dof;

    // ---- load state -------------------------------------------------------------------------
    float q = 0.f, qd = 0.f, effort = 0.f, target = 0.f;
    if (is_joint) {
        q = A.q[env * D + dofi];
        qd = A.qd[env * D + dofi];
        effort = A.actions[env * D + dofi] * P.joint_gears[dofi] * P.power_scale;  // locomotion.py:109
        if (A.targets != nullptr) target = A.targets[env * D + dofi];
    }
    V3 rpos = {0.f, 0.f, 0.f};
    Q4 rquat = {1.f, 0.f, 0.f, 0.f};
    V6 rvel = {{0.f, 0.f, 0.f, 0.f, 0.f, 0.f}};  // (ang, lin)
    if (is_root) {
        rpos = {A.pos[env * 3], A.pos[env * 3 + 1], A.pos[env * 3 + 2]};
        const float4 t = *reinterpret_cast<const float4*>(A.quat + env * 4);
        rquat = {t.x, t.y, t.z, t.w};
        const float* vv = A.vel + env * 6;
        rvel = {{vv[3], vv[4], vv[5], vv[0], vv[1], vv[2]}};
    }

    // per-lane constants
    const V3 jpos = {M.jpos[0][li], M.jpos[1][li], M.jpos[2][li]};
    const Q4 jquat = {M.jquat[0][li], M.jquat[1][li], M.jquat[2][li], M.jquat[3][li]};
    const V3 axis = {M.axis[0][li], M.axis[1][li], M.axis[2][li]};
    const V3 grav = {M.gravity[0], M.gravity[1], M.gravity[2]};
    const int max_depth = M.max_depth;

    for (int step = 0; step < A.substeps; ++step) {
        const bool last = (step == A.substeps - 1);
        const float root_z = shfl(rpos.z, ebase);

        // ---- pass 1: kinematics by POINTER JUMPING (log2(depth) rounds instead of depth levels) -------
        // Every lane starts with its transform relative to its parent, T_i = (R_i, p_i); round k
        // composes it with the transform of its 2^k-th ancestor (T_i <- T_anc o T_i), so after
        // ceil(log2(depth)) rounds all transforms are relative to the root.  The root carries the
        // identity, so links that already reached it are left unchanged.
        M3 R;
        V3 p = {0.f, 0.f, 0.f};
        {
            Q4 ql = {1.f, 0.f, 0.f, 0.f};
            if (is_joint) {
                float sh, ch;
                sincos_half_angle(prismatic ? 0.f : 0.5f * q, &sh, &ch);
                ql = qmul(jquat, Q4{ch, sh * axis.x, sh * axis.y, sh * axis.z});
                p = jpos;
            }
            R = qmat(ql);
            if (prismatic) p = p + q * mulv(R, axis);  // slide along the joint axis (fixed in the parent frame)
        }
        for (int k = 0; k < M.jump_rounds; ++k) {
            const int an = active ? ebase + M.anc[k][li] : lane;
            M3 Ra;
            V3 pa;
#pragma unroll
            for (int j = 0; j < 9; ++j) Ra.m[j] = shfl(R.m[j], an);
            pa = {shfl(p.x, an), shfl(p.y, an), shfl(p.z, an)};
            p = pa + mulv(Ra, p);
            M3 Rn;
#pragma unroll
            for (int r = 0; r < 3; ++r)
#pragma unroll
                for (int cidx = 0; cidx < 3; ++cidx)
                    Rn.m[3 * r + cidx] = Ra.m[3 * r] * R.m[cidx] + Ra.m[3 * r + 1] * R.m[3 + cidx] + Ra.m[3 * r + 2] * R.m[6 + cidx];
            R = Rn;
        }
        {   // root-relative -> world-aligned axes about O (the root origin)
            const M3 R0l = qmat(rquat);
            M3 R0;
#pragma unroll
            for (int j = 0; j < 9; ++j) R0.m[j] = shfl(R0l.m[j], ebase);
            p = mulv(R0, p);
            M3 Rn;
#pragma unroll
            for (int r = 0; r < 3; ++r)
#pragma unroll
                for (int cidx = 0; cidx < 3; ++cidx)
                    Rn.m[3 * r + cidx] = R0.m[3 * r] * R.m[cidx] + R0.m[3 * r + 1] * R.m[3 + cidx] + R0.m[3 * r + 2] * R.m[6 + cidx];
            R = Rn;
        }
        // joint motion subspace, velocities (prefix sums of s_j qd_j over the ancestor path, same jumps)
        V6 s = {{0.f, 0.f, 0.f, 0.f, 0.f, 0.f}};
        V6 v = {{0.f, 0.f, 0.f, 0.f, 0.f, 0.f}};
        if (is_joint) {
            s = joint_subspace(prismatic, mulv(R, axis), p);
#pragma unroll
            for (int k = 0; k < 6; ++k) v.a[k] = s.a[k] * qd;
        }
        for (int k = 0; k < M.jump_rounds; ++k) {
            const int an = active ? ebase + M.anc[k][li] : lane;
#pragma unroll
            for (int j = 0; j < 6; ++j) {
                const float t = shfl(v.a[j], an);
                v.a[j] += (an != lane) ? t : 0.f;
            }
        }
        V6 v0;
#pragma unroll
        for (int j = 0; j < 6; ++j) {
            v0.a[j] = shfl(rvel.a[j], ebase);
            v.a[j] += v0.a[j];
        }
        // velocity-product accelerations c_i and their path sums b_i (b_root = -(0, w0 x v0))
        V6 c = {{0.f, 0.f, 0.f, 0.f, 0.f, 0.f}};
        if (is_joint) c = velocity_product(v, s, qd);
        V6 b = c;
        for (int k = 0; k < M.jump_rounds; ++k) {
            const int an = active ? ebase + M.anc[k][li] : lane;
#pragma unroll
            for (int j = 0; j < 6; ++j) {
                const float t = shfl(b.a[j], an);
                b.a[j] += (an != lane) ? t : 0.f;
            }
        }
        {
            const V3 wxv = cross(ang(v0), lin(v0));
            b.a[3] -= wxv.x; b.a[4] -= wxv.y; b.a[5] -= wxv.z;
        }

        // ---- per-link spatial inertia, bias force, contacts (all lanes in parallel) -------------
        S6 IA;
        V6 pA;
        link_inertia_bias(M, li, R, p, v, grav, IA, pA);
        link_contacts(M, li, R, p, v, b, root_z, dt, IA, pA, M.max_sph_per_link);

        // ---- joint-space force and implicit diagonal ----------------------------------------------
        float tau0 = 0.f, dextra = 1.f;
        if (is_joint) joint_force(M, li, q, qd, effort, dt, A.targets != nullptr, target, tau0, dextra);

        // ---- pass 2: articulated inertia, bottom-up -------------------------------------------------
        V6 U = {{0.f, 0.f, 0.f, 0.f, 0.f, 0.f}};
        float dinv = 0.f, u = 0.f;
        for (int l = max_depth; l >= 1; --l) {
            if (level == l) {
                U = s6mul(IA, s);
                const float d = dot6(s, U) + dextra;
                dinv = 1.f / d;
                u = tau0 - dot6(s, pA);
                // I^a = I^A - U U^T / d   (in place)
#pragma unroll
                for (int i = 0; i < 6; ++i) {
                    const float ui = U.a[i] * dinv;
#pragma unroll
                    for (int j = i; j < 6; ++j) IA.a[sidx(i, j)] = fmaf(-ui, U.a[j], IA.a[sidx(i, j)]);
                }
                // p^a = p^A + I^a c + U u / d
                const V6 Iac = s6mul(IA, c);
                const float ud = u * dinv;
#pragma unroll
                for (int i = 0; i < 6; ++i) pA.a[i] = pA.a[i] + Iac.a[i] + U.a[i] * ud;
            }
            const int nch = M.maxch_at_level[l - 1];
            for (int k = 0; k < nch; ++k) {
                const int ch = (level == l - 1) ? M.child[k][li] : -1;
                const bool valid = ch >= 0;
                const int src = valid ? ebase + ch : lane;
#pragma unroll
                for (int j = 0; j < 21; ++j) {
                    const float t = shfl(IA.a[j], src);
                    if (valid) IA.a[j] += t;
                }
#pragma unroll
                for (int j = 0; j < 6; ++j) {
                    const float t = shfl(pA.a[j], src);
                    if (valid) pA.a[j] += t;
                }
            }
        }

        // ---- root: solve I^A a = -p^A (6x6 Cholesky, root lane) -------------------------------------
        V6 a = {{0.f, 0.f, 0.f, 0.f, 0.f, 0.f}};
        if (is_root && !fixed_base) chol6_solve(IA, pA, a);  // a welded base does not accelerate

        // ---- pass 3: accelerations, top-down -----------------------------------------------------------
        float qdd = 0.f;
        for (int l = 1; l <= max_depth; ++l) {
            const bool mine = (level == l);
            const int src = mine ? parent : lane;
            V6 pa;
#pragma unroll
            for (int k = 0; k < 6; ++k) pa.a[k] = shfl(a.a[k], src);
            if (mine) {
#pragma unroll
                for (int k = 0; k < 6; ++k) pa.a[k] += c.a[k];
                qdd = (u - dot6(U, pa)) * dinv;
#pragma unroll
                for (int k = 0; k < 6; ++k) a.a[k] = fmaf(s.a[k], qdd, pa.a[k]);
            }
        }

        // ---- foot force sensors: net ground wrench on the sensor link, in the link frame ----------------
        if (last && A.sens != nullptr) {
            const int sidx_ = active ? M.sensor_of_link[li] : -1;
            if (sidx_ >= 0) sensor_wrench(M, li, R, p, v, b, a, root_z, dt, A.sens + (env * M.num_sensors + sidx_) * 6);
        }

        // ---- integrate (semi-implicit Euler) ----------------------------------------------------------------
        if (is_joint) {
            qd = fmaf(dt, qdd, qd);
            qd = clamp_qd(M, li, qd);
            q = fmaf(dt, qd, q);
        }
        if (is_root && !fixed_base) integrate_root(M, a, dt, rpos, rquat, rvel);
    }

    // ---- non-finite guard: park the env below the termination height so K1 flags it for reset ----
    bool bad = false;
    if (is_joint) bad = !(isfinite(q) && isfinite(qd));
    if (is_root) {
        bad = !(isfinite(rpos.x) && isfinite(rpos.y) && isfinite(rpos.z) && isfinite(rquat.w) && isfinite(rquat.x) &&
                isfinite(rquat.y) && isfinite(rquat.z));
#pragma unroll
        for (int k = 0; k < 6; ++k) bad = bad || !isfinite(rvel.a[k]);
    }
    const unsigned bad_lanes = __ballot_sync(0xffffffffu, bad);
    const unsigned env_lanes = active ? (((L >= 32 ? 0u : (1u << L)) - 1u) << ebase) : 0u;
    if (bad_lanes & env_lanes) {  // any lane of THIS env
        if (is_root) atomicAdd(&g_nonfinite_envs, 1ull);  // one lane per env
        q = 0.f; qd = 0.f;
        if (!fixed_base) {  // a welded base stays where it is; only the joints are zeroed
            rpos = {0.f, 0.f, -1.f};
            rquat = {1.f, 0.f, 0.f, 0.f};
            rvel = {{0.f, 0.f, 0.f, 0.f, 0.f, 0.f}};
        }
        if (A.sens != nullptr)
            for (int k = li; k < M.num_sensors * 6; k += L) A.sens[env * M.num_sensors * 6 + k] = 0.f;
    }

    // ---- store state ---------------------------------------------------------------------------------------
    if (is_joint) {
        A.q[env * D + dofi] = q;
        A.qd[env * D + dofi] = qd;
    }
    if (is_root) {
        A.pos[env * 3] = rpos.x; A.pos[env * 3 + 1] = rpos.y; A.pos[env * 3 + 2] = rpos.z;
        *reinterpret_cast<float4*>(A.quat + env * 4) = make_float4(rquat.w, rquat.x, rquat.y, rquat.z);
        float* vv = A.vel + env * 6;
        vv[0] = rvel.a[3]; vv[1] = rvel.a[4]; vv[2] = rvel.a[5];
        vv[3] = rvel.a[0]; vv[4] = rvel.a[1]; vv[5] = rvel.a[2];
    }
}

// ================================================================================================
// K2-T: the same articulation step with ONE THREAD PER ENVIRONMENT (lane = env).
//
// The warp-per-env kernel above is latency-optimal for few environments, but a lane-per-link warp
// issues every instruction of the tree recursion for all 32 lanes while 1-4 links are active, so its
// issue cost is ~4500 warp-instructions per env-substep.  Here every lane does useful work: a warp
// advances 32 envs at ~17k instructions per substep (~530 per env).  All lanes of a warp visit the
// links in the same order, so the model-table reads are shared-memory broadcasts and the control
// flow is uniform except for contact activation.  Per-link intermediates that must survive between
// the passes live in per-thread local memory (interleaved by the hardware, hence coalesced):
//   fk[i]  = quat(4) p(3) v(6) b(6)           forward kinematics, written in pass 1, read in passes 2, 3
//   pu[i]  = U(6) 1/d (u - U.c)               written in pass 2, read in pass 3
// (27 floats per link: the arrays are write-once/read-once per substep, so their size IS the L2/HBM
// traffic of the kernel; rotation matrices, s and c are recomputed instead of stored.)
// Articulated inertias travel up a heavy path in REGISTERS (links are visited in post-order with the
// heavy child last, so a link directly follows its heavy child); only light children (chain heads)
// go through a small slot array.  Same equations, same contact model, same integrator as above.
// ================================================================================================
template <int MAXL, int MINB>
__global__ void __launch_bounds__(128, MINB)
articulation_step_tpe_kernel(const DevModel* __restrict__ gmodel, const __grid_constant__ OigeTaskParams P,
                             const __grid_constant__ K2Args A) {
    __shared__ __align__(16) DevModel M;
    {
        const int4* src = reinterpret_cast<const int4*>(gmodel);
        int4* dst = reinterpret_cast<int4*>(&M);
        for (int i = threadIdx.x; i < (int)(sizeof(DevModel) / 16); i += blockDim.x) dst[i] = src[i];
    }
    __syncthreads();
    const long long env = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (env >= A.n) return;

    const int L = M.num_links, D = P.num_dof;
    const float dt = A.dt;
    const V3 grav = {M.gravity[0], M.gravity[1], M.gravity[2]};

    float fk[MAXL][19];
    float pu[MAXL][8];
    float slotI[kMaxSlots][27];
    float slotA[kMaxSlots][6];
    float qv[MAXL], qdv[MAXL], eff[MAXL], tgt[MAXL];

    for (int i = 1; i < L; ++i) {
        const int d = M.dof[i];
        qv[i] = A.q[env * D + d];
        qdv[i] = A.qd[env * D + d];
        eff[i] = A.actions[env * D + d] * P.joint_gears[d] * P.power_scale;  // locomotion.py:109
        tgt[i] = A.targets != nullptr ? A.targets[env * D + d] : 0.f;
    }
    V3 rpos = {A.pos[env * 3], A.pos[env * 3 + 1], A.pos[env * 3 + 2]};
    Q4 rquat;
    {
        const float4 t = *reinterpret_cast<const float4*>(A.quat + env * 4);
        rquat = {t.x, t.y, t.z, t.w};
    }
    V6 rvel;
    {
        const float* vv = A.vel + env * 6;
        rvel = {{vv[3], vv[4], vv[5], vv[0], vv[1], vv[2]}};
    }

    for (int step = 0; step < A.substeps; ++step) {
        const bool last = (step == A.substeps - 1);
        const float root_z = rpos.z;
        const M3 R0 = qmat(rquat);
        V6 b0;
        {
            const V3 wxv = cross(ang(rvel), lin(rvel));
            b0 = {{0.f, 0.f, 0.f, -wxv.x, -wxv.y, -wxv.z}};
        }
        // ---- pass 1: kinematics in index order (parent[i] < i) ----------------------------------
        for (int i = 1; i < L; ++i) {
            const int pa = M.parent[i];
            Q4 qp;
            V3 pp;
            V6 pv, pb;
            if (pa == 0) {
                qp = rquat; pp = {0.f, 0.f, 0.f}; pv = rvel; pb = b0;
            } else {
                qp = {fk[pa][0], fk[pa][1], fk[pa][2], fk[pa][3]};
                pp = {fk[pa][4], fk[pa][5], fk[pa][6]};
#pragma unroll
                for (int k = 0; k < 6; ++k) { pv.a[k] = fk[pa][7 + k]; pb.a[k] = fk[pa][13 + k]; }
            }
            const M3 Rp = qmat(qp);
            const V3 axis = {M.axis[0][i], M.axis[1][i], M.axis[2][i]};
            float sh, ch;
            const bool prismatic = M.jtype[i] == OIGE_JOINT_PRISMATIC;
            sincos_half_angle(prismatic ? 0.f : 0.5f * qv[i], &sh, &ch);
            const Q4 ql = qmul(Q4{M.jquat[0][i], M.jquat[1][i], M.jquat[2][i], M.jquat[3][i]},
                               Q4{ch, sh * axis.x, sh * axis.y, sh * axis.z});
            const Q4 qi = qmul(qp, ql);
            V3 p = pp + mulv(Rp, V3{M.jpos[0][i], M.jpos[1][i], M.jpos[2][i]});
            const M3 R = qmat(qi);
            const V3 aw = mulv(R, axis);
            if (prismatic) p = p + qv[i] * aw;
            const V6 s = joint_subspace(prismatic, aw, p);
            const float qd = qdv[i];
            V6 v;
#pragma unroll
            for (int k = 0; k < 6; ++k) v.a[k] = fmaf(s.a[k], qd, pv.a[k]);
            const V3 w = ang(v), vl = lin(v);
            const V3 ca = cross(w, ang(s));
            const V3 cl = cross(w, lin(s)) + cross(vl, ang(s));
            fk[i][0] = qi.w; fk[i][1] = qi.x; fk[i][2] = qi.y; fk[i][3] = qi.z;
            fk[i][4] = p.x; fk[i][5] = p.y; fk[i][6] = p.z;
#pragma unroll
            for (int k = 0; k < 6; ++k) fk[i][7 + k] = v.a[k];
            fk[i][13] = pb.a[0] + qd * ca.x; fk[i][14] = pb.a[1] + qd * ca.y; fk[i][15] = pb.a[2] + qd * ca.z;
            fk[i][16] = pb.a[3] + qd * cl.x; fk[i][17] = pb.a[4] + qd * cl.y; fk[i][18] = pb.a[5] + qd * cl.z;
        }

        // ---- pass 2: articulated inertia, post-order with the heavy child last ---------------------
        S6 carryI;
        V6 carryP;
#pragma unroll
        for (int k = 0; k < 21; ++k) carryI.a[k] = 0.f;
#pragma unroll
        for (int k = 0; k < 6; ++k) carryP.a[k] = 0.f;
        for (int k2 = 0; k2 <= L - 1; ++k2) {
            const int i = (k2 < L - 1) ? M.order2[k2] : 0;  // the root closes the sweep
            M3 R;
            V3 p;
            V6 v, b;
            if (i == 0) {
                R = R0; p = {0.f, 0.f, 0.f}; v = rvel; b = b0;
            } else {
                R = qmat(Q4{fk[i][0], fk[i][1], fk[i][2], fk[i][3]});
                p = {fk[i][4], fk[i][5], fk[i][6]};
#pragma unroll
                for (int k = 0; k < 6; ++k) { v.a[k] = fk[i][7 + k]; b.a[k] = fk[i][13 + k]; }
            }
            S6 IA;
            V6 pA;
            link_inertia_bias(M, i, R, p, v, grav, IA, pA);
            link_contacts(M, i, R, p, v, b, root_z, dt, IA, pA);
            const int flags = M.tflags[i];
            if (flags & kHasHeavy) {
#pragma unroll
                for (int k = 0; k < 21; ++k) IA.a[k] += carryI.a[k];
#pragma unroll
                for (int k = 0; k < 6; ++k) pA.a[k] += carryP.a[k];
            }
            const int myslot = M.light_slot[i];
            if (myslot >= 0) {  // light children finished earlier and left their I^a here
#pragma unroll
                for (int k = 0; k < 21; ++k) IA.a[k] += slotI[myslot][k];
#pragma unroll
                for (int k = 0; k < 6; ++k) pA.a[k] += slotI[myslot][21 + k];
            }
            if (i == 0) {
                carryI = IA; carryP = pA;  // root: solved below
                break;
            }
            const V3 axis = {M.axis[0][i], M.axis[1][i], M.axis[2][i]};
            const V6 s = joint_subspace(M.jtype[i] == OIGE_JOINT_PRISMATIC, mulv(R, axis), p);
            const float q = qv[i], qd = qdv[i];
            const V6 c = velocity_product(v, s, qd);
            float tau0, dextra;
            joint_force(M, i, q, qd, eff[i], dt, A.targets != nullptr, tgt[i], tau0, dextra);
            const V6 U = s6mul(IA, s);
            const float dinv = 1.f / (dot6(s, U) + dextra);
            const float u = tau0 - dot6(s, pA);
#pragma unroll
            for (int a1 = 0; a1 < 6; ++a1) {
                const float ui = U.a[a1] * dinv;
#pragma unroll
                for (int a2 = a1; a2 < 6; ++a2) IA.a[sidx(a1, a2)] = fmaf(-ui, U.a[a2], IA.a[sidx(a1, a2)]);
            }
            const V6 Iac = s6mul(IA, c);
            const float ud = u * dinv;
#pragma unroll
            for (int k = 0; k < 6; ++k) pA.a[k] = pA.a[k] + Iac.a[k] + U.a[k] * ud;
#pragma unroll
            for (int k = 0; k < 6; ++k) pu[i][k] = U.a[k];
            pu[i][6] = dinv; pu[i][7] = u - dot6(U, c);  // pass 3: qdd = (u - U.(a_parent + c)) / d
            if (flags & kIsHead) {  // light child: park I^a in the parent's slot
                const int ps = M.light_slot[M.parent[i]];
                if (flags & kFirstLight) {
#pragma unroll
                    for (int k = 0; k < 21; ++k) slotI[ps][k] = IA.a[k];
#pragma unroll
                    for (int k = 0; k < 6; ++k) slotI[ps][21 + k] = pA.a[k];
                } else {
#pragma unroll
                    for (int k = 0; k < 21; ++k) slotI[ps][k] += IA.a[k];
#pragma unroll
                    for (int k = 0; k < 6; ++k) slotI[ps][21 + k] += pA.a[k];
                }
            } else {
                carryI = IA; carryP = pA;
            }
        }

        // ---- root: solve I^A a = -p^A (6x6 Cholesky) ----------------------------------------------
        V6 a0 = {{0.f, 0.f, 0.f, 0.f, 0.f, 0.f}};
        if (!M.fixed_base) chol6_solve(carryI, carryP, a0);  // a welded base does not accelerate
        if (M.light_slot[0] >= 0) {
#pragma unroll
            for (int k = 0; k < 6; ++k) slotA[M.light_slot[0]][k] = a0.a[k];
        }

        // ---- pass 3: accelerations in reverse post-order (parent, then its heavy child) ---------------
        V6 acarry = a0;
        for (int k2 = L - 2; k2 >= 0; --k2) {
            const int i = M.order2[k2];
            const int flags = M.tflags[i];
            V6 ap;
            if (flags & kIsHead) {
                const int ps = M.light_slot[M.parent[i]];
#pragma unroll
                for (int k = 0; k < 6; ++k) ap.a[k] = slotA[ps][k];
            } else {
                ap = acarry;
            }
            // recompute s and c from the stored kinematics (cheaper than storing 12 more floats per link)
            const M3 R = qmat(Q4{fk[i][0], fk[i][1], fk[i][2], fk[i][3]});
            const V3 p = {fk[i][4], fk[i][5], fk[i][6]};
            V6 v;
#pragma unroll
            for (int k = 0; k < 6; ++k) v.a[k] = fk[i][7 + k];
            const V6 s = joint_subspace(M.jtype[i] == OIGE_JOINT_PRISMATIC,
                                        mulv(R, V3{M.axis[0][i], M.axis[1][i], M.axis[2][i]}), p);
            const float qd0 = qdv[i];
            V6 U;
#pragma unroll
            for (int k = 0; k < 6; ++k) U.a[k] = pu[i][k];
            const float qdd = (pu[i][7] - dot6(U, ap)) * pu[i][6];
            V6 a;
            {
                const V6 c = velocity_product(v, s, qd0);
#pragma unroll
                for (int k = 0; k < 6; ++k) a.a[k] = fmaf(s.a[k], qdd, ap.a[k] + c.a[k]);
            }
            const int myslot = M.light_slot[i];
            if (myslot >= 0) {
#pragma unroll
                for (int k = 0; k < 6; ++k) slotA[myslot][k] = a.a[k];
            }
            if (last && A.sens != nullptr && M.sensor_of_link[i] >= 0) {
                V6 b;
#pragma unroll
                for (int k = 0; k < 6; ++k) b.a[k] = fk[i][13 + k];
                sensor_wrench(M, i, R, p, v, b, a, root_z, dt, A.sens + (env * M.num_sensors + M.sensor_of_link[i]) * 6);
            }
            // integrate this joint (semi-implicit Euler); q/qd of link i are not read again this substep
            float qd = fmaf(dt, qdd, qdv[i]);
            qd = clamp_qd(M, i, qd);
            qdv[i] = qd;
            qv[i] = fmaf(dt, qd, qv[i]);
            acarry = a;
        }
        if (!M.fixed_base) integrate_root(M, a0, dt, rpos, rquat, rvel);
    }

    // ---- non-finite guard + store ---------------------------------------------------------------
    bool bad = !(isfinite(rpos.x) && isfinite(rpos.y) && isfinite(rpos.z) && isfinite(rquat.w) && isfinite(rquat.x) &&
                 isfinite(rquat.y) && isfinite(rquat.z));
#pragma unroll
    for (int k = 0; k < 6; ++k) bad = bad || !isfinite(rvel.a[k]);
    for (int i = 1; i < L; ++i) bad = bad || !(isfinite(qv[i]) && isfinite(qdv[i]));
    if (bad && !M.fixed_base) {
        rpos = {0.f, 0.f, -1.f};
        rquat = {1.f, 0.f, 0.f, 0.f};
        rvel = {{0.f, 0.f, 0.f, 0.f, 0.f, 0.f}};
    }
    if (bad) {
        atomicAdd(&g_nonfinite_envs, 1ull);
        if (A.sens != nullptr)
            for (int k = 0; k < M.num_sensors * 6; ++k) A.sens[env * M.num_sensors * 6 + k] = 0.f;
    }
    for (int i = 1; i < L; ++i) {
        const int d = M.dof[i];
        A.q[env * D + d] = bad ? 0.f : qv[i];
        A.qd[env * D + d] = bad ? 0.f : qdv[i];
    }
    A.pos[env * 3] = rpos.x; A.pos[env * 3 + 1] = rpos.y; A.pos[env * 3 + 2] = rpos.z;
    *reinterpret_cast<float4*>(A.quat + env * 4) = make_float4(rquat.w, rquat.x, rquat.y, rquat.z);
    float* vv = A.vel + env * 6;
    vv[0] = rvel.a[3]; vv[1] = rvel.a[4]; vv[2] = rvel.a[5];
    vv[3] = rvel.a[0]; vv[4] = rvel.a[1]; vv[5] = rvel.a[2];
}

#include "env_step_role.cuh"  // K2-F: role-split whole-step kernel (same device functions)

}  // namespace

// ---- run-time knobs (OIGE_K2_MODE, OIGE_K2_MINB, OIGE_K2_EPC): read ONCE, at the first use; oige_reload_env()
// re-reads them (the test-suite switches kernels between tests) ------------------------------------------------
struct K2Knobs {
    int mode;   // 0 auto, 'w' warp-per-env, 't' thread-per-env, 'r' role split
    int minb;   // 0 auto, 2, 4: register-budget variant of K2-T
    int epc;    // 0 auto, 32, 128: envs per CTA of K2-F
};
static K2Knobs g_knobs = {0, 0, 0};
static std::once_flag g_knobs_once;
static void read_knobs() {
    K2Knobs k = {0, 0, 0};
    const char* e = getenv("OIGE_K2_MODE");
    if (e && (e[0] == 'w' || e[0] == 't' || e[0] == 'r')) k.mode = e[0];
    const char* mb = getenv("OIGE_K2_MINB");
    if (mb && (mb[0] == '2' || mb[0] == '4')) k.minb = mb[0] - '0';
    const char* ep = getenv("OIGE_K2_EPC");
    if (ep) { const int v = atoi(ep); if (v == 32 || v == 128) k.epc = v; }
    g_knobs = k;
}
static const K2Knobs& knobs() {
    std::call_once(g_knobs_once, read_knobs);
    return g_knobs;
}
void k2_reload_env() { knobs(); read_knobs(); }

// Kernel choice: OIGE_K2_MODE=warp|thread forces one; otherwise thread-per-env from kTpeMinEnvs envs on.
static bool use_thread_per_env(int64_t n) {
    const int m = knobs().mode;
    if (m == 'w') return false;
    if (m == 't') return true;
    return n >= kTpeMinEnvs;
}

// Role-split kernel (K2-F): from the point where the warp-per-env kernel runs out of issue slots.  OIGE_K2_MODE=role
// forces it; warp|thread exclude it.
static bool use_role_split(int64_t n, int num_links) {
    const int m = knobs().mode;
    if (m == 'r') return true;
    if (m == 'w' || m == 't') return false;
    return n >= (num_links <= 12 ? kRoleMinEnvsSmall : kRoleMinEnvs);
}

bool k2_uses_role_kernel(const OigeModel* model, int64_t decision_n) {
    if (model->has_sc) {  // self-collision is integrated by K2-F only, at every env count (a forced warp / thread mode fails)
        const int m = knobs().mode;
        return model->role_ok && m != 'w' && m != 't';
    }
    if (!(model->role_ok && use_role_split(decision_n, model->num_links))) return false;
    // The four-role variants (Ant, Anymal: 128-thread CTAs of 32 envs at 128 registers, four per SM) hold 128 envs per SM
    // and wave; from the third wave on the thread-per-env path K3 -> K2-T -> K1 (512+ envs per SM) has the higher
    // throughput for these small trees (profiles/r2_env_sweep_1gpu.txt: Ant 65536 envs 154 vs 129 us per step).  The
    // Humanoid's three-role variant wins at every count (317 vs 280 M env-steps/s at 1 Mi envs).
    if (model->nroles == 4 && knobs().mode != 'r' && decision_n > 2ll * 4 * 32 * sm_count()) return false;
    return true;
}

namespace {

// dynamic shared memory of one K2-F CTA: the groups' scratch, aliased by the input tile and by the output tile
size_t role_smem_bytes(const OigeModel* model, int epc, int D, int S) {
    const int OBS = 12 + 3 * D + 6 * S;
    const size_t scratch = (size_t)(epc / 32) * model->group_floats * 32;
    const size_t tile_in = (size_t)epc * (3 * D + 13 + 2);  // q, qd, actions, root state, reset flags (int64)
    const size_t tile_out = (size_t)epc * (2 * D + 6 * S + 13) + (size_t)epc * (D + OBS + 6 + 1 + 12) + 128;
    size_t fl = scratch > tile_in ? scratch : tile_in;
    if (tile_out > fl) fl = tile_out;
    return fl * sizeof(float);
}

template <typename K>
int role_launch(K kern, int variant, int nthreads, int epc, const OigeModel* model, const OigeTaskParams* P,
                const FArgs& A, cudaStream_t stream) {
    static std::atomic<int> configured[8][64];  // [variant][device]: max dynamic smem the kernel was configured for
    const size_t smem = role_smem_bytes(model, epc, A.D, A.S);
    {   // a group's output tile (32 envs) aliases the group's own scratch
        const int OBS = 12 + 3 * A.D + 6 * A.S;
        const size_t group_tile = (size_t)32 * (2 * A.D + 6 * A.S + 13) + (size_t)32 * (A.D + OBS + 6 + 1 + 12) + 128;
        if (epc > 32 && group_tile > (size_t)model->group_floats * 32)
            return fail(OIGE_E_MODEL, "env_step_role_kernel: a group's output tile (%zu floats) does not fit in its scratch (%d)",
                        group_tile, model->group_floats * 32);
    }
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 0 || dev >= 64) dev = 0;
    if ((size_t)configured[variant][dev].load(std::memory_order_acquire) < smem) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return fail((int)e, "cudaFuncSetAttribute(env_step_role_kernel, %zu B): %s", smem, cudaGetErrorString(e));
        configured[variant][dev].store((int)smem, std::memory_order_release);
    }
    const unsigned grid = (unsigned)((A.n + epc - 1) / epc);
    launch_pdl(kern, grid, (unsigned)nthreads, smem, stream, reinterpret_cast<const RoleModel*>(model->dev_role), *P, A);
    count_launch();
    return check_launch("env_step_role_kernel");
}

// `dn`: env count of the whole step (chunked callers), decides the CTA shape so that chunks == one shot bit for bit
int role_dispatch(const OigeModel* model, const OigeTaskParams* P, const FArgs& A, cudaStream_t stream, int64_t dn) {
    const bool tgt = A.targets != nullptr;
    if (model->nroles == 3 && model->maxc == 6 && !tgt) {
        // up to three 32-env CTAs per SM fit in shared memory; beyond that one 128-env CTA per SM (the model table is
        // staged once per CTA, and four groups is what one wave of 16384 envs needs on 148 SMs)
        const int force = knobs().epc;
        const bool big = force ? force == 128 : (dn + 31) / 32 > 3ll * sm_count();
        if (model->has_sc) {  // self-collision pairs (Humanoid.yaml:80): the variant with the pair phase
            if (big) return role_launch(env_step_role_kernel<6, 3, false, 128, 1, true, true>, 4, 3 * 128, 128, model, P, A, stream);
            return role_launch(env_step_role_kernel<6, 3, false, 32, 3, true, true>, 5, 3 * 32, 32, model, P, A, stream);
        }
        if (big) return role_launch(env_step_role_kernel<6, 3, false, 128, 1, true, false>, 0, 3 * 128, 128, model, P, A, stream);
        return role_launch(env_step_role_kernel<6, 3, false, 32, 3, true, false>, 1, 3 * 32, 32, model, P, A, stream);
    }
    if (model->nroles == 4 && model->maxc == 3 && !model->has_sc) {
        // up to three CTAs per SM cover the launch: 168 instead of 128 registers per thread (no spills)
        // (OIGE_K2_MINB=4 pins the 128-register shape, as it pins K2-T's: the two shapes are different ptxas schedules
        // and agree to rounding, not bit for bit)
        const int mb = knobs().minb;
        const bool roomy = mb ? mb < 4 : (dn + 31) / 32 <= 3ll * sm_count();
        if (tgt) {
            if (roomy) return role_launch(env_step_role_kernel<3, 4, true, 32, 3, false, false>, 6, 4 * 32, 32, model, P, A, stream);
            return role_launch(env_step_role_kernel<3, 4, true, 32, 4, false, false>, 2, 4 * 32, 32, model, P, A, stream);
        }
        if (roomy) return role_launch(env_step_role_kernel<3, 4, false, 32, 3, false, false>, 7, 4 * 32, 32, model, P, A, stream);
        return role_launch(env_step_role_kernel<3, 4, false, 32, 4, false, false>, 3, 4 * 32, 32, model, P, A, stream);
    }
    return fail(OIGE_E_MODEL, "env_step_role_kernel: no instantiation for %d roles x %d links%s", model->nroles,
                model->maxc, tgt ? " with position targets" : "");
}

bool role_variant_exists(const OigeModel* model, bool tgt) {
    return (model->nroles == 3 && model->maxc == 6 && !tgt) || (model->nroles == 4 && model->maxc == 3 && !model->has_sc);
}

inline bool al16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

}  // namespace

// Whole env step in one launch (K2-F with prologue and epilogue).  Returns OIGE_E_MODEL + 1000 ("not applicable") when
// the model / env count is not served by the role kernel -- the caller then runs the K3 -> K2 -> K1 sequence.
// Side channel of the host-buffer step (capi.cu): device-mapped pinned host rows for the reward / reset flags of the env
// range of the NEXT k2_fused_step call of this thread (consumed by it).
static thread_local float* t_rew_host = nullptr;
static thread_local long long* t_reset_host = nullptr;
void k2_set_host_mirror(float* rew_host, long long* reset_host) { t_rew_host = rew_host; t_reset_host = reset_host; }
bool k2_fused_applicable(const OigeModel* model, int64_t decision_n, bool with_targets) {
    return k2_uses_role_kernel(model, decision_n) && role_variant_exists(model, with_targets);
}

int k2_fused_step(const OigeModel* model, const OigeTaskParams* P, const OigeState* st, const OigeTaskBuffers* b,
                  const OigeResetInit* init, const float* actions_in, float clip, uint64_t seed, uint64_t counter,
                  int64_t env_id_offset, int64_t n, int32_t substeps, float dt, cudaStream_t stream, int64_t decision_n,
                  bool bump_counter) {
    float* const rew_host = t_rew_host;          // (consumed whatever happens below)
    long long* const reset_host = t_reset_host;
    t_rew_host = nullptr; t_reset_host = nullptr;
    const int64_t dn = decision_n > 0 ? decision_n : n;
    if (!k2_uses_role_kernel(model, dn) || !role_variant_exists(model, st->dof_pos_targets != nullptr)) return 1000;
    if (P->num_dof != model->num_dof)
        return fail(OIGE_E_SHAPE, "env_step: params.num_dof %d != model dofs %d", P->num_dof, model->num_dof);
    if (P->num_sensors != model->num_sensors || P->num_obs != 12 + 3 * P->num_dof + 6 * P->num_sensors)
        return fail(OIGE_E_SHAPE, "env_step: params (sensors %d, obs %d) do not match the model (%d sensors)",
                    P->num_sensors, P->num_obs, model->num_sensors);
    FArgs A{};
    A.pos = st->root_pos; A.quat = st->root_quat; A.vel = st->root_vel; A.q = st->dof_pos; A.qd = st->dof_vel;
    A.sens = st->sensors; A.targets = st->dof_pos_targets; A.actions = b->actions; A.n = n; A.substeps = substeps; A.dt = dt;
    A.D = P->num_dof; A.S = P->num_sensors; A.OBS = P->num_obs;
    A.pro = 1; A.epi = 1; A.bump = bump_counter ? 1 : 0;
    A.actions_in = actions_in; A.actions_out = b->actions; A.clip = clip;
    A.init_pos = init->initial_root_pos; A.init_rot = init->initial_root_rot; A.init_q = init->initial_dof_pos;
    A.rand_pos = nullptr; A.rand_vel = nullptr; A.seed = seed; A.counter = counter;
    A.counter_dev = reinterpret_cast<unsigned long long*>(b->reset_counter);
    A.env_id_offset = env_id_offset;
    A.obs = b->obs; A.rew = b->rew; A.pot = b->potentials; A.prev_pot = b->prev_potentials; A.up = b->up_vec;
    A.head = b->heading_vec; A.reset = reinterpret_cast<long long*>(b->reset_buf);
    A.progress = reinterpret_cast<long long*>(b->progress_buf); A.stats = b->episode_stats;
    A.ep_ret = b->episode_stats != nullptr ? b->episode_return : nullptr;
    if (rew_host != nullptr && reset_host != nullptr) { A.rew_host = rew_host; A.reset_host = reset_host; }
    const void* ptrs[] = {A.pos, A.quat, A.vel, A.q, A.qd, A.sens, A.actions_in, A.actions_out, A.init_pos, A.init_rot,
                          A.init_q, A.obs, A.rew, A.pot, A.prev_pot, A.up, A.head, A.reset, A.progress};
    A.use_bulk = 1;
    for (const void* p : ptrs) {
        if (p == nullptr) return fail(OIGE_E_NULL, "oige_env_step_f32: null device pointer");
        if (!al16(p)) A.use_bulk = 0;  // plain coalesced path for oddly offset views
    }
    if (reinterpret_cast<uintptr_t>(A.quat) & 15u) return fail(OIGE_E_ALIGN, "env_step: root_quat not 16B aligned");
    if (reinterpret_cast<uintptr_t>(A.init_rot) & 15u) return fail(OIGE_E_ALIGN, "env_step: initial_root_rot not 16B aligned");
    return role_dispatch(model, P, A, stream, dn);
}

// `decision_n`: env count the kernel choice is based on (the whole step's count when the caller
// processes the envs in chunks, so that chunked and one-shot steps run the SAME kernel and stay
// bit-identical); <= 0 means n.
int k2_step(const OigeModel* model, const OigeTaskParams* P, const OigeState* st, const float* actions, int64_t n,
            int32_t substeps, float dt, cudaStream_t stream, int64_t decision_n) {
    if (P->num_dof != model->num_dof)
        return fail(OIGE_E_SHAPE, "articulation_step: params.num_dof %d != model dofs %d", P->num_dof, model->num_dof);
    K2Args A;
    A.pos = st->root_pos; A.quat = st->root_quat; A.vel = st->root_vel; A.q = st->dof_pos; A.qd = st->dof_vel;
    A.sens = st->sensors; A.targets = st->dof_pos_targets; A.actions = actions; A.n = n; A.substeps = substeps; A.dt = dt;
    if (!A.pos || !A.quat || !A.vel || !A.q || !A.qd) return fail(OIGE_E_NULL, "articulation_step: null state pointer");
    if (reinterpret_cast<uintptr_t>(A.quat) & 15u) return fail(OIGE_E_ALIGN, "articulation_step: root_quat not 16B aligned");
    if (model->num_sensors > 0 && P->num_sensors != model->num_sensors)
        return fail(OIGE_E_SHAPE, "articulation_step: params.num_sensors %d != model %d", P->num_sensors, model->num_sensors);
    const int64_t dn = decision_n > 0 ? decision_n : n;
    if (k2_uses_role_kernel(model, dn) && role_variant_exists(model, A.targets != nullptr)) {
        FArgs F{};  // physics only: the same instantiation as the fused step with prologue / epilogue switched off
        F.pos = A.pos; F.quat = A.quat; F.vel = A.vel; F.q = A.q; F.qd = A.qd; F.sens = A.sens; F.actions = actions;
        F.targets = A.targets; F.n = n; F.substeps = substeps; F.dt = dt;
        F.D = model->num_dof; F.S = model->num_sensors; F.OBS = 12 + 3 * F.D + 6 * F.S;
        F.use_bulk = 1;
        const void* ptrs[] = {F.pos, F.quat, F.vel, F.q, F.qd, F.actions, F.sens};
        for (const void* p : ptrs)
            if (p != nullptr && !al16(p)) F.use_bulk = 0;
        return role_dispatch(model, P, F, stream, dn);
    }
    if (model->has_sc)
        return fail(OIGE_E_MODEL, "articulation_step: the model has self-collision pairs, which only the role-split kernel "
                                  "(K2-F) integrates, and that kernel does not serve this tree%s",
                    A.targets != nullptr ? " with position targets" : "");
    if (model->tpe_ok && use_thread_per_env(dn)) {
        const int bs = 128;
        const unsigned tb = (unsigned)((n + bs - 1) / bs);
        const DevModel* dm = reinterpret_cast<const DevModel*>(model->dev_table);
        // Up to two CTAs per SM cover the launch in one wave: the register budget per thread is then 255 instead of
        // 126 and ptxas schedules for latency (the regime where one warp per scheduler runs alone).  Same PTX; ptxas
        // schedules (and may fuse mul+add) differently, so the two variants agree to rounding, not bit for bit.
        // (decided on the WHOLE step's env count, like the kernel choice: chunked and one-shot steps must stay
        // bit-identical)
        const int mb = knobs().minb;
        const bool roomy = mb ? mb == 2 : (dn + bs - 1) / bs <= 2ll * sm_count();
        if (roomy) {
            if (model->num_links <= 12) articulation_step_tpe_kernel<12, 2><<<tb, bs, 0, stream>>>(dm, *P, A);
            else if (model->num_links <= 24) articulation_step_tpe_kernel<24, 2><<<tb, bs, 0, stream>>>(dm, *P, A);
            else articulation_step_tpe_kernel<32, 2><<<tb, bs, 0, stream>>>(dm, *P, A);
        } else {
            if (model->num_links <= 12) articulation_step_tpe_kernel<12, TPE_MINBLOCKS><<<tb, bs, 0, stream>>>(dm, *P, A);
            else if (model->num_links <= 24) articulation_step_tpe_kernel<24, TPE_MINBLOCKS><<<tb, bs, 0, stream>>>(dm, *P, A);
            else articulation_step_tpe_kernel<32, TPE_MINBLOCKS><<<tb, bs, 0, stream>>>(dm, *P, A);
        }
        count_launch();
        return check_launch("articulation_step_tpe_kernel");
    }
    const int pack = 32 / model->num_links;
    const long long warps = (n + pack - 1) / pack;
    const unsigned blocks = (unsigned)((warps + kWarpsPerBlock - 1) / kWarpsPerBlock);
    if (pack > 1)
        articulation_step_kernel<true><<<blocks, kWarpsPerBlock * 32, 0, stream>>>(
            reinterpret_cast<const DevModel*>(model->dev_table), *P, A);
    else
        articulation_step_kernel<false><<<blocks, kWarpsPerBlock * 32, 0, stream>>>(
            reinterpret_cast<const DevModel*>(model->dev_table), *P, A);
    count_launch();
    return check_launch("articulation_step_kernel");
}

#ifdef OIGE_STAMPS
// diagnostic build only (tools/k2f_stamps.py): the (tag, clock64) pairs of group 0 of CTA 0 of the last K2-F launch
extern "C" int oige_debug_read_stamps(long long* out) {
    return (int)cudaMemcpyFromSymbol(out, g_stamps, sizeof(long long) * 4 * 128 * 2);
}
extern "C" int oige_debug_clear_stamps(void) {
    static long long zero[4 * 128 * 2];
    return (int)cudaMemcpyToSymbol(g_stamps, zero, sizeof(zero));
}
#endif

uint64_t k2_nonfinite_count() {
    unsigned long long v = 0;
    if (cudaMemcpyFromSymbol(&v, g_nonfinite_envs, sizeof(v)) != cudaSuccess) { cudaGetLastError(); return 0; }
    return v;
}

OigeModel* k2_model_create(const OigeModelDesc* d) {
    const int L = d->num_links;
    if (L < 1 || L > OIGE_MAX_LINKS || d->num_spheres < 0 || d->num_spheres > OIGE_MAX_SPHERES ||
        d->num_sensors < 0 || d->num_sensors > OIGE_MAX_SENSORS) {
        set_error("oige_model_create: counts out of range (links %d spheres %d sensors %d)", L, d->num_spheres,
                  d->num_sensors);
        return nullptr;
    }
    DevModel* h = new DevModel();
    memset(h, 0, sizeof(DevModel));
    h->num_links = L; h->num_spheres = d->num_spheres; h->num_sensors = d->num_sensors;
    h->fixed_base = d->fixed_base != 0;
    for (int k = 0; k < kMaxChildren; ++k)
        for (int i = 0; i < 32; ++i) h->child[k][i] = -1;
    int nchild[32] = {0};
    bool ok = d->parent[0] == -1;
    for (int i = 0; i < L && ok; ++i) {
        h->parent[i] = d->parent[i];
        h->dof[i] = d->dof_index[i];
        h->jtype[i] = d->joint_type[i];
        h->sensor_of_link[i] = -1;
        if (i == 0) { h->level[0] = 0; continue; }
        const int p = d->parent[i];
        if (p < 0 || p >= i) { ok = false; break; }
        if (d->dof_index[i] < 0 || d->dof_index[i] >= L - 1) { ok = false; break; }
        if (d->joint_type[i] != OIGE_JOINT_HINGE && d->joint_type[i] != OIGE_JOINT_PRISMATIC) { ok = false; break; }
        h->level[i] = h->level[p] + 1;
        if (h->level[i] >= kMaxLevels || nchild[p] >= kMaxChildren) { ok = false; break; }
        h->child[nchild[p]++][p] = i;
        if (h->level[i] > h->max_depth) h->max_depth = h->level[i];
    }
    if (!ok) {
        set_error("oige_model_create: malformed tree (need parent[0]=-1, parent[i]<i, <=%d children, depth<%d)",
                  kMaxChildren, kMaxLevels);
        delete h;
        return nullptr;
    }
    for (int i = 0; i < L; ++i)
        if (nchild[i] > h->maxch_at_level[h->level[i]]) h->maxch_at_level[h->level[i]] = nchild[i];
    for (int i = 0; i < 32; ++i) h->anc[0][i] = (i < L && i > 0) ? h->parent[i] : (i < L ? 0 : i);
    for (int k = 1; k < kMaxJump; ++k)
        for (int i = 0; i < 32; ++i) h->anc[k][i] = i < L ? h->anc[k - 1][h->anc[k - 1][i]] : i;
    h->pack = 32 / L;  // environments carried side by side by one warp
    {   // heavy-path structure for the thread-per-env kernel
        int height[32] = {0}, heavy[32], nslots = 0, cnt = 0;
        for (int i = 0; i < 32; ++i) { heavy[i] = -1; h->light_slot[i] = -1; h->tflags[i] = 0; h->order2[i] = 0; }
        for (int i = L - 1; i >= 1; --i) {  // parent[i] < i: children are final when i is visited
            const int p = h->parent[i], cand = height[i] + 1;
            if (cand > height[p] || (cand == height[p] && i < heavy[p])) { height[p] = cand; heavy[p] = i; }
        }
        // iterative post-order, light children first (index order), heavy child last
        int stack[64], state[64], sp = 0;
        stack[0] = 0; state[0] = 0; sp = 1;
        while (sp > 0) {
            const int node = stack[sp - 1];
            int& st = state[sp - 1];
            int next = -1;
            for (; st < L; ++st) {  // light children in index order
                if (st > 0 && h->parent[st] == node && st != heavy[node]) { next = st++; break; }
            }
            if (next < 0 && st == L) {  // then the heavy child, once
                st = L + 1;
                if (heavy[node] >= 0) next = heavy[node];
            }
            if (next >= 0) { stack[sp] = next; state[sp] = 0; ++sp; continue; }
            if (node != 0) h->order2[cnt++] = node;
            --sp;
        }
        h->tpe_ok = (cnt == L - 1) ? 1 : 0;
        int seen_light[32] = {0};
        for (int k = 0; k < cnt; ++k) {
            const int i = h->order2[k], p = h->parent[i];
            if (heavy[i] >= 0) h->tflags[i] |= kHasHeavy;
            if (heavy[p] != i) {
                h->tflags[i] |= kIsHead;
                if (h->light_slot[p] < 0) {
                    if (nslots >= kMaxSlots) { h->tpe_ok = 0; break; }
                    h->light_slot[p] = nslots++;
                }
                if (!seen_light[p]) { h->tflags[i] |= kFirstLight; seen_light[p] = 1; }
            }
        }
        if (heavy[0] >= 0) h->tflags[0] |= kHasHeavy;
    }
    // ---- role split (K2-F): trunk = root + ancestors-or-self of every branching link; the rest are plain chains ----
    RoleModel* rm = new RoleModel();
    memset(rm, 0, sizeof(RoleModel));
    int role_ok = 0, role_nroles = 0, role_maxc = 0;
    {
        bool trunk[32] = {false};
        trunk[0] = true;
        for (int i = 0; i < L; ++i)
            if (nchild[i] >= 2)
                for (int a = i; a > 0 && !trunk[a]; a = h->parent[a]) trunk[a] = true;
        int tidx[32], nt = 0, nh = 0, heads[32], hlen[32], tip[32];
        bool fits = L <= kRL && d->num_spheres <= OIGE_MAX_SPHERES;
        const bool has_sc = d->num_collision_pairs > 0;
        for (int i = 0; i < 32; ++i) tidx[i] = -1;
        for (int i = 0; i < L && fits; ++i) {
            if (trunk[i]) {
                if (nt >= kRT) { fits = false; break; }
                // the trunk must be a PATH: link nt-1 is the parent of link nt
                if (nt > 0 && h->parent[i] != rm->tlink[nt - 1]) { fits = false; break; }
                tidx[i] = nt; rm->tlink[nt++] = (unsigned char)i;
            } else if (trunk[h->parent[i]]) {
                if (nh >= 32) { fits = false; break; }
                heads[nh++] = i;
            }
        }
        for (int c = 0; c < nh && fits; ++c) {  // chain below each head (no branching by construction)
            int len = 1, cur = heads[c];
            while (nchild[cur] == 1) { cur = h->child[0][cur]; ++len; }
            hlen[c] = len; tip[c] = cur;
        }
        // smallest number of roles that reaches the best critical path (longest-processing-time dealing)
        int best_r = 1, best_load = 1 << 30, owner[32], order[32];
        for (int c = 0; c < nh; ++c) order[c] = c;
        for (int a = 0; a < nh; ++a)
            for (int b = a + 1; b < nh; ++b)
                if (hlen[order[b]] > hlen[order[a]]) { int t = order[a]; order[a] = order[b]; order[b] = t; }
        for (int r = 1; r <= kMaxRoles && fits; ++r) {
            int load[kMaxRoles] = {0}, mx = 0;
            for (int a = 0; a < nh; ++a) {
                int w = 0;
                for (int q = 1; q < r; ++q) if (load[q] < load[w]) w = q;
                load[w] += hlen[order[a]];
            }
            for (int q = 0; q < r; ++q) if (load[q] > mx) mx = load[q];
            if (mx < best_load) { best_load = mx; best_r = r; }
        }
        int load[kMaxRoles] = {0};
        for (int a = 0; a < nh; ++a) {
            int w = 0;
            for (int q = 1; q < best_r; ++q) if (load[q] < load[w]) w = q;
            load[w] += hlen[order[a]];
            owner[order[a]] = w;
        }
        // instantiated combinations: 3 roles x <= 6 chain links (Humanoid), 4 roles x <= 3 (Ant, Anymal)
        const int maxc = (best_r == 3 && best_load <= 6) ? 6 : (best_r == 4 && best_load <= 3) ? 3 : 0;
        fits = fits && maxc > 0 && nh > 0;
        if (fits) {
            rm->num_links = L; rm->num_sensors = d->num_sensors; rm->fixed_base = d->fixed_base != 0; rm->nroles = best_r;
            rm->ntrunk = nt; rm->maxc = maxc;
            for (int i = 0; i < kRT; ++i) { rm->tpub[i] = 255; rm->town[i] = 0; }
            // hand-over slots: one per (role, attach trunk link)
            int slot_of[kMaxRoles][kRT];
            for (int r = 0; r < kMaxRoles; ++r) for (int t = 0; t < kRT; ++t) slot_of[r][t] = -1;
            int nslots = 0;
            for (int a = 0; a < nh && fits; ++a) {
                const int c = order[a], r = owner[c], t = tidx[h->parent[heads[c]]];
                if (slot_of[r][t] < 0) {
                    if (nslots >= kRSlots) { fits = false; break; }
                    slot_of[r][t] = nslots++;
                }
            }
            rm->nslots = nslots;
            int npub = 0;
            for (int t = 0; t < nt; ++t) {  // a' is published for the trunk links that chains attach to
                bool any = false;
                for (int r = 0; r < best_r; ++r) any = any || slot_of[r][t] >= 0;
                if (any) rm->tpub[t] = (unsigned char)npub++;
            }
            rm->npub = npub;
            // The OWN inertia / bias / contacts of a trunk link are taken off role 0's serial trunk sweep: they go to the
            // least loaded role that has a hand-over slot on the link (one link per role), or -- for a trunk link
            // without chains -- to a role without such a job, through an extra slot while shared memory allows
            // (4 groups of a 128-env CTA must fit in 227 KB).  Massless intermediate links have nothing to evaluate.
            auto is_massless = [&](int i) {
                bool zero = i > 0 && d->mass[i] == 0.f;
                for (int k = 0; k < 6; ++k) zero = zero && d->inertia[i][k] == 0.f;
                for (int s2 = 0; s2 < d->num_spheres; ++s2) zero = zero && d->sphere_link[s2] != i;
                for (int s2 = 0; s2 < d->num_sensors; ++s2) zero = zero && d->sensor_link[s2] != i;
                return zero;
            };
            const int fixed_floats = npub * 6 + nt * kKin + (nt - 1) * 4 + 2 + (has_sc ? best_r : 0) + best_r * maxc * kPu;
            const int max_group_floats = (int)((227 * 1024 - sizeof(RoleModel) - 256) / (4 * 32 * sizeof(float)));
            int own_t[kMaxRoles] = {-1, -1, -1, -1};
            for (int t = 0; t < nt; ++t) {
                if (is_massless(rm->tlink[t])) { rm->town[t] = 1; continue; }
                int pick = -1;
                for (int r = best_r - 1; r >= 0; --r)  // ties go to the higher role: role 0 carries the trunk sweep already
                    if (slot_of[r][t] >= 0 && own_t[r] < 0 && (pick < 0 || load[r] < load[pick])) pick = r;
                if (pick >= 0) { own_t[pick] = t; rm->town[t] = 1; }
            }
            for (int t = 0; t < nt; ++t) {
                if (rm->town[t]) continue;
                int pick = -1;
                for (int r = best_r - 1; r >= 1; --r)  // not role 0: the point is to shorten ITS critical path
                    if (own_t[r] < 0 && (pick < 0 || load[r] < load[pick])) pick = r;
                if (pick < 0 || nslots >= kRSlots || fixed_floats + (nslots + 1) * kHand > max_group_floats) continue;
                slot_of[pick][t] = nslots++;
                own_t[pick] = t; rm->town[t] = 1;
            }
            rm->nslots = nslots;
            int pos = 0;
            for (int t = 0; t < nt; ++t) {
                rm->ts_begin[t] = (unsigned char)pos;
                for (int r = 0; r < best_r; ++r)
                    if (slot_of[r][t] >= 0) rm->ts_slot[pos++] = (unsigned char)slot_of[r][t];
            }
            rm->ts_begin[nt] = (unsigned char)pos;
            // per-role programs.  Segments in execution order:
            //   role 0:  trunk kinematics | barrier K | [own inertia of a trunk link] chains ... | barrier H | trunk sweep
            //   role r:  barrier K | [own inertia of a trunk link] chains ... | barrier H
            int sensor_chains[kMaxRoles] = {0};
            int hb_roles[kRT] = {0, 0, 0, 0};
            bool slot_written[kRSlots] = {false};
            bool sensor_seen[OIGE_MAX_SENSORS] = {false};
            for (int r = 0; r < best_r && fits; ++r) {
                int k = 0, ns = 0;
                auto add_seg = [&](int kind, int first, int last, int flags, int att, int slot) {
                    if (ns >= kRSeg) { fits = false; return; }
                    rm->seg_kind[r][ns] = (unsigned char)kind; rm->seg_first[r][ns] = (unsigned char)first;
                    rm->seg_last[r][ns] = (unsigned char)last; rm->seg_flag[r][ns] = (unsigned char)flags;
                    rm->seg_att[r][ns] = (unsigned char)att; rm->seg_slot[r][ns] = (unsigned char)slot;
                    ++ns;
                };
                if (r == 0) add_seg(kSegTrunkKin, 1, nt - 1, kSegBarAfter, 0, 0);
                const int first_work = ns;
                // chains of this role by position (head first)
                int nchains = 0, ck0[8], ck1[8], cfl[8], catt[8], csl[8];
                for (int a = 0; a < nh && fits; ++a) {
                    const int c = order[a];
                    if (owner[c] != r) continue;
                    if (nchains >= 8) { fits = false; break; }
                    const int t = tidx[h->parent[heads[c]]];
                    int sens = -1;
                    for (int s2 = 0; s2 < d->num_sensors; ++s2) if (d->sensor_link[s2] == tip[c]) sens = s2;
                    if (sens >= 0) { ++sensor_chains[r]; sensor_seen[sens] = true; }
                    int links[32], n = 0;
                    for (int cur = tip[c];; cur = h->parent[cur]) { links[n++] = cur; if (cur == heads[c]) break; }
                    ck0[nchains] = k;
                    for (int j = n - 1; j >= 0; --j, ++k) rm->plink[r][k] = (unsigned char)links[j];  // head first
                    ck1[nchains] = k - 1; cfl[nchains] = sens >= 0 ? kSegSensor : 0; catt[nchains] = t;
                    csl[nchains] = slot_of[r][t];
                    ++nchains;
                }
                rm->npos[r] = (unsigned char)k;
                if (nchains == 0 || sensor_chains[r] > 1) fits = false;  // every role works; one sensor wrench per role
                if (!fits) break;
                if (has_sc) {
                    // self-collision program: every kinematics sweep first (capsule end points and link velocities published
                    // | barrier C), the pair phase (| barrier P), then the inertia sweeps
                    for (int c = 0; c < nchains; ++c) add_seg(kSegChain, ck0[c], ck1[c], cfl[c] | kSegKinOnly, catt[c], csl[c]);
                    if (fits) rm->seg_flag[r][ns - 1] |= kSegBarAfter;
                    add_seg(kSegPairs, 0, 0, kSegBarAfter, 0, 0);
                }
                const int first_up = ns;
                if (own_t[r] >= 0) {
                    const int sl = slot_of[r][own_t[r]];
                    add_seg(kSegOwn, own_t[r], own_t[r], slot_written[sl] ? 0 : kSegSlotFirst, own_t[r], sl);
                    slot_written[sl] = true;
                }
                for (int c = 0; c < nchains; ++c) {
                    const int fl = cfl[c] | (slot_written[csl[c]] ? 0 : kSegSlotFirst) | (has_sc ? kSegUpOnly : 0);
                    slot_written[csl[c]] = true;
                    add_seg(kSegChain, ck0[c], ck1[c], fl, catt[c], csl[c]);
                }
                if (!fits) break;
                (void)first_up;
                if (r == 0) {
                    add_seg(kSegTrunkUp, 0, nt - 1, 0, 0, 0);  // (waits per trunk link, see hb_id)
                } else {
                    rm->seg_flag[r][first_work] |= kSegBarBefore;
                    // after its last sweep into a slot on trunk link t the role ARRIVES at t's hand-over barrier
                    for (int t = 0; t < nt; ++t) {
                        int lastseg = -1;
                        for (int q = first_up; q < ns; ++q)
                            if ((rm->seg_kind[r][q] == kSegChain || rm->seg_kind[r][q] == kSegOwn) && rm->seg_att[r][q] == t) lastseg = q;
                        if (lastseg >= 0) { rm->seg_flag[r][lastseg] |= kSegArrive; hb_roles[t] |= 1 << r; }
                    }
                }
                rm->nseg[r] = (unsigned char)ns;
            }
            {   // hand-over barriers: one per trunk link that other roles hand over to (two ids per group)
                int nb = 0;
                for (int t = 0; t < nt && fits; ++t) {
                    if (!hb_roles[t]) continue;
                    if (++nb > 2) { fits = false; break; }
                    int cnt = 1;
                    for (int r = 1; r < best_r; ++r) cnt += (hb_roles[t] >> r) & 1;
                    rm->hb_id[t] = (unsigned char)nb; rm->hb_cnt[t] = (unsigned char)(32 * cnt);
                }
            }
            for (int s2 = 0; s2 < d->num_sensors; ++s2) fits = fits && sensor_seen[s2];  // sensors sit on chain tips
            // scratch layout of a 32-env group (float offsets, x32 lanes)
            rm->oI = 0;
            // (with self-collision the hand-over region also holds the capsule end points between the kinematics and the
            // inertia sweeps: 6 floats per capsule)
            int ncaps = 0;
            for (int i = 0; i < L; ++i) if (d->capsule_radius[i] > 0.f) ++ncaps;
            const int handover_floats = has_sc && 6 * ncaps > nslots * kHand ? 6 * ncaps : nslots * kHand;
            rm->oA = rm->oI + handover_floats;
            rm->oK = rm->oA + npub * 6;
            rm->oT = rm->oK + nt * kKin;
            rm->oS = rm->oT + (nt - 1) * 4;
            rm->oM = rm->oS + (best_r == 3 ? 0 : d->num_sensors * 7);  // the 3-role variants keep the sensor frame in registers
            rm->oH = rm->oM + 2;                       // one hit word per role (self-collision)
            rm->oL = rm->oH + (has_sc ? best_r : 0);
            rm->group_floats = rm->oL + best_r * maxc * kPu;
            if (getenv("OIGE_DUMP_ROLES") != nullptr) {  // the trunk and the roles' segment programs (tools/dump_roles.py)
                fprintf(stderr, "[oige] trunk:");
                for (int t = 0; t < nt; ++t)
                    fprintf(stderr, " t%d=link%d(own %s)", t, rm->tlink[t], rm->town[t] ? "handed over" : "role 0");
                fprintf(stderr, "; %d hand-over slots, group scratch %d floats per lane (max %d)\n", rm->nslots, rm->group_floats,
                        max_group_floats);
                for (int r = 0; r < best_r; ++r) {
                    fprintf(stderr, "[oige] program of role %d:", r);
                    for (int q = 0; q < rm->nseg[r]; ++q)
                        fprintf(stderr, " [kind %d links %d..%d flags 0x%x att t%d slot %d]", rm->seg_kind[r][q], rm->seg_first[r][q],
                                rm->seg_last[r][q], rm->seg_flag[r][q], rm->seg_att[r][q], rm->seg_slot[r][q]);
                    fprintf(stderr, "\n");
                }
            }
            // ---- self-collision tables ----
            rm->has_sc = has_sc ? 1 : 0;
            if (has_sc && fits) {
                int owner_of_link[32], sphn[32];
                for (int i = 0; i < L; ++i) { owner_of_link[i] = tidx[i] >= 0 ? 0 : -1; sphn[i] = 0; }
                for (int i = 0; i < L; ++i) {
                    if (tidx[i] >= 0) { rm->w_idx[i] = (unsigned char)tidx[i]; rm->rec_off[i] = (unsigned short)(rm->oK + tidx[i] * kKin); }
                }
                // a trunk link's pairs (and wrench) belong to the role that evaluates its own inertia
                for (int r = 0; r < best_r; ++r)
                    if (own_t[r] >= 0) owner_of_link[rm->tlink[own_t[r]]] = r;
                for (int r = 0; r < best_r; ++r)
                    for (int k2 = 0; k2 < rm->npos[r]; ++k2) {
                        const int li = rm->plink[r][k2];
                        owner_of_link[li] = r;
                        rm->w_idx[li] = (unsigned char)(kRT + k2);
                        rm->rec_off[li] = (unsigned short)(rm->oL + (r * maxc + k2) * kPu);
                    }
                // one capsule per link that takes part: slot in the published end points, bounding radius about its midpoint
                int nc2 = 0;
                for (int i = 0; i < L; ++i) {
                    sphn[i] = d->capsule_radius[i] > 0.f ? 1 : 0;  // "has a self-collision shape"
                    if (!sphn[i]) continue;
                    rm->cap_idx[i] = (unsigned char)nc2++;
                    float hl = 0.f;
                    for (int q = 0; q < 3; ++q) {
                        rm->cap_a[q][i] = d->capsule_a[i][q]; rm->cap_b[q][i] = d->capsule_b[i][q];
                        hl += (d->capsule_b[i][q] - d->capsule_a[i][q]) * (d->capsule_b[i][q] - d->capsule_a[i][q]);
                    }
                    rm->cap_r[i] = d->capsule_radius[i];
                    rm->cap_il2[i] = hl > 1e-12f ? 1.f / hl : 0.f;
                    rm->brad[i] = (0.5f * sqrtf(hl) + d->capsule_radius[i]) * 1.0001f + 1e-6f;  // (never cuts a touching pair)
                }
                // Every pair goes into the list of ONE role: the owner of both links, else one of the two owners, dealt so
                // that the roles' expected NARROW-phase work is even (the phases are separated by group barriers, the
                // slowest role sets the pace).  A pair whose capsules are near each other at the rest pose (q = 0) is a
                // narrow-phase candidate most of the time (the two thighs, a thigh and the lower waist, ...) and weighs
                // kNearWeight, any other pair kFarWeight (its broad-phase test plus an occasional narrow test).  A list
                // holds at most 32 entries (one candidate bit each).  Lists are sorted by the first link.
                double Rw[32][9], pw[32][3];
                for (int i = 0; i < L; ++i) {
                    double Rl[9];
                    {
                        const double w = i == 0 ? 1.0 : d->joint_quat[i][0], x = i == 0 ? 0.0 : d->joint_quat[i][1];
                        const double y = i == 0 ? 0.0 : d->joint_quat[i][2], z = i == 0 ? 0.0 : d->joint_quat[i][3];
                        const double m[9] = {1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y),
                                             2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x),
                                             2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)};
                        memcpy(Rl, m, sizeof(m));
                    }
                    if (i == 0) {
                        memcpy(Rw[0], Rl, sizeof(Rl));
                        pw[0][0] = pw[0][1] = pw[0][2] = 0.0;
                        continue;
                    }
                    const int pa = h->parent[i];
                    for (int r2 = 0; r2 < 3; ++r2) {
                        pw[i][r2] = pw[pa][r2];
                        for (int c2 = 0; c2 < 3; ++c2) {
                            pw[i][r2] += Rw[pa][3 * r2 + c2] * d->joint_pos[i][c2];
                            Rw[i][3 * r2 + c2] = 0.0;
                            for (int k2 = 0; k2 < 3; ++k2) Rw[i][3 * r2 + c2] += Rw[pa][3 * r2 + k2] * Rl[3 * k2 + c2];
                        }
                    }
                }
                auto rest_gap = [&](int la, int lb) {  // distance between the two capsule surfaces at the rest pose (sampled)
                    double best = 1e30;
                    for (int ia = 0; ia <= 8; ++ia)
                        for (int ib = 0; ib <= 8; ++ib) {
                            double dd = 0.0;
                            for (int r2 = 0; r2 < 3; ++r2) {
                                double xa = pw[la][r2], xb = pw[lb][r2];
                                for (int c2 = 0; c2 < 3; ++c2) {
                                    xa += Rw[la][3 * r2 + c2] * (d->capsule_a[la][c2] + (d->capsule_b[la][c2] - d->capsule_a[la][c2]) * ia / 8.0);
                                    xb += Rw[lb][3 * r2 + c2] * (d->capsule_a[lb][c2] + (d->capsule_b[lb][c2] - d->capsule_a[lb][c2]) * ib / 8.0);
                                }
                                dd += (xa - xb) * (xa - xb);
                            }
                            if (dd < best) best = dd;
                        }
                    return sqrt(best) - d->capsule_radius[la] - d->capsule_radius[lb];
                };
                const double kNearGap = 0.12, kNearWeight = 9.0, kFarWeight = 2.0;
                int pair_role[OIGE_MAX_PAIRS], sc_count[kMaxRoles] = {0, 0, 0, 0};
                double sc_load[kMaxRoles] = {0.0, 0.0, 0.0, 0.0};
                double pair_w[OIGE_MAX_PAIRS];
                for (int k2 = 0; k2 < d->num_collision_pairs && fits; ++k2) {
                    const int la = d->pair_link_a[k2], lb = d->pair_link_b[k2];
                    if (la < 0 || la >= L || lb < 0 || lb >= L || la == lb || sphn[la] == 0 || sphn[lb] == 0 ||
                        owner_of_link[la] < 0 || owner_of_link[lb] < 0) { fits = false; break; }
                    pair_w[k2] = rest_gap(la, lb) < kNearGap ? kNearWeight : kFarWeight;
                    pair_role[k2] = -1;
                    if (owner_of_link[la] == owner_of_link[lb]) {
                        pair_role[k2] = owner_of_link[la]; sc_load[pair_role[k2]] += pair_w[k2]; ++sc_count[pair_role[k2]];
                    }
                }
                for (int pass = 0; pass < 2 && fits; ++pass)  // heavy pairs first, each to the lighter of its two owners
                    for (int k2 = 0; k2 < d->num_collision_pairs; ++k2) {
                        if (pair_role[k2] >= 0 || (pass == 0) != (pair_w[k2] > kFarWeight)) continue;
                        const int ra = owner_of_link[d->pair_link_a[k2]], rb = owner_of_link[d->pair_link_b[k2]];
                        int pick = sc_load[ra] < sc_load[rb] || (sc_load[ra] == sc_load[rb] && ra > rb) ? ra : rb;
                        if (sc_count[pick] >= 32) pick = pick == ra ? rb : ra;
                        pair_role[k2] = pick;
                        sc_load[pick] += pair_w[k2]; ++sc_count[pick];
                    }
                int npairs = 0;
                for (int r = 0; r < best_r && fits; ++r) {
                    rm->sc_off[r] = (unsigned char)npairs;
                    for (int a = 0; a < L && fits; ++a) {
                        if (owner_of_link[a] != r) continue;
                        for (int k2 = 0; k2 < d->num_collision_pairs; ++k2) {
                            int la = d->pair_link_a[k2], lb = d->pair_link_b[k2];
                            if (pair_role[k2] != r) continue;
                            const bool both = owner_of_link[la] == r && owner_of_link[lb] == r;
                            if (both && lb < la) { const int t2 = la; la = lb; lb = t2; }  // first link = the lower one
                            if (!both && owner_of_link[la] != r) { const int t2 = la; la = lb; lb = t2; }
                            if (la != a) continue;
                            if (npairs >= kSCMax) { fits = false; break; }
                            rm->sc_a[npairs] = (unsigned char)la; rm->sc_b[npairs] = (unsigned char)(lb | (both ? 128 : 0));
                            ++npairs;
                        }
                    }
                }
                for (int r = best_r; r < 8; ++r) rm->sc_off[r] = (unsigned char)npairs;
                for (int r = 0; r < best_r; ++r) fits = fits && rm->sc_off[r + 1] - rm->sc_off[r] <= 32;  // candidate bit mask
                fits = fits && npairs == d->num_collision_pairs && best_r <= 4;
                for (int r = 0; r < best_r && fits; ++r)
                    for (int k2 = rm->sc_off[r]; k2 < rm->sc_off[r + 1]; ++k2) {
                        const int ob = owner_of_link[rm->sc_b[k2] & 127];
                        if (ob != r) rm->sc_to[r][ob] |= 1u << (k2 - rm->sc_off[r]);
                    }
                if (getenv("OIGE_DUMP_ROLES") != nullptr) {
                    for (int r = 0; r < best_r; ++r) {
                        fprintf(stderr, "[oige] self-collision list of role %d (load %.0f):", r, sc_load[r]);
                        for (int k2 = rm->sc_off[r]; k2 < rm->sc_off[r + 1]; ++k2)
                            fprintf(stderr, " %d-%d%s", rm->sc_a[k2], rm->sc_b[k2] & 127, (rm->sc_b[k2] & 128) ? "*" : "");
                        fprintf(stderr, "\n");
                    }
                }
                // Bounding centres (3 floats per link with spheres) live in memory that is dead between the kinematics and
                // the inertia sweeps: floats 7..12 of the records of the role's own massless intermediate links (two
                // centres each), then the published-acceleration rows and the tail of the sphere-centre region.
                {
                    int pool[64], npool = 0, used = 0;
                    auto is_massless2 = [&](int i) {
                        bool zero = i > 0 && d->mass[i] == 0.f && sphn[i] == 0;
                        for (int s2 = 0; s2 < d->num_spheres; ++s2) zero = zero && d->sphere_link[s2] != i;
                        for (int k2 = 0; k2 < 6; ++k2) zero = zero && d->inertia[i][k2] == 0.f;
                        for (int s2 = 0; s2 < d->num_sensors; ++s2) zero = zero && d->sensor_link[s2] != i;
                        return zero;
                    };
                    for (int r = 0; r < best_r; ++r)
                        for (int k2 = 0; k2 < rm->npos[r]; ++k2)
                            if (is_massless2(rm->plink[r][k2]) && npool + 2 <= 64) {
                                const int base = rm->oL + (r * maxc + k2) * kPu;
                                pool[npool++] = base + 7; pool[npool++] = base + 10;
                            }
                    for (int q = 0; q + 3 <= npub * 6 && npool < 64; q += 3) pool[npool++] = rm->oA + q;
                    for (int q = 6 * ncaps; q + 3 <= handover_floats && npool < 64; q += 3) pool[npool++] = rm->oI + q;
                    for (int i = 0; i < L; ++i) {
                        if (sphn[i] == 0) continue;
                        if (used >= npool) { fits = false; break; }
                        rm->ctr_off[i] = (unsigned short)pool[used++];
                    }
                }
                for (int k2 = 0; k2 < npairs && fits; ++k2) {
                    const int la = rm->sc_a[k2], lb = rm->sc_b[k2] & 127;
                    const float rr = rm->brad[la] + rm->brad[lb], rr2 = rr * rr;
                    unsigned bits;
                    memcpy(&bits, &rr2, sizeof(bits));
                    const unsigned oca = (unsigned)rm->ctr_off[la] * 128u, ocb = (unsigned)rm->ctr_off[lb] * 128u;
                    if (oca > 0xffffu || ocb > 0xffffu) { fits = false; break; }
                    rm->sc_brd[k2].x = oca << 16 | ocb; rm->sc_brd[k2].y = bits;
                    const int lo = la < lb ? la : lb, hi = la < lb ? lb : la;
                    const unsigned olo = (unsigned)rm->cap_idx[lo] * 6u * 128u, ohi = (unsigned)rm->cap_idx[hi] * 6u * 128u;
                    if (olo > 0xffffu || ohi > 0xffffu) { fits = false; break; }
                    const float cr = rm->cap_r[la] + rm->cap_r[lb], cr2 = cr * cr;
                    rm->sc_nar[k2].x = olo << 16 | ohi;
                    memcpy(&rm->sc_nar[k2].y, &rm->cap_il2[lo], 4);
                    memcpy(&rm->sc_nar[k2].z, &rm->cap_il2[hi], 4);
                    memcpy(&rm->sc_nar[k2].w, &cr2, 4);
                }
                // the capsule end points live in the hand-over slots' memory during the kinematics / pair phases
                if (kRT + maxc > 32 || rm->group_floats > max_group_floats) fits = false;
            }
        }
        role_ok = (fits && h->tpe_ok) ? 1 : 0;
        role_nroles = best_r; role_maxc = maxc;
    }
    h->jump_rounds = 0;
    while ((1 << h->jump_rounds) < h->max_depth) ++h->jump_rounds;
    for (int i = 0; i < L; ++i) {
        for (int k = 0; k < 3; ++k) { h->jpos[k][i] = d->joint_pos[i][k]; h->axis[k][i] = d->joint_axis[i][k]; h->com[k][i] = d->com[i][k]; }
        for (int k = 0; k < 4; ++k) h->jquat[k][i] = d->joint_quat[i][k];
        for (int k = 0; k < 6; ++k) h->inertia[k][i] = d->inertia[i][k];
        h->mass[i] = d->mass[i]; h->armature[i] = d->armature[i]; h->damping[i] = d->damping[i];
        h->stiffness[i] = d->stiffness[i]; h->sref[i] = d->spring_ref[i]; h->lo[i] = d->limit_lower[i]; h->hi[i] = d->limit_upper[i];
        h->drive_kp[i] = d->drive_stiffness[i]; h->drive_kd[i] = d->drive_damping[i]; h->drive_fmax[i] = d->drive_max_force[i];
    }
    h->jquat[0][0] = 1.f;
    // spheres sorted by link (stable)
    int pos = 0;
    for (int i = 0; i < L; ++i) {
        h->sph_start[i] = pos;
        for (int s = 0; s < d->num_spheres; ++s) {
            if (d->sphere_link[s] != i) continue;
            for (int k = 0; k < 3; ++k) h->sph_pos[k][pos] = d->sphere_pos[s][k];
            h->sph_r[pos] = d->sphere_radius[s];
            ++pos;
        }
        h->sph_count[i] = pos - h->sph_start[i];
        if (h->sph_count[i] > h->max_sph_per_link) h->max_sph_per_link = h->sph_count[i];
    }
    if (pos != d->num_spheres) {
        set_error("oige_model_create: sphere_link out of range");
        delete h;
        delete rm;
        return nullptr;
    }
    for (int s = 0; s < d->num_sensors; ++s) {
        if (d->sensor_link[s] < 0 || d->sensor_link[s] >= L) {
            set_error("oige_model_create: sensor_link out of range");
            delete h;
            delete rm;
            return nullptr;
        }
        h->sensor_of_link[d->sensor_link[s]] = s;
    }
    for (int k = 0; k < 3; ++k) h->gravity[k] = d->gravity[k];
    h->k_c = d->contact_stiffness; h->d_c = d->contact_damping; h->mu = d->friction_mu; h->c_visc = d->friction_viscous;
    h->k_lim = d->limit_stiffness; h->d_lim = d->limit_damping; h->ang_damp = d->angular_damping;
    h->max_w = d->max_angular_velocity > 0.f ? d->max_angular_velocity : 1e30f;
    h->max_qd_lin = d->max_joint_velocity > 0.f ? d->max_joint_velocity : 1e30f;
    h->max_qd = fminf(h->max_qd_lin, h->max_w);
    h->pen_cap_v = d->max_depenetration_velocity > 0.f ? d->max_depenetration_velocity : 1e30f;

    const int tpe_ok = h->tpe_ok;
    void* dev_role = nullptr;
    if (role_ok) {  // compact copy of the tables for K2-F
        for (int k = 0; k < 4; ++k) rm->gravity[k] = h->gravity[k];
        rm->k_c = h->k_c; rm->d_c = h->d_c; rm->mu = h->mu; rm->c_visc = h->c_visc; rm->k_lim = h->k_lim;
        rm->d_lim = h->d_lim; rm->max_qd = h->max_qd; rm->ang_damp = h->ang_damp;
        rm->max_qd_lin = h->max_qd_lin; rm->max_w = h->max_w; rm->pen_cap_v = h->pen_cap_v;
        rm->k_self = d->self_contact_stiffness; rm->d_self = d->self_contact_damping;
        for (int i = 0; i < L; ++i) {
            for (int k = 0; k < 3; ++k) { rm->jpos[k][i] = h->jpos[k][i]; rm->axis[k][i] = h->axis[k][i]; rm->com[k][i] = h->com[k][i]; }
            for (int k = 0; k < 4; ++k) rm->jquat[k][i] = h->jquat[k][i];
            for (int k = 0; k < 6; ++k) rm->inertia[k][i] = h->inertia[k][i];
            rm->mass[i] = h->mass[i]; rm->armature[i] = h->armature[i]; rm->damping[i] = h->damping[i];
            rm->stiffness[i] = h->stiffness[i]; rm->sref[i] = h->sref[i]; rm->lo[i] = h->lo[i]; rm->hi[i] = h->hi[i];
            rm->drive_kp[i] = h->drive_kp[i]; rm->drive_kd[i] = h->drive_kd[i]; rm->drive_fmax[i] = h->drive_fmax[i];
            rm->dof[i] = (unsigned char)(i > 0 ? h->dof[i] : 0); rm->jtype[i] = (unsigned char)h->jtype[i];
            rm->sph_start[i] = (unsigned char)h->sph_start[i]; rm->sph_count[i] = (unsigned char)h->sph_count[i];
            rm->sensor_of_link[i] = (signed char)h->sensor_of_link[i];
            // intermediate link of a multi-dof joint: no mass, no spheres, no sensor -> its own inertia is skipped
            bool zero = i > 0 && d->mass[i] == 0.f && h->sph_count[i] == 0 && h->sensor_of_link[i] < 0;
            for (int k = 0; k < 6; ++k) zero = zero && d->inertia[i][k] == 0.f;
            rm->massless[i] = zero ? 1 : 0;
        }
        for (int s2 = 0; s2 < d->num_spheres; ++s2) {
            for (int k = 0; k < 3; ++k) rm->sph_pos[k][s2] = h->sph_pos[k][s2];
            rm->sph_r[s2] = h->sph_r[s2];
        }
    }
    const int group_floats = rm->group_floats;
    void* dev = nullptr;
    cudaError_t e = cudaMalloc(&dev, sizeof(DevModel));
    if (e == cudaSuccess) e = cudaMemcpy(dev, h, sizeof(DevModel), cudaMemcpyHostToDevice);
    if (e == cudaSuccess && role_ok) {
        e = cudaMalloc(&dev_role, sizeof(RoleModel));
        if (e == cudaSuccess) e = cudaMemcpy(dev_role, rm, sizeof(RoleModel), cudaMemcpyHostToDevice);
    }
    delete h;
    delete rm;
    if (e != cudaSuccess) {
        set_error("oige_model_create: %s", cudaGetErrorString(e));
        if (dev) cudaFree(dev);
        if (dev_role) cudaFree(dev_role);
        return nullptr;
    }
    OigeModel* m = new OigeModel();
    m->dev_table = dev; m->dev_role = dev_role; m->num_links = L; m->num_dof = L - 1; m->num_sensors = d->num_sensors;
    m->tpe_ok = tpe_ok; m->role_ok = role_ok; m->nroles = role_nroles; m->maxc = role_maxc; m->group_floats = group_floats;
    m->has_sc = d->num_collision_pairs > 0 ? 1 : 0;
    return m;
}

void k2_model_destroy(OigeModel* m) {
    if (m->dev_table) cudaFree(m->dev_table);
    if (m->dev_role) cudaFree(m->dev_role);
    delete m;
}

}  // namespace oige
