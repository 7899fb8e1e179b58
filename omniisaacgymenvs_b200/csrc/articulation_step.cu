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
//   articulation_step_role_kernel one CTA per 32 envs, lane = env, warp = ROLE (trunk + limb chains of the tree on
//                                 separate warps, shared-memory hand-over, named barriers; articulation_role.cuh):
//                                 about half the dependency chain per warp -- used while the launch fits one wave.
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
#include <stdlib.h>

#include "oige_common.cuh"

struct OigeModel {
    void* dev_table;      // DevModel in device memory
    int num_links;
    int num_dof;
    int num_sensors;
    int tpe_ok;
    int role_ok;      // K2-R tables valid (>= 2 roles)
    int nroles;
    int role_smem;    // dynamic shared memory of the role-split kernel
};

namespace oige {
namespace {

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
constexpr int kMaxRoles = 4, kMaxTrunk = 8, kMaxHeads = 8;
constexpr int kRTip = 1, kRHead = 2, kRTrunk = 4, kRMassless = 8;

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
    // role-split kernel (K2-R): trunk = root + every link with a branching descendant-or-self; what hangs off the
    // trunk are plain chains, dealt to `nroles` warps; role 0 also owns the trunk
    int nroles, ntrunk, nheads, role_ok;
    int rflags[32];                  // kRTip | kRHead | kRTrunk
    int trunk_idx[32];               // index among the trunk links (root = 0), else -1
    int head_slot[32];               // hand-over slot of a chain head, else -1
    int trunk_link[kMaxTrunk];       // trunk links in increasing index order
    int th_begin[kMaxTrunk + 4];     // heads attached to trunk link t: th_heads[th_begin[t] .. th_begin[t+1])
    int th_heads[kMaxHeads];
    int n1[kMaxRoles], n2[kMaxRoles], sync2[kMaxRoles], sync3[kMaxRoles];
    int list1[kMaxRoles][32];        // kinematics order: trunk (non-root) + own links, increasing index
    int list2[kMaxRoles][32];        // inertia order: own chains tip->head, then (role 0) trunk downwards, root last
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
__device__ __forceinline__ bool contact_eval(const DevModel& M, int s, const M3& R, V3 p, const V6& v, const V6& b,
                                             float root_z, float dt, Contact& c) {
    const V3 local = {M.sph_pos[0][s], M.sph_pos[1][s], M.sph_pos[2][s]};
    const float rad = M.sph_r[s];
    const V3 x = p + mulv(R, local);
    const float pen = rad - (root_z + x.z);
    if (!(pen > 0.f)) return false;
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
__device__ __forceinline__ void link_inertia_bias(const DevModel& M, int li, const M3& R, V3 p, const V6& v, V3 grav,
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
}

// Adds the linearly-implicit ground contacts of link li: P^T C P to the inertia, -wrench(F0) to the bias.
// `nloop` = loop bound (uniform across a warp-per-env warp: the model's max spheres per link).
__device__ __forceinline__ void link_contacts(const DevModel& M, int li, const M3& R, V3 p, const V6& v, const V6& b,
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
__device__ __forceinline__ void joint_force(const DevModel& M, int li, float q, float qd, float effort, float dt,
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
__device__ __forceinline__ void sensor_wrench(const DevModel& M, int li, const M3& R, V3 p, const V6& v, const V6& b,
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
__device__ __forceinline__ void integrate_root(const DevModel& M, const V6& a, float dt, V3& rpos, Q4& rquat, V6& rvel) {
    const V3 w0 = ang(rvel), v0 = lin(rvel);
    const V3 vdot = lin(a) + cross(w0, v0);  // classical acceleration of the root origin
    float damp = 1.f - dt * M.ang_damp;
    damp = damp < 0.f ? 0.f : damp;
    const V3 w1 = damp * (w0 + dt * ang(a));
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
                sincosf(prismatic ? 0.f : 0.5f * q, &sh, &ch);
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
            qd = fminf(fmaxf(qd, -M.max_qd), M.max_qd);
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
            sincosf(prismatic ? 0.f : 0.5f * qv[i], &sh, &ch);
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
            qd = fminf(fmaxf(qd, -M.max_qd), M.max_qd);
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

#include "articulation_role.cuh"  // K2-R: role-split kernel (same device functions)

}  // namespace

// Kernel choice: OIGE_K2_MODE=warp|thread forces one; otherwise thread-per-env from kTpeMinEnvs envs on.
static bool use_thread_per_env(int64_t n) {
    const char* e = getenv("OIGE_K2_MODE");  // read per call so tests can switch it
    if (e && e[0] == 'w') return false;
    if (e && e[0] == 't') return true;
    return n >= kTpeMinEnvs;
}

// Role-split kernel (K2-R): between the warp-per-env kernel's range and the point where one thread per env fills the
// machine by itself.  OIGE_K2_MODE=role forces it; warp|thread exclude it.
static bool use_role_split(int64_t n, int num_links) {
    const char* e = getenv("OIGE_K2_MODE");
    if (e && e[0] == 'r') return true;
    if (e && (e[0] == 'w' || e[0] == 't')) return false;
    // upper end: TWO waves at 4 CTAs (of 32 envs) per SM -- 37888 envs on a 148-SM B200.  Humanoid, K2-R vs K2-T:
    // 16384: 82 vs 110 us, 20480 (2nd wave starts): 132 vs 136, 36864: 152 vs 183, 40960 (3rd wave): 202 vs 210,
    // 57344: 253 vs 241 -> K2-T from there (profiles/r1_k2_modes.txt)
    const long long two_waves = 2ll * 4ll * sm_count() * 32;
    return n >= (num_links <= 12 ? kRoleMinEnvsSmall : kRoleMinEnvs) && n <= two_waves;
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
    if (model->role_ok && use_role_split(decision_n > 0 ? decision_n : n, model->num_links)) {
        const unsigned tb = (unsigned)((n + 31) / 32);
        const DevModel* dm = reinterpret_cast<const DevModel*>(model->dev_table);
        const int bs = model->nroles * 32;
        // One wave with up to two CTAs per SM: 255-register budget (latency schedule); up to four: a budget sized for
        // the model's warp count.  Decided on the WHOLE step's env count (chunked == one-shot stays bit-identical).
        const long long dn = decision_n > 0 ? decision_n : n;
        const char* mb = getenv("OIGE_K2_MINB");
        const bool roomy = mb ? mb[0] == '2' : (dn + 31) / 32 <= 2ll * sm_count();
        const int sm = model->role_smem;
#define OIGE_ROLE_LAUNCH(MAXL)                                                                                  \
    do {                                                                                                        \
        if (roomy) launch_pdl(articulation_step_role_kernel<MAXL, kMaxRoles, 2>, tb, bs, sm, stream, dm, *P, A);        \
        else if (model->nroles == 2) launch_pdl(articulation_step_role_kernel<MAXL, 2, 4>, tb, bs, sm, stream, dm, *P, A); \
        else if (model->nroles == 3) launch_pdl(articulation_step_role_kernel<MAXL, 3, 4>, tb, bs, sm, stream, dm, *P, A); \
        else launch_pdl(articulation_step_role_kernel<MAXL, 4, 4>, tb, bs, sm, stream, dm, *P, A);                      \
    } while (0)
        if (model->num_links <= 12) OIGE_ROLE_LAUNCH(12);
        else OIGE_ROLE_LAUNCH(24);
#undef OIGE_ROLE_LAUNCH
        count_launch();
        return check_launch("articulation_step_role_kernel");
    }
    if (model->tpe_ok && use_thread_per_env(decision_n > 0 ? decision_n : n)) {
        const int bs = 128;
        const unsigned tb = (unsigned)((n + bs - 1) / bs);
        const DevModel* dm = reinterpret_cast<const DevModel*>(model->dev_table);
        // Up to two CTAs per SM cover the launch in one wave: the register budget per thread is then 255 instead of
        // 126 and ptxas schedules for latency (the regime where one warp per scheduler runs alone).  Same PTX; ptxas
        // schedules (and may fuse mul+add) differently, so the two variants agree to rounding, not bit for bit.
        const char* mb = getenv("OIGE_K2_MINB");
        // (decided on the WHOLE step's env count, like the kernel choice: ptxas may contract mul+add differently in
        // the two instantiations, and chunked and one-shot steps must stay bit-identical)
        const long long dn = decision_n > 0 ? decision_n : n;
        const bool roomy = mb ? mb[0] == '2' : (dn + bs - 1) / bs <= 2ll * sm_count();
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
    {   // role split (K2-R): trunk = root + ancestors-or-self of every branching link; the rest are chains
        bool trunk[32] = {false};
        trunk[0] = true;
        for (int i = 0; i < L; ++i)
            if (nchild[i] >= 2)
                for (int a = i; a > 0 && !trunk[a]; a = h->parent[a]) trunk[a] = true;
        for (int i = 0; i < 32; ++i) { h->trunk_idx[i] = -1; h->head_slot[i] = -1; h->rflags[i] = 0; }
        int nt = 0, nh = 0, heads[32], hlen[32], tip[32];
        bool fits = true;
        for (int i = 0; i < L; ++i) {
            if (trunk[i]) {
                if (nt >= kMaxTrunk) { fits = false; break; }
                h->trunk_idx[i] = nt; h->trunk_link[nt++] = i; h->rflags[i] |= kRTrunk;
            } else if (trunk[h->parent[i]]) {
                if (nh >= kMaxHeads) { fits = false; break; }
                heads[nh++] = i;
            }
        }
        for (int c = 0; c < nh && fits; ++c) {  // chain below each head (no branching by construction)
            int len = 1, cur = heads[c];
            while (nchild[cur] == 1) { cur = h->child[0][cur]; ++len; }
            hlen[c] = len; tip[c] = cur;
            h->head_slot[heads[c]] = c; h->rflags[heads[c]] |= kRHead; h->rflags[cur] |= kRTip;
        }
        // smallest number of roles that reaches the best critical path (longest-processing-time dealing)
        int best_r = 1, best_load = 1 << 30, owner[32];
        int order[32];
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
        {
            int load[kMaxRoles] = {0};
            for (int a = 0; a < nh; ++a) {
                int w = 0;
                for (int q = 1; q < best_r; ++q) if (load[q] < load[w]) w = q;
                load[w] += hlen[order[a]];
                owner[order[a]] = w;
            }
        }
        h->nroles = best_r; h->ntrunk = nt; h->nheads = nh;
        h->role_ok = (fits && best_r >= 2 && h->tpe_ok && L <= 24) ? 1 : 0;
        if (h->role_ok) {
            for (int t = 0, pos = 0; t <= nt; ++t) {  // heads attached to each trunk link
                h->th_begin[t] = pos;
                if (t == nt) break;
                for (int c = 0; c < nh; ++c)
                    if (h->parent[heads[c]] == h->trunk_link[t]) h->th_heads[pos++] = c;
            }
            for (int r = 0; r < best_r; ++r) {
                bool mine[32] = {false};
                int c2 = 0;
                for (int a = 0; a < nh; ++a) {  // chains in dealing order, each tip -> head
                    const int c = order[a];
                    if (owner[c] != r) continue;
                    for (int cur = tip[c];; cur = h->parent[cur]) {
                        h->list2[r][c2++] = cur; mine[cur] = true;
                        if (cur == heads[c]) break;
                    }
                }
                const int nchain = c2;
                if (r == 0) {
                    for (int t = nt - 1; t >= 1; --t) h->list2[0][c2++] = h->trunk_link[t];
                    h->list2[0][c2++] = 0;
                }
                h->n2[r] = c2;
                h->sync2[r] = (r == 0) ? nchain : c2;
                h->sync3[r] = nchain - 1;  // role 0 without chains: -1 = barrier after the sweep
                int c1 = 0;
                for (int i = 1; i < L; ++i)
                    if (trunk[i] || mine[i]) h->list1[r][c1++] = i;
                h->n1[r] = c1;
            }
        }
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
    for (int i = 1; i < L; ++i) {  // K2-R skips the own-inertia work of massless, sphere-less links
        bool zero = d->mass[i] == 0.f;
        for (int k = 0; k < 6; ++k) zero = zero && d->inertia[i][k] == 0.f;
        for (int s = 0; s < d->num_spheres; ++s) zero = zero && d->sphere_link[s] != i;
        for (int s = 0; s < d->num_sensors; ++s) zero = zero && d->sensor_link[s] != i;
        if (zero) h->rflags[i] |= kRMassless;
    }
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
        return nullptr;
    }
    for (int s = 0; s < d->num_sensors; ++s) {
        if (d->sensor_link[s] < 0 || d->sensor_link[s] >= L) {
            set_error("oige_model_create: sensor_link out of range");
            delete h;
            return nullptr;
        }
        h->sensor_of_link[d->sensor_link[s]] = s;
    }
    for (int k = 0; k < 3; ++k) h->gravity[k] = d->gravity[k];
    h->k_c = d->contact_stiffness; h->d_c = d->contact_damping; h->mu = d->friction_mu; h->c_visc = d->friction_viscous;
    h->k_lim = d->limit_stiffness; h->d_lim = d->limit_damping; h->max_qd = d->max_joint_velocity; h->ang_damp = d->angular_damping;

    const int tpe_ok = h->tpe_ok, role_ok = h->role_ok, nroles = h->nroles;
    const int role_smem = (h->nheads * 27 + h->ntrunk * 7 + 1) * 32 * (int)sizeof(float);
    void* dev = nullptr;
    cudaError_t e = cudaMalloc(&dev, sizeof(DevModel));
    if (e == cudaSuccess) e = cudaMemcpy(dev, h, sizeof(DevModel), cudaMemcpyHostToDevice);
    delete h;
    if (e != cudaSuccess) {
        set_error("oige_model_create: %s", cudaGetErrorString(e));
        if (dev) cudaFree(dev);
        return nullptr;
    }
    OigeModel* m = new OigeModel();
    m->dev_table = dev; m->num_links = L; m->num_dof = L - 1; m->num_sensors = d->num_sensors; m->tpe_ok = tpe_ok;
    m->role_ok = role_ok; m->nroles = nroles; m->role_smem = role_smem;
    return m;
}

void k2_model_destroy(OigeModel* m) {
    if (m->dev_table) cudaFree(m->dev_table);
    delete m;
}

}  // namespace oige
