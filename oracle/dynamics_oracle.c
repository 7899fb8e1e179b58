/*
 * TEST INFRASTRUCTURE ONLY -- float64 CPU reference integrator for the articulation step (K2).
 *
 * SUBSTITUTION NOTICE: the reference steps physics with PhysX 5 (World.step,
 * envs/vec_env_rlgames.py:64-66), which is closed-source, absent from /root/reference and not
 * installable here.  Per BASELINE.json's north_star the GPU kernel is therefore validated against
 * THIS float64 integrator of the same mechanical model (DESIGN.md "K2").  Parity vs PhysX is
 * unpinned and stated as such.
 *
 * Deliberately a DIFFERENT algorithm from the GPU kernel so that shared derivation errors cannot
 * hide: the kernel runs a world-frame articulated-body algorithm (O(n), spatial 6-vectors about a
 * moving reference point); this file assembles the dense joint-space equations with the
 * projection (Kane / Jacobian) method in absolute world coordinates with classical 3-vector
 * accelerations,
 *        (M + sum_p Jp^T C_p Jp + diag(arm + dt D + dt^2 K)) x = tau0 - h + sum_p Jp^T F0_p ,
 * x = (vdot_root, alphadot_root, qdd), solves it by Cholesky and applies the same
 * linearly-implicit Euler update.  Only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs may call it.
 *
 * Build: gcc -O2 -fopenmp -shared -fPIC -I include oracle/dynamics_oracle.c -o oracle/_build/liboracle_dyn.so -lm
 */
#include <math.h>
#include <stdint.h>
#include <string.h>

#include "oige_b200.h"

#define NV_MAX (6 + OIGE_MAX_DOF)

typedef struct { double x, y, z; } v3;
static v3 V(double x, double y, double z) { v3 r = {x, y, z}; return r; }
static v3 add(v3 a, v3 b) { return V(a.x + b.x, a.y + b.y, a.z + b.z); }
static v3 sub(v3 a, v3 b) { return V(a.x - b.x, a.y - b.y, a.z - b.z); }
static v3 scl(double s, v3 a) { return V(s * a.x, s * a.y, s * a.z); }
static v3 crs(v3 a, v3 b) { return V(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); }
static double dt3(v3 a, v3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
static double comp(v3 a, int k) { return k == 0 ? a.x : (k == 1 ? a.y : a.z); }

typedef struct { double m[3][3]; } m3;
static m3 quat_mat(const double q[4]) {
    double w = q[0], x = q[1], y = q[2], z = q[3];
    m3 R = {{{1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y)},
             {2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x)},
             {2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)}}};
    return R;
}
static m3 mmul(m3 a, m3 b) {
    m3 c;
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) c.m[i][j] = a.m[i][0] * b.m[0][j] + a.m[i][1] * b.m[1][j] + a.m[i][2] * b.m[2][j];
    return c;
}
static v3 mv(m3 a, v3 v) {
    return V(a.m[0][0] * v.x + a.m[0][1] * v.y + a.m[0][2] * v.z, a.m[1][0] * v.x + a.m[1][1] * v.y + a.m[1][2] * v.z,
             a.m[2][0] * v.x + a.m[2][1] * v.y + a.m[2][2] * v.z);
}
static v3 mtv(m3 a, v3 v) {
    return V(a.m[0][0] * v.x + a.m[1][0] * v.y + a.m[2][0] * v.z, a.m[0][1] * v.x + a.m[1][1] * v.y + a.m[2][1] * v.z,
             a.m[0][2] * v.x + a.m[1][2] * v.y + a.m[2][2] * v.z);
}
static m3 axis_angle(v3 a, double th) { /* Rodrigues */
    double c = cos(th), s = sin(th), t = 1 - c;
    m3 R = {{{c + a.x * a.x * t, a.x * a.y * t - a.z * s, a.x * a.z * t + a.y * s},
             {a.y * a.x * t + a.z * s, c + a.y * a.y * t, a.y * a.z * t - a.x * s},
             {a.z * a.x * t - a.y * s, a.z * a.y * t + a.x * s, c + a.z * a.z * t}}};
    return R;
}

/* generalized velocity layout: [0:3) root linear (classical), [3:6) root angular, [6:6+D) joints by LINK index-1 */
typedef struct {
    v3 jw[NV_MAX]; /* angular Jacobian columns */
    v3 jv[NV_MAX]; /* linear Jacobian columns (of a given point) */
} jac;

static void point_jacobian(const OigeModelDesc* M, int link, v3 point, const v3* p, const v3* axw, jac* J, int nv) {
    for (int c = 0; c < nv; ++c) { J->jw[c] = V(0, 0, 0); J->jv[c] = V(0, 0, 0); }
    for (int k = 0; k < 3; ++k) {
        v3 e = V(k == 0, k == 1, k == 2);
        J->jv[k] = e;
        J->jw[3 + k] = e;
        J->jv[3 + k] = crs(e, sub(point, p[0]));
    }
    for (int j = link; j > 0; j = M->parent[j]) {
        if (M->joint_type[j] == OIGE_JOINT_PRISMATIC) { /* pure translation of everything downstream */
            J->jv[6 + j - 1] = axw[j];
        } else {
            J->jw[6 + j - 1] = axw[j];
            J->jv[6 + j - 1] = crs(axw[j], sub(point, p[j]));
        }
    }
}

static int cholesky_solve(double* A, double* b, int n, int lda) { /* in place; A lower used */
    for (int j = 0; j < n; ++j) {
        double d = A[j * lda + j];
        for (int k = 0; k < j; ++k) d -= A[j * lda + k] * A[j * lda + k];
        if (!(d > 0)) return -1;
        d = sqrt(d);
        A[j * lda + j] = d;
        for (int i = j + 1; i < n; ++i) {
            double t = A[i * lda + j];
            for (int k = 0; k < j; ++k) t -= A[i * lda + k] * A[j * lda + k];
            A[i * lda + j] = t / d;
        }
    }
    for (int i = 0; i < n; ++i) {
        double t = b[i];
        for (int k = 0; k < i; ++k) t -= A[i * lda + k] * b[k];
        b[i] = t / A[i * lda + i];
    }
    for (int i = n - 1; i >= 0; --i) {
        double t = b[i];
        for (int k = i + 1; k < n; ++k) t -= A[k * lda + i] * b[k];
        b[i] = t / A[i * lda + i];
    }
    return 0;
}

typedef struct { int active; v3 r; v3 F0; v3 C; } contact;

static contact contact_eval(const OigeModelDesc* M, int s, m3 R, v3 p, v3 vlin, v3 w, double dt) {
    contact c;
    memset(&c, 0, sizeof c);
    v3 x = add(p, mv(R, V(M->sphere_pos[s][0], M->sphere_pos[s][1], M->sphere_pos[s][2])));
    double rad = M->sphere_radius[s];
    double pen = rad - x.z;
    if (!(pen > 0)) return c;
    if (M->max_depenetration_velocity > 0) pen = fmin(pen, (double)M->max_depenetration_velocity * dt);
    c.r = V(x.x, x.y, x.z - rad);
    v3 vp = add(vlin, crs(w, sub(c.r, p)));
    double vn = vp.z;
    double Gn = dt * (double)M->contact_stiffness + (vn < 0 ? (double)M->contact_damping : 0.0);
    double fn0 = (double)M->contact_stiffness * pen - Gn * vn;
    if (!(fn0 > 0)) return c;
    double vt = sqrt(vp.x * vp.x + vp.y * vp.y);
    double ct = fmin((double)M->friction_viscous, (double)M->friction_mu * fn0 / (vt + 1e-6));
    c.C = V(dt * ct, dt * ct, dt * Gn);
    c.F0 = V(-ct * vp.x, -ct * vp.y, fn0);
    c.active = 1;
    return c;
}

static int step_one(const OigeModelDesc* M, double* pos, double* quat, double* vel, double* q, double* qd,
                    const double* eff, const double* tgt, double* sens, int last, double dt) {
    const int L = M->num_links, D = L - 1, nv = 6 + D;
    m3 R[OIGE_MAX_LINKS];
    v3 p[OIGE_MAX_LINKS], axw[OIGE_MAX_LINKS], w[OIGE_MAX_LINKS], vo[OIGE_MAX_LINKS], al0[OIGE_MAX_LINKS], ac0[OIGE_MAX_LINKS];
    double A[NV_MAX * NV_MAX], rhs[NV_MAX];
    memset(A, 0, sizeof A);
    memset(rhs, 0, sizeof rhs);
    /* 1. kinematics */
    R[0] = quat_mat(quat);
    p[0] = V(pos[0], pos[1], pos[2]);
    vo[0] = V(vel[0], vel[1], vel[2]);
    w[0] = V(vel[3], vel[4], vel[5]);
    al0[0] = V(0, 0, 0);
    ac0[0] = V(0, 0, 0);
    axw[0] = V(0, 0, 0);
    for (int i = 1; i < L; ++i) {
        int pa = M->parent[i], d = M->dof_index[i];
        v3 ax = V(M->joint_axis[i][0], M->joint_axis[i][1], M->joint_axis[i][2]);
        double jq[4] = {M->joint_quat[i][0], M->joint_quat[i][1], M->joint_quat[i][2], M->joint_quat[i][3]};
        v3 r = mv(R[pa], V(M->joint_pos[i][0], M->joint_pos[i][1], M->joint_pos[i][2]));
        if (M->joint_type[i] == OIGE_JOINT_PRISMATIC) {
            /* the link slides along the axis (fixed in the parent): origin offset r + axw q, relative
             * velocity axw qd, hence the Coriolis term 2 w x (axw qd) in the origin's acceleration */
            R[i] = mmul(R[pa], quat_mat(jq));
            axw[i] = mv(R[i], ax);
            r = add(r, scl(q[d], axw[i]));
            p[i] = add(p[pa], r);
            w[i] = w[pa];
            vo[i] = add(vo[pa], add(crs(w[pa], r), scl(qd[d], axw[i])));
            al0[i] = al0[pa];
            ac0[i] = add(ac0[pa], add(add(crs(al0[pa], r), crs(w[pa], crs(w[pa], r))),
                                      scl(2.0 * qd[d], crs(w[pa], axw[i]))));
            continue;
        }
        p[i] = add(p[pa], r);
        R[i] = mmul(mmul(R[pa], quat_mat(jq)), axis_angle(ax, q[d]));
        axw[i] = mv(R[i], ax);
        w[i] = add(w[pa], scl(qd[d], axw[i]));
        vo[i] = add(vo[pa], crs(w[pa], r));
        al0[i] = add(al0[pa], scl(qd[d], crs(w[pa], axw[i])));
        ac0[i] = add(ac0[pa], add(crs(al0[pa], r), crs(w[pa], crs(w[pa], r))));
    }
    /* 2. inertia matrix and bias by projection */
    v3 g = V(M->gravity[0], M->gravity[1], M->gravity[2]);
    jac J;
    for (int i = 0; i < L; ++i) {
        double m = M->mass[i];
        if (m == 0.0) continue;
        v3 rc = mv(R[i], V(M->com[i][0], M->com[i][1], M->com[i][2]));
        v3 cw = add(p[i], rc);
        m3 Il = {{{M->inertia[i][0], M->inertia[i][3], M->inertia[i][4]},
                  {M->inertia[i][3], M->inertia[i][1], M->inertia[i][5]},
                  {M->inertia[i][4], M->inertia[i][5], M->inertia[i][2]}}};
        m3 Rt;
        for (int a = 0; a < 3; ++a) for (int b = 0; b < 3; ++b) Rt.m[a][b] = R[i].m[b][a];
        m3 Iw = mmul(mmul(R[i], Il), Rt);
        v3 a0c = add(ac0[i], add(crs(al0[i], rc), crs(w[i], crs(w[i], rc))));
        v3 fb = scl(m, sub(a0c, g));                         /* linear bias incl. gravity */
        v3 tb = add(mv(Iw, al0[i]), crs(w[i], mv(Iw, w[i]))); /* angular bias */
        if (i > 0) tb = add(tb, scl((double)M->angular_damping, mv(Iw, w[i]))); /* damping torque -c I w (root: step 7) */
        point_jacobian(M, i, cw, p, axw, &J, nv);
        for (int a = 0; a < nv; ++a) {
            v3 Iwa = mv(Iw, J.jw[a]);
            rhs[a] -= dt3(J.jv[a], fb) + dt3(J.jw[a], tb);
            for (int b = 0; b <= a; ++b) A[a * NV_MAX + b] += m * dt3(J.jv[a], J.jv[b]) + dt3(Iwa, J.jw[b]);
        }
    }
    /* 3. joints: efforts, passive spring-damper, limit penalty; implicit diagonal */
    for (int i = 1; i < L; ++i) {
        int d = M->dof_index[i], c = 6 + i - 1;
        double Ks = M->stiffness[i], Dj = M->damping[i];
        double qn = q[d] + dt * qd[d];
        double Kimp = Ks, Dimp = Dj;
        double tau = eff[d] - Ks * (qn - (double)M->spring_ref[i]) - Dj * qd[d];
        double lo = M->limit_lower[i], hi = M->limit_upper[i];
        if (q[d] < lo) { tau += -(double)M->limit_stiffness * (qn - lo) - (double)M->limit_damping * qd[d]; Kimp += M->limit_stiffness; Dimp += M->limit_damping; }
        else if (q[d] > hi) { tau += -(double)M->limit_stiffness * (qn - hi) - (double)M->limit_damping * qd[d]; Kimp += M->limit_stiffness; Dimp += M->limit_damping; }
        if (tgt) { /* PD position drive, implicit; past the force limit both gains are scaled by fmax/|torque| */
            double kp = M->drive_stiffness[i], kd = M->drive_damping[i], fmax = M->drive_max_force[i];
            double tpd = kp * (tgt[d] - qn) - kd * qd[d];
            double alpha = (fmax > 0 && fabs(tpd) > fmax) ? fmax / fabs(tpd) : 1.0;
            tau += alpha * tpd; Kimp += alpha * kp; Dimp += alpha * kd;
        }
        rhs[c] += tau;
        A[c * NV_MAX + c] += (double)M->armature[i] + dt * (Dimp + dt * Kimp);
    }
    /* 4. ground contacts */
    contact cons[OIGE_MAX_SPHERES];
    for (int s = 0; s < M->num_spheres; ++s) {
        int i = M->sphere_link[s];
        cons[s] = contact_eval(M, s, R[i], p[i], vo[i], w[i], dt);
        if (!cons[s].active) continue;
        point_jacobian(M, i, cons[s].r, p, axw, &J, nv);
        for (int a = 0; a < nv; ++a) {
            rhs[a] += dt3(J.jv[a], cons[s].F0);
            for (int b = 0; b <= a; ++b)
                A[a * NV_MAX + b] += cons[s].C.x * J.jv[a].x * J.jv[b].x + cons[s].C.y * J.jv[a].y * J.jv[b].y +
                                     cons[s].C.z * J.jv[a].z * J.jv[b].z;
        }
    }
    /* 4b. self-collision: explicit capsule-capsule penalty between the links of each collision pair.  One contact per
     * pair at the closest points of the two segments; equal and opposite forces at the midpoint of the overlap; damping
     * only while approaching; no friction. */
    for (int k = 0; k < M->num_collision_pairs; ++k) {
        int la = M->pair_link_a[k], lb = M->pair_link_b[k];
        double ra = M->capsule_radius[la], rb = M->capsule_radius[lb];
        if (!(ra > 0) || !(rb > 0)) continue;
        v3 P1 = add(p[la], mv(R[la], V(M->capsule_a[la][0], M->capsule_a[la][1], M->capsule_a[la][2])));
        v3 Q1 = add(p[la], mv(R[la], V(M->capsule_b[la][0], M->capsule_b[la][1], M->capsule_b[la][2])));
        v3 P2 = add(p[lb], mv(R[lb], V(M->capsule_a[lb][0], M->capsule_a[lb][1], M->capsule_a[lb][2])));
        v3 Q2 = add(p[lb], mv(R[lb], V(M->capsule_b[lb][0], M->capsule_b[lb][1], M->capsule_b[lb][2])));
        /* closest points of two segments (Ericson, Real-Time Collision Detection 5.1.9) */
        v3 d1 = sub(Q1, P1), d2 = sub(Q2, P2), r = sub(P1, P2);
        double a = dt3(d1, d1), e = dt3(d2, d2), f = dt3(d2, r), c = dt3(d1, r), b = dt3(d1, d2);
        double sp = 0, tp = 0, eps = 1e-12;
        if (a <= eps && e <= eps) { sp = tp = 0; }
        else if (a <= eps) { sp = 0; tp = fmin(fmax(f / e, 0.0), 1.0); }
        else if (e <= eps) { tp = 0; sp = fmin(fmax(-c / a, 0.0), 1.0); }
        else {
            double den = a * e - b * b;
            sp = den > 1e-9 * a * e ? fmin(fmax((b * f - c * e) / den, 0.0), 1.0) : 0.0;
            tp = (b * sp + f) / e;
            if (tp < 0) { tp = 0; sp = fmin(fmax(-c / a, 0.0), 1.0); }
            else if (tp > 1) { tp = 1; sp = fmin(fmax((b - c) / a, 0.0), 1.0); }
        }
        v3 c1 = add(P1, scl(sp, d1)), c2 = add(P2, scl(tp, d2));
        v3 d = sub(c1, c2);
        double dist = sqrt(dt3(d, d)), pen = ra + rb - dist;
        if (!(pen > 0) || !(dist > 1e-9)) continue;
        v3 n = scl(1.0 / dist, d); /* from b to a */
        v3 pc = add(c2, scl(rb - 0.5 * pen, n));
        v3 va = add(vo[la], crs(w[la], sub(pc, p[la])));
        v3 vb = add(vo[lb], crs(w[lb], sub(pc, p[lb])));
        double vn = dt3(sub(va, vb), n);
        double penc = pen;
        if (M->max_depenetration_velocity > 0) penc = fmin(penc, (double)M->max_depenetration_velocity * dt);
        double fm = (double)M->self_contact_stiffness * penc - (vn < 0 ? (double)M->self_contact_damping * vn : 0.0);
        if (!(fm > 0)) continue;
        v3 F = scl(fm, n);
        point_jacobian(M, la, pc, p, axw, &J, nv);
        for (int q2 = 0; q2 < nv; ++q2) rhs[q2] += dt3(J.jv[q2], F);
        point_jacobian(M, lb, pc, p, axw, &J, nv);
        for (int q2 = 0; q2 < nv; ++q2) rhs[q2] -= dt3(J.jv[q2], F);
    }
    /* 5. solve (a welded base: its six rows/columns become the identity with zero right-hand side) */
    if (M->fixed_base) {
        for (int a = 0; a < 6; ++a) {
            for (int b = 0; b <= a; ++b) A[a * NV_MAX + b] = (a == b) ? 1.0 : 0.0;
            for (int c = 6; c < nv; ++c) A[c * NV_MAX + a] = 0.0;
            rhs[a] = 0.0;
        }
    }
    if (cholesky_solve(A, rhs, nv, NV_MAX) != 0) return -1;
    const double* x = rhs;
    /* 6. sensors: net ground wrench on each sensor link, in the link frame, torque about the link origin */
    if (last && sens) {
        for (int sidx = 0; sidx < M->num_sensors; ++sidx) {
            int i = M->sensor_link[sidx];
            v3 F = V(0, 0, 0), T = V(0, 0, 0);
            for (int s = 0; s < M->num_spheres; ++s) {
                if (M->sphere_link[s] != i || !cons[s].active) continue;
                point_jacobian(M, i, cons[s].r, p, axw, &J, nv);
                v3 dv = V(0, 0, 0);
                for (int a = 0; a < nv; ++a) dv = add(dv, scl(x[a], J.jv[a]));
                v3 f = V(cons[s].F0.x - cons[s].C.x * dv.x, cons[s].F0.y - cons[s].C.y * dv.y, cons[s].F0.z - cons[s].C.z * dv.z);
                F = add(F, f);
                T = add(T, crs(sub(cons[s].r, p[i]), f));
            }
            v3 Fl = mtv(R[i], F), Tl = mtv(R[i], T);
            double* o = sens + sidx * 6;
            o[0] = Fl.x; o[1] = Fl.y; o[2] = Fl.z; o[3] = Tl.x; o[4] = Tl.y; o[5] = Tl.z;
        }
    }
    /* 7. integrate */
    for (int i = 1; i < L; ++i) {
        int d = M->dof_index[i];
        double v = qd[d] + dt * x[6 + i - 1];
        double mq = M->max_joint_velocity; /* JointAPI limit; a hinge's relative angular velocity also obeys the body limit */
        if (M->joint_type[i] != OIGE_JOINT_PRISMATIC && M->max_angular_velocity > 0) mq = fmin(mq, (double)M->max_angular_velocity);
        v = fmin(fmax(v, -mq), mq);
        qd[d] = v;
        q[d] += dt * v;
    }
    if (M->fixed_base) return 0;
    double damp = 1.0 - dt * (double)M->angular_damping;
    if (damp < 0) damp = 0;
    v3 w1 = scl(damp, add(w[0], scl(dt, V(x[3], x[4], x[5]))));
    if (M->max_angular_velocity > 0) { /* RigidBodyAPI max angular velocity of the root body */
        double wn2 = dt3(w1, w1), wm = M->max_angular_velocity;
        if (wn2 > wm * wm) w1 = scl(wm / sqrt(wn2), w1);
    }
    v3 v1 = add(vo[0], scl(dt, V(x[0], x[1], x[2])));
    vel[0] = v1.x; vel[1] = v1.y; vel[2] = v1.z; vel[3] = w1.x; vel[4] = w1.y; vel[5] = w1.z;
    pos[0] += dt * v1.x; pos[1] += dt * v1.y; pos[2] += dt * v1.z;
    double wn = sqrt(dt3(w1, w1)), th = 0.5 * dt * wn;
    double k = wn > 1e-9 ? sin(th) / wn : 0.5 * dt;
    double dq[4] = {cos(th), k * w1.x, k * w1.y, k * w1.z};
    double a[4] = {quat[0], quat[1], quat[2], quat[3]};
    double nq[4] = {dq[0] * a[0] - dq[1] * a[1] - dq[2] * a[2] - dq[3] * a[3],
                    dq[0] * a[1] + dq[1] * a[0] + dq[2] * a[3] - dq[3] * a[2],
                    dq[0] * a[2] - dq[1] * a[3] + dq[2] * a[0] + dq[3] * a[1],
                    dq[0] * a[3] + dq[1] * a[2] - dq[2] * a[1] + dq[3] * a[0]};
    double inv = 1.0 / sqrt(nq[0] * nq[0] + nq[1] * nq[1] + nq[2] * nq[2] + nq[3] * nq[3]);
    for (int c = 0; c < 4; ++c) quat[c] = nq[c] * inv;
    (void)comp;
    return 0;
}

/* Steps n_envs environments `substeps` times.  vel is (lin xyz, ang xyz) like ArticulationView.
 * efforts [N,D] are joint torques in DoF order (actions * gears * power_scale, formed by the caller).
 * Returns 0, or the number of envs whose system matrix was not positive definite. */
int oracle_articulation_step_f64(const OigeModelDesc* M, int64_t n_envs, double* pos, double* quat, double* vel,
                                 double* q, double* qd, const double* efforts, const double* targets,
                                 double* sensors, int substeps, double dt, int num_threads) {
    const int D = M->num_links - 1, S = M->num_sensors;
    int bad = 0;
    (void)num_threads;
#pragma omp parallel for schedule(static) num_threads(num_threads) reduction(+ : bad)
    for (int64_t e = 0; e < n_envs; ++e) {
        for (int s = 0; s < substeps; ++s) {
            int rc = step_one(M, pos + 3 * e, quat + 4 * e, vel + 6 * e, q + D * e, qd + D * e, efforts + D * e,
                              targets ? targets + D * e : 0, sensors ? sensors + (int64_t)S * 6 * e : 0, s == substeps - 1, dt);
            if (rc != 0) { bad += 1; break; }
        }
    }
    return bad;
}

/* Total mechanical energy (kinetic + gravitational potential + joint springs) -- used by invariants tests. */
double oracle_energy_f64(const OigeModelDesc* M, const double* pos, const double* quat, const double* vel,
                         const double* q, const double* qd) {
    const int L = M->num_links;
    m3 R[OIGE_MAX_LINKS];
    v3 p[OIGE_MAX_LINKS], w[OIGE_MAX_LINKS], vo[OIGE_MAX_LINKS];
    R[0] = quat_mat(quat); p[0] = V(pos[0], pos[1], pos[2]); vo[0] = V(vel[0], vel[1], vel[2]); w[0] = V(vel[3], vel[4], vel[5]);
    double E = 0;
    for (int i = 0; i < L; ++i) {
        if (i > 0) {
            int pa = M->parent[i], d = M->dof_index[i];
            v3 ax = V(M->joint_axis[i][0], M->joint_axis[i][1], M->joint_axis[i][2]);
            double jq[4] = {M->joint_quat[i][0], M->joint_quat[i][1], M->joint_quat[i][2], M->joint_quat[i][3]};
            v3 r = mv(R[pa], V(M->joint_pos[i][0], M->joint_pos[i][1], M->joint_pos[i][2]));
            if (M->joint_type[i] == OIGE_JOINT_PRISMATIC) {
                R[i] = mmul(R[pa], quat_mat(jq));
                v3 a = mv(R[i], ax);
                r = add(r, scl(q[d], a));
                p[i] = add(p[pa], r);
                w[i] = w[pa];
                vo[i] = add(vo[pa], add(crs(w[pa], r), scl(qd[d], a)));
            } else {
                p[i] = add(p[pa], r);
                R[i] = mmul(mmul(R[pa], quat_mat(jq)), axis_angle(ax, q[d]));
                w[i] = add(w[pa], scl(qd[d], mv(R[i], ax)));
                vo[i] = add(vo[pa], crs(w[pa], r));
            }
            E += 0.5 * (double)M->armature[i] * qd[d] * qd[d];
            double dq = q[d] - (double)M->spring_ref[i];
            E += 0.5 * (double)M->stiffness[i] * dq * dq;
        }
        double m = M->mass[i];
        if (m == 0.0) continue;
        v3 rc = mv(R[i], V(M->com[i][0], M->com[i][1], M->com[i][2]));
        v3 vc = add(vo[i], crs(w[i], rc));
        v3 wl = mtv(R[i], w[i]);
        double Iw = M->inertia[i][0] * wl.x * wl.x + M->inertia[i][1] * wl.y * wl.y + M->inertia[i][2] * wl.z * wl.z +
                    2 * (M->inertia[i][3] * wl.x * wl.y + M->inertia[i][4] * wl.x * wl.z + M->inertia[i][5] * wl.y * wl.z);
        E += 0.5 * m * dt3(vc, vc) + 0.5 * Iw;
        v3 g = V(M->gravity[0], M->gravity[1], M->gravity[2]);
        E -= m * dt3(g, add(p[i], rc));
    }
    return E;
}
