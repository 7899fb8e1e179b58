// K2-R `articulation_step_role_kernel` -- the thread-per-env articulation step split over the warps of a CTA.
// (Included by articulation_step.cu inside its anonymous namespace: same DevModel, same device functions.)
//
// K2-T runs one thread per env; at 16384 envs that is a single warp per scheduler walking ~50 k dependent
// instructions per step (latency floor ~100 us, 28 % issue utilisation).  Here a CTA carries 32 envs and `nroles`
// warps; lane = env, warp = ROLE.  The tree is cut into a TRUNK (root + every link that has a branching link at or
// below it) and the plain CHAINS hanging off it (limbs); the chains are dealt to the roles, role 0 also owns the trunk.
// Per substep:
//   phase A (all roles, no communication): kinematics of the trunk (recomputed by every role: a few links) and of
//            the role's own chains; articulated inertia up each own chain (carried in registers); the chain head's
//            I^A, p^A (27 floats) go to a shared-memory slot [slot][field][lane].
//   barrier 1
//   phase B (role 0): trunk bottom-up (own inertia + attached heads' slots + trunk children), 6x6 Cholesky at the
//            root, trunk accelerations top-down; root / trunk accelerations and trunk joint accelerations published.
//   barrier 2
//   phase C (all roles): accelerations down the own chains, joint integration, foot sensors; every role integrates
//            the root and the trunk joints itself (same inputs, same code: bit-identical copies).
// Control flow is uniform inside a warp (role is warp-uniform); the barriers are PTX named barriers (`bar.sync 1, n`)
// reached from role-dependent program points, the usual warp-specialisation pattern.

__device__ __forceinline__ void role_barrier(int nthreads) {
    asm volatile("bar.sync 1, %0;" ::"r"(nthreads) : "memory");
}

// NR = warps per CTA the register budget is sized for (>= the model's roles), MINB = CTAs per SM to fit.
template <int MAXL, int NR, int MINB>
__global__ void __launch_bounds__(NR * 32, MINB)
articulation_step_role_kernel(const DevModel* __restrict__ gmodel, const __grid_constant__ OigeTaskParams P,
                              const __grid_constant__ K2Args A) {
    __shared__ __align__(16) DevModel M;
    extern __shared__ __align__(16) float dynsm[];
    {
        const int4* src = reinterpret_cast<const int4*>(gmodel);
        int4* dst = reinterpret_cast<int4*>(&M);
        for (int i = threadIdx.x; i < (int)(sizeof(DevModel) / 16); i += blockDim.x) dst[i] = src[i];
    }
    __syncthreads();
    pdl_wait();     // state / actions are written by the preceding kernel (reset) -- the model table is not
    pdl_trigger();  // the observation kernel's CTAs may take the SM slots this launch frees as it drains
    const int role = threadIdx.x >> 5, lane = threadIdx.x & 31, nthr = blockDim.x;
    float* xI = dynsm;                                  // [nheads][27][32]
    float* xA = xI + M.nheads * 27 * 32;                // [ntrunk][7][32]: acceleration (6), joint acceleration
    int* xbad = reinterpret_cast<int*>(xA + M.ntrunk * 7 * 32);  // [32]
    if (role == 0) xbad[lane] = 0;
    const long long env0 = (long long)blockIdx.x * 32 + lane;
    const bool valid = env0 < A.n;
    const long long env = valid ? env0 : A.n - 1;  // idle lanes shadow the last env and store nothing

    const int L = M.num_links, D = P.num_dof;
    const float dt = A.dt;
    const V3 grav = {M.gravity[0], M.gravity[1], M.gravity[2]};
    const int n1 = M.n1[role];
    const int n2 = M.n2[role];  // role 0: includes the trunk and, last, the root
    const int sync2 = M.sync2[role], sync3 = M.sync3[role];

    float fk[MAXL][19];
    float pu[MAXL][8];
    float accT[kMaxTrunk][27];
    float aT[kMaxTrunk][6];
    float qv[MAXL], qdv[MAXL], eff[MAXL], tgt[MAXL];

    for (int k = 0; k < n1; ++k) {
        const int i = M.list1[role][k];
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
        // ---- phase A.1: kinematics of the trunk + own chains, increasing index (parent[i] < i) ---------------
        for (int k = 0; k < n1; ++k) {
            const int i = M.list1[role][k];
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
                for (int j = 0; j < 6; ++j) { pv.a[j] = fk[pa][7 + j]; pb.a[j] = fk[pa][13 + j]; }
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
            for (int j = 0; j < 6; ++j) v.a[j] = fmaf(s.a[j], qd, pv.a[j]);
            const V3 w = ang(v), vl = lin(v);
            const V3 ca = cross(w, ang(s));
            const V3 cl = cross(w, lin(s)) + cross(vl, ang(s));
            fk[i][0] = qi.w; fk[i][1] = qi.x; fk[i][2] = qi.y; fk[i][3] = qi.z;
            fk[i][4] = p.x; fk[i][5] = p.y; fk[i][6] = p.z;
#pragma unroll
            for (int j = 0; j < 6; ++j) fk[i][7 + j] = v.a[j];
            fk[i][13] = pb.a[0] + qd * ca.x; fk[i][14] = pb.a[1] + qd * ca.y; fk[i][15] = pb.a[2] + qd * ca.z;
            fk[i][16] = pb.a[3] + qd * cl.x; fk[i][17] = pb.a[4] + qd * cl.y; fk[i][18] = pb.a[5] + qd * cl.z;
        }
        // ---- phase A.2 / B.1: articulated inertia: own chains tip -> head, then (role 0) the trunk ------------
        if (role == 0) {
            for (int t = 0; t < M.ntrunk; ++t) {
#pragma unroll
                for (int j = 0; j < 27; ++j) accT[t][j] = 0.f;
            }
        }
        S6 carryI;
        V6 carryP;
#pragma unroll
        for (int j = 0; j < 21; ++j) carryI.a[j] = 0.f;
#pragma unroll
        for (int j = 0; j < 6; ++j) carryP.a[j] = 0.f;
        for (int k2 = 0; k2 < n2; ++k2) {
            if (k2 == sync2) role_barrier(nthr);  // role 0: every chain head is in its slot from here on
            const int i = M.list2[role][k2];
            M3 R;
            V3 p;
            V6 v, b;
            if (i == 0) {
                R = R0; p = {0.f, 0.f, 0.f}; v = rvel; b = b0;
            } else {
                R = qmat(Q4{fk[i][0], fk[i][1], fk[i][2], fk[i][3]});
                p = {fk[i][4], fk[i][5], fk[i][6]};
#pragma unroll
                for (int j = 0; j < 6; ++j) { v.a[j] = fk[i][7 + j]; b.a[j] = fk[i][13 + j]; }
            }
            S6 IA;
            V6 pA;
            const int flags = M.rflags[i];
            if (flags & kRMassless) {  // intermediate link of a multi-dof joint: no inertia, no bias, no contacts
#pragma unroll
                for (int j = 0; j < 21; ++j) IA.a[j] = 0.f;
#pragma unroll
                for (int j = 0; j < 6; ++j) pA.a[j] = 0.f;
            } else {
                link_inertia_bias(M, i, R, p, v, grav, IA, pA);
                link_contacts(M, i, R, p, v, b, root_z, dt, IA, pA);
            }
            if (flags & kRTrunk) {
                const int t = M.trunk_idx[i];
#pragma unroll
                for (int j = 0; j < 21; ++j) IA.a[j] += accT[t][j];
#pragma unroll
                for (int j = 0; j < 6; ++j) pA.a[j] += accT[t][21 + j];
                for (int h = M.th_begin[t]; h < M.th_begin[t + 1]; ++h) {
                    const float* src = xI + (M.th_heads[h] * 27) * 32 + lane;
#pragma unroll
                    for (int j = 0; j < 21; ++j) IA.a[j] += src[j * 32];
#pragma unroll
                    for (int j = 0; j < 6; ++j) pA.a[j] += src[(21 + j) * 32];
                }
            } else if (!(flags & kRTip)) {
#pragma unroll
                for (int j = 0; j < 21; ++j) IA.a[j] += carryI.a[j];
#pragma unroll
                for (int j = 0; j < 6; ++j) pA.a[j] += carryP.a[j];
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
            for (int j = 0; j < 6; ++j) pA.a[j] = pA.a[j] + Iac.a[j] + U.a[j] * ud;
#pragma unroll
            for (int j = 0; j < 6; ++j) pu[i][j] = U.a[j];
            pu[i][6] = dinv; pu[i][7] = u - dot6(U, c);  // phase C: qdd = (u - U.(a_parent + c)) / d
            if (flags & kRTrunk) {  // trunk link: add to the trunk parent's accumulator
                const int tp = M.trunk_idx[M.parent[i]];
#pragma unroll
                for (int j = 0; j < 21; ++j) accT[tp][j] += IA.a[j];
#pragma unroll
                for (int j = 0; j < 6; ++j) accT[tp][21 + j] += pA.a[j];
            } else if (flags & kRHead) {  // chain head: hand I^a, p^a to the trunk through shared memory
                float* dst = xI + (M.head_slot[i] * 27) * 32 + lane;
#pragma unroll
                for (int j = 0; j < 21; ++j) dst[j * 32] = IA.a[j];
#pragma unroll
                for (int j = 0; j < 6; ++j) dst[(21 + j) * 32] = pA.a[j];
            } else {
                carryI = IA; carryP = pA;
            }
        }
        if (sync2 >= n2) role_barrier(nthr);  // roles without trunk work arrive after their last chain
        // ---- phase B.2: root solve I^A a = -p^A (6x6 Cholesky), published for every role ---------------------
        if (role == 0) {
            V6 a0 = {{0.f, 0.f, 0.f, 0.f, 0.f, 0.f}};
            if (!M.fixed_base) chol6_solve(carryI, carryP, a0);  // a welded base does not accelerate
#pragma unroll
            for (int j = 0; j < 6; ++j) { aT[0][j] = a0.a[j]; xA[j * 32 + lane] = a0.a[j]; }
        }
        // ---- phase B.3 / C: accelerations: (role 0) trunk top-down, then own chains head -> tip ------------------
        V6 acarry = {{0.f, 0.f, 0.f, 0.f, 0.f, 0.f}};
        for (int k2 = (role == 0 ? n2 - 2 : n2 - 1); k2 >= 0; --k2) {
            if (k2 == sync3) role_barrier(nthr);  // trunk accelerations are published from here on
            const int i = M.list2[role][k2];
            const int flags = M.rflags[i];
            V6 ap;
            if (flags & kRTrunk) {
                const int tp = M.trunk_idx[M.parent[i]];
#pragma unroll
                for (int j = 0; j < 6; ++j) ap.a[j] = aT[tp][j];
            } else if (flags & kRHead) {
                const float* src = xA + (M.trunk_idx[M.parent[i]] * 7) * 32 + lane;
#pragma unroll
                for (int j = 0; j < 6; ++j) ap.a[j] = src[j * 32];
            } else {
                ap = acarry;
            }
            // recompute s and c from the stored kinematics (cheaper than storing 12 more floats per link)
            const M3 R = qmat(Q4{fk[i][0], fk[i][1], fk[i][2], fk[i][3]});
            const V3 p = {fk[i][4], fk[i][5], fk[i][6]};
            V6 v;
#pragma unroll
            for (int j = 0; j < 6; ++j) v.a[j] = fk[i][7 + j];
            const V6 s = joint_subspace(M.jtype[i] == OIGE_JOINT_PRISMATIC,
                                        mulv(R, V3{M.axis[0][i], M.axis[1][i], M.axis[2][i]}), p);
            const float qd0 = qdv[i];
            V6 U;
#pragma unroll
            for (int j = 0; j < 6; ++j) U.a[j] = pu[i][j];
            const float qdd = (pu[i][7] - dot6(U, ap)) * pu[i][6];
            V6 a;
            {
                const V6 c = velocity_product(v, s, qd0);
#pragma unroll
                for (int j = 0; j < 6; ++j) a.a[j] = fmaf(s.a[j], qdd, ap.a[j] + c.a[j]);
            }
            if (last && A.sens != nullptr && M.sensor_of_link[i] >= 0 && valid) {
                V6 b;
#pragma unroll
                for (int j = 0; j < 6; ++j) b.a[j] = fk[i][13 + j];
                sensor_wrench(M, i, R, p, v, b, a, root_z, dt, A.sens + (env * M.num_sensors + M.sensor_of_link[i]) * 6);
            }
            if (flags & kRTrunk) {  // trunk joints are integrated by every role after the barrier
                const int t = M.trunk_idx[i];
                float* dst = xA + (t * 7) * 32 + lane;
#pragma unroll
                for (int j = 0; j < 6; ++j) { aT[t][j] = a.a[j]; dst[j * 32] = a.a[j]; }
                dst[6 * 32] = qdd;
            } else {
                // integrate this joint (semi-implicit Euler); q/qd of link i are not read again this substep
                float qd = fmaf(dt, qdd, qdv[i]);
                qd = fminf(fmaxf(qd, -M.max_qd), M.max_qd);
                qdv[i] = qd;
                qv[i] = fmaf(dt, qd, qv[i]);
                acarry = a;
            }
        }
        if (sync3 < 0) role_barrier(nthr);
        // every role advances its own copy of the trunk joints and of the root
        for (int t = 1; t < M.ntrunk; ++t) {
            const int i = M.trunk_link[t];
            const float qdd = xA[(t * 7 + 6) * 32 + lane];
            float qd = fmaf(dt, qdd, qdv[i]);
            qd = fminf(fmaxf(qd, -M.max_qd), M.max_qd);
            qdv[i] = qd;
            qv[i] = fmaf(dt, qd, qv[i]);
        }
        if (!M.fixed_base) {
            V6 a0;
#pragma unroll
            for (int j = 0; j < 6; ++j) a0.a[j] = xA[j * 32 + lane];
            integrate_root(M, a0, dt, rpos, rquat, rvel);
        }
    }
    // ---- non-finite guard (agreed across the roles of an env) + store ---------------------------------------
    bool bad = !(isfinite(rpos.x) && isfinite(rpos.y) && isfinite(rpos.z) && isfinite(rquat.w) && isfinite(rquat.x) &&
                 isfinite(rquat.y) && isfinite(rquat.z));
#pragma unroll
    for (int j = 0; j < 6; ++j) bad = bad || !isfinite(rvel.a[j]);
    for (int k = 0; k < n1; ++k) {
        const int i = M.list1[role][k];
        bad = bad || !(isfinite(qv[i]) && isfinite(qdv[i]));
    }
    if (bad) xbad[lane] = 1;
    role_barrier(nthr);
    bad = xbad[lane] != 0;
    if (!valid) return;
    if (bad && !M.fixed_base) {
        rpos = {0.f, 0.f, -1.f};
        rquat = {1.f, 0.f, 0.f, 0.f};
        rvel = {{0.f, 0.f, 0.f, 0.f, 0.f, 0.f}};
    }
    // role 0 stores the root, the trunk joints and its chains; the other roles their chains
    for (int k = 0; k < n1; ++k) {
        const int i = M.list1[role][k];
        if (role != 0 && (M.rflags[i] & kRTrunk)) continue;
        const int d = M.dof[i];
        A.q[env * D + d] = bad ? 0.f : qv[i];
        A.qd[env * D + d] = bad ? 0.f : qdv[i];
    }
    if (role != 0) return;
    if (bad) {
        if (A.sens != nullptr)
            for (int k = 0; k < M.num_sensors * 6; ++k) A.sens[env * M.num_sensors * 6 + k] = 0.f;
    }
    A.pos[env * 3] = rpos.x; A.pos[env * 3 + 1] = rpos.y; A.pos[env * 3 + 2] = rpos.z;
    *reinterpret_cast<float4*>(A.quat + env * 4) = make_float4(rquat.w, rquat.x, rquat.y, rquat.z);
    float* vv = A.vel + env * 6;
    vv[0] = rvel.a[3]; vv[1] = rvel.a[4]; vv[2] = rvel.a[5];
    vv[3] = rvel.a[0]; vv[4] = rvel.a[1]; vv[5] = rvel.a[2];
}
