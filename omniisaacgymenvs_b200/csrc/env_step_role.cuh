// K2-F `env_step_role_kernel` -- the whole VecEnvRLGames.step of a locomotion task in ONE launch:
//   prologue  action clamp + masked reset_idx                (envs/vec_env_rlgames.py:57, tasks/shared/locomotion.py:103-145)
//   main      `substeps` articulation substeps                (vec_env_rlgames.py:64-66; PhysX stand-in)
//   epilogue  progress += 1, observations, reward, done       (tasks/base/rl_task.py:242, locomotion.py:190-321)
// (Included by articulation_step.cu inside its anonymous namespace: same device functions as K2-W / K2-T.)
// The prologue and the epilogue are switched by run-time flags of the SAME instantiation, so the physics of the fused
// step and of a stand-alone oige_articulation_step_f32 call are the same machine code (bit-identical results); the
// epilogue evaluates locomotion_math.cuh, the arithmetic K1 evaluates, on the tile it just staged for the write-back.
//
// Mapping.  lane = env, warp = ROLE, a GROUP = 32 envs x `nroles` warps, a CTA = EPC/32 groups.  The tree is cut on the
// host into a TRUNK (root + every link with a branching link at or below it; must be a path) and the plain CHAINS
// hanging off it, dealt to the roles; role 0 also owns the trunk.  Per substep (named barriers of the group; the
// self-collision variant adds the pair phase between the kinematics and the inertia sweeps, see below):
//   S0  role 0: kinematics of the trunk -> shared memory (xK)                                        | barrier K
//   S1  every role, chain by chain: kinematics head -> tip (frames into the chain's link slots), then the articulated
//       inertia tip -> head in registers; each link leaves (s, U/d, u''') in its slot; the head's I^a, p^a are added to
//       the role's hand-over slot (xI).  A role with a slot on trunk link t may also take t's OWN inertia / bias /
//       contacts off role 0's hands.                  | barriers H_t: one per trunk link t with hand-over slots -- role 0
//       waits at t only for the roles that hand over at t, which arrive (bar.arrive) and go on to barrier A
//   S2  role 0: trunk tip -> root (own inertia + hand-over slots + carry), 6x6 Cholesky at the root, trunk accelerations
//       root -> tip, trunk joints and root integrated, accelerations of the attach links published (xA) | barrier A
//   S3  every role: accelerations down its chains -- 13 FMAs per link: qdd = u''' - (U/d).a', a' += s qdd -- and joint
//       integration; the foot-sensor wrench in the last substep.
// Everything a link must remember between the sweeps lives in SHARED memory, 13 floats per link:
//   after the kinematics sweep   quat (4), p (3)              -- v, b are rebuilt on the way up (v_parent = v - s qd)
//   after the inertia sweep      s (6), U/d (6), u''' (1)     -- with a' = a - b (acceleration minus the velocity-
//                                                                product path sum) the downward sweep needs nothing else
// and the joint state q, qd, effort (target) of a role's <= MAXC chain links stays in REGISTERS (select chains instead
// of dynamic indexing).  No per-thread local memory, hence none of the L1-miss / DRAM spill traffic of K2-T / the
// former K2-R (profiles/r1_ncu_k2r_16384.txt: 37.8 MB of local-memory writes per launch).
constexpr int kRL = 24;      // links
constexpr int kRPos = 8;     // chain links per role (>= MAXC)
constexpr int kRT = 4;       // trunk links incl. the root
constexpr int kRSlots = 8;   // hand-over slots
constexpr int kRSeg = 8;     // segments per role
constexpr int kSegTrunkKin = 0, kSegOwn = 1, kSegChain = 2, kSegTrunkUp = 3, kSegPairs = 4;
constexpr int kSegBarBefore = 1, kSegBarAfter = 2, kSegSlotFirst = 4, kSegSensor = 8, kSegKinOnly = 16, kSegUpOnly = 32;
constexpr int kSegArrive = 64;  // after the sweep: arrive at the hand-over barrier of the segment's attach trunk link
constexpr int kSCMax = 96;   // self-collision link pairs (Humanoid: 66), each listed ONCE, with the role that evaluates it
constexpr int kKin = 19;     // quat 4, p 3, v 6, b 6
constexpr int kPu = 13;      // s 6, U/d 6, u'''
constexpr int kHand = 27;    // I^a 21, p^a 6

struct alignas(16) RoleModel {
    int num_links, num_sensors, fixed_base, nroles;
    int ntrunk, nslots, npub, group_floats;
    int oI, oA, oK, oT, oS, oM, oL, maxc;   // float offsets (x32 lanes) of the regions of a group's scratch
    float gravity[4];
    float k_c, d_c, mu, c_visc, k_lim, d_lim, max_qd, ang_damp;
    float max_qd_lin, max_w, pen_cap_v, k_self, d_self;
    int has_sc, pad_i[2];
    float jpos[3][kRL];
    float jquat[4][kRL];
    float axis[3][kRL];
    float mass[kRL];
    float com[3][kRL];
    float inertia[6][kRL];
    float armature[kRL], damping[kRL], stiffness[kRL], sref[kRL], lo[kRL], hi[kRL];
    float drive_kp[kRL], drive_kd[kRL], drive_fmax[kRL];
    float sph_pos[3][OIGE_MAX_SPHERES];
    float sph_r[OIGE_MAX_SPHERES];
    unsigned char dof[kRL], jtype[kRL], sph_start[kRL], sph_count[kRL], massless[kRL];
    signed char sensor_of_link[kRL];
    unsigned char npos[4], nseg[4];
    unsigned char plink[4][kRPos];                        // chain links of role r by position (chains concatenated head -> tip)
    // the role's segments, in execution order (see the kernel): kind, link range (positions for a chain, trunk
    // indices otherwise), flags, attach trunk link, hand-over slot
    unsigned char seg_kind[4][kRSeg], seg_first[4][kRSeg], seg_last[4][kRSeg], seg_flag[4][kRSeg];
    unsigned char seg_att[4][kRSeg], seg_slot[4][kRSeg];
    unsigned char tlink[kRT], town[kRT], tpub[kRT];      // trunk idx -> link, own inertia handed to a role?, xA index (255: none)
    unsigned char ts_begin[kRT + 1], ts_slot[kRSlots];   // hand-over slots attached to trunk link t
    // ---- self-collision (enable_self_collisions): every link pair is listed once, in the list of ONE of the two roles
    // that own its links (a = a link of that role, b = its partner); the other role learns of a contact through the
    // evaluating role's hit word (xH) ----
    unsigned char sc_off[8];                              // role r evaluates entries [sc_off[r], sc_off[r + 1]) (<= 32)
    unsigned char sc_a[kSCMax], sc_b[kSCMax];             // (bit 7 of sc_b: both links belong to the evaluating role)
    unsigned char w_idx[kRL];                             // link -> entry of its role's wrench accumulator
    unsigned short rec_off[kRL];                          // link -> float offset (x32 lanes) of its (quat, p, v) record
    float brad[kRL];                                      // bounding radius about the midpoint of its capsule axis
    unsigned short ctr_off[kRL];                          // link -> float offset (x32 lanes) of its bounding centre (3 floats)
    unsigned char cap_idx[kRL];                           // link -> its capsule's slot in the published end points
    // hand-over barriers: role 0 waits at trunk link t only for the roles with a slot on t (hb_id[t]: 0 none, else the
    // barrier's offset from the group's first id; hb_cnt[t]: threads taking part, role 0's warp included)
    unsigned char hb_id[kRT], hb_cnt[kRT];
    unsigned char pad[7];
    int oH, pad_h[3];                                     // float offset (x32 lanes) of the roles' hit words
    float cap_a[3][kRL], cap_b[3][kRL], cap_r[kRL];       // self-collision capsule of a link (link frame); r = 0: none
    float cap_il2[kRL];                                   // 1 / |b - a|^2 (0 for a sphere)
    // broad phase of entry k: x = byte offset (lane 0) of a's bounding centre << 16 | b's, y = bits of
    // (brad[a] + brad[b])^2; zero-filled past the last entry (a threshold of 0 never passes), 4 entries of slack
    uint2 sc_brd[kSCMax + 4];
    // narrow phase of entry k, links in canonical order lo < hi: x = byte offset of lo's capsule end points << 16 | hi's,
    // y, z = bits of 1 / length^2 of lo's, hi's axis, w = bits of (r_lo + r_hi)^2
    uint4 sc_nar[kSCMax];
    unsigned sc_to[4][4];                                 // [src role][dst role]: entries of src's list whose partner link is dst's
};
static_assert(sizeof(RoleModel) % 16 == 0, "RoleModel must be int4-copyable");

struct FArgs {
    float* pos; float* quat; float* vel; float* q; float* qd; float* sens;
    const float* actions;   // [N,D] what the efforts are formed from (fused step: the clamped copy in actions_out)
    const float* targets;   // [N,D] PD position targets or null
    long long n;
    int substeps;
    float dt;
    int D, S, OBS;
    int pro, epi, use_bulk, bump;
    // prologue (masked reset + clamp)
    const float* actions_in; float* actions_out; float clip;
    const float* init_pos; const float* init_rot; const float* init_q;
    const float* rand_pos; const float* rand_vel;
    unsigned long long seed, counter;
    unsigned long long* counter_dev;   // [2]: Philox reset counter, CTA ticket (see OigeTaskBuffers.reset_counter)
    long long env_id_offset;
    // epilogue
    float* obs; float* rew; float* pot; float* prev_pot; float* up; float* head;
    long long* reset; long long* progress;
    double* stats;
    float* ep_ret;   // [N] running return of the current episode (OigeTaskBuffers.episode_return) or null
    // host-buffer step, zero-copy mode: reward and reset flags are ALSO stored straight into the caller's pinned host
    // buffers (device-mapped pointers; null otherwise), so no copy-engine transfer follows the kernel
    float* rew_host; long long* reset_host;
};

// g_nonfinite_envs (envs parked by the non-finite guard since load) is defined in articulation_step.cu: K2-W / K2-T count too

// Phase timing (diagnostic build only, -DOIGE_STAMPS -> tools/k2f_stamps.py): (tag, clock64) pairs of group 0 of CTA 0.
#ifdef OIGE_STAMPS
__device__ long long g_stamps[4][128][2];
#define STAMP(tag)                                                                                         \
    do {                                                                                                   \
        if (blockIdx.x == 0 && grp == 0 && lane == 0 && stamp_i < 128) {                                   \
            g_stamps[role][stamp_i][0] = (tag);                                                            \
            g_stamps[role][stamp_i][1] = clock64();                                                        \
            ++stamp_i;                                                                                     \
        }                                                                                                  \
    } while (0)
#else
#define STAMP(tag) do { } while (0)
#endif

__device__ __forceinline__ void group_barrier(int id, int nthreads) {
    asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}
__device__ __forceinline__ void group_arrive(int id, int nthreads) {  // producer side: does not wait
    asm volatile("bar.arrive %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

// register arrays with a run-time index: select chains keep them in registers (no local memory)
template <int N>
__device__ __forceinline__ float rsel(const float (&r)[N], int k) {
    float v = r[0];
#pragma unroll
    for (int j = 1; j < N; ++j) v = (k == j) ? r[j] : v;
    return v;
}
template <int N>
__device__ __forceinline__ void rput(float (&r)[N], int k, float v) {
#pragma unroll
    for (int j = 0; j < N; ++j) r[j] = (k == j) ? v : r[j];
}

struct KinCtx {  // frame, spatial velocity and velocity-product path sum of the link the sweep stands on
    Q4 q;
    M3 R;
    V3 p;
    V6 v, b;
};

// parent context -> this link's context (same arithmetic as pass 1 of K2-T)
template <class MT>
__device__ __forceinline__ void kin_step(const MT& M, int i, float qj, float qdj, KinCtx& c) {
    const V3 axis = {M.axis[0][i], M.axis[1][i], M.axis[2][i]};
    const bool prismatic = M.jtype[i] == OIGE_JOINT_PRISMATIC;
    float sh, ch;
    sincos_half_angle(prismatic ? 0.f : 0.5f * qj, &sh, &ch);
    const Q4 ql = qmul(Q4{M.jquat[0][i], M.jquat[1][i], M.jquat[2][i], M.jquat[3][i]},
                       Q4{ch, sh * axis.x, sh * axis.y, sh * axis.z});
    const Q4 qi = qmul(c.q, ql);
    V3 p = c.p + mulv(c.R, V3{M.jpos[0][i], M.jpos[1][i], M.jpos[2][i]});
    const M3 R = qmat(qi);
    const V3 aw = mulv(R, axis);
    if (prismatic) p = p + qj * aw;
    const V6 s = joint_subspace(prismatic, aw, p);
    V6 v;
#pragma unroll
    for (int k = 0; k < 6; ++k) v.a[k] = fmaf(s.a[k], qdj, c.v.a[k]);
    const V6 cc = velocity_product(v, s, qdj);
#pragma unroll
    for (int k = 0; k < 6; ++k) c.b.a[k] += cc.a[k];
    c.q = qi; c.R = R; c.p = p; c.v = v;
}

// own spatial inertia, bias force and linearly-implicit ground contacts of link i (zero for a massless,
// sphere-less intermediate link of a multi-dof joint)
template <class MT>
__device__ __forceinline__ void own_inertia(const MT& M, int i, const M3& R, V3 p, const V6& v, const V6& b,
                                            float root_z, float dt, V3 grav, S6& IA, V6& pA) {
    if (M.massless[i]) {
#pragma unroll
        for (int j = 0; j < 21; ++j) IA.a[j] = 0.f;
#pragma unroll
        for (int j = 0; j < 6; ++j) pA.a[j] = 0.f;
    } else {
        link_inertia_bias(M, i, R, p, v, grav, IA, pA);
        link_contacts(M, i, R, p, v, b, root_z, dt, IA, pA);
    }
}

// One joint of the bottom-up sweep: (I^A, p^A) of link i -> (I^a, p^a) seen by its parent; leaves (s, U/d, u''') in
// `pu` (stride 32) and turns (v, b) into the parent's.
template <class MT>
__device__ __forceinline__ void joint_reduce(const MT& M, int i, const M3& R, V3 p, V6& v, V6& b, float qj, float qdj,
                                             float eff, bool has_tgt, float tgt, float dt, S6& IA, V6& pA, float* pu) {
    const V3 axis = {M.axis[0][i], M.axis[1][i], M.axis[2][i]};
    const V6 s = joint_subspace(M.jtype[i] == OIGE_JOINT_PRISMATIC, mulv(R, axis), p);
    const V6 c = velocity_product(v, s, qdj);
    float tau0, dextra;
    joint_force(M, i, qj, qdj, eff, dt, has_tgt, tgt, tau0, dextra);
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
    // downward sweep: qdd = (u - U.(a_parent + c)) / d with a = a' + b (b = path sum of the c's), i.e.
    // qdd = u''' - (U/d).a'_parent, u''' = (u - U.b_i) / d
#pragma unroll
    for (int j = 0; j < 6; ++j) {
        pu[j * 32] = s.a[j];
        pu[(6 + j) * 32] = U.a[j] * dinv;
    }
    pu[12 * 32] = (u - dot6(U, b)) * dinv;
#pragma unroll
    for (int j = 0; j < 6; ++j) {
        v.a[j] = fmaf(-s.a[j], qdj, v.a[j]);
        b.a[j] -= c.a[j];
    }
}

// ---- self-collision (enable_self_collisions, cfg/task/Humanoid.yaml:80) -------------------------------------------------
// Explicit capsule-capsule penalty between the links of a collision pair (one capsule per body), evaluated at the state
// the substep starts from (an implicit pair force would couple two branches of the tree, i.e. close a kinematic loop,
// and break the O(n) recursion).  Every pair is TESTED once, by the role whose list holds it (broad phase on the bounding
// spheres, narrow phase capsule against capsule); the pairs found in contact are published as one hit word per role, and
// each of the two roles that own a link of such a pair evaluates the contact for ITS link -- the same arithmetic on the
// same operands (segment order fixed by the link numbers), so the two forces are equal and opposite to the bit.  One
// contact per pair at the closest points of the two segments, force along their connection at the midpoint of the overlap:
// f = k pen - d vn (damping while approaching only), no friction.

// world end points of link `link`'s capsule -> C[slot][P xyz, Q xyz] (stride 32), its midpoint (centre of the bounding
// sphere of radius M.brad) -> ctr[xyz]
template <class MT>
__device__ __forceinline__ void publish_capsule(const MT& M, int link, const M3& R, V3 p, float* C, float* ctr) {
    if (!(M.cap_r[link] > 0.f)) return;
    const V3 P = p + mulv(R, V3{M.cap_a[0][link], M.cap_a[1][link], M.cap_a[2][link]});
    const V3 Q = p + mulv(R, V3{M.cap_b[0][link], M.cap_b[1][link], M.cap_b[2][link]});
    float* c = C + (int)M.cap_idx[link] * 6 * 32;
    c[0] = P.x; c[32] = P.y; c[64] = P.z; c[96] = Q.x; c[128] = Q.y; c[160] = Q.z;
    ctr[0] = 0.5f * (P.x + Q.x); ctr[32] = 0.5f * (P.y + Q.y); ctr[64] = 0.5f * (P.z + Q.z);
}
// Closest points of the capsule axes of a pair, links in canonical order lo < hi (Ericson, Real-Time Collision Detection
// 5.1.9), branch-free.  `nar` = the pair's narrow-phase record (RoleModel::sc_nar), Cb = this lane's column of the
// published end points.  Returns the squared distance; c1 on lo's axis, c2 on hi's.
__device__ __forceinline__ float capsule_closest(const char* Cb, uint4 nar, V3& c1, V3& c2) {
    const float* s1 = reinterpret_cast<const float*>(Cb + (nar.x >> 16));
    const float* s2 = reinterpret_cast<const float*>(Cb + (nar.x & 0xffffu));
    const V3 P1 = {s1[0], s1[32], s1[64]}, Q1 = {s1[96], s1[128], s1[160]};
    const V3 P2 = {s2[0], s2[32], s2[64]}, Q2 = {s2[96], s2[128], s2[160]};
    const V3 d1 = Q1 - P1, d2 = Q2 - P2, r = P1 - P2;
    const float a = dot(d1, d1), e = dot(d2, d2), f = dot(d2, r), c = dot(d1, r), b = dot(d1, d2);
    const float ia = __uint_as_float(nar.y), ie = __uint_as_float(nar.z);  // 1 / length^2: 0 for a sphere (zero-length axis)
    const float den = a * e - b * b;
    float sp = den > 1e-9f * a * e ? __saturatef((b * f - c * e) / den) : 0.f;
    float tp = (b * sp + f) * ie;
    if (tp < 0.f || ie == 0.f) sp = __saturatef(-c * ia);
    else if (tp > 1.f) sp = __saturatef((b - c) * ia);
    tp = __saturatef(tp);
    c1 = P1 + sp * d1;
    c2 = P2 + tp * d2;
    const V3 d = c1 - c2;
    return dot(d, d);
}
// The contact of the pair (a, b), list entry `nar`: force on link a (and, BOTH: the opposite one on link b, which this
// role owns too) about O -> W (first touch stores, later ones add).
template <class MT, int NW>
__device__ __forceinline__ void capsule_contact(const MT& M, const char* Cb, const float* G0, uint4 nar, int a, int b,
                                                float dt, bool both, float (&W)[NW][6], unsigned& wmask) {
    const bool a_lo = a < b;
    const int lo = a_lo ? a : b, hi = a_lo ? b : a;
    V3 clo, chi;
    const float d2 = capsule_closest(Cb, nar, clo, chi);
    const float rlo_ = M.cap_r[lo], rhi_ = M.cap_r[hi];
    if (!(d2 > 1e-18f)) return;
    const float dist = sqrtf(d2), pen = (rlo_ + rhi_) - dist;
    if (!(pen > 0.f)) return;
    const V3 n = (1.f / dist) * (clo - chi);  // from the higher-numbered link's capsule to the lower one's
    const V3 pc = chi + (rhi_ - 0.5f * pen) * n;
    const float* rlo = G0 + (int)M.rec_off[lo] * 32;
    const float* rhi = G0 + (int)M.rec_off[hi] * 32;
    const V3 wlo = {rlo[7 * 32], rlo[8 * 32], rlo[9 * 32]}, ulo = {rlo[10 * 32], rlo[11 * 32], rlo[12 * 32]};
    const V3 whi = {rhi[7 * 32], rhi[8 * 32], rhi[9 * 32]}, uhi = {rhi[10 * 32], rhi[11 * 32], rhi[12 * 32]};
    const V3 vrel = (ulo + cross(wlo, pc)) - (uhi + cross(whi, pc));
    const float vn = dot(vrel, n);
    const float f = M.k_self * fminf(pen, M.pen_cap_v * dt) - (vn < 0.f ? M.d_self * vn : 0.f);
    if (!(f > 0.f)) return;
    const V3 F = (a_lo ? f : -f) * n;  // on link a
    const V3 tq = cross(pc, F);
    const float wr[6] = {tq.x, tq.y, tq.z, F.x, F.y, F.z};
    // (W lives in local memory, and the L1 left beside 220 KB of shared memory is small: a first touch must not load)
    {
        const int w = M.w_idx[a];
        if ((wmask >> w) & 1u) {
#pragma unroll
            for (int k = 0; k < 6; ++k) W[w][k] += wr[k];
        } else {
#pragma unroll
            for (int k = 0; k < 6; ++k) W[w][k] = wr[k];
        }
        wmask |= 1u << w;
    }
    if (both) {
        const int w = M.w_idx[b];
        if ((wmask >> w) & 1u) {
#pragma unroll
            for (int k = 0; k < 6; ++k) W[w][k] -= wr[k];
        } else {
#pragma unroll
            for (int k = 0; k < 6; ++k) W[w][k] = -wr[k];
        }
        wmask |= 1u << w;
    }
}

// SREG: the sensor link's frame waits in registers (Humanoid: shared memory is full) or in shared memory (xS; the
// four-role variants run at 128 registers per thread)
// SC: the model has self-collision pairs (per substep three more group barriers: capsules published | hit words | pair forces)
template <int MAXC, int NR, bool TGT, int EPC, int MINB, bool SREG, bool SC>
__global__ void __launch_bounds__(NR * EPC, MINB)
env_step_role_kernel(const RoleModel* __restrict__ gmodel, const __grid_constant__ OigeTaskParams P,
                     const __grid_constant__ FArgs A) {
    __shared__ __align__(16) RoleModel M;
    __shared__ __align__(8) uint64_t s_bar;
    __shared__ unsigned long long s_counter;
    extern __shared__ __align__(128) float dynsm[];
    constexpr int NT = NR * EPC;
    constexpr int GT = NR * 32;  // threads of a group
    const int tid = threadIdx.x;
    // (warp w = 3 grp + role runs on scheduler w mod 4, so the four role-0 warps of a 128-env CTA -- the critical path --
    // sit on four different schedulers.  Dealing the warps so that a scheduler holds ONE role, for its private L0
    // instruction cache, was measured and is worse: 92.8 -> 96.5 us; three always-busy warps then share an issue port.)
    const int grp = tid / GT, role = (tid / 32) % NR, lane = tid & 31;
    const int bar_id = 1 + grp * 3;  // + 1, + 2: the group's hand-over barriers
    const int D = A.D;
    const long long e0 = (long long)blockIdx.x * EPC;
    const int cnt = (int)((A.n - e0) < (long long)EPC ? (A.n - e0) : (long long)EPC);
    const bool bulk = A.use_bulk && cnt == EPC;
    const int e_raw = grp * 32 + lane;
    const bool valid = e_raw < cnt;
    const int e = valid ? e_raw : cnt - 1;  // idle lanes shadow the tile's last env and store nothing
    const long long env = e0 + e;
    const float dt = A.dt;

    // ================= stage the tile: q, qd, actions, root state (+ reset flags) ==========================
    float* s_q = dynsm;
    float* s_qd = s_q + EPC * D;
    float* s_act = s_qd + EPC * D;
    float* s_pos = s_act + EPC * D;
    float* s_quat = s_pos + EPC * 3;
    float* s_vel = s_quat + EPC * 4;
    long long* s_flag = reinterpret_cast<long long*>(s_vel + EPC * 6);  // [EPC] reset_buf rows (fused step only)
    const float* act_src = A.pro ? A.actions_in : A.actions;
    pdl_wait();  // state / actions may be written by a preceding kernel
    // The tile's bulk loads are issued FIRST, so that the model-table copy below, the reset counter and the ticket
    // overlap their DRAM round trip instead of queueing behind it (the step starts with a cold L2 in the benchmark).
    if (bulk && tid == 0) {
        mbar_init(&s_bar, 1);
        mbar_fence_init();
        mbar_expect_tx(&s_bar, (uint32_t)EPC * ((3u * D + 13u) * 4u + (A.pro ? 8u : 0u)));
        bulk_g2s(s_q, A.q + e0 * D, EPC * D * 4, &s_bar);
        bulk_g2s(s_qd, A.qd + e0 * D, EPC * D * 4, &s_bar);
        bulk_g2s(s_act, act_src + e0 * D, EPC * D * 4, &s_bar);
        bulk_g2s(s_pos, A.pos + e0 * 3, EPC * 12, &s_bar);
        bulk_g2s(s_quat, A.quat + e0 * 4, EPC * 16, &s_bar);
        bulk_g2s(s_vel, A.vel + e0 * 6, EPC * 24, &s_bar);
        if (A.pro) bulk_g2s(s_flag, A.reset + e0, EPC * 8, &s_bar);
    }
    if (tid == 0) {
        // Philox reset counter of this step.  In a captured (CUDA-graph) step it lives in device memory: every CTA
        // reads it, THEN takes a ticket; the CTA that draws the last ticket knows everyone has read it and advances it
        // for the next replay -- here, off the critical path, not at the end of the kernel.
        unsigned long long c = A.counter;
        if (A.pro && A.counter_dev != nullptr) {
            c = *reinterpret_cast<volatile unsigned long long*>(A.counter_dev);
            if (A.bump) {
                __threadfence();
                const unsigned long long t = atomicAdd(A.counter_dev + 1, 1ull);
                if (t == (unsigned long long)gridDim.x - 1) {
                    A.counter_dev[1] = 0ull;
                    A.counter_dev[0] = c + 1ull;
                }
            }
        }
        s_counter = c;
    }
    {
        const int4* src = reinterpret_cast<const int4*>(gmodel);
        int4* dst = reinterpret_cast<int4*>(&M);
        for (int i = tid; i < (int)(sizeof(RoleModel) / 16); i += NT) dst[i] = src[i];
    }
    if (!bulk) {
        _Pragma("unroll 1") for (int i = tid; i < cnt * D; i += NT) { s_q[i] = A.q[e0 * D + i]; s_qd[i] = A.qd[e0 * D + i]; s_act[i] = act_src[e0 * D + i]; }
        _Pragma("unroll 1") for (int i = tid; i < cnt * 3; i += NT) s_pos[i] = A.pos[e0 * 3 + i];
        _Pragma("unroll 1") for (int i = tid; i < cnt * 4; i += NT) s_quat[i] = A.quat[e0 * 4 + i];
        _Pragma("unroll 1") for (int i = tid; i < cnt * 6; i += NT) s_vel[i] = A.vel[e0 * 6 + i];
        if (A.pro) { _Pragma("unroll 1") for (int i = tid; i < cnt; i += NT) s_flag[i] = A.reset[e0 + i]; }
    }
    __syncthreads();  // model table staged, barrier initialised, counter published
    if (bulk) mbar_wait(&s_bar, 0);
    pdl_trigger();
    if (A.pro) {
        // VecEnvRLGames.step's torch.clamp(actions, -clip, clip) (envs/vec_env_rlgames.py:57): the clamped copy the
        // task keeps (locomotion.py:108) goes to actions_out, coalesced; the tile copy feeds the efforts below
        const float c = A.clip;
        _Pragma("unroll 1") for (int i = tid; i < cnt * D; i += NT) {
            const float v = s_act[i];
            const float w = v < -c ? -c : (v > c ? c : v);
            s_act[i] = w;
            A.actions_out[e0 * D + i] = w;
        }
        // masked reset_idx (locomotion.py:116-145; same arithmetic as reset_idx.cu) applied to the TILE: one warp per
        // flagged env, lane = dof.  A random policy flags ~4 % of the envs per step; every warp ballots the same flags
        // and takes every NW-th flagged env, so the rows' DRAM round trips (initial state) overlap across the warps.
        const unsigned long long rcounter = s_counter;
        int taken = 0;
        for (int base = 0; base < cnt; base += 32) {
            const int f0 = base + lane;
            unsigned m = __ballot_sync(0xffffffffu, f0 < cnt && s_flag[f0] != 0);
            for (; m; m &= m - 1, ++taken) {
                if (taken % (NT / 32) != (tid >> 5)) continue;
                const int f = base + __ffs(m) - 1;
                const long long genv = e0 + f;
                if (lane < D) {
                    const uint64_t gid = (uint64_t)(A.env_id_offset + genv);
                    const float up = A.rand_pos ? A.rand_pos[genv * D + lane] : reset_uniform(A.seed, rcounter, gid, lane, 0);
                    const float uv = A.rand_vel ? A.rand_vel[genv * D + lane] : reset_uniform(A.seed, rcounter, gid, lane, 1);
                    const float noise = __fadd_rn(__fmul_rn(0.2f - (-0.2f), up), -0.2f);          // locomotion.py:120
                    float qv = __fadd_rn(A.init_q[genv * D + lane], noise);
                    qv = fmaxf(fminf(qv, P.dof_limits_upper[lane]), P.dof_limits_lower[lane]);    // locomotion.py:121-123
                    s_q[f * D + lane] = qv;
                    s_qd[f * D + lane] = __fadd_rn(__fmul_rn(0.1f - (-0.1f), uv), -0.1f);         // locomotion.py:124
                }
                if (lane < 3) s_pos[f * 3 + lane] = A.init_pos[genv * 3 + lane];
                if (lane < 4) s_quat[f * 4 + lane] = A.init_rot[genv * 4 + lane];
                if (lane < 6) s_vel[f * 6 + lane] = 0.0f;
                if (lane == 0) {
                    // potentials = -|| (targets - initial_root_pos)_xy || / dt         (locomotion.py:136-139)
                    const float tx = __fsub_rn(P.targets[0], A.init_pos[genv * 3]);
                    const float ty = __fsub_rn(P.targets[1], A.init_pos[genv * 3 + 1]);
                    const float nrm = sqrtf(fmaf(0.0f, 0.0f, fmaf(ty, ty, __fmul_rn(tx, tx))));
                    const float pp = __fdiv_rn(-nrm, P.dt);
                    A.prev_pot[genv] = pp;
                    A.pot[genv] = pp;
                    A.reset[genv] = 0;     // locomotion.py:142-143
                    A.progress[genv] = 0;
                }
            }
        }
        __syncthreads();
    }
    // joint state of this role's chain links -> registers (position k in the role's list); the ACTION is kept, the
    // effort a * gear * power_scale (locomotion.py:109) is formed where it is used
    const int np = M.npos[role];
    float qr[MAXC], qdr[MAXC], ar[MAXC], tgtr[TGT ? MAXC : 1];
    float tq0[kRT - 1], tq1[kRT - 1], tq2[kRT - 1], tq3[kRT - 1];  // role 0: trunk q, qd, action, target on their way to xT
#pragma unroll
    for (int k = 0; k < MAXC; ++k) {
        qr[k] = 0.f; qdr[k] = 0.f; ar[k] = 0.f;
        if (TGT) tgtr[k] = 0.f;
        if (k < np) {
            const int d = M.dof[M.plink[role][k]];
            qr[k] = s_q[e * D + d]; qdr[k] = s_qd[e * D + d]; ar[k] = s_act[e * D + d];
            if (TGT) tgtr[k] = A.targets != nullptr ? A.targets[env * D + d] : 0.f;
        }
    }
    V3 rpos = {0.f, 0.f, 0.f};
    Q4 rquat = {1.f, 0.f, 0.f, 0.f};
    V6 rvel = {{0.f, 0.f, 0.f, 0.f, 0.f, 0.f}};
    if (role == 0) {
#pragma unroll
        for (int t = 1; t < kRT; ++t) {
            tq0[t - 1] = 0.f; tq1[t - 1] = 0.f; tq2[t - 1] = 0.f; tq3[t - 1] = 0.f;
            if (t < M.ntrunk) {
                const int d = M.dof[M.tlink[t]];
                tq0[t - 1] = s_q[e * D + d]; tq1[t - 1] = s_qd[e * D + d]; tq2[t - 1] = s_act[e * D + d];
                tq3[t - 1] = (TGT && A.targets != nullptr) ? A.targets[env * D + d] : 0.f;
            }
        }
        rpos = {s_pos[e * 3], s_pos[e * 3 + 1], s_pos[e * 3 + 2]};
        const float4 t = *reinterpret_cast<const float4*>(s_quat + e * 4);
        rquat = {t.x, t.y, t.z, t.w};
        const float* vv = s_vel + e * 6;
        rvel = {{vv[3], vv[4], vv[5], vv[0], vv[1], vv[2]}};
    }
    __syncthreads();  // the staging tile is dead from here on: the scratch regions below alias it

    // ================= scratch of this group ============================================================
    float* const G0 = dynsm + (size_t)grp * M.group_floats * 32 + lane;
    float* const xI = G0 + M.oI * 32;   // [nslots][27]   hand-over: I^a, p^a of the chains (+ own inertia of a trunk link)
    float* const xA = G0 + M.oA * 32;   // [npub][6]      a' of the trunk links that chains attach to
    float* const xK = G0 + M.oK * 32;   // [ntrunk][19]   trunk kinematics; rows t >= 1 become (s, U/d, u3) in the trunk sweep
    float* const xT = G0 + M.oT * 32;   // [ntrunk-1][4]  trunk q, qd, action, target
    float* const xS = G0 + M.oS * 32;   // [nsensors][7]  frame of the sensor links (last substep; !SREG)
    float* const xM = G0 + M.oM * 32;   // [2]            root height, non-finite flag
    float* const LS = G0 + (M.oL + role * MAXC * kPu) * 32;  // [MAXC][13] link slots of this role
    const int ntrunk = M.ntrunk;
    const bool fixed_base = M.fixed_base != 0;
    const V3 grav = {M.gravity[0], M.gravity[1], M.gravity[2]};
    const bool has_tgt = TGT && A.targets != nullptr;
    const float power = P.power_scale;
    if (role == 0) {
#pragma unroll
        for (int t = 1; t < kRT; ++t) {
            if (t < ntrunk) {
                xT[((t - 1) * 4 + 0) * 32] = tq0[t - 1];
                xT[((t - 1) * 4 + 1) * 32] = tq1[t - 1];
                xT[((t - 1) * 4 + 2) * 32] = tq2[t - 1];
                xT[((t - 1) * 4 + 3) * 32] = tq3[t - 1];
            }
        }
        xM[32] = 0.f;
    }
    float sw[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};  // sensor wrench of this role's sensor chain
    int sw_idx = -1;
    V6 vatt = {{0.f, 0.f, 0.f, 0.f, 0.f, 0.f}}, batt = {{0.f, 0.f, 0.f, 0.f, 0.f, 0.f}};
    Q4 sq = {1.f, 0.f, 0.f, 0.f};  // frame of the sensor link at the start of the last substep
    V3 sp = {0.f, 0.f, 0.f};
    const int nseg = M.nseg[role];
    // self-collision: capsule end points live in the hand-over slots' memory (dead until the first chain hands over), the
    // wrenches on this role's own links in a per-thread array that is only touched when a pair is in contact
    float* const xC = xI;
    float W[SC ? kRT + MAXC : 1][6];
    unsigned wmask = 0u;
    V6 vkeep = {{0.f, 0.f, 0.f, 0.f, 0.f, 0.f}}, bkeep = vkeep;  // velocity, velocity-product sum at the tip of the chain swept last
    int keep_last = -1;

#ifdef OIGE_STAMPS
    int stamp_i = 0;
#endif
    STAMP(1);
    for (int step = 0; step < A.substeps; ++step) {
        const bool last = (step == A.substeps - 1);
        STAMP(10);
        float root_z = 0.f;
        S6 carryI;
        V6 carryP;
        if (SC) wmask = 0u;
        // ---- S0 / S1 / S2a: the role's SEGMENTS.  One loop body serves the trunk kinematics (role 0), the chains
        // (kinematics head -> tip, then articulated inertia tip -> head), the own inertia of a trunk link taken off
        // role 0's hands, and the trunk's bottom-up sweep (role 0): ONE copy of every sweep in the instruction stream
        // (the first version inlined them two to three times and stalled on instruction fetch, ncu no_instruction 1.5
        // per issue).
        for (int sg = 0; sg < nseg; ++sg) {
            const int kind = M.seg_kind[role][sg], sflag = M.seg_flag[role][sg];
            const int first = M.seg_first[role][sg], lastl = M.seg_last[role][sg];
            STAMP(20 + sg);
            if (sflag & kSegBarBefore) { group_barrier(bar_id, GT); STAMP(30 + sg); }
            if (SC && kind == kSegPairs) {
                // ---- pair phase (every link pair is in exactly one role's list) ---------------------------------
                // 1. broad phase, warp-uniform loop over the role's list, grouped by the first link: bounding spheres
                //    about the capsule midpoints, one candidate bit per entry.
                // 2. narrow phase: every lane pops ITS OWN next candidate and runs the capsule-capsule distance on it
                //    -- lanes work on different pairs at the same time (the scratch is laid out [float][lane], so
                //    lane-dependent indices stay conflict-free); the warp iterates max-over-lanes times.  The pairs in
                //    contact go to the role's hit word xH[role].                                    | barrier
                // 3. contacts (rare, lane-divergent): the role's own hits, then the hits of the other roles' entries
                //    whose partner link is this role's (sc_to) -- the wrench on the OWN link is kept.
                const int k0 = M.sc_off[role];
                const char* const Gb = reinterpret_cast<const char*>(G0);
                const char* const Cb = reinterpret_cast<const char*>(xC);
                unsigned cand = 0u;
                {
                    // four entries per trip, loads first: the tests are independent, their latencies overlap (a trip may
                    // run into the next role's entries or the zero fill; those bits are masked off)
                    const int nk = M.sc_off[role + 1] - k0;
                    for (int k = 0; k < nk; k += 4) {
                        uint2 rec[4];
                        float ax[4], ay[4], az[4], bx[4], by[4], bz[4];
#pragma unroll
                        for (int u = 0; u < 4; ++u) rec[u] = M.sc_brd[k0 + k + u];
#pragma unroll
                        for (int u = 0; u < 4; ++u) {
                            const float* ca = reinterpret_cast<const float*>(Gb + (rec[u].x >> 16));
                            const float* cb = reinterpret_cast<const float*>(Gb + (rec[u].x & 0xffffu));
                            ax[u] = ca[0]; ay[u] = ca[32]; az[u] = ca[64];
                            bx[u] = cb[0]; by[u] = cb[32]; bz[u] = cb[64];
                        }
#pragma unroll
                        for (int u = 0; u < 4; ++u) {
                            const float dx = ax[u] - bx[u], dy = ay[u] - by[u], dz = az[u] - bz[u];
                            if (fmaf(dz, dz, fmaf(dy, dy, dx * dx)) < __uint_as_float(rec[u].y)) cand |= 1u << (k + u);
                        }
                    }
                    cand &= nk >= 32 ? 0xffffffffu : (1u << nk) - 1u;
                }
                STAMP(40);
                // two candidates per trip (the two distance computations are independent dependency chains)
                unsigned hits = 0u;
                while (__any_sync(0xffffffffu, cand != 0u)) {
                    if (cand != 0u) {
                        const int kb1 = __ffs((int)cand) - 1;
                        cand &= cand - 1u;
                        const int kb2 = cand != 0u ? __ffs((int)cand) - 1 : kb1;
                        cand &= cand - 1u;  // (0 stays 0)
                        const uint4 nar1 = M.sc_nar[k0 + kb1], nar2 = M.sc_nar[k0 + kb2];
                        V3 c1, c2, c3, c4;
                        const float da = capsule_closest(Cb, nar1, c1, c2), db = capsule_closest(Cb, nar2, c3, c4);
                        if (da < __uint_as_float(nar1.w)) hits |= 1u << kb1;
                        if (db < __uint_as_float(nar2.w)) hits |= 1u << kb2;
                    }
                }
                G0[(M.oH + role) * 32] = __uint_as_float(hits);
                STAMP(41);
                group_barrier(bar_id, GT);
                STAMP(42);
                // one pop loop over the own hits and the other roles' hits on this role's links (the warp iterates
                // max-over-lanes of the lane's TOTAL, not once per source list)
                unsigned hsrc[NR];
                hsrc[0] = hits;
#pragma unroll
                for (int dr = 1; dr < NR; ++dr) {
                    const int src = role + dr < NR ? role + dr : role + dr - NR;
                    hsrc[dr] = __float_as_uint(G0[(M.oH + src) * 32]) & M.sc_to[src][role];
                }
                for (;;) {
                    unsigned any = 0u;
#pragma unroll
                    for (int dr = 0; dr < NR; ++dr) any |= hsrc[dr];
                    if (!__any_sync(0xffffffffu, any != 0u)) break;
                    if (any != 0u) {
                        int dsel = 0;
#pragma unroll
                        for (int dr = NR - 1; dr >= 1; --dr) dsel = hsrc[dr] != 0u ? dr : dsel;
                        dsel = hsrc[0] != 0u ? 0 : dsel;
                        unsigned h = hsrc[0];
#pragma unroll
                        for (int dr = 1; dr < NR; ++dr) h = dsel == dr ? hsrc[dr] : h;
                        const int src = role + dsel < NR ? role + dsel : role + dsel - NR;
                        const int k = M.sc_off[src] + __ffs((int)h) - 1;
                        h &= h - 1u;
#pragma unroll
                        for (int dr = 0; dr < NR; ++dr) hsrc[dr] = dsel == dr ? h : hsrc[dr];
                        const int la = M.sc_a[k], lb = M.sc_b[k];
                        // (a foreign entry: this role's link is the entry's partner b)
                        capsule_contact(M, Cb, G0, M.sc_nar[k], dsel == 0 ? la : (lb & 127), dsel == 0 ? (lb & 127) : la, dt,
                                        dsel == 0 && (lb & 128) != 0, W, wmask);
                    }
                }
                STAMP(43);
                if (sflag & kSegBarAfter) group_barrier(bar_id, GT);
                STAMP(44);
                continue;
            }
            if (kind != kSegTrunkKin) root_z = xM[0];
            const bool chain = kind == kSegChain;
            const bool do_kin = kind == kSegTrunkKin || (chain && !(SC && (sflag & kSegUpOnly)));
            const bool do_up = kind != kSegTrunkKin && !(SC && (sflag & kSegKinOnly));
            V6 v, b;
            // ---- kinematics sweep: parent -> child -----------------------------------------------------------
            if (do_kin) {
                KinCtx c;
                if (chain) {
                    const float* k = xK + M.seg_att[role][sg] * kKin * 32;
                    c.q = {k[0], k[32], k[64], k[96]};
                    c.p = {k[4 * 32], k[5 * 32], k[6 * 32]};
#pragma unroll
                    for (int j = 0; j < 6; ++j) { c.v.a[j] = k[(7 + j) * 32]; c.b.a[j] = k[(13 + j) * 32]; }
                } else {
                    c.q = rquat; c.p = {0.f, 0.f, 0.f}; c.v = rvel;
                    const V3 wxv = cross(ang(rvel), lin(rvel));
                    c.b = {{0.f, 0.f, 0.f, -wxv.x, -wxv.y, -wxv.z}};
                    xM[0] = rpos.z;
                    float* k = xK;  // trunk row 0 = the root itself
                    k[0] = c.q.w; k[32] = c.q.x; k[64] = c.q.y; k[96] = c.q.z;
                    k[4 * 32] = 0.f; k[5 * 32] = 0.f; k[6 * 32] = 0.f;
#pragma unroll
                    for (int j = 0; j < 6; ++j) { k[(7 + j) * 32] = c.v.a[j]; k[(13 + j) * 32] = c.b.a[j]; }
                }
                c.R = qmat(c.q);
                if (SC && !chain) publish_capsule(M, 0, c.R, c.p, xC, G0 + (int)M.ctr_off[0] * 32);  // the root body's capsule
                for (int i = first; i <= lastl; ++i) {
                    int link;
                    float qj, qdj;
                    if (chain) { link = M.plink[role][i]; qj = rsel(qr, i); qdj = rsel(qdr, i); }
                    else { link = M.tlink[i]; qj = xT[((i - 1) * 4 + 0) * 32]; qdj = xT[((i - 1) * 4 + 1) * 32]; }
                    kin_step(M, link, qj, qdj, c);
                    float* dst = chain ? LS + i * kPu * 32 : xK + i * kKin * 32;
                    dst[0] = c.q.w; dst[32] = c.q.x; dst[64] = c.q.y; dst[96] = c.q.z;
                    dst[4 * 32] = c.p.x; dst[5 * 32] = c.p.y; dst[6 * 32] = c.p.z;
                    // (self-collision reads a body's velocity from its record; a massless intermediate link's floats
                    // 7..12 may hold the bounding centres of two bodies instead)
                    if (!chain || (SC && !M.massless[link])) {
#pragma unroll
                        for (int j = 0; j < 6; ++j) dst[(7 + j) * 32] = c.v.a[j];
                    }
                    if (!chain) {
#pragma unroll
                        for (int j = 0; j < 6; ++j) dst[(13 + j) * 32] = c.b.a[j];
                    }
                    if (SC) publish_capsule(M, link, c.R, c.p, xC, G0 + (int)M.ctr_off[link] * 32);
                }
                if (SC && chain) { vkeep = c.v; bkeep = c.b; keep_last = lastl; }
                if (chain && last && (sflag & kSegSensor)) {  // the foot's frame is needed again after the downward sweep
                    if (SREG) {
                        sq = c.q; sp = c.p;  // (registers: they only stay live across barrier waits of the chain roles)
                    } else {
                        float* f = xS + M.sensor_of_link[M.plink[role][lastl]] * 7 * 32;
                        f[0] = c.q.w; f[32] = c.q.x; f[64] = c.q.y; f[96] = c.q.z;
                        f[4 * 32] = c.p.x; f[5 * 32] = c.p.y; f[6 * 32] = c.p.z;
                    }
                }
                v = c.v; b = c.b;  // of the tip: walked back up with the inertia sweep
            }
            STAMP(50 + sg);
            // ---- articulated-inertia sweep: child -> parent, I^a / p^a carried in registers ---------------------
            if (do_up) {
                const bool trunk_up = kind == kSegTrunkUp;
                if (SC && chain) {
                    // the kinematics ran in an earlier segment: the tip's velocity and velocity-product sum are still in
                    // registers for the chain swept last, else they are rebuilt from the attach link down the stored frames
                    if (keep_last == lastl) {
                        v = vkeep; b = bkeep;
                    } else {
                        const float* k = xK + M.seg_att[role][sg] * kKin * 32;
#pragma unroll
                        for (int j = 0; j < 6; ++j) { v.a[j] = k[(7 + j) * 32]; b.a[j] = k[(13 + j) * 32]; }
                        for (int i = first; i <= lastl; ++i) {
                            const int link = M.plink[role][i];
                            const float* r2 = LS + i * kPu * 32;
                            const M3 R = qmat(Q4{r2[0], r2[32], r2[64], r2[96]});
                            const V3 p = {r2[4 * 32], r2[5 * 32], r2[6 * 32]};
                            const V3 axis = {M.axis[0][link], M.axis[1][link], M.axis[2][link]};
                            const V6 sj = joint_subspace(M.jtype[link] == OIGE_JOINT_PRISMATIC, mulv(R, axis), p);
                            const float qdj = rsel(qdr, i);
#pragma unroll
                            for (int j = 0; j < 6; ++j) v.a[j] = fmaf(sj.a[j], qdj, v.a[j]);
                            const V6 cc = velocity_product(v, sj, qdj);
#pragma unroll
                            for (int j = 0; j < 6; ++j) b.a[j] += cc.a[j];
                        }
                    }
                }
                for (int i = lastl; i >= first; --i) {
                    const int link = chain ? M.plink[role][i] : M.tlink[i];
                    float* src = chain ? LS + i * kPu * 32 : xK + i * kKin * 32;
                    const M3 R = qmat(Q4{src[0], src[32], src[64], src[96]});
                    const V3 p = {src[4 * 32], src[5 * 32], src[6 * 32]};
                    if (!chain) {
#pragma unroll
                        for (int j = 0; j < 6; ++j) { v.a[j] = src[(7 + j) * 32]; b.a[j] = src[(13 + j) * 32]; }
                    }
                    S6 IA;
                    V6 pA;
                    if (trunk_up && M.town[i]) {  // a chain role already put this link's own inertia into its slot
#pragma unroll
                        for (int j = 0; j < 21; ++j) IA.a[j] = 0.f;
#pragma unroll
                        for (int j = 0; j < 6; ++j) pA.a[j] = 0.f;
                    } else {
                        own_inertia(M, link, R, p, v, b, root_z, dt, grav, IA, pA);
                    }
                    if (SC) {  // self-collision wrench on this link (held by the role that evaluates its pairs): an external force
                        const int w = M.w_idx[link];
                        if ((wmask >> w) & 1u) {
#pragma unroll
                            for (int j = 0; j < 6; ++j) pA.a[j] -= W[w][j];
                        }
                    }
                    if (i != lastl) {
#pragma unroll
                        for (int j = 0; j < 21; ++j) IA.a[j] += carryI.a[j];
#pragma unroll
                        for (int j = 0; j < 6; ++j) pA.a[j] += carryP.a[j];
                    }
                    if (trunk_up) {
                        // the chains (and own inertias) handed over at this trunk link: wait for THEIR roles only
                        if (M.hb_id[i]) group_barrier(bar_id + M.hb_id[i], M.hb_cnt[i]);
                        _Pragma("unroll 1")  // (ptxas unrolled this 1..3-trip loop fourfold: 580 instructions of code)
                        for (int h = M.ts_begin[i]; h < M.ts_begin[i + 1]; ++h) {
                            const float* hs = xI + M.ts_slot[h] * kHand * 32;
#pragma unroll
                            for (int j = 0; j < 21; ++j) IA.a[j] += hs[j * 32];
#pragma unroll
                            for (int j = 0; j < 6; ++j) pA.a[j] += hs[(21 + j) * 32];
                        }
                    }
                    if (chain || (trunk_up && i > 0)) {
                        float qj, qdj, aj, tj;
                        if (chain) { qj = rsel(qr, i); qdj = rsel(qdr, i); aj = rsel(ar, i); tj = TGT ? rsel(tgtr, i) : 0.f; }
                        else { const float* x = xT + (i - 1) * 4 * 32; qj = x[0]; qdj = x[32]; aj = x[64]; tj = x[96]; }
                        const float eff = aj * P.joint_gears[M.dof[link]] * power;  // locomotion.py:109
                        joint_reduce(M, link, R, p, v, b, qj, qdj, eff, has_tgt, tj, dt, IA, pA, src);
                    }
                    carryI = IA; carryP = pA;
                }
                if (!trunk_up) {  // hand the head's I^a, p^a (or a trunk link's own inertia) to the trunk
                    float* dst = xI + M.seg_slot[role][sg] * kHand * 32;
                    if (sflag & kSegSlotFirst) {
#pragma unroll
                        for (int j = 0; j < 21; ++j) dst[j * 32] = carryI.a[j];
#pragma unroll
                        for (int j = 0; j < 6; ++j) dst[(21 + j) * 32] = carryP.a[j];
                    } else {
#pragma unroll
                        for (int j = 0; j < 21; ++j) dst[j * 32] += carryI.a[j];
#pragma unroll
                        for (int j = 0; j < 6; ++j) dst[(21 + j) * 32] += carryP.a[j];
                    }
                    if (chain && last && (sflag & kSegSensor)) { vatt = v; batt = b; }  // = the attach link's v, b
                }
            }
            STAMP(60 + sg);
            if (sflag & kSegArrive) {
                const int t = M.seg_att[role][sg];
                __threadfence_block();
                group_arrive(bar_id + M.hb_id[t], M.hb_cnt[t]);
            }
            if (sflag & kSegBarAfter) { group_barrier(bar_id, GT); STAMP(70 + sg); }
        }

        // ---- S2b (role 0): root solve, trunk accelerations root -> tip, trunk joints and root integrated ----------
        if (role == 0) {
            V6 a0 = {{0.f, 0.f, 0.f, 0.f, 0.f, 0.f}};
            if (!fixed_base) chol6_solve(carryI, carryP, a0);  // a welded base does not accelerate
            V6 ap = a0;  // a' = a - b; b_root = -(0, w x v)
            {
                const V3 wxv = cross(ang(rvel), lin(rvel));
                ap.a[3] += wxv.x; ap.a[4] += wxv.y; ap.a[5] += wxv.z;
            }
            for (int t = 0; t < ntrunk; ++t) {
                if (t > 0) {
                    const float* k = xK + t * kKin * 32;
                    float acc = k[12 * 32];
#pragma unroll
                    for (int j = 0; j < 6; ++j) acc = fmaf(-k[(6 + j) * 32], ap.a[j], acc);
#pragma unroll
                    for (int j = 0; j < 6; ++j) ap.a[j] = fmaf(k[j * 32], acc, ap.a[j]);
                    float* x = xT + (t - 1) * 4 * 32;
                    float qd = fmaf(dt, acc, x[32]);
                    qd = clamp_qd(M, M.tlink[t], qd);
                    x[32] = qd;
                    x[0] = fmaf(dt, qd, x[0]);
                }
                if (M.tpub[t] != 255) {
                    float* dst = xA + M.tpub[t] * 6 * 32;
#pragma unroll
                    for (int j = 0; j < 6; ++j) dst[j * 32] = ap.a[j];
                }
            }
            if (!fixed_base) integrate_root(M, a0, dt, rpos, rquat, rvel);
        }
        STAMP(80);
        group_barrier(bar_id, GT);  // A
        STAMP(81);

        // ---- S3: accelerations down the own chains, joint integration, foot sensors ---------------------------
        for (int sg = 0; sg < nseg; ++sg) {
            if (M.seg_kind[role][sg] != kSegChain) continue;
            const int sflag = M.seg_flag[role][sg];
            if (SC && (sflag & kSegKinOnly)) continue;  // (a chain appears twice in a self-collision program)
            const int first = M.seg_first[role][sg], lastl = M.seg_last[role][sg];
            const bool sens_chain = last && (sflag & kSegSensor) && A.sens != nullptr;
            V6 ap;
            {
                const float* src = xA + M.tpub[M.seg_att[role][sg]] * 6 * 32;
#pragma unroll
                for (int j = 0; j < 6; ++j) ap.a[j] = src[j * 32];
            }
            V6 v = vatt, b = batt;
            for (int i = first; i <= lastl; ++i) {
                const float* slot = LS + i * kPu * 32;
                V6 s;
#pragma unroll
                for (int j = 0; j < 6; ++j) s.a[j] = slot[j * 32];
                float qdd = slot[12 * 32];
#pragma unroll
                for (int j = 0; j < 6; ++j) qdd = fmaf(-slot[(6 + j) * 32], ap.a[j], qdd);
#pragma unroll
                for (int j = 0; j < 6; ++j) ap.a[j] = fmaf(s.a[j], qdd, ap.a[j]);
                const float qd0 = rsel(qdr, i);
                if (sens_chain) {  // rebuild the start-of-substep v, b down to the foot
#pragma unroll
                    for (int j = 0; j < 6; ++j) v.a[j] = fmaf(s.a[j], qd0, v.a[j]);
                    const V6 cc = velocity_product(v, s, qd0);
#pragma unroll
                    for (int j = 0; j < 6; ++j) b.a[j] += cc.a[j];
                }
                float qd = fmaf(dt, qdd, qd0);
                qd = clamp_qd(M, M.plink[role][i], qd);
                rput(qdr, i, qd);
                rput(qr, i, fmaf(dt, qd, rsel(qr, i)));
            }
            if (sens_chain) {
                const int link = M.plink[role][lastl];
                sw_idx = M.sensor_of_link[link];
                if (!SREG) {
                    const float* f = xS + sw_idx * 7 * 32;
                    sq = {f[0], f[32], f[64], f[96]};
                    sp = {f[4 * 32], f[5 * 32], f[6 * 32]};
                }
                const M3 R = qmat(sq);
                const V3 p = sp;
                V6 a;
#pragma unroll
                for (int j = 0; j < 6; ++j) a.a[j] = ap.a[j] + b.a[j];
                sensor_wrench(M, link, R, p, v, b, a, xM[0], dt, sw);
            }
        }
    }

    STAMP(90);
    // ================= non-finite guard (agreed across the roles of an env) ================================
    bool bad = false;
#pragma unroll
    for (int k = 0; k < MAXC; ++k) bad = bad || !(isfinite(qr[k]) && isfinite(qdr[k]));
    float tq[kRT - 1], tqd[kRT - 1], ta[kRT - 1];
    if (role == 0) {
        bad = bad || !(isfinite(rpos.x) && isfinite(rpos.y) && isfinite(rpos.z) && isfinite(rquat.w) &&
                       isfinite(rquat.x) && isfinite(rquat.y) && isfinite(rquat.z));
#pragma unroll
        for (int j = 0; j < 6; ++j) bad = bad || !isfinite(rvel.a[j]);
#pragma unroll
        for (int t = 1; t < kRT; ++t) {
            tq[t - 1] = 0.f; tqd[t - 1] = 0.f; ta[t - 1] = 0.f;
            if (t < ntrunk) {
                tq[t - 1] = xT[((t - 1) * 4 + 0) * 32];
                tqd[t - 1] = xT[((t - 1) * 4 + 1) * 32];
                ta[t - 1] = xT[((t - 1) * 4 + 2) * 32];
                bad = bad || !(isfinite(tq[t - 1]) && isfinite(tqd[t - 1]));
            }
        }
    }
    if (bad) xM[32] = 1.f;  // (zeroed by role 0 before the first barrier K)
    group_barrier(bar_id, GT);
    bad = xM[32] != 0.f;
    if (bad && !fixed_base) {  // parked below the termination height so that the observation pass flags it for reset
        rpos = {0.f, 0.f, -1.f};
        rquat = {1.f, 0.f, 0.f, 0.f};
        rvel = {{0.f, 0.f, 0.f, 0.f, 0.f, 0.f}};
    }
    if (bad && role == 0 && valid) atomicAdd(&g_nonfinite_envs, 1ull);
    // From here on every GROUP works alone: its output tile (32 envs) aliases its OWN scratch, so a group that is done
    // with its substeps does not wait for the CTA's slowest group (4.7 k cycles at 16384 Humanoids) -- group barriers
    // only, one lane per group issues the group's bulk stores.
    group_barrier(bar_id, GT);  // every role of the group is done with the scratch
    STAMP(91);

    // ================= output tile ============================================================================
    const int S6n = A.S * 6, OBS = A.OBS;
    constexpr int GE = 32;                  // envs of a group
    const int le = lane;                    // env index inside the group's tile
    const int gtid = role * 32 + lane;      // thread index inside the group
    const int gcnt = cnt - grp * GE < 0 ? 0 : (cnt - grp * GE > GE ? GE : cnt - grp * GE);  // envs of this group
    const long long ge0 = e0 + grp * GE;    // first global env of the group
    float* o_q = dynsm + (size_t)grp * M.group_floats * 32;
    float* o_qd = o_q + GE * D;
    float* o_sens = o_qd + GE * D;
    float* o_pos = o_sens + GE * S6n;
    float* o_quat = o_pos + GE * 3;
    float* o_vel = o_quat + GE * 4;
    float* o_act = o_vel + GE * 6;          // epilogue only from here on
    float* o_obs = o_act + GE * D;
    float* o_up = o_obs + GE * OBS;
    float* o_head = o_up + GE * 3;
    float* o_pot = o_head + GE * 3;
    float* s_part = o_pot + GE;             // [3 reductions][4 quarters][GE]
    float* s_lo = s_part + 12 * GE;         // per-dof constants
    float* s_hi = s_lo + 32;
    float* s_ratio = s_hi + 32;
    float* s_rcp = s_ratio + 32;
    if (valid) {
#pragma unroll
        for (int k = 0; k < MAXC; ++k) {
            if (k < np) {
                const int d = M.dof[M.plink[role][k]];
                o_q[le * D + d] = bad ? 0.f : qr[k];
                o_qd[le * D + d] = bad ? 0.f : qdr[k];
                o_act[le * D + d] = ar[k];
            }
        }
        if (role == 0) {
#pragma unroll
            for (int t = 1; t < kRT; ++t) {
                if (t < ntrunk) {
                    const int d = M.dof[M.tlink[t]];
                    o_q[le * D + d] = bad ? 0.f : tq[t - 1];
                    o_qd[le * D + d] = bad ? 0.f : tqd[t - 1];
                    o_act[le * D + d] = ta[t - 1];
                }
            }
            o_pos[le * 3] = rpos.x; o_pos[le * 3 + 1] = rpos.y; o_pos[le * 3 + 2] = rpos.z;
            *reinterpret_cast<float4*>(o_quat + le * 4) = make_float4(rquat.w, rquat.x, rquat.y, rquat.z);
            float* vv = o_vel + le * 6;
            vv[0] = rvel.a[3]; vv[1] = rvel.a[4]; vv[2] = rvel.a[5];
            vv[3] = rvel.a[0]; vv[4] = rvel.a[1]; vv[5] = rvel.a[2];
        }
        if (sw_idx >= 0) {
#pragma unroll
            for (int j = 0; j < 6; ++j) o_sens[le * S6n + sw_idx * 6 + j] = bad ? 0.f : sw[j];
        }
    }
    if (A.epi) {
        if (gtid < 32) {
            s_lo[gtid] = gtid < D ? P.dof_limits_lower[gtid] : 0.0f;
            s_hi[gtid] = gtid < D ? P.dof_limits_upper[gtid] : 0.0f;
            s_ratio[gtid] = gtid < D ? P.motor_effort_ratio[gtid] : 0.0f;
            s_rcp[gtid] = gtid < D ? 1.0f / (P.dof_limits_upper[gtid] - P.dof_limits_lower[gtid]) : 0.0f;
        }
    }
    group_barrier(bar_id, GT);
    STAMP(92);

    // ================= epilogue: observations, reward, termination (locomotion_math.cuh = K1's arithmetic) ==
    float st_rew = 0.0f, st_done = 0.0f, st_len = 0.0f;
    if (A.epi) {
        const bool closed = P.closed_form_angles != 0;
        const float clip = P.clip_obs;  // the stored rows are clamped; the metrics see the unclamped values
        // inputs of the reward / termination thread, requested before the arithmetic below hides their latency
        float prev_pot_in = 0.f;
        long long prog_in = 0, reset_in = 0;
        if (role == 0 && valid) { prev_pot_in = A.pot[env]; prog_in = A.progress[env]; reset_in = A.reset[env]; }
        if (valid) {
            float* row = o_obs + le * OBS;
            // quarter r of env le = K1's thread (le, r).  Quarters are dealt so that role 0, which also evaluates the reward
            // below, gets the light one (2: heading, potentials) and the last role the two around it (0: roll / up, 3:
            // local velocities); who evaluates a quarter does not change a bit of the result.
#pragma unroll 1
            for (int r = NR - 1 - role; r < 4; r += NR) {
                lm::Partial part;
                lm::partial_init(part);
                for (int k = 0; k < (D + 3) / 4; ++k) {
                    const int d = lm::owned_index_rt(D, r, k, le);
                    if (d < D) {
                        const float qs = lm::unscale(o_q[le * D + d], s_lo[d], s_hi[d], s_rcp[d]);
                        const float dv = lm::mul(o_qd[le * D + d], P.dof_vel_scale);
                        const float a = o_act[le * D + d];
                        row[12 + d] = lm::clampf(qs, clip);
                        row[12 + D + d] = lm::clampf(dv, clip);
                        row[12 + 2 * D + S6n + d] = lm::clampf(a, clip);
                        lm::partial_add(P, part, qs, dv, a, s_ratio[d]);
                    }
                }
                lm::partial_finish(P, part);
                s_part[(0 * 4 + r) * GE + le] = part.limit;
                s_part[(1 * 4 + r) * GE + le] = part.act;
                s_part[(2 * 4 + r) * GE + le] = part.elec;
                for (int k = 0; k < (S6n + 3) / 4; ++k) {
                    const int j = lm::owned_index_elem_rt(S6n, r, k, le);
                    if (j < S6n) row[12 + 2 * D + j] = lm::clampf(lm::mul(o_sens[le * S6n + j], P.contact_force_scale), clip);
                }
                const float4 q4 = *reinterpret_cast<const float4*>(o_quat + 4 * le);
                const float rot[4] = {q4.x, q4.y, q4.z, q4.w};
                float tqq[4];
                lm::torso_quat(P, rot, tqq);
                const float px = o_pos[3 * le], py = o_pos[3 * le + 1], pz = o_pos[3 * le + 2];
                // columns 0 (height), 10 (up projection) and 11 (heading projection) stay unclamped until the reward
                // thread has read them; every other column is clamped by the thread that wrote it
                if (r == 0) {
                    lm::root_part0(P, tqq, closed, row, o_up + 3 * le);
                    row[8] = lm::clampf(row[8], clip);
                } else if (r == 1) {
                    lm::root_part1(P, tqq, px, py, pz, closed, row);
                    row[7] = lm::clampf(row[7], clip);
                    row[9] = lm::clampf(row[9], clip);
                } else if (r == 2) {
                    lm::root_part2(P, tqq, px, py, row, o_head + 3 * le, o_pot + le);
                } else {
                    const float lin3[3] = {o_vel[6 * le], o_vel[6 * le + 1], o_vel[6 * le + 2]};
                    const float ang3[3] = {o_vel[6 * le + 3], o_vel[6 * le + 4], o_vel[6 * le + 5]};
                    lm::root_part3(P, tqq, pz, lin3, ang3, row);
#pragma unroll
                    for (int j = 1; j < 7; ++j) row[j] = lm::clampf(row[j], clip);
                }
            }
        }
        group_barrier(bar_id, GT);
        STAMP(93);
        if (role == 0 && valid) {  // reward + termination: one thread per env, coalesced [N] accesses
            float* row = o_obs + le * OBS;
            float red[3];
#pragma unroll
            for (int c = 0; c < 3; ++c)
                red[c] = lm::combine4(s_part[(c * 4 + 0) * GE + le], s_part[(c * 4 + 1) * GE + le],
                                      s_part[(c * 4 + 2) * GE + le], s_part[(c * 4 + 3) * GE + le]);
            const float prev = prev_pot_in;
            float total;
            long long prog, rst;
            lm::reward_done(P, row[0], row[10], row[11], o_pot[le], prev, red, prog_in, reset_in, total, prog, rst);
            row[0] = lm::clampf(row[0], clip);
            row[10] = lm::clampf(row[10], clip);
            row[11] = lm::clampf(row[11], clip);
            A.progress[env] = prog;
            A.reset[env] = rst;
            A.rew[env] = total;
            if (A.rew_host != nullptr) { A.rew_host[env] = total; A.reset_host[env] = rst; }
            A.prev_pot[env] = prev;
            A.pot[env] = o_pot[le];
            st_rew = total;
            st_done = rst != 0 ? 1.0f : 0.0f;
            st_len = rst != 0 ? (float)prog : 0.0f;
            if (A.ep_ret != nullptr) {  // episodic return: published (and started over) in the step that ends the episode
                const float ret = A.ep_ret[env] + total;
                A.ep_ret[env] = rst != 0 ? 0.0f : ret;
                st_rew = rst != 0 ? ret : 0.0f;
            }
        }
        if (A.stats != nullptr && role == 0) {  // warp-uniform: one striped atomic set per group
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                st_rew += __shfl_xor_sync(0xffffffffu, st_rew, o);
                st_done += __shfl_xor_sync(0xffffffffu, st_done, o);
                st_len += __shfl_xor_sync(0xffffffffu, st_len, o);
            }
            if (lane == 0 && gcnt > 0) {
                double* slot = A.stats + ((blockIdx.x * (EPC / 32) + grp) % OIGE_STATS_SLOTS) * 4;
                atomicAdd(slot + 0, (double)st_rew);
                atomicAdd(slot + 1, (double)st_done);
                atomicAdd(slot + 2, (double)st_len);
                atomicAdd(slot + 3, (double)gcnt);
            }
        }
    }
    STAMP(94);

    // ================= drain (per group) ======================================================================
    if (A.use_bulk && gcnt == GE) {
        fence_async_smem();
        group_barrier(bar_id, GT);
        STAMP(95);
        if (gtid == 0) {
            bulk_s2g(A.q + ge0 * D, o_q, GE * D * 4);
            bulk_s2g(A.qd + ge0 * D, o_qd, GE * D * 4);
            if (A.sens != nullptr && S6n > 0) bulk_s2g(A.sens + ge0 * S6n, o_sens, GE * S6n * 4);
            bulk_s2g(A.pos + ge0 * 3, o_pos, GE * 12);
            bulk_s2g(A.quat + ge0 * 4, o_quat, GE * 16);
            bulk_s2g(A.vel + ge0 * 6, o_vel, GE * 24);
            if (A.epi) {
                bulk_s2g(A.obs + ge0 * OBS, o_obs, GE * OBS * 4);
                bulk_s2g(A.up + ge0 * 3, o_up, GE * 12);
                bulk_s2g(A.head + ge0 * 3, o_head, GE * 12);
            }
            bulk_commit();
            STAMP(99);
            bulk_wait_read0();  // smem must stay alive until the TMA engine has read it
        }
    } else {
        group_barrier(bar_id, GT);
        _Pragma("unroll 1") for (int i = gtid; i < gcnt * D; i += GT) { A.q[ge0 * D + i] = o_q[i]; A.qd[ge0 * D + i] = o_qd[i]; }
        if (A.sens != nullptr)
            _Pragma("unroll 1") for (int i = gtid; i < gcnt * S6n; i += GT) A.sens[ge0 * S6n + i] = o_sens[i];
        _Pragma("unroll 1") for (int i = gtid; i < gcnt * 3; i += GT) A.pos[ge0 * 3 + i] = o_pos[i];
        _Pragma("unroll 1") for (int i = gtid; i < gcnt * 4; i += GT) A.quat[ge0 * 4 + i] = o_quat[i];
        _Pragma("unroll 1") for (int i = gtid; i < gcnt * 6; i += GT) A.vel[ge0 * 6 + i] = o_vel[i];
        if (A.epi) {
            _Pragma("unroll 1") for (int i = gtid; i < gcnt * OBS; i += GT) A.obs[ge0 * OBS + i] = o_obs[i];
            _Pragma("unroll 1") for (int i = gtid; i < gcnt * 3; i += GT) { A.up[ge0 * 3 + i] = o_up[i]; A.head[ge0 * 3 + i] = o_head[i]; }
        }
    }
}
