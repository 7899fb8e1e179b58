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

#include <atomic>

#include "locomotion_math.cuh"
#include "oige_common.cuh"

namespace oige {

namespace {

using namespace lm;  // exact-order arithmetic shared with the whole-step kernel (locomotion_math.cuh)

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
    float* ep_ret;  // [N] running return of the current episode (OigeTaskBuffers.episode_return) or null
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
        Partial part;
        partial_init(part);
#pragma unroll
        for (int k = 0; k < (D + 3) / 4; ++k) {
            const int d = owned_index<D>(r, k, e);
            if (d < D) {
                const float qs = unscale(s_q[e * D + d], s_lo[d], s_hi[d], s_rcp[d]);  // maths.unscale
                const float dv = mul(s_qd[e * D + d], P.dof_vel_scale);
                const float a = s_act[e * D + d];
                row[12 + d] = qs;
                row[12 + D + d] = dv;
                row[12 + 2 * D + 6 * S + d] = a;
                partial_add(P, part, qs, dv, a, s_ratio[d]);
            }
        }
        partial_finish(P, part);
        s_part[(0 * 4 + r) * E + e] = part.limit;
        s_part[(1 * 4 + r) * E + e] = part.act;
        s_part[(2 * 4 + r) * E + e] = part.elec;
#pragma unroll
        for (int k = 0; k < (6 * S + 3) / 4; ++k) {
            const int j = owned_index_elem<6 * S>(r, k, e);
            if (j < 6 * S) row[12 + 2 * D + j] = mul(s_sens[e * 6 * S + j], P.contact_force_scale);
        }

        const float4 q4 = *reinterpret_cast<const float4*>(s_quat + 4 * e);
        const float rot[4] = {q4.x, q4.y, q4.z, q4.w};
        float tq[4];
        torso_quat(P, rot, tq);
        const float px = s_pos[3 * e], py = s_pos[3 * e + 1], pz = s_pos[3 * e + 2];
        if (r == 0) {
            root_part0(P, tq, closed, row, o_up + 3 * e);
        } else if (r == 1) {
            root_part1(P, tq, px, py, pz, closed, row);
        } else if (r == 2) {
            root_part2(P, tq, px, py, row, o_head + 3 * e, o_pot + e);
        } else {
            const float lin[3] = {s_vel[6 * e], s_vel[6 * e + 1], s_vel[6 * e + 2]};
            const float ang[3] = {s_vel[6 * e + 3], s_vel[6 * e + 4], s_vel[6 * e + 5]};
            root_part3(P, tq, pz, lin, ang, row);
        }
    }
    __syncthreads();

    // ---- reward + termination: thread r == 0 of each env ------------------------------------------
    float st_rew = 0.0f, st_done = 0.0f, st_len = 0.0f;
    if (live && r == 0) {
        const float* row = o_obs + e * OBS;
        // the four partial sums of each reduction are combined in a fixed order
        float red[3];
#pragma unroll
        for (int c = 0; c < 3; ++c)
            red[c] = combine4(s_part[(c * 4 + 0) * E + e], s_part[(c * 4 + 1) * E + e], s_part[(c * 4 + 2) * E + e],
                              s_part[(c * 4 + 3) * E + e]);
        float total;
        long long prog, rst;
        reward_done(P, row[0], row[10], row[11], o_pot[e], s_pot[e], red, s_prog[e], s_reset[e], total, prog, rst);
        s_prog[e] = prog;
        s_reset[e] = rst;
        o_rew[e] = total;
        st_rew = total;
        st_done = rst != 0 ? 1.0f : 0.0f;
        st_len = rst != 0 ? (float)prog : 0.0f;
        if (A.ep_ret != nullptr) {  // episodic return: published (and started over) in the pass that ends the episode
            const float ret = A.ep_ret[e0 + e] + total;
            A.ep_ret[e0 + e] = rst != 0 ? 0.0f : ret;
            st_rew = rst != 0 ? ret : 0.0f;
        }
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
    static std::atomic<bool> configured[64];  // per device: the attribute belongs to the device's copy of the function
    constexpr size_t smem = k1_smem_bytes<D, S, E>();
    auto kern = obs_reward_reset_kernel<D, S, E>;
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) dev = 0;
    if (!configured[dev].load(std::memory_order_acquire)) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return fail((int)e, "cudaFuncSetAttribute(K1): %s", cudaGetErrorString(e));
        configured[dev].store(true, std::memory_order_release);
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
    A.ep_ret = b->episode_stats != nullptr ? b->episode_return : nullptr;
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
