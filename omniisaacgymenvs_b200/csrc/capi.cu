// C ABI of liboige_b200.so (declared in include/oige_b200.h).  Thin argument checking + dispatch;
// the kernels live in obs_reward_reset.cu (K1), articulation_step.cu (K2), reset_idx.cu (K3).
#include <atomic>
#include <mutex>
#include <cstdarg>
#include <cstdio>
#include <cstring>

#include "oige_common.cuh"

namespace oige {

static thread_local char g_err[512] = "";
static std::atomic<uint64_t> g_launches{0};

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

int fail(int code, const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return code;
}

int check_launch(const char* what) {
    cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess) return OIGE_OK;
    set_error("%s: %s", what, cudaGetErrorString(e));
    return (int)e;
}

void count_launch(uint64_t n) { g_launches.fetch_add(n, std::memory_order_relaxed); }

bool pdl_enabled() {
    static std::atomic<int> v{-1};
    int cur = v.load(std::memory_order_acquire);
    if (cur < 0) { const char* e = getenv("OIGE_PDL"); cur = (e && e[0] == '0') ? 0 : 1; v.store(cur, std::memory_order_release); }
    return cur != 0;
}

int sm_count() {  // of the CURRENT device (a process may drive several GPUs)
    static std::atomic<int> cached[64];
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
    int n = cached[dev].load(std::memory_order_acquire);
    if (n <= 0) {
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) return 148;
        cached[dev].store(n, std::memory_order_release);
    }
    return n;
}

// kernels' host-side dispatchers
int k1_dispatch(const OigeTaskParams*, const OigeState*, const OigeTaskBuffers*, int64_t, cudaStream_t, int, bool);
int k3_masked(const OigeTaskParams*, const OigeState*, const OigeTaskBuffers*, const OigeResetInit*, const float*,
              const float*, uint64_t, uint64_t, int64_t, int64_t, cudaStream_t, const float* actions_in, float clip);
int k3_idx(const OigeTaskParams*, const OigeState*, const OigeTaskBuffers*, const OigeResetInit*, const int64_t*,
           int64_t, const float*, const float*, uint64_t, uint64_t, int64_t, cudaStream_t);
int k2_step(const OigeModel*, const OigeTaskParams*, const OigeState*, const float*, int64_t, int32_t, float,
            cudaStream_t, int64_t decision_n);
int k2_fused_step(const OigeModel*, const OigeTaskParams*, const OigeState*, const OigeTaskBuffers*, const OigeResetInit*,
                  const float* actions_in, float clip, uint64_t seed, uint64_t counter, int64_t env_id_offset, int64_t n,
                  int32_t substeps, float dt, cudaStream_t, int64_t decision_n, bool bump_counter);
void k2_reload_env();
void k2_set_host_mirror(float* rew_host, long long* reset_host);
bool k2_fused_applicable(const OigeModel*, int64_t decision_n, bool with_targets);
uint64_t k2_nonfinite_count();
OigeModel* k2_model_create(const OigeModelDesc*);
int dr_apply(const OigeNoiseSpec*, const OigeNoiseSpec*, const OigeNoiseState*, const float*, float*, const int64_t*,
             int64_t, int32_t, float, float, const float*, const float*, uint64_t, uint64_t, const uint64_t*, int32_t,
             int64_t, cudaStream_t);
int anymal_post(const OigeAnymalParams*, const OigeState*, const OigeAnymalBuffers*, int64_t, cudaStream_t, bool);
int anymal_pre(const OigeAnymalParams*, const OigeState*, const OigeAnymalBuffers*, const OigeResetInit*, const float*,
               float, const float*, const float*, uint64_t, uint64_t, int64_t, int64_t, cudaStream_t, bool);
int anymal_reset_idx(const OigeAnymalParams*, const OigeState*, const OigeAnymalBuffers*, const OigeResetInit*,
                     const int64_t*, int64_t, const float*, const float*, uint64_t, uint64_t, int64_t, cudaStream_t);
int cartpole_post(const OigeCartpoleParams*, const OigeState*, const OigeCartpoleBuffers*, int64_t, cudaStream_t, bool);
int cartpole_pre(const OigeCartpoleParams*, const OigeState*, const OigeCartpoleBuffers*, const float*, float,
                 const float*, uint64_t, uint64_t, int64_t, int64_t, cudaStream_t, bool);
int cartpole_reset_idx(const OigeCartpoleParams*, const OigeState*, const OigeCartpoleBuffers*, const int64_t*, int64_t,
                       const float*, uint64_t, uint64_t, int64_t, cudaStream_t);
void k2_model_destroy(OigeModel*);
void host_knobs_reload();

static int check_common(const char* who, const OigeTaskParams* p, const OigeState* s, const OigeTaskBuffers* b,
                        int64_t n) {
    if (!p || !s || !b) return fail(OIGE_E_NULL, "%s: null struct pointer", who);
    if (n < 0 || n > (int64_t)1 << 31) return fail(OIGE_E_RANGE, "%s: n_envs %lld out of range", who, (long long)n);
    if (p->num_dof < 1 || p->num_dof > OIGE_MAX_DOF) return fail(OIGE_E_SHAPE, "%s: num_dof %d", who, p->num_dof);
    return OIGE_OK;
}

// lazily created per-device copy/compute pipeline of oige_env_step_host_f32
constexpr int kHostChunks = 8;  // streams / events; the chunk count used is host_chunk_count()
struct HostPipe {
    cudaStream_t streams[kHostChunks];
    cudaEvent_t start, done[kHostChunks];
    cudaEvent_t t0, t[kHostChunks][3];  // OIGE_HOST_TRACE=1: after H2D, after the kernels, after D2H (timing events)
    bool trace, pending;
    int last_used;
};
// run-time knobs of the host-buffer step, read once (oige_reload_env() re-reads them)
struct HostKnobs { int chunks, zerocopy, trace; };
static HostKnobs g_hk = {4, 0, 0};
static std::once_flag g_hk_once;
static void host_knobs_read() {
    HostKnobs k = {4, 0, 0};
    if (const char* e = getenv("OIGE_HOST_CHUNKS")) { const int v = atoi(e); k.chunks = v < 1 ? 1 : (v > kHostChunks ? kHostChunks : v); }
    if (const char* e = getenv("OIGE_HOST_ZEROCOPY")) k.zerocopy = atoi(e);
    if (const char* e = getenv("OIGE_HOST_TRACE")) k.trace = e[0] == '1';
    g_hk = k;
}
static const HostKnobs& host_knobs() { std::call_once(g_hk_once, host_knobs_read); return g_hk; }
void host_knobs_reload() { host_knobs(); host_knobs_read(); }
static int host_chunk_count() { return host_knobs().chunks; }  // OIGE_HOST_CHUNKS=1..8 (default 4, tools/e2e_breakdown.py)
static HostPipe* host_pipe() {
    static HostPipe* pipes[64] = {nullptr};
    static std::mutex mu;  // creation is serialised: the C ABI may be entered from several host threads
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return nullptr;
    std::lock_guard<std::mutex> lock(mu);
    if (!pipes[dev]) {
        HostPipe* p = new HostPipe();
        bool ok = cudaEventCreateWithFlags(&p->start, cudaEventDisableTiming) == cudaSuccess;
        for (int i = 0; i < kHostChunks && ok; ++i) {
            ok = cudaStreamCreateWithFlags(&p->streams[i], cudaStreamNonBlocking) == cudaSuccess &&
                 cudaEventCreateWithFlags(&p->done[i], cudaEventDisableTiming) == cudaSuccess;
        }
        p->trace = host_knobs().trace != 0;
        p->pending = false;
        if (p->trace) {
            ok = ok && cudaEventCreate(&p->t0) == cudaSuccess;
            for (int i = 0; i < kHostChunks && ok; ++i)
                for (int k = 0; k < 3 && ok; ++k) ok = cudaEventCreate(&p->t[i][k]) == cudaSuccess;
        }
        if (!ok) { delete p; return nullptr; }
        pipes[dev] = p;
    }
    return pipes[dev];
}

__global__ void bump_counter_kernel(unsigned long long* c) { *c += 1; }

}  // namespace oige

using namespace oige;

extern "C" {

int oige_abi_version(void) { return OIGE_ABI_VERSION; }
void oige_reload_env(void) { k2_reload_env(); host_knobs_reload(); }
uint64_t oige_nonfinite_env_count(void) { return k2_nonfinite_count(); }
const char* oige_last_error(void) { return g_err; }
uint64_t oige_kernel_launch_count(void) { return g_launches.load(std::memory_order_relaxed); }

int oige_obs_reward_reset_f32(const OigeTaskParams* params, const OigeState* state, const OigeTaskBuffers* buf,
                              int64_t n_envs, void* stream) {
    if (int rc = check_common("oige_obs_reward_reset_f32", params, state, buf, n_envs)) return rc;
    if (n_envs == 0) return OIGE_OK;
    return k1_dispatch(params, state, buf, n_envs, (cudaStream_t)stream, 0, false);
}

// same as above with an explicit tile size (32/64/128) -- used by the benchmark's tile sweep
int oige_obs_reward_reset_tile_f32(const OigeTaskParams* params, const OigeState* state, const OigeTaskBuffers* buf,
                                   int64_t n_envs, int32_t tile, void* stream) {
    if (int rc = check_common("oige_obs_reward_reset_tile_f32", params, state, buf, n_envs)) return rc;
    if (n_envs == 0) return OIGE_OK;
    return k1_dispatch(params, state, buf, n_envs, (cudaStream_t)stream, tile, false);
}

int oige_reset_masked_f32(const OigeTaskParams* params, const OigeState* state, const OigeTaskBuffers* buf,
                          const OigeResetInit* init, const float* rand_pos, const float* rand_vel, uint64_t seed,
                          uint64_t counter, int64_t env_id_offset, int64_t n_envs, void* stream) {
    if (int rc = check_common("oige_reset_masked_f32", params, state, buf, n_envs)) return rc;
    if (n_envs == 0) return OIGE_OK;
    return k3_masked(params, state, buf, init, rand_pos, rand_vel, seed, counter, env_id_offset, n_envs,
                     (cudaStream_t)stream, nullptr, 0.0f);
}

int oige_reset_idx_f32(const OigeTaskParams* params, const OigeState* state, const OigeTaskBuffers* buf,
                       const OigeResetInit* init, const int64_t* env_ids, int64_t k, const float* rand_pos,
                       const float* rand_vel, uint64_t seed, uint64_t counter, int64_t env_id_offset, void* stream) {
    if (int rc = check_common("oige_reset_idx_f32", params, state, buf, k)) return rc;
    if (k == 0) return OIGE_OK;
    return k3_idx(params, state, buf, init, env_ids, k, rand_pos, rand_vel, seed, counter, env_id_offset,
                  (cudaStream_t)stream);
}

OigeModel* oige_model_create(const OigeModelDesc* desc) {
    if (!desc) {
        set_error("oige_model_create: null desc");
        return nullptr;
    }
    return k2_model_create(desc);
}

void oige_model_destroy(OigeModel* model) {
    if (model) k2_model_destroy(model);
}

int oige_articulation_step_f32(const OigeModel* model, const OigeTaskParams* params, const OigeState* state,
                               const float* actions, int64_t n_envs, int32_t substeps, float dt, void* stream) {
    if (!model || !params || !state || !actions) return fail(OIGE_E_NULL, "oige_articulation_step_f32: null argument");
    if (n_envs < 0 || n_envs > (int64_t)1 << 26) return fail(OIGE_E_RANGE, "n_envs %lld", (long long)n_envs);
    if (substeps < 1 || substeps > 64) return fail(OIGE_E_RANGE, "substeps %d not in [1,64]", substeps);
    if (!(dt > 0.0f)) return fail(OIGE_E_RANGE, "dt must be > 0");
    if (n_envs == 0) return OIGE_OK;
    return k2_step(model, params, state, actions, n_envs, substeps, dt, (cudaStream_t)stream, 0);
}

}  // extern "C" (reopened below)

static int env_step_impl(const OigeModel* model, const OigeTaskParams* params, const OigeState* state,
                      const OigeTaskBuffers* buf, const OigeResetInit* init, const float* actions_in,
                      float clip_actions, uint64_t seed, uint64_t counter, int64_t env_id_offset, int64_t n_envs,
                      int32_t substeps, float dt, void* stream, int64_t decision_n) {
    if (int rc = check_common("oige_env_step_f32", params, state, buf, n_envs)) return rc;
    if (!model || !init || !actions_in) return fail(OIGE_E_NULL, "oige_env_step_f32: null argument");
    if (substeps < 1 || substeps > 64) return fail(OIGE_E_RANGE, "substeps %d not in [1,64]", substeps);
    if (n_envs == 0) return OIGE_OK;
    cudaStream_t s = (cudaStream_t)stream;
    // VecEnvRLGames.step: clamp -> pre_physics_step (reset flagged envs, store actions, efforts)
    //                     -> control_frequency_inv x world.step -> post_physics_step
    // ONE launch when the role-split kernel serves this model and env count (K2-F: prologue + substeps + epilogue) ...
    {
        const int rc = k2_fused_step(model, params, state, buf, init, actions_in, clip_actions, seed, counter,
                                     env_id_offset, n_envs, substeps, dt, s, decision_n, decision_n <= 0);
        if (rc != 1000) return rc;
    }
    // ... otherwise three: K3 (reset + clamp) -> K2-W / K2-T -> K1
    if (int rc = k3_masked(params, state, buf, init, nullptr, nullptr, seed, counter, env_id_offset, n_envs, s,
                           actions_in, clip_actions))
        return rc;
    if (int rc = k2_step(model, params, state, buf->actions, n_envs, substeps, dt, s, decision_n)) return rc;
    return k1_dispatch(params, state, buf, n_envs, s, 0, decision_n <= 0);  // chunked callers bump the counter themselves
}

extern "C" {

int oige_env_step_f32(const OigeModel* model, const OigeTaskParams* params, const OigeState* state,
                      const OigeTaskBuffers* buf, const OigeResetInit* init, const float* actions_in,
                      float clip_actions, uint64_t seed, uint64_t counter, int64_t env_id_offset, int64_t n_envs,
                      int32_t substeps, float dt, void* stream) {
    return env_step_impl(model, params, state, buf, init, actions_in, clip_actions, seed, counter, env_id_offset,
                         n_envs, substeps, dt, stream, 0);
}

int oige_env_step_host_f32(const OigeModel* model, const OigeTaskParams* params, const OigeState* state,
                           const OigeTaskBuffers* buf, const OigeResetInit* init, const float* actions_host,
                           float* actions_dev_staging, float* obs_host, float* rew_host, int64_t* reset_host,
                           float clip_actions, uint64_t seed, uint64_t counter, int64_t env_id_offset,
                           int64_t n_envs, int32_t substeps, float dt, void* stream) {
    if (int rc = check_common("oige_env_step_host_f32", params, state, buf, n_envs)) return rc;
    if (!actions_host || !actions_dev_staging || !obs_host || !rew_host || !reset_host || !init || !model)
        return fail(OIGE_E_NULL, "oige_env_step_host_f32: null host/staging buffer");
    cudaStream_t user = (cudaStream_t)stream;
    HostPipe* hp = host_pipe();  // created on the first call, also for n_envs == 0 (lets a caller warm up before a capture)
    if (!hp) return fail((int)cudaErrorUnknown, "oige_env_step_host_f32: cannot create copy/compute streams");
    if (n_envs == 0) return OIGE_OK;
    // Envs are independent, so the step is cut into kHostChunks contiguous env ranges, each on its own
    // stream: H2D(actions) -> K3 -> K2 -> K1 -> D2H(obs, rew, resets).  Chunk c's D2H overlaps chunk c+1's
    // kernels and chunk c+2's H2D (separate copy engines per direction).  Results are identical to the
    // one-launch step: every kernel is row-wise and reset noise is keyed by the global env id.
    const int D = params->num_dof, S = params->num_sensors, OBS = params->num_obs;
    // Opt-in (OIGE_HOST_ZEROCOPY=1) zero-copy observations: when obs_host is pinned (device-accessible under UVA) K1
    // stores its observation tiles straight into it over PCIe (bulk shared->global stores) instead of a D2H copy per
    // chunk; buf->obs on the device is then NOT written by this call.  Measured 272 vs 284 us per 16384-env step.
    float* obs_host_dev = nullptr;
    const float* actions_host_dev = nullptr;
    {
        cudaPointerAttributes at;
        const int zmode = host_knobs().zerocopy;  // bit 0: observations written straight to the host, bit 1: actions read from it
        if ((zmode & 1) && cudaPointerGetAttributes(&at, obs_host) == cudaSuccess &&
            at.type == cudaMemoryTypeHost && at.devicePointer != nullptr)
            obs_host_dev = static_cast<float*>(at.devicePointer);
        else if (zmode & 1)
            cudaGetLastError();  // unregistered host memory reports an error on older drivers: clear it
        // zero-copy actions: K3's clamp pass reads the pinned host rows over PCIe (coalesced) and writes the clamped
        // copy to HBM -- no H2D copy-engine transfer, no staging buffer, one stream operation less per chunk
        if ((zmode & 2) && cudaPointerGetAttributes(&at, actions_host) == cudaSuccess &&
            at.type == cudaMemoryTypeHost && at.devicePointer != nullptr)
            actions_host_dev = static_cast<const float*>(at.devicePointer);
        else if (zmode & 2)
            cudaGetLastError();
    }
    // ... and, in the same mode, the reward / reset rows when the whole-step kernel serves the call: it stores them into
    // the pinned buffers itself (besides the device copies the next step needs), and the two D2H copies below go away
    float* rew_host_dev = nullptr;
    long long* reset_host_dev = nullptr;
    if (obs_host_dev && k2_fused_applicable(model, n_envs, state->dof_pos_targets != nullptr)) {
        cudaPointerAttributes at;
        if (cudaPointerGetAttributes(&at, rew_host) == cudaSuccess && at.type == cudaMemoryTypeHost && at.devicePointer)
            rew_host_dev = static_cast<float*>(at.devicePointer);
        else cudaGetLastError();
        if (cudaPointerGetAttributes(&at, reset_host) == cudaSuccess && at.type == cudaMemoryTypeHost && at.devicePointer)
            reset_host_dev = static_cast<long long*>(at.devicePointer);
        else cudaGetLastError();
        if (!rew_host_dev || !reset_host_dev) { rew_host_dev = nullptr; reset_host_dev = nullptr; }
    }
    const int nchunks = host_chunk_count();
    int64_t per = ((n_envs + nchunks - 1) / nchunks + 31) / 32 * 32;  // tile-aligned chunk size
    if (n_envs < 4096) per = n_envs;                                           // too small to be worth splitting
    cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
    cudaStreamIsCapturing(user, &cap);
    if (hp->trace && hp->pending && cap == cudaStreamCaptureStatusNone) {  // diagnostic: timeline of the PREVIOUS call
        cudaEventSynchronize(hp->t[hp->last_used - 1][2]);
        for (int c = 0; c < hp->last_used; ++c) {
            float a = 0, b = 0, d = 0;
            cudaEventElapsedTime(&a, hp->t0, hp->t[c][0]); cudaEventElapsedTime(&b, hp->t0, hp->t[c][1]);
            cudaEventElapsedTime(&d, hp->t0, hp->t[c][2]);
            fprintf(stderr, "[oige host step] chunk %d: H2D done %.1f us, kernels done %.1f us, D2H done %.1f us\n", c,
                    a * 1e3f, b * 1e3f, d * 1e3f);
        }
    }
    cudaError_t e = cudaEventRecord(hp->start, user);
    if (e != cudaSuccess) return fail((int)e, "event record: %s", cudaGetErrorString(e));
    if (hp->trace) { cudaEventRecord(hp->t0, user); hp->pending = true; }
    int used = 0;
    for (int64_t off = 0; off < n_envs; off += per, ++used) {
        const int64_t cnt = (n_envs - off) < per ? (n_envs - off) : per;
        cudaStream_t s = hp->streams[used];
        cudaStreamWaitEvent(s, hp->start, 0);
        OigeState st = *state;
        st.root_pos += off * 3; st.root_quat += off * 4; st.root_vel += off * 6;
        st.dof_pos += off * D; st.dof_vel += off * D;
        if (st.sensors) st.sensors += off * 6 * S;
        if (st.dof_pos_targets) st.dof_pos_targets += off * D;
        OigeTaskBuffers bf = *buf;
        bf.obs += off * OBS; bf.rew += off; bf.reset_buf += off; bf.progress_buf += off; bf.potentials += off;
        bf.prev_potentials += off; bf.up_vec += off * 3; bf.heading_vec += off * 3; bf.actions += off * D;
        if (bf.episode_return) bf.episode_return += off;
        if (obs_host_dev) bf.obs = obs_host_dev + off * OBS;
        // chunks share one counter: the host-provided one, or -- when buf->reset_counter is set (the call is being
        // captured in a CUDA graph) -- the device scalar, which K3 only reads; it is bumped once after the last chunk
        OigeResetInit in = *init;
        in.initial_root_pos += off * 3; in.initial_root_rot += off * 4; in.initial_dof_pos += off * D;
        const float* stage = actions_dev_staging + off * D;
        if (actions_host_dev) {
            stage = actions_host_dev + off * D;
        } else {
            e = cudaMemcpyAsync(actions_dev_staging + off * D, actions_host + off * D, (size_t)cnt * D * sizeof(float),
                                cudaMemcpyHostToDevice, s);
            if (e != cudaSuccess) return fail((int)e, "H2D actions: %s", cudaGetErrorString(e));
        }
        if (hp->trace) cudaEventRecord(hp->t[used][0], s);
        if (rew_host_dev) k2_set_host_mirror(rew_host_dev + off, reset_host_dev + off);
        if (int rc = env_step_impl(model, params, &st, &bf, &in, stage, clip_actions, seed, counter,
                                   env_id_offset + off, cnt, substeps, dt, (void*)s, n_envs))
            return rc;
        if (hp->trace) cudaEventRecord(hp->t[used][1], s);
        // only the big observation block is copied per chunk: every copy costs ~5 us of fixed overhead on the copy
        // engine, so the small rew / reset blocks of ALL chunks go out as two copies after the last chunk
        if (!obs_host_dev) {
            e = cudaMemcpyAsync(obs_host + off * OBS, bf.obs, (size_t)cnt * OBS * sizeof(float), cudaMemcpyDeviceToHost, s);
            if (e != cudaSuccess) return fail((int)e, "D2H obs: %s", cudaGetErrorString(e));
        }
        if (hp->trace) cudaEventRecord(hp->t[used][2], s);
        cudaEventRecord(hp->done[used], s);
        cudaStreamWaitEvent(user, hp->done[used], 0);  // the caller's stream observes the whole step
    }
    hp->last_used = used;
    if (!rew_host_dev) {
        e = cudaMemcpyAsync(rew_host, buf->rew, (size_t)n_envs * sizeof(float), cudaMemcpyDeviceToHost, user);
        if (e == cudaSuccess)
            e = cudaMemcpyAsync(reset_host, buf->reset_buf, (size_t)n_envs * sizeof(int64_t), cudaMemcpyDeviceToHost, user);
        if (e != cudaSuccess) return fail((int)e, "D2H rew/reset: %s", cudaGetErrorString(e));
    }
    if (buf->reset_counter != nullptr) {
        bump_counter_kernel<<<1, 1, 0, user>>>(reinterpret_cast<unsigned long long*>(buf->reset_counter));
        count_launch();
        return check_launch("bump_counter_kernel");
    }
    return OIGE_OK;
}

// Diagnostic (OIGE_HOST_TRACE=1): timeline of the last completed oige_env_step_host_f32 call or graph replay of it.
int oige_host_trace_dump(void) {
    HostPipe* hp = host_pipe();
    if (!hp || !hp->trace || hp->last_used <= 0) return fail(OIGE_E_RANGE, "oige_host_trace_dump: no trace (set OIGE_HOST_TRACE=1)");
    if (cudaEventSynchronize(hp->t[hp->last_used - 1][2]) != cudaSuccess) { cudaGetLastError(); return fail(OIGE_E_RANGE, "trace events not recorded"); }
    for (int c = 0; c < hp->last_used; ++c) {
        float a = 0, b = 0, d = 0;
        cudaEventElapsedTime(&a, hp->t0, hp->t[c][0]); cudaEventElapsedTime(&b, hp->t0, hp->t[c][1]);
        cudaEventElapsedTime(&d, hp->t0, hp->t[c][2]);
        fprintf(stderr, "[oige host step] chunk %d: H2D done %.1f us, kernels done %.1f us, D2H done %.1f us\n", c,
                a * 1e3f, b * 1e3f, d * 1e3f);
    }
    cudaGetLastError();
    return OIGE_OK;
}

// ---- Cartpole (tasks/cartpole.py) ------------------------------------------------------------------------
static int cartpole_range(const char* who, int64_t n) {
    if (n < 0 || n > (int64_t)1 << 31) return fail(OIGE_E_RANGE, "%s: n_envs %lld out of range", who, (long long)n);
    return OIGE_OK;
}

int oige_cartpole_obs_reward_reset_f32(const OigeCartpoleParams* params, const OigeState* state,
                                       const OigeCartpoleBuffers* buf, int64_t n_envs, void* stream) {
    if (int rc = cartpole_range("oige_cartpole_obs_reward_reset_f32", n_envs)) return rc;
    if (n_envs == 0) return (params && state && buf) ? OIGE_OK : fail(OIGE_E_NULL, "oige_cartpole_obs_reward_reset_f32: null struct pointer");
    return cartpole_post(params, state, buf, n_envs, (cudaStream_t)stream, false);
}

int oige_cartpole_reset_masked_f32(const OigeCartpoleParams* params, const OigeState* state,
                                   const OigeCartpoleBuffers* buf, const float* actions_in, float clip_actions,
                                   const float* rand, uint64_t seed, uint64_t counter, int64_t env_id_offset,
                                   int64_t n_envs, void* stream) {
    if (int rc = cartpole_range("oige_cartpole_reset_masked_f32", n_envs)) return rc;
    if (n_envs == 0) return (params && state && buf) ? OIGE_OK : fail(OIGE_E_NULL, "oige_cartpole_reset_masked_f32: null struct pointer");
    return cartpole_pre(params, state, buf, actions_in, clip_actions, rand, seed, counter, env_id_offset, n_envs,
                        (cudaStream_t)stream, false);
}

int oige_cartpole_reset_idx_f32(const OigeCartpoleParams* params, const OigeState* state,
                                const OigeCartpoleBuffers* buf, const int64_t* env_ids, int64_t k, const float* rand,
                                uint64_t seed, uint64_t counter, int64_t env_id_offset, void* stream) {
    if (int rc = cartpole_range("oige_cartpole_reset_idx_f32", k)) return rc;
    if (k == 0) return (params && state && buf) ? OIGE_OK : fail(OIGE_E_NULL, "oige_cartpole_reset_idx_f32: null struct pointer");
    return cartpole_reset_idx(params, state, buf, env_ids, k, rand, seed, counter, env_id_offset, (cudaStream_t)stream);
}

int oige_cartpole_env_step_f32(const OigeModel* model, const OigeCartpoleParams* params, const OigeState* state,
                               const OigeCartpoleBuffers* buf, const float* actions_in, float clip_actions,
                               uint64_t seed, uint64_t counter, int64_t env_id_offset, int64_t n_envs,
                               int32_t substeps, float dt, void* stream) {
    if (int rc = cartpole_range("oige_cartpole_env_step_f32", n_envs)) return rc;
    if (!model || !params || !state || !buf || !actions_in)
        return fail(OIGE_E_NULL, "oige_cartpole_env_step_f32: null argument");
    if (substeps < 1 || substeps > 64) return fail(OIGE_E_RANGE, "substeps %d not in [1,64]", substeps);
    if (n_envs == 0) return OIGE_OK;
    cudaStream_t s = (cudaStream_t)stream;
    // VecEnvRLGames.step: clamp -> pre_physics_step (reset flagged envs, efforts) -> substeps x world.step
    //                     -> post_physics_step
    if (int rc = cartpole_pre(params, state, buf, actions_in, clip_actions, nullptr, seed, counter, env_id_offset,
                              n_envs, s, buf->reset_counter != nullptr))
        return rc;
    OigeTaskParams unit;  // K2 reads num_dof, joint_gears and power_scale: the efforts are already formed
    memset(&unit, 0, sizeof(unit));
    unit.num_dof = 2;
    unit.joint_gears[0] = unit.joint_gears[1] = 1.0f;
    unit.power_scale = 1.0f;
    if (int rc = k2_step(model, &unit, state, buf->efforts, n_envs, substeps, dt, s, 0)) return rc;
    return cartpole_post(params, state, buf, n_envs, s, true);
}


// ---- Anymal (tasks/anymal.py) ------------------------------------------------------------------------------
int oige_anymal_obs_reward_reset_f32(const OigeAnymalParams* params, const OigeState* state,
                                     const OigeAnymalBuffers* buf, int64_t n_envs, void* stream) {
    if (int rc = cartpole_range("oige_anymal_obs_reward_reset_f32", n_envs)) return rc;
    if (n_envs == 0) return (params && state && buf) ? OIGE_OK : fail(OIGE_E_NULL, "oige_anymal_obs_reward_reset_f32: null struct pointer");
    return anymal_post(params, state, buf, n_envs, (cudaStream_t)stream, false);
}

int oige_anymal_reset_masked_f32(const OigeAnymalParams* params, const OigeState* state, const OigeAnymalBuffers* buf,
                                 const OigeResetInit* init, const float* actions_in, float clip_actions,
                                 const float* rand_vel, const float* rand_cmd, uint64_t seed, uint64_t counter,
                                 int64_t env_id_offset, int64_t n_envs, void* stream) {
    if (int rc = cartpole_range("oige_anymal_reset_masked_f32", n_envs)) return rc;
    if (n_envs == 0) return (params && state && buf) ? OIGE_OK : fail(OIGE_E_NULL, "oige_anymal_reset_masked_f32: null struct pointer");
    return anymal_pre(params, state, buf, init, actions_in, clip_actions, rand_vel, rand_cmd, seed, counter,
                      env_id_offset, n_envs, (cudaStream_t)stream, false);
}

int oige_anymal_reset_idx_f32(const OigeAnymalParams* params, const OigeState* state, const OigeAnymalBuffers* buf,
                              const OigeResetInit* init, const int64_t* env_ids, int64_t k, const float* rand_vel,
                              const float* rand_cmd, uint64_t seed, uint64_t counter, int64_t env_id_offset,
                              void* stream) {
    if (int rc = cartpole_range("oige_anymal_reset_idx_f32", k)) return rc;
    if (k == 0) return (params && state && buf) ? OIGE_OK : fail(OIGE_E_NULL, "oige_anymal_reset_idx_f32: null struct pointer");
    return anymal_reset_idx(params, state, buf, init, env_ids, k, rand_vel, rand_cmd, seed, counter, env_id_offset,
                            (cudaStream_t)stream);
}

int oige_anymal_env_step_f32(const OigeModel* model, const OigeAnymalParams* params, const OigeState* state,
                             const OigeAnymalBuffers* buf, const OigeResetInit* init, const float* actions_in,
                             float clip_actions, uint64_t seed, uint64_t counter, int64_t env_id_offset,
                             int64_t n_envs, int32_t substeps, float dt, void* stream) {
    if (int rc = cartpole_range("oige_anymal_env_step_f32", n_envs)) return rc;
    if (!model || !params || !state || !buf || !init || !actions_in)
        return fail(OIGE_E_NULL, "oige_anymal_env_step_f32: null argument");
    if (substeps < 1 || substeps > 64) return fail(OIGE_E_RANGE, "substeps %d not in [1,64]", substeps);
    if (state->dof_pos_targets == nullptr || state->dof_pos_targets != buf->current_targets)
        return fail(OIGE_E_NULL, "oige_anymal_env_step_f32: state.dof_pos_targets must be buf.current_targets");
    if (n_envs == 0) return OIGE_OK;
    cudaStream_t s = (cudaStream_t)stream;
    if (int rc = anymal_pre(params, state, buf, init, actions_in, clip_actions, nullptr, nullptr, seed, counter,
                            env_id_offset, n_envs, s, buf->reset_counter != nullptr))
        return rc;
    OigeTaskParams unit;  // K2 reads num_dof, joint_gears and power_scale; zero gears = no effort actuation,
    memset(&unit, 0, sizeof(unit));  // the joints are driven by the PD position drive toward current_targets
    unit.num_dof = OIGE_ANYMAL_DOF;
    unit.power_scale = 1.0f;
    if (int rc = k2_step(model, &unit, state, buf->actions, n_envs, substeps, dt, s, 0)) return rc;
    return anymal_post(params, state, buf, n_envs, s, true);
}

// ---- observation / action noise (utils/domain_randomization/randomize.py:212-305) -------------------------------
static int noise_spec_ok(const char* who, const OigeNoiseSpec* sp, bool interval) {
    if (!sp || !sp->enabled) return OIGE_OK;
    if (sp->operation != OIGE_DR_ADDITIVE && sp->operation != OIGE_DR_SCALING)
        return fail(OIGE_E_RANGE, "%s: operation %d is not additive/scaling", who, sp->operation);
    if (sp->distribution < OIGE_DR_GAUSSIAN || sp->distribution > OIGE_DR_LOGUNIFORM)
        return fail(OIGE_E_RANGE, "%s: distribution %d is not gaussian/uniform/loguniform", who, sp->distribution);
    if (sp->distribution == OIGE_DR_LOGUNIFORM && !(sp->p0 > 0.f && sp->p1 > 0.f))
        return fail(OIGE_E_RANGE, "%s: loguniform bounds must be positive", who);
    if (interval && sp->frequency_interval < 0) return fail(OIGE_E_RANGE, "%s: negative frequency_interval", who);
    return OIGE_OK;
}

int oige_dr_apply_f32(const OigeNoiseSpec* on_reset, const OigeNoiseSpec* on_interval, const OigeNoiseState* state,
                      const float* in, float* out, const int64_t* reset_buf, int64_t n_envs, int32_t dim, float clip_in,
                      float clip_out, const float* noise_reset, const float* noise_interval, uint64_t seed, uint64_t counter,
                      const uint64_t* counter_dev, int32_t stream_id, int64_t env_id_offset, void* stream) {
    const char* who = "oige_dr_apply_f32";
    if (n_envs < 0 || n_envs > (int64_t)1 << 31) return fail(OIGE_E_RANGE, "%s: n_envs %lld out of range", who, (long long)n_envs);
    if (dim < 1 || dim > 4096) return fail(OIGE_E_RANGE, "%s: dim %d not in [1,4096]", who, dim);
    if (!state || !state->counter || !in || !out || !reset_buf) return fail(OIGE_E_NULL, "%s: null argument", who);
    if (int rc = noise_spec_ok(who, on_reset, false)) return rc;
    if (int rc = noise_spec_ok(who, on_interval, true)) return rc;
    if (on_reset && on_reset->enabled && !state->correlated)
        return fail(OIGE_E_NULL, "%s: on_reset noise needs state.correlated", who);
    if (!(clip_in >= 0.f) || !(clip_out >= 0.f)) return fail(OIGE_E_RANGE, "%s: clip_in / clip_out must be >= 0 (INFINITY = no clamp)", who);
    if (stream_id != OIGE_DR_STREAM_ACTIONS && stream_id != OIGE_DR_STREAM_OBSERVATIONS)
        return fail(OIGE_E_RANGE, "%s: stream_id %d", who, stream_id);
    if (n_envs == 0) return OIGE_OK;
    return dr_apply(on_reset, on_interval, state, in, out, reset_buf, n_envs, dim, clip_in, clip_out, noise_reset, noise_interval, seed,
                    counter, counter_dev, stream_id, env_id_offset, (cudaStream_t)stream);
}

}  // extern "C"
