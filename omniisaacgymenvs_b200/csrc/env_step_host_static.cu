// oige_env_step_host_static_f32: the host-buffer env step for callers that bring NO device scratch.
//
// oige_env_step_host_f32 (capi.cu) asks its caller for `actions_dev_staging`, an [N,D] device buffer that is neither
// an input nor an output of the step: the host's action rows land there before the step kernel reads them.  The side
// of a trainer that calls the host-buffer step is the side that lives in host memory -- a policy loop holding numpy /
// pinned arrays behind a ctypes or cgo binding, with no CUDA runtime of its own to ask for device memory -- and even
// a torch caller should not have this library call cudaMalloc on its behalf in the step loop: torch's caching
// allocator may already hold the HBM (a raw cudaMalloc then fails) and cudaMalloc stalls every stream of the device.
// Here the staging rows live in a fixed array inside the library image, so the call allocates nothing, anywhere.
//
// What that costs, and the rules that follow from it:
//   * OIGE_STATIC_STAGING_BYTES (8 MiB) of HBM are taken on every GPU this library is loaded on, until the process
//     exits, whether or not this entry point is ever called.
//   * n_envs * num_dof may not exceed OIGE_STATIC_STAGING_FLOATS; larger calls are refused (OIGE_E_RANGE), never
//     truncated.
//   * There is ONE array per GPU and every call on that GPU uses it.  Two calls whose GPU work can overlap (different
//     streams, different host threads, two envs stepped concurrently) overwrite each other's action rows: keep calls
//     on one GPU ordered, or use oige_env_step_host_f32 with a staging buffer per env.
//   * The array keeps whatever the previous call left in it.  Nothing in it has to start at zero: the rows
//     [0, n_envs * num_dof) are overwritten by this call's H2D copy before the step kernel reads them, and the kernel
//     reads no row beyond them.
//   * The FIRST call on each GPU looks up the array's device address (cudaGetSymbolAddress, with that GPU current)
//     and caches it; that look-up may initialise the CUDA context / load the module, so make the first call on a GPU
//     -- n_envs = 0 is enough -- before any stream capture is begun.
//
// The step itself is the existing one: after the H2D copy this file calls oige_env_step_f32 on the caller's stream
// and copies obs / rew / reset_buf back, so the bits are those of oige_env_step_host_f32 and oige_env_step_f32.
// Unlike oige_env_step_host_f32 it does not split the env range over library-owned streams (creating those would be
// an allocation too): copies and kernel run back to back on `stream`.
#include <mutex>

#include "oige_common.cuh"

namespace oige {

// The staging rows.  In HOST code the identifier `g_static_staging` is only a handle for the symbol, not a device
// address -- handing it to a kernel or to cudaMemcpy as a pointer would address the wrong memory.  Host code below
// therefore only ever uses the address cudaGetSymbolAddress returns for the current device.
static __device__ __align__(256) float g_static_staging[OIGE_STATIC_STAGING_FLOATS];

// device address of g_static_staging on the CURRENT device, looked up once per device
static int static_staging_ptr(float** out) {
    static float* ptrs[64] = {nullptr};
    static std::mutex mu;  // the C ABI may be entered from several host threads
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return fail((int)e, "oige_env_step_host_static_f32: cudaGetDevice: %s", cudaGetErrorString(e));
    if (dev < 0 || dev >= 64) return fail(OIGE_E_NODEVICE, "oige_env_step_host_static_f32: device ordinal %d not in [0,64)", dev);
    std::lock_guard<std::mutex> lock(mu);
    if (!ptrs[dev]) {
        void* p = nullptr;
        e = cudaGetSymbolAddress(&p, g_static_staging);  // resolves the symbol in the CURRENT device's copy of the module
        if (e != cudaSuccess || p == nullptr) {
            cudaGetLastError();
            return fail(e != cudaSuccess ? (int)e : (int)cudaErrorInvalidSymbol,
                        "oige_env_step_host_static_f32: cudaGetSymbolAddress(static staging) on device %d: %s", dev,
                        cudaGetErrorString(e));
        }
        ptrs[dev] = static_cast<float*>(p);
    }
    *out = ptrs[dev];
    return OIGE_OK;
}

}  // namespace oige

using namespace oige;

extern "C" int oige_env_step_host_static_f32(const OigeModel* model, const OigeTaskParams* params, const OigeState* state,
                                             const OigeTaskBuffers* buf, const OigeResetInit* init,
                                             const float* actions_host, float* obs_host, float* rew_host,
                                             int64_t* reset_host, float clip_actions, uint64_t seed, uint64_t counter,
                                             int64_t env_id_offset, int64_t n_envs, int32_t substeps, float dt,
                                             void* stream) {
    if (!model || !params || !state || !buf || !init)
        return fail(OIGE_E_NULL, "oige_env_step_host_static_f32: null struct pointer");
    if (!actions_host || !obs_host || !rew_host || !reset_host)
        return fail(OIGE_E_NULL, "oige_env_step_host_static_f32: null host buffer");
    if (params->num_dof < 1 || params->num_dof > OIGE_MAX_DOF)
        return fail(OIGE_E_SHAPE, "oige_env_step_host_static_f32: num_dof %d", params->num_dof);
    if (n_envs < 0) return fail(OIGE_E_RANGE, "oige_env_step_host_static_f32: n_envs %lld out of range", (long long)n_envs);
    const int64_t D = params->num_dof, OBS = params->num_obs;
    if (n_envs > (int64_t)OIGE_STATIC_STAGING_FLOATS / D)
        return fail(OIGE_E_RANGE,
                    "oige_env_step_host_static_f32: %lld envs x %lld DoF = %lld action floats exceed the library's static "
                    "staging array (OIGE_STATIC_STAGING_FLOATS = %lld, i.e. at most %lld envs of this robot per call); "
                    "use oige_env_step_host_f32 with a caller-allocated staging buffer",
                    (long long)n_envs, (long long)D, (long long)(n_envs * D), (long long)OIGE_STATIC_STAGING_FLOATS,
                    (long long)(OIGE_STATIC_STAGING_FLOATS / D));
    float* staging = nullptr;
    if (int rc = static_staging_ptr(&staging)) return rc;  // first call on this device: address look-up (see the header)
    if (n_envs == 0) return OIGE_OK;
    cudaStream_t s = (cudaStream_t)stream;
    cudaError_t e = cudaMemcpyAsync(staging, actions_host, (size_t)(n_envs * D) * sizeof(float), cudaMemcpyHostToDevice, s);
    if (e != cudaSuccess) return fail((int)e, "oige_env_step_host_static_f32: H2D actions: %s", cudaGetErrorString(e));
    if (int rc = oige_env_step_f32(model, params, state, buf, init, staging, clip_actions, seed, counter, env_id_offset,
                                   n_envs, substeps, dt, stream))
        return rc;
    e = cudaMemcpyAsync(obs_host, buf->obs, (size_t)(n_envs * OBS) * sizeof(float), cudaMemcpyDeviceToHost, s);
    if (e == cudaSuccess) e = cudaMemcpyAsync(rew_host, buf->rew, (size_t)n_envs * sizeof(float), cudaMemcpyDeviceToHost, s);
    if (e == cudaSuccess)
        e = cudaMemcpyAsync(reset_host, buf->reset_buf, (size_t)n_envs * sizeof(int64_t), cudaMemcpyDeviceToHost, s);
    if (e != cudaSuccess) return fail((int)e, "oige_env_step_host_static_f32: D2H obs/rew/reset: %s", cudaGetErrorString(e));
    return OIGE_OK;
}
