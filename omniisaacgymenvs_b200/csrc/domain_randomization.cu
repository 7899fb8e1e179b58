// Observation / action noise (reference: utils/domain_randomization/randomize.py:212-305, hooks in
// envs/vec_env_rlgames.py:59-60,70-71).  One launch per randomised buffer replaces, per hook, the reference's
// nonzero() host sync, counter bookkeeping, 1-2 RNG launches, gathers/scatters and the whole-buffer add/mul:
//
//   env_ids = reset_buf.nonzero(); counter[env_ids] = 0; counter += 1                     (:213-215)
//   on_reset   : correlated[env_ids] = draw(k, dim);  buffer (+= | *=) correlated          (:283-305)
//   on_interval: ids = counter >= frequency_interval; counter[ids] = 0;
//                buffer[ids] (+= | *=) draw(len(ids), dim)                                  (:227-236, :267-281)
//
// HBM-bound.  Algorithmic bytes per env-step with both blocks active: buffer read + written (8*dim), correlated read
// (4*dim, + 4*dim written on reset), reset flag 8, counter 4 + 4  =>  12*dim + 16 B  (1060 B for 87 observations,
// 112 B for 8 actions).  A CTA owns a tile of E consecutive envs = one contiguous, 16-byte aligned chunk of every
// [N,dim] array (row length need not be a multiple of 4); a thread moves 16-byte vectors of that chunk and one
// Philox4x32-10 call yields the 4 noise values of a vector.
#include <math_constants.h>

#include "oige_common.cuh"

namespace oige {
namespace {

constexpr int E = 64;     // envs per tile
constexpr int NT = 256;   // threads per CTA

struct DrArgs {
    OigeNoiseSpec on_reset, on_interval;
    int* env_counter; float* correlated;
    const float* in; float* out;
    const long long* reset_buf;
    const float* noise_reset; const float* noise_interval;
    const unsigned long long* counter_dev;
    unsigned long long seed, counter;
    long long env_id_offset, n;
    int dim, stream_id;
    float clip_in, clip_out;
    int vec_ok;  // every [N,dim] pointer is 16-byte aligned
};

// The 4 noise values of one GROUP = 4 consecutive elements of the flattened GLOBAL [N_global, dim] array
// (element (env, col) has flat index gid * dim + col, group = flat >> 2, component = flat & 3), so a value depends only
// on (seed, counter, stream, global env id, column) -- not on sharding or tiling -- and a 16-byte vector of the
// buffer needs exactly one Philox call.  Mirrors oracle/dr_oracle.py:philox_noise.
__device__ __forceinline__ void philox_group(const OigeNoiseSpec& sp, unsigned long long seed, unsigned long long ctr,
                                             unsigned long long group, int stream, float v[4]) {
    uint32_t r[4];
    philox4x32_10((uint32_t)group, (uint32_t)(group >> 32), (uint32_t)ctr, (uint32_t)(ctr >> 32), (uint32_t)seed,
                  (uint32_t)(seed >> 32) ^ 0x44520000u ^ (uint32_t)stream, r);
    if (sp.distribution == OIGE_DR_GAUSSIAN) {
#pragma unroll
        for (int h = 0; h < 2; ++h) {  // Box-Muller, radius uniform in (0, 1]
            const float u1 = ((float)(r[2 * h] >> 8) + 1.0f) * (1.0f / 16777216.0f);
            const float u2 = u01(r[2 * h + 1]);
            const float rad = sqrtf(-2.0f * logf(u1));
            const float ang = 6.2831853071795864f * u2;
            float s, c;
            sincosf(ang, &s, &c);
            v[2 * h] = sp.p0 + sp.p1 * (rad * c);
            v[2 * h + 1] = sp.p0 + sp.p1 * (rad * s);
        }
    } else if (sp.distribution == OIGE_DR_UNIFORM) {
#pragma unroll
        for (int j = 0; j < 4; ++j) v[j] = (sp.p1 - sp.p0) * u01(r[j]) + sp.p0;
    } else {
        const float l0 = logf(sp.p0), l1 = logf(sp.p1);
#pragma unroll
        for (int j = 0; j < 4; ++j) v[j] = expf((l1 - l0) * u01(r[j]) + l0);
    }
}

__device__ __forceinline__ float apply_op(int op, float b, float noise) { return op == OIGE_DR_ADDITIVE ? b + noise : b * noise; }
__device__ __forceinline__ float clampf(float x, float c) { return fminf(fmaxf(x, -c), c); }

// one element, any alignment: used for ragged tails and for launches whose global flat index is not 16-byte phased
__device__ __forceinline__ float dr_element(const DrArgs& A, unsigned long long ctr, long long idx, unsigned flag,
                                            unsigned long long gflat) {
    float b = clampf(A.in[idx], A.clip_in);
    if (A.on_reset.enabled) {
        float corr;
        if (flag & 1u) {
            if (A.noise_reset != nullptr) {
                corr = A.noise_reset[idx];
            } else {
                float v[4];
                philox_group(A.on_reset, A.seed, ctr, gflat >> 2, A.stream_id, v);
                corr = v[gflat & 3];
            }
            A.correlated[idx] = corr;
        } else {
            corr = A.correlated[idx];
        }
        b = apply_op(A.on_reset.operation, b, corr);
    }
    if (flag & 2u) {
        float nz;
        if (A.noise_interval != nullptr) {
            nz = A.noise_interval[idx];
        } else {
            float v[4];
            philox_group(A.on_interval, A.seed, ctr, gflat >> 2, A.stream_id + 1, v);
            nz = v[gflat & 3];
        }
        b = apply_op(A.on_interval.operation, b, nz);
    }
    return clampf(b, A.clip_out);
}

__global__ void __launch_bounds__(NT, 6) dr_apply_kernel(const __grid_constant__ DrArgs A) {
    __shared__ unsigned char s_flag[E];  // bit 0: env was reset, bit 1: interval noise due
    const int tid = threadIdx.x;
    const long long e0 = (long long)blockIdx.x * E;
    const int cnt = (int)min((long long)E, A.n - e0);
    if (tid < cnt) {
        const long long e = e0 + tid;
        const bool rst = A.reset_buf[e] != 0;
        int c = rst ? 0 : A.env_counter[e];
        c += 1;
        bool due = false;
        if (A.on_interval.enabled && c >= A.on_interval.frequency_interval) { due = true; c = 0; }
        A.env_counter[e] = c;
        s_flag[tid] = (unsigned char)((rst ? 1 : 0) | (due ? 2 : 0));
    }
    __syncthreads();
    const unsigned long long ctr = A.counter_dev != nullptr ? *A.counter_dev : A.counter;
    const int dim = A.dim;
    const long long base = e0 * dim;                                            // first element of the tile
    const unsigned long long gbase = (unsigned long long)(A.env_id_offset + e0) * (unsigned long long)dim;
    const int F = cnt * dim;
    // 16-byte path: the tile starts on a 16-byte boundary of every array (E % 4 == 0) and, if the shard offset keeps
    // the global flat index a multiple of 4 as well, vector v of the tile is exactly Philox group (gbase >> 2) + v
    const int n4 = (A.vec_ok && (gbase & 3ull) == 0ull) ? (F >> 2) : 0;
    const float4* in4 = reinterpret_cast<const float4*>(A.in + base);
    float4* corr4 = reinterpret_cast<float4*>(A.correlated + base);
    float4* out4 = reinterpret_cast<float4*>(A.out + base);
    const float4* nr4 = reinterpret_cast<const float4*>(A.noise_reset + base);
    const float4* ni4 = reinterpret_cast<const float4*>(A.noise_interval + base);
    const float rdim = 1.0f / (float)dim;
    for (int v = tid; v < n4; v += NT) {
        {
            const float4 bi = in4[v];
            const int i0 = 4 * v;
            int el = (int)((float)i0 * rdim);            // env of element i0 (float estimate, fixed up below)
            el += (el + 1) * dim <= i0 ? 1 : 0;
            el -= el * dim > i0 ? 1 : 0;
            const int row_end = (el + 1) * dim;           // elements >= row_end belong to env el + 1
            unsigned fl[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) fl[j] = s_flag[min(el + (i0 + j >= row_end ? 1 : 0), cnt - 1)];
            const unsigned any = fl[0] | fl[1] | fl[2] | fl[3];
            float b[4] = {clampf(bi.x, A.clip_in), clampf(bi.y, A.clip_in), clampf(bi.z, A.clip_in), clampf(bi.w, A.clip_in)};
            const unsigned long long group = (gbase >> 2) + (unsigned long long)v;
            if (A.on_reset.enabled) {
                const float4 c4 = corr4[v];
                float corr[4] = {c4.x, c4.y, c4.z, c4.w};
                if (any & 1u) {
                    float nv[4];
                    if (A.noise_reset != nullptr) {
                        const float4 t = nr4[v];
                        nv[0] = t.x; nv[1] = t.y; nv[2] = t.z; nv[3] = t.w;
                    } else {
                        philox_group(A.on_reset, A.seed, ctr, group, A.stream_id, nv);
                    }
#pragma unroll
                    for (int j = 0; j < 4; ++j) corr[j] = (fl[j] & 1u) ? nv[j] : corr[j];
                    corr4[v] = make_float4(corr[0], corr[1], corr[2], corr[3]);
                }
#pragma unroll
                for (int j = 0; j < 4; ++j) b[j] = apply_op(A.on_reset.operation, b[j], corr[j]);
            }
            if (any & 2u) {
                float nz[4];
                if (A.noise_interval != nullptr) {
                    const float4 t = ni4[v];
                    nz[0] = t.x; nz[1] = t.y; nz[2] = t.z; nz[3] = t.w;
                } else {
                    philox_group(A.on_interval, A.seed, ctr, group, A.stream_id + 1, nz);
                }
#pragma unroll
                for (int j = 0; j < 4; ++j) b[j] = (fl[j] & 2u) ? apply_op(A.on_interval.operation, b[j], nz[j]) : b[j];
            }
            out4[v] = make_float4(clampf(b[0], A.clip_out), clampf(b[1], A.clip_out), clampf(b[2], A.clip_out),
                                  clampf(b[3], A.clip_out));
        }
    }
    for (int i = 4 * n4 + tid; i < F; i += NT) {  // ragged tail of the last tile, or the whole tile when unphased
        const int el = i / dim;
        A.out[base + i] = dr_element(A, ctr, base + i, s_flag[el], gbase + (unsigned long long)i);
    }
}

}  // namespace

int dr_apply(const OigeNoiseSpec* on_reset, const OigeNoiseSpec* on_interval, const OigeNoiseState* st, const float* in,
             float* out, const int64_t* reset_buf, int64_t n, int32_t dim, float clip_in, float clip_out, const float* noise_reset,
             const float* noise_interval, uint64_t seed, uint64_t counter, const uint64_t* counter_dev, int32_t stream_id,
             int64_t env_id_offset, cudaStream_t stream) {
    DrArgs A{};
    if (on_reset) A.on_reset = *on_reset;
    if (on_interval) A.on_interval = *on_interval;
    A.env_counter = st->counter; A.correlated = st->correlated;
    A.in = in; A.out = out; A.reset_buf = (const long long*)reset_buf;
    A.noise_reset = noise_reset; A.noise_interval = noise_interval;
    A.counter_dev = (const unsigned long long*)counter_dev;
    A.seed = seed; A.counter = counter; A.env_id_offset = env_id_offset; A.n = n;
    A.dim = dim; A.stream_id = stream_id; A.clip_in = clip_in; A.clip_out = clip_out;
    uintptr_t bits = reinterpret_cast<uintptr_t>(in) | reinterpret_cast<uintptr_t>(out) |
                     reinterpret_cast<uintptr_t>(st->correlated) | reinterpret_cast<uintptr_t>(noise_reset) |
                     reinterpret_cast<uintptr_t>(noise_interval);
    A.vec_ok = (bits & 15u) == 0;
    const unsigned grid = (unsigned)((n + E - 1) / E);
    dr_apply_kernel<<<grid, NT, 0, stream>>>(A);
    count_launch();
    return check_launch("dr_apply_kernel");
}

}  // namespace oige
