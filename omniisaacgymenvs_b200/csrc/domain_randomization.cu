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
// 112 B for 8 actions).  A CTA owns a tile of E consecutive envs = one contiguous chunk of every [N,dim] array;
// a thread owns (env, block of 4 columns): one Philox4x32-10 call yields the 4 values of a block.
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
};

// the 4 noise values of (env, block) for one stream; mirrors oracle/dr_oracle.py:philox_noise
__device__ __forceinline__ void philox_block(const OigeNoiseSpec& sp, unsigned long long seed, unsigned long long ctr,
                                             unsigned long long gid, int blk, int stream, float v[4]) {
    uint32_t r[4];
    philox4x32_10((uint32_t)gid, (uint32_t)(gid >> 32), (uint32_t)ctr, ((uint32_t)(ctr >> 32) << 8) ^ (uint32_t)blk,
                  (uint32_t)seed, (uint32_t)(seed >> 32) ^ 0x44520000u ^ (uint32_t)stream, r);
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

__global__ void __launch_bounds__(NT) dr_apply_kernel(const __grid_constant__ DrArgs A) {
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
    const int dim = A.dim, nblk = (dim + 3) >> 2;
    for (int item = tid; item < cnt * nblk; item += NT) {
        const int el = item / nblk, blk = item - el * nblk;
        const int c0 = blk * 4, w = min(4, dim - c0);
        const long long base = (e0 + el) * dim + c0;
        const unsigned flag = s_flag[el];
        float b[4];
#pragma unroll
        for (int j = 0; j < 4; ++j)
            if (j < w) b[j] = fminf(fmaxf(A.in[base + j], -A.clip_in), A.clip_in);
        if (A.on_reset.enabled) {
            float corr[4];
            if (flag & 1u) {
                if (A.noise_reset != nullptr) {
#pragma unroll
                    for (int j = 0; j < 4; ++j)
                        if (j < w) corr[j] = A.noise_reset[base + j];
                } else {
                    philox_block(A.on_reset, A.seed, ctr, (unsigned long long)(A.env_id_offset + e0 + el), blk, A.stream_id, corr);
                }
#pragma unroll
                for (int j = 0; j < 4; ++j)
                    if (j < w) A.correlated[base + j] = corr[j];
            } else {
#pragma unroll
                for (int j = 0; j < 4; ++j)
                    if (j < w) corr[j] = A.correlated[base + j];
            }
#pragma unroll
            for (int j = 0; j < 4; ++j) b[j] = apply_op(A.on_reset.operation, b[j], corr[j]);
        }
        if (flag & 2u) {
            float nz[4];
            if (A.noise_interval != nullptr) {
#pragma unroll
                for (int j = 0; j < 4; ++j)
                    if (j < w) nz[j] = A.noise_interval[base + j];
            } else {
                philox_block(A.on_interval, A.seed, ctr, (unsigned long long)(A.env_id_offset + e0 + el), blk, A.stream_id + 1, nz);
            }
#pragma unroll
            for (int j = 0; j < 4; ++j) b[j] = apply_op(A.on_interval.operation, b[j], nz[j]);
        }
#pragma unroll
        for (int j = 0; j < 4; ++j)
            if (j < w) A.out[base + j] = fminf(fmaxf(b[j], -A.clip_out), A.clip_out);
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
    const unsigned grid = (unsigned)((n + E - 1) / E);
    dr_apply_kernel<<<grid, NT, 0, stream>>>(A);
    count_launch();
    return check_launch("dr_apply_kernel");
}

}  // namespace oige
