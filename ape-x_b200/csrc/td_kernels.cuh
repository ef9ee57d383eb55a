// td_kernels.cuh — K5: the arithmetic of utils.compute_loss after the three Q forwards
// (utils.py:71-80), producing the priorities that feed update_priorities (ApeX.py:58,63).
//
// float32 throughout, one IEEE operation per reference operation (torch eager issues one
// kernel per op, so nothing is fused there either): explicit __fmul_rn/__fadd_rn keep
// nvcc from contracting them into FMAs.
#pragma once
#include "common.cuh"

namespace perb {

struct TdArgs {
    const float *q, *q_next, *q_tgt;  // [batch, nA] row-major
    int64_t nA;
    const int64_t *act;
    const float *rew, *done, *w;
    float gamma_n;
    int64_t batch;
    float *out_td;    // |y - q_a|                     (scratch or caller's)
    float *out_prio;  // 0.9*max(td) + 0.1*td + 1e-6   (scratch or caller's)
    float *out_loss;  // mean(huber(td) * w), may be null
    float *red;       // [0] running max of td, [1] running sum of loss terms
    u64 *status;      // device status word (PER_STATUS_BAD_INDEX for an action outside [0, nA)); may be null
};

// The action column indexes the Q row (q.gather(1, actions), utils.py:71): an action outside [0, nA) would read
// another row; it is reported through the status word and row 0's column is used instead.
__device__ __forceinline__ int64_t td_action(const TdArgs &a, int64_t b) {
    const int64_t act = a.act[b];
    if (act < 0 || act >= a.nA) {
        if (a.status) atomicOr(a.status, (u64)PER_STATUS_BAD_INDEX);
        return 0;
    }
    return act;
}

// |td| of row b  (utils.py:71-76)
__device__ __forceinline__ float td_row(const TdArgs &a, int64_t b) {
    const float *qn = a.q_next + b * a.nA;
    int64_t best = 0;
    float bv = qn[0];
    for (int64_t i = 1; i < a.nA; ++i) {  // first maximum wins, like argmax
        float x = qn[i];
        if (x > bv) { bv = x; best = i; }
    }
    const float q_a = a.q[b * a.nA + td_action(a, b)];
    const float boot = a.q_tgt[b * a.nA + best];
    const float y = __fadd_rn(a.rew[b], __fmul_rn(__fmul_rn(a.gamma_n, boot), __fsub_rn(1.0f, a.done[b])));
    return fabsf(__fsub_rn(y, q_a));
}

__device__ __forceinline__ float td_prio(float td, float mx) {  // utils.py:77
    return __fadd_rn(__fadd_rn(__fmul_rn(0.9f, mx), __fmul_rn(0.1f, td)), 1e-6f);
}
__device__ __forceinline__ float td_huber(float td) {  // utils.py:79
    return td < 1.0f ? __fmul_rn(__fmul_rn(0.5f, td), td) : __fsub_rn(td, 0.5f);
}

// max that keeps a NaN, like torch.max (utils.py:77: one NaN TD error makes every priority of the batch NaN, and
// update_priorities then refuses them, memory.py:315); fmaxf alone would drop it
__device__ __forceinline__ float max_keep_nan(float x, float y) { return (x != x) ? x : ((y != y) ? y : fmaxf(x, y)); }
__device__ __forceinline__ float warp_max_f32(float x) {
    for (int o = 16; o; o >>= 1) x = max_keep_nan(x, __shfl_xor_sync(0xffffffffu, x, o));
    return x;
}
__device__ __forceinline__ float warp_sum_f32(float x) {
    for (int o = 16; o; o >>= 1) x += __shfl_xor_sync(0xffffffffu, x, o);
    return x;
}

// pass 1: td per row + batch max (grid-wide, through an integer atomicMax on the bits)
__global__ void k_td_pass1(TdArgs a) {
    float mx = 0.0f;
    for (int64_t b = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; b < a.batch; b += (int64_t)gridDim.x * blockDim.x) {
        float td = td_row(a, b);
        a.out_td[b] = td;
        mx = max_keep_nan(mx, td);
    }
    mx = warp_max_f32(mx);
    if ((threadIdx.x & 31) == 0) atomic_max_pos_f32(a.red, mx);
}

// pass 2: priorities + loss terms
__global__ void k_td_pass2(TdArgs a) {
    const float mx = __ldcg(a.red);
    float acc = 0.0f;
    for (int64_t b = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; b < a.batch; b += (int64_t)gridDim.x * blockDim.x) {
        float td = a.out_td[b];
        a.out_prio[b] = td_prio(td, mx);
        if (a.out_loss) acc += __fmul_rn(td_huber(td), a.w[b]);
    }
    if (a.out_loss) {
        acc = warp_sum_f32(acc);
        if ((threadIdx.x & 31) == 0) atomicAdd(a.red + 1, acc);
    }
}

__global__ void k_td_finish(TdArgs a) {
    if (a.out_loss) a.out_loss[0] = a.red[1] / (float)a.batch;  // .mean()  (utils.py:80)
}

__global__ void k_td_reset(float *red) {
    red[0] = 0.0f;
    red[1] = 0.0f;
}

}  // namespace perb
