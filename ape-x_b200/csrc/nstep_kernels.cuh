// nstep_kernels.cuh — K7: batched n-step assembly, the arithmetic of BatchStorage
// (memory.py:393-478) for many actor streams at once.
//
// A stream is one actor's consecutive environment steps (reward, action, done, Q row); the
// reference feeds them one at a time to BatchStorage.add (memory.py:415-440).  What add()
// emits at step t depends only on
//   * len  = how many steps have been appended since the last `done` (capped at n):
//            emit iff len == n or done[t]                                   (memory.py:416)
//   * the first step of that window, t - len                                (memory.py:417-420)
//   * the return  sum_k r[t-len+k] * gamma**k  (k = 0..len-1)  +  r[t] * gamma**len — len+1
//     terms, added in that order from 0.0                                   (memory.py:418,471-475)
//   * the OLDEST entry of the Q window, which unlike the other windows is NOT cleared on
//     done and skips done steps: the n-th most recent appended step, or the oldest one
//     appended so far                                                       (memory.py:420,432-440)
// so one warp walks a stream in tiles of 32 steps (lane = step) and gets all of it from a
// ballot of the done flags plus three carried integers; the actor priority
// |R + gamma**n * max_a q[t] * (1 - done) - q0[action]| + 1e-6 (memory.py:451-464) follows with
// numpy's promotion rules: the bootstrap product in float32, the rest in float64.
//
// gamma**k comes from a host table (CPython's float pow), so returns are bit-identical.
#pragma once
#include "common.cuh"

namespace perb {

constexpr int NSTEP_MAX_N = 64;
constexpr int NSTEP_RING = 128;  // >= NSTEP_MAX_N + 32

struct NstepArgs {
    int64_t n_streams;
    const int64_t *offsets;    // [n_streams + 1] step ranges
    const int64_t *emit_from;  // [n_streams] first step (absolute) whose emissions are reported; null = all
    const double *reward;      // [T]
    const int64_t *action;     // [T]
    const uint8_t *done;       // [T]
    const float *q;            // [T, nA]
    int64_t nA;
    int n;                     // n_steps
    double gpow[NSTEP_MAX_N + 1];  // gamma ** k, k = 0..n
    // outputs, one slot per step of the stream at most (slot = offsets[s] + emission number)
    int64_t *out_first;        // step index of the window's first state  (states[t0])
    int64_t *out_next;         // step index of the bootstrap state       (next_states = state at t)
    int64_t *out_action;
    double *out_return;
    float *out_done;
    double *out_prio;
    int64_t *out_count;        // [n_streams]
    u64 *status;               // PER_STATUS_NSTEP_EMPTY: done arrived on an empty window (IndexError in the reference)
};

__global__ void __launch_bounds__(128) k_nstep(NstepArgs a) {
    __shared__ int64_t s_ring[4][NSTEP_RING];  // per warp: step index of the Q-window entries by append ordinal
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    const int64_t s = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
    if (s >= a.n_streams) return;
    int64_t *ring = s_ring[wib];
    const int64_t lo = a.offsets[s], hi = a.offsets[s + 1];
    const int64_t from = a.emit_from ? a.emit_from[s] : lo;
    int64_t appended = 0;       // Q-window appends so far (non-done steps)
    int64_t last_clear = lo;    // first step after the most recent done
    int64_t emitted = 0;
    const float disc = (float)a.gpow[a.n];  // numpy multiplies the float32 Q by float32(gamma**n)
    for (int64_t base = lo; base < hi; base += 32) {
        const int64_t t = base + lane;
        const bool live = t < hi;
        const bool d = live && a.done[t] != 0;
        const unsigned dmask = __ballot_sync(0xffffffffu, d);
        const unsigned below = (1u << lane) - 1u;
        // steps appended since the last done, before t
        const unsigned dprev = dmask & below;
        const int64_t clear_at = dprev ? base + (31 - __clz(dprev)) + 1 : last_clear;
        int64_t len = t - clear_at;
        if (len > a.n) len = a.n;
        // Q window: publish this tile's appends, then look up the oldest entry
        const unsigned live_mask = __ballot_sync(0xffffffffu, live);
        const unsigned amask = ~dmask & live_mask;  // appended steps of this tile
        const int64_t ord = appended + __popc(amask & below);  // appends before t
        if (live && !d) ring[ord % NSTEP_RING] = t;
        __syncwarp();
        const int64_t back = ord < a.n ? ord : a.n;  // entries in the Q window when t arrives
        const bool emit = live && (len == a.n || d);
        const bool empty = emit && (len == 0 || back == 0);
        if (empty) atomicOr(a.status, (u64)PER_STATUS_NSTEP_EMPTY);
        const bool report = emit && !empty && t >= from;
        const unsigned rmask = __ballot_sync(0xffffffffu, report);
        if (report) {
            const int64_t t0 = t - len;
            const int64_t tq = ring[(ord - back) % NSTEP_RING];
            double ret = 0.0;
            for (int k = 0; k < (int)len; ++k) ret = __dadd_rn(ret, __dmul_rn(a.reward[t0 + k], a.gpow[k]));
            ret = __dadd_rn(ret, __dmul_rn(a.reward[t], a.gpow[len]));
            const float *qt = a.q + t * a.nA;
            float mx = qt[0];
            for (int64_t i = 1; i < a.nA; ++i) mx = fmaxf(mx, qt[i]);
            const int64_t act = a.action[t0];
            const float dn = d ? 1.0f : 0.0f;
            const float boot = __fmul_rn(__fmul_rn(disc, mx), __fsub_rn(1.0f, dn));
            const double y = __dadd_rn(ret, (double)boot);
            const double prio = __dadd_rn(fabs(__dsub_rn(y, (double)a.q[tq * a.nA + act])), 1e-6);
            const int64_t slot = lo + emitted + __popc(rmask & below);
            a.out_first[slot] = t0;
            a.out_next[slot] = t;
            a.out_action[slot] = act;
            a.out_return[slot] = ret;
            a.out_done[slot] = dn;
            a.out_prio[slot] = prio;
        }
        __syncwarp();
        emitted += __popc(rmask);
        appended += __popc(amask);
        if (dmask) last_clear = base + (31 - __clz(dmask)) + 1;
    }
    if (lane == 0) a.out_count[s] = emitted;
}

}  // namespace perb
