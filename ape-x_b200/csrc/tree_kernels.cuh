// tree_kernels.cuh — float64 sum / min segment trees resident in HBM.
//
// Layout: the reference's heap (memory.py:36) — `v[1]` root, children of i at 2i, 2i+1,
// leaf k at v[cap + k]; depth d holds nodes 2^d .. 2^(d+1)-1, contiguous.
//
// Invariant kept by every kernel here (and by memory.py:82-87): each internal node equals
// op(left child, right child) of the current children.  The node array is therefore a pure
// function of the leaves, and recomputing a clean node rewrites the same bits.  All tree
// arithmetic is IEEE float64 round-to-nearest with explicit __dadd_rn/__dmul_rn so that
// nvcc cannot contract a*b+c*d into an FMA (memory.py:247 is two products and a sum).
#pragma once
#include "common.cuh"
#include "td_kernels.cuh"

namespace perb {

constexpr int TOP_LEVELS = 10;            // depths 0..TOP_LEVELS are rebuilt densely in shared memory
constexpr int TOP_NODES = 1 << TOP_LEVELS;
constexpr int UPD_BLOCK_MAX_ITEMS = 4096;  // largest batch the one-CTA update kernel takes

enum LeafSrc : int {
    LEAF_GIVEN = 0,     // float64 leaves supplied by the caller
    LEAF_PRIO_F64 = 1,  // float64 priorities, leaf = p ** alpha
    LEAF_PRIO_F32 = 2,  // float32 priorities, leaf = double(p) ** alpha
    LEAF_MAXPRIO = 3,   // leaf = max_priority ** alpha  (memory.py:239-240)
};

struct TreeRef {
    double *sum;  // may be null (stand-alone MinSegmentTree)
    double *mnn;  // may be null (stand-alone SumSegmentTree)
    int64_t cap;
    int L;  // log2(cap)
};

struct UpdArgs {
    TreeRef t;
    const int64_t *idx;       // null: ring mode, idx_j = (ring_start + j) % ring_size
    int64_t ring_start, ring_size;
    const void *src;          // leaves / priorities, see src_kind
    int src_kind;
    double alpha;
    int64_t n;
    int64_t len_limit;        // valid indices are [0, len_limit)
    u64 *stamps;              // cap entries; null when indices are known distinct
    u64 epoch_hi;             // call counter << 32
    u64 *status;
    double *max_prio;         // device scalar `_max_priority`; may be null (stand-alone trees)
    double batch_max;         // LEAF_GIVEN: raw-priority maximum of the batch (<=0: none)
};

__device__ __forceinline__ int64_t upd_index(const UpdArgs &a, int64_t j) {
    if (a.idx) return a.idx[j];
    int64_t k = a.ring_start + j;
    return k >= a.ring_size ? k - a.ring_size : k;
}

// Raw priority (for the running max) and leaf value of item j.  Returns false for an
// item the reference would have rejected (memory.py:315-316).
__device__ __forceinline__ bool upd_item(const UpdArgs &a, int64_t j, int64_t &k, double &leaf, double &raw) {
    k = upd_index(a, j);
    raw = 0.0;
    unsigned bad = 0;
    if (k < 0 || k >= a.len_limit) bad |= PER_STATUS_BAD_INDEX;
    switch (a.src_kind) {
        case LEAF_GIVEN:
            leaf = static_cast<const double *>(a.src)[j];
            break;
        case LEAF_PRIO_F64:
            raw = static_cast<const double *>(a.src)[j];
            if (!(raw > 0.0)) bad |= PER_STATUS_BAD_PRIORITY;
            leaf = pow_alpha(raw, a.alpha);
            break;
        case LEAF_PRIO_F32:
            raw = (double)static_cast<const float *>(a.src)[j];
            if (!(raw > 0.0)) bad |= PER_STATUS_BAD_PRIORITY;
            leaf = pow_alpha(raw, a.alpha);
            break;
        default:  // LEAF_MAXPRIO
            leaf = pow_alpha(ldcg_f64(a.max_prio), a.alpha);
            break;
    }
    if (bad) {
        atomicOr(a.status, (u64)bad);
        return false;
    }
    return true;
}

// Lift S levels: recompute the S generations of ancestors that sit above the 2^S
// contiguous nodes  (g << S) .. (g << S) + 2^S - 1 , where g is their common ancestor.
// Threads that share ancestors write identical bits, so duplicates are harmless.
template <int S, bool IS_SUM>
__device__ __forceinline__ void lift(double *v, int64_t g) {
    constexpr int W = 1 << S;
    double c[W];
    const double *base = v + (g << S);
#pragma unroll
    for (int i = 0; i < W; i += 2) {
        double2 x = ldcg_f64x2(base + i);
        c[i] = x.x;
        c[i + 1] = x.y;
    }
#pragma unroll
    for (int s = S - 1; s >= 0; --s) {
        double *out = v + (g << s);
#pragma unroll
        for (int i = 0; i < (1 << s); ++i) {
            c[i] = IS_SUM ? __dadd_rn(c[2 * i], c[2 * i + 1]) : fmin(c[2 * i], c[2 * i + 1]);
            out[i] = c[i];
        }
    }
}

template <int S>
__device__ __forceinline__ void lift_both(const TreeRef &t, int64_t g) {
    if (t.sum) lift<S, true>(t.sum, g);
    if (t.mnn) lift<S, false>(t.mnn, g);
}

__device__ __forceinline__ void lift_dyn(const TreeRef &t, int64_t g, int s) {
    switch (s) {
        case 1: lift_both<1>(t, g); break;
        case 2: lift_both<2>(t, g); break;
        case 3: lift_both<3>(t, g); break;
        default: lift_both<4>(t, g); break;
    }
}

// Dense rebuild, by one CTA, of the k generations above the 2^k contiguous nodes that
// descend from node `a` (k <= TOP_LEVELS).  smem: 2 * 2^k doubles per tree, addressed as a
// relative heap (generation j, position t at s[(1<<j) + t]).
__device__ __forceinline__ void dense_lift_block(const TreeRef &t, int64_t a, int k, double *s_sum, double *s_min) {
    const int n = 1 << k;
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        if (t.sum) s_sum[n + i] = ldcg_f64(t.sum + (a << k) + i);
        if (t.mnn) s_min[n + i] = ldcg_f64(t.mnn + (a << k) + i);
    }
    __syncthreads();
    for (int j = k - 1; j >= 0; --j) {
        const int m = 1 << j;
        for (int i = threadIdx.x; i < m; i += blockDim.x) {
            if (t.sum) {
                double2 c = *reinterpret_cast<double2 *>(s_sum + 2 * (m + i));
                double x = __dadd_rn(c.x, c.y);
                s_sum[m + i] = x;
                t.sum[(a << j) + i] = x;
            }
            if (t.mnn) {
                double2 c = *reinterpret_cast<double2 *>(s_min + 2 * (m + i));
                double x = fmin(c.x, c.y);
                s_min[m + i] = x;
                t.mnn[(a << j) + i] = x;
            }
        }
        __syncthreads();
    }
}

// ---------------------------------------------------------------------------------------
// K1, one-CTA form (n <= UPD_BLOCK_MAX_ITEMS): batched leaf update with sequential
// semantics + level re-reduction, all in one launch.
//   phase 1  last-write-wins arbitration: atomicMax(stamp[idx], epoch|j)   (memory.py:314-318)
//   phase 2  winners write the leaves of both trees; block max -> _max_priority (memory.py:320)
//   phase 3  sparse lifts, <=4 generations per L2 round trip, until depth TOP_LEVELS
//   phase 4  dense rebuild of depths TOP_LEVELS..0 in shared memory
// ---------------------------------------------------------------------------------------
constexpr int UPD_ITEMS_PER_THREAD = 8;
constexpr int UPD_BLOCK_THREADS = 512;

// Optional prologue (fused_td != 0): K5, the learner's TD-error -> priority arithmetic
// (utils.py:71-80), for the same batch; its float32 priorities (written to td.out_prio)
// are then the source of the leaf update below, so ApeX.py:58-63 is one launch.
__global__ void __launch_bounds__(UPD_BLOCK_THREADS, 1) k_update_block(UpdArgs a, TdArgs td, int fused_td, int skip_update) {
    __shared__ double s_sum[2 * TOP_NODES];
    __shared__ double s_min[2 * TOP_NODES];
    __shared__ double s_red[32];

    const int T = blockDim.x;
    if (fused_td) {
        float *f_red = reinterpret_cast<float *>(s_red);  // [0..31] warp partials, [32] result
        float tdv[UPD_ITEMS_PER_THREAD];
        float mx = 0.0f;
#pragma unroll
        for (int it = 0; it < UPD_ITEMS_PER_THREAD; ++it) {
            int64_t j = (int64_t)it * T + threadIdx.x;
            tdv[it] = 0.0f;
            if (j < td.batch) {
                tdv[it] = td_row(td, j);
                mx = fmaxf(mx, tdv[it]);
            }
        }
        mx = warp_max_f32(mx);
        if ((threadIdx.x & 31) == 0) f_red[threadIdx.x >> 5] = mx;
        __syncthreads();
        if (threadIdx.x < 32) {
            float m = threadIdx.x < (T + 31) / 32 ? f_red[threadIdx.x] : 0.0f;
            m = warp_max_f32(m);
            if (threadIdx.x == 0) f_red[32] = m;
        }
        __syncthreads();
        mx = f_red[32];
        float acc = 0.0f;
#pragma unroll
        for (int it = 0; it < UPD_ITEMS_PER_THREAD; ++it) {
            int64_t j = (int64_t)it * T + threadIdx.x;
            if (j < td.batch) {
                td.out_prio[j] = td_prio(tdv[it], mx);
                if (td.out_td) td.out_td[j] = tdv[it];
                if (td.out_loss) acc += __fmul_rn(td_huber(tdv[it]), td.w[j]);
            }
        }
        if (td.out_loss) {
            __syncthreads();
            acc = warp_sum_f32(acc);
            if ((threadIdx.x & 31) == 0) f_red[threadIdx.x >> 5] = acc;
            __syncthreads();
            if (threadIdx.x < 32) {
                float m = threadIdx.x < (T + 31) / 32 ? f_red[threadIdx.x] : 0.0f;
                m = warp_sum_f32(m);
                if (threadIdx.x == 0) td.out_loss[0] = m / (float)td.batch;  // .mean(), utils.py:80
            }
        }
        if (skip_update) return;
        __syncthreads();  // s_red is reused below
    }
    int64_t pos[UPD_ITEMS_PER_THREAD];  // leaf node of the items this thread won, -1 otherwise
    double leafv[UPD_ITEMS_PER_THREAD];
    double my_max = 0.0;

    // phase 1
#pragma unroll
    for (int it = 0; it < UPD_ITEMS_PER_THREAD; ++it) {
        int64_t j = (int64_t)it * T + threadIdx.x;
        pos[it] = -1;
        if (j < a.n) {
            int64_t k;
            double raw;
            if (upd_item(a, j, k, leafv[it], raw)) {
                pos[it] = a.t.cap + k;
                my_max = fmax(my_max, raw);
                if (a.stamps) atomicMax(a.stamps + k, a.epoch_hi | (u64)j);
            }
        }
    }
    if (a.stamps) {
        __syncthreads();
#pragma unroll
        for (int it = 0; it < UPD_ITEMS_PER_THREAD; ++it) {
            if (pos[it] >= 0) {
                int64_t j = (int64_t)it * T + threadIdx.x;
                u64 w = __ldcg(a.stamps + (pos[it] - a.t.cap));
                if (w != (a.epoch_hi | (u64)j)) pos[it] = -1;  // an item later in the batch overwrites this one
            }
        }
    }
    // phase 2
#pragma unroll
    for (int it = 0; it < UPD_ITEMS_PER_THREAD; ++it) {
        if (pos[it] >= 0) {
            if (a.t.sum) a.t.sum[pos[it]] = leafv[it];
            if (a.t.mnn) a.t.mnn[pos[it]] = leafv[it];
        }
    }
    if (a.max_prio) {
        if (a.src_kind == LEAF_GIVEN) my_max = (threadIdx.x == 0) ? a.batch_max : 0.0;
        for (int o = 16; o; o >>= 1) my_max = fmax(my_max, shfl_f64(my_max, (threadIdx.x & 31) ^ o));
        if ((threadIdx.x & 31) == 0) s_red[threadIdx.x >> 5] = my_max;
        __syncthreads();
        if (threadIdx.x < 32) {
            double m = threadIdx.x < (T + 31) / 32 ? s_red[threadIdx.x] : 0.0;
            for (int o = 16; o; o >>= 1) m = fmax(m, shfl_f64(m, threadIdx.x ^ o));
            if (threadIdx.x == 0 && m > 0.0) atomic_max_pos_f64(a.max_prio, m);
        }
    }
    __syncthreads();
    // phase 3
    const int top = a.t.L < TOP_LEVELS ? a.t.L : TOP_LEVELS;
    int done = 0;  // generations already lifted above the leaves
    while (a.t.L - done > top) {
        int s = a.t.L - done - top;
        if (s > 4) s = 4;
#pragma unroll
        for (int it = 0; it < UPD_ITEMS_PER_THREAD; ++it)
            if (pos[it] >= 0) lift_dyn(a.t, pos[it] >> (done + s), s);
        done += s;
        __syncthreads();
    }
    // phase 4
    dense_lift_block(a.t, 1, top, s_sum, s_min);
}

// ---------------------------------------------------------------------------------------
// K1, multi-CTA form (any n): the same phases as separate launches.
// ---------------------------------------------------------------------------------------
__global__ void k_upd_stamp(UpdArgs a) {
    for (int64_t j = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; j < a.n; j += (int64_t)gridDim.x * blockDim.x) {
        int64_t k = upd_index(a, j);
        if (k >= 0 && k < a.len_limit) atomicMax(a.stamps + k, a.epoch_hi | (u64)j);
    }
}

__device__ __forceinline__ bool upd_is_winner(const UpdArgs &a, int64_t j, int64_t k) {
    return !a.stamps || __ldcg(a.stamps + k) == (a.epoch_hi | (u64)j);
}

__global__ void k_upd_leaf(UpdArgs a) {
    double my_max = 0.0;
    for (int64_t j = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; j < a.n; j += (int64_t)gridDim.x * blockDim.x) {
        int64_t k;
        double leaf, raw;
        if (!upd_item(a, j, k, leaf, raw)) continue;
        my_max = fmax(my_max, raw);
        if (!upd_is_winner(a, j, k)) continue;
        if (a.t.sum) a.t.sum[a.t.cap + k] = leaf;
        if (a.t.mnn) a.t.mnn[a.t.cap + k] = leaf;
    }
    if (a.max_prio) {
        if (a.src_kind == LEAF_GIVEN) my_max = (blockIdx.x == 0 && threadIdx.x == 0) ? a.batch_max : 0.0;
        for (int o = 16; o; o >>= 1) my_max = fmax(my_max, shfl_f64(my_max, (threadIdx.x & 31) ^ o));
        if ((threadIdx.x & 31) == 0 && my_max > 0.0) atomic_max_pos_f64(a.max_prio, my_max);
    }
}

// Lift `s` generations above depth (L - done) for every valid item.
__global__ void k_upd_sparse(UpdArgs a, int done, int s) {
    for (int64_t j = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; j < a.n; j += (int64_t)gridDim.x * blockDim.x) {
        int64_t k = upd_index(a, j);
        if (k < 0 || k >= a.len_limit) continue;
        lift_dyn(a.t, (a.t.cap + k) >> (done + s), s);
    }
}

// CTA b rebuilds the k generations above the 2^k nodes below node (2^(d-k) + b).
__global__ void __launch_bounds__(256) k_dense_mid(TreeRef t, int d, int k) {
    __shared__ double s_sum[2 * TOP_NODES];
    __shared__ double s_min[2 * TOP_NODES];
    dense_lift_block(t, (int64_t(1) << (d - k)) + blockIdx.x, k, s_sum, s_min);
}

__global__ void __launch_bounds__(1024) k_dense_top(TreeRef t) {
    __shared__ double s_sum[2 * TOP_NODES];
    __shared__ double s_min[2 * TOP_NODES];
    dense_lift_block(t, 1, t.L < TOP_LEVELS ? t.L : TOP_LEVELS, s_sum, s_min);
}

// ---------------------------------------------------------------------------------------
// Range reductions with the reference's association (memory.py:39-52).
// ---------------------------------------------------------------------------------------

// Warp-cooperative  sum over leaves [0, need)  =  t1 + (t2 + (... + tk)) : the node walk
// depends only on `need`, so lane s replays s integer steps, all loads go out together,
// and the chain is folded innermost-first through shuffles.  Every lane returns the total.
__device__ __forceinline__ double warp_prefix_total(const double *v, int64_t cap, int64_t need) {
    const int lane = threadIdx.x & 31;
    int64_t node = 1, width = cap, rem = need, tnode = 0;
    bool emit = false, finished = false;
    for (int s = 0; s <= lane; ++s) {
        emit = false;
        if (finished) continue;
        if (rem == width) {
            emit = true;
            tnode = node;
            finished = true;
        } else {
            int64_t half = width >> 1;
            if (rem <= half) {
                node = 2 * node;
            } else {
                emit = true;
                tnode = 2 * node;
                node = 2 * node + 1;
                rem -= half;
            }
            width = half;
        }
    }
    double term = emit ? v[tnode] : 0.0;
    unsigned mask = __ballot_sync(0xffffffffu, emit);
    double acc = 0.0;
    bool first = true;
    while (mask) {
        int s = 31 - __clz(mask);
        mask &= ~(1u << s);
        double t = shfl_f64(term, s);
        acc = first ? t : __dadd_rn(t, acc);
        first = false;
    }
    return acc;
}

// General [lo, hi] (inclusive) by one thread.  At the node where the range splits the
// result is op(suffix part of the left child, prefix part of the right child); each part
// is a chain folded from its deepest exactly-covered node upwards.
template <bool IS_SUM>
__device__ double range_fold(const double *v, int64_t cap, int64_t lo, int64_t hi) {
    auto op = [](double x, double y) { return IS_SUM ? __dadd_rn(x, y) : fmin(x, y); };
    int64_t node = 1, nlo = 0, width = cap;
    while (true) {  // walk down while the range sits inside one child
        if (lo == nlo && hi == nlo + width - 1) return v[node];
        int64_t half = width >> 1, mid = nlo + half;
        if (hi < mid) { node = 2 * node; width = half; }
        else if (lo >= mid) { node = 2 * node + 1; nlo = mid; width = half; }
        else break;
    }
    const int64_t half = width >> 1, mid = nlo + half;
    // suffix of the left child: leaves [lo, mid)
    double terms[64];
    int nt = 0;
    {
        int64_t n = 2 * node, s_lo = nlo, w = half;
        while (!(lo == s_lo)) {
            int64_t h = w >> 1;
            if (lo >= s_lo + h) { n = 2 * n + 1; s_lo += h; }
            else { terms[nt++] = v[2 * n + 1]; n = 2 * n; }
            w = h;
        }
        double acc = v[n];
        for (int i = nt - 1; i >= 0; --i) acc = op(acc, terms[i]);
        terms[0] = acc;
    }
    const double left_part = terms[0];
    // prefix of the right child: leaves [mid, hi]
    nt = 0;
    double right_part;
    {
        int64_t n = 2 * node + 1, p_lo = mid, w = half;
        while (!(hi == p_lo + w - 1)) {
            int64_t h = w >> 1;
            if (hi < p_lo + h) { n = 2 * n; }
            else { terms[nt++] = v[2 * n]; n = 2 * n + 1; p_lo += h; }
            w = h;
        }
        double acc = v[n];
        for (int i = nt - 1; i >= 0; --i) acc = op(terms[i], acc);
        right_part = acc;
    }
    return op(left_part, right_part);
}

__global__ void k_range_reduce(const double *v, int64_t cap, int64_t lo, int64_t hi, int is_sum, double *out) {
    if (threadIdx.x == 0 && blockIdx.x == 0)
        out[0] = is_sum ? range_fold<true>(v, cap, lo, hi) : range_fold<false>(v, cap, lo, hi);
}

// ---------------------------------------------------------------------------------------
// K2 + K3: stratified prefix-sum descent, one warp per query, and importance weights.
//
// Descent (memory.py:122-129) reads only LEFT children.  Below node i, the left children
// of the next 5 generations number 1+2+4+8+16 = 31, so lanes 0..30 fetch them in one
// round trip (generation j, pair q -> node i*2^j + 2q, lane 2^(j-1)-1+q) and the five
// compare / subtract steps then run out of registers through shuffles: ceil(L/5)
// dependent memory round trips per query instead of L.  The subtraction sequence is the
// reference's, so the index is bit-exact.
// ---------------------------------------------------------------------------------------
struct SampleArgs {
    const double *sum;
    const double *mnn;     // null: no weights (stand-alone descent)
    int64_t cap;
    int L;
    const double *u;       // uniforms (per_sample_indices) or masses (seg_find_prefixsum)
    int direct_mass;       // 1: `u` holds masses already
    int64_t batch;
    int64_t n_rows;        // len(_storage)
    double beta;
    int64_t *out_idx;
    double *out_w64;
    float *out_w32;
    u64 *status;
};

__global__ void __launch_bounds__(128) k_sample(SampleArgs a) {
    const int lane = threadIdx.x & 31;
    const int64_t q = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
    if (q >= a.batch) return;  // whole warp leaves together

    double mass;
    const double root_sum = a.sum[1];
    if (a.direct_mass) {
        mass = a.u[q];
        if (lane == 0 && !(mass >= 0.0 && mass <= __dadd_rn(root_sum, 1e-5))) atomicOr(a.status, (u64)PER_STATUS_BAD_PREFIXSUM);
    } else {
        const double total = warp_prefix_total(a.sum, a.cap, a.n_rows - 1);  // memory.py:244
        const double seg = __ddiv_rn(total, (double)a.batch);                // memory.py:245
        mass = __dadd_rn(__dmul_rn(a.u[q], seg), __dmul_rn((double)q, seg)); // memory.py:247
    }

    // lane -> (generation j in 1..5, pair q_in_gen)
    const int j = 32 - __clz(lane + 1);           // lane 0 -> 1, lanes 1-2 -> 2, 3-6 -> 3, 7-14 -> 4, 15-30 -> 5
    const int qg = lane + 1 - (1 << (j - 1));
    int64_t node = 1;
    int depth = 0;
    while (depth < a.L) {
        const int s = a.L - depth < 5 ? a.L - depth : 5;
        double left = 0.0;
        if (j <= s && lane < 31) left = a.sum[(node << j) + 2 * qg];
        int m = 0;  // path inside this 5-generation block
#pragma unroll
        for (int g = 1; g <= 5; ++g) {
            if (g <= s) {
                double lv = shfl_f64(left, (1 << (g - 1)) - 1 + m);
                if (lv > mass) {
                    m = 2 * m;
                } else {
                    mass = __dsub_rn(mass, lv);
                    m = 2 * m + 1;
                }
            }
        }
        node = (node << s) + m;
        depth += s;
    }
    const int64_t k = node - a.cap;
    if (lane == 0) {
        a.out_idx[q] = k;
        if (a.mnn && (a.out_w64 || a.out_w32)) {
            // memory.py:289-295
            const double n = (double)a.n_rows;
            const double p_min = __ddiv_rn(a.mnn[1], root_sum);
            const double max_w = pow(__dmul_rn(p_min, n), -a.beta);
            const double p = __ddiv_rn(a.sum[a.cap + k], root_sum);
            const double w = __ddiv_rn(pow(__dmul_rn(p, n), -a.beta), max_w);
            if (a.out_w64) a.out_w64[q] = w;
            if (a.out_w32) a.out_w32[q] = (float)w;
        }
    }
}

}  // namespace perb
