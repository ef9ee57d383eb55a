// update_kernels.cuh — K1 (+K5 prologue): batched leaf update with the reference's
// sequential semantics and re-reduction of every touched ancestor (memory.py:76-87, 300-320).
//
// What bounds it (profiles/r01_notes.md): a batch of 512 leaves dirties ~500 distinct 128-B
// lines per generation in each tree.  One SM retires about one 32-B sector per cycle for
// scattered accesses, so the first, single-CTA version (every thread re-reading its own 16
// nodes with 8 LDG.128 and writing 15 STG.64 per tree and round) needed ~95 k sector
// operations on one SM = 55 us.  This version
//   * spreads the batch over a thread-block CLUSTER of up to 8 CTAs (8 SMs) that meet at
//     hardware cluster barriers between phases, and
//   * lets 16 adjacent lanes share one item: each lane loads 16 B of the two 128-B lines below the item
//     (one wavefront per 4 items instead of 8 per item) and the 4 generations are reduced
//     with shuffles, left operand always the lower-numbered child.
//
// The update of exactly the sample just drawn (ascending indices, <= 128 items per CTA) takes the SORTED FAST
// PATH further down instead of the level-synchronous rounds: per-CTA register climb over prefetched siblings,
// one cluster barrier, border chains in CTA 0 (see the comment in front of SortSmem).
#pragma once
#include <cooperative_groups.h>

#include "common.cuh"
#include "td_kernels.cuh"

namespace perb {
namespace cg = cooperative_groups;

constexpr int TOP_LEVELS = 10;            // the dense top rebuild covers at most this many generations
constexpr int TOP_NODES = 1 << TOP_LEVELS;
constexpr int LIFT = 5;                    // generations recomputed per sparse round (32 nodes = two 128-B lines)
constexpr int LIFT_LANES = 16;             // lanes that share one item in a sparse round: 2^(LIFT-1), 16 B each
constexpr int UPD_CLUSTER_MAX_ITEMS = 4096;  // largest batch the one-launch (cluster) form takes
constexpr int UPD_SLICE_MAX = UPD_CLUSTER_MAX_ITEMS / 8;  // items per CTA at most (8 CTAs above 448 items)
constexpr int UPD_STAGE_FLOATS = 12288;      // DYNAMIC shared memory: staging for Q rows (48 KB; its first 32 KB double as the top-of-tree
                                             // buffers, the stage-2 border staging sits behind them)
constexpr int UPD_SMEM_BYTES = UPD_STAGE_FLOATS * (int)sizeof(float);
constexpr int UPD_THREADS = 1024;           // 64 items x 16 lanes per sparse pass
constexpr int UPD_MAX_CTAS = 16;           // largest cluster the kernel is written for (non-portable size, opted into by the host;
                                           // used for sorted batches of 1 025 .. 2 048 items and on request, see run_update)

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
    u64 *stamps;              // cap entries, all zero between calls; null when indices are known distinct.  Arbitration of
                              // duplicates: atomicMax(stamps[k], j + 1); the item whose stamp survives wins and puts the
                              // entry back to zero with its leaf write, so nothing outlives the call (a launch replayed from
                              // a CUDA graph, or another tree sharing the array, starts from zeros again)
    u64 *status;
    double *max_prio;         // device scalar `_max_priority`; may be null (stand-alone trees)
    double batch_max;         // LEAF_GIVEN: raw-priority maximum of the batch (<=0: none)
    int64_t idx_base;         // sharded replay: global index of this shard's leaf 0
    int skip_foreign;         // sharded replay: idx == -1 marks another shard's item, skipped silently
    double *rec_out;          // sharded replay: refresh this shard's 4-double record after the update (may be null)
    int64_t rec_need;         //   leaves of this slice inside the global prefix [0, len-1)
    int64_t rec_len;          //   local len
    int cluster_max;          // host side: plain updates above this many items use the multi-launch form (<= 0: the built-in limit)
    int max_ctas;             // host side: cap on the cluster size of the one-launch form (<= 0: none)
    int sorted_fast;          // sorted_idx: take the per-CTA climb (sorted fast path) instead of the level-synchronous rounds
    int sorted_idx;           // the indices are non-decreasing (they are the stratified sample just drawn): equal
                              // indices are adjacent, so "a later duplicate wins" needs one look at idx[j+1], no atomics
    u64 *trace;               // diagnostics: CTA 0 / thread 0 stores globaltimer ns at phase boundaries (null: off)
    double *const *publish_to; // sharded replay, early publish: the mailboxes of all ranks (null: the next sample publishes)
    int n_shards, shard_rank;  //   this launch's tail stores the refreshed record, tagged with the NEXT exchange epoch
    const u64 *d_epoch;        //   (completed exchange samples + 1), straight into them
    u64 *tl;                  // timeline (diagnostics), see tl_stamp: this launch's first entry
    const u64 *range;         // sharded replay: {epoch, -, lo tag, hi tag} left by the last per_sample_sharded; when set, only
                              // the strata [lo, hi) that this shard owns are visited (null: all n items)
    unsigned *dense_counter;  // per-buffer word for k_upd_dense_sorted's "last CTA" election (null: that form is not offered)
    int dense_lmax;           // deepest tree (log2 leaves) whose bound updates always take the dense form (0: form off)
    int late_release;         // k_update_cluster as the TD launch in front of k_upd_dense_sorted: dependents are released behind the wait
    int early_search;         // k_upd_dense_sorted behind such a launch: the item runs are searched before the wait
};

// Owned strata of the last sharded sample: tags are (epoch << 32) | value, written with atomicMax
// by the sampling kernel (lo is stored as 0xffffffff - lo), valid only if their epoch is current.
__device__ __forceinline__ void owned_range(const u64 *range, int64_t n, int &lo, int &hi) {
    lo = 0;
    hi = (int)n;
    if (!range) return;
    const u64 epoch = __ldcg(range), tl = __ldcg(range + 2), th = __ldcg(range + 3);
    if ((tl >> 32) != (epoch & 0xffffffffu) || (th >> 32) != (epoch & 0xffffffffu)) {
        lo = hi = 0;  // this shard owned no stratum of that sample
        return;
    }
    lo = (int)(0xffffffffu - (unsigned)(tl & 0xffffffffu));
    hi = (int)(th & 0xffffffffu);
}

constexpr int64_t IDX_SKIP = INT64_MIN;

// Depth at which the dense top rebuild starts: the largest value <= TOP_LEVELS that leaves
// a multiple of LIFT generations below it.
__host__ __device__ __forceinline__ int top_depth(int L) {
    if (L <= TOP_LEVELS) return L;
    const int r = (L - TOP_LEVELS) % LIFT;
    return r == 0 ? TOP_LEVELS : TOP_LEVELS - (LIFT - r);  // >= TOP_LEVELS - LIFT + 1
}

__device__ __forceinline__ int64_t upd_index(const UpdArgs &a, int64_t j) {
    if (a.idx) {
        const int64_t g = a.idx[j];
        return (a.skip_foreign && g == -1) ? IDX_SKIP : g - a.idx_base;
    }
    int64_t k = a.ring_start + j;
    return k >= a.ring_size ? k - a.ring_size : k;
}

// Would the reference accept item j (memory.py:315-316)?  No status side effects, no pow: what the stamp pass of the
// multi-launch form needs, so that a rejected item never holds a stamp.
__device__ __forceinline__ bool upd_item_ok(const UpdArgs &a, int64_t j, int64_t &k) {
    k = upd_index(a, j);
    if (k == IDX_SKIP || k < 0 || k >= a.len_limit) return false;
    if (a.src_kind == LEAF_PRIO_F64) return static_cast<const double *>(a.src)[j] > 0.0;
    if (a.src_kind == LEAF_PRIO_F32) return static_cast<const float *>(a.src)[j] > 0.0f;
    return true;
}

// Raw priority (for the running max) and leaf value of item j.  Returns false for an
// item the reference would have rejected (memory.py:315-316).  One pow() call site.
__device__ __forceinline__ bool upd_item(const UpdArgs &a, int64_t j, int64_t &k, double &leaf, double &raw) {
    k = upd_index(a, j);
    raw = 0.0;
    if (k == IDX_SKIP) return false;
    unsigned bad = (k < 0 || k >= a.len_limit) ? PER_STATUS_BAD_INDEX : 0u;
    double base = 1.0;
    bool powered = true, checked = false;
    if (a.src_kind == LEAF_GIVEN) {
        base = static_cast<const double *>(a.src)[j];
        powered = false;
    } else if (a.src_kind == LEAF_PRIO_F64) {
        base = raw = static_cast<const double *>(a.src)[j];
        checked = true;
    } else if (a.src_kind == LEAF_PRIO_F32) {
        base = raw = (double)static_cast<const float *>(a.src)[j];
        checked = true;
    } else {
        base = ldcg_f64(a.max_prio);
    }
    if (checked && !(raw > 0.0)) bad |= PER_STATUS_BAD_PRIORITY;
    if (bad) {
        atomicOr(a.status, (u64)bad);
        return false;
    }
    leaf = powered ? pow_alpha(base, a.alpha) : base;
    return true;
}

// One sparse round for one item, executed by LIFT_LANES adjacent lanes (sub = lane % LIFT_LANES):
// recompute the LIFT generations of ancestors above the 2^LIFT contiguous nodes below node g.
// Lane `sub` loads 16 B (two nodes); generation 1 is one add per lane, generations 2..LIFT halve
// the owning lanes through xor-shuffles (the owning lane always holds the LEFT child).
// Every lane of the warp must call this (shuffles); `valid` guards the memory traffic.
template <bool IS_SUM>
__device__ __forceinline__ void lift8(double *v, int64_t g, int sub, bool valid) {
    double x = 0.0, y = 0.0;
    if (valid) {
        double2 c = ldcg_f64x2(v + (g << LIFT) + 2 * sub);
        x = c.x;
        y = c.y;
    }
    double r = IS_SUM ? __dadd_rn(x, y) : fmin(x, y);  // generation 1: node (g << (LIFT-1)) + sub
    if (valid) v[(g << (LIFT - 1)) + sub] = r;
#pragma unroll
    for (int s = 1; s < LIFT; ++s) {  // lanes with the low s bits clear own a node of generation s+1
        const double other = __shfl_xor_sync(0xffffffffu, r, 1 << (s - 1));
        r = IS_SUM ? __dadd_rn(r, other) : fmin(r, other);
        if (valid && (sub & ((1 << s) - 1)) == 0) v[(g << (LIFT - 1 - s)) + (sub >> s)] = r;
    }
}

__device__ __forceinline__ void lift8_both(const TreeRef &t, int64_t g, int sub, bool valid) {
    if (t.sum) lift8<true>(t.sum, g, sub, valid);
    if (t.mnn) lift8<false>(t.mnn, g, sub, valid);
}

// Dense rebuild, by one CTA, of the k generations above the 2^k contiguous nodes that
// descend from node `a` (k <= TOP_LEVELS).  smem: 2 * 2^k doubles per tree, addressed as a
// relative heap (generation j, position t at s[(1<<j) + t]).
__device__ __forceinline__ void dense_load(const TreeRef &t, int64_t a, int k, double *s_sum, double *s_min) {
    const int n = 1 << k;
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        if (t.sum) s_sum[n + i] = ldcg_f64(t.sum + (a << k) + i);
        if (t.mnn) s_min[n + i] = ldcg_f64(t.mnn + (a << k) + i);
    }
}
// The 10-generation rebuild of a 1 024-node row by a 1 024-thread CTA, shaped for what bounds it: instruction issue
// (all 32 warps walking the generations in lock step issue ~5 000 warp instructions for 2 046 useful operations and meet
// at ten barriers) and shared-memory banks (a thread that walks 16 consecutive nodes alone makes every access a 32-way
// conflict).  Here 8 warps take the sum tree and 8 the min tree; a thread loads 4 consecutive row nodes (two 16-byte
// loads, conflict-free), folds two generations in registers, and the warp folds five more through xor-shuffles (the
// lane whose bit is clear holds the LEFT child; both lanes form left (op) right): seven generations without a barrier.
// The 8 nodes that are left per tree go to the root in one warp each.  Every node is op(left, right) of its children in
// that order (memory.py:82-87): same bits as the level-by-level form below.
template <bool IS_SUM>
__device__ __forceinline__ void dense_fast10_tree(double *s, const int w /* warp of this tree, 0..7 */, const int lane) {
    auto op = [](double x, double y) { return IS_SUM ? __dadd_rn(x, y) : fmin(x, y); };
    const int i = w * 32 + lane;  // 4 row nodes per thread: 1024 + 4 i ..
    const double2 c0 = *reinterpret_cast<const double2 *>(s + 1024 + 4 * i);
    const double2 c1 = *reinterpret_cast<const double2 *>(s + 1024 + 4 * i + 2);
    const double g1a = op(c0.x, c0.y), g1b = op(c1.x, c1.y);
    *reinterpret_cast<double2 *>(s + 512 + 2 * i) = make_double2(g1a, g1b);
    double v = op(g1a, g1b);
    s[256 + i] = v;
#pragma unroll
    for (int g = 1; g <= 5; ++g) {  // generations 3..7: row 256 -> 8
        const double o = __shfl_xor_sync(0xffffffffu, v, 1 << (g - 1));
        const bool left = ((lane >> (g - 1)) & 1) == 0;
        v = left ? op(v, o) : op(o, v);
        if ((lane & ((1 << g) - 1)) == 0) s[(256 >> g) + (i >> g)] = v;
    }
}
template <bool IS_SUM>
__device__ __forceinline__ void dense_fast10_top(double *s, const int lane) {  // row of 8 at s[8..15] -> root
    auto op = [](double x, double y) { return IS_SUM ? __dadd_rn(x, y) : fmin(x, y); };
    double v = lane < 8 ? s[8 + lane] : 0.0;
#pragma unroll
    for (int g = 1; g <= 3; ++g) {
        const double o = __shfl_xor_sync(0xffffffffu, v, 1 << (g - 1));
        const bool left = ((lane >> (g - 1)) & 1) == 0;
        v = left ? op(v, o) : op(o, v);
        if (lane < 8 && (lane & ((1 << g) - 1)) == 0) s[(8 >> g) + (lane >> g)] = v;
    }
}
__device__ __forceinline__ void dense_build_fast10(const TreeRef &t, double *s_sum, double *s_min) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (warp < 8) {
        if (t.sum) dense_fast10_tree<true>(s_sum, warp, lane);
    } else if (warp < 16) {
        if (t.mnn) dense_fast10_tree<false>(s_min, warp - 8, lane);
    }
    __syncthreads();
    if (warp == 0) {
        if (t.sum) dense_fast10_top<true>(s_sum, lane);
    } else if (warp == 8) {
        if (t.mnn) dense_fast10_top<false>(s_min, lane);
    }
    __syncthreads();
}

// Expects the bottom row in s_sum / s_min [n, 2n) and a CTA barrier behind the stores that filled it; ends with a barrier.
__device__ __forceinline__ void dense_build(const TreeRef &t, int64_t a, int k, double *s_sum, double *s_min) {
    if (k == 10 && blockDim.x == 1024) {
        dense_build_fast10(t, s_sum, s_min);
        return;
    }
    // all generations in shared memory first (CTA barriers down to 32 nodes, then one warp) ...
#pragma unroll 1
    for (int j = k - 1; j >= 5; --j) {
        const int m = 1 << j;
        for (int i = threadIdx.x; i < m; i += blockDim.x) {
            if (t.sum) {
                double2 c = *reinterpret_cast<double2 *>(s_sum + 2 * (m + i));
                s_sum[m + i] = __dadd_rn(c.x, c.y);
            }
            if (t.mnn) {
                double2 c = *reinterpret_cast<double2 *>(s_min + 2 * (m + i));
                s_min[m + i] = fmin(c.x, c.y);
            }
        }
        __syncthreads();
    }
    if (threadIdx.x < 32) {
#pragma unroll 1
        for (int j = (k - 1 < 4 ? k - 1 : 4); j >= 0; --j) {
            const int m = 1 << j;
            const int i = threadIdx.x;
            if (i < m) {
                if (t.sum) s_sum[m + i] = __dadd_rn(s_sum[2 * (m + i)], s_sum[2 * (m + i) + 1]);
                if (t.mnn) s_min[m + i] = fmin(s_min[2 * (m + i)], s_min[2 * (m + i) + 1]);
            }
            __syncwarp();
        }
    }
    __syncthreads();
}
// ... then one coalesced pass writes them out: relative heap slot (1<<j)+i  <->  node (a<<j)+i
__device__ __forceinline__ void dense_write(const TreeRef &t, int64_t a, int k, const double *s_sum, const double *s_min) {
    const int n = 1 << k;
    for (int p = 1 + threadIdx.x; p < n; p += blockDim.x) {
        const int j = 31 - __clz(p);
        const int64_t node = (a << j) + (p - (1 << j));
        if (t.sum) t.sum[node] = s_sum[p];
        if (t.mnn) t.mnn[node] = s_min[p];
    }
}
__device__ __forceinline__ void dense_build_write(const TreeRef &t, int64_t a, int k, double *s_sum, double *s_min) {
    dense_build(t, a, k, s_sum, s_min);
    dense_write(t, a, k, s_sum, s_min);
}
__device__ __forceinline__ void dense_lift_block(const TreeRef &t, int64_t a, int k, double *s_sum, double *s_min) {
    dense_load(t, a, k, s_sum, s_min);
    __syncthreads();
    dense_build_write(t, a, k, s_sum, s_min);
}

// ---------------------------------------------------------------------------------------
// Sorted fast path of the one-launch form.  When the indices are the stratified sample just drawn
// they ascend, each CTA of the cluster owns a contiguous run of leaves, and an ancestor can only be
// shared between CTAs along the <= 7 borders between their runs.  So
//   stage 1  every CTA lifts its own items through all generations below the dense top WITHOUT
//            meeting the others: one thread per surviving item climbs level by level; the sibling at
//            each level is either clean (prefetched from global memory for all levels in ONE round
//            trip) or the neighbouring run of the same CTA (its value comes through shared memory,
//            a named barrier per level).  Nodes that have a leaf of another CTA below them are left out;
//   one cluster barrier;
//   stage 2  one warp of CTA 0 recomputes the left-out nodes, a chain per border from the level where
//            the two border leaves meet, children prefetched in one round trip beside the loads of
//            the dense top, whose bottom row is then patched in shared memory.
// Three level-synchronous phases with five cluster barriers become one barrier and two round trips.
// Every node is still op(left child, right child) of final children: same bits (memory.py:82-87).
// ---------------------------------------------------------------------------------------
constexpr int SORT_SLICE_MAX = 128;  // items per CTA
constexpr int SORT_CLIMB_MAX = 16;   // generations between the leaves and the dense top
struct SortSmem {
    double sib[SORT_CLIMB_MAX][SORT_SLICE_MAX][2];  // sibling (sum, min) of the level-l node above slice item t (by slice position)
    double v[SORT_SLICE_MAX][2];                    // published value of the node holding survivor i (read where two runs meet)
    double leaf[SORT_SLICE_MAX];
    int k[SORT_SLICE_MAX];                          // surviving leaves of this CTA, ascending
    short fl[SORT_SLICE_MAX][2];                    // published first / last survivor of the run below that node
    short orig[SORT_SLICE_MAX];                     // slice position of survivor i
    int warp_cnt[4];
};
enum { SRT_KEY_MIN = 0, SRT_KEY_MAX, SRT_NB_LEFT, SRT_NB_RIGHT, SRT_MERGE_MASK, SRT_CHAINS, SRT_WORDS = 8 };
static_assert(sizeof(SortSmem) <= UPD_STAGE_FLOATS * sizeof(float), "SortSmem must fit the staging buffer");
constexpr int S2_OFF_DOUBLES = 4 * TOP_NODES;  // stage-2 child staging sits behind the two dense-top buffers
static_assert((S2_OFF_DOUBLES + UPD_MAX_CTAS * SORT_CLIMB_MAX * 6) * sizeof(double) <= UPD_STAGE_FLOATS * sizeof(float), "stage-2 staging");

// level (generations above the leaves) of the lowest common ancestor of two leaves; equal leaves count as 1
__device__ __forceinline__ int lca_level(int64_t x, int64_t y) { return x == y ? 1 : 64 - __clzll((long long)(x ^ y)); }

// ---------------------------------------------------------------------------------------
// One-launch form (n <= UPD_CLUSTER_MAX_ITEMS): a cluster of 1..8 CTAs.
//   phase 0  (optional, fused_td) K5: TD error -> priority for the batch (utils.py:71-80);
//            each CTA evaluates its own slice of rows (Q rows staged through shared memory
//            with coalesced loads), the batch-wide max and the loss terms meet in CTA 0's
//            shared memory (DSMEM) at one cluster barrier
//   phase 1  last-write-wins arbitration: atomicMax(stamp[idx], epoch|j)    (memory.py:314-318)
//   phase 2  winners write the leaves of both trees; max -> _max_priority   (memory.py:320)
//   phase 3  sparse rounds (lift8) down to depth top_depth(L), cluster barrier between rounds
//   phase 4  CTA 0: dense rebuild of depths top_depth(L)..0 in shared memory
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(UPD_THREADS, 1) k_update_cluster(UpdArgs a, TdArgs td, int fused_td, int skip_update) {
    extern __shared__ __align__(16) float s_raw[];           // UPD_STAGE_FLOATS floats. phase 0: staged Q rows; phases 1-2: leaf values; phase 4: tree tops
    __shared__ int s_k[UPD_SLICE_MAX];                       // leaf index of each surviving item of the slice, else -1
    __shared__ float s_tdv[UPD_SLICE_MAX];                   // |td| of the slice
    double *s_tree = reinterpret_cast<double *>(s_raw);
    __shared__ double s_red[34];
    __shared__ float s_clu[2 * UPD_MAX_CTAS];               // CTA 0's copy collects per-CTA max / loss partials
    __shared__ int s_srt[SRT_WORDS];                        // sorted path: slice key range, neighbouring leaves, levels at which runs meet
    __shared__ long long s_bound[4 * UPD_MAX_CTAS];         // sorted path, CTA 0's copy: {smallest, largest leaf} per CTA; then {border leaf, level} per border
    double *s_sum = s_tree, *s_min = s_tree + 2 * TOP_NODES;

    cg::cluster_group cl = cg::this_cluster();
    const int ncta = (int)cl.num_blocks();
    const int rank = (int)cl.block_rank();
    const int T = blockDim.x, tid = threadIdx.x;
    const int nwarp = (T + 31) >> 5;
    if (tid == 0) {
        s_srt[SRT_KEY_MIN] = INT_MAX;
        s_srt[SRT_KEY_MAX] = -1;
        s_srt[SRT_NB_LEFT] = s_srt[SRT_NB_RIGHT] = -1;
        s_srt[SRT_MERGE_MASK] = 0;
    }
#define UPD_TRACE(slot)                                                      \
    do {                                                                     \
        if (a.trace && rank == 0 && tid == 0) {                              \
            u64 t__;                                                         \
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t__));          \
            a.trace[slot] = t__;                                             \
        }                                                                    \
    } while (0)
    // With early publish: wait first, release second.  The sampling kernel behind this one then starts its record
    // exchange before its own wait (tree_kernels.cuh) and relies on every kernel up to the previous sample having
    // retired by then.  Otherwise the dependents are released at once.
    const bool tl_on = a.tl && rank == 0 && tid == 0;
    if (tl_on) tl_stamp(a.tl, 0, 7);
    const bool release_late = a.publish_to || a.late_release;
    if (!release_late) pdl_launch_dependents();  // (early publish, TD in front of the dense form: only behind the wait)
    pdl_wait();  // indices / Q tables / the tree itself come from the kernels in front of this one
    if (release_late) pdl_launch_dependents();
    if (tl_on) tl_stamp(a.tl, 1, 8);
    UPD_TRACE(0);
    int first, last;  // items this launch visits: all of them, or the strata this shard owns
    owned_range(a.range, a.n, first, last);
    const int n = last - first;
    const int per = (n + ncta - 1) / ncta;  // slice of this CTA: items [lo, hi)
    const int lo = first + (rank * per < n ? rank * per : n);
    const int hi = lo + per < last ? lo + per : last;
    const int cnt = hi - lo;
    const int climb = a.t.L > TOP_LEVELS ? a.t.L - TOP_LEVELS : 0;
    // uniform over the cluster: `per`, the tree and the flags are the same in every CTA
    const bool sorted_path = a.sorted_idx && a.sorted_fast && per <= SORT_SLICE_MAX && climb <= SORT_CLIMB_MAX && a.t.sum && a.t.mnn;
    const int top = sorted_path ? (a.t.L < TOP_LEVELS ? a.t.L : TOP_LEVELS) : top_depth(a.t.L);
    // sorted path: the index of slice item tid % 128 is requested now, so that the sibling loads behind the TD phase
    // can go out at once
    const int64_t early_ki = (sorted_path && (tid & (SORT_SLICE_MAX - 1)) < cnt) ? upd_index(a, lo + (tid & (SORT_SLICE_MAX - 1))) : IDX_SKIP;

    if (fused_td) {  // rows of the WHOLE batch (the max is batch-wide), partitioned over the CTAs
        float *f_red = reinterpret_cast<float *>(s_red);
        float *s_q = s_raw;  // staged q_next rows of the slice
        float *s_td = s_tdv;
        const int nA = (int)td.nA;
        const int nb = (int)td.batch;
        const int tper = (nb + ncta - 1) / ncta;
        const int lo = rank * tper < nb ? rank * tper : nb;  // shadows the item slice inside this block
        const int hi = lo + tper < nb ? lo + tper : nb;
        const int cnt = hi - lo;
        const bool staged = (int64_t)cnt * nA <= UPD_STAGE_FLOATS;
        // cnt <= UPD_SLICE_MAX == blockDim: one row per thread.  The scalar columns and Q(s, a) are
        // requested first so their round trips overlap the staging of the Q(s', .) rows.
        const bool have = tid < cnt;
        const int64_t b = lo + tid;
        int64_t act_b = 0;
        float rew_b = 0.0f, done_b = 0.0f, w_b = 0.0f, q_a = 0.0f;
        if (have) {
            act_b = td_action(td, b);
            rew_b = td.rew[b];
            done_b = td.done[b];
            if (td.out_loss) w_b = td.w[b];
        }
        // the target network's rows as well when both fit: Q_target(s', argmax) then needs no dependent global load
        const bool staged_t = staged && (int64_t)2 * cnt * nA <= UPD_STAGE_FLOATS;
        if (staged) {
            // The slice's rows are one contiguous block of cnt * nA floats per table: 16-byte loads, up to three per
            // thread and table requested before the first one is stored (a slice of 512 rows, the 8-GPU case, in one
            // round trip instead of one per 1 024 floats).
            const float *src = td.q_next + (int64_t)lo * nA;
            const float *src_t = td.q_tgt + (int64_t)lo * nA;
            const int nf = cnt * nA;
            if ((nf & 3) == 0 && ((((uintptr_t)src) | ((uintptr_t)src_t)) & 15) == 0) {
                const int nv = nf >> 2;
                const float4 *s4 = reinterpret_cast<const float4 *>(src), *t4 = reinterpret_cast<const float4 *>(src_t);
                float4 *d4 = reinterpret_cast<float4 *>(s_q), *dt4 = reinterpret_cast<float4 *>(s_q + nf);
#pragma unroll 1
                for (int base = tid; base < nv; base += 3 * T) {
                    float4 x[3], y[3];
#pragma unroll
                    for (int u = 0; u < 3; ++u) {
                        const int e = base + u * T;
                        if (e < nv) {
                            x[u] = s4[e];
                            if (staged_t) y[u] = t4[e];
                        }
                    }
#pragma unroll
                    for (int u = 0; u < 3; ++u) {
                        const int e = base + u * T;
                        if (e < nv) {
                            d4[e] = x[u];
                            if (staged_t) dt4[e] = y[u];
                        }
                    }
                }
            } else {
#pragma unroll 1
                for (int e = tid; e < nf; e += T) {
                    s_q[e] = src[e];
                    if (staged_t) s_q[nf + e] = src_t[e];
                }
            }
        }
        if (have) q_a = td.q[b * nA + act_b];
        __syncthreads();
        UPD_TRACE(6);
        float mx = 0.0f, acc = 0.0f;
        if (have) {
            const float *qn = staged ? s_q + tid * nA : td.q_next + b * nA;
            int best = 0;
            float bv = qn[0];
#pragma unroll 1
            for (int i = 1; i < nA; ++i) {  // first maximum wins, like argmax (utils.py:72)
                const float x = qn[i];
                if (x > bv) { bv = x; best = i; }
            }
            const float boot = staged_t ? s_q[cnt * nA + tid * nA + best] : td.q_tgt[b * nA + best];
            const float y = __fadd_rn(rew_b, __fmul_rn(__fmul_rn(td.gamma_n, boot), __fsub_rn(1.0f, done_b)));
            const float x = fabsf(__fsub_rn(y, q_a));
            s_td[tid] = x;
            mx = x;
            if (td.out_loss) acc = __fmul_rn(td_huber(x), w_b);
        }
        mx = warp_max_f32(mx);
        acc = warp_sum_f32(acc);
        if ((tid & 31) == 0) {
            f_red[tid >> 5] = mx;
            f_red[32 + (tid >> 5)] = acc;
        }
        __syncthreads();
        UPD_TRACE(7);
        if (tid < 32) {
            float m = tid < nwarp ? f_red[tid] : 0.0f;
            float l = tid < nwarp ? f_red[32 + tid] : 0.0f;
            m = warp_max_f32(m);
            l = warp_sum_f32(l);
            if (tid == 0) {
                float *home = cl.map_shared_rank(s_clu, 0);
                home[rank] = m;
                home[UPD_MAX_CTAS + rank] = l;
            }
        }
        cl.sync();
        UPD_TRACE(8);
        if (tid < 32) {
            // one warp per CTA fetches the per-CTA partials from CTA 0 (DSMEM moves ~17 B/cycle: 4096 threads
            // reading them all would cost microseconds) and leaves the batch max in local shared memory
            const float *home = cl.map_shared_rank(s_clu, 0);
            float m = tid < ncta ? home[tid] : 0.0f;
            float l = tid < ncta ? home[UPD_MAX_CTAS + tid] : 0.0f;
            m = warp_max_f32(m);  // keeps a NaN (see max_keep_nan)
            if (tid == 0) f_red[32] = m;
            if (td.out_loss && rank == 0) {
                float acc_l = 0.0f;
                for (int r = 0; r < ncta; ++r) acc_l += __shfl_sync(0xffffffffu, l, r);  // fixed order: deterministic
                if (tid == 0) td.out_loss[0] = acc_l / (float)td.batch;               // .mean(), utils.py:80
            }
        }
        __syncthreads();
        UPD_TRACE(9);
        mx = f_red[32];
#pragma unroll 1
        for (int i = tid; i < cnt; i += T) {
            const float x = s_td[i];
            td.out_prio[lo + i] = td_prio(x, mx);
            if (td.out_td) td.out_td[lo + i] = x;
        }
        // out_prio must be complete before the items below read it: they are this CTA's own rows unless the
        // owned-range mode re-partitions the items (then another CTA may have produced them)
        if (a.range || skip_update) cl.sync();
        else __syncthreads();
        if (skip_update) return;
    }
    if (tl_on) tl_stamp(a.tl, 2, 12);  // TD arithmetic done

    UPD_TRACE(1);
    double my_max = 0.0;
    if (sorted_path) {
        SortSmem &S = *reinterpret_cast<SortSmem *>(s_raw);
        const int lane = tid & 31, warp = tid >> 5;
        if (!fused_td) __syncthreads();  // s_srt was initialised at kernel entry; the TD phase ends with a barrier of its own
        // ---- siblings first: they depend on nothing but the indices, so their round trip runs beside the pow() below.
        // Thread t takes slice item t % 128 at levels t / 128 + 1 and + 9 (everything is loaded; shared or joined
        // levels are simply not used).
        {
            const int it = tid & (SORT_SLICE_MAX - 1), lg = tid / SORT_SLICE_MAX;
            if (it < cnt) {
                const int64_t ki = early_ki;
                if (ki != IDX_SKIP && ki >= 0 && ki < a.len_limit) {
#pragma unroll
                    for (int r = 0; r < 2; ++r) {
                        const int l = lg + 1 + r * (UPD_THREADS / SORT_SLICE_MAX);
                        if (l <= climb) {
                            const int64_t sibn = ((a.t.cap + ki) >> (l - 1)) ^ 1;
                            S.sib[l - 1][it][0] = ldcg_f64(a.t.sum + sibn);
                            S.sib[l - 1][it][1] = ldcg_f64(a.t.mnn + sibn);
                        }
                    }
                }
            }
        }
        // ---- leaves: one thread per item of the slice (cnt <= 128) ----
        int kk = -1, kraw = -1;
        double leaf = 0.0;
        if (tid < cnt) {
            const int64_t ki = early_ki;
            if (ki != IDX_SKIP && ki >= 0 && ki < a.len_limit) kraw = (int)ki;
            int64_t k;
            double raw;
            if (upd_item(a, lo + tid, k, leaf, raw)) {
                kk = (int)k;
                my_max = raw;
                if (lo + tid + 1 < (int)a.n && a.idx[lo + tid + 1] == a.idx[lo + tid]) kk = -1;  // the next item overwrites this one
            }
        } else if (tid == SORT_SLICE_MAX) {  // nearest index-valid leaf in front of the slice ...
            for (int j = lo - 1; j >= first; --j) {
                const int64_t ki = upd_index(a, j);
                if (ki != IDX_SKIP && ki >= 0 && ki < a.len_limit) {
                    s_srt[SRT_NB_LEFT] = (int)ki;
                    break;
                }
            }
        } else if (tid == SORT_SLICE_MAX + 32) {  // ... and behind it
            for (int j = hi; j < last; ++j) {
                const int64_t ki = upd_index(a, j);
                if (ki != IDX_SKIP && ki >= 0 && ki < a.len_limit) {
                    s_srt[SRT_NB_RIGHT] = (int)ki;
                    break;
                }
            }
        }
        unsigned alive = 0;
        if (tid < SORT_SLICE_MAX) {
            // keys ascend: the smallest / largest index-valid leaf of the slice, one shared-memory atomic per warp
            const int wmin = __reduce_min_sync(0xffffffffu, kraw >= 0 ? kraw : INT_MAX), wmax = __reduce_max_sync(0xffffffffu, kraw);
            alive = __ballot_sync(0xffffffffu, kk >= 0);
            // The path relies on ascending indices (they are the sampler's own output).  A caller that scribbled on
            // the buffer in between is told so: neighbours inside a warp, and the slice against the leaves around it
            // (checked below), must not descend.
            const int prev = __shfl_up_sync(0xffffffffu, kraw, 1);
            if (lane > 0 && kraw >= 0 && prev > kraw) atomicOr(a.status, (u64)PER_STATUS_BAD_INDEX);
            if (lane == 0) {
                S.warp_cnt[warp] = __popc(alive);
                if (wmax >= 0) {
                    atomicMin(&s_srt[SRT_KEY_MIN], wmin);
                    atomicMax(&s_srt[SRT_KEY_MAX], wmax);
                }
            }
            if (a.max_prio) {
                if (a.src_kind == LEAF_GIVEN) my_max = (tid == 0 && rank == 0) ? a.batch_max : 0.0;
                for (int o = 16; o; o >>= 1) my_max = fmax(my_max, shfl_f64(my_max, lane ^ o));
                if (lane == 0) s_red[warp] = my_max;
            }
        }
        __syncthreads();
        UPD_TRACE(10);
        int m = 0;  // survivors of this CTA
        for (int w = 0; w < SORT_SLICE_MAX / 32; ++w) m += S.warp_cnt[w];
        if (tid < SORT_SLICE_MAX && kk >= 0) {  // compact: surviving leaves, still ascending
            int pos = __popc(alive & ((1u << lane) - 1u));
            for (int w = 0; w < warp; ++w) pos += S.warp_cnt[w];
            S.k[pos] = kk;
            S.leaf[pos] = leaf;
            S.orig[pos] = (short)tid;
            S.v[pos][0] = S.v[pos][1] = leaf;
            S.fl[pos][0] = S.fl[pos][1] = (short)pos;
            a.t.sum[a.t.cap + kk] = leaf;
            a.t.mnn[a.t.cap + kk] = leaf;
        }
        if (tid == 0) {  // CTA 0 learns every CTA's key range for the borders of stage 2
            if (s_srt[SRT_KEY_MAX] >= 0 && ((s_srt[SRT_NB_LEFT] > s_srt[SRT_KEY_MIN]) ||
                                            (s_srt[SRT_NB_RIGHT] >= 0 && s_srt[SRT_NB_RIGHT] < s_srt[SRT_KEY_MAX])))
                atomicOr(a.status, (u64)PER_STATUS_BAD_INDEX);  // not ascending across slices
            long long *home = cl.map_shared_rank(s_bound, 0);
            home[2 * rank] = s_srt[SRT_KEY_MAX] >= 0 ? (long long)s_srt[SRT_KEY_MIN] : -1;
            home[2 * rank + 1] = (long long)s_srt[SRT_KEY_MAX];
            if (a.max_prio) {
                double mx = 0.0;
                for (int w = 0; w < SORT_SLICE_MAX / 32; ++w) mx = fmax(mx, s_red[w]);
                if (mx > 0.0) atomic_max_pos_f64(a.max_prio, mx);
            }
        }
        UPD_TRACE(2);
        __syncthreads();
        // ---- stage 1 ----
        if (tid < SORT_SLICE_MAX) {
            // One thread per survivor keeps, in registers, the value of the node that holds its item and the first /
            // last survivor of the run of items below that node.  Two runs meet at the level of the lowest common
            // ancestor of the adjacent leaves between them (mL / mR): only there does a thread need a neighbour's
            // value, which that neighbour published one level earlier; the four warps meet at a named barrier only
            // at the levels where some pair of runs of this CTA meets (rare below the dense top for a spread-out sample).
            const int64_t nbl = s_srt[SRT_NB_LEFT], nbr = s_srt[SRT_NB_RIGHT];
            const int i = tid;
            const bool have = i < m;
            const int ck = have ? (int)a.t.cap + S.k[i] : 0;  // my leaf's node number (cap <= 2^26 on this path)
            const int org = have ? S.orig[i] : 0;
            int stop = climb + 1;  // first level whose node another CTA shares
            if (have && nbl >= 0) stop = min(stop, lca_level(S.k[i], nbl));
            if (have && nbr >= 0) stop = min(stop, lca_level(S.k[i], nbr));
            double vs = have ? S.leaf[i] : 0.0, vm = vs;
            int F = i, Lr = i;
            int mR = (have && i + 1 < m) ? lca_level(S.k[i], S.k[i + 1]) : 99;  // level where the run behind mine joins
            int mL = (have && i > 0) ? lca_level(S.k[i - 1], S.k[i]) : 99;      // ... and the run in front of mine
            unsigned meet = 0;
            if (mR < stop) meet |= 1u << mR;
            if (mL < stop) meet |= 1u << mL;
            meet = __reduce_or_sync(0xffffffffu, meet);
            if (lane == 0 && meet) atomicOr(reinterpret_cast<unsigned *>(&s_srt[SRT_MERGE_MASK]), meet);
            asm volatile("bar.sync 1, %0;" ::"n"(SORT_SLICE_MAX) : "memory");
            meet = (unsigned)s_srt[SRT_MERGE_MASK];
#pragma unroll 1
            for (int l = 1; l <= climb; ++l) {
                const bool act = have && l < stop;
                const bool sync_here = (meet >> l) & 1u;
                if (sync_here) asm volatile("bar.sync 1, %0;" ::"n"(SORT_SLICE_MAX) : "memory");  // level l-1 values are published
                const int c = ck >> (l - 1);
                if (act) {
                    double2 o = *reinterpret_cast<const double2 *>(S.sib[l - 1][org]);  // clean sibling unless a run joins
                    if (mR == l) {  // my node is the left child, the run behind mine fills the right one
                        const int j = Lr + 1;
                        o = *reinterpret_cast<const double2 *>(S.v[j]);
                        Lr = S.fl[j][1];
                        mR = Lr + 1 < m ? lca_level(S.k[Lr], S.k[Lr + 1]) : 99;
                    } else if (mL == l) {
                        const int j = F - 1;
                        o = *reinterpret_cast<const double2 *>(S.v[j]);
                        F = S.fl[j][0];
                        mL = F > 0 ? lca_level(S.k[F - 1], S.k[F]) : 99;
                    }
                    vs = (c & 1) ? __dadd_rn(o.x, vs) : __dadd_rn(vs, o.x);  // left operand = left child
                    vm = fmin(vm, o.y);
                    if (i == F) {  // one writer per node
                        a.t.sum[c >> 1] = vs;
                        a.t.mnn[c >> 1] = vm;
                    }
                }
                if (sync_here) asm volatile("bar.sync 1, %0;" ::"n"(SORT_SLICE_MAX) : "memory");  // everybody has read level l-1
                if (act && (mR == l + 1 || mL == l + 1)) {  // somebody reads me one level up
                    *reinterpret_cast<double2 *>(S.v[i]) = make_double2(vs, vm);
                    *reinterpret_cast<short2 *>(S.fl[i]) = make_short2((short)F, (short)Lr);
                }
            }
        }
        UPD_TRACE(11);
        cl.sync();  // every CTA's leaves and unshared ancestors are in global memory
        UPD_TRACE(3);
        UPD_TRACE(4);
    } else {
    // phase 1
    double *s_leaf = s_sum;
#pragma unroll 1
    for (int i = tid; i < cnt; i += T) {
        int64_t k;
        double leaf, raw;
        int kk = -1;
        if (upd_item(a, lo + i, k, leaf, raw)) {
            kk = (int)k;
            s_leaf[i] = leaf;
            my_max = fmax(my_max, raw);
            if (a.sorted_idx) {
                if (lo + i + 1 < (int)a.n && a.idx[lo + i + 1] == a.idx[lo + i]) kk = -1;  // the next item overwrites this one
            } else if (a.stamps) {
                atomicMax(a.stamps + k, (u64)(lo + i) + 1);
            }
        }
        s_k[i] = kk;
    }
    if (a.stamps && !a.sorted_idx) {
        cl.sync();
#pragma unroll 1
        for (int i = tid; i < cnt; i += T) {
            const int kk = s_k[i];
            if (kk >= 0) {
                if (__ldcg(a.stamps + kk) != (u64)(lo + i) + 1) s_k[i] = -1;  // a later item wins
                else a.stamps[kk] = 0;  // the winner clears the entry (losers that look later still see "not mine")
            }
        }
    }
    UPD_TRACE(2);
    // phase 2 (each thread revisits the items it handled itself)
#pragma unroll 1
    for (int i = tid; i < cnt; i += T) {
        const int kk = s_k[i];
        if (kk >= 0) {
            const double leaf = s_leaf[i];
            if (a.t.sum) a.t.sum[a.t.cap + kk] = leaf;
            if (a.t.mnn) a.t.mnn[a.t.cap + kk] = leaf;
        }
    }
    if (a.max_prio) {
        if (a.src_kind == LEAF_GIVEN) my_max = (tid == 0 && rank == 0) ? a.batch_max : 0.0;
        for (int o = 16; o; o >>= 1) my_max = fmax(my_max, shfl_f64(my_max, (tid & 31) ^ o));
        if ((tid & 31) == 0) s_red[tid >> 5] = my_max;
        __syncthreads();
        if (tid < 32) {
            double m = tid < nwarp ? s_red[tid] : 0.0;
            for (int o = 16; o; o >>= 1) m = fmax(m, shfl_f64(m, tid ^ o));
            if (tid == 0 && m > 0.0) atomic_max_pos_f64(a.max_prio, m);
        }
    }
    cl.sync();
    UPD_TRACE(3);
    // phase 3
    const int sub = tid & (LIFT_LANES - 1);
#pragma unroll 1
    for (int done = 0; a.t.L - done > top; done += LIFT) {
#pragma unroll 1
        for (int base = 0; base < cnt; base += T / LIFT_LANES) {
            const int i = base + tid / LIFT_LANES;
            const int kk = i < cnt ? s_k[i] : -1;
            lift8_both(a.t, (a.t.cap + (int64_t)(kk < 0 ? 0 : kk)) >> (done + LIFT), sub, kk >= 0);
        }
        cl.sync();
    }
    UPD_TRACE(4);
    }
    // phase 4
    if (rank == 0) {
        // The shard record (roots + prefix chain) needs nothing from global memory that the dense rebuild writes:
        // the chain's terms above the dense top are taken from shared memory afterwards, the ones below it are
        // final since the last cluster barrier (sorted path: unless stage 2 recomputes them, see there) and are
        // requested now, so their round trip hides behind the rebuild.
        const bool want_rec = a.rec_out || a.publish_to;
        const bool want_partial = want_rec && a.rec_need > 0 && a.rec_need < a.t.cap;
        const bool emit = want_partial && tid < 32 && ((a.rec_need >> tid) & 1);
        const int64_t term_node = prefix_term_node(a.t.cap, a.rec_need, tid & 31);
        const bool term_low = a.t.L - (tid & 31) > top;  // the node lies below the dense top
        double term = (emit && term_low) ? __ldcg(a.t.sum + term_node) : 0.0;
        if (!sorted_path) {
            dense_load(a.t, 1, top, s_sum, s_min);
            __syncthreads();
        } else {
            // ---- stage 2: the nodes shared between CTAs, one chain per border, next to the dense top ----
            double *s2 = s_tree + S2_OFF_DOUBLES;  // [border][level - 1][sum L, sum R, min L, min R]
            double *s_chain_v = s2 + UPD_MAX_CTAS * SORT_CLIMB_MAX * 4;                      // [level - 1][border] fresh sums
            int *s_chain_n = reinterpret_cast<int *>(s_chain_v + SORT_CLIMB_MAX * UPD_MAX_CTAS);  // ... and their nodes
            int ckR = -1;  // warp 0, lane p: node number of the first leaf behind border p (borders in ascending order)
            int sh = 0;    // level of the lowest node the border's two leaves share
            int nbord = 0, l_first = climb + 1;
            if (tid < 32) {
                const int my_min = tid < ncta ? (int)s_bound[2 * tid] : -1, my_max_k = tid < ncta ? (int)s_bound[2 * tid + 1] : -1;
                int kL = __shfl_up_sync(0xffffffffu, my_max_k, 1);  // keys ascend with the CTA rank: the largest leaf in front
                if (tid == 0) kL = -1;                              // of mine is the maximum over the lanes below me
                for (int o = 1; o < UPD_MAX_CTAS; o <<= 1) {
                    const int other = __shfl_up_sync(0xffffffffu, kL, o);
                    if (tid >= o) kL = max(kL, other);
                }
                const bool border = my_min >= 0 && kL >= 0;
                const int my_sh = border ? lca_level(kL, my_min) : 0;
                const int my_ck = border ? (int)a.t.cap + my_min : -1;
                const unsigned mask = __ballot_sync(0xffffffffu, border);
                nbord = __popc(mask);
                const unsigned src = __fns(mask, 0, tid + 1);  // lane of the (tid+1)-th border
                const int g_ck = __shfl_sync(0xffffffffu, my_ck, src & 31), g_sh = __shfl_sync(0xffffffffu, my_sh, src & 31);
                if (tid < nbord) {
                    ckR = g_ck;
                    sh = g_sh;
                }
                l_first = __reduce_min_sync(0xffffffffu, tid < nbord ? sh : climb + 1);
                if (tid < UPD_MAX_CTAS) {
                    s_bound[2 * UPD_MAX_CTAS + 2 * tid] = ckR;
                    s_bound[2 * UPD_MAX_CTAS + 2 * tid + 1] = sh;
                }
                if (tid == 0) s_srt[SRT_CHAINS] = l_first <= climb;
            }
            dense_load(a.t, 1, top, s_sum, s_min);
            __syncthreads();
            UPD_TRACE(12);
            const bool chains = s_srt[SRT_CHAINS] != 0;  // some node below the dense top is shared between CTAs (a sample that crowds)
            if (chains) {
                if (tid < UPD_MAX_CTAS * SORT_CLIMB_MAX) {  // children of every chain node: one round trip
                    const int b = tid / SORT_CLIMB_MAX, l = tid % SORT_CLIMB_MAX + 1;
                    const int bck = (int)s_bound[2 * UPD_MAX_CTAS + 2 * b];
                    const int bsh = (int)s_bound[2 * UPD_MAX_CTAS + 2 * b + 1];
                    if (bck >= 0 && l >= bsh && l <= climb) {
                        const int64_t X = bck >> l;
                        const double2 cs = ldcg_f64x2(a.t.sum + 2 * X), cm = ldcg_f64x2(a.t.mnn + 2 * X);
                        double *d = s2 + (b * SORT_CLIMB_MAX + (l - 1)) * 4;
                        d[0] = cs.x;
                        d[1] = cs.y;
                        d[2] = cm.x;
                        d[3] = cm.y;
                    }
                }
                __syncthreads();
            }
            if (chains && tid < 32) {
                // Chains that have reached the same node form a run of adjacent lanes (borders ascend).  Inside a run the
                // lanes whose border lies in the left child come first, then at most one lane whose border is the split
                // itself, then the lanes of the right child: if a child is shared at all, the run's first lane holds the
                // left child's fresh value and its last lane the right child's.
                int pn = -1;  // my chain's node one level down and its fresh value
                double pvs = 0.0, pvm = 0.0;
                for (int t = tid; t < (l_first - 1) * UPD_MAX_CTAS; t += 32) s_chain_n[t] = -1;  // levels no chain reaches
#pragma unroll 1
                for (int l = l_first; l <= climb; ++l) {
                    const bool on = tid < nbord && l >= sh;
                    const int X = on ? (ckR >> l) : -1;
                    double ls = 0.0, rs = 0.0, lm = 0.0, rm = 0.0;
                    if (on) {
                        const double *d = s2 + (tid * SORT_CLIMB_MAX + (l - 1)) * 4;
                        ls = d[0];
                        rs = d[1];
                        lm = d[2];
                        rm = d[3];
                    }
                    const int Xn = __shfl_down_sync(0xffffffffu, X, 1);
                    const unsigned ends = __ballot_sync(0xffffffffu, !(on && tid < 31 && Xn == X));  // bit p: a run ends at lane p
                    const int g1 = tid + __ffs(ends >> tid) - 1;                                     // (bit 31 is always set)
                    const unsigned before = tid ? (ends & ((1u << tid) - 1u)) : 0u;
                    const int g0 = before ? 32 - __clz(before) : 0;
                    const int n0 = __shfl_sync(0xffffffffu, pn, g0), n1 = __shfl_sync(0xffffffffu, pn, g1);
                    const double s0 = shfl_f64(pvs, g0), m0 = shfl_f64(pvm, g0);
                    const double s1 = shfl_f64(pvs, g1), m1 = shfl_f64(pvm, g1);
                    if (on) {
                        if (n0 == 2 * X) { ls = s0; lm = m0; }
                        if (n1 == 2 * X + 1) { rs = s1; rm = m1; }
                    }
                    const double nvs = __dadd_rn(ls, rs), nvm = fmin(lm, rm);
                    if (on) {
                        a.t.sum[X] = nvs;
                        a.t.mnn[X] = nvm;
                        if (l == climb) {  // bottom row of the dense top: heap slot == node number
                            s_sum[X] = nvs;
                            s_min[X] = nvm;
                        }
                    }
                    if (tid < UPD_MAX_CTAS) {
                        s_chain_n[(l - 1) * UPD_MAX_CTAS + tid] = X;
                        s_chain_v[(l - 1) * UPD_MAX_CTAS + tid] = nvs;
                    }
                    pn = X;
                    pvs = nvs;
                    pvm = nvm;
                }
                __syncwarp();
                if (emit && term_low && tid >= 1 && tid <= climb) {  // a term of the record's chain that stage 2 has just recomputed
                    for (int c2 = 0; c2 < UPD_MAX_CTAS; ++c2)
                        if (s_chain_n[(tid - 1) * UPD_MAX_CTAS + c2] == (int)term_node) term = s_chain_v[(tid - 1) * UPD_MAX_CTAS + c2];
                }
            }
            if (chains) __syncthreads();
        }
        // the top in shared memory; the record leaves BEFORE the write-out (its readers need none of the node stores,
        // and this rank's own next sample waits for this kernel to retire anyway): eight self-validating words per
        // peer, no fence (see emit_record_values)
        dense_build(a.t, 1, top, s_sum, s_min);
        UPD_TRACE(5);
        if (want_rec && tid < 32) {
            if (emit && !term_low) term = s_sum[term_node];
            const double partial = want_partial ? warp_prefix_fold(term, a.rec_need) : 0.0;
            emit_record_values(s_sum[1], s_min[1], partial, a.rec_len, a.rec_out, a.publish_to, a.n_shards, a.shard_rank,
                               a.publish_to ? __ldcg(a.d_epoch) + 1 : 0);
        }
        dense_write(a.t, 1, top, s_sum, s_min);
        UPD_TRACE(13);
        if (tl_on) tl_stamp(a.tl, 3, 9);
    }
}

// ---------------------------------------------------------------------------------------
// Dense form for LARGE sorted batches (the update of the sample just drawn): when a batch dirties a sizeable share of the
// tree, rebuilding whole 1 024-leaf chunks costs less than chasing every dirty path through barriers — a pair of
// 2^20-leaf trees is 16 MB of leaves in and 16 MB of nodes out, a few microseconds of L2 streaming.  ONE launch; the
// tree is cut into chunks of 1 024 leaves and CTA b owns the contiguous range of chunks [b C / G, (b + 1) C / G):
//   1. the items that fall into that range are a contiguous run of the ascending index array: its two ends are found by
//      two warps side by side (32-ary searches);
//   2. the run is walked in pieces of 128 items: the items of the piece that belong to the chunk of its first item write
//      their leaves (equal indices are adjacent: the last of a run wins, memory.py:314-318) and mark them in a 1 024-bit
//      map; when the chunk's last item has gone by, the chunk is lifted.  Chunks without items are never touched;
//   3. dense_lift(): every thread takes 8 consecutive leaves of each tree in two 32-byte loads, folds three generations in
//      registers, five more by warp shuffles, the last two through 64 bytes of shared memory; nodes go out as 32-, 16- and
//      8-byte stores.  A thread whose 8 leaves are clean fetches the node three generations up instead, and a node is
//      stored only if something below it changed.  No shared-memory staging of the chunk: 10 resident CTAs per SM (the
//      first version staged the chunk in 32 KB: 6 CTAs per SM, 1.15 waves at 2^20 leaves, 70 instructions per leaf —
//      23-27 us where this one takes 12-13, profiles/r02_dense_update_ncu.csv);
//   4. chunks are counted per GROUP of 1 024 chunks (self-resetting counters); the CTA that completes a group lifts the
//      ten generations above its chunk roots, and the CTA that completes the last group lifts what is left above the
//      group roots (trees above 2^20 leaves) and writes the shard record.
// Same node values as every other form: the node array is a function of the leaves alone.
// ---------------------------------------------------------------------------------------
constexpr int DENSE_CHUNK_LOG2 = 10;  // leaves per chunk = 8 per thread
constexpr int DENSE_THREADS = 128;
constexpr int DENSE_GROUP_LOG2 = 10;  // chunks per group: one dense_lift above their roots
constexpr int DENSE_L_MIN = 12, DENSE_L_MAX = 26;  // tree depths the form is offered for (run_update narrows it further)
constexpr int DENSE_MAX_GROUPS = 1 << (DENSE_L_MAX - DENSE_CHUNK_LOG2 - DENSE_GROUP_LOG2);  // 64 group counters + 1 for the groups

// first position j in [lo, hi) with key(j) >= want (hi if none), keys ascending; all lanes of the warp return it
__device__ __forceinline__ int64_t dense_lower_bound(const UpdArgs &a, int64_t lo, int64_t hi, int64_t want, int lane) {
    while (hi - lo > 0) {
        const int64_t n = hi - lo, step = (n + 31) >> 5;
        int64_t j = lo + (lane + 1) * step - 1;  // last position of this lane's chunk
        if (j > hi - 1) j = hi - 1;
        const bool ge = upd_index(a, j) >= want;
        const unsigned m = __ballot_sync(0xffffffffu, ge);
        if (m == 0) return hi;
        const int f = __ffs(m) - 1;
        const int64_t cb = lo + f * step;
        int64_t ce = cb + step - 1;
        if (ce > hi - 1) ce = hi - 1;
        lo = cb;
        hi = ce;  // position ce satisfies the predicate; [cb, ce) undecided
        if (hi == lo) return lo;
    }
    return lo;
}

// 32-byte accesses that are coherent at L2 (the nodes read were written by other threads of this launch)
__device__ __forceinline__ void ld_cg_f64x4(const double *p, double &a, double &b, double &c, double &d) {
    asm volatile("ld.global.cg.v4.f64 {%0, %1, %2, %3}, [%4];" : "=d"(a), "=d"(b), "=d"(c), "=d"(d) : "l"(p) : "memory");
}
__device__ __forceinline__ void st_f64x4(double *p, double a, double b, double c, double d) {
    asm volatile("st.global.v4.f64 [%0], {%1, %2, %3, %4};" ::"l"(p), "d"(a), "d"(b), "d"(c), "d"(d) : "memory");
}

// Rebuilds, in both trees, the g generations (1 <= g <= 10) above the 2^g consecutive nodes below node `sub_root`
// (children of generation g: nodes (sub_root << g) + i).  Called by all DENSE_THREADS threads of the CTA; the nodes read
// must be visible to the CTA (written before a __syncthreads() by this CTA, or fenced by their writers).  s_w: 8 doubles.
// Rebuilds, in both trees, the g generations (1 <= g <= 10) above the 2^g consecutive nodes below node `sub_root`
// (children of generation g: nodes (sub_root << g) + i).  Called by all DENSE_THREADS threads of the CTA; the nodes read
// must be visible to the CTA (written before a __syncthreads() by this CTA, or fenced by their writers).  s_w: 8 doubles.
// dirty8: which of this thread's 8 nodes changed (0xff where that is not known).  A thread whose 8 nodes are clean
// fetches the node three generations above them instead (it is right already) and a node goes out only if something
// below it changed: a chunk with a handful of new leaves costs 2 KB of reads and the dirty paths, not 32 KB each way.
__device__ __forceinline__ void dense_lift(const TreeRef &t, const int64_t sub_root, const int g, double *s_w, const unsigned dirty8 = 0xffu) {
    const int tid = threadIdx.x, lane = tid & 31;
    if (g < 3) {  // at most four nodes: one thread, level by level (it re-reads what it has just written)
        if (tid == 0) {
            for (int lvl = g - 1; lvl >= 0; --lvl)
                for (int i = 0; i < (1 << lvl); ++i) {
                    const int64_t node = (sub_root << lvl) + i;
                    const double2 cs = ldcg_f64x2(t.sum + 2 * node), cm = ldcg_f64x2(t.mnn + 2 * node);
                    t.sum[node] = __dadd_rn(cs.x, cs.y);
                    t.mnn[node] = fmin(cm.x, cm.y);
                }
        }
        return;
    }
    const bool on = tid < (1 << (g - 3));  // threads with 8 nodes each
    double vs = 0.0, vm = 0.0;
    bool dirty = on && dirty8 != 0;
    if (on && !dirty) {
        const int64_t n0 = (sub_root << g) + 8 * tid;
        vs = ldcg_f64(t.sum + (n0 >> 3));
        vm = ldcg_f64(t.mnn + (n0 >> 3));
    }
    if (dirty) {
        const int64_t n0 = (sub_root << g) + 8 * tid;
        double x0, x1, x2, x3, x4, x5, x6, x7, y0, y1, y2, y3, y4, y5, y6, y7;
        ld_cg_f64x4(t.sum + n0, x0, x1, x2, x3);
        ld_cg_f64x4(t.sum + n0 + 4, x4, x5, x6, x7);
        ld_cg_f64x4(t.mnn + n0, y0, y1, y2, y3);
        ld_cg_f64x4(t.mnn + n0 + 4, y4, y5, y6, y7);
        const double a0 = __dadd_rn(x0, x1), a1 = __dadd_rn(x2, x3), a2 = __dadd_rn(x4, x5), a3 = __dadd_rn(x6, x7);
        const double b0 = fmin(y0, y1), b1 = fmin(y2, y3), b2 = fmin(y4, y5), b3 = fmin(y6, y7);
        st_f64x4(t.sum + (n0 >> 1), a0, a1, a2, a3);
        st_f64x4(t.mnn + (n0 >> 1), b0, b1, b2, b3);
        const double c0 = __dadd_rn(a0, a1), c1 = __dadd_rn(a2, a3), d0 = fmin(b0, b1), d1 = fmin(b2, b3);
        *reinterpret_cast<double2 *>(t.sum + (n0 >> 2)) = make_double2(c0, c1);
        *reinterpret_cast<double2 *>(t.mnn + (n0 >> 2)) = make_double2(d0, d1);
        vs = __dadd_rn(c0, c1);
        vm = fmin(d0, d1);
        t.sum[n0 >> 3] = vs;
        t.mnn[n0 >> 3] = vm;
    }
    const int64_t n3 = (sub_root << (g - 3)) + tid;  // this thread's node three generations up
    const int shuf = g - 3 < 5 ? g - 3 : 5;
#pragma unroll
    for (int s = 1; s <= 5; ++s) {
        if (s <= shuf) {  // the same for the whole CTA
            const double ps = shfl_f64(vs, (lane + (1 << (s - 1))) & 31), pm = shfl_f64(vm, (lane + (1 << (s - 1))) & 31);
            const bool pd = __shfl_sync(0xffffffffu, (int)dirty, (lane + (1 << (s - 1))) & 31) != 0;
            if ((lane & ((1 << s) - 1)) == 0) {  // left operand here, right operand 2^(s-1) lanes up
                vs = __dadd_rn(vs, ps);
                vm = fmin(vm, pm);
                dirty = dirty || pd;
                if (on && dirty) {
                    t.sum[n3 >> s] = vs;
                    t.mnn[n3 >> s] = vm;
                }
            }
        }
    }
    if (g - 3 > 5) {  // two or four warps hold one subtree root each (callers lift only what has a dirty node: all stored)
        if (lane == 0 && on) {
            s_w[tid >> 5] = vs;
            s_w[4 + (tid >> 5)] = vm;
        }
        __syncthreads();
        if (tid == 0) {
            if (g == 10) {
                const double l0 = __dadd_rn(s_w[0], s_w[1]), l1 = __dadd_rn(s_w[2], s_w[3]);
                const double m0 = fmin(s_w[4], s_w[5]), m1 = fmin(s_w[6], s_w[7]);
                *reinterpret_cast<double2 *>(t.sum + 2 * sub_root) = make_double2(l0, l1);
                *reinterpret_cast<double2 *>(t.mnn + 2 * sub_root) = make_double2(m0, m1);
                t.sum[sub_root] = __dadd_rn(l0, l1);
                t.mnn[sub_root] = fmin(m0, m1);
            } else {
                t.sum[sub_root] = __dadd_rn(s_w[0], s_w[1]);
                t.mnn[sub_root] = fmin(s_w[4], s_w[5]);
            }
        }
        __syncthreads();  // s_w is free again
    }
}

__global__ void __launch_bounds__(DENSE_THREADS, 8) k_upd_dense_sorted(UpdArgs a, unsigned *counter) {
    constexpr int CHUNK = 1 << DENSE_CHUNK_LOG2;
    __shared__ int64_t s_run[2];
    __shared__ int64_t s_chunk;
    __shared__ int s_lift[2];  // whether this CTA's count completed the group at hand / the last group of all
    __shared__ double s_max[DENSE_THREADS / 32];
    __shared__ double s_w[8];
    __shared__ unsigned s_dirty[CHUNK / 32];  // leaves of the chunk that got a new value
    // Behind the TD launch of the same update (early_search) the item run is searched BEFORE the wait: that launch
    // releases its dependents only behind its own wait, so the sample that wrote the indices and the owned range has
    // retired; the priorities it is still computing are read behind the wait.  With early publish the sampling kernel
    // behind this one starts its exchange before its wait and relies on every kernel up to the previous sample having
    // retired: dependents are then released behind the wait (as k_update_cluster does).
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    bool waited = !a.early_search;
    if (waited) {
        if (!a.publish_to) pdl_launch_dependents();
        pdl_wait();
        if (a.publish_to) pdl_launch_dependents();
    }
#define DENSE_TRACE(slot, any_cta)                                        \
    do {                                                                  \
        if (a.trace && tid == 0 && ((any_cta) || blockIdx.x == 0)) {      \
            u64 t__;                                                      \
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t__));       \
            a.trace[slot] = t__;                                          \
        }                                                                 \
    } while (0)
    if (!a.early_search) DENSE_TRACE(0, false);
    int first, last;
    owned_range(a.range, a.n, first, last);
    const int64_t chunks = a.t.cap >> DENSE_CHUNK_LOG2;
    const int64_t c0 = chunks * blockIdx.x / gridDim.x, c1 = chunks * (blockIdx.x + 1) / gridDim.x;  // this CTA's chunks
    if (warp == 0) {
        const int64_t lo = dense_lower_bound(a, first, last, c0 << DENSE_CHUNK_LOG2, lane);
        if (lane == 0) s_run[0] = lo;
    } else if (warp == 1) {
        const int64_t hi = dense_lower_bound(a, first, last, c1 << DENSE_CHUNK_LOG2, lane);
        if (lane == 0) s_run[1] = hi;
    } else if (warp == 2) {
        s_dirty[lane] = 0;
    }
    DENSE_TRACE(14, false);  // run found
    if (!waited) {
        pdl_wait();
        pdl_launch_dependents();
    }
    DENSE_TRACE(1, false);  // priorities ready
    __syncthreads();
    const int64_t lo = s_run[0], hi = s_run[1];
    double my_max = 0.0;
    int64_t j = lo;
    int64_t open_chunk = -1;  // chunk with leaves written and not lifted yet (the same in every thread)
    while (j < hi) {  // pieces of up to DENSE_THREADS items, all of one chunk
        const bool have = j + tid < hi;
        const int64_t key = have ? upd_index(a, j + tid) : -1;  // (valid or not: what the run was searched by)
        if (tid == 0) s_chunk = key >> DENSE_CHUNK_LOG2;
        __syncthreads();
        const int64_t c = s_chunk;
        if (c != open_chunk && open_chunk >= 0) {  // the chunk in front is complete: lift it (ends with a CTA barrier)
            const unsigned dirty8 = (s_dirty[tid >> 2] >> ((tid & 3) * 8)) & 0xffu;
            dense_lift(a.t, (a.t.cap >> DENSE_CHUNK_LOG2) + open_chunk, DENSE_CHUNK_LOG2, s_w, dirty8);
            if (tid < CHUNK / 32) s_dirty[tid] = 0;
            __syncthreads();
        }
        open_chunk = c;
        const bool mine = have && (key >> DENSE_CHUNK_LOG2) == c;
        if (c < c0 || c >= c1) {  // cannot happen for an ascending array
            if (tid == 0) atomicOr(a.status, (u64)PER_STATUS_BAD_INDEX);
            open_chunk = -1;
            break;
        }
        int64_t k = -1;
        double leaf = 0.0, raw = 0.0;
        if (mine && upd_item(a, j + tid, k, leaf, raw)) {  // (pow only for the items of this chunk)
            my_max = fmax(my_max, raw);
            const int64_t nxt = j + tid + 1 < hi ? upd_index(a, j + tid + 1) : INT64_MAX;
            if (nxt < k) atomicOr(a.status, (u64)PER_STATUS_BAD_INDEX);  // not ascending: not the sample's buffer any more
            if (nxt != k) {
                a.t.sum[a.t.cap + k] = leaf;
                a.t.mnn[a.t.cap + k] = leaf;
                atomicOr(&s_dirty[(k & (CHUNK - 1)) >> 5], 1u << (k & 31));
            }
        }
        j += __syncthreads_count(mine);  // (also: the leaves just written are visible to the CTA's loads)
    }
    DENSE_TRACE(10, false);  // leaves written
    if (open_chunk >= 0) {
        const unsigned dirty8 = (s_dirty[tid >> 2] >> ((tid & 3) * 8)) & 0xffu;
        dense_lift(a.t, (a.t.cap >> DENSE_CHUNK_LOG2) + open_chunk, DENSE_CHUNK_LOG2, s_w, dirty8);
    }
    DENSE_TRACE(11, false);  // chunks lifted
    if (a.max_prio) {
        if (a.src_kind == LEAF_GIVEN) my_max = (tid == 0 && blockIdx.x == 0) ? a.batch_max : 0.0;
        for (int o = 16; o; o >>= 1) my_max = fmax(my_max, shfl_f64(my_max, lane ^ o));
        if (lane == 0) s_max[warp] = my_max;
    }
    // Count this CTA's chunks into their groups.  A CTA that completes a group lifts the generations above the group's
    // chunk roots; the one that completes the last group lifts what is above the group roots.
    __threadfence();
    __syncthreads();
    DENSE_TRACE(3, false);  // fenced
    if (tid == 0 && a.max_prio) {
        double m = 0.0;
        for (int w = 0; w < DENSE_THREADS / 32; ++w) m = fmax(m, s_max[w]);
        if (m > 0.0) atomic_max_pos_f64(a.max_prio, m);
    }
    const int top_gens = a.t.L - DENSE_CHUNK_LOG2;                                  // generations above the chunk roots
    const int group_gens = top_gens < DENSE_GROUP_LOG2 ? top_gens : DENSE_GROUP_LOG2;  // ... of which a group lift takes these
    const int64_t n_groups = chunks >> group_gens;
    bool closes_all = false;
    for (int64_t g = c0 >> group_gens; g <= (c1 - 1) >> group_gens && c1 > c0; ++g) {
        if (tid == 0) {
            const int64_t g_lo = g << group_gens, g_hi = (g + 1) << group_gens;
            const unsigned mine = (unsigned)((c1 < g_hi ? c1 : g_hi) - (c0 > g_lo ? c0 : g_lo));
            const unsigned before = atomicAdd(counter + g, mine);
            s_lift[0] = before + mine == (1u << group_gens);
            s_lift[1] = 0;
            if (s_lift[0]) {
                counter[g] = 0;  // ready for the next launch
                __threadfence();
            }
        }
        __syncthreads();
        if (s_lift[0]) {  // (the same for the whole CTA)
            DENSE_TRACE(12, true);  // a group completed
            dense_lift(a.t, (int64_t(1) << (top_gens - group_gens)) + g, group_gens, s_w);
            __syncthreads();
            if (n_groups > 1) {
                __threadfence();
                __syncthreads();
            }
            if (n_groups == 1) {  // (trees up to 2^20 leaves: nothing above the group)
                closes_all = true;
                break;
            }
            if (tid == 0) {
                const unsigned before = atomicAdd(counter + DENSE_MAX_GROUPS, 1u);
                s_lift[1] = before + 1 == (unsigned)n_groups;
                if (s_lift[1]) {
                    counter[DENSE_MAX_GROUPS] = 0;
                    __threadfence();
                }
            }
            __syncthreads();
            closes_all = closes_all || s_lift[1];
        }
        __syncthreads();  // s_lift is rewritten by the next round
    }
    if (!closes_all) return;
    if (top_gens > group_gens) {
        dense_lift(a.t, 1, top_gens - group_gens, s_w);  // at most 6 generations above the group roots
        __syncthreads();
    }
    DENSE_TRACE(5, true);  // top lifted
    if (a.rec_out || a.publish_to) {  // sharded replay: the record of the refreshed tree, for this rank and (early publish) its peers
        if (tid < 32)     // (reads nodes this CTA wrote before the barrier above, or fenced by the CTAs that have counted)
            emit_record(a.t.sum, a.t.mnn, a.t.cap, a.rec_need, a.rec_len, a.rec_out, a.publish_to, a.n_shards, a.shard_rank,
                        a.publish_to ? __ldcg(a.d_epoch) + 1 : 0);
    }
    DENSE_TRACE(13, true);
#undef DENSE_TRACE
    if (a.tl && tid == 0) tl_stamp(a.tl, 3, 9);  // (the CTA that finishes, not CTA 0: this is the end of the update)
}

// ---------------------------------------------------------------------------------------
// Multi-launch form (any n): the same phases as separate grid-wide launches.
// ---------------------------------------------------------------------------------------
__global__ void k_upd_stamp(UpdArgs a) {
    for (int64_t j = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; j < a.n; j += (int64_t)gridDim.x * blockDim.x) {
        int64_t k;
        if (upd_item_ok(a, j, k)) atomicMax(a.stamps + k, (u64)j + 1);
    }
}

// The winner of a leaf clears its stamp on the way (see UpdArgs::stamps).
__device__ __forceinline__ bool upd_is_winner(const UpdArgs &a, int64_t j, int64_t k) {
    if (!a.stamps) return true;
    if (__ldcg(a.stamps + k) != (u64)j + 1) return false;
    a.stamps[k] = 0;
    return true;
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

// One sparse round above depth (L - done) for every valid item, LIFT_LANES lanes per item.
__global__ void __launch_bounds__(256) k_upd_sparse(UpdArgs a, int done) {
    const int sub = threadIdx.x & (LIFT_LANES - 1);
    const int64_t group = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) / LIFT_LANES;
    const int64_t groups = ((int64_t)gridDim.x * blockDim.x) / LIFT_LANES;
    const int64_t rounds = (a.n + groups - 1) / groups;  // same trip count for every lane of a warp
    for (int64_t r = 0; r < rounds; ++r) {
        const int64_t j = r * groups + group;
        int64_t k = j < a.n ? upd_index(a, j) : -1;
        const bool valid = k >= 0 && k < a.len_limit;
        lift8_both(a.t, (a.t.cap + (valid ? k : 0)) >> (done + LIFT), sub, valid);
    }
}

// CTA b rebuilds the k generations above the 2^k nodes below node (2^(d-k) + b).
__global__ void __launch_bounds__(256) k_dense_mid(TreeRef t, int d, int k) {
    __shared__ __align__(16) double s_sum[2 * TOP_NODES];
    __shared__ __align__(16) double s_min[2 * TOP_NODES];
    dense_lift_block(t, (int64_t(1) << (d - k)) + blockIdx.x, k, s_sum, s_min);
}

__global__ void __launch_bounds__(1024) k_dense_top(TreeRef t, int top, double *rec_out, int64_t rec_need, int64_t rec_len) {
    __shared__ __align__(16) double s_sum[2 * TOP_NODES];
    __shared__ __align__(16) double s_min[2 * TOP_NODES];
    dense_lift_block(t, 1, top, s_sum, s_min);
    if (rec_out) {  // the record's prefix chain reads nodes other threads have just written
        __threadfence();
        __syncthreads();
    }
    if (rec_out && threadIdx.x < 32) write_shard_record(t.sum, t.mnn, t.cap, rec_need, rec_len, rec_out);
}

}  // namespace perb
