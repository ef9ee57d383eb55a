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
//
// This file: range reductions and the stratified descent (K2+K3) in three forms — k_sample (one warp per
// query, small batches), k_sample_wide (one thread per query, batches >= 2048), k_sample_sharded (one slice of a
// replay cut over several GPUs, with the record exchange through peer-memory mailboxes).  The batched update (K1)
// lives in update_kernels.cuh.  Both are latency-bound (a batch of 512 touches ~0.4 MB), so
// they are written for a short dependent chain and a small instruction footprint (rolled
// loops, one pow() call site per kernel).
#pragma once
#include "common.cuh"

namespace perb {

// ---------------------------------------------------------------------------------------
// Range reductions with the reference's association (memory.py:39-52).
// ---------------------------------------------------------------------------------------

// Warp-cooperative  sum over leaves [0, need)  =  t1 + (t2 + (... + tk)).  Walking the
// reference's recursion (memory.py:39-52) from the root with start = 0, a term is emitted for
// every set bit b of `need`, most significant first: the aligned block of 2^b leaves that
// starts where the higher bits of `need` end, i.e. node (cap >> b) + (need >> (b+1) << 1).
// Lane b owns bit b, so all loads go out together; the chain is then folded innermost
// (lowest bit) first.  Every lane returns the total.  Requires cap <= 2^31.
__device__ __forceinline__ int64_t prefix_term_node(int64_t cap, int64_t need, int b) { return (cap >> b) + ((need >> (b + 1)) << 1); }
// lane b holds the term of bit b (anything where the bit is clear); every lane returns t1 + (t2 + (... + tk))
__device__ __forceinline__ double warp_prefix_fold(const double term, const int64_t need) {
    double acc = 0.0;
    bool first = true;
#pragma unroll
    for (int b = 0; b < 32; ++b) {
        const double t = shfl_f64(term, b);  // independent of acc: the shuffles pipeline
        if ((need >> b) & 1) {
            acc = first ? t : __dadd_rn(t, acc);
            first = false;
        }
    }
    return acc;
}
__device__ __forceinline__ double warp_prefix_total(const double *v, int64_t cap, int64_t need) {
    const int lane = threadIdx.x & 31;
    const bool emit = (need >> lane) & 1;
    const double term = emit ? __ldcg(v + prefix_term_node(cap, need, lane)) : 0.0;
    return warp_prefix_fold(term, need);
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
    double terms[64];
    int nt = 0;
    double left_part;
    {  // suffix of the left child: leaves [lo, mid)
        int64_t n = 2 * node, s_lo = nlo, w = half;
        while (!(lo == s_lo)) {
            int64_t h = w >> 1;
            if (lo >= s_lo + h) { n = 2 * n + 1; s_lo += h; }
            else { terms[nt++] = v[2 * n + 1]; n = 2 * n; }
            w = h;
        }
        double acc = v[n];
        for (int i = nt - 1; i >= 0; --i) acc = op(acc, terms[i]);
        left_part = acc;
    }
    nt = 0;
    double right_part;
    {  // prefix of the right child: leaves [mid, hi]
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
// of the next 7 generations number 1+2+...+64 = 127 (generation j, pair q -> node i*2^j + 2q):
// the warp fetches them with four independent loads per lane in one round trip and the seven
// compare / subtract steps then run out of registers through shuffles: ceil(L/7) dependent
// memory round trips per query instead of L (three for a 2^20-leaf tree, the first of them
// issued before the query's mass is known).  The subtraction sequence is the reference's, so
// the index is bit-exact.
// ---------------------------------------------------------------------------------------
struct SampleArgs {
    const double *sum;
    const double *mnn;     // null: no weights (stand-alone descent)
    int64_t cap;
    int L;
    const double *u;       // uniforms (per_sample_indices) or masses (seg_find_prefixsum)
    int direct_mass;       // 1: `u` holds masses already
    int64_t batch;
    int64_t n_rows;        // len(_storage)  (sharded: of the whole replay)
    double beta;
    int64_t *out_idx;
    double *out_w64;
    float *out_w32;
    u64 *status;
    // Sharded replay: this tree is slice `rank` of a virtual tree of
    // n_shards * cap leaves whose depth-log2(n_shards) nodes are the shard roots.
    const double *recs;    // null: not sharded.  n_shards records {root_sum, root_min, boundary_partial, len}
    int n_shards;          // power of two, <= 16
    int rank;
    // Record exchange through peer memory instead of a collective (epoch != 0): this kernel
    // publishes its own record into every rank's mailbox and waits for the others'.
    double *mailbox;             // this rank's mailbox: [2 parities][n_shards] slots of MAIL_SLOT doubles
    double *const *peer_mailbox; // device array [n_shards] of the mailboxes of all ranks (peer-mapped)
    u64 *d_epoch;                // null: no exchange.  Device counter of completed exchange samples; this launch is
                                 // sample number *d_epoch + 1 (slot parity = that & 1) and bumps it when its last warp retires,
                                 // so the same captured launch can be replayed from a CUDA graph
    unsigned *d_done;            // warps of this launch that have retired
    u64 *d_range;                // {lo tag, hi tag}: strata owned by this shard in this sample (see owned_range)
    int use_mailbox;             // 1: exchange through the mailboxes; 0: records come in `recs`
    u64 *d_wait_stats;           // {waits, total ns, max ns} spent by CTA 0 between kernel start and the last record's arrival
    int64_t rec_need, rec_len;   // this shard's part of the global prefix, local len
    u64 *tl;                     // timeline (diagnostics), see tl_stamp: this launch's first entry
    int skip_publish;            // 1: this epoch's record is in the mailboxes already (published by the update kernel)
};

constexpr int SHARD_REC = 4;   // doubles per shard record
constexpr int MAIL_SLOT = 8;   // 64-bit words per mailbox slot (64 B): the record's 4 doubles as 8 self-validating words, see below

__device__ __forceinline__ u64 global_timer_ns() {
    u64 t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

// One warp: computes this shard's record once and stores it where asked: into `rec_out` (4 doubles,
// the all-gather path) and / or into the slot [epoch & 1][rank] of every rank's mailbox (lane p -> rank p;
// plain stores to peer-mapped addresses travel over NVLink).
//
// Mailbox wire format — no flag, no fence.  A slot is 8 words of 64 bits; word 2 i carries the HIGH 32 bits of the record's
// i-th double in its upper half and word 2 i + 1 the LOW 32 bits, each with the low 32 bits of the sample's epoch in its lower
// half.  Every word is one aligned 8-byte store, which is single-copy atomic, so a reader that sees the epoch it waits for in
// a word has that word's payload whole, whatever order the eight stores arrive in: it polls until all eight carry the epoch.
// A release store of a separate flag would first have to wait for the record's stores to be acknowledged by the peer — an
// NVLink round trip (~2 us) at the end of the update kernel, in front of every sample of every rank.  Slots are double-
// buffered by epoch parity, so a word of epoch e is overwritten only by epoch e + 2, which no reader confuses with e.
// The record (one per shard and sample call): the two roots, and the shard-local part of the
// reference's  sum(0, len-1)  chain (memory.py:244): the prefix over this shard's first `need_local`
// leaves when the global prefix ends inside this shard.  Nodes are read through L2 (the caller may
// have just written them).
__device__ __forceinline__ void st_relaxed_sys_u64(u64 *p, u64 v) {
    asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ u64 ld_relaxed_sys_u64(const u64 *p) {
    u64 v;
    asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void emit_record_values(const double rs, const double rm, const double partial, int64_t len_local,
                                                   double *rec_out, double *const *peer_mailbox, int n_shards, int rank, u64 epoch) {
    const int lane = threadIdx.x & 31;
    if (rec_out && lane == 0) {
        rec_out[0] = rs;
        rec_out[1] = rm;
        rec_out[2] = partial;
        rec_out[3] = (double)len_local;
    }
    if (peer_mailbox && lane < n_shards) {
        u64 *slot = reinterpret_cast<u64 *>(peer_mailbox[lane]) + ((epoch & 1) * n_shards + rank) * MAIL_SLOT;
        const u64 tag = epoch & 0xffffffffull;
        const double vals[SHARD_REC] = {rs, rm, partial, (double)len_local};
#pragma unroll
        for (int i = 0; i < SHARD_REC; ++i) {
            const u64 bits = (u64)__double_as_longlong(vals[i]);
            st_relaxed_sys_u64(slot + 2 * i, (bits & 0xffffffff00000000ull) | tag);
            st_relaxed_sys_u64(slot + 2 * i + 1, (bits << 32) | tag);
        }
    }
}
__device__ __forceinline__ void emit_record(const double *sum, const double *mnn, int64_t cap, int64_t need_local, int64_t len_local,
                                            double *rec_out, double *const *peer_mailbox, int n_shards, int rank, u64 epoch) {
    double partial = 0.0;
    if (need_local > 0 && need_local < cap) partial = warp_prefix_total(sum, cap, need_local);
    emit_record_values(__ldcg(sum + 1), __ldcg(mnn + 1), partial, len_local, rec_out, peer_mailbox, n_shards, rank, epoch);
}
__device__ __forceinline__ void publish_record(const double *sum, const double *mnn, int64_t cap, int64_t need_local,
                                               int64_t len_local, double *const *peer_mailbox, int n_shards, int rank, u64 epoch) {
    emit_record(sum, mnn, cap, need_local, len_local, nullptr, peer_mailbox, n_shards, rank, epoch);
}

// All lanes of the calling warp return once every rank's slot of this epoch has arrived in the local mailbox (or after
// 2 s, with PER_STATUS_EXCHANGE_TIMEOUT set — never a hang); the decoded records (4 doubles per rank) are left in `out`
// (shared memory of the calling CTA, n_shards * SHARD_REC doubles).  Lane l polls the words l, l + 32, ... of the
// n_shards * 8 words of this epoch's half of the mailbox.
__device__ __forceinline__ void await_records(const double *mailbox, int n_shards, u64 epoch, u64 *status, double *out) {
    const int lane = threadIdx.x & 31;
    const u64 *half = reinterpret_cast<const u64 *>(mailbox) + (epoch & 1) * n_shards * MAIL_SLOT;
    const u64 tag = epoch & 0xffffffffull;
    const int nwords = n_shards * MAIL_SLOT;  // <= 128: at most 4 words per lane
    u64 w[4] = {0, 0, 0, 0};
    u64 t0 = 0;
    bool started = false;
    while (true) {
        bool all = true;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int k = lane + 32 * j;
            if (k < nwords && (w[j] & 0xffffffffull) != tag) {
                w[j] = ld_relaxed_sys_u64(half + k);
                all = all && (w[j] & 0xffffffffull) == tag;
            }
        }
        if (__all_sync(0xffffffffu, all)) break;
        if (!started) {
            t0 = global_timer_ns();
            started = true;
        }
        __nanosleep(100);  // back off: thousands of pollers on one L2 line delay the very write they wait for
        if (global_timer_ns() - t0 > 2000000000ull) {
            if (lane == 0) atomicOr(status, (u64)PER_STATUS_EXCHANGE_TIMEOUT);
            break;
        }
    }
    // word k = 8 r + 2 i + h: rank r, value i, half h (0 = high 32 bits).  Pairs (h = 0, 1) sit in adjacent lanes.
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const int k = lane + 32 * j;
        const u64 other = __shfl_down_sync(0xffffffffu, w[j], 1);
        if (k < nwords && (k & 1) == 0) {
            const u64 bits = (w[j] & 0xffffffff00000000ull) | (other >> 32);
            out[k >> 1] = __longlong_as_double((long long)bits);
        }
    }
    __syncwarp();
}

__global__ void k_shard_publish(const double *sum, const double *mnn, int64_t cap, int64_t need_local, int64_t len_local,
                                double *const *peer_mailbox, int n_shards, int rank, const u64 *d_epoch) {
    publish_record(sum, mnn, cap, need_local, len_local, peer_mailbox, n_shards, rank, __ldcg(d_epoch) + 1);
}

__device__ __forceinline__ void write_shard_record(const double *sum, const double *mnn, int64_t cap, int64_t need_local,
                                                   int64_t len_local, double *out) {
    emit_record(sum, mnn, cap, need_local, len_local, out, nullptr, 0, 0, 0);
}

__global__ void k_shard_record(const double *sum, const double *mnn, int64_t cap, int64_t need_local, int64_t len_local,
                               double *out) {
    write_shard_record(sum, mnn, cap, need_local, len_local, out);
}

// One block of the descent: the LEFT children of the next (up to) seven generations below `node`,
// 1+2+4+8+16+32+64 = 127 of them, fetched by the warp in one round trip.  Register r0 of lane p holds
// generations 1..5 in the order (generation j, pair q) -> p = 2^(j-1)-1+q (31 values); r1 holds generation 6
// (pair q at lane q); r2 / r3 hold generation 7 (pair q at lane q & 31 of register 2 + (q >> 5)).
struct DescentBlock {
    double r0, r1, r2, r3;
};
constexpr int DESCENT_GENS = 7;

__device__ __forceinline__ DescentBlock load_block(const double *v, const int64_t node, const int s, const int lane) {
    const int j = 32 - __clz(lane + 1);  // lane -> generation 1..5 (lane 31 unused)
    const int qg = lane + 1 - (1 << (j - 1));
    DescentBlock b;
    b.r0 = (j <= s && lane < 31) ? v[(node << j) + 2 * qg] : 0.0;
    b.r1 = s >= 6 ? v[(node << 6) + 2 * lane] : 0.0;
    b.r2 = s >= 7 ? v[(node << 7) + 2 * lane] : 0.0;
    b.r3 = s >= 7 ? v[(node << 7) + 2 * (lane + 32)] : 0.0;
    return b;
}

// Walks the s generations of one block (memory.py:122-129: left child if its sum exceeds the mass, else subtract
// and go right); returns the path inside the block.
__device__ __forceinline__ int walk_block(const DescentBlock &b, const int s, double &mass) {
    int m = 0;
#pragma unroll
    for (int g = 1; g <= DESCENT_GENS; ++g) {
        if (g <= s) {
            double lv;
            if (g <= 5) {
                lv = shfl_f64(b.r0, (1 << (g - 1)) - 1 + m);
            } else if (g == 6) {
                lv = shfl_f64(b.r1, m);
            } else {
                const double lo = shfl_f64(b.r2, m & 31), hi = shfl_f64(b.r3, m & 31);
                lv = (m >> 5) ? hi : lo;
            }
            if (lv > mass) {
                m = 2 * m;
            } else {
                mass = __dsub_rn(mass, lv);
                m = 2 * m + 1;
            }
        }
    }
    return m;
}

// The block below the root: the same for every query, so callers load it before they know the query's mass.
__device__ __forceinline__ DescentBlock root_block(const SampleArgs &a, const int lane) {
    return load_block(a.sum, 1, a.L < DESCENT_GENS ? a.L : DESCENT_GENS, lane);
}

// Value of the left child of generation g, pair q, out of a loaded block (every lane asks for the same q).
__device__ __forceinline__ double block_left(const DescentBlock &b, const int g, const int q) {
    if (g <= 5) return shfl_f64(b.r0, (1 << (g - 1)) - 1 + q);
    if (g == 6) return shfl_f64(b.r1, q);
    const double lo = shfl_f64(b.r2, q & 31), hi = shfl_f64(b.r3, q & 31);
    return (q >> 5) ? hi : lo;
}

// The local descent of one query (all 32 lanes) from the root of this tree, then the index and
// the importance weight.  `mass` is the prefix sum looked for inside this tree, `blk` = root_block().
// ceil(L / 7) dependent round trips, the first of them hoisted by the caller; the block that reaches
// the leaves also brings the RIGHT children of its last generation, so the sampled leaf's value (the
// weight's numerator) needs no round trip of its own.
__device__ __forceinline__ void descend_and_emit(const SampleArgs &a, const int64_t q, const int lane, double mass, DescentBlock blk,
                                                 const double root_sum, const double root_min, const int64_t idx_base) {
    const bool want_w = a.mnn && (a.out_w64 || a.out_w32);
    int64_t node = 1;
    int depth = 0;
    double leaf_v = 0.0;
#pragma unroll 1
    while (depth < a.L) {
        const int s = a.L - depth < DESCENT_GENS ? a.L - depth : DESCENT_GENS;
        const bool last = depth + s == a.L;
        if (depth) blk = load_block(a.sum, node, s, lane);
        double rr0 = 0.0, rr1 = 0.0;  // right children of generation s (the leaves' right halves)
        if (last && want_w) {
            if (lane < (1 << (s - 1))) rr0 = a.sum[(node << s) + 2 * lane + 1];
            if (s == 7) rr1 = a.sum[(node << 7) + 2 * (lane + 32) + 1];
        }
        const int m = walk_block(blk, s, mass);
        if (last && want_w) {
            const int pm = m >> 1;
            const double lv = block_left(blk, s, pm);
            const double r_lo = shfl_f64(rr0, pm & 31), r_hi = shfl_f64(rr1, pm & 31);
            leaf_v = (m & 1) ? ((pm >> 5) ? r_hi : r_lo) : lv;
        }
        node = (node << s) + m;
        depth += s;
    }
    const int64_t k = node - a.cap;
    if (a.tl && blockIdx.x == 0 && threadIdx.x == 0) tl_stamp(a.tl, 4, 21);  // descent done (first query of CTA 0)
    if (lane == 0) a.out_idx[q] = idx_base + k;
    if (want_w) {
        // memory.py:289-295.
        // w = (p N)^-beta / (p_min N)^-beta = 2^(-beta (log2(p N) - log2(p_min N))): the two logarithms side by side
        // on lanes 0 / 1 and one exp2, about a third of the dependent instructions of two pow() and a division
        // (agrees with the reference's float64 expression to ~1e-15 relative; the bar is 1e-5)
        const double n = (double)a.n_rows;
        const double num = lane == 1 ? root_min : leaf_v;
        const double lg = log2(__dmul_rn(__ddiv_rn(num, root_sum), n));
        const double lg_min = shfl_f64(lg, 1);
        if (lane == 0) {
            const double w = exp2(-a.beta * (lg - lg_min));
            if (a.out_w64) a.out_w64[q] = w;
            if (a.out_w32) a.out_w32[q] = (float)w;
        }
    }
}

// What every query of a launch shares — the block under the root, the two roots, the prefix total — is fetched by ONE
// warp per CTA and handed to the others through shared memory: with every warp of the grid fetching it, 512 warps queue
// up at the same ~50 L2 sectors before their first descent step.
struct SampleShared {
    double blk[4][32];  // DescentBlock of the root, by register and lane
    double root_sum, root_min, total;
};
__device__ __forceinline__ DescentBlock block_from_shared(const SampleShared &sh, const int lane) {
    DescentBlock b;
    b.r0 = sh.blk[0][lane];
    b.r1 = sh.blk[1][lane];
    b.r2 = sh.blk[2][lane];
    b.r3 = sh.blk[3][lane];
    return b;
}
__device__ __forceinline__ void block_to_shared(SampleShared &sh, const DescentBlock &b, const int lane) {
    sh.blk[0][lane] = b.r0;
    sh.blk[1][lane] = b.r1;
    sh.blk[2][lane] = b.r2;
    sh.blk[3][lane] = b.r3;
}

// Unsharded: one warp per query.
__global__ void __launch_bounds__(128) k_sample(SampleArgs a) {
    __shared__ SampleShared sh;
    const bool tl_on = a.tl && blockIdx.x == 0 && threadIdx.x == 0;
    if (tl_on) tl_stamp(a.tl, 0, 1);
    pdl_launch_dependents();
    pdl_wait();  // the tree is written by the kernel in front of this one
    if (tl_on) tl_stamp(a.tl, 1, 2);
    const int lane = threadIdx.x & 31;
    const int64_t q = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
    const double uq = (q < a.batch && lane == 0) ? a.u[q] : 0.0;  // requested beside the shared part
    if (threadIdx.x < 32) {
        const DescentBlock left = root_block(a, lane);  // in flight together with the loads of the prefix total
        const double rs = a.sum[1], rm = a.mnn ? a.mnn[1] : 0.0;
        const double total = a.direct_mass ? 0.0 : warp_prefix_total(a.sum, a.cap, a.n_rows - 1);  // memory.py:244
        block_to_shared(sh, left, lane);
        if (lane == 0) {
            sh.root_sum = rs;
            sh.root_min = rm;
            sh.total = total;
        }
    }
    __syncthreads();
    if (tl_on) tl_stamp(a.tl, 3, 20);  // shared part in
    if (q >= a.batch) return;  // whole warps take the same side
    const DescentBlock left = block_from_shared(sh, lane);
    const double root_sum = sh.root_sum;
    double mass = shfl_f64(uq, 0);
    if (a.direct_mass) {
        if (lane == 0 && !(mass >= 0.0 && mass <= __dadd_rn(root_sum, 1e-5))) atomicOr(a.status, (u64)PER_STATUS_BAD_PREFIXSUM);
    } else {
        const double seg = __ddiv_rn(sh.total, (double)a.batch);            // memory.py:245
        mass = __dadd_rn(__dmul_rn(mass, seg), __dmul_rn((double)q, seg));  // memory.py:247
    }
    descend_and_emit(a, q, lane, mass, left, root_sum, sh.root_min, 0);
    if (tl_on) tl_stamp(a.tl, 2, 3);
}

// Unsharded, large batches (thousands of queries): one THREAD per query.  A warp per query spends ~750 warp
// instructions on a query (shuffle-resolved walk, pow on two lanes); with enough queries to fill the machine the
// dependent loads of a per-thread walk hide behind each other and the instruction count per query drops 30-fold.
// THREE generations per round trip: below node n the left children the walk can meet are 2n, {4n, 4n+2} and
// {8n, 8n+2, 8n+4, 8n+6} — one 8-byte and three 32-byte loads whose addresses depend on n alone, four sectors instead of
// three, a third of the dependent round trips.  L mod 3 single generations come first (the top of the tree: L1 hits), so the
// last round ends at the leaves and its 32-byte loads have brought the sampled leaf itself (left or right child): the
// weight's numerator needs no load of its own.  Same comparisons and subtractions in the same order: same bits.
// p_total and the minimum's logarithm are computed once per CTA.
constexpr int SAMPLE_WIDE_THREADS = 128;  // measured at 2^20 leaves, 4 096 / 16 384 / 65 536 queries: 32, 64, 128, 256 threads -> 128 best (5.75 / 6.19 / 6.65 us)
__device__ __forceinline__ void ld_f64x4(const double *p, double &a, double &b, double &c, double &d) {
    asm volatile("ld.global.v4.f64 {%0, %1, %2, %3}, [%4];" : "=d"(a), "=d"(b), "=d"(c), "=d"(d) : "l"(p));
}
__device__ __forceinline__ bool walk_step(const double lv, double &mass) {  // memory.py:122-129; true: right child
    if (lv > mass) return false;
    mass = __dsub_rn(mass, lv);
    return true;
}
// The descent of one query by one thread (memory.py:122-129), three generations per round trip; returns the leaf's slot
// and leaves its value in leaf_v.
__device__ __forceinline__ int64_t thread_descent(const double *sum, const int64_t cap, const int L, double mass, double &leaf_v) {
    int64_t node = 1;
    leaf_v = sum[1];  // (a tree of one leaf)
    int d = 0;
    for (const int singles = L % 3; d < singles; ++d) {
        const double lv = sum[2 * node];
        const bool right = walk_step(lv, mass);
        if (d + 1 == L) leaf_v = right ? sum[2 * node + 1] : lv;  // (trees of 2 or 4 leaves)
        node = 2 * node + right;
    }
#pragma unroll 1
    for (; d < L; d += 3) {
        double g0, g1, g2, g3, h0, h1, h2, h3, h4, h5, h6, h7;
        const double c = sum[2 * node];
        ld_f64x4(sum + 4 * node, g0, g1, g2, g3);
        ld_f64x4(sum + 8 * node, h0, h1, h2, h3);
        ld_f64x4(sum + 8 * node + 4, h4, h5, h6, h7);
        const bool b1 = walk_step(c, mass);
        const bool b2 = walk_step(b1 ? g2 : g0, mass);
        const double hl = b1 ? (b2 ? h6 : h4) : (b2 ? h2 : h0);  // left child of the node two steps down
        const bool b3 = walk_step(hl, mass);
        if (b3) leaf_v = b1 ? (b2 ? h7 : h5) : (b2 ? h3 : h1);   // (used after the last round only)
        else leaf_v = hl;
        node = 8 * node + 4 * b1 + 2 * b2 + b3;
    }
    return node - cap;
}

__global__ void __launch_bounds__(SAMPLE_WIDE_THREADS) k_sample_wide(SampleArgs a) {
    __shared__ double s_seg, s_lg_min;
    pdl_launch_dependents();
    pdl_wait();
    const int64_t q = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    const bool want_w = a.mnn && (a.out_w64 || a.out_w32);
    const double uq = q < a.batch ? a.u[q] : 0.0;  // requested beside the shared part
    const double root_sum = a.sum[1];
    if (threadIdx.x < 32) {
        if (!a.direct_mass) {
            const double total = warp_prefix_total(a.sum, a.cap, a.n_rows - 1);  // memory.py:244
            if (threadIdx.x == 0) s_seg = __ddiv_rn(total, (double)a.batch);     // memory.py:245
        }
        if (threadIdx.x == 0 && want_w) s_lg_min = log2(__dmul_rn(__ddiv_rn(a.mnn[1], root_sum), (double)a.n_rows));
    }
    __syncthreads();
    if (q >= a.batch) return;
    double mass;
    if (a.direct_mass) {
        mass = uq;
        if (!(mass >= 0.0 && mass <= __dadd_rn(root_sum, 1e-5))) atomicOr(a.status, (u64)PER_STATUS_BAD_PREFIXSUM);
    } else {
        const double seg = s_seg;
        mass = __dadd_rn(__dmul_rn(uq, seg), __dmul_rn((double)q, seg));  // memory.py:247
    }
    double leaf_v;
    const int64_t k = thread_descent(a.sum, a.cap, a.L, mass, leaf_v);
    a.out_idx[q] = k;
    if (want_w) {  // memory.py:289-295
        const double w = exp2(-a.beta * (log2(__dmul_rn(__ddiv_rn(leaf_v, root_sum), (double)a.n_rows)) - s_lg_min));
        if (a.out_w64) a.out_w64[q] = w;
        if (a.out_w32) a.out_w32[q] = (float)w;
    }
}

// ---------------------------------------------------------------------------------------
// Sharded form.  The view of the virtual tree above the shard roots is built ONCE per CTA (not
// per query): wait for the records, top tree, p_total, segment width.  Stratum masses grow with
// the stratum number and the top descent is monotone in the mass, so the strata one shard owns
// are a contiguous range [lo, hi); warp 0 finds it with the exact ownership test (32 probes
// around the position the shard-root prefix predicts, one round trip; a 32-ary search otherwise)
// and only those strata are descended.  Everything else gets index -1.
// ---------------------------------------------------------------------------------------
struct ShardView {
    double top;       // lane i: node i of the top tree (lanes N..2N-1 the shard roots)
    double root_sum, root_min, seg;
};

__device__ __forceinline__ ShardView shard_view(const SampleArgs &a, const double *mail_recs, const int lane) {
    const int N = a.n_shards;
    // the records: the caller's all-gathered array, or what await_records decoded out of the mailbox
    const double *recs = a.use_mailbox ? mail_recs : a.recs;
    const int rec_stride = SHARD_REC;
    ShardView v;
    // Top of the virtual tree, one node per lane: lanes N..2N-1 hold the shard roots,
    // lanes 1..N-1 their pairwise sums built level by level like memory.py:82-87.
    double top = (lane >= N && lane < 2 * N) ? recs[(lane - N) * rec_stride] : 0.0;
    double mn = lane < N ? recs[lane * rec_stride + 1] : __longlong_as_double(0x7ff0000000000000LL);
    // p_total = sum over global leaves [0, need): shard-aligned blocks come from the top
    // tree (one per set bit of need / cap), the rest from the owning shard's record.
    const int64_t need = a.n_rows - 1;
    const int64_t need_hi = need >> a.L;  // whole shards covered
    const int64_t need_lo = need & (a.cap - 1);
    double acc = 0.0;
    bool first = true;
    if (need_lo) {
        acc = recs[need_hi * rec_stride + 2];
        first = false;
    }
    for (int w = N >> 1; w >= 1; w >>= 1) {
        const double l = shfl_f64(top, (2 * lane) & 31), r = shfl_f64(top, (2 * lane + 1) & 31);
        if (lane >= w && lane < 2 * w) top = __dadd_rn(l, r);
    }
    for (int o = 16; o; o >>= 1) mn = fmin(mn, shfl_f64(mn, lane ^ o));
#pragma unroll 1
    for (int b = 0; (int64_t(1) << b) <= need_hi; ++b) {
        if ((need_hi >> b) & 1) {
            const double t = shfl_f64(top, (int)((N >> b) + ((need_hi >> (b + 1)) << 1)));
            acc = first ? t : __dadd_rn(t, acc);
            first = false;
        }
    }
    v.top = top;
    v.root_sum = shfl_f64(top, 1);
    v.root_min = mn;
    v.seg = __ddiv_rn(acc, (double)a.batch);
    return v;
}

// Mass of stratum q (memory.py:247) and its descent through the top tree: returns the owning
// shard and leaves in `mass` the prefix sum inside that shard.  Every lane may ask for a different q.
__device__ __forceinline__ int shard_owner(const SampleArgs &a, const double top, const double seg, const int64_t q, double &mass) {
    const int N = a.n_shards;
    mass = __dadd_rn(__dmul_rn(a.u[q], seg), __dmul_rn((double)q, seg));
    int node = 1;
    while (node < N) {  // log2(N) rounds on every lane
        const double lv = shfl_f64(top, 2 * node);
        if (lv > mass) {
            node = 2 * node;
        } else {
            mass = __dsub_rn(mass, lv);
            node = 2 * node + 1;
        }
    }
    return node - N;
}

// First stratum whose owner is >= r (a.batch if there is none) by 32-ary search; all lanes return it.
__device__ __forceinline__ int64_t first_stratum_search(const SampleArgs &a, const ShardView &v, const int r, const int lane) {
    if (r <= 0) return 0;
    if (r >= a.n_shards) return a.batch;
    double mass;
    int64_t base = 0, n = a.batch;  // answer in [base, base + n]
    while (n > 0) {
        const int64_t step = (n + 31) >> 5;
        int64_t q = base + (lane + 1) * step - 1;  // last stratum of this lane's chunk
        if (q > base + n - 1) q = base + n - 1;
        const bool p = shard_owner(a, v.top, v.seg, q, mass) >= r;
        const unsigned m = __ballot_sync(0xffffffffu, p);
        if (m == 0) {
            base += n;
            break;
        }
        const int f = __ffs(m) - 1;
        const int64_t cb = base + f * step;
        int64_t ce = cb + step - 1;
        if (ce > base + n - 1) ce = base + n - 1;
        base = cb;  // strata cb .. ce-1 undecided, ce satisfies the predicate
        n = ce - cb;
    }
    return base;
}

// The strata [lo, hi) this shard owns.  One round trip in the expected case: the prefix of the shard
// roots below rank (lanes 0..15) and below rank + 1 (lanes 16..31), in units of the segment width,
// predicts each border to within a stratum; 16 strata around each prediction take the exact ownership
// test, and a border counts as found only if it is bracketed (a false probe below a true one).
__device__ __forceinline__ void owned_strata(const SampleArgs &a, const ShardView &v, const int lane, int64_t &lo, int64_t &hi) {
    const int N = a.n_shards;
    const int half = lane >> 4;            // 0: border below this shard, 1: border above it
    const int r = a.rank + half;           // the predicate of this half-warp: owner >= r
    double below = 0.0;                    // sum of the shard roots 0 .. r-1 (prediction only: any order will do)
    for (int sidx = 0; sidx < N; ++sidx) {
        const double root = shfl_f64(v.top, N + sidx);
        if (sidx < r) below += root;
    }
    const double guess = below / v.seg;
    const bool usable = guess >= 0.0 && guess <= (double)a.batch;  // false for NaN (seg == 0)
    const int64_t g0 = usable ? (int64_t)guess : 0;
    const int64_t q = g0 - 8 + (lane & 15);
    const int64_t qc = q < 0 ? 0 : (q >= a.batch ? a.batch - 1 : q);
    double mass;
    const bool own = shard_owner(a, v.top, v.seg, qc, mass) >= r;
    const bool p = r >= N ? q >= a.batch : (r <= 0 ? q >= 0 : (q >= a.batch || (q >= 0 && own)));
    const unsigned m = __ballot_sync(0xffffffffu, p);
    const unsigned ok = __ballot_sync(0xffffffffu, usable);
    const unsigned m_lo = m & 0xffffu, m_hi = m >> 16;
    const int64_t g_lo = __shfl_sync(0xffffffffu, g0, 0), g_hi = __shfl_sync(0xffffffffu, g0, 16);
    if ((ok & 1u) && m_lo != 0 && !(m_lo & 1u)) lo = g_lo - 8 + (__ffs(m_lo) - 1);
    else lo = first_stratum_search(a, v, a.rank, lane);
    if ((ok & 0x10000u) && m_hi != 0 && !(m_hi & 1u)) hi = g_hi - 8 + (__ffs(m_hi) - 1);
    else hi = first_stratum_search(a, v, a.rank + 1, lane);
}

// Learner-ward delivery of a sharded sample's scalars: the strata this shard owns (idx != -1) are copied into the
// learner GPU's full-batch index / weight arrays (peer-mapped: plain stores over NVLink).  Strata of other shards are
// left alone, their owners write them.
__global__ void k_shard_deliver(const int64_t *idx, const float *w32, int64_t n, int64_t *dst_idx, float *dst_w32) {
    for (int64_t q = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; q < n; q += (int64_t)gridDim.x * blockDim.x) {
        const int64_t k = idx[q];
        if (k == -1) continue;
        if (dst_idx) dst_idx[q] = k;
        if (dst_w32 && w32) dst_w32[q] = w32[q];
    }
}

constexpr int SAMPLE_THREADS = 128;  // one CTA per 4 owned strata: the descents spread over ~160 SMs (512-thread CTAs: +1 us per step, A/B in profiles/r02_notes.md)

// WIDE: one THREAD per owned stratum instead of one warp (shards that own thousands of strata, like k_sample_wide).
template <bool WIDE>
__global__ void __launch_bounds__(SAMPLE_THREADS) k_sample_sharded(SampleArgs a) {
    __shared__ double s_top[32];
    __shared__ SampleShared sh;                // the block under this shard's root: fetched by warp 0 for the whole CTA
    __shared__ double s_recs[16 * SHARD_REC];  // the records of this sample, decoded out of the mailbox by warp 0
    __shared__ double s_scal[3];
    __shared__ int64_t s_range[2];
    const bool tl_on = a.tl && blockIdx.x == 0 && threadIdx.x == 0;
    if (tl_on) tl_stamp(a.tl, 0, 1);
    pdl_launch_dependents();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    // With early publish the records of this sample come from update kernels (this rank's and the peers'), through
    // release / acquire flags: that part does not depend on the completion of the kernel in front, so warp 0 runs it
    // BEFORE the programmatic-launch wait, while that kernel (normally this rank's update, which releases its
    // dependents only after its own wait, i.e. after the previous sample has retired and advanced the epoch) is
    // still running.  Everything that reads the tree or the caller's inputs comes after the wait.
    const bool early = a.use_mailbox && a.skip_publish;
    if (!early) pdl_wait();
    DescentBlock left;
    if (warp == 0) {
        // only this warp reads the epoch counter, so that it can be advanced as soon as every CTA's warp 0 has it
        const u64 epoch = __ldcg(a.d_epoch) + 1;
        if (a.use_mailbox) {
            // exchange through peer memory: the first warp of the grid publishes this shard's record to every
            // rank (unless the update kernel in front has done so); warp 0 of every CTA waits for all ranks' records
            const u64 t0 = (blockIdx.x == 0 && lane == 0) ? global_timer_ns() : 0;
            if (blockIdx.x == 0 && !a.skip_publish)
                publish_record(a.sum, a.mnn, a.cap, a.rec_need, a.rec_len, a.peer_mailbox, a.n_shards, a.rank, epoch);
            await_records(a.mailbox, a.n_shards, epoch, a.status, s_recs);
            if (blockIdx.x == 0 && lane == 0 && a.d_wait_stats) {
                const u64 dt = global_timer_ns() - t0;
                atomicAdd(a.d_wait_stats, 1ull);
                atomicAdd(a.d_wait_stats + 1, dt);
                atomicMax(a.d_wait_stats + 2, dt);
            }
        }
        const ShardView v = shard_view(a, s_recs, lane);
        if (tl_on) tl_stamp(a.tl, 1, 10);  // records in, view built
        if (early) pdl_wait();
        if (tl_on) tl_stamp(a.tl, 2, 2);   // the kernel in front has retired
        left = root_block(a, lane);  // in flight while the owned strata are searched
        int64_t lo, hi;
        owned_strata(a, v, lane, lo, hi);
        block_to_shared(sh, left, lane);
        s_top[lane] = v.top;
        if (lane == 0) {
            s_scal[0] = v.root_sum;
            s_scal[1] = v.root_min;
            s_scal[2] = v.seg;
            s_range[0] = lo;
            s_range[1] = hi;
            if (blockIdx.x == 0) {  // tags for the gather / update kernels that follow (owned_range)
                a.d_range[0] = (epoch << 32) | (u64)(0xffffffffu - (unsigned)lo);
                a.d_range[1] = (epoch << 32) | (u64)hi;
            }
        }
    } else {
        if (early) pdl_wait();
    }
    if (tl_on) tl_stamp(a.tl, 3, 11);  // owned strata known
    __syncthreads();
    if (warp != 0) left = block_from_shared(sh, lane);  // the same for every query of the launch
    if (threadIdx.x == 32) {
        // warp 0 of this CTA has consumed the epoch: the last CTA to say so advances it (nothing in this launch
        // reads it again; the kernels behind this one see the new value)
        if (atomicAdd(a.d_done, 1u) == gridDim.x - 1) {
            *a.d_done = 0;
            __threadfence();
            atomicAdd(a.d_epoch, 1ull);
        }
    }
    const double top = s_top[lane], root_sum = s_scal[0], root_min = s_scal[1], seg = s_scal[2];
    const int64_t lo = s_range[0], hi = s_range[1];
    // strata of other shards
    for (int64_t q = blockIdx.x * (int64_t)SAMPLE_THREADS + threadIdx.x; q < a.batch; q += (int64_t)gridDim.x * SAMPLE_THREADS) {
        if (q < lo || q >= hi) {
            a.out_idx[q] = -1;
            if (a.out_w64) a.out_w64[q] = 0.0;
            if (a.out_w32) a.out_w32[q] = 0.0f;
        }
    }
    const int64_t idx_base = (int64_t)a.rank * a.cap;
    if (WIDE) {  // own strata, one thread per query: the top walk out of shared memory, then thread_descent()
        const bool want_w = a.mnn && (a.out_w64 || a.out_w32);
        const double lg_min = want_w ? log2(__dmul_rn(__ddiv_rn(root_min, root_sum), (double)a.n_rows)) : 0.0;
        for (int64_t q = lo + blockIdx.x * (int64_t)SAMPLE_THREADS + threadIdx.x; q < hi; q += (int64_t)gridDim.x * SAMPLE_THREADS) {
            double mass = __dadd_rn(__dmul_rn(a.u[q], seg), __dmul_rn((double)q, seg));  // memory.py:247
            for (int node = 1; node < a.n_shards;) {  // == shard_owner(): ends at this shard by construction of [lo, hi)
                const double lv = s_top[2 * node];
                if (lv > mass) {
                    node = 2 * node;
                } else {
                    mass = __dsub_rn(mass, lv);
                    node = 2 * node + 1;
                }
            }
            double leaf_v;
            const int64_t k = thread_descent(a.sum, a.cap, a.L, mass, leaf_v);
            a.out_idx[q] = idx_base + k;
            if (want_w) {  // memory.py:289-295, as in descend_and_emit()
                const double w = exp2(-a.beta * (log2(__dmul_rn(__ddiv_rn(leaf_v, root_sum), (double)a.n_rows)) - lg_min));
                if (a.out_w64) a.out_w64[q] = w;
                if (a.out_w32) a.out_w32[q] = (float)w;
            }
        }
        if (tl_on) tl_stamp(a.tl, 4, 3);
        return;
    }
    // own strata, one warp per query
    const int64_t warps = (int64_t)gridDim.x * (SAMPLE_THREADS / 32);
    for (int64_t q = lo + blockIdx.x * (SAMPLE_THREADS / 32) + warp; q < hi; q += warps) {
        double mass;
        shard_owner(a, top, seg, q, mass);  // == a.rank by construction of [lo, hi)
        descend_and_emit(a, q, lane, mass, left, root_sum, root_min, idx_base);
    }
    if (tl_on) tl_stamp(a.tl, 4, 3);
}

}  // namespace perb