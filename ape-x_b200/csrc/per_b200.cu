// per_b200.cu — C ABI (include/per_b200.h) over the sm_100a kernels.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -shared -Xcompiler -fPIC
#include <math.h>
#include <string.h>

#include <vector>

#include "common.cuh"
#include "gather_kernels.cuh"
#include "aql_kernels.cuh"
#include "nstep_kernels.cuh"
#include "td_kernels.cuh"
#include "tree_kernels.cuh"
#include "update_kernels.cuh"

using namespace perb;

// ------------------------------------------------------------------------------------------
// handles
// ------------------------------------------------------------------------------------------
struct seg_tree_s {
    int device;
    int kind;
    TreeRef t;        // exactly one of sum / mnn is set
    double *v;
    u64 *stamps;      // last-write-wins arbitration scratch, all zero between calls
    u64 *status;      // device word
    int sm_count;
};

struct per_buffer_s {
    int device;
    int64_t ring_size;
    double alpha;
    int64_t obs_bytes, obs_stride, extra_bytes, extra_stride;
    TreeRef t;
    u64 *stamps;
    // device scalars: word 0 status (u64), word 1 _max_priority (f64), word 2 two floats for K5,
    // word 3 sharded-sample epoch, word 4 retired-warp counter, words 5-6 owned-strata range tags
    // ... words 48 and up: the chunk counters of k_upd_dense_sorted (DENSE_MAX_GROUPS + 1 unsigned)
    u64 *scal;
    uint8_t *obs, *next, *extra;
    int64_t *act;
    float *rew, *done;
    float *td_scratch;       // out_td | out_prio when the caller passes none
    int64_t td_scratch_cap;  // rows
    int64_t len, cursor;
    int sm_count;
    GatherTune tune;
    int sharded;               // created by per_create_shard: samples leave the owned-strata range on the device
    u64 *timeline;             // per_tune("timeline", entries): kernel time stamps, read by per_timeline
    unsigned timeline_cap;     //   entries; launch number timeline_seq owns entries [seq * TL_SLOTS, +TL_SLOTS)
    unsigned timeline_seq;
    u64 *tl_next() {           //   first entry of the launch being issued, or null
        if (!timeline) return nullptr;
        const unsigned launches = timeline_cap / TL_SLOTS;
        u64 *p = timeline + (size_t)(timeline_seq % launches) * TL_SLOTS;
        ++timeline_seq;
        return p;
    }
    // host-buffer entry points (per_sample_host / per_update_host): device staging for one batch and a pinned
    // block for the small uploads; grown on demand, reused from call to call
    struct HostPath {
        int64_t cap;  // rows
        double *d_u, *d_w64;     // d_w64 .. d_done are the parts of d_small
        int64_t *d_idx, *d_act;
        float *d_rew, *d_done;
        uint8_t *d_small;        // [cap] float64 weights | [cap] int64 indices | [cap] int64 actions | [cap] float rewards | [cap] float dones
        uint8_t *h_small;        // pinned mirror of d_small: the per-row scalars of a sample come down in ONE copy
        uint8_t *d_obs, *d_next, *d_extra;
        uint8_t *d_upd;   // [cap] int64 indices | [cap] float64 leaves
        uint8_t *h_pin;   // pinned: [cap] float64 uniforms | [cap] int64 indices | [cap] float64 leaves
        cudaEvent_t up_done;  // the last upload out of h_pin
        bool have_event;
    } hp;
    int early_publish;         // per_tune("early_publish", 1): updates publish the record of the next sample (see header)
    bool rec_published;        //   ... and one has done so since the last sample
    int pdl;                   // per_tune("pdl", bits): programmatic dependent launch for 1 sampling, 2 gather, 4 update kernels
    int upd_cluster_max;       // per_tune("update_cluster_max", items): larger plain updates take the multi-launch form
    int upd_max_ctas;          // per_tune("update_ctas", n): cap on the cluster size of the one-launch update (0: none)
    int sorted_update;         // per_tune("sorted_update", 0/1): updates of the sample just drawn take the sorted fast path
    int dense_lmax;            // per_tune("dense_update_lmax", log2 leaves): bound updates of trees up to this depth take the dense form (0: never)
    int update_trace;          // per_tune("update_trace", 1): phase timestamps of the update kernel at scal[11..16]
    const int64_t *last_sample_idx;  // sharded replay: output buffer and size of the last per_sample_sharded; gather /
    int64_t last_sample_batch;       //   update calls that pass the same buffer visit only the strata this shard owns
    double *mailbox;           // sharded replay, peer-memory exchange: this rank's mailbox (may be null)
    double **d_peer_mailbox;   //   device array [n_shards] of all ranks' mailboxes
    double *rec_out;           // sharded replay: record refreshed by every tree-modifying call (may be null)
    int64_t global_len;        //   len(_storage) of the whole replay, as last told by the caller
    int shard_rank, n_shards;  // sharded replay: this buffer is slice shard_rank of n_shards (1: not sharded)
    int64_t idx_base;          // shard_rank * tree capacity
    u64 *status() const { return scal; }
    double *max_prio() const { return reinterpret_cast<double *>(scal + 1); }
    float *td_red() const { return reinterpret_cast<float *>(scal + 2); }
    u64 *xchg_epoch() const { return scal + 3; }                                   // exchange samples completed
    unsigned *xchg_done() const { return reinterpret_cast<unsigned *>(scal + 4); }  // retired warps of the running sample
};

namespace {

__global__ void k_fill_f64(double *p, int64_t n, double x) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) p[i] = x;
}

constexpr int SCAL_WORDS = 48 + (DENSE_MAX_GROUPS + 1 + 1) / 2;  // per_buffer::scal, in u64 words
constexpr int64_t SAMPLE_WIDE_MIN = 2048;  // per_sample_indices: batches from here on use k_sample_wide

bool is_pow2(int64_t x) { return x > 0 && (x & (x - 1)) == 0; }

int ceil_log2(int64_t x) {
    int l = 0;
    while ((int64_t(1) << l) < x) ++l;
    return l;
}

int alloc_tree(double **out, int64_t cap, double neutral, cudaStream_t s) {
    PER_CUDA(cudaMalloc(out, sizeof(double) * 2 * cap));
    if (neutral == 0.0) {
        PER_CUDA(cudaMemsetAsync(*out, 0, sizeof(double) * 2 * cap, s));
    } else {
        k_fill_f64<<<(int)((2 * cap + 255) / 256 > 4096 ? 4096 : (2 * cap + 255) / 256), 256, 0, s>>>(*out, 2 * cap, neutral);
        PER_CUDA(cudaGetLastError());
    }
    return PER_OK;
}

int grid_for(int64_t n, int threads, int sm_count) {
    int64_t g = (n + threads - 1) / threads;
    int64_t cap = (int64_t)sm_count * 8;
    if (g > cap) g = cap;
    if (g < 1) g = 1;
    return (int)g;
}

// The dense form (k_upd_dense_sorted) serves updates of the ascending sample just drawn that are too large for the
// per-CTA climb of k_update_cluster (measured, tools/tree_sweep.py and profiles/r02_notes.md): more than 1 024 items on
// trees up to 2^20 leaves, more than 2 048 on deeper ones, as long as no CTA gets more than ~7 dirty chunks.  Smaller
// batches stay on the climb: in isolation the dense form is faster for them too on trees up to 2^18 leaves, but its one
// CTA per 1 024 leaves occupies the whole GPU, and inside a sample / gather / update step that costs the neighbouring
// kernels more than it saves (2^20 leaves, 512 items: step 28.3 -> 29.8 us).  per_tune("dense_update_lmax", L) sends
// every bound update of trees up to 2^L leaves there.
bool dense_update_ok(const UpdArgs &a) {
    if (!(a.dense_counter && a.sorted_idx && a.sorted_fast && a.t.sum && a.t.mnn && a.t.L >= DENSE_L_MIN && a.t.L <= DENSE_L_MAX)) return false;
    if (a.t.L <= a.dense_lmax) return true;
    const int64_t n = (a.range && a.n_shards > 1) ? a.n / a.n_shards : a.n;  // a shard applies its own strata only
    if (n <= (a.t.L <= 20 ? 1024 : 2048)) return false;  // (up to 2 048 items a 16-CTA cluster keeps the per-CTA climb)
    // a CTA lifts its dirty chunks one after the other (~4 us each): past ~7 of them the multi-launch form's grid-wide
    // passes win (measured at 2^24 / 2^26 leaves: 8 192 items 37 / 38 us against 40 / 55, 16 384 items 56 / 60 against
    // 43 / 60; 1 184 = the resident CTAs of a B200 at 64 registers)
    const int64_t chunks = a.t.cap >> DENSE_CHUNK_LOG2;
    return (n < chunks ? n : chunks) <= 7 * 1184;
}

static int launch_update_cluster(const UpdArgs &a, const TdArgs *td, int skip_update, cudaStream_t s, bool pdl) {
    TdArgs none;
    memset(&none, 0, sizeof none);
    // one launch: a cluster of CTAs (at most 16 where the device offers the non-portable size, else 8)
    static int max_ctas = 0;
    if (!max_ctas) {
        PER_CUDA(cudaFuncSetAttribute(k_update_cluster, cudaFuncAttributeMaxDynamicSharedMemorySize, UPD_SMEM_BYTES));
        max_ctas = 8;
        if (cudaFuncSetAttribute(k_update_cluster, cudaFuncAttributeNonPortableClusterSizeAllowed, 1) == cudaSuccess) {
            cudaLaunchConfig_t probe;
            memset(&probe, 0, sizeof probe);
            probe.gridDim = dim3(UPD_MAX_CTAS);
            probe.blockDim = dim3(UPD_THREADS);
            probe.dynamicSmemBytes = UPD_SMEM_BYTES;
            cudaLaunchAttribute pa[1];
            pa[0].id = cudaLaunchAttributeClusterDimension;
            pa[0].val.clusterDim.x = UPD_MAX_CTAS;
            pa[0].val.clusterDim.y = 1;
            pa[0].val.clusterDim.z = 1;
            probe.attrs = pa;
            probe.numAttrs = 1;
            int nclusters = 0;
            if (cudaOccupancyMaxActiveClusters(&nclusters, k_update_cluster, &probe) == cudaSuccess && nclusters >= 1) max_ctas = UPD_MAX_CTAS;
        }
        cudaGetLastError();
    }
    // 8 CTAs by default: A/B on one box, 16 CTAs cost +0.3 us per step at 512 items (the phases are latency chains, not
    // sector throughput; the larger cluster's barriers cost more than the halved per-SM work saves).  16 where it keeps a
    // sorted batch of 1 025 .. 2 048 items on the per-CTA climb (<= 128 items per CTA), or when asked for (per_tune).
    int limit = 8;
    if (a.max_ctas > 0) limit = a.max_ctas < max_ctas ? a.max_ctas : max_ctas;
    else if (a.sorted_idx && a.sorted_fast && a.n > 8 * SORT_SLICE_MAX && a.n <= 16 * SORT_SLICE_MAX) limit = max_ctas;
    int ncta = (int)((a.n + UPD_THREADS / LIFT_LANES - 1) / (UPD_THREADS / LIFT_LANES));  // 64 items per CTA and round
    if (ncta < 1) ncta = 1;
    if (ncta > limit) ncta = limit;
    cudaLaunchConfig_t cfg;
    memset(&cfg, 0, sizeof cfg);
    cfg.gridDim = dim3(ncta);
    cfg.blockDim = dim3(UPD_THREADS);
    cfg.dynamicSmemBytes = UPD_SMEM_BYTES;
    cfg.stream = s;
    cudaLaunchAttribute attr[2];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = ncta;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;  // the kernel starts with pdl_wait()
    attr[1].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl ? 2 : 1;
    PER_CUDA(cudaLaunchKernelEx(&cfg, k_update_cluster, a, td ? *td : none, td ? 1 : 0, skip_update));
    return PER_OK;
}

// Batched leaf update + re-reduction (K1).  `td` non-null puts K5 in front: fused into the one-launch cluster form, or as
// a TD launch of that kernel followed by the dense form.
int run_update(UpdArgs &a, const TdArgs *td, int skip_update, int sm_count, cudaStream_t s, bool pdl = false) {
    if (!skip_update && dense_update_ok(a) && (!td || td->batch <= UPD_CLUSTER_MAX_ITEMS)) {
        static int slots = 0;  // resident CTAs of the device
        if (!slots) {
            int per_sm = 0;
            PER_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_upd_dense_sorted, DENSE_THREADS, 0));
            slots = (per_sm > 0 ? per_sm : 1) * sm_count;
        }
        UpdArgs d = a;
        if (td) {
            // the TD launch writes td->out_prio; the dense launch behind it searches its item runs meanwhile
            UpdArgs t = a;
            t.late_release = 1;
            t.n = td->batch;
            t.sorted_idx = 0;  // (the cluster size then follows the generic rule: the loss is summed in the same grouping as the fused form's)
            const int rc = launch_update_cluster(t, td, 1, s, pdl);
            if (rc) return rc;
            d.src = td->out_prio;
            d.src_kind = LEAF_PRIO_F32;
            d.early_search = 1;
        }
        const int64_t chunks = a.t.cap >> DENSE_CHUNK_LOG2;
        PER_CUDA(launch_pdl(k_upd_dense_sorted, dim3((unsigned)(chunks < slots ? chunks : slots)), dim3(DENSE_THREADS), 0, s, pdl || td, d,
                            a.dense_counter));
        return PER_OK;
    }
    if (a.n <= (td || a.range || a.cluster_max <= 0 ? UPD_CLUSTER_MAX_ITEMS : a.cluster_max))
        return launch_update_cluster(a, td, skip_update, s, pdl);
    if (td) return fail(PER_ERR_ARG, "internal: fused TD needs the one-launch update form");
    const int g = grid_for(a.n, 256, sm_count);
    if (a.stamps) k_upd_stamp<<<g, 256, 0, s>>>(a);
    k_upd_leaf<<<g, 256, 0, s>>>(a);
    const int L = a.t.L;
    const int top = L < TOP_LEVELS ? L : TOP_LEVELS;
    int want = ceil_log2(a.n) + 2;  // below this depth a dense rebuild touches fewer nodes than n sparse lifts
    if (want > L) want = L;
    if (want < top) want = top;
    const int d_dense = L - LIFT * ((L - want) / LIFT);  // sparse rounds lift LIFT generations each
    for (int done = 0; L - done > d_dense; done += LIFT) k_upd_sparse<<<g, 256, 0, s>>>(a, done);
    int d = d_dense;
    while (d > top) {
        int k = d - top < TOP_LEVELS ? d - top : TOP_LEVELS;
        k_dense_mid<<<(unsigned)(int64_t(1) << (d - k)), 256, 0, s>>>(a.t, d, k);
        d -= k;
    }
    k_dense_top<<<1, 1024, 0, s>>>(a.t, top, a.rec_out, a.rec_need, a.rec_len);
    PER_CUDA(cudaGetLastError());
    return PER_OK;
}

int64_t shard_need(int64_t global_len, int64_t idx_base, int64_t cap) {
    int64_t need = (global_len - 1) - idx_base;  // leaves of this slice inside the global prefix [0, len-1)
    if (need < 0) need = 0;
    if (need > cap) need = cap;
    return need;
}

int read_status(u64 *d_status, int64_t *h_out, cudaStream_t s) {
    u64 w = 0;
    PER_CUDA(cudaMemcpyAsync(&w, d_status, sizeof w, cudaMemcpyDeviceToHost, s));
    PER_CUDA(cudaMemsetAsync(d_status, 0, sizeof w, s));
    PER_CUDA(cudaStreamSynchronize(s));
    if (h_out) *h_out = (int64_t)w;
    return PER_OK;
}

}  // namespace

extern "C" {

const char *per_last_error(void) { return err_slot().c_str(); }
int per_destroy(per_buffer *b);
int seg_destroy(seg_tree *t);
int per_version(void) { return 100; }

// ---------------------------------------------------------------------------------- trees
int seg_create(seg_tree **out, int device, int64_t capacity, int kind) {
    if (!out) return fail(PER_ERR_ARG, "seg_create: null out");
    if (!is_pow2(capacity)) return fail(PER_ERR_ASSERT, "capacity must be positive and a power of 2.");
    if (kind != PER_TREE_SUM && kind != PER_TREE_MIN) return fail(PER_ERR_ARG, "seg_create: bad kind %d", kind);
    DeviceGuard g(device);
    if (!g.ok) return fail(PER_ERR_CUDA, "seg_create: cannot select device %d", device);
    seg_tree *t = new seg_tree();
    memset(t, 0, sizeof *t);
    t->device = device;
    t->kind = kind;
    cudaDeviceGetAttribute(&t->sm_count, cudaDevAttrMultiProcessorCount, device);
    int rc = alloc_tree(&t->v, capacity, kind == PER_TREE_SUM ? 0.0 : (double)INFINITY, 0);
    if (rc) { delete t; return rc; }
    t->t.cap = capacity;
    t->t.L = ilog2_exact(capacity);
    (kind == PER_TREE_SUM ? t->t.sum : t->t.mnn) = t->v;
    cudaError_t e = cudaMalloc(&t->stamps, sizeof(u64) * capacity);
    if (e == cudaSuccess) e = cudaMemsetAsync(t->stamps, 0, sizeof(u64) * capacity, 0);
    if (e == cudaSuccess) e = cudaMalloc(&t->status, sizeof(u64));
    if (e == cudaSuccess) e = cudaMemsetAsync(t->status, 0, sizeof(u64), 0);
    if (e == cudaSuccess) e = cudaStreamSynchronize(0);
    if (e != cudaSuccess) {  // nothing of a half-built tree stays behind
        cudaGetLastError();
        seg_destroy(t);
        return fail(e == cudaErrorMemoryAllocation ? PER_ERR_NOMEM : PER_ERR_CUDA, "seg_create: %s", cudaGetErrorString(e));
    }
    *out = t;
    return PER_OK;
}

int seg_destroy(seg_tree *t) {
    if (!t) return PER_OK;
    DeviceGuard g(t->device);
    cudaFree(t->v);
    cudaFree(t->stamps);
    cudaFree(t->status);
    delete t;
    return PER_OK;
}

int64_t seg_capacity(const seg_tree *t) { return t ? t->t.cap : -1; }

int seg_set(seg_tree *t, const int64_t *d_idx, const double *d_val, int64_t n, void *stream) {
    if (!t || n < 0 || (n && (!d_idx || !d_val))) return fail(PER_ERR_ARG, "seg_set: bad argument");
    if (n == 0) return PER_OK;
    DeviceGuard g(t->device);
    cudaStream_t s = (cudaStream_t)stream;
    UpdArgs a;
    memset(&a, 0, sizeof a);
    a.t = t->t;
    a.idx = d_idx;
    a.src = d_val;
    a.src_kind = LEAF_GIVEN;
    a.alpha = 1.0;
    a.n = n;
    a.len_limit = t->t.cap;
    a.stamps = t->stamps;
    a.status = t->status;
    return run_update(a, nullptr, 0, t->sm_count, s);
}

int seg_reduce(seg_tree *t, int64_t start, int64_t end, double *d_out, void *stream) {
    if (!t || !d_out) return fail(PER_ERR_ARG, "seg_reduce: bad argument");
    if (end < 0) end += t->t.cap;  // memory.py:71-72
    if (start < 0 || end > t->t.cap || end - 1 < start)
        return fail(PER_ERR_ASSERT, "seg_reduce: empty or out-of-range [%lld, %lld)", (long long)start, (long long)end);
    DeviceGuard g(t->device);
    k_range_reduce<<<1, 32, 0, (cudaStream_t)stream>>>(t->v, t->t.cap, start, end - 1, t->kind == PER_TREE_SUM, d_out);
    PER_CUDA(cudaGetLastError());
    return PER_OK;
}

int seg_find_prefixsum(seg_tree *t, const double *d_mass, int64_t n, int64_t *d_out_idx, void *stream) {
    if (!t || n < 0 || (n && (!d_mass || !d_out_idx))) return fail(PER_ERR_ARG, "seg_find_prefixsum: bad argument");
    if (t->kind != PER_TREE_SUM) return fail(PER_ERR_ARG, "seg_find_prefixsum: not a sum tree");
    if (n == 0) return PER_OK;
    DeviceGuard g(t->device);
    SampleArgs a;
    memset(&a, 0, sizeof a);
    a.sum = t->v;
    a.cap = t->t.cap;
    a.L = t->t.L;
    a.u = d_mass;
    a.direct_mass = 1;
    a.batch = n;
    a.out_idx = d_out_idx;
    a.status = t->status;
    PER_CUDA(launch_pdl(k_sample, dim3((unsigned)((n * 32 + 127) / 128)), dim3(128), 0, (cudaStream_t)stream, false, a));
    PER_CUDA(cudaGetLastError());
    return PER_OK;
}

int seg_read(seg_tree *t, double *d_out, void *stream) {
    if (!t || !d_out) return fail(PER_ERR_ARG, "seg_read: bad argument");
    DeviceGuard g(t->device);
    PER_CUDA(cudaMemcpyAsync(d_out, t->v, sizeof(double) * 2 * t->t.cap, cudaMemcpyDeviceToDevice, (cudaStream_t)stream));
    return PER_OK;
}

static int read_leaf(const double *nodes, int64_t cap, int64_t idx, double *h_out, cudaStream_t s) {
    if (idx < 0 || idx >= cap) return fail(PER_ERR_ASSERT, "0 <= idx < capacity");  // memory.py:90
    PER_CUDA(cudaMemcpyAsync(h_out, nodes + cap + idx, sizeof(double), cudaMemcpyDeviceToHost, s));
    PER_CUDA(cudaStreamSynchronize(s));
    return PER_OK;
}

int seg_get(seg_tree *t, int64_t idx, double *h_out, void *stream) {
    if (!t || !h_out) return fail(PER_ERR_ARG, "seg_get: bad argument");
    DeviceGuard g(t->device);
    return read_leaf(t->v, t->t.cap, idx, h_out, (cudaStream_t)stream);
}

int seg_status(seg_tree *t, int64_t *h_out, void *stream) {
    if (!t) return fail(PER_ERR_ARG, "seg_status: null handle");
    DeviceGuard g(t->device);
    return read_status(t->status, h_out, (cudaStream_t)stream);
}

// ------------------------------------------------------------------------------ replay buffer
static int create_impl(per_buffer **out, int device, int64_t ring_size, int64_t tree_capacity, double alpha, int64_t obs_bytes,
                       int64_t extra_bytes, int rank, int n_shards) {
    if (!out) return fail(PER_ERR_ARG, "per_create: null out");
    if (!(alpha >= 0)) return fail(PER_ERR_ASSERT, "alpha >= 0");
    if (ring_size < (n_shards > 1 ? 0 : 1) || obs_bytes < 0 || extra_bytes < 0) return fail(PER_ERR_ARG, "per_create: bad size");
    DeviceGuard g(device);
    if (!g.ok) return fail(PER_ERR_CUDA, "per_create: cannot select device %d", device);
    per_buffer *b = new per_buffer();
    memset(b, 0, sizeof *b);
    b->device = device;
    b->pdl = 3;
    b->sorted_update = 1;
    b->dense_lmax = 0;
    b->ring_size = ring_size;
    b->alpha = alpha;
    b->obs_bytes = obs_bytes;
    b->obs_stride = round_up(obs_bytes, 16);
    b->extra_bytes = extra_bytes;
    b->extra_stride = round_up(extra_bytes, 16);
    cudaDeviceGetAttribute(&b->sm_count, cudaDevAttrMultiProcessorCount, device);
    int64_t cap = 1;
    while (cap < ring_size) cap *= 2;  // memory.py:227-229
    if (tree_capacity > 0) cap = tree_capacity;
    b->shard_rank = rank;
    b->n_shards = n_shards;
    b->idx_base = (int64_t)rank * cap;
    b->t.cap = cap;
    b->t.L = ilog2_exact(cap);
    int rc = alloc_tree(&b->t.sum, cap, 0.0, 0);
    if (!rc) rc = alloc_tree(&b->t.mnn, cap, (double)INFINITY, 0);
    if (rc) { per_destroy(b); return rc; }
#define TRY_ALLOC(ptr, bytes)                                                                      \
    do {                                                                                           \
        if ((bytes) > 0 && cudaMalloc(&(ptr), (bytes)) != cudaSuccess) {                           \
            cudaGetLastError();                                                                    \
            per_destroy(b);                                                                        \
            return fail(PER_ERR_NOMEM, "per_create: cudaMalloc of %lld bytes failed", (long long)(bytes)); \
        }                                                                                          \
    } while (0)
    TRY_ALLOC(b->stamps, sizeof(u64) * cap);
    TRY_ALLOC(b->scal, sizeof(u64) * SCAL_WORDS);
    TRY_ALLOC(b->obs, b->obs_stride * ring_size);
    TRY_ALLOC(b->next, b->obs_stride * ring_size);
    TRY_ALLOC(b->extra, b->extra_stride * ring_size);
    TRY_ALLOC(b->act, sizeof(int64_t) * ring_size);
    TRY_ALLOC(b->rew, sizeof(float) * ring_size);
    TRY_ALLOC(b->done, sizeof(float) * ring_size);
    b->td_scratch_cap = UPD_CLUSTER_MAX_ITEMS;
    TRY_ALLOC(b->td_scratch, sizeof(float) * 2 * b->td_scratch_cap);
#undef TRY_ALLOC
    PER_CUDA(cudaMemsetAsync(b->stamps, 0, sizeof(u64) * cap, 0));
    PER_CUDA(cudaMemsetAsync(b->scal, 0, sizeof(u64) * SCAL_WORDS, 0));
    const double one = 1.0;  // _max_priority = 1.0 (memory.py:233)
    PER_CUDA(cudaMemcpyAsync(b->max_prio(), &one, sizeof one, cudaMemcpyHostToDevice, 0));
    PER_CUDA(cudaFuncSetAttribute(k_gather_tma, cudaFuncAttributeMaxDynamicSharedMemorySize, GATHER_SMEM_MAX));
    PER_CUDA(cudaFuncSetAttribute(k_gather_cast, cudaFuncAttributeMaxDynamicSharedMemorySize, GATHER_SMEM_MAX));
    b->tune = GatherTune();
    PER_CUDA(cudaStreamSynchronize(0));
    *out = b;
    return PER_OK;
}

int per_create(per_buffer **out, int device, int64_t ring_size, double alpha, int64_t obs_bytes, int64_t extra_bytes) {
    return create_impl(out, device, ring_size, 0, alpha, obs_bytes, extra_bytes, 0, 1);
}

int per_create_shard(per_buffer **out, int device, int64_t global_size, int rank, int n_shards, double alpha,
                     int64_t obs_bytes, int64_t extra_bytes) {
    if (n_shards < 1 || n_shards > 16 || (n_shards & (n_shards - 1)) || rank < 0 || rank >= n_shards)
        return fail(PER_ERR_ARG, "per_create_shard: n_shards must be a power of two <= 16 and 0 <= rank < n_shards");
    if (global_size < 2) return fail(PER_ERR_ARG, "per_create_shard: bad size");
    int64_t cap_g = 1;
    while (cap_g < global_size) cap_g *= 2;  // memory.py:227-229 on the whole replay
    if (cap_g < n_shards) return fail(PER_ERR_ARG, "per_create_shard: fewer leaves than shards");
    const int64_t cap_l = cap_g / n_shards;
    int64_t ring_l = global_size - (int64_t)rank * cap_l;  // ring slots that fall into this slice
    if (ring_l < 0) ring_l = 0;
    if (ring_l > cap_l) ring_l = cap_l;
    int rc = create_impl(out, device, ring_l, cap_l, alpha, obs_bytes, extra_bytes, rank, n_shards);
    if (!rc) {
        (*out)->sharded = 1;
        if (n_shards > 1) (*out)->pdl = 7;  // sharded steps gain from chaining the update as well (A/B on 2 GPUs: 37.5 -> 36.8 us per step)
    }
    return rc;
}


static void free_host_path(per_buffer *b) {
    per_buffer::HostPath &h = b->hp;
    cudaFree(h.d_u); cudaFree(h.d_small);
    cudaFree(h.d_obs); cudaFree(h.d_next); cudaFree(h.d_extra); cudaFree(h.d_upd);
    if (h.h_pin) cudaFreeHost(h.h_pin);
    if (h.h_small) cudaFreeHost(h.h_small);
    if (h.have_event) cudaEventDestroy(h.up_done);
    memset(&h, 0, sizeof h);
}

// Staging for `rows` rows (device) and the pinned upload block.  Called with the handle's device current.
static int ensure_host_path(per_buffer *b, int64_t rows, cudaStream_t s) {
    per_buffer::HostPath &h = b->hp;
    if (rows <= h.cap) return PER_OK;
    PER_CUDA(cudaStreamSynchronize(s));
    free_host_path(b);
    int64_t cap = 512;
    while (cap < rows) cap *= 2;
    cudaError_t e = cudaSuccess;
    auto grab = [&e](void **p, size_t bytes) {
        if (e == cudaSuccess && bytes) e = cudaMalloc(p, bytes);
    };
    grab((void **)&h.d_u, sizeof(double) * cap);
    grab((void **)&h.d_small, 32 * (size_t)cap);
    grab((void **)&h.d_obs, (size_t)b->obs_bytes * cap);
    grab((void **)&h.d_next, (size_t)b->obs_bytes * cap);
    grab((void **)&h.d_extra, (size_t)b->extra_bytes * cap);
    grab((void **)&h.d_upd, 16 * (size_t)cap);
    if (e == cudaSuccess) e = cudaHostAlloc(&h.h_pin, 24 * (size_t)cap, cudaHostAllocDefault);
    if (e == cudaSuccess) e = cudaHostAlloc(&h.h_small, 32 * (size_t)cap, cudaHostAllocDefault);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&h.up_done, cudaEventDisableTiming);
    if (e != cudaSuccess) {  // all or nothing: a later, smaller request starts from a clean slate
        cudaGetLastError();
        free_host_path(b);
        return fail(e == cudaErrorMemoryAllocation ? PER_ERR_NOMEM : PER_ERR_CUDA, "host-buffer staging for %lld rows: %s",
                    (long long)cap, cudaGetErrorString(e));
    }
    h.d_w64 = reinterpret_cast<double *>(h.d_small);
    h.d_idx = reinterpret_cast<int64_t *>(h.d_small + 8 * (size_t)cap);
    h.d_act = reinterpret_cast<int64_t *>(h.d_small + 16 * (size_t)cap);
    h.d_rew = reinterpret_cast<float *>(h.d_small + 24 * (size_t)cap);
    h.d_done = reinterpret_cast<float *>(h.d_small + 28 * (size_t)cap);
    h.have_event = true;
    h.cap = cap;
    return PER_OK;
}

int per_destroy(per_buffer *b) {
    if (!b) return PER_OK;
    DeviceGuard g(b->device);
    cudaFree(b->t.sum);
    cudaFree(b->t.mnn);
    cudaFree(b->stamps);
    cudaFree(b->scal);
    cudaFree(b->timeline);
    free_host_path(b);
    cudaFree(b->obs);
    cudaFree(b->next);
    cudaFree(b->extra);
    cudaFree(b->act);
    cudaFree(b->rew);
    cudaFree(b->done);
    cudaFree(b->td_scratch);
    cudaFree(b->d_peer_mailbox);
    delete b;
    return PER_OK;
}

int64_t per_capacity(const per_buffer *b) { return b ? b->t.cap : -1; }
int64_t per_len(const per_buffer *b) { return b ? b->len : -1; }
int64_t per_cursor(const per_buffer *b) { return b ? b->cursor : -1; }

int per_set_cursor(per_buffer *b, int64_t cursor, double max_priority, void *stream) {
    if (!b || cursor < 0 || cursor >= (b->ring_size > 0 ? b->ring_size : 1)) return fail(PER_ERR_ARG, "per_set_cursor: bad cursor");
    DeviceGuard g(b->device);
    b->cursor = cursor;
    if (max_priority > 0.0)
        PER_CUDA(cudaMemcpyAsync(b->max_prio(), &max_priority, sizeof(double), cudaMemcpyHostToDevice, (cudaStream_t)stream));
    PER_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
    return PER_OK;
}

int per_max_priority(per_buffer *b, double *h_out, void *stream) {
    if (!b || !h_out) return fail(PER_ERR_ARG, "per_max_priority: bad argument");
    DeviceGuard g(b->device);
    PER_CUDA(cudaMemcpyAsync(h_out, b->max_prio(), sizeof(double), cudaMemcpyDeviceToHost, (cudaStream_t)stream));
    PER_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
    return PER_OK;
}

int per_status(per_buffer *b, int64_t *h_out, void *stream) {
    if (!b) return fail(PER_ERR_ARG, "per_status: null handle");
    DeviceGuard g(b->device);
    return read_status(b->status(), h_out, (cudaStream_t)stream);
}

int per_add_batch(per_buffer *b, const void *d_obs, const void *d_next, const int64_t *d_act, const float *d_rew,
                  const float *d_done, const void *d_extra, const double *d_prio, const double *d_leaf,
                  double batch_max_priority, int64_t n, void *stream) {
    if (!b || n < 0) return fail(PER_ERR_ARG, "per_add_batch: bad argument");
    if (n > b->ring_size) return fail(PER_ERR_ARG, "per_add_batch: n=%lld exceeds the ring size %lld", (long long)n, (long long)b->ring_size);
    if (n == 0) return PER_OK;
    if (b->rec_published) return fail(PER_ERR_ARG, "early publish: the record of the next sample is already in the peers' mailboxes; sample before adding");
    DeviceGuard g(b->device);
    cudaStream_t s = (cudaStream_t)stream;
    AddArgs r;
    memset(&r, 0, sizeof r);
    r.src[0] = (const uint8_t *)d_obs;   r.dst[0] = b->obs;   r.row_bytes[0] = b->obs_bytes;   r.dst_stride[0] = b->obs_stride;
    r.src[1] = (const uint8_t *)d_next;  r.dst[1] = b->next;  r.row_bytes[1] = b->obs_bytes;   r.dst_stride[1] = b->obs_stride;
    r.src[2] = (const uint8_t *)d_extra; r.dst[2] = b->extra; r.row_bytes[2] = b->extra_bytes; r.dst_stride[2] = b->extra_stride;
    r.act = d_act; r.rew = d_rew; r.done = d_done;
    r.ring_act = b->act; r.ring_rew = b->rew; r.ring_done = b->done;
    r.start = b->cursor; r.ring_size = b->ring_size; r.n = n;
    int64_t work = n * ((b->obs_bytes * 2 + b->extra_bytes) / 16 + 1);
    k_ring_add<<<grid_for(work, 256, b->sm_count * 2), 256, 0, s>>>(r);
    PER_CUDA(cudaGetLastError());

    UpdArgs a;
    memset(&a, 0, sizeof a);
    a.t = b->t;
    a.idx = nullptr;
    a.ring_start = b->cursor;
    a.ring_size = b->ring_size;
    a.src = d_leaf ? d_leaf : d_prio;
    a.src_kind = d_leaf ? LEAF_GIVEN : (d_prio ? LEAF_PRIO_F64 : LEAF_MAXPRIO);
    a.batch_max = batch_max_priority;
    a.alpha = b->alpha;
    a.n = n;
    a.len_limit = b->ring_size;
    a.stamps = nullptr;  // ring slots of one batch are distinct
    a.status = b->status();
    a.max_prio = b->max_prio();
    int rc = run_update(a, nullptr, 0, b->sm_count, s);
    if (rc) return rc;
    b->last_sample_idx = nullptr;                                      // rows changed: no sample is bound any more
    b->cursor = (b->cursor + n) % b->ring_size;                        // memory.py:169
    b->len = b->len + n > b->ring_size ? b->ring_size : b->len + n;    // memory.py:165-168
    return PER_OK;
}

int per_sample_indices(per_buffer *b, const double *d_u, int64_t batch, double beta, int64_t *d_out_idx,
                       double *d_out_w64, float *d_out_w32, void *stream) {
    if (!b || batch <= 0 || !d_u || !d_out_idx) return fail(PER_ERR_ARG, "per_sample_indices: bad argument");
    if (!(beta > 0)) return fail(PER_ERR_ASSERT, "beta > 0");
    if (b->len < 2) return fail(PER_ERR_ASSERT, "sample needs at least 2 stored transitions (the reference's range sum over [0, len-2] does not terminate for len == 1)");
    DeviceGuard g(b->device);
    SampleArgs a;
    memset(&a, 0, sizeof a);
    a.sum = b->t.sum;
    a.mnn = b->t.mnn;
    a.cap = b->t.cap;
    a.L = b->t.L;
    a.u = d_u;
    a.batch = batch;
    a.n_rows = b->len;
    a.beta = beta;
    a.out_idx = d_out_idx;
    a.out_w64 = d_out_w64;
    a.out_w32 = d_out_w32;
    a.status = b->status();
    a.tl = b->tl_next();
    if (batch >= SAMPLE_WIDE_MIN)  // enough queries to fill the machine: one thread per query
        PER_CUDA(launch_pdl(k_sample_wide, dim3((unsigned)((batch + SAMPLE_WIDE_THREADS - 1) / SAMPLE_WIDE_THREADS)),
                            dim3(SAMPLE_WIDE_THREADS), 0, (cudaStream_t)stream, (b->pdl & 1) != 0, a));
    else
        PER_CUDA(launch_pdl(k_sample, dim3((unsigned)((batch * 32 + 127) / 128)), dim3(128), 0, (cudaStream_t)stream, (b->pdl & 1) != 0, a));
    PER_CUDA(cudaGetLastError());
    b->last_sample_idx = d_out_idx;
    b->last_sample_batch = batch;
    return PER_OK;
}

int per_shard_record(per_buffer *b, int64_t global_len, double *d_out, void *stream) {
    if (!b || !d_out || global_len < 0) return fail(PER_ERR_ARG, "per_shard_record: bad argument");
    DeviceGuard g(b->device);
    const int64_t need_local = shard_need(global_len, b->idx_base, b->t.cap);
    k_shard_record<<<1, 32, 0, (cudaStream_t)stream>>>(b->t.sum, b->t.mnn, b->t.cap, need_local, b->len, d_out);
    PER_CUDA(cudaGetLastError());
    return PER_OK;
}

int per_shard_attach_record(per_buffer *b, double *d_record, int64_t global_len) {
    if (!b || global_len < 0) return fail(PER_ERR_ARG, "per_shard_attach_record: bad argument");
    if (b->rec_published && global_len != b->global_len)
        return fail(PER_ERR_ARG, "early publish: the record of the next sample is already in the peers' mailboxes; sample before changing len");
    b->rec_out = d_record;
    b->global_len = global_len;
    return PER_OK;
}

// ---- CUDA IPC helpers: buffers another process's GPU can write into / read from over NVLink ----
int per_ipc_alloc(int device, int64_t bytes, void **d_ptr, uint8_t *h_handle64) {
    if (!d_ptr || !h_handle64 || bytes <= 0) return fail(PER_ERR_ARG, "per_ipc_alloc: bad argument");
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
    DeviceGuard g(device);
    if (!g.ok) return fail(PER_ERR_CUDA, "per_ipc_alloc: cannot select device %d", device);
    PER_CUDA(cudaMalloc(d_ptr, (size_t)bytes));
    PER_CUDA(cudaMemset(*d_ptr, 0, (size_t)bytes));
    cudaIpcMemHandle_t hnd;
    PER_CUDA(cudaIpcGetMemHandle(&hnd, *d_ptr));
    memcpy(h_handle64, &hnd, 64);
    return PER_OK;
}

int per_ipc_open(int device, const uint8_t *h_handle64, void **d_ptr) {
    if (!d_ptr || !h_handle64) return fail(PER_ERR_ARG, "per_ipc_open: bad argument");
    DeviceGuard g(device);
    if (!g.ok) return fail(PER_ERR_CUDA, "per_ipc_open: cannot select device %d", device);
    cudaIpcMemHandle_t hnd;
    memcpy(&hnd, h_handle64, 64);
    PER_CUDA(cudaIpcOpenMemHandle(d_ptr, hnd, cudaIpcMemLazyEnablePeerAccess));
    return PER_OK;
}

int per_ipc_close(int device, void *d_ptr) {
    DeviceGuard g(device);
    if (d_ptr) PER_CUDA(cudaIpcCloseMemHandle(d_ptr));
    return PER_OK;
}

int per_ipc_free(int device, void *d_ptr) {
    DeviceGuard g(device);
    if (d_ptr) PER_CUDA(cudaFree(d_ptr));
    return PER_OK;
}

int per_shard_attach_exchange(per_buffer *b, double *d_mailbox, void *const *h_peer_mailboxes) {
    if (!b || !d_mailbox || !h_peer_mailboxes) return fail(PER_ERR_ARG, "per_shard_attach_exchange: bad argument");
    DeviceGuard g(b->device);
    if (!b->d_peer_mailbox) PER_CUDA(cudaMalloc(&b->d_peer_mailbox, sizeof(double *) * 16));
    PER_CUDA(cudaMemcpy(b->d_peer_mailbox, h_peer_mailboxes, sizeof(double *) * b->n_shards, cudaMemcpyHostToDevice));
    b->mailbox = d_mailbox;
    PER_CUDA(cudaMemset(b->xchg_epoch(), 0, sizeof(u64) * 2));
    return PER_OK;
}

int per_shard_exchange_stats(per_buffer *b, int64_t *h_out3, void *stream) {
    if (!b || !h_out3) return fail(PER_ERR_ARG, "per_shard_exchange_stats: bad argument");
    DeviceGuard g(b->device);
    PER_CUDA(cudaMemcpyAsync(h_out3, b->scal + 8, sizeof(u64) * 3, cudaMemcpyDeviceToHost, (cudaStream_t)stream));
    PER_CUDA(cudaMemsetAsync(b->scal + 8, 0, sizeof(u64) * 3, (cudaStream_t)stream));
    PER_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
    return PER_OK;
}

int per_timeline(per_buffer *b, int64_t *h_out, int64_t max_entries, int64_t *h_count, void *stream) {
    if (!b || !h_out || !h_count || max_entries < 0) return fail(PER_ERR_ARG, "per_timeline: bad argument");
    *h_count = 0;
    if (!b->timeline) return PER_OK;
    DeviceGuard g(b->device);
    cudaStream_t s = (cudaStream_t)stream;
    std::vector<u64> host(b->timeline_cap);
    PER_CUDA(cudaMemcpyAsync(host.data(), b->timeline, sizeof(u64) * b->timeline_cap, cudaMemcpyDeviceToHost, s));
    PER_CUDA(cudaMemsetAsync(b->timeline, 0, sizeof(u64) * b->timeline_cap, s));
    PER_CUDA(cudaStreamSynchronize(s));
    int64_t n = 0;
    for (u64 e : host)
        if (e && n < max_entries) h_out[n++] = (int64_t)e;
    *h_count = n;
    return PER_OK;
}

int per_update_trace(per_buffer *b, int64_t *h_out16, void *stream) {
    if (!b || !h_out16) return fail(PER_ERR_ARG, "per_update_trace: bad argument");
    DeviceGuard g(b->device);
    PER_CUDA(cudaMemcpyAsync(h_out16, b->scal + 11, sizeof(u64) * 16, cudaMemcpyDeviceToHost, (cudaStream_t)stream));
    PER_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
    return PER_OK;
}

int per_shard_publish(per_buffer *b, int64_t global_len, void *stream) {
    if (!b || !b->mailbox) return fail(PER_ERR_ARG, "per_shard_publish: no exchange attached");
    DeviceGuard g(b->device);
    k_shard_publish<<<1, 32, 0, (cudaStream_t)stream>>>(b->t.sum, b->t.mnn, b->t.cap, shard_need(global_len, b->idx_base, b->t.cap),
                                                        b->len, b->d_peer_mailbox, b->n_shards, b->shard_rank, b->xchg_epoch());
    PER_CUDA(cudaGetLastError());
    return PER_OK;
}

int per_shard_deliver(per_buffer *b, const int64_t *d_idx, const float *d_w32, int64_t batch, int64_t *d_dst_idx,
                      float *d_dst_w32, void *stream) {
    if (!b || !d_idx || batch < 0 || (!d_dst_idx && !d_dst_w32)) return fail(PER_ERR_ARG, "per_shard_deliver: bad argument");
    if (batch == 0) return PER_OK;
    DeviceGuard g(b->device);
    k_shard_deliver<<<grid_for(batch, 256, b->sm_count), 256, 0, (cudaStream_t)stream>>>(d_idx, d_w32, batch, d_dst_idx, d_dst_w32);
    PER_CUDA(cudaGetLastError());
    return PER_OK;
}

int per_sample_sharded(per_buffer *b, const double *d_records, const double *d_u, int64_t batch, double beta,
                       int64_t global_len, int64_t *d_out_idx, double *d_out_w64, float *d_out_w32, void *stream) {
    if (!b || batch <= 0 || !d_u || !d_out_idx || (!d_records && !b->mailbox))
        return fail(PER_ERR_ARG, "per_sample_sharded: bad argument");
    if (!(beta > 0)) return fail(PER_ERR_ASSERT, "beta > 0");
    if (global_len < 2) return fail(PER_ERR_ASSERT, "sample needs at least 2 stored transitions");
    DeviceGuard g(b->device);
    SampleArgs a;
    memset(&a, 0, sizeof a);
    a.sum = b->t.sum;
    a.mnn = b->t.mnn;
    a.cap = b->t.cap;
    a.L = b->t.L;
    a.u = d_u;
    a.batch = batch;
    a.n_rows = global_len;
    a.beta = beta;
    a.out_idx = d_out_idx;
    a.out_w64 = d_out_w64;
    a.out_w32 = d_out_w32;
    a.status = b->status();
    a.recs = d_records;
    a.n_shards = b->n_shards;
    a.rank = b->shard_rank;
    a.d_epoch = b->xchg_epoch();
    a.d_done = b->xchg_done();
    a.d_range = b->scal + 5;
    a.tl = b->tl_next();
    if (!d_records) {  // records travel through the peer-memory mailboxes
        a.use_mailbox = 1;
        a.d_wait_stats = b->scal + 8;
        a.mailbox = b->mailbox;
        a.peer_mailbox = b->d_peer_mailbox;
        a.rec_need = shard_need(global_len, b->idx_base, b->t.cap);
        a.rec_len = b->len;
        if (b->rec_published) {
            if (global_len != b->global_len)
                return fail(PER_ERR_ARG, "early publish: the published record was computed for global_len %lld, not %lld", (long long)b->global_len, (long long)global_len);
            a.skip_publish = 1;
        }
    }
    b->rec_published = false;
    b->last_sample_idx = d_out_idx;
    b->last_sample_batch = batch;
    // one warp per OWNED stratum in the expected case (batch / n_shards of them, plus slack); a shard that owns
    // more loops over them
    int64_t warps = batch / b->n_shards + batch / (4 * b->n_shards) + 4;
    if (warps > batch) warps = batch;
    if (batch / b->n_shards >= SAMPLE_WIDE_MIN) {  // thousands of own strata: one thread each (the slack of the warp form, in threads)
        const int64_t ctas = (warps + SAMPLE_THREADS - 1) / SAMPLE_THREADS;
        PER_CUDA(launch_pdl(k_sample_sharded<true>, dim3((unsigned)ctas), dim3(SAMPLE_THREADS), 0, (cudaStream_t)stream, (b->pdl & 1) != 0, a));
        PER_CUDA(cudaGetLastError());
        return PER_OK;
    }
    const int64_t ctas = (warps * 32 + SAMPLE_THREADS - 1) / SAMPLE_THREADS;
    PER_CUDA(launch_pdl(k_sample_sharded<false>, dim3((unsigned)ctas), dim3(SAMPLE_THREADS), 0, (cudaStream_t)stream, (b->pdl & 1) != 0, a));
    PER_CUDA(cudaGetLastError());
    return PER_OK;
}

int per_gather(per_buffer *b, const int64_t *d_idx, int64_t batch, void *d_out_obs, void *d_out_next, int64_t *d_out_act,
               float *d_out_rew, float *d_out_done, void *d_out_extra, void *stream) {
    if (!b || batch < 0 || (batch && !d_idx)) return fail(PER_ERR_ARG, "per_gather: bad argument");
    if (batch == 0) return PER_OK;
    DeviceGuard g(b->device);
    cudaStream_t s = (cudaStream_t)stream;
    GatherArgs a;
    memset(&a, 0, sizeof a);
    const uint8_t *srcs[3] = {b->obs, b->next, b->extra};
    void *dsts[3] = {d_out_obs, d_out_next, d_out_extra};
    int64_t rb[3] = {b->obs_bytes, b->obs_bytes, b->extra_bytes};
    int64_t ss[3] = {b->obs_stride, b->obs_stride, b->extra_stride};
    GatherArgs slow = a;
    const GatherTune &tn = b->tune;
    int per_query = 0;
    bool any_slow = false;
    for (int c = 0; c < 3; ++c) {
        RowClass rc;
        rc.src = srcs[c];
        rc.dst = (uint8_t *)dsts[c];
        rc.src_stride = ss[c];
        rc.dst_stride = rb[c];
        rc.row_bytes = rb[c];
        rc.pieces = 0;
        RowClass off = rc;
        off.dst = nullptr;
        const bool wanted = rc.dst && rb[c] > 0;
        const bool tma_ok = wanted && (rb[c] % 16 == 0) && (((uintptr_t)rc.dst) % 16 == 0);
        if (tma_ok) {
            rc.pieces = (int)((rb[c] + tn.piece - 1) / tn.piece);
            per_query += rc.pieces;
            a.rc[c] = rc;
            slow.rc[c] = off;
        } else {
            a.rc[c] = off;
            slow.rc[c] = wanted ? rc : off;
            any_slow |= wanted;
        }
    }
    a.idx = slow.idx = d_idx;
    a.batch = slow.batch = batch;
    a.n_rows = slow.n_rows = b->len;
    a.act = b->act; a.rew = b->rew; a.done = b->done;
    a.out_act = d_out_act; a.out_rew = d_out_rew; a.out_done = d_out_done;
    a.status = slow.status = b->status();
    a.idx_base = slow.idx_base = b->idx_base;
    a.skip_foreign = slow.skip_foreign = b->n_shards > 1;
    a.range = (b->sharded && b->last_sample_idx && d_idx == b->last_sample_idx && batch == b->last_sample_batch) ? b->scal + 3 : nullptr;
    a.piece = slow.piece = tn.piece;
    a.stages = slow.stages = tn.stages;
    a.tl = b->tl_next();
    const int64_t items = batch * per_query;
    if (tn.mode == 2) {
        // balanced mover: vectors of VB bytes per class (classes that cannot be cut into them stay with the slow mover)
        int vb = tn.vbytes;
        for (int c = 0; c < 3; ++c)
            if (a.rc[c].dst && (rb[c] % 32 || ss[c] % 32 || ((uintptr_t)a.rc[c].dst) % 32 || ((uintptr_t)a.rc[c].src) % 32)) vb = 16;
        int v[3];
        for (int c = 0; c < 3; ++c) v[c] = a.rc[c].dst ? (int)(rb[c] / vb) : 0;
        const int64_t vecs = batch * ((int64_t)v[0] + v[1] + v[2]);
        int64_t grid = (int64_t)b->sm_count * tn.ctas_per_sm;
        const int64_t enough = (vecs + 8 * 32 * 4 - 1) / (8 * 32 * 4);  // at least four vectors per lane
        if (grid > enough) grid = enough;
        if (grid < 1) grid = 1;
        const bool pdl = (b->pdl & 2) != 0;
        // the grid is exactly the CTA slots this instantiation gets (its register count decides; asked once per shape)
#define BAL_LAUNCH(VB, D, P)                                                                                              \
    do {                                                                                                                 \
        static int occ = 0;                                                                                              \
        if (!occ) PER_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_gather_bal<VB, D, P>, 256, 0));         \
        const int64_t slots = (int64_t)b->sm_count * (occ < tn.ctas_per_sm ? occ : tn.ctas_per_sm);                      \
        if (grid > slots) grid = slots;                                                                                  \
        PER_CUDA(launch_pdl(k_gather_bal<VB, D, P>, dim3((unsigned)grid), dim3(256), 0, s, pdl, a, v[0], v[1], v[2]));   \
    } while (0)
#define BAL_DEPTHS(VB, D0, D1, D2, D3)                                                   \
    do {                                                                                 \
        const int d = tn.depth;                                                          \
        if (tn.pipe) {                                                                   \
            if (d <= D0) BAL_LAUNCH(VB, D0, true); else if (d <= D1) BAL_LAUNCH(VB, D1, true); \
            else if (d <= D2) BAL_LAUNCH(VB, D2, true); else BAL_LAUNCH(VB, D3, true);   \
        } else {                                                                         \
            if (d <= D0) BAL_LAUNCH(VB, D0, false); else if (d <= D1) BAL_LAUNCH(VB, D1, false); \
            else if (d <= D2) BAL_LAUNCH(VB, D2, false); else BAL_LAUNCH(VB, D3, false); \
        }                                                                                \
    } while (0)
        if (vb == 32) BAL_DEPTHS(32, 2, 4, 6, 8);
        else BAL_DEPTHS(16, 4, 8, 12, 16);
#undef BAL_DEPTHS
#undef BAL_LAUNCH
    } else if (tn.mode == 0) {
        int64_t grid = (items + tn.warps - 1) / tn.warps;
        if (grid > (int64_t)b->sm_count * tn.ctas_per_sm) grid = (int64_t)b->sm_count * tn.ctas_per_sm;  // persistent CTAs
        if (grid < 1) grid = 1;
        const size_t smem = per_query ? (size_t)round_up(tn.warps * tn.stages * 8, 128) + (size_t)tn.warps * tn.stages * tn.piece : 0;
        PER_CUDA(launch_pdl(k_gather_tma, dim3((unsigned)grid), dim3(tn.warps * 32), smem, s, (b->pdl & 2) != 0, a, per_query));
    } else {
        int64_t grid = (items + 7) / 8;
        if (grid > (int64_t)b->sm_count * tn.ctas_per_sm) grid = (int64_t)b->sm_count * tn.ctas_per_sm;
        if (grid < 1) grid = 1;
        PER_CUDA(launch_pdl(k_gather_ldg<16>, dim3((unsigned)grid), dim3(256), 0, s, (b->pdl & 2) != 0, a, per_query));
    }
    if (any_slow) k_gather_bytes<<<grid_for(batch * 3 * 32, 256, b->sm_count), 256, 0, s>>>(slow);
    PER_CUDA(cudaGetLastError());
    return PER_OK;
}

int per_gather_cast(per_buffer *b, const int64_t *d_idx, int64_t batch, int out_dtype, float scale, void *d_out_obs,
                    void *d_out_next, int64_t *d_out_act, float *d_out_rew, float *d_out_done, void *stream) {
    if (!b || batch < 0 || (batch && !d_idx) || out_dtype < 0 || out_dtype > 2) return fail(PER_ERR_ARG, "per_gather_cast: bad argument");
    if (b->obs_bytes % 16) return fail(PER_ERR_ARG, "per_gather_cast: observation rows must be a multiple of 16 bytes");
    if (batch == 0) return PER_OK;
    DeviceGuard g(b->device);
    cudaStream_t s = (cudaStream_t)stream;
    const GatherTune &tn = b->tune;
    GatherArgs a;
    memset(&a, 0, sizeof a);
    void *outs[2] = {d_out_obs, d_out_next};
    const uint8_t *srcs[2] = {b->obs, b->next};
    int per_query = 0;
    for (int c = 0; c < 2; ++c) {
        if (!outs[c]) continue;
        a.rc[c].src = srcs[c];
        a.rc[c].dst = (uint8_t *)outs[c];
        a.rc[c].src_stride = b->obs_stride;
        a.rc[c].dst_stride = b->obs_bytes;
        a.rc[c].row_bytes = b->obs_bytes;
        a.rc[c].pieces = (int)((b->obs_bytes + tn.piece - 1) / tn.piece);
        a.cast_out[c] = (uint8_t *)outs[c];
        per_query += a.rc[c].pieces;
    }
    a.idx = d_idx;
    a.batch = batch;
    a.n_rows = b->len;
    a.act = b->act; a.rew = b->rew; a.done = b->done;
    a.out_act = d_out_act; a.out_rew = d_out_rew; a.out_done = d_out_done;
    a.status = b->status();
    a.idx_base = b->idx_base;
    a.skip_foreign = b->n_shards > 1;
    a.range = (b->sharded && b->last_sample_idx && d_idx == b->last_sample_idx && batch == b->last_sample_batch) ? b->scal + 3 : nullptr;
    a.piece = tn.piece;
    a.stages = tn.stages;
    int64_t grid = (batch * per_query + tn.warps - 1) / tn.warps;
    if (grid > (int64_t)b->sm_count * tn.ctas_per_sm) grid = (int64_t)b->sm_count * tn.ctas_per_sm;
    if (grid < 1) grid = 1;
    const size_t smem = per_query ? (size_t)round_up(tn.warps * tn.stages * 8, 128) + (size_t)tn.warps * tn.stages * tn.piece : 0;
    k_gather_cast<<<(unsigned)grid, tn.warps * 32, smem, s>>>(a, per_query, out_dtype, scale);
    PER_CUDA(cudaGetLastError());
    return PER_OK;
}

int per_gather_cast_nhwc(per_buffer *b, const int64_t *d_idx, int64_t batch, int out_dtype, float scale, int channels,
                         void *d_out_obs, void *d_out_next, int64_t *d_out_act, float *d_out_rew, float *d_out_done, void *stream) {
    if (!b || batch < 0 || (batch && !d_idx) || out_dtype < 0 || out_dtype > 2) return fail(PER_ERR_ARG, "per_gather_cast_nhwc: bad argument");
    if (channels != 4) return fail(PER_ERR_ARG, "per_gather_cast_nhwc: 4 channels (the Atari frame stack) only, got %d", channels);
    if (b->obs_bytes % 16 || b->obs_bytes == 0) return fail(PER_ERR_ARG, "per_gather_cast_nhwc: an observation must be 4 planes of a multiple of 4 bytes");
    if (batch == 0) return PER_OK;
    DeviceGuard g(b->device);
    GatherArgs a;
    memset(&a, 0, sizeof a);
    void *outs[2] = {d_out_obs, d_out_next};
    const uint8_t *srcs[2] = {b->obs, b->next};
    for (int c = 0; c < 2; ++c) {
        a.rc[c].src = srcs[c];
        a.rc[c].src_stride = b->obs_stride;
        a.rc[c].row_bytes = b->obs_bytes;
        a.cast_out[c] = (uint8_t *)outs[c];
    }
    a.idx = d_idx;
    a.batch = batch;
    a.n_rows = b->len;
    a.act = b->act; a.rew = b->rew; a.done = b->done;
    a.out_act = d_out_act; a.out_rew = d_out_rew; a.out_done = d_out_done;
    a.status = b->status();
    a.idx_base = b->idx_base;
    a.skip_foreign = b->n_shards > 1;
    a.range = (b->sharded && b->last_sample_idx && d_idx == b->last_sample_idx && batch == b->last_sample_batch) ? b->scal + 3 : nullptr;
    const int plane = (int)(b->obs_bytes / 4);
    const int64_t work = batch * (plane / 4);
    int64_t grid = (work + 255) / 256;
    if (grid > (int64_t)b->sm_count * 8) grid = (int64_t)b->sm_count * 8;
    cudaStream_t s = (cudaStream_t)stream;
    if (out_dtype == CAST_F32) k_gather_cast_nhwc<float4><<<(unsigned)grid, 256, 0, s>>>(a, plane, scale);
    else if (out_dtype == CAST_BF16) k_gather_cast_nhwc<BF16x4><<<(unsigned)grid, 256, 0, s>>>(a, plane, scale);
    else k_gather_cast_nhwc<F16x4><<<(unsigned)grid, 256, 0, s>>>(a, plane, scale);
    PER_CUDA(cudaGetLastError());
    return PER_OK;
}

int per_tune(per_buffer *b, const char *key, int64_t value) {
    if (!b || !key) return fail(PER_ERR_ARG, "per_tune: bad argument");
    GatherTune t = b->tune;
    if (!strcmp(key, "update_trace")) {
        b->update_trace = value != 0;
        return PER_OK;
    }
    if (!strcmp(key, "timeline")) {
        DeviceGuard g(b->device);
        PER_CUDA(cudaDeviceSynchronize());
        cudaFree(b->timeline);
        b->timeline = nullptr;
        b->timeline_cap = 0;
        b->timeline_seq = 0;
        if (value >= TL_SLOTS) {
            PER_CUDA(cudaMalloc(&b->timeline, sizeof(u64) * (size_t)value));
            PER_CUDA(cudaMemset(b->timeline, 0, sizeof(u64) * (size_t)value));
            b->timeline_cap = (unsigned)value;
        }
        return PER_OK;
    }
    if (!strcmp(key, "early_publish")) {
        if (b->rec_published) return fail(PER_ERR_ARG, "early publish: a published record is pending; sample first");
        b->early_publish = value != 0;
        return PER_OK;
    }
    if (!strcmp(key, "update_cluster_max")) {
        b->upd_cluster_max = (int)value;
        return PER_OK;
    }
    if (!strcmp(key, "update_ctas")) {  // largest cluster of the one-launch update (0: what the device gives, 16 on B200)
        b->upd_max_ctas = (int)value;
        return PER_OK;
    }
    if (!strcmp(key, "dense_update_lmax")) {
        if (value < 0 || value > DENSE_L_MAX) return fail(PER_ERR_ARG, "per_tune: dense_update_lmax must be 0..%d", DENSE_L_MAX);
        b->dense_lmax = (int)value;
        return PER_OK;
    }
    if (!strcmp(key, "sorted_update")) {
        b->sorted_update = value != 0;
        return PER_OK;
    }
    if (!strcmp(key, "pdl")) {  // bit 0: sampling kernels, bit 1: gather, bit 2: update
        b->pdl = (int)value;
        return PER_OK;
    }
    if (!strcmp(key, "gather_mode")) t.mode = (int)value;
    else if (!strcmp(key, "gather_depth")) t.depth = (int)value;
    else if (!strcmp(key, "gather_pipe")) t.pipe = value != 0;
    else if (!strcmp(key, "gather_vbytes")) t.vbytes = (int)value;
    else if (!strcmp(key, "gather_warps")) t.warps = (int)value;
    else if (!strcmp(key, "gather_stages")) t.stages = (int)value;
    else if (!strcmp(key, "gather_piece")) t.piece = (int)value;
    else if (!strcmp(key, "gather_ctas_per_sm")) t.ctas_per_sm = (int)value;
    else return fail(PER_ERR_ARG, "per_tune: unknown key %s", key);
    const int64_t smem = round_up(t.warps * t.stages * 8, 128) + (int64_t)t.warps * t.stages * t.piece;
    if (t.mode < 0 || t.mode > 2 || t.depth < 1 || t.depth > 16 || (t.vbytes != 16 && t.vbytes != 32) || t.warps < 1 || t.warps > GATHER_MAX_WARPS || t.stages < 2 || t.stages > 64 ||
        t.piece < 16 || t.piece % 16 || t.ctas_per_sm < 1 || t.ctas_per_sm > 8 || smem > GATHER_SMEM_MAX ||
        (t.mode == 0 && (smem + 1024) * t.ctas_per_sm > 228 * 1024))
        return fail(PER_ERR_ARG, "per_tune: %s=%lld gives an invalid launch shape (smem %lld B x %d CTAs/SM)", key,
                    (long long)value, (long long)smem, t.ctas_per_sm);
    b->tune = t;
    return PER_OK;
}

static int update_common(per_buffer *b, const int64_t *d_idx, const void *src, int kind, double batch_max, int64_t n,
                         const TdArgs *td, cudaStream_t s) {
    UpdArgs a;
    memset(&a, 0, sizeof a);
    a.t = b->t;
    a.idx = d_idx;
    a.src = src;
    a.src_kind = kind;
    a.alpha = b->alpha;
    a.n = n;
    a.len_limit = b->len;
    a.stamps = b->stamps;
    a.status = b->status();
    a.max_prio = b->max_prio();
    a.batch_max = batch_max;
    a.idx_base = b->idx_base;
    a.skip_foreign = b->n_shards > 1;
    a.rec_out = b->rec_out;
    a.rec_need = shard_need(b->global_len, b->idx_base, b->t.cap);
    a.rec_len = b->len;
    a.range = (b->sharded && b->last_sample_idx && d_idx == b->last_sample_idx && n == b->last_sample_batch) ? b->scal + 3 : nullptr;
    a.n_shards = b->n_shards;
    if (b->early_publish && b->mailbox && n <= UPD_CLUSTER_MAX_ITEMS) {
        if (b->rec_published) return fail(PER_ERR_ARG, "early publish: the record of the next sample is already in the peers' mailboxes; sample before updating again");
        a.publish_to = b->d_peer_mailbox;
        a.n_shards = b->n_shards;
        a.shard_rank = b->shard_rank;
        a.d_epoch = b->xchg_epoch();
        b->rec_published = true;
    }
    a.trace = b->update_trace ? b->scal + 11 : nullptr;
    a.tl = b->tl_next();
    // the buffer the last sample wrote its (ascending) indices into: duplicates are adjacent
    a.sorted_idx = b->last_sample_idx && d_idx == b->last_sample_idx && n == b->last_sample_batch;
    a.sorted_fast = b->sorted_update;
    a.cluster_max = b->upd_cluster_max;
    a.max_ctas = b->upd_max_ctas;
    a.dense_counter = reinterpret_cast<unsigned *>(b->scal + 48);
    a.dense_lmax = b->dense_lmax;
    if (a.sorted_idx) b->last_sample_idx = nullptr;  // the binding to that sample ends with its first update (see per_sample_sharded)
    return run_update(a, td, 0, b->sm_count, s, (b->pdl & 4) != 0);
}

int per_update_priorities(per_buffer *b, const int64_t *d_idx, const void *d_prio, int prio_dtype, int64_t n, void *stream) {
    if (!b || n < 0 || (n && (!d_idx || !d_prio))) return fail(PER_ERR_ARG, "per_update_priorities: bad argument");
    if (prio_dtype != PER_PRIO_F64 && prio_dtype != PER_PRIO_F32) return fail(PER_ERR_ARG, "per_update_priorities: bad dtype");
    if (n == 0) return PER_OK;
    DeviceGuard g(b->device);
    return update_common(b, d_idx, d_prio, prio_dtype == PER_PRIO_F64 ? LEAF_PRIO_F64 : LEAF_PRIO_F32, 0.0, n, nullptr,
                         (cudaStream_t)stream);
}

int per_update_leaves(per_buffer *b, const int64_t *d_idx, const double *d_leaf, double batch_max_priority, int64_t n,
                      void *stream) {
    if (!b || n < 0 || (n && (!d_idx || !d_leaf))) return fail(PER_ERR_ARG, "per_update_leaves: bad argument");
    if (n == 0) return PER_OK;
    DeviceGuard g(b->device);
    return update_common(b, d_idx, d_leaf, LEAF_GIVEN, batch_max_priority, n, nullptr, (cudaStream_t)stream);
}

// ---- host-buffer entry points: what a binding without device pointers calls (memory.py's sample() and
// update_priorities() go through these; bench.py's e2e number is measured on them) ----
int per_sample_host(per_buffer *b, const double *h_u, int64_t batch, double beta, void *h_obs, void *h_next, int64_t *h_act,
                    float *h_rew, float *h_done, void *h_extra, double *h_w64, int64_t *h_idx, void *stream) {
    if (!b || !h_u || batch <= 0) return fail(PER_ERR_ARG, "per_sample_host: bad argument");
    DeviceGuard g(b->device);
    cudaStream_t s = (cudaStream_t)stream;
    int rc = ensure_host_path(b, batch, s);
    if (rc) return rc;
    per_buffer::HostPath &h = b->hp;
    PER_CUDA(cudaEventSynchronize(h.up_done));  // the pinned block is free again (no-op before the first upload)
    memcpy(h.h_pin, h_u, sizeof(double) * batch);
    PER_CUDA(cudaMemcpyAsync(h.d_u, h.h_pin, sizeof(double) * batch, cudaMemcpyHostToDevice, s));
    PER_CUDA(cudaEventRecord(h.up_done, s));
    rc = per_sample_indices(b, h.d_u, batch, beta, h.d_idx, h_w64 ? h.d_w64 : nullptr, nullptr, stream);
    if (rc) return rc;
    const bool rows = h_obs || h_next || h_act || h_rew || h_done || h_extra;
    if (rows) {
        rc = per_gather(b, h.d_idx, batch, h_obs ? h.d_obs : nullptr, h_next ? h.d_next : nullptr, h_act ? h.d_act : nullptr,
                        h_rew ? h.d_rew : nullptr, h_done ? h.d_done : nullptr, (h_extra && b->extra_bytes) ? h.d_extra : nullptr, stream);
        if (rc) return rc;
    }
    // the two frame batches first: they are the PCIe time, everything else hides behind them
    if (h_obs) PER_CUDA(cudaMemcpyAsync(h_obs, h.d_obs, (size_t)b->obs_bytes * batch, cudaMemcpyDeviceToHost, s));
    if (h_next) PER_CUDA(cudaMemcpyAsync(h_next, h.d_next, (size_t)b->obs_bytes * batch, cudaMemcpyDeviceToHost, s));
    if (h_extra && b->extra_bytes) PER_CUDA(cudaMemcpyAsync(h_extra, h.d_extra, (size_t)b->extra_bytes * batch, cudaMemcpyDeviceToHost, s));
    // the per-row scalars (32 bytes per row) in one copy into the pinned mirror, handed out from there
    PER_CUDA(cudaMemcpyAsync(h.h_small, h.d_small, 32 * (size_t)h.cap, cudaMemcpyDeviceToHost, s));
    PER_CUDA(cudaStreamSynchronize(s));
    const size_t cap = (size_t)h.cap;
    if (h_w64) memcpy(h_w64, h.h_small, sizeof(double) * batch);
    if (h_idx) memcpy(h_idx, h.h_small + 8 * cap, sizeof(int64_t) * batch);
    if (h_act) memcpy(h_act, h.h_small + 16 * cap, sizeof(int64_t) * batch);
    if (h_rew) memcpy(h_rew, h.h_small + 24 * cap, sizeof(float) * batch);
    if (h_done) memcpy(h_done, h.h_small + 28 * cap, sizeof(float) * batch);
    return PER_OK;
}

int per_update_host(per_buffer *b, const int64_t *h_idx, const double *h_prio, int64_t n, double *h_batch_max, void *stream) {
    if (!b || n < 0 || (n && (!h_idx || !h_prio))) return fail(PER_ERR_ARG, "per_update_host: bad argument");
    if (h_batch_max) *h_batch_max = 0.0;
    if (n == 0) return PER_OK;
    double bmax = 0.0;
    for (int64_t i = 0; i < n; ++i) {  // the reference asserts pair by pair (memory.py:315-316); here before anything is applied
        if (!(h_prio[i] > 0)) return fail(PER_ERR_ASSERT, "priority > 0");
        if (h_idx[i] < 0 || h_idx[i] >= b->len) return fail(PER_ERR_ASSERT, "0 <= idx < len(self._storage)");
        if (h_prio[i] > bmax) bmax = h_prio[i];
    }
    DeviceGuard g(b->device);
    cudaStream_t s = (cudaStream_t)stream;
    int rc = ensure_host_path(b, n, s);
    if (rc) return rc;
    per_buffer::HostPath &h = b->hp;
    PER_CUDA(cudaEventSynchronize(h.up_done));
    int64_t *p_idx = reinterpret_cast<int64_t *>(h.h_pin + 8 * (size_t)h.cap);
    double *p_leaf = reinterpret_cast<double *>(h.h_pin + 16 * (size_t)h.cap);
    memcpy(p_idx, h_idx, sizeof(int64_t) * n);
    for (int64_t i = 0; i < n; ++i) p_leaf[i] = pow(h_prio[i], b->alpha);  // libm: the bits CPython produces at memory.py:317
    int64_t *d_idx = reinterpret_cast<int64_t *>(h.d_upd);
    double *d_leaf = reinterpret_cast<double *>(h.d_upd + 8 * (size_t)h.cap);
    // indices and leaves are neighbours in the pinned block and in d_upd: one copy ([n, cap) of the index part is stale, unused)
    PER_CUDA(cudaMemcpyAsync(d_idx, p_idx, 8 * (size_t)h.cap + sizeof(double) * n, cudaMemcpyHostToDevice, s));
    PER_CUDA(cudaEventRecord(h.up_done, s));
    if (h_batch_max) *h_batch_max = bmax;
    return per_update_leaves(b, d_idx, d_leaf, bmax, n, stream);
}

int per_pow_alpha_host(const double *h_prio, double alpha, double *h_out, int64_t n) {
    if (n < 0 || (n && (!h_prio || !h_out))) return fail(PER_ERR_ARG, "per_pow_alpha_host: bad argument");
    for (int64_t i = 0; i < n; ++i) h_out[i] = pow(h_prio[i], alpha);
    return PER_OK;
}

int per_td_priority_update(per_buffer *b, const float *d_q, const float *d_q_next, const float *d_q_tgt, int64_t n_actions,
                           const int64_t *d_act, const float *d_rew, const float *d_done, const float *d_w, float gamma_n,
                           const int64_t *d_idx, int64_t batch, float *d_out_prio, float *d_out_td, float *d_out_loss,
                           void *stream) {
    if (!b || batch <= 0 || n_actions <= 0 || !d_q || !d_q_next || !d_q_tgt || !d_act || !d_rew || !d_done)
        return fail(PER_ERR_ARG, "per_td_priority_update: bad argument");
    if (d_out_loss && !d_w) return fail(PER_ERR_ARG, "per_td_priority_update: loss needs weights");
    DeviceGuard g(b->device);
    cudaStream_t s = (cudaStream_t)stream;
    if (batch > b->td_scratch_cap && (!d_out_prio || !d_out_td)) {
        PER_CUDA(cudaStreamSynchronize(s));
        cudaFree(b->td_scratch);
        b->td_scratch = nullptr;
        PER_CUDA(cudaMalloc(&b->td_scratch, sizeof(float) * 2 * batch));
        b->td_scratch_cap = batch;
    }
    TdArgs t;
    memset(&t, 0, sizeof t);
    t.q = d_q; t.q_next = d_q_next; t.q_tgt = d_q_tgt;
    t.nA = n_actions;
    t.act = d_act; t.rew = d_rew; t.done = d_done; t.w = d_w;
    t.gamma_n = gamma_n;
    t.batch = batch;
    t.out_td = d_out_td ? d_out_td : b->td_scratch;
    t.out_prio = d_out_prio ? d_out_prio : b->td_scratch + b->td_scratch_cap;
    t.out_loss = d_out_loss;
    t.red = b->td_red();
    t.status = b->status();
    if (batch <= UPD_CLUSTER_MAX_ITEMS) {
        if (d_idx) return update_common(b, d_idx, t.out_prio, LEAF_PRIO_F32, 0.0, batch, &t, s);
        UpdArgs a;
        memset(&a, 0, sizeof a);
        a.t = b->t;
        a.n = batch;
        return run_update(a, &t, 1, b->sm_count, s, (b->pdl & 4) != 0);
    }
    const int grid = grid_for(batch, 256, b->sm_count);
    k_td_reset<<<1, 1, 0, s>>>(t.red);
    k_td_pass1<<<grid, 256, 0, s>>>(t);
    k_td_pass2<<<grid, 256, 0, s>>>(t);
    k_td_finish<<<1, 1, 0, s>>>(t);
    PER_CUDA(cudaGetLastError());
    if (!d_idx) return PER_OK;
    return update_common(b, d_idx, t.out_prio, LEAF_PRIO_F32, 0.0, batch, nullptr, s);
}

int per_tree_read(per_buffer *b, int which, double *d_out, void *stream) {
    if (!b || !d_out || (which != PER_TREE_SUM && which != PER_TREE_MIN)) return fail(PER_ERR_ARG, "per_tree_read: bad argument");
    DeviceGuard g(b->device);
    PER_CUDA(cudaMemcpyAsync(d_out, which == PER_TREE_SUM ? b->t.sum : b->t.mnn, sizeof(double) * 2 * b->t.cap,
                             cudaMemcpyDeviceToDevice, (cudaStream_t)stream));
    return PER_OK;
}

int per_tree_get(per_buffer *b, int which, int64_t idx, double *h_out, void *stream) {
    if (!b || !h_out || (which != PER_TREE_SUM && which != PER_TREE_MIN)) return fail(PER_ERR_ARG, "per_tree_get: bad argument");
    DeviceGuard g(b->device);
    return read_leaf(which == PER_TREE_SUM ? b->t.sum : b->t.mnn, b->t.cap, idx, h_out, (cudaStream_t)stream);
}

int per_tree_reduce(per_buffer *b, int which, int64_t start, int64_t end, double *d_out, void *stream) {
    if (!b || !d_out || (which != PER_TREE_SUM && which != PER_TREE_MIN)) return fail(PER_ERR_ARG, "per_tree_reduce: bad argument");
    if (end < 0) end += b->t.cap;  // memory.py:71-72
    if (start < 0 || end > b->t.cap || end - 1 < start)
        return fail(PER_ERR_ASSERT, "per_tree_reduce: empty or out-of-range [%lld, %lld)", (long long)start, (long long)end);
    DeviceGuard g(b->device);
    k_range_reduce<<<1, 32, 0, (cudaStream_t)stream>>>(which == PER_TREE_SUM ? b->t.sum : b->t.mnn, b->t.cap, start, end - 1,
                                                       which == PER_TREE_SUM, d_out);
    PER_CUDA(cudaGetLastError());
    return PER_OK;
}

// ------------------------------------------------------------------------------ AQL candidate Q
struct aql_plan_s {
    int device, continuous, A, sm_count;
    uint8_t *img;        // swizzled TF32 hi/lo weight image
    AqlSmall small;      // small parameters (host copy; travels to the kernel as a by-value parameter)
    unsigned long long *trace;  // 8 words of phase cycle counters (diagnostics)
    bool trace_on;       // set by the first aql_trace call
    float *sv;           // [B, 64] state halves
    int64_t sv_cap;
    bool have_weights;
};

int aql_create(aql_plan **out, int device, int continuous, int action_dim) {
    if (!out) return fail(PER_ERR_ARG, "aql_create: null out");
    if (continuous && (action_dim < 1 || action_dim > AQL_MAX_A))
        return fail(PER_ERR_ARG, "aql_create: action_dim must be in [1, %d]", AQL_MAX_A);
    DeviceGuard g(device);
    if (!g.ok) return fail(PER_ERR_CUDA, "aql_create: cannot select device %d", device);
    aql_plan *p = new aql_plan();
    memset(p, 0, sizeof *p);
    p->device = device;
    p->continuous = continuous ? 1 : 0;
    p->A = continuous ? action_dim : 1;
    cudaDeviceGetAttribute(&p->sm_count, cudaDevAttrMultiProcessorCount, device);
    if (cudaMalloc(&p->trace, 64) == cudaSuccess) cudaMemset(p->trace, 0, 64);
    if (cudaMalloc(&p->img, AQL_IMG_BYTES) != cudaSuccess) {
        cudaGetLastError();
        delete p;
        return fail(PER_ERR_NOMEM, "aql_create: cudaMalloc failed");
    }
    *out = p;
    return PER_OK;
}

int aql_destroy(aql_plan *p) {
    if (!p) return PER_OK;
    DeviceGuard g(p->device);
    cudaFree(p->img);
    cudaFree(p->sv);
    delete p;
    return PER_OK;
}

int aql_set_weights(aql_plan *p, const float *d_wa1, const float *d_ba1, const float *d_wa2, const float *d_ba2,
                    const float *d_wadv1, const float *d_badv1, const float *d_wadv2, const float *d_badv2, void *stream) {
    if (!p || !d_wa1 || !d_ba1 || !d_wadv1 || !d_badv1 || !d_wadv2 || !d_badv2 || (p->continuous && (!d_wa2 || !d_ba2)))
        return fail(PER_ERR_ARG, "aql_set_weights: null weight");
    DeviceGuard g(p->device);
    cudaStream_t s = (cudaStream_t)stream;
    const int e1 = p->continuous ? AQL_E1 : AQL_H;
    memset(&p->small, 0, sizeof(AqlSmall));
    PER_CUDA(cudaMemcpyAsync(p->small.wa1, d_wa1, sizeof(float) * e1 * p->A, cudaMemcpyDeviceToHost, s));
    PER_CUDA(cudaMemcpyAsync(p->small.ba1, d_ba1, sizeof(float) * e1, cudaMemcpyDeviceToHost, s));
    if (p->continuous) PER_CUDA(cudaMemcpyAsync(p->small.ba2, d_ba2, sizeof(float) * AQL_H, cudaMemcpyDeviceToHost, s));
    PER_CUDA(cudaMemcpyAsync(p->small.w2, d_wadv2, sizeof(float) * AQL_H, cudaMemcpyDeviceToHost, s));
    PER_CUDA(cudaMemcpyAsync(&p->small.b2, d_badv2, sizeof(float), cudaMemcpyDeviceToHost, s));
    PER_CUDA(cudaStreamSynchronize(s));  // the small weights ride in the kernel's parameter block from now on
    k_aql_prep_weights<<<(AQL_H * 48 + 127) / 128, 128, 0, s>>>(p->continuous ? d_wa2 : nullptr, d_wadv1, p->img);
    PER_CUDA(cudaGetLastError());
    p->have_weights = true;
    return PER_OK;
}

int aql_candidate_q(aql_plan *p, const float *d_wadv1, const float *d_badv1, const float *d_f, const void *d_a_mu, int64_t B,
                    int64_t K, float *d_q, int64_t *d_argmax, void *stream) {
    if (!p || !d_wadv1 || !d_badv1 || !d_f || !d_a_mu || !d_q || B <= 0 || K <= 0) return fail(PER_ERR_ARG, "aql_candidate_q: bad argument");
    if (!p->have_weights) return fail(PER_ERR_ARG, "aql_candidate_q: call aql_set_weights first");
    DeviceGuard g(p->device);
    cudaStream_t s = (cudaStream_t)stream;
    if (B > p->sv_cap) {
        PER_CUDA(cudaStreamSynchronize(s));
        cudaFree(p->sv);
        p->sv = nullptr;
        PER_CUDA(cudaMalloc(&p->sv, sizeof(float) * B * AQL_H));
        p->sv_cap = B;
    }
    k_aql_state_half<<<(unsigned)((B * AQL_H + 255) / 256), 256, 0, s>>>(d_wadv1, d_badv1, d_f, B, p->sv);
    AqlArgs a;
    memset(&a, 0, sizeof a);
    a.img = p->img;
    a.sv = p->sv;
    a.a_mu = d_a_mu;
    a.K = K;
    a.total = B * K;
    a.q = d_q;
    a.trace = p->trace_on ? p->trace : nullptr;
    int64_t tiles = (a.total + AQL_TILE - 1) / AQL_TILE;
    const int64_t slots = (int64_t)p->sm_count * AQL_CTAS_PER_SM;
    int64_t grid = tiles < slots ? tiles : slots;
    // The three launches are chained with programmatic dependent launch: the state-half kernel releases its
    // dependents at once, so the Q kernel's prologue (barriers, TMEM, 96 KB weight image) and its first two
    // layers run beside it and only the first read of sv waits (griddepcontrol.wait); likewise the argmax grid
    // is resident and waiting when the Q kernel retires.
    cudaLaunchAttribute pdl[1];
    pdl[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    pdl[0].val.programmaticStreamSerializationAllowed = 1;
    cudaLaunchConfig_t cfg;
    memset(&cfg, 0, sizeof cfg);
    cfg.stream = s;
    cfg.attrs = pdl;
    cfg.numAttrs = 1;
    cfg.gridDim = dim3((unsigned)grid);
    cfg.blockDim = dim3(AQL_THREADS);
    cfg.dynamicSmemBytes = AQL_SMEM_BYTES;
#define AQL_LAUNCH(AA, CC)                                                                                          \
    do {                                                                                                           \
        PER_CUDA(cudaFuncSetAttribute(k_aql_q<AA, CC>, cudaFuncAttributeMaxDynamicSharedMemorySize, AQL_SMEM_BYTES)); \
        PER_CUDA(cudaLaunchKernelEx(&cfg, k_aql_q<AA, CC>, a, p->small));                                          \
    } while (0)
    if (!p->continuous) AQL_LAUNCH(1, false);
    else switch (p->A) {
        case 1: AQL_LAUNCH(1, true); break;
        case 2: AQL_LAUNCH(2, true); break;
        case 3: AQL_LAUNCH(3, true); break;
        case 4: AQL_LAUNCH(4, true); break;
        case 5: AQL_LAUNCH(5, true); break;
        case 6: AQL_LAUNCH(6, true); break;
        case 7: AQL_LAUNCH(7, true); break;
        default: AQL_LAUNCH(8, true); break;
    }
#undef AQL_LAUNCH
    if (d_argmax) {
        cfg.gridDim = dim3((unsigned)B);
        cfg.blockDim = dim3(AQL_ARGMAX_THREADS);
        cfg.dynamicSmemBytes = 0;
        PER_CUDA(cudaLaunchKernelEx(&cfg, k_aql_argmax, (const float *)d_q, B, K, d_argmax));
    }
    PER_CUDA(cudaGetLastError());
    return PER_OK;
}

int aql_trace(aql_plan *p, int64_t *h_out8, void *stream) {
    if (!p || !h_out8 || !p->trace) return fail(PER_ERR_ARG, "aql_trace: bad argument");
    p->trace_on = true;  // the stamps cost the traced CTA a global read-modify-write per phase: off until asked for
    DeviceGuard g(p->device);
    PER_CUDA(cudaMemcpyAsync(h_out8, p->trace, 64, cudaMemcpyDeviceToHost, (cudaStream_t)stream));
    PER_CUDA(cudaMemsetAsync(p->trace, 0, 64, (cudaStream_t)stream));
    PER_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
    return PER_OK;
}

// ------------------------------------------------------------------------------ n-step assembly
int nstep_assemble(int device, int64_t n_streams, const int64_t *d_offsets, const int64_t *d_emit_from, const double *d_reward,
                   const int64_t *d_action, const uint8_t *d_done, const float *d_q, int64_t n_actions, int n_steps,
                   const double *h_gamma_pow, int64_t *d_out_first, int64_t *d_out_next, int64_t *d_out_action,
                   double *d_out_return, float *d_out_done, double *d_out_prio, int64_t *d_out_count, uint64_t *d_status,
                   void *stream) {
    if (n_streams < 0 || n_actions <= 0 || !h_gamma_pow || !d_status) return fail(PER_ERR_ARG, "nstep_assemble: bad argument");
    if (n_steps < 1 || n_steps > NSTEP_MAX_N) return fail(PER_ERR_ARG, "nstep_assemble: n_steps must be in [1, %d]", NSTEP_MAX_N);
    if (n_streams == 0) return PER_OK;
    if (!d_offsets || !d_reward || !d_action || !d_done || !d_q || !d_out_first || !d_out_next || !d_out_action ||
        !d_out_return || !d_out_done || !d_out_prio || !d_out_count)
        return fail(PER_ERR_ARG, "nstep_assemble: null buffer");
    DeviceGuard g(device);
    if (!g.ok) return fail(PER_ERR_CUDA, "nstep_assemble: cannot select device %d", device);
    NstepArgs a;
    memset(&a, 0, sizeof a);
    a.n_streams = n_streams;
    a.offsets = d_offsets;
    a.emit_from = d_emit_from;
    a.reward = d_reward;
    a.action = d_action;
    a.done = d_done;
    a.q = d_q;
    a.nA = n_actions;
    a.n = n_steps;
    for (int k = 0; k <= n_steps; ++k) a.gpow[k] = h_gamma_pow[k];
    a.out_first = d_out_first;
    a.out_next = d_out_next;
    a.out_action = d_out_action;
    a.out_return = d_out_return;
    a.out_done = d_out_done;
    a.out_prio = d_out_prio;
    a.out_count = d_out_count;
    a.status = reinterpret_cast<u64 *>(d_status);
    k_nstep<<<(unsigned)((n_streams + 3) / 4), 128, 0, (cudaStream_t)stream>>>(a);
    PER_CUDA(cudaGetLastError());
    return PER_OK;
}

// ------------------------------------------------------------------------------ no-allocation entry
// seg_set_noalloc: the batched tree write of seg_set for callers that may not allocate device
// memory.  seg_set keeps its last-write-wins arbitration stamps (8 bytes per leaf) in memory
// that seg_create obtained with cudaMalloc; here the stamps live in the file-scope array
// below, which every GPU that loads this library carries from module load until the process
// exits (8 MiB + 256 B).  Host code must not pass `g_noalloc_scratch` itself to a kernel (in
// host code that name is only a stand-in for the symbol): the device address is looked up
// once per device with cudaGetSymbolAddress, with that device current, and cached.
#define SEG_NOALLOC_HEADER_WORDS 32  /* 256 B: word 0 is the fallback status word */
static __device__ __align__(256) unsigned long long g_noalloc_scratch[SEG_NOALLOC_HEADER_WORDS + SEG_NOALLOC_MAX_CAPACITY];

namespace {
constexpr int NOALLOC_MAX_DEVICES = 64;
struct NoallocDevice {
    u64 *base = nullptr;  // device address of g_noalloc_scratch on this device
    int sm_count = 0;
};
NoallocDevice g_noalloc_dev[NOALLOC_MAX_DEVICES];
}  // namespace

int seg_set_noalloc(int device, double *d_sum_nodes, double *d_min_nodes, int64_t capacity, const int64_t *d_idx,
                    const double *d_val, int64_t n, uint64_t *d_status, void *stream) {
    if (device < 0 || device >= NOALLOC_MAX_DEVICES) return fail(PER_ERR_ARG, "seg_set_noalloc: bad device %d", device);
    if ((!d_sum_nodes && !d_min_nodes) || n < 0 || (n && (!d_idx || !d_val)))
        return fail(PER_ERR_ARG, "seg_set_noalloc: bad argument");
    if (!is_pow2(capacity)) return fail(PER_ERR_ASSERT, "capacity must be positive and a power of 2.");
    if (capacity > SEG_NOALLOC_MAX_CAPACITY)
        return fail(PER_ERR_ARG,
                    "seg_set_noalloc: capacity %lld exceeds the built-in scratch (SEG_NOALLOC_MAX_CAPACITY = %lld leaves); "
                    "use seg_create/seg_set for larger trees",
                    (long long)capacity, (long long)SEG_NOALLOC_MAX_CAPACITY);
    if (n >= (int64_t(1) << 32)) return fail(PER_ERR_ARG, "seg_set_noalloc: n too large");
    if (n == 0) return PER_OK;
    DeviceGuard g(device);
    if (!g.ok) return fail(PER_ERR_CUDA, "seg_set_noalloc: cannot select device %d", device);
    cudaStream_t s = (cudaStream_t)stream;
    NoallocDevice &nd = g_noalloc_dev[device];
    if (!nd.base) {  // first call on this device: resolve the symbol while the device is current
        void *p = nullptr;
        PER_CUDA(cudaGetSymbolAddress(&p, g_noalloc_scratch));
        PER_CUDA(cudaDeviceGetAttribute(&nd.sm_count, cudaDevAttrMultiProcessorCount, device));
        nd.base = static_cast<u64 *>(p);
    }
    UpdArgs a;
    memset(&a, 0, sizeof a);
    a.t.sum = d_sum_nodes;
    a.t.mnn = d_min_nodes;
    a.t.cap = capacity;
    a.t.L = ilog2_exact(capacity);
    a.idx = d_idx;
    a.src = d_val;
    a.src_kind = LEAF_GIVEN;
    a.alpha = 1.0;
    a.n = n;
    a.len_limit = capacity;
    a.stamps = nd.base + SEG_NOALLOC_HEADER_WORDS;
    a.status = d_status ? reinterpret_cast<u64 *>(d_status) : nd.base;
    return run_update(a, nullptr, 0, nd.sm_count, s);
}

}  // extern "C"
