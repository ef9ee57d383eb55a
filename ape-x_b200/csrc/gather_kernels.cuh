// gather_kernels.cuh — K4: gather sampled transitions out of the HBM-resident ring into
// dense batch arrays (memory.py:171-182 `_encode_sample`), and the ring insert (memory.py:162-169).
//
// A transition row is two observations (uint8[4,84,84] = 28 224 B each on Atari), an int64
// action, float32 reward and done, and an optional extra row (AQL's a_mu).  Rows are padded
// to a multiple of 16 B so that every row start is 16-B aligned.  Three movers (per_tune "gather_mode"):
//   k_gather_bal  (2, default) the batch as one linear space of 16-byte vectors cut into equal shares, one per warp of a
//                 grid sized to exactly the resident CTA slots; register-staged streaming loads / stores;
//   k_gather_ldg  (1) each row cut into pieces of at most PIECE_BYTES, one warp per piece, sixteen 16-byte streaming
//                 loads in flight per lane, then the stores;
//   k_gather_tma  (0) pieces move HBM -> shared memory -> HBM as two TMA bulk copies (cp.async.bulk, SASS UBLKCP) issued
//                 by ONE lane per warp through an mbarrier ring: no registers are staged, the LSU is idle.
// Alone: 9.3 / 10.6 / 10.5 us per 512 rows (profiles/r02_gather_sweep.txt); inside the learner step the TMA mover's
// 226 KB/SM of dynamic shared memory also slows the tree kernels before and after it (profiles/r01_notes.md).
// k_gather_cast (gather + uint8 -> float widening) stages through shared memory like k_gather_tma; k_gather_cast_nhwc
// writes channels-last rows.
// Algorithmic traffic per sampled transition: read 2*28 224+16 B, write the same.
#pragma once
#include <cuda_bf16.h>
#include <cuda_fp16.h>

#include "common.cuh"

namespace perb {

constexpr int GATHER_MAX_WARPS = 16;    // independent copy engines per CTA (upper bound)
constexpr int GATHER_SMEM_MAX = 200 * 1024;

// Launch shape of the gather; defaults chosen by the sweeps in profiles/ (see DESIGN.md section 3).
struct GatherTune {
    int mode = 2;        // 0: TMA bulk copies through shared memory (k_gather_tma), 1: one piece per warp through registers
                         // (k_gather_ldg), 2: balanced shares through registers (k_gather_bal; profiles/r02_gather_sweep.txt)
    int warps = 4;       // mode 0: issuing warps per CTA
    int stages = 2;      // mode 0: pieces in flight per warp
    int piece = 7056;    // modes 0/1: bytes per piece (multiple of 16); 7056 = one 84x84 frame
    int ctas_per_sm = 4; // CTAs per SM at most (mode 0: 4 x (4 x 2 x 7056 B) = 226 KB of shared memory per SM; mode 2: the
                         // kernel's own occupancy caps it, 3 at depth 8)
    int depth = 8;       // mode 2: vectors per lane and stage
    int pipe = 0;        // mode 2: 1 = the next stage's loads go out before this stage's stores (costs registers, hence CTAs per SM:
                         // measured slower at every depth)
    int vbytes = 16;     // mode 2: bytes per vector (16: LDG.128, 32: LDG.256 with L2 eviction hints)
};

struct RowClass {
    const uint8_t *src;  // ring base
    uint8_t *dst;        // batch base (null: skip)
    int64_t src_stride;  // padded row bytes in the ring
    int64_t dst_stride;  // row bytes in the batch array (unpadded)
    int64_t row_bytes;   // bytes to move per row (multiple of 16 when moved by TMA)
    int pieces;          // ceil(row_bytes / piece)
};

struct GatherArgs {
    RowClass rc[3];  // obs, next_obs, extra
    const int64_t *idx;
    int64_t batch;
    int64_t n_rows;  // rows present (for clamping bad indices)
    const int64_t *act; const float *rew; const float *done;
    int64_t *out_act; float *out_rew; float *out_done;
    u64 *status;
    int piece;   // bytes per piece
    int stages;  // TMA ring depth per warp
    int64_t idx_base;  // sharded replay: global index of this shard's row 0
    int skip_foreign;  // sharded replay: idx == -1 = another shard's row, skipped silently
    const u64 *range;  // sharded replay: owned strata of the last sample (see owned_range); null = all queries
    u64 *tl;           // timeline (diagnostics), see tl_stamp: this launch's first entry
    uint8_t *cast_out[2];  // k_gather_cast: widened obs / next_obs batches (rc[0..1].dst then only define offsets)
};

__device__ __forceinline__ uint32_t smem_u32(const void *p) {
    return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "W_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra D_%=;\n"
        "bra W_%=;\n"
        "D_%=:\n"
        "}\n" ::"r"(bar), "r"(parity)
        : "memory");
}
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void *src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
                 "l"(src), "r"(bytes), "r"(bar)
                 : "memory");
}
__device__ __forceinline__ void bulk_s2g(void *dst, uint32_t src, uint32_t bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(src), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ int round_up_dev(int x, int m) { return (x + m - 1) / m * m; }
template <int N>
__device__ __forceinline__ void bulk_wait_read() {
    asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}

// Decodes work item w -> source / destination / size.  Items are ordered row-major over
// (query, class, piece) so that consecutive warps touch consecutive pieces of one row.
struct Piece {
    const uint8_t *src;
    uint8_t *dst;
    uint32_t bytes;
    int cls;       // row class (0 obs, 1 next_obs, 2 extra)
    int64_t elem;  // byte offset of the piece inside the (uint8) batch array of its class
};
__device__ __forceinline__ Piece decode_piece(const GatherArgs &a, int64_t w, int per_query, int64_t first_q = 0) {
    int64_t q = w / per_query;
    int r = (int)(w - q * per_query);
    q += first_q;
    int c = 0;
    while (r >= a.rc[c].pieces) {
        r -= a.rc[c].pieces;
        ++c;
    }
    int64_t k = a.idx[q];
    const bool foreign = a.skip_foreign && k == -1;
    k -= a.idx_base;
    if (k < 0 || k >= a.n_rows) k = 0;  // flagged by the scalar pass; keep the copy in bounds
    const RowClass &rc = a.rc[c];
    int64_t off = (int64_t)r * a.piece;
    int64_t left = rc.row_bytes - off;
    Piece p;
    p.cls = c;
    p.elem = q * rc.dst_stride + off;
    if (foreign) {
        p.src = nullptr;
        p.dst = nullptr;
        p.bytes = 0;
        return p;
    }
    p.src = rc.src + k * rc.src_stride + off;
    p.dst = rc.dst + q * rc.dst_stride + off;
    p.bytes = (uint32_t)(left < a.piece ? left : a.piece);
    return p;
}

// Owned strata of a sharded sample (see owned_range in update_kernels.cuh).  ONE thread per CTA reads the three tag words
// and hands the range to the others through shared memory: with every warp of the grid reading them (3 552 warps x 3
// loads of one 32-byte sector) the requests queue up at a single L2 slice and the whole gather starts ~1.7 us late
// (measured: bound gather 11.05 -> see profiles/r02_notes.md).  Must be called by all threads of the CTA.
__device__ __forceinline__ void gather_range(const GatherArgs &a, int64_t &first, int64_t &last) {
    first = 0;
    last = a.batch;
    if (!a.range) return;  // uniform over the grid
    __shared__ int64_t s_range[2];
    if (threadIdx.x == 0) {
        const u64 epoch = __ldcg(a.range), tl = __ldcg(a.range + 2), th = __ldcg(a.range + 3);
        int64_t f = 0, l = 0;  // tags of another epoch: this shard owned no stratum of that sample
        if ((tl >> 32) == (epoch & 0xffffffffu) && (th >> 32) == (epoch & 0xffffffffu)) {
            f = (int64_t)(0xffffffffu - (unsigned)(tl & 0xffffffffu));
            l = (int64_t)(th & 0xffffffffu);
        }
        s_range[0] = f;
        s_range[1] = l;
    }
    __syncthreads();
    first = s_range[0];
    last = s_range[1];
}

__device__ __forceinline__ void gather_scalars(const GatherArgs &a, int64_t first, int64_t last) {
    // scalar columns (action / reward / done): plain coalesced loads, all threads
    for (int64_t q = first + blockIdx.x * (int64_t)blockDim.x + threadIdx.x; q < last; q += (int64_t)gridDim.x * blockDim.x) {
        int64_t k = a.idx[q];
        if (a.skip_foreign && k == -1) continue;  // another shard's row
        k -= a.idx_base;
        if (k < 0 || k >= a.n_rows) {
            atomicOr(a.status, (u64)PER_STATUS_BAD_INDEX);
            k = 0;
        }
        if (a.out_act) a.out_act[q] = a.act[k];
        if (a.out_rew) a.out_rew[q] = a.rew[k];
        if (a.out_done) a.out_done[q] = a.done[k];
    }
}

// Dynamic shared memory: [warps*stages mbarriers, padded to 128 B][warps*stages pieces].
__global__ void __launch_bounds__(GATHER_MAX_WARPS * 32) k_gather_tma(GatherArgs a, int per_query) {
    extern __shared__ __align__(128) uint8_t g_smem[];
    const bool tl_on = a.tl && blockIdx.x == 0 && threadIdx.x == 0;
    if (tl_on) tl_stamp(a.tl, 0, 4);
    pdl_launch_dependents();
    pdl_wait();  // indices and range tags come from the sampling kernel in front of this one
    if (tl_on) tl_stamp(a.tl, 1, 5);
    int64_t first_q, last_q;
    gather_range(a, first_q, last_q);
    gather_scalars(a, first_q, last_q);
    if (per_query == 0) return;

    const int warp = threadIdx.x >> 5, warps = blockDim.x >> 5;
    if ((threadIdx.x & 31) != 0) return;  // one issuing lane per warp; nothing below is a CTA barrier

    const int NS = a.stages;
    const uint32_t bar0 = smem_u32(g_smem) + (uint32_t)(warp * NS) * 8u;
    const uint32_t buf0 = smem_u32(g_smem) + (uint32_t)round_up_dev(warps * NS * 8, 128) + (uint32_t)(warp * NS) * (uint32_t)a.piece;
    for (int s = 0; s < NS; ++s) mbar_init(bar0 + 8u * s, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");

    const int64_t total = (last_q - first_q) * per_query;
    const int64_t first = (int64_t)blockIdx.x * warps + warp;
    const int64_t stride = (int64_t)gridDim.x * warps;
    const int64_t mine = first < total ? (total - first + stride - 1) / stride : 0;

    // prologue: fill the ring
    for (int64_t i = 0; i < NS && i < mine; ++i) {
        Piece p = decode_piece(a, first + i * stride, per_query, first_q);
        mbar_expect_tx(bar0 + 8u * (uint32_t)i, p.bytes);  // 0 bytes (skipped row): the phase completes at once
        if (p.bytes) bulk_g2s(buf0 + (uint32_t)i * (uint32_t)a.piece, p.src, p.bytes, bar0 + 8u * (uint32_t)i);
    }
    int s = 0;
    uint32_t parity = 0;
    for (int64_t i = 0; i < mine; ++i) {
        Piece p = decode_piece(a, first + i * stride, per_query, first_q);
        mbar_wait(bar0 + 8u * s, parity);
        if (p.bytes) bulk_s2g(p.dst, buf0 + (uint32_t)s * (uint32_t)a.piece, p.bytes);
        bulk_commit();
        // refill the stage whose store was committed one iteration ago
        if (i >= 1 && i - 1 + NS < mine) {
            bulk_wait_read<1>();  // all but the newest store have finished reading shared memory
            const int sp = s == 0 ? NS - 1 : s - 1;
            Piece n = decode_piece(a, first + (i - 1 + NS) * stride, per_query, first_q);
            mbar_expect_tx(bar0 + 8u * sp, n.bytes);
            if (n.bytes) bulk_g2s(buf0 + (uint32_t)sp * (uint32_t)a.piece, n.src, n.bytes, bar0 + 8u * sp);
        }
        if (++s == NS) { s = 0; parity ^= 1u; }
    }
    bulk_wait_read<0>();
    if (tl_on) tl_stamp(a.tl, 2, 6);
    if (a.tl) {  // timeline on: the latest finish over all issuing lanes of the grid (diagnostic runs only)
        unsigned long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        atomicMax(a.tl + 3, (13ull << 56) | (t & ((1ull << 56) - 1)));
    }
}

// Register-staged variant: one warp per piece, up to 16 x 16-B loads in flight per lane,
// streaming cache hints (rows are touched once).
__device__ __forceinline__ int4 ld_stream(const int4 *p) {
    int4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.s32 {%0, %1, %2, %3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}
__device__ __forceinline__ void st_stream(int4 *p, const int4 &v) {
    asm volatile("st.global.L1::no_allocate.v4.s32 [%0], {%1, %2, %3, %4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

// DEPTH = 16-byte loads in flight per lane.  16 (96 registers, two 256-thread CTAs per SM, a whole 7 056-byte piece in
// flight per warp) measured better inside the step than 8 (64 registers, four CTAs per SM, 4 KB per warp) at every
// piece size tried (profiles/r01_notes.md): bytes in flight per warp matter more than the number of warps.
template <int DEPTH>
__global__ void __launch_bounds__(256, DEPTH <= 8 ? 4 : 2) k_gather_ldg(GatherArgs a, int per_query) {
    const bool tl_on = a.tl && blockIdx.x == 0 && threadIdx.x == 0;
    if (tl_on) tl_stamp(a.tl, 0, 4);
    pdl_launch_dependents();
    pdl_wait();  // indices and range tags come from the sampling kernel in front of this one
    if (tl_on) tl_stamp(a.tl, 1, 5);
    int64_t first_q, last_q;
    gather_range(a, first_q, last_q);
    gather_scalars(a, first_q, last_q);
    if (per_query == 0) return;
    const int lane = threadIdx.x & 31;
    const int64_t warps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    const int64_t total = (last_q - first_q) * per_query;
    for (int64_t w = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5; w < total; w += warps) {
        Piece p = decode_piece(a, w, per_query, first_q);
        const int4 *src = reinterpret_cast<const int4 *>(p.src);
        int4 *dst = reinterpret_cast<int4 *>(p.dst);
        const int nv = (int)(p.bytes >> 4);
        for (int base = 0; base < nv; base += 32 * DEPTH) {
            int4 v[DEPTH];
#pragma unroll
            for (int j = 0; j < DEPTH; ++j) {
                int e = base + j * 32 + lane;
                if (e < nv) v[j] = ld_stream(src + e);
            }
#pragma unroll
            for (int j = 0; j < DEPTH; ++j) {
                int e = base + j * 32 + lane;
                if (e < nv) st_stream(dst + e, v[j]);
            }
        }
    }
    if (tl_on) tl_stamp(a.tl, 2, 6);
    if (a.tl && lane == 0) {  // timeline on: the latest finish over all warps of the grid (diagnostic runs only)
        unsigned long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        atomicMax(a.tl + 3, (13ull << 56) | (t & ((1ull << 56) - 1)));
    }
}

// ---------------------------------------------------------------------------------------
// Balanced register-staged mover (gather_mode 2).  The batch is one linear space of VB-byte vectors
// (query-major: obs row, next_obs row, extra row) that is cut into equal contiguous shares, one per warp
// of a grid sized to exactly the resident CTA slots (ctas_per_sm x SMs): no wave tail — with one 7 056-byte
// piece per warp the 4 096 pieces of a 512-row batch make 1.73 waves of the 296 slots — and every warp
// finishes together.  A warp walks its share in stages of 32 x DEPTH vectors; with PIPE the loads of stage
// s+1 are issued before the stores of stage s, so a warp always has loads in flight.
// VB = 16: LDG.128 / STG.128.  VB = 32: LDG.256 / STG.256 (sm_100), which is also the only form that takes the
// L2 eviction-priority hints: rows are read once (evict_first), the batch is re-read by the learner (evict_last).
// ---------------------------------------------------------------------------------------
template <int VB> struct GVec;
template <> struct GVec<16> { int4 a; };
template <> struct GVec<32> { int4 a, b; };

__device__ __forceinline__ void gv_load(GVec<16> &v, const uint8_t *p) { v.a = ld_stream(reinterpret_cast<const int4 *>(p)); }
__device__ __forceinline__ void gv_store(uint8_t *p, const GVec<16> &v) { st_stream(reinterpret_cast<int4 *>(p), v.a); }
__device__ __forceinline__ void gv_load(GVec<32> &v, const uint8_t *p) {
    asm volatile("ld.global.nc.L1::no_allocate.L2::evict_first.v8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                 : "=r"(v.a.x), "=r"(v.a.y), "=r"(v.a.z), "=r"(v.a.w), "=r"(v.b.x), "=r"(v.b.y), "=r"(v.b.z), "=r"(v.b.w)
                 : "l"(p));
}
__device__ __forceinline__ void gv_store(uint8_t *p, const GVec<32> &v) {
    asm volatile("st.global.L1::no_allocate.L2::evict_last.v8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"l"(p), "r"(v.a.x),
                 "r"(v.a.y), "r"(v.a.z), "r"(v.a.w), "r"(v.b.x), "r"(v.b.y), "r"(v.b.z), "r"(v.b.w)
                 : "memory");
}

// Position of one lane inside the linear vector space: query q, vector r of the query's V vectors, and the
// ring row k of that query (-1: not this shard's row, nothing to move).
struct VecCursor {
    int64_t q;
    int r;
    int64_t k;
};
__device__ __forceinline__ int64_t gather_row_of(const GatherArgs &a, int64_t q) {
    int64_t k = a.idx[q];
    if (a.skip_foreign && k == -1) return -1;
    k -= a.idx_base;
    return (k < 0 || k >= a.n_rows) ? 0 : k;  // flagged by the scalar pass; keep the copy in bounds
}

template <int VB, int DEPTH>
struct GatherStage {
    GVec<VB> v[DEPTH];
    uint8_t *dst[DEPTH];  // null: nothing to store
};

// Loads the next DEPTH vectors of this lane (vectors g, g + 32, ...), advancing the cursor.
template <int VB, int DEPTH>
__device__ __forceinline__ void gather_stage_load(const GatherArgs &a, GatherStage<VB, DEPTH> &st, VecCursor &c, int64_t &g,
                                                  const int64_t end, const int64_t last_q, const int v0, const int v01, const int V) {
#pragma unroll
    for (int j = 0; j < DEPTH; ++j) {
        st.dst[j] = nullptr;
        if (g < end) {
            if (c.k >= 0) {
                const int cls = c.r < v0 ? 0 : (c.r < v01 ? 1 : 2);
                const int o = c.r - (cls == 0 ? 0 : (cls == 1 ? v0 : v01));
                const RowClass &rc = a.rc[cls];
                gv_load(st.v[j], rc.src + c.k * rc.src_stride + (int64_t)o * VB);
                st.dst[j] = rc.dst + c.q * rc.dst_stride + (int64_t)o * VB;
            }
            g += 32;
            c.r += 32;
            if (c.r >= V) {
                do {
                    c.r -= V;
                    ++c.q;
                } while (c.r >= V);
                c.k = c.q < last_q ? gather_row_of(a, c.q) : -1;
            }
        }
    }
}
template <int VB, int DEPTH>
__device__ __forceinline__ void gather_stage_store(const GatherStage<VB, DEPTH> &st) {
#pragma unroll
    for (int j = 0; j < DEPTH; ++j)
        if (st.dst[j]) gv_store(st.dst[j], st.v[j]);
}

template <int VB, int DEPTH, bool PIPE>
__global__ void __launch_bounds__(256) k_gather_bal(GatherArgs a, int v0, int v1, int v2) {
    const bool tl_on = a.tl && blockIdx.x == 0 && threadIdx.x == 0;
    if (tl_on) tl_stamp(a.tl, 0, 4);
    pdl_launch_dependents();
    pdl_wait();  // indices and range tags come from the sampling kernel in front of this one
    if (tl_on) tl_stamp(a.tl, 1, 5);
    int64_t first_q, last_q;
    gather_range(a, first_q, last_q);
    gather_scalars(a, first_q, last_q);
    const int V = v0 + v1 + v2, v01 = v0 + v1;
    if (V == 0) return;
    const int lane = threadIdx.x & 31;
    const int64_t warps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    const int64_t w = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
    const int64_t total = (last_q - first_q) * V;
    const int64_t share = (((total + warps - 1) / warps) + 31) & ~int64_t(31);  // whole warp-rows of 32 vectors
    const int64_t start = w * share;
    const int64_t end = start + share < total ? start + share : total;
    int64_t g = start + lane;
    if (start < end) {
        VecCursor c;
        c.q = g / V;
        c.r = (int)(g - c.q * V);
        c.q += first_q;
        c.k = (g < end) ? gather_row_of(a, c.q) : -1;
        if (PIPE) {
            GatherStage<VB, DEPTH> s0, s1;
            gather_stage_load(a, s0, c, g, end, last_q, v0, v01, V);
            while (true) {
                const bool more1 = __any_sync(0xffffffffu, g < end);
                if (more1) gather_stage_load(a, s1, c, g, end, last_q, v0, v01, V);
                gather_stage_store(s0);
                if (!more1) break;
                const bool more0 = __any_sync(0xffffffffu, g < end);
                if (more0) gather_stage_load(a, s0, c, g, end, last_q, v0, v01, V);
                gather_stage_store(s1);
                if (!more0) break;
            }
        } else {
            while (__any_sync(0xffffffffu, g < end)) {
                GatherStage<VB, DEPTH> s0;
                gather_stage_load(a, s0, c, g, end, last_q, v0, v01, V);
                gather_stage_store(s0);
            }
        }
    }
    if (tl_on) tl_stamp(a.tl, 2, 6);
    if (a.tl && lane == 0) {  // timeline on: the latest finish over all warps of the grid (diagnostic runs only)
        unsigned long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        atomicMax(a.tl + 3, (13ull << 56) | (t & ((1ull << 56) - 1)));
    }
}

// Gather + widen: the learner turns every sampled uint8 frame stack into
// float32 before the first conv (`torch.FloatTensor(states).to(device)`, ApeX.py:50,53 — a 4x
// inflation done on the host in the reference).  This variant writes float32 / bfloat16 / float16
// rows (value * scale) straight from the ring: pieces arrive by TMA bulk copy as above, the warp
// that owns the piece converts them out of shared memory (lane l takes word k*32+l, so the 16-byte
// stores of a warp cover 512 contiguous bytes) — the uint8 batch is never written or re-read.
enum GatherCastDtype : int { CAST_F32 = 0, CAST_BF16 = 1, CAST_F16 = 2 };

__device__ __forceinline__ float u8_to_f32(uint32_t word, int byte) {
    // 0x4B000000 | b is the float 8388608 + b: exact for 0..255, no I2F
    return __uint_as_float(0x4B000000u | ((word >> (8 * byte)) & 0xffu)) - 8388608.0f;
}

__global__ void __launch_bounds__(GATHER_MAX_WARPS * 32) k_gather_cast(GatherArgs a, int per_query, int dtype, float scale) {
    extern __shared__ __align__(128) uint8_t g_smem[];
    int64_t first_q, last_q;
    gather_range(a, first_q, last_q);
    gather_scalars(a, first_q, last_q);
    if (per_query == 0) return;
    const int warp = threadIdx.x >> 5, warps = blockDim.x >> 5, lane = threadIdx.x & 31;
    const int NS = a.stages;
    const uint32_t bar0 = smem_u32(g_smem) + (uint32_t)(warp * NS) * 8u;
    const uint32_t data0 = (uint32_t)round_up_dev(warps * NS * 8, 128) + (uint32_t)(warp * NS) * (uint32_t)a.piece;
    if (lane == 0) {
        for (int s = 0; s < NS; ++s) mbar_init(bar0 + 8u * s, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();
    const int64_t total = (last_q - first_q) * per_query;
    const int64_t first = (int64_t)blockIdx.x * warps + warp;
    const int64_t stride = (int64_t)gridDim.x * warps;
    const int64_t mine = first < total ? (total - first + stride - 1) / stride : 0;
    const int esz = dtype == CAST_F32 ? 4 : 2;
    if (lane == 0) {
        for (int64_t i = 0; i < NS && i < mine; ++i) {
            Piece p = decode_piece(a, first + i * stride, per_query, first_q);
            mbar_expect_tx(bar0 + 8u * (uint32_t)i, p.bytes);
            if (p.bytes) bulk_g2s(smem_u32(g_smem) + data0 + (uint32_t)i * (uint32_t)a.piece, p.src, p.bytes, bar0 + 8u * (uint32_t)i);
        }
    }
    int s = 0;
    uint32_t parity = 0;
    for (int64_t i = 0; i < mine; ++i) {
        Piece p = decode_piece(a, first + i * stride, per_query, first_q);
        mbar_wait(bar0 + 8u * s, parity);  // every lane waits: the data must be visible to all of them
        if (p.bytes) {
            const uint32_t *src = reinterpret_cast<const uint32_t *>(g_smem + data0 + (size_t)s * a.piece);
            uint8_t *out = a.cast_out[p.cls] + p.elem * esz;  // same element offset, esz bytes per element
            const int words = (int)(p.bytes >> 2);
            for (int w = lane; w < words; w += 32) {
                const uint32_t x = src[w];
                const float f0 = u8_to_f32(x, 0) * scale, f1 = u8_to_f32(x, 1) * scale, f2 = u8_to_f32(x, 2) * scale,
                            f3 = u8_to_f32(x, 3) * scale;
                if (dtype == CAST_F32) {
                    reinterpret_cast<float4 *>(out)[w] = make_float4(f0, f1, f2, f3);
                } else if (dtype == CAST_BF16) {
                    __nv_bfloat162 lo = __floats2bfloat162_rn(f0, f1), hi = __floats2bfloat162_rn(f2, f3);
                    reinterpret_cast<uint2 *>(out)[w] = make_uint2(*reinterpret_cast<uint32_t *>(&lo), *reinterpret_cast<uint32_t *>(&hi));
                } else {
                    __half2 lo = __floats2half2_rn(f0, f1), hi = __floats2half2_rn(f2, f3);
                    reinterpret_cast<uint2 *>(out)[w] = make_uint2(*reinterpret_cast<uint32_t *>(&lo), *reinterpret_cast<uint32_t *>(&hi));
                }
            }
        }
        __syncwarp();  // all lanes are done with this stage
        if (lane == 0 && i + NS < mine) {
            Piece n = decode_piece(a, first + (i + NS) * stride, per_query, first_q);
            mbar_expect_tx(bar0 + 8u * s, n.bytes);
            if (n.bytes) bulk_g2s(smem_u32(g_smem) + data0 + (uint32_t)s * (uint32_t)a.piece, n.src, n.bytes, bar0 + 8u * s);
        }
        if (++s == NS) { s = 0; parity ^= 1u; }
    }
}

// Gather + widen into CHANNELS-LAST batches: the learner's first layer is a cuDNN convolution
// (model.py:33, fed by ApeX.py:50,53 / origin_repo/learner.py:102-110), which runs fastest on NHWC tensors.  A stored
// observation is C planes of `plane` bytes (uint8[4,84,84]: C = 4, plane = 7 056); the output row q holds, for every
// pixel p, the C widened channel values next to each other: out[q][p][c] = scale * ring[k][c][p].  One thread takes 4
// consecutive pixels: one 32-bit load per plane (a warp reads 128 contiguous bytes of each plane), four C-wide vector
// stores (a warp writes 2 KB contiguous at float32).  C == 4 only (the Atari frame stack).
template <typename OutVec4>
__device__ __forceinline__ OutVec4 cast4(float a, float b, float c, float d);
template <>
__device__ __forceinline__ float4 cast4<float4>(float a, float b, float c, float d) { return make_float4(a, b, c, d); }
struct BF16x4 { __nv_bfloat162 lo, hi; };
struct F16x4 { __half2 lo, hi; };
template <>
__device__ __forceinline__ BF16x4 cast4<BF16x4>(float a, float b, float c, float d) {
    BF16x4 r;
    r.lo = __floats2bfloat162_rn(a, b);
    r.hi = __floats2bfloat162_rn(c, d);
    return r;
}
template <>
__device__ __forceinline__ F16x4 cast4<F16x4>(float a, float b, float c, float d) {
    F16x4 r;
    r.lo = __floats2half2_rn(a, b);
    r.hi = __floats2half2_rn(c, d);
    return r;
}

template <typename OutVec4>
__global__ void __launch_bounds__(256) k_gather_cast_nhwc(GatherArgs a, int plane, float scale) {
    int64_t first_q, last_q;
    gather_range(a, first_q, last_q);
    gather_scalars(a, first_q, last_q);
    const int quads = plane >> 2;                       // 4-pixel groups per plane
    const int64_t per_class = (last_q - first_q) * quads;
    const int64_t tid = blockIdx.x * (int64_t)blockDim.x + threadIdx.x, nthreads = (int64_t)gridDim.x * blockDim.x;
#pragma unroll 1
    for (int c = 0; c < 2; ++c) {
        if (!a.cast_out[c]) continue;
        const RowClass &rc = a.rc[c];
        OutVec4 *out = reinterpret_cast<OutVec4 *>(a.cast_out[c]);
        for (int64_t w = tid; w < per_class; w += nthreads) {
            const int64_t q = first_q + w / quads;
            const int g = (int)(w % quads);
            const int64_t k = gather_row_of(a, q);
            if (k < 0) continue;
            const uint32_t *src = reinterpret_cast<const uint32_t *>(rc.src + k * rc.src_stride) + g;
            const uint32_t p0 = __ldg(src), p1 = __ldg(src + quads), p2 = __ldg(src + 2 * quads), p3 = __ldg(src + 3 * quads);
            OutVec4 *dst = out + (q * plane + 4 * (int64_t)g);  // one OutVec4 per pixel
#pragma unroll
            for (int px = 0; px < 4; ++px)
                dst[px] = cast4<OutVec4>(u8_to_f32(p0, px) * scale, u8_to_f32(p1, px) * scale, u8_to_f32(p2, px) * scale,
                                         u8_to_f32(p3, px) * scale);
        }
    }
}

// Fallback mover for row classes whose size is not a multiple of 16 B (small non-Atari
// observations, e.g. CartPole's float32[4]): byte-granular, one warp per row.
__global__ void k_gather_bytes(GatherArgs a) {
    const int lane = threadIdx.x & 31;
    const int64_t warp = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t w = warp; w < a.batch * 3; w += nwarps) {
        int64_t q = w / 3;
        const RowClass &rc = a.rc[w - q * 3];
        if (!rc.dst || rc.row_bytes == 0) continue;
        int64_t k = a.idx[q];
        if (a.skip_foreign && k == -1) continue;
        k -= a.idx_base;
        if (k < 0 || k >= a.n_rows) k = 0;
        const uint8_t *s = rc.src + k * rc.src_stride;
        uint8_t *d = rc.dst + q * rc.dst_stride;
        for (int64_t i = lane; i < rc.row_bytes; i += 32) d[i] = s[i];
    }
}

// Ring insert: n rows land in slots (start + j) mod ring_size.  Contiguous in the ring
// except for one wrap, so this is a streaming copy: 16-B vectors, grid-stride.
struct AddArgs {
    const uint8_t *src[3];  // obs, next, extra batches (unpadded rows), may be null
    uint8_t *dst[3];        // ring bases
    int64_t row_bytes[3];
    int64_t dst_stride[3];
    const int64_t *act; const float *rew; const float *done;
    int64_t *ring_act; float *ring_rew; float *ring_done;
    int64_t start, ring_size, n;
};

__global__ void k_ring_add(AddArgs a) {
    const int64_t tid = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    const int64_t nthreads = (int64_t)gridDim.x * blockDim.x;
    for (int64_t j = tid; j < a.n; j += nthreads) {
        int64_t k = a.start + j;
        if (k >= a.ring_size) k -= a.ring_size;
        if (a.act) a.ring_act[k] = a.act[j];
        if (a.rew) a.ring_rew[k] = a.rew[j];
        if (a.done) a.ring_done[k] = a.done[j];
    }
    for (int c = 0; c < 3; ++c) {
        if (!a.src[c] || a.row_bytes[c] == 0) continue;
        const int64_t rb = a.row_bytes[c];
        if ((rb & 15) == 0 && ((uintptr_t)a.src[c] & 15) == 0) {
            const int64_t vec_per_row = rb >> 4;
            const int64_t total = a.n * vec_per_row;
            for (int64_t e = tid; e < total; e += nthreads) {
                int64_t j = e / vec_per_row, o = e - j * vec_per_row;
                int64_t k = a.start + j;
                if (k >= a.ring_size) k -= a.ring_size;
                int4 x = __ldcs(reinterpret_cast<const int4 *>(a.src[c] + j * rb) + o);
                reinterpret_cast<int4 *>(a.dst[c] + k * a.dst_stride[c])[o] = x;
            }
        } else {
            const int64_t total = a.n * rb;
            for (int64_t e = tid; e < total; e += nthreads) {
                int64_t j = e / rb, o = e - j * rb;
                int64_t k = a.start + j;
                if (k >= a.ring_size) k -= a.ring_size;
                a.dst[c][k * a.dst_stride[c] + o] = a.src[c][j * rb + o];
            }
        }
    }
}

}  // namespace perb
