// aql_kernels.cuh — K6: AQL candidate-action Q evaluation + argmax (model.py:294-335,
// NoisyLinear.forward model.py:132-143), fused so that the [B, K, 128] concat tensor the
// reference materialises (model.py:301-304, 131 MB at B=256, K=1000) never exists.
//
// Per candidate (state b, candidate k), float32 in the reference:
//     e  = relu(Wa2 . relu(Wa1 . a_bk + ba1) + ba2)          action_out, model.py:252-257 (continuous)
//          relu(Wa1 . a_bk + ba1)                             model.py:260-263 (discrete, Linear(1,64))
//     h  = relu(W1[:, :64] . e  +  W1[:, 64:] . f_b + b1)     advantage1 on cat[e, f_b], model.py:303-304,273
//     q  = w2 . h + b2                                        advantage2, model.py:274
// The state half  sv_b = W1[:, 64:] . f_b + b1  is per state and is hoisted (k_aql_state_half).
//
// One CTA owns a tile of 128 candidates (TMEM lane = candidate).  The two dense layers run on
// the 5th-generation tensor cores: tcgen05.mma kind::tf32, M=128, N=64, K=8, operands in
// shared memory in the K-major 128-byte-swizzled canonical layout, accumulators in TMEM.
// float32 parity (the reference is cuBLAS SGEMM without TF32) comes from the 3xTF32 split:
// x = hi + lo with both parts representable in TF32, and hi.hi + hi.lo + lo.hi accumulated in
// fp32, which leaves an error of a few 2^-22 — inside the 1e-5 bar.  Weights are split and
// swizzled once (k_aql_prep_weights) into a 96 KB image that each CTA pulls in with one TMA
// bulk copy; activations are split on the fly by the thread that owns the row and handed to
// the MMAs through TMEM.
#pragma once
#include "common.cuh"
#include "gather_kernels.cuh"  // mbarrier / bulk-copy wrappers

namespace perb {

constexpr int AQL_TILE = 128;                 // candidates per tile (UMMA M)
constexpr int AQL_H = 64;                     // a_out_unit = feature_out_unit = hidden of advantage1
constexpr int AQL_E1 = 128;                   // hidden of action_out (continuous)
constexpr int AQL_MAX_A = 8;
// One K-atom (32 input columns) of a weight matrix: 128 rows of 128 bytes — rows 0..63 hold the
// TF32 hi part of output row n, rows 64..127 the lo part, so that ONE N=128 MMA gives
// a_hi.W_hi (accumulator columns 0..63) and a_hi.W_lo (columns 64..127), and an N=64 MMA on the
// same descriptor gives a_lo.W_hi.
constexpr int AQL_ATOM_W = 2 * AQL_H * 128;   // 16 KB
constexpr int AQL_IMG_W2 = 0;                 // Wa2: 4 atoms
constexpr int AQL_IMG_W1 = 4 * AQL_ATOM_W;    // W1[:, :64]: 2 atoms
constexpr int AQL_IMG_BYTES = 6 * AQL_ATOM_W;                  // 98 304
constexpr int AQL_SMEM_BYTES = AQL_IMG_BYTES + 1024 + 1024;    // weight image + small area + alignment slack (activations live in TMEM)
constexpr int AQL_CTAS_PER_SM = 2;            // two CTAs share an SM: 2 x 256 TMEM columns, 2 x 98 KB shared memory
constexpr int AQL_COMPUTE_THREADS = 256;      // two threads per candidate row: each owns half of the columns
constexpr int AQL_ISSUER_WARP = AQL_COMPUTE_THREADS / 32;
constexpr int AQL_THREADS = AQL_COMPUTE_THREADS + 32;  // + one warp whose lane 0 issues every MMA

struct AqlSmall {  // small parameters; passed to the kernel by value, i.e. through the constant bank
    float wa1[AQL_E1 * AQL_MAX_A];  // continuous: [128, A]; discrete: [64, 1]
    float ba1[AQL_E1];
    float ba2[AQL_H];
    float w2[AQL_H];
    float b2;
};

__device__ __forceinline__ float tf32_hi(float x) {
    uint32_t r;
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
    return __uint_as_float(r);
}

// Round-to-nearest (ties away, like cvt.rna.tf32) on the bit pattern: two integer instructions
// where the cvt expands to a NaN-guarded sequence of five.  Inputs here are finite (post-ReLU
// activations); Inf stays Inf.
__device__ __forceinline__ float tf32_rn_bits(float x) { return __uint_as_float((__float_as_uint(x) + 0x1000u) & 0xffffe000u); }
__device__ __forceinline__ void tf32_split(float x, float &hi, float &lo) {
    hi = tf32_rn_bits(x);
    lo = tf32_rn_bits(x - hi);
}

// Byte offset of 16-byte chunk `c16` (0..7) of row `r` inside one K-atom of the canonical
// K-major SWIZZLE_128B layout: 8-row groups of 1024 B, chunk index XOR-ed with the row in group.
__device__ __forceinline__ uint32_t sw128(int r, int c16) {
    return (uint32_t)((r >> 3) * 1024 + (r & 7) * 128 + ((c16 ^ (r & 7)) << 4));
}

// UMMA shared-memory descriptor, K-major, SWIZZLE_128B: start >> 4, LBO = 1 (unused),
// SBO = 1024 B between 8-row groups, version 1, layout type 2.
__device__ __forceinline__ uint64_t umma_desc(uint32_t smem_addr) {
    return (uint64_t)((smem_addr >> 4) & 0x3fff) | ((uint64_t)1 << 16) | ((uint64_t)(1024 >> 4) << 32) | ((uint64_t)1 << 46) |
           ((uint64_t)2 << 61);
}
// kind::tf32, D = f32, A/B = tf32 K-major, M = 128, N = n
__host__ __device__ constexpr uint32_t aql_idesc(uint32_t n) {
    return (1u << 4) | (2u << 7) | (2u << 10) | ((n >> 3) << 17) | ((128u >> 4) << 24);
}

__device__ __forceinline__ void umma_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

#define AQL_LD32_REGS(r)                                                                                                           \
    "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]),        \
        "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]),       \
        "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]),       \
        "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
__device__ __forceinline__ void tmem_ld32_nowait(uint32_t taddr, float *v) {
    uint32_t *r = reinterpret_cast<uint32_t *>(v);
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
        : AQL_LD32_REGS(r)
        : "r"(taddr));
}
// This thread's 32 columns of the two partial accumulators (AQL_H columns apart), summed.
__device__ __forceinline__ void tmem_ld32_sum2(uint32_t taddr, float *acc) {
    float t[32];
    tmem_ld32_nowait(taddr, acc);
    tmem_ld32_nowait(taddr + AQL_H, t);
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int i = 0; i < 32; ++i) acc[i] += t[i];
}

// tcgen05.st: this thread's TMEM lane, 16 consecutive 32-bit columns starting at taddr
__device__ __forceinline__ void tmem_st16(uint32_t taddr, const float *v) {
    const uint32_t *r = reinterpret_cast<const uint32_t *>(v);
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};" ::"r"(taddr),
        "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]), "r"(r[10]),
        "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15])
        : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

// A operand from TMEM (lane = row, 8 consecutive columns = the K slice), B from shared memory
template <uint32_t N>
__device__ __forceinline__ void umma_tf32_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t db, uint32_t accumulate) {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n\t"
        "}\n" ::"r"(tmem_d),
        "r"(tmem_a), "l"(db), "r"(aql_idesc(N)), "r"(accumulate)
        : "memory");
}

// One K-atom (32 columns) of the 3xTF32 product with A in TMEM, 8 MMAs: per K = 8 slice
//   D[:, 0:128] (+)= a_hi . [W_hi ; W_lo]^T        (N = 128)
//   D[:, 0:64]   += a_lo . W_hi^T                  (N = 64, same descriptor: rows 0..63 of the atom)
// so columns 0..63 hold hi.hi + lo.hi and columns 64..127 hold hi.lo; the epilogue adds the two.
__device__ __forceinline__ void issue_atom_ts(uint32_t tmem_d, uint32_t a_hi, uint32_t a_lo, uint32_t w_atom, bool first) {
#pragma unroll
    for (int k4 = 0; k4 < 4; ++k4) {
        const uint64_t b = umma_desc(w_atom + k4 * 32);
        umma_tf32_ts<128>(tmem_d, a_hi + k4 * 8, b, (first && k4 == 0) ? 0u : 1u);
        umma_tf32_ts<64>(tmem_d, a_lo + k4 * 8, b, 1u);
    }
}

// ---- one-off preparation ---------------------------------------------------------------
// Splits Wa2 [64,128] and W1[:, :64] [64,64] into TF32 hi / lo and lays them out as the
// swizzled shared-memory image.  One thread per (matrix row n, 16-byte chunk).
__global__ void k_aql_prep_weights(const float *wa2, const float *wadv1, uint8_t *img) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    // Wa2: 64 rows x 32 chunks; W1a: 64 rows x 16 chunks
    const int n2 = AQL_H * 32, n1 = AQL_H * 16;
    if (t >= n2 + n1) return;
    const bool second = t >= n2;
    const int u = second ? t - n2 : t;
    const int chunks = second ? 16 : 32;
    const int n = u / chunks, cc = u % chunks;
    const int kb = cc >> 3, c16 = cc & 7;
    float hi[4], lo[4];
    for (int e = 0; e < 4; ++e) {
        const int k = cc * 4 + e;
        const float x = second ? wadv1[n * (2 * AQL_H) + k] : (wa2 ? wa2[n * AQL_E1 + k] : 0.0f);
        hi[e] = tf32_hi(x);
        lo[e] = tf32_hi(x - hi[e]);
    }
    uint8_t *atom = img + (second ? AQL_IMG_W1 : AQL_IMG_W2) + kb * AQL_ATOM_W;
    *reinterpret_cast<float4 *>(atom + sw128(n, c16)) = make_float4(hi[0], hi[1], hi[2], hi[3]);
    *reinterpret_cast<float4 *>(atom + sw128(AQL_H + n, c16)) = make_float4(lo[0], lo[1], lo[2], lo[3]);
}

// sv[b][j] = b1[j] + sum_i W1[j][64 + i] * f[b][i]   (f is post-ReLU already, model.py:244-249)
// One thread per output; the 16 + 16 float4 loads are independent, the FMA chain runs in column order.
__global__ void k_aql_state_half(const float *wadv1, const float *badv1, const float *f, int64_t B, float *sv) {
    pdl_launch_dependents();  // the Q kernel may start its prologue and first layers now (it waits before reading sv)
    const int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (t >= B * AQL_H) return;
    const int64_t b = t / AQL_H;
    const int j = (int)(t % AQL_H);
    const float4 *w4 = reinterpret_cast<const float4 *>(wadv1 + j * (2 * AQL_H) + AQL_H);
    const float4 *f4 = reinterpret_cast<const float4 *>(f + b * AQL_H);
    float4 w[AQL_H / 4], x[AQL_H / 4];
#pragma unroll
    for (int i = 0; i < AQL_H / 4; ++i) {
        w[i] = w4[i];
        x[i] = f4[i];
    }
    float acc = 0.0f;
#pragma unroll
    for (int i = 0; i < AQL_H / 4; ++i) {
        acc = fmaf(w[i].x, fmaxf(x[i].x, 0.0f), acc);
        acc = fmaf(w[i].y, fmaxf(x[i].y, 0.0f), acc);
        acc = fmaf(w[i].z, fmaxf(x[i].z, 0.0f), acc);
        acc = fmaf(w[i].w, fmaxf(x[i].w, 0.0f), acc);
    }
    sv[t] = acc + badv1[j];
}

// ---- the fused forward ---------------------------------------------------------------------
struct AqlArgs {
    const uint8_t *img;      // swizzled weight image (AQL_IMG_BYTES)
    const float *sv;         // [B, 64]
    const void *a_mu;        // continuous: float [B*K, A]; discrete: int64 [B*K]
    int64_t K, total;        // candidates per state, B*K
    float *q;                // [B*K]
    unsigned long long *trace;  // diagnostics (may be null): CTA 0 / thread 0 accumulates clock cycles per phase
};

// TMEM columns of one CTA (256 allocated, so two CTAs fit an SM):
//   0..127   layer-1 activations, a ring of two K-atoms: slot s = [hi 32 | lo 32] at 64 s;
//            later the layer-2 activations (A operand of layer 3): hi 0..63, lo 64..127
//   128..255 accumulator of the layer in flight (two partials, see issue_atom_ts)
constexpr uint32_t AQL_T_RING = 0, AQL_T_A2H = 0, AQL_T_A2L = 64, AQL_T_D = 128, AQL_T_COLS = 256;
// mbarriers
enum { AQL_BAR_W = 0, AQL_BAR_FULL = 1, AQL_BAR_EMPTY = 3, AQL_BAR_FULL2 = 5, AQL_BAR_DONE2 = 6, AQL_BAR_DONE3 = 7, AQL_BARS = 8 };

// A = action dimension (continuous) at compile time so that layer 1 unrolls into FFMAs whose
// weight operands come straight from the constant bank (`small` is a __grid_constant__ kernel
// parameter, 5 KB); CONT = false is the discrete path (Linear(1,64), no second action layer).
//
// Warp roles: warps 0..7 (two threads per candidate row) run the CUDA-core parts — layer 1, the
// two epilogues — and hand activations over through TMEM (tcgen05.st; the MMAs take A from
// TMEM, B = weights from the swizzled image in shared memory).  Warp 8 does nothing but issue
// MMAs: it waits on a "staged" mbarrier (256 arrivals), issues the 8 MMAs of that K-atom and
// commits to the barrier the consumers of the result wait on.  Nobody else ever stalls behind an
// MMA issue, and with two CTAs per SM the CUDA-core phases of one overlap the MMAs of the other.
template <int A, bool CONT>
__global__ void __launch_bounds__(AQL_THREADS, AQL_CTAS_PER_SM) k_aql_q(const AqlArgs a, const __grid_constant__ AqlSmall small) {
    extern __shared__ uint8_t aql_raw[];
    uint8_t *sm = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(aql_raw) + 1023) & ~uintptr_t(1023));
    float *s_part = reinterpret_cast<float *>(sm + AQL_IMG_BYTES);          // [128] partial dot of the upper column half
    uint64_t *bars = reinterpret_cast<uint64_t *>(s_part + AQL_TILE);
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(bars + AQL_BARS);
    const uint32_t bar0 = smem_u32(bars);
#define AQL_BAR(i) (bar0 + 8u * (uint32_t)(i))

    const long long t_entry = clock64();
    const int tid = threadIdx.x;
    const int r = tid & (AQL_TILE - 1);   // row of the tile = TMEM lane
    const int half = (tid >> 7) & 1;      // which half of the columns this thread owns
    const int warp = tid >> 5;
    if (tid == 0) {
        mbar_init(AQL_BAR(AQL_BAR_W), 1);
        for (int s = 0; s < 2; ++s) {
            mbar_init(AQL_BAR(AQL_BAR_FULL + s), AQL_COMPUTE_THREADS);
            mbar_init(AQL_BAR(AQL_BAR_EMPTY + s), 1);
        }
        mbar_init(AQL_BAR(AQL_BAR_FULL2), AQL_COMPUTE_THREADS);
        mbar_init(AQL_BAR(AQL_BAR_DONE2), 1);
        mbar_init(AQL_BAR(AQL_BAR_DONE3), 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        mbar_expect_tx(AQL_BAR(AQL_BAR_W), AQL_IMG_BYTES);
        bulk_g2s(smem_u32(sm), a.img, AQL_IMG_BYTES, AQL_BAR(AQL_BAR_W));
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "n"(AQL_T_COLS));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *tmem_slot;
    const int64_t tiles = (a.total + AQL_TILE - 1) / AQL_TILE;
    pdl_launch_dependents();  // the argmax grid may become resident; it waits for this grid to retire before reading q

    if (warp == AQL_ISSUER_WARP) {
        // ---- MMA issuer ------------------------------------------------------------------------
        const bool leader = (tid & 31) == 0;
        const uint32_t w2 = smem_u32(sm + AQL_IMG_W2), w1 = smem_u32(sm + AQL_IMG_W1);
        mbar_wait(AQL_BAR(AQL_BAR_W), 0);
        uint32_t it = 0;
        for (int64_t tile = blockIdx.x; tile < tiles; tile += gridDim.x, ++it) {
            if (CONT) {
#pragma unroll 1
                for (int kb = 0; kb < 4; ++kb) {
                    const int s = kb & 1;
                    mbar_wait(AQL_BAR(AQL_BAR_FULL + s), (uint32_t)(kb >> 1));  // a slot fills twice per tile
                    tc_fence_after();
                    if (leader) {
                        issue_atom_ts(tmem + AQL_T_D, tmem + AQL_T_RING + s * 64, tmem + AQL_T_RING + s * 64 + 32, w2 + kb * AQL_ATOM_W,
                                      kb == 0);
                        if (kb < 2) umma_commit(AQL_BAR(AQL_BAR_EMPTY + s));  // slot may be refilled once these MMAs are done
                        else if (kb == 3) umma_commit(AQL_BAR(AQL_BAR_DONE2));
                    }
                    __syncwarp();
                }
            }
            mbar_wait(AQL_BAR(AQL_BAR_FULL2), it & 1u);
            tc_fence_after();
            if (leader) {
                issue_atom_ts(tmem + AQL_T_D, tmem + AQL_T_A2H, tmem + AQL_T_A2L, w1, true);
                issue_atom_ts(tmem + AQL_T_D, tmem + AQL_T_A2H + 32, tmem + AQL_T_A2L + 32, w1 + AQL_ATOM_W, false);
                umma_commit(AQL_BAR(AQL_BAR_DONE3));
            }
            __syncwarp();
        }
    } else {
        // ---- CUDA-core side: 256 threads ---------------------------------------------------------
        const uint32_t lane_base = tmem + ((uint32_t)((warp & 3) * 32) << 16);  // this warp's 32 TMEM lanes, column 0
        uint32_t it = 0;
        float act_next[A];  // continuous actions of this thread's row in the tile about to be processed
        if (CONT) {
            const int64_t c0 = (int64_t)blockIdx.x * AQL_TILE + r;
            const float *ap0 = static_cast<const float *>(a.a_mu) + (c0 < a.total ? c0 : a.total - 1) * A;
#pragma unroll
            for (int j = 0; j < A; ++j) act_next[j] = ap0[j];
        }
        if (a.trace && blockIdx.x == 0 && tid == 0) a.trace[6] += (unsigned long long)(clock64() - t_entry);
        for (int64_t tile = blockIdx.x; tile < tiles; tile += gridDim.x, ++it) {
            const uint32_t tp = it & 1u;
            long long t_mark = clock64();
#define AQL_TRACE(slot)                                                     \
    do {                                                                    \
        if (a.trace && blockIdx.x == 0 && tid == 0) {                       \
            const long long now__ = clock64();                              \
            a.trace[slot] += (unsigned long long)(now__ - t_mark);          \
            t_mark = now__;                                                 \
        }                                                                   \
    } while (0)
            const int64_t c = tile * AQL_TILE + r;
            const bool live = c < a.total;
            const int64_t ci = live ? c : a.total - 1;  // dead rows recompute the last candidate; never stored
            float hi[16], lo[16];
            if (CONT) {
                // layer 1 on CUDA cores (K = A): h1 = relu(Wa1 . a + ba1), one 32-column K-atom at a time (16 columns per
                // thread) into ring slot kb & 1; the issuer starts on an atom as soon as all 256 threads have stored it
                float act[A];
#pragma unroll
                for (int j = 0; j < A; ++j) act[j] = act_next[j];  // requested one tile ago
                {   // this row's action of the NEXT tile: its round trip hides behind this tile's three layers
                    const int64_t cn = (tile + gridDim.x) * AQL_TILE + r;
                    const int64_t cin = cn < a.total ? cn : a.total - 1;
                    const float *apn = static_cast<const float *>(a.a_mu) + cin * A;
#pragma unroll
                    for (int j = 0; j < A; ++j) act_next[j] = apn[j];
                }
#pragma unroll
                for (int kb = 0; kb < 4; ++kb) {
#pragma unroll
                    for (int e = 0; e < 16; ++e) {
                        const int o = kb * 32 + half * 16 + e;
                        float s = small.ba1[o];
#pragma unroll
                        for (int j = 0; j < A; ++j) s = fmaf(small.wa1[o * A + j], act[j], s);
                        tf32_split(fmaxf(s, 0.0f), hi[e], lo[e]);
                    }
                    const int s = kb & 1;
                    if (kb >= 2) {  // the MMAs that read atom kb - 2 out of this slot must have retired
                        mbar_wait(AQL_BAR(AQL_BAR_EMPTY + s), tp);
                        tc_fence_after();
                    }
                    tmem_st16(lane_base + AQL_T_RING + s * 64 + half * 16, hi);
                    tmem_st16(lane_base + AQL_T_RING + s * 64 + 32 + half * 16, lo);
                    tmem_st_wait();
                    tc_fence_before();
                    mbar_arrive(AQL_BAR(AQL_BAR_FULL + s));
                }
                AQL_TRACE(0);  // layer 1
                mbar_wait(AQL_BAR(AQL_BAR_DONE2), tp);
                tc_fence_after();
                AQL_TRACE(1);  // wait for layer-2 MMAs
                // layer 2 epilogue: e = relu(D2 + ba2) -> A operand of layer 3 (K = 64), 32 columns per thread
                float acc[32];
                tmem_ld32_sum2(lane_base + AQL_T_D + half * 32, acc);
#pragma unroll
                for (int g = 0; g < 2; ++g) {
#pragma unroll
                    for (int e = 0; e < 16; ++e) {
                        tf32_split(fmaxf(acc[g * 16 + e] + small.ba2[half * 32 + g * 16 + e], 0.0f), hi[e], lo[e]);
                    }
                    tmem_st16(lane_base + AQL_T_A2H + half * 32 + g * 16, hi);
                    tmem_st16(lane_base + AQL_T_A2L + half * 32 + g * 16, lo);
                }
            } else {
                // discrete: e = relu(Wa1[:,0] * float(a) + ba1), 64 wide (model.py:319-320); 32 outputs per thread
                const float av = (float)static_cast<const int64_t *>(a.a_mu)[ci];
#pragma unroll
                for (int g = 0; g < 2; ++g) {
#pragma unroll
                    for (int e = 0; e < 16; ++e) {
                        const int o = half * 32 + g * 16 + e;
                        tf32_split(fmaxf(fmaf(small.wa1[o], av, small.ba1[o]), 0.0f), hi[e], lo[e]);
                    }
                    tmem_st16(lane_base + AQL_T_A2H + half * 32 + g * 16, hi);
                    tmem_st16(lane_base + AQL_T_A2L + half * 32 + g * 16, lo);
                }
            }
            tmem_st_wait();
            tc_fence_before();
            mbar_arrive(AQL_BAR(AQL_BAR_FULL2));
            AQL_TRACE(2);  // layer-2 epilogue (or discrete layer 1)
            const float4 *svb = reinterpret_cast<const float4 *>(a.sv + (ci / a.K) * AQL_H + half * 32);
            float4 s4[8];
            if (it == 0) pdl_wait();  // sv comes from the state-half kernel launched just before this one
#pragma unroll
            for (int i4 = 0; i4 < 8; ++i4) s4[i4] = svb[i4];  // in flight while the layer-3 MMAs run
            mbar_wait(AQL_BAR(AQL_BAR_DONE3), tp);
            tc_fence_after();
            AQL_TRACE(3);  // wait for layer-3 MMAs
            // layer 3 epilogue: q = w2 . relu(D3 + sv_b) + b2; each thread sums its 32 columns
            float qv = 0.0f;
            {
                float acc[32];
                tmem_ld32_sum2(lane_base + AQL_T_D + half * 32, acc);
#pragma unroll
                for (int i4 = 0; i4 < 8; ++i4) {
                    qv = fmaf(small.w2[half * 32 + i4 * 4 + 0], fmaxf(acc[i4 * 4 + 0] + s4[i4].x, 0.0f), qv);
                    qv = fmaf(small.w2[half * 32 + i4 * 4 + 1], fmaxf(acc[i4 * 4 + 1] + s4[i4].y, 0.0f), qv);
                    qv = fmaf(small.w2[half * 32 + i4 * 4 + 2], fmaxf(acc[i4 * 4 + 2] + s4[i4].z, 0.0f), qv);
                    qv = fmaf(small.w2[half * 32 + i4 * 4 + 3], fmaxf(acc[i4 * 4 + 3] + s4[i4].w, 0.0f), qv);
                }
            }
            if (half) s_part[r] = qv;
            // the next tile overwrites the TMEM activations and the accumulator: this thread's loads are complete
            // (wait::ld) and ordered before its next "staged" arrival by the fence
            tc_fence_before();
            asm volatile("bar.sync 1, %0;" ::"n"(AQL_COMPUTE_THREADS) : "memory");
            if (!half && live) a.q[c] = (qv + s_part[r]) + small.b2;
            AQL_TRACE(4);  // layer-3 epilogue
            if (a.trace && blockIdx.x == 0 && tid == 0) a.trace[5] += 1;
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "n"(AQL_T_COLS));
    if (a.trace && blockIdx.x == 0 && tid == 0) a.trace[7] += (unsigned long long)(clock64() - t_entry);
#undef AQL_BAR
}

// argmax over the K candidates of each state, one CTA of AQL_ARGMAX_THREADS per state; the first
// maximum wins (torch.argmax, model.py:330-332).
constexpr int AQL_ARGMAX_THREADS = 128;
__device__ __forceinline__ void argmax_merge(float &bv, int64_t &bi, float ov, int64_t oi) {
    if (ov > bv || (ov == bv && oi < bi)) {
        bv = ov;
        bi = oi;
    }
}
__global__ void __launch_bounds__(AQL_ARGMAX_THREADS) k_aql_argmax(const float *q, int64_t B, int64_t K, int64_t *out) {
    __shared__ float s_v[AQL_ARGMAX_THREADS / 32];
    __shared__ int64_t s_i[AQL_ARGMAX_THREADS / 32];
    const int64_t b = blockIdx.x;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const float *row = q + b * K;
    pdl_wait();  // q is complete and visible once the Q kernel has retired
    float bv = -INFINITY;
    int64_t bi = K;
    for (int64_t k0 = 0; k0 < K; k0 += 8 * AQL_ARGMAX_THREADS) {
        float x[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) {  // eight independent loads in flight per thread
            const int64_t k = k0 + u * AQL_ARGMAX_THREADS + tid;
            x[u] = k < K ? row[k] : -INFINITY;
        }
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            const int64_t k = k0 + u * AQL_ARGMAX_THREADS + tid;
            if (k < K && x[u] > bv) {  // ascending k per thread: strict > keeps the first maximum
                bv = x[u];
                bi = k;
            }
        }
    }
    for (int o = 16; o; o >>= 1) argmax_merge(bv, bi, __shfl_xor_sync(0xffffffffu, bv, o), __shfl_xor_sync(0xffffffffu, bi, o));
    if (lane == 0) {
        s_v[warp] = bv;
        s_i[warp] = bi;
    }
    __syncthreads();
    if (tid == 0) {
        for (int w = 1; w < AQL_ARGMAX_THREADS / 32; ++w) argmax_merge(bv, bi, s_v[w], s_i[w]);
        out[b] = bi;
    }
}

}  // namespace perb
