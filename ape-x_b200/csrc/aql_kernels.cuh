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
// bulk copy; activations are split on the fly by the thread that owns the row.
#pragma once
#include "common.cuh"
#include "gather_kernels.cuh"  // mbarrier / bulk-copy wrappers

namespace perb {

constexpr int AQL_TILE = 128;                 // candidates per tile (UMMA M)
constexpr int AQL_H = 64;                     // a_out_unit = feature_out_unit = hidden of advantage1
constexpr int AQL_E1 = 128;                   // hidden of action_out (continuous)
constexpr int AQL_MAX_A = 8;
constexpr int AQL_ATOM_A = AQL_TILE * 128;    // bytes of one 32-float K-atom of a 128-row operand (16 KB)
constexpr int AQL_ATOM_W = AQL_H * 128;       // same for a 64-row operand (8 KB)
// weight image: [Wa2 hi: 4 atoms][Wa2 lo: 4 atoms][W1a hi: 2 atoms][W1a lo: 2 atoms]
constexpr int AQL_IMG_W2_HI = 0;
constexpr int AQL_IMG_W2_LO = 4 * AQL_ATOM_W;
constexpr int AQL_IMG_W1_HI = 8 * AQL_ATOM_W;
constexpr int AQL_IMG_W1_LO = 10 * AQL_ATOM_W;
constexpr int AQL_IMG_BYTES = 12 * AQL_ATOM_W;                 // 98 304
constexpr int AQL_SM_A_HI = AQL_IMG_BYTES;                     // activations, up to 4 atoms each
constexpr int AQL_SM_A_LO = AQL_SM_A_HI + 4 * AQL_ATOM_A;
constexpr int AQL_SM_SMALL = AQL_SM_A_LO + 4 * AQL_ATOM_A;     // 229 376: ba2[64] w2[64] barriers tmem slot
constexpr int AQL_SMEM_BYTES = AQL_IMG_BYTES + 1024 + 1024;     // weight image + small area + alignment slack (activations live in TMEM)

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
// kind::tf32, D = f32, A/B = tf32 K-major, N = 64, M = 128
constexpr uint32_t AQL_IDESC = (1u << 4) | (2u << 7) | (2u << 10) | ((64u >> 3) << 17) | ((128u >> 4) << 24);

__device__ __forceinline__ void umma_tf32(uint32_t tmem_d, uint64_t da, uint64_t db, uint32_t accumulate) {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t"
        "}\n" ::"r"(tmem_d),
        "l"(da), "l"(db), "r"(AQL_IDESC), "r"(accumulate)
        : "memory");
}
__device__ __forceinline__ void umma_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, float *v) {
    uint32_t *r = reinterpret_cast<uint32_t *>(v);
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]),
          "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]),
          "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
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
    const uint32_t off = (uint32_t)kb * AQL_ATOM_W + sw128(n, c16);
    *reinterpret_cast<float4 *>(img + (second ? AQL_IMG_W1_HI : AQL_IMG_W2_HI) + off) = make_float4(hi[0], hi[1], hi[2], hi[3]);
    *reinterpret_cast<float4 *>(img + (second ? AQL_IMG_W1_LO : AQL_IMG_W2_LO) + off) = make_float4(lo[0], lo[1], lo[2], lo[3]);
}

// sv[b][j] = b1[j] + sum_i W1[j][64 + i] * f[b][i]   (f is post-ReLU already, model.py:244-249)
__global__ void k_aql_state_half(const float *wadv1, const float *badv1, const float *f, int64_t B, float *sv) {
    const int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (t >= B * AQL_H) return;
    const int64_t b = t / AQL_H;
    const int j = (int)(t % AQL_H);
    float acc = 0.0f;
    for (int i = 0; i < AQL_H; ++i) acc = fmaf(wadv1[j * (2 * AQL_H) + AQL_H + i], fmaxf(f[b * AQL_H + i], 0.0f), acc);
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

__device__ __forceinline__ void store_split(uint8_t *a_hi, uint8_t *a_lo, int r, int cc, const float *v) {
    const uint32_t off = (uint32_t)(cc >> 3) * AQL_ATOM_A + sw128(r, cc & 7);
    float h[4], l[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
        h[e] = tf32_hi(v[e]);
        l[e] = tf32_hi(v[e] - h[e]);
    }
    *reinterpret_cast<float4 *>(a_hi + off) = make_float4(h[0], h[1], h[2], h[3]);
    *reinterpret_cast<float4 *>(a_lo + off) = make_float4(l[0], l[1], l[2], l[3]);
}

// Issues hi.lo + lo.hi + hi.hi for K = 32 * atoms into the accumulator at tmem_d (one thread).
__device__ __forceinline__ void issue_3xtf32(uint32_t tmem_d, uint32_t a_hi, uint32_t a_lo, uint32_t w_hi, uint32_t w_lo, int atoms) {
    uint32_t acc = 0;
#pragma unroll 1
    for (int p = 0; p < 3; ++p) {
        const uint32_t a = p == 0 ? a_lo : a_hi;
        const uint32_t w = p == 1 ? w_lo : w_hi;
#pragma unroll 1
        for (int kb = 0; kb < atoms; ++kb) {
#pragma unroll
            for (int k4 = 0; k4 < 4; ++k4) {
                umma_tf32(tmem_d, umma_desc(a + kb * AQL_ATOM_A + k4 * 32), umma_desc(w + kb * AQL_ATOM_W + k4 * 32), acc);
                acc = 1;
            }
        }
    }
}

// One K-atom (32 columns) of the 3xTF32 product, 12 MMAs.  MMAs that accumulate into the same TMEM
// columns form a dependent chain (K is only 8 per instruction), so the three partial products
// lo.hi / hi.lo / hi.hi go to three different accumulators (AQL_H columns apart, starting at
// tmem_d) and are interleaved; the epilogue adds the partials.  `first`: overwrite, do not accumulate.
__device__ __forceinline__ void issue_atom(uint32_t tmem_d, uint32_t a_hi, uint32_t a_lo, uint32_t w_hi, uint32_t w_lo, bool first) {
#pragma unroll
    for (int k4 = 0; k4 < 4; ++k4) {
        const uint32_t acc = (first && k4 == 0) ? 0u : 1u;
        umma_tf32(tmem_d + 0 * AQL_H, umma_desc(a_lo + k4 * 32), umma_desc(w_hi + k4 * 32), acc);
        umma_tf32(tmem_d + 1 * AQL_H, umma_desc(a_hi + k4 * 32), umma_desc(w_lo + k4 * 32), acc);
        umma_tf32(tmem_d + 2 * AQL_H, umma_desc(a_hi + k4 * 32), umma_desc(w_hi + k4 * 32), acc);
    }
}

// Sum of the AQL_PARTS partial accumulators (32 columns of this thread's row each).
constexpr int AQL_PARTS = 6;  // 3 products x 2 K groups
__device__ __forceinline__ void tmem_ld32_sum(uint32_t taddr, float *acc) {
    tmem_ld32(taddr, acc);
#pragma unroll 1
    for (int part = 1; part < AQL_PARTS; ++part) {
        float t[32];
        tmem_ld32(taddr + part * AQL_H, t);
#pragma unroll
        for (int i = 0; i < 32; ++i) acc[i] += t[i];
    }
}

constexpr int AQL_THREADS = 256;  // two threads per candidate row: each owns half of the columns

// tcgen05.st: this thread's TMEM lane, N consecutive 32-bit columns starting at taddr
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
__device__ __forceinline__ void umma_tf32_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t db, uint32_t accumulate) {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n\t"
        "}\n" ::"r"(tmem_d),
        "r"(tmem_a), "l"(db), "r"(AQL_IDESC), "r"(accumulate)
        : "memory");
}

// One K-atom (32 columns) of the 3xTF32 product with A in TMEM: lo.hi + hi.lo + hi.hi, 12 MMAs.
__device__ __forceinline__ void issue_atom_ts(uint32_t tmem_d, uint32_t a_hi, uint32_t a_lo, uint32_t w_hi, uint32_t w_lo, bool first) {
#pragma unroll
    for (int k4 = 0; k4 < 4; ++k4) {
        const uint64_t bh = umma_desc(w_hi + k4 * 32), bl = umma_desc(w_lo + k4 * 32);
        umma_tf32_ts(tmem_d, a_lo + k4 * 8, bh, (first && k4 == 0) ? 0u : 1u);
        umma_tf32_ts(tmem_d, a_hi + k4 * 8, bl, 1u);
        umma_tf32_ts(tmem_d, a_hi + k4 * 8, bh, 1u);
    }
}

// TMEM columns (512 allocated): activations of layer 1 (hi | lo, 128 each), accumulators of the two
// dense layers, activations of layer 2 (hi | lo, 64 each)
constexpr uint32_t AQL_T_A1H = 0, AQL_T_A1L = 128, AQL_T_D2 = 256, AQL_T_D3 = 320, AQL_T_A2H = 384, AQL_T_A2L = 448;

// A = action dimension (continuous) at compile time so that layer 1 unrolls into FFMAs whose
// weight operands come straight from the constant bank (`small` is a __grid_constant__ kernel
// parameter, 5 KB); CONT = false is the discrete path (Linear(1,64), no second action layer).
// Activations never touch shared memory: the thread that owns a row writes its TF32 hi / lo
// halves into its TMEM lane (tcgen05.st) and the MMAs take A from TMEM, B (weights) from the
// swizzled image in shared memory — the SS form re-read 6 KB of shared memory per K=8 MMA
// and was bound by that (profiles/r01_notes.md).
template <int A, bool CONT>
__global__ void __launch_bounds__(AQL_THREADS, 1) k_aql_q(const AqlArgs a, const __grid_constant__ AqlSmall small) {
    extern __shared__ uint8_t aql_raw[];
    uint8_t *sm = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(aql_raw) + 1023) & ~uintptr_t(1023));
    float *s_part = reinterpret_cast<float *>(sm + AQL_IMG_BYTES);          // [128] partial dot of the upper column half
    uint64_t *bar_w = reinterpret_cast<uint64_t *>(s_part + AQL_TILE);      // weights landed
    uint64_t *bar_m = bar_w + 1;                                            // MMAs of one layer done
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(bar_m + 1);

    const int tid = threadIdx.x;
    const int r = tid & (AQL_TILE - 1);   // row of the tile = TMEM lane
    const int half = tid >> 7;            // which half of the columns this thread owns
    const int warp = tid >> 5;
    if (tid == 0) {
        mbar_init(smem_u32(bar_w), 1);
        mbar_init(smem_u32(bar_m), 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        mbar_expect_tx(smem_u32(bar_w), AQL_IMG_BYTES);
        bulk_g2s(smem_u32(sm), a.img, AQL_IMG_BYTES, smem_u32(bar_w));
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "n"(512));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem = *tmem_slot;
    const uint32_t lane_base = tmem + ((uint32_t)((warp & 3) * 32) << 16);  // this warp's 32 TMEM lanes, column 0
    const uint32_t w2h = smem_u32(sm + AQL_IMG_W2_HI), w2l = smem_u32(sm + AQL_IMG_W2_LO);
    const uint32_t w1h = smem_u32(sm + AQL_IMG_W1_HI), w1l = smem_u32(sm + AQL_IMG_W1_LO);
    mbar_wait(smem_u32(bar_w), 0);
    uint32_t phase = 0;

    const int64_t tiles = (a.total + AQL_TILE - 1) / AQL_TILE;
    for (int64_t tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
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
            // thread); the 12 MMAs of an atom are issued as soon as it sits in TMEM
            float act[A];
            const float *ap = static_cast<const float *>(a.a_mu) + ci * A;
#pragma unroll
            for (int j = 0; j < A; ++j) act[j] = ap[j];
#pragma unroll
            for (int kb = 0; kb < 4; ++kb) {
#pragma unroll
                for (int e = 0; e < 16; ++e) {
                    const int o = kb * 32 + half * 16 + e;
                    float s = small.ba1[o];
#pragma unroll
                    for (int j = 0; j < A; ++j) s = fmaf(small.wa1[o * A + j], act[j], s);
                    s = fmaxf(s, 0.0f);
                    hi[e] = tf32_hi(s);
                    lo[e] = tf32_hi(s - hi[e]);
                }
                tmem_st16(lane_base + AQL_T_A1H + kb * 32 + half * 16, hi);
                tmem_st16(lane_base + AQL_T_A1L + kb * 32 + half * 16, lo);
                tmem_st_wait();
                asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
                __syncthreads();
                if (tid == 0) {
                    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                    issue_atom_ts(tmem + AQL_T_D2, tmem + AQL_T_A1H + kb * 32, tmem + AQL_T_A1L + kb * 32, w2h + kb * AQL_ATOM_W,
                                  w2l + kb * AQL_ATOM_W, kb == 0);
                    if (kb == 3) umma_commit(smem_u32(bar_m));
                }
            }
            AQL_TRACE(0);  // layer 1 + MMA issue
            mbar_wait(smem_u32(bar_m), phase);
            phase ^= 1u;
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            AQL_TRACE(1);  // wait for layer-2 MMAs
            // layer 2 epilogue: e = relu(D2 + ba2) -> A operand of layer 3 (K = 64), 32 columns per thread
            float acc[32];
            tmem_ld32(lane_base + AQL_T_D2 + half * 32, acc);
#pragma unroll
            for (int g = 0; g < 2; ++g) {
#pragma unroll
                for (int e = 0; e < 16; ++e) {
                    const float s = fmaxf(acc[g * 16 + e] + small.ba2[half * 32 + g * 16 + e], 0.0f);
                    hi[e] = tf32_hi(s);
                    lo[e] = tf32_hi(s - hi[e]);
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
                    const float s = fmaxf(fmaf(small.wa1[o], av, small.ba1[o]), 0.0f);
                    hi[e] = tf32_hi(s);
                    lo[e] = tf32_hi(s - hi[e]);
                }
                tmem_st16(lane_base + AQL_T_A2H + half * 32 + g * 16, hi);
                tmem_st16(lane_base + AQL_T_A2L + half * 32 + g * 16, lo);
            }
        }
        tmem_st_wait();
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
        __syncthreads();
        if (tid == 0) {
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            for (int kb = 0; kb < 2; ++kb)
                issue_atom_ts(tmem + AQL_T_D3, tmem + AQL_T_A2H + kb * 32, tmem + AQL_T_A2L + kb * 32, w1h + kb * AQL_ATOM_W,
                              w1l + kb * AQL_ATOM_W, kb == 0);
            umma_commit(smem_u32(bar_m));
        }
        AQL_TRACE(2);  // layer-2 epilogue (or discrete layer 1) + MMA issue
        mbar_wait(smem_u32(bar_m), phase);
        phase ^= 1u;
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        AQL_TRACE(3);  // wait for layer-3 MMAs
        // layer 3 epilogue: q = w2 . relu(D3 + sv_b) + b2; each thread sums its 32 columns
        const float4 *svb = reinterpret_cast<const float4 *>(a.sv + (ci / a.K) * AQL_H + half * 32);
        float qv = 0.0f;
        {
            float acc[32];
            tmem_ld32(lane_base + AQL_T_D3 + half * 32, acc);
#pragma unroll
            for (int i4 = 0; i4 < 8; ++i4) {
                const float4 s4 = svb[i4];
                qv = fmaf(small.w2[half * 32 + i4 * 4 + 0], fmaxf(acc[i4 * 4 + 0] + s4.x, 0.0f), qv);
                qv = fmaf(small.w2[half * 32 + i4 * 4 + 1], fmaxf(acc[i4 * 4 + 1] + s4.y, 0.0f), qv);
                qv = fmaf(small.w2[half * 32 + i4 * 4 + 2], fmaxf(acc[i4 * 4 + 2] + s4.z, 0.0f), qv);
                qv = fmaf(small.w2[half * 32 + i4 * 4 + 3], fmaxf(acc[i4 * 4 + 3] + s4.w, 0.0f), qv);
            }
        }
        if (half) s_part[r] = qv;
        // the next tile overwrites the TMEM activations and accumulators: everyone must be done reading
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
        __syncthreads();
        if (!half && live) a.q[c] = (qv + s_part[r]) + small.b2;
        AQL_TRACE(4);  // layer-3 epilogue
        if (a.trace && blockIdx.x == 0 && tid == 0) a.trace[5] += 1;
    }
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "n"(512));
}

// argmax over the K candidates of each state, one warp per state; the first maximum wins
// (torch.argmax, model.py:330-332).
__global__ void k_aql_argmax(const float *q, int64_t B, int64_t K, int64_t *out) {
    const int lane = threadIdx.x & 31;
    const int64_t b = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
    if (b >= B) return;
    float bv = -INFINITY;
    int64_t bi = K;
    for (int64_t k = lane; k < K; k += 32) {
        const float x = q[b * K + k];
        if (x > bv) { bv = x; bi = k; }
    }
    for (int o = 16; o; o >>= 1) {
        const float ov = __shfl_xor_sync(0xffffffffu, bv, o);
        const int64_t oi = __shfl_xor_sync(0xffffffffu, bi, o);
        if (ov > bv || (ov == bv && oi < bi)) { bv = ov; bi = oi; }
    }
    if (lane == 0) out[b] = bi;
}

}  // namespace perb
