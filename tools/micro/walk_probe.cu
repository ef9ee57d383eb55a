// Micro-benchmark (one warp): cycles of the 7-generation block walk of the sampler, three formulations.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o walk_probe walk_probe.cu && ./walk_probe
#include <cstdio>
#include <cuda_runtime.h>
__device__ __forceinline__ double shfl_f64(double v, int src) {
    int lo = __double2loint(v), hi = __double2hiint(v);
    lo = __shfl_sync(0xffffffffu, lo, src);
    hi = __shfl_sync(0xffffffffu, hi, src);
    return __hiloint2double(hi, lo);
}
struct Blk { double r0, r1, r2, r3; };
__device__ __forceinline__ double block_left(const Blk &b, const int g, const int q) {
    if (g <= 5) return shfl_f64(b.r0, (1 << (g - 1)) - 1 + q);
    if (g == 6) return shfl_f64(b.r1, q);
    const double lo = shfl_f64(b.r2, q & 31), hi = shfl_f64(b.r3, q & 31);
    return (q >> 5) ? hi : lo;
}
__device__ __forceinline__ int walk_a(const Blk &b, double &mass) {  // shuffle behind the comparison
    int m = 0;
#pragma unroll
    for (int g = 1; g <= 7; ++g) {
        const double lv = block_left(b, g, m);
        if (lv > mass) m = 2 * m;
        else { mass = __dsub_rn(mass, lv); m = 2 * m + 1; }
    }
    return m;
}
__device__ __forceinline__ int walk_b(const Blk &b, double &mass) {  // both candidates of the next generation fetched early
    int m = 0;
    double lv = block_left(b, 1, 0);
#pragma unroll
    for (int g = 1; g <= 7; ++g) {
        double nl = 0.0, nr = 0.0;
        if (g < 7) { nl = block_left(b, g + 1, 2 * m); nr = block_left(b, g + 1, 2 * m + 1); }
        const bool right = !(lv > mass);
        const double less = __dsub_rn(mass, lv);
        mass = right ? less : mass;
        m = 2 * m + (right ? 1 : 0);
        lv = right ? nr : nl;
    }
    return m;
}
__device__ __forceinline__ int walk_c(const Blk &b, double &mass) {  // as b, integer comparison of the (non-negative) bit patterns
    int m = 0;
    double lv = block_left(b, 1, 0);
#pragma unroll
    for (int g = 1; g <= 7; ++g) {
        double nl = 0.0, nr = 0.0;
        if (g < 7) { nl = block_left(b, g + 1, 2 * m); nr = block_left(b, g + 1, 2 * m + 1); }
        const bool right = !(__double_as_longlong(lv) > __double_as_longlong(mass));
        const double less = __dsub_rn(mass, lv);
        mass = right ? less : mass;
        m = 2 * m + (right ? 1 : 0);
        lv = right ? nr : nl;
    }
    return m;
}
template <int V> __global__ void k(const double *v, double mass0, long long *out, int *sink) {
    const int lane = threadIdx.x;
    Blk b{v[lane], v[32 + lane], v[64 + lane], v[96 + lane]};
    double mass = mass0;
    int acc = 0;
    long long t0 = clock64();
#pragma unroll 1
    for (int r = 0; r < 64; ++r) {
        int m = V == 0 ? walk_a(b, mass) : (V == 1 ? walk_b(b, mass) : walk_c(b, mass));
        acc += m;
        mass = mass0 + 1e-9 * (m & 3);  // next walk depends on this one
    }
    long long t1 = clock64();
    if (lane == 0) { out[V] = (t1 - t0) / 64; sink[V] = acc; }
}
int main() {
    double h[128];
    for (int i = 0; i < 128; ++i) h[i] = 0.5 + 0.001 * i;
    double *d; long long *o; int *s;
    cudaMalloc(&d, sizeof h); cudaMalloc(&o, 64); cudaMalloc(&s, 64);
    cudaMemcpy(d, h, sizeof h, cudaMemcpyHostToDevice);
    for (int rep = 0; rep < 2; ++rep) { k<0><<<1, 32>>>(d, 3.7, o, s); k<1><<<1, 32>>>(d, 3.7, o, s); k<2><<<1, 32>>>(d, 3.7, o, s); }
    long long ho[3]; int hs[3];
    cudaMemcpy(ho, o, sizeof ho, cudaMemcpyDeviceToHost); cudaMemcpy(hs, s, sizeof hs, cudaMemcpyDeviceToHost);
    printf("cycles per 7-generation walk: shuffle behind compare %lld | candidates early %lld | + integer compare %lld  (paths %d %d %d)\n", ho[0], ho[1], ho[2], hs[0], hs[1], hs[2]);
    return 0;
}
