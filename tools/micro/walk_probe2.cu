// The sampler's real walk_block / load_block (tree_kernels.cuh) with a runtime generation count, 4 warps per CTA.
#include <cstdio>
#include <cuda_runtime.h>
#include "tree_kernels.cuh"
using namespace perb;
__global__ void k(const double *v, int64_t node0, int s, double mass0, long long *out, int *sink) {
    const int lane = threadIdx.x & 31;
    double mass = mass0 + threadIdx.x * 1e-12;
    int acc = 0;
    long long t_load = 0, t_walk = 0;
#pragma unroll 1
    for (int r = 0; r < 32; ++r) {
        const long long t0 = clock64();
        DescentBlock blk = load_block(v, node0 + (acc & 1), s, lane);
        asm volatile("" ::"d"(blk.r0), "d"(blk.r1), "d"(blk.r2), "d"(blk.r3));
        const long long t1 = clock64();
        const int m = walk_block(blk, s, mass);
        const long long t2 = clock64();
        acc += m;
        mass = mass0 + 1e-9 * (m & 3);
        t_load += t1 - t0;
        t_walk += t2 - t1;
    }
    if (threadIdx.x == 0) { out[0] = t_load / 32; out[1] = t_walk / 32; sink[0] = acc; }
}
int main() {
    const int n = 1 << 16;
    double *h = new double[n];
    for (int i = 0; i < n; ++i) h[i] = 0.5 + 1e-6 * i;
    double *d; long long *o; int *s;
    cudaMalloc(&d, n * 8); cudaMalloc(&o, 64); cudaMalloc(&s, 64);
    cudaMemcpy(d, h, n * 8, cudaMemcpyHostToDevice);
    for (int threads : {32, 128}) for (int gens : {7, 6}) {
        for (int rep = 0; rep < 2; ++rep) k<<<1, threads>>>(d, 3, gens, 3.7, o, s);
        long long ho[2]; cudaMemcpy(ho, o, sizeof ho, cudaMemcpyDeviceToHost);
        printf("threads %d gens %d: load_block %lld cycles, walk_block %lld cycles\n", threads, gens, ho[0], ho[1]);
    }
    return 0;
}
