// Microbenchmark: throughput of divergent record gathers (what a BVH node fetch is) on B200.
// Each thread walks a pseudo-random chain through a table of REC-byte records, reading each record with
// LDG.128 or LDG.256 instructions; 32 lanes of a warp hit 32 different records. Reports records/s and the
// implied L1 wavefronts. Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o gather gather.cu
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
#include <cstdlib>

__device__ __forceinline__ void ld256(const void *p, uint32_t (&v)[8]) {
    asm volatile("ld.global.nc.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
        : "=r"(v[0]),"=r"(v[1]),"=r"(v[2]),"=r"(v[3]),"=r"(v[4]),"=r"(v[5]),"=r"(v[6]),"=r"(v[7]) : "l"(p));
}
__device__ __forceinline__ void ld128(const void *p, uint32_t (&v)[4]) {
    asm volatile("ld.global.nc.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v[0]),"=r"(v[1]),"=r"(v[2]),"=r"(v[3]) : "l"(p));
}

// REC: record bytes; W: 16 or 32 (load width in bytes); steps: chain length
template <int REC, int W>
__global__ void __launch_bounds__(128) k_gather(const char *table, uint32_t nRec, int steps, uint32_t *out) {
    uint32_t x = (blockIdx.x * blockDim.x + threadIdx.x) * 2654435761u + 12345u;
    uint32_t acc = 0;
    for (int s = 0; s < steps; ++s) {
        const uint32_t idx = x % nRec;
        const char *p = table + (size_t) idx * REC;
        if (W == 32) {
#pragma unroll
            for (int o = 0; o < REC; o += 32) { uint32_t v[8]; ld256(p + o, v); for (int j = 0; j < 8; ++j) acc += v[j]; }
        } else {
#pragma unroll
            for (int o = 0; o < REC; o += 16) { uint32_t v[4]; ld128(p + o, v); for (int j = 0; j < 4; ++j) acc += v[j]; }
        }
        x = x * 1664525u + 1013904223u + (acc & 1u);     // next index depends on the data: a dependent chain like traversal
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}

template <int REC, int W>
void run(const char *table, size_t tableBytes, uint32_t *out, int blocks, int steps) {
    const uint32_t nRec = (uint32_t) (tableBytes / REC);
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    k_gather<REC, W><<<blocks, 128>>>(table, nRec, steps, out);
    cudaEventRecord(a);
    k_gather<REC, W><<<blocks, 128>>>(table, nRec, steps, out);
    cudaEventRecord(b); cudaEventSynchronize(b);
    float ms; cudaEventElapsedTime(&ms, a, b);
    const double recs = (double) blocks * 128 * steps;
    printf("table %6.0f MB  rec %3d B  LDG.%d x%d : %8.3f ms  %7.2f Grec/s  %7.1f GB/s useful\n", tableBytes / 1e6, REC, W * 8, REC / W, ms,
        recs / ms * 1e-6, recs * REC / ms * 1e-6);
}

int main() {
    const size_t maxBytes = (size_t) 1 << 30;
    char *table; uint32_t *out;
    cudaMalloc(&table, maxBytes); cudaMemset(table, 1, maxBytes);
    const int blocks = 148 * 16, steps = 256;
    cudaMalloc(&out, (size_t) blocks * 128 * 4);
    for (size_t tb : { (size_t) 32 << 20, (size_t) 180 << 20, (size_t) 1 << 30 }) {
        run<128, 16>(table, tb, out, blocks, steps);
        run<128, 32>(table, tb, out, blocks, steps);
        run<96, 16>(table, tb, out, blocks, steps);
        run<96, 32>(table, tb, out, blocks, steps);
        run<64, 16>(table, tb, out, blocks, steps);
        run<64, 32>(table, tb, out, blocks, steps);
        run<32, 16>(table, tb, out, blocks, steps);
        run<32, 32>(table, tb, out, blocks, steps);
    }
    cudaError_t e = cudaDeviceSynchronize();
    printf("status: %s\n", cudaGetErrorString(e));
    return e != cudaSuccess;
}
