// lat.cu — dependent-issue latencies of the instructions the Rice bit chain is made of (sm_100a), one warp at a time.
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o lat lat.cu && ./lat
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

#define N_IT 4096

__global__ void k_flo(uint32_t seed, uint32_t *out, long long *cyc) {
    uint32_t v = seed | 0x00010000u;
    long long t0 = clock64();
#pragma unroll 16
    for (int i = 0; i < N_IT; ++i) {
        uint32_t z;
        asm volatile("bfind.shiftamt.u32 %0, %1;" : "=r"(z) : "r"(v));
        v = z | 0x00010000u;  // LOP3 (4) + FLO
    }
    long long t1 = clock64();
    out[threadIdx.x] = v;
    if (threadIdx.x == 0) cyc[0] = t1 - t0;
}
__global__ void k_lop(uint32_t seed, uint32_t *out, long long *cyc) {
    uint32_t v = seed;
    long long t0 = clock64();
#pragma unroll 16
    for (int i = 0; i < N_IT; ++i) {
        asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(v) : "r"(seed), "r"(i));
        asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(v) : "r"(seed), "r"(i));
    }
    long long t1 = clock64();
    out[threadIdx.x] = v;
    if (threadIdx.x == 0) cyc[1] = t1 - t0;
}
__global__ void k_shf(uint32_t seed, uint32_t *out, long long *cyc) {
    uint32_t v = seed, w = seed * 3;
    long long t0 = clock64();
#pragma unroll 16
    for (int i = 0; i < N_IT; ++i) {
        asm volatile("shf.l.wrap.b32 %0, %1, %0, %2;" : "+r"(v) : "r"(w), "r"(i));
        asm volatile("shf.l.wrap.b32 %0, %1, %0, %2;" : "+r"(v) : "r"(w), "r"(i));
    }
    long long t1 = clock64();
    out[threadIdx.x] = v;
    if (threadIdx.x == 0) cyc[2] = t1 - t0;
}
__global__ void k_lds(uint32_t seed, uint32_t *out, long long *cyc) {
    __shared__ uint32_t ring[32 * 64];
    for (int i = threadIdx.x; i < 32 * 64; i += 32) ring[i] = ((i / 32 + 1) % 64) * 128 + (i % 32) * 4;  // next address
    __syncwarp();
    uint32_t a = threadIdx.x * 4;
    const uint32_t base = (uint32_t)__cvta_generic_to_shared(ring);
    long long t0 = clock64();
#pragma unroll 16
    for (int i = 0; i < N_IT; ++i) asm volatile("ld.shared.u32 %0, [%1];" : "=r"(a) : "r"(base + a));
    long long t1 = clock64();
    out[threadIdx.x] = a;
    if (threadIdx.x == 0) cyc[3] = t1 - t0;
}
__global__ void k_lds64(uint32_t seed, uint32_t *out, long long *cyc) {
    __shared__ uint2 ring[32 * 64];
    for (int i = threadIdx.x; i < 32 * 64; i += 32) ring[i] = make_uint2(((i / 32 + 1) % 64) * 256 + (i % 32) * 8, i);
    __syncwarp();
    uint32_t a = threadIdx.x * 8, b = 0;
    const uint32_t base = (uint32_t)__cvta_generic_to_shared(ring);
    long long t0 = clock64();
#pragma unroll 16
    for (int i = 0; i < N_IT; ++i) asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(a), "=r"(b) : "r"(base + a));
    long long t1 = clock64();
    out[threadIdx.x] = a + b;
    if (threadIdx.x == 0) cyc[4] = t1 - t0;
}
__global__ void k_imad(uint32_t seed, uint32_t *out, long long *cyc) {
    uint32_t v = seed;
    long long t0 = clock64();
#pragma unroll 16
    for (int i = 0; i < N_IT; ++i) {
        asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(v) : "r"(seed), "r"(i));
        asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(v) : "r"(seed), "r"(i));
    }
    long long t1 = clock64();
    out[threadIdx.x] = v;
    if (threadIdx.x == 0) cyc[5] = t1 - t0;
}
// alternating ALU -> FMA pipe
__global__ void k_mix(uint32_t seed, uint32_t *out, long long *cyc) {
    uint32_t v = seed;
    long long t0 = clock64();
#pragma unroll 16
    for (int i = 0; i < N_IT; ++i) {
        asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(v) : "r"(seed), "r"(i));
        asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(v) : "r"(seed), "r"(i));
    }
    long long t1 = clock64();
    out[threadIdx.x] = v;
    if (threadIdx.x == 0) cyc[6] = t1 - t0;
}
// throughput: 8 independent LOP3 chains per thread, W warps per SMSP -> cycles per warp-instruction
__global__ void k_alu_tput(uint32_t seed, uint32_t *out, long long *cyc, int slot) {
    uint32_t v[8];
    for (int j = 0; j < 8; ++j) v[j] = seed + j + threadIdx.x;
    long long t0 = clock64();
#pragma unroll 4
    for (int i = 0; i < N_IT; ++i) {
#pragma unroll
        for (int j = 0; j < 8; ++j) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(v[j]) : "r"(seed), "r"(i));
    }
    long long t1 = clock64();
    uint32_t s = 0;
    for (int j = 0; j < 8; ++j) s += v[j];
    out[threadIdx.x] = s;
    if (threadIdx.x == 0) cyc[slot] = t1 - t0;
}
__global__ void k_flo_tput(uint32_t seed, uint32_t *out, long long *cyc, int slot) {
    uint32_t v[8];
    for (int j = 0; j < 8; ++j) v[j] = seed + j + threadIdx.x;
    long long t0 = clock64();
#pragma unroll 4
    for (int i = 0; i < N_IT; ++i) {
#pragma unroll
        for (int j = 0; j < 8; ++j) asm volatile("bfind.shiftamt.u32 %0, %0;" : "+r"(v[j]));
    }
    long long t1 = clock64();
    uint32_t s = 0;
    for (int j = 0; j < 8; ++j) s += v[j];
    out[threadIdx.x] = s;
    if (threadIdx.x == 0) cyc[slot] = t1 - t0;
}

int main() {
    uint32_t *out;
    long long *cyc;
    cudaMalloc(&out, 4096 * 4);
    cudaMallocManaged(&cyc, 64 * 8);
    k_flo<<<1, 32>>>(12345, out, cyc);
    k_lop<<<1, 32>>>(12345, out, cyc);
    k_shf<<<1, 32>>>(12345, out, cyc);
    k_lds<<<1, 32>>>(12345, out, cyc);
    k_lds64<<<1, 32>>>(12345, out, cyc);
    k_imad<<<1, 32>>>(12345, out, cyc);
    k_mix<<<1, 32>>>(12345, out, cyc);
    k_alu_tput<<<1, 32>>>(1, out, cyc, 8);     // 1 warp on one SMSP
    k_alu_tput<<<1, 128>>>(1, out, cyc, 9);    // 1 warp per SMSP
    k_alu_tput<<<1, 256>>>(1, out, cyc, 10);   // 2 warps per SMSP
    k_flo_tput<<<1, 32>>>(1, out, cyc, 11);
    k_flo_tput<<<1, 256>>>(1, out, cyc, 12);
    cudaDeviceSynchronize();
    printf("FLO+LOP3 chain: %.2f cyc/iter (minus LOP3)\n", (double)cyc[0] / N_IT);
    printf("LOP3 chain:     %.2f cyc/op\n", (double)cyc[1] / N_IT / 2);
    printf("SHF chain:      %.2f cyc/op\n", (double)cyc[2] / N_IT / 2);
    printf("LDS.32 chain:   %.2f cyc/op (incl. IADD)\n", (double)cyc[3] / N_IT);
    printf("LDS.64 chain:   %.2f cyc/op (incl. IADD)\n", (double)cyc[4] / N_IT);
    printf("IMAD chain:     %.2f cyc/op\n", (double)cyc[5] / N_IT / 2);
    printf("IMAD<->LOP3:    %.2f cyc/op\n", (double)cyc[6] / N_IT / 2);
    printf("LOP3 tput 1 warp:        %.2f cyc/warp-inst\n", (double)cyc[8] / N_IT / 8);
    printf("LOP3 tput 1 warp/SMSP:   %.2f cyc/warp-inst\n", (double)cyc[9] / N_IT / 8);
    printf("LOP3 tput 2 warps/SMSP:  %.2f cyc/warp-inst per warp\n", (double)cyc[10] / N_IT / 8);
    printf("FLO tput 1 warp:         %.2f cyc/warp-inst\n", (double)cyc[11] / N_IT / 8);
    printf("FLO tput 2 warps/SMSP:   %.2f cyc/warp-inst per warp\n", (double)cyc[12] / N_IT / 8);
    printf("err=%s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
