// chain2.cu — variants of the K2a row chain (chain only, no global traffic), one warp per SMSP and two.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
constexpr int RING = 64;
constexpr int BODIES = 2048;
template <int V>
__global__ void __launch_bounds__(32) k(uint32_t k_, long long *cyc, uint32_t *sink) {
    __shared__ __align__(16) uint2 ring[RING * 32];
    const uint32_t lane = threadIdx.x, lane8 = lane * 8;
    for (int i = lane; i < RING * 32; i += 32) {
        uint32_t h = (i * 2654435761u) ^ (i >> 3);
        ring[i] = make_uint2(h | 0x01010101u, (h * 40503u) | 0x10101010u);
    }
    __syncwarp();
    const uint32_t C = (V == 1 ? 150 : 159) + k_;
    uint32_t Q = 0u - lane * 13u - 64u, acc = 0;
    uint32_t mask, magic;
    asm("mov.b32 %0, 0x7FFFFF;" : "=r"(mask));
    asm("mov.b32 %0, 0x4B000000;" : "=r"(magic));
    const long long t0 = clock64();
#pragma unroll 1
    for (int b = 0; b < BODIES; ++b) {
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            if (V == 3) {  // no LDS: window from registers (lower bound of the ALU part)
                const uint32_t win = __funnelshift_r(acc | 0x01010101u, Q | 0x10101010u, Q);
                const uint32_t f = __float_as_uint(__uint2float_rz(win));
                Q = (Q - C) + (f >> 23);
                acc += f;
                continue;
            }
            const uint32_t off = ((255u - 8u * Q) & ((RING - 1u) << 8)) | lane8;
            const uint2 e = *reinterpret_cast<const uint2 *>(reinterpret_cast<const char *>(ring) + off);
            const uint32_t win = __funnelshift_r(e.y, e.x, Q);
            uint32_t f;
            if (V == 0) f = __float_as_uint(__uint2float_rz(win));
            else if (V == 1) {
                uint32_t m;
                asm("lop3.b32 %0, %1, %2, %3, 0xEA;" : "=r"(m) : "r"(win), "r"(mask), "r"(magic));
                f = __float_as_uint(__fadd_rn(__uint_as_float(m), -8388608.0f));
            } else {  // V == 2: FLO
                uint32_t z;
                asm("bfind.shiftamt.u32 %0, %1;" : "=r"(z) : "r"(win));
                f = (158u - z) << 23;
            }
            Q = (Q - C) + (f >> 23);
            acc += f;
        }
    }
    const long long t1 = clock64();
    sink[blockIdx.x * 32 + lane] = acc + Q;
    if (lane == 0) cyc[blockIdx.x] = t1 - t0;
}
template <int V> void run(const char *name, int w, long long *cyc, uint32_t *sink) {
    const int grid = 148 * 4 * w;
    k<V><<<grid, 32>>>(9, cyc, sink);
    cudaDeviceSynchronize();
    double s = 0;
    for (int i = 0; i < grid; ++i) s += cyc[i];
    printf("%-28s W=%d: %.1f cycles/row  (%s)\n", name, w, s / grid / BODIES / 8, cudaGetErrorString(cudaGetLastError()));
}
int main() {
    long long *cyc; uint32_t *sink;
    cudaMallocManaged(&cyc, 148 * 16 * 8);
    cudaMalloc(&sink, 148 * 16 * 32 * 4);
    for (int w = 1; w <= 4; w *= 2) {
        run<0>("LDS + I2FP", w, cyc, sink);
        run<1>("LDS + LOP3 + FADD", w, cyc, sink);
        run<2>("LDS + FLO", w, cyc, sink);
        run<3>("no LDS, I2FP", w, cyc, sink);
    }
}
