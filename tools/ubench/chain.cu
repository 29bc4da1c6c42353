// chain.cu — replica of the narrow K2a body (pos_reader.cuh) on synthetic data: what does a row cost, alone and with
// the kernel's global traffic (one scattered 16-byte store per body and lane, 64 scattered bytes loaded per six bodies)?
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o chain chain.cu && ./chain
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

constexpr int RING = 64;
constexpr int BODIES = 2048;

template <int MODE>  // bit 0: scattered STG.128 per body; bit 1: scattered 4 x LDG.128 per 6 bodies (+16 STS.64)
__global__ void __launch_bounds__(32) k_chain(const uint4 *__restrict__ in, uint4 *__restrict__ out, uint32_t k, long long *cyc) {
    __shared__ __align__(16) uint2 ring[RING * 32];
    const uint32_t lane = threadIdx.x, lane8 = lane * 8;
    for (int i = lane; i < RING * 32; i += 32) {
        uint32_t h = (i * 2654435761u) ^ (i >> 3);
        ring[i] = make_uint2(h | 0x01010101u, (h * 40503u) | 0x10101010u);  // dense in ones: short unary parts
    }
    __syncwarp();
    const uint32_t C = 159 + k, sh = 23 - k, nmask = ~((1u << k) - 1u), C2 = 158u << k;
    uint32_t Q = 0u - lane * 13u;
    const size_t region = (size_t)(blockIdx.x * 32 + lane) * 4096;  // 64 KB per lane: scattered sectors
    uint4 q[4] = {};
    uint32_t acc = 0;
    const long long t0 = clock64();
#pragma unroll 1
    for (int b = 0; b < BODIES; ++b) {
        uint32_t f[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            const uint32_t off = ((255u - 8u * Q) & ((RING - 1u) << 8)) | lane8;
            const uint2 e = *reinterpret_cast<const uint2 *>(reinterpret_cast<const char *>(ring) + off);
            const uint32_t win = __funnelshift_r(e.y, e.x, Q);
            f[u] = __float_as_uint(__uint2float_rz(win));
            Q = (Q - C) + (f[u] >> 23);
        }
        uint32_t r[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            const uint32_t x = f[u] >> sh, y = x & nmask, cw = (x + C2) - 2u * y;
            r[u] = (cw >> 1) ^ (0u - (cw & 1u));
        }
        if (MODE & 1) {
            uint4 v;
            v.x = __byte_perm(r[0], r[1], 0x5410); v.y = __byte_perm(r[2], r[3], 0x5410);
            v.z = __byte_perm(r[4], r[5], 0x5410); v.w = __byte_perm(r[6], r[7], 0x5410);
            out[region + (b & 4095)] = v;
        } else {
#pragma unroll
            for (int u = 0; u < 8; ++u) acc += r[u];
        }
        if ((MODE & 2) && (b % 6) == 5) {
            char *row = reinterpret_cast<char *>(ring) + ((((b / 6) * 16) & (RING - 1)) << 8 | lane8);
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                *reinterpret_cast<uint2 *>(row + 256 * (4 * j + 0)) = make_uint2(q[j].x | 0x01010101u, q[j].y | 0x10101010u);
                *reinterpret_cast<uint2 *>(row + 256 * (4 * j + 1)) = make_uint2(q[j].y | 0x01010101u, q[j].z | 0x10101010u);
                *reinterpret_cast<uint2 *>(row + 256 * (4 * j + 2)) = make_uint2(q[j].z | 0x01010101u, q[j].w | 0x10101010u);
                *reinterpret_cast<uint2 *>(row + 256 * (4 * j + 3)) = make_uint2(q[j].w | 0x01010101u, q[j].x | 0x10101010u);
            }
            const uint4 *p = in + region + (size_t)(b / 6) * 4;
#pragma unroll
            for (int j = 0; j < 4; ++j) q[j] = __ldg(p + j);
        }
    }
    const long long t1 = clock64();
    if (acc == 0x12345678u) out[0] = make_uint4(acc, Q, 0, 0);
    if (lane == 0) cyc[blockIdx.x] = t1 - t0;
}

template <int MODE>
void run(const char *name, int warps_per_smsp, const uint4 *in, uint4 *out, long long *cyc, uint32_t k) {
    const int grid = 148 * 4 * warps_per_smsp;
    k_chain<MODE><<<grid, 32>>>(in, out, k, cyc);
    cudaDeviceSynchronize();
    double s = 0;
    long long mx = 0;
    for (int i = 0; i < grid; ++i) { s += cyc[i]; mx = cyc[i] > mx ? cyc[i] : mx; }
    printf("%-34s W=%d: %.1f cycles/row avg, %.1f max   (%s)\n", name, warps_per_smsp, s / grid / BODIES / 8, (double)mx / BODIES / 8,
           cudaGetErrorString(cudaGetLastError()));
}

int main() {
    uint4 *in, *out;
    long long *cyc;
    const size_t n = (size_t)148 * 4 * 4 * 32 * 4096;  // 16-byte units: 64 KB per lane
    cudaMalloc(&in, n * 16);
    cudaMalloc(&out, n * 16);
    cudaMemset(in, 0x5a, n * 16);
    cudaMallocManaged(&cyc, 148 * 4 * 4 * 8);
    for (int w = 1; w <= 4; w *= 2) {
        run<0>("chain only", w, in, out, cyc, 9);
        run<1>("chain + scattered STG.128 / body", w, in, out, cyc, 9);
        run<2>("chain + refill LDG / 6 bodies", w, in, out, cyc, 9);
        run<3>("chain + both", w, in, out, cyc, 9);
    }
    return 0;
}
