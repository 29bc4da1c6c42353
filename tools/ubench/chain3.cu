// chain3.cu — the K2a chain with the ring refilled (a) through registers by scattered LDG.128 (today), (b) by one
// cp.async.bulk (TMA) of 64 or 128 bytes per lane into a raw staging row + conversion to pair entries.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
constexpr int RING = 64;
constexpr int BODIES = 1536;
// MODE 0: chain only; 1: LDG refill (4 x LDG.128 per lane per 6 bodies); 2: TMA 64 B per lane per 6 bodies; 3: TMA 128 B per 12 bodies
template <int MODE>
__global__ void __launch_bounds__(32) k(const uint8_t *__restrict__ in, uint32_t k_, long long *cyc, uint32_t *sink) {
    __shared__ __align__(16) uint2 ring[RING * 32];
    __shared__ __align__(16) uint8_t raw[32][144];
    __shared__ __align__(8) uint64_t mbar[32];
    const uint32_t lane = threadIdx.x, lane8 = lane * 8;
    for (int i = lane; i < RING * 32; i += 32) {
        uint32_t h = (i * 2654435761u) ^ (i >> 3);
        ring[i] = make_uint2(h | 0x01010101u, (h * 40503u) | 0x10101010u);
    }
    const uint32_t mb = (uint32_t)__cvta_generic_to_shared(&mbar[lane]);
    const uint32_t rawp = (uint32_t)__cvta_generic_to_shared(&raw[lane][0]);
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(mb));
    asm volatile("fence.mbarrier_init.release.cluster;");
    __syncwarp();
    const uint32_t C = 159 + k_;
    uint32_t Q = 0u - lane * 13u - 64u, acc = 0, phase = 0, staged = 0;
    const uint8_t *g = in + (size_t)(blockIdx.x * 32 + lane) * 65536;  // every lane its own 64 KB: scattered
    uint4 q[4] = {};
    constexpr int PER = MODE == 3 ? 12 : 6;
    constexpr int BYTES = MODE == 3 ? 128 : 64;
    if (MODE >= 2) {
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(mb), "r"(BYTES) : "memory");
        asm volatile("cp.async.bulk.shared::cta.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(rawp), "l"(g), "r"(BYTES), "r"(mb) : "memory");
    }
    const long long t0 = clock64();
#pragma unroll 1
    for (int b = 0; b < BODIES; ++b) {
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            const uint32_t off = ((255u - 8u * Q) & ((RING - 1u) << 8)) | lane8;
            const uint2 e = *reinterpret_cast<const uint2 *>(reinterpret_cast<const char *>(ring) + off);
            const uint32_t win = __funnelshift_r(e.y, e.x, Q);
            const uint32_t f = __float_as_uint(__uint2float_rz(win));
            Q = (Q - C) + (f >> 23);
            acc += f;
        }
        if (MODE != 0 && (b % PER) == PER - 1) {
            char *row = reinterpret_cast<char *>(ring) + (((staged & (RING - 1)) << 8) | lane8);
            if (MODE == 1) {
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const uint32_t w0 = __byte_perm(q[j].x, 0, 0x0123) | 0x01010101u, w1 = __byte_perm(q[j].y, 0, 0x0123) | 0x10101010u;
                    const uint32_t w2 = __byte_perm(q[j].z, 0, 0x0123) | 0x01010101u, w3 = __byte_perm(q[j].w, 0, 0x0123) | 0x10101010u;
                    *reinterpret_cast<uint2 *>(row + 256 * (4 * j + 0)) = make_uint2(w3, w0);
                    *reinterpret_cast<uint2 *>(row + 256 * (4 * j + 1)) = make_uint2(w0, w1);
                    *reinterpret_cast<uint2 *>(row + 256 * (4 * j + 2)) = make_uint2(w1, w2);
                    *reinterpret_cast<uint2 *>(row + 256 * (4 * j + 3)) = make_uint2(w2, w3);
                }
                staged += 16;
                const uint4 *p = reinterpret_cast<const uint4 *>(g + (size_t)staged * 4);
#pragma unroll
                for (int j = 0; j < 4; ++j) q[j] = __ldg(p + j);
            } else {
                uint32_t done = 0;
                while (!done)
                    asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }" : "=r"(done) : "r"(mb), "r"(phase) : "memory");
                phase ^= 1;
#pragma unroll
                for (int j = 0; j < BYTES / 16; ++j) {
                    const uint4 v = *reinterpret_cast<const uint4 *>(&raw[lane][16 * j]);
                    const uint32_t w0 = __byte_perm(v.x, 0, 0x0123) | 0x01010101u, w1 = __byte_perm(v.y, 0, 0x0123) | 0x10101010u;
                    const uint32_t w2 = __byte_perm(v.z, 0, 0x0123) | 0x01010101u, w3 = __byte_perm(v.w, 0, 0x0123) | 0x10101010u;
                    char *r2 = reinterpret_cast<char *>(ring) + ((((staged + 4 * j) & (RING - 1)) << 8) | lane8);
                    *reinterpret_cast<uint2 *>(r2) = make_uint2(w3, w0);
                    *reinterpret_cast<uint2 *>(r2 + 256) = make_uint2(w0, w1);
                    *reinterpret_cast<uint2 *>(r2 + 512) = make_uint2(w1, w2);
                    *reinterpret_cast<uint2 *>(r2 + 768) = make_uint2(w2, w3);
                }
                staged += BYTES / 4;
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // generic reads of raw[] before the async proxy overwrites it
                asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(mb), "r"(BYTES) : "memory");
                asm volatile("cp.async.bulk.shared::cta.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(rawp), "l"(g + (size_t)staged * 4), "r"(BYTES), "r"(mb) : "memory");
            }
        }
    }
    const long long t1 = clock64();
    sink[blockIdx.x * 32 + lane] = acc + Q + q[0].x;
    if (lane == 0) cyc[blockIdx.x] = t1 - t0;
}
template <int MODE> void run(const char *name, int w, const uint8_t *in, long long *cyc, uint32_t *sink) {
    const int grid = 148 * 4 * w;
    k<MODE><<<grid, 32>>>(in, 9, cyc, sink);
    cudaDeviceSynchronize();
    double s = 0;
    for (int i = 0; i < grid; ++i) s += cyc[i];
    printf("%-36s W=%d: %.1f cycles/row  (%s)\n", name, w, s / grid / BODIES / 8, cudaGetErrorString(cudaGetLastError()));
}
int main() {
    long long *cyc; uint32_t *sink; uint8_t *in;
    cudaMallocManaged(&cyc, 148 * 16 * 8);
    cudaMalloc(&sink, 148 * 16 * 32 * 4);
    const size_t n = (size_t)148 * 4 * 4 * 32 * 65536;
    cudaMalloc(&in, n);
    cudaMemset(in, 0x5a, n);
    for (int w = 1; w <= 4; w *= 2) {
        run<0>("chain only", w, in, cyc, sink);
        run<1>("refill: scattered LDG.128 x4 / 6 bodies", w, in, cyc, sink);
        run<2>("refill: TMA 64 B per lane / 6 bodies", w, in, cyc, sink);
        run<3>("refill: TMA 128 B per lane / 12 bodies", w, in, cyc, sink);
    }
}
