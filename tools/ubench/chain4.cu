// chain4.cu — cumulative replica of the two-phase narrow K2a segment: which part costs what?
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
constexpr int RING = 64;
constexpr int SEGS = 384;   // segments of 32 rows
template <int MODE>  // 0 chain; 1 +tile STS; 2 +extract phase (LDS+ALU); 3 +coalesced STG.128 per body; 4 +scattered LDG refill per 6 bodies
__global__ void __launch_bounds__(32) k(const uint8_t *__restrict__ in, uint4 *__restrict__ out, uint32_t k_, long long *cyc, uint32_t *sink) {
    __shared__ __align__(16) uint2 ring[RING * 32];
    __shared__ uint32_t tile[32 * 32];
    const uint32_t lane = threadIdx.x, lane8 = lane * 8;
    for (int i = lane; i < RING * 32; i += 32) {
        uint32_t h = (i * 2654435761u) ^ (i >> 3);
        ring[i] = make_uint2(h | 0x01010101u, (h * 40503u) | 0x10101010u);
    }
    __syncwarp();
    const uint32_t C = 159 + k_, sh = 23 - k_, nmask = ~((1u << k_) - 1u), C2 = 158u << k_;
    uint32_t Q = 0u - lane * 13u - 64u, acc = 0, staged = 0, body = 0;
    const uint8_t *g = in + (size_t)(blockIdx.x * 32 + lane) * 65536;
    uint4 *o = out + (size_t)blockIdx.x * 32 * 4096 + lane;  // warp-interleaved rows: 512 contiguous bytes per store
    uint4 q[4] = {};
    const long long t0 = clock64();
#pragma unroll 1
    for (int s = 0; s < SEGS; ++s) {
        uint32_t fmin = 0xFFFFFFFFu;
#pragma unroll 1
        for (int r0 = 0; r0 < 32; r0 += 8) {
            if (MODE >= 4 && (body % 6) == 5) {
                char *row = reinterpret_cast<char *>(ring) + (((staged & (RING - 1)) << 8) | lane8);
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
                const uint4 *p = reinterpret_cast<const uint4 *>(g + (size_t)(staged & 8191) * 4);
#pragma unroll
                for (int j = 0; j < 4; ++j) q[j] = __ldg(p + j);
            }
            ++body;
#pragma unroll
            for (int u = 0; u < 8; ++u) {
                const uint32_t off = ((255u - 8u * Q) & ((RING - 1u) << 8)) | lane8;
                const uint2 e = *reinterpret_cast<const uint2 *>(reinterpret_cast<const char *>(ring) + off);
                const uint32_t win = __funnelshift_r(e.y, e.x, Q);
                const uint32_t f = __float_as_uint(__uint2float_rz(win));
                Q = (Q - C) + (f >> 23);
                fmin = fmin < f ? fmin : f;
                if (MODE >= 1) tile[(r0 + u) * 32 + lane] = f; else acc += f;
            }
        }
        if (MODE >= 2 && fmin != 12345u) {
#pragma unroll 1
            for (int r0 = 0; r0 < 32; r0 += 8) {
                int32_t r[8];
#pragma unroll
                for (int u = 0; u < 8; ++u) {
                    const uint32_t f = tile[(r0 + u) * 32 + lane];
                    const uint32_t x = f >> sh, y = x & nmask, cw = (x + C2) - 2u * y;
                    const uint32_t b = cw & 1u, h = cw >> 1;
                    r[u] = (int32_t)(h - b * cw);
                }
                uint4 v;
                v.x = __byte_perm(r[0], r[1], 0x5410); v.y = __byte_perm(r[2], r[3], 0x5410);
                v.z = __byte_perm(r[4], r[5], 0x5410); v.w = __byte_perm(r[6], r[7], 0x5410);
                if (MODE >= 3) o[(size_t)((s * 4 + (r0 >> 3)) & 4095) * 32] = v;
                else acc += v.x ^ v.y ^ v.z ^ v.w;
            }
        }
    }
    const long long t1 = clock64();
    sink[blockIdx.x * 32 + lane] = acc + Q + q[0].x;
    if (lane == 0) cyc[blockIdx.x] = t1 - t0;
}
template <int MODE> void run(const char *name, int w, const uint8_t *in, uint4 *out, long long *cyc, uint32_t *sink) {
    const int grid = 148 * 4 * w;
    k<MODE><<<grid, 32>>>(in, out, 9, cyc, sink);
    cudaDeviceSynchronize();
    double s = 0;
    for (int i = 0; i < grid; ++i) s += cyc[i];
    printf("%-44s W=%d: %.1f cycles/row  (%s)\n", name, w, s / grid / SEGS / 32, cudaGetErrorString(cudaGetLastError()));
}
int main() {
    long long *cyc; uint32_t *sink; uint8_t *in; uint4 *out;
    cudaMallocManaged(&cyc, 148 * 16 * 8);
    cudaMalloc(&sink, 148 * 16 * 32 * 4);
    const size_t n = (size_t)148 * 4 * 2 * 32 * 65536;
    cudaMalloc(&in, n);
    cudaMalloc(&out, n);
    cudaMemset(in, 0x5a, n);
    for (int w = 1; w <= 2; ++w) {
        run<0>("chain", w, in, out, cyc, sink);
        run<1>("+ tile STS per row", w, in, out, cyc, sink);
        run<2>("+ extract phase (LDS, ALU)", w, in, out, cyc, sink);
        run<3>("+ coalesced STG.128 per body", w, in, out, cyc, sink);
        run<4>("+ scattered LDG refill per 6 bodies", w, in, out, cyc, sink);
    }
}
