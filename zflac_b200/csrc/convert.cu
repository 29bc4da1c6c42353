// convert.cu — output conventions the reference lists as undecided (Readme.md:53, src/zflac.zig:282-292): the decoded PCM
// stays in the reference's container (s8 / s16 / s32, MSB-aligned) and this pass re-expresses it on the way out:
//   F32        v * 2^-(8 * container_bytes - 1)                 (int -> float round-to-nearest, then an exact scaling)
//   S24_PACKED the top 24 bits of an s32 sample, 3 bytes each, little-endian (v >> 8, arithmetic)
//   U8         s8 +% 128, the commented-out loop of :288-292
// One streaming pass, HBM-bound: every thread takes 4 samples (one 4/8/16-byte load, one 4/12/16-byte store).
#include "kernels.cuh"

namespace zf {

template <int CB> __device__ __forceinline__ void load4(const uint8_t *p, size_t i4, int32_t v[4]);
template <> __device__ __forceinline__ void load4<1>(const uint8_t *p, size_t i4, int32_t v[4]) {
    const char4 q = reinterpret_cast<const char4 *>(p)[i4];
    v[0] = q.x; v[1] = q.y; v[2] = q.z; v[3] = q.w;
}
template <> __device__ __forceinline__ void load4<2>(const uint8_t *p, size_t i4, int32_t v[4]) {
    const short4 q = reinterpret_cast<const short4 *>(p)[i4];
    v[0] = q.x; v[1] = q.y; v[2] = q.z; v[3] = q.w;
}
template <> __device__ __forceinline__ void load4<4>(const uint8_t *p, size_t i4, int32_t v[4]) {
    const int4 q = reinterpret_cast<const int4 *>(p)[i4];
    v[0] = q.x; v[1] = q.y; v[2] = q.z; v[3] = q.w;
}
template <int CB> __device__ __forceinline__ int32_t load1(const uint8_t *p, size_t i) {
    if (CB == 1) return reinterpret_cast<const int8_t *>(p)[i];
    if (CB == 2) return reinterpret_cast<const int16_t *>(p)[i];
    return reinterpret_cast<const int32_t *>(p)[i];
}

template <int CB, int FMT>
__global__ void __launch_bounds__(256) k_convert_pcm(const uint8_t *__restrict__ pcm, size_t n, uint8_t *__restrict__ out) {
    constexpr float kScale = CB == 1 ? 1.0f / 128.0f : CB == 2 ? 1.0f / 32768.0f : 1.0f / 2147483648.0f;
    const size_t tid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t nthreads = (size_t)gridDim.x * blockDim.x;
    const size_t n4 = n >> 2;
    for (size_t i = tid; i < n4; i += nthreads) {
        int32_t v[4];
        load4<CB>(pcm, i, v);
        if (FMT == kFmtF32) {
            reinterpret_cast<float4 *>(out)[i] = make_float4(__int2float_rn(v[0]) * kScale, __int2float_rn(v[1]) * kScale,
                                                              __int2float_rn(v[2]) * kScale, __int2float_rn(v[3]) * kScale);
        } else if (FMT == kFmtS24) {
            const uint32_t a = (uint32_t)v[0] >> 8, b = (uint32_t)v[1] >> 8, c = (uint32_t)v[2] >> 8, d = (uint32_t)v[3] >> 8;
            uint32_t *o = reinterpret_cast<uint32_t *>(out) + 3 * i;  // 12 bytes: a0 a1 a2 b0 | b1 b2 c0 c1 | c2 d0 d1 d2
            o[0] = a | (b << 24);
            o[1] = (b >> 8) | (c << 16);
            o[2] = (c >> 16) | (d << 8);
        } else {
            reinterpret_cast<uint32_t *>(out)[i] = reinterpret_cast<const uint32_t *>(pcm)[i] ^ 0x80808080u;
        }
    }
    for (size_t i = (n4 << 2) + tid; i < n; i += nthreads) {
        const int32_t v = load1<CB>(pcm, i);
        if (FMT == kFmtF32) {
            reinterpret_cast<float *>(out)[i] = __int2float_rn(v) * kScale;
        } else if (FMT == kFmtS24) {
            out[3 * i] = (uint8_t)(v >> 8);
            out[3 * i + 1] = (uint8_t)(v >> 16);
            out[3 * i + 2] = (uint8_t)(v >> 24);
        } else {
            out[i] = (uint8_t)(v + 128);
        }
    }
}

// pcm and out: 16-byte aligned device pointers. Returns cudaErrorInvalidValue for a pair the format does not define.
cudaError_t launch_convert_pcm(const void *pcm, uint64_t n_samples, uint32_t container_bytes, int fmt, void *out, cudaStream_t st) {
    if (n_samples == 0) return cudaSuccess;
    const uint32_t blocks = (uint32_t)std::min<uint64_t>(((n_samples >> 2) + 255) / 256 + 1, 148ull * 16);
    const uint8_t *p = (const uint8_t *)pcm;
    uint8_t *o = (uint8_t *)out;
    ZF_COUNT_LAUNCH();
    if (fmt == kFmtF32 && container_bytes == 1) k_convert_pcm<1, kFmtF32><<<blocks, 256, 0, st>>>(p, n_samples, o);
    else if (fmt == kFmtF32 && container_bytes == 2) k_convert_pcm<2, kFmtF32><<<blocks, 256, 0, st>>>(p, n_samples, o);
    else if (fmt == kFmtF32 && container_bytes == 4) k_convert_pcm<4, kFmtF32><<<blocks, 256, 0, st>>>(p, n_samples, o);
    else if (fmt == kFmtS24 && container_bytes == 4) k_convert_pcm<4, kFmtS24><<<blocks, 256, 0, st>>>(p, n_samples, o);
    else if (fmt == kFmtU8 && container_bytes == 1) k_convert_pcm<1, kFmtU8><<<blocks, 256, 0, st>>>(p, n_samples, o);
    else return cudaErrorInvalidValue;
    return cudaGetLastError();
}

}  // namespace zf
