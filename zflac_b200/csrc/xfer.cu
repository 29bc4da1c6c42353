// xfer.cu — small transfers that must not queue behind the bulk copies (sm_100a).
//
// The decode calls keep both PCIe copy engines busy with the stream bytes going up and the PCM coming
// back. A cudaMemcpyAsync of a few bytes (a candidate count, per-stream descriptors, MD5 jobs) is
// served by the same engines in submission order, so it waits for every bulk copy queued before it —
// measured r02b: the first sub-batch of a batch call took 25 ms instead of 6.6 ms, exactly the upload
// time of the seven sub-batches behind it. These few bytes therefore travel by load/store from the
// SMs: the host side of every small transfer is pinned (and so device-addressable under unified
// addressing), and this kernel copies between it and device memory on the pipeline's own stream.
#include "kernels.cuh"

namespace zf {

__global__ void __launch_bounds__(256) k_copy_small(uint8_t *__restrict__ dst, const uint8_t *__restrict__ src, size_t bytes) {
    const size_t tid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t nthreads = (size_t)gridDim.x * blockDim.x;
    if (((reinterpret_cast<uintptr_t>(dst) | reinterpret_cast<uintptr_t>(src)) & 15u) == 0) {
        const size_t n16 = bytes >> 4;
        const uint4 *s = reinterpret_cast<const uint4 *>(src);
        uint4 *d = reinterpret_cast<uint4 *>(dst);
        for (size_t i = tid; i < n16; i += nthreads) d[i] = s[i];
        for (size_t i = (n16 << 4) + tid; i < bytes; i += nthreads) dst[i] = src[i];
    } else {
        for (size_t i = tid; i < bytes; i += nthreads) dst[i] = src[i];
    }
}

cudaError_t launch_copy_small(void *dst, const void *src, size_t bytes, cudaStream_t st) {
    if (bytes == 0) return cudaSuccess;
    const size_t units = (bytes + 15) >> 4;
    const uint32_t blocks = (uint32_t)std::min<size_t>((units + 255) / 256, 296);
    ZF_COUNT_LAUNCH();
    k_copy_small<<<blocks, 256, 0, st>>>((uint8_t *)dst, (const uint8_t *)src, bytes);
    return cudaGetLastError();
}

}  // namespace zf
