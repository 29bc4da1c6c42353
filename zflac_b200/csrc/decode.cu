// decode.cu — K2 decode_frames (sm_100a): one lane per frame, frame bytes staged through a
// per-lane shared-memory word ring that cp.async (LDGSTS) keeps ahead of the bit cursor.
//
// The per-frame decoder itself (subframe loop, Rice, CONSTANT/VERBATIM/FIXED/LPC; reference
// src/zflac.zig:425-551, :614-666, src/bit_reader.zig) lives in decode_core.cuh; this file is the
// device word source and the kernel.
//
// Staging. Lane L of a CTA owns column L of a ring of kRingWords rows x NT columns of 32-bit
// words: stream word w of the lane's frame sits at ring[w % kRingWords][L]. Every shared load of
// the bit reader is therefore bank-conflict free whatever the lanes' cursors are (a lane only
// ever touches bank L % 32), and the fill is a 4-byte cp.async.ca per word — L1 keeps the 32-byte
// sector, so DRAM sees every compressed byte once. Fills happen at a uniform cadence (before each
// 8-row body) so the warp does not diverge on "my queue is empty"; completion is tracked with
// cp.async groups: a commit per refill, boundaries of the three newest groups remembered, and a
// wait_group cascade that only blocks on groups old enough to have landed.
#include "decode_core.cuh"
#include "kernels.cuh"
#include "ring_src.cuh"

namespace zf {

// ST = int32_t: streams whose frames carry <= 31 bits per sample (subframe depth <= 32).
// ST = int64_t: 32 bps streams.
template <typename ST>
__global__ void __launch_bounds__(kDecodeThreads)
k_decode_frames(const uint8_t *__restrict__ s, const FrameRec *__restrict__ frames, uint32_t n_frames,
                ST *__restrict__ scratch, FrameOut *__restrict__ outs, int write_samples, int inter64) {
    const uint32_t f = blockIdx.x * kDecodeThreads + threadIdx.x;
    if (f >= n_frames) return;
    const FrameRec fr = frames[f];
    if (fr.block_size == 0) return;  // dropped candidate slot
    const uint64_t start = fr.offset + fr.header_len;  // first subframe header
    const uint64_t origin = start & ~3ull;             // word 0 of the lane's source
    BitReader<RingSrc<kDecodeThreads>> br;
    br.src.init(s + origin, fr.stream_end > origin ? fr.stream_end - origin : 0, threadIdx.x);
    br.init((uint32_t)(start & 3));
    const uint64_t len_bits = (fr.stream_end > origin ? fr.stream_end - origin : 0) * 8;
    FrameOut o;
    decode_frame<ST>(br, fr, scratch, len_bits, write_samples != 0, origin, o, inter64 != 0);
    asm volatile("cp.async.wait_group 0;" ::: "memory");  // nothing may land in the ring after the CTA retires
    outs[f] = o;
}

cudaError_t launch_decode_frames(const uint8_t *d_stream, const FrameRec *d_frames, uint32_t n_frames, void *d_scratch,
                                 bool wide_scratch, FrameOut *d_outs, bool write_samples, bool inter64, cudaStream_t st) {
    if (n_frames == 0) return cudaSuccess;
    const uint32_t blocks = (n_frames + kDecodeThreads - 1) / kDecodeThreads;
    ZF_COUNT_LAUNCH();
    if (wide_scratch)
        k_decode_frames<int64_t><<<blocks, kDecodeThreads, kRingWords * kDecodeThreads * sizeof(uint32_t), st>>>(d_stream, d_frames, n_frames, (int64_t *)d_scratch, d_outs,
                                                                    write_samples ? 1 : 0, inter64 ? 1 : 0);
    else
        k_decode_frames<int32_t><<<blocks, kDecodeThreads, kRingWords * kDecodeThreads * sizeof(uint32_t), st>>>(d_stream, d_frames, n_frames, (int32_t *)d_scratch, d_outs,
                                                                    write_samples ? 1 : 0, inter64 ? 1 : 0);
    return cudaGetLastError();
}

}  // namespace zf
