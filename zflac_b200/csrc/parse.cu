// parse.cu — K2a parse_frames (sm_100a): pass A of the two-pass subframe decoder. One lane per
// frame walks the frame's serial bit chain (subframe headers, warm-up samples, coefficients, Rice
// partitions; reference src/zflac.zig:425-551, :614-666 and src/bit_reader.zig) and leaves the raw
// material of every subframe in HBM: residuals / warm-up / verbatim samples in the planar int32
// scratch, one SubRec per subframe. The predictor recurrence itself runs in pass B (predict.cu),
// one lane per subframe. The logic is parse_core.cuh; the word source is StagedSrc of ring_src.cuh.
#include "kernels.cuh"
#include "parse_core.cuh"
#include "parse_narrow.cuh"
#include "ring_src.cuh"

namespace zf {

// SRC = StagedSrc (16-word warp-synchronous refills, 64-word ring, 116 registers: 512 lanes per SM, one-warp CTAs) or
// QuadSrc (4-word refills, 32-word ring, 72 registers: 896 lanes per SM, 128-thread CTAs). The kernel is latency-bound,
// so what counts is that ALL frames of a launch are resident at once and evenly spread: the launcher picks QuadSrc when
// the frames do not fit one wave of the first.
template <class SRC>
__global__ void __launch_bounds__(SRC::kThreads, SRC::kMinBlocks)
k_parse_frames(const uint8_t *__restrict__ s, const FrameRec *__restrict__ frames, uint32_t n_frames,
               int32_t *__restrict__ scratch, SubRec *__restrict__ subs, FrameOut *__restrict__ outs,
               uint32_t *__restrict__ big_orders, int write_samples) {
    const uint32_t f = blockIdx.x * SRC::kThreads + threadIdx.x;
    if (f >= n_frames) return;
    const FrameRec fr = frames[f];
    if (fr.block_size == 0) return;  // dropped candidate slot
    const uint64_t start = fr.offset + fr.header_len;  // first subframe header
    // word 0 of the lane's source: the 16-byte aligned ADDRESS at or below the first subframe header
    const uint32_t lead = (uint32_t)(reinterpret_cast<uintptr_t>(s + start) & 15u);
    const uint64_t origin = start - lead;  // may reach up to 15 bytes below offset 0 of an unaligned buffer (same allocation)
    const uint64_t span = fr.stream_end > start ? fr.stream_end - start + lead : 0;
    BitReader<SRC> br;
    br.src.init(s + start - lead, span, threadIdx.x);
    br.init_at(lead);
    const uint64_t len_bits = span * 8;
    FrameOut o;
    uint32_t max_order = 0;
    parse_frame(br, fr, scratch, subs + (size_t)f * 8, len_bits, write_samples != 0, origin, o, max_order);
    if (max_order > 12 && write_samples) *big_orders = 1u;  // pass B then also runs its 32-tap kernel
    outs[f] = o;
}

// Narrow K2a (parse_narrow.cuh): streams of <= 16 bits per sample. One-warp CTAs, one lane per frame, PosReader over a
// PairSrc ring in static shared memory; the scratch is uint16 (folded residuals as coded). flags[0]: some subframe has an
// order above 12 (pass B then also runs its 32-tap kernel); flags[1]: a value did not fit 16 bits (the host redoes the
// launch group on the int32 path).
// WIDE: the same kernel with int32 rows (parse_narrow.cuh) for streams of 17..31 bits per sample, and for launch groups
// of narrower streams that reported a value beyond 16 bits.
template <int RING, bool WIDE = false>
__global__ void __launch_bounds__(32, RING >= 64 ? 14 : 20)
k_parse_frames_narrow(const uint8_t *__restrict__ s, const FrameRec *__restrict__ frames, uint32_t n_frames,
                      uint16_t *__restrict__ scratch, SubRec *__restrict__ subs, FrameOut *__restrict__ outs,
                      uint32_t *__restrict__ flags, int write_samples) {
    const uint32_t f = blockIdx.x * 32u + threadIdx.x;
    FrameRec fr;
    fr.block_size = 0;
    fr.channels = 0;
    fr.scratch_base = 0;
    if (f < n_frames) fr = frames[f];
    // layout of the warp's rows (parse_narrow.cuh): interleaved when its 32 frames are alike and their regions in order
    // (one channel count, 32 live frames, regions in stream order, and together at least as large as 32 columns of the
    // LONGEST frame's rows: 16 * C * roundup8(max block size) elements — true when most frames have the stream's block
    // size, e.g. a warp that holds the short last frames of a few files of a batch)
    const uint32_t c0 = __shfl_sync(0xFFFFFFFFu, (uint32_t)fr.channels, 0);
    const uint32_t bs_max = __reduce_max_sync(0xFFFFFFFFu, fr.block_size);
    const uint64_t base0 = __shfl_sync(0xFFFFFFFFu, fr.scratch_base, 0);
    const uint64_t end31 = __shfl_sync(0xFFFFFFFFu, fr.scratch_base + (uint64_t)fr.channels * ((fr.block_size + 7u) & ~7u), 31);
    const bool alike = fr.block_size != 0 && fr.channels == c0 && fr.scratch_base >= base0;
    const bool interleave = __all_sync(0xFFFFFFFFu, alike) && end31 >= base0 + (WIDE ? 32ull : 16ull) * c0 * ((bs_max + 7u) & ~7u);
    if (fr.block_size == 0) return;  // dropped candidate slot / past the end
    NarrowSink sink;
    if (interleave) sink.set_warp(scratch, base0, threadIdx.x, bs_max, WIDE);
    else sink.set_frame(scratch, fr.scratch_base, fr.channels, WIDE);
    const uint64_t start = fr.offset + fr.header_len;  // first subframe header
    const uint32_t lead = (uint32_t)(reinterpret_cast<uintptr_t>(s + start) & 15u);
    const uint64_t origin = start - lead;
    const uint64_t span = fr.stream_end > start ? fr.stream_end - start + lead : 0;
    PosReader<PairSrc<RING>> br;
    br.src.init(s + start - lead, span, threadIdx.x);
    br.init_at(lead);
    const uint64_t len_bits = span * 8;
    FrameOut o;
    uint32_t max_order = 0, overflow = 0;
    parse_frame_narrow<PairSrc<RING>, WIDE>(br, fr, sink, subs + (size_t)f * 8, len_bits, write_samples != 0, origin, o, max_order, overflow);
    if (max_order > 12 && write_samples) flags[0] = 1u;
    if (overflow && write_samples) flags[1] = 1u;
    outs[f] = o;
}

cudaError_t launch_parse_frames_narrow(const uint8_t *d_stream, const FrameRec *d_frames, uint32_t n_frames, uint16_t *d_scratch,
                                       SubRec *d_subs, FrameOut *d_outs, uint32_t *d_flags, bool write_samples, cudaStream_t st) {
    if (n_frames == 0) return cudaSuccess;
    ZF_COUNT_LAUNCH();
    static int sms = 0;
    if (sms == 0) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        if (sms <= 0) sms = 148;
    }
    const uint32_t warps = (n_frames + 31u) / 32u;
    // the deep ring (64 entries: a refill every other body never runs dry) while one wave of it holds every frame
    g_pass_a_launches[warps <= (uint32_t)sms * 14u ? kPassANarrow64 : kPassANarrow32].fetch_add(1, std::memory_order_relaxed);
    if (warps <= (uint32_t)sms * 14u)
        k_parse_frames_narrow<64><<<warps, 32, 0, st>>>(d_stream, d_frames, n_frames, d_scratch, d_subs, d_outs, d_flags, write_samples ? 1 : 0);
    else
        k_parse_frames_narrow<32><<<warps, 32, 0, st>>>(d_stream, d_frames, n_frames, d_scratch, d_subs, d_outs, d_flags, write_samples ? 1 : 0);
    return cudaGetLastError();
}

cudaError_t launch_parse_frames(const uint8_t *d_stream, const FrameRec *d_frames, uint32_t n_frames, int32_t *d_scratch,
                                SubRec *d_subs, FrameOut *d_outs, uint32_t *d_big_orders, bool write_samples, cudaStream_t st,
                                bool word_pipeline) {
    if (n_frames == 0) return cudaSuccess;
    const uint32_t blocks = (n_frames + kDecodeThreads - 1) / kDecodeThreads;
    ZF_COUNT_LAUNCH();
    static int sms = 0;
    if (sms == 0) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        if (sms <= 0) sms = 148;
    }
    if (!word_pipeline) {  // the position reader with int32 rows (parse_narrow.cuh, WIDE): the default
        const uint32_t warps = (n_frames + 31u) / 32u;
        const bool deep = warps <= (uint32_t)sms * 14u;
        g_pass_a_launches[deep ? kPassAWide64 : kPassAWide32].fetch_add(1, std::memory_order_relaxed);
        if (deep)
            k_parse_frames_narrow<64, true><<<warps, 32, 0, st>>>(d_stream, d_frames, n_frames, reinterpret_cast<uint16_t *>(d_scratch), d_subs,
                                                                  d_outs, d_big_orders, write_samples ? 1 : 0);
        else
            k_parse_frames_narrow<32, true><<<warps, 32, 0, st>>>(d_stream, d_frames, n_frames, reinterpret_cast<uint16_t *>(d_scratch), d_subs,
                                                                  d_outs, d_big_orders, write_samples ? 1 : 0);
        return cudaGetLastError();
    }
    // One-warp CTAs for the StagedSrc kernel: with 128-thread CTAs a 1-hour stream is 303 CTAs on 148 SMs, seven SMs
    // carry three CTAs (3 warps per scheduler instead of 2) and the whole launch waits 26 % longer for them
    // (sm__cycles_active max 1.77 M vs avg 1.40 M, r02m); 1 212 one-warp CTAs spread the 28 extra warps over 28 SMs.
    constexpr int kStagedThreads = 32;
    g_pass_a_launches[blocks <= (uint32_t)sms * 4u ? kPassAStaged : kPassAQuad].fetch_add(1, std::memory_order_relaxed);
    if (blocks <= (uint32_t)sms * 4u) {
        const uint32_t warps = (n_frames + kStagedThreads - 1) / kStagedThreads;
        k_parse_frames<StagedSrc<kStagedThreads>><<<warps, kStagedThreads, kStagedRing * kStagedThreads * sizeof(uint32_t), st>>>(
            d_stream, d_frames, n_frames, d_scratch, d_subs, d_outs, d_big_orders, write_samples ? 1 : 0);
    } else {
        // (one-warp CTAs were slower here, r02p: C5 3.83 -> 4.30 ms; 860 CTAs of 128 already spread to within one CTA per SM)
        k_parse_frames<QuadSrc<kDecodeThreads>><<<blocks, kDecodeThreads, kRingWords * kDecodeThreads * sizeof(uint32_t), st>>>(
            d_stream, d_frames, n_frames, d_scratch, d_subs, d_outs, d_big_orders, write_samples ? 1 : 0);
    }
    return cudaGetLastError();
}

}  // namespace zf
