// chain.cu — device-side frame chain for the fast path (sm_100a): k_hint_chain, k_verify_chain.
//
// The reference finds frame k+1 only after decoding frame k (src/zflac.zig:340-581): the chain is
// serial. Here every candidate header of the sync scan is turned into a frame record IN PARALLEL
// from a hint the reference parses and drops (Q6): the coded frame / sample number (:354,
// :203-214) says which samples the frame holds, so its PCM offset is known without a prefix sum
// and K2 can decode all candidates at once. Hints decide nothing: after K2, k_verify_chain checks
// the reference's own chain rule with the DECODED frame ends — first frame at the audio offset,
// frame k+1 starts at the byte where frame k ended (:546-548) and at the sample where it ended
// (:580), all frames consistent with the first (:376-392), block size 1 only last (:405), chain
// covers exactly the stated total (:341) — and the host falls back to walking the chain itself
// whenever one of them does not hold (variable streams the hints misjudge, unknown totals,
// damaged or faulty streams, every error case).
#include "common.cuh"
#include "kernels.cuh"

namespace zf {

__constant__ uint8_t c_ch_count[16] = {1, 2, 3, 4, 5, 6, 7, 8, 2, 2, 2, 0, 0, 0, 0, 0};
__constant__ uint8_t c_bps_of_code[8] = {0, 8, 12, 0, 16, 20, 24, 32};

constexpr int kChainThreads = 128;

__device__ __forceinline__ uint32_t find_stream(const StreamDesc *__restrict__ streams, uint32_t n_streams, uint64_t offset) {
    uint32_t lo = 0, hi = n_streams;  // last stream with begin <= offset
    while (hi - lo > 1) {
        const uint32_t mid = (lo + hi) >> 1;
        if (streams[mid].begin <= offset) lo = mid;
        else hi = mid;
    }
    return lo;
}

__device__ __forceinline__ uint32_t lower_bound_cand(const Cand *__restrict__ cands, uint32_t n, uint64_t offset) {
    uint32_t lo = 0, hi = n;
    while (lo < hi) {
        const uint32_t mid = (lo + hi) >> 1;
        if (cands[mid].offset < offset) lo = mid + 1;
        else hi = mid;
    }
    return lo;
}

// Is candidate c consistent with the stream's first frame, and which per-channel sample does it start at?
__device__ __forceinline__ bool hint_of(const Cand &c, const Cand &ref, const StreamDesc &sd, uint64_t &start) {
    const uint32_t rate_c = c.sr_code == 0 ? sd.si_sample_rate : c.sr_value;      // :376-379
    const uint32_t rate_r = ref.sr_code == 0 ? sd.si_sample_rate : ref.sr_value;
    if (rate_c != rate_r || c_ch_count[c.ch_code] != c_ch_count[ref.ch_code] || c.bd_code != ref.bd_code) return false;  // :391
    if (c.strategy != ref.strategy || c.number < ref.number) return false;
    const uint64_t d = c.number - ref.number;
    start = c.strategy ? d : d * ref.block_size;
    return start + c.block_size <= sd.total;
}

__global__ void __launch_bounds__(kChainThreads)
k_hint_chain(const Cand *__restrict__ cands, uint32_t first, uint32_t n, const StreamDesc *__restrict__ streams, uint32_t n_streams,
             FrameRec *__restrict__ recs, uint32_t window, unsigned long long *__restrict__ pcm_hi, RangeHint rh) {
    const uint32_t i = first + blockIdx.x * kChainThreads + threadIdx.x;
    if (i >= n) return;
    const Cand c = cands[i];
    FrameRec fr;
    fr.offset = c.offset;
    fr.block_size = 0;  // dropped unless proven plausible below
    fr.stream_end = 0;
    fr.sample_base = fr.scratch_base = 0;
    fr.scratch_pitch = 0;
    fr.header_len = c.header_len;
    fr.ch_code = c.ch_code;
    fr.channels = 0;
    fr.bps = 0;
    const uint32_t sid = n_streams == 1 ? 0 : find_stream(streams, n_streams, c.offset);
    fr.stream_id = sid;
    const StreamDesc sd = streams[sid];
    const uint32_t r = n_streams == 1 ? 0 : lower_bound_cand(cands, n, sd.audio_begin);
    bool keep = false;
    uint64_t start = 0;
    if (c.offset >= sd.audio_begin && c.offset < sd.end && r < n && sd.total > 0) {
        const Cand ref = cands[r];
        const uint32_t bps = ref.bd_code == 0 ? sd.si_depth : c_bps_of_code[ref.bd_code];
        if (ref.offset == sd.audio_begin && ref.bd_code != 3 && c_ch_count[ref.ch_code] == sd.si_channels &&  // :386
            hint_of(c, ref, sd, start)) {
            if (c.offset == sd.audio_begin) {
                keep = start == 0;
            } else {  // some earlier candidate must end, by its own hint, where this one starts
                const uint32_t floor_j = r > rh.window_floor ? r : rh.window_floor;  // range decode: slot 0 is a stand-in
                const uint32_t lo = i - floor_j > window ? i - window : floor_j;
                for (uint32_t j = i; j-- > lo && !keep;) {
                    const Cand p = cands[j];
                    uint64_t ps;
                    keep = hint_of(p, ref, sd, ps) && ps + p.block_size == start;
                }
            }
            // range decode: candidates below the shard are predecessors for the rule above only
            if (c.offset < rh.decode_lo) keep = false;
            if (keep) {
                const uint32_t C = sd.si_channels;
                fr.stream_end = sd.end;
                fr.block_size = c.block_size;
                fr.scratch_pitch = round_up_u32(c.block_size, 8);
                fr.sample_base = sd.pcm_base + start * C;
                // planes of different frames never overlap: C * roundup8(start) grows by at least
                // C * (pitch - 8) per frame and the 64 * i term adds the missing 8 * C (C <= 8)
                fr.scratch_base = sd.scratch_base + (uint64_t)C * ((start + 7) & ~7ull) + 64ull * i;
                fr.channels = (uint8_t)C;
                fr.bps = (uint8_t)bps;
                if (pcm_hi) atomicMax(pcm_hi, (unsigned long long)(fr.sample_base + (uint64_t)c.block_size * C));
                if (rh.lo_out) {  // range decode, first pass: where the shard starts, in samples and in bytes
                    atomicMin(rh.lo_out, (unsigned long long)fr.sample_base);
                    atomicMin(rh.lo_out + 1, (unsigned long long)c.offset);
                    atomicAdd(rh.lo_out + 2, 1ull);
                }
            }
        }
    }
    recs[i] = fr;
}

__global__ void __launch_bounds__(kChainThreads)
k_verify_chain(const Cand *__restrict__ cands, uint32_t n, const StreamDesc *__restrict__ streams, uint32_t n_streams,
               const FrameRec *__restrict__ recs, const FrameOut *__restrict__ outs, StreamSummary *__restrict__ sums,
               uint32_t window, uint32_t open_end) {
    const uint32_t i = blockIdx.x * kChainThreads + threadIdx.x;
    if (i >= n) return;
    const FrameRec fr = recs[i];
    if (fr.block_size == 0) return;
    const uint32_t sid = fr.stream_id;
    const StreamDesc sd = streams[sid];
    const FrameOut o = outs[i];
    const uint32_t C = fr.channels;
    const uint64_t s_end = fr.sample_base + (uint64_t)fr.block_size * C;
    const uint64_t total_end = sd.pcm_base + sd.total * C;
    bool ok = o.status == ST_OK;
    // next frame record of the same stream
    uint32_t j = i + 1;
    const uint32_t stop = n - i - 1 > window ? i + 1 + window : n;
    while (j < stop && (recs[j].block_size == 0 || recs[j].stream_id != sid)) {
        if (recs[j].offset >= sd.end) { j = n; break; }  // ran into the next stream: i is this stream's last record
        ++j;
    }
    if (j < stop) {
        ok = ok && o.end_offset == recs[j].offset && recs[j].sample_base == s_end;  // :546-548, :580
        if (fr.block_size == 1) ok = false;  // block size 1 before the end of the stream (:405)
    } else if (j < n && cands[j].offset < sd.end) {
        ok = false;  // more candidates than the window covers: cannot tell, let the host walk
    } else {
        if (!open_end) ok = ok && s_end == total_end;  // the chain must cover exactly the stated total (:341)
        sums[sid].end_offset = o.end_offset;        // (range decode: the host checks the ends of adjacent shards instead)
    }
    if (fr.offset == sd.audio_begin) {  // head of the chain: latch what DecodedFLAC reports (:376-388)
        const Cand c = cands[i];
        StreamSummary &s = sums[sid];
        s.sample_rate = (c.sr_code == 0 ? sd.si_sample_rate : c.sr_value) & 0xFFFFFFu;
        s.bps = fr.bps;
        s.channels = fr.channels;
        s.first_ok = (fr.sample_base == sd.pcm_base) ? 1u : 0u;
    }
    // one atomic per warp and stream instead of one per frame (a 1-hour stream: 38 760 frames on ONE summary)
    const uint32_t active = __activemask();
    const uint32_t peers = __match_any_sync(active, sid);
    const int leader = __ffs(peers) - 1;
    const int lane = threadIdx.x & 31;
    uint32_t bad = ok ? 0u : 1u, frames = 1u;
    unsigned long long samples = (unsigned long long)fr.block_size * C;
    for (uint32_t rest = peers & ~(1u << leader); rest; rest &= rest - 1) {  // leader gathers its peers' values
        const int src = __ffs(rest) - 1;
        const uint32_t b2 = __shfl_sync(peers, bad, src);
        const unsigned long long s2 = __shfl_sync(peers, samples, src);
        if (lane == leader) {
            bad |= b2;
            frames += 1u;
            samples += s2;
        }
    }
    if (lane == leader) {
        if (bad) atomicOr(&sums[sid].fail, 1u);
        atomicAdd(&sums[sid].n_frames, frames);
        atomicAdd((unsigned long long *)&sums[sid].n_samples, samples);
    }
}

cudaError_t launch_hint_chain(const Cand *d_cands, uint32_t n, const StreamDesc *d_streams, uint32_t n_streams,
                              FrameRec *d_recs, uint32_t window, cudaStream_t st) {
    return launch_hint_chain_range(d_cands, 0, n, d_streams, n_streams, d_recs, window, nullptr, st);
}

cudaError_t launch_hint_chain_range(const Cand *d_cands, uint32_t first, uint32_t n, const StreamDesc *d_streams, uint32_t n_streams,
                                    FrameRec *d_recs, uint32_t window, unsigned long long *d_pcm_hi, cudaStream_t st) {
    return launch_hint_chain_shard(d_cands, first, n, d_streams, n_streams, d_recs, window, d_pcm_hi, RangeHint{}, st);
}

cudaError_t launch_hint_chain_shard(const Cand *d_cands, uint32_t first, uint32_t n, const StreamDesc *d_streams, uint32_t n_streams,
                                    FrameRec *d_recs, uint32_t window, unsigned long long *d_pcm_hi, RangeHint rh, cudaStream_t st) {
    if (n <= first) return cudaSuccess;
    ZF_COUNT_LAUNCH();
    k_hint_chain<<<(n - first + kChainThreads - 1) / kChainThreads, kChainThreads, 0, st>>>(d_cands, first, n, d_streams, n_streams,
                                                                                            d_recs, window, d_pcm_hi, rh);
    return cudaGetLastError();
}

cudaError_t launch_verify_chain(const Cand *d_cands, uint32_t n, const StreamDesc *d_streams, uint32_t n_streams,
                                const FrameRec *d_recs, const FrameOut *d_outs, StreamSummary *d_sums, uint32_t window,
                                cudaStream_t st, bool open_end) {
    if (n == 0) return cudaSuccess;
    ZF_COUNT_LAUNCH();
    k_verify_chain<<<(n + kChainThreads - 1) / kChainThreads, kChainThreads, 0, st>>>(d_cands, n, d_streams, n_streams, d_recs,
                                                                                      d_outs, d_sums, window, open_end ? 1u : 0u);
    return cudaGetLastError();
}

}  // namespace zf
