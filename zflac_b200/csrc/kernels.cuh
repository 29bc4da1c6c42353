// kernels.cuh — host-side launchers of the sm_100a kernels (one translation unit each).
#pragma once
#include <atomic>

#include "common.cuh"

namespace zf {

// every kernel launch of the library is counted here (zflacb200_kernel_launches)
extern std::atomic<unsigned long long> g_kernel_launches;
#define ZF_COUNT_LAUNCH() (::zf::g_kernel_launches.fetch_add(1, std::memory_order_relaxed))
// which instantiation of pass A a launch took (zflacb200_pass_a_launches): the parity tests make every variant run
enum { kPassAStaged = 0, kPassAQuad = 1, kPassANarrow64 = 2, kPassANarrow32 = 3, kPassAWide64 = 4, kPassAWide32 = 5, kPassAVariants = 6 };
extern std::atomic<unsigned long long> g_pass_a_launches[kPassAVariants];

// scan.cu
cudaError_t scan_init_constants();
void set_last_error(const char *msg);  // api.cu: what zflacb200_last_error() reports for the calling thread
uint32_t scan_tile_count(uint64_t scan_begin, uint64_t len);
cudaError_t launch_sync_scan(const uint8_t *d_stream, uint64_t scan_begin, uint64_t len, Cand *d_cands, uint32_t cap,
                             uint32_t *d_total, unsigned long long *d_tile_state, uint32_t *d_ticket, cudaStream_t st);
cudaError_t launch_sync_scan_range(const uint8_t *d_stream, uint64_t scan_begin, uint64_t scan_end, uint64_t len, Cand *d_cands,
                                   uint32_t cap, uint32_t *d_total, unsigned long long *d_tile_state, uint32_t *d_ticket,
                                   cudaStream_t st);
cudaError_t launch_frame_link(const uint8_t *d_stream, uint64_t len, const Cand *d_cands, uint32_t n, uint32_t *d_link,
                              uint32_t window, uint64_t max_frame_bytes, cudaStream_t st);
cudaError_t launch_crc16_ranges(const uint8_t *d_stream, const uint64_t *d_begin, const uint64_t *d_end, uint32_t n,
                                uint16_t *d_crc, cudaStream_t st);

cudaError_t launch_crc16_frames(const uint8_t *d_stream, const FrameRec *d_recs, const FrameOut *d_outs, uint32_t n,
                                StreamSummary *d_sums, uint8_t *d_frame_ok, cudaStream_t st);

// chain.cu
cudaError_t launch_hint_chain(const Cand *d_cands, uint32_t n, const StreamDesc *d_streams, uint32_t n_streams,
                              FrameRec *d_recs, uint32_t window, cudaStream_t st);
// Chunked variant: records for candidates [first, n) only (earlier ones exist already); *d_pcm_hi is raised to the
// largest PCM element end among the kept records (what may be copied out once these frames are decoded).
cudaError_t launch_hint_chain_range(const Cand *d_cands, uint32_t first, uint32_t n, const StreamDesc *d_streams, uint32_t n_streams,
                                    FrameRec *d_recs, uint32_t window, unsigned long long *d_pcm_hi, cudaStream_t st);
// Range decode of ONE stream (a shard of its frames): slot 0 of the candidate list is a stand-in for the
// stream's first frame (window_floor = 1 keeps it out of the predecessor search), candidates below decode_lo only serve
// as predecessors, and lo_out (optional) receives {min sample_base, min offset, count} of the records kept.
struct RangeHint {
    uint64_t decode_lo = 0;
    uint32_t window_floor = 0;
    unsigned long long *lo_out = nullptr;
};
cudaError_t launch_hint_chain_shard(const Cand *d_cands, uint32_t first, uint32_t n, const StreamDesc *d_streams, uint32_t n_streams,
                                    FrameRec *d_recs, uint32_t window, unsigned long long *d_pcm_hi, RangeHint rh, cudaStream_t st);
// open_end: the last record of the chain need not end at the stream's total (range decode: the host checks adjacent shards)
cudaError_t launch_verify_chain(const Cand *d_cands, uint32_t n, const StreamDesc *d_streams, uint32_t n_streams,
                                const FrameRec *d_recs, const FrameOut *d_outs, StreamSummary *d_sums, uint32_t window,
                                cudaStream_t st, bool open_end = false);

// decode.cu
cudaError_t launch_decode_frames(const uint8_t *d_stream, const FrameRec *d_frames, uint32_t n_frames, void *d_scratch,
                                 bool wide_scratch, FrameOut *d_outs, bool write_samples, bool inter64, cudaStream_t st);

// parse.cu + predict.cu: the two-pass decoder for streams with <= 31 bits per sample
struct SubRec;  // parse_core.cuh
constexpr size_t kSubRecBytes = 80;  // sizeof(SubRec); 8 slots per frame record
cudaError_t launch_parse_frames(const uint8_t *d_stream, const FrameRec *d_frames, uint32_t n_frames, int32_t *d_scratch,
                                SubRec *d_subs, FrameOut *d_outs, uint32_t *d_big_orders, bool write_samples, cudaStream_t st,
                                bool word_pipeline = false);
// narrow K2a (streams of <= 16 bits per sample): uint16 scratch of folded residuals; d_flags = {big orders, 16-bit overflow}
cudaError_t launch_parse_frames_narrow(const uint8_t *d_stream, const FrameRec *d_frames, uint32_t n_frames, uint16_t *d_scratch,
                                       SubRec *d_subs, FrameOut *d_outs, uint32_t *d_flags, bool write_samples, cudaStream_t st);
struct PredictParams {
    const FrameRec *frames;
    const FrameOut *outs;      // frames whose status is not ST_OK are skipped
    uint32_t n_frames;
    uint32_t max_channels;     // largest channel count among the frames (sizes the lane groups)
    const void *scratch;       // int32 planes written by launch_parse_frames
    const void *subs;          // SubRec[8 * n_frames]
    void *pcm;
    uint32_t container_bytes;  // 1, 2, 4
    uint32_t msb_shift;
    uint32_t *lossy_flag;
    const uint32_t *big_orders;  // written by launch_parse_frames: != 0 iff some subframe has an order above 12
    bool narrow = false;         // scratch is uint16 folded residuals (launch_parse_frames_narrow)
};
cudaError_t launch_predict_epilogue(const PredictParams &p, cudaStream_t st);

// epilogue.cu
struct EpilogueParams {
    const FrameRec *frames;
    uint32_t n_frames;
    uint32_t max_block_size;
    const void *scratch;
    bool wide_scratch;
    void *pcm;
    uint32_t container_bytes;  // 1, 2, 4
    uint32_t msb_shift;        // left shift applied to every container value (zflac.zig:287-306)
    uint32_t *lossy_flag;      // set to 1 if a shifted value no longer shifts back (MD5 must then hash pre-shift data)
};
cudaError_t launch_epilogue(const EpilogueParams &p, cudaStream_t st);
// xfer.cu: copies `bytes` between device memory and PINNED host memory (either direction) from the SMs.
cudaError_t launch_copy_small(void *dst, const void *src, size_t bytes, cudaStream_t st);
cudaError_t launch_shift_inplace(void *pcm, uint64_t n_samples, uint32_t container_bytes, uint32_t shift, cudaStream_t st);

struct Md5Job {
    const void *pcm;        // container-typed samples
    uint64_t n_samples;
    uint32_t container_bytes;
    uint32_t hash_bytes;    // bytes of each sample that enter the hash: 1, 2, 3 or 4 (Q4)
    uint32_t unshift;       // arithmetic right shift that undoes the MSB-align shift before hashing
    uint32_t pad;
};
cudaError_t launch_md5(const Md5Job *d_jobs, uint32_t n_jobs, uint8_t *d_digests, cudaStream_t st);

// convert.cu: output conventions (zflacb200_read_as). Values match zflacb200_format in include/zflacb200.h.
constexpr int kFmtNative = 0, kFmtF32 = 1, kFmtS24 = 2, kFmtU8 = 3;
cudaError_t launch_convert_pcm(const void *pcm, uint64_t n_samples, uint32_t container_bytes, int fmt, void *out, cudaStream_t st);

}  // namespace zf
