// common.cuh — records shared by the host pipeline and the sm_100a kernels.
//
// Pipeline:
//   K1  sync_scan      coalesced scan of the compressed stream -> frame candidates
//                      (header grammar + CRC-8), ordered compaction (ballot + decoupled look-back)
//   K1b frame_link     CRC-16 over the bytes between neighbouring candidates -> verified frame
//                      ends without entropy decoding (the reference ignores the CRCs, Q1; here they
//                      are accelerators, the chain stays the truth)
//   K2  decode_frames  one thread per frame: Rice + CONSTANT/VERBATIM/FIXED/LPC -> planar scratch
//   K3  epilogue       decorrelation + MSB-align shift + narrow + interleave -> PCM
//   K4  md5            STREAMINFO MD5 gate over the PCM
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace zf {

// --- status codes: numerically identical to include/zflacb200.h ---
enum : uint32_t {
    ST_OK = 0,
    ST_INVALID_FRAME_HEADER = 6,
    ST_INVALID_CODED_NUMBER = 7,
    ST_INCONSISTENT_PARAMETERS = 8,
    ST_INVALID_SUBFRAME_HEADER = 9,
    ST_INVALID_RESIDUAL_CODING_METHOD = 10,
    ST_END_OF_STREAM = 11,
    ST_UNDEFINED_IN_REFERENCE = 90,
};

constexpr uint32_t kNoLink = 0xFFFFFFFFu;

// K1 output: one grammar- and CRC-8-valid frame header found in the stream.
struct Cand {
    uint64_t offset;      // byte offset of the sync code
    uint64_t number;      // coded frame / sample number (the reference parses and drops it, Q6; here a HINT only)
    uint32_t block_size;  // decoded block size (1..65535)
    uint32_t sr_value;    // frame sample rate as the reference computes it (0 = "from STREAMINFO")
    uint8_t header_len;   // bytes including the CRC-8
    uint8_t ch_code;      // 4-bit channel assignment
    uint8_t bd_code;      // 3-bit bit depth code
    uint8_t sr_code;      // 4-bit sample rate code
    uint8_t strategy;     // blocking strategy bit: 0 = number counts frames, 1 = number counts samples
    uint8_t pad[3];
};
static_assert(sizeof(Cand) == 32, "Cand layout");

// One stream of a launch (single decode: exactly one). Host-filled from the metadata walk.
struct StreamDesc {
    uint64_t begin;          // byte offset of the stream in the device buffer
    uint64_t end;            // one past its last byte
    uint64_t audio_begin;    // offset of the first frame
    uint64_t total;          // STREAMINFO total samples per channel (fast path needs > 0)
    uint64_t pcm_base;       // element offset of the stream's PCM in the output buffer
    uint64_t scratch_base;   // element offset of the stream's planes in the scratch
    uint32_t si_sample_rate;
    uint8_t si_channels;
    uint8_t si_depth;
    uint8_t pad[2];
};
static_assert(sizeof(StreamDesc) == 56, "StreamDesc layout");

// Per-stream result of the device-side chain check (k_verify_chain). The hinted chain is accepted
// only if fail == 0, first_ok == 1 and n_samples == total * channels; otherwise the host walks the
// chain itself (robust path) — the hints never decide what is returned.
struct StreamSummary {
    uint32_t fail;          // a chain link, a frame status or a stream-level rule did not hold
    uint32_t n_frames;      // frames on the chain
    uint64_t end_offset;    // just past the last frame
    uint64_t n_samples;     // interleaved samples covered by the chain
    uint32_t sample_rate;   // latched from the first frame (zflac.zig:376-388)
    uint8_t bps;
    uint8_t channels;
    uint8_t pad[2];
    uint32_t n_crc16_mismatch;
    uint32_t first_ok;      // the candidate at audio_begin exists and heads the chain
};
static_assert(sizeof(StreamSummary) == 40, "StreamSummary layout");

// K2 input: one frame of the chain. Offsets are absolute in the device buffer that holds the
// stream (batch mode: the packed buffer of all streams), so one launch can mix streams.
struct FrameRec {
    uint64_t offset;        // byte offset of the sync code
    uint64_t stream_end;    // end of the stream the frame belongs to (the reader's EndOfStream bound)
    uint64_t sample_base;   // interleaved-sample offset of the frame in the PCM buffer
    uint64_t scratch_base;  // element offset of the frame in the planar scratch
    uint32_t block_size;    // 0 = slot not in use (dropped candidate): K2/K3 skip it
    uint32_t scratch_pitch; // elements per channel plane (block_size rounded up to 8)
    uint8_t header_len;
    uint8_t ch_code;
    uint8_t channels;
    uint8_t bps;            // bits per sample latched from the first frame (zflac.zig:381-384)
    uint32_t stream_id;     // batch mode: which stream the frame belongs to
};
static_assert(sizeof(FrameRec) == 48, "FrameRec layout");

// K2 output.
struct FrameOut {
    uint64_t end_offset;  // byte offset just past the CRC-16 (valid when status == ST_OK)
    uint32_t status;
    uint32_t stored_crc16;
};
static_assert(sizeof(FrameOut) == 16, "FrameOut layout");

__host__ __device__ inline uint32_t round_up_u32(uint32_t v, uint32_t m) { return (v + m - 1) / m * m; }

}  // namespace zf
