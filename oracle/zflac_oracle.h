/*
 * zflac_oracle.h — TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * CPU restatement (C++17, C ABI) of Senryoku/zflac's decode path, used only as
 * the parity checker and as the timed CPU baseline:
 *   - tests/                         (checker)
 *   - __graft_entry__.smoke()        (checker)
 *   - bench.py cpu_baseline / --impl reference   (timed baseline)
 * Nothing under zflac_b200/ may include, link, import or execute this.
 *
 * It follows /root/reference/src/zflac.zig:187-666 and
 * /root/reference/src/bit_reader.zig:1-126 function for function (each
 * function in zflac_oracle.cpp cites the lines it restates). The reference is
 * Zig 0.14.1; no Zig toolchain exists in this image, so the reference itself
 * cannot be compiled here (no oracle/_ref). Parity is PINNED instead by the
 * reference's own tests: the 3 RFC 9639 known-answer vectors of
 * tests/basic.zig, the 39 available whole-stream golden .raw files of
 * tests/expected_samples, the STREAMINFO MD5 of every decodable corpus file and
 * the 12 expected error names of tests/std_uncommon.zig / tests/std_faulty.zig
 * (see tests/test_oracle_*.py and tests/golden/manifest.json).
 */
#ifndef ZFLAC_ORACLE_H
#define ZFLAC_ORACLE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* One value per Zig error name reachable from decode(). The
 * numeric values are shared with include/zflacb200.h so tests can compare
 * either numbers or names. */
enum zo_status {
    ZO_OK = 0,
    ZO_INVALID_SIGNATURE = 1,             /* zflac.zig:220 */
    ZO_INVALID_METADATA_HEADER = 2,       /* :248 */
    ZO_UNIMPLEMENTED = 3,                 /* :263 (unreachable in practice) */
    ZO_INVALID_CHECKSUM = 4,              /* :280 */
    ZO_MISSING_STREAMINFO = 5,            /* :309 */
    ZO_INVALID_FRAME_HEADER = 6,          /* :352,357,361,372,405 */
    ZO_INVALID_CODED_NUMBER = 7,          /* :206 */
    ZO_INCONSISTENT_PARAMETERS = 8,       /* :386,391 */
    ZO_INVALID_SUBFRAME_HEADER = 9,       /* :431,471,542 */
    ZO_INVALID_RESIDUAL_CODING_METHOD = 10, /* :618 */
    ZO_END_OF_STREAM = 11,                /* reader error, e.g. :344 */
    ZO_OUT_OF_MEMORY = 12,
    /* The reference hits `unreachable`, an out-of-range @intCast, an
     * out-of-bounds slice or an integer underflow: a panic in Debug, undefined
     * behaviour in ReleaseFast. No parity is defined for such inputs. */
    ZO_UNDEFINED_IN_REFERENCE = 90
};

/* One decoded frame, in stream order (diagnostic trace; lets tests check the
 * GPU frame index and chain against the oracle). */
typedef struct zo_frame {
    uint64_t offset;        /* byte offset of the sync code */
    uint64_t end_offset;    /* byte offset just past the CRC-16 */
    uint64_t sample_offset; /* interleaved-sample offset of the frame's first sample */
    uint32_t block_size;
    uint8_t header_len;     /* bytes, including the CRC-8 */
    uint8_t channel_code;   /* 4-bit channel assignment field */
    uint8_t bit_depth_code; /* 3-bit field */
    uint8_t sample_rate_code;
} zo_frame;

typedef struct zo_result {
    int32_t status;           /* enum zo_status */
    uint8_t channels;         /* DecodedFLAC.channels   (zflac.zig:19) */
    uint8_t bits_per_sample;  /* DecodedFLAC.bits_per_sample (:21), from the first frame header (:381-384) */
    uint8_t container_bytes;  /* 1 = s8, 2 = s16, 4 = s32 (union tag, :259-264) */
    uint8_t overflow_flag;    /* 1 if an arithmetic result left the reference's type (UB in ReleaseFast, panic in Debug) */
    uint32_t sample_rate;     /* DecodedFLAC.sample_rate (:20), u24 */
    uint64_t n_samples;       /* samples.len: interleaved sample count */
    void *samples;            /* 32-byte aligned, n_samples * container_bytes bytes, post MSB-align shift (:287-306) */
    uint8_t md5_computed[16]; /* digest the reference computes at :267-277 (valid when all frames parsed) */
    uint8_t md5_streaminfo[16];
    uint64_t n_frames;
    zo_frame *frames;         /* n_frames entries, or NULL if tracing was off */
    uint64_t audio_offset;    /* byte offset of the first frame */
    /* timing breakdown of the last call, seconds (steady_clock) */
    double t_frames, t_md5, t_shift;
} zo_result;

#define ZO_FLAG_TRACE_FRAMES 1u  /* fill result->frames */
#define ZO_FLAG_KEEP_ON_MD5_MISMATCH 2u /* still return samples on InvalidChecksum (diagnostics) */

/* Restates `pub fn decode(allocator, reader) !DecodedFLAC` (zflac.zig:217).
 * Returns result->status. On error no samples are returned (errdefer :265,332). */
int zo_decode(const uint8_t *data, size_t len, uint32_t flags, zo_result *result);
void zo_free(zo_result *result);
const char *zo_status_name(int status);

/* helpers exposed for tests of the checksums the reference reads but ignores */
uint8_t zo_crc8(const uint8_t *p, size_t n);   /* poly x^8+x^2+x+1, init 0   (zflac.zig:409-410) */
uint16_t zo_crc16(const uint8_t *p, size_t n); /* poly x^16+x^15+x^2+1, init 0 (zflac.zig:551) */
void zo_md5(const uint8_t *p, size_t n, uint8_t out[16]); /* RFC 1321 (std.crypto.hash.Md5, Zig 0.14.1 std) */

#ifdef __cplusplus
}
#endif
#endif
