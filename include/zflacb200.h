/*
 * zflacb200.h — C ABI of the B200-native FLAC decode path (libzflacb200.so).
 *
 * This is the drop-in boundary for Senryoku/zflac's one public entry point,
 *
 *     pub fn decode(allocator: std.mem.Allocator, reader: anytype) !DecodedFLAC
 *                                                   (reference src/zflac.zig:217)
 *
 * and its result type DecodedFLAC { channels, sample_rate, bits_per_sample,
 * samples: union{s8,s16,s32}, _samples_backing } (src/zflac.zig:12-28). A Zig
 * shim (zig/zflac.zig + zig/build.zig.fragment) keeps that signature and calls the
 * functions below; nothing in the signatures is a CUDA or torch type.
 *
 * Everything behind these calls runs on the GPU (sm_100a kernels). There is no
 * CPU decode fallback: if CUDA is unavailable the calls fail with
 * ZFLACB200_CUDA_ERROR.
 *
 * Ownership follows the reference ("Caller owns the returned memory",
 * src/zflac.zig:216; buffer from the caller's allocator, 32-byte aligned,
 * :331): decoding is two-phase so the CALLER allocates the PCM buffer:
 *
 *     zflacb200_open()   parse + upload + all kernels; PCM stays on the device;
 *                        returns the DecodedFLAC header fields and sizes
 *     (caller allocates info.pcm_bytes with its own allocator, align 32)
 *     zflacb200_read()   device -> caller buffer
 *     zflacb200_close()
 */
#ifndef ZFLACB200_H
#define ZFLACB200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ZFLACB200_ABI_VERSION 2

/* One value per Zig error name decode() can return; the shim maps
 * them 1:1 onto the Zig error set. */
typedef enum zflacb200_status {
    ZFLACB200_OK = 0,
    ZFLACB200_INVALID_SIGNATURE = 1,               /* error.InvalidSignature            src/zflac.zig:220 */
    ZFLACB200_INVALID_METADATA_HEADER = 2,         /* error.InvalidMetadataHeader       :248 */
    ZFLACB200_UNIMPLEMENTED = 3,                   /* error.Unimplemented               :263 */
    ZFLACB200_INVALID_CHECKSUM = 4,                /* error.InvalidChecksum             :280 */
    ZFLACB200_MISSING_STREAMINFO = 5,              /* error.MissingStreaminfo           :309 */
    ZFLACB200_INVALID_FRAME_HEADER = 6,            /* error.InvalidFrameHeader          :352,357,361,372,405 */
    ZFLACB200_INVALID_CODED_NUMBER = 7,            /* error.InvalidCodedNumber          :206 */
    ZFLACB200_INCONSISTENT_PARAMETERS = 8,         /* error.InconsistentParameters      :386,391 */
    ZFLACB200_INVALID_SUBFRAME_HEADER = 9,         /* error.InvalidSubframeHeader       :431,471,542 */
    ZFLACB200_INVALID_RESIDUAL_CODING_METHOD = 10, /* error.InvalidResidualCodingMethod :618 */
    ZFLACB200_END_OF_STREAM = 11,                  /* error.EndOfStream (reader)        e.g. :344 */
    ZFLACB200_OUT_OF_MEMORY = 12,                  /* error.OutOfMemory */
    /* input on which the reference itself is undefined (unreachable / @intCast
     * out of range / slice out of bounds: panic in Debug, UB in ReleaseFast) */
    ZFLACB200_UNDEFINED_IN_REFERENCE = 90,
    /* not Zig errors: */
    ZFLACB200_CUDA_ERROR = 100,      /* no device, kernel launch or copy failed; see zflacb200_last_error() */
    ZFLACB200_INVALID_ARGUMENT = 101
} zflacb200_status;

/* zflacb200_options.flags */
#define ZFLACB200_VERIFY_MD5 0x1u      /* default: reproduce the reference's MD5 gate (src/zflac.zig:267-280) */
#define ZFLACB200_SKIP_MD5 0x2u        /* additive: do not hash; never returns InvalidChecksum */
#define ZFLACB200_INPUT_ON_DEVICE 0x4u /* `data` is a device pointer (bench: inputs resident in HBM) */
#define ZFLACB200_NO_CRC_ACCEL 0x8u    /* diagnostics: ignore CRC-16 links, probe every frame end (slow path) */
#define ZFLACB200_NO_FAST_PATH 0x10u   /* diagnostics: skip the device-side hinted chain, walk the chain on the host */
#define ZFLACB200_SKIP_FRAME_CRC16 0x80u /* additive: do not compute the per-frame CRC-16 diagnostic (the reference reads the field and
                                          * ignores it, :551): info.n_crc16_mismatch stays 0, zflacb200_frame.crc16_ok says 2 = not checked */
#define ZFLACB200_NO_PIPELINE 0x40u    /* diagnostics: zflacb200_decode uploads, decodes and reads back one after the other */
#define ZFLACB200_SINGLE_PASS 0x20u    /* diagnostics: fused single-pass subframe decoder (K2 + K3) instead of parse + predict */
#define ZFLACB200_WORD_PIPELINE 0x400u /* diagnostics: pass A of 17..31-bit streams by the first bit reader (word pipeline, parse_core.cuh) instead of
                                        * the position reader (parse_narrow.cuh, WIDE): the A/B of the two, and what keeps both kernels tested */
#define ZFLACB200_MD5_ON_DEVICE 0x100u /* diagnostics: run a single stream's MD5 chain on the GPU (k_md5, 119 MB/s) instead of on a host
                                        * thread over the PCM as it lands in host memory (the default for streams >= 256 KiB of PCM;
                                        * batches always hash on the device, one chain per lane) */

#define ZFLACB200_RANGE_INTO_WHOLE 0x200u /* zflacb200_decode_range: `dst` is the buffer of the WHOLE stream's PCM (dst_capacity its size) and the
                                          * shard is written at its own place, byte range.first_sample * channels * container: the ranks of a
                                          * sharded decode fill ONE gathered buffer (e.g. shared pinned memory) with no copy afterwards */

typedef struct zflacb200_options {
    uint32_t struct_size; /* sizeof(zflacb200_options) */
    uint32_t flags;       /* 0 = defaults (VERIFY_MD5) */
    int32_t device;       /* CUDA device ordinal, -1 = current */
    void *stream;         /* cudaStream_t to launch on (NULL = the context's own stream) */
} zflacb200_options;

/* What DecodedFLAC carries (src/zflac.zig:18-24) plus sizes and diagnostics. */
typedef struct zflacb200_info {
    uint8_t channels;        /* DecodedFLAC.channels */
    uint8_t bits_per_sample; /* DecodedFLAC.bits_per_sample: from the first frame header (:381-384) */
    uint8_t container_bytes; /* union tag: 1 = .s8, 2 = .s16, 4 = .s32 (from STREAMINFO, :256-264) */
    uint8_t md5_checked;     /* 1 if the MD5 gate ran */
    uint32_t sample_rate;    /* DecodedFLAC.sample_rate (u24) */
    uint64_t n_samples;      /* samples.len: interleaved sample count */
    uint64_t pcm_bytes;      /* n_samples * container_bytes == _samples_backing.len */
    /* diagnostics (not part of the reference's result) */
    uint64_t n_frames;
    uint64_t audio_offset;       /* byte offset of the first frame */
    uint64_t audio_bytes;        /* bytes of the frames decoded */
    uint64_t n_candidates;       /* frame-index entries the sync scan emitted */
    uint64_t n_probed_frames;    /* frames whose end had to be found by decoding (no CRC-16 link) */
    uint64_t n_crc16_mismatch;   /* decoded frames whose CRC-16 does not match (reference ignores it, Q1) */
    uint64_t fast_path;          /* 1: device-side hinted chain accepted; 0: chain walked on the host */
    uint8_t md5_computed[16];
    uint8_t md5_streaminfo[16];
    /* device time of the last open(), ms, CUDA events on the launch stream */
    float ms_scan, ms_link, ms_decode, ms_epilogue, ms_md5, ms_total;
} zflacb200_info;

typedef struct zflacb200_stream zflacb200_stream; /* opaque handle of one decoded stream */

/* Library / device lifecycle. zflacb200_init is optional (open() does it lazily). */
int zflacb200_abi_version(void);
int zflacb200_init(int device);
int zflacb200_device_count(void);
void zflacb200_shutdown(void);
void zflacb200_trim(void); /* hands the library's idle device blocks (its block cache) back to the driver; handles stay valid */
const char *zflacb200_status_name(int status); /* the Zig error name, e.g. "InvalidFrameHeader" */
const char *zflacb200_last_error(void);        /* thread-local detail for CUDA_ERROR / INVALID_ARGUMENT */
uint64_t zflacb200_kernel_launches(void);      /* kernels this library has launched since load (bench: gpu_launches) */
/* diagnostics: launches of pass A (the frame-serial bit-chain kernel) by instantiation since load. variant 0:
 * k_parse_frames<StagedSrc> (int32 scratch, one wave), 1: k_parse_frames<QuadSrc> (int32 scratch, more frames than one
 * wave holds), 2: k_parse_frames_narrow<64> (streams of <= 16 bits), 3: k_parse_frames_narrow<32> (the same, large launches),
 * 4 / 5: k_parse_frames_narrow<64 / 32, WIDE> (int32 rows: streams of 17..31 bits; 0 / 1 then only run with ZFLACB200_WORD_PIPELINE). */
uint64_t zflacb200_pass_a_launches(int variant);

/* Replaces the body of decode() (src/zflac.zig:217-310): signature + metadata walk,
 * frame index, subframe decode, decorrelation, MD5 gate, MSB-align shift.
 * `data`/`len` is the whole stream the reader would yield. On success *out is a
 * handle holding the PCM on the device and *info is filled; on any error no
 * handle is returned (the reference returns no data on error, errdefer :265,:332). */
int zflacb200_open(const uint8_t *data, size_t len, const zflacb200_options *opt,
                   zflacb200_stream **out, zflacb200_info *info);

/* Copies the PCM (interleaved, container_bytes each, post MSB-align shift —
 * exactly DecodedFLAC._samples_backing) into `dst` (host memory, >= info.pcm_bytes). */
int zflacb200_read(zflacb200_stream *s, void *dst, size_t dst_bytes);

/* Device pointer of the PCM (valid until close); for callers that keep audio on the GPU. */
int zflacb200_device_pcm(zflacb200_stream *s, void **dev_ptr, size_t *bytes);

void zflacb200_close(zflacb200_stream *s);

/* One-call convenience used by hosts that own a plain malloc-style allocator:
 * open + read into `dst` if it is large enough (else returns pcm_bytes needed in info and
 * ZFLACB200_OUT_OF_MEMORY) + close. */
int zflacb200_decode(const uint8_t *data, size_t len, const zflacb200_options *opt,
                     void *dst, size_t dst_capacity, zflacb200_info *info);

/* Host only, no device: the bytes of PCM STREAMINFO promises (total samples x channels x container), so that a caller can
 * allocate `dst` for zflacb200_decode() up front the way the reference sizes its buffer (src/zflac.zig:329-334). *bytes = 0
 * when the stream states no total (:329 falls back to 4096 samples and grows) or the walk over the metadata fails — the
 * caller then takes the two-phase open / read path, which also reports the error. A stream may decode to MORE than it
 * states (:394-402); zflacb200_decode() then answers ZFLACB200_OUT_OF_MEMORY with info.pcm_bytes set. */
int zflacb200_pcm_bound(const uint8_t *data, size_t len, uint64_t *bytes);

/* File-parallel batch (BASELINE config 5): n independent streams decoded in ONE pipeline pass
 * on one device — one sync scan over the packed bytes, one K2 / K3 launch over the frames of all
 * streams, one MD5 chain per stream running side by side. Each stream is what one call of the
 * reference's decode() (src/zflac.zig:217) would get; statuses[i] / infos[i] are what that call
 * would return. `blob` holds the streams back to back (any gaps allowed): stream i is
 * blob[offsets[i] .. offsets[i] + lens[i]). `blob` is host memory (pinned memory makes the upload
 * run at full PCIe speed); `device_blob`, if not NULL, is a device copy of the same bytes and
 * skips the upload (bench: inputs resident in HBM). Streams the device-side chain cannot take
 * (no stated total, inconsistent numbering, damage, every error case) are decoded one by one
 * through the same path as zflacb200_open(). Returns ZFLACB200_OK if the batch ran. */
typedef struct zflacb200_batch zflacb200_batch;
int zflacb200_batch_open(const uint8_t *blob, const uint8_t *device_blob, const uint64_t *offsets,
                         const uint64_t *lens, size_t n, const zflacb200_options *opt,
                         zflacb200_batch **out, zflacb200_info *infos, int *statuses);
/* Bytes zflacb200_batch_read_all() writes: the PCM of all streams that decoded, 64-byte aligned each. */
int zflacb200_batch_pcm_bytes(zflacb200_batch *b, uint64_t *bytes);
/* Copies every decoded stream's PCM into `dst` (host) with a few large device-to-host copies;
 * pcm_offsets[i] receives the byte offset of stream i inside dst (UINT64_MAX if it did not decode). */
int zflacb200_batch_read_all(zflacb200_batch *b, void *dst, size_t dst_bytes, uint64_t *pcm_offsets);
/* One stream's PCM (host destination) / its device pointer (valid until close). */
int zflacb200_batch_read_stream(zflacb200_batch *b, size_t i, void *dst, size_t dst_bytes);
int zflacb200_batch_device_pcm(zflacb200_batch *b, size_t i, void **dev_ptr, size_t *bytes);
void zflacb200_batch_close(zflacb200_batch *b);

/* One-call variant, host blob in, host PCM out (the batch counterpart of zflacb200_decode): the
 * batch is cut into a few contiguous sub-batches; while sub-batch k is decoded, sub-batch k+1 is
 * still being uploaded and the PCM of k-1 already travels back — both PCIe directions and the SMs
 * busy at once. pcm_offsets[i] receives the byte offset of stream i inside dst (UINT64_MAX if it did
 * not decode; ZFLACB200_OUT_OF_MEMORY in statuses[i] if dst had no room left for it).
 * zflacb200_batch_pcm_bound() says how large dst must be, from the STREAMINFO totals (host only;
 * streams without a stated total count as 0: give them extra room or use the two-phase calls). */
int zflacb200_batch_pcm_bound(const uint8_t *blob, const uint64_t *offsets, const uint64_t *lens, size_t n, uint64_t *bytes);
int zflacb200_batch_decode(const uint8_t *blob, const uint64_t *offsets, const uint64_t *lens, size_t n,
                           const zflacb200_options *opt, void *dst, size_t dst_bytes, uint64_t *pcm_offsets,
                           zflacb200_info *infos, int *statuses);

/* Frame-range decode of ONE stream (several GPUs on one long stream): decodes the frames whose
 * header starts in [byte_lo, byte_hi) of `data` — the part of decode_frames()' loop
 * (src/zflac.zig:340-581) that falls into that byte range — and writes their PCM (interleaved,
 * post MSB-align shift) to `dst`. `data`/`len` is the whole stream in host memory; only the range
 * plus one maximum frame on either side is uploaded. Nothing links the shards on the device:
 * a frame's sample position comes from its coded number, the reference's chain rule (frame k+1
 * starts at the byte and at the sample where frame k ended, :546-548, :580) is checked INSIDE the
 * shard with the decoded frame ends, and zflacb200_stitch() checks it ACROSS shards from the
 * zflacb200_range each call returns, plus "first frame at the audio offset" and "covers exactly
 * the stated total" (:341). A caller splits [0, len) into contiguous ranges, runs one call per
 * GPU, gathers the n small structs and stitches; if a shard says accepted == 0 or the stitch
 * fails (no stated total, numbering the hints misjudge, damage, any error case), the stream is
 * decoded whole with zflacb200_decode(), which reports the reference's error. The MD5 gate is one
 * serial chain over the whole PCM and is not run here (info.md5_checked = 0).
 * Returns a metadata error exactly as zflacb200_open() would (every shard sees the same header),
 * ZFLACB200_OUT_OF_MEMORY if dst is too small (range.n_samples says what is needed). */
typedef struct zflacb200_range {
    uint32_t struct_size;        /* sizeof(zflacb200_range), set by the caller */
    uint32_t accepted;           /* 1: the shard decoded and its inner chain holds; 0: decode the stream whole */
    uint64_t first_frame_offset; /* byte offset of the first frame decoded */
    uint64_t end_offset;         /* one past the last byte of the last frame decoded */
    uint64_t first_sample;       /* per-channel index of the first sample decoded */
    uint64_t n_samples;          /* samples per channel decoded (dst receives n_samples * channels elements) */
    uint64_t n_frames;           /* 0: no frame starts in the range (the shard is skipped by the stitch) */
    uint64_t total_samples;      /* STREAMINFO total per channel, for the stitch */
} zflacb200_range;
int zflacb200_decode_range(const uint8_t *data, size_t len, const zflacb200_options *opt, uint64_t byte_lo, uint64_t byte_hi,
                           void *dst, size_t dst_capacity, zflacb200_range *range, zflacb200_info *info);
/* 1 if the n ranges (in stream order) are all accepted and link up into the one chain the reference walks, else 0. */
int zflacb200_stitch(const zflacb200_range *ranges, size_t n, uint64_t audio_offset);

/* ---- Raw frame streams / "sync to next frame" (additive; the reference's Readme.md:54 TODO) -------------
 * `data` needs no "fLaC" signature and no metadata (flac-test-files uncommon/10-11, which zflac.decode rejects with
 * InvalidSignature, tests/std_uncommon.zig:53-59), and `byte_offset` may point anywhere, into the middle of a frame too.
 * The sync scan runs from byte_offset; decoding starts at the first frame header (grammar of src/zflac.zig:343-374 and a
 * valid CRC-8) whose CRC-16 reaches another such header or the end of the data, else at the first header found. That
 * frame's fields stand in for STREAMINFO exactly as the reference latches them from its first frame (:376-388): channel
 * count, bits per sample (code 0 = "from STREAMINFO" -> MissingStreaminfo) and sample rate (code 0 -> 0). No total is known,
 * so the frame loop follows the reference's rules for a stream that states none (:341, :394-402): it runs until fewer than 4
 * bytes remain, and anything after the last frame that is not a frame is an InvalidFrameHeader. There is no MD5 to check
 * (info.md5_checked = 0). info.audio_offset says where decoding began; no frame found at all = EndOfStream.
 * Everything else (read, device_pcm, frames, close) is the same handle API as zflacb200_open(). */
int zflacb200_open_frames(const uint8_t *data, size_t len, const zflacb200_options *opt, uint64_t byte_offset,
                          zflacb200_stream **out, zflacb200_info *info);

/* ---- Output conventions (additive) ------------------------------------------------------------------
 * The reference leaves the output convention open (Readme.md:53 "auto convert to u8"; src/zflac.zig:282-292: "output
 * conventions differ quite a bit ... let the caller deal with it") and returns s8 / s16 / s32, MSB-aligned for 9-15 and
 * 17-31 bit streams. zflacb200_read() returns exactly that. zflacb200_read_as() re-expresses the same decoded PCM with one
 * streaming pass on the device before the read-back:
 *   F32         float, v * 2^-7 / 2^-15 / 2^-31 for the s8 / s16 / s32 container (int -> float rounds to nearest even)
 *   S24_PACKED  s32 container only: the top 24 bits of every sample (v >> 8), 3 bytes little-endian — for a 24-bit stream
 *               that is the sample as coded; a 32-bit stream loses its low 8 bits
 *   U8          s8 container only: v +% 128, the commented-out loop at :288-292
 * A format the stream's container does not define is ZFLACB200_INVALID_ARGUMENT. */
typedef enum zflacb200_format {
    ZFLACB200_FMT_NATIVE = 0, /* what zflacb200_read() writes */
    ZFLACB200_FMT_F32 = 1,
    ZFLACB200_FMT_S24_PACKED = 2,
    ZFLACB200_FMT_U8 = 3
} zflacb200_format;
int zflacb200_format_bytes(zflacb200_stream *s, int format, uint64_t *bytes); /* what zflacb200_read_as() writes */
int zflacb200_read_as(zflacb200_stream *s, int format, void *dst, size_t dst_bytes);

/* ---- Metadata surface (additive, host only, no device needed) ----------------------------------------
 * The reference walks the metadata blocks and skips everything but STREAMINFO (src/zflac.zig:243-247). These calls run the
 * same walk — same errors in the same order (InvalidSignature :220, EndOfStream, InvalidMetadataHeader :248,
 * MissingStreaminfo :309) — and report what it stepped over. `data` is a HOST pointer. Each call fills at most `capacity`
 * entries of `dst` (NULL allowed) and always sets the count, so a caller can size the array with a first call. */
typedef struct zflacb200_metadata_block {
    uint64_t offset;       /* of the block's data, just past its 4-byte header */
    uint32_t length;       /* the 24-bit length field as stored */
    uint32_t stored_bytes; /* what the walk advances by: `length`, except STREAMINFO = 34 whatever the field says (:224-240) */
    uint8_t type;          /* 0 STREAMINFO, 1 PADDING, 2 APPLICATION, 3 SEEKTABLE, 4 VORBIS_COMMENT, 5 CUESHEET, 6 PICTURE */
    uint8_t is_last;
    uint8_t pad[6];
} zflacb200_metadata_block;
int zflacb200_metadata(const uint8_t *data, size_t len, zflacb200_metadata_block *dst, size_t capacity, size_t *n_blocks,
                       uint64_t *audio_offset);

/* SEEKTABLE (RFC 9639 8.5): stream_offset counts from the first frame (add audio_offset). */
typedef struct zflacb200_seekpoint {
    uint64_t sample_number; /* first sample (per channel) of the target frame; 0xFFFFFFFFFFFFFFFF = placeholder */
    uint64_t stream_offset;
    uint32_t frame_samples;
    uint32_t is_placeholder;
} zflacb200_seekpoint;
int zflacb200_seektable(const uint8_t *data, size_t len, zflacb200_seekpoint *dst, size_t capacity, size_t *n_points);

/* VORBIS_COMMENT (RFC 9639 8.6): byte spans inside `data` of the vendor string and of each "NAME=value" field (UTF-8, not
 * NUL-terminated). A block whose internal lengths overrun it is InvalidMetadataHeader. */
typedef struct zflacb200_span {
    uint64_t offset, length;
} zflacb200_span;
int zflacb200_vorbis_comments(const uint8_t *data, size_t len, zflacb200_span *vendor, zflacb200_span *dst, size_t capacity,
                              size_t *n_comments);

/* PICTURE blocks (RFC 9639 8.8), one entry per block in stream order: the fixed fields and the byte spans of the MIME
 * type, the description (UTF-8) and the picture data inside `data`. A block whose internal lengths overrun it is
 * InvalidMetadataHeader (the reference skips the block unread, src/zflac.zig:243-247: decode() is not affected). */
typedef struct zflacb200_picture {
    uint32_t picture_type; /* 0..20, 3 = front cover */
    uint32_t width, height, depth, colors;
    uint32_t pad;
    zflacb200_span mime, description, image;
} zflacb200_picture;
int zflacb200_pictures(const uint8_t *data, size_t len, zflacb200_picture *dst, size_t capacity, size_t *n_pictures);

/* CUESHEET (RFC 9639 8.7): the sheet's fixed fields and its tracks; index points of track t are
 * indices[tracks[t].first_index .. + tracks[t].n_indices). Offsets are in samples (per channel). */
typedef struct zflacb200_cuesheet {
    char media_catalog_number[129]; /* NUL-terminated */
    uint8_t is_cd;
    uint8_t pad[6];
    uint64_t lead_in_samples;
    uint32_t n_tracks, n_indices;
} zflacb200_cuesheet;
typedef struct zflacb200_cue_track {
    uint64_t offset_samples;
    char isrc[13]; /* NUL-terminated */
    uint8_t number, is_audio, pre_emphasis;
    uint32_t first_index, n_indices;
} zflacb200_cue_track;
typedef struct zflacb200_cue_index {
    uint64_t offset_samples; /* relative to the track's offset */
    uint32_t number;
    uint32_t pad;
} zflacb200_cue_index;
/* *found = 0 when the stream has no CUESHEET block. tracks / indices may be NULL (counts only, in *sheet). */
int zflacb200_cuesheet_read(const uint8_t *data, size_t len, zflacb200_cuesheet *sheet, zflacb200_cue_track *tracks,
                            size_t track_capacity, zflacb200_cue_index *indices, size_t index_capacity, int *found);

/* Frame index of an opened stream (diagnostics / tests): one entry per decoded frame. */
typedef struct zflacb200_frame {
    uint64_t offset;        /* byte offset of the sync code */
    uint64_t end_offset;    /* just past the CRC-16 */
    uint64_t sample_offset; /* interleaved-sample offset */
    uint32_t block_size;
    uint8_t header_len;
    uint8_t channel_code;
    uint8_t crc16_ok;
    uint8_t probed;
} zflacb200_frame;
int zflacb200_frames(zflacb200_stream *s, zflacb200_frame *dst, size_t capacity, size_t *n_frames);

/* Frame-header scan for callers that may not allocate. NO COUNTERPART IN THE REFERENCE: added
 * for code under CUDA graph capture, tight serving loops and harnesses that forbid cudaMalloc. It is the
 * pipeline's first kernel (K1, k_sync_scan: every position whose bytes are a grammar-valid frame header with a matching
 * CRC-8, in stream order) with its scratch — one look-back word per 64 KB tile and a ticket — in a file-scope device array
 * of the library (128 KiB + 256 B, on every GPU that loads the library, until the process exits) instead of the block
 * cache. The call makes no cudaMalloc / cudaFree and no allocating library call: it enqueues two memsets of the scratch,
 * one of *d_count and the kernel on `stream`, and returns without synchronising.
 *   d_stream, len   the stream's bytes IN DEVICE MEMORY, 16-byte aligned; bytes at or past len are never read
 *   scan_begin      first byte position to test (the audio offset, or 0: a false match inside metadata is the caller's to drop)
 *   d_headers       DEVICE array of `capacity` records, 16-byte aligned; record i = the i-th header found
 *   d_count         DEVICE word: headers found; when it exceeds `capacity` only the first `capacity` were stored
 *   device, stream  CUDA device ordinal (-1 = current; the caller's current device is restored) and cudaStream_t
 * At most ZFLACB200_SCAN_NO_ALLOC_MAX_BYTES from scan_begin to len per call: beyond that ZFLACB200_INVALID_ARGUMENT (split
 * the range, or use zflacb200_open). EVERY call on a device shares the one array, which keeps the last call's contents
 * (the call clears what it uses): calls that overlap on the GPU — two streams, two threads — race and corrupt each
 * other's ranks; order them on one stream or with events. Candidates are hints, as everywhere in this library: the
 * frames of a stream are those zflacb200_frames() reports after the chain rules were checked. */
#define ZFLACB200_SCAN_NO_ALLOC_MAX_BYTES (1ull << 30)
typedef struct zflacb200_frame_header {
    uint64_t offset;       /* byte offset of the sync code */
    uint64_t number;       /* coded frame number (fixed block size) or first sample number (variable) */
    uint32_t block_size;   /* 1..65535 */
    uint32_t sample_rate;  /* as the frame header states it, 0 = "see STREAMINFO" */
    uint8_t header_len;    /* bytes, CRC-8 included */
    uint8_t channel_code;  /* 4-bit channel assignment */
    uint8_t depth_code;    /* 3-bit sample size code */
    uint8_t rate_code;     /* 4-bit sample rate code */
    uint8_t variable;      /* blocking strategy bit */
    uint8_t pad[3];
} zflacb200_frame_header;
int zflacb200_scan_no_alloc(const void *d_stream, uint64_t len, uint64_t scan_begin, zflacb200_frame_header *d_headers,
                            uint32_t capacity, uint32_t *d_count, int device, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* ZFLACB200_H */
