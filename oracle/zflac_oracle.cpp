/*
 * zflac_oracle.cpp — TEST INFRASTRUCTURE, NOT PRODUCT CODE (see zflac_oracle.h).
 *
 * CPU restatement of Senryoku/zflac (Zig 0.14.1) in C++17. It keeps the
 * reference's algorithmic structure on purpose, so that timing it is a fair
 * stand-in for the reference on this box: a byte-at-a-time MSB-first bit reader
 * with an end-of-stream check per byte, strictly sequential frames, per-container
 * intermediate widths (i8->i16, i16->i32, i32->i64), a separate MD5 pass and a
 * separate MSB-align shift pass, CRC-8/CRC-16 read but never checked.
 *
 * "ref:" comments give the lines of /root/reference/src/zflac.zig (or
 * bit_reader.zig where stated) that the code below restates.
 *
 * Where the reference has undefined behaviour in ReleaseFast (a panic in Debug)
 * the oracle either wraps like two's-complement hardware and raises
 * overflow_flag, or — where no sensible continuation exists — returns
 * ZO_UNDEFINED_IN_REFERENCE.
 */
#include "zflac_oracle.h"

#include <algorithm>
#include <chrono>
#include <cstdlib>
#include <cstring>
#include <type_traits>
#include <vector>

namespace {

#define ZO_TRY(expr)                 \
    do {                             \
        int _e = (expr);             \
        if (_e != ZO_OK) return _e;  \
    } while (0)

// ---------------------------------------------------------------------------
// Byte reader: the `reader: anytype` of decode(); models std.io reader calls
// readByte / readInt / readBytesNoEof / skipBytes with error.EndOfStream.
// ---------------------------------------------------------------------------
struct ByteReader {
    const uint8_t *p;
    size_t len;
    size_t pos = 0;

    inline int read_byte(uint8_t &out) {
        if (pos >= len) return ZO_END_OF_STREAM;
        out = p[pos++];
        return ZO_OK;
    }
    // readInt(T, .big): reads sizeof(T) bytes with readNoEof
    int read_be(unsigned nbytes, uint32_t &out) {
        if (len - pos < nbytes) {
            pos = len;
            return ZO_END_OF_STREAM;
        }
        uint32_t v = 0;
        for (unsigned i = 0; i < nbytes; ++i) v = (v << 8) | p[pos++];
        out = v;
        return ZO_OK;
    }
    int skip(size_t n) {
        if (len - pos < n) {
            pos = len;
            return ZO_END_OF_STREAM;
        }
        pos += n;
        return ZO_OK;
    }
};

// ---------------------------------------------------------------------------
// ref: bit_reader.zig:7-126. MSB-first, at most 8 buffered bits, one
// read_byte() per byte consumed.
// ---------------------------------------------------------------------------
struct BitReader {
    ByteReader &r;
    uint8_t bits = 0;   // ref bit_reader.zig:10
    uint8_t count = 0;  // ref bit_reader.zig:11

    explicit BitReader(ByteReader &reader) : r(reader) {}

    static inline uint8_t low_mask(unsigned n) { return (uint8_t)((1u << n) - 1u); }

    // ref: bit_reader.zig:25-70 readBitsNoEof (U = 64 bits is a superset of every T)
    inline int read_bits(unsigned num, uint64_t &out) {
        if (num <= count) {  // ref :31 / removeBits :72-81
            unsigned keep = count - num;
            out = (uint64_t)(bits >> keep);
            bits &= low_mask(keep);
            count = (uint8_t)keep;
            return ZO_OK;
        }
        unsigned bits_left = num - count;  // ref :33
        uint64_t acc = bits;               // ref :35 flush()
        bits = 0;
        count = 0;
        unsigned full = bits_left / 8;  // ref :46-52
        for (unsigned i = 0; i < full; ++i) {
            uint8_t b;
            ZO_TRY(r.read_byte(b));
            acc = (acc << 8) | b;
        }
        bits_left %= 8;  // ref :54
        if (bits_left == 0) {
            out = acc;
            return ZO_OK;
        }
        uint8_t fin;  // ref :61-69
        ZO_TRY(r.read_byte(fin));
        unsigned keep = 8 - bits_left;
        acc = (acc << bits_left) | (uint64_t)(fin >> keep);
        bits = fin & low_mask(keep);
        count = (uint8_t)keep;
        out = acc;
        return ZO_OK;
    }

    // ref: bit_reader.zig:90-93
    inline void align_to_byte() {
        bits = 0;
        count = 0;
    }

    // ref: bit_reader.zig:95-120 readUnary / readUnaryFromEmptyBuffer
    inline int read_unary(uint32_t &out) {
        if (bits == 0) {
            uint32_t n = count;
            uint8_t b;
            ZO_TRY(r.read_byte(b));
            while (b == 0) {
                n += 8;
                ZO_TRY(r.read_byte(b));
            }
            unsigned lz = (unsigned)__builtin_clz((unsigned)b) - 24u;
            count = (uint8_t)(7 - lz);
            bits = b & low_mask(count);
            out = n + lz;
            return ZO_OK;
        }
        unsigned lz = ((unsigned)__builtin_clz((unsigned)bits) - 24u) - (8u - count);
        count = (uint8_t)(count - 1 - lz);
        bits &= low_mask(count);
        out = lz;
        return ZO_OK;
    }
};

// ---------------------------------------------------------------------------
// Type plumbing: SampleType -> InterType (ref :314-319) and unsigned twins.
// ---------------------------------------------------------------------------
template <typename S> struct Inter;
template <> struct Inter<int8_t> { using type = int16_t; };
template <> struct Inter<int16_t> { using type = int32_t; };
template <> struct Inter<int32_t> { using type = int64_t; };

template <typename T> constexpr unsigned bits_of() { return sizeof(T) * 8; }

// wrap an int64 result into T the way two's-complement hardware does
// ZO_TIMED (the build bench.py times as the CPU arm): the checks that only feed `overflow_flag` — where a Zig Debug build
// would panic and ReleaseFast, which is what gets benchmarked, computes on — are compiled out, and the LPC loop is
// instantiated per order like the reference's comptime unroll (ref :526-533, @Vector(32) for order 32 :604-612). The
// instrumented build stays the one the tests use; both must decode every fixture to the same samples.
template <typename T> inline T wrap_to(int64_t v, uint8_t &ovf) {
    T t = (T)(typename std::make_unsigned<T>::type)(uint64_t)v;
#ifndef ZO_TIMED
    if ((int64_t)t != v) ovf = 1;
#else
    (void)ovf;
#endif
    return t;
}

#ifdef ZO_TIMED
// ref :522-534 with `order` a compile-time constant: the inner loop unrolls (and vectorises for 32, ref :604-612)
template <typename I, typename UI, unsigned ORDER>
static void lpc_restore(I *w, const I *coef, unsigned shift, uint32_t block_size) {
    for (uint32_t i = ORDER; i < block_size; ++i) {
        UI pred = 0;
#pragma GCC unroll 32
        for (unsigned o = 0; o < ORDER; ++o) pred = (UI)(pred + (UI)((UI)w[i - ORDER + o] * (UI)coef[o]));
        w[i] = (I)((UI)w[i] + (UI)(I)((I)pred >> shift));
    }
}
template <typename I, typename UI>
static void lpc_restore_any(I *w, const I *coef, unsigned order, unsigned shift, uint32_t block_size) {
    switch (order) {
#define ZO_ORDER(n) case n: lpc_restore<I, UI, n>(w, coef, shift, block_size); break;
        ZO_ORDER(1) ZO_ORDER(2) ZO_ORDER(3) ZO_ORDER(4) ZO_ORDER(5) ZO_ORDER(6) ZO_ORDER(7) ZO_ORDER(8)
        ZO_ORDER(9) ZO_ORDER(10) ZO_ORDER(11) ZO_ORDER(12) ZO_ORDER(13) ZO_ORDER(14) ZO_ORDER(15) ZO_ORDER(16)
        ZO_ORDER(17) ZO_ORDER(18) ZO_ORDER(19) ZO_ORDER(20) ZO_ORDER(21) ZO_ORDER(22) ZO_ORDER(23) ZO_ORDER(24)
        ZO_ORDER(25) ZO_ORDER(26) ZO_ORDER(27) ZO_ORDER(28) ZO_ORDER(29) ZO_ORDER(30) ZO_ORDER(31) ZO_ORDER(32)
#undef ZO_ORDER
        default: break;
    }
}
#endif

// ref: :188-196 read_signed_integer(T, bit_reader, bit_depth)
template <typename T> inline int read_signed_integer(BitReader &br, unsigned bit_depth, T &out) {
    if (bit_depth == 0 || bit_depth > bits_of<T>()) return ZO_UNDEFINED_IN_REFERENCE;  // assert :189
    uint64_t raw;
    ZO_TRY(br.read_bits(bit_depth, raw));
    unsigned shift = 64 - bit_depth;
    out = (T)((int64_t)(raw << shift) >> shift);
    return ZO_OK;
}

// ref: :198-201 read_unencoded_sample(SampleType, bit_reader, wasted_bits, bits_per_sample)
// reads through InterType and narrows with @intCast.
template <typename S>
inline int read_unencoded_sample(BitReader &br, unsigned wasted_bits, unsigned bits_per_sample, S &out, uint8_t &ovf) {
    using I = typename Inter<S>::type;
    if (wasted_bits > bits_per_sample) return ZO_UNDEFINED_IN_REFERENCE;  // u6 underflow
    I v;
    ZO_TRY(read_signed_integer<I>(br, bits_per_sample - wasted_bits, v));
    out = wrap_to<S>((int64_t)v, ovf);
    return ZO_OK;
}

// ref: :203-214 read_coded_number
int read_coded_number(ByteReader &r, uint64_t &out) {
    uint8_t first;
    ZO_TRY(r.read_byte(first));
    unsigned inv = (uint8_t)(first ^ 0xFF);
    unsigned byte_count = inv == 0 ? 8u : ((unsigned)__builtin_clz(inv) - 24u);
    if (first == 0xFF || byte_count == 1) return ZO_INVALID_CODED_NUMBER;
    if (byte_count == 0) {
        out = first;
        return ZO_OK;
    }
    uint64_t v = first & (0x7Fu >> byte_count);
    for (unsigned i = 0; i + 1 < byte_count; ++i) {
        uint8_t b;
        ZO_TRY(r.read_byte(b));
        v = (v << 6) | (b & 0x3F);
    }
    out = v;
    return ZO_OK;
}

struct Streaminfo {  // ref :45-55
    uint32_t min_block_size, max_block_size, min_frame_size, max_frame_size;
    uint32_t sample_rate;       // u20
    uint32_t channel_count;     // u3: channels - 1
    uint32_t sample_bit_depth;  // u5: bits - 1
    uint64_t number_of_samples; // u36
    uint8_t md5[16];
};

// ref: :635-666 decode_residual_partition
template <typename R>
int decode_residual_partition(unsigned coding_method, R *residuals, uint32_t count, BitReader &br, uint8_t &ovf) {
    using U = typename std::make_unsigned<R>::type;
    const unsigned param_bits = coding_method == 0 ? 4 : 5;  // ref :637-640
    uint64_t rice_parameter;
    ZO_TRY(br.read_bits(param_bits, rice_parameter));
    if (rice_parameter == (1u << param_bits) - 1u) {  // ref :646-654 escape
        uint64_t bit_depth;
        ZO_TRY(br.read_bits(5, bit_depth));
        if (bit_depth == 0) {
            std::memset(residuals, 0, sizeof(R) * count);
        } else {
            for (uint32_t i = 0; i < count; ++i) ZO_TRY(read_signed_integer<R>(br, (unsigned)bit_depth, residuals[i]));
        }
        return ZO_OK;
    }
    const unsigned k = (unsigned)rice_parameter;
    if (k >= bits_of<U>()) return ZO_UNDEFINED_IN_REFERENCE;  // ref :656 unreachable
    for (uint32_t i = 0; i < count; ++i) {                     // ref :657-663
        uint32_t q;
        ZO_TRY(br.read_unary(q));
        uint64_t rem;
        ZO_TRY(br.read_bits(k, rem));
#ifndef ZO_TIMED
        if ((uint64_t)q > (uint64_t)(U)~(U)0) ovf = 1;  // @intCast(quotient)
#endif
        U quotient = (U)q;
        U shifted = (U)(quotient << k);  // `<<` discards
        U zz = (U)(shifted + (U)rem);
#ifndef ZO_TIMED
        if ((U)(zz - shifted) != (U)rem || zz < shifted) ovf = 1;
#endif
        residuals[i] = (R)((U)(zz >> 1) ^ (U)(0 - (U)(zz & 1)));
    }
    return ZO_OK;
}

// ref: :614-633 decode_residuals
template <typename R>
int decode_residuals(R *residuals, size_t capacity, uint32_t block_size, unsigned order, BitReader &br, uint8_t &ovf) {
    uint64_t coding_method, partition_order;
    ZO_TRY(br.read_bits(2, coding_method));
    if (coding_method >= 2) return ZO_INVALID_RESIDUAL_CODING_METHOD;
    ZO_TRY(br.read_bits(4, partition_order));
    uint32_t start = 0;
    const uint32_t n_part = 1u << partition_order;
    for (uint32_t part = 0; part < n_part; ++part) {
        uint32_t count = block_size >> partition_order;  // u16
        if (part == 0) {
            if (count < order) return ZO_UNDEFINED_IN_REFERENCE;  // u16 underflow (Q16)
            count -= order;
        }
        if ((size_t)start + count > capacity) return ZO_UNDEFINED_IN_REFERENCE;  // slice bounds
        ZO_TRY(decode_residual_partition<R>((unsigned)coding_method, residuals + start, count, br, ovf));
        start += count;
    }
    return ZO_OK;
}

void *aligned_alloc32(size_t bytes) {
    void *p = nullptr;
    if (posix_memalign(&p, 32, bytes ? bytes : 32) != 0) return nullptr;
    return p;
}

template <typename S> struct Frames {
    S *samples = nullptr;  // _samples_backing viewed as SampleType
    size_t len = 0;        // samples.len
    uint8_t channels = 0;
    uint32_t sample_rate = 0;
    uint8_t bits_per_sample = 0;
};

const uint32_t kRateTable[16] = {0, 88200, 176400, 192000, 8000, 16000, 22050, 24000,
                                 32000, 44100, 48000, 96000, 0, 0, 0, 0};  // ref :75-90
const uint8_t kChannelCount[16] = {1, 2, 3, 4, 5, 6, 7, 8, 2, 2, 2, 0, 0, 0, 0, 0};  // ref :107-122
const uint8_t kBps[8] = {0, 8, 12, 0, 16, 20, 24, 32};  // ref :135-145

// ref: :312-602 decode_frames(SampleType, allocator, stream_info, reader)
template <typename S>
int decode_frames(const Streaminfo &si, ByteReader &reader, Frames<S> &out, uint8_t &ovf, std::vector<zo_frame> *trace) {
    using I = typename Inter<S>::type;
    using UI = typename std::make_unsigned<I>::type;

    bool first_frame = true;
    uint32_t sample_rate = 0;
    unsigned channel_count = 0;
    unsigned bit_depth_code = 0;
    unsigned bits_per_sample = 0;

    bool valid_total = si.number_of_samples > 0;  // ref :327
    const size_t expected_channel_count = (size_t)si.channel_count + 1;
    const size_t total_sample_count = expected_channel_count * (valid_total ? (size_t)si.number_of_samples : 4096);  // ref :330
    S *samples = (S *)aligned_alloc32(sizeof(S) * total_sample_count);
    if (!samples) return ZO_OUT_OF_MEMORY;
    size_t samples_len = total_sample_count;

    std::vector<I> work(si.max_block_size > 0 ? si.max_block_size : 4096);  // ref :336

    auto fail = [&](int code) {
        free(samples);
        return code;
    };

    size_t frame_sample_offset = 0;
    for (;;) {
        if (valid_total && frame_sample_offset >= total_sample_count) break;  // ref :341

        const size_t frame_start = reader.pos;
        uint32_t hdr;
        int e = reader.read_be(4, hdr);  // ref :343-350
        if (e != ZO_OK) {
            if (valid_total) return fail(e);
            break;  // EndOfStream on a frame boundary with unknown total
        }
        if ((hdr >> 17) != (0xFFF8u >> 1)) return fail(ZO_INVALID_FRAME_HEADER);  // ref :351
        const unsigned bs_code = (hdr >> 12) & 0xF, sr_code = (hdr >> 8) & 0xF;
        const unsigned ch_code = (hdr >> 4) & 0xF, bd_code = (hdr >> 1) & 0x7;

        uint64_t coded_number;
        if ((e = read_coded_number(reader, coded_number)) != ZO_OK) return fail(e);  // ref :354
        (void)coded_number;

        uint32_t block_size;  // ref :356-365
        if (bs_code == 0) return fail(ZO_INVALID_FRAME_HEADER);
        else if (bs_code == 6) {
            uint32_t v;
            if ((e = reader.read_be(1, v)) != ZO_OK) return fail(e);
            block_size = v + 1;
        } else if (bs_code == 7) {
            uint32_t v;
            if ((e = reader.read_be(2, v)) != ZO_OK) return fail(e);
            if (v == 0xFFFF) return fail(ZO_INVALID_FRAME_HEADER);
            block_size = v + 1;
        } else if (bs_code == 1) block_size = 192;
        else if (bs_code <= 5) block_size = 144u << bs_code;
        else block_size = 1u << bs_code;

        uint32_t frame_sample_rate;  // ref :367-374
        if (sr_code == 0) frame_sample_rate = si.sample_rate;
        else if (sr_code == 12) {
            if ((e = reader.read_be(1, frame_sample_rate)) != ZO_OK) return fail(e);  // NOTE: kHz value returned as is (Q7)
        } else if (sr_code == 13) {
            if ((e = reader.read_be(2, frame_sample_rate)) != ZO_OK) return fail(e);
        } else if (sr_code == 14) {
            uint32_t v;
            if ((e = reader.read_be(2, v)) != ZO_OK) return fail(e);
            frame_sample_rate = 10 * v;
        } else if (sr_code == 15) return fail(ZO_INVALID_FRAME_HEADER);
        else frame_sample_rate = kRateTable[sr_code];

        if (first_frame) {  // ref :376-388
            sample_rate = frame_sample_rate;
            channel_count = kChannelCount[ch_code];
            bit_depth_code = bd_code;
            if (bd_code == 0) bits_per_sample = si.sample_bit_depth + 1;
            else if (bd_code == 3) return fail(ZO_UNDEFINED_IN_REFERENCE);  // `else => unreachable` :143
            else bits_per_sample = kBps[bd_code];
            if (channel_count != expected_channel_count) return fail(ZO_INCONSISTENT_PARAMETERS);
            first_frame = false;
        } else {  // ref :389-392
            if (sample_rate != frame_sample_rate || channel_count != kChannelCount[ch_code] || bit_depth_code != bd_code)
                return fail(ZO_INCONSISTENT_PARAMETERS);
        }

        const size_t expected_samples = frame_sample_offset + (size_t)block_size * channel_count;  // ref :394-402
        if (samples_len < expected_samples) {
            const size_t new_size = std::max(2 * samples_len, expected_samples);
            S *grown = (S *)aligned_alloc32(new_size * sizeof(S));
            if (!grown) return fail(ZO_OUT_OF_MEMORY);
            std::memcpy(grown, samples, frame_sample_offset * sizeof(S));
            free(samples);
            samples = grown;
            samples_len = new_size;
            valid_total = false;
        }

        // ref :405
        if (block_size == 1 && (valid_total && frame_sample_offset + (size_t)channel_count * block_size < total_sample_count))
            return fail(ZO_INVALID_FRAME_HEADER);

        uint32_t frame_header_crc;  // ref :407 (never checked, Q1)
        if ((e = reader.read_be(1, frame_header_crc)) != ZO_OK) return fail(e);
        const size_t header_len = reader.pos - frame_start;

        BitReader br(reader);  // ref :423

        for (unsigned channel = 0; channel < channel_count; ++channel) {  // ref :425
            uint64_t zero, type, wasted_flag;
            if ((e = br.read_bits(1, zero)) != ZO_OK) return fail(e);
            if ((e = br.read_bits(6, type)) != ZO_OK) return fail(e);
            if ((e = br.read_bits(1, wasted_flag)) != ZO_OK) return fail(e);
            if (zero != 0) return fail(ZO_INVALID_SUBFRAME_HEADER);  // ref :431

            unsigned wasted_bits = 0;  // ref :433
            if (wasted_flag) {
                uint32_t u;
                if ((e = br.read_unary(u)) != ZO_OK) return fail(e);
                if (u + 1 > 63) return fail(ZO_UNDEFINED_IN_REFERENCE);  // @intCast to u6
                wasted_bits = u + 1;
            }

            unsigned depth = bits_per_sample;  // ref :436-441
            if ((ch_code == 8 && channel == 1) || (ch_code == 9 && channel == 0) || (ch_code == 10 && channel == 1)) depth += 1;

            S *sub = samples + frame_sample_offset + channel;  // ref :443, stride = channel_count
            const size_t stride = channel_count;

            if (type == 0) {  // ref :445-454 CONSTANT — reads bits_per_sample, not depth (Q8)
                S v;
                if ((e = read_unencoded_sample<S>(br, wasted_bits, bits_per_sample, v, ovf)) != ZO_OK) return fail(e);
                v = (S)((typename std::make_unsigned<S>::type)v << wasted_bits);
                for (uint32_t i = 0; i < block_size; ++i) sub[stride * i] = v;
            } else if (type == 1) {  // ref :455-465 VERBATIM
                for (uint32_t i = 0; i < block_size; ++i) {
                    S v;
                    if ((e = read_unencoded_sample<S>(br, wasted_bits, depth, v, ovf)) != ZO_OK) return fail(e);
                    sub[stride * i] = (S)((typename std::make_unsigned<S>::type)v << wasted_bits);
                }
            } else if (type >= 8 && type <= 12) {  // ref :466-498 FIXED
                if (work.size() < block_size) work.resize(block_size);
                const unsigned order = (unsigned)(type & 7);
                if (order > block_size) return fail(ZO_UNDEFINED_IN_REFERENCE);
                for (unsigned i = 0; i < order; ++i) {
                    S v;
                    if ((e = read_unencoded_sample<S>(br, wasted_bits, depth, v, ovf)) != ZO_OK) return fail(e);
                    work[i] = v;
                }
                if ((e = decode_residuals<I>(work.data() + order, work.size() - order, block_size, order, br, ovf)) != ZO_OK) return fail(e);
                I *w = work.data();
                for (uint32_t i = order; i < block_size; ++i) {  // ref :481-490, InterType arithmetic (wraps)
                    int64_t p = 0;
                    switch (order) {
                        case 0: p = 0; break;
                        case 1: p = (int64_t)w[i - 1]; break;
                        case 2: p = 2 * (int64_t)w[i - 1] - (int64_t)w[i - 2]; break;
                        case 3: p = 3 * (int64_t)w[i - 1] - 3 * (int64_t)w[i - 2] + (int64_t)w[i - 3]; break;
                        case 4: p = 4 * (int64_t)w[i - 1] - 6 * (int64_t)w[i - 2] + 4 * (int64_t)w[i - 3] - (int64_t)w[i - 4]; break;
                    }
                    w[i] = wrap_to<I>((int64_t)w[i] + p, ovf);
                }
                for (uint32_t i = 0; i < block_size; ++i) {  // ref :493-497 interleave + wasted shift
                    S v = wrap_to<S>((int64_t)w[i], ovf);
                    sub[stride * i] = (S)((typename std::make_unsigned<S>::type)v << wasted_bits);
                }
            } else if (type >= 32) {  // ref :499-541 LPC
                if (work.size() < block_size) work.resize(block_size);
                const unsigned order = (unsigned)type - 31;
                if (order > block_size) return fail(ZO_UNDEFINED_IN_REFERENCE);
                for (unsigned i = 0; i < order; ++i) {
                    S v;
                    if ((e = read_unencoded_sample<S>(br, wasted_bits, depth, v, ovf)) != ZO_OK) return fail(e);
                    work[i] = v;
                }
                uint64_t prec_m1, shift;
                if ((e = br.read_bits(4, prec_m1)) != ZO_OK) return fail(e);  // ref :508
                if ((e = br.read_bits(5, shift)) != ZO_OK) return fail(e);    // ref :510 unsigned (Q17)
                const unsigned precision = (unsigned)prec_m1 + 1;
                I coef[32] = {0};  // ref :512-514, stored reversed
                for (unsigned i = 0; i < order; ++i) {
                    I c;  // read_unencoded_sample(InterType, .., 0, precision): 1..16 bits always fit InterType
                    if ((e = read_signed_integer<I>(br, precision, c)) != ZO_OK) return fail(e);
                    coef[order - 1 - i] = c;
                }
                if ((e = decode_residuals<I>(work.data() + order, work.size() - order, block_size, order, br, ovf)) != ZO_OK) return fail(e);
                if (shift >= bits_of<I>()) return fail(ZO_UNDEFINED_IN_REFERENCE);  // @intCast of the shift amount :531
                I *w = work.data();
#ifdef ZO_TIMED
                lpc_restore_any<I, UI>(w, coef, order, (unsigned)shift, block_size);
                for (uint32_t i = block_size; i < block_size; ++i) {  // (the instrumented loop below, compiled out)
#else
                for (uint32_t i = order; i < block_size; ++i) {  // ref :522-534 and :604-612
#endif
                    UI pred = 0;  // InterType accumulate, wrapping like the hardware
                    int64_t wide = 0;
                    for (unsigned o = 0; o < order; ++o) {
                        pred = (UI)(pred + (UI)((uint64_t)(UI)w[i - order + o] * (uint64_t)(UI)coef[o]));
                        wide += (int64_t)w[i - order + o] * (int64_t)coef[o];
                    }
                    if (sizeof(I) < 8 && (int64_t)(I)pred != wide) ovf = 1;
                    I shifted = (I)((I)pred >> shift);
                    w[i] = wrap_to<I>((int64_t)w[i] + (int64_t)shifted, ovf);
                }
                for (uint32_t i = 0; i < block_size; ++i) {  // ref :536-540
                    S v = wrap_to<S>((int64_t)w[i], ovf);
                    sub[stride * i] = (S)((typename std::make_unsigned<S>::type)v << wasted_bits);
                }
            } else {  // ref :542 reserved types
                return fail(ZO_INVALID_SUBFRAME_HEADER);
            }
        }

        br.align_to_byte();  // ref :546
        uint32_t frame_crc;   // ref :548 (never checked, Q1)
        if ((e = reader.read_be(2, frame_crc)) != ZO_OK) return fail(e);

        // ref :553-578 stereo decorrelation, in place on the interleaved buffer
        S *fs = samples + frame_sample_offset;
        if (ch_code == 8) {
            for (uint32_t i = 0; i < block_size; ++i) {
                S *p = fs + (size_t)channel_count * i;
                p[1] = wrap_to<S>((int64_t)p[0] - (int64_t)p[1], ovf);
            }
        } else if (ch_code == 9) {
            for (uint32_t i = 0; i < block_size; ++i) {
                S *p = fs + (size_t)channel_count * i;
                p[0] = wrap_to<S>((int64_t)p[0] + (int64_t)p[1], ovf);
            }
        } else if (ch_code == 10) {
            for (uint32_t i = 0; i < block_size; ++i) {
                S *p = fs + (size_t)channel_count * i;
                I mid = (I)((UI)(I)p[0] << 1);
                const S side = p[1];
                mid = (I)(mid + (I)(side & 1));
                p[0] = wrap_to<S>(((int64_t)mid + (int64_t)side) >> 1, ovf);
                p[1] = wrap_to<S>(((int64_t)mid - (int64_t)side) >> 1, ovf);
            }
        }

        if (trace) {
            zo_frame f;
            f.offset = frame_start;
            f.end_offset = reader.pos;
            f.sample_offset = frame_sample_offset;
            f.block_size = block_size;
            f.header_len = (uint8_t)header_len;
            f.channel_code = (uint8_t)ch_code;
            f.bit_depth_code = (uint8_t)bd_code;
            f.sample_rate_code = (uint8_t)sr_code;
            trace->push_back(f);
        }
        frame_sample_offset += (size_t)channel_count * block_size;  // ref :580
    }

    // ref :583-588 trim
    out.samples = samples;
    out.len = frame_sample_offset;
    out.channels = (uint8_t)channel_count;
    out.sample_rate = sample_rate & 0xFFFFFF;
    out.bits_per_sample = (uint8_t)bits_per_sample;
    return ZO_OK;
}

// --------------------------------- MD5 (RFC 1321) ---------------------------
struct Md5 {
    uint32_t a = 0x67452301, b = 0xefcdab89, c = 0x98badcfe, d = 0x10325476;
    uint64_t total = 0;
    uint8_t buf[64];
    unsigned fill = 0;

    static inline uint32_t rotl(uint32_t x, unsigned s) { return (x << s) | (x >> (32 - s)); }

    void block(const uint8_t *p) {
        static const uint32_t K[64] = {
            0xd76aa478, 0xe8c7b756, 0x242070db, 0xc1bdceee, 0xf57c0faf, 0x4787c62a, 0xa8304613, 0xfd469501,
            0x698098d8, 0x8b44f7af, 0xffff5bb1, 0x895cd7be, 0x6b901122, 0xfd987193, 0xa679438e, 0x49b40821,
            0xf61e2562, 0xc040b340, 0x265e5a51, 0xe9b6c7aa, 0xd62f105d, 0x02441453, 0xd8a1e681, 0xe7d3fbc8,
            0x21e1cde6, 0xc33707d6, 0xf4d50d87, 0x455a14ed, 0xa9e3e905, 0xfcefa3f8, 0x676f02d9, 0x8d2a4c8a,
            0xfffa3942, 0x8771f681, 0x6d9d6122, 0xfde5380c, 0xa4beea44, 0x4bdecfa9, 0xf6bb4b60, 0xbebfbc70,
            0x289b7ec6, 0xeaa127fa, 0xd4ef3085, 0x04881d05, 0xd9d4d039, 0xe6db99e5, 0x1fa27cf8, 0xc4ac5665,
            0xf4292244, 0x432aff97, 0xab9423a7, 0xfc93a039, 0x655b59c3, 0x8f0ccc92, 0xffeff47d, 0x85845dd1,
            0x6fa87e4f, 0xfe2ce6e0, 0xa3014314, 0x4e0811a1, 0xf7537e82, 0xbd3af235, 0x2ad7d2bb, 0xeb86d391};
        static const uint8_t S[64] = {7, 12, 17, 22, 7, 12, 17, 22, 7, 12, 17, 22, 7, 12, 17, 22,
                                      5, 9,  14, 20, 5, 9,  14, 20, 5, 9,  14, 20, 5, 9,  14, 20,
                                      4, 11, 16, 23, 4, 11, 16, 23, 4, 11, 16, 23, 4, 11, 16, 23,
                                      6, 10, 15, 21, 6, 10, 15, 21, 6, 10, 15, 21, 6, 10, 15, 21};
        uint32_t m[16];
        for (int i = 0; i < 16; ++i)
            m[i] = (uint32_t)p[4 * i] | ((uint32_t)p[4 * i + 1] << 8) | ((uint32_t)p[4 * i + 2] << 16) | ((uint32_t)p[4 * i + 3] << 24);
        uint32_t A = a, B = b, C = c, D = d;
        for (int i = 0; i < 64; ++i) {
            uint32_t f;
            int g;
            if (i < 16) { f = (B & C) | (~B & D); g = i; }
            else if (i < 32) { f = (D & B) | (~D & C); g = (5 * i + 1) & 15; }
            else if (i < 48) { f = B ^ C ^ D; g = (3 * i + 5) & 15; }
            else { f = C ^ (B | ~D); g = (7 * i) & 15; }
            uint32_t t = D;
            D = C;
            C = B;
            B = B + rotl(A + f + K[i] + m[g], S[i]);
            A = t;
        }
        a += A; b += B; c += C; d += D;
    }
    void update(const uint8_t *p, size_t n) {
        total += n;
        if (fill) {
            size_t take = std::min<size_t>(64 - fill, n);
            std::memcpy(buf + fill, p, take);
            fill += (unsigned)take; p += take; n -= take;
            if (fill == 64) { block(buf); fill = 0; }
        }
        while (n >= 64) { block(p); p += 64; n -= 64; }
        if (n) { std::memcpy(buf, p, n); fill = (unsigned)n; }
    }
    void final(uint8_t out[16]) {
        uint64_t bitlen = total * 8;
        uint8_t pad[72] = {0x80};
        size_t padlen = (fill < 56) ? (56 - fill) : (120 - fill);
        update(pad, padlen);
        uint8_t lenb[8];
        for (int i = 0; i < 8; ++i) lenb[i] = (uint8_t)(bitlen >> (8 * i));
        update(lenb, 8);
        const uint32_t st[4] = {a, b, c, d};
        for (int i = 0; i < 4; ++i)
            for (int j = 0; j < 4; ++j) out[4 * i + j] = (uint8_t)(st[i] >> (8 * j));
    }
};

double now_s() {
    return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

template <typename S>
int run_container(const Streaminfo &si, ByteReader &reader, uint32_t flags, zo_result *res, unsigned aligned_bits) {
    Frames<S> fr;
    std::vector<zo_frame> trace;
    uint8_t ovf = 0;
    double t0 = now_s();
    int e = decode_frames<S>(si, reader, fr, ovf, (flags & ZO_FLAG_TRACE_FRAMES) ? &trace : nullptr);
    double t1 = now_s();
    res->t_frames = t1 - t0;
    res->overflow_flag = ovf;
    if (e != ZO_OK) return e;

    // ref :267-277 MD5 over the decoded samples, 3 bytes per sample for the 24-bit container
    Md5 md;
    if (aligned_bits == 24) {
        for (size_t i = 0; i < fr.len; ++i) md.update((const uint8_t *)&fr.samples[i], 3);
    } else {
        md.update((const uint8_t *)fr.samples, fr.len * sizeof(S));
    }
    md.final(res->md5_computed);
    double t2 = now_s();
    res->t_md5 = t2 - t1;

    const bool md5_ok = std::memcmp(res->md5_computed, si.md5, 16) == 0;
    if (!md5_ok && !(flags & ZO_FLAG_KEEP_ON_MD5_MISMATCH)) {  // ref :279-280
        free(fr.samples);
        return ZO_INVALID_CHECKSUM;
    }

    // ref :287-306 MSB-align shift, keyed on the STREAMINFO depth (Q2, Q5)
    const unsigned depth = si.sample_bit_depth + 1;
    if constexpr (sizeof(S) == 2) {
        if (depth >= 9 && depth <= 15)
            for (size_t i = 0; i < fr.len; ++i) fr.samples[i] = (S)((uint16_t)fr.samples[i] << (16 - depth));
    }
    if constexpr (sizeof(S) == 4) {
        if (depth >= 17 && depth <= 31)
            for (size_t i = 0; i < fr.len; ++i) fr.samples[i] = (S)((uint32_t)fr.samples[i] << (32 - depth));
    }
    res->t_shift = now_s() - t2;

    res->channels = fr.channels;
    res->sample_rate = fr.sample_rate;
    res->bits_per_sample = fr.bits_per_sample;
    res->container_bytes = sizeof(S);
    res->n_samples = fr.len;
    res->samples = fr.samples;
    if (flags & ZO_FLAG_TRACE_FRAMES) {
        res->n_frames = trace.size();
        res->frames = (zo_frame *)malloc(sizeof(zo_frame) * std::max<size_t>(trace.size(), 1));
        if (!trace.empty()) std::memcpy(res->frames, trace.data(), sizeof(zo_frame) * trace.size());
    }
    return md5_ok ? ZO_OK : ZO_INVALID_CHECKSUM;
}

}  // namespace

extern "C" {

// ref: :217-310 decode(allocator, reader)
int zo_decode(const uint8_t *data, size_t len, uint32_t flags, zo_result *res) {
    std::memset(res, 0, sizeof(*res));
    ByteReader reader{data, len};
    auto done = [&](int code) {
        res->status = code;
        return code;
    };

    uint32_t sig;
    int e = reader.read_be(4, sig);  // ref :218-220
    if (e != ZO_OK) return done(e);
    if (sig != 0x664C6143u) return done(ZO_INVALID_SIGNATURE);

    bool have_si = false;
    Streaminfo si{};
    for (;;) {  // ref :223-253
        uint32_t h;
        if ((e = reader.read_be(4, h)) != ZO_OK) return done(e);  // readStruct(MetadataHeader)
        const unsigned type = (h >> 24) & 0x7F;
        const bool last = (h >> 31) & 1;
        const uint32_t length = h & 0xFFFFFF;
        if (type == 0) {  // Streaminfo: length field is not used (Q13)
            BitReader br(reader);
            uint64_t v;
            if ((e = br.read_bits(16, v)) != ZO_OK) return done(e);
            si.min_block_size = (uint32_t)v;
            if ((e = br.read_bits(16, v)) != ZO_OK) return done(e);
            si.max_block_size = (uint32_t)v;
            if ((e = br.read_bits(24, v)) != ZO_OK) return done(e);
            si.min_frame_size = (uint32_t)v;
            if ((e = br.read_bits(24, v)) != ZO_OK) return done(e);
            si.max_frame_size = (uint32_t)v;
            if ((e = br.read_bits(20, v)) != ZO_OK) return done(e);
            si.sample_rate = (uint32_t)v;
            if ((e = br.read_bits(3, v)) != ZO_OK) return done(e);
            si.channel_count = (uint32_t)v;
            if ((e = br.read_bits(5, v)) != ZO_OK) return done(e);
            si.sample_bit_depth = (uint32_t)v;
            if ((e = br.read_bits(36, v)) != ZO_OK) return done(e);
            si.number_of_samples = v;
            if (reader.len - reader.pos < 16) return done(ZO_END_OF_STREAM);
            std::memcpy(si.md5, reader.p + reader.pos, 16);
            reader.pos += 16;
            have_si = true;
        } else if (type >= 1 && type <= 6) {  // ref :243-247
            if ((e = reader.skip(length)) != ZO_OK) return done(e);
        } else {
            return done(ZO_INVALID_METADATA_HEADER);  // ref :248
        }
        if (last) break;
    }
    if (!have_si) return done(ZO_MISSING_STREAMINFO);  // ref :309
    std::memcpy(res->md5_streaminfo, si.md5, 16);
    res->audio_offset = reader.pos;

    const unsigned depth = si.sample_bit_depth + 1;   // ref :256
    const unsigned aligned = (depth + 7u) & ~7u;      // ref :257
    switch (aligned) {                                 // ref :259-264
        case 8: e = run_container<int8_t>(si, reader, flags, res, aligned); break;
        case 16: e = run_container<int16_t>(si, reader, flags, res, aligned); break;
        case 24:
        case 32: e = run_container<int32_t>(si, reader, flags, res, aligned); break;
        default: e = ZO_UNIMPLEMENTED; break;
    }
    return done(e);
}

void zo_free(zo_result *res) {
    if (!res) return;
    free(res->samples);
    free(res->frames);
    res->samples = nullptr;
    res->frames = nullptr;
}

const char *zo_status_name(int status) {
    switch (status) {
        case ZO_OK: return "OK";
        case ZO_INVALID_SIGNATURE: return "InvalidSignature";
        case ZO_INVALID_METADATA_HEADER: return "InvalidMetadataHeader";
        case ZO_UNIMPLEMENTED: return "Unimplemented";
        case ZO_INVALID_CHECKSUM: return "InvalidChecksum";
        case ZO_MISSING_STREAMINFO: return "MissingStreaminfo";
        case ZO_INVALID_FRAME_HEADER: return "InvalidFrameHeader";
        case ZO_INVALID_CODED_NUMBER: return "InvalidCodedNumber";
        case ZO_INCONSISTENT_PARAMETERS: return "InconsistentParameters";
        case ZO_INVALID_SUBFRAME_HEADER: return "InvalidSubframeHeader";
        case ZO_INVALID_RESIDUAL_CODING_METHOD: return "InvalidResidualCodingMethod";
        case ZO_END_OF_STREAM: return "EndOfStream";
        case ZO_OUT_OF_MEMORY: return "OutOfMemory";
        case ZO_UNDEFINED_IN_REFERENCE: return "UndefinedInReference";
        default: return "Unknown";
    }
}

uint8_t zo_crc8(const uint8_t *p, size_t n) {
    uint8_t crc = 0;
    for (size_t i = 0; i < n; ++i) {
        crc ^= p[i];
        for (int b = 0; b < 8; ++b) crc = (uint8_t)((crc & 0x80) ? ((crc << 1) ^ 0x07) : (crc << 1));
    }
    return crc;
}

uint16_t zo_crc16(const uint8_t *p, size_t n) {
    uint16_t crc = 0;
    for (size_t i = 0; i < n; ++i) {
        crc ^= (uint16_t)((uint16_t)p[i] << 8);
        for (int b = 0; b < 8; ++b) crc = (uint16_t)((crc & 0x8000) ? ((crc << 1) ^ 0x8005) : (crc << 1));
    }
    return crc;
}

void zo_md5(const uint8_t *p, size_t n, uint8_t out[16]) {
    Md5 m;
    m.update(p, n);
    m.final(out);
}

}  // extern "C"
