// decode.cu — K2 decode_frames (sm_100a): one thread per frame.
//
// Restates, per frame, the reference's subframe loop (src/zflac.zig:425-544): subframe header
// and wasted bits (:426-441), CONSTANT (:445-454), VERBATIM (:455-465), FIXED (:466-498),
// LPC (:499-541, :604-612), decode_residuals / decode_residual_partition (:614-666) and the
// MSB-first bit reader (src/bit_reader.zig:25-120), then align + CRC-16 read (:546-551).
//
// Why one THREAD per frame: subframe c+1 starts at the bit where subframe c ends, and a Rice
// codeword's length depends on its own value, so parsing is a serial chain per frame. A warp
// cooperating on one chain (pointer jumping over unary terminators) costs ~4-5x the issue slots
// of 32 lanes each walking their own frame, and this kernel is issue-bound, not HBM-bound. So
// lanes own frames: a 64-bit MSB-aligned window refilled from a 4-word register queue that is
// loaded with one 128-bit LDG per 16 stream bytes; predictor coefficients and history live in
// registers (statically indexed, unrolled by 8); decoded samples leave as 128-bit stores into a
// planar per-frame scratch that K3 transposes into interleaved PCM.
//
// Arithmetic: the reference accumulates in i16/i32/i64 depending on the container (Q10). Here a
// subframe runs in int32 when depth + precision + ceil(log2(order)) <= 32 proves the sum cannot
// overflow, else in int64 (IMAD.WIDE) — both are exact wherever the reference is defined.
#include "common.cuh"
#include "kernels.cuh"

namespace zf {

__device__ __forceinline__ uint32_t shr_clamp(uint32_t v, uint32_t n) {  // n >= 32 -> 0 (PTX semantics)
    uint32_t r;
    asm("shr.b32 %0, %1, %2;" : "=r"(r) : "r"(v), "r"(n));
    return r;
}
__device__ __forceinline__ uint32_t shl_clamp(uint32_t v, uint32_t n) {
    uint32_t r;
    asm("shl.b32 %0, %1, %2;" : "=r"(r) : "r"(v), "r"(n));
    return r;
}

// ---------------------------------------------------------------------------------------
// Bit cursor. Invariants: the `cnt` valid bits sit at the top of `w`, everything below is 0;
// after refill() cnt >= 33. Bytes at or beyond `limit` (the stream length rounded up to 16)
// are never touched and read as zero; running past the real end is detected from the bit
// position (pos_bits() > 8*len), which is exactly when the reference's byte reader would have
// returned error.EndOfStream.
// ---------------------------------------------------------------------------------------
struct BitCursor {
    const uint8_t *base;
    uint64_t next_vec;  // byte offset of the next 16-byte vector to load
    uint64_t limit;     // loads at offsets >= limit return zeros
    uint64_t w;
    uint32_t q0, q1, q2, q3;
    int qn;
    int cnt;

    __device__ __forceinline__ void load_vec() {
        uint4 v = make_uint4(0, 0, 0, 0);
        if (next_vec < limit) v = __ldg(reinterpret_cast<const uint4 *>(base + next_vec));
        next_vec += 16;
        q0 = __byte_perm(v.x, 0, 0x0123);
        q1 = __byte_perm(v.y, 0, 0x0123);
        q2 = __byte_perm(v.z, 0, 0x0123);
        q3 = __byte_perm(v.w, 0, 0x0123);
        qn = 4;
    }
    __device__ __forceinline__ uint32_t pop_word() {
        if (qn == 0) load_vec();
        const uint32_t r = q0;
        q0 = q1; q1 = q2; q2 = q3;
        --qn;
        return r;
    }
    __device__ __forceinline__ void refill() {
        if (cnt <= 32) {
            w |= (uint64_t)pop_word() << (32 - cnt);
            cnt += 32;
        }
    }
    __device__ void init(const uint8_t *stream, uint64_t len, uint64_t byte_off) {
        base = stream;
        limit = (len + 15) & ~15ull;
        next_vec = byte_off & ~15ull;
        w = 0;
        cnt = 0;
        qn = 0;
        q0 = q1 = q2 = q3 = 0;
        uint32_t skip = (uint32_t)(byte_off & 15);
        load_vec();
        while (skip >= 4) {
            pop_word();
            skip -= 4;
        }
        refill();
        refill();
        if (skip) {
            w <<= 8 * skip;
            cnt -= 8 * skip;
        }
    }
    __device__ __forceinline__ uint64_t pos_bits() const { return next_vec * 8 - (uint64_t)qn * 32 - (uint64_t)cnt; }

    // n in [0, 32]
    __device__ __forceinline__ uint32_t read(uint32_t n) {
        refill();
        const uint32_t v = shr_clamp((uint32_t)(w >> 32), 32 - n);
        w <<= n;
        cnt -= (int)n;
        return v;
    }
    // n in [1, 32]
    __device__ __forceinline__ int32_t read_signed(uint32_t n) {
        const uint32_t v = read(n);
        return (int32_t)shl_clamp(v, 32 - n) >> (32 - n);
    }
    // n in [1, 64]
    __device__ __forceinline__ int64_t read_signed64(uint32_t n) {
        if (n <= 32) return (int64_t)read_signed(n);
        const uint32_t hi = read(n - 32);
        const uint32_t lo = read(32);
        const uint64_t v = ((uint64_t)hi << 32) | lo;
        return (int64_t)(v << (64 - n)) >> (64 - n);
    }
    // zeros before the next 1 bit; consumes the 1. Returns 0xFFFFFFFF if the stream ran out.
    __device__ uint32_t read_unary_slow(uint64_t len_bits) {
        uint32_t q = 0;
        for (;;) {
            refill();
            const uint32_t hi = (uint32_t)(w >> 32);
            if (hi != 0) {
                const uint32_t z = __clz(hi);
                q += z;
                w <<= (z + 1);
                cnt -= (int)(z + 1);
                return q;
            }
            q += 32;
            w <<= 32;
            cnt -= 32;
            if (pos_bits() > len_bits) return 0xFFFFFFFFu;
        }
    }
    __device__ __forceinline__ uint32_t read_unary(uint64_t len_bits) {
        refill();
        const uint32_t hi = (uint32_t)(w >> 32);
        if (hi != 0) {
            const uint32_t z = __clz(hi);
            w <<= (z + 1);
            cnt -= (int)(z + 1);
            return z;
        }
        return read_unary_slow(len_bits);
    }
    __device__ __forceinline__ void align_byte() {
        const uint32_t drop = (8u - (uint32_t)(pos_bits() & 7)) & 7u;
        read(drop);
    }
};

// ---------------------------------------------------------------------------------------
// Residual decoder state (decode_residuals / decode_residual_partition, :614-666). Partition
// headers are read lazily, when the first residual of the partition is requested: nothing
// else is read in between, so the bit order equals the reference's.
// ---------------------------------------------------------------------------------------
struct Rice {
    uint32_t k;          // Rice parameter of the current partition
    uint32_t esc_bits;   // escape width when escaped
    uint32_t remaining;  // residuals left in the current partition
    uint32_t part_size;
    uint32_t parts_left;
    uint32_t first_sub;  // predictor order, taken off the first partition
    uint32_t param_bits; // 4 (Rice) or 5 (Rice2)
    bool escaped;
    uint64_t len_bits;
    uint32_t error;

    __device__ void begin(BitCursor &bc, uint32_t block_size, uint32_t order, uint64_t lb) {
        len_bits = lb;
        error = ST_OK;
        const uint32_t method = bc.read(2);
        if (method >= 2) {
            error = ST_INVALID_RESIDUAL_CODING_METHOD;
            return;
        }
        param_bits = 4 + method;
        const uint32_t po = bc.read(4);
        part_size = block_size >> po;
        // not divisible / first partition shorter than the order: the reference under-fills or
        // underflows a u16 (Q16) -> undefined there, refused here
        if ((part_size << po) != block_size || part_size < order) {
            error = ST_UNDEFINED_IN_REFERENCE;
            return;
        }
        parts_left = 1u << po;
        remaining = 0;
        first_sub = order;
        escaped = false;
        k = 0;
        esc_bits = 0;
    }
    __device__ __forceinline__ void open_partition(BitCursor &bc) {
        const uint32_t p = bc.read(param_bits);
        if (p == (1u << param_bits) - 1u) {
            escaped = true;
            esc_bits = bc.read(5);
        } else {
            escaped = false;
            k = p;
        }
        remaining = part_size - first_sub;
        first_sub = 0;
        --parts_left;
    }
    __device__ __forceinline__ int32_t next(BitCursor &bc) {
        while (remaining == 0 && parts_left > 0) open_partition(bc);
        --remaining;
        if (!escaped) {
            bc.refill();
            const uint32_t hi = (uint32_t)(bc.w >> 32);
            const uint32_t z = __clz(hi);  // 32 when hi == 0
            const uint32_t n = z + 1 + k;
            uint32_t u;
            if (n <= 32) {  // whole codeword inside the top word: the common case
                const uint32_t r = shr_clamp(shl_clamp(hi, z + 1), 32 - k);
                bc.w <<= n;
                bc.cnt -= (int)n;
                u = (z << k) | r;
            } else {
                const uint32_t q = bc.read_unary(len_bits);
                if (q == 0xFFFFFFFFu) {
                    error = ST_END_OF_STREAM;
                    remaining = 0;
                    parts_left = 0;
                    return 0;
                }
                const uint32_t r = bc.read(k);
                u = (q << k) + r;
            }
            return (int32_t)((u >> 1) ^ (0u - (u & 1u)));
        }
        return esc_bits ? bc.read_signed(esc_bits) : 0;
    }
    // zero-length trailing partitions still carry a header in the stream
    __device__ void finish(BitCursor &bc) {
        while (parts_left > 0 && error == ST_OK) {
            open_partition(bc);
            if (remaining != 0) break;  // cannot happen on a consistent subframe
        }
    }
};

template <typename ST> __device__ __forceinline__ void store1(ST *out, uint32_t i, int64_t v, bool wr) {
    if (wr) out[i] = (ST)v;
}

// ---------------------------------------------------------------------------------------
// Predictor run for int32 samples: P = register window length (>= order), ACC = int32_t or
// int64_t. c[j] multiplies s[i-1-j]; h[P-1] is the most recent sample. Unused taps are zero.
// ---------------------------------------------------------------------------------------
template <int P, typename ACC>
__device__ __forceinline__ void predict_run(BitCursor &bc, Rice &rc, int32_t *__restrict__ out, uint32_t order,
                                            uint32_t block_size, const int32_t (&c)[P], int32_t (&h)[P],
                                            uint32_t shift, uint32_t wasted, bool wr) {
    uint32_t i = order;
    const uint32_t head_end = min(block_size, (order + 7u) & ~7u);
#pragma unroll 1
    for (int phase = 0; phase < 2; ++phase) {
        const uint32_t lim = phase == 0 ? head_end : block_size;
#pragma unroll 1
        for (; i < lim; ++i) {  // one sample at a time: head up to an 8-aligned index, and the tail
            ACC acc = 0;
#pragma unroll
            for (int j = 0; j < P; ++j) acc += (ACC)c[j] * (ACC)h[P - 1 - j];
            const int32_t v = rc.next(bc) + (int32_t)(acc >> shift);
#pragma unroll
            for (int j = 0; j < P - 1; ++j) h[j] = h[j + 1];
            h[P - 1] = v;
            if (wr) out[i] = (int32_t)((uint32_t)v << wasted);
        }
        if (phase == 0) {
#pragma unroll 1
            for (; i + 8 <= block_size; i += 8) {  // 8 samples per trip, window indices static
                int32_t x[P + 8];
#pragma unroll
                for (int j = 0; j < P; ++j) x[j] = h[j];
#pragma unroll
                for (int u = 0; u < 8; ++u) {
                    ACC acc = 0;
#pragma unroll
                    for (int j = 0; j < P; ++j) acc += (ACC)c[j] * (ACC)x[P + u - 1 - j];
                    x[P + u] = rc.next(bc) + (int32_t)(acc >> shift);
                }
#pragma unroll
                for (int j = 0; j < P; ++j) h[j] = x[j + 8];
                if (wr) {
                    int4 a, b;
                    a.x = (int32_t)((uint32_t)x[P + 0] << wasted);
                    a.y = (int32_t)((uint32_t)x[P + 1] << wasted);
                    a.z = (int32_t)((uint32_t)x[P + 2] << wasted);
                    a.w = (int32_t)((uint32_t)x[P + 3] << wasted);
                    b.x = (int32_t)((uint32_t)x[P + 4] << wasted);
                    b.y = (int32_t)((uint32_t)x[P + 5] << wasted);
                    b.z = (int32_t)((uint32_t)x[P + 6] << wasted);
                    b.w = (int32_t)((uint32_t)x[P + 7] << wasted);
                    *reinterpret_cast<int4 *>(out + i) = a;
                    *reinterpret_cast<int4 *>(out + i + 4) = b;
                }
            }
        }
    }
}

__constant__ int32_t c_fixed_coef[5][4] = {{0, 0, 0, 0}, {1, 0, 0, 0}, {2, -1, 0, 0}, {3, -3, 1, 0}, {4, -6, 4, -1}};

// One FIXED or LPC subframe with int32 samples, in stream order: warm-up samples, then (LPC)
// precision / shift / coefficients, then the residual (:466-541). Window slots are chosen with
// predicated selects so every register index stays static.
template <int P, typename ACC>
__device__ uint32_t predicted_subframe(BitCursor &bc, int32_t *__restrict__ out, uint32_t order, bool is_lpc,
                                       uint32_t eff_bits, uint32_t block_size, uint32_t wasted, uint64_t len_bits,
                                       bool wr) {
    int32_t h[P], c[P];
#pragma unroll
    for (int j = 0; j < P; ++j) {
        h[j] = 0;
        c[j] = 0;
    }
    // warm-up samples s[0..order-1]; s[t] belongs at h[P - order + t]
#pragma unroll 1
    for (uint32_t t = 0; t < order; ++t) {
        const int32_t v = bc.read_signed(eff_bits);
        const int slot = P - (int)order + (int)t;
#pragma unroll
        for (int j = 0; j < P; ++j) h[j] = (j == slot) ? v : h[j];
        if (wr) out[t] = (int32_t)((uint32_t)v << wasted);
    }
    uint32_t shift = 0;
    if (is_lpc) {
        const uint32_t precision = bc.read(4) + 1;  // :508 (0b1111 is not rejected by the reference)
        shift = bc.read(5);                         // :510 unsigned (Q17)
#pragma unroll 1
        for (uint32_t t = 0; t < order; ++t) {      // t-th coefficient multiplies s[i-1-t] (:513-514)
            const int32_t v = bc.read_signed(precision);
#pragma unroll
            for (int j = 0; j < P; ++j) c[j] = (j == (int)t) ? v : c[j];
        }
    } else {
#pragma unroll
        for (int j = 0; j < (P < 4 ? P : 4); ++j) c[j] = c_fixed_coef[order][j];
    }
    Rice rc;
    rc.begin(bc, block_size, order, len_bits);
    if (rc.error != ST_OK) return rc.error;
    predict_run<P, ACC>(bc, rc, out, order, block_size, c, h, shift, wasted, wr);
    if (rc.error == ST_OK) rc.finish(bc);
    return rc.error;
}

__device__ __forceinline__ uint32_t ceil_log2(uint32_t v) { return v <= 1 ? 0 : 32 - __clz(v - 1); }

// Dispatch on register-window length and accumulator width.
__device__ uint32_t decode_predicted32(BitCursor &bc, int32_t *__restrict__ out, uint32_t order, bool is_lpc,
                                       uint32_t depth, uint32_t eff_bits, uint32_t block_size, uint32_t wasted,
                                       uint64_t len_bits, bool wr) {
    if (!is_lpc) {
        // |sum| <= 15 * 2^(depth-1) < 2^(depth+3)
        if (depth + 4 <= 32) return predicted_subframe<4, int32_t>(bc, out, order, false, eff_bits, block_size, wasted, len_bits, wr);
        return predicted_subframe<4, int64_t>(bc, out, order, false, eff_bits, block_size, wasted, len_bits, wr);
    }
    // The accumulator width depends on the coefficient precision, which sits after the warm-up
    // samples: peek at it with a copy of the cursor.
    BitCursor peek = bc;
#pragma unroll 1
    for (uint32_t t = 0; t < order; ++t) peek.read(eff_bits);
    const uint32_t precision = peek.read(4) + 1;
    const bool wide = depth + precision + ceil_log2(order) > 32;
#define ZF_RUN(PP)                                                                                                  \
    return wide ? predicted_subframe<PP, int64_t>(bc, out, order, true, eff_bits, block_size, wasted, len_bits, wr) \
                : predicted_subframe<PP, int32_t>(bc, out, order, true, eff_bits, block_size, wasted, len_bits, wr)
    if (order <= 8) ZF_RUN(8);
    if (order <= 12) ZF_RUN(12);
    ZF_RUN(32);
#undef ZF_RUN
}

// Generic path for 32-bit-per-sample streams (side channel needs 33 bits): everything in int64,
// history in a small local array. Rare; correctness over speed.
__device__ uint32_t decode_predicted64(BitCursor &bc, int64_t *__restrict__ out, uint32_t order, bool is_lpc,
                                       uint32_t eff_bits, uint32_t block_size, uint32_t wasted, uint64_t len_bits,
                                       bool wr) {
    int64_t hist[32];
    int32_t coef[32];
    for (uint32_t t = 0; t < order; ++t) {
        hist[t] = bc.read_signed64(eff_bits);
        if (wr) out[t] = (int64_t)((uint64_t)hist[t] << wasted);
    }
    uint32_t shift = 0;
    if (is_lpc) {
        const uint32_t precision = bc.read(4) + 1;
        shift = bc.read(5);
        for (uint32_t t = 0; t < order; ++t) coef[t] = bc.read_signed(precision);
    } else {
        for (uint32_t t = 0; t < 4; ++t) coef[t] = c_fixed_coef[order][t];
    }
    Rice rc;
    rc.begin(bc, block_size, order, len_bits);
    if (rc.error != ST_OK) return rc.error;
    // hist is a ring: sample i lives at hist[i & 31]
    for (uint32_t i = order; i < block_size; ++i) {
        int64_t acc = 0;
        for (uint32_t t = 0; t < order; ++t) acc += (int64_t)coef[t] * hist[(i - 1 - t) & 31];
        // Rice residuals are decoded in 32 bits (RFC 9639 limits them to int32); escape codes too
        const int64_t v = (int64_t)rc.next(bc) + (acc >> shift);
        hist[i & 31] = v;
        if (wr) out[i] = (int64_t)((uint64_t)v << wasted);
    }
    if (rc.error == ST_OK) rc.finish(bc);
    return rc.error;
}

template <typename ST> struct IsWide { static constexpr bool value = false; };
template <> struct IsWide<int64_t> { static constexpr bool value = true; };

constexpr int kDecodeThreads = 128;

// ST = int32_t: streams whose frames carry <= 31 bits per sample (subframe depth <= 32).
// ST = int64_t: 32 bps streams.
template <typename ST>
__global__ void __launch_bounds__(kDecodeThreads)
k_decode_frames(const uint8_t *__restrict__ s, uint64_t len, const FrameRec *__restrict__ frames, uint32_t n_frames,
                ST *__restrict__ scratch, FrameOut *__restrict__ outs, int write_samples) {
    const uint32_t f = blockIdx.x * blockDim.x + threadIdx.x;
    if (f >= n_frames) return;
    const FrameRec fr = frames[f];
    const bool wr = write_samples != 0;
    const uint64_t len_bits = len * 8;
    BitCursor bc;
    bc.init(s, len, fr.offset + fr.header_len);
    uint32_t status = ST_OK;
    const uint32_t bs = fr.block_size;
    const uint32_t bps = fr.bps;

#pragma unroll 1
    for (uint32_t ch = 0; ch < fr.channels && status == ST_OK; ++ch) {
        const uint32_t hdr = bc.read(8);  // zero:1 type:6 wasted_flag:1 (:426-430)
        if (hdr & 0x80u) {
            status = ST_INVALID_SUBFRAME_HEADER;
            break;
        }
        const uint32_t type = (hdr >> 1) & 0x3Fu;
        uint32_t wasted = 0;
        if (hdr & 1u) {
            const uint32_t u = bc.read_unary(len_bits);
            if (u == 0xFFFFFFFFu) { status = ST_END_OF_STREAM; break; }
            if (u + 1 > 63) { status = ST_UNDEFINED_IN_REFERENCE; break; }
            wasted = u + 1;  // :433
        }
        uint32_t depth = bps;  // :436-441
        if ((fr.ch_code == 8 && ch == 1) || (fr.ch_code == 9 && ch == 0) || (fr.ch_code == 10 && ch == 1)) depth += 1;
        ST *out = scratch + fr.scratch_base + (uint64_t)ch * fr.scratch_pitch;

        if (type == 0) {  // CONSTANT (:445-454): reads bps - wasted bits even on a side channel (Q8)
            if (wasted >= bps) { status = ST_UNDEFINED_IN_REFERENCE; break; }
            int64_t v;
            if (IsWide<ST>::value) v = bc.read_signed64(bps - wasted);
            else v = bc.read_signed(bps - wasted);
            v = (int64_t)((uint64_t)v << wasted);
            if (wr)
                for (uint32_t i = 0; i < bs; ++i) out[i] = (ST)v;
        } else if (type == 1) {  // VERBATIM (:455-465)
            if (wasted >= depth) { status = ST_UNDEFINED_IN_REFERENCE; break; }
            const uint32_t n = depth - wasted;
#pragma unroll 1
            for (uint32_t i = 0; i < bs; ++i) {
                int64_t v;
                if (IsWide<ST>::value) v = bc.read_signed64(n);
                else v = bc.read_signed(n);
                store1(out, i, (int64_t)((uint64_t)v << wasted), wr);
                if ((i & 255u) == 255u && bc.pos_bits() > len_bits) break;  // ran off the stream
            }
        } else if ((type >= 8 && type <= 12) || type >= 32) {  // FIXED (:466-498) / LPC (:499-541)
            const bool is_lpc = type >= 32;
            const uint32_t order = is_lpc ? type - 31 : type - 8;
            if (wasted >= depth || order > bs) { status = ST_UNDEFINED_IN_REFERENCE; break; }
            const uint32_t eff = depth - wasted;
            if (IsWide<ST>::value)
                status = decode_predicted64(bc, (int64_t *)out, order, is_lpc, eff, bs, wasted, len_bits, wr);
            else
                status = decode_predicted32(bc, (int32_t *)out, order, is_lpc, eff, eff, bs, wasted, len_bits, wr);
        } else {  // reserved (:542)
            status = ST_INVALID_SUBFRAME_HEADER;
        }
    }
    uint32_t stored_crc = 0;
    if (status == ST_OK) {
        bc.align_byte();            // :546
        stored_crc = bc.read(16);   // :548, not checked by the reference
    }
    const uint64_t pos = bc.pos_bits();
    // Whatever went wrong, if the cursor is past the end the reference's reader failed first.
    if (pos > len_bits) status = ST_END_OF_STREAM;
    FrameOut o;
    o.end_offset = pos >> 3;
    o.status = status;
    o.stored_crc16 = stored_crc;
    outs[f] = o;
}

cudaError_t launch_decode_frames(const uint8_t *d_stream, uint64_t len, const FrameRec *d_frames, uint32_t n_frames,
                                 void *d_scratch, bool wide_scratch, FrameOut *d_outs, bool write_samples,
                                 cudaStream_t st) {
    if (n_frames == 0) return cudaSuccess;
    const uint32_t blocks = (n_frames + kDecodeThreads - 1) / kDecodeThreads;
    ZF_COUNT_LAUNCH();
    if (wide_scratch)
        k_decode_frames<int64_t><<<blocks, kDecodeThreads, 0, st>>>(d_stream, len, d_frames, n_frames, (int64_t *)d_scratch,
                                                                    d_outs, write_samples ? 1 : 0);
    else
        k_decode_frames<int32_t><<<blocks, kDecodeThreads, 0, st>>>(d_stream, len, d_frames, n_frames, (int32_t *)d_scratch,
                                                                    d_outs, write_samples ? 1 : 0);
    return cudaGetLastError();
}

}  // namespace zf
