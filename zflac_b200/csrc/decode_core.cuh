// decode_core.cuh — the per-frame subframe decoder of K2, written against an abstract word
// source so that the same code runs
//   * on the device, fed by a per-lane shared-memory word ring that cp.async keeps full
//     (RingSrc in decode.cu), and
//   * on the host inside tests/host_core/ (HostSrc), where the bit-level logic is unit-tested
//     against the CPU restatement without a GPU. libzflacb200.so never instantiates the host source:
//     the product has no CPU decode path.
//
// Restates, per frame, the reference's subframe loop (src/zflac.zig:425-544): subframe header
// and wasted bits (:426-441), CONSTANT (:445-454), VERBATIM (:455-465), FIXED (:466-498),
// LPC (:499-541, :604-612), decode_residuals / decode_residual_partition (:614-666) and the
// MSB-first bit reader (src/bit_reader.zig:25-120), then align + CRC-16 read (:546-551).
//
// Why one LANE per frame: subframe c+1 starts at the bit where subframe c ends and a Rice
// codeword's length depends on its own value, so parsing is one serial chain per frame; 32
// lanes walking 32 frames cost a quarter of the issue slots of a warp cooperating on one
// chain. All lanes of a warp execute the same row (sample index) of their own frame.
//
// Bit reader: two 32-bit words A:B plus a prefetched third (N) and a bit offset o < 32. The
// next 32 stream bits are one funnel shift; consuming n <= 32 bits is `o += n` and, when o
// wraps, a predicated shift of the word pipeline fed by ONE shared-memory load. No 64-bit
// shifts, no per-lane queue management, no global load on the critical path.
//
// Arithmetic: the reference accumulates in i16/i32/i64 depending on the container (Q10). Here a
// subframe runs in int32 when depth + precision + ceil(log2(order)) <= 32 proves the sum cannot
// overflow for samples of `depth` bits, else in int64; in a 32-bit container (`inter64`: the reference's
// i64) every shifted sum runs in int64, because a malformed stream may carry samples wider than `depth`
// there without leaving the reference's defined range — exact wherever the reference is defined.
#pragma once
#include "common.cuh"

#define ZF_HD __host__ __device__ __forceinline__
// Everything that takes the bit reader by reference is force-inlined: a real call would make the
// reader's address escape and push its registers (A, B, N, o, ...) into local memory.
#define ZF_HDI __host__ __device__ __forceinline__
#if defined(__CUDACC__)
#define ZF_NOINLINE __noinline__
#else
#define ZF_NOINLINE __attribute__((noinline))
#endif

namespace zf {

ZF_HD uint32_t clz32(uint32_t v) {
#ifdef __CUDA_ARCH__
    return (uint32_t)__clz((int)v);
#else
    return v ? (uint32_t)__builtin_clz(v) : 32u;
#endif
}
// leading zeros as a shift amount; 0xFFFFFFFF for v == 0 (one FLO.SH instead of FLO + subtract)
ZF_HD uint32_t clz_sh(uint32_t v) {
#ifdef __CUDA_ARCH__
    uint32_t r;
    asm("bfind.shiftamt.u32 %0, %1;" : "=r"(r) : "r"(v));
    return r;
#else
    return v ? (uint32_t)__builtin_clz(v) : 0xFFFFFFFFu;
#endif
}
ZF_HD uint32_t shr_c(uint32_t v, uint32_t n) {  // n >= 32 -> 0
#ifdef __CUDA_ARCH__
    uint32_t r;
    asm("shr.b32 %0, %1, %2;" : "=r"(r) : "r"(v), "r"(n));
    return r;
#else
    return n >= 32 ? 0u : v >> n;
#endif
}
ZF_HD uint32_t shl_c(uint32_t v, uint32_t n) {  // n >= 32 -> 0
#ifdef __CUDA_ARCH__
    uint32_t r;
    asm("shl.b32 %0, %1, %2;" : "=r"(r) : "r"(v), "r"(n));
    return r;
#else
    return n >= 32 ? 0u : v << n;
#endif
}
// top 32 bits of (hi:lo) << (s & 31)
ZF_HD uint32_t funnel_hi(uint32_t hi, uint32_t lo, uint32_t s) {
#ifdef __CUDA_ARCH__
    return __funnelshift_l(lo, hi, s);
#else
    s &= 31;
    return s ? (hi << s) | (lo >> (32 - s)) : hi;
#endif
}
ZF_HD uint32_t ceil_log2(uint32_t v) { return v <= 1 ? 0 : 32 - clz32(v - 1); }
ZF_HD int32_t zigzag(uint32_t u) {
#ifdef __CUDA_ARCH__
    int32_t m;  // sign-extend bit 0: one SGXT instead of AND + negate
    asm("bfe.s32 %0, %1, 0, 1;" : "=r"(m) : "r"(u));
    return (int32_t)(u >> 1) ^ m;
#else
    return (int32_t)((u >> 1) ^ (0u - (u & 1u)));
#endif
}

// ---------------------------------------------------------------------------------------
// Src concept (word index w counts 32-bit big-endian stream words from the source origin):
//   uint32_t word(w)           word w in the source's raw byte order; the caller guarantees it is resident
//   uint32_t word_checked(w)   same, but makes the word resident first (slow paths)
//   static uint32_t fix(raw)   raw -> MSB-first
//   void     top_up(w)         words < w are consumed; refill, and guarantee that words
//                              [w, w + kFastWords) are resident (before a fast 8-row body)
// Words at or past the end of the stream read as zero; running off the stream is detected
// from the bit position, exactly when the reference's byte reader returns EndOfStream.
// ---------------------------------------------------------------------------------------
constexpr uint32_t kFastWords = 9;  // a fast body re-reads word next-1 and pulls at most 8 more: next-1 .. next+7

template <class Src>
struct BitReader {
    Src src;
    uint32_t A, B, N;  // words next-3, next-2, next-1 (N still in the source's raw byte order: the swap is
                       // applied when it moves into B, so the shared load is never consumed right after issue)
    uint32_t o;        // bits of A already consumed, 0..31
    uint32_t next;     // index of the word after N

    ZF_HD void init(uint32_t skip_bytes) {  // skip_bytes: 0..3
        A = Src::fix(src.word_checked(0));
        B = Src::fix(src.word_checked(1));
        N = src.word_checked(2);
        next = 3;
        o = 8 * skip_bytes;
    }
    ZF_HD void init_at(uint32_t skip_bytes) {  // any byte offset from word 0
        const uint32_t w = skip_bytes >> 2;
        A = Src::fix(src.word_checked(w));
        B = Src::fix(src.word_checked(w + 1));
        N = src.word_checked(w + 2);
        next = w + 3;
        o = 8 * (skip_bytes & 3u);
    }
    ZF_HD uint64_t pos_bits() const { return (uint64_t)(next - 3) * 32 + o; }
    ZF_HD uint32_t peek() const { return funnel_hi(A, B, o); }
    template <bool CHECKED> ZF_HD void skip(uint32_t n) {  // n <= 32
        o += n;
        if (CHECKED) {
            if (o >= 32) {
                o -= 32;
                A = B;
                B = Src::fix(N);
                N = src.word_checked(next);
                ++next;
            }
        } else {
            // Branch-free (fast 8-row bodies): the word pipeline advances under a predicate and N is
            // RE-loaded from its own (possibly advanced) index every row, so no select ever sits on a
            // value fresh out of shared memory and the compiler can schedule across rows.
            const bool wrap = o >= 32;
            o &= 31u;
            A = wrap ? B : A;
            B = wrap ? Src::fix(N) : B;
            next = wrap ? next + 1 : next;
            N = src.word(next - 1);
        }
    }
    ZF_HD uint32_t read(uint32_t n) {  // n in [0, 32]
        const uint32_t v = shr_c(peek(), 32 - n);
        skip<true>(n);
        return v;
    }
    ZF_HD int32_t read_signed(uint32_t n) {  // n in [1, 32]
        const uint32_t v = peek();
        skip<true>(n);
        return (int32_t)v >> (32 - n);
    }
    ZF_HD int64_t read_signed64(uint32_t n) {  // n in [1, 64]
        if (n <= 32) return (int64_t)read_signed(n);
        const uint32_t hi = read(n - 32);
        const uint32_t lo = read(32);
        const uint64_t v = ((uint64_t)hi << 32) | lo;
        return (int64_t)(v << (64 - n)) >> (64 - n);
    }
    // zeros before the next 1 bit; consumes the 1. 0xFFFFFFFF if the stream ran out first.
    ZF_HDI uint32_t read_unary(uint64_t len_bits) {
        uint32_t q = 0;
        for (;;) {
            const uint32_t w = peek();
            if (w != 0) {
                const uint32_t z = clz32(w);
                skip<true>(z + 1);
                return q + z;
            }
            q += 32;
            skip<true>(32);
            if (pos_bits() > len_bits) return 0xFFFFFFFFu;
        }
    }
    // Branch-free advance of the fast 8-row bodies, n < 2^31 (garbage lengths of a body that is rolled
    // back included): one word step at most, every update a predicated single instruction. Inside a
    // body the word cursor is carried pre-scaled for the source's ring (ws = (next - 1) * Src::kWordScale):
    // enter_fast() / leave_fast() convert.
    ZF_HD uint32_t enter_fast() const { return (next - 1) * Src::kWordScale; }
    ZF_HD void leave_fast(uint32_t ws) { next = ws / Src::kWordScale + 1; }
    ZF_HD void skip_fast(uint32_t n, uint32_t &ws) {
        o += n;
#ifdef __CUDA_ARCH__
        asm("{\n\t.reg .pred p;\n\t"
            "setp.gt.u32 p, %0, 31;\n\t"
            "and.b32 %0, %0, 31;\n\t"
            "@p mov.b32 %1, %2;\n\t"
            "@p prmt.b32 %2, %3, 0, 0x0123;\n\t"
            "@p add.u32 %4, %4, %5;\n\t}"
            : "+r"(o), "+r"(A), "+r"(B), "+r"(N), "+r"(ws)
            : "n"(Src::kWordScale));
#else
        const bool wrap = o >= 32;
        o &= 31u;
        A = wrap ? B : A;
        B = wrap ? Src::fix(N) : B;
        ws = wrap ? ws + Src::kWordScale : ws;
#endif
        N = src.word_scaled(ws);
    }
    ZF_HD void align_byte() { skip<true>((8u - (o & 7u)) & 7u); }  // word starts are byte aligned
    // N (word next-1) is re-read by the branch-free skip: it still counts as live
    ZF_HD void prepare_fast() { src.top_up(next - 1); }
    // `tick`: a value converged lanes share (the body index): sources that refill every other body key on it
    ZF_HD void prepare_fast(uint32_t tick) { src.top_up(next - 1, tick); }
};

ZF_HD int32_t fixed_coef(uint32_t order, int j) {  // :481-490 {1; 2,-1; 3,-3,1; 4,-6,4,-1}
    const int32_t t0 = order == 1 ? 1 : order == 2 ? 2 : order == 3 ? 3 : order == 4 ? 4 : 0;
    const int32_t t1 = order == 2 ? -1 : order == 3 ? -3 : order == 4 ? -6 : 0;
    const int32_t t2 = order == 3 ? 1 : order == 4 ? 4 : 0;
    const int32_t t3 = order == 4 ? -1 : 0;
    return j == 0 ? t0 : j == 1 ? t1 : j == 2 ? t2 : j == 3 ? t3 : 0;
}

// One Rice codeword with parameter k through the slow (checked, unbounded unary) path. Kept out
// of line (it is rare: a codeword longer than 32 bits); the caller hands over a COPY of the
// reader so that only the copy's address escapes.
template <class Src>
__host__ __device__ ZF_NOINLINE uint32_t rice_slow_outlined(BitReader<Src> *br, uint32_t k, uint64_t len_bits, uint32_t *err) {
    const uint32_t q = br->read_unary(len_bits);
    if (q == 0xFFFFFFFFu) {
        *err = ST_END_OF_STREAM;
        return 0;
    }
    const uint32_t r = br->read(k);
    return (q << k) + r;
}
// The same with a 64-bit codeword: in a 32-bit container the reference's residuals are i64 (the InterType, :614-666), and a
// full-scale 32 bps signal has residuals of 33 bits whatever RFC 9639 says about encoders.
template <class Src>
__host__ __device__ ZF_NOINLINE uint64_t rice_slow64_outlined(BitReader<Src> *br, uint32_t k, uint64_t len_bits, uint32_t *err) {
    const uint32_t q = br->read_unary(len_bits);
    if (q == 0xFFFFFFFFu) {
        *err = ST_END_OF_STREAM;
        return 0;
    }
    const uint32_t r = br->read(k);
    return ((uint64_t)q << k) + r;
}
template <class Src>
ZF_HDI uint32_t rice_slow(BitReader<Src> &br, uint32_t k, uint64_t len_bits, uint32_t &err) {
    BitReader<Src> t = br;
    uint32_t e = ST_OK;
    const uint32_t cw = rice_slow_outlined(&t, k, len_bits, &e);
    br = t;
    if (e != ST_OK) err = e;
    return cw;
}

// ---------------------------------------------------------------------------------------
// FIXED / LPC subframe with int32 samples, in stream order: warm-up samples, (LPC) precision,
// shift, coefficients, then the partitioned residual (:466-541, :614-666). P = register window
// length (>= order); c[j] multiplies s[i-1-j]; h[P-1] is the most recent sample; unused taps are
// zero. Window slots are picked with predicated selects so every register index stays static.
// ---------------------------------------------------------------------------------------
template <int P, typename ACC, class Src>
ZF_HDI uint32_t predicted_subframe(BitReader<Src> &br, int32_t *__restrict__ out, const int32_t *warm, uint32_t order,
                                   bool is_lpc, uint32_t precision, uint32_t shift, uint32_t block_size, uint32_t wasted,
                                   uint64_t len_bits, bool wr) {
    int32_t h[P], c[P];
#pragma unroll
    for (int j = 0; j < P; ++j) {
        h[j] = 0;
        c[j] = 0;
    }
#pragma unroll 1
    for (uint32_t t = 0; t < order; ++t) {  // warm-up s[0..order-1] (already read); s[t] belongs at h[P - order + t]
        const int32_t v = warm[t];
        const int slot = P - (int)order + (int)t;
#pragma unroll
        for (int j = 0; j < P; ++j) h[j] = (j == slot) ? v : h[j];
    }
    if (is_lpc) {
#pragma unroll 1
        for (uint32_t t = 0; t < order; ++t) {      // t-th coefficient multiplies s[i-1-t] (:513-514)
            const int32_t v = br.read_signed(precision);
#pragma unroll
            for (int j = 0; j < P; ++j) c[j] = (j == (int)t) ? v : c[j];
        }
    } else {
#pragma unroll
        for (int j = 0; j < (P < 4 ? P : 4); ++j) c[j] = fixed_coef(order, j);
    }

    // ---- residual (:614-633)
    const uint32_t method = br.read(2);
    if (method >= 2) return ST_INVALID_RESIDUAL_CODING_METHOD;
    const uint32_t pbits = 4 + method;
    const uint32_t po = br.read(4);
    const uint32_t psize = block_size >> po;
    // not divisible / first partition shorter than the order: the reference under-fills or
    // underflows a u16 (Q16) -> undefined there, refused here
    if ((psize << po) != block_size || psize < order) return ST_UNDEFINED_IN_REFERENCE;
    uint32_t err = ST_OK;
    uint32_t i = order;
#pragma unroll 1
    for (uint32_t part = 0; part < (1u << po); ++part) {
        const uint32_t param = br.read(pbits);
        const uint32_t end = (part + 1) * psize;
        if (param == (1u << pbits) - 1u) {  // escape: raw signed samples of esc bits (:646-654)
            const uint32_t esc = br.read(5);
#pragma unroll 1
            for (; i < end; ++i) {
                const int32_t r = esc ? br.read_signed(esc) : 0;
                ACC acc = 0;
#pragma unroll
                for (int j = 0; j < P; ++j) acc += (ACC)c[j] * (ACC)h[P - 1 - j];
                const int32_t v = r + (int32_t)(acc >> shift);
#pragma unroll
                for (int j = 0; j < P - 1; ++j) h[j] = h[j + 1];
                h[P - 1] = v;
                if (wr) out[i] = (int32_t)((uint32_t)v << wasted);
            }
        } else {
            const uint32_t k = param, kp1 = param + 1, rsh = 32 - param;
            uint32_t force_single = 0;  // rows that must take the checked one-row path (after a rolled-back body)
#pragma unroll 1
            while (i < end) {
                if (force_single == 0 && (i & 7u) == 0 && i + 8 <= end) {
                    // ---- 8 rows, speculative: straight-line code, unchecked word loads, static window
                    // indices. A codeword longer than the 32-bit window (rare) only raises `bad`; the
                    // body is then rolled back and its rows are redone one by one below.
                    br.prepare_fast();
                    const uint32_t sA = br.A, sB = br.B, sN = br.N, so = br.o, snext = br.next;
                    uint32_t bad = 0;
                    int32_t x[P + 8];
#pragma unroll
                    for (int j = 0; j < P; ++j) x[j] = h[j];
#pragma unroll
                    for (int u = 0; u < 8; ++u) {
                        const uint32_t win = br.peek();
                        const uint32_t z = clz32(win);
                        const uint32_t n = z + kp1;
                        bad |= (n > 32u) ? 1u : 0u;
                        const uint32_t cw = (z << k) | shr_c(shl_c(win, z + 1), rsh);
                        br.template skip<false>(n & 63u);
                        ACC acc = 0;
#pragma unroll
                        for (int j = P - 1; j >= 0; --j) acc += (ACC)c[j] * (ACC)x[P + u - 1 - j];  // newest tap last
                        x[P + u] = zigzag(cw) + (int32_t)(acc >> shift);
                    }
                    if (!bad) {
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
                        i += 8;
                    } else {
                        br.A = sA; br.B = sB; br.N = sN; br.o = so; br.next = snext;
                        br.src.note_rollback();
                        force_single = 8;
                    }
                } else {
                    // ---- one row (head up to an 8-aligned index, partition tails, rolled-back bodies): checked reads
                    const uint32_t win = br.peek();
                    const uint32_t z = clz32(win);
                    const uint32_t n = z + kp1;
                    uint32_t cw;
                    if (n <= 32) {
                        cw = (z << k) | shr_c(shl_c(win, z + 1), rsh);
                        br.template skip<true>(n);
                    } else {
                        cw = rice_slow(br, k, len_bits, err);
                    }
                    ACC acc = 0;
#pragma unroll
                    for (int j = 0; j < P; ++j) acc += (ACC)c[j] * (ACC)h[P - 1 - j];
                    const int32_t v = zigzag(cw) + (int32_t)(acc >> shift);
#pragma unroll
                    for (int j = 0; j < P - 1; ++j) h[j] = h[j + 1];
                    h[P - 1] = v;
                    if (wr) out[i] = (int32_t)((uint32_t)v << wasted);
                    ++i;
                    if (force_single) --force_single;
                    if (err != ST_OK) return err;
                }
            }
        }
        if (br.pos_bits() > len_bits) return ST_END_OF_STREAM;
    }
    return ST_OK;
}

// Warm-up samples, (LPC) precision and shift, then dispatch on register-window length and
// accumulator width (the width depends on the precision, which sits after the warm-up samples).
template <class Src>
ZF_HDI uint32_t decode_predicted32(BitReader<Src> &br, int32_t *__restrict__ out, uint32_t order, bool is_lpc,
                                   uint32_t depth, uint32_t eff_bits, uint32_t block_size, uint32_t wasted,
                                   uint64_t len_bits, bool wr, bool inter64) {
    int32_t warm[32];
#pragma unroll 1
    for (uint32_t t = 0; t < order; ++t) {
        const int32_t v = br.read_signed(eff_bits);
        warm[t] = v;
        if (wr) out[t] = (int32_t)((uint32_t)v << wasted);
    }
    if (!is_lpc) {
        // |sum| <= 15 * 2^(depth-1) < 2^(depth+3)
        if (depth + 4 <= 32) return predicted_subframe<4, int32_t>(br, out, warm, order, false, 0, 0, block_size, wasted, len_bits, wr);
        return predicted_subframe<4, int64_t>(br, out, warm, order, false, 0, 0, block_size, wasted, len_bits, wr);
    }
    const uint32_t precision = br.read(4) + 1;  // :508 (0b1111 is not rejected by the reference)
    const uint32_t shift = br.read(5);          // :510 unsigned (Q17)
    // int32 sums are the reference's own arithmetic in a 16-bit container (InterType i32, :314-319) and cannot overflow
    // while the samples keep to `depth` bits. In a 32-bit container the reference sums in i64 and stays defined when a
    // malformed stream drives samples past `depth` bits (still inside i32): there a shifted sum is always taken in int64
    // (without a shift the low 32 bits are all that is kept, and those are exact either way).
    const bool wide = depth + precision + ceil_log2(order) > 32 || (inter64 && shift != 0);
#define ZF_RUN(PP)                                                                                                                  \
    return wide ? predicted_subframe<PP, int64_t>(br, out, warm, order, true, precision, shift, block_size, wasted, len_bits, wr) \
                : predicted_subframe<PP, int32_t>(br, out, warm, order, true, precision, shift, block_size, wasted, len_bits, wr)
    if (order <= 8) ZF_RUN(8);
    if (order <= 12) ZF_RUN(12);
    ZF_RUN(32);
#undef ZF_RUN
}

// Generic path for 32-bit-per-sample streams (side channel needs 33 bits): everything in int64,
// history in a small local ring. Rare; correctness over speed.
template <class Src>
ZF_HDI uint32_t decode_predicted64(BitReader<Src> &br, int64_t *__restrict__ out, uint32_t order, bool is_lpc,
                                   uint32_t eff_bits, uint32_t block_size, uint32_t wasted, uint64_t len_bits, bool wr) {
    int64_t hist[32];
    int32_t coef[32];
    for (uint32_t t = 0; t < order; ++t) {
        hist[t] = br.read_signed64(eff_bits);
        if (wr) out[t] = (int64_t)((uint64_t)hist[t] << wasted);
    }
    uint32_t shift = 0;
    if (is_lpc) {
        const uint32_t precision = br.read(4) + 1;
        shift = br.read(5);
        for (uint32_t t = 0; t < order; ++t) coef[t] = br.read_signed(precision);
    } else {
        for (uint32_t t = 0; t < 4; ++t) coef[t] = fixed_coef(order, (int)t);
    }
    const uint32_t method = br.read(2);
    if (method >= 2) return ST_INVALID_RESIDUAL_CODING_METHOD;
    const uint32_t pbits = 4 + method;
    const uint32_t po = br.read(4);
    const uint32_t psize = block_size >> po;
    if ((psize << po) != block_size || psize < order) return ST_UNDEFINED_IN_REFERENCE;
    uint32_t err = ST_OK;
    uint32_t i = order;
    for (uint32_t part = 0; part < (1u << po); ++part) {
        const uint32_t param = br.read(pbits);
        const uint32_t end = (part + 1) * psize;
        const bool escaped = param == (1u << pbits) - 1u;
        const uint32_t esc = escaped ? br.read(5) : 0;
        for (; i < end; ++i) {  // hist is a ring: sample i lives at hist[i & 31]
            int64_t r;
            if (escaped) {
                r = esc ? br.read_signed(esc) : 0;  // at most 31 bits (:646-654)
            } else {  // 64-bit codeword: see rice_slow64_outlined
                BitReader<Src> t = br;
                const uint64_t cw = rice_slow64_outlined(&t, param, len_bits, &err);
                br = t;
                r = (int64_t)(cw >> 1) ^ -(int64_t)(cw & 1u);
            }
            if (err != ST_OK) return err;
            int64_t acc = 0;
            for (uint32_t t = 0; t < order; ++t) acc += (int64_t)coef[t] * hist[(i - 1 - t) & 31];
            const int64_t v = r + (acc >> shift);
            hist[i & 31] = v;
            if (wr) out[i] = (int64_t)((uint64_t)v << wasted);
        }
        if (br.pos_bits() > len_bits) return ST_END_OF_STREAM;
    }
    return ST_OK;
}

template <typename ST> struct IsWide { static constexpr bool value = false; };
template <> struct IsWide<int64_t> { static constexpr bool value = true; };

// One whole frame: subframes in channel order, then align + CRC-16 (:425-551). `br` is positioned
// on the first subframe header. ST = int32_t: frames with <= 31 bits per sample (subframe depth
// <= 32); ST = int64_t: 32 bps streams. len_bits is relative to the source origin. inter64: the stream's
// container is wider than 16 bits (see "Arithmetic" above).
template <typename ST, class Src>
ZF_HDI void decode_frame(BitReader<Src> &br, const FrameRec &fr, ST *__restrict__ scratch, uint64_t len_bits, bool wr,
                         uint64_t origin_byte, FrameOut &o, bool inter64 = false) {
    uint32_t status = ST_OK;
    const uint32_t bs = fr.block_size;
    const uint32_t bps = fr.bps;
#pragma unroll 1
    for (uint32_t ch = 0; ch < fr.channels && status == ST_OK; ++ch) {
        const uint32_t hdr = br.read(8);  // zero:1 type:6 wasted_flag:1 (:426-430)
        if (hdr & 0x80u) {
            status = ST_INVALID_SUBFRAME_HEADER;
            break;
        }
        const uint32_t type = (hdr >> 1) & 0x3Fu;
        uint32_t wasted = 0;
        if (hdr & 1u) {
            const uint32_t u = br.read_unary(len_bits);
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
            if (IsWide<ST>::value) v = br.read_signed64(bps - wasted);
            else v = br.read_signed(bps - wasted);
            v = (int64_t)((uint64_t)v << wasted);
            if (wr)
                for (uint32_t i = 0; i < bs; ++i) out[i] = (ST)v;
        } else if (type == 1) {  // VERBATIM (:455-465)
            if (wasted >= depth) { status = ST_UNDEFINED_IN_REFERENCE; break; }
            const uint32_t n = depth - wasted;
#pragma unroll 1
            for (uint32_t i = 0; i < bs; ++i) {
                int64_t v;
                if (IsWide<ST>::value) v = br.read_signed64(n);
                else v = br.read_signed(n);
                if (wr) out[i] = (ST)(int64_t)((uint64_t)v << wasted);
                if ((i & 255u) == 255u && br.pos_bits() > len_bits) break;  // ran off the stream
            }
        } else if ((type >= 8 && type <= 12) || type >= 32) {  // FIXED (:466-498) / LPC (:499-541)
            const bool is_lpc = type >= 32;
            const uint32_t order = is_lpc ? type - 31 : type - 8;
            if (wasted >= depth || order > bs) { status = ST_UNDEFINED_IN_REFERENCE; break; }
            const uint32_t eff = depth - wasted;
            if (IsWide<ST>::value)
                status = decode_predicted64(br, (int64_t *)out, order, is_lpc, eff, bs, wasted, len_bits, wr);
            else
                status = decode_predicted32(br, (int32_t *)out, order, is_lpc, eff, eff, bs, wasted, len_bits, wr, inter64);
        } else {  // reserved (:542)
            status = ST_INVALID_SUBFRAME_HEADER;
        }
    }
    uint32_t stored_crc = 0;
    if (status == ST_OK) {
        br.align_byte();            // :546
        stored_crc = br.read(16);   // :548, not checked by the reference
    }
    const uint64_t pos = br.pos_bits();
    // Whatever went wrong, if the cursor is past the end the reference's reader failed first.
    if (pos > len_bits) status = ST_END_OF_STREAM;
    o.end_offset = origin_byte + (pos >> 3);
    o.status = status;
    o.stored_crc16 = stored_crc;
}

}  // namespace zf
