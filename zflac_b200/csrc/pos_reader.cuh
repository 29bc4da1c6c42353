// pos_reader.cuh — bit reader of the "narrow" pass A (parse_narrow.cuh): the serial bit chain of a frame carried as ONE
// register, the bit position, with the 32-bit window re-read from shared memory for every codeword.
//
// Why. The first pass A (decode_core.cuh's BitReader) keeps a word pipeline A:B:N in registers and moves it under
// predicates; measured on a B200 that costs 26-28 instructions and ~150 cycles per codeword and warp, because almost
// every instruction of a row depends on the one before it (in-order issue) and the unary count is a FLO — 19 cycles of
// latency on the XU pipe (tools/ubench/lat.cu: FLO 19, LOP3/SHF/IMAD 4.3, LDS 19-21 cycles, ALU pipe 2 cycles per warp
// instruction). Here a row is
//
//     address = f(Q)            IMAD + LOP3                     Q = -(bit position), the only state of the chain
//     {hi, lo} = ring[address]  one LDS.64: ring entry j holds stream words (j - 1, j), already MSB-first
//     win = (hi:lo) >> (Q & 31) SHF.R.W — a right funnel by -P is a left funnel by P
//     f   = float(win), RZ      I2FP on the ALU pipe (4 cycles): exponent = 158 - leading zeros, mantissa = the bits
//                               after the stop bit — unary count AND remainder from one instruction, no FLO
//     Q'  = (Q - C) + (f >> 23) LEA.HI, C = 158 + k + 1: the position of the next codeword
//
// six dependent instructions (~44 cycles, half of it the shared load) and seven more off the chain for the codeword
// value (x = f >> (23 - k) = e * 2^k + r, cw = 158 * 2^k + x - 2 * (x & ~(2^k - 1))), 13 in all. No word pipeline, no
// predicated moves, a rolled-back body restores one register.
//
// Ring entries are PAIRS so that one aligned 8-byte load yields the 64 stream bits around any position: entry j =
// (word j - 1, word j), at ring[(j % RING)][lane] — a lane only ever touches its own 8-byte column, so the loads are
// conflict free whatever the cursors are. The ring is filled through registers sixteen words at a time like StagedSrc
// (ring_src.cuh); the byte swap to MSB-first happens once per word at fill time.
//
// Host-testable (tests/host_core/): HostPairSrc there implements the same concept over plain memory.
#pragma once
#include <cstring>

#include "decode_core.cuh"

namespace zf {

// low 32 bits of (hi:lo) >> (s & 31)
ZF_HD uint32_t funnel_r_wrap(uint32_t lo, uint32_t hi, uint32_t s) {
#ifdef __CUDA_ARCH__
    return __funnelshift_r(lo, hi, s);
#else
    s &= 31;
    return s ? (hi << (32 - s)) | (lo >> s) : lo;
#endif
}
// bit pattern of (float)v rounded toward zero: exponent field = 127 + floor(log2 v) (0 for v == 0), mantissa = the 23
// bits after the leading one
ZF_HD uint32_t float_bits_rz(uint32_t v) {
#ifdef __CUDA_ARCH__
    return __float_as_uint(__uint2float_rz(v));
#else
    if (v == 0) return 0;
    const uint32_t lz = (uint32_t)__builtin_clz(v);
    const uint32_t norm = v << lz;  // leading one at bit 31
    return ((158u - lz) << 23) | ((norm >> 8) & 0x7FFFFFu);
#endif
}

// ---------------------------------------------------------------------------------------
// PairSrc concept (entry index j counts 32-bit stream words from the source origin; entry j = {hi: word j-1, lo: word j},
// both MSB-first; words at or past the end of the stream read as zero):
//   uint2 entry(j)            entry j, the caller guarantees it is resident
//   uint2 entry_at(Q)         entry (31 - Q) >> 5 = the entry that holds the 32 bits at position P = -Q; resident
//   void  ensure(j)           make entry j resident (slow paths); entries below j - 1 may be dropped
//   void  top_up(j0, tick)    entries < j0 are consumed; refill, and guarantee [j0, j0 + kFastEntries) resident
//   void  body_done(ok)       a speculative body ended (test hook; nothing on the device)
// ---------------------------------------------------------------------------------------
constexpr uint32_t kFastEntries = 9;  // an 8-row body starts in entry j0 and moves at most 8 words on

template <class Src>
struct PosReader {
    Src src;
    uint32_t Q;  // -(bit position from word 0 of the source), mod 2^32

    ZF_HD void init_at(uint32_t skip_bytes) { Q = 0u - 8u * skip_bytes; }
    ZF_HD uint32_t pos() const { return 0u - Q; }
    ZF_HD uint64_t pos_bits() const { return (uint64_t)(0u - Q); }
    ZF_HD uint32_t entry_index() const { return (31u - Q) >> 5; }
    ZF_HD uint32_t peek() {
        const uint32_t j = entry_index();
        src.ensure(j);
        const uint2 e = src.entry(j);
        return funnel_r_wrap(e.y, e.x, Q);
    }
    ZF_HD void skip(uint32_t n) { Q -= n; }
    ZF_HD uint32_t read(uint32_t n) {  // n in [0, 32]
        const uint32_t v = shr_c(peek(), 32 - n);
        Q -= n;
        return v;
    }
    ZF_HD int32_t read_signed(uint32_t n) {  // n in [1, 32]
        const uint32_t v = peek();
        Q -= n;
        return (int32_t)v >> (32 - n);
    }
    // zeros before the next 1 bit; consumes the 1. 0xFFFFFFFF if the stream ran out first (or the run is absurd:
    // 2^30 zeros, where the reference's own arithmetic is long past its defined range)
    ZF_HDI uint32_t read_unary(uint64_t len_bits) {
        uint32_t q = 0;
        for (;;) {
            const uint32_t w = peek();
            if (w != 0) {
                const uint32_t z = clz32(w);
                Q -= z + 1;
                return q + z;
            }
            q += 32;
            Q -= 32;
            if (pos_bits() > len_bits || q >= (1u << 30)) return 0xFFFFFFFFu;
        }
    }
    ZF_HD void align_byte() { Q &= ~7u; }  // -P rounded down to a multiple of 8 = P rounded up
};

// ---------------------------------------------------------------------------------------
// Eight Rice codewords with parameter k <= 16, speculative and branch free; a body is rejected when a codeword did not
// fit the 32-bit window or might not fit 16 bits, and the caller then redoes its rows one by one.
//
// What was measured on the way here (B200, tools/ubench/chain*.cu, cycles per row of one warp, chain only unless said):
//   this chain 54.5 (31 of them address + LDS.64); FLO instead of I2FP 75; exponent by LOP3 + FADD (2^23 magic) 60.5;
//   the same chain with the ring refilled by one cp.async.bulk per lane 89-113 (an elect loop of 9 instructions per
//   lane and TMA's per-request cost); value extraction in a second phase through a shared-memory tile +25 (+37 with two
//   warps per scheduler): with ~21 instructions a row, most on the half-rate ALU pipe, two warps per scheduler are
//   bound by instruction throughput, not by the chain's latency — so the work below is counted in instructions.
//   Moving shifts to the FMA pipe as IMAD.HI (f >> 23 and f >> sh as the high word of a product with an opaque power of
//   two: 113 -> 97 ALU-pipe instructions a body) made the kernel SLOWER, 0.575 -> 0.591 ms on C2, with either one of the
//   two alone as well (r06j): IMAD.HI is no half-rate instruction here.
// ---------------------------------------------------------------------------------------
struct RiceParams {
    uint32_t C;      // 158 + k + 1: (Q - C) + exponent = Q - (zeros + 1 + k)
    uint32_t sh;     // 23 - k
    uint32_t nmask;  // ~(2^k - 1)
    uint32_t C2;     // 158 << k
    uint32_t fmin;   // smallest float pattern a body accepts: window fit (zeros <= 31 - k) and cw < 65536 guaranteed
    // k <= 16 for int16 rows; k <= 23 for int32 rows (`wide`: the mantissa of a float holds 23 bits after the stop bit, and
    // a body is accepted on the window fit alone)
    ZF_HD void set(uint32_t k, bool wide = false) {
        C = 159u + k;
        sh = 23u - k;
        nmask = ~((1u << k) - 1u);
        C2 = 158u << k;
        // window: zeros <= 31 - k.  16 bits: (zeros + 1) << k <= 65536  <=>  zeros <= (65536 >> k) - 1
        const uint32_t zfit = 31u - k;
        const uint32_t z16 = wide ? zfit : (65536u >> k) - 1u;
        const uint32_t zlim = zfit < z16 ? zfit : z16;
        fmin = (158u - zlim) << 23;
    }
};

// The chain of eight codewords: the float pattern of every window (unary count and remainder bits in one word, see the
// header) into f[], the position after the eighth into q_out. Returns whether every codeword fitted the window and is
// known to fit 16 bits; br itself is not moved. Nothing here depends on anything but Q: the caller overlaps the
// value extraction of the PREVIOUS body with this chain.
template <int N, class Src>
ZF_HDI bool rice_chain(PosReader<Src> &br, const RiceParams &rp, uint32_t (&f)[N], uint32_t &q_out) {
    uint32_t Q = br.Q;
    uint32_t fmin = 0xFFFFFFFFu;
#pragma unroll
    for (int u = 0; u < N; ++u) {
        const uint2 e = br.src.entry_at(Q);
        const uint32_t win = funnel_r_wrap(e.y, e.x, Q);
        f[u] = float_bits_rz(win);
        const uint32_t qmc = Q - rp.C;
        Q = qmc + (f[u] >> 23);  // LEA.HI
        fmin = fmin < f[u] ? fmin : f[u];
    }
    // a rejected body may have read entries far beyond the guarantee (a window without a stop bit moves the cursor by up
    // to 175 bits): harmless on the device — ring addresses wrap inside the ring and the values are thrown away — and
    // told to the source so that the host stand-in only counts unchecked reads of ACCEPTED bodies
    const bool ok = fmin >= rp.fmin;
    br.src.body_done(ok);
    q_out = Q;
    return ok;
}

template <class Src>
ZF_HDI bool rice_chain8(PosReader<Src> &br, const RiceParams &rp, uint32_t (&f)[8], uint32_t &q_out) {
    return rice_chain<8>(br, rp, f, q_out);
}

// The residual a float pattern stands for: x = f >> (23 - k) = exponent * 2^k + r, codeword (158 - exponent) * 2^k + r,
// unfolded ((cw >> 1) ^ -(cw & 1), src/zflac.zig:659-663). (Written as h - (cw & 1) * cw to get an IMAD, the compiler
// makes a compare and a select of it: slower, r2l.)
ZF_HD int32_t rice_value(uint32_t f, const RiceParams &rp) {
    const uint32_t x = f >> rp.sh;
    const uint32_t y = x & rp.nmask;
    return zigzag((x + rp.C2) - 2u * y);
}

}  // namespace zf
