// parse_core.cuh — pass A of the two-pass subframe decoder (K2a): the serial part only.
//
// A FLAC frame is one serial bit chain: subframe c+1 starts where subframe c ends and a Rice
// codeword's length depends on its value (reference: src/zflac.zig:425-551, :614-666 walks it with
// one byte reader). What is NOT tied to that chain is the predictor recurrence (:466-541,
// :604-612): once the residuals, warm-up samples and coefficients of a subframe are known, its
// samples depend on nothing else. So pass A (this file) walks the chain with one lane per FRAME
// and writes, per subframe, the raw material — warm-up samples and residuals into the planar
// scratch, order/shift/coefficients/wasted bits into a SubRec — and pass B (predict.cu) runs the
// recurrence with one lane per SUBFRAME (channels x more chains, 32 IMADs per sample off the
// bit-chain's critical path) fused with the decorrelation/interleave epilogue.
//
// Same word-source concept and bit reader as decode_core.cuh (host-testable: tests/host_core/).
// 32 bps streams (33-bit side channel) stay on the single-pass int64 decoder of decode_core.cuh.
#pragma once
#include "decode_core.cuh"

namespace zf {

// What pass B needs to rebuild one subframe. Indexed [frame slot * 8 + channel].
struct SubRec {
    uint8_t kind;      // kSubRaw: scratch holds the samples (VERBATIM); kSubConst: `constant`; kSubPred: warm-up + residuals
    uint8_t order;     // predictor order, 0..32 (0 for raw / constant)
    uint8_t shift;     // LPC quantisation shift (:510); 0 for FIXED
    uint8_t wasted;    // wasted bits: every sample is shifted left by this much on output (:433, :494, :537)
    uint8_t wide_acc;  // 1: depth + precision + log2(order) > 32, the predictor sum needs int64
    uint8_t pad[3];
    int32_t constant;  // kind == kSubConst: the (unshifted) value
    int32_t pad2;
    int16_t coef[32];  // coef[j] multiplies s[i-1-j] (:513-514 stores them reversed; FIXED: :481-490)
};
static_assert(sizeof(SubRec) == 80, "SubRec layout");
enum : uint8_t { kSubRaw = 0, kSubConst = 1, kSubPred = 2 };

// Scratch layout of the two-pass path: per frame, groups of 8 time steps; inside a group the
// channels one after another, 8 consecutive values each:
//     element(frame, ch, i) = scratch_base + (i / 8) * 8 * C + ch * 8 + (i % 8)
// Pass A writes 32 bytes (one sector) per lane and 8-row body either way; in pass B the C lanes of a
// frame then read C adjacent sectors per step instead of C far-apart ones. Same size as C planes of
// pitch roundup8(block_size).
ZF_HD uint32_t sub_idx(uint32_t i, uint32_t c8) { return (i >> 3) * c8 + (i & 7u); }

// Residual of a FIXED / LPC subframe (:614-666) into out[order .. block_size): the Rice loop of
// decode_core.cuh without the predictor.
template <class Src>
ZF_HDI uint32_t parse_residual(BitReader<Src> &br, int32_t *__restrict__ out, uint32_t c8, uint32_t order, uint32_t block_size,
                               uint64_t len_bits, bool wr) {
    const uint32_t method = br.read(2);
    if (method >= 2) return ST_INVALID_RESIDUAL_CODING_METHOD;
    const uint32_t pbits = 4 + method;
    const uint32_t po = br.read(4);
    const uint32_t psize = block_size >> po;
    if ((psize << po) != block_size || psize < order) return ST_UNDEFINED_IN_REFERENCE;  // Q16
    uint32_t err = ST_OK;
    uint32_t i = order;
    // ONE loop over partitions and rows. With a partition loop around a row loop, the lanes of a warp meet again at
    // the end of every partition: a lane with 16 partitions of 256 next to a lane with one partition of 4096 would
    // idle through 4096 rows and then run its other 15 partitions alone (1.9x the work of either). Here a partition
    // boundary is a short detour inside the row loop and the 8-row bodies of all lanes run together whatever their
    // partition layout is.
    const uint32_t nparts = 1u << po;
    uint32_t part = 0, end = order;  // "partition -1" ends where the residual starts
    uint32_t k = 0, kp1 = 1, pk = 1;
    uint32_t force_single = 0;
    uint32_t tick = 0;  // body count: warp-uniform, so the word source's every-other-body refill stays in step
#pragma unroll 1
    for (;;) {
        if (i >= end) {  // partition boundary
            if (part && br.pos_bits() > len_bits) return ST_END_OF_STREAM;
            if (part == nparts) break;
            const uint32_t param = br.read(pbits);
            end = (part + 1) * psize;
            ++part;
            if (param == (1u << pbits) - 1u) {  // escape: raw signed residuals of esc bits (:646-654)
                const uint32_t esc = br.read(5);
#pragma unroll 1
                for (; i < end; ++i) {
                    const int32_t r = esc ? br.read_signed(esc) : 0;
                    if (wr) out[sub_idx(i, c8)] = r;
                    if ((i & 255u) == 255u && br.pos_bits() > len_bits) return ST_END_OF_STREAM;
                }
                continue;
            }
            k = param;
            kp1 = param + 1;
            pk = 1u << param;
#ifdef __CUDA_ARCH__
            asm("" : "+r"(pk));  // keep z * pk an IMAD (FMA pipe): the ALU pipe carries the bit chain
#endif
            if (i >= end) continue;  // a first partition that holds warm-up samples only
        }
        if (force_single == 0 && (i & 7u) == 0 && i + 8 <= end) {
            // 8 codewords, speculative and branch-free; a codeword longer than the 32-bit window
            // (rare) only raises zmax: the body is rolled back and redone row by row
            br.prepare_fast(tick++);
            const uint32_t sA = br.A, sB = br.B, sN = br.N, so = br.o;
            uint32_t ws = br.enter_fast();
            uint32_t zmax = 0;
            int32_t r[8];
#pragma unroll
            for (int u = 0; u < 8; ++u) {
                const uint32_t win = br.peek();
                const uint32_t z = clz_sh(win);                                   // unary part (0xFFFFFFFF: no stop bit in sight)
                zmax = zmax > z ? zmax : z;
                const uint32_t low = funnel_hi(0u, shl_c(win, z + 1), k);        // the k bits after the stop bit
                const uint32_t cw = z * pk + low;                                 // (z << k) | low on the IMAD pipe
                br.skip_fast(z + kp1, ws);
                r[u] = zigzag(cw);
            }
            if (zmax <= 31u - k) {  // every codeword fitted the 32-bit window
                if (wr) {
                    int32_t *dst = out + (size_t)(i >> 3) * c8;  // i is a multiple of 8
                    *reinterpret_cast<int4 *>(dst) = make_int4(r[0], r[1], r[2], r[3]);
                    *reinterpret_cast<int4 *>(dst + 4) = make_int4(r[4], r[5], r[6], r[7]);
                }
                br.leave_fast(ws);
                i += 8;
            } else {
                br.A = sA; br.B = sB; br.N = sN; br.o = so;
                br.src.note_rollback();
                force_single = 8;
            }
        } else {
            const uint32_t win = br.peek();
            const uint32_t z = clz32(win);
            const uint32_t n = z + kp1;
            uint32_t cw;
            if (n <= 32) {
                cw = (z << k) | shr_c(shl_c(win, z + 1), 32 - k);
                br.template skip<true>(n);
            } else {
                cw = rice_slow(br, k, len_bits, err);
            }
            if (wr) out[sub_idx(i, c8)] = zigzag(cw);
            ++i;
            if (force_single) --force_single;
            if (err != ST_OK) return err;
        }
    }
    return ST_OK;
}

// One whole frame, subframes in channel order, then align + CRC-16 (:425-551). Scratch planes are
// int32 (frames with <= 31 bits per sample: subframe depth <= 32). `subs` points at the frame's 8
// SubRec slots. Error sites and precedence are those of decode_core.cuh's decode_frame.
template <class Src>
ZF_HDI void parse_frame(BitReader<Src> &br, const FrameRec &fr, int32_t *__restrict__ scratch, SubRec *__restrict__ subs,
                        uint64_t len_bits, bool wr, uint64_t origin_byte, FrameOut &o, uint32_t &max_order) {
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
        int32_t *out = scratch + fr.scratch_base + ch * 8u;  // sub_idx() layout
        const uint32_t c8 = 8u * fr.channels;
        SubRec *sub = subs + ch;
        uint32_t kind = kSubRaw, order = 0, shift = 0, wide_acc = 0;
        int32_t constant = 0;

        if (type == 0) {  // CONSTANT (:445-454): reads bps - wasted bits even on a side channel (Q8)
            if (wasted >= bps) { status = ST_UNDEFINED_IN_REFERENCE; break; }
            constant = br.read_signed(bps - wasted);
            kind = kSubConst;
        } else if (type == 1) {  // VERBATIM (:455-465)
            if (wasted >= depth) { status = ST_UNDEFINED_IN_REFERENCE; break; }
            const uint32_t n = depth - wasted;
#pragma unroll 1
            for (uint32_t i = 0; i < bs; ++i) {
                const int32_t v = br.read_signed(n);
                if (wr) out[sub_idx(i, c8)] = v;
                if ((i & 255u) == 255u && br.pos_bits() > len_bits) break;  // ran off the stream
            }
        } else if ((type >= 8 && type <= 12) || type >= 32) {  // FIXED (:466-498) / LPC (:499-541)
            const bool is_lpc = type >= 32;
            order = is_lpc ? type - 31 : type - 8;
            if (wasted >= depth || order > bs) { status = ST_UNDEFINED_IN_REFERENCE; break; }
            const uint32_t eff = depth - wasted;
            kind = kSubPred;
            max_order = max_order > order ? max_order : order;
#pragma unroll 1
            for (uint32_t t = 0; t < order; ++t) {  // warm-up samples (:473-479, :501-507)
                const int32_t v = br.read_signed(eff);
                if (wr) out[sub_idx(t, c8)] = v;
            }
            if (is_lpc) {
                const uint32_t precision = br.read(4) + 1;  // :508 (0b1111 is not rejected by the reference)
                shift = br.read(5);                         // :510 unsigned (Q17)
                wide_acc = eff + precision + ceil_log2(order) > 32 ? 1u : 0u;
#pragma unroll 1
                for (uint32_t t = 0; t < order; ++t) {      // t-th coefficient multiplies s[i-1-t] (:513-514)
                    const int32_t v = br.read_signed(precision);
                    if (wr) sub->coef[t] = (int16_t)v;
                }
            } else {
                wide_acc = eff + 4 > 32 ? 1u : 0u;  // |sum| <= 15 * 2^(depth-1)
                if (wr) {
#pragma unroll
                    for (int j = 0; j < 4; ++j) sub->coef[j] = (int16_t)fixed_coef(order, j);
                }
            }
            status = parse_residual(br, out, c8, order, bs, len_bits, wr);
        } else {  // reserved (:542)
            status = ST_INVALID_SUBFRAME_HEADER;
        }
        if (wr) {
            sub->kind = (uint8_t)kind;
            sub->order = (uint8_t)order;
            sub->shift = (uint8_t)shift;
            sub->wasted = (uint8_t)wasted;
            sub->wide_acc = (uint8_t)wide_acc;
            sub->pad[0] = 0;  // kLayoutFrame (parse_narrow.cuh): pass B finds the rows in the frame's own region, sub_idx()
            sub->pad[1] = 0;
            sub->pad[2] = 0;
            sub->constant = constant;
        }
    }
    uint32_t stored_crc = 0;
    if (status == ST_OK) {
        br.align_byte();            // :546
        stored_crc = br.read(16);   // :548, not checked by the reference
    }
    const uint64_t pos = br.pos_bits();
    if (pos > len_bits) status = ST_END_OF_STREAM;  // whatever went wrong, the reference's reader failed first
    o.end_offset = origin_byte + (pos >> 3);
    o.status = status;
    o.stored_crc16 = stored_crc;
}

}  // namespace zf
