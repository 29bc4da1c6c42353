// parse_narrow.cuh — pass A of the two-pass subframe decoder for streams of <= 16 bits per sample ("narrow" K2a).
//
// Same job and same error sites as parse_core.cuh's parse_frame (reference src/zflac.zig:425-551, :614-666): one lane
// walks the serial bit chain of one frame and leaves, per subframe, a SubRec and the raw material of the predictor.
// What differs:
//   * the bit reader is PosReader (pos_reader.cuh): one register of state, I2FP instead of FLO, 13 instructions a row;
//   * the scratch holds int16 values — residuals, warm-up samples, VERBATIM samples — half the bytes pass A writes and
//     pass B (predict.cu, NARROW) reads back. A row of eight is one 16-byte store (layouts below).
//   * a value that does not fit 16 bits raises `overflow`: the host then decodes the launch group again on the
//     int32 path (parse_core.cuh). For audio of <= 16 bits that takes a residual beyond +-32767, i.e. never in practice.
#pragma once
#include "parse_core.cuh"
#include "pos_reader.cuh"

namespace zf {

ZF_HD uint32_t misfit16(int32_t v) { return v != (int32_t)(int16_t)v ? 1u : 0u; }
ZF_HD uint32_t pack16(int32_t lo, int32_t hi) {  // two int16 in one word (PRMT)
#ifdef __CUDA_ARCH__
    return __byte_perm((uint32_t)lo, (uint32_t)hi, 0x5410);
#else
    return ((uint32_t)lo & 0xFFFFu) | ((uint32_t)hi << 16);
#endif
}

// Narrow scratch layouts. A lane writes one 16-byte row (eight int16) per body, a pass-B lane reads one per step; what
// the layouts decide is how the rows of the 32 frames of a pass-A warp lie next to each other in memory, i.e. how many
// memory requests one store / load instruction of a warp becomes. Measured (tools/ubench/chain.cu, r2g): the L1 handles
// about one 32-byte-sector request per two cycles per SM, and a warp storing 32 rows into 32 frames' own regions costs
// 32 of them — the scattered rows, not the bit chain, were what bounded K2a.
//   kLayoutWarp  the 32 frames of the warp are INTERLEAVED row by row: row t of channel c of the frame in lane L at
//                warp_base + ((c * nb + t) * 32 + L) * 16 bytes, nb = ceil(longest block of the warp / 8). Lanes run their
//                bodies in step, so a store instruction writes 512 contiguous bytes; pass B's lanes (consecutive frames x
//                channels) read runs of 16-byte rows too. Needs a warp of 32 live frames with one channel count whose
//                regions together hold 32 columns of the longest frame (so: most frames of the stream's block size);
//                decided per warp, flagged in SubRec.pad[0] with nb in pad[1..2].
//   kLayoutFrame the frame's own region, groups of sixteen steps, channels one after another inside a group (one
//                32-byte sector per channel and group, written by the same lane in two halves):
//                frame_base + (t / 2) * 32 * C + c * 32 + (t % 2) * 16.
// Both live inside the region the int32 layout would use (scratch_base counts 4-byte elements): room to spare.
//
// WIDE = true: the same pass A for streams of 17..31 bits per sample. Rows are eight int32 = two 16-byte halves, no
// value can overflow, bodies are accepted for k <= 23 on the window fit alone. kLayoutFrame is then exactly the int32
// layout of parse_core.cuh (sub_idx(): 32 bytes per channel and row, channels one after another); kLayoutWarp keeps the
// two halves of a row 512 bytes apart so that each of a body's two store instructions writes 512 contiguous bytes:
// half h of row t of channel c of lane L at warp_base + (((c * nb + t) * 2 + h) * 32 + L) * 16.
enum : uint8_t { kLayoutFrame = 0, kLayoutWarp = 1 };

struct NarrowSink {
    uint8_t *base;       // row 0 of channel 0 of this lane's frame
    uint32_t ch_stride;  // bytes from a channel's row 0 to the next channel's
    uint32_t row_shift;  // row t lives at (t >> row_shift) * row_stride + (t & row_mask) * 16
    uint32_t row_mask;
    uint32_t row_stride;
    uint32_t half_off;   // WIDE: bytes from the first half of a row (values 0..3) to the second (values 4..7)
    uint16_t rows;       // kLayoutWarp: rows per channel of the warp's columns = ceil(longest block / 8), in SubRec.pad[1..2]
    uint8_t layout;
    ZF_HD void set_frame(void *scratch, uint64_t scratch_base, uint32_t channels, bool wide = false) {
        base = reinterpret_cast<uint8_t *>(scratch) + 4u * scratch_base;
        rows = 0;
        ch_stride = 32u;
        row_shift = wide ? 0u : 1u;
        row_mask = wide ? 0u : 1u;
        row_stride = 32u * channels;
        half_off = 16u;
        layout = kLayoutFrame;
    }
    ZF_HD void set_warp(void *scratch, uint64_t warp_scratch_base, uint32_t lane, uint32_t max_block_size, bool wide = false) {
        base = reinterpret_cast<uint8_t *>(scratch) + 4u * warp_scratch_base + 16u * lane;
        rows = (uint16_t)((max_block_size + 7u) >> 3);  // <= 8192
        row_stride = wide ? 1024u : 512u;
        ch_stride = rows * row_stride;
        row_shift = 0u;
        row_mask = 0u;
        half_off = 512u;
        layout = kLayoutWarp;
    }
};
struct NarrowOut {  // one channel's rows
    uint8_t *base;
    uint32_t row_shift, row_mask, row_stride, half_off;
    ZF_HD uint8_t *row(uint32_t i) const {  // the row that holds step i (WIDE: its first half)
        const uint32_t t = i >> 3;
        return base + (size_t)(t >> row_shift) * row_stride + (t & row_mask) * 16u;
    }
    template <bool WIDE>
    ZF_HD void put(uint32_t i, int32_t v) const {
        if (WIDE) reinterpret_cast<int32_t *>(row(i) + ((i >> 2) & 1u) * half_off)[i & 3u] = v;
        else reinterpret_cast<uint16_t *>(row(i))[i & 7u] = (uint16_t)v;
    }
    // four values, i a multiple of 4: half a row
    template <bool WIDE>
    ZF_HD void put4(uint32_t i, const int32_t (&r)[4]) const {
        uint8_t *p = row(i);
        if (WIDE) *reinterpret_cast<int4 *>(p + ((i >> 2) & 1u) * half_off) = make_int4(r[0], r[1], r[2], r[3]);
        else *reinterpret_cast<uint2 *>(p + ((i >> 2) & 1u) * 8u) = make_uint2(pack16(r[0], r[1]), pack16(r[2], r[3]));
    }
    // the eight values of a body, i a multiple of 8
    template <bool WIDE>
    ZF_HD void put8(uint32_t i, const int32_t (&r)[8]) const {
        uint8_t *p = row(i);
        if (WIDE) {
            *reinterpret_cast<int4 *>(p) = make_int4(r[0], r[1], r[2], r[3]);
            *reinterpret_cast<int4 *>(p + half_off) = make_int4(r[4], r[5], r[6], r[7]);
        } else {
            uint4 v;
            v.x = pack16(r[0], r[1]);
            v.y = pack16(r[2], r[3]);
            v.z = pack16(r[4], r[5]);
            v.w = pack16(r[6], r[7]);
            *reinterpret_cast<uint4 *>(p) = v;
        }
    }
};

template <class Src, bool WIDE>
ZF_HDI uint32_t parse_residual_narrow(PosReader<Src> &br, const NarrowOut &out, uint32_t order,
                                      uint32_t block_size, uint64_t len_bits, bool wr, uint32_t &overflow) {
    const uint32_t method = br.read(2);
    if (method >= 2) return ST_INVALID_RESIDUAL_CODING_METHOD;
    const uint32_t pbits = 4 + method;
    const uint32_t po = br.read(4);
    const uint32_t psize = block_size >> po;
    if ((psize << po) != block_size || psize < order) return ST_UNDEFINED_IN_REFERENCE;  // Q16
    uint32_t i = order;
    const uint32_t nparts = 1u << po;
    uint32_t part = 0, end = order;  // "partition -1" ends where the residual starts
    uint32_t k = 0;
    RiceParams rp;
    rp.set(0);
    bool body_ok = false;       // k <= 16: a body can be accepted at all
    uint32_t force_single = 0;  // rows that must take the one-row path (after a rejected body)
    uint32_t nbody = 0;         // whole 8-row bodies to run before anything else needs a look
    // The value extraction and the store of a body run one body late, inside the next body's chain: the chain is a string
    // of dependent instructions with a shared-memory load in every row, the extraction of the previous body's eight float
    // patterns is independent work the scheduler drops into its gaps. fp / ip: patterns and first row of the pending body.
    uint32_t fp[8];
    uint32_t ip = 0;
    bool pend = false;
#pragma unroll
    for (int u = 0; u < 8; ++u) fp[u] = 0;
    uint32_t j_stale = br.entry_index();  // ring entry of the cursor one body ago: the refill decision never waits for the newest position
    // ONE loop over partitions and rows (see parse_core.cuh): lanes of a warp run their bodies together whatever their
    // partition layouts are. The common iteration tests one counter; everything else (partition boundary, rows up to
    // an 8-aligned index, partition tails, rows after a rejected body) is a detour that ends by recounting it.
#pragma unroll 1
    for (;;) {
        if (nbody != 0) {
            if (Src::kStaleCursor) {
                br.src.top_up_stale(j_stale);
                j_stale = br.entry_index();
            } else {
                br.src.top_up(br.entry_index(), 0u);
            }
            uint32_t fn[8], qn;
            const bool ok = rice_chain8(br, rp, fn, qn);
            {   // the pending body (straight-line code: only the store is predicated)
                int32_t r[8];
#pragma unroll
                for (int u = 0; u < 8; ++u) r[u] = rice_value(fp[u], rp);
                if (pend && wr) out.template put8<WIDE>(ip, r);  // ip is a multiple of 8
            }
#pragma unroll
            for (int u = 0; u < 8; ++u) fp[u] = fn[u];
            pend = ok;
            ip = i;
            br.Q = ok ? qn : br.Q;
            i = ok ? i + 8 : i;
            nbody = ok ? nbody - 1 : 0u;
            force_single = ok ? force_single : 8u;
            continue;
        }
        if (pend) {  // leaving the run of bodies: finish the last one (k is still its partition's)
            if (wr) {
                int32_t r[8];
#pragma unroll
                for (int u = 0; u < 8; ++u) r[u] = rice_value(fp[u], rp);
                out.template put8<WIDE>(ip, r);
            }
            pend = false;
        }
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
                    if (!WIDE) overflow |= misfit16(r);
                    if (wr) out.template put<WIDE>(i, r);
                    if ((i & 255u) == 255u && br.pos_bits() > len_bits) return ST_END_OF_STREAM;
                }
                continue;
            }
            k = param;
            body_ok = k <= (WIDE ? 23u : 16u);
            if (body_ok) rp.set(k, WIDE);
        } else if (WIDE && body_ok && force_single == 0 && (i & 3u) == 0 && end - i >= 4u) {
            // half a body: the four rows between a partition boundary and the next multiple of eight (or the last four of
            // a partition) — Rice partitions of 36 or 12 samples would otherwise spend a third of their time on single rows
            // (C3, 36-sample partitions: 2.78 -> 2.29 ms). WIDE only: with this branch in them the int16 kernels, which
            // never take it on C2, ran 0.575 -> 0.827 ms there (r06n) — the compiler's layout of the hot loop, not the work.
            if (Src::kStaleCursor) br.src.top_up_stale(j_stale);
            else br.src.top_up(br.entry_index(), 0u);
            uint32_t f4[4], qn;
            const bool ok = rice_chain<4>(br, rp, f4, qn);
            if (ok) {
                int32_t r[4];
#pragma unroll
                for (int u = 0; u < 4; ++u) r[u] = rice_value(f4[u], rp);
                if (wr) out.template put4<WIDE>(i, r);
                br.Q = qn;
                i += 4;
            } else {
                force_single = 4u;
            }
        } else {  // one row
            const uint32_t win = br.peek();
            const uint32_t z = clz32(win);
            const uint32_t n = z + k + 1;
            uint32_t cw;
            if (n <= 32) {
                cw = (z << k) | shr_c(shl_c(win, z + 1), 32 - k);
                br.skip(n);
            } else {  // a codeword longer than the window: unbounded unary part, then k bits
                const uint32_t q = br.read_unary(len_bits);
                if (q == 0xFFFFFFFFu) return ST_END_OF_STREAM;
                const uint32_t r = br.read(k);
                cw = (q << k) + r;
                if (!WIDE && (q >> 16) != 0) overflow |= 1u;  // q << k wrapped or is far beyond 16 bits either way
            }
            if (!WIDE) overflow |= cw >> 16;
            if (wr) out.template put<WIDE>(i, zigzag(cw));
            ++i;
            if (force_single) --force_single;
        }
        nbody = (body_ok && force_single == 0 && (i & 7u) == 0) ? (end > i ? (end - i) >> 3 : 0u) : 0u;
        j_stale = br.entry_index();  // exact after a detour; inside a run of bodies it trails by one body
    }
    return ST_OK;
}

// One whole frame, subframes in channel order, then align + CRC-16 (:425-551). `sink` says where the frame's rows go.
template <class Src, bool WIDE = false>
ZF_HDI void parse_frame_narrow(PosReader<Src> &br, const FrameRec &fr, const NarrowSink &sink, SubRec *__restrict__ subs,
                               uint64_t len_bits, bool wr, uint64_t origin_byte, FrameOut &o, uint32_t &max_order, uint32_t &overflow) {
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
        NarrowOut out;
        out.base = sink.base + (size_t)ch * sink.ch_stride;
        out.row_shift = sink.row_shift;
        out.row_mask = sink.row_mask;
        out.row_stride = sink.row_stride;
        out.half_off = sink.half_off;
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
            uint32_t i = 0;
            // rows of eight: fixed-width fields have no chain — eight independent windows, one row store (a sample at a
            // time is ~50 dependent cycles each, and the lanes of a warp that hold Rice-coded subframes wait meanwhile)
#pragma unroll 1
            for (; i + 8u <= bs; i += 8u) {
                if (Src::kStaleCursor) br.src.top_up_stale(br.entry_index());
                else br.src.top_up(br.entry_index(), 0u);
                int32_t r[8];
                uint32_t bad = 0;
#pragma unroll
                for (int u = 0; u < 8; ++u) {
                    const uint32_t q = br.Q - (uint32_t)u * n;
                    const uint2 e = br.src.entry_at(q);
                    r[u] = (int32_t)funnel_r_wrap(e.y, e.x, q) >> (32u - n);
                    if (!WIDE) bad |= misfit16(r[u]);
                }
                br.src.body_done(true);
                overflow |= bad;
                if (wr) out.template put8<WIDE>(i, r);
                br.Q -= 8u * n;
                if ((i & 255u) == 248u && br.pos_bits() > len_bits) break;  // ran off the stream
            }
#pragma unroll 1
            for (; i < bs; ++i) {
                const int32_t v = br.read_signed(n);
                if (!WIDE) overflow |= misfit16(v);
                if (wr) out.template put<WIDE>(i, v);
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
                if (!WIDE) overflow |= misfit16(v);
                if (wr) out.template put<WIDE>(t, v);
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
            status = parse_residual_narrow<Src, WIDE>(br, out, order, bs, len_bits, wr, overflow);
        } else {  // reserved (:542)
            status = ST_INVALID_SUBFRAME_HEADER;
        }
        if (wr) {
            sub->kind = (uint8_t)kind;
            sub->order = (uint8_t)order;
            sub->shift = (uint8_t)shift;
            sub->wasted = (uint8_t)wasted;
            sub->wide_acc = (uint8_t)wide_acc;
            sub->pad[0] = sink.layout;
            sub->pad[1] = (uint8_t)(sink.rows & 0xFFu);
            sub->pad[2] = (uint8_t)(sink.rows >> 8);
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
