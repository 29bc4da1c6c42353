// host_core.cu — TEST INFRASTRUCTURE. Instantiates the K2 per-frame decoder
// (zflac_b200/csrc/decode_core.cuh) on the HOST with a plain-memory word source so that the
// bit-level logic (funnel reader, Rice fast/slow paths, partition walk, predictor windows) can be
// unit-tested against the oracle without a GPU. Built into tests/host_core/_build/ only; it is
// never linked into libzflacb200.so and the product package never loads it.
#include <cstring>

#include "../../zflac_b200/csrc/decode_core.cuh"
#include "../../zflac_b200/csrc/parse_core.cuh"
#include "../../zflac_b200/csrc/parse_narrow.cuh"

namespace {

struct HostSrc {
    const uint8_t *g;
    uint64_t avail;
    // poison tracking: the device ring only guarantees kFastWords resident words after top_up();
    // mimic that so an unchecked read beyond the guarantee is caught here, not on the GPU
    uint32_t guaranteed;
    uint32_t consumed_below;  // words below this index may already be overwritten in the device ring
    int violations;
    int rollbacks;  // fast bodies that met a codeword longer than the 32-bit window
    __host__ __device__ uint32_t raw(uint32_t w) const {
        uint32_t v = 0;
        for (int i = 0; i < 4; ++i) {
            const uint64_t off = (uint64_t)w * 4 + i;
            v = (v << 8) | (off < avail ? g[off] : 0u);
        }
        return __builtin_bswap32(v);
    }
    __host__ __device__ uint32_t word(uint32_t w) {
        if (w >= guaranteed || w < consumed_below) ++violations;
        return raw(w);
    }
    static constexpr uint32_t kWordScale = 512;  // like the device ring of a 128-thread CTA
    __host__ __device__ uint32_t word_scaled(uint32_t ws) { return word(ws / kWordScale); }
    __host__ __device__ uint32_t word_checked(uint32_t w) {
        if (w > consumed_below) consumed_below = w;  // the device refills with `words < w are consumed`
        return raw(w);
    }
    __host__ __device__ void top_up(uint32_t cur, uint32_t) { top_up(cur); }
    __host__ __device__ void top_up(uint32_t cur) {
        guaranteed = cur + zf::kFastWords;
        if (cur > consumed_below) consumed_below = cur;
    }
    // raw order here = byte-reversed, like the little-endian shared-memory word of the device ring
    __host__ __device__ void note_rollback() { ++rollbacks; }
    static __host__ __device__ uint32_t fix(uint32_t raw) { return __builtin_bswap32(raw); }
};

// Host stand-in for PairSrc (ring_src.cuh): entry j = {word j - 1, word j}, MSB-first, over plain memory, with the
// device ring's residency rules mimicked so that an unchecked entry read outside the guarantee is caught here.
struct HostPairSrc {
    const uint8_t *g;
    uint64_t avail;
    uint32_t guaranteed_lo, guaranteed_hi;  // entries [lo, hi) may be read unchecked (set by top_up)
    uint32_t ensured;                       // the one entry ensure() made resident
    int violations;
    uint32_t word_be(uint32_t w) const {
        uint32_t v = 0;
        for (int i = 0; i < 4; ++i) {
            const uint64_t off = (uint64_t)w * 4 + i;
            v = (v << 8) | (off < avail ? g[off] : 0u);
        }
        return v;
    }
    uint2 make(uint32_t j) const { return make_uint2(j ? word_be(j - 1) : 0u, word_be(j)); }
    uint2 entry(uint32_t j) {
        if (j != ensured && !(j >= guaranteed_lo && j < guaranteed_hi)) ++violations;
        return make(j);
    }
    int speculative = 0;  // unchecked reads outside the guarantee inside the current body: they count if it is accepted
    uint2 entry_at(uint32_t Q) {
        const uint32_t j = (31u - Q) >> 5;
        if (!(j >= guaranteed_lo && j < guaranteed_hi)) ++speculative;
        return make(j);
    }
    static constexpr bool kStaleCursor = true;
    void body_done(bool ok) {
        if (ok) violations += speculative;
        speculative = 0;
    }
    void ensure(uint32_t j) { ensured = j; guaranteed_lo = guaranteed_hi = 0; }
    void top_up(uint32_t j0, uint32_t) { guaranteed_lo = j0; guaranteed_hi = j0 + zf::kFastEntries; ensured = 0xFFFFFFFFu; }
    // cursor of one body ago: the device keeps [j_old, j_old + 8 + kFastEntries) resident
    // (and the entry below j_old: a body reads its windows nine bits early)
    void top_up_stale(uint32_t j_old) { guaranteed_lo = j_old ? j_old - 1 : 0; guaranteed_hi = j_old + 8 + zf::kFastEntries; ensured = 0xFFFFFFFFu; }
};

}  // namespace

static long long *g_rollbacks = nullptr;

extern "C" {

// Decodes one frame on the host. scratch: int32 (wide == 0) or int64 planes; inter64: the stream's container is wider than
// 16 bits (decode_core.cuh, "Arithmetic"). Returns the number of unchecked-read violations (must be 0).
int zf_host_decode_frame(const uint8_t *stream, const zf::FrameRec *fr, void *scratch, int wide, int write_samples,
                         zf::FrameOut *out, int inter64) {
    const uint64_t start = fr->offset + fr->header_len;
    const uint64_t origin = start & ~3ull;
    const uint64_t bytes = fr->stream_end > origin ? fr->stream_end - origin : 0;
    zf::BitReader<HostSrc> br;
    br.src.g = stream + origin;
    br.src.avail = bytes;
    br.src.guaranteed = 0;
    br.src.consumed_below = 0;
    br.src.violations = 0;
    br.src.rollbacks = 0;
    br.init((uint32_t)(start & 3));
    if (wide) zf::decode_frame<int64_t>(br, *fr, (int64_t *)scratch, bytes * 8, write_samples != 0, origin, *out);
    else zf::decode_frame<int32_t>(br, *fr, (int32_t *)scratch, bytes * 8, write_samples != 0, origin, *out, inter64 != 0);
    if (g_rollbacks) *g_rollbacks += br.src.rollbacks;
    return br.src.violations;
}

// Pass A of the two-pass decoder on the host: residuals / warm-up / verbatim samples into the int32
// scratch, one SubRec per channel into subs[0..8). Returns the unchecked-read violations (must be 0).
int zf_host_parse_frame(const uint8_t *stream, const zf::FrameRec *fr, int32_t *scratch, zf::SubRec *subs, int write_samples,
                        zf::FrameOut *out) {
    const uint64_t start = fr->offset + fr->header_len;
    const uint64_t origin = start & ~3ull;
    const uint64_t bytes = fr->stream_end > origin ? fr->stream_end - origin : 0;
    zf::BitReader<HostSrc> br;
    br.src.g = stream + origin;
    br.src.avail = bytes;
    br.src.guaranteed = 0;
    br.src.consumed_below = 0;
    br.src.violations = 0;
    br.src.rollbacks = 0;
    br.init((uint32_t)(start & 3));
    uint32_t max_order = 0;
    zf::parse_frame(br, *fr, scratch, subs, bytes * 8, write_samples != 0, origin, *out, max_order);
    if (g_rollbacks) *g_rollbacks += br.src.rollbacks;
    return br.src.violations;
}

// Narrow pass A (parse_narrow.cuh) on the host: int16 rows into `scratch`. layout 0: the frame's own region
// (kLayoutFrame, scratch_base 0); layout 1: the warp-interleaved one (kLayoutWarp) as lane `lane` of a warp whose region
// starts at scratch. Returns the unchecked-read violations (must be 0); *overflow != 0 when a value did not fit 16 bits.
// wide != 0: int32 rows (parse_narrow.cuh, WIDE) — no overflow is ever reported.
int zf_host_parse_frame_narrow(const uint8_t *stream, const zf::FrameRec *fr, uint16_t *scratch, zf::SubRec *subs, int write_samples,
                               zf::FrameOut *out, uint32_t *overflow, int layout, int lane, int wide) {
    const uint64_t start = fr->offset + fr->header_len;
    const uint64_t origin = start & ~15ull;  // the device source starts at a 16-byte aligned address
    const uint64_t bytes = fr->stream_end > origin ? fr->stream_end - origin : 0;
    zf::PosReader<HostPairSrc> br;
    br.src.g = stream + origin;
    br.src.avail = bytes;
    br.src.guaranteed_lo = br.src.guaranteed_hi = 0;
    br.src.ensured = 0xFFFFFFFFu;
    br.src.violations = 0;
    br.init_at((uint32_t)(start - origin));
    zf::NarrowSink sink;
    if (layout == zf::kLayoutWarp) sink.set_warp(scratch, 0, (uint32_t)lane, fr->block_size, wide != 0);
    else sink.set_frame(scratch, 0, fr->channels, wide != 0);
    uint32_t max_order = 0;
    *overflow = 0;
    if (wide) zf::parse_frame_narrow<HostPairSrc, true>(br, *fr, sink, subs, bytes * 8, write_samples != 0, origin, *out, max_order, *overflow);
    else zf::parse_frame_narrow<HostPairSrc, false>(br, *fr, sink, subs, bytes * 8, write_samples != 0, origin, *out, max_order, *overflow);
    return br.src.violations;
}

// Scalar restatement of what pass B (predict.cu) does to one scratch plane: predictor recurrence in
// place, then the wasted-bits shift. Test-side stand-in for the CUDA kernel (the -m gpu tests pin the kernel).
// `frame` is the frame's scratch in the sub_idx() layout of parse_core.cuh; the result goes to plane[0..block_size).
void zf_host_predict_plane(const zf::SubRec *sub, const int32_t *frame, uint32_t ch, uint32_t channels, uint32_t block_size,
                           int32_t *plane) {
    const uint32_t c8 = 8 * channels;
    for (uint32_t i = 0; i < block_size; ++i) {
        int32_t v;
        if (sub->kind == zf::kSubConst) {
            v = sub->constant;
        } else {
            v = frame[ch * 8 + zf::sub_idx(i, c8)];
            if (sub->kind == zf::kSubPred && i >= sub->order) {
                int64_t acc = 0;
                for (uint32_t j = 0; j < sub->order; ++j) acc += (int64_t)sub->coef[j] * (int64_t)plane[i - 1 - j];
                v += (int32_t)(acc >> sub->shift);
            }
        }
        plane[i] = v;
    }
    for (uint32_t i = 0; i < block_size; ++i) plane[i] = (int32_t)((uint32_t)plane[i] << sub->wasted);
}

int zf_host_subrec_size(void) { return (int)sizeof(zf::SubRec); }

void zf_host_set_rollback_counter(long long *p) { g_rollbacks = p; }

int zf_host_framerec_size(void) { return (int)sizeof(zf::FrameRec); }
}
