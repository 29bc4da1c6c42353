// ring_src.cuh — the device word source of the frame-serial kernels (K2 decode_frames, K2a parse_frames).
//
// Lane L of a CTA owns column L of a ring of kRingWords rows x NT columns of 32-bit words: stream
// word w of the lane's frame sits at ring[w % kRingWords][L]. Every shared load of the bit reader
// is therefore bank-conflict free whatever the lanes' cursors are (a lane only ever touches bank
// L % 32), and the fill is a 4-byte cp.async.ca per word — L1 keeps the 32-byte sector, so DRAM
// sees every compressed byte once. Fills happen at a uniform cadence (before each 8-row body) so
// the warp does not diverge on "my queue is empty"; completion is tracked with cp.async groups: a
// commit per refill, boundaries of the three newest groups remembered, and a wait_group cascade
// that only blocks on groups old enough to have landed.
#pragma once
#include "decode_core.cuh"
#include "pos_reader.cuh"

namespace zf {

constexpr int kDecodeThreads = 128;
constexpr uint32_t kRingWords = 32;  // 128 B per lane ahead of the cursor (~130 rows of 16-bit audio)

// Block until words < need have landed, waiting on as few (= as old) cp.async groups as possible.
// By-value in and out so that the ring bookkeeping stays in registers at the call sites.
static __device__ __noinline__ uint32_t ring_wait(uint32_t need, uint32_t ready, uint32_t h0, uint32_t h1, uint32_t h2, uint32_t filled) {
    asm volatile("cp.async.wait_group 3;" ::: "memory");
    ready = max(ready, h2);
    if (ready >= need) return ready;
    asm volatile("cp.async.wait_group 2;" ::: "memory");
    ready = max(ready, h1);
    if (ready >= need) return ready;
    asm volatile("cp.async.wait_group 1;" ::: "memory");
    ready = max(ready, h0);
    if (ready >= need) return ready;
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    return filled;
}

extern __shared__ uint32_t zf_ring[];  // kRingWords x NT words

template <int NT>
struct RingSrc {
    const uint8_t *g;    // global address of word 0 (4-byte aligned)
    uint32_t avail;      // bytes readable from g (stream end), clamped
    uint32_t col;        // shared-space byte address of zf_ring[lane column]: word w lives (w % kRingWords) * NT words further
    uint32_t filled;     // words < filled have been issued
    uint32_t ready;      // words < ready have landed
    uint32_t h0, h1, h2; // `filled` before the newest / 2nd newest / 3rd newest committed group

    __device__ __forceinline__ void init(const uint8_t *base, uint64_t bytes, uint32_t column) {
        g = base;
        avail = bytes > 0xFFFFFFF0ull ? 0xFFFFFFF0u : (uint32_t)bytes;
        col = (uint32_t)__cvta_generic_to_shared(&zf_ring[column]);
        filled = ready = h0 = h1 = h2 = 0;
    }
    static constexpr uint32_t kWordScale = NT * 4;  // byte distance of consecutive words of a lane in the ring
    __device__ __forceinline__ uint32_t word_scaled(uint32_t ws) const {  // ws = word index * kWordScale
        uint32_t v;
        asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(col + (ws & ((kRingWords - 1) * kWordScale))));
        return v;
    }
    __device__ __forceinline__ uint32_t slot(uint32_t w) const { return col + (w & (kRingWords - 1)) * (NT * 4); }
    __device__ __forceinline__ void issue(uint32_t w) {
        const uint32_t dst = slot(w);
        if (w < (avail >> 2)) {  // a whole word inside the stream: the hot case
            asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(dst), "l"(g + (size_t)w * 4u) : "memory");
        } else {                 // bytes past the stream end are zero-filled
            const uint32_t off = w * 4u;
            const uint32_t n = off < avail ? avail - off : 0u;  // 0..3
            const uint8_t *src = g + (n ? off : 0u);
            asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"(dst), "l"(src), "r"(n) : "memory");
        }
    }
    // words < cur are consumed: issue everything the ring has room for
    __device__ __forceinline__ void fill(uint32_t cur) {
        if (filled < cur + kRingWords) {
            const uint32_t before = filled;
#pragma unroll 1
            do {
                issue(filled);
                ++filled;
            } while (filled < cur + kRingWords);
            asm volatile("cp.async.commit_group;" ::: "memory");
            h2 = h1;
            h1 = h0;
            h0 = before;
        }
    }
    __device__ __forceinline__ void wait_for(uint32_t need) { ready = ring_wait(need, ready, h0, h1, h2, filled); }
    __device__ __forceinline__ void top_up(uint32_t cur) {
        fill(cur);
        if (ready < cur + kFastWords) wait_for(cur + kFastWords);
    }
    __device__ __forceinline__ uint32_t word(uint32_t w) const {
        uint32_t v;  // volatile: stays ordered against the cp.async / wait_group statements
        asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(slot(w)));
        return v;
    }
    __device__ __forceinline__ void note_rollback() {}
    static __device__ __forceinline__ uint32_t fix(uint32_t raw) { return __byte_perm(raw, 0, 0x0123); }
    __device__ __forceinline__ uint32_t word_checked(uint32_t w) {
        if (w >= ready) {
            fill(w);
            wait_for(w + 1);
        }
        return word(w);
    }
};

// the last bytes of a stream (rare): byte-wise, zero past the end
static __device__ __noinline__ uint4 staged_tail_quad(const uint8_t *g, uint32_t off, uint32_t avail) {
    uint32_t v[4] = {0, 0, 0, 0};
    for (uint32_t b = 0; b < 16; ++b)
        if (off + b < avail) v[b >> 2] |= (uint32_t)g[off + b] << (8 * (b & 3));
    return make_uint4(v[0], v[1], v[2], v[3]);
}

// ---------------------------------------------------------------------------------------------
// StagedSrc — word source of K2a parse_frames. Ring of kStagedRing words per lane (word w of lane L at
// ring[w % kStagedRing][L], bank-conflict free), filled 64 bytes at a time THROUGH REGISTERS: four
// 128-bit global loads per sixteen words, issued one refill ahead of their sixteen shared stores.
//
// The scoreboard is per warp REGISTER: a shared store of the staging registers waits for the newest
// global load into them by ANY lane. So refills are not left to each lane's own rhythm (some lane
// would refill before almost every body and every body would wait for the previous body's loads):
// all lanes refill on the same bodies — every other one, keyed on the body index converged lanes
// share — which puts two bodies of work between a load and the store that needs it. A lane whose
// ring has no room just skips that turn. 16 words per turn cover the worst case of 8 words per body.
// Word 0 sits at a 16-byte aligned global address.
// ---------------------------------------------------------------------------------------------
constexpr uint32_t kStagedRing = 64;

template <int NT>
struct StagedSrc {
    static constexpr int kThreads = NT;
    static constexpr int kMinBlocks = 512 / NT;  // CTAs per SM the kernel is compiled for (512 lanes per SM: <= 128 registers)
    const uint8_t *g;    // global address of word 0 (16-byte aligned)
    uint32_t avail;      // bytes readable from g (stream end), clamped
    uint32_t col;        // shared-space byte address of zf_ring[lane column]
    uint32_t staged;     // words < staged are in the ring; words [staged, staged + 16) are in q[] (possibly still in flight)
    uint4 q[4];

    static constexpr uint32_t kWordScale = NT * 4;  // byte distance of consecutive words of a lane in the ring
    __device__ __forceinline__ uint32_t word_scaled(uint32_t ws) const {  // ws = word index * kWordScale
        uint32_t v;
        asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(col + (ws & ((kStagedRing - 1) * kWordScale))) : "memory");
        return v;
    }
    __device__ __forceinline__ uint32_t slot(uint32_t w) const { return col + (w & (kStagedRing - 1)) * (NT * 4); }
    __device__ __forceinline__ void load16(uint32_t w) {  // words [w, w + 16) -> q; bytes past the stream end read as zero
        const uint32_t off = w * 4u;
        if (off + 64u <= avail) {
            const uint4 *p = reinterpret_cast<const uint4 *>(g + off);
#pragma unroll
            for (int k = 0; k < 4; ++k) q[k] = __ldg(p + k);
            // DRAM latency here is ~3000 cycles, more than the two bodies between a load and its store: ask L2 for
            // the bytes ~8 refills ahead so that the loads above are L2 hits
            if (off + 1024u <= avail) asm volatile("prefetch.global.L2 [%0];" ::"l"(g + off + 576u));
        } else {
#pragma unroll
            for (int k = 0; k < 4; ++k) q[k] = staged_tail_quad(g, off + 16u * k, avail);
        }
    }
    __device__ __forceinline__ void store16() {  // q -> slots of words [staged, staged + 16); staged % 16 == 0: no wrap inside
        const uint32_t a = slot(staged);
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            asm volatile("st.shared.u32 [%0], %1;" ::"r"(a + NT * 4 * (4 * k + 0)), "r"(q[k].x) : "memory");
            asm volatile("st.shared.u32 [%0], %1;" ::"r"(a + NT * 4 * (4 * k + 1)), "r"(q[k].y) : "memory");
            asm volatile("st.shared.u32 [%0], %1;" ::"r"(a + NT * 4 * (4 * k + 2)), "r"(q[k].z) : "memory");
            asm volatile("st.shared.u32 [%0], %1;" ::"r"(a + NT * 4 * (4 * k + 3)), "r"(q[k].w) : "memory");
        }
    }
    __device__ __forceinline__ void flush_and_load() {
        store16();
        staged += 16;
        load16(staged);
    }
    __device__ __forceinline__ void init(const uint8_t *base, uint64_t bytes, uint32_t column) {
        g = base;
        avail = bytes > 0xFFFFFF80ull ? 0xFFFFFF80u : (uint32_t)bytes;
        col = (uint32_t)__cvta_generic_to_shared(&zf_ring[column]);
        staged = 0;
        load16(0);
#pragma unroll 1
        for (int k = 0; k < 3; ++k) flush_and_load();  // a frame starts with 48 words in the ring and 16 on their way
    }
    // words < cur are consumed; afterwards words [cur, cur + kFastWords) are resident
    __device__ __forceinline__ void top_up(uint32_t cur, uint32_t tick) {
        if ((tick & 1u) == 0 && staged + 16u <= cur + kStagedRing) flush_and_load();
#pragma unroll 1
        while (staged < cur + kFastWords) flush_and_load();  // only after slow-path rows that ran ahead of the refills
    }
    __device__ __forceinline__ void top_up(uint32_t cur) { top_up(cur, 0u); }
    __device__ __forceinline__ uint32_t word(uint32_t w) const {
        uint32_t v;  // volatile: stays ordered against the shared stores of the refills
        asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(slot(w)) : "memory");
        return v;
    }
    __device__ __forceinline__ void note_rollback() {}
    static __device__ __forceinline__ uint32_t fix(uint32_t raw) { return __byte_perm(raw, 0, 0x0123); }
    __device__ __forceinline__ uint32_t word_checked(uint32_t w) {
#pragma unroll 1
        while (w >= staged) flush_and_load();
        return word(w);
    }
};

// ---------------------------------------------------------------------------------------------
// QuadSrc — word source of K2a parse_frames for launches with more frames than one wave of the StagedSrc kernel holds
// (fewer registers: 7 CTAs per SM instead of 4). Same ring (word w of lane L at ring[w % kRingWords][L],
// bank-conflict free), filled 16 bytes at a time THROUGH REGISTERS: one 128-bit global load per four
// words, issued one refill ahead of its four shared stores, so the bit chain never waits on it and
// no cp.async group bookkeeping is needed (a lane's shared loads follow its own shared stores in
// program order). Word 0 sits at a 16-byte aligned global address.
// ---------------------------------------------------------------------------------------------
template <int NT>
struct QuadSrc {
    static constexpr int kThreads = NT;
    static constexpr int kMinBlocks = 896 / NT;  // CTAs per SM the kernel is compiled for (896 lanes per SM: <= 72 registers)
    const uint8_t *g;    // global address of word 0 (16-byte aligned)
    uint32_t avail;      // bytes readable from g (stream end), clamped
    uint32_t col;        // shared-space byte address of zf_ring[lane column]
    uint32_t staged;     // words < staged are in the ring; words [staged, staged + 4) are in q (possibly still in flight)
    uint4 q;

    static constexpr uint32_t kWordScale = NT * 4;  // byte distance of consecutive words of a lane in the ring
    __device__ __forceinline__ uint32_t word_scaled(uint32_t ws) const {  // ws = word index * kWordScale
        uint32_t v;
        asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(col + (ws & ((kRingWords - 1) * kWordScale))) : "memory");
        return v;
    }
    __device__ __forceinline__ uint32_t slot(uint32_t w) const { return col + (w & (kRingWords - 1)) * (NT * 4); }
    __device__ __forceinline__ void load_quad(uint32_t w) {  // words [w, w + 4) -> q; bytes past the stream end read as zero
        const uint32_t off = w * 4u;
        if (off + 16u <= avail) {
            q = __ldg(reinterpret_cast<const uint4 *>(g + off));
            // a lane streams through its frame: pull the line two ahead into L1 once per 128-byte line, so the
            // quad loads above are L1 hits when their turn comes (the ring itself only looks 16..32 bytes ahead)
            if ((w & 31u) == 0 && off + 384u <= avail) asm volatile("prefetch.global.L1 [%0];" ::"l"(g + off + 256u));
        } else {
            q = staged_tail_quad(g, off, avail);
        }
    }
    __device__ __forceinline__ void init(const uint8_t *base, uint64_t bytes, uint32_t column) {
        g = base;
        avail = bytes > 0xFFFFFFE0ull ? 0xFFFFFFE0u : (uint32_t)bytes;
        col = (uint32_t)__cvta_generic_to_shared(&zf_ring[column]);
        staged = 0;
        if (avail >= 256u) asm volatile("prefetch.global.L1 [%0];" ::"l"(g + 128u));
        load_quad(0);
    }
    __device__ __forceinline__ void flush_and_load() {  // q -> ring, then start loading the next four words
        const uint32_t a = slot(staged);  // staged is a multiple of 4: the four slots do not wrap
        asm volatile("st.shared.u32 [%0], %1;" ::"r"(a), "r"(q.x) : "memory");
        asm volatile("st.shared.u32 [%0], %1;" ::"r"(a + NT * 4), "r"(q.y) : "memory");
        asm volatile("st.shared.u32 [%0], %1;" ::"r"(a + NT * 8), "r"(q.z) : "memory");
        asm volatile("st.shared.u32 [%0], %1;" ::"r"(a + NT * 12), "r"(q.w) : "memory");
        staged += 4;
        load_quad(staged);
    }
    // words < cur are consumed; afterwards words [cur, cur + kFastWords) are resident
    __device__ __forceinline__ void top_up(uint32_t cur, uint32_t) { top_up(cur); }
    __device__ __forceinline__ void top_up(uint32_t cur) {
        if (staged + 4 <= cur + kRingWords && staged < cur + 24) flush_and_load();
        if (staged + 4 <= cur + kRingWords && staged < cur + 16) flush_and_load();
#pragma unroll 1
        while (staged < cur + kFastWords) flush_and_load();
    }
    __device__ __forceinline__ uint32_t word(uint32_t w) const {
        uint32_t v;  // volatile: stays ordered against the shared stores of flush_and_load
        asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(slot(w)) : "memory");
        return v;
    }
    __device__ __forceinline__ void note_rollback() {}
    static __device__ __forceinline__ uint32_t fix(uint32_t raw) { return __byte_perm(raw, 0, 0x0123); }
    __device__ __forceinline__ uint32_t word_checked(uint32_t w) {
#pragma unroll 1
        while (w >= staged) flush_and_load();
        return word(w);
    }
};


// ---------------------------------------------------------------------------------------------
// PairSrc — word source of the narrow K2a (pos_reader.cuh, parse_narrow.cuh). One-warp CTAs; ring entry j of lane L
// at ring[j % RING][L] holds stream words (j - 1, j) as {x, y}, both MSB-first, so ONE aligned 8-byte shared load
// yields the 64 stream bits around any position. Filled through registers like StagedSrc: four 128-bit global loads
// per sixteen words, issued one refill ahead of their sixteen 8-byte shared stores, every lane refilling on the same
// (even) bodies. The ring is STATIC shared memory, so its base is an immediate of the LDS and the address of an entry
// is two instructions: IMAD (255 - 8 * Q) and LOP3 ((that & ring mask) | lane * 8).
// Word 0 sits at a 16-byte aligned global address.
// ---------------------------------------------------------------------------------------------
template <int RING>
struct PairSrc {
    static constexpr int kThreads = 32;
    static constexpr uint32_t kRing = RING;
    const uint8_t *g;   // global address of word 0 (16-byte aligned)
    uint32_t avail;     // bytes readable from g (stream end), clamped
    uint32_t staged;    // entries < staged are in the ring; words [staged, staged + 16) are in q[] (possibly still in flight)
    uint32_t carry;     // word staged - 1, MSB-first
    uint32_t lane8;     // lane * 8: byte offset of the lane's column inside a ring row
    uint4 q[4];

    static __device__ __forceinline__ uint2 *ring() {
        __shared__ __align__(16) uint2 r[RING * 32];
        return r;
    }
    __device__ __forceinline__ uint2 entry_at(uint32_t Q) const {  // entry (31 - Q) >> 5; byte offset of its row = ((255 - 8Q) & mask)
        const uint32_t off = ((255u - 8u * Q) & ((RING - 1u) << 8)) | lane8;
        return *reinterpret_cast<const uint2 *>(reinterpret_cast<const char *>(ring()) + off);
    }
    __device__ __forceinline__ uint2 entry(uint32_t j) const {
        const uint32_t off = ((j & (RING - 1u)) << 8) | lane8;
        return *reinterpret_cast<const uint2 *>(reinterpret_cast<const char *>(ring()) + off);
    }
    __device__ __forceinline__ void load16(uint32_t w) {  // words [w, w + 16) -> q; bytes past the stream end read as zero
        const uint32_t off = w * 4u;
        if (off + 64u <= avail) {
            const uint4 *p = reinterpret_cast<const uint4 *>(g + off);
#pragma unroll
            for (int k = 0; k < 4; ++k) q[k] = __ldg(p + k);
            if (off + 1024u <= avail) asm volatile("prefetch.global.L2 [%0];" ::"l"(g + off + 576u));  // see StagedSrc
        } else {
#pragma unroll
            for (int k = 0; k < 4; ++k) q[k] = staged_tail_quad(g, off + 16u * k, avail);
        }
    }
    __device__ __forceinline__ void flush_and_load() {  // q -> entries [staged, staged + 16); staged % 16 == 0: no wrap inside
        char *row = reinterpret_cast<char *>(ring()) + (((staged & (RING - 1u)) << 8) | lane8);
        uint32_t prev = carry;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const uint32_t w0 = __byte_perm(q[k].x, 0, 0x0123), w1 = __byte_perm(q[k].y, 0, 0x0123);
            const uint32_t w2 = __byte_perm(q[k].z, 0, 0x0123), w3 = __byte_perm(q[k].w, 0, 0x0123);
            *reinterpret_cast<uint2 *>(row + 256 * (4 * k + 0)) = make_uint2(prev, w0);
            *reinterpret_cast<uint2 *>(row + 256 * (4 * k + 1)) = make_uint2(w0, w1);
            *reinterpret_cast<uint2 *>(row + 256 * (4 * k + 2)) = make_uint2(w1, w2);
            *reinterpret_cast<uint2 *>(row + 256 * (4 * k + 3)) = make_uint2(w2, w3);
            prev = w3;
        }
        carry = prev;
        staged += 16;
        load16(staged);
    }
    __device__ __forceinline__ void init(const uint8_t *base, uint64_t bytes, uint32_t lane) {
        g = base;
        avail = bytes > 0xFFFFFF80ull ? 0xFFFFFF80u : (uint32_t)bytes;
        lane8 = lane * 8u;
        staged = 0;
        carry = 0;
        load16(0);
#pragma unroll 1
        for (int k = 0; k < (RING >= 64 ? 3 : 1); ++k) flush_and_load();  // a frame starts with 48 (16) entries in the ring and 16 words on their way
    }
    // entries < j0 are consumed; afterwards entries [j0, j0 + kFastEntries) are resident.
    // WHEN a lane refills is a warp decision. The scoreboard is per warp register: a refill's shared stores wait for the
    // newest global load into q[] by ANY lane, so lanes refilling each on its own schedule (a quarter of them on every
    // other body, r2c: 70 % of all stall samples sat here) make every refill wait out a DRAM round trip. Instead nobody
    // refills until some lane runs low, and then every lane that has room does: the lanes of a warp consume at similar
    // rates, so refills come once per ~16 consumed words (six bodies of 16-bit audio) and the loads of one refill have
    // that long to land before the next one stores them.
    static constexpr uint32_t kLow = RING >= 64 ? 20u : 12u;
    __device__ __forceinline__ void top_up(uint32_t j0, uint32_t) {
        const uint32_t ahead = staged - j0;
        if (__any_sync(__activemask(), ahead < kLow) && ahead + 16u <= RING) flush_and_load();
#pragma unroll 1
        while (staged < j0 + kFastEntries) flush_and_load();  // only when the refills above fell behind (long codewords, slow-path rows)
    }
    // The same with a cursor that is one 8-row body old (the true one is at most 8 entries further): the refill decision
    // does not wait for the newest position, so it is off the bit chain's critical path.
    __device__ __forceinline__ void top_up_stale(uint32_t j_old) {
        const uint32_t ahead = staged - j_old;  // true ahead >= ahead - 8
        if (__any_sync(__activemask(), ahead < kLow + 8u) && ahead + 16u <= RING) flush_and_load();
#pragma unroll 1
        while (staged < j_old + 8u + kFastEntries) flush_and_load();
    }
    static constexpr bool kStaleCursor = RING >= 64;  // a 32-entry ring has no room for the 8 entries of slack a stale cursor costs
    __device__ __forceinline__ void ensure(uint32_t j) {
#pragma unroll 1
        while (j >= staged) flush_and_load();
    }
    __device__ __forceinline__ void body_done(bool) {}
};


}  // namespace zf
