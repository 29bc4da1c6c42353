// scan.cu — K1 sync_scan and K1b frame_link (sm_100a).
//
// K1 restates, as a data-parallel predicate over every byte position, the frame-header
// grammar the reference walks sequentially (src/zflac.zig:343-374 FrameHeader bitfields
// :165-173, read_coded_number :203-214, block-size / sample-rate escapes :356-374), plus
// the CRC-8 the reference reads and ignores (:407-410). Candidates leave the kernel in
// stream order: per-tile ballot/scan ranks + a decoupled look-back across tiles.
//
// K1b computes the CRC-16 (:551, also ignored by the reference) of the bytes between
// neighbouring candidates. A match proves "candidate k is a frame that ends where candidate j
// starts" without entropy-decoding it, which is what lets K2 decode every frame in parallel
// with its output offset already known. A mismatch proves nothing (the reference would
// still decode the frame, Q1): the host then probes that frame's end with K2 instead.
#include <mutex>
#include <string>

#include "../../include/zflacb200.h"
#include "common.cuh"
#include "kernels.cuh"

namespace zf {

// ------------------------------------------------------------------ CRC helpers
// CRC-8 (poly 0x07, init 0) one byte per table look-up: the header check runs with a single active lane, so its
// instruction count is paid by the whole warp (the bitwise form was ~40 % of K1's issue slots).
struct Crc8Table {
    uint8_t t[256];
    constexpr Crc8Table() : t() {
        for (int v = 0; v < 256; ++v) {
            unsigned crc = (unsigned)v;
            for (int i = 0; i < 8; ++i) crc = (crc & 0x80u) ? ((crc << 1) ^ 0x07u) & 0xFFu : (crc << 1) & 0xFFu;
            t[v] = (uint8_t)crc;
        }
    }
};
__constant__ Crc8Table c_crc8 = Crc8Table();
__device__ __forceinline__ uint32_t crc8_byte(uint32_t crc, uint32_t b) { return c_crc8.t[(crc ^ b) & 0xFFu]; }

__constant__ uint32_t c_rate_table[16] = {0, 88200, 176400, 192000, 8000, 16000, 22050, 24000,
                                          32000, 44100, 48000, 96000, 0, 0, 0, 0};

// The 16 stream bytes at position p as two little-endian 64-bit words (bytes at or past len read as zero). Where the
// two 16-byte groups around p are inside the stream they come as two aligned vector loads (L1 / L2 hits: the scan has just
// read them) and funnel shifts — the check of a filter hit used to be 33 dependent byte loads, which is what the other
// warps of the tile waited for at the barrier (r04q).
__device__ __forceinline__ void fetch16(const uint8_t *__restrict__ s, uint64_t p, uint64_t len, uint64_t &lo, uint64_t &hi) {
    const uint64_t a = p & ~15ull;
    if (a + 32 <= len) {
        const uint4 x = __ldg(reinterpret_cast<const uint4 *>(s + a)), y = __ldg(reinterpret_cast<const uint4 *>(s + a + 16));
        const uint32_t wi = (uint32_t)(p & 15u) >> 2, sh = (uint32_t)(p & 3u) * 8u;
        // the five words that hold bytes p .. p + 15
        const uint32_t w0 = wi == 0 ? x.x : wi == 1 ? x.y : wi == 2 ? x.z : x.w;
        const uint32_t w1 = wi == 0 ? x.y : wi == 1 ? x.z : wi == 2 ? x.w : y.x;
        const uint32_t w2 = wi == 0 ? x.z : wi == 1 ? x.w : wi == 2 ? y.x : y.y;
        const uint32_t w3 = wi == 0 ? x.w : wi == 1 ? y.x : wi == 2 ? y.y : y.z;
        const uint32_t w4 = wi == 0 ? y.x : wi == 1 ? y.y : wi == 2 ? y.z : y.w;
        const uint32_t o0 = __funnelshift_r(w0, w1, sh), o1 = __funnelshift_r(w1, w2, sh);
        const uint32_t o2 = __funnelshift_r(w2, w3, sh), o3 = __funnelshift_r(w3, w4, sh);
        lo = (uint64_t)o0 | ((uint64_t)o1 << 32);
        hi = (uint64_t)o2 | ((uint64_t)o3 << 32);
        return;
    }
    lo = 0;
    hi = 0;
    for (int i = 0; i < 16 && p + i < len; ++i) {
        const uint64_t v = s[p + i];
        if (i < 8) lo |= v << (8 * i);
        else hi |= v << (8 * (i - 8));
    }
}
__device__ __forceinline__ uint32_t byte_of(uint64_t lo, uint64_t hi, int i) {  // i in [0, 16)
    return (uint32_t)((i < 8 ? lo >> (8 * i) : hi >> (8 * (i - 8))) & 0xFFu);
}

// Full header check at byte position p. Returns true and fills `c` when the bytes at p are a
// frame header the reference would accept (grammar) AND the CRC-8 matches.
__device__ __noinline__ bool parse_header_at(const uint8_t *__restrict__ s, uint64_t p, uint64_t len, Cand &c) {
    if (p + 6 > len) return false;
    const int avail = (len - p) < 16 ? (int)(len - p) : 16;
    uint64_t lo, hi;
    fetch16(s, p, len, lo, hi);
    const uint32_t b0 = (uint32_t)lo & 0xFFu, b1 = (uint32_t)(lo >> 8) & 0xFFu, b2 = (uint32_t)(lo >> 16) & 0xFFu, b3 = (uint32_t)(lo >> 24) & 0xFFu;
    if (b0 != 0xFF || (b1 & 0xFE) != 0xF8) return false;
    const uint32_t bs_code = b2 >> 4, sr_code = b2 & 15, ch_code = b3 >> 4, bd_code = (b3 >> 1) & 7;
    if (bs_code == 0 || sr_code == 15 || ch_code > 10 || bd_code == 3) return false;
    const uint32_t c0 = (uint32_t)(lo >> 32) & 0xFFu;
    const uint32_t ones = __clz((int)((~c0 & 0xFFu) << 24 | 0x00800000u));  // leading ones of c0 (0..8)
    if (c0 == 0xFF || ones == 1) return false;
    int n = 4 + (ones == 0 ? 1 : (int)ones);
    // value of the UTF-8-like number (:203-214 checks the lead byte only; continuation bytes are taken as they are)
    uint64_t number = ones == 0 ? c0 : (c0 & (0x7Fu >> ones));
    for (int i = 5; i < n; ++i) number = (number << 6) | (byte_of(lo, hi, i) & 0x3Fu);
    uint32_t block_size;
    if (bs_code == 6) {
        block_size = byte_of(lo, hi, n) + 1;
        n += 1;
    } else if (bs_code == 7) {
        const uint32_t v = (byte_of(lo, hi, n) << 8) | byte_of(lo, hi, n + 1);
        if (v == 0xFFFF) return false;
        block_size = v + 1;
        n += 2;
    } else if (bs_code == 1) {
        block_size = 192;
    } else if (bs_code <= 5) {
        block_size = 144u << bs_code;
    } else {
        block_size = 1u << bs_code;
    }
    uint32_t sr_value = c_rate_table[sr_code];
    if (sr_code == 12) {
        sr_value = byte_of(lo, hi, n);
        n += 1;
    } else if (sr_code == 13) {
        sr_value = (byte_of(lo, hi, n) << 8) | byte_of(lo, hi, n + 1);
        n += 2;
    } else if (sr_code == 14) {
        sr_value = 10u * ((byte_of(lo, hi, n) << 8) | byte_of(lo, hi, n + 1));
        n += 2;
    }
    if (n + 1 > avail) return false;
    uint32_t crc = 0;
    for (int i = 0; i < n; ++i) crc = crc8_byte(crc, byte_of(lo, hi, i));
    if (crc != byte_of(lo, hi, n)) return false;
    c.offset = p;
    c.block_size = block_size;
    c.sr_value = sr_value;
    c.header_len = (uint8_t)(n + 1);
    c.ch_code = (uint8_t)ch_code;
    c.bd_code = (uint8_t)bd_code;
    c.sr_code = (uint8_t)sr_code;
    c.number = number;
    c.strategy = (uint8_t)(b1 & 1u);
    c.pad[0] = c.pad[1] = c.pad[2] = 0;
    return true;
}

// ------------------------------------------------------------------ K1
// One CTA scans one 64 KB tile as 16 rows of 256 threads x 16 bytes, every global load a coalesced 128-bit vector,
// the loads of the next four rows in flight while four rows are tested.
//
// The kernel is a streaming read and was instruction bound (r02z: 70 M warp instructions for 284 MB, 23 % of the HBM
// peak): every row paid the exact test, per-row counts and five 64-bit shuffle steps of an ordered scan although a
// candidate turns up once per ~7 KB. Now a row costs a FILTER — four instructions per four stream positions that say
// whether any byte is 0xFF and followed by a byte >= 0xF8 — and nothing else; a thread whose filter fires (2^-9 per
// row) runs the exact sync test and the header check (grammar + CRC-8), and a validated header sets ONE BIT in a
// bitmap of the tile in shared memory. Order comes for free afterwards: thread t owns bits [256 t, 256 t + 256), a
// block scan of the population counts ranks them, one decoupled look-back across tiles places the tile, and the few
// headers are parsed once more where they are written.
constexpr int kScanThreads = 256;
constexpr int kScanRows = 16;
constexpr int kScanTileBytes = kScanThreads * 16 * kScanRows;  // 65536

constexpr unsigned long long kFlagAgg = 1ull << 62;
constexpr unsigned long long kFlagPrefix = 2ull << 62;
constexpr unsigned long long kValueMask = (1ull << 62) - 1;

// 0x80 in every byte i where byte i of cur == 0xFF and byte i of nxt (the following stream byte) is 0xF8 or 0xF9:
// t has a zero byte exactly there, and ~(((t & 0x7F..) + 0x7F..) | t | 0x7F..) is the exact zero-byte test (no borrow
// between bytes).
__device__ __forceinline__ uint32_t sync_hits4(uint32_t cur, uint32_t nxt) {
    const uint32_t t = ~cur | ((nxt | 0x01010101u) ^ 0xF9F9F9F9u);
    return ~(((t & 0x7F7F7F7Fu) + 0x7F7F7F7Fu) | t | 0x7F7F7F7Fu);
}
__device__ __forceinline__ uint32_t hits_to_mask4(uint32_t z) {  // 4-bit mask, bit i <-> byte i
    return (((z >> 7) & 0x01010101u) * 0x01020408u) >> 24;
}
// Filter: non-zero (after the caller's & 0x80808080) iff some byte i has cur == 0xFF and nxt in {0xF8, 0xF9}: the sync
// code. t = cur & ((nxt | 1) ^ 6) has an 0xFF byte exactly there ((0xF9 ^ 0x06) == 0xFF), and (~t - 0x01..) & t & 0x80..
// is non-zero iff ~t has a zero byte. Four instructions per four stream positions (c1 / c6 live in registers so that
// (nxt | c1) ^ c6 is one LOP3); WHICH byte it was is left to the slow path (borrows may flag bytes above the real one).
__device__ __forceinline__ uint32_t sync_filter4(uint32_t cur, uint32_t nxt, uint32_t c1, uint32_t c6) {
    const uint32_t t = cur & ((nxt | c1) ^ c6);
    return (0u - (t + 0x01010102u)) & t;
}

// The exact test + header check of one 16-byte chunk whose filter fired (rare): sets the bit of every validated header.
// Chunks are queued by the scanning threads and checked afterwards, one chunk per thread, so that a hit costs the
// scanning warp a handful of instructions instead of a header check executed for one lane of it.
// The headers a tile validated, kept in shared memory so that writing them out in order does not parse them again
// (the second parse was a tenth of the kernel's instructions, executed for one lane at a time).
constexpr uint32_t kTileCands = 24;  // a 64 KB tile of audio has ~10; beyond: parsed again where they are written
struct TileCands {
    Cand c[kTileCands];
    uint16_t rel[kTileCands];
    uint32_t n;
};

static __device__ __noinline__ void scan_check_chunk(const uint8_t *__restrict__ s, uint64_t pos, uint64_t tile_base, uint64_t scan_begin,
                                                     uint64_t scan_end, uint64_t len, uint32_t *bits, TileCands *cache) {
    uint32_t w[5] = {0, 0, 0, 0, 0};
    if (pos + 32 <= len) {  // pos is a multiple of 16
        const uint4 v = __ldg(reinterpret_cast<const uint4 *>(s + pos));
        w[0] = v.x; w[1] = v.y; w[2] = v.z; w[3] = v.w;
        w[4] = __ldg(reinterpret_cast<const uint32_t *>(s + pos + 16));
    } else {
        for (int i = 0; i < 17 && pos + i < len; ++i) w[i >> 2] |= (uint32_t)s[pos + i] << (8 * (i & 3));
    }
    const uint32_t z0 = sync_hits4(w[0], __funnelshift_r(w[0], w[1], 8)), z1 = sync_hits4(w[1], __funnelshift_r(w[1], w[2], 8));
    const uint32_t z2 = sync_hits4(w[2], __funnelshift_r(w[2], w[3], 8)), z3 = sync_hits4(w[3], __funnelshift_r(w[3], w[4], 8));
    uint32_t raw = hits_to_mask4(z0) | (hits_to_mask4(z1) << 4) | (hits_to_mask4(z2) << 8) | (hits_to_mask4(z3) << 12);
    while (raw) {
        const int b = __ffs(raw) - 1;
        raw &= raw - 1;
        const uint64_t p = pos + b;
        Cand c;
        if (p >= scan_begin && p < scan_end && parse_header_at(s, p, len, c)) {
            const uint32_t rel = (uint32_t)(p - tile_base);
            atomicOr(&bits[rel >> 5], 1u << (rel & 31u));
            const uint32_t k = atomicAdd(&cache->n, 1u);
            if (k < kTileCands) {
                cache->c[k] = c;
                cache->rel[k] = (uint16_t)rel;
            }
        }
    }
}
constexpr uint32_t kScanQueue = 512;  // chunks a tile can queue (a tile of audio has ~10 frames and ~2 chance matches); beyond: checked in place

__global__ void __launch_bounds__(kScanThreads, 3)
k_sync_scan(const uint8_t *__restrict__ s, uint64_t scan_begin, uint64_t scan_end, uint64_t len, uint64_t tile0, uint32_t n_tiles,
            Cand *__restrict__ out, uint32_t cap, uint32_t *__restrict__ total_out,
            unsigned long long *__restrict__ tile_state, uint32_t *__restrict__ ticket) {
    __shared__ uint32_t s_tile;
    __shared__ __align__(16) uint32_t s_bits[kScanTileBytes / 32];  // one bit per byte of the tile: a validated header starts there
    __shared__ uint32_t s_wsum[kScanThreads / 32];
    __shared__ unsigned long long s_base;
    __shared__ uint32_t s_nq;
    __shared__ uint16_t s_queue[kScanQueue];  // chunk index (tile byte / 16) of every chunk whose filter fired
    __shared__ __align__(16) TileCands s_cands;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) {
        s_tile = atomicAdd(ticket, 1u);  // tiles start in order: look-back below cannot deadlock
        s_nq = 0;
        s_cands.n = 0;
    }
    uint32_t c1, c6;  // filter constants the compiler must keep in registers (see sync_filter4)
    asm("mov.b32 %0, 0x01010101;" : "=r"(c1));
    asm("mov.b32 %0, 0x06060606;" : "=r"(c6));
    {
        uint4 *z = reinterpret_cast<uint4 *>(s_bits) + tid * 2;  // thread t clears the eight words it will own
        z[0] = make_uint4(0, 0, 0, 0);
        z[1] = make_uint4(0, 0, 0, 0);
    }
    __syncthreads();
    const uint32_t tile = s_tile;
    if (tile >= n_tiles) return;
    const uint64_t tile_base = tile0 + (uint64_t)tile * kScanTileBytes;
    const bool interior = tile_base + kScanTileBytes + 16 <= len;

    auto queue_chunk = [&](uint64_t pos) {  // rare
        const uint32_t q = atomicAdd(&s_nq, 1u);
        if (q < kScanQueue) s_queue[q] = (uint16_t)((pos - tile_base) >> 4);
        else scan_check_chunk(s, pos, tile_base, scan_begin, scan_end, len, s_bits, &s_cands);
    };
    if (interior) {  // the whole tile and the 16 bytes after it are inside the stream: no length checks
        // warp w scans the CONTIGUOUS 8 KB [8192 w, 8192 w + 8192) of the tile, 16 rows of 512 bytes: the byte after
        // lane 31's sixteen is lane 0's first byte of the next row, which is in registers already (rows are
        // prefetched four ahead) — one extra load per warp and tile instead of one per row
        const uint64_t warp_base = tile_base + (uint64_t)warp * (kScanRows * 512);
        const uint4 *src = reinterpret_cast<const uint4 *>(s + warp_base) + lane;
        uint4 v[4], nx[4];
#pragma unroll
        for (int r = 0; r < 4; ++r) v[r] = __ldg(src + r * 32);
        const uint32_t after = __ldg(reinterpret_cast<const uint32_t *>(s + warp_base + kScanRows * 512));  // the next warp's (tile's) first bytes
#pragma unroll 1
        for (int r0 = 0; r0 < kScanRows; r0 += 4) {
#pragma unroll
            for (int r = 0; r < 4; ++r) nx[r] = (r0 + 4 < kScanRows) ? __ldg(src + (r0 + 4 + r) * 32) : make_uint4(after, 0, 0, 0);
#pragma unroll
            for (int r = 0; r < 4; ++r) {
                const uint32_t next_row0 = __shfl_sync(0xFFFFFFFFu, r < 3 ? v[r + 1].x : nx[0].x, 0);
                uint32_t nb = __shfl_down_sync(0xFFFFFFFFu, v[r].x, 1);  // the byte after my 16: the next lane's first
                if (lane == 31) nb = next_row0;
                const uint32_t w0 = v[r].x, w1 = v[r].y, w2 = v[r].z, w3 = v[r].w;
                const uint32_t f = (sync_filter4(w0, __funnelshift_r(w0, w1, 8), c1, c6) | sync_filter4(w1, __funnelshift_r(w1, w2, 8), c1, c6) |
                                    sync_filter4(w2, __funnelshift_r(w2, w3, 8), c1, c6) | sync_filter4(w3, __funnelshift_r(w3, nb, 8), c1, c6)) &
                                   0x80808080u;
                if (f) queue_chunk(warp_base + (uint64_t)(r0 + r) * 512 + (uint64_t)lane * 16);
            }
#pragma unroll
            for (int r = 0; r < 4; ++r) v[r] = nx[r];
        }
    } else {
#pragma unroll 1
        for (int r = 0; r < kScanRows; ++r) {
            const uint64_t pos = tile_base + (uint64_t)r * (kScanThreads * 16) + (uint64_t)tid * 16;
            uint4 v = make_uint4(0, 0, 0, 0);
            if (pos + 16 <= len) {
                v = __ldg(reinterpret_cast<const uint4 *>(s + pos));
            } else if (pos < len) {  // stream tail: bytes past len read as zero
                uint32_t t[4] = {0, 0, 0, 0};
                for (int i = 0; i < 16 && pos + i < len; ++i) t[i >> 2] |= (uint32_t)s[pos + i] << (8 * (i & 3));
                v = make_uint4(t[0], t[1], t[2], t[3]);
            }
            uint32_t nb = __shfl_down_sync(0xFFFFFFFFu, v.x, 1);
            if (lane == 31) nb = (pos + 16 < len) ? s[pos + 16] : 0u;
            const uint32_t f = (sync_filter4(v.x, __funnelshift_r(v.x, v.y, 8), c1, c6) | sync_filter4(v.y, __funnelshift_r(v.y, v.z, 8), c1, c6) |
                                sync_filter4(v.z, __funnelshift_r(v.z, v.w, 8), c1, c6) | sync_filter4(v.w, __funnelshift_r(v.w, nb, 8), c1, c6)) &
                               0x80808080u;
            if (f) queue_chunk(pos);
        }
    }
    __syncthreads();
    // ---- the queued chunks, one per thread: exact sync test, header grammar, CRC-8 -> bits
    {
        const uint32_t nq = s_nq < kScanQueue ? s_nq : kScanQueue;
        for (uint32_t q = tid; q < nq; q += kScanThreads)
            scan_check_chunk(s, tile_base + (uint64_t)s_queue[q] * 16, tile_base, scan_begin, scan_end, len, s_bits, &s_cands);
    }
    __syncthreads();

    // ---- ranks: thread t owns tile bytes [256 t, 256 t + 256) = bitmap words [8 t, 8 t + 8)
    uint32_t mine[8];
    {
        const uint4 a = reinterpret_cast<const uint4 *>(s_bits)[tid * 2], b = reinterpret_cast<const uint4 *>(s_bits)[tid * 2 + 1];
        mine[0] = a.x; mine[1] = a.y; mine[2] = a.z; mine[3] = a.w;
        mine[4] = b.x; mine[5] = b.y; mine[6] = b.z; mine[7] = b.w;
    }
    uint32_t cnt = 0;
#pragma unroll
    for (int k = 0; k < 8; ++k) cnt += __popc(mine[k]);
    uint32_t incl = cnt;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const uint32_t t = __shfl_up_sync(0xFFFFFFFFu, incl, d);
        if (lane >= d) incl += t;
    }
    if (lane == 31) s_wsum[warp] = incl;
    __syncthreads();

    // ---- decoupled look-back by warp 0: tile t publishes its aggregate, then 32 predecessors at a
    // time are inspected until one of them carries an inclusive prefix
    if (warp == 0) {
        uint32_t running = 0;
#pragma unroll
        for (int w = 0; w < kScanThreads / 32; ++w) running += s_wsum[w];
        unsigned long long base = 0;
        volatile unsigned long long *st = tile_state;
        if (tile > 0) {
            if (lane == 0) atomicExch(&tile_state[tile], kFlagAgg | (unsigned long long)running);
            for (int64_t hi = (int64_t)tile - 1;; hi -= 32) {
                const int64_t idx = hi - lane;
                unsigned long long val = kFlagPrefix;  // before tile 0: an empty prefix
                if (idx >= 0) {
                    do {
                        val = st[idx];
                    } while ((val >> 62) == 0);
                }
                const uint32_t has_prefix = __ballot_sync(0xFFFFFFFFu, (val >> 62) == 2);
                const int first = has_prefix ? __ffs(has_prefix) - 1 : 32;  // nearest predecessor with a prefix
                unsigned long long contrib = lane <= first ? (val & kValueMask) : 0ull;
#pragma unroll
                for (int d = 16; d > 0; d >>= 1) contrib += __shfl_xor_sync(0xFFFFFFFFu, contrib, d);
                base += contrib;
                if (has_prefix) break;
            }
        }
        if (lane == 0) {
            atomicExch(&tile_state[tile], kFlagPrefix | (base + running));
            s_base = base;
            if (tile == n_tiles - 1) *total_out = (uint32_t)min(base + running, (unsigned long long)0xFFFFFFFFull);
        }
    }
    __syncthreads();
    if (cnt == 0) return;

    // ---- the tile's headers in stream order
    unsigned long long slot = s_base + (incl - cnt);
    for (int w = 0; w < warp; ++w) slot += s_wsum[w];
#pragma unroll 1
    for (int k = 0; k < 8; ++k) {
        uint32_t m = mine[k];
        while (m) {
            const int b = __ffs(m) - 1;
            m &= m - 1;
            const uint32_t rel = (uint32_t)tid * 256u + (uint32_t)k * 32u + (uint32_t)b;
            if (slot < cap) {
                const uint32_t nc = s_cands.n < kTileCands ? s_cands.n : kTileCands;
                uint32_t j = 0;
                while (j < nc && s_cands.rel[j] != rel) ++j;
                if (j < nc) {
                    const uint4 *src = reinterpret_cast<const uint4 *>(&s_cands.c[j]);
                    uint4 *dst = reinterpret_cast<uint4 *>(out + slot);
                    dst[0] = src[0];
                    dst[1] = src[1];
                } else {
                    Cand c;
                    parse_header_at(s, tile_base + rel, len, c);  // validated above
                    out[slot] = c;
                }
            }
            ++slot;
        }
    }
}

// ------------------------------------------------------------------ K1b
// CRC-16 (poly 0x8005, init 0, MSB first) as arithmetic in GF(2)[x]/P: the state after a
// message M is M(x)*x^16 mod P, so state(A||B) = state(A)*x^(8|B|) + state(B). A warp covers a
// byte range with coalesced 16-byte groups: lane L takes groups L, L+32, ... (Horner step
// x^(8*512)), the lanes are then aligned with x^(128*t) and XOR-reduced.
__constant__ uint16_t c_pow_x8[40];    // x^(8*2^i) mod P
__constant__ uint16_t c_pow_x128[32];  // x^(128*t) mod P
__constant__ uint16_t c_x3968;         // x^(8*(512-16)) mod P

__host__ __device__ inline uint32_t gf16_mul(uint32_t a, uint32_t b) {
    uint32_t r = 0;
    for (int i = 15; i >= 0; --i) {
        r = ((r << 1) ^ ((r & 0x8000u) ? 0x8005u : 0u)) & 0xFFFFu;
        if ((b >> i) & 1u) r ^= a;
    }
    return r;
}

struct CrcTables {
    uint16_t byte[256];   // state update for one byte
    uint16_t mul_hi[256]; // (v<<8) * x^3968
    uint16_t mul_lo[256]; // v * x^3968
};

// built once per device by k_init_crc_tables (scan_init_constants); every CTA copies the 1.5 KB into shared memory
__device__ CrcTables g_crc_tables;

__global__ void k_init_crc_tables() {
    const int v = threadIdx.x;
    uint32_t crc = (uint32_t)v << 8;
    for (int i = 0; i < 8; ++i) crc = (crc & 0x8000u) ? ((crc << 1) ^ 0x8005u) & 0xFFFFu : (crc << 1) & 0xFFFFu;
    g_crc_tables.byte[v] = (uint16_t)crc;
    g_crc_tables.mul_hi[v] = (uint16_t)gf16_mul((uint32_t)v << 8, c_x3968);
    g_crc_tables.mul_lo[v] = (uint16_t)gf16_mul((uint32_t)v, c_x3968);
}

__device__ void build_crc_tables(CrcTables &t) {
    static_assert(sizeof(CrcTables) % 4 == 0, "word copy");
    const uint32_t *src = reinterpret_cast<const uint32_t *>(&g_crc_tables);
    uint32_t *dst = reinterpret_cast<uint32_t *>(&t);
    for (int i = threadIdx.x; i < (int)(sizeof(CrcTables) / 4); i += blockDim.x) dst[i] = src[i];
    __syncthreads();
}

__device__ __forceinline__ uint32_t crc16_absorb_word(const CrcTables &t, uint32_t st, uint32_t w_le) {
    // w_le holds 4 stream bytes, first byte in bits 0..7
    st = ((st << 8) & 0xFFFFu) ^ t.byte[(st >> 8) ^ (w_le & 0xFFu)];
    st = ((st << 8) & 0xFFFFu) ^ t.byte[(st >> 8) ^ ((w_le >> 8) & 0xFFu)];
    st = ((st << 8) & 0xFFFFu) ^ t.byte[(st >> 8) ^ ((w_le >> 16) & 0xFFu)];
    st = ((st << 8) & 0xFFFFu) ^ t.byte[(st >> 8) ^ (w_le >> 24)];
    return st;
}

__device__ __forceinline__ uint32_t gf16_shift_bytes(uint32_t st, uint64_t nbytes) {
    for (int i = 0; nbytes; ++i, nbytes >>= 1)
        if (nbytes & 1) st = gf16_mul(st, c_pow_x8[i]);
    return st;
}

// CRC-16 state of s[a, e), computed by a whole warp; result valid in every lane.
__device__ uint32_t warp_crc16(const CrcTables &t, const uint8_t *__restrict__ s, uint64_t a, uint64_t e) {
    const int lane = threadIdx.x & 31;
    if (e <= a) return 0;
    const uint64_t a0 = a & ~15ull, e0 = e & ~15ull;
    uint32_t st = 0;
    uint64_t tail_from = a;
    if (e0 > a0) {
        const uint64_t m = (e0 - a0) >> 4;  // full 16-byte groups in [a0, e0)
        uint64_t last = 0;
        bool any = false;
        for (uint64_t g = lane; g < m; g += 32) {
            uint4 v = __ldg(reinterpret_cast<const uint4 *>(s + a0 + (g << 4)));
            if (g == 0 && a > a0) {  // leading bytes before `a` count as zeros (init 0: no effect)
                const uint32_t skip = (uint32_t)(a - a0);
                uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const int lo = i * 4;
                    if ((int)skip >= lo + 4) w[i] = 0;
                    else if ((int)skip > lo) w[i] &= 0xFFFFFFFFu << (8 * (skip - lo));
                }
                v = make_uint4(w[0], w[1], w[2], w[3]);
            }
            st = t.mul_hi[st >> 8] ^ t.mul_lo[st & 0xFFu];  // * x^(8*496)
            st = crc16_absorb_word(t, st, v.x);             // * x^128 + group
            st = crc16_absorb_word(t, st, v.y);
            st = crc16_absorb_word(t, st, v.z);
            st = crc16_absorb_word(t, st, v.w);
            last = g;
            any = true;
        }
        if (any) st = gf16_mul(st, c_pow_x128[(m - 1 - last) & 31]);
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) st ^= __shfl_xor_sync(0xFFFFFFFFu, st, d);
        tail_from = e0;
    }
    // tail (< 16 bytes, or a range that never reaches a 16-byte boundary): lane 0, bytewise
    if (lane == 0) {
        for (uint64_t p = tail_from; p < e; ++p) st = ((st << 8) & 0xFFFFu) ^ t.byte[(st >> 8) ^ s[p]];
    }
    return __shfl_sync(0xFFFFFFFFu, st, 0);
}

constexpr int kLinkThreads = 256;

__global__ void __launch_bounds__(kLinkThreads)
k_frame_link(const uint8_t *__restrict__ s, uint64_t len, const Cand *__restrict__ cands, uint32_t n,
             uint32_t *__restrict__ link, uint32_t window, uint64_t max_frame_bytes) {
    __shared__ CrcTables tab;
    build_crc_tables(tab);
    const uint32_t k = blockIdx.x * (kLinkThreads / 32) + (threadIdx.x >> 5);
    if (k >= n) return;
    const uint64_t a = cands[k].offset;
    uint32_t acc = 0, found = kNoLink;
    uint64_t done_to = a;
    for (uint32_t t = 1; t <= window; ++t) {
        const uint32_t j = k + t;
        if (j > n) break;
        const uint64_t b = (j < n) ? cands[j].offset : len;  // j == n: the end of the stream
        if (b - a > max_frame_bytes) break;
        if (b < a + 8) continue;  // shorter than any frame (header 6 + 1 + CRC-16... rounded)
        const uint64_t upto = b - 2;
        const uint32_t seg = warp_crc16(tab, s, done_to, upto);
        acc = gf16_shift_bytes(acc, upto - done_to) ^ seg;
        done_to = upto;
        const uint32_t stored = ((uint32_t)s[b - 2] << 8) | s[b - 1];
        if (acc == stored) {
            found = j;
            break;
        }
    }
    if ((threadIdx.x & 31) == 0) link[k] = found;
}

// CRC-16 of explicit [begin, end) ranges (diagnostic check of decoded frames).
__global__ void __launch_bounds__(kLinkThreads)
k_crc16_ranges(const uint8_t *__restrict__ s, const uint64_t *__restrict__ begin, const uint64_t *__restrict__ end,
               uint32_t n, uint16_t *__restrict__ crc_out) {
    __shared__ CrcTables tab;
    build_crc_tables(tab);
    const uint32_t k = blockIdx.x * (kLinkThreads / 32) + (threadIdx.x >> 5);
    if (k >= n) return;
    const uint32_t st = warp_crc16(tab, s, begin[k], end[k]);
    if ((threadIdx.x & 31) == 0) crc_out[k] = (uint16_t)st;
}

// CRC-16 of every decoded frame of a launch against the value K2 read at its end (:548-551; the
// reference reads it and moves on, Q1 — here it is counted per stream as a diagnostic).
__global__ void __launch_bounds__(kLinkThreads)
k_crc16_frames(const uint8_t *__restrict__ s, const FrameRec *__restrict__ recs, const FrameOut *__restrict__ outs, uint32_t n,
               StreamSummary *__restrict__ sums, uint8_t *__restrict__ frame_ok) {
    __shared__ CrcTables tab;
    build_crc_tables(tab);
    const uint32_t k = blockIdx.x * (kLinkThreads / 32) + (threadIdx.x >> 5);
    if (k >= n) return;
    const FrameRec fr = recs[k];
    if (fr.block_size == 0) return;
    const FrameOut o = outs[k];
    if (o.status != ST_OK || o.end_offset < fr.offset + 2) return;
    const uint32_t st = warp_crc16(tab, s, fr.offset, o.end_offset - 2);
    if ((threadIdx.x & 31) == 0) {
        const bool ok = st == (o.stored_crc16 & 0xFFFFu);
        if (frame_ok) frame_ok[k] = ok ? 1 : 0;
        if (!ok) atomicAdd(&sums[fr.stream_id].n_crc16_mismatch, 1u);
    }
}

// ------------------------------------------------------------------ host launchers
static uint32_t host_pow_x(uint64_t bits) {  // x^bits mod P
    uint32_t r = 1, base = 2;  // base = x
    while (bits) {
        if (bits & 1) r = gf16_mul(r, base);
        base = gf16_mul(base, base);
        bits >>= 1;
    }
    return r;
}

cudaError_t scan_init_constants() {
    uint16_t p8[40], p128[32];
    for (int i = 0; i < 40; ++i) p8[i] = (uint16_t)host_pow_x(8ull << i);
    for (int t = 0; t < 32; ++t) p128[t] = (uint16_t)host_pow_x(128ull * t);
    const uint16_t x3968 = (uint16_t)host_pow_x(3968);
    cudaError_t e;
    if ((e = cudaMemcpyToSymbol(c_pow_x8, p8, sizeof(p8))) != cudaSuccess) return e;
    if ((e = cudaMemcpyToSymbol(c_pow_x128, p128, sizeof(p128))) != cudaSuccess) return e;
    if ((e = cudaMemcpyToSymbol(c_x3968, &x3968, sizeof(x3968))) != cudaSuccess) return e;
    k_init_crc_tables<<<1, 256>>>();
    if ((e = cudaGetLastError()) != cudaSuccess) return e;
    return cudaDeviceSynchronize();
}

uint32_t scan_tile_count(uint64_t scan_begin, uint64_t len) {
    if (len <= scan_begin) return 0;
    const uint64_t tile0 = scan_begin & ~15ull;
    return (uint32_t)((len - tile0 + kScanTileBytes - 1) / kScanTileBytes);
}

cudaError_t launch_sync_scan(const uint8_t *d_stream, uint64_t scan_begin, uint64_t len, Cand *d_cands, uint32_t cap,
                             uint32_t *d_total, unsigned long long *d_tile_state, uint32_t *d_ticket,
                             cudaStream_t st) {
    return launch_sync_scan_range(d_stream, scan_begin, len, len, d_cands, cap, d_total, d_tile_state, d_ticket, st);
}

// Candidates whose sync code starts in [scan_begin, scan_end); bytes up to `len` may be read (a header
// that starts just before scan_end extends past it).
cudaError_t launch_sync_scan_range(const uint8_t *d_stream, uint64_t scan_begin, uint64_t scan_end, uint64_t len, Cand *d_cands,
                                   uint32_t cap, uint32_t *d_total, unsigned long long *d_tile_state, uint32_t *d_ticket,
                                   cudaStream_t st) {
    if (scan_end > len) scan_end = len;
    const uint32_t n_tiles = scan_tile_count(scan_begin, scan_end);
    cudaError_t e;
    if ((e = cudaMemsetAsync(d_total, 0, sizeof(uint32_t), st)) != cudaSuccess) return e;
    if (n_tiles == 0) return cudaSuccess;
    if ((e = cudaMemsetAsync(d_tile_state, 0, sizeof(unsigned long long) * n_tiles, st)) != cudaSuccess) return e;
    if ((e = cudaMemsetAsync(d_ticket, 0, sizeof(uint32_t), st)) != cudaSuccess) return e;
    ZF_COUNT_LAUNCH();
    k_sync_scan<<<n_tiles, kScanThreads, 0, st>>>(d_stream, scan_begin, scan_end, len, scan_begin & ~15ull, n_tiles, d_cands, cap,
                                                  d_total, d_tile_state, d_ticket);
    return cudaGetLastError();
}

cudaError_t launch_frame_link(const uint8_t *d_stream, uint64_t len, const Cand *d_cands, uint32_t n, uint32_t *d_link,
                              uint32_t window, uint64_t max_frame_bytes, cudaStream_t st) {
    if (n == 0) return cudaSuccess;
    const uint32_t per_block = kLinkThreads / 32;
    ZF_COUNT_LAUNCH();
    k_frame_link<<<(n + per_block - 1) / per_block, kLinkThreads, 0, st>>>(d_stream, len, d_cands, n, d_link, window,
                                                                          max_frame_bytes);
    return cudaGetLastError();
}

cudaError_t launch_crc16_ranges(const uint8_t *d_stream, const uint64_t *d_begin, const uint64_t *d_end, uint32_t n,
                                uint16_t *d_crc, cudaStream_t st) {
    if (n == 0) return cudaSuccess;
    const uint32_t per_block = kLinkThreads / 32;
    ZF_COUNT_LAUNCH();
    k_crc16_ranges<<<(n + per_block - 1) / per_block, kLinkThreads, 0, st>>>(d_stream, d_begin, d_end, n, d_crc);
    return cudaGetLastError();
}

cudaError_t launch_crc16_frames(const uint8_t *d_stream, const FrameRec *d_recs, const FrameOut *d_outs, uint32_t n,
                                StreamSummary *d_sums, uint8_t *d_frame_ok, cudaStream_t st) {
    if (n == 0) return cudaSuccess;
    const uint32_t per_block = kLinkThreads / 32;
    ZF_COUNT_LAUNCH();
    k_crc16_frames<<<(n + per_block - 1) / per_block, kLinkThreads, 0, st>>>(d_stream, d_recs, d_outs, n, d_sums, d_frame_ok);
    return cudaGetLastError();
}

}  // namespace zf

// ---------------------------------------------------------------------------------------------------------------
// zflacb200_scan_no_alloc — K1 for callers that may not allocate (no counterpart in the reference; include/zflacb200.h).
//
// k_sync_scan needs scratch that every other path takes from the library's block cache (cudaMalloc on a miss): one
// look-back word per 64 KB tile and a ticket counter. Here that scratch is a file-scope device array, part of the
// library's image on every GPU that loads it (128 KiB + 256 B, held until the process exits), so the call allocates
// nothing: two stream-ordered memsets and the launch.
//
// g_scan_scratch below is the DEVICE symbol. In host code the name is only a stand-in (its host address means nothing
// to a kernel), so the device address is fetched with cudaGetSymbolAddress, once per device, with that device current.
// ---------------------------------------------------------------------------------------------------------------
namespace {

constexpr uint32_t kNoAllocMaxTiles = 16384;  // x 64 KB = ZFLACB200_SCAN_NO_ALLOC_MAX_BYTES (1 GiB) of stream per call
constexpr uint32_t kNoAllocWords = kNoAllocMaxTiles + 32;  // tile states, then 256 bytes of counters (word 0 of them: the ticket)

static __device__ __align__(256) unsigned long long g_scan_scratch[kNoAllocWords];

std::mutex g_scan_scratch_mutex;
unsigned long long *g_scan_scratch_ptr[64];  // per device ordinal: the array's address ON THAT DEVICE (nullptr = not asked yet)

// the caller's device is current again when this goes out of scope
struct DeviceSwitch {
    int prev = -1;
    bool switched = false;
    cudaError_t to(int device) {
        cudaError_t e = cudaGetDevice(&prev);
        if (e != cudaSuccess) return e;
        if (prev == device) return cudaSuccess;
        e = cudaSetDevice(device);
        switched = e == cudaSuccess;
        return e;
    }
    ~DeviceSwitch() {
        if (switched) cudaSetDevice(prev);
    }
};

}  // namespace

extern "C" int zflacb200_scan_no_alloc(const void *d_stream, uint64_t len, uint64_t scan_begin, zflacb200_frame_header *d_headers,
                                       uint32_t capacity, uint32_t *d_count, int device, void *stream) {
    static_assert(sizeof(zflacb200_frame_header) == sizeof(zf::Cand), "zflacb200_frame_header is the scan's candidate record");
    static_assert(ZFLACB200_SCAN_NO_ALLOC_MAX_BYTES == (uint64_t)kNoAllocMaxTiles * zf::kScanTileBytes, "stated maximum");
    if (!d_stream || !d_count || (!d_headers && capacity != 0)) {
        zf::set_last_error("zflacb200_scan_no_alloc: null argument");
        return ZFLACB200_INVALID_ARGUMENT;
    }
    if ((reinterpret_cast<uintptr_t>(d_stream) & 15u) != 0 || (reinterpret_cast<uintptr_t>(d_headers) & 15u) != 0) {
        zf::set_last_error("zflacb200_scan_no_alloc: d_stream and d_headers must be 16-byte aligned device pointers");
        return ZFLACB200_INVALID_ARGUMENT;
    }
    if (scan_begin > len) scan_begin = len;
    if (len - (scan_begin & ~15ull) > ZFLACB200_SCAN_NO_ALLOC_MAX_BYTES) {
        zf::set_last_error("zflacb200_scan_no_alloc: more than ZFLACB200_SCAN_NO_ALLOC_MAX_BYTES (1 GiB) to scan in one call: "
                           "split the range or use zflacb200_open");
        return ZFLACB200_INVALID_ARGUMENT;
    }
    int dev = device;
    if (dev < 0 && cudaGetDevice(&dev) != cudaSuccess) {
        cudaGetLastError();
        zf::set_last_error("zflacb200_scan_no_alloc: no current CUDA device");
        return ZFLACB200_CUDA_ERROR;
    }
    if (dev < 0 || dev >= 64) {
        zf::set_last_error("zflacb200_scan_no_alloc: device ordinal out of range");
        return ZFLACB200_INVALID_ARGUMENT;
    }
    DeviceSwitch sw;
    cudaError_t e = sw.to(dev);
    unsigned long long *scratch = nullptr;
    if (e == cudaSuccess) {
        std::lock_guard<std::mutex> lk(g_scan_scratch_mutex);
        if (!g_scan_scratch_ptr[dev]) {  // set-up, once per device, with that device current
            void *p = nullptr;
            e = cudaGetSymbolAddress(&p, g_scan_scratch);
            if (e == cudaSuccess) g_scan_scratch_ptr[dev] = static_cast<unsigned long long *>(p);
        }
        scratch = g_scan_scratch_ptr[dev];
    }
    if (e == cudaSuccess) {
        // the array keeps the last call's contents: launch_sync_scan_range clears the tile states it uses, the ticket
        // and *d_count on `stream` ahead of the kernel
        e = zf::launch_sync_scan_range(static_cast<const uint8_t *>(d_stream), scan_begin, len, len, reinterpret_cast<zf::Cand *>(d_headers),
                                       capacity, d_count, scratch, reinterpret_cast<uint32_t *>(scratch + kNoAllocMaxTiles),
                                       static_cast<cudaStream_t>(stream));
    }
    if (e != cudaSuccess) {
        cudaGetLastError();
        zf::set_last_error((std::string("zflacb200_scan_no_alloc: ") + cudaGetErrorString(e)).c_str());
        return ZFLACB200_CUDA_ERROR;
    }
    return ZFLACB200_OK;
}
