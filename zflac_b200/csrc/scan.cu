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

// Full header check at byte position p. Returns true and fills `c` when the bytes at p are a
// frame header the reference would accept (grammar) AND the CRC-8 matches.
__device__ __noinline__ bool parse_header_at(const uint8_t *__restrict__ s, uint64_t p, uint64_t len, Cand &c) {
    if (p + 6 > len) return false;
    uint8_t b[16];
    const int avail = (len - p) < 16 ? (int)(len - p) : 16;
#pragma unroll
    for (int i = 0; i < 16; ++i) b[i] = (i < avail) ? s[p + i] : 0;
    if (b[0] != 0xFF || (b[1] & 0xFE) != 0xF8) return false;
    const uint32_t bs_code = b[2] >> 4, sr_code = b[2] & 15, ch_code = b[3] >> 4, bd_code = (b[3] >> 1) & 7;
    if (bs_code == 0 || sr_code == 15 || ch_code > 10 || bd_code == 3) return false;
    const uint32_t c0 = b[4];
    const uint32_t ones = __clz((int)((~c0 & 0xFFu) << 24 | 0x00800000u));  // leading ones of c0 (0..8)
    if (c0 == 0xFF || ones == 1) return false;
    int n = 4 + (ones == 0 ? 1 : (int)ones);
    // value of the UTF-8-like number (:203-214 checks the lead byte only; continuation bytes are taken as they are)
    uint64_t number = ones == 0 ? c0 : (c0 & (0x7Fu >> ones));
    for (int i = 5; i < n; ++i) number = (number << 6) | (b[i] & 0x3Fu);
    uint32_t block_size;
    if (bs_code == 6) {
        block_size = (uint32_t)b[n] + 1;
        n += 1;
    } else if (bs_code == 7) {
        const uint32_t v = ((uint32_t)b[n] << 8) | b[n + 1];
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
        sr_value = b[n];
        n += 1;
    } else if (sr_code == 13) {
        sr_value = ((uint32_t)b[n] << 8) | b[n + 1];
        n += 2;
    } else if (sr_code == 14) {
        sr_value = 10u * (((uint32_t)b[n] << 8) | b[n + 1]);
        n += 2;
    }
    if (n + 1 > avail) return false;
    uint32_t crc = 0;
    for (int i = 0; i < n; ++i) crc = crc8_byte(crc, b[i]);
    if (crc != b[n]) return false;
    c.offset = p;
    c.block_size = block_size;
    c.sr_value = sr_value;
    c.header_len = (uint8_t)(n + 1);
    c.ch_code = (uint8_t)ch_code;
    c.bd_code = (uint8_t)bd_code;
    c.sr_code = (uint8_t)sr_code;
    c.number = number;
    c.strategy = (uint8_t)(b[1] & 1u);
    c.pad[0] = c.pad[1] = c.pad[2] = 0;
    return true;
}

// ------------------------------------------------------------------ K1
// One CTA scans one 64 KB tile as 4 groups of 4 rows of 256 threads x 16 bytes, every global load a
// coalesced 128-bit vector, four of them in flight per thread. Sync codes are found four bytes at a
// time with plain integer arithmetic; only the (rare) hits run the full header check. Candidates
// leave in stream order: per group a warp scan of the four packed row counts, ONE block barrier for
// the whole tile, one decoupled look-back across tiles. (With 16 KB tiles the two barriers and the
// look-back per tile were the kernel: barrier stalls 10 per issued instruction, r02g.)
constexpr int kScanThreads = 256;
constexpr int kScanRows = 4;     // rows per group
constexpr int kScanGroups = 4;   // groups per tile
constexpr int kScanGroupBytes = kScanThreads * 16 * kScanRows;
constexpr int kScanTileBytes = kScanGroupBytes * kScanGroups;

constexpr unsigned long long kFlagAgg = 1ull << 62;
constexpr unsigned long long kFlagPrefix = 2ull << 62;
constexpr unsigned long long kValueMask = (1ull << 62) - 1;

// 0x80 in every byte i where byte i of cur == 0xFF and byte i of nxt (the following stream byte) is 0xF8 or 0xF9:
// t has a zero byte exactly there, and ~(((t & 0x7F..) + 0x7F..) | t | 0x7F..) is the exact zero-byte test (no borrow
// between bytes). Six plain integer instructions per four stream positions.
__device__ __forceinline__ uint32_t sync_hits4(uint32_t cur, uint32_t nxt) {
    const uint32_t t = ~cur | ((nxt | 0x01010101u) ^ 0xF9F9F9F9u);
    return ~(((t & 0x7F7F7F7Fu) + 0x7F7F7F7Fu) | t | 0x7F7F7F7Fu);
}
__device__ __forceinline__ uint32_t hits_to_mask4(uint32_t z) {  // 4-bit mask, bit i <-> byte i
    return (((z >> 7) & 0x01010101u) * 0x01020408u) >> 24;
}

__global__ void __launch_bounds__(kScanThreads, 4)
k_sync_scan(const uint8_t *__restrict__ s, uint64_t scan_begin, uint64_t scan_end, uint64_t len, uint64_t tile0, uint32_t n_tiles,
            Cand *__restrict__ out, uint32_t cap, uint32_t *__restrict__ total_out,
            unsigned long long *__restrict__ tile_state, uint32_t *__restrict__ ticket) {
    __shared__ uint32_t s_tile;
    __shared__ unsigned long long s_warp_sum[kScanGroups][kScanThreads / 32];  // 4 x 16-bit row counts per warp and group
    __shared__ unsigned long long s_base;
    __shared__ Cand s_first[kScanThreads];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) s_tile = atomicAdd(ticket, 1u);  // tiles start in order: look-back below cannot deadlock
    __syncthreads();
    const uint32_t tile = s_tile;
    if (tile >= n_tiles) return;
    const uint64_t tile_base = tile0 + (uint64_t)tile * kScanTileBytes;
    const bool interior = tile_base + kScanTileBytes + 16 <= len;

    uint64_t first_off = ~0ull;  // stream offset of the candidate kept in s_first[tid]
    uint32_t valid[kScanGroups][kScanRows / 2];   // per row a 16-bit mask of validated candidates in my 16 bytes, two rows per word
    unsigned long long incl[kScanGroups];         // inclusive warp scan of the packed row counts (16 bits per row)
    unsigned long long mine[kScanGroups];
#pragma unroll
    for (int g = 0; g < kScanGroups; ++g) {
        const uint64_t group_base = tile_base + (uint64_t)g * kScanGroupBytes;
        // ---- all loads of the group first (4 independent 128-bit loads in flight per thread)
        uint4 v[kScanRows];
        if (interior) {  // the whole tile and the byte after it are inside the stream: no per-row length checks
            const uint4 *src = reinterpret_cast<const uint4 *>(s + group_base) + tid;
#pragma unroll
            for (int r = 0; r < kScanRows; ++r) v[r] = __ldg(src + r * kScanThreads);
        } else {
#pragma unroll
            for (int r = 0; r < kScanRows; ++r) {
                const uint64_t pos = group_base + (uint64_t)r * (kScanThreads * 16) + (uint64_t)tid * 16;
                v[r] = make_uint4(0, 0, 0, 0);
                if (pos + 16 <= len) {
                    v[r] = __ldg(reinterpret_cast<const uint4 *>(s + pos));
                } else if (pos < len) {  // stream tail: bytes past len read as zero
                    uint32_t t[4] = {0, 0, 0, 0};
                    for (int i = 0; i < 16 && pos + i < len; ++i) t[i >> 2] |= (uint32_t)s[pos + i] << (8 * (i & 3));
                    v[r] = make_uint4(t[0], t[1], t[2], t[3]);
                }
            }
        }
        unsigned long long counts = 0;   // 16 bits per row
        valid[g][0] = valid[g][1] = 0;
#pragma unroll
        for (int r = 0; r < kScanRows; ++r) {
            const uint64_t pos = group_base + (uint64_t)r * (kScanThreads * 16) + (uint64_t)tid * 16;
            // the byte after my 16: the next lane's first byte; lane 31 fetches it (L1/L2 hit: a neighbour loads that line)
            uint32_t nb;
            if (interior) {  // every lane reads it (an L1 hit: the neighbour loaded that line): no shuffle, no one-lane branch
                nb = __ldg(s + pos + 16);
            } else {
                nb = __shfl_down_sync(0xFFFFFFFFu, v[r].x, 1) & 0xFFu;
                if (lane == 31) nb = (pos + 16 < len) ? s[pos + 16] : 0u;
            }
            const uint32_t w0 = v[r].x, w1 = v[r].y, w2 = v[r].z, w3 = v[r].w;
            const uint32_t z0 = sync_hits4(w0, __funnelshift_r(w0, w1, 8)), z1 = sync_hits4(w1, __funnelshift_r(w1, w2, 8));
            const uint32_t z2 = sync_hits4(w2, __funnelshift_r(w2, w3, 8)), z3 = sync_hits4(w3, __funnelshift_r(w3, nb, 8));
            uint32_t raw = 0;
            if (z0 | z1 | z2 | z3)  // a sync pattern in my 16 bytes: 2^-11; in the warp's 512 bytes: 1.6 %
                raw = hits_to_mask4(z0) | (hits_to_mask4(z1) << 4) | (hits_to_mask4(z2) << 8) | (hits_to_mask4(z3) << 12);
            uint32_t ok = 0;
            while (raw) {  // rare: ~2^-15 per byte on compressed data
                const int b = __ffs(raw) - 1;
                raw &= raw - 1;
                const uint64_t p = pos + b;
                Cand c;
                if (p >= scan_begin && p < scan_end && parse_header_at(s, p, len, c)) {
                    ok |= 1u << b;
                    if (first_off == ~0ull) {  // almost every thread that has a candidate has exactly one: keep it for the output pass
                        first_off = p;
                        s_first[tid] = c;
                    }
                }
            }
            valid[g][r >> 1] |= ok << (16 * (r & 1));
            counts |= (unsigned long long)__popc(ok) << (16 * r);
        }
        // ---- warp-level inclusive scan of the four row counts at once (row counts <= 4096 fit 16 bits)
        unsigned long long in = counts;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const unsigned long long t = __shfl_up_sync(0xFFFFFFFFu, in, d);
            if (lane >= d) in += t;
        }
        if (lane == 31) s_warp_sum[g][warp] = in;
        incl[g] = in;
        mine[g] = counts;
    }
    __syncthreads();

    // ---- candidates in the whole tile (what the look-back publishes)
    uint32_t running = 0;
#pragma unroll
    for (int g = 0; g < kScanGroups; ++g) {
        unsigned long long all = 0;
#pragma unroll
        for (int w = 0; w < kScanThreads / 32; ++w) all += s_warp_sum[g][w];
        running += (uint32_t)(all & 0xFFFFu) + (uint32_t)((all >> 16) & 0xFFFFu) + (uint32_t)((all >> 32) & 0xFFFFu) + (uint32_t)(all >> 48);
    }

    // ---- decoupled look-back by warp 0: tile t publishes its aggregate, then 32 predecessors at a
    // time are inspected until one of them carries an inclusive prefix
    if (warp == 0) {
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
    const unsigned long long base = s_base;

    // ---- ranks inside the tile: groups in order, rows in order inside a group, threads in order inside a row
    uint32_t before = 0;  // candidates in earlier groups / rows of this tile
#pragma unroll
    for (int g = 0; g < kScanGroups; ++g) {
        unsigned long long warp_off = 0, all = 0;
#pragma unroll
        for (int w = 0; w < kScanThreads / 32; ++w) {
            const unsigned long long t = s_warp_sum[g][w];
            if (w < warp) warp_off += t;
            all += t;
        }
        const unsigned long long excl = warp_off + incl[g] - mine[g];
#pragma unroll
        for (int r = 0; r < kScanRows; ++r) {
            uint32_t ok = (valid[g][r >> 1] >> (16 * (r & 1))) & 0xFFFFu;
            unsigned long long slot = base + before + (uint32_t)((excl >> (16 * r)) & 0xFFFFu);
            before += (uint32_t)((all >> (16 * r)) & 0xFFFFu);
            const uint64_t pos = tile_base + (uint64_t)g * kScanGroupBytes + (uint64_t)r * (kScanThreads * 16) + (uint64_t)tid * 16;
            while (ok) {
                const int b = __ffs(ok) - 1;
                ok &= ok - 1;
                Cand c;
                if (first_off == pos + b) c = s_first[tid];
                else parse_header_at(s, pos + b, len, c);
                if (slot < cap) out[slot] = c;
                ++slot;
            }
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
