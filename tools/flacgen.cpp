// flacgen.cpp — synthetic FLAC (RFC 9639) stream generator. TEST / BENCH INFRASTRUCTURE, not product code.
//
// BASELINE.json asks for test inputs "produced offline by a small in-repo Zig FLAC encoder that
// writes synthetic LPC/Rice-coded signals of the named shapes". There is no Zig toolchain in this
// image, so the encoder is C++ (same role, same outputs). It writes complete streams — fLaC
// marker, STREAMINFO with a correct MD5 and total, frames with valid CRC-8 / CRC-16 — for the
// shapes of BASELINE.json configs 2-5 and stays inside the behaviour the reference defines
// (SURVEY.md §7 Q8-Q10, Q16): no CONSTANT subframe on a side channel, side values that fit the
// container, predictor sums that fit the reference's accumulator type.
//
// Everything is a pure function of fg_params (seeded xoshiro256**, per-frame streams), so the
// same parameters give the same bytes on any machine and with any thread count.
//
// Build: g++ -O3 -ffp-contract=off -std=c++17 -shared -fPIC -pthread (tools/flacgen.py does it). The CUDA generator
// (flacgen_cuda.cu) writes the same bytes: what the two share, and why FMA contraction is off, is in flacgen_common.h.
#include <algorithm>
#include <atomic>
#include <chrono>
#include <cstdio>
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <thread>
#include <vector>

#include "flacgen_common.h"

extern "C" {


typedef struct fg_result {
    uint8_t *flac;
    uint64_t flac_bytes;
    void *pcm;            // interleaved, in the container the reference returns (i8/i16/i32), MSB-aligned (Q2)
    uint64_t pcm_bytes;
    uint64_t n_frames;
    uint64_t audio_offset;
    uint32_t container_bytes;
    uint32_t pad;
    uint8_t md5[16];
} fg_result;

int fg_generate(const fg_params *p, int want_pcm, fg_result *out);
int fg_generate_batch(const fg_params *ps, uint32_t n, int want_pcm, fg_result *outs, uint32_t n_threads);
void fg_free(fg_result *r);
void fg_set_trace(uint32_t *buf, uint64_t n_records);
}

namespace {

using fgc::Xo;
using fgc::zigzag;
using fgc::bits_needed_signed;
using fgc::container_of;

// ------------------------------------------------------------------ checksums
uint16_t g_crc16[8][256];
struct TableInit {
    TableInit() { fgc::crc16_make_tables(g_crc16); }
} g_table_init;

uint16_t crc16(const uint8_t *p, size_t n) {
    uint16_t c = 0;
    size_t i = 0;
    for (; i + 8 <= n; i += 8) {  // slice-by-8
        const uint8_t b0 = p[i] ^ (uint8_t)(c >> 8), b1 = p[i + 1] ^ (uint8_t)c;
        c = g_crc16[7][b0] ^ g_crc16[6][b1] ^ g_crc16[5][p[i + 2]] ^ g_crc16[4][p[i + 3]] ^ g_crc16[3][p[i + 4]] ^
            g_crc16[2][p[i + 5]] ^ g_crc16[1][p[i + 6]] ^ g_crc16[0][p[i + 7]];
    }
    for (; i < n; ++i) c = (uint16_t)((c << 8) ^ g_crc16[0][(c >> 8) ^ p[i]]);
    return c;
}

using fgc::Md5;

// ------------------------------------------------------------------ bit writer (MSB first)
struct BitWriter {
    std::vector<uint8_t> &out;
    uint64_t acc = 0;
    int n = 0;  // bits held in acc (low n bits)
    explicit BitWriter(std::vector<uint8_t> &o) : out(o) {}
    void put(uint64_t v, int bits) {  // bits in [0, 32]
        if (bits == 0) return;
        acc = (acc << bits) | (v & ((bits == 64) ? ~0ull : ((1ull << bits) - 1)));
        n += bits;
        while (n >= 8) {
            out.push_back((uint8_t)(acc >> (n - 8)));
            n -= 8;
        }
    }
    void put_signed(int64_t v, int bits) {
        if (bits > 32) {
            put((uint64_t)v >> 32, bits - 32);
            put((uint64_t)v & 0xFFFFFFFFull, 32);
        } else put((uint64_t)v, bits);
    }
    void put_unary(uint32_t q) {  // q zeros then a one
        while (q >= 32) { put(0, 32); q -= 32; }
        put(1, (int)q + 1);
    }
    void align() {
        if (n) put(0, 8 - n);
    }
};

// ------------------------------------------------------------------ frame plan
struct FramePlan {
    uint64_t first_sample;
    uint32_t block_size;
    uint64_t index;  // frame number
};

std::vector<FramePlan> plan_frames(const fg_params &p) {
    std::vector<FramePlan> v;
    uint64_t pos = 0, idx = 0;
    while (pos < p.n_samples) {
        uint32_t bs = p.variable_blocksize ? fgc::var_block_size(idx) : p.block_size;
        if ((uint64_t)bs > p.n_samples - pos) bs = (uint32_t)(p.n_samples - pos);
        v.push_back({pos, bs, idx});
        pos += bs;
        ++idx;
    }
    // the reference accepts a block size of 1 only on the last frame (zflac.zig:405): true by construction
    return v;
}

// ------------------------------------------------------------------ signal
// Samples of one frame, all channels, as exact integers of `bps` bits.
void synth_frame(const fg_params &p, const FramePlan &f, std::vector<std::vector<int64_t>> &ch) {
    const uint32_t C = p.channels, n = f.block_size;
    const fgc::SignalConsts k = fgc::signal_consts(p);
    const double sigma = fgc::frame_sigma(p, k, f.index);
    Xo rng(fgc::frame_rng_seed(p.seed, f.index));
    ch.assign(C, std::vector<int64_t>(n));
    std::vector<double> base(n);
    for (uint32_t c = 0; c < C; ++c) {
        const bool follow = (C == 2 && c == 1);  // R = 0.8 L + its own noise: gives mid/side something to win
        const fgc::ToneStep st = fgc::tone_step(p, c);
        fgc::ChannelSynth cs;
        cs.init(fgc::tone_state(p, c, f.first_sample), sigma, follow);
        for (uint32_t i = 0; i < n; ++i) {
            const double v = cs.next(rng, st, k, follow, base[i]);
            if (c == 0) base[i] = v;
            ch[c][i] = fgc::quantize_sample(v, k.limit);
        }
    }
    {   // some frames of a subframe_mix stream carry a constant channel
        const uint32_t c = fgc::constant_channel_of(p, f.index);
        if (c < C) std::fill(ch[c].begin(), ch[c].end(), ch[c][0]);
    }
    if (const uint32_t w = fgc::wasted_of_frame(p, f.index))  // low bits cleared on every channel of the frame
        for (auto &s : ch)
            for (auto &x : s) x = (x >> w) * ((int64_t)1 << w);
}

// ------------------------------------------------------------------ subframe encoding
struct RicePlan {
    uint32_t order;               // partition order
    std::vector<uint32_t> param;  // per partition; 0xFFFFFFFF = escaped
    std::vector<uint32_t> esc_bits;
    uint64_t bits;
};

// Cheapest partitioning of the residual res[0..n) (first `pred_order` entries absent).
RicePlan plan_rice(const fg_params &p, const std::vector<int64_t> &res, uint32_t bs, uint32_t pred_order, uint64_t salt) {
    const uint32_t kmax = fgc::rice_kmax(p);
    const uint32_t pbits = fgc::rice_pbits(p);
    RicePlan best;
    best.bits = ~0ull;
    auto valid = [&](uint32_t po) { return fgc::partition_order_valid(bs, po, pred_order); };
    uint32_t lo, hi;
    fgc::partition_order_range(p, bs, pred_order, lo, hi);
    for (uint32_t po = lo; po <= hi; ++po) {
        if (!valid(po)) break;
        RicePlan rp;
        rp.order = po;
        rp.bits = 0;
        const uint32_t np = 1u << po, psz = bs >> po;
        rp.param.resize(np);
        rp.esc_bits.assign(np, 0);
        for (uint32_t q = 0; q < np; ++q) {
            const uint32_t a = (q == 0) ? pred_order : q * psz, b = (q + 1) * psz;
            const uint32_t cnt = b - a;
            uint64_t sum = 0;
            int64_t mn = 0, mx = 0;
            for (uint32_t i = a; i < b; ++i) {
                sum += zigzag(res[i]);
                mn = std::min(mn, res[i]);
                mx = std::max(mx, res[i]);
            }
            uint32_t k = fgc::rice_k_of_mean(sum, cnt);
            const bool force_esc = fgc::partition_forced_escape(p, salt, q);
            if (k > kmax || force_esc) {
                rp.param[q] = 0xFFFFFFFFu;
                const int eb = (mn == 0 && mx == 0) ? 0 : bits_needed_signed(mn, mx);
                rp.esc_bits[q] = (uint32_t)eb;
                rp.bits += pbits + 5 + (uint64_t)eb * cnt;
            } else {
                // exact cost for k and k+1, keep the cheaper
                uint64_t bk[2] = {0, 0};
                for (int t = 0; t < 2; ++t) {
                    const uint32_t kk = std::min(kmax, k + (uint32_t)t);
                    uint64_t s = 0;
                    for (uint32_t i = a; i < b; ++i) s += (zigzag(res[i]) >> kk) + 1 + kk;
                    bk[t] = s;
                }
                if (bk[1] < bk[0] && k + 1 <= kmax) { k += 1; rp.bits += bk[1]; } else rp.bits += bk[0];
                rp.param[q] = k;
                rp.bits += pbits;
            }
        }
        if (rp.bits < best.bits) best = std::move(rp);
    }
    return best;
}

void write_residual(const fg_params &p, BitWriter &bw, const RicePlan &rp, const std::vector<int64_t> &res, uint32_t bs,
                    uint32_t pred_order) {
    const uint32_t pbits = fgc::rice_pbits(p);
    bw.put(p.rice_method ? 1 : 0, 2);
    bw.put(rp.order, 4);
    const uint32_t np = 1u << rp.order, psz = bs >> rp.order;
    for (uint32_t q = 0; q < np; ++q) {
        const uint32_t a = (q == 0) ? pred_order : q * psz, b = (q + 1) * psz;
        if (rp.param[q] == 0xFFFFFFFFu) {
            bw.put((1u << pbits) - 1, (int)pbits);
            bw.put(rp.esc_bits[q], 5);
            if (rp.esc_bits[q])
                for (uint32_t i = a; i < b; ++i) bw.put_signed(res[i], (int)rp.esc_bits[q]);
        } else {
            const uint32_t k = rp.param[q];
            bw.put(k, (int)pbits);
            for (uint32_t i = a; i < b; ++i) {
                const uint64_t u = zigzag(res[i]);
                const uint64_t qv = u >> k;
                if (qv + 1 + k <= 32) {
                    bw.put((1ull << k) | (u & ((1ull << k) - 1)), (int)(qv + 1 + k));
                } else {
                    bw.put_unary((uint32_t)qv);
                    bw.put(u & ((1ull << k) - 1), (int)k);
                }
            }
        }
    }
}

// Quantised LPC coefficients for s[0..n) (libFLAC-style: windowed autocorrelation, Levinson-Durbin,
// shift chosen so the largest coefficient uses the full precision, error feedback on rounding).
bool fit_lpc(const std::vector<int64_t> &s, uint32_t n, uint32_t order, uint32_t precision, int32_t *qcoef, uint32_t &shift) {
    std::vector<double> w(n);
    for (uint32_t i = 0; i < n; ++i) w[i] = fgc::hann(i, n) * (double)s[i];
    double ac[33];
    for (uint32_t lag = 0; lag <= order; ++lag) {
        double a = 0;
        for (uint32_t i = lag; i < n; ++i) a += w[i] * w[i - lag];
        ac[lag] = a;
    }
    return fgc::lpc_from_autocorr(ac, order, precision, qcoef, shift);
}

// Encodes one subframe of `depth` bits. `allow_constant` is false on side channels (Q8).
// acc_bits: width of the reference's accumulator type (16/32/64, Q10): predictors whose worst-case sum
// does not fit are not used.
// Debug trace (fg_set_trace): eight words per subframe — type code, wasted bits, precision, shift, partition order,
// bits written, a checksum of the subframe's samples, a checksum of the quantised coefficients. The CUDA generator
// fills the same records, so a byte mismatch between the two can be pinned to a stage.
uint32_t *g_trace = nullptr;
uint64_t g_trace_records = 0;

void encode_subframe(const fg_params &p, BitWriter &bw, const std::vector<int64_t> &sig, uint32_t bs, uint32_t depth,
                     bool allow_constant, uint32_t acc_bits, uint64_t salt, uint32_t *tr) {
    uint32_t tr_dummy[8];
    if (!tr) tr = tr_dummy;
    memset(tr, 0, 32);
    struct BitsAtExit {  // bits this call wrote
        BitWriter &bw; uint32_t *tr; uint64_t start;
        ~BitsAtExit() { tr[5] = (uint32_t)(bw.out.size() * 8 + (uint64_t)bw.n - start); }
    } at_exit{bw, tr, bw.out.size() * 8 + (uint64_t)bw.n};
    // wasted bits
    uint32_t wasted = 0;
    {
        uint64_t orv = 0;
        for (uint32_t i = 0; i < bs; ++i) orv |= (uint64_t)sig[i];
        if (orv != 0) {
            while (((orv >> wasted) & 1) == 0) ++wasted;
        }
        if (wasted >= depth) wasted = 0;
    }
    std::vector<int64_t> s(bs);
    for (uint32_t i = 0; i < bs; ++i) s[i] = sig[i] >> wasted;
    const uint32_t eff = depth - wasted;
    bool constant = true;
    for (uint32_t i = 1; i < bs && constant; ++i) constant = s[i] == s[0];

    tr[1] = wasted;
    for (uint32_t i = 0; i < bs; ++i) tr[6] += (uint32_t)s[i] * (i + 1);
    auto header = [&](uint32_t type) {
        tr[0] = type;
        bw.put(0, 1);
        bw.put(type, 6);
        bw.put(wasted ? 1 : 0, 1);
        if (wasted) bw.put_unary(wasted - 1);
    };
    auto verbatim = [&]() {
        header(1);
        for (uint32_t i = 0; i < bs; ++i) bw.put_signed(s[i], (int)eff);
    };

    if (constant && allow_constant) {
        // the reference reads bps - wasted bits here, which equals eff only off the side channel (Q8)
        header(0);
        bw.put_signed(s[0], (int)eff);
        return;
    }
    // choose the predictor
    const fgc::Choice choice = fgc::choose_predictor(p, salt, bs, eff, acc_bits);
    uint32_t kind = choice.kind, order = choice.order;
    if (kind == fgc::kVerbatim) { verbatim(); return; }

    int32_t qc[32] = {0};
    uint32_t shift = 0;
    if (kind == fgc::kLpc) {
        const uint32_t prec = choice.prec;
        if (!fit_lpc(s, bs, order, prec, qc, shift)) { kind = fgc::kFixed; order = std::min<uint32_t>(2, order); }
        else {
            // store the precision actually used
            std::vector<int64_t> res(bs, 0);
            bool ok = true;
            for (uint32_t i = order; i < bs && ok; ++i) {
                int64_t acc = 0;
                for (uint32_t j = 0; j < order; ++j) acc += (int64_t)qc[j] * s[i - 1 - j];
                const int64_t r = s[i] - (acc >> shift);
                if (r > 0x3FFFFFFFll || r < -0x40000000ll) ok = false;
                res[i] = r;
            }
            if (!ok) { verbatim(); return; }
            header(32 + order - 1);
            tr[2] = prec;
            tr[3] = shift;
            for (uint32_t j = 0; j < order; ++j) tr[7] += (uint32_t)qc[j] * (j + 1);
            for (uint32_t i = 0; i < order; ++i) bw.put_signed(s[i], (int)eff);
            bw.put(prec - 1, 4);
            bw.put(shift, 5);
            for (uint32_t j = 0; j < order; ++j) bw.put_signed(qc[j], (int)prec);
            const RicePlan rp = plan_rice(p, res, bs, order, salt);
            tr[4] = rp.order;
            write_residual(p, bw, rp, res, bs, order);
            return;
        }
    }
    // fixed predictor
    if (!fgc::fixed_allowed(eff, acc_bits)) { verbatim(); return; }
    std::vector<int64_t> res(bs, 0);
    bool ok = true;
    for (uint32_t i = order; i < bs && ok; ++i) {
        int64_t acc = 0;
        for (uint32_t j = 0; j < order; ++j) acc += (int64_t)fgc::fixed_coef(order, j) * s[i - 1 - j];
        const int64_t r = s[i] - acc;
        if (r > 0x3FFFFFFFll || r < -0x40000000ll) ok = false;
        res[i] = r;
    }
    if (!ok) { verbatim(); return; }
    header(8 + order);
    for (uint32_t i = 0; i < order; ++i) bw.put_signed(s[i], (int)eff);
    const RicePlan rp = plan_rice(p, res, bs, order, salt);
    tr[4] = rp.order;
    write_residual(p, bw, rp, res, bs, order);
}

// One whole frame -> bytes appended to `out`; PCM (container-typed, MSB aligned) -> pcm if non-null.
void encode_frame(const fg_params &p, const FramePlan &f, std::vector<uint8_t> &out, uint8_t *pcm_base,
                  std::vector<std::vector<int64_t>> &ch) {
    synth_frame(p, f, ch);
    const uint32_t C = p.channels, bs = f.block_size;
    const uint32_t cb = container_of(p.bps);
    if (pcm_base) {
        const uint32_t msb = fgc::msb_shift_of(p.bps);
        uint8_t *dst = pcm_base + f.first_sample * C * cb;
        for (uint32_t i = 0; i < bs; ++i)
            for (uint32_t c = 0; c < C; ++c) {
                const int64_t v = ch[c][i] * ((int64_t)1 << msb);
                if (cb == 1) ((int8_t *)dst)[(size_t)i * C + c] = (int8_t)v;
                else if (cb == 2) ((int16_t *)dst)[(size_t)i * C + c] = (int16_t)v;
                else ((int32_t *)dst)[(size_t)i * C + c] = (int32_t)v;
            }
    }
    // channel assignment
    const uint32_t mode = fgc::stereo_mode_of(p, f.index);
    std::vector<int64_t> a, b;
    uint32_t side_channel = 99;
    if (mode == 1) {  // left/side
        b.resize(bs);
        for (uint32_t i = 0; i < bs; ++i) b[i] = ch[0][i] - ch[1][i];
        side_channel = 1;
    } else if (mode == 2) {  // side/right
        a.resize(bs);
        for (uint32_t i = 0; i < bs; ++i) a[i] = ch[0][i] - ch[1][i];
        side_channel = 0;
    } else if (mode == 3) {  // mid/side
        a.resize(bs); b.resize(bs);
        for (uint32_t i = 0; i < bs; ++i) {
            a[i] = (ch[0][i] + ch[1][i]) >> 1;
            b[i] = ch[0][i] - ch[1][i];
        }
        side_channel = 1;
    }
    const size_t start = out.size();
    // header
    const fgc::FrameHeaderBytes hdr = fgc::frame_header(p, f.index, f.first_sample, bs);
    out.insert(out.end(), hdr.b, hdr.b + hdr.n);
    // subframes
    BitWriter bw(out);
    const uint32_t acc_bits = cb == 1 ? 16 : cb == 2 ? 32 : 64;
    for (uint32_t c = 0; c < C; ++c) {
        const std::vector<int64_t> *sig = &ch[c];
        if (mode == 1 && c == 1) sig = &b;
        if (mode == 2 && c == 0) sig = &a;
        if (mode == 3) sig = c == 0 ? &a : &b;
        const bool side = c == side_channel;
        encode_subframe(p, bw, *sig, bs, p.bps + (side ? 1 : 0), !side, acc_bits, fgc::subframe_salt(p, f.index, c),
                        (g_trace && f.index * C + c < g_trace_records) ? g_trace + 8 * (f.index * C + c) : nullptr);
    }
    bw.align();
    const uint16_t crc = crc16(out.data() + start, out.size() - start);
    out.push_back((uint8_t)(crc >> 8));
    out.push_back((uint8_t)crc);
}

int generate(const fg_params &p, int want_pcm, fg_result *out, uint32_t threads) {
    memset(out, 0, sizeof(*out));
    if (p.channels < 1 || p.channels > 8 || p.bps < 4 || p.bps > 32 || p.n_samples == 0) return 1;
    if (!p.variable_blocksize && (p.block_size < 1 || p.block_size > 65535)) return 1;
    if (p.lpc_order > 32 || p.qlp_precision < 1 || p.qlp_precision > 15) return 1;
    const std::vector<FramePlan> plan = plan_frames(p);
    const size_t nf = plan.size();
    const uint32_t cb = container_of(p.bps);
    const uint64_t pcm_bytes = p.n_samples * p.channels * cb;
    uint8_t *pcm = (uint8_t *)aligned_alloc(64, (pcm_bytes + 63) & ~63ull);
    if (!pcm) return 2;
    if (threads == 0) threads = std::max(1u, std::thread::hardware_concurrency());
    threads = (uint32_t)std::min<size_t>(threads, std::max<size_t>(1, nf / 4));
    // frames are encoded in contiguous chunks, several per thread, stitched in order afterwards
    const size_t n_chunks = std::min<size_t>(nf, (size_t)threads * 8);
    std::vector<std::vector<uint8_t>> chunk_bytes(n_chunks);
    std::vector<uint32_t> frame_len(nf);
    std::atomic<size_t> next{0};
    auto worker = [&]() {
        std::vector<std::vector<int64_t>> ch;
        for (;;) {
            const size_t k = next.fetch_add(1);
            if (k >= n_chunks) break;
            const size_t f0 = nf * k / n_chunks, f1 = nf * (k + 1) / n_chunks;
            std::vector<uint8_t> &o = chunk_bytes[k];
            o.reserve((f1 - f0) * (size_t)plan[f0].block_size * p.channels * ((p.bps + 7) / 8) / 2 + 1024);
            for (size_t f = f0; f < f1; ++f) {
                const size_t before = o.size();
                encode_frame(p, plan[f], o, pcm, ch);
                frame_len[f] = (uint32_t)(o.size() - before);
            }
        }
    };
    const auto t_a = std::chrono::steady_clock::now();
    if (threads <= 1) worker();
    else {
        std::vector<std::thread> ts;
        for (uint32_t t = 0; t < threads; ++t) ts.emplace_back(worker);
        for (auto &t : ts) t.join();
    }
    const auto t_b = std::chrono::steady_clock::now();
    // MD5 over the unshifted little-endian samples, bytes per sample = ceil(bps/8) (zflac.zig:267-277, Q4)
    Md5 md5;
    {
        const uint32_t hb = (p.bps + 7) / 8;
        const uint32_t msb = fgc::msb_shift_of(p.bps);
        const uint64_t n = p.n_samples * p.channels;
        if (hb == cb && msb == 0) md5.update(pcm, pcm_bytes);
        else {
            std::vector<uint8_t> tmp;
            tmp.reserve(1 << 20);
            for (uint64_t i = 0; i < n; ++i) {
                int32_t v = cb == 1 ? ((int8_t *)pcm)[i] : cb == 2 ? ((int16_t *)pcm)[i] : ((int32_t *)pcm)[i];
                v >>= msb;
                for (uint32_t k = 0; k < hb; ++k) tmp.push_back((uint8_t)((uint32_t)v >> (8 * k)));
                if (tmp.size() >= (1 << 20) - 8) { md5.update(tmp.data(), tmp.size()); tmp.clear(); }
            }
            md5.update(tmp.data(), tmp.size());
        }
        md5.finish(out->md5);
    }
    const auto t_c = std::chrono::steady_clock::now();
    uint64_t audio = 0;
    uint32_t min_f = 0xFFFFFF, max_f = 0, min_b = 65535, max_b = 0;
    for (size_t f = 0; f < nf; ++f) {
        audio += frame_len[f];
        min_f = std::min(min_f, frame_len[f]);
        max_f = std::max(max_f, frame_len[f]);
        max_b = std::max(max_b, plan[f].block_size);
        if (f + 1 < nf || nf == 1) min_b = std::min(min_b, plan[f].block_size);
    }
    if (max_f > 0xFFFFFF) max_f = 0;
    const uint64_t total_bytes = 4 + 4 + 34 + audio;
    uint8_t *fl = (uint8_t *)aligned_alloc(64, (total_bytes + 64 + 63) & ~63ull);
    if (!fl) { free(pcm); return 2; }
    uint8_t *w = fl;
    memcpy(w, "fLaC", 4); w += 4;
    *w++ = 0x80; *w++ = 0; *w++ = 0; *w++ = 34;  // last block, STREAMINFO, 34 bytes
    *w++ = (uint8_t)(min_b >> 8); *w++ = (uint8_t)min_b;
    *w++ = (uint8_t)(max_b >> 8); *w++ = (uint8_t)max_b;
    *w++ = (uint8_t)(min_f >> 16); *w++ = (uint8_t)(min_f >> 8); *w++ = (uint8_t)min_f;
    *w++ = (uint8_t)(max_f >> 16); *w++ = (uint8_t)(max_f >> 8); *w++ = (uint8_t)max_f;
    const uint64_t total = p.set_total ? p.n_samples : 0;
    const uint64_t packed = ((uint64_t)p.sample_rate << 44) | ((uint64_t)(p.channels - 1) << 41) | ((uint64_t)(p.bps - 1) << 36) | (total & 0xFFFFFFFFFull);
    for (int i = 7; i >= 0; --i) *w++ = (uint8_t)(packed >> (8 * i));
    memcpy(w, out->md5, 16); w += 16;
    for (auto &cbv : chunk_bytes) {
        memcpy(w, cbv.data(), cbv.size());
        w += cbv.size();
        std::vector<uint8_t>().swap(cbv);
    }
    memset(w, 0, 64);
    if (getenv("FLACGEN_TIMING")) {
        const auto t_d = std::chrono::steady_clock::now();
        fprintf(stderr, "flacgen: encode %.3f s (%u threads), md5 %.3f s, stitch %.3f s\n", std::chrono::duration<double>(t_b - t_a).count(), threads,
                std::chrono::duration<double>(t_c - t_b).count(), std::chrono::duration<double>(t_d - t_c).count());
    }
    out->flac = fl;
    out->flac_bytes = total_bytes;
    out->n_frames = nf;
    out->audio_offset = 42;
    out->container_bytes = cb;
    out->pcm_bytes = pcm_bytes;
    if (want_pcm) out->pcm = pcm;
    else { free(pcm); out->pcm = nullptr; }
    return 0;
}

}  // namespace

extern "C" {

int fg_generate(const fg_params *p, int want_pcm, fg_result *out) {
    if (!p || !out || p->struct_size != sizeof(fg_params)) return 1;
    return generate(*p, want_pcm, out, p->n_threads);
}

int fg_generate_batch(const fg_params *ps, uint32_t n, int want_pcm, fg_result *outs, uint32_t n_threads) {
    if (!ps || !outs) return 1;
    if (n_threads == 0) n_threads = std::max(1u, std::thread::hardware_concurrency());
    std::atomic<uint32_t> next{0};
    std::atomic<int> rc{0};
    auto worker = [&]() {
        for (;;) {
            const uint32_t i = next.fetch_add(1);
            if (i >= n) break;
            const int e = (ps[i].struct_size == sizeof(fg_params)) ? generate(ps[i], want_pcm, &outs[i], 1) : 1;
            if (e) rc = e;
        }
    };
    std::vector<std::thread> ts;
    for (uint32_t t = 0; t < std::min(n_threads, std::max(1u, n)); ++t) ts.emplace_back(worker);
    for (auto &t : ts) t.join();
    return rc;
}

// Debug: subframe records of the NEXT single-stream fg_generate call go to buf (8 words per subframe, frame-major);
// nullptr switches the trace off.
void fg_set_trace(uint32_t *buf, uint64_t n_records) {
    g_trace = buf;
    g_trace_records = buf ? n_records : 0;
}

void fg_free(fg_result *r) {
    if (!r) return;
    free(r->flac);
    free(r->pcm);
    r->flac = nullptr;
    r->pcm = nullptr;
}

}  // extern "C"
