// flacgen_cuda.cu — the synthetic FLAC generator of tools/flacgen.cpp as CUDA kernels.
// TEST / BENCH INFRASTRUCTURE, not product code: it builds the inputs of BASELINE.json's synthetic configs in HBM so
// that bench.py and the full-size tests do not spend seconds of every GPU lease encoding on the host.
//
// Contract: for the same fg_params, the same bytes as flacgen.cpp (stream AND PCM), bit for bit —
// tests/test_gpu_flacgen.py holds it to that for every preset. The reference has no encoder; what is inverted here is
// its decoder: Rice partitions /root/reference/src/zflac.zig:614-666, LPC :499-541, FIXED :462-497, mid/side :553-578,
// the MD5 packing :267-277.
//
// Pipeline (a batch of streams is one flat list of frames; every kernel is over frames or subframes):
//   host      frame plan, per-frame tone phases and Hann windows (libm calls: flacgen_common.h says why)
//   k_synth     one THREAD per frame: the sample recurrences (RNG, phasors, AR(2) noise) are serial per frame
//   k_channels  one block per frame: interleaved PCM out, stereo decorrelation, wasted bits, CONSTANT detection,
//               the predictor choice
//   k_lpc       one block per LPC subframe: windowed autocorrelation (one lag per thread, summed in index order —
//               the order is part of the contract), Levinson-Durbin and quantisation by one thread
//   k_residual  one block per subframe: residual, then the Rice plan (every partition order, exact cost of k and k+1)
//   host      frame lengths -> byte offsets, STREAMINFO
//   k_write     one warp per subframe: every lane owns a run of eight residuals; code lengths, a warp scan, then each
//               lane writes its bits (whole words it owns by plain stores, shared edge words by atomicOr)
//   k_crc16     one thread per frame
//   host      MD5 of the PCM (a serial chain: ~0.8 GB/s on one core, the floor of the whole generator)
//
// Build (tools/flacgen.py): nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -fmad=false
//                           -Xcompiler -fPIC,-ffp-contract=off -shared
#include <cuda_runtime.h>

#include <algorithm>
#include <atomic>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <map>
#include <string>
#include <thread>
#include <vector>

#include "flacgen_common.h"

namespace {

using namespace fgc;

struct StreamDev {
    fg_params p;
    SignalConsts k;
    ToneStep step[8];
    uint64_t pcm_off;      // byte offset of the stream's PCM in the PCM blob
    uint32_t cb, msb, acc_bits, pad;
};
struct FrameJob {
    uint64_t first_sample, index;
    uint64_t sig_base;     // element offset of the frame's planar block: channel c at sig_base + c * chs
    uint64_t out_byte;     // byte offset of the frame in the FLAC blob (known after k_residual)
    double sigma;
    uint32_t stream, block_size, chs, win_off;
    uint32_t phase_off;    // index of ToneState[channels] of this frame
    uint32_t sub0;         // index of the SubInfo of channel 0
    uint32_t frame_bytes, pad;
};
struct SubInfo {
    uint32_t type;         // subframe header code: 0 CONSTANT, 1 VERBATIM, 8 + order FIXED, 31 + order LPC
    uint32_t kind, order, wasted, eff, prec, shift, po;
    uint64_t bits;         // of the whole subframe
    uint32_t frame, channel;
    int32_t qc[32];
};

constexpr int kMaxCh = 8;

// ------------------------------------------------------------------------------------------------ k_synth
__global__ void k_synth(const StreamDev *streams, const FrameJob *frames, uint32_t n_frames, const ToneState *phases,
                        int32_t *sig, double *base, uint32_t base_rows) {
    const uint32_t f = blockIdx.x * blockDim.x + threadIdx.x;
    if (f >= n_frames) return;
    const FrameJob fr = frames[f];
    const StreamDev &sd = streams[fr.stream];
    const fg_params &p = sd.p;
    const uint32_t C = p.channels, n = fr.block_size;
    Xo rng(frame_rng_seed(p.seed, fr.index));
    const uint32_t wasted = wasted_of_frame(p, fr.index);
    const uint32_t cconst = constant_channel_of(p, fr.index);
    // the first channel's unclamped samples, for the channel that follows it: [warp][i][lane]
    double *my_base = base ? base + ((size_t)(f >> 5) * base_rows * 32 + (f & 31)) : nullptr;
    for (uint32_t c = 0; c < C; ++c) {
        const bool follow = (C == 2 && c == 1);
        const ToneStep st = sd.step[c];
        ChannelSynth cs;
        cs.init(phases[fr.phase_off + c], fr.sigma, follow);
        int32_t *dst = sig + fr.sig_base + (size_t)c * fr.chs;
        int64_t first = 0;
        int32_t q4[4];
        for (uint32_t i = 0; i < n; ++i) {
            const double b = follow ? my_base[(size_t)i * 32] : 0.0;
            const double v = cs.next(rng, st, sd.k, follow, b);
            if (c == 0 && C == 2) my_base[(size_t)i * 32] = v;
            int64_t q = quantize_sample(v, sd.k.limit);
            if (i == 0) first = q;
            if (c == cconst) q = first;
            if (wasted) q = (q >> wasted) * ((int64_t)1 << wasted);
            q4[i & 3] = (int32_t)q;
            if ((i & 3) == 3) *reinterpret_cast<int4 *>(dst + (i & ~3u)) = make_int4(q4[0], q4[1], q4[2], q4[3]);
        }
        if (n & 3) {
            for (uint32_t j = n & 3; j < 4; ++j) q4[j] = 0;
            *reinterpret_cast<int4 *>(dst + (n & ~3u)) = make_int4(q4[0], q4[1], q4[2], q4[3]);
        }
    }
}

// ------------------------------------------------------------------------------------------------ k_channels
__device__ __forceinline__ uint32_t sub_prologue_bits(const SubInfo &s) {  // header byte + the unary wasted-bits count
    return 8u + s.wasted;
}

__global__ void __launch_bounds__(256) k_channels(const StreamDev *streams, const FrameJob *frames, const int32_t *sig,
                                                  int32_t *sub, uint8_t *pcm, SubInfo *subs) {
    const uint32_t f = blockIdx.x;
    const FrameJob &fr = frames[f];
    const StreamDev &sd = streams[fr.stream];
    const fg_params &p = sd.p;
    const uint32_t C = p.channels, n = fr.block_size, chs = fr.chs;
    const uint32_t mode = stereo_mode_of(p, fr.index);
    const int32_t *src = sig + fr.sig_base;
    int32_t *dst = sub + fr.sig_base;
    uint8_t *out = pcm + sd.pcm_off + fr.first_sample * C * sd.cb;
    __shared__ uint32_t s_or[kMaxCh], s_diff[kMaxCh];
    if (threadIdx.x < kMaxCh) {
        s_or[threadIdx.x] = 0;
        s_diff[threadIdx.x] = 0;
    }
    __syncthreads();
    int32_t first[kMaxCh];
    uint32_t orv[kMaxCh], diff[kMaxCh];
#pragma unroll
    for (int c = 0; c < kMaxCh; ++c) {
        orv[c] = 0;
        diff[c] = 0;
        first[c] = 0;
    }
    auto transform = [&](int32_t *v) {  // channel samples -> subframe signals
        if (mode == 1) v[1] = v[0] - v[1];                    // left / side
        else if (mode == 2) v[0] = v[0] - v[1];               // side / right
        else if (mode == 3) {                                 // mid / side
            const int32_t m = (v[0] + v[1]) >> 1, s = v[0] - v[1];
            v[0] = m;
            v[1] = s;
        }
    };
    {
        int32_t v[kMaxCh];
#pragma unroll
        for (int c = 0; c < kMaxCh; ++c) v[c] = c < (int)C ? src[(size_t)c * chs] : 0;
        transform(v);
#pragma unroll
        for (int c = 0; c < kMaxCh; ++c) first[c] = v[c];
    }
    for (uint32_t i = threadIdx.x; i < n; i += blockDim.x) {
        int32_t v[kMaxCh];
#pragma unroll
        for (int c = 0; c < kMaxCh; ++c) v[c] = c < (int)C ? src[(size_t)c * chs + i] : 0;
#pragma unroll
        for (int c = 0; c < kMaxCh; ++c)
            if (c < (int)C) {
                const int64_t w = (int64_t)v[c] * ((int64_t)1 << sd.msb);
                const size_t o = (size_t)i * C + c;
                if (sd.cb == 1) reinterpret_cast<int8_t *>(out)[o] = (int8_t)w;
                else if (sd.cb == 2) reinterpret_cast<int16_t *>(out)[o] = (int16_t)w;
                else reinterpret_cast<int32_t *>(out)[o] = (int32_t)w;
            }
        transform(v);
#pragma unroll
        for (int c = 0; c < kMaxCh; ++c)
            if (c < (int)C) {
                dst[(size_t)c * chs + i] = v[c];
                orv[c] |= (uint32_t)v[c];
                diff[c] |= (uint32_t)(v[c] != first[c]);
            }
    }
#pragma unroll
    for (int c = 0; c < kMaxCh; ++c)
        if (c < (int)C) {
            const uint32_t o = __reduce_or_sync(0xFFFFFFFFu, orv[c]), d = __reduce_or_sync(0xFFFFFFFFu, diff[c]);
            if ((threadIdx.x & 31) == 0) {
                if (o) atomicOr(&s_or[c], o);
                if (d) atomicOr(&s_diff[c], d);
            }
        }
    __syncthreads();
    __shared__ uint32_t s_wasted[kMaxCh];
    if (threadIdx.x < C) {
        const uint32_t c = threadIdx.x;
        const uint32_t side = (mode == 1 || mode == 3) ? 1u : mode == 2 ? 0u : 99u;
        const uint32_t depth = p.bps + (c == side ? 1 : 0);
        uint32_t wasted = 0;
        if (s_or[c] != 0) wasted = (uint32_t)__ffs((int)s_or[c]) - 1;
        if (wasted >= depth) wasted = 0;
        s_wasted[c] = wasted;
        SubInfo si;
        memset(&si, 0, sizeof(si));
        si.frame = f;
        si.channel = c;
        si.wasted = wasted;
        si.eff = depth - wasted;
        const bool constant = s_diff[c] == 0;
        if (constant && c != side) {  // the reference reads bps - wasted bits here, which equals eff only off the side channel (Q8)
            si.kind = kConstant;
            si.type = 0;
            si.bits = sub_prologue_bits(si) + si.eff;
        } else {
            Choice ch = choose_predictor(p, subframe_salt(p, fr.index, c), n, si.eff, sd.acc_bits);
            if (ch.kind == kFixed && !fixed_allowed(si.eff, sd.acc_bits)) ch.kind = kVerbatim;
            si.kind = ch.kind;
            si.order = ch.kind == kVerbatim ? 0 : ch.order;
            si.prec = ch.prec;
            si.type = 1;  // k_lpc / k_residual settle the rest
            si.bits = sub_prologue_bits(si) + (uint64_t)n * si.eff;
        }
        subs[fr.sub0 + c] = si;
    }
    __syncthreads();
#pragma unroll 1
    for (uint32_t c = 0; c < C; ++c) {
        const uint32_t w = s_wasted[c];
        if (!w) continue;
        for (uint32_t i = threadIdx.x; i < n; i += blockDim.x) dst[(size_t)c * chs + i] >>= w;
    }
}

// ------------------------------------------------------------------------------------------------ k_lpc
// Autocorrelation of the Hann-windowed subframe, one lag per thread, every sum in index order (flacgen.cpp's fit_lpc:
// a[lag] = sum_{i = lag}^{n-1} w[i] * w[i - lag]); then Levinson-Durbin and the quantisation by thread 0 — the shared
// code of flacgen_common.h.
constexpr int kLpcTile = 256, kLpcHist = 32, kLpcThreads = 64;

__global__ void __launch_bounds__(kLpcThreads) k_lpc(const StreamDev *streams, const FrameJob *frames, const int32_t *sub,
                                                      const double *windows, SubInfo *subs) {
    SubInfo &si = subs[blockIdx.x];
    if (si.kind != kLpc) return;
    const FrameJob &fr = frames[si.frame];
    const StreamDev &sd = streams[fr.stream];
    const uint32_t n = fr.block_size, order = si.order, lag = threadIdx.x;
    const int32_t *s = sub + fr.sig_base + (size_t)si.channel * fr.chs;
    const double *win = windows + fr.win_off;
    __shared__ double sw[kLpcHist + kLpcTile];
    __shared__ double s_ac[33];
    double a = 0;
    for (uint32_t t0 = 0; t0 < n; t0 += kLpcTile) {
        for (uint32_t j = threadIdx.x; j < kLpcHist + kLpcTile; j += kLpcThreads) {
            const int64_t i = (int64_t)t0 - kLpcHist + j;
            sw[j] = (i >= 0 && i < (int64_t)n) ? __dmul_rn(win[i], (double)s[i]) : 0.0;
        }
        __syncthreads();
        if (lag <= order) {
            const uint32_t end = min(n - t0, (uint32_t)kLpcTile);
            uint32_t j = lag > t0 ? lag - t0 : 0;  // i = t0 + j >= lag
            for (; j < end; ++j) a = __dadd_rn(a, __dmul_rn(sw[kLpcHist + j], sw[kLpcHist + j - lag]));
        }
        __syncthreads();
    }
    if (lag <= order) s_ac[lag] = a;
    __syncthreads();
    if (threadIdx.x == 0) {
        double ac[33];
        for (uint32_t i = 0; i <= order; ++i) ac[i] = s_ac[i];
        int32_t qc[32];
        for (int i = 0; i < 32; ++i) qc[i] = 0;
        uint32_t shift = 0;
        if (lpc_from_autocorr(ac, order, si.prec, qc, shift)) {
            si.shift = shift;
            for (int i = 0; i < 32; ++i) si.qc[i] = qc[i];
        } else {  // no usable fit: FIXED min(2, order), VERBATIM if the reference's accumulator could not hold that either
            si.kind = fixed_allowed(si.eff, sd.acc_bits) ? (uint32_t)kFixed : (uint32_t)kVerbatim;
            si.order = si.kind == kFixed ? min(2u, order) : 0u;
        }
    }
}

// ------------------------------------------------------------------------------------------------ k_residual
// Residual of a FIXED / LPC subframe, then the cheapest Rice partitioning (flacgen.cpp's plan_rice): every partition
// order of the allowed range, per partition k ~ log2(mean) and k + 1 costed exactly, escapes where forced or k too large.
// plan[] of a subframe: kPlanStride entries; the partitions of order po at [2^po, 2^(po+1)): low byte the Rice
// parameter or 0xFF = escaped, high byte the escape width.
constexpr int kResThreads = 256;

__device__ __forceinline__ uint32_t zz32(int32_t r) { return ((uint32_t)r << 1) ^ (uint32_t)(r >> 31); }

__global__ void __launch_bounds__(kResThreads) k_residual(const StreamDev *streams, const FrameJob *frames, const int32_t *sub,
                                                           int32_t *res, uint16_t *plan, uint32_t plan_stride, SubInfo *subs) {
    SubInfo &si = subs[blockIdx.x];
    if (si.kind != kFixed && si.kind != kLpc) return;  // CONSTANT / VERBATIM: k_channels has set type and bits
    const FrameJob &fr = frames[si.frame];
    const fg_params &p = streams[fr.stream].p;
    const uint32_t n = fr.block_size, order = si.order;
    const int32_t *s = sub + fr.sig_base + (size_t)si.channel * fr.chs;
    int32_t *r = res + fr.sig_base + (size_t)si.channel * fr.chs;
    __shared__ int32_t s_qc[32];
    __shared__ unsigned long long s_level_bits, s_best_bits;
    __shared__ uint32_t s_best_po;
    if (threadIdx.x < 32) s_qc[threadIdx.x] = si.kind == kLpc ? si.qc[threadIdx.x] : (threadIdx.x < 4 ? fixed_coef(order, threadIdx.x) : 0);
    __syncthreads();
    const uint32_t shift = si.kind == kLpc ? si.shift : 0;
    int bad = 0;
    for (uint32_t i = order + threadIdx.x; i < n; i += kResThreads) {
        int64_t acc = 0;
        for (uint32_t j = 0; j < order; ++j) acc += (int64_t)s_qc[j] * s[i - 1 - j];
        const int64_t v = (int64_t)s[i] - (acc >> shift);
        if (v > 0x3FFFFFFFll || v < -0x40000000ll) bad = 1;
        r[i] = (int32_t)v;
    }
    if (__syncthreads_or(bad)) {  // a residual the format cannot carry: VERBATIM
        if (threadIdx.x == 0) {
            si.kind = kVerbatim;
            si.order = 0;
            si.type = 1;
            si.bits = sub_prologue_bits(si) + (uint64_t)n * si.eff;
        }
        return;
    }
    const uint64_t salt = subframe_salt(p, fr.index, si.channel);
    const uint32_t kmax = rice_kmax(p), pbits = rice_pbits(p);
    uint32_t lo, hi;
    partition_order_range(p, n, order, lo, hi);
    uint16_t *my_plan = plan + (size_t)blockIdx.x * plan_stride;
    const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) {
        s_best_bits = ~0ull;
        s_best_po = lo;
    }
    for (uint32_t po = lo; po <= hi; ++po) {
        if (!partition_order_valid(n, po, order)) break;
        if (threadIdx.x == 0) s_level_bits = 0;
        __syncthreads();
        const uint32_t np = 1u << po, psz = n >> po;
        for (uint32_t q = warp; q < np; q += kResThreads / 32) {
            const uint32_t a = q == 0 ? order : q * psz, b = (q + 1) * psz, cnt = b - a;
            unsigned long long sum = 0;
            int32_t mn = 0, mx = 0;
            for (uint32_t i = a + lane; i < b; i += 32) {
                const int32_t v = r[i];
                sum += zz32(v);
                mn = min(mn, v);
                mx = max(mx, v);
            }
            for (int o = 16; o; o >>= 1) sum += __shfl_xor_sync(0xFFFFFFFFu, sum, o);
            mn = __reduce_min_sync(0xFFFFFFFFu, mn);
            mx = __reduce_max_sync(0xFFFFFFFFu, mx);
            uint32_t k = rice_k_of_mean(sum, cnt);
            unsigned long long bits;
            uint32_t entry;
            if (k > kmax || partition_forced_escape(p, salt, q)) {
                const uint32_t eb = (mn == 0 && mx == 0) ? 0u : (uint32_t)bits_needed_signed(mn, mx);
                bits = pbits + 5 + (unsigned long long)eb * cnt;
                entry = 0xFFu | (eb << 8);
            } else {  // exact cost for k and k + 1, keep the cheaper
                const uint32_t k0 = min(kmax, k), k1 = min(kmax, k + 1);
                unsigned long long b0 = 0, b1 = 0;
                for (uint32_t i = a + lane; i < b; i += 32) {
                    const uint32_t z = zz32(r[i]);
                    b0 += (z >> k0) + 1 + k0;
                    b1 += (z >> k1) + 1 + k1;
                }
                for (int o = 16; o; o >>= 1) {
                    b0 += __shfl_xor_sync(0xFFFFFFFFu, b0, o);
                    b1 += __shfl_xor_sync(0xFFFFFFFFu, b1, o);
                }
                if (b1 < b0 && k + 1 <= kmax) {
                    k += 1;
                    bits = b1;
                } else bits = b0;
                bits += pbits;
                entry = k;
            }
            if (lane == 0) {
                my_plan[np + q] = (uint16_t)entry;
                atomicAdd(&s_level_bits, bits);
            }
        }
        __syncthreads();
        if (threadIdx.x == 0 && s_level_bits < s_best_bits) {
            s_best_bits = s_level_bits;
            s_best_po = po;
        }
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        si.po = s_best_po;
        si.type = si.kind == kLpc ? 32 + order - 1 : 8 + order;
        si.bits = sub_prologue_bits(si) + (uint64_t)order * si.eff + (si.kind == kLpc ? 4 + 5 + (uint64_t)order * si.prec : 0) + 2 + 4 +
                  s_best_bits;
    }
}

// frame_bytes of every frame: header + subframes (padded to a byte) + CRC-16
__global__ void k_frame_len(const StreamDev *streams, FrameJob *frames, uint32_t n_frames, const SubInfo *subs, uint32_t *frame_len) {
    const uint32_t f = blockIdx.x * blockDim.x + threadIdx.x;
    if (f >= n_frames) return;
    FrameJob &fr = frames[f];
    const fg_params &p = streams[fr.stream].p;
    uint64_t bits = 0;
    for (uint32_t c = 0; c < p.channels; ++c) bits += subs[fr.sub0 + c].bits;
    const uint32_t hdr = frame_header(p, fr.index, fr.first_sample, fr.block_size).n;
    fr.frame_bytes = hdr + (uint32_t)((bits + 7) >> 3) + 2;
    frame_len[f] = fr.frame_bytes;
}

// ------------------------------------------------------------------------------------------------ k_write
// MSB-first bit writer of one lane into a zeroed blob. The lane owns the bit range [own_lo, own_hi): a 32-bit word that
// lies wholly inside it is stored plainly, a word it shares with a neighbour (another lane, subframe or frame) is OR-ed
// in atomically. Zero bits are never written (unary runs cost nothing).
struct Emitter {
    uint32_t *out;
    uint64_t pos, own_lo, own_hi;
    uint32_t cur;
    __device__ __forceinline__ void begin(uint32_t *blob, uint64_t bit_lo, uint64_t bit_hi) {
        out = blob;
        pos = own_lo = bit_lo;
        own_hi = bit_hi;
        cur = 0;
    }
    __device__ __forceinline__ void flush(uint64_t word) {
        if (cur) {
            const uint32_t w = __byte_perm(cur, 0, 0x0123);  // the stream is big-endian
            if (word * 32 >= own_lo && word * 32 + 32 <= own_hi) out[word] = w;
            else atomicOr(out + word, w);
        }
        cur = 0;
    }
    __device__ __forceinline__ void put(uint32_t v, uint32_t nbits) {  // nbits in [0, 32]
        if (!nbits) return;
        if (nbits < 32) v &= (1u << nbits) - 1u;
        const uint32_t room = 32u - (uint32_t)(pos & 31);
        if (nbits < room) {
            cur |= v << (room - nbits);
        } else {
            cur |= v >> (nbits - room);
            flush(pos >> 5);
            const uint32_t rest = nbits - room;
            cur = rest ? v << (32u - rest) : 0u;
        }
        pos += nbits;
    }
    __device__ __forceinline__ void skip(uint64_t zeros) {
        const uint32_t room = 32u - (uint32_t)(pos & 31);
        if (zeros >= room) flush(pos >> 5);
        pos += zeros;
    }
    __device__ __forceinline__ void finish() {
        if (pos & 31) flush(pos >> 5);
    }
};

constexpr int kWriteWarps = 8, kRun = 8;

__global__ void __launch_bounds__(kWriteWarps * 32) k_write(const StreamDev *streams, const FrameJob *frames, const int32_t *sub,
                                                             const int32_t *res, const uint16_t *plan, uint32_t plan_stride,
                                                             const SubInfo *subs, uint32_t n_subs, uint32_t *blob) {
    const uint32_t sidx = blockIdx.x * kWriteWarps + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (sidx >= n_subs) return;
    const SubInfo &si = subs[sidx];
    const FrameJob &fr = frames[si.frame];
    const fg_params &p = streams[fr.stream].p;
    const uint32_t n = fr.block_size, order = si.order, eff = si.eff, pbits = rice_pbits(p);
    const int32_t *s = sub + fr.sig_base + (size_t)si.channel * fr.chs;
    const int32_t *r = res + fr.sig_base + (size_t)si.channel * fr.chs;
    const uint16_t *my_plan = plan + (size_t)sidx * plan_stride;
    const bool coded = si.kind == kFixed || si.kind == kLpc;
    const uint32_t np = 1u << si.po, psz = coded ? n >> si.po : n;
    const FrameHeaderBytes hdr = frame_header(p, fr.index, fr.first_sample, n);
    uint64_t bit0 = (fr.out_byte + hdr.n) * 8;
    for (uint32_t c = 0; c < si.channel; ++c) bit0 += subs[fr.sub0 + c].bits;
    // what lane 0 writes ahead of the per-sample part
    uint64_t prologue = 8 + si.wasted;
    if (si.kind == kConstant) prologue += eff;
    if (coded) {
        prologue += (uint64_t)order * eff + (si.kind == kLpc ? 9 + (uint64_t)order * si.prec : 0) + 6;
        prologue += pbits + ((my_plan[np] & 0xFF) == 0xFF ? 5 : 0);
    }
    if (lane == 0) {
        Emitter em;
        if (si.channel == 0) {
            em.begin(blob, fr.out_byte * 8, (fr.out_byte + hdr.n) * 8);
            for (uint32_t i = 0; i < hdr.n; ++i) em.put(hdr.b[i], 8);
            em.finish();
        }
        em.begin(blob, bit0, bit0 + prologue);
        em.put(0, 1);
        em.put(si.type, 6);
        em.put(si.wasted ? 1 : 0, 1);
        if (si.wasted) {
            em.skip(si.wasted - 1);
            em.put(1, 1);
        }
        if (si.kind == kConstant) em.put((uint32_t)s[0], eff);
        if (coded) {
            for (uint32_t i = 0; i < order; ++i) em.put((uint32_t)s[i], eff);
            if (si.kind == kLpc) {
                em.put(si.prec - 1, 4);
                em.put(si.shift, 5);
                for (uint32_t j = 0; j < order; ++j) em.put((uint32_t)si.qc[j], si.prec);
            }
            em.put(p.rice_method ? 1 : 0, 2);
            em.put(si.po, 4);
            const uint32_t e = my_plan[np];
            if ((e & 0xFF) == 0xFF) {
                em.put((1u << pbits) - 1, pbits);
                em.put(e >> 8, 5);
            } else em.put(e, pbits);
        }
        em.finish();
    }
    if (si.kind == kConstant) return;
    // the per-sample part: residuals [order, n) of a coded subframe (a partition's parameter goes with its first
    // residual), samples [0, n) of a VERBATIM one
    uint64_t pos_base = bit0 + prologue;
    const uint32_t first = coded ? order : 0;
    for (uint32_t c0 = first; c0 < n; c0 += 32 * kRun) {
        const uint32_t i0 = min(n, c0 + lane * kRun), i1 = min(n, i0 + kRun);
        Emitter em;
        uint64_t start = 0;
        for (int pass = 0; pass < 2; ++pass) {
            uint64_t len = 0;
            // the partition of the residual BEFORE the run: a run that starts on a partition's first residual writes its parameter
            uint32_t q = (coded && i0) ? (i0 - 1) / psz : 0, next_q = coded ? (q + 1) * psz : 0xFFFFFFFFu;
            uint32_t e = coded ? my_plan[np + q] : 0;
            for (uint32_t i = i0; i < i1; ++i) {
                if (!coded) {
                    if (pass) em.put((uint32_t)s[i], eff);
                    else len += eff;
                    continue;
                }
                if (i == next_q) {  // a new partition starts here: its parameter first
                    ++q;
                    next_q += psz;
                    e = my_plan[np + q];
                    if ((e & 0xFF) == 0xFF) {
                        if (pass) {
                            em.put((1u << pbits) - 1, pbits);
                            em.put(e >> 8, 5);
                        } else len += pbits + 5;
                    } else {
                        if (pass) em.put(e, pbits);
                        else len += pbits;
                    }
                }
                const int32_t v = r[i];
                if ((e & 0xFF) == 0xFF) {
                    if (pass) em.put((uint32_t)v, e >> 8);
                    else len += e >> 8;
                } else {
                    const uint32_t k = e, z = zz32(v), qv = z >> k;
                    if (pass) {
                        em.skip(qv);
                        em.put((1u << k) | (z & ((1u << k) - 1u)), k + 1);
                    } else len += qv + 1 + k;
                }
            }
            if (pass == 0) {
                uint64_t incl = len;
                for (int o = 1; o < 32; o <<= 1) {
                    const uint64_t up = __shfl_up_sync(0xFFFFFFFFu, incl, o);
                    if (lane >= (uint32_t)o) incl += up;
                }
                start = pos_base + incl - len;
                pos_base += __shfl_sync(0xFFFFFFFFu, incl, 31);
                em.begin(blob, start, start + len);
            } else em.finish();
        }
    }
}

// ------------------------------------------------------------------------------------------------ k_crc16
__global__ void __launch_bounds__(128) k_crc16(const FrameJob *frames, uint32_t n_frames, const uint16_t *tables, uint8_t *blob) {
    __shared__ uint16_t t[8][256];
    for (uint32_t i = threadIdx.x; i < 8 * 256; i += blockDim.x) (&t[0][0])[i] = tables[i];
    __syncthreads();
    const uint32_t f = blockIdx.x * blockDim.x + threadIdx.x;
    if (f >= n_frames) return;
    const uint8_t *b = blob + frames[f].out_byte;
    const size_t n = frames[f].frame_bytes - 2;
    uint16_t c = 0;
    size_t i = 0;
    for (; i < n && ((uintptr_t)(b + i) & 7); ++i) c = (uint16_t)((c << 8) ^ t[0][(c >> 8) ^ b[i]]);
    for (; i + 8 <= n; i += 8) {  // slice-by-8
        const uint2 w = *reinterpret_cast<const uint2 *>(b + i);
        const uint32_t x = w.x ^ (uint32_t)(c >> 8) ^ ((uint32_t)(c & 0xFF) << 8), y = w.y;
        c = t[7][x & 0xFF] ^ t[6][(x >> 8) & 0xFF] ^ t[5][(x >> 16) & 0xFF] ^ t[4][x >> 24] ^ t[3][y & 0xFF] ^ t[2][(y >> 8) & 0xFF] ^
            t[1][(y >> 16) & 0xFF] ^ t[0][y >> 24];
    }
    for (; i < n; ++i) c = (uint16_t)((c << 8) ^ t[0][(c >> 8) ^ b[i]]);
    blob[frames[f].out_byte + n] = (uint8_t)(c >> 8);
    blob[frames[f].out_byte + n + 1] = (uint8_t)c;
}

// debug trace: the same eight words per subframe flacgen.cpp's fg_set_trace writes
__global__ void k_trace(const FrameJob *frames, const int32_t *sub, const SubInfo *subs, uint32_t *trace) {
    const SubInfo &si = subs[blockIdx.x];
    const FrameJob &fr = frames[si.frame];
    const int32_t *s = sub + fr.sig_base + (size_t)si.channel * fr.chs;
    uint32_t sum = 0;
    for (uint32_t i = threadIdx.x; i < fr.block_size; i += blockDim.x) sum += (uint32_t)s[i] * (i + 1);
    __shared__ uint32_t s_sum;
    if (threadIdx.x == 0) s_sum = 0;
    __syncthreads();
    atomicAdd(&s_sum, sum);
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t *t = trace + 8 * (size_t)blockIdx.x;
        t[0] = si.type;
        t[1] = si.wasted;
        t[2] = si.kind == kLpc ? si.prec : 0;
        t[3] = si.kind == kLpc ? si.shift : 0;
        t[4] = (si.kind == kLpc || si.kind == kFixed) ? si.po : 0;
        t[5] = (uint32_t)si.bits;
        t[6] = s_sum;
        uint32_t q = 0;
        if (si.kind == kLpc)
            for (uint32_t j = 0; j < si.order; ++j) q += (uint32_t)si.qc[j] * (j + 1);
        t[7] = q;
    }
}

}  // namespace

// ================================================================================================ host side
namespace {

// MD5 of every stream's PCM by one device thread per stream: a batch of short files hashes in parallel (a single long
// stream does not: see hash_on_host). Message = the samples unshifted, little-endian, ceil(bps / 8) bytes each
// (/root/reference/src/zflac.zig:267-277).
__global__ void k_md5_streams(const StreamDev *streams, uint32_t n_streams, const uint8_t *pcm, uint8_t *md5_out) {
    const uint32_t sidx = blockIdx.x * blockDim.x + threadIdx.x;
    if (sidx >= n_streams) return;
    const StreamDev &sd = streams[sidx];
    const uint32_t hb = (sd.p.bps + 7) / 8;
    const uint64_t n = sd.p.n_samples * sd.p.channels;
    const uint8_t *src = pcm + sd.pcm_off;
    Md5 h;
    uint8_t buf[192];
    uint32_t fill = 0;
    for (uint64_t i = 0; i < n; ++i) {
        int32_t v = sd.cb == 1 ? reinterpret_cast<const int8_t *>(src)[i]
                  : sd.cb == 2 ? reinterpret_cast<const int16_t *>(src)[i] : reinterpret_cast<const int32_t *>(src)[i];
        v >>= sd.msb;
        for (uint32_t k = 0; k < hb; ++k) buf[fill++] = (uint8_t)((uint32_t)v >> (8 * k));
        if (fill >= 128) {
            h.update(buf, 128);
            for (uint32_t j = 128; j < fill; ++j) buf[j - 128] = buf[j];
            fill -= 128;
        }
    }
    h.update(buf, fill);
    h.finish(md5_out + 16 * (size_t)sidx);
}

// the 42 bytes ahead of the first frame ("fLaC" + STREAMINFO), one thread per stream
__global__ void k_put_heads(const uint8_t *heads, const uint64_t *flac_off, uint32_t n_streams, uint8_t *blob) {
    const uint32_t sidx = blockIdx.x * blockDim.x + threadIdx.x;
    if (sidx >= n_streams) return;
    for (int i = 0; i < 42; ++i) blob[flac_off[sidx] + i] = heads[42 * (size_t)sidx + i];
}

thread_local std::string g_err;

struct StreamOut {  // fgc_stream in the C interface below
    uint64_t flac_off, flac_bytes, pcm_off, pcm_bytes, n_frames;
    uint32_t container_bytes, pad;
    uint8_t md5[16];
};

struct Handle {
    int device = 0;
    uint32_t n_streams = 0, n_frames = 0, n_subs = 0;
    std::vector<StreamOut> outs;
    uint8_t *d_flac = nullptr, *d_pcm = nullptr;
    uint64_t flac_bytes = 0, pcm_bytes = 0;
    FrameJob *d_frames = nullptr;  // kept with FGC_KEEP_TRACE only
    SubInfo *d_subs = nullptr;
    int32_t *d_sub = nullptr;
    double t[8] = {0};
};

struct DevMem {  // temporaries: freed when the call ends, whichever way
    std::vector<void *> v;
    template <class T>
    bool get(T **p, size_t count) {
        void *q = nullptr;
        const cudaError_t e = cudaMalloc(&q, std::max<size_t>(count * sizeof(T), 256));
        if (e != cudaSuccess) {
            g_err = std::string("cudaMalloc: ") + cudaGetErrorString(e);
            cudaGetLastError();
            return false;
        }
        v.push_back(q);
        *p = reinterpret_cast<T *>(q);
        return true;
    }
    void *release(void *p) {  // the caller keeps p
        v.erase(std::remove(v.begin(), v.end(), p), v.end());
        return p;
    }
    ~DevMem() {
        for (void *p : v) cudaFree(p);
    }
};

#define FGC_CK(call)                                                                                         \
    do {                                                                                                     \
        const cudaError_t e_ = (call);                                                                       \
        if (e_ != cudaSuccess) {                                                                             \
            g_err = std::string(#call) + ": " + cudaGetErrorString(e_);                                      \
            cudaGetLastError();                                                                              \
            return 3;                                                                                        \
        }                                                                                                    \
    } while (0)

double now_s() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

// MD5 of one stream's PCM on the host: chunks come over through two pinned buffers while the previous one is hashed.
int hash_on_host(const StreamDev &sd, const uint8_t *d_pcm, uint8_t out[16]) {
    constexpr size_t kChunk = 16u << 20;
    uint8_t *stage[2] = {nullptr, nullptr};
    cudaStream_t st;
    cudaEvent_t ev[2];
    FGC_CK(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
    for (int i = 0; i < 2; ++i) {
        FGC_CK(cudaHostAlloc((void **)&stage[i], kChunk, cudaHostAllocDefault));
        FGC_CK(cudaEventCreateWithFlags(&ev[i], cudaEventDisableTiming));
    }
    const uint32_t hb = (sd.p.bps + 7) / 8;
    const uint64_t bytes = sd.p.n_samples * sd.p.channels * sd.cb;
    const bool plain = hb == sd.cb && sd.msb == 0;
    std::vector<uint8_t> packed(plain ? 0 : kChunk);
    Md5 h;
    const uint64_t n_chunks = (bytes + kChunk - 1) / kChunk;
    auto issue = [&](uint64_t k) {
        const size_t off = k * kChunk, len = (size_t)std::min<uint64_t>(kChunk, bytes - off);
        cudaMemcpyAsync(stage[k & 1], d_pcm + sd.pcm_off + off, len, cudaMemcpyDeviceToHost, st);
        cudaEventRecord(ev[k & 1], st);
    };
    if (n_chunks) issue(0);
    for (uint64_t k = 0; k < n_chunks; ++k) {
        FGC_CK(cudaEventSynchronize(ev[k & 1]));
        if (k + 1 < n_chunks) issue(k + 1);
        const size_t len = (size_t)std::min<uint64_t>(kChunk, bytes - k * kChunk);
        if (plain) h.update(stage[k & 1], len);
        else {
            size_t w = 0;
            const size_t ns = len / sd.cb;
            for (size_t i = 0; i < ns; ++i) {
                int32_t v = sd.cb == 1 ? ((int8_t *)stage[k & 1])[i] : sd.cb == 2 ? ((int16_t *)stage[k & 1])[i] : ((int32_t *)stage[k & 1])[i];
                v >>= sd.msb;
                for (uint32_t b = 0; b < hb; ++b) packed[w++] = (uint8_t)((uint32_t)v >> (8 * b));
            }
            h.update(packed.data(), w);
        }
    }
    h.finish(out);
    for (int i = 0; i < 2; ++i) {
        cudaFreeHost(stage[i]);
        cudaEventDestroy(ev[i]);
    }
    cudaStreamDestroy(st);
    return 0;
}

}  // namespace

extern "C" {

typedef struct fgc_stream {
    uint64_t flac_off, flac_bytes;  // in the FLAC blob (streams start at multiples of `align`)
    uint64_t pcm_off, pcm_bytes;    // in the PCM blob: interleaved, container-typed, MSB aligned — what zflac.decode returns
    uint64_t n_frames;
    uint32_t container_bytes, pad;
    uint8_t md5[16];
} fgc_stream;

enum { FGC_KEEP_TRACE = 1 };

const char *fgc_last_error(void) { return g_err.c_str(); }

void fgc_free(void *handle) {
    Handle *h = reinterpret_cast<Handle *>(handle);
    if (!h) return;
    cudaSetDevice(h->device);
    cudaFree(h->d_flac);
    cudaFree(h->d_pcm);
    cudaFree(h->d_frames);
    cudaFree(h->d_subs);
    cudaFree(h->d_sub);
    delete h;
}

// Generates the n streams ps[0..n) on `device`. The results stay in device memory inside *handle until fgc_free.
// Returns 0, 1 (bad parameters), 2 (host memory), 3 (CUDA error: fgc_last_error()).
int fgc_generate(const fg_params *ps, uint32_t n, int device, uint32_t align, uint32_t flags, void **handle) {
    if (!ps || !n || !handle) return 1;
    *handle = nullptr;
    if (align == 0) align = 1;
    const double t_start = now_s();
    FGC_CK(cudaSetDevice(device));
    // ---- plan (host)
    std::vector<StreamDev> streams(n);
    std::vector<FrameJob> frames;
    std::vector<ToneState> phases;
    std::vector<double> windows;
    std::map<uint32_t, uint32_t> win_of;  // block size -> offset in windows
    std::vector<uint32_t> first_frame(n + 1);
    uint64_t sig_total = 0, pcm_total = 0;
    uint32_t n_subs = 0, max_bs = 0, max_po = 0;
    bool any_stereo = false;
    for (uint32_t s = 0; s < n; ++s) {
        const fg_params &p = ps[s];
        if (p.struct_size != sizeof(fg_params)) return 1;
        if (p.channels < 1 || p.channels > 8 || p.bps < 4 || p.bps > 32 || p.n_samples == 0) return 1;
        if (!p.variable_blocksize && (p.block_size < 1 || p.block_size > 65535)) return 1;
        if (p.lpc_order > 32 || p.qlp_precision < 1 || p.qlp_precision > 15) return 1;
        StreamDev &sd = streams[s];
        memset(&sd, 0, sizeof(sd));
        sd.p = p;
        sd.k = signal_consts(p);
        for (uint32_t c = 0; c < p.channels; ++c) sd.step[c] = tone_step(p, c);
        sd.cb = container_of(p.bps);
        sd.msb = msb_shift_of(p.bps);
        sd.acc_bits = sd.cb == 1 ? 16 : sd.cb == 2 ? 32 : 64;
        pcm_total = (pcm_total + 15) & ~15ull;
        sd.pcm_off = pcm_total;
        pcm_total += p.n_samples * p.channels * sd.cb;
        any_stereo |= p.channels == 2;
        first_frame[s] = (uint32_t)frames.size();
        uint64_t pos = 0, idx = 0;
        while (pos < p.n_samples) {
            uint32_t bs = p.variable_blocksize ? var_block_size(idx) : p.block_size;
            if ((uint64_t)bs > p.n_samples - pos) bs = (uint32_t)(p.n_samples - pos);
            FrameJob f;
            memset(&f, 0, sizeof(f));
            f.first_sample = pos;
            f.index = idx;
            f.stream = s;
            f.block_size = bs;
            f.chs = (bs + 3) & ~3u;
            f.sig_base = sig_total;
            sig_total += (uint64_t)f.chs * p.channels;
            f.sigma = frame_sigma(p, sd.k, idx);
            f.sub0 = f.phase_off = n_subs;
            n_subs += p.channels;
            auto it = win_of.find(bs);
            if (it == win_of.end()) {
                it = win_of.emplace(bs, (uint32_t)windows.size()).first;
                for (uint32_t i = 0; i < bs; ++i) windows.push_back(hann(i, bs));
            }
            f.win_off = it->second;
            for (uint32_t c = 0; c < p.channels; ++c) {
                ToneState ts;
                memset(&ts, 0, sizeof(ts));
                if (!(p.channels == 2 && c == 1)) ts = tone_state(p, c, pos);  // the following channel has no tones of its own
                phases.push_back(ts);
            }
            max_bs = std::max(max_bs, bs);
            uint32_t hi = std::min<uint32_t>(p.max_partition_order, 15), tz = 0;
            while (tz < hi && !((bs >> tz) & 1)) ++tz;  // a partition order must divide the block
            max_po = std::max(max_po, std::min(hi, tz));
            frames.push_back(f);
            pos += bs;
            ++idx;
        }
    }
    first_frame[n] = (uint32_t)frames.size();
    const uint32_t n_frames = (uint32_t)frames.size();
    const uint32_t plan_stride = 2u << max_po;
    uint16_t crc_tables[8][256];
    crc16_make_tables(crc_tables);
    const double t_plan = now_s();

    // ---- device buffers
    DevMem mem;
    StreamDev *d_streams;
    FrameJob *d_frames;
    ToneState *d_phases;
    double *d_windows, *d_base = nullptr;
    int32_t *d_sig, *d_sub, *d_res;
    uint16_t *d_plan, *d_crc;
    SubInfo *d_subs;
    uint32_t *d_frame_len;
    uint8_t *d_pcm;
    const size_t base_rows = (max_bs + 3) & ~3u;
    const bool timing = getenv("FLACGEN_TIMING") != nullptr;
    if (!mem.get(&d_streams, n) || !mem.get(&d_frames, n_frames) || !mem.get(&d_phases, phases.size()) ||
        !mem.get(&d_windows, windows.size()) || !mem.get(&d_sig, sig_total) ||
        !mem.get(&d_plan, (size_t)n_subs * plan_stride) || !mem.get(&d_crc, 8 * 256) ||
        !mem.get(&d_subs, n_subs) || !mem.get(&d_frame_len, n_frames) || !mem.get(&d_pcm, pcm_total + 64))
        return 3;
    d_res = d_sig;  // the residuals overwrite the channel samples, which nothing reads after k_channels
    const size_t base_count = any_stereo ? (size_t)((n_frames + 31) / 32) * base_rows * 32 : 0;
    if (any_stereo && !mem.get(&d_base, base_count)) return 3;
    // the subframe signals take over the base buffer (dead once k_synth is through) when it is large enough
    if (base_count * sizeof(double) >= sig_total * sizeof(int32_t)) d_sub = reinterpret_cast<int32_t *>(d_base);
    else if (!mem.get(&d_sub, sig_total)) return 3;
    if (timing) fprintf(stderr, "flacgen_cuda: cudaMalloc %.3f s\n", now_s() - t_plan);
    FGC_CK(cudaMemcpy(d_streams, streams.data(), n * sizeof(StreamDev), cudaMemcpyHostToDevice));
    FGC_CK(cudaMemcpy(d_frames, frames.data(), n_frames * sizeof(FrameJob), cudaMemcpyHostToDevice));
    FGC_CK(cudaMemcpy(d_phases, phases.data(), phases.size() * sizeof(ToneState), cudaMemcpyHostToDevice));
    FGC_CK(cudaMemcpy(d_windows, windows.data(), windows.size() * sizeof(double), cudaMemcpyHostToDevice));
    FGC_CK(cudaMemcpy(d_crc, crc_tables, sizeof(crc_tables), cudaMemcpyHostToDevice));
    FGC_CK(cudaMemsetAsync(d_pcm, 0, pcm_total + 64));
    const double t_up = now_s();
    if (timing) fprintf(stderr, "flacgen_cuda: plan %.3f s, malloc + upload %.3f s\n", t_plan - t_start, t_up - t_plan);

    // ---- signal, channels, predictors, residuals, Rice plans
    k_synth<<<(n_frames + 63) / 64, 64>>>(d_streams, d_frames, n_frames, d_phases, d_sig, d_base, (uint32_t)base_rows);
    k_channels<<<n_frames, 256>>>(d_streams, d_frames, d_sig, d_sub, d_pcm, d_subs);
    k_lpc<<<n_subs, kLpcThreads>>>(d_streams, d_frames, d_sub, d_windows, d_subs);
    k_residual<<<n_subs, kResThreads>>>(d_streams, d_frames, d_sub, d_res, d_plan, plan_stride, d_subs);
    k_frame_len<<<(n_frames + 255) / 256, 256>>>(d_streams, d_frames, n_frames, d_subs, d_frame_len);
    FGC_CK(cudaGetLastError());
    std::vector<uint32_t> frame_len(n_frames);
    FGC_CK(cudaMemcpy(frame_len.data(), d_frame_len, n_frames * sizeof(uint32_t), cudaMemcpyDeviceToHost));
    const double t_enc = now_s();

    // ---- layout of the blob, STREAMINFO (MD5 left for later)
    Handle *h = new (std::nothrow) Handle;
    if (!h) return 2;
    struct Guard {
        Handle *h;
        ~Guard() { if (h) fgc_free(h); }
    } guard{h};
    h->device = device;
    h->n_streams = n;
    h->n_frames = n_frames;
    h->n_subs = n_subs;
    h->outs.resize(n);
    std::vector<uint8_t> heads(42 * (size_t)n);
    std::vector<uint64_t> flac_off(n);
    uint64_t at = 0;
    for (uint32_t s = 0; s < n; ++s) {
        const fg_params &p = ps[s];
        at = (at + align - 1) / align * align;
        flac_off[s] = at;
        uint64_t audio = 0;
        uint32_t min_f = 0xFFFFFF, max_f = 0, min_b = 65535, max_b = 0;
        const uint32_t f0 = first_frame[s], nf = first_frame[s + 1] - f0;
        for (uint32_t k = 0; k < nf; ++k) {
            FrameJob &f = frames[f0 + k];
            f.frame_bytes = frame_len[f0 + k];
            f.out_byte = at + 42 + audio;
            audio += f.frame_bytes;
            min_f = std::min(min_f, f.frame_bytes);
            max_f = std::max(max_f, f.frame_bytes);
            max_b = std::max(max_b, f.block_size);
            if (k + 1 < nf || nf == 1) min_b = std::min(min_b, f.block_size);
        }
        if (max_f > 0xFFFFFF) max_f = 0;
        uint8_t *w = heads.data() + 42 * (size_t)s;
        memcpy(w, "fLaC", 4); w += 4;
        *w++ = 0x80; *w++ = 0; *w++ = 0; *w++ = 34;  // last block, STREAMINFO, 34 bytes
        *w++ = (uint8_t)(min_b >> 8); *w++ = (uint8_t)min_b;
        *w++ = (uint8_t)(max_b >> 8); *w++ = (uint8_t)max_b;
        *w++ = (uint8_t)(min_f >> 16); *w++ = (uint8_t)(min_f >> 8); *w++ = (uint8_t)min_f;
        *w++ = (uint8_t)(max_f >> 16); *w++ = (uint8_t)(max_f >> 8); *w++ = (uint8_t)max_f;
        const uint64_t total = p.set_total ? p.n_samples : 0;
        const uint64_t packed = ((uint64_t)p.sample_rate << 44) | ((uint64_t)(p.channels - 1) << 41) | ((uint64_t)(p.bps - 1) << 36) | (total & 0xFFFFFFFFFull);
        for (int i = 7; i >= 0; --i) *w++ = (uint8_t)(packed >> (8 * i));
        StreamOut &o = h->outs[s];
        memset(&o, 0, sizeof(o));
        o.flac_off = at;
        o.flac_bytes = 42 + audio;
        o.pcm_off = streams[s].pcm_off;
        o.pcm_bytes = p.n_samples * p.channels * streams[s].cb;
        o.n_frames = nf;
        o.container_bytes = streams[s].cb;
        at += 42 + audio;
    }
    h->flac_bytes = at;
    h->pcm_bytes = pcm_total;
    uint8_t *d_flac;
    if (!mem.get(&d_flac, ((at + 3) & ~3ull) + 64)) return 3;
    FGC_CK(cudaMemsetAsync(d_flac, 0, ((at + 3) & ~3ull) + 64));
    FGC_CK(cudaMemcpy(d_frames, frames.data(), n_frames * sizeof(FrameJob), cudaMemcpyHostToDevice));

    // ---- bits, CRC-16
    k_write<<<(n_subs + kWriteWarps - 1) / kWriteWarps, kWriteWarps * 32>>>(d_streams, d_frames, d_sub, d_res, d_plan, plan_stride, d_subs,
                                                                           n_subs, reinterpret_cast<uint32_t *>(d_flac));
    k_crc16<<<(n_frames + 127) / 128, 128>>>(d_frames, n_frames, d_crc, d_flac);
    FGC_CK(cudaGetLastError());
    FGC_CK(cudaDeviceSynchronize());
    const double t_write = now_s();

    // ---- MD5 of the PCM: per stream on the device for a batch, on the host for a few long streams
    if (n >= 32) {
        uint8_t *d_md5;
        if (!mem.get(&d_md5, 16 * (size_t)n)) return 3;
        k_md5_streams<<<(n + 31) / 32, 32>>>(d_streams, n, d_pcm, d_md5);
        std::vector<uint8_t> md5s(16 * (size_t)n);
        FGC_CK(cudaMemcpy(md5s.data(), d_md5, md5s.size(), cudaMemcpyDeviceToHost));
        for (uint32_t s = 0; s < n; ++s) memcpy(h->outs[s].md5, md5s.data() + 16 * (size_t)s, 16);
    } else {
        for (uint32_t s = 0; s < n; ++s)
            if (const int rc = hash_on_host(streams[s], d_pcm, h->outs[s].md5)) return rc;
    }
    for (uint32_t s = 0; s < n; ++s) memcpy(heads.data() + 42 * (size_t)s + 26, h->outs[s].md5, 16);
    {
        uint8_t *d_heads;
        uint64_t *d_off;
        if (!mem.get(&d_heads, heads.size()) || !mem.get(&d_off, n)) return 3;
        FGC_CK(cudaMemcpy(d_heads, heads.data(), heads.size(), cudaMemcpyHostToDevice));
        FGC_CK(cudaMemcpy(d_off, flac_off.data(), n * sizeof(uint64_t), cudaMemcpyHostToDevice));
        k_put_heads<<<(n + 127) / 128, 128>>>(d_heads, d_off, n, d_flac);
        FGC_CK(cudaDeviceSynchronize());
    }
    const double t_md5 = now_s();
    h->d_flac = reinterpret_cast<uint8_t *>(mem.release(d_flac));
    h->d_pcm = reinterpret_cast<uint8_t *>(mem.release(d_pcm));
    if (flags & FGC_KEEP_TRACE) {
        h->d_frames = reinterpret_cast<FrameJob *>(mem.release(d_frames));
        h->d_subs = reinterpret_cast<SubInfo *>(mem.release(d_subs));
        h->d_sub = reinterpret_cast<int32_t *>(mem.release(d_sub));
    }
    h->t[0] = t_plan - t_start;   // host: frame plan, phases, windows
    h->t[1] = t_up - t_plan;      // allocation + upload
    h->t[2] = t_enc - t_up;       // k_synth .. k_frame_len
    h->t[3] = t_write - t_enc;    // layout, k_write, k_crc16
    h->t[4] = t_md5 - t_write;    // MD5 + heads
    h->t[5] = t_md5 - t_start;    // all of it
    guard.h = nullptr;
    *handle = h;
    return 0;
}

// sizes[0] = FLAC blob bytes, [1] = PCM blob bytes, [2] = frames, [3] = subframes; seconds[0..6) as in fgc_generate
int fgc_info(void *handle, fgc_stream *outs, uint64_t *sizes, double *seconds) {
    Handle *h = reinterpret_cast<Handle *>(handle);
    if (!h) return 1;
    static_assert(sizeof(fgc_stream) == sizeof(StreamOut), "layout");
    if (outs) memcpy(outs, h->outs.data(), h->n_streams * sizeof(StreamOut));
    if (sizes) {
        sizes[0] = h->flac_bytes;
        sizes[1] = h->pcm_bytes;
        sizes[2] = h->n_frames;
        sizes[3] = h->n_subs;
    }
    if (seconds) memcpy(seconds, h->t, 6 * sizeof(double));
    return 0;
}

// Copies the blobs out: to device memory (D2D, *_on_device != 0) or host memory. A null destination is skipped.
int fgc_copy(void *handle, void *dst_flac, int flac_on_device, void *dst_pcm, int pcm_on_device, void *stream) {
    Handle *h = reinterpret_cast<Handle *>(handle);
    if (!h) return 1;
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    FGC_CK(cudaSetDevice(h->device));
    if (dst_flac) FGC_CK(cudaMemcpyAsync(dst_flac, h->d_flac, h->flac_bytes, flac_on_device ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost, st));
    if (dst_pcm) FGC_CK(cudaMemcpyAsync(dst_pcm, h->d_pcm, h->pcm_bytes, pcm_on_device ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost, st));
    FGC_CK(cudaStreamSynchronize(st));
    return 0;
}

// Debug (FGC_KEEP_TRACE): eight words per subframe, frame-major — what flacgen.cpp's fg_set_trace records.
int fgc_trace(void *handle, uint32_t *host_buf) {
    Handle *h = reinterpret_cast<Handle *>(handle);
    if (!h || !h->d_subs || !host_buf) return 1;
    FGC_CK(cudaSetDevice(h->device));
    uint32_t *d;
    FGC_CK(cudaMalloc((void **)&d, 32 * (size_t)h->n_subs));
    k_trace<<<h->n_subs, 128>>>(h->d_frames, h->d_sub, h->d_subs, d);
    const cudaError_t e = cudaMemcpy(host_buf, d, 32 * (size_t)h->n_subs, cudaMemcpyDeviceToHost);
    cudaFree(d);
    FGC_CK(e);
    return 0;
}

}  // extern "C"
