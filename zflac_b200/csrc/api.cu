// api.cu — host side of libzflacb200.so: the C ABI of include/zflacb200.h.
//
// Host work is what the reference does once per stream and never per sample: the signature
// and metadata walk (src/zflac.zig:217-253), container choice (:256-264), the stream-level
// frame-loop rules (first-frame latch :376-392, buffer growth / "stop trusting the total"
// :394-402, block-size-1 rule :405, termination :341-350, trim :583-588), error precedence
// ("first error in stream order wins") and the MD5 comparison (:279-280). Everything per byte
// or per sample runs in the kernels of scan.cu / decode.cu / epilogue.cu. There is no CPU
// decode path: without a CUDA device every entry point returns ZFLACB200_CUDA_ERROR.
#include "../../include/zflacb200.h"

#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <thread>
#include <mutex>
#include <string>
#include <unordered_map>
#include <vector>

#include "common.cuh"
#include "devcache.cuh"
#include "hostmd5.h"
#include "kernels.cuh"

using namespace zf;

namespace zf {
std::atomic<unsigned long long> g_kernel_launches{0};
std::atomic<unsigned long long> g_pass_a_launches[kPassAVariants];
}

namespace {

thread_local std::string g_last_error;
// One lock per DEVICE (not per process): calls on different devices run side by side, calls on one device take turns
// (they share its streams, pinned arena and mailbox). The reference's decode() is re-entrant; so is this.
std::mutex g_dev_mutex[64];

}  // namespace

namespace zf {
void set_last_error(const char *msg) { g_last_error = msg ? msg : ""; }  // for entry points that live in other files
}

namespace {

struct DeviceCtx {
    bool ready = false;
    cudaStream_t stream = nullptr;
    cudaStream_t side = nullptr;  // diagnostics that overlap the main pipeline (per-frame CRC-16)
    cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
    // zflacb200_decode pipeline: upload / read-back streams (both PCIe directions at once) and a pinned mailbox
    cudaStream_t copy_in = nullptr, copy_out = nullptr;
    static constexpr int kLanes = 8;        // chunks in flight: a chunk's kernels are latency-bound, several run side by side
    cudaStream_t lane[kLanes] = {};
    unsigned long long *mailbox = nullptr;  // pinned host: [0] candidates of the chunk, [2 + i] PCM elements ready after chunk i
    // pinned staging for the small transfers of one call (xfer.cu): bump-allocated, rewound by the public entry points
    std::vector<std::pair<uint8_t *, size_t>> pin_blocks;
    size_t pin_used = 0;
    // host-side MD5 gate of a device-resident stream (hostmd5.h): two pinned staging buffers the PCM passes through
    static constexpr size_t kMd5Stage = 8u << 20;
    uint8_t *md5_stage[2] = {nullptr, nullptr};
    cudaEvent_t ev_md5[2] = {nullptr, nullptr};
};
DeviceCtx g_ctx[64];

#define CU_CHECK(expr)                                                                          \
    do {                                                                                        \
        cudaError_t _e = (expr);                                                                \
        if (_e != cudaSuccess) {                                                                \
            char _b[512];                                                                       \
            snprintf(_b, sizeof(_b), "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__); \
            g_last_error = _b;                                                                  \
            return ZFLACB200_CUDA_ERROR;                                                        \
        }                                                                                       \
    } while (0)

// Device memory: the library's own block cache (devcache.cuh); ZFLACB200_ALLOC=pool selects the driver's
// stream-ordered pool instead (A/B runs).
zf::DevCache g_cache;
bool use_driver_pool() {
    static const bool v = [] {
        const char *e = getenv("ZFLACB200_ALLOC");
        return e && strcmp(e, "pool") == 0;
    }();
    return v;
}
cudaError_t dev_alloc(void **p, size_t bytes, cudaStream_t st) {
    return use_driver_pool() ? cudaMallocAsync(p, bytes, st) : g_cache.alloc(p, bytes, st);
}
void dev_free(void *p, cudaStream_t st) {
    if (!p) return;
    if (use_driver_pool()) cudaFreeAsync(p, st);
    else g_cache.free(p, st);
}

// ---- small transfers (xfer.cu): host side in the per-device pinned arena, moved by a kernel on the caller's stream
void pin_rewind(DeviceCtx &c) {  // only the newest (largest) block survives a call
    while (c.pin_blocks.size() > 1) {
        cudaFreeHost(c.pin_blocks.front().first);
        c.pin_blocks.erase(c.pin_blocks.begin());
    }
    c.pin_used = 0;
}
void *pin_take(DeviceCtx &c, size_t bytes) {
    bytes = (bytes + 63) & ~(size_t)63;
    if (c.pin_blocks.empty() || c.pin_used + bytes > c.pin_blocks.back().second) {
        const size_t last = c.pin_blocks.empty() ? 0 : c.pin_blocks.back().second;
        const size_t cap = std::max<size_t>({bytes, 2 * last, (size_t)4 << 20});
        uint8_t *p = nullptr;
        if (cudaHostAlloc((void **)&p, cap, cudaHostAllocDefault) != cudaSuccess) {
            cudaGetLastError();
            return nullptr;
        }
        c.pin_blocks.emplace_back(p, cap);
        c.pin_used = 0;
    }
    void *r = c.pin_blocks.back().first + c.pin_used;
    c.pin_used += bytes;
    return r;
}
thread_local bool g_alloc_oom = false;  // an allocation the device could not satisfy (DevBuf::alloc): OutOfMemory, not CudaError

// Scope of one C-ABI call: resolves the device the call runs on, takes that device's lock, rewinds its pinned arena,
// and puts the CALLER's current device back when the call ends (the library switches devices inside).
struct ApiScope {
    int device = 0, prev = -1;
    std::unique_lock<std::mutex> lk;
    explicit ApiScope(int want) {
        if (cudaGetDevice(&prev) != cudaSuccess) {
            cudaGetLastError();
            prev = -1;
        }
        device = want >= 0 ? want : prev;
        if (device < 0 || device >= 64) device = 0;
        lk = std::unique_lock<std::mutex>(g_dev_mutex[device]);
        g_last_error.clear();
        g_alloc_oom = false;
        if (g_ctx[device].ready) pin_rewind(g_ctx[device]);
    }
    ~ApiScope() {
        if (prev >= 0) cudaSetDevice(prev);
    }
    int done(int rc) const { return rc == ZFLACB200_CUDA_ERROR && g_alloc_oom ? (int)ZFLACB200_OUT_OF_MEMORY : rc; }
};
int opt_device(const zflacb200_options *opt) { return opt ? opt->device : -1; }
// host -> device; `src` may be a temporary (it is copied into the arena before this returns)
cudaError_t put_small(DeviceCtx &c, void *d_dst, const void *src, size_t bytes, cudaStream_t st) {
    if (bytes == 0) return cudaSuccess;
    void *slot = pin_take(c, bytes);
    if (!slot) return cudaErrorMemoryAllocation;
    memcpy(slot, src, bytes);
    return zf::launch_copy_small(d_dst, slot, bytes, st);
}
// device -> host: returns where the bytes will be once `st` has been synchronised (nullptr: out of pinned memory)
const void *get_small(DeviceCtx &c, const void *d_src, size_t bytes, cudaStream_t st) {
    void *slot = pin_take(c, bytes ? bytes : 1);
    if (!slot) return nullptr;
    if (zf::launch_copy_small(slot, d_src, bytes, st) != cudaSuccess) return nullptr;
    return slot;
}
// STREAMINFO's total is untrusted input and the fast path sizes its buffers from it: a frame holds at most 65535 samples
// per channel and takes at least 10 bytes (header 6, one subframe byte, CRC-16, ...), so a total beyond what the bytes
// could hold sends the stream to the robust path, which sizes everything from the frames it actually finds (and reports
// EndOfStream like the reference, :341-350) — instead of asking the driver for terabytes.
bool plausible_total(uint64_t total, uint64_t audio_bytes) { return total <= 65535ull * (audio_bytes / 10 + 1); }

// (g_alloc_oom: an allocation the device could not satisfy, remembered per thread so that callers can tell it from a
// real CUDA error: fast path -> robust path, and OutOfMemory rather than CudaError at the C ABI)
constexpr int kFastOom = -2;

// ---- where one stream's MD5 chain runs (hostmd5.h): on a host thread unless the stream is short or the caller insists
bool md5_on_host(uint32_t flags, uint64_t pcm_bytes) {
    const char *e = getenv("ZFLACB200_HOST_MD5_MIN_BYTES");  // test knob, read on every call
    const uint64_t min_bytes = e ? strtoull(e, nullptr, 0) : (256ull << 10);
    return (flags & ZFLACB200_MD5_ON_DEVICE) == 0 && pcm_bytes >= min_bytes;
}

#define GET_SMALL(var, ctx, d_src, bytes, st)                                  \
    const void *var = get_small(ctx, d_src, bytes, st);                        \
    if (!var) {                                                                \
        g_last_error = "small device->host transfer failed (" #d_src ")";      \
        return ZFLACB200_CUDA_ERROR;                                           \
    }

// MD5 of PCM that stays in HBM, on the calling host thread (hostmd5.h): the bytes pass through two pinned staging
// buffers, the copy of piece i + 1 running under the hash of piece i. Everything queued on `st` so far is waited for.
int host_md5_of_device_pcm(DeviceCtx &ctx, const void *d_pcm, uint64_t n_samples, uint32_t cb, uint32_t hb, uint32_t unshift,
                           cudaStream_t st, uint8_t out[16], float *ms_host) {
    for (int k = 0; k < 2; ++k) {
        if (!ctx.md5_stage[k]) CU_CHECK(cudaHostAlloc((void **)&ctx.md5_stage[k], DeviceCtx::kMd5Stage, cudaHostAllocDefault));
        if (!ctx.ev_md5[k]) CU_CHECK(cudaEventCreateWithFlags(&ctx.ev_md5[k], cudaEventDisableTiming));
    }
    const auto t0 = std::chrono::steady_clock::now();
    CU_CHECK(cudaEventRecord(ctx.ev_fork, st));
    CU_CHECK(cudaStreamWaitEvent(ctx.copy_out, ctx.ev_fork, 0));
    zf::PcmHasher h;
    h.container_bytes = cb;
    h.hash_bytes = hb;
    h.unshift = unshift;
    const uint64_t per = DeviceCtx::kMd5Stage / cb;
    uint64_t queued = 0, done = 0, cnt[2] = {0, 0};
    auto queue = [&](int slot) -> int {
        const uint64_t m = std::min<uint64_t>(per, n_samples - queued);
        CU_CHECK(cudaMemcpyAsync(ctx.md5_stage[slot], (const uint8_t *)d_pcm + queued * cb, m * cb, cudaMemcpyDeviceToHost, ctx.copy_out));
        CU_CHECK(cudaEventRecord(ctx.ev_md5[slot], ctx.copy_out));
        cnt[slot] = m;
        queued += m;
        return ZFLACB200_OK;
    };
    int cur = 0;
    if (n_samples && queue(0) != ZFLACB200_OK) return ZFLACB200_CUDA_ERROR;
    while (done < n_samples) {
        if (queued < n_samples && queue(cur ^ 1) != ZFLACB200_OK) return ZFLACB200_CUDA_ERROR;
        CU_CHECK(cudaEventSynchronize(ctx.ev_md5[cur]));
        h.update(ctx.md5_stage[cur], cnt[cur]);
        done += cnt[cur];
        cur ^= 1;
    }
    if (n_samples == 0) CU_CHECK(cudaStreamSynchronize(st));
    h.finish(out);
    if (ms_host) *ms_host = std::chrono::duration<float, std::milli>(std::chrono::steady_clock::now() - t0).count();
    return ZFLACB200_OK;
}

// Host-side phase timer: ZFLACB200_TRACE=1 prints where the HOST spends its time inside a call (stderr).
struct HostTrace {
    bool on;
    const char *what;
    std::chrono::steady_clock::time_point t0, last;
    std::string log;
    explicit HostTrace(const char *w) : on(getenv("ZFLACB200_TRACE") != nullptr), what(w) {
        if (on) t0 = last = std::chrono::steady_clock::now();
    }
    void mark(const char *name) {
        if (!on) return;
        const auto t = std::chrono::steady_clock::now();
        char b[96];
        snprintf(b, sizeof(b), " %s=%.3f", name, std::chrono::duration<double, std::milli>(t - last).count());
        log += b;
        last = t;
    }
    ~HostTrace() {
        if (!on) return;
        const auto t = std::chrono::steady_clock::now();
        fprintf(stderr, "[zflacb200 trace] %s total=%.3f ms:%s\n", what, std::chrono::duration<double, std::milli>(t - t0).count(), log.c_str());
    }
};

// Diagnostics: ZFLACB200_EXTRA_FLAGS=<hex> is OR-ed into every call's option flags (A/B runs of the bench).
uint32_t env_flags() {
    static const uint32_t v = [] {
        const char *e = getenv("ZFLACB200_EXTRA_FLAGS");
        return e ? (uint32_t)strtoul(e, nullptr, 16) : 0u;
    }();
    return v;
}

int ensure_device(int device, int *resolved) {
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0) {
        g_last_error = std::string("no CUDA device: ") + cudaGetErrorString(e) + " (there is no CPU fallback)";
        return ZFLACB200_CUDA_ERROR;
    }
    if (device < 0) CU_CHECK(cudaGetDevice(&device));
    if (device >= count || device >= 64) {
        g_last_error = "device ordinal out of range";
        return ZFLACB200_INVALID_ARGUMENT;
    }
    CU_CHECK(cudaSetDevice(device));
    DeviceCtx &c = g_ctx[device];
    if (!c.ready) {
        cudaDeviceProp prop;
        CU_CHECK(cudaGetDeviceProperties(&prop, device));
        if (prop.major < 10) {
            g_last_error = "libzflacb200 is built for sm_100a (B200) only; device is sm_" + std::to_string(prop.major) + std::to_string(prop.minor);
            return ZFLACB200_CUDA_ERROR;
        }
        CU_CHECK(cudaStreamCreateWithFlags(&c.stream, cudaStreamNonBlocking));
        CU_CHECK(cudaStreamCreateWithFlags(&c.side, cudaStreamNonBlocking));
        CU_CHECK(cudaEventCreateWithFlags(&c.ev_fork, cudaEventDisableTiming));
        CU_CHECK(cudaEventCreateWithFlags(&c.ev_join, cudaEventDisableTiming));
        CU_CHECK(cudaStreamCreateWithFlags(&c.copy_in, cudaStreamNonBlocking));
        CU_CHECK(cudaStreamCreateWithFlags(&c.copy_out, cudaStreamNonBlocking));
        for (auto &l : c.lane) CU_CHECK(cudaStreamCreateWithFlags(&l, cudaStreamNonBlocking));
        CU_CHECK(cudaHostAlloc((void **)&c.mailbox, 8 * (2 + 4096), cudaHostAllocDefault));
        CU_CHECK(scan_init_constants());
        if (use_driver_pool()) {  // ZFLACB200_ALLOC=pool (A/B runs) only: the process's default pool is not ours to retune otherwise
            cudaMemPool_t pool;
            CU_CHECK(cudaDeviceGetDefaultMemPool(&pool, device));
            uint64_t thresh = ~0ull;  // keep freed blocks cached: open() allocates the same sizes again and again
            CU_CHECK(cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &thresh));
        }
        c.ready = true;
    }
    *resolved = device;
    return ZFLACB200_OK;
}

// ------------------------------------------------------------------ byte access for host logic
struct ByteSource {
    const uint8_t *host = nullptr;
    const uint8_t *dev = nullptr;
    uint64_t len = 0;
    // device-resident stream: its head, fetched with ONE copy — the metadata walk and the first frame header read a few
    // bytes at a time, and every small synchronous copy costs a host round trip (~15 us each, 5 % of a C2 step together)
    std::vector<uint8_t> head;
    void set_device(const uint8_t *d, uint64_t n) {
        dev = d;
        len = n;
        head.resize((size_t)std::min<uint64_t>(n, 16384));
        if (!head.empty() && cudaMemcpy(head.data(), d, head.size(), cudaMemcpyDeviceToHost) != cudaSuccess) {
            cudaGetLastError();
            head.clear();
        }
    }
    // copies [off, off+n) into dst; bytes past len are not written. Returns bytes available.
    uint64_t get(uint64_t off, uint64_t n, uint8_t *dst) const {
        if (off >= len) return 0;
        const uint64_t m = std::min<uint64_t>(n, len - off);
        if (host) memcpy(dst, host + off, m);
        else if (off + m <= head.size()) memcpy(dst, head.data() + off, m);
        else if (cudaMemcpy(dst, dev + off, m, cudaMemcpyDeviceToHost) != cudaSuccess) return 0;
        return m;
    }
};

struct Streaminfo {
    uint32_t min_block = 0, max_block = 0, min_frame = 0, max_frame = 0;
    uint32_t sample_rate = 0, channels_m1 = 0, depth_m1 = 0;
    uint64_t total = 0;
    uint8_t md5[16] = {0};
};

// src/zflac.zig:217-253 + :309: signature, metadata blocks, STREAMINFO.
int parse_metadata(const ByteSource &src, Streaminfo &si, uint64_t &audio_offset) {
    uint8_t b[40];
    if (src.get(0, 4, b) < 4) return ZFLACB200_END_OF_STREAM;
    if (!(b[0] == 0x66 && b[1] == 0x4C && b[2] == 0x61 && b[3] == 0x43)) return ZFLACB200_INVALID_SIGNATURE;
    uint64_t pos = 4;
    bool have = false;
    for (;;) {
        if (src.get(pos, 4, b) < 4) return ZFLACB200_END_OF_STREAM;
        pos += 4;
        const unsigned type = b[0] & 0x7F;
        const bool last = (b[0] & 0x80) != 0;
        const uint32_t length = ((uint32_t)b[1] << 16) | ((uint32_t)b[2] << 8) | b[3];
        if (type == 0) {  // STREAMINFO: always 34 bytes, the length field is not consulted (Q13)
            if (src.get(pos, 34, b) < 34) return ZFLACB200_END_OF_STREAM;
            pos += 34;
            si.min_block = ((uint32_t)b[0] << 8) | b[1];
            si.max_block = ((uint32_t)b[2] << 8) | b[3];
            si.min_frame = ((uint32_t)b[4] << 16) | ((uint32_t)b[5] << 8) | b[6];
            si.max_frame = ((uint32_t)b[7] << 16) | ((uint32_t)b[8] << 8) | b[9];
            si.sample_rate = ((uint32_t)b[10] << 12) | ((uint32_t)b[11] << 4) | (b[12] >> 4);
            si.channels_m1 = (b[12] >> 1) & 7;
            si.depth_m1 = ((b[12] & 1u) << 4) | (b[13] >> 4);
            si.total = ((uint64_t)(b[13] & 15) << 32) | ((uint64_t)b[14] << 24) | ((uint64_t)b[15] << 16) | ((uint64_t)b[16] << 8) | b[17];
            memcpy(si.md5, b + 18, 16);
            have = true;
        } else if (type >= 1 && type <= 6) {
            if (src.len - pos < length) return ZFLACB200_END_OF_STREAM;
            pos += length;
        } else {
            return ZFLACB200_INVALID_METADATA_HEADER;
        }
        if (last) break;
    }
    if (!have) return ZFLACB200_MISSING_STREAMINFO;
    audio_offset = pos;
    return ZFLACB200_OK;
}

const uint32_t kRate[16] = {0, 88200, 176400, 192000, 8000, 16000, 22050, 24000, 32000, 44100, 48000, 96000, 0, 0, 0, 0};
const uint8_t kChCount[16] = {1, 2, 3, 4, 5, 6, 7, 8, 2, 2, 2, 0, 0, 0, 0, 0};
const uint8_t kBps[8] = {0, 8, 12, 0, 16, 20, 24, 32};

struct Hdr {
    uint32_t block_size = 0, sr_value = 0;
    uint8_t n_before_crc = 0, ch_code = 0, bd_code = 0, sr_code = 0;
};

// Frame header at a position the scan did not emit: same checks in the same order as the
// reference (:343-374). `n_before_crc` excludes the CRC-8 byte (read later, :407).
int host_parse_header(const ByteSource &src, uint64_t pos, Hdr &h) {
    uint8_t b[16] = {0};
    const uint64_t avail = src.get(pos, 16, b);
    if (avail < 4) return ZFLACB200_END_OF_STREAM;
    if (!(b[0] == 0xFF && (b[1] & 0xFE) == 0xF8)) return ZFLACB200_INVALID_FRAME_HEADER;
    const unsigned bs_code = b[2] >> 4, sr_code = b[2] & 15;
    h.ch_code = b[3] >> 4;
    h.bd_code = (b[3] >> 1) & 7;
    h.sr_code = (uint8_t)sr_code;
    if (avail < 5) return ZFLACB200_END_OF_STREAM;
    const unsigned c0 = b[4];
    unsigned ones = 0;
    while (ones < 8 && (c0 & (0x80u >> ones))) ++ones;
    if (c0 == 0xFF || ones == 1) return ZFLACB200_INVALID_CODED_NUMBER;
    unsigned n = 4 + (ones == 0 ? 1 : ones);
    if (n > avail) return ZFLACB200_END_OF_STREAM;
    if (bs_code == 0) return ZFLACB200_INVALID_FRAME_HEADER;
    if (bs_code == 6) {
        if (n + 1 > avail) return ZFLACB200_END_OF_STREAM;
        h.block_size = (uint32_t)b[n] + 1;
        n += 1;
    } else if (bs_code == 7) {
        if (n + 2 > avail) return ZFLACB200_END_OF_STREAM;
        const uint32_t v = ((uint32_t)b[n] << 8) | b[n + 1];
        if (v == 0xFFFF) return ZFLACB200_INVALID_FRAME_HEADER;
        h.block_size = v + 1;
        n += 2;
    } else if (bs_code == 1) h.block_size = 192;
    else if (bs_code <= 5) h.block_size = 144u << bs_code;
    else h.block_size = 1u << bs_code;
    h.sr_value = kRate[sr_code];
    if (sr_code == 12) {
        if (n + 1 > avail) return ZFLACB200_END_OF_STREAM;
        h.sr_value = b[n];
        n += 1;
    } else if (sr_code == 13 || sr_code == 14) {
        if (n + 2 > avail) return ZFLACB200_END_OF_STREAM;
        h.sr_value = (((uint32_t)b[n] << 8) | b[n + 1]) * (sr_code == 14 ? 10u : 1u);
        n += 2;
    } else if (sr_code == 15) return ZFLACB200_INVALID_FRAME_HEADER;
    h.n_before_crc = (uint8_t)n;
    return ZFLACB200_OK;
}

struct ChainFrame {
    FrameRec rec;
    uint64_t expect_end;
    bool probed;
};

struct Chain {
    std::vector<ChainFrame> frames;
    int status = ZFLACB200_OK;  // error that ends the stream, reported unless an earlier frame fails in K2
    uint64_t n_samples = 0;
    uint64_t scratch_elems = 0;
    uint32_t max_block = 0;
    uint8_t channels = 0, bps = 0;
    uint32_t sample_rate = 0;
    uint64_t n_probed = 0;
};

struct Pipeline {
    const uint8_t *d_stream = nullptr;
    uint64_t len = 0;
    cudaStream_t st = nullptr;
    ByteSource src;
    Streaminfo si;
    uint64_t audio_offset = 0;
    std::vector<Cand> cands;
    std::vector<uint32_t> link;
    std::unordered_map<uint64_t, FrameOut> known;  // frame offset -> end/status found by decoding it
    bool use_links = true;
    bool wide = false;  // 32 bps stream: int64 scratch
    // one-frame probe buffers
    FrameRec *d_probe_rec = nullptr;
    FrameOut *d_probe_out = nullptr;

    int probe(const FrameRec &rec, FrameOut &out) {
        CU_CHECK(cudaMemcpyAsync(d_probe_rec, &rec, sizeof(rec), cudaMemcpyHostToDevice, st));
        CU_CHECK(launch_decode_frames(d_stream, d_probe_rec, 1, nullptr, wide, d_probe_out, false, false, st));
        CU_CHECK(cudaMemcpyAsync(&out, d_probe_out, sizeof(out), cudaMemcpyDeviceToHost, st));
        CU_CHECK(cudaStreamSynchronize(st));
        return ZFLACB200_OK;
    }

    // The reference's frame loop (:340-581) minus the per-sample work. Returns a CUDA-level
    // status; the stream-level outcome is chain.status.
    int build_chain(Chain &ch) {
        ch = Chain();
        bool first = true;
        uint32_t sample_rate = 0, channel_count = 0, bd_code = 0, bps = 0;
        bool valid_total = si.total > 0;
        const uint64_t expected_channels = (uint64_t)si.channels_m1 + 1;
        const uint64_t total = expected_channels * (valid_total ? si.total : 4096);
        uint64_t samples_len = total;
        uint64_t pos = audio_offset, off = 0;
        size_t ci = 0;  // moving index into cands (sorted by offset)
        for (;;) {
            if (valid_total && off >= total) break;  // :341
            if (len - std::min(pos, len) < 4) {        // readInt(u32) fails (:343-350)
                if (valid_total) ch.status = ZFLACB200_END_OF_STREAM;
                break;
            }
            while (ci < cands.size() && cands[ci].offset < pos) ++ci;
            const bool is_cand = ci < cands.size() && cands[ci].offset == pos;
            Hdr h;
            if (is_cand) {
                const Cand &c = cands[ci];
                h.block_size = c.block_size;
                h.sr_value = c.sr_value;
                h.n_before_crc = c.header_len - 1;
                h.ch_code = c.ch_code;
                h.bd_code = c.bd_code;
                h.sr_code = c.sr_code;
            } else {
                const int e = host_parse_header(src, pos, h);
                if (e != ZFLACB200_OK) {
                    ch.status = e;
                    break;
                }
            }
            const uint32_t frame_rate = (h.sr_code == 0) ? si.sample_rate : h.sr_value;
            if (first) {  // :376-388
                sample_rate = frame_rate;
                channel_count = kChCount[h.ch_code];
                bd_code = h.bd_code;
                if (bd_code == 0) bps = si.depth_m1 + 1;
                else if (bd_code == 3) { ch.status = ZFLACB200_UNDEFINED_IN_REFERENCE; break; }
                else bps = kBps[bd_code];
                if (channel_count != expected_channels) { ch.status = ZFLACB200_INCONSISTENT_PARAMETERS; break; }
                first = false;
                ch.channels = (uint8_t)channel_count;
                ch.bps = (uint8_t)bps;
                ch.sample_rate = sample_rate & 0xFFFFFF;
            } else if (sample_rate != frame_rate || channel_count != kChCount[h.ch_code] || bd_code != h.bd_code) {
                ch.status = ZFLACB200_INCONSISTENT_PARAMETERS;  // :391
                break;
            }
            const uint64_t expected = off + (uint64_t)h.block_size * channel_count;  // :394-402
            if (samples_len < expected) {
                samples_len = std::max(2 * samples_len, expected);
                valid_total = false;
            }
            if (h.block_size == 1 && valid_total && off + channel_count < total) {  // :405
                ch.status = ZFLACB200_INVALID_FRAME_HEADER;
                break;
            }
            if (pos + h.n_before_crc + 1 > len) {  // CRC-8 byte (:407)
                ch.status = ZFLACB200_END_OF_STREAM;
                break;
            }
            ChainFrame cf;
            cf.rec.offset = pos;
            cf.rec.stream_end = len;
            cf.rec.sample_base = off;
            cf.rec.scratch_base = ch.scratch_elems;
            cf.rec.block_size = h.block_size;
            cf.rec.scratch_pitch = round_up_u32(h.block_size, 8);
            cf.rec.header_len = h.n_before_crc + 1;
            cf.rec.ch_code = h.ch_code;
            cf.rec.channels = (uint8_t)channel_count;
            cf.rec.bps = (uint8_t)bps;
            cf.rec.stream_id = 0;
            cf.probed = false;
            // where does this frame end?
            FrameOut fo;
            auto it = known.find(pos);
            if (it != known.end()) {
                fo = it->second;
                cf.probed = true;
            } else if (use_links && is_cand && link[ci] != kNoLink) {
                fo.status = ST_OK;
                fo.end_offset = link[ci] < cands.size() ? cands[link[ci]].offset : len;
                fo.stored_crc16 = 0;
            } else {
                const int e = probe(cf.rec, fo);
                if (e != ZFLACB200_OK) return e;
                known[pos] = fo;
                cf.probed = true;
                ++ch.n_probed;
            }
            if (fo.status != ST_OK) {
                ch.status = (int)fo.status;
                break;
            }
            cf.expect_end = fo.end_offset;
            ch.frames.push_back(cf);
            ch.scratch_elems += (uint64_t)cf.rec.scratch_pitch * channel_count;
            ch.max_block = std::max(ch.max_block, h.block_size);
            off += (uint64_t)h.block_size * channel_count;  // :580
            pos = fo.end_offset;
        }
        ch.n_samples = off;
        return ZFLACB200_OK;
    }
};

struct Events {
    cudaEvent_t e[8];
    bool made = false;
    int init() {
        for (auto &x : e) CU_CHECK(cudaEventCreate(&x));
        made = true;
        return ZFLACB200_OK;
    }
    ~Events() {
        if (made)
            for (auto &x : e) cudaEventDestroy(x);
    }
};

float ev_ms(cudaEvent_t a, cudaEvent_t b) {
    float ms = 0;
    cudaEventElapsedTime(&ms, a, b);
    return ms;
}

}  // namespace

struct zflacb200_stream {
    int device = 0;
    void *d_pcm = nullptr;
    uint64_t pcm_bytes = 0;
    zflacb200_info info;
    std::vector<zflacb200_frame> frames;
    // fast path: the frame table stays on the device until zflacb200_frames() asks for it
    FrameRec *d_recs = nullptr;
    FrameOut *d_outs = nullptr;
    uint8_t *d_frame_ok = nullptr;
    uint32_t n_recs = 0;
    bool frames_on_device = false;
};

struct zflacb200_batch {
    int device = 0;
    struct Group { void *d_pcm = nullptr; uint64_t bytes = 0; };
    struct Item {
        int group = -1;              // launch group whose PCM buffer holds the stream, -1 = none
        uint64_t byte_off = 0, bytes = 0;
        zflacb200_stream *solo = nullptr;  // stream decoded on its own (robust path)
    };
    std::vector<Group> groups;
    std::vector<Item> items;
    uint64_t read_all_bytes = 0;
};

namespace {

struct DevBuf {  // stream-ordered allocation released on scope exit
    void *p = nullptr;
    cudaStream_t st = nullptr;
    int alloc(size_t bytes, cudaStream_t s) {
        if (p) {
            dev_free(p, st ? st : s);
            p = nullptr;
        }
        st = s;
        const cudaError_t e = dev_alloc(&p, bytes ? bytes : 16, s);
        if (e == cudaErrorMemoryAllocation) {
            cudaGetLastError();  // not sticky
            p = nullptr;
            g_alloc_oom = true;
            g_last_error = "device out of memory (" + std::to_string(bytes) + " bytes requested)";
            return ZFLACB200_OUT_OF_MEMORY;
        }
        CU_CHECK(e);
        return ZFLACB200_OK;
    }
    void *release() {
        void *r = p;
        p = nullptr;
        return r;
    }
    ~DevBuf() {
        if (p) dev_free(p, st);
    }
};

// Subframe decode + epilogue over a set of frame records. Streams with <= 31 bits per sample run the
// two-pass decoder (K2a parse_frames: the serial bit chain, one lane per frame; K3p predict_epilogue:
// predictor recurrence + decorrelation + interleave, one lane per subframe); 32 bps streams (int64
// scratch) and ZFLACB200_SINGLE_PASS run the fused single-pass K2 + K3.
struct SubframeStage {
    bool two_pass = false;
    bool narrow = false;  // uint16 scratch of folded residuals (parse_narrow.cuh): streams of <= 16 bits per sample
    bool word_pipeline = false;  // diagnostics (ZFLACB200_WORD_PIPELINE): int32 rows by the first pass A (parse_core.cuh) instead of the position reader
    bool inter64 = false;  // container wider than 16 bits: the reference's sums are i64 (decode_core.cuh, "Arithmetic")
    DevBuf b_subs, b_big;  // b_big: {some order above 12, a value did not fit the narrow scratch}
    // depth: bits per sample by STREAMINFO (largest of a batch group); 0 or allow_narrow = false keeps the int32 scratch
    int prepare(bool wide, uint32_t flags, uint32_t n_recs, cudaStream_t st, uint32_t depth = 0, bool allow_narrow = true) {
        two_pass = !wide && (flags & ZFLACB200_SINGLE_PASS) == 0;
        narrow = false;
        word_pipeline = (flags & ZFLACB200_WORD_PIPELINE) != 0;
        inter64 = depth > 16;
        if (!two_pass) return ZFLACB200_OK;
        static const bool env_no_narrow = getenv("ZFLACB200_NO_NARROW") != nullptr;  // diagnostics: A/B runs
        narrow = allow_narrow && !env_no_narrow && depth >= 1 && depth <= 16;
        if (b_subs.alloc(kSubRecBytes * 8 * (size_t)std::max<uint32_t>(n_recs, 1), st) || b_big.alloc(2 * sizeof(uint32_t), st))
            return ZFLACB200_CUDA_ERROR;
        CU_CHECK(cudaMemsetAsync(b_big.p, 0, 2 * sizeof(uint32_t), st));
        return ZFLACB200_OK;
    }
    const uint32_t *overflow_flag() const { return (const uint32_t *)b_big.p + 1; }
    // d_recs / d_outs point at record `first` of the prepared range (chunked decode), n records from there
    int decode(const uint8_t *d_stream, const FrameRec *d_recs, uint32_t n, void *d_scratch, bool wide, FrameOut *d_outs, bool write,
               cudaStream_t st, uint32_t first = 0) {
        if (two_pass && narrow)
            CU_CHECK(launch_parse_frames_narrow(d_stream, d_recs, n, (uint16_t *)d_scratch,
                                                (SubRec *)((uint8_t *)b_subs.p + kSubRecBytes * 8 * (size_t)first), d_outs,
                                                (uint32_t *)b_big.p, write, st));
        else if (two_pass)
            CU_CHECK(launch_parse_frames(d_stream, d_recs, n, (int32_t *)d_scratch,
                                         (SubRec *)((uint8_t *)b_subs.p + kSubRecBytes * 8 * (size_t)first), d_outs, (uint32_t *)b_big.p,
                                         write, st, word_pipeline));
        else
            CU_CHECK(launch_decode_frames(d_stream, d_recs, n, d_scratch, wide, d_outs, write, inter64, st));
        return ZFLACB200_OK;
    }
    int epilogue(const EpilogueParams &ep, const FrameOut *d_outs, uint32_t max_channels, cudaStream_t st, uint32_t first = 0) {
        if (!two_pass) {
            CU_CHECK(launch_epilogue(ep, st));
            return ZFLACB200_OK;
        }
        PredictParams pp;
        pp.frames = ep.frames;
        pp.outs = d_outs;
        pp.n_frames = ep.n_frames;
        pp.max_channels = max_channels;
        pp.scratch = ep.scratch;
        pp.subs = (const uint8_t *)b_subs.p + kSubRecBytes * 8 * (size_t)first;
        pp.pcm = ep.pcm;
        pp.container_bytes = ep.container_bytes;
        pp.msb_shift = ep.msb_shift;
        pp.lossy_flag = ep.lossy_flag;
        pp.big_orders = (const uint32_t *)b_big.p;
        pp.narrow = narrow;
        CU_CHECK(launch_predict_epilogue(pp, st));
        return ZFLACB200_OK;
    }
};


// One launch group of the fast path: streams that share the scratch element type (wide), the
// container type and the MSB-align shift, packed in one device buffer.
struct FastGroup {
    int device = 0;
    cudaStream_t st = nullptr;
    const uint8_t *d_blob = nullptr;
    uint64_t blob_len = 0;
    const Cand *d_cands = nullptr;
    uint32_t n_cand = 0;
    bool wide = false;
    uint32_t container_bytes = 2, msb_shift = 0, max_block = 4096;
    bool do_md5 = true;
    uint32_t flags = 0;                   // zflacb200_options.flags of the call
    std::vector<StreamDesc> streams;      // pcm_base / scratch_base filled by plan()
    std::vector<uint32_t> hash_bytes;     // per stream: bytes of each sample that enter the MD5 (Q4)
    uint64_t pcm_elems = 0, scratch_elems = 0;
    // results
    std::vector<StreamSummary> sums;
    std::vector<uint8_t> digests;         // 16 bytes per stream
    DevBuf b_pcm, b_recs, b_outs, b_ok;
    float ms_hint = 0, ms_decode = 0, ms_epilogue = 0, ms_md5 = 0;
    bool md5_host = false;  // ms_md5 is host time (hostmd5.h), not part of the event span e[0]..e[7]
    bool allow_narrow = true, narrow_overflow = false;  // uint16 scratch (parse_narrow.cuh) and its escape hatch

    void plan() {  // output and scratch slots from the STREAMINFO totals (host-known)
        pcm_elems = scratch_elems = 0;
        for (auto &sd : streams) {
            const uint64_t C = sd.si_channels;
            sd.pcm_base = pcm_elems;
            sd.scratch_base = scratch_elems;
            pcm_elems += (sd.total * C + 15) & ~15ull;  // 16-element slots keep the vector stores aligned
            scratch_elems += C * (((sd.total + 7) & ~7ull) + 8);
        }
        scratch_elems += 64ull * n_cand + 64;
    }
    bool accepted(size_t i) const {
        const StreamSummary &m = sums[i];
        const StreamDesc &sd = streams[i];
        return !m.fail && m.first_ok && m.n_samples == sd.total * sd.si_channels && m.channels == sd.si_channels;
    }
};

// Fast path: hinted frame records for every candidate -> K2 over all of them -> K3,
// the chain rules verified on the device from the decoded frame ends, the per-frame CRC-16 on a side
// stream, K4 per stream. One host synchronisation at the very end (two when an MSB-aligned stream
// is hashed). Returns a CUDA-level status; per-stream acceptance is g.accepted(i).
int run_fast_group_once(FastGroup &g, Events &ev);
int run_fast_group(FastGroup &g, Events &ev) {
    int r = run_fast_group_once(g, ev);
    if (r == ZFLACB200_OK && g.narrow_overflow) {  // a residual beyond 16 bits: once more on the int32 scratch
        g.allow_narrow = false;
        g.narrow_overflow = false;
        r = run_fast_group_once(g, ev);
    }
    return r;
}
int run_fast_group_once(FastGroup &g, Events &ev) {
    cudaStream_t st = g.st;
    DeviceCtx &ctx = g_ctx[g.device];
    const uint32_t n = g.n_cand;
    const uint32_t ns = (uint32_t)g.streams.size();
    const uint32_t window = 16;
    HostTrace tr("run_fast_group");
    g.plan();
    tr.mark("plan");

    DevBuf b_sd, b_scratch, b_sum, b_flag, b_jobs, b_digest;
    SubframeStage stage;
    uint32_t max_channels = 1, max_depth = 0;
    for (const auto &sd : g.streams) {
        max_channels = std::max<uint32_t>(max_channels, sd.si_channels);
        max_depth = std::max<uint32_t>(max_depth, sd.si_depth);
    }
    g_alloc_oom = false;
    if (stage.prepare(g.wide, g.flags, n, st, max_depth, g.allow_narrow) != ZFLACB200_OK) return g_alloc_oom ? kFastOom : ZFLACB200_CUDA_ERROR;
    if (b_sd.alloc(sizeof(StreamDesc) * ns, st) || g.b_recs.alloc(sizeof(FrameRec) * (size_t)n, st) ||
        g.b_outs.alloc(sizeof(FrameOut) * (size_t)n, st) || b_scratch.alloc(g.scratch_elems * (g.wide ? 8 : 4), st) ||
        g.b_pcm.alloc(g.pcm_elems * g.container_bytes + 64, st) || b_sum.alloc(sizeof(StreamSummary) * ns, st) ||
        g.b_ok.alloc(n, st) || b_flag.alloc(sizeof(uint32_t), st) || b_jobs.alloc(sizeof(Md5Job) * ns, st) ||
        b_digest.alloc(16 * (size_t)ns, st))
        return g_alloc_oom ? kFastOom : ZFLACB200_CUDA_ERROR;
    tr.mark("alloc");
    CU_CHECK(put_small(ctx, b_sd.p, g.streams.data(), sizeof(StreamDesc) * ns, st));
    CU_CHECK(cudaMemsetAsync(b_sum.p, 0, sizeof(StreamSummary) * ns, st));
    CU_CHECK(cudaMemsetAsync(g.b_ok.p, 2, n, st));  // 2 = not checked
    CU_CHECK(cudaMemsetAsync(b_flag.p, 0, sizeof(uint32_t), st));
    CU_CHECK(cudaMemsetAsync(g.b_outs.p, 0xFF, sizeof(FrameOut) * (size_t)n, st));  // dropped slots: status != OK
    const StreamDesc *d_sd = (const StreamDesc *)b_sd.p;
    FrameRec *d_recs = (FrameRec *)g.b_recs.p;
    FrameOut *d_outs = (FrameOut *)g.b_outs.p;
    StreamSummary *d_sum = (StreamSummary *)b_sum.p;

    CU_CHECK(launch_hint_chain(g.d_cands, n, d_sd, ns, d_recs, window, st));
    CU_CHECK(cudaEventRecord(ev.e[3], st));
    CU_CHECK(cudaEventRecord(ev.e[4], st));
    if (stage.decode(g.d_blob, d_recs, n, b_scratch.p, g.wide, d_outs, true, st) != ZFLACB200_OK) return ZFLACB200_CUDA_ERROR;
    CU_CHECK(cudaEventRecord(ev.e[5], st));

    // side stream: CRC-16 of every decoded frame, overlapping K3 (+K4). A diagnostic only — the reference reads the
    // field and ignores it (:551) — but it re-reads the whole stream while K3p wants the HBM: ZFLACB200_SKIP_FRAME_CRC16
    const bool frame_crc = (g.flags & ZFLACB200_SKIP_FRAME_CRC16) == 0;
    if (frame_crc) {
        CU_CHECK(cudaEventRecord(ctx.ev_fork, st));
        CU_CHECK(cudaStreamWaitEvent(ctx.side, ctx.ev_fork, 0));
        CU_CHECK(launch_crc16_frames(g.d_blob, d_recs, d_outs, n, d_sum, (uint8_t *)g.b_ok.p, ctx.side));
        CU_CHECK(cudaEventRecord(ctx.ev_join, ctx.side));
    }

    CU_CHECK(launch_verify_chain(g.d_cands, n, d_sd, ns, d_recs, d_outs, d_sum, window, st));
    EpilogueParams ep;
    ep.frames = d_recs;
    ep.n_frames = n;
    ep.max_block_size = g.max_block;  // sizes the grid only (block sizes are known on the device); any value is correct
    ep.scratch = b_scratch.p;
    ep.wide_scratch = g.wide;
    ep.pcm = g.b_pcm.p;
    ep.container_bytes = g.container_bytes;
    ep.msb_shift = g.msb_shift;
    ep.lossy_flag = (uint32_t *)b_flag.p;
    if (stage.epilogue(ep, d_outs, max_channels, st) != ZFLACB200_OK) return ZFLACB200_CUDA_ERROR;
    CU_CHECK(cudaEventRecord(ev.e[6], st));
    tr.mark("launches");

    g.digests.assign(16 * (size_t)ns, 0);
    const void *h_digests = nullptr;
    if (g.do_md5) {
        uint32_t lossy = 0;
        if (g.msb_shift) {  // a sample wider than the STREAMINFO depth loses bits in the shift (Q5): hash before shifting
            GET_SMALL(h_lossy, ctx, b_flag.p, sizeof(uint32_t), st);
            CU_CHECK(cudaStreamSynchronize(st));
            lossy = *(const uint32_t *)h_lossy;
            if (lossy) {
                ep.msb_shift = 0;
                if (stage.epilogue(ep, d_outs, max_channels, st) != ZFLACB200_OK) return ZFLACB200_CUDA_ERROR;
            }
        }
        if (ns == 1 && md5_on_host(g.flags, g.streams[0].total * g.streams[0].si_channels * g.container_bytes)) {
            // one long stream: the chain runs on the host (hostmd5.h); the verdict of the device-side chain check is
            // read first so that a refused stream does not pay for a hash nobody will look at
            const StreamDesc &sd = g.streams[0];
            CU_CHECK(cudaEventRecord(ev.e[7], st));
            if (frame_crc) CU_CHECK(cudaStreamWaitEvent(st, ctx.ev_join, 0));  // its counters land in the same summary
            GET_SMALL(h_sums0, ctx, b_sum.p, sizeof(StreamSummary), st);
            GET_SMALL(h_ovf0, ctx, stage.two_pass ? (const void *)stage.overflow_flag() : b_flag.p, sizeof(uint32_t), st);
            CU_CHECK(cudaStreamSynchronize(st));
            g.sums.resize(1);
            memcpy(g.sums.data(), h_sums0, sizeof(StreamSummary));
            if (stage.narrow && *(const uint32_t *)h_ovf0) {
                g.narrow_overflow = true;
                return ZFLACB200_OK;
            }
            float ms_host = 0;
            if (g.accepted(0)) {
                const int r = host_md5_of_device_pcm(ctx, (const uint8_t *)g.b_pcm.p + sd.pcm_base * g.container_bytes,
                                                     sd.total * sd.si_channels, g.container_bytes, g.hash_bytes[0],
                                                     lossy ? 0 : g.msb_shift, st, g.digests.data(), &ms_host);
                if (r != ZFLACB200_OK) return r;
                if (lossy) {
                    CU_CHECK(launch_shift_inplace(g.b_pcm.p, g.pcm_elems, g.container_bytes, g.msb_shift, st));
                    CU_CHECK(cudaStreamSynchronize(st));
                }
            }
            tr.mark("host_md5");
            g.ms_hint = ev_ms(ev.e[2], ev.e[3]);
            g.ms_decode = ev_ms(ev.e[4], ev.e[5]);
            g.ms_epilogue = ev_ms(ev.e[5], ev.e[6]);
            g.ms_md5 = ms_host;
            g.md5_host = true;
            return ZFLACB200_OK;
        }
        std::vector<Md5Job> jobs(ns);
        for (uint32_t i = 0; i < ns; ++i) {
            const StreamDesc &sd = g.streams[i];
            jobs[i].pcm = (const uint8_t *)g.b_pcm.p + sd.pcm_base * g.container_bytes;
            jobs[i].n_samples = sd.total * sd.si_channels;
            jobs[i].container_bytes = g.container_bytes;
            jobs[i].hash_bytes = g.hash_bytes[i];
            jobs[i].unshift = lossy ? 0 : g.msb_shift;
            jobs[i].pad = 0;
        }
        CU_CHECK(put_small(ctx, b_jobs.p, jobs.data(), sizeof(Md5Job) * ns, st));
        CU_CHECK(launch_md5((const Md5Job *)b_jobs.p, ns, (uint8_t *)b_digest.p, st));
        if (lossy) CU_CHECK(launch_shift_inplace(g.b_pcm.p, g.pcm_elems, g.container_bytes, g.msb_shift, st));
        h_digests = get_small(ctx, b_digest.p, 16 * (size_t)ns, st);
        if (!h_digests) {
            g_last_error = "small device->host transfer failed (digests)";
            return ZFLACB200_CUDA_ERROR;
        }
    }
    CU_CHECK(cudaEventRecord(ev.e[7], st));
    if (frame_crc) CU_CHECK(cudaStreamWaitEvent(st, ctx.ev_join, 0));
    g.sums.resize(ns);
    GET_SMALL(h_sums, ctx, b_sum.p, sizeof(StreamSummary) * ns, st);
    GET_SMALL(h_ovf, ctx, stage.two_pass ? (const void *)stage.overflow_flag() : b_flag.p, sizeof(uint32_t), st);
    tr.mark("md5_setup");
    CU_CHECK(cudaStreamSynchronize(st));
    if (stage.narrow && *(const uint32_t *)h_ovf) g.narrow_overflow = true;
    if (h_digests) memcpy(g.digests.data(), h_digests, 16 * (size_t)ns);
    memcpy(g.sums.data(), h_sums, sizeof(StreamSummary) * ns);
    tr.mark("final_sync");
    g.ms_hint = ev_ms(ev.e[2], ev.e[3]);
    g.ms_decode = ev_ms(ev.e[4], ev.e[5]);
    g.ms_epilogue = ev_ms(ev.e[5], ev.e[6]);
    g.ms_md5 = ev_ms(ev.e[6], ev.e[7]);
    return ZFLACB200_OK;
}

// raw_from != kNotRaw: zflacb200_open_frames() — no signature, no metadata; the stream parameters come from the first
// frame the sync scan and the CRC-16 links agree on at or after `raw_from` (the robust path only: no stated total).
constexpr uint64_t kNotRaw = ~0ull;

int open_impl(const uint8_t *data, size_t len, const zflacb200_options *opt, zflacb200_stream **out, zflacb200_info *info,
              uint64_t raw_from = kNotRaw) {
    const bool raw = raw_from != kNotRaw;
    if ((!data && len) || !out || !info || (raw && raw_from > len)) {  // an empty slice may carry a null pointer: EndOfStream below
        g_last_error = "null argument";
        return ZFLACB200_INVALID_ARGUMENT;
    }
    *out = nullptr;
    memset(info, 0, sizeof(*info));
    const uint32_t flags = (opt ? opt->flags : 0) | env_flags();
    int device = -1;
    {
        const int e = ensure_device(opt ? opt->device : -1, &device);
        if (e != ZFLACB200_OK) return e;
    }
    cudaStream_t st = (opt && opt->stream) ? (cudaStream_t)opt->stream : g_ctx[device].stream;
    const bool on_device = (flags & ZFLACB200_INPUT_ON_DEVICE) != 0;

    Pipeline pl;
    pl.len = len;
    pl.st = st;
    pl.src.len = len;
    if (on_device) pl.src.set_device(data, len);
    else pl.src.host = data;
    pl.use_links = (flags & ZFLACB200_NO_CRC_ACCEL) == 0;

    // ---- signature + metadata (host, a few dozen bytes)
    if (raw) {
        pl.audio_offset = raw_from;  // moved to the first frame once the scan has found it
    } else {
        const int e = parse_metadata(pl.src, pl.si, pl.audio_offset);
        if (e != ZFLACB200_OK) return e;
    }
    uint32_t depth = pl.si.depth_m1 + 1;            // :256
    uint32_t aligned = (depth + 7u) & ~7u;          // :257
    uint32_t container_bytes = aligned == 8 ? 1 : aligned == 16 ? 2 : 4;  // :259-264
    memcpy(info->md5_streaminfo, pl.si.md5, 16);
    info->audio_offset = pl.audio_offset;
    info->container_bytes = (uint8_t)container_bytes;

    Events ev;
    if (ev.init() != ZFLACB200_OK) return ZFLACB200_CUDA_ERROR;
    CU_CHECK(cudaEventRecord(ev.e[0], st));

    // ---- stream bytes in HBM
    DevBuf b_stream;
    if (on_device) {
        pl.d_stream = data;
    } else {
        if (b_stream.alloc(((len + 15) & ~15ull) + 64, st) != ZFLACB200_OK) return ZFLACB200_CUDA_ERROR;
        CU_CHECK(cudaMemcpyAsync(b_stream.p, data, len, cudaMemcpyHostToDevice, st));
        pl.d_stream = (const uint8_t *)b_stream.p;
    }

    DevBuf b_probe_rec, b_probe_out;
    if (b_probe_rec.alloc(sizeof(FrameRec), st) || b_probe_out.alloc(sizeof(FrameOut), st)) return ZFLACB200_CUDA_ERROR;
    pl.d_probe_rec = (FrameRec *)b_probe_rec.p;
    pl.d_probe_out = (FrameOut *)b_probe_out.p;

    // ---- K1: sync scan -> candidates in stream order
    CU_CHECK(cudaEventRecord(ev.e[1], st));
    uint32_t n_cand = 0;
    DevBuf b_cands, b_total, b_tiles, b_ticket, b_link;
    {
        const uint32_t n_tiles = scan_tile_count(pl.audio_offset, len);
        uint32_t cap = (uint32_t)std::min<uint64_t>(len / 256 + 4096, 0x7FFFFFFFull);
        if (b_total.alloc(sizeof(uint32_t), st) || b_ticket.alloc(sizeof(uint32_t), st) ||
            b_tiles.alloc(sizeof(unsigned long long) * std::max<uint32_t>(n_tiles, 1), st))
            return ZFLACB200_CUDA_ERROR;
        for (int attempt = 0; attempt < 2; ++attempt) {
            if (b_cands.p) { dev_free(b_cands.p, st); b_cands.p = nullptr; }
            if (b_cands.alloc(sizeof(Cand) * (size_t)cap, st)) return ZFLACB200_CUDA_ERROR;
            CU_CHECK(launch_sync_scan(pl.d_stream, pl.audio_offset, len, (Cand *)b_cands.p, cap, (uint32_t *)b_total.p,
                                      (unsigned long long *)b_tiles.p, (uint32_t *)b_ticket.p, st));
            CU_CHECK(cudaMemcpyAsync(&n_cand, b_total.p, sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
            CU_CHECK(cudaStreamSynchronize(st));
            if (n_cand <= cap) break;
            cap = n_cand;  // pathological density of valid headers: rerun with room for all of them
        }
    }
    CU_CHECK(cudaEventRecord(ev.e[2], st));
    info->n_candidates = n_cand;

    {   // a 32 bps stream (side channel: 33 bits) runs on the int64 scratch; bps is latched by the first frame
        Hdr h0;
        if (host_parse_header(pl.src, pl.audio_offset, h0) == ZFLACB200_OK)
            pl.wide = (h0.bd_code == 0 ? depth : kBps[h0.bd_code]) == 32;
    }
    uint32_t msb_shift = 0;  // :287-306, keyed on the STREAMINFO depth
    if (depth >= 9 && depth <= 15) msb_shift = 16 - depth;
    else if (depth >= 17 && depth <= 31) msb_shift = 32 - depth;
    const bool do_md5 = (flags & ZFLACB200_SKIP_MD5) == 0 && !raw;  // a raw frame stream states no MD5

    // ---- fast path: device-side hinted chain, verified after decoding (no host walk, no CRC-16 pre-pass)
    if (n_cand > 0 && pl.si.total > 0 && (flags & (ZFLACB200_NO_FAST_PATH | ZFLACB200_NO_CRC_ACCEL)) == 0 &&
        pl.si.total * (pl.si.channels_m1 + 1) < (1ull << 40) && plausible_total(pl.si.total, len - std::min<uint64_t>(len, pl.audio_offset)) &&
        pl.si.total <= 65535ull * n_cand) {
        FastGroup g;
        g.device = device;
        g.st = st;
        g.d_blob = pl.d_stream;
        g.blob_len = len;
        g.d_cands = (const Cand *)b_cands.p;
        g.n_cand = n_cand;
        g.wide = pl.wide;
        g.container_bytes = container_bytes;
        g.msb_shift = msb_shift;
        g.max_block = pl.si.max_block >= 16 ? pl.si.max_block : 4096;
        g.do_md5 = do_md5;
        g.flags = flags;
        StreamDesc sd;
        memset(&sd, 0, sizeof(sd));
        sd.begin = 0;
        sd.end = len;
        sd.audio_begin = pl.audio_offset;
        sd.total = pl.si.total;
        sd.si_sample_rate = pl.si.sample_rate;
        sd.si_channels = (uint8_t)(pl.si.channels_m1 + 1);
        sd.si_depth = (uint8_t)depth;
        g.streams.push_back(sd);
        g.hash_bytes.push_back(aligned == 24 ? 3 : container_bytes);  // :268-277
        const int fast_rc = run_fast_group(g, ev);
        if (fast_rc != ZFLACB200_OK && fast_rc != kFastOom) return fast_rc;
        if (fast_rc == ZFLACB200_OK && g.accepted(0)) {
            const StreamSummary &sum = g.sums[0];
            info->ms_scan = ev_ms(ev.e[1], ev.e[2]);
            info->ms_link = g.ms_hint;
            info->ms_decode = g.ms_decode;
            info->ms_epilogue = g.ms_epilogue;
            info->ms_md5 = g.ms_md5;
            info->ms_total = ev_ms(ev.e[0], ev.e[7]) + (g.md5_host ? g.ms_md5 : 0.f);
            info->md5_checked = do_md5 ? 1 : 0;
            memcpy(info->md5_computed, g.digests.data(), 16);
            if (do_md5 && memcmp(info->md5_computed, pl.si.md5, 16) != 0) return ZFLACB200_INVALID_CHECKSUM;  // :279-280
            info->channels = sum.channels;
            info->bits_per_sample = sum.bps;
            info->sample_rate = sum.sample_rate;
            info->n_samples = sum.n_samples;
            info->pcm_bytes = sum.n_samples * container_bytes;
            info->n_frames = sum.n_frames;
            info->n_probed_frames = 0;
            info->fast_path = 1;
            info->n_crc16_mismatch = sum.n_crc16_mismatch;
            info->audio_bytes = sum.end_offset - pl.audio_offset;
            zflacb200_stream *h = new zflacb200_stream();
            h->device = device;
            h->pcm_bytes = info->pcm_bytes;
            h->d_pcm = g.b_pcm.release();
            h->d_recs = (FrameRec *)g.b_recs.release();
            h->d_outs = (FrameOut *)g.b_outs.release();
            h->d_frame_ok = (uint8_t *)g.b_ok.release();
            h->n_recs = n_cand;
            h->frames_on_device = true;
            h->info = *info;
            *out = h;
            return ZFLACB200_OK;
        }
        info->fast_path = 0;
    }

    // ---- robust path: the host walks the chain like the reference's frame loop
    // ---- K1b: CRC-16 links between neighbouring candidates
    if (b_link.alloc(sizeof(uint32_t) * std::max<uint32_t>(n_cand, 1), st)) return ZFLACB200_CUDA_ERROR;
    pl.cands.resize(n_cand);
    pl.link.assign(n_cand, kNoLink);
    if (n_cand) {
        if (pl.use_links)
            CU_CHECK(launch_frame_link(pl.d_stream, len, (const Cand *)b_cands.p, n_cand, (uint32_t *)b_link.p, 4, 16ull << 20, st));
        CU_CHECK(cudaMemcpyAsync(pl.cands.data(), b_cands.p, sizeof(Cand) * n_cand, cudaMemcpyDeviceToHost, st));
        if (pl.use_links)
            CU_CHECK(cudaMemcpyAsync(pl.link.data(), b_link.p, sizeof(uint32_t) * n_cand, cudaMemcpyDeviceToHost, st));
    }
    CU_CHECK(cudaEventRecord(ev.e[3], st));
    CU_CHECK(cudaStreamSynchronize(st));
    info->n_candidates = n_cand;

    if (raw) {  // "sync to next frame": the first header whose CRC-16 reaches another header (or the end of the data),
                // else the first header; its fields stand in for STREAMINFO the way the reference latches them (:376-388)
        if (n_cand == 0) return ZFLACB200_END_OF_STREAM;  // no frame header before the data ran out
        uint32_t c0 = 0;
        for (uint32_t i = 0; i < n_cand; ++i)
            if (pl.link[i] != kNoLink) { c0 = i; break; }
        const Cand &c = pl.cands[c0];
        if (c.bd_code == 0 || kBps[c.bd_code] == 0) return ZFLACB200_MISSING_STREAMINFO;  // "take the depth from STREAMINFO"
        if (kChCount[c.ch_code] == 0) return ZFLACB200_INVALID_FRAME_HEADER;
        pl.audio_offset = c.offset;
        pl.si.channels_m1 = kChCount[c.ch_code] - 1u;
        pl.si.depth_m1 = kBps[c.bd_code] - 1u;
        pl.si.sample_rate = c.sr_code == 0 ? 0 : c.sr_value;
        pl.si.total = 0;  // unknown: the reference's rules for a stream that states no total (:341, :394-402)
        depth = pl.si.depth_m1 + 1;
        aligned = (depth + 7u) & ~7u;
        container_bytes = aligned == 8 ? 1 : aligned == 16 ? 2 : 4;
        pl.wide = depth == 32;
        msb_shift = 0;
        if (depth >= 9 && depth <= 15) msb_shift = 16 - depth;
        else if (depth >= 17 && depth <= 31) msb_shift = 32 - depth;
        info->audio_offset = pl.audio_offset;
        info->container_bytes = (uint8_t)container_bytes;
    }

    // ---- chain + K2, repeated only if a decoded frame end contradicts a CRC-16 link
    Chain chain;
    DevBuf b_frames, b_outs, b_scratch, b_pcm;
    SubframeStage stage;
    std::vector<FrameOut> outs;
    int final_status = ZFLACB200_OK;
    bool decoded_ok = false;
    bool allow_narrow = true;
    for (int round = 0; round < 64 && !decoded_ok; ++round) {
        if (round == 63) pl.use_links = false;  // give up on links: probe everything (always consistent)
        {
            const int e = pl.build_chain(chain);
            if (e != ZFLACB200_OK) return e;
        }
        const uint32_t nf = (uint32_t)chain.frames.size();
        const bool write = chain.status == ZFLACB200_OK;
        std::vector<FrameRec> recs(nf);
        for (uint32_t i = 0; i < nf; ++i) recs[i] = chain.frames[i].rec;
        if (b_frames.p) { dev_free(b_frames.p, st); b_frames.p = nullptr; }
        if (b_outs.p) { dev_free(b_outs.p, st); b_outs.p = nullptr; }
        if (b_scratch.p) { dev_free(b_scratch.p, st); b_scratch.p = nullptr; }
        if (b_frames.alloc(sizeof(FrameRec) * std::max<uint32_t>(nf, 1), st) || b_outs.alloc(sizeof(FrameOut) * std::max<uint32_t>(nf, 1), st))
            return ZFLACB200_CUDA_ERROR;
        if (write && b_scratch.alloc(chain.scratch_elems * (pl.wide ? 8 : 4), st)) return ZFLACB200_CUDA_ERROR;
        if (stage.prepare(pl.wide, write ? flags : (flags | ZFLACB200_SINGLE_PASS), nf, st, depth, allow_narrow) != ZFLACB200_OK)
            return ZFLACB200_CUDA_ERROR;
        uint32_t narrow_overflow = 0;
        outs.resize(nf);
        if (nf) {
            CU_CHECK(cudaMemcpyAsync(b_frames.p, recs.data(), sizeof(FrameRec) * nf, cudaMemcpyHostToDevice, st));
            CU_CHECK(cudaEventRecord(ev.e[4], st));
            if (stage.decode(pl.d_stream, (const FrameRec *)b_frames.p, nf, b_scratch.p, pl.wide, (FrameOut *)b_outs.p, write, st) !=
                ZFLACB200_OK)
                return ZFLACB200_CUDA_ERROR;
            CU_CHECK(cudaEventRecord(ev.e[5], st));
            CU_CHECK(cudaMemcpyAsync(outs.data(), b_outs.p, sizeof(FrameOut) * nf, cudaMemcpyDeviceToHost, st));
            if (stage.narrow) CU_CHECK(cudaMemcpyAsync(&narrow_overflow, stage.overflow_flag(), sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
            CU_CHECK(cudaStreamSynchronize(st));
        } else {
            CU_CHECK(cudaEventRecord(ev.e[4], st));
            CU_CHECK(cudaEventRecord(ev.e[5], st));
        }
        if (narrow_overflow) {  // a residual beyond 16 bits: once more on the int32 scratch
            allow_narrow = false;
            continue;
        }
        decoded_ok = true;
        for (uint32_t i = 0; i < nf; ++i) {
            if (outs[i].status != ST_OK) {  // first failing frame in stream order wins
                final_status = (int)outs[i].status;
                break;
            }
            if (outs[i].end_offset != chain.frames[i].expect_end) {  // link disagreed with the decode: decode is the truth
                pl.known[chain.frames[i].rec.offset] = outs[i];
                decoded_ok = false;
                break;
            }
        }
        if (final_status != ZFLACB200_OK) return final_status;
    }
    if (!decoded_ok) {
        g_last_error = "frame chain did not converge";
        return ZFLACB200_CUDA_ERROR;
    }
    if (chain.status != ZFLACB200_OK) return chain.status;

    const uint32_t nf = (uint32_t)chain.frames.size();
    const uint64_t pcm_bytes = chain.n_samples * container_bytes;
    if (b_pcm.alloc(pcm_bytes + 64, st)) return ZFLACB200_CUDA_ERROR;

    // ---- K3: decorrelate + shift + interleave
    DevBuf b_flag, b_jobs, b_digest;
    if (b_flag.alloc(sizeof(uint32_t), st) || b_jobs.alloc(sizeof(Md5Job), st) || b_digest.alloc(16, st)) return ZFLACB200_CUDA_ERROR;
    CU_CHECK(cudaMemsetAsync(b_flag.p, 0, sizeof(uint32_t), st));
    EpilogueParams ep;
    ep.frames = (const FrameRec *)b_frames.p;
    ep.n_frames = nf;
    ep.max_block_size = chain.max_block;
    ep.scratch = b_scratch.p;
    ep.wide_scratch = pl.wide;
    ep.pcm = b_pcm.p;
    ep.container_bytes = container_bytes;
    ep.msb_shift = msb_shift;
    ep.lossy_flag = (uint32_t *)b_flag.p;
    if (stage.epilogue(ep, (const FrameOut *)b_outs.p, chain.channels, st) != ZFLACB200_OK) return ZFLACB200_CUDA_ERROR;
    CU_CHECK(cudaEventRecord(ev.e[6], st));

    // ---- K4: MD5 gate
    uint32_t lossy = 0;
    bool md5_host = false;
    float ms_md5_host = 0;
    if (do_md5) {
        if (msb_shift) {
            CU_CHECK(cudaMemcpyAsync(&lossy, b_flag.p, sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
            CU_CHECK(cudaStreamSynchronize(st));
            if (lossy) {  // a sample exceeds the STREAMINFO depth (Q5): hash the unshifted values, shift afterwards
                ep.msb_shift = 0;
                if (stage.epilogue(ep, (const FrameOut *)b_outs.p, chain.channels, st) != ZFLACB200_OK) return ZFLACB200_CUDA_ERROR;
            }
        }
        if (md5_on_host(flags, pcm_bytes)) {  // one long stream: the chain runs on the host (hostmd5.h)
            CU_CHECK(cudaEventRecord(ev.e[7], st));
            const int r = host_md5_of_device_pcm(g_ctx[device], b_pcm.p, chain.n_samples, container_bytes,
                                                 aligned == 24 ? 3 : container_bytes, lossy ? 0 : msb_shift, st, info->md5_computed,
                                                 &ms_md5_host);
            if (r != ZFLACB200_OK) return r;
            md5_host = true;
            if (lossy) CU_CHECK(launch_shift_inplace(b_pcm.p, chain.n_samples, container_bytes, msb_shift, st));
        } else {
            Md5Job job;
            job.pcm = b_pcm.p;
            job.n_samples = chain.n_samples;
            job.container_bytes = container_bytes;
            job.hash_bytes = aligned == 24 ? 3 : container_bytes;  // :268-277
            job.unshift = lossy ? 0 : msb_shift;
            job.pad = 0;
            CU_CHECK(cudaMemcpyAsync(b_jobs.p, &job, sizeof(job), cudaMemcpyHostToDevice, st));
            CU_CHECK(launch_md5((const Md5Job *)b_jobs.p, 1, (uint8_t *)b_digest.p, st));
            if (lossy) CU_CHECK(launch_shift_inplace(b_pcm.p, chain.n_samples, container_bytes, msb_shift, st));
            CU_CHECK(cudaMemcpyAsync(info->md5_computed, b_digest.p, 16, cudaMemcpyDeviceToHost, st));
        }
    }
    if (!md5_host) CU_CHECK(cudaEventRecord(ev.e[7], st));
    CU_CHECK(cudaStreamSynchronize(st));

    info->ms_scan = ev_ms(ev.e[1], ev.e[2]);
    info->ms_link = ev_ms(ev.e[2], ev.e[3]);
    info->ms_decode = ev_ms(ev.e[4], ev.e[5]);
    info->ms_epilogue = ev_ms(ev.e[5], ev.e[6]);
    info->ms_md5 = md5_host ? ms_md5_host : ev_ms(ev.e[6], ev.e[7]);
    info->ms_total = ev_ms(ev.e[0], ev.e[7]) + (md5_host ? ms_md5_host : 0.f);
    info->md5_checked = do_md5 ? 1 : 0;
    if (do_md5 && memcmp(info->md5_computed, pl.si.md5, 16) != 0) return ZFLACB200_INVALID_CHECKSUM;  // :279-280

    info->channels = chain.channels;
    info->bits_per_sample = chain.bps;
    info->sample_rate = chain.sample_rate;
    info->n_samples = chain.n_samples;
    info->pcm_bytes = pcm_bytes;
    info->n_frames = nf;
    info->n_probed_frames = 0;  // frames whose end came from decoding them, in whichever round of the loop above
    for (const auto &cf : chain.frames) info->n_probed_frames += cf.probed ? 1 : 0;
    info->audio_bytes = nf ? chain.frames.back().expect_end - pl.audio_offset : 0;

    zflacb200_stream *h = new zflacb200_stream();
    h->device = device;
    h->pcm_bytes = pcm_bytes;
    h->frames.resize(nf);
    for (uint32_t i = 0; i < nf; ++i) {
        zflacb200_frame &f = h->frames[i];
        const ChainFrame &cf = chain.frames[i];
        f.offset = cf.rec.offset;
        f.end_offset = cf.expect_end;
        f.sample_offset = cf.rec.sample_base;
        f.block_size = cf.rec.block_size;
        f.header_len = cf.rec.header_len;
        f.channel_code = cf.rec.ch_code;
        f.crc16_ok = cf.probed ? 2 : 1;  // 2 = not verified by a link (see n_crc16_mismatch)
        f.probed = cf.probed ? 1 : 0;
    }
    // CRC-16 diagnostic for frames that were not linked (the reference ignores the CRC, Q1)
    {
        std::vector<uint64_t> rb, re;
        std::vector<uint32_t> idx;
        for (uint32_t i = 0; i < nf; ++i)
            if (chain.frames[i].probed) {
                rb.push_back(chain.frames[i].rec.offset);
                re.push_back(chain.frames[i].expect_end - 2);
                idx.push_back(i);
            }
        if (!rb.empty()) {
            DevBuf b_rb, b_re, b_crc;
            const size_t m = rb.size();
            if (b_rb.alloc(8 * m, st) || b_re.alloc(8 * m, st) || b_crc.alloc(2 * m, st)) return ZFLACB200_CUDA_ERROR;
            std::vector<uint16_t> crc(m);
            CU_CHECK(cudaMemcpyAsync(b_rb.p, rb.data(), 8 * m, cudaMemcpyHostToDevice, st));
            CU_CHECK(cudaMemcpyAsync(b_re.p, re.data(), 8 * m, cudaMemcpyHostToDevice, st));
            CU_CHECK(launch_crc16_ranges(pl.d_stream, (const uint64_t *)b_rb.p, (const uint64_t *)b_re.p, (uint32_t)m, (uint16_t *)b_crc.p, st));
            CU_CHECK(cudaMemcpyAsync(crc.data(), b_crc.p, 2 * m, cudaMemcpyDeviceToHost, st));
            CU_CHECK(cudaStreamSynchronize(st));
            for (size_t k = 0; k < m; ++k) {
                const bool ok = crc[k] == (uint16_t)outs[idx[k]].stored_crc16;
                h->frames[idx[k]].crc16_ok = ok ? 1 : 0;
                if (!ok) ++info->n_crc16_mismatch;
            }
        }
    }
    h->d_pcm = b_pcm.release();
    h->info = *info;
    *out = h;
    return ZFLACB200_OK;
}


// File-parallel batch: metadata walk per stream on the host, ONE sync scan over the packed bytes,
// then one fast-path launch group per (scratch type, container, MSB shift) class — usually one.
int batch_open_impl(const uint8_t *blob, const uint8_t *device_blob, const uint64_t *offsets, const uint64_t *lens, size_t n,
                    const zflacb200_options *opt, zflacb200_batch **out, zflacb200_info *infos, int *statuses) {
    if (!blob || !offsets || !lens || !out || !infos || !statuses) {
        g_last_error = "null argument";
        return ZFLACB200_INVALID_ARGUMENT;
    }
    *out = nullptr;
    const uint32_t flags = (opt ? opt->flags : 0) | env_flags();
    int device = -1;
    {
        const int e = ensure_device(opt ? opt->device : -1, &device);
        if (e != ZFLACB200_OK) return e;
    }
    cudaStream_t st = (opt && opt->stream) ? (cudaStream_t)opt->stream : g_ctx[device].stream;
    const bool do_md5 = (flags & ZFLACB200_SKIP_MD5) == 0;

    struct Item {
        Streaminfo si;
        uint64_t audio_offset = 0;  // relative to the stream
        uint32_t depth = 0, aligned = 0, container = 0, msb_shift = 0;
        bool wide = false, eligible = false;
        int group = -1;
        uint32_t slot = 0;  // index inside the group
    };
    HostTrace tr("batch_open");
    std::vector<Item> items(n);
    uint64_t blob_len = 0, scan_begin = ~0ull;
    std::vector<uint32_t> keys;  // distinct (channels, wide, container, shift) classes

    // ---- K1 ahead of the metadata walk. The scan needs nothing the walk finds (it starts at the first stream's first byte:
    // the metadata of every later stream was always scanned, a chance match in there is no frame of any chain), and the
    // walk of 10 000 streams is a millisecond of host time the GPU used to wait out (r07d). Only when some stream looks
    // like a fast-path candidate (signature + a stated total), so that a batch of other streams uploads nothing twice.
    Events ev;
    DevBuf b_blob, b_cands, b_total, b_tiles, b_ticket;
    const uint8_t *d_blob = device_blob;
    uint32_t scan_cap = 0;
    const void *h_total_early = nullptr;
    bool early = false;
    {
        uint64_t first_off = ~0ull, len_pre = 0;
        bool maybe_fast = false;
        for (size_t i = 0; i < n; ++i) {
            len_pre = std::max(len_pre, offsets[i] + lens[i]);
            first_off = std::min(first_off, offsets[i]);
            if (!maybe_fast && lens[i] >= 42) {
                const uint8_t *p = blob + offsets[i];  // STREAMINFO is the first block: its 36-bit total is at bytes 21..25
                maybe_fast = p[0] == 0x66 && p[1] == 0x4C && p[2] == 0x61 && p[3] == 0x43 && ((p[21] & 0x0F) | p[22] | p[23] | p[24] | p[25]) != 0;
            }
        }
        early = maybe_fast && n > 0 && (flags & (ZFLACB200_NO_FAST_PATH | ZFLACB200_NO_CRC_ACCEL)) == 0;
        if (early) {
            if (ev.init() != ZFLACB200_OK) return ZFLACB200_CUDA_ERROR;
            CU_CHECK(cudaEventRecord(ev.e[0], st));
            if (!d_blob) {
                if (b_blob.alloc(((len_pre + 15) & ~15ull) + 64, st) != ZFLACB200_OK) return ZFLACB200_CUDA_ERROR;
                CU_CHECK(cudaMemcpyAsync(b_blob.p, blob, len_pre, cudaMemcpyHostToDevice, st));
                d_blob = (const uint8_t *)b_blob.p;
            }
            CU_CHECK(cudaEventRecord(ev.e[1], st));
            scan_begin = first_off;
            scan_cap = (uint32_t)std::min<uint64_t>(len_pre / 256 + 4096 + 4 * n, 0x7FFFFFFFull);
            if (b_total.alloc(sizeof(uint32_t), st) || b_ticket.alloc(sizeof(uint32_t), st) ||
                b_tiles.alloc(sizeof(unsigned long long) * std::max<uint32_t>(scan_tile_count(scan_begin, len_pre), 1), st) ||
                b_cands.alloc(sizeof(Cand) * (size_t)scan_cap, st))
                return ZFLACB200_CUDA_ERROR;
            CU_CHECK(launch_sync_scan(d_blob, scan_begin, len_pre, (Cand *)b_cands.p, scan_cap, (uint32_t *)b_total.p,
                                      (unsigned long long *)b_tiles.p, (uint32_t *)b_ticket.p, st));
            h_total_early = get_small(g_ctx[device], b_total.p, sizeof(uint32_t), st);
            if (!h_total_early) {
                g_last_error = "small device->host transfer failed (candidate count)";
                return ZFLACB200_CUDA_ERROR;
            }
            CU_CHECK(cudaEventRecord(ev.e[2], st));  // ms_scan stays the scan's time, not the walk's
            tr.mark("scan_enqueue");
        }
    }
    const uint64_t early_scan_begin = scan_begin;
    // The walk of one stream touches nothing but that stream's slots, so a large batch is walked by a few host threads
    // (10 000 streams: ~1.0 ms on one core); what orders the streams — the class list and the scan start — follows serially.
    auto walk = [&](size_t i) {
        memset(&infos[i], 0, sizeof(zflacb200_info));
        Item &it = items[i];
        ByteSource src;
        src.host = blob + offsets[i];
        src.len = lens[i];
        statuses[i] = parse_metadata(src, it.si, it.audio_offset);
        if (statuses[i] != ZFLACB200_OK) return;
        it.depth = it.si.depth_m1 + 1;
        it.aligned = (it.depth + 7u) & ~7u;
        it.container = it.aligned == 8 ? 1 : it.aligned == 16 ? 2 : 4;
        if (it.depth >= 9 && it.depth <= 15) it.msb_shift = 16 - it.depth;
        else if (it.depth >= 17 && it.depth <= 31) it.msb_shift = 32 - it.depth;
        memcpy(infos[i].md5_streaminfo, it.si.md5, 16);
        infos[i].audio_offset = it.audio_offset;
        infos[i].container_bytes = (uint8_t)it.container;
        Hdr h0;
        if (host_parse_header(src, it.audio_offset, h0) == ZFLACB200_OK)
            it.wide = (h0.bd_code == 0 ? it.depth : kBps[h0.bd_code]) == 32;
        it.eligible = it.si.total > 0 && it.si.total * (it.si.channels_m1 + 1) < (1ull << 40) &&
                      plausible_total(it.si.total, lens[i] - std::min<uint64_t>(lens[i], it.audio_offset)) &&
                      (flags & (ZFLACB200_NO_FAST_PATH | ZFLACB200_NO_CRC_ACCEL)) == 0;
    };
    {
        const unsigned hw = std::thread::hardware_concurrency();
        const size_t n_threads = n >= 4096 ? std::min<size_t>(4, hw > 1 ? hw : 1) : 1;
        if (n_threads <= 1) {
            for (size_t i = 0; i < n; ++i) walk(i);
        } else {
            std::vector<std::thread> pool;
            for (size_t t = 1; t < n_threads; ++t)
                pool.emplace_back([&, t] { for (size_t i = n * t / n_threads; i < n * (t + 1) / n_threads; ++i) walk(i); });
            for (size_t i = 0; i < n / n_threads; ++i) walk(i);
            for (auto &th : pool) th.join();
        }
    }
    for (size_t i = 0; i < n; ++i) {
        blob_len = std::max(blob_len, offsets[i] + lens[i]);
        Item &it = items[i];
        if (statuses[i] != ZFLACB200_OK || !it.eligible) continue;
        const uint32_t key = ((uint32_t)it.si.channels_m1 << 20) | (it.wide ? 1u << 16 : 0u) | (it.container << 8) | it.msb_shift;
        size_t g = 0;
        while (g < keys.size() && keys[g] != key) ++g;
        if (g == keys.size()) keys.push_back(key);
        it.group = (int)g;
        scan_begin = std::min(scan_begin, offsets[i] + it.audio_offset);
    }

    tr.mark("metadata");
    zflacb200_batch *b = new zflacb200_batch();
    b->device = device;
    b->items.resize(n);
    struct Cleanup {  // on an early CUDA error nothing is returned
        zflacb200_batch *b;
        bool armed = true;
        ~Cleanup() {
            if (armed) {
                cudaDeviceSynchronize();  // error path: nothing may still be writing the blocks released below
                for (auto &it : b->items)
                    if (it.solo) { dev_free(it.solo->d_pcm, nullptr); delete it.solo; }
                for (auto &g : b->groups)
                    if (g.d_pcm) dev_free(g.d_pcm, nullptr);
                delete b;
            }
        }
    } cleanup{b};

    if (early) scan_begin = early_scan_begin;  // what the scan that is already running covers
    if (early && keys.empty()) CU_CHECK(cudaStreamSynchronize(st));  // nothing will use the scan: let it finish before its buffers go
    if (!keys.empty()) {
        if (!early) {
            if (ev.init() != ZFLACB200_OK) return ZFLACB200_CUDA_ERROR;
            CU_CHECK(cudaEventRecord(ev.e[0], st));
            if (!d_blob) {
                if (b_blob.alloc(((blob_len + 15) & ~15ull) + 64, st) != ZFLACB200_OK) return ZFLACB200_CUDA_ERROR;
                tr.mark("blob_alloc");
                CU_CHECK(cudaMemcpyAsync(b_blob.p, blob, blob_len, cudaMemcpyHostToDevice, st));
                tr.mark("h2d_enqueue");
                if (tr.on) { cudaStreamSynchronize(st); tr.mark("h2d_wait"); }
                d_blob = (const uint8_t *)b_blob.p;
            }
            CU_CHECK(cudaEventRecord(ev.e[1], st));
        }
        // ---- K1 once over the packed bytes (already on its way when `early`)
        uint32_t n_cand = 0;
        {
            uint32_t cap = early ? scan_cap : (uint32_t)std::min<uint64_t>(blob_len / 256 + 4096 + 4 * n, 0x7FFFFFFFull);
            if (!early) {
                const uint32_t n_tiles = scan_tile_count(scan_begin, blob_len);
                if (b_total.alloc(sizeof(uint32_t), st) || b_ticket.alloc(sizeof(uint32_t), st) ||
                    b_tiles.alloc(sizeof(unsigned long long) * std::max<uint32_t>(n_tiles, 1), st))
                    return ZFLACB200_CUDA_ERROR;
            }
            bool rescanned = false;
            for (int attempt = 0; attempt < 2; ++attempt) {
                const void *h_total = h_total_early;
                if (!(early && attempt == 0)) {
                    rescanned = true;
                    if (b_cands.p) { dev_free(b_cands.p, st); b_cands.p = nullptr; }
                    if (b_cands.alloc(sizeof(Cand) * (size_t)cap, st)) return ZFLACB200_CUDA_ERROR;
                    CU_CHECK(launch_sync_scan(d_blob, scan_begin, blob_len, (Cand *)b_cands.p, cap, (uint32_t *)b_total.p,
                                              (unsigned long long *)b_tiles.p, (uint32_t *)b_ticket.p, st));
                    h_total = get_small(g_ctx[device], b_total.p, sizeof(uint32_t), st);
                    if (!h_total) {
                        g_last_error = "small device->host transfer failed (candidate count)";
                        return ZFLACB200_CUDA_ERROR;
                    }
                }
                CU_CHECK(cudaStreamSynchronize(st));
                n_cand = *(const uint32_t *)h_total;
                if (n_cand <= cap) break;
                cap = n_cand;
            }
            if (!early || rescanned) CU_CHECK(cudaEventRecord(ev.e[2], st));
        }
        tr.mark("h2d+scan");

        // ---- one launch group per class
        for (size_t gi = 0; gi < keys.size() && n_cand > 0; ++gi) {
            FastGroup g;
            g.device = device;
            g.st = st;
            g.d_blob = d_blob;
            g.blob_len = blob_len;
            g.d_cands = (const Cand *)b_cands.p;
            g.n_cand = n_cand;
            g.wide = (keys[gi] >> 16) & 1;
            g.container_bytes = (keys[gi] >> 8) & 0xFF;
            g.msb_shift = keys[gi] & 0xFF;
            g.do_md5 = do_md5;
            g.flags = flags;
            g.max_block = 16;
            std::vector<size_t> members;
            for (size_t i = 0; i < n; ++i) {  // streams in blob order (the device searches them by offset)
                if (items[i].group != (int)gi) continue;
                members.push_back(i);
            }
            std::sort(members.begin(), members.end(), [&](size_t x, size_t y) { return offsets[x] < offsets[y]; });
            for (size_t i : members) {
                Item &it = items[i];
                StreamDesc sd;
                memset(&sd, 0, sizeof(sd));
                sd.begin = offsets[i];
                sd.end = offsets[i] + lens[i];
                sd.audio_begin = offsets[i] + it.audio_offset;
                sd.total = it.si.total;
                sd.si_sample_rate = it.si.sample_rate;
                sd.si_channels = (uint8_t)(it.si.channels_m1 + 1);
                sd.si_depth = (uint8_t)it.depth;
                it.slot = (uint32_t)g.streams.size();
                g.streams.push_back(sd);
                g.hash_bytes.push_back(it.aligned == 24 ? 3 : it.container);
                g.max_block = std::max<uint32_t>(g.max_block, it.si.max_block >= 16 ? it.si.max_block : 4096);
            }
            tr.mark("group_setup");
            {
                const int r = run_fast_group(g, ev);
                if (r == kFastOom) {  // the group's buffers do not fit: its streams are decoded one by one on the robust path
                    for (size_t i : members) items[i].eligible = false;
                    continue;
                }
                if (r != ZFLACB200_OK) return r;
            }
            tr.mark("run_fast_group");
            const float ms_total = ev_ms(ev.e[0], ev.e[7]);
            const float ms_scan = ev_ms(ev.e[1], ev.e[2]);  // once: an event query per stream costs more than the decode
            zflacb200_batch::Group grp;
            grp.bytes = g.pcm_elems * g.container_bytes;
            grp.d_pcm = g.b_pcm.release();
            b->groups.push_back(grp);
            for (size_t i : members) {
                Item &it = items[i];
                if (!g.accepted(it.slot)) {  // the host walks this one
                    it.eligible = false;
                    continue;
                }
                const StreamSummary &sum = g.sums[it.slot];
                const StreamDesc &sd = g.streams[it.slot];
                zflacb200_info &inf = infos[i];
                inf.md5_checked = do_md5 ? 1 : 0;
                memcpy(inf.md5_computed, g.digests.data() + 16 * (size_t)it.slot, 16);
                inf.ms_scan = ms_scan;
                inf.ms_link = g.ms_hint;
                inf.ms_decode = g.ms_decode;
                inf.ms_epilogue = g.ms_epilogue;
                inf.ms_md5 = g.ms_md5;
                inf.ms_total = ms_total;
                if (do_md5 && memcmp(inf.md5_computed, it.si.md5, 16) != 0) {
                    statuses[i] = ZFLACB200_INVALID_CHECKSUM;  // :279-280: no data is returned
                    continue;
                }
                inf.channels = sum.channels;
                inf.bits_per_sample = sum.bps;
                inf.sample_rate = sum.sample_rate;
                inf.n_samples = sum.n_samples;
                inf.pcm_bytes = sum.n_samples * g.container_bytes;
                inf.n_frames = sum.n_frames;
                inf.n_candidates = n_cand;
                inf.fast_path = 1;
                inf.n_crc16_mismatch = sum.n_crc16_mismatch;
                inf.audio_bytes = sum.end_offset - sd.audio_begin;
                zflacb200_batch::Item &bi = b->items[i];
                bi.group = (int)b->groups.size() - 1;
                bi.byte_off = sd.pcm_base * g.container_bytes;
                bi.bytes = inf.pcm_bytes;
            }
        }
    }

    tr.mark("collect");
    // ---- everything the device-side chain did not take: one by one, chain walked on the host
    zflacb200_options solo_opt;
    solo_opt.struct_size = sizeof(solo_opt);
    solo_opt.flags = (flags | ZFLACB200_NO_FAST_PATH) & ~ZFLACB200_INPUT_ON_DEVICE;
    solo_opt.device = device;
    solo_opt.stream = opt ? opt->stream : nullptr;
    for (size_t i = 0; i < n; ++i) {
        if (statuses[i] != ZFLACB200_OK || b->items[i].group >= 0) continue;
        zflacb200_stream *h = nullptr;
        g_alloc_oom = false;
        statuses[i] = open_impl(blob + offsets[i], lens[i], &solo_opt, &h, &infos[i]);
        if (statuses[i] == ZFLACB200_CUDA_ERROR && g_alloc_oom) {  // this stream does not fit the device: the others go on
            statuses[i] = ZFLACB200_OUT_OF_MEMORY;
            g_alloc_oom = false;
            continue;
        }
        if (statuses[i] == ZFLACB200_CUDA_ERROR) return ZFLACB200_CUDA_ERROR;
        if (statuses[i] == ZFLACB200_OK) {
            b->items[i].solo = h;
            b->items[i].bytes = h->pcm_bytes;
        }
    }
    uint64_t total = 0;
    for (auto &g : b->groups) total += (g.bytes + 63) & ~63ull;
    for (auto &it : b->items)
        if (it.solo) total += (it.bytes + 63) & ~63ull;
    b->read_all_bytes = total;
    cleanup.armed = false;
    *out = b;
    return ZFLACB200_OK;
}


// ------------------------------------------------------------------ zflacb200_decode, pipelined
// Host stream in, host PCM out: the stream is uploaded in chunks and every chunk runs the fast path
// (scan -> hinted frame records -> K2a/K3p) as soon as it and its successor have landed, while later
// chunks are still on the wire and the PCM of earlier chunks already travels back — both PCIe
// directions and the SMs busy at once. The hinted chain makes chunks independent (a frame's PCM offset
// comes from its coded number); the chain rules are verified over ALL frames at the end exactly as in
// run_fast_group, and anything unexpected (chain not accepted, more candidates than planned, a lossy
// MSB shift, a destination that is too small) returns kNotPipelined: the caller then decodes the
// stream the plain way, which overwrites whatever was already copied out.
constexpr int kNotPipelined = -1;

struct EventList {
    std::vector<cudaEvent_t> v;
    int make(size_t n) {
        v.assign(n, nullptr);
        for (auto &e : v) CU_CHECK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
        return ZFLACB200_OK;
    }
    ~EventList() {
        for (auto e : v)
            if (e) cudaEventDestroy(e);
    }
};

// Two forms: host stream in / host PCM out (zflacb200_decode: `dst`), and device stream in / PCM left in HBM behind a
// handle (zflacb200_open with ZFLACB200_INPUT_ON_DEVICE: `handle_out`). The second has no copies to overlap, but the
// chunks still pay: chunk i's frame-serial kernels start as soon as ITS scan is done instead of after the scan of the
// whole stream, and the HBM-bound K3p of early chunks runs under the latency-bound K2a of later ones. Measured (r02j,
// C2): 2.04 ms with 3-4 chunks, 2.12 with 6, 2.67 with 12 against 1.96 ms unchunked — every chunk costs a host round trip
// for its candidate count (~0.1 ms with the launches behind it), which only the PCIe-bound host form hides. So the
// device form is opt-in (ZFLACB200_DEVICE_CHUNKS=n) until the count stays on the device.
int decode_pipelined(const uint8_t *data, size_t len, const zflacb200_options *opt, void *dst, size_t dst_capacity,
                     zflacb200_info *info, zflacb200_stream **handle_out = nullptr) {
    const uint32_t flags = (opt ? opt->flags : 0) | env_flags();
    if (flags & (ZFLACB200_NO_FAST_PATH | ZFLACB200_NO_CRC_ACCEL | ZFLACB200_NO_PIPELINE)) return kNotPipelined;
    const bool on_device = (flags & ZFLACB200_INPUT_ON_DEVICE) != 0;
    if (on_device != (handle_out != nullptr)) return kNotPipelined;
    // test knobs: ZFLACB200_PIPELINE_MIN_BYTES (default 48 MiB), ZFLACB200_PIPELINE_CHUNK (default: len / 24, >= 8 MiB)
    static const uint64_t min_bytes = [] { const char *e = getenv("ZFLACB200_PIPELINE_MIN_BYTES"); return e ? strtoull(e, nullptr, 0) : (48ull << 20); }();
    static const uint64_t env_chunk = [] { const char *e = getenv("ZFLACB200_PIPELINE_CHUNK"); return e ? strtoull(e, nullptr, 0) : 0ull; }();
    static const uint64_t dev_chunks = [] { const char *e = getenv("ZFLACB200_DEVICE_CHUNKS"); return e ? strtoull(e, nullptr, 0) : 0ull; }();  // off by default, see below
    if (!data || (!dst && !on_device) || !info || len < min_bytes || len < 4096) return kNotPipelined;  // short streams: nothing to overlap
    if (on_device && dev_chunks < 2) return kNotPipelined;
    int device = -1;
    if (ensure_device(opt ? opt->device : -1, &device) != ZFLACB200_OK) return kNotPipelined;
    DeviceCtx &ctx = g_ctx[device];
    cudaStream_t st = (opt && opt->stream) ? (cudaStream_t)opt->stream : ctx.stream;
    HostTrace tr("decode_pipelined");

    ByteSource src;
    if (on_device) src.set_device(data, len);
    else src.host = data;
    src.len = len;
    Streaminfo si;
    uint64_t audio_offset = 0;
    if (parse_metadata(src, si, audio_offset) != ZFLACB200_OK) return kNotPipelined;  // the plain path reports the error
    const uint32_t C = si.channels_m1 + 1;
    const uint32_t depth = si.depth_m1 + 1, aligned = (depth + 7u) & ~7u;
    const uint32_t cb = aligned == 8 ? 1 : aligned == 16 ? 2 : 4;
    uint32_t msb_shift = 0;
    if (depth >= 9 && depth <= 15) msb_shift = 16 - depth;
    else if (depth >= 17 && depth <= 31) msb_shift = 32 - depth;
    if (si.total == 0 || si.total * C >= (1ull << 40)) return kNotPipelined;
    const uint64_t n_elems = si.total * C, pcm_bytes = n_elems * cb;
    if (!on_device && dst_capacity < pcm_bytes) return kNotPipelined;
    bool wide = false;
    {
        Hdr h0;
        if (host_parse_header(src, audio_offset, h0) != ZFLACB200_OK) return kNotPipelined;
        wide = (h0.bd_code == 0 ? depth : kBps[h0.bd_code]) == 32;
    }
    const bool do_md5 = (flags & ZFLACB200_SKIP_MD5) == 0;

    // ---- chunks: >= 8 MiB (a frame is < 2.1 MB: one that starts in chunk i ends inside chunk i + 1), about 24 of them
    uint64_t chunk = std::max<uint64_t>(8ull << 20, (len / (on_device ? dev_chunks : 24) + 0xFFFF) & ~0xFFFFull);
    if (env_chunk >= 4096) chunk = (env_chunk + 15) & ~15ull;  // a frame longer than a chunk only costs the fallback
    // The first chunks are smaller (1/8, 1/8, 1/4, 1/2 of a chunk): nothing travels back until the first chunk is uploaded
    // AND decoded, and a chunk's kernels take ~0.9 ms whatever its size (they are latency-bound) — with full-size chunks the
    // read-back, the long pole of the call, started 1.3 ms late. No chunk is smaller than the largest frame the format
    // allows (65535 samples x 8 channels x 33 bits < 2.25 MiB), so "a frame that starts in chunk i ends inside chunk i + 1"
    // holds whatever STREAMINFO claims. Only for host streams (the device form has no copies to overlap), not under the test knob.
    static const bool no_ramp = getenv("ZFLACB200_PIPELINE_NO_RAMP") != nullptr;  // diagnostics: A/B runs
    const uint64_t small = std::max<uint64_t>((chunk / 8 + 0xFFFF) & ~0xFFFFull, 0x240000ull);
    const bool ramp = !no_ramp && !on_device && env_chunk < 4096 && audio_offset < small && 2 * small <= chunk && len > 4 * chunk;
    std::vector<uint64_t> b;
    b.push_back(audio_offset);
    {
        uint64_t at = 0;
        if (ramp) {
            const uint64_t steps[4] = {small, small, 2 * small, 4 * small};
            for (uint64_t sz : steps) {
                at += sz;
                b.push_back(at);
            }
        }
        while (at < len) {
            at = std::min<uint64_t>(len, at + chunk);
            b.push_back(at);
        }
    }
    if (audio_offset >= b[1]) return kNotPipelined;
    const uint32_t K = (uint32_t)b.size() - 1;

    // ---- device buffers for the whole stream
    const uint32_t min_bs = std::max<uint32_t>(si.min_block, 16);
    const uint64_t cap64 = (si.total / min_bs + 64) * 5 / 4 + 4096;  // frames by STREAMINFO + room for false candidates
    if (cap64 > 0x3FFFFFFFull || K > 4096) return kNotPipelined;
    const uint32_t cap = (uint32_t)cap64;
    constexpr int S = DeviceCtx::kLanes;
    const uint32_t chunk_tiles = scan_tile_count(0, chunk + 64) + 2;
    const uint64_t scratch_elems = (uint64_t)C * (((si.total + 7) & ~7ull) + 8) + 64ull * cap + 64;
    DevBuf b_stream, b_cands, b_total, b_tiles, b_ticket, b_sd, b_recs, b_outs, b_ok, b_scratch, b_pcm, b_sum, b_flag, b_hi, b_jobs, b_digest;
    SubframeStage stage;
    if ((!on_device && b_stream.alloc(((len + 15) & ~15ull) + 64, st)) || b_cands.alloc(sizeof(Cand) * (size_t)cap, st) ||
        b_total.alloc(sizeof(uint32_t), st) || b_tiles.alloc(sizeof(unsigned long long) * chunk_tiles, st) ||
        b_ticket.alloc(sizeof(uint32_t), st) || b_sd.alloc(sizeof(StreamDesc), st) || b_recs.alloc(sizeof(FrameRec) * (size_t)cap, st) ||
        b_outs.alloc(sizeof(FrameOut) * (size_t)cap, st) || b_ok.alloc(cap, st) || b_scratch.alloc(scratch_elems * (wide ? 8 : 4), st) ||
        b_pcm.alloc(((n_elems + 15) & ~15ull) * cb + 64, st) || b_sum.alloc(sizeof(StreamSummary), st) ||
        b_flag.alloc(sizeof(uint32_t), st) || b_hi.alloc(sizeof(unsigned long long) * K, st) || b_jobs.alloc(sizeof(Md5Job), st) ||
        b_digest.alloc(16, st) || stage.prepare(wide, flags, cap, st, depth) != ZFLACB200_OK)
        return ZFLACB200_CUDA_ERROR;
    // Declared after the buffers, so it runs before they are released: an early return (a CUDA error, a bail-out) must not
    // hand blocks back to the cache while a lane still decodes into them or a copy still reads them.
    struct Quiesce {
        DeviceCtx &ctx;
        bool armed = true;
        ~Quiesce() {
            if (!armed) return;
            for (auto &l : ctx.lane) cudaStreamSynchronize(l);
            cudaStreamSynchronize(ctx.copy_in);
            cudaStreamSynchronize(ctx.copy_out);
        }
    } quiesce{ctx};
    StreamDesc sd;
    memset(&sd, 0, sizeof(sd));
    sd.begin = 0;
    sd.end = len;
    sd.audio_begin = audio_offset;
    sd.total = si.total;
    sd.si_sample_rate = si.sample_rate;
    sd.si_channels = (uint8_t)C;
    sd.si_depth = (uint8_t)depth;
    EventList ev_in, ev_hint, ev_mid, ev_done, ev_out;
    Events ev;
    if (ev_in.make(K) || ev_hint.make(K) || ev_mid.make(K) || ev_done.make(K) || ev_out.make(K + 1) || ev.init()) return ZFLACB200_CUDA_ERROR;
    // the MD5 gate of a host-output decode runs on a host thread over the PCM pieces as they land in `dst` (hostmd5.h)
    std::unique_ptr<zf::HashWorker> hasher;
    if (do_md5 && !on_device && md5_on_host(flags, pcm_bytes))
        hasher.reset(new zf::HashWorker(device, cb, aligned == 24 ? 3 : cb, msb_shift));
    uint32_t n_out = 0;
    CU_CHECK(put_small(ctx, b_sd.p, &sd, sizeof(sd), st));
    CU_CHECK(cudaMemsetAsync(b_sum.p, 0, sizeof(StreamSummary), st));
    CU_CHECK(cudaMemsetAsync(b_ok.p, 2, cap, st));
    CU_CHECK(cudaMemsetAsync(b_flag.p, 0, sizeof(uint32_t), st));
    CU_CHECK(cudaMemsetAsync(b_hi.p, 0, sizeof(unsigned long long) * K, st));
    CU_CHECK(cudaEventRecord(ev.e[0], st));
    CU_CHECK(cudaEventRecord(ctx.ev_fork, st));  // the buffers above are stream-ordered allocations of `st`
    CU_CHECK(cudaStreamWaitEvent(ctx.copy_in, ctx.ev_fork, 0));
    CU_CHECK(cudaStreamWaitEvent(ctx.copy_out, ctx.ev_fork, 0));
    for (auto &l : ctx.lane) CU_CHECK(cudaStreamWaitEvent(l, ctx.ev_fork, 0));
    for (uint32_t i = 0; i < K && !on_device; ++i) {  // the whole upload is queued at once: the wire never waits for the host
        const uint64_t lo = i == 0 ? 0 : b[i];
        CU_CHECK(cudaMemcpyAsync((uint8_t *)b_stream.p + lo, data + lo, b[i + 1] - lo, cudaMemcpyHostToDevice, ctx.copy_in));
        CU_CHECK(cudaEventRecord(ev_in.v[i], ctx.copy_in));
    }
    tr.mark("setup");

    const uint8_t *d_stream = on_device ? data : (const uint8_t *)b_stream.p;
    Cand *d_cands = (Cand *)b_cands.p;
    FrameRec *d_recs = (FrameRec *)b_recs.p;
    FrameOut *d_outs = (FrameOut *)b_outs.p;
    StreamSummary *d_sum = (StreamSummary *)b_sum.p;
    unsigned long long *d_hi = (unsigned long long *)b_hi.p;
    const uint32_t window = 16;
    uint32_t n = 0;          // candidates so far
    uint64_t copied = 0;     // PCM elements already queued for the trip back
    bool bail = false;
    // PCM [copied, upto) leaves once chunk_i is decoded (it holds frames of chunk_i only: earlier chunks end below `copied`)
    auto copy_out = [&](uint32_t chunk_i, uint64_t upto) -> int {
        if (on_device) return ZFLACB200_OK;  // the PCM stays in HBM
        if (upto > n_elems) upto = n_elems;
        if (upto <= copied) return ZFLACB200_OK;
        CU_CHECK(cudaStreamWaitEvent(ctx.copy_out, ev_done.v[chunk_i], 0));
        CU_CHECK(cudaMemcpyAsync((uint8_t *)dst + copied * cb, (const uint8_t *)b_pcm.p + copied * cb, (upto - copied) * cb,
                                 cudaMemcpyDeviceToHost, ctx.copy_out));
        if (hasher && n_out < ev_out.v.size()) {
            CU_CHECK(cudaEventRecord(ev_out.v[n_out], ctx.copy_out));
            hasher->push(ev_out.v[n_out], (const uint8_t *)dst + copied * cb, upto - copied);
            ++n_out;
        }
        copied = upto;
        return ZFLACB200_OK;
    };
    uint32_t launched = 0;  // chunks whose kernels are queued
    for (uint32_t i = 0; i < K && !bail; ++i) {
        // scan + hinted records on `st` (short, the host waits for the count); the frame-serial kernels of the
        // chunk on one of S lanes: they are latency-bound (one lane per frame), so several chunks run side by side
        cudaStream_t cs = ctx.lane[i % S];
        if (!on_device) CU_CHECK(cudaStreamWaitEvent(st, ev_in.v[std::min(i + 1, K - 1)], 0));
        const uint64_t readable = i + 2 >= K ? len : b[i + 2];
        CU_CHECK(launch_sync_scan_range(d_stream, b[i], b[i + 1], readable, d_cands + n, cap - n, (uint32_t *)b_total.p,
                                        (unsigned long long *)b_tiles.p, (uint32_t *)b_ticket.p, st));
        CU_CHECK(launch_copy_small(&ctx.mailbox[0], b_total.p, sizeof(uint32_t), st));  // not a memcpy: see xfer.cu
        CU_CHECK(cudaStreamSynchronize(st));  // the host needs the count to place the next chunk's records
        const uint32_t m = (uint32_t)(ctx.mailbox[0] & 0xFFFFFFFFull);
        if (m > cap - n) {  // denser in valid headers than STREAMINFO suggests
            bail = true;
            break;
        }
        if (m) {
            CU_CHECK(launch_hint_chain_range(d_cands, n, n + m, (const StreamDesc *)b_sd.p, 1, d_recs, window, d_hi + i, st));
            CU_CHECK(cudaMemsetAsync(d_outs + n, 0xFF, sizeof(FrameOut) * (size_t)m, st));  // dropped slots: status != OK
        }
        CU_CHECK(launch_copy_small(&ctx.mailbox[2 + i], d_hi + i, sizeof(unsigned long long), st));
        CU_CHECK(cudaEventRecord(ev_hint.v[i], st));
        CU_CHECK(cudaStreamWaitEvent(cs, ev_hint.v[i], 0));
        if (i == 0) CU_CHECK(cudaEventRecord(ev.e[4], cs));  // the first K2a starts
        if (m) {
            if (stage.decode(d_stream, d_recs + n, m, b_scratch.p, wide, d_outs + n, true, cs, n) != ZFLACB200_OK) return ZFLACB200_CUDA_ERROR;
            CU_CHECK(cudaEventRecord(ev_mid.v[i], cs));
            EpilogueParams ep;
            ep.frames = d_recs + n;
            ep.n_frames = m;
            ep.max_block_size = si.max_block >= 16 ? si.max_block : 4096;
            ep.scratch = b_scratch.p;
            ep.wide_scratch = wide;
            ep.pcm = b_pcm.p;
            ep.container_bytes = cb;
            ep.msb_shift = msb_shift;
            ep.lossy_flag = (uint32_t *)b_flag.p;
            if (stage.epilogue(ep, d_outs + n, C, cs, n) != ZFLACB200_OK) return ZFLACB200_CUDA_ERROR;
        }
        CU_CHECK(cudaEventRecord(ev_done.v[i], cs));
        launched = i + 1;
        n += m;
        // chunk i-1's PCM range is known since the stream synchronisation above (its records were built before this scan)
        if (i > 0 && copy_out(i - 1, ctx.mailbox[2 + i - 1]) != ZFLACB200_OK) return ZFLACB200_CUDA_ERROR;
    }
    tr.mark("chunks");
    CU_CHECK(cudaEventRecord(ev.e[3], st));  // every scan and hint pass is done
    // (an event that was never recorded counts as complete: chunks without candidates)
    for (uint32_t i = launched > (uint32_t)S ? launched - S : 0; i < launched; ++i) CU_CHECK(cudaStreamWaitEvent(st, ev_mid.v[i], 0));
    CU_CHECK(cudaEventRecord(ev.e[5], st));  // every K2a is done
    for (uint32_t i = launched > (uint32_t)S ? launched - S : 0; i < launched; ++i) CU_CHECK(cudaStreamWaitEvent(st, ev_done.v[i], 0));
    CU_CHECK(cudaEventRecord(ev.e[6], st));  // every K3p is done
    uint32_t lossy = 0;
    StreamSummary sum;
    memset(&sum, 0, sizeof(sum));
    if (!bail && n > 0) {
        if ((flags & ZFLACB200_SKIP_FRAME_CRC16) == 0)
            CU_CHECK(launch_crc16_frames(d_stream, d_recs, d_outs, n, d_sum, (uint8_t *)b_ok.p, st));
        CU_CHECK(launch_verify_chain(d_cands, n, (const StreamDesc *)b_sd.p, 1, d_recs, d_outs, d_sum, window, st));
        GET_SMALL(h_sum, ctx, b_sum.p, sizeof(sum), st);
        GET_SMALL(h_lossy, ctx, b_flag.p, sizeof(lossy), st);
        GET_SMALL(h_ovf, ctx, stage.two_pass ? (const void *)stage.overflow_flag() : b_flag.p, sizeof(uint32_t), st);
        CU_CHECK(cudaEventRecord(ev.e[7], st));
        if (copy_out(K - 1, n_elems) != ZFLACB200_OK) return ZFLACB200_CUDA_ERROR;
        CU_CHECK(cudaStreamSynchronize(st));
        memcpy(&sum, h_sum, sizeof(sum));
        lossy = *(const uint32_t *)h_lossy;
        if (stage.narrow && *(const uint32_t *)h_ovf) bail = true;  // a residual beyond 16 bits: the plain path redoes it on the int32 scratch
    } else {
        CU_CHECK(cudaStreamSynchronize(st));
    }
    const bool accepted = !bail && n > 0 && !sum.fail && sum.first_ok && sum.n_samples == n_elems && sum.channels == C &&
                          !(msb_shift && lossy);
    memset(info, 0, sizeof(*info));
    memcpy(info->md5_streaminfo, si.md5, 16);
    info->audio_offset = audio_offset;
    info->container_bytes = (uint8_t)cb;
    int result = ZFLACB200_OK;
    if (accepted && do_md5 && hasher) {
        const auto t0 = std::chrono::steady_clock::now();
        const bool ok = hasher->finish(info->md5_computed) && copied == n_elems;
        info->ms_md5 = std::chrono::duration<float, std::milli>(std::chrono::steady_clock::now() - t0).count();  // the tail the copies did not hide
        if (!ok) {
            g_last_error = "host MD5 worker failed";
            return ZFLACB200_CUDA_ERROR;
        }
        info->md5_checked = 1;
        if (memcmp(info->md5_computed, si.md5, 16) != 0) result = ZFLACB200_INVALID_CHECKSUM;  // :279-280
    } else if (accepted && do_md5 && md5_on_host(flags, pcm_bytes)) {  // device form: PCM stays in HBM, staged through pinned buffers
        const int r = host_md5_of_device_pcm(ctx, b_pcm.p, n_elems, cb, aligned == 24 ? 3 : cb, msb_shift, st, info->md5_computed, &info->ms_md5);
        if (r != ZFLACB200_OK) return r;
        info->md5_checked = 1;
        if (memcmp(info->md5_computed, si.md5, 16) != 0) result = ZFLACB200_INVALID_CHECKSUM;  // :279-280
    } else if (accepted && do_md5) {  // K4 over the device copy while the last PCM chunk is still on its way out
        Md5Job job;
        job.pcm = b_pcm.p;
        job.n_samples = n_elems;
        job.container_bytes = cb;
        job.hash_bytes = aligned == 24 ? 3 : cb;  // :268-277
        job.unshift = msb_shift;
        job.pad = 0;
        CU_CHECK(put_small(ctx, b_jobs.p, &job, sizeof(job), st));
        CU_CHECK(launch_md5((const Md5Job *)b_jobs.p, 1, (uint8_t *)b_digest.p, st));
        GET_SMALL(h_digest, ctx, b_digest.p, 16, st);
        CU_CHECK(cudaStreamSynchronize(st));
        memcpy(info->md5_computed, h_digest, 16);
        info->md5_checked = 1;
        if (memcmp(info->md5_computed, si.md5, 16) != 0) result = ZFLACB200_INVALID_CHECKSUM;  // :279-280
    }
    CU_CHECK(cudaStreamSynchronize(ctx.copy_out));
    CU_CHECK(cudaStreamSynchronize(ctx.copy_in));
    quiesce.armed = false;  // every lane was joined into `st` above and the copies are through
    tr.mark("tail");
    if (!accepted) return kNotPipelined;
    info->ms_total = ev_ms(ev.e[0], ev.e[7]);
    if (launched) {  // the chunks overlap: these are SPANS (first start to last end), they do not add up to ms_total
        info->ms_scan = ev_ms(ev.e[0], ev.e[3]);      // K1 + K1h of all chunks, host round trips included
        info->ms_decode = ev_ms(ev.e[4], ev.e[5]);    // first K2a start .. last K2a end (K3p of early chunks runs inside)
        info->ms_epilogue = ev_ms(ev.e[5], ev.e[6]);  // what is left of K3p after the last K2a
    }
    if (result != ZFLACB200_OK) return result;
    info->channels = sum.channels;
    info->bits_per_sample = sum.bps;
    info->sample_rate = sum.sample_rate;
    info->n_samples = sum.n_samples;
    info->pcm_bytes = pcm_bytes;
    info->n_frames = sum.n_frames;
    info->n_candidates = n;
    info->fast_path = 1;
    info->n_crc16_mismatch = sum.n_crc16_mismatch;
    info->audio_bytes = sum.end_offset - audio_offset;
    if (handle_out) {  // the PCM and the frame table stay on the device behind a handle, as zflacb200_open() leaves them
        zflacb200_stream *h = new zflacb200_stream();
        h->device = device;
        h->pcm_bytes = info->pcm_bytes;
        h->d_pcm = b_pcm.release();
        h->d_recs = (FrameRec *)b_recs.release();
        h->d_outs = (FrameOut *)b_outs.release();
        h->d_frame_ok = (uint8_t *)b_ok.release();
        h->n_recs = n;
        h->frames_on_device = true;
        h->info = *info;
        *handle_out = h;
    }
    return ZFLACB200_OK;
}

// ---- frame-range decode of one stream: the shard's frames through the same K1 / K1h / K2a / K3p / K1v
// kernels as the whole-stream fast path. What changes is bookkeeping only: slot 0 of the candidate list is the stream's
// first frame (scanned from a 4 KiB head upload) so every shard judges its candidates against the same reference
// header; the scan starts one maximum frame below the shard so the first frame of the shard finds its predecessor;
// PCM and scratch positions are made relative to the shard's first sample by a first pass of the hint kernel.
int decode_range_impl(const uint8_t *data, size_t len, const zflacb200_options *opt, uint64_t byte_lo, uint64_t byte_hi, void *dst,
                      size_t dst_capacity, zflacb200_range *out, zflacb200_info *info) {
    if (!data || !out || !info || out->struct_size < sizeof(zflacb200_range) || (!dst && dst_capacity)) {
        g_last_error = "null argument or zflacb200_range.struct_size not set";
        return ZFLACB200_INVALID_ARGUMENT;
    }
    memset(info, 0, sizeof(*info));
    memset((uint8_t *)out + 4, 0, sizeof(*out) - 4);
    const uint32_t flags = (opt ? opt->flags : 0) | env_flags();
    if (flags & ZFLACB200_INPUT_ON_DEVICE) {
        g_last_error = "zflacb200_decode_range takes host input";
        return ZFLACB200_INVALID_ARGUMENT;
    }
    int device = -1;
    {
        const int e = ensure_device(opt ? opt->device : -1, &device);
        if (e != ZFLACB200_OK) return e;
    }
    DeviceCtx &ctx = g_ctx[device];
    cudaStream_t st = (opt && opt->stream) ? (cudaStream_t)opt->stream : ctx.stream;
    HostTrace tr("decode_range");

    ByteSource src;
    src.host = data;
    src.len = len;
    Streaminfo si;
    uint64_t audio_offset = 0;
    {
        const int e = parse_metadata(src, si, audio_offset);
        if (e != ZFLACB200_OK) return e;  // every shard reports what decode() reports (:220, :248, :309)
    }
    const uint32_t C = si.channels_m1 + 1;
    const uint32_t depth = si.depth_m1 + 1, aligned = (depth + 7u) & ~7u;
    const uint32_t cb = aligned == 8 ? 1 : aligned == 16 ? 2 : 4;
    uint32_t msb_shift = 0;
    if (depth >= 9 && depth <= 15) msb_shift = 16 - depth;
    else if (depth >= 17 && depth <= 31) msb_shift = 32 - depth;
    info->audio_offset = audio_offset;
    info->container_bytes = (uint8_t)cb;
    memcpy(info->md5_streaminfo, si.md5, 16);
    out->total_samples = si.total;
    Hdr h0;
    if (si.total == 0 || si.total * C >= (1ull << 40) || host_parse_header(src, audio_offset, h0) != ZFLACB200_OK)
        return ZFLACB200_OK;  // accepted = 0: the whole-stream path decides what this stream is
    const uint32_t bps0 = h0.bd_code == 0 ? depth : kBps[h0.bd_code];
    const bool wide = bps0 == 32;
    info->channels = (uint8_t)C;
    info->bits_per_sample = (uint8_t)bps0;                                                 // :381-384
    info->sample_rate = (h0.sr_code == 0 ? si.sample_rate : h0.sr_value) & 0xFFFFFFu;     // :376-379

    // ---- what to scan and what to upload
    const uint64_t lo = std::min<uint64_t>(std::max<uint64_t>(byte_lo, audio_offset), len);
    const uint64_t hi = std::min<uint64_t>(std::max<uint64_t>(byte_hi, lo), len);
    const uint32_t max_bs = si.max_block >= 16 ? si.max_block : 65535;
    uint64_t max_frame = si.max_frame ? (uint64_t)si.max_frame + 16 : 32 + (uint64_t)max_bs * C * ((depth + 1 + 7) / 8 + 1);
    max_frame = std::min<uint64_t>(std::max<uint64_t>(max_frame, 64u << 10), 4u << 20);
    const uint64_t scan_lo = std::max<uint64_t>(audio_offset, lo > max_frame ? lo - max_frame : 0);
    const uint64_t up_lo = scan_lo & ~4095ull;
    const uint64_t up_hi = std::min<uint64_t>(len, hi + max_frame + 4096);
    const uint64_t head_lo = audio_offset & ~4095ull;
    const uint64_t head_hi = std::min<uint64_t>(len, audio_offset + 64);
    if (hi <= lo) {  // nothing starts here
        out->accepted = 1;
        out->first_frame_offset = out->end_offset = lo;
        return ZFLACB200_OK;
    }
    const uint64_t cap64 = (hi - scan_lo) / 256 + 4096;
    if (cap64 > 0x3FFFFFFFull) return ZFLACB200_OK;
    const uint32_t cap = (uint32_t)cap64;
    const uint32_t n_tiles = scan_tile_count(scan_lo, hi) + 2;

    DevBuf b_stream, b_head, b_cands, b_total, b_tiles, b_ticket, b_sd, b_recs, b_outs, b_ok, b_sum, b_flag, b_lohi, b_scratch, b_pcm;
    SubframeStage stage;
    if (b_stream.alloc(((up_hi - up_lo + 15) & ~15ull) + 64, st) || b_head.alloc(4096 + 64 + 64, st) ||
        b_cands.alloc(sizeof(Cand) * (size_t)cap, st) || b_total.alloc(sizeof(uint32_t), st) ||
        b_tiles.alloc(sizeof(unsigned long long) * n_tiles, st) || b_ticket.alloc(sizeof(uint32_t), st) ||
        b_sd.alloc(sizeof(StreamDesc), st) || b_recs.alloc(sizeof(FrameRec) * (size_t)cap, st) ||
        b_outs.alloc(sizeof(FrameOut) * (size_t)cap, st) || b_ok.alloc(cap, st) || b_sum.alloc(sizeof(StreamSummary), st) ||
        b_flag.alloc(sizeof(uint32_t), st) || b_lohi.alloc(4 * sizeof(unsigned long long), st))
        return ZFLACB200_CUDA_ERROR;
    Events ev;
    if (ev.init() != ZFLACB200_OK) return ZFLACB200_CUDA_ERROR;
    CU_CHECK(cudaEventRecord(ev.e[0], st));
    CU_CHECK(cudaMemcpyAsync(b_head.p, data + head_lo, head_hi - head_lo, cudaMemcpyHostToDevice, st));
    CU_CHECK(cudaMemcpyAsync(b_stream.p, data + up_lo, up_hi - up_lo, cudaMemcpyHostToDevice, st));
    CU_CHECK(cudaEventRecord(ev.e[1], st));  // upload done
    const uint8_t *d_head = (const uint8_t *)b_head.p - head_lo;    // absolute stream offsets index both buffers
    const uint8_t *d_stream = (const uint8_t *)b_stream.p - up_lo;
    Cand *d_cands = (Cand *)b_cands.p;
    FrameRec *d_recs = (FrameRec *)b_recs.p;
    FrameOut *d_outs = (FrameOut *)b_outs.p;
    StreamSummary *d_sum = (StreamSummary *)b_sum.p;
    unsigned long long *d_lohi = (unsigned long long *)b_lohi.p;

    // ---- K1 twice: the stream's first header into slot 0, the shard's candidates behind it
    CU_CHECK(launch_sync_scan_range(d_head, audio_offset, audio_offset + 1, head_hi, d_cands, 1, (uint32_t *)b_total.p,
                                    (unsigned long long *)b_tiles.p, (uint32_t *)b_ticket.p, st));
    GET_SMALL(h_m0, ctx, b_total.p, sizeof(uint32_t), st);
    CU_CHECK(launch_sync_scan_range(d_stream, scan_lo, hi, up_hi, d_cands + 1, cap - 1, (uint32_t *)b_total.p,
                                    (unsigned long long *)b_tiles.p, (uint32_t *)b_ticket.p, st));
    GET_SMALL(h_m, ctx, b_total.p, sizeof(uint32_t), st);
    CU_CHECK(cudaEventRecord(ev.e[2], st));
    CU_CHECK(cudaStreamSynchronize(st));
    const uint32_t m0 = *(const uint32_t *)h_m0, m = *(const uint32_t *)h_m;
    tr.mark("upload+scan");
    if (m0 != 1 || m > cap - 1) return ZFLACB200_OK;  // accepted = 0
    const uint32_t n = 1 + m;
    info->n_candidates = m;

    // ---- K1h, first pass: which records stay, and where the shard starts (samples, bytes)
    StreamDesc sd;
    memset(&sd, 0, sizeof(sd));
    sd.begin = 0;
    sd.end = up_hi;  // nothing above the uploaded bytes is readable: a frame that runs past them fails the shard
    sd.audio_begin = audio_offset;
    sd.total = si.total;
    sd.si_sample_rate = si.sample_rate;
    sd.si_channels = (uint8_t)C;
    sd.si_depth = (uint8_t)depth;
    const uint32_t window = 16;
    const unsigned long long lohi0[4] = {~0ull, ~0ull, 0ull, 0ull};  // min sample_base, min offset, kept, max sample end
    CU_CHECK(put_small(ctx, b_sd.p, &sd, sizeof(sd), st));
    CU_CHECK(put_small(ctx, d_lohi, lohi0, sizeof(lohi0), st));
    CU_CHECK(cudaMemsetAsync(d_recs, 0, sizeof(FrameRec), st));  // slot 0 is never a record
    RangeHint rh;
    rh.decode_lo = lo;
    rh.window_floor = 1;
    rh.lo_out = d_lohi;
    CU_CHECK(launch_hint_chain_shard(d_cands, 1, n, (const StreamDesc *)b_sd.p, 1, d_recs, window, d_lohi + 3, rh, st));
    GET_SMALL(h_lohi, ctx, d_lohi, sizeof(lohi0), st);
    CU_CHECK(cudaStreamSynchronize(st));
    unsigned long long lohi[4];
    memcpy(lohi, h_lohi, sizeof(lohi));
    if (lohi[2] == 0) {  // no frame starts in [lo, hi): a shard thinner than one frame
        out->accepted = 1;
        out->first_frame_offset = out->end_offset = lo;
        return ZFLACB200_OK;
    }
    if (lohi[0] % C != 0 || lohi[3] <= lohi[0]) return ZFLACB200_OK;
    const uint64_t first_sample = lohi[0] / C, n_elems = lohi[3] - lohi[0];
    out->first_sample = first_sample;
    out->n_samples = n_elems / C;
    out->first_frame_offset = lohi[1];
    info->n_samples = n_elems;
    info->pcm_bytes = n_elems * cb;
    const uint64_t dst_off = (flags & ZFLACB200_RANGE_INTO_WHOLE) ? lohi[0] * cb : 0;  // the shard's place in the whole stream's PCM
    if (dst_capacity < dst_off + n_elems * cb) {
        g_last_error = "destination too small for the shard's PCM";
        return ZFLACB200_OUT_OF_MEMORY;
    }

    // ---- second pass with positions relative to the shard (unsigned wrap-around: base + absolute = relative)
    sd.pcm_base = 0ull - lohi[0];
    sd.scratch_base = 0ull - (uint64_t)C * ((first_sample + 7) & ~7ull);
    const uint64_t scratch_elems = (uint64_t)C * (((out->n_samples + 7) & ~7ull) + 16) + 64ull * n + 64;
    if (b_scratch.alloc(scratch_elems * (wide ? 8 : 4), st) || b_pcm.alloc(((n_elems + 15) & ~15ull) * cb + 64, st) ||
        stage.prepare(wide, flags, n, st, sd.si_depth) != ZFLACB200_OK)
        return ZFLACB200_CUDA_ERROR;
    CU_CHECK(put_small(ctx, b_sd.p, &sd, sizeof(sd), st));
    rh.lo_out = nullptr;
    CU_CHECK(launch_hint_chain_shard(d_cands, 1, n, (const StreamDesc *)b_sd.p, 1, d_recs, window, nullptr, rh, st));
    CU_CHECK(cudaMemsetAsync(b_sum.p, 0, sizeof(StreamSummary), st));
    CU_CHECK(cudaMemsetAsync(b_ok.p, 2, n, st));
    CU_CHECK(cudaMemsetAsync(b_flag.p, 0, sizeof(uint32_t), st));
    CU_CHECK(cudaMemsetAsync(d_outs, 0xFF, sizeof(FrameOut) * (size_t)n, st));
    CU_CHECK(cudaEventRecord(ev.e[3], st));
    if (stage.decode(d_stream, d_recs, n, b_scratch.p, wide, d_outs, true, st) != ZFLACB200_OK) return ZFLACB200_CUDA_ERROR;
    CU_CHECK(cudaEventRecord(ev.e[4], st));
    if ((flags & ZFLACB200_SKIP_FRAME_CRC16) == 0)
        CU_CHECK(launch_crc16_frames(d_stream, d_recs, d_outs, n, d_sum, (uint8_t *)b_ok.p, st));
    CU_CHECK(launch_verify_chain(d_cands, n, (const StreamDesc *)b_sd.p, 1, d_recs, d_outs, d_sum, window, st, true));
    EpilogueParams ep;
    ep.frames = d_recs;
    ep.n_frames = n;
    ep.max_block_size = max_bs;
    ep.scratch = b_scratch.p;
    ep.wide_scratch = wide;
    ep.pcm = b_pcm.p;
    ep.container_bytes = cb;
    ep.msb_shift = msb_shift;
    ep.lossy_flag = (uint32_t *)b_flag.p;
    if (stage.epilogue(ep, d_outs, C, st) != ZFLACB200_OK) return ZFLACB200_CUDA_ERROR;
    CU_CHECK(cudaEventRecord(ev.e[5], st));
    GET_SMALL(h_sum, ctx, b_sum.p, sizeof(StreamSummary), st);
    GET_SMALL(h_lossy, ctx, b_flag.p, sizeof(uint32_t), st);
    GET_SMALL(h_ovf, ctx, stage.two_pass ? (const void *)stage.overflow_flag() : b_flag.p, sizeof(uint32_t), st);
    CU_CHECK(cudaMemcpyAsync((uint8_t *)dst + dst_off, b_pcm.p, n_elems * cb, cudaMemcpyDeviceToHost, st));  // an unaccepted shard's bytes are garbage
    CU_CHECK(cudaEventRecord(ev.e[7], st));
    CU_CHECK(cudaStreamSynchronize(st));
    tr.mark("decode+readback");
    StreamSummary sum;
    memcpy(&sum, h_sum, sizeof(sum));
    const uint32_t lossy = *(const uint32_t *)h_lossy;
    info->ms_scan = ev_ms(ev.e[1], ev.e[2]);      // the kernels only: upload (e0..e1) and read-back (e5..e7) are in ms_total
    info->ms_link = ev_ms(ev.e[2], ev.e[3]);      // both hint passes, with the host round trip between them
    info->ms_decode = ev_ms(ev.e[3], ev.e[4]);    // K2a
    info->ms_epilogue = ev_ms(ev.e[4], ev.e[5]);  // CRC-16 + chain check + K3p
    info->ms_total = ev_ms(ev.e[0], ev.e[7]);
    info->n_frames = sum.n_frames;
    info->n_crc16_mismatch = sum.n_crc16_mismatch;
    info->audio_bytes = sum.end_offset > lohi[1] ? sum.end_offset - lohi[1] : 0;
    out->n_frames = sum.n_frames;
    out->end_offset = sum.end_offset;
    const bool narrow_overflow = stage.narrow && *(const uint32_t *)h_ovf != 0;  // the whole-stream path redoes it on the int32 scratch
    const bool accepted = !sum.fail && sum.n_frames == lohi[2] && sum.n_samples == n_elems && !(msb_shift && lossy) &&
                          sum.end_offset > lohi[1] && !narrow_overflow;
    out->accepted = accepted ? 1u : 0u;
    info->fast_path = accepted ? 1 : 0;
    return ZFLACB200_OK;
}
}  // namespace

extern "C" {

int zflacb200_abi_version(void) { return ZFLACB200_ABI_VERSION; }

int zflacb200_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) return 0;
    return n;
}

int zflacb200_init(int device) {
    ApiScope api(device);
    int d;
    return ensure_device(api.device, &d);
}

void zflacb200_shutdown(void) {
    int prev = -1;
    if (cudaGetDevice(&prev) != cudaSuccess) prev = -1;
    for (int d = 0; d < 64; ++d) {
        std::lock_guard<std::mutex> lock(g_dev_mutex[d]);
        if (g_ctx[d].ready) {
            cudaSetDevice(d);
            cudaStreamDestroy(g_ctx[d].stream);
            cudaStreamDestroy(g_ctx[d].side);
            cudaStreamDestroy(g_ctx[d].copy_in);
            cudaStreamDestroy(g_ctx[d].copy_out);
            for (auto &l : g_ctx[d].lane) cudaStreamDestroy(l);
            cudaFreeHost(g_ctx[d].mailbox);
            for (int k = 0; k < 2; ++k) {
                if (g_ctx[d].md5_stage[k]) cudaFreeHost(g_ctx[d].md5_stage[k]);
                if (g_ctx[d].ev_md5[k]) cudaEventDestroy(g_ctx[d].ev_md5[k]);
            }
            for (auto &pb : g_ctx[d].pin_blocks) cudaFreeHost(pb.first);
            cudaEventDestroy(g_ctx[d].ev_fork);
            cudaEventDestroy(g_ctx[d].ev_join);
            g_ctx[d] = DeviceCtx();
        }
    }
    g_cache.release_all();  // after the streams are gone: every block is idle
    if (prev >= 0) cudaSetDevice(prev);
}

void zflacb200_trim(void) { g_cache.release_all(); }

const char *zflacb200_status_name(int status) {
    switch (status) {
        case ZFLACB200_OK: return "OK";
        case ZFLACB200_INVALID_SIGNATURE: return "InvalidSignature";
        case ZFLACB200_INVALID_METADATA_HEADER: return "InvalidMetadataHeader";
        case ZFLACB200_UNIMPLEMENTED: return "Unimplemented";
        case ZFLACB200_INVALID_CHECKSUM: return "InvalidChecksum";
        case ZFLACB200_MISSING_STREAMINFO: return "MissingStreaminfo";
        case ZFLACB200_INVALID_FRAME_HEADER: return "InvalidFrameHeader";
        case ZFLACB200_INVALID_CODED_NUMBER: return "InvalidCodedNumber";
        case ZFLACB200_INCONSISTENT_PARAMETERS: return "InconsistentParameters";
        case ZFLACB200_INVALID_SUBFRAME_HEADER: return "InvalidSubframeHeader";
        case ZFLACB200_INVALID_RESIDUAL_CODING_METHOD: return "InvalidResidualCodingMethod";
        case ZFLACB200_END_OF_STREAM: return "EndOfStream";
        case ZFLACB200_OUT_OF_MEMORY: return "OutOfMemory";
        case ZFLACB200_UNDEFINED_IN_REFERENCE: return "UndefinedInReference";
        case ZFLACB200_CUDA_ERROR: return "CudaError";
        case ZFLACB200_INVALID_ARGUMENT: return "InvalidArgument";
        default: return "Unknown";
    }
}

const char *zflacb200_last_error(void) { return g_last_error.c_str(); }

uint64_t zflacb200_kernel_launches(void) { return g_kernel_launches.load(); }

uint64_t zflacb200_pass_a_launches(int variant) {
    return variant >= 0 && variant < kPassAVariants ? g_pass_a_launches[variant].load() : 0;
}

int zflacb200_open(const uint8_t *data, size_t len, const zflacb200_options *opt, zflacb200_stream **out, zflacb200_info *info) {
    ApiScope api(opt_device(opt));
    if (out && info && (((opt ? opt->flags : 0) | env_flags()) & ZFLACB200_INPUT_ON_DEVICE)) {  // long device-resident stream: chunked
        *out = nullptr;
        const int p = decode_pipelined(data, len, opt, nullptr, 0, info, out);
        if (p != kNotPipelined) return api.done(p);
        if (*out) { zflacb200_stream *h = *out; *out = nullptr; dev_free(h->d_pcm, g_ctx[api.device].stream); delete h; }
    }
    return api.done(open_impl(data, len, opt, out, info));
}

int zflacb200_open_frames(const uint8_t *data, size_t len, const zflacb200_options *opt, uint64_t byte_offset,
                          zflacb200_stream **out, zflacb200_info *info) {
    ApiScope api(opt_device(opt));
    if (byte_offset == kNotRaw || byte_offset > len) {
        g_last_error = "byte_offset past the end of the data";
        return ZFLACB200_INVALID_ARGUMENT;
    }
    return api.done(open_impl(data, len, opt, out, info, byte_offset));
}

int zflacb200_read(zflacb200_stream *s, void *dst, size_t dst_bytes) {
    if (!s || (!dst && s->pcm_bytes)) return ZFLACB200_INVALID_ARGUMENT;
    if (dst_bytes < s->pcm_bytes) {
        g_last_error = "destination smaller than pcm_bytes";
        return ZFLACB200_INVALID_ARGUMENT;
    }
    ApiScope api(s->device);
    CU_CHECK(cudaSetDevice(s->device));
    if (s->pcm_bytes) CU_CHECK(cudaMemcpy(dst, s->d_pcm, s->pcm_bytes, cudaMemcpyDeviceToHost));
    return ZFLACB200_OK;
}

static int format_bytes(const zflacb200_stream *s, int format, uint64_t *bytes) {
    const uint64_t n = s->info.n_samples;
    const uint32_t cb = s->info.container_bytes;
    switch (format) {
        case ZFLACB200_FMT_NATIVE: *bytes = s->pcm_bytes; return ZFLACB200_OK;
        case ZFLACB200_FMT_F32: *bytes = n * 4; return ZFLACB200_OK;
        case ZFLACB200_FMT_S24_PACKED:
            if (cb != 4) break;
            *bytes = n * 3;
            return ZFLACB200_OK;
        case ZFLACB200_FMT_U8:
            if (cb != 1) break;
            *bytes = n;
            return ZFLACB200_OK;
        default: break;
    }
    g_last_error = "format not defined for this stream's container";
    return ZFLACB200_INVALID_ARGUMENT;
}

int zflacb200_format_bytes(zflacb200_stream *s, int format, uint64_t *bytes) {
    if (!s || !bytes) return ZFLACB200_INVALID_ARGUMENT;
    return format_bytes(s, format, bytes);
}

int zflacb200_read_as(zflacb200_stream *s, int format, void *dst, size_t dst_bytes) {
    if (!s) return ZFLACB200_INVALID_ARGUMENT;
    if (format == ZFLACB200_FMT_NATIVE) return zflacb200_read(s, dst, dst_bytes);
    uint64_t need = 0;
    {
        const int e = format_bytes(s, format, &need);
        if (e != ZFLACB200_OK) return e;
    }
    if ((!dst && need) || dst_bytes < need) {
        g_last_error = "destination smaller than zflacb200_format_bytes()";
        return ZFLACB200_INVALID_ARGUMENT;
    }
    if (need == 0) return ZFLACB200_OK;
    ApiScope api(s->device);
    CU_CHECK(cudaSetDevice(s->device));
    cudaStream_t st = g_ctx[s->device].stream;
    DevBuf b_out;
    if (b_out.alloc(need + 64, st) != ZFLACB200_OK) return ZFLACB200_CUDA_ERROR;
    CU_CHECK(launch_convert_pcm(s->d_pcm, s->info.n_samples, s->info.container_bytes, format, b_out.p, st));
    CU_CHECK(cudaMemcpyAsync(dst, b_out.p, need, cudaMemcpyDeviceToHost, st));
    CU_CHECK(cudaStreamSynchronize(st));
    return ZFLACB200_OK;
}

int zflacb200_device_pcm(zflacb200_stream *s, void **dev_ptr, size_t *bytes) {
    if (!s || !dev_ptr || !bytes) return ZFLACB200_INVALID_ARGUMENT;
    *dev_ptr = s->d_pcm;
    *bytes = s->pcm_bytes;
    return ZFLACB200_OK;
}

void zflacb200_close(zflacb200_stream *s) {
    if (!s) return;
    ApiScope api(s->device);
    cudaSetDevice(s->device);
    if (s->d_pcm) dev_free(s->d_pcm, g_ctx[s->device].stream);
    if (s->d_recs) dev_free(s->d_recs, g_ctx[s->device].stream);
    if (s->d_outs) dev_free(s->d_outs, g_ctx[s->device].stream);
    if (s->d_frame_ok) dev_free(s->d_frame_ok, g_ctx[s->device].stream);
    delete s;
}

int zflacb200_decode(const uint8_t *data, size_t len, const zflacb200_options *opt, void *dst, size_t dst_capacity, zflacb200_info *info) {
    {
        ApiScope api(opt_device(opt));
        const int p = decode_pipelined(data, len, opt, dst, dst_capacity, info);
        if (p != kNotPipelined) return api.done(p);
    }
    zflacb200_stream *h = nullptr;
    const int e = zflacb200_open(data, len, opt, &h, info);
    if (e != ZFLACB200_OK) return e;
    int r = ZFLACB200_OK;
    if (dst_capacity < info->pcm_bytes) r = ZFLACB200_OUT_OF_MEMORY;
    else r = zflacb200_read(h, dst, dst_capacity);
    zflacb200_close(h);
    return r;
}

int zflacb200_pcm_bound(const uint8_t *data, size_t len, uint64_t *bytes) {
    if ((!data && len) || !bytes) return ZFLACB200_INVALID_ARGUMENT;
    *bytes = 0;
    ByteSource src;
    src.host = data;
    src.len = len;
    Streaminfo si;
    uint64_t audio_offset = 0;
    if (parse_metadata(src, si, audio_offset) != ZFLACB200_OK) return ZFLACB200_OK;  // the decode call reports it
    const uint32_t aligned = (si.depth_m1 + 1 + 7u) & ~7u;
    const uint64_t cb = aligned == 8 ? 1 : aligned == 16 ? 2 : 4;
    const uint64_t n = si.total * (si.channels_m1 + 1);
    if (si.total == 0 || !plausible_total(si.total, len - std::min<uint64_t>(len, audio_offset)) || n >= (1ull << 40)) return ZFLACB200_OK;
    *bytes = n * cb;
    return ZFLACB200_OK;
}

int zflacb200_decode_range(const uint8_t *data, size_t len, const zflacb200_options *opt, uint64_t byte_lo, uint64_t byte_hi,
                           void *dst, size_t dst_capacity, zflacb200_range *range, zflacb200_info *info) {
    ApiScope api(opt_device(opt));
    return api.done(decode_range_impl(data, len, opt, byte_lo, byte_hi, dst, dst_capacity, range, info));
}

// The reference's chain rule across shards: the first frame sits at the audio offset and holds sample 0 (:340), every
// shard starts at the byte and at the sample where the one before ended (:546-548, :580), the chain covers exactly the
// stated total (:341). Shards in which no frame starts are skipped.
int zflacb200_stitch(const zflacb200_range *r, size_t n, uint64_t audio_offset) {
    if (!r || n == 0) return 0;
    uint64_t next_offset = audio_offset, next_sample = 0;
    bool any = false;
    for (size_t i = 0; i < n; ++i) {
        if (!r[i].accepted || r[i].total_samples != r[0].total_samples) return 0;
        if (r[i].n_frames == 0) continue;
        if (r[i].first_frame_offset != next_offset || r[i].first_sample != next_sample) return 0;
        next_offset = r[i].end_offset;
        next_sample = r[i].first_sample + r[i].n_samples;
        any = true;
    }
    return any && next_sample == r[0].total_samples ? 1 : 0;
}

int zflacb200_batch_open(const uint8_t *blob, const uint8_t *device_blob, const uint64_t *offsets, const uint64_t *lens,
                         size_t n, const zflacb200_options *opt, zflacb200_batch **out, zflacb200_info *infos, int *statuses) {
    ApiScope api(opt_device(opt));
    return batch_open_impl(blob, device_blob, offsets, lens, n, opt, out, infos, statuses);
}

int zflacb200_batch_pcm_bound(const uint8_t *blob, const uint64_t *offsets, const uint64_t *lens, size_t n, uint64_t *bytes) {
    if (!blob || !offsets || !lens || !bytes) return ZFLACB200_INVALID_ARGUMENT;
    uint64_t total = 0;
    for (size_t i = 0; i < n; ++i) {
        ByteSource src;
        src.host = blob + offsets[i];
        src.len = lens[i];
        Streaminfo si;
        uint64_t audio_offset = 0;
        if (parse_metadata(src, si, audio_offset) != ZFLACB200_OK) continue;
        const uint32_t aligned = (si.depth_m1 + 1 + 7u) & ~7u;
        const uint32_t cb = aligned == 8 ? 1 : aligned == 16 ? 2 : 4;
        const uint64_t elems = (si.total * (si.channels_m1 + 1) + 15) & ~15ull;  // the device slots are 16-element aligned
        total += (elems * cb + 63) & ~63ull;
    }
    *bytes = total + 4096;
    return ZFLACB200_OK;
}

int zflacb200_batch_decode(const uint8_t *blob, const uint64_t *offsets, const uint64_t *lens, size_t n,
                           const zflacb200_options *opt, void *dst, size_t dst_bytes, uint64_t *pcm_offsets,
                           zflacb200_info *infos, int *statuses) {
    if (!blob || !offsets || !lens || !pcm_offsets || !infos || !statuses || (!dst && dst_bytes)) return ZFLACB200_INVALID_ARGUMENT;
    ApiScope api(opt_device(opt));
    HostTrace tr("batch_decode");
    int device = -1;
    {
        const int e = ensure_device(opt ? opt->device : -1, &device);
        if (e != ZFLACB200_OK) return e;
    }
    DeviceCtx &ctx = g_ctx[device];
    cudaStream_t st = (opt && opt->stream) ? (cudaStream_t)opt->stream : ctx.stream;
    uint64_t blob_len = 0;
    bool ordered = true;  // sub-batches are contiguous byte ranges: needs streams in blob order
    for (size_t i = 0; i < n; ++i) {
        blob_len = std::max(blob_len, offsets[i] + lens[i]);
        if (i && offsets[i] < offsets[i - 1] + lens[i - 1]) ordered = false;
        pcm_offsets[i] = ~0ull;
    }
    // ---- sub-batches: a sub-batch pays the fixed latency of the frame-serial kernels and of the per-file MD5 chain
    // once (C5: 3.5 ms for an eighth of the batch, 6.3 ms for all of it), so few and large ones: at most 8
    // (r07e, C5 end to end: 4 parts 46.5 ms, 6 45.0, 8 44.3, 12 48.3, 16 61.6; with the slower kernels of round 1 it was 4)
    static const uint64_t max_parts = [] {
        const char *e = getenv("ZFLACB200_BATCH_PARTS");
        const uint64_t v = e ? strtoull(e, nullptr, 10) : 8;
        return std::min<uint64_t>(std::max<uint64_t>(v, 1), 64);
    }();
    // ... and no part below ~96 MiB of stream: a rank's eighth of C5 (150 MB) in 8 parts was 8 x 3.5 ms of fixed latency (r08 N = 8)
    uint64_t part_min = 96ull << 20;
    if (const char *e = getenv("ZFLACB200_BATCH_PART_MIN_BYTES")) part_min = std::max<uint64_t>(strtoull(e, nullptr, 0), 4096);  // test knob, read per call
    size_t parts = ordered ? (size_t)std::min<uint64_t>(max_parts, std::max<uint64_t>(1, blob_len / part_min)) : 1;
    if (parts > n) parts = n ? n : 1;
    std::vector<size_t> first(parts + 1, n);
    first[0] = 0;
    for (size_t k = 1, i = 0; k < parts; ++k) {
        const uint64_t target = blob_len / parts * k;
        while (i < n && offsets[i] < target) ++i;
        first[k] = std::max(i, first[k - 1]);
    }
    DevBuf b_blob;
    if (b_blob.alloc(((blob_len + 15) & ~15ull) + 64, st) != ZFLACB200_OK) return ZFLACB200_CUDA_ERROR;
    EventList ev_in, ev_free;
    if (ev_in.make(parts) || ev_free.make(1)) return ZFLACB200_CUDA_ERROR;
    CU_CHECK(cudaEventRecord(ctx.ev_fork, st));
    CU_CHECK(cudaStreamWaitEvent(ctx.copy_in, ctx.ev_fork, 0));
    CU_CHECK(cudaStreamWaitEvent(ctx.copy_out, ctx.ev_fork, 0));
    for (size_t k = 0; k < parts; ++k) {  // the whole upload is queued at once
        const uint64_t lo = first[k] < n ? offsets[first[k]] : blob_len;
        const uint64_t hi = (k + 1 < parts && first[k + 1] < n) ? offsets[first[k + 1]] : blob_len;
        const uint64_t a = k == 0 ? 0 : lo;
        if (hi > a) CU_CHECK(cudaMemcpyAsync((uint8_t *)b_blob.p + a, blob + a, hi - a, cudaMemcpyHostToDevice, ctx.copy_in));
        CU_CHECK(cudaEventRecord(ev_in.v[k], ctx.copy_in));
    }
    zflacb200_options sub_opt;
    sub_opt.struct_size = sizeof(sub_opt);
    sub_opt.flags = (opt ? opt->flags : 0) & ~ZFLACB200_INPUT_ON_DEVICE;
    sub_opt.device = device;
    sub_opt.stream = (void *)st;
    std::vector<zflacb200_batch *> handles;
    struct CloseAll {
        std::vector<zflacb200_batch *> &h;
        int device;
        ~CloseAll() {
            for (auto *b : h) {
                cudaStream_t fs = g_ctx[device].stream;  // all copies were synchronised (or the call failed) before this runs
                for (auto &it : b->items)
                    if (it.solo) {
                        zflacb200_stream *s = it.solo;
                        if (s->d_pcm) dev_free(s->d_pcm, fs);
                        if (s->d_recs) dev_free(s->d_recs, fs);
                        if (s->d_outs) dev_free(s->d_outs, fs);
                        if (s->d_frame_ok) dev_free(s->d_frame_ok, fs);
                        delete s;
                    }
                for (auto &g : b->groups)
                    if (g.d_pcm) dev_free(g.d_pcm, fs);
                delete b;
            }
        }
    } close_all{handles, device};
    uint64_t pos = 0;  // write cursor in dst
    int result = ZFLACB200_OK;
    for (size_t k = 0; k < parts; ++k) {
        const size_t i0 = first[k], i1 = k + 1 < parts ? first[k + 1] : n;
        if (i1 <= i0) continue;
        CU_CHECK(cudaStreamWaitEvent(st, ev_in.v[k], 0));
        zflacb200_batch *b = nullptr;
        const int e = batch_open_impl(blob, (const uint8_t *)b_blob.p, offsets + i0, lens + i0, i1 - i0, &sub_opt, &b, infos + i0, statuses + i0);
        if (e != ZFLACB200_OK) {
            result = e;
            break;
        }
        handles.push_back(b);
        // ---- this sub-batch's PCM starts its trip back while the next one is decoded
        CU_CHECK(cudaEventRecord(ev_free.v[0], st));
        CU_CHECK(cudaStreamWaitEvent(ctx.copy_out, ev_free.v[0], 0));
        std::vector<uint64_t> group_pos(b->groups.size(), ~0ull);
        for (size_t gi = 0; gi < b->groups.size(); ++gi) {
            const uint64_t bytes = b->groups[gi].bytes;
            if (pos + bytes > dst_bytes) continue;  // no room: its streams report OutOfMemory below
            group_pos[gi] = pos;
            if (bytes) CU_CHECK(cudaMemcpyAsync((uint8_t *)dst + pos, b->groups[gi].d_pcm, bytes, cudaMemcpyDeviceToHost, ctx.copy_out));
            pos += (bytes + 63) & ~63ull;
        }
        for (size_t j = 0; j < b->items.size(); ++j) {
            const zflacb200_batch::Item &it = b->items[j];
            if (statuses[i0 + j] != ZFLACB200_OK) continue;
            if (it.group >= 0) {
                if (group_pos[it.group] == ~0ull) statuses[i0 + j] = ZFLACB200_OUT_OF_MEMORY;
                else pcm_offsets[i0 + j] = group_pos[it.group] + it.byte_off;
            } else if (it.solo) {
                if (pos + it.bytes > dst_bytes) {
                    statuses[i0 + j] = ZFLACB200_OUT_OF_MEMORY;
                    continue;
                }
                pcm_offsets[i0 + j] = pos;
                if (it.bytes) CU_CHECK(cudaMemcpyAsync((uint8_t *)dst + pos, it.solo->d_pcm, it.bytes, cudaMemcpyDeviceToHost, ctx.copy_out));
                pos += (it.bytes + 63) & ~63ull;
            }
        }
    }
    CU_CHECK(cudaStreamSynchronize(ctx.copy_out));
    CU_CHECK(cudaStreamSynchronize(ctx.copy_in));
    return result;
}

int zflacb200_batch_pcm_bytes(zflacb200_batch *b, uint64_t *bytes) {
    if (!b || !bytes) return ZFLACB200_INVALID_ARGUMENT;
    *bytes = b->read_all_bytes;
    return ZFLACB200_OK;
}

int zflacb200_batch_read_all(zflacb200_batch *b, void *dst, size_t dst_bytes, uint64_t *pcm_offsets) {
    if (!b || !pcm_offsets || (!dst && b->read_all_bytes)) return ZFLACB200_INVALID_ARGUMENT;
    if (dst_bytes < b->read_all_bytes) {
        g_last_error = "destination smaller than zflacb200_batch_pcm_bytes()";
        return ZFLACB200_INVALID_ARGUMENT;
    }
    ApiScope api(b->device);
    CU_CHECK(cudaSetDevice(b->device));
    cudaStream_t st = g_ctx[b->device].stream;
    uint64_t pos = 0;
    std::vector<uint64_t> group_pos(b->groups.size());
    for (size_t gi = 0; gi < b->groups.size(); ++gi) {  // one large copy per launch group
        group_pos[gi] = pos;
        if (b->groups[gi].bytes)
            CU_CHECK(cudaMemcpyAsync((uint8_t *)dst + pos, b->groups[gi].d_pcm, b->groups[gi].bytes, cudaMemcpyDeviceToHost, st));
        pos += (b->groups[gi].bytes + 63) & ~63ull;
    }
    for (size_t i = 0; i < b->items.size(); ++i) {
        const zflacb200_batch::Item &it = b->items[i];
        if (it.group >= 0) {
            pcm_offsets[i] = group_pos[it.group] + it.byte_off;
        } else if (it.solo) {
            pcm_offsets[i] = pos;
            if (it.bytes) CU_CHECK(cudaMemcpyAsync((uint8_t *)dst + pos, it.solo->d_pcm, it.bytes, cudaMemcpyDeviceToHost, st));
            pos += (it.bytes + 63) & ~63ull;
        } else {
            pcm_offsets[i] = ~0ull;
        }
    }
    CU_CHECK(cudaStreamSynchronize(st));
    return ZFLACB200_OK;
}

static int batch_item_ptr(zflacb200_batch *b, size_t i, void **p, size_t *bytes) {
    if (!b || i >= b->items.size()) return ZFLACB200_INVALID_ARGUMENT;
    const zflacb200_batch::Item &it = b->items[i];
    if (it.group >= 0) *p = (uint8_t *)b->groups[it.group].d_pcm + it.byte_off;
    else if (it.solo) *p = it.solo->d_pcm;
    else return ZFLACB200_INVALID_ARGUMENT;
    *bytes = it.bytes;
    return ZFLACB200_OK;
}

int zflacb200_batch_read_stream(zflacb200_batch *b, size_t i, void *dst, size_t dst_bytes) {
    void *p = nullptr;
    size_t bytes = 0;
    const int e = batch_item_ptr(b, i, &p, &bytes);
    if (e != ZFLACB200_OK) return e;
    if (dst_bytes < bytes || (!dst && bytes)) return ZFLACB200_INVALID_ARGUMENT;
    ApiScope api(b->device);
    CU_CHECK(cudaSetDevice(b->device));
    if (bytes) CU_CHECK(cudaMemcpy(dst, p, bytes, cudaMemcpyDeviceToHost));
    return ZFLACB200_OK;
}

int zflacb200_batch_device_pcm(zflacb200_batch *b, size_t i, void **dev_ptr, size_t *bytes) {
    if (!dev_ptr || !bytes) return ZFLACB200_INVALID_ARGUMENT;
    return batch_item_ptr(b, i, dev_ptr, bytes);
}

void zflacb200_batch_close(zflacb200_batch *b) {
    if (!b) return;
    for (auto &it : b->items)
        if (it.solo) zflacb200_close(it.solo);
    ApiScope api(b->device);
    cudaSetDevice(b->device);
    for (auto &g : b->groups)
        if (g.d_pcm) dev_free(g.d_pcm, g_ctx[b->device].stream);
    delete b;
}

// ---- metadata surface (host only): the blocks the reference's walk skips over (src/zflac.zig:243-247)
static int walk_metadata(const uint8_t *data, size_t len, std::vector<zflacb200_metadata_block> &blocks, uint64_t &audio_offset) {
    // same walk, same errors, same order as parse_metadata() / :217-253 + :309
    if (len < 4) return ZFLACB200_END_OF_STREAM;
    if (!(data[0] == 0x66 && data[1] == 0x4C && data[2] == 0x61 && data[3] == 0x43)) return ZFLACB200_INVALID_SIGNATURE;
    uint64_t pos = 4;
    bool have = false;
    for (;;) {
        if (len - pos < 4) return ZFLACB200_END_OF_STREAM;
        const uint8_t *b = data + pos;
        pos += 4;
        zflacb200_metadata_block mb;
        memset(&mb, 0, sizeof(mb));
        mb.type = b[0] & 0x7F;
        mb.is_last = (b[0] & 0x80) ? 1 : 0;
        mb.length = ((uint32_t)b[1] << 16) | ((uint32_t)b[2] << 8) | b[3];
        mb.offset = pos;
        if (mb.type == 0) {  // STREAMINFO is read as 34 bytes whatever its length field says (Q13)
            if (len - pos < 34) return ZFLACB200_END_OF_STREAM;
            mb.stored_bytes = 34;
            have = true;
        } else if (mb.type >= 1 && mb.type <= 6) {
            if (len - pos < mb.length) return ZFLACB200_END_OF_STREAM;
            mb.stored_bytes = mb.length;
        } else {
            return ZFLACB200_INVALID_METADATA_HEADER;
        }
        pos += mb.stored_bytes;
        blocks.push_back(mb);
        if (mb.is_last) break;
    }
    if (!have) return ZFLACB200_MISSING_STREAMINFO;
    audio_offset = pos;
    return ZFLACB200_OK;
}

static uint64_t be_uint(const uint8_t *p, int n) {
    uint64_t v = 0;
    for (int i = 0; i < n; ++i) v = (v << 8) | p[i];
    return v;
}

int zflacb200_metadata(const uint8_t *data, size_t len, zflacb200_metadata_block *dst, size_t capacity, size_t *n_blocks,
                       uint64_t *audio_offset) {
    if (!data || !n_blocks) return ZFLACB200_INVALID_ARGUMENT;
    std::vector<zflacb200_metadata_block> blocks;
    uint64_t ao = 0;
    const int e = walk_metadata(data, len, blocks, ao);
    *n_blocks = blocks.size();
    for (size_t i = 0; dst && i < blocks.size() && i < capacity; ++i) dst[i] = blocks[i];
    if (e != ZFLACB200_OK) return e;
    if (audio_offset) *audio_offset = ao;
    return ZFLACB200_OK;
}

int zflacb200_seektable(const uint8_t *data, size_t len, zflacb200_seekpoint *dst, size_t capacity, size_t *n_points) {
    if (!data || !n_points) return ZFLACB200_INVALID_ARGUMENT;
    *n_points = 0;
    std::vector<zflacb200_metadata_block> blocks;
    uint64_t ao = 0;
    const int e = walk_metadata(data, len, blocks, ao);
    if (e != ZFLACB200_OK) return e;
    for (const zflacb200_metadata_block &mb : blocks) {
        if (mb.type != 3) continue;
        const size_t n = mb.stored_bytes / 18;  // 64-bit sample number, 64-bit offset from the first frame, 16-bit frame samples
        *n_points = n;
        for (size_t i = 0; dst && i < n && i < capacity; ++i) {
            const uint8_t *p = data + mb.offset + 18 * i;
            dst[i].sample_number = be_uint(p, 8);
            dst[i].stream_offset = be_uint(p + 8, 8);
            dst[i].frame_samples = (uint32_t)be_uint(p + 16, 2);
            dst[i].is_placeholder = dst[i].sample_number == ~0ull ? 1u : 0u;
        }
        break;  // a stream has at most one SEEKTABLE; the first one counts
    }
    return ZFLACB200_OK;
}

int zflacb200_vorbis_comments(const uint8_t *data, size_t len, zflacb200_span *vendor, zflacb200_span *dst, size_t capacity,
                              size_t *n_comments) {
    if (!data || !n_comments) return ZFLACB200_INVALID_ARGUMENT;
    *n_comments = 0;
    if (vendor) vendor->offset = vendor->length = 0;
    std::vector<zflacb200_metadata_block> blocks;
    uint64_t ao = 0;
    const int e = walk_metadata(data, len, blocks, ao);
    if (e != ZFLACB200_OK) return e;
    for (const zflacb200_metadata_block &mb : blocks) {
        if (mb.type != 4) continue;
        // little-endian (Vorbis convention): vendor length + vendor, count, then count x (length + "NAME=value")
        const uint64_t end = mb.offset + mb.stored_bytes;
        uint64_t pos = mb.offset;
        auto le32 = [&](uint64_t at, uint32_t &v) {
            if (end - at < 4) return false;
            v = (uint32_t)data[at] | ((uint32_t)data[at + 1] << 8) | ((uint32_t)data[at + 2] << 16) | ((uint32_t)data[at + 3] << 24);
            return true;
        };
        uint32_t vlen = 0, count = 0;
        if (!le32(pos, vlen) || end - (pos + 4) < vlen) return ZFLACB200_INVALID_METADATA_HEADER;
        if (vendor) { vendor->offset = pos + 4; vendor->length = vlen; }
        pos += 4 + (uint64_t)vlen;
        if (!le32(pos, count)) return ZFLACB200_INVALID_METADATA_HEADER;
        pos += 4;
        size_t got = 0;
        for (uint32_t i = 0; i < count; ++i) {
            uint32_t clen = 0;
            if (!le32(pos, clen) || end - (pos + 4) < clen) return ZFLACB200_INVALID_METADATA_HEADER;
            if (dst && got < capacity) { dst[got].offset = pos + 4; dst[got].length = clen; }
            ++got;
            pos += 4 + (uint64_t)clen;
        }
        *n_comments = got;
        break;
    }
    return ZFLACB200_OK;
}

int zflacb200_pictures(const uint8_t *data, size_t len, zflacb200_picture *dst, size_t capacity, size_t *n_pictures) {
    if (!data || !n_pictures) return ZFLACB200_INVALID_ARGUMENT;
    *n_pictures = 0;
    std::vector<zflacb200_metadata_block> blocks;
    uint64_t ao = 0;
    const int e = walk_metadata(data, len, blocks, ao);
    if (e != ZFLACB200_OK) return e;
    size_t got = 0;
    for (const zflacb200_metadata_block &mb : blocks) {
        if (mb.type != 6) continue;
        const uint64_t end = mb.offset + mb.stored_bytes;
        uint64_t pos = mb.offset;
        auto be32 = [&](uint32_t &v) {  // big-endian, like every FLAC-native field
            if (end - pos < 4) return false;
            v = (uint32_t)be_uint(data + pos, 4);
            pos += 4;
            return true;
        };
        auto span = [&](zflacb200_span &sp) {
            uint32_t n = 0;
            if (!be32(n) || end - pos < n) return false;
            sp.offset = pos;
            sp.length = n;
            pos += n;
            return true;
        };
        zflacb200_picture p;
        memset(&p, 0, sizeof(p));
        if (!be32(p.picture_type) || !span(p.mime) || !span(p.description) || !be32(p.width) || !be32(p.height) || !be32(p.depth) ||
            !be32(p.colors) || !span(p.image))
            return ZFLACB200_INVALID_METADATA_HEADER;
        if (dst && got < capacity) dst[got] = p;
        ++got;
    }
    *n_pictures = got;
    return ZFLACB200_OK;
}

int zflacb200_cuesheet_read(const uint8_t *data, size_t len, zflacb200_cuesheet *sheet, zflacb200_cue_track *tracks,
                            size_t track_capacity, zflacb200_cue_index *indices, size_t index_capacity, int *found) {
    if (!data || !sheet || !found) return ZFLACB200_INVALID_ARGUMENT;
    *found = 0;
    memset(sheet, 0, sizeof(*sheet));
    std::vector<zflacb200_metadata_block> blocks;
    uint64_t ao = 0;
    const int e = walk_metadata(data, len, blocks, ao);
    if (e != ZFLACB200_OK) return e;
    for (const zflacb200_metadata_block &mb : blocks) {
        if (mb.type != 5) continue;
        // 128 catalog + 8 lead-in + 1 flags + 258 reserved + 1 track count = 396 bytes, then per track
        // 8 offset + 1 number + 12 ISRC + 1 flags + 13 reserved + 1 index count = 36, then per index 8 + 1 + 3 reserved = 12
        const uint64_t end = mb.offset + mb.stored_bytes;
        uint64_t pos = mb.offset;
        if (end - pos < 396) return ZFLACB200_INVALID_METADATA_HEADER;
        memcpy(sheet->media_catalog_number, data + pos, 128);
        sheet->media_catalog_number[128] = 0;
        sheet->lead_in_samples = be_uint(data + pos + 128, 8);
        sheet->is_cd = data[pos + 136] >> 7;
        const uint32_t nt = data[pos + 395];
        pos += 396;
        uint32_t ni = 0;
        for (uint32_t t = 0; t < nt; ++t) {
            if (end - pos < 36) return ZFLACB200_INVALID_METADATA_HEADER;
            zflacb200_cue_track tr;
            memset(&tr, 0, sizeof(tr));
            tr.offset_samples = be_uint(data + pos, 8);
            tr.number = data[pos + 8];
            memcpy(tr.isrc, data + pos + 9, 12);
            tr.is_audio = (data[pos + 21] >> 7) ? 0 : 1;
            tr.pre_emphasis = (data[pos + 21] >> 6) & 1;
            tr.first_index = ni;
            tr.n_indices = data[pos + 35];
            pos += 36;
            if (end - pos < 12ull * tr.n_indices) return ZFLACB200_INVALID_METADATA_HEADER;
            for (uint32_t k = 0; k < tr.n_indices; ++k, pos += 12) {
                if (indices && ni + k < index_capacity) {
                    indices[ni + k].offset_samples = be_uint(data + pos, 8);
                    indices[ni + k].number = data[pos + 8];
                    indices[ni + k].pad = 0;
                }
            }
            ni += tr.n_indices;
            if (tracks && t < track_capacity) tracks[t] = tr;
        }
        sheet->n_tracks = nt;
        sheet->n_indices = ni;
        *found = 1;
        break;
    }
    return ZFLACB200_OK;
}

int zflacb200_frames(zflacb200_stream *s, zflacb200_frame *dst, size_t capacity, size_t *n_frames) {
    if (!s || !n_frames) return ZFLACB200_INVALID_ARGUMENT;
    if (s->frames_on_device) {  // fast path: build the table from the device records on first use
        ApiScope api(s->device);
        CU_CHECK(cudaSetDevice(s->device));
        std::vector<FrameRec> recs(s->n_recs);
        std::vector<FrameOut> outs(s->n_recs);
        std::vector<uint8_t> ok(s->n_recs);
        CU_CHECK(cudaMemcpy(recs.data(), s->d_recs, sizeof(FrameRec) * s->n_recs, cudaMemcpyDeviceToHost));
        CU_CHECK(cudaMemcpy(outs.data(), s->d_outs, sizeof(FrameOut) * s->n_recs, cudaMemcpyDeviceToHost));
        CU_CHECK(cudaMemcpy(ok.data(), s->d_frame_ok, s->n_recs, cudaMemcpyDeviceToHost));
        s->frames.clear();
        for (uint32_t i = 0; i < s->n_recs; ++i) {
            if (recs[i].block_size == 0) continue;
            zflacb200_frame f;
            f.offset = recs[i].offset;
            f.end_offset = outs[i].end_offset;
            f.sample_offset = recs[i].sample_base;
            f.block_size = recs[i].block_size;
            f.header_len = recs[i].header_len;
            f.channel_code = recs[i].ch_code;
            f.crc16_ok = ok[i];
            f.probed = 0;
            s->frames.push_back(f);
        }
        s->frames_on_device = false;
    }
    *n_frames = s->frames.size();
    if (dst) memcpy(dst, s->frames.data(), sizeof(zflacb200_frame) * std::min(capacity, s->frames.size()));
    return ZFLACB200_OK;
}

}  // extern "C"
