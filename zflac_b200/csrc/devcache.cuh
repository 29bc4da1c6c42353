// devcache.cuh — the library's device-memory cache (host code).
//
// Every call of the decode path allocates the same handful of large buffers (stream bytes, planar
// scratch, PCM, frame records) and frees them when the call or the handle ends; the shapes repeat
// from call to call. cudaMallocAsync's pool re-maps physical memory when the request sizes of one
// call are served from blocks another call shaped (measured r02b: 8 ms + ~15 ms of host time per C5
// batch step). This cache keeps whole blocks: a freed block goes on a per-device free list together
// with an event recorded on the freeing stream; a request takes the smallest cached block that fits
// (and is not wastefully large) and, if it was freed on another stream, makes the new stream wait
// for that event. Only a miss calls cudaMalloc. 180 GB of HBM3e: the cache is allowed to hold on to
// ZFLACB200_CACHE_MAX_MIB (default: a quarter of the device's memory) of idle blocks; above that, and
// whenever cudaMalloc fails, idle blocks are returned to the driver; zflacb200_trim() returns all of them.
#pragma once
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdlib>
#include <iterator>
#include <map>
#include <mutex>
#include <unordered_map>

namespace zf {

class DevCache {
  public:
    // Stream-ordered: the block may be used by work enqueued on `st` after this returns.
    cudaError_t alloc(void **out, size_t bytes, cudaStream_t st) {
        *out = nullptr;
        int dev = 0;
        cudaError_t e = cudaGetDevice(&dev);
        if (e != cudaSuccess) return e;
        const size_t want = round_up(bytes);
        std::lock_guard<std::mutex> lock(m_);
        PerDevice &d = dev_[dev];
        auto it = d.idle.lower_bound(want);
        // at most 25 % of slack, plus 1 MiB for large requests (their sizes wobble from call to call) or 4 KiB for small ones
        if (it != d.idle.end() && it->first <= want + want / 4 + (want >= (1u << 20) ? (1u << 20) : 4096u)) {
            Block b = it->second;
            d.idle.erase(it);
            d.idle_bytes -= b.bytes;
            if (b.freed_on != st) {
                e = cudaStreamWaitEvent(st, b.ev, 0);
                if (e != cudaSuccess) return e;
            }
            d.live[b.p] = b;
            *out = b.p;
            ++hits_;
            return cudaSuccess;
        }
        Block b;
        b.bytes = want;
        e = cudaMalloc(&b.p, want);
        if (e != cudaSuccess) {  // give the idle blocks back and try once more
            cudaGetLastError();
            trim(d, 0);
            e = cudaMalloc(&b.p, want);
            if (e != cudaSuccess) return e;
        }
        e = cudaEventCreateWithFlags(&b.ev, cudaEventDisableTiming);
        if (e != cudaSuccess) {
            cudaFree(b.p);
            return e;
        }
        d.live[b.p] = b;
        *out = b.p;
        ++misses_;
        return cudaSuccess;
    }

    // Stream-ordered: work already enqueued on `st` may still use the block.
    void free(void *p, cudaStream_t st) {
        if (!p) return;
        int dev = 0;
        cudaGetDevice(&dev);
        std::lock_guard<std::mutex> lock(m_);
        PerDevice *d = &dev_[dev];
        auto it = d->live.find(p);
        if (it == d->live.end()) {  // freed with another device current: look the owner up
            d = nullptr;
            for (auto &kv : dev_) {
                it = kv.second.live.find(p);
                if (it != kv.second.live.end()) { d = &kv.second; break; }
            }
            if (!d) return;  // not ours
        }
        Block b = it->second;
        d->live.erase(it);
        if (cudaEventRecord(b.ev, st) != cudaSuccess) {  // cannot order the reuse: hand it back for good
            cudaGetLastError();
            cudaFree(b.p);
            cudaEventDestroy(b.ev);
            return;
        }
        b.freed_on = st;
        d->idle.emplace(b.bytes, b);
        d->idle_bytes += b.bytes;
        if (d->idle_bytes > limit()) trim(*d, limit() / 2);
    }

    // Returns every idle block of every device to the driver (zflacb200_shutdown).
    void release_all() {
        std::lock_guard<std::mutex> lock(m_);
        int cur = 0;
        cudaGetDevice(&cur);
        for (auto &kv : dev_) {
            cudaSetDevice(kv.first);
            trim(kv.second, 0);
        }
        cudaSetDevice(cur);
    }

    unsigned long long hits() const { return hits_; }
    unsigned long long misses() const { return misses_; }

  private:
    struct Block {
        void *p = nullptr;
        size_t bytes = 0;
        cudaEvent_t ev = nullptr;
        cudaStream_t freed_on = nullptr;
    };
    struct PerDevice {
        std::multimap<size_t, Block> idle;
        std::unordered_map<void *, Block> live;
        size_t idle_bytes = 0;
    };

    static size_t round_up(size_t bytes) {
        if (bytes < 512) return 512;
        if (bytes < (1u << 20)) return (bytes + 511) & ~(size_t)511;
        return (bytes + (2u << 20) - 1) & ~(size_t)((2u << 20) - 1);
    }
    // idle blocks the cache may sit on: ZFLACB200_CACHE_MAX_MIB, else a quarter of the device's memory (45 GiB on a B200) —
    // invisible to a co-resident allocator, so not more; zflacb200_trim() / zflacb200_shutdown() hand everything back
    static size_t limit() {
        static const size_t v = [] {
            const char *e = getenv("ZFLACB200_CACHE_MAX_MIB");
            if (e) return (size_t)strtoull(e, nullptr, 10) << 20;
            size_t free_b = 0, total_b = 0;
            if (cudaMemGetInfo(&free_b, &total_b) != cudaSuccess || total_b == 0) return (size_t)16 << 30;
            return total_b / 4;
        }();
        return v;
    }
    // Largest idle blocks first, until at most `keep` bytes stay cached. cudaFree waits for the device, so every
    // recorded event has fired by the time the memory goes away.
    void trim(PerDevice &d, size_t keep) {
        while (d.idle_bytes > keep && !d.idle.empty()) {
            auto it = std::prev(d.idle.end());
            cudaFree(it->second.p);
            cudaEventDestroy(it->second.ev);
            d.idle_bytes -= it->second.bytes;
            d.idle.erase(it);
        }
    }

    std::mutex m_;
    std::map<int, PerDevice> dev_;
    unsigned long long hits_ = 0, misses_ = 0;
};

}  // namespace zf
