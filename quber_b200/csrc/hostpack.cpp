// Host side of the packed mask transfer (end-to-end path only).
//
// The reference hands MaskRefinerPredictor.predict its initial masks as a [K,H,W] uint8 array, 0 outside and one value
// inside (maskrefiner/predictor.py:304-310), and the batched path here moved it to the GPU byte for byte.  At C2 that
// is 295 MB per 32 frames and the host link, not the GPU, sets the end-to-end rate.  These functions turn the bytes
// into one bit per pixel on the host cores (persistent worker threads, AVX2 compare + movemask, 32 pixels per word), so
// 1/8 of the bytes cross the link; quber_unpack_masks (fusion.cu) expands them again in HBM.  The one value the set
// pixels carry (1, or 255 as eval/refiner_model.py:230-231 makes them) is returned beside the words; a chunk with more
// than one non-zero value is reported (value -1) and the caller sends its bytes as they are.
//
// One job at a time: quber_host_pack_begin cuts the buffer into the caller's chunks and those into 256 KB blocks that the
// workers take in address order from one counter, so the chunks finish in order and the caller (which keeps enqueueing
// GPU work meanwhile) picks them up with quber_host_pack_wait — no wake-up per chunk, no idle tail on small chunks; a
// caller that has to wait packs blocks itself until its chunk is complete.
#include <immintrin.h>
#include <stddef.h>
#include <stdint.h>

#include <atomic>
#include <condition_variable>
#include <mutex>
#include <thread>
#include <vector>

#include "../../include/quber_b200.h"

namespace {

// bit i of dst[w] = src[32 w + i] != 0; also the largest byte and the smallest NON-ZERO byte (255 if none) of the
// range: the range is "binary" when the two agree (every set pixel carries the same value).
struct Range { uint8_t mx = 0, mn = 255; };

Range pack_scalar(const uint8_t *src, size_t n_words, uint32_t *dst) {
    Range r;
    for (size_t w = 0; w < n_words; ++w) {
        uint32_t bits = 0;
        for (int i = 0; i < 32; ++i) {
            const uint8_t v = src[32 * w + i];
            if (v > r.mx) r.mx = v;
            if (v && v < r.mn) r.mn = v;
            bits |= (uint32_t)(v != 0) << i;
        }
        dst[w] = bits;
    }
    return r;
}

__attribute__((target("avx2"))) Range pack_avx2(const uint8_t *src, size_t n_words, uint32_t *dst) {
    const __m256i zero = _mm256_setzero_si256();
    __m256i mx = zero, mn = _mm256_set1_epi8((char)0xff);
    size_t w = 0;
    for (; w + 4 <= n_words; w += 4) {
        const __m256i a = _mm256_loadu_si256(reinterpret_cast<const __m256i *>(src + 32 * w));
        const __m256i b = _mm256_loadu_si256(reinterpret_cast<const __m256i *>(src + 32 * w + 32));
        const __m256i c = _mm256_loadu_si256(reinterpret_cast<const __m256i *>(src + 32 * w + 64));
        const __m256i d = _mm256_loadu_si256(reinterpret_cast<const __m256i *>(src + 32 * w + 96));
        const __m256i za = _mm256_cmpeq_epi8(a, zero), zb = _mm256_cmpeq_epi8(b, zero);
        const __m256i zc = _mm256_cmpeq_epi8(c, zero), zd = _mm256_cmpeq_epi8(d, zero);
        mx = _mm256_max_epu8(mx, _mm256_max_epu8(_mm256_max_epu8(a, b), _mm256_max_epu8(c, d)));
        // zero bytes become 255 (v | 0xff) and drop out of the minimum
        mn = _mm256_min_epu8(mn, _mm256_min_epu8(_mm256_min_epu8(_mm256_or_si256(a, za), _mm256_or_si256(b, zb)),
                                                 _mm256_min_epu8(_mm256_or_si256(c, zc), _mm256_or_si256(d, zd))));
        dst[w] = ~(uint32_t)_mm256_movemask_epi8(za);
        dst[w + 1] = ~(uint32_t)_mm256_movemask_epi8(zb);
        dst[w + 2] = ~(uint32_t)_mm256_movemask_epi8(zc);
        dst[w + 3] = ~(uint32_t)_mm256_movemask_epi8(zd);
    }
    alignas(32) uint8_t lx[32], ln[32];
    _mm256_store_si256(reinterpret_cast<__m256i *>(lx), mx);
    _mm256_store_si256(reinterpret_cast<__m256i *>(ln), mn);
    Range r;
    for (int i = 0; i < 32; ++i) {
        if (lx[i] > r.mx) r.mx = lx[i];
        if (ln[i] < r.mn) r.mn = ln[i];
    }
    if (w < n_words) {
        const Range t = pack_scalar(src + 32 * w, n_words - w, dst + w);
        if (t.mx > r.mx) r.mx = t.mx;
        if (t.mn < r.mn) r.mn = t.mn;
    }
    return r;
}

constexpr int MAX_CHUNKS = 256, MAX_THREADS = 64;
constexpr size_t BLOCK_WORDS = 8192;                        // 256 KB of mask bytes per block

struct Chunk {
    size_t w0 = 0, w1 = 0;                                  // word range
    size_t b0 = 0;                                          // first block id
    std::atomic<long> remaining{0};                         // blocks not finished yet
    std::atomic<uint32_t> mx{0}, mn{255};
};

struct Pool {
    std::mutex m;
    std::condition_variable cv_go, cv_done;
    std::vector<std::thread> workers;
    const uint8_t *src = nullptr;
    uint32_t *dst = nullptr;
    Chunk chunks[MAX_CHUNKS];
    int n_chunks = 0, parts = 0;
    size_t n_blocks = 0;
    std::atomic<size_t> next{0};
    std::atomic<long> blocks_left{0};
    uint64_t epoch = 0;
    int busy = 0;                                           // workers inside the current job
    bool avx2 = false, stop = false;
    std::mutex entry;                                       // serialises begin / synchronous callers

    // until no block is left — or, for a waiting caller that lends a hand, until its chunk is complete
    void run_blocks(const Chunk *until = nullptr) {
        for (;;) {
            if (until && until->remaining.load(std::memory_order_acquire) == 0) return;
            const size_t b = next.fetch_add(1, std::memory_order_relaxed);
            if (b >= n_blocks) return;
            int c = 0;
            while (c + 1 < n_chunks && chunks[c + 1].b0 <= b) ++c;
            Chunk &ch = chunks[c];
            const size_t lo = ch.w0 + (b - ch.b0) * BLOCK_WORDS, hi = lo + BLOCK_WORDS < ch.w1 ? lo + BLOCK_WORDS : ch.w1;
            const Range r = avx2 ? pack_avx2(src + 32 * lo, hi - lo, dst + lo) : pack_scalar(src + 32 * lo, hi - lo, dst + lo);
            uint32_t cur = ch.mx.load(std::memory_order_relaxed);
            while (r.mx > cur && !ch.mx.compare_exchange_weak(cur, r.mx, std::memory_order_relaxed)) {}
            cur = ch.mn.load(std::memory_order_relaxed);
            while (r.mn < cur && !ch.mn.compare_exchange_weak(cur, r.mn, std::memory_order_relaxed)) {}
            const long left = ch.remaining.fetch_sub(1, std::memory_order_acq_rel) - 1;
            const long all = blocks_left.fetch_sub(1, std::memory_order_acq_rel) - 1;
            if (left == 0 || all == 0) {                    // somebody may be waiting for this chunk / for the job
                std::lock_guard<std::mutex> lk(m);
                cv_done.notify_all();
            }
        }
    }
    void worker(int id) {
        uint64_t seen = 0;
        for (;;) {
            {
                std::unique_lock<std::mutex> lk(m);
                cv_go.wait(lk, [&] { return stop || (epoch != seen && id < parts); });
                if (stop) return;
                seen = epoch;
                ++busy;
            }
            run_blocks();
            {
                std::lock_guard<std::mutex> lk(m);
                if (--busy == 0) cv_done.notify_all();
            }
        }
    }
    void ensure(int n) {
        while ((int)workers.size() < n) {
            const int id = (int)workers.size();
            workers.emplace_back([this, id] { worker(id); });
        }
    }
    // previous job fully drained: every block done and no worker still inside run_blocks
    void drain() {
        std::unique_lock<std::mutex> lk(m);
        cv_done.wait(lk, [&] { return blocks_left.load(std::memory_order_acquire) == 0 && busy == 0; });
    }
    ~Pool() {
        {
            std::lock_guard<std::mutex> lk(m);
            stop = true;
        }
        cv_go.notify_all();
        for (auto &t : workers) t.join();
    }
};

Pool &pool() {
    static Pool p;
    return p;
}

int begin_locked(Pool &p, const uint8_t *src, size_t n_bytes, uint32_t *dst, int n_threads, const size_t *chunk_end_bytes,
                 int n_chunks, bool caller_helps) {
    if (!src || !dst || n_bytes == 0 || n_bytes % 32 != 0 || n_chunks < 1 || n_chunks > MAX_CHUNKS || !chunk_end_bytes)
        return QUBER_E_INVALID;
    size_t prev = 0;
    for (int c = 0; c < n_chunks; ++c) {
        if (chunk_end_bytes[c] <= prev || chunk_end_bytes[c] % 32 != 0) return QUBER_E_INVALID;
        prev = chunk_end_bytes[c];
    }
    if (prev != n_bytes) return QUBER_E_INVALID;
    p.drain();
    int parts = n_threads < 1 ? 1 : (n_threads > MAX_THREADS ? MAX_THREADS : n_threads);
    const size_t total_blocks = (n_bytes / 32 + BLOCK_WORDS - 1) / BLOCK_WORDS;
    if ((size_t)parts > total_blocks) parts = (int)total_blocks;
    const int helpers = caller_helps ? parts - 1 : parts;  // the synchronous entry works too
    p.avx2 = __builtin_cpu_supports("avx2");
    p.ensure(helpers);
    {
        std::lock_guard<std::mutex> lk(p.m);
        p.src = src; p.dst = dst; p.n_chunks = n_chunks; p.parts = helpers;
        size_t w = 0, b = 0;
        for (int c = 0; c < n_chunks; ++c) {
            Chunk &ch = p.chunks[c];
            ch.w0 = w; ch.w1 = chunk_end_bytes[c] / 32; ch.b0 = b;
            const size_t nb = (ch.w1 - ch.w0 + BLOCK_WORDS - 1) / BLOCK_WORDS;
            ch.remaining.store((long)nb, std::memory_order_relaxed);
            ch.mx.store(0, std::memory_order_relaxed); ch.mn.store(255, std::memory_order_relaxed);
            w = ch.w1; b += nb;
        }
        p.n_blocks = b;
        p.next.store(0, std::memory_order_relaxed);
        p.blocks_left.store((long)b, std::memory_order_release);
        ++p.epoch;
    }
    if (helpers > 0) p.cv_go.notify_all();
    return QUBER_OK;
}

int wait_chunk(Pool &p, int chunk, int *value) {
    if (chunk < 0 || chunk >= p.n_chunks || !value) return QUBER_E_INVALID;
    Chunk &ch = p.chunks[chunk];
    if (ch.remaining.load(std::memory_order_acquire) != 0) {   // not there yet: pack blocks too instead of sleeping
        {
            std::lock_guard<std::mutex> lk(p.m);
            ++p.busy;
        }
        p.run_blocks(&ch);
        {
            std::lock_guard<std::mutex> lk(p.m);
            if (--p.busy == 0) p.cv_done.notify_all();
        }
    }
    for (int spin = 0; spin < 2000 && ch.remaining.load(std::memory_order_acquire) != 0; ++spin) _mm_pause();
    if (ch.remaining.load(std::memory_order_acquire) != 0) {
        std::unique_lock<std::mutex> lk(p.m);
        p.cv_done.wait(lk, [&] { return ch.remaining.load(std::memory_order_acquire) == 0; });
    }
    const uint32_t mx = ch.mx.load(std::memory_order_relaxed), mn = ch.mn.load(std::memory_order_relaxed);
    *value = mx == 0 ? 0 : (mx == mn ? (int)mx : -1);
    return QUBER_OK;
}

}  // namespace

extern "C" int quber_host_pack_begin(const uint8_t *src, size_t n_bytes, uint32_t *dst, int n_threads,
                                     const size_t *chunk_end_bytes, int n_chunks) {
    Pool &p = pool();
    std::lock_guard<std::mutex> entry(p.entry);
    return begin_locked(p, src, n_bytes, dst, n_threads, chunk_end_bytes, n_chunks, false);
}

extern "C" int quber_host_pack_wait(int chunk, int *value) { return wait_chunk(pool(), chunk, value); }

extern "C" int quber_host_pack_masks(const uint8_t *src, size_t n_bytes, uint32_t *dst, int n_threads, int *value) {
    if (!value) return QUBER_E_INVALID;
    Pool &p = pool();
    std::lock_guard<std::mutex> entry(p.entry);
    const size_t end = n_bytes;
    if (int rc = begin_locked(p, src, n_bytes, dst, n_threads, &end, 1, true)) return rc;
    p.run_blocks();                                         // the caller takes blocks as well
    return wait_chunk(p, 0, value);
}
