// Host side of the packed mask transfer (end-to-end path only).
//
// The reference hands MaskRefinerPredictor.predict its initial masks as a [K,H,W] uint8 array, 0 outside and one value
// inside (maskrefiner/predictor.py:304-310), and the batched path here moves it to the GPU byte for byte.  At C2 that is 295 MB per 32 frames and the
// host link, not the GPU, sets the end-to-end rate.  quber_host_pack_masks turns the bytes into one bit per pixel on the
// host cores (persistent worker threads, AVX2 compare + movemask, 32 pixels per word), so 1/8 of the bytes cross the
// link; quber_unpack_masks (fusion.cu) expands them again in HBM.  The one value the set pixels carry (1, or 255 as
// eval/refiner_model.py:230-231 makes them) is returned beside the words; a buffer with more than one non-zero value is
// reported (*value = -1) and the caller sends the bytes as they are.
#include <immintrin.h>
#include <stddef.h>
#include <stdint.h>

#include <condition_variable>
#include <mutex>
#include <thread>
#include <vector>

#include "../../include/quber_b200.h"

namespace {

// bit i of dst[w] = src[32 w + i] != 0; also the largest byte and the smallest NON-ZERO byte (255 if none) of the
// range: the buffer is "binary" when the two agree (every set pixel carries the same value).
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

struct Pool {
    std::mutex m;
    std::condition_variable cv_go, cv_done;
    std::vector<std::thread> workers;
    // one job at a time (the public entry serialises callers)
    const uint8_t *src = nullptr;
    uint32_t *dst = nullptr;
    size_t n_words = 0;
    int parts = 0;
    uint64_t epoch = 0;
    int pending = 0;
    Range part_range[64];
    bool avx2 = false, stop = false;
    std::mutex entry;

    void run_part(int part) {
        const size_t per = ((n_words + parts - 1) / parts + 3) & ~size_t(3);
        const size_t lo = (size_t)part * per, hi = lo + per < n_words ? lo + per : n_words;
        part_range[part] = Range();
        if (lo >= hi) return;
        part_range[part] = avx2 ? pack_avx2(src + 32 * lo, hi - lo, dst + lo) : pack_scalar(src + 32 * lo, hi - lo, dst + lo);
    }
    void worker(int id) {
        uint64_t seen = 0;
        for (;;) {
            {
                std::unique_lock<std::mutex> lk(m);
                cv_go.wait(lk, [&] { return stop || (epoch != seen && id < parts); });
                if (stop) return;
                seen = epoch;
            }
            run_part(id);
            {
                std::lock_guard<std::mutex> lk(m);
                if (--pending == 0) cv_done.notify_one();
            }
        }
    }
    void ensure(int n) {                                   // workers 1..n-1; the caller is part 0
        while ((int)workers.size() < n - 1) {
            const int id = (int)workers.size() + 1;
            workers.emplace_back([this, id] { worker(id); });
        }
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

}  // namespace

extern "C" int quber_host_pack_masks(const uint8_t *src, size_t n_bytes, uint32_t *dst, int n_threads, int *value) {
    if (!src || !dst || !value || n_bytes == 0 || n_bytes % 32 != 0) return QUBER_E_INVALID;
    Pool &p = pool();
    std::lock_guard<std::mutex> entry(p.entry);
    const size_t n_words = n_bytes / 32;
    int parts = n_threads < 1 ? 1 : (n_threads > 64 ? 64 : n_threads);
    if (n_words < (size_t)parts * 4096) parts = 1;         // small buffers: not worth waking anybody
    p.avx2 = __builtin_cpu_supports("avx2");
    p.ensure(parts);
    {
        std::lock_guard<std::mutex> lk(p.m);
        p.src = src; p.dst = dst; p.n_words = n_words; p.parts = parts;
        p.pending = parts - 1;
        ++p.epoch;
    }
    if (parts > 1) p.cv_go.notify_all();
    p.run_part(0);
    if (parts > 1) {
        std::unique_lock<std::mutex> lk(p.m);
        p.cv_done.wait(lk, [&] { return p.pending == 0; });
    }
    Range r;
    for (int i = 0; i < parts; ++i) {
        if (p.part_range[i].mx > r.mx) r.mx = p.part_range[i].mx;
        if (p.part_range[i].mn < r.mn) r.mn = p.part_range[i].mn;
    }
    *value = r.mx == 0 ? 0 : (r.mx == r.mn ? (int)r.mx : -1);
    return QUBER_OK;
}
