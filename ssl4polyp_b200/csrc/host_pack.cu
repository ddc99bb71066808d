// Host side of the frame upload: gather N rectangles of uint8 rows (the crop boxes of decoded frames, or whole frames) into one
// pinned staging buffer with plain memcpy on a few threads.  Replaces the per-image work the reference leaves to DataLoader
// worker processes (classification/data/transforms.py:233-258 runs per sample in `__getitem__`; main_pretrain.py:156-175).
// One call for the whole batch: a ctypes / cgo caller gives up its interpreter lock once, so the parameter draws of the same
// batch run beside the copies.  No CUDA calls in here.
#include "common.cuh"
#include "ssl4p.h"
#include <cstring>
#include <thread>
#include <vector>

namespace {

struct Rect { const uint8_t* src; long long pitch, rows, row_bytes; uint8_t* dst; };

// rows [r0, r1) of one rectangle
inline void copy_rows(const Rect& q, long long r0, long long r1) {
    if (q.pitch == q.row_bytes) { memcpy(q.dst + r0 * q.row_bytes, q.src + r0 * q.pitch, (size_t)((r1 - r0) * q.row_bytes)); return; }
    for (long long r = r0; r < r1; ++r) memcpy(q.dst + r * q.row_bytes, q.src + r * q.pitch, (size_t)q.row_bytes);
}

}  // namespace

extern "C" int ssl4p_host_pack_rects(const void* const* src, const long long* src_pitch, const long long* rows, const long long* row_bytes,
                                     void* dst, const long long* dst_offset, int n, int threads) {
    SSL4P_REQUIRE(n >= 0 && threads >= 1 && threads <= 256, "host_pack_rects: bad n / threads");
    if (n == 0) return SSL4P_OK;
    SSL4P_REQUIRE(src && src_pitch && rows && row_bytes && dst && dst_offset, "host_pack_rects: null pointer");
    std::vector<Rect> rects((size_t)n);
    long long total = 0;
    for (int i = 0; i < n; ++i) {
        SSL4P_REQUIRE(src[i] != nullptr && rows[i] >= 0 && row_bytes[i] >= 0 && dst_offset[i] >= 0 && (src_pitch[i] >= row_bytes[i] || rows[i] <= 1),
                      "host_pack_rects: rectangle %d is malformed", i);
        rects[(size_t)i] = Rect{(const uint8_t*)src[i], src_pitch[i], rows[i], row_bytes[i], (uint8_t*)dst + dst_offset[i]};
        total += rows[i] * row_bytes[i];
    }
    // thread t takes the byte range [t, t+1) * total / threads of the concatenated rectangles, cut at row boundaries
    auto work = [&rects, n, total, threads](int t) {
        const long long lo = total * t / threads, hi = total * (t + 1) / threads;
        long long base = 0;
        for (int i = 0; i < n && base < hi; ++i) {
            const Rect& q = rects[(size_t)i];
            const long long bytes = q.rows * q.row_bytes;
            if (bytes > 0 && base + bytes > lo) {
                // first row whose start offset is >= lo - base, first row whose start offset is >= hi - base
                const long long a = lo > base ? (lo - base + q.row_bytes - 1) / q.row_bytes : 0;
                const long long b = hi < base + bytes ? (hi - base + q.row_bytes - 1) / q.row_bytes : q.rows;
                if (b > a) copy_rows(q, a, b);
            }
            base += bytes;
        }
    };
    if (threads == 1 || total < (1 << 20)) {
        for (int i = 0; i < n; ++i) copy_rows(rects[(size_t)i], 0, rects[(size_t)i].rows);
        return SSL4P_OK;
    }
    std::vector<std::thread> pool;
    pool.reserve((size_t)threads - 1);
    for (int t = 1; t < threads; ++t) pool.emplace_back(work, t);
    work(0);
    for (auto& th : pool) th.join();
    return SSL4P_OK;
}
