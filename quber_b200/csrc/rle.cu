// COCO run-length encoded masks on the device (the wire format of the reference's datasets).
//
// Reference call sites (pycocotools.mask, a third-party dependency of the reference; its algorithm is
// cocoapi/common/maskApi.c: rleEncode / rleDecode / rleToString / rleFrString):
//   mask_util.decode   maskrefiner/data/dataset_mappers/perturbed_panoptic_dataset_mapper.py:195-214 (initial masks and
//                      the eight t*/f* error maps), tools/ours/panoptic2eee.py:95-97
//   mask_util.encode   tools/ours/panoptic2eee.py:61-64,125-132, tools/ours/perturbate_masks.py:222
//   frPyObjects        maskrefiner/data/datasets/register_uoais_sim_instance.py:207-209 (uncompressed -> compressed)
//
// An RLE is the list of run lengths of the mask read in COLUMN-major order, starting with a (possibly empty) run of
// zeros.  Everything else in this library is row-major, so decode is "which run holds column-major index x*H + y" for a
// row-major pixel (y, x), and encode walks columns.  The initial masks of a 640x480 frame are ~150 runs each: 30 masks
// are ~20 KB of counts instead of 9.2 MB of bytes, which is what takes the host link out of the end-to-end call.
//
// decode:  rle_scan_kernel (inclusive prefix sums of the counts, one CTA per mask)
//          rle_fill_runs_kernel (bit planes: cleared, then every run of ones ORs its own bits in)
//          rle_fill_kernel (byte planes: lane = column, binary search for the run holding the first row of the chunk,
//                           then a walk down the column)
// encode:  rle_count_kernel -> rle_colscan_kernel -> rle_offsets_kernel -> rle_positions_kernel -> rle_diff_kernel
#include "common.cuh"

namespace quber {

constexpr int RLE_ROWS = 32;        // rows per CTA of the fill kernel
constexpr int RLE_THREADS = 256;    // 8 warps = 256 columns

__global__ void __launch_bounds__(1024)
rle_scan_kernel(const uint32_t *__restrict__ counts, const int64_t *__restrict__ offsets, uint32_t *__restrict__ ends,
                unsigned area, int32_t *__restrict__ status) {
    __shared__ unsigned warp_tot[32];
    __shared__ unsigned s_carry;
    const int i = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int64_t o0 = offsets[i], o1 = offsets[i + 1];
    if (tid == 0) s_carry = 0u;
    __syncthreads();
    for (int64_t base = o0; base < o1; base += 1024) {
        const int64_t j = base + tid;
        const unsigned v = j < o1 ? __ldg(counts + j) : 0u;
        unsigned inc = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const unsigned t = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += t;
        }
        if (lane == 31) warp_tot[warp] = inc;
        __syncthreads();
        unsigned wbase = 0, tot = 0;
        for (int w = 0; w < 32; ++w) {
            const unsigned t = warp_tot[w];
            if (w < warp) wbase += t;
            tot += t;
        }
        const unsigned carry = s_carry;
        if (j < o1) ends[j] = carry + wbase + inc;
        __syncthreads();
        if (tid == 0) s_carry = carry + tot;
        __syncthreads();
    }
    // maskApi.c rleDecode trusts the counts; here a sum different from H*W is reported (status bit 0) and the mask is
    // decoded as far as the runs go (pixels past the last run are 0, runs past H*W are cut)
    if (tid == 0 && s_carry != area) atomicOr(status, 1);
}

// Byte output: lane = column, binary search for the run holding the first row of the chunk, then a walk down the column.
__global__ void __launch_bounds__(RLE_THREADS)
rle_fill_kernel(const uint32_t *__restrict__ ends, const int64_t *__restrict__ offsets, int H, int W, int value,
                uint8_t *__restrict__ bytes) {
    const int i = blockIdx.z, tid = threadIdx.x;
    const int x = blockIdx.x * RLE_THREADS + tid, y0 = blockIdx.y * RLE_ROWS;
    const int rows = min(RLE_ROWS, H - y0);
    const uint32_t *e = ends + offsets[i];
    const int m = (int)(offsets[i + 1] - offsets[i]);
    if (x >= W) return;
    // j = index of the run holding column-major index idx = x*H + y0: the first run whose inclusive end exceeds idx
    unsigned idx = (unsigned)x * (unsigned)H + (unsigned)y0;
    int lo = 0, hi = m;
    while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (__ldg(e + mid) > idx) hi = mid; else lo = mid + 1;
    }
    int j = lo;
    unsigned end = j < m ? __ldg(e + j) : 0xffffffffu;
    const size_t plane = (size_t)i * H;
    for (int r = 0; r < rows; ++r, ++idx) {
        while (idx >= end) {                              // zero-length runs are stepped over here as well
            ++j;
            end = j < m ? __ldg(e + j) : 0xffffffffu;
        }
        bytes[(plane + y0 + r) * (size_t)W + x] = (j < m && (j & 1)) ? (uint8_t)value : (uint8_t)0;
    }
}

// Packed output, run-major: the planes are cleared by one memset and every run of ones sets its own bits (column-major
// index idx -> pixel (idx % H, idx / H) -> bit x % 32 of word (y, x / 32)) with atomicOr.  Work is proportional to the set
// pixels (~4 000 per mask at C2) instead of to H*W, and there is no per-chunk binary search over the runs.  Thread =
// one count of the flat counts array; its mask is found by bisection of the n + 1 offsets; even runs (zeros) leave.
__global__ void __launch_bounds__(256)
rle_fill_runs_kernel(const uint32_t *__restrict__ ends, const int64_t *__restrict__ offsets, int n, size_t total, int H, int W,
                     uint32_t *__restrict__ packed) {
    const size_t g = (size_t)blockIdx.x * 256 + threadIdx.x;
    if (g >= total) return;
    int lo = 0, hi = n;                                 // the mask i with offsets[i] <= g < offsets[i + 1]
    while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if ((size_t)__ldg(offsets + mid) <= g) lo = mid; else hi = mid;
    }
    const int i = lo;
    const size_t o0 = (size_t)__ldg(offsets + i);
    if (g >= (size_t)__ldg(offsets + i + 1)) return;    // counts past the last mask / empty masks in between
    const size_t j = g - o0;
    if (!(j & 1)) return;                               // odd runs are the ones
    const unsigned area = (unsigned)H * (unsigned)W;
    const unsigned a = min(__ldg(ends + g - 1), area), b = min(__ldg(ends + g), area);   // runs past H*W are cut
    if (a >= b) return;
    const int WB = W / 32;
    uint32_t *plane = packed + (size_t)i * H * WB;
    unsigned x = a / (unsigned)H, y = a - x * (unsigned)H;
    for (unsigned idx = a; idx < b; ++idx) {
        atomicOr(plane + (size_t)y * WB + (x >> 5), 1u << (x & 31));
        if (++y == (unsigned)H) { y = 0; ++x; }
    }
}

// ---- encode ----------------------------------------------------------------------------------------------------
// value of the pixel preceding (y = 0, x) in column-major order: (H - 1, x - 1), 0 in front of the first pixel
__device__ __forceinline__ unsigned rle_prev_of_column(const uint8_t *__restrict__ mask, int H, int W, int x) {
    return x > 0 ? (unsigned)(mask[(size_t)(H - 1) * W + x - 1] != 0) : 0u;
}

// WRITE = false: transitions per column -> colcnt[i][x].  WRITE = true: their column-major positions -> pos[...].
template <bool WRITE>
__global__ void __launch_bounds__(256)
rle_walk_kernel(const uint8_t *__restrict__ masks, int n, int H, int W, uint32_t *__restrict__ colcnt,
                const int64_t *__restrict__ offsets, uint32_t *__restrict__ pos, size_t cap) {
    const int i = blockIdx.y, x = blockIdx.x * 256 + threadIdx.x;
    if (x >= W) return;
    const uint8_t *mask = masks + (size_t)i * H * W;
    unsigned p = rle_prev_of_column(mask, H, W, x);
    unsigned c = 0;
    size_t o = 0;
    if (WRITE) o = (size_t)offsets[i] + colcnt[(size_t)i * W + x];      // colcnt holds the exclusive scan by now
    const unsigned base = (unsigned)x * (unsigned)H;
    for (int y = 0; y < H; ++y) {
        const unsigned v = mask[(size_t)y * W + x] != 0;
        if (v != p) {
            if (WRITE) { if (o + c < cap) pos[o + c] = base + (unsigned)y; }
            ++c;
            p = v;
        }
    }
    if (!WRITE) colcnt[(size_t)i * W + x] = c;
}

// exclusive scan of the column counts of one mask (in place) and its number of runs m = transitions + 1
__global__ void __launch_bounds__(1024)
rle_colscan_kernel(uint32_t *__restrict__ colcnt, int W, int32_t *__restrict__ runs) {
    __shared__ unsigned warp_tot[32];
    __shared__ unsigned s_carry;
    const int i = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    uint32_t *c = colcnt + (size_t)i * W;
    if (tid == 0) s_carry = 0u;
    __syncthreads();
    for (int base = 0; base < W; base += 1024) {
        const int j = base + tid;
        const unsigned v = j < W ? c[j] : 0u;
        unsigned inc = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const unsigned t = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += t;
        }
        if (lane == 31) warp_tot[warp] = inc;
        __syncthreads();
        unsigned wbase = 0, tot = 0;
        for (int w = 0; w < 32; ++w) {
            const unsigned t = warp_tot[w];
            if (w < warp) wbase += t;
            tot += t;
        }
        const unsigned carry = s_carry;
        if (j < W) c[j] = carry + wbase + inc - v;
        __syncthreads();
        if (tid == 0) s_carry = carry + tot;
        __syncthreads();
    }
    if (tid == 0) runs[i] = (int32_t)s_carry + 1;
}

__global__ void __launch_bounds__(1024)
rle_offsets_kernel(const int32_t *__restrict__ runs, int n, int64_t *__restrict__ offsets, size_t cap, int32_t *__restrict__ status) {
    __shared__ long long warp_tot[32];
    __shared__ long long s_carry;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) s_carry = 0;
    __syncthreads();
    for (int base = 0; base < n; base += 1024) {
        const int j = base + tid;
        const long long v = j < n ? runs[j] : 0;
        long long inc = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const long long t = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += t;
        }
        if (lane == 31) warp_tot[warp] = inc;
        __syncthreads();
        long long wbase = 0, tot = 0;
        for (int w = 0; w < 32; ++w) {
            const long long t = warp_tot[w];
            if (w < warp) wbase += t;
            tot += t;
        }
        const long long carry = s_carry;
        if (j < n) offsets[j] = carry + wbase + inc - v;
        __syncthreads();
        if (tid == 0) s_carry = carry + tot;
        __syncthreads();
    }
    if (tid == 0) {
        offsets[n] = s_carry;
        if ((unsigned long long)s_carry > (unsigned long long)cap) atomicOr(status, 2);      // counts buffer too small
    }
}

// counts[j] = pos[j] - pos[j-1]; the first run starts at 0, the last one ends at H*W (maskApi.c rleEncode)
__global__ void __launch_bounds__(256)
rle_diff_kernel(const uint32_t *__restrict__ pos, const int64_t *__restrict__ offsets, unsigned area, size_t cap,
                uint32_t *__restrict__ counts) {
    const int i = blockIdx.y;
    const int64_t o0 = offsets[i], m = offsets[i + 1] - o0;
    for (int64_t j = (int64_t)blockIdx.x * 256 + threadIdx.x; j < m; j += (int64_t)gridDim.x * 256) {
        if ((size_t)(o0 + j) >= cap) return;
        const unsigned hi = j < m - 1 ? pos[o0 + j] : area;
        const unsigned lo = j > 0 ? pos[o0 + j - 1] : 0u;
        counts[o0 + j] = hi - lo;
    }
}

}  // namespace quber

extern "C" size_t quber_rle_decode_workspace_bytes(size_t total_counts) {
    return quber::align_up((total_counts > 0 ? total_counts : 1) * sizeof(uint32_t), 256);
}

extern "C" int quber_rle_decode(const uint32_t *counts, const int64_t *offsets, size_t total_counts, int n, int H, int W,
                                int value, uint32_t *packed_out, uint8_t *masks_out, void *workspace, size_t workspace_bytes,
                                int32_t *status, void *stream) {
    using namespace quber;
    if (!counts || !offsets || !status || (!packed_out && !masks_out)) return QUBER_E_INVALID;
    if (n <= 0 || H <= 0 || W <= 0 || value < 1 || value > 255 || n > 65535) return QUBER_E_INVALID;
    if ((long long)H * W >= (1ll << 32) - 1) return QUBER_E_UNSUPPORTED;
    if (packed_out && W % 32 != 0) return QUBER_E_UNSUPPORTED;
    if (!workspace || !aligned16(workspace) || workspace_bytes < quber_rle_decode_workspace_bytes(total_counts)) return QUBER_E_WORKSPACE;
    cudaStream_t st = (cudaStream_t)stream;
    uint32_t *ends = static_cast<uint32_t *>(workspace);
    QUBER_CUDA_TRY(cudaMemsetAsync(status, 0, sizeof(int32_t), st));
    rle_scan_kernel<<<n, 1024, 0, st>>>(counts, offsets, ends, (unsigned)((long long)H * W), status);
    if (int rc = check_launch("rle_scan_kernel")) return rc;
    const dim3 grid(cdiv(W, RLE_THREADS), cdiv(H, RLE_ROWS), n);
    if (packed_out) {
        QUBER_CUDA_TRY(cudaMemsetAsync(packed_out, 0, (size_t)n * H * (W / 32) * sizeof(uint32_t), st));
        if (total_counts > 0)
            rle_fill_runs_kernel<<<(unsigned)((total_counts + 255) / 256), 256, 0, st>>>(ends, offsets, n, total_counts, H, W, packed_out);
        if (int rc = check_launch("rle_fill_runs_kernel")) return rc;
    }
    if (masks_out) {
        rle_fill_kernel<<<grid, RLE_THREADS, 0, st>>>(ends, offsets, H, W, value, masks_out);
        if (int rc = check_launch("rle_fill_kernel<bytes>")) return rc;
    }
    return QUBER_OK;
}

extern "C" size_t quber_rle_encode_workspace_bytes(int n, int H, int W, size_t cap) {
    if (n <= 0 || H <= 0 || W <= 0) return 0;
    return quber::align_up((size_t)n * W * sizeof(uint32_t), 256) + quber::align_up((size_t)n * sizeof(int32_t), 256) +
           quber::align_up((cap > 0 ? cap : 1) * sizeof(uint32_t), 256);
}

extern "C" int quber_rle_encode(const uint8_t *masks, int n, int H, int W, uint32_t *counts_out, size_t cap, int64_t *offsets_out,
                                void *workspace, size_t workspace_bytes, int32_t *status, void *stream) {
    using namespace quber;
    if (!masks || !counts_out || !offsets_out || !status) return QUBER_E_INVALID;
    if (n <= 0 || H <= 0 || W <= 0 || n > 65535 || cap == 0) return QUBER_E_INVALID;
    if ((long long)H * W >= (1ll << 32) - 1) return QUBER_E_UNSUPPORTED;
    if (!workspace || !aligned16(workspace) || workspace_bytes < quber_rle_encode_workspace_bytes(n, H, W, cap)) return QUBER_E_WORKSPACE;
    cudaStream_t st = (cudaStream_t)stream;
    char *w = static_cast<char *>(workspace);
    uint32_t *colcnt = reinterpret_cast<uint32_t *>(w);
    int32_t *runs = reinterpret_cast<int32_t *>(w + align_up((size_t)n * W * sizeof(uint32_t), 256));
    uint32_t *pos = reinterpret_cast<uint32_t *>(reinterpret_cast<char *>(runs) + align_up((size_t)n * sizeof(int32_t), 256));
    QUBER_CUDA_TRY(cudaMemsetAsync(status, 0, sizeof(int32_t), st));
    const dim3 wgrid(cdiv(W, 256), n);
    rle_walk_kernel<false><<<wgrid, 256, 0, st>>>(masks, n, H, W, colcnt, nullptr, nullptr, 0);
    if (int rc = check_launch("rle_walk_kernel<count>")) return rc;
    rle_colscan_kernel<<<n, 1024, 0, st>>>(colcnt, W, runs);
    if (int rc = check_launch("rle_colscan_kernel")) return rc;
    rle_offsets_kernel<<<1, 1024, 0, st>>>(runs, n, offsets_out, cap, status);
    if (int rc = check_launch("rle_offsets_kernel")) return rc;
    rle_walk_kernel<true><<<wgrid, 256, 0, st>>>(masks, n, H, W, colcnt, offsets_out, pos, cap);
    if (int rc = check_launch("rle_walk_kernel<positions>")) return rc;
    rle_diff_kernel<<<dim3(8, n), 256, 0, st>>>(pos, offsets_out, (unsigned)((long long)H * W), cap, counts_out);
    return check_launch("rle_diff_kernel");
}
