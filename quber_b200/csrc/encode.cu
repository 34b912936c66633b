// K0: initial-mask encode (centre heat-map + normalised offsets).
//
// Reference: maskrefiner/predictor.py:304-357 (MaskRefinerPredictor.predict), Gaussian table
// predictor.py:246-251; twin maskrefiner/data/dataset_mappers/target_generator.py:229-276.
//
// Two passes, both one CTA per 16x128-pixel tile of one image:
//   reduce: reads the K mask planes once (8 pixels / thread / plane, 64-bit streaming loads),
//           accumulates per-mask (count, sum y, sum x) — integers, so the float64 centroid is the
//           same correctly-rounded division numpy makes — and records per pixel the LAST mask that
//           covers it (the reference's overwrite order).
//   emit:   per pixel, offsets from that last mask's centroid in float64 (as numpy: float64
//           arithmetic, float32 store) and the max over the clipped 63x63 Gaussian pastes.
// HBM traffic = K*HW (masks) + HW (last-mask map, written then read) + 12*HW (output).
#include <limits.h>
#include <math.h>

#include "common.cuh"

namespace quber {

constexpr int E_TW = QUBER_OCC_TILE_W, E_TH = QUBER_OCC_TILE_H, E_THREADS = 256;   // 128 x 16 pixel tiles, 8 px per thread
constexpr uint8_t E_NONE = 0xff;

struct EncSums { unsigned long long n, sy, sx; };

// Pass 1 (reduce): streams the K planes of a 128x16 tile once with 64-bit loads (a fetch group of four planes at a time,
// the plane loop unrolled by the group), accumulates per mask (count, sum y, sum x) as integers — warp sums by
// redux.sync, shared-memory atomics per CTA, one global atomic per CTA and mask — and records per pixel the LAST mask
// covering it (the reference's overwrite order, predictor.py:337-356).
// PACKED: `masks` holds one bit per pixel (quber_host_pack_masks / quber_rle_decode: bit i of word w = pixel 32 w + i of
// the flattened [B,K,H,W] array); with `unpacked` the kernel also writes the byte planes (0 / value) — the expansion
// rides on the pass that reads every plane anyway.  quber_mask_boundary_bits does not need them.
// DB (not instantiated in the shipped library: measured equal, 12 more registers): the next group's loads issued before
// the current group is processed.
template <bool PACKED, bool DB>
__global__ void __launch_bounds__(E_THREADS)
encode_reduce_unrolled_kernel(const uint8_t *__restrict__ masks, int K, int H, int W, EncSums *__restrict__ sums,
                              uint8_t *__restrict__ lastk, uint8_t *__restrict__ unpacked, uint32_t value,
                              uint8_t *__restrict__ occ) {
    extern __shared__ unsigned int acc[];   // [K][3]
    const int tid = threadIdx.x, b = blockIdx.z;
    for (int i = tid; i < 3 * K; i += E_THREADS) acc[i] = 0;
    __syncthreads();
    const int y = blockIdx.y * E_TH + (tid >> 4), x0 = blockIdx.x * E_TW + (tid & 15) * 8;
    const bool inb = y < H && x0 < W;
    const size_t HW = (size_t)H * W;
    const size_t pix = (size_t)y * W + x0;
    uint32_t last_lo = 0xffffffffu, last_hi = 0xffffffffu;
    const size_t plane0 = (size_t)b * K * HW + pix;
    const uint8_t *base = masks + (PACKED ? plane0 / 8 : plane0);
    constexpr int PF = 4;
    auto fetch = [&](int k) -> uint2 {
        uint2 v = make_uint2(0, 0);
        if (inb && k < K) {
            if (PACKED) {
                const uint32_t bits = __ldg(base + (size_t)k * (HW / 8));
                v = make_uint2((((bits & 0xfu) * 0x00204081u) & 0x01010101u) * value,
                               (((bits >> 4) * 0x00204081u) & 0x01010101u) * value);
                if (unpacked) st_stream_u2(unpacked + plane0 + (size_t)k * HW, v);
            } else {
                v = ld_stream_u2(base + (size_t)k * HW);
            }
        }
        return v;
    };
    auto process = [&](int k, uint2 v) {
        const bool nz = (v.x | v.y) != 0;
        if (!__any_sync(0xffffffffu, nz)) return;
        unsigned n = 0, sx = 0;
        if (nz) {
            const uint32_t mlo = __vcmpne4(v.x, 0u), mhi = __vcmpne4(v.y, 0u);
            const uint32_t kk = (uint32_t)k * 0x01010101u;
            last_lo = (last_lo & ~mlo) | (kk & mlo);
            last_hi = (last_hi & ~mhi) | (kk & mhi);
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const unsigned a = (mlo >> (8 * j)) & 1u, c = (mhi >> (8 * j)) & 1u;
                n += a + c;
                sx += a * (unsigned)(x0 + j) + c * (unsigned)(x0 + 4 + j);
            }
        }
        unsigned sy = n * (unsigned)y;
        n = __reduce_add_sync(0xffffffffu, n);
        sy = __reduce_add_sync(0xffffffffu, sy);
        sx = __reduce_add_sync(0xffffffffu, sx);
        if ((tid & 31) == 0) {
            atomicAdd(&acc[3 * k], n); atomicAdd(&acc[3 * k + 1], sy); atomicAdd(&acc[3 * k + 2], sx);
        }
    };
    uint2 nxt[PF];
    if (DB) {
#pragma unroll
        for (int j = 0; j < PF; ++j) nxt[j] = fetch(j);
    }
    for (int k0 = 0; k0 < K; k0 += PF) {
        uint2 cur[PF];
#pragma unroll
        for (int j = 0; j < PF; ++j) cur[j] = DB ? nxt[j] : fetch(k0 + j);
        if (DB) {
#pragma unroll
            for (int j = 0; j < PF; ++j) nxt[j] = fetch(k0 + PF + j);
        }
#pragma unroll
        for (int j = 0; j < PF; ++j) process(k0 + j, cur[j]);    // planes past K were fetched as zeros: skipped by the vote
    }
    if (inb) st_stream_u2(lastk + (size_t)b * HW + pix, make_uint2(last_lo, last_hi));
    __syncthreads();
    for (int k = tid; k < K; k += E_THREADS) {
        // occupancy [B][K][tile rows][tile columns]: does this tile hold a set pixel of mask k — the boundary pass skips
        // the planes that have none near its tile (csrc/boundary.cu)
        if (occ) occ[(((size_t)b * K + k) * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x] = acc[3 * k] != 0;
        if (acc[3 * k]) {
            EncSums *s = sums + (size_t)b * K + k;
            atomicAdd(&s->n, (unsigned long long)acc[3 * k]);
            atomicAdd(&s->sy, (unsigned long long)acc[3 * k + 1]);
            atomicAdd(&s->sx, (unsigned long long)acc[3 * k + 2]);
        }
    }
}

struct EncCentre { double cy, cx; int uly, ulx; };

// EARLY: the last-mask bytes of this thread's pixels are requested before the per-CTA centre set-up, so that their
// DRAM latency runs under the set-up (sums load, float64 divisions, barrier) instead of after the paste loop; MINB =
// resident CTAs per SM asked of the compiler (QUBER_ENCODE_EMIT=0|1|2 selects old / early / early + 5 CTAs for A/B runs).
template <bool EARLY, int MINB>
__global__ void __launch_bounds__(E_THREADS, MINB)
encode_emit_kernel(const EncSums *__restrict__ sums, const uint8_t *__restrict__ lastk, int K, int H, int W, int norm_w, int sigma,
                   const float *__restrict__ gauss, float *__restrict__ out) {
    extern __shared__ __align__(16) unsigned char smraw[];
    EncCentre *ctr = reinterpret_cast<EncCentre *>(smraw);            // [K]
    int *list = reinterpret_cast<int *>(ctr + K);                     // [K] masks whose paste touches this tile
    __shared__ int nlist;
    const int tid = threadIdx.x, b = blockIdx.z;
    const int gs = 6 * sigma + 3, half = 3 * sigma + 1;
    const int ty0 = blockIdx.y * E_TH, tx0 = blockIdx.x * E_TW;
    const int y = ty0 + (tid >> 4), x0 = tx0 + (tid & 15) * 8;
    const size_t HW = (size_t)H * W, pix = (size_t)y * W + x0;
    const bool inside = y < H && x0 < W;
    uint2 lk = make_uint2(0u, 0u);
    if (EARLY && inside) lk = ld_stream_u2(lastk + (size_t)b * HW + pix);
    if (tid == 0) nlist = 0;
    __syncthreads();
    for (int k = tid; k < K; k += E_THREADS) {
        const EncSums s = sums[(size_t)b * K + k];
        EncCentre c;
        c.cy = c.cx = 0.0; c.uly = c.ulx = INT_MIN / 2;
        if (s.n) {
            c.cy = (double)s.sy / (double)s.n;                       // np.mean of exact integer sums
            c.cx = (double)s.sx / (double)s.n;
            c.uly = __double2int_rn(c.cy) - half;                    // round(): half to even (predictor.py:324-327)
            c.ulx = __double2int_rn(c.cx) - half;
            if (c.uly < ty0 + E_TH && c.uly + gs > ty0 && c.ulx < tx0 + E_TW && c.ulx + gs > tx0)
                list[atomicAdd(&nlist, 1)] = k;
        }
        ctr[k] = c;
    }
    __syncthreads();
    if (!inside) return;
    float cen[8], oy[8], ox[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) cen[j] = 0.f;
    const int n = nlist;
    for (int t = 0; t < n; ++t) {                                     // max is order-independent
        const EncCentre c = ctr[list[t]];
        const int gy = y - c.uly;
        if ((unsigned)gy >= (unsigned)gs) continue;
        const float *grow = gauss + gy * gs;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const int gx = x0 + j - c.ulx;
            if ((unsigned)gx < (unsigned)gs) cen[j] = fmaxf(cen[j], __ldg(grow + gx));
        }
    }
    if (!EARLY) lk = *reinterpret_cast<const uint2 *>(lastk + (size_t)b * HW + pix);
    const double dH = (double)H, dW = (double)norm_w;                 // W is the row pitch, norm_w the frame's width
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        const uint32_t k = ((j < 4 ? lk.x : lk.y) >> (8 * (j & 3))) & 0xffu;
        if (k == E_NONE) { oy[j] = 0.f; ox[j] = 0.f; continue; }
        const EncCentre c = ctr[k];
        oy[j] = (float)((c.cy - (double)y) / dH);                    // predictor.py:345-346
        ox[j] = (float)((c.cx - (double)(x0 + j)) / dW);
    }
    float *o = out + (size_t)b * 3 * HW + pix;
    st_stream_f4(o, make_float4(cen[0], cen[1], cen[2], cen[3]));
    st_stream_f4(o + 4, make_float4(cen[4], cen[5], cen[6], cen[7]));
    st_stream_f4(o + HW, make_float4(oy[0], oy[1], oy[2], oy[3]));
    st_stream_f4(o + HW + 4, make_float4(oy[4], oy[5], oy[6], oy[7]));
    st_stream_f4(o + 2 * HW, make_float4(ox[0], ox[1], ox[2], ox[3]));
    st_stream_f4(o + 2 * HW + 4, make_float4(ox[4], ox[5], ox[6], ox[7]));
}

}  // namespace quber

extern "C" void quber_gaussian_table(int sigma, float *host_table) {
    // predictor.py:246-251 (float64 exp, observed through a float32 heat-map)
    const int size = 6 * sigma + 3;
    const double c0 = 3 * sigma + 1;
    for (int y = 0; y < size; ++y)
        for (int x = 0; x < size; ++x) {
            const double dx = (double)x - c0, dy = (double)y - c0;
            host_table[y * size + x] = (float)exp(-(dx * dx + dy * dy) / (double)(2 * sigma * sigma));
        }
}

extern "C" size_t quber_occupancy_bytes(int B, int K, int H, int W) {
    if (B <= 0 || K <= 0 || H <= 0 || W <= 0) return 0;
    return (size_t)B * K * quber::cdiv(H, QUBER_OCC_TILE_H) * quber::cdiv(W, QUBER_OCC_TILE_W);
}

extern "C" size_t quber_encode_workspace_bytes(int B, int K, int H, int W) {
    if (B <= 0 || K <= 0 || H <= 0 || W <= 0) return 0;
    return quber::align_up((size_t)B * K * sizeof(quber::EncSums), 256) + quber::align_up((size_t)B * H * W, 256);
}

static int encode_launch(const uint8_t *masks, bool packed, int value, uint8_t *unpacked, int B, int K, int H, int W, int valid_w, int sigma,
                         const float *gauss, float *out, uint8_t *occupancy, void *workspace, size_t workspace_bytes, void *stream) {
    using namespace quber;
    if (!masks || !gauss || !out || B <= 0 || K <= 0 || K > 254 || H <= 0 || W <= 0 || sigma < 1) return QUBER_E_INVALID;
    if (valid_w < 0 || valid_w > W) return QUBER_E_INVALID;
    if (W % 16 != 0 || !aligned16(out) || (reinterpret_cast<uintptr_t>(masks) & (packed ? 3u : 15u))) return QUBER_E_ALIGN;
    if (packed && (value < 1 || value > 255)) return QUBER_E_INVALID;
    if (packed && unpacked && !aligned16(unpacked)) return QUBER_E_ALIGN;
    if (!workspace || workspace_bytes < quber_encode_workspace_bytes(B, K, H, W) || !aligned16(workspace)) return QUBER_E_WORKSPACE;
    cudaStream_t st = (cudaStream_t)stream;
    EncSums *sums = reinterpret_cast<EncSums *>(workspace);
    const size_t sums_bytes = align_up((size_t)B * K * sizeof(EncSums), 256);
    uint8_t *lastk = reinterpret_cast<uint8_t *>(workspace) + sums_bytes;
    QUBER_CUDA_TRY(cudaMemsetAsync(sums, 0, sums_bytes, st));
    dim3 grid(cdiv(W, E_TW), cdiv(H, E_TH), B);
    const size_t rsm = 3 * K * sizeof(unsigned);
    const uint32_t val = packed ? (uint32_t)value : 0u;
    uint8_t *unp = packed ? unpacked : nullptr;
#define QUBER_REDUCE_LAUNCH(KERNEL) KERNEL<<<grid, E_THREADS, rsm, st>>>(masks, K, H, W, sums, lastk, unp, val, occupancy)
    if (packed) QUBER_REDUCE_LAUNCH((encode_reduce_unrolled_kernel<true, false>));
    else QUBER_REDUCE_LAUNCH((encode_reduce_unrolled_kernel<false, false>));
#undef QUBER_REDUCE_LAUNCH
    if (int rc = check_launch("encode_reduce_kernel")) return rc;
    const size_t esm = K * (sizeof(EncCentre) + sizeof(int));
    encode_emit_kernel<true, 4><<<grid, E_THREADS, esm, st>>>(sums, lastk, K, H, W, valid_w > 0 ? valid_w : W, sigma, gauss, out);
    return check_launch("encode_emit_kernel");
}

extern "C" int quber_encode_initial_masks(const uint8_t *masks, int B, int K, int H, int W, int valid_w, int sigma, const float *gauss,
                                          float *out, uint8_t *occupancy, void *workspace, size_t workspace_bytes, void *stream) {
    return encode_launch(masks, false, 0, nullptr, B, K, H, W, valid_w, sigma, gauss, out, occupancy, workspace, workspace_bytes, stream);
}

extern "C" int quber_encode_initial_masks_packed(const uint32_t *packed, int value, uint8_t *masks_out, int B, int K, int H, int W,
                                                 int valid_w, int sigma, const float *gauss, float *out, uint8_t *occupancy,
                                                 void *workspace, size_t workspace_bytes, void *stream) {
    return encode_launch(reinterpret_cast<const uint8_t *>(packed), true, value, masks_out, B, K, H, W, valid_w, sigma, gauss, out,
                         occupancy, workspace, workspace_bytes, stream);
}

// ------------------------------------------------------------------------------------------------
// quber_encode_initial_masks without any allocation anywhere: the scratch (per-mask sums + the last-mask map) lives in a
// file-scope device array instead of a caller-provided workspace.  No counterpart in the reference; added
// for callers that cannot allocate (CUDA graph capture, serving loops, harnesses that forbid
// cudaMalloc).  Every GPU that loads the library holds QUBER_ENCODE_STATIC_SCRATCH_BYTES until the process exits.  All
// calls on a device share the array: calls that overlap on the GPU (different streams) race — serialise them.  The array
// keeps the previous call's contents; the sums, the only part that must start at zero, are cleared by every call (the
// same cudaMemsetAsync as the workspace path), the last-mask map is fully rewritten for the tiles the call covers.
// ------------------------------------------------------------------------------------------------
namespace quber {
static __device__ __align__(256) unsigned char g_encode_scratch[QUBER_ENCODE_STATIC_SCRATCH_BYTES];

// In host code `g_encode_scratch` is only a handle for the symbol, not a device address: the address is asked of the
// runtime once per device, with that device current.
static int encode_scratch_address(void **ptr) {
    constexpr int MAX_DEV = 64;
    static void *addr[MAX_DEV] = {};
    int dev = -1;
    QUBER_CUDA_TRY(cudaGetDevice(&dev));
    if (dev < 0 || dev >= MAX_DEV) return QUBER_E_UNSUPPORTED;
    void *p = __atomic_load_n(&addr[dev], __ATOMIC_ACQUIRE);
    if (!p) {
        QUBER_CUDA_TRY(cudaGetSymbolAddress(&p, g_encode_scratch));     // the same answer from every thread: a benign race
        if (!p) return QUBER_E_CUDA;
        __atomic_store_n(&addr[dev], p, __ATOMIC_RELEASE);
    }
    *ptr = p;
    return QUBER_OK;
}
}  // namespace quber

extern "C" size_t quber_encode_static_scratch_bytes(void) { return QUBER_ENCODE_STATIC_SCRATCH_BYTES; }

extern "C" int quber_encode_initial_masks_static(const uint8_t *masks, int B, int K, int H, int W, int valid_w, int sigma,
                                                 const float *gauss, float *out, uint8_t *occupancy, void *stream) {
    using namespace quber;
    if (B <= 0 || K <= 0 || H <= 0 || W <= 0) return QUBER_E_INVALID;
    const size_t need = quber_encode_workspace_bytes(B, K, H, W);
    if (need > (size_t)QUBER_ENCODE_STATIC_SCRATCH_BYTES) {             // checked before any CUDA call
        char msg[200];
        snprintf(msg, sizeof(msg), "quber_encode_initial_masks_static: B=%d K=%d H=%d W=%d needs %zu bytes of scratch, the static array holds %zu",
                 B, K, H, W, need, (size_t)QUBER_ENCODE_STATIC_SCRATCH_BYTES);
        set_last_cuda_error(cudaErrorMemoryAllocation, msg);
        return QUBER_E_WORKSPACE;
    }
    void *scratch = nullptr;
    if (int rc = encode_scratch_address(&scratch)) return rc;
    return encode_launch(masks, false, 0, nullptr, B, K, H, W, valid_w, sigma, gauss, out, occupancy, scratch,
                         QUBER_ENCODE_STATIC_SCRATCH_BYTES, stream);
}
