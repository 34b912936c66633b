// GroupNorm (+ ReLU), inference forward, for the network side of the path.
//
// Reference: every `norm="GN"` convolution of the decoder / heads is detectron2 Conv2d -> nn.GroupNorm(32, C) ->
// F.relu (maskrefiner/modeling/mask_refiner/model.py:386-403, 431-451 via get_norm), and the RGB-D fusion blocks are
// nn.Conv2d -> nn.GroupNorm(32, C) -> nn.ReLU (maskrefiner/modeling/backbone/resnet.py:466-479).  PyTorch runs that as
// RowwiseMoments + an elementwise apply + a separate clamp_min (plus fp32 <-> bf16 casts under autocast): 3 reads and
// 2 writes of the activation.  Here: one read for the statistics, one read + one write for normalise + affine + ReLU,
// for NCHW and channels-last (NHWC) tensors, fp32 or bf16 storage, fp32 arithmetic.
//
// Deterministic: per-tile partial sums are combined in a fixed order (no floating-point atomics).  Sums are taken
// relative to a per-(sample, group) pivot (the group's first element), so var = (S2 - S1^2/n)/n does not cancel when
// |mean| >> std; the final combine over tiles is in double.  HBM-bound: 2 reads + 1 write of the tensor.
#include <cuda_bf16.h>

#include "common.cuh"

namespace quber {

constexpr int GN_THREADS = 256;
// NHWC: pixels per CTA, chosen per call (gn_tile_pix): large enough to amortise the per-CTA prologue (pivot loads) and the
// serial fixed-order combine, small enough for >= ~4 CTAs per SM
constexpr int GN_TILE_ELEMS = 16384;   // NCHW: elements of one (sample, group) slice per CTA

template <typename T> struct Vec;      // one 16-byte access
template <> struct Vec<float> {
    static constexpr int N = 4;
    static __device__ __forceinline__ void load(const float *p, float *v) {
        const float4 r = *reinterpret_cast<const float4 *>(p);
        v[0] = r.x; v[1] = r.y; v[2] = r.z; v[3] = r.w;
    }
    static __device__ __forceinline__ void store(float *p, const float *v) {
        *reinterpret_cast<float4 *>(p) = make_float4(v[0], v[1], v[2], v[3]);
    }
    static __device__ __forceinline__ float one(const float *p) { return *p; }
};
template <> struct Vec<__nv_bfloat16> {
    static constexpr int N = 8;
    static __device__ __forceinline__ void load(const __nv_bfloat16 *p, float *v) {
        const uint4 r = *reinterpret_cast<const uint4 *>(p);
        const uint32_t w[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            v[2 * i] = __uint_as_float(w[i] << 16);
            v[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u);
        }
    }
    static __device__ __forceinline__ void store(__nv_bfloat16 *p, const float *v) {
        uint32_t w[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const __nv_bfloat162 h = __floats2bfloat162_rn(v[2 * i], v[2 * i + 1]);
            w[i] = *reinterpret_cast<const uint32_t *>(&h);
        }
        *reinterpret_cast<uint4 *>(p) = make_uint4(w[0], w[1], w[2], w[3]);
    }
    static __device__ __forceinline__ float one(const __nv_bfloat16 *p) { return __bfloat162float(*p); }
};

// fixed-order block reduction of (a, b); result valid in thread 0
__device__ __forceinline__ void block_sum2(float &a, float &b, float *sm /* [2 * GN_THREADS / 32] */) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        a += __shfl_down_sync(0xffffffffu, a, o);
        b += __shfl_down_sync(0xffffffffu, b, o);
    }
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
    if (l == 0) { sm[2 * w] = a; sm[2 * w + 1] = b; }
    __syncthreads();
    if (threadIdx.x == 0) {
        a = 0.f; b = 0.f;
        for (int i = 0; i < GN_THREADS / 32; ++i) { a += sm[2 * i]; b += sm[2 * i + 1]; }
    }
}

// ------------------------------------------------------------------------------------------------ NCHW
// grid (T, N*G): partial (S1, S2) of tile t of the contiguous slice of (n, g), relative to the slice's first element
template <typename T>
__global__ void __launch_bounds__(GN_THREADS)
gn_stats_nchw_kernel(const T *__restrict__ x, const float *__restrict__ pb, size_t slice, int HW, int cg, int G, int tiles,
                     float *__restrict__ part) {
    constexpr int V = Vec<T>::N;
    __shared__ float sm[2 * GN_THREADS / 32];
    const size_t ng = blockIdx.y;
    const int c0 = (int)(ng % G) * cg;
    const T *p = x + ng * slice;
    const float K = Vec<T>::one(p) + (pb ? __ldg(pb + c0) : 0.f);
    const size_t lo = (size_t)blockIdx.x * GN_TILE_ELEMS, hi = min(slice, lo + (size_t)GN_TILE_ELEMS);
    float s1 = 0.f, s2 = 0.f;
#pragma unroll 4
    for (size_t i = lo + (size_t)threadIdx.x * V; i < hi; i += (size_t)GN_THREADS * V) {     // slice % V == 0
        float v[V];
        Vec<T>::load(p + i, v);
        const float kb = K - (pb ? __ldg(pb + c0 + (int)(i / HW)) : 0.f);
#pragma unroll
        for (int j = 0; j < V; ++j) { const float d = v[j] - kb; s1 += d; s2 = fmaf(d, d, s2); }
    }
    block_sum2(s1, s2, sm);
    if (threadIdx.x == 0) {
        part[(ng * tiles + blockIdx.x) * 2] = s1;
        part[(ng * tiles + blockIdx.x) * 2 + 1] = s2;
    }
}

template <typename T>
__global__ void __launch_bounds__(GN_THREADS)
gn_apply_nchw_kernel(const T *__restrict__ x, T *__restrict__ y, const float *__restrict__ pb, const float *__restrict__ gamma,
                     const float *__restrict__ beta, const float *__restrict__ stats /* [N*G][2] mean, rstd */, size_t slice, int HW,
                     int cg, int G, int relu) {
    constexpr int V = Vec<T>::N;
    const size_t ng = blockIdx.y;
    const int g = (int)(ng % G);
    const float mean = stats[2 * ng], rstd = stats[2 * ng + 1];
    const T *p = x + ng * slice;
    T *q = y + ng * slice;
    const size_t lo = (size_t)blockIdx.x * GN_TILE_ELEMS, hi = min(slice, lo + (size_t)GN_TILE_ELEMS);
#pragma unroll 4
    for (size_t i = lo + (size_t)threadIdx.x * V; i < hi; i += (size_t)GN_THREADS * V) {     // HW % V == 0: one channel per vector
        const int c = g * cg + (int)(i / HW);
        const float a = rstd * __ldg(gamma + c), b = fmaf((pb ? __ldg(pb + c) : 0.f) - mean, a, __ldg(beta + c));
        float v[V];
        Vec<T>::load(p + i, v);
#pragma unroll
        for (int j = 0; j < V; ++j) {
            v[j] = fmaf(v[j], a, b);
            if (relu) v[j] = fmaxf(v[j], 0.f);
        }
        Vec<T>::store(q + i, v);
    }
}

// ------------------------------------------------------------------------------------------------ NHWC
// grid (T, N): CTA = tile_pix pixels x C channels.  Thread t owns the channel vectors {t, t + 256, ...} (< C / V)
// when C / V >= 256, else vector t % (C / V) on pixel rows t / (C / V), t / (C / V) + rows, ...
template <typename T>
__global__ void __launch_bounds__(GN_THREADS)
gn_stats_nhwc_kernel(const T *__restrict__ x, const float *__restrict__ pb, int HW, int C, int cg, int tiles, int tile_pix,
                     float *__restrict__ part /* [N][G][tiles][2] */) {
    constexpr int V = Vec<T>::N;
    extern __shared__ float acc[];                       // [rows][C][2]
    const int n = blockIdx.y, t = blockIdx.x, tid = threadIdx.x;
    const int TP = C / V;                                // vectors per pixel
    const int rows = TP >= GN_THREADS ? 1 : GN_THREADS / TP;
    const int row = TP >= GN_THREADS ? 0 : tid / TP;
    const T *xs = x + (size_t)n * HW * C;
    const int p0 = t * tile_pix, p1 = min(HW, p0 + tile_pix);
    const int G = C / cg;
    for (int ch = (TP >= GN_THREADS ? tid : tid % TP); ch < TP; ch += GN_THREADS) {
        if (TP < GN_THREADS && tid >= rows * TP) break;  // 256 % TP != 0: surplus threads idle
        float K[V], s1[V], s2[V];
#pragma unroll
        for (int j = 0; j < V; ++j) {
            const int c = ch * V + j, cf = (c / cg) * cg;
            // pivot: the group's first element of the sample (bias included); the lane's own bias is folded into it
            K[j] = Vec<T>::one(xs + cf) + (pb ? __ldg(pb + cf) - __ldg(pb + c) : 0.f);
            s1[j] = 0.f; s2[j] = 0.f;
        }
#pragma unroll 4
        for (int p = p0 + row; p < p1; p += rows) {
            float v[V];
            Vec<T>::load(xs + (size_t)p * C + ch * V, v);
#pragma unroll
            for (int j = 0; j < V; ++j) { const float d = v[j] - K[j]; s1[j] += d; s2[j] = fmaf(d, d, s2[j]); }
        }
#pragma unroll
        for (int j = 0; j < V; ++j) {
            acc[((size_t)row * C + ch * V + j) * 2] = s1[j];
            acc[((size_t)row * C + ch * V + j) * 2 + 1] = s2[j];
        }
        if (TP < GN_THREADS) break;
    }
    __syncthreads();
    const int used_rows = TP >= GN_THREADS ? 1 : rows;
    for (int g = tid; g < G; g += GN_THREADS) {          // fixed-order combine: rows, then the group's channels
        float a = 0.f, b = 0.f;
        for (int r = 0; r < used_rows; ++r)
            for (int c = g * cg; c < (g + 1) * cg; ++c) { a += acc[((size_t)r * C + c) * 2]; b += acc[((size_t)r * C + c) * 2 + 1]; }
        float *o = part + (((size_t)n * G + g) * tiles + t) * 2;
        o[0] = a; o[1] = b;
    }
}

template <typename T>
__global__ void __launch_bounds__(GN_THREADS)
gn_apply_nhwc_kernel(const T *__restrict__ x, T *__restrict__ y, const float *__restrict__ pb, const float *__restrict__ gamma,
                     const float *__restrict__ beta, const float *__restrict__ stats, int HW, int C, int cg, int relu, int tile_pix) {
    constexpr int V = Vec<T>::N;
    const int n = blockIdx.y, t = blockIdx.x, tid = threadIdx.x;
    const int TP = C / V;
    const int rows = TP >= GN_THREADS ? 1 : GN_THREADS / TP;
    const int row = TP >= GN_THREADS ? 0 : tid / TP;
    const int G = C / cg;
    const size_t base = (size_t)n * HW * C;
    const int p0 = t * tile_pix, p1 = min(HW, p0 + tile_pix);
    for (int ch = (TP >= GN_THREADS ? tid : tid % TP); ch < TP; ch += GN_THREADS) {
        if (TP < GN_THREADS && tid >= rows * TP) break;
        float a[V], b[V];
#pragma unroll
        for (int j = 0; j < V; ++j) {
            const int c = ch * V + j, g = c / cg;
            const float mean = stats[((size_t)n * G + g) * 2], rstd = stats[((size_t)n * G + g) * 2 + 1];
            a[j] = rstd * __ldg(gamma + c);
            b[j] = fmaf((pb ? __ldg(pb + c) : 0.f) - mean, a[j], __ldg(beta + c));
        }
#pragma unroll 4
        for (int p = p0 + row; p < p1; p += rows) {
            float v[V];
            Vec<T>::load(x + base + (size_t)p * C + ch * V, v);
#pragma unroll
            for (int j = 0; j < V; ++j) {
                v[j] = fmaf(v[j], a[j], b[j]);
                if (relu) v[j] = fmaxf(v[j], 0.f);
            }
            Vec<T>::store(y + base + (size_t)p * C + ch * V, v);
        }
        if (TP < GN_THREADS) break;
    }
}

// one thread per (n, g): combine the tile partials in double, in tile order
template <typename T>
__global__ void gn_finalize_kernel(const T *__restrict__ x, const float *__restrict__ pb, int cg, const float *__restrict__ part, int NG,
                                   int G, int tiles, double count, size_t pivot_stride_n, size_t pivot_stride_g, float eps,
                                   float *__restrict__ stats) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= NG) return;
    const int n = i / G, g = i % G;
    const double K = (double)(Vec<T>::one(x + (size_t)n * pivot_stride_n + (size_t)g * pivot_stride_g) + (pb ? pb[g * cg] : 0.f));
    double s1 = 0.0, s2 = 0.0;
    for (int t = 0; t < tiles; ++t) { s1 += (double)part[((size_t)i * tiles + t) * 2]; s2 += (double)part[((size_t)i * tiles + t) * 2 + 1]; }
    const double m = s1 / count;
    double var = s2 / count - m * m;                    // biased variance, as torch.nn.GroupNorm
    if (var < 0.0) var = 0.0;
    stats[2 * i] = (float)(K + m);
    stats[2 * i + 1] = (float)(1.0 / sqrt(var + (double)eps));
}

static int gn_tile_pix(int N, int HW) {
    long long t = ((long long)HW * N / 592 + 63) / 64 * 64;          // ~4 CTAs per SM over the launch
    return (int)(t < 64 ? 64 : t > 512 ? 512 : t);
}

static int gn_tiles(int N, int C, int HW, int G, int layout) {
    if (layout == 0) return (int)(((size_t)(C / G) * HW + GN_TILE_ELEMS - 1) / GN_TILE_ELEMS);
    const int tp = gn_tile_pix(N, HW);
    return (HW + tp - 1) / tp;
}

template <typename T>
static int gn_run(const void *xv, void *yv, const float *pb, const float *gamma, const float *beta, int N, int C, int HW, int G, float eps, int relu,
                  int layout, float *ws, cudaStream_t st) {
    constexpr int V = Vec<T>::N;
    const T *x = static_cast<const T *>(xv);
    T *y = static_cast<T *>(yv);
    const int cg = C / G, tiles = gn_tiles(N, C, HW, G, layout), tile_pix = gn_tile_pix(N, HW);
    float *part = ws, *stats = ws + (size_t)N * G * tiles * 2;
    const double count = (double)cg * HW;
    if (layout == 0) {
        if (HW % V != 0) return QUBER_E_UNSUPPORTED;
        const size_t slice = (size_t)cg * HW;
        dim3 grid(tiles, N * G);
        gn_stats_nchw_kernel<T><<<grid, GN_THREADS, 0, st>>>(x, pb, slice, HW, cg, G, tiles, part);
        if (int rc = check_launch("gn_stats_nchw_kernel")) return rc;
        gn_finalize_kernel<T><<<cdiv(N * G, 128), 128, 0, st>>>(x, pb, cg, part, N * G, G, tiles, count, (size_t)C * HW, slice, eps, stats);
        if (int rc = check_launch("gn_finalize_kernel")) return rc;
        gn_apply_nchw_kernel<T><<<grid, GN_THREADS, 0, st>>>(x, y, pb, gamma, beta, stats, slice, HW, cg, G, relu);
        return check_launch("gn_apply_nchw_kernel");
    }
    if (C % V != 0) return QUBER_E_UNSUPPORTED;
    const int TP = C / V;
    if (TP > GN_THREADS && TP % GN_THREADS != 0) return QUBER_E_UNSUPPORTED;
    const int rows = TP >= GN_THREADS ? 1 : GN_THREADS / TP;
    const size_t smem = (size_t)rows * C * 2 * sizeof(float);
    if (smem > 48 * 1024) {
        QUBER_CUDA_TRY(cudaFuncSetAttribute(gn_stats_nhwc_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    }
    dim3 grid(tiles, N);
    gn_stats_nhwc_kernel<T><<<grid, GN_THREADS, smem, st>>>(x, pb, HW, C, cg, tiles, tile_pix, part);
    if (int rc = check_launch("gn_stats_nhwc_kernel")) return rc;
    gn_finalize_kernel<T><<<cdiv(N * G, 128), 128, 0, st>>>(x, pb, cg, part, N * G, G, tiles, count, (size_t)C * HW, (size_t)cg, eps, stats);
    if (int rc = check_launch("gn_finalize_kernel")) return rc;
    gn_apply_nhwc_kernel<T><<<grid, GN_THREADS, 0, st>>>(x, y, pb, gamma, beta, stats, HW, C, cg, relu, tile_pix);
    return check_launch("gn_apply_nhwc_kernel");
}

// ------------------------------------------------------------------------------------------------ upsample + concat
// DeepLabV3+ decoder step (detectron2 DeepLabV3PlusHead.layers, called from model.py:713):
//     y = F.interpolate(y, size=proj_x.shape[-2:], mode="bilinear", align_corners=False);  cat([proj_x, y], dim=1)
// as ONE pass over channels-last tensors: out[n,y,x,:] = [ a[n,y,x,:Ca] | bilinear(b)[n,y,x,:Cb] ].  PyTorch's source
// index and lambda arithmetic (area_pixel_compute_source_index, fp32 accumulation for bf16).
template <typename T>
__global__ void __launch_bounds__(256)
upsample_concat_kernel(const T *__restrict__ a, const T *__restrict__ b, T *__restrict__ out, size_t total, int Ca, int Cb, int H,
                       int W, int h, int w, float rh, float rw) {
    constexpr int V = Vec<T>::N;
    const int TPa = Ca / V, TPo = (Ca + Cb) / V;
    const size_t idx = (size_t)blockIdx.x * 256 + threadIdx.x;
    if (idx >= total) return;
    const int vec = (int)(idx % TPo);
    const size_t pix = idx / TPo;                        // n * H * W + y * W + x
    T *o = out + pix * (Ca + Cb) + (size_t)vec * V;
    if (vec < TPa) {
        *reinterpret_cast<uint4 *>(o) = *reinterpret_cast<const uint4 *>(a + pix * Ca + (size_t)vec * V);
        return;
    }
    const int x = (int)(pix % W), y = (int)((pix / W) % H);
    const size_t n = pix / ((size_t)W * H);
    float sy = rh * ((float)y + 0.5f) - 0.5f, sx = rw * ((float)x + 0.5f) - 0.5f;
    sy = sy < 0.f ? 0.f : sy;
    sx = sx < 0.f ? 0.f : sx;
    const int y1 = min((int)sy, h - 1), x1 = min((int)sx, w - 1);
    const int yp = y1 < h - 1 ? 1 : 0, xp = x1 < w - 1 ? 1 : 0;
    const float ly = sy - (float)y1, lx = sx - (float)x1, hy = 1.f - ly, hx = 1.f - lx;
    const T *s = b + ((n * h + y1) * (size_t)w + x1) * Cb + (size_t)(vec - TPa) * V;
    float v00[V], v01[V], v10[V], v11[V], r[V];
    Vec<T>::load(s, v00);
    Vec<T>::load(s + (size_t)xp * Cb, v01);
    Vec<T>::load(s + (size_t)yp * w * Cb, v10);
    Vec<T>::load(s + ((size_t)yp * w + xp) * Cb, v11);
#pragma unroll
    for (int j = 0; j < V; ++j) r[j] = hy * (hx * v00[j] + lx * v01[j]) + ly * (hx * v10[j] + lx * v11[j]);
    Vec<T>::store(o, r);
}

// ------------------------------------------------------------------------------------------------ max-pool 3x3 / 2
// F.max_pool2d(x, kernel_size=3, stride=2, padding=1) of the ResNet stems (maskrefiner/modeling/backbone/resnet.py:76,125)
// on channels-last tensors: one thread per (output pixel, 16-byte channel vector), nine vector loads (L1-resident
// neighbours), padding = -inf as PyTorch.
template <typename T>
__global__ void __launch_bounds__(256)
maxpool3x3s2_nhwc_kernel(const T *__restrict__ x, T *__restrict__ y, size_t total, int C, int H, int W, int Ho, int Wo) {
    constexpr int V = Vec<T>::N;
    const int TP = C / V;
    const size_t idx = (size_t)blockIdx.x * 256 + threadIdx.x;
    if (idx >= total) return;
    const int vec = (int)(idx % TP);
    const size_t pix = idx / TP;
    const int ox = (int)(pix % Wo), oy = (int)((pix / Wo) % Ho);
    const size_t n = pix / ((size_t)Wo * Ho);
    float m[V];
#pragma unroll
    for (int j = 0; j < V; ++j) m[j] = -INFINITY;
#pragma unroll
    for (int dy = -1; dy <= 1; ++dy) {
        const int iy = 2 * oy + dy;
        if (iy < 0 || iy >= H) continue;
#pragma unroll
        for (int dx = -1; dx <= 1; ++dx) {
            const int ix = 2 * ox + dx;
            if (ix < 0 || ix >= W) continue;
            float v[V];
            Vec<T>::load(x + ((n * H + iy) * (size_t)W + ix) * C + (size_t)vec * V, v);
#pragma unroll
            for (int j = 0; j < V; ++j) m[j] = fmaxf(m[j], v[j]);        // NaN-free activations (post-ReLU)
        }
    }
    Vec<T>::store(y + pix * C + (size_t)vec * V, m);
}

// ------------------------------------------------------------------------------------------------ residual + LayerNorm
// MSDeformAttnTransformerEncoderLayer.forward (msdeformattn.py:122-131): src = norm1(src + dropout(src2)) and
// forward_ffn's src = norm2(src + src2) — torch runs an add and a LayerNorm pass (0.09 + 0.35 ms for 32 x 6300 x 256);
// here one pass: one warp per row, the row held in registers, mean then centred variance (two-pass, as accurate as
// torch's), fixed-order shuffle reductions.  Optional second output in bf16 (the next GEMM's A operand).
template <int VPL>                                      // float4 vectors per lane: C = 128 * VPL
__global__ void __launch_bounds__(256)
add_layer_norm_kernel(const float *__restrict__ a, const float *__restrict__ b, const float *__restrict__ gamma,
                      const float *__restrict__ beta, size_t rows, float eps, float *__restrict__ y, __nv_bfloat16 *__restrict__ y16,
                      const float *__restrict__ pos, void *__restrict__ qout, int q_bf16) {
    constexpr int C = 128 * VPL;
    const size_t row = (size_t)blockIdx.x * 8 + (threadIdx.x >> 5);
    if (row >= rows) return;
    const int lane = threadIdx.x & 31;
    float v[4 * VPL];
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < VPL; ++i) {
        const size_t o = row * C + (size_t)(i * 32 + lane) * 4;
        const float4 x = *reinterpret_cast<const float4 *>(a + o);
        float4 r = make_float4(0.f, 0.f, 0.f, 0.f);
        if (b) r = *reinterpret_cast<const float4 *>(b + o);
        v[4 * i] = x.x + r.x; v[4 * i + 1] = x.y + r.y; v[4 * i + 2] = x.z + r.z; v[4 * i + 3] = x.w + r.w;
        s += (v[4 * i] + v[4 * i + 1]) + (v[4 * i + 2] + v[4 * i + 3]);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    const float mean = s * (1.0f / C);
    float q = 0.f;
#pragma unroll
    for (int i = 0; i < 4 * VPL; ++i) { const float d = v[i] - mean; q = fmaf(d, d, q); }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) q += __shfl_xor_sync(0xffffffffu, q, o);
    const float rstd = rsqrtf(q * (1.0f / C) + eps);
#pragma unroll
    for (int i = 0; i < VPL; ++i) {
        const int c = (i * 32 + lane) * 4;
        const float4 g = *reinterpret_cast<const float4 *>(gamma + c), bb = *reinterpret_cast<const float4 *>(beta + c);
        float4 r;
        r.x = fmaf((v[4 * i] - mean) * rstd, g.x, bb.x);
        r.y = fmaf((v[4 * i + 1] - mean) * rstd, g.y, bb.y);
        r.z = fmaf((v[4 * i + 2] - mean) * rstd, g.z, bb.z);
        r.w = fmaf((v[4 * i + 3] - mean) * rstd, g.w, bb.w);
        *reinterpret_cast<float4 *>(y + row * C + c) = r;
        if (y16) {
            const __nv_bfloat162 lo = __floats2bfloat162_rn(r.x, r.y), hi = __floats2bfloat162_rn(r.z, r.w);
            *reinterpret_cast<uint2 *>(y16 + row * C + c) = make_uint2(*reinterpret_cast<const uint32_t *>(&lo), *reinterpret_cast<const uint32_t *>(&hi));
        }
        if (qout) {                                      // the next layer's query: with_pos_embed(y, pos), msdeformattn.py:124
            const float4 p = *reinterpret_cast<const float4 *>(pos + row * C + c);
            const float4 t = make_float4(r.x + p.x, r.y + p.y, r.z + p.z, r.w + p.w);
            if (q_bf16) {
                const __nv_bfloat162 lo = __floats2bfloat162_rn(t.x, t.y), hi = __floats2bfloat162_rn(t.z, t.w);
                *reinterpret_cast<uint2 *>(static_cast<__nv_bfloat16 *>(qout) + row * C + c) =
                    make_uint2(*reinterpret_cast<const uint32_t *>(&lo), *reinterpret_cast<const uint32_t *>(&hi));
            } else {
                *reinterpret_cast<float4 *>(static_cast<float *>(qout) + row * C + c) = t;
            }
        }
    }
}

// ------------------------------------------------------------------------------------------------ MSDA operand preparation
// MSDeformAttn.forward between the two Linear layers and the sampling op (OPS/modules/ms_deform_attn.py:102-117):
//   sampling_offsets(q).view(N,Q,M,L,P,2), softmax(attention_weights(q).view(N,Q,M,L*P)), and
//   loc = reference_points[:, :, None, :, None, :] + offsets / (W_l, H_l)
// from the ONE GEMM output row [offsets (M*L*P*2) | logits (M*L*P)] -> loc [N,Q,M,L,P,2], attn [N,Q,M,L,P], one pass
// (torch: a softmax, a stack, a division, an add and two .contiguous() passes).  One thread per (row, head).
__global__ void __launch_bounds__(256)
msda_prepare_kernel(const float *__restrict__ oa, int ld, const float *__restrict__ ref /* [R, L, 2] */, const int64_t *__restrict__ shapes,
                    size_t rows, int M, int L, int P, float *__restrict__ loc, float *__restrict__ attn) {
    const size_t idx = (size_t)blockIdx.x * 256 + threadIdx.x;
    if (idx >= rows * M) return;
    const size_t row = idx / M;
    const int m = (int)(idx % M), LP = L * P;
    const float *off = oa + row * ld + (size_t)m * LP * 2;
    const float *lg = oa + row * ld + (size_t)M * LP * 2 + (size_t)m * LP;
    float mx = -INFINITY;
    for (int t = 0; t < LP; ++t) mx = fmaxf(mx, __ldg(lg + t));
    float sum = 0.f;
    for (int t = 0; t < LP; ++t) sum += expf(__ldg(lg + t) - mx);
    const float inv = 1.0f / sum;
    float *lo = loc + idx * LP * 2, *ao = attn + idx * LP;
    for (int l = 0; l < L; ++l) {
        const float w = (float)shapes[2 * l + 1], h = (float)shapes[2 * l];
        const float rx = __ldg(ref + (row * L + l) * 2), ry = __ldg(ref + (row * L + l) * 2 + 1);
        for (int p = 0; p < P; ++p) {
            const int t = l * P + p;
            lo[2 * t] = rx + __ldg(off + 2 * t) / w;                       // ms_deform_attn.py:106-109
            lo[2 * t + 1] = ry + __ldg(off + 2 * t + 1) / h;
            ao[t] = expf(__ldg(lg + t) - mx) * inv;
        }
    }
}

// P == 4 (every config of the path): 16-byte accesses, the head's logits and offsets held in registers.
template <int L>
__global__ void __launch_bounds__(256)
msda_prepare_p4_kernel(const float *__restrict__ oa, int ld, const float *__restrict__ ref, const int64_t *__restrict__ shapes,
                       size_t rows, int M, float *__restrict__ loc, float *__restrict__ attn) {
    constexpr int LP = L * 4;
    const size_t idx = (size_t)blockIdx.x * 256 + threadIdx.x;
    if (idx >= rows * M) return;
    const size_t row = idx / M;
    const int m = (int)(idx % M);
    const float4 *off4 = reinterpret_cast<const float4 *>(oa + row * ld + (size_t)m * LP * 2);
    const float4 *lg4 = reinterpret_cast<const float4 *>(oa + row * ld + (size_t)M * LP * 2 + (size_t)m * LP);
    float4 lgv[L];
    float mx = -INFINITY;
#pragma unroll
    for (int l = 0; l < L; ++l) {
        lgv[l] = __ldg(lg4 + l);
        mx = fmaxf(fmaxf(mx, fmaxf(lgv[l].x, lgv[l].y)), fmaxf(lgv[l].z, lgv[l].w));
    }
    float sum = 0.f;
#pragma unroll
    for (int l = 0; l < L; ++l) {                        // same summation order as the scalar kernel
        lgv[l].x = expf(lgv[l].x - mx); sum += lgv[l].x;
        lgv[l].y = expf(lgv[l].y - mx); sum += lgv[l].y;
        lgv[l].z = expf(lgv[l].z - mx); sum += lgv[l].z;
        lgv[l].w = expf(lgv[l].w - mx); sum += lgv[l].w;
    }
    const float inv = 1.0f / sum;
    float4 *lo4 = reinterpret_cast<float4 *>(loc + idx * LP * 2), *ao4 = reinterpret_cast<float4 *>(attn + idx * LP);
#pragma unroll
    for (int l = 0; l < L; ++l) {
        const float w = (float)shapes[2 * l + 1], h = (float)shapes[2 * l];
        const float2 r = __ldg(reinterpret_cast<const float2 *>(ref + (row * L + l) * 2));
        const float4 o0 = __ldg(off4 + 2 * l), o1 = __ldg(off4 + 2 * l + 1);
        lo4[2 * l] = make_float4(r.x + o0.x / w, r.y + o0.y / h, r.x + o0.z / w, r.y + o0.w / h);
        lo4[2 * l + 1] = make_float4(r.x + o1.x / w, r.y + o1.y / h, r.x + o1.z / w, r.y + o1.w / h);
        ao4[l] = make_float4(lgv[l].x * inv, lgv[l].y * inv, lgv[l].z * inv, lgv[l].w * inv);
    }
}

}  // namespace quber

extern "C" size_t quber_group_norm_workspace_bytes(int N, int C, int HW, int G, int layout) {
    if (N <= 0 || C <= 0 || HW <= 0 || G <= 0 || C % G != 0 || (layout != 0 && layout != 1)) return 0;
    return ((size_t)N * G * quber::gn_tiles(N, C, HW, G, layout) * 2 + (size_t)N * G * 2) * sizeof(float);
}

extern "C" int quber_group_norm_relu(const void *x, void *y, const float *pre_bias, const float *gamma, const float *beta, int N, int C, int HW, int G,
                                     float eps, int relu, int layout, int dtype, void *workspace, size_t workspace_bytes,
                                     void *stream) {
    using namespace quber;
    if (!x || !y || !gamma || !beta || !workspace) return QUBER_E_INVALID;
    if (N <= 0 || C <= 0 || HW <= 0 || G <= 0 || C % G != 0 || (layout != 0 && layout != 1) || (dtype != 0 && dtype != 1))
        return QUBER_E_INVALID;
    if (!aligned16(x) || !aligned16(y) || !aligned16(workspace)) return QUBER_E_ALIGN;
    if (workspace_bytes < quber_group_norm_workspace_bytes(N, C, HW, G, layout)) return QUBER_E_WORKSPACE;
    if (N > 65535 || (layout == 0 && (size_t)N * G > 65535)) return QUBER_E_UNSUPPORTED;
    cudaStream_t st = (cudaStream_t)stream;
    float *ws = static_cast<float *>(workspace);
    if (dtype == 0) return gn_run<float>(x, y, pre_bias, gamma, beta, N, C, HW, G, eps, relu, layout, ws, st);
    return gn_run<__nv_bfloat16>(x, y, pre_bias, gamma, beta, N, C, HW, G, eps, relu, layout, ws, st);
}

extern "C" int quber_upsample_concat(const void *a, const void *b, void *out, int N, int Ca, int Cb, int H, int W, int h, int w,
                                     int dtype, void *stream) {
    using namespace quber;
    if (!b || !out || (Ca > 0 && !a) || N <= 0 || Ca < 0 || Cb <= 0 || H <= 0 || W <= 0 || h <= 0 || w <= 0 || (dtype != 0 && dtype != 1))
        return QUBER_E_INVALID;
    const int V = dtype == 0 ? 4 : 8;
    if (Ca % V != 0 || Cb % V != 0) return QUBER_E_UNSUPPORTED;
    if ((a && !aligned16(a)) || !aligned16(b) || !aligned16(out)) return QUBER_E_ALIGN;
    const size_t total = (size_t)N * H * W * ((Ca + Cb) / V);
    if ((total + 255) / 256 > 0x7fffffffull) return QUBER_E_UNSUPPORTED;
    const float rh = (float)h / (float)H, rw = (float)w / (float)W;      // area_pixel_compute_scale, align_corners=False
    cudaStream_t st = (cudaStream_t)stream;
    const unsigned grid = (unsigned)((total + 255) / 256);
    if (dtype == 0)
        upsample_concat_kernel<float><<<grid, 256, 0, st>>>(static_cast<const float *>(a), static_cast<const float *>(b),
                                                          static_cast<float *>(out), total, Ca, Cb, H, W, h, w, rh, rw);
    else
        upsample_concat_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>(static_cast<const __nv_bfloat16 *>(a),
                                                                  static_cast<const __nv_bfloat16 *>(b),
                                                                  static_cast<__nv_bfloat16 *>(out), total, Ca, Cb, H, W, h, w, rh, rw);
    return check_launch("upsample_concat_kernel");
}

extern "C" int quber_maxpool3x3s2(const void *x, void *y, int N, int C, int H, int W, int dtype, void *stream) {
    using namespace quber;
    if (!x || !y || N <= 0 || C <= 0 || H <= 0 || W <= 0 || (dtype != 0 && dtype != 1)) return QUBER_E_INVALID;
    const int V = dtype == 0 ? 4 : 8;
    if (C % V != 0) return QUBER_E_UNSUPPORTED;
    if (!aligned16(x) || !aligned16(y)) return QUBER_E_ALIGN;
    const int Ho = (H + 2 - 3) / 2 + 1, Wo = (W + 2 - 3) / 2 + 1;
    const size_t total = (size_t)N * Ho * Wo * (C / V);
    if ((total + 255) / 256 > 0x7fffffffull) return QUBER_E_UNSUPPORTED;
    const unsigned grid = (unsigned)((total + 255) / 256);
    cudaStream_t st = (cudaStream_t)stream;
    if (dtype == 0)
        maxpool3x3s2_nhwc_kernel<float><<<grid, 256, 0, st>>>(static_cast<const float *>(x), static_cast<float *>(y), total, C, H, W, Ho, Wo);
    else
        maxpool3x3s2_nhwc_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>(static_cast<const __nv_bfloat16 *>(x),
                                                                    static_cast<__nv_bfloat16 *>(y), total, C, H, W, Ho, Wo);
    return check_launch("maxpool3x3s2_nhwc_kernel");
}

extern "C" int quber_add_layer_norm(const float *a, const float *b, const float *gamma, const float *beta, size_t rows, int C,
                                    float eps, float *y, void *y_bf16, const float *pos, void *q, int q_bf16, void *stream) {
    using namespace quber;
    if (!a || !gamma || !beta || !y || rows == 0 || C <= 0 || (q && !pos)) return QUBER_E_INVALID;
    if (C % 128 != 0 || C > 1024) return QUBER_E_UNSUPPORTED;
    if ((pos && !aligned16(pos)) || (q && !aligned16(q))) return QUBER_E_ALIGN;
    if (!aligned16(a) || (b && !aligned16(b)) || !aligned16(gamma) || !aligned16(beta) || !aligned16(y) || (y_bf16 && !aligned16(y_bf16)))
        return QUBER_E_ALIGN;
    if ((rows + 7) / 8 > 0x7fffffffull) return QUBER_E_UNSUPPORTED;
    const unsigned grid = (unsigned)((rows + 7) / 8);
    cudaStream_t st = (cudaStream_t)stream;
    __nv_bfloat16 *y16 = static_cast<__nv_bfloat16 *>(y_bf16);
    switch (C / 128) {
        case 1: add_layer_norm_kernel<1><<<grid, 256, 0, st>>>(a, b, gamma, beta, rows, eps, y, y16, pos, q, q_bf16); break;
        case 2: add_layer_norm_kernel<2><<<grid, 256, 0, st>>>(a, b, gamma, beta, rows, eps, y, y16, pos, q, q_bf16); break;
        case 3: add_layer_norm_kernel<3><<<grid, 256, 0, st>>>(a, b, gamma, beta, rows, eps, y, y16, pos, q, q_bf16); break;
        case 4: add_layer_norm_kernel<4><<<grid, 256, 0, st>>>(a, b, gamma, beta, rows, eps, y, y16, pos, q, q_bf16); break;
        case 6: add_layer_norm_kernel<6><<<grid, 256, 0, st>>>(a, b, gamma, beta, rows, eps, y, y16, pos, q, q_bf16); break;
        case 8: add_layer_norm_kernel<8><<<grid, 256, 0, st>>>(a, b, gamma, beta, rows, eps, y, y16, pos, q, q_bf16); break;
        default: return QUBER_E_UNSUPPORTED;
    }
    return check_launch("add_layer_norm_kernel");
}

extern "C" int quber_msda_prepare(const float *oa, int ld, const float *reference_points, const int64_t *shapes, size_t rows, int M,
                                  int L, int P, float *loc, float *attn, void *stream) {
    using namespace quber;
    if (!oa || !reference_points || !shapes || !loc || !attn || rows == 0 || M <= 0 || L <= 0 || P <= 0 || ld < M * L * P * 3)
        return QUBER_E_INVALID;
    const size_t n = rows * (size_t)M;
    if ((n + 255) / 256 > 0x7fffffffull) return QUBER_E_UNSUPPORTED;
    const unsigned grid = (unsigned)((n + 255) / 256);
    cudaStream_t st = (cudaStream_t)stream;
    const bool vec = P == 4 && ld % 4 == 0 && aligned16(oa) && aligned16(loc) && aligned16(attn) &&
                     (reinterpret_cast<uintptr_t>(reference_points) & 7u) == 0;
    if (vec && L == 3) msda_prepare_p4_kernel<3><<<grid, 256, 0, st>>>(oa, ld, reference_points, shapes, rows, M, loc, attn);
    else if (vec && L == 4) msda_prepare_p4_kernel<4><<<grid, 256, 0, st>>>(oa, ld, reference_points, shapes, rows, M, loc, attn);
    else if (vec && L == 1) msda_prepare_p4_kernel<1><<<grid, 256, 0, st>>>(oa, ld, reference_points, shapes, rows, M, loc, attn);
    else msda_prepare_kernel<<<grid, 256, 0, st>>>(oa, ld, reference_points, shapes, rows, M, L, P, loc, attn);
    return check_launch("msda_prepare_kernel");
}
