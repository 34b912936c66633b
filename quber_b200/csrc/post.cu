// K3/K4: centre NMS + top-k, then the fused decode / grouping / panoptic merge / instance pass.
//
// Reference (all of it replaced by quber_post_process, with no host round trip):
//   maskrefiner/modeling/mask_refiner/model.py:689-708   x4 bilinear up-sampling of every head output
//   model.py:292                                         r.sigmoid().round()
//   maskrefiner/modeling/mask_refiner/post_processing.py:9-41    find_instance_center
//   post_processing.py:44-76                                     group_pixels
//   post_processing.py:79-162                                    get_instance_segmentation, merge_semantic_and_instance
//   model.py:313-356                                     per-instance masks / scores / boxes
//   eval/eval_utils.py:308-328                           error-map argmax -> one-hot
//
// Launch sequence on one stream (B images per launch, no host synchronisation):
//   memset(zeroed workspace) -> center_nms -> center_select -> group -> eee -> relabel -> offsets -> emit -> finalize
// The 1/4-resolution head outputs are the only float inputs read from HBM; the up-sampled maps are
// recomputed where they are consumed (the bilinear formula is deterministic) and written once when
// the caller asks for them (sem_seg, eee_full).
#include <limits.h>

#include "common.cuh"

namespace quber {

constexpr int P_TW = 128, P_TH = 16, P_THREADS = 256;   // pixel tile of group/emit: 8 px per thread
constexpr int N_TW = 64, N_TH = 32;                     // pixel tile of the NMS kernel
constexpr int SEL_THREADS = 1024;
constexpr int SEL_SMEM_VALS = 4096;
constexpr unsigned INS_STUFF = 0xffffu;                 // foreground pixel of an image without centres

struct InstStat {                 // per (image, instance index); zero-initialised, integer atomics only
    unsigned long long sig;       // sum of sigmoid(fg logit) in 2^-32 fixed point (deterministic)
    unsigned long long sy, sx;
    unsigned int cnt, maxx, maxnx, maxy, maxny, pad;
};

struct PostWs {
    uint32_t *bitmask;   // [B,H,WB]
    uint16_t *ins;       // [B,H,W]
    InstStat *stats;     // [B,top_k+1]  per GROUPED id (index 0: foreground pixels of a centre-less image)   zeroed
    uint32_t *newid;     // [B,top_k+1]  grouped id -> instance index + 1 (0 = dropped)
    uint32_t *oldid;     // [B,top_k+1]  instance index -> grouped id
    int32_t *n_inst;     // [B]
    int32_t *has_stuff;  // [B]
    size_t zero_bytes;
    size_t total;
};

// low-resolution input size and logical width of the frames (quber_post_params.in_h / in_w / valid_w, 0 = derived)
__host__ __device__ inline UpSampler up_of(const quber_post_params &p) {
    return UpSampler(p.in_h > 0 ? p.in_h : p.H / p.scale, p.in_w > 0 ? p.in_w : p.W / p.scale, p.scale, true);
}
__host__ __device__ inline int valid_w_of(const quber_post_params &p) { return p.valid_w > 0 ? p.valid_w : p.W; }

static PostWs carve(const quber_post_params &p, void *base) {
    PostWs w;
    const size_t B = p.B, H = p.H, W = p.W, WB = (W + 31) / 32, K1 = (size_t)p.top_k + 1;
    size_t off = 0;
    auto take = [&](size_t bytes) { size_t o = off; off += align_up(bytes, 256); return o; };
    char *b = static_cast<char *>(base);
    const size_t o_stats = take(B * K1 * sizeof(InstStat));
    w.zero_bytes = off;
    const size_t o_bit = take(B * H * WB * 4);
    const size_t o_ins = take(B * H * W * 2);
    const size_t o_new = take(B * K1 * 4);
    const size_t o_old = take(B * K1 * 4);
    const size_t o_ni = take(B * 4);
    const size_t o_hs = take(B * 4);
    w.total = off;
    w.stats = reinterpret_cast<InstStat *>(b + o_stats);
    w.bitmask = reinterpret_cast<uint32_t *>(b + o_bit);
    w.ins = reinterpret_cast<uint16_t *>(b + o_ins);
    w.newid = reinterpret_cast<uint32_t *>(b + o_new);
    w.oldid = reinterpret_cast<uint32_t *>(b + o_old);
    w.n_inst = reinterpret_cast<int32_t *>(b + o_ni);
    w.has_stuff = reinterpret_cast<int32_t *>(b + o_hs);
    return w;
}

// ------------------------------------------------------------------------------------------------
// Eight consecutive full-resolution pixels (x0 .. x0+7, x0 % 8 == 0) of row y, sampled from a 1/4-resolution plane
// with EXACTLY UpSampler::at's arithmetic (same products, same sums, same rounding), but with the per-row and
// per-column set-up shared by every plane and 8 loads per plane instead of 32: at scale 4 the eight pixels only
// touch source columns 2m-1 .. 2m+2 (m = x0 / 8).  Column c[i] is clamped into the plane, which reproduces the
// reference's `+ (x1 < w - 1)` step at the right edge; at the left edge (m == 0) the clamped source coordinate
// makes pixels 0 and 1 read columns (0, 1) instead of (-1, 0).
// ------------------------------------------------------------------------------------------------
struct Row8 {
    int top, bot;            // element offsets of the two source rows
    int c0, c1, c2, c3;      // the four source columns
    float ly, hy;
    float lx[8];
    bool left;
    __device__ __forceinline__ Row8(const UpSampler &up, int y, int x0) {
        float sy = __fsub_rn(__fmul_rn(up.r, __fadd_rn((float)y, 0.5f)), 0.5f);
        sy = sy < 0.f ? 0.f : sy;
        const int y1 = (int)sy;
        top = y1 * up.w;
        bot = top + ((y1 < up.h - 1) ? up.w : 0);
        ly = __fsub_rn(sy, (float)y1);
        hy = __fsub_rn(1.f, ly);
        const int m2 = x0 >> 2;                          // 2m
        left = m2 == 0;
        c0 = max(m2 - 1, 0); c1 = m2; c2 = min(m2 + 1, up.w - 1); c3 = min(m2 + 2, up.w - 1);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            float sx = __fsub_rn(__fmul_rn(up.r, __fadd_rn((float)(x0 + j), 0.5f)), 0.5f);
            sx = sx < 0.f ? 0.f : sx;
            const int x1 = (int)sx;
            lx[j] = __fsub_rn(sx, (float)x1);
        }
    }
    __device__ __forceinline__ void sample(const float *__restrict__ plane, float *out) const {
        const float *t = plane + top, *b = plane + bot;
        const float t0 = __ldg(t + c0), t1 = __ldg(t + c1), t2 = __ldg(t + c2), t3 = __ldg(t + c3);
        const float b0 = __ldg(b + c0), b1 = __ldg(b + c1), b2 = __ldg(b + c2), b3 = __ldg(b + c3);
        // pixel j reads source columns (x1, x1 + 1): j = 0,1 -> (2m-1, 2m); 2..5 -> (2m, 2m+1); 6,7 -> (2m+1, 2m+2)
        const float ta[8] = {left ? t1 : t0, left ? t1 : t0, t1, t1, t1, t1, t2, t2};
        const float tb[8] = {left ? t2 : t1, left ? t2 : t1, t2, t2, t2, t2, t3, t3};
        const float ba[8] = {left ? b1 : b0, left ? b1 : b0, b1, b1, b1, b1, b2, b2};
        const float bb[8] = {left ? b2 : b1, left ? b2 : b1, b2, b2, b2, b2, b3, b3};
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const float hx = __fsub_rn(1.f, lx[j]);
            const float tp = __fadd_rn(__fmul_rn(hx, ta[j]), __fmul_rn(lx[j], tb[j]));
            const float bt = __fadd_rn(__fmul_rn(hx, ba[j]), __fmul_rn(lx[j], bb[j]));
            out[j] = __fadd_rn(__fmul_rn(hy, tp), __fmul_rn(ly, bt));
        }
    }
};

// ------------------------------------------------------------------------------------------------
// find_instance_center, part 1: threshold + kxk max-pool NMS -> candidate bit mask.
// post_processing.py:22-29.  Candidate <=> thresholded value equals its window maximum and is > 0
// (only positive survivors can pass the final `> max(kth, 0)` test at :37/:41).
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(P_THREADS)
center_nms_kernel(const float *__restrict__ center, quber_post_params pp, uint32_t *__restrict__ bitmask) {
    extern __shared__ float nsm[];
    const int H = pp.H, W = pp.W, scale = pp.scale, nms_kernel = pp.nms_kernel, Wv = valid_w_of(pp);
    const float threshold = pp.threshold;
    const int p = (nms_kernel - 1) / 2;
    const int TWp = N_TW + nms_kernel - 1, THp = N_TH + nms_kernel - 1;
    float *T = nsm, *RM = nsm + THp * TWp;        // RM: [THp][N_TW] (moved behind the wider T on the fast path)
    const UpSampler up = up_of(pp);
    const int b = blockIdx.z, ty0 = blockIdx.y * N_TH, tx0 = blockIdx.x * N_TW;
    const float *plane = center + (size_t)b * up.h * up.w;
    const int tid = threadIdx.x;
    if (threshold > 0.f && scale > 1) {
        // A bilinear sample is a convex combination of its four low-resolution neighbours (up to rounding), so a tile
        // whose low-resolution footprint (a superset: one cell beyond on every side) stays below the threshold by a
        // relative 1e-4 has no candidate: every thresholded value is -1.  Typical frames: more than half of the tiles.
        const int ly0 = max(0, (ty0 - p) / scale - 1), ly1 = min(up.h - 1, (ty0 + N_TH - 1 + p) / scale + 1);
        const int lx0 = max(0, (tx0 - p) / scale - 1), lx1 = min(up.w - 1, (tx0 + N_TW - 1 + p) / scale + 1);
        const int fw = lx1 - lx0 + 1, fn = max(0, (ly1 - ly0 + 1) * fw);
        const float cut = threshold * 0.9999f;
        int any = 0;
        for (int idx = tid; idx < fn; idx += P_THREADS) {
            const int r = idx / fw, c = idx - r * fw;
            any |= !(__ldg(plane + (size_t)(ly0 + r) * up.w + lx0 + c) <= cut);   // NaN counts as "may be a candidate"
        }
        if (!__syncthreads_or(any)) {
            const int WB0 = (W + 31) / 32;
            if (tid < N_TH * 2) {
                const int gy = ty0 + (tid >> 1), wx = (tx0 >> 5) + (tid & 1);
                if (gy < H && wx < WB0) bitmask[((size_t)b * H + gy) * WB0 + wx] = 0u;
            }
            return;
        }
    }
    // x4 maps with a halo of at most 8: the staged rows start 8 columns left of the tile, so that every thread up-samples
    // 8 consecutive, 8-aligned pixels of one row with the shared set-up of Row8 (the same arithmetic, operation for
    // operation, as UpSampler::at) instead of one pixel at a time
    const bool fast = scale == 4 && p <= 8;
    const int TWs = fast ? N_TW + 16 : TWp, xoff = fast ? 8 - p : 0;      // row pitch of T, column of the window's first tap
    if (fast) {
        RM = nsm + THp * TWs;
        constexpr int UPR = (N_TW + 16) / 8;
        for (int u = tid; u < THp * UPR; u += P_THREADS) {
            const int r = u / UPR, cu = u - r * UPR;
            const int gy = ty0 - p + r, gx0 = tx0 - 8 + 8 * cu;
            float v[8];
            const bool rowin = gy >= 0 && gy < H && gx0 >= 0 && gx0 < Wv;
            if (rowin) {
                const Row8 row(up, gy, gx0);
                row.sample(plane, v);
            }
#pragma unroll
            for (int j = 0; j < 8; ++j)
                T[r * TWs + 8 * cu + j] = (rowin && gx0 + j < Wv) ? (v[j] > threshold ? v[j] : -1.0f) : -INFINITY;
        }
    } else {
        for (int idx = tid; idx < THp * TWp; idx += P_THREADS) {
            const int r = idx / TWp, c = idx - r * TWp;
            const int gy = ty0 - p + r, gx = tx0 - p + c;
            float t = -INFINITY;                         // max_pool2d pads with -inf (columns past the cropped width too)
            if (gy >= 0 && gy < H && gx >= 0 && gx < Wv) {
                const float v = up.at(plane, gy, gx);
                t = v > threshold ? v : -1.0f;          // F.threshold(x, thr, -1)
            }
            T[idx] = t;
        }
    }
    __syncthreads();
    for (int idx = tid; idx < THp * N_TW; idx += P_THREADS) {
        const int r = idx / N_TW, c = idx - r * N_TW;
        float m = T[r * TWs + xoff + c];
        for (int j = 1; j < nms_kernel; ++j) m = fmaxf(m, T[r * TWs + xoff + c + j]);
        RM[idx] = m;
    }
    __syncthreads();
    const int warp = tid >> 5, lane = tid & 31, WB = (W + 31) / 32;
    for (int wi = warp; wi < N_TH * 2; wi += P_THREADS / 32) {
        const int row = wi >> 1, half = wi & 1, xx = half * 32 + lane;
        const int gy = ty0 + row, gx = tx0 + xx;
        float m = RM[row * N_TW + xx];
        for (int j = 1; j < nms_kernel; ++j) m = fmaxf(m, RM[(row + j) * N_TW + xx]);
        const float v = T[(row + p) * TWs + xoff + xx + p];
        const bool cand = gy < H && gx < Wv && v == m && v > 0.f;
        const uint32_t word = __ballot_sync(0xffffffffu, cand);
        const int wx = (tx0 >> 5) + half;
        if (lane == 0 && gy < H && wx < WB) bitmask[((size_t)b * H + gy) * WB + wx] = word;
    }
}

// block-wide exclusive scan of one int per thread (blockDim.x == SEL_THREADS); returns the block total via `total`
__device__ __forceinline__ int block_excl_scan(int v, int *warp_tot, int &total) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
    int inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += t;
    }
    __syncthreads();                 // warp_tot may still be read from a previous call
    if (lane == 31) warp_tot[warp] = inc;
    __syncthreads();
    int wbase = 0, tot = 0;
    for (int i = 0; i < nw; ++i) {
        const int t = warp_tot[i];
        if (i < warp) wbase += t;
        tot += t;
    }
    total = tot;
    return wbase + inc - v;
}

// ------------------------------------------------------------------------------------------------
// find_instance_center, part 2: k-th largest candidate (radix select on the float bits), strict `>`,
// raster-ordered emission.  post_processing.py:36-41.  One CTA per image.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(SEL_THREADS)
center_select_kernel(const float *__restrict__ center, quber_post_params pp, int no_topk,
                     const uint32_t *__restrict__ bitmask, int32_t *__restrict__ centers_yx, int32_t *__restrict__ n_centers,
                     int32_t *__restrict__ overflow) {
    const int H = pp.H, W = pp.W, top_k = pp.top_k;
    __shared__ float vals[SEL_SMEM_VALS];
    __shared__ int hist[256];
    __shared__ int warp_tot[32];
    __shared__ int s_n, s_fill, s_bin, s_remaining;
    const UpSampler up = up_of(pp);
    const int b = blockIdx.x, tid = threadIdx.x, WB = (W + 31) / 32, nwords = H * WB;
    const float *plane = center + (size_t)b * up.h * up.w;
    const uint32_t *bm = bitmask + (size_t)b * nwords;
    if (tid == 0) { s_n = 0; s_fill = 0; }
    __syncthreads();
    int local = 0;
    for (int i = tid; i < nwords; i += SEL_THREADS) local += __popc(bm[i]);
#pragma unroll
    for (int o = 16; o; o >>= 1) local += __shfl_xor_sync(0xffffffffu, local, o);
    if ((tid & 31) == 0 && local) atomicAdd(&s_n, local);
    __syncthreads();
    const int n = s_n;
    float cut = 0.f;                                  // `> 0` when fewer than top_k candidates (kth is a -1 filler, clamped)
    if (n >= top_k && !no_topk) {
        const bool in_smem = n <= SEL_SMEM_VALS;
        if (in_smem) {
            for (int i = tid; i < nwords; i += SEL_THREADS) {
                uint32_t wbits = bm[i];
                const int y = i / WB, xb = (i - y * WB) * 32;
                while (wbits) {
                    const int bit = __ffs(wbits) - 1;
                    wbits &= wbits - 1;
                    vals[atomicAdd(&s_fill, 1)] = up.at(plane, y, xb + bit);
                }
            }
            __syncthreads();
        }
        uint32_t prefix = 0, pmask = 0;
        if (tid == 0) s_remaining = top_k;
        for (int pass = 0; pass < 4; ++pass) {
            const int shift = 24 - 8 * pass;
            for (int i = tid; i < 256; i += SEL_THREADS) hist[i] = 0;
            __syncthreads();
            if (in_smem) {
                for (int i = tid; i < n; i += SEL_THREADS) {
                    const uint32_t key = __float_as_uint(vals[i]);
                    if ((key & pmask) == prefix) atomicAdd(&hist[(key >> shift) & 255], 1);
                }
            } else {
                for (int i = tid; i < nwords; i += SEL_THREADS) {
                    uint32_t wbits = bm[i];
                    const int y = i / WB, xb = (i - y * WB) * 32;
                    while (wbits) {
                        const int bit = __ffs(wbits) - 1;
                        wbits &= wbits - 1;
                        const uint32_t key = __float_as_uint(up.at(plane, y, xb + bit));
                        if ((key & pmask) == prefix) atomicAdd(&hist[(key >> shift) & 255], 1);
                    }
                }
            }
            __syncthreads();
            if (tid == 0) {
                int cum = 0, rem = s_remaining, bin = 255;
                for (; bin >= 0; --bin) {
                    if (cum + hist[bin] >= rem) break;
                    cum += hist[bin];
                }
                s_bin = bin < 0 ? 0 : bin;
                s_remaining = rem - cum;
            }
            __syncthreads();
            prefix |= (uint32_t)s_bin << shift;
            pmask |= 255u << shift;
        }
        cut = __uint_as_float(prefix);                // the top_k-th largest score (> 0)
    }
    // raster-ordered emission of candidates with value > cut
    int base = 0;
    for (int i0 = 0; i0 < nwords; i0 += SEL_THREADS) {
        const int i = i0 + tid;
        uint32_t wbits = i < nwords ? bm[i] : 0u;
        const int y = i / WB, xb = (i - y * WB) * 32;
        if (cut > 0.f && wbits) {
            uint32_t keep = 0, t = wbits;
            while (t) {
                const int bit = __ffs(t) - 1;
                t &= t - 1;
                if (up.at(plane, y, xb + bit) > cut) keep |= 1u << bit;
            }
            wbits = keep;
        }
        int total;
        int pos = base + block_excl_scan(__popc(wbits), warp_tot, total);
        while (wbits) {
            const int bit = __ffs(wbits) - 1;
            wbits &= wbits - 1;
            if (pos < top_k) {
                centers_yx[((size_t)b * top_k + pos) * 2] = y;
                centers_yx[((size_t)b * top_k + pos) * 2 + 1] = xb + bit;
            }
            ++pos;
        }
        base += total;
    }
    if (tid == 0) {
        n_centers[b] = base < top_k ? base : top_k;
        if (base > top_k) atomicOr(overflow, 2);
    }
}

// Samples 8 pixels of one plane: the shared-set-up route at scale 4, UpSampler::at otherwise.
template <bool S4>
__device__ __forceinline__ void sample8(const UpSampler &up, const Row8 &r8, const float *__restrict__ plane, int y, int x0,
                                        float *out) {
    if (S4) { r8.sample(plane, out); return; }
#pragma unroll
    for (int j = 0; j < 8; ++j) out[j] = up.at(plane, y, x0 + j);
}

// ------------------------------------------------------------------------------------------------
// Fused per-pixel pass 1: up-sample, foreground test, nearest-centre grouping, per-id statistics (area, score sums,
// centroid sums, box), error-map up-sample + argmax one-hot.   model.py:689-708,292,318-353; post_processing.py:44-76,
// 103-107; eval/eval_utils.py:311-313.
//
// Grouping = argmin_k sqrt(fma(dx,dx,dy*dy)) with the first minimum winning (torch.norm + argmin,
// post_processing.py:71-73).  Three levels of exact pruning keep the K' x HW distance evaluations of the reference
// (12 K' HW bytes materialised there) down to a few centres per pixel:
//   CTA     the voted positions of the tile's foreground pixels lie in a box; a centre whose minimum distance to the box
//           exceeds the smallest maximum distance of any centre to it can not win (or tie) for any pixel of the tile.
//           Survivors are a bit set in shared memory, walked in ascending order (the first-minimum rule needs that order).
//   thread  the same argument on the thread's eight pixels against their current bests.
//   pixel   SQUARED distances are compared: sqrt is monotone, so a candidate with d2 >= best_d2 never wins, and one with
//           d2 < best_d2 * (1 - 2^-21) always does (the two roots are then more than one ulp apart).  Only in the 4-ulp band
//           in between can both roots round to one float — a tie the reference gives to the earlier centre — and there the
//           two correctly rounded roots are compared like the reference does.
// The statistics of model.py:318-353 are accumulated here, per GROUPED id and in integers (the sigmoid in 2^-32 fixed
// point), while the up-sampled foreground logits are still in registers; relabel_kernel later maps ids to instances.
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ int f2ord(float f) { const int i = __float_as_int(f); return i >= 0 ? i : i ^ 0x7fffffff; }   // order-preserving
__device__ __forceinline__ float ord2f(int i) { return __int_as_float(i >= 0 ? i : i ^ 0x7fffffff); }

template <bool S4>
__global__ void __launch_bounds__(P_THREADS, S4 ? 4 : 2)
group_kernel(quber_post_params p, quber_post_io io, const int32_t *__restrict__ n_centers, uint16_t *__restrict__ ins,
             InstStat *__restrict__ stats) {
    extern __shared__ __align__(16) unsigned char gsm_raw[];
    const int b = blockIdx.z, tid = threadIdx.x, H = p.H, W = p.W;
    const int n = n_centers[b];
    InstStat *sst = reinterpret_cast<InstStat *>(gsm_raw);                           // [n+1] (0 = stuff)
    float2 *cyx = reinterpret_cast<float2 *>(sst + (p.top_k + 1));                    // [top_k] (y, x)
    uint32_t *cand = reinterpret_cast<uint32_t *>(cyx + p.top_k);                     // [(top_k+31)/32] surviving centres
    __shared__ int s_box[4];          // ordered-int min y, max y, min x, max x of the tile's voted positions
    __shared__ int s_bound;           // ordered-int of the smallest maximum distance
    __shared__ int s_noprune;
    const int nwords = (n + 31) >> 5;
    for (int i = tid; i < n; i += P_THREADS)
        cyx[i] = make_float2((float)io.centers_yx[((size_t)b * p.top_k + i) * 2], (float)io.centers_yx[((size_t)b * p.top_k + i) * 2 + 1]);
    for (int i = tid; i < (n + 1) * (int)(sizeof(InstStat) / 4); i += P_THREADS) reinterpret_cast<uint32_t *>(sst)[i] = 0;
    for (int i = tid; i < nwords; i += P_THREADS) cand[i] = 0;
    if (tid == 0) { s_box[0] = s_box[2] = INT_MAX; s_box[1] = s_box[3] = INT_MIN; s_bound = INT_MAX; s_noprune = 0; }
    __syncthreads();
    const UpSampler up = up_of(p);
    const int Wv = valid_w_of(p);
    const size_t hw = (size_t)up.h * up.w, HW = (size_t)H * W;
    // a warp covers 8 rows x 32 pixels of the tile (not 2 x 128): the compact footprint is what lets the thread-level
    // pruning skip a centre for the WHOLE warp
    const int wrp = tid >> 5, ln = tid & 31;
    const int y = blockIdx.y * P_TH + (wrp >> 2) * 8 + (ln >> 2), x0 = blockIdx.x * P_TW + ((wrp & 3) * 4 + (ln & 3)) * 8;
    const bool inb = y < H && x0 < W;
    const size_t pix = (size_t)y * W + x0;
    float sem[8], ly[8], lx[8];
    unsigned id[8];
    unsigned fgm = 0;
#pragma unroll
    for (int j = 0; j < 8; ++j) { id[j] = 0; sem[j] = 0.f; ly[j] = lx[j] = 0.f; }
    if (inb) {
        const float *fgp = io.foreground + (size_t)b * hw;
        const Row8 row(up, y, x0);
        sample8<S4>(up, row, fgp, y, x0, sem);
#pragma unroll
        for (int j = 0; j < 8; ++j)
            if (sem[j] > QUBER_FG_LOGIT_CUT && x0 + j < Wv) fgm |= 1u << j;   // r.sigmoid().round() == 1 (inside the crop)
        if (io.sem_seg) {
            float *o = io.sem_seg + (size_t)b * HW + pix;
            st_stream_f4(o, make_float4(sem[0], sem[1], sem[2], sem[3]));
            st_stream_f4(o + 4, make_float4(sem[4], sem[5], sem[6], sem[7]));
        }
        if (fgm && n > 0) {
            const float *oyp = io.offset + (size_t)b * 2 * hw, *oxp = oyp + hw;
            sample8<S4>(up, row, oyp, y, x0, ly);
            sample8<S4>(up, row, oxp, y, x0, lx);
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                ly[j] = __fadd_rn((float)y, __fmul_rn(ly[j], p.offset_mul));
                lx[j] = __fadd_rn((float)(x0 + j), __fmul_rn(lx[j], p.offset_mul));
            }
        }
    }
    // ---- CTA-level pruning: box of the voted positions of the tile's foreground pixels
    if (n > 1) {
        float mny = INFINITY, mxy = -INFINITY, mnx = INFINITY, mxx = -INFINITY;
        bool weird = false;
#pragma unroll
        for (int j = 0; j < 8; ++j)
            if (fgm >> j & 1) {
                weird |= !(fabsf(ly[j]) < 1e30f) || !(fabsf(lx[j]) < 1e30f);          // NaN / inf offsets: no pruning at all
                mny = fminf(mny, ly[j]); mxy = fmaxf(mxy, ly[j]);
                mnx = fminf(mnx, lx[j]); mxx = fmaxf(mxx, lx[j]);
            }
        const unsigned any = __ballot_sync(0xffffffffu, fgm != 0);
        if (any) {
            const int a0 = __reduce_min_sync(0xffffffffu, f2ord(mny)), a1 = __reduce_max_sync(0xffffffffu, f2ord(mxy));
            const int a2 = __reduce_min_sync(0xffffffffu, f2ord(mnx)), a3 = __reduce_max_sync(0xffffffffu, f2ord(mxx));
            const unsigned wd = __ballot_sync(0xffffffffu, weird);
            if (ln == 0) {
                atomicMin(&s_box[0], a0); atomicMax(&s_box[1], a1); atomicMin(&s_box[2], a2); atomicMax(&s_box[3], a3);
                if (wd) s_noprune = 1;
            }
        }
        __syncthreads();
        const bool have = s_box[0] != INT_MAX;
        const float by0 = ord2f(s_box[0]), by1 = ord2f(s_box[1]), bx0 = ord2f(s_box[2]), bx1 = ord2f(s_box[3]);
        if (have && !s_noprune) {
            for (int k = tid; k < n; k += P_THREADS) {          // largest distance from centre k to the box: at a corner
                const float2 c = cyx[k];
                const float dy = fmaxf(fabsf(c.x - by0), fabsf(c.x - by1)), dx = fmaxf(fabsf(c.y - bx0), fabsf(c.y - bx1));
                atomicMin(&s_bound, f2ord(sqrtf(dy * dy + dx * dx)));
            }
        }
        __syncthreads();
        if (have) {
            // keep k iff its smallest distance to the box does not exceed the bound; slack (relative 1e-4, absolute 1e-2 px) is
            // orders of magnitude above the fp32 rounding of these distances, so a dropped centre never wins or ties
            const float bound = s_noprune ? INFINITY : ord2f(s_bound) * 1.0001f + 1e-2f;
            for (int k = tid; k < n; k += P_THREADS) {
                const float2 c = cyx[k];
                const float dy = fmaxf(fmaxf(by0 - c.x, c.x - by1), 0.f), dx = fmaxf(fmaxf(bx0 - c.y, c.y - bx1), 0.f);
                if (!(sqrtf(dy * dy + dx * dx) > bound)) atomicOr(&cand[k >> 5], 1u << (k & 31));
            }
        }
        __syncthreads();
    } else if (n == 1) {
        if (tid == 0) cand[0] = 1u;
        __syncthreads();
    }
    if (inb) {
        if (fgm && n == 0) {
#pragma unroll
            for (int j = 0; j < 8; ++j) if (fgm >> j & 1) id[j] = INS_STUFF;
        } else if (fgm) {
            float bd2[8];
            int best[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) { bd2[j] = INFINITY; best[j] = 0; }
            // Thread-level pruning (exact): the eight voted positions of a thread lie in a small box (midpoint m, half
            // diagonal r).  A centre at distance D from m is at least D - r from each of them, so it cannot beat any current
            // best when D > max_j best_j + r; the bound carries a relative 2e-5 + absolute 1e-3 slack.
            float mny = ly[0], mxy = ly[0], mnx = lx[0], mxx = lx[0];
#pragma unroll
            for (int j = 1; j < 8; ++j) {
                mny = fminf(mny, ly[j]); mxy = fmaxf(mxy, ly[j]);
                mnx = fminf(mnx, lx[j]); mxx = fmaxf(mxx, lx[j]);
            }
            const float my = 0.5f * (mny + mxy), mx = 0.5f * (mnx + mxx);
            const float rad = sqrtf((mxy - my) * (mxy - my) + (mxx - mx) * (mxx - mx)) * 1.0001f + 1e-3f;
            float cut2 = INFINITY;
            bool first = true;
            for (int wd = 0; wd < nwords; ++wd) {
                uint32_t bits = cand[wd];
                while (bits) {
                    const int k = 32 * wd + __ffs(bits) - 1;
                    bits &= bits - 1;
                    const float2 c = cyx[k];
                    if (!first) {
                        const float my_ = c.x - my, mx_ = c.y - mx;
                        if (my_ * my_ + mx_ * mx_ > cut2) continue;                            // NaN bound: never skips
                    }
                    bool changed = false;
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        const float dy = __fsub_rn(c.x, ly[j]), dx = __fsub_rn(c.y, lx[j]);
                        const float d2 = __fmaf_rn(dx, dx, __fmul_rn(dy, dy));
                        if (first) { bd2[j] = d2; best[j] = k; changed = true; }
                        else if (d2 < bd2[j]) {
                            bool better = true;
                            if (!(d2 < bd2[j] * 0.99999952316284179688f))                      // 1 - 2^-21: the 4-ulp band,
                                better = __fsqrt_rn(d2) < __fsqrt_rn(bd2[j]);                  // compare the roots
                            if (better) { bd2[j] = d2; best[j] = k; changed = true; }
                        }
                    }
                    first = false;
                    if (changed) {
                        float bb = bd2[0];
#pragma unroll
                        for (int j = 1; j < 8; ++j) bb = fmaxf(bb, bd2[j]);
                        const float s_ = sqrtf(bb) + rad;
                        cut2 = s_ * s_ * 1.00002f + 1e-3f;
                    }
                }
            }
#pragma unroll
            for (int j = 0; j < 8; ++j) if (fgm >> j & 1) id[j] = best[j] + 1;
        }
        st_stream_u4(ins + (size_t)b * HW + pix,
                     make_uint4(id[0] | (id[1] << 16), id[2] | (id[3] << 16), id[4] | (id[5] << 16), id[6] | (id[7] << 16)));
        if (io.ins_seg) {
            int32_t *o = io.ins_seg + (size_t)b * HW + pix;
            unsigned v[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) v[j] = id[j] == INS_STUFF ? 0u : id[j];
            st_stream_u4(o, make_uint4(v[0], v[1], v[2], v[3]));
            st_stream_u4(o + 4, make_uint4(v[4], v[5], v[6], v[7]));
        }
    }
    // ---- statistics per grouped id (slot 0 = stuff).  A thread whose eight pixels share one id (all but the threads on an
    // instance border) joins the other such lanes of its warp: match.any groups the lanes by id and redux.sync adds /
    // maximises inside the group, so one lane issues the shared-memory atomics.  All sums are integers (the sigmoid in 32.32
    // fixed point), so the result does not depend on the grouping or the order.
    {
        int slot[8];
        bool uni = true, anyfg = false;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            slot[j] = id[j] == 0 ? -1 : (id[j] == INS_STUFF ? 0 : (int)id[j]);
            uni &= slot[j] == (id[0] == 0 ? -1 : (id[0] == INS_STUFF ? 0 : (int)id[0]));
            anyfg |= slot[j] >= 0;
        }
        unsigned long long q[8];
        if (anyfg) {
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const float sg = __fdividef(1.0f, 1.0f + __expf(-sem[j]));     // model.py:319 sigmoid
                q[j] = __float2ull_rn(sg * 4294967296.0f);
            }
        }
        const bool fast = uni && slot[0] >= 0;
        const unsigned act = __ballot_sync(0xffffffffu, fast);
        if (fast) {
            unsigned long long sig = 0;
#pragma unroll
            for (int j = 0; j < 8; ++j) sig += q[j];
            const unsigned grp = __match_any_sync(act, slot[0]);
            const unsigned shi = __reduce_add_sync(grp, (unsigned)(sig >> 16)), slo = __reduce_add_sync(grp, (unsigned)(sig & 0xffffu));
            const unsigned cn = __reduce_add_sync(grp, 8u), sys = __reduce_add_sync(grp, 8u * (unsigned)y);
            const unsigned sxs = __reduce_add_sync(grp, 8u * (unsigned)x0 + 28u);
            const unsigned mx = __reduce_max_sync(grp, (unsigned)(x0 + 7)), mnx = __reduce_max_sync(grp, (unsigned)(W - 1 - x0));
            const unsigned my = __reduce_max_sync(grp, (unsigned)y), mny = __reduce_max_sync(grp, (unsigned)(H - 1 - y));
            if ((tid & 31) == __ffs(grp) - 1) {
                InstStat *s = sst + slot[0];
                atomicAdd(&s->sig, ((unsigned long long)shi << 16) + slo);
                atomicAdd(&s->sy, (unsigned long long)sys);
                atomicAdd(&s->sx, (unsigned long long)sxs);
                atomicAdd(&s->cnt, cn);
                atomicMax(&s->maxx, mx);
                atomicMax(&s->maxnx, mnx);
                atomicMax(&s->maxy, my);
                atomicMax(&s->maxny, mny);
            }
        } else if (anyfg) {                                                    // instance border: run-length compressed per thread
            int cur = -1;
            unsigned cnt = 0, sx = 0, mnx = 0, mxx = 0;
            unsigned long long sig = 0;
            auto flush = [&]() {
                if (cur < 0 || cnt == 0) return;
                InstStat *s = sst + cur;
                atomicAdd(&s->sig, sig);
                atomicAdd(&s->sy, (unsigned long long)cnt * (unsigned)y);
                atomicAdd(&s->sx, (unsigned long long)sx);
                atomicAdd(&s->cnt, cnt);
                atomicMax(&s->maxx, mxx);
                atomicMax(&s->maxnx, (unsigned)(W - 1) - mnx);
                atomicMax(&s->maxy, (unsigned)y);
                atomicMax(&s->maxny, (unsigned)(H - 1 - y));
            };
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                if (slot[j] != cur) { flush(); cur = slot[j]; cnt = 0; sx = 0; sig = 0; mnx = x0 + j; }
                if (cur >= 0) { sig += q[j]; ++cnt; sx += x0 + j; mxx = x0 + j; }
            }
            flush();
        }
    }
    __syncthreads();
    for (int i = tid; i <= n; i += P_THREADS) {
        const InstStat s = sst[i];
        if (!s.cnt) continue;
        InstStat *g = stats + (size_t)b * (p.top_k + 1) + i;
        atomicAdd(&g->sig, s.sig); atomicAdd(&g->sy, s.sy); atomicAdd(&g->sx, s.sx);
        atomicAdd(&g->cnt, s.cnt);
        atomicMax(&g->maxx, s.maxx); atomicMax(&g->maxnx, s.maxnx);
        atomicMax(&g->maxy, s.maxy); atomicMax(&g->maxny, s.maxny);
    }
}

// Error maps on their own: x4 up-sample of the C error-head planes (model.py:704-708), argmax -> one-hot
// (eval/eval_utils.py:311-313).  A pure streaming pass (16 + 4 bytes written per pixel and channel, 0.6 MB read per frame),
// kept out of the grouping kernel so that both run at their own occupancy.
template <bool S4>
__global__ void __launch_bounds__(P_THREADS, 4)
eee_kernel(quber_post_params p, quber_post_io io) {
    const int b = blockIdx.z, tid = threadIdx.x, H = p.H, W = p.W, C = p.eee_channels;
    const UpSampler up = up_of(p);
    const size_t hw = (size_t)up.h * up.w, HW = (size_t)H * W;
    const int y = blockIdx.y * P_TH + (tid >> 4), x0 = blockIdx.x * P_TW + (tid & 15) * 8;
    if (y >= H || x0 >= W) return;
    const size_t pix = (size_t)y * W + x0;
    float bestv[8];
    int bestc[8];
    const Row8 rowe(up, y, x0);
    for (int c = 0; c < C; ++c) {
        const float *ep = io.eee + ((size_t)b * C + c) * hw;
        float v[8];
        sample8<S4>(up, rowe, ep, y, x0, v);
#pragma unroll
        for (int j = 0; j < 8; ++j)
            if (c == 0 || v[j] > bestv[j]) { bestv[j] = v[j]; bestc[j] = c; }   // first maximum wins
        if (io.eee_full) {
            float *o = io.eee_full + ((size_t)b * C + c) * HW + pix;
            st_stream_f4(o, make_float4(v[0], v[1], v[2], v[3]));
            st_stream_f4(o + 4, make_float4(v[4], v[5], v[6], v[7]));
        }
    }
    if (io.eee_onehot)
        for (int c = 0; c < C; ++c) {
            uint32_t lo = 0, hi = 0;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                lo |= (uint32_t)(bestc[j] == c) << (8 * j);
                hi |= (uint32_t)(bestc[4 + j] == c) << (8 * j);
            }
            st_stream_u2(io.eee_onehot + ((size_t)b * C + c) * HW + pix, make_uint2(lo, hi));
        }
}

// merge_semantic_and_instance's id loop (post_processing.py:137-150): ids ascending, drop area < min_area,
// running new id; stuff paste (:152-160) reachable only for a centre-less image.  One CTA per image.
__global__ void __launch_bounds__(SEL_THREADS)
relabel_kernel(quber_post_params p, const int32_t *__restrict__ n_centers, const InstStat *__restrict__ stats,
               uint32_t *__restrict__ newid, uint32_t *__restrict__ oldid, int32_t *__restrict__ n_inst,
               int32_t *__restrict__ has_stuff) {
    __shared__ int warp_tot[32];
    const int b = blockIdx.x, tid = threadIdx.x, n = n_centers[b];
    const InstStat *a = stats + (size_t)b * (p.top_k + 1);       // .cnt = area of the grouped id
    uint32_t *nid = newid + (size_t)b * (p.top_k + 1), *oid = oldid + (size_t)b * (p.top_k + 1);
    int base = 0;
    for (int i0 = 0; i0 < n; i0 += SEL_THREADS) {
        const int id = i0 + tid + 1;
        const int flag = (id <= n) && a[id].cnt > 0 && (int)a[id].cnt >= p.min_area;
        int total;
        const int ex = block_excl_scan(flag, warp_tot, total);
        if (id <= n) nid[id] = flag ? base + ex + 1 : 0;
        if (flag) oid[base + ex] = (uint32_t)id;
        base += total;
    }
    if (tid == 0) {
        nid[0] = 0;
        int hs = 0;
        if (n == 0) {
            oid[0] = 0;                                          // the stuff paste's statistics live in slot 0
            const uint32_t s = a[0].cnt;
            hs = s > 0 && (long long)s >= (long long)p.stuff_area;
        }
        has_stuff[b] = hs;
        n_inst[b] = hs ? 1 : base;
    }
}

__global__ void __launch_bounds__(SEL_THREADS)
offsets_kernel(int B, int max_instances, const int32_t *__restrict__ n_inst, int32_t *__restrict__ inst_offsets,
               int32_t *__restrict__ status) {
    __shared__ int warp_tot[32];
    int base = 0;
    for (int i0 = 0; i0 < B; i0 += SEL_THREADS) {
        const int i = i0 + threadIdx.x;
        const int v = i < B ? n_inst[i] : 0;
        int total;
        const int ex = block_excl_scan(v, warp_tot, total);
        if (i < B) inst_offsets[i] = base + ex;
        base += total;
    }
    if (threadIdx.x == 0) {
        inst_offsets[B] = base;
        if (base > max_instances) atomicOr(status, 1);
    }
}

// Fused per-pixel pass 2: panoptic labels, compact per-instance masks, label map and the integer
// statistics the scores / boxes are made of.  post_processing.py:129,150,160; model.py:318-353.
__global__ void __launch_bounds__(P_THREADS)
emit_kernel(quber_post_params p, quber_post_io io, const uint16_t *__restrict__ ins, const uint32_t *__restrict__ newid,
            const int32_t *__restrict__ n_inst, const int32_t *__restrict__ has_stuff) {
    extern __shared__ __align__(16) unsigned char esm[];
    const int b = blockIdx.z, tid = threadIdx.x, H = p.H, W = p.W;
    const int ninst = n_inst[b], hs = has_stuff[b], nc = io.n_centers[b];
    uint32_t *nid = reinterpret_cast<uint32_t *>(esm);                         // [nc+1]
    for (int i = tid; i <= nc; i += P_THREADS) nid[i] = newid[(size_t)b * (p.top_k + 1) + i];
    __syncthreads();
    const size_t HW = (size_t)H * W;
    const int y = blockIdx.y * P_TH + (tid >> 4), x0 = blockIdx.x * P_TW + (tid & 15) * 8;
    const bool inb = y < H && x0 < W;
    int idx[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) idx[j] = -1;
    const size_t pix = (size_t)y * W + x0;
    if (inb) {
        const uint4 raw = *reinterpret_cast<const uint4 *>(ins + (size_t)b * HW + pix);
        const uint32_t rw[4] = {raw.x, raw.y, raw.z, raw.w};
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const unsigned v = (rw[j >> 1] >> (16 * (j & 1))) & 0xffffu;
            int t = -1;
            if (v == INS_STUFF) t = hs ? 0 : -1;
            else if (v) t = (int)nid[v] - 1;
            idx[j] = t;
        }
        if (io.panoptic) {
            float pv[8];
#pragma unroll
            for (int j = 0; j < 8; ++j)
                pv[j] = idx[j] < 0 ? p.void_label : (float)(p.label_divisor + (hs ? 0 : idx[j] + 1));
            float *o = io.panoptic + (size_t)b * HW + pix;
            st_stream_f4(o, make_float4(pv[0], pv[1], pv[2], pv[3]));
            st_stream_f4(o + 4, make_float4(pv[4], pv[5], pv[6], pv[7]));
        }
        if (io.label_map && (p.flags & QUBER_POST_LABEL_U8)) {               // ids <= top_k <= 255 (post_validate)
            uint32_t lo = 0, hi = 0;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                lo |= (uint32_t)(idx[j] + 1) << (8 * j);
                hi |= (uint32_t)(idx[4 + j] + 1) << (8 * j);
            }
            st_stream_u2(static_cast<uint8_t *>(io.label_map) + (size_t)b * HW + pix, make_uint2(lo, hi));
        } else if (io.label_map) {
            int32_t *o = static_cast<int32_t *>(io.label_map) + (size_t)b * HW + pix;
            st_stream_u4(o, make_uint4(idx[0] + 1, idx[1] + 1, idx[2] + 1, idx[3] + 1));
            st_stream_u4(o + 4, make_uint4(idx[4] + 1, idx[5] + 1, idx[6] + 1, idx[7] + 1));
        }
        if (io.masks) {
            const int off = io.inst_offsets[b];
            const int ne = min(ninst, p.max_instances - off);
            uint8_t *mo = io.masks + (size_t)off * HW + pix;
            if (ninst <= 254) {
                // ids as bytes (0 = none), one SIMD byte compare per plane and 4 pixels instead of 4 compares + shifts
                uint32_t plo = 0, phi = 0;
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    plo |= (uint32_t)(idx[j] + 1) << (8 * j);
                    phi |= (uint32_t)(idx[4 + j] + 1) << (8 * j);
                }
                for (int i = 0; i < ne; ++i, mo += HW) {
                    const uint32_t key = (uint32_t)(i + 1) * 0x01010101u;
                    st_stream_u2(mo, make_uint2(__vcmpeq4(plo, key) & 0x01010101u, __vcmpeq4(phi, key) & 0x01010101u));
                }
            } else {
                for (int i = 0; i < ne; ++i, mo += HW) {
                    uint32_t lo = 0, hi = 0;
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        lo |= (uint32_t)(idx[j] == i) << (8 * j);
                        hi |= (uint32_t)(idx[4 + j] == i) << (8 * j);
                    }
                    st_stream_u2(mo, make_uint2(lo, hi));
                }
            }
        }
    }
}

// model.py:322-353: label, score = mean(sigmoid(fg)[mask]) * centre[int(cy), int(cx)], box.
__global__ void __launch_bounds__(P_THREADS)
finalize_kernel(quber_post_params p, quber_post_io io, const int32_t *__restrict__ n_inst,
                const int32_t *__restrict__ has_stuff, const InstStat *__restrict__ stats, const uint32_t *__restrict__ oldid) {
    const int b = blockIdx.x;
    const int ninst = n_inst[b], hs = has_stuff[b], off = io.inst_offsets[b];
    const UpSampler up = up_of(p);
    const float *cplane = io.center + (size_t)b * up.h * up.w;
    for (int i = threadIdx.x; i < ninst; i += blockDim.x) {
        const int row = off + i;
        if (row >= p.max_instances) break;
        const InstStat s = stats[(size_t)b * (p.top_k + 1) + oldid[(size_t)b * (p.top_k + 1) + i]];
        io.labels[row] = p.label_divisor + (hs ? 0 : i + 1);
        const double cnt = (double)s.cnt;
        const float sem = (float)((double)s.sig * (1.0 / 4294967296.0) / cnt);
        const int cy = (int)((double)s.sy / cnt), cx = (int)((double)s.sx / cnt);
        io.scores[row] = __fmul_rn(sem, up.at(cplane, cy, cx));
        io.boxes[4 * row + 0] = (float)(p.W - 1 - (int)s.maxnx);
        io.boxes[4 * row + 1] = (float)(p.H - 1 - (int)s.maxny);
        io.boxes[4 * row + 2] = (float)(s.maxx + 1);
        io.boxes[4 * row + 3] = (float)(s.maxy + 1);
    }
}

// ------------------------------------------------------------------------------------------------
// Caller-side filter of the refined instances (eval/refiner_model.py:271-288):
//   keep instance j  <=>  sum(mask_j & fg) / sum(mask_j) > thr        (thr = 0.3; numpy int64 / int64 -> float64)
//   OCID / DoPose: kept masks lose their zero-depth pixels            (refined_mask[zero_depth] = False)
// on the label-map form of the instances (0 = void, j + 1 = j-th instance): one counting pass (integer atomics,
// deterministic), one tiny per-frame pass that decides and renumbers, one relabel pass.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
filter_count_kernel(const int32_t *__restrict__ lmap, const uint8_t *__restrict__ fg, int HW, int max_per,
                    uint32_t *__restrict__ cnt /* [B][max_per][2]: area, overlap */) {
    extern __shared__ uint32_t fsm[];                    // [max_per][2]
    const int b = blockIdx.y;
    for (int i = threadIdx.x; i < 2 * max_per; i += 256) fsm[i] = 0;
    __syncthreads();
    const size_t base = (size_t)b * HW;
    for (int i = (blockIdx.x * 256 + threadIdx.x) * 4; i < HW; i += gridDim.x * 256 * 4) {     // HW % 4 == 0
        const int4 l = *reinterpret_cast<const int4 *>(lmap + base + i);
        uint32_t f = 0x01010101u;
        if (fg) f = *reinterpret_cast<const uint32_t *>(fg + base + i);
        const int ls[4] = {l.x, l.y, l.z, l.w};
#pragma unroll
        for (int j = 0; j < 4; ++j)
            if (ls[j] > 0 && ls[j] <= max_per) {
                atomicAdd(&fsm[2 * (ls[j] - 1)], 1u);
                if ((f >> (8 * j)) & 0xffu) atomicAdd(&fsm[2 * (ls[j] - 1) + 1], 1u);
            }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < 2 * max_per; i += 256)
        if (fsm[i]) atomicAdd(&cnt[(size_t)b * 2 * max_per + i], fsm[i]);
}

__global__ void filter_decide_kernel(const uint32_t *__restrict__ cnt, const int32_t *__restrict__ counts, int B, int max_per, double thr,
                                     int use_fg, uint8_t *__restrict__ keep, int32_t *__restrict__ newid /* [B][max_per] */,
                                     int32_t *__restrict__ new_counts) {
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= B) return;
    const int n = min(counts[b], max_per);
    int m = 0;
    for (int j = 0; j < max_per; ++j) {
        const uint32_t area = cnt[((size_t)b * max_per + j) * 2], ov = cnt[((size_t)b * max_per + j) * 2 + 1];
        // numpy: int64 / int64 -> float64 true division; 0 / 0 = nan fails the comparison
        const bool k = j < n && (!use_fg || (area > 0 && (double)ov / (double)area > thr));
        keep[(size_t)b * max_per + j] = k;
        newid[(size_t)b * max_per + j] = k ? ++m : 0;
    }
    new_counts[b] = m;
}

__global__ void __launch_bounds__(256)
filter_relabel_kernel(const int32_t *__restrict__ lmap, const uint8_t *__restrict__ zero_depth, const int32_t *__restrict__ newid,
                      int HW, int max_per, int32_t *__restrict__ out) {
    const int b = blockIdx.y;
    const size_t base = (size_t)b * HW;
    const int32_t *nid = newid + (size_t)b * max_per;
    for (int i = (blockIdx.x * 256 + threadIdx.x) * 4; i < HW; i += gridDim.x * 256 * 4) {
        const int4 l = *reinterpret_cast<const int4 *>(lmap + base + i);
        uint32_t z = 0u;
        if (zero_depth) z = *reinterpret_cast<const uint32_t *>(zero_depth + base + i);
        const int ls[4] = {l.x, l.y, l.z, l.w};
        int o[4];
#pragma unroll
        for (int j = 0; j < 4; ++j)
            o[j] = (ls[j] > 0 && ls[j] <= max_per && !((z >> (8 * j)) & 0xffu)) ? __ldg(nid + ls[j] - 1) : 0;
        *reinterpret_cast<int4 *>(out + base + i) = make_int4(o[0], o[1], o[2], o[3]);
    }
}

}  // namespace quber

static int post_validate(const quber_post_params *p) {
    if (!p) return QUBER_E_INVALID;
    if (p->B <= 0 || p->H <= 0 || p->W <= 0 || p->scale < 1 || p->top_k < 1 || p->top_k > 1024 ||
        p->nms_kernel < 1 || p->nms_kernel > 15 || (p->nms_kernel & 1) == 0 || p->max_instances < 0 ||
        (p->eee_channels != 0 && (p->eee_channels < 2 || p->eee_channels > 4)) ||
        ((p->flags & QUBER_POST_LABEL_U8) && p->top_k > 255))
        return QUBER_E_INVALID;
    if (p->in_h < 0 || p->in_w < 0 || p->valid_w < 0 || p->valid_w > p->W) return QUBER_E_INVALID;
    if (p->W % 8 != 0) return QUBER_E_UNSUPPORTED;
    if (p->in_h == 0 && p->H % p->scale != 0) return QUBER_E_UNSUPPORTED;       // derived low-resolution size must be exact
    if (p->in_w == 0 && p->W % p->scale != 0) return QUBER_E_UNSUPPORTED;
    // a cropped frame must lie inside the up-sampled map
    if (p->in_h > 0 && (long long)p->in_h * p->scale < p->H) return QUBER_E_INVALID;
    if (p->in_w > 0 && (long long)p->in_w * p->scale < (p->valid_w > 0 ? p->valid_w : p->W)) return QUBER_E_INVALID;
    return QUBER_OK;
}

namespace quber {
// The error-map pass (eee_kernel) depends on nothing but the call's inputs, and it is a pure store stream, while the centre
// chain beside it (NMS -> select -> grouping -> relabel -> offsets) is compute- and latency-bound with several one-CTA-per-
// image steps that leave the GPU almost idle.  It is therefore forked onto a low-priority side stream at the start of the
// call and joined at its end (event fork / join: legal under CUDA graph capture, the side stream joins the capture).
// One side stream and two events per host thread and device, created on first use.
struct PostFork { cudaStream_t side = nullptr; cudaEvent_t fork = nullptr, join = nullptr; };
static int post_fork(cudaStream_t st, PostFork **out) {
    constexpr int MAX_DEV = 64;
    static thread_local PostFork pf[MAX_DEV];
    int dev = -1;
    *out = nullptr;                                                          // no fork: everything on the caller's stream
    QUBER_CUDA_TRY(cudaGetDevice(&dev));
    if (dev < 0 || dev >= MAX_DEV) return QUBER_OK;
    PostFork &f = pf[dev];
    if (!f.side) {
        // streams and events must not be created while the caller's stream is being captured: the first call of a thread
        // under capture stays on one stream
        cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
        QUBER_CUDA_TRY(cudaStreamIsCapturing(st, &cs));
        if (cs != cudaStreamCaptureStatusNone) return QUBER_OK;
        int lo = 0, hi = 0;
        QUBER_CUDA_TRY(cudaDeviceGetStreamPriorityRange(&lo, &hi));
        cudaStream_t s = nullptr;
        QUBER_CUDA_TRY(cudaStreamCreateWithPriority(&s, cudaStreamNonBlocking, lo));
        QUBER_CUDA_TRY(cudaEventCreateWithFlags(&f.fork, cudaEventDisableTiming));
        QUBER_CUDA_TRY(cudaEventCreateWithFlags(&f.join, cudaEventDisableTiming));
        f.side = s;
    }
    *out = &f;
    return QUBER_OK;
}
}  // namespace quber

extern "C" size_t quber_post_workspace_bytes(const quber_post_params *p) {
    if (post_validate(p) != QUBER_OK) return 0;
    return quber::carve(*p, reinterpret_cast<void *>(4096)).total;
}

extern "C" int quber_post_process(const quber_post_params *pp, const quber_post_io *iop, void *workspace,
                                  size_t workspace_bytes, void *stream) {
    using namespace quber;
    if (int rc = post_validate(pp)) return rc;
    if (!iop) return QUBER_E_INVALID;
    const quber_post_params p = *pp;
    const quber_post_io io = *iop;
    if (!io.foreground || !io.center || !io.offset || !io.centers_yx || !io.n_centers || !io.inst_offsets || !io.labels ||
        !io.scores || !io.boxes || !io.status)
        return QUBER_E_INVALID;
    if (p.eee_channels > 0 && !io.eee) return QUBER_E_INVALID;
    const void *al[] = {io.sem_seg, io.eee_full, io.eee_onehot, io.panoptic, io.label_map, io.masks, io.ins_seg};
    for (const void *q : al)
        if (q && !aligned16(q)) return QUBER_E_ALIGN;
    if (!workspace || !aligned16(workspace) || workspace_bytes < carve(p, reinterpret_cast<void *>(4096)).total) return QUBER_E_WORKSPACE;
    const PostWs ws = carve(p, workspace);
    cudaStream_t st = (cudaStream_t)stream;
    QUBER_CUDA_TRY(cudaMemsetAsync(workspace, 0, ws.zero_bytes, st));
    QUBER_CUDA_TRY(cudaMemsetAsync(io.status, 0, sizeof(int32_t), st));

    dim3 pgrid(cdiv(p.W, P_TW), cdiv(p.H, P_TH), p.B);
    const bool has_eee = p.eee_channels > 0 && io.eee && (io.eee_full || io.eee_onehot);
    PostFork *fk = nullptr;
    if (has_eee) {
        // the error maps: a pure streaming pass of their own (measured, C2: fused with grouping 172 us; 99 + 66 us split),
        // forked beside the centre chain
        if (int rc = post_fork(st, &fk)) return rc;
        cudaStream_t es = st;
        if (fk) {
            QUBER_CUDA_TRY(cudaEventRecord(fk->fork, st));
            QUBER_CUDA_TRY(cudaStreamWaitEvent(fk->side, fk->fork, 0));
            es = fk->side;
        }
        if (p.scale == 4) eee_kernel<true><<<pgrid, P_THREADS, 0, es>>>(p, io);
        else eee_kernel<false><<<pgrid, P_THREADS, 0, es>>>(p, io);
        if (int rc = check_launch("eee_kernel")) return rc;
        if (fk) QUBER_CUDA_TRY(cudaEventRecord(fk->join, fk->side));
    }
    if (!(p.flags & QUBER_POST_GIVEN_CENTERS)) {
        const int TWp = N_TW + p.nms_kernel - 1, THp = N_TH + p.nms_kernel - 1;
        const size_t smem = ((size_t)THp * (TWp > N_TW + 16 ? TWp : N_TW + 16) + (size_t)THp * N_TW) * sizeof(float);
        dim3 grid(cdiv(p.W, N_TW), cdiv(p.H, N_TH), p.B);
        center_nms_kernel<<<grid, P_THREADS, smem, st>>>(io.center, p, ws.bitmask);
        if (int rc = check_launch("center_nms_kernel")) return rc;
        center_select_kernel<<<p.B, SEL_THREADS, 0, st>>>(io.center, p, (p.flags & QUBER_POST_NO_TOPK) != 0,
                                                         ws.bitmask, io.centers_yx, io.n_centers, io.status);
        if (int rc = check_launch("center_select_kernel")) return rc;
    }
    {
        const size_t gsmem = ((size_t)p.top_k + 1) * sizeof(InstStat) + (size_t)p.top_k * sizeof(float2) + (((size_t)p.top_k + 31) / 32) * 4;
        if (gsmem > 48 * 1024) {
            QUBER_CUDA_TRY(cudaFuncSetAttribute(group_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)gsmem));
            QUBER_CUDA_TRY(cudaFuncSetAttribute(group_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)gsmem));
        }
        if (p.scale == 4) group_kernel<true><<<pgrid, P_THREADS, gsmem, st>>>(p, io, io.n_centers, ws.ins, ws.stats);
        else group_kernel<false><<<pgrid, P_THREADS, gsmem, st>>>(p, io, io.n_centers, ws.ins, ws.stats);
        if (int rc = check_launch("group_kernel")) return rc;
    }
    relabel_kernel<<<p.B, SEL_THREADS, 0, st>>>(p, io.n_centers, ws.stats, ws.newid, ws.oldid, ws.n_inst, ws.has_stuff);
    if (int rc = check_launch("relabel_kernel")) return rc;
    offsets_kernel<<<1, SEL_THREADS, 0, st>>>(p.B, p.max_instances, ws.n_inst, io.inst_offsets, io.status);
    if (int rc = check_launch("offsets_kernel")) return rc;
    {
        const size_t smem = ((size_t)p.top_k + 1) * 4;
        emit_kernel<<<pgrid, P_THREADS, smem, st>>>(p, io, ws.ins, ws.newid, ws.n_inst, ws.has_stuff);
        if (int rc = check_launch("emit_kernel")) return rc;
    }
    finalize_kernel<<<p.B, P_THREADS, 0, st>>>(p, io, ws.n_inst, ws.has_stuff, ws.stats, ws.oldid);
    if (int rc = check_launch("finalize_kernel")) return rc;
    if (fk) QUBER_CUDA_TRY(cudaStreamWaitEvent(st, fk->join, 0));        // join: the error maps belong to this call's results
    return QUBER_OK;
}

extern "C" size_t quber_filter_workspace_bytes(int B, int max_per) {
    if (B <= 0 || max_per <= 0) return 0;
    return quber::align_up((size_t)B * max_per * 2 * sizeof(uint32_t), 256) + quber::align_up((size_t)B * max_per * sizeof(int32_t), 256);
}

extern "C" int quber_filter_instances(const int32_t *label_map, const int32_t *counts, const uint8_t *fg, const uint8_t *zero_depth,
                                      int B, int H, int W, int max_per, double thr, int32_t *out_label_map, uint8_t *keep,
                                      int32_t *new_counts, void *workspace, size_t workspace_bytes, void *stream) {
    using namespace quber;
    if (!label_map || !counts || !out_label_map || !keep || !new_counts || !workspace) return QUBER_E_INVALID;
    if (B <= 0 || H <= 0 || W <= 0 || max_per <= 0 || max_per > 4096) return QUBER_E_INVALID;
    const long long HWl = (long long)H * W;
    if (HWl % 4 != 0 || HWl > 0x7fffffffll || B > 65535) return QUBER_E_UNSUPPORTED;
    if (!aligned16(label_map) || !aligned16(out_label_map) || (fg && (reinterpret_cast<uintptr_t>(fg) & 3u)) ||
        (zero_depth && (reinterpret_cast<uintptr_t>(zero_depth) & 3u)) || !aligned16(workspace))
        return QUBER_E_ALIGN;
    if (workspace_bytes < quber_filter_workspace_bytes(B, max_per)) return QUBER_E_WORKSPACE;
    const int HW = (int)HWl;
    cudaStream_t st = (cudaStream_t)stream;
    uint32_t *cnt = static_cast<uint32_t *>(workspace);
    int32_t *newid = reinterpret_cast<int32_t *>(static_cast<char *>(workspace) + align_up((size_t)B * max_per * 2 * sizeof(uint32_t), 256));
    QUBER_CUDA_TRY(cudaMemsetAsync(cnt, 0, (size_t)B * max_per * 2 * sizeof(uint32_t), st));
    const int gx = max(1, min(cdiv(HW, 256 * 4 * 4), 64));
    dim3 grid(gx, B);
    filter_count_kernel<<<grid, 256, (size_t)max_per * 2 * sizeof(uint32_t), st>>>(label_map, fg, HW, max_per, cnt);
    if (int rc = check_launch("filter_count_kernel")) return rc;
    filter_decide_kernel<<<cdiv(B, 64), 64, 0, st>>>(cnt, counts, B, max_per, thr, fg != nullptr, keep, newid, new_counts);
    if (int rc = check_launch("filter_decide_kernel")) return rc;
    filter_relabel_kernel<<<grid, 256, 0, st>>>(label_map, zero_depth, newid, HW, max_per, out_label_map);
    return check_launch("filter_relabel_kernel");
}
