// Evaluation metrics of refined masks, the per-pixel part.
//
// Reference: eval/evaluation.py:22-55 (boundary_overlap), :57-274 (multilabel_metrics), eval/utilities.py:672-696
// (seg2bmap).  For ONE image the reference loops over every (gt label, predicted label) pair and runs, per pair, two
// cv2.findContours + drawContours, two cv2.dilate with a disk and four full-frame reductions — K x K' x ~10 passes over
// the frame on the CPU.  Here a batch of (prediction, gt) label-map pairs goes through five launches and everything
// that is O(pixels) happens on the device; what is left for the host is O(K x K') float64 arithmetic on the integer
// counts (quber_b200/evaluation.py), which is why the metrics come out bit-identical to the reference's.
//
//   eval_presence_kernel   label census: one bit per label value (0 .. 65535) and map
//   eval_labels_kernel     sorted label lists (np.unique minus background 0, evaluation.py:83-89)
//   eval_dense_kernel      dense index maps (0 = background, k + 1 = k-th label), the (gt, pred) co-occurrence
//                          histogram = every pair's intersection and every mask's area (:180-197), bounding boxes
//   eval_boundary_kernel   seg2bmap of EVERY mask, one CTA per (image, map, label): cv2.findContours(RETR_EXTERNAL,
//                          CHAIN_APPROX_NONE) + drawContours(thickness 1) = the mask pixels with a 4-neighbour in the
//                          background component that contains the image frame (hole borders and objects nested in holes
//                          do not count) — a flood fill of the complement from the frame, done on bit rows in shared memory
//                          inside the label's bounding box (+1), alternating full-row and full-column sweeps
//   eval_match_kernel      boundary_overlap of every pair at once: a predicted boundary pixel counts for gt label i iff a
//                          boundary pixel of i lies inside skimage's disk(bound_pix) around it (== inside cv2.dilate of that
//                          boundary), and vice versa (:41-55)
#include <limits.h>

#include "common.cuh"

namespace quber {

constexpr int EV_WORDS = 65536 / 32;      // presence bitmap words per (image, map)
constexpr int EV_MAXL = 255;

struct EvalWs {
    uint32_t *presence;   // [2][B][EV_WORDS]                       zeroed
    int32_t *bbox;        // [2][B][ML][4]  (min y, min x, -max y, -max x as atomicMin targets)   set to INT_MAX
    uint8_t *dense;       // [2][B][H][W]
    uint8_t *bmap;        // [2][B][H][W]                           zeroed
    size_t presence_bytes, bbox_bytes, dense_bytes, total;
};

static EvalWs eval_carve(int B, int H, int W, int ML, void *base) {
    EvalWs w;
    char *b = static_cast<char *>(base);
    size_t off = 0;
    auto take = [&](size_t bytes) { size_t o = off; off += align_up(bytes, 256); return o; };
    w.presence_bytes = (size_t)2 * B * EV_WORDS * 4;
    w.bbox_bytes = (size_t)2 * B * ML * 4 * 4;
    w.dense_bytes = (size_t)2 * B * H * W;
    const size_t o_p = take(w.presence_bytes), o_m = take(w.dense_bytes), o_b = take(w.bbox_bytes), o_d = take(w.dense_bytes);
    w.presence = reinterpret_cast<uint32_t *>(b + o_p);
    w.bmap = reinterpret_cast<uint8_t *>(b + o_m);
    w.bbox = reinterpret_cast<int32_t *>(b + o_b);
    w.dense = reinterpret_cast<uint8_t *>(b + o_d);
    w.total = off;
    return w;
}

// which = 0: prediction, 1: gt
__global__ void __launch_bounds__(256)
eval_presence_kernel(const int32_t *__restrict__ pred, const int32_t *__restrict__ gt, size_t HW, int B, uint32_t *__restrict__ presence,
                     int32_t *__restrict__ status) {
    const int which = blockIdx.z, b = blockIdx.y;
    const int32_t *map = (which ? gt : pred) + (size_t)b * HW;
    uint32_t *pr = presence + ((size_t)which * B + b) * EV_WORDS;
    bool bad = false;
    for (size_t i = (size_t)blockIdx.x * 256 + threadIdx.x; i < HW; i += (size_t)gridDim.x * 256) {
        const int32_t v = __ldg(map + i);
        if (v == 0) continue;                                   // BACKGROUND_LABEL
        if (v < 0 || v > 65535) { bad = true; continue; }
        const uint32_t bit = 1u << (v & 31);
        if (!(pr[v >> 5] & bit)) atomicOr(&pr[v >> 5], bit);    // racy pre-check: at most a few redundant atomics per label
    }
    if (__syncthreads_or(bad) && threadIdx.x == 0) atomicOr(status, 2);
}

__global__ void __launch_bounds__(256)
eval_labels_kernel(const uint32_t *__restrict__ presence, int B, int ML, int32_t *__restrict__ pred_labels, int32_t *__restrict__ n_pred,
                   int32_t *__restrict__ gt_labels, int32_t *__restrict__ n_gt, int32_t *__restrict__ status) {
    __shared__ int wtot[8];
    const int which = blockIdx.y, b = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t *pr = presence + ((size_t)which * B + b) * EV_WORDS;
    int32_t *labels = (which ? gt_labels : pred_labels) + (size_t)b * ML;
    uint32_t w[EV_WORDS / 256];                                 // 8 consecutive words per thread: ascending label order
    int cnt = 0;
#pragma unroll
    for (int k = 0; k < EV_WORDS / 256; ++k) { w[k] = pr[tid * (EV_WORDS / 256) + k]; cnt += __popc(w[k]); }
    int inc = cnt;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += t; }
    if (lane == 31) wtot[warp] = inc;
    __syncthreads();
    int base = 0, total = 0;
    for (int i = 0; i < 8; ++i) { if (i < warp) base += wtot[i]; total += wtot[i]; }
    int pos = base + inc - cnt;
#pragma unroll
    for (int k = 0; k < EV_WORDS / 256; ++k) {
        uint32_t bits = w[k];
        while (bits) {
            const int bit = __ffs(bits) - 1;
            bits &= bits - 1;
            if (pos < ML) labels[pos] = (tid * (EV_WORDS / 256) + k) * 32 + bit;
            ++pos;
        }
    }
    if (tid == 0) {
        (which ? n_gt : n_pred)[b] = total < ML ? total : ML;
        if (total > ML) atomicOr(status, 1);
    }
}

__device__ __forceinline__ int ev_find(const int32_t *__restrict__ s_labels, int n, int32_t v) {     // dense index + 1, 0 = none
    int lo = 0, hi = n;
    while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (s_labels[mid] < v) lo = mid + 1; else hi = mid;
    }
    return (lo < n && s_labels[lo] == v) ? lo + 1 : 0;
}

// Dense maps, co-occurrence histogram inter_full[b][gi][pj] (index 0 = background) and bounding boxes.
__global__ void __launch_bounds__(256)
eval_dense_kernel(const int32_t *__restrict__ pred, const int32_t *__restrict__ gt, int H, int W, int B, int ML,
                  const int32_t *__restrict__ pred_labels, const int32_t *__restrict__ n_pred, const int32_t *__restrict__ gt_labels,
                  const int32_t *__restrict__ n_gt, uint8_t *__restrict__ dense, int32_t *__restrict__ bbox,
                  int32_t *__restrict__ inter_full, int hist_in_smem) {
    extern __shared__ int32_t esm[];
    const int b = blockIdx.y, tid = threadIdx.x;
    const int np_ = n_pred[b], ng = n_gt[b];
    int32_t *s_pl = esm, *s_gl = esm + ML;
    int32_t *s_bb = esm + 2 * ML;                               // [2][ML][4]
    int32_t *s_hist = esm + 2 * ML + 8 * ML;                    // [(ng+1)][(np+1)] when it fits
    for (int i = tid; i < np_; i += 256) s_pl[i] = pred_labels[(size_t)b * ML + i];
    for (int i = tid; i < ng; i += 256) s_gl[i] = gt_labels[(size_t)b * ML + i];
    for (int i = tid; i < 8 * ML; i += 256) s_bb[i] = INT_MAX;
    const int hn = (ng + 1) * (np_ + 1);
    if (hist_in_smem) for (int i = tid; i < hn; i += 256) s_hist[i] = 0;
    __syncthreads();
    const size_t HW = (size_t)H * W, base = (size_t)b * HW;
    uint8_t *dp = dense + base, *dg = dense + ((size_t)B + b) * HW;
    int32_t *gh = inter_full + (size_t)b * (ML + 1) * (ML + 1);
    for (size_t i = (size_t)blockIdx.x * 256 + tid; i < HW; i += (size_t)gridDim.x * 256) {
        const int32_t vp = __ldg(pred + base + i), vg = __ldg(gt + base + i);
        const int pj = (vp > 0 && vp <= 65535) ? ev_find(s_pl, np_, vp) : 0;
        const int gi = (vg > 0 && vg <= 65535) ? ev_find(s_gl, ng, vg) : 0;
        dp[i] = (uint8_t)pj;
        dg[i] = (uint8_t)gi;
        if (pj | gi) {
            if (hist_in_smem) atomicAdd(&s_hist[gi * (np_ + 1) + pj], 1);
            else atomicAdd(&gh[(size_t)gi * (ML + 1) + pj], 1);
            const int y = (int)(i / W), x = (int)(i - (size_t)y * W);
            if (pj) { int32_t *q = s_bb + (pj - 1) * 4; atomicMin(q, y); atomicMin(q + 1, x); atomicMin(q + 2, -y); atomicMin(q + 3, -x); }
            if (gi) { int32_t *q = s_bb + (ML + gi - 1) * 4; atomicMin(q, y); atomicMin(q + 1, x); atomicMin(q + 2, -y); atomicMin(q + 3, -x); }
        }
    }
    __syncthreads();
    if (hist_in_smem)
        for (int i = tid; i < hn; i += 256)
            if (s_hist[i]) atomicAdd(&gh[(size_t)(i / (np_ + 1)) * (ML + 1) + i % (np_ + 1)], s_hist[i]);
    for (int i = tid; i < 8 * ML; i += 256) {
        if (s_bb[i] == INT_MAX) continue;
        const int which = i / (4 * ML), r = i % (4 * ML);
        atomicMin(&bbox[((size_t)which * B + b) * ML * 4 + r], s_bb[i]);
    }
}

// seg2bmap of one mask = label `k + 1` of one dense map, inside its bounding box grown by one pixel.
// Bits: bit c of word w of row r = region pixel (r, 32 w + c).  `free` = not the mask (pixels outside the image are free:
// they are the frame), `reach` = free pixels 4-connected to the region's border ring, which lies outside the bounding box and
// is therefore free and connected to the frame (the complement of a rectangle is connected through the frame).
__global__ void __launch_bounds__(256)
eval_boundary_kernel(const uint8_t *__restrict__ dense, const int32_t *__restrict__ bbox, const int32_t *__restrict__ n_pred,
                     const int32_t *__restrict__ n_gt, int H, int W, int B, int ML, unsigned smem_bytes, uint8_t *__restrict__ bmap,
                     int32_t *__restrict__ bcount_pred, int32_t *__restrict__ bcount_gt, int32_t *__restrict__ status) {
    extern __shared__ uint32_t bsm[];
    const int k = blockIdx.x, which = blockIdx.y, b = blockIdx.z, tid = threadIdx.x;
    if (k >= (which ? n_gt : n_pred)[b]) return;
    const int32_t *bb = bbox + (((size_t)which * B + b) * ML + k) * 4;
    const int y0 = bb[0], x0 = bb[1], y1 = -bb[2], x1 = -bb[3];
    const int ry0 = y0 - 1, rx0 = x0 - 1, R = y1 - y0 + 3, Cpx = x1 - x0 + 3, Cw = (Cpx + 31) >> 5;
    const size_t words = (size_t)R * Cw;
    if (words * 4 > smem_bytes) { if (tid == 0) atomicOr(status, 4); return; }          // bounding box beyond what one SM holds
    const bool free_in_smem = 2 * words * 4 <= smem_bytes;      // else `free` is rebuilt from the dense map on every use
    uint32_t *reach = bsm, *freeb = bsm + words;
    const uint8_t *dm = dense + ((size_t)which * B + b) * H * W;
    const uint8_t lab = (uint8_t)(k + 1);
    auto free_word = [&](int r, int w) -> uint32_t {            // built from the dense map (global / L2) when it is not staged
        const int y = ry0 + r;
        uint32_t f = 0;
        for (int c = 0; c < 32; ++c) {
            const int cx = 32 * w + c;
            if (cx >= Cpx) break;
            const int x = rx0 + cx;
            const bool inside = y >= 0 && y < H && x >= 0 && x < W;
            if (!inside || dm[(size_t)y * W + x] != lab) f |= 1u << c;
        }
        return f;
    };
    for (size_t i = tid; i < words; i += 256) {
        const int r = (int)(i / Cw), w = (int)(i % Cw);
        const uint32_t f = free_word(r, w);
        if (free_in_smem) freeb[i] = f;
        uint32_t ring = 0;
        if (r == 0 || r == R - 1) ring = 0xffffffffu;
        else {
            if (w == 0) ring |= 1u;
            if (w == (Cpx - 1) >> 5) ring |= 1u << ((Cpx - 1) & 31);
        }
        reach[i] = f & ring;
    }
    __syncthreads();
    auto FREE = [&](int r, int w) -> uint32_t { return free_in_smem ? freeb[(size_t)r * Cw + w] : free_word(r, w); };
    // alternate full-row sweeps (one thread per row, left->right then right->left with carries) and full-column sweeps
    // (one thread per word column, top->bottom then bottom->top) until nothing changes
    for (int iter = 0; iter < 4096; ++iter) {
        int changed = 0;
        for (int r = tid; r < R; r += 256) {
            uint32_t carry = 0;
            for (int w = 0; w < Cw; ++w) {                     // left -> right: bit c reaches bit c+1 through free pixels
                const uint32_t f = FREE(r, w);
                uint32_t x = reach[(size_t)r * Cw + w];
                const uint32_t x0_ = x;
                x |= f & carry;                                 // carry into bit 0
                uint32_t m = f;
                x |= m & (x << 1); m &= m << 1;
                x |= m & (x << 2); m &= m << 2;
                x |= m & (x << 4); m &= m << 4;
                x |= m & (x << 8); m &= m << 8;
                x |= m & (x << 16);
                carry = x >> 31;
                if (x != x0_) { reach[(size_t)r * Cw + w] = x; changed = 1; }
            }
            carry = 0;
            for (int w = Cw - 1; w >= 0; --w) {                 // right -> left
                const uint32_t f = FREE(r, w);
                uint32_t x = reach[(size_t)r * Cw + w];
                const uint32_t x0_ = x;
                x |= f & (carry << 31);
                uint32_t m = f;
                x |= m & (x >> 1); m &= m >> 1;
                x |= m & (x >> 2); m &= m >> 2;
                x |= m & (x >> 4); m &= m >> 4;
                x |= m & (x >> 8); m &= m >> 8;
                x |= m & (x >> 16);
                carry = x & 1u;
                if (x != x0_) { reach[(size_t)r * Cw + w] = x; changed = 1; }
            }
        }
        __syncthreads();
        for (int w = tid; w < Cw; w += 256) {
            uint32_t prev = reach[w];
            for (int r = 1; r < R; ++r) {                       // top -> bottom
                uint32_t x = reach[(size_t)r * Cw + w];
                const uint32_t n = x | (FREE(r, w) & prev);
                if (n != x) { reach[(size_t)r * Cw + w] = n; changed = 1; }
                prev = n;
            }
            for (int r = R - 2; r >= 0; --r) {                  // bottom -> top
                uint32_t x = reach[(size_t)r * Cw + w];
                const uint32_t n = x | (FREE(r, w) & prev);
                if (n != x) { reach[(size_t)r * Cw + w] = n; changed = 1; }
                prev = n;
            }
        }
        if (!__syncthreads_or(changed)) break;
    }
    // boundary = mask pixels with a reachable 4-neighbour (region rows 1 .. R-2 hold the bounding box)
    int count = 0;
    uint8_t *out = bmap + ((size_t)which * B + b) * H * W;
    for (size_t i = tid; i < words; i += 256) {
        const int r = (int)(i / Cw), w = (int)(i % Cw);
        if (r == 0 || r == R - 1) continue;
        const uint32_t cur = reach[i];
        const uint32_t lft = (cur << 1) | (w > 0 ? reach[i - 1] >> 31 : 0u);
        const uint32_t rgt = (cur >> 1) | (w + 1 < Cw ? reach[i + 1] << 31 : 0u);
        const uint32_t nb = reach[i - Cw] | reach[i + Cw] | lft | rgt;
        uint32_t valid = 0xffffffffu;
        if (32 * w + 32 > Cpx) valid = (Cpx - 32 * w) > 0 ? ((1u << (Cpx - 32 * w)) - 1u) : 0u;
        uint32_t bnd = ~FREE(r, w) & nb & valid;
        count += __popc(bnd);
        while (bnd) {
            const int c = __ffs(bnd) - 1;
            bnd &= bnd - 1;
            out[(size_t)(ry0 + r) * W + rx0 + 32 * w + c] = lab;
        }
    }
#pragma unroll
    for (int o = 16; o; o >>= 1) count += __shfl_xor_sync(0xffffffffu, count, o);
    if ((tid & 31) == 0 && count) atomicAdd(&(which ? bcount_gt : bcount_pred)[(size_t)b * ML + k], count);
}

// boundary_overlap of every (gt i, pred j) pair: btp[b][i][j][0] += predicted boundary pixels of j with a boundary pixel of i
// inside the disk; [1] the other way round.
__global__ void __launch_bounds__(256)
eval_match_kernel(const uint8_t *__restrict__ bmap, int H, int W, int B, int ML, int radius, int32_t *__restrict__ btp) {
    const int b = blockIdx.z;
    const int x = blockIdx.x * 64 + (threadIdx.x & 63), y = blockIdx.y * 4 + (threadIdx.x >> 6);
    if (x >= W || y >= H) return;
    const size_t HW = (size_t)H * W;
    const uint8_t *Bp = bmap + (size_t)b * HW, *Bg = bmap + ((size_t)B + b) * HW;
    const int pj = Bp[(size_t)y * W + x], gi = Bg[(size_t)y * W + x];
    if (!(pj | gi)) return;
    uint32_t near_g[8], near_p[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) near_g[i] = near_p[i] = 0;
    const int r2 = radius * radius;
    for (int dy = -radius; dy <= radius; ++dy) {
        const int yy = y + dy;
        if (yy < 0 || yy >= H) continue;
        for (int dx = -radius; dx <= radius; ++dx) {
            if (dx * dx + dy * dy > r2) continue;               // skimage.morphology.disk(radius)
            const int xx = x + dx;
            if (xx < 0 || xx >= W) continue;
            if (pj) { const int g = Bg[(size_t)yy * W + xx]; if (g) near_g[(g - 1) >> 5] |= 1u << ((g - 1) & 31); }
            if (gi) { const int p = Bp[(size_t)yy * W + xx]; if (p) near_p[(p - 1) >> 5] |= 1u << ((p - 1) & 31); }
        }
    }
    int32_t *m = btp + (size_t)b * ML * ML * 2;
#pragma unroll
    for (int wd = 0; wd < 8; ++wd) {
        uint32_t bits = near_g[wd];
        while (bits) {
            const int i = 32 * wd + __ffs(bits) - 1;
            bits &= bits - 1;
            atomicAdd(&m[((size_t)i * ML + (pj - 1)) * 2], 1);
        }
        bits = near_p[wd];
        while (bits) {
            const int j = 32 * wd + __ffs(bits) - 1;
            bits &= bits - 1;
            atomicAdd(&m[((size_t)(gi - 1) * ML + j) * 2 + 1], 1);
        }
    }
}

__global__ void eval_fill_kernel(int32_t *p, size_t n, int32_t v) {
    for (size_t i = (size_t)blockIdx.x * 256 + threadIdx.x; i < n; i += (size_t)gridDim.x * 256) p[i] = v;
}

}  // namespace quber

extern "C" size_t quber_eval_workspace_bytes(int B, int H, int W, int max_labels) {
    if (B <= 0 || H <= 0 || W <= 0 || max_labels <= 0 || max_labels > quber::EV_MAXL) return 0;
    return quber::eval_carve(B, H, W, max_labels, reinterpret_cast<void *>(4096)).total;
}

extern "C" int quber_eval_counts(const int32_t *pred, const int32_t *gt, int B, int H, int W, int max_labels, int radius, int boundary,
                                 int32_t *pred_labels, int32_t *n_pred, int32_t *gt_labels, int32_t *n_gt, int32_t *inter_full,
                                 int32_t *bcount_pred, int32_t *bcount_gt, int32_t *btp, uint8_t *bmaps_out, void *workspace,
                                 size_t workspace_bytes, int32_t *status, void *stream) {
    using namespace quber;
    if (!pred || !gt || !pred_labels || !n_pred || !gt_labels || !n_gt || !inter_full || !status) return QUBER_E_INVALID;
    if (B <= 0 || H <= 0 || W <= 0 || max_labels <= 0 || max_labels > EV_MAXL || radius < 0 || radius > 64) return QUBER_E_INVALID;
    if (boundary && (!bcount_pred || !bcount_gt || !btp)) return QUBER_E_INVALID;
    if (B > 65535 || (long long)H * W > 0x7fffffffll) return QUBER_E_UNSUPPORTED;
    if (!workspace || !aligned16(workspace) || workspace_bytes < quber_eval_workspace_bytes(B, H, W, max_labels)) return QUBER_E_WORKSPACE;
    const EvalWs ws = eval_carve(B, H, W, max_labels, workspace);
    cudaStream_t st = (cudaStream_t)stream;
    const int ML = max_labels;
    const size_t HW = (size_t)H * W;
    QUBER_CUDA_TRY(cudaMemsetAsync(status, 0, sizeof(int32_t), st));
    QUBER_CUDA_TRY(cudaMemsetAsync(ws.presence, 0, ws.presence_bytes, st));
    QUBER_CUDA_TRY(cudaMemsetAsync(inter_full, 0, (size_t)B * (ML + 1) * (ML + 1) * sizeof(int32_t), st));
    eval_fill_kernel<<<64, 256, 0, st>>>(ws.bbox, (size_t)2 * B * ML * 4, INT_MAX);
    const int gx = (int)((HW + 256 * 8 - 1) / (256 * 8)) < 64 ? (int)((HW + 256 * 8 - 1) / (256 * 8)) : 64;
    eval_presence_kernel<<<dim3(gx, B, 2), 256, 0, st>>>(pred, gt, HW, B, ws.presence, status);
    if (int rc = check_launch("eval_presence_kernel")) return rc;
    eval_labels_kernel<<<dim3(B, 2), 256, 0, st>>>(ws.presence, B, ML, pred_labels, n_pred, gt_labels, n_gt, status);
    if (int rc = check_launch("eval_labels_kernel")) return rc;
    {
        const size_t fixed = (size_t)(2 * ML + 8 * ML) * 4, hist = (size_t)(ML + 1) * (ML + 1) * 4;
        const int in_smem = fixed + hist <= 96 * 1024;
        const size_t smem = fixed + (in_smem ? hist : 0);
        if (smem > 48 * 1024)
            QUBER_CUDA_TRY(cudaFuncSetAttribute(eval_dense_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        eval_dense_kernel<<<dim3(gx, B), 256, smem, st>>>(pred, gt, H, W, B, ML, pred_labels, n_pred, gt_labels, n_gt, ws.dense, ws.bbox,
                                                          inter_full, in_smem);
        if (int rc = check_launch("eval_dense_kernel")) return rc;
    }
    if (!boundary) return QUBER_OK;
    QUBER_CUDA_TRY(cudaMemsetAsync(ws.bmap, 0, ws.dense_bytes, st));
    QUBER_CUDA_TRY(cudaMemsetAsync(bcount_pred, 0, (size_t)B * ML * sizeof(int32_t), st));
    QUBER_CUDA_TRY(cudaMemsetAsync(bcount_gt, 0, (size_t)B * ML * sizeof(int32_t), st));
    QUBER_CUDA_TRY(cudaMemsetAsync(btp, 0, (size_t)B * ML * ML * 2 * sizeof(int32_t), st));
    {
        // worst-case region = the whole frame grown by one pixel; both bit planes in shared memory when they fit
        const size_t words = (size_t)(H + 2) * ((W + 2 + 31) / 32);
        size_t smem = 2 * words * 4;                            // both bit planes of a full-frame box ...
        if (smem > 200 * 1024) smem = 200 * 1024;               // ... or what an SM has: larger boxes keep only `reach` there (status bit 2 if even that does not fit)
        QUBER_CUDA_TRY(cudaFuncSetAttribute(eval_boundary_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        eval_boundary_kernel<<<dim3(ML, 2, B), 256, smem, st>>>(ws.dense, ws.bbox, n_pred, n_gt, H, W, B, ML, (unsigned)smem, ws.bmap,
                                                                bcount_pred, bcount_gt, status);
        if (int rc = check_launch("eval_boundary_kernel")) return rc;
    }
    eval_match_kernel<<<dim3(cdiv(W, 64), cdiv(H, 4), B), 256, 0, st>>>(ws.bmap, H, W, B, ML, radius, btp);
    if (int rc = check_launch("eval_match_kernel")) return rc;
    if (bmaps_out) QUBER_CUDA_TRY(cudaMemcpyAsync(bmaps_out, ws.bmap, ws.dense_bytes, cudaMemcpyDeviceToDevice, st));
    return QUBER_OK;
}
