// K1: multi-scale deformable attention, forward.
//
// Reference: OPS/src/cuda/ms_deform_im2col_cuda.cuh:242-304 (ms_deformable_im2col_gpu_kernel) with the
// bilinear tap of :38-89, host OPS/src/cuda/ms_deform_attn_cuda.cu:25-85, Python
// OPS/functions/ms_deform_attn_func.py:32-39 (OPS = ext_modules/UnseenObjectsWithMeanShift/MSMFormer/
// meanshiftformer/modeling/pixel_decoder/ops).
//
// The reference runs one thread per output scalar: 48 scalar gathers each and the sampling
// locations re-read (and the bilinear set-up recomputed) by all D channel threads.  Here D/4 lanes own
// one (query, head): every corner of every tap is ONE 128-bit load per lane (a head's D floats are
// contiguous, value is [B,S,M,D]), the set-up of a tap is done once and warp-shuffled to the group,
// and a CTA works on 32 CONSECUTIVE queries of ONE head so that overlapping bilinear footprints of
// neighbouring queries hit in L1.  Accumulation order over (level, point) and the w1..w4 arithmetic are the
// reference's, so fp32 results agree with it to rounding.
#include <cuda_bf16.h>

#include "common.cuh"

namespace quber {

constexpr int MSDA_MAX_L = 8;

template <typename VT> struct Vec4;
template <> struct Vec4<float> {
    static __device__ __forceinline__ float4 load(const float *p) { return __ldg(reinterpret_cast<const float4 *>(p)); }
    static __device__ __forceinline__ void store(float *p, float4 v) { st_stream_f4(p, v); }
};
template <> struct Vec4<__nv_bfloat16> {
    static __device__ __forceinline__ float4 load(const __nv_bfloat16 *p) {
        const uint2 r = __ldg(reinterpret_cast<const uint2 *>(p));
        return make_float4(__uint_as_float(r.x << 16), __uint_as_float(r.x & 0xffff0000u), __uint_as_float(r.y << 16),
                           __uint_as_float(r.y & 0xffff0000u));
    }
    static __device__ __forceinline__ void store(__nv_bfloat16 *p, float4 v) {
        const __nv_bfloat162 a = __floats2bfloat162_rn(v.x, v.y), b = __floats2bfloat162_rn(v.z, v.w);
        st_stream_u2(p, make_uint2(*reinterpret_cast<const uint32_t *>(&a), *reinterpret_cast<const uint32_t *>(&b)));
    }
};

// D/4 lanes per (query, head); blockDim.x = 256; blockIdx = (query block, head, batch).
//
// Two phases per round of LPG taps.  SETUP: lane t of the group prepares tap (s0 + t) — loads its
// (x, y, weight), does the floor / bounds / corner-weight arithmetic ONCE (the reference repeats it in
// all D channel threads) and turns the four corners into clamped element offsets with zeroed weights
// for corners outside the map (== the reference's `v = 0`, cuh:58-81, without predicated loads).
// GATHER: the prepared tap is broadcast inside the group with width-LPG shuffles and every lane issues
// four unconditional 128-bit loads and the reference's w1..w4 arithmetic on its four channels.
template <int D, typename VT>
__global__ void __launch_bounds__(256, 2)
msda_forward_kernel(const VT *__restrict__ value, const int64_t *__restrict__ shapes, const int64_t *__restrict__ starts,
                    const float *__restrict__ loc, const float *__restrict__ attn, int S, int M, int L, int Q, int P,
                    VT *__restrict__ out) {
    constexpr int LPG = D / 4;              // lanes per group
    constexpr int GPB = 256 / LPG;          // (query, head) groups per CTA
    const int lane = threadIdx.x % LPG, grp = threadIdx.x / LPG;
    const int m = blockIdx.y, b = blockIdx.z;
    int q = blockIdx.x * GPB + grp;
    const bool active = q < Q;
    if (!active) q = Q - 1;                 // stay in the shuffles; nothing is stored
    const size_t qm = ((size_t)b * Q + q) * M + m;
    const int LP = L * P;
    const float2 *lp = reinterpret_cast<const float2 *>(loc) + qm * LP;
    const float *ap = attn + qm * LP;
    const int MD = M * D;
    const VT *vb = value + (size_t)b * S * MD + m * D + lane * 4;
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int s0 = 0; s0 < LP; s0 += LPG) {
        // ---- setup: this lane's tap ----
        const int tap = s0 + lane;
        int o1 = 0, o3 = 0;
        float w1 = 0.f, w2 = 0.f, w3 = 0.f, w4 = 0.f, a = 0.f;
        if (tap < LP) {
            int l = 0;
            for (int i = 1; i < L; ++i) l += tap >= i * P;                       // tap / P without an integer division
            const int h = (int)__ldg(shapes + 2 * l), w = (int)__ldg(shapes + 2 * l + 1);
            const int start = (int)__ldg(starts + l);
            const float2 xy = __ldg(lp + tap);
            const float him = xy.y * h - 0.5f, wim = xy.x * w - 0.5f;            // cuh:290-291
            if (him > -1.f && wim > -1.f && him < (float)h && wim < (float)w) {  // cuh:293
                a = __ldg(ap + tap);
                const int hl = (int)floorf(him), wl = (int)floorf(wim);
                const float lh = him - hl, lw = wim - wl, hh = 1.f - lh, hw = 1.f - lw;
                const bool y0 = hl >= 0, y1 = hl + 1 <= h - 1, x0 = wl >= 0, x1 = wl + 1 <= w - 1;
                const int ry0 = max(hl, 0), ry1 = min(hl + 1, h - 1), cx0 = max(wl, 0);
                o1 = ((start + ry0 * w + cx0) * MD) | (int)(x0 && x1);           // bit 0: the x+1 corner is a different pixel
                o3 = (start + ry1 * w + cx0) * MD;
                w1 = (y0 && x0) ? hh * hw : 0.f;                                 // cuh:85 (w1..w4), 0 where the reference reads v = 0
                w2 = (y0 && x1) ? hh * lw : 0.f;
                w3 = (y1 && x0) ? lh * hw : 0.f;
                w4 = (y1 && x1) ? lh * lw : 0.f;
            }
        }
        // ---- gather: all lanes of the group walk the prepared taps in the reference's (l, p) order ----
        const int nt = min(LPG, LP - s0);
#pragma unroll
        for (int j = 0; j < LPG; ++j) {
            if (j < nt) {
                int bo1 = __shfl_sync(0xffffffffu, o1, j, LPG);
                const int bo3 = __shfl_sync(0xffffffffu, o3, j, LPG);
                const float b1 = __shfl_sync(0xffffffffu, w1, j, LPG), b2 = __shfl_sync(0xffffffffu, w2, j, LPG);
                const float b3 = __shfl_sync(0xffffffffu, w3, j, LPG), b4 = __shfl_sync(0xffffffffu, w4, j, LPG);
                const float ba = __shfl_sync(0xffffffffu, a, j, LPG);
                const int dx = (bo1 & 1) ? MD : 0;
                bo1 &= ~1;
                const float4 v1 = Vec4<VT>::load(vb + bo1), v2 = Vec4<VT>::load(vb + bo1 + dx);
                const float4 v3 = Vec4<VT>::load(vb + bo3), v4 = Vec4<VT>::load(vb + bo3 + dx);
                acc.x += (b1 * v1.x + b2 * v2.x + b3 * v3.x + b4 * v4.x) * ba;   // cuh:87,295
                acc.y += (b1 * v1.y + b2 * v2.y + b3 * v3.y + b4 * v4.y) * ba;
                acc.z += (b1 * v1.z + b2 * v2.z + b3 * v3.z + b4 * v4.z) * ba;
                acc.w += (b1 * v1.w + b2 * v2.w + b3 * v3.w + b4 * v4.w) * ba;
            }
        }
    }
    if (active) Vec4<VT>::store(out + qm * D + lane * 4, acc);
}

// Any D (the reference's own mapping: one thread per output scalar), fp32 only.
__global__ void __launch_bounds__(256)
msda_forward_scalar_kernel(const float *__restrict__ value, const int64_t *__restrict__ shapes,
                           const int64_t *__restrict__ starts, const float *__restrict__ loc, const float *__restrict__ attn,
                           size_t n, int S, int M, int D, int L, int Q, int P, float *__restrict__ out) {
    const size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= n) return;
    const int c = idx % D;
    const size_t qm = idx / D;
    const int m = qm % M;
    const size_t b = qm / M / Q;
    const int MD = M * D;
    float col = 0.f;
    for (int l = 0; l < L; ++l) {
        const int h = (int)shapes[2 * l], w = (int)shapes[2 * l + 1];
        const float *v = value + (b * S + (size_t)starts[l]) * MD + m * D + c;
        for (int pt = 0; pt < P; ++pt) {
            const size_t lpi = (qm * L + l) * P + pt;
            const float lw_ = loc[2 * lpi], lh_ = loc[2 * lpi + 1], a = attn[lpi];
            const float him = lh_ * h - 0.5f, wim = lw_ * w - 0.5f;
            if (!(him > -1.f && wim > -1.f && him < (float)h && wim < (float)w)) continue;
            const int hl = (int)floorf(him), wl = (int)floorf(wim);
            const float lh = him - hl, lw = wim - wl, hh = 1.f - lh, hw = 1.f - lw;
            float v1 = 0, v2 = 0, v3 = 0, v4 = 0;
            if (hl >= 0 && wl >= 0) v1 = v[((size_t)hl * w + wl) * MD];
            if (hl >= 0 && wl + 1 <= w - 1) v2 = v[((size_t)hl * w + wl + 1) * MD];
            if (hl + 1 <= h - 1 && wl >= 0) v3 = v[((size_t)(hl + 1) * w + wl) * MD];
            if (hl + 1 <= h - 1 && wl + 1 <= w - 1) v4 = v[((size_t)(hl + 1) * w + wl + 1) * MD];
            col += (hh * hw * v1 + hh * lw * v2 + lh * hw * v3 + lh * lw * v4) * a;
        }
    }
    out[idx] = col;
}

template <int D, typename VT>
static int launch_vec(const void *value, const int64_t *shapes, const int64_t *starts, const float *loc, const float *attn,
                      int B, int S, int M, int L, int Q, int P, void *out, cudaStream_t st) {
    constexpr int GPB = 256 / (D / 4);
    dim3 grid(cdiv(Q, GPB), M, B);
    msda_forward_kernel<D, VT><<<grid, 256, 0, st>>>(static_cast<const VT *>(value), shapes, starts, loc, attn, S, M, L, Q, P,
                                                    static_cast<VT *>(out));
    return check_launch("msda_forward_kernel");
}

}  // namespace quber

extern "C" int quber_msda_forward(const void *value, const int64_t *shapes, const int64_t *starts, const float *loc,
                                  const float *attn, int B, int S, int M, int D, int L, int Q, int P, int dtype, void *out,
                                  void *stream) {
    using namespace quber;
    if (!value || !shapes || !starts || !loc || !attn || !out) return QUBER_E_INVALID;
    if (B <= 0 || S <= 0 || M <= 0 || D <= 0 || L <= 0 || L > MSDA_MAX_L || Q <= 0 || P <= 0) return QUBER_E_INVALID;
    if (dtype != 0 && dtype != 1) return QUBER_E_INVALID;
    if (M > 65535 || B > 65535) return QUBER_E_UNSUPPORTED;
    cudaStream_t st = (cudaStream_t)stream;
    const bool vec_ok = aligned16(value) && aligned16(out) && ((reinterpret_cast<uintptr_t>(loc) & 7u) == 0);
    if (dtype == 1) {
        if (!vec_ok) return QUBER_E_ALIGN;
        switch (D) {
            case 8: return launch_vec<8, __nv_bfloat16>(value, shapes, starts, loc, attn, B, S, M, L, Q, P, out, st);
            case 16: return launch_vec<16, __nv_bfloat16>(value, shapes, starts, loc, attn, B, S, M, L, Q, P, out, st);
            case 32: return launch_vec<32, __nv_bfloat16>(value, shapes, starts, loc, attn, B, S, M, L, Q, P, out, st);
            case 64: return launch_vec<64, __nv_bfloat16>(value, shapes, starts, loc, attn, B, S, M, L, Q, P, out, st);
            default: return QUBER_E_UNSUPPORTED;
        }
    }
    if (vec_ok) {
        switch (D) {
            case 8: return launch_vec<8, float>(value, shapes, starts, loc, attn, B, S, M, L, Q, P, out, st);
            case 16: return launch_vec<16, float>(value, shapes, starts, loc, attn, B, S, M, L, Q, P, out, st);
            case 32: return launch_vec<32, float>(value, shapes, starts, loc, attn, B, S, M, L, Q, P, out, st);
            case 64: return launch_vec<64, float>(value, shapes, starts, loc, attn, B, S, M, L, Q, P, out, st);
            default: break;
        }
    }
    const size_t n = (size_t)B * Q * M * D;
    msda_forward_scalar_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(
        static_cast<const float *>(value), shapes, starts, loc, attn, n, S, M, D, L, Q, P, static_cast<float *>(out));
    return check_launch("msda_forward_scalar_kernel");
}
