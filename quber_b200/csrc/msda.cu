// K1: multi-scale deformable attention, forward.
//
// Reference: OPS/src/cuda/ms_deform_im2col_cuda.cuh:242-304 (ms_deformable_im2col_gpu_kernel) with the
// bilinear tap of :38-89, host OPS/src/cuda/ms_deform_attn_cuda.cu:25-85, Python
// OPS/functions/ms_deform_attn_func.py:32-39 (OPS = ext_modules/UnseenObjectsWithMeanShift/MSMFormer/
// meanshiftformer/modeling/pixel_decoder/ops).
//
// The reference runs one thread per output scalar: 48 scalar gathers each and the sampling
// locations re-read by all D channel threads.  Here D/4 lanes own one (query, head): every corner of
// every tap is ONE 128-bit load per lane (a head's D floats are contiguous, value is [B,S,M,D]), the
// (x, y, weight) triple of a tap is a broadcast load shared by the lane group, and a CTA works on
// 32 CONSECUTIVE queries of ONE head so that overlapping bilinear footprints of neighbouring
// queries hit in L1.  Accumulation order over (level, point) and the w1..w4 arithmetic are the
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
template <int D, typename VT>
__global__ void __launch_bounds__(256)
msda_forward_kernel(const VT *__restrict__ value, const int64_t *__restrict__ shapes, const int64_t *__restrict__ starts,
                    const float *__restrict__ loc, const float *__restrict__ attn, int S, int M, int L, int Q, int P,
                    VT *__restrict__ out) {
    constexpr int LPG = D / 4;              // lanes per group
    constexpr int GPB = 256 / LPG;          // (query, head) groups per CTA
    const int lane = threadIdx.x % LPG, grp = threadIdx.x / LPG;
    const int m = blockIdx.y, b = blockIdx.z;
    const int q = blockIdx.x * GPB + grp;
    if (q >= Q) return;
    const size_t qm = ((size_t)b * Q + q) * M + m;
    const float *lp = loc + qm * L * P * 2;
    const float *ap = attn + qm * L * P;
    const int MD = M * D;
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int l = 0; l < L; ++l) {
        const int h = (int)__ldg(shapes + 2 * l), w = (int)__ldg(shapes + 2 * l + 1);
        const VT *vbase = value + ((size_t)b * S + (size_t)__ldg(starts + l)) * MD + m * D + lane * 4;
#pragma unroll 4
        for (int pt = 0; pt < P; ++pt) {
            const float2 xy = __ldg(reinterpret_cast<const float2 *>(lp) + l * P + pt);
            const float a = __ldg(ap + l * P + pt);
            const float him = xy.y * h - 0.5f, wim = xy.x * w - 0.5f;          // cuh:290-291
            if (him > -1.f && wim > -1.f && him < (float)h && wim < (float)w) { // cuh:293
                const int hl = (int)floorf(him), wl = (int)floorf(wim);
                const float lh = him - hl, lw = wim - wl, hh = 1.f - lh, hw = 1.f - lw;
                const bool y0 = hl >= 0, y1 = hl + 1 <= h - 1, x0 = wl >= 0, x1 = wl + 1 <= w - 1;
                const VT *p00 = vbase + ((size_t)hl * w + wl) * MD;
                float4 v1 = make_float4(0, 0, 0, 0), v2 = v1, v3 = v1, v4 = v1;
                if (y0 && x0) v1 = Vec4<VT>::load(p00);
                if (y0 && x1) v2 = Vec4<VT>::load(p00 + MD);
                if (y1 && x0) v3 = Vec4<VT>::load(p00 + (size_t)w * MD);
                if (y1 && x1) v4 = Vec4<VT>::load(p00 + (size_t)w * MD + MD);
                const float w1 = hh * hw, w2 = hh * lw, w3 = lh * hw, w4 = lh * lw;   // cuh:85
                acc.x += (w1 * v1.x + w2 * v2.x + w3 * v3.x + w4 * v4.x) * a;        // cuh:87,295
                acc.y += (w1 * v1.y + w2 * v2.y + w3 * v3.y + w4 * v4.y) * a;
                acc.z += (w1 * v1.z + w2 * v2.z + w3 * v3.z + w4 * v4.z) * a;
                acc.w += (w1 * v1.w + w2 * v2.w + w3 * v3.w + w4 * v4.w) * a;
            }
        }
    }
    Vec4<VT>::store(out + qm * D + lane * 4, acc);
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
