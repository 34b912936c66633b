// K1: multi-scale deformable attention, forward.
//
// Reference: OPS/src/cuda/ms_deform_im2col_cuda.cuh:242-304 (ms_deformable_im2col_gpu_kernel) with the
// bilinear tap of :38-89, host OPS/src/cuda/ms_deform_attn_cuda.cu:25-85, Python
// OPS/functions/ms_deform_attn_func.py:32-39 (OPS = ext_modules/UnseenObjectsWithMeanShift/MSMFormer/
// meanshiftformer/modeling/pixel_decoder/ops).
//
// The reference runs one thread per output scalar: 48 scalar gathers each and the sampling
// locations re-read (and the bilinear set-up recomputed) by all D channel threads.  Here a group of
// D/CPL lanes owns one (query, head) and every corner of every tap is ONE 128-bit load per lane (a head's
// D channels are contiguous, value is [B,S,M,D]; CPL = 4 fp32 or 8 bf16 channels per lane).  A CTA
// works on 256/(D/CPL) CONSECUTIVE queries of ONE head so that overlapping bilinear footprints of
// neighbouring queries hit in L1 (ncu: 68 % L1 sector hit rate, value read from DRAM exactly once).
//
// What bounds it (ncu --set full, profiles/): the L1 data pipe, not HBM.  Every gathered corner is a
// 128-byte line that must be written back to registers at 128 B/clk/SM: 48 lines x B*Q*M = 9.9 GB per
// launch at C2 = 266 us at 1.965 GHz, before the per-tap broadcast of the set-up (one more write-back
// cycle per 32-bit value).  l1tex__lsu_writeback_active sits at 72 % and l1tex__data_pipe_lsu_wavefronts
// at 75 %; 54 % of warp samples wait in long_scoreboard on that pipe.  Measured alternatives, all slower
// at C2 (fp32, us/launch): two of the four corners through the texture pipe (tex1Dfetch of raw 16-byte texels on a
// linear texture of `value`) 507 against 497, all four 665 — TEX shares the L1 data stage, it is not a second path;
// set-up recomputed by every lane 594; 7-value shuffle broadcast 568; 16-byte
// shared-memory tap records read with LDS.128 648; 4-value broadcast + per-lane weights 587; register
// software pipelining of the corner loads 681 (3 CTAs/SM) and 903 (2 CTAs/SM) — throughput follows
// resident warps, not loads in flight; 2-D query patches 585.  Kept: the variant with the fewest
// instructions per gathered byte (558).
#include <cuda_bf16.h>

#include "common.cuh"

namespace quber {

constexpr int MSDA_MAX_L = 8;

template <typename VT> struct Lane;     // CPL channels of one pixel of one head, as ONE 128-bit access
template <> struct Lane<float> {
    static constexpr int CPL = 4;
    float c[4];
    __device__ __forceinline__ void load(const float *p) {
        const float4 v = __ldg(reinterpret_cast<const float4 *>(p));
        c[0] = v.x; c[1] = v.y; c[2] = v.z; c[3] = v.w;
    }
    static __device__ __forceinline__ void store(float *p, const float *a) { st_stream_f4(p, make_float4(a[0], a[1], a[2], a[3])); }
    static __device__ __forceinline__ uint4 load_raw(const float *p) { return __ldg(reinterpret_cast<const uint4 *>(p)); }
    static __device__ __forceinline__ void unpack(const uint4 &r, float *c) {
        c[0] = __uint_as_float(r.x); c[1] = __uint_as_float(r.y); c[2] = __uint_as_float(r.z); c[3] = __uint_as_float(r.w);
    }
};
template <> struct Lane<__nv_bfloat16> {
    static constexpr int CPL = 8;
    float c[8];
    __device__ __forceinline__ void load(const __nv_bfloat16 *p) {
        const uint4 r = __ldg(reinterpret_cast<const uint4 *>(p));
        const uint32_t w[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            c[2 * i] = __uint_as_float(w[i] << 16);
            c[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u);
        }
    }
    static __device__ __forceinline__ void store(__nv_bfloat16 *p, const float *a) {
        uint32_t w[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const __nv_bfloat162 h = __floats2bfloat162_rn(a[2 * i], a[2 * i + 1]);
            w[i] = *reinterpret_cast<const uint32_t *>(&h);
        }
        st_stream_u4(p, make_uint4(w[0], w[1], w[2], w[3]));
    }
    static __device__ __forceinline__ uint4 load_raw(const __nv_bfloat16 *p) { return __ldg(reinterpret_cast<const uint4 *>(p)); }
    static __device__ __forceinline__ void unpack(const uint4 &r, float *c) {
        const uint32_t w[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            c[2 * i] = __uint_as_float(w[i] << 16);
            c[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u);
        }
    }
};

// D/CPL lanes per (query, head); blockDim.x = 256; blockIdx = (query block, head, batch).
//
// Two phases per round of LPG taps.  SETUP: lane t of the group prepares tap (s0 + t) — its (x, y, weight)
// were fetched one round ahead with streaming loads — doing the floor / bounds / corner-weight arithmetic
// ONCE (the reference repeats it in all D channel threads): clamped element offsets, and the four corner
// weights already multiplied by the attention weight, zero for corners outside the map (== the
// reference's `v = 0`, cuh:58-81, without predicated loads).  GATHER: the prepared tap is broadcast inside
// the group with six width-LPG shuffles and every lane issues four unconditional 128-bit loads and four FMAs
// per channel.  (w_i * a) * v_i differs from the reference's (sum_i w_i v_i) * a by rounding only
// (measured: <= 6e-9 abs on the C2 inputs against the oracle, tolerance 1e-3).
template <int D, typename VT, int MINB, bool LDSB>
__global__ void __launch_bounds__(256, MINB)
msda_forward_kernel(const VT *__restrict__ value, const int64_t *__restrict__ shapes, const int64_t *__restrict__ starts,
                          const float *__restrict__ loc, const float *__restrict__ attn, int S, int M, int L, int Q, int P,
                          VT *__restrict__ out) {
    constexpr int CPL = Lane<VT>::CPL;
    constexpr int LPG = D / CPL;
    constexpr int GPB = 256 / LPG;
    __shared__ int s_h[MSDA_MAX_L], s_w[MSDA_MAX_L], s_start[MSDA_MAX_L];
    // LDSB: the prepared taps are broadcast through shared memory (one LDS.128 + one LDS.64 per tap, every lane of a
    // group reading the same address) instead of six shuffles — fewer L1 data-pipe wavefronts, the limiter of this
    // kernel (fp32, D = 32: 527 -> 497 us at C2).  Per warp and group g: recA[g * (LPG + 1) + t] = (o1, o3, w1, w2),
    // recB[g * (LPG + 1) + t] = (w3, w4); the +1 pad keeps the per-group broadcast addresses in distinct banks (an unpadded
    // recB cost one bank conflict per tap: ncu l1tex__data_bank_conflicts_pipe_lsu_mem_shared_op_ld, r01 v7).
    constexpr int GPW = 32 / LPG;                       // groups per warp
    __shared__ __align__(16) uint4 s_recA[LDSB ? 8 * GPW * (LPG + 1) : 1];
    __shared__ __align__(8) float2 s_recB[LDSB ? 8 * GPW * (LPG + 1) : 1];
    const int tid = threadIdx.x;
    const int lane = tid % LPG, grp = tid / LPG;
    const int m = blockIdx.y, b = blockIdx.z;
    const int LP = L * P, MD = M * D;
    if (tid < L) {
        s_h[tid] = (int)__ldg(shapes + 2 * tid);
        s_w[tid] = (int)__ldg(shapes + 2 * tid + 1);
        s_start[tid] = (int)__ldg(starts + tid);
    }
    __syncthreads();
    const int qreal = blockIdx.x * GPB + grp;
    const int q = min(qreal, Q - 1);
    const size_t qm = ((size_t)b * Q + q) * M + m;
    const float *lp = loc + qm * LP * 2;
    const float *ap = attn + qm * LP;
    // Gather addresses are 32-bit offsets in 16-byte units (one lane access) from a per-lane 64-bit base, so that each
    // corner costs ONE IMAD.WIDE.U32 instead of a 64-bit add pair.  quber_msda_forward guarantees S*M*D < 2^31.
    const uint4 *vlane = reinterpret_cast<const uint4 *>(value + (size_t)b * S * MD + m * D) + lane;
    asm volatile("" : "+l"(vlane));                     // opaque: keeps (base + lane) in one register pair, see above
    const uint32_t MDu = MD / CPL;                      // pixel pitch in 16-byte units
    float acc[CPL];
#pragma unroll
    for (int c = 0; c < CPL; ++c) acc[c] = 0.f;
    uint2 raw_xy = make_uint2(0u, 0u);
    uint32_t raw_a = 0u;
    if (lane < LP) { raw_xy = ld_stream_u2(lp + 2 * lane); raw_a = ld_stream_u1(ap + lane); }
    for (int s0 = 0; s0 < LP; s0 += LPG) {
        const int tap = s0 + lane;
        uint32_t o1 = 0, o3 = 0;
        float w1 = 0.f, w2 = 0.f, w3 = 0.f, w4 = 0.f;
        if (tap < LP) {
            const int l = tap / P;
            const int h = s_h[l], w = s_w[l];
            const float him = __uint_as_float(raw_xy.y) * h - 0.5f, wim = __uint_as_float(raw_xy.x) * w - 0.5f;   // cuh:290-291
            if (him > -1.f && wim > -1.f && him < (float)h && wim < (float)w) {          // cuh:293
                const float a = __uint_as_float(raw_a);
                const int hl = (int)floorf(him), wl = (int)floorf(wim);
                const float lh = him - hl, lw = wim - wl, hh = 1.f - lh, hw = 1.f - lw;
                const bool y0 = hl >= 0, y1 = hl + 1 <= h - 1, x0 = wl >= 0, x1 = wl + 1 <= w - 1;
                const int ry0 = max(hl, 0), ry1 = min(hl + 1, h - 1), cx0 = max(wl, 0);
                const int start = s_start[l];
                o1 = ((uint32_t)(start + ry0 * w + cx0) * MDu) | ((x0 && x1) ? 0x80000000u : 0u);   // bit 31: the x+1 corner is a different pixel
                o3 = (uint32_t)(start + ry1 * w + cx0) * MDu;
                w1 = (y0 && x0) ? hh * hw * a : 0.f;                                     // cuh:85, 0 where the reference reads v = 0
                w2 = (y0 && x1) ? hh * lw * a : 0.f;
                w3 = (y1 && x0) ? lh * hw * a : 0.f;
                w4 = (y1 && x1) ? lh * lw * a : 0.f;
            }
        }
        if (tap + LPG < LP) { raw_xy = ld_stream_u2(lp + 2 * (tap + LPG)); raw_a = ld_stream_u1(ap + tap + LPG); }
        const int nt = min(LPG, LP - s0);
        uint4 *recA = s_recA + (LDSB ? (tid >> 5) * (GPW * (LPG + 1)) + ((tid & 31) / LPG) * (LPG + 1) : 0);
        float2 *recB = s_recB + (LDSB ? (tid >> 5) * (GPW * (LPG + 1)) + ((tid & 31) / LPG) * (LPG + 1) : 0);
        if (LDSB) {
            recA[lane] = make_uint4(o1, o3, __float_as_uint(w1), __float_as_uint(w2));
            recB[lane] = make_float2(w3, w4);
            __syncwarp();
        }
#pragma unroll
        for (int j = 0; j < LPG; ++j) {
            if (j < nt) {
                uint32_t bo1, bo3;
                float b1, b2, b3, b4;
                if (LDSB) {
                    const uint4 ra = recA[j];
                    const float2 rb = recB[j];
                    bo1 = ra.x; bo3 = ra.y; b1 = __uint_as_float(ra.z); b2 = __uint_as_float(ra.w); b3 = rb.x; b4 = rb.y;
                } else {
                    bo1 = __shfl_sync(0xffffffffu, o1, j, LPG);
                    bo3 = __shfl_sync(0xffffffffu, o3, j, LPG);
                    b1 = __shfl_sync(0xffffffffu, w1, j, LPG); b2 = __shfl_sync(0xffffffffu, w2, j, LPG);
                    b3 = __shfl_sync(0xffffffffu, w3, j, LPG); b4 = __shfl_sync(0xffffffffu, w4, j, LPG);
                }
                const uint32_t dx = (uint32_t)((int32_t)bo1 >> 31) & MDu;
                uint32_t a1 = bo1 & 0x7fffffffu;
                asm volatile("" : "+r"(a1));            // opaque, or the mask is folded into a 64-bit shift/add sequence
                float v1[CPL], v2[CPL], v3[CPL], v4[CPL];
                const uint4 r1 = __ldg(vlane + a1), r2 = __ldg(vlane + (a1 + dx));
                const uint4 r3 = __ldg(vlane + bo3), r4 = __ldg(vlane + (bo3 + dx));
                Lane<VT>::unpack(r1, v1); Lane<VT>::unpack(r2, v2); Lane<VT>::unpack(r3, v3); Lane<VT>::unpack(r4, v4);
#pragma unroll
                for (int c = 0; c < CPL; ++c)
                    acc[c] = fmaf(b4, v4[c], fmaf(b3, v3[c], fmaf(b2, v2[c], fmaf(b1, v1[c], acc[c]))));
            }
        }
        if (LDSB) __syncwarp();
    }
    if (qreal < Q) Lane<VT>::store(out + qm * D + lane * CPL, acc);
}

// P == 4 variant with ONE 16-byte record per tap.  Every 32-bit value a warp broadcasts costs one L1 data-pipe
// wavefront (ncu: 5.95 shared wavefronts per tap step with the 6-value records above), so the record is cut to four
// words: (offset | flags, A, B, f) with
//   corner weights  w1 = A (1 - f), w2 = A f, w3 = B (1 - f), w4 = B f,
//   A = hh a g, B = lh a g (0 for a row outside the map), f = lw when both columns are inside, else 0 with the surviving
//   column's weight folded into g (both loads then read that one pixel), bit 31 = "x + 1 is a different pixel",
//   bit 30 = "y + 1 is a different row" (the row pitch of the tap's level is read once per round of LPG taps).
// The products are associated differently from the reference's hh * hw * v (cuh:85): rounding only.
// VB = bytes per lane and corner: 16 (LDG.128) or 32 (LDG.256, sm_100: a group is D*sizeof(VT)/32 lanes, so one
// broadcast serves twice as many (query, head) pairs per warp and the gather still moves whole 128-byte lines).
template <int VB> __device__ __forceinline__ void ld_corner(const char *p, uint32_t *w);
template <> __device__ __forceinline__ void ld_corner<16>(const char *p, uint32_t *w) {
    const uint4 r = __ldg(reinterpret_cast<const uint4 *>(p));
    w[0] = r.x; w[1] = r.y; w[2] = r.z; w[3] = r.w;
}
template <> __device__ __forceinline__ void ld_corner<32>(const char *p, uint32_t *w) {
    asm volatile("ld.global.nc.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(w[0]), "=r"(w[1]), "=r"(w[2]), "=r"(w[3]), "=r"(w[4]), "=r"(w[5]), "=r"(w[6]), "=r"(w[7]) : "l"(p));
}
template <typename VT, int NW> __device__ __forceinline__ void fma_corner(float wgt, const uint32_t *w, float *acc) {
    if constexpr (sizeof(VT) == 4) {
#pragma unroll
        for (int i = 0; i < NW; ++i) acc[i] = fmaf(wgt, __uint_as_float(w[i]), acc[i]);
    } else {
#pragma unroll
        for (int i = 0; i < NW; ++i) {
            acc[2 * i] = fmaf(wgt, __uint_as_float(w[i] << 16), acc[2 * i]);
            acc[2 * i + 1] = fmaf(wgt, __uint_as_float(w[i] & 0xffff0000u), acc[2 * i + 1]);
        }
    }
}

template <int D, typename VT, int MINB, int VB>
__global__ void __launch_bounds__(256, MINB)
msda_forward_p4_kernel(const VT *__restrict__ value, const int64_t *__restrict__ shapes, const int64_t *__restrict__ starts,
                       const float *__restrict__ loc, const float *__restrict__ attn, int S, int M, int L, int Q,
                       VT *__restrict__ out) {
    constexpr int P = 4;
    constexpr int NW = VB / 4;                          // 32-bit words per lane and corner
    constexpr int CPL = VB / (int)sizeof(VT);
    constexpr int LPG = D / CPL;
    static_assert(LPG >= 1 && LPG <= 32 && (LPG % P == 0 || P % LPG == 0), "a round of LPG taps must stay level-aligned");
    constexpr int NLR = LPG >= P ? LPG / P : 1;         // levels per round
    constexpr int GPB = 256 / LPG;
    constexpr int GPW = 32 / LPG;
    __shared__ int s_h[MSDA_MAX_L], s_w[MSDA_MAX_L], s_start[MSDA_MAX_L];
    __shared__ __align__(16) uint4 s_rec[8 * GPW * (LPG + 1)];
    const int tid = threadIdx.x;
    const int lane = tid % LPG, grp = tid / LPG;
    const int m = blockIdx.y, b = blockIdx.z;
    const int LP = L * P, MD = M * D;
    if (tid < L) {
        s_h[tid] = (int)__ldg(shapes + 2 * tid);
        s_w[tid] = (int)__ldg(shapes + 2 * tid + 1);
        s_start[tid] = (int)__ldg(starts + tid);
    }
    __syncthreads();
    const int qreal = blockIdx.x * GPB + grp;
    const int q = min(qreal, Q - 1);
    const size_t qm = ((size_t)b * Q + q) * M + m;
    const float *lp = loc + qm * LP * 2;
    const float *ap = attn + qm * LP;
    const char *vlane = reinterpret_cast<const char *>(value + (size_t)b * S * MD + m * D) + lane * VB;
    asm volatile("" : "+l"(vlane));                     // opaque: (base + lane) stays one register pair
    const uint32_t MDu = MD / CPL;                      // pixel pitch in VB-byte units
    float acc[CPL];
#pragma unroll
    for (int c = 0; c < CPL; ++c) acc[c] = 0.f;
    uint2 raw_xy = make_uint2(0u, 0u);
    uint32_t raw_a = 0u;
    if (lane < LP) { raw_xy = ld_stream_u2(lp + 2 * lane); raw_a = ld_stream_u1(ap + lane); }
    uint4 *rec = s_rec + (tid >> 5) * (GPW * (LPG + 1)) + ((tid & 31) / LPG) * (LPG + 1);
    for (int s0 = 0; s0 < LP; s0 += LPG) {
        const int tap = s0 + lane;
        uint32_t o1 = 0;
        float A = 0.f, Bw = 0.f, f = 0.f;
        if (tap < LP) {
            const int l = tap / P;
            const int h = s_h[l], w = s_w[l];
            const float him = __uint_as_float(raw_xy.y) * h - 0.5f, wim = __uint_as_float(raw_xy.x) * w - 0.5f;   // cuh:290-291
            if (him > -1.f && wim > -1.f && him < (float)h && wim < (float)w) {          // cuh:293
                const float a = __uint_as_float(raw_a);
                const int hl = (int)floorf(him), wl = (int)floorf(wim);
                const float lh = him - hl, lw = wim - wl, hh = 1.f - lh, hw = 1.f - lw;
                const bool y0 = hl >= 0, y1 = hl + 1 <= h - 1, x0 = wl >= 0, x1 = wl + 1 <= w - 1;
                const int ry0 = max(hl, 0), cx0 = max(wl, 0);
                const bool xb = x0 && x1, yb = y0 && y1;
                const float ag = xb ? a : a * (x0 ? hw : lw);                            // one column left: its weight goes into g
                f = xb ? lw : 0.f;
                A = y0 ? hh * ag : 0.f;                                                  // 0 where the reference reads v = 0 (cuh:58-81)
                Bw = y1 ? lh * ag : 0.f;
                o1 = ((uint32_t)(s_start[l] + ry0 * w + cx0) * MDu) | (xb ? 0x80000000u : 0u) | (yb ? 0x40000000u : 0u);
            }
        }
        if (tap + LPG < LP) { raw_xy = ld_stream_u2(lp + 2 * (tap + LPG)); raw_a = ld_stream_u1(ap + tap + LPG); }
        const int nt = min(LPG, LP - s0);
        uint32_t pitch[NLR];
#pragma unroll
        for (int r = 0; r < NLR; ++r) pitch[r] = (uint32_t)s_w[min(s0 / P + r, L - 1)] * MDu;
        rec[lane] = make_uint4(o1, __float_as_uint(A), __float_as_uint(Bw), __float_as_uint(f));
        __syncwarp();
#pragma unroll
        for (int j = 0; j < LPG; ++j) {
            if (j < nt) {
                const uint4 ra = rec[j];
                const uint32_t dx = (uint32_t)((int32_t)ra.x >> 31) & MDu;
                const uint32_t dy = (uint32_t)((int32_t)(ra.x << 1) >> 31) & pitch[j / P];
                uint32_t a1 = ra.x & 0x3fffffffu;
                asm volatile("" : "+r"(a1));
                const uint32_t a3 = a1 + dy;
                const float bf = __uint_as_float(ra.w), bg = 1.f - bf;
                const float b1 = __uint_as_float(ra.y) * bg, b2 = __uint_as_float(ra.y) * bf;
                const float b3 = __uint_as_float(ra.z) * bg, b4 = __uint_as_float(ra.z) * bf;
                uint32_t r1[NW], r2[NW], r3[NW], r4[NW];
                ld_corner<VB>(vlane + (size_t)a1 * VB, r1);
                ld_corner<VB>(vlane + (size_t)(a1 + dx) * VB, r2);
                ld_corner<VB>(vlane + (size_t)a3 * VB, r3);
                ld_corner<VB>(vlane + (size_t)(a3 + dx) * VB, r4);
                fma_corner<VT, NW>(b1, r1, acc); fma_corner<VT, NW>(b2, r2, acc);
                fma_corner<VT, NW>(b3, r3, acc); fma_corner<VT, NW>(b4, r4, acc);
            }
        }
        __syncwarp();
    }
    if (qreal < Q) {
        VT *o = out + qm * D + lane * CPL;
        if constexpr (sizeof(VT) == 4) {
#pragma unroll
            for (int c = 0; c < CPL; c += 4) st_stream_f4(o + c, make_float4(acc[c], acc[c + 1], acc[c + 2], acc[c + 3]));
        } else {
#pragma unroll
            for (int c = 0; c < CPL; c += 8) Lane<__nv_bfloat16>::store(o + c, acc + c);
        }
    }
}


// ------------------------------------------------------------------------------------------------
// MSDeformAttn.forward between its Linear layers and the output projection in ONE kernel
// (OPS/modules/ms_deform_attn.py:102-117 + OPS/src/cuda/ms_deform_im2col_cuda.cuh:242-304): the set-up phase of the
// one-record kernel above reads the fused sampling_offsets | attention_weights GEMM output row directly
//   oa [B*Q, ld] = [offsets (M, L, P, 2) | logits (M, L, P)]
// does the softmax over the L*P logits of its (query, head) inside the lane group (two shuffle reductions) and
//   loc = reference_points[:, :, None, :, None, :] + offsets / (W_l, H_l)                       (:106-109)
// with the module's own arithmetic (IEEE division, then the add), then samples as above.  The separate preparation
// pass (quber_msda_prepare: 464 MB of traffic and 110 us per layer at C2) and the loc / attn tensors disappear; the
// bytes this kernel reads are the same (288 floats per row either way).
// ------------------------------------------------------------------------------------------------
struct LevelTab { int h[4], w[4], start[4]; };

template <int D, typename VT, int MINB, int L, bool CLV, bool ALLREC = false>
__global__ void __launch_bounds__(256, MINB)
msda_fused_p4_kernel(const VT *__restrict__ value, const int64_t *__restrict__ shapes, const int64_t *__restrict__ starts,
                     const float *__restrict__ oa, int ld, const float *__restrict__ ref, int S, int M, int Q,
                     VT *__restrict__ out, const __grid_constant__ LevelTab lv) {
    constexpr int P = 4, VB = 16;
    constexpr int NW = VB / 4;
    constexpr int CPL = VB / (int)sizeof(VT);
    constexpr int LPG = D / CPL;
    static_assert(LPG >= 1 && LPG <= 32 && (LPG % P == 0 || P % LPG == 0), "a round of LPG taps must stay level-aligned");
    constexpr int LP = L * P;
    constexpr int NR = (LP + LPG - 1) / LPG;            // rounds of LPG taps
    constexpr int NLR = LPG >= P ? LPG / P : 1;         // levels per round
    constexpr int GPB = 256 / LPG;
    constexpr int GPW = 32 / LPG;
    __shared__ int s_h[MSDA_MAX_L], s_w[MSDA_MAX_L], s_start[MSDA_MAX_L];
    constexpr int RPG = ALLREC ? LP + 1 : LPG + 1;      // record slots per group (+1: bank spread)
    __shared__ __align__(16) uint4 s_rec[8 * GPW * RPG];
    const int tid = threadIdx.x;
    const int lane = tid % LPG, grp = tid / LPG;
    const int m = blockIdx.y, b = blockIdx.z;
    const int MD = M * D;
    // CLV: the caller knows the level table on the host (it built spatial_shapes itself) and passes it by value: it then
    // sits in the constant bank and the per-round look-ups cost no shared-memory wavefront (2.5 of the 20 set-up
    // wavefronts per (query, head) that share the L1 data pipe with the 48 gathered lines).
    if (!CLV) {
        if (tid < L) {
            s_h[tid] = (int)__ldg(shapes + 2 * tid);
            s_w[tid] = (int)__ldg(shapes + 2 * tid + 1);
            s_start[tid] = (int)__ldg(starts + tid);
        }
        __syncthreads();
    }
    auto lv_h = [&](int l) { return CLV ? lv.h[l] : s_h[l]; };
    auto lv_w = [&](int l) { return CLV ? lv.w[l] : s_w[l]; };
    auto lv_start = [&](int l) { return CLV ? lv.start[l] : s_start[l]; };
    const int qreal = blockIdx.x * GPB + grp;
    const int q = min(qreal, Q - 1);
    const size_t row = (size_t)b * Q + q;
    const float *orow = oa + row * (size_t)ld;
    const float *offp = orow + (size_t)m * LP * 2;                   // (l, p, xy) of this head
    const float *lgp = orow + (size_t)M * LP * 2 + (size_t)m * LP;   // (l, p) logits of this head
    const float *refp = ref + row * L * 2;
    const char *vlane = reinterpret_cast<const char *>(value + (size_t)b * S * MD + m * D) + lane * VB;
    asm volatile("" : "+l"(vlane));
    const uint32_t MDu = MD / CPL;
    // first round's offsets and reference point are requested before the softmax's shuffles
    uint2 raw_off = make_uint2(0u, 0u), raw_ref = make_uint2(0u, 0u);
    if (lane < LP) { raw_off = ld_stream_u2(offp + 2 * lane); raw_ref = ld_stream_u2(refp + 2 * (lane / P)); }
    // softmax over the L*P logits of (query, head): lane t holds taps t, t + LPG, ...   (ms_deform_attn.py:104-105)
    float aw[NR];
    float mx = -INFINITY;
#pragma unroll
    for (int r = 0; r < NR; ++r) {
        const int tap = r * LPG + lane;
        aw[r] = tap < LP ? __uint_as_float(ld_stream_u1(lgp + tap)) : -INFINITY;
        mx = fmaxf(mx, aw[r]);
    }
#pragma unroll
    for (int o = LPG / 2; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o, LPG));
    float sum = 0.f;
#pragma unroll
    for (int r = 0; r < NR; ++r) {
        aw[r] = (r * LPG + lane < LP) ? expf(aw[r] - mx) : 0.f;
        sum += aw[r];
    }
#pragma unroll
    for (int o = LPG / 2; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o, LPG);
    const float inv = 1.0f / sum;
    float acc[CPL];
#pragma unroll
    for (int c = 0; c < CPL; ++c) acc[c] = 0.f;
    uint4 *rec = s_rec + (tid >> 5) * (GPW * RPG) + ((tid & 31) / LPG) * RPG;
#pragma unroll
    for (int r = 0; r < NR; ++r) {
        const int s0 = r * LPG;
        const int tap = s0 + lane;
        uint32_t o1 = 0;
        float A = 0.f, Bw = 0.f, f = 0.f;
        if (tap < LP) {
            const int l = tap / P;
            const int h = lv_h(l), w = lv_w(l);
            const float lx = __fadd_rn(__uint_as_float(raw_ref.x), __fdiv_rn(__uint_as_float(raw_off.x), (float)w));   // :106-109
            const float ly = __fadd_rn(__uint_as_float(raw_ref.y), __fdiv_rn(__uint_as_float(raw_off.y), (float)h));
            const float him = ly * h - 0.5f, wim = lx * w - 0.5f;                              // cuh:290-291
            if (him > -1.f && wim > -1.f && him < (float)h && wim < (float)w) {               // cuh:293
                const float a = aw[r] * inv;
                const int hl = (int)floorf(him), wl = (int)floorf(wim);
                const float lh = him - hl, lw = wim - wl, hh = 1.f - lh, hw = 1.f - lw;
                const bool y0 = hl >= 0, y1 = hl + 1 <= h - 1, x0 = wl >= 0, x1 = wl + 1 <= w - 1;
                const int ry0 = max(hl, 0), cx0 = max(wl, 0);
                const bool xb = x0 && x1, yb = y0 && y1;
                const float ag = xb ? a : a * (x0 ? hw : lw);
                f = xb ? lw : 0.f;
                A = y0 ? hh * ag : 0.f;
                Bw = y1 ? lh * ag : 0.f;
                o1 = ((uint32_t)(lv_start(l) + ry0 * w + cx0) * MDu) | (xb ? 0x80000000u : 0u) | (yb ? 0x40000000u : 0u);
            }
        }
        if (r + 1 < NR && tap + LPG < LP) {
            raw_off = ld_stream_u2(offp + 2 * (tap + LPG));
            raw_ref = ld_stream_u2(refp + 2 * ((tap + LPG) / P));
        }
        const int nt = min(LPG, LP - s0);
        uint32_t pitch[NLR];
#pragma unroll
        for (int k = 0; k < NLR; ++k) pitch[k] = (uint32_t)lv_w(min(s0 / P + k, L - 1)) * MDu;
        if (ALLREC) {                                    // all rounds' records first, then one uninterrupted gather phase
            if (tap < LP) rec[tap] = make_uint4(o1, __float_as_uint(A), __float_as_uint(Bw), __float_as_uint(f));
            if (r + 1 < NR) continue;
            __syncwarp();
#pragma unroll
            for (int j = 0; j < LP; ++j) {
                const uint4 ra = rec[j];
                const uint32_t dx = (uint32_t)((int32_t)ra.x >> 31) & MDu;
                const uint32_t dy = (uint32_t)((int32_t)(ra.x << 1) >> 31) & ((uint32_t)lv_w(j / P) * MDu);
                uint32_t a1 = ra.x & 0x3fffffffu;
                asm volatile("" : "+r"(a1));
                const uint32_t a3 = a1 + dy;
                const float bf = __uint_as_float(ra.w), bg = 1.f - bf;
                const float b1 = __uint_as_float(ra.y) * bg, b2 = __uint_as_float(ra.y) * bf;
                const float b3 = __uint_as_float(ra.z) * bg, b4 = __uint_as_float(ra.z) * bf;
                uint32_t r1[NW], r2[NW], r3[NW], r4[NW];
                ld_corner<VB>(vlane + (size_t)a1 * VB, r1);
                ld_corner<VB>(vlane + (size_t)(a1 + dx) * VB, r2);
                ld_corner<VB>(vlane + (size_t)a3 * VB, r3);
                ld_corner<VB>(vlane + (size_t)(a3 + dx) * VB, r4);
                fma_corner<VT, NW>(b1, r1, acc); fma_corner<VT, NW>(b2, r2, acc);
                fma_corner<VT, NW>(b3, r3, acc); fma_corner<VT, NW>(b4, r4, acc);
            }
            continue;
        }
        rec[lane] = make_uint4(o1, __float_as_uint(A), __float_as_uint(Bw), __float_as_uint(f));
        __syncwarp();
#pragma unroll
        for (int j = 0; j < LPG; ++j) {
            if (j < nt) {
                const uint4 ra = rec[j];
                const uint32_t dx = (uint32_t)((int32_t)ra.x >> 31) & MDu;
                const uint32_t dy = (uint32_t)((int32_t)(ra.x << 1) >> 31) & pitch[j / P];
                uint32_t a1 = ra.x & 0x3fffffffu;
                asm volatile("" : "+r"(a1));
                const uint32_t a3 = a1 + dy;
                const float bf = __uint_as_float(ra.w), bg = 1.f - bf;
                const float b1 = __uint_as_float(ra.y) * bg, b2 = __uint_as_float(ra.y) * bf;
                const float b3 = __uint_as_float(ra.z) * bg, b4 = __uint_as_float(ra.z) * bf;
                uint32_t r1[NW], r2[NW], r3[NW], r4[NW];
                ld_corner<VB>(vlane + (size_t)a1 * VB, r1);
                ld_corner<VB>(vlane + (size_t)(a1 + dx) * VB, r2);
                ld_corner<VB>(vlane + (size_t)a3 * VB, r3);
                ld_corner<VB>(vlane + (size_t)(a3 + dx) * VB, r4);
                fma_corner<VT, NW>(b1, r1, acc); fma_corner<VT, NW>(b2, r2, acc);
                fma_corner<VT, NW>(b3, r3, acc); fma_corner<VT, NW>(b4, r4, acc);
            }
        }
        __syncwarp();
    }
    if (qreal < Q) {
        VT *o = out + (row * M + m) * D + lane * CPL;
        if constexpr (sizeof(VT) == 4) {
#pragma unroll
            for (int c = 0; c < CPL; c += 4) st_stream_f4(o + c, make_float4(acc[c], acc[c + 1], acc[c + 2], acc[c + 3]));
        } else {
#pragma unroll
            for (int c = 0; c < CPL; c += 8) Lane<__nv_bfloat16>::store(o + c, acc + c);
        }
    }
}

template <int D, typename VT>
static int launch_fused(const void *value, const int64_t *shapes, const int64_t *starts, const float *oa, int ld, const float *ref,
                        int B, int S, int M, int L, int Q, void *out, const int32_t *levels_host, cudaStream_t st) {
    constexpr int GPB = 256 / (D * (int)sizeof(VT) / 16);
    const dim3 grid(cdiv(Q, GPB), M, B);
    const VT *v = static_cast<const VT *>(value);
    VT *o = static_cast<VT *>(out);
    LevelTab lv = {};
    if (levels_host)
        for (int l = 0; l < L; ++l) { lv.h[l] = levels_host[3 * l]; lv.w[l] = levels_host[3 * l + 1]; lv.start[l] = levels_host[3 * l + 2]; }
    // Parameter-bank variant: all L*P tap records are prepared first and the gathers run as one uninterrupted phase
    // (ALLREC); without the shared-memory level table the fp32 kernel fits 40 registers = 6 CTAs per SM.  Measured at C2,
    // us per launch: fp32 474 -> 463 (6 CTAs; 5 CTAs with ALLREC 491, 6 without 487), bf16 350 -> 343 (5 CTAs; 6: 352).
    constexpr int CLV_MINB = sizeof(VT) == 4 ? 6 : 5;
#define QUBER_FUSED(LL)                                                                                                         \
    case LL:                                                                                                                    \
        if (levels_host) msda_fused_p4_kernel<D, VT, CLV_MINB, LL, true, true><<<grid, 256, 0, st>>>(v, shapes, starts, oa, ld, ref, S, M, Q, o, lv);  \
        else msda_fused_p4_kernel<D, VT, 5, LL, false><<<grid, 256, 0, st>>>(v, shapes, starts, oa, ld, ref, S, M, Q, o, lv);          \
        break;
    switch (L) {
        QUBER_FUSED(1) QUBER_FUSED(2) QUBER_FUSED(3) QUBER_FUSED(4)
        default: return QUBER_E_UNSUPPORTED;
    }
#undef QUBER_FUSED
    return check_launch("msda_fused_p4_kernel");
}

// Any D, any P, float32 or float64 (the reference dispatches both, ms_deform_attn_cuda.cu:69): the reference's own
// mapping, one thread per output scalar, evaluated entirely in T.
template <typename T>
__global__ void __launch_bounds__(256)
msda_forward_scalar_kernel(const T *__restrict__ value, const int64_t *__restrict__ shapes, const int64_t *__restrict__ starts,
                           const T *__restrict__ loc, const T *__restrict__ attn, size_t n, int S, int M, int D, int L, int Q, int P,
                           T *__restrict__ out) {
    const size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= n) return;
    const int c = idx % D;
    const size_t qm = idx / D;
    const int m = qm % M;
    const size_t b = qm / M / Q;
    const int MD = M * D;
    T col = 0;
    for (int l = 0; l < L; ++l) {
        const int h = (int)shapes[2 * l], w = (int)shapes[2 * l + 1];
        const T *v = value + (b * S + (size_t)starts[l]) * MD + m * D + c;
        for (int pt = 0; pt < P; ++pt) {
            const size_t lpi = (qm * L + l) * P + pt;
            const T lw_ = loc[2 * lpi], lh_ = loc[2 * lpi + 1], a = attn[lpi];
            const T him = lh_ * h - T(0.5), wim = lw_ * w - T(0.5);                        // cuh:290-291
            if (!(him > T(-1) && wim > T(-1) && him < (T)h && wim < (T)w)) continue;       // cuh:293
            const int hl = (int)floor(him), wl = (int)floor(wim);
            const T lh = him - hl, lw = wim - wl, hh = T(1) - lh, hw = T(1) - lw;
            T v1 = 0, v2 = 0, v3 = 0, v4 = 0;
            if (hl >= 0 && wl >= 0) v1 = v[((size_t)hl * w + wl) * MD];
            if (hl >= 0 && wl + 1 <= w - 1) v2 = v[((size_t)hl * w + wl + 1) * MD];
            if (hl + 1 <= h - 1 && wl >= 0) v3 = v[((size_t)(hl + 1) * w + wl) * MD];
            if (hl + 1 <= h - 1 && wl + 1 <= w - 1) v4 = v[((size_t)(hl + 1) * w + wl + 1) * MD];
            col += (hh * hw * v1 + hh * lw * v2 + lh * hw * v3 + lh * lw * v4) * a;        // cuh:85, :297
        }
    }
    out[idx] = col;
}

// One variant per shape class in the shipped library (the A/B switches of round 1 live in git history and
// profiles/r01_msda_variants_*.log): P == 4 and a head of >= 64 bytes -> the one-record kernel (fp32 486 us, bf16
// 354 us per launch at C2); anything else with a vectorisable head -> the two-record shared-memory broadcast.
// 5 resident CTAs per SM: the L1 data pipe is the limit and throughput follows the number of resident warps.
template <int D, typename VT>
static int launch_vec(const void *value, const int64_t *shapes, const int64_t *starts, const float *loc, const float *attn,
                      int B, int S, int M, int L, int Q, int P, void *out, cudaStream_t st) {
    constexpr int GPB = 256 / (D / Lane<VT>::CPL);
    const dim3 grid(cdiv(Q, GPB), M, B);
    const VT *v = static_cast<const VT *>(value);
    VT *o = static_cast<VT *>(out);
    constexpr int DB = D * (int)sizeof(VT);             // bytes of one head's channels
    if constexpr (DB % 64 == 0) {
        if (P == 4) {
            msda_forward_p4_kernel<D, VT, 5, 16><<<grid, 256, 0, st>>>(v, shapes, starts, loc, attn, S, M, L, Q, o);
            return check_launch("msda_forward_p4_kernel");
        }
    }
    msda_forward_kernel<D, VT, 5, true><<<grid, 256, 0, st>>>(v, shapes, starts, loc, attn, S, M, L, Q, P, o);
    return check_launch("msda_forward_kernel");
}

}  // namespace quber

extern "C" int quber_msda_forward(const void *value, const int64_t *shapes, const int64_t *starts, const float *loc,
                                  const float *attn, int B, int S, int M, int D, int L, int Q, int P, int dtype, void *out,
                                  void *stream) {
    using namespace quber;
    if (!value || !shapes || !starts || !loc || !attn || !out) return QUBER_E_INVALID;
    if (B <= 0 || S <= 0 || M <= 0 || D <= 0 || L <= 0 || L > MSDA_MAX_L || Q <= 0 || P <= 0) return QUBER_E_INVALID;
    if (dtype != 0 && dtype != 1 && dtype != 2) return QUBER_E_INVALID;
    if (M > 65535 || B > 65535 || (long long)S * M * D >= (1ll << 31)) return QUBER_E_UNSUPPORTED;
    cudaStream_t st = (cudaStream_t)stream;
    if (dtype == 2) {                                    // float64: value, loc, attn and out are all double
        const size_t n = (size_t)B * Q * M * D;
        msda_forward_scalar_kernel<double><<<(unsigned)((n + 255) / 256), 256, 0, st>>>(
            static_cast<const double *>(value), shapes, starts, reinterpret_cast<const double *>(loc),
            reinterpret_cast<const double *>(attn), n, S, M, D, L, Q, P, static_cast<double *>(out));
        return check_launch("msda_forward_scalar_kernel<double>");
    }
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
    msda_forward_scalar_kernel<float><<<(unsigned)((n + 255) / 256), 256, 0, st>>>(
        static_cast<const float *>(value), shapes, starts, loc, attn, n, S, M, D, L, Q, P, static_cast<float *>(out));
    return check_launch("msda_forward_scalar_kernel");
}

extern "C" int quber_msda_forward_fused(const void *value, const int64_t *shapes, const int64_t *starts, const float *oa, int ld,
                                        const float *reference_points, int B, int S, int M, int D, int L, int Q, int P,
                                        int dtype, void *out, const int32_t *levels_host, void *stream) {
    using namespace quber;
    if (!value || !shapes || !starts || !oa || !reference_points || !out) return QUBER_E_INVALID;
    if (B <= 0 || S <= 0 || M <= 0 || D <= 0 || L <= 0 || Q <= 0 || P <= 0 || ld < M * L * P * 3) return QUBER_E_INVALID;
    if (dtype != 0 && dtype != 1) return QUBER_E_INVALID;
    if (P != 4 || L > 4) return QUBER_E_UNSUPPORTED;                  // callers fall back to quber_msda_prepare + quber_msda_forward
    if (M > 65535 || B > 65535 || (long long)S * M * D >= (1ll << 31)) return QUBER_E_UNSUPPORTED;
    if (!aligned16(value) || !aligned16(out) || (reinterpret_cast<uintptr_t>(oa) & 7u) || (ld & 1) ||
        (reinterpret_cast<uintptr_t>(reference_points) & 7u))
        return QUBER_E_ALIGN;
    cudaStream_t st = (cudaStream_t)stream;
    if (dtype == 1) {
        switch (D) {
            case 32: return launch_fused<32, __nv_bfloat16>(value, shapes, starts, oa, ld, reference_points, B, S, M, L, Q, out, levels_host, st);
            case 64: return launch_fused<64, __nv_bfloat16>(value, shapes, starts, oa, ld, reference_points, B, S, M, L, Q, out, levels_host, st);
            default: return QUBER_E_UNSUPPORTED;
        }
    }
    switch (D) {
        case 16: return launch_fused<16, float>(value, shapes, starts, oa, ld, reference_points, B, S, M, L, Q, out, levels_host, st);
        case 32: return launch_fused<32, float>(value, shapes, starts, oa, ld, reference_points, B, S, M, L, Q, out, levels_host, st);
        case 64: return launch_fused<64, float>(value, shapes, starts, oa, ld, reference_points, B, S, M, L, Q, out, levels_host, st);
        default: return QUBER_E_UNSUPPORTED;
    }
}
