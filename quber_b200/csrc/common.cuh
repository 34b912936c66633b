// Shared helpers for the sm_100a kernels of libquber_b200.so.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include "../../include/quber_b200.h"

namespace quber {

void set_last_cuda_error(cudaError_t e, const char *where);

// Checks the launch just issued (no synchronisation: capturable, asynchronous).
inline int check_launch(const char *where) {
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        set_last_cuda_error(e, where);
        return QUBER_E_CUDA;
    }
    return QUBER_OK;
}

#define QUBER_CUDA_TRY(expr)                                  \
    do {                                                      \
        cudaError_t _e = (expr);                              \
        if (_e != cudaSuccess) {                              \
            ::quber::set_last_cuda_error(_e, #expr);          \
            return QUBER_E_CUDA;                              \
        }                                                     \
    } while (0)

inline bool aligned16(const void *p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }
inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }
__host__ __device__ inline int cdiv(int a, int b) { return (a + b - 1) / b; }
// SMs of the current device (asked once per process: every device of a box is the same part)
inline int sm_count() {
    static int n = [] {
        int dev = 0, sms = 148;
        if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        return sms > 0 ? sms : 148;
    }();
    return n;
}

// Streaming 128-bit accesses: inputs that are read once bypass L1 allocation, outputs
// that are never re-read by the kernel are written with the same hint.
__device__ __forceinline__ uint4 ld_stream_u4(const void *p) {
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}
__device__ __forceinline__ uint2 ld_stream_u2(const void *p) {
    uint2 r;
    asm volatile("ld.global.nc.L1::no_allocate.v2.u32 {%0,%1}, [%2];" : "=r"(r.x), "=r"(r.y) : "l"(p));
    return r;
}
__device__ __forceinline__ uint32_t ld_stream_u1(const void *p) {
    uint32_t r;
    asm volatile("ld.global.nc.L1::no_allocate.u32 %0, [%1];" : "=r"(r) : "l"(p));
    return r;
}
__device__ __forceinline__ void st_stream_u4(void *p, uint4 v) {
    asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ void st_stream_u2(void *p, uint2 v) {
    asm volatile("st.global.L1::no_allocate.v2.u32 [%0], {%1,%2};" ::"l"(p), "r"(v.x), "r"(v.y) : "memory");
}
__device__ __forceinline__ void st_stream_u1(void *p, uint32_t v) {
    asm volatile("st.global.L1::no_allocate.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void st_stream_f4(float *p, float4 v) {
    asm volatile("st.global.L1::no_allocate.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}

// Bilinear sample of a low-resolution plane at full-resolution pixel (oy, ox), written as
// PyTorch's upsample_bilinear2d CUDA kernel writes it (align_corners=False) with every product and
// sum rounded on its own — the oracle (oracle/quber_oracle.c: up_sample_at) does the same, so the
// two agree bit for bit.  scale == 1 reads the plane directly.
struct UpSampler {
    int h, w, scale;
    float r;
    __host__ __device__ UpSampler(int H, int W, int s) : h(H / s), w(W / s), scale(s), r(1.0f / (float)s) {}
    // explicit low-resolution size: the full-resolution map is then a CROP of the up-sampled one (sem_seg_postprocess'
    // `result[:, :img_h, :img_w]`, detectron2 as called at model.py:266-289), e.g. 188 x 334 -> 750 x 1333
    __host__ __device__ UpSampler(int h_, int w_, int s, bool) : h(h_), w(w_), scale(s), r(1.0f / (float)s) {}
    __device__ __forceinline__ float at(const float *__restrict__ plane, int oy, int ox) const {
        if (scale == 1) return __ldg(plane + (size_t)oy * w + ox);
        float sy = __fsub_rn(__fmul_rn(r, __fadd_rn((float)oy, 0.5f)), 0.5f);
        float sx = __fsub_rn(__fmul_rn(r, __fadd_rn((float)ox, 0.5f)), 0.5f);
        sy = sy < 0.f ? 0.f : sy;
        sx = sx < 0.f ? 0.f : sx;
        const int y1 = (int)sy, x1 = (int)sx;
        const int yp = (y1 < h - 1) ? w : 0, xp = (x1 < w - 1) ? 1 : 0;
        const float ly = __fsub_rn(sy, (float)y1), lx = __fsub_rn(sx, (float)x1);
        const float hy = __fsub_rn(1.f, ly), hx = __fsub_rn(1.f, lx);
        const float *q = plane + (size_t)y1 * w + x1;
        const float a = __ldg(q), b = __ldg(q + xp), c = __ldg(q + yp), d = __ldg(q + yp + xp);
        const float top = __fadd_rn(__fmul_rn(hx, a), __fmul_rn(lx, b));
        const float bot = __fadd_rn(__fmul_rn(hx, c), __fmul_rn(lx, d));
        return __fadd_rn(__fmul_rn(hy, top), __fmul_rn(ly, bot));
    }
};

// model.py:292 `r.sigmoid().round()` == 1  <=>  x > 1.5 * 2^-24 (measured on torch 2.11, see oracle).
#define QUBER_FG_LOGIT_CUT 8.940696716308594e-08f

}  // namespace quber
