// Elementwise glue of the network that is pure HBM bandwidth, fused into single passes:
//
//  quber_pack_input   — MaskRefiner.forward's input stage (maskrefiner/modeling/mask_refiner/model.py:137-153):
//                       uint8 image -> float, (x - PIXEL_MEAN) / PIXEL_STD, channel-concat with the encoded
//                       initial-mask offsets.  Three torch kernels + a cat in the reference; one pass here.
//  quber_fuse_concat  — the "error-guided fusion" input of every hierarchy stage
//                       (model.py:745-759): y' = cat(y, head features of the previous stage...,
//                       softmax(eee logits) | sigmoid(other logits)...) along channels.  The reference
//                       materialises every softmax / sigmoid and re-copies the growing tensor once per torch.cat
//                       (about 2 x n_sources x the final size in traffic); here every source plane is read once
//                       and the result written once, either as NCHW fp32 (drop-in for the torch.cat result) or
//                       as channels-last rows [B*h*w, pitch] in fp32 / bf16, which is the A operand of
//                       quber_gemm_bias_act for the 1x1 convolution that follows (model.py:430-440).
#include <cuda_bf16.h>

#include "common.cuh"

namespace quber {

// ---------------------------------------------------------------------------------------------- pack_input
__global__ void __launch_bounds__(256)
pack_input_kernel(const uint8_t *__restrict__ image, const float *__restrict__ offsets, const float *__restrict__ mean,
                  const float *__restrict__ stdv, int C, int CO, size_t HW, float *__restrict__ out) {
    const int c = blockIdx.y, b = blockIdx.z;
    const size_t i = ((size_t)blockIdx.x * 256 + threadIdx.x) * 16;
    if (i >= HW) return;
    float *o = out + ((size_t)b * (C + CO) + c) * HW + i;
    if (c < C) {
        const float m = __ldg(mean + c), s = __ldg(stdv + c);
        const uint4 v = ld_stream_u4(image + ((size_t)b * C + c) * HW + i);     // 16 pixels
        const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            float4 r;
            r.x = __fdiv_rn(__fsub_rn((float)(w[k] & 0xffu), m), s);
            r.y = __fdiv_rn(__fsub_rn((float)((w[k] >> 8) & 0xffu), m), s);
            r.z = __fdiv_rn(__fsub_rn((float)((w[k] >> 16) & 0xffu), m), s);
            r.w = __fdiv_rn(__fsub_rn((float)(w[k] >> 24), m), s);
            st_stream_f4(o + 4 * k, r);
        }
    } else {
        const float *src = offsets + ((size_t)b * CO + (c - C)) * HW + i;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const uint4 v = ld_stream_u4(src + 4 * k);
            st_stream_u4(o + 4 * k, v);
        }
    }
}

// ---------------------------------------------------------------------------------------------- split_rgbd_input
// RGBDFusionBackbone.forward's input stage (maskrefiner/modeling/backbone/resnet.py:489-493): x[:, :3] | x[:, 3:6] |
// x[:, 6:] -> cat([rgb, offsets], 1) for the RGB stream and cat([depth, offsets], 1) for the depth stream.  On the bf16
// inference plan both stems want channels-last bf16: the two cats, two layout changes and two casts are one pass here
// (nine fp32 planes read once, two [B,H,W,6] bf16 tensors written once).  Four pixels per thread: 48-byte stores.
template <int PITCH>     // channels per pixel in the outputs: 6, or 8 with two zero channels (16-byte pixels, what cuDNN's tensor-core kernels want)
__global__ void __launch_bounds__(256)
split_rgbd_input_kernel(const float *__restrict__ x, size_t HW, __nv_bfloat16 *__restrict__ out_rgb, __nv_bfloat16 *__restrict__ out_depth) {
    const int b = blockIdx.y;
    const size_t i = ((size_t)blockIdx.x * 256 + threadIdx.x) * 4;
    if (i >= HW) return;
    float v[9][4];
#pragma unroll
    for (int c = 0; c < 9; ++c) {
        const uint4 r = ld_stream_u4(x + ((size_t)b * 9 + c) * HW + i);
        v[c][0] = __uint_as_float(r.x); v[c][1] = __uint_as_float(r.y); v[c][2] = __uint_as_float(r.z); v[c][3] = __uint_as_float(r.w);
    }
#pragma unroll
    for (int s = 0; s < 2; ++s) {
        uint32_t w[4 * PITCH / 2];                       // 4 pixels x PITCH channels, two bf16 per word
#pragma unroll
        for (int px = 0; px < 4; ++px)
#pragma unroll
            for (int cp = 0; cp < PITCH / 2; ++cp) {
                const int c0 = 2 * cp, c1 = 2 * cp + 1;  // output channels; 0-2 = the stream's image, 3-5 = the offsets, 6-7 = 0
                uint32_t word = 0u;
                if (cp < 3) {
                    const float f0 = v[c0 < 3 ? 3 * s + c0 : 3 + c0][px], f1 = v[c1 < 3 ? 3 * s + c1 : 3 + c1][px];
                    const __nv_bfloat162 h = __floats2bfloat162_rn(f0, f1);
                    word = *reinterpret_cast<const uint32_t *>(&h);
                }
                w[px * (PITCH / 2) + cp] = word;
            }
        __nv_bfloat16 *o = (s == 0 ? out_rgb : out_depth) + ((size_t)b * HW + i) * PITCH;
#pragma unroll
        for (int q = 0; q < PITCH / 2; ++q) st_stream_u4(o + 8 * q, make_uint4(w[4 * q], w[4 * q + 1], w[4 * q + 2], w[4 * q + 3]));
    }
}

// ---------------------------------------------------------------------------------------------- fuse_concat
constexpr int F_PIX = 64;            // pixels per CTA
constexpr int F_THREADS = 256;

struct FuseArgs {
    const float *src[QUBER_FUSE_MAX_SOURCES];
    int channels[QUBER_FUSE_MAX_SOURCES];
    int op[QUBER_FUSE_MAX_SOURCES];
    int n;
};

__device__ __forceinline__ float sigmoid_f(float x) { return 1.0f / (1.0f + __expf(-x)); }

// layout 0: NCHW fp32;  1: channels-last fp32 rows;  2: channels-last bf16 rows (pitch elements per pixel, zero padded)
template <int LAYOUT>
__global__ void __launch_bounds__(F_THREADS)
fuse_concat_kernel(FuseArgs a, int ctot, int pitch, size_t hw, void *__restrict__ out) {
    extern __shared__ float tile[];                      // [F_PIX][pitch + 1] for the channels-last layouts
    const int b = blockIdx.y;
    const size_t p0 = (size_t)blockIdx.x * F_PIX;
    const int p = threadIdx.x & (F_PIX - 1), cs = threadIdx.x / F_PIX;   // 4 channel lanes x 64 pixels
    const bool inb = p0 + p < hw;
    const int tp = pitch + 1;
    int coff = 0;
    for (int s = 0; s < a.n; ++s) {
        const int C = a.channels[s];
        const float *src = a.src[s] + (size_t)b * C * hw + p0 + p;
        if (a.op[s] == 2) {
            // softmax over the source's channels (model.py:756, eee logits: 2..4 channels), one pixel per thread
            if (cs == 0) {
                float v[QUBER_FUSE_MAX_SOFTMAX];
                float mx = -INFINITY;
#pragma unroll
                for (int c = 0; c < QUBER_FUSE_MAX_SOFTMAX; ++c)
                    if (c < C) { v[c] = inb ? __ldg(src + (size_t)c * hw) : 0.f; mx = fmaxf(mx, v[c]); }
                float sum = 0.f;
#pragma unroll
                for (int c = 0; c < QUBER_FUSE_MAX_SOFTMAX; ++c)
                    if (c < C) { v[c] = __expf(v[c] - mx); sum += v[c]; }
                const float inv = 1.0f / sum;
#pragma unroll
                for (int c = 0; c < QUBER_FUSE_MAX_SOFTMAX; ++c)
                    if (c < C) {
                        const float r = v[c] * inv;
                        if (LAYOUT == 0) { if (inb) static_cast<float *>(out)[((size_t)b * ctot + coff + c) * hw + p0 + p] = r; }
                        else tile[p * tp + coff + c] = r;
                    }
            }
        } else {
            const bool sg = a.op[s] == 1;
            for (int c = cs; c < C; c += F_THREADS / F_PIX) {
                float r = inb ? __ldg(src + (size_t)c * hw) : 0.f;
                if (sg) r = sigmoid_f(r);
                if (LAYOUT == 0) { if (inb) static_cast<float *>(out)[((size_t)b * ctot + coff + c) * hw + p0 + p] = r; }
                else tile[p * tp + coff + c] = r;
            }
        }
        coff += C;
    }
    if (LAYOUT == 0) return;
    for (int c = ctot + cs; c < pitch; c += F_THREADS / F_PIX) tile[p * tp + c] = 0.f;     // K padding for the GEMM
    __syncthreads();
    // the CTA's F_PIX rows are one contiguous block of F_PIX * pitch elements
    const size_t row0 = (size_t)b * hw + p0;
    const int nrow = (int)min((size_t)F_PIX, hw - p0);
    for (int i = threadIdx.x; i < nrow * pitch; i += F_THREADS) {
        const int r = i / pitch, c = i - r * pitch;
        const float v = tile[r * tp + c];
        if (LAYOUT == 1) static_cast<float *>(out)[row0 * pitch + i] = v;
        else static_cast<__nv_bfloat16 *>(out)[row0 * pitch + i] = __float2bfloat16_rn(v);
    }
}

// ---------------------------------------------------------------------------------------------- fuse_concat, NHWC sources
// Same operation as fuse_concat_kernel for sources that are already channels-last rows [rows, C_s] in fp32 or bf16 (what
// the cuDNN tensor-core convolutions produce): one warp per pixel row, lanes across channels, no layout copies.
struct FuseRowsArgs {
    const void *src[QUBER_FUSE_MAX_SOURCES];
    int channels[QUBER_FUSE_MAX_SOURCES];
    int op[QUBER_FUSE_MAX_SOURCES];
    int bf16[QUBER_FUSE_MAX_SOURCES];
    int n;
};

template <bool OUT_BF16>
__global__ void __launch_bounds__(256)
fuse_concat_rows_kernel(FuseRowsArgs a, int ctot, int pitch, size_t rows, void *__restrict__ out) {
    const size_t row = (size_t)blockIdx.x * 8 + (threadIdx.x >> 5);
    if (row >= rows) return;
    const int lane = threadIdx.x & 31;
    auto put = [&](int c, float v) {
        if (OUT_BF16) static_cast<__nv_bfloat16 *>(out)[row * pitch + c] = __float2bfloat16_rn(v);
        else static_cast<float *>(out)[row * pitch + c] = v;
    };
    int coff = 0;
    for (int s = 0; s < a.n; ++s) {
        const int C = a.channels[s];
        auto get = [&](int c) -> float {
            return a.bf16[s] ? __bfloat162float(static_cast<const __nv_bfloat16 *>(a.src[s])[row * C + c])
                             : __ldg(static_cast<const float *>(a.src[s]) + row * C + c);
        };
        if (a.op[s] == 2) {                               // softmax over C <= 8 channels, lanes 0..C-1
            const float v = lane < C ? get(lane) : -INFINITY;
            float mx = v;
#pragma unroll
            for (int o = 4; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
            const float e = lane < C ? __expf(v - mx) : 0.f;
            float sum = e;
#pragma unroll
            for (int o = 4; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
            if (lane < C) put(coff + lane, e * (1.0f / sum));
        } else {
            const bool sg = a.op[s] == 1;
            for (int c = lane; c < C; c += 32) {
                float v = get(c);
                if (sg) v = sigmoid_f(v);
                put(coff + c, v);
            }
        }
        coff += C;
    }
    for (int c = ctot + lane; c < pitch; c += 32) put(c, 0.f);
}

}  // namespace quber

namespace quber {
// Expansion of bit-packed masks (quber_host_pack_masks, hostpack.cpp): bit i of word w -> byte 32 w + i = `value`.
// One word per thread, two 128-bit streaming stores.
__global__ void __launch_bounds__(256)
unpack_masks_kernel(const uint32_t *__restrict__ packed, size_t n_words, uint32_t value, uint8_t *__restrict__ out) {
    const size_t w = (size_t)blockIdx.x * 256 + threadIdx.x;
    if (w >= n_words) return;
    const uint32_t bits = __ldg(packed + w);
    uint32_t o[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        const uint32_t nib = (bits >> (4 * j)) & 0xfu;
        // spread the four bits of a nibble to the low bit of four bytes: multiply puts bit k at 7k + k, mask keeps bit 8k
        o[j] = ((nib * 0x00204081u) & 0x01010101u) * value;
    }
    st_stream_u4(out + 32 * w, make_uint4(o[0], o[1], o[2], o[3]));
    st_stream_u4(out + 32 * w + 16, make_uint4(o[4], o[5], o[6], o[7]));
}
}  // namespace quber

extern "C" int quber_unpack_masks(const uint32_t *packed, size_t n_words, int value, uint8_t *out, void *stream) {
    using namespace quber;
    if (!packed || !out || n_words == 0 || value < 0 || value > 255) return QUBER_E_INVALID;
    if (!aligned16(out) || (reinterpret_cast<uintptr_t>(packed) & 3u)) return QUBER_E_ALIGN;
    unpack_masks_kernel<<<(unsigned)cdiv(n_words, (size_t)256), 256, 0, (cudaStream_t)stream>>>(packed, n_words, (uint32_t)value, out);
    return check_launch("unpack_masks_kernel");
}

extern "C" int quber_pack_input(const uint8_t *image, const float *offsets, const float *mean, const float *stdv, int B, int C,
                                int CO, int H, int W, float *out, void *stream) {
    using namespace quber;
    if (!image || !mean || !stdv || !out || B <= 0 || C <= 0 || CO < 0 || H <= 0 || W <= 0) return QUBER_E_INVALID;
    if (CO > 0 && !offsets) return QUBER_E_INVALID;
    const size_t HW = (size_t)H * W;
    if (HW % 16 != 0) return QUBER_E_UNSUPPORTED;
    if (!aligned16(image) || !aligned16(out) || (offsets && !aligned16(offsets))) return QUBER_E_ALIGN;
    if (B > 65535 || C + CO > 65535) return QUBER_E_UNSUPPORTED;
    dim3 grid((unsigned)((HW / 16 + 255) / 256), C + CO, B);
    pack_input_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(image, offsets, mean, stdv, C, CO, HW, out);
    return check_launch("pack_input_kernel");
}

extern "C" int quber_split_rgbd_input(const float *x, int B, int H, int W, int pitch, void *out_rgb, void *out_depth, void *stream) {
    using namespace quber;
    if (!x || !out_rgb || !out_depth || B <= 0 || H <= 0 || W <= 0) return QUBER_E_INVALID;
    const size_t HW = (size_t)H * W;
    if (pitch != 6 && pitch != 8) return QUBER_E_INVALID;
    if (HW % 4 != 0 || B > 65535) return QUBER_E_UNSUPPORTED;
    if (!aligned16(x) || !aligned16(out_rgb) || !aligned16(out_depth)) return QUBER_E_ALIGN;
    dim3 grid((unsigned)((HW / 4 + 255) / 256), B);
    __nv_bfloat16 *o1 = static_cast<__nv_bfloat16 *>(out_rgb), *o2 = static_cast<__nv_bfloat16 *>(out_depth);
    if (pitch == 6) split_rgbd_input_kernel<6><<<grid, 256, 0, (cudaStream_t)stream>>>(x, HW, o1, o2);
    else split_rgbd_input_kernel<8><<<grid, 256, 0, (cudaStream_t)stream>>>(x, HW, o1, o2);
    return check_launch("split_rgbd_input_kernel");
}

extern "C" int quber_fuse_concat(const float *const *sources, const int *channels, const int *ops, int n_sources, int B, int h,
                                 int w, int layout, int pitch, void *out, void *stream) {
    using namespace quber;
    if (!sources || !channels || !ops || !out || n_sources <= 0 || n_sources > QUBER_FUSE_MAX_SOURCES || B <= 0 || h <= 0 || w <= 0)
        return QUBER_E_INVALID;
    if (layout < 0 || layout > 2) return QUBER_E_INVALID;
    FuseArgs a;
    a.n = n_sources;
    int ctot = 0;
    for (int s = 0; s < n_sources; ++s) {
        if (!sources[s] || channels[s] <= 0 || ops[s] < 0 || ops[s] > 2) return QUBER_E_INVALID;
        if (ops[s] == 2 && channels[s] > QUBER_FUSE_MAX_SOFTMAX) return QUBER_E_UNSUPPORTED;
        a.src[s] = sources[s]; a.channels[s] = channels[s]; a.op[s] = ops[s];
        ctot += channels[s];
    }
    if (layout == 0) pitch = ctot;
    if (pitch < ctot) return QUBER_E_INVALID;
    if (B > 65535) return QUBER_E_UNSUPPORTED;
    const size_t hw = (size_t)h * w;
    dim3 grid((unsigned)((hw + F_PIX - 1) / F_PIX), B);
    cudaStream_t st = (cudaStream_t)stream;
    const size_t smem = layout == 0 ? 0 : (size_t)F_PIX * (pitch + 1) * sizeof(float);
    if (layout == 0) {
        fuse_concat_kernel<0><<<grid, F_THREADS, 0, st>>>(a, ctot, pitch, hw, out);
    } else if (layout == 1) {
        if (smem > 48 * 1024) QUBER_CUDA_TRY(cudaFuncSetAttribute(fuse_concat_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        fuse_concat_kernel<1><<<grid, F_THREADS, smem, st>>>(a, ctot, pitch, hw, out);
    } else {
        if (smem > 48 * 1024) QUBER_CUDA_TRY(cudaFuncSetAttribute(fuse_concat_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        fuse_concat_kernel<2><<<grid, F_THREADS, smem, st>>>(a, ctot, pitch, hw, out);
    }
    return check_launch("fuse_concat_kernel");
}

extern "C" int quber_fuse_concat_rows(const void *const *sources, const int *channels, const int *ops, const int *src_bf16,
                                      int n_sources, size_t rows, int pitch, int out_bf16, void *out, void *stream) {
    using namespace quber;
    if (!sources || !channels || !ops || !src_bf16 || !out || n_sources <= 0 || n_sources > QUBER_FUSE_MAX_SOURCES || rows == 0)
        return QUBER_E_INVALID;
    FuseRowsArgs a;
    a.n = n_sources;
    int ctot = 0;
    for (int s = 0; s < n_sources; ++s) {
        if (!sources[s] || channels[s] <= 0 || ops[s] < 0 || ops[s] > 2) return QUBER_E_INVALID;
        if (ops[s] == 2 && channels[s] > QUBER_FUSE_MAX_SOFTMAX) return QUBER_E_UNSUPPORTED;
        a.src[s] = sources[s]; a.channels[s] = channels[s]; a.op[s] = ops[s]; a.bf16[s] = src_bf16[s] != 0;
        ctot += channels[s];
    }
    if (pitch < ctot) return QUBER_E_INVALID;
    if ((rows + 7) / 8 > 0x7fffffffull) return QUBER_E_UNSUPPORTED;
    const unsigned grid = (unsigned)((rows + 7) / 8);
    cudaStream_t st = (cudaStream_t)stream;
    if (out_bf16) fuse_concat_rows_kernel<true><<<grid, 256, 0, st>>>(a, ctot, pitch, rows, out);
    else fuse_concat_rows_kernel<false><<<grid, 256, 0, st>>>(a, ctot, pitch, rows, out);
    return check_launch("fuse_concat_rows_kernel");
}
