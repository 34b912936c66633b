// Size-changing steps either side of the path (SURVEY §8-a A7, A1's twin):
//
//   quber_upsample_crop_resize   MaskRefinerInsEmbedHead's xS bilinear up-sampling (model.py:689-708) followed by
//       detectron2's sem_seg_postprocess as called at model.py:266-289 — crop to the image size, bilinear resize to the
//       requested output size — in ONE pass from the 1/S-resolution head outputs: an output pixel takes its four
//       neighbours in the cropped up-sampled map, each of which is recomputed from the low-resolution plane with
//       PyTorch's own operation order (common.cuh::UpSampler).  Nothing full-resolution is materialised in between.
//       The reference runs 5 F.interpolate + 5 crops + 5 more F.interpolate per image.
//   quber_resize_nearest_u8      cv2.resize(mask, ..., interpolation=cv2.INTER_NEAREST) on the initial masks
//       (maskrefiner/data/dataset_mappers/target_generator.py:238, perturbed_panoptic_dataset_mapper.py:197,
//       eval/refiner_model.py:226-236 for armbench).
#include "common.cuh"

namespace quber {

// one up-sampled (and post-scaled) value of the cropped map; mul is applied after the sample, rounded on its own
__device__ __forceinline__ float up_mul(const UpSampler &up, const float *__restrict__ plane, int y, int x, float mul) {
    const float v = up.at(plane, y, x);
    return mul == 1.0f ? v : __fmul_rn(v, mul);
}

__global__ void __launch_bounds__(256)
upsample_crop_resize_kernel(const float *__restrict__ src, int C, int h, int w, int scale, float mul, int crop_h, int crop_w,
                            int out_h, int out_w, float *__restrict__ dst) {
    const int x = blockIdx.x * 64 + (threadIdx.x & 63), y = blockIdx.y * 4 + (threadIdx.x >> 6);
    if (x >= out_w || y >= out_h) return;
    const UpSampler up(h, w, scale, true);
    const bool same = out_h == crop_h && out_w == crop_w;             // same-size bilinear resize is the identity
    // aten UpSample.h: r = (float)in / out, source index = r * (dst + 0.5) - 0.5 (one FMA), clamped at 0
    const float rh = (float)crop_h / (float)out_h, rw = (float)crop_w / (float)out_w;
    float sy = fmaf(rh, (float)y + 0.5f, -0.5f), sx = fmaf(rw, (float)x + 0.5f, -0.5f);
    sy = sy < 0.f ? 0.f : sy;
    sx = sx < 0.f ? 0.f : sx;
    const int y1 = (int)sy, x1 = (int)sx;
    const int yp = (y1 < crop_h - 1) ? 1 : 0, xp = (x1 < crop_w - 1) ? 1 : 0;
    const float ly = __fsub_rn(sy, (float)y1), lx = __fsub_rn(sx, (float)x1);
    const float hy = __fsub_rn(1.f, ly), hx = __fsub_rn(1.f, lx);
    const size_t n_planes = (size_t)gridDim.z * C;                     // gridDim.z = B
    (void)n_planes;
    for (int c = 0; c < C; ++c) {
        const size_t pl = (size_t)blockIdx.z * C + c;
        const float *plane = src + pl * h * w;
        float v;
        if (same) {
            v = up_mul(up, plane, y, x, mul);
        } else {
            const float a = up_mul(up, plane, y1, x1, mul), b = up_mul(up, plane, y1, x1 + xp, mul);
            const float cc = up_mul(up, plane, y1 + yp, x1, mul), d = up_mul(up, plane, y1 + yp, x1 + xp, mul);
            const float top = __fadd_rn(__fmul_rn(hx, a), __fmul_rn(lx, b));
            const float bot = __fadd_rn(__fmul_rn(hx, cc), __fmul_rn(lx, d));
            v = __fadd_rn(__fmul_rn(hy, top), __fmul_rn(ly, bot));
        }
        dst[(pl * out_h + y) * (size_t)out_w + x] = v;
    }
}

__global__ void __launch_bounds__(256)
resize_nearest_u8_kernel(const uint8_t *__restrict__ src, int h, int w, int out_h, int out_w, uint8_t *__restrict__ dst) {
    const int x = blockIdx.x * 64 + (threadIdx.x & 63), y = blockIdx.y * 4 + (threadIdx.x >> 6);
    if (x >= out_w || y >= out_h) return;
    // OpenCV resizeNN: min(cvFloor(dst * (double)src / dst_size), src - 1)
    const double fy = (double)h / out_h, fx = (double)w / out_w;
    int sy = (int)floor(y * fy), sx = (int)floor(x * fx);
    sy = sy > h - 1 ? h - 1 : sy;
    sx = sx > w - 1 ? w - 1 : sx;
    const size_t n = blockIdx.z;
    dst[(n * out_h + y) * (size_t)out_w + x] = src[(n * h + sy) * (size_t)w + sx];
}

}  // namespace quber

extern "C" int quber_upsample_crop_resize(const float *src, int B, int C, int h, int w, int scale, float mul, int crop_h, int crop_w,
                                          int out_h, int out_w, float *dst, void *stream) {
    using namespace quber;
    if (!src || !dst || B <= 0 || C <= 0 || h <= 0 || w <= 0 || scale < 1 || out_h <= 0 || out_w <= 0) return QUBER_E_INVALID;
    if (crop_h <= 0 || crop_w <= 0 || crop_h > h * scale || crop_w > w * scale) return QUBER_E_INVALID;
    if (B > 65535 || cdiv(out_h, 4) > 65535) return QUBER_E_UNSUPPORTED;
    const dim3 grid(cdiv(out_w, 64), cdiv(out_h, 4), B);
    upsample_crop_resize_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(src, C, h, w, scale, mul, crop_h, crop_w, out_h, out_w, dst);
    return check_launch("upsample_crop_resize_kernel");
}

extern "C" int quber_resize_nearest_u8(const uint8_t *src, int n, int h, int w, int out_h, int out_w, uint8_t *dst, void *stream) {
    using namespace quber;
    if (!src || !dst || n <= 0 || h <= 0 || w <= 0 || out_h <= 0 || out_w <= 0) return QUBER_E_INVALID;
    if (n > 65535 || cdiv(out_h, 4) > 65535) return QUBER_E_UNSUPPORTED;
    const dim3 grid(cdiv(out_w, 64), cdiv(out_h, 4), n);
    resize_nearest_u8_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(src, h, w, out_h, out_w, dst);
    return check_launch("resize_nearest_u8_kernel");
}
