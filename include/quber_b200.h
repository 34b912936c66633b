/*
 * quber_b200.h — C ABI of libquber_b200.so: the B200 (sm_100a) implementation of
 * QuBER's maskrefiner inference hot path.
 *
 * Conventions (all entry points):
 *   - plain C, no C++/torch types; every pointer marked "dev" is a CUDA device
 *     pointer owned by the caller, "host" a host pointer; nothing is allocated
 *     or freed inside the library; scratch comes from the caller through a
 *     `*_workspace_bytes` query;
 *   - `stream` is a cudaStream_t passed as void*; calls are asynchronous on it,
 *     re-entrant, hold no global state and are CUDA-graph capturable (no host
 *     synchronisation, no host read-back);
 *   - return value: QUBER_OK or a negative QUBER_E_* code (never an exception).
 *     The reference reports the same conditions through ATen asserts
 *     (OPS/src/cuda/ms_deform_attn_cuda.cu:33-57) or Python ValueError
 *     (maskrefiner/modeling/mask_refiner/post_processing.py:199-204); the Python
 *     shim in quber_b200/_lib.py maps the codes back to those exception types.
 *
 * Reference paths are relative to the QuBER checkout; OPS =
 * ext_modules/UnseenObjectsWithMeanShift/MSMFormer/meanshiftformer/modeling/pixel_decoder/ops.
 * The reference has no C ABI of its own (its native surface is a pybind/ATen
 * module, OPS/src/vision.cpp:18-21); each entry point below names the reference
 * interface it stands in for.
 */
#ifndef QUBER_B200_H
#define QUBER_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define QUBER_ABI_VERSION 3

enum {
    QUBER_OK = 0,
    QUBER_E_INVALID = -1,    /* bad argument (null pointer, non-positive size, unsupported value) */
    QUBER_E_WORKSPACE = -2,  /* workspace pointer null / too small */
    QUBER_E_CUDA = -3,       /* a CUDA runtime call or kernel launch failed */
    QUBER_E_ALIGN = -4,      /* pointer not aligned as documented */
    QUBER_E_UNSUPPORTED = -5 /* shape outside what the kernels implement (documented per call) */
};

/* Library / ABI version (QUBER_ABI_VERSION it was built with). */
int quber_version(void);
/* Static string for a QUBER_* code. */
const char *quber_error_string(int code);
/* cudaGetLastError-style text of the last CUDA failure seen by this thread (static buffer). */
const char *quber_last_cuda_error(void);

/* ---------------------------------------------------------------------------------------------
 * K0  initial-mask encode — replaces the NumPy loop of MaskRefinerPredictor.predict
 *     (maskrefiner/predictor.py:304-357; twin: PerturbedInputGenerator.__call__,
 *     maskrefiner/data/dataset_mappers/target_generator.py:229-276).
 *   masks  dev [B,K,H,W] uint8, non-zero = inside;  K <= 254
 *   gauss  dev [(6*sigma+3)^2] float32 — quber_gaussian_table() fills a host copy
 *   out    dev [B,3,H,W] float32 = (centre heat-map, offset_y/H, offset_x/W)
 *   Later masks overwrite earlier ones in the offset planes, empty masks are skipped, centroid
 *   arithmetic is float64 exactly as in the reference.  W must be a multiple of 16 and `masks`
 *   16-byte aligned (QUBER_E_ALIGN otherwise).  valid_w (0 = W): frames whose width is not a multiple of 16
 *   (armbench 750 x 1333) are stored with a zero-padded row pitch W and their own width valid_w, which is what the
 *   x-offset plane is normalised by (predictor.py:346); the zero columns change neither a centroid nor a paste.
 * ------------------------------------------------------------------------------------------- */
void quber_gaussian_table(int sigma, float *host_table);
size_t quber_encode_workspace_bytes(int B, int K, int H, int W);
/* occupancy (optional, NULL = not wanted): dev uint8 [B][K][ceil(H/16)][ceil(W/128)], quber_occupancy_bytes() bytes —
 * 1 where the 128 x 16 pixel tile holds a set pixel of mask k.  A by-product of the pass that reads every mask byte;
 * quber_mask_boundary takes it to skip the (tile, plane) pairs that have no mask pixel nearby (87 % of them at C2). */
#define QUBER_OCC_TILE_W 128
#define QUBER_OCC_TILE_H 16
size_t quber_occupancy_bytes(int B, int K, int H, int W);
int quber_encode_initial_masks(const uint8_t *masks, int B, int K, int H, int W, int valid_w, int sigma,
                               const float *gauss, float *out, uint8_t *occupancy, void *workspace, size_t workspace_bytes,
                               void *stream);
/* quber_encode_initial_masks for callers that cannot allocate and have no workspace to pass (CUDA graph capture, tight
 * serving loops, harnesses that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * The scratch (quber_encode_workspace_bytes(B, K, H, W) bytes) lives in a static device array of
 * QUBER_ENCODE_STATIC_SCRATCH_BYTES inside the library — every GPU that loads the library holds it until the process
 * exits; larger problems are refused with QUBER_E_WORKSPACE (quber_last_cuda_error() names the sizes; C2's 32 frames of
 * 640 x 480 with 30 masks need 9.9 MB).  The call allocates nothing.  All calls on a device share the array: calls that
 * overlap on the GPU race, so keep them on one stream or order them with events.  Arguments and results as
 * quber_encode_initial_masks. */
#define QUBER_ENCODE_STATIC_SCRATCH_BYTES (16u << 20)
size_t quber_encode_static_scratch_bytes(void);
int quber_encode_initial_masks_static(const uint8_t *masks, int B, int K, int H, int W, int valid_w, int sigma,
                                      const float *gauss, float *out, uint8_t *occupancy, void *stream);
/* The same from bit-packed masks (quber_host_pack_masks / quber_rle_decode): packed dev uint32 [B*K*H*W / 32], set pixels
 * carry `value` (1..255).  masks_out (optional, NULL = not wanted): dev [B,K,H,W] uint8 receives the byte planes
 * (0 / value) — what quber_unpack_masks would give — on the way; quber_mask_boundary_bits works from the bits. */
int quber_encode_initial_masks_packed(const uint32_t *packed, int value, uint8_t *masks_out, int B, int K, int H, int W,
                                      int valid_w, int sigma, const float *gauss, float *out, uint8_t *occupancy,
                                      void *workspace, size_t workspace_bytes, void *stream);

/* ---------------------------------------------------------------------------------------------
 * K2  boundary extraction by erosion — replaces mask_to_boundary / masks_to_boundary /
 *     masks_to_fg_mask (explicit_error_estimation/util.py:62-99; duplicates
 *     tools/ours/panoptic2eee.py:23-59).
 *   masks         dev [B,K,H,W] uint8 (any values; erosion is a true uint8 min filter)
 *   d             erosion iterations = quber_boundary_dilation(H, W, ratio); 1 <= d <= 64
 *   per_instance  dev [B,K,H,W] uint8 or NULL: mask - erode_d(mask)      (util.py:72-90)
 *   boundary      dev [B,H,W] uint8 or NULL: (uint8 sum over k of per-instance) > 0   (util.py:92-99)
 *   fg            dev [B,H,W] uint8 or NULL: (uint8 sum over k of masks) > 0          (util.py:62-68)
 *   occupancy     dev uint8 or NULL: the tile occupancy written by quber_encode_initial_masks[_packed] for the SAME masks
 *                 (layout there).  Planes without a set pixel within d of a tile are not loaded for that tile and get the
 *                 zero tile; results are identical with and without it.  Ignored when K > 512.
 *   status        dev int32 [1] or NULL: cleared by the call; bit 0 is set if a TMA tile load never arrived (the kernel
 *                 gives up instead of hanging the GPU) — the outputs of THIS call are then invalid.  Always 0 on a
 *                 healthy device; the Python shim copies it out with the results and raises.
 *   W must be a multiple of 16 and the pointers 16-byte aligned.
 * quber_quadruple_maps: TP/TN/FP/FN of two binary maps (tools/ours/panoptic2eee.py:110-123),
 *   gt, pred dev [B,n] uint8 (non-zero = set) -> out dev [B,4,n] uint8 in (TP,TN,FP,FN) order
 *   (maskrefiner/modeling/mask_refiner/model.py:210-213).
 * ------------------------------------------------------------------------------------------- */
int quber_boundary_dilation(int H, int W, double ratio);
int quber_mask_boundary(const uint8_t *masks, int B, int K, int H, int W, int d, uint8_t *per_instance,
                        uint8_t *boundary, uint8_t *fg, const uint8_t *occupancy, int32_t *status, void *stream);
/* quber_mask_boundary from bit planes (packed dev uint32 [B*K*H*W / 32], bit i of word w = pixel 32 w + i; set pixels
 * carry `value` in the outputs): same three outputs, no byte planes read or needed, no TMA load (nothing to time out).
 * W % 32 == 0.  occupancy as above (from quber_encode_initial_masks_packed on the same planes) or NULL. */
int quber_mask_boundary_bits(const uint32_t *packed, int value, int B, int K, int H, int W, int d, uint8_t *per_instance,
                             uint8_t *boundary, uint8_t *fg, const uint8_t *occupancy, void *stream);
int quber_quadruple_maps(const uint8_t *gt, const uint8_t *pred, int B, size_t n, uint8_t *out, void *stream);
/* Diagnostic, NOT on the hot path (synchronises the device): number of TMA tile loads that
 * quber_mask_boundary gave up waiting for since the library was loaded (process-wide counter; the per-call signal is
 * `status` above), or -1 on a CUDA error. */
int quber_tma_timeouts(void);

/* ---------------------------------------------------------------------------------------------
 * K3/K4  post-processing — replaces, in ONE call and with no host round trip:
 *     MaskRefinerInsEmbedHead.forward's x4 bilinear up-sampling (model.py:689-708),
 *     `r.sigmoid().round()` (model.py:292), find_instance_center / group_pixels /
 *     get_instance_segmentation / merge_semantic_and_instance / get_panoptic_segmentation
 *     (maskrefiner/modeling/mask_refiner/post_processing.py:9-221), the per-instance loop of
 *     MaskRefiner.forward (model.py:313-356) and the caller-side error-map decode
 *     (eval/eval_utils.py:308-328).
 * ------------------------------------------------------------------------------------------- */
typedef struct quber_post_params {
    int32_t B, H, W;        /* output (full) resolution */
    int32_t scale;          /* 1: inputs are full-resolution maps (already up-sampled / post-processed
                               by the caller); s>1: inputs are [.., H/s, W/s] head outputs and the
                               kernels up-sample by s on the fly (bilinear, align_corners=False) */
    float offset_mul;       /* multiplies the offset after up-sampling (model.py:698: 4.0 when scale=4) */
    int32_t eee_channels;   /* 0 (no error head) or 2..4 */
    float threshold;        /* MODEL.PANOPTIC_DEEPLAB.CENTER_THRESHOLD */
    int32_t nms_kernel;     /* odd, <= 15 */
    int32_t top_k;          /* 1..1024 */
    int32_t min_area;       /* 512: post_processing.py:145 */
    int32_t stuff_area;     /* MODEL.PANOPTIC_DEEPLAB.STUFF_AREA */
    int32_t label_divisor;  /* metadata label_divisor (1000) */
    float void_label;       /* -1 */
    int32_t max_instances;  /* capacity (rows) of masks/labels/scores/boxes over the whole batch */
    int32_t flags;          /* QUBER_POST_* bits */
    /* Cropped frames — detectron2's sem_seg_postprocess as called at model.py:266-289 crops the up-sampled maps to the
     * image size (`result[:, :img_h, :img_w]`) before the (identity) resize: armbench frames are 750 x 1333 while the
     * head outputs are 188 x 334 (x4 = 752 x 1336).  All three 0 = no crop (in_h = H / scale, in_w = W / scale). */
    int32_t in_h, in_w;     /* low-resolution input size; in_h * scale >= H, in_w * scale >= valid_w */
    int32_t valid_w;        /* logical frame width <= W; W stays the row pitch of every output (W % 8 == 0) and the
                               columns valid_w .. W-1 are void (never foreground, never a centre) */
} quber_post_params;

/* centers_yx / n_centers are INPUTS (group_pixels with caller-supplied centres, post_processing.py:44);
 * the NMS / top-k kernels are skipped */
#define QUBER_POST_GIVEN_CENTERS 1
/* find_instance_center(top_k=None) (post_processing.py:36-37): every positive NMS survivor is a
 * centre; top_k is then only the capacity of centers_yx (status bit 2 is set if it overflowed) */
#define QUBER_POST_NO_TOPK 2
/* label_map points to uint8 [B,H,W] instead of int32: a quarter of the bytes to write, to send to the
 * host and to all-gather.  Needs top_k <= 255 (an image never has more instances than centres), else
 * QUBER_E_INVALID. */
#define QUBER_POST_LABEL_U8 4

typedef struct quber_post_io {
    /* inputs, dev float32; low resolution [B,C,H/scale,W/scale] */
    const float *foreground; /* [B,1,h,w] logits */
    const float *center;     /* [B,1,h,w] */
    const float *offset;     /* [B,2,h,w] (y,x) */
    const float *eee;        /* [B,C,h,w] or NULL */
    /* outputs, dev; any of the optional ones may be NULL */
    float *sem_seg;          /* opt [B,1,H,W]  up-sampled foreground logits ("sem_seg") */
    float *eee_full;         /* opt [B,C,H,W]  up-sampled error logits ("eee_boundary") */
    uint8_t *eee_onehot;     /* opt [B,C,H,W]  argmax one-hot (TP,TN,FP,FN) */
    float *panoptic;         /* opt [B,H,W]    void_label / label_divisor (+ running id) */
    void *label_map;         /* opt [B,H,W] int32 (uint8 with QUBER_POST_LABEL_U8)  0 = void, j+1 = j-th instance of
                                the image (collective payload) */
    int32_t *ins_seg;        /* opt [B,H,W]    foreground * grouped centre id (get_instance_segmentation, :107) */
    uint8_t *masks;          /* opt [max_instances,H,W] 0/1, images back to back (inst_offsets) */
    int32_t *centers_yx;     /* req [B,top_k,2] */
    int32_t *n_centers;      /* req [B] */
    int32_t *inst_offsets;   /* req [B+1] exclusive prefix of instances per image */
    int32_t *labels;         /* req [max_instances] panoptic label of each instance */
    float *scores;           /* req [max_instances] */
    float *boxes;            /* req [max_instances,4] x0,y0,x1,y1 */
    int32_t *status;         /* req [1] bit 0: max_instances exceeded (extra instances dropped);
                                        bit 1: more than top_k centres with QUBER_POST_NO_TOPK (extra dropped) */
} quber_post_io;

size_t quber_post_workspace_bytes(const quber_post_params *p);
int quber_post_process(const quber_post_params *p, const quber_post_io *io, void *workspace,
                       size_t workspace_bytes, void *stream);

/* ---------------------------------------------------------------------------------------------
 * K1  MSDeformAttn forward — replaces MultiScaleDeformableAttention.ms_deform_attn_forward
 *     (OPS/src/vision.cpp:18-21 -> OPS/src/cuda/ms_deform_attn_cuda.cu:25-85 ->
 *     OPS/src/cuda/ms_deform_im2col_cuda.cuh:242-304).
 *   value   dev [B,S,M,D]      float32 (dtype 0), bfloat16 (dtype 1) or float64 (dtype 2: loc, attn and out are then
 *                              float64 as well — the reference dispatches both floating types, ms_deform_attn_cuda.cu:69)
 *   shapes  dev [L,2] int64 (H_l, W_l);  starts dev [L] int64  — int64 as in the reference API
 *   loc     dev [B,Q,M,L,P,2]  float32 (x,y) in [0,1];  attn dev [B,Q,M,L,P] float32
 *   out     dev [B,Q,M*D]      float32 (dtype 0) / bfloat16 (dtype 1)
 *   L <= 8.  Vector path for D in {8,16,32,64} (16-byte aligned value/out), scalar path otherwise.
 *   `im2col_step` of the reference API is accepted by the Python shim and ignored: the whole batch
 *   is one launch.
 * ------------------------------------------------------------------------------------------- */
int quber_msda_forward(const void *value, const int64_t *shapes, const int64_t *starts, const float *loc,
                       const float *attn, int B, int S, int M, int D, int L, int Q, int P, int dtype,
                       void *out, void *stream);
/* The same sampling with MSDeformAttn.forward's operand arithmetic (OPS/modules/ms_deform_attn.py:102-117) inside the
 * kernel's set-up phase: no loc / attn tensors, no preparation pass.
 *   oa                dev [B*Q, ld] float32 — the output rows of ONE GEMM over the concatenated sampling_offsets |
 *                     attention_weights Linear layers: columns [0, M*L*P*2) = offsets in (m, l, p, xy) order,
 *                     [M*L*P*2, M*L*P*3) = attention logits in (m, l, p) order (ld >= M*L*P*3, even)
 *   reference_points  dev [B*Q, L, 2] float32 (x, y) in [0,1]
 *   loc = reference_points + offsets / (W_l, H_l) (:106-109), attention = softmax over L*P (:104-105), then as
 *   quber_msda_forward.  P == 4, L <= 4 and D*sizeof(value) a multiple of 64 bytes; QUBER_E_UNSUPPORTED otherwise
 *   (callers then use quber_msda_prepare + quber_msda_forward).
 *   levels_host       HOST int32 [L][3] = (H_l, W_l, start_l) or NULL — the same level table as shapes / starts, for
 *                     callers that built it on the host (the pixel decoder does): the kernel then takes it from its
 *                     parameter bank instead of shared memory (fewer wavefronts on the L1 data pipe that bounds it).
 *                     Results are identical either way; it must agree with shapes / starts. */
int quber_msda_forward_fused(const void *value, const int64_t *shapes, const int64_t *starts, const float *oa, int ld,
                             const float *reference_points, int B, int S, int M, int D, int L, int Q, int P, int dtype,
                             void *out, const int32_t *levels_host, void *stream);

/* ---------------------------------------------------------------------------------------------
 * Network glue that is pure bandwidth (csrc/fusion.cu).
 * quber_pack_input — replaces the input stage of MaskRefiner.forward
 *     (maskrefiner/modeling/mask_refiner/model.py:137-153): (uint8 image - PIXEL_MEAN) / PIXEL_STD and the
 *     channel concat with the encoded initial-mask offsets.
 *   image dev [B,C,H,W] uint8; offsets dev [B,CO,H,W] float32 (NULL iff CO == 0); mean, stdv dev [C] float32
 *   out dev [B,C+CO,H,W] float32.  H*W % 16 == 0, 16-byte aligned pointers.  Bit-exact with torch.
 * quber_fuse_concat — replaces the softmax / sigmoid / torch.cat chain in front of a hierarchy stage
 *     (model.py:745-759).  sources[i] dev [B,channels[i],h,w] float32, ops[i] 0: copy, 1: sigmoid,
 *     2: softmax over that source's channels (<= QUBER_FUSE_MAX_SOFTMAX); host arrays of n_sources entries.
 *   layout 0: out dev [B,sum(channels),h,w] float32 (what torch.cat returns)
 *   layout 1 / 2: out dev [B*h*w, pitch] float32 / bfloat16 channels-last rows, columns >= sum(channels)
 *     zero-filled — the A operand of quber_gemm_bias_act for the 1x1 convolution that follows (model.py:430-440)
 * ------------------------------------------------------------------------------------------- */
#define QUBER_FUSE_MAX_SOURCES 12
#define QUBER_FUSE_MAX_SOFTMAX 8
int quber_pack_input(const uint8_t *image, const float *offsets, const float *mean, const float *stdv, int B, int C,
                     int CO, int H, int W, float *out, void *stream);
/* RGBDFusionBackbone.forward's input stage for the bf16 inference plan (maskrefiner/modeling/backbone/resnet.py:489-493):
 *   x dev [B,9,H,W] float32 = (rgb 3 | depth 3 | encoded offsets 3), the output of quber_pack_input
 *   out_rgb / out_depth dev [B,H,W,pitch] bfloat16 (channels-last): cat([rgb, offsets], 1) / cat([depth, offsets], 1);
 *   pitch = 6, or 8 with two zero channels behind them (16-byte pixels: the stem convolution then runs with a weight
 *   zero-padded to 8 input channels — same sums — on cuDNN's tensor-core kernels instead of its 6-channel fall-backs)
 * — the two torch.cat, layout changes and casts in one pass.  H*W must be a multiple of 4. */
int quber_split_rgbd_input(const float *x, int B, int H, int W, int pitch, void *out_rgb, void *out_depth, void *stream);
int quber_fuse_concat(const float *const *sources, const int *channels, const int *ops, int n_sources, int B, int h,
                      int w, int layout, int pitch, void *out, void *stream);
/* quber_fuse_concat_rows — the same chain (model.py:745-759) for sources that are already channels-last rows:
 *   sources[i] dev [rows, channels[i]] float32 (src_bf16[i] == 0) or bfloat16 (!= 0), rows = B*h*w;
 *   out dev [rows, pitch] float32 | bfloat16 (out_bf16), columns >= sum(channels) zero-filled. */
int quber_fuse_concat_rows(const void *const *sources, const int *channels, const int *ops, const int *src_bf16,
                           int n_sources, size_t rows, int pitch, int out_bf16, void *out, void *stream);

/* ---------------------------------------------------------------------------------------------
 * Dense projections on tcgen05 / TMEM:  C[M,N] = act(A[M,K] * W[N,K]^T + bias[N])
 *     — the nn.Linear layers of MSDeformAttn (OPS/modules/ms_deform_attn.py:98-124: value_proj,
 *     sampling_offsets, attention_weights, output_proj), the encoder FFN
 *     (MSMF/modeling/pixel_decoder/msdeformattn.py:116-131) and, on channels-last activations, the 1x1
 *     convolutions of FusionLayers / the RGB-D level fusion (maskrefiner/modeling/mask_refiner/model.py:430-452,
 *     maskrefiner/modeling/backbone/resnet.py:473-485).  In the reference these are cuBLAS / cuDNN calls
 *     made by PyTorch.
 *   A     dev [M,K] row-major, row pitch lda elements;  W dev [N,K] row-major (PyTorch Linear weight), pitch ldw
 *   in_dtype   1: bfloat16 (tcgen05 kind::f16)   2: float32 multiplied as TF32 (kind::tf32)
 *   bias  dev [N] float32 or NULL;   act  0: none, 1: ReLU
 *   C     dev [M,N] row-major, pitch ldc;  out_dtype 0: float32, 1: bfloat16.  Accumulation is fp32 in TMEM.
 *   A, W, C 16-byte aligned with row pitches that are multiples of 16 bytes (QUBER_E_ALIGN otherwise).
 *   Any M, N, K: tiles past a matrix edge are zero-filled on load and clipped on store by TMA.
 * quber_gemm_timeouts: diagnostic like quber_tma_timeouts (pipeline waits that gave up; 0 when healthy).
 * ------------------------------------------------------------------------------------------- */
int quber_gemm_bias_act(const void *A, const void *W, const float *bias, void *C, int M, int N, int K, int lda,
                        int ldw, int ldc, int in_dtype, int out_dtype, int act, void *stream);
int quber_gemm_timeouts(void);
/* fp32-accurate variant ("3xTF32", split operands): the reference runs these layers in fp32
 * (`autocast(enabled=False)`, MSMF/modeling/pixel_decoder/msdeformattn.py:314; nn.Linear -> cuBLAS sgemm).  Each fp32
 * operand is split into hi (its top 19 bits, what kind::tf32 reads) and lo = x - hi, and
 * A W^T ~= A_hi W_hi^T + A_hi W_lo^T + A_lo W_hi^T is accumulated in one fp32 TMEM tile (three tcgen05.mma per k-step);
 * relative error ~2^-21 of sum |a||w|, against 2^-10 for a single TF32 product.
 * quber_split_tf32: w dev [n] float32 -> hi, lo dev [n] float32 (once per weight).
 * quber_gemm_bias_act_x3: as quber_gemm_bias_act with float32 A [M,K] (split inside the kernel, in shared memory) and
 *   the pre-split weight W_hi / W_lo dev [N,K] (pitch ldw). */
/* quber_gemm_bias_act over the K-concatenation A = [A1 | A2] (A1 dev [M,K1], A2 dev [M,K2], W dev [N, K1+K2]) without
 * materialising it — the 1x1 convolution behind the RGB-D "concat" fusion (torch.cat([rgb, depth], 1) -> Conv2d(2c, c, 1),
 * maskrefiner/modeling/backbone/resnet.py:466-479,497-501) on channels-last rows.  K1 must be a multiple of 64 (bf16) /
 * 32 (fp32) elements; everything else as quber_gemm_bias_act. */
int quber_gemm_bias_act_cat2(const void *A1, const void *A2, const void *W, const float *bias, void *C, int M, int N, int K1,
                             int K2, int lda1, int lda2, int ldw, int ldc, int in_dtype, int out_dtype, int act, void *stream);
int quber_split_tf32(const float *w, size_t n, float *hi, float *lo, void *stream);
int quber_gemm_bias_act_x3(const float *A, const float *W_hi, const float *W_lo, const float *bias, void *C, int M, int N, int K,
                           int lda, int ldw, int ldc, int out_dtype, int act, void *stream);

/* ---------------------------------------------------------------------------------------------
 * GroupNorm (+ ReLU), inference forward — replaces nn.GroupNorm(G, C) followed by ReLU wherever the network uses
 *     norm "GN": detectron2 Conv2d(norm=GN, activation=relu) in the decoder and prediction heads
 *     (maskrefiner/modeling/mask_refiner/model.py:386-403, 431-451) and the nn.Conv2d -> nn.GroupNorm -> nn.ReLU
 *     fusion blocks of the RGB-D backbone (maskrefiner/modeling/backbone/resnet.py:466-479).  PyTorch runs it as
 *     native_group_norm + clamp_min (+ dtype casts under autocast).
 *   x, y   dev [N,C,H*W] (layout 0, NCHW) or [N,H*W,C] (layout 1, channels-last); dtype 0: float32, 1: bfloat16;
 *          y may alias x.  gamma, beta dev [C] float32.  pre_bias dev [C] float32 or NULL: the convolution's bias,
 *          added to x before the statistics (GroupNorm(conv(x) + b) without a separate bias pass).  Statistics and arithmetic in fp32 (combined in double),
 *          biased variance, eps as nn.GroupNorm.  relu != 0 applies max(., 0).
 *   Needs H*W % (16 / sizeof(dtype)) == 0 for layout 0, C % (16 / sizeof(dtype)) == 0 for layout 1
 *   (QUBER_E_UNSUPPORTED otherwise), 16-byte aligned x, y, workspace.  Deterministic (no floating-point atomics).
 * ------------------------------------------------------------------------------------------- */
size_t quber_group_norm_workspace_bytes(int N, int C, int HW, int G, int layout);
int quber_group_norm_relu(const void *x, void *y, const float *pre_bias, const float *gamma, const float *beta, int N, int C, int HW, int G,
                          float eps, int relu, int layout, int dtype, void *workspace, size_t workspace_bytes,
                          void *stream);

/* ---------------------------------------------------------------------------------------------
 * quber_upsample_concat — one decoder step of detectron2's DeepLabV3PlusHead.layers (called at model.py:713):
 *     y = F.interpolate(y, size=proj_x.shape[-2:], mode="bilinear", align_corners=False); torch.cat([proj_x, y], 1)
 *   a dev [N,H,W,Ca] (proj_x; NULL iff Ca == 0), b dev [N,h,w,Cb] (y), out dev [N,H,W,Ca+Cb]: channels-last,
 *   dtype 0: float32, 1: bfloat16 (fp32 blend).  Ca, Cb multiples of 16 / sizeof(dtype) (QUBER_E_UNSUPPORTED otherwise).
 * ------------------------------------------------------------------------------------------- */
int quber_upsample_concat(const void *a, const void *b, void *out, int N, int Ca, int Cb, int H, int W, int h, int w,
                          int dtype, void *stream);

/* ---------------------------------------------------------------------------------------------
 * quber_filter_instances — the caller-side filter of the refined instances, eval/refiner_model.py:271-288:
 *     keep instance j  <=>  sum(mask_j & fg_mask) / sum(mask_j) > thr (0.3, float64 division as numpy),
 *     then (OCID / DoPose) every kept mask loses its zero-depth pixels.
 *   label_map      dev [B,H,W] int32: 0 = void, j + 1 = j-th refined instance of the frame (quber_post_process)
 *   counts         dev [B] int32 instances per frame (labels above min(counts[b], max_per) are dropped)
 *   fg             dev [B,H,W] uint8 foreground mask (non-zero = foreground) or NULL (no overlap test)
 *   zero_depth     dev [B,H,W] uint8 (non-zero = pixel without depth) or NULL
 *   out_label_map  dev [B,H,W] int32: kept instances renumbered 1..new_counts[b] in their original order
 *   keep           dev [B,max_per] uint8, new_counts dev [B] int32.  H*W % 4 == 0.
 * ------------------------------------------------------------------------------------------- */
size_t quber_filter_workspace_bytes(int B, int max_per);
int quber_filter_instances(const int32_t *label_map, const int32_t *counts, const uint8_t *fg, const uint8_t *zero_depth,
                           int B, int H, int W, int max_per, double thr, int32_t *out_label_map, uint8_t *keep,
                           int32_t *new_counts, void *workspace, size_t workspace_bytes, void *stream);

/* ---------------------------------------------------------------------------------------------
 * quber_maxpool3x3s2 — F.max_pool2d(x, kernel_size=3, stride=2, padding=1) closing the ResNet stems
 *     (maskrefiner/modeling/backbone/resnet.py:76, :125), channels-last.
 *   x dev [N,H,W,C], y dev [N,Ho,Wo,C] with Ho = (H - 1) / 2 + 1, Wo = (W - 1) / 2 + 1; dtype 0: float32, 1: bfloat16;
 *   C a multiple of 16 / sizeof(dtype).  Bit-exact with PyTorch on NaN-free input.
 * ------------------------------------------------------------------------------------------- */
int quber_maxpool3x3s2(const void *x, void *y, int N, int C, int H, int W, int dtype, void *stream);

/* ---------------------------------------------------------------------------------------------
 * Glue of the MSDeformAttn transformer encoder layer (MSMF/modeling/pixel_decoder/msdeformattn.py:116-131,
 * OPS/modules/ms_deform_attn.py:102-117) between the tcgen05 GEMMs and quber_msda_forward:
 * quber_add_layer_norm — y = LayerNorm(a + b) * gamma + beta over rows [rows, C] float32 (b may be NULL): the
 *     `src = norm(src + src2)` pairs (msdeformattn.py:126-127, :119-120).  C a multiple of 128, <= 1024.
 *     y_bf16 (optional, dev [rows, C] bfloat16) receives a rounded copy, the A operand of the next bf16 GEMM.
 *     q (optional, with pos dev [rows, C] float32): q = y + pos, the next layer's query (with_pos_embed,
 *     msdeformattn.py:124), float32 or bfloat16 (q_bf16).
 * quber_msda_prepare — from the fused sampling_offsets | attention_weights GEMM output rows
 *     oa dev [rows, ld] float32 (first M*L*P*2 offsets, then M*L*P logits) and reference_points dev [rows, L, 2]:
 *     loc dev [rows, M, L, P, 2] = reference + offset / (W_l, H_l)  (ms_deform_attn.py:106-109) and
 *     attn dev [rows, M, L, P] = softmax over L*P (:104-105); shapes dev [L,2] int64 (H_l, W_l).
 * ------------------------------------------------------------------------------------------- */
int quber_add_layer_norm(const float *a, const float *b, const float *gamma, const float *beta, size_t rows, int C,
                         float eps, float *y, void *y_bf16, const float *pos, void *q, int q_bf16, void *stream);
int quber_msda_prepare(const float *oa, int ld, const float *reference_points, const int64_t *shapes, size_t rows, int M,
                       int L, int P, float *loc, float *attn, void *stream);

/* ---------------------------------------------------------------------------------------------
 * Packed mask transfer for the host-buffer (end-to-end) path.  The reference moves the [K,H,W] uint8 initial masks
 * to the GPU byte for byte (maskrefiner/predictor.py:304-310, `torch.as_tensor(...).to(device)` in eval/refiner_model.py);
 * here the host packs them to one bit per pixel and the device expands them again.
 * quber_host_pack_masks — HOST function (no CUDA call): bit i of dst[w] = (src[32 w + i] != 0), n_bytes a multiple of 32,
 *     on up to n_threads persistent worker threads (AVX2 when the CPU has it).  *value = the one value every set pixel
 *     carries (1, or 255 as eval/refiner_model.py:230-231 makes the masks), 0 for an all-zero buffer, -1 when the buffer
 *     holds more than one non-zero value (dst is still written; the caller should send the bytes instead, since the
 *     per-instance boundary keeps the mask's own values, explicit_error_estimation/util.py:72-90).
 * quber_unpack_masks — packed dev [n_words] uint32 -> out dev [32 n_words] uint8 (0 / value), out 16-byte aligned.
 * ------------------------------------------------------------------------------------------- */
int quber_host_pack_masks(const uint8_t *src, size_t n_bytes, uint32_t *dst, int n_threads, int *value);
/* The same, asynchronous and chunked: _begin returns at once; n_threads pool threads pack the chunks
 * [chunk_end_bytes[c-1], chunk_end_bytes[c]) (ascending, multiples of 32, the last one == n_bytes; n_chunks <= 256) in
 * order; _wait blocks until chunk c of the job in flight is packed and returns its *value as above.  One job at a time:
 * a second _begin first waits for the previous job to finish. */
int quber_host_pack_begin(const uint8_t *src, size_t n_bytes, uint32_t *dst, int n_threads, const size_t *chunk_end_bytes,
                          int n_chunks);
int quber_host_pack_wait(int chunk, int *value);
int quber_unpack_masks(const uint32_t *packed, size_t n_words, int value, uint8_t *out, void *stream);

/* ---------------------------------------------------------------------------------------------
 * COCO run-length encoded masks — replaces pycocotools.mask.decode / encode / frPyObjects (a
 *     third-party dependency of the reference, cocoapi common/maskApi.c) at the reference's call sites:
 *     mask_util.decode  maskrefiner/data/dataset_mappers/perturbed_panoptic_dataset_mapper.py:195-214,
 *                       tools/ours/panoptic2eee.py:95-97
 *     mask_util.encode  tools/ours/panoptic2eee.py:61-64,125-132
 *     frPyObjects       maskrefiner/data/datasets/register_uoais_sim_instance.py:207-209
 *     and the dataset fields maskrefiner/data/datasets/register_uoais_sim_panoptic.py:121-144 carries in this form.
 *   An RLE = run lengths of the mask in COLUMN-major order, starting with a (possibly empty) run of zeros.
 * quber_rle_decode: counts dev uint32 [total_counts] (the n masks' counts back to back), offsets dev int64 [n+1]
 *     (mask i owns counts[offsets[i] .. offsets[i+1])) -> ROW-major masks:
 *       packed_out dev uint32 [n*H*W/32] (bit i of word w = pixel 32 w + i; needs W % 32 == 0; the input format of
 *                  quber_encode_initial_masks_packed) and / or masks_out dev uint8 [n,H,W] (0 / value); either may be NULL.
 *     status dev int32 [1]: bit 0 = some mask's counts do not sum to H*W (decoded as far as they go, rest 0).
 * quber_rle_encode: masks dev uint8 [n,H,W] (non-zero = set) -> counts_out dev uint32 [cap], offsets_out dev int64 [n+1];
 *     status bit 1 = cap too small (offsets_out is still complete: offsets_out[n] is the capacity needed).
 * quber_rle_from_string / quber_rle_to_string: HOST, the compressed ASCII form (maskApi.c rleFrString / rleToString);
 *     return the number of counts / characters (which exceeds `cap` when the buffer was too small), or a negative
 *     QUBER_E_* code for a malformed string.
 * ------------------------------------------------------------------------------------------- */
size_t quber_rle_decode_workspace_bytes(size_t total_counts);
int quber_rle_decode(const uint32_t *counts, const int64_t *offsets, size_t total_counts, int n, int H, int W, int value,
                     uint32_t *packed_out, uint8_t *masks_out, void *workspace, size_t workspace_bytes, int32_t *status,
                     void *stream);
size_t quber_rle_encode_workspace_bytes(int n, int H, int W, size_t cap);
int quber_rle_encode(const uint8_t *masks, int n, int H, int W, uint32_t *counts_out, size_t cap, int64_t *offsets_out,
                     void *workspace, size_t workspace_bytes, int32_t *status, void *stream);
long long quber_rle_from_string(const char *s, size_t len, uint32_t *counts, size_t cap);
long long quber_rle_to_string(const uint32_t *counts, size_t m, char *s, size_t cap);

/* ---------------------------------------------------------------------------------------------
 * Size-changing steps either side of the path (csrc/resize.cu).
 * quber_upsample_crop_resize — MaskRefinerInsEmbedHead's x`scale` bilinear up-sampling (model.py:689-708, `mul` = the
 *     offset's post-scale, :698) + detectron2 sem_seg_postprocess as called at model.py:266-289 (crop to the image size
 *     `result[:, :crop_h, :crop_w]`, bilinear resize to (out_h, out_w), align_corners=False) in one pass:
 *     src dev [B,C,h,w] float32 -> dst dev [B,C,out_h,out_w] float32.  out == crop is the identity resize (the crop only).
 *     Bit-exact with the oracle; within 2e-6 of PyTorch's two F.interpolate calls on unit-scale maps.
 * quber_resize_nearest_u8 — cv2.resize(mask, (out_w, out_h), interpolation=cv2.INTER_NEAREST) on n uint8 planes
 *     (maskrefiner/data/dataset_mappers/target_generator.py:238, eval/refiner_model.py:226-236): bit-exact with OpenCV.
 * ------------------------------------------------------------------------------------------- */
int quber_upsample_crop_resize(const float *src, int B, int C, int h, int w, int scale, float mul, int crop_h, int crop_w,
                               int out_h, int out_w, float *dst, void *stream);
int quber_resize_nearest_u8(const uint8_t *src, int n, int h, int w, int out_h, int out_w, uint8_t *dst, void *stream);

/* ---------------------------------------------------------------------------------------------
 * Evaluation metrics, the per-pixel part — replaces the loops of multilabel_metrics
 *     (eval/evaluation.py:57-274) over every (gt label, predicted label) pair: np.unique, the pairwise
 *     intersection / area counts (:180-197), seg2bmap of every mask (eval/utilities.py:672-696 = cv2.findContours
 *     RETR_EXTERNAL + drawContours) and boundary_overlap of every pair (:22-55: cv2.dilate with skimage's disk(radius)).
 *     The K x K' float64 ratios and the Hungarian assignment stay on the host (quber_b200/evaluation.py).
 *   pred, gt        dev [B,H,W] int32 label maps, labels in [0, 65535], 0 = background
 *   max_labels      ML <= 255: capacity of the per-image label lists / matrices
 *   radius          bound_pix of evaluation.py:34-35;  boundary != 0 computes the boundary counts as well
 *   pred_labels, gt_labels dev [B,ML] int32 ascending (np.unique minus 0);  n_pred, n_gt dev [B] int32
 *   inter_full      dev [B,ML+1,ML+1] int32: co-occurrence counts of (gt index, pred index), index 0 = background, so
 *                   inter_full[b][i+1][j+1] = |gt_i & pred_j| and the row / column sums are the mask areas
 *   bcount_pred/gt  dev [B,ML] int32: boundary pixels of every mask;  btp dev [B,ML,ML,2] int32: (precision tps, recall tps)
 *   bmaps_out       dev [2,B,H,W] uint8 or NULL: the boundary maps themselves (k+1 on the boundary of the k-th label; [0] = pred)
 *   status          dev int32 [1]: bit 0 more than ML labels in a map, bit 1 a label outside [0, 65535], bit 2 a mask whose
 *                   bounding box does not fit one SM's shared memory as a bit plane (> ~1.6 M pixels)
 * ------------------------------------------------------------------------------------------- */
size_t quber_eval_workspace_bytes(int B, int H, int W, int max_labels);
int quber_eval_counts(const int32_t *pred, const int32_t *gt, int B, int H, int W, int max_labels, int radius, int boundary,
                      int32_t *pred_labels, int32_t *n_pred, int32_t *gt_labels, int32_t *n_gt, int32_t *inter_full,
                      int32_t *bcount_pred, int32_t *bcount_gt, int32_t *btp, uint8_t *bmaps_out, void *workspace,
                      size_t workspace_bytes, int32_t *status, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* QUBER_B200_H */
