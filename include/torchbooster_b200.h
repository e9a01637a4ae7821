/*
 * torchbooster_b200 -- C ABI of the B200-native batch-transform + metrics hot path.
 *
 * The reference (yliess86/TorchBooster) is pure Python and has no FFI of its own; each entry
 * point below names the reference code it stands in for (paths relative to the reference
 * checkout; "torchvision/..." = the third-party code the reference's examples call).
 *
 * Conventions (every function):
 *   - returns 0 on success, a negative TB_E* code for argument errors, a positive cudaError_t for
 *     CUDA errors; tb_last_error() gives a thread-local message for the last non-zero return;
 *   - never throws, never synchronises the device, only enqueues on `stream`
 *     (a cudaStream_t; NULL = legacy default stream);
 *   - the caller owns every buffer and keeps it alive until the enqueued work has completed;
 *   - works on the current CUDA device of the calling thread.  What the library keeps between calls, all of it
 *     thread-safe and none of it visible in results: a launch counter (tb_launch_count), per-device caches of
 *     device attributes and of "kernel attribute already set" flags (atomics; racing first calls write the same
 *     values), the table of LMDB maps registered with CUDA (mutex), and the fixed device array of
 *     tb_to_image_stretch_u8_static (shared by every call on a device: see there);
 *   - accumulating outputs (counters, confusion, running sums) are ADDED TO: the caller zeroes them.
 */
#ifndef TORCHBOOSTER_B200_H
#define TORCHBOOSTER_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define TB_ABI_VERSION 1

#define TB_OK        0
#define TB_EINVAL  (-22)   /* bad argument (shape, alignment, enum) */
#define TB_ENOTSUP (-95)   /* valid request this build does not cover */
#define TB_ENOENT   (-2)   /* LMDB: file or key not found */

/* element types of outputs / logits */
#define TB_F32  0
#define TB_BF16 1
#define TB_F16  2
#define TB_U8   3   /* the resize entry points only: bytes out (lut = uint8 [C][256]), the hand-off to tb_augment */

/* padding modes of torchvision F.pad (torchvision/transforms/_functional_pil.py:147-223) */
#define TB_PAD_NONE      0  /* no padding declared; reads outside the image still give 0 */
#define TB_PAD_CONSTANT  1  /* fill = 0 */
#define TB_PAD_REFLECT   2  /* numpy "reflect": no edge repeat */
#define TB_PAD_EDGE      3
#define TB_PAD_SYMMETRIC 4

/* lookup-table kinds for tb_build_lut */
#define TB_LUT_TO_TENSOR  0  /* fl(x)/255                          ToTensor   (torchvision functional.py:127-178) */
#define TB_LUT_NORMALIZE  1  /* (fl(x)/255 - mean) / std           Normalize  (_functional_tensor.py:905-928)     */
#define TB_LUT_ONE_MINUS  2  /* 1 - fl(x)/255                      examples/img_gen/gan/gan.py:99, vae/vae.py:107 */
#define TB_LUT_NORMALIZE_ONE_MINUS 3  /* 1 - (fl(x)/255 - mean)/std */

/* flags */
/* tb_augment operation codes: word 0 of an 8-word int32 record; words 1..6 are its parameters */
#define TB_AUG_IDENTITY     0
#define TB_AUG_AFFINE       1   /* words 1-6: Pillow's 16.16 fixed-point a0 a1 a2 a3 a4 a5 (half-pixel folded into a2, a5) */
#define TB_AUG_BRIGHTNESS   2   /* word 1: float32 bits of the enhancement factor (likewise 3, 4, 5) */
#define TB_AUG_COLOR        3
#define TB_AUG_CONTRAST     4
#define TB_AUG_SHARPNESS    5
#define TB_AUG_POSTERIZE    6   /* word 1: byte mask */
#define TB_AUG_SOLARIZE     7   /* word 1: integer threshold t: v -> v < t ? v : 255 - v */
#define TB_AUG_AUTOCONTRAST 8
#define TB_AUG_EQUALIZE     9
#define TB_AUG_INVERT       10
#define TB_AUG_MAX_OPS      8

/* tb_to_image_u8 modes */
#define TB_IMAGE_PLAIN   0   /* Normalize -> ToPILImage */
#define TB_IMAGE_CLAMP   1   /* Normalize -> x.clamp(0, 1) -> ToPILImage            (adain.py:168-171) */
#define TB_IMAGE_STRETCH 2   /* Normalize -> (x - x.min()) / (x.max() - x.min()) -> ToPILImage  (online.py:169-174) */
#define TB_IMAGE_STATIC_MAX_CHANNELS 4   /* channels the fixed extrema array of tb_to_image_stretch_u8_static holds */

#define TB_FLAG_FORCE_GENERIC 1  /* tests: skip the streaming path, use the one-gather-per-pixel kernel */
#define TB_FLAG_FORCE_STREAM  2  /* tests: take the streaming path even where the gather kernel is the better launch */
#define TB_FLAG_NO_PDL        4  /* launch without programmatic stream serialization */
#define TB_FLAG_TWO_PHASE    16  /* tests: the two-phase (shared-memory intermediate) resize kernel instead of the sliding-window one */
#define TB_FLAG_SLIDE        32  /* tests: the sliding-window resize kernel where the ring (bulk-copy staged) kernel would run */
#define TB_FLAG_SRC_PLANAR    8  /* tb_augment: the source is uint8 [B][C][H][W] (src_row_pitch = bytes per row of one plane) */
/* Bits 8 and up of `flags` are development knobs of single kernels (group / cluster / thread counts, row split); they
   change how a launch is cut, never its result.  Callers pass 0 there. */

typedef struct CUstream_st* tb_stream_t;   /* == cudaStream_t */

#if defined(__GNUC__)
#define TB_API __attribute__((visibility("default")))
#else
#define TB_API
#endif

TB_API int         tb_version(void);
TB_API const char* tb_last_error(void);
/* number of kernels this library has launched from the calling process (bench.py's gpu_launches) */
TB_API int64_t     tb_launch_count(void);

/*
 * Per-channel 256-entry table of the pointwise part of the pipeline, built ON THE DEVICE with
 * IEEE round-to-nearest sub/div in the reference's operation order (mean/std rounded to fp32
 * first, no FMA contraction), then rounded RNE to `dst_dtype`.
 * Replaces: T.ToTensor + T.Normalize (examples/img_cls/resnet/resnet.py:94,101;
 * examples/img_stt/adain/adain.py:167-170) and the `1.0 - X` of examples/img_gen/gan/gan.py:99.
 *   mean, std : HOST pointers, C floats each (ignored for kinds without normalise; may be NULL)
 *   lut       : DEVICE, [C][256] elements of dst_dtype (TB_F32 or TB_BF16 or TB_F16)
 */
TB_API int tb_build_lut(int kind, const float* mean, const float* std, int C, int dst_dtype,
                 void* lut, tb_stream_t stream);

/*
 * Fused pad -> crop -> horizontal flip -> table lookup -> cast, uint8 NHWC -> NCHW.
 * Replaces, per batch instead of per sample: T.RandomCrop(size, padding, padding_mode) +
 * T.RandomHorizontalFlip + T.ToTensor + T.Normalize + default collate
 * (examples/img_cls/resnet/resnet.py:94-103; torchvision/transforms/transforms.py:634-722;
 * torchbooster/config.py:371-379) and the fp32 host->device copy of torchbooster/config.py:174-175.
 *
 *   dst[b,c,y,x] = lut[c][ src[b, u, v, c] ],  xs = flip[b] ? out_w-1-x : x,
 *   u = crop_top[b] + y - pad_y,  v = crop_left[b] + xs - pad_x,  folded by pad_mode; 0 outside.
 *
 *   src              : DEVICE uint8, element (b,h,w,c) at src + b*src_batch_stride + h*src_row_pitch + w*C + c
 *   dst              : DEVICE contiguous [B][C][out_h][out_w] of dst_dtype (TB_F32 / TB_BF16 / TB_F16)
 *   crop_top/left    : DEVICE int32[B] in padded coordinates (what F.crop receives); NULL = 0
 *   flip             : DEVICE uint8[B]; NULL = no flip
 *   lut              : DEVICE [C][256] of dst_dtype
 * 1 <= C <= 4.  For reflect pad_y <= H-1, pad_x <= W-1; for symmetric pad <= size.
 */
TB_API int tb_transform_u8_nhwc_to_nchw(
    const uint8_t* src, int B, int H, int W, int C, int64_t src_batch_stride, int64_t src_row_pitch,
    void* dst, int dst_dtype, int out_h, int out_w,
    const int32_t* crop_top, const int32_t* crop_left, const uint8_t* flip,
    int pad_y, int pad_x, int pad_mode,
    const void* lut, int flags, tb_stream_t stream);

/*
 * The same transform over records at arbitrary offsets from one base: image b starts at
 * src + offsets[b] (DEVICE int64[B], multiples of 16 for the streaming path).  Meant for an LMDB map
 * registered with tb_lmdb_cuda_register: the kernel reads the records in place, straight out of the
 * page cache over PCIe.  Replaces, in addition to the above, the host-side collate
 * (torchbooster/config.py:371-379).
 */
TB_API int tb_transform_u8_records_to_nchw(
    const uint8_t* src, const int64_t* offsets, int B, int H, int W, int C, int64_t src_row_pitch,
    void* dst, int dst_dtype, int out_h, int out_w,
    const int32_t* crop_top, const int32_t* crop_left, const uint8_t* flip,
    int pad_y, int pad_x, int pad_mode,
    const void* lut, int flags, tb_stream_t stream);

/*
 * Pillow's antialiased bilinear filter taps for one axis (host computation, float64 exactly as
 * libImaging/Resample.c precompute_coeffs + normalize_coeffs_8bpc, 22-bit fixed point).
 * Replaces the coefficient set-up inside T.Resize (examples/img_stt/adain/adain.py:170 ->
 * torchvision/transforms/_functional_pil.py:242-253 -> PIL.Image.resize).
 *   tb_resize_ksize : taps per output position
 *   bounds : HOST int32 [out_size][2] = (first input index, tap count)
 *   taps   : HOST int32 [out_size][ksize]
 */
TB_API int tb_resize_ksize(int in_size, int out_size);
TB_API int tb_resize_taps(int in_size, int out_size, int32_t* bounds, int32_t* taps);

/*
 * Pillow-exact bilinear resize (horizontal pass -> uint8-rounded intermediate -> vertical pass)
 * fused with crop / centre-crop (zero fill outside), flip, table lookup and cast.
 * Replaces T.Resize + T.CenterCrop + T.ToTensor + T.Normalize
 * (examples/img_stt/adain/adain.py:170; online/online.py:171-173; offline/offline.py:110).
 *
 *   R = resize(src[b]) of size [rs_h][rs_w][C]  (uint8; lives in shared memory only)
 *   dst[b,c,y,x] = lut[c][ R[crop_top[b] + off_y + y, crop_left[b] + off_x + xs, c] ]  (0 outside R)
 *
 *   bounds_x,taps_x : DEVICE copies of tb_resize_taps(W, rs_w, ..) ; ks_x = tb_resize_ksize(W, rs_w)
 *   bounds_y,taps_y : DEVICE copies of tb_resize_taps(H, rs_h, ..) ; ks_y likewise
 *   off_y, off_x    : static offsets (centre crop: int(round((rs_h - out_h)/2.0)), may be negative)
 * C in {1,3}: Pillow premultiplies alpha before resampling RGBA, which this path does not restate.
 */
TB_API int tb_resize_bilinear_u8_nhwc_to_nchw(
    const uint8_t* src, int B, int H, int W, int C, int64_t src_batch_stride, int64_t src_row_pitch,
    int rs_h, int rs_w,
    const int32_t* bounds_x, const int32_t* taps_x, int ks_x,
    const int32_t* bounds_y, const int32_t* taps_y, int ks_y,
    void* dst, int dst_dtype, int out_h, int out_w,
    const int32_t* crop_top, const int32_t* crop_left, const uint8_t* flip, int off_y, int off_x,
    const void* lut, int flags, tb_stream_t stream);

/*
 * T.RandomResizedCrop: every sample's own source box is resized (Pillow-exact bilinear, taps
 * computed on the device in float64 exactly as Pillow does) to (out_h, out_w), then flip, table
 * lookup and cast.  Replaces T.RandomResizedCrop + T.ToTensor + T.Normalize
 * (examples/img_stt/online/online.py:173 -> torchvision/transforms/transforms.py RandomResizedCrop.forward
 * -> F.resized_crop = F.crop then F.resize).
 *   box : DEVICE int32 [B][4] = top, left, height, width of the source box (RandomResizedCrop.get_params);
 *         a box outside the image yields zeros for that sample
 * C in {1,3}.  The source may be at most four times the output size per axis (TB_ENOTSUP otherwise;
 * four DP4A taps up to 2x, eight up to 4x) and
 * its base / row pitch / batch stride must be multiples of 4 bytes.
 */
TB_API int tb_resized_crop_bilinear_u8_nhwc_to_nchw(
    const uint8_t* src, int B, int H, int W, int C, int64_t src_batch_stride, int64_t src_row_pitch,
    const int32_t* box, void* dst, int dst_dtype, int out_h, int out_w,
    const uint8_t* flip, const void* lut, int flags, tb_stream_t stream);

/*
 * T.RandomRotation / T.RandAugment fused into the crop / flip / normalise launch: pad -> crop ->
 * flip (exactly tb_transform_u8_nhwc_to_nchw's arguments), then the sample's own list of n_ops
 * operations on the cropped uint8 image, then table lookup and cast.  One CTA keeps one image in
 * shared memory for the whole chain.  Replaces, per sample on PIL images,
 * T.RandomCrop + T.RandomHorizontalFlip + T.RandomRotation(15) + T.RandAugment() + T.ToTensor +
 * T.Normalize (examples/img_cls/resnet/resnet.py:96-103 -> torchvision/transforms/autoaugment.py:13-90
 * _apply_op, transforms.py RandomRotation.forward -> Pillow Image.rotate / Image.transform /
 * ImageEnhance / ImageOps); every operation reproduces Pillow's result bit for bit.
 *   ops : DEVICE int32 [B][n_ops][8], records of TB_AUG_* (torchbooster_b200/augment.py encodes
 *         (name, magnitude) pairs); unknown codes are skipped.  Fill is 0, expand is False.
 * C in {1,3}.  Images larger than about 24 KB are owned by a thread-block cluster of 2, 4 or 8
 * CTAs (distributed shared memory); 2 * C * out_h * out_w bytes must fit in the shared memory of 8
 * SMs and C * out_h * out_w < 2^20 (TB_ENOTSUP otherwise; 512 x 512 RGB fits).
 * flags: bits 8-11 force a cluster size (development knob).
 */
TB_API int tb_augment_u8_nhwc_to_nchw(
    const uint8_t* src, int B, int H, int W, int C, int64_t src_batch_stride, int64_t src_row_pitch,
    void* dst, int dst_dtype, int out_h, int out_w,
    const int32_t* crop_top, const int32_t* crop_left, const uint8_t* flip,
    int pad_y, int pad_x, int pad_mode, const int32_t* ops, int n_ops,
    const void* lut, int flags, tb_stream_t stream);

/*
 * Nearest-neighbour affine of whole uint8 images, NHWC -> NHWC (dense), one TB_AUG_AFFINE /
 * TB_AUG_IDENTITY record per sample (records[b][8], as in tb_augment): T.RandomRotation at the HEAD
 * of a pipeline, before a resized crop (examples/img_stt/online/online.py:173:
 * T.RandomRotation(180), T.RandomResizedCrop(size), T.RandAugment()), where the image is too large
 * to live in shared memory.  Pillow's 16.16 fixed-point rule, outside = 0.  C in [1,4].
 */
TB_API int tb_affine_nearest_u8_nhwc(
    const uint8_t* src, int B, int H, int W, int C, int64_t src_batch_stride, int64_t src_row_pitch,
    const int32_t* records, uint8_t* dst, tb_stream_t stream);

/*
 * Host helpers of tb_augment (no GPU work).  tb_augment_rotation_records: angles (degrees) ->
 * records of PIL.Image.rotate (expand False, centre (w/2, h/2), nearest), float64 arithmetic in
 * CPython's evaluation order.  tb_augment_params_walk: one pass over torch's mt19937 words making,
 * per sample, the draws of T.RandomCrop.get_params (word % range, when crop_range_* > 0),
 * T.RandomHorizontalFlip (flip_p >= 0), T.RandomRotation.get_params (uniform_ with rot_lo,
 * rot_span = hi - lo in float32; fused = torch's CPU kernel contracts the multiply-add) and
 * T.RandAugment.forward (word % n_choices, then word % 2 for signed operations;
 * torchvision/transforms/autoaugment.py:356-362), writing records[batch][has_rotation + n_ops][8]
 * from op_table[n_choices][2][8] (positive / negative magnitude).  Returns the number of words
 * consumed, or a negative TB_E* code.
 */
TB_API int tb_augment_rotation_records(const double* angles, int64_t n, int height, int width, int32_t* records);
TB_API int64_t tb_augment_params_walk(
    const uint64_t* words, int64_t n_words, int64_t batch, int crop_range_h, int crop_range_w, double flip_p,
    int has_rotation, float rot_lo, float rot_span, int fused, int height, int width,
    int n_ops, int n_choices, const int32_t* op_table, const uint8_t* op_signed,
    int32_t* top, int32_t* left, uint8_t* flip, int32_t* records);

/*
 * Model output -> displayable image: float32 [images][C][H][W] -> uint8 [images][H][W][C], replacing
 * the reference's `utransform` = T.Compose([T.Normalize(mean=-mean/std, std=1.0/std), clamp | hdr,
 * T.ToPILImage()]) (examples/img_stt/adain/adain.py:168-171, online/online.py:169-174,
 * offline/offline.py:108-111): (x - mean[c]) / std[c] in float32 (IEEE sub, div), the clamp or the
 * min-max stretch, * 255, truncation; the same operations in the same order as torchvision, so the
 * bytes are identical.
 *   mean, std : HOST float[C], the T.Normalize arguments as float32 (NULL, NULL = no Normalize)
 *   minmax    : DEVICE float[C][2], per-channel (min, max) of src, for TB_IMAGE_STRETCH only.
 *               tb_channel_minmax ACCUMULATES into it (the caller initialises it to +inf, -inf).
 *               -0.0 and +0.0 order as torch.min / max order them; NaNs are skipped (torch would
 *               propagate them and stretch every pixel to NaN: undefined bytes either way).
 */
TB_API int tb_channel_minmax(const float* src, int64_t images, int C, int64_t plane, float* minmax, tb_stream_t stream);
TB_API int tb_to_image_u8(const float* src, int64_t images, int C, int H, int W, const float* mean, const float* std,
                          int mode, const float* minmax, uint8_t* dst, tb_stream_t stream);

/*
 * tb_channel_minmax + tb_to_image_u8(TB_IMAGE_STRETCH) in one call that allocates nothing, here or
 * in the caller: the per-channel extrema between the two passes live in a fixed device array of
 * this library (TB_IMAGE_STATIC_MAX_CHANNELS x 2 floats; C beyond that is rejected with
 * TB_EINVAL), reset to (+inf, -inf) on the stream at the start of every call.  No reference
 * counterpart (the reference's tensors come from torch's allocator); same bytes as the two-call
 * form.  The array is ONE per device: calls that overlap in time on the same device (different
 * streams or threads) race on it -- keep them on one stream or serialise them.  The array's
 * address is looked up once per device, on the first call for that device; make that first call
 * before any stream capture begins.
 */
TB_API int tb_to_image_stretch_u8_static(const float* src, int64_t images, int C, int H, int W, const float* mean,
                                         const float* std, uint8_t* dst, tb_stream_t stream);

/*
 * The resize / resized-crop path over a RAGGED batch: images of different sizes (what the reference's
 * datasets hand to T.Resize(size) + T.CenterCrop(size) -- examples/img_stt/adain/adain.py:170 on
 * paintings and COCO photographs -- or to T.RandomResizedCrop, examples/img_stt/online/online.py:173),
 * one launch, fixed output size.  Image b starts at src + offsets[b] with row pitch pitches[b]
 * (all multiples of 4 bytes).  geometry[b] = {top, left, height, width of the source region;
 * resized height, width of that region; y, x of the output window inside the resized region (may
 * be negative or overhang: zero fill, T.CenterCrop's padding); image height, width}.  Taps are
 * computed in the kernel exactly as Pillow does for (region size -> resized size).  max_taps: 4
 * when every region is at most twice its resized size per axis, 8 when at most four times
 * (a sample beyond that yields zeros; the host wrapper refuses such batches).
 *   offsets : DEVICE int64[B]   pitches : DEVICE int32[B]   geometry : DEVICE int32[B][10]
 */
TB_API int tb_resize_crop_ragged_u8_to_nchw(
    const uint8_t* src, const int64_t* offsets, const int32_t* pitches, int B, int C, const int32_t* geometry, int max_taps,
    void* dst, int dst_dtype, int out_h, int out_w, const uint8_t* flip, const void* lut, int flags, tb_stream_t stream);

/*
 * One pass over a logits matrix: top-1 and top-k correct counts and confusion-matrix bins.
 * Replaces torchbooster.metrics.accuracy (torchbooster/metrics.py:11-27: argmax == labels, sum)
 * -- the count, before the division -- and adds the top-k / confusion accumulators the north
 * star names (no reference counterpart; defined by the rank rule).
 *
 *   pred_n = first index of the row maximum, NaN counting as maximal (torch.argmax)
 *   rank_n = #{j : l[n,j] > l[n,y_n]  or  (l[n,j] == l[n,y_n] and j < y_n)}
 *   counters[0] += N ; [1] += #{rank_n < 1} ; [2] += #{rank_n < k} ; [3] += #{y_n outside [0,C)}
 *   confusion[y_n*C + pred_n] += 1   (skipped for labels outside [0,C))
 *
 *   logits   : DEVICE [N][C], row n at logits + n*row_stride elements, dtype TB_F32/TB_BF16/TB_F16
 *   labels   : DEVICE int64[N]
 *   counters : DEVICE int64[4];  confusion : DEVICE int64[C*C] or NULL
 */
TB_API int tb_metrics_update(const void* logits, int logits_dtype, int64_t N, int C, int64_t row_stride,
                      const int64_t* labels, int k,
                      int64_t* counters, int64_t* confusion, tb_stream_t stream);

/*
 * Device-resident running mean: state[0] += sum(x[0..n)), state[1] += n   (float64).
 * Replaces the two `.item()` syncs per step that feed torchbooster.metrics.RunningAverage
 * (torchbooster/metrics.py:55-75; call sites examples/img_cls/resnet/resnet.py:66-67).
 * The sum over x is taken in a fixed order (deterministic for a given n).
 *   x : DEVICE, n elements of dtype TB_F32/TB_BF16/TB_F16 ;  state : DEVICE double[2]
 */
TB_API int tb_running_sum_update(const void* x, int dtype, int64_t n, double* state, tb_stream_t stream);

/*
 * int32 wire format of the packed metric state for the cross-rank all_reduce (the north star's
 * metric reduction; the reference itself offers only gather / barrier, torchbooster/distributed.py:41-68, on the
 * process group it creates at :174-179).  Half the bytes of the int64 state on the wire, same totals:
 *   state : DEVICE int64 [4 + bins] = n, correct_1, correct_k, invalid_labels, confusion bins (tb_metrics_update)
 *   wire  : DEVICE int32 [12 + bins] = every header value as three 21-bit limbs (summing over up to 1024 ranks carries
 *           nothing from limb to limb, so unpack rebuilds the exact int64 totals), then the bins narrowed to int32.
 * pack before the all_reduce(SUM) of `wire`, unpack after it (unpack OVERWRITES state with the totals).  Exact while
 * every bin TOTAL stays below 2^31, i.e. while the ranks together have scored fewer than 2^31 rows since the last
 * reset; the caller guards that (MetricState counts the rows it was fed).
 */
TB_API int tb_metric_state_pack_i32(const int64_t* state, int64_t bins, int32_t* wire, tb_stream_t stream);
TB_API int tb_metric_state_unpack_i32(const int32_t* wire, int64_t bins, int64_t* state, tb_stream_t stream);

/*
 * Read-only LMDB reader (host side; liblmdb 0.9 data format, memcmp key order, main DB only).
 * Replaces py-lmdb under torchbooster.lmdb.LMDBReader (torchbooster/lmdb.py:48-83:
 * lmdb.open(path, readonly=True, lock=False, readahead=False) + txn.get(key)).
 *   tb_lmdb_open   : path = LMDB directory (holding data.mdb) or the data file; TB_ENOENT if missing
 *   tb_lmdb_get    : *data points into the read-only map (valid until close); TB_ENOENT if absent
 *   tb_lmdb_gather : batched `reader[idx]` (key = decimal idx, torchbooster/lmdb.py:96-97) of n
 *                    records of exactly record_bytes each into dst + i*record_bytes (host memory,
 *                    ideally pinned) on `threads` host threads
 */
typedef struct tb_lmdb tb_lmdb;
TB_API int     tb_lmdb_open(const char* path, tb_lmdb** out);
TB_API int     tb_lmdb_close(tb_lmdb* db);
TB_API int64_t tb_lmdb_entries(const tb_lmdb* db);
TB_API int     tb_lmdb_get(const tb_lmdb* db, const void* key, int64_t klen, const void** data, int64_t* len);
TB_API int     tb_lmdb_gather(const tb_lmdb* db, const int64_t* indices, int64_t n, int64_t record_bytes,
                              void* dst, int threads);
/* Records of different sizes.  dst == NULL: sizes[i] <- length of record str(indices[i]).  Otherwise record i
 * is also copied to dst + offsets[i] (offsets laid out by the caller from a first call's sizes; capacity =
 * bytes available at dst).  TB_ENOENT for a missing key. */
TB_API int     tb_lmdb_gather_ragged(const tb_lmdb* db, const int64_t* indices, int64_t n, int64_t* sizes,
                                     const int64_t* offsets, void* dst, int64_t capacity, int threads);

/*
 * In-place GPU access to an LMDB database: tb_lmdb_cuda_register page-locks the read-only map and
 * maps it into the device address space (cudaHostRegister); *device_base is the device address of
 * the first byte of the map.  tb_lmdb_record_offsets gives the byte offsets of fixed-size records
 * `str(idx)` from that base (HOST arrays).  Together with tb_transform_u8_records_to_nchw the
 * kernel reads the records straight out of the page cache over PCIe: no host-side gather.
 * Replaces the per-sample `txn.get` + decode + collate of the reference's workers
 * (torchbooster/lmdb.py:72-97, torchbooster/config.py:371-379).  TB_ENOTSUP when the platform cannot
 * register the mapping (callers fall back to tb_lmdb_gather).
 */
TB_API int tb_lmdb_cuda_register(tb_lmdb* db, void** device_base);
TB_API int tb_lmdb_cuda_unregister(tb_lmdb* db);
TB_API int tb_lmdb_record_offsets(const tb_lmdb* db, const int64_t* indices, int64_t n, int64_t record_bytes,
                                  int64_t* offsets);

/*
 * Host helper of the T.RandomResizedCrop parameter replay: the sequential accept / reject walk of
 * RandomResizedCrop.get_params (torchvision/transforms/transforms.py; used at
 * examples/img_stt/online/online.py:173) over pre-drawn generator words and their per-word
 * candidates (computed by the caller with torch's own float32 arithmetic).  All pointers HOST.
 * Returns the number of words consumed, or TB_EINVAL when n_words is too small.
 */
TB_API int64_t tb_rrc_params_walk(const uint64_t* words, const double* area_cand, const double* aspect_cand,
                                  int64_t n_words, int64_t batch, int height, int width, double ratio_min,
                                  double ratio_max, double flip_p, int32_t* box, uint8_t* flip);
/* The same walk over images of different sizes: scale_cand[w] is the uniform draw itself, multiplied
 * here by sample b's own area heights[b] * widths[b] (host int32 arrays). */
TB_API int64_t tb_rrc_params_walk_ragged(
    const uint64_t* words, const double* scale_cand, const double* aspect_cand, int64_t n_words, int64_t batch,
    const int32_t* heights, const int32_t* widths, double ratio_min, double ratio_max, double flip_p,
    int32_t* box, uint8_t* flip);

#ifdef __cplusplus
}
#endif
#endif /* TORCHBOOSTER_B200_H */
