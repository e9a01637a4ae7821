// Model output back to a displayable image: Normalize(mean', std') -> clamp(0, 1) | min-max stretch
// -> ToPILImage, float32 NCHW -> uint8 NHWC, on the device.
//
// Replaces the `utransform` of the reference's style-transfer examples
// (examples/img_stt/adain/adain.py:168-171: T.Normalize(mean=-mean/std, std=1.0/std), clamp(0, 1),
// T.ToPILImage(); examples/img_stt/online/online.py:169-174 and offline/offline.py:108-111 use the
// stretch hdr(x) = (x - x.min()) / (x.max() - x.min()) instead of the clamp).  torchvision's chain
// is tensor.sub(mean).div(std) in float32 (_functional_tensor.py:905-928), the Lambda, then
// pic.mul(255).byte() and a CHW -> HWC transpose (functional.py to_pil_image): five passes over
// the tensor.  Here: one pass (plus one reduction pass for the stretch), every float operation the
// same IEEE round-to-nearest operation in the same order, so the bytes are identical.
//
// The stretch needs the extrema of the NORMALISED tensor.  (x - m) / s with s > 0 is monotone in x
// under IEEE rounding, so they are the images of the per-channel extrema of the input:
// tb_channel_minmax accumulates those (atomics on the ordered-integer view of the floats).
#include "tb_common.cuh"

namespace {

struct ImgParams {
    const float* src;
    uint8_t* dst;
    int64_t images;     // B
    int C;
    int64_t plane;      // H * W
    float mean[4], std[4];
    int mode;
    const float* minmax;    // device [C][2]: per-channel min, max of src (TB_IMAGE_STRETCH)
};

// Float extrema through integer atomics on the bit pattern.  The branch is on the SIGN BIT, not on
// `v >= 0`: -0.0f compares equal to zero but its pattern is 0x80000000 (INT_MIN as a signed int), which
// on the signed route would wipe out a negative minimum already stored and could never raise a maximum.
// Sign clear: signed order == float order (and every sign-set pattern is a negative int, so it wins a min
// and loses a max as it should).  Sign set: unsigned order is the reverse float order, and every sign-set
// pattern is above every sign-clear one.  NaNs never get here (fminf / fmaxf drop them; torch.min / max
// would propagate them and the stretched image would be all-NaN, i.e. undefined bytes, either way).
__device__ __forceinline__ void atomic_min_float(float* addr, float v) {
    if (__float_as_int(v) >= 0) atomicMin((int*)addr, __float_as_int(v));
    else atomicMax((unsigned int*)addr, __float_as_uint(v));
}
__device__ __forceinline__ void atomic_max_float(float* addr, float v) {
    if (__float_as_int(v) >= 0) atomicMax((int*)addr, __float_as_int(v));
    else atomicMin((unsigned int*)addr, __float_as_uint(v));
}

__global__ void __launch_bounds__(256) tb_channel_minmax_kernel(const float* __restrict__ src, int64_t images, int C,
                                                                 int64_t plane, int chunks, float* minmax) {
    tb_grid_dependency_wait();
    const int64_t bc = blockIdx.x / chunks;          // (image, channel) plane
    const int chunk = blockIdx.x - (int)(bc * chunks);
    const int c = (int)(bc % C);
    const float* __restrict__ pl = src + bc * plane;
    const int64_t per = (plane + chunks - 1) / chunks;
    const int64_t lo = chunk * per, hi = min(plane, lo + per);
    float mn = __int_as_float(0x7f800000), mx = __int_as_float(0xff800000);
    for (int64_t i = lo + threadIdx.x; i < hi; i += 256) {
        const float v = __ldg(pl + i);
        mn = fminf(mn, v);
        mx = fmaxf(mx, v);
    }
#pragma unroll
    for (int d = 16; d; d >>= 1) {
        mn = fminf(mn, __shfl_xor_sync(0xffffffffu, mn, d));
        mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, d));
    }
    if ((threadIdx.x & 31) == 0 && lo < hi) {
        atomic_min_float(minmax + 2 * c, mn);
        atomic_max_float(minmax + 2 * c + 1, mx);
    }
    (void)images;
}

__device__ __forceinline__ float normalised(float x, float m, float s) { return __fdiv_rn(__fsub_rn(x, m), s); }

template <int C>
__global__ void __launch_bounds__(256) tb_to_image_kernel(ImgParams p) {
    tb_grid_dependency_wait();
    float lo = 0.0f, span = 1.0f;
    if (p.mode == TB_IMAGE_STRETCH) {
        // x.min() / x.max() of the normalised tensor (see the header of this file)
        float hi = __int_as_float(0xff800000);
        lo = __int_as_float(0x7f800000);
#pragma unroll
        for (int c = 0; c < C; ++c) {
            lo = fminf(lo, normalised(p.minmax[2 * c], p.mean[c], p.std[c]));
            hi = fmaxf(hi, normalised(p.minmax[2 * c + 1], p.mean[c], p.std[c]));
        }
        span = __fsub_rn(hi, lo);
    }
    const int64_t total = p.images * p.plane;
    for (int64_t q = (int64_t)blockIdx.x * 256 + threadIdx.x; q < total; q += (int64_t)gridDim.x * 256) {
        const int64_t b = q / p.plane;
        const float* __restrict__ s = p.src + (b * C) * p.plane + (q - b * p.plane);
        uint8_t out[C];
#pragma unroll
        for (int c = 0; c < C; ++c) {
            float y = normalised(__ldg(s + c * p.plane), p.mean[c], p.std[c]);
            if (p.mode == TB_IMAGE_CLAMP) y = fminf(fmaxf(y, 0.0f), 1.0f);           // x.clamp(0, 1)
            else if (p.mode == TB_IMAGE_STRETCH) y = __fdiv_rn(__fsub_rn(y, lo), span);
            // pic.mul(255).byte(): truncation; out-of-range values wrap like the C cast of the int value
            out[c] = (uint8_t)(int)__fmul_rn(y, 255.0f);
        }
        uint8_t* d = p.dst + q * C;
#pragma unroll
        for (int c = 0; c < C; ++c) d[c] = out[c];
    }
}

}  // namespace

extern "C" int tb_channel_minmax(const float* src, int64_t images, int C, int64_t plane, float* minmax, tb_stream_t stream) {
    TB_REQUIRE(images >= 0 && C >= 1 && C <= 4 && plane >= 1, "tb_channel_minmax: bad shape");
    if (images == 0) return TB_OK;
    TB_REQUIRE(src && minmax, "tb_channel_minmax: NULL argument");
    TbDevice dev;
    int rc = tb_device(&dev);
    if (rc) return rc;
    const int64_t planes = images * C;
    int chunks = (int)((plane + 16383) / 16384);
    while ((int64_t)chunks * planes < 4 * dev.sm_count && chunks * 1024 < plane) chunks *= 2;
    TB_REQUIRE(planes * chunks < ((int64_t)1 << 31), "tb_channel_minmax: grid too large");
    tb_channel_minmax_kernel<<<(unsigned)(planes * chunks), 256, 0, (cudaStream_t)stream>>>(src, images, C, plane, chunks, minmax);
    TB_LAUNCH_CHECK("tb_channel_minmax_kernel");
    return TB_OK;
}

// Every argument rule of the display transform in one place, so that an entry point that enqueues several
// kernels (the static stretch below) can refuse a bad call BEFORE its first launch.
static int check_to_image_args(const char* fn, const float* src, int64_t images, int C, int H, int W, const float* mean,
                               const float* std, int mode, bool minmax_given, const uint8_t* dst) {
    TB_REQUIRE(images >= 0 && H >= 1 && W >= 1, "%s: bad shape", fn);
    TB_REQUIRE(C == 1 || C == 3 || C == 4, "%s: C=%d is not a PIL mode (1, 3 or 4 channels)", fn, C);
    TB_REQUIRE(mode >= TB_IMAGE_PLAIN && mode <= TB_IMAGE_STRETCH, "%s: unknown mode %d", fn, mode);
    TB_REQUIRE((mean == nullptr) == (std == nullptr), "%s: mean and std come together", fn);
    if (images == 0) return TB_OK;          // an empty batch is a no-op: nothing below is read
    for (int c = 0; c < C && mean; ++c) {
        // torchvision raises for a zero std (_functional_tensor.py:923-924); the stretch relies on std > 0
        TB_REQUIRE(std[c] != 0.0f, "%s: std evaluated to zero, leading to division by zero", fn);
        TB_REQUIRE(mode != TB_IMAGE_STRETCH || std[c] > 0.0f, "%s: the stretch needs std > 0", fn);
    }
    TB_REQUIRE(src && dst, "%s: NULL src/dst", fn);
    TB_REQUIRE(mode != TB_IMAGE_STRETCH || minmax_given, "%s: the stretch needs the per-channel extrema (tb_channel_minmax)", fn);
    return TB_OK;
}

extern "C" int tb_to_image_u8(const float* src, int64_t images, int C, int H, int W, const float* mean, const float* std,
                              int mode, const float* minmax, uint8_t* dst, tb_stream_t stream) {
    int rc = check_to_image_args("tb_to_image_u8", src, images, C, H, W, mean, std, mode, minmax != nullptr, dst);
    if (rc) return rc;
    if (images == 0) return TB_OK;
    ImgParams p;
    p.src = src; p.dst = dst; p.images = images; p.C = C; p.plane = (int64_t)H * W; p.mode = mode; p.minmax = minmax;
    for (int c = 0; c < 4; ++c) { p.mean[c] = 0.0f; p.std[c] = 1.0f; }
    for (int c = 0; c < C && mean; ++c) {
        p.mean[c] = mean[c];
        p.std[c] = std[c];
    }
    TbDevice dev;
    rc = tb_device(&dev);
    if (rc) return rc;
    const int64_t total = images * p.plane;
    const int64_t want = (total + 255) / 256, cap = (int64_t)dev.sm_count * 16;
    const int64_t blocks = want < cap ? want : cap;
    cudaStream_t s = (cudaStream_t)stream;
    if (C == 1) tb_to_image_kernel<1><<<(unsigned)blocks, 256, 0, s>>>(p);
    else if (C == 3) tb_to_image_kernel<3><<<(unsigned)blocks, 256, 0, s>>>(p);
    else tb_to_image_kernel<4><<<(unsigned)blocks, 256, 0, s>>>(p);
    TB_LAUNCH_CHECK("tb_to_image_kernel");
    return TB_OK;
}

// ----------------------------------------------------------------------------------------------
// The stretch without any allocation by anyone: the per-channel extrema live in a fixed array in
// device memory that belongs to this library.
// ----------------------------------------------------------------------------------------------
namespace {

static __device__ __align__(256) float g_stretch_extrema[TB_IMAGE_STATIC_MAX_CHANNELS * 2];

// (min, max) slots back to the identities of min / max
__global__ void tb_reset_extrema_kernel(float* minmax, int C) {
    tb_grid_dependency_wait();
    const int i = threadIdx.x;
    if (i < 2 * C) minmax[i] = __int_as_float((i & 1) ? 0xff800000 : 0x7f800000);
}

constexpr int kMaxDevices = 64;
float* g_extrema_addr[kMaxDevices];

// Address of this device's copy of g_stretch_extrema, looked up once per device.
int stretch_extrema(float** out) {
    int dev = 0;
    TB_CUDA(cudaGetDevice(&dev));
    TB_REQUIRE(dev >= 0 && dev < kMaxDevices, "tb_to_image_stretch_u8_static: device ordinal %d beyond %d", dev, kMaxDevices);
    float* a = __atomic_load_n(&g_extrema_addr[dev], __ATOMIC_ACQUIRE);
    if (!a) {
        void* sym = nullptr;
        TB_CUDA(cudaGetSymbolAddress(&sym, g_stretch_extrema));
        a = (float*)sym;
        __atomic_store_n(&g_extrema_addr[dev], a, __ATOMIC_RELEASE);       // racing first calls store the same address
    }
    *out = a;
    return TB_OK;
}

}  // namespace

extern "C" int tb_to_image_stretch_u8_static(const float* src, int64_t images, int C, int H, int W, const float* mean,
                                             const float* std, uint8_t* dst, tb_stream_t stream) {
    TB_REQUIRE(C >= 1 && C <= TB_IMAGE_STATIC_MAX_CHANNELS, "tb_to_image_stretch_u8_static: C=%d beyond the %d channels the fixed array holds",
               C, TB_IMAGE_STATIC_MAX_CHANNELS);
    // every argument rule of the two passes, checked before anything is enqueued: a refused call leaves
    // neither kernels on the stream nor the shared array touched
    int rc = check_to_image_args("tb_to_image_stretch_u8_static", src, images, C, H, W, mean, std, TB_IMAGE_STRETCH, true, dst);
    if (rc) return rc;
    float* extrema = nullptr;
    rc = stretch_extrema(&extrema);
    if (rc) return rc;
    if (images == 0) return TB_OK;
    tb_reset_extrema_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(extrema, C);
    TB_LAUNCH_CHECK("tb_reset_extrema_kernel");
    rc = tb_channel_minmax(src, images, C, (int64_t)H * W, extrema, stream);
    if (rc) return rc;
    return tb_to_image_u8(src, images, C, H, W, mean, std, TB_IMAGE_STRETCH, extrema, dst, stream);
}
