// Per-channel 256-entry lookup tables of the pointwise stage (ToTensor / Normalize / 1-X / cast).
//
// The table is built on the device with IEEE round-to-nearest intrinsics in the reference's
// operation order, so every entry equals what torchvision's CPU path produces for that byte:
//   ToTensor   : fl(x) / 255                       torchvision/transforms/functional.py:127-178
//   Normalize  : (t - mean32) / std32              torchvision/transforms/_functional_tensor.py:905-928
//   1.0 - X    :                                   reference examples/img_gen/gan/gan.py:99
// __fdiv_rn / __fsub_rn are never contracted into FMAs and ignore -use_fast_math; this file is
// additionally compiled with -fmad=false.  A fused x*scale+bias would differ by up to 338 ulp
// (SURVEY.md F8), which is why the hot kernels look bytes up instead of computing.
#include "tb_common.cuh"

struct LutParams {
    float mean[4];
    float std[4];
};

template <typename T>
__device__ __forceinline__ T tb_cast_rn(float v);
template <>
__device__ __forceinline__ float tb_cast_rn<float>(float v) { return v; }
template <>
__device__ __forceinline__ __nv_bfloat16 tb_cast_rn<__nv_bfloat16>(float v) { return __float2bfloat16_rn(v); }
template <>
__device__ __forceinline__ __half tb_cast_rn<__half>(float v) { return __float2half_rn(v); }

template <typename T>
__global__ void tb_build_lut_kernel(int kind, LutParams p, int C, T* __restrict__ lut) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= C * 256) return;
    int c = i >> 8;
    float t = __fdiv_rn((float)(i & 255), 255.0f);
    if (kind == TB_LUT_NORMALIZE || kind == TB_LUT_NORMALIZE_ONE_MINUS)
        t = __fdiv_rn(__fsub_rn(t, p.mean[c]), p.std[c]);
    if (kind == TB_LUT_ONE_MINUS || kind == TB_LUT_NORMALIZE_ONE_MINUS)
        t = __fsub_rn(1.0f, t);
    lut[i] = tb_cast_rn<T>(t);
}

extern "C" int tb_build_lut(int kind, const float* mean, const float* std, int C, int dst_dtype, void* lut,
                            tb_stream_t stream) {
    TB_REQUIRE(kind >= TB_LUT_TO_TENSOR && kind <= TB_LUT_NORMALIZE_ONE_MINUS, "tb_build_lut: unknown kind %d", kind);
    TB_REQUIRE(C >= 1 && C <= 4, "tb_build_lut: C=%d not in [1,4]", C);
    TB_REQUIRE(lut != nullptr, "tb_build_lut: lut is NULL");
    TB_REQUIRE(tb_dtype_size(dst_dtype) != 0, "tb_build_lut: unknown dst_dtype %d", dst_dtype);
    LutParams p;
    for (int c = 0; c < 4; ++c) { p.mean[c] = 0.f; p.std[c] = 1.f; }
    if (kind == TB_LUT_NORMALIZE || kind == TB_LUT_NORMALIZE_ONE_MINUS) {
        TB_REQUIRE(mean != nullptr && std != nullptr, "tb_build_lut: normalise needs mean and std");
        for (int c = 0; c < C; ++c) {
            // torchvision raises ValueError for a zero std (_functional_tensor.py:923-924)
            TB_REQUIRE(std[c] != 0.f, "tb_build_lut: std evaluated to zero, leading to division by zero");
            p.mean[c] = mean[c];
            p.std[c] = std[c];
        }
    }
    cudaStream_t s = (cudaStream_t)stream;
    int n = C * 256, threads = 256, blocks = (n + threads - 1) / threads;
    if (dst_dtype == TB_F32)
        tb_build_lut_kernel<float><<<blocks, threads, 0, s>>>(kind, p, C, (float*)lut);
    else if (dst_dtype == TB_BF16)
        tb_build_lut_kernel<__nv_bfloat16><<<blocks, threads, 0, s>>>(kind, p, C, (__nv_bfloat16*)lut);
    else
        tb_build_lut_kernel<__half><<<blocks, threads, 0, s>>>(kind, p, C, (__half*)lut);
    TB_LAUNCH_CHECK("tb_build_lut_kernel");
    return TB_OK;
}
