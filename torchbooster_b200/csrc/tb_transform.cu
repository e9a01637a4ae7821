// Fused pad -> crop -> horizontal flip -> table lookup -> cast, uint8 NHWC -> NCHW (sm_100a).
//
// Replaces the per-sample torchvision chain of the reference's examples
// (examples/img_cls/resnet/resnet.py:94-103: RandomCrop(32, 4, "reflect"), RandomHorizontalFlip,
// ToTensor, Normalize; collate at torchbooster/config.py:371-379) with one launch per batch.
//
// Two kernels, same index map (DESIGN.md "transform"):
//   * tb_xf_generic_kernel : one gather per output pixel, any C in [1,4], any geometry / padding.
//   * tb_xf_staged_kernel  : the HBM-bound path for C == 3.  Persistent CTAs (one per SM); a
//     producer warp streams the crop rows of a tile into a shared-memory ring with 1-D bulk
//     async copies (TMA engine, cp.async.bulk -> SASS UBLKCP) signalled on mbarriers; consumer
//     warps de-interleave / flip out of shared memory, look every byte up in a bank-conflict-free
//     (32x lane-replicated) table and write 16-byte streaming stores to the three planes.
//     Tiles whose crop window leaves the image (padding) take the gather route inside the same
//     launch, so any parameter values are safe.
// The arithmetic is a table lookup on purpose: see tb_lut.cu.
#include "tb_common.cuh"

namespace {

struct XfParams {
    const uint8_t* src;
    int B, H, W, C;
    int64_t bs, pitch;
    void* dst;
    int out_h, out_w;
    const int32_t* top;
    const int32_t* left;
    const uint8_t* flip;
    int pad_y, pad_x, pad_mode;
    const void* lut;
};

// Source coordinate of padded coordinate i for an axis of length n; `inside` cleared when the
// element is fill (0).  numpy.pad semantics (torchvision/transforms/_functional_pil.py:200-223).
__device__ __forceinline__ int tb_fold(int i, int n, int mode, bool& inside) {
    if ((unsigned)i < (unsigned)n) return i;
    if (mode == TB_PAD_REFLECT)
        i = i < 0 ? -i : 2 * (n - 1) - i;
    else if (mode == TB_PAD_SYMMETRIC)
        i = i < 0 ? -i - 1 : 2 * n - 1 - i;
    else if (mode == TB_PAD_EDGE)
        i = i < 0 ? 0 : n - 1;
    else
        i = -1;
    if ((unsigned)i >= (unsigned)n) {
        inside = false;
        return 0;
    }
    return i;
}

// ------------------------------------------------------------------------------------------
// generic: one thread per output pixel, all channels
// ------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(256) tb_xf_generic_kernel(XfParams p, int blocks_per_image) {
    __shared__ T s_lut[4 * 256];
    const T* __restrict__ glut = (const T*)p.lut;
    for (int i = threadIdx.x; i < p.C * 256; i += blockDim.x) s_lut[i] = glut[i];
    __syncthreads();

    const int b = blockIdx.x / blocks_per_image;
    const int pix = (blockIdx.x - b * blocks_per_image) * blockDim.x + threadIdx.x;
    if (pix >= p.out_h * p.out_w) return;
    const int y = pix / p.out_w;
    const int x = pix - y * p.out_w;
    const int top = (p.top ? p.top[b] : 0) - p.pad_y;
    const int left = (p.left ? p.left[b] : 0) - p.pad_x;
    const bool flip = p.flip ? (p.flip[b] != 0) : false;
    const int xs = flip ? p.out_w - 1 - x : x;
    bool inside = true;
    const int u = tb_fold(top + y, p.H, p.pad_mode, inside);
    const int v = tb_fold(left + xs, p.W, p.pad_mode, inside);
    const uint8_t* __restrict__ s = p.src + (int64_t)b * p.bs + (int64_t)u * p.pitch + (int64_t)v * p.C;
    T* __restrict__ d = (T*)p.dst + ((int64_t)b * p.C * p.out_h + y) * p.out_w + x;
    const int64_t plane = (int64_t)p.out_h * p.out_w;
    for (int c = 0; c < p.C; ++c) {
        const int e = inside ? (int)s[c] : 0;
        d[c * plane] = s_lut[c * 256 + e];
    }
}

// ------------------------------------------------------------------------------------------
// staged: C == 3, persistent, bulk-copy ring + replicated table
// ------------------------------------------------------------------------------------------
constexpr int kXfThreads = 512;                 // warp 0 = producer, warps 1..15 = consumers
constexpr int kXfConsumers = kXfThreads - 32;
constexpr int kXfStages = 4;
constexpr int kXfMaxRows = 256;
constexpr int kXfStageBudget = 24 * 1024;       // target bytes per ring stage
constexpr int kXfLutWords = 3 * 256 * 32;       // [c][entry][lane] 32-bit words = 96 KB

struct XfStagedGeom {
    int rows;          // output rows per tile
    int rowbuf;        // bytes per staged row (multiple of 16)
    int contiguous;    // crop rows are whole source rows back to back: one bulk copy per tile
    int tiles_per_image;
    int total_tiles;
    int gpr;           // 16-byte output groups per row
    uint32_t gpr_magic;
};

template <typename T>
struct XfPack;  // PX pixels of one channel -> one 16-byte store
template <>
struct XfPack<float> {
    static constexpr int PX = 4;
    __device__ static __forceinline__ uint4 pack(const uint32_t* v) { return make_uint4(v[0], v[1], v[2], v[3]); }
    __device__ static __forceinline__ uint32_t widen(float f) { return __float_as_uint(f); }
    __device__ static __forceinline__ float narrow(uint32_t w) { return __uint_as_float(w); }
};
template <>
struct XfPack<uint16_t> {   // bf16 and fp16: the table already holds the rounded 16-bit pattern
    static constexpr int PX = 8;
    __device__ static __forceinline__ uint4 pack(const uint32_t* v) {
        return make_uint4(__byte_perm(v[0], v[1], 0x5410), __byte_perm(v[2], v[3], 0x5410),
                          __byte_perm(v[4], v[5], 0x5410), __byte_perm(v[6], v[7], 0x5410));
    }
    __device__ static __forceinline__ uint32_t widen(uint16_t h) { return h; }
    __device__ static __forceinline__ uint16_t narrow(uint32_t w) { return (uint16_t)w; }
};

template <typename T>
__global__ void __launch_bounds__(kXfThreads, 1) tb_xf_staged_kernel(XfParams p, XfStagedGeom g) {
    constexpr int PX = XfPack<T>::PX;
    constexpr int NB = 3 * PX;            // source bytes per unit
    constexpr int NW = NB / 4;            // aligned words once shifted
    extern __shared__ __align__(128) uint8_t smem[];
    uint32_t* lut32 = reinterpret_cast<uint32_t*>(smem);
    uint8_t* ring = smem + kXfLutWords * 4;
    uint64_t* full = reinterpret_cast<uint64_t*>(ring + (size_t)kXfStages * g.rows * g.rowbuf);
    uint64_t* empty = full + kXfStages;
    __shared__ T s_plain[3 * 256];

    const int tid = threadIdx.x;
    const int warp = tid >> 5, lane = tid & 31;

    if (tid == 0) {
        for (int s = 0; s < kXfStages; ++s) {
            tb_mbar_init(&full[s], 1);
            tb_mbar_init(&empty[s], kXfConsumers / 32);
        }
        tb_mbar_fence_init();
    }
    for (int i = tid; i < 3 * 256; i += kXfThreads) s_plain[i] = ((const T*)p.lut)[i];
    __syncthreads();

    const int stage_bytes = g.rows * g.rowbuf;

    if (warp == 0) {
        // ===== producer: one bulk copy per tile row, issued by lane r =====
        int it = 0;
        for (int t = blockIdx.x; t < g.total_tiles; t += gridDim.x, ++it) {
            const int s = it % kXfStages;
            const uint32_t ph = (it / kXfStages) & 1;
            const int b = t / g.tiles_per_image;
            const int y0 = (t - b * g.tiles_per_image) * g.rows;
            const int nrows = min(g.rows, p.out_h - y0);
            const int top = (p.top ? p.top[b] : 0) - p.pad_y;
            const int left = (p.left ? p.left[b] : 0) - p.pad_x;
            const bool inside = top >= 0 && top + p.out_h <= p.H && left >= 0 && left + p.out_w <= p.W;
            tb_mbar_wait(&empty[s], ph ^ 1);
            if (!inside) {
                if (lane == 0) tb_mbar_arrive(&full[s]);
                continue;
            }
            const int boff = left * 3;
            const uint8_t* gsrc = p.src + (int64_t)b * p.bs + (int64_t)(top + y0) * p.pitch + (boff & ~15);
            uint8_t* sdst = ring + (size_t)s * stage_bytes;
            if (g.contiguous) {
                if (lane == 0) {
                    const uint32_t len = (uint32_t)nrows * (uint32_t)g.rowbuf;
                    tb_mbar_arrive_expect_tx(&full[s], len);
                    tb_bulk_g2s(sdst, gsrc, len, &full[s]);
                }
            } else {
                const uint32_t len = (uint32_t)(((boff & 15) + p.out_w * 3 + 15) & ~15);
                if (lane == 0) tb_mbar_arrive_expect_tx(&full[s], len * (uint32_t)nrows);
                __syncwarp();
                for (int r = lane; r < nrows; r += 32)
                    tb_bulk_g2s(sdst + (size_t)r * g.rowbuf, gsrc + (int64_t)r * p.pitch, len, &full[s]);
            }
        }
        return;
    }

    // ===== consumers =====
    const int ctid = tid - 32;
    // replicate the table: lut32[(c*256+e)*32 + lane] = entry, every lane reads its own bank
    for (int i = ctid; i < kXfLutWords; i += kXfConsumers) lut32[i] = XfPack<T>::widen(s_plain[i >> 5]);
    asm volatile("bar.sync 1, %0;" ::"n"(kXfConsumers) : "memory");
    const uint32_t* lutl = lut32 + lane;

    int it = 0;
    for (int t = blockIdx.x; t < g.total_tiles; t += gridDim.x, ++it) {
        const int s = it % kXfStages;
        const uint32_t ph = (it / kXfStages) & 1;
        const int b = t / g.tiles_per_image;
        const int y0 = (t - b * g.tiles_per_image) * g.rows;
        const int nrows = min(g.rows, p.out_h - y0);
        const int top = (p.top ? p.top[b] : 0) - p.pad_y;
        const int left = (p.left ? p.left[b] : 0) - p.pad_x;
        const bool flip = p.flip ? (p.flip[b] != 0) : false;
        const bool inside = top >= 0 && top + p.out_h <= p.H && left >= 0 && left + p.out_w <= p.W;
        T* __restrict__ dimg = (T*)p.dst + (int64_t)b * 3 * p.out_h * p.out_w;
        const int64_t plane = (int64_t)p.out_h * p.out_w;

        tb_mbar_wait(&full[s], ph);

        if (inside) {
            const uint8_t* stage = ring + (size_t)s * stage_bytes;
            const int a = (left * 3) & 15;
            const int units = nrows * g.gpr;
            for (int u = ctid; u < units; u += kXfConsumers) {
                const int r = (int)__umulhi((uint32_t)u, g.gpr_magic);
                const int gq = u - r * g.gpr;
                // first source byte of this unit inside the staged row
                const int start = a + 3 * (flip ? (p.out_w - PX - gq * PX) : gq * PX);
                const uint32_t* wp = reinterpret_cast<const uint32_t*>(stage + (size_t)r * g.rowbuf + (start & ~3));
                const int sh = (start & 3) * 8;
                uint32_t raw[NW + 1];
#pragma unroll
                for (int i = 0; i <= NW; ++i) raw[i] = wp[i];
                uint32_t q[NW];
#pragma unroll
                for (int i = 0; i < NW; ++i) q[i] = __funnelshift_r(raw[i], raw[i + 1], sh);
                T* drow = dimg + (int64_t)(y0 + r) * p.out_w + gq * PX;
#pragma unroll
                for (int c = 0; c < 3; ++c) {
                    uint32_t v[PX];
                    if (flip) {
#pragma unroll
                        for (int j = 0; j < PX; ++j) {
                            const int bi = 3 * (PX - 1 - j) + c;
                            const uint32_t e = (q[bi >> 2] >> (8 * (bi & 3))) & 0xFFu;
                            v[j] = lutl[(c * 256 + e) * 32];
                        }
                    } else {
#pragma unroll
                        for (int j = 0; j < PX; ++j) {
                            const int bi = 3 * j + c;
                            const uint32_t e = (q[bi >> 2] >> (8 * (bi & 3))) & 0xFFu;
                            v[j] = lutl[(c * 256 + e) * 32];
                        }
                    }
                    tb_st_cs_v4(drow + c * plane, XfPack<T>::pack(v));
                }
            }
        } else {
            // crop window touches padding: gather route for this tile
            const int npix = nrows * p.out_w;
            for (int i = ctid; i < npix; i += kXfConsumers) {
                const int r = i / p.out_w;
                const int x = i - r * p.out_w;
                const int xs = flip ? p.out_w - 1 - x : x;
                bool in = true;
                const int uu = tb_fold(top + y0 + r, p.H, p.pad_mode, in);
                const int vv = tb_fold(left + xs, p.W, p.pad_mode, in);
                const uint8_t* sp = p.src + (int64_t)b * p.bs + (int64_t)uu * p.pitch + (int64_t)vv * 3;
                T* d = dimg + (int64_t)(y0 + r) * p.out_w + x;
#pragma unroll
                for (int c = 0; c < 3; ++c) {
                    const uint32_t e = in ? (uint32_t)sp[c] : 0u;
                    d[c * plane] = XfPack<T>::narrow(lutl[(c * 256 + e) * 32]);
                }
            }
        }
        __syncwarp();
        if (lane == 0) tb_mbar_arrive(&empty[s]);
    }
}

template <typename T>
int launch_generic(const XfParams& p, cudaStream_t s) {
    const int threads = 256;
    const int64_t bpi = ((int64_t)p.out_h * p.out_w + threads - 1) / threads;
    const int64_t blocks = bpi * p.B;
    TB_REQUIRE(blocks < (int64_t)1 << 31, "tb_transform: grid too large");
    tb_xf_generic_kernel<T><<<(unsigned)blocks, threads, 0, s>>>(p, (int)bpi);
    TB_LAUNCH_CHECK("tb_xf_generic_kernel");
    return TB_OK;
}

template <typename T>
int launch_staged(const XfParams& p, const TbDevice& dev, cudaStream_t s, bool* taken) {
    constexpr int PX = XfPack<T>::PX;
    *taken = false;
    XfStagedGeom g;
    // Whole-row crops of a dense source land as one copy per tile; otherwise one copy per row
    // covering the 16-byte aligned superset of the crop row.
    g.contiguous = (p.W == p.out_w && p.pitch == (int64_t)p.W * 3) ? 1 : 0;
    g.rowbuf = g.contiguous ? (int)p.pitch : ((p.out_w * 3 + 15 + 15) & ~15);
    g.gpr = p.out_w / PX;
    const int fixed = kXfLutWords * 4 + 2 * kXfStages * 8 + 4096 /* static table copy + slack */;
    const int avail = dev.max_smem_optin - fixed;
    if (avail < kXfStages * g.rowbuf) return TB_OK;
    int max_rows = avail / (kXfStages * g.rowbuf);
    const int budget_rows = kXfStageBudget / g.rowbuf > 1 ? kXfStageBudget / g.rowbuf : 1;
    if (max_rows > budget_rows) max_rows = budget_rows;
    if (max_rows > kXfMaxRows) max_rows = kXfMaxRows;
    if (max_rows > p.out_h) max_rows = p.out_h;
    // Rows per tile: fewest consumer "rounds" per image (a round = every consumer thread takes one
    // 16-byte unit), then fewest tiles.
    int rows = 1;
    int64_t best_rounds = -1;
    for (int r = max_rows; r >= 1; --r) {
        const int full_tiles = p.out_h / r, tail = p.out_h - full_tiles * r;
        const int64_t rounds = (int64_t)full_tiles * ((r * g.gpr + kXfConsumers - 1) / kXfConsumers) +
                               (tail ? (tail * g.gpr + kXfConsumers - 1) / kXfConsumers : 0);
        if (best_rounds < 0 || rounds < best_rounds) { best_rounds = rounds; rows = r; }
    }
    g.rows = rows;
    g.tiles_per_image = (p.out_h + rows - 1) / rows;
    const int64_t total = (int64_t)g.tiles_per_image * p.B;
    if (total >= (int64_t)1 << 31) return TB_OK;
    g.total_tiles = (int)total;
    g.gpr_magic = (uint32_t)(((uint64_t)1 << 32) / (uint32_t)g.gpr) + 1u;
    if ((int64_t)rows * g.gpr >= 60000) return TB_OK;  // keeps the magic division exact
    const size_t smem = (size_t)kXfLutWords * 4 + (size_t)kXfStages * rows * g.rowbuf + 2 * kXfStages * 8;
    static bool attr_set[64] = {};
    if (!attr_set[dev.device]) {
        TB_CUDA(cudaFuncSetAttribute(tb_xf_staged_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     dev.max_smem_optin - 4096 + 2 * kXfStages * 8));
        attr_set[dev.device] = true;
    }
    const int grid = (int)(total < dev.sm_count ? total : dev.sm_count);
    tb_xf_staged_kernel<T><<<grid, kXfThreads, smem, s>>>(p, g);
    TB_LAUNCH_CHECK("tb_xf_staged_kernel");
    *taken = true;
    return TB_OK;
}

}  // namespace

extern "C" int tb_transform_u8_nhwc_to_nchw(const uint8_t* src, int B, int H, int W, int C, int64_t src_batch_stride,
                                            int64_t src_row_pitch, void* dst, int dst_dtype, int out_h, int out_w,
                                            const int32_t* crop_top, const int32_t* crop_left, const uint8_t* flip,
                                            int pad_y, int pad_x, int pad_mode, const void* lut, int flags,
                                            tb_stream_t stream) {
    TB_REQUIRE(B >= 0 && H >= 1 && W >= 1, "tb_transform: bad source shape B=%d H=%d W=%d", B, H, W);
    TB_REQUIRE(C >= 1 && C <= 4, "tb_transform: C=%d not in [1,4]", C);
    TB_REQUIRE(out_h >= 1 && out_w >= 1, "tb_transform: bad output size %dx%d", out_h, out_w);
    TB_REQUIRE(tb_dtype_size(dst_dtype) != 0, "tb_transform: unknown dst_dtype %d", dst_dtype);
    TB_REQUIRE(pad_mode >= TB_PAD_NONE && pad_mode <= TB_PAD_SYMMETRIC, "tb_transform: unknown pad_mode %d", pad_mode);
    TB_REQUIRE(pad_y >= 0 && pad_x >= 0, "tb_transform: negative padding");
    if (pad_mode == TB_PAD_REFLECT)
        TB_REQUIRE(pad_y <= H - 1 && pad_x <= W - 1, "tb_transform: reflect padding must be smaller than the image");
    if (pad_mode == TB_PAD_SYMMETRIC)
        TB_REQUIRE(pad_y <= H && pad_x <= W, "tb_transform: symmetric padding must not exceed the image");
    TB_REQUIRE(src_row_pitch >= (int64_t)W * C, "tb_transform: src_row_pitch %lld < W*C", (long long)src_row_pitch);
    TB_REQUIRE(src_batch_stride >= 0, "tb_transform: negative src_batch_stride");
    if (B == 0) return TB_OK;
    TB_REQUIRE(src && dst && lut, "tb_transform: NULL src/dst/lut");
    TB_REQUIRE(((uintptr_t)dst % tb_dtype_size(dst_dtype)) == 0, "tb_transform: dst misaligned for its dtype");

    XfParams p;
    p.src = src; p.B = B; p.H = H; p.W = W; p.C = C; p.bs = src_batch_stride; p.pitch = src_row_pitch;
    p.dst = dst; p.out_h = out_h; p.out_w = out_w;
    p.top = crop_top; p.left = crop_left; p.flip = flip;
    p.pad_y = pad_y; p.pad_x = pad_x; p.pad_mode = pad_mode; p.lut = lut;
    cudaStream_t s = (cudaStream_t)stream;

    TbDevice dev;
    int rc = tb_device(&dev);
    if (rc) return rc;

    const int px = 16 / tb_dtype_size(dst_dtype);
    const bool fast_ok = !(flags & TB_FLAG_FORCE_GENERIC) && C == 3 && (out_w % px) == 0 &&
                         ((uintptr_t)dst % 16) == 0 && ((uintptr_t)src % 16) == 0 && (src_row_pitch % 16) == 0 &&
                         (src_batch_stride % 16) == 0;
    if (fast_ok) {
        bool taken = false;
        rc = (dst_dtype == TB_F32) ? launch_staged<float>(p, dev, s, &taken) : launch_staged<uint16_t>(p, dev, s, &taken);
        if (rc) return rc;
        if (taken) return TB_OK;
    }
    if (dst_dtype == TB_F32) return launch_generic<float>(p, s);
    return launch_generic<uint16_t>(p, s);
}
