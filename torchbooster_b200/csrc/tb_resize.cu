// Pillow-exact antialiased bilinear resize fused with crop / flip / table lookup / cast (sm_100a).
//
// Replaces T.Resize + T.CenterCrop + T.ToTensor + T.Normalize of the reference's style-transfer
// examples (examples/img_stt/adain/adain.py:170, online/online.py:171-173, offline/offline.py:110),
// i.e. torchvision's PIL path -> PIL.Image.resize(BILINEAR) -> libImaging ImagingResample:
// 22-bit fixed-point taps, horizontal pass first into a uint8-ROUNDED intermediate, then the
// vertical pass.  Both roundings are reproduced, so the result is bit-identical to Pillow.
//
// One launch, no global intermediate: a CTA owns a tile of final output pixels; phase 1 runs the
// horizontal pass for exactly the input rows / resized columns that tile needs into shared
// memory, phase 2 runs the vertical pass out of shared memory, looks the byte up and stores.
#include <math.h>
#include <type_traits>
#include <vector>

#include "tb_common.cuh"

// ----------------------------------------------------------------------------------------------
// host: filter taps (libImaging/Resample.c precompute_coeffs + normalize_coeffs_8bpc, bilinear)
// ----------------------------------------------------------------------------------------------
static const int kPrecisionBits = 32 - 8 - 2;

extern "C" int tb_resize_ksize(int in_size, int out_size) {
    if (in_size < 1 || out_size < 1) return tb_fail(TB_EINVAL, "tb_resize_ksize: sizes must be >= 1");
    double scale = (double)in_size / (double)out_size;
    double filterscale = scale < 1.0 ? 1.0 : scale;
    double support = 1.0 * filterscale;
    return (int)ceil(support) * 2 + 1;
}

extern "C" int tb_resize_taps(int in_size, int out_size, int32_t* bounds, int32_t* taps) {
    TB_REQUIRE(in_size >= 1 && out_size >= 1, "tb_resize_taps: sizes must be >= 1");
    TB_REQUIRE(bounds && taps, "tb_resize_taps: NULL output");
    const double scale = (double)in_size / (double)out_size;
    const double filterscale = scale < 1.0 ? 1.0 : scale;
    const double support = 1.0 * filterscale;
    const int ksize = (int)ceil(support) * 2 + 1;
    const double ss = 1.0 / filterscale;
    std::vector<double> w(ksize);
    for (int xx = 0; xx < out_size; ++xx) {
        const double center = (xx + 0.5) * scale;
        int xmin = (int)(center - support + 0.5);
        if (xmin < 0) xmin = 0;
        int xmax = (int)(center + support + 0.5);
        if (xmax > in_size) xmax = in_size;
        xmax -= xmin;
        double ww = 0.0;
        for (int x = 0; x < xmax; ++x) {
            double a = (x + xmin - center + 0.5) * ss;
            if (a < 0.0) a = -a;
            const double v = a < 1.0 ? 1.0 - a : 0.0;
            w[x] = v;
            ww += v;
        }
        for (int x = 0; x < xmax; ++x)
            if (ww != 0.0) w[x] /= ww;
        for (int x = xmax; x < ksize; ++x) w[x] = 0.0;
        for (int x = 0; x < ksize; ++x) {
            const double v = w[x];
            taps[(size_t)xx * ksize + x] =
                v < 0 ? (int)(-0.5 + v * (1 << kPrecisionBits)) : (int)(0.5 + v * (1 << kPrecisionBits));
        }
        bounds[2 * xx] = xmin;
        bounds[2 * xx + 1] = xmax;
    }
    return TB_OK;
}

// ----------------------------------------------------------------------------------------------
// device
// ----------------------------------------------------------------------------------------------
namespace {

struct RsParams {
    const uint8_t* src;
    int B, H, W;
    int64_t bs, pitch;
    int rs_h, rs_w;
    const int32_t* bx; const int32_t* kx; int ksx;
    const int32_t* by; const int32_t* ky; int ksy;
    void* dst;
    int out_h, out_w;
    const int32_t* top; const int32_t* left; const uint8_t* flip;
    int off_y, off_x;
    const void* lut;
    int TR, TC, tiles_x, tiles_y;
    int ty_per_cta, chunks_y;   // fast kernel: row tiles per CTA and CTAs per (sample, column tile)
    int row_bytes;     // bytes per intermediate row in shared memory (multiple of 4)
    int words_ok;      // source rows may be read as aligned 32-bit words
    const int32_t* box;  // resized-crop mode: per-sample source box [B][4] = top, left, height, width
    int box_words;       // 4, or 10 for a ragged batch: + resized h, w, crop offset y, x, image h, w
    const int64_t* offsets;   // ragged batch: byte offset of every image from src (else b * bs)
    const int32_t* pitches;   // ragged batch: row pitch of every image in bytes (else pitch)
};

__device__ __forceinline__ uint8_t tb_clip8(int acc) {
    int v = acc >> kPrecisionBits;
    v = v < 0 ? 0 : (v > 255 ? 255 : v);
    return (uint8_t)v;
}

// C bytes at p (any alignment) packed little-endian into one word.
template <int C>
__device__ __forceinline__ uint32_t tb_load_px(const uint8_t* p, bool words_ok) {
    if (C == 1) return *p;
    if (words_ok) {
        const uint32_t* wp = reinterpret_cast<const uint32_t*>((uintptr_t)p & ~(uintptr_t)3);
        const int sh = (int)((uintptr_t)p & 3);
        uint32_t lo = __ldg(wp);
        if (sh + C > 4) {
            const uint32_t hi = __ldg(wp + 1);
            lo = __funnelshift_r(lo, hi, sh * 8);
        } else {
            lo >>= sh * 8;
        }
        return lo;
    }
    uint32_t w = 0;
#pragma unroll
    for (int c = 0; c < C; ++c) w |= (uint32_t)p[c] << (8 * c);
    return w;
}

template <typename T, int C>
__global__ void __launch_bounds__(256) tb_resize_fused_kernel(RsParams p) {
    extern __shared__ __align__(16) uint8_t inter[];   // [nin][row_bytes]
    __shared__ T s_lut[C * 256];
    for (int i = threadIdx.x; i < C * 256; i += blockDim.x) s_lut[i] = ((const T*)p.lut)[i];

    int t = blockIdx.x;
    const int tx = t % p.tiles_x; t /= p.tiles_x;
    const int ty = t % p.tiles_y;
    const int b = t / p.tiles_y;
    const int y0 = ty * p.TR, x0 = tx * p.TC;
    const int nr = min(p.TR, p.out_h - y0), nc = min(p.TC, p.out_w - x0);
    const int oy = (p.top ? p.top[b] : 0) + p.off_y;
    const int ox = (p.left ? p.left[b] : 0) + p.off_x;
    const bool flip = p.flip ? (p.flip[b] != 0) : false;

    // resized rows / columns this tile touches
    const int ry0 = max(oy + y0, 0), ry1 = min(oy + y0 + nr, p.rs_h);
    const int xs0 = flip ? p.out_w - x0 - nc : x0;
    const int rx0 = max(ox + xs0, 0), rx1 = min(ox + xs0 + nc, p.rs_w);
    const bool any = ry1 > ry0 && rx1 > rx0;
    const bool vpass = p.rs_h != p.H, hpass = p.rs_w != p.W;

    int iy0 = 0, nin = 0;
    if (any) {
        if (vpass) {
            iy0 = p.by[2 * ry0];
            nin = p.by[2 * (ry1 - 1)] + p.by[2 * (ry1 - 1) + 1] - iy0;
        } else {
            iy0 = ry0;
            nin = ry1 - ry0;
        }
    }
    const int ncols = any ? rx1 - rx0 : 0;
    const uint8_t* simg = p.src + (int64_t)b * p.bs;

    // ---- phase 1: horizontal pass (or copy) of input rows [iy0, iy0+nin), columns [rx0, rx1)
    for (int i = threadIdx.x; i < nin * ncols; i += blockDim.x) {
        const int r = i / ncols;
        const int cx = i - r * ncols;
        const int rx = rx0 + cx;
        const uint8_t* row = simg + (int64_t)(iy0 + r) * p.pitch;
        uint8_t* o = inter + (size_t)r * p.row_bytes + cx * C;
        if (hpass) {
            const int xmin = p.bx[2 * rx], n = p.bx[2 * rx + 1];
            const int32_t* k = p.kx + (size_t)rx * p.ksx;
            int acc[C];
#pragma unroll
            for (int c = 0; c < C; ++c) acc[c] = 1 << (kPrecisionBits - 1);
            for (int j = 0; j < n; ++j) {
                const uint32_t w = tb_load_px<C>(row + (size_t)(xmin + j) * C, p.words_ok != 0);
                const int kk = __ldg(k + j);
#pragma unroll
                for (int c = 0; c < C; ++c) acc[c] += (int)((w >> (8 * c)) & 0xffu) * kk;
            }
#pragma unroll
            for (int c = 0; c < C; ++c) o[c] = tb_clip8(acc[c]);
        } else {
            const uint32_t w = tb_load_px<C>(row + (size_t)rx * C, p.words_ok != 0);
#pragma unroll
            for (int c = 0; c < C; ++c) o[c] = (uint8_t)(w >> (8 * c));
        }
    }
    __syncthreads();

    // ---- phase 2: vertical pass out of shared memory, table lookup, NCHW store
    T* dimg = (T*)p.dst + (int64_t)b * C * p.out_h * p.out_w;
    const int64_t plane = (int64_t)p.out_h * p.out_w;
    for (int i = threadIdx.x; i < nr * nc; i += blockDim.x) {
        const int yy = i / nc;
        const int xx = i - yy * nc;
        const int y = y0 + yy, x = x0 + xx;
        const int ry = oy + y;
        const int rx = ox + (flip ? p.out_w - 1 - x : x);
        uint32_t e[C];
#pragma unroll
        for (int c = 0; c < C; ++c) e[c] = 0;
        if (ry >= 0 && ry < p.rs_h && rx >= 0 && rx < p.rs_w) {
            const uint8_t* col = inter + (size_t)(rx - rx0) * C;
            if (vpass) {
                const int ymin = p.by[2 * ry] - iy0, n = p.by[2 * ry + 1];
                const int32_t* k = p.ky + (size_t)ry * p.ksy;
                int acc[C];
#pragma unroll
                for (int c = 0; c < C; ++c) acc[c] = 1 << (kPrecisionBits - 1);
                for (int j = 0; j < n; ++j) {
                    const uint8_t* q = col + (size_t)(ymin + j) * p.row_bytes;
                    const int kk = __ldg(k + j);
#pragma unroll
                    for (int c = 0; c < C; ++c) acc[c] += (int)q[c] * kk;
                }
#pragma unroll
                for (int c = 0; c < C; ++c) e[c] = tb_clip8(acc[c]);
            } else {
                const uint8_t* q = col + (size_t)(ry - iy0) * p.row_bytes;
#pragma unroll
                for (int c = 0; c < C; ++c) e[c] = q[c];
            }
        }
        T* d = dimg + (int64_t)y * p.out_w + x;
#pragma unroll
        for (int c = 0; c < C; ++c) d[c * plane] = s_lut[c * 256 + e[c]];
    }
}


// ----------------------------------------------------------------------------------------------
// fast path: both passes active, scale factor <= 4 on both axes (so every output touches at most
// NT = 4 input samples for scales up to 2, NT = 8 up to 4: n <= ceil(2 * support)), source rows
// readable as aligned 32-bit words.
//
// Four 8-bit samples times four 22-bit taps is three DP4A instructions: split each tap into its
// three bytes, k = k0 + 2^8 k1 + 2^16 k2, then sum(p*k) = D0 + 2^8 D1 + 2^16 D2 with
// Dm = dp4a(samples, {km of the 4 taps}); every Dm < 2^18, the chain is exact (eight taps: two
// DP4A per byte plane).  The rounding constant 2^21 rides in as 32 << 16.  The result needs no
// clip: taps are >= 0 and sum to at most 2^22 + NT/2, so (2^21 + 255 * sum) >> 22 <= 255.
//
//   phase 1  horizontal pass, thread <-> resized column for all input rows of the tile: its taps
//            are packed registers; a row's NT*C-byte window comes as aligned words + funnel
//            shift, two PRMTs per channel and tap group gather the samples; results go to shared
//            memory TRANSPOSED and planar, inter[c][column][row], so that ...
//   phase 2  ... the vertical pass finds its NT samples (consecutive rows) in one unaligned word
//            group (LDS + funnel shift), runs the same DP4A chain, looks the byte up and stores.
// The column pitch is 4 * odd bytes: lanes (= consecutive columns) never share a bank.
// ----------------------------------------------------------------------------------------------
template <int G>   // G tap groups of four
__device__ __forceinline__ uint32_t tb_dot(const uint32_t* px, const uint32_t (*kb)[3]) {
    uint32_t a = 32u;
#pragma unroll
    for (int g = 0; g < G; ++g) a = __dp4a(px[g], kb[g][2], a);
    a <<= 8;
#pragma unroll
    for (int g = 0; g < G; ++g) a = __dp4a(px[g], kb[g][1], a);
    a <<= 8;
#pragma unroll
    for (int g = 0; g < G; ++g) a = __dp4a(px[g], kb[g][0], a);
    return a >> kPrecisionBits;
}
// byte m of each tap of a group of four -> one word per m
template <int G>
__device__ __forceinline__ void tb_pack_taps(const int* k, uint32_t (*kb)[3]) {
#pragma unroll
    for (int g = 0; g < G; ++g)
#pragma unroll
        for (int m = 0; m < 3; ++m)
            kb[g][m] = ((uint32_t)(k[4 * g] >> (8 * m)) & 0xffu) | (((uint32_t)(k[4 * g + 1] >> (8 * m)) & 0xffu) << 8) |
                       (((uint32_t)(k[4 * g + 2] >> (8 * m)) & 0xffu) << 16) |
                       (((uint32_t)(k[4 * g + 3] >> (8 * m)) & 0xffu) << 24);
}

constexpr int kRsCtasPerSm = 16;     // a CTA takes several row tiles only while the grid keeps this many CTAs per SM

// One output position's taps, computed on the device exactly as tb_resize_taps does on the host
// (and Pillow in C): float64, explicit round-to-nearest operations in Pillow's evaluation order,
// no FMA contraction.  Valid while at most NT taps are non-zero (support <= NT / 2).
template <int NT>
struct TapN {
    int xmin, n;
    int k[NT];
};
template <int NT>
__device__ __forceinline__ TapN<NT> tb_bilinear_tap(int in_size, int out_size, int xx) {
    const double scale = __ddiv_rn((double)in_size, (double)out_size);
    const double fs = scale < 1.0 ? 1.0 : scale;      // filterscale; support = 1.0 * filterscale
    const double center = __dmul_rn((double)xx + 0.5, scale);
    const double ss = __ddiv_rn(1.0, fs);
    TapN<NT> t;
    t.xmin = (int)__dadd_rn(__dsub_rn(center, fs), 0.5);
    if (t.xmin < 0) t.xmin = 0;
    int xmax = (int)__dadd_rn(__dadd_rn(center, fs), 0.5);
    if (xmax > in_size) xmax = in_size;
    t.n = xmax - t.xmin;
    double w[NT], ww = 0.0;
#pragma unroll
    for (int x = 0; x < NT; ++x) {
        double a = __dmul_rn(__dadd_rn(__dsub_rn((double)(x + t.xmin), center), 0.5), ss);
        if (a < 0.0) a = -a;
        w[x] = (x < t.n && a < 1.0) ? __dsub_rn(1.0, a) : 0.0;
        if (x < t.n) ww = __dadd_rn(ww, w[x]);
    }
#pragma unroll
    for (int x = 0; x < NT; ++x) {
        double v = w[x];
        if (x < t.n && ww != 0.0) v = __ddiv_rn(v, ww);
        t.k[x] = x < t.n ? (int)__dadd_rn(0.5, __dmul_rn(v, 4194304.0)) : 0;
    }
    return t;
}

// BOX = false: T.Resize of the whole image, taps from the host-built tables, then crop / flip.
// BOX = true : T.RandomResizedCrop -- every sample resizes its own source box to the output size;
//              taps are computed in the kernel (tb_bilinear_tap), no tables, no extra crop.
template <typename T, int C, bool BOX, int NT>
__global__ void __launch_bounds__(256) tb_resize_fast_kernel(RsParams p) {
    constexpr int G = NT / 4;
    extern __shared__ __align__(16) uint8_t inter[];  // [C][TC][col_pitch]
    __shared__ T s_lut[C * 256];
    __shared__ uint32_t s_kyb[32][G][3];
    __shared__ int s_ymin[32], s_yn[32];
    tb_grid_dependency_wait();
    for (int i = threadIdx.x; i < C * 256; i += 256) s_lut[i] = ((const T*)p.lut)[i];

    // CTA <-> (sample b, column tile tx, a run of row tiles): the per-column set-up (taps, source
    // window) is done once and reused for every row tile of the run
    int t = blockIdx.x;
    const int tx = t % p.tiles_x; t /= p.tiles_x;
    const int chunk = t % p.chunks_y;
    const int b = t / p.chunks_y;
    const int x0 = tx * p.TC;
    const int nc = min(p.TC, p.out_w - x0);
    // source box of this sample (whole image unless BOX) and placement of the output in the resized image
    int box_top = 0, box_left = 0, in_h = p.H, in_w = p.W;
    if constexpr (BOX) {
        box_top = p.box[4 * b]; box_left = p.box[4 * b + 1]; in_h = p.box[4 * b + 2]; in_w = p.box[4 * b + 3];
        // a box the host did not vouch for (outside the image, or too large for NT taps) yields zeros
        if (box_top < 0 || box_left < 0 || in_h < 1 || in_w < 1 || box_top + in_h > p.H || box_left + in_w > p.W ||
            in_h > (NT / 2) * p.rs_h || in_w > (NT / 2) * p.rs_w)
            in_h = in_w = 0;
    }
    const int oy = BOX ? 0 : (p.top ? p.top[b] : 0) + p.off_y;
    const int ox = BOX ? 0 : (p.left ? p.left[b] : 0) + p.off_x;
    const bool flip = p.flip ? (p.flip[b] != 0) : false;
    const int xs0 = flip ? p.out_w - x0 - nc : x0;
    const int rx0 = max(ox + xs0, 0), rx1 = min(ox + xs0 + nc, p.rs_w);
    const int ncols = (rx1 > rx0 && in_h > 0) ? rx1 - rx0 : 0;      // resized columns this CTA needs (<= TC <= 256)
    const int cp = p.row_bytes;                      // column pitch (bytes), 4 * odd
    const int plane_b = p.TC * cp;                   // bytes per channel plane of the intermediate
    const uint8_t* simg = p.src + (int64_t)b * p.bs;
    const int pitch_w = (int)(p.pitch >> 2);
    constexpr int NQ = NT * C / 4;                   // aligned words of an NT-sample window after the shift

    // ---- this thread's column in the horizontal pass
    const int ngrp1 = ncols ? max(1, 256 / ncols) : 1;
    const int grp1 = ncols ? threadIdx.x / ncols : 0;
    const bool active1 = ncols > 0 && grp1 < ngrp1;
    uint32_t kbx[G][3];
#pragma unroll
    for (int g = 0; g < G; ++g) kbx[g][0] = kbx[g][1] = kbx[g][2] = 0;
    int a1 = 0, nw1 = 0;
    const uint32_t* wp0 = nullptr;
    uint8_t* o0 = inter;
    if (active1) {
        const int cx = threadIdx.x - grp1 * ncols;
        const int rx = rx0 + cx;
        TapN<NT> tp;
        if constexpr (BOX) {
            tp = tb_bilinear_tap<NT>(in_w, p.rs_w, rx);
        } else {
            tp.xmin = p.bx[2 * rx]; tp.n = p.bx[2 * rx + 1];
#pragma unroll
            for (int j = 0; j < NT; ++j) tp.k[j] = (j < p.ksx && j < tp.n) ? __ldg(p.kx + (size_t)rx * p.ksx + j) : 0;
        }
        tb_pack_taps<G>(tp.k, kbx);
        const int boff = (box_left + tp.xmin) * C;
        a1 = boff & 3;
        nw1 = (a1 + tp.n * C + 3) >> 2;                  // words that hold this window (<= NQ + 1)
        wp0 = reinterpret_cast<const uint32_t*>(simg + (boff - a1)) + (size_t)box_top * pitch_w;
        o0 = inter + (size_t)cx * cp;
    }
    // ---- this thread's column in the vertical pass
    const int ngrp2 = max(1, 256 / nc);
    const int grp2 = threadIdx.x / nc;
    const bool active2 = grp2 < ngrp2;
    const int x = x0 + threadIdx.x - grp2 * nc;
    const int rx2 = ox + (flip ? p.out_w - 1 - x : x);
    const bool colin = active2 && ncols > 0 && rx2 >= rx0 && rx2 < rx1;
    const uint32_t col = tb_smem_addr(inter) + (uint32_t)((colin ? rx2 - rx0 : 0) * cp);
    T* dcol = (T*)p.dst + (int64_t)b * C * p.out_h * p.out_w + x;
    const int64_t plane = (int64_t)p.out_h * p.out_w;

    const int ty_end = min(p.tiles_y, (chunk + 1) * p.ty_per_cta);
    for (int ty = chunk * p.ty_per_cta; ty < ty_end; ++ty) {
        const int y0 = ty * p.TR;
        const int nr = min(p.TR, p.out_h - y0);
        const int ry0 = max(oy + y0, 0), ry1 = min(oy + y0 + nr, p.rs_h);
        const bool any = ncols > 0 && ry1 > ry0;

        // vertical taps of the resized rows this tile touches, packed for DP4A
        if (any && threadIdx.x < ry1 - ry0) {
            const int ry = ry0 + threadIdx.x;
            TapN<NT> tp;
            if constexpr (BOX) {
                tp = tb_bilinear_tap<NT>(in_h, p.rs_h, ry);
            } else {
                tp.xmin = p.by[2 * ry]; tp.n = p.by[2 * ry + 1];
#pragma unroll
                for (int j = 0; j < NT; ++j) tp.k[j] = (j < p.ksy && j < tp.n) ? p.ky[(size_t)ry * p.ksy + j] : 0;
            }
            tb_pack_taps<G>(tp.k, s_kyb[threadIdx.x]);
            s_ymin[threadIdx.x] = tp.xmin;
            s_yn[threadIdx.x] = tp.n;
        }
        __syncthreads();
        const int iy0 = any ? s_ymin[0] : 0;                                            // first input row of the tile
        const int nin = any ? s_ymin[ry1 - ry0 - 1] + s_yn[ry1 - ry0 - 1] - iy0 : 0;    // input rows the tile touches

        // ---- phase 1: horizontal pass; thread <-> column, looping over the input rows
        if (any && active1) {
            constexpr int UR = NT == 4 ? 4 : 2;              // rows in flight per thread
            const uint32_t* wp = wp0 + (size_t)(iy0 + grp1) * pitch_w;
            uint8_t* o = o0 + grp1;
            for (int r = grp1; r < nin; r += UR * ngrp1, wp += (size_t)UR * ngrp1 * pitch_w, o += UR * ngrp1) {
                uint32_t w[UR][NQ + 1];
#pragma unroll
                for (int u = 0; u < UR; ++u) {
                    const bool ok = r + u * ngrp1 < nin;
#pragma unroll
                    for (int i = 0; i <= NQ; ++i) w[u][i] = (ok && i < nw1) ? __ldg(wp + (size_t)u * ngrp1 * pitch_w + i) : 0u;
                }
#pragma unroll
                for (int u = 0; u < UR; ++u) {
                    if (r + u * ngrp1 >= nin) break;
                    uint32_t q[NQ];
#pragma unroll
                    for (int i = 0; i < NQ; ++i) q[i] = __funnelshift_r(w[u][i], w[u][i + 1], a1 * 8);
                    uint8_t* ou = o + u * ngrp1;
                    if (C == 3) {
                        // samples of channel c sit at bytes c, c+3, c+6, c+9 of each 12-byte group of four pixels
                        uint32_t p0[G], p1[G], p2[G];
#pragma unroll
                        for (int g = 0; g < G; ++g) {
                            p0[g] = __byte_perm(__byte_perm(q[3 * g], q[3 * g + 1], 0x0630), q[3 * g + 2], 0x5210);
                            p1[g] = __byte_perm(__byte_perm(q[3 * g], q[3 * g + 1], 0x0741), q[3 * g + 2], 0x6210);
                            p2[g] = __byte_perm(__byte_perm(q[3 * g], q[3 * g + 1], 0x0052), q[3 * g + 2], 0x7410);
                        }
                        ou[0] = (uint8_t)tb_dot<G>(p0, kbx);
                        ou[plane_b] = (uint8_t)tb_dot<G>(p1, kbx);
                        ou[2 * plane_b] = (uint8_t)tb_dot<G>(p2, kbx);
                    } else {
                        ou[0] = (uint8_t)tb_dot<G>(q, kbx);
                    }
                }
            }
        }
        __syncthreads();

        // ---- phase 2: vertical pass out of shared memory; thread <-> output column, looping over rows
        if (active2) {
            for (int yy = grp2; yy < nr; yy += ngrp2) {
                const int y = y0 + yy;
                const int ry = oy + y;
                uint32_t e[C];
#pragma unroll
                for (int c = 0; c < C; ++c) e[c] = 0;
                if (any && colin && ry >= ry0 && ry < ry1) {
                    const int rr = ry - ry0;
                    uint32_t kb[G][3];
#pragma unroll
                    for (int g = 0; g < G; ++g) { kb[g][0] = s_kyb[rr][g][0]; kb[g][1] = s_kyb[rr][g][1]; kb[g][2] = s_kyb[rr][g][2]; }
                    const int ymin = s_ymin[rr] - iy0;
                    const uint32_t wa = col + (uint32_t)(ymin & ~3);
                    const int sh = (ymin & 3) * 8;
#pragma unroll
                    for (int c = 0; c < C; ++c) {
                        uint32_t w[G + 1], px[G];
#pragma unroll
                        for (int i = 0; i <= G; ++i) asm("ld.shared.u32 %0, [%1];" : "=r"(w[i]) : "r"(wa + c * plane_b + 4 * i));
#pragma unroll
                        for (int g = 0; g < G; ++g) px[g] = __funnelshift_r(w[g], w[g + 1], sh);
                        e[c] = tb_dot<G>(px, kb);
                    }
                }
                T* d = dcol + (int64_t)y * p.out_w;
#pragma unroll
                for (int c = 0; c < C; ++c) d[c * plane] = s_lut[c * 256 + e[c]];
            }
        }
        __syncthreads();     // the next row tile overwrites the intermediate and the tap tables
    }
}


// ----------------------------------------------------------------------------------------------
// sliding-window variant of the fast path: NO shared-memory intermediate.
//
// A thread owns one output column and walks down the input rows of its strip.  For every input
// row it runs the horizontal pass for its column (same DP4A chain) and pushes the uint8 result
// into a per-channel register window holding the last NT rows -- exactly the NT samples the
// vertical pass of the next output row needs, already packed for DP4A.  Whenever the window's
// newest row is the last row an output row depends on (s_ytop), that output row is emitted:
// vertical DP4A chain on the window registers, table lookup, coalesced store.  All threads of
// the CTA see the same row schedule, so there is no divergence, no __syncthreads in the loop,
// and none of the transposed stores / unaligned window loads of the two-phase kernel.
// Works for every scale the NT taps cover, up- and down-scaling (an output row is emitted after
// 0, 1 or several pushes).
// ----------------------------------------------------------------------------------------------
constexpr int kSlideMaxTR = 64;

// v = *p when take, else v keeps its bits (no zeroing, no branch)
__device__ __forceinline__ void tb_ldg_if(uint32_t& v, const uint32_t* p, bool take) {
    asm("{\n\t.reg .pred q;\n\tsetp.ne.b32 q, %2, 0;\n\t@q ld.global.nc.u32 %0, [%1];\n\t}" : "+r"(v) : "l"(p), "r"((int)take));
}

template <typename T, int C, bool BOX, int NT>
__global__ void __launch_bounds__(256) tb_resize_slide_kernel(RsParams p) {
    constexpr int G = NT / 4;
    constexpr int NQ = NT * C / 4;                   // aligned words of an NT-sample window after the shift
    constexpr int UR = NT == 4 ? 4 : 2;              // input rows per register buffer (two buffers)
    __shared__ T s_lut[C * 256];
    __shared__ uint32_t s_kyb[kSlideMaxTR][G][3];
    __shared__ int s_ytop[kSlideMaxTR];              // last input row an output row needs; -1: outside the resized image
    tb_grid_dependency_wait();
    for (int i = threadIdx.x; i < C * 256; i += blockDim.x) s_lut[i] = ((const T*)p.lut)[i];

    int t = blockIdx.x;
    const int tx = t % p.tiles_x; t /= p.tiles_x;
    const int ty = t % p.tiles_y;
    const int b = t / p.tiles_y;
    const int x0 = tx * p.TC, y0 = ty * p.TR;
    const int nc = min(p.TC, p.out_w - x0), nr = min(p.TR, p.out_h - y0);
    int box_top = 0, box_left = 0, in_h = p.H, in_w = p.W, rs_h = p.rs_h, rs_w = p.rs_w;
    int oy = BOX ? 0 : (p.top ? p.top[b] : 0) + p.off_y;
    int ox = BOX ? 0 : (p.left ? p.left[b] : 0) + p.off_x;
    int frame_h = p.H, frame_w = p.W;
    if constexpr (BOX) {
        const int32_t* g = p.box + (int64_t)p.box_words * b;
        box_top = g[0]; box_left = g[1]; in_h = g[2]; in_w = g[3];
        if (p.box_words == 10) {         // ragged batch: this sample's own resized size, crop offset and image size
            rs_h = g[4]; rs_w = g[5]; oy = g[6]; ox = g[7]; frame_h = g[8]; frame_w = g[9];
        }
        // a box the host did not vouch for (outside the image, or too large for NT taps) yields zeros
        if (box_top < 0 || box_left < 0 || in_h < 1 || in_w < 1 || box_top + in_h > frame_h || box_left + in_w > frame_w ||
            rs_h < 1 || rs_w < 1 || in_h > (NT / 2) * rs_h || in_w > (NT / 2) * rs_w)
            in_h = in_w = 0;
    }
    const bool flip = p.flip ? (p.flip[b] != 0) : false;
    const uint8_t* simg = p.src + (p.offsets ? p.offsets[b] : (int64_t)b * p.bs);
    const int64_t pitch = p.pitches ? (int64_t)p.pitches[b] : p.pitch;

    // ---- vertical taps of this strip's output rows
    if ((int)threadIdx.x < nr) {
        const int ry = oy + y0 + (int)threadIdx.x;
        int top = -1;
        if (ry >= 0 && ry < rs_h && in_h > 0) {
            TapN<NT> tp;
            if constexpr (BOX) {
                tp = tb_bilinear_tap<NT>(in_h, rs_h, ry);
            } else {
                tp.xmin = p.by[2 * ry]; tp.n = p.by[2 * ry + 1];
#pragma unroll
                for (int j = 0; j < NT; ++j) tp.k[j] = (j < p.ksy && j < tp.n) ? p.ky[(size_t)ry * p.ksy + j] : 0;
            }
            tb_pack_taps<G>(tp.k, s_kyb[threadIdx.x]);
            top = tp.xmin + NT - 1;
        }
        s_ytop[threadIdx.x] = top;
    }
    // ---- this thread's column: horizontal taps and source window
    const bool active = (int)threadIdx.x < nc;
    const int x = x0 + (int)threadIdx.x;
    const int rx = ox + (flip ? p.out_w - 1 - x : x);
    const bool colin = active && in_w > 0 && rx >= 0 && rx < rs_w;
    // A column outside the resized image keeps all-zero taps: its horizontal results, its window
    // and its outputs are then 0 without any special casing below.
    uint32_t kbx[G][3];
#pragma unroll
    for (int g = 0; g < G; ++g) kbx[g][0] = kbx[g][1] = kbx[g][2] = 0;
    int a1 = 0, nw1 = 1;
    const int64_t pitch_w = pitch >> 2;
    const uint32_t* wbase = reinterpret_cast<const uint32_t*>(simg);      // always a readable word
    if (colin) {
        TapN<NT> tp;
        if constexpr (BOX) {
            tp = tb_bilinear_tap<NT>(in_w, rs_w, rx);
        } else {
            tp.xmin = p.bx[2 * rx]; tp.n = p.bx[2 * rx + 1];
#pragma unroll
            for (int j = 0; j < NT; ++j) tp.k[j] = (j < p.ksx && j < tp.n) ? __ldg(p.kx + (size_t)rx * p.ksx + j) : 0;
        }
        tb_pack_taps<G>(tp.k, kbx);
        const int boff = (box_left + tp.xmin) * C;
        a1 = boff & 3;
        nw1 = (a1 + tp.n * C + 3) >> 2;                  // words that hold this window (1 .. NQ + 1)
        wbase = reinterpret_cast<const uint32_t*>(simg + (boff - a1)) + (int64_t)box_top * pitch_w;
    }
    const int sh1 = a1 * 8;
    const uint32_t pitch_b = (uint32_t)pitch;
    const int64_t plane = (int64_t)p.out_h * p.out_w;
    T* d[C];
#pragma unroll
    for (int c = 0; c < C; ++c) d[c] = (T*)p.dst + ((int64_t)b * C + c) * plane + (int64_t)y0 * p.out_w + x;
    __syncthreads();

    uint32_t win[C][G];
#pragma unroll
    for (int c = 0; c < C; ++c)
#pragma unroll
        for (int g = 0; g < G; ++g) win[c][g] = 0;

    int yy = 0;
    auto emit = [&](bool valid) {       // output row yy from the window (zeros when the row lies outside the resized image)
        uint32_t e[C];
#pragma unroll
        for (int c = 0; c < C; ++c) e[c] = 0;
        if (valid) {
            uint32_t kb[G][3];
#pragma unroll
            for (int g = 0; g < G; ++g) { kb[g][0] = s_kyb[yy][g][0]; kb[g][1] = s_kyb[yy][g][1]; kb[g][2] = s_kyb[yy][g][2]; }
#pragma unroll
            for (int c = 0; c < C; ++c) e[c] = tb_dot<G>(win[c], kb);
        }
#pragma unroll
        for (int c = 0; c < C; ++c) {
            if (active) *d[c] = s_lut[c * 256 + e[c]];
            d[c] += p.out_w;
        }
        ++yy;
    };

    while (yy < nr && s_ytop[yy] < 0) emit(false);                // rows above the resized image (centre-crop padding)
    if (yy < nr) {
        int nv = nr;
        while (s_ytop[nv - 1] < 0) --nv;                         // rows below it
        const int first = s_ytop[yy] - (NT - 1), last = s_ytop[nv - 1];
        // Rows past `last` (the batch's tail) or past the source (their taps are zero) re-read row
        // `rlim`: every load stays inside the image and what it returns is never weighted.
        const int rlim = min(last, in_h - 1);
        int next_top = s_ytop[yy];
        // Two register buffers of UR rows each: the loads of the next rows are in flight while the
        // current ones are consumed, so a warp does not sit out a DRAM round trip per batch.
        uint32_t wa[UR][NQ + 1], wb[UR][NQ + 1];
#pragma unroll
        for (int u = 0; u < UR; ++u)
#pragma unroll
            for (int i = 0; i <= NQ; ++i) wa[u][i] = wb[u][i] = 0;
        auto load = [&](uint32_t (&w)[UR][NQ + 1], int r0) {
#pragma unroll
            for (int u = 0; u < UR; ++u) {
                // one 32 x 32 -> 64-bit multiply-add per row (the host keeps the pitch below 4 GiB)
                const uint32_t* wp = reinterpret_cast<const uint32_t*>(
                    reinterpret_cast<const char*>(wbase) + (uint64_t)(uint32_t)min(r0 + u, rlim) * pitch_b);
                w[u][0] = __ldg(wp);
#pragma unroll
                for (int i = 1; i <= NQ; ++i) tb_ldg_if(w[u][i], wp + i, i < nw1);     // words past the window keep stale bits: zero taps
            }
        };
        auto consume = [&](const uint32_t (&w)[UR][NQ + 1], int r0) {
#pragma unroll
            for (int u = 0; u < UR; ++u) {
                const int r = r0 + u;
                if (r <= last) {
                    uint32_t q[NQ];
#pragma unroll
                    for (int i = 0; i < NQ; ++i) q[i] = __funnelshift_r(w[u][i], w[u][i + 1], sh1);
                    uint32_t h[C];
                    if (C == 3) {
                        // samples of channel c sit at bytes c, c+3, c+6, c+9 of each 12-byte group of four pixels
                        uint32_t p0[G], p1[G], p2[G];
#pragma unroll
                        for (int g = 0; g < G; ++g) {
                            p0[g] = __byte_perm(__byte_perm(q[3 * g], q[3 * g + 1], 0x0630), q[3 * g + 2], 0x5210);
                            p1[g] = __byte_perm(__byte_perm(q[3 * g], q[3 * g + 1], 0x0741), q[3 * g + 2], 0x6210);
                            p2[g] = __byte_perm(__byte_perm(q[3 * g], q[3 * g + 1], 0x0052), q[3 * g + 2], 0x7410);
                        }
                        h[0] = tb_dot<G>(p0, kbx);
                        h[1 % C] = tb_dot<G>(p1, kbx);
                        h[2 % C] = tb_dot<G>(p2, kbx);
                    } else {
                        h[0] = tb_dot<G>(q, kbx);
                    }
                    // push: the window keeps rows r - NT + 1 .. r, oldest in byte 0
#pragma unroll
                    for (int c = 0; c < C; ++c) {
#pragma unroll
                        for (int g = 0; g + 1 < G; ++g) win[c][g] = __funnelshift_r(win[c][g], win[c][g + 1], 8);
                        win[c][G - 1] = (win[c][G - 1] >> 8) | (h[c] << 24);
                    }
                    while (next_top == r) {
                        emit(true);
                        next_top = yy < nv ? s_ytop[yy] : 0x7fffffff;
                    }
                }
            }
        };
        load(wa, first);
        for (int r0 = first; r0 <= last; r0 += 2 * UR) {
            load(wb, r0 + UR);
            consume(wa, r0);
            load(wa, r0 + 2 * UR);
            consume(wb, r0 + UR);
        }
        while (yy < nr) emit(false);
    }
}


template <typename T, int C>
int launch_slide(const RsParams& p, int64_t blocks, int threads, int nt, cudaStream_t s) {
    if (p.box) {
        if (nt == 4) tb_resize_slide_kernel<T, C, true, 4><<<(unsigned)blocks, threads, 0, s>>>(p);
        else tb_resize_slide_kernel<T, C, true, 8><<<(unsigned)blocks, threads, 0, s>>>(p);
    } else {
        if (nt == 4) tb_resize_slide_kernel<T, C, false, 4><<<(unsigned)blocks, threads, 0, s>>>(p);
        else tb_resize_slide_kernel<T, C, false, 8><<<(unsigned)blocks, threads, 0, s>>>(p);
    }
    TB_LAUNCH_CHECK("tb_resize_slide_kernel");
    return TB_OK;
}

// Strip geometry of the sliding-window kernel: CTA = TC output columns (one thread each) x TR output rows.
// Taller strips recompute fewer halo rows (2 * TR + 2 input rows per TR outputs at scale 2), so TR is
// the tallest that still puts a full GPU's worth of threads in the grid; small grids are then cut
// into narrower column tiles until every SM has a few CTAs (measured on C4: batch 256 -> 32 x 256,
// batch 32 -> 16 x 128; 512 resized crops of 224 x 224 -> 64 x 224).
inline void slide_geometry(RsParams& p, int B, int sm_count, int flags, int64_t* blocks, int* threads) {
    const int64_t want_threads = (int64_t)sm_count * 1024 * 3 / 4;
    int TR = 32;
    while (TR > 8 && (int64_t)B * ((p.out_h + TR - 1) / TR) * p.out_w < want_threads) TR >>= 1;
    // very large grids afford 64-row strips (halo and per-strip set-up halve again) while >= 12 CTAs per SM remain
    if (TR == 32 && (int64_t)B * ((p.out_h + 63) / 64) * ((p.out_w + 255) / 256) >= 12 * (int64_t)sm_count) TR = 64;
    int TC = p.out_w < 256 ? p.out_w : 256;
    while (TC > 64 && (int64_t)B * ((p.out_h + TR - 1) / TR) * ((p.out_w + TC - 1) / TC) < 4 * (int64_t)sm_count) TC = ((TC / 2 + 31) / 32) * 32;
    if ((flags >> 24) & 0xff) TR = (flags >> 24) & 0xff;             // development knobs
    if ((flags >> 16) & 0xff) TC = 32 * ((flags >> 16) & 0xff);
    if (TR > kSlideMaxTR) TR = kSlideMaxTR;
    if (TR > p.out_h) TR = p.out_h;
    if (TC > 256) TC = 256;
    if (TC > p.out_w) TC = p.out_w;
    p.TR = TR; p.TC = TC; p.tiles_x = (p.out_w + TC - 1) / TC; p.tiles_y = (p.out_h + TR - 1) / TR;
    *blocks = (int64_t)B * p.tiles_x * p.tiles_y;
    *threads = ((TC + 31) / 32) * 32;
}


// ----------------------------------------------------------------------------------------------
// ring variant of the sliding-window kernel (C == 3, float / 16-bit outputs, 16-byte aligned rows).
//
// Same arithmetic and the same per-thread register window as tb_resize_slide_kernel; what changes is
// how the bytes arrive and how the work is cut:
//   * the input rows of a strip are fetched ONCE per column tile by 1-D bulk async copies
//     (cp.async.bulk, TMA engine, SASS UBLKCP) into a shared-memory ring of row slots, completion on
//     mbarriers; the column threads read their 16-byte windows out of the ring (LDS) instead of every
//     thread pulling its own overlapping window through L1.  A dedicated producer warp runs ahead of the
//     consumers by the depth of the ring (full / empty mbarrier pairs per slot), so no thread ever sits
//     out a DRAM round trip;
//   * one persistent CTA per SM holds the lookup table LANE-REPLICATED (word = value, row of 32 lanes
//     per byte value and channel, 96 KiB): lane l reads word l of its row, so the lookups never collide
//     on a bank whatever the 32 byte values are (the plain [c][256] table cost 35 % extra wavefronts);
//   * the flattened (column tile, image, output row) space is split EVENLY over the groups of the grid
//     (a group = tc column threads), in chunks of at most kRingMaxRows rows that never straddle an image:
//     no wave quantisation (2048 strips over 592 resident CTAs ran 4 waves for 3.46 waves of work);
//   * tap byte planes that are zero for a whole warp (horizontal) or output row (vertical) are skipped:
//     an exact 2x / 4x reduction has taps k * 2^-3 / 2^-4, i.e. only the top byte plane, and its dot
//     product is one DP4A instead of three.  Same integers either way.
// ----------------------------------------------------------------------------------------------
constexpr int kRingMaxRows = 256;     // output rows per chunk (their vertical taps sit in shared memory); at most tc
constexpr int kRingRG = 4;            // input rows per ring slot (one full / empty barrier pair per slot)
constexpr int kRingMaxGroups = 3;
constexpr uint32_t kRingTabBytes = 3u * 256u * 32u * 4u;

struct RingGeom {
    int ng;                 // groups per CTA
    int tc;                 // column threads per group (multiple of 32)
    int slots;              // ring slots per group
    int seg_pitch;          // bytes per staged row (multiple of 16, >= longest segment + 16)
    int tiles_x;
    int max_rows;           // output rows per chunk
    int total_rows;         // tiles_x * B * out_h (< 2^31: the kernel's row arithmetic is 32-bit, 64-bit divisions cost ~0.3 us each)
    int gpu;                // > 0: groups per (image, column tile) unit -- no group straddles two units; 0: even split of all rows
    int rows_q, rows_r;     // even split: total_rows = rows_q * groups + rows_r
    uint32_t ring_off, bar_off, kyb_off, ytop_off, dyn_bytes;
};

__device__ __forceinline__ void tb_prefetch_l1(const void* p) { asm volatile("prefetch.global.L1 [%0];" ::"l"(p)); }
__device__ __forceinline__ void tb_named_barrier(int id, int nthreads) {
    asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}
// Shared-memory loads by 32-bit window address (no generic-address arithmetic in the loop).  volatile keeps them
// ordered against the mbarrier wait / arrive that guard the ring; ALU code moves freely around them.
template <int OFF>
__device__ __forceinline__ uint32_t tb_lds_at(uint32_t addr) {
    uint32_t v;
    asm volatile("ld.shared.u32 %0, [%1+%2];" : "=r"(v) : "r"(addr), "n"(OFF));
    return v;
}
__device__ __forceinline__ uint4 tb_lds128(uint32_t addr) {
    uint4 v;
    asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr));
    return v;
}
// base + idx as ONE 32 x 32 + 64-bit multiply-add (IMAD.WIDE, the FMA pipe): left to itself the compiler builds the
// 64-bit address from three or four integer-ALU instructions per store, and that pipe is what bounds the ring kernel
template <typename T>
__device__ __forceinline__ T* tb_elem_ptr(T* base, uint32_t idx) {
    uint64_t a;
    asm("mad.wide.u32 %0, %1, %2, %3;" : "=l"(a) : "r"(idx), "r"((uint32_t)sizeof(T)), "l"(base));
    return reinterpret_cast<T*>(a);
}
template <typename T>
__device__ __forceinline__ void tb_st_cs(T* p, uint32_t bits);
template <>
__device__ __forceinline__ void tb_st_cs<float>(float* p, uint32_t bits) {
    asm volatile("st.global.cs.b32 [%0], %1;" ::"l"(p), "r"(bits));       // no "memory": dst is never read back
}
template <>
__device__ __forceinline__ void tb_st_cs<uint16_t>(uint16_t* p, uint32_t bits) {
    asm volatile("st.global.cs.b16 [%0], %1;" ::"l"(p), "h"((uint16_t)bits));
}

// Dot product of G groups of four samples with their taps, in one of three tap encodings:
//   MODE 0  taps scaled by 4 (24 bits, none equal to 2^22) and only their top byte plane non-zero (an exact 2x / 4x
//           reduction): one DP4A per group; the rounded result is BYTE 1 of the accumulator;
//   MODE 1  taps scaled by 4, three byte planes: the result is BYTE 3 of the accumulator (no final shift);
//   MODE 2  plain 22-bit taps (a tap of exactly 1.0 = 2^22 does not fit 24 bits once scaled): the result is the value.
// Scaled: sum(p * 4k) + 2^23 = 4 * (sum(p * k) + 2^21), so bits 24-31 are Pillow's (2^21 + sum) >> 22; the sum stays
// below 2^32 because the taps sum to at most 2^22 + NT / 2.
template <int G, int MODE>
__device__ __forceinline__ uint32_t tb_dot_mode(const uint32_t* px, const uint32_t (*kb)[3]) {
    if (MODE == 2) return tb_dot<G>(px, kb);
    uint32_t a = 128u;
#pragma unroll
    for (int g = 0; g < G; ++g) a = __dp4a(px[g], kb[g][2], a);
    if (MODE == 0) return a;
    a <<= 8;
#pragma unroll
    for (int g = 0; g < G; ++g) a = __dp4a(px[g], kb[g][1], a);
    a <<= 8;
#pragma unroll
    for (int g = 0; g < G; ++g) a = __dp4a(px[g], kb[g][0], a);
    return a;
}
// newest sample into the top byte of a window word, the oldest drops out of byte 0
template <int MODE>
__device__ __forceinline__ uint32_t tb_push_mode(uint32_t win, uint32_t a) {
    if (MODE == 0) return __byte_perm(win, a, 0x5321);
    if (MODE == 1) return __byte_perm(win, a, 0x7321);
    return (win >> 8) | (a << 24);
}
// 128 * result (the byte offset of a table row) from an accumulator
template <int MODE>
__device__ __forceinline__ uint32_t tb_row7_mode(uint32_t a) {
    if (MODE == 0) return (a >> 1) & 0x7f80u;
    if (MODE == 1) return (a >> 17) & 0x7f80u;
    return a << 7;
}
// taps -> byte planes; scaled by 4 unless one of them is 2^22.  Returns 1 when the plain encoding was used.
template <int G, int NT>
__device__ __forceinline__ uint32_t tb_pack_taps_scaled(const int* k, uint32_t (*kb)[3]) {
    bool unity = false;
#pragma unroll
    for (int j = 0; j < NT; ++j) unity = unity || k[j] >= (1 << kPrecisionBits);
    int k4[NT];
#pragma unroll
    for (int j = 0; j < NT; ++j) k4[j] = unity ? k[j] : k[j] << 2;
    tb_pack_taps<G>(k4, kb);
    return unity ? 1u : 0u;
}

// One chunk of a group's share of the flattened row space: output rows [y0, y0 + nr) of column tile tx of
// image b, the input rows [first, last] they depend on and the byte segment [s0, s0 + seg_len) of every
// such row that the tile's columns read.  Producer and consumers derive it independently from (F, F1).
struct RingChunk {
    int b, tx, y0, nr;
    int oy, ox, flip;
    int first, last;        // last < first: nothing to fetch (all rows / columns outside the resized image)
    int s0, seg_len;
};
// Output rows [F, F1) of group `me` of `ngt`.  Even split of all rows, or -- when the host found that every unit (one
// image x one column tile, out_h rows) can be given g.gpu whole groups at little loss -- an even split of each unit
// among its own groups: a group that straddles two units pays a second chunk set-up and window warm-up.
__device__ __forceinline__ void tb_ring_range(const RingGeom& g, int out_h, int me, int& F, int& F1) {
    if (g.gpu > 0) {
        const int unit = (int)((unsigned)me / (unsigned)g.gpu);
        const int j = me - unit * g.gpu;
        F = F1 = 0;
        if ((long long)unit * out_h >= g.total_rows) return;
        F = unit * out_h + (int)((unsigned)(j * out_h) / (unsigned)g.gpu);
        F1 = unit * out_h + (int)((unsigned)((j + 1) * out_h) / (unsigned)g.gpu);
        return;
    }
    F = me * g.rows_q + min(me, g.rows_r);
    F1 = F + g.rows_q + (me < g.rows_r ? 1 : 0);
}

template <int NT>
__device__ __forceinline__ bool tb_ring_next(const RsParams& p, const RingGeom& g, int& F, int F1, RingChunk& c) {
    if (F >= F1) return false;
    const int unit = (int)((unsigned)F / (unsigned)p.out_h);
    c.y0 = F - unit * p.out_h;
    c.b = (int)((unsigned)unit / (unsigned)g.tiles_x);
    c.tx = unit - c.b * g.tiles_x;
    c.nr = min(min(p.out_h - c.y0, g.max_rows), F1 - F);
    F += c.nr;
    c.oy = (p.top ? p.top[c.b] : 0) + p.off_y;
    c.ox = (p.left ? p.left[c.b] : 0) + p.off_x;
    c.flip = p.flip ? (p.flip[c.b] != 0) : 0;
    c.first = 0; c.last = -1; c.s0 = 0; c.seg_len = 0;
    const int ia = max(0, -(c.oy + c.y0)), ib = min(c.nr, p.rs_h - (c.oy + c.y0)) - 1;      // valid output rows of the chunk
    const int x0 = c.tx * g.tc, nc = min(g.tc, p.out_w - x0);
    int ra = c.ox + (c.flip ? p.out_w - x0 - nc : x0), rb = ra + nc - 1;                   // resized columns of the tile
    ra = max(ra, 0); rb = min(rb, p.rs_w - 1);
    if (ia <= ib && ra <= rb) {
        c.first = p.by[2 * (c.oy + c.y0 + ia)];
        c.last = p.by[2 * (c.oy + c.y0 + ib)] + NT - 1;
        c.s0 = (p.bx[2 * ra] * 3) & ~15;
        const int end = (p.bx[2 * rb] + p.bx[2 * rb + 1]) * 3;
        c.seg_len = ((end + 15) & ~15) - c.s0;
        if (c.seg_len + 16 > g.seg_pitch) asm volatile("trap;");      // the host's bound on a tile's segment is wrong
    }
    return true;
}

template <typename T, int NT>
__global__ void __launch_bounds__(kRingMaxGroups * (256 + 32), 1) tb_resize_ring_kernel(RsParams p, RingGeom g) {
    constexpr int C = 3;
    constexpr int G = NT / 4;
    constexpr int NQ = NT * C / 4;
    constexpr int RG = kRingRG;
    extern __shared__ __align__(128) uint8_t smem[];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int wpg = g.tc >> 5;
    const int n_cons = g.ng * wpg;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + g.bar_off);        // [ng][slots][2]: full, empty

    if (tid == 0) {
        for (int i = 0; i < g.ng * g.slots; ++i) {
            tb_mbar_init(&bars[2 * i], 1);                      // full: the producer's arrive.expect_tx
            tb_mbar_init(&bars[2 * i + 1], (uint32_t)wpg);      // empty: one arrival per consumer warp of the group
        }
        tb_mbar_fence_init();
    }
    __syncthreads();                // the barriers exist before anyone arrives on or waits for them
    tb_launch_dependents();
    // the group's share of the rows: arithmetic only, done while the previous launch drains
    int F, F1;
    tb_ring_range(g, p.out_h, (int)blockIdx.x * g.ng + (warp >= n_cons ? warp - n_cons : warp / wpg), F, F1);
    tb_grid_dependency_wait();
    if (warp < n_cons) {
        // the head of the four tap tables into L1 (one 128-byte line per thread and table): the chunk set-up below then
        // reads them at L1 latency instead of paying two more dependent trips to L2 / DRAM after the table fill
        {
            const size_t line = (size_t)tid * 128;
            if (line < (size_t)p.rs_h * 8) tb_prefetch_l1((const char*)p.by + line);
            if (line < (size_t)p.rs_h * p.ksy * 4) tb_prefetch_l1((const char*)p.ky + line);
            if (line < (size_t)p.rs_w * 8) tb_prefetch_l1((const char*)p.bx + line);
            if (line < (size_t)p.rs_w * p.ksx * 4) tb_prefetch_l1((const char*)p.kx + line);
        }
        // lane-replicated table: [c][e][lane] words, filled by the consumer warps while the producer warps already work
        // out their first chunk and start its copies (they never touch the table).  A dozen values of the thread's share
        // at a time: one round trip to L2 for all of them at the usual 512 / 768 consumer threads.
        const int nthr = n_cons * 32;
        uint4* t4 = reinterpret_cast<uint4*>(smem);
        constexpr int kPer = 12;
        for (int i0 = tid; i0 < 3 * 256 * 8; i0 += kPer * nthr) {
            uint32_t tv[kPer];
#pragma unroll
            for (int k = 0; k < kPer; ++k) {
                const int i = i0 + k * nthr;
                tv[k] = 0;
                if (i < 3 * 256 * 8) {
                    if (sizeof(T) == 4) tv[k] = __float_as_uint(__ldg((const float*)p.lut + (i >> 3)));
                    else tv[k] = __ldg((const uint16_t*)p.lut + (i >> 3));
                }
            }
#pragma unroll
            for (int k = 0; k < kPer; ++k) {
                const int i = i0 + k * nthr;
                if (i < 3 * 256 * 8) t4[i] = make_uint4(tv[k], tv[k], tv[k], tv[k]);
            }
        }
        tb_named_barrier(15, nthr);
    }

    if (warp >= n_cons) {
        // ------------------------------------------------------------------ producer warps: one per group
        // (one warp polling every ring topped out near 1.5 TB/s of reads: tools/micro/bulk_copy_rate.cu shows a
        // single warp issuing four 1.5 KB copies per step saturates at ~4 TB/s even with nothing else to do)
        const int gi = warp - n_cons;
        int slot = 0;
        uint32_t phase = 1;                       // a fresh barrier passes a wait on parity 1
        RingChunk ch;
        while (tb_ring_next<NT>(p, g, F, F1, ch)) {
            const uint8_t* gsrc0 = p.src + (int64_t)ch.b * p.bs + ch.s0 + (int64_t)lane * p.pitch;
            for (int r = ch.first; r <= ch.last; r += RG) {
                uint64_t* full = &bars[2 * (gi * g.slots + slot)];
                tb_mbar_wait(full + 1, phase);                            // every consumer warp has drained the slot
                const int nrows = min(RG, ch.last - r + 1);
                const int ncopy = max(0, min(nrows, p.H - r));            // rows past the source are never weighted: not fetched
                if (lane == 0) tb_mbar_arrive_expect_tx(full, (uint32_t)(ncopy * ch.seg_len));
                __syncwarp();
                if (lane < ncopy) {
                    uint8_t* sdst = smem + g.ring_off + (size_t)((gi * g.slots + slot) * RG + lane) * g.seg_pitch;
                    tb_bulk_g2s(sdst, gsrc0 + (int64_t)r * p.pitch, (uint32_t)ch.seg_len, full);
                }
                if (++slot == g.slots) { slot = 0; phase ^= 1u; }
            }
        }
        return;
    }

    // ---------------------------------------------------------------------- consumers
    const int grp = warp / wpg;
    const int gt = tid - grp * g.tc;                 // column thread within the group
    const uint32_t s_base = tb_smem_addr(smem);
    uint32_t (*s_kyb)[4] = reinterpret_cast<uint32_t (*)[4]>(smem + g.kyb_off) + grp * kRingMaxRows * G;     // [row][G] x {k0, k1, k2, -}
    int* s_ytop = reinterpret_cast<int*>(smem + g.ytop_off) + grp * kRingMaxRows;
    const uint32_t kyb_addr = s_base + g.kyb_off + (uint32_t)(grp * kRingMaxRows * G * 16);
    const uint32_t ytop_addr = s_base + g.ytop_off + (uint32_t)(grp * kRingMaxRows * 4);
    const uint32_t tab_addr = s_base + (uint32_t)lane * 4u;
    const uint32_t ring_addr = s_base + g.ring_off + (uint32_t)(grp * g.slots * RG) * (uint32_t)g.seg_pitch;
    const uint32_t slot_bytes = (uint32_t)(RG * g.seg_pitch);
    const int64_t plane = (int64_t)p.out_h * p.out_w;
    int slot = 0;
    uint32_t phase = 0;
    RingChunk c;
    while (tb_ring_next<NT>(p, g, F, F1, c)) {
        // Every global load of the set-up first, none waiting for another: the chunk's bounds (in tb_ring_next), this
        // thread's row taps and its column taps -- one trip to L2 / DRAM instead of three in a row (a trace showed 0.9 us
        // for the bounds and 0.9 us for the column taps).  Taps are read whole and masked by their count afterwards.
        const int ry = c.oy + c.y0 + gt;
        const bool rowok = gt < c.nr && ry >= 0 && ry < p.rs_h;
        int kyr[NT], ymin = 0, yn = 0;
#pragma unroll
        for (int j = 0; j < NT; ++j) kyr[j] = 0;
        if (rowok) {
            ymin = p.by[2 * ry]; yn = p.by[2 * ry + 1];
#pragma unroll
            for (int j = 0; j < NT; ++j)
                if (j < p.ksy) kyr[j] = p.ky[(size_t)ry * p.ksy + j];
        }
        const int x0 = c.tx * g.tc;
        const int nc = min(g.tc, p.out_w - x0);
        const bool active = gt < nc;
        const int x = x0 + gt;
        const int rx = c.ox + (c.flip ? p.out_w - 1 - x : x);
        const bool colok = active && rx >= 0 && rx < p.rs_w;
        int kx[NT], xmin_c = 0, xn = 0;
#pragma unroll
        for (int j = 0; j < NT; ++j) kx[j] = 0;
        if (colok) {
            xmin_c = p.bx[2 * rx]; xn = p.bx[2 * rx + 1];
#pragma unroll
            for (int j = 0; j < NT; ++j)
                if (j < p.ksx) kx[j] = __ldg(p.kx + (size_t)rx * p.ksx + j);
        }
        tb_named_barrier(1 + grp, g.tc);             // every thread of the group is done with the previous chunk's tables
        if (gt < c.nr) {
            int top = -1;
            if (rowok && c.last >= c.first) {
                int k[NT];
                const int xmin = ymin, n = yn;
#pragma unroll
                for (int j = 0; j < NT; ++j) k[j] = j < n ? kyr[j] : 0;
                // vertical taps: mode word 1 = plain 22-bit taps (one of them is 1.0); 2 = taps scaled by TWO with only their
                // top byte plane non-zero (k a multiple of 2^15: exact 2x / 4x) -- the rounded row lands in bits 7..14 of
                // the accumulator, so the byte offset of its table row is one AND; 0 = taps scaled by four, three planes
                uint32_t kb[G][3];
                uint32_t vmode = tb_pack_taps_scaled<G, NT>(k, kb);
                bool top_only = vmode == 0;
#pragma unroll
                for (int j = 0; j < NT; ++j) top_only = top_only && (k[j] & 0x7fff) == 0;
                if (top_only) {
                    int k2[NT];
#pragma unroll
                    for (int j = 0; j < NT; ++j) k2[j] = k[j] << 1;
                    tb_pack_taps<G>(k2, kb);
                    vmode = 2;
                }
#pragma unroll
                for (int q = 0; q < G; ++q) { s_kyb[gt * G + q][0] = kb[q][0]; s_kyb[gt * G + q][1] = kb[q][1]; s_kyb[gt * G + q][2] = kb[q][2]; s_kyb[gt * G + q][3] = vmode; }
                top = xmin + NT - 1;
            }
            s_ytop[gt] = top;
        }
        const bool colin = colok && c.last >= c.first;
        int woff = 0, sh1 = 0;
#pragma unroll
        for (int j = 0; j < NT; ++j) kx[j] = (colin && j < xn) ? kx[j] : 0;
        if (colin) {
            const int off = xmin_c * 3 - c.s0;
            sh1 = (off & 3) * 8;
            woff = off & ~3;
        }
        // one tap encoding per warp: plain if any of its columns has a tap of exactly 1.0, else scaled (top plane
        // only when that holds for the whole warp)
        bool unity = false;
#pragma unroll
        for (int j = 0; j < NT; ++j) unity = unity || kx[j] >= (1 << kPrecisionBits);
        unity = __any_sync(0xffffffffu, unity);
        uint32_t kbx[G][3];
        {
            int k4[NT];
#pragma unroll
            for (int j = 0; j < NT; ++j) k4[j] = unity ? kx[j] : kx[j] << 2;
            tb_pack_taps<G>(k4, kbx);
        }
        bool lowzero = true;
#pragma unroll
        for (int q = 0; q < G; ++q) lowzero = lowzero && (kbx[q][0] | kbx[q][1]) == 0;
        lowzero = __all_sync(0xffffffffu, lowzero);
        const int hmode = unity ? 2 : (lowzero ? 0 : 1);
        T* dbase[C];
#pragma unroll
        for (int ch = 0; ch < C; ++ch) dbase[ch] = (T*)p.dst + ((int64_t)c.b * C + ch) * plane + (int64_t)c.y0 * p.out_w + x;
        uint32_t eo = 0;                             // element offset of output row yy from dbase
        tb_named_barrier(1 + grp, g.tc);

        uint32_t win[C][G];
#pragma unroll
        for (int ch = 0; ch < C; ++ch)
#pragma unroll
            for (int q = 0; q < G; ++q) win[ch][q] = 0;
        int yy = 0;
        const int nr = c.nr;
        auto ytop = [&](int i) { return (int)tb_lds_at<0>(ytop_addr + 4u * (uint32_t)i); };
        auto store_row = [&](const uint32_t (&row7)[C]) {
            uint32_t v[C];
            v[0] = tb_lds_at<0>(tab_addr + row7[0]);
            v[1] = tb_lds_at<32768>(tab_addr + row7[1]);
            v[2] = tb_lds_at<65536>(tab_addr + row7[2]);
            if (active) {
#pragma unroll
                for (int ch = 0; ch < C; ++ch) tb_st_cs<T>(tb_elem_ptr<T>(dbase[ch], eo), v[ch]);
            }
            eo += (uint32_t)p.out_w;
            ++yy;
        };
        auto emit_zero = [&]() {
            const uint32_t row7[C] = {0u, 0u, 0u};
            store_row(row7);
        };
        auto emit = [&]() {          // output row yy from the register window
            uint32_t kb[G][3];
            uint32_t vmode = 0;
#pragma unroll
            for (int q = 0; q < G; ++q) {
                const uint4 v = tb_lds128(kyb_addr + (uint32_t)((yy * G + q) * 16));
                kb[q][0] = v.x; kb[q][1] = v.y; kb[q][2] = v.z;
                vmode = v.w;
            }
            uint32_t row7[C];
            if (vmode == 2) {
#pragma unroll
                for (int ch = 0; ch < C; ++ch) {
                    uint32_t a = 64u;                       // 2^21 at the scale of these taps (2 k >> 16)
#pragma unroll
                    for (int q = 0; q < G; ++q) a = __dp4a(win[ch][q], kb[q][2], a);
                    row7[ch] = a & 0x7f80u;                 // ((a >> 7) & 255) << 7
                }
            } else if (vmode == 1) {
#pragma unroll
                for (int ch = 0; ch < C; ++ch) row7[ch] = tb_row7_mode<2>(tb_dot_mode<G, 2>(win[ch], kb));
            } else {
#pragma unroll
                for (int ch = 0; ch < C; ++ch) row7[ch] = tb_row7_mode<1>(tb_dot_mode<G, 1>(win[ch], kb));
            }
            store_row(row7);
        };
        while (yy < nr && ytop(yy) < 0) emit_zero();               // rows above the resized image (centre-crop padding)
        if (yy < nr) {
            int nv = nr;
            while (ytop(nv - 1) < 0) --nv;
            const int last = c.last;
            int next_top = ytop(yy);
            auto rows = [&](auto mode_c) {
                constexpr int MODE = decltype(mode_c)::value;
                // one ring slot: RG input rows.  FULL: all of them are rows of this chunk (no per-row test).
                auto slot_rows = [&](int r0, auto full_c) {
                    constexpr bool FULL = decltype(full_c)::value;
                    uint64_t* full = &bars[2 * (grp * g.slots + slot)];
                    tb_mbar_wait(full, phase);
                    // every window word of the slot's RG rows first (rows past `last` hold stale bytes that are
                    // computed on but never pushed), then RG * 3 independent dot products: the loads' latency and
                    // the DP4A chains overlap instead of running row after row
                    const uint32_t rowa = ring_addr + (uint32_t)slot * slot_bytes + (uint32_t)woff;
                    uint32_t w[RG][NQ + 1];
#pragma unroll
                    for (int u = 0; u < RG; ++u) {
                        const uint32_t ra = rowa + (uint32_t)(u * g.seg_pitch);
                        w[u][0] = tb_lds_at<0>(ra);
                        w[u][1] = tb_lds_at<4>(ra);
                        w[u][2] = tb_lds_at<8>(ra);
                        w[u][3] = tb_lds_at<12>(ra);
                        if (NQ >= 6) { w[u][4] = tb_lds_at<16>(ra); w[u][5] = tb_lds_at<20>(ra); w[u][NQ] = tb_lds_at<24>(ra); }
                    }
                    uint32_t h[RG][C];
#pragma unroll
                    for (int u = 0; u < RG; ++u) {
                        uint32_t q[NQ];
#pragma unroll
                        for (int i = 0; i < NQ; ++i) q[i] = __funnelshift_r(w[u][i], w[u][i + 1], sh1);
                        uint32_t p0[G], p1[G], p2[G];
#pragma unroll
                        for (int t = 0; t < G; ++t) {
                            p0[t] = __byte_perm(__byte_perm(q[3 * t], q[3 * t + 1], 0x0630), q[3 * t + 2], 0x5210);
                            p1[t] = __byte_perm(__byte_perm(q[3 * t], q[3 * t + 1], 0x0741), q[3 * t + 2], 0x6210);
                            p2[t] = __byte_perm(__byte_perm(q[3 * t], q[3 * t + 1], 0x0052), q[3 * t + 2], 0x7410);
                        }
                        h[u][0] = tb_dot_mode<G, MODE>(p0, kbx);
                        h[u][1] = tb_dot_mode<G, MODE>(p1, kbx);
                        h[u][2] = tb_dot_mode<G, MODE>(p2, kbx);
                    }
#pragma unroll
                    for (int u = 0; u < RG; ++u) {
                        const int r = r0 + u;
                        if (FULL || r <= last) {
#pragma unroll
                            for (int ch = 0; ch < C; ++ch) {
#pragma unroll
                                for (int t = 0; t + 1 < G; ++t) win[ch][t] = __funnelshift_r(win[ch][t], win[ch][t + 1], 8);
                                win[ch][G - 1] = tb_push_mode<MODE>(win[ch][G - 1], h[u][ch]);
                            }
                            while (next_top == r) {
                                emit();
                                next_top = yy < nv ? ytop(yy) : 0x7fffffff;
                            }
                        }
                    }
                    // The slot is drained by this warp (every loaded word has been consumed above): hand it back.
                    // No proxy fence: the reads were performed before the arrive, the refill is ordered after it.
                    __syncwarp();
                    if (lane == 0) tb_mbar_arrive(full + 1);
                    if (++slot == g.slots) { slot = 0; phase ^= 1u; }
                };
                int r0 = c.first;
                for (; r0 + RG - 1 <= last; r0 += RG) slot_rows(r0, std::true_type{});
                if (r0 <= last) slot_rows(r0, std::false_type{});
            };
            if (hmode == 0) rows(std::integral_constant<int, 0>{});
            else if (hmode == 1) rows(std::integral_constant<int, 1>{});
            else rows(std::integral_constant<int, 2>{});
            while (yy < nr) emit_zero();
        }
    }
}

// Geometry of the ring kernel, or false when this launch does not qualify (the sliding-window kernel runs then).
inline bool ring_geometry(const RsParams& p, int B, int nt, const TbDevice& dev, int flags, RingGeom* out, int* grid, int* threads) {
    RingGeom g = {};
    g.tc = p.out_w >= 256 ? 256 : ((p.out_w + 31) / 32) * 32;
    g.tiles_x = (p.out_w + g.tc - 1) / g.tc;
    // longest byte segment a column tile reads from one row: its resized columns span at most
    // (tc - 1) * scale + NT + 2 source pixels; + the 16-byte alignment on both sides + one window of slack
    const double scale = (double)p.W / (double)p.rs_w;
    int64_t span = (int64_t)((g.tc - 1) * scale) + nt + 3;
    if (span > p.W) span = p.W;
    int64_t seg = ((span * 3 + 15 + 15) & ~(int64_t)15);
    if (seg > (int64_t)p.W * 3) seg = (int64_t)p.W * 3;
    g.seg_pitch = (int)(seg + 16);
    const int G = nt / 4;
    const int64_t avail = (int64_t)dev.max_smem_optin - 1024 - kRingTabBytes;
    int ng = (flags >> 8) & 0xff;
    auto fits = [&](int n, int slots) {
        return (int64_t)n * ((int64_t)slots * kRingRG * g.seg_pitch + slots * 16 + kRingMaxRows * G * 16 + kRingMaxRows * 4) + 256 <= avail;
    };
    const long long total64 = (long long)g.tiles_x * B * p.out_h;
    if (total64 >= ((long long)1 << 30)) return false;      // 32-bit row arithmetic in the kernel
    g.total_rows = (int)total64;
    // three groups (24 consumer warps) once every group of a full grid owns 16 rows or more (measured on C4: batch 256 ->
    // 90 us with three groups, 104 us with two; batch 64 25.8 / 28.1 us; batch 32, 18-20 rows per group, 15.3 / 15.7 us)
    if (ng == 0) ng = (fits(3, 4) && g.total_rows >= 3LL * dev.sm_count * 16) ? 3 : 2;
    if (ng > kRingMaxGroups) ng = kRingMaxGroups;
    int slots = 8;
    while (slots > 3 && !fits(ng, slots)) --slots;
    if (!fits(ng, slots)) {
        ng = 1; slots = 8;
        while (slots > 3 && !fits(ng, slots)) --slots;
        if (!fits(ng, slots)) return false;
    }
    g.ng = ng; g.slots = slots;
    g.max_rows = kRingMaxRows < p.out_h ? kRingMaxRows : p.out_h;
    if (g.max_rows > g.tc) g.max_rows = g.tc;          // one thread of the group sets up one row of a chunk
    // Small batches are latency, not throughput: rather more SMs with few rows each than few SMs with many, down to
    // 8 rows per group (4 while that still leaves most SMs empty), although a chunk's halo is NT input rows.  Measured on
    // C4 (512^2 -> 256^2), us per launch with 24 / 8 / 4 rows at least: batch 1 12.1 / 7.2 / 5.9, batch 2 - / 7.3 / 6.1,
    // batch 8 14.7 / 7.9 / 8.6.
    long long min_rows = (g.total_rows <= 2LL * ng * dev.sm_count) ? 4 : 8;
    if ((flags >> 17) & 0x7f) min_rows = (flags >> 17) & 0x7f;      // (development knob)
    long long ctas = (g.total_rows + min_rows * ng - 1) / (min_rows * ng);
    if (ctas > dev.sm_count) ctas = dev.sm_count;
    if (ctas < 1) ctas = 1;
    // whole groups per unit when that costs little: the longest group then has ceil(out_h / gpu) rows against the even
    // split's total / groups, but none pays for a second chunk (measured on C4 at batch 32: see DESIGN.md 4.3)
    g.gpu = 0;
    {
        const long long groups = ctas * ng, units = g.total_rows / p.out_h;
        const long long per = units > 0 ? groups / units : 0;
        if (per >= 2 && per <= 64 && !((flags >> 16) & 1)) {
            const long long longest = (p.out_h + per - 1) / per;
            if (longest >= 8 && (double)g.total_rows / ((double)groups * (double)longest) >= 0.92) g.gpu = (int)per;
        }
    }
    g.rows_q = (int)(g.total_rows / (ctas * ng));
    g.rows_r = (int)(g.total_rows % (ctas * ng));
    uint32_t off = kRingTabBytes;
    g.ring_off = off; off += (uint32_t)(ng * slots * kRingRG * g.seg_pitch);
    off = (off + 15u) & ~15u;
    g.bar_off = off; off += (uint32_t)(ng * slots * 16);
    g.kyb_off = off; off += (uint32_t)(ng * kRingMaxRows * G * 16);
    g.ytop_off = off; off += (uint32_t)(ng * kRingMaxRows * 4);
    g.dyn_bytes = off;
    *out = g;
    *grid = (int)ctas;
    *threads = ng * (g.tc + 32);
    return true;
}

template <typename T>
int launch_ring(const RsParams& p, const RingGeom& g, int grid, int threads, int nt, bool pdl, cudaStream_t s) {
    static int attr_set[64][2] = {};
    int dev = 0;
    TB_CUDA(cudaGetDevice(&dev));
    auto kern = nt == 4 ? tb_resize_ring_kernel<T, 4> : tb_resize_ring_kernel<T, 8>;
    if (dev >= 0 && dev < 64 && !__atomic_load_n(&attr_set[dev][nt == 8], __ATOMIC_ACQUIRE)) {
        int max_optin = 0;
        TB_CUDA(cudaDeviceGetAttribute(&max_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev));
        TB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, max_optin));
        __atomic_store_n(&attr_set[dev][nt == 8], 1, __ATOMIC_RELEASE);
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)grid);
    cfg.blockDim = dim3((unsigned)threads);
    cfg.dynamicSmemBytes = g.dyn_bytes;
    cfg.stream = s;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl ? 1 : 0;
    TB_CUDA(cudaLaunchKernelEx(&cfg, kern, p, g));
    TB_LAUNCH_CHECK("tb_resize_ring_kernel");
    return TB_OK;
}

template <typename T, int C>
int launch_resize(const RsParams& p, size_t smem, int64_t blocks, int max_smem, bool fast, int nt, cudaStream_t s) {
    static int attr_set[64] = {};            // per device, atomic: racing first calls set the same attribute
    int dev = 0;
    TB_CUDA(cudaGetDevice(&dev));
    if (smem > 40 * 1024 && dev >= 0 && dev < 64 && !__atomic_load_n(&attr_set[dev], __ATOMIC_ACQUIRE)) {
        TB_CUDA(cudaFuncSetAttribute(tb_resize_fused_kernel<T, C>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     max_smem - 8192));
        __atomic_store_n(&attr_set[dev], 1, __ATOMIC_RELEASE);
    }
    if (fast && smem <= 44 * 1024) {
        if (p.box) {
            if (nt == 4) tb_resize_fast_kernel<T, C, true, 4><<<(unsigned)blocks, 256, smem, s>>>(p);
            else tb_resize_fast_kernel<T, C, true, 8><<<(unsigned)blocks, 256, smem, s>>>(p);
        } else {
            if (nt == 4) tb_resize_fast_kernel<T, C, false, 4><<<(unsigned)blocks, 256, smem, s>>>(p);
            else tb_resize_fast_kernel<T, C, false, 8><<<(unsigned)blocks, 256, smem, s>>>(p);
        }
        TB_LAUNCH_CHECK("tb_resize_fast_kernel");
        return TB_OK;
    }
    if (p.box) return tb_fail(TB_ENOTSUP, "tb_resized_crop: tile does not fit the fast path");
    tb_resize_fused_kernel<T, C><<<(unsigned)blocks, 256, smem, s>>>(p);
    TB_LAUNCH_CHECK("tb_resize_fused_kernel");
    return TB_OK;
}

}  // namespace

extern "C" int tb_resize_bilinear_u8_nhwc_to_nchw(
    const uint8_t* src, int B, int H, int W, int C, int64_t src_batch_stride, int64_t src_row_pitch, int rs_h,
    int rs_w, const int32_t* bounds_x, const int32_t* taps_x, int ks_x, const int32_t* bounds_y,
    const int32_t* taps_y, int ks_y, void* dst, int dst_dtype, int out_h, int out_w, const int32_t* crop_top,
    const int32_t* crop_left, const uint8_t* flip, int off_y, int off_x, const void* lut, int flags,
    tb_stream_t stream) {
    TB_REQUIRE(B >= 0 && H >= 1 && W >= 1, "tb_resize: bad source shape B=%d H=%d W=%d", B, H, W);
    TB_REQUIRE(C == 1 || C == 3, "tb_resize: C=%d not supported (Pillow premultiplies alpha for RGBA); use C in {1,3}", C);
    TB_REQUIRE(rs_h >= 1 && rs_w >= 1 && out_h >= 1 && out_w >= 1, "tb_resize: bad output sizes");
    TB_REQUIRE(dst_dtype == TB_U8 || tb_dtype_size(dst_dtype) != 0, "tb_resize: unknown dst_dtype %d", dst_dtype);
    TB_REQUIRE(src_row_pitch >= (int64_t)W * C && src_batch_stride >= 0, "tb_resize: bad source strides");
    if (rs_w != W) {
        TB_REQUIRE(bounds_x && taps_x, "tb_resize: horizontal taps missing");
        TB_REQUIRE(ks_x == tb_resize_ksize(W, rs_w), "tb_resize: ks_x=%d does not match tb_resize_ksize(%d,%d)", ks_x, W, rs_w);
    }
    if (rs_h != H) {
        TB_REQUIRE(bounds_y && taps_y, "tb_resize: vertical taps missing");
        TB_REQUIRE(ks_y == tb_resize_ksize(H, rs_h), "tb_resize: ks_y=%d does not match tb_resize_ksize(%d,%d)", ks_y, H, rs_h);
    }
    if (B == 0) return TB_OK;
    TB_REQUIRE(src && dst && lut, "tb_resize: NULL src/dst/lut");
    TB_REQUIRE(dst_dtype == TB_U8 || ((uintptr_t)dst % tb_dtype_size(dst_dtype)) == 0, "tb_resize: dst misaligned for its dtype");

    TbDevice dev;
    int rc = tb_device(&dev);
    if (rc) return rc;

    RsParams p = {};
    p.src = src; p.B = B; p.H = H; p.W = W; p.bs = src_batch_stride; p.pitch = src_row_pitch;
    p.rs_h = rs_h; p.rs_w = rs_w;
    p.bx = bounds_x; p.kx = taps_x; p.ksx = ks_x; p.by = bounds_y; p.ky = taps_y; p.ksy = ks_y;
    p.dst = dst; p.out_h = out_h; p.out_w = out_w;
    p.top = crop_top; p.left = crop_left; p.flip = flip; p.off_y = off_y; p.off_x = off_x; p.lut = lut;
    p.box = nullptr;
    p.words_ok = (((uintptr_t)src & 3) == 0 && (src_row_pitch & 3) == 0 && (src_batch_stride & 3) == 0) ? 1 : 0;

    // Tile: TC output columns x TR output rows; the intermediate holds every input row the TR
    // resized rows can touch: at most ceil(TR * scale_y) + ks_y rows.
    const double scale_y = rs_h != H ? (double)H / rs_h : 1.0;
    const int budget = 40 * 1024;
    // scale <= 4 on both axes: at most 8 taps are ever non-zero (n <= ceil(2 * support)); 4 up to scale 2
    bool fast = !(flags & TB_FLAG_FORCE_GENERIC) && p.words_ok && rs_w != W && rs_h != H &&
                (int64_t)W <= 4 * (int64_t)rs_w && (int64_t)H <= 4 * (int64_t)rs_h;
    const int nt = ((int64_t)W <= 2 * (int64_t)rs_w && (int64_t)H <= 2 * (int64_t)rs_h) ? 4 : 8;
    // Ring kernel: rows staged once per column tile by bulk copies (16-byte aligned rows), persistent CTAs.
    if (fast && !(flags & (TB_FLAG_TWO_PHASE | TB_FLAG_SLIDE)) && C == 3 && dst_dtype != TB_U8 && ((uintptr_t)src & 15) == 0 &&
        (src_row_pitch & 15) == 0 && (src_batch_stride & 15) == 0 && (((int64_t)W * 3) & 15) == 0 && src_row_pitch < ((int64_t)1 << 31)) {
        RingGeom g;
        int grid = 0, threads = 0;
        if (ring_geometry(p, B, nt, dev, flags, &g, &grid, &threads)) {
            const bool pdl = !(flags & TB_FLAG_NO_PDL);
            cudaStream_t s = (cudaStream_t)stream;
            return dst_dtype == TB_F32 ? launch_ring<float>(p, g, grid, threads, nt, pdl, s)
                                       : launch_ring<uint16_t>(p, g, grid, threads, nt, pdl, s);
        }
    }
    if (fast && !(flags & TB_FLAG_TWO_PHASE) && src_row_pitch < ((int64_t)1 << 32)) {
        int64_t blocks = 0;
        int threads = 0;
        slide_geometry(p, B, dev.sm_count, flags, &blocks, &threads);
        TB_REQUIRE(blocks < (int64_t)1 << 31, "tb_resize: grid too large");
        cudaStream_t s = (cudaStream_t)stream;
        if (dst_dtype == TB_U8)
            return C == 3 ? launch_slide<uint8_t, 3>(p, blocks, threads, nt, s) : launch_slide<uint8_t, 1>(p, blocks, threads, nt, s);
        if (dst_dtype == TB_F32)
            return C == 3 ? launch_slide<float, 3>(p, blocks, threads, nt, s) : launch_slide<float, 1>(p, blocks, threads, nt, s);
        return C == 3 ? launch_slide<uint16_t, 3>(p, blocks, threads, nt, s) : launch_slide<uint16_t, 1>(p, blocks, threads, nt, s);
    }
    int TC = out_w < 256 ? out_w : 256;
    // taller tiles recompute fewer halo rows; keep at least ~4 CTAs per SM in the grid
    int TR = ((int64_t)B * ((out_h + 15) / 16) * ((out_w + TC - 1) / TC) >= 4 * (int64_t)dev.sm_count) ? 16 : 8;
    auto rows_for = [&](int tr) { return rs_h != H ? (int)ceil(tr * scale_y) + ks_y + 1 : tr; };
    auto rowb = [&](int tc) { return (tc * C + 3) & ~3; };
    while (TR > 1 && (int64_t)rows_for(TR) * rowb(TC) > budget) TR >>= 1;
    while (TC > 8 && (int64_t)rows_for(TR) * rowb(TC) > budget) TC = (TC + 1) / 2;
    int64_t smem = (int64_t)rows_for(TR) * rowb(TC);
    TB_REQUIRE(smem <= dev.max_smem_optin - 8192, "tb_resize: a %dx%d tile needs %lld bytes of shared memory",
               TR, TC, (long long)smem);
    int col_pitch = 0;
    if (fast) {
        // transposed planar intermediate [C][TC][col_pitch]: a column holds the tile's input rows plus
        // room for the NT-sample window of the last row; pitch = 4 * odd keeps lanes on distinct banks
        auto pitch_for = [&](int tr) {
            int cpb = (rows_for(tr) + nt + 3) & ~3;
            if (((cpb >> 2) & 1) == 0) cpb += 4;
            return cpb;
        };
        while (TR > 1 && (int64_t)C * TC * pitch_for(TR) + 16 > 44 * 1024) TR >>= 1;
        col_pitch = pitch_for(TR);
        smem = (int64_t)C * TC * col_pitch + 16;   // + slack: the last window may poke past the last column
        if (smem > 44 * 1024) fast = false;
    }
    if (!fast) smem = (int64_t)rows_for(TR) * rowb(TC);
    if (TR > out_h) TR = out_h;
    p.TR = TR; p.TC = TC;
    p.tiles_x = (out_w + TC - 1) / TC;
    p.tiles_y = (out_h + TR - 1) / TR;
    p.row_bytes = fast ? col_pitch : rowb(TC);
    p.ty_per_cta = 1;
    p.chunks_y = p.tiles_y;
    if (fast) {
        // several row tiles per CTA amortise the per-column set-up, as long as the grid stays >= 4 CTAs per SM
        int64_t per = (int64_t)B * p.tiles_x * p.tiles_y / (kRsCtasPerSm * (int64_t)dev.sm_count);
        if ((flags >> 16) & 0xff) per = (flags >> 16) & 0xff;      // development knob
        if (per < 1) per = 1;
        if (per > p.tiles_y) per = p.tiles_y;
        p.ty_per_cta = (int)per;
        p.chunks_y = (p.tiles_y + p.ty_per_cta - 1) / p.ty_per_cta;
    }
    const int64_t blocks = fast ? (int64_t)p.tiles_x * p.chunks_y * B : (int64_t)p.tiles_x * p.tiles_y * B;
    TB_REQUIRE(blocks < (int64_t)1 << 31, "tb_resize: grid too large");
    cudaStream_t s = (cudaStream_t)stream;
    const bool f32 = dst_dtype == TB_F32;
    if (dst_dtype == TB_U8)
        return C == 3 ? launch_resize<uint8_t, 3>(p, smem, blocks, dev.max_smem_optin, fast, nt, s)
                      : launch_resize<uint8_t, 1>(p, smem, blocks, dev.max_smem_optin, fast, nt, s);
    if (C == 3) return f32 ? launch_resize<float, 3>(p, smem, blocks, dev.max_smem_optin, fast, nt, s)
                           : launch_resize<uint16_t, 3>(p, smem, blocks, dev.max_smem_optin, fast, nt, s);
    return f32 ? launch_resize<float, 1>(p, smem, blocks, dev.max_smem_optin, fast, nt, s)
               : launch_resize<uint16_t, 1>(p, smem, blocks, dev.max_smem_optin, fast, nt, s);
}

// T.RandomResizedCrop: every sample resizes its own source box to (out_h, out_w).
extern "C" int tb_resized_crop_bilinear_u8_nhwc_to_nchw(const uint8_t* src, int B, int H, int W, int C,
                                                        int64_t src_batch_stride, int64_t src_row_pitch,
                                                        const int32_t* box, void* dst, int dst_dtype, int out_h,
                                                        int out_w, const uint8_t* flip, const void* lut, int flags,
                                                        tb_stream_t stream) {
    TB_REQUIRE(B >= 0 && H >= 1 && W >= 1, "tb_resized_crop: bad source shape B=%d H=%d W=%d", B, H, W);
    TB_REQUIRE(C == 1 || C == 3, "tb_resized_crop: C=%d not supported (Pillow premultiplies alpha for RGBA); use C in {1,3}", C);
    TB_REQUIRE(out_h >= 1 && out_w >= 1, "tb_resized_crop: bad output size");
    TB_REQUIRE(dst_dtype == TB_U8 || tb_dtype_size(dst_dtype) != 0, "tb_resized_crop: unknown dst_dtype %d", dst_dtype);
    TB_REQUIRE(src_row_pitch >= (int64_t)W * C && src_batch_stride >= 0, "tb_resized_crop: bad source strides");
    if (B == 0) return TB_OK;
    TB_REQUIRE(src && dst && lut && box, "tb_resized_crop: NULL src/dst/lut/box");
    TB_REQUIRE(dst_dtype == TB_U8 || ((uintptr_t)dst % tb_dtype_size(dst_dtype)) == 0, "tb_resized_crop: dst misaligned for its dtype");
    // every box is at most the whole image: a source within 4x of the output keeps every scale <= 4
    if ((int64_t)H > 4 * (int64_t)out_h || (int64_t)W > 4 * (int64_t)out_w)
        return tb_fail(TB_ENOTSUP, "tb_resized_crop: source %dx%d is more than four times the output %dx%d; "
                       "reductions beyond 4x are not built for per-sample boxes", H, W, out_h, out_w);
    const int nt = ((int64_t)H <= 2 * (int64_t)out_h && (int64_t)W <= 2 * (int64_t)out_w) ? 4 : 8;
    if (((uintptr_t)src & 3) || (src_row_pitch & 3) || (src_batch_stride & 3))
        return tb_fail(TB_ENOTSUP, "tb_resized_crop: source base, row pitch and batch stride must be multiples of 4 bytes");

    TbDevice dev;
    int rc = tb_device(&dev);
    if (rc) return rc;
    RsParams p = {};
    p.src = src; p.B = B; p.H = H; p.W = W; p.bs = src_batch_stride; p.pitch = src_row_pitch;
    p.rs_h = out_h; p.rs_w = out_w;
    p.dst = dst; p.out_h = out_h; p.out_w = out_w;
    p.flip = flip; p.lut = lut; p.box = box; p.box_words = 4; p.words_ok = 1;
    if (!(flags & TB_FLAG_TWO_PHASE) && src_row_pitch < ((int64_t)1 << 32)) {
        int64_t blocks = 0;
        int threads = 0;
        slide_geometry(p, B, dev.sm_count, flags, &blocks, &threads);
        TB_REQUIRE(blocks < (int64_t)1 << 31, "tb_resized_crop: grid too large");
        cudaStream_t s = (cudaStream_t)stream;
        if (dst_dtype == TB_U8)
            return C == 3 ? launch_slide<uint8_t, 3>(p, blocks, threads, nt, s) : launch_slide<uint8_t, 1>(p, blocks, threads, nt, s);
        if (dst_dtype == TB_F32)
            return C == 3 ? launch_slide<float, 3>(p, blocks, threads, nt, s) : launch_slide<float, 1>(p, blocks, threads, nt, s);
        return C == 3 ? launch_slide<uint16_t, 3>(p, blocks, threads, nt, s) : launch_slide<uint16_t, 1>(p, blocks, threads, nt, s);
    }
    const int TC = out_w < 256 ? out_w : 256;
    int TR = ((int64_t)B * ((out_h + 15) / 16) * ((out_w + TC - 1) / TC) >= 4 * (int64_t)dev.sm_count) ? 16 : 8;
    // up to nt/2 input rows per output row + the tap overhang
    auto pitch_for = [&](int tr) {
        int cpb = ((nt / 2) * tr + nt + 2 + nt + 3) & ~3;
        if (((cpb >> 2) & 1) == 0) cpb += 4;
        return cpb;
    };
    while (TR > 1 && (int64_t)C * TC * pitch_for(TR) + 16 > 44 * 1024) TR >>= 1;
    const int col_pitch = pitch_for(TR);
    p.TR = TR > out_h ? out_h : TR; p.TC = TC;
    p.tiles_x = (out_w + TC - 1) / TC;
    p.tiles_y = (out_h + p.TR - 1) / p.TR;
    p.row_bytes = col_pitch;
    {
        int64_t per = (int64_t)B * p.tiles_x * p.tiles_y / (kRsCtasPerSm * (int64_t)dev.sm_count);
        if ((flags >> 16) & 0xff) per = (flags >> 16) & 0xff;      // development knob
        if (per < 1) per = 1;
        if (per > p.tiles_y) per = p.tiles_y;
        p.ty_per_cta = (int)per;
        p.chunks_y = (p.tiles_y + p.ty_per_cta - 1) / p.ty_per_cta;
    }
    const int64_t smem = (int64_t)C * TC * col_pitch + 16;
    const int64_t blocks = (int64_t)p.tiles_x * p.chunks_y * B;
    TB_REQUIRE(blocks < (int64_t)1 << 31, "tb_resized_crop: grid too large");
    cudaStream_t s = (cudaStream_t)stream;
    const bool f32 = dst_dtype == TB_F32;
    if (dst_dtype == TB_U8)     // bytes out: the resized crop hands a uint8 image on (to tb_augment)
        return C == 3 ? launch_resize<uint8_t, 3>(p, smem, blocks, dev.max_smem_optin, true, nt, s)
                      : launch_resize<uint8_t, 1>(p, smem, blocks, dev.max_smem_optin, true, nt, s);
    if (C == 3) return f32 ? launch_resize<float, 3>(p, smem, blocks, dev.max_smem_optin, true, nt, s)
                           : launch_resize<uint16_t, 3>(p, smem, blocks, dev.max_smem_optin, true, nt, s);
    return f32 ? launch_resize<float, 1>(p, smem, blocks, dev.max_smem_optin, true, nt, s)
               : launch_resize<uint16_t, 1>(p, smem, blocks, dev.max_smem_optin, true, nt, s);
}

// Ragged batch: every sample has its own image size, source region, resized size and crop offset
// (T.Resize(int) + T.CenterCrop, or T.RandomResizedCrop, over images of different sizes).
extern "C" int tb_resize_crop_ragged_u8_to_nchw(const uint8_t* src, const int64_t* offsets, const int32_t* pitches, int B,
                                                int C, const int32_t* geometry, int max_taps, void* dst, int dst_dtype,
                                                int out_h, int out_w, const uint8_t* flip, const void* lut, int flags,
                                                tb_stream_t stream) {
    TB_REQUIRE(B >= 0, "tb_resize_crop_ragged: negative B");
    TB_REQUIRE(C == 1 || C == 3, "tb_resize_crop_ragged: C=%d not supported (Pillow premultiplies alpha for RGBA); use C in {1,3}", C);
    TB_REQUIRE(out_h >= 1 && out_w >= 1, "tb_resize_crop_ragged: bad output size");
    TB_REQUIRE(max_taps == 4 || max_taps == 8, "tb_resize_crop_ragged: max_taps is 4 (every scale <= 2) or 8 (<= 4)");
    TB_REQUIRE(dst_dtype == TB_U8 || tb_dtype_size(dst_dtype) != 0, "tb_resize_crop_ragged: unknown dst_dtype %d", dst_dtype);
    if (B == 0) return TB_OK;
    TB_REQUIRE(src && offsets && pitches && geometry && dst && lut, "tb_resize_crop_ragged: NULL argument");
    TB_REQUIRE(((uintptr_t)src & 3) == 0, "tb_resize_crop_ragged: src must be 4-byte aligned (and every offset / pitch a multiple of 4)");
    TB_REQUIRE(dst_dtype == TB_U8 || ((uintptr_t)dst % tb_dtype_size(dst_dtype)) == 0, "tb_resize_crop_ragged: dst misaligned for its dtype");
    TbDevice dev;
    int rc = tb_device(&dev);
    if (rc) return rc;
    RsParams p = {};
    p.src = src; p.B = B; p.offsets = offsets; p.pitches = pitches;
    p.rs_h = out_h; p.rs_w = out_w;
    p.dst = dst; p.out_h = out_h; p.out_w = out_w;
    p.flip = flip; p.lut = lut; p.box = geometry; p.box_words = 10; p.words_ok = 1;
    int64_t blocks = 0;
    int threads = 0;
    slide_geometry(p, B, dev.sm_count, flags, &blocks, &threads);
    TB_REQUIRE(blocks < (int64_t)1 << 31, "tb_resize_crop_ragged: grid too large");
    cudaStream_t s = (cudaStream_t)stream;
    if (dst_dtype == TB_U8)
        return C == 3 ? launch_slide<uint8_t, 3>(p, blocks, threads, max_taps, s) : launch_slide<uint8_t, 1>(p, blocks, threads, max_taps, s);
    if (dst_dtype == TB_F32)
        return C == 3 ? launch_slide<float, 3>(p, blocks, threads, max_taps, s) : launch_slide<float, 1>(p, blocks, threads, max_taps, s);
    return C == 3 ? launch_slide<uint16_t, 3>(p, blocks, threads, max_taps, s) : launch_slide<uint16_t, 1>(p, blocks, threads, max_taps, s);
}
