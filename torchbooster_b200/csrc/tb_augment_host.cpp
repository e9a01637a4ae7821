// Host side of tb_augment: one sequential walk over torch's mt19937 words that makes, per sample,
// the draws of RandomCrop -> RandomHorizontalFlip -> RandomRotation -> RandAugment
// (examples/img_cls/resnet/resnet.py:96-100; torchvision/transforms/autoaugment.py:340-365) and
// writes the kernel's operation records directly.  The Python statement of the same walk is
// torchbooster_b200/augment.py (AugmentParams.draw + encode_ops); this one exists because a
// Python loop per sample would cap the pipeline far below the kernel.
//
// Compiled with -ffp-contract=off: the matrix arithmetic below must round exactly like CPython's
// float expressions in PIL.Image.rotate.
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#include "tb_common.cuh"

namespace {

constexpr int kWords = 8;

inline int32_t wrap32(int64_t v) { return (int32_t)(uint32_t)(uint64_t)v; }
inline int64_t fix16(double v) { return (int64_t)floor(v * 65536.0 + 0.5); }

// CPython's round(x, 15) for |x| <= 1: the decimal with 15 digits after the point nearest to the
// exact value of x (ties to even), read back as a double.  q = that decimal's integer numerator:
// x * 1e15 split exactly into a rounded product and its FMA residual, rounded half-even; q / 1e15
// is then ONE correctly rounded division of two exact doubles, which is what strtod returns.
inline double round15(double x) {
    if (!(fabs(x) <= 1.0)) {                 // not reached for cos / sin; the slow exact route
        char buf[64];
        snprintf(buf, sizeof buf, "%.15f", x);
        return strtod(buf, nullptr);
    }
    const double p = x * 1e15;
    const double err = fma(x, 1e15, -p);     // exact: x * 1e15 == p + err
    double q = nearbyint(p);                 // half-even on p
    const double d = (p - q) + err;          // exact distance of the true product from q (|p| < 2^53)
    if (d > 0.5 || (d == 0.5 && fmod(q, 2.0) != 0.0)) q += 1.0;
    else if (d < -0.5 || (d == -0.5 && fmod(q, 2.0) != 0.0)) q -= 1.0;
    return q / 1e15;
}

// PIL.Image.rotate(angle), expand False, centre (w/2, h/2), nearest -> Pillow's 16.16 coefficients.
bool rotation_record(double angle, int height, int width, int32_t* rec) {
    for (int i = 0; i < kWords; ++i) rec[i] = 0;
    double a = fmod(angle, 360.0);          // Python's float %: the sign of the divisor
    if (a != 0.0 && a < 0.0) a += 360.0;
    if (a == 0.0) return true;              // TB_AUG_IDENTITY
    const int64_t one = 65536, half = 32768;
    rec[0] = TB_AUG_AFFINE;
    if (a == 180.0) {
        rec[1] = wrap32(-one); rec[3] = wrap32((int64_t)(width - 1) * one + half);
        rec[5] = wrap32(-one); rec[6] = wrap32((int64_t)(height - 1) * one + half);
        return true;
    }
    if (a == 90.0 && width == height) {
        rec[2] = wrap32(-one); rec[3] = wrap32((int64_t)(width - 1) * one + half);
        rec[4] = wrap32(one); rec[6] = wrap32(half);
        return true;
    }
    if (a == 270.0 && width == height) {
        rec[2] = wrap32(one); rec[3] = wrap32(half);
        rec[4] = wrap32(-one); rec[6] = wrap32((int64_t)(height - 1) * one + half);
        return true;
    }
    const double cx = width / 2.0, cy = height / 2.0;
    const double r = -(a * (M_PI / 180.0));                 // -math.radians(a)
    const double rc = round15(cos(r)), rs = round15(sin(r));     // round15 is odd: round15(-s) == -round15(s)
    double m[6] = {rc, rs, 0.0, -rs, rc, 0.0};
    const double t2 = m[0] * -cx + m[1] * -cy + m[2];
    const double t5 = m[3] * -cx + m[4] * -cy + m[5];
    m[2] = t2 + cx;
    m[5] = t5 + cy;
    if (m[1] == 0.0 && m[3] == 0.0) return false;           // Pillow's scale path: not a rotation we encode
    const double xs[4] = {0.0, (double)width, 0.0, (double)width}, ys[4] = {0.0, (double)height, (double)height, 0.0};
    for (int k = 0; k < 4; ++k)
        if (!(fabs(xs[k] * m[0] + ys[k] * m[1] + m[2]) < 32768.0 && fabs(xs[k] * m[3] + ys[k] * m[4] + m[5]) < 32768.0))
            return false;
    rec[1] = wrap32(fix16(m[0]));
    rec[2] = wrap32(fix16(m[1]));
    rec[3] = wrap32(fix16(m[2] + m[0] * 0.5 + m[1] * 0.5));
    rec[4] = wrap32(fix16(m[3]));
    rec[5] = wrap32(fix16(m[4]));
    rec[6] = wrap32(fix16(m[5] + m[3] * 0.5 + m[4] * 0.5));
    return true;
}

}  // namespace

extern "C" int tb_augment_rotation_records(const double* angles, int64_t n, int height, int width, int32_t* records) {
    TB_REQUIRE(n >= 0 && height >= 1 && width >= 1 && (n == 0 || (angles && records)), "tb_augment_rotation_records: bad argument");
    for (int64_t i = 0; i < n; ++i)
        if (!rotation_record(angles[i], height, width, records + i * kWords))
            return tb_fail(TB_ENOTSUP, "tb_augment_rotation_records: angle %.17g is outside Pillow's fixed-point path", angles[i]);
    return TB_OK;
}

extern "C" int64_t tb_augment_params_walk(const uint64_t* words, int64_t n_words, int64_t batch, int crop_range_h,
                                          int crop_range_w, double flip_p, int has_rotation, float rot_lo,
                                          float rot_span, int fused, int height, int width, int n_ops,
                                          int n_choices, const int32_t* op_table, const uint8_t* op_signed,
                                          int32_t* top, int32_t* left, uint8_t* flip, int32_t* records) {
    if (!words || batch < 0 || n_ops < 0 || height < 1 || width < 1 || !top || !left || !flip ||
        ((has_rotation || n_ops) && !records) || (n_ops && (!op_table || !op_signed || n_choices < 1)))
        return tb_fail(TB_EINVAL, "tb_augment_params_walk: bad argument");
    const bool crop = crop_range_h > 0 && crop_range_w > 0;
    const int per = (has_rotation ? 1 : 0) + n_ops;
    const int64_t most = (crop ? 2 : 0) + (flip_p >= 0.0 ? 1 : 0) + (has_rotation ? 1 : 0) + 2 * (int64_t)n_ops;
    if (batch * most > n_words) return tb_fail(TB_EINVAL, "tb_augment_params_walk: not enough words");
    int64_t pos = 0;
    for (int64_t b = 0; b < batch; ++b) {
        top[b] = left[b] = 0;
        flip[b] = 0;
        if (crop) {
            top[b] = (int32_t)(words[pos] % (uint64_t)crop_range_h);
            left[b] = (int32_t)(words[pos + 1] % (uint64_t)crop_range_w);
            pos += 2;
        }
        if (flip_p >= 0.0) {
            const float u = (float)(words[pos] & 0xFFFFFFull) * 5.9604644775390625e-08f;    // 2^-24
            flip[b] = u < (float)flip_p ? 1 : 0;
            pos += 1;
        }
        int32_t* rec = records ? records + b * per * kWords : nullptr;
        if (has_rotation) {
            // torch's uniform_: x * (to - from) + from in float32, one FMA in its vectorised CPU
            // kernels, two roundings in the DEFAULT one
            const float x = (float)(words[pos] & 0xFFFFFFull) * 5.9604644775390625e-08f;
            float angle;
            if (fused) {
                angle = fmaf(x, rot_span, rot_lo);
            } else {
                volatile float prod = x * rot_span;
                angle = prod + rot_lo;
            }
            pos += 1;
            if (!rotation_record((double)angle, height, width, rec))
                return tb_fail(TB_ENOTSUP, "tb_augment_params_walk: angle %.9g is outside Pillow's fixed-point path", (double)angle);
            rec += kWords;
        }
        for (int o = 0; o < n_ops; ++o, rec += kWords) {
            const int which = (int)(words[pos] % (uint64_t)n_choices);
            pos += 1;
            int negative = 0;
            if (op_signed[which]) {
                negative = (int)(words[pos] % 2u);
                pos += 1;
            }
            const int32_t* src = op_table + ((int64_t)which * 2 + negative) * kWords;
            for (int i = 0; i < kWords; ++i) rec[i] = src[i];
        }
    }
    return pos;
}
