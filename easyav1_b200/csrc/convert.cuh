// convert.cuh — output conversion of a finished picture (formats and formulas: include/easyav1_cuda.h).
// One thread per luma sample; for the 4:2:0 semi-planar formats the thread of every even (x, y) also writes the
// interleaved chroma pair. Device code shared with the host emulation (tests/emu).
#pragma once
#include "dev_common.cuh"

namespace eav1 {

struct ConvParams {
    Plane src[3];
    int w, h, ss_hor, ss_ver, mono, bpc, format;
    int ky, krv, kgu, kgv, kbu, y_off, c_off;
    uint8_t *dst;
    long long dst_stride;
};

// the coefficients of the RGBA8 formula (host side, double -> rounded integers)
inline void conv_coefficients(ConvParams &P, int matrix, int full_range)
{
    const double kr = matrix == 1 ? 0.2126 : matrix == 9 ? 0.2627 : 0.299;
    const double kb = matrix == 1 ? 0.0722 : matrix == 9 ? 0.0593 : 0.114;
    const double kg = 1.0 - kr - kb;
    const int sh = P.bpc - 8;
    const double yr = (double)((full_range ? 255 : 219) << sh), cr = (double)((full_range ? 255 : 224) << sh);
    auto q = [](double v) { return (int)(v + 0.5); };
    P.ky = q(8192.0 * 255.0 / yr);
    P.krv = q(8192.0 * 255.0 * 2.0 * (1.0 - kr) / cr);
    P.kgu = q(8192.0 * 255.0 * 2.0 * (1.0 - kb) * kb / kg / cr);
    P.kgv = q(8192.0 * 255.0 * 2.0 * (1.0 - kr) * kr / kg / cr);
    P.kbu = q(8192.0 * 255.0 * 2.0 * (1.0 - kb) / cr);
    P.y_off = full_range ? 0 : 16 << sh;
    P.c_off = 128 << sh;
}

template <typename pixel>
EAV1_DEVH void convert_sample(const ConvParams &P, int x, int y)
{
    const int Y = *px_at<pixel>(P.src[0], x, y);
    if (P.format == 3) {   // RGBA8
        int u = 0, v = 0;
        if (!P.mono) {
            u = (int)*px_at<pixel>(P.src[1], x >> P.ss_hor, y >> P.ss_ver) - P.c_off;
            v = (int)*px_at<pixel>(P.src[2], x >> P.ss_hor, y >> P.ss_ver) - P.c_off;
        }
        const int yy = P.ky * (Y - P.y_off) + 4096;
        const int r = iclip((yy + P.krv * v) >> 13, 0, 255);
        const int g = iclip((yy - P.kgu * u - P.kgv * v) >> 13, 0, 255);
        const int b = iclip((yy + P.kbu * u) >> 13, 0, 255);
        *reinterpret_cast<uint32_t *>(P.dst + y * P.dst_stride + 4 * (long long)x) =
            (uint32_t)r | ((uint32_t)g << 8) | ((uint32_t)b << 16) | 0xff000000u;
        return;
    }
    const bool pair = !(x & 1) && !(y & 1);
    if (P.format == 1) {   // NV12
        P.dst[y * P.dst_stride + x] = (uint8_t)Y;
        if (pair) {
            uint8_t *c = P.dst + (P.h + (y >> 1)) * P.dst_stride + x;
            c[0] = (uint8_t)*px_at<pixel>(P.src[1], x >> 1, y >> 1);
            c[1] = (uint8_t)*px_at<pixel>(P.src[2], x >> 1, y >> 1);
        }
    } else {               // P010
        const int up = 16 - P.bpc;
        *reinterpret_cast<uint16_t *>(P.dst + y * P.dst_stride + 2 * (long long)x) = (uint16_t)(Y << up);
        if (pair) {
            uint16_t *c = reinterpret_cast<uint16_t *>(P.dst + (P.h + (y >> 1)) * P.dst_stride + 2 * (long long)x);
            c[0] = (uint16_t)(*px_at<pixel>(P.src[1], x >> 1, y >> 1) << up);
            c[1] = (uint16_t)(*px_at<pixel>(P.src[2], x >> 1, y >> 1) << up);
        }
    }
}

// argument checks and output size shared by the engine and the emulation; returns the bytes needed or 0 if invalid
inline size_t conv_needed(int format, int layout, int bpc, int w, int h, size_t stride)
{
    if (format == 1) return (layout == 1 && bpc == 8 && stride >= (size_t)((w + 1) & ~1)) ? stride * (size_t)(h + (h + 1) / 2) : 0;
    if (format == 2) return (layout == 1 && bpc > 8 && stride >= 2 * (size_t)((w + 1) & ~1)) ? stride * (size_t)(h + (h + 1) / 2) : 0;
    if (format == 3) return stride >= 4 * (size_t)w ? stride * (size_t)h : 0;
    return 0;
}

} // namespace eav1
