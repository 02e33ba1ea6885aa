/* condition_oracle.cpp — CPU ORACLE (test infrastructure only; see orb_oracle.h for the rules).
 * Input conditioning in front of the extractor: restates the OpenCV calls of
 *   /root/reference/Examples/Stereo-Inertial/stereo_inertial_euroc.cc:201-202  cv::remap(..., cv::INTER_LINEAR), CV_32FC1 maps
 *   /root/reference/src/Tracking.cc:1303-1328                                  cv::cvtColor(..., *2GRAY)
 *   /root/reference/Examples/Stereo-Inertial/stereo_inertial_tum_vi.cc:128-138 cv::createCLAHE(3.0, cv::Size(8, 8))->apply(im, im)
 * OpenCV is not vendored in the reference; the arithmetic is pinned against cv2 4.13.0 (tools/make_goldens.py ->
 * tests/golden/conditioning.npz). */
#include <cmath>
#include <cstdint>
#include <cstddef>

extern "C" {

/* cv::remap, 8UC1, INTER_LINEAR, BORDER_CONSTANT(0): maps to 5-bit fixed point with cvRound, INTER_REMAP_COEF_SCALE = 2^15. */
void orc_remap_linear_u8(const uint8_t* src, int w, int h, int sstep, const float* mapx, const float* mapy, uint8_t* dst, int dstep) {
    for (int y = 0; y < h; y++)
        for (int x = 0; x < w; x++) {
            const int sx = (int)lrintf(mapx[(size_t)y * w + x] * 32.0f), sy = (int)lrintf(mapy[(size_t)y * w + x] * 32.0f);
            int ix = sx >> 5, iy = sy >> 5;
            ix = ix < -32768 ? -32768 : (ix > 32767 ? 32767 : ix); /* saturate_cast<short> */
            iy = iy < -32768 ? -32768 : (iy > 32767 ? 32767 : iy);
            const int fx = sx & 31, fy = sy & 31;
            int acc = 0;
            for (int dy = 0; dy < 2; dy++)
                for (int dx = 0; dx < 2; dx++) {
                    const int xx = ix + dx, yy = iy + dy;
                    const int tap = (xx >= 0 && xx < w && yy >= 0 && yy < h) ? src[(ptrdiff_t)yy * sstep + xx] : 0;
                    acc += tap * ((dx ? fx : 32 - fx) * (dy ? fy : 32 - fy) * 32);
                }
            dst[(ptrdiff_t)y * dstep + x] = (uint8_t)((acc + 16384) >> 15);
        }
}

/* cv::cvtColor to gray, 8U: (R*9798 + G*19235 + B*3735 + 2^14) >> 15. */
void orc_to_gray(const uint8_t* src, long long npix, int channels, int rgb, uint8_t* dst) {
    for (long long i = 0; i < npix; i++) {
        const uint8_t* p = src + i * channels;
        const int r = rgb ? p[0] : p[2], b = rgb ? p[2] : p[0];
        dst[i] = (uint8_t)((r * 9798 + p[1] * 19235 + b * 3735 + 16384) >> 15);
    }
}

/* cv::CLAHE::apply, 8UC1 (OpenCV imgproc/src/clahe.cpp, CLAHE_CalcLut_Body + CLAHE_Interpolation_Body).
 * An image whose size is not a multiple of the tile grid is extended to the right/bottom by tiles - size % tiles pixels
 * (a whole `tiles` in the dimension that does divide) with BORDER_REFLECT_101 for the histograms only.
 * lut: optional [tiles_y][tiles_x][256] output for stage tests. Returns 0, or -1 for sizes the restatement does not cover. */
int orc_clahe_u8(const uint8_t* src, int w, int h, int sstep, double clip, int tiles_x, int tiles_y, uint8_t* dst, int dstep, uint8_t* lut_out) {
    if (w < 1 || h < 1 || tiles_x < 1 || tiles_y < 1 || tiles_x > 64 || tiles_y > 64) return -1;
    const bool whole = w % tiles_x == 0 && h % tiles_y == 0;
    const int ew = whole ? w : w + tiles_x - w % tiles_x, eh = whole ? h : h + tiles_y - h % tiles_y;
    if (ew - w >= w || eh - h >= h) return -1; /* reflection would wrap more than once */
    const int tw = ew / tiles_x, th = eh / tiles_y, area = tw * th;
    int clipLimit = 0;
    if (clip > 0.0) { clipLimit = (int)(clip * area / 256); if (clipLimit < 1) clipLimit = 1; }
    const float lutScale = 255.0f / (float)area;
    uint8_t* lut = new uint8_t[(size_t)tiles_x * tiles_y * 256];
    for (int tyi = 0; tyi < tiles_y; tyi++)
        for (int txi = 0; txi < tiles_x; txi++) {
            int hist[256] = {0};
            for (int y = tyi * th; y < (tyi + 1) * th; y++) {
                const int sy = y < h ? y : 2 * h - 2 - y;
                for (int x = txi * tw; x < (txi + 1) * tw; x++) hist[src[(ptrdiff_t)sy * sstep + (x < w ? x : 2 * w - 2 - x)]]++;
            }
            if (clipLimit > 0) {
                int clipped = 0;
                for (int k = 0; k < 256; k++)
                    if (hist[k] > clipLimit) { clipped += hist[k] - clipLimit; hist[k] = clipLimit; }
                const int redist = clipped / 256;
                int residual = clipped - redist * 256;
                for (int k = 0; k < 256; k++) hist[k] += redist;
                if (residual != 0) {
                    const int step = 256 / residual > 1 ? 256 / residual : 1;
                    for (int k = 0; k < 256 && residual > 0; k += step, residual--) hist[k]++;
                }
            }
            uint8_t* l = lut + ((size_t)tyi * tiles_x + txi) * 256;
            int sum = 0;
            for (int k = 0; k < 256; k++) {
                sum += hist[k];
                const long r = lrintf((float)sum * lutScale);
                l[k] = (uint8_t)(r < 0 ? 0 : (r > 255 ? 255 : r));
            }
        }
    const float inv_tw = 1.0f / (float)tw, inv_th = 1.0f / (float)th;
    for (int y = 0; y < h; y++) {
        const float tyf = (float)y * inv_th - 0.5f;
        int ty1 = (int)floorf(tyf), ty2 = ty1 + 1;
        const float ya = tyf - (float)ty1, ya1 = 1.0f - ya;
        ty1 = ty1 < 0 ? 0 : ty1; ty2 = ty2 > tiles_y - 1 ? tiles_y - 1 : ty2;
        for (int x = 0; x < w; x++) {
            const float txf = (float)x * inv_tw - 0.5f;
            int tx1 = (int)floorf(txf), tx2 = tx1 + 1;
            const float xa = txf - (float)tx1, xa1 = 1.0f - xa;
            tx1 = tx1 < 0 ? 0 : tx1; tx2 = tx2 > tiles_x - 1 ? tiles_x - 1 : tx2;
            const int v = src[(ptrdiff_t)y * sstep + x];
            const float a = lut[((size_t)ty1 * tiles_x + tx1) * 256 + v], b = lut[((size_t)ty1 * tiles_x + tx2) * 256 + v];
            const float c = lut[((size_t)ty2 * tiles_x + tx1) * 256 + v], d = lut[((size_t)ty2 * tiles_x + tx2) * 256 + v];
            const float res = (a * xa1 + b * xa) * ya1 + (c * xa1 + d * xa) * ya;
            const long r = lrintf(res);
            dst[(ptrdiff_t)y * dstep + x] = (uint8_t)(r < 0 ? 0 : (r > 255 ? 255 : r));
        }
    }
    if (lut_out) for (size_t i = 0; i < (size_t)tiles_x * tiles_y * 256; i++) lut_out[i] = lut[i];
    delete[] lut;
    return 0;
}

}  /* extern "C" */
