/* condition_oracle.cpp — CPU ORACLE (test infrastructure only; see orb_oracle.h for the rules).
 * Input conditioning in front of the extractor: restates the OpenCV calls of
 *   /root/reference/Examples/Stereo-Inertial/stereo_inertial_euroc.cc:201-202  cv::remap(..., cv::INTER_LINEAR), CV_32FC1 maps
 *   /root/reference/src/Tracking.cc:1303-1328                                  cv::cvtColor(..., *2GRAY)
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

}  /* extern "C" */
