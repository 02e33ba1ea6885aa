/* orbfe_condition.cu — input conditioning in front of the extractor, so that frames uploaded once
 * stay on the GPU until the keypoints come back:
 *   k_remap_linear   cv::remap(im, imRect, M1, M2, cv::INTER_LINEAR) with CV_32FC1 maps, BORDER_CONSTANT(0): the stereo
 *                    rectification of the example drivers (/root/reference/Examples/Stereo-Inertial/stereo_inertial_euroc.cc:201-202,
 *                    maps from cv::initUndistortRectifyMap(..., CV_32F, ...) at :150-151)
 *   k_to_gray        cv::cvtColor(im, gray, RGB2GRAY / BGR2GRAY / RGBA2GRAY / BGRA2GRAY) of Tracking::GrabImage*
 *                    (/root/reference/src/Tracking.cc:1303-1328)
 * Arithmetic (verified bit-exact against cv2 4.13.0, tests/golden/conditioning.npz):
 *   remap: sx = cvRound(mapx*32), ix = sx >> 5, fx = sx & 31 (same for y, ix/iy saturated to int16); weights
 *          (32-fx)(32-fy)*32, fx(32-fy)*32, (32-fx)fy*32, fx*fy*32 (sum 2^15); dst = (sum w*tap + 2^14) >> 15; taps outside
 *          the image read 0.
 *   gray:  (R*9798 + G*19235 + B*3735 + 2^14) >> 15. */
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>

#include <new>
#include <vector>

#include "orbfe_internal.h"

struct orbfe_rectifier {
    int device, W, H;
    uint2* dMap;     /* per output pixel: (ix & 0xFFFF) | iy << 16,  fx | fy << 8 */
    uint8_t* dSrc;   /* staging for the host entry point */
    uint8_t* dDst;
    cudaStream_t stream;
};

/* one thread = four adjacent output pixels (one 32-bit store) */
__global__ void __launch_bounds__(256) k_remap_linear(const uint2* __restrict__ map, const uint8_t* __restrict__ src, uint8_t* __restrict__ dst,
                                                      int W, int H, int groupsPerRow) {
    const int g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= groupsPerRow * H) return;
    const int y = g / groupsPerRow, x0 = (g - y * groupsPerRow) * 4;
    const size_t frame = (size_t)blockIdx.y * W * H;
    const uint8_t* s = src + frame;
    uint32_t out = 0;
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const int x = x0 + k;
        if (x >= W) break;
        const uint2 m = __ldg(map + (size_t)y * W + x);
        const int ix = (int)(short)(m.x & 0xFFFF), iy = (int)(short)(m.x >> 16);
        const int fx = m.y & 0xFF, fy = (m.y >> 8) & 0xFF;
        const bool x0in = ix >= 0 && ix < W, x1in = ix + 1 >= 0 && ix + 1 < W;
        const bool y0in = iy >= 0 && iy < H, y1in = iy + 1 >= 0 && iy + 1 < H;
        const uint8_t* p = s + (ptrdiff_t)iy * W + ix;
        const int t00 = (x0in && y0in) ? p[0] : 0, t01 = (x1in && y0in) ? p[1] : 0;
        const int t10 = (x0in && y1in) ? p[W] : 0, t11 = (x1in && y1in) ? p[W + 1] : 0;
        const int w00 = (32 - fx) * (32 - fy) * 32, w01 = fx * (32 - fy) * 32, w10 = (32 - fx) * fy * 32, w11 = fx * fy * 32;
        out |= (uint32_t)((t00 * w00 + t01 * w01 + t10 * w10 + t11 * w11 + 16384) >> 15) << (8 * k);
    }
    uint8_t* d = dst + frame + (size_t)y * W + x0;
    if (x0 + 4 <= W && (((size_t)y * W + x0) & 3) == 0 && ((reinterpret_cast<uintptr_t>(dst) + frame) & 3) == 0) *reinterpret_cast<uint32_t*>(d) = out;
    else
        for (int k = 0; k < 4 && x0 + k < W; k++) d[k] = (uint8_t)(out >> (8 * k));
}

/* one thread = one pixel; channels = 3 or 4; rgb != 0: the first channel is R */
__global__ void __launch_bounds__(256) k_to_gray(const uint8_t* __restrict__ src, uint8_t* __restrict__ dst, long long npix, int channels, int rgb) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= npix) return;
    const uint8_t* p = src + i * channels;
    const int c0 = p[0], c1 = p[1], c2 = p[2];
    const int r = rgb ? c0 : c2, b = rgb ? c2 : c0;
    dst[i] = (uint8_t)((r * 9798 + c1 * 19235 + b * 3735 + 16384) >> 15);
}

extern "C" {

void orbfe_rectifier_destroy(orbfe_rectifier* r) {
    if (!r) return;
    cudaSetDevice(r->device);
    if (r->stream) cudaStreamSynchronize(r->stream);
    cudaFree(r->dMap); cudaFree(r->dSrc); cudaFree(r->dDst);
    if (r->stream) cudaStreamDestroy(r->stream);
    delete r;
}

int orbfe_rectifier_create(int device, int width, int height, const float* map_x, const float* map_y, orbfe_rectifier** out) {
    if (!out || !map_x || !map_y || width < 1 || height < 1 || width > 16384 || height > 16384) { orbfe_set_error("bad argument"); return ORBFE_E_INVALID; }
    *out = nullptr;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || device < 0 || device >= ndev) {
        orbfe_set_error("no usable CUDA device %d (found %d); this library has no CPU path", device, ndev);
        return ORBFE_E_CUDA;
    }
    ORBFE_CUDA(cudaSetDevice(device));
    orbfe_rectifier* r = new (std::nothrow) orbfe_rectifier();
    if (!r) return ORBFE_E_INVALID;
    r->device = device; r->W = width; r->H = height; r->dMap = nullptr; r->dSrc = nullptr; r->dDst = nullptr; r->stream = nullptr;
    /* float maps -> OpenCV's 5-bit fixed point, once (cv::remap does this per call, per block) */
    std::vector<uint2> m((size_t)width * height);
    for (size_t i = 0; i < m.size(); i++) {
        const int sx = (int)lrintf(map_x[i] * 32.0f), sy = (int)lrintf(map_y[i] * 32.0f);
        int ix = sx >> 5, iy = sy >> 5;
        ix = ix < -32768 ? -32768 : (ix > 32767 ? 32767 : ix);
        iy = iy < -32768 ? -32768 : (iy > 32767 ? 32767 : iy);
        m[i] = make_uint2((uint32_t)(ix & 0xFFFF) | ((uint32_t)(iy & 0xFFFF) << 16), (uint32_t)(sx & 31) | ((uint32_t)(sy & 31) << 8));
    }
    const size_t npix = (size_t)width * height;
    cudaError_t e = cudaStreamCreateWithFlags(&r->stream, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaMalloc((void**)&r->dMap, npix * sizeof(uint2));
    if (e == cudaSuccess) e = cudaMalloc((void**)&r->dSrc, npix);
    if (e == cudaSuccess) e = cudaMalloc((void**)&r->dDst, npix);
    if (e == cudaSuccess) e = cudaMemcpy(r->dMap, m.data(), npix * sizeof(uint2), cudaMemcpyHostToDevice);
    if (e != cudaSuccess) { const int rc = orbfe_cuda_fail(e, "rectifier allocation"); orbfe_rectifier_destroy(r); return rc; }
    *out = r;
    return ORBFE_OK;
}

int orbfe_rectify_batch_device(orbfe_rectifier* r, const uint8_t* d_src, uint8_t* d_dst, int nframes, void* stream) {
    if (!r || !d_src || !d_dst || nframes < 1 || nframes > 65535) { orbfe_set_error("bad argument"); return ORBFE_E_INVALID; }
    ORBFE_CUDA(cudaSetDevice(r->device));
    const int gpr = (r->W + 3) / 4;
    k_remap_linear<<<dim3((gpr * r->H + 255) / 256, nframes), 256, 0, stream ? (cudaStream_t)stream : r->stream>>>(r->dMap, d_src, d_dst, r->W, r->H, gpr);
    ORBFE_CUDA(cudaGetLastError());
    return ORBFE_OK;
}

int orbfe_rectify(orbfe_rectifier* r, const uint8_t* src, size_t sstep, uint8_t* dst, size_t dstep) {
    if (!r || !src || !dst || sstep < (size_t)r->W || dstep < (size_t)r->W) { orbfe_set_error("bad argument"); return ORBFE_E_INVALID; }
    ORBFE_CUDA(cudaSetDevice(r->device));
    ORBFE_CUDA(cudaMemcpy2DAsync(r->dSrc, r->W, src, sstep, r->W, r->H, cudaMemcpyHostToDevice, r->stream));
    int rc = orbfe_rectify_batch_device(r, r->dSrc, r->dDst, 1, r->stream);
    if (rc) return rc;
    ORBFE_CUDA(cudaMemcpy2DAsync(dst, dstep, r->dDst, r->W, r->W, r->H, cudaMemcpyDeviceToHost, r->stream));
    ORBFE_CUDA(cudaStreamSynchronize(r->stream));
    return ORBFE_OK;
}

int orbfe_to_gray_device(int device, const uint8_t* d_src, int channels, int rgb_order, int width, int height, int nframes, uint8_t* d_dst, void* stream) {
    if (!d_src || !d_dst || (channels != 3 && channels != 4) || width < 1 || height < 1 || nframes < 1) { orbfe_set_error("bad argument"); return ORBFE_E_INVALID; }
    ORBFE_CUDA(cudaSetDevice(device));
    const long long npix = (long long)width * height * nframes;
    k_to_gray<<<(unsigned)((npix + 255) / 256), 256, 0, (cudaStream_t)stream>>>(d_src, d_dst, npix, channels, rgb_order ? 1 : 0);
    ORBFE_CUDA(cudaGetLastError());
    return ORBFE_OK;
}

}  /* extern "C" */
