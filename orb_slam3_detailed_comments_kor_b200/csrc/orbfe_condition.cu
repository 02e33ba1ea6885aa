/* orbfe_condition.cu — input conditioning in front of the extractor, so that frames uploaded once
 * stay on the GPU until the keypoints come back:
 *   k_remap_linear   cv::remap(im, imRect, M1, M2, cv::INTER_LINEAR) with CV_32FC1 maps, BORDER_CONSTANT(0): the stereo
 *                    rectification of the example drivers (/root/reference/Examples/Stereo-Inertial/stereo_inertial_euroc.cc:201-202,
 *                    maps from cv::initUndistortRectifyMap(..., CV_32F, ...) at :150-151)
 *   k_to_gray4, k_to_gray   cv::cvtColor(im, gray, RGB2GRAY / BGR2GRAY / RGBA2GRAY / BGRA2GRAY) of Tracking::GrabImage*
 *                    (/root/reference/src/Tracking.cc:1303-1328)
 *   k_clahe_lut, k_clahe_apply_band, k_clahe_apply   cv::createCLAHE(3.0, cv::Size(8, 8))->apply(im, im) of the TUM-VI drivers
 *                    (/root/reference/Examples/Stereo-Inertial/stereo_inertial_tum_vi.cc:128-138, Examples/Monocular/mono_tum_vi.cc)
 * Arithmetic (verified bit-exact against cv2 4.13.0, tests/golden/conditioning.npz):
 *   remap: sx = cvRound(mapx*32), ix = sx >> 5, fx = sx & 31 (same for y, ix/iy saturated to int16); weights
 *          (32-fx)(32-fy)*32, fx(32-fy)*32, (32-fx)fy*32, fx*fy*32 (sum 2^15); dst = (sum w*tap + 2^14) >> 15; taps outside
 *          the image read 0.
 *   gray:  (R*9798 + G*19235 + B*3735 + 2^14) >> 15.
 *   clahe: per tile a 256-bin histogram (the image extended right/bottom with REFLECT_101 when its size is not a multiple of
 *          the grid), bins clipped to max(int(clip*area/256), 1), the excess spread as excess/256 to every bin plus one more
 *          to bins 0, s, 2s, ... (s = max(256/(excess%256), 1)), lut[k] = sat(cvRound(float(cumsum[k]) * (255.f/area)));
 *          then each pixel blends the four neighbouring tiles' luts in fp32 without FMA:
 *          (l11*xa1 + l12*xa)*ya1 + (l21*xa1 + l22*xa)*ya, xa = x*(1.f/tw) - .5f - floor(.), rounded with cvRound. */
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

/* one thread = four adjacent pixels: their 12 or 16 source bytes as three words or one uint4, one word out (src 4- / 16-byte aligned,
 * dst 4-byte aligned; the caller sends what is left over, or a misaligned batch, to k_to_gray) */
__device__ __forceinline__ uint32_t gray_of(uint32_t c0, uint32_t c1, uint32_t c2, int rgb) {
    const uint32_t r = rgb ? c0 : c2, b = rgb ? c2 : c0;
    return (r * 9798u + c1 * 19235u + b * 3735u + 16384u) >> 15;
}
__global__ void __launch_bounds__(256) k_to_gray4(const uint32_t* __restrict__ src, uint32_t* __restrict__ dst, long long ngroups, int channels, int rgb) {
    const long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= ngroups) return;
    uint32_t out;
    if (channels == 3) {
        const uint32_t w0 = __ldg(src + 3 * g), w1 = __ldg(src + 3 * g + 1), w2 = __ldg(src + 3 * g + 2);
        out = gray_of(w0 & 255, (w0 >> 8) & 255, (w0 >> 16) & 255, rgb) | gray_of(w0 >> 24, w1 & 255, (w1 >> 8) & 255, rgb) << 8 |
              gray_of((w1 >> 16) & 255, w1 >> 24, w2 & 255, rgb) << 16 | gray_of((w2 >> 8) & 255, (w2 >> 16) & 255, w2 >> 24, rgb) << 24;
    } else {
        const uint4 w = __ldg(reinterpret_cast<const uint4*>(src) + g);
        out = gray_of(w.x & 255, (w.x >> 8) & 255, (w.x >> 16) & 255, rgb) | gray_of(w.y & 255, (w.y >> 8) & 255, (w.y >> 16) & 255, rgb) << 8 |
              gray_of(w.z & 255, (w.z >> 8) & 255, (w.z >> 16) & 255, rgb) << 16 | gray_of(w.w & 255, (w.w >> 8) & 255, (w.w >> 16) & 255, rgb) << 24;
    }
    dst[g] = out;
}

struct orbfe_clahe {
    int device, W, H, tilesX, tilesY, tw, th, clipLimit, maxBatch;
    float lutScale, invTw, invTh;
    uint8_t* dLut;   /* [maxBatch][tilesY][tilesX][256] */
    uint8_t* dImg;   /* staging for the host entry point */
    cudaStream_t stream;
};

/* one CTA = one tile of one frame: histogram (one private copy per warp), clip, redistribute, prefix sum, lut */
__global__ void __launch_bounds__(256) k_clahe_lut(const uint8_t* __restrict__ src, uint8_t* __restrict__ lut, int W, int H, int tilesX, int tw, int th,
                                                   int clipLimit, float lutScale) {
    __shared__ int hist[8][256];
    __shared__ int wsum[8];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int tile = blockIdx.x, txi = tile % tilesX, tyi = tile / tilesX;
    const uint8_t* s = src + (size_t)blockIdx.y * W * H;
#pragma unroll
    for (int k = 0; k < 8; k++) hist[k][tid] = 0;
    __syncthreads();
    for (int r = warp; r < th; r += 8) {
        const int y = tyi * th + r, sy = y < H ? y : 2 * H - 2 - y;
        const uint8_t* row = s + (size_t)sy * W;
        for (int c = lane; c < tw; c += 32) {
            const int x = txi * tw + c;
            atomicAdd(&hist[warp][row[x < W ? x : 2 * W - 2 - x]], 1);
        }
    }
    __syncthreads();
    int h = 0;
#pragma unroll
    for (int k = 0; k < 8; k++) h += hist[k][tid];
    if (clipLimit > 0) {
        int excess = max(h - clipLimit, 0);
        h = min(h, clipLimit);
#pragma unroll
        for (int o = 16; o; o >>= 1) excess += __shfl_xor_sync(0xffffffffu, excess, o);
        if (lane == 0) wsum[warp] = excess;
        __syncthreads();
        int clipped = 0;
#pragma unroll
        for (int k = 0; k < 8; k++) clipped += wsum[k];
        const int redist = clipped >> 8, residual = clipped & 255;
        h += redist;
        if (residual) {
            const int step = max(256 / residual, 1);
            if (tid % step == 0 && tid / step < residual) h++;
        }
        __syncthreads();
    }
    int sum = h; /* inclusive prefix sum over the 256 bins */
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, sum, o);
        if (lane >= o) sum += t;
    }
    if (lane == 31) wsum[warp] = sum;
    __syncthreads();
    for (int k = 0; k < warp; k++) sum += wsum[k];
    const int v = __float2int_rn(__fmul_rn((float)sum, lutScale));
    lut[((size_t)blockIdx.y * gridDim.x + tile) * 256 + tid] = (uint8_t)min(max(v, 0), 255);
}

/* one thread = four adjacent pixels of one row; src == dst is allowed (each pixel depends on itself and the luts only) */
__global__ void __launch_bounds__(256) k_clahe_apply(const uint8_t* __restrict__ src, const uint8_t* __restrict__ lut, uint8_t* __restrict__ dst, int W, int H,
                                                     int tilesX, int tilesY, float invTw, float invTh, int groupsPerRow) {
    const int g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= groupsPerRow * H) return;
    const int y = g / groupsPerRow, x0 = (g - y * groupsPerRow) * 4;
    const size_t frame = (size_t)blockIdx.y * W * H;
    const uint8_t* l = lut + (size_t)blockIdx.y * tilesX * tilesY * 256;
    const float tyf = __fsub_rn(__fmul_rn((float)y, invTh), 0.5f);
    int ty1 = __float2int_rd(tyf);
    const float ya = __fsub_rn(tyf, (float)ty1), ya1 = __fsub_rn(1.0f, ya);
    const int ty2 = min(ty1 + 1, tilesY - 1);
    ty1 = max(ty1, 0);
    const uint8_t* l1 = l + (size_t)ty1 * tilesX * 256;
    const uint8_t* l2 = l + (size_t)ty2 * tilesX * 256;
    const uint8_t* sp = src + frame + (size_t)y * W + x0;
    uint8_t* dp = dst + frame + (size_t)y * W + x0;
    const bool word = x0 + 4 <= W && ((reinterpret_cast<uintptr_t>(sp) | reinterpret_cast<uintptr_t>(dp)) & 3) == 0;
    uint32_t in = 0, out = 0;
    if (word) in = *reinterpret_cast<const uint32_t*>(sp);
    else
        for (int k = 0; k < 4 && x0 + k < W; k++) in |= (uint32_t)sp[k] << (8 * k);
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const int x = x0 + k;
        if (x >= W) break;
        const float txf = __fsub_rn(__fmul_rn((float)x, invTw), 0.5f);
        int tx1 = __float2int_rd(txf);
        const float xa = __fsub_rn(txf, (float)tx1), xa1 = __fsub_rn(1.0f, xa);
        const int tx2 = min(tx1 + 1, tilesX - 1);
        tx1 = max(tx1, 0);
        const int v = (in >> (8 * k)) & 255;
        const float a = (float)l1[tx1 * 256 + v], b = (float)l1[tx2 * 256 + v], c = (float)l2[tx1 * 256 + v], d = (float)l2[tx2 * 256 + v];
        const float top = __fadd_rn(__fmul_rn(a, xa1), __fmul_rn(b, xa)), bot = __fadd_rn(__fmul_rn(c, xa1), __fmul_rn(d, xa));
        const int r = __float2int_rn(__fadd_rn(__fmul_rn(top, ya1), __fmul_rn(bot, ya)));
        out |= (uint32_t)min(max(r, 0), 255) << (8 * k);
    }
    if (word) *reinterpret_cast<uint32_t*>(dp) = out;
    else
        for (int k = 0; k < 4 && x0 + k < W; k++) dp[k] = (uint8_t)(out >> (8 * k));
}

/* The same blend for the usual shapes (width a multiple of 4, word-aligned frames, at most 16 tiles across, tiles at least 32 rows high),
 * arranged so that a pixel costs its arithmetic and little else.  One CTA = 256 columns x 32 rows.  It first copies the tables it can
 * touch into shared memory, one 256-byte slot per (tile row, tile column) with the clamping of the reference's tx1/tx2/ty1/ty2 built
 * into the copy: slot (r, c) holds tile (clamp(ty0 + r), clamp(c - 1)), so that the four tables of a pixel are at a, a + 256, a + STRIDE,
 * a + STRIDE + 256 for one address a — four loads with constant offsets.  A thread owns four adjacent columns (their x weights and table
 * columns are computed once) and walks down eight rows. */
#define CL_BW 64            /* threads across: 64 words = 256 columns */
#define CL_BH 4             /* threads down */
#define CL_ROWS 32          /* rows per CTA */
#define CL_MAXTX 16
#define CL_STRIDE ((CL_MAXTX + 2) * 256)   /* slot columns 0 .. tilesX + 1: raw tx1 = -1 .. tilesX - 1 and the column after it */
#define CL_SLOTROWS 4       /* tile rows a 32-row band can touch (tiles >= 32 rows high: at most three), plus one */
__global__ void __launch_bounds__(CL_BW * CL_BH) k_clahe_apply_band(const uint8_t* __restrict__ src, const uint8_t* __restrict__ lut, uint8_t* __restrict__ dst,
                                                                    int W, int H, int tilesX, int tilesY, float invTw, float invTh) {
    __shared__ __align__(16) uint8_t s_lut[CL_SLOTROWS * CL_STRIDE];
    const int tid = threadIdx.y * CL_BW + threadIdx.x;
    const int yb = blockIdx.y * CL_ROWS, yend = min(yb + CL_ROWS, H);
    const size_t frame = (size_t)blockIdx.z * W * H;
    const uint8_t* l = lut + (size_t)blockIdx.z * tilesX * tilesY * 256;
    const int ty0 = __float2int_rd(__fsub_rn(__fmul_rn((float)yb, invTh), 0.5f));             /* raw ty1 of the band's first row, >= -1 */
    const int nr = __float2int_rd(__fsub_rn(__fmul_rn((float)(yend - 1), invTh), 0.5f)) + 1 - ty0 + 1;   /* slot rows in use */
    for (int i = tid; i < nr * (tilesX + 2) * 16; i += CL_BW * CL_BH) {
        const int q = i & 15, sc = (i >> 4) % (tilesX + 2), sr = (i >> 4) / (tilesX + 2);
        const int ty = min(max(ty0 + sr, 0), tilesY - 1), tx = min(max(sc - 1, 0), tilesX - 1);
        reinterpret_cast<uint4*>(s_lut + sr * CL_STRIDE + sc * 256)[q] = __ldg(reinterpret_cast<const uint4*>(l + ((size_t)ty * tilesX + tx) * 256) + q);
    }
    const int x0 = (blockIdx.x * CL_BW + threadIdx.x) * 4;
    int col[4];
    float xa[4], xa1[4];
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const float txf = __fsub_rn(__fmul_rn((float)(x0 + k), invTw), 0.5f);
        const int tx1 = __float2int_rd(txf);
        xa[k] = __fsub_rn(txf, (float)tx1);
        xa1[k] = __fsub_rn(1.0f, xa[k]);
        col[k] = (min(tx1, tilesX - 1) + 1) * 256;      /* columns beyond the image are never used; keep the address inside the row of slots */
    }
    __syncthreads();
    if (x0 >= W) return;
    for (int y = yb + threadIdx.y; y < yend; y += CL_BH) {
        const float tyf = __fsub_rn(__fmul_rn((float)y, invTh), 0.5f);
        const int ty1 = __float2int_rd(tyf);
        const float ya = __fsub_rn(tyf, (float)ty1), ya1 = __fsub_rn(1.0f, ya);
        const uint8_t* row = s_lut + (ty1 - ty0) * CL_STRIDE;
        const uint32_t in = *reinterpret_cast<const uint32_t*>(src + frame + (size_t)y * W + x0);
        uint32_t out = 0;
#pragma unroll
        for (int k = 0; k < 4; k++) {
            const uint8_t* t = row + col[k] + ((in >> (8 * k)) & 255);
            const float a = (float)t[0], b = (float)t[256], c = (float)t[CL_STRIDE], d = (float)t[CL_STRIDE + 256];
            const float top = __fadd_rn(__fmul_rn(a, xa1[k]), __fmul_rn(b, xa[k])), bot = __fadd_rn(__fmul_rn(c, xa1[k]), __fmul_rn(d, xa[k]));
            const int r = __float2int_rn(__fadd_rn(__fmul_rn(top, ya1), __fmul_rn(bot, ya)));
            out |= (uint32_t)min(max(r, 0), 255) << (8 * k);
        }
        *reinterpret_cast<uint32_t*>(dst + frame + (size_t)y * W + x0) = out;
    }
}

extern "C" {

void orbfe_clahe_destroy(orbfe_clahe* c) {
    if (!c) return;
    cudaSetDevice(c->device);
    if (c->stream) cudaStreamSynchronize(c->stream);
    cudaFree(c->dLut); cudaFree(c->dImg);
    if (c->stream) cudaStreamDestroy(c->stream);
    delete c;
}

int orbfe_clahe_create(int device, int width, int height, double clip_limit, int tiles_x, int tiles_y, int max_batch, orbfe_clahe** out) {
    if (!out || width < 1 || height < 1 || width > 16384 || height > 16384 || tiles_x < 1 || tiles_y < 1 || tiles_x > 64 || tiles_y > 64 ||
        max_batch < 1 || max_batch > 65535) { orbfe_set_error("bad argument"); return ORBFE_E_INVALID; }
    *out = nullptr;
    const bool whole = width % tiles_x == 0 && height % tiles_y == 0;
    const int ew = whole ? width : width + tiles_x - width % tiles_x, eh = whole ? height : height + tiles_y - height % tiles_y;
    if (ew - width >= width || eh - height >= height) { orbfe_set_error("image %dx%d too small for a %dx%d tile grid", width, height, tiles_x, tiles_y); return ORBFE_E_INVALID; }
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || device < 0 || device >= ndev) {
        orbfe_set_error("no usable CUDA device %d (found %d); this library has no CPU path", device, ndev);
        return ORBFE_E_CUDA;
    }
    ORBFE_CUDA(cudaSetDevice(device));
    orbfe_clahe* c = new (std::nothrow) orbfe_clahe();
    if (!c) return ORBFE_E_INVALID;
    c->device = device; c->W = width; c->H = height; c->tilesX = tiles_x; c->tilesY = tiles_y; c->maxBatch = max_batch;
    c->tw = ew / tiles_x; c->th = eh / tiles_y;
    const int area = c->tw * c->th;
    c->clipLimit = 0;
    if (clip_limit > 0.0) { c->clipLimit = (int)(clip_limit * area / 256); if (c->clipLimit < 1) c->clipLimit = 1; }
    c->lutScale = 255.0f / (float)area; c->invTw = 1.0f / (float)c->tw; c->invTh = 1.0f / (float)c->th;
    c->dLut = nullptr; c->dImg = nullptr; c->stream = nullptr;
    cudaError_t e = cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaMalloc((void**)&c->dLut, (size_t)max_batch * tiles_x * tiles_y * 256);
    if (e == cudaSuccess) e = cudaMalloc((void**)&c->dImg, (size_t)width * height);
    if (e != cudaSuccess) { const int rc = orbfe_cuda_fail(e, "clahe allocation"); orbfe_clahe_destroy(c); return rc; }
    *out = c;
    return ORBFE_OK;
}

int orbfe_clahe_apply_batch_device(orbfe_clahe* c, const uint8_t* d_src, uint8_t* d_dst, int nframes, void* stream) {
    if (!c || !d_src || !d_dst || nframes < 1) { orbfe_set_error("bad argument"); return ORBFE_E_INVALID; }
    if (nframes > c->maxBatch) { orbfe_set_error("batch of %d frames exceeds max_batch %d", nframes, c->maxBatch); return ORBFE_E_CAPACITY; }
    ORBFE_CUDA(cudaSetDevice(c->device));
    cudaStream_t st = stream ? (cudaStream_t)stream : c->stream;
    k_clahe_lut<<<dim3(c->tilesX * c->tilesY, nframes), 256, 0, st>>>(d_src, c->dLut, c->W, c->H, c->tilesX, c->tw, c->th, c->clipLimit, c->lutScale);
    const int gpr = (c->W + 3) / 4;
    if ((c->W & 3) == 0 && (((uintptr_t)d_src | (uintptr_t)d_dst) & 3) == 0 && c->tilesX <= CL_MAXTX && c->th >= CL_ROWS && nframes <= 65535)
        k_clahe_apply_band<<<dim3((gpr + CL_BW - 1) / CL_BW, (c->H + CL_ROWS - 1) / CL_ROWS, nframes), dim3(CL_BW, CL_BH), 0, st>>>(
            d_src, c->dLut, d_dst, c->W, c->H, c->tilesX, c->tilesY, c->invTw, c->invTh);
    else
        k_clahe_apply<<<dim3((gpr * c->H + 255) / 256, nframes), 256, 0, st>>>(d_src, c->dLut, d_dst, c->W, c->H, c->tilesX, c->tilesY, c->invTw, c->invTh, gpr);
    ORBFE_CUDA(cudaGetLastError());
    return ORBFE_OK;
}

int orbfe_clahe_apply(orbfe_clahe* c, const uint8_t* src, size_t sstep, uint8_t* dst, size_t dstep) {
    if (!c || !src || !dst || sstep < (size_t)c->W || dstep < (size_t)c->W) { orbfe_set_error("bad argument"); return ORBFE_E_INVALID; }
    ORBFE_CUDA(cudaSetDevice(c->device));
    ORBFE_CUDA(cudaMemcpy2DAsync(c->dImg, c->W, src, sstep, c->W, c->H, cudaMemcpyHostToDevice, c->stream));
    int rc = orbfe_clahe_apply_batch_device(c, c->dImg, c->dImg, 1, c->stream);
    if (rc) return rc;
    ORBFE_CUDA(cudaMemcpy2DAsync(dst, dstep, c->dImg, c->W, c->W, c->H, cudaMemcpyDeviceToHost, c->stream));
    ORBFE_CUDA(cudaStreamSynchronize(c->stream));
    return ORBFE_OK;
}

int orbfe_clahe_lut(orbfe_clahe* c, int frame, uint8_t* out) {
    if (!c || !out || frame < 0 || frame >= c->maxBatch) { orbfe_set_error("bad argument"); return ORBFE_E_INVALID; }
    ORBFE_CUDA(cudaSetDevice(c->device));
    /* tables of the last apply call: on the object's stream unless the caller brought its own (then the caller has synchronised it) */
    ORBFE_CUDA(cudaMemcpyAsync(out, c->dLut + (size_t)frame * c->tilesX * c->tilesY * 256, (size_t)c->tilesX * c->tilesY * 256, cudaMemcpyDeviceToHost, c->stream));
    ORBFE_CUDA(cudaStreamSynchronize(c->stream));
    return ORBFE_OK;
}

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
    long long done = 0;
    if (((uintptr_t)d_src & (channels == 4 ? 15 : 3)) == 0 && ((uintptr_t)d_dst & 3) == 0 && npix >= 4) {
        const long long ngroups = npix / 4;
        k_to_gray4<<<(unsigned)((ngroups + 255) / 256), 256, 0, (cudaStream_t)stream>>>((const uint32_t*)d_src, (uint32_t*)d_dst, ngroups, channels, rgb_order ? 1 : 0);
        ORBFE_CUDA(cudaGetLastError());
        done = ngroups * 4;
    }
    if (done < npix)
        k_to_gray<<<(unsigned)((npix - done + 255) / 256), 256, 0, (cudaStream_t)stream>>>(d_src + done * channels, d_dst + done, npix - done, channels, rgb_order ? 1 : 0);
    ORBFE_CUDA(cudaGetLastError());
    return ORBFE_OK;
}

}  /* extern "C" */
