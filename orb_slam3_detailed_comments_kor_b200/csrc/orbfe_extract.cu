/* orbfe_extract.cu — sm_100a kernels of the ORB extractor (replaces ORBextractor::operator(),
 * /root/reference/src/ORBextractor.cc:1068-1150, and everything it calls).
 *
 * Stage            kernel              reference lines (src/ORBextractor.cc)       roofline
 *  pyramid         k_pyramid_level     1152-1177 (+ cv::resize, copyMakeBorder)     HBM
 *  FAST per cell   k_fast_stream, k_fast_nms, k_fast_cells_list
 *                                      763-854   (+ cv::FAST 9/16, NMS, 20/7 retry) integer ALU
 *  quadtree        k_quadtree          479-761                                      latency (L2 resident)
 *  ordering        k_assemble          856-872, 1100-1149                           latency
 *  blur            k_blur7             1114-1115 (+ cv::GaussianBlur 7x7 s=2)       HBM
 *  angle + rBRIEF  k_orient_describe   75-145, 470-477, 1059-1066                   L2 gathers
 *
 * All frames of a batch go through one launch per stage (gridDim.y / .z = frame); frames never interact.
 * Integer stages are bit-exact by construction; float stages use round-to-nearest intrinsics so that no FMA
 * contraction can change a bit relative to the oracle (built with -ffp-contract=off). */
#include <math.h>
#include <string.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include <mutex>

#include "orbfe_internal.h"
#include "orbfe_tma.cuh"
#include "orbfe_trig.cuh"

__constant__ int c_umax[16];
__device__ __align__(16) float g_patternf[1024];   /* bit_pattern_31_ as floats: (x0,y0,x1,y1) per test */
/* IC_Angle weights for k_orient_describe: [s = patch alignment 0..3][step 0..8][lane] -> (u weights, v weights) as four signed bytes
 * each, zero outside the radius-15 disc.  Steps 0..7: lane = (row & 3) * 8 + word, row = 4*step + (lane >> 3); step 8: lane = row,
 * ninth word of the row.  Byte t of a row's 36-byte window is patch column u = t - s - 15. */
__device__ __align__(16) uint2 g_icw[4 * 9 * 32];

static __device__ __forceinline__ int reflect101(int p, int n) {
    /* |overshoot| <= 19 < n: one reflection is enough (cv::copyMakeBorder BORDER_REFLECT_101) */
    p = p < 0 ? -p : p;
    return p >= n ? 2 * (n - 1) - p : p;
}

/* ------------------------------------------------------------------------------------------------ pyramid */
/* Level 0: the interior of copyMakeBorder(image, 19, REFLECT_101) (src/ORBextractor.cc:1170-1173).  One thread = one
 * 16-byte chunk of an image row: a 128-bit load/store when the chunk is whole and the source is 16-byte aligned, byte by
 * byte otherwise (odd widths, unaligned inputs).  The border comes from k_pyramid_fringe / k_pyramid_borders. */
__global__ void __launch_bounds__(256) k_pyramid_copy0(const __grid_constant__ ExDesc d, ExBuffers b, const uint8_t* __restrict__ input) {
    const LevelDesc& L = d.lv[0];
    const int chunks = (L.w + 15) >> 4;
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= chunks * L.h) return;
    const int row = idx / chunks, px0 = (idx - row * chunks) * 16;
    const size_t f = blockIdx.y;
    const uint8_t* src = input + f * (size_t)d.W * d.H + (size_t)row * d.W + px0;
    uint8_t* dst = b.pyr + f * d.pyrSlab + L.pyrOff + (size_t)(row + ORBFE_EDGE) * L.pitch + ORBFE_ROW_LEAD + px0;
    if (px0 + 16 <= L.w && ((reinterpret_cast<uintptr_t>(src) & 15) == 0)) {
        *reinterpret_cast<uint4*>(dst) = __ldg(reinterpret_cast<const uint4*>(src));
        return;
    }
    if (px0 + 16 <= L.w) {
        /* odd widths (KITTI: 1241) leave most rows unaligned: five aligned words around the chunk, funnel-shifted into place,
         * as long as those words lie inside the batch */
        const uintptr_t a = reinterpret_cast<uintptr_t>(src);
        const uint32_t* q = reinterpret_cast<const uint32_t*>(a & ~(uintptr_t)3);
        const uint8_t* end = input + (size_t)gridDim.y * d.W * d.H;
        if (reinterpret_cast<const uint8_t*>(q) >= input && reinterpret_cast<const uint8_t*>(q + 5) <= end) {
            const uint32_t sh = (uint32_t)(a & 3) * 8;
            const uint32_t w0 = __ldg(q), w1 = __ldg(q + 1), w2 = __ldg(q + 2), w3 = __ldg(q + 3), w4 = __ldg(q + 4);
            *reinterpret_cast<uint4*>(dst) = make_uint4(__funnelshift_r(w0, w1, sh), __funnelshift_r(w1, w2, sh), __funnelshift_r(w2, w3, sh), __funnelshift_r(w3, w4, sh));
            return;
        }
    }
    for (int k = 0; k < 16 && px0 + k < L.w; k++) dst[k] = __ldg(src + k);
}

/* Levels >= 1: cv::resize(INTER_LINEAR) from the level below + copyMakeBorder in one pass
 * (src/ORBextractor.cc:1163-1168).  One thread = four adjacent output pixels (one 32-bit store, lanes adjacent, so
 * a warp writes 128 contiguous bytes).  Interior groups fetch each source row as three aligned words, funnel-shift
 * them to the first tap, and form every horizontal tap pair with one PRMT + one DP2A
 * (c0*S[sx] + c1*S[sx+1], 16-bit coefficients x 8-bit pixels); groups that touch the border recompute the reflected
 * interior pixel byte-wise.  Arithmetic is OpenCV's 11-bit fixed point. */

#define PYR_WARPS 2   /* warps per CTA (they are independent: small CTAs free their slots sooner) */
#define PYR_ROWS 4   /* output rows per thread: the horizontal tables, selectors and shifts are shared between them */
/* Geometry of one resize launch, passed by value so that every field sits at a fixed constant-bank address. */
struct ResizeArgs {
    int w, h, pitch, pyrOff;      /* destination level */
    int spitch, spyrOff;          /* source level */
    int tabX, tabY;
};

/* vertical blend of OpenCV's fixed-point bilinear: (((b0*(h0>>4))>>16) + ((b1*(h1>>4))>>16) + 2) >> 2, with the
 * >>16 taken as the high half of a product by b<<16 */
static __device__ __forceinline__ uint32_t resize_vblend(uint32_t h0, uint32_t h1, uint32_t b0s, uint32_t b1s) {
    return (__umulhi(h0 >> 4, b0s) + __umulhi(h1 >> 4, b1s) + 2u) >> 2;
}

__global__ void __launch_bounds__(256) k_pyramid_level(const __grid_constant__ ExDesc d, ExBuffers b, const __grid_constant__ ResizeArgs a) {
    /* block = 32 x PYR_WARPS: a warp works on PYR_ROWS consecutive interior rows; a thread on one group of four pixels that
     * lies fully inside the image.  The border pixels the extractor reads (and the last 1-3 columns when w % 4 != 0) are filled
     * afterwards for all levels at once by k_pyramid_fringe: the next level's resize never reads a border pixel with a non-zero weight. */
    const int row0 = (blockIdx.y * PYR_WARPS + threadIdx.y) * PYR_ROWS;   /* interior row */
    if (row0 >= a.h) return;
    const int tx = blockIdx.x * 32 + threadIdx.x;
    if (tx >= ((a.w + 3) >> 2)) return;   /* a partial last group (w % 4 != 0) is computed like a full one, see the x table */
    uint8_t* slab = b.pyr + (size_t)blockIdx.z * d.pyrSlab;
    const uint8_t* sint = slab + (uint32_t)(a.spyrOff + ORBFE_EDGE * a.spitch + ORBFE_ROW_LEAD);
    const int px0 = tx * 4;
    uint8_t* dstp = slab + (uint32_t)(a.pyrOff + (row0 + ORBFE_EDGE) * a.pitch + ORBFE_ROW_LEAD + px0);
    const uint2* tabY = b.tabs + a.tabY;
    const uint4* tp = reinterpret_cast<const uint4*>(b.tabs + a.tabX + px0); /* tabX is 4-entry aligned */
    const uint4 ta = __ldg(tp), tb = __ldg(tp + 1); /* entries (sx, c0|c1<<16) of the four pixels */
    uint2 yt[PYR_ROWS];
#pragma unroll
    for (int rr = 0; rr < PYR_ROWS; rr++) yt[rr] = __ldg(tabY + min(row0 + rr, a.h - 1));
    const uint32_t sx0 = ta.x;
    const uint32_t base = sx0 & ~3u, sh = (sx0 & 3u) * 8;
    const uint32_t cfs[4] = {ta.y, ta.w, tb.y, tb.w};
    /* byte selectors (off, off+1) -> low half; off = sx - sx0 in 0..6 (host-checked): four taps within 8 bytes */
    const uint32_t sel[4] = {0x10u, (ta.z - sx0) * 0x11u + 0x10u, (tb.x - sx0) * 0x11u + 0x10u, (tb.z - sx0) * 0x11u + 0x10u};
    uint32_t wsrc[PYR_ROWS][6];   /* all source words of all rows are requested before the arithmetic starts */
#pragma unroll
    for (int rr = 0; rr < PYR_ROWS; rr++) {
        /* row yt.x+1 and column sx+1 may be the source's own border: their coefficient is then 0 */
        const uint32_t* q0 = reinterpret_cast<const uint32_t*>(sint + (uint32_t)((int)yt[rr].x * a.spitch) + base);
        const uint32_t* q1 = reinterpret_cast<const uint32_t*>(reinterpret_cast<const uint8_t*>(q0) + a.spitch);
        wsrc[rr][0] = q0[0]; wsrc[rr][1] = q0[1]; wsrc[rr][2] = q0[2];
        wsrc[rr][3] = q1[0]; wsrc[rr][4] = q1[1]; wsrc[rr][5] = q1[2];
    }
    asm volatile("" ::: "memory");
#pragma unroll
    for (int rr = 0; rr < PYR_ROWS; rr++) {
        if (row0 + rr >= a.h) break;
        const uint32_t b0s = (yt[rr].y & 0xFFFFu) << 16, b1s = yt[rr].y & 0xFFFF0000u;
        const uint32_t u0 = __funnelshift_r(wsrc[rr][0], wsrc[rr][1], sh), u1 = __funnelshift_r(wsrc[rr][1], wsrc[rr][2], sh); /* bytes sx0 .. sx0+7 */
        const uint32_t v0 = __funnelshift_r(wsrc[rr][3], wsrc[rr][4], sh), v1 = __funnelshift_r(wsrc[rr][4], wsrc[rr][5], sh);
        uint32_t o[4];
#pragma unroll
        for (int k = 0; k < 4; k++)
            o[k] = resize_vblend(__dp2a_lo(cfs[k], __byte_perm(u0, u1, sel[k]), 0u), __dp2a_lo(cfs[k], __byte_perm(v0, v1, sel[k]), 0u), b0s, b1s);
        *reinterpret_cast<uint32_t*>(dstp + rr * a.pitch) = __byte_perm(__byte_perm(o[0], o[1], 0x0040), __byte_perm(o[2], o[3], 0x0040), 0x5410);
    }
}

/* The part of the REFLECT_101 border that the extractor itself reads: the Gaussian blur reaches 3 px past the image edge,
 * nothing else leaves the interior (FAST stays 16 px inside, the orientation patch and the descriptor taps stay inside by
 * construction, the next level's resize gives border pixels weight 0).  Per level: the word left of every interior row, the
 * word(s) holding px w .. w+2 right of it, and three rows above and below over that same span.  One thread = one 4-byte
 * group; all levels in one launch (gridDim.y = level).  The full 19 px border of the public pyramid
 * (ORBextractor::mvImagePyramid lives in it) is written by k_pyramid_borders when somebody asks for it. */
#define FRINGE 3
__global__ void __launch_bounds__(128) k_pyramid_fringe(const __grid_constant__ ExDesc d, ExBuffers b) {
    const int level = blockIdx.y;
    const LevelDesc& L = d.lv[level];
    const int gRight0 = L.w >> 2;                                 /* first group with a pixel >= w (or the partial one) */
    const int gRight1 = (L.w + FRINGE - 1) >> 2;                  /* last group with a pixel <= w+2 */
    const int nRight = gRight1 - gRight0 + 1, side = 1 + nRight;  /* fringe groups of an interior row */
    const int span = gRight1 + 2;                                 /* groups -1 .. gRight1 of a band row */
    int idx = blockIdx.x * blockDim.x + threadIdx.x;
    int row, g;                                                   /* interior row coordinate, group index (px = 4g) */
    if (idx < L.h * side) {
        row = idx / side;
        const int j = idx - row * side;
        g = j == 0 ? -1 : gRight0 + j - 1;
    } else {
        idx -= L.h * side;
        if (idx >= 2 * FRINGE * span) return;
        const int r = idx / span;
        g = idx - r * span - 1;
        row = r < FRINGE ? r - FRINGE : L.h + r - FRINGE;
    }
    uint8_t* slab = b.pyr + (size_t)blockIdx.z * d.pyrSlab;
    uint8_t* lint = slab + (uint32_t)(L.pyrOff + ORBFE_EDGE * L.pitch + ORBFE_ROW_LEAD);   /* this level, interior (0,0) */
    const uint8_t* srow = lint + reflect101(row, L.h) * L.pitch;
    const int px0 = g * 4;
    uint32_t out;
    if (px0 >= 0 && px0 + 4 <= L.w) {
        out = *reinterpret_cast<const uint32_t*>(srow + px0);
    } else {
        out = 0;
#pragma unroll
        for (int k = 0; k < 4; k++) out |= (uint32_t)srow[reflect101(px0 + k, L.w)] << (8 * k);   /* |overshoot| <= 4 */
    }
    *reinterpret_cast<uint32_t*>(lint + row * L.pitch + px0) = out;
}

/* The whole REFLECT_101 border (copyMakeBorder, src/ORBextractor.cc:1167-1168, 1172-1173), all levels in one launch
 * (gridDim.y = level).  One thread = one 4-byte group that is not a full interior group.  A border pixel is a copy
 * of the interior pixel it mirrors (all interior pixels are final when this kernel starts).  Not on the extraction path:
 * launched by orbfe_materialize_borders for the callers that read the padded pyramid. */
__global__ void __launch_bounds__(256) k_pyramid_borders(const __grid_constant__ ExDesc d, ExBuffers b) {
    const int level = blockIdx.y;
    const LevelDesc& L = d.lv[level];
    const int G = L.pitch >> 2, nInt = L.w >> 2, gFirst = ORBFE_ROW_LEAD / 4;
    const int side = G - nInt;                                  /* border groups of an interior row */
    const int nband = ORBFE_EDGE * G;                           /* groups of the top (or bottom) band */
    int idx = blockIdx.x * blockDim.x + threadIdx.x;
    int row, g;
    if (idx < nband) { row = idx / G; g = idx - row * G; }
    else if (idx < 2 * nband) { idx -= nband; row = idx / G; g = idx - row * G; row += ORBFE_EDGE + L.h; }
    else {
        idx -= 2 * nband;
        if (idx >= L.h * side) return;
        row = idx / side;
        const int j = idx - row * side;
        g = j < gFirst ? j : nInt + j;
        row += ORBFE_EDGE;
    }
    uint8_t* slab = b.pyr + (size_t)blockIdx.z * d.pyrSlab;
    const uint8_t* lint = slab + (uint32_t)(L.pyrOff + ORBFE_EDGE * L.pitch + ORBFE_ROW_LEAD);   /* this level, interior (0,0) */
    const int yi = reflect101(row - ORBFE_EDGE, L.h);
    const int px0 = g * 4 - ORBFE_ROW_LEAD;
    uint32_t out = 0;
    const uint8_t* srow = lint + yi * L.pitch;
    if (px0 >= 0 && px0 + 4 <= L.w) {
        out = *reinterpret_cast<const uint32_t*>(srow + px0);                   /* top / bottom band above interior columns */
    } else {
#pragma unroll
        for (int k = 0; k < 4; k++) {
            const int px = px0 + k;
            if (px < -ORBFE_EDGE || px >= L.w + ORBFE_EDGE) continue;
            out |= (uint32_t)srow[reflect101(px, L.w)] << (8 * k);              /* interior bytes of a partial group copy themselves */
        }
    }
    *reinterpret_cast<uint32_t*>(slab + (uint32_t)(L.pyrOff + row * L.pitch + g * 4)) = out;
}

/* --------------------------------------------------------------------------------------------------- blur */
/* 7x7 sigma-2 Gaussian as OpenCV's 8U fixed point: taps {18,34,48,56,48,34,18}/256 horizontally (Q8.8, exact in
 * 16 bits), the same vertically (Q16.16), + 2^15 >> 16.  The padded pyramid already holds the REFLECT_101 border,
 * so no edge logic.  Horizontal pass: the 7 taps of a pixel are two DP4A on funnel-shifted source words.  Its results
 * are stored as vertically adjacent 16-bit pairs (rows 2j, 2j+1 of the tile in one 32-bit word), so the vertical
 * pass is four DP2A (16-bit x 8-bit pairs) per pixel and one thread produces a 4 x 2 block from five 128-bit loads. */
#define BLUR_TW 64
#define BLUR_TH 56
#define BLUR_PAIRS (BLUR_TH / 2 + 4)   /* tile rows y0-4 .. y0+TH+3 as pairs */
__global__ void __launch_bounds__(256) k_blur7(const __grid_constant__ ExDesc d, ExBuffers b) {
    __shared__ __align__(16) uint32_t s_h[BLUR_PAIRS][BLUR_TW];
    /* one launch covers every level: blockIdx.x walks the 64x56 tiles of level 0, then level 1, ... (d.blurTile0[l] =
     * first tile of level l) */
    int level = 0;
    while (level + 1 < d.nlevels && (int)blockIdx.x >= d.blurTile0[level + 1]) level++;
    const LevelDesc& L = d.lv[level];
    const size_t f = blockIdx.y;
    const int tilesX = (L.w + BLUR_TW - 1) / BLUR_TW, tile = blockIdx.x - d.blurTile0[level];
    const int ty = tile / tilesX, tx = tile - ty * tilesX;
    const int x0 = tx * BLUR_TW, y0 = ty * BLUR_TH;
    const uint8_t* pad = b.pyr + f * d.pyrSlab + L.pyrOff + ORBFE_ROW_LEAD + x0 - 4;
    const uint32_t KLO = 18u | (34u << 8) | (48u << 16) | (56u << 24);  /* taps -3..0 */
    const uint32_t KHI = 48u | (34u << 8) | (18u << 16);                /* taps +1..+3 */
    const int ymax = L.h + ORBFE_EDGE - 1;
    static_assert(BLUR_PAIRS * (BLUR_TW / 4) == 512 && BLUR_TW == 64, "horizontal pass assumes two whole sweeps of 256 threads");
    {
        const int q = threadIdx.x & 15;
        uint32_t w[2][2][3];   /* [sweep][row of the pair][word]: all twelve loads are issued before any is used */
#pragma unroll
        for (int sweep = 0; sweep < 2; sweep++)
#pragma unroll
            for (int rr = 0; rr < 2; rr++) {
                const int j = (threadIdx.x >> 4) + 16 * sweep;
                const int gy = min(y0 - 4 + 2 * j + rr, ymax);
                const uint32_t* p = reinterpret_cast<const uint32_t*>(pad + (uint32_t)((gy + ORBFE_EDGE) * L.pitch + 4 * q));
                w[sweep][rr][0] = __ldg(p); w[sweep][rr][1] = __ldg(p + 1); w[sweep][rr][2] = __ldg(p + 2);   /* px x-4..x-1 | x..x+3 | x+4..x+7 */
            }
        asm volatile("" ::: "memory");
#pragma unroll
        for (int sweep = 0; sweep < 2; sweep++) {
            const int j = (threadIdx.x >> 4) + 16 * sweep;   /* row pair j, pixel group q (4 px) */
            uint32_t hrow[2][4];
#pragma unroll
            for (int rr = 0; rr < 2; rr++) {
                const uint32_t w0 = w[sweep][rr][0], w1 = w[sweep][rr][1], w2 = w[sweep][rr][2];
                hrow[rr][0] = __dp4a(__funnelshift_r(w0, w1, 8), KLO, __dp4a(__funnelshift_r(w1, w2, 8), KHI, 0u));
                hrow[rr][1] = __dp4a(__funnelshift_r(w0, w1, 16), KLO, __dp4a(__funnelshift_r(w1, w2, 16), KHI, 0u));
                hrow[rr][2] = __dp4a(__funnelshift_r(w0, w1, 24), KLO, __dp4a(__funnelshift_r(w1, w2, 24), KHI, 0u));
                hrow[rr][3] = __dp4a(w1, KLO, __dp4a(w2, KHI, 0u));
            }
            *reinterpret_cast<uint4*>(&s_h[j][4 * q]) = make_uint4(hrow[0][0] | (hrow[1][0] << 16), hrow[0][1] | (hrow[1][1] << 16),
                                                                   hrow[0][2] | (hrow[1][2] << 16), hrow[0][3] | (hrow[1][3] << 16));
        }
    }
    __syncthreads();
    uint8_t* dst = b.blur + f * d.blurSlab + L.blurOff;
    /* one item = output rows (y0+2m, y0+2m+1) x 4 pixels; needs pairs m .. m+4 (rows y0+2m-4 .. y0+2m+5) */
#pragma unroll
    for (int sweep = 0; sweep < (BLUR_TH / 2 + 15) / 16; sweep++) {
    const int m = (threadIdx.x >> 4) + 16 * sweep, q = threadIdx.x & 15;
    const int gy = y0 + 2 * m, gx = x0 + 4 * q;
    if (m < BLUR_TH / 2 && gy < L.h && gx < L.bpitch) {
        uint4 pr[5];
#pragma unroll
        for (int k = 0; k < 5; k++) pr[k] = *reinterpret_cast<const uint4*>(&s_h[m + k][4 * q]);
        /* even row y: rows y-3..y+3 = (pair0.hi) (pair1) (pair2) (pair3); odd row y+1: (pair1) (pair2) (pair3) (pair4.lo) */
        const uint32_t E0 = 18u << 8, E1 = 34u | (48u << 8), E2 = 56u | (48u << 8), E3 = 34u | (18u << 8);
        const uint32_t O1 = 18u | (34u << 8), O2 = 48u | (56u << 8), O3 = 48u | (34u << 8), O4 = 18u;
        uint32_t oe = 0, oo = 0;
        const uint32_t* P0 = &pr[0].x; const uint32_t* P1 = &pr[1].x; const uint32_t* P2 = &pr[2].x;
        const uint32_t* P3 = &pr[3].x; const uint32_t* P4 = &pr[4].x;
#pragma unroll
        for (int k = 0; k < 4; k++) {
            const uint32_t ve = __dp2a_lo(P0[k], E0, __dp2a_lo(P1[k], E1, __dp2a_lo(P2[k], E2, __dp2a_lo(P3[k], E3, 32768u))));
            const uint32_t vo = __dp2a_lo(P1[k], O1, __dp2a_lo(P2[k], O2, __dp2a_lo(P3[k], O3, __dp2a_lo(P4[k], O4, 32768u))));
            oe |= (ve >> 16) << (8 * k);
            oo |= (vo >> 16) << (8 * k);
        }
        uint8_t* o = dst + (uint32_t)(gy * L.bpitch + gx);
        *reinterpret_cast<uint32_t*>(o) = oe;
        if (gy + 1 < L.h) *reinterpret_cast<uint32_t*>(o + L.bpitch) = oo;
    }
    }
}

/* The same blur with the tile brought in by TMA: one thread issues a 3-D tensor copy (x, padded row, frame) of the
 * 96 x 64 byte box around the tile (box rows beyond the padded level are zero-filled by the hardware; they only feed
 * outputs below the image, which are not stored), everybody waits on the mbarrier, and both passes run from shared
 * memory.  No per-thread address arithmetic, clamping or global loads remain; the tile (level, tx, ty) comes from a host
 * table instead of a search over the levels and a division.  Tensor maps: one per level over this extractor's own pyramid
 * slab (orbfe_host.cu), never over memory of another GPU, no multicast. */
#define BLUR_SROWS (2 * BLUR_PAIRS)   /* 64 tile rows: y0-4 .. y0+59 */
#define BLUR_SPITCH 96                /* bytes per tile row: px x0-16 .. x0+79 (the box must start on a 16-byte boundary of the row and be a multiple of
                                         16 bytes wide; px x0-4 .. x0+67 are used) */
__global__ void __launch_bounds__(256) k_blur7_tma(const __grid_constant__ ExDesc d, ExBuffers b, const CUtensorMap* __restrict__ maps, int frame0) {
    __shared__ __align__(128) uint8_t s_tile[BLUR_SROWS * BLUR_SPITCH];
    __shared__ __align__(16) uint32_t s_h[BLUR_PAIRS][BLUR_TW];
    __shared__ __align__(8) uint64_t s_bar;
    const uint32_t t = __ldg(b.blurTiles + blockIdx.x);
    const int level = t >> 24, y0 = ((t >> 12) & 0xFFF) * BLUR_TH, x0 = (t & 0xFFF) * BLUR_TW;
    const LevelDesc& L = d.lv[level];
    if (threadIdx.x == 0) mbar_init(&s_bar, 1);
    __syncthreads();
    if (threadIdx.x == 0) {
        mbar_expect_tx(&s_bar, BLUR_SROWS * BLUR_SPITCH);
        tma_load_3d(s_tile, maps + level, &s_bar, ORBFE_ROW_LEAD + x0 - 16, y0 - 4 + ORBFE_EDGE, frame0 + (int)blockIdx.y);
    }
    const uint32_t KLO = 18u | (34u << 8) | (48u << 16) | (56u << 24);  /* taps -3..0 */
    const uint32_t KHI = 48u | (34u << 8) | (18u << 16);                /* taps +1..+3 */
    const int q = threadIdx.x & 15;
    const int jbase = threadIdx.x >> 4;   /* the half-warps of a warp read rows two apart: 48 words = bank offset 16, conflict-free */
    mbar_wait(&s_bar, 0);
#pragma unroll
    for (int sweep = 0; sweep < 2; sweep++) {
        const int j = jbase + 16 * sweep;   /* row pair j, pixel group q (4 px) */
        uint32_t hrow[2][4];
#pragma unroll
        for (int rr = 0; rr < 2; rr++) {
            const uint32_t* p = reinterpret_cast<const uint32_t*>(s_tile + (2 * j + rr) * BLUR_SPITCH + 12) + q;
            const uint32_t w0 = p[0], w1 = p[1], w2 = p[2];   /* px x-4..x-1 | x..x+3 | x+4..x+7 */
            hrow[rr][0] = __dp4a(__funnelshift_r(w0, w1, 8), KLO, __dp4a(__funnelshift_r(w1, w2, 8), KHI, 0u));
            hrow[rr][1] = __dp4a(__funnelshift_r(w0, w1, 16), KLO, __dp4a(__funnelshift_r(w1, w2, 16), KHI, 0u));
            hrow[rr][2] = __dp4a(__funnelshift_r(w0, w1, 24), KLO, __dp4a(__funnelshift_r(w1, w2, 24), KHI, 0u));
            hrow[rr][3] = __dp4a(w1, KLO, __dp4a(w2, KHI, 0u));
        }
        *reinterpret_cast<uint4*>(&s_h[j][4 * q]) = make_uint4(__byte_perm(hrow[0][0], hrow[1][0], 0x5410), __byte_perm(hrow[0][1], hrow[1][1], 0x5410),
                                                               __byte_perm(hrow[0][2], hrow[1][2], 0x5410), __byte_perm(hrow[0][3], hrow[1][3], 0x5410));
    }
    __syncthreads();
    uint8_t* dst = b.blur + (size_t)blockIdx.y * d.blurSlab + L.blurOff;
    /* one item = output rows (y0+2m, y0+2m+1) x 4 pixels; needs pairs m .. m+4 (rows y0+2m-4 .. y0+2m+5) */
#pragma unroll
    for (int sweep = 0; sweep < (BLUR_TH / 2 + 15) / 16; sweep++) {
        const int m = (threadIdx.x >> 4) + 16 * sweep;
        const int gy = y0 + 2 * m, gx = x0 + 4 * q;
        if (m < BLUR_TH / 2 && gy < L.h && gx < L.bpitch) {
            uint4 pr[5];
#pragma unroll
            for (int k = 0; k < 5; k++) pr[k] = *reinterpret_cast<const uint4*>(&s_h[m + k][4 * q]);
            /* even row y: rows y-3..y+3 = (pair0.hi) (pair1) (pair2) (pair3); odd row y+1: (pair1) (pair2) (pair3) (pair4.lo) */
            const uint32_t E0 = 18u << 8, E1 = 34u | (48u << 8), E2 = 56u | (48u << 8), E3 = 34u | (18u << 8);
            const uint32_t O1 = 18u | (34u << 8), O2 = 48u | (56u << 8), O3 = 48u | (34u << 8), O4 = 18u;
            const uint32_t* P0 = &pr[0].x; const uint32_t* P1 = &pr[1].x; const uint32_t* P2 = &pr[2].x;
            const uint32_t* P3 = &pr[3].x; const uint32_t* P4 = &pr[4].x;
            uint32_t ve[4], vo[4];
#pragma unroll
            for (int k = 0; k < 4; k++) {
                ve[k] = __dp2a_lo(P0[k], E0, __dp2a_lo(P1[k], E1, __dp2a_lo(P2[k], E2, __dp2a_lo(P3[k], E3, 32768u))));
                vo[k] = __dp2a_lo(P1[k], O1, __dp2a_lo(P2[k], O2, __dp2a_lo(P3[k], O3, __dp2a_lo(P4[k], O4, 32768u))));
            }
            /* byte 2 of each Q16.16 sum is the output pixel (sums < 2^24) */
            const uint32_t oe = __byte_perm(__byte_perm(ve[0], ve[1], 0x0062), __byte_perm(ve[2], ve[3], 0x0062), 0x5410);
            const uint32_t oo = __byte_perm(__byte_perm(vo[0], vo[1], 0x0062), __byte_perm(vo[2], vo[3], 0x0062), 0x5410);
            uint8_t* o = dst + (uint32_t)(gy * L.bpitch + gx);
            *reinterpret_cast<uint32_t*>(o) = oe;
            if (gy + 1 < L.h) *reinterpret_cast<uint32_t*>(o + L.bpitch) = oo;
        }
    }
}

/* --------------------------------------------------------------------------------------------------- FAST */
#define FAST_TP 80   /* smem tile pitch: ROI width <= 76 (wCell <= 70, +6) + 3 bytes of alignment lead */
#define FAST_RH 76
#define FAST_NT 128
/* ring offsets of cv::FAST 9/16: (dx,dy) for k = 0..15 */
#define RING_OFF(k, TP)                                                                              \
    ((k) == 0 ? 3 * TP : (k) == 1 ? 3 * TP + 1 : (k) == 2 ? 2 * TP + 2 : (k) == 3 ? TP + 3 : (k) == 4 ? 3 : \
     (k) == 5 ? -TP + 3 : (k) == 6 ? -2 * TP + 2 : (k) == 7 ? -3 * TP + 1 : (k) == 8 ? -3 * TP :           \
     (k) == 9 ? -3 * TP - 1 : (k) == 10 ? -2 * TP - 2 : (k) == 11 ? -TP - 3 : (k) == 12 ? -3 :              \
     (k) == 13 ? TP - 3 : (k) == 14 ? 2 * TP - 2 : 3 * TP - 1)

#define RING_DX(k) ((k) == 0 ? 0 : (k) == 1 ? 1 : (k) == 2 ? 2 : (k) == 3 ? 3 : (k) == 4 ? 3 : (k) == 5 ? 3 : (k) == 6 ? 2 : (k) == 7 ? 1 : \
                    (k) == 8 ? 0 : (k) == 9 ? -1 : (k) == 10 ? -2 : (k) == 11 ? -3 : (k) == 12 ? -3 : (k) == 13 ? -3 : (k) == 14 ? -2 : -1)
#define RING_DY(k) ((k) == 0 ? 3 : (k) == 1 ? 3 : (k) == 2 ? 2 : (k) == 3 ? 1 : (k) == 4 ? 0 : (k) == 5 ? -1 : (k) == 6 ? -2 : (k) == 7 ? -3 : \
                    (k) == 8 ? -3 : (k) == 9 ? -3 : (k) == 10 ? -2 : (k) == 11 ? -1 : (k) == 12 ? 0 : (k) == 13 ? 1 : (k) == 14 ? 2 : 3)

static __host__ __device__ __forceinline__ bool run9(uint32_t m) {
    m |= m << 16;
    uint32_t x = m & (m >> 1);
    x &= x >> 2;
    x &= x >> 4;
    x &= m >> 8;
    return (x & 0xFFFFu) != 0;
}

/* corner at threshold t  <=>  9 contiguous ring pixels all > v+t or all < v-t.
 * Masks are gathered with one funnel shift per ring pixel: (B << 1) | sign(v + t - p) collects "p > v+t" (the bit
 * order comes out reversed, which a circular run test does not care about). */
static __host__ __device__ __forceinline__ bool fast_is_corner_vals(int v, const int (&r)[16], int t) {
    const int hi = v + t, lo = v - t;
    uint32_t B = 0, D = 0;
#pragma unroll
    for (int k = 0; k < 16; k++) {
#ifdef __CUDA_ARCH__
        B = __funnelshift_l((uint32_t)(hi - r[k]), B, 1);   /* sign bit set <=> r > hi */
        D = __funnelshift_l((uint32_t)(r[k] - lo), D, 1);   /* sign bit set <=> r < lo */
#else
        B = (B << 1) | ((uint32_t)(hi - r[k]) >> 31);
        D = (D << 1) | ((uint32_t)(r[k] - lo) >> 31);
#endif
    }
    return run9(B) || run9(D);
}

/* cv::FAST cornerScore = m - 1 with m = max over the 16 arcs of max(min_k (v - r_k), min_k (r_k - v)).
 * Evaluated on the raw ring values: min_k (v - r_k) = v - max_k r_k and min_k (r_k - v) = min_k r_k - v, so
 * m = max(v - min_arcs(max_arc r), max_arcs(min_arc r) - v).  Besides being cheaper this keeps every min/max
 * operand a plain non-negated register: nvcc 12.9 folds a negation into VIMNMX3 on sm_100a and the result
 * loses the sign (found by tools/fast_device_test3.cu: max(m, max(lo, -hi)) returned max(m, lo, hi)). */
#ifdef __CUDA_ARCH__
#define MIN3(a, b, c) __vimin3_s32((a), (b), (c))   /* one VIMNMX3; written as min(min()) nvcc 12.9 emits two VIMNMX for most of them */
#define MAX3(a, b, c) __vimax3_s32((a), (b), (c))
#else
#define MIN3(a, b, c) min(min((a), (b)), (c))
#define MAX3(a, b, c) max(max((a), (b)), (c))
#endif
static __host__ __device__ __forceinline__ int fast_score_vals(int v, const int (&r)[16]) {
    /* arcs of nine as three runs of three: 16 + 16 three-input min (max) per polarity, then a three-input reduction */
    int lo3[16], hi3[16];
#pragma unroll
    for (int k = 0; k < 16; k++) {
        lo3[k] = MIN3(r[k], r[(k + 1) & 15], r[(k + 2) & 15]);
        hi3[k] = MAX3(r[k], r[(k + 1) & 15], r[(k + 2) & 15]);
    }
    int lo9[16], hi9[16];
#pragma unroll
    for (int k = 0; k < 16; k++) {
        lo9[k] = MIN3(lo3[k], lo3[(k + 3) & 15], lo3[(k + 6) & 15]);
        hi9[k] = MAX3(hi3[k], hi3[(k + 3) & 15], hi3[(k + 6) & 15]);
    }
    int a[6], b[6];
#pragma unroll
    for (int k = 0; k < 5; k++) {
        a[k] = MAX3(lo9[3 * k], lo9[3 * k + 1], lo9[3 * k + 2]);
        b[k] = MIN3(hi9[3 * k], hi9[3 * k + 1], hi9[3 * k + 2]);
    }
    a[5] = lo9[15]; b[5] = hi9[15];
    const int maxOfMin = max(MAX3(a[0], a[1], a[2]), MAX3(a[3], a[4], a[5]));
    const int minOfMax = min(MIN3(b[0], b[1], b[2]), MIN3(b[3], b[4], b[5]));
    const int dark = v - minOfMax, bright = maxOfMin - v;
    return (dark > bright ? dark : bright) - 1;
}

/* the same on a pixel of a FAST_TP-pitch tile */
static __host__ __device__ __forceinline__ bool fast_is_corner(const uint8_t* c, int t) {
    int r[16];
#pragma unroll
    for (int k = 0; k < 16; k++) r[k] = (int)c[RING_OFF(k, FAST_TP)];
    return fast_is_corner_vals(c[0], r, t);
}
static __host__ __device__ __forceinline__ int fast_score(const uint8_t* c) {
    int r[16];
#pragma unroll
    for (int k = 0; k < 16; k++) r[k] = (int)c[RING_OFF(k, FAST_TP)];
    return fast_score_vals(c[0], r);
}

/* Sieve for four horizontally adjacent pixels at once (one 32-bit word of the tile): with a = |I(x,y+3) - I|,
 * b = |I(x,y-3) - I|, r = |I(x+3,y) - I|, l = |I(x-3,y) - I| (VABSDIFF4, one instruction per four pixels), a corner at
 * threshold t needs (a > t or b > t) and (r > t or l > t): every 9-arc of the ring holds one pixel of each opposite
 * pair.  Returns bit 7 of byte k set iff pixel k passes.  gt(x) = bytes of x that exceed t, exact per byte (t <= 127). */
static __device__ __forceinline__ uint32_t fast_sieve4(uint32_t c, uint32_t top, uint32_t bot, uint32_t lft, uint32_t rgt, uint32_t k127mt) {
    const uint32_t a = __vabsdiffu4(top, c), b = __vabsdiffu4(bot, c), r = __vabsdiffu4(rgt, c), l = __vabsdiffu4(lft, c);
    const uint32_t ga = ((a & 0x7F7F7F7Fu) + k127mt) | a, gb = ((b & 0x7F7F7F7Fu) + k127mt) | b;
    const uint32_t gr = ((r & 0x7F7F7F7Fu) + k127mt) | r, gl = ((l & 0x7F7F7F7Fu) + k127mt) | l;
    return (ga | gb) & (gr | gl) & 0x80808080u;
}

#define FAST_LISTCAP (70 * 70)
/* Fallback path of the cell stage.  One CTA = one 35-px-grid cell of one level of one frame, processed from its pixels:
 * cv::FAST(roi, th) with NMS seeing only this cell's scores at that threshold (SURVEY.md §7 hard part 2).  It runs only
 * for the cells that came out empty at iniTh (k_fast_stream + k_fast_nms) and redoes them at minTh
 * (src/ORBextractor.cc:825-828); persistent CTAs walk the list those kernels left in b.fbList.
 * Work is funnelled through compacted lists so that each stage runs dense:
 *   sieve (4 px per thread-step, all candidates) -> list1 -> exact 9-of-16 test -> list2 -> score -> NMS -> bitmask.
 * Output: the cell's keypoints in (y, x) order, packed, at the cell's slot range. */
static __device__ void fast_cell_from_pixels(const ExDesc& d, const ExBuffers& b, const int cellIdx, const size_t f, int th, const int npass) {
    /* dynamic smem, sized for the largest cell of this extractor: tile | score tile | list1 | list2 | mask */
    extern __shared__ __align__(16) uint8_t s_dyn[];
    uint8_t* s_img = s_dyn;
    uint8_t* s_score = s_img + d.fastRows * FAST_TP;
    uint16_t* s_list1 = reinterpret_cast<uint16_t*>(s_score + d.fastRows * FAST_TP);
    uint16_t* s_list2 = s_list1 + d.fastPix;
    uint32_t* s_mask = reinterpret_cast<uint32_t*>(s_list2 + d.fastPix);
    __shared__ int s_n1, s_n2, s_total;
    __shared__ int s_wsum[FAST_NT / 32];

    const CellDesc cell = b.cells[cellIdx];
    int* countOut = b.cellCount + f * d.nCells + cellIdx;
    const int cw = cell.rw - 6, ch = cell.rh - 6;
    if (cw <= 0 || ch <= 0) {
        if (threadIdx.x == 0) *countOut = 0;
        return;
    }
    const LevelDesc& L = d.lv[cell.level];
    const int ox = cell.x0 & 3;
    {
        /* tile load + score clear: 8 rows x 16 words per sweep, no index arithmetic beyond adds */
        const int tr = threadIdx.x >> 4, tc = threadIdx.x & 15;
        const int pitch = L.pitch, nw = cell.nw;
        const uint8_t* g = b.pyr + f * d.pyrSlab + L.pyrOff + (size_t)(cell.y0 + ORBFE_EDGE + tr) * pitch + ORBFE_ROW_LEAD + (cell.x0 - ox) + 4 * tc;
        uint32_t* ti = reinterpret_cast<uint32_t*>(s_img) + tr * (FAST_TP / 4) + tc;
        uint32_t* ts = reinterpret_cast<uint32_t*>(s_score) + tr * (FAST_TP / 4) + tc;
        for (int r = tr; r < cell.rh; r += FAST_NT / 16, g += (FAST_NT / 16) * pitch, ti += (FAST_NT / 16) * (FAST_TP / 4), ts += (FAST_NT / 16) * (FAST_TP / 4)) {
            if (tc < nw) { ti[0] = __ldg(reinterpret_cast<const uint32_t*>(g)); ts[0] = 0; }
            if (tc + 16 < nw) { ti[16] = __ldg(reinterpret_cast<const uint32_t*>(g + 64)); ts[16] = 0; }
        }
    }
    const int npix = cw * ch;
    const int nwords = (npix + 31) >> 5;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t invcw = cell.invcw; /* p / cw == (p * invcw) >> 24 for p < 70*70 */
    /* candidate columns are tile columns [cx0, cx0+cw); the sieve walks the tile words that overlap them */
    const int cx0 = ox + 3, w0 = cx0 >> 2, ngw = cell.ngw;
    const uint32_t invgw = cell.invgw;
    const int nitems = ch * ngw;
    for (int pass = 0; pass < npass; pass++) {
        if (threadIdx.x == 0) { s_n1 = 0; s_n2 = 0; }
        for (int i = threadIdx.x; i < nwords; i += FAST_NT) s_mask[i] = 0;
        __syncthreads();
        /* 1. sieve */
        const uint32_t k127mt = (uint32_t)(127 - th) * 0x01010101u;
        for (int i0 = 0; i0 < nitems; i0 += FAST_NT) {
            const int i = i0 + threadIdx.x;
            uint32_t hit = 0;
            int r = 0, col0 = 0;
            if (i < nitems) {
                r = (int)(((uint32_t)i * invgw) >> 24);
                const int wi = w0 + (i - r * ngw);
                const uint32_t* row = reinterpret_cast<const uint32_t*>(s_img + (r + 3) * FAST_TP);
                const uint32_t c = row[wi], lw = row[wi - 1], rw = row[wi + 1];
                const uint32_t top = row[wi + 3 * (FAST_TP / 4)], bot = row[wi - 3 * (FAST_TP / 4)];
                hit = fast_sieve4(c, top, bot, __byte_perm(lw, c, 0x4321), __byte_perm(c, rw, 0x6543), k127mt);
                col0 = 4 * wi - cx0; /* candidate x of byte 0 (may be negative at the left edge) */
                /* drop the bytes of the first / last word that lie outside the candidate columns */
                const int gi = wi - w0;
                if (gi == 0) hit &= cell.edgeL;
                if (gi == ngw - 1) hit &= cell.edgeR;
            }
            const int cnt = __popc(hit);
            int incl = cnt;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int t = __shfl_up_sync(0xFFFFFFFFu, incl, o);
                if (lane >= o) incl += t;
            }
            const int wtot = __shfl_sync(0xFFFFFFFFu, incl, 31);
            if (wtot) {
                int base = 0;
                if (lane == 31) base = atomicAdd(&s_n1, wtot);
                base = __shfl_sync(0xFFFFFFFFu, base, 31) + incl - cnt;
#pragma unroll
                for (int k = 0; k < 4; k++)
                    if (hit & (0x80u << (8 * k))) s_list1[base++] = (uint16_t)(r * cw + col0 + k);
            }
        }
        __syncthreads();
        const int n1 = s_n1;
        /* 2. exact test on the sieve's survivors */
        for (int e0 = 0; e0 < n1; e0 += FAST_NT) {
            const int e = e0 + threadIdx.x;
            bool isc = false;
            int p = 0;
            if (e < n1) {
                p = s_list1[e];
                const int py = (int)(((uint32_t)p * invcw) >> 24), px = p - py * cw;
                isc = fast_is_corner(s_img + (py + 3) * FAST_TP + cx0 + px, th);
            }
            const uint32_t bal = __ballot_sync(0xFFFFFFFFu, isc);
            if (bal) {
                int base = 0;
                if (lane == 0) base = atomicAdd(&s_n2, __popc(bal));
                base = __shfl_sync(0xFFFFFFFFu, base, 0);
                if (isc) s_list2[base + __popc(bal & ((1u << lane) - 1))] = (uint16_t)p;
            }
        }
        __syncthreads();
        const int n2 = s_n2;
        /* 3. scores */
        for (int e = threadIdx.x; e < n2; e += FAST_NT) {
            const int p = s_list2[e];
            const int py = (int)(((uint32_t)p * invcw) >> 24), px = p - py * cw;
            const int o = (py + 3) * FAST_TP + cx0 + px;
            s_score[o] = (uint8_t)fast_score(s_img + o);
        }
        __syncthreads();
        /* 4. non-maximum suppression: strictly greater than the 8 neighbours (non-corners count 0) */
        for (int e = threadIdx.x; e < n2; e += FAST_NT) {
            const int p = s_list2[e];
            const int py = (int)(((uint32_t)p * invcw) >> 24), px = p - py * cw;
            const uint8_t* sc = s_score + (py + 3) * FAST_TP + cx0 + px;
            const int v = sc[0];
            const int nb = max(max(max((int)sc[-FAST_TP - 1], (int)sc[-FAST_TP]), max((int)sc[-FAST_TP + 1], (int)sc[-1])),
                               max(max((int)sc[1], (int)sc[FAST_TP - 1]), max((int)sc[FAST_TP], (int)sc[FAST_TP + 1])));
            if (v > nb) atomicOr(&s_mask[p >> 5], 1u << (p & 31));
        }
        __syncthreads();
        /* 5. ordered count: each thread owns a contiguous run of mask words, so thread order = (y,x) order */
        const int per = (nwords + FAST_NT - 1) / FAST_NT;
        const int wbeg = min((int)threadIdx.x * per, nwords), wend = min(wbeg + per, nwords);
        int mine = 0;
        for (int i = wbeg; i < wend; i++) mine += __popc(s_mask[i]);
        int incl = mine;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_up_sync(0xFFFFFFFFu, incl, o);
            if (lane >= o) incl += t;
        }
        if (lane == 31) s_wsum[warp] = incl;
        __syncthreads();
        int wbase = 0;
        for (int w = 0; w < warp; w++) wbase += s_wsum[w];
        if (threadIdx.x == FAST_NT - 1) s_total = wbase + incl;
        __syncthreads();
        const int total = s_total;
        if (total > 0 || pass == npass - 1) {
            int k = wbase + incl - mine;
            uint32_t* out = b.cand + f * d.candCap + cell.candBase;
            for (int i = wbeg; i < wend; i++) {
                uint32_t m = s_mask[i];
                while (m) {
                    const int bit = __ffs(m) - 1;
                    m &= m - 1;
                    const int p = i * 32 + bit;
                    const int py = (int)(((uint32_t)p * invcw) >> 24), px = p - py * cw;
                    const int sc = s_score[(py + 3) * FAST_TP + cx0 + px];
                    out[k++] = (uint32_t)(px + 3 + cell.xoff) | ((uint32_t)(py + 3 + cell.yoff) << 12) | ((uint32_t)sc << 24);
                }
            }
            if (threadIdx.x == 0) *countOut = total;
            return;
        }
        /* empty at iniTh: wipe the scores written so far and retry with minTh (src/ORBextractor.cc:825-828) */
        for (int e = threadIdx.x; e < n2; e += FAST_NT) {
            const int p = s_list2[e];
            const int py = (int)(((uint32_t)p * invcw) >> 24), px = p - py * cw;
            s_score[(py + 3) * FAST_TP + cx0 + px] = 0;
        }
        th = d.minTh;
        __syncthreads();
    }
}

/* Persistent CTAs re-detect, at minTh, the cells listed in b.fbList (entry = frame * nCells + cell). */
__global__ void __launch_bounds__(FAST_NT, 8) k_fast_cells_list(const __grid_constant__ ExDesc d, ExBuffers b) {
    const int n = *b.fbCount;
    for (int i = blockIdx.x; i < n; i += gridDim.x) {
        const int id = b.fbList[i];
        const int f = id / d.nCells;
        fast_cell_from_pixels(d, b, id - f * d.nCells, (size_t)f, d.minTh, 1);
        __syncthreads();
    }
}

/* ---- streaming detector: the iniTh pass over whole levels ------------------------------------------------------ */
#ifndef FS_WARPS
#define FS_WARPS 1
#endif
#ifndef FS_R
#define FS_R 8          /* rows per compaction chunk: a multiple of 8 (one 32-bit survivor mask per lane and 8 rows) */
#endif
#define FS_H (FS_R / 8)
#define FS_COLS 120     /* candidate columns per strip (lanes 1..30; lanes 0 and 31 carry the 4 px halos) */

/* One warp = one strip (120 columns x up to 32 rows; one strip = one CTA) of one level of one frame; warps are independent (no block
 * barrier).  Each lane owns four adjacent columns.  Per row the warp loads the row three below (coalesced 128 B),
 * keeps the last 16 rows in a shared-memory ring, and runs the 4-pixel sieve on registers (left/right neighbours come
 * by shuffle).  Every FS_R rows the sieve's survivors are compacted (per-lane bit masks, one warp scan) into a queue,
 * the exact 9-of-16 test runs densely over the queue (corners compacted in place), scores are computed for the corners
 * and each corner (x, y, score) is appended to the raw list of the 35-px cell that contains it (one global atomic).
 * k_fast_nms then does the cell-local non-maximum suppression and ordering. */
#define FS_BROWS (FS_R + 6)   /* rows yc-3 .. yc+FS_R+2 of the current chunk */
#define FS_PW 33               /* words per buffered row: 32 + 1 so that equal columns of different rows hit different banks */
#ifndef FS_MINB
#define FS_MINB (32 / FS_WARPS)
#endif
__global__ void __launch_bounds__(FS_WARPS * 32, FS_MINB) k_fast_stream(const __grid_constant__ ExDesc d, ExBuffers b, int stripBase, int nStrips) {
    __shared__ __align__(16) uint32_t s_buf[FS_WARPS][FS_BROWS][FS_PW];
    __shared__ uint16_t s_q[FS_WARPS][FS_R * FS_COLS + 32];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const size_t f = blockIdx.y;
    if (blockIdx.x == 0 && threadIdx.x == 0 && f == 0) *b.fbCount = 0; /* list of empty cells restarts every batch */
    const int sidx = blockIdx.x * FS_WARPS + warp;
    if (sidx >= nStrips) return;
    const StripDesc sd = b.strips[stripBase + sidx];
    const LevelDesc& L = d.lv[sd.level];
    const int pitch = L.pitch, th = d.iniTh;
    const int col = sd.xs - 4 + 4 * lane;                       /* level x of my word's byte 0 */
    const int colc = min(col, (L.w + ORBFE_EDGE - 4) & ~3);     /* keep the load inside the padded row */
    const int y0 = sd.y0, nrows = sd.nrows;
    /* p walks down the level: it always points at the next row to fetch (row yc + r + 3) */
    const uint8_t* p = b.pyr + f * d.pyrSlab + (uint32_t)(L.pyrOff + (ORBFE_EDGE + y0 - 3) * pitch + ORBFE_ROW_LEAD + colc);
    uint32_t* buf = &s_buf[warp][0][0];
    const uint8_t* bufb = reinterpret_cast<const uint8_t*>(buf) + 3 * (4 * FS_PW);   /* row yc of the chunk */
    uint16_t* q = s_q[warp];
    uint32_t vmask = 0;
#pragma unroll
    for (int k = 0; k < 4; k++)
        if (lane >= 1 && lane <= 30 && col + k >= ORBFE_EDGE && col + k < L.w - ORBFE_EDGE) vmask |= 0x80u << (8 * k);
#pragma unroll
    for (int r = 0; r < 6; r++, p += pitch) buf[r * FS_PW + lane] = __ldg(reinterpret_cast<const uint32_t*>(p));   /* rows y0-3 .. y0+2 */
    const uint32_t k127mt = (uint32_t)(127 - th) * 0x01010101u;
    int* rawCount = b.rawCount + f * d.nCells;
    uint32_t* raw = b.raw + f * (size_t)d.rawCap;
    /* the fetch of the next row is always one step ahead of its use (rows beyond the strip stay inside the padded level) */
    uint32_t nxt = __ldg(reinterpret_cast<const uint32_t*>(p));
    for (int c0 = 0; c0 < nrows; c0 += FS_R) {
        const int rows = min(FS_R, nrows - c0);
        __syncwarp();
        /* 1. sieve, one row per step: buffer rows r (y-3), r+3 (y), r+6 (y+3, fetched one step earlier).  The survivors of
         *    my four columns collect in hm, one nibble per row: the high half of h * 0x02040810 holds bits 7, 15, 23, 31 of h
         *    in its low nibble, and the funnel shift keeps exactly that nibble */
        uint32_t hmv[FS_H];
        int cnt = 0;
#pragma unroll
        for (int g = 0; g < FS_H; g++) {
            const int gr = min(8, rows - 8 * g);   /* rows of this group of eight (<= 0: none) */
            uint32_t hm = 0;
            for (int rr = 0; rr < gr; rr++) {
                const int r = 8 * g + rr;
                const uint32_t top = nxt;
                p += pitch;
                nxt = __ldg(reinterpret_cast<const uint32_t*>(p));
                buf[(r + 6) * FS_PW + lane] = top;
                const uint32_t cen = buf[(r + 3) * FS_PW + lane], bot = buf[r * FS_PW + lane];
                const uint32_t lw = __shfl_up_sync(0xFFFFFFFFu, cen, 1), rw = __shfl_down_sync(0xFFFFFFFFu, cen, 1);
                const uint32_t h = fast_sieve4(cen, top, bot, __byte_perm(lw, cen, 0x4321), __byte_perm(cen, rw, 0x6543), k127mt) & vmask;
                hm = __funnelshift_r(hm, __umulhi(h, 0x02040810u), 4);
            }
            hm = gr > 0 ? hm >> (4 * (8 - gr)) : 0u;   /* bit 4r + k: row r of the group, column k of my word */
            hmv[g] = hm;
            cnt += __popc(hm);
        }
        __syncwarp();
        /* 2. compaction: one warp scan of the per-lane counts gives every lane its run of queue slots */
        int incl = cnt;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_up_sync(0xFFFFFFFFu, incl, o);
            if (lane >= o) incl += t;
        }
        const int n = __shfl_sync(0xFFFFFFFFu, incl, 31);
        {
            int pos = incl - cnt;
#pragma unroll
            for (int g = 0; g < FS_H; g++) {
                uint32_t m = hmv[g];
                while (m) {
                    const int bit = __ffs(m) - 1;
                    m &= m - 1;
                    q[pos++] = (uint16_t)((8 * g + (bit >> 2)) * (4 * FS_PW) + 4 * lane + (bit & 3));   /* byte offset from bufb */
                }
            }
        }
        if (n > 0) {
            __syncwarp();
            /* 3. exact test over the queue; corners are written back at the front of the queue */
            int m = 0;
            for (int e0 = 0; e0 < n; e0 += 32) {
                const int e = e0 + lane;
                bool isc = false;
                uint32_t ent = 0;
                if (e < n) {
                    ent = q[e];
                    const uint8_t* c = bufb + ent;
                    int rv[16];
#pragma unroll
                    for (int k = 0; k < 16; k++) rv[k] = c[RING_OFF(k, (4 * FS_PW))];
                    isc = fast_is_corner_vals(c[0], rv, th);
                }
                const uint32_t bal = __ballot_sync(0xFFFFFFFFu, isc);
                if (isc) q[m + __popc(bal & ((1u << lane) - 1))] = (uint16_t)ent;   /* m <= e0: never ahead of the reads */
                m += __popc(bal);
                __syncwarp();
            }
            /* 4. scores, and hand every corner to its cell */
            for (int e = lane; e < m; e += 32) {
                const uint32_t ent = q[e];
                const uint8_t* c = bufb + ent;
                int rv[16];
#pragma unroll
                for (int k = 0; k < 16; k++) rv[k] = c[RING_OFF(k, (4 * FS_PW))];
                const int sc = fast_score_vals(c[0], rv);
                const int er = (int)ent / (4 * FS_PW);
                const int x = sd.xs - 4 + ((int)ent - er * (4 * FS_PW)), y = y0 + c0 + er;
                const int cj = (int)(((uint32_t)(x - ORBFE_EDGE) * L.invW) >> 24), ci = (int)(((uint32_t)(y - ORBFE_EDGE) * L.invH) >> 24);
                const int cell = ci * L.nCols + cj;
                const int slot = atomicAdd(&rawCount[L.cellBase + cell], 1);
                raw[L.rawBase + cell * L.rawStride + slot] = (uint32_t)(x - ORBFE_MIN_BORDER) | ((uint32_t)(y - ORBFE_MIN_BORDER) << 12) | ((uint32_t)sc << 24);
            }
            __syncwarp();
        }
        /* 5. the last six rows of this chunk are the first six of the next */
        if (c0 + FS_R < nrows) {
            uint32_t keep[6];
#pragma unroll
            for (int r = 0; r < 6; r++) keep[r] = buf[(FS_R + r) * FS_PW + lane];
#pragma unroll
            for (int r = 0; r < 6; r++) buf[r * FS_PW + lane] = keep[r];
        }
    }
}

/* One warp = one cell (four cells per CTA, no block barrier): scatter the cell's raw corners into a zeroed score tile,
 * keep the strict local maxima (cv::FAST nonmaxSuppression: neighbours outside the cell's candidate area or below the
 * threshold count as 0), write them in (y, x) order.  A cell that ends up empty is queued for the minTh fallback
 * (k_fast_cells_list). */
#ifndef FN_WARPS
#define FN_WARPS 1
#endif
#define FN_TP 72
__global__ void __launch_bounds__(FN_WARPS * 32) k_fast_nms(const __grid_constant__ ExDesc d, ExBuffers b) {
    __shared__ __align__(16) uint8_t s_scAll[FN_WARPS][FN_TP * 80];   /* (ch+2) rows x tp <= 72 x 80 */
    __shared__ uint32_t s_maskAll[FN_WARPS][(FAST_LISTCAP + 31) / 32 + 1];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int cellIdx = blockIdx.x * FN_WARPS + warp;
    if (cellIdx >= d.nCells) return;
    const size_t f = blockIdx.y;
    uint8_t* s_sc = s_scAll[warp];
    uint32_t* s_mask = s_maskAll[warp];
    const CellDesc cell = b.cells[cellIdx];
    int* countOut = b.cellCount + f * d.nCells + cellIdx;
    const int cw = cell.rw - 6, ch = cell.rh - 6;
    if (cw <= 0 || ch <= 0) {
        if (lane == 0) *countOut = 0;
        return;
    }
    int* rawCountP = b.rawCount + f * d.nCells + cellIdx;
    const int n = *rawCountP;
    int total = 0;
    if (n > 0) {
        const LevelDesc& L = d.lv[cell.level];
        const uint32_t* raw = b.raw + f * (size_t)d.rawCap + L.rawBase + (cellIdx - L.cellBase) * L.rawStride;
        const int tp = (cw + 2 + 15) & ~15, nwords = (cw * ch + 31) >> 5;
        /* the first raw entries are fetched while the tile is being cleared */
        uint32_t e0 = lane < n ? __ldg(raw + lane) : 0u, e1 = lane + 32 < n ? __ldg(raw + lane + 32) : 0u;
        for (int i = lane; i < (ch + 2) * tp / 16; i += 32) reinterpret_cast<uint4*>(s_sc)[i] = make_uint4(0, 0, 0, 0);
        for (int i = lane; i < nwords; i += 32) s_mask[i] = 0;
        __syncwarp();
        const int bx = 3 + cell.xoff - 1, by = 3 + cell.yoff - 1; /* raw x,y are relative to the 16 px border; tile has a 1 px apron */
        for (int e = lane; e < n; e += 32) {
            const uint32_t p = e < 32 ? e0 : (e < 64 ? e1 : __ldg(raw + e));
            s_sc[(int)(((p >> 12) & 0xFFF) - by) * tp + (int)((p & 0xFFF) - bx)] = (uint8_t)(p >> 24);
        }
        __syncwarp();
        for (int e = lane; e < n; e += 32) {
            const uint32_t p = e < 32 ? e0 : (e < 64 ? e1 : __ldg(raw + e));
            const int px = (int)(p & 0xFFF) - bx, py = (int)((p >> 12) & 0xFFF) - by;   /* 1-based inside the tile */
            const uint8_t* sc = s_sc + py * tp + px;
            const int v = sc[0];
            const int nb = max(max(max((int)sc[-tp - 1], (int)sc[-tp]), max((int)sc[-tp + 1], (int)sc[-1])),
                               max(max((int)sc[1], (int)sc[tp - 1]), max((int)sc[tp], (int)sc[tp + 1])));
            if (v > nb) {
                const int pi = (py - 1) * cw + (px - 1);
                atomicOr(&s_mask[pi >> 5], 1u << (pi & 31));
            }
        }
        __syncwarp();
        /* ordered output: lane owns a contiguous run of mask words */
        const int per = (nwords + 31) >> 5;
        const int wbeg = min(lane * per, nwords), wend = min(wbeg + per, nwords);
        int mine = 0;
        for (int i = wbeg; i < wend; i++) mine += __popc(s_mask[i]);
        int incl = mine;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_up_sync(0xFFFFFFFFu, incl, o);
            if (lane >= o) incl += t;
        }
        total = __shfl_sync(0xFFFFFFFFu, incl, 31);
        int k = incl - mine;
        uint32_t* out = b.cand + f * d.candCap + cell.candBase;
        const uint32_t invcw = cell.invcw;
        for (int i = wbeg; i < wend; i++) {
            uint32_t mm = s_mask[i];
            while (mm) {
                const int bit = __ffs(mm) - 1;
                mm &= mm - 1;
                const int pi = i * 32 + bit;
                const int py = (int)(((uint32_t)pi * invcw) >> 24), px = pi - py * cw;
                out[k++] = (uint32_t)(px + 3 + cell.xoff) | ((uint32_t)(py + 3 + cell.yoff) << 12) | ((uint32_t)s_sc[(py + 1) * tp + px + 1] << 24);
            }
        }
    }
    if (lane == 0) {
        *countOut = total;
        if (n > 0) *rawCountP = 0;                        /* ready for the next batch */
        if (total == 0) b.fbList[atomicAdd(b.fbCount, 1)] = (int)f * d.nCells + cellIdx;
    }
}

/* ----------------------------------------------------------------------------------------------- quadtree */
#ifndef QT_NT
#define QT_NT 256
#endif
#define QT_NOKEY 0xFFFFFFFFFFFFFFFFull   /* sorts behind every real key */

/* In-place exclusive scan of data[0..n) by the whole CTA; returns the total. tmp: QT_NT/32+1 ints. */
static __device__ int block_exscan(int* data, int n, int* tmp) {
    const int per = (n + QT_NT - 1) / QT_NT;
    const int beg = min(threadIdx.x * per, n), end = min(beg + per, n);
    int sum = 0;
    for (int i = beg; i < end; i++) sum += data[i];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int incl = sum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(0xFFFFFFFFu, incl, o);
        if (lane >= o) incl += t;
    }
    __syncthreads(); /* tmp may still be read from a previous call */
    if (lane == 31) tmp[warp] = incl;
    __syncthreads();
    if (warp == 0) {
        int w = lane < QT_NT / 32 ? tmp[lane] : 0;
        int wi = w;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_up_sync(0xFFFFFFFFu, wi, o);
            if (lane >= o) wi += t;
        }
        if (lane < QT_NT / 32) tmp[lane] = wi - w;
        if (lane == QT_NT / 32 - 1) tmp[QT_NT / 32] = wi;
    }
    __syncthreads();
    int run = tmp[warp] + incl - sum;
    for (int i = beg; i < end; i++) {
        const int v = data[i];
        data[i] = run;
        run += v;
    }
    const int total = tmp[QT_NT / 32];
    __syncthreads();
    return total;
}

static __device__ __forceinline__ int quadrant(short4 bx, int x, int y) {
    /* ExtractorNode::DivideNode, src/ORBextractor.cc:481-523: halves are ceil((x1-x0)/2) */
    const int mx = bx.x + ((bx.z - bx.x + 1) >> 1), my = bx.y + ((bx.w - bx.y + 1) >> 1);
    return (x < mx ? 0 : 1) + (y < my ? 0 : 2);
}
static __device__ __forceinline__ short4 child_box(short4 bx, int c) {
    const int mx = bx.x + ((bx.z - bx.x + 1) >> 1), my = bx.y + ((bx.w - bx.y + 1) >> 1);
    short4 r;
    r.x = (c & 1) ? mx : bx.x;
    r.z = (c & 1) ? bx.z : mx;
    r.y = (c & 2) ? my : bx.y;
    r.w = (c & 2) ? bx.w : my;
    return r;
}

/* One CTA = DistributeOctTree for one level of one frame (src/ORBextractor.cc:537-761).  The std::list is an
 * array kept in list order; one "round" rebuilds it: children of the nodes expanded this round go to the front
 * in reverse processing order (push_front), untouched nodes keep their relative order behind them.  Ties between
 * equal-sized nodes in the final partial rounds are broken by creation order (latest first) = ascending list
 * position among the nodes created in the previous round.
 * Dynamic smem layout: see the carve-up below (cap = d.nodeCap, a power of two). */
template <bool SCRATCH>   /* SCRATCH: the node arrays live in b.qtScratch (global) instead of shared memory */
__global__ void __launch_bounds__(QT_NT) k_quadtree(const __grid_constant__ ExDesc d, ExBuffers b) {
    extern __shared__ __align__(16) uint8_t qt_dyn[];
    const int cap = d.nodeCap;
    /* the node arrays: shared memory, or this CTA's block of b.qtScratch when a level may hold more than 2048 nodes */
    uint8_t* smem = SCRATCH ? b.qtScratch + ((size_t)blockIdx.y * d.nlevels + blockIdx.x) * (size_t)d.qtScratchBytes : qt_dyn;
    short4* box0 = reinterpret_cast<short4*>(smem);
    short4* box1 = box0 + cap;
    int* cnt0 = reinterpret_cast<int*>(box1 + cap);
    int* cnt1 = cnt0 + cap;
    int* childCnt = cnt1 + cap;            /* [cap][4] */
    int* scanA = childCnt + 4 * cap;       /* [cap] S: children before me in processing order */
    int* scanB = scanA + cap;              /* [cap] unexpanded-before-me */
    uint64_t* sortK = reinterpret_cast<uint64_t*>(scanB + cap); /* [cap] (~size << 32 | list position): any node count, any node size */
    uint16_t* childPos = reinterpret_cast<uint16_t*>(sortK + cap); /* [cap][4] */
    uint16_t* newPos = childPos + 4 * cap; /* [cap] */
    uint8_t* expand = reinterpret_cast<uint8_t*>(newPos + cap); /* [cap] */
    __shared__ int s_tmp[QT_NT / 32 + 2];
    __shared__ int s_nexp, s_cut;

    const int level = blockIdx.x;
    const size_t f = blockIdx.y;
    const LevelDesc& L = d.lv[level];
    const int N = L.quota;
    const int tid = threadIdx.x;
    uint32_t* keys = b.keys + f * d.candCap + L.keyBase;
    uint16_t* keyNode = b.keyNode + f * d.candCap + L.keyBase;
    int* selCount = b.selCount + f * d.nlevels + level;

    /* ---- gather the level's candidates in (cell row, cell col, y, x) order = vToDistributeKeys */
    const int ncell = L.nCols * L.nRows;
    const int* cellCount = b.cellCount + f * d.nCells + L.cellBase;
    int nKeys;
    {
        /* scanA doubles as the per-cell prefix; levels with more cells than node slots go through it in slices of `cap` cells */
        int* pre = scanA;
        const uint32_t* cand = b.cand + f * d.candCap;
        const int warp = tid >> 5, lane = tid & 31;
        nKeys = 0;
        for (int c0 = 0; c0 < ncell; c0 += cap) {
            const int m = min(cap, ncell - c0);
            for (int i = tid; i < m; i += QT_NT) pre[i] = cellCount[c0 + i];
            __syncthreads();
            const int tot = block_exscan(pre, m, s_tmp);
            for (int c = warp; c < m; c += QT_NT / 32) {
                const int n = cellCount[c0 + c];
                const uint32_t* src = cand + b.cells[L.cellBase + c0 + c].candBase;
                for (int i = lane; i < n; i += 32) keys[nKeys + pre[c] + i] = src[i];
            }
            nKeys += tot;
            __syncthreads();
        }
    }
    if (nKeys == 0) {
        if (tid == 0) *selCount = 0;
        return;
    }

    /* ---- initial nodes (src/ORBextractor.cc:541-586) */
    short4* box = box0;
    short4* boxN = box1;
    int* cnt = cnt0;
    int* cntN = cnt1;
    const int nIni = L.nIni;
    for (int i = tid; i < nIni; i += QT_NT) {
        cnt1[i] = 0;
        short4 bx;
        bx.x = (short)(int)__fmul_rn(L.hX, (float)i);
        bx.z = (short)(int)__fmul_rn(L.hX, (float)(i + 1));
        bx.y = 0;
        bx.w = (short)(L.h - 2 * ORBFE_MIN_BORDER);
        box1[i] = bx;
    }
    __syncthreads();
    for (int k = tid; k < nKeys; k += QT_NT) {
        const int x = keys[k] & 0xFFF;
        const int n = (int)__fdiv_rn((float)x, L.hX);
        keyNode[k] = (uint16_t)n;
        atomicAdd(&cnt1[n], 1);
    }
    __syncthreads();
    for (int i = tid; i < nIni; i += QT_NT) scanA[i] = cnt1[i] > 0;
    __syncthreads();
    int nCur = block_exscan(scanA, nIni, s_tmp);
    for (int i = tid; i < nIni; i += QT_NT)
        if (cnt1[i] > 0) { box[scanA[i]] = box1[i]; cnt[scanA[i]] = cnt1[i]; }
    for (int k = tid; k < nKeys; k += QT_NT) keyNode[k] = (uint16_t)scanA[keyNode[k]];
    __syncthreads();

    bool partial = false;
    int nNew = 0; /* nodes created by the previous round: positions [0,nNew) */
    for (;;) {
        const int prevSize = nCur;
        /* 1. who may be expanded this round */
        for (int i = tid; i < nCur; i += QT_NT) {
            expand[i] = partial ? (i < nNew && cnt[i] > 1) : (cnt[i] > 1);
            childCnt[4 * i] = childCnt[4 * i + 1] = childCnt[4 * i + 2] = childCnt[4 * i + 3] = 0;
        }
        __syncthreads();
        /* 2. children sizes */
        for (int k = tid; k < nKeys; k += QT_NT) {
            const int i = keyNode[k];
            if (expand[i]) {
                const uint32_t kk = keys[k];
                atomicAdd(&childCnt[4 * i + quadrant(box[i], kk & 0xFFF, (kk >> 12) & 0xFFF)], 1);
            }
        }
        __syncthreads();
        /* 3. processing order and (partial rounds) the cut where the list reaches N */
        int C = 0;
        if (!partial) {
            for (int i = tid; i < nCur; i += QT_NT) {
                const int nch = (childCnt[4 * i] > 0) + (childCnt[4 * i + 1] > 0) + (childCnt[4 * i + 2] > 0) + (childCnt[4 * i + 3] > 0);
                scanA[i] = expand[i] ? nch : 0;
            }
            __syncthreads();
        } else {
            /* sort candidates by (size desc, list position asc) = the reference's descending walk of the sorted
             * (size, pointer) vector with later-created nodes at higher addresses */
            int n2 = 1;
            while (n2 < nNew) n2 <<= 1;
            for (int i = tid; i < n2; i += QT_NT)
                sortK[i] = (i < nNew && expand[i]) ? (((uint64_t)(~(uint32_t)cnt[i]) << 32) | (uint32_t)i) : QT_NOKEY;
            __syncthreads();
            for (int k2 = 2; k2 <= n2; k2 <<= 1)
                for (int j = k2 >> 1; j > 0; j >>= 1) {
                    for (int i = tid; i < n2; i += QT_NT) {
                        const int ixj = i ^ j;
                        if (ixj > i) {
                            const uint64_t a = sortK[i], c = sortK[ixj];
                            const bool up = (i & k2) == 0;
                            if ((a > c) == up) { sortK[i] = c; sortK[ixj] = a; }
                        }
                    }
                    __syncthreads();
                }
            /* scanB[j] = list growth when the j-th candidate is divided (children - 1) */
            for (int j = tid; j < nNew; j += QT_NT) {
                int g = 0;
                if (sortK[j] != QT_NOKEY) {
                    const int i = (int)(uint32_t)sortK[j];
                    g = (childCnt[4 * i] > 0) + (childCnt[4 * i + 1] > 0) + (childCnt[4 * i + 2] > 0) + (childCnt[4 * i + 3] > 0) - 1;
                }
                scanB[j] = g;
            }
            if (tid == 0) s_cut = nNew;
            __syncthreads();
            block_exscan(scanB, nNew, s_tmp);
            /* first j whose inclusive growth brings the list to N: everything after it is left alone */
            for (int j = tid; j < nNew; j += QT_NT) {
                if (sortK[j] == QT_NOKEY) continue;
                const int i = (int)(uint32_t)sortK[j];
                const int g = (childCnt[4 * i] > 0) + (childCnt[4 * i + 1] > 0) + (childCnt[4 * i + 2] > 0) + (childCnt[4 * i + 3] > 0) - 1;
                if (prevSize + scanB[j] + g >= N) atomicMin(&s_cut, j);
            }
            __syncthreads();
            const int cut = s_cut;
            for (int i = tid; i < nCur; i += QT_NT) { expand[i] = 0; scanA[i] = 0; }
            __syncthreads();
            /* children of the candidates that do get divided, prefixed in processing (= sorted) order */
            for (int j = tid; j < nNew; j += QT_NT) {
                int nch = 0;
                if (sortK[j] != QT_NOKEY && j <= cut) {
                    const int i = (int)(uint32_t)sortK[j];
                    expand[i] = 1;
                    nch = (childCnt[4 * i] > 0) + (childCnt[4 * i + 1] > 0) + (childCnt[4 * i + 2] > 0) + (childCnt[4 * i + 3] > 0);
                }
                scanB[j] = nch;
            }
            __syncthreads();
            C = block_exscan(scanB, nNew, s_tmp);
            for (int j = tid; j < nNew; j += QT_NT)
                if (sortK[j] != QT_NOKEY && j <= cut) scanA[(uint32_t)sortK[j]] = scanB[j];
            __syncthreads();
        }
        if (!partial) C = block_exscan(scanA, nCur, s_tmp);
        for (int i = tid; i < nCur; i += QT_NT) scanB[i] = expand[i] ? 0 : 1;
        __syncthreads();
        const int nKeep = block_exscan(scanB, nCur, s_tmp);
        const int nNext = C + nKeep;
        if (tid == 0) s_nexp = 0;
        __syncthreads();
        /* 4. build the next list */
        for (int i = tid; i < nCur; i += QT_NT) {
            if (expand[i]) {
                int k = 0, big = 0;
#pragma unroll
                for (int c = 0; c < 4; c++) {
                    const int n = childCnt[4 * i + c];
                    if (n > 0) {
                        const int pos = C - 1 - (scanA[i] + k);
                        boxN[pos] = child_box(box[i], c);
                        cntN[pos] = n;
                        childPos[4 * i + c] = (uint16_t)pos;
                        big += n > 1;
                        k++;
                    }
                }
                if (big) atomicAdd(&s_nexp, big);
            } else {
                const int pos = C + scanB[i];
                boxN[pos] = box[i];
                cntN[pos] = cnt[i];
                newPos[i] = (uint16_t)pos;
            }
        }
        __syncthreads();
        /* 5. move the keys */
        for (int k = tid; k < nKeys; k += QT_NT) {
            const int i = keyNode[k];
            if (expand[i]) {
                const uint32_t kk = keys[k];
                keyNode[k] = childPos[4 * i + quadrant(box[i], kk & 0xFFF, (kk >> 12) & 0xFFF)];
            } else {
                keyNode[k] = newPos[i];
            }
        }
        const int nToExpand = s_nexp;
        __syncthreads();
        { short4* t = box; box = boxN; boxN = t; }
        { int* t = cnt; cnt = cntN; cntN = t; }
        nCur = nNext;
        nNew = C;
        /* 6. termination (src/ORBextractor.cc:655-735) */
        if (nCur >= N || nCur == prevSize) break;
        if (!partial) {
            if (nCur + nToExpand * 3 > N) partial = true;
        }
    }

    /* ---- keep the best response of every node; ties -> first in vToDistributeKeys order (src/ORBextractor.cc:740-758) */
    uint32_t* best = reinterpret_cast<uint32_t*>(childCnt);
    for (int i = tid; i < nCur; i += QT_NT) best[i] = 0;
    __syncthreads();
    for (int k = tid; k < nKeys; k += QT_NT)
        atomicMax(&best[keyNode[k]], ((keys[k] >> 24) << 24) | (uint32_t)(0xFFFFFF - k));
    __syncthreads();
    uint32_t* sel = b.sel + f * d.kpCap + L.selBase;
    for (int i = tid; i < nCur; i += QT_NT) sel[i] = keys[0xFFFFFF - (best[i] & 0xFFFFFF)];
    if (tid == 0) *selCount = nCur;
}

/* ----------------------------------------------------------------------------------- ordering (K7 of SURVEY) */
#define AS_NT 256
/* One CTA = one frame: level-ordered keypoints -> final slots.  Keypoints inside the lapping interval fill the
 * output from the back, the others from the front (src/ORBextractor.cc:1104,1135-1144). */
__global__ void __launch_bounds__(AS_NT) k_assemble(const __grid_constant__ ExDesc d, ExBuffers b, int lap0, int lap1) {
    __shared__ int s_start[ORBFE_MAX_LEVELS + 1];
    __shared__ int s_w[AS_NT / 32];
    __shared__ int s_carry;
    const size_t f = blockIdx.x;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) {
        int run = 0;
        for (int l = 0; l < d.nlevels; l++) { s_start[l] = run; run += b.selCount[f * d.nlevels + l]; }
        s_start[d.nlevels] = run;
        s_carry = 0;
    }
    __syncthreads();
    const int n = s_start[d.nlevels];
    const float flap0 = (float)lap0, flap1 = (float)lap1;
    for (int j0 = 0; j0 < n; j0 += AS_NT) {
        const int j = j0 + tid;
        int l = 0, st = 0;
        float xs = 0.f, ys = 0.f, resp = 0.f;
        uint32_t x = 0, y = 0;
        if (j < n) {
            while (j >= s_start[l + 1]) l++;
            const uint32_t p = b.sel[f * d.kpCap + d.lv[l].selBase + (j - s_start[l])];
            x = (p & 0xFFF) + ORBFE_MIN_BORDER;
            y = ((p >> 12) & 0xFFF) + ORBFE_MIN_BORDER;
            resp = (float)(p >> 24);
            xs = (float)x;
            ys = (float)y;
            if (l) { xs = __fmul_rn(xs, d.lv[l].scale); ys = __fmul_rn(ys, d.lv[l].scale); }
            st = (xs >= flap0 && xs <= flap1) ? 1 : 0;
        }
        int incl = st;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_up_sync(0xFFFFFFFFu, incl, o);
            if (lane >= o) incl += t;
        }
        if (lane == 31) s_w[warp] = incl;
        __syncthreads();
        int before = s_carry + incl - st;
        for (int w = 0; w < warp; w++) before += s_w[w];
        if (j < n) {
            const int slot = st ? (n - 1 - before) : (j - before);
            orbfe_keypoint kp;
            kp.x = xs; kp.y = ys; kp.size = d.lv[l].kpSize; kp.angle = -1.f; kp.response = resp;
            kp.octave = l; kp.class_id = -1;
            b.kps[f * d.kpCap + slot] = kp;
            b.work[f * d.kpCap + j] = make_uint2(x | (y << 16), (uint32_t)l | ((uint32_t)slot << 8));
        }
        __syncthreads();
        if (tid == AS_NT - 1) s_carry = before + st;
        __syncthreads();
    }
    if (tid == 0) {
        b.count[f] = n;
        b.mono[f] = n - s_carry;
    }
}

/* --------------------------------------------------------------------------------- orientation + descriptor */
/* cv::fastAtan2: every operation rounded to binary32, no contraction */
static __device__ __forceinline__ float fast_atan2_deg(float y, float x) {
    const float s = (float)(180.0 / 3.14159265358979323846);
    const float p1 = __fmul_rn(0.9997878412794807f, s), p3 = __fmul_rn(-0.3258083974640975f, s);
    const float p5 = __fmul_rn(0.1555786518463281f, s), p7 = __fmul_rn(-0.04432655554792128f, s);
    const float eps = (float)2.2204460492503131e-16;
    const float ax = fabsf(x), ay = fabsf(y);
    float a, c, c2;
    if (ax >= ay) {
        c = __fdiv_rn(ay, __fadd_rn(ax, eps));
        c2 = __fmul_rn(c, c);
        a = __fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(p7, c2), p5), c2), p3), c2), p1), c);
    } else {
        c = __fdiv_rn(ax, __fadd_rn(ay, eps));
        c2 = __fmul_rn(c, c);
        a = __fsub_rn(90.f, __fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(p7, c2), p5), c2), p3), c2), p1), c));
    }
    if (x < 0) a = __fsub_rn(180.f, a);
    if (y < 0) a = __fsub_rn(360.f, a);
    return a;
}

#define OD_NT 256
#define OD_PW 10   /* words per row of the staged 37 x 37 descriptor box (3 bytes of alignment + 37) */
/* One warp = one keypoint.  IC_Angle: lane u-15 walks the 31 rows of the radius-15 disc on the unblurred level
 * (a per-lane row mask says which rows belong to the disc), moments reduced by shuffles.  rBRIEF: lane i builds
 * descriptor byte i from 16 rotated taps of the blurred level. */
/* eight CTAs per SM (32 registers): the kernel waits on memory most of the time, so the resident warps matter more than the dozen
 * spilled bytes (40 registers / six CTAs: 0.75 ms instead of 0.54 ms per 512 frames) */
#ifndef OD_MINB
#define OD_MINB 8
#endif
__global__ void __launch_bounds__(OD_NT, OD_MINB) k_orient_describe(const __grid_constant__ ExDesc d, ExBuffers b) {
    __shared__ __align__(16) float s_pat[1024];
    reinterpret_cast<float4*>(s_pat)[threadIdx.x] = __ldg(reinterpret_cast<const float4*>(g_patternf) + threadIdx.x);
    __syncthreads();
    const size_t f = blockIdx.y;
    const int lane = threadIdx.x & 31;
    const int j = blockIdx.x * (OD_NT / 32) + (threadIdx.x >> 5);
    if (j >= b.count[f]) return;
    const uint2 wk = b.work[f * d.kpCap + j];
    const int x = wk.x & 0xFFFF, y = wk.x >> 16;
    const int l = wk.y & 0xFF, slot = wk.y >> 8;
    const LevelDesc& L = d.lv[l];

    /* The 37 x 37 box of the blurred level that the rotated taps can reach (|rotated coordinate| <= 18) is staged in shared
     * memory with coalesced word loads (ten aligned words per row, issued here so that they are in flight during the angle
     * computation): 512 scattered byte gathers from global memory touch some twenty cache lines per request and saturate the
     * load/store unit; from shared memory the same gathers cost a few bank conflicts each. */
    __shared__ uint32_t s_patch[OD_NT / 32][37 * OD_PW];
    uint32_t* patch = s_patch[threadIdx.x >> 5];
    const int bpitch = L.bpitch;
    const int bx0 = (x - 18) & ~3, bdx = (x - 18) & 3;
    uint32_t pw[12];
    {
        const uint8_t* pb = b.blur + f * d.blurSlab + (uint32_t)(L.blurOff + (y - 18) * bpitch + bx0);
#pragma unroll
        for (int k = 0; k < 12; k++) {
            const int e = min(lane + 32 * k, 37 * OD_PW - 1), row = (e * 205) >> 11, word = e - row * OD_PW;   /* e / 10 for e < 1029 */
            pw[k] = __ldg(reinterpret_cast<const uint32_t*>(pb + row * bpitch) + word);
        }
    }
    /* IC_Angle, src/ORBextractor.cc:75-102: m10 = sum u*I, m01 = sum v*I over the disc.  The 31 x 31 box is read as aligned
     * words, four rows of eight words per step (a request covers four 32-byte runs) plus one step for the ninth word of every
     * row; each word meets its two weight words (u and v per byte, 0 outside the disc) in two DP4A. */
    const int pitch = L.pitch;
    const int xb = ORBFE_ROW_LEAD + x - 15, s = xb & 3;
    const uint8_t* p = b.pyr + f * d.pyrSlab + (uint32_t)(L.pyrOff + (y + ORBFE_EDGE - 15 + (lane >> 3)) * pitch + (xb - s) + 4 * (lane & 7));
    const uint2* wt = g_icw + s * (9 * 32) + lane;
    uint32_t px[9];
#pragma unroll
    for (int k = 0; k < 8; k++) px[k] = __ldg(reinterpret_cast<const uint32_t*>(p + k * 4 * pitch));   /* rows 4k + (lane >> 3); row 31 has weight 0 */
    px[8] = __ldg(reinterpret_cast<const uint32_t*>(b.pyr + f * d.pyrSlab + (uint32_t)(L.pyrOff + (y + ORBFE_EDGE - 15 + min(lane, 30)) * pitch + (xb - s) + 32)));
    int m10 = 0, m01 = 0;
#pragma unroll
    for (int k = 0; k < 9; k++) {
        const uint2 w = __ldg(wt + 32 * k);
        asm("dp4a.s32.u32 %0, %1, %2, %0;" : "+r"(m10) : "r"(w.x), "r"(px[k]));
        asm("dp4a.s32.u32 %0, %1, %2, %0;" : "+r"(m01) : "r"(w.y), "r"(px[k]));
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        m10 += __shfl_xor_sync(0xFFFFFFFFu, m10, o);
        m01 += __shfl_xor_sync(0xFFFFFFFFu, m01, o);
    }
    const float angle = fast_atan2_deg((float)m01, (float)m10);

    /* computeOrbDescriptor, src/ORBextractor.cc:106-145; a = cosf, b = sinf exactly as the host libm (orbfe_trig.cuh) */
    float ca, sa;
    orbfe_rotation(angle, &ca, &sa);
#pragma unroll
    for (int k = 0; k < 12; k++)
        if (lane + 32 * k < 37 * OD_PW) patch[lane + 32 * k] = pw[k];
    __syncwarp();
    const uint8_t* cb = reinterpret_cast<const uint8_t*>(patch) + 18 * (4 * OD_PW) + 18 + bdx;   /* the keypoint inside the staged box */
    const float4* pt = reinterpret_cast<const float4*>(s_pat) + lane;   /* [test pair k][lane]: conflict-free */
    int off[16];
#pragma unroll
    for (int k = 0; k < 8; k++) {
        const float4 t = pt[32 * k]; /* x0, y0, x1, y1 of tests 16*lane + 2k, 2k+1 */
        const int r0 = __float2int_rn(__fadd_rn(__fmul_rn(t.x, sa), __fmul_rn(t.y, ca)));
        const int q0 = __float2int_rn(__fsub_rn(__fmul_rn(t.x, ca), __fmul_rn(t.y, sa)));
        const int r1 = __float2int_rn(__fadd_rn(__fmul_rn(t.z, sa), __fmul_rn(t.w, ca)));
        const int q1 = __float2int_rn(__fsub_rn(__fmul_rn(t.z, ca), __fmul_rn(t.w, sa)));
        off[2 * k] = r0 * (4 * OD_PW) + q0;
        off[2 * k + 1] = r1 * (4 * OD_PW) + q1;
    }
    int tv[16];
#pragma unroll
    for (int k = 0; k < 16; k++) tv[k] = cb[off[k]];
    uint32_t val = 0;
#pragma unroll
    for (int k = 0; k < 8; k++) val |= (uint32_t)(tv[2 * k] < tv[2 * k + 1]) << k;
    b.desc[(f * d.kpCap + slot) * 32 + lane] = (uint8_t)val;
    if (lane == 0) b.kps[f * d.kpCap + slot].angle = angle;
}

/* --------------------------------------------------------------------------------------------- host driver */
int orbfe_upload_constants(const int* umax16, const int8_t* pattern1024) {
    float pf[1024];   /* device layout [k = 0..7][lane = 0..31][x0,y0,x1,y1]: descriptor byte `lane`, bit k */
    for (int lane = 0; lane < 32; lane++)
        for (int k = 0; k < 8; k++)
            for (int c = 0; c < 4; c++) pf[(k * 32 + lane) * 4 + c] = (float)pattern1024[(lane * 8 + k) * 4 + c];
    ORBFE_CUDA(cudaMemcpyToSymbol(c_umax, umax16, 16 * sizeof(int)));
    ORBFE_CUDA(cudaMemcpyToSymbol(g_patternf, pf, sizeof pf));
    static uint2 icw[4 * 9 * 32];
    for (int sft = 0; sft < 4; sft++)
        for (int k = 0; k < 9; k++)
            for (int lane = 0; lane < 32; lane++) {
                const int row = k < 8 ? 4 * k + (lane >> 3) : lane, word = k < 8 ? (lane & 7) : 8;
                uint32_t wu = 0, wv = 0;
                for (int bt = 0; bt < 4; bt++) {
                    const int u = 4 * word + bt - sft - 15, v = row - 15;
                    if (row > 30 || u < -15 || u > 15 || (u < 0 ? -u : u) > umax16[v < 0 ? -v : v]) continue;
                    wu |= (uint32_t)(uint8_t)(int8_t)u << (8 * bt);
                    wv |= (uint32_t)(uint8_t)(int8_t)v << (8 * bt);
                }
                icw[(sft * 9 + k) * 32 + lane] = make_uint2(wu, wv);
            }
    ORBFE_CUDA(cudaMemcpyToSymbol(g_icw, icw, sizeof icw));
    ORBFE_CUDA(orbfe_trig_init());
    return 0;
}

/* Host only, no device needed.  The 96 angles of orbfe_trig_exceptions.inc carry the cosf / sinf of glibc 2.39 (x86-64); a host whose
 * libm rounds those differently would see its own CPU build of the reference disagree with this library on a tap of such a keypoint.
 * Returns how many of the listed (angle, cosf, sinf) triples this host's libm does not reproduce (0 = the table is this host's). */
extern "C" int orbfe_check_host_libm(int* listed) {
    static const OrbfeTrigEx host_table[] = {
#include "orbfe_trig_exceptions.inc"
    };
    const float factorPI = (float)(3.14159265358979323846 / 180.0);   /* src/ORBextractor.cc:104 */
    int bad = 0, n = 0;
    for (const OrbfeTrigEx& e : host_table) {
        float angle, c, s;
        memcpy(&angle, &e.angle, 4);
        volatile float x = angle * factorPI;
        c = cosf(x); s = sinf(x);
        uint32_t cb, sb;
        memcpy(&cb, &c, 4); memcpy(&sb, &s, 4);
        bad += (cb != e.c || sb != e.s) ? 1 : 0;
        n++;
    }
    if (listed) *listed = n;
    return bad;
}

int orbfe_quadtree_smem_bytes(int cap) {
    /* box 2x8 + cnt 2x4 + childCnt 16 + scanA 4 + scanB 4 + sortK 8 + childPos 8 + newPos 2 + expand 1 */
    return cap * (16 + 8 + 16 + 4 + 4 + 8 + 8 + 2 + 1) + 64;
}

int orbfe_collect_stage_times(orbfe_extractor* ex);

int orbfe_blur_tiles(int w, int h) { return ((w + BLUR_TW - 1) / BLUR_TW) * ((h + BLUR_TH - 1) / BLUR_TH); }

/* The dynamic shared-memory limit is a property of the kernel on a device, not of an extractor object: extractors of
 * different sizes live side by side (Tracking holds mpIniORBextractor with 5 x nFeatures next to the per-frame one,
 * /root/reference/src/Tracking.cc:1151-1157), so the limit only ever grows to the largest request seen on the device. */
int orbfe_configure_kernels(int quadSmem, int fastSmem) {
    static std::mutex mu;
    static int quadMax[ORBFE_MAX_DEVICES], fastMax[ORBFE_MAX_DEVICES];
    int dev = 0;
    ORBFE_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= ORBFE_MAX_DEVICES) { orbfe_set_error("device ordinal %d out of range", dev); return ORBFE_E_INVALID; }
    std::lock_guard<std::mutex> lock(mu);
    static bool carved[ORBFE_MAX_DEVICES];
    if (!carved[dev]) {
        /* Shared-memory carve-out per kernel, in percent of the SM's 228 KB.  Left to the driver the split is chosen from what ran
         * before, and a step measured 3.34 or 3.49 ms depending on which call an object saw first; fixed, every kernel gets the
         * split it was fastest with in a sweep over 25 / 50 / 75 / 100 (the table is in DESIGN.md section 4) and kernels of the two
         * images that meet on an SM agree on it. */
        ORBFE_CUDA(cudaFuncSetAttribute(k_pyramid_copy0, cudaFuncAttributePreferredSharedMemoryCarveout, 25));
        ORBFE_CUDA(cudaFuncSetAttribute(k_pyramid_level, cudaFuncAttributePreferredSharedMemoryCarveout, 25));
        ORBFE_CUDA(cudaFuncSetAttribute(k_pyramid_fringe, cudaFuncAttributePreferredSharedMemoryCarveout, 25));
        ORBFE_CUDA(cudaFuncSetAttribute(k_fast_stream, cudaFuncAttributePreferredSharedMemoryCarveout, 50));
        ORBFE_CUDA(cudaFuncSetAttribute(k_fast_nms, cudaFuncAttributePreferredSharedMemoryCarveout, 100));
        ORBFE_CUDA(cudaFuncSetAttribute(k_fast_cells_list, cudaFuncAttributePreferredSharedMemoryCarveout, 100));
        ORBFE_CUDA(cudaFuncSetAttribute(k_blur7_tma, cudaFuncAttributePreferredSharedMemoryCarveout, 75));
        ORBFE_CUDA(cudaFuncSetAttribute(k_blur7, cudaFuncAttributePreferredSharedMemoryCarveout, 75));
        ORBFE_CUDA(cudaFuncSetAttribute(k_quadtree<false>, cudaFuncAttributePreferredSharedMemoryCarveout, 75));
        ORBFE_CUDA(cudaFuncSetAttribute(k_quadtree<true>, cudaFuncAttributePreferredSharedMemoryCarveout, 75));
        ORBFE_CUDA(cudaFuncSetAttribute(k_assemble, cudaFuncAttributePreferredSharedMemoryCarveout, 75));
        ORBFE_CUDA(cudaFuncSetAttribute(k_orient_describe, cudaFuncAttributePreferredSharedMemoryCarveout, 50));
        carved[dev] = true;
    }
    if (quadSmem > quadMax[dev]) {
        ORBFE_CUDA(cudaFuncSetAttribute(k_quadtree<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, quadSmem));
        quadMax[dev] = quadSmem;
    }
    if (fastSmem > fastMax[dev]) {
        ORBFE_CUDA(cudaFuncSetAttribute(k_fast_cells_list, cudaFuncAttributeMaxDynamicSharedMemorySize, fastSmem));
        fastMax[dev] = fastSmem;
    }
    return 0;
}

/* The per-frame buffers as seen by a launch whose frame 0 is the object's frame f0 (chunked host pipeline). */
ExBuffers orbfe_shift_frames(const ExBuffers& b, const ExDesc& d, int f0) {
    ExBuffers s = b;
    const size_t f = (size_t)f0;
    s.pyr += f * d.pyrSlab; s.blur += f * d.blurSlab;
    s.cand += f * d.candCap; s.cellCount += f * d.nCells; s.raw += f * (size_t)d.rawCap; s.rawCount += f * d.nCells;
    s.keys += f * d.candCap; s.keyNode += f * d.candCap;
    s.sel += f * d.kpCap; s.selCount += f * d.nlevels; s.work += f * d.kpCap;
    if (s.qtScratch) s.qtScratch += f * d.nlevels * (size_t)d.qtScratchBytes;
    s.kps += f * d.kpCap; s.desc += f * d.kpCap * 32; s.count += f; s.mono += f;
    s.uRight += f * d.kpCap; s.depth += f * d.kpCap; s.sadBest += f * d.kpCap;
    s.stTable += f * d.kpCap; s.stBand += f * (d.stBands + 1);
    return s;
}

/* Enqueue the whole extractor for `nframes` device-resident frames on `st`; results land at the object's frames
 * frame0 .. frame0+nframes-1 (frame0 != 0 only in the chunked host pipeline, which runs its chunks in stream order). */
int orbfe_launch_extract(orbfe_extractor* ex, const uint8_t* d_images, int nframes, int lap0, int lap1, cudaStream_t st, int frame0) {
    const ExDesc& d = ex->d;
    const ExBuffers xb = frame0 ? orbfe_shift_frames(ex->b, d, frame0) : ex->b;
    const bool prof = ex->profiling != 0 && frame0 == 0;
    int e = 0;
#define STAGE_MARK() do { if (prof) ORBFE_CUDA(cudaEventRecord(ex->ev[e++], st)); } while (0)
    if (prof && ex->stagePending) { int rc = orbfe_collect_stage_times(ex); if (rc) return rc; }
    STAGE_MARK();
    {
        const LevelDesc& L = d.lv[0];
        const int items = ((L.w + 15) >> 4) * L.h;
        k_pyramid_copy0<<<dim3((items + 255) / 256, nframes), 256, 0, st>>>(d, xb, d_images);
    }
    for (int l = 1; l < d.nlevels; l++) {
        const LevelDesc& L = d.lv[l];
        k_pyramid_level<<<dim3(((L.w + 3) / 4 + 31) / 32, (L.h + PYR_WARPS * PYR_ROWS - 1) / (PYR_WARPS * PYR_ROWS), nframes), dim3(32, PYR_WARPS), 0, st>>>(
            d, xb, ResizeArgs{L.w, L.h, L.pitch, L.pyrOff, d.lv[l - 1].pitch, d.lv[l - 1].pyrOff, L.tabX, L.tabY});
    }
    {
        int items = 0;
        for (int l = 0; l < d.nlevels; l++) {
            const LevelDesc& L = d.lv[l];
            const int n = L.h * (2 + ((L.w + FRINGE - 1) >> 2) - (L.w >> 2)) + 2 * FRINGE * (((L.w + FRINGE - 1) >> 2) + 2);
            items = items > n ? items : n;
        }
        k_pyramid_fringe<<<dim3((items + 127) / 128, d.nlevels, nframes), 128, 0, st>>>(d, xb);
    }
    if (frame0 == 0) ex->batchFrames = 0;
    ex->batchFrames = ex->batchFrames > frame0 + nframes ? ex->batchFrames : frame0 + nframes;
    ex->bordersValid = 0;
    STAGE_MARK();
    {
        const bool few = nframes < 8;   /* few frames: the 16-row cut puts three times as many warps in flight */
        const int sb = few ? d.nStrips : 0, sn = few ? d.nStripsShort : d.nStrips;
        k_fast_stream<<<dim3((sn + FS_WARPS - 1) / FS_WARPS, nframes), FS_WARPS * 32, 0, st>>>(d, xb, sb, sn);
    }
    k_fast_nms<<<dim3((d.nCells + FN_WARPS - 1) / FN_WARPS, nframes), FN_WARPS * 32, 0, st>>>(d, xb);
    k_fast_cells_list<<<148 * 8, FAST_NT, ex->fastSmem, st>>>(d, xb);
    STAGE_MARK();
    /* The blur needs only the pyramid, quadtree + assemble only the corners: they run side by side (the quadtree is a few long
     * latency-bound CTAs per frame, the blur fills the issue slots they leave empty) and meet again in front of k_orient_describe.
     * With stage profiling on everything stays on `st` so that the per-stage event pairs keep their meaning. */
    const bool fork = ex->forkBlur && !prof;
    const cudaStream_t bs = fork ? ex->side : st;
    if (fork) {
        ORBFE_CUDA(cudaEventRecord(ex->evFork, st));
        ORBFE_CUDA(cudaStreamWaitEvent(bs, ex->evFork, 0));
    }
    if (ex->blurLdg) k_blur7<<<dim3(d.blurTile0[d.nlevels], nframes), 256, 0, bs>>>(d, xb);
    else k_blur7_tma<<<dim3(d.blurTile0[d.nlevels], nframes), 256, 0, bs>>>(d, xb, ex->blurMaps, frame0);
    if (fork) ORBFE_CUDA(cudaEventRecord(ex->evJoin, bs));
    STAGE_MARK();
    if (d.qtScratchBytes) k_quadtree<true><<<dim3(d.nlevels, nframes), QT_NT, 0, st>>>(d, xb);
    else k_quadtree<false><<<dim3(d.nlevels, nframes), QT_NT, ex->quadSmem, st>>>(d, xb);
    STAGE_MARK();
    k_assemble<<<nframes, AS_NT, 0, st>>>(d, xb, lap0, lap1);
    STAGE_MARK();
    if (fork) ORBFE_CUDA(cudaStreamWaitEvent(st, ex->evJoin, 0));
    k_orient_describe<<<dim3((d.kpCap + OD_NT / 32 - 1) / (OD_NT / 32), nframes), OD_NT, 0, st>>>(d, xb);
    STAGE_MARK();
#undef STAGE_MARK
    if (prof) ex->stagePending = 1;
    ex->launches += d.nlevels + 8;
    ORBFE_CUDA(cudaGetLastError());
    return 0;
}

/* The full 19 px border of every level of the frames of the last batch (idempotent; cleared by the next extraction). */
int orbfe_materialize_borders(orbfe_extractor* ex, cudaStream_t st) {
    if (ex->bordersValid || ex->batchFrames <= 0) return 0;
    const ExDesc& d = ex->d;
    int maxBorder = 0;
    for (int l = 0; l < d.nlevels; l++) {
        const LevelDesc& L = d.lv[l];
        const int nInt = L.w >> 2, G = L.pitch >> 2, items = 2 * ORBFE_EDGE * G + L.h * (G - nInt);
        maxBorder = maxBorder > items ? maxBorder : items;
    }
    k_pyramid_borders<<<dim3((maxBorder + 255) / 256, d.nlevels, ex->batchFrames), 256, 0, st>>>(d, ex->b);
    ORBFE_CUDA(cudaGetLastError());
    ex->bordersValid = 1;
    return 0;
}

/* Fold the event pairs of the last profiled launch into stageMs (synchronises on the last event). */
int orbfe_collect_stage_times(orbfe_extractor* ex) {
    if (ex->stagePending) {
        ORBFE_CUDA(cudaEventSynchronize(ex->ev[ORBFE_NUM_STAGES]));
        for (int i = 0; i < ORBFE_NUM_STAGES; i++) {
            float ms = 0.f;
            ORBFE_CUDA(cudaEventElapsedTime(&ms, ex->ev[i], ex->ev[i + 1]));
            ex->stageMs[i] += ms;
        }
        ex->stagePending = 0;
    }
    if (ex->stereoPending) {
        float ms = 0.f;
        ORBFE_CUDA(cudaEventSynchronize(ex->ev[ORBFE_NUM_STAGES + 2]));
        ORBFE_CUDA(cudaEventElapsedTime(&ms, ex->ev[ORBFE_NUM_STAGES + 1], ex->ev[ORBFE_NUM_STAGES + 2]));
        ex->stageMs[ORBFE_NUM_STAGES] += ms;
        ex->stereoPending = 0;
    }
    return 0;
}
