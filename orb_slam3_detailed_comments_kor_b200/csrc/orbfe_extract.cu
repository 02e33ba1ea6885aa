/* orbfe_extract.cu — sm_100a kernels of the ORB extractor (replaces ORBextractor::operator(),
 * /root/reference/src/ORBextractor.cc:1068-1150, and everything it calls).
 *
 * Stage            kernel              reference lines (src/ORBextractor.cc)       roofline
 *  pyramid         k_pyramid_level     1152-1177 (+ cv::resize, copyMakeBorder)     HBM
 *  FAST per cell   k_fast_cells        763-854   (+ cv::FAST 9/16, NMS, 20/7 retry) HBM/ALU
 *  quadtree        k_quadtree          479-761                                      latency (L2 resident)
 *  ordering        k_assemble          856-872, 1100-1149                           latency
 *  blur            k_blur7             1114-1115 (+ cv::GaussianBlur 7x7 s=2)       HBM
 *  angle + rBRIEF  k_orient_describe   75-145, 470-477, 1059-1066                   L2 gathers
 *
 * All frames of a batch go through one launch per stage (gridDim.y / .z = frame); frames never interact.
 * Integer stages are bit-exact by construction; float stages use round-to-nearest intrinsics so that no FMA
 * contraction can change a bit relative to the oracle (built with -ffp-contract=off). */
#include <cuda_runtime.h>
#include <stdint.h>

#include "orbfe_internal.h"

__constant__ int c_umax[16];
__constant__ __align__(16) int8_t c_pattern[1024];

static __device__ __forceinline__ int reflect101(int p, int n) {
    /* |overshoot| <= 19 < n: one reflection is enough (cv::copyMakeBorder BORDER_REFLECT_101) */
    p = p < 0 ? -p : p;
    return p >= n ? 2 * (n - 1) - p : p;
}

/* ------------------------------------------------------------------------------------------------ pyramid */
/* One thread = one 16-byte chunk of one padded output row.  Border bytes are recomputed through the
 * reflected interior coordinate, so resize and copyMakeBorder are one pass with 128-bit stores only. */
__global__ void __launch_bounds__(256) k_pyramid_level(const __grid_constant__ ExDesc d, ExBuffers b,
                                                       const uint8_t* __restrict__ input, int level) {
    const LevelDesc& L = d.lv[level];
    const int chunks = L.pitch >> 4;
    const int rows = L.h + 2 * ORBFE_EDGE;
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= chunks * rows) return;
    const int row = idx / chunks, chunk = idx - row * chunks;
    const size_t f = blockIdx.y;
    uint8_t* slab = b.pyr + f * d.pyrSlab;
    const int yi = reflect101(row - ORBFE_EDGE, L.h);
    uint32_t out[4] = {0, 0, 0, 0};
    const int px0 = chunk * 16 - ORBFE_ROW_LEAD;
    if (level == 0) {
        const uint8_t* src = input + f * (size_t)d.W * d.H + (size_t)yi * d.W;
#pragma unroll
        for (int k = 0; k < 16; k++) {
            const int px = px0 + k;
            uint32_t v = 0;
            if (px >= -ORBFE_EDGE && px < L.w + ORBFE_EDGE) v = __ldg(src + reflect101(px, L.w));
            out[k >> 2] |= v << (8 * (k & 3));
        }
    } else {
        const LevelDesc& P = d.lv[level - 1];
        const uint8_t* sint = slab + P.pyrOff + (size_t)ORBFE_EDGE * P.pitch + ORBFE_ROW_LEAD;
        const uint2 yt = __ldg(b.tabs + L.tabY + yi);
        /* row yt.x+1 and column sx+1 may be the source's own border: their coefficient is then 0 */
        const uint8_t* r0 = sint + (size_t)yt.x * P.pitch;
        const uint8_t* r1 = r0 + P.pitch;
        const int b0 = yt.y & 0xFFFF, b1 = yt.y >> 16;
#pragma unroll
        for (int k = 0; k < 16; k++) {
            const int px = px0 + k;
            uint32_t v = 0;
            if (px >= -ORBFE_EDGE && px < L.w + ORBFE_EDGE) {
                const uint2 xt = __ldg(b.tabs + L.tabX + reflect101(px, L.w));
                const int c0 = xt.y & 0xFFFF, c1 = xt.y >> 16;
                const int h0 = r0[xt.x] * c0 + r0[xt.x + 1] * c1;
                const int h1 = r1[xt.x] * c0 + r1[xt.x + 1] * c1;
                v = (uint32_t)((((b0 * (h0 >> 4)) >> 16) + ((b1 * (h1 >> 4)) >> 16) + 2) >> 2);
            }
            out[k >> 2] |= v << (8 * (k & 3));
        }
    }
    *reinterpret_cast<uint4*>(slab + L.pyrOff + (size_t)row * L.pitch + chunk * 16) = make_uint4(out[0], out[1], out[2], out[3]);
}

/* --------------------------------------------------------------------------------------------------- blur */
/* 7x7 sigma-2 Gaussian as OpenCV's 8U fixed point: taps {18,34,48,56,48,34,18}/256 horizontally (Q8.8, kept
 * exact in 16 bits), the same vertically (Q16.16), + 2^15 >> 16.  The padded pyramid already holds the
 * REFLECT_101 border, so the tile load needs no edge logic. */
#define BLUR_TW 64
#define BLUR_TH 32
__global__ void __launch_bounds__(256) k_blur7(const __grid_constant__ ExDesc d, ExBuffers b, int level) {
    __shared__ uint32_t s_src[BLUR_TH + 6][18];
    __shared__ __align__(8) uint16_t s_h[BLUR_TH + 6][BLUR_TW];
    const LevelDesc& L = d.lv[level];
    const size_t f = blockIdx.z;
    const int x0 = blockIdx.x * BLUR_TW, y0 = blockIdx.y * BLUR_TH;
    const uint8_t* pad = b.pyr + f * d.pyrSlab + L.pyrOff;
    for (int i = threadIdx.x; i < (BLUR_TH + 6) * 18; i += 256) {
        const int r = i / 18, c = i - r * 18;
        int gy = y0 - 3 + r;
        gy = min(gy, L.h + ORBFE_EDGE - 1);
        const uint8_t* p = pad + (size_t)(gy + ORBFE_EDGE) * L.pitch + ORBFE_ROW_LEAD + x0 - 4 + 4 * c;
        s_src[r][c] = *reinterpret_cast<const uint32_t*>(p);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < (BLUR_TH + 6) * 16; i += 256) {
        const int r = i >> 4, q = i & 15;
        const uint32_t w0 = s_src[r][q], w1 = s_src[r][q + 1], w2 = s_src[r][q + 2];
        int v[10];
        v[0] = (w0 >> 8) & 0xFF; v[1] = (w0 >> 16) & 0xFF; v[2] = w0 >> 24;
        v[3] = w1 & 0xFF; v[4] = (w1 >> 8) & 0xFF; v[5] = (w1 >> 16) & 0xFF; v[6] = w1 >> 24;
        v[7] = w2 & 0xFF; v[8] = (w2 >> 8) & 0xFF; v[9] = (w2 >> 16) & 0xFF;
        uint32_t o[4];
#pragma unroll
        for (int j = 0; j < 4; j++)
            o[j] = 18 * (v[j] + v[j + 6]) + 34 * (v[j + 1] + v[j + 5]) + 48 * (v[j + 2] + v[j + 4]) + 56 * v[j + 3];
        *reinterpret_cast<uint2*>(&s_h[r][4 * q]) = make_uint2(o[0] | (o[1] << 16), o[2] | (o[3] << 16));
    }
    __syncthreads();
    uint8_t* dst = b.blur + f * d.blurSlab + L.blurOff;
    for (int i = threadIdx.x; i < BLUR_TH * 16; i += 256) {
        const int r = i >> 4, q = i & 15;
        const int gy = y0 + r, gx = x0 + 4 * q;
        if (gy >= L.h || gx >= L.bpitch) continue;
        uint32_t acc[4] = {0, 0, 0, 0};
        const int K[7] = {18, 34, 48, 56, 48, 34, 18};
#pragma unroll
        for (int k = 0; k < 7; k++) {
            const uint2 hv = *reinterpret_cast<const uint2*>(&s_h[r + k][4 * q]);
            acc[0] += K[k] * (hv.x & 0xFFFF); acc[1] += K[k] * (hv.x >> 16);
            acc[2] += K[k] * (hv.y & 0xFFFF); acc[3] += K[k] * (hv.y >> 16);
        }
        const uint32_t o = ((acc[0] + 32768) >> 16) | (((acc[1] + 32768) >> 16) << 8) |
                           (((acc[2] + 32768) >> 16) << 16) | (((acc[3] + 32768) >> 16) << 24);
        *reinterpret_cast<uint32_t*>(dst + (size_t)gy * L.bpitch + gx) = o;
    }
}

/* --------------------------------------------------------------------------------------------------- FAST */
#define FAST_TP 80   /* smem tile pitch: ROI width <= 76 (wCell <= 70, +6) + 3 bytes of alignment lead */
#define FAST_RH 76
#define FAST_NT 256
/* ring offsets of cv::FAST 9/16: (dx,dy) for k = 0..15 */
#define RING_OFF(k, TP)                                                                              \
    ((k) == 0 ? 3 * TP : (k) == 1 ? 3 * TP + 1 : (k) == 2 ? 2 * TP + 2 : (k) == 3 ? TP + 3 : (k) == 4 ? 3 : \
     (k) == 5 ? -TP + 3 : (k) == 6 ? -2 * TP + 2 : (k) == 7 ? -3 * TP + 1 : (k) == 8 ? -3 * TP :           \
     (k) == 9 ? -3 * TP - 1 : (k) == 10 ? -2 * TP - 2 : (k) == 11 ? -TP - 3 : (k) == 12 ? -3 :              \
     (k) == 13 ? TP - 3 : (k) == 14 ? 2 * TP - 2 : 3 * TP - 1)

static __host__ __device__ __forceinline__ bool run9(uint32_t m) {
    m |= m << 16;
    uint32_t x = m & (m >> 1);
    x &= x >> 2;
    x &= x >> 4;
    x &= m >> 8;
    return (x & 0xFFFFu) != 0;
}

/* corner at threshold t  <=>  9 contiguous ring pixels all > v+t or all < v-t */
static __host__ __device__ __forceinline__ bool fast_is_corner(const uint8_t* c, int t) {
    const int v = c[0];
    const int hi = v + t, lo = v - t;
    const int p0 = c[RING_OFF(0, FAST_TP)], p8 = c[RING_OFF(8, FAST_TP)];
    const int p4 = c[RING_OFF(4, FAST_TP)], p12 = c[RING_OFF(12, FAST_TP)];
    /* any 9-arc contains one pixel of every opposite pair (k, k+8) */
    const bool mayB = (p0 > hi || p8 > hi) && (p4 > hi || p12 > hi);
    const bool mayD = (p0 < lo || p8 < lo) && (p4 < lo || p12 < lo);
    if (!(mayB || mayD)) return false;
    uint32_t B = 0, D = 0;
#pragma unroll
    for (int k = 0; k < 16; k++) {
        const int p = c[RING_OFF(k, FAST_TP)];
        B |= (uint32_t)(p > hi) << k;
        D |= (uint32_t)(p < lo) << k;
    }
    return run9(B) || run9(D);
}

/* cv::FAST cornerScore = m - 1 with m = max over the 16 arcs of max(min_k (v - r_k), min_k (r_k - v)).
 * Evaluated on the raw ring values: min_k (v - r_k) = v - max_k r_k and min_k (r_k - v) = min_k r_k - v, so
 * m = max(v - min_arcs(max_arc r), max_arcs(min_arc r) - v).  Besides being cheaper this keeps every min/max
 * operand a plain non-negated register: nvcc 12.9 folds a negation into VIMNMX3 on sm_100a and the result
 * loses the sign (found by tools/fast_device_test3.cu: max(m, max(lo, -hi)) returned max(m, lo, hi)). */
static __host__ __device__ __forceinline__ int fast_score(const uint8_t* c) {
    const int v = c[0];
    int r[16];
#pragma unroll
    for (int k = 0; k < 16; k++) r[k] = (int)c[RING_OFF(k, FAST_TP)];
    int lo2[16], hi2[16], lo4[16], hi4[16];
#pragma unroll
    for (int k = 0; k < 16; k++) { lo2[k] = min(r[k], r[(k + 1) & 15]); hi2[k] = max(r[k], r[(k + 1) & 15]); }
#pragma unroll
    for (int k = 0; k < 16; k++) { lo4[k] = min(lo2[k], lo2[(k + 2) & 15]); hi4[k] = max(hi2[k], hi2[(k + 2) & 15]); }
    int minOfMax = 255, maxOfMin = 0;
#pragma unroll
    for (int k = 0; k < 16; k++) {
        const int lo9 = min(min(lo4[k], lo4[(k + 4) & 15]), r[(k + 8) & 15]);
        const int hi9 = max(max(hi4[k], hi4[(k + 4) & 15]), r[(k + 8) & 15]);
        minOfMax = min(minOfMax, hi9);
        maxOfMin = max(maxOfMin, lo9);
    }
    const int dark = v - minOfMax, bright = maxOfMin - v;
    return (dark > bright ? dark : bright) - 1;
}

/* One CTA = one 35-px-grid cell of one level of one frame: cv::FAST(roi, iniTh) -> if empty cv::FAST(roi, minTh),
 * with NMS seeing only this cell's scores at the current threshold (SURVEY.md §7 hard part 2).
 * Output: the cell's keypoints in (y, x) order, packed, at the cell's slot range. */
__global__ void __launch_bounds__(FAST_NT) k_fast_cells(const __grid_constant__ ExDesc d, ExBuffers b) {
    __shared__ __align__(16) uint8_t s_img[FAST_RH * FAST_TP];
    __shared__ __align__(16) uint8_t s_score[FAST_RH * FAST_TP];
    __shared__ uint16_t s_list[70 * 70];
    __shared__ uint32_t s_mask[(70 * 70 + 31) / 32 + 1];
    __shared__ int s_nlist, s_total;
    __shared__ int s_wsum[FAST_NT / 32];

    const CellDesc cell = b.cells[blockIdx.x];
    const size_t f = blockIdx.y;
    int* countOut = b.cellCount + f * d.nCells + blockIdx.x;
    const int cw = cell.rw - 6, ch = cell.rh - 6;
    if (cw <= 0 || ch <= 0) {
        if (threadIdx.x == 0) *countOut = 0;
        return;
    }
    const LevelDesc& L = d.lv[cell.level];
    const int ox = cell.x0 & 3;
    {
        const uint8_t* g = b.pyr + f * d.pyrSlab + L.pyrOff + (size_t)(cell.y0 + ORBFE_EDGE) * L.pitch + ORBFE_ROW_LEAD + (cell.x0 - ox);
        const int nw = (ox + cell.rw + 3) >> 2;
        for (int i = threadIdx.x; i < cell.rh * nw; i += FAST_NT) {
            const int r = i / nw, c = i - r * nw;
            reinterpret_cast<uint32_t*>(s_img + r * FAST_TP)[c] = *reinterpret_cast<const uint32_t*>(g + (size_t)r * L.pitch + 4 * c);
        }
        for (int i = threadIdx.x; i < FAST_RH * FAST_TP / 4; i += FAST_NT) reinterpret_cast<uint32_t*>(s_score)[i] = 0;
    }
    const int npix = cw * ch;
    const int nwords = (npix + 31) >> 5;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int th = d.iniTh;
    for (int pass = 0; pass < 2; pass++) {
        if (threadIdx.x == 0) s_nlist = 0;
        for (int i = threadIdx.x; i < nwords; i += FAST_NT) s_mask[i] = 0;
        __syncthreads();
        /* 1. threshold test on every candidate pixel; corners are queued so the expensive score is dense work */
        for (int p0 = 0; p0 < npix; p0 += FAST_NT) {
            const int p = p0 + threadIdx.x;
            bool isc = false;
            if (p < npix) {
                const int py = p / cw, px = p - py * cw;
                isc = fast_is_corner(s_img + (py + 3) * FAST_TP + ox + px + 3, th);
            }
            const uint32_t bal = __ballot_sync(0xFFFFFFFFu, isc);
            if (bal) {
                int base = 0;
                if (lane == 0) base = atomicAdd(&s_nlist, __popc(bal));
                base = __shfl_sync(0xFFFFFFFFu, base, 0);
                if (isc) s_list[base + __popc(bal & ((1u << lane) - 1))] = (uint16_t)p;
            }
        }
        __syncthreads();
        const int nl = s_nlist;
        /* 2. scores */
        for (int e = threadIdx.x; e < nl; e += FAST_NT) {
            const int p = s_list[e];
            const int py = p / cw, px = p - py * cw;
            const int o = (py + 3) * FAST_TP + ox + px + 3;
            s_score[o] = (uint8_t)fast_score(s_img + o);
        }
        __syncthreads();
        /* 3. non-maximum suppression: strictly greater than the 8 neighbours (non-corners count 0) */
        for (int e = threadIdx.x; e < nl; e += FAST_NT) {
            const int p = s_list[e];
            const int py = p / cw, px = p - py * cw;
            const uint8_t* s = s_score + (py + 3) * FAST_TP + ox + px + 3;
            const int v = s[0];
            const int nb = max(max(max(s[-FAST_TP - 1], s[-FAST_TP]), max(s[-FAST_TP + 1], s[-1])),
                               max(max(s[1], s[FAST_TP - 1]), max(s[FAST_TP], s[FAST_TP + 1])));
            if (v > nb) atomicOr(&s_mask[p >> 5], 1u << (p & 31));
        }
        __syncthreads();
        /* 4. ordered count: exclusive prefix over the mask words (row-major = (y,x) order) */
        int mine = 0;
        for (int i = threadIdx.x; i < nwords; i += FAST_NT) mine += __popc(s_mask[i]);
        /* nwords <= 154 < FAST_NT, so each thread owns at most one word: prefix over threads */
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
        if (total > 0 || pass == 1) {
            if (threadIdx.x < nwords) {
                uint32_t m = s_mask[threadIdx.x];
                int k = wbase + incl - mine;
                uint32_t* out = b.cand + f * d.candCap + cell.candBase;
                while (m) {
                    const int bit = __ffs(m) - 1;
                    m &= m - 1;
                    const int p = threadIdx.x * 32 + bit;
                    const int py = p / cw, px = p - py * cw;
                    const int sc = s_score[(py + 3) * FAST_TP + ox + px + 3];
                    out[k++] = (uint32_t)(px + 3 + cell.xoff) | ((uint32_t)(py + 3 + cell.yoff) << 12) | ((uint32_t)sc << 24);
                }
            }
            if (threadIdx.x == 0) *countOut = total;
            return;
        }
        /* empty at iniTh: wipe the scores written so far and retry with minTh (src/ORBextractor.cc:825-828) */
        for (int e = threadIdx.x; e < nl; e += FAST_NT) {
            const int p = s_list[e];
            const int py = p / cw, px = p - py * cw;
            s_score[(py + 3) * FAST_TP + ox + px + 3] = 0;
        }
        th = d.minTh;
        __syncthreads();
    }
}

/* ----------------------------------------------------------------------------------------------- quadtree */
#define QT_NT 512

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
__global__ void __launch_bounds__(QT_NT) k_quadtree(const __grid_constant__ ExDesc d, ExBuffers b) {
    extern __shared__ __align__(16) uint8_t smem[];
    const int cap = d.nodeCap;
    short4* box0 = reinterpret_cast<short4*>(smem);
    short4* box1 = box0 + cap;
    int* cnt0 = reinterpret_cast<int*>(box1 + cap);
    int* cnt1 = cnt0 + cap;
    int* childCnt = cnt1 + cap;            /* [cap][4] */
    int* scanA = childCnt + 4 * cap;       /* [cap] S: children before me in processing order */
    int* scanB = scanA + cap;              /* [cap] unexpanded-before-me */
    uint32_t* sortK = reinterpret_cast<uint32_t*>(scanB + cap); /* [cap] */
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
        /* scanA doubles as the per-cell prefix when ncell <= cap, otherwise cells are walked serially */
        int* pre = scanA;
        if (ncell <= cap) {
            for (int i = tid; i < ncell; i += QT_NT) pre[i] = cellCount[i];
            __syncthreads();
            nKeys = block_exscan(pre, ncell, s_tmp);
            const uint32_t* cand = b.cand + f * d.candCap;
            const int warp = tid >> 5, lane = tid & 31;
            for (int c = warp; c < ncell; c += QT_NT / 32) {
                const int n = cellCount[c];
                const uint32_t* src = cand + b.cells[L.cellBase + c].candBase;
                for (int i = lane; i < n; i += 32) keys[pre[c] + i] = src[i];
            }
        } else {
            if (tid == 0) {
                int run = 0;
                const uint32_t* cand = b.cand + f * d.candCap;
                for (int c = 0; c < ncell; c++) {
                    const int n = cellCount[c];
                    const uint32_t* src = cand + b.cells[L.cellBase + c].candBase;
                    for (int i = 0; i < n; i++) keys[run + i] = src[i];
                    run += n;
                }
                s_nexp = run;
            }
            __syncthreads();
            nKeys = s_nexp;
        }
        __syncthreads();
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
                sortK[i] = (i < nNew && expand[i]) ? (((uint32_t)(524287 - min(cnt[i], 524287)) << 13) | (uint32_t)i) : 0xFFFFFFFFu;
            __syncthreads();
            for (int k2 = 2; k2 <= n2; k2 <<= 1)
                for (int j = k2 >> 1; j > 0; j >>= 1) {
                    for (int i = tid; i < n2; i += QT_NT) {
                        const int ixj = i ^ j;
                        if (ixj > i) {
                            const uint32_t a = sortK[i], c = sortK[ixj];
                            const bool up = (i & k2) == 0;
                            if ((a > c) == up) { sortK[i] = c; sortK[ixj] = a; }
                        }
                    }
                    __syncthreads();
                }
            /* scanB[j] = list growth when the j-th candidate is divided (children - 1) */
            for (int j = tid; j < nNew; j += QT_NT) {
                int g = 0;
                if (sortK[j] != 0xFFFFFFFFu) {
                    const int i = sortK[j] & 0x1FFF;
                    g = (childCnt[4 * i] > 0) + (childCnt[4 * i + 1] > 0) + (childCnt[4 * i + 2] > 0) + (childCnt[4 * i + 3] > 0) - 1;
                }
                scanB[j] = g;
            }
            if (tid == 0) s_cut = nNew;
            __syncthreads();
            block_exscan(scanB, nNew, s_tmp);
            /* first j whose inclusive growth brings the list to N: everything after it is left alone */
            for (int j = tid; j < nNew; j += QT_NT) {
                if (sortK[j] == 0xFFFFFFFFu) continue;
                const int i = sortK[j] & 0x1FFF;
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
                if (sortK[j] != 0xFFFFFFFFu && j <= cut) {
                    const int i = sortK[j] & 0x1FFF;
                    expand[i] = 1;
                    nch = (childCnt[4 * i] > 0) + (childCnt[4 * i + 1] > 0) + (childCnt[4 * i + 2] > 0) + (childCnt[4 * i + 3] > 0);
                }
                scanB[j] = nch;
            }
            __syncthreads();
            C = block_exscan(scanB, nNew, s_tmp);
            for (int j = tid; j < nNew; j += QT_NT)
                if (sortK[j] != 0xFFFFFFFFu && j <= cut) scanA[sortK[j] & 0x1FFF] = scanB[j];
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
/* One warp = one keypoint.  IC_Angle: lane u-15 walks the 31 rows of the radius-15 disc on the unblurred level,
 * moments reduced by shuffles.  rBRIEF: lane i builds descriptor byte i from 16 rotated taps of the blurred level. */
__global__ void __launch_bounds__(OD_NT) k_orient_describe(const __grid_constant__ ExDesc d, ExBuffers b) {
    __shared__ int8_t s_pat[1024];
    reinterpret_cast<uint32_t*>(s_pat)[threadIdx.x] = reinterpret_cast<const uint32_t*>(c_pattern)[threadIdx.x];
    __syncthreads();
    const size_t f = blockIdx.y;
    const int lane = threadIdx.x & 31;
    const int j = blockIdx.x * (OD_NT / 32) + (threadIdx.x >> 5);
    if (j >= b.count[f]) return;
    const uint2 wk = b.work[f * d.kpCap + j];
    const int x = wk.x & 0xFFFF, y = wk.x >> 16;
    const int l = wk.y & 0xFF, slot = wk.y >> 8;
    const LevelDesc& L = d.lv[l];

    /* IC_Angle, src/ORBextractor.cc:75-102 */
    const uint8_t* c = b.pyr + f * d.pyrSlab + L.pyrOff + (size_t)(y + ORBFE_EDGE) * L.pitch + ORBFE_ROW_LEAD + x;
    int m10 = 0, m01 = 0;
    const int u = lane - 15;
    if (lane < 31) {
        const int au = u < 0 ? -u : u;
#pragma unroll 1
        for (int v = -15; v <= 15; v++) {
            const int av = v < 0 ? -v : v;
            if (au <= c_umax[av]) {
                const int val = c[v * L.pitch + u];
                m10 += u * val;
                m01 += v * val;
            }
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        m10 += __shfl_xor_sync(0xFFFFFFFFu, m10, o);
        m01 += __shfl_xor_sync(0xFFFFFFFFu, m01, o);
    }
    const float angle = fast_atan2_deg((float)m01, (float)m10);

    /* computeOrbDescriptor, src/ORBextractor.cc:106-145.  cosf/sinf of the host libm are reproduced by the
     * correctly-rounded value (double evaluation rounded once); see DESIGN.md "float reproducibility". */
    const float factorPI = (float)(3.14159265358979323846 / 180.0);
    const float ang = __fmul_rn(angle, factorPI);
    const float ca = (float)cos((double)ang), sa = (float)sin((double)ang);
    const uint8_t* cb = b.blur + f * d.blurSlab + L.blurOff + (size_t)y * L.bpitch + x;
    const int8_t* p = s_pat + 32 * lane;
    uint32_t val = 0;
#pragma unroll
    for (int k = 0; k < 8; k++) {
        const float x0 = (float)p[4 * k], y0 = (float)p[4 * k + 1], x1 = (float)p[4 * k + 2], y1 = (float)p[4 * k + 3];
        const int r0 = __float2int_rn(__fadd_rn(__fmul_rn(x0, sa), __fmul_rn(y0, ca)));
        const int q0 = __float2int_rn(__fsub_rn(__fmul_rn(x0, ca), __fmul_rn(y0, sa)));
        const int r1 = __float2int_rn(__fadd_rn(__fmul_rn(x1, sa), __fmul_rn(y1, ca)));
        const int q1 = __float2int_rn(__fsub_rn(__fmul_rn(x1, ca), __fmul_rn(y1, sa)));
        const int t0 = cb[r0 * L.bpitch + q0], t1 = cb[r1 * L.bpitch + q1];
        val |= (uint32_t)(t0 < t1) << k;
    }
    b.desc[(f * d.kpCap + slot) * 32 + lane] = (uint8_t)val;
    if (lane == 0) b.kps[f * d.kpCap + slot].angle = angle;
}

/* --------------------------------------------------------------------------------------------- host driver */
int orbfe_upload_constants(const int* umax16, const int8_t* pattern1024) {
    ORBFE_CUDA(cudaMemcpyToSymbol(c_umax, umax16, 16 * sizeof(int)));
    ORBFE_CUDA(cudaMemcpyToSymbol(c_pattern, pattern1024, 1024));
    return 0;
}

int orbfe_quadtree_smem_bytes(int cap) {
    /* box 2x8 + cnt 2x4 + childCnt 16 + scanA 4 + scanB 4 + sortK 4 + childPos 8 + newPos 2 + expand 1 */
    return cap * (16 + 8 + 16 + 4 + 4 + 4 + 8 + 2 + 1) + 64;
}

int orbfe_collect_stage_times(orbfe_extractor* ex);

int orbfe_configure_kernels(int quadSmem) {
    ORBFE_CUDA(cudaFuncSetAttribute(k_quadtree, cudaFuncAttributeMaxDynamicSharedMemorySize, quadSmem));
    return 0;
}

/* Enqueue the whole extractor for `nframes` device-resident frames on `st`. */
int orbfe_launch_extract(orbfe_extractor* ex, const uint8_t* d_images, int nframes, int lap0, int lap1, cudaStream_t st) {
    const ExDesc& d = ex->d;
    const bool prof = ex->profiling != 0;
    int e = 0;
#define STAGE_MARK() do { if (prof) ORBFE_CUDA(cudaEventRecord(ex->ev[e++], st)); } while (0)
    if (prof && ex->stagePending) { int rc = orbfe_collect_stage_times(ex); if (rc) return rc; }
    STAGE_MARK();
    for (int l = 0; l < d.nlevels; l++) {
        const LevelDesc& L = d.lv[l];
        const int items = (L.pitch >> 4) * (L.h + 2 * ORBFE_EDGE);
        k_pyramid_level<<<dim3((items + 255) / 256, nframes), 256, 0, st>>>(d, ex->b, d_images, l);
    }
    STAGE_MARK();
    k_fast_cells<<<dim3(d.nCells, nframes), FAST_NT, 0, st>>>(d, ex->b);
    STAGE_MARK();
    for (int l = 0; l < d.nlevels; l++) {
        const LevelDesc& L = d.lv[l];
        k_blur7<<<dim3((L.w + BLUR_TW - 1) / BLUR_TW, (L.h + BLUR_TH - 1) / BLUR_TH, nframes), 256, 0, st>>>(d, ex->b, l);
    }
    STAGE_MARK();
    k_quadtree<<<dim3(d.nlevels, nframes), QT_NT, ex->quadSmem, st>>>(d, ex->b);
    STAGE_MARK();
    k_assemble<<<nframes, AS_NT, 0, st>>>(d, ex->b, lap0, lap1);
    STAGE_MARK();
    k_orient_describe<<<dim3((d.kpCap + OD_NT / 32 - 1) / (OD_NT / 32), nframes), OD_NT, 0, st>>>(d, ex->b);
    STAGE_MARK();
#undef STAGE_MARK
    if (prof) ex->stagePending = 1;
    ex->launches += 2 * d.nlevels + 4;
    ORBFE_CUDA(cudaGetLastError());
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
