/* orbfe_match.cu — sm_100a kernels + C ABI of the descriptor-matching core.
 *
 *  k_top2_partial / k_top2_merge   top-2 Hamming search, /root/reference/src/ORBmatcher.cc:311-375, 870-908
 *                                  (distance = ORBmatcher::DescriptorDistance, :2591-2607); also serves
 *                                  BFMatcher::knnMatch(k=2) + Lowe test of src/Frame.cc:1137-1144
 *  k_stereo_match / k_stereo_median Frame::ComputeStereoMatches, src/Frame.cc:797-967
 *
 * The Hamming work is XOR + POPC on the integer pipe (8 x 32-bit per 256-bit pair); it is deliberately not
 * reshaped into a tensor-core contraction (BASELINE.json north_star). */
#include <cuda_runtime.h>
#include <dlfcn.h>
#include <limits.h>
#include <stdint.h>
#include <string.h>

#include <algorithm>
#include <mutex>
#include <new>

#include "orbfe_internal.h"
#include "orbfe_kb8.cuh"
#include "orbfe_tma.cuh"

int orbfe_collect_stage_times(orbfe_extractor* ex);

/* ------------------------------------------------------------------------------------------ top-2 search */
#define T2_NT 256
#define T2_QPT 4                    /* queries held in registers per thread */
#define T2_QTILE (T2_NT * T2_QPT)   /* queries per CTA */
#define T2_CHUNK 256                /* database rows per shared-memory stage (8 KB) */

/* 256-bit Hamming distance = DescriptorDistance (src/ORBmatcher.cc:2591-2607).  POPC issues on the XU pipe at a
 * quarter of the logic rate; LOP3/IADD go to the ALU pipe.  ncu on the 4096 x 2^20 search: eight plain population
 * counts -> XU pipe 97 % busy, ALU half idle (5.58e11 pairs/s); a full carry-save tree (4 POPC + 14 LOP3) -> ALU pipe
 * 94 % busy, XU 49 % (5.61e11); two full adders in front of six population counts balance the two pipes (6.78e11). */
static __device__ __forceinline__ int ham256(const uint32_t (&q)[8], const uint4 a, const uint4 b) {
    const uint32_t x0 = q[0] ^ a.x, x1 = q[1] ^ a.y, x2 = q[2] ^ a.z, x3 = q[3] ^ a.w;
    const uint32_t x4 = q[4] ^ b.x, x5 = q[5] ^ b.y, x6 = q[6] ^ b.z, x7 = q[7] ^ b.w;
    const uint32_t s0 = x0 ^ x1 ^ x2, c0 = (x0 & x1) | (x2 & (x0 ^ x1));   /* x0 + x1 + x2 = s0 + 2*c0, bitwise */
    const uint32_t s1 = x3 ^ x4 ^ x5, c1 = (x3 & x4) | (x5 & (x3 ^ x4));
    return __popc(s0) + __popc(s1) + __popc(x6) + __popc(x7) + 2 * (__popc(c0) + __popc(c1));
}


/* grid = (query tiles, database slices).  Each thread keeps T2_QPT queries in registers and scans its CTA's
 * slice, streamed through a double-buffered bulk-copy (TMA) stage that the whole CTA reads by broadcast.
 * Scan order inside a thread is ascending row, so strict `<` keeps the lowest row on ties, exactly like the
 * reference's sequential loop.  Output: one orbfe_match per (slice, query) in `parts` (idx is global). */
__global__ void __launch_bounds__(T2_NT) k_top2_partial(const uint4* __restrict__ q, int nq, const uint4* __restrict__ db,
                                                        long long ndb, long long indexBase, int rowsPerSlice, int initDist,
                                                        orbfe_match* __restrict__ parts) {
    __shared__ __align__(16) uint4 s_db[2][T2_CHUNK * 2];
    const int slice = blockIdx.y;
    const long long r0 = (long long)slice * rowsPerSlice;
    const int nrows = (int)max(0LL, min((long long)rowsPerSlice, ndb - r0));
    uint32_t qr[T2_QPT][8];
    int b1[T2_QPT], b2[T2_QPT], bi[T2_QPT];
#pragma unroll
    for (int i = 0; i < T2_QPT; i++) {
        const int qi = blockIdx.x * T2_QTILE + i * T2_NT + threadIdx.x;
        uint4 a = make_uint4(0, 0, 0, 0), b = a;
        if (qi < nq) { a = __ldg(q + 2 * (size_t)qi); b = __ldg(q + 2 * (size_t)qi + 1); }
        qr[i][0] = a.x; qr[i][1] = a.y; qr[i][2] = a.z; qr[i][3] = a.w;
        qr[i][4] = b.x; qr[i][5] = b.y; qr[i][6] = b.z; qr[i][7] = b.w;
        b1[i] = initDist; b2[i] = initDist; bi[i] = -1;
    }
    const uint4* src = db + 2 * r0;
    const int nchunks = (nrows + T2_CHUNK - 1) / T2_CHUNK;
    /* the slice arrives as linear bulk copies (cp.async.bulk: one instruction per 8 KB stage, issued by one thread, completion on
     * an mbarrier per buffer) instead of one 16-byte cp.async per thread */
    __shared__ __align__(8) uint64_t s_bar[2];
    if (threadIdx.x == 0) { mbar_init(&s_bar[0], 1); mbar_init(&s_bar[1], 1); }
    __syncthreads();
    auto stage = [&](int c, int buf) {
        if (threadIdx.x == 0) {
            const uint32_t bytes = (uint32_t)min(T2_CHUNK, nrows - c * T2_CHUNK) * 32u;
            mbar_expect_tx(&s_bar[buf], bytes);
            bulk_load(&s_db[buf][0], src + (size_t)c * T2_CHUNK * 2, bytes, &s_bar[buf]);
        }
    };
    if (nchunks > 0) stage(0, 0);
    for (int c = 0; c < nchunks; c++) {
        if (c + 1 < nchunks) stage(c + 1, (c + 1) & 1);   /* its buffer was released by the barrier that ended iteration c-1 */
        mbar_wait(&s_bar[c & 1], (uint32_t)(c >> 1) & 1u);
        const uint4* s = s_db[c & 1];
        const int rows = min(T2_CHUNK, nrows - c * T2_CHUNK);
        const int rowBase = (int)(indexBase + r0) + c * T2_CHUNK;
#pragma unroll 2
        for (int r = 0; r < rows; r++) {
            const uint4 a = s[2 * r], b = s[2 * r + 1];
#pragma unroll
            for (int i = 0; i < T2_QPT; i++) {
                const int dist = ham256(qr[i], a, b);
                if (dist < b2[i]) {
                    if (dist < b1[i]) { b2[i] = b1[i]; b1[i] = dist; bi[i] = rowBase + r; }
                    else b2[i] = dist;
                }
            }
        }
        __syncthreads();
    }
#pragma unroll
    for (int i = 0; i < T2_QPT; i++) {
        const int qi = blockIdx.x * T2_QTILE + i * T2_NT + threadIdx.x;
        if (qi < nq) {
            orbfe_match m;
            m.best1 = b1[i]; m.best2 = b2[i]; m.idx = bi[i]; m.accepted = 0;
            parts[(size_t)slice * nq + qi] = m;
        }
    }
}

/* mode 0: best1 <= th && (float)best1 < ratio*(float)best2   (KF-Frame, src/ORBmatcher.cc:373-375)
 * mode 1: best1 <  th && ...                                    (KF-KF,    src/ORBmatcher.cc:906-908)
 * mode 2: knn-2 + Lowe in double: (double)(float)d0 < (double)(float)d1 * ratio, needs two rows (src/Frame.cc:1144) */
__global__ void k_top2_merge(const orbfe_match* __restrict__ parts, int nshards, int nq, int th, float ratio, double dratio,
                             int mode, int haveTwo, int initDist, orbfe_match* __restrict__ out) {
    const int qi = blockIdx.x * blockDim.x + threadIdx.x;
    if (qi >= nq) return;
    int b1 = initDist, b2 = initDist, bi = -1;
    for (int s = 0; s < nshards; s++) {
        const orbfe_match m = parts[(size_t)s * nq + qi];
        if (m.best1 < b1) { b2 = b1; b1 = m.best1; bi = m.idx; }
        else if (m.best1 < b2) b2 = m.best1;
        if (m.best2 < b2) b2 = m.best2; /* m.best2 >= m.best1 >= b1 here */
    }
    int ok;
    if (mode == 2) ok = haveTwo && ((double)(float)b1 < (double)(float)b2 * dratio);
    else ok = (mode == 1 ? (b1 < th) : (b1 <= th)) && ((float)b1 < __fmul_rn(ratio, (float)b2));
    orbfe_match r;
    r.best1 = b1; r.best2 = b2; r.idx = bi; r.accepted = ok;
    out[qi] = r;
}

/* Matching step of Frame::ComputeStereoFishEyeMatches (src/Frame.cc:1119-1144) for a batch of frames straight from the two
 * extractors' device buffers: the queries of frame f are the left keypoints inside the lapping area (rows monoL .. nL-1, where
 * ORBextractor::operator() put them), the train rows the right ones (monoR .. nR-1); BFMatcher(NORM_HAMMING).knnMatch(k=2) keeps
 * the two smallest distances with ties to the lowest train row, and the match survives iff there are two rows and
 * d0 < d1*ratio with the float distances promoted to double by the 0.7 literal.  One thread = one query, a CTA's 256 queries
 * share the train rows, which stream through shared memory as bulk copies.  match_lr[f*cap + row] = right row or -1. */
#define KB_NT 256
__global__ void __launch_bounds__(KB_NT) k_fisheye_knn2(const __grid_constant__ ExDesc d, ExBuffers Lb, ExBuffers Rb, double ratio, int32_t* __restrict__ matchLR,
                                                        int32_t* __restrict__ dist01) {
    __shared__ __align__(16) uint4 s_db[2][T2_CHUNK * 2];
    __shared__ __align__(8) uint64_t s_bar[2];
    const size_t f = blockIdx.y;
    const int nL = Lb.count[f], monoL = Lb.mono[f], nR = Rb.count[f], monoR = Rb.mono[f];
    const int nq = nL - monoL, nt = nR - monoR;
    int32_t* out = matchLR + f * d.kpCap;
    if (blockIdx.x == 0)
        for (int i = threadIdx.x; i < monoL; i += KB_NT) out[i] = -1;     /* outside the lapping area: never matched */
    if ((int)blockIdx.x * KB_NT >= nq) return;
    if (threadIdx.x == 0) { mbar_init(&s_bar[0], 1); mbar_init(&s_bar[1], 1); }
    __syncthreads();
    const int qi = blockIdx.x * KB_NT + threadIdx.x;
    uint32_t qr[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    if (qi < nq) {
        const uint4* q = reinterpret_cast<const uint4*>(Lb.desc + (f * d.kpCap + monoL + qi) * 32);
        const uint4 a = __ldg(q), b = __ldg(q + 1);
        qr[0] = a.x; qr[1] = a.y; qr[2] = a.z; qr[3] = a.w; qr[4] = b.x; qr[5] = b.y; qr[6] = b.z; qr[7] = b.w;
    }
    const uint4* src = reinterpret_cast<const uint4*>(Rb.desc + (f * d.kpCap + monoR) * 32);
    const int nchunks = (nt + T2_CHUNK - 1) / T2_CHUNK;
    auto stage = [&](int c, int buf) {
        if (threadIdx.x == 0) {
            const uint32_t bytes = (uint32_t)min(T2_CHUNK, nt - c * T2_CHUNK) * 32u;
            mbar_expect_tx(&s_bar[buf], bytes);
            bulk_load(&s_db[buf][0], src + (size_t)c * T2_CHUNK * 2, bytes, &s_bar[buf]);
        }
    };
    int b1 = 257, b2 = 257, bi = -1;
    if (nchunks > 0) stage(0, 0);
    for (int c = 0; c < nchunks; c++) {
        if (c + 1 < nchunks) stage(c + 1, (c + 1) & 1);
        mbar_wait(&s_bar[c & 1], (uint32_t)(c >> 1) & 1u);
        const uint4* s = s_db[c & 1];
        const int rows = min(T2_CHUNK, nt - c * T2_CHUNK);
#pragma unroll 4
        for (int r = 0; r < rows; r++) {
            const int dist = ham256(qr, s[2 * r], s[2 * r + 1]);
            if (dist < b2) {
                if (dist < b1) { b2 = b1; b1 = dist; bi = c * T2_CHUNK + r; }
                else b2 = dist;
            }
        }
        __syncthreads();
    }
    if (qi < nq) {
        const bool ok = nt >= 2 && ((double)(float)b1 < (double)(float)b2 * ratio);
        out[monoL + qi] = ok ? monoR + bi : -1;
        if (dist01) { dist01[2 * (f * d.kpCap + monoL + qi)] = b1; dist01[2 * (f * d.kpCap + monoL + qi) + 1] = b2; }
    }
}

#define TR_NT 128
__global__ void __launch_bounds__(TR_NT) k_fisheye_triangulate(const __grid_constant__ ExDesc d, ExBuffers Lb, ExBuffers Rb, const __grid_constant__ FisheyeRig rig,
                                                               int32_t* __restrict__ matchLR, int32_t* __restrict__ matchRL, float* __restrict__ depthOut,
                                                               float* __restrict__ p3dOut) {
    const size_t f = blockIdx.y;
    const int i = blockIdx.x * TR_NT + threadIdx.x;
    if (i >= Lb.count[f]) return;
    const size_t row = f * d.kpCap + i;
    const int t = matchLR[row];
    float depth = -1.f, x3D[3] = {0.f, 0.f, 0.f};
    if (t >= 0) {
        const orbfe_keypoint k1 = Lb.kps[row], k2 = Rb.kps[f * d.kpCap + t];
        const float s1 = d.lv[k1.octave].scale * d.lv[k1.octave].scale, s2 = d.lv[k2.octave].scale * d.lv[k2.octave].scale;   /* mvLevelSigma2 */
        depth = kb8_triangulate_pair(rig, k1.x, k1.y, k2.x, k2.y, s1, s2, x3D);
        if (depth > 0.0001f) atomicMax(&matchRL[f * d.kpCap + t], i);   /* src/Frame.cc:1151-1153: the last left keypoint that took a right one keeps it */
        else { matchLR[row] = -1; depth = -1.f; x3D[0] = x3D[1] = x3D[2] = 0.f; }
    }
    depthOut[row] = depth;
    p3dOut[3 * row] = x3D[0]; p3dOut[3 * row + 1] = x3D[1]; p3dOut[3 * row + 2] = x3D[2];
}

/* the same function on plain arrays of keypoint pairs (what a caller with its own match lists uses, and what the parity test
 * feeds the golden pairs through): depth = TriangulateMatches' return value, -1 when it rejects */
__global__ void __launch_bounds__(TR_NT) k_kb8_triangulate_pairs(const __grid_constant__ FisheyeRig rig, const float2* __restrict__ uv1, const float2* __restrict__ uv2,
                                                                 const float* __restrict__ sigma1, const float* __restrict__ sigma2, int n,
                                                                 float* __restrict__ depthOut, float* __restrict__ p3dOut) {
    const int i = blockIdx.x * TR_NT + threadIdx.x;
    if (i >= n) return;
    float x3D[3] = {0.f, 0.f, 0.f};
    const float2 a = uv1[i], b = uv2[i];
    const float depth = kb8_triangulate_pair(rig, a.x, a.y, b.x, b.y, sigma1[i], sigma2[i], x3D);
    depthOut[i] = depth;
    const bool ok = depth > 0;
    p3dOut[3 * i] = ok ? x3D[0] : 0.f; p3dOut[3 * i + 1] = ok ? x3D[1] : 0.f; p3dOut[3 * i + 2] = ok ? x3D[2] : 0.f;
}

__global__ void k_pair_distances(const uint4* __restrict__ a, const uint4* __restrict__ b, long long n, int* __restrict__ out) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint4 a0 = a[2 * i], a1 = a[2 * i + 1], b0 = b[2 * i], b1 = b[2 * i + 1];
    out[i] = __popc(a0.x ^ b0.x) + __popc(a0.y ^ b0.y) + __popc(a0.z ^ b0.z) + __popc(a0.w ^ b0.w) +
             __popc(a1.x ^ b1.x) + __popc(a1.y ^ b1.y) + __popc(a1.z ^ b1.z) + __popc(a1.w ^ b1.w);
}

struct orbfe_matcher {
    int device;
    int maxQ;
    long long maxDb;
    uint8_t* dDbOwned;     /* storage for orbfe_matcher_set_db */
    const uint8_t* dDb;    /* current database (owned or caller's) */
    long long ndb, indexBase;
    uint8_t* dQ;
    orbfe_match* dParts;   /* [maxSlices][maxQ] scratch of the partial kernel */
    orbfe_match* dOut;
    int maxSlices;
    cudaStream_t stream;
    orbfe_match* dGather;  /* [gatherRanks][maxQ] all-gathered per-shard records of orbfe_hamming_top2_sharded, allocated on first use */
    orbfe_match* dLocal;   /* [maxQ] this shard's records (the all-gather's send buffer) */
    int gatherRanks;
};

static int top2_plan(long long ndb, int nq, int maxSlices, int* rowsPerSlice) {
    /* enough CTAs to fill 148 SMs several times over, but slices of at least one smem stage */
    const int qTiles = (nq + T2_QTILE - 1) / T2_QTILE;
    int slices = std::max(1, std::min(maxSlices, (148 * 4 + qTiles - 1) / qTiles));
    long long rps = (ndb + slices - 1) / slices;
    rps = std::max<long long>(rps, T2_CHUNK);
    rps = (rps + 7) / 8 * 8;
    slices = (int)std::max<long long>(1, (ndb + rps - 1) / rps);
    *rowsPerSlice = (int)rps;
    return slices;
}

static int launch_top2(const uint8_t* dQ, int nq, const uint8_t* dDb, long long ndb, long long indexBase, orbfe_match* dParts,
                       int maxSlices, int th, float ratio, double dratio, int mode, orbfe_match* dOut, cudaStream_t st) {
    const int initDist = mode == 2 ? 257 : 256;
    int rps = 0;
    const int slices = top2_plan(ndb, nq, maxSlices, &rps);
    const int qTiles = (nq + T2_QTILE - 1) / T2_QTILE;
    k_top2_partial<<<dim3(qTiles, slices), T2_NT, 0, st>>>((const uint4*)dQ, nq, (const uint4*)dDb, ndb, indexBase, rps, initDist, dParts);
    k_top2_merge<<<(nq + 127) / 128, 128, 0, st>>>(dParts, slices, nq, th, ratio, dratio, mode, ndb >= 2, initDist, dOut);
    ORBFE_CUDA(cudaGetLastError());
    return 0;
}

/* --------------------------------------------------------------------------------------------------- stereo */
#define ST_NT 64
#define ST_BANDSHIFT 3
/* One CTA = one frame: bucket the right keypoints by the 8-row band of their y (counting sort in shared memory), so that
 * the matcher only walks the bands that can hold candidates for a left keypoint's row instead of every right keypoint.
 * Entry = (x bits, floor(y-r) | ceil(y+r) << 12 | octave << 24, original index, 0); r = 2*scale (src/Frame.cc:814-824). */
__global__ void __launch_bounds__(256) k_stereo_bucket(const __grid_constant__ ExDesc d, ExBuffers R) {
    extern __shared__ int s_cnt[];   /* [stBands + 1] */
    const size_t f = blockIdx.x;
    const int nR = R.count[f], nb = d.stBands;
    for (int i = threadIdx.x; i <= nb; i += 256) s_cnt[i] = 0;
    __syncthreads();
    const orbfe_keypoint* kR = R.kps + f * d.kpCap;
    for (int i = threadIdx.x; i < nR; i += 256) atomicAdd(&s_cnt[min(nb - 1, max(0, (int)kR[i].y) >> ST_BANDSHIFT)], 1);
    __syncthreads();
    if (threadIdx.x == 0) {
        int run = 0;
        for (int i = 0; i < nb; i++) { const int c = s_cnt[i]; s_cnt[i] = run; run += c; }
        s_cnt[nb] = run;
    }
    __syncthreads();
    int* band = R.stBand + f * (nb + 1);
    for (int i = threadIdx.x; i <= nb; i += 256) band[i] = s_cnt[i];
    __syncthreads();
    uint4* tab = R.stTable + f * d.kpCap;
    for (int i = threadIdx.x; i < nR; i += 256) {
        const orbfe_keypoint k = kR[i];
        const float r = __fmul_rn(2.0f, d.lv[k.octave].scale);
        const int maxr = max(0, min(4095, (int)ceilf(__fadd_rn(k.y, r)))), minr = max(0, min(4095, (int)floorf(__fsub_rn(k.y, r))));
        const int slot = atomicAdd(&s_cnt[min(nb - 1, max(0, (int)k.y) >> ST_BANDSHIFT)], 1);
        tab[slot] = make_uint4(__float_as_uint(k.x), (uint32_t)minr | ((uint32_t)maxr << 12) | ((uint32_t)k.octave << 24), (uint32_t)i, 0u);
    }
}

#define ST_KPB 2   /* left keypoints per CTA = warps per CTA */
/* One warp = one left keypoint.  Lanes stride over the right keypoints of the bands whose rows can reach this keypoint's
 * row (a right keypoint covers rows [floor(y-r), ceil(y+r)] with r <= 2*maxScale), apply the reference's gates (row band,
 * octave +-1, disparity range), keep (distance, original index) minima — the reference scans in index order with a
 * strict `<`, i.e. the lowest index wins ties — then the warp does the 11-shift SAD and the parabola fit.  Float steps
 * use round-to-nearest intrinsics (no contraction). */
__global__ void __launch_bounds__(ST_NT) k_stereo_match(const __grid_constant__ ExDesc d, ExBuffers L, ExBuffers R, float mbf,
                                                        float minD, float maxD) {
    const size_t f = blockIdx.y;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int n = L.count[f];
    {
        const int iL = blockIdx.x * ST_KPB + warp;
        if (iL >= n) return;
        const size_t o = f * d.kpCap + iL;
        float outU = -1.f, outD = -1.f;
        int outS = -1;
        const orbfe_keypoint kpL = L.kps[o];
        const int levelL = kpL.octave;
        const float vL = kpL.y, uL = kpL.x;
        const int row = (int)vL;
        const float minU = __fsub_rn(uL, maxD), maxU = __fsub_rn(uL, minD);
        int best = 100, bestR = 0x7FFFFFFF; /* TH_HIGH, src/ORBmatcher.cc:36 */
        float bestX = 0.f;
        if (!(maxU < 0)) {
            const uint4* dL = reinterpret_cast<const uint4*>(L.desc + o * 32);
            const uint4 dl0 = dL[0], dl1 = dL[1];
            /* |(int)y - row| <= r + 2 with r <= 2 * scale of the coarsest level */
            const int reach = (int)ceilf(__fmul_rn(2.0f, d.lv[d.nlevels - 1].scale)) + 2;
            const int nb = d.stBands;
            const int* band = R.stBand + f * (nb + 1);
            const int b0 = max(0, (row - reach) >> ST_BANDSHIFT), b1 = min(nb - 1, (row + reach) >> ST_BANDSHIFT);
            const int e0 = band[b0], e1 = band[b1 + 1];
            const uint4* tab = R.stTable + f * d.kpCap;
            for (int e = e0 + lane; e < e1; e += 32) {
                const uint4 k = tab[e];
                const int minr = k.y & 0xFFF, maxr = (k.y >> 12) & 0xFFF, oct = k.y >> 24;
                if (row < minr || row > maxr) continue;
                if (oct < levelL - 1 || oct > levelL + 1) continue;
                const float uR = __uint_as_float(k.x);
                if (uR >= minU && uR <= maxU) {
                    const int iR = (int)k.z;
                    const uint4* dR = reinterpret_cast<const uint4*>(R.desc + (f * d.kpCap + iR) * 32);
                    const uint4 c0 = dR[0], c1 = dR[1];
                    const int dist = __popc(dl0.x ^ c0.x) + __popc(dl0.y ^ c0.y) + __popc(dl0.z ^ c0.z) + __popc(dl0.w ^ c0.w) +
                                     __popc(dl1.x ^ c1.x) + __popc(dl1.y ^ c1.y) + __popc(dl1.z ^ c1.z) + __popc(dl1.w ^ c1.w);
                    if (dist < best || (dist == best && iR < bestR)) { best = dist; bestR = iR; bestX = uR; }
                }
            }
#pragma unroll
            for (int s = 16; s > 0; s >>= 1) {
                const int ob = __shfl_xor_sync(0xFFFFFFFFu, best, s), oi = __shfl_xor_sync(0xFFFFFFFFu, bestR, s);
                const float ox = __shfl_xor_sync(0xFFFFFFFFu, bestX, s);
                if (ob < best || (ob == best && oi < bestR)) { best = ob; bestR = oi; bestX = ox; }
            }
        }
        if (best < 75 && !(maxU < 0)) { /* thOrbDist = (TH_HIGH+TH_LOW)/2, src/Frame.cc:802 */
            const LevelDesc& lv = d.lv[levelL];
            const float uR0 = bestX;
            const float sf = __fdiv_rn(1.0f, lv.scale); /* mvInvScaleFactors[octave] */
            const float suL = roundf(__fmul_rn(kpL.x, sf)), svL = roundf(__fmul_rn(kpL.y, sf)), suR0 = roundf(__fmul_rn(uR0, sf));
            const float iniu = __fsub_rn(__fadd_rn(suR0, 5.f), 5.f), endu = __fadd_rn(__fadd_rn(__fadd_rn(suR0, 5.f), 5.f), 1.f);
            if (!(iniu < 0 || endu >= (float)lv.w)) {
                const int y0 = (int)__fsub_rn(svL, 5.f), xl0 = (int)__fsub_rn(suL, 5.f), xr0 = (int)suR0 - 5;
                const uint8_t* pl = L.pyr + f * d.pyrSlab + lv.pyrOff + (size_t)(y0 + ORBFE_EDGE) * lv.pitch + ORBFE_ROW_LEAD + xl0;
                const uint8_t* pr = R.pyr + f * d.pyrSlab + lv.pyrOff + (size_t)(y0 + ORBFE_EDGE) * lv.pitch + ORBFE_ROW_LEAD + xr0;
                int sad[11];
#pragma unroll
                for (int k = 0; k < 11; k++) sad[k] = 0;
                for (int p = lane; p < 121; p += 32) {
                    const int dy = p / 11, dx = p - dy * 11;
                    const int a = pl[dy * lv.pitch + dx];
                    const uint8_t* rr = pr + dy * lv.pitch + dx - 5;
#pragma unroll
                    for (int k = 0; k < 11; k++) sad[k] += abs(a - (int)rr[k]);
                }
#pragma unroll
                for (int k = 0; k < 11; k++)
#pragma unroll
                    for (int s = 16; s > 0; s >>= 1) sad[k] += __shfl_xor_sync(0xFFFFFFFFu, sad[k], s);
                int bestS = 0x7FFFFFFF, bestInc = 0;
#pragma unroll
                for (int k = 0; k < 11; k++)
                    if (sad[k] < bestS) { bestS = sad[k]; bestInc = k - 5; }
                if (bestInc != -5 && bestInc != 5) {
                    float d1 = 0.f, d2 = 0.f, d3 = 0.f;
#pragma unroll
                    for (int k = 1; k < 10; k++)
                        if (k == bestInc + 5) { d1 = (float)sad[k - 1]; d2 = (float)sad[k]; d3 = (float)sad[k + 1]; }
                    const float deltaR = __fdiv_rn(__fsub_rn(d1, d3), __fmul_rn(2.0f, __fsub_rn(__fadd_rn(d1, d3), __fmul_rn(2.0f, d2))));
                    if (!(deltaR < -1 || deltaR > 1)) {
                        float bestuR = __fmul_rn(lv.scale, __fadd_rn(__fadd_rn(suR0, (float)bestInc), deltaR));
                        float disparity = __fsub_rn(uL, bestuR);
                        if (disparity >= minD && disparity < maxD) {
                            if (disparity <= 0) {
                                disparity = 0.01f;
                                bestuR = (float)((double)uL - 0.01);
                            }
                            outD = __fdiv_rn(mbf, disparity);
                            outU = bestuR;
                            outS = bestS;
                        }
                    }
                }
            }
        }
        if (lane == 0) {
            L.uRight[o] = outU;
            L.depth[o] = outD;
            L.sadBest[o] = outS;
        }
    }
}

/* One CTA = one frame: median of the accepted SAD scores (element size/2 of the sorted list, src/Frame.cc:953-955)
 * by a two-level histogram, then drop matches with SAD >= 1.5*1.4*median (src/Frame.cc:957-966). */
__global__ void __launch_bounds__(256) k_stereo_median(const __grid_constant__ ExDesc d, ExBuffers L) {
    __shared__ int s_hist[256];
    __shared__ int s_sel, s_below, s_m;
    const size_t f = blockIdx.x;
    const int n = L.count[f];
    const int* sad = L.sadBest + f * d.kpCap;
    const int tid = threadIdx.x;
    s_hist[tid] = 0;
    if (tid == 0) s_m = 0;
    __syncthreads();
    int mine = 0;
    for (int i = tid; i < n; i += 256) {
        const int v = sad[i];
        if (v >= 0) { mine++; atomicAdd(&s_hist[min(v >> 8, 255)], 1); }
    }
    if (mine) atomicAdd(&s_m, mine);
    __syncthreads();
    const int m = s_m;
    if (m == 0) return; /* defined deviation: the reference indexes an empty vector here */
    const int target = m / 2;
    if (tid == 0) {
        int run = 0, b = 0;
        for (; b < 256; b++) { if (run + s_hist[b] > target) break; run += s_hist[b]; }
        s_sel = b; s_below = run;
    }
    __syncthreads();
    const int hi = s_sel, below = s_below;
    __syncthreads();
    s_hist[tid] = 0;
    __syncthreads();
    for (int i = tid; i < n; i += 256) {
        const int v = sad[i];
        if (v >= 0 && min(v >> 8, 255) == hi) atomicAdd(&s_hist[v & 255], 1);
    }
    __syncthreads();
    if (tid == 0) {
        int run = below, b = 0;
        for (; b < 256; b++) { if (run + s_hist[b] > target) break; run += s_hist[b]; }
        s_sel = (hi << 8) | b;
    }
    __syncthreads();
    const float median = (float)s_sel;
    const float thDist = __fmul_rn(__fmul_rn(1.5f, 1.4f), median);
    for (int i = tid; i < n; i += 256) {
        const int v = sad[i];
        if (v >= 0 && !((float)v < thDist)) {
            L.uRight[f * d.kpCap + i] = -1.f;
            L.depth[f * d.kpCap + i] = -1.f;
        }
    }
}

static bool same_geometry(const orbfe_extractor* a, const orbfe_extractor* b) {
    return a->device == b->device && a->d.nlevels == b->d.nlevels && a->d.W == b->d.W && a->d.H == b->d.H &&
           a->d.kpCap == b->d.kpCap && a->d.pyrSlab == b->d.pyrSlab && a->params.scale_factor == b->params.scale_factor;
}

/* MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:349-423) for many map points at once: point p owns the descriptor
 * rows [off[p], off[p+1]) (its observations); the representative is the row with the least median distance to all rows of
 * the point (its own 0 included), median = sorted[int(0.5*(N-1))], first row on ties.  One warp = one map point, one lane =
 * one row at a time.  The median of a row is the smallest v with #(d <= v) > rank, found by bisection on v in [0, 256]: nine
 * passes over the point's rows, no sort, no N x N matrix. */
#define DD_WARPS 4
__global__ void __launch_bounds__(DD_WARPS * 32) k_distinctive(const uint4* __restrict__ desc, const int* __restrict__ off, int npoints,
                                                               int* __restrict__ bestIdx, int* __restrict__ bestMedian) {
    const int p = blockIdx.x * DD_WARPS + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (p >= npoints) return;
    const int r0 = off[p], n = off[p + 1] - r0;
    if (n <= 0) {
        if (lane == 0) { bestIdx[p] = -1; if (bestMedian) bestMedian[p] = INT_MAX; }
        return;
    }
    const int rank = (int)(0.5 * (n - 1));
    const uint4* d = desc + (size_t)r0 * 2;
    unsigned best = 0xFFFFFFFFu;   /* median << 16 | row */
    for (int i = lane; i < n; i += 32) {
        uint32_t q[8];
        { const uint4 a = __ldg(d + 2 * i), b = __ldg(d + 2 * i + 1); q[0] = a.x; q[1] = a.y; q[2] = a.z; q[3] = a.w; q[4] = b.x; q[5] = b.y; q[6] = b.z; q[7] = b.w; }
        int lo = 0, hi = 256;      /* the median lies in [lo, hi] */
        while (lo < hi) {
            const int mid = (lo + hi) >> 1;
            int cnt = 0;
            for (int j = 0; j < n; j++) cnt += ham256(q, __ldg(d + 2 * j), __ldg(d + 2 * j + 1)) <= mid ? 1 : 0;
            if (cnt > rank) hi = mid; else lo = mid + 1;
        }
        best = min(best, ((unsigned)lo << 16) | (unsigned)min(i, 0xFFFF));
    }
#pragma unroll
    for (int o = 16; o; o >>= 1) best = min(best, __shfl_xor_sync(0xFFFFFFFFu, best, o));
    if (lane == 0) { bestIdx[p] = (int)(best & 0xFFFFu); if (bestMedian) bestMedian[p] = (int)(best >> 16); }
}

int orbfe_launch_stereo(orbfe_extractor* exL, orbfe_extractor* exR, int nframes, int frame0, float mbf, float minD, float maxD, cudaStream_t st) {
    const ExBuffers bl = frame0 ? orbfe_shift_frames(exL->b, exL->d, frame0) : exL->b;
    const ExBuffers br = frame0 ? orbfe_shift_frames(exR->b, exR->d, frame0) : exR->b;
    k_stereo_bucket<<<nframes, 256, (exR->d.stBands + 1) * sizeof(int), st>>>(exR->d, br);
    k_stereo_match<<<dim3((exL->d.kpCap + ST_KPB - 1) / ST_KPB, nframes), ST_NT, 0, st>>>(exL->d, bl, br, mbf, minD, maxD);
    k_stereo_median<<<nframes, 256, 0, st>>>(exL->d, bl);
    exL->launches += 3;
    ORBFE_CUDA(cudaGetLastError());
    return ORBFE_OK;
}

/* knn-2 + ratio, then TriangulateMatches and the output vectors, for frames frame0 .. frame0+nframes-1 of the two objects; the four
 * output pointers are those of the launch's first frame ([nframes][kpCap] each, p3d [nframes][kpCap][3]). */
int orbfe_launch_fisheye(orbfe_extractor* exL, orbfe_extractor* exR, int nframes, int frame0, double ratio, const orbfe_fisheye_rig* rig, int32_t* lr, int32_t* rl,
                         float* depth, float* p3d, cudaStream_t st) {
    const ExBuffers bl = frame0 ? orbfe_shift_frames(exL->b, exL->d, frame0) : exL->b;
    const ExBuffers br = frame0 ? orbfe_shift_frames(exR->b, exR->d, frame0) : exR->b;
    FisheyeRig r;
    static_assert(sizeof(FisheyeRig) == sizeof(orbfe_fisheye_rig), "rig layout");
    memcpy(&r, rig, sizeof r);
    k_fisheye_knn2<<<dim3((exL->d.kpCap + KB_NT - 1) / KB_NT, nframes), KB_NT, 0, st>>>(exL->d, bl, br, ratio, lr, nullptr);
    ORBFE_CUDA(cudaMemsetAsync(rl, 0xFF, (size_t)nframes * exL->d.kpCap * sizeof(int32_t), st));
    k_fisheye_triangulate<<<dim3((exL->d.kpCap + TR_NT - 1) / TR_NT, nframes), TR_NT, 0, st>>>(exL->d, bl, br, r, lr, rl, depth, p3d);
    exL->launches += 2;
    ORBFE_CUDA(cudaGetLastError());
    return ORBFE_OK;
}

extern "C" {

int orbfe_stereo_match_batch_device(orbfe_extractor* exL, orbfe_extractor* exR, int nframes, float mbf, float minD, float maxD, void* stream) {
    if (!exL || !exR) { orbfe_set_error("null extractor"); return ORBFE_E_INVALID; }
    if (!same_geometry(exL, exR)) { orbfe_set_error("left/right extractors differ in geometry or device"); return ORBFE_E_INVALID; }
    if (nframes < 1 || nframes > exL->lastBatch || nframes > exR->lastBatch) { orbfe_set_error("nframes exceeds the last extracted batch"); return ORBFE_E_INVALID; }
    ORBFE_CUDA(cudaSetDevice(exL->device));
    cudaStream_t st = stream ? (cudaStream_t)stream : exL->stream;
    if (!stream) ORBFE_CUDA(cudaStreamSynchronize(exR->stream)); /* right extraction ran on its own stream */
    const bool prof = exL->profiling != 0;
    if (prof && exL->stereoPending) { int rc = orbfe_collect_stage_times(exL); if (rc) return rc; }
    cudaEvent_t* ev = exL->ev + ORBFE_NUM_STAGES + 1; /* two private events: the extractor's own stay untouched */
    if (prof) ORBFE_CUDA(cudaEventRecord(ev[0], st));
    { int rc = orbfe_launch_stereo(exL, exR, nframes, 0, mbf, minD, maxD, st); if (rc) return rc; }
    if (prof) {
        /* stereo is short: time both kernels as one span, reported under "stereo match" */
        ORBFE_CUDA(cudaEventRecord(ev[1], st));
        exL->stereoPending = 1;
    }
    { int rc = orbfe_mark_last(exL, st); if (!rc) rc = orbfe_mark_last(exR, st); if (rc) return rc; }
    return ORBFE_OK;
}

int orbfe_fisheye_match_batch_device(orbfe_extractor* exL, orbfe_extractor* exR, int nframes, double ratio, int32_t* d_match_lr, int32_t* d_dist01, void* stream) {
    if (!exL || !exR || !d_match_lr) { orbfe_set_error("null argument"); return ORBFE_E_INVALID; }
    if (!same_geometry(exL, exR)) { orbfe_set_error("left/right extractors differ in geometry or device"); return ORBFE_E_INVALID; }
    if (nframes < 1 || nframes > exL->lastBatch || nframes > exR->lastBatch) { orbfe_set_error("nframes exceeds the last extracted batch"); return ORBFE_E_INVALID; }
    ORBFE_CUDA(cudaSetDevice(exL->device));
    cudaStream_t st = stream ? (cudaStream_t)stream : exL->stream;
    if (!stream) ORBFE_CUDA(cudaStreamSynchronize(exR->stream)); /* right extraction ran on its own stream */
    k_fisheye_knn2<<<dim3((exL->d.kpCap + KB_NT - 1) / KB_NT, nframes), KB_NT, 0, st>>>(exL->d, exL->b, exR->b, ratio, d_match_lr, d_dist01);
    ORBFE_CUDA(cudaGetLastError());
    { int rc = orbfe_mark_last(exL, st); if (!rc) rc = orbfe_mark_last(exR, st); if (rc) return rc; }
    return ORBFE_OK;
}

int orbfe_fisheye_triangulate_batch_device(orbfe_extractor* exL, orbfe_extractor* exR, int nframes, const orbfe_fisheye_rig* rig, int32_t* d_match_lr,
                                           int32_t* d_match_rl, float* d_depth, float* d_p3d, void* stream) {
    if (!exL || !exR || !rig || !d_match_lr || !d_match_rl || !d_depth || !d_p3d) { orbfe_set_error("null argument"); return ORBFE_E_INVALID; }
    if (!same_geometry(exL, exR)) { orbfe_set_error("left/right extractors differ in geometry or device"); return ORBFE_E_INVALID; }
    if (nframes < 1 || nframes > exL->lastBatch || nframes > exR->lastBatch) { orbfe_set_error("nframes exceeds the last extracted batch"); return ORBFE_E_INVALID; }
    ORBFE_CUDA(cudaSetDevice(exL->device));
    cudaStream_t st = stream ? (cudaStream_t)stream : exL->stream;
    FisheyeRig r;
    static_assert(sizeof(FisheyeRig) == sizeof(orbfe_fisheye_rig), "rig layout");
    memcpy(&r, rig, sizeof r);
    ORBFE_CUDA(cudaMemsetAsync(d_match_rl, 0xFF, (size_t)nframes * exL->d.kpCap * sizeof(int32_t), st));   /* -1 everywhere */
    k_fisheye_triangulate<<<dim3((exL->d.kpCap + TR_NT - 1) / TR_NT, nframes), TR_NT, 0, st>>>(exL->d, exL->b, exR->b, r, d_match_lr, d_match_rl, d_depth, d_p3d);
    ORBFE_CUDA(cudaGetLastError());
    { int rc = orbfe_mark_last(exL, st); if (!rc) rc = orbfe_mark_last(exR, st); if (rc) return rc; }
    return ORBFE_OK;
}

int orbfe_kb8_triangulate_pairs_device(int device, const orbfe_fisheye_rig* rig, const float* d_uv1, const float* d_uv2, const float* d_sigma1, const float* d_sigma2,
                                       int n, float* d_depth, float* d_p3d, void* stream) {
    if (!rig || !d_uv1 || !d_uv2 || !d_sigma1 || !d_sigma2 || !d_depth || !d_p3d || n < 1) { orbfe_set_error("bad argument"); return ORBFE_E_INVALID; }
    ORBFE_CUDA(cudaSetDevice(device));
    FisheyeRig r;
    memcpy(&r, rig, sizeof r);
    k_kb8_triangulate_pairs<<<(n + TR_NT - 1) / TR_NT, TR_NT, 0, (cudaStream_t)stream>>>(r, (const float2*)d_uv1, (const float2*)d_uv2, d_sigma1, d_sigma2, n, d_depth, d_p3d);
    ORBFE_CUDA(cudaGetLastError());
    return ORBFE_OK;
}

int orbfe_stereo_result(orbfe_extractor* exL, int frame, float* u_right, float* depth, int cap, int* n) {
    if (!exL || frame < 0 || frame >= exL->lastBatch) { orbfe_set_error("bad frame index"); return ORBFE_E_INVALID; }
    ORBFE_CUDA(cudaSetDevice(exL->device));
    { const int rc = orbfe_wait_last(exL); if (rc) return rc; }
    int cnt = 0;
    ORBFE_D2H(exL, &cnt, exL->b.count + frame, sizeof(int));
    ORBFE_D2H_DONE(exL);
    if (n) *n = cnt;
    if (cnt > cap) { orbfe_set_error("caller capacity %d < %d", cap, cnt); return ORBFE_E_CAPACITY; }
    if (u_right && cnt) ORBFE_D2H(exL, u_right, exL->b.uRight + (size_t)frame * exL->d.kpCap, cnt * sizeof(float));
    if (depth && cnt) ORBFE_D2H(exL, depth, exL->b.depth + (size_t)frame * exL->d.kpCap, cnt * sizeof(float));
    ORBFE_D2H_DONE(exL);
    return ORBFE_OK;
}

int orbfe_stereo_result_batch(orbfe_extractor* exL, int nframes, float* u_right, float* depth, int cap) {
    if (!exL || nframes < 1 || nframes > exL->lastBatch) { orbfe_set_error("bad frame count"); return ORBFE_E_INVALID; }
    if (cap < exL->d.kpCap) { orbfe_set_error("cap %d < keypoint capacity %d", cap, exL->d.kpCap); return ORBFE_E_CAPACITY; }
    ORBFE_CUDA(cudaSetDevice(exL->device));
    { const int rc = orbfe_wait_last(exL); if (rc) return rc; }
    const size_t w = (size_t)exL->d.kpCap * sizeof(float);
    if (u_right) ORBFE_D2H_2D(exL, u_right, (size_t)cap * sizeof(float), exL->b.uRight, w, w, nframes);
    if (depth) ORBFE_D2H_2D(exL, depth, (size_t)cap * sizeof(float), exL->b.depth, w, w, nframes);
    ORBFE_D2H_DONE(exL);
    return ORBFE_OK;
}

int orbfe_stereo_device_views(orbfe_extractor* exL, const float** u, const float** dd) {
    if (!exL) return ORBFE_E_INVALID;
    if (u) *u = exL->b.uRight;
    if (dd) *dd = exL->b.depth;
    return ORBFE_OK;
}

/* ------------------------------------------------------------------------------------------------- matcher */
void orbfe_matcher_destroy(orbfe_matcher* m) {
    if (!m) return;
    cudaSetDevice(m->device);
    if (m->stream) cudaStreamSynchronize(m->stream);
    cudaFree(m->dDbOwned);
    cudaFree(m->dQ);
    cudaFree(m->dParts);
    cudaFree(m->dOut);
    cudaFree(m->dGather);
    cudaFree(m->dLocal);
    if (m->stream) cudaStreamDestroy(m->stream);
    delete m;
}

int orbfe_matcher_create(int device, int max_queries, int64_t max_db_rows, orbfe_matcher** out) {
    if (!out || max_queries < 1 || max_db_rows < 0) { orbfe_set_error("bad argument"); return ORBFE_E_INVALID; }
    *out = nullptr;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || device < 0 || device >= ndev) {
        orbfe_set_error("no usable CUDA device %d (found %d); this library has no CPU path", device, ndev);
        return ORBFE_E_CUDA;
    }
    ORBFE_CUDA(cudaSetDevice(device));
    orbfe_matcher* m = new (std::nothrow) orbfe_matcher();
    if (!m) return ORBFE_E_INVALID;
    memset(m, 0, sizeof *m);
    m->device = device; m->maxQ = max_queries; m->maxDb = max_db_rows;
    m->maxSlices = 148 * 4;
    cudaError_t e = cudaStreamCreateWithFlags(&m->stream, cudaStreamNonBlocking);
    if (e == cudaSuccess && max_db_rows > 0) e = cudaMalloc((void**)&m->dDbOwned, (size_t)max_db_rows * 32);
    if (e == cudaSuccess) e = cudaMalloc((void**)&m->dQ, (size_t)max_queries * 32);
    if (e == cudaSuccess) e = cudaMalloc((void**)&m->dParts, (size_t)m->maxSlices * max_queries * sizeof(orbfe_match));
    if (e == cudaSuccess) e = cudaMalloc((void**)&m->dOut, (size_t)max_queries * sizeof(orbfe_match));
    if (e != cudaSuccess) { const int rc = orbfe_cuda_fail(e, "matcher allocation"); orbfe_matcher_destroy(m); return rc; }
    *out = m;
    return ORBFE_OK;
}

int orbfe_matcher_set_db(orbfe_matcher* m, const uint8_t* h_db, int64_t rows, int64_t index_base) {
    if (!m || (rows > 0 && !h_db) || rows < 0) { orbfe_set_error("bad argument"); return ORBFE_E_INVALID; }
    if (rows > m->maxDb) { orbfe_set_error("database of %lld rows exceeds capacity %lld", (long long)rows, m->maxDb); return ORBFE_E_CAPACITY; }
    ORBFE_CUDA(cudaSetDevice(m->device));
    if (rows) ORBFE_CUDA(cudaMemcpyAsync(m->dDbOwned, h_db, (size_t)rows * 32, cudaMemcpyHostToDevice, m->stream));
    ORBFE_CUDA(cudaStreamSynchronize(m->stream));
    m->dDb = m->dDbOwned; m->ndb = rows; m->indexBase = index_base;
    return ORBFE_OK;
}

int orbfe_matcher_set_db_device(orbfe_matcher* m, const uint8_t* d_db, int64_t rows, int64_t index_base) {
    if (!m || (rows > 0 && !d_db) || rows < 0) { orbfe_set_error("bad argument"); return ORBFE_E_INVALID; }
    if (((uintptr_t)d_db & 15) != 0) { orbfe_set_error("database pointer must be 16-byte aligned"); return ORBFE_E_INVALID; }
    m->dDb = d_db; m->ndb = rows; m->indexBase = index_base;
    return ORBFE_OK;
}

int orbfe_matcher_top2_device(orbfe_matcher* m, const uint8_t* d_queries, int nq, int th, float ratio, int strict_lt,
                              orbfe_match* d_out, void* stream) {
    if (!m || !d_queries || !d_out || nq < 1) { orbfe_set_error("bad argument"); return ORBFE_E_INVALID; }
    if (nq > m->maxQ) { orbfe_set_error("%d queries exceed capacity %d", nq, m->maxQ); return ORBFE_E_CAPACITY; }
    if (((uintptr_t)d_queries & 15) != 0) { orbfe_set_error("query pointer must be 16-byte aligned"); return ORBFE_E_INVALID; }
    ORBFE_CUDA(cudaSetDevice(m->device));
    return launch_top2(d_queries, nq, m->dDb, m->ndb, m->indexBase, m->dParts, m->maxSlices, th, ratio, 0.0, strict_lt ? 1 : 0,
                       d_out, stream ? (cudaStream_t)stream : m->stream);
}

/* ---- the sharded search: local top-2, NCCL all-gather of the per-shard records, merge — one stream, no host
 * round trip in between.  NCCL is looked up at run time (dlopen of the libnccl.so.2 the process already uses, e.g. the one torch
 * loaded, else the system one), so liborbfe.so itself has no link-time dependency on it. */
namespace {
typedef int (*nccl_allgather_fn)(const void*, void*, size_t, int /*ncclDataType_t*/, void* /*ncclComm_t*/, cudaStream_t);
typedef const char* (*nccl_errstr_fn)(int);
typedef int (*nccl_uid_fn)(void*);
typedef struct { char internal[128]; } nccl_uid;   /* ncclUniqueId, nccl.h: NCCL_UNIQUE_ID_BYTES = 128 */
typedef int (*nccl_init_fn)(void**, int, nccl_uid, int);
typedef int (*nccl_destroy_fn)(void*);
struct NcclApi {
    void* lib = nullptr;
    nccl_allgather_fn allGather = nullptr;
    nccl_errstr_fn errStr = nullptr;
    nccl_uid_fn uniqueId = nullptr;
    nccl_init_fn initRank = nullptr;
    nccl_destroy_fn destroy = nullptr;
};
std::mutex g_nccl_mu;
NcclApi g_nccl;
int nccl_api(NcclApi** out) {
    std::lock_guard<std::mutex> lk(g_nccl_mu);
    if (!g_nccl.lib) {
        const char* names[] = {getenv("ORBFE_NCCL_LIB"), "libnccl.so.2", "libnccl.so"};
        for (const char* n : names) {
            if (!n || !*n) continue;
            g_nccl.lib = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
            if (g_nccl.lib) break;
        }
        if (!g_nccl.lib) { orbfe_set_error("NCCL library not found (libnccl.so.2; set ORBFE_NCCL_LIB): %s", dlerror()); return ORBFE_E_NCCL; }
        g_nccl.allGather = (nccl_allgather_fn)dlsym(g_nccl.lib, "ncclAllGather");
        g_nccl.errStr = (nccl_errstr_fn)dlsym(g_nccl.lib, "ncclGetErrorString");
        g_nccl.uniqueId = (nccl_uid_fn)dlsym(g_nccl.lib, "ncclGetUniqueId");
        g_nccl.initRank = (nccl_init_fn)dlsym(g_nccl.lib, "ncclCommInitRank");
        g_nccl.destroy = (nccl_destroy_fn)dlsym(g_nccl.lib, "ncclCommDestroy");
        if (!g_nccl.allGather || !g_nccl.errStr || !g_nccl.uniqueId || !g_nccl.initRank || !g_nccl.destroy) {
            g_nccl.lib = nullptr;
            orbfe_set_error("the NCCL library lacks ncclAllGather / ncclCommInitRank");
            return ORBFE_E_NCCL;
        }
    }
    *out = &g_nccl;
    return 0;
}
int nccl_fail(NcclApi* api, int rc, const char* what) {
    orbfe_set_error("NCCL error %d (%s) in %s", rc, api->errStr ? api->errStr(rc) : "?", what);
    return ORBFE_E_NCCL;
}
}  // namespace

int orbfe_nccl_unique_id(uint8_t id128[128]) {
    NcclApi* api = nullptr;
    int rc = nccl_api(&api);
    if (rc) return rc;
    if (!id128) { orbfe_set_error("bad argument"); return ORBFE_E_INVALID; }
    nccl_uid u;
    rc = api->uniqueId(&u);
    if (rc) return nccl_fail(api, rc, "ncclGetUniqueId");
    memcpy(id128, u.internal, 128);
    return ORBFE_OK;
}

int orbfe_nccl_comm_create(int device, int nranks, int rank, const uint8_t id128[128], void** comm) {
    NcclApi* api = nullptr;
    int rc = nccl_api(&api);
    if (rc) return rc;
    if (!id128 || !comm || nranks < 1 || rank < 0 || rank >= nranks) { orbfe_set_error("bad argument"); return ORBFE_E_INVALID; }
    ORBFE_CUDA(cudaSetDevice(device));
    nccl_uid u;
    memcpy(u.internal, id128, 128);
    *comm = nullptr;
    rc = api->initRank(comm, nranks, u, rank);
    if (rc) return nccl_fail(api, rc, "ncclCommInitRank");
    return ORBFE_OK;
}

int orbfe_nccl_comm_destroy(void* comm) {
    NcclApi* api = nullptr;
    int rc = nccl_api(&api);
    if (rc) return rc;
    if (comm) { rc = api->destroy(comm); if (rc) return nccl_fail(api, rc, "ncclCommDestroy"); }
    return ORBFE_OK;
}

int orbfe_hamming_top2_sharded(orbfe_matcher* m, void* nccl_comm, int nranks, const uint8_t* d_queries, int nq, int th, float ratio, int strict_lt,
                               orbfe_match* d_out, void* stream) {
    if (!m || !d_queries || !d_out || nq < 1 || nranks < 1 || (nranks > 1 && !nccl_comm)) { orbfe_set_error("bad argument"); return ORBFE_E_INVALID; }
    if (nq > m->maxQ) { orbfe_set_error("%d queries exceed capacity %d", nq, m->maxQ); return ORBFE_E_CAPACITY; }
    if (((uintptr_t)d_queries & 15) != 0) { orbfe_set_error("query pointer must be 16-byte aligned"); return ORBFE_E_INVALID; }
    ORBFE_CUDA(cudaSetDevice(m->device));
    cudaStream_t st = stream ? (cudaStream_t)stream : m->stream;
    if (nranks == 1 && !nccl_comm) return launch_top2(d_queries, nq, m->dDb, m->ndb, m->indexBase, m->dParts, m->maxSlices, th, ratio, 0.0, strict_lt ? 1 : 0, d_out, st);
    NcclApi* api = nullptr;
    int rc = nccl_api(&api);
    if (rc) return rc;
    if (m->gatherRanks < nranks) {   /* first sharded call (or a larger communicator): the gather buffers */
        ORBFE_CUDA(cudaStreamSynchronize(st));
        cudaFree(m->dGather); m->dGather = nullptr;
        if (!m->dLocal) ORBFE_CUDA(cudaMalloc((void**)&m->dLocal, (size_t)m->maxQ * sizeof(orbfe_match)));
        ORBFE_CUDA(cudaMalloc((void**)&m->dGather, (size_t)nranks * m->maxQ * sizeof(orbfe_match)));
        m->gatherRanks = nranks;
    }
    /* this shard's top-2 (global row indices; the threshold verdict is recomputed by the merge) */
    rc = launch_top2(d_queries, nq, m->dDb, m->ndb, m->indexBase, m->dParts, m->maxSlices, th, ratio, 0.0, strict_lt ? 1 : 0, m->dLocal, st);
    if (rc) return rc;
    /* rank r's nq records land at dGather[r*nq ..]: ascending rank = ascending global row, the order the merge scans.
     * (Measured and dropped: the queries in four groups with the all-gather + merge of one group on a side stream under the search
     * of the next — 3.76 ms against 3.20 ms on two GPUs; four short searches fill the GPU worse than the hidden exchange saves.) */
    const int nrc = api->allGather(m->dLocal, m->dGather, (size_t)nq * sizeof(orbfe_match), 0 /* ncclInt8 */, nccl_comm, st);
    if (nrc) return nccl_fail(api, nrc, "ncclAllGather");
    k_top2_merge<<<(nq + 127) / 128, 128, 0, st>>>(m->dGather, nranks, nq, th, ratio, 0.0, strict_lt ? 1 : 0, 1, 256, d_out);
    ORBFE_CUDA(cudaGetLastError());
    return ORBFE_OK;
}

int orbfe_matcher_top2(orbfe_matcher* m, const uint8_t* h_queries, int nq, int th, float ratio, int strict_lt, orbfe_match* h_out) {
    if (!m || !h_queries || !h_out || nq < 1) { orbfe_set_error("bad argument"); return ORBFE_E_INVALID; }
    if (nq > m->maxQ) { orbfe_set_error("%d queries exceed capacity %d", nq, m->maxQ); return ORBFE_E_CAPACITY; }
    ORBFE_CUDA(cudaSetDevice(m->device));
    ORBFE_CUDA(cudaMemcpyAsync(m->dQ, h_queries, (size_t)nq * 32, cudaMemcpyHostToDevice, m->stream));
    int rc = launch_top2(m->dQ, nq, m->dDb, m->ndb, m->indexBase, m->dParts, m->maxSlices, th, ratio, 0.0, strict_lt ? 1 : 0, m->dOut, m->stream);
    if (rc) return rc;
    ORBFE_CUDA(cudaMemcpyAsync(h_out, m->dOut, (size_t)nq * sizeof(orbfe_match), cudaMemcpyDeviceToHost, m->stream));
    ORBFE_CUDA(cudaStreamSynchronize(m->stream));
    return ORBFE_OK;
}

int orbfe_top2_merge_device(int device, const orbfe_match* d_parts, int nshards, int nq, int th, float ratio, int strict_lt,
                            orbfe_match* d_out, void* stream) {
    if (!d_parts || !d_out || nshards < 1 || nq < 1) { orbfe_set_error("bad argument"); return ORBFE_E_INVALID; }
    ORBFE_CUDA(cudaSetDevice(device));
    k_top2_merge<<<(nq + 127) / 128, 128, 0, (cudaStream_t)stream>>>(d_parts, nshards, nq, th, ratio, 0.0, strict_lt ? 1 : 0, 1, 256, d_out);
    ORBFE_CUDA(cudaGetLastError());
    return ORBFE_OK;
}

int orbfe_hamming_top2(int device, const uint8_t* q, int nq, const uint8_t* db, int64_t ndb, int th, float ratio, int strict_lt, orbfe_match* out) {
    orbfe_matcher* m = nullptr;
    int rc = orbfe_matcher_create(device, std::max(nq, 1), std::max<int64_t>(ndb, 1), &m);
    if (rc) return rc;
    rc = orbfe_matcher_set_db(m, db, ndb, 0);
    if (!rc) rc = orbfe_matcher_top2(m, q, nq, th, ratio, strict_lt, out);
    orbfe_matcher_destroy(m);
    return rc;
}

int orbfe_knn2_ratio(int device, const uint8_t* q, int nq, const uint8_t* t, int nt, double ratio, int32_t* match, int32_t* d0, int32_t* d1) {
    if (!q || !match || nq < 1 || nt < 0 || (nt > 0 && !t)) { orbfe_set_error("bad argument"); return ORBFE_E_INVALID; }
    orbfe_matcher* m = nullptr;
    int rc = orbfe_matcher_create(device, nq, std::max(nt, 1), &m);
    if (rc) return rc;
    rc = orbfe_matcher_set_db(m, t, nt, 0);
    orbfe_match* h = new (std::nothrow) orbfe_match[nq];
    if (!rc && h) {
        cudaSetDevice(device);
        cudaError_t e = cudaMemcpyAsync(m->dQ, q, (size_t)nq * 32, cudaMemcpyHostToDevice, m->stream);
        if (e == cudaSuccess) {
            rc = launch_top2(m->dQ, nq, m->dDb, m->ndb, 0, m->dParts, m->maxSlices, 0, 0.f, ratio, 2, m->dOut, m->stream);
            if (!rc) e = cudaMemcpyAsync(h, m->dOut, (size_t)nq * sizeof(orbfe_match), cudaMemcpyDeviceToHost, m->stream);
            if (!rc && e == cudaSuccess) e = cudaStreamSynchronize(m->stream);
        }
        if (e != cudaSuccess) rc = orbfe_cuda_fail(e, "knn2");
        if (!rc)
            for (int i = 0; i < nq; i++) {
                match[i] = h[i].accepted ? h[i].idx : -1;
                if (d0) d0[i] = nt >= 1 ? h[i].best1 : -1;
                if (d1) d1[i] = nt >= 2 ? h[i].best2 : -1;
            }
    }
    delete[] h;
    orbfe_matcher_destroy(m);
    return rc;
}

int orbfe_descriptor_distances(int device, const uint8_t* a, const uint8_t* b, int64_t n, int32_t* out) {
    if (!a || !b || !out || n < 1) { orbfe_set_error("bad argument"); return ORBFE_E_INVALID; }
    ORBFE_CUDA(cudaSetDevice(device));
    const size_t bIn = ((size_t)n * 32 + 255) & ~(size_t)255;
    if (2 * bIn + (size_t)n * 4 <= ORBFE_STAGING_MAX) {                          /* the usual size: through the thread's staging block */
        OrbfeStaging ws;
        const int rc = orbfe_thread_staging(device, 2 * bIn + (size_t)n * 4, &ws);
        if (rc != ORBFE_OK) return rc;
        memcpy(ws.h, a, (size_t)n * 32);
        memcpy(ws.h + bIn, b, (size_t)n * 32);
        cudaError_t e = cudaMemcpyAsync(ws.d, ws.h, 2 * bIn, cudaMemcpyHostToDevice, ws.stream);
        if (e == cudaSuccess) {
            k_pair_distances<<<(unsigned)((n + 255) / 256), 256, 0, ws.stream>>>((const uint4*)ws.d, (const uint4*)(ws.d + bIn), n, (int*)(ws.d + 2 * bIn));
            e = cudaGetLastError();
        }
        if (e == cudaSuccess) e = cudaMemcpyAsync(ws.h + 2 * bIn, ws.d + 2 * bIn, (size_t)n * 4, cudaMemcpyDeviceToHost, ws.stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(ws.stream);
        if (e != cudaSuccess) return orbfe_cuda_fail(e, "descriptor_distances");
        memcpy(out, ws.h + 2 * bIn, (size_t)n * 4);
        return ORBFE_OK;
    }
    uint8_t *da = nullptr, *db = nullptr;
    int* dout = nullptr;
    cudaError_t e = cudaMalloc((void**)&da, (size_t)n * 32);
    if (e == cudaSuccess) e = cudaMalloc((void**)&db, (size_t)n * 32);
    if (e == cudaSuccess) e = cudaMalloc((void**)&dout, (size_t)n * 4);
    if (e == cudaSuccess) e = cudaMemcpy(da, a, (size_t)n * 32, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(db, b, (size_t)n * 32, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) {
        k_pair_distances<<<(unsigned)((n + 255) / 256), 256>>>((const uint4*)da, (const uint4*)db, n, dout);
        e = cudaGetLastError();
    }
    if (e == cudaSuccess) e = cudaMemcpy(out, dout, (size_t)n * 4, cudaMemcpyDeviceToHost);
    cudaFree(da); cudaFree(db); cudaFree(dout);
    if (e != cudaSuccess) return orbfe_cuda_fail(e, "descriptor_distances");
    return ORBFE_OK;
}

int orbfe_distinctive_descriptors_device(int device, const uint8_t* d_desc, const int32_t* d_offsets, int npoints, int32_t* d_best_idx,
                                         int32_t* d_best_median, void* stream) {
    if (!d_desc || !d_offsets || !d_best_idx || npoints < 1) { orbfe_set_error("bad argument"); return ORBFE_E_INVALID; }
    if ((uintptr_t)d_desc & 15) { orbfe_set_error("descriptor array must be 16-byte aligned"); return ORBFE_E_INVALID; }
    ORBFE_CUDA(cudaSetDevice(device));
    k_distinctive<<<(npoints + DD_WARPS - 1) / DD_WARPS, DD_WARPS * 32, 0, (cudaStream_t)stream>>>((const uint4*)d_desc, d_offsets, npoints, d_best_idx, d_best_median);
    ORBFE_CUDA(cudaGetLastError());
    return ORBFE_OK;
}

int orbfe_distinctive_descriptors(int device, const uint8_t* desc, const int32_t* offsets, int npoints, int32_t* best_idx, int32_t* best_median) {
    if (!desc || !offsets || !best_idx || npoints < 1) { orbfe_set_error("bad argument"); return ORBFE_E_INVALID; }
    for (int p = 0; p < npoints; p++)
        if (offsets[p + 1] < offsets[p] || offsets[p + 1] - offsets[p] > 65535) { orbfe_set_error("map point %d: bad row range (at most 65535 observations)", p); return ORBFE_E_INVALID; }
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || device < 0 || device >= ndev) {
        orbfe_set_error("no usable CUDA device %d (found %d); this library has no CPU path", device, ndev);
        return ORBFE_E_CUDA;
    }
    ORBFE_CUDA(cudaSetDevice(device));
    const size_t rows = (size_t)offsets[npoints], bDesc = (rows * 32 + 255) & ~(size_t)255, bOff = ((size_t)(npoints + 1) * 4 + 255) & ~(size_t)255;
    const size_t bOut = ((size_t)npoints * 4 + 255) & ~(size_t)255;
    if (bDesc + bOff + 2 * bOut <= ORBFE_STAGING_MAX) {                           /* the usual size: through the thread's staging block */
        OrbfeStaging ws;
        int rc = orbfe_thread_staging(device, bDesc + bOff + 2 * bOut, &ws);
        if (rc != ORBFE_OK) return rc;
        if (rows) memcpy(ws.h, desc, rows * 32);
        memcpy(ws.h + bDesc, offsets, (size_t)(npoints + 1) * 4);
        cudaError_t e = cudaMemcpyAsync(ws.d, ws.h, bDesc + bOff, cudaMemcpyHostToDevice, ws.stream);
        if (e == cudaSuccess)
            rc = orbfe_distinctive_descriptors_device(device, ws.d, (const int32_t*)(ws.d + bDesc), npoints, (int32_t*)(ws.d + bDesc + bOff),
                                                      (int32_t*)(ws.d + bDesc + bOff + bOut), ws.stream);
        if (rc == ORBFE_OK && e == cudaSuccess) e = cudaMemcpyAsync(ws.h + bDesc + bOff, ws.d + bDesc + bOff, 2 * bOut, cudaMemcpyDeviceToHost, ws.stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(ws.stream);
        if (rc != ORBFE_OK) return rc;
        if (e != cudaSuccess) return orbfe_cuda_fail(e, "distinctive_descriptors");
        memcpy(best_idx, ws.h + bDesc + bOff, (size_t)npoints * 4);
        if (best_median) memcpy(best_median, ws.h + bDesc + bOff + bOut, (size_t)npoints * 4);
        return ORBFE_OK;
    }
    uint8_t* d = nullptr;
    ORBFE_CUDA(cudaMalloc((void**)&d, bDesc + bOff + 2 * bOut + 256));
    cudaError_t e = rows ? cudaMemcpy(d, desc, rows * 32, cudaMemcpyHostToDevice) : cudaSuccess;
    if (e == cudaSuccess) e = cudaMemcpy(d + bDesc, offsets, (size_t)(npoints + 1) * 4, cudaMemcpyHostToDevice);
    int rc = ORBFE_OK;
    if (e == cudaSuccess) rc = orbfe_distinctive_descriptors_device(device, d, (const int32_t*)(d + bDesc), npoints, (int32_t*)(d + bDesc + bOff), (int32_t*)(d + bDesc + bOff + bOut), nullptr);
    if (rc == ORBFE_OK && e == cudaSuccess) e = cudaMemcpy(best_idx, d + bDesc + bOff, (size_t)npoints * 4, cudaMemcpyDeviceToHost);
    if (rc == ORBFE_OK && e == cudaSuccess && best_median) e = cudaMemcpy(best_median, d + bDesc + bOff + bOut, (size_t)npoints * 4, cudaMemcpyDeviceToHost);
    cudaFree(d);
    if (rc != ORBFE_OK) return rc;
    if (e != cudaSuccess) return orbfe_cuda_fail(e, "distinctive_descriptors");
    return ORBFE_OK;
}

}  /* extern "C" */

/* ------------------------------------------------------------------------------------- POPC throughput probe */
/* What the integer roofline of the Hamming search is read against (SURVEY.md 8d asks for a measured figure, not the nominal
 * 16 POPC/clk/SM): every thread runs eight independent POPC chains; the adds that keep the chains alive go to the ALU pipe. */
__global__ void __launch_bounds__(256) k_probe_popc(uint32_t* sink, int iters) {
    uint32_t x[8], acc[8];
#pragma unroll
    for (int j = 0; j < 8; j++) { x[j] = (threadIdx.x + 1u) * 2654435761u + j * 40503u + blockIdx.x; acc[j] = 0; }
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int j = 0; j < 8; j++) { acc[j] += __popc(x[j]); x[j] += acc[j]; }
    }
    uint32_t t = 0;
#pragma unroll
    for (int j = 0; j < 8; j++) t ^= acc[j] + x[j];
    if (t == 0x12345678u) sink[0] = t;   /* never true in practice; keeps the chains from being removed */
}

extern "C" int orbfe_probe_popc(int device, double* popc_per_second) {
    if (!popc_per_second) { orbfe_set_error("bad argument"); return ORBFE_E_INVALID; }
    ORBFE_CUDA(cudaSetDevice(device));
    uint32_t* sink = nullptr;
    ORBFE_CUDA(cudaMalloc((void**)&sink, 4));
    cudaEvent_t e0, e1;
    ORBFE_CUDA(cudaEventCreate(&e0));
    ORBFE_CUDA(cudaEventCreate(&e1));
    const int iters = 4096, blocks = 148 * 8;
    k_probe_popc<<<blocks, 256>>>(sink, iters);   /* warm-up */
    double best = 0.0;
    for (int rep = 0; rep < 5; rep++) {
        cudaEventRecord(e0);
        k_probe_popc<<<blocks, 256>>>(sink, iters);
        cudaEventRecord(e1);
        cudaError_t e = cudaEventSynchronize(e1);
        float ms = 0.f;
        if (e == cudaSuccess) e = cudaEventElapsedTime(&ms, e0, e1);
        if (e != cudaSuccess) { cudaFree(sink); return orbfe_cuda_fail(e, "popc probe"); }
        best = std::max(best, (double)blocks * 256.0 * iters * 8.0 / (ms * 1e-3));
    }
    cudaEventDestroy(e0); cudaEventDestroy(e1); cudaFree(sink);
    *popc_per_second = best;
    return ORBFE_OK;
}

/* ------------------------------------------------------------------------- top-2 search without any allocation */
/* orbfe_hamming_top2_static: the same top-2 search as orbfe_matcher_top2_device for callers that must not, or
 * cannot, allocate device memory.  In ORB-SLAM3 terms that is the per-frame Tracking thread: it has a frame-period
 * deadline and shares the GPU with the LocalMapping and LoopClosing threads, while cudaMalloc/cudaFree take a
 * device-wide lock and cudaFree waits for the whole device — an allocation on that thread stalls the others.  It is
 * also any host process whose memory pool already owns all of the device memory, where cudaMalloc simply fails.
 * The search needs scratch memory (one partial top-2 record per (database slice, query)); here that scratch is a
 * fixed array that lives in the library image instead of being allocated: 64 slices x 4096 queries x 16 bytes
 * = 4 MiB, held by every GPU that loads liborbfe.so until the process exits.
 *
 * g_static_parts is a device symbol.  In host code its name is only a stand-in, not a device address, so the
 * launch code below never passes it to a kernel directly: once per device, with that device current, it asks the
 * runtime for the real address (cudaGetSymbolAddress) and caches it. */
#define ORBFE_STATIC_MAX_QUERIES 4096
#define ORBFE_STATIC_MAX_SLICES 64
static __device__ __align__(256) orbfe_match g_static_parts[(size_t)ORBFE_STATIC_MAX_SLICES * ORBFE_STATIC_MAX_QUERIES];

static std::mutex g_static_mu;
static orbfe_match* g_static_addr[64]; /* per device ordinal, filled on first use */

extern "C" int orbfe_hamming_top2_static(int device, const uint8_t* d_queries, int nq, const uint8_t* d_db, int64_t ndb,
                                         int64_t index_base, int th, float ratio, int strict_lt, orbfe_match* d_out, void* stream) {
    if (!d_queries || !d_out || nq < 1 || ndb < 0 || (ndb > 0 && !d_db)) { orbfe_set_error("bad argument"); return ORBFE_E_INVALID; }
    if (nq > ORBFE_STATIC_MAX_QUERIES) {
        orbfe_set_error("orbfe_hamming_top2_static: %d queries exceed the built-in scratch (max %d); use orbfe_matcher_top2_device", nq,
                        ORBFE_STATIC_MAX_QUERIES);
        return ORBFE_E_CAPACITY;
    }
    if ((((uintptr_t)d_queries) | ((uintptr_t)d_db)) & 15) { orbfe_set_error("query/database pointers must be 16-byte aligned"); return ORBFE_E_INVALID; }
    if (device < 0 || device >= 64) { orbfe_set_error("device ordinal %d out of range", device); return ORBFE_E_INVALID; }
    ORBFE_CUDA(cudaSetDevice(device));
    orbfe_match* parts = nullptr;
    {
        std::lock_guard<std::mutex> lk(g_static_mu);
        if (!g_static_addr[device]) {
            /* first call on this device: resolve the symbol's address here, with `device` current */
            orbfe_match* p = nullptr;
            ORBFE_CUDA(cudaGetSymbolAddress((void**)&p, g_static_parts));
            g_static_addr[device] = p;
        }
        parts = g_static_addr[device];
    }
    /* Every (slice, query) record the merge reads is written by the search kernel of the same call, so nothing in
     * the shared array has to start at zero and nothing is cleared; what a previous call left behind is overwritten
     * or never read. */
    return launch_top2(d_queries, nq, d_db, ndb, index_base, parts, ORBFE_STATIC_MAX_SLICES, th, ratio, 0.0, strict_lt ? 1 : 0, d_out,
                       (cudaStream_t)stream);
}
