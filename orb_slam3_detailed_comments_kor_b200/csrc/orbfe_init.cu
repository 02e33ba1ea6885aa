/* orbfe_init.cu — ORBmatcher::SearchForInitialization (/root/reference/src/ORBmatcher.cc:706-821), the matcher of the monocular
 * initialiser (Tracking::MonocularInitialization, src/Tracking.cc:2462), for batches of frame pairs; with the grid walk of
 * Frame::GetFeaturesInArea (src/Frame.cc:643-709) over a grid built like Frame::AssignFeaturesToGrid (:380-411, 711-721).
 *
 * The reference loop is serial in a stronger way than the projection searches: a keypoint of F2 that is already matched stays
 * visible to later keypoints of F1, but only at distances below the one it is held at (vMatchedDistance, :745), and a better match
 * takes it away from its holder (:764-768).  What keypoint q of F1 sees therefore depends on what every earlier keypoint did.
 *
 * The kernel computes the same result as a fixed point.  State: for every level-0 keypoint q of F1 the F2 keypoint it is accepted
 * for and at which distance (or nothing).  One round recomputes every q against the state of the round before: the distance a
 * candidate i2 is held at when q's turn comes is the distance of the last writer q' < q of i2 (writers of one i2 have strictly
 * decreasing distances, so the last one is also the smallest).  By induction the first r keypoints are right after r rounds and
 * never change again, so the iteration ends at the serial result; in practice a handful of rounds (a change can only propagate to
 * later keypoints, so a round starts behind the first keypoint that changed in the round before).  One CTA per pair.
 * Afterwards: vnMatches12[q] survives iff q is the last writer of its keypoint; the rotation histogram counts every writer, displaced
 * ones included, as the reference does (:783 pushes at acceptance time and never removes). */
#include <cuda_runtime.h>
#include <limits.h>
#include <stdint.h>

#include "orbfe_internal.h"

#define SI_NT 1024
#define SI_MAXKP 8192   /* Tracking's initial extractor asks for 5 x nFeatures (src/Tracking.cc:1151-1157) */
#define SI_COLS 64
#define SI_ROWS 48
#define SI_CELLS (SI_COLS * SI_ROWS)
#define SI_CACHE 1536
#define SI_G 8          /* lanes that share one keypoint of F1 in a round */

struct SiSmem {
    int cellEnd[SI_CELLS];       /* end of each cell's run in `sorted` */
    int wEnd[SI_MAXKP];          /* per F2 keypoint: end of its run of writers in `wq` (start = end of the keypoint before) */
    uint32_t state[2][SI_MAXKP]; /* per F1 keypoint: 0 = not accepted, else (distance << 16) | (F2 row + 1); two copies (round r-1, round r) */
    uint16_t sorted[SI_MAXKP];   /* level-0 rows of F2 by cell, ascending inside a cell */
    uint16_t wq[SI_MAXKP];       /* writers (F1 rows) grouped by the F2 row they write, ascending inside a group */
    uint16_t qList[SI_MAXKP];    /* the level-0 rows of F1, ascending */
    int warpSum[SI_NT / 32];
    int rot[32];
    /* the level-0 keypoints of F2 in `sorted` order, when there are at most SI_CACHE of them (the usual case: a fifth of the keypoints are
     * on level 0): every round reads every candidate's position and descriptor, 150 candidates per query, and from global memory that
     * is L2 bandwidth (8 MB per round and pair) */
    uint4 cDesc[SI_CACHE][2];
    float2 cXy[SI_CACHE];
};

static __device__ __forceinline__ int si_ham(const uint4 qa, const uint4 qb, const uint4 a, const uint4 b) {
    return __popc(qa.x ^ a.x) + __popc(qa.y ^ a.y) + __popc(qa.z ^ a.z) + __popc(qa.w ^ a.w) + __popc(qb.x ^ b.x) + __popc(qb.y ^ b.y) +
           __popc(qb.z ^ b.z) + __popc(qb.w ^ b.w);
}

/* exclusive scan of a[0..n) in place (n <= SI_MAXKP, multiple-of-SI_NT padding not required); all threads call it */
static __device__ void si_exclusive_scan(int* a, int n, int* warpSum) {
    const int tid = threadIdx.x;
    const int per = (n + SI_NT - 1) / SI_NT, base = tid * per;
    int sum = 0;
    for (int i = 0; i < per; i++) sum += (base + i < n) ? a[base + i] : 0;
    int incl = sum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, incl, o);
        if ((tid & 31) >= o) incl += t;
    }
    if ((tid & 31) == 31) warpSum[tid >> 5] = incl;
    __syncthreads();
    int run = incl - sum;
    for (int w = 0; w < (tid >> 5); w++) run += warpSum[w];
    for (int i = 0; i < per && base + i < n; i++) {
        const int c = a[base + i];
        a[base + i] = run;
        run += c;
    }
    __syncthreads();
}

/* counting sort: item k (0..nItems) with bucket key(k) >= 0 goes into out[], grouped by bucket, ascending item value inside a
 * bucket; ends[b] becomes the end of bucket b's run.  value(k) is what is stored. */
template <class Key, class Val>
static __device__ void si_group(int* ends, int nBuckets, uint16_t* out, int nItems, int* warpSum, Key&& key, Val&& value) {
    const int tid = threadIdx.x;
    for (int b = tid; b < nBuckets; b += SI_NT) ends[b] = 0;
    __syncthreads();
    for (int k = tid; k < nItems; k += SI_NT) {
        const int b = key(k);
        if (b >= 0) atomicAdd(&ends[b], 1);
    }
    __syncthreads();
    si_exclusive_scan(ends, nBuckets, warpSum);
    for (int k = tid; k < nItems; k += SI_NT) {
        const int b = key(k);
        if (b >= 0) out[atomicAdd(&ends[b], 1)] = (uint16_t)value(k);
    }
    __syncthreads();
    for (int b = tid; b < nBuckets; b += SI_NT) {   /* runs are a few entries long */
        const int beg = b ? ends[b - 1] : 0, end = ends[b];
        for (int i = beg + 1; i < end; i++) {
            const uint16_t v = out[i];
            int j = i - 1;
            while (j >= beg && out[j] > v) { out[j + 1] = out[j]; j--; }
            out[j + 1] = v;
        }
    }
    __syncthreads();
}

__global__ void __launch_bounds__(SI_NT) k_search_for_initialization(const orbfe_keypoint* __restrict__ kps1All, const uint8_t* __restrict__ desc1All,
                                                                     const int32_t* __restrict__ off1, const orbfe_keypoint* __restrict__ kps2All,
                                                                     const uint8_t* __restrict__ desc2All, const int32_t* __restrict__ off2,
                                                                     const orbfe_frame_grid* __restrict__ grids, float* __restrict__ prevAll, float window,
                                                                     float nnratio, int checkOri, int thLow, int32_t* __restrict__ matchesAll,
                                                                     int32_t* __restrict__ nmatchesAll, int32_t* __restrict__ roundsAll) {
    extern __shared__ __align__(16) unsigned char si_raw[];
    SiSmem& s = *reinterpret_cast<SiSmem*>(si_raw);
    __shared__ int s_nq, s_minChanged, s_nm;
    const int tid = threadIdx.x, f = blockIdx.x;
    const int a0 = off1[f], n1 = off1[f + 1] - a0, b0 = off2[f], n2 = off2[f + 1] - b0;
    const orbfe_keypoint* kps1 = kps1All + a0;
    const orbfe_keypoint* kps2 = kps2All + b0;
    const uint4* desc1 = reinterpret_cast<const uint4*>(desc1All) + (size_t)a0 * 2;
    const uint4* desc2 = reinterpret_cast<const uint4*>(desc2All) + (size_t)b0 * 2;
    float* prev = prevAll + (size_t)a0 * 2;
    int32_t* matches = matchesAll + a0;
    if (n1 < 0 || n2 < 0 || n1 > SI_MAXKP || n2 > SI_MAXKP) {   /* refused: the caller sees -1 */
        if (tid == 0) { nmatchesAll[f] = -1; if (roundsAll) roundsAll[f] = 0; }
        return;
    }
    const orbfe_frame_grid g = grids[f];

    /* ---- F2's grid, level-0 keypoints only (GetFeaturesInArea is called with minLevel = maxLevel = 0, :726) ---- */
    si_group(s.cellEnd, SI_CELLS, s.sorted, n2, s.warpSum,
             [&](int i) {
                 if (__ldg(&kps2[i].octave) != 0) return -1;
                 const int px = (int)roundf((__ldg(&kps2[i].x) - g.min_x) * g.inv_w), py = (int)roundf((__ldg(&kps2[i].y) - g.min_y) * g.inv_h);
                 return (px >= 0 && px < SI_COLS && py >= 0 && py < SI_ROWS) ? px * SI_ROWS + py : -1;
             },
             [](int i) { return i; });
    const int n2lvl0 = s.cellEnd[SI_CELLS - 1];
    const bool cached = n2lvl0 <= SI_CACHE;
    if (cached)
        for (int j = tid; j < n2lvl0; j += SI_NT) {
            const int i2 = s.sorted[j];
            s.cDesc[j][0] = __ldg(desc2 + 2 * i2); s.cDesc[j][1] = __ldg(desc2 + 2 * i2 + 1);
            s.cXy[j] = make_float2(__ldg(&kps2[i2].x), __ldg(&kps2[i2].y));
        }
    /* ---- the level-0 keypoints of F1, in order ---- */
    if (tid == 0) { s_nq = 0; s_minChanged = -1; }
    for (int i = tid; i < n1; i += SI_NT) { s.state[0][i] = 0; s.state[1][i] = 0; matches[i] = -1; }
    __syncthreads();
    {   /* ordered compaction: chunks of SI_NT rows */
        for (int c0 = 0; c0 < n1; c0 += SI_NT) {
            const int i = c0 + tid;
            const bool is0 = i < n1 && __ldg(&kps1[i].octave) <= 0;   /* `if(level1>0) continue;` */
            const unsigned bal = __ballot_sync(0xffffffffu, is0);
            if ((tid & 31) == 0) s.warpSum[tid >> 5] = __popc(bal);
            __syncthreads();
            int base = s_nq;
            for (int w = 0; w < (tid >> 5); w++) base += s.warpSum[w];
            if (is0) s.qList[base + __popc(bal & ((1u << (tid & 31)) - 1))] = (uint16_t)i;
            __syncthreads();
            if (tid == 0) { int t = 0; for (int w = 0; w < SI_NT / 32; w++) t += s.warpSum[w]; s_nq += t; }
            __syncthreads();
        }
    }
    const int nq = s_nq;
    const float r = window;
    int cur = 0, rounds = 0;
    for (;;) {
        const uint32_t* st = s.state[cur];
        uint32_t* nx = s.state[cur ^ 1];
        /* writers of every F2 keypoint under the current state, ascending */
        si_group(s.wEnd, n2, s.wq, nq, s.warpSum,
                 [&](int k) { const uint32_t v = st[s.qList[k]]; return v ? (int)(v & 0xFFFFu) - 1 : -1; },
                 [&](int k) { return (int)s.qList[k]; });
        const int firstDirty = s_minChanged;   /* keypoints up to here cannot change in this round */
        __syncthreads();
        if (tid == 0) s_minChanged = INT_MAX;
        __syncthreads();
        /* SI_G lanes share a keypoint of F1: its window spans a dozen or more grid columns with very unequal numbers of keypoints, and with a
         * thread per keypoint most lanes of a warp wait for its longest window.  Each lane takes every SI_G-th column; the reference's update
         * rule (`dist < bestDist`: the old best becomes the second; `else if dist < bestDist2`) keeps the two smallest candidates in the order
         * (distance, position in the walk), so a lane keeps two keys distance << 16 | position and the group merges them. */
        int myMin = INT_MAX;
        const int gl = tid & (SI_G - 1), leader = tid & 31 & ~(SI_G - 1);
        const unsigned gmask = ((1u << SI_G) - 1u) << leader;
        for (int k = tid / SI_G; k < nq; k += SI_NT / SI_G) {
            const int q = s.qList[k];
            if (rounds > 0 && q <= firstDirty) { if (gl == 0) nx[q] = st[q]; continue; }
            const float x = prev[2 * q], y = prev[2 * q + 1];
            unsigned k1 = 0xFFFFFFFFu, k2 = 0xFFFFFFFFu;
            const float ax = x - g.min_x, ay = y - g.min_y;
            const int x0 = max(0, (int)floorf((ax - r) * g.inv_w)), x1 = min(SI_COLS - 1, (int)ceilf((ax + r) * g.inv_w));
            const int y0 = max(0, (int)floorf((ay - r) * g.inv_h)), y1 = min(SI_ROWS - 1, (int)ceilf((ay + r) * g.inv_h));
            if (x0 < SI_COLS && x1 >= 0 && y0 < SI_ROWS && y1 >= 0) {
                const uint4 qa = __ldg(desc1 + 2 * q), qb = __ldg(desc1 + 2 * q + 1);
                for (int ix = x0 + gl; ix <= x1; ix += SI_G) {
                    const int c0 = ix * SI_ROWS + y0, c1 = ix * SI_ROWS + y1;
                    const int beg = c0 ? s.cellEnd[c0 - 1] : 0, end = s.cellEnd[c1];
                    for (int j = beg; j < end; j++) {
                        const int i2 = s.sorted[j];
                        const float2 xy = cached ? s.cXy[j] : make_float2(__ldg(&kps2[i2].x), __ldg(&kps2[i2].y));
                        if (!(fabsf(xy.x - x) < r && fabsf(xy.y - y) < r)) continue;
                        const int dist = cached ? si_ham(qa, qb, s.cDesc[j][0], s.cDesc[j][1]) : si_ham(qa, qb, __ldg(desc2 + 2 * i2), __ldg(desc2 + 2 * i2 + 1));
                        int held = INT_MAX;   /* vMatchedDistance[i2] when q's turn comes */
                        for (int w = i2 ? s.wEnd[i2 - 1] : 0, we = s.wEnd[i2]; w < we; w++) {
                            const int qw = s.wq[w];
                            if (qw >= q) break;
                            held = (int)(st[qw] >> 16);
                        }
                        if (held <= dist) continue;
                        const unsigned key = ((unsigned)dist << 16) | (unsigned)j;
                        if (key < k1) { k2 = k1; k1 = key; }
                        else if (key < k2) k2 = key;
                    }
                }
            }
#pragma unroll
            for (int o = SI_G >> 1; o; o >>= 1) {
                const unsigned o1 = __shfl_xor_sync(gmask, k1, o), o2 = __shfl_xor_sync(gmask, k2, o);
                const unsigned lo = min(k1, o1), hi = max(k1, o1);
                k2 = min(hi, min(k2, o2));
                k1 = lo;
            }
            if (gl == 0) {
                const int bestDist = k1 == 0xFFFFFFFFu ? INT_MAX : (int)(k1 >> 16), bestDist2 = k2 == 0xFFFFFFFFu ? INT_MAX : (int)(k2 >> 16);
                uint32_t v = 0;
                if (k1 != 0xFFFFFFFFu && bestDist <= thLow && (float)bestDist < __fmul_rn((float)bestDist2, nnratio))
                    v = ((uint32_t)bestDist << 16) | (uint32_t)(s.sorted[k1 & 0xFFFFu] + 1);
                nx[q] = v;
                if (v != st[q]) myMin = min(myMin, q);
            }
        }
        if (myMin != INT_MAX) atomicMin(&s_minChanged, myMin);
        __syncthreads();
        rounds++;
        if (s_minChanged == INT_MAX) break;   /* state[cur] == state[cur ^ 1]; wq / wEnd describe it */
        cur ^= 1;
        __syncthreads();
    }
    /* ---- vnMatches12, nmatches, rotation histogram ---- */
    const uint32_t* st = s.state[cur];
    if (tid < 32) s.rot[tid] = 0;
    if (tid == 0) s_nm = 0;
    __syncthreads();
    const float factor = 1.0f / 30.0f;
    int mine = 0;
    for (int k = tid; k < nq; k += SI_NT) {
        const int q = s.qList[k];
        const uint32_t v = st[q];
        if (!v) continue;
        const int i2 = (int)(v & 0xFFFFu) - 1;
        if (s.wq[s.wEnd[i2] - 1] == q) { matches[q] = i2; mine++; }   /* the last writer holds the keypoint */
        if (checkOri) {
            float rot = __ldg(&kps1[q].angle) - __ldg(&kps2[i2].angle);
            if (rot < 0.0f) rot += 360.0f;
            int bin = (int)roundf(rot * factor);
            if (bin == 30) bin = 0;
            if (bin >= 0 && bin < 30) atomicAdd(&s.rot[bin], 1);
        }
    }
    if (mine) atomicAdd(&s_nm, mine);
    __syncthreads();
    if (checkOri) {
        if (tid == 0) {   /* ComputeThreeMaxima (ORBmatcher.cc:2545-2586) */
            int ind1 = -1, ind2 = -1, ind3 = -1, max1 = 0, max2 = 0, max3 = 0;
            for (int i = 0; i < 30; i++) {
                const int c = s.rot[i];
                if (c > max1) { max3 = max2; max2 = max1; max1 = c; ind3 = ind2; ind2 = ind1; ind1 = i; }
                else if (c > max2) { max3 = max2; max2 = c; ind3 = ind2; ind2 = i; }
                else if (c > max3) { max3 = c; ind3 = i; }
            }
            if ((float)max2 < 0.1f * (float)max1) { ind2 = -1; ind3 = -1; }
            else if ((float)max3 < 0.1f * (float)max1) { ind3 = -1; }
            s.rot[30] = ind1; s.rot[31] = (ind2 & 0xFF) | ((ind3 & 0xFF) << 8);
        }
        __syncthreads();
        const int ind1 = s.rot[30], ind2 = (int)(signed char)(s.rot[31] & 0xFF), ind3 = (int)(signed char)((s.rot[31] >> 8) & 0xFF);
        int removed = 0;
        for (int k = tid; k < nq; k += SI_NT) {
            const int q = s.qList[k];
            const uint32_t v = st[q];
            if (!v || matches[q] < 0) continue;
            float rot = __ldg(&kps1[q].angle) - __ldg(&kps2[(int)(v & 0xFFFFu) - 1].angle);
            if (rot < 0.0f) rot += 360.0f;
            int bin = (int)roundf(rot * factor);
            if (bin == 30) bin = 0;
            if (bin >= 0 && bin < 30 && bin != ind1 && bin != ind2 && bin != ind3) { matches[q] = -1; removed++; }
        }
        if (removed) atomicSub(&s_nm, removed);
        __syncthreads();
    }
    /* ---- vbPrevMatched (:816-818) ---- */
    for (int k = tid; k < nq; k += SI_NT) {
        const int q = s.qList[k], m = matches[q];
        if (m >= 0) { prev[2 * q] = __ldg(&kps2[m].x); prev[2 * q + 1] = __ldg(&kps2[m].y); }
    }
    if (tid == 0) { nmatchesAll[f] = s_nm; if (roundsAll) roundsAll[f] = rounds; }
}

extern "C" {

int orbfe_search_for_initialization_device(int device, int npairs, const orbfe_keypoint* d_kps1, const uint8_t* d_desc1, const int32_t* d_off1,
                                           const orbfe_keypoint* d_kps2, const uint8_t* d_desc2, const int32_t* d_off2, const orbfe_frame_grid* d_grids,
                                           float* d_prev_matched, int window, float nnratio, int check_orientation, int th_low, int32_t* d_matches12,
                                           int32_t* d_nmatches, int32_t* d_rounds, void* stream) {
    if (npairs < 1 || !d_kps1 || !d_desc1 || !d_off1 || !d_kps2 || !d_desc2 || !d_off2 || !d_grids || !d_prev_matched || !d_matches12 || !d_nmatches) {
        orbfe_set_error("bad argument");
        return ORBFE_E_INVALID;
    }
    if (((uintptr_t)d_desc1 | (uintptr_t)d_desc2) & 15) { orbfe_set_error("descriptor arrays must be 16-byte aligned"); return ORBFE_E_INVALID; }
    if (th_low < 0 || th_low > 256) { orbfe_set_error("th_low outside [0, 256]"); return ORBFE_E_INVALID; }
    ORBFE_CUDA(cudaSetDevice(device));
    static thread_local int configured = -1;
    if (configured != device) {
        ORBFE_CUDA(cudaFuncSetAttribute(k_search_for_initialization, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(SiSmem)));
        configured = device;
    }
    k_search_for_initialization<<<npairs, SI_NT, sizeof(SiSmem), (cudaStream_t)stream>>>(d_kps1, d_desc1, d_off1, d_kps2, d_desc2, d_off2, d_grids, d_prev_matched,
                                                                                        (float)window, nnratio, check_orientation ? 1 : 0, th_low, d_matches12,
                                                                                        d_nmatches, d_rounds);
    ORBFE_CUDA(cudaGetLastError());
    return ORBFE_OK;
}

int orbfe_search_for_initialization(int device, const orbfe_keypoint* kps1, const uint8_t* desc1, int n1, const orbfe_keypoint* kps2, const uint8_t* desc2, int n2,
                                    const orbfe_frame_grid* grid, float* prev_matched, int window, float nnratio, int check_orientation, int th_low,
                                    int32_t* matches12, int* nmatches) {
    if (!grid || n1 < 0 || n2 < 0 || (n1 && (!kps1 || !desc1 || !prev_matched || !matches12)) || (n2 && (!kps2 || !desc2))) { orbfe_set_error("bad argument"); return ORBFE_E_INVALID; }
    if (n1 > SI_MAXKP || n2 > SI_MAXKP) { orbfe_set_error("at most %d keypoints per frame", SI_MAXKP); return ORBFE_E_CAPACITY; }
    if (nmatches) *nmatches = 0;
    if (n1 == 0) return ORBFE_OK;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || device < 0 || device >= ndev) {
        orbfe_set_error("no usable CUDA device %d (found %d); this library has no CPU path", device, ndev);
        return ORBFE_E_CUDA;
    }
    /* the thread's staging block: inputs | vbPrevMatched (in and out) | outputs */
    size_t off = 0;
    auto take = [&](size_t bytes) { const size_t o = off; off += (bytes + 255) & ~(size_t)255; return o; };
    const size_t oK1 = take((size_t)n1 * sizeof(orbfe_keypoint)), oD1 = take((size_t)n1 * 32), oK2 = take((size_t)n2 * sizeof(orbfe_keypoint)), oD2 = take((size_t)n2 * 32);
    const size_t oGrid = take(sizeof(orbfe_frame_grid)), oOff = take(4 * sizeof(int32_t)), oPrev = take((size_t)n1 * 8);
    const size_t inEnd = off;
    const size_t oM = take((size_t)n1 * 4), oRes = take(2 * sizeof(int32_t));
    OrbfeStaging ws;
    int rc = orbfe_thread_staging(device, off, &ws);
    if (rc != ORBFE_OK) return rc;
    uint8_t* h = ws.h;
    uint8_t* d = ws.d;
    memcpy(h + oK1, kps1, (size_t)n1 * sizeof(orbfe_keypoint));
    memcpy(h + oD1, desc1, (size_t)n1 * 32);
    if (n2) { memcpy(h + oK2, kps2, (size_t)n2 * sizeof(orbfe_keypoint)); memcpy(h + oD2, desc2, (size_t)n2 * 32); }
    memcpy(h + oGrid, grid, sizeof(orbfe_frame_grid));
    const int32_t offs[4] = {0, n1, 0, n2};
    memcpy(h + oOff, offs, sizeof(offs));
    memcpy(h + oPrev, prev_matched, (size_t)n1 * 8);
    cudaError_t e = cudaMemcpyAsync(d, h, inEnd, cudaMemcpyHostToDevice, ws.stream);
    if (e == cudaSuccess)
        rc = orbfe_search_for_initialization_device(device, 1, (const orbfe_keypoint*)(d + oK1), d + oD1, (const int32_t*)(d + oOff), (const orbfe_keypoint*)(d + oK2),
                                                    d + oD2, (const int32_t*)(d + oOff) + 2, (const orbfe_frame_grid*)(d + oGrid), (float*)(d + oPrev), window, nnratio,
                                                    check_orientation, th_low, (int32_t*)(d + oM), (int32_t*)(d + oRes), (int32_t*)(d + oRes) + 1, ws.stream);
    if (rc == ORBFE_OK && e == cudaSuccess) e = cudaMemcpyAsync(h + oPrev, d + oPrev, off - oPrev, cudaMemcpyDeviceToHost, ws.stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(ws.stream);
    if (rc != ORBFE_OK) return rc;
    if (e != cudaSuccess) return orbfe_cuda_fail(e, "search_for_initialization");
    memcpy(matches12, h + oM, (size_t)n1 * 4);
    memcpy(prev_matched, h + oPrev, (size_t)n1 * 8);
    if (nmatches) *nmatches = *(const int32_t*)(h + oRes);
    return ORBFE_OK;
}

}  /* extern "C" */
