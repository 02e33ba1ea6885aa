/* orbfe_project.cu — projection-gated descriptor search: the searches of Tracking::SearchLocalPoints and
 * Tracking::TrackWithMotionModel,
 *   ORBmatcher::SearchByProjection(Frame&, const vector<MapPoint*>&, th, bFarPoints, thFarPoints)   /root/reference/src/ORBmatcher.cc:44-214
 *   ORBmatcher::SearchByProjection(Frame &CurrentFrame, const Frame &LastFrame, th, bMono)          /root/reference/src/ORBmatcher.cc:2193-2419
 *     (best match only, then the rotation-consistency histogram with ComputeThreeMaxima, :2545-2586)
 * together with the frame grid they walk,
 *   Frame::AssignFeaturesToGrid / PosInGrid   /root/reference/src/Frame.cc:380-411, 711-721   (64 x 48 cells, include/Frame.h:37-38)
 *   Frame::GetFeaturesInArea                  /root/reference/src/Frame.cc:643-709
 *
 * One CTA = one frame; frames of a batch are independent.  The grid (both cameras' halves for a fisheye rig) is built in
 * shared memory as a counting sort whose cell lists are ascending keypoint rows, i.e. the reference's push_back order.
 *
 * The reference loop is serial: a map point that gets matched occupies its keypoint, and the map points after it skip that
 * keypoint.  The kernel reproduces that result exactly with rounds of speculation.  In every round each unresolved query
 *   A. stamps its index (atomicMin) on every keypoint it could still take: the candidates of its window and their stereo
 *      partners;
 *   B. scans its window against the keypoints taken so far, with the reference's best/second update rule, and becomes final
 *      iff no keypoint that can change its outcome — a candidate whose distance is <= its second-best distance, or the
 *      partner of its best — carries the stamp of an earlier unresolved query (the outcome of the scan depends only on the
 *      candidates at or below the second-best distance).  The keypoints it takes become visible in the next round.
 * The earliest unresolved query always becomes final, so the loop ends; with the small windows of tracking (radius 2.5-4 px
 * times the level scale) almost everything resolves in the first rounds. */
#include <cuda_runtime.h>
#include <limits.h>
#include <stdint.h>

#include "orbfe_internal.h"

#define PJ_MAXKP 4096
#define PJ_COLS 64
#define PJ_ROWS 48
#define PJ_CELLS (PJ_COLS * PJ_ROWS)
#define PJ_PENDING (-1)
#define PJ_FREE16 0xFFFFu
#define PJ_MAXQ 65534       /* queries per frame: their indices + 1 live in 16 bits */
/* occupied as far as query q is concerned: before the search, or by a query in front of q (queries become final out of order) */
#define PJ_TAKEN_FOR(s, row, q) (((s).taken[row] & 0xFFFFu) <= (unsigned)(q))
#define PJ_FRONT 2048   /* only the earliest PJ_FRONT unresolved queries take part in a round */

__device__ int g_pjRoundsMax, g_pjRoundsSum;   /* diagnostics: rounds of the last launches (orbfe_debug_projection_rounds) */

/* Shared memory of one CTA, carved from the dynamic block by the launch's max_keypoints (the caller's promise about the frames): a
 * tracking-sized frame (1200-2000 keypoints) needs 36-44 KB, so that the two 512-thread CTAs an SM holds fit with room to spare; the full 4096 need 64 KB. */
struct PjSmem {
    int* cellEnd;        /* [PJ_CELLS, twice with rigs in the batch] end of each cell's run in `sorted` (start = end of the cell before) */
    int* owner;          /* [maxKp] */
    uint32_t* taken;     /* [maxKp] low half: 1 + the first query that took the keypoint (0: occupied before the search), PJ_FREE16 = nobody;
                            high half: the same for the takes of the running round, which become visible with the next one */
    uint16_t* sorted;    /* [maxKp] */
    int* warpSum;        /* [32] */
    /* CACHE: what a window walk reads of a keypoint, in the order of `sorted` (a window column is one contiguous run), so that the walks —
     * three per query and round, every step a dependent load — go to shared memory instead of L2 */
    uint4* descS;        /* [2 * maxKp] */
    float2* xyS;         /* [maxKp] */
    int* octS;           /* [maxKp] */
    float* urS;          /* [maxKp] mvuRight, -1 without one */
};
#define PJ_CACHE_BYTES_PER_KP 48
static __host__ __device__ inline size_t pj_smem_bytes(int maxKp, bool rigs, bool cache) {
    return (size_t)(rigs ? 2 : 1) * PJ_CELLS * 4 + (size_t)((maxKp + 3) & ~3) * (10 + (cache ? PJ_CACHE_BYTES_PER_KP : 0)) + 32 * 4;
}
/* a stereo partner row is overwritten unconditionally (src/ORBmatcher.cc:130-133), so a row can be taken more than once: what counts for a
 * query is whether ANY query before it did, i.e. the smallest taker */
static __device__ __forceinline__ void pj_take(const PjSmem& s, int row, int q) {
    uint32_t old = s.taken[row];
    for (;;) {
        const uint32_t hi = min(old >> 16, (uint32_t)(q + 1));
        const uint32_t seen = atomicCAS(&s.taken[row], old, (hi << 16) | (old & 0xFFFFu));
        if (seen == old) break;
        old = seen;
    }
}

struct PjFrame {
    const orbfe_keypoint* kps;
    const uint4* desc;
    const float* uRight;
    const int32_t* partner;
    int n, nLeft;
    float minx, miny, invw, invh;
    const float* invSigma2;
};

/* the per-candidate gate on mvuRight: |xr - mvuRight| <= radius for the SearchByProjection overloads (rectified rigs only), the
 * chi-square reprojection test of ORBmatcher::Fuse for ORBFE_PQ_ERROR_GATE queries.  true = the candidate is skipped. */
template <bool CACHE>
static __device__ __forceinline__ bool pj_gated(const PjFrame& fr, const PjSmem& s, const orbfe_proj_query& p, int row, int j, int oct, bool gateRight) {
    if ((p.flags & ORBFE_PQ_ERROR_GATE) && fr.invSigma2) {
        const float kx = CACHE ? s.xyS[j].x : __ldg(&fr.kps[row].x), ky = CACHE ? s.xyS[j].y : __ldg(&fr.kps[row].y);
        const float ex = __fsub_rn(p.x, kx), ey = __fsub_rn(p.y, ky);
        const float is2 = __ldg(fr.invSigma2 + oct);
        const float ur = CACHE ? s.urS[j] : (fr.uRight ? __ldg(fr.uRight + row) : -1.0f);
        if (ur >= 0) {
            const float er = __fsub_rn(p.xr, ur);
            const float e2 = __fadd_rn(__fadd_rn(__fmul_rn(ex, ex), __fmul_rn(ey, ey)), __fmul_rn(er, er));
            return (double)__fmul_rn(e2, is2) > 7.8;
        }
        const float e2 = __fadd_rn(__fmul_rn(ex, ex), __fmul_rn(ey, ey));
        return (double)__fmul_rn(e2, is2) > 5.99;
    }
    if (gateRight) {
        const float ur = CACHE ? s.urS[j] : __ldg(fr.uRight + row);
        if (ur > 0 && fabsf(p.xr - ur) > p.radius) return true;
    }
    return false;
}

static __device__ __forceinline__ int pj_ham(const uint4 qa, const uint4 qb, const uint4 a, const uint4 b) {
    return __popc(qa.x ^ a.x) + __popc(qa.y ^ a.y) + __popc(qa.z ^ a.z) + __popc(qa.w ^ a.w) + __popc(qb.x ^ b.x) + __popc(qb.y ^ b.y) +
           __popc(qb.z ^ b.z) + __popc(qb.w ^ b.w);
}

/* Frame::GetFeaturesInArea: visit(row, octave, position in `sorted`) for every keypoint of the window, in the reference's order (ascending
 * position in `sorted`: column by column, rows ascending inside a cell) */
template <bool CACHE, class F>
static __device__ __forceinline__ void pj_window(const PjFrame& fr, const PjSmem& s, const orbfe_proj_query& p, int colFirst, int colStep, F&& visit) {
    const float r = p.radius, ax = p.x - fr.minx, ay = p.y - fr.miny;
    const int x0 = max(0, (int)floorf((ax - r) * fr.invw));
    if (x0 >= PJ_COLS) return;
    const int x1 = min(PJ_COLS - 1, (int)ceilf((ax + r) * fr.invw));
    if (x1 < 0) return;
    const int y0 = max(0, (int)floorf((ay - r) * fr.invh));
    if (y0 >= PJ_ROWS) return;
    const int y1 = min(PJ_ROWS - 1, (int)ceilf((ay + r) * fr.invh));
    if (y1 < 0) return;
    if ((p.flags & ORBFE_PQ_RIGHT) && fr.nLeft < 0) return;   /* no right camera: mGridRight is empty (and not even allocated here) */
    const int side = (p.flags & ORBFE_PQ_RIGHT) ? PJ_CELLS : 0;
    for (int ix = x0 + colFirst; ix <= x1; ix += colStep) {   /* a group of colStep lanes shares a window: lane colFirst takes every colStep-th column */
        const int c0 = side + ix * PJ_ROWS + y0, c1 = side + ix * PJ_ROWS + y1;
        /* the cells iy = y0..y1 of one column are adjacent in the sort: one run */
        const int beg = c0 ? s.cellEnd[c0 - 1] : 0, end = s.cellEnd[c1];
        for (int j = beg; j < end; j++) {
            const int row = s.sorted[j];
            if (CACHE) {
                const int oct = s.octS[j];
                if (oct < p.min_level || (p.max_level >= 0 && oct > p.max_level)) continue;
                const float2 xy = s.xyS[j];
                if (fabsf(xy.x - p.x) < r && fabsf(xy.y - p.y) < r) visit(row, oct, j);
            } else {
                const orbfe_keypoint* kp = fr.kps + row;
                const int oct = __ldg(&kp->octave);
                if (oct < p.min_level || (p.max_level >= 0 && oct > p.max_level)) continue;
                if (fabsf(__ldg(&kp->x) - p.x) < r && fabsf(__ldg(&kp->y) - p.y) < r) visit(row, oct, j);
            }
        }
    }
}

/* NT threads per frame: 512 with two CTAs per SM for batches that fill the machine, 1024 for a call with a few frames, where one CTA's
 * latency is the whole cost (orbfe_search_by_projection_device picks).  Loading the window's keypoints and descriptors in groups of four
 * ahead of their use was measured too: the registers it takes cost more than the waits it saves, in both regimes. */
template <int NT, int MINB, bool CACHE>
__global__ void __launch_bounds__(NT, MINB) k_search_by_projection(int maxKp, const orbfe_keypoint* __restrict__ kpsAll, const uint8_t* __restrict__ descAll,
                                                                const int32_t* __restrict__ kpOff, const int32_t* __restrict__ nLeftAll,
                                                                const float* __restrict__ uRightAll, const int32_t* __restrict__ partnerAll,
                                                                const uint8_t* __restrict__ takenIn, const orbfe_frame_grid* __restrict__ grids,
                                                                const orbfe_proj_query* __restrict__ queries, const uint8_t* __restrict__ qdescAll,
                                                                const int32_t* __restrict__ qOff, int thHigh, float nnratio, int checkOri,
                                                                const float* __restrict__ invSigma2, orbfe_proj_match* out, uint8_t* __restrict__ takenOut) {
    extern __shared__ __align__(16) unsigned char pj_raw[];
    const int nCells = nLeftAll ? 2 * PJ_CELLS : PJ_CELLS;   /* the second grid (mGridRight) only exists with fisheye rigs in the batch */
    PjSmem s;
    {
        const int cap = (maxKp + 3) & ~3;
        unsigned char* base = pj_raw;
        s.descS = reinterpret_cast<uint4*>(base);
        s.xyS = reinterpret_cast<float2*>(base + (CACHE ? (size_t)cap * 32 : 0));
        s.octS = reinterpret_cast<int*>(base + (CACHE ? (size_t)cap * 40 : 0));
        s.urS = reinterpret_cast<float*>(base + (CACHE ? (size_t)cap * 44 : 0));
        if (CACHE) base += (size_t)cap * PJ_CACHE_BYTES_PER_KP;
        s.cellEnd = reinterpret_cast<int*>(base);
        s.owner = s.cellEnd + nCells;
        s.taken = reinterpret_cast<uint32_t*>(s.owner + cap);
        s.sorted = reinterpret_cast<uint16_t*>(s.taken + cap);
        s.warpSum = reinterpret_cast<int*>(s.sorted + cap);
    }
    const int tid = threadIdx.x, f = blockIdx.x;
    const int k0 = kpOff[f], q0 = qOff[f], nq = qOff[f + 1] - q0;
    PjFrame fr;
    fr.n = kpOff[f + 1] - k0;
    fr.kps = kpsAll + k0;
    fr.desc = reinterpret_cast<const uint4*>(descAll) + (size_t)k0 * 2;
    fr.uRight = uRightAll ? uRightAll + k0 : nullptr;
    fr.partner = partnerAll ? partnerAll + k0 : nullptr;
    fr.nLeft = nLeftAll ? nLeftAll[f] : -1;
    const orbfe_frame_grid g = grids[f];
    fr.minx = g.min_x; fr.miny = g.min_y; fr.invw = g.inv_w; fr.invh = g.inv_h;
    fr.invSigma2 = invSigma2;
    const orbfe_proj_query* qs = queries + q0;
    const uint4* qdesc = reinterpret_cast<const uint4*>(qdescAll) + (size_t)q0 * 2;
    orbfe_proj_match* res = out + q0;
    volatile int* status = &res[0].status; /* stride 7 ints; PJ_PENDING while unresolved */
    __shared__ int s_rot[32];

    if (fr.n > maxKp || fr.n > PJ_MAXKP || fr.n < 0 || nq > PJ_MAXQ) { /* the caller's promise about the frame size does not hold */
        for (int q = tid; q < nq; q += NT) {
            orbfe_proj_match m;
            m.best_idx = -1; m.best_dist = 256; m.best_level = -1; m.second_dist = 256; m.second_level = -1; m.accepted = 0; m.status = ORBFE_PS_CAPACITY;
            res[q] = m;
        }
        return;
    }
    const bool gateRight = fr.nLeft < 0 && fr.uRight != nullptr;
    const bool stereoRig = fr.nLeft >= 0 && fr.partner != nullptr;

    /* ---- grid: AssignFeaturesToGrid as a counting sort ---- */
    for (int i = tid; i < nCells; i += NT) s.cellEnd[i] = 0;
    for (int i = tid; i < fr.n; i += NT) s.taken[i] = 0xFFFFFFFFu;
    __syncthreads();
    for (int i = tid; i < fr.n; i += NT) {
        const float x = __ldg(&fr.kps[i].x), y = __ldg(&fr.kps[i].y);
        const int px = (int)roundf((x - fr.minx) * fr.invw), py = (int)roundf((y - fr.miny) * fr.invh);
        int c = -1;
        if (px >= 0 && px < PJ_COLS && py >= 0 && py < PJ_ROWS) {
            c = ((fr.nLeft >= 0 && i >= fr.nLeft) ? PJ_CELLS : 0) + px * PJ_ROWS + py;
            atomicAdd(&s.cellEnd[c], 1);
        }
        s.owner[i] = c;
        if (takenIn && takenIn[k0 + i]) s.taken[i] = 0xFFFF0000u;
    }
    __syncthreads();
    {   /* exclusive scan of the 6144 counts: 24 consecutive cells per thread */
        const int per = nCells / NT, base = tid * per;
        int sum = 0;
        for (int i = 0; i < per; i++) sum += s.cellEnd[base + i];
        int incl = sum;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, incl, o);
            if ((tid & 31) >= o) incl += t;
        }
        if ((tid & 31) == 31) s.warpSum[tid >> 5] = incl;
        __syncthreads();
        int run = incl - sum;
        for (int w = 0; w < (tid >> 5); w++) run += s.warpSum[w];
        for (int i = 0; i < per; i++) {
            const int c = s.cellEnd[base + i];
            s.cellEnd[base + i] = run;
            run += c;
        }
    }
    __syncthreads();
    for (int i = tid; i < fr.n; i += NT) {
        const int c = s.owner[i];
        if (c >= 0) s.sorted[atomicAdd(&s.cellEnd[c], 1)] = (uint16_t)i;   /* cellEnd[c] ends up as the end of cell c */
    }
    __syncthreads();
    for (int c = tid; c < nCells; c += NT) {   /* ascending rows inside each cell (the lists are a few entries long) */
        const int beg = c ? s.cellEnd[c - 1] : 0, end = s.cellEnd[c];
        for (int i = beg + 1; i < end; i++) {
            const uint16_t v = s.sorted[i];
            int j = i - 1;
            while (j >= beg && s.sorted[j] > v) { s.sorted[j + 1] = s.sorted[j]; j--; }
            s.sorted[j + 1] = v;
        }
    }
    for (int q = tid; q < nq; q += NT) {
        orbfe_proj_match m;
        m.best_idx = -1; m.best_dist = 256; m.best_level = -1; m.second_dist = 256; m.second_level = -1; m.accepted = 0;
        m.status = (qs[q].flags & ORBFE_PQ_SKIP) ? ORBFE_PS_SKIPPED : PJ_PENDING;
        res[q] = m;
    }
    __syncthreads();
    if (CACHE) {
        const int ns = s.cellEnd[nCells - 1];
        for (int j = tid; j < ns; j += NT) {
            const int row = s.sorted[j];
            const orbfe_keypoint* kp = fr.kps + row;
            s.xyS[j] = make_float2(__ldg(&kp->x), __ldg(&kp->y));
            s.octS[j] = __ldg(&kp->octave);
            s.urS[j] = fr.uRight ? __ldg(fr.uRight + row) : -1.0f;
            s.descS[2 * j] = __ldg(fr.desc + 2 * row);
            s.descS[2 * j + 1] = __ldg(fr.desc + 2 * row + 1);
        }
        __syncthreads();
    }
#define PJ_DESC_A(row, j) (CACHE ? s.descS[2 * (j)] : __ldg(fr.desc + 2 * (row)))
#define PJ_DESC_B(row, j) (CACHE ? s.descS[2 * (j) + 1] : __ldg(fr.desc + 2 * (row) + 1))

    /* ---- rounds ----
     * A round costs (participants x window); a query behind a long chain of earlier ones over the same keypoints would pay that
     * every round while it waits.  So only the earliest PJ_FRONT unresolved queries take part: the rule of step B needs the stamps
     * of all EARLIER unresolved queries, and those are inside the front by construction. */
    __shared__ int s_front, s_walked;
    /* How many lanes share a query.  With the small windows of SearchLocalPoints (a handful of keypoints) one thread per query is the
     * efficient form.  With the wide windows of the frame-to-frame search (th = 7 or 15 times the level scale: tens to hundreds of
     * keypoints, and very unequal) a thread per query leaves most lanes of a warp waiting for the longest window, so eight lanes share
     * a query, each taking every eighth grid column, and merge their results.  Decided per frame from the mean number of keypoints in
     * the columns its queries walk. */
    if (tid == 0) s_walked = 0;
    __syncthreads();
    {
        int mine = 0;
        for (int q = tid; q < nq; q += NT) {
            const orbfe_proj_query p = qs[q];
            if (p.flags & ORBFE_PQ_SKIP) continue;
            const float r = p.radius, ax = p.x - fr.minx, ay = p.y - fr.miny;
            const int x0 = max(0, (int)floorf((ax - r) * fr.invw)), x1 = min(PJ_COLS - 1, (int)ceilf((ax + r) * fr.invw));
            const int y0 = max(0, (int)floorf((ay - r) * fr.invh)), y1 = min(PJ_ROWS - 1, (int)ceilf((ay + r) * fr.invh));
            if (x0 >= PJ_COLS || x1 < 0 || y0 >= PJ_ROWS || y1 < 0 || ((p.flags & ORBFE_PQ_RIGHT) && fr.nLeft < 0)) continue;
            const int side = (p.flags & ORBFE_PQ_RIGHT) ? PJ_CELLS : 0;
            for (int ix = x0; ix <= x1; ix++) {
                const int c0 = side + ix * PJ_ROWS + y0, c1 = side + ix * PJ_ROWS + y1;
                mine += s.cellEnd[c1] - (c0 ? s.cellEnd[c0 - 1] : 0);
            }
        }
#pragma unroll
        for (int o = 16; o; o >>= 1) mine += __shfl_xor_sync(0xffffffffu, mine, o);
        if ((tid & 31) == 0 && mine) atomicAdd(&s_walked, mine);
    }
    __syncthreads();
    const int G = (nq > 0 && s_walked > 12 * nq) ? 8 : 1;       /* lanes per query */
    const int gl = tid & (G - 1), NG = NT / G;                 /* my lane in the group, groups in the CTA */
    const unsigned gmask = G == 1 ? (1u << (tid & 31)) : (0xFFu << (tid & 24));
    const int leader = tid & 31 & ~(G - 1);
    int myFirst = tid / G;   /* my group's first query that may still be unresolved (it owns myFirst, myFirst + NG, ...) */
    int rounds = 0;
    for (;;) {
        for (int i = tid; i < fr.n; i += NT) {
            s.owner[i] = INT_MAX;
            const uint32_t w = s.taken[i];
            if ((w >> 16) != PJ_FREE16) s.taken[i] = 0xFFFF0000u | min(w & 0xFFFFu, w >> 16);
        }
        if (tid == 0) s_front = INT_MAX;
        __syncthreads();
        while (myFirst < nq && status[myFirst * 7] != PJ_PENDING) myFirst += NG;
        if (myFirst < nq) atomicMin(&s_front, myFirst);
        __syncthreads();
        const int limit = s_front > INT_MAX - PJ_FRONT ? INT_MAX : s_front + PJ_FRONT;
        /* A. stamps */
        for (int q = myFirst; q < nq && q < limit; q += NG) {
            if (status[q * 7] != PJ_PENDING) continue;
            const orbfe_proj_query p = qs[q];
            if ((p.flags & (ORBFE_PQ_AFTER_LEFT | ORBFE_PQ_NEEDS_LEFT_WINDOW)) && gl == 0)
                res[q].accepted = (q > 0 && status[(q - 1) * 7] == PJ_PENDING) ? 0 : 1;   /* may I look at q-1 in step B? */
            if (!(p.flags & ORBFE_PQ_BLOCKS)) continue;   /* a map point without observations occupies nothing: no stamps */
            /* what a query takes is its best candidate, and that has to be within thHigh: only those rows (and their partners) get its
             * stamp.  With stamps on the whole window, queries whose windows overlap resolved one per round whether or not they competed. */
            const uint4 sa = __ldg(qdesc + 2 * q), sb = __ldg(qdesc + 2 * q + 1);
            pj_window<CACHE>(fr, s, p, gl, G, [&](int row, int oct, int j) {
                if (PJ_TAKEN_FOR(s, row, q)) return;
                if (pj_gated<CACHE>(fr, s, p, row, j, oct, gateRight)) return;
                if (pj_ham(sa, sb, PJ_DESC_A(row, j), PJ_DESC_B(row, j)) > thHigh) return;
                atomicMin(&s.owner[row], q);
                if (stereoRig) {
                    const int pt = __ldg(fr.partner + row);
                    if (pt >= 0 && pt < fr.n) atomicMin(&s.owner[pt], q);
                }
            });
        }
        __syncthreads();
        /* B. scan, decide.  The reference's update rule (`d < bestDist`: the old best becomes the second; `else if d < bestDist2`) keeps
         * the two smallest candidates in the order (distance, position in the walk), so a lane keeps two keys distance << 16 | position
         * and the lanes of a group merge theirs. */
        int pending = 0;
        for (int q = myFirst; q < nq; q += NG) {
            if (q >= limit) { pending += gl == 0; break; }   /* behind the front: waits */
            if (status[q * 7] != PJ_PENDING) continue;
            const orbfe_proj_query p = qs[q];
            if (p.flags & (ORBFE_PQ_AFTER_LEFT | ORBFE_PQ_NEEDS_LEFT_WINDOW)) {
                int may = gl == 0 ? res[q].accepted : 0;
                may = __shfl_sync(gmask, may, leader);
                if (may == 0) { pending += gl == 0; continue; }
                if (gl == 0) res[q].accepted = 0;
                const int prev = q > 0 && !(qs[q - 1].flags & ORBFE_PQ_RIGHT) ? status[(q - 1) * 7] : -2;
                if (((p.flags & ORBFE_PQ_AFTER_LEFT) && prev == ORBFE_PS_RATIO_REJECTED) ||        /* `continue` at ORBmatcher.cc:117-118 */
                    ((p.flags & ORBFE_PQ_NEEDS_LEFT_WINDOW) && prev == ORBFE_PS_EMPTY_WINDOW)) {   /* `continue` at ORBmatcher.cc:2262-2263 */
                    __syncwarp(gmask);                   /* everyone of the group has read the status of q before it changes */
                    if (gl == 0) status[q * 7] = ORBFE_PS_DROPPED;
                    continue;
                }
            }
            const uint4 qa = __ldg(qdesc + 2 * q), qb = __ldg(qdesc + 2 * q + 1);
            unsigned k1 = 0xFFFFFFFFu, k2 = 0xFFFFFFFFu;
            int seen = 0;
            int foreign = 257;   /* the closest candidate that carries the stamp of an earlier unresolved query */
            pj_window<CACHE>(fr, s, p, gl, G, [&](int row, int oct, int j) {
                seen = 1;
                if (PJ_TAKEN_FOR(s, row, q)) return;
                if (pj_gated<CACHE>(fr, s, p, row, j, oct, gateRight)) return;
                const int dist = pj_ham(qa, qb, PJ_DESC_A(row, j), PJ_DESC_B(row, j));
                const unsigned key = ((unsigned)dist << 16) | (unsigned)j;
                if (key < k1) { k2 = k1; k1 = key; }
                else if (key < k2) k2 = key;
                if (s.owner[row] < q) foreign = min(foreign, dist);
            });
            for (int o = G >> 1; o; o >>= 1) {
                const unsigned o1 = __shfl_xor_sync(gmask, k1, o), o2 = __shfl_xor_sync(gmask, k2, o);
                const unsigned lo = min(k1, o1), hi = max(k1, o1);
                k2 = min(hi, min(k2, o2));
                k1 = lo;
                foreign = min(foreign, __shfl_xor_sync(gmask, foreign, o));
                seen |= __shfl_xor_sync(gmask, seen, o);
            }
            int bestDist = 256, bestLevel = -1, bestDist2 = 256, bestLevel2 = -1, bestIdx = -1;
            if (k1 != 0xFFFFFFFFu) {
                const int j1 = (int)(k1 & 0xFFFFu);
                bestDist = (int)(k1 >> 16); bestIdx = s.sorted[j1];
                bestLevel = CACHE ? s.octS[j1] : __ldg(&fr.kps[bestIdx].octave);
            }
            if (k2 != 0xFFFFFFFFu) {
                const int j2 = (int)(k2 & 0xFFFFu);
                bestDist2 = (int)(k2 >> 16);
                bestLevel2 = CACHE ? s.octS[j2] : __ldg(&fr.kps[s.sorted[j2]].octave);
            }
            int st = seen ? ORBFE_PS_NO_CANDIDATES : ORBFE_PS_EMPTY_WINDOW;
            bool blocked = false;
            int partnerRow = -1;
            if (bestIdx >= 0) {
                const bool bestOnly = (p.flags & ORBFE_PQ_BEST_ONLY) != 0;
                st = bestDist > thHigh ? ORBFE_PS_OVER_THRESHOLD
                     : (!bestOnly && bestLevel == bestLevel2 && (float)bestDist > nnratio * (float)bestDist2) ? ORBFE_PS_RATIO_REJECTED : ORBFE_PS_ACCEPTED;
                blocked = foreign <= bestDist2;   /* candidates above the second best cannot change any reported field */
                if (st == ORBFE_PS_ACCEPTED && stereoRig) {
                    const int pt = __ldg(fr.partner + bestIdx);
                    if (pt >= 0 && pt < fr.n) { partnerRow = pt; if (s.owner[pt] < q) blocked = true; }
                }
            }
            if (blocked) { pending += gl == 0; continue; }
            __syncwarp(gmask);                           /* everyone of the group has read the status of q before it changes */
            if (gl == 0) {
                orbfe_proj_match m;
                m.best_idx = bestIdx; m.best_dist = bestDist; m.best_level = bestLevel; m.second_dist = bestDist2; m.second_level = bestLevel2;
                m.accepted = st == ORBFE_PS_ACCEPTED ? 1 : 0; m.status = st;
                res[q] = m;
                if (st == ORBFE_PS_ACCEPTED && (p.flags & ORBFE_PQ_BLOCKS)) {
                    pj_take(s, bestIdx, q);
                    if (partnerRow >= 0) pj_take(s, partnerRow, q);
                }
            }
        }
        rounds++;
        if (__syncthreads_count(pending) == 0) break;
    }
    if (tid == 0) { atomicMax(&g_pjRoundsMax, rounds); atomicAdd(&g_pjRoundsSum, rounds); }
    /* ---- rotation consistency (ORBmatcher.cc:2305-2322, 2391-2412): histogram of kpLF.angle - kpCF.angle over the matches ---- */
    if (checkOri) {
        const float factor = 1.0f / 30.0f;   /* 1.0f/HISTO_LENGTH */
        if (tid < 32) s_rot[tid] = 0;
        __syncthreads();
        for (int q = tid; q < nq; q += NT) {
            if (status[q * 7] != ORBFE_PS_ACCEPTED) continue;
            float rot = qs[q].angle - __ldg(&fr.kps[res[q].best_idx].angle);
            if (rot < 0.0f) rot += 360.0f;
            int bin = (int)roundf(rot * factor);
            if (bin == 30) bin = 0;
            if (bin >= 0 && bin < 30) atomicAdd(&s_rot[bin], 1);
        }
        __syncthreads();
        if (tid == 0) {   /* ComputeThreeMaxima (ORBmatcher.cc:2545-2586) */
            int ind1 = -1, ind2 = -1, ind3 = -1, max1 = 0, max2 = 0, max3 = 0;
            for (int i = 0; i < 30; i++) {
                const int c = s_rot[i];
                if (c > max1) { max3 = max2; max2 = max1; max1 = c; ind3 = ind2; ind2 = ind1; ind1 = i; }
                else if (c > max2) { max3 = max2; max2 = c; ind3 = ind2; ind2 = i; }
                else if (c > max3) { max3 = c; ind3 = i; }
            }
            if ((float)max2 < 0.1f * (float)max1) { ind2 = -1; ind3 = -1; }
            else if ((float)max3 < 0.1f * (float)max1) { ind3 = -1; }
            s_rot[30] = ind1; s_rot[31] = (ind2 & 0xFF) | ((ind3 & 0xFF) << 8);
        }
        __syncthreads();
        const int ind1 = s_rot[30], ind2 = (int)(signed char)(s_rot[31] & 0xFF), ind3 = (int)(signed char)((s_rot[31] >> 8) & 0xFF);
        for (int q = tid; q < nq; q += NT) {
            if (status[q * 7] != ORBFE_PS_ACCEPTED) continue;
            const int row = res[q].best_idx;
            float rot = qs[q].angle - __ldg(&fr.kps[row].angle);
            if (rot < 0.0f) rot += 360.0f;
            int bin = (int)roundf(rot * factor);
            if (bin == 30) bin = 0;
            if (bin >= 0 && bin < 30 && bin != ind1 && bin != ind2 && bin != ind3) {
                res[q].accepted = 0;
                status[q * 7] = ORBFE_PS_ROTATION;
                s.taken[row] = 0xFFFFFFFFu;      /* mvpMapPoints[row] = NULL */
            }
        }
        __syncthreads();
    }
    if (takenOut)
        for (int i = tid; i < fr.n; i += NT) takenOut[k0 + i] = s.taken[i] != 0xFFFFFFFFu ? 1 : 0;
}

extern "C" {

/* diagnostics: maximum and sum over the CTAs (frames) of the rounds k_search_by_projection needed since the last call; resets both */
int orbfe_debug_projection_rounds(int* max_rounds, int* sum_rounds) {
    int z = 0, a = 0, b = 0;
    ORBFE_CUDA(cudaDeviceSynchronize());
    ORBFE_CUDA(cudaMemcpyFromSymbol(&a, g_pjRoundsMax, sizeof a));
    ORBFE_CUDA(cudaMemcpyFromSymbol(&b, g_pjRoundsSum, sizeof b));
    ORBFE_CUDA(cudaMemcpyToSymbol(g_pjRoundsMax, &z, sizeof z));
    ORBFE_CUDA(cudaMemcpyToSymbol(g_pjRoundsSum, &z, sizeof z));
    if (max_rounds) *max_rounds = a;
    if (sum_rounds) *sum_rounds = b;
    return ORBFE_OK;
}

int orbfe_search_by_projection_device(int device, int nframes, int max_keypoints, const orbfe_keypoint* d_kps, const uint8_t* d_desc,
                                      const int32_t* d_kp_offsets, const int32_t* d_n_left, const float* d_u_right, const int32_t* d_partner,
                                      const uint8_t* d_taken_in, const orbfe_frame_grid* d_grids, const orbfe_proj_query* d_queries,
                                      const uint8_t* d_qdesc, const int32_t* d_q_offsets, int th_high, float nnratio, int check_orientation,
                                      const float* d_inv_level_sigma2, orbfe_proj_match* d_out, uint8_t* d_taken_out, void* stream) {
    if (nframes < 1 || !d_kps || !d_desc || !d_kp_offsets || !d_grids || !d_queries || !d_qdesc || !d_q_offsets || !d_out) { orbfe_set_error("bad argument"); return ORBFE_E_INVALID; }
    if (max_keypoints < 0 || max_keypoints > PJ_MAXKP) { orbfe_set_error("at most %d keypoints per frame (both cameras together)", PJ_MAXKP); return ORBFE_E_CAPACITY; }
    if (((uintptr_t)d_desc | (uintptr_t)d_qdesc) & 15) { orbfe_set_error("descriptor arrays must be 16-byte aligned"); return ORBFE_E_INVALID; }
    ORBFE_CUDA(cudaSetDevice(device));
    typedef void (*PjKernel)(int, const orbfe_keypoint*, const uint8_t*, const int32_t*, const int32_t*, const float*, const int32_t*, const uint8_t*,
                             const orbfe_frame_grid*, const orbfe_proj_query*, const uint8_t*, const int32_t*, int, float, int, const float*, orbfe_proj_match*, uint8_t*);
    int sms = 0, smemMax = 0;
    ORBFE_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device));
    ORBFE_CUDA(cudaDeviceGetAttribute(&smemMax, cudaDevAttrMaxSharedMemoryPerBlockOptin, device));
    smemMax -= 1024;                                           /* the kernel's few static words, and slack */
    const bool few = nframes <= 2 * sms;                      /* not enough frames to fill the SMs with 512-thread CTAs */
    const bool rigs = d_n_left != nullptr;
    /* the keypoints' window data in shared memory when it fits: for a batch twice per SM beside each other, for a few frames once */
    const bool cache = pj_smem_bytes(max_keypoints, rigs, true) <= (size_t)(few ? smemMax : smemMax / 2 - 1024);
    const PjKernel fn = few ? (cache ? (PjKernel)k_search_by_projection<1024, 1, true> : (PjKernel)k_search_by_projection<1024, 1, false>)
                            : (cache ? (PjKernel)k_search_by_projection<512, 2, true> : (PjKernel)k_search_by_projection<512, 2, false>);
    const int nt = few ? 1024 : 512;
    static thread_local int configured = -1;
    if (configured != device) {
        ORBFE_CUDA(cudaFuncSetAttribute(k_search_by_projection<1024, 1, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smemMax));
        ORBFE_CUDA(cudaFuncSetAttribute(k_search_by_projection<512, 2, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smemMax));
        ORBFE_CUDA(cudaFuncSetAttribute(k_search_by_projection<1024, 1, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pj_smem_bytes(PJ_MAXKP, true, false)));
        ORBFE_CUDA(cudaFuncSetAttribute(k_search_by_projection<512, 2, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pj_smem_bytes(PJ_MAXKP, true, false)));
        configured = device;
    }
    fn<<<nframes, nt, pj_smem_bytes(max_keypoints, rigs, cache), (cudaStream_t)stream>>>(max_keypoints, d_kps, d_desc, d_kp_offsets, d_n_left, d_u_right, d_partner,
                                                                         d_taken_in, d_grids, d_queries, d_qdesc, d_q_offsets, th_high, nnratio,
                                                                         check_orientation ? 1 : 0, d_inv_level_sigma2, d_out, d_taken_out);
    ORBFE_CUDA(cudaGetLastError());
    return ORBFE_OK;
}

int orbfe_search_by_projection(int device, const orbfe_keypoint* kps, const uint8_t* desc, int n, int n_left, const float* u_right, const int32_t* partner,
                               uint8_t* taken, const orbfe_frame_grid* grid, const orbfe_proj_query* queries, const uint8_t* qdesc, int nq, int th_high,
                               float nnratio, int check_orientation, const float* inv_level_sigma2, orbfe_proj_match* out, int* nmatches) {
    if (!kps || !desc || !grid || !queries || !qdesc || !out || n < 0 || nq < 0) { orbfe_set_error("bad argument"); return ORBFE_E_INVALID; }
    if (n > PJ_MAXKP) { orbfe_set_error("at most %d keypoints per frame (both cameras together)", PJ_MAXKP); return ORBFE_E_CAPACITY; }
    if (nmatches) *nmatches = 0;
    if (nq == 0) return ORBFE_OK;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || device < 0 || device >= ndev) {
        orbfe_set_error("no usable CUDA device %d (found %d); this library has no CPU path", device, ndev);
        return ORBFE_E_CUDA;
    }
    for (int i = 0; i < nq; i++)
        if ((queries[i].flags & ORBFE_PQ_ERROR_GATE) && !inv_level_sigma2) { orbfe_set_error("ORBFE_PQ_ERROR_GATE needs inv_level_sigma2"); return ORBFE_E_INVALID; }
    /* the thread's staging block: [kps | desc | qdesc | queries | uRight | partner | grid | offsets | sigma] [taken: in and out] [out] */
    size_t off = 0;
    auto take = [&](size_t bytes) { const size_t o = off; off += (bytes + 255) & ~(size_t)255; return o; };
    const size_t oKps = take((size_t)n * sizeof(orbfe_keypoint)), oDesc = take((size_t)n * 32), oQd = take((size_t)nq * 32);
    const size_t oQ = take((size_t)nq * sizeof(orbfe_proj_query)), oUr = take((size_t)n * 4), oPt = take((size_t)n * 4), oGrid = take(sizeof(orbfe_frame_grid));
    const size_t oOff = take(5 * sizeof(int32_t)), oSig = take(16 * sizeof(float)), oTk = take((size_t)n);
    const size_t inEnd = off;
    const size_t oOut = take((size_t)nq * sizeof(orbfe_proj_match));
    OrbfeStaging ws;
    int rc = orbfe_thread_staging(device, off, &ws);
    if (rc != ORBFE_OK) return rc;
    uint8_t* h = ws.h;
    uint8_t* d = ws.d;
    const int32_t offs[5] = {0, n, 0, nq, n_left};
    if (n) { memcpy(h + oKps, kps, (size_t)n * sizeof(orbfe_keypoint)); memcpy(h + oDesc, desc, (size_t)n * 32); }
    memcpy(h + oQd, qdesc, (size_t)nq * 32);
    memcpy(h + oQ, queries, (size_t)nq * sizeof(orbfe_proj_query));
    if (u_right && n) memcpy(h + oUr, u_right, (size_t)n * 4);
    if (partner && n) memcpy(h + oPt, partner, (size_t)n * 4);
    if (taken && n) memcpy(h + oTk, taken, (size_t)n);
    memcpy(h + oGrid, grid, sizeof(orbfe_frame_grid));
    memcpy(h + oOff, offs, sizeof(offs));
    if (inv_level_sigma2) memcpy(h + oSig, inv_level_sigma2, 16 * sizeof(float));
    cudaError_t e = cudaMemcpyAsync(d, h, inEnd, cudaMemcpyHostToDevice, ws.stream);
    if (e == cudaSuccess)
        rc = orbfe_search_by_projection_device(device, 1, n > 0 ? n : 1, (const orbfe_keypoint*)(d + oKps), d + oDesc, (const int32_t*)(d + oOff),
                                               (const int32_t*)(d + oOff) + 4, u_right ? (const float*)(d + oUr) : nullptr,
                                               partner ? (const int32_t*)(d + oPt) : nullptr, taken ? d + oTk : nullptr, (const orbfe_frame_grid*)(d + oGrid),
                                               (const orbfe_proj_query*)(d + oQ), d + oQd, (const int32_t*)(d + oOff) + 2, th_high, nnratio, check_orientation,
                                               inv_level_sigma2 ? (const float*)(d + oSig) : nullptr, (orbfe_proj_match*)(d + oOut), taken ? d + oTk : nullptr, ws.stream);
    if (rc == ORBFE_OK && e == cudaSuccess) e = cudaMemcpyAsync(h + oTk, d + oTk, off - oTk, cudaMemcpyDeviceToHost, ws.stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(ws.stream);
    if (rc != ORBFE_OK) return rc;
    if (e != cudaSuccess) return orbfe_cuda_fail(e, "search_by_projection");
    memcpy(out, h + oOut, (size_t)nq * sizeof(orbfe_proj_match));
    if (taken && n > 0) memcpy(taken, h + oTk, (size_t)n);
    if (nmatches) {
        int c = 0;
        for (int i = 0; i < nq; i++)
            if (out[i].accepted) c += 1 + ((n_left >= 0 && partner && partner[out[i].best_idx] >= 0) ? 1 : 0);
        *nmatches = c;
    }
    return ORBFE_OK;
}

}  /* extern "C" */
