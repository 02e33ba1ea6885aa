/* orbfe_bow.cu — bag-of-words assignment: Frame::ComputeBoW,
 *   mpORBvocabulary->transform(vCurrentDesc, mBowVec, mFeatVec, 4)                    /root/reference/src/Frame.cc:724-731
 *   TemplatedVocabulary::transform(features, BowVector&, FeatureVector&, levelsup)    Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1125-1193
 *   TemplatedVocabulary::transform(feature, word_id, weight, nid, levelsup)           Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1217-1260
 * for batches of frames whose descriptors are already on the device (the extractor's own output buffers can be passed as they are).
 *
 *   k_bow_descend    one thread = one descriptor: walk the vocabulary tree from the root, at every node take the child with the
 *                    smallest Hamming distance (first child on ties, like the reference's strict `<`), remember the node passed
 *                    at level L - levelsup, stop at a leaf -> (word id, node id, leaf node).
 *   k_bow_assemble   one CTA = one frame: the std::map<WordId, WordValue> and std::map<NodeId, vector<unsigned>> of the reference
 *                    are two shared-memory bitonic sorts of (id << 32 | feature index).  A word's value is its weight added once per
 *                    feature in double precision (every addend of a word is the same number, so the order of the features does not
 *                    matter); the L1 / L2 norm is accumulated by one thread in ascending word order, exactly the reference's
 *                    loop over the map, then every entry is divided by it. */
#include <cuda_runtime.h>
#include <limits.h>
#include <math.h>
#include <stdint.h>
#include <string.h>

#include <algorithm>
#include <new>
#include <vector>

#include "orbfe_internal.h"
#include "orbfe_kb8.cuh"

#define BOW_NT 256
#define BOW_MAXF 4096

struct orbfe_vocabulary {
    int device, k, L, scoring, weighting, nnodes, nwords, maxFeatures, maxBatch;
    /* tree (device): children of node i are childList[childStart[i] .. childStart[i+1]) */
    int* childStart; int* childList; uint4* desc; double* weight; int* wordId;
    /* results of the last batch (device), frame stride = maxFeatures */
    int *featWord, *featNode, *featLeaf;
    int *bowWords, *bowCount, *fvNodes, *fvStarts, *fvFeats, *fvCount;
    cudaEvent_t evLast;   /* behind the last transform, on whichever stream ran it: the getter waits for this, not for the device */
    double* bowValues;
    uint8_t* stage; int* stageCount;   /* staging for the host entry point */
    cudaStream_t stream;
};

static __device__ __forceinline__ int bow_ham(const uint4 qa, const uint4 qb, const uint4 a, const uint4 b) {
    return __popc(qa.x ^ a.x) + __popc(qa.y ^ a.y) + __popc(qa.z ^ a.z) + __popc(qa.w ^ a.w) + __popc(qb.x ^ b.x) + __popc(qb.y ^ b.y) +
           __popc(qb.z ^ b.z) + __popc(qb.w ^ b.w);
}

__global__ void __launch_bounds__(BOW_NT) k_bow_descend(const uint4* __restrict__ features, int rowStride, const int* __restrict__ counts, int maxF,
                                                        const int* __restrict__ childStart, const int* __restrict__ childList,
                                                        const uint4* __restrict__ ndesc, const int* __restrict__ wordId, int nidLevel,
                                                        int* __restrict__ featWord, int* __restrict__ featNode, int* __restrict__ featLeaf) {
    const int f = blockIdx.y, i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= min(counts[f], maxF)) return;
    const uint4* q = features + ((size_t)f * rowStride + i) * 2;
    const uint4 qa = __ldg(q), qb = __ldg(q + 1);
    int node = 0, level = 0, nid = 0;
    for (;;) {
        const int c0 = __ldg(childStart + node), c1 = __ldg(childStart + node + 1);
        if (c0 == c1) break;                       /* isLeaf() */
        ++level;
        int best = __ldg(childList + c0);
        int bestD = bow_ham(qa, qb, __ldg(ndesc + 2 * best), __ldg(ndesc + 2 * best + 1));
        for (int c = c0 + 1; c < c1; c++) {
            const int id = __ldg(childList + c);
            const int d = bow_ham(qa, qb, __ldg(ndesc + 2 * id), __ldg(ndesc + 2 * id + 1));
            if (d < bestD) { bestD = d; best = id; }
        }
        node = best;
        if (level == nidLevel) nid = node;
    }
    const size_t o = (size_t)f * maxF + i;
    featWord[o] = __ldg(wordId + node);
    featNode[o] = nid;
    featLeaf[o] = node;
}

/* ascending bitonic sort of P (power of two) 64-bit keys in shared memory by the whole CTA */
static __device__ void bow_sort(unsigned long long* key, int P) {
    for (int k = 2; k <= P; k <<= 1)
        for (int j = k >> 1; j > 0; j >>= 1) {
            __syncthreads();
            for (int t = threadIdx.x; t < (P >> 1); t += BOW_NT) {
                const int lo = ((t & ~(j - 1)) << 1) | (t & (j - 1)), hi = lo | j;
                const unsigned long long a = key[lo], b = key[hi];
                const bool up = (lo & k) == 0;
                if ((a > b) == up) { key[lo] = b; key[hi] = a; }
            }
        }
    __syncthreads();
}

/* Runs of equal ids in the sorted keys -> entry index of every run head; returns the number of runs.  entry[j] receives the
 * index for heads and -1 otherwise. */
static __device__ int bow_runs(const unsigned long long* key, int nvalid, int P, int* entry, int* warpTmp) {
    const int per = (P + BOW_NT - 1) / BOW_NT, beg = threadIdx.x * per, end = min(beg + per, nvalid);
    int cnt = 0;
    for (int j = beg; j < end; j++) cnt += (j == 0 || (key[j] >> 32) != (key[j - 1] >> 32)) ? 1 : 0;
    int incl = cnt;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += t;
    }
    if (lane == 31) warpTmp[warp] = incl;
    __syncthreads();
    int base = incl - cnt, total = 0;
    for (int w = 0; w < BOW_NT / 32; w++) { if (w < warp) base += warpTmp[w]; total += warpTmp[w]; }
    for (int j = beg; j < end; j++) {
        const bool head = j == 0 || (key[j] >> 32) != (key[j - 1] >> 32);
        entry[j] = head ? base++ : -1;
    }
    __syncthreads();
    return total;
}

__global__ void __launch_bounds__(BOW_NT) k_bow_assemble(const int* __restrict__ counts, int maxF, const double* __restrict__ nodeWeight, int weighting,
                                                         int norm /* 0 none, 1 L1, 2 L2 */, const int* __restrict__ featWord, const int* __restrict__ featNode,
                                                         const int* __restrict__ featLeaf, int* __restrict__ bowWords, double* __restrict__ bowValues,
                                                         int* __restrict__ bowCount, int* __restrict__ fvNodes, int* __restrict__ fvStarts,
                                                         int* __restrict__ fvFeats, int* __restrict__ fvCount) {
    extern __shared__ __align__(16) unsigned char bow_raw[];
    unsigned long long* key = reinterpret_cast<unsigned long long*>(bow_raw);      /* P entries */
    __shared__ int warpTmp[BOW_NT / 32];
    __shared__ double s_norm;
    const int f = blockIdx.x, tid = threadIdx.x;
    const int n = min(counts[f], maxF);
    int P = 32;
    while (P < n) P <<= 1;
    int* entry = reinterpret_cast<int*>(key + P);                                   /* P entries */
    const size_t fo = (size_t)f * maxF;
    /* ---- BowVector ---- */
    int myValid = 0;
    for (int i = tid; i < P; i += BOW_NT) {
        unsigned long long kk = ~0ull;
        if (i < n && nodeWeight[featLeaf[fo + i]] > 0) { kk = ((unsigned long long)(unsigned)featWord[fo + i] << 32) | (unsigned)i; myValid++; }
        key[i] = kk;
    }
    /* block sum of myValid */
    int v = myValid;
#pragma unroll
    for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if ((tid & 31) == 0) warpTmp[tid >> 5] = v;
    __syncthreads();
    int nv = 0;
    for (int w = 0; w < BOW_NT / 32; w++) nv += warpTmp[w];
    __syncthreads();
    bow_sort(key, P);
    const int nb = bow_runs(key, nv, P, entry, warpTmp);
    for (int j = tid; j < nv; j += BOW_NT) {
        const int e = entry[j];
        if (e < 0) continue;
        const unsigned word = (unsigned)(key[j] >> 32);
        const double w = nodeWeight[featLeaf[fo + (unsigned)(key[j] & 0xffffffffu)]];
        double val = w;
        if (weighting <= 1)                                           /* TF_IDF, TF: addWeight once per feature of the word */
            for (int t = j + 1; t < nv && (unsigned)(key[t] >> 32) == word; t++) val += w;
        if (weighting <= 1 && norm == 0) val /= (double)nb;           /* the `!must` branch: divide by v.size() */
        bowWords[fo + e] = (int)word;
        bowValues[fo + e] = val;
    }
    __syncthreads();
    if (norm != 0) {
        if (tid == 0) {
            double s = 0.0;
            if (norm == 1) for (int e = 0; e < nb; e++) s += fabs(bowValues[fo + e]);
            else { for (int e = 0; e < nb; e++) s += bowValues[fo + e] * bowValues[fo + e]; s = sqrt(s); }
            s_norm = s;
        }
        __syncthreads();
        const double s = s_norm;
        if (s > 0.0)
            for (int e = tid; e < nb; e += BOW_NT) bowValues[fo + e] = bowValues[fo + e] / s;
    }
    if (tid == 0) bowCount[f] = nb;
    __syncthreads();
    /* ---- FeatureVector ---- */
    for (int i = tid; i < P; i += BOW_NT) {
        unsigned long long kk = ~0ull;
        if (i < n && nodeWeight[featLeaf[fo + i]] > 0) kk = ((unsigned long long)(unsigned)featNode[fo + i] << 32) | (unsigned)i;
        key[i] = kk;
    }
    bow_sort(key, P);
    const int nn = bow_runs(key, nv, P, entry, warpTmp);
    const size_t so = (size_t)f * (maxF + 1);
    for (int j = tid; j < nv; j += BOW_NT) {
        fvFeats[fo + j] = (int)(unsigned)(key[j] & 0xffffffffu);
        const int e = entry[j];
        if (e >= 0) { fvNodes[fo + e] = (int)(unsigned)(key[j] >> 32); fvStarts[so + e] = j; }
    }
    if (tid == 0) { fvStarts[so + nn] = nv; fvCount[f] = nn; }
    /* more descriptors than the object holds: DBoW2 would use them all, so the truncated vectors are marked unusable */
    if (tid == 0 && counts[f] > maxF) { bowCount[f] = -1; fvCount[f] = -1; }
}

/* ------------------------------------------------------------------------------------------------- SearchByBoW */
/* ORBmatcher::SearchByBoW(KeyFrame* pKF, Frame &F, vector<MapPoint*> &vpMapPointMatches) (/root/reference/src/ORBmatcher.cc:269-470),
 * the F.Nleft == -1 path: for every vocabulary node present in both feature vectors, every keyframe feature of the node that has
 * a map point scans the frame's features of the node that are still unmatched (best / second best with the strict `<` rule),
 * and takes the best one if bestDist1 <= TH_LOW and (float)bestDist1 < mfNNratio * (float)bestDist2; then the rotation histogram
 * (ComputeThreeMaxima) removes the matches outside the three fullest bins.
 * One CTA = one (keyframe, frame) pair; one warp = one vocabulary node at a time (the warps of the CTA draw nodes from a counter).
 * The serial "already matched" rule only ties together the queries of ONE node — a frame feature sits in exactly one node of the frame's
 * feature vector, so queries of different nodes never share a candidate — which makes the reference's loop over a node a job for a
 * warp that simply runs it in order: for each keyframe feature of the node in turn, the lanes scan the node's frame features in
 * parallel (lane l takes candidates l, l + 32, ...), a butterfly merges the lanes' (best, first index of the best, second best) — the
 * serial scan's strict `<` updates give exactly the smallest distance, its first position, and the second smallest value of the
 * multiset — and lane 0 decides and marks the taken feature before the next query starts.  When the node's run in the frame has at
 * most 32 features (the usual case: ~12) every lane keeps its candidate's descriptor in registers for the whole node.
 * (Round 2 first ran this as speculative rounds over all queries of the pair, as orbfe_project.cu does: there, windows overlap
 * arbitrarily; here the independence of the nodes gives the serial order for free, and a pair costs microseconds instead of rounds of
 * barriers: one pair of 1500 x 1500 features with 30-deep chains 0.55 -> 0.057 ms, 512 such pairs 2.0 -> 0.2 ms.)
 * 1024 threads for a call with few pairs (up to two per SM), 512 for batches (256 measured: 5 % slower). */
#define SB_NT 256

template <int SBW_NT>
__global__ void __launch_bounds__(SBW_NT) k_search_by_bow(int maxF, const int* __restrict__ pairKf, const int* __restrict__ pairF, const uint4* __restrict__ desc,
                                                         const orbfe_keypoint* __restrict__ kps, int rowStride, const int* __restrict__ fvNodes,
                                                         const int* __restrict__ fvStarts, const int* __restrict__ fvFeats, const int* __restrict__ fvCount,
                                                         const uint8_t* __restrict__ hasMp, const int* __restrict__ nLeftAll, int thLow, float nnratio,
                                                         int checkOri, int kfkf, int* __restrict__ matchF, orbfe_match* __restrict__ kfOut,
                                                         int* __restrict__ nmatches) {
    __shared__ uint8_t s_taken[BOW_MAXF];     /* per frame feature: matched already (vpMapPointMatches[realIdxF] != NULL) */
    __shared__ int s_rot[32];
    __shared__ int s_next;
    const int tid = threadIdx.x, lane = tid & 31, pr = blockIdx.x;
    const int a = pairKf[pr], b = pairF[pr];
    const int* nodesA = fvNodes + (size_t)a * maxF; const int* startsA = fvStarts + (size_t)a * (maxF + 1); const int* featsA = fvFeats + (size_t)a * maxF;
    const int* nodesB = fvNodes + (size_t)b * maxF; const int* startsB = fvStarts + (size_t)b * (maxF + 1); const int* featsB = fvFeats + (size_t)b * maxF;
    const int nnA = fvCount[a], nnB = fvCount[b];
    const int nLeftF = (nLeftAll && !kfkf) ? nLeftAll[b] : -1;   /* F.Nleft: rows >= it belong to the right camera of a fisheye rig */
    const uint4* descA = desc + (size_t)a * rowStride * 2; const uint4* descB = desc + (size_t)b * rowStride * 2;
    const uint8_t* mpA = hasMp ? hasMp + (size_t)a * rowStride : nullptr;
    int* mF = matchF + (size_t)pr * rowStride;
    orbfe_match* ko = kfOut ? kfOut + (size_t)pr * rowStride : nullptr;
    for (int i = tid; i < rowStride; i += SBW_NT) { mF[i] = -1; if (ko) { orbfe_match m; m.best1 = 256; m.best2 = 256; m.idx = -1; m.accepted = 0; ko[i] = m; } }
    /* keyframe-keyframe variant: rows of the second keyframe without a map point are never candidates (:886-892) */
    for (int i = tid; i < BOW_MAXF; i += SBW_NT) s_taken[i] = (kfkf && hasMp && i < rowStride && !hasMp[(size_t)b * rowStride + i]) ? 1 : 0;
    if (tid == 0) s_next = 0;
    __syncthreads();
    /* (best1 << 16 | position of the first best, best2) of two candidate sets -> of their union */
#define SB_MERGE(k1, b2, ok1, ob2)                                                \
    do {                                                                          \
        const unsigned lo_ = min(k1, ok1), hi_ = max(k1, ok1);                     \
        b2 = min(min(b2, ob2), (int)(hi_ >> 16));                                  \
        k1 = lo_;                                                                  \
    } while (0)
    for (;;) {
        int e = 0;
        if (lane == 0) e = atomicAdd(&s_next, 1);
        e = __shfl_sync(0xffffffffu, e, 0);
        if (e >= nnA) break;
        /* the lockstep walk of :290-420 is a merge join: the node's run in the frame's feature vector by binary search */
        const int node = nodesA[e];
        int lo = 0, hi = nnB;
        while (lo < hi) { const int mid = (lo + hi) >> 1; if (nodesB[mid] < node) lo = mid + 1; else hi = mid; }
        if (lo >= nnB || nodesB[lo] != node) continue;
        const int j0 = startsB[lo], run = startsB[lo + 1] - j0;
        const int qBeg = startsA[e], qEnd = startsA[e + 1];
        if (run <= 0) continue;
        /* my candidate when the run fits the warp: row and descriptor stay in registers for all the node's queries */
        const bool small = run <= 32;
        int myRow = -1;
        uint4 ca = make_uint4(0, 0, 0, 0), cb = ca;
        if (small && lane < run) { myRow = featsB[j0 + lane]; ca = __ldg(descB + 2 * myRow); cb = __ldg(descB + 2 * myRow + 1); }
        for (int q = qBeg; q < qEnd; q++) {
            const int rowA = featsA[q];
            if (mpA && !mpA[rowA]) continue;                               /* pMP && !pMP->isBad() */
            const uint4 qa = __ldg(descA + 2 * rowA), qb = __ldg(descA + 2 * rowA + 1);
            unsigned k1 = 0xFFFFFFFFu, k1R = 0xFFFFFFFFu;                   /* 256 << 16 | 0xFFFF would do; all ones = nothing seen */
            int b2 = 256, b2R = 256;
            if (small) {
                if (myRow >= 0 && !s_taken[myRow]) {                       /* if(vpMapPointMatches[realIdxF]) continue; */
                    const unsigned key = ((unsigned)bow_ham(qa, qb, ca, cb) << 16) | (unsigned)lane;
                    if (nLeftF < 0 || myRow < nLeftF) k1 = key; else k1R = key;
                }
            } else {
                for (int c = lane; c < run; c += 32) {
                    const int row = featsB[j0 + c];
                    if (s_taken[row]) continue;
                    const int dist = bow_ham(qa, qb, __ldg(descB + 2 * row), __ldg(descB + 2 * row + 1));
                    const unsigned key = ((unsigned)dist << 16) | (unsigned)c;
                    if (nLeftF < 0 || row < nLeftF) { if (key < k1) { b2 = min(b2, (int)(k1 >> 16)); k1 = key; } else b2 = min(b2, dist); }
                    else { if (key < k1R) { b2R = min(b2R, (int)(k1R >> 16)); k1R = key; } else b2R = min(b2R, dist); }
                }
            }
#pragma unroll
            for (int o = 16; o; o >>= 1) {
                const unsigned ok1 = __shfl_xor_sync(0xffffffffu, k1, o);
                const int ob2 = __shfl_xor_sync(0xffffffffu, b2, o);
                SB_MERGE(k1, b2, ok1, ob2);
            }
            if (nLeftF >= 0) {
#pragma unroll
                for (int o = 16; o; o >>= 1) {
                    const unsigned ok1 = __shfl_xor_sync(0xffffffffu, k1R, o);
                    const int ob2 = __shfl_xor_sync(0xffffffffu, b2R, o);
                    SB_MERGE(k1R, b2R, ok1, ob2);
                }
            }
            /* all lanes hold the same result; lane 0 writes */
            const int best1 = min((int)(k1 >> 16), 256), best2 = min(b2, 256);
            const int bestIdx = k1 == 0xFFFFFFFFu ? -1 : featsB[j0 + (int)(k1 & 0xFFFFu)];
            const int best1R = min((int)(k1R >> 16), 256);
            const int bestIdxR = k1R == 0xFFFFFFFFu ? -1 : featsB[j0 + (int)(k1R & 0xFFFFu)];
            const bool th1 = bestIdx >= 0 && (kfkf ? best1 < thLow : best1 <= thLow);
            const bool ok = th1 && (float)best1 < nnratio * (float)best2;
            /* :388-416: inside `if(bestDist1<=TH_LOW)`, the right-camera best is taken whenever it is <= TH_LOW (`|| true`) */
            const bool okR = !kfkf && nLeftF >= 0 && th1 && bestIdxR >= 0 && best1R <= thLow;
            if (lane == 0) {
                if (ko) { orbfe_match m; m.best1 = best1; m.best2 = best2; m.idx = bestIdx; m.accepted = ok ? 1 : 0; ko[rowA] = m; }
                if (ok) { mF[bestIdx] = rowA; s_taken[bestIdx] = 1; }
                if (okR) { mF[bestIdxR] = rowA; s_taken[bestIdxR] = 1; }
            }
            __syncwarp();
        }
    }
#undef SB_MERGE
    __syncthreads();
    /* rotation consistency (:368-384, 440-458) */
    if (checkOri) {
        const float factor = 1.0f / 30.0f;
        const orbfe_keypoint* kA = kps + (size_t)a * rowStride; const orbfe_keypoint* kB = kps + (size_t)b * rowStride;
        if (tid < 32) s_rot[tid] = 0;
        __syncthreads();
        for (int i = tid; i < rowStride; i += SBW_NT) {
            const int ra = mF[i];
            if (ra < 0) continue;
            float rot = __ldg(&kA[ra].angle) - __ldg(&kB[i].angle);
            if (rot < 0.0f) rot += 360.0f;
            int bin = (int)roundf(rot * factor);
            if (bin == 30) bin = 0;
            if (bin >= 0 && bin < 30) atomicAdd(&s_rot[bin], 1);
        }
        __syncthreads();
        if (tid == 0) {
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
        for (int i = tid; i < rowStride; i += SBW_NT) {
            const int ra = mF[i];
            if (ra < 0) continue;
            float rot = __ldg(&kA[ra].angle) - __ldg(&kB[i].angle);
            if (rot < 0.0f) rot += 360.0f;
            int bin = (int)roundf(rot * factor);
            if (bin == 30) bin = 0;
            if (bin >= 0 && bin < 30 && bin != ind1 && bin != ind2 && bin != ind3) {
                mF[i] = -1;                                       /* vpMapPointMatches[...] = NULL; nmatches-- */
                if (ko) ko[ra].accepted = 2;
            }
        }
        __syncthreads();
    }
    int cnt = 0;
    for (int i = tid; i < rowStride; i += SBW_NT) cnt += mF[i] >= 0 ? 1 : 0;
#pragma unroll
    for (int o = 16; o; o >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
    __shared__ int s_total;
    if (tid == 0) s_total = 0;
    __syncthreads();
    if ((tid & 31) == 0) atomicAdd(&s_total, cnt);
    __syncthreads();
    if (tid == 0 && nmatches) nmatches[pr] = s_total;
}

/* ---------------------------------------------------------------------------------------- SearchForTriangulation */
/* ORBmatcher::SearchForTriangulation(KeyFrame *pKF1, KeyFrame *pKF2, cv::Mat F12, vMatchedPairs, bOnlyStereo, bCoarse)
 * (/root/reference/src/ORBmatcher.cc:965-1206), pinhole rigs (no mpCamera2): for every vocabulary node present in both feature
 * vectors, every feature of keyframe 1 WITHOUT a map point looks, among the features of keyframe 2 in that node that have no map
 * point either, for the smallest descriptor distance <= TH_LOW whose keypoint is not too close to the epipole (both monocular:
 * |ep - kp2|^2 >= 100 * mvScaleFactors[octave2]) and lies on the epipolar line of kp1 (Pinhole::epipolarConstrain,
 * src/CameraModels/Pinhole.cpp:135-157: dsqr < 3.84 * mvLevelSigma2[octave2]) or bCoarse; `dist > bestDist` skips, so among equal
 * distances the LAST candidate wins.  This version of the reference never sets vbMatched2, so the features of keyframe 1 are
 * independent of each other: one thread = one feature, no rounds.  Then the rotation histogram as in SearchByBoW. */
/* RIG: both keyframes come from a two-camera (fisheye) rig (pKF->mpCamera2 != NULL, :990-1003, 1102-1132): rows >= NLeft belong to the right
 * camera, nothing is "stereo", there is no epipole test, and the constraint is KannalaBrandt8::epipolarConstrain = TriangulateMatches(...) >
 * 0.0001f with the cameras and the relative pose of the two sides the keypoints sit on. */
struct TriRig { Kb8Cam cam[4]; float R[4][9], t[4][3]; int nLeft1, nLeft2; };   /* cam: 1 left, 1 right, 2 left, 2 right; poses: ll, lr, rl, rr */
template <bool RIG>
__global__ void __launch_bounds__(SB_NT) k_search_for_triangulation(int maxF, const int* __restrict__ pair1, const int* __restrict__ pair2,
                                                                    const uint4* __restrict__ desc, const orbfe_keypoint* __restrict__ kps,
                                                                    const float* __restrict__ uRight, int rowStride, const int* __restrict__ fvNodes,
                                                                    const int* __restrict__ fvStarts, const int* __restrict__ fvFeats,
                                                                    const int* __restrict__ fvCount, const uint8_t* __restrict__ hasMp,
                                                                    const float* __restrict__ f12All, const float* __restrict__ epAll,
                                                                    const float* __restrict__ scaleFactors, const float* __restrict__ levelSigma2, int thLow,
                                                                    int onlyStereo, int coarse, int* __restrict__ match12,
                                                                    const TriRig* __restrict__ rigs) {
    const int tid = threadIdx.x, pr = blockIdx.y;
    const int a = pair1[pr], b = pair2[pr];
    const int* nodesA = fvNodes + (size_t)a * maxF; const int* startsA = fvStarts + (size_t)a * (maxF + 1); const int* featsA = fvFeats + (size_t)a * maxF;
    const int* nodesB = fvNodes + (size_t)b * maxF; const int* startsB = fvStarts + (size_t)b * (maxF + 1); const int* featsB = fvFeats + (size_t)b * maxF;
    const int nnA = fvCount[a], nnB = fvCount[b];
    const uint4* descA = desc + (size_t)a * rowStride * 2; const uint4* descB = desc + (size_t)b * rowStride * 2;
    const orbfe_keypoint* kA = kps + (size_t)a * rowStride; const orbfe_keypoint* kB = kps + (size_t)b * rowStride;
    const float* urA = uRight ? uRight + (size_t)a * rowStride : nullptr; const float* urB = uRight ? uRight + (size_t)b * rowStride : nullptr;
    const uint8_t* mpA = hasMp ? hasMp + (size_t)a * rowStride : nullptr; const uint8_t* mpB = hasMp ? hasMp + (size_t)b * rowStride : nullptr;
    const float* F = f12All + (size_t)pr * 9;
    const float epx = epAll[2 * pr], epy = epAll[2 * pr + 1];
    int* m12 = match12 + (size_t)pr * rowStride;                                /* set to -1 by the caller's memset */
    /* one warp walks nodes of keyframe 1 (binary search for the node in keyframe 2), its lanes take the node's features; the nodes of a pair
     * are dealt out over the gridDim.x CTAs of its row, so that a single pair still fills the machine */
    for (int e = blockIdx.x * (SB_NT / 32) + (tid >> 5); e < nnA; e += gridDim.x * (SB_NT / 32)) {
        const int node = nodesA[e];
        int lo = 0, hi = nnB;
        while (lo < hi) { const int mid = (lo + hi) >> 1; if (nodesB[mid] < node) lo = mid + 1; else hi = mid; }
        if (lo >= nnB || nodesB[lo] != node) continue;
        const int fb0 = startsB[lo], fb1 = startsB[lo + 1];
        for (int j = startsA[e] + (tid & 31); j < startsA[e + 1]; j += 32) {
            const int idx1 = featsA[j];
            if (mpA && mpA[idx1]) continue;                                   /* if(pMP1) continue; */
            const bool stereo1 = !RIG && urA && __ldg(urA + idx1) >= 0;      /* bStereo1 = !pKF1->mpCamera2 && mvuRight[idx1] >= 0 */
            if (onlyStereo && !stereo1) continue;
            const float x1 = __ldg(&kA[idx1].x), y1 = __ldg(&kA[idx1].y);
            /* epipolar line of kp1 in image 2: l = x1' F12 */
            const float la = __fadd_rn(__fadd_rn(__fmul_rn(x1, F[0]), __fmul_rn(y1, F[3])), F[6]);
            const float lb = __fadd_rn(__fadd_rn(__fmul_rn(x1, F[1]), __fmul_rn(y1, F[4])), F[7]);
            const float lc = __fadd_rn(__fadd_rn(__fmul_rn(x1, F[2]), __fmul_rn(y1, F[5])), F[8]);
            const float den = __fadd_rn(__fmul_rn(la, la), __fmul_rn(lb, lb));
            const uint4 qa = __ldg(descA + 2 * idx1), qb = __ldg(descA + 2 * idx1 + 1);
            int bestDist = thLow, bestIdx2 = -1;
            if (RIG && !coarse) {
                /* The constraint is a pure function of the pair, so the loop of :1060-1140 returns the candidate of smallest distance <= TH_LOW
                 * (the last one among equals) that passes it.  Candidates are therefore tried in that order, (distance, later first), and the
                 * first that triangulates wins: never more triangulations than the reference's order makes, and the lanes of the warp
                 * triangulate together instead of one after another. */
                const TriRig& rg = rigs[pr];
                const int r1 = idx1 >= rg.nLeft1 ? 1 : 0;
                const float s1 = __ldg(levelSigma2 + __ldg(&kA[idx1].octave));
                unsigned prevKey = 0;
                for (;;) {
                    unsigned best = 0xFFFFFFFFu;
                    for (int c = fb0; c < fb1; c++) {
                        const int idx2 = featsB[c];
                        if (mpB && mpB[idx2]) continue;
                        const int dist = bow_ham(qa, qb, __ldg(descB + 2 * idx2), __ldg(descB + 2 * idx2 + 1));
                        if (dist > thLow) continue;
                        const unsigned key = ((unsigned)dist << 16) | (unsigned)(0xFFFF - (c - fb0));
                        if (key > prevKey && key < best) best = key;
                    }
                    if (best == 0xFFFFFFFFu) break;
                    const int idx2 = featsB[fb0 + (0xFFFF - (int)(best & 0xFFFFu))];
                    const int r2 = idx2 >= rg.nLeft2 ? 1 : 0, combo = 2 * r1 + r2;   /* ll, lr, rl, rr */
                    FisheyeRig fr;
                    fr.c1 = rg.cam[r1]; fr.c2 = rg.cam[2 + r2];
#pragma unroll
                    for (int k = 0; k < 9; k++) fr.R12[k] = rg.R[combo][k];
#pragma unroll
                    for (int k = 0; k < 3; k++) fr.t12[k] = rg.t[combo][k];
                    float x3D[3];
                    if (kb8_triangulate_pair(fr, x1, y1, __ldg(&kB[idx2].x), __ldg(&kB[idx2].y), s1, __ldg(levelSigma2 + __ldg(&kB[idx2].octave)), x3D) > 0.0001f) {
                        bestIdx2 = idx2;
                        break;
                    }
                    prevKey = best;
                }
            } else
            for (int c = fb0; c < fb1; c++) {
                const int idx2 = featsB[c];
                if (mpB && mpB[idx2]) continue;                               /* if(vbMatched2[idx2] || pMP2) continue; (vbMatched2 is never set) */
                const bool stereo2 = !RIG && urB && __ldg(urB + idx2) >= 0;
                if (onlyStereo && !stereo2) continue;
                const int dist = bow_ham(qa, qb, __ldg(descB + 2 * idx2), __ldg(descB + 2 * idx2 + 1));
                if (dist > thLow || dist > bestDist) continue;
                const float x2 = __ldg(&kB[idx2].x), y2 = __ldg(&kB[idx2].y);
                const int oct2 = __ldg(&kB[idx2].octave);
                if (!RIG && !stereo1 && !stereo2) {
                    const float dex = __fsub_rn(epx, x2), dey = __fsub_rn(epy, y2);
                    if (__fadd_rn(__fmul_rn(dex, dex), __fmul_rn(dey, dey)) < __fmul_rn(100.0f, __ldg(scaleFactors + oct2))) continue;
                }
                bool ok = coarse != 0;
                if (!RIG && !ok && den != 0.0f) {
                    const float num = __fadd_rn(__fadd_rn(__fmul_rn(la, x2), __fmul_rn(lb, y2)), lc);
                    const float dsqr = __fdiv_rn(__fmul_rn(num, num), den);
                    ok = (double)dsqr < 3.84 * (double)__ldg(levelSigma2 + oct2);
                }
                if (ok) { bestIdx2 = idx2; bestDist = dist; }
            }
            if (bestIdx2 >= 0) m12[idx1] = bestIdx2;
        }
    }
}

/* The rotation histogram over a pair's matches (ORBmatcher.cc:1146-1185: only the three fullest 30-degree bins stay) and the count. */
__global__ void __launch_bounds__(SB_NT) k_triangulation_rotation(const int* __restrict__ pair1, const int* __restrict__ pair2,
                                                                  const orbfe_keypoint* __restrict__ kps, int rowStride, int checkOri,
                                                                  int* __restrict__ match12, int* __restrict__ nmatches) {
    __shared__ int s_rot[32];
    __shared__ int s_total;
    const int tid = threadIdx.x, pr = blockIdx.x;
    const orbfe_keypoint* kA = kps + (size_t)pair1[pr] * rowStride; const orbfe_keypoint* kB = kps + (size_t)pair2[pr] * rowStride;
    int* m12 = match12 + (size_t)pr * rowStride;
    if (tid < 32) s_rot[tid] = 0;
    if (tid == 0) s_total = 0;
    __syncthreads();
    if (checkOri) {
        for (int i = tid; i < rowStride; i += SB_NT) {
            const int r2 = m12[i];
            if (r2 < 0) continue;
            float rot = __ldg(&kA[i].angle) - __ldg(&kB[r2].angle);
            if (rot < 0.0f) rot += 360.0f;
            int bin = (int)roundf(rot * (1.0f / 30.0f));
            if (bin == 30) bin = 0;
            if (bin >= 0 && bin < 30) atomicAdd(&s_rot[bin], 1);
        }
        __syncthreads();
        if (tid == 0) {
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
        for (int i = tid; i < rowStride; i += SB_NT) {
            const int r2 = m12[i];
            if (r2 < 0) continue;
            float rot = __ldg(&kA[i].angle) - __ldg(&kB[r2].angle);
            if (rot < 0.0f) rot += 360.0f;
            int bin = (int)roundf(rot * (1.0f / 30.0f));
            if (bin == 30) bin = 0;
            if (bin >= 0 && bin < 30 && bin != ind1 && bin != ind2 && bin != ind3) m12[i] = -1;
        }
        __syncthreads();
    }
    int cnt = 0;
    for (int i = tid; i < rowStride; i += SB_NT) cnt += m12[i] >= 0 ? 1 : 0;
#pragma unroll
    for (int o = 16; o; o >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
    if ((tid & 31) == 0) atomicAdd(&s_total, cnt);
    __syncthreads();
    if (tid == 0 && nmatches) nmatches[pr] = s_total;
}

/* a host feature vector must be well formed before it goes to a kernel that indexes with it: node ids strictly ascending (the kernels find a
 * node by binary search), offsets non-decreasing from 0, at most n rows in all, every row inside the frame */
static bool fv_well_formed(const int32_t* nodes, const int32_t* starts, const int32_t* feats, int nfv, int n) {
    if (nfv == 0) return true;
    if (starts[0] != 0) return false;
    for (int e = 0; e < nfv; e++) {
        if (starts[e + 1] < starts[e] || starts[e + 1] > n) return false;
        if (e && nodes[e] <= nodes[e - 1]) return false;
    }
    for (int j = 0; j < starts[nfv]; j++)
        if (feats[j] < 0 || feats[j] >= n) return false;
    return true;
}

static int bow_norm_of(int scoring) { return scoring == 5 ? 0 : (scoring == 1 ? 2 : 1); }   /* ScoringObject.h:73-90 */

extern "C" {

void orbfe_vocabulary_destroy(orbfe_vocabulary* v) {
    if (!v) return;
    cudaSetDevice(v->device);
    if (v->stream) cudaStreamSynchronize(v->stream);
    void* bufs[] = {v->childStart, v->childList, v->desc, v->weight, v->wordId, v->featWord, v->featNode, v->featLeaf, v->bowWords, v->bowCount,
                    v->fvNodes, v->fvStarts, v->fvFeats, v->fvCount, v->bowValues, v->stage, v->stageCount};
    for (void* b : bufs) cudaFree(b);
    if (v->evLast) cudaEventDestroy(v->evLast);
    if (v->stream) cudaStreamDestroy(v->stream);
    delete v;
}

int orbfe_vocabulary_create(int device, int k, int L, int scoring, int weighting, int nrows, const int32_t* parent, const uint8_t* is_leaf,
                            const uint8_t* desc, const double* weight, int max_features, int max_batch, orbfe_vocabulary** out) {
    if (!out || !parent || !is_leaf || !desc || !weight || nrows < 1 || k < 0 || k > 20 || L < 1 || L > 10 || scoring < 0 || scoring > 5 ||
        weighting < 0 || weighting > 3 || max_batch < 1) { orbfe_set_error("bad argument"); return ORBFE_E_INVALID; }   /* limits of loadFromTextFile (:1359) */
    if (max_features < 1 || max_features > BOW_MAXF) { orbfe_set_error("at most %d features per frame", BOW_MAXF); return ORBFE_E_CAPACITY; }
    *out = nullptr;
    const int nn = nrows + 1;
    std::vector<int> cnt(nn + 1, 0), start(nn + 1, 0), list(nrows), wid(nn, 0);
    for (int r = 0; r < nrows; r++) {
        if (parent[r] < 0 || parent[r] > r) { orbfe_set_error("row %d: parent %d does not precede the node", r, parent[r]); return ORBFE_E_INVALID; }
        cnt[parent[r]]++;
    }
    for (int i = 0; i < nn; i++) start[i + 1] = start[i] + cnt[i];
    std::vector<int> cur(start.begin(), start.end() - 1);
    int nwords = 0;
    for (int r = 0; r < nrows; r++) {
        list[cur[parent[r]]++] = r + 1;
        if ((is_leaf[r] != 0) != (cnt[r + 1] == 0)) { orbfe_set_error("row %d: leaf flag and child list disagree", r); return ORBFE_E_INVALID; }
        if (is_leaf[r]) wid[r + 1] = nwords++;
    }
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || device < 0 || device >= ndev) {
        orbfe_set_error("no usable CUDA device %d (found %d); this library has no CPU path", device, ndev);
        return ORBFE_E_CUDA;
    }
    ORBFE_CUDA(cudaSetDevice(device));
    orbfe_vocabulary* v = new (std::nothrow) orbfe_vocabulary();
    if (!v) return ORBFE_E_INVALID;
    memset(v, 0, sizeof(*v));
    v->device = device; v->k = k; v->L = L; v->scoring = scoring; v->weighting = weighting; v->nnodes = nn; v->nwords = nwords;
    v->maxFeatures = max_features; v->maxBatch = max_batch;
    std::vector<uint8_t> nd((size_t)nn * 32, 0);
    memcpy(nd.data() + 32, desc, (size_t)nrows * 32);
    std::vector<double> nw(nn, 0.0);
    memcpy(nw.data() + 1, weight, (size_t)nrows * sizeof(double));
    const size_t fr = (size_t)max_batch * max_features;
    cudaError_t e = cudaStreamCreateWithFlags(&v->stream, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&v->evLast, cudaEventDisableTiming);
#define BOW_ALLOC(p, bytes) if (e == cudaSuccess) e = cudaMalloc((void**)&(p), (bytes))
    BOW_ALLOC(v->childStart, (size_t)(nn + 1) * 4); BOW_ALLOC(v->childList, (size_t)std::max(nrows, 1) * 4); BOW_ALLOC(v->desc, (size_t)nn * 32);
    BOW_ALLOC(v->weight, (size_t)nn * 8); BOW_ALLOC(v->wordId, (size_t)nn * 4);
    BOW_ALLOC(v->featWord, fr * 4); BOW_ALLOC(v->featNode, fr * 4); BOW_ALLOC(v->featLeaf, fr * 4);
    BOW_ALLOC(v->bowWords, fr * 4); BOW_ALLOC(v->bowValues, fr * 8); BOW_ALLOC(v->bowCount, (size_t)max_batch * 4);
    BOW_ALLOC(v->fvNodes, fr * 4); BOW_ALLOC(v->fvStarts, (size_t)max_batch * (max_features + 1) * 4); BOW_ALLOC(v->fvFeats, fr * 4);
    BOW_ALLOC(v->fvCount, (size_t)max_batch * 4); BOW_ALLOC(v->stage, (size_t)max_features * 32); BOW_ALLOC(v->stageCount, 4);
#undef BOW_ALLOC
    if (e == cudaSuccess) e = cudaMemcpy(v->childStart, start.data(), (size_t)(nn + 1) * 4, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(v->childList, list.data(), (size_t)nrows * 4, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(v->desc, nd.data(), (size_t)nn * 32, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(v->weight, nw.data(), (size_t)nn * 8, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(v->wordId, wid.data(), (size_t)nn * 4, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(k_bow_assemble, cudaFuncAttributeMaxDynamicSharedMemorySize, BOW_MAXF * 12);
    if (e != cudaSuccess) { const int rc = orbfe_cuda_fail(e, "vocabulary allocation"); orbfe_vocabulary_destroy(v); return rc; }
    *out = v;
    return ORBFE_OK;
}

int orbfe_vocabulary_words(const orbfe_vocabulary* v) { return v ? v->nwords : 0; }

int orbfe_bow_transform_batch_device(orbfe_vocabulary* v, const uint8_t* d_desc, int row_stride, const int32_t* d_counts, int nframes, int levelsup,
                                     void* stream) {
    if (!v || !d_desc || !d_counts || nframes < 1 || row_stride < 1) { orbfe_set_error("bad argument"); return ORBFE_E_INVALID; }
    if (nframes > v->maxBatch) { orbfe_set_error("batch of %d frames exceeds max_batch %d", nframes, v->maxBatch); return ORBFE_E_CAPACITY; }
    if ((uintptr_t)d_desc & 15) { orbfe_set_error("descriptor array must be 16-byte aligned"); return ORBFE_E_INVALID; }
    ORBFE_CUDA(cudaSetDevice(v->device));
    cudaStream_t st = stream ? (cudaStream_t)stream : v->stream;
    const int F = v->maxFeatures;
    k_bow_descend<<<dim3((F + BOW_NT - 1) / BOW_NT, nframes), BOW_NT, 0, st>>>((const uint4*)d_desc, row_stride, d_counts, F, v->childStart, v->childList, v->desc,
                                                                               v->wordId, v->L - levelsup, v->featWord, v->featNode, v->featLeaf);
    int P = 32;
    while (P < F) P <<= 1;
    k_bow_assemble<<<nframes, BOW_NT, (size_t)P * 12, st>>>(d_counts, F, v->weight, v->weighting, bow_norm_of(v->scoring), v->featWord, v->featNode, v->featLeaf,
                                                           v->bowWords, v->bowValues, v->bowCount, v->fvNodes, v->fvStarts, v->fvFeats, v->fvCount);
    ORBFE_CUDA(cudaGetLastError());
    ORBFE_CUDA(cudaEventRecord(v->evLast, st));
    return ORBFE_OK;
}

int orbfe_bow_result(orbfe_vocabulary* v, int frame, int nfeatures, int32_t* word_ids, int32_t* node_ids, int32_t* bow_words, double* bow_values,
                     int32_t* nbow, int32_t* fv_nodes, int32_t* fv_starts, int32_t* fv_feats, int32_t* nfv) {
    if (!v || frame < 0 || frame >= v->maxBatch || nfeatures < 0 || nfeatures > v->maxFeatures) { orbfe_set_error("bad argument"); return ORBFE_E_INVALID; }
    ORBFE_CUDA(cudaSetDevice(v->device));
    ORBFE_CUDA(cudaEventSynchronize(v->evLast));
    /* everything the frame produced in one round trip: the arrays are copied at the sizes known beforehand (nfeatures, and max_features for
     * the vectors, a few tens of KB) into the calling thread's pinned block, one synchronisation, and the parts in use are handed out */
    const size_t F = (size_t)v->maxFeatures, fo = (size_t)frame * F, so = (size_t)frame * (F + 1);
    size_t off = 0;
    auto take = [&](size_t bytes) { const size_t o = off; off += (bytes + 255) & ~(size_t)255; return o; };
    const size_t oCnt = take(8), oWid = take((size_t)nfeatures * 4), oNid = take((size_t)nfeatures * 4), oBw = take(F * 4), oBv = take(F * 8);
    const size_t oFn = take(F * 4), oFs = take((F + 1) * 4), oFf = take(F * 4);
    OrbfeStaging ws;
    const int rc = orbfe_thread_staging(v->device, off, &ws);
    if (rc != ORBFE_OK) return rc;
    uint8_t* h = ws.h;
#define BOW_D2H(dst, src, bytes) ORBFE_CUDA(cudaMemcpyAsync((dst), (src), (bytes), cudaMemcpyDeviceToHost, v->stream))
    BOW_D2H(h + oCnt, v->bowCount + frame, 4);
    BOW_D2H(h + oCnt + 4, v->fvCount + frame, 4);
    if (word_ids && nfeatures) BOW_D2H(h + oWid, v->featWord + fo, (size_t)nfeatures * 4);
    if (node_ids && nfeatures) BOW_D2H(h + oNid, v->featNode + fo, (size_t)nfeatures * 4);
    if (bow_words) BOW_D2H(h + oBw, v->bowWords + fo, F * 4);
    if (bow_values) BOW_D2H(h + oBv, v->bowValues + fo, F * 8);
    if (fv_nodes) BOW_D2H(h + oFn, v->fvNodes + fo, F * 4);
    if (fv_starts || fv_feats) BOW_D2H(h + oFs, v->fvStarts + so, (F + 1) * 4);
    if (fv_feats) BOW_D2H(h + oFf, v->fvFeats + fo, F * 4);
    ORBFE_CUDA(cudaStreamSynchronize(v->stream));
    const int nb = ((const int32_t*)(h + oCnt))[0], nn = ((const int32_t*)(h + oCnt))[1];
    if (nb < 0 || nn < 0) {
        orbfe_set_error("frame %d of the batch has more than max_features = %d descriptors: its BoW vectors were not computed", frame, v->maxFeatures);
        return ORBFE_E_CAPACITY;
    }
    if (nbow) *nbow = nb;
    if (nfv) *nfv = nn;
    if (word_ids && nfeatures) memcpy(word_ids, h + oWid, (size_t)nfeatures * 4);
    if (node_ids && nfeatures) memcpy(node_ids, h + oNid, (size_t)nfeatures * 4);
    if (bow_words && nb) memcpy(bow_words, h + oBw, (size_t)nb * 4);
    if (bow_values && nb) memcpy(bow_values, h + oBv, (size_t)nb * 8);
    if (fv_nodes && nn) memcpy(fv_nodes, h + oFn, (size_t)nn * 4);
    if (fv_starts) memcpy(fv_starts, h + oFs, (size_t)(nn + 1) * 4);
    if (fv_feats && nn) {
        const int total = ((const int32_t*)(h + oFs))[nn];
        if (total > 0) memcpy(fv_feats, h + oFf, (size_t)std::min<size_t>((size_t)total, F) * 4);
    }
#undef BOW_D2H
    return ORBFE_OK;
}

int orbfe_bow_transform(orbfe_vocabulary* v, const uint8_t* desc, int n, int levelsup, int32_t* word_ids, int32_t* node_ids, int32_t* bow_words,
                        double* bow_values, int32_t* nbow, int32_t* fv_nodes, int32_t* fv_starts, int32_t* fv_feats, int32_t* nfv) {
    if (!v || (n > 0 && !desc) || n < 0) { orbfe_set_error("bad argument"); return ORBFE_E_INVALID; }
    if (n > v->maxFeatures) { orbfe_set_error("%d features exceed max_features %d", n, v->maxFeatures); return ORBFE_E_CAPACITY; }
    ORBFE_CUDA(cudaSetDevice(v->device));
    if (n) ORBFE_CUDA(cudaMemcpyAsync(v->stage, desc, (size_t)n * 32, cudaMemcpyHostToDevice, v->stream));
    ORBFE_CUDA(cudaMemcpyAsync(v->stageCount, &n, 4, cudaMemcpyHostToDevice, v->stream));
    int rc = orbfe_bow_transform_batch_device(v, v->stage, v->maxFeatures, v->stageCount, 1, levelsup, v->stream);
    if (rc) return rc;
    ORBFE_CUDA(cudaStreamSynchronize(v->stream));
    return orbfe_bow_result(v, 0, n, word_ids, node_ids, bow_words, bow_values, nbow, fv_nodes, fv_starts, fv_feats, nfv);
}

int orbfe_bow_device_views(orbfe_vocabulary* v, const int32_t** d_feat_word, const int32_t** d_feat_node, const int32_t** d_bow_words,
                           const double** d_bow_values, const int32_t** d_bow_count, const int32_t** d_fv_nodes, const int32_t** d_fv_starts,
                           const int32_t** d_fv_feats, const int32_t** d_fv_count) {
    if (!v) return ORBFE_E_INVALID;
    if (d_feat_word) *d_feat_word = v->featWord;
    if (d_feat_node) *d_feat_node = v->featNode;
    if (d_bow_words) *d_bow_words = v->bowWords;
    if (d_bow_values) *d_bow_values = v->bowValues;
    if (d_bow_count) *d_bow_count = v->bowCount;
    if (d_fv_nodes) *d_fv_nodes = v->fvNodes;
    if (d_fv_starts) *d_fv_starts = v->fvStarts;
    if (d_fv_feats) *d_fv_feats = v->fvFeats;
    if (d_fv_count) *d_fv_count = v->fvCount;
    return ORBFE_OK;
}

int orbfe_search_by_bow_device(int device, int npairs, int max_features, const int32_t* d_pair_kf, const int32_t* d_pair_f, const uint8_t* d_desc,
                               const orbfe_keypoint* d_kps, int row_stride, const int32_t* d_fv_nodes, const int32_t* d_fv_starts, const int32_t* d_fv_feats,
                               const int32_t* d_fv_count, const uint8_t* d_has_map_point, const int32_t* d_n_left, int th_low, float nnratio,
                               int check_orientation, int kf_kf, int32_t* d_match_f, orbfe_match* d_kf_out, int32_t* d_nmatches, void* stream) {
    if (npairs < 1 || !d_pair_kf || !d_pair_f || !d_desc || !d_fv_nodes || !d_fv_starts || !d_fv_feats || !d_fv_count || !d_match_f || row_stride < 1 ||
        (check_orientation && !d_kps)) { orbfe_set_error("bad argument"); return ORBFE_E_INVALID; }
    if (max_features < 1 || max_features > BOW_MAXF || row_stride > BOW_MAXF) { orbfe_set_error("at most %d features per frame", BOW_MAXF); return ORBFE_E_CAPACITY; }
    if ((uintptr_t)d_desc & 15) { orbfe_set_error("descriptor array must be 16-byte aligned"); return ORBFE_E_INVALID; }
    ORBFE_CUDA(cudaSetDevice(device));
    int sms = 0;
    ORBFE_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device));
    const bool few = npairs <= 2 * sms;
    (few ? k_search_by_bow<1024> : k_search_by_bow<512>)<<<npairs, few ? 1024 : 512, 0, (cudaStream_t)stream>>>(max_features, d_pair_kf, d_pair_f, (const uint4*)d_desc, d_kps, row_stride, d_fv_nodes, d_fv_starts,
                                                                d_fv_feats, d_fv_count, d_has_map_point, d_n_left, th_low, nnratio, check_orientation ? 1 : 0, kf_kf ? 1 : 0, d_match_f,
                                                                d_kf_out, d_nmatches);
    ORBFE_CUDA(cudaGetLastError());
    return ORBFE_OK;
}

/* one (keyframe, frame) pair over host memory: the feature vectors are the fv_* arrays orbfe_bow_result returns */
int orbfe_search_by_bow(int device, const uint8_t* desc_kf, const orbfe_keypoint* kps_kf, int n_kf, const int32_t* fv_nodes_kf, const int32_t* fv_starts_kf,
                        const int32_t* fv_feats_kf, int nfv_kf, const uint8_t* has_map_point, const uint8_t* desc_f, const orbfe_keypoint* kps_f, int n_f,
                        const int32_t* fv_nodes_f, const int32_t* fv_starts_f, const int32_t* fv_feats_f, int nfv_f, const uint8_t* has_map_point_f,
                        int n_left_f, int th_low, float nnratio, int check_orientation, int kf_kf, int32_t* match_f, orbfe_match* kf_out,
                        int* nmatches) {
    if (n_kf < 0 || n_f < 0 || nfv_kf < 0 || nfv_f < 0 || !match_f || (n_kf && !desc_kf) || (n_f && !desc_f) || (check_orientation && (!kps_kf || !kps_f)) ||
        (nfv_kf && (!fv_nodes_kf || !fv_starts_kf || !fv_feats_kf)) || (nfv_f && (!fv_nodes_f || !fv_starts_f || !fv_feats_f))) { orbfe_set_error("bad argument"); return ORBFE_E_INVALID; }
    const int S = std::max(std::max(n_kf, n_f), 1);
    if (S > BOW_MAXF) { orbfe_set_error("at most %d features per frame", BOW_MAXF); return ORBFE_E_CAPACITY; }
    if (nmatches) *nmatches = 0;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || device < 0 || device >= ndev) {
        orbfe_set_error("no usable CUDA device %d (found %d); this library has no CPU path", device, ndev);
        return ORBFE_E_CUDA;
    }
    if (nfv_kf > S || nfv_f > S || !fv_well_formed(fv_nodes_kf, fv_starts_kf, fv_feats_kf, nfv_kf, n_kf) || !fv_well_formed(fv_nodes_f, fv_starts_f, fv_feats_f, nfv_f, n_f)) {
        orbfe_set_error("malformed feature vector (node ids ascending, offsets from 0, rows inside the frame)");
        return ORBFE_E_INVALID;
    }
    /* two frames (0 = keyframe, 1 = frame) in the strided layout of the batch entry point, laid out in the thread's staging block: inputs, then
     * outputs */
    size_t off = 0;
    auto take = [&](size_t bytes) { const size_t o = off; off += (bytes + 255) & ~(size_t)255; return o; };
    const size_t oDesc = take((size_t)2 * S * 32), oKps = take((size_t)2 * S * sizeof(orbfe_keypoint)), oMp = take((size_t)2 * S), oNodes = take((size_t)2 * S * 4);
    const size_t oStarts = take((size_t)2 * (S + 1) * 4), oFeats = take((size_t)2 * S * 4), oCount = take(8), oPair = take(8), oLeft = take(8);
    const size_t inEnd = off;
    const size_t oMatch = take((size_t)S * 4), oKf = take((size_t)S * sizeof(orbfe_match)), oNm = take(4);
    OrbfeStaging ws;
    int rc = orbfe_thread_staging(device, off, &ws);
    if (rc != ORBFE_OK) return rc;
    uint8_t* h = ws.h;
    uint8_t* d = ws.d;
    memset(h, 0, inEnd);
    memset(h + oMp, 1, (size_t)2 * S);
    orbfe_keypoint* hKps = (orbfe_keypoint*)(h + oKps);
    int32_t* hNodes = (int32_t*)(h + oNodes); int32_t* hStarts = (int32_t*)(h + oStarts); int32_t* hFeats = (int32_t*)(h + oFeats);
    if (n_kf) memcpy(h + oDesc, desc_kf, (size_t)n_kf * 32);
    if (n_f) memcpy(h + oDesc + (size_t)S * 32, desc_f, (size_t)n_f * 32);
    if (kps_kf && n_kf) memcpy(hKps, kps_kf, (size_t)n_kf * sizeof(orbfe_keypoint));
    if (kps_f && n_f) memcpy(hKps + S, kps_f, (size_t)n_f * sizeof(orbfe_keypoint));
    if (has_map_point && n_kf) memcpy(h + oMp, has_map_point, (size_t)n_kf);
    if (has_map_point_f && n_f) memcpy(h + oMp + S, has_map_point_f, (size_t)n_f);
    if (nfv_kf) { memcpy(hNodes, fv_nodes_kf, (size_t)nfv_kf * 4); memcpy(hStarts, fv_starts_kf, (size_t)(nfv_kf + 1) * 4); memcpy(hFeats, fv_feats_kf, (size_t)fv_starts_kf[nfv_kf] * 4); }
    if (nfv_f) { memcpy(hNodes + S, fv_nodes_f, (size_t)nfv_f * 4); memcpy(hStarts + S + 1, fv_starts_f, (size_t)(nfv_f + 1) * 4); memcpy(hFeats + S, fv_feats_f, (size_t)fv_starts_f[nfv_f] * 4); }
    ((int32_t*)(h + oCount))[0] = nfv_kf; ((int32_t*)(h + oCount))[1] = nfv_f;
    ((int32_t*)(h + oPair))[0] = 0; ((int32_t*)(h + oPair))[1] = 1;
    ((int32_t*)(h + oLeft))[0] = -1; ((int32_t*)(h + oLeft))[1] = n_left_f < 0 ? -1 : n_left_f;
    cudaError_t e = cudaMemcpyAsync(d, h, inEnd, cudaMemcpyHostToDevice, ws.stream);
    if (e == cudaSuccess)
        rc = orbfe_search_by_bow_device(device, 1, S, (const int32_t*)(d + oPair), (const int32_t*)(d + oPair) + 1, d + oDesc, (const orbfe_keypoint*)(d + oKps), S,
                                        (const int32_t*)(d + oNodes), (const int32_t*)(d + oStarts), (const int32_t*)(d + oFeats), (const int32_t*)(d + oCount),
                                        d + oMp, (const int32_t*)(d + oLeft), th_low, nnratio, check_orientation, kf_kf, (int32_t*)(d + oMatch), (orbfe_match*)(d + oKf),
                                        (int32_t*)(d + oNm), ws.stream);
    if (rc == ORBFE_OK && e == cudaSuccess) e = cudaMemcpyAsync(h + inEnd, d + inEnd, off - inEnd, cudaMemcpyDeviceToHost, ws.stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(ws.stream);
    if (rc != ORBFE_OK) return rc;
    if (e != cudaSuccess) return orbfe_cuda_fail(e, "search_by_bow");
    if (n_f) memcpy(match_f, h + oMatch, (size_t)n_f * 4);
    if (kf_out && n_kf) memcpy(kf_out, h + oKf, (size_t)n_kf * sizeof(orbfe_match));
    if (nmatches) *nmatches = *(const int32_t*)(h + oNm);
    return ORBFE_OK;
}

int orbfe_search_for_triangulation_device(int device, int npairs, int max_features, const int32_t* d_pair_1, const int32_t* d_pair_2, const uint8_t* d_desc,
                                          const orbfe_keypoint* d_kps, const float* d_u_right, int row_stride, const int32_t* d_fv_nodes,
                                          const int32_t* d_fv_starts, const int32_t* d_fv_feats, const int32_t* d_fv_count, const uint8_t* d_has_map_point,
                                          const float* d_f12, const float* d_epipole, const float* d_scale_factors, const float* d_level_sigma2, int th_low,
                                          int only_stereo, int coarse, int check_orientation, int32_t* d_match_12, int32_t* d_nmatches, void* stream,
                                          const orbfe_triangulation_rig* d_rigs) {
    if (npairs < 1 || !d_pair_1 || !d_pair_2 || !d_desc || !d_kps || !d_fv_nodes || !d_fv_starts || !d_fv_feats || !d_fv_count || !d_f12 || !d_epipole ||
        !d_scale_factors || !d_level_sigma2 || !d_match_12 || row_stride < 1) { orbfe_set_error("bad argument"); return ORBFE_E_INVALID; }
    if (max_features < 1 || max_features > BOW_MAXF || row_stride > BOW_MAXF) { orbfe_set_error("at most %d features per frame", BOW_MAXF); return ORBFE_E_CAPACITY; }
    if ((uintptr_t)d_desc & 15) { orbfe_set_error("descriptor array must be 16-byte aligned"); return ORBFE_E_INVALID; }
    if (npairs > 65535) { orbfe_set_error("at most 65535 keyframe pairs per call"); return ORBFE_E_CAPACITY; }
    ORBFE_CUDA(cudaSetDevice(device));
    static_assert(sizeof(TriRig) == sizeof(orbfe_triangulation_rig), "rig layout");
    /* matching over (split, pair) CTAs — a pair's nodes are dealt out over `split` CTAs, enough of them to fill the SMs — then the rotation
     * check, one CTA per pair */
    int sms = 0;
    ORBFE_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device));
    const int split = std::min(std::max((5 * sms + npairs - 1) / npairs, 1), 16);   /* five 256-thread CTAs fit an SM (44 registers) */
    const dim3 grid(split, npairs);
    ORBFE_CUDA(cudaMemsetAsync(d_match_12, 0xFF, (size_t)npairs * row_stride * sizeof(int32_t), (cudaStream_t)stream));
    if (d_rigs)
        k_search_for_triangulation<true><<<grid, SB_NT, 0, (cudaStream_t)stream>>>(max_features, d_pair_1, d_pair_2, (const uint4*)d_desc, d_kps, d_u_right, row_stride,
                                                                                   d_fv_nodes, d_fv_starts, d_fv_feats, d_fv_count, d_has_map_point, d_f12, d_epipole,
                                                                                   d_scale_factors, d_level_sigma2, th_low, only_stereo ? 1 : 0, coarse ? 1 : 0,
                                                                                   d_match_12, (const TriRig*)d_rigs);
    else
        k_search_for_triangulation<false><<<grid, SB_NT, 0, (cudaStream_t)stream>>>(max_features, d_pair_1, d_pair_2, (const uint4*)d_desc, d_kps, d_u_right, row_stride,
                                                                                    d_fv_nodes, d_fv_starts, d_fv_feats, d_fv_count, d_has_map_point, d_f12, d_epipole,
                                                                                    d_scale_factors, d_level_sigma2, th_low, only_stereo ? 1 : 0, coarse ? 1 : 0,
                                                                                    d_match_12, nullptr);
    k_triangulation_rotation<<<npairs, SB_NT, 0, (cudaStream_t)stream>>>(d_pair_1, d_pair_2, d_kps, row_stride, check_orientation ? 1 : 0, d_match_12, d_nmatches);
    ORBFE_CUDA(cudaGetLastError());
    return ORBFE_OK;
}

int orbfe_search_for_triangulation(int device, const uint8_t* desc_1, const orbfe_keypoint* kps_1, const float* u_right_1, const uint8_t* has_map_point_1, int n_1,
                                   const int32_t* fv_nodes_1, const int32_t* fv_starts_1, const int32_t* fv_feats_1, int nfv_1, const uint8_t* desc_2,
                                   const orbfe_keypoint* kps_2, const float* u_right_2, const uint8_t* has_map_point_2, int n_2, const int32_t* fv_nodes_2,
                                   const int32_t* fv_starts_2, const int32_t* fv_feats_2, int nfv_2, const float* f12, const float* epipole,
                                   const float* scale_factors, const float* level_sigma2, int nlevels, int th_low, int only_stereo, int coarse,
                                   int check_orientation, int32_t* match_12, int* nmatches, const orbfe_triangulation_rig* rig) {
    if (n_1 < 0 || n_2 < 0 || nfv_1 < 0 || nfv_2 < 0 || !match_12 || !f12 || !epipole || !scale_factors || !level_sigma2 || nlevels < 1 || nlevels > 16 ||
        (n_1 && (!desc_1 || !kps_1)) || (n_2 && (!desc_2 || !kps_2)) || (nfv_1 && (!fv_nodes_1 || !fv_starts_1 || !fv_feats_1)) ||
        (nfv_2 && (!fv_nodes_2 || !fv_starts_2 || !fv_feats_2))) { orbfe_set_error("bad argument"); return ORBFE_E_INVALID; }
    const int S = std::max(std::max(n_1, n_2), 1);
    if (S > BOW_MAXF) { orbfe_set_error("at most %d features per frame", BOW_MAXF); return ORBFE_E_CAPACITY; }
    if (nfv_1 > S || nfv_2 > S || !fv_well_formed(fv_nodes_1, fv_starts_1, fv_feats_1, nfv_1, n_1) || !fv_well_formed(fv_nodes_2, fv_starts_2, fv_feats_2, nfv_2, n_2)) {
        orbfe_set_error("malformed feature vector (node ids ascending, offsets from 0, rows inside the frame)");
        return ORBFE_E_INVALID;
    }
    if (nmatches) *nmatches = 0;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || device < 0 || device >= ndev) {
        orbfe_set_error("no usable CUDA device %d (found %d); this library has no CPU path", device, ndev);
        return ORBFE_E_CUDA;
    }
    /* two keyframes (0 and 1) in the strided layout of the batch entry point, laid out in the thread's staging block: inputs, then outputs */
    size_t off = 0;
    auto take = [&](size_t bytes) { const size_t o = off; off += (bytes + 255) & ~(size_t)255; return o; };
    const size_t oDesc = take((size_t)2 * S * 32), oKps = take((size_t)2 * S * sizeof(orbfe_keypoint)), oUr = take((size_t)2 * S * 4), oMp = take((size_t)2 * S);
    const size_t oNodes = take((size_t)2 * S * 4), oStarts = take((size_t)2 * (S + 1) * 4), oFeats = take((size_t)2 * S * 4), oPar = take(43 * 4);
    const size_t oMisc = take(16), oRig = take(sizeof(orbfe_triangulation_rig));
    const size_t inEnd = off;
    const size_t oMatch = take((size_t)S * 4), oNm = take(4);
    OrbfeStaging ws;
    int rc = orbfe_thread_staging(device, off, &ws);
    if (rc != ORBFE_OK) return rc;
    uint8_t* h = ws.h;
    uint8_t* d = ws.d;
    memset(h, 0, inEnd);
    float* hUr = (float*)(h + oUr);
    for (int i = 0; i < 2 * S; i++) hUr[i] = -1.0f;
    orbfe_keypoint* hKps = (orbfe_keypoint*)(h + oKps);
    int32_t* hNodes = (int32_t*)(h + oNodes); int32_t* hStarts = (int32_t*)(h + oStarts); int32_t* hFeats = (int32_t*)(h + oFeats);
    float* hPar = (float*)(h + oPar);
    int32_t* hMisc = (int32_t*)(h + oMisc);
    if (n_1) { memcpy(h + oDesc, desc_1, (size_t)n_1 * 32); memcpy(hKps, kps_1, (size_t)n_1 * sizeof(orbfe_keypoint)); }
    if (n_2) { memcpy(h + oDesc + (size_t)S * 32, desc_2, (size_t)n_2 * 32); memcpy(hKps + S, kps_2, (size_t)n_2 * sizeof(orbfe_keypoint)); }
    if (u_right_1 && n_1) memcpy(hUr, u_right_1, (size_t)n_1 * 4);
    if (u_right_2 && n_2) memcpy(hUr + S, u_right_2, (size_t)n_2 * 4);
    if (has_map_point_1 && n_1) memcpy(h + oMp, has_map_point_1, (size_t)n_1);
    if (has_map_point_2 && n_2) memcpy(h + oMp + S, has_map_point_2, (size_t)n_2);
    if (nfv_1) { memcpy(hNodes, fv_nodes_1, (size_t)nfv_1 * 4); memcpy(hStarts, fv_starts_1, (size_t)(nfv_1 + 1) * 4); memcpy(hFeats, fv_feats_1, (size_t)fv_starts_1[nfv_1] * 4); }
    if (nfv_2) { memcpy(hNodes + S, fv_nodes_2, (size_t)nfv_2 * 4); memcpy(hStarts + S + 1, fv_starts_2, (size_t)(nfv_2 + 1) * 4); memcpy(hFeats + S, fv_feats_2, (size_t)fv_starts_2[nfv_2] * 4); }
    memcpy(hPar, f12, 9 * 4); memcpy(hPar + 9, epipole, 2 * 4);
    memcpy(hPar + 11, scale_factors, (size_t)nlevels * 4); memcpy(hPar + 27, level_sigma2, (size_t)nlevels * 4);
    hMisc[0] = nfv_1; hMisc[1] = nfv_2; hMisc[2] = 0; hMisc[3] = 1;
    if (rig) memcpy(h + oRig, rig, sizeof(orbfe_triangulation_rig));
    cudaError_t e = cudaMemcpyAsync(d, h, inEnd, cudaMemcpyHostToDevice, ws.stream);
    const float* dp = (const float*)(d + oPar);
    if (e == cudaSuccess)
        rc = orbfe_search_for_triangulation_device(device, 1, S, (const int32_t*)(d + oMisc) + 2, (const int32_t*)(d + oMisc) + 3, d + oDesc, (const orbfe_keypoint*)(d + oKps),
                                                   (const float*)(d + oUr), S, (const int32_t*)(d + oNodes), (const int32_t*)(d + oStarts), (const int32_t*)(d + oFeats),
                                                   (const int32_t*)(d + oMisc), d + oMp, dp, dp + 9, dp + 11, dp + 27, th_low, only_stereo, coarse, check_orientation,
                                                   (int32_t*)(d + oMatch), (int32_t*)(d + oNm), ws.stream, rig ? (const orbfe_triangulation_rig*)(d + oRig) : nullptr);
    if (rc == ORBFE_OK && e == cudaSuccess) e = cudaMemcpyAsync(h + inEnd, d + inEnd, off - inEnd, cudaMemcpyDeviceToHost, ws.stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(ws.stream);
    if (rc != ORBFE_OK) return rc;
    if (e != cudaSuccess) return orbfe_cuda_fail(e, "search_for_triangulation");
    if (n_1) memcpy(match_12, h + oMatch, (size_t)n_1 * 4);
    if (nmatches) *nmatches = *(const int32_t*)(h + oNm);
    return ORBFE_OK;
}

/* ---- many pairs over host memory: the frames packed once into the strided layout of the batch entry points, in the thread's staging block ---- */
struct FramesPack {
    int S;                                            /* rows per frame in the strided layout (the largest n, at least 1) */
    size_t oDesc, oKps, oUr, oMp, oNodes, oStarts, oFeats, oCount, oLeft, end;
};

static size_t frames_max_rows(const orbfe_host_frame* frames, int nframes) {
    int S = 1;
    for (int f = 0; frames && f < nframes; f++) S = std::max(S, std::min(frames[f].n, BOW_MAXF));
    return (size_t)S;
}

/* validates the frames and the pair lists, lays the block out from offset 0 and fills its host side; mpDefault = what a NULL has_map_point means */
static int frames_pack(const orbfe_host_frame* frames, int nframes, const int32_t* pa, const int32_t* pb, int npairs, bool needKps, uint8_t mpDefault,
                       FramesPack* fp, size_t extraBytes, int device, OrbfeStaging* ws, size_t* extraOff) {
    if (!frames || nframes < 1 || !pa || !pb || npairs < 1) { orbfe_set_error("bad argument"); return ORBFE_E_INVALID; }
    int S = 1;
    for (int f = 0; f < nframes; f++) {
        const orbfe_host_frame& fr = frames[f];
        if (fr.n < 0 || fr.nfv < 0 || (fr.n && (!fr.desc || (needKps && !fr.kps))) || (fr.nfv && (!fr.fv_nodes || !fr.fv_starts || !fr.fv_feats))) {
            orbfe_set_error("frame %d: bad argument", f);
            return ORBFE_E_INVALID;
        }
        if (fr.n > BOW_MAXF) { orbfe_set_error("at most %d features per frame", BOW_MAXF); return ORBFE_E_CAPACITY; }
        if (fr.nfv > std::max(fr.n, 1) || !fv_well_formed(fr.fv_nodes, fr.fv_starts, fr.fv_feats, fr.nfv, fr.n)) {
            orbfe_set_error("frame %d: malformed feature vector (node ids ascending, offsets from 0, rows inside the frame)", f);
            return ORBFE_E_INVALID;
        }
        S = std::max(S, fr.n);
    }
    for (int p = 0; p < npairs; p++)
        if (pa[p] < 0 || pa[p] >= nframes || pb[p] < 0 || pb[p] >= nframes) { orbfe_set_error("pair %d names a frame outside the list", p); return ORBFE_E_INVALID; }
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || device < 0 || device >= ndev) {
        orbfe_set_error("no usable CUDA device %d (found %d); this library has no CPU path", device, ndev);
        return ORBFE_E_CUDA;
    }
    size_t off = 0;
    auto take = [&](size_t bytes) { const size_t o = off; off += (bytes + 255) & ~(size_t)255; return o; };
    const size_t N = (size_t)nframes;
    fp->S = S;
    fp->oDesc = take(N * S * 32); fp->oKps = take(N * S * sizeof(orbfe_keypoint)); fp->oUr = take(N * S * 4); fp->oMp = take(N * S);
    fp->oNodes = take(N * S * 4); fp->oStarts = take(N * (S + 1) * 4); fp->oFeats = take(N * S * 4); fp->oCount = take(N * 4); fp->oLeft = take(N * 4);
    fp->end = off;
    *extraOff = off;
    if (off + extraBytes > ((size_t)1 << 30)) { orbfe_set_error("the frames of one call must fit 1 GB"); return ORBFE_E_CAPACITY; }
    const int rc = orbfe_thread_staging(device, off + extraBytes, ws);
    if (rc != ORBFE_OK) return rc;
    uint8_t* h = ws->h;
    for (int f = 0; f < nframes; f++) {
        const orbfe_host_frame& fr = frames[f];
        const size_t r0 = (size_t)f * S;
        if (fr.n) {
            memcpy(h + fp->oDesc + r0 * 32, fr.desc, (size_t)fr.n * 32);
            if (fr.kps) memcpy(h + fp->oKps + r0 * sizeof(orbfe_keypoint), fr.kps, (size_t)fr.n * sizeof(orbfe_keypoint));
            else memset(h + fp->oKps + r0 * sizeof(orbfe_keypoint), 0, (size_t)fr.n * sizeof(orbfe_keypoint));
            float* ur = (float*)(h + fp->oUr) + r0;
            if (fr.u_right) memcpy(ur, fr.u_right, (size_t)fr.n * 4);
            else for (int i = 0; i < fr.n; i++) ur[i] = -1.0f;
            if (fr.has_map_point) memcpy(h + fp->oMp + r0, fr.has_map_point, (size_t)fr.n);
            else memset(h + fp->oMp + r0, mpDefault, (size_t)fr.n);
        }
        if (fr.nfv) {
            memcpy((int32_t*)(h + fp->oNodes) + r0, fr.fv_nodes, (size_t)fr.nfv * 4);
            memcpy((int32_t*)(h + fp->oStarts) + (size_t)f * (S + 1), fr.fv_starts, (size_t)(fr.nfv + 1) * 4);
            memcpy((int32_t*)(h + fp->oFeats) + r0, fr.fv_feats, (size_t)fr.fv_starts[fr.nfv] * 4);
        }
        ((int32_t*)(h + fp->oCount))[f] = fr.nfv;
        ((int32_t*)(h + fp->oLeft))[f] = fr.n_left < 0 ? -1 : fr.n_left;
    }
    return ORBFE_OK;
}

int orbfe_search_for_triangulation_frames(int device, const orbfe_host_frame* frames, int nframes, const int32_t* pair_1, const int32_t* pair_2, int npairs,
                                          const float* f12, const float* epipole, const float* scale_factors, const float* level_sigma2, int nlevels, int th_low,
                                          int only_stereo, int coarse, int check_orientation, int32_t* match_12, int match_stride, int32_t* nmatches,
                                          const orbfe_triangulation_rig* rigs) {
    if (!f12 || !epipole || !scale_factors || !level_sigma2 || nlevels < 1 || nlevels > 16 || !match_12 || match_stride < 1) { orbfe_set_error("bad argument"); return ORBFE_E_INVALID; }
    FramesPack fp;
    OrbfeStaging ws;
    size_t x = 0;
    /* behind the frames: pair lists, F12, epipoles, the two level tables, the rigs | outputs: matches, counts */
    const size_t P = (size_t)(npairs > 0 ? npairs : 0);
    auto up = [](size_t b) { return (b + 255) & ~(size_t)255; };
    const size_t bPairs = up(P * 8), bF = up(P * 36), bE = up(P * 8), bLv = up(32 * 4), bRig = up(P * sizeof(orbfe_triangulation_rig));
    int rc = frames_pack(frames, nframes, pair_1, pair_2, npairs, true, 0, &fp, bPairs + bF + bE + bLv + bRig + up(P * frames_max_rows(frames, nframes) * 4) + up(P * 4), device, &ws, &x);
    if (rc != ORBFE_OK) return rc;
    for (int p = 0; p < npairs; p++)
        if (frames[pair_1[p]].n > match_stride) { orbfe_set_error("pair %d: match_stride %d below the keyframe's %d rows", p, match_stride, frames[pair_1[p]].n); return ORBFE_E_INVALID; }
    const int S = fp.S;
    const size_t oPairs = x, oF = oPairs + bPairs, oE = oF + bF, oLv = oE + bE, oRig = oLv + bLv, inEnd = oRig + bRig;
    const size_t oMatch = inEnd, oNm = oMatch + up(P * S * 4), total = oNm + up(P * 4);
    uint8_t* h = ws.h;
    uint8_t* d = ws.d;
    memcpy(h + oPairs, pair_1, P * 4); memcpy(h + oPairs + P * 4, pair_2, P * 4);
    memcpy(h + oF, f12, P * 36); memcpy(h + oE, epipole, P * 8);
    memset(h + oLv, 0, 32 * 4);
    memcpy(h + oLv, scale_factors, (size_t)nlevels * 4); memcpy(h + oLv + 64, level_sigma2, (size_t)nlevels * 4);
    if (rigs) memcpy(h + oRig, rigs, P * sizeof(orbfe_triangulation_rig));
    cudaError_t e = cudaMemcpyAsync(d, h, inEnd, cudaMemcpyHostToDevice, ws.stream);
    if (e == cudaSuccess)
        rc = orbfe_search_for_triangulation_device(device, npairs, S, (const int32_t*)(d + oPairs), (const int32_t*)(d + oPairs) + P, d + fp.oDesc,
                                                   (const orbfe_keypoint*)(d + fp.oKps), (const float*)(d + fp.oUr), S, (const int32_t*)(d + fp.oNodes),
                                                   (const int32_t*)(d + fp.oStarts), (const int32_t*)(d + fp.oFeats), (const int32_t*)(d + fp.oCount), d + fp.oMp,
                                                   (const float*)(d + oF), (const float*)(d + oE), (const float*)(d + oLv), (const float*)(d + oLv) + 16, th_low,
                                                   only_stereo, coarse, check_orientation, (int32_t*)(d + oMatch), (int32_t*)(d + oNm), ws.stream,
                                                   rigs ? (const orbfe_triangulation_rig*)(d + oRig) : nullptr);
    if (rc == ORBFE_OK && e == cudaSuccess) e = cudaMemcpyAsync(h + oMatch, d + oMatch, total - oMatch, cudaMemcpyDeviceToHost, ws.stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(ws.stream);
    if (rc != ORBFE_OK) return rc;
    if (e != cudaSuccess) return orbfe_cuda_fail(e, "search_for_triangulation_frames");
    for (int p = 0; p < npairs; p++) {
        const int n1 = frames[pair_1[p]].n;
        if (n1) memcpy(match_12 + (size_t)p * match_stride, h + oMatch + (size_t)p * S * 4, (size_t)n1 * 4);
        if (nmatches) nmatches[p] = ((const int32_t*)(h + oNm))[p];
    }
    return ORBFE_OK;
}

int orbfe_search_by_bow_frames(int device, const orbfe_host_frame* frames, int nframes, const int32_t* pair_kf, const int32_t* pair_f, int npairs, int th_low,
                               float nnratio, int check_orientation, int kf_kf, int32_t* match_f, int match_stride, orbfe_match* kf_out, int32_t* nmatches) {
    if (!match_f || match_stride < 1) { orbfe_set_error("bad argument"); return ORBFE_E_INVALID; }
    FramesPack fp;
    OrbfeStaging ws;
    size_t x = 0;
    const size_t P = (size_t)(npairs > 0 ? npairs : 0);
    auto up = [](size_t b) { return (b + 255) & ~(size_t)255; };
    const size_t bPairs = up(P * 8);
    int rc = frames_pack(frames, nframes, pair_kf, pair_f, npairs, check_orientation != 0, 1, &fp,
                         bPairs + up(P * frames_max_rows(frames, nframes) * 4) + up(P * frames_max_rows(frames, nframes) * sizeof(orbfe_match)) + up(P * 4), device, &ws, &x);
    if (rc != ORBFE_OK) return rc;
    for (int p = 0; p < npairs; p++)
        if (frames[pair_f[p]].n > match_stride || (kf_out && frames[pair_kf[p]].n > match_stride)) {
            orbfe_set_error("pair %d: match_stride %d below the rows of its frames", p, match_stride);
            return ORBFE_E_INVALID;
        }
    const int S = fp.S;
    const size_t oPairs = x, inEnd = oPairs + bPairs, oMatch = inEnd, oKf = oMatch + up(P * S * 4), oNm = oKf + up(P * S * sizeof(orbfe_match)), total = oNm + up(P * 4);
    uint8_t* h = ws.h;
    uint8_t* d = ws.d;
    memcpy(h + oPairs, pair_kf, P * 4); memcpy(h + oPairs + P * 4, pair_f, P * 4);
    cudaError_t e = cudaMemcpyAsync(d, h, inEnd, cudaMemcpyHostToDevice, ws.stream);
    if (e == cudaSuccess)
        rc = orbfe_search_by_bow_device(device, npairs, S, (const int32_t*)(d + oPairs), (const int32_t*)(d + oPairs) + P, d + fp.oDesc, (const orbfe_keypoint*)(d + fp.oKps), S,
                                        (const int32_t*)(d + fp.oNodes), (const int32_t*)(d + fp.oStarts), (const int32_t*)(d + fp.oFeats), (const int32_t*)(d + fp.oCount),
                                        d + fp.oMp, (const int32_t*)(d + fp.oLeft), th_low, nnratio, check_orientation, kf_kf, (int32_t*)(d + oMatch),
                                        kf_out ? (orbfe_match*)(d + oKf) : nullptr, (int32_t*)(d + oNm), ws.stream);
    if (rc == ORBFE_OK && e == cudaSuccess) e = cudaMemcpyAsync(h + oMatch, d + oMatch, total - oMatch, cudaMemcpyDeviceToHost, ws.stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(ws.stream);
    if (rc != ORBFE_OK) return rc;
    if (e != cudaSuccess) return orbfe_cuda_fail(e, "search_by_bow_frames");
    for (int p = 0; p < npairs; p++) {
        const int nf = frames[pair_f[p]].n, nk = frames[pair_kf[p]].n;
        if (nf) memcpy(match_f + (size_t)p * match_stride, h + oMatch + (size_t)p * S * 4, (size_t)nf * 4);
        if (kf_out && nk) memcpy(kf_out + (size_t)p * match_stride, h + oKf + (size_t)p * S * sizeof(orbfe_match), (size_t)nk * sizeof(orbfe_match));
        if (nmatches) nmatches[p] = ((const int32_t*)(h + oNm))[p];
    }
    return ORBFE_OK;
}

}  /* extern "C" */
