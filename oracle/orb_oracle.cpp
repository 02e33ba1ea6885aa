/* orb_oracle.cpp — CPU ORACLE (test infrastructure only; see orb_oracle.h for the rules).
 * Extractor side: restates /root/reference/src/ORBextractor.cc and the OpenCV primitives it calls.
 * Build: g++ -O3 -march=x86-64-v3 -ffp-contract=off -std=c++17 (oracle/Makefile).
 * Pinned: bit-exact against the cv2-4.13.0 pipeline of tools/make_goldens.py (tests/golden/, tests/test_oracle_golden.py). */
#include "orb_oracle.h"

#include <algorithm>
#include <cfloat>
#include <cmath>
#include <cstring>
#include <list>
#include <utility>
#include <vector>

namespace {

const int PATCH_SIZE = 31;       /* ORBextractor.cc:70 */
const int HALF_PATCH_SIZE = 15;  /* ORBextractor.cc:71 */
const int EDGE_THRESHOLD = 19;   /* ORBextractor.cc:72 */

const int8_t kPattern[1024] = {
#include "../include/orbfe_pattern31.inc"
};

/* cvRound(float/double): round half to even (SSE cvtss2si under the default rounding mode). */
inline int cv_round(float v) { return (int)lrintf(v); }
inline int cv_round(double v) { return (int)lrint(v); }

struct Image {
    int w = 0, h = 0, step = 0;
    std::vector<uint8_t> buf;
    uint8_t* data = nullptr; /* points at (0,0) of the w x h view inside buf */
    const uint8_t* row(int y) const { return data + (ptrdiff_t)y * step; }
    uint8_t* row(int y) { return data + (ptrdiff_t)y * step; }
};

/* ---------------------------------------------------------------- cv::resize INTER_LINEAR, 8UC1 --- */
struct LinCoef {
    std::vector<int> idx;
    std::vector<int> c0, c1;
};

LinCoef linear_coefs(int ssize, int dsize) {
    LinCoef t;
    t.idx.resize(dsize);
    t.c0.resize(dsize);
    t.c1.resize(dsize);
    const double inv_scale = (double)dsize / ssize;
    const double scale = 1.0 / inv_scale;
    for (int d = 0; d < dsize; d++) {
        float f = (float)((d + 0.5) * scale - 0.5);
        int s = (int)std::floor(f);
        f -= s;
        if (s < 0) { s = 0; f = 0.f; }
        if (s >= ssize - 1) { s = ssize - 1; f = 0.f; }
        t.idx[d] = s;
        t.c0[d] = cv_round((1.f - f) * 2048.f);
        t.c1[d] = cv_round(f * 2048.f);
    }
    return t;
}

void resize_linear_u8(const uint8_t* src, int sw, int sh, int sstep, uint8_t* dst, int dw, int dh, int dstep) {
    LinCoef xc = linear_coefs(sw, dw), yc = linear_coefs(sh, dh);
    std::vector<int> h0(dw), h1(dw);
    for (int y = 0; y < dh; y++) {
        const int sy0 = yc.idx[y], sy1 = std::min(sy0 + 1, sh - 1);
        const uint8_t* r0 = src + (ptrdiff_t)sy0 * sstep;
        const uint8_t* r1 = src + (ptrdiff_t)sy1 * sstep;
        for (int x = 0; x < dw; x++) {
            const int sx0 = xc.idx[x], sx1 = std::min(sx0 + 1, sw - 1);
            h0[x] = r0[sx0] * xc.c0[x] + r0[sx1] * xc.c1[x];
            h1[x] = r1[sx0] * xc.c0[x] + r1[sx1] * xc.c1[x];
        }
        const int b0 = yc.c0[y], b1 = yc.c1[y];
        uint8_t* d = dst + (ptrdiff_t)y * dstep;
        for (int x = 0; x < dw; x++)
            d[x] = (uint8_t)((((b0 * (h0[x] >> 4)) >> 16) + ((b1 * (h1[x] >> 4)) >> 16) + 2) >> 2);
    }
}

/* ------------------------------------------------------- copyMakeBorder(BORDER_REFLECT_101), 19 px --- */
inline int reflect101(int p, int n) {
    if (n == 1) return 0;
    while (p < 0 || p >= n) {
        if (p < 0) p = -p;
        else p = 2 * (n - 1) - p;
    }
    return p;
}

/* Allocates a padded buffer (w+38)x(h+38); the view is the interior. */
void alloc_padded(Image& im, int w, int h) {
    im.w = w;
    im.h = h;
    im.step = w + 2 * EDGE_THRESHOLD;
    im.buf.assign((size_t)im.step * (h + 2 * EDGE_THRESHOLD), 0);
    im.data = im.buf.data() + (size_t)EDGE_THRESHOLD * im.step + EDGE_THRESHOLD;
}

void fill_border(Image& im) {
    const int B = EDGE_THRESHOLD;
    for (int y = -B; y < im.h + B; y++) {
        const int sy = reflect101(y, im.h);
        uint8_t* d = im.data + (ptrdiff_t)y * im.step;
        const uint8_t* s = im.data + (ptrdiff_t)sy * im.step;
        for (int x = -B; x < im.w + B; x++) {
            if (y >= 0 && y < im.h && x >= 0 && x < im.w) continue;
            d[x] = s[reflect101(x, im.w)];
        }
    }
}

/* --------------------------------------------------------------- cv::GaussianBlur 7x7 sigma 2, 8U --- */
void gaussian7_u8(const uint8_t* src, int w, int h, int sstep, uint8_t* dst, int dstep) {
    static const int K[7] = {18, 34, 48, 56, 48, 34, 18};
    std::vector<int> hor((size_t)w * h);
    for (int y = 0; y < h; y++) {
        const uint8_t* s = src + (ptrdiff_t)y * sstep;
        for (int x = 0; x < w; x++) {
            int acc = 0;
            for (int k = 0; k < 7; k++) acc += K[k] * s[reflect101(x + k - 3, w)];
            hor[(size_t)y * w + x] = acc;
        }
    }
    for (int y = 0; y < h; y++) {
        uint8_t* d = dst + (ptrdiff_t)y * dstep;
        for (int x = 0; x < w; x++) {
            int acc = 0;
            for (int k = 0; k < 7; k++) acc += K[k] * hor[(size_t)reflect101(y + k - 3, h) * w + x];
            d[x] = (uint8_t)((acc + 32768) >> 16);
        }
    }
}

/* ------------------------------------------------------------------- cv::FAST(roi, t, true), 9/16 --- */
const int kRingX[16] = {0, 1, 2, 3, 3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1};
const int kRingY[16] = {3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1, 0, 1, 2, 3};

/* m = max over the 16 arcs of 9 consecutive ring pixels of max(min d, min -d), d = I(p) - I(ring). */
inline int fast_arc_strength(const uint8_t* p, int step) {
    int d[25];
    for (int k = 0; k < 16; k++) d[k] = (int)p[0] - (int)p[kRingY[k] * step + kRingX[k]];
    for (int k = 0; k < 9; k++) d[16 + k] = d[k];
    int m = -256;
    for (int k = 0; k < 16; k++) {
        int lo = d[k], hi = d[k];
        for (int j = 1; j < 9; j++) {
            lo = std::min(lo, d[k + j]);
            hi = std::max(hi, d[k + j]);
        }
        m = std::max(m, std::max(lo, -hi));
    }
    return m;
}

struct RawKp {
    int x, y, score;
};

void fast9_16(const uint8_t* roi, int w, int h, int step, int t, std::vector<RawKp>& out) {
    out.clear();
    if (w < 7 || h < 7) return;
    std::vector<int> score((size_t)w * h, 0);
    for (int y = 3; y < h - 3; y++)
        for (int x = 3; x < w - 3; x++) {
            /* cheap necessary condition first (any 9-arc holds one pixel of each opposite pair), as OpenCV's
               own early-outs do; the arc strength below decides */
            const uint8_t* p = roi + (ptrdiff_t)y * step + x;
            const int hi = p[0] + t, lo = p[0] - t;
            const int a = p[3 * step], b = p[-3 * step], c = p[3], d = p[-3];
            if (!(((a > hi || b > hi) && (c > hi || d > hi)) || ((a < lo || b < lo) && (c < lo || d < lo)))) continue;
            const int m = fast_arc_strength(p, step);
            if (m > t) score[(size_t)y * w + x] = m - 1;
        }
    for (int y = 3; y < h - 3; y++)
        for (int x = 3; x < w - 3; x++) {
            const int s = score[(size_t)y * w + x];
            if (s <= 0) continue;
            bool keep = true;
            for (int dy = -1; dy <= 1 && keep; dy++)
                for (int dx = -1; dx <= 1; dx++) {
                    if (!dx && !dy) continue;
                    if (score[(size_t)(y + dy) * w + (x + dx)] >= s) { keep = false; break; }
                }
            if (keep) out.push_back({x, y, s});
        }
}

/* --------------------------------------------------------------------------------- cv::fastAtan2 --- */
float fast_atan2(float y, float x) {
    const float s = (float)(180.0 / 3.14159265358979323846);
    const float p1 = 0.9997878412794807f * s, p3 = -0.3258083974640975f * s;
    const float p5 = 0.1555786518463281f * s, p7 = -0.04432655554792128f * s;
    const float ax = std::fabs(x), ay = std::fabs(y);
    float a, c, c2;
    if (ax >= ay) {
        c = ay / (ax + (float)DBL_EPSILON);
        c2 = c * c;
        a = (((p7 * c2 + p5) * c2 + p3) * c2 + p1) * c;
    } else {
        c = ax / (ay + (float)DBL_EPSILON);
        c2 = c * c;
        a = 90.f - (((p7 * c2 + p5) * c2 + p3) * c2 + p1) * c;
    }
    if (x < 0) a = 180.f - a;
    if (y < 0) a = 360.f - a;
    return a;
}

/* --------------------------------------------------------------------------------- the extractor --- */
struct Key {
    float x, y, response; /* what the quadtree looks at */
    float size, angle;
    int octave;
};

struct Node { /* ExtractorNode, include/ORBextractor.h:30-41 */
    std::vector<Key> keys;
    int ULx, ULy, URx, URy, BLx, BLy, BRx, BRy;
    std::list<Node>::iterator lit;
    bool noMore = false;
    long created = 0; /* creation order: stands in for the heap address in the reference's sort */
};

}  // namespace

struct orc_extractor {
    int nfeatures;
    double scaleFactor; /* the reference stores it as double (ORBextractor.h:94) */
    int nlevels, iniTh, minTh;
    std::vector<float> scale, invScale, sigma2, invSigma2;
    std::vector<int> quota, umax;
    std::vector<Image> pyr, blurred;
    std::vector<std::vector<Key>> cand;    /* vToDistributeKeys per level */
    std::vector<std::vector<Key>> levelKp; /* allKeypoints per level (after orientation) */
    long nodeCounter = 0;

    void divide(const Node& n, Node& n1, Node& n2, Node& n3, Node& n4);
    std::vector<Key> distribute(const std::vector<Key>& in, int minX, int maxX, int minY, int maxY, int N);
    void pyramid(const uint8_t* img, int rows, int cols, int step);
    bool keypoints();
};

/* ORBextractor.cc:479-535 */
void orc_extractor::divide(const Node& n, Node& n1, Node& n2, Node& n3, Node& n4) {
    const int halfX = (int)std::ceil(static_cast<float>(n.URx - n.ULx) / 2);
    const int halfY = (int)std::ceil(static_cast<float>(n.BRy - n.ULy) / 2);
    n1.ULx = n.ULx;           n1.ULy = n.ULy;
    n1.URx = n.ULx + halfX;   n1.URy = n.ULy;
    n1.BLx = n.ULx;           n1.BLy = n.ULy + halfY;
    n1.BRx = n.ULx + halfX;   n1.BRy = n.ULy + halfY;
    n2.ULx = n1.URx;          n2.ULy = n1.URy;
    n2.URx = n.URx;           n2.URy = n.URy;
    n2.BLx = n1.BRx;          n2.BLy = n1.BRy;
    n2.BRx = n.URx;           n2.BRy = n.ULy + halfY;
    n3.ULx = n1.BLx;          n3.ULy = n1.BLy;
    n3.URx = n1.BRx;          n3.URy = n1.BRy;
    n3.BLx = n.BLx;           n3.BLy = n.BLy;
    n3.BRx = n1.BRx;          n3.BRy = n.BLy;
    n4.ULx = n3.URx;          n4.ULy = n3.URy;
    n4.URx = n2.BRx;          n4.URy = n2.BRy;
    n4.BLx = n3.BRx;          n4.BLy = n3.BRy;
    n4.BRx = n.BRx;           n4.BRy = n.BRy;
    for (const Key& k : n.keys) {
        if (k.x < n1.URx) {
            if (k.y < n1.BRy) n1.keys.push_back(k);
            else n3.keys.push_back(k);
        } else if (k.y < n1.BRy) n2.keys.push_back(k);
        else n4.keys.push_back(k);
    }
    Node* ch[4] = {&n1, &n2, &n3, &n4};
    for (Node* c : ch)
        if (c->keys.size() == 1) c->noMore = true;
}

/* ORBextractor.cc:537-761 */
std::vector<Key> orc_extractor::distribute(const std::vector<Key>& in, int minX, int maxX, int minY, int maxY, int N) {
    const int nIni = (int)std::round(static_cast<float>(maxX - minX) / (maxY - minY));
    const float hX = static_cast<float>(maxX - minX) / nIni;
    std::list<Node> nodes;
    std::vector<Node*> ini(nIni);
    for (int i = 0; i < nIni; i++) {
        Node ni;
        ni.ULx = (int)(hX * static_cast<float>(i));     ni.ULy = 0;
        ni.URx = (int)(hX * static_cast<float>(i + 1)); ni.URy = 0;
        ni.BLx = ni.ULx; ni.BLy = maxY - minY;
        ni.BRx = ni.URx; ni.BRy = maxY - minY;
        ni.created = nodeCounter++;
        nodes.push_back(ni);
        ini[i] = &nodes.back();
    }
    for (const Key& k : in) ini[(size_t)(k.x / hX)]->keys.push_back(k);

    for (auto lit = nodes.begin(); lit != nodes.end();) {
        if (lit->keys.size() == 1) { lit->noMore = true; ++lit; }
        else if (lit->keys.empty()) lit = nodes.erase(lit);
        else ++lit;
    }

    typedef std::pair<int, std::pair<long, Node*>> SizeNode; /* (size, (creation order, node)) */
    std::vector<SizeNode> sizeAndNode;
    bool finish = false;

    auto adopt = [&](Node& c, int* nToExpand) {
        if (c.keys.empty()) return;
        c.created = nodeCounter++;
        nodes.push_front(c);
        if (c.keys.size() > 1) {
            if (nToExpand) ++*nToExpand;
            sizeAndNode.push_back({(int)c.keys.size(), {nodes.front().created, &nodes.front()}});
            nodes.front().lit = nodes.begin();
        }
    };

    while (!finish) {
        int prevSize = (int)nodes.size();
        int nToExpand = 0;
        sizeAndNode.clear();
        for (auto lit = nodes.begin(); lit != nodes.end();) {
            if (lit->noMore) { ++lit; continue; }
            Node n1, n2, n3, n4;
            divide(*lit, n1, n2, n3, n4);
            adopt(n1, &nToExpand);
            adopt(n2, &nToExpand);
            adopt(n3, &nToExpand);
            adopt(n4, &nToExpand);
            lit = nodes.erase(lit);
        }
        if ((int)nodes.size() >= N || (int)nodes.size() == prevSize) {
            finish = true;
        } else if (((int)nodes.size() + nToExpand * 3) > N) {
            while (!finish) {
                prevSize = (int)nodes.size();
                std::vector<SizeNode> prev = sizeAndNode;
                sizeAndNode.clear();
                std::sort(prev.begin(), prev.end()); /* (size, creation order) ascending; walked from the back */
                for (int j = (int)prev.size() - 1; j >= 0; j--) {
                    Node* p = prev[j].second.second;
                    Node n1, n2, n3, n4;
                    divide(*p, n1, n2, n3, n4);
                    adopt(n1, nullptr);
                    adopt(n2, nullptr);
                    adopt(n3, nullptr);
                    adopt(n4, nullptr);
                    nodes.erase(p->lit);
                    if ((int)nodes.size() >= N) break;
                }
                if ((int)nodes.size() >= N || (int)nodes.size() == prevSize) finish = true;
            }
        }
    }

    std::vector<Key> result;
    result.reserve(nodes.size());
    for (const Node& n : nodes) {
        const Key* best = &n.keys[0];
        float maxResponse = best->response;
        for (size_t k = 1; k < n.keys.size(); k++)
            if (n.keys[k].response > maxResponse) {
                best = &n.keys[k];
                maxResponse = n.keys[k].response;
            }
        result.push_back(*best);
    }
    return result;
}

/* ORBextractor.cc:1152-1177 */
void orc_extractor::pyramid(const uint8_t* img, int rows, int cols, int step) {
    for (int level = 0; level < nlevels; level++) {
        const float s = invScale[level];
        const int w = cv_round((float)cols * s), h = cv_round((float)rows * s);
        alloc_padded(pyr[level], w, h);
        if (level != 0) {
            const Image& prev = pyr[level - 1];
            resize_linear_u8(prev.data, prev.w, prev.h, prev.step, pyr[level].data, w, h, pyr[level].step);
        } else {
            for (int y = 0; y < rows; y++) memcpy(pyr[0].row(y), img + (ptrdiff_t)y * step, cols);
        }
        fill_border(pyr[level]);
    }
}

/* IC_Angle, ORBextractor.cc:75-102 */
static float ic_angle(const Image& im, float px, float py, const std::vector<int>& umax) {
    int m01 = 0, m10 = 0;
    const uint8_t* c = im.row(cv_round(py)) + cv_round(px);
    for (int u = -HALF_PATCH_SIZE; u <= HALF_PATCH_SIZE; ++u) m10 += u * c[u];
    const int step = im.step;
    for (int v = 1; v <= HALF_PATCH_SIZE; ++v) {
        int vsum = 0;
        const int d = umax[v];
        for (int u = -d; u <= d; ++u) {
            const int vp = c[u + v * step], vm = c[u - v * step];
            vsum += (vp - vm);
            m10 += u * (vp + vm);
        }
        m01 += v * vsum;
    }
    return fast_atan2((float)m01, (float)m10);
}

/* ComputeKeyPointsOctTree, ORBextractor.cc:763-878.  Returns false if a level is too small for a cell grid. */
bool orc_extractor::keypoints() {
    const float W = 35;
    for (int level = 0; level < nlevels; ++level) {
        const Image& im = pyr[level];
        const int minBX = EDGE_THRESHOLD - 3, minBY = minBX;
        const int maxBX = im.w - EDGE_THRESHOLD + 3, maxBY = im.h - EDGE_THRESHOLD + 3;
        std::vector<Key>& toDist = cand[level];
        toDist.clear();
        const float width = (float)(maxBX - minBX), height = (float)(maxBY - minBY);
        const int nCols = (int)(width / W), nRows = (int)(height / W);
        if (nCols < 1 || nRows < 1) return false;
        const int wCell = (int)std::ceil(width / nCols), hCell = (int)std::ceil(height / nRows);
        std::vector<RawKp> cell;
        for (int i = 0; i < nRows; i++) {
            const float iniY = (float)(minBY + i * hCell);
            float maxY = iniY + hCell + 6;
            if (iniY >= maxBY - 3) continue;
            if (maxY > maxBY) maxY = (float)maxBY;
            for (int j = 0; j < nCols; j++) {
                const float iniX = (float)(minBX + j * wCell);
                float maxX = iniX + wCell + 6;
                if (iniX >= maxBX - 6) continue;
                if (maxX > maxBX) maxX = (float)maxBX;
                const int x0 = (int)iniX, x1 = (int)maxX, y0 = (int)iniY, y1 = (int)maxY;
                const uint8_t* roi = im.row(y0) + x0;
                fast9_16(roi, x1 - x0, y1 - y0, im.step, iniTh, cell);
                if (cell.empty()) fast9_16(roi, x1 - x0, y1 - y0, im.step, minTh, cell);
                for (const RawKp& r : cell) {
                    Key k;
                    k.x = (float)r.x + j * wCell;
                    k.y = (float)r.y + i * hCell;
                    k.response = (float)r.score;
                    k.size = 7.f;
                    k.angle = -1.f;
                    k.octave = 0;
                    toDist.push_back(k);
                }
            }
        }
        std::vector<Key>& kps = levelKp[level];
        kps = distribute(toDist, minBX, maxBX, minBY, maxBY, quota[level]);
        const int scaledPatchSize = (int)(PATCH_SIZE * scale[level]);
        for (Key& k : kps) {
            k.x += minBX;
            k.y += minBY;
            k.octave = level;
            k.size = (float)scaledPatchSize;
        }
    }
    for (int level = 0; level < nlevels; ++level)
        for (Key& k : levelKp[level]) k.angle = ic_angle(pyr[level], k.x, k.y, umax);
    return true;
}

/* computeOrbDescriptor, ORBextractor.cc:106-145 */
static void orb_descriptor(const Key& kp, const Image& img, uint8_t* desc) {
    const float factorPI = (float)(3.14159265358979323846 / 180.f);
    const float angle = kp.angle * factorPI;
    const float a = cosf(angle), b = sinf(angle);
    const uint8_t* c = img.row(cv_round(kp.y)) + cv_round(kp.x);
    const int step = img.step;
    const int8_t* pat = kPattern;
    for (int i = 0; i < 32; ++i, pat += 32) {
        int val = 0;
        for (int k = 0; k < 8; k++) {
            const int8_t* p = pat + 4 * k; /* (x0,y0,x1,y1): points 2k, 2k+1 of this byte's 16 */
            const int t0 = c[cv_round(p[0] * b + p[1] * a) * step + cv_round(p[0] * a - p[1] * b)];
            const int t1 = c[cv_round(p[2] * b + p[3] * a) * step + cv_round(p[2] * a - p[3] * b)];
            val |= (t0 < t1) << k;
        }
        desc[i] = (uint8_t)val;
    }
}

extern "C" {

orc_extractor* orc_extractor_create(int nfeatures, float scaleFactor, int nlevels, int iniTh, int minTh) {
    if (nlevels < 1 || nfeatures < 1) return nullptr;
    orc_extractor* e = new orc_extractor;
    e->nfeatures = nfeatures;
    e->scaleFactor = scaleFactor;
    e->nlevels = nlevels;
    e->iniTh = iniTh;
    e->minTh = minTh;
    e->scale.resize(nlevels);
    e->sigma2.resize(nlevels);
    e->scale[0] = 1.0f;
    e->sigma2[0] = 1.0f;
    for (int i = 1; i < nlevels; i++) {
        e->scale[i] = (float)(e->scale[i - 1] * e->scaleFactor);
        e->sigma2[i] = e->scale[i] * e->scale[i];
    }
    e->invScale.resize(nlevels);
    e->invSigma2.resize(nlevels);
    for (int i = 0; i < nlevels; i++) {
        e->invScale[i] = 1.0f / e->scale[i];
        e->invSigma2[i] = 1.0f / e->sigma2[i];
    }
    e->pyr.resize(nlevels);
    e->blurred.resize(nlevels);
    e->cand.resize(nlevels);
    e->levelKp.resize(nlevels);
    e->quota.resize(nlevels);
    const float factor = (float)(1.0f / e->scaleFactor);
    float nDesired = nfeatures * (1 - factor) / (1 - (float)pow((double)factor, (double)nlevels));
    int sum = 0;
    for (int level = 0; level < nlevels - 1; level++) {
        e->quota[level] = cv_round(nDesired);
        sum += e->quota[level];
        nDesired *= factor;
    }
    e->quota[nlevels - 1] = std::max(nfeatures - sum, 0);

    e->umax.resize(HALF_PATCH_SIZE + 1);
    int v, v0;
    const int vmax = (int)std::floor(HALF_PATCH_SIZE * std::sqrt(2.f) / 2 + 1);
    const int vmin = (int)std::ceil(HALF_PATCH_SIZE * std::sqrt(2.f) / 2);
    const double hp2 = HALF_PATCH_SIZE * HALF_PATCH_SIZE;
    for (v = 0; v <= vmax; ++v) e->umax[v] = cv_round(std::sqrt(hp2 - v * v));
    for (v = HALF_PATCH_SIZE, v0 = 0; v >= vmin; --v) {
        while (e->umax[v0] == e->umax[v0 + 1]) ++v0;
        e->umax[v] = v0;
        ++v0;
    }
    return e;
}

void orc_extractor_destroy(orc_extractor* e) { delete e; }

int orc_extract(orc_extractor* e, const uint8_t* img, int rows, int cols, int step, int lap0, int lap1,
                orc_keypoint* kps, uint8_t* desc, int cap, int* n) {
    if (n) *n = 0;
    if (!img || rows <= 0 || cols <= 0) return -1; /* ORBextractor.cc:1072-1073 */
    e->pyramid(img, rows, cols, step);
    if (!e->keypoints()) return -2;
    int total = 0;
    for (int l = 0; l < e->nlevels; l++) total += (int)e->levelKp[l].size();
    if (n) *n = total;
    if (total > cap) return -2;

    int monoIndex = 0, stereoIndex = total - 1;
    uint8_t d[32];
    for (int level = 0; level < e->nlevels; ++level) {
        std::vector<Key>& lk = e->levelKp[level];
        Image& bl = e->blurred[level];
        const Image& src = e->pyr[level];
        if (lk.empty()) { bl.w = bl.h = 0; continue; }
        bl.w = src.w; bl.h = src.h; bl.step = src.w;
        bl.buf.resize((size_t)src.w * src.h);
        bl.data = bl.buf.data();
        gaussian7_u8(src.data, src.w, src.h, src.step, bl.data, bl.step);
        const float scale = e->scale[level];
        for (const Key& k0 : lk) {
            orb_descriptor(k0, bl, d);
            Key k = k0;
            if (level != 0) { k.x *= scale; k.y *= scale; }
            int slot;
            if (k.x >= lap0 && k.x <= lap1) slot = stereoIndex--;
            else slot = monoIndex++;
            orc_keypoint& o = kps[slot];
            o.x = k.x; o.y = k.y; o.size = k.size; o.angle = k.angle; o.response = k.response;
            o.octave = k.octave; o.class_id = -1;
            memcpy(desc + (size_t)slot * 32, d, 32);
        }
    }
    return monoIndex;
}

int orc_num_levels(const orc_extractor* e) { return e->nlevels; }
float orc_scale_factor(const orc_extractor* e, int level) { return e->scale[level]; }
int orc_features_per_level(const orc_extractor* e, int level) { return e->quota[level]; }
int orc_umax(const orc_extractor* e, int v) { return e->umax[v]; }

int orc_level_size(const orc_extractor* e, int level, int* w, int* h) {
    if (level < 0 || level >= e->nlevels) return -1;
    *w = e->pyr[level].w;
    *h = e->pyr[level].h;
    return 0;
}

int orc_pyramid_level(const orc_extractor* e, int level, int padded, uint8_t* dst) {
    if (level < 0 || level >= e->nlevels) return -1;
    const Image& im = e->pyr[level];
    const int B = padded ? EDGE_THRESHOLD : 0;
    const int W = im.w + 2 * B;
    for (int y = -B; y < im.h + B; y++) memcpy(dst + (size_t)(y + B) * W, im.data + (ptrdiff_t)y * im.step - B, W);
    return 0;
}

int orc_blurred_level(const orc_extractor* e, int level, uint8_t* dst) {
    if (level < 0 || level >= e->nlevels) return -1;
    const Image& im = e->blurred[level];
    if (im.w == 0) return 1;
    memcpy(dst, im.data, (size_t)im.w * im.h);
    return 0;
}

static int copy_keys(const std::vector<Key>& v, orc_keypoint* out, int cap) {
    if (out)
        for (int i = 0; i < (int)v.size() && i < cap; i++) {
            out[i].x = v[i].x; out[i].y = v[i].y; out[i].size = v[i].size; out[i].angle = v[i].angle;
            out[i].response = v[i].response; out[i].octave = v[i].octave; out[i].class_id = -1;
        }
    return (int)v.size();
}

int orc_fast_candidates(const orc_extractor* e, int level, orc_keypoint* out, int cap) {
    return copy_keys(e->cand[level], out, cap);
}
int orc_level_keypoints(const orc_extractor* e, int level, orc_keypoint* out, int cap) {
    return copy_keys(e->levelKp[level], out, cap);
}

void orc_resize_linear_u8(const uint8_t* src, int sw, int sh, int sstep, uint8_t* dst, int dw, int dh, int dstep) {
    resize_linear_u8(src, sw, sh, sstep, dst, dw, dh, dstep);
}
void orc_gaussian7_u8(const uint8_t* src, int w, int h, int sstep, uint8_t* dst, int dstep) {
    gaussian7_u8(src, w, h, sstep, dst, dstep);
}
int orc_fast9_16(const uint8_t* roi, int w, int h, int step, int threshold, orc_keypoint* out, int cap) {
    std::vector<RawKp> v;
    fast9_16(roi, w, h, step, threshold, v);
    for (int i = 0; i < (int)v.size() && i < cap; i++) {
        out[i].x = (float)v[i].x; out[i].y = (float)v[i].y; out[i].size = 7.f; out[i].angle = -1.f;
        out[i].response = (float)v[i].score; out[i].octave = 0; out[i].class_id = -1;
    }
    return (int)v.size();
}
float orc_fast_atan2(float y, float x) { return fast_atan2(y, x); }

/* Pyramid access for the stereo oracle (match_oracle.cpp). */
const uint8_t* orc__pyr_data(const orc_extractor* e, int level, int* w, int* h, int* step) {
    const Image& im = e->pyr[level];
    *w = im.w; *h = im.h; *step = im.step;
    return im.data;
}
float orc__inv_scale(const orc_extractor* e, int level) { return e->invScale[level]; }

}  /* extern "C" */
