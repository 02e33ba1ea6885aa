/* synth.cpp — deterministic synthetic stereo frames for tests and bench.py (SURVEY.md §8d "corners-N").
 *
 * Integer-only, so every platform produces the same bytes.  PRNG: splitmix64 seeded 0x0B5EED00 + frame index.
 * Scene: mid-grey + 5x5-box-filtered uniform noise (background plane whose disparity grows with y from 2 to
 * 20 px) + ~400 per 752x480 filled convex shapes (axis rectangles, parallelograms, triangles) each on its own
 * fronto-parallel plane with disparity in [background, 60] px, painted far-to-near; the right view is the same
 * scene shifted by the per-plane disparity; both views get independent +-2 sensor noise.
 * Not part of the reference: the reference ships no synthetic data (datasets are external downloads). */
#include <stdint.h>
#include <stdlib.h>
#include <algorithm>
#include <vector>

namespace {
struct Rng {
    uint64_t s;
    uint64_t next() {
        uint64_t z = (s += 0x9E3779B97F4A7C15ull);
        z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
        z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
        return z ^ (z >> 31);
    }
    int below(int n) { return (int)(next() % (uint64_t)n); }
    int range(int lo, int hi) { return lo + below(hi - lo + 1); }
};

struct Shape {
    int nv;
    int vx[4], vy[4];
    int grey, disp, order;
};

inline int clamp255(int v) { return v < 0 ? 0 : (v > 255 ? 255 : v); }

void paint(uint8_t* img, int W, int H, const Shape& s, int shift) {
    int x0 = s.vx[0], x1 = s.vx[0], y0 = s.vy[0], y1 = s.vy[0];
    for (int i = 1; i < s.nv; i++) {
        x0 = std::min(x0, s.vx[i]); x1 = std::max(x1, s.vx[i]);
        y0 = std::min(y0, s.vy[i]); y1 = std::max(y1, s.vy[i]);
    }
    long area2 = 0;
    for (int i = 0; i < s.nv; i++) {
        const int j = (i + 1) % s.nv;
        area2 += (long)s.vx[i] * s.vy[j] - (long)s.vx[j] * s.vy[i];
    }
    if (area2 == 0) return;
    const int sgn = area2 > 0 ? 1 : -1;
    for (int y = std::max(y0, 0); y <= std::min(y1, H - 1); y++)
        for (int x = x0; x <= x1; x++) {
            const int xs = x - shift;
            if (xs < 0 || xs >= W) continue;
            bool in = true;
            for (int i = 0; i < s.nv && in; i++) {
                const int j = (i + 1) % s.nv;
                const long e = (long)(s.vx[j] - s.vx[i]) * (y - s.vy[i]) - (long)(s.vy[j] - s.vy[i]) * (x - s.vx[i]);
                if (e * sgn < 0) in = false;
            }
            if (in) img[(size_t)y * W + xs] = (uint8_t)s.grey;
        }
}
}  // namespace

extern "C" {

/* left / right: dense W*H u8 buffers; right may be NULL.  Returns 0. */
int orbfe_synth_stereo(uint64_t frame_index, int W, int H, uint8_t* left, uint8_t* right) {
    if (W < 8 || H < 8 || !left) return -1;
    Rng rng{0x0B5EED00ull + frame_index};
    const int PADX = 32, MAXBG = 20;
    const int WN = W + PADX + MAXBG + 8;
    std::vector<int16_t> raw((size_t)H * WN), lp((size_t)H * WN);
    for (size_t i = 0; i < raw.size(); i += 4) {
        uint64_t r = rng.next();
        for (int k = 0; k < 4 && i + k < raw.size(); k++, r >>= 16) raw[i + k] = (int16_t)((int)(r & 0xFFFF) % 121 - 60);
    }
    /* 5x5 box filter with clamped borders */
    std::vector<int> hs((size_t)H * WN);
    for (int y = 0; y < H; y++)
        for (int x = 0; x < WN; x++) {
            int a = 0;
            for (int k = -2; k <= 2; k++) a += raw[(size_t)y * WN + std::min(std::max(x + k, 0), WN - 1)];
            hs[(size_t)y * WN + x] = a;
        }
    for (int y = 0; y < H; y++)
        for (int x = 0; x < WN; x++) {
            int a = 0;
            for (int k = -2; k <= 2; k++) a += hs[(size_t)std::min(std::max(y + k, 0), H - 1) * WN + x];
            lp[(size_t)y * WN + x] = (int16_t)((a + 25 * 1000 + 12) / 25 - 1000);
        }
    for (int y = 0; y < H; y++) {
        const int dbg = 2 + (y * (MAXBG - 2)) / H;
        for (int x = 0; x < W; x++) {
            left[(size_t)y * W + x] = (uint8_t)clamp255(128 + lp[(size_t)y * WN + x + PADX]);
            if (right) right[(size_t)y * W + x] = (uint8_t)clamp255(128 + lp[(size_t)y * WN + x + PADX + dbg]);
        }
    }
    const int nshape = (int)((400LL * W * H) / 360960);
    std::vector<Shape> shapes(nshape);
    for (int i = 0; i < nshape; i++) {
        Shape& s = shapes[i];
        const int type = rng.below(3);
        const int cx = rng.below(W), cy = rng.below(H);
        const int a = rng.range(6, 60), b = rng.range(6, 60);
        s.grey = rng.below(256);
        const int dbg = 2 + (cy * (MAXBG - 2)) / H;
        s.disp = rng.range(dbg, 60);
        s.order = i;
        if (type == 0) {
            s.nv = 4;
            s.vx[0] = cx - a / 2; s.vy[0] = cy - b / 2;
            s.vx[1] = cx + a / 2; s.vy[1] = cy - b / 2;
            s.vx[2] = cx + a / 2; s.vy[2] = cy + b / 2;
            s.vx[3] = cx - a / 2; s.vy[3] = cy + b / 2;
        } else if (type == 1) {
            const int dx = rng.range(-30, 30), dy = rng.range(4, 30);
            const int ex = -dy * b / 32, ey = dx * b / 32; /* roughly perpendicular second edge */
            s.nv = 4;
            s.vx[0] = cx;           s.vy[0] = cy;
            s.vx[1] = cx + dx;      s.vy[1] = cy + dy;
            s.vx[2] = cx + dx + ex; s.vy[2] = cy + dy + ey;
            s.vx[3] = cx + ex;      s.vy[3] = cy + ey;
        } else {
            s.nv = 3;
            s.vx[0] = cx;                      s.vy[0] = cy - b / 2;
            s.vx[1] = cx + rng.range(-a, a);   s.vy[1] = cy + b / 2;
            s.vx[2] = cx + rng.range(-a, a);   s.vy[2] = cy + rng.range(-b / 2, b / 2);
        }
    }
    std::stable_sort(shapes.begin(), shapes.end(), [](const Shape& p, const Shape& q) { return p.disp < q.disp; });
    for (const Shape& s : shapes) {
        paint(left, W, H, s, 0);
        if (right) paint(right, W, H, s, s.disp);
    }
    for (int v = 0; v < (right ? 2 : 1); v++) {
        uint8_t* img = v ? right : left;
        const size_t n = (size_t)W * H;
        for (size_t i = 0; i < n; i += 8) {
            uint64_t r = rng.next();
            for (int k = 0; k < 8 && i + k < n; k++, r >>= 8) img[i + k] = (uint8_t)clamp255(img[i + k] + (int)(r & 0xFF) % 5 - 2);
        }
    }
    return 0;
}

/* Descriptor database for config 5: i.i.d. uniform bytes; 25 % of the queries are database rows with
 * k in [0,60] random bit flips so that both the TH_LOW and the ratio test accept and reject. */
int orbfe_synth_descriptors(uint64_t seed, uint8_t* db, int64_t ndb, uint8_t* q, int nq) {
    if (!db || !q || ndb < 2 || nq < 1) return -1;
    Rng rng{0x0B5EED00ull + seed};
    uint64_t* d64 = (uint64_t*)db;
    for (int64_t i = 0; i < ndb * 4; i++) d64[i] = rng.next();
    uint64_t* q64 = (uint64_t*)q;
    for (int64_t i = 0; i < (int64_t)nq * 4; i++) q64[i] = rng.next();
    for (int i = 0; i < nq; i += 4) {
        const int64_t row = (int64_t)(rng.next() % (uint64_t)ndb);
        for (int k = 0; k < 32; k++) q[(size_t)i * 32 + k] = db[(size_t)row * 32 + k];
        const int flips = rng.below(61);
        for (int f = 0; f < flips; f++) {
            const int bit = rng.below(256);
            q[(size_t)i * 32 + bit / 8] ^= (uint8_t)(1u << (bit % 8));
        }
    }
    return 0;
}

}  /* extern "C" */
