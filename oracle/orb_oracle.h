/* orb_oracle.h — CPU ORACLE for the ORB front-end hot path.  TEST INFRASTRUCTURE ONLY.
 *
 * This is a plain C++17 restatement (no OpenCV, no CUDA) of the reference algorithm
 *   /root/reference/src/ORBextractor.cc:70-145,408-878,1059-1177   (extractor)
 *   /root/reference/src/ORBmatcher.cc:36-38,311-375,870-908,2591-2607 (Hamming / top-2 core)
 *   /root/reference/src/Frame.cc:797-967,1119-1145                  (stereo matching)
 * plus the integer/float arithmetic of the OpenCV primitives those files call (OpenCV is a
 * third-party dependency that is NOT vendored in the reference: CMakeLists.txt:36-42
 * `find_package(OpenCV 4.0)`; the arithmetic restated here is pinned against cv2 4.13.0 by
 * tools/make_goldens.py -> tests/golden/, see tests/test_oracle_golden.py).
 *
 * Parity is PINNED twice: the OpenCV primitives against the real cv2 4.13 (above), and the extractor, the
 * stereo matcher, every matcher function of the path and the camera-model members they call against the
 * reference's own sources compiled in this container (oracle/Makefile -> oracle/_ref/liborb_ref.so;
 * tests/test_ref_pinned.py, test_ref_triangulation.py, test_initialization.py): identical keypoints, angles,
 * descriptors, pyramids, uRight and depth on every golden case, identical matches on every search.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
 * load this library.  The product (liborbfe.so) never links, imports or calls it.
 *
 * Defined deviations from the reference:
 *  (i)   quadtree equal-size tie-break = node creation order (the reference sorts by heap address, which is
 *        creation order under a monotonic allocator -- the one oracle/_ref is run with -- and differs under
 *        glibc malloc, where the reference's own output changes with the state of the heap);
 *  (ii)  stereo maxD is an explicit argument (the reference computes mbf/mb, src/Frame.cc:827-829, with mb = mbf/fx
 *        from the Frame constructor, :173; callers pass that float, which is fx up to one rounding);
 *  (iii) empty stereo match list => no median cut (the reference indexes an empty vector);
 *  (iv)  float arithmetic is IEEE binary32 without FMA contraction (-ffp-contract=off).
 */
#ifndef ORB_ORACLE_H
#define ORB_ORACLE_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* Layout-identical to cv::KeyPoint (7 x 4 bytes). */
typedef struct {
    float x, y;
    float size;
    float angle;
    float response;
    int32_t octave;
    int32_t class_id;
} orc_keypoint;

typedef struct {
    int32_t best1;    /* smallest distance (256 if no candidate) */
    int32_t best2;    /* second smallest of the multiset (256 if <2 candidates) */
    int32_t idx;      /* index of best1, lowest index on ties, -1 if none */
    int32_t accepted; /* TH + ratio test outcome */
} orc_match;

typedef struct orc_extractor orc_extractor;

/* ORBextractor::ORBextractor, src/ORBextractor.cc:408-468 */
orc_extractor* orc_extractor_create(int nfeatures, float scaleFactor, int nlevels, int iniThFAST, int minThFAST);
void orc_extractor_destroy(orc_extractor*);

/* ORBextractor::operator(), src/ORBextractor.cc:1068-1150.  Returns monoIndex, -1 on an empty image,
 * -2 if cap is too small or the image is too small for the cell grid.  *n receives the keypoint count. */
int orc_extract(orc_extractor*, const uint8_t* img, int rows, int cols, int step, int lap0, int lap1,
                orc_keypoint* kps, uint8_t* desc, int cap, int* n);

/* Introspection used by the stage-level parity tests (state of the last orc_extract call). */
int orc_num_levels(const orc_extractor*);
float orc_scale_factor(const orc_extractor*, int level);
int orc_features_per_level(const orc_extractor*, int level);
int orc_umax(const orc_extractor*, int v);
int orc_level_size(const orc_extractor*, int level, int* w, int* h);
/* padded != 0: copy the (h+38)x(w+38) bordered buffer, else the h x w interior. dst is dense row-major. */
int orc_pyramid_level(const orc_extractor*, int level, int padded, uint8_t* dst);
int orc_blurred_level(const orc_extractor*, int level, uint8_t* dst);
/* vToDistributeKeys of a level (coords relative to minBorder=16), in (cell row, cell col, y, x) order. */
int orc_fast_candidates(const orc_extractor*, int level, orc_keypoint* out, int cap);
/* Post-quadtree + orientation keypoints of a level (level coordinates, before the scale multiply). */
int orc_level_keypoints(const orc_extractor*, int level, orc_keypoint* out, int cap);

/* OpenCV primitive restatements exposed one by one so tools/make_goldens.py can pin them to cv2. */
void orc_resize_linear_u8(const uint8_t* src, int sw, int sh, int sstep, uint8_t* dst, int dw, int dh, int dstep);
void orc_gaussian7_u8(const uint8_t* src, int w, int h, int sstep, uint8_t* dst, int dstep);
int orc_fast9_16(const uint8_t* roi, int w, int h, int step, int threshold, orc_keypoint* out, int cap);
float orc_fast_atan2(float y, float x);
/* cv::SVD::compute on a row-major 4x4 float matrix: singular values and vt (triangulate_oracle.cpp; bit-exact against cv2.SVDecomp 4.13) */
void orc_svd4x4_vt(const float* A, float* w, float* vt);

/* ORBmatcher::DescriptorDistance, src/ORBmatcher.cc:2591-2607 */
int orc_descriptor_distance(const uint8_t* a, const uint8_t* b);

/* Top-2 core of SearchByBoW (src/ORBmatcher.cc:311-375 KF-Frame `<=TH`, 870-908 KF-KF `<TH`) as a brute-force
 * scan of the whole database in index order.  strict_lt selects `best1 < th` (KF-KF) vs `best1 <= th`. */
void orc_hamming_top2(const uint8_t* q, int nq, const uint8_t* db, int64_t ndb, int th, float ratio,
                      int strict_lt, orc_match* out);

/* Frame::ComputeStereoMatches, src/Frame.cc:797-967.  The pyramids are those held by exL / exR after
 * their last orc_extract call.  uRight/depth get -1 where unmatched.  Returns the number of matches kept. */
int orc_stereo_match(const orc_extractor* exL, const orc_extractor* exR,
                     const orc_keypoint* kpsL, const uint8_t* descL, int nL,
                     const orc_keypoint* kpsR, const uint8_t* descR, int nR,
                     float mbf, float minD, float maxD, float* uRight, float* depth);

/* Matching core of Frame::ComputeStereoFishEyeMatches, src/Frame.cc:1137-1144: BFMatcher(NORM_HAMMING)
 * knnMatch(k=2) + `d0 < d1*ratio` evaluated in double.  match[i] = train index or -1.  Also returns d0/d1. */
void orc_knn2_ratio(const uint8_t* q, int nq, const uint8_t* t, int nt, double ratio,
                    int32_t* match, int32_t* d0, int32_t* d1);

#ifdef __cplusplus
}
#endif
#endif
