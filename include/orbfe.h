/* orbfe.h — C ABI of liborbfe.so: the B200 (sm_100a) ORB front-end.
 *
 * This is the drop-in boundary for ORB-SLAM3's data-parallel hot path.  The reference has no FFI of its
 * own: the boundary is its C++ class surface, so every entry point below names the reference member
 * function (file:line under /root/reference) whose work it takes over.  The C++ shims in
 * include/orbfe_shim/ re-declare those classes with the reference's signatures and forward here.
 *
 * Conventions: plain pointers and sizes only; every function returns 0 on success or a negative
 * ORBFE_E_* code (orbfe_last_error() gives the text, per thread); nothing here falls back to the CPU —
 * if no CUDA device is usable the call fails.  `stream` arguments are cudaStream_t values passed as
 * void* (NULL = the object's own stream).  "device" pointers must belong to the object's GPU.
 */
#ifndef ORBFE_H
#define ORBFE_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ORBFE_OK 0
#define ORBFE_E_EMPTY (-1)    /* empty image: ORBextractor::operator() returns -1 (src/ORBextractor.cc:1072-1073) */
#define ORBFE_E_INVALID (-2)  /* bad argument / unsupported shape */
#define ORBFE_E_CUDA (-3)     /* CUDA runtime error (text in orbfe_last_error) */
#define ORBFE_E_CAPACITY (-4) /* caller buffer or object capacity too small */
#define ORBFE_E_NCCL (-5)

/* Layout-identical to cv::KeyPoint (pt.x, pt.y, size, angle, response, octave, class_id): 28 bytes. */
typedef struct orbfe_keypoint {
    float x, y;
    float size;
    float angle;
    float response;
    int32_t octave;
    int32_t class_id;
} orbfe_keypoint;

/* The five constructor arguments of ORBextractor (include/ORBextractor.h:49-50). */
typedef struct orbfe_params {
    int32_t nfeatures;
    float scale_factor;
    int32_t nlevels;
    int32_t ini_th_fast;
    int32_t min_th_fast;
} orbfe_params;

/* Result of the top-2 Hamming search for one query (src/ORBmatcher.cc:311-375, 870-908). */
typedef struct orbfe_match {
    int32_t best1;    /* bestDist1 (256 when the database is empty) */
    int32_t best2;    /* bestDist2 */
    int32_t idx;      /* bestIdx, global database row, lowest row on ties; -1 when empty */
    int32_t accepted; /* 1 iff best1 passes TH_LOW and the nnratio test */
} orbfe_match;

typedef struct orbfe_extractor orbfe_extractor;
typedef struct orbfe_matcher orbfe_matcher;

const char* orbfe_last_error(void);
/* The one-frame host entry points of the matchers (orbfe_search_by_projection, orbfe_search_by_bow, orbfe_search_for_triangulation,
 * orbfe_search_for_initialization) stage their arguments through a block of pinned host and device memory that belongs to the calling
 * thread; it grows to the largest call seen and is freed when the thread ends, or here. */
void orbfe_thread_release(void);
int orbfe_device_count(void);
const char* orbfe_version(void);

/* ------------------------------------------------------------------------------------------------ extractor */

/* Replaces ORBextractor::ORBextractor (src/ORBextractor.cc:408-468): scale tables, per-level quotas, umax,
 * pattern upload — plus everything a GPU needs once: the image shape and the largest batch are fixed here so
 * that no call below allocates.  width/height are the input image size; max_batch >= 1.
 * Limits (ORBFE_E_INVALID / ORBFE_E_CAPACITY at creation, never a wrong result): width, height <= 4000; a per-level feature quota
 * <= 16381 (nfeatures up to about 75000 at scale 1.2; up to a quota of 2045 the quadtree of a level lives in shared memory, beyond
 * in global scratch); every level at least 67 px wide and high (the reference divides by zero below that,
 * src/ORBextractor.cc:782-785). */
int orbfe_extractor_create(const orbfe_params* params, int device, int width, int height, int max_batch,
                           orbfe_extractor** out);
void orbfe_extractor_destroy(orbfe_extractor* ex);

/* Getters of ORBextractor (include/ORBextractor.h:61-81): GetLevels, GetScaleFactors, GetInverseScaleFactors,
 * GetScaleSigmaSquares, GetInverseScaleSigmaSquares; plus mnFeaturesPerLevel and umax for the parity tests. */
int orbfe_get_levels(const orbfe_extractor* ex);
int orbfe_get_scale_factors(const orbfe_extractor* ex, float* scale, float* inv_scale, float* sigma2, float* inv_sigma2);
int orbfe_get_features_per_level(const orbfe_extractor* ex, int32_t* quota);
int orbfe_get_umax(const orbfe_extractor* ex, int32_t* umax16);
/* Upper bound of keypoints one frame can return (sum over levels of quota+3, src/ORBextractor.cc:655-735). */
int orbfe_keypoint_capacity(const orbfe_extractor* ex);
int orbfe_level_size(const orbfe_extractor* ex, int level, int* width, int* height);

/* Replaces int ORBextractor::operator()(image, mask, keypoints, descriptors, vLappingArea)
 * (src/ORBextractor.cc:1068-1150) for one host image.  step = bytes per image row.  kps/desc receive *n
 * entries (desc is n x 32 bytes); *mono_index is the reference's return value.  rows/cols must equal the
 * shape given at creation.  Returns ORBFE_E_EMPTY for a NULL/0-sized image like the reference's -1. */
int orbfe_extract(orbfe_extractor* ex, const uint8_t* image, int rows, int cols, size_t step, int lap0, int lap1,
                  orbfe_keypoint* kps, uint8_t* desc, int cap, int* n, int* mono_index);

/* The same operator() over a batch of frames that already sit in device memory, densely packed
 * (frame f at d_images + f*rows*cols).  Asynchronous on `stream`; results stay on the device until fetched.
 * This is the call the roofline numbers are quoted on (frames are independent: SURVEY.md §8e). */
int orbfe_extract_batch_device(orbfe_extractor* ex, const uint8_t* d_images, int nframes, int lap0, int lap1,
                               void* stream);
/* Frame's stereo constructor (src/Frame.cc:104-170) for a batch of pairs that already sit in device memory: the reference runs
 * the two ExtractORB calls on two threads (:119-122) and ComputeStereoMatches when both are back (:136); here the left images go
 * through exL on its own stream, the right ones through exR on its own stream, the matching follows on the left stream.  `stream`
 * (may be NULL) is the caller's: the work starts after what that stream holds now and is complete, for that stream, after the
 * call (fork / join by events, the host is never blocked).  Results as after orbfe_extract_batch_device + orbfe_stereo_match_batch_device. */
int orbfe_stereo_frames_batch_device(orbfe_extractor* exL, orbfe_extractor* exR, const uint8_t* d_left, const uint8_t* d_right, int nframes, int lap0,
                                     int lap1, float mbf, float min_d, float max_d, void* stream);
/* The same without the matcher: the two extractions of Frame's fisheye-rig constructor (src/Frame.cc:1051-1054), side by side; the
 * caller continues on `stream` with orbfe_fisheye_match_batch_device / orbfe_fisheye_triangulate_batch_device (or any other consumer). */
int orbfe_extract_pair_batch_device(orbfe_extractor* exL, orbfe_extractor* exR, const uint8_t* d_left, const uint8_t* d_right, int nframes, int lap0,
                                    int lap1, void* stream);
/* Same, from host memory (pinned or pageable): host->device copy, extraction, device->host copy of the results.
 * kps: nframes*cap entries, desc: nframes*cap*32 bytes, n / mono_index: nframes ints.  Synchronous. */
int orbfe_extract_batch_host(orbfe_extractor* ex, const uint8_t* h_images, int nframes, int lap0, int lap1,
                             orbfe_keypoint* kps, uint8_t* desc, int cap, int32_t* n, int32_t* mono_index);
/* Frame's stereo constructor for a batch of pairs in one call (src/Frame.cc:104-178: the two ExtractORB threads, then
 * ComputeStereoMatches): uploads, both extractions, the stereo matching and the downloads are pipelined over chunks of frames on
 * the two objects' streams, from one host thread.  Outputs as orbfe_extract_batch_host for each side (cap >= the keypoint
 * capacity; any output pointer may be NULL), plus mvuRight / mvDepth of every pair in u_right / depth [nframes*cap] (-1 = no match;
 * entries beyond a frame's keypoint count are unspecified).  mbf, min_d, max_d as orbfe_stereo_match_batch_device.  Synchronous. */
int orbfe_stereo_frames_batch_host(orbfe_extractor* exL, orbfe_extractor* exR, const uint8_t* h_left, const uint8_t* h_right, int nframes, int lap0,
                                   int lap1, float mbf, float min_d, float max_d, orbfe_keypoint* kps_l, uint8_t* desc_l, int32_t* n_l, int32_t* mono_l,
                                   orbfe_keypoint* kps_r, uint8_t* desc_r, int32_t* n_r, int32_t* mono_r, int cap, float* u_right, float* depth);
/* The same batch call split in two, so that a stream of batches keeps the copy engines and the SMs busy at the same time (the
 * reference has no counterpart: its Frame constructor is synchronous).  _submit enqueues the uploads, both extractions, the stereo
 * matching and the downloads of one batch and returns at once; _wait blocks until the OLDEST submitted batch is complete on the
 * host and then fills its counts.  Up to two batches may be in flight (a third _submit is refused with ORBFE_E_CAPACITY): the
 * upload of batch k+1 overlaps the kernels of batch k, whose downloads overlap the kernels of batch k+1.  Typical loop:
 *     submit(0); for k = 1..: { submit(k); wait(k-1); consume(k-1); }  wait(last);
 * The caller passes a different set of (pinned) output arrays to batches that are in flight together and keeps the input arrays
 * untouched until their batch has been waited for.  Between a _submit and its _wait the two extractors must not be used through
 * any other entry point.  orbfe_stereo_frames_batch_host is _submit followed by _wait. */
int orbfe_stereo_frames_submit(orbfe_extractor* exL, orbfe_extractor* exR, const uint8_t* h_left, const uint8_t* h_right, int nframes, int lap0, int lap1,
                               float mbf, float min_d, float max_d, orbfe_keypoint* kps_l, uint8_t* desc_l, orbfe_keypoint* kps_r, uint8_t* desc_r, int cap,
                               float* u_right, float* depth);
int orbfe_stereo_frames_wait(orbfe_extractor* exL, orbfe_extractor* exR, int32_t* n_l, int32_t* mono_l, int32_t* n_r, int32_t* mono_r);
/* Fetch what the last batch call left on the device (synchronises the object's stream). */
int orbfe_batch_counts(orbfe_extractor* ex, int nframes, int32_t* n, int32_t* mono_index);
int orbfe_batch_result(orbfe_extractor* ex, int frame, orbfe_keypoint* kps, uint8_t* desc, int cap, int* n,
                       int* mono_index);
/* Device views of the last batch's results (for callers that keep everything on the GPU):
 * keypoints[frame*capacity + i], descriptors[(frame*capacity + i)*32], counts[frame], mono[frame]. */
int orbfe_batch_device_views(orbfe_extractor* ex, const orbfe_keypoint** d_kps, const uint8_t** d_desc,
                             const int32_t** d_counts, const int32_t** d_mono);

/* Host mirrors of ORBextractor::mvImagePyramid (include/ORBextractor.h:83; read by src/Frame.cc:804,894-909)
 * and stage-level views used by the parity tests.  padded != 0 returns the (h+38)x(w+38) bordered buffer. */
int orbfe_pyramid_level(orbfe_extractor* ex, int frame, int level, int padded, uint8_t* dst);
/* All levels of one frame in one call (what the shim's mvImagePyramid mirrors need after every operator()): dst_levels[l] receives
 * level l like orbfe_pyramid_level would (NULL entries are skipped); the border kernel and the copies share one wait. */
int orbfe_pyramid_levels(orbfe_extractor* ex, int frame, int padded, uint8_t* const* dst_levels);
/* Page-lock / release a host buffer the caller keeps reusing as a download target (cudaHostRegister / cudaHostUnregister): the
 * copies into it then run at link speed and without a staging copy.  The shim pins its mvImagePyramid mirrors with it. */
int orbfe_pin_host(void* p, size_t bytes);
int orbfe_unpin_host(void* p);
int orbfe_blurred_level(orbfe_extractor* ex, int frame, int level, uint8_t* dst);
/* vToDistributeKeys of one level (src/ORBextractor.cc:846-850): x, y (relative to the 16 px border), response. */
int orbfe_fast_candidates(orbfe_extractor* ex, int frame, int level, orbfe_keypoint* out, int cap, int* n);

/* Per-stage device timing (replaces the REGISTER_TIMES chrono brackets of src/Frame.cc:116-127,136-144): when
 * enabled, CUDA events on the launch stream bracket every stage; orbfe_stage_times accumulates the finished
 * launches into ms[8] = {pyramid, fast, blur, quadtree, assemble, orient+describe, stereo match, stereo median}
 * (adds to the caller's values; synchronises) and returns the number of kernels launched so far in *launches. */
int orbfe_set_profiling(orbfe_extractor* ex, int enable);
int orbfe_stage_times(orbfe_extractor* ex, double* ms8, int64_t* launches);

/* --------------------------------------------------------------------------------------------------- stereo */

/* Replaces Frame::ComputeStereoMatches (src/Frame.cc:797-967) for the frames both extractors processed last:
 * left keypoints/descriptors/pyramids come from exL, right ones from exR (device resident, nothing is
 * re-uploaded).  minD = 0 and maxD = fx in the reference (maxD must be passed: the reference reads an
 * uninitialised member, src/Frame.cc:827).  Asynchronous on `stream` (NULL = exL's stream). */
int orbfe_stereo_match_batch_device(orbfe_extractor* exL, orbfe_extractor* exR, int nframes, float mbf, float minD,
                                    float maxD, void* stream);
/* mvuRight / mvDepth of one frame (include/Frame.h:172-173), -1 where unmatched. */
int orbfe_stereo_result(orbfe_extractor* exL, int frame, float* u_right, float* depth, int cap, int* n);
/* All frames at once: u_right / depth are nframes*cap floats (row f = frame f). Synchronous. */
int orbfe_stereo_result_batch(orbfe_extractor* exL, int nframes, float* u_right, float* depth, int cap);
/* Device views: u_right[frame*capacity + i], depth[...]. */
int orbfe_stereo_device_views(orbfe_extractor* exL, const float** d_u_right, const float** d_depth);

/* ------------------------------------------------------------------------------------------------- matching */

/* Replaces the distance loop of ORBmatcher::DescriptorDistance (src/ORBmatcher.cc:2591-2607) for n row pairs
 * held on the host: out[i] = Hamming(a[i], b[i]).  (The one-pair static member itself is an inline in the
 * shim header: a kernel launch per 32-byte pair would be slower than the eight popcounts.) */
int orbfe_descriptor_distances(int device, const uint8_t* a, const uint8_t* b, int64_t n, int32_t* out);

/* A database of 256-bit descriptors resident on one GPU + the top-2 search over it: the brute-force core of
 * ORBmatcher::SearchByBoW (src/ORBmatcher.cc:311-375 `<=TH_LOW`, 870-908 `<TH_LOW`).  index_base is added
 * to the returned row indices (the shard's first global row). */
int orbfe_matcher_create(int device, int max_queries, int64_t max_db_rows, orbfe_matcher** out);
void orbfe_matcher_destroy(orbfe_matcher* m);
int orbfe_matcher_set_db(orbfe_matcher* m, const uint8_t* h_db, int64_t rows, int64_t index_base);
int orbfe_matcher_set_db_device(orbfe_matcher* m, const uint8_t* d_db, int64_t rows, int64_t index_base);
/* Host queries in, host results out (copies inside). th = TH_LOW (50), ratio = mfNNratio, strict_lt selects
 * `best1 < th` (KF-KF variant) instead of `best1 <= th`. */
int orbfe_matcher_top2(orbfe_matcher* m, const uint8_t* h_queries, int nq, int th, float ratio, int strict_lt,
                       orbfe_match* h_out);
/* Device queries in, device results out, asynchronous on `stream` (NULL = the object's stream).  The per-slice scratch belongs to the
 * object: one search per matcher may be in flight at a time — calls on different streams of the same matcher must be ordered by the
 * caller (events), or use one matcher per stream. */
int orbfe_matcher_top2_device(orbfe_matcher* m, const uint8_t* d_queries, int nq, int th, float ratio,
                              int strict_lt, orbfe_match* d_out, void* stream);
/* Merge step of the sharded database: d_parts holds nshards x nq orbfe_match records (one
 * block per shard, as gathered over NCCL); writes the global top-2 + acceptance into d_out[nq].  New; the
 * reference is single-process. */
int orbfe_top2_merge_device(int device, const orbfe_match* d_parts, int nshards, int nq, int th, float ratio,
                            int strict_lt, orbfe_match* d_out, void* stream);
/* The sharded search in one call (new: the reference is single-process; this is what a multi-GPU
 * LoopClosing::DetectCommonRegionsFromBoW-style search, src/LoopClosing.cc:681-725, calls from C++): `m` holds THIS rank's shard
 * (rows [index_base, index_base + rows), set with orbfe_matcher_set_db*); the call searches it, all-gathers the nq 16-byte records of
 * every rank with ncclAllGather on `nccl_comm` (an ncclComm_t of `nranks` ranks whose rank order is ascending index_base) and
 * merges them into d_out[nq] exactly as one scan of the whole database would (ties to the lowest global row): search, collective
 * and merge are enqueued on `stream` (NULL = the object's stream) with no host synchronisation in between.  nranks == 1 needs no
 * communicator (with one, the one-rank collective still runs).  The first call allocates the gather buffers (nranks x max_queries records).  NCCL is resolved with dlopen
 * (libnccl.so.2, or $ORBFE_NCCL_LIB); ORBFE_E_NCCL when it is missing or a collective fails.  Every rank must make the call. */
int orbfe_hamming_top2_sharded(orbfe_matcher* m, void* nccl_comm, int nranks, const uint8_t* d_queries, int nq, int th, float ratio,
                               int strict_lt, orbfe_match* d_out, void* stream);
/* Convenience for hosts without their own NCCL bootstrap: rank 0 calls orbfe_nccl_unique_id, ships the 128 bytes to the other
 * ranks by any means, every rank calls orbfe_nccl_comm_create (collective) and later orbfe_nccl_comm_destroy. */
int orbfe_nccl_unique_id(uint8_t id128[128]);
int orbfe_nccl_comm_create(int device, int nranks, int rank, const uint8_t id128[128], void** comm);
int orbfe_nccl_comm_destroy(void* comm);
/* The top-2 search for callers that cannot allocate (no counterpart in the reference; it has no GPU memory to
 * manage).  Meant for ORB-SLAM3's per-frame Tracking thread, which runs against a frame deadline while LocalMapping
 * and LoopClosing use the same GPU: cudaMalloc/cudaFree lock the device and cudaFree waits for all of it, so that
 * thread must not allocate; and for processes whose own pool has already taken all device memory, where an
 * allocation would just fail.  Same inputs, outputs and results as orbfe_matcher_top2_device, but no object: the
 * per-slice scratch is a fixed 4 MiB array inside the library image (each GPU that loads the library keeps it until
 * exit).  Limits: nq <= 4096 (more is refused with ORBFE_E_CAPACITY); pointers are device pointers on `device`,
 * 16-byte aligned; `stream` is used as given (NULL = the default stream).  The call performs no cudaMalloc/cudaFree
 * and calls no allocating library routine.  The first call on each device looks up the array's device address
 * (cudaGetSymbolAddress) — make that first call before starting any stream capture.  All calls on a device share the
 * one array, which keeps whatever the last call wrote (nothing needs clearing: each call rewrites every record it
 * reads); two calls whose kernels can overlap on the GPU (different streams, or different threads) race on it —
 * keep them on one stream or otherwise ordered. */
int orbfe_hamming_top2_static(int device, const uint8_t* d_queries, int nq, const uint8_t* d_db, int64_t ndb,
                              int64_t index_base, int th, float ratio, int strict_lt, orbfe_match* d_out, void* stream);
/* Measures the device's POPC issue rate (32-bit population counts per second, whole GPU) with a small kernel; bench.py reads
 * the Hamming search's integer-pipe roofline against it.  No counterpart in the reference. */
int orbfe_probe_popc(int device, double* popc_per_second);
/* One-shot convenience over host memory (creates and destroys a matcher). */
int orbfe_hamming_top2(int device, const uint8_t* q, int nq, const uint8_t* db, int64_t ndb, int th, float ratio,
                       int strict_lt, orbfe_match* out);

/* Replaces BFMatcher(NORM_HAMMING).knnMatch(k=2) + the Lowe test `d0 < d1*0.7` evaluated in double of
 * Frame::ComputeStereoFishEyeMatches (src/Frame.cc:1137-1144).  match[i] = train row or -1; d0/d1 optional. */
int orbfe_knn2_ratio(int device, const uint8_t* q, int nq, const uint8_t* t, int nt, double ratio, int32_t* match,
                     int32_t* d0, int32_t* d1);

/* The same matching step for a batch of fisheye stereo frames that the two extractors have just processed (device resident,
 * vLappingArea given to both): queries = the left keypoints inside the lapping area (rows mono_l .. n_l-1 of each frame, where
 * operator() placed them, src/ORBextractor.cc:1135-1139), train rows = the right ones; d_match_lr[f*cap + row] = matched right row
 * (index into the right frame's keypoints, i.e. trainIdx + monoRight as in src/Frame.cc:1151) or -1, cap = keypoint capacity;
 * d_dist01 (optional, [nframes*cap][2]) receives the two smallest distances.  Asynchronous on `stream`. */
int orbfe_fisheye_match_batch_device(orbfe_extractor* exL, orbfe_extractor* exR, int nframes, double ratio, int32_t* d_match_lr, int32_t* d_dist01,
                                     void* stream);

/* The rest of Frame::ComputeStereoFishEyeMatches (src/Frame.cc:1145-1157) on the device: KannalaBrandt8::TriangulateMatches
 * (src/CameraModels/KannalaBrandt8.cpp:338-407: unproject both keypoints, parallax test, Triangulate's 4x4 SVD, depth signs, the
 * two reprojection tests at 5.991 * mvLevelSigma2[octave]) for every match that orbfe_fisheye_match_batch_device left in d_match_lr,
 * and the `depth > 0.0001f` acceptance: rejected matches are reset to -1 in d_match_lr, accepted ones fill
 *   d_match_rl [nframes*cap]   mvRightToLeftMatch (the last left keypoint that matched a right one keeps it, as the reference's loop),
 *   d_depth    [nframes*cap]   mvDepth (-1 = none),
 *   d_p3d      [nframes*cap*3] mvStereo3Dpoints (camera-1 coordinates; zeros where there is none).
 * rig: mvParameters of mpCamera and mpCamera2 (fx, fy, cx, cy, k1..k4), mRlr row-major and mtlr (Frame.h, from Tlr of the settings).
 * Arithmetic: OpenCV's as pinned by oracle/triangulate_oracle.cpp (4x4 SVD = cv2.SVDecomp bit for bit on the host); the device's
 * tanf / atan2f / cosf / sinf differ from the host libm by an ulp or two, so depths agree to 1e-4 relative, not to the bit. */
typedef struct {
    float cam1[8], cam2[8];
    float R12[9], t12[3];
} orbfe_fisheye_rig;
int orbfe_fisheye_triangulate_batch_device(orbfe_extractor* exL, orbfe_extractor* exR, int nframes, const orbfe_fisheye_rig* rig, int32_t* d_match_lr,
                                           int32_t* d_match_rl, float* d_depth, float* d_p3d, void* stream);
/* KannalaBrandt8::TriangulateMatches (src/CameraModels/KannalaBrandt8.cpp:338-407) for n keypoint pairs given as plain device arrays:
 * d_uv1 / d_uv2 [n][2] pixel coordinates in camera 1 / 2, d_sigma1 / d_sigma2 [n] = mvLevelSigma2 of their octaves;
 * d_depth[i] = the function's return value (-1 when it rejects the pair), d_p3d [n][3] = p3D.  Asynchronous on `stream`. */
/* Frame's fisheye-rig constructor (src/Frame.cc:1023-1101) for a batch of pairs from host memory, the pipelined counterpart of
 * orbfe_stereo_frames_submit: uploads, both extractions with the lapping interval, ComputeStereoFishEyeMatches (knn-2 + ratio 0.7 in the
 * reference, TriangulateMatches, the depth test) and the downloads overlap chunk by chunk; at most two batches in flight, collected
 * in order by orbfe_stereo_frames_wait.  match_lr / match_rl / depth [nframes*cap] = mvLeftToRightMatch / mvRightToLeftMatch (right
 * rows) / mvDepth, p3d [nframes*cap*3] = mvStereo3Dpoints; any output pointer may be NULL; pinned host memory keeps the copies async. */
int orbfe_fisheye_frames_submit(orbfe_extractor* exL, orbfe_extractor* exR, const uint8_t* h_left, const uint8_t* h_right, int nframes, int lap0, int lap1,
                                const orbfe_fisheye_rig* rig, double ratio, orbfe_keypoint* kps_l, uint8_t* desc_l, orbfe_keypoint* kps_r, uint8_t* desc_r, int cap,
                                int32_t* match_lr, int32_t* match_rl, float* depth, float* p3d);
/* Replaces Frame::ComputeStereoFishEyeMatches (src/Frame.cc:1119-1159) as a whole for the frames both extractors processed last (the single
 * frame of the two operator() calls, or a batch), results on the host: match_lr [nframes*cap] = mvLeftToRightMatch (right rows), match_rl =
 * mvRightToLeftMatch, depth = mvDepth (-1: none), p3d [nframes*cap*3] = mvStereo3Dpoints; entries beyond a frame's keypoint counts are
 * unspecified; any output may be NULL.  ratio = 0.7 in the reference.  Synchronous. */
int orbfe_fisheye_stereo_frames(orbfe_extractor* exL, orbfe_extractor* exR, int nframes, const orbfe_fisheye_rig* rig, double ratio, int32_t* match_lr,
                                int32_t* match_rl, float* depth, float* p3d, int cap);
int orbfe_kb8_triangulate_pairs_device(int device, const orbfe_fisheye_rig* rig, const float* d_uv1, const float* d_uv2, const float* d_sigma1,
                                       const float* d_sigma2, int n, float* d_depth, float* d_p3d, void* stream);

/* ------------------------------------------------------------------------------------ bag-of-words assignment (N2) */
/* Replaces Frame::ComputeBoW / KeyFrame::ComputeBoW (src/Frame.cc:724-731, src/KeyFrame.cc:105-114):
 *   mpORBvocabulary->transform(vCurrentDesc, mBowVec, mFeatVec, 4)
 * i.e. DBoW2's TemplatedVocabulary<FORB::TDescriptor, FORB>::transform(features, BowVector&, FeatureVector&, levelsup)
 * (Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1125-1193) with the per-feature tree descent of :1217-1260.
 * The vocabulary is handed over as the rows of the text file that ORBVocabulary::loadFromTextFile reads (:1338-1426), in file
 * order: row r is node r+1 (node 0 is the root); parent[r], is_leaf[r], 32 descriptor bytes, weight[r]; words are numbered in
 * the order their rows appear.  k, L, scoring (ScoringType 0..5), weighting (WeightingType 0..3) are the four numbers of the
 * file's first line.  max_features (<= 4096) / max_batch size the result buffers. */
typedef struct orbfe_vocabulary orbfe_vocabulary;
int orbfe_vocabulary_create(int device, int k, int L, int scoring, int weighting, int nrows, const int32_t* parent, const uint8_t* is_leaf,
                            const uint8_t* desc, const double* weight, int max_features, int max_batch, orbfe_vocabulary** out);
void orbfe_vocabulary_destroy(orbfe_vocabulary* v);
int orbfe_vocabulary_words(const orbfe_vocabulary* v);
/* A batch of frames on the device: frame f owns d_counts[f] descriptors at d_desc + (f*row_stride + i)*32 — the layout of
 * orbfe_batch_device_views (row_stride = keypoint capacity), so the extractor's buffers can be passed as they are.
 * Asynchronous on `stream` (NULL = the object's stream); results stay in the object until the next call.
 * A frame with d_counts[f] > max_features (the extractor's capacity can exceed 4096 for large nfeatures) is not silently
 * truncated: its bow / feature-vector counts are set to -1 on the device and orbfe_bow_result returns ORBFE_E_CAPACITY for it. */
int orbfe_bow_transform_batch_device(orbfe_vocabulary* v, const uint8_t* d_desc, int row_stride, const int32_t* d_counts, int nframes,
                                     int levelsup, void* stream);
/* Results of frame `frame` of the last batch, to the host (any pointer may be NULL):
 *   word_ids / node_ids [nfeatures]: per feature the WordId and the NodeId at level L - levelsup (0 when that level is not
 *     reached; the reference leaves *nid unset then);
 *   bow_words / bow_values [*nbow]: mBowVec in key order (features whose word weight is 0 are left out, :1157);
 *   fv_nodes [*nfv], fv_starts [*nfv + 1], fv_feats: mFeatVec in key order, node e owns fv_feats[fv_starts[e] .. fv_starts[e+1]). */
int orbfe_bow_result(orbfe_vocabulary* v, int frame, int nfeatures, int32_t* word_ids, int32_t* node_ids, int32_t* bow_words, double* bow_values,
                     int32_t* nbow, int32_t* fv_nodes, int32_t* fv_starts, int32_t* fv_feats, int32_t* nfv);
/* One frame over host memory (copies inside, synchronous). */
int orbfe_bow_transform(orbfe_vocabulary* v, const uint8_t* desc, int n, int levelsup, int32_t* word_ids, int32_t* node_ids, int32_t* bow_words,
                        double* bow_values, int32_t* nbow, int32_t* fv_nodes, int32_t* fv_starts, int32_t* fv_feats, int32_t* nfv);
/* Device views of the last batch (frame stride max_features; fv_starts: max_features + 1), for callers that keep going on the GPU. */
int orbfe_bow_device_views(orbfe_vocabulary* v, const int32_t** d_feat_word, const int32_t** d_feat_node, const int32_t** d_bow_words,
                           const double** d_bow_values, const int32_t** d_bow_count, const int32_t** d_fv_nodes, const int32_t** d_fv_starts,
                           const int32_t** d_fv_feats, const int32_t** d_fv_count);

/* Replaces ORBmatcher::SearchByBoW(KeyFrame* pKF, Frame &F, vector<MapPoint*> &vpMapPointMatches) (src/ORBmatcher.cc:269-470, the
 * F.Nleft == -1 path) for batches of (keyframe, frame) pairs: the walk over the vocabulary nodes present in both feature vectors,
 * the best / second-best scan of each keyframe feature over the frame's still unmatched features of the node, TH_LOW and the
 * mfNNratio test, and the rotation-consistency histogram (ComputeThreeMaxima, :2545-2586) when check_orientation is set.  The
 * result is the reference's, including its serial rule that a matched frame feature is skipped by the keyframe features after it.
 * All frames live in strided arrays: descriptors / keypoints / has_map_point at (frame*row_stride + row), the feature vectors in
 * the layout of orbfe_bow_device_views (stride max_features, starts max_features + 1); pair p matches keyframe d_pair_kf[p] against
 * frame d_pair_f[p].  d_has_map_point: nonzero iff vpMapPointsKF[row] && !isBad() (NULL = all).  d_match_f[p*row_stride + i] = the
 * keyframe row whose map point goes to vpMapPointMatches[i], or -1; d_kf_out (may be NULL) [p*row_stride + row]: best1 / best2 /
 * idx of that keyframe row's scan, accepted = 1 kept, 2 removed by the rotation check; d_nmatches[p] (may be NULL) = return value.
 * d_n_left (per frame, NULL = all -1) is F.Nleft: on a fisheye rig (>= 0) the frame's rows >= Nleft are the right camera's and compete
 * separately (:330-358); the right-camera best is then assigned too whenever it and the left best are <= TH_LOW (:388-416).
 * th_low = ORBmatcher::TH_LOW (50).  kf_kf != 0 selects ORBmatcher::SearchByBoW(KeyFrame *pKF1, KeyFrame *pKF2, vpMatches12)
 * (src/ORBmatcher.cc:823-963, the NLeft == -1 path): rows of the second keyframe without a map point are no candidates either
 * (d_has_map_point is read for both frames) and the threshold is bestDist1 < TH_LOW; vpMatches12[row1] = vpMapPoints2[d_kf_out[row1].idx]
 * where accepted == 1.  At most 4096 features per frame. */
int orbfe_search_by_bow_device(int device, int npairs, int max_features, const int32_t* d_pair_kf, const int32_t* d_pair_f, const uint8_t* d_desc,
                               const orbfe_keypoint* d_kps, int row_stride, const int32_t* d_fv_nodes, const int32_t* d_fv_starts, const int32_t* d_fv_feats,
                               const int32_t* d_fv_count, const uint8_t* d_has_map_point, const int32_t* d_n_left, int th_low, float nnratio,
                               int check_orientation, int kf_kf, int32_t* d_match_f, orbfe_match* d_kf_out, int32_t* d_nmatches, void* stream);
/* One pair over host memory; the fv_* arrays are those orbfe_bow_result / orbfe_bow_transform return for the two frames
 * (has_map_point_f is read only with kf_kf; NULL = all; n_left_f = F.Nleft, -1 for one camera). */
int orbfe_search_by_bow(int device, const uint8_t* desc_kf, const orbfe_keypoint* kps_kf, int n_kf, const int32_t* fv_nodes_kf, const int32_t* fv_starts_kf,
                        const int32_t* fv_feats_kf, int nfv_kf, const uint8_t* has_map_point, const uint8_t* desc_f, const orbfe_keypoint* kps_f, int n_f,
                        const int32_t* fv_nodes_f, const int32_t* fv_starts_f, const int32_t* fv_feats_f, int nfv_f, const uint8_t* has_map_point_f,
                        int n_left_f, int th_low, float nnratio, int check_orientation, int kf_kf, int32_t* match_f, orbfe_match* kf_out,
                        int* nmatches);

/* Two-camera (fisheye) rigs in SearchForTriangulation (src/ORBmatcher.cc:990-1003, 1102-1132): when both keyframes have an mpCamera2, rows
 * >= NLeft of a keyframe are right-camera keypoints (mvKeysRight), no keypoint counts as stereo, the epipole test is off, and the geometric
 * test is KannalaBrandt8::epipolarConstrain = TriangulateMatches(...) > 0.0001f (src/CameraModels/KannalaBrandt8.cpp:234-237) with the cameras
 * and the relative pose of the two sides: cam = mvParameters of pKF1->mpCamera, pKF1->mpCamera2, pKF2->mpCamera, pKF2->mpCamera2; R / t =
 * Rll, Rlr, Rrl, Rrr and tll, tlr, trl, trr exactly as the reference computes them (row-major).  Pass one per pair (device entry) or one
 * (host entry) in the last argument; NULL = pinhole keyframes, F12 / epipole are then used as before (they are ignored with a rig). */
typedef struct {
    float cam[4][8];
    float R[4][9], t[4][3];
    int32_t n_left_1, n_left_2;
} orbfe_triangulation_rig;
/* Replaces ORBmatcher::SearchForTriangulation(KeyFrame *pKF1, KeyFrame *pKF2, cv::Mat F12, vMatchedPairs, bOnlyStereo, bCoarse)
 * (src/ORBmatcher.cc:965-1206; LocalMapping::CreateNewMapPoints) for pinhole keyframes without a second camera, for batches of
 * keyframe pairs in the strided layout of orbfe_search_by_bow_device: features of keyframe 1 without a map point against the features
 * of keyframe 2 without one, inside the common vocabulary nodes, smallest distance <= th_low (TH_LOW) that keeps away from the epipole
 * (:1084-1092; d_epipole = ep of pair p, 2 floats; d_scale_factors = mvScaleFactors, 16 floats) and satisfies
 * Pinhole::epipolarConstrain (src/CameraModels/Pinhole.cpp:135-157) for d_f12 (F12 of pair p, 9 floats row-major; d_level_sigma2 =
 * pKF2->mvLevelSigma2, 16 floats) or coarse; rotation histogram with check_orientation.  d_u_right (mvuRight, may be NULL) decides
 * bStereo1/2; d_has_map_point as in orbfe_search_by_bow_device (here a row WITH a map point is left out).
 * d_match_12[p*row_stride + i] = vMatches12[i]; d_nmatches[p] = the return value.  At most 65535 pairs per call; a pair's vocabulary
 * nodes are dealt out over up to 16 CTAs, so that the few pairs of one CreateNewMapPoints call still spread over the SMs. */
int orbfe_search_for_triangulation_device(int device, int npairs, int max_features, const int32_t* d_pair_1, const int32_t* d_pair_2, const uint8_t* d_desc,
                                          const orbfe_keypoint* d_kps, const float* d_u_right, int row_stride, const int32_t* d_fv_nodes,
                                          const int32_t* d_fv_starts, const int32_t* d_fv_feats, const int32_t* d_fv_count, const uint8_t* d_has_map_point,
                                          const float* d_f12, const float* d_epipole, const float* d_scale_factors, const float* d_level_sigma2, int th_low,
                                          int only_stereo, int coarse, int check_orientation, int32_t* d_match_12, int32_t* d_nmatches, void* stream,
                                          const orbfe_triangulation_rig* d_rigs);
/* One pair over host memory (scale_factors / level_sigma2: nlevels entries). */
int orbfe_search_for_triangulation(int device, const uint8_t* desc_1, const orbfe_keypoint* kps_1, const float* u_right_1, const uint8_t* has_map_point_1, int n_1,
                                   const int32_t* fv_nodes_1, const int32_t* fv_starts_1, const int32_t* fv_feats_1, int nfv_1, const uint8_t* desc_2,
                                   const orbfe_keypoint* kps_2, const float* u_right_2, const uint8_t* has_map_point_2, int n_2, const int32_t* fv_nodes_2,
                                   const int32_t* fv_starts_2, const int32_t* fv_feats_2, int nfv_2, const float* f12, const float* epipole,
                                   const float* scale_factors, const float* level_sigma2, int nlevels, int th_low, int only_stereo, int coarse,
                                   int check_orientation, int32_t* match_12, int* nmatches, const orbfe_triangulation_rig* rig);

/* Many pairs over host memory in ONE call.  The reference calls these matchers in loops whose iterations do not depend on each other:
 * LocalMapping::CreateNewMapPoints runs SearchForTriangulation of the current keyframe against each of its 10-20 neighbours
 * (src/LocalMapping.cc:556-621), Tracking::Relocalization and LoopClosing run SearchByBoW against every candidate keyframe
 * (src/Tracking.cc:3784, src/LoopClosing.cc:725).  One pair is too little work for a GPU (a call costs about what one CPU thread needs);
 * the loop as one call costs little more than one pair.  frames[] lists every frame or keyframe once, the pairs index into it. */
typedef struct orbfe_host_frame {
    const uint8_t* desc;            /* n x 32 */
    const orbfe_keypoint* kps;      /* n; may be NULL for SearchByBoW without check_orientation */
    const float* u_right;           /* n, or NULL: no stereo keypoints */
    const uint8_t* has_map_point;   /* n, or NULL: every row has one (SearchByBoW) / no row has one (SearchForTriangulation) */
    int32_t n;
    const int32_t* fv_nodes;        /* the FeatureVector: nfv node ids ascending, */
    const int32_t* fv_starts;       /* nfv + 1 offsets into fv_feats, */
    const int32_t* fv_feats;        /* feature rows, in the map's order */
    int32_t nfv;
    int32_t n_left;                 /* F.Nleft for the frame side of SearchByBoW; -1 without a second camera */
} orbfe_host_frame;
/* orbfe_search_for_triangulation for pairs (frames[pair_1[p]], frames[pair_2[p]]), p < npairs: f12 = 9 floats and epipole = 2 floats per pair, rigs =
 * one orbfe_triangulation_rig per pair or NULL.  match_12[p * match_stride + i] = vMatches12[i] of pair p (match_stride >= the n of every first
 * keyframe), nmatches[p] = its return value. */
int orbfe_search_for_triangulation_frames(int device, const orbfe_host_frame* frames, int nframes, const int32_t* pair_1, const int32_t* pair_2, int npairs,
                                          const float* f12, const float* epipole, const float* scale_factors, const float* level_sigma2, int nlevels, int th_low,
                                          int only_stereo, int coarse, int check_orientation, int32_t* match_12, int match_stride, int32_t* nmatches,
                                          const orbfe_triangulation_rig* rigs);
/* orbfe_search_by_bow for pairs (keyframe frames[pair_kf[p]], frame frames[pair_f[p]]): match_f[p * match_stride + i] = the keyframe row matched to
 * row i of the frame or -1 (match_stride >= the n of every frame side), kf_out (may be NULL) = npairs x match_stride records as in
 * orbfe_search_by_bow (match_stride >= the n of every keyframe side too when it is asked for), nmatches[p] = the return value. */
int orbfe_search_by_bow_frames(int device, const orbfe_host_frame* frames, int nframes, const int32_t* pair_kf, const int32_t* pair_f, int npairs, int th_low,
                               float nnratio, int check_orientation, int kf_kf, int32_t* match_f, int match_stride, orbfe_match* kf_out, int32_t* nmatches);

/* Replaces the computation of MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:349-423) for many map points at once
 * (SURVEY.md §8f N4): map point p owns the descriptor rows [offsets[p], offsets[p+1]) (the rows of vDescriptors, :352-384);
 * best_idx[p] = BestIdx, the row (relative to offsets[p]) with the least median distance to all rows of the point,
 * median = sorted distances[int(0.5*(N-1))], first row on ties (:404-417); -1 for a point without rows.  best_median (may be
 * NULL) receives BestMedian.  At most 65535 rows per point. */
int orbfe_distinctive_descriptors_device(int device, const uint8_t* d_desc, const int32_t* d_offsets, int npoints, int32_t* d_best_idx,
                                         int32_t* d_best_median, void* stream);
int orbfe_distinctive_descriptors(int device, const uint8_t* desc, const int32_t* offsets, int npoints, int32_t* best_idx,
                                  int32_t* best_median);

/* ------------------------------------------------------------------------------- projection-gated search (N1) */
/* Replaces the searches of Tracking::SearchLocalPoints (src/Tracking.cc:3430-3524) and Tracking::TrackWithMotionModel:
 *   int ORBmatcher::SearchByProjection(Frame &F, const vector<MapPoint*> &vpMapPoints, const float th, const bool bFarPoints,
 *                                      const float thFarPoints)                                   (src/ORBmatcher.cc:44-214)
 *   int ORBmatcher::SearchByProjection(Frame &CurrentFrame, const Frame &LastFrame, const float th, const bool bMono)
 *                                      (src/ORBmatcher.cc:2193-2419; queries flagged ORBFE_PQ_BEST_ONLY, check_orientation = mbCheckOrientation)
 * including the frame grid they walk: Frame::AssignFeaturesToGrid / PosInGrid (src/Frame.cc:380-411, 711-721; 64 x 48 cells,
 * include/Frame.h:37-38) and Frame::GetFeaturesInArea (src/Frame.cc:643-709).  The result is the reference's, including its
 * serial rule that a matched map point occupies its keypoint for the map points after it. */
typedef struct orbfe_frame_grid {   /* Frame::mnMinX, mnMinY, mfGridElementWidthInv, mfGridElementHeightInv */
    float min_x, min_y, inv_w, inv_h;
} orbfe_frame_grid;

#define ORBFE_PQ_RIGHT 1       /* search the right camera's keypoints (bRight; rows >= n_left) */
#define ORBFE_PQ_BLOCKS 2      /* pMP->Observations() > 0: once matched, later map points skip the keypoint (:91-93) */
#define ORBFE_PQ_SKIP 4        /* not in view / bad / too far (:52-59): no search */
#define ORBFE_PQ_AFTER_LEFT 8  /* right-camera half (:145-211) of the map point of the query before: not run when that one
                                  failed the ratio test (the `continue` at :117-118 leaves the whole iteration) */
#define ORBFE_PQ_BEST_ONLY 16  /* frame-to-frame rule (:2267-2300): accepted iff bestDist <= TH_HIGH, no second-best test */
#define ORBFE_PQ_ERROR_GATE 64  /* ORBmatcher::Fuse (:1643-1841): a candidate is skipped when its reprojection error fails the chi-square
                                  test — (ex^2+ey^2+er^2)*mvInvLevelSigma2[octave] > 7.8 if mvuRight[row] >= 0 (er = xr - mvuRight[row]), else
                                  (ex^2+ey^2)*mvInvLevelSigma2[octave] > 5.99 (:1714-1740) — instead of the |xr - mvuRight| gate */
#define ORBFE_PQ_NEEDS_LEFT_WINDOW 32 /* right-camera half (:2325-2386) of the query before: not run when that one's window holds
                                  no keypoint (`if(vIndices2.empty()) continue;` at :2262-2263) */
typedef struct orbfe_proj_query {   /* one (map point, camera) search, in the reference's loop order */
    float x, y;        /* mTrackProjX, mTrackProjY (mTrackProjXR, mTrackProjYR with ORBFE_PQ_RIGHT) */
    float radius;      /* RadiusByViewingCos(mTrackViewCos) [* th] * F.mvScaleFactors[nPredictedLevel] (:67-73) */
    float xr;          /* mTrackProjXR, for the |xr - mvuRight| gate (:95-100); read only when u_right is given */
    int32_t min_level; /* nPredictedLevel - 1        (frame-to-frame: nLastOctave / 0 / nLastOctave-1, :2253-2258) */
    int32_t max_level; /* nPredictedLevel; -1 = open (frame-to-frame: -1 / nLastOctave / nLastOctave+1) */
    int32_t flags;     /* ORBFE_PQ_* */
    float angle;       /* kpLF.angle (:2314), read only with check_orientation */
} orbfe_proj_query;

#define ORBFE_PS_NO_CANDIDATES 0   /* every keypoint of the window is occupied or gated out by mvuRight */
#define ORBFE_PS_ACCEPTED 1
#define ORBFE_PS_OVER_THRESHOLD 2  /* bestDist > TH_HIGH */
#define ORBFE_PS_RATIO_REJECTED 3  /* bestLevel == bestLevel2 && bestDist > mfNNratio * bestDist2 */
#define ORBFE_PS_SKIPPED 4         /* ORBFE_PQ_SKIP */
#define ORBFE_PS_CAPACITY 5        /* the frame holds more keypoints than promised: nothing was searched */
#define ORBFE_PS_DROPPED 6         /* ORBFE_PQ_AFTER_LEFT / ORBFE_PQ_NEEDS_LEFT_WINDOW and the left half ended the iteration */
#define ORBFE_PS_ROTATION 7        /* matched, then removed by the rotation-consistency histogram (:2391-2412) */
#define ORBFE_PS_EMPTY_WINDOW 8    /* GetFeaturesInArea returned nothing */
typedef struct orbfe_proj_match {
    int32_t best_idx;      /* bestIdx as a row of the frame's keypoint array (right camera: local index + n_left); -1 if none */
    int32_t best_dist, best_level, second_dist, second_level;   /* bestDist, bestLevel, bestDist2, bestLevel2 (256 / -1 when unset) */
    int32_t accepted;      /* 1: F.mvpMapPoints[best_idx] = pMP (and its stereo partner row, if any) */
    int32_t status;        /* ORBFE_PS_* */
} orbfe_proj_match;

/* A batch of frames, everything on the device, asynchronous on `stream`, no allocation.
 *   d_kps / d_desc: the frames' keypoints (pt and octave are read: mvKeysUn for a rectified rig, mvKeys followed by mvKeysRight
 *     for a fisheye rig) and 32-byte descriptors, frame f = rows [d_kp_offsets[f], d_kp_offsets[f+1]); at most max_keypoints
 *     (<= 4096) rows per frame, a frame with more gets ORBFE_PS_CAPACITY in every result (the CTA's shared memory is sized by
 *     max_keypoints: pass what the frames really hold, not the limit); at most 65534 queries per frame;
 *   d_n_left: per frame F.Nleft (-1: one grid), or NULL for all -1;   d_u_right: mvuRight per row, or NULL;
 *   d_partner: per row the row matched in the other camera (mvLeftToRightMatch[i] + Nleft / mvRightToLeftMatch[i - Nleft]) or
 *     -1, or NULL;   d_taken_in: per row, nonzero iff F.mvpMapPoints[row] has observations (NULL: none);
 *   d_queries / d_qdesc (pMP->GetDescriptor()): frame f = entries [d_q_offsets[f], d_q_offsets[f+1]);
 *   th_high = ORBmatcher::TH_HIGH (100), nnratio = mfNNratio;   d_taken_out (may be NULL or d_taken_in): occupancy afterwards;
 *   d_inv_level_sigma2: mvInvLevelSigma2 (16 floats, one per pyramid level), read for ORBFE_PQ_ERROR_GATE queries (else may be NULL).
 *     ORBmatcher::Fuse(pKF, vpMapPoints, th, bRight) is: ORBFE_PQ_BEST_ONLY | ORBFE_PQ_ERROR_GATE queries without ORBFE_PQ_BLOCKS,
 *     th_high = TH_LOW, no d_taken_in; best_idx of an accepted query is the keypoint to fuse with or to add the observation to;
 *   check_orientation != 0: after the searches, the rotation histogram of (query angle - keypoint angle) over the frame's matches
 *     (30 bins, ComputeThreeMaxima, src/ORBmatcher.cc:2545-2586) removes the matches outside the three fullest bins. */
int orbfe_search_by_projection_device(int device, int nframes, int max_keypoints, const orbfe_keypoint* d_kps, const uint8_t* d_desc,
                                      const int32_t* d_kp_offsets, const int32_t* d_n_left, const float* d_u_right, const int32_t* d_partner,
                                      const uint8_t* d_taken_in, const orbfe_frame_grid* d_grids, const orbfe_proj_query* d_queries,
                                      const uint8_t* d_qdesc, const int32_t* d_q_offsets, int th_high, float nnratio, int check_orientation,
                                      const float* d_inv_level_sigma2, orbfe_proj_match* d_out, uint8_t* d_taken_out, void* stream);
/* Diagnostics (synchronises the device): maximum and sum over the frames of the speculative rounds the projection kernel needed since the last call. */
int orbfe_debug_projection_rounds(int* max_rounds, int* sum_rounds);
/* One frame over host memory (copies inside, synchronous).  taken is updated in place (may be NULL); *nmatches receives the
 * reference's return value (an accepted row with a stereo partner counts twice). */
int orbfe_search_by_projection(int device, const orbfe_keypoint* kps, const uint8_t* desc, int n, int n_left, const float* u_right,
                               const int32_t* partner, uint8_t* taken, const orbfe_frame_grid* grid, const orbfe_proj_query* queries,
                               const uint8_t* qdesc, int nq, int th_high, float nnratio, int check_orientation,
                               const float* inv_level_sigma2, orbfe_proj_match* out, int* nmatches);

/* Replaces int ORBmatcher::SearchForInitialization(Frame &F1, Frame &F2, vector<cv::Point2f> &vbPrevMatched, vector<int> &vnMatches12,
 * int windowSize) (src/ORBmatcher.cc:706-821; caller Tracking::MonocularInitialization, src/Tracking.cc:2462, with windowSize 100 and
 * ORBmatcher(0.9, true)) for a batch of frame pairs on the device.  kps1 / kps2 are F1.mvKeysUn / F2.mvKeysUn (pt, octave >= 0 and angle
 * are read), pair p = rows [d_off1[p], d_off1[p+1]) and [d_off2[p], d_off2[p+1]), at most 8192 rows per frame (a pair with more gets
 * nmatches = -1 and is not searched); d_grids[p] is F2's grid geometry; d_prev_matched [rows of F1][2] is vbPrevMatched, read and
 * updated (:816-818); d_matches12 [rows of F1] receives vnMatches12, d_nmatches[p] the return value; d_rounds (may be NULL) the number of
 * rounds the kernel needed.  th_low = ORBmatcher::TH_LOW (50).  The result is the reference's, including the order dependence of
 * vMatchedDistance (:745) and of the displaced matches (:764-768).  Asynchronous on `stream`, no allocation. */
int orbfe_search_for_initialization_device(int device, int npairs, const orbfe_keypoint* d_kps1, const uint8_t* d_desc1, const int32_t* d_off1,
                                           const orbfe_keypoint* d_kps2, const uint8_t* d_desc2, const int32_t* d_off2, const orbfe_frame_grid* d_grids,
                                           float* d_prev_matched, int window, float nnratio, int check_orientation, int th_low, int32_t* d_matches12,
                                           int32_t* d_nmatches, int32_t* d_rounds, void* stream);
/* One pair over host memory (copies inside, synchronous). */
int orbfe_search_for_initialization(int device, const orbfe_keypoint* kps1, const uint8_t* desc1, int n1, const orbfe_keypoint* kps2, const uint8_t* desc2, int n2,
                                    const orbfe_frame_grid* grid, float* prev_matched, int window, float nnratio, int check_orientation, int th_low,
                                    int32_t* matches12, int* nmatches);

/* Host-only self check (no device needed): rBRIEF's rotation uses cosf / sinf of the host libm in the reference
 * (src/ORBextractor.cc:110-111); the device evaluates them in double and corrects the 96 binary32 angles where that moves a tap with a
 * table of glibc 2.39 (x86-64) values.  Returns how many of the table's entries this host's libm does NOT reproduce (0: descriptors equal
 * this host's own CPU build for every possible angle); *listed (may be NULL) receives the table size. */
int orbfe_check_host_libm(int* listed);

/* ------------------------------------------------------------------------------------------ input conditioning */
/* The steps the example drivers and Tracking run on each image before the extractor, so that an
 * uploaded frame can stay on the GPU.  Bit-exact with OpenCV 4.13's 8-bit fixed-point paths. */
typedef struct orbfe_rectifier orbfe_rectifier;

/* Replaces cv::remap(im, imRect, M1, M2, cv::INTER_LINEAR) (Examples/Stereo-Inertial/stereo_inertial_euroc.cc:201-202; also the
 * other stereo drivers) for CV_32FC1 maps as produced by cv::initUndistortRectifyMap(..., CV_32F, M1, M2) (:150-151).
 * map_x / map_y: width*height floats on the host, converted once to OpenCV's 5-bit fixed point.  Border: constant 0. */
int orbfe_rectifier_create(int device, int width, int height, const float* map_x, const float* map_y, orbfe_rectifier** out);
void orbfe_rectifier_destroy(orbfe_rectifier* r);
/* nframes dense width*height u8 frames, device to device (may feed orbfe_extract_batch_device); asynchronous on `stream`. */
int orbfe_rectify_batch_device(orbfe_rectifier* r, const uint8_t* d_src, uint8_t* d_dst, int nframes, void* stream);
/* one host image in, one host image out (steps in bytes); synchronous */
int orbfe_rectify(orbfe_rectifier* r, const uint8_t* src, size_t sstep, uint8_t* dst, size_t dstep);

/* Replaces cvtColor(im, gray, cv::COLOR_RGB2GRAY / BGR2GRAY / RGBA2GRAY / BGRA2GRAY) of Tracking::GrabImageStereo / RGBD /
 * Monocular (src/Tracking.cc:1303-1328, 1360-1372, 1399-1411).  channels = 3 or 4 (interleaved), rgb_order != 0 when the first
 * channel is red (mbRGB).  Device to device, asynchronous on `stream` (NULL = default stream). */
int orbfe_to_gray_device(int device, const uint8_t* d_src, int channels, int rgb_order, int width, int height, int nframes,
                         uint8_t* d_dst, void* stream);

/* Replaces cv::createCLAHE(clip_limit, cv::Size(tiles_x, tiles_y)) + clahe->apply(im, im) of the TUM-VI drivers
 * (Examples/Stereo-Inertial/stereo_inertial_tum_vi.cc:128 and :137-138; also Examples/Stereo/stereo_tum_vi.cc,
 * Examples/Monocular/mono_tum_vi.cc, Examples/Monocular-Inertial/mono_inertial_tum_vi.cc: clip 3.0, 8 x 8 tiles).  8-bit single
 * channel; sizes that do not divide by the grid follow OpenCV (REFLECT_101 extension for the histograms).
 * ORBFE_E_INVALID when the image is smaller than that extension.  max_batch sizes the per-frame lut storage. */
typedef struct orbfe_clahe orbfe_clahe;
int orbfe_clahe_create(int device, int width, int height, double clip_limit, int tiles_x, int tiles_y, int max_batch, orbfe_clahe** out);
void orbfe_clahe_destroy(orbfe_clahe* c);
/* nframes dense width*height u8 frames, device to device; d_dst may equal d_src (the reference applies in place).  Asynchronous on
 * `stream` (NULL = the object's stream).  ORBFE_E_CAPACITY when nframes > max_batch. */
int orbfe_clahe_apply_batch_device(orbfe_clahe* c, const uint8_t* d_src, uint8_t* d_dst, int nframes, void* stream);
/* one host image in, one host image out (steps in bytes; dst may be src); synchronous */
int orbfe_clahe_apply(orbfe_clahe* c, const uint8_t* src, size_t sstep, uint8_t* dst, size_t dstep);
/* stage test hook: the [tiles_y][tiles_x][256] lookup tables of frame `frame` of the last batch, to the host */
int orbfe_clahe_lut(orbfe_clahe* c, int frame, uint8_t* out);

#ifdef __cplusplus
}
#endif
#endif /* ORBFE_H */
