/* orbfe_internal.h — shared declarations of liborbfe.so (not installed; the public surface is include/orbfe.h). */
#ifndef ORBFE_INTERNAL_H
#define ORBFE_INTERNAL_H
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <string>
#include <vector>

#include "../../include/orbfe.h"

#define ORBFE_MAX_LEVELS 16
#define ORBFE_MAX_DEVICES 64  /* per-device bookkeeping tables (kernel attributes, static scratch addresses) */
#define ORBFE_EDGE 19        /* EDGE_THRESHOLD, src/ORBextractor.cc:72 */
#define ORBFE_MIN_BORDER 16  /* EDGE_THRESHOLD-3, src/ORBextractor.cc:771 */
#define ORBFE_NUM_STAGES 6   /* pyramid, fast, blur, quadtree, assemble, orient+describe (+2 stereo) */
#define ORBFE_HOST_CHUNK 64   /* frames per stage of orbfe_extract_batch_host's copy/compute pipeline */
#define ORBFE_ROW_LEAD 32    /* byte of a padded row where interior x = 0 sits (16-byte aligned) */

/* Geometry of one pyramid level; identical for every frame of a batch. */
struct LevelDesc {
    int w, h;        /* interior size */
    int pitch;       /* bytes per padded row; padded x=-19 is byte ORBFE_ROW_LEAD-19 */
    int pyrOff;      /* byte offset of padded row y=-19 inside a frame's pyramid slab */
    int bpitch;      /* bytes per row of the blurred level (unpadded) */
    int blurOff;     /* byte offset inside a frame's blur slab */
    int nCols, nRows, wCell, hCell; /* FAST cell grid, src/ORBextractor.cc:782-785 */
    int cellBase;    /* first cell of this level in the per-frame cell table */
    int keyBase;     /* first candidate slot of this level in the per-frame candidate array */
    int keyCap;      /* candidate slots of this level */
    int quota;       /* mnFeaturesPerLevel[level] */
    int nIni;        /* initial quadtree nodes, src/ORBextractor.cc:541 */
    float hX;        /* src/ORBextractor.cc:543 */
    int selBase;     /* first selected-keypoint slot of this level in the per-frame arrays */
    int selCap;      /* max keypoints the quadtree can return for this level */
    int tabX, tabY;  /* resize coefficient tables (uint2 entries) for producing this level from level-1 */
    float scale;     /* mvScaleFactor[level] */
    float kpSize;    /* (float)(int)(31*scale), src/ORBextractor.cc:862 */
    unsigned invW, invH; /* ceil(2^24 / wCell), ceil(2^24 / hCell): cell of a candidate pixel without a division */
    int rawBase;     /* first raw-corner slot of this level (per frame) */
    int rawStride;   /* raw-corner slots per cell = wCell * hCell */
};

struct ExDesc {
    int nlevels;
    int W, H;
    int nCells;       /* cells per frame, all levels */
    int candCap;      /* candidate slots per frame, all levels */
    int kpCap;        /* selected keypoints per frame, all levels (= sum selCap) */
    int nodeCap;      /* quadtree node capacity (max over levels of selCap, padded) */
    int qtScratchBytes; /* 0: the node arrays of a level fit in shared memory; else bytes per (frame, level) of b.qtScratch */
    int iniTh, minTh;
    int blurTile0[ORBFE_MAX_LEVELS + 1]; /* first blur tile of each level in the one blur launch; [nlevels] = total */
    int stBands;      /* 8-row bands of the stereo candidate table: (H >> 3) + 1 */
    int nStrips;      /* strips of k_fast_stream per frame (32-row cut) */
    int nStripsShort; /* 16-row cut, stored after the first in b.strips; used for launches of a few frames */
    int rawCap;       /* raw-corner slots per frame */
    int fastRows;     /* tallest FAST cell ROI (rows of the smem tile) */
    int fastPix;      /* most candidate pixels in one cell, rounded up to a multiple of 8 */
    size_t pyrSlab;   /* bytes per frame */
    size_t blurSlab;
    LevelDesc lv[ORBFE_MAX_LEVELS];
};

/* One FAST cell (src/ORBextractor.cc:787-854): ROI origin in level coordinates and clipped size. */
struct CellDesc {
    short x0, y0;     /* iniX, iniY */
    short rw, rh;     /* maxX-iniX, maxY-iniY (0 when the reference skips the cell) */
    short level;
    short xoff, yoff; /* j*wCell, i*hCell added to the keypoint coordinates */
    short pad;
    int candBase;     /* first candidate slot of this cell (per frame) */
    int candCap;
    /* derived constants of the FAST kernel (host-computed so the kernel does no integer division) */
    unsigned invcw;   /* ceil(2^24 / cw): p / cw == (p * invcw) >> 24 for p < 4900 */
    unsigned invgw;   /* same for ngw */
    short ngw;        /* tile words per row that overlap the candidate columns */
    short nw;         /* tile words per row to load */
    unsigned edgeL, edgeR; /* byte-flag masks of the first / last of those words */
};

/* One warp-sized piece of a level for k_fast_stream: candidate columns [xs, xs+120), rows [y0, y0+nrows). */
struct StripDesc {
    short level, xs, y0, nrows;
};

/* Per-frame device buffers, all sized for max_batch at creation. */
struct ExBuffers {
    uint8_t* pyr;        /* [B][pyrSlab] */
    uint8_t* blur;       /* [B][blurSlab] */
    uint32_t* cand;      /* [B][candCap]  packed x | y<<12 | score<<24, coordinates relative to minBorder */
    int* cellCount;      /* [B][nCells] */
    uint32_t* raw;       /* [B][rawCap]   corners at iniTh before NMS, per cell at rawBase + cell*rawStride */
    int* rawCount;       /* [B][nCells]   zero between batches (k_fast_nms resets what it consumes) */
    int* fbList;         /* [B*nCells]    cells that came out empty at iniTh */
    int* fbCount;        /* [1] */
    const StripDesc* strips;
    const uint32_t* blurTiles; /* [blurTile0[nlevels]] level << 24 | tile row << 12 | tile column, in launch order */
    uint32_t* keys;      /* [B][candCap]  quadtree scratch: a level's candidates gathered in cell order */
    uint16_t* keyNode;   /* [B][candCap]  quadtree scratch: node of each candidate */
    uint32_t* sel;       /* [B][kpCap]    quadtree output, packed like cand, per level at selBase */
    int* selCount;       /* [B][nlevels] */
    uint8_t* qtScratch;  /* [B][nlevels][qtScratchBytes] node arrays of the quadtree when they do not fit on chip, else NULL */
    uint2* work;         /* [B][kpCap]    (x | y<<16 in level coords, level | slot<<8) in level order */
    orbfe_keypoint* kps; /* [B][kpCap] */
    uint8_t* desc;       /* [B][kpCap][32] */
    int* count;          /* [B] */
    int* mono;           /* [B] */
    float* uRight;       /* [B][kpCap] */
    float* depth;        /* [B][kpCap] */
    int* sadBest;        /* [B][kpCap] stereo scratch: best SAD per left keypoint, -1 = no match */
    uint4* stTable;      /* [B][kpCap] stereo scratch: this extractor's keypoints bucketed by 8-row band (x, band rows | octave, index) */
    int* stBand;         /* [B][stBands+1] first table entry of each band */
    const CellDesc* cells;
    const uint2* tabs;   /* resize tables */
    const uint8_t* input; /* [B][H*W] staging for host-image entry points */
};

struct orbfe_extractor {
    int device;
    int maxBatch;
    int lastBatch;
    orbfe_params params;
    ExDesc d;
    ExBuffers b;
    uint8_t* inputStage;  /* device copy of host images */
    cudaStream_t stream;
    cudaStream_t copyIn, copyOut;          /* host-batch pipeline: uploads and downloads overlap the kernels */
    cudaStream_t side;                     /* the blur runs here, beside quadtree + assemble (they only meet again in k_orient_describe) */
    cudaEvent_t evFork, evJoin;
    int forkBlur;                          /* 0 with ORBFE_NO_FORK=1 */
    int32_t *fishLR, *fishRL;              /* fisheye host pipeline: mvLeftToRightMatch / mvRightToLeftMatch per frame, allocated on first use */
    float* fishP3d;                        /* mvStereo3Dpoints */
    std::vector<cudaEvent_t> evIn, evDone; /* one pair per chunk of hostChunk frames */
    std::vector<cudaEvent_t> evOut;        /* per chunk: the download of its results has finished (next batch may rewrite them) */
    std::vector<char> evOutValid;
    cudaEvent_t evLast;                    /* recorded behind the last work enqueued for this object, on whichever stream took it */
    cudaEvent_t evBatch[2];                /* per submit slot: everything of that batch is on the host */
    uint8_t* inputStage2;                  /* second image staging buffer (slot 1 of orbfe_stereo_frames_submit), allocated on first use */
    int pendingHead, pendingCount, pendingFrames[2];
    int hostChunk;                         /* frames per pipeline stage (ORBFE_HOST_CHUNK, or env ORBFE_HOST_CHUNK) */
    std::vector<float> scale, invScale, sigma2, invSigma2;
    std::vector<int> quota, umax;
    std::vector<void*> allocs;
    int fastSmem, quadSmem;
    CUtensorMap* blurMaps;   /* device array, one 3-D map (x bytes, padded rows, frames) per level over b.pyr, for k_blur7_tma */
    int blurLdg;             /* ORBFE_BLUR_LDG=1: the pre-TMA blur kernel (A/B measurements) */
    /* optional per-stage timing (orbfe_set_profiling): events bracket each stage of the last launch */
    int profiling;
    cudaEvent_t ev[ORBFE_NUM_STAGES + 3];
    double stageMs[ORBFE_NUM_STAGES + 2];
    int stagePending, stereoPending;
    long long launches;   /* kernels launched so far by this object */
    int batchFrames;      /* frames of the last batch (chunks of the host pipeline accumulate) */
    int bordersValid;     /* the 19 px pyramid border of those frames has been written (orbfe_materialize_borders) */
    cudaGraphExec_t graphExec;   /* the single-frame path (orbfe_extract) as a graph, captured per lapping interval */
    int graphLap0, graphLap1, graphLaunches, singleGraph;
    /* pinned host staging for the host entry points */
    orbfe_keypoint* hKps;
    uint8_t* hDesc;
    int* hCount;
    int* hMono;
};

/* views of the per-frame buffers that start at frame f0 (kernels index frames from blockIdx) */
ExBuffers orbfe_shift_frames(const ExBuffers& b, const ExDesc& d, int f0);
/* Frame::ComputeStereoMatches kernels for frames [frame0, frame0 + nframes) of two extractors, enqueued on st */
int orbfe_launch_fisheye(orbfe_extractor* exL, orbfe_extractor* exR, int nframes, int frame0, double ratio, const orbfe_fisheye_rig* rig, int32_t* lr, int32_t* rl,
                         float* depth, float* p3d, cudaStream_t st);
int orbfe_launch_stereo(orbfe_extractor* exL, orbfe_extractor* exR, int nframes, int frame0, float mbf, float minD, float maxD, cudaStream_t st);

/* Getters wait for the object's own last work (not for the whole device: Tracking, LocalMapping and LoopClosing share the GPU)
 * and copy on the object's stream. */
int orbfe_mark_last(orbfe_extractor* ex, cudaStream_t st);
int orbfe_wait_last(orbfe_extractor* ex);
#define ORBFE_D2H(ex, dst, src, bytes) ORBFE_CUDA(cudaMemcpyAsync((dst), (src), (bytes), cudaMemcpyDeviceToHost, (ex)->stream))
#define ORBFE_D2H_2D(ex, dst, dpitch, src, spitch, w, h) ORBFE_CUDA(cudaMemcpy2DAsync((dst), (dpitch), (src), (spitch), (w), (h), cudaMemcpyDeviceToHost, (ex)->stream))
#define ORBFE_D2H_DONE(ex) ORBFE_CUDA(cudaStreamSynchronize((ex)->stream))

void orbfe_set_error(const char* fmt, ...);
int orbfe_cuda_fail(cudaError_t e, const char* what);
#define ORBFE_CUDA(call)                                          \
    do {                                                          \
        cudaError_t e__ = (call);                                 \
        if (e__ != cudaSuccess) return orbfe_cuda_fail(e__, #call); \
    } while (0)

/* The calling thread's staging block on a device: pinned host memory and device memory of the same size, and a stream.  It grows when a
 * call needs more and stays for the life of the thread (or until orbfe_thread_release), so that a one-frame host entry point — the
 * matchers behind the shims, called once per frame or keyframe pair — costs one copy up, its kernels and one copy down rather than an
 * allocation, a stream and a dozen pageable copies per call. */
struct OrbfeStaging { uint8_t* d; uint8_t* h; cudaStream_t stream; };
#define ORBFE_STAGING_MAX ((size_t)32 << 20)   /* entry points whose size is open-ended stage through it up to here, beyond it they allocate per call */
int orbfe_thread_staging(int device, size_t bytes, OrbfeStaging* out);

#endif
