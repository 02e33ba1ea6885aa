/* orbfe_host.cu — host side of the extractor part of the C ABI (include/orbfe.h): geometry tables that
 * ORBextractor::ORBextractor and the per-frame prologues of the reference compute on the CPU
 * (/root/reference/src/ORBextractor.cc:408-468, 537-545, 763-806, 1152-1160), buffer ownership, copies. */
#include <math.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <new>
#include <vector>

#include "orbfe_internal.h"

int orbfe_upload_constants(const int* umax16, const int8_t* pattern1024);
int orbfe_quadtree_smem_bytes(int cap);
int orbfe_configure_kernels(int quadSmem, int fastSmem);
int orbfe_blur_tiles(int w, int h);
int orbfe_launch_extract(orbfe_extractor* ex, const uint8_t* d_images, int nframes, int lap0, int lap1, cudaStream_t st, int frame0);
int orbfe_collect_stage_times(orbfe_extractor* ex);
int orbfe_materialize_borders(orbfe_extractor* ex, cudaStream_t st);

static thread_local char g_err[512] = "";

void orbfe_set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof g_err, fmt, ap);
    va_end(ap);
}

int orbfe_cuda_fail(cudaError_t e, const char* what) {
    orbfe_set_error("CUDA error %d (%s) in %s", (int)e, cudaGetErrorString(e), what);
    return ORBFE_E_CUDA;
}

namespace {
struct StagingSlot { int device; uint8_t* d; uint8_t* h; size_t cap; cudaStream_t stream; };
struct ThreadStaging {
    std::vector<StagingSlot> slots;
    void release() {
        for (StagingSlot& s : slots) {
            if (cudaSetDevice(s.device) != cudaSuccess) continue;
            if (s.stream) { cudaStreamSynchronize(s.stream); cudaStreamDestroy(s.stream); }
            cudaFree(s.d);
            cudaFreeHost(s.h);
        }
        slots.clear();
    }
    ~ThreadStaging() { release(); }
};
thread_local ThreadStaging t_staging;
}  // namespace

int orbfe_thread_staging(int device, size_t bytes, OrbfeStaging* out) {
    ORBFE_CUDA(cudaSetDevice(device));
    StagingSlot* s = nullptr;
    for (StagingSlot& x : t_staging.slots)
        if (x.device == device) s = &x;
    if (!s) {
        cudaStream_t st = nullptr;
        ORBFE_CUDA(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
        t_staging.slots.push_back(StagingSlot{device, nullptr, nullptr, 0, st});
        s = &t_staging.slots.back();
    }
    if (s->cap < bytes) {
        ORBFE_CUDA(cudaStreamSynchronize(s->stream));
        cudaFree(s->d); cudaFreeHost(s->h);
        s->d = s->h = nullptr; s->cap = 0;
        const size_t want = std::max(bytes + bytes / 2, (size_t)1 << 20);
        void* d = nullptr; void* h = nullptr;
        ORBFE_CUDA(cudaMalloc(&d, want));
        const cudaError_t e = cudaMallocHost(&h, want);
        if (e != cudaSuccess) { cudaFree(d); return orbfe_cuda_fail(e, "cudaMallocHost (thread staging)"); }
        s->d = (uint8_t*)d; s->h = (uint8_t*)h; s->cap = want;
    }
    out->d = s->d; out->h = s->h; out->stream = s->stream;
    return ORBFE_OK;
}

static const int8_t kPattern[1024] = {
#include "../../include/orbfe_pattern31.inc"
};

static inline int cv_round_f(float v) { return (int)lrintf(v); }
static inline int cv_round_d(double v) { return (int)lrint(v); }
static inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

/* cv::resize INTER_LINEAR coefficient tables: entry = (source index, c0 | c1 << 16). */
static void linear_table(int ssize, int dsize, std::vector<uint2>& out) {
    const double scale = 1.0 / ((double)dsize / ssize);
    for (int dd = 0; dd < dsize; dd++) {
        float f = (float)((dd + 0.5) * scale - 0.5);
        int s = (int)floorf(f);
        f -= s;
        if (s < 0) { s = 0; f = 0.f; }
        if (s >= ssize - 1) { s = ssize - 1; f = 0.f; }
        const int c0 = cv_round_f((1.f - f) * 2048.f), c1 = cv_round_f(f * 2048.f);
        out.push_back(make_uint2((unsigned)s, (unsigned)c0 | ((unsigned)c1 << 16)));
    }
}

/* cuTensorMapEncodeTiled through the runtime's driver entry point lookup (the library links no libcuda itself). */
typedef CUresult (*orbfe_encode_tiled_fn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                          const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static int encode_tiled_u8_3d(CUtensorMap* out, void* base, uint64_t dim0, uint64_t dim1, uint64_t dim2, uint64_t stride1, uint64_t stride2, uint32_t box0,
                              uint32_t box1) {
    static orbfe_encode_tiled_fn fn = nullptr;
    if (!fn) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        ORBFE_CUDA(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q));
        if (!p || q != cudaDriverEntryPointSuccess) { orbfe_set_error("the CUDA driver has no cuTensorMapEncodeTiled"); return ORBFE_E_CUDA; }
        fn = (orbfe_encode_tiled_fn)p;
    }
    const cuuint64_t dims[3] = {dim0, dim1, dim2}, strides[2] = {stride1, stride2};
    const cuuint32_t box[3] = {box0, box1, 1}, estr[3] = {1, 1, 1};
    const CUresult r = fn(out, CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, base, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                          CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { orbfe_set_error("cuTensorMapEncodeTiled failed (%d)", (int)r); return ORBFE_E_CUDA; }
    return 0;
}

int orbfe_mark_last(orbfe_extractor* ex, cudaStream_t st) {
    ORBFE_CUDA(cudaEventRecord(ex->evLast, st));
    return 0;
}
int orbfe_wait_last(orbfe_extractor* ex) {
    ORBFE_CUDA(cudaEventSynchronize(ex->evLast));
    return 0;
}

template <typename T>
static int dev_alloc(orbfe_extractor* ex, T** p, size_t count) {
    void* v = nullptr;
    ORBFE_CUDA(cudaMalloc(&v, std::max<size_t>(count, 1) * sizeof(T)));
    ex->allocs.push_back(v);
    *p = (T*)v;
    return 0;
}

extern "C" {

const char* orbfe_last_error(void) { return g_err; }

void orbfe_thread_release(void) { t_staging.release(); }
const char* orbfe_version(void) { return "orbfe 0.1 (sm_100a)"; }

int orbfe_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) return 0;
    return n;
}

void orbfe_extractor_destroy(orbfe_extractor* ex) {
    if (!ex) return;
    cudaSetDevice(ex->device);
    if (ex->stream) cudaStreamSynchronize(ex->stream);
    if (ex->graphExec) cudaGraphExecDestroy(ex->graphExec);
    for (void* p : ex->allocs) cudaFree(p);
    if (ex->hKps) cudaFreeHost(ex->hKps);
    if (ex->hDesc) cudaFreeHost(ex->hDesc);
    if (ex->hCount) cudaFreeHost(ex->hCount);
    if (ex->hMono) cudaFreeHost(ex->hMono);
    for (cudaEvent_t ev : ex->ev) if (ev) cudaEventDestroy(ev);
    for (cudaEvent_t ev : ex->evIn) if (ev) cudaEventDestroy(ev);
    for (cudaEvent_t ev : ex->evDone) if (ev) cudaEventDestroy(ev);
    for (cudaEvent_t ev : ex->evOut) if (ev) cudaEventDestroy(ev);
    for (cudaEvent_t ev : ex->evBatch) if (ev) cudaEventDestroy(ev);
    if (ex->evLast) cudaEventDestroy(ex->evLast);
    if (ex->evFork) cudaEventDestroy(ex->evFork);
    if (ex->evJoin) cudaEventDestroy(ex->evJoin);
    if (ex->side) cudaStreamDestroy(ex->side);
    if (ex->copyIn) cudaStreamDestroy(ex->copyIn);
    if (ex->copyOut) cudaStreamDestroy(ex->copyOut);
    if (ex->stream) cudaStreamDestroy(ex->stream);
    delete ex;
}

int orbfe_extractor_create(const orbfe_params* p, int device, int width, int height, int max_batch, orbfe_extractor** out) {
    if (!p || !out) { orbfe_set_error("null argument"); return ORBFE_E_INVALID; }
    *out = nullptr;
    if (p->nlevels < 1 || p->nlevels > ORBFE_MAX_LEVELS || p->nfeatures < 1 || !(p->scale_factor > 1.0f) || max_batch < 1 ||
        p->ini_th_fast < 1 || p->min_th_fast < 1 || p->ini_th_fast > 127 || p->min_th_fast > 127) {
        orbfe_set_error("unsupported extractor parameters");
        return ORBFE_E_INVALID;
    }
    if (width < 1 || height < 1 || width > 4000 || height > 4000) {
        orbfe_set_error("unsupported image size %dx%d", width, height);
        return ORBFE_E_INVALID;
    }
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || device < 0 || device >= ndev) {
        orbfe_set_error("no usable CUDA device %d (found %d); this library has no CPU path", device, ndev);
        return ORBFE_E_CUDA;
    }
    ORBFE_CUDA(cudaSetDevice(device));

    orbfe_extractor* ex = new (std::nothrow) orbfe_extractor();
    if (!ex) return ORBFE_E_INVALID;
    ex->device = device;
    ex->maxBatch = max_batch;
    ex->lastBatch = 0;
    ex->params = *p;
    ex->stream = nullptr; ex->copyIn = nullptr; ex->copyOut = nullptr; ex->side = nullptr; ex->evFork = nullptr; ex->evJoin = nullptr; ex->forkBlur = 1; ex->fishLR = ex->fishRL = nullptr; ex->fishP3d = nullptr;
    ex->hKps = nullptr; ex->hDesc = nullptr; ex->hCount = nullptr; ex->hMono = nullptr;
    ex->profiling = 0; ex->stagePending = 0; ex->stereoPending = 0; ex->launches = 0; ex->batchFrames = 0; ex->bordersValid = 0;
    ex->graphExec = nullptr; ex->graphLap0 = ex->graphLap1 = 0; ex->graphLaunches = 0; ex->singleGraph = 1;
    if (const char* e = getenv("ORBFE_SINGLE_FRAME_GRAPH")) ex->singleGraph = atoi(e) != 0;
    memset(ex->ev, 0, sizeof ex->ev);
    memset(ex->stageMs, 0, sizeof ex->stageMs);
    memset(&ex->d, 0, sizeof ex->d);
    memset(&ex->b, 0, sizeof ex->b);
    const int nl = p->nlevels;

    /* ORBextractor.cc:413-444 — the member scaleFactor is a double initialised from the float argument */
    const double sf = (double)p->scale_factor;
    ex->scale.assign(nl, 1.f); ex->sigma2.assign(nl, 1.f);
    for (int i = 1; i < nl; i++) {
        ex->scale[i] = (float)(ex->scale[i - 1] * sf);
        ex->sigma2[i] = ex->scale[i] * ex->scale[i];
    }
    ex->invScale.resize(nl); ex->invSigma2.resize(nl);
    for (int i = 0; i < nl; i++) { ex->invScale[i] = 1.0f / ex->scale[i]; ex->invSigma2[i] = 1.0f / ex->sigma2[i]; }
    ex->quota.resize(nl);
    {
        const float factor = (float)(1.0f / sf);
        float nDesired = p->nfeatures * (1 - factor) / (1 - (float)pow((double)factor, (double)nl));
        int sum = 0;
        for (int l = 0; l < nl - 1; l++) {
            ex->quota[l] = cv_round_f(nDesired);
            sum += ex->quota[l];
            nDesired *= factor;
        }
        ex->quota[nl - 1] = std::max(p->nfeatures - sum, 0);
    }
    /* ORBextractor.cc:452-467 */
    ex->umax.assign(16, 0);
    {
        int v, v0;
        const int vmax = (int)floorf(15 * sqrtf(2.f) / 2 + 1), vmin = (int)ceilf(15 * sqrtf(2.f) / 2);
        for (v = 0; v <= vmax; ++v) ex->umax[v] = cv_round_d(sqrt(225.0 - v * v));
        for (v = 15, v0 = 0; v >= vmin; --v) {
            while (ex->umax[v0] == ex->umax[v0 + 1]) ++v0;
            ex->umax[v] = v0;
            ++v0;
        }
    }

    ExDesc& d = ex->d;
    d.nlevels = nl; d.W = width; d.H = height; d.iniTh = p->ini_th_fast; d.minTh = p->min_th_fast;
    std::vector<CellDesc> cells;
    std::vector<StripDesc> strips, stripsShort;
    int rawTotal = 0;
    std::vector<uint2> tabs;
    size_t pyrOff = 0, blurOff = 0;
    int keyTotal = 0, selTotal = 0, nodeCap = 64;
    for (int l = 0; l < nl; l++) {
        LevelDesc& L = d.lv[l];
        L.w = cv_round_f((float)width * ex->invScale[l]);  /* ORBextractor.cc:1156-1157 */
        L.h = cv_round_f((float)height * ex->invScale[l]);
        const float fw = (float)(L.w - 2 * ORBFE_MIN_BORDER), fh = (float)(L.h - 2 * ORBFE_MIN_BORDER);
        L.nCols = fw > 0 ? (int)(fw / 35.f) : 0;
        L.nRows = fh > 0 ? (int)(fh / 35.f) : 0;
        if (L.nCols < 1 || L.nRows < 1) {
            orbfe_set_error("level %d is %dx%d: too small for the 35 px FAST grid (ORBextractor.cc:782-785 would divide by zero)", l, L.w, L.h);
            delete ex;
            return ORBFE_E_INVALID;
        }
        L.wCell = (int)ceilf(fw / L.nCols);
        L.hCell = (int)ceilf(fh / L.nRows);
        L.nIni = (int)roundf(fw / (float)(L.h - 2 * ORBFE_MIN_BORDER)); /* ORBextractor.cc:541 */
        if (L.nIni < 1) {
            orbfe_set_error("level %d aspect ratio gives 0 initial quadtree nodes (ORBextractor.cc:541)", l);
            delete ex;
            return ORBFE_E_INVALID;
        }
        L.hX = fw / L.nIni;
        L.pitch = (int)align_up(ORBFE_ROW_LEAD + L.w + ORBFE_EDGE, 16);
        L.pyrOff = (int)pyrOff;
        pyrOff = align_up(pyrOff + (size_t)L.pitch * (L.h + 2 * ORBFE_EDGE), 256);
        L.bpitch = (int)align_up(L.w, 16);
        L.blurOff = (int)blurOff;
        blurOff = align_up(blurOff + (size_t)L.bpitch * L.h, 256);
        L.quota = ex->quota[l];
        L.scale = ex->scale[l];
        L.kpSize = (float)(int)(31 * ex->scale[l]); /* ORBextractor.cc:862 */
        L.selBase = selTotal;
        L.selCap = std::max(L.quota + 3, 4 * L.nIni);
        selTotal += L.selCap;
        L.cellBase = (int)cells.size();
        L.keyBase = keyTotal;
        const int minB = ORBFE_MIN_BORDER, maxBX = L.w - ORBFE_EDGE + 3, maxBY = L.h - ORBFE_EDGE + 3;
        for (int i = 0; i < L.nRows; i++)
            for (int j = 0; j < L.nCols; j++) {
                CellDesc c;
                memset(&c, 0, sizeof c);
                const int iniY = minB + i * L.hCell, iniX = minB + j * L.wCell;
                int maxY = iniY + L.hCell + 6, maxX = iniX + L.wCell + 6;
                const bool skip = (iniY >= maxBY - 3) || (iniX >= maxBX - 6); /* ORBextractor.cc:792,801 */
                if (maxY > maxBY) maxY = maxBY;
                if (maxX > maxBX) maxX = maxBX;
                c.x0 = (short)iniX; c.y0 = (short)iniY;
                c.rw = (short)(skip ? 0 : maxX - iniX);
                c.rh = (short)(skip ? 0 : maxY - iniY);
                c.level = (short)l;
                c.xoff = (short)(j * L.wCell); c.yoff = (short)(i * L.hCell);
                const int cw = std::max(c.rw - 6, 0), ch = std::max(c.rh - 6, 0);
                c.candBase = keyTotal;
                c.candCap = ((cw + 1) / 2) * ((ch + 1) / 2); /* NMS survivors are never 8-adjacent */
                keyTotal += c.candCap;
                if (cw > 0 && ch > 0) {
                    const int ox = iniX & 3, cx0 = ox + 3, w0 = cx0 >> 2, wl = (cx0 + cw - 1) >> 2;
                    c.ngw = (short)(wl - w0 + 1);
                    c.nw = (short)((ox + c.rw + 3) >> 2);
                    c.invcw = (16777216u + cw - 1) / cw;
                    c.invgw = (16777216u + c.ngw - 1) / c.ngw;
                    c.edgeL = 0x80808080u << (8 * (cx0 & 3));
                    c.edgeR = 0x80808080u >> (8 * (3 - ((cx0 + cw - 1) & 3)));
                }
                if (c.rw > 76 || c.rh > 76) { orbfe_set_error("internal: cell larger than the FAST tile"); delete ex; return ORBFE_E_INVALID; }
                cells.push_back(c);
            }
        L.keyCap = keyTotal - L.keyBase;
        L.invW = (16777216u + L.wCell - 1) / L.wCell;
        L.invH = (16777216u + L.hCell - 1) / L.hCell;
        L.rawBase = rawTotal;
        L.rawStride = L.wCell * L.hCell;
        rawTotal += L.nCols * L.nRows * L.rawStride;
        /* two cuts of the level into strips: 32 rows (batches: less preload overhead; 48 balanced worse, 24 measured the same) and 16 rows (a few frames: three
           times as many warps in flight, lower latency) */
        for (int cut = 0; cut < 2; cut++) {
            const int sh = cut ? 16 : 32;
            std::vector<StripDesc>& dst = cut ? stripsShort : strips;
            for (int y0 = ORBFE_EDGE; y0 < L.h - ORBFE_EDGE; y0 += sh)
                for (int xs = ORBFE_MIN_BORDER; xs < L.w - ORBFE_EDGE; xs += 120) {
                    StripDesc sd;
                    sd.level = (short)l; sd.xs = (short)xs; sd.y0 = (short)y0;
                    sd.nrows = (short)std::min(sh, L.h - ORBFE_EDGE - y0);
                    dst.push_back(sd);
                }
        }
        nodeCap = std::max(nodeCap, std::max(L.selCap, L.nIni));
        if (l > 0) {
            while (tabs.size() % 4) tabs.push_back(make_uint2(0, 0)); /* k_pyramid_level reads tabX four entries at a time */
            L.tabX = (int)tabs.size();
            linear_table(d.lv[l - 1].w, L.w, tabs);
            /* the last column group may be partial (w % 4 != 0): it runs through the same four-pixel path with the last
               entry repeated; the 1-3 surplus bytes fall in the border, which k_pyramid_fringe rewrites afterwards */
            while ((tabs.size() - L.tabX) % 4) tabs.push_back(tabs.back());
            for (int x = 0; x + 3 < L.w; x += 4)
                if (tabs[L.tabX + x + 3].x - tabs[L.tabX + x].x > 6) {
                    orbfe_set_error("scale factor too large for the resize kernel's 8-byte source window");
                    delete ex;
                    return ORBFE_E_INVALID;
                }
            L.tabY = (int)tabs.size();
            linear_table(d.lv[l - 1].h, L.h, tabs);
        }
    }
    int cap2 = 64;
    while (cap2 < nodeCap) cap2 <<= 1;
    if (cap2 > 16384) {
        orbfe_set_error("the quadtree holds 16384 nodes per level: needs %d here (per-level feature quota above 16381)", nodeCap);
        delete ex;
        return ORBFE_E_CAPACITY;
    }
    d.nodeCap = cap2;
    /* up to 2048 nodes the node arrays of a level live in the CTA's shared memory; beyond (5 x nFeatures initialisation extractors of
       the monocular KITTI settings: 10000 features) the same arrays are carved out of a global scratch block per (frame, level) */
    d.qtScratchBytes = cap2 > 2048 ? (int)align_up(orbfe_quadtree_smem_bytes(cap2), 256) : 0;
    d.nCells = (int)cells.size();
    d.candCap = keyTotal;
    d.nStrips = (int)strips.size();
    d.nStripsShort = (int)stripsShort.size();
    d.blurTile0[0] = 0;
    for (int l = 0; l < nl; l++) d.blurTile0[l + 1] = d.blurTile0[l] + orbfe_blur_tiles(d.lv[l].w, d.lv[l].h);
    d.stBands = (height >> 3) + 1;
    d.rawCap = rawTotal;
    d.kpCap = selTotal;
    d.pyrSlab = align_up(pyrOff + 256, 256);
    d.blurSlab = align_up(blurOff + 256, 256);
    ex->quadSmem = d.qtScratchBytes ? 0 : orbfe_quadtree_smem_bytes(cap2);
    {
        int rows = 8, pix = 8;
        for (const CellDesc& c : cells) {
            rows = std::max(rows, (int)c.rh);
            pix = std::max(pix, std::max(c.rw - 6, 0) * std::max(c.rh - 6, 0));
        }
        d.fastRows = rows;
        d.fastPix = (pix + 7) / 8 * 8;
        ex->fastSmem = 2 * rows * 80 + 2 * d.fastPix * 2 + ((d.fastPix + 31) / 32 + 1) * 4 + 16;
    }

    int rc = 0;
    const size_t B = (size_t)max_batch;
    ExBuffers& b = ex->b;
    CellDesc* dCells = nullptr;
    StripDesc* dStrips = nullptr;
    uint2* dTabs = nullptr;
#define TRY(x) do { rc = (x); if (rc) { orbfe_extractor_destroy(ex); return rc; } } while (0)
#define TRYC(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { rc = orbfe_cuda_fail(e_, #x); orbfe_extractor_destroy(ex); return rc; } } while (0)
    TRYC(cudaStreamCreateWithFlags(&ex->stream, cudaStreamNonBlocking));
    TRYC(cudaStreamCreateWithFlags(&ex->copyIn, cudaStreamNonBlocking));
    TRYC(cudaStreamCreateWithFlags(&ex->copyOut, cudaStreamNonBlocking));
    TRYC(cudaStreamCreateWithFlags(&ex->side, cudaStreamNonBlocking));
    TRYC(cudaEventCreateWithFlags(&ex->evFork, cudaEventDisableTiming));
    TRYC(cudaEventCreateWithFlags(&ex->evJoin, cudaEventDisableTiming));
    if (const char* e = getenv("ORBFE_NO_FORK")) ex->forkBlur = atoi(e) ? 0 : 1;
    ex->hostChunk = ORBFE_HOST_CHUNK;
    if (const char* e = getenv("ORBFE_HOST_CHUNK")) { const int v = atoi(e); if (v >= 1) ex->hostChunk = v; }
    ex->evIn.assign((max_batch + ex->hostChunk - 1) / ex->hostChunk, nullptr);
    ex->evDone.assign(ex->evIn.size(), nullptr);
    for (cudaEvent_t& ev : ex->evIn) TRYC(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
    for (cudaEvent_t& ev : ex->evDone) TRYC(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
    ex->evOut.assign(ex->evIn.size(), nullptr);
    ex->evOutValid.assign(ex->evIn.size(), 0);
    for (cudaEvent_t& ev : ex->evOut) TRYC(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
    for (cudaEvent_t& ev : ex->evBatch) TRYC(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
    TRYC(cudaEventCreateWithFlags(&ex->evLast, cudaEventDisableTiming));
    for (cudaEvent_t& ev : ex->ev) TRYC(cudaEventCreate(&ev));
    TRY(dev_alloc(ex, &b.pyr, B * d.pyrSlab + 256));
    TRY(dev_alloc(ex, &b.blur, B * d.blurSlab + 256));
    TRY(dev_alloc(ex, &b.cand, B * d.candCap));
    TRY(dev_alloc(ex, &b.cellCount, B * d.nCells));
    TRY(dev_alloc(ex, &b.keys, B * d.candCap));
    TRY(dev_alloc(ex, &b.keyNode, B * d.candCap));
    TRY(dev_alloc(ex, &b.sel, B * d.kpCap));
    TRY(dev_alloc(ex, &b.selCount, B * nl));
    TRY(dev_alloc(ex, &b.work, B * d.kpCap));
    b.qtScratch = nullptr;
    if (d.qtScratchBytes) TRY(dev_alloc(ex, &b.qtScratch, B * nl * (size_t)d.qtScratchBytes));
    TRY(dev_alloc(ex, &b.kps, B * d.kpCap));
    TRY(dev_alloc(ex, &b.desc, B * d.kpCap * 32));
    TRY(dev_alloc(ex, &b.count, B));
    TRY(dev_alloc(ex, &b.mono, B));
    TRY(dev_alloc(ex, &b.uRight, B * d.kpCap));
    TRY(dev_alloc(ex, &b.depth, B * d.kpCap));
    TRY(dev_alloc(ex, &b.sadBest, B * d.kpCap));
    TRY(dev_alloc(ex, &b.stTable, B * d.kpCap));
    TRY(dev_alloc(ex, &b.stBand, B * (d.stBands + 1)));
    TRY(dev_alloc(ex, &dCells, cells.size()));
    TRY(dev_alloc(ex, &dStrips, strips.size() + stripsShort.size()));
    TRY(dev_alloc(ex, &b.raw, B * (size_t)d.rawCap));
    TRY(dev_alloc(ex, &b.rawCount, B * d.nCells));
    TRY(dev_alloc(ex, &b.fbList, B * d.nCells));
    TRY(dev_alloc(ex, &b.fbCount, 1));
    TRY(dev_alloc(ex, &dTabs, tabs.size()));
    TRY(dev_alloc(ex, &ex->inputStage, B * (size_t)width * height));
    b.cells = dCells;
    b.strips = dStrips;
    TRYC(cudaMemcpy(dStrips, strips.data(), strips.size() * sizeof(StripDesc), cudaMemcpyHostToDevice));
    TRYC(cudaMemcpy(dStrips + strips.size(), stripsShort.data(), stripsShort.size() * sizeof(StripDesc), cudaMemcpyHostToDevice));
    {
        /* the blur launch: tile list in launch order, and one tensor map per level over this object's pyramid slab
           (x = bytes of a padded row, y = padded rows, z = frame) with a 96 x 64 byte box */
        std::vector<uint32_t> tiles;
        std::vector<CUtensorMap> maps(nl);
        for (int l = 0; l < nl; l++) {
            const LevelDesc& L = d.lv[l];
            const int tilesX = (L.w + 63) / 64, n = orbfe_blur_tiles(L.w, L.h);
            for (int t = 0; t < n; t++) tiles.push_back(((uint32_t)l << 24) | ((uint32_t)(t / tilesX) << 12) | (uint32_t)(t % tilesX));
            TRY(encode_tiled_u8_3d(&maps[l], b.pyr + L.pyrOff, (uint64_t)L.pitch, (uint64_t)(L.h + 2 * ORBFE_EDGE), (uint64_t)B, (uint64_t)L.pitch, (uint64_t)d.pyrSlab, 96, 64));
        }
        uint32_t* dTiles = nullptr;
        TRY(dev_alloc(ex, &dTiles, tiles.size()));
        TRY(dev_alloc(ex, &ex->blurMaps, maps.size()));
        TRYC(cudaMemcpy(dTiles, tiles.data(), tiles.size() * sizeof(uint32_t), cudaMemcpyHostToDevice));
        TRYC(cudaMemcpy(ex->blurMaps, maps.data(), maps.size() * sizeof(CUtensorMap), cudaMemcpyHostToDevice));
        b.blurTiles = dTiles;
        const char* e = getenv("ORBFE_BLUR_LDG");
        ex->blurLdg = e && atoi(e) != 0;
    }
    TRYC(cudaMemset(b.rawCount, 0, B * d.nCells * sizeof(int)));
    TRYC(cudaMemset(b.fbCount, 0, sizeof(int)));
    b.tabs = dTabs;
    b.input = ex->inputStage;
    TRYC(cudaMemcpy(dCells, cells.data(), cells.size() * sizeof(CellDesc), cudaMemcpyHostToDevice));
    if (!tabs.empty()) TRYC(cudaMemcpy(dTabs, tabs.data(), tabs.size() * sizeof(uint2), cudaMemcpyHostToDevice));
    TRYC(cudaMemset(b.pyr, 0, B * d.pyrSlab + 256));
    TRYC(cudaMemset(b.blur, 0, B * d.blurSlab + 256));
    TRYC(cudaMemset(b.count, 0, B * sizeof(int)));
    TRYC(cudaMemset(b.mono, 0, B * sizeof(int)));
    TRYC(cudaMallocHost((void**)&ex->hKps, B * d.kpCap * sizeof(orbfe_keypoint)));
    TRYC(cudaMallocHost((void**)&ex->hDesc, B * d.kpCap * 32));
    TRYC(cudaMallocHost((void**)&ex->hCount, 2 * B * sizeof(int)));   /* two slots: orbfe_stereo_frames_submit keeps two batches in flight */
    TRYC(cudaMallocHost((void**)&ex->hMono, 2 * B * sizeof(int)));
    TRY(orbfe_upload_constants(ex->umax.data(), kPattern));
    TRY(orbfe_configure_kernels(ex->quadSmem, ex->fastSmem));
#undef TRY
#undef TRYC
    *out = ex;
    return ORBFE_OK;
}

int orbfe_set_profiling(orbfe_extractor* ex, int enable) {
    if (!ex) return ORBFE_E_INVALID;
    ORBFE_CUDA(cudaSetDevice(ex->device));
    if (!enable && (ex->stagePending || ex->stereoPending)) { int rc = orbfe_collect_stage_times(ex); if (rc) return rc; }
    ex->profiling = enable ? 1 : 0;
    return ORBFE_OK;
}

int orbfe_stage_times(orbfe_extractor* ex, double* ms8, int64_t* launches) {
    if (!ex) return ORBFE_E_INVALID;
    ORBFE_CUDA(cudaSetDevice(ex->device));
    int rc = orbfe_collect_stage_times(ex);
    if (rc) return rc;
    if (ms8) for (int i = 0; i < ORBFE_NUM_STAGES + 2; i++) { ms8[i] += ex->stageMs[i]; ex->stageMs[i] = 0.0; }
    if (launches) *launches = ex->launches;
    return ORBFE_OK;
}

int orbfe_get_levels(const orbfe_extractor* ex) { return ex ? ex->d.nlevels : ORBFE_E_INVALID; }

int orbfe_get_scale_factors(const orbfe_extractor* ex, float* s, float* is, float* s2, float* is2) {
    if (!ex) return ORBFE_E_INVALID;
    for (int i = 0; i < ex->d.nlevels; i++) {
        if (s) s[i] = ex->scale[i];
        if (is) is[i] = ex->invScale[i];
        if (s2) s2[i] = ex->sigma2[i];
        if (is2) is2[i] = ex->invSigma2[i];
    }
    return ORBFE_OK;
}

int orbfe_get_features_per_level(const orbfe_extractor* ex, int32_t* q) {
    if (!ex || !q) return ORBFE_E_INVALID;
    for (int i = 0; i < ex->d.nlevels; i++) q[i] = ex->quota[i];
    return ORBFE_OK;
}

int orbfe_get_umax(const orbfe_extractor* ex, int32_t* u) {
    if (!ex || !u) return ORBFE_E_INVALID;
    for (int i = 0; i < 16; i++) u[i] = ex->umax[i];
    return ORBFE_OK;
}

int orbfe_keypoint_capacity(const orbfe_extractor* ex) { return ex ? ex->d.kpCap : ORBFE_E_INVALID; }

int orbfe_level_size(const orbfe_extractor* ex, int level, int* w, int* h) {
    if (!ex || level < 0 || level >= ex->d.nlevels) return ORBFE_E_INVALID;
    if (w) *w = ex->d.lv[level].w;
    if (h) *h = ex->d.lv[level].h;
    return ORBFE_OK;
}

int orbfe_extract_batch_device(orbfe_extractor* ex, const uint8_t* d_images, int nframes, int lap0, int lap1, void* stream) {
    if (!ex || !d_images) { orbfe_set_error("null argument"); return ORBFE_E_INVALID; }
    if (nframes < 1 || nframes > ex->maxBatch) {
        orbfe_set_error("nframes %d outside [1, max_batch=%d]", nframes, ex->maxBatch);
        return ORBFE_E_CAPACITY;
    }
    ORBFE_CUDA(cudaSetDevice(ex->device));
    ex->lastBatch = nframes;
    cudaStream_t st = stream ? (cudaStream_t)stream : ex->stream;
    const int rc = orbfe_launch_extract(ex, d_images, nframes, lap0, lap1, st, 0);
    return rc ? rc : orbfe_mark_last(ex, st);
}

/* The stereo Frame constructors (src/Frame.cc:104-170 rectified stereo, :1023-1101 fisheye rig) for a batch of pairs already on
 * the device: the reference extracts the two images on two threads (:119-122, :1051-1054) and matches when both are back; here the
 * left images go through exL on its stream, the right ones through exR on its stream, and (match != 0) ComputeStereoMatches
 * follows on the left stream once the right one is done.  `user` (may be null) is the caller's stream: everything enqueued here
 * starts after what it holds now and is finished, for that stream, when the call returns (fork / join by events; nothing blocks
 * the host). */
static int pair_batch_device(orbfe_extractor* exL, orbfe_extractor* exR, const uint8_t* d_left, const uint8_t* d_right, int nframes, int lap0, int lap1,
                             int match, float mbf, float min_d, float max_d, cudaStream_t user) {
    if (!exL || !exR || !d_left || !d_right) { orbfe_set_error("null argument"); return ORBFE_E_INVALID; }
    if (exL == exR || exL->device != exR->device || exL->d.W != exR->d.W || exL->d.H != exR->d.H || exL->d.kpCap != exR->d.kpCap || exL->d.nlevels != exR->d.nlevels) {
        orbfe_set_error("left/right extractors must be two objects of the same geometry on one device");
        return ORBFE_E_INVALID;
    }
    if (nframes < 1 || nframes > exL->maxBatch || nframes > exR->maxBatch) {
        orbfe_set_error("nframes %d outside [1, max_batch=%d]", nframes, exL->maxBatch < exR->maxBatch ? exL->maxBatch : exR->maxBatch);
        return ORBFE_E_CAPACITY;
    }
    if (exL->pendingCount) { orbfe_set_error("batches submitted with orbfe_stereo_frames_submit are still in flight"); return ORBFE_E_INVALID; }
    ORBFE_CUDA(cudaSetDevice(exL->device));
    cudaStream_t sl = exL->stream, sr = exR->stream;
    if (user) {
        ORBFE_CUDA(cudaEventRecord(exL->evIn[0], user));
        ORBFE_CUDA(cudaStreamWaitEvent(sl, exL->evIn[0], 0));
        ORBFE_CUDA(cudaStreamWaitEvent(sr, exL->evIn[0], 0));
    }
    exL->lastBatch = exR->lastBatch = nframes;
    int rc = orbfe_launch_extract(exL, d_left, nframes, lap0, lap1, sl, 0);
    if (rc) return rc;
    rc = orbfe_launch_extract(exR, d_right, nframes, lap0, lap1, sr, 0);
    if (rc) return rc;
    ORBFE_CUDA(cudaEventRecord(exR->evDone[0], sr));
    ORBFE_CUDA(cudaStreamWaitEvent(sl, exR->evDone[0], 0));
    if (match) {
        rc = orbfe_launch_stereo(exL, exR, nframes, 0, mbf, min_d, max_d, sl);
        if (rc) return rc;
    }
    /* the right object's next batch must not overwrite what a matcher on the left stream (or on the caller's) still reads */
    ORBFE_CUDA(cudaEventRecord(exL->evDone[0], sl));
    ORBFE_CUDA(cudaStreamWaitEvent(sr, exL->evDone[0], 0));
    if (user) {
        ORBFE_CUDA(cudaStreamWaitEvent(user, exL->evDone[0], 0));
        rc = orbfe_mark_last(exL, user);
        return rc ? rc : orbfe_mark_last(exR, user);
    }
    rc = orbfe_mark_last(exL, sl);
    return rc ? rc : orbfe_mark_last(exR, sl);
}

int orbfe_stereo_frames_batch_device(orbfe_extractor* exL, orbfe_extractor* exR, const uint8_t* d_left, const uint8_t* d_right, int nframes, int lap0, int lap1,
                                     float mbf, float min_d, float max_d, void* stream) {
    return pair_batch_device(exL, exR, d_left, d_right, nframes, lap0, lap1, 1, mbf, min_d, max_d, (cudaStream_t)stream);
}

int orbfe_extract_pair_batch_device(orbfe_extractor* exL, orbfe_extractor* exR, const uint8_t* d_left, const uint8_t* d_right, int nframes, int lap0, int lap1,
                                    void* stream) {
    return pair_batch_device(exL, exR, d_left, d_right, nframes, lap0, lap1, 0, 0.f, 0.f, 0.f, (cudaStream_t)stream);
}

int orbfe_batch_counts(orbfe_extractor* ex, int nframes, int32_t* n, int32_t* mono) {
    if (!ex || nframes < 1 || nframes > ex->lastBatch) { orbfe_set_error("bad frame count"); return ORBFE_E_INVALID; }
    ORBFE_CUDA(cudaSetDevice(ex->device));
    { const int rc = orbfe_wait_last(ex); if (rc) return rc; }
    if (n) ORBFE_D2H(ex, n, ex->b.count, nframes * sizeof(int));
    if (mono) ORBFE_D2H(ex, mono, ex->b.mono, nframes * sizeof(int));
    ORBFE_D2H_DONE(ex);
    return ORBFE_OK;
}

int orbfe_batch_result(orbfe_extractor* ex, int frame, orbfe_keypoint* kps, uint8_t* desc, int cap, int* n, int* mono) {
    if (!ex || frame < 0 || frame >= ex->lastBatch) { orbfe_set_error("bad frame index"); return ORBFE_E_INVALID; }
    ORBFE_CUDA(cudaSetDevice(ex->device));
    { const int rc = orbfe_wait_last(ex); if (rc) return rc; }
    int cnt = 0, mi = 0;
    ORBFE_D2H(ex, &cnt, ex->b.count + frame, sizeof(int));
    ORBFE_D2H(ex, &mi, ex->b.mono + frame, sizeof(int));
    ORBFE_D2H_DONE(ex);
    if (n) *n = cnt;
    if (mono) *mono = mi;
    if (cnt > cap) { orbfe_set_error("caller capacity %d < %d keypoints", cap, cnt); return ORBFE_E_CAPACITY; }
    if (kps && cnt) ORBFE_D2H(ex, kps, ex->b.kps + (size_t)frame * ex->d.kpCap, (size_t)cnt * sizeof(orbfe_keypoint));
    if (desc && cnt) ORBFE_D2H(ex, desc, ex->b.desc + (size_t)frame * ex->d.kpCap * 32, (size_t)cnt * 32);
    ORBFE_D2H_DONE(ex);
    return ORBFE_OK;
}

int orbfe_batch_device_views(orbfe_extractor* ex, const orbfe_keypoint** k, const uint8_t** de, const int32_t** c, const int32_t** m) {
    if (!ex) return ORBFE_E_INVALID;
    if (k) *k = ex->b.kps;
    if (de) *de = ex->b.desc;
    if (c) *c = ex->b.count;
    if (m) *m = ex->b.mono;
    return ORBFE_OK;
}

int orbfe_extract_batch_host(orbfe_extractor* ex, const uint8_t* h_images, int nframes, int lap0, int lap1,
                             orbfe_keypoint* kps, uint8_t* desc, int cap, int32_t* n, int32_t* mono) {
    if (!ex || !h_images) { orbfe_set_error("null argument"); return ORBFE_E_INVALID; }
    if (nframes < 1 || nframes > ex->maxBatch) { orbfe_set_error("nframes %d outside [1, max_batch=%d]", nframes, ex->maxBatch); return ORBFE_E_CAPACITY; }
    ORBFE_CUDA(cudaSetDevice(ex->device));
    const size_t frameBytes = (size_t)ex->d.W * ex->d.H;
    const int kc = ex->d.kpCap;
    const bool direct = cap >= kc;   /* caller rows are at least as wide as ours: strided copies straight into the caller's arrays */
    cudaStream_t st = ex->stream;
    ex->lastBatch = nframes;
    /* three-stage pipeline over chunks of frames: H2D of chunk k+1 (copyIn), kernels of chunk k (st), D2H of the results
     * of chunk k-1 (copyOut) overlap; the chunks themselves run in order on `st` */
    const int C = ex->hostChunk;
    for (int k = 0, f0 = 0; f0 < nframes; k++, f0 += C) {
        const int nf = std::min(C, nframes - f0);
        ORBFE_CUDA(cudaMemcpyAsync(ex->inputStage + (size_t)f0 * frameBytes, h_images + (size_t)f0 * frameBytes, frameBytes * nf, cudaMemcpyHostToDevice, ex->copyIn));
        ORBFE_CUDA(cudaEventRecord(ex->evIn[k], ex->copyIn));
        ORBFE_CUDA(cudaStreamWaitEvent(st, ex->evIn[k], 0));
        int rc = orbfe_launch_extract(ex, ex->inputStage + (size_t)f0 * frameBytes, nf, lap0, lap1, st, f0);
        if (rc) return rc;
        ORBFE_CUDA(cudaEventRecord(ex->evDone[k], st));
        ORBFE_CUDA(cudaStreamWaitEvent(ex->copyOut, ex->evDone[k], 0));
        ORBFE_CUDA(cudaMemcpyAsync(ex->hCount + f0, ex->b.count + f0, nf * sizeof(int), cudaMemcpyDeviceToHost, ex->copyOut));
        ORBFE_CUDA(cudaMemcpyAsync(ex->hMono + f0, ex->b.mono + f0, nf * sizeof(int), cudaMemcpyDeviceToHost, ex->copyOut));
        const orbfe_keypoint* dk = ex->b.kps + (size_t)f0 * kc;
        const uint8_t* dd = ex->b.desc + (size_t)f0 * kc * 32;
        if (direct) {
            if (kps) ORBFE_CUDA(cudaMemcpy2DAsync(kps + (size_t)f0 * cap, (size_t)cap * sizeof(orbfe_keypoint), dk, (size_t)kc * sizeof(orbfe_keypoint),
                                                 (size_t)kc * sizeof(orbfe_keypoint), nf, cudaMemcpyDeviceToHost, ex->copyOut));
            if (desc) ORBFE_CUDA(cudaMemcpy2DAsync(desc + (size_t)f0 * cap * 32, (size_t)cap * 32, dd, (size_t)kc * 32, (size_t)kc * 32, nf, cudaMemcpyDeviceToHost, ex->copyOut));
        } else {
            if (kps) ORBFE_CUDA(cudaMemcpyAsync(ex->hKps + (size_t)f0 * kc, dk, (size_t)nf * kc * sizeof(orbfe_keypoint), cudaMemcpyDeviceToHost, ex->copyOut));
            if (desc) ORBFE_CUDA(cudaMemcpyAsync(ex->hDesc + (size_t)f0 * kc * 32, dd, (size_t)nf * kc * 32, cudaMemcpyDeviceToHost, ex->copyOut));
        }
    }
    ORBFE_CUDA(cudaStreamSynchronize(ex->copyOut));
    ORBFE_CUDA(cudaStreamSynchronize(st));
    if (!direct)
        for (int f = 0; f < nframes; f++) {
            if (ex->hCount[f] > cap) { orbfe_set_error("caller capacity %d < %d keypoints (frame %d)", cap, ex->hCount[f], f); return ORBFE_E_CAPACITY; }
            if (kps) memcpy(kps + (size_t)f * cap, ex->hKps + (size_t)f * kc, (size_t)ex->hCount[f] * sizeof(orbfe_keypoint));
            if (desc) memcpy(desc + (size_t)f * cap * 32, ex->hDesc + (size_t)f * kc * 32, (size_t)ex->hCount[f] * 32);
        }
    for (int f = 0; f < nframes; f++) {
        if (n) n[f] = ex->hCount[f];
        if (mono) mono[f] = ex->hMono[f];
    }
    return ORBFE_OK;
}

/* One batch of stereo pairs through upload -> both extractions -> stereo matching -> download, enqueued chunk by chunk and not
 * waited for.  `slot` (0/1) selects the input staging buffers and the pinned count arrays, so that the upload of one batch can
 * run while the previous batch is still in the kernels. */
/* device buffers of the fisheye-rig stereo outputs (match vectors and 3-D points; the depths share b.depth): allocated on first use */
static int fisheye_buffers(orbfe_extractor* exL) {
    if (exL->fishLR) return ORBFE_OK;
    const size_t n = (size_t)exL->maxBatch * exL->d.kpCap;
    ORBFE_CUDA(cudaMalloc((void**)&exL->fishLR, n * sizeof(int32_t))); exL->allocs.push_back(exL->fishLR);
    ORBFE_CUDA(cudaMalloc((void**)&exL->fishRL, n * sizeof(int32_t))); exL->allocs.push_back(exL->fishRL);
    ORBFE_CUDA(cudaMalloc((void**)&exL->fishP3d, n * 3 * sizeof(float))); exL->allocs.push_back(exL->fishP3d);
    return ORBFE_OK;
}

/* Frame::ComputeStereoFishEyeMatches (src/Frame.cc:1119-1159) for the frames both extractors processed last, results on the host:
 * knn-2 + ratio, TriangulateMatches, the depth test.  match_lr / depth: per frame `cap` entries of which the first n_left are written,
 * match_rl: the first n_right, p3d: 3 * n_left (cap >= the keypoint capacity).  Synchronous. */
extern "C" int orbfe_fisheye_stereo_frames(orbfe_extractor* exL, orbfe_extractor* exR, int nframes, const orbfe_fisheye_rig* rig, double ratio, int32_t* match_lr,
                                           int32_t* match_rl, float* depth, float* p3d, int cap) {
    if (!exL || !exR || !rig) { orbfe_set_error("null argument"); return ORBFE_E_INVALID; }
    if (exL == exR || exL->device != exR->device || exL->d.kpCap != exR->d.kpCap) { orbfe_set_error("left/right extractors must be two objects of the same geometry on one device"); return ORBFE_E_INVALID; }
    if (nframes < 1 || nframes > exL->lastBatch || nframes > exR->lastBatch) { orbfe_set_error("nframes exceeds the last extracted batch"); return ORBFE_E_INVALID; }
    if (cap < exL->d.kpCap) { orbfe_set_error("cap %d < keypoint capacity %d", cap, exL->d.kpCap); return ORBFE_E_CAPACITY; }
    if (exL->pendingCount) { orbfe_set_error("batches submitted with orbfe_stereo_frames_submit are still in flight"); return ORBFE_E_INVALID; }
    ORBFE_CUDA(cudaSetDevice(exL->device));
    { int rc = fisheye_buffers(exL); if (rc) return rc; }
    { int rc = orbfe_wait_last(exL); if (!rc) rc = orbfe_wait_last(exR); if (rc) return rc; }
    cudaStream_t st = exL->stream;
    const int kc = exL->d.kpCap;
    int rc = orbfe_launch_fisheye(exL, exR, nframes, 0, ratio, rig, exL->fishLR, exL->fishRL, exL->b.depth, exL->fishP3d, st);
    if (rc) return rc;
    const size_t w = (size_t)kc * 4;
    if (match_lr) ORBFE_CUDA(cudaMemcpy2DAsync(match_lr, (size_t)cap * 4, exL->fishLR, w, w, nframes, cudaMemcpyDeviceToHost, st));
    if (match_rl) ORBFE_CUDA(cudaMemcpy2DAsync(match_rl, (size_t)cap * 4, exL->fishRL, w, w, nframes, cudaMemcpyDeviceToHost, st));
    if (depth) ORBFE_CUDA(cudaMemcpy2DAsync(depth, (size_t)cap * 4, exL->b.depth, w, w, nframes, cudaMemcpyDeviceToHost, st));
    if (p3d) ORBFE_CUDA(cudaMemcpy2DAsync(p3d, (size_t)cap * 12, exL->fishP3d, 3 * w, 3 * w, nframes, cudaMemcpyDeviceToHost, st));
    ORBFE_CUDA(cudaStreamSynchronize(st));
    rc = orbfe_mark_last(exL, st);
    return rc ? rc : orbfe_mark_last(exR, st);
}

struct FishArgs {   /* the fisheye-rig variant of the batch: ComputeStereoFishEyeMatches instead of ComputeStereoMatches */
    const orbfe_fisheye_rig* rig;
    double ratio;
    int32_t *lr, *rl;   /* host outputs [nframes][cap] (may be NULL) */
    float* p3d;         /* host output [nframes][cap][3] (may be NULL) */
};

static int stereo_frames_enqueue(orbfe_extractor* exL, orbfe_extractor* exR, const uint8_t* h_left, const uint8_t* h_right, int nframes, int lap0, int lap1,
                                 float mbf, float min_d, float max_d, orbfe_keypoint* kps_l, uint8_t* desc_l, orbfe_keypoint* kps_r, uint8_t* desc_r, int cap,
                                 float* u_right, float* depth, int slot, int chunk, const FishArgs* fish) {
    const int kc = exL->d.kpCap;
    if (fish) { const int rc = fisheye_buffers(exL); if (rc) return rc; }
    const size_t frameBytes = (size_t)exL->d.W * exL->d.H;
    exL->lastBatch = exR->lastBatch = nframes;
    cudaStream_t sl = exL->stream, sr = exR->stream, in = exL->copyIn, out = exL->copyOut;
    orbfe_extractor* exs[2] = {exL, exR};
    for (orbfe_extractor* ex : exs)
        if (slot && !ex->inputStage2) {   /* second staging buffer: allocated the first time two batches are in flight */
            ORBFE_CUDA(cudaMalloc((void**)&ex->inputStage2, (size_t)ex->maxBatch * frameBytes));
            ex->allocs.push_back(ex->inputStage2);
        }
    uint8_t* stageL = slot ? exL->inputStage2 : exL->inputStage;
    uint8_t* stageR = slot ? exR->inputStage2 : exR->inputStage;
    const size_t hs = (size_t)slot * exL->maxBatch, hsR = (size_t)slot * exR->maxBatch;
    /* per chunk: upload L then R on one copy stream (so the left extraction can start after half the upload), extract L and R on
     * their own streams, match on the left stream once the right one is done, download everything on the other copy stream */
    const int C = chunk;
    for (int k = 0, f0 = 0; f0 < nframes; k++, f0 += C) {
        const int nf = std::min(C, nframes - f0);
        ORBFE_CUDA(cudaMemcpyAsync(stageL + (size_t)f0 * frameBytes, h_left + (size_t)f0 * frameBytes, frameBytes * nf, cudaMemcpyHostToDevice, in));
        ORBFE_CUDA(cudaEventRecord(exL->evIn[k], in));
        ORBFE_CUDA(cudaMemcpyAsync(stageR + (size_t)f0 * frameBytes, h_right + (size_t)f0 * frameBytes, frameBytes * nf, cudaMemcpyHostToDevice, in));
        ORBFE_CUDA(cudaEventRecord(exR->evIn[k], in));
        ORBFE_CUDA(cudaStreamWaitEvent(sl, exL->evIn[k], 0));
        if (exL->evOutValid[k]) {   /* the previous batch's download of these frames' results must be over before they are rewritten */
            ORBFE_CUDA(cudaStreamWaitEvent(sl, exL->evOut[k], 0));
            ORBFE_CUDA(cudaStreamWaitEvent(sr, exL->evOut[k], 0));
        }
        int rc = orbfe_launch_extract(exL, stageL + (size_t)f0 * frameBytes, nf, lap0, lap1, sl, f0);
        if (rc) return rc;
        ORBFE_CUDA(cudaStreamWaitEvent(sr, exR->evIn[k], 0));
        rc = orbfe_launch_extract(exR, stageR + (size_t)f0 * frameBytes, nf, lap0, lap1, sr, f0);
        if (rc) return rc;
        ORBFE_CUDA(cudaEventRecord(exR->evDone[k], sr));
        ORBFE_CUDA(cudaStreamWaitEvent(sl, exR->evDone[k], 0));
        if (fish) rc = orbfe_launch_fisheye(exL, exR, nf, f0, fish->ratio, fish->rig, exL->fishLR + (size_t)f0 * kc, exL->fishRL + (size_t)f0 * kc,
                                            exL->b.depth + (size_t)f0 * kc, exL->fishP3d + (size_t)f0 * kc * 3, sl);
        else rc = orbfe_launch_stereo(exL, exR, nf, f0, mbf, min_d, max_d, sl);
        if (rc) return rc;
        ORBFE_CUDA(cudaEventRecord(exL->evDone[k], sl));
        ORBFE_CUDA(cudaStreamWaitEvent(out, exL->evDone[k], 0));
        orbfe_keypoint* kk[2] = {kps_l, kps_r};
        uint8_t* dd[2] = {desc_l, desc_r};
        for (int s = 0; s < 2; s++) {
            orbfe_extractor* ex = exs[s];
            const size_t h0 = (s ? hsR : hs) + f0;
            ORBFE_CUDA(cudaMemcpyAsync(ex->hCount + h0, ex->b.count + f0, nf * sizeof(int), cudaMemcpyDeviceToHost, out));
            ORBFE_CUDA(cudaMemcpyAsync(ex->hMono + h0, ex->b.mono + f0, nf * sizeof(int), cudaMemcpyDeviceToHost, out));
            if (kk[s]) ORBFE_CUDA(cudaMemcpy2DAsync(kk[s] + (size_t)f0 * cap, (size_t)cap * sizeof(orbfe_keypoint), ex->b.kps + (size_t)f0 * kc, (size_t)kc * sizeof(orbfe_keypoint),
                                                   (size_t)kc * sizeof(orbfe_keypoint), nf, cudaMemcpyDeviceToHost, out));
            if (dd[s]) ORBFE_CUDA(cudaMemcpy2DAsync(dd[s] + (size_t)f0 * cap * 32, (size_t)cap * 32, ex->b.desc + (size_t)f0 * kc * 32, (size_t)kc * 32, (size_t)kc * 32, nf,
                                                   cudaMemcpyDeviceToHost, out));
        }
        const size_t w = (size_t)kc * sizeof(float);
        if (u_right) ORBFE_CUDA(cudaMemcpy2DAsync(u_right + (size_t)f0 * cap, (size_t)cap * sizeof(float), exL->b.uRight + (size_t)f0 * kc, w, w, nf, cudaMemcpyDeviceToHost, out));
        if (depth) ORBFE_CUDA(cudaMemcpy2DAsync(depth + (size_t)f0 * cap, (size_t)cap * sizeof(float), exL->b.depth + (size_t)f0 * kc, w, w, nf, cudaMemcpyDeviceToHost, out));
        if (fish) {
            if (fish->lr) ORBFE_CUDA(cudaMemcpy2DAsync(fish->lr + (size_t)f0 * cap, (size_t)cap * 4, exL->fishLR + (size_t)f0 * kc, w, w, nf, cudaMemcpyDeviceToHost, out));
            if (fish->rl) ORBFE_CUDA(cudaMemcpy2DAsync(fish->rl + (size_t)f0 * cap, (size_t)cap * 4, exL->fishRL + (size_t)f0 * kc, w, w, nf, cudaMemcpyDeviceToHost, out));
            if (fish->p3d) ORBFE_CUDA(cudaMemcpy2DAsync(fish->p3d + (size_t)f0 * cap * 3, (size_t)cap * 12, exL->fishP3d + (size_t)f0 * kc * 3, 3 * w, 3 * w, nf, cudaMemcpyDeviceToHost, out));
        }
        ORBFE_CUDA(cudaEventRecord(exL->evOut[k], out));
        exL->evOutValid[k] = 1;
    }
    ORBFE_CUDA(cudaEventRecord(exL->evBatch[slot], out));   /* the downloads are the last work of a batch */
    return ORBFE_OK;
}

static int stereo_frames_check(orbfe_extractor* exL, orbfe_extractor* exR, const uint8_t* h_left, const uint8_t* h_right, int nframes, int cap) {
    if (!exL || !exR || !h_left || !h_right) { orbfe_set_error("null argument"); return ORBFE_E_INVALID; }
    if (exL == exR || exL->device != exR->device || exL->d.W != exR->d.W || exL->d.H != exR->d.H || exL->d.kpCap != exR->d.kpCap ||
        exL->d.nlevels != exR->d.nlevels || exL->hostChunk != exR->hostChunk) { orbfe_set_error("left/right extractors must be two objects of the same geometry on one device"); return ORBFE_E_INVALID; }
    if (nframes < 1 || nframes > exL->maxBatch || nframes > exR->maxBatch) { orbfe_set_error("nframes %d outside [1, max_batch]", nframes); return ORBFE_E_CAPACITY; }
    if (cap < exL->d.kpCap) { orbfe_set_error("cap %d < keypoint capacity %d", cap, exL->d.kpCap); return ORBFE_E_CAPACITY; }
    return ORBFE_OK;
}

static int stereo_frames_submit(orbfe_extractor* exL, orbfe_extractor* exR, const uint8_t* h_left, const uint8_t* h_right, int nframes, int lap0, int lap1,
                                float mbf, float min_d, float max_d, orbfe_keypoint* kps_l, uint8_t* desc_l, orbfe_keypoint* kps_r, uint8_t* desc_r, int cap,
                                float* u_right, float* depth, int chunk, const FishArgs* fish = nullptr) {
    int rc = stereo_frames_check(exL, exR, h_left, h_right, nframes, cap);
    if (rc) return rc;
    if (exL->pendingCount >= 2) { orbfe_set_error("two batches are already in flight: call orbfe_stereo_frames_wait first"); return ORBFE_E_CAPACITY; }
    ORBFE_CUDA(cudaSetDevice(exL->device));
    const int slot = (exL->pendingHead + exL->pendingCount) & 1;
    rc = stereo_frames_enqueue(exL, exR, h_left, h_right, nframes, lap0, lap1, mbf, min_d, max_d, kps_l, desc_l, kps_r, desc_r, cap, u_right, depth, slot, chunk, fish);
    if (rc) return rc;
    exL->pendingFrames[slot] = nframes;
    exL->pendingCount++;
    return ORBFE_OK;
}

int orbfe_stereo_frames_submit(orbfe_extractor* exL, orbfe_extractor* exR, const uint8_t* h_left, const uint8_t* h_right, int nframes, int lap0, int lap1,
                               float mbf, float min_d, float max_d, orbfe_keypoint* kps_l, uint8_t* desc_l, orbfe_keypoint* kps_r, uint8_t* desc_r, int cap,
                               float* u_right, float* depth) {
    /* with a second batch in flight the copies of one batch hide behind the kernels of the other, so the chunks can be twice as
     * large as in the synchronous call (fewer, fuller launches): measured 125.8 k -> 128.4 k frames/s on 256 EuRoC pairs */
    return stereo_frames_submit(exL, exR, h_left, h_right, nframes, lap0, lap1, mbf, min_d, max_d, kps_l, desc_l, kps_r, desc_r, cap, u_right, depth,
                                exL ? 2 * exL->hostChunk : 0);
}

/* Frame's fisheye-rig constructor (src/Frame.cc:1023-1101) for a batch of pairs from host memory, pipelined like
 * orbfe_stereo_frames_submit: both extractions with the lapping interval, then ComputeStereoFishEyeMatches (knn-2 + ratio,
 * TriangulateMatches, the four output vectors).  Collected by orbfe_stereo_frames_wait. */
int orbfe_fisheye_frames_submit(orbfe_extractor* exL, orbfe_extractor* exR, const uint8_t* h_left, const uint8_t* h_right, int nframes, int lap0, int lap1,
                                const orbfe_fisheye_rig* rig, double ratio, orbfe_keypoint* kps_l, uint8_t* desc_l, orbfe_keypoint* kps_r, uint8_t* desc_r, int cap,
                                int32_t* match_lr, int32_t* match_rl, float* depth, float* p3d) {
    if (!rig) { orbfe_set_error("null rig"); return ORBFE_E_INVALID; }
    const FishArgs fish = {rig, ratio, match_lr, match_rl, p3d};
    return stereo_frames_submit(exL, exR, h_left, h_right, nframes, lap0, lap1, 0.f, 0.f, 0.f, kps_l, desc_l, kps_r, desc_r, cap, nullptr, depth,
                                exL ? 2 * exL->hostChunk : 0, &fish);
}

int orbfe_stereo_frames_wait(orbfe_extractor* exL, orbfe_extractor* exR, int32_t* n_l, int32_t* mono_l, int32_t* n_r, int32_t* mono_r) {
    if (!exL || !exR) { orbfe_set_error("null argument"); return ORBFE_E_INVALID; }
    if (exL->pendingCount < 1) { orbfe_set_error("no batch in flight"); return ORBFE_E_INVALID; }
    ORBFE_CUDA(cudaSetDevice(exL->device));
    const int slot = exL->pendingHead;
    ORBFE_CUDA(cudaEventSynchronize(exL->evBatch[slot]));
    const int nframes = exL->pendingFrames[slot];
    const size_t hs = (size_t)slot * exL->maxBatch, hsR = (size_t)slot * exR->maxBatch;
    for (int f = 0; f < nframes; f++) {
        if (n_l) n_l[f] = exL->hCount[hs + f];
        if (mono_l) mono_l[f] = exL->hMono[hs + f];
        if (n_r) n_r[f] = exR->hCount[hsR + f];
        if (mono_r) mono_r[f] = exR->hMono[hsR + f];
    }
    exL->pendingHead ^= 1;
    exL->pendingCount--;
    return ORBFE_OK;
}

int orbfe_stereo_frames_batch_host(orbfe_extractor* exL, orbfe_extractor* exR, const uint8_t* h_left, const uint8_t* h_right, int nframes, int lap0,
                                   int lap1, float mbf, float min_d, float max_d, orbfe_keypoint* kps_l, uint8_t* desc_l, int32_t* n_l, int32_t* mono_l,
                                   orbfe_keypoint* kps_r, uint8_t* desc_r, int32_t* n_r, int32_t* mono_r, int cap, float* u_right, float* depth) {
    int rc = stereo_frames_check(exL, exR, h_left, h_right, nframes, cap);
    if (rc) return rc;
    if (exL->pendingCount) { orbfe_set_error("batches submitted with orbfe_stereo_frames_submit are still in flight"); return ORBFE_E_INVALID; }
    rc = stereo_frames_submit(exL, exR, h_left, h_right, nframes, lap0, lap1, mbf, min_d, max_d, kps_l, desc_l, kps_r, desc_r, cap, u_right, depth, exL->hostChunk);
    if (rc) return rc;
    return orbfe_stereo_frames_wait(exL, exR, n_l, mono_l, n_r, mono_r);
}

int orbfe_extract(orbfe_extractor* ex, const uint8_t* image, int rows, int cols, size_t step, int lap0, int lap1,
                  orbfe_keypoint* kps, uint8_t* desc, int cap, int* n, int* mono_index) {
    if (n) *n = 0;
    if (mono_index) *mono_index = -1;
    if (!ex) { orbfe_set_error("null extractor"); return ORBFE_E_INVALID; }
    if (!image || rows <= 0 || cols <= 0) { orbfe_set_error("empty image"); return ORBFE_E_EMPTY; }
    if (rows != ex->d.H || cols != ex->d.W || step < (size_t)cols) {
        orbfe_set_error("image %dx%d does not match the extractor shape %dx%d", cols, rows, ex->d.W, ex->d.H);
        return ORBFE_E_INVALID;
    }
    ORBFE_CUDA(cudaSetDevice(ex->device));
    cudaStream_t st = ex->stream;
    ORBFE_CUDA(cudaMemcpy2DAsync(ex->inputStage, cols, image, step, cols, rows, cudaMemcpyHostToDevice, st));
    ex->lastBatch = 1;
    /* one frame is sixteen short kernels: launch-bound.  They and the result downloads are captured once per lapping interval
     * into a CUDA graph and replayed (not while per-stage profiling is on: its events belong to no graph). */
    const bool useGraph = ex->singleGraph != 0 && ex->profiling == 0;
    if (useGraph && (!ex->graphExec || ex->graphLap0 != lap0 || ex->graphLap1 != lap1)) {
        if (ex->graphExec) { cudaGraphExecDestroy(ex->graphExec); ex->graphExec = nullptr; }
        cudaGraph_t g = nullptr;
        ORBFE_CUDA(cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal));
        int rc = orbfe_launch_extract(ex, ex->inputStage, 1, lap0, lap1, st, 0);
        cudaError_t e = cudaSuccess;
        if (!rc) e = cudaMemcpyAsync(ex->hCount, ex->b.count, sizeof(int), cudaMemcpyDeviceToHost, st);
        if (!rc && e == cudaSuccess) e = cudaMemcpyAsync(ex->hMono, ex->b.mono, sizeof(int), cudaMemcpyDeviceToHost, st);
        if (!rc && e == cudaSuccess) e = cudaMemcpyAsync(ex->hKps, ex->b.kps, (size_t)ex->d.kpCap * sizeof(orbfe_keypoint), cudaMemcpyDeviceToHost, st);
        if (!rc && e == cudaSuccess) e = cudaMemcpyAsync(ex->hDesc, ex->b.desc, (size_t)ex->d.kpCap * 32, cudaMemcpyDeviceToHost, st);
        const cudaError_t e2 = cudaStreamEndCapture(st, &g);
        if (rc) { if (g) cudaGraphDestroy(g); return rc; }
        if (e != cudaSuccess || e2 != cudaSuccess) { if (g) cudaGraphDestroy(g); return orbfe_cuda_fail(e != cudaSuccess ? e : e2, "graph capture of the single-frame path"); }
        e = cudaGraphInstantiate(&ex->graphExec, g, 0);
        cudaGraphDestroy(g);
        if (e != cudaSuccess) { ex->graphExec = nullptr; return orbfe_cuda_fail(e, "cudaGraphInstantiate"); }
        ex->graphLap0 = lap0; ex->graphLap1 = lap1;
        ex->graphLaunches = ex->d.nlevels + 8;
        ex->launches -= ex->graphLaunches;   /* the capture counted them once without running them */
    }
    if (useGraph) {
        ORBFE_CUDA(cudaGraphLaunch(ex->graphExec, st));
        ex->launches += ex->graphLaunches;
        ex->batchFrames = 1; ex->bordersValid = 0;
    } else {
        int rc = orbfe_launch_extract(ex, ex->inputStage, 1, lap0, lap1, st, 0);
        if (rc) return rc;
        ORBFE_CUDA(cudaMemcpyAsync(ex->hCount, ex->b.count, sizeof(int), cudaMemcpyDeviceToHost, st));
        ORBFE_CUDA(cudaMemcpyAsync(ex->hMono, ex->b.mono, sizeof(int), cudaMemcpyDeviceToHost, st));
        ORBFE_CUDA(cudaMemcpyAsync(ex->hKps, ex->b.kps, (size_t)ex->d.kpCap * sizeof(orbfe_keypoint), cudaMemcpyDeviceToHost, st));
        ORBFE_CUDA(cudaMemcpyAsync(ex->hDesc, ex->b.desc, (size_t)ex->d.kpCap * 32, cudaMemcpyDeviceToHost, st));
    }
    ORBFE_CUDA(cudaStreamSynchronize(st));
    const int cnt = ex->hCount[0];
    if (n) *n = cnt;
    if (mono_index) *mono_index = ex->hMono[0];
    if (cnt > cap) { orbfe_set_error("caller capacity %d < %d keypoints", cap, cnt); return ORBFE_E_CAPACITY; }
    if (kps) memcpy(kps, ex->hKps, (size_t)cnt * sizeof(orbfe_keypoint));
    if (desc) memcpy(desc, ex->hDesc, (size_t)cnt * 32);
    return ORBFE_OK;
}

int orbfe_pyramid_level(orbfe_extractor* ex, int frame, int level, int padded, uint8_t* dst) {
    if (!ex || !dst || level < 0 || level >= ex->d.nlevels || frame < 0 || frame >= ex->maxBatch) return ORBFE_E_INVALID;
    ORBFE_CUDA(cudaSetDevice(ex->device));
    { const int rc = orbfe_wait_last(ex); if (rc) return rc; }
    if (padded) {   /* the extraction writes only the 3 px of border it reads itself; the rest is filled here, once per batch */
        int rc = orbfe_materialize_borders(ex, ex->stream);
        if (rc) return rc;
    }
    const LevelDesc& L = ex->d.lv[level];
    const int B = padded ? ORBFE_EDGE : 0;
    const uint8_t* src = ex->b.pyr + (size_t)frame * ex->d.pyrSlab + L.pyrOff + (size_t)(ORBFE_EDGE - B) * L.pitch + ORBFE_ROW_LEAD - B;
    ORBFE_D2H_2D(ex, dst, L.w + 2 * B, src, L.pitch, L.w + 2 * B, L.h + 2 * B);   /* stream order: behind the border kernel */
    ORBFE_D2H_DONE(ex);
    return ORBFE_OK;
}

int orbfe_pyramid_levels(orbfe_extractor* ex, int frame, int padded, uint8_t* const* dst_levels) {
    if (!ex || !dst_levels || frame < 0 || frame >= ex->maxBatch) return ORBFE_E_INVALID;
    ORBFE_CUDA(cudaSetDevice(ex->device));
    { const int rc = orbfe_wait_last(ex); if (rc) return rc; }
    if (padded) { const int rc = orbfe_materialize_borders(ex, ex->stream); if (rc) return rc; }
    const int B = padded ? ORBFE_EDGE : 0;
    for (int l = 0; l < ex->d.nlevels; l++) {
        if (!dst_levels[l]) continue;
        const LevelDesc& L = ex->d.lv[l];
        const uint8_t* src = ex->b.pyr + (size_t)frame * ex->d.pyrSlab + L.pyrOff + (size_t)(ORBFE_EDGE - B) * L.pitch + ORBFE_ROW_LEAD - B;
        ORBFE_D2H_2D(ex, dst_levels[l], L.w + 2 * B, src, L.pitch, L.w + 2 * B, L.h + 2 * B);
    }
    ORBFE_D2H_DONE(ex);   /* one wait for the border kernel and all the copies */
    return ORBFE_OK;
}

int orbfe_pin_host(void* p, size_t bytes) {
    if (!p || !bytes) return ORBFE_E_INVALID;
    ORBFE_CUDA(cudaHostRegister(p, bytes, cudaHostRegisterDefault));
    return ORBFE_OK;
}

int orbfe_unpin_host(void* p) {
    if (!p) return ORBFE_E_INVALID;
    ORBFE_CUDA(cudaHostUnregister(p));
    return ORBFE_OK;
}

int orbfe_blurred_level(orbfe_extractor* ex, int frame, int level, uint8_t* dst) {
    if (!ex || !dst || level < 0 || level >= ex->d.nlevels || frame < 0 || frame >= ex->maxBatch) return ORBFE_E_INVALID;
    ORBFE_CUDA(cudaSetDevice(ex->device));
    { const int rc = orbfe_wait_last(ex); if (rc) return rc; }
    const LevelDesc& L = ex->d.lv[level];
    ORBFE_D2H_2D(ex, dst, L.w, ex->b.blur + (size_t)frame * ex->d.blurSlab + L.blurOff, L.bpitch, L.w, L.h);
    ORBFE_D2H_DONE(ex);
    return ORBFE_OK;
}

int orbfe_fast_candidates(orbfe_extractor* ex, int frame, int level, orbfe_keypoint* out, int cap, int* n) {
    if (!ex || level < 0 || level >= ex->d.nlevels || frame < 0 || frame >= ex->maxBatch) return ORBFE_E_INVALID;
    ORBFE_CUDA(cudaSetDevice(ex->device));
    { const int rc = orbfe_wait_last(ex); if (rc) return rc; }
    const LevelDesc& L = ex->d.lv[level];
    const int ncell = L.nCols * L.nRows;
    std::vector<int> counts(ncell);
    std::vector<CellDesc> cells(ncell);
    ORBFE_D2H(ex, counts.data(), ex->b.cellCount + (size_t)frame * ex->d.nCells + L.cellBase, ncell * sizeof(int));
    ORBFE_D2H(ex, cells.data(), ex->b.cells + L.cellBase, ncell * sizeof(CellDesc));
    std::vector<uint32_t> keys(std::max(L.keyCap, 1));
    ORBFE_D2H(ex, keys.data(), ex->b.cand + (size_t)frame * ex->d.candCap + L.keyBase, (size_t)L.keyCap * 4);
    ORBFE_D2H_DONE(ex);
    int total = 0;
    for (int c = 0; c < ncell; c++)
        for (int i = 0; i < counts[c]; i++, total++) {
            if (!out || total >= cap) continue;
            const uint32_t p = keys[cells[c].candBase - L.keyBase + i];
            orbfe_keypoint& o = out[total];
            o.x = (float)(p & 0xFFF); o.y = (float)((p >> 12) & 0xFFF); o.size = 7.f; o.angle = -1.f;
            o.response = (float)(p >> 24); o.octave = 0; o.class_id = -1;
        }
    if (n) *n = total;
    return ORBFE_OK;
}

}  /* extern "C" */
