/* orbfe_trig.cuh — a = cosf(angle*pi/180), b = sinf(angle*pi/180) of computeOrbDescriptor
 * (/root/reference/src/ORBextractor.cc:110-111) on the device.
 *
 * The reference takes cosf/sinf from the host libm (glibc), which is not correctly rounded; the device evaluates
 * cos/sin in double and rounds once.  tools/verify_trig.cu compared the two for every binary32 angle in [0, 360] on
 * the GPU box (1,135,869,953 angles): they differ by 1 ulp for 0.04 % (cos) / 0.09 % (sin) of the angles, and for 96
 * angles the difference moves one of the 512 rotated taps to another pixel.  Those 96 angles are listed, with the host
 * values, in orbfe_trig_exceptions.inc and are looked up here, so the tap coordinates — the only way a and b reach the
 * descriptor — are identical to the reference's for every possible angle (glibc 2.39, x86-64). */
#ifndef ORBFE_TRIG_CUH
#define ORBFE_TRIG_CUH
#include <stdint.h>

struct OrbfeTrigEx { uint32_t angle, c, s; };
__constant__ OrbfeTrigEx c_trigEx[] = {
#include "orbfe_trig_exceptions.inc"
};
#define ORBFE_N_TRIGEX ((int)(sizeof(c_trigEx) / sizeof(c_trigEx[0])))
__constant__ uint32_t c_trigBloom[32];   /* 1024-bit filter over the table's angle keys (filled by orbfe_trig_init) */
#define ORBFE_TRIG_HASH(key) (((key) * 2654435761u) >> 22)

static __device__ __forceinline__ void orbfe_rotation(float angle, float* ca, float* sa) {
    const float ang = __fmul_rn(angle, (float)(3.14159265358979323846 / 180.0));   /* kpt.angle * factorPI */
    double sd, cd;
    sincos((double)ang, &sd, &cd);
    float c = (float)cd, s = (float)sd;
    const uint32_t key = __float_as_uint(angle);
    const uint32_t h = ORBFE_TRIG_HASH(key);
    int lo = 0, hi = ((c_trigBloom[h >> 5] >> (h & 31)) & 1u) ? ORBFE_N_TRIGEX - 1 : -1;   /* almost always: not listed */
    while (lo <= hi) {                       /* sorted table, warp-uniform key: constant-cache broadcasts */
        const int mid = (lo + hi) >> 1;
        const uint32_t k = c_trigEx[mid].angle;
        if (k == key) { c = __uint_as_float(c_trigEx[mid].c); s = __uint_as_float(c_trigEx[mid].s); break; }
        if (k < key) lo = mid + 1; else hi = mid - 1;
    }
    *ca = c;
    *sa = s;
}

/* host: fill the filter on the current device (once per device, from orbfe_extractor_create) */
static inline cudaError_t orbfe_trig_init() {
    static const OrbfeTrigEx host_table[] = {
#include "orbfe_trig_exceptions.inc"
    };
    uint32_t bloom[32] = {0};
    for (const OrbfeTrigEx& e : host_table) {
        const uint32_t h = ORBFE_TRIG_HASH(e.angle);
        bloom[h >> 5] |= 1u << (h & 31);
    }
    return cudaMemcpyToSymbol(c_trigBloom, bloom, sizeof bloom);
}
#endif
