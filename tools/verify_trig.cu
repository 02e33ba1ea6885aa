// verify_trig.cu — exhaustive check of the one float step of the descriptor kernel that is not reproduced operation by
// operation: a = cosf(angle*pi/180), b = sinf(angle*pi/180) (reference src/ORBextractor.cc:110-111, host libm) versus the
// kernel's (float)cos((double)x), (float)sin((double)x).  Every binary32 angle in [0, 360] (the range of cv::fastAtan2) is
// evaluated on the GPU and on the host; where a or b differ, the 512 rotated tap coordinates
// cvRound(x*b + y*a), cvRound(x*a - y*b) are compared too, because only they reach the descriptor.
//   nvcc -O2 -std=c++17 --fmad=false -gencode arch=compute_100a,code=sm_100a -o tools/verify_trig.bin tools/verify_trig.cu
//   ./tools/verify_trig.bin      (on the GPU box; prints one summary line and writes gpurun_out/orbfe_trig_exceptions.inc)
// With -DORBFE_APPLY_TRIG_EXCEPTIONS it checks the kernel's own routine (csrc/orbfe_trig.cuh, table applied): expected 0 taps.
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <thread>
#include <vector>
#ifdef ORBFE_APPLY_TRIG_EXCEPTIONS
#include "../orb_slam3_detailed_comments_kor_b200/csrc/orbfe_trig.cuh"   /* the kernel's own routine, exception table included */
#endif
static const int8_t kPattern[1024] = {
#include "../include/orbfe_pattern31.inc"
};
__global__ void k_trig(uint32_t first, uint32_t count, float* ca, float* sa) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= count) return;
    const float angle = __uint_as_float(first + i);
#ifdef ORBFE_APPLY_TRIG_EXCEPTIONS
    orbfe_rotation(angle, &ca[i], &sa[i]);
#else
    const float ang = __fmul_rn(angle, (float)(3.14159265358979323846 / 180.0));
    double sd, cd;
    sincos((double)ang, &sd, &cd);
    ca[i] = (float)cd;
    sa[i] = (float)sd;
#endif
}
#include <mutex>
#include <algorithm>
#include <array>
static std::mutex g_mu;
static std::vector<std::array<uint32_t, 3>> g_ex;   /* angle bits, host cosf bits, host sinf bits */
static inline int cvr(float v) { return (int)lrintf(v); }
int main() {
    uint32_t lastBits;
    const float last = 360.0f;
    memcpy(&lastBits, &last, 4);
    const uint32_t CH = 1u << 24;
    float *dca, *dsa;
#ifdef ORBFE_APPLY_TRIG_EXCEPTIONS
    orbfe_trig_init();
#endif
    cudaMalloc(&dca, CH * 4); cudaMalloc(&dsa, CH * 4);
    std::vector<float> hca(CH), hsa(CH);
    const int NT = std::max(1u, std::thread::hardware_concurrency());
    unsigned long long total = 0, diffA = 0, diffB = 0, anglesWithTapDiff = 0, tapDiffs = 0;
    uint32_t firstBad = 0;
    for (uint64_t first = 0; first <= lastBits; first += CH) {
        const uint32_t count = (uint32_t)std::min<uint64_t>(CH, (uint64_t)lastBits + 1 - first);
        k_trig<<<(count + 255) / 256, 256>>>((uint32_t)first, count, dca, dsa);
        cudaMemcpy(hca.data(), dca, (size_t)count * 4, cudaMemcpyDeviceToHost);
        cudaMemcpy(hsa.data(), dsa, (size_t)count * 4, cudaMemcpyDeviceToHost);
        std::vector<unsigned long long> dA(NT, 0), dB(NT, 0), dT(NT, 0), dN(NT, 0);
        std::vector<uint32_t> fb(NT, 0);
        std::vector<std::thread> th;
        for (int t = 0; t < NT; t++)
            th.emplace_back([&, t]() {
                for (uint32_t i = t; i < count; i += NT) {
                    uint32_t bits = (uint32_t)first + i;
                    float angle; memcpy(&angle, &bits, 4);
                    const float ang = angle * (float)(3.14159265358979323846 / 180.f);
                    const float a = cosf(ang), b = sinf(ang);
                    const bool da = a != hca[i], db = b != hsa[i];
                    if (!da && !db) continue;
                    dA[t] += da; dB[t] += db;
                    int nd = 0;
                    for (int p = 0; p < 512; p++) {
                        const float x = kPattern[2 * p], y = kPattern[2 * p + 1];
                        nd += cvr(x * b + y * a) != cvr(x * hsa[i] + y * hca[i]);
                        nd += cvr(x * a - y * b) != cvr(x * hca[i] - y * hsa[i]);
                    }
                    if (nd) {
                        dT[t] += nd; dN[t]++; if (!fb[t]) fb[t] = bits;
                        uint32_t ab, bb; memcpy(&ab, &a, 4); memcpy(&bb, &b, 4);
                        std::lock_guard<std::mutex> lk(g_mu);
                        g_ex.push_back({bits, ab, bb});
                    }
                }
            });
        for (auto& x : th) x.join();
        for (int t = 0; t < NT; t++) { diffA += dA[t]; diffB += dB[t]; tapDiffs += dT[t]; anglesWithTapDiff += dN[t]; if (!firstBad && fb[t]) firstBad = fb[t]; }
        total += count;
    }
    printf("{\"angles_checked\": %llu, \"cos_differs\": %llu, \"sin_differs\": %llu, \"angles_with_a_different_tap\": %llu, "
           "\"different_tap_coordinates\": %llu, \"first_bad_angle_bits\": \"0x%08x\", \"cuda_error\": \"%s\"}\n",
           total, diffA, diffB, anglesWithTapDiff, tapDiffs, firstBad, cudaGetErrorString(cudaGetLastError()));
    std::sort(g_ex.begin(), g_ex.end());
    FILE* f = g_ex.empty() ? nullptr : fopen("gpurun_out/orbfe_trig_exceptions.inc", "w");
    if (f) {
        fprintf(f, "/* angles (binary32 bits) for which (float)cos((double)x), (float)sin((double)x) rotate at least one rBRIEF tap to a\n"
                   "   different pixel than the host libm's cosf/sinf do, with the host values; generated by tools/verify_trig.cu on the\n"
                   "   GPU box (exhaustive over [0, 360]).  {angle, cosf, sinf} */\n");
        for (auto& e : g_ex) fprintf(f, "{0x%08xu, 0x%08xu, 0x%08xu},\n", e[0], e[1], e[2]);
        fclose(f);
    }
    return 0;
}
