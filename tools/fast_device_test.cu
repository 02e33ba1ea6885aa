// Development aid: device-side FAST primitives vs a host reference on random tiles.
#include "../orb_slam3_detailed_comments_kor_b200/csrc/orbfe_extract.cu"
#include <cstdio>
#include <cstdlib>
#include <algorithm>
__global__ void k_test(const uint8_t* tile, int* outc, int* outs, int t) {
    int x = 3 + threadIdx.x, y = 3 + blockIdx.x;
    __shared__ uint8_t s[FAST_RH * FAST_TP];
    for (int i = threadIdx.x; i < FAST_RH * FAST_TP; i += blockDim.x) s[i] = tile[i];
    __syncthreads();
    if (x >= FAST_TP - 3) return;
    const uint8_t* c = s + y * FAST_TP + x;
    outc[y * FAST_TP + x] = fast_is_corner(c, t);
    outs[y * FAST_TP + x] = fast_score(c);
}
static int ref_m(const uint8_t* p, int step) {
    const int X[16] = {0, 1, 2, 3, 3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1};
    const int Y[16] = {3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1, 0, 1, 2, 3};
    int d[25];
    for (int k = 0; k < 16; k++) d[k] = (int)p[0] - (int)p[Y[k] * step + X[k]];
    for (int k = 0; k < 9; k++) d[16 + k] = d[k];
    int m = -256;
    for (int k = 0; k < 16; k++) { int lo = d[k], hi = d[k]; for (int j = 1; j < 9; j++) { lo = std::min(lo, d[k + j]); hi = std::max(hi, d[k + j]); } m = std::max(m, std::max(lo, -hi)); }
    return m;
}
int main() {
    static uint8_t tile[FAST_RH * FAST_TP];
    uint8_t* dt; int *dc, *ds;
    cudaMalloc(&dt, sizeof tile); cudaMalloc(&dc, sizeof(int) * FAST_RH * FAST_TP); cudaMalloc(&ds, sizeof(int) * FAST_RH * FAST_TP);
    static int hc[FAST_RH * FAST_TP], hs[FAST_RH * FAST_TP];
    srand(1);
    int bad = 0, n = 0;
    for (int it = 0; it < 50; it++) {
        for (int i = 0; i < FAST_RH * FAST_TP; i++) tile[i] = (uint8_t)((rand() % 4 == 0) ? rand() % 256 : 100 + rand() % 20);
        cudaMemcpy(dt, tile, sizeof tile, cudaMemcpyHostToDevice);
        k_test<<<FAST_RH - 6, 96>>>(dt, dc, ds, 20);
        cudaMemcpy(hc, dc, sizeof hc, cudaMemcpyDeviceToHost); cudaMemcpy(hs, ds, sizeof hs, cudaMemcpyDeviceToHost);
        for (int y = 3; y < FAST_RH - 3; y++) for (int x = 3; x < FAST_TP - 3; x++) {
            int m = ref_m(tile + y * FAST_TP + x, FAST_TP); n++;
            if (hc[y * FAST_TP + x] != (m > 20)) { if (bad < 5) printf("corner mismatch m=%d got %d\n", m, hc[y * FAST_TP + x]); bad++; }
            if (hs[y * FAST_TP + x] != m - 1) { if (bad < 5) printf("score mismatch m=%d got %d\n", m, hs[y * FAST_TP + x]); bad++; }
        }
    }
    printf("checked %d bad %d err %s\n", n, bad, cudaGetErrorString(cudaGetLastError()));
    return 0;
}
