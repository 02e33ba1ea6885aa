// Development aid: which formulation of the FAST score survives the device compiler.
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <algorithm>
#define TP 80
#define RH 76
__device__ __host__ inline int offk(int k) {
    const int X[16] = {0, 1, 2, 3, 3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1};
    const int Y[16] = {3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1, 0, 1, 2, 3};
    return Y[k] * TP + X[k];
}
__device__ int score_a(const uint8_t* c) {
    const int v = c[0];
    int d[16];
#pragma unroll
    for (int k = 0; k < 16; k++) d[k] = v - (int)c[offk(k)];
    int lo2[16], hi2[16], lo4[16], hi4[16];
#pragma unroll
    for (int k = 0; k < 16; k++) { lo2[k] = min(d[k], d[(k + 1) & 15]); hi2[k] = max(d[k], d[(k + 1) & 15]); }
#pragma unroll
    for (int k = 0; k < 16; k++) { lo4[k] = min(lo2[k], lo2[(k + 2) & 15]); hi4[k] = max(hi2[k], hi2[(k + 2) & 15]); }
    int m = -256;
#pragma unroll
    for (int k = 0; k < 16; k++) {
        const int lo9 = min(min(lo4[k], lo4[(k + 4) & 15]), d[(k + 8) & 15]);
        const int hi9 = max(max(hi4[k], hi4[(k + 4) & 15]), d[(k + 8) & 15]);
        m = max(m, max(lo9, -hi9));
    }
    return m - 1;
}
__device__ int score_b(const uint8_t* c) {
    const int v = c[0];
    int d[25];
    for (int k = 0; k < 16; k++) d[k] = v - (int)c[offk(k)];
    for (int k = 0; k < 9; k++) d[16 + k] = d[k];
    int m = -256;
    for (int k = 0; k < 16; k++) { int lo = d[k], hi = d[k]; for (int j = 1; j < 9; j++) { lo = min(lo, d[k + j]); hi = max(hi, d[k + j]); } m = max(m, max(lo, -hi)); }
    return m - 1;
}
__global__ void k_test(const uint8_t* tile, int* oa, int* ob, int* od) {
    int x = 3 + threadIdx.x, y = 3 + blockIdx.x;
    __shared__ uint8_t s[RH * TP];
    for (int i = threadIdx.x; i < RH * TP; i += blockDim.x) s[i] = tile[i];
    __syncthreads();
    if (x >= TP - 3) return;
    const uint8_t* c = s + y * TP + x;
    oa[y * TP + x] = score_a(c);
    ob[y * TP + x] = score_b(c);
    if (x == 10 && y == 10) for (int k = 0; k < 16; k++) od[k] = (int)c[0] - (int)c[offk(k)];
}
static int ref_m(const uint8_t* p) {
    int d[25];
    for (int k = 0; k < 16; k++) d[k] = (int)p[0] - (int)p[offk(k)];
    for (int k = 0; k < 9; k++) d[16 + k] = d[k];
    int m = -256;
    for (int k = 0; k < 16; k++) { int lo = d[k], hi = d[k]; for (int j = 1; j < 9; j++) { lo = std::min(lo, d[k + j]); hi = std::max(hi, d[k + j]); } m = std::max(m, std::max(lo, -hi)); }
    return m;
}
int main() {
    static uint8_t tile[RH * TP];
    uint8_t* dt; int *da, *db, *dd;
    cudaMalloc(&dt, sizeof tile); cudaMalloc(&da, 4 * RH * TP); cudaMalloc(&db, 4 * RH * TP); cudaMalloc(&dd, 64);
    static int ha[RH * TP], hb[RH * TP], hd[16];
    srand(1);
    for (int i = 0; i < RH * TP; i++) tile[i] = (uint8_t)((rand() % 4 == 0) ? rand() % 256 : 100 + rand() % 20);
    cudaMemcpy(dt, tile, sizeof tile, cudaMemcpyHostToDevice);
    k_test<<<RH - 6, 96>>>(dt, da, db, dd);
    cudaMemcpy(ha, da, sizeof ha, cudaMemcpyDeviceToHost); cudaMemcpy(hb, db, sizeof hb, cudaMemcpyDeviceToHost); cudaMemcpy(hd, dd, 64, cudaMemcpyDeviceToHost);
    int bada = 0, badb = 0;
    for (int y = 3; y < RH - 3; y++) for (int x = 3; x < TP - 3; x++) {
        int m = ref_m(tile + y * TP + x);
        bada += ha[y * TP + x] != m - 1; badb += hb[y * TP + x] != m - 1;
    }
    printf("bad a %d bad b %d\n", bada, badb);
    const uint8_t* c = tile + 10 * TP + 10;
    printf("device d: "); for (int k = 0; k < 16; k++) printf("%d ", hd[k]); printf("\nhost   d: "); for (int k = 0; k < 16; k++) printf("%d ", (int)c[0] - (int)c[offk(k)]);
    printf("\nhost m %d dev a %d dev b %d\n", ref_m(c), ha[10 * TP + 10], hb[10 * TP + 10]);
    return 0;
}
