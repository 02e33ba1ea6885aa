#include <cstdio>
#include <cstdint>
__global__ void k(const int* din, int* out) {
    int d[25];
    for (int k = 0; k < 16; k++) d[k] = din[k];
    for (int k = 0; k < 9; k++) d[16 + k] = d[k];
    int m = -256;
    for (int k = 0; k < 16; k++) {
        int lo = d[k], hi = d[k];
        for (int j = 1; j < 9; j++) { lo = min(lo, d[k + j]); hi = max(hi, d[k + j]); }
        out[2 * k] = lo; out[2 * k + 1] = hi;
        m = max(m, max(lo, -hi));
    }
    out[32] = m;
    out[33] = min(-15, 25); out[34] = max(-15, 25); out[35] = min(din[0], din[1]); out[36] = max(din[0], din[1]);
}
int main() {
    int hd[16] = {25, -15, 0, -1, -8, -13, -6, 99, -12, -11, -41, -6, -1, 0, -84, 0};
    int *dd, *dout, ho[40];
    cudaMalloc(&dd, 64); cudaMalloc(&dout, 160);
    cudaMemcpy(dd, hd, 64, cudaMemcpyHostToDevice);
    k<<<1, 1>>>(dd, dout);
    cudaMemcpy(ho, dout, 160, cudaMemcpyDeviceToHost);
    for (int k = 0; k < 16; k++) printf("arc %d lo %d hi %d\n", k, ho[2 * k], ho[2 * k + 1]);
    printf("m %d | min(-15,25) %d max %d | rt min %d max %d | err %s\n", ho[32], ho[33], ho[34], ho[35], ho[36], cudaGetErrorString(cudaGetLastError()));
}
