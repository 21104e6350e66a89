#include <cstdio>
#include "../mam3slam_b200/csrc/orbx_kernels_image.cuh"
using namespace orbx;
__global__ void k(const int* in, int* out) {
    int r[16]; for (int i = 0; i < 16; ++i) r[i] = in[1 + i];
    out[0] = fast_strength(in[0], r);
}
int main() {
    int h[17] = {123, 78,128,101,87,114,99,101,115,95,111,129,76,83,87,54,83};
    int *d, *o; cudaMalloc(&d, sizeof(h)); cudaMalloc(&o, 16); cudaMemcpy(d, h, sizeof(h), cudaMemcpyHostToDevice);
    k<<<1,1>>>(d, o); int r[1]; cudaMemcpy(r, o, 4, cudaMemcpyDeviceToHost); printf("s=%d (expect 0)\n", r[0]);
}
