// prints the shared-window offset of the dynamic shared memory base for a few launch shapes (is it 1024-byte aligned?)
#include <cstdio>
#include <cstdint>
extern __shared__ uint8_t smem_raw[];
__global__ void k(int *out) { if (threadIdx.x == 0 && blockIdx.x == 0) *out = (int)__cvta_generic_to_shared(smem_raw); }
int main() {
    int *d; cudaMalloc(&d, 4);
    for (int bytes : {1024, 65536, 200000, 232448}) {
        cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
        k<<<148, 640, bytes>>>(d);
        int h = -1; cudaMemcpy(&h, d, 4, cudaMemcpyDeviceToHost);
        printf("dynamic smem %d bytes: base offset %d (mod 1024 = %d) err=%s\n", bytes, h, h & 1023, cudaGetErrorString(cudaGetLastError()));
    }
    return 0;
}
