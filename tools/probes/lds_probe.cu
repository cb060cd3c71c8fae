// lds_probe.cu -- how many cycles a 64-bit shared-memory gather costs per warp instruction on this GPU, by address pattern.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o lds_probe lds_probe.cu ; run on the GPU box.
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

__global__ void probe(const unsigned int* __restrict__ offs, int n_off, int iters, float* out, long long* cycles)
{
    extern __shared__ float2 s[];
    for (int i = threadIdx.x; i < 16384; i += blockDim.x) s[i] = make_float2((float)i, 1.0f);
    __syncthreads();
    const int lane = threadIdx.x & 31;
    unsigned int off[8];
    for (int k = 0; k < 8; ++k) off[k] = offs[(k * 32 + lane) % n_off];
    float a = 0.f, b = 0.f;
    __syncthreads();
    const long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            float2 h;
            const unsigned int addr = (unsigned int)__cvta_generic_to_shared(s) + off[k];
            asm volatile("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(h.x), "=f"(h.y) : "r"(addr));
            a += h.x; b += h.y;
        }
    }
    const long long t1 = clock64();
    if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
    out[blockIdx.x * blockDim.x + threadIdx.x] = a + b;
}

int main()
{
    const int threads = 512, iters = 2000;
    unsigned int h[256];
    unsigned int* d_off; float* d_out; long long* d_cyc;
    cudaMalloc(&d_off, sizeof(h)); cudaMalloc(&d_out, 148 * threads * sizeof(float)); cudaMalloc(&d_cyc, 148 * sizeof(long long));
    cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 131072);
    const char* names[] = {"linear: lane l -> 8 l (conflict-free)", "lanes l and l+16 on the same bank pair, different rows ([col][16] replica)",
                           "random columns, plain interleaved row", "all lanes one address (broadcast)", "lane l -> 8 (l & 15): halves share addresses",
                           "random columns, [col][16] replica", "lanes l and l+16 same bank pair via [col][32] float (32 copies, 32-bit)"};
    srand(1);
    for (int p = 0; p < 6; ++p) {
        for (int k = 0; k < 8; ++k)
            for (int l = 0; l < 32; ++l) {
                const unsigned int col = (unsigned int)(rand() % 784);
                unsigned int o = 0;
                switch (p) {
                    case 0: o = 8u * l + 256u * k; break;
                    case 1: o = 8u * (l & 15) + 128u * (unsigned)(l >> 4) + 256u * k; break;
                    case 2: o = 8u * col; break;
                    case 3: o = 64u * k; break;
                    case 4: o = 8u * (l & 15) + 256u * k; break;
                    case 5: o = 128u * col + 8u * (l & 15); break;
                }
                h[k * 32 + l] = o;
            }
        cudaMemcpy(d_off, h, sizeof(h), cudaMemcpyHostToDevice);
        probe<<<148, threads, 131072>>>(d_off, 256, iters, d_out, d_cyc);
        cudaDeviceSynchronize();
        long long c[148];
        cudaMemcpy(c, d_cyc, sizeof(c), cudaMemcpyDeviceToHost);
        double avg = 0; for (int i = 0; i < 148; ++i) avg += (double)c[i]; avg /= 148;
        // per SM: 16 warps x iters x 8 loads
        printf("%-90s %.2f cycles per 64-bit warp load per SM (16 warps resident)\n", names[p], avg / ((double)iters * 8 * 16));
    }
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
