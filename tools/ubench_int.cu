// ubench_int.cu — integer pipe throughput on sm_100a: IMAD vs IDP.4A vs IDP.2A vs PRMT/SHF,
// plus LDS.128 bandwidth.  Decides how the MC filters are expressed.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/ubench_int tools/ubench_int.cu
#include <cstdio>
#include <cuda_runtime.h>

#define ITER 4096
template <int OP>
__global__ void k(int *out, int a0, int b0)
{
    int acc[8];
    int a = a0 + threadIdx.x, b = b0;
#pragma unroll
    for (int i = 0; i < 8; i++) acc[i] = i + threadIdx.x;
    for (int it = 0; it < ITER; it++) {
#pragma unroll
        for (int i = 0; i < 8; i++) {
            if (OP == 0) acc[i] = acc[i] * b + a;                       // IMAD (dependent per chain)
            if (OP == 1) acc[i] = __dp4a(a, acc[i], acc[i]);            // IDP.4A
            if (OP == 2) acc[i] = __dp2a_lo(a, b, acc[i]);              // IDP.2A
            if (OP == 3) acc[i] = __byte_perm(acc[i], a, 0x5410 + (b & 0x1111)); // PRMT
            if (OP == 4) acc[i] = __funnelshift_r(acc[i], a, 16) ^ b;   // SHF (+LOP)
            if (OP == 5) { acc[i] = __dp2a_lo(a, b, acc[i]); a = __byte_perm(a, acc[(i + 4) & 7], 0x7632); } // mixed
            if (OP == 6) acc[i] = a * b + acc[i];                       // IMAD accumulate form
            if (OP == 7) acc[i] = min(max(acc[i] + a, 0), 1023);        // clip
        }
    }
    int s = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) s += acc[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

__global__ void lds_k(int *out)
{
    __shared__ uint4 sm[1024];
    for (int i = threadIdx.x; i < 1024; i += blockDim.x) sm[i] = make_uint4(i, i, i, i);
    __syncthreads();
    uint4 acc = make_uint4(0, 0, 0, 0);
    int idx = threadIdx.x;
    for (int it = 0; it < ITER; it++) {
#pragma unroll
        for (int i = 0; i < 8; i++) {
            uint4 v = sm[(idx + i * 32) & 1023];
            acc.x += v.x; acc.y ^= v.y; acc.z += v.z; acc.w ^= v.w;
        }
        idx += acc.x & 1;
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc.x + acc.y + acc.z + acc.w;
}

template <int OP> void run(const char *name, int *d)
{
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    const int blocks = 148 * 8, threads = 256;
    k<OP><<<blocks, threads>>>(d, 3, 5);
    cudaDeviceSynchronize();
    cudaEventRecord(e0);
    k<OP><<<blocks, threads>>>(d, 3, 5);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    double ops = (double)blocks * threads * ITER * 8;
    printf("%-28s %8.3f ms  %8.2f Tops/s  (%.1f ops/clk/SM at 1.9 GHz)\n", name, ms, ops / ms / 1e9,
           ops / (ms * 1e-3) / 148 / 1.9e9);
}

int main()
{
    int *d; cudaMalloc(&d, 148 * 8 * 256 * 4);
    run<0>("IMAD (chain a*b+c)", d);
    run<6>("IMAD (accumulate)", d);
    run<1>("IDP.4A", d);
    run<2>("IDP.2A", d);
    run<3>("PRMT", d);
    run<4>("SHF+LOP", d);
    run<5>("IDP.2A + PRMT interleaved", d);
    run<7>("IADD+VIMNMX clip", d);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    lds_k<<<148 * 8, 256>>>(d); cudaDeviceSynchronize();
    cudaEventRecord(e0); lds_k<<<148 * 8, 256>>>(d); cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    double bytes = (double)148 * 8 * 256 * ITER * 8 * 16;
    printf("LDS.128                      %8.3f ms  %8.2f TB/s  (%.1f B/clk/SM at 1.9 GHz)\n", ms, bytes / ms / 1e9,
           bytes / (ms * 1e-3) / 148 / 1.9e9);
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
