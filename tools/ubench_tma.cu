// ubench_tma.cu — how fast does TMA (cp.async.bulk.tensor.2d) fetch many SMALL boxes at
// arbitrary (unaligned) coordinates of a 4K u16 plane?  This is the access pattern of HEVC
// motion compensation (one reference window per PU).  Tensor maps live in GLOBAL memory
// (indexed by box class), written by the host before the launch.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/ubench_tma tools/ubench_tma.cu
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <vector>
#include <cuda.h>
#include <cuda_runtime.h>

typedef CUresult (*EncodeTiled)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t *b, int count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(b)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *b, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(b)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *b, uint32_t parity)
{
    asm volatile(
        "{\n.reg .pred p;\nWAIT:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE;\nbra WAIT;\nDONE:\n}" ::"r"(smem_u32(b)), "r"(parity) : "memory");
}
__device__ __forceinline__ void tma_load_2d(void *dst, const CUtensorMap *map, int x, int y, uint64_t *bar)
{
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
        ::"r"(smem_u32(dst)), "l"(map), "r"(x), "r"(y), "r"(smem_u32(bar)) : "memory");
}

__device__ __forceinline__ void tmap_fence(const CUtensorMap *map)
{
    asm volatile("fence.proxy.tensormap::generic.acquire.sys [%0], 128;" ::"l"(map) : "memory");
}

// same check with the map passed as a kernel parameter (the documented default)
__global__ void check_param(const __grid_constant__ CUtensorMap map, int bw, int bh, int x, int y, const uint16_t *plane,
                            int pitch_el, int *bad)
{
    extern __shared__ __align__(128) uint8_t pool[];
    __shared__ __align__(8) uint64_t bar;
    if (threadIdx.x == 0) {
        mbar_init(&bar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        mbar_expect_tx(&bar, bw * bh * 2);
        tma_load_2d(pool, &map, x, y, &bar);
    }
    __syncthreads();
    mbar_wait(&bar, 0);
    const uint16_t *s = (const uint16_t *)pool;
    for (int i = threadIdx.x; i < bw * bh; i += blockDim.x) {
        const int r = i / bw, c = i % bw;
        if (s[i] != plane[(size_t)(y + r) * pitch_el + x + c]) atomicAdd(bad, 1);
    }
}

// Each CTA: `iters` rounds; in each round `nbox` threads issue one box load each into their own
// smem slot; everyone waits on the mbarrier; checksum a word so the data is really consumed.
__global__ void __launch_bounds__(256) k(const CUtensorMap *maps, int cls, int boxbytes, int nbox, int iters, int W, int H,
                                          int bw, int bh, unsigned *out, int fence_every)
{
    extern __shared__ __align__(128) uint8_t pool[];
    __shared__ __align__(8) uint64_t bar;
    if (threadIdx.x == 0) {
        mbar_init(&bar, nbox);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    unsigned rng = blockIdx.x * 9781u + threadIdx.x * 6271u + 1u, acc = 0;
    for (int it = 0; it < iters; it++) {
        if (threadIdx.x < nbox) {
            rng = rng * 1664525u + 1013904223u;
            const int x = ((rng >> 8) % (W - bw)) & ~7, y = (rng >> 4) % (H - bh);
            if (fence_every || it == 0) tmap_fence(maps + cls);
            mbar_expect_tx(&bar, boxbytes);
            tma_load_2d(pool + (size_t)threadIdx.x * boxbytes, maps + cls, x, y, &bar);
        }
        mbar_wait(&bar, it & 1);
        acc += ((const unsigned *)pool)[(threadIdx.x * 37 + it) % (nbox * boxbytes / 4)];
        __syncthreads();
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}

// correctness: one box at (x, y) compared element by element with plain loads
__global__ void check(const CUtensorMap *maps, int cls, int bw, int bh, int x, int y, const uint16_t *plane, int pitch_el,
                      int *bad)
{
    extern __shared__ __align__(128) uint8_t pool[];
    __shared__ __align__(8) uint64_t bar;
    if (threadIdx.x == 0) {
        mbar_init(&bar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        tmap_fence(maps + cls);
        mbar_expect_tx(&bar, bw * bh * 2);
        tma_load_2d(pool, maps + cls, x, y, &bar);
    }
    __syncthreads();
    mbar_wait(&bar, 0);
    const uint16_t *s = (const uint16_t *)pool;
    for (int i = threadIdx.x; i < bw * bh; i += blockDim.x) {
        const int r = i / bw, c = i % bw;
        if (s[i] != plane[(size_t)(y + r) * pitch_el + x + c]) atomicAdd(bad, 1);
    }
}

int main()
{
    const int W = 4000, H = 2320, pitch = 8192;      // padded 4K luma plane, u16, pitch bytes
    uint16_t *plane;
    cudaMalloc(&plane, (size_t)pitch * H);
    std::vector<uint16_t> h((size_t)pitch / 2 * H);
    for (size_t i = 0; i < h.size(); i++) h[i] = (uint16_t)((i * 2654435761u) >> 20);
    cudaMemcpy(plane, h.data(), h.size() * 2, cudaMemcpyHostToDevice);

    EncodeTiled enc = nullptr;
    cudaDriverEntryPointQueryResult qr;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", (void **)&enc, cudaEnableDefault, &qr) != cudaSuccess || !enc) {
        printf("no cuTensorMapEncodeTiled\n");
        return 1;
    }
    const int boxes[][2] = {{16, 12}, {16, 16}, {24, 24}, {40, 40}, {72, 72}, {8, 8}, {16, 8}, {32, 8}, {32, 24}, {48, 16}, {48, 40}, {80, 8}, {80, 16}, {80, 24}, {80, 40}, {80, 72}, {24, 72}, {16, 72}};
    const int ncls = sizeof(boxes) / sizeof(boxes[0]);
    std::vector<CUtensorMap> maps(ncls);
    for (int i = 0; i < ncls; i++) {
        cuuint64_t gdim[2] = {(cuuint64_t)W, (cuuint64_t)H};
        cuuint64_t gstr[1] = {(cuuint64_t)pitch};
        cuuint32_t box[2] = {(cuuint32_t)boxes[i][0], (cuuint32_t)boxes[i][1]};
        cuuint32_t estr[2] = {1, 1};
        CUresult r = enc(&maps[i], CU_TENSOR_MAP_DATA_TYPE_UINT16, 2, plane, gdim, gstr, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                         CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS) { printf("encode %d failed: %d\n", i, (int)r); return 1; }
    }
    CUtensorMap *dmaps;
    cudaMalloc(&dmaps, sizeof(CUtensorMap) * ncls);
    cudaMemcpy(dmaps, maps.data(), sizeof(CUtensorMap) * ncls, cudaMemcpyHostToDevice);
    unsigned *out;
    cudaMalloc(&out, 148 * 8 * 256 * 4);
    int *bad;
    cudaMalloc(&bad, 4);
    cudaMemset(bad, 0, 4);
    cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);

    int hbad = -1;
    for (int i = 0; i < ncls; i++)
        check_param<<<1, 128, boxes[i][0] * boxes[i][1] * 2>>>(maps[i], boxes[i][0], boxes[i][1], 1000 + 8 * i, 777 + 3 * i, plane, pitch / 2, bad);
    cudaError_t e = cudaDeviceSynchronize();
    cudaMemcpy(&hbad, bad, 4, cudaMemcpyDeviceToHost);
    printf("param-map correctness: %d mismatching elements (%s)\n", hbad, cudaGetErrorString(e));
    if (e != cudaSuccess) return 1;
    for (int i = 0; i < ncls; i++)
        check<<<1, 128, boxes[i][0] * boxes[i][1] * 2>>>(dmaps, i, boxes[i][0], boxes[i][1], 1000 + 8 * i, 777 + 3 * i, plane, pitch / 2, bad);
    e = cudaDeviceSynchronize();
    cudaMemcpy(&hbad, bad, 4, cudaMemcpyDeviceToHost);
    printf("global-map + fence correctness: %d mismatching elements (%s)\n", hbad, cudaGetErrorString(e));
    if (e != cudaSuccess) return 1;
    for (int i = 0; i < ncls; i++) {
        cudaMemset(bad, 0, 4);
        check<<<1, 128, boxes[i][0] * boxes[i][1] * 2>>>(dmaps, i, boxes[i][0], boxes[i][1], 1000 + 8 * i, 777 + 3 * i, plane, pitch / 2, bad);
        e = cudaDeviceSynchronize();
        cudaMemcpy(&hbad, bad, 4, cudaMemcpyDeviceToHost);
        printf("box %dx%d: %d bad (%s)\n", boxes[i][0], boxes[i][1], hbad, cudaGetErrorString(e));
        if (e != cudaSuccess) return 1;
    }
    return 0;

    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    for (int fence_every = 0; fence_every < 2; fence_every++)
    for (int occ = 1; occ <= 4; occ *= 4)
        for (int i = 0; i < ncls; i++) {
            const int bw = boxes[i][0], bh = boxes[i][1], bytes = bw * bh * 2;
            int nbox = 32 * 1024 / bytes;
            if (nbox > 64) nbox = 64;
            if (nbox < 1) nbox = 1;
            const int iters = 200, ctas = 148 * occ;
            const size_t smem = (size_t)nbox * bytes;
            k<<<ctas, 256, smem>>>(dmaps, i, bytes, nbox, iters, W, H, bw, bh, out, fence_every);
            cudaDeviceSynchronize();
            cudaEventRecord(e0);
            k<<<ctas, 256, smem>>>(dmaps, i, bytes, nbox, iters, W, H, bw, bh, out, fence_every);
            cudaEventRecord(e1);
            cudaEventSynchronize(e1);
            float ms;
            cudaEventElapsedTime(&ms, e0, e1);
            const double nb = (double)ctas * nbox * iters;
            if (cudaDeviceSynchronize() != cudaSuccess) { printf("fault: %s\n", cudaGetErrorString(cudaGetLastError())); return 1; }
            printf("fence_every %d ctas/SM %d box %2dx%2d u16 (%5d B) x%2d per round: %7.3f ms  %7.1f Mbox/s  %6.1f clk/box/SM  %7.1f GB/s  round %.2f us  (%s)\n",
                   fence_every, occ, bw, bh, bytes, nbox, ms, nb / ms / 1e3, ms * 1e-3 * 1.9e9 / (nb / 148), nb * bytes / ms / 1e6,
                   ms * 1e3 / iters, cudaGetErrorString(cudaGetLastError()));
        }
    return 0;
}
