// ubench_mc_mma.cu — the HEVC luma qpel interpolation (8-tap horizontal, then 8-tap vertical, hevcdsp_template.c
// put_hevc_qpel_hv + put_unweighted_pred) of one 16x16 block as two CHAINED integer tensor-core products, measured against
// the question DESIGN.md section 6 leaves open: can mc_kernel's multiply-adds leave the issue-bound integer pipes?
//
//   pass 1   T^T (16 columns x 8 rows)  =  H (16 x 32, taps as a banded matrix, s8)  x  X^T (32 columns x 8 rows, samples)
//   pass 2   out (16 rows x 8 columns)  =  V (16 x 32, taps, s8)                     x  T   (24 rows  x 8 columns)
// mma.sync.m16n8k16 s8/u8 -> s32.  16-bit samples and the 16-bit intermediate enter as two byte planes (hi * 256 + lo).
// The accumulator layout of pass 1 IS a B-fragment layout of pass 2 once the contraction index of pass 2 is permuted
// (K slot 4t + b  <->  T row {2t, 2t+1, 8+2t, 8+2t+1}): no shuffle, no shared memory between the passes.  The window is
// read with aligned 8-byte loads; its 0..3-sample misalignment is folded into the tap matrix (table per phase).
// One warp per block.  Checked bit for bit against a scalar kernel; timed with CUDA events.
// build: nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo -o ubench_mc_mma tools/ubench_mc_mma.cu
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("%s -> %s\n", #x, cudaGetErrorString(e_)); exit(1); } } while (0)

static const int8_t TAPS[4][8] = {{0, 0, 0, 64, 0, 0, 0, 0}, {-1, 4, -10, 58, 17, -5, 1, 0}, {-1, 4, -11, 40, 40, -11, 4, -1}, {0, 1, -5, 17, 58, -10, 4, -1}};
__constant__ int8_t c_taps[4][8];

constexpr int BD = 10, W = 3840, H = 2160, PAD = 64, PITCH = W + 2 * PAD;       // samples
struct Blk { int16_t dx, dy; uint8_t mx, my; };                                  // integer displacement, quarter-pel phases

// ------------------------------------------------------------------ scalar reference
__global__ void fir_ref_kernel(const uint16_t *ref, const Blk *blk, uint16_t *out)
{
    const int bx = blockIdx.x, by = blockIdx.y, x = threadIdx.x, y = threadIdx.y;
    const Blk b = blk[by * (W / 16) + bx];
    const uint16_t *src = ref + (size_t)(PAD + by * 16 + b.dy) * PITCH + PAD + bx * 16 + b.dx;
    int v = 0;
    for (int j = 0; j < 8; j++) {
        int t = 0;
        for (int i = 0; i < 8; i++) t += c_taps[b.mx][i] * src[(y + j - 3) * PITCH + x + i - 3];
        v += c_taps[b.my][j] * (t >> (BD - 8));
    }
    v >>= 6;
    v = (v + (1 << (13 - BD))) >> (14 - BD);
    out[(size_t)(by * 16 + y) * W + bx * 16 + x] = (uint16_t)min(max(v, 0), (1 << BD) - 1);
}

// ------------------------------------------------------------------ tensor-core version
template <bool SA, bool SB>
__device__ __forceinline__ void imma(int (&d)[4], uint32_t a0, uint32_t a1, uint32_t b)
{
    if (SB)
        asm volatile("mma.sync.aligned.m16n8k16.row.col.s32.s8.s8.s32 {%0,%1,%2,%3}, {%4,%5}, {%6}, {%0,%1,%2,%3};"
                     : "+r"(d[0]), "+r"(d[1]), "+r"(d[2]), "+r"(d[3]) : "r"(a0), "r"(a1), "r"(b));
    else
        asm volatile("mma.sync.aligned.m16n8k16.row.col.s32.s8.u8.s32 {%0,%1,%2,%3}, {%4,%5}, {%6}, {%0,%1,%2,%3};"
                     : "+r"(d[0]), "+r"(d[1]), "+r"(d[2]), "+r"(d[3]) : "r"(a0), "r"(a1), "r"(b));
}

// A fragments: tabH[mx][phase][kstep][lane] = (a0, a1), tabV[my][kstep][lane] = (a0, a1)
__global__ void __launch_bounds__(256) fir_mma_kernel(const uint16_t *__restrict__ ref, const Blk *__restrict__ blk, uint16_t *__restrict__ out,
                                                      const uint2 *__restrict__ tabH, const uint2 *__restrict__ tabV, int nblk)
{
    const int lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
    const int id = blockIdx.x * 8 + (threadIdx.x >> 5);
    if (id >= nblk) return;
    const int bx = id % (W / 16), by = id / (W / 16);
    const Blk b = blk[id];
    const int xs = PAD + bx * 16 + b.dx - 3, ys = PAD + by * 16 + b.dy - 3, xa = xs & ~3, ph = xs - xa;
    const uint2 h0 = __ldg(&tabH[((b.mx * 4 + ph) * 2 + 0) * 32 + lane]), h1 = __ldg(&tabH[((b.mx * 4 + ph) * 2 + 1) * 32 + lane]);
    const uint2 v0 = __ldg(&tabV[(b.my * 2 + 0) * 32 + lane]), v1 = __ldg(&tabV[(b.my * 2 + 1) * 32 + lane]);
    // ---- the window: 24 rows x 32 columns, 8 bytes per thread and (row tile, K step), all in flight
    const uint16_t *p = ref + (size_t)(ys + g) * PITCH + xa + 4 * t;
    uint2 w[3][2];
#pragma unroll
    for (int nt = 0; nt < 3; nt++)
#pragma unroll
        for (int s = 0; s < 2; s++) w[nt][s] = __ldg((const uint2 *)(p + (size_t)(8 * nt) * PITCH + 16 * s));
    // ---- pass 1: T^T tile nt = columns g / g + 8, rows 8 nt + 2t, + 1
    int T[3][4];
#pragma unroll
    for (int nt = 0; nt < 3; nt++) {
        int acc[4] = {0, 0, 0, 0};
        imma<true, false>(acc, h0.x, h0.y, __byte_perm(w[nt][0].x, w[nt][0].y, 0x7531));
        imma<true, false>(acc, h1.x, h1.y, __byte_perm(w[nt][1].x, w[nt][1].y, 0x7531));
#pragma unroll
        for (int k = 0; k < 4; k++) acc[k] <<= 8;
        imma<true, false>(acc, h0.x, h0.y, __byte_perm(w[nt][0].x, w[nt][0].y, 0x6420));
        imma<true, false>(acc, h1.x, h1.y, __byte_perm(w[nt][1].x, w[nt][1].y, 0x6420));
#pragma unroll
        for (int k = 0; k < 4; k++) T[nt][k] = acc[k] >> (BD - 8);
    }
    // ---- pass 2: per column tile (c = 0: columns g, c = 1: columns g + 8) B = T in the permuted K order
    uint32_t res[2][2];                                  // [column tile][row g / g + 8] = two packed samples (columns 2t, 2t + 1)... see below
#pragma unroll
    for (int c = 0; c < 2; c++) {
        const int a = T[0][2 * c], bb = T[0][2 * c + 1], cc = T[1][2 * c], dd = T[1][2 * c + 1], e = T[2][2 * c], f = T[2][2 * c + 1];
        // K step 0: rows 2t, 2t+1, 8+2t, 8+2t+1; K step 1: rows 16+2t, 16+2t+1, -, -
        const uint32_t lo0 = __byte_perm(__byte_perm(a, bb, 0x0040), __byte_perm(cc, dd, 0x0040), 0x5410);
        const uint32_t hi0 = __byte_perm(__byte_perm(a, bb, 0x0051), __byte_perm(cc, dd, 0x0051), 0x5410);
        const uint32_t lo1 = __byte_perm(e, f, 0x0040) & 0xffffu, hi1 = __byte_perm(e, f, 0x0051) & 0xffffu;
        int acc[4] = {0, 0, 0, 0};
        imma<true, true>(acc, v0.x, v0.y, hi0);
        imma<true, true>(acc, v1.x, v1.y, hi1);
#pragma unroll
        for (int k = 0; k < 4; k++) acc[k] <<= 8;
        imma<true, false>(acc, v0.x, v0.y, lo0);
        imma<true, false>(acc, v1.x, v1.y, lo1);
        // acc[0], acc[1]: row g, columns (2t, 2t+1) OF THE N TILE = block columns ... N index = column within tile c
#pragma unroll
        for (int k = 0; k < 4; k++) {
            int v = acc[k] >> 6;
            v = (v + (1 << (13 - BD))) >> (14 - BD);
            acc[k] = min(max(v, 0), (1 << BD) - 1);
        }
        res[c][0] = (uint32_t)acc[0] | ((uint32_t)acc[1] << 16);
        res[c][1] = (uint32_t)acc[2] | ((uint32_t)acc[3] << 16);
    }
    // pass 2's N index is the block column inside the tile: n = 2t, 2t + 1 -> columns 8 c + 2t, rows g and g + 8
    uint16_t *o = out + (size_t)(by * 16 + g) * W + bx * 16 + 2 * t;
    *(uint32_t *)(o) = res[0][0];
    *(uint32_t *)(o + 8) = res[1][0];
    *(uint32_t *)(o + (size_t)8 * W) = res[0][1];
    *(uint32_t *)(o + (size_t)8 * W + 8) = res[1][1];
}

int main()
{
    const int nbx = W / 16, nby = H / 16, nblk = nbx * nby;
    std::vector<uint16_t> ref((size_t)PITCH * (H + 2 * PAD));
    uint64_t s = 88172645463325252ull;
    auto rnd = [&]() { s ^= s << 13; s ^= s >> 7; s ^= s << 17; return (uint32_t)(s >> 11); };
    for (auto &v : ref) v = rnd() & 1023;
    std::vector<Blk> blk(nblk);
    for (auto &b : blk) { b.dx = (int16_t)(rnd() % 41) - 20; b.dy = (int16_t)(rnd() % 41) - 20; b.mx = 1 + rnd() % 3; b.my = 1 + rnd() % 3; }
    // ---- A fragment tables
    std::vector<uint2> tabH(4 * 4 * 2 * 32), tabV(4 * 2 * 32);
    auto tap = [&](int m, int i) -> int { return (i >= 0 && i < 8) ? TAPS[m][i] : 0; };
    for (int m = 0; m < 4; m++)
        for (int ph = 0; ph < 4; ph++)
            for (int ks = 0; ks < 2; ks++)
                for (int lane = 0; lane < 32; lane++) {
                    const int g = lane >> 2, t = lane & 3;
                    uint32_t a[2] = {0, 0};
                    for (int r = 0; r < 2; r++)
                        for (int b = 0; b < 4; b++) {
                            const int x = g + 8 * r, j = 16 * ks + 4 * t + b;
                            a[r] |= (uint32_t)(uint8_t)tap(m, j - ph - x) << (8 * b);
                        }
                    tabH[((m * 4 + ph) * 2 + ks) * 32 + lane] = make_uint2(a[0], a[1]);
                }
    for (int m = 0; m < 4; m++)
        for (int ks = 0; ks < 2; ks++)
            for (int lane = 0; lane < 32; lane++) {
                const int g = lane >> 2, t = lane & 3;
                uint32_t a[2] = {0, 0};
                for (int r = 0; r < 2; r++)
                    for (int b = 0; b < 4; b++) {
                        const int y = g + 8 * r;
                        const int row = ks == 0 ? (b < 2 ? 2 * t + b : 8 + 2 * t + (b - 2)) : (b < 2 ? 16 + 2 * t + b : -100);
                        a[r] |= (uint32_t)(uint8_t)(row < 0 ? 0 : tap(m, row - y)) << (8 * b);
                    }
                tabV[(m * 2 + ks) * 32 + lane] = make_uint2(a[0], a[1]);
            }
    uint16_t *d_ref, *d_out, *d_out2; Blk *d_blk; uint2 *d_tabH, *d_tabV;
    CK(cudaMalloc(&d_ref, ref.size() * 2)); CK(cudaMalloc(&d_out, (size_t)W * H * 2)); CK(cudaMalloc(&d_out2, (size_t)W * H * 2));
    CK(cudaMalloc(&d_blk, blk.size() * sizeof(Blk))); CK(cudaMalloc(&d_tabH, tabH.size() * 8)); CK(cudaMalloc(&d_tabV, tabV.size() * 8));
    CK(cudaMemcpy(d_ref, ref.data(), ref.size() * 2, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(d_blk, blk.data(), blk.size() * sizeof(Blk), cudaMemcpyHostToDevice));
    CK(cudaMemcpy(d_tabH, tabH.data(), tabH.size() * 8, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(d_tabV, tabV.data(), tabV.size() * 8, cudaMemcpyHostToDevice));
    CK(cudaMemcpyToSymbol(c_taps, TAPS, sizeof(TAPS)));
    fir_ref_kernel<<<dim3(nbx, nby), dim3(16, 16)>>>(d_ref, d_blk, d_out2);
    fir_mma_kernel<<<(nblk + 7) / 8, 256>>>(d_ref, d_blk, d_out, d_tabH, d_tabV, nblk);
    CK(cudaDeviceSynchronize());
    std::vector<uint16_t> o1((size_t)W * H), o2((size_t)W * H);
    CK(cudaMemcpy(o1.data(), d_out, o1.size() * 2, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(o2.data(), d_out2, o2.size() * 2, cudaMemcpyDeviceToHost));
    size_t bad = 0;
    for (size_t i = 0; i < o1.size(); i++) bad += o1[i] != o2[i];
    printf("mma vs scalar: %zu of %zu samples differ\n", bad, o1.size());
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    for (int i = 0; i < 5; i++) fir_mma_kernel<<<(nblk + 7) / 8, 256>>>(d_ref, d_blk, d_out, d_tabH, d_tabV, nblk);
    CK(cudaEventRecord(e0));
    const int N = 50;
    for (int i = 0; i < N; i++) fir_mma_kernel<<<(nblk + 7) / 8, 256>>>(d_ref, d_blk, d_out, d_tabH, d_tabV, nblk);
    CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
    float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
    ms /= N;
    const double bytes = (double)W * H * 2 * 2;          // every sample read once + written once (uni-prediction, one list)
    printf("fir_mma_kernel: %.4f ms per 3840x2160 luma plane (uni, hv), %d blocks of 16x16, %.0f GB/s algorithmic\n", ms, nblk, bytes / ms / 1e6);
    return bad != 0;
}
