// picture.cu — whole-picture helpers around the two hot kernels: the replicated halo of a
// finished DPB picture (what ff_emulated_edge_mc, videodsp_template.c:22-100, builds per
// block in the reference), a plane copy for partial-stage launches, and the L2 flush the
// benchmark uses between timed frames.
#include "ohcu_internal.h"

namespace ohcu {

template <int BD>
__global__ void copy_picture_kernel(uint8_t *dst, int dpitch, const uint8_t *src, int spitch, int wbytes, int h)
{
    const int x = (blockIdx.x * blockDim.x + threadIdx.x) * 16;
    const int y = blockIdx.y;
    if (x >= wbytes || y >= h) return;
    if (x + 16 <= wbytes)
        *(uint4 *)(dst + (size_t)y * dpitch + x) = *(const uint4 *)(src + (size_t)y * spitch + x);
    else
        for (int i = x; i < wbytes; i++) dst[(size_t)y * dpitch + i] = src[(size_t)y * spitch + i];
}

// Replicated halo of a DPB picture (PAD_X / PAD_Y, ohcu_internal.h).  Equivalent of
// ff_emulated_edge_mc (videodsp_template.c:22-100) applied once to the whole picture.
// 16-byte stores throughout (plane origins, PAD_X samples and the pitches are multiples of
// 16 bytes).  Work items of one plane: first the rows above / below the picture, one CTA each
// (the border row repeated with its own margins), then the picture's rows, one WARP each
// (left and right margin only).
struct PadPlane { uint8_t *p; int pitch, pw, ph; };
struct PadArgs { PadPlane pl[3]; int nplanes; int cta0[4]; };     // cta0[c]: first CTA of plane c

template <typename pixel>
__device__ __forceinline__ uint4 splat16(pixel v)
{
    uint32_t w = sizeof(pixel) == 2 ? (uint32_t)v * 0x00010001u : (uint32_t)v * 0x01010101u;
    return make_uint4(w, w, w, w);
}

template <typename pixel>
__global__ void __launch_bounds__(256) pad_kernel(const PadArgs a)
{
    constexpr int P = sizeof(pixel), MARGIN16 = PAD_X * P / 16;          // 16-byte stores per margin
    int c = 0;
    while (c + 1 < a.nplanes && (int)blockIdx.x >= a.cta0[c + 1]) c++;
    const PadPlane pl = a.pl[c];
    const int cta = blockIdx.x - a.cta0[c];
    const int tid = threadIdx.x;
    if (cta < 2 * PAD_Y) {
        // a row above (cta < PAD_Y) or below the picture
        const int y = cta < PAD_Y ? cta - PAD_Y : pl.ph + cta - PAD_Y;
        const pixel *src = (const pixel *)(pl.p + (ptrdiff_t)(cta < PAD_Y ? 0 : pl.ph - 1) * pl.pitch);
        uint8_t *dst = pl.p + (ptrdiff_t)y * pl.pitch;
        const int wbytes = pl.pw * P, mid16 = wbytes / 16;
        const uint4 l = splat16<pixel>(src[0]), r = splat16<pixel>(src[pl.pw - 1]);
        for (int i = tid; i < MARGIN16; i += 256) *(uint4 *)(dst - PAD_X * P + 16 * i) = l;
        for (int i = tid; i < mid16; i += 256) *(uint4 *)(dst + 16 * i) = *(const uint4 *)((const uint8_t *)src + 16 * i);
        // a picture row that does not end on a 16-byte boundary: the tail and the right margin sample by sample
        if (wbytes & 15) {
            for (int x = mid16 * (16 / P) + tid; x < pl.pw + PAD_X; x += 256) ((pixel *)dst)[x] = src[min(x, pl.pw - 1)];
        } else {
            for (int i = tid; i < MARGIN16; i += 256) *(uint4 *)(dst + wbytes + 16 * i) = r;
        }
    } else {
        // eight picture rows, one per warp: lanes store the two margins
        const int y = (cta - 2 * PAD_Y) * 8 + (tid >> 5), lane = tid & 31;
        if (y >= pl.ph) return;
        uint8_t *row = pl.p + (ptrdiff_t)y * pl.pitch;
        const int wbytes = pl.pw * P;
        const pixel lv = ((const pixel *)row)[0], rv = ((const pixel *)row)[pl.pw - 1];
        for (int i = lane; i < MARGIN16; i += 32) *(uint4 *)(row - PAD_X * P + 16 * i) = splat16<pixel>(lv);
        if (wbytes & 15) {
            for (int x = lane; x < PAD_X; x += 32) ((pixel *)row)[pl.pw + x] = rv;
        } else {
            for (int i = lane; i < MARGIN16; i += 32) *(uint4 *)(row + wbytes + 16 * i) = splat16<pixel>(rv);
        }
    }
}

int launch_pad_picture(const FrameDev &f, const PlaneSet &pic, cudaStream_t st)
{
    static_assert((PAD_X % 16) == 0, "margins are whole 16-byte stores for 8- and 16-bit samples");
    PadArgs a{};
    a.nplanes = f.nplanes;
    int n = 0;
    for (int c = 0; c < f.nplanes; c++) {
        a.pl[c].p = pic.p[c]; a.pl[c].pitch = pic.pitch[c]; a.pl[c].pw = f.pw[c]; a.pl[c].ph = f.ph[c];
        a.cta0[c] = n;
        n += 2 * PAD_Y + (f.ph[c] + 7) / 8;
    }
    a.cta0[f.nplanes] = n;
    if (f.bd > 8) pad_kernel<uint16_t><<<n, 256, 0, st>>>(a);
    else          pad_kernel<uint8_t><<<n, 256, 0, st>>>(a);
    return 1;
}

int launch_copy_picture(const FrameDev &f, const PlaneSet &dst, const PlaneSet &src, cudaStream_t st)
{
    const int P = f.bd > 8 ? 2 : 1;
    for (int c = 0; c < f.nplanes; c++) {
        const int wbytes = f.pw[c] * P;
        dim3 grid((wbytes / 16 + 128) / 128, f.ph[c]);
        copy_picture_kernel<8><<<grid, 128, 0, st>>>(dst.p[c], dst.pitch[c], src.p[c], src.pitch[c], wbytes, f.ph[c]);
    }
    return f.nplanes;
}

// Cropped copy-out (ohcu_slot_output_cropped): the window of every plane, packed without row
// gaps into one device buffer that crosses the bus as linear copies.  What the reference does
// with a pointer offset (hevc_refs.c:335-341) and one memcpy per row (openhevc.c:722-735).
// One CTA per 4 rows of a plane, a warp per row; 16-byte loads and stores when the source
// window and the row length allow it, else sample by sample.
struct CropPlane { const uint8_t *src; uint8_t *dst; int pitch, wbytes, rows; };
struct CropArgs { CropPlane pl[3]; int nplanes; int cta0[4]; };

__global__ void __launch_bounds__(128) crop_pack_kernel(const CropArgs a)
{
    int c = 0;
    while (c + 1 < a.nplanes && (int)blockIdx.x >= a.cta0[c + 1]) c++;
    const CropPlane pl = a.pl[c];
    const int y = (blockIdx.x - a.cta0[c]) * 4 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (y >= pl.rows) return;
    const uint8_t *s = pl.src + (size_t)y * pl.pitch;
    uint8_t *d = pl.dst + (size_t)y * pl.wbytes;
    if ((((uintptr_t)s | (uintptr_t)d | (uintptr_t)pl.wbytes) & 15) == 0) {
        for (int i = lane; i < pl.wbytes / 16; i += 32) ((uint4 *)d)[i] = __ldcs((const uint4 *)s + i);
    } else {
        // bytes up to the first 4-byte boundary of the destination, whole words after that
        // (funnel-shifted from the two source words they straddle), then the tail
        int i = min((int)(-(uintptr_t)d & 3), pl.wbytes);
        if (lane < i) d[lane] = s[lane];
        const int nw = (pl.wbytes - i) >> 2;
        const uint8_t *sw = s + i;
        const int sh = (int)((uintptr_t)sw & 3) * 8;
        const uint32_t *sa = (const uint32_t *)(sw - ((uintptr_t)sw & 3));
        for (int k = lane; k < nw; k += 32) {
            const uint32_t lo = sa[k], hi = sh ? sa[k + 1] : 0;
            ((uint32_t *)(d + i))[k] = __funnelshift_r(lo, hi, sh);
        }
        const int t0 = i + 4 * nw;
        if (t0 + lane < pl.wbytes) d[t0 + lane] = s[t0 + lane];
    }
}

int launch_crop_pack(const PlaneSet &pic, int nplanes, int P, const int left[3], const int top[3], const int w[3],
                     const int h[3], uint8_t *const dst[3], cudaStream_t st)
{
    CropArgs a{};
    a.nplanes = nplanes;
    int n = 0;
    for (int c = 0; c < nplanes; c++) {
        a.pl[c].src = pic.p[c] + (ptrdiff_t)top[c] * pic.pitch[c] + (ptrdiff_t)left[c] * P;
        a.pl[c].dst = dst[c]; a.pl[c].pitch = pic.pitch[c]; a.pl[c].wbytes = w[c] * P; a.pl[c].rows = h[c];
        a.cta0[c] = n;
        n += (h[c] + 3) / 4;
    }
    a.cta0[nplanes] = n;
    if (n) crop_pack_kernel<<<n, 128, 0, st>>>(a);
    return n ? 1 : 0;
}

__global__ void l2_flush_kernel(uint4 *buf, size_t n)
{
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (; i < n; i += stride) { uint4 v = buf[i]; v.x += 1; buf[i] = v; }
}

int launch_l2_flush(void *buf, size_t bytes, cudaStream_t st)
{
    l2_flush_kernel<<<148 * 8, 256, 0, st>>>((uint4 *)buf, bytes / 16);
    return 1;
}

}  // namespace ohcu

// ================================================================= picture hash (verify_md5, hevcdec.c:4035-4098)
// The SEI picture hash is one RFC 1321 MD5 per plane over its rows (w << pixel_shift bytes each, little endian).
// MD5 is a serial chain of 64-byte blocks, so a plane is ONE thread's work: a warp per plane stages 512 bytes (eight
// blocks) per step with 16-byte loads, lane 0 runs the chain out of shared memory.  This exists so that a conformance
// run can verify pictures without moving them across the bus, and so that the cost is a measured number
// (DESIGN.md section 4.6): a chain step is ~5 dependent ALU operations, nothing inside a plane runs in parallel.
namespace ohcu {

__constant__ uint32_t c_md5_k[64] = {
    0xd76aa478, 0xe8c7b756, 0x242070db, 0xc1bdceee, 0xf57c0faf, 0x4787c62a, 0xa8304613, 0xfd469501, 0x698098d8, 0x8b44f7af,
    0xffff5bb1, 0x895cd7be, 0x6b901122, 0xfd987193, 0xa679438e, 0x49b40821, 0xf61e2562, 0xc040b340, 0x265e5a51, 0xe9b6c7aa,
    0xd62f105d, 0x02441453, 0xd8a1e681, 0xe7d3fbc8, 0x21e1cde6, 0xc33707d6, 0xf4d50d87, 0x455a14ed, 0xa9e3e905, 0xfcefa3f8,
    0x676f02d9, 0x8d2a4c8a, 0xfffa3942, 0x8771f681, 0x6d9d6122, 0xfde5380c, 0xa4beea44, 0x4bdecfa9, 0xf6bb4b60, 0xbebfbc70,
    0x289b7ec6, 0xeaa127fa, 0xd4ef3085, 0x04881d05, 0xd9d4d039, 0xe6db99e5, 0x1fa27cf8, 0xc4ac5665, 0xf4292244, 0x432aff97,
    0xab9423a7, 0xfc93a039, 0x655b59c3, 0x8f0ccc92, 0xffeff47d, 0x85845dd1, 0x6fa87e4f, 0xfe2ce6e0, 0xa3014314, 0x4e0811a1,
    0xf7537e82, 0xbd3af235, 0x2ad7d2bb, 0xeb86d391};
__constant__ uint8_t c_md5_s[64] = {7, 12, 17, 22, 7, 12, 17, 22, 7, 12, 17, 22, 7, 12, 17, 22, 5, 9, 14, 20, 5, 9, 14, 20, 5, 9, 14, 20, 5, 9, 14, 20,
                                    4, 11, 16, 23, 4, 11, 16, 23, 4, 11, 16, 23, 4, 11, 16, 23, 6, 10, 15, 21, 6, 10, 15, 21, 6, 10, 15, 21, 6, 10, 15, 21};

__device__ __forceinline__ void md5_block(uint32_t (&h)[4], const uint32_t *m)
{
    uint32_t a = h[0], b = h[1], c = h[2], d = h[3];
#pragma unroll
    for (int i = 0; i < 64; i++) {
        uint32_t f;
        int g;
        if (i < 16) { f = (b & c) | (~b & d); g = i; }
        else if (i < 32) { f = (d & b) | (~d & c); g = (5 * i + 1) & 15; }
        else if (i < 48) { f = b ^ c ^ d; g = (3 * i + 5) & 15; }
        else { f = c ^ (b | ~d); g = (7 * i) & 15; }
        f += a + c_md5_k[i] + m[g];
        a = d; d = c; c = b;
        b += __funnelshift_l(f, f, c_md5_s[i]);
    }
    h[0] += a; h[1] += b; h[2] += c; h[3] += d;
}

struct Md5Args { const uint8_t *p[3]; int pitch[3], wbytes[3], h[3]; uint8_t *out; };

// one warp per plane (blockIdx.x)
__global__ void __launch_bounds__(32) md5_kernel(const Md5Args a)
{
    __shared__ __align__(16) uint32_t buf[2][128];             // two stages of 512 bytes
    const int pl = blockIdx.x, lane = threadIdx.x;
    const uint8_t *base = a.p[pl];
    const int pitch = a.pitch[pl], wb = a.wbytes[pl];
    const unsigned long long total = (unsigned long long)wb * a.h[pl];
    uint32_t h[4] = {0x67452301u, 0xefcdab89u, 0x98badcfeu, 0x10325476u};
    // message bytes [off, off + 16) of the plane: rows are multiples of 4 bytes, a 16-byte group may straddle a row end
    auto load16 = [&](unsigned long long off, uint32_t *dst) {
        int row = (int)(off / wb), col = (int)(off - (unsigned long long)row * wb);
#pragma unroll
        for (int k = 0; k < 4; k++) {
            uint32_t v = 0;
            if (off + 4 * k < total) v = *(const uint32_t *)(base + (size_t)row * pitch + col);
            dst[k] = v;
            col += 4;
            if (col >= wb) { col = 0; row++; }
        }
    };
    const unsigned long long nfull = total / 64;               // whole blocks
    unsigned long long blk = 0;
    int stage = 0;
    if (nfull) load16((unsigned long long)lane * 16, &buf[0][lane * 4]);
    __syncwarp();
    while (blk < nfull) {
        const int nb = (int)min((unsigned long long)8, nfull - blk);
        // prefetch the next eight blocks into the other stage while lane 0 chews on these
        if (blk + 8 < nfull) load16((blk + 8) * 64 + (unsigned long long)lane * 16, &buf[stage ^ 1][lane * 4]);
        if (lane == 0)
            for (int i = 0; i < nb; i++) md5_block(h, &buf[stage][16 * i]);
        __syncwarp();
        blk += nb;
        stage ^= 1;
    }
    if (lane == 0) {
        // the tail: remaining bytes, 0x80, zeros, the bit length
        uint32_t m[32];
        const int rem = (int)(total - nfull * 64);
        for (int i = 0; i < 32; i++) m[i] = 0;
        for (int i = 0; i < rem; i += 4) {
            const unsigned long long off = nfull * 64 + i;
            const int row = (int)(off / wb), col = (int)(off - (unsigned long long)row * wb);
            m[i >> 2] = *(const uint32_t *)(base + (size_t)row * pitch + col);
        }
        m[rem >> 2] = 0x80u;                                    // rem is a multiple of 4
        const int words = rem + 9 > 64 ? 32 : 16;
        m[words - 2] = (uint32_t)(total << 3);
        m[words - 1] = (uint32_t)(total >> 29);
        md5_block(h, m);
        if (words == 32) md5_block(h, m + 16);
        for (int i = 0; i < 4; i++) ((uint32_t *)(a.out + 16 * pl))[i] = h[i];
    }
}

int launch_md5(const PlaneSet &pic, int nplanes, int P, const int pw[3], const int ph[3], uint8_t *out48, cudaStream_t st)
{
    Md5Args a{};
    for (int i = 0; i < nplanes; i++) { a.p[i] = pic.p[i]; a.pitch[i] = pic.pitch[i]; a.wbytes[i] = pw[i] * P; a.h[i] = ph[i]; }
    a.out = out48;
    md5_kernel<<<nplanes, 32, 0, st>>>(a);
    return 1;
}

}  // namespace ohcu
