// Batched lidar scan -> bird's-eye bitmap rasteriser for sm_100a.
//
// Replaces weap_util.lidar.lidar_to_bitmap (weap_util/weap_util/lidar.py:4-154; copies in src/bitmap.py and
// src/SAL.py:274-385), which this fork's RL loop calls on every step's scan (src/SAL.py:76-77,119-120).  The reference
// draws with OpenCV (fillPoly / polylines / line / rectangle); the output must match it byte for byte, so the
// kernel evaluates the same integer algorithms (the test-side CPU restatement pins them against cv2):
//   * cv::clipLine         Cohen-Sutherland with truncating double-precision intersections
//   * cv::LineIterator     8-connected, left to right; closed form of the error recurrence per pixel
//   * cv::fillPoly         boundary lines + scan-line fill of 16.16 fixed-point edges.  Instead of the sequential
//                          active-edge list, every edge toggles one parity bit per row it crosses; a row's filled
//                          pixels are (exclusive prefix XOR of the toggles) | (edges exactly on a pixel corner).
// One CTA per image.  The image is built as bit masks in shared memory (3 x H x W/32 words) and expanded to bytes
// only in the final, fully coalesced 16-byte stores: the kernel is bound by the HBM write of H*W bytes per image
// (64 KB at 256 x 256; 8 KB with the packed 1-bit output).
// Compiled with -fmad=false: the pixel position is rint(c + (s * d) * cos) in the reference's rounding order.
#include <cuda_runtime.h>
#include <stdint.h>
#include "f110_bitmap_dev.h"

#define XY_SHIFT 16
#define XY_ONE (1 << XY_SHIFT)

// cv::clipLine(Size(w, h), p1, p2); returns false when nothing of the segment is inside
static __device__ __forceinline__ bool clip_line(int w, int h, long long &x1, long long &y1, long long &x2, long long &y2)
{
    const long long right = w - 1, bottom = h - 1;
    int c1 = (x1 < 0) + (x1 > right) * 2 + (y1 < 0) * 4 + (y1 > bottom) * 8;
    int c2 = (x2 < 0) + (x2 > right) * 2 + (y2 < 0) * 4 + (y2 > bottom) * 8;
    if ((c1 & c2) == 0 && (c1 | c2) != 0) {
        long long a;
        if (c1 & 12) {
            a = c1 < 8 ? 0 : bottom;
            x1 += (long long)((double)(a - y1) * (double)(x2 - x1) / (double)(y2 - y1));
            y1 = a;
            c1 = (x1 < 0) + (x1 > right) * 2;
        }
        if (c2 & 12) {
            a = c2 < 8 ? 0 : bottom;
            x2 += (long long)((double)(a - y2) * (double)(x2 - x1) / (double)(y2 - y1));
            y2 = a;
            c2 = (x2 < 0) + (x2 > right) * 2;
        }
        if ((c1 & c2) == 0 && (c1 | c2) != 0) {
            if (c1) {
                a = c1 == 1 ? 0 : right;
                y1 += (long long)((double)(a - x1) * (double)(y2 - y1) / (double)(x2 - x1));
                x1 = a;
                c1 = 0;
            }
            if (c2) {
                a = c2 == 1 ? 0 : right;
                y2 += (long long)((double)(a - x2) * (double)(y2 - y1) / (double)(x2 - x1));
                x2 = a;
                c2 = 0;
            }
        }
    }
    return (c1 | c2) == 0;
}

static __device__ __forceinline__ void set_px(unsigned *mask, int wpr, int x, int y)
{
    atomicOr(mask + y * wpr + (x >> 5), 1u << (x & 31));
}

// cv::Line(img, p1, p2, color, LINE_8), thickness 1
static __device__ void draw_line(unsigned *mask, int wpr, int w, int h, int ax, int ay, int bx, int by)
{
    long long x1 = ax, y1 = ay, x2 = bx, y2 = by;
    if ((unsigned)ax >= (unsigned)w || (unsigned)bx >= (unsigned)w || (unsigned)ay >= (unsigned)h || (unsigned)by >= (unsigned)h)
        if (!clip_line(w, h, x1, y1, x2, y2)) return;
    int dx = (int)(x2 - x1), dy = (int)(y2 - y1), sy = 1;
    int px = (int)x1, py = (int)y1;
    if (dx < 0) { dx = -dx; dy = -dy; px = (int)x2; py = (int)y2; }       // left to right
    if (dy < 0) { dy = -dy; sy = -1; }
    const bool vert = dy > dx;
    if (vert) { int t = dx; dx = dy; dy = t; }
    // LineIterator: the minor coordinate advances when err < 0; err starts at dx - 2 dy, loses 2 dy per step and
    // gains 2 dx per minor step
    int err = dx - 2 * dy, m = 0;
    for (int i = 0; i <= dx; i++) {
        if (vert) set_px(mask, wpr, px + m, py + sy * i);
        else set_px(mask, wpr, px + i, py + sy * m);
        const bool neg = err < 0;
        err += -2 * dy + (neg ? 2 * dx : 0);
        m += neg ? 1 : 0;
    }
}

// cv::rectangle(img, (x - 2, y - 2), (x + 2, y + 2), thickness = -1), set or clear
static __device__ void fill_box(unsigned *mask, int wpr, int w, int h, int cx, int cy, bool value)
{
    for (int y = max(cy - 2, 0); y <= min(cy + 2, h - 1); y++)
        for (int x = max(cx - 2, 0); x <= min(cx + 2, w - 1); x++) {
            if (value) atomicOr(mask + y * wpr + (x >> 5), 1u << (x & 31));
            else atomicAnd(mask + y * wpr + (x >> 5), ~(1u << (x & 31)));
        }
}

// One polygon edge p0 -> p1 for the scan-line fill (cv::CollectPolyEdges + the per-row part of FillEdgeCollection):
// toggles the parity bit right of the edge in every row it crosses
static __device__ void fill_edge(unsigned *par, unsigned *exact, unsigned *carry, int wpr, int w, int h,
                                 int p0x, int p0y, int p1x, int p1y)
{
    if (p0y == p1y) return;
    long long c0x = p0x, c0y = p0y, c1x = p1x, c1y = p1y;
    if ((unsigned)p0x >= (unsigned)w || (unsigned)p1x >= (unsigned)w || (unsigned)p0y >= (unsigned)h || (unsigned)p1y >= (unsigned)h) {
        long long q0x = p0x, q0y = p0y, q1x = p1x, q1y = p1y;
        clip_line(w, h, q0x, q0y, q1x, q1y);
        c0x = q0x; c1x = q1x;                                   // x always from the clipped end points,
        if (q0y != q1y) { c0y = q0y; c1y = q1y; }               // y only when the clipped piece spans rows
    }
    const long long dx = ((c1x - c0x) * XY_ONE) / (c1y - c0y);      // C division: truncates, like the reference
    int y0, y1;
    long long x;
    if (p0y < p1y) { y0 = p0y; y1 = p1y; x = c0x * XY_ONE + (p0y - c0y) * dx; }
    else { y0 = p1y; y1 = p0y; x = c1x * XY_ONE + (p1y - c1y) * dx; }
    const int ya = max(y0, 0), yb = min(y1, h);
    x += (long long)(ya - y0) * dx;
    for (int y = ya; y < yb; y++, x += dx) {
        const long long a = x >> XY_SHIFT;
        if (a < 0) atomicXor(carry + y, 1u);
        else if (a < w) {
            if (a + 1 < w) atomicXor(par + y * wpr + ((a + 1) >> 5), 1u << ((a + 1) & 31));
            if ((x & (XY_ONE - 1)) == 0) atomicOr(exact + y * wpr + (a >> 5), 1u << (a & 31));
        }
    }
}

// dynamic shared memory: pts[n_points] (int2) | mask | par | exact (height * wpr words each) | carry[height]
__global__ void __launch_bounds__(256) k_lidar_bitmap(const F110BitmapDev d, const void *__restrict__ scans, long long n,
                                                      uint8_t *__restrict__ out, int packed)
{
    extern __shared__ unsigned smem[];
    const int T = d.n_points, H = d.height, W = d.width, wpr = d.wpr;
    int2 *pts = (int2 *)smem;
    unsigned *mask = smem + 2 * T;
    unsigned *par = mask + H * wpr;
    unsigned *exact = par + H * wpr;
    unsigned *carry = exact + H * wpr;
    const int tid = threadIdx.x, nt = blockDim.x;
    for (long long img = blockIdx.x; img < n; img += gridDim.x) {
        // polar -> pixel (lidar.py:71-83): rint(c + (scaling_factor * data) * cos(angle))
        for (int t = tid; t < T; t += nt) {
            const size_t src = (size_t)img * d.n_beams + d.indices[t];
            double sd;
            if (d.scan_is_f32) sd = (double)(d.scale32 * ((const float *)scans)[src]);
            else sd = d.scale * ((const double *)scans)[src];
            pts[t] = make_int2((int)rint((double)d.cx + sd * d.cosa[t]), (int)rint((double)d.cy + sd * d.sina[t]));
        }
        for (int i = tid; i < 3 * H * wpr + H; i += nt) mask[i] = 0u;
        __syncthreads();
        if (d.draw_mode == F110_BM_RAYS) {
            for (int t = tid; t < T; t += nt) {
                draw_line(mask, wpr, W, H, d.cx, d.cy, pts[t].x, pts[t].y);
                fill_box(mask, wpr, W, H, pts[t].x, pts[t].y, true);
            }
        } else {
            for (int t = tid; t < T; t += nt) {
                const int2 p0 = pts[t ? t - 1 : T - 1], p1 = pts[t];
                draw_line(mask, wpr, W, H, p0.x, p0.y, p1.x, p1.y);
                if (d.draw_mode == F110_BM_FILL) fill_edge(par, exact, carry, wpr, W, H, p0.x, p0.y, p1.x, p1.y);
            }
        }
        __syncthreads();
        if (d.draw_mode == F110_BM_FILL) {
            // per row: inclusive prefix XOR of the toggles (a toggle sits one pixel right of its edge)
            for (int y = tid; y < H; y += nt) {
                unsigned c = (carry[y] & 1u) ? 0xffffffffu : 0u;
                for (int k = 0; k < wpr; k++) {
                    unsigned v = par[y * wpr + k];
                    v ^= v << 1; v ^= v << 2; v ^= v << 4; v ^= v << 8; v ^= v << 16;
                    v ^= c;
                    c = (v >> 31) ? 0xffffffffu : 0u;
                    mask[y * wpr + k] |= v | exact[y * wpr + k];
                }
            }
            __syncthreads();
        }
        if (d.draw_center && tid == 0) fill_box(mask, wpr, W, H, d.cx, d.cy, d.draw_mode != F110_BM_FILL);
        __syncthreads();
        if (packed) {
            // 1 bit per pixel, row-major, bit b of word k of a row = pixel 32 k + b (set = drawn)
            unsigned *o = (unsigned *)out + (size_t)img * H * wpr;
            for (int i = tid; i < H * wpr; i += nt) {
                unsigned v = mask[i];
                const int k = i % wpr;
                if (k == wpr - 1 && (W & 31)) v &= (1u << (W & 31)) - 1u;
                o[i] = v;
            }
        } else if ((W & 15) == 0) {
            uint4 *o = (uint4 *)(out + (size_t)img * H * W);
            const unsigned fg = d.fg * 0x01010101u, bg = d.bg * 0x01010101u;
            for (int i = tid; i < H * (W >> 4); i += nt) {
                const int y = i / (W >> 4), g = i % (W >> 4);
                const unsigned bits = (mask[y * wpr + (g >> 1)] >> ((g & 1) * 16)) & 0xffffu;
                uint4 v;
                unsigned q[4];
#pragma unroll
                for (int j = 0; j < 4; j++) {
                    const unsigned b4 = (bits >> (4 * j)) & 15u;
                    // spread 4 bits to 4 byte lanes of 0x00 / 0xff
                    unsigned m = ((b4 & 1u) ? 0x000000ffu : 0u) | ((b4 & 2u) ? 0x0000ff00u : 0u) |
                                 ((b4 & 4u) ? 0x00ff0000u : 0u) | ((b4 & 8u) ? 0xff000000u : 0u);
                    q[j] = (fg & m) | (bg & ~m);
                }
                v.x = q[0]; v.y = q[1]; v.z = q[2]; v.w = q[3];
                o[i] = v;
            }
        } else {
            uint8_t *o = out + (size_t)img * H * W;
            for (int i = tid; i < H * W; i += nt) {
                const int y = i / W, x = i % W;
                o[i] = ((mask[y * wpr + (x >> 5)] >> (x & 31)) & 1u) ? d.fg : d.bg;
            }
        }
        __syncthreads();
    }
}

size_t f110_bitmap_smem_bytes(const F110BitmapDev &d)
{
    return (size_t)4 * (2 * (size_t)d.n_points + 3 * (size_t)d.height * d.wpr + d.height);
}

void f110_launch_bitmap(const F110BitmapDev &d, const void *scans, long long n, uint8_t *out, int packed, cudaStream_t st)
{
    if (n <= 0) return;
    const size_t smem = f110_bitmap_smem_bytes(d);
    // every SM holds as many CTAs as shared memory allows; a grid-stride loop covers the rest
    long long grid = n < 148LL * 8 ? n : 148LL * 8;
    if (smem > 48 * 1024) cudaFuncSetAttribute(k_lidar_bitmap, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    k_lidar_bitmap<<<(unsigned)grid, 256, smem, st>>>(d, scans, n, out, packed);
}
