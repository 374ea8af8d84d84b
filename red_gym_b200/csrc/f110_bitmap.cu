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
#include "f110_dev.h"      // F110_CHECK / F110_V_*: the checked build tests every pixel, row and queue slot below

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

// what the checked build needs to test a pixel; in the normal build nothing reads it and it is compiled away
struct BmChk { unsigned long long *v; int w, h; };

static __device__ __forceinline__ void set_px(unsigned *mask, int wpr, int x, int y, const BmChk &chk)
{
    F110_CHECK(chk.v, (unsigned)x < (unsigned)chk.w && (unsigned)y < (unsigned)chk.h, F110_V_BEAM);     // clipping is the only bounds test a line has
    atomicOr(mask + y * wpr + (x >> 5), 1u << (x & 31));
}

// cv::Line(img, p1, p2, color, LINE_8), thickness 1, split in two so that a long line can be drawn by a whole warp:
// line_setup clips and normalises (left to right, major axis first), line_span draws pixels [i0, i1) of it.
struct LineJob { int px, py, dx, dy, sy; bool vert; };

static __device__ __forceinline__ bool line_setup(int w, int h, int ax, int ay, int bx, int by, LineJob &j)
{
    long long x1 = ax, y1 = ay, x2 = bx, y2 = by;
    if ((unsigned)ax >= (unsigned)w || (unsigned)bx >= (unsigned)w || (unsigned)ay >= (unsigned)h || (unsigned)by >= (unsigned)h)
        if (!clip_line(w, h, x1, y1, x2, y2)) return false;
    int dx = (int)(x2 - x1), dy = (int)(y2 - y1);
    j.sy = 1;
    j.px = (int)x1; j.py = (int)y1;
    if (dx < 0) { dx = -dx; dy = -dy; j.px = (int)x2; j.py = (int)y2; }       // left to right
    if (dy < 0) { dy = -dy; j.sy = -1; }
    j.vert = dy > dx;
    if (j.vert) { int t = dx; dx = dy; dy = t; }
    j.dx = dx; j.dy = dy;
    return true;
}

// LineIterator: the minor coordinate advances when err < 0; err starts at dx - 2 dy, loses 2 dy per step and gains 2 dx
// per minor step, so before pixel i the minor offset is m_i = floor((2 dy i + dx - 1) / (2 dx)) and
// err_i = dx - 2 dy (i + 1) + 2 dx m_i: a span can start anywhere.
static __device__ __forceinline__ void line_span(unsigned *mask, int wpr, const LineJob &j, int i0, int i1, const BmChk &chk)
{
    if (i0 >= i1) return;
    int m = 0;
    if (i0 > 0 && j.dx > 0) {
        if (j.dx < 16384) m = (2 * j.dy * i0 + j.dx - 1) / (2 * j.dx);      // dy <= dx, i0 <= dx: below 2^30
        else m = (int)((2ll * j.dy * i0 + j.dx - 1) / (2ll * j.dx));
    }
    int err = j.dx - 2 * j.dy * (i0 + 1) + 2 * j.dx * m;
    for (int i = i0; i < i1; i++) {
        if (j.vert) set_px(mask, wpr, j.px + m, j.py + j.sy * i, chk);
        else set_px(mask, wpr, j.px + i, j.py + j.sy * m, chk);
        const bool neg = err < 0;
        err += -2 * j.dy + (neg ? 2 * j.dx : 0);
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

// One polygon edge p0 -> p1 of the scan-line fill (cv::CollectPolyEdges + the per-row part of FillEdgeCollection), also
// split: edge_setup gives the 16.16 fixed-point x at row ya and the slope, edge_rows toggles the parity bit right of the
// edge in rows [y0, y1).
struct EdgeJob { int ya, yb; long long x, dx; };

static __device__ __forceinline__ bool edge_setup(int w, int h, int p0x, int p0y, int p1x, int p1y, EdgeJob &e)
{
    if (p0y == p1y) return false;
    long long c0x = p0x, c0y = p0y, c1x = p1x, c1y = p1y;
    if ((unsigned)p0x >= (unsigned)w || (unsigned)p1x >= (unsigned)w || (unsigned)p0y >= (unsigned)h || (unsigned)p1y >= (unsigned)h) {
        long long q0x = p0x, q0y = p0y, q1x = p1x, q1y = p1y;
        clip_line(w, h, q0x, q0y, q1x, q1y);
        c0x = q0x; c1x = q1x;                                   // x always from the clipped end points,
        if (q0y != q1y) { c0y = q0y; c1y = q1y; }               // y only when the clipped piece spans rows
    }
    {   // C division: truncates, like the reference; 32-bit when the operands fit (a 64-bit division is ~4x the instructions)
        const long long ddx = c1x - c0x, ddy = c1y - c0y;
        if (ddx > -32768 && ddx < 32768 && ddy > -32768 && ddy < 32768) e.dx = (long long)(((int)ddx * XY_ONE) / (int)ddy);
        else e.dx = (ddx * XY_ONE) / ddy;
    }
    int y0, y1;
    long long x;
    if (p0y < p1y) { y0 = p0y; y1 = p1y; x = c0x * XY_ONE + (p0y - c0y) * e.dx; }
    else { y0 = p1y; y1 = p0y; x = c1x * XY_ONE + (p1y - c1y) * e.dx; }
    e.ya = max(y0, 0); e.yb = min(y1, h);
    e.x = x + (long long)(e.ya - y0) * e.dx;
    return e.ya < e.yb;
}

static __device__ __forceinline__ void edge_rows(unsigned *par, unsigned *exact, unsigned *carry, int wpr, int w, const EdgeJob &e,
                                                 int y0, int y1, const BmChk &chk)
{
    long long x = e.x + (long long)(y0 - e.ya) * e.dx;
    for (int y = y0; y < y1; y++, x += e.dx) {
        F110_CHECK(chk.v, (unsigned)y < (unsigned)chk.h, F110_V_BEAM);
        const long long a = x >> XY_SHIFT;
        if (a < 0) atomicXor(carry + y, 1u);
        else if (a < w) {
            if (a + 1 < w) atomicXor(par + y * wpr + ((a + 1) >> 5), 1u << ((a + 1) & 31));
            if ((x & (XY_ONE - 1)) == 0) atomicOr(exact + y * wpr + (a >> 5), 1u << (a & 31));
        }
    }
}

// Items longer than this many pixels / rows are not finished by the thread that owns them (the other 255 threads of the
// CTA would wait at the barrier) but queued and drawn by whole warps, 32 pixels / rows at a time.
#define BM_SHORT 12

// dynamic shared memory: pts[n_points] (int2) | mask | par | exact (height * wpr words each) | carry[height] |
// queue[2 * n_points] | queue counter (4 words) | colour table (256 x 8 bytes)
__global__ void __launch_bounds__(256, 4) k_lidar_bitmap(const F110BitmapDev d, const void *__restrict__ scans, long long n,
                                                      uint8_t *__restrict__ out, int packed)
{
    extern __shared__ unsigned smem[];
    const int T = d.n_points, H = d.height, W = d.width, wpr = d.wpr;
    int2 *pts = (int2 *)smem;
    unsigned *mask = smem + 2 * T;
    unsigned *par = mask + H * wpr;
    unsigned *exact = par + H * wpr;
    unsigned *carry = exact + H * wpr;
    unsigned *queue = carry + H;              // entries: point index t, bit 31 set = fill edge (else line)
    unsigned *qn = queue + 2 * T;
    uint2 *lut = (uint2 *)(smem + ((2 * T + 3 * H * wpr + H + 2 * T + 4 + 1) & ~1));   // 256 entries: 8 mask bits -> 8 bytes of bg / fg
    const int tid = threadIdx.x, nt = blockDim.x, lane = tid & 31, wid = tid >> 5, nw = nt >> 5;
    const bool rays = d.draw_mode == F110_BM_RAYS, fill = d.draw_mode == F110_BM_FILL;
    const BmChk chk = { d.violations, W, H };
    for (int b = tid; b < 256; b += nt) {
        const unsigned bg = d.bg * 0x01010101u, fx = (d.fg ^ d.bg) * 0x01010101u;
        const unsigned m0 = (((b & 15u) * 0x00204081u) & 0x01010101u) * 0xffu, m1 = ((((b >> 4) & 15u) * 0x00204081u) & 0x01010101u) * 0xffu;
        lut[b] = make_uint2(bg ^ (fx & m0), bg ^ (fx & m1));
    }
    for (long long img = blockIdx.x; img < n; img += gridDim.x) {
        // polar -> pixel (lidar.py:71-83): rint(c + (scaling_factor * data) * cos(angle))
        for (int t = tid; t < T; t += nt) {
            F110_CHECK(d.violations, d.indices[t] >= 0 && d.indices[t] < d.n_beams, F110_V_BEAM);
            const size_t src = (size_t)img * d.n_beams + d.indices[t];
            double sd;
            if (d.scan_is_f32) sd = (double)(d.scale32 * ((const float *)scans)[src]);
            else sd = d.scale * ((const double *)scans)[src];
            pts[t] = make_int2((int)rint((double)d.cx + sd * d.cosa[t]), (int)rint((double)d.cy + sd * d.sina[t]));
        }
        for (int i = tid; i < 3 * H * wpr + H; i += nt) mask[i] = 0u;
        if (tid == 0) { qn[0] = 0u; qn[1] = 0u; }     // queue fill, next unclaimed point of pass 1
        __syncthreads();
        // pass 1: one thread per line / edge, claimed one at a time (their cost varies: clipping, 0-12 pixels / rows);
        // short ones are finished here, long ones queued
        for (int t = (int)atomicAdd(qn + 1, 1u); t < T; t = (int)atomicAdd(qn + 1, 1u)) {
            const int2 p1 = pts[t];
            const int2 p0 = rays ? make_int2(d.cx, d.cy) : pts[t ? t - 1 : T - 1];
            LineJob lj;
            if (line_setup(W, H, p0.x, p0.y, p1.x, p1.y, lj)) {
                if (lj.dx < BM_SHORT) line_span(mask, wpr, lj, 0, lj.dx + 1, chk);
                else {
                    const unsigned slot = atomicAdd(qn, 1u);
                    F110_CHECK(d.violations, slot < 2u * (unsigned)T, F110_V_QUEUE);
                    queue[slot] = (unsigned)t;
                }
            }
            if (rays) fill_box(mask, wpr, W, H, p1.x, p1.y, true);
            if (fill) {
                EdgeJob ej;
                if (edge_setup(W, H, p0.x, p0.y, p1.x, p1.y, ej)) {
                    if (ej.yb - ej.ya <= BM_SHORT) edge_rows(par, exact, carry, wpr, W, ej, ej.ya, ej.yb, chk);
                    else {
                        const unsigned slot = atomicAdd(qn, 1u);
                        F110_CHECK(d.violations, slot < 2u * (unsigned)T, F110_V_QUEUE);
                        queue[slot] = (unsigned)t | 0x80000000u;
                    }
                }
            }
        }
        __syncthreads();
        // pass 2: the queued long items, one warp per item, lanes over pixels / rows
        const int nq = (int)*qn;
        for (int k = wid; k < nq; k += nw) {
            const unsigned e = queue[k];
            const int t = (int)(e & 0x7fffffffu);
            F110_CHECK(d.violations, t < T, F110_V_QUEUE);
            const int2 p1 = pts[t];
            const int2 p0 = rays ? make_int2(d.cx, d.cy) : pts[t ? t - 1 : T - 1];
            if (e & 0x80000000u) {
                EdgeJob ej;
                edge_setup(W, H, p0.x, p0.y, p1.x, p1.y, ej);
                const int rows = ej.yb - ej.ya, per = (rows + 31) >> 5;
                const int y0 = ej.ya + lane * per;
                edge_rows(par, exact, carry, wpr, W, ej, min(y0, ej.yb), min(y0 + per, ej.yb), chk);
            } else {
                LineJob lj;
                line_setup(W, H, p0.x, p0.y, p1.x, p1.y, lj);
                const int per = (lj.dx + 1 + 31) >> 5;
                line_span(mask, wpr, lj, min(lane * per, lj.dx + 1), min(lane * per + per, lj.dx + 1), chk);
            }
        }
        __syncthreads();
        if (fill) {
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
        if (d.draw_center && tid < 25) {      // cv2.rectangle(center - 2, center + 2, thickness = -1): 5 x 5 pixels
            const int x = d.cx - 2 + tid % 5, y = d.cy - 2 + tid / 5;
            if ((unsigned)x < (unsigned)W && (unsigned)y < (unsigned)H) {
                if (!fill) atomicOr(mask + y * wpr + (x >> 5), 1u << (x & 31));
                else atomicAnd(mask + y * wpr + (x >> 5), ~(1u << (x & 31)));
            }
        }
        __syncthreads();
        if (packed) {
            // 1 bit per pixel, row-major, bit b of word k of a row = pixel 32 k + b (set = drawn)
            unsigned *o = (unsigned *)out + (size_t)img * H * wpr;
            const unsigned last = (W & 31) ? (1u << (W & 31)) - 1u : 0xffffffffu;
            for (int i = tid; i < H * wpr; i += nt) o[i] = ((i + 1) % wpr == 0) ? (mask[i] & last) : mask[i];
        } else if ((W & 31) == 0) {
            // one mask word -> 32 pixels -> two 16-byte stores; byte -> 8 colour bytes through the table
            uint4 *o = (uint4 *)(out + (size_t)img * H * W);
            for (int i = tid; i < H * wpr; i += nt) {
                const unsigned v = mask[i];
                const uint2 a0 = lut[v & 255u], a1 = lut[(v >> 8) & 255u], a2 = lut[(v >> 16) & 255u], a3 = lut[v >> 24];
                o[2 * i] = make_uint4(a0.x, a0.y, a1.x, a1.y);
                o[2 * i + 1] = make_uint4(a2.x, a2.y, a3.x, a3.y);
            }
        } else {
            uint8_t *o = out + (size_t)img * H * W;
            for (int y = wid; y < H; y += nw)
                for (int x = lane; x < W; x += 32)
                    o[(size_t)y * W + x] = ((mask[y * wpr + (x >> 5)] >> (x & 31)) & 1u) ? d.fg : d.bg;
        }
        __syncthreads();
    }
}

size_t f110_bitmap_smem_bytes(const F110BitmapDev &d)
{
    return (size_t)4 * (2 * (size_t)d.n_points + 3 * (size_t)d.height * d.wpr + d.height + 2 * (size_t)d.n_points + 4 + 2 + 512);
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
