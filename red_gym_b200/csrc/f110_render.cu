// Headless renderer for sm_100a: F110Env.render's scene (gym/f110_gym/envs/rendering.py) rasterised
// by a CUDA kernel into an RGB8 device buffer.  The reference path is pyglet/OpenGL, which cannot run on this box; the
// window, swap chain, text labels and mouse camera are left out, the scene and the projection are the reference's:
// clear colour, one point per zero-valued map pixel at 50 x world coordinates, one quad per car (ego / others in the
// reference's colours) drawn over the map, orthographic camera (fixed, or following the ego car like
// examples/waypoint_follow.py:259-269).  A point lights the pixel that contains it, a quad the pixels whose centre it
// contains.  One thread per output pixel: the map is sampled by inverting the (separable) point projection and checking
// the candidate cells with the forward formula, so the image does not depend on how many obstacle pixels the map has.
// Compiled with -fmad=false (same rounding as the numpy restatement the tests compare against).
#include <math.h>
#include "f110_render_dev.h"
#include "f110_dev.h"      // F110_CHECK / F110_V_*

#define CAR_LENGTH 0.58
#define CAR_WIDTH 0.31
#define MAX_A 16

static __device__ __forceinline__ void car_vertices(double x, double y, double th, double v[4][2])
{
    // collision_models.get_vertices (:238-260): rl, rr, fr, fl through the homogeneous transform of get_trmtx
    const double c = cos(th), s = sin(th);
    const double lx[4] = { -CAR_LENGTH / 2, -CAR_LENGTH / 2, CAR_LENGTH / 2, CAR_LENGTH / 2 };
    const double ly[4] = { CAR_WIDTH / 2, -CAR_WIDTH / 2, -CAR_WIDTH / 2, CAR_WIDTH / 2 };
#pragma unroll
    for (int k = 0; k < 4; k++) {
        v[k][0] = 50.0 * (c * lx[k] + -s * ly[k] + x);
        v[k][1] = 50.0 * (s * lx[k] + c * ly[k] + y);
    }
}

// does any obstacle cell project into pixel (px, py)?  forward: pixel = floor((50 (i res + o) - lo) / (hi - lo) * n)
static __device__ __forceinline__ void cell_range(double lo, double hi, int npx, int p, double res, double o, int ncell, int &c0, int &c1)
{
    const double span = hi - lo;
    int c = (int)floor(((lo + (double)p * span / npx) / 50.0 - o) / res) - 1;      // a cell safely left of the pixel
    if (c < 0) c = 0;
    while (c < ncell && floor((50.0 * (c * res + o) - lo) / span * npx) < (double)p) c++;
    c0 = c;
    while (c < ncell && floor((50.0 * (c * res + o) - lo) / span * npx) == (double)p) c++;
    c1 = c;
}

__global__ void __launch_bounds__(256) k_render(const F110RenderDev d, const double *__restrict__ xs, const double *__restrict__ ys,
                                                const double *__restrict__ thetas, uint8_t *__restrict__ out)
{
    __shared__ double verts[MAX_A][4][2];
    __shared__ double cam[4];
    const long long env = blockIdx.x;
    const int A = d.n_agents;
    if (threadIdx.x < A) {
        const long long ia = env * A + threadIdx.x;
        car_vertices(xs[ia], ys[ia], thetas[ia], verts[threadIdx.x]);
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        if (d.follow_margin >= 0.0) {
            double xl = verts[d.ego_idx][0][0], xh = xl, yl = verts[d.ego_idx][0][1], yh = yl;
            for (int k = 1; k < 4; k++) {
                xl = fmin(xl, verts[d.ego_idx][k][0]); xh = fmax(xh, verts[d.ego_idx][k][0]);
                yl = fmin(yl, verts[d.ego_idx][k][1]); yh = fmax(yh, verts[d.ego_idx][k][1]);
            }
            cam[0] = xl - d.follow_margin; cam[1] = xh + d.follow_margin; cam[2] = yl - d.follow_margin; cam[3] = yh + d.follow_margin;
        } else { cam[0] = d.left; cam[1] = d.right; cam[2] = d.bottom; cam[3] = d.top; }
    }
    __syncthreads();
    const int pix = blockIdx.y * blockDim.x + threadIdx.x;
    if (pix >= d.width * d.height) return;
    const int row = pix / d.width, px = pix % d.width;      // row 0 = top of the window
    const int py = d.height - 1 - row;
    const double left = cam[0], right = cam[1], bottom = cam[2], top = cam[3];
    unsigned char r = 9, g = 32, b = 87;                    // glClearColor (rendering.py:72)
    int c0, c1, r0, r1;
    cell_range(left, right, d.width, px, d.res, d.ox, d.map_w, c0, c1);
    cell_range(bottom, top, d.height, py, d.res, d.oy, d.map_h, r0, r1);
    bool hit = false;
    F110_CHECK(d.violations, r0 >= 0 && r1 <= d.map_h && c0 >= 0 && c1 <= d.map_w, F110_V_EDT64);      // cell_range clamps to the map
    for (int rr = r0; rr < r1 && !hit; rr++)
        for (int cc = c0; cc < c1; cc++)
            if (d.obstacle[(size_t)rr * d.map_w + cc]) { hit = true; break; }
    if (hit) { r = 183; g = 193; b = 222; }
    const double X = left + ((double)px + 0.5) * ((right - left) / d.width);
    const double Y = bottom + ((double)py + 0.5) * ((top - bottom) / d.height);
    for (int a = 0; a < A; a++) {
        bool pos = true, neg = true;
#pragma unroll
        for (int k = 0; k < 4; k++) {
            const double x0 = verts[a][k][0], y0 = verts[a][k][1], x1 = verts[a][(k + 1) & 3][0], y1 = verts[a][(k + 1) & 3][1];
            const double cr = (x1 - x0) * (Y - y0) - (y1 - y0) * (X - x0);
            pos = pos && cr >= 0;
            neg = neg && cr <= 0;
        }
        if (pos || neg) {
            if (a == d.ego_idx) { r = 172; g = 97; b = 185; } else { r = 99; g = 52; b = 94; }
        }
    }
    uint8_t *o = out + ((size_t)env * d.width * d.height + pix) * 3;
    o[0] = r; o[1] = g; o[2] = b;
}

void f110_launch_render(const F110RenderDev &d, const double *x, const double *y, const double *theta, long long n_envs,
                        uint8_t *out, cudaStream_t st)
{
    if (n_envs <= 0) return;
    dim3 grid((unsigned)n_envs, (unsigned)((d.width * d.height + 255) / 256));
    k_render<<<grid, 256, 0, st>>>(d, x, y, theta, out);
}
