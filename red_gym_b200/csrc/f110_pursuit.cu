// Batched pure-pursuit planner for sm_100a: closes the loop of examples/waypoint_follow.py
// (BASELINE config 1) on the device, one thread per car, no host round trip between step and plan.
//
// Same operations in the same order as the reference's njit helpers (examples/waypoint_follow.py):
//   nearest_point_on_trajectory :15-46, first_point_on_trajectory_intersecting_circle :48-127 (wrap = True),
//   get_actuation :129-141, PurePursuitPlanner._get_current_waypoint / plan :183-217 -- including its quirks: the
//   look-ahead target is the START waypoint of the intersected segment, its speed comes from the NEAREST segment's
//   waypoint, segment ends are shifted by 1e-6, no target -> (speed 4.0, steer 0.0).
// fp64, compiled with -fmad=false.  The polyline (x, y, v) sits in shared memory.
//
// nearest_point_on_trajectory is an argmin over every segment (782 on example_waypoints); a car is near a handful.
// Segments are taken in groups of 16 with a bounding circle each; a group whose circle is farther away than a
// segment distance already known cannot hold the minimum and is skipped, the others are evaluated exactly, in
// index order (np.argmin returns the FIRST minimum).  Same result, ~50 circle tests + ~50 exact segments per car
// instead of 782 exact segments with an IEEE division each.
#include <cuda_runtime.h>
#include <math.h>
#include <math_constants.h>
#include <stdint.h>
#include "f110_pursuit_dev.h"
#include "f110_dev.h"      // F110_CHECK / F110_V_*

static __device__ __forceinline__ bool circle_hit(const double *wx, const double *wy, int i0, int i1, double px, double py,
                                                  double radius, double &t1, double &t2)
{
    const double sx = wx[i0], sy = wy[i0];
    const double ex = wx[i1] + 1e-6, ey = wy[i1] + 1e-6;
    const double vx = ex - sx, vy = ey - sy;
    const double a = vx * vx + vy * vy;
    const double b = 2.0 * (vx * (sx - px) + vy * (sy - py));
    const double c = (sx * sx + sy * sy) + (px * px + py * py) - 2.0 * (sx * px + sy * py) - radius * radius;
    double disc = b * b - 4 * a * c;
    if (disc < 0) return false;
    disc = sqrt(disc);
    t1 = (-b - disc) / (2.0 * a);
    t2 = (-b + disc) / (2.0 * a);
    return true;
}

#define PP_GROUP 16

__global__ void __launch_bounds__(128) k_pure_pursuit(const F110PursuitDev d, const double *__restrict__ xs,
                                                      const double *__restrict__ ys, const double *__restrict__ thetas,
                                                      long long n, long long stride, double lookahead, double vgain,
                                                      const double *__restrict__ lookahead_q, const double *__restrict__ vgain_q,
                                                      double *__restrict__ actions, int *__restrict__ picked)
{
    extern __shared__ double sm[];
    const int nw = d.n_waypoints, nseg = nw - 1, ng = (nseg + PP_GROUP - 1) / PP_GROUP;
    double *wx = sm, *wy = sm + nw, *wv = sm + 2 * nw;
    double *sdx = sm + 3 * nw, *sdy = sm + 4 * nw, *sl2 = sm + 5 * nw;       // diffs and l2s of the reference (:27-28)
    double *gcx = sm + 6 * nw, *gcy = gcx + ng, *gr = gcy + ng;              // bounding circle of each group of segments
    for (int i = threadIdx.x; i < nw; i += blockDim.x) { wx[i] = d.wx[i]; wy[i] = d.wy[i]; wv[i] = d.wv[i]; }
    __syncthreads();
    for (int i = threadIdx.x; i < nseg; i += blockDim.x) {
        const double dx = wx[i + 1] - wx[i], dy = wy[i + 1] - wy[i];
        sdx[i] = dx; sdy[i] = dy; sl2[i] = dx * dx + dy * dy;
    }
    for (int g = threadIdx.x; g < ng; g += blockDim.x) {
        const int k0 = g * PP_GROUP, k1 = min(k0 + PP_GROUP, nseg);             // way points k0 .. k1 span the group's segments
        double cx = 0.0, cy = 0.0;
        for (int k = k0; k <= k1; k++) { cx += wx[k]; cy += wy[k]; }
        cx /= (double)(k1 - k0 + 1); cy /= (double)(k1 - k0 + 1);
        double r2 = 0.0;
        for (int k = k0; k <= k1; k++) r2 = fmax(r2, (wx[k] - cx) * (wx[k] - cx) + (wy[k] - cy) * (wy[k] - cy));
        gcx[g] = cx; gcy[g] = cy; gr[g] = sqrt(r2) * (1.0 + 1e-9) + 1e-12;      // NaN way points: NaN radius, never skipped
    }
    __syncthreads();
    const long long q = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= n) return;
    const double px = xs[q * stride], py = ys[q * stride], th = thetas[q * stride];
    const double ld = lookahead_q ? lookahead_q[q] : lookahead;
    const double vg = vgain_q ? vgain_q[q] : vgain;
    // an upper bound of the minimum: the closest way point of the group whose centre is closest
    int gs = 0;
    double gd = CUDART_INF;
    for (int g = 0; g < ng; g++) {
        const double ex = px - gcx[g], ey = py - gcy[g];
        const double c2 = ex * ex + ey * ey;
        if (c2 < gd) { gd = c2; gs = g; }
    }
    double bound2 = CUDART_INF;
    for (int k = gs * PP_GROUP; k <= min(gs * PP_GROUP + PP_GROUP, nseg); k++) {
        const double ex = px - wx[k], ey = py - wy[k];
        bound2 = fmin(bound2, ex * ex + ey * ey);
    }
    double reach = sqrt(bound2) * (1.0 + 1e-9) + 1e-12;                          // groups farther than this are out
    // nearest_point_on_trajectory: first minimum of the point-to-segment distances (np.argmin)
    int bi = 0;
    bool none = true;
    double bd = 0.0, bt = 0.0, bd2 = 0.0;
    for (int g = 0; g < ng; g++) {
        {
            const double ex = px - gcx[g], ey = py - gcy[g];
            const double lim = gr[g] + reach;
            if (ex * ex + ey * ey > lim * lim) continue;                         // false for NaN: evaluated
        }
        const int i1 = min(g * PP_GROUP + PP_GROUP, nseg);
        for (int i = g * PP_GROUP; i < i1; i++) {
            const double dx = sdx[i], dy = sdy[i];
            double t = ((px - wx[i]) * dx + (py - wy[i]) * dy) / sl2[i];
            if (t < 0.0) t = 0.0;
            if (t > 1.0) t = 1.0;
            const double ex = px - (wx[i] + t * dx), ey = py - (wy[i] + t * dy);
            const double d2 = ex * ex + ey * ey;
            // sqrt only for candidates: d2 >= best d2 cannot give a strictly smaller rounded root
            if (none || d2 < bd2 || d2 != d2) {
                const double dist = sqrt(d2);
                if (none || dist < bd || (dist != dist && bd == bd)) {
                    bd = dist; bt = t; bi = i; bd2 = d2; none = false;
                    reach = fmin(reach, dist * (1.0 + 1e-9) + 1e-12);
                }
            }
        }
    }
    bool have = false;
    int i2 = -2;
    double tx = 0.0, ty = 0.0;
    if (bd < ld) {
        const double tt = (double)bi + bt;
        const int start_i = (int)tt;
        const double start_t = tt - floor(tt);            // t % 1.0, t >= 0
        double t1, t2;
        F110_CHECK(d.violations, start_i >= 0 && start_i < nw, F110_V_TABLE);
        for (int i = start_i; i < nw - 1 && !have; i++) {
            if (!circle_hit(wx, wy, i, i + 1, px, py, ld, t1, t2)) continue;
            if (i == start_i) {
                if ((t1 >= 0.0 && t1 <= 1.0 && t1 >= start_t) || (t2 >= 0.0 && t2 <= 1.0 && t2 >= start_t)) { have = true; i2 = i; }
            } else if ((t1 >= 0.0 && t1 <= 1.0) || (t2 >= 0.0 && t2 <= 1.0)) { have = true; i2 = i; }
        }
        for (int i = -1; i < start_i && !have; i++) {     // wrap around to the beginning of the trajectory
            const int i0 = i < 0 ? i + nw : i % nw, i1 = (i + 1) % nw;
            F110_CHECK(d.violations, i0 >= 0 && i0 < nw && i1 >= 0 && i1 < nw, F110_V_TABLE);
            if (!circle_hit(wx, wy, i0, i1, px, py, ld, t1, t2)) continue;
            if ((t1 >= 0.0 && t1 <= 1.0) || (t2 >= 0.0 && t2 <= 1.0)) { have = true; i2 = i; }
        }
        if (have) {      // wpts[i2] (python index -1 = last)
            const int k = i2 < 0 ? i2 + nw : i2;
            F110_CHECK(d.violations, k >= 0 && k < nw, F110_V_TABLE);
            tx = wx[k]; ty = wy[k];
        }
    } else if (bd < d.max_reacquire) {
        have = true; i2 = bi; tx = wx[bi]; ty = wy[bi];
    }
    double speed = 4.0, steer = 0.0;
    F110_CHECK(d.violations, bi >= 0 && bi < nw, F110_V_TABLE);
    if (have) {
        // get_actuation
        const double wyl = sin(-th) * (tx - px) + cos(-th) * (ty - py);
        speed = vg * wv[bi];
        if (!(fabs(wyl) < 1e-6)) {
            const double radius = 1 / (2.0 * wyl / (ld * ld));
            steer = atan(d.wheelbase / radius);
        }
    }
    actions[2 * q] = steer;        // env action layout: [steer, speed] (waypoint_follow.py:302)
    actions[2 * q + 1] = speed;
    if (picked) { picked[2 * q] = bi; picked[2 * q + 1] = i2; }
}

void f110_launch_pursuit(const F110PursuitDev &d, const double *x, const double *y, const double *theta, long long n,
                         long long stride, double lookahead, double vgain, const double *lookahead_q, const double *vgain_q,
                         double *actions, int *picked, cudaStream_t st)
{
    if (n <= 0) return;
    const int ng = (d.n_waypoints - 1 + PP_GROUP - 1) / PP_GROUP;
    const size_t smem = ((size_t)6 * d.n_waypoints + 3 * (size_t)(ng > 0 ? ng : 0)) * sizeof(double);
    if (smem > 48 * 1024) cudaFuncSetAttribute(k_pure_pursuit, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    k_pure_pursuit<<<(unsigned)((n + 127) / 128), 128, smem, st>>>(d, x, y, theta, n, stride, lookahead, vgain, lookahead_q, vgain_q,
                                                                   actions, picked);
}
