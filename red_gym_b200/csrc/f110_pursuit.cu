// Batched pure-pursuit planner for sm_100a: closes the loop of examples/waypoint_follow.py
// (BASELINE config 1) on the device, one thread per car, no host round trip between step and plan.
//
// Same operations in the same order as the reference's njit helpers (examples/waypoint_follow.py):
//   nearest_point_on_trajectory :15-46, first_point_on_trajectory_intersecting_circle :48-127 (wrap = True),
//   get_actuation :129-141, PurePursuitPlanner._get_current_waypoint / plan :183-217 -- including its quirks: the
//   look-ahead target is the START waypoint of the intersected segment, its speed comes from the NEAREST segment's
//   waypoint, segment ends are shifted by 1e-6, no target -> (speed 4.0, steer 0.0).
// fp64, compiled with -fmad=false.  The polyline (x, y, v) sits in shared memory.
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>
#include "f110_pursuit_dev.h"

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

__global__ void __launch_bounds__(128) k_pure_pursuit(const F110PursuitDev d, const double *__restrict__ xs,
                                                      const double *__restrict__ ys, const double *__restrict__ thetas,
                                                      long long n, long long stride, double lookahead, double vgain,
                                                      const double *__restrict__ lookahead_q, const double *__restrict__ vgain_q,
                                                      double *__restrict__ actions, int *__restrict__ picked)
{
    extern __shared__ double sm[];
    const int nw = d.n_waypoints;
    double *wx = sm, *wy = sm + nw, *wv = sm + 2 * nw;
    for (int i = threadIdx.x; i < nw; i += blockDim.x) { wx[i] = d.wx[i]; wy[i] = d.wy[i]; wv[i] = d.wv[i]; }
    __syncthreads();
    const long long q = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= n) return;
    const double px = xs[q * stride], py = ys[q * stride], th = thetas[q * stride];
    const double ld = lookahead_q ? lookahead_q[q] : lookahead;
    const double vg = vgain_q ? vgain_q[q] : vgain;
    // nearest_point_on_trajectory: first minimum of the point-to-segment distances (np.argmin)
    int bi = 0;
    double bd = 0.0, bt = 0.0, bd2 = 0.0;
    for (int i = 0; i + 1 < nw; i++) {
        const double dx = wx[i + 1] - wx[i], dy = wy[i + 1] - wy[i];
        const double l2 = dx * dx + dy * dy;
        double t = ((px - wx[i]) * dx + (py - wy[i]) * dy) / l2;
        if (t < 0.0) t = 0.0;
        if (t > 1.0) t = 1.0;
        const double ex = px - (wx[i] + t * dx), ey = py - (wy[i] + t * dy);
        const double d2 = ex * ex + ey * ey;
        // sqrt only for candidates: d2 >= best d2 cannot give a strictly smaller rounded root
        if (i == 0 || d2 < bd2 || d2 != d2) {
            const double dist = sqrt(d2);
            if (i == 0 || dist < bd || (dist != dist && bd == bd)) { bd = dist; bt = t; bi = i; bd2 = d2; }
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
        for (int i = start_i; i < nw - 1 && !have; i++) {
            if (!circle_hit(wx, wy, i, i + 1, px, py, ld, t1, t2)) continue;
            if (i == start_i) {
                if ((t1 >= 0.0 && t1 <= 1.0 && t1 >= start_t) || (t2 >= 0.0 && t2 <= 1.0 && t2 >= start_t)) { have = true; i2 = i; }
            } else if ((t1 >= 0.0 && t1 <= 1.0) || (t2 >= 0.0 && t2 <= 1.0)) { have = true; i2 = i; }
        }
        for (int i = -1; i < start_i && !have; i++) {     // wrap around to the beginning of the trajectory
            const int i0 = i < 0 ? i + nw : i % nw, i1 = (i + 1) % nw;
            if (!circle_hit(wx, wy, i0, i1, px, py, ld, t1, t2)) continue;
            if ((t1 >= 0.0 && t1 <= 1.0) || (t2 >= 0.0 && t2 <= 1.0)) { have = true; i2 = i; }
        }
        if (have) { const int k = i2 < 0 ? i2 + nw : i2; tx = wx[k]; ty = wy[k]; }     // wpts[i2] (python index -1 = last)
    } else if (bd < d.max_reacquire) {
        have = true; i2 = bi; tx = wx[bi]; ty = wy[bi];
    }
    double speed = 4.0, steer = 0.0;
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
    const size_t smem = (size_t)3 * d.n_waypoints * sizeof(double);
    if (smem > 48 * 1024) cudaFuncSetAttribute(k_pure_pursuit, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    k_pure_pursuit<<<(unsigned)((n + 127) / 128), 128, smem, st>>>(d, x, y, theta, n, stride, lookahead, vgain, lookahead_q, vgain_q,
                                                                   actions, picked);
}
