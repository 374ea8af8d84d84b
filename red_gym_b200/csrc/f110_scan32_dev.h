// Per-scan preparation of the fp32 lidar, shared by the ray-cast kernels (f110_kernels_fp32.cu) and the spatial sort
// (f110_order.cu): pose -> origin cell and its distance, beam table index, noise key.  Large launches are sorted anyway
// (f110_order.cu reads every agent's pose); the sort then also prepares the scan and hands the ray-cast a 32-byte record per
// scan IN LAUNCH ORDER, so that a warp starts marching after one coalesced load instead of waiting at a CTA barrier for eight
// threads to walk a chain of dependent loads, an fp64 fmod and a division (5 % of the kernel's stall samples, ncu r2a).
// Every double-precision step is written with explicit rounding intrinsics: the same source must give the same bits in both
// translation units, whatever the compiler would contract where.
#pragma once
#include <cuda_runtime.h>
#include "f110_dev.h"

struct Scan32 {
    float gx0, gy0, d0, vel, vel_abs;
    unsigned long long t0_fix;        // 32.32 fixed-point table index of beam 0
    unsigned ia, step;                // noise key: global agent id, noise step | episode << 20
    float r_gate;                     // iTTC can only fire below this range: max(side_dist) + 1.2*thresh*|v|
    bool noisy;
};

// What one thread prepares for one scan (shared memory, one entry per warp of the CTA).
struct ScanJob {
    Scan32 c;
    int ia;              // agent index within this handle; -1: nothing to do (past the end, or masked out of this step)
    unsigned nlook0;     // the lookup at the scan origin (COUNT)
};

// EDT32 encoding: distance in cells; cells with distance <= eps (obstacles: an exact EDT is 0 or >= 1 cell) hold
// SENT instead.  Adding SENT pushes the accumulated range past every limit, so the marching loop needs a single
// compare (total <= limit); the caller subtracts SENT again (error < 1e-4 cell).
// SENT = d.sent32: the power of two above max_range in cells + 2 (at least 1024); free cells are capped below it.

// general (checked) lookup: out of the map -> value of the last cell, like the reference (laser_models.py:79-81)
template <bool COUNT>
static __device__ __forceinline__ float edt_checked(const F110Dev &d, float gx, float gy, unsigned &nlook)
{
    int c = __float2int_rd(gx), r = __float2int_rd(gy);
    if (COUNT) nlook++;
    if ((unsigned)c >= (unsigned)d.map_w || (unsigned)r >= (unsigned)d.map_h) return d.oob32;
    return __ldg(d.edt32 + (unsigned)r * (unsigned)d.map_w + (unsigned)c);
}

static __device__ __forceinline__ void scan_setup(const F110Dev &d, double px, double py, double yaw, Scan32 &c)
{
    const double xt = __dsub_rn(px, d.ox), yt = __dsub_rn(py, d.oy);
    c.gx0 = (float)__dmul_rn(__dadd_rn(__dmul_rn(xt, d.oc), __dmul_rn(yt, d.os)), d.inv_res);
    c.gy0 = (float)__dmul_rn(__dsub_rn(__dmul_rn(yt, d.oc), __dmul_rn(xt, d.os)), d.inv_res);
    // start of the float beam index (laser_models.py:166-172), then 32.32 fixed point
    const double td = (double)d.theta_dis;
    double t = __ddiv_rn(__dmul_rn(td, __dsub_rn(yaw, __dmul_rn(d.fov, 0.5))), 6.283185307179586);
    t = fmod(t, td);
    while (t < 0) t = __dadd_rn(t, td);
    c.t0_fix = (unsigned long long)__dmul_rn(t, 4294967296.0);
}

// iTTC gate: 0 <= (r - side) / (v cos) < thresh is impossible for r >= max(side) + 1.2 thresh |v|;  vel == 0: no iTTC (laser_models.py:206)
static __device__ __forceinline__ float gate32(const F110Dev &d, float vel)
{
    return (vel != 0.0f) ? fmaf(1.2f * d.ttc_thresh32, fabsf(vel), d.side_max32) : -1.0f;
}

// per-scan preparation of the env kernels: pose -> origin cell and its distance, beam table index, iTTC gate, noise key
template <bool COUNT>
static __device__ __forceinline__ void prepare_env_scan(const F110Dev &d, int ia, ScanJob &j)
{
    const int NA = d.n_envs * d.n_agents;
    const int e = ia / d.n_agents;
    j.ia = -1;
    j.nlook0 = 0;
    if (d.step_mask && !d.step_mask[e]) return;
    j.ia = ia;
    Scan32 &c = j.c;
    const double vel = d.state[3 * (size_t)NA + ia];
    scan_setup(d, d.state[ia], d.state[(size_t)NA + ia], d.state[4 * (size_t)NA + ia], c);
    c.d0 = edt_checked<COUNT>(d, c.gx0, c.gy0, j.nlook0);
    c.vel = (float)vel;
    c.vel_abs = fabsf(c.vel);
    // (a velocity that is not zero in fp64 but rounds to 0 in fp32 -- below 1e-45 m/s -- takes no iTTC either)
    c.r_gate = gate32(d, c.vel);
    c.noisy = d.noise_mode == 2;
    c.ia = (unsigned)(ia + d.env_offset * d.n_agents);   // global agent id: the noise stream does not depend on the sharding
    c.step = (unsigned)d.noise_step[e] + ((unsigned)d.episode_count[e] << 20);
}

// The 32-byte record the sort leaves for the ray-cast: two float4, in launch order.
static __device__ __forceinline__ void scan_rec_store(float4 *__restrict__ rec, const ScanJob &j)
{
    rec[0] = make_float4(j.c.gx0, j.c.gy0, j.c.d0, j.c.vel);
    rec[1] = make_float4(__uint_as_float((unsigned)j.c.t0_fix), __uint_as_float((unsigned)(j.c.t0_fix >> 32)), __uint_as_float(j.c.step),
                         __int_as_float(j.ia));
}
static __device__ __forceinline__ void scan_rec_load(const F110Dev &d, const float4 *__restrict__ rec, ScanJob &j)
{
    const float4 a = __ldg(rec), b = __ldg(rec + 1);
    Scan32 &c = j.c;
    c.gx0 = a.x; c.gy0 = a.y; c.d0 = a.z; c.vel = a.w;
    c.vel_abs = fabsf(a.w);
    c.r_gate = gate32(d, a.w);
    c.t0_fix = (unsigned long long)__float_as_uint(b.x) | ((unsigned long long)__float_as_uint(b.y) << 32);
    c.step = __float_as_uint(b.z);
    j.ia = __float_as_int(b.w);
    j.nlook0 = 0;
    c.noisy = d.noise_mode == 2;
    c.ia = (unsigned)(j.ia + d.env_offset * d.n_agents);
}
