// fp64 kernels of the batched F1TENTH step for sm_100a.  Compiled with -fmad=false: the reference
// (Numba) emits no FMA, and the parity contract needs the same roundings in the same order.
//
//   k_dynamics   thread per (env, agent): steer FIFO + pid + RK4/Euler + yaw wrap   (base_classes.py:254-402)
//   k_raycast64  warp per (env, agent) scan: sphere tracing on the fp64 EDT, noise, iTTC + zeroing
//                                                                          (laser_models.py:107-217, base_classes.py:227-252)
//   k_agents     warp per (env, agent), A >= 2: GJK vs every other agent + opponent ray-cast
//                                                                          (collision_models.py:114-260, laser_models.py:250-346)
//   k_finalize   thread per env: collision flags, lap bookkeeping, done      (f110_env.py:202-244,292-299)
//   k_reset      thread per (env, agent)                                     (base_classes.py:181-202, f110_env.py:318-329)
#include <cuda_runtime.h>
#include <math_constants.h>
#include <string.h>
#include "f110_dev.h"

#define FULL 0xffffffffu
static __device__ __forceinline__ double TWO_PI() { return 6.283185307179586; }
static __device__ __forceinline__ double PI_() { return 3.141592653589793; }

// ------------------------------------------------------------------------------------------------
// dynamics (dynamic_models.py)
// ------------------------------------------------------------------------------------------------
// dynamic_models.py:29-60
static __device__ __forceinline__ double accl_constraints(double vel, double accl, double v_switch, double a_max,
                                                          double v_min, double v_max)
{
    double pos_limit = (vel > v_switch) ? a_max * v_switch / vel : a_max;
    if ((vel <= v_min && accl <= 0) || (vel >= v_max && accl >= 0)) accl = 0.;
    else if (accl <= -a_max) accl = -a_max;
    else if (accl >= pos_limit) accl = pos_limit;
    return accl;
}
// dynamic_models.py:62-87
static __device__ __forceinline__ double steering_constraint(double steer, double sv, double s_min, double s_max,
                                                             double sv_min, double sv_max)
{
    if ((steer <= s_min && sv <= 0) || (steer >= s_max && sv >= 0)) sv = 0.;
    else if (sv <= sv_min) sv = sv_min;
    else if (sv >= sv_max) sv = sv_max;
    return sv;
}

struct VehP {
    double mu, C_Sf, C_Sr, lf, lr, h, m, I, s_min, s_max, sv_min, sv_max, v_switch, a_max, v_min, v_max;
};

// dynamic_models.py:123-176 (kinematic branch :152-160 -> :90-121 re-applies the constraints)
static __device__ __forceinline__ void dynamics_st(const double x[7], double sv_in, double ac_in, const VehP &p,
                                                   double f[7])
{
    const double g = 9.81;
    double u0 = steering_constraint(x[2], sv_in, p.s_min, p.s_max, p.sv_min, p.sv_max);
    double u1 = accl_constraints(x[3], ac_in, p.v_switch, p.a_max, p.v_min, p.v_max);
    const double lf = p.lf, lr = p.lr, mu = p.mu, m = p.m, I = p.I, h = p.h, C_Sf = p.C_Sf, C_Sr = p.C_Sr;
    if (fabs(x[3]) < 0.5) {
        double lwb = lf + lr;
        double k0 = steering_constraint(x[2], u0, p.s_min, p.s_max, p.sv_min, p.sv_max);
        double k1 = accl_constraints(x[3], u1, p.v_switch, p.a_max, p.v_min, p.v_max);
        double s4, c4;
        sincos(x[4], &s4, &c4);
        double t2 = tan(x[2]);
        double c2 = cos(x[2]);
        f[0] = x[3] * c4;
        f[1] = x[3] * s4;
        f[2] = k0;
        f[3] = k1;
        f[4] = x[3] / lwb * t2;
        f[5] = u1 / lwb * t2 + x[3] / (lwb * (c2 * c2)) * u0;
        f[6] = 0.;
    } else {
        double s4, c4;
        sincos(x[6] + x[4], &s4, &c4);
        f[0] = x[3] * c4;
        f[1] = x[3] * s4;
        f[2] = u0;
        f[3] = u1;
        f[4] = x[5];
        f[5] = -mu * m / (x[3] * I * (lr + lf)) * (lf * lf * C_Sf * (g * lr - u1 * h) + lr * lr * C_Sr * (g * lf + u1 * h)) * x[5]
             + mu * m / (I * (lr + lf)) * (lr * C_Sr * (g * lf + u1 * h) - lf * C_Sf * (g * lr - u1 * h)) * x[6]
             + mu * m / (I * (lr + lf)) * lf * C_Sf * (g * lr - u1 * h) * x[2];
        f[6] = (mu / (x[3] * x[3] * (lr + lf)) * (C_Sr * (g * lf + u1 * h) * lr - C_Sf * (g * lr - u1 * h) * lf) - 1) * x[5]
             - mu / (x[3] * (lr + lf)) * (C_Sr * (g * lf + u1 * h) + C_Sf * (g * lr - u1 * h)) * x[6]
             + mu / (x[3] * (lr + lf)) * (C_Sf * (g * lr - u1 * h)) * x[2];
    }
}

// pid (dynamic_models.py:178-221): speed / steer set points -> acceleration, steering velocity
static __device__ __forceinline__ void pid_ctrl(double speed, double steer, double cur_speed, double cur_steer, double max_sv,
                                                double max_a, double max_v, double min_v, double &accl, double &sv)
{
    double steer_diff = steer - cur_steer;
    if (fabs(steer_diff) > 1e-4) sv = (steer_diff / fabs(steer_diff)) * max_sv;
    else sv = 0.0;
    double vel_diff = speed - cur_speed;
    double kp;
    if (cur_speed > 0.) kp = (vel_diff > 0) ? 10.0 * max_a / max_v : 10.0 * max_a / (-min_v);
    else kp = (vel_diff > 0) ? 2.0 * max_a / max_v : 2.0 * max_a / (-min_v);
    accl = kp * vel_diff;
}

static __device__ __forceinline__ unsigned pool_index(int e, long long episode, int n_pool)
{
    unsigned long long z = (unsigned long long)e * 0x9E3779B97F4A7C15ull + (unsigned long long)episode * 0xBF58476D1CE4E5B9ull;
    z ^= z >> 31; z *= 0x94D049BB133111EBull; z ^= z >> 29;
    return (unsigned)(z % (unsigned long long)n_pool);
}

// RaceCar.update_pose without the scan (base_classes.py:254-402) for agent ia: steering FIFO, pid, RK4 / Euler, yaw wrap
// cmd: the agent's (steer, velocity) command when the caller already holds it (the single-env step gets it as a kernel
// argument instead of reading the caller's host buffer over PCIe), else null -> d.actions
static __device__ __forceinline__ void dynamics_agent(const F110Dev &d, int ia, const double *cmd = nullptr)
{
    const int NA = d.n_envs * d.n_agents;
    int e = ia / d.n_agents, a = ia - e * d.n_agents;
    if (d.step_mask && !d.step_mask[e]) return;

    double x[7];
    double raw_steer = cmd ? cmd[0] : d.actions[2 * ia + 0];
    double vel_cmd = cmd ? cmd[1] : d.actions[2 * ia + 1];
    int slen = d.steer_len[ia];
    if (d.auto_reset && d.reset_pending[e]) {
        // F110Env.reset on the device: the pending env restarts from a pool pose and this step is
        // the reset's zero-action step (f110_env.py:318-336)
        unsigned k = pool_index(e + d.env_offset, d.episode_count[e], d.n_pool);
        const double *pp = d.pool + ((size_t)k * d.n_agents + a) * 3;
        x[0] = pp[0]; x[1] = pp[1]; x[2] = 0.; x[3] = 0.; x[4] = pp[2]; x[5] = 0.; x[6] = 0.;
        raw_steer = 0.; vel_cmd = 0.; slen = 0;
        d.near_start[ia] = 1; d.toggles[ia] = 0; d.start_x[ia] = pp[0]; d.start_y[ia] = pp[1];
        if (a == 0) {
            const double *pe = d.pool + ((size_t)k * d.n_agents + d.ego_idx) * 3;
            double sn, cs;
            sincos(-pe[2], &sn, &cs);
            double *R = d.start_rot + 4 * (size_t)e;
            R[0] = cs; R[1] = -sn; R[2] = sn; R[3] = cs;
            d.cur_time[e] = 0.0;
            d.noise_step[e] = 0;
        }
    } else {
#pragma unroll
        for (int i = 0; i < 7; i++) x[i] = d.state[(size_t)i * NA + ia];
    }

    // every load of the step before its first store (nothing tells the compiler that the arrays do not alias, and a load
    // it must keep behind a store is one more trip to L2 on the single-env path, where thread 0 of k_step_fused64 runs this alone)
    const double *pp = d.params + F110_NP * a;
    VehP p = { pp[0], pp[1], pp[2], pp[3], pp[4], pp[5], pp[6], pp[7], pp[8], pp[9], pp[10], pp[11], pp[12], pp[13], pp[14], pp[15] };

    // steering delay FIFO of depth 2 (base_classes.py:268-276)
    double b0 = d.steer_buf[ia], b1 = d.steer_buf[NA + ia];
    double steer;
    if (slen < 2) { steer = 0.; slen += 1; }
    else { steer = b1; }
    d.steer_buf[NA + ia] = b0;
    d.steer_buf[ia] = raw_steer;
    d.steer_len[ia] = (uint8_t)slen;

    double sv, accl;
    pid_ctrl(vel_cmd, steer, x[3], x[2], p.sv_max, p.a_max, p.v_max, p.v_min, accl, sv);

    const double dt = d.timestep;
    if (d.integrator == 1) {   // RK4 (base_classes.py:282-371)
        double k1[7], k2[7], k3[7], k4[7], xs[7];
        dynamics_st(x, sv, accl, p, k1);
#pragma unroll
        for (int i = 0; i < 7; i++) xs[i] = x[i] + dt * (k1[i] / 2);
        dynamics_st(xs, sv, accl, p, k2);
#pragma unroll
        for (int i = 0; i < 7; i++) xs[i] = x[i] + dt * (k2[i] / 2);
        dynamics_st(xs, sv, accl, p, k3);
#pragma unroll
        for (int i = 0; i < 7; i++) xs[i] = x[i] + dt * k3[i];
        dynamics_st(xs, sv, accl, p, k4);
        const double w = dt * (1.0 / 6.0);
#pragma unroll
        for (int i = 0; i < 7; i++) x[i] = x[i] + w * (((k1[i] + 2 * k2[i]) + 2 * k3[i]) + k4[i]);
    } else {                   // Euler (:373-393)
        double f[7];
        dynamics_st(x, sv, accl, p, f);
#pragma unroll
        for (int i = 0; i < 7; i++) x[i] = x[i] + dt * f[i];
    }
    // single yaw wrap (:398-402)
    if (x[4] > TWO_PI()) x[4] = x[4] - TWO_PI();
    else if (x[4] < 0) x[4] = x[4] + TWO_PI();

#pragma unroll
    for (int i = 0; i < 7; i++) d.state[(size_t)i * NA + ia] = x[i];
    d.pose_pre[ia] = x[0];
    d.pose_pre[NA + ia] = x[1];
    d.pose_pre[2 * (size_t)NA + ia] = x[4];
}

// 8 CTAs of 128 threads per SM (64 registers, 500 bytes of spill) instead of the compiler's 128 registers / 4 CTAs: round 1 left the
// spilling version untimed; measured at 2^20 agents the step's non-ray-cast part is 0.298 ms against 0.329 (6 CTAs 0.305, 10 CTAs
// 0.325; profiles/r2g_dynamics_occupancy_ab.txt) -- the dependent FP64 chain wants resident warps more than registers
__global__ void __launch_bounds__(128, 8) k_dynamics(const F110Dev d)
{
    int ia = blockIdx.x * blockDim.x + threadIdx.x;
    if (ia < d.n_envs * d.n_agents) dynamics_agent(d, ia);
}

void f110_launch_dynamics(const F110Dev &d, cudaStream_t st)
{
    int NA = d.n_envs * d.n_agents;
    k_dynamics<<<(NA + 127) / 128, 128, 0, st>>>(d);
}

// ------------------------------------------------------------------------------------------------
// lidar: sphere tracing against the EDT (laser_models.py:56-186), fp64 parity path
// ------------------------------------------------------------------------------------------------
// int(x_rot / resolution) of xy_2_rc (laser_models.py:83-84), bit-exact, for 0 <= v / res < 2^22.
// Resolution a power of two: v / res is an exact scaling, and floor of it is what v + 2^52 * res rounded toward
// zero leaves in its low mantissa word -- one FP64 instruction.  Otherwise multiply by the rounded reciprocal and
// add 2^22 (round toward zero): the mantissa then holds floor(q * 2^30), i.e. the cell in bits 30.. and 30 fraction
// bits below, read with integer instructions (the FP64 pipe is what bounds this kernel).  The product and the true
// IEEE quotient can truncate differently only when q sits within q * 2^-52 < 2^-30 of an integer: those lookups
// (fraction bits < 2 or > 2^30 - 3) redo the division.
template <bool POW2>
static __device__ __forceinline__ int cell_index(double v, const F110Dev &d)
{
    if (POW2) return __double2loint(__dadd_rz(v, d.magic64));
    // one fused instruction: the exact product plus 2^22, truncated.  Where that differs from truncating the ROUNDED
    // product (the last of the 30 fraction bits, rarely) the value is inside the redo band below anyway.
    double t = __fma_rz(v, d.inv_res, 4194304.0);
    unsigned lo = (unsigned)__double2loint(t), hi = (unsigned)__double2hiint(t);
    int c = (int)(__funnelshift_l(lo, hi, 2) & 0x3FFFFFu);
    if (((lo + 2u) & 0x3FFFFFFFu) < 4u) c = (int)__ddiv_rn(v, d.res);
    return c;
}

// distance_transform(x, y) (laser_models.py:89-104).  CHECKED: with the bounds test of xy_2_rc
// (out of bounds -> (-1,-1) -> last cell); unchecked: caller proved the point is inside the map.
template <bool ROT, bool POW2, bool CHECKED>
static __device__ __forceinline__ double edt_lookup(const F110Dev &d, double x, double y)
{
    double xt = x - d.ox, yt = y - d.oy;
    double xr, yr;
    if (ROT) {
        xr = xt * d.oc + yt * d.os;
        yr = -xt * d.os + yt * d.oc;
    } else {   // origin yaw == 0: x*1 + y*0 is x (bit-exact for finite inputs)
        xr = xt; yr = yt;
    }
    if (CHECKED) {
        if (xr < 0 || xr >= d.xmax || yr < 0 || yr >= d.ymax) return d.oob64;
    }
    int c = cell_index<POW2>(xr, d);
    int r = cell_index<POW2>(yr, d);
    F110_CHECK(d.violations, (unsigned)r < (unsigned)d.map_h && (unsigned)c < (unsigned)d.map_w, F110_V_EDT64);     // the hoisted bounds test must have been right
    return __ldg(d.edt64 + ((unsigned)r * (unsigned)d.map_w + (unsigned)c));
}

// Start of the float beam index of get_scan (laser_models.py:166-172)
static __device__ __forceinline__ double theta_start(const F110Dev &d, double yaw)
{
    double td = (double)d.theta_dis;
    double t = __ddiv_rn(td * (yaw - d.fov / 2.), 2. * PI_());
    t = fmod(t, td);
    while (t < 0) t += td;
    return t;
}

// Per-scan data, one row per warp in shared memory (at 40 registers a thread has no room for them, and ptxas
// would otherwise recompute them beam after beam).  x_lo .. y_hi hoist the bounds test out of the loop: distances
// (metres, map frame) from the scan origin to the map border minus a 2-cell margin.  While the accumulated range of
// a ray is below the range at which it crosses that margin, every lookup is inside the map and the test of
// xy_2_rc (:79) is provably false -- skipping it changes no result.
struct __align__(16) Scan64 {
    float x_lo, x_hi, y_lo, y_hi;
    double px, py, d0, t0, vel, r_gate;
    double *out;
    const double *noise;
    float max_fast, pad_;
};

// Table index of beam i.  The reference accumulates theta_index += increment beam after beam
// (laser_models.py:180-184) and truncates; the closed form t0 + i*inc differs from that running
// sum by < n_beams * 2^-43 (2.5e-10 for 2160 beams), so int() can only differ within a band around the integers
// that grows with the beam count (theta_band, 2.8e-9 up to 2160 beams) -- those beams replay the sequential sum exactly.
// The reference's running sum itself (theta_index += increment; wrap), for the beams whose closed form is too close to
// an integer to trust.  Not as rare as 4e-9 per beam suggests: a car stopped by iTTC has yaw == 0 exactly (check_ttc
// zeroes it), which puts the LAST beam of a 2 pi scan on an integer -- every scan of every crashed car replays B - 1
// additions.  The additions are the reference's, in its order; the wrap test after each of them is skipped for stretches
// in which it cannot fire (t + n * inc stays 2 increments below theta_dis), which leaves a bare chain of DADDs.
// Out of line: it must not cost the beam loop registers.
static __device__ __noinline__ int theta_index_replay(double t, int i, double inc, double td)
{
    int j = 0;
    while (j < i) {
        int n = (int)__ddiv_rn(td - t, inc) - 2;
        n = min(n, i - j);
        for (int k = 0; k < n; k++) t += inc;
        if (n > 0) j += n;
        if (j < i) {
            t += inc;
            while (t >= td) t -= td;
            j++;
        }
    }
    return (int)t;
}

static __device__ __forceinline__ int theta_index_of(const F110Dev &d, const Scan64 *c, int i, double a)
{
    // a ~ t0 + i*inc (closed form, accumulated per round: a few 1e-12 off at most)
    double td = (double)d.theta_dis;
    // integer part and 30 fraction bits of a, as in cell_index
    double tz = __dadd_rz(a, 4194304.0);
    unsigned lo = (unsigned)__double2loint(tz), hi = (unsigned)__double2hiint(tz);
    int k = (int)(__funnelshift_l(lo, hi, 2) & 0x3FFFFFu);
    // band = d.theta_band units of 2^-30 on either side of an integer: covers the drift of the reference's running sum
    // (<= half an ulp of 2048 = 1.1e-13 per addition) for the handle's beam count, see f110_create
    if (((lo + d.theta_band) & 0x3FFFFFFFu) < 2u * d.theta_band) return theta_index_replay(c->t0, i, d.theta_inc, td);
    while (k >= d.theta_dis) k -= d.theta_dis;
    return k;
}

// `writer`: this thread stores the row (lane 0 when a whole warp prepares one scan, every calling thread when
// threads prepare one scan each).  The caller synchronises before the row is read.
template <bool ROT, bool POW2, bool STEP>
static __device__ __forceinline__ void scan_setup64(const F110Dev &d, double px, double py, double yaw, double vel,
                                                    double *out, const double *noise_row, Scan64 *c, bool writer)
{
    double xt = px - d.ox, yt = py - d.oy;
    double xr = ROT ? xt * d.oc + yt * d.os : xt;
    double yr = ROT ? -xt * d.os + yt * d.oc : yt;
    float margin = 2.0f * (float)d.res;
    float x_lo = margin - (float)xr, x_hi = (float)d.xmax - margin - (float)xr;
    float y_lo = margin - (float)yr, y_hi = (float)d.ymax - margin - (float)yr;
    double d0 = edt_lookup<ROT, POW2, true>(d, px, py);
    double t0 = theta_start(d, yaw);
    if (writer) {
        c->x_lo = x_lo; c->x_hi = x_hi; c->y_lo = y_lo; c->y_hi = y_hi;
        c->max_fast = (x_lo <= 0.f && x_hi >= 0.f && y_lo <= 0.f && y_hi >= 0.f) ? (float)d.max_range : -1.0f;
        c->px = px; c->py = py; c->d0 = d0; c->t0 = t0; c->vel = vel;
        // iTTC (laser_models.py:208-213) needs 0 <= (r - side)/(v cos) < thresh, impossible for r >= side_max + 1.2*thresh*|v|
        c->r_gate = (STEP && vel != 0.0) ? (double)d.side_max32 * 1.0001 + 1.2 * d.ttc_thresh * fabs(vel) : -1.0;
        c->out = out; c->noise = noise_row;
    }
}

// trace_ray (laser_models.py:107-146) given the first lookup d0 at the scan origin
template <bool ROT, bool POW2, bool COUNT>
static __device__ __forceinline__ double trace_ray(const F110Dev &d, const Scan64 *c, int ti, unsigned &nlook)
{
    F110_CHECK(d.violations, ti >= 0 && ti < d.theta_dis, F110_V_TABLE);
    const double sn = __ldg(d.sines + ti), cs = __ldg(d.cosines + ti);
    // fp32 estimate (conservative through the margin) of the range where the ray nears the border
    const float4 dir = __ldg(d.dir32 + ti);
    const float4 lim = *reinterpret_cast<const float4 *>(c);
    float tx = (dir.x > 0.f ? lim.y : lim.x) * dir.z;
    float ty = (dir.y > 0.f ? lim.w : lim.z) * dir.w;
    const double t_fast = (double)fminf(fminf(tx, ty), c->max_fast);
    double x = c->px, y = c->py, dist = c->d0, total = dist;
    // The two tests of the fast loop compare high words with integer instructions (off the FP64 pipe): both are
    // one-sided -- hi(dist) > hi(eps) implies dist > eps, hi(total) < hi(t_fast) implies total < t_fast (negative
    // values have a negative high word) -- and whatever they leave undecided the exact loop below decides.
    const int eps_hi = __double2hiint(d.eps), fast_hi = __double2hiint(t_fast);
    while (__double2hiint(dist) > eps_hi && __double2hiint(total) < fast_hi) {
        x += dist * cs;
        y += dist * sn;
        dist = edt_lookup<ROT, POW2, false>(d, x, y);
        total += dist;
        if (COUNT) nlook++;
    }
    while (dist > d.eps && total <= d.max_range) {
        x += dist * cs;
        y += dist * sn;
        dist = edt_lookup<ROT, POW2, true>(d, x, y);
        total += dist;
        if (COUNT) nlook++;
    }
    if (total > d.max_range) total = d.max_range;
    return total;
}

// iTTC test of one beam (laser_models.py:208-213).  The IEEE division is only issued when the
// quotient can be inside [0, thresh): |num| > 1.2*thresh*|den| already decides "no".
static __device__ __forceinline__ bool ttc_hit(double range, double side, double vel, double bcos, double thresh)
{
    double proj = vel * bcos;
    double num = range - side;
    if (fabs(num) > 1.2 * thresh * fabs(proj)) return false;
    double ttc = __ddiv_rn(num, proj);
    return (ttc < thresh) && (ttc >= 0.0);
}

// One warp per scan, 32 consecutive beams in lockstep per round (adjacent beams hit adjacent EDT
// cells, so a warp-level gather touches few cache lines); lanes whose ray has converged idle until
// the round's slowest ray is done.
// A scan may also be split over the warps of a CTA: below a few thousand scans one warp per scan leaves the GPU waiting
// on 365 dependent gathers in a row per warp.  The warps then claim rounds one at a time from a shared counter
// (`next`): rounds differ in length, and the CTA ends with its slowest warp.
template <bool ROT, bool POW2, bool STEP, bool COUNT>
static __device__ __forceinline__ void scan_beam(const F110Dev &d, const Scan64 *c, int i, double a, bool &hit_any, unsigned &nlook)
{
    int ti = theta_index_of(d, c, i, a);
    double rng = trace_ray<ROT, POW2, COUNT>(d, c, ti, nlook);
    const double *np_ = c->noise;
    F110_CHECK(d.violations, i >= 0 && i < d.n_beams, F110_V_BEAM);
    F110_CHECK(d.violations, !np_ || (np_ >= d.noise && np_ + d.n_beams <= d.noise + d.noise_rows * d.n_beams), F110_V_NOISE);
    if (np_) rng += __ldg(np_ + i);
    c->out[i] = rng;
    if (STEP && rng < c->r_gate)
        hit_any |= ttc_hit(rng, __ldg(d.side_dist + i), c->vel, __ldg(d.beam_cos + i), d.ttc_thresh);
}

// a = closed-form beam index t0 + i*inc (theta_index_of replays the reference's running sum where it matters)
template <bool ROT, bool POW2, bool STEP, bool COUNT>
static __device__ __forceinline__ void scan_rounds(const F110Dev &d, const Scan64 *c, bool &hit_any, unsigned &nlook)
{
    const int lane = threadIdx.x & 31;
    if (COUNT) nlook += (lane == 0);
    double a = c->t0 + (double)lane * d.theta_inc;
    for (int i = lane; i < d.n_beams; i += 32) {
        scan_beam<ROT, POW2, STEP, COUNT>(d, c, i, a, hit_any, nlook);
        a += 32.0 * d.theta_inc;
    }
}

// the CTA-per-scan version: warps claim rounds from `next`
template <bool ROT, bool POW2, bool STEP, bool COUNT>
static __device__ __forceinline__ void scan_rounds_claimed(const F110Dev &d, const Scan64 *c, bool &hit_any, unsigned &nlook, int *next,
                                                           bool counts_origin = (threadIdx.x == 0))
{
    const int lane = threadIdx.x & 31;
    if (COUNT) nlook += counts_origin;
    const int rounds = (d.n_beams + 31) >> 5;
    for (;;) {
        int r = 0;
        if (lane == 0) r = atomicAdd(next, 1);
        r = __shfl_sync(FULL, r, 0);
        if (r >= rounds) break;
        const int i = lane + 32 * r;
        if (i < d.n_beams) scan_beam<ROT, POW2, STEP, COUNT>(d, c, i, c->t0 + (double)i * d.theta_inc, hit_any, nlook);
    }
}

// per 262144 scans (example_map, power-of-two resolution): 64 regs 4.34 ms, 48 regs 4.08 ms, 40 regs 3.92 ms with the per-scan
// constants in registers; with the shared-memory row 40 regs 3.37 ms, 32 regs (MINB = 8, no spill) 3.09 ms
// SPLIT: one CTA (8 warps) per scan instead of one warp per scan
// the env kernels' per-scan preparation: state -> Scan64 row
template <bool ROT, bool POW2>
static __device__ __forceinline__ void prepare_env_scan64(const F110Dev &d, int ia, Scan64 *c, bool writer)
{
    const int NA = d.n_envs * d.n_agents;
    const int e = ia / d.n_agents;
    double px = d.state[ia], py = d.state[(size_t)NA + ia];
    double vel = d.state[3 * (size_t)NA + ia], yaw = d.state[4 * (size_t)NA + ia];
    const double *noise_row = nullptr;
    if (d.noise_mode == 1) {
        long long row = d.noise_step[e];
        if (row >= d.noise_rows) row %= d.noise_rows;
        noise_row = d.noise + row * d.n_beams;
    }
    scan_setup64<ROT, POW2, true>(d, px, py, yaw, vel, (double *)d.scans + (size_t)ia * d.n_beams, noise_row, c, writer);
}

// One warp per scan (!SPLIT): threads 0..7 of the CTA prepare its 8 scans side by side (fp64 division, fmod and the
// origin lookup: a few hundred instructions that 32 identical lanes per warp would otherwise repeat 8 times over).
// SPLIT: a scan is shared by `wps` warps (8, 4 or 2: the CTA holds 8 / wps scans) that claim its rounds one at a time.
// Small launches pick wps so that the launch is about one wave of CTAs (f110_split_wps): with fewer, longer-lived
// warps per scan more scans are in flight per SM at a time -- neighbours on the track, whose rays walk the same EDT
// lines (ncu at 4096 scans, 8 warps per scan: L1 hit rate 52 % against 92 % in large launches) -- and no CTA of a last,
// partly filled wave is left running alone.  The arithmetic per beam does not depend on wps.
template <bool ROT, bool POW2, bool COUNT, int MINB, bool SPLIT>
__global__ void __launch_bounds__(256, MINB) k_raycast64(const F110Dev d, int env_begin, int env_count, int wps)
{
    const int NA = d.n_envs * d.n_agents;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    __shared__ Scan64 rows[8];
    __shared__ int row_ia[8];
    __shared__ int next_round[8], hit_any[8];
    const int per_cta = SPLIT ? 8 / wps : 8;                  // scans of this CTA
    const int slot = SPLIT ? warp / wps : warp;
    if (threadIdx.x < per_cta) {
        const int total = env_count * d.n_agents;
        int cta = (int)blockIdx.x;
        if (SPLIT) cta = f110_split_scan(cta, (total + per_cta - 1) / per_cta, d.n_sm);
        const int w = cta * per_cta + (int)threadIdx.x;
        int a = -1;
        if (w < total && (!SPLIT || cta < (total + per_cta - 1) / per_cta)) {
            a = env_begin * d.n_agents + w;
            if (d.order) a = d.order[a];      // spatial launch order (f110_order.cu)
            if (d.step_mask && !d.step_mask[a / d.n_agents]) a = -1;
            else prepare_env_scan64<ROT, POW2>(d, a, rows + threadIdx.x, true);
        }
        row_ia[threadIdx.x] = a;
        next_round[threadIdx.x] = 0;
        hit_any[threadIdx.x] = 0;
    }
    __syncthreads();
    const int ia = row_ia[slot];
    Scan64 *c = rows + slot;
    bool hit = false;
    unsigned nlook = 0;
    if (SPLIT) {
        // own scan first; a warp whose scan has no rounds left takes rounds of the CTA's other scans (scans differ by 2 x in
        // their number of lookups, and the CTA ends with its slowest one).  Which warp casts a round changes no result.
        const int n_steal = d.no_steal ? 1 : per_cta;
        for (int k = 0; k < n_steal; k++) {
            int s2 = slot + k;
            if (s2 >= per_cta) s2 -= per_cta;
            if (row_ia[s2] < 0) continue;
            bool h = false;
            scan_rounds_claimed<ROT, POW2, true, COUNT>(d, rows + s2, h, nlook, &next_round[s2], k == 0 && lane == 0 && warp % wps == 0);
            if (__any_sync(FULL, h) && lane == 0) atomicOr(&hit_any[s2], 1);
        }
        __syncthreads();                        // every warp has read the state by now
        hit = hit_any[slot] != 0;
        if (ia < 0) return;
    } else {
        if (ia < 0) return;
        scan_rounds<ROT, POW2, true, COUNT>(d, c, hit, nlook);
        hit = __any_sync(FULL, hit);
    }
    // RaceCar.check_ttc (base_classes.py:241-250): state[3:] = 0 -- velocity, YAW, yaw rate, slip
    if (lane == 0 && (!SPLIT || warp % wps == 0)) {
        if (hit) {
            d.state[3 * (size_t)NA + ia] = 0.; d.state[4 * (size_t)NA + ia] = 0.;
            d.state[5 * (size_t)NA + ia] = 0.; d.state[6 * (size_t)NA + ia] = 0.;
        }
        d.in_collision[ia] = hit ? 1 : 0;
    }
    if (COUNT) {
#pragma unroll
        for (int o = 16; o; o >>= 1) nlook += __shfl_xor_sync(FULL, nlook, o);
        if (lane == 0) atomicAdd(d.lookups, (unsigned long long)nlook);
    }
}

template <bool ROT, bool POW2, bool COUNT, bool SPLIT>
__global__ void __launch_bounds__(256) k_scan_only64(const F110Dev d, const double *__restrict__ poses, int n,
                                                     double *__restrict__ out, const int *__restrict__ order)
{
    __shared__ Scan64 rows[8];
    __shared__ int row_ok[8];
    __shared__ int next_round;
    Scan64 *c;
    if (SPLIT) {
        int w = f110_split_scan((int)blockIdx.x, n, d.n_sm);
        if (w >= n) return;
        if (order) w = order[w];
        c = rows;
        if (threadIdx.x < 32)
            scan_setup64<ROT, POW2, false>(d, poses[3 * w], poses[3 * w + 1], poses[3 * w + 2], 0.0, out + (size_t)w * d.n_beams, nullptr, c,
                                           threadIdx.x == 0);
        if (threadIdx.x == 0) next_round = 0;
        __syncthreads();
    } else {
        if (threadIdx.x < 8) {
            int w = (int)blockIdx.x * 8 + (int)threadIdx.x;
            row_ok[threadIdx.x] = w < n;
            if (w < n) {
                if (order) w = order[w];
                scan_setup64<ROT, POW2, false>(d, poses[3 * w], poses[3 * w + 1], poses[3 * w + 2], 0.0, out + (size_t)w * d.n_beams, nullptr,
                                               rows + threadIdx.x, true);
            }
        }
        __syncthreads();
        if (!row_ok[threadIdx.x >> 5]) return;
        c = rows + (threadIdx.x >> 5);
    }
    bool hit = false;
    unsigned nlook = 0;
    if (SPLIT) scan_rounds_claimed<ROT, POW2, false, COUNT>(d, c, hit, nlook, &next_round);
    else scan_rounds<ROT, POW2, false, COUNT>(d, c, hit, nlook);
    if (COUNT) {
#pragma unroll
        for (int o = 16; o; o >>= 1) nlook += __shfl_xor_sync(FULL, nlook, o);
        if ((threadIdx.x & 31) == 0) atomicAdd(d.lookups, (unsigned long long)nlook);
    }
}


#define DISPATCH64(KERNEL, SPLIT, GRID, ...)                                                      \
    do {                                                                                          \
        bool cnt = d.count_lookups != 0;                                                          \
        if (d.has_rot) {                                                                          \
            if (cnt) KERNEL<true, false, true, SPLIT><<<GRID, 256, 0, st>>>(__VA_ARGS__);         \
            else KERNEL<true, false, false, SPLIT><<<GRID, 256, 0, st>>>(__VA_ARGS__);            \
        } else if (d.res_pow2) {                                                                  \
            if (cnt) KERNEL<false, true, true, SPLIT><<<GRID, 256, 0, st>>>(__VA_ARGS__);         \
            else KERNEL<false, true, false, SPLIT><<<GRID, 256, 0, st>>>(__VA_ARGS__);            \
        } else {                                                                                  \
            if (cnt) KERNEL<false, false, true, SPLIT><<<GRID, 256, 0, st>>>(__VA_ARGS__);        \
            else KERNEL<false, false, false, SPLIT><<<GRID, 256, 0, st>>>(__VA_ARGS__);           \
        }                                                                                         \
    } while (0)

void f110_launch_raycast64(const F110Dev &d, int env_begin, int env_count, int variant, cudaStream_t st)
{
    long long warps = (long long)env_count * d.n_agents;
    if (warps == 0) return;
    const bool split = (warps < F110_SPLIT_BELOW64 || (variant & 4096)) && !(variant & 128);     // bit 12: split at any size
    const int wps = (variant & 2048) ? 8 : f110_split_wps((int)warps, d.n_sm);      // variant bit 11: always a CTA per scan
    int grid = split ? f110_split_grid((int)((warps * wps + 7) / 8), d.n_sm) : (int)((warps * 32 + 255) / 256);
    const bool cnt = d.count_lookups != 0;
#define RC64(R, P, MB, ALT)                                                                                         \
    do {                                                                                                          \
        if (split) {                                                                                              \
            if (cnt) k_raycast64<R, P, true, 1, true><<<grid, 256, 0, st>>>(d, env_begin, env_count, wps);             \
            else k_raycast64<R, P, false, MB, true><<<grid, 256, 0, st>>>(d, env_begin, env_count, wps);               \
        } else if (cnt) k_raycast64<R, P, true, 1, false><<<grid, 256, 0, st>>>(d, env_begin, env_count, 1);         \
        else if (variant & 8) k_raycast64<R, P, false, ALT, false><<<grid, 256, 0, st>>>(d, env_begin, env_count, 1); \
        else k_raycast64<R, P, false, MB, false><<<grid, 256, 0, st>>>(d, env_begin, env_count, 1);                  \
    } while (0)
    // registers per thread: 32 (8 CTAs per SM) where the kernel fits without spilling, else 40; variant bit 3 swaps the two
    if (d.has_rot) RC64(true, false, 6, 8);
    else if (d.res_pow2) RC64(false, true, 8, 6);
    else RC64(false, false, 6, 8);
#undef RC64
}

void f110_launch_scan_only64(const F110Dev &d, const double *poses, int n, double *out, const int *order, int variant, cudaStream_t st)
{
    if (n <= 0) return;
    if (n < F110_SPLIT_BELOW64 && !(variant & 128)) {
        DISPATCH64(k_scan_only64, true, f110_split_grid(n, d.n_sm), d, poses, n, out, order);
    } else {
        int grid = (int)(((long long)n * 32 + 255) / 256);
        DISPATCH64(k_scan_only64, false, grid, d, poses, n, out, order);
    }
}

// ------------------------------------------------------------------------------------------------
// multi-agent: GJK + opponent ray-cast (collision_models.py, laser_models.py:220-346)
// ------------------------------------------------------------------------------------------------
// collision_models.py:219-260: corners rl, rr, fr, fl of the car rectangle
static __device__ __forceinline__ void get_vertices(double x, double y, double th, double length, double width,
                                                    double v[4][2])
{
    double sn, c;
    sincos(th, &sn, &c);
    const double hl = length / 2, hw = width / 2;
    const double lx[4] = { -hl, -hl, hl, hl };
    const double ly[4] = { hw, -hw, -hw, hw };
#pragma unroll
    for (int k = 0; k < 4; k++) {
        v[k][0] = (c * lx[k] + (-sn) * ly[k]) + x;
        v[k][1] = (sn * lx[k] + c * ly[k]) + y;
    }
}

// np.argmax(vertices.dot(d)) -- first maximum (collision_models.py:82-92)
static __device__ __forceinline__ int furthest(const double v[4][2], double dx, double dy)
{
    int best = 0;
    double bv = v[0][0] * dx + v[0][1] * dy;
#pragma unroll
    for (int k = 1; k < 4; k++) {
        double t = v[k][0] * dx + v[k][1] * dy;
        if (t > bv) { bv = t; best = k; }
    }
    return best;
}
static __device__ __forceinline__ void support(const double v1[4][2], const double v2[4][2], double dx, double dy,
                                               double &ox, double &oy)
{
    int i = furthest(v1, dx, dy), j = furthest(v2, -dx, -dy);
    // register-resident select instead of dynamic indexing
    double ax = v1[0][0], ay = v1[0][1], bx = v2[0][0], by = v2[0][1];
#pragma unroll
    for (int k = 1; k < 4; k++) {
        if (i == k) { ax = v1[k][0]; ay = v1[k][1]; }
        if (j == k) { bx = v2[k][0]; by = v2[k][1]; }
    }
    ox = ax - bx; oy = ay - by;
}
// tripleProduct(a, b, c) = b*(a.c) - a*(b.c)   (collision_models.py:52-64)
static __device__ __forceinline__ void triple(double ax, double ay, double bx, double by, double cx, double cy,
                                              double &ox, double &oy)
{
    double ac = ax * cx + ay * cy, bc = bx * cx + by * cy;
    ox = bx * ac - ax * bc;
    oy = by * ac - ay * bc;
}

// collision (GJK, collision_models.py:114-182); simplex kept in registers as s0, s1 (+ the new point a)
static __device__ bool gjk_collision(const double v1[4][2], const double v2[4][2])
{
    double p1x = (((v1[0][0] + v1[1][0]) + v1[2][0]) + v1[3][0]) / 4, p1y = (((v1[0][1] + v1[1][1]) + v1[2][1]) + v1[3][1]) / 4;
    double p2x = (((v2[0][0] + v2[1][0]) + v2[2][0]) + v2[3][0]) / 4, p2y = (((v2[0][1] + v2[1][1]) + v2[2][1]) + v2[3][1]) / 4;
    double dx = p1x - p2x, dy = p1y - p2y;
    if (dx == 0 && dy == 0) dx = 1.0;
    double ax, ay;
    support(v1, v2, dx, dy, ax, ay);
    double s0x = ax, s0y = ay, s1x = 0, s1y = 0;
    if (dx * ax + dy * ay <= 0) return false;
    dx = -ax; dy = -ay;
    int index = 0, iter_count = 0;
    while (iter_count < 1000) {
        support(v1, v2, dx, dy, ax, ay);
        index += 1;
        if (dx * ax + dy * ay <= 0) return false;
        double aox = -ax, aoy = -ay;
        if (index < 2) {
            s1x = ax; s1y = ay;                 // simplex[1] = a
            double abx = s0x - ax, aby = s0y - ay;
            triple(abx, aby, aox, aoy, abx, aby, dx, dy);
            if (sqrt(dx * dx + dy * dy) < 1e-10) { dx = aby; dy = -1 * abx; }
            continue;
        }
        // index == 2: simplex = {s0 (c), s1 (b), a}
        double abx = s1x - ax, aby = s1y - ay, acx = s0x - ax, acy = s0y - ay;
        double px, py;
        triple(abx, aby, acx, acy, acx, acy, px, py);          // acperp
        if (px * aox + py * aoy >= 0) {
            dx = px; dy = py;
        } else {
            triple(acx, acy, abx, aby, abx, aby, px, py);      // abperp
            if (px * aox + py * aoy < 0) return true;
            s0x = s1x; s0y = s1y;                               // simplex[0] = simplex[1]
            dx = px; dy = py;
        }
        s1x = ax; s1y = ay;                                     // simplex[1] = simplex[2]
        index -= 1;
        iter_count += 1;
    }
    return false;
}

// get_range (laser_models.py:250-280)
static __device__ __forceinline__ double get_range(double ox_, double oy_, double v3x, double v3y, double vax,
                                                   double vay, double vbx, double vby)
{
    double v1x = ox_ - vax, v1y = oy_ - vay;
    double v2x = vbx - vax, v2y = vby - vay;
    double denom = v2x * v3x + v2y * v3y;
    double distance = CUDART_INF;
    if (fabs(denom) > 0.0) {
        // d1 = n1 / denom >= 0 and 0 <= d2 = n2 / denom <= 1 are decided on signs and magnitudes (an IEEE quotient is
        // >= 0 iff the numerator is 0 or has the denominator's sign, and <= 1 iff |n2| <= |denom|): the division
        // is only issued for the edge the beam does hit
        double n1 = v2x * v1y - v2y * v1x;
        double n2 = v1x * v3x + v1y * v3y;
        bool pos = denom > 0.0;
        bool ok1 = (n1 == 0.0) || (pos ? n1 > 0.0 : n1 < 0.0);
        bool ok2 = (n2 == 0.0) || ((pos ? n2 > 0.0 : n2 < 0.0) && fabs(n2) <= fabs(denom));
        if (ok1 && ok2) distance = __ddiv_rn(n1, denom);
    } else {
        double bax = vax - ox_, bay = vay - oy_, cax = ox_ - vbx, cay = oy_ - vby;
        if (fabs(bax * cay - bay * cax) < 1e-8) {
            double da = sqrt(bax * bax + bay * bay);
            double db = sqrt(cax * cax + cay * cay);
            distance = da < db ? da : db;
        }
    }
    return distance;
}

// int(np.argmin(|scan_angles - target|)) (laser_models.py:310-313): scan_angles is an arithmetic
// progression, so the minimum is next to round((target - a0)/inc); the table decides ties exactly.
static __device__ __forceinline__ int nearest_beam(const F110Dev &d, double target)
{
    if (target != target) return 0;                 // np.argmin returns the first NaN
    const int B = d.n_beams;
    double a0 = __ldg(d.scan_angles);
    double inc = d.fov / (double)(B - 1);
    double g = (target - a0) / inc;
    int k = (g <= 0.) ? 0 : (g >= (double)(B - 1) ? B - 1 : (int)(g + 0.5));
    int lo = max(0, k - 2), hi = min(B - 1, k + 2);
    int best = lo;
    double bv = fabs(__ldg(d.scan_angles + lo) - target);
    for (int i = lo + 1; i <= hi; i++) {
        double t = fabs(__ldg(d.scan_angles + i) - target);
        if (t < bv) { bv = t; best = i; }
    }
    return best;
}

// Simulator.check_collision + RaceCar.ray_cast_agents for all agents (base_classes.py:204-225, 529-543).
// The per-agent part is scalar fp64 work (vertices, GJK, four atan2 per opponent): one THREAD per agent does it,
// 32 agents per warp, and leaves per opponent a record (pose, opponent corners, first/last blocked beam) in shared
// memory.  The beams need lanes, but far fewer than the reference visits: it casts every beam between the first and
// last blocked index -- the whole scan when the opponent is behind the car, where the two indices clamp to the two
// ends of the field of view.  A beam can only change if
//   (a) its ray points at the opponent's bounding circle: direction within asin(R / distance) of the bearing of the
//       centre, R = half diagonal; or
//   (b) it runs exactly parallel to an edge that lies on its line BEHIND the origin: get_range (:250-280) has no
//       forward test in its denom == 0 branch.  denom = |edge| sin(angle between beam and edge) is far from 0 (and
//       from its rounding error) unless the beam is within 2e-3 rad of an edge direction.
// Both tests cost a few fp32 instructions per beam (margins 2e-3 rad against their own rounding).  The warp walks
// its 32 records, visits only the index windows where (a) or (b) can hold, and queues the (record, beam) pairs that
// pass; the four fp64 edge intersections then run on full batches of 32 pairs.  Results are identical to casting
// every beam; 131072 agents: 1.50 ms (warp per agent, all beams) -> DESIGN.md.
#define AG_THREADS 128
// one (record, beam) pair: item = record slot in the warp << 16 | beam.  span[5] holds the scanning agent of a record.
// ATOMIC: records of the same agent are cast by different warps at the same time (k_step_fused_multi64) -- the minimum is
// taken on the bit pattern as a signed integer, which orders doubles like the comparison does whenever the new range is
// not negative (a range never is; a scan value with noise may be, and then stays, as in the reference)
template <typename SCAN_T, bool ATOMIC>
static __device__ __forceinline__ void agents_cast(const F110Dev &d, const double (*rec)[AG_THREADS], const int (*span)[AG_THREADS], int w0, int item)
{
    const int s = w0 + (item >> 16), i = item & 0xffff;
    const double qx = rec[0][s], qy = rec[1][s];
    double v3s, v3c;
    sincos((rec[2][s] + __ldg(d.scan_angles + i)) + PI_() / 2., &v3s, &v3c);
    SCAN_T *scan = (SCAN_T *)d.scans + (size_t)span[5][s] * d.n_beams + i;
    double r = (double)*scan;
    const double r0 = r;
    double ax = rec[9][s], ay = rec[10][s];        // corner 3, then edges 3-0, 0-1, 1-2, 2-3: min() does not care about the order
#pragma unroll
    for (int j = 0; j < 4; j++) {
        const double bx = rec[3 + 2 * j][s], by = rec[4 + 2 * j][s];
        double rr = get_range(qx, qy, v3c, v3s, ax, ay, bx, by);
        if (rr < r) r = rr;
        ax = bx; ay = by;
    }
    if (r < r0) {
        if (ATOMIC) atomicMin((long long *)scan, __double_as_longlong(r));
        else *scan = (SCAN_T)r;
    }
}

// Simulator.check_collision for agent a of env e (base_classes.py:529-543) on the pre-iTTC poses with the construct-time car
// size: pair (min, max) of every partner; collision_idx = last writer = largest partner that collides
static __device__ __forceinline__ void agent_collision(const F110Dev &d, int ia, int e, int a)
{
    const int NA = d.n_envs * d.n_agents, A = d.n_agents;
    const size_t base = (size_t)e * A;
    const double gw = d.params[F110_NP * A + 0], gl = d.params[F110_NP * A + 1];
    double va[4][2], vo[4][2];
    get_vertices(d.pose_pre[ia], d.pose_pre[(size_t)NA + ia], d.pose_pre[2 * (size_t)NA + ia], gl, gw, va);
    int last = -1;
    for (int o = 0; o < A; o++) {
        if (o == a) continue;
        get_vertices(d.pose_pre[base + o], d.pose_pre[NA + base + o], d.pose_pre[2 * (size_t)NA + base + o], gl, gw, vo);
        if (a < o ? gjk_collision(va, vo) : gjk_collision(vo, va)) last = o;
    }
    d.collision_idx[ia] = (double)last;
    d.collisions[ia] = (last >= 0 || d.in_collision[ia]) ? 1.0 : 0.0;
}

// RaceCar.ray_cast_agents, the part per (agent, opponent): the record of opponent o (agent base + o) as seen by agent ia at
// (px, py, yaw) -- own pose AFTER the iTTC zeroing, the opponent's from before it, drawn with THIS agent's length / width --
// into column t: rec = px, py, yaw, corner x / y * 4; span = first / last blocked beam (get_blocked_view_indices,
// laser_models.py:283-315), float bits of bearing offset, cone, edge offset; scanning agent
// get_blocked_view_indices (laser_models.py:283-315), one corner: the beam nearest to its bearing
static __device__ __forceinline__ int corner_beam(const F110Dev &d, double cx, double cy, double px, double py, double ego_ang)
{
    double vx = cx - px, vy = cy - py;
    double nrm = sqrt(vx * vx + vy * vy);
    double angle = ego_ang - atan2(vy / nrm, vx / nrm);
    if (angle > PI_()) angle = angle - 2 * PI_();
    else if (angle < -PI_()) angle = angle + 2 * PI_();
    return nearest_beam(d, -angle);
}

// the rest of a record once the corners v and the first / last blocked beam are known: pose, float bits of bearing offset,
// cone and edge offset for the beam tests (a) and (b) above, scanning agent
static __device__ __forceinline__ void pair_record_tail(const F110Dev &d, double (*rec)[AG_THREADS], int (*span)[AG_THREADS], int t, int ia,
                                                        const double v[4][2], int mn, int mx, double yaw_o, double px, double py, double yaw,
                                                        double len, double wid)
{
    rec[0][t] = px; rec[1][t] = py; rec[2][t] = yaw;
    // (a): bearing of the centre relative to scan angle 0 and the half angle of the bounding circle; (b): edge directions
    const double cx = 0.25 * (((v[0][0] + v[1][0]) + v[2][0]) + v[3][0]) - px;
    const double cy = 0.25 * (((v[0][1] + v[1][1]) + v[2][1]) + v[3][1]) - py;
    const float dist = sqrtf((float)(cx * cx + cy * cy));
    const float R = 0.5f * sqrtf((float)(len * len + wid * wid)) * 1.0001f;
    double off = yaw - (double)atan2f((float)cy, (float)cx);
    off -= 2 * PI_() * rint(off / (2 * PI_()));
    double par = yaw - yaw_o;
    par -= PI_() / 2 * rint(par / (PI_() / 2));
    float cone = (dist > 1.001f * R) ? asinf(R / dist) + 2e-3f : 4.0f;   // inside the circle (or NaN): every beam
    if (!(cone == cone) || !(off == off)) { cone = 4.0f; off = 0.0; }
    // Both tests take a beam's direction to be yaw + scan_angle.  The reference rounds that sum (pose[2] + scan_angles[i],
    // laser_models.py:343), and the yaw is only wrapped by 2 pi once per step (base_classes.py:398-402): a car whose model
    // has blown up (yaw rate 1e17 with a harsh parameter set and dt = 0.02 in tools/step_fuzz.py) has a yaw of 1e15, where
    // doubles are 0.125 rad apart, six neighbouring beams point the same way and hit what the exact directions miss.
    // Far outside any yaw a wrapped heading can have, no beam is skipped: the reference's full window is cast.
    if (!(fabs(yaw) < 1e6)) { cone = 4.0f; off = 0.0; }
    if (!(fabs(yaw) < 1e6) || !(fabs(yaw_o) < 1e6)) par = CUDART_NAN;
    span[0][t] = mn; span[1][t] = mx;
    span[2][t] = __float_as_int((float)off);
    span[3][t] = __float_as_int(cone);
    span[4][t] = __float_as_int((float)par);      // NaN: every beam counts as parallel
    span[5][t] = ia;
}

// RaceCar.ray_cast_agents, the part per (agent, opponent): the record of opponent o (agent base + o) as seen by agent ia at
// (px, py, yaw) -- own pose AFTER the iTTC zeroing, the opponent's from before it, drawn with THIS agent's length / width --
// into column t: rec = px, py, yaw, corner x / y * 4; span = first / last blocked beam, the beam tests' constants, agent
static __device__ __forceinline__ void agent_pair_record(const F110Dev &d, double (*rec)[AG_THREADS], int (*span)[AG_THREADS], int t, int ia,
                                                         size_t base, int o, double px, double py, double yaw, double ego_ang, double len, double wid)
{
    const int NA = d.n_envs * d.n_agents;
    double v[4][2];
    const double yaw_o = d.pose_pre[2 * (size_t)NA + base + o];
    get_vertices(d.pose_pre[base + o], d.pose_pre[NA + base + o], yaw_o, len, wid, v);
    int mn = 0x7fffffff, mx = -1;
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const int b = corner_beam(d, v[k][0], v[k][1], px, py, ego_ang);
        mn = min(mn, b); mx = max(mx, b);
        rec[3 + 2 * k][t] = v[k][0];
        rec[4 + 2 * k][t] = v[k][1];
    }
    pair_record_tail(d, rec, span, t, ia, v, mn, mx, yaw_o, px, py, yaw, len, wid);
}

// the same by four neighbouring lanes (lane & 3 = corner; `group` = the ballot of the lanes that take part, whole groups of
// four): the four corner bearings -- square root, two divisions, atan2 and the nearest-beam search each, 3 500 cycles -- side
// by side instead of one after the other (19 000 of the 38 000 cycles of a two-agent single-env tick)
static __device__ __forceinline__ void agent_pair_record4(const F110Dev &d, double (*rec)[AG_THREADS], int (*span)[AG_THREADS], int t, int ia,
                                                          size_t base, int o, double px, double py, double yaw, double ego_ang, double len, double wid,
                                                          unsigned group)
{
    const int NA = d.n_envs * d.n_agents, k = threadIdx.x & 3;
    double v[4][2];
    const double yaw_o = d.pose_pre[2 * (size_t)NA + base + o];
    get_vertices(d.pose_pre[base + o], d.pose_pre[NA + base + o], yaw_o, len, wid, v);
    const double cx = k == 0 ? v[0][0] : k == 1 ? v[1][0] : k == 2 ? v[2][0] : v[3][0];
    const double cy = k == 0 ? v[0][1] : k == 1 ? v[1][1] : k == 2 ? v[2][1] : v[3][1];
    const int b = corner_beam(d, cx, cy, px, py, ego_ang);
    int mn = b, mx = b;
    mn = min(mn, __shfl_xor_sync(group, mn, 1)); mx = max(mx, __shfl_xor_sync(group, mx, 1));
    mn = min(mn, __shfl_xor_sync(group, mn, 2)); mx = max(mx, __shfl_xor_sync(group, mx, 2));
    rec[3 + 2 * k][t] = cx;
    rec[4 + 2 * k][t] = cy;
    if (k == 0) pair_record_tail(d, rec, span, t, ia, v, mn, mx, yaw_o, px, py, yaw, len, wid);
}

// ray_cast (laser_models.py:318-346) over the candidate beams of record s (item code `code`: s = w0 + code): the calling warp
// queues them in queue_w (`fill` entries so far) and casts every full batch of 32.  The reference casts every beam between the
// first and last blocked index; only beams that pass test (a) or (b) above can change.  n_units > 1: the walk is shared by
// n_units warps, each taking the 32-beam steps whose running number `unit` is its own (my_unit) modulo n_units.
template <typename SCAN_T, bool ATOMIC>
static __device__ __forceinline__ void agents_walk_record(const F110Dev &d, const double (*rec)[AG_THREADS], const int (*span)[AG_THREADS],
                                                          int *queue_w, int &fill, int s, int code, int w0, int lane, float a032, float inc32,
                                                          float inv_inc32, int &unit, int n_units, int my_unit, int cast_at = 32)
{   // cast_at: queue fill at which a batch of 32 is cast on the spot (the queue holds cast_at + 32 entries)
    const int first = span[0][s], lastb = span[1][s];
    if (first > lastb) return;
    const float off = __int_as_float(span[2][s]), cone = __int_as_float(span[3][s]), par = __int_as_float(span[4][s]);
    // windows 2 (cone + 2 beams) wide, pi apart: disjoint under these limits; otherwise one pass over [first, last]
    const bool windows = cone < 1.2f && inc32 < 0.05f && inc32 > 0.0015f && d.fov <= 6.2831854;
    const float half = cone * inv_inc32 + 2.0f;
    // passes: -1, 0, 1 = beams within `cone` of the bearing (- 2 pi, + 0, + 2 pi); 2 = the (at most five) beam
    // directions parallel to an edge, one lane per candidate and neighbour (2e-3 rad < 1.5 beams)
    for (int n = windows ? -1 : 0; n <= (windows ? 2 : 0); n++) {
        int lo = first, hi = lastb;
        if (windows && n < 2) {
            const float c = ((float)n * 6.28318531f - off - a032) * inv_inc32;
            lo = max(first, (int)ceilf(c - half));
            hi = min(lastb, (int)floorf(c + half));
        }
        if (n == 2) { lo = 0; hi = 14; }
        for (int i0 = lo; i0 <= hi; i0 += 32) {
            if (n_units > 1 && (unit++ % n_units) != my_unit) continue;
            int i = i0 + lane;
            bool inside = i <= hi;
            if (n == 2) {
                const int m = lane / 3;
                i = __float2int_rn(((float)(m - 2) * 1.57079633f - par - a032) * inv_inc32) + (lane - 3 * m - 1);
                inside = lane < 15 && i >= first && i <= lastb;
            }
            float rel = fmaf((float)i, inc32, a032) + off;
            rel -= 6.28318531f * rintf(rel * 0.159154943f);
            float rp = fmaf((float)i, inc32, a032) + par;
            rp -= 1.57079633f * rintf(rp * 0.636619772f);
            const bool towards = !(fabsf(rel) > cone);
            const bool away = !(3.14159265f - fabsf(rel) > cone) && !(fabsf(rp) > 2e-3f);
            const bool take = inside && (n == 2 ? (away && !towards) : (windows ? towards : (towards || away)));
            const unsigned m = __ballot_sync(FULL, take);
            F110_CHECK(d.violations, !take || (fill + __popc(m & ((1u << lane) - 1u)) < cast_at + 32 && i >= 0 && i < d.n_beams), F110_V_QUEUE);
            if (take) queue_w[fill + __popc(m & ((1u << lane) - 1u))] = (code << 16) | i;
            fill += __popc(m);
            __syncwarp();
            if (fill >= cast_at) {
                agents_cast<SCAN_T, ATOMIC>(d, rec, span, w0, queue_w[lane]);
                __syncwarp();
                for (int q0 = 0; q0 + 32 < fill; q0 += 32) {       // move the rest down, 32 entries at a time (cast_at = 32: once)
                    const int q = q0 + lane;
                    const int rest = q + 32 < fill ? queue_w[32 + q] : 0;
                    __syncwarp();
                    if (q + 32 < fill) queue_w[q] = rest;
                    __syncwarp();
                }
                fill -= 32;
                __syncwarp();
            }
        }
    }
}

// Simulator.check_collision + RaceCar.ray_cast_agents for all agents (base_classes.py:204-225, 529-543), thread t of a CTA
// of AG_THREADS threads = agent ia: the scalar fp64 work per agent and per opponent (vertices, GJK, four atan2) by its
// thread, 32 agents per warp, which leaves a record per opponent in shared memory; the warp then walks its 32 records.
template <typename SCAN_T>
static __device__ __forceinline__ void agents_body(const F110Dev &d, double (*rec)[AG_THREADS], int (*span)[AG_THREADS], int (*queue)[64],
                                                   const int t, const int ia, const bool in_range)
{
    const int NA = d.n_envs * d.n_agents, A = d.n_agents;
    const int lane = t & 31, w0 = t & ~31, wq = t >> 5;
    int e = 0, a = 0;
    bool live = in_range;
    if (live) {
        e = ia / A; a = ia - e * A;
        if (d.step_mask && !d.step_mask[e]) live = false;
    }
    const size_t base = (size_t)e * A;
    double px = 0, py = 0, yaw = 0, ego_ang = 0, len = 0, wid = 0;
    if (live) {
        agent_collision(d, ia, e, a);
        px = d.state[ia]; py = d.state[(size_t)NA + ia]; yaw = d.state[4 * (size_t)NA + ia];
        len = d.params[F110_NP * a + F110_P_LENGTH]; wid = d.params[F110_NP * a + F110_P_WIDTH];
        double esn, ecs;
        sincos(yaw, &esn, &ecs);
        ego_ang = atan2(esn, ecs);
    }
    const float a032 = (float)__ldg(d.scan_angles), inc32 = (float)(d.fov / (double)(d.n_beams - 1)), inv_inc32 = 1.0f / inc32;
    for (int o = 0; o < A; o++) {
        if (live && o != a) agent_pair_record(d, rec, span, t, ia, base, o, px, py, yaw, ego_ang, len, wid);
        else { span[0][t] = 1; span[1][t] = 0; }
        __syncwarp();
        int fill = 0, unit = 0;
        for (int k = 0; k < 32; k++)
            agents_walk_record<SCAN_T, false>(d, rec, span, queue[wq], fill, w0 + k, k, w0, lane, a032, inc32, inv_inc32, unit, 1, 0);
        if (lane < fill) agents_cast<SCAN_T, false>(d, rec, span, w0, queue[wq][lane]);
        __syncwarp();
    }
}

template <typename SCAN_T>
// 8 CTAs of 128 threads per SM (64 registers, 240 bytes of spill): measured on config 4 (131 072 agents) against the compiler's own
// choice (96 registers, 5 CTAs) and 6 / 10 / 12 CTAs: step 1.887 ms against 1.937 / 1.951 / 1.937 / 1.991 (profiles/r2g_agents_occupancy_ab.json)
__global__ void __launch_bounds__(AG_THREADS, 8) k_agents(const F110Dev d)
{
    __shared__ double rec[11][AG_THREADS];
    __shared__ int span[6][AG_THREADS];
    __shared__ int queue[AG_THREADS / 32][64];
    const int ia = blockIdx.x * AG_THREADS + threadIdx.x;
    agents_body<SCAN_T>(d, rec, span, queue, threadIdx.x, ia, ia < d.n_envs * d.n_agents);
}

void f110_launch_agents(const F110Dev &d, cudaStream_t st)
{
    long long NA = (long long)d.n_envs * d.n_agents;
    int grid = (int)((NA + AG_THREADS - 1) / AG_THREADS);
    if (d.scan_is_f32) k_agents<float><<<grid, AG_THREADS, 0, st>>>(d);
    else k_agents<double><<<grid, AG_THREADS, 0, st>>>(d);
}

// ------------------------------------------------------------------------------------------------
// F110Env.step tail: time, _check_done (f110_env.py:202-244, 292-299)
// ------------------------------------------------------------------------------------------------
// the lap counter of one agent (f110_env.py:202-244): (dx, dy) = position - start, R = the start frame's rotation
static __device__ __forceinline__ void lap_update(double r0, double r1, double r2, double r3, double dx, double dy, int &tg, bool &near)
{
    double d0 = r0 * dx + r1 * dy;
    double ty = r2 * dx + r3 * dy;
    if (ty > 2) ty -= 2;
    else if (ty < -2) ty = -2 - ty;
    else ty = 0;
    double dist2 = d0 * d0 + ty * ty;
    bool closes = dist2 <= 0.1;
    if (closes && !near) { near = true; tg += 1; }
    else if (!closes && near) { near = false; tg += 1; }
}

static __device__ __forceinline__ void finalize_env(const F110Dev &d, int e)
{
    if (d.step_mask && !d.step_mask[e]) return;
    const int A = d.n_agents, NA = d.n_envs * A;
    // loads first, stores last (per env and per agent): with the stores in between every load would wait for the one before it
    const double t = d.cur_time[e] + d.timestep;
    const long long ns = d.noise_step[e];
    const double *R = d.start_rot + 4 * (size_t)e;
    const double r0 = R[0], r1 = R[1], r2 = R[2], r3 = R[3];
    const long long episodes = d.auto_reset ? d.episode_count[e] : 0;
    // collisions of the ego: with several agents k_agents wrote them (GJK | iTTC); alone it is the iTTC flag (base_classes.py:581-582)
    bool ego_collides = A > 1 && d.collisions[(size_t)e * A + d.ego_idx] != 0.0;
    bool all_done = true;
    for (int a = 0; a < A; a++) {
        const size_t ia = (size_t)e * A + a;
        const bool hit = A == 1 && d.in_collision[ia];
        const double dx = d.state[ia] - d.start_x[ia];
        const double dy = d.state[(size_t)NA + ia] - d.start_y[ia];
        int tg = d.toggles[ia];
        bool near = d.near_start[ia] != 0;
        if (A == 1) {
            d.collisions[ia] = hit ? 1.0 : 0.0;
            d.collision_idx[ia] = -1.0;
            ego_collides = hit;
        }
        lap_update(r0, r1, r2, r3, dx, dy, tg, near);
        d.near_start[ia] = near ? 1 : 0;
        d.toggles[ia] = tg;
        d.checkpoint_done[ia] = tg >= 4 ? 1 : 0;
        d.lap_counts[ia] = (double)(tg / 2);
        if (tg < 4) d.lap_times[ia] = t;
        if (tg < 4) all_done = false;
    }
    d.cur_time[e] = t;
    d.noise_step[e] = ns + 1;
    const bool done = ego_collides || all_done;
    d.done[e] = done ? 1 : 0;
    if (d.auto_reset) {
        d.reset_pending[e] = done ? 1 : 0;
        if (done) d.episode_count[e] = episodes + 1;
    }
}

__global__ void __launch_bounds__(128) k_finalize(const F110Dev d)
{
    int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e < d.n_envs) finalize_env(d, e);
}

__global__ void k_check_selftest(unsigned long long *viol)
{
    F110_CHECK(viol, threadIdx.x != 0, F110_V_SELFTEST);      // one deliberate violation: proves the checks are compiled in and read back
}
void f110_launch_check_selftest(unsigned long long *viol, cudaStream_t st) { k_check_selftest<<<1, 32, 0, st>>>(viol); }

void f110_launch_finalize(const F110Dev &d, cudaStream_t st)
{
    k_finalize<<<(d.n_envs + 127) / 128, 128, 0, st>>>(d);
}

// ------------------------------------------------------------------------------------------------
// The whole tick of a single-agent env in ONE kernel (north_star: "dynamics, integration and TTC fused into one per-env
// kernel"), for a handful of envs: a CTA per env.  Thread 0 integrates the car (dynamics_agent) and prepares the scan, the
// eight warps cast it (rounds claimed from a shared counter, as in the split kernel), thread 0 applies the iTTC outcome
// and the lap bookkeeping (finalize_env's arithmetic for one agent: lap_update), and -- when the caller's host buffers are
// mapped -- the CTA writes the env's observation straight into them.  Same device functions as k_dynamics / k_raycast64 /
// k_finalize, so the results are bit-identical to the three-kernel step (tests/test_gpu_step.py); what goes away is three
// launches and their gaps, which dominate a single-env step (one env: 46 us of kernels in an 80 us call before).
// ------------------------------------------------------------------------------------------------
struct F110HostOut {
    double *obs;            // [10][N] or null
    double *scans;          // [N][B] or null
    uint8_t *done;          // [N] or null
    double cmd[16];         // one env of up to 8 agents: their (steer, velocity) commands, read by the host (has_cmd) -- saves the kernel a PCIe read
    int has_cmd;
};

// what the step's tail needs from global memory, fetched by a thread of another warp while the scan is cast, so that thread 0's
// tail is shared-memory reads, arithmetic and stores (0.4 us; tools/phase_probe.py -- the single-env kernel is 8 us of
// thread 0's RK4, 1 us of scan preparation and 7.5 us until the slowest round of the scan is in)
struct FusedTail {
    double x[7];            // the car's state after the dynamics
    double cur_time, r0, r1, r2, r3, start_x, start_y, lap_time;
    long long noise_step, episodes;
    int toggles, near;
};

#ifdef F110_PHASES      // diagnostic build (build.py --phases, tools/phase_probe.py): clock64 stamps of env 0's step by thread 0
// [0..7] %globaltimer (ns; read at the first and the last stamp only, for the clock rate), [8..15] clock64 at stamp i.
// (A stamp placed right after __syncthreads is taken BEFORE the barrier resolves -- BAR.SYNC.DEFER_BLOCKING lets the warp run on
// to the first instruction that needs the barrier -- hence stamp 4 sits inside thread 0's branch behind it.)
__device__ unsigned long long f110_phase[16];
extern "C" int f110_debug_phases(unsigned long long *out16) { return (int)cudaMemcpyFromSymbol(out16, f110_phase, sizeof(f110_phase)); }
#define PHASE(i) do { if (blockIdx.x == 0 && threadIdx.x == 0) { unsigned long long t_ = 0, c_; if ((i) == 0 || (i) == 6) asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_) :: "memory"); asm volatile("mov.u64 %0, %%clock64;" : "=l"(c_) :: "memory"); f110_phase[i] = t_; f110_phase[8 + i] = c_; } } while (0)
#else
#define PHASE(i) do { } while (0)
#endif

// NT threads per CTA: 256, or 1024 when there are no more envs than SMs -- a lone scan is 34 rounds of dependent gathers, and with
// 32 warps every warp casts about one round instead of four (the 1024-thread build is capped at 64 registers: thread 0's RK4 spills,
// which costs less than it saves -- see k_dynamics).
template <bool ROT, bool POW2, bool COUNT, int NT>
__global__ void __launch_bounds__(NT) k_step_fused64(const F110Dev d, const F110HostOut ho)
{
    __shared__ Scan64 row;
    __shared__ FusedTail tail;
    __shared__ int next_round, hit_any, active;
    const int e = (int)blockIdx.x, lane = threadIdx.x & 31;      // one agent per env: ia == e
    const int N = d.n_envs;
    PHASE(0);
    if (threadIdx.x == 0) {
        active = !(d.step_mask && !d.step_mask[e]);
        if (active) {
            dynamics_agent(d, e, ho.has_cmd ? ho.cmd : nullptr);
            PHASE(1);
            prepare_env_scan64<ROT, POW2>(d, e, &row, true);
        }
        next_round = 0;
        hit_any = 0;
    }
    PHASE(2);
    __syncthreads();
    if (!active) return;
    if (threadIdx.x == 32) {      // after the barrier: what dynamics_agent wrote (also an auto-reset's new start frame) is visible
#pragma unroll
        for (int k = 0; k < 7; k++) tail.x[k] = d.state[(size_t)k * N + e];
        const double *R = d.start_rot + 4 * (size_t)e;
        tail.r0 = R[0]; tail.r1 = R[1]; tail.r2 = R[2]; tail.r3 = R[3];
        tail.cur_time = d.cur_time[e];
        tail.noise_step = d.noise_step[e];
        tail.start_x = d.start_x[e];
        tail.start_y = d.start_y[e];
        tail.lap_time = d.lap_times[e];
        tail.toggles = d.toggles[e];
        tail.near = d.near_start[e];
        tail.episodes = d.auto_reset ? d.episode_count[e] : 0;
    }
    bool hit = false;
    unsigned nlook = 0;
    scan_rounds_claimed<ROT, POW2, true, COUNT>(d, &row, hit, nlook, &next_round, threadIdx.x == 0);
    PHASE(3);
    if (__any_sync(FULL, hit) && lane == 0) atomicOr(&hit_any, 1);
    __syncthreads();
    if (threadIdx.x == 0) {
        PHASE(4);
        // RaceCar.check_ttc (base_classes.py:241-250), then finalize_env for the env's only agent, from `tail` instead of global memory
        const bool h = hit_any != 0;
        double x[7];
#pragma unroll
        for (int k = 0; k < 7; k++) x[k] = tail.x[k];
        if (h) {
            x[3] = 0.; x[4] = 0.; x[5] = 0.; x[6] = 0.;
            d.state[3 * (size_t)N + e] = 0.; d.state[4 * (size_t)N + e] = 0.; d.state[5 * (size_t)N + e] = 0.; d.state[6 * (size_t)N + e] = 0.;
        }
        d.in_collision[e] = h ? 1 : 0;
        const double t = tail.cur_time + d.timestep;
        int tg = tail.toggles;
        bool near = tail.near != 0;
        lap_update(tail.r0, tail.r1, tail.r2, tail.r3, x[0] - tail.start_x, x[1] - tail.start_y, tg, near);
        const double coll = h ? 1.0 : 0.0, laps = (double)(tg / 2), lap_time = tg < 4 ? t : tail.lap_time;
        const bool done = h || tg >= 4;
        d.collisions[e] = coll;
        d.collision_idx[e] = -1.0;
        d.near_start[e] = near ? 1 : 0;
        d.toggles[e] = tg;
        d.checkpoint_done[e] = tg >= 4 ? 1 : 0;
        d.lap_counts[e] = laps;
        if (tg < 4) d.lap_times[e] = t;
        d.cur_time[e] = t;
        d.noise_step[e] = tail.noise_step + 1;
        d.done[e] = done ? 1 : 0;
        if (d.auto_reset) {
            d.reset_pending[e] = done ? 1 : 0;
            if (done) d.episode_count[e] = tail.episodes + 1;
        }
        if (ho.done) ho.done[e] = done ? 1 : 0;
        if (ho.obs) {
#pragma unroll
            for (int k = 0; k < 7; k++) ho.obs[(size_t)k * N + e] = x[k];
            ho.obs[7 * (size_t)N + e] = coll;
            ho.obs[8 * (size_t)N + e] = lap_time;
            ho.obs[9 * (size_t)N + e] = laps;
        }
        PHASE(5);
    }
    if (ho.scans) {       // the scan this CTA just wrote (visible CTA-wide after the barrier) -> the caller's host buffer
        const double *src = (const double *)d.scans + (size_t)e * d.n_beams;
        double *dst = ho.scans + (size_t)e * d.n_beams;
        for (int i = threadIdx.x; i < d.n_beams; i += NT) dst[i] = src[i];
    }
    PHASE(6);
    if (COUNT) {
#pragma unroll
        for (int o = 16; o; o >>= 1) nlook += __shfl_xor_sync(FULL, nlook, o);
        if (lane == 0) atomicAdd(d.lookups, (unsigned long long)nlook);
    }
}

// The same for a handful of envs of 2 .. F110_FUSED_MULTI_MAX_AGENTS agents (the reference's head-to-head races through the drop-in
// F110Env: five launches and their gaps were 80 us per step of a two-agent env): a CTA of 32 warps per env.  Threads 0 .. A-1
// integrate the agents and prepare their scans, the warps cast the A scans (rounds claimed per scan, a warp starts on scan
// warp % A and moves on to the others), threads 0 .. A-1 apply the iTTC outcomes, k_agents' work is spread over the CTA (GJK
// on the pre-iTTC poses per pair, opponent records by four threads each, their candidate beams cast by all warps: see below),
// thread 0 does the lap bookkeeping, and the CTA writes the observation to the caller's mapped host buffers.  Order and
// device functions are those of k_dynamics -> k_raycast64 -> k_agents -> k_finalize (46 us per call for two agents).
template <bool ROT, bool POW2>
__global__ void __launch_bounds__(1024) k_step_fused_multi64(const F110Dev d, const F110HostOut ho)
{
    __shared__ Scan64 rows[F110_FUSED_MULTI_MAX_AGENTS];
    __shared__ int next_round[F110_FUSED_MULTI_MAX_AGENTS], hit_any[F110_FUSED_MULTI_MAX_AGENTS], active;
    __shared__ double rec[11][AG_THREADS];      // a column per (agent, opponent) pair: A (A - 1) <= 56
    __shared__ int span[6][AG_THREADS];
    __shared__ int queue[32][128], fills[32];
    __shared__ unsigned char pair_hit[F110_FUSED_MULTI_MAX_AGENTS][F110_FUSED_MULTI_MAX_AGENTS];
    const int e = (int)blockIdx.x, A = d.n_agents, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const size_t NA = (size_t)d.n_envs * A;
    PHASE(0);
    if (tid == 0) active = !(d.step_mask && !d.step_mask[e]);
    if (tid < A) {
        const int ia = e * A + tid;
        if (!(d.step_mask && !d.step_mask[e])) {
            dynamics_agent(d, ia, ho.has_cmd ? ho.cmd + 2 * tid : nullptr);
            prepare_env_scan64<ROT, POW2>(d, ia, &rows[tid], true);
        }
        next_round[tid] = 0;
        hit_any[tid] = 0;
    }
    __syncthreads();
    if (!active) return;
    unsigned nlook = 0;
    for (int k = 0; k < A; k++) {
        int s2 = warp % A + k;
        if (s2 >= A) s2 -= A;
        bool h = false;
        scan_rounds_claimed<ROT, POW2, true, false>(d, rows + s2, h, nlook, &next_round[s2], false);
        if (__any_sync(FULL, h) && lane == 0) atomicOr(&hit_any[s2], 1);
    }
    PHASE(3);
    __syncthreads();
    if (tid < A) {       // RaceCar.check_ttc (base_classes.py:241-250)
        PHASE(4);
        const size_t ia = (size_t)e * A + tid;
        const bool h = hit_any[tid] != 0;
        if (h) { d.state[3 * NA + ia] = 0.; d.state[4 * NA + ia] = 0.; d.state[5 * NA + ia] = 0.; d.state[6 * NA + ia] = 0.; }
        d.in_collision[ia] = h ? 1 : 0;
    }
    __syncthreads();
    // k_agents' work spread over the CTA.  Its scalar part is a chain of fp64 divisions, atan2 and GJK per agent and opponent
    // (33 us for two agents in the lanes of one warp): here four threads per (agent, opponent) pair write that pair's record,
    // threads 256 .. run GJK on the A (A - 1) / 2 pairs beside them, then a warp per record queues its candidate beams and all 32
    // warps cast them, a batch of 32 each -- records of one agent at the same time, hence the atomic minimum.
    {
        const int P = A * (A - 1);
        const unsigned group = __ballot_sync(FULL, tid < 4 * P);        // four lanes per pair: threads 0 .. 4 P - 1 (at most 224)
        if (tid < 4 * P) {
            const int p = tid >> 2, a = p / (A - 1);
            int o = p - a * (A - 1);
            if (o >= a) o++;
            const size_t ia = (size_t)e * A + a;
            const double px = d.state[ia], py = d.state[NA + ia], yaw = d.state[4 * NA + ia];
            const double len = d.params[F110_NP * a + F110_P_LENGTH], wid = d.params[F110_NP * a + F110_P_WIDTH];
            double esn, ecs;
            sincos(yaw, &esn, &ecs);
            agent_pair_record4(d, rec, span, p, (int)ia, (size_t)e * A, o, px, py, yaw, atan2(esn, ecs), len, wid, group);
        } else if (tid >= 256 && tid < 256 + P / 2) {
            // Simulator.check_collision (base_classes.py:529-543): a thread per unordered pair i < j, as collision_multiple
            // visits them (collision_models.py:185-212), instead of agent_collision's loop over every partner per agent
            int q = tid - 256, i = 0;
            while (q >= A - 1 - i) { q -= A - 1 - i; i++; }
            const int j = i + 1 + q;
            const size_t ii = (size_t)e * A + i, jj = (size_t)e * A + j;
            const double gw = d.params[F110_NP * A + 0], gl = d.params[F110_NP * A + 1];
            double vi[4][2], vj[4][2];
            get_vertices(d.pose_pre[ii], d.pose_pre[NA + ii], d.pose_pre[2 * NA + ii], gl, gw, vi);
            get_vertices(d.pose_pre[jj], d.pose_pre[NA + jj], d.pose_pre[2 * NA + jj], gl, gw, vj);
            const bool c = gjk_collision(vi, vj);
            pair_hit[i][j] = c; pair_hit[j][i] = c;
        }
        __syncthreads();
        PHASE(1);
        if (tid < A) {      // collision_idx = last writer = largest partner that collides
            int last = -1;
            for (int o = 0; o < A; o++)
                if (o != tid && pair_hit[tid][o]) last = o;
            const size_t ia = (size_t)e * A + tid;
            d.collision_idx[ia] = (double)last;
            d.collisions[ia] = (last >= 0 || d.in_collision[ia]) ? 1.0 : 0.0;
        }
        const float a032 = (float)__ldg(d.scan_angles), inc32 = (float)(d.fov / (double)(d.n_beams - 1)), inv_inc32 = 1.0f / inc32;
        int fill = 0, unit = 0;
        for (int p = warp; p < P; p += 32)       // a warp per record (every warp walking every record cost 2 700 cycles per record)
            agents_walk_record<double, true>(d, rec, span, queue[warp], fill, p, p, 0, lane, a032, inc32, inv_inc32, unit, 1, 0, 96);
        PHASE(2);
        // the candidates sit in the queues of the warps that walked a record: level them -- batch b of the CTA's candidates, in
        // queue order, is cast by warp b mod 32
        if (lane == 0) fills[warp] = fill;
        __syncthreads();
        int total = 0;
        for (int w = 0; w < 32; w++) total += fills[w];
        for (int g = 32 * warp + lane; g < ((total + 31) & ~31); g += 1024) {
            int w = 0, before = 0;
            while (w < 31 && before + fills[w] <= g) { before += fills[w]; w++; }
            if (g < total) agents_cast<double, true>(d, rec, span, 0, queue[w][g - before]);
        }
    }
    PHASE(5);
    __syncthreads();
    if (tid == 0) {
        PHASE(7);
        finalize_env(d, e);
        if (ho.done) ho.done[e] = d.done[e];
    }
    if (ho.scans) {
        const double *src = (const double *)d.scans + (size_t)e * A * d.n_beams;
        double *dst = ho.scans + (size_t)e * A * d.n_beams;
        for (int i = tid; i < A * d.n_beams; i += 1024) dst[i] = src[i];
    }
    if (ho.obs) {
        __syncthreads();      // collisions, lap times and counts of finalize_env
        for (int i = tid; i < 10 * A; i += 1024) {
            const size_t at = (size_t)(i / A) * NA + (size_t)e * A + (size_t)(i % A);
            ho.obs[at] = d.state[at];          // obs_block = state | collisions | lap_times | lap_counts
        }
    }
    PHASE(6);
}

void f110_launch_step_fused64(const F110Dev &d, double *obs_host, double *scans_host, uint8_t *done_host, const double *cmd_host, int variant,
                              cudaStream_t st)
{
    F110HostOut ho;
    memset(&ho, 0, sizeof(ho));
    ho.obs = obs_host; ho.scans = scans_host; ho.done = done_host;
    if (cmd_host && d.n_envs == 1 && d.n_agents <= 8) {
        for (int k = 0; k < 2 * d.n_agents; k++) ho.cmd[k] = cmd_host[k];
        ho.has_cmd = 1;
    }
    if (d.n_agents > 1) {
        if (d.has_rot) k_step_fused_multi64<true, false><<<d.n_envs, 1024, 0, st>>>(d, ho);
        else if (d.res_pow2) k_step_fused_multi64<false, true><<<d.n_envs, 1024, 0, st>>>(d, ho);
        else k_step_fused_multi64<false, false><<<d.n_envs, 1024, 0, st>>>(d, ho);
        return;
    }
    const bool cnt = d.count_lookups != 0;
    const bool wide = d.n_envs <= d.n_sm && !(variant & 1);        // variant bit 0: always 256 threads per env
#define FUSED(R, P)                                                                   \
    do {                                                                              \
        if (cnt) k_step_fused64<R, P, true, 256><<<d.n_envs, 256, 0, st>>>(d, ho);    \
        else if (wide) k_step_fused64<R, P, false, 1024><<<d.n_envs, 1024, 0, st>>>(d, ho); \
        else k_step_fused64<R, P, false, 256><<<d.n_envs, 256, 0, st>>>(d, ho);       \
    } while (0)
    if (d.has_rot) FUSED(true, false);
    else if (d.res_pow2) FUSED(false, true);
    else FUSED(false, false);
#undef FUSED
}

// ------------------------------------------------------------------------------------------------
// F110Env.reset / RaceCar.reset (f110_env.py:318-332, base_classes.py:181-202); the zero-action
// step that follows (:335-336) is issued by the host wrapper.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128) k_reset(const F110Dev d, const double *__restrict__ poses,
                                               const double *__restrict__ rot, const uint8_t *__restrict__ mask)
{
    const int A = d.n_agents, NA = d.n_envs * A;
    int ia = blockIdx.x * blockDim.x + threadIdx.x;
    if (ia >= NA) return;
    int e = ia / A, a = ia - e * A;
    if (mask && !mask[e]) return;
    const double *p = poses + 3 * (size_t)ia;
    d.state[ia] = p[0];
    d.state[(size_t)NA + ia] = p[1];
    d.state[2 * (size_t)NA + ia] = 0.;
    d.state[3 * (size_t)NA + ia] = 0.;
    d.state[4 * (size_t)NA + ia] = p[2];
    d.state[5 * (size_t)NA + ia] = 0.;
    d.state[6 * (size_t)NA + ia] = 0.;
    d.steer_len[ia] = 0;
    d.in_collision[ia] = 0;
    d.collisions[ia] = 0.;
    d.near_start[ia] = 1;
    d.toggles[ia] = 0;
    d.checkpoint_done[ia] = 0;
    d.start_x[ia] = p[0];
    d.start_y[ia] = p[1];
    if (a == 0) {
        double *R = d.start_rot + 4 * (size_t)e;
        if (rot) {
            R[0] = rot[4 * e]; R[1] = rot[4 * e + 1]; R[2] = rot[4 * e + 2]; R[3] = rot[4 * e + 3];
        } else {
            double sn, cs;
            sincos(-poses[3 * ((size_t)e * A + d.ego_idx) + 2], &sn, &cs);
            R[0] = cs; R[1] = -sn; R[2] = sn; R[3] = cs;
        }
        d.cur_time[e] = 0.0;
        d.noise_step[e] = 0;
        d.done[e] = 0;
        d.reset_pending[e] = 0;
        d.episode_count[e] += 1;
    }
}

void f110_launch_reset(const F110Dev &d, const double *poses_dev, const double *rot_dev, const uint8_t *mask_dev,
                       cudaStream_t st)
{
    int NA = d.n_envs * d.n_agents;
    k_reset<<<(NA + 127) / 128, 128, 0, st>>>(d, poses_dev, rot_dev, mask_dev);
}

// ------------------------------------------------------------------------------------------------
// fp32 copy of the EDT in CELL units (metres / resolution) for the fast path
// ------------------------------------------------------------------------------------------------
// free cells: min(distance in cells, cap) with cap > max_range in cells (a larger value ends the ray at max range all the same);
// obstacle cells (dist <= eps, where trace_ray stops): the sentinel `sent` > cap (see the fp32 kernels)
__global__ void k_edt32(const double *__restrict__ src, float *__restrict__ dst, size_t cells, double inv_res, double eps,
                        float cap, float sent)
{
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < cells) {
        double v = src[i];
        dst[i] = (v > eps) ? fminf((float)(v * inv_res), cap) : sent;
    }
}
void f110_launch_edt32(const double *src, float *dst, size_t cells, double inv_res, double eps, float cap, float sent, cudaStream_t st)
{
    k_edt32<<<(unsigned)((cells + 255) / 256), 256, 0, st>>>(src, dst, cells, inv_res, eps, cap, sent);
}

// ------------------------------------------------------------------------------------------------
// The model functions of the step on their own (dynamic_models.py, collision_models.py), batched: what the reference
// exposes as f110_gym.envs.dynamic_models / collision_models and pins in its unit tests (f_st_gt, GJK known answers).
// They call the very device functions the step kernels use.
// ------------------------------------------------------------------------------------------------
// vehicle_dynamics_st (:123-176), x [n][7], u [n][2] -> f [n][7]; ks = 1: vehicle_dynamics_ks (:91-121), 5 states
__global__ void __launch_bounds__(128) k_model_dynamics(const double *__restrict__ x, const double *__restrict__ u, VehP p,
                                                        long long n, int ks, double *__restrict__ f)
{
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    if (ks) {
        const double *xi = x + 5 * i;
        double u0 = steering_constraint(xi[2], u[2 * i], p.s_min, p.s_max, p.sv_min, p.sv_max);
        double u1 = accl_constraints(xi[3], u[2 * i + 1], p.v_switch, p.a_max, p.v_min, p.v_max);
        double s4, c4;
        sincos(xi[4], &s4, &c4);
        double *fi = f + 5 * i;
        fi[0] = xi[3] * c4; fi[1] = xi[3] * s4; fi[2] = u0; fi[3] = u1; fi[4] = xi[3] / (p.lf + p.lr) * tan(xi[2]);
        return;
    }
    double xs[7], fs[7];
#pragma unroll
    for (int k = 0; k < 7; k++) xs[k] = x[7 * i + k];
    dynamics_st(xs, u[2 * i], u[2 * i + 1], p, fs);
#pragma unroll
    for (int k = 0; k < 7; k++) f[7 * i + k] = fs[k];
}
// pid (:178-221): in [n][4] = speed, steer, current_speed, current_steer -> out [n][2] = accl, sv
__global__ void __launch_bounds__(128) k_model_pid(const double *__restrict__ in, double max_sv, double max_a, double max_v, double min_v,
                                                   long long n, double *__restrict__ out)
{
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    double accl, sv;
    pid_ctrl(in[4 * i], in[4 * i + 1], in[4 * i + 2], in[4 * i + 3], max_sv, max_a, max_v, min_v, accl, sv);
    out[2 * i] = accl; out[2 * i + 1] = sv;
}
// accl_constraints (:30-60) / steering_constraint (:63-87) on their own: in [n][2] -> out [n]
__global__ void __launch_bounds__(128) k_model_constraint(const double *__restrict__ in, double a, double b, double c, double e_,
                                                          long long n, int steering, double *__restrict__ out)
{
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    out[i] = steering ? steering_constraint(in[2 * i], in[2 * i + 1], a, b, c, e_) : accl_constraints(in[2 * i], in[2 * i + 1], a, b, c, e_);
}
// get_vertices (collision_models.py:238-260): poses [n][3] -> corners [n][4][2]
__global__ void __launch_bounds__(128) k_model_vertices(const double *__restrict__ poses, double length, double width, long long n,
                                                        double *__restrict__ verts)
{
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    double v[4][2];
    get_vertices(poses[3 * i], poses[3 * i + 1], poses[3 * i + 2], length, width, v);
#pragma unroll
    for (int k = 0; k < 4; k++) { verts[8 * i + 2 * k] = v[k][0]; verts[8 * i + 2 * k + 1] = v[k][1]; }
}
// collision (GJK, :114-182): quads a, b [n][4][2] -> hit [n]
__global__ void __launch_bounds__(128) k_model_collision(const double *__restrict__ a, const double *__restrict__ b, long long n,
                                                         uint8_t *__restrict__ hit)
{
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    double va[4][2], vb[4][2];
#pragma unroll
    for (int k = 0; k < 4; k++) {
        va[k][0] = a[8 * i + 2 * k]; va[k][1] = a[8 * i + 2 * k + 1];
        vb[k][0] = b[8 * i + 2 * k]; vb[k][1] = b[8 * i + 2 * k + 1];
    }
    hit[i] = gjk_collision(va, vb) ? 1 : 0;
}

void f110_launch_model_dynamics(const double *x, const double *u, const double *p16, long long n, int ks, double *f, cudaStream_t st)
{
    if (n <= 0) return;
    VehP p = { p16[0], p16[1], p16[2], p16[3], p16[4], p16[5], p16[6], p16[7], p16[8], p16[9], p16[10], p16[11], p16[12], p16[13],
               p16[14], p16[15] };
    k_model_dynamics<<<(unsigned)((n + 127) / 128), 128, 0, st>>>(x, u, p, n, ks, f);
}
void f110_launch_model_pid(const double *in, double max_sv, double max_a, double max_v, double min_v, long long n, double *out, cudaStream_t st)
{
    if (n > 0) k_model_pid<<<(unsigned)((n + 127) / 128), 128, 0, st>>>(in, max_sv, max_a, max_v, min_v, n, out);
}
void f110_launch_model_constraint(const double *in, double a, double b, double c, double e_, long long n, int steering, double *out, cudaStream_t st)
{
    if (n > 0) k_model_constraint<<<(unsigned)((n + 127) / 128), 128, 0, st>>>(in, a, b, c, e_, n, steering, out);
}
void f110_launch_model_vertices(const double *poses, double length, double width, long long n, double *verts, cudaStream_t st)
{
    if (n > 0) k_model_vertices<<<(unsigned)((n + 127) / 128), 128, 0, st>>>(poses, length, width, n, verts);
}
void f110_launch_model_collision(const double *a, const double *b, long long n, uint8_t *hit, cudaStream_t st)
{
    if (n > 0) k_model_collision<<<(unsigned)((n + 127) / 128), 128, 0, st>>>(a, b, n, hit);
}
