/*
 * f110_oracle.c -- CPU restatement of the reference hot path (TEST INFRASTRUCTURE, NOT PRODUCT).
 *
 * Plain C restatement of WE-Autopilot/red_gym's f110_gym `Simulator.step` path, written from the
 * reference's algorithm (never copied from its sources) so the CUDA product can be checked
 * against it anywhere (the GPU box has no reference tree).  Every function cites the reference
 * file:line it follows (paths relative to /root/reference/gym/f110_gym/envs/).
 *
 * PARITY PINNING: this oracle is pinned against outputs of the UNMODIFIED reference run in the
 * build container (tests/golden/ (.npz), produced by oracle/gen_golden.py) by tests/test_oracle_golden.py.
 * It uses the same libm (glibc sin/cos/tan/atan2) that Numba lowers to, is compiled with
 * -ffp-contract=off (Numba emits no FMA / no fast-math), and keeps the reference's evaluation
 * order, so states agree bit-for-bit on the build machine.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
 * load this library.  The product (red_gym_b200) never does.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <pthread.h>
#include <unistd.h>

#define F110O_NPARAMS 18
/* params order == keys of the reference params dict (f110_env.py:128) */
enum { P_MU, P_CSF, P_CSR, P_LF, P_LR, P_H, P_M, P_I, P_SMIN, P_SMAX, P_SVMIN, P_SVMAX,
       P_VSWITCH, P_AMAX, P_VMIN, P_VMAX, P_WIDTH, P_LENGTH };

typedef struct f110o {
    int n_envs, n_agents, n_beams, integrator /*1 RK4, 2 Euler: base_classes.py:40-42*/, ego_idx, theta_dis;
    double timestep, fov, eps, max_range, theta_index_increment, ttc_thresh;
    /* map (laser_models.py:383-427) */
    int map_h, map_w;
    double res, ox, oy, oc, os;
    double *dt;
    /* tables */
    double *sines, *cosines;                     /* theta_dis  (laser_models.py:379-381) */
    double *scan_angles, *beam_cos, *side_dist;  /* n_beams    (base_classes.py:120-156) */
    double *params;                              /* n_agents * 18 */
    /* per (env, agent) */
    double *state;      /* 7: x, y, steer, v, yaw, yaw_rate, slip (base_classes.py:95-96) */
    double *steer_buf;  /* 2, index 0 = newest */
    int *steer_len;
    double *accel, *sv;
    unsigned char *in_collision;
    double *scans;      /* n_beams */
    double *collisions, *collision_idx;
    /* lap bookkeeping (f110_env.py:166-187) */
    double *toggles, *lap_counts, *lap_times, *start_x, *start_y;
    unsigned char *near_start;
    /* per env */
    double *start_rot;  /* 4 */
    double *cur_time;
    int64_t *noise_step;
    unsigned char *done;
    uint64_t lookups;   /* EDT lookups since last clear (for the algorithmic-bytes figure) */
} f110o;

/* ------------------------------------------------------------------------------------------ */
/* laser_models.py:56-86 xy_2_rc + :89-104 distance_transform.  Out of bounds -> (r,c)=(-1,-1), */
/* which numpy resolves to the LAST cell dt[h-1, w-1].                                          */
static inline double edt_lookup(const f110o *s, double x, double y)
{
    double x_trans = x - s->ox;
    double y_trans = y - s->oy;
    double x_rot = x_trans * s->oc + y_trans * s->os;
    double y_rot = -x_trans * s->os + y_trans * s->oc;
    long r, c;
    if (x_rot < 0 || x_rot >= s->map_w * s->res || y_rot < 0 || y_rot >= s->map_h * s->res) {
        c = s->map_w - 1;
        r = s->map_h - 1;
    } else {
        c = (long)(x_rot / s->res);
        r = (long)(y_rot / s->res);
    }
    return s->dt[r * (long)s->map_w + c];
}

/* laser_models.py:107-146 trace_ray */
static inline double trace_ray(const f110o *s, double x, double y, double theta_index, uint64_t *lookups)
{
    long ti = (long)theta_index;
    double sn = s->sines[ti];
    double cs = s->cosines[ti];
    double d = edt_lookup(s, x, y);
    double total = d;
    uint64_t n = 1;
    while (d > s->eps && total <= s->max_range) {
        x += d * cs;
        y += d * sn;
        d = edt_lookup(s, x, y);
        total += d;
        n++;
    }
    if (total > s->max_range) total = s->max_range;
    *lookups += n;
    return total;
}

/* laser_models.py:149-186 get_scan */
static void get_scan(const f110o *s, double px, double py, double yaw, double *scan, uint64_t *lookups)
{
    double td = (double)s->theta_dis;
    double theta_index = td * (yaw - s->fov / 2.) / (2. * M_PI);
    theta_index = fmod(theta_index, td);
    while (theta_index < 0) theta_index += td;
    for (int i = 0; i < s->n_beams; i++) {
        scan[i] = trace_ray(s, px, py, theta_index, lookups);
        theta_index += s->theta_index_increment;
        while (theta_index >= td) theta_index -= td;
    }
}

/* laser_models.py:188-217 check_ttc_jit (error_model='numpy': x/0 -> inf/nan, compares false) */
static int check_ttc(const f110o *s, const double *scan, double vel)
{
    if (vel != 0.0) {
        for (int i = 0; i < s->n_beams; i++) {
            double proj_vel = vel * s->beam_cos[i];
            double ttc = (scan[i] - s->side_dist[i]) / proj_vel;
            if (ttc < s->ttc_thresh && ttc >= 0.0) return 1;
        }
    }
    return 0;
}

/* collision_models.py:219-260 get_trmtx + get_vertices: corners rl, rr, fr, fl.              */
/* H.dot([lx, ly, 0, 1]) row 0 = cos*lx + (-sin)*ly + 0*0 + x*1, accumulated left to right;   */
/* the homogeneous divide by 1.0 is exact.                                                    */
static void get_vertices(double x, double y, double th, double length, double width, double v[4][2])
{
    double c = cos(th), sn = sin(th);
    const double lx[4] = { -length / 2, -length / 2, length / 2, length / 2 };
    const double ly[4] = { width / 2, -width / 2, -width / 2, width / 2 };
    for (int k = 0; k < 4; k++) {
        v[k][0] = ((c * lx[k] + (-sn) * ly[k]) + 0. * 0.) + x * 1.;
        v[k][1] = ((sn * lx[k] + c * ly[k]) + 0. * 0.) + y * 1.;
    }
}

/* collision_models.py:82-110 indexOfFurthestPoint (np.argmax: first maximum) + support */
static inline int furthest(const double v[4][2], double dx, double dy)
{
    int best = 0;
    double bv = v[0][0] * dx + v[0][1] * dy;
    for (int k = 1; k < 4; k++) {
        double t = v[k][0] * dx + v[k][1] * dy;
        if (t > bv) { bv = t; best = k; }
    }
    return best;
}
static inline void support(const double v1[4][2], const double v2[4][2], double dx, double dy, double out[2])
{
    int i = furthest(v1, dx, dy);
    int j = furthest(v2, -dx, -dy);
    out[0] = v1[i][0] - v2[j][0];
    out[1] = v1[i][1] - v2[j][1];
}
/* collision_models.py:52-64 tripleProduct(a,b,c) = b*(a.c) - a*(b.c) */
static inline void triple(const double a[2], const double b[2], const double c[2], double out[2])
{
    double ac = a[0] * c[0] + a[1] * c[1];
    double bc = b[0] * c[0] + b[1] * c[1];
    out[0] = b[0] * ac - a[0] * bc;
    out[1] = b[1] * ac - a[1] * bc;
}

/* collision_models.py:114-182 collision (GJK).  iter_count only advances on the triangle branch. */
static int gjk_collision(const double v1[4][2], const double v2[4][2])
{
    int index = 0;
    double simplex[3][2];
    double p1[2] = { (((v1[0][0] + v1[1][0]) + v1[2][0]) + v1[3][0]) / 4, (((v1[0][1] + v1[1][1]) + v1[2][1]) + v1[3][1]) / 4 };
    double p2[2] = { (((v2[0][0] + v2[1][0]) + v2[2][0]) + v2[3][0]) / 4, (((v2[0][1] + v2[1][1]) + v2[2][1]) + v2[3][1]) / 4 };
    double d[2] = { p1[0] - p2[0], p1[1] - p2[1] };
    double a[2];
    if (d[0] == 0 && d[1] == 0) d[0] = 1.0;
    support(v1, v2, d[0], d[1], a);
    simplex[0][0] = a[0]; simplex[0][1] = a[1];
    if (d[0] * a[0] + d[1] * a[1] <= 0) return 0;
    d[0] = -a[0]; d[1] = -a[1];
    int iter_count = 0;
    while (iter_count < 1000) {
        support(v1, v2, d[0], d[1], a);
        index += 1;
        simplex[index][0] = a[0]; simplex[index][1] = a[1];
        if (d[0] * a[0] + d[1] * a[1] <= 0) return 0;
        double ao[2] = { -a[0], -a[1] };
        if (index < 2) {
            double ab[2] = { simplex[0][0] - a[0], simplex[0][1] - a[1] };
            triple(ab, ao, ab, d);
            if (sqrt(d[0] * d[0] + d[1] * d[1]) < 1e-10) {   /* perpendicular(ab) :35-48 */
                d[0] = ab[1];
                d[1] = -1 * ab[0];
            }
            continue;
        }
        double ab[2] = { simplex[1][0] - a[0], simplex[1][1] - a[1] };
        double ac[2] = { simplex[0][0] - a[0], simplex[0][1] - a[1] };
        double acperp[2];
        triple(ab, ac, ac, acperp);
        if (acperp[0] * ao[0] + acperp[1] * ao[1] >= 0) {
            d[0] = acperp[0]; d[1] = acperp[1];
        } else {
            double abperp[2];
            triple(ac, ab, ab, abperp);
            if (abperp[0] * ao[0] + abperp[1] * ao[1] < 0) return 1;
            simplex[0][0] = simplex[1][0]; simplex[0][1] = simplex[1][1];
            d[0] = abperp[0]; d[1] = abperp[1];
        }
        simplex[1][0] = simplex[2][0]; simplex[1][1] = simplex[2][1];
        index -= 1;
        iter_count += 1;
    }
    return 0;
}

/* collision_models.py:185-212 collision_multiple: all pairs i<j, last writer wins for idx */
static void collision_multiple(int n, const double (*verts)[4][2], double *collisions, double *collision_idx)
{
    for (int i = 0; i < n; i++) { collisions[i] = 0.; collision_idx[i] = -1.; }
    for (int i = 0; i < n - 1; i++)
        for (int j = i + 1; j < n; j++)
            if (gjk_collision(verts[i], verts[j])) {
                collisions[i] = 1.; collisions[j] = 1.;
                collision_idx[i] = j; collision_idx[j] = i;
            }
}

/* ------------------------------------------------------------------------------------------ */
/* laser_models.py:250-280 get_range (+ :220-247 cross / are_collinear) */
static double get_range(double ox_, double oy_, double beam_theta, const double va[2], const double vb[2])
{
    double v1[2] = { ox_ - va[0], oy_ - va[1] };
    double v2[2] = { vb[0] - va[0], vb[1] - va[1] };
    double v3[2] = { cos(beam_theta + M_PI / 2.), sin(beam_theta + M_PI / 2.) };
    double denom = v2[0] * v3[0] + v2[1] * v3[1];
    double distance = INFINITY;
    if (fabs(denom) > 0.0) {
        double d1 = (v2[0] * v1[1] - v2[1] * v1[0]) / denom;
        double d2 = (v1[0] * v3[0] + v1[1] * v3[1]) / denom;
        if (d1 >= 0.0 && d2 >= 0.0 && d2 <= 1.0) distance = d1;
    } else {
        /* are_collinear(o, va, vb): ba = va - o, ca = o - vb, |cross(ba, ca)| < 1e-8 */
        double ba[2] = { va[0] - ox_, va[1] - oy_ };
        double ca[2] = { ox_ - vb[0], oy_ - vb[1] };
        if (fabs(ba[0] * ca[1] - ba[1] * ca[0]) < 1e-8) {
            double da = sqrt((va[0] - ox_) * (va[0] - ox_) + (va[1] - oy_) * (va[1] - oy_));
            double db = sqrt((vb[0] - ox_) * (vb[0] - ox_) + (vb[1] - oy_) * (vb[1] - oy_));
            distance = da < db ? da : db;
        }
    }
    return distance;
}

/* laser_models.py:283-315 get_blocked_view_indices: np.argmin(|scan_angles - a|) = first minimum */
static void blocked_view(const f110o *s, double px, double py, double yaw, const double verts[4][2], int *lo, int *hi)
{
    double ex = cos(yaw), ey = sin(yaw);
    double ego_ang = atan2(ey, ex);
    int mn = 0, mx = 0;
    for (int k = 0; k < 4; k++) {
        double vx = verts[k][0] - px, vy = verts[k][1] - py;
        double nrm = sqrt(vx * vx + vy * vy);
        double ux = vx / nrm, uy = vy / nrm;
        double angle = ego_ang - atan2(uy, ux);
        if (angle > M_PI) angle = angle - 2 * M_PI;
        else if (angle < -M_PI) angle = angle + 2 * M_PI;
        double target = -angle;
        int best = 0;
        double bv = fabs(s->scan_angles[0] - target);
        for (int i = 1; i < s->n_beams; i++) {
            double t = fabs(s->scan_angles[i] - target);
            if (t < bv) { bv = t; best = i; }   /* NaN never wins, like np.argmin's first-min... */
        }
        if (isnan(target)) best = 0;            /* ...np.argmin returns the first NaN = index 0 */
        if (k == 0) { mn = mx = best; }
        else { if (best < mn) mn = best; if (best > mx) mx = best; }
    }
    *lo = mn; *hi = mx;
}

/* laser_models.py:318-346 ray_cast: in place, edges (v0,v1),(v1,v2),(v2,v3),(v3,v0) */
static void ray_cast(const f110o *s, double px, double py, double yaw, double *scan, const double verts[4][2])
{
    int lo, hi;
    blocked_view(s, px, py, yaw, verts, &lo, &hi);
    for (int i = lo; i <= hi; i++)
        for (int j = 0; j < 4; j++) {
            double r = get_range(px, py, yaw + s->scan_angles[i], verts[j], verts[(j + 1) & 3]);
            if (r < scan[i]) scan[i] = r;
        }
}

/* ------------------------------------------------------------------------------------------ */
/* dynamic_models.py:29-60 accl_constraints */
static inline double accl_constraints(double vel, double accl, double v_switch, double a_max, double v_min, double v_max)
{
    double pos_limit;
    if (vel > v_switch) pos_limit = a_max * v_switch / vel;
    else pos_limit = a_max;
    if ((vel <= v_min && accl <= 0) || (vel >= v_max && accl >= 0)) accl = 0.;
    else if (accl <= -a_max) accl = -a_max;
    else if (accl >= pos_limit) accl = pos_limit;
    return accl;
}
/* dynamic_models.py:62-87 steering_constraint */
static inline double steering_constraint(double steer, double sv, double s_min, double s_max, double sv_min, double sv_max)
{
    if ((steer <= s_min && sv <= 0) || (steer >= s_max && sv >= 0)) sv = 0.;
    else if (sv <= sv_min) sv = sv_min;
    else if (sv >= sv_max) sv = sv_max;
    return sv;
}

/* dynamic_models.py:123-176 vehicle_dynamics_st (kinematic branch :152-160 calls :90-121,   */
/* which re-applies both constraints to the already constrained input)                        */
void f110o_dynamics_st(const double *x, const double *u_init, const double *p, double *f)
{
    const double mu = p[P_MU], C_Sf = p[P_CSF], C_Sr = p[P_CSR], lf = p[P_LF], lr = p[P_LR], h = p[P_H],
                 m = p[P_M], I = p[P_I];
    const double g = 9.81;
    double u0 = steering_constraint(x[2], u_init[0], p[P_SMIN], p[P_SMAX], p[P_SVMIN], p[P_SVMAX]);
    double u1 = accl_constraints(x[3], u_init[1], p[P_VSWITCH], p[P_AMAX], p[P_VMIN], p[P_VMAX]);
    if (fabs(x[3]) < 0.5) {
        double lwb = lf + lr;
        double k0 = steering_constraint(x[2], u0, p[P_SMIN], p[P_SMAX], p[P_SVMIN], p[P_SVMAX]);
        double k1 = accl_constraints(x[3], u1, p[P_VSWITCH], p[P_AMAX], p[P_VMIN], p[P_VMAX]);
        f[0] = x[3] * cos(x[4]);
        f[1] = x[3] * sin(x[4]);
        f[2] = k0;
        f[3] = k1;
        f[4] = x[3] / lwb * tan(x[2]);
        double c2 = cos(x[2]);
        f[5] = u1 / lwb * tan(x[2]) + x[3] / (lwb * (c2 * c2)) * u0;
        f[6] = 0.;
    } else {
        f[0] = x[3] * cos(x[6] + x[4]);
        f[1] = x[3] * sin(x[6] + x[4]);
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

/* dynamic_models.py:178-221 pid -> (accl, sv) */
void f110o_pid(double speed, double steer, double cur_speed, double cur_steer, double max_sv, double max_a,
               double max_v, double min_v, double *accl, double *sv)
{
    double steer_diff = steer - cur_steer;
    if (fabs(steer_diff) > 1e-4) *sv = (steer_diff / fabs(steer_diff)) * max_sv;
    else *sv = 0.0;
    double vel_diff = speed - cur_speed;
    double kp;
    if (cur_speed > 0.) {
        if (vel_diff > 0) kp = 10.0 * max_a / max_v;
        else kp = 10.0 * max_a / (-min_v);
    } else {
        if (vel_diff > 0) kp = 2.0 * max_a / max_v;
        else kp = 2.0 * max_a / (-min_v);
    }
    *accl = kp * vel_diff;
}

/* base_classes.py:254-407 RaceCar.update_pose up to (not including) the scan */
static void update_pose(f110o *s, int e, int a, double raw_steer, double vel)
{
    int ia = e * s->n_agents + a;
    double *st = s->state + 7 * ia;
    double *buf = s->steer_buf + 2 * ia;
    const double *p = s->params + F110O_NPARAMS * a;
    double steer;
    if (s->steer_len[ia] < 2) {           /* :270-272 np.append(raw, buf): newest first */
        steer = 0.;
        buf[1] = buf[0];
        buf[0] = raw_steer;
        s->steer_len[ia] += 1;
    } else {                              /* :274-276 pop the oldest, push the newest */
        steer = buf[1];
        buf[1] = buf[0];
        buf[0] = raw_steer;
    }
    double accl, sv;
    f110o_pid(vel, steer, st[3], st[2], p[P_SVMAX], p[P_AMAX], p[P_VMAX], p[P_VMIN], &accl, &sv);
    double u[2] = { sv, accl };
    double dt = s->timestep;
    if (s->integrator == 1) {             /* RK4 :282-371 */
        double k1[7], k2[7], k3[7], k4[7], xs[7];
        f110o_dynamics_st(st, u, p, k1);
        for (int i = 0; i < 7; i++) xs[i] = st[i] + dt * (k1[i] / 2);
        f110o_dynamics_st(xs, u, p, k2);
        for (int i = 0; i < 7; i++) xs[i] = st[i] + dt * (k2[i] / 2);
        f110o_dynamics_st(xs, u, p, k3);
        for (int i = 0; i < 7; i++) xs[i] = st[i] + dt * k3[i];
        f110o_dynamics_st(xs, u, p, k4);
        double w = dt * (1.0 / 6.0);
        for (int i = 0; i < 7; i++) st[i] = st[i] + w * (((k1[i] + 2 * k2[i]) + 2 * k3[i]) + k4[i]);
    } else {                              /* Euler :373-393 */
        double f[7];
        f110o_dynamics_st(st, u, p, f);
        for (int i = 0; i < 7; i++) st[i] = st[i] + dt * f[i];
    }
    if (st[4] > 2 * M_PI) st[4] = st[4] - 2 * M_PI;   /* :398-402 single wrap */
    else if (st[4] < 0) st[4] = st[4] + 2 * M_PI;
}

/* ------------------------------------------------------------------------------------------ */
/* base_classes.py:546-605 Simulator.step + f110_env.py:261-302 F110Env.step/_check_done for   */
/* one env.  noise_row: n_beams doubles (rng.normal(0, .01, B), laser_models.py:450-452), or   */
/* NULL for no noise.  All agents of an env see the same row (same seed, base_classes.py:202). */
static void step_env(f110o *s, int e, const double *actions, const double *noise_row, uint64_t *lookups)
{
    const int A = s->n_agents, B = s->n_beams;
    double poses[32][3];
    double verts[32][4][2];
    /* :561-567 every agent integrates and scans at its NEW pose */
    for (int a = 0; a < A; a++) {
        int ia = e * A + a;
        update_pose(s, e, a, actions[2 * a + 0], actions[2 * a + 1]);
        double *st = s->state + 7 * ia;
        double *scan = s->scans + (size_t)ia * B;
        get_scan(s, st[0], st[1], st[4], scan, lookups);
        if (noise_row) for (int i = 0; i < B; i++) scan[i] += noise_row[i];
        poses[a][0] = st[0]; poses[a][1] = st[1]; poses[a][2] = st[4];
    }
    /* :570 -> :529-543 GJK on post-dynamics, pre-TTC poses, with the Simulator's construct-time params */
    const double *p0 = s->params;   /* see f110o_set_params: slot -1 semantics handled by caller */
    for (int a = 0; a < A; a++)
        get_vertices(poses[a][0], poses[a][1], poses[a][2], s->params[F110O_NPARAMS * A + 1], s->params[F110O_NPARAMS * A + 0], verts[a]);
    (void)p0;
    collision_multiple(A, (const double (*)[4][2])verts, s->collisions + e * A, s->collision_idx + e * A);
    /* :572-582 iTTC on the noisy static-map scan, THEN opponents are ray-cast into it */
    for (int a = 0; a < A; a++) {
        int ia = e * A + a;
        double *st = s->state + 7 * ia;
        double *scan = s->scans + (size_t)ia * B;
        int hit = check_ttc(s, scan, st[3]);          /* :227-252 */
        if (hit) {
            st[3] = st[4] = st[5] = st[6] = 0.;       /* state[3:] = 0 -- yaw included */
            s->accel[ia] = 0.; s->sv[ia] = 0.;
        }
        s->in_collision[ia] = (unsigned char)hit;
        const double *pa = s->params + F110O_NPARAMS * a;
        for (int o = 0; o < A; o++) {                 /* :204-225 opponents in index order */
            if (o == a) continue;
            double ov[4][2];
            get_vertices(poses[o][0], poses[o][1], poses[o][2], pa[P_LENGTH], pa[P_WIDTH], ov);
            ray_cast(s, st[0], st[1], st[4], scan, (const double (*)[2])ov);
        }
        if (hit) s->collisions[ia] = 1.;
    }
    /* f110_env.py:292-299 */
    s->cur_time[e] = s->cur_time[e] + s->timestep;
    const double *R = s->start_rot + 4 * e;
    int all_done = 1;
    for (int a = 0; a < A; a++) {                     /* :202-244 _check_done */
        int ia = e * A + a;
        const double *st = s->state + 7 * ia;
        double dx = st[0] - s->start_x[ia];
        double dy = st[1] - s->start_y[ia];
        double d0 = R[0] * dx + R[1] * dy;
        double ty = R[2] * dx + R[3] * dy;
        if (ty > 2) ty -= 2;
        else if (ty < -2) ty = -2 - ty;
        else ty = 0;
        double dist2 = d0 * d0 + ty * ty;
        int closes = dist2 <= 0.1;
        if (closes && !s->near_start[ia]) { s->near_start[ia] = 1; s->toggles[ia] += 1; }
        else if (!closes && s->near_start[ia]) { s->near_start[ia] = 0; s->toggles[ia] += 1; }
        s->lap_counts[ia] = floor(s->toggles[ia] / 2);
        if (s->toggles[ia] < 4) s->lap_times[ia] = s->cur_time[e];
        if (!(s->toggles[ia] >= 4)) all_done = 0;
    }
    s->done[e] = (unsigned char)((s->collisions[e * A + s->ego_idx] != 0.) || all_done);
}

/* ------------------------------------------------------------------------------------------ */
/* C API (ctypes).  params slot n_agents holds {width,length} used by Simulator.check_collision */
f110o *f110o_create(int n_envs, int n_agents, int n_beams, int integrator, int ego_idx, int theta_dis,
                    double timestep, double fov, double eps, double max_range, double ttc_thresh)
{
    if (n_agents > 32 || n_agents < 1) return NULL;
    f110o *s = (f110o *)calloc(1, sizeof(f110o));
    s->n_envs = n_envs; s->n_agents = n_agents; s->n_beams = n_beams; s->integrator = integrator;
    s->ego_idx = ego_idx; s->theta_dis = theta_dis; s->timestep = timestep; s->fov = fov; s->eps = eps;
    s->max_range = max_range; s->ttc_thresh = ttc_thresh;
    /* laser_models.py:367-368 */
    double angle_increment = fov / (n_beams - 1);
    s->theta_index_increment = theta_dis * angle_increment / (2. * M_PI);
    size_t na = (size_t)n_envs * n_agents;
    s->sines = calloc(theta_dis, 8); s->cosines = calloc(theta_dis, 8);
    s->scan_angles = calloc(n_beams, 8); s->beam_cos = calloc(n_beams, 8); s->side_dist = calloc(n_beams, 8);
    s->params = calloc((size_t)(n_agents + 1) * F110O_NPARAMS, 8);
    s->state = calloc(na * 7, 8); s->steer_buf = calloc(na * 2, 8); s->steer_len = calloc(na, sizeof(int));
    s->accel = calloc(na, 8); s->sv = calloc(na, 8); s->in_collision = calloc(na, 1);
    s->scans = calloc(na * n_beams, 8);
    s->collisions = calloc(na, 8); s->collision_idx = calloc(na, 8);
    s->toggles = calloc(na, 8); s->lap_counts = calloc(na, 8); s->lap_times = calloc(na, 8);
    s->start_x = calloc(na, 8); s->start_y = calloc(na, 8); s->near_start = calloc(na, 1);
    s->start_rot = calloc((size_t)n_envs * 4, 8); s->cur_time = calloc(n_envs, 8);
    s->noise_step = calloc(n_envs, sizeof(int64_t)); s->done = calloc(n_envs, 1);
    for (size_t i = 0; i < na; i++) { s->near_start[i] = 1; s->collision_idx[i] = -1.; }
    for (int e = 0; e < n_envs; e++) { s->start_rot[4 * e] = 1.; s->start_rot[4 * e + 3] = 1.; }
    return s;
}

void f110o_destroy(f110o *s)
{
    if (!s) return;
    free(s->dt); free(s->sines); free(s->cosines); free(s->scan_angles); free(s->beam_cos); free(s->side_dist);
    free(s->params); free(s->state); free(s->steer_buf); free(s->steer_len); free(s->accel); free(s->sv);
    free(s->in_collision); free(s->scans); free(s->collisions); free(s->collision_idx); free(s->toggles);
    free(s->lap_counts); free(s->lap_times); free(s->start_x); free(s->start_y); free(s->near_start);
    free(s->start_rot); free(s->cur_time); free(s->noise_step); free(s->done); free(s);
}

void f110o_set_map(f110o *s, const double *dt, int h, int w, double res, double ox, double oy, double oc, double os)
{
    free(s->dt);
    s->dt = malloc((size_t)h * w * 8);
    memcpy(s->dt, dt, (size_t)h * w * 8);
    s->map_h = h; s->map_w = w; s->res = res; s->ox = ox; s->oy = oy; s->oc = oc; s->os = os;
}

void f110o_set_tables(f110o *s, const double *sines, const double *cosines, const double *scan_angles,
                      const double *beam_cos, const double *side_dist)
{
    memcpy(s->sines, sines, (size_t)s->theta_dis * 8); memcpy(s->cosines, cosines, (size_t)s->theta_dis * 8);
    memcpy(s->scan_angles, scan_angles, (size_t)s->n_beams * 8); memcpy(s->beam_cos, beam_cos, (size_t)s->n_beams * 8);
    memcpy(s->side_dist, side_dist, (size_t)s->n_beams * 8);
}

/* agent_idx in [0, n_agents): that agent (base_classes.py:522-524); agent_idx == n_agents: the   */
/* Simulator-level {width,length} that check_collision keeps from construction (:477, :542).      */
void f110o_set_params(f110o *s, const double *params18, int agent_idx)
{
    if (agent_idx == s->n_agents) {
        s->params[F110O_NPARAMS * s->n_agents + 0] = params18[P_WIDTH];
        s->params[F110O_NPARAMS * s->n_agents + 1] = params18[P_LENGTH];
    } else {
        memcpy(s->params + F110O_NPARAMS * agent_idx, params18, F110O_NPARAMS * 8);
    }
}

/* f110_env.py:304-347 reset (+ base_classes.py:181-202, :607-623), then the zero-action step. */
/* mask: n_envs bytes or NULL (= all).  noise as in f110o_step.                                */
void f110o_step_impl(f110o *s, const double *actions, const double *noise, int64_t noise_rows, const unsigned char *mask, int nthreads);

void f110o_reset(f110o *s, const double *poses, const unsigned char *mask, const double *noise, int64_t noise_rows, int nthreads)
{
    const int A = s->n_agents;
    for (int e = 0; e < s->n_envs; e++) {
        if (mask && !mask[e]) continue;
        s->cur_time[e] = 0.0;
        s->noise_step[e] = 0;
        double th = poses[((size_t)e * A + s->ego_idx) * 3 + 2];
        double *R = s->start_rot + 4 * e;
        R[0] = cos(-th); R[1] = -sin(-th); R[2] = sin(-th); R[3] = cos(-th);   /* :329 */
        for (int a = 0; a < A; a++) {
            int ia = e * A + a;
            const double *p = poses + (size_t)ia * 3;
            s->collisions[ia] = 0.; s->near_start[ia] = 1; s->toggles[ia] = 0.;
            s->start_x[ia] = p[0]; s->start_y[ia] = p[1];
            s->accel[ia] = 0.; s->sv[ia] = 0.; s->in_collision[ia] = 0;
            double *st = s->state + 7 * ia;
            memset(st, 0, 7 * 8);
            st[0] = p[0]; st[1] = p[1]; st[4] = p[2];
            s->steer_len[ia] = 0;
        }
    }
    double *zero = calloc((size_t)s->n_envs * A * 2, 8);
    f110o_step_impl(s, zero, noise, noise_rows, mask, nthreads);
    free(zero);
}

/* minimal pthread parallel-for (dynamic chunks of 16 envs) -- no OpenMP runtime needed */
typedef struct {
    f110o *s; const double *actions; const double *noise; int64_t noise_rows; const unsigned char *mask;
    const double *poses; double *out; uint64_t *per_scan; int n; int mode; volatile int next; uint64_t lookups;
    pthread_mutex_t mu;
} pf_job;

static void *pf_worker(void *arg)
{
    pf_job *j = (pf_job *)arg;
    f110o *s = j->s;
    uint64_t lookups = 0;
    for (;;) {
        int b = __sync_fetch_and_add(&j->next, 16);
        if (b >= j->n) break;
        int e_end = b + 16 < j->n ? b + 16 : j->n;
        for (int e = b; e < e_end; e++) {
            if (j->mode == 0) {
                if (j->mask && !j->mask[e]) continue;
                const double *row = NULL;
                if (j->noise) {
                    int64_t r = s->noise_step[e];
                    if (r >= j->noise_rows) r = r % j->noise_rows;
                    row = j->noise + r * s->n_beams;
                }
                step_env(s, e, j->actions + (size_t)e * s->n_agents * 2, row, &lookups);
                s->noise_step[e] += 1;
            } else {
                uint64_t l = 0;
                get_scan(s, j->poses[3 * e], j->poses[3 * e + 1], j->poses[3 * e + 2], j->out + (size_t)e * s->n_beams, &l);
                if (j->per_scan) j->per_scan[e] = l;
                lookups += l;
            }
        }
    }
    pthread_mutex_lock(&j->mu);
    j->lookups += lookups;
    pthread_mutex_unlock(&j->mu);
    return NULL;
}

static void pf_run(pf_job *j, int nthreads)
{
    if (nthreads <= 0) nthreads = (int)sysconf(_SC_NPROCESSORS_ONLN);
    if (nthreads > 256) nthreads = 256;
    if (nthreads > (j->n + 15) / 16) nthreads = (j->n + 15) / 16;
    if (nthreads < 1) nthreads = 1;
    pthread_mutex_init(&j->mu, NULL);
    j->next = 0; j->lookups = 0;
    if (nthreads == 1) { pf_worker(j); }
    else {
        pthread_t th[256];
        for (int t = 0; t < nthreads; t++) pthread_create(&th[t], NULL, pf_worker, j);
        for (int t = 0; t < nthreads; t++) pthread_join(th[t], NULL);
    }
    pthread_mutex_destroy(&j->mu);
    j->s->lookups += j->lookups;
}

void f110o_step_impl(f110o *s, const double *actions, const double *noise, int64_t noise_rows, const unsigned char *mask, int nthreads)
{
    pf_job j; memset(&j, 0, sizeof(j));
    j.s = s; j.actions = actions; j.noise = noise; j.noise_rows = noise_rows; j.mask = mask; j.n = s->n_envs; j.mode = 0;
    pf_run(&j, nthreads);
}

void f110o_step(f110o *s, const double *actions, const double *noise, int64_t noise_rows, int nthreads)
{
    f110o_step_impl(s, actions, noise, noise_rows, NULL, nthreads);
}

/* raycast only (ScanSimulator2D.scan(pose, None), laser_models.py:429-454) over n poses */
void f110o_scan(f110o *s, const double *poses, int n, double *out, uint64_t *lookups_per_scan, int nthreads)
{
    pf_job j; memset(&j, 0, sizeof(j));
    j.s = s; j.poses = poses; j.out = out; j.per_scan = lookups_per_scan; j.n = n; j.mode = 1;
    pf_run(&j, nthreads);
}

/* stand-alone kernels for known-answer tests */
int f110o_gjk(const double *v1, const double *v2) { return gjk_collision((const double (*)[2])v1, (const double (*)[2])v2); }
void f110o_collision_multiple(int n, const double *verts, double *collisions, double *collision_idx)
{ collision_multiple(n, (const double (*)[4][2])verts, collisions, collision_idx); }
void f110o_get_vertices(const double *pose, double length, double width, double *out)
{ get_vertices(pose[0], pose[1], pose[2], length, width, (double (*)[2])out); }
void f110o_ray_cast(f110o *s, const double *pose, double *scan, const double *verts)
{ ray_cast(s, pose[0], pose[1], pose[2], scan, (const double (*)[2])verts); }
int f110o_check_ttc(f110o *s, const double *scan, double vel) { return check_ttc(s, scan, vel); }

/* field access: pointers into the handle (numpy wraps them without copying) */
double *f110o_state(f110o *s) { return s->state; }
double *f110o_scans(f110o *s) { return s->scans; }
double *f110o_collisions(f110o *s) { return s->collisions; }
double *f110o_collision_idx(f110o *s) { return s->collision_idx; }
double *f110o_toggles(f110o *s) { return s->toggles; }
double *f110o_lap_counts(f110o *s) { return s->lap_counts; }
double *f110o_lap_times(f110o *s) { return s->lap_times; }
double *f110o_cur_time(f110o *s) { return s->cur_time; }
unsigned char *f110o_done(f110o *s) { return s->done; }
unsigned char *f110o_in_collision(f110o *s) { return s->in_collision; }
int64_t *f110o_noise_step(f110o *s) { return s->noise_step; }
uint64_t f110o_lookups(f110o *s, int clear) { uint64_t v = s->lookups; if (clear) s->lookups = 0; return v; }
