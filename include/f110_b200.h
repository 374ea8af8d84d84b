/*
 * f110_b200.h -- C ABI of the B200-native batched F1TENTH simulator (libf110_b200.so).
 *
 * Drop-in boundary for ONE hot path of WE-Autopilot/red_gym: `Simulator.step` of the fork's
 * f110_gym (dynamics -> 1080-beam EDT sphere-traced lidar -> iTTC -> GJK -> lap bookkeeping)
 * over N independent environments.  The reference has no FFI of its own (pure Python + Numba,
 * one env per process); the entry points below are what a binding for that path binds, each
 * citing the reference interface it replaces (paths relative to gym/f110_gym/envs/).
 *
 * Conventions
 *   - plain C: pointers + sizes only, no C++/torch types.  Every function returns 0 on success
 *     or a negative f110_status; f110_last_error() gives the message (thread-local).
 *   - `stream` is a cudaStream_t passed as void* (NULL = default stream).  Calls are
 *     asynchronous on that stream unless stated; one handle per GPU, not thread-safe.
 *   - the library owns all device memory; f110_get_buffers() exposes the device pointers so a
 *     host framework (PyTorch, CuPy, ...) can wrap them without copies.
 *   - layouts: per-agent arrays are [n_envs][n_agents] row-major, scans are
 *     [n_envs][n_agents][n_beams]; the 7 state fields are SoA: state[field][n_envs*n_agents].
 *   - there is NO CPU fallback: every entry point fails with F110_ERR_CUDA when no device works.
 */
#ifndef F110_B200_H
#define F110_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define F110_ABI_VERSION 2
#define F110_NUM_PARAMS 18   /* order: mu C_Sf C_Sr lf lr h m I s_min s_max sv_min sv_max v_switch a_max
                                v_min v_max width length  (keys of the params dict, f110_env.py:128) */
#define F110_MAX_AGENTS 32

typedef enum {
    F110_OK = 0,
    F110_ERR_INVALID = -1,   /* bad argument (ValueError / IndexError on the Python side) */
    F110_ERR_CUDA = -2,      /* CUDA runtime failure, incl. "no device" */
    F110_ERR_STATE = -3,     /* call order: e.g. step before set_map (laser_models.py:445-446) */
    F110_ERR_NOMEM = -4
} f110_status;

/* Integrator enum values of the reference (base_classes.py:40-42) */
#define F110_INTEGRATOR_RK4 1
#define F110_INTEGRATOR_EULER 2

/* Arithmetic of the lidar path.  FP64 = parity path: fp64 EDT, same op order as the reference,
 * no FMA contraction, noise rows supplied by the host (numpy PCG64 stream).  FP32 = fast path:
 * fp32 EDT and scans, ray marching in cell units with FMA, Philox noise on the device;
 * dynamics/GJK/lap logic stay fp64 in both. */
#define F110_PRECISION_FP64 0
#define F110_PRECISION_FP32 1

/* Storage format of the scans (f110_config.scan_format).  NATIVE: double (FP64 path) / float (FP32 path) metres, what
 * the reference returns (base_classes.py:587-603, converted to the path's type).  U16: opt-in compact observation of the
 * FP32 path with one agent per env: uint16 at 0.5 mm per LSB (range = value / 2000 m; max_range 30 m = 60000),
 * converted in the ray-cast epilogue -- half the HBM and PCIe bytes of a float scan; |error| <= 0.25 mm, i.e. 40 x below
 * the lidar noise of 1 cm; a noisy range below zero is stored as 0. */
#define F110_SCAN_NATIVE 0
#define F110_SCAN_U16 1
#define F110_SCAN_U16_PER_METRE 2000.0

/* Noise source */
#define F110_NOISE_OFF 0      /* ScanSimulator2D.scan(pose, rng=None) */
#define F110_NOISE_TABLE 1    /* host-generated rows: row k = k-th draw after reset (base_classes.py:202) */
#define F110_NOISE_PHILOX 2   /* device counter-based N(0, std) keyed by (seed, env, agent, step, beam) */

typedef struct f110_config {
    int32_t abi_version;      /* F110_ABI_VERSION */
    int32_t n_envs;           /* N independent copies of the reference env */
    int32_t n_agents;         /* kwargs['num_agents'] (f110_env.py:132), <= F110_MAX_AGENTS */
    int32_t n_beams;          /* ScanSimulator2D.num_beams (1080 through the env, base_classes.py:68) */
    int32_t integrator;       /* F110_INTEGRATOR_* (f110_env.py:149-151) */
    int32_t ego_idx;          /* f110_env.py:143 */
    int32_t theta_dis;        /* 2000 (laser_models.py:360) */
    int32_t precision;        /* F110_PRECISION_* */
    int32_t noise_mode;       /* F110_NOISE_* */
    int32_t auto_reset;       /* 0: like the reference (keep stepping after done); 1: a done env restarts from
                                 the reset-pose pool on its next step (that step IS the reset's zero-action step) */
    int32_t raycast_variant;  /* 0 = default; >0 selects an experimental kernel variant (bench only) */
    int32_t env_offset;       /* global id of this handle's env 0 (env sharding): keys the Philox noise and the reset-pool
                                 choice, so per-env results do not depend on how envs are split over GPUs */
    int32_t scan_format;      /* F110_SCAN_* */
    int32_t reserved0;        /* 0 */
    double timestep;          /* f110_env.py:137 */
    double fov;               /* f110_env.py:154-157 */
    double eps;               /* 1e-4  (laser_models.py:360) */
    double max_range;         /* 30.0 */
    double ttc_thresh;        /* 0.005 (base_classes.py:113) */
    double noise_std;         /* 0.01  (laser_models.py:429), used by F110_NOISE_PHILOX */
    uint64_t seed;            /* f110_env.py:103 */
} f110_config;

typedef struct f110_sim f110_sim;

/* Device pointers owned by the handle (valid until f110_destroy). */
typedef struct f110_buffers {
    double *state;          /* [7][N*A]: x, y, steer_angle, v, yaw, yaw_rate, slip (base_classes.py:95) */
    double *obs_block;      /* [10][N*A] contiguous: the 7 state fields, then collisions, lap_times, lap_counts */
    double *collisions;     /* [N][A] 0/1 (base_classes.py:480,582) */
    double *collision_idx;  /* [N][A] -1 or partner (collision_models.py:197) */
    double *lap_times;      /* [N][A] (f110_env.py:175) */
    double *lap_counts;     /* [N][A] */
    int32_t *toggles;       /* [N][A] toggle_list (f110_env.py:183) */
    uint8_t *done;          /* [N] */
    uint8_t *in_collision;  /* [N][A] iTTC flag (base_classes.py:250) */
    void *scans;            /* [N][A][B] double (FP64), float (FP32) or uint16 (F110_SCAN_U16) */
    double *cur_time;       /* [N] */
    int64_t *noise_step;    /* [N] scans drawn since reset */
    double *actions;        /* [N][A][2] device staging for f110_step_host (may also be written by the caller) */
    int64_t *episode_count; /* [N] number of (auto-)resets */
    uint64_t *lookups;      /* [1] EDT lookups counted when count_lookups was enabled */
    uint8_t *checkpoint_done; /* [N][A] toggle_list >= 4 (info['checkpoint_done'], f110_env.py:244,300) */
} f110_buffers;

const char *f110_last_error(void);
int f110_abi_version(void);
int f110_device_count(void);

/* F110Env.__init__ / Simulator.__init__ (f110_env.py:100-194, base_classes.py:459-490).
 * device: CUDA device ordinal. */
int f110_create(const f110_config *cfg, int device, f110_sim **out);
int f110_destroy(f110_sim *sim);

/* ScanSimulator2D.set_map (laser_models.py:383-427) after the host has loaded the image and
 * built the EDT (host pointers; copied).  edt: [h][w] metres, row 0 = bottom row of the image. */
int f110_set_map(f110_sim *sim, const double *edt, int32_t h, int32_t w, double resolution,
                 double origin_x, double origin_y, double origin_c, double origin_s, void *stream);

/* The same, with get_dt (laser_models.py:40-53: resolution * scipy.ndimage.distance_transform_edt(bitmap)) done on the
 * device: free_px is the HOST [h][w] thresholded image (non-zero = free, pixel > 128; row 0 = bottom row), the exact EDT
 * is built by the library (bit-identical to scipy's).  For update_map / per-episode map randomisation. */
int f110_set_map_image(f110_sim *sim, const uint8_t *free_px, int32_t h, int32_t w, double resolution,
                       double origin_x, double origin_y, double origin_c, double origin_s, void *stream);
/* get_dt alone, device to device: free_dev DEVICE [h][w] bytes -> edt_dev DEVICE [h][w] doubles (metres). */
int f110_edt_build(const uint8_t *free_dev, int32_t h, int32_t w, double resolution, double *edt_dev, int device, void *stream);
/* f110_edt_build for callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).
 * No counterpart in the reference.  Same inputs, same output bits; the
 * column-distance scratch that f110_edt_build allocates per call is a fixed static device array inside the library
 * (F110_EDT_NOALLOC_MAX_CELLS * 4 bytes = 16 MiB on every GPU that loads the library, until exit), so the call allocates
 * and frees nothing.  Limits: h * w <= F110_EDT_NOALLOC_MAX_CELLS and w <= F110_EDT_NOALLOC_MAX_WIDTH, else F110_ERR_INVALID.
 * Asynchronous on `stream`.  All calls on a device share the scratch: it keeps the last call's contents (every cell is
 * rewritten before it is read, nothing needs clearing), and calls that overlap on the GPU (different streams without an
 * ordering between them) race -- serialise them. */
#define F110_EDT_NOALLOC_MAX_CELLS (2048 * 2048)
#define F110_EDT_NOALLOC_MAX_WIDTH 16384
int f110_edt_build_noalloc(const uint8_t *free_dev, int32_t h, int32_t w, double resolution, double *edt_dev, int device, void *stream);

/* ---- The model functions of the step on their own, batched (what the reference exposes as
 * f110_gym.envs.dynamic_models / collision_models and pins in its unit tests).  All pointers but params16 are DEVICE
 * pointers on `device`; the calls are asynchronous on `stream`.  params16: host, the first 16 entries of the params
 * vector (mu, C_Sf, C_Sr, lf, lr, h, m, I, s_min, s_max, sv_min, sv_max, v_switch, a_max, v_min, v_max). ----
 * vehicle_dynamics_st (dynamic_models.py:123-176): x [n][7], u [n][2] (steering velocity, acceleration) -> f [n][7] */
int f110_model_dynamics_st(const double *x_dev, const double *u_dev, const double *params16, int64_t n, double *f_dev,
                           int device, void *stream);
/* vehicle_dynamics_ks (dynamic_models.py:91-121): x [n][5], u [n][2] -> f [n][5] */
int f110_model_dynamics_ks(const double *x_dev, const double *u_dev, const double *params16, int64_t n, double *f_dev,
                           int device, void *stream);
/* pid (dynamic_models.py:178-221): in [n][4] = speed, steer, current_speed, current_steer -> out [n][2] = accl, sv */
int f110_model_pid(const double *in_dev, double max_sv, double max_a, double max_v, double min_v, int64_t n, double *out_dev,
                   int device, void *stream);
/* get_vertices (collision_models.py:238-260): poses [n][3] -> corners [n][4][2] (rl, rr, fr, fl) */
int f110_model_vertices(const double *poses_dev, double length, double width, int64_t n, double *verts_dev, int device,
                        void *stream);
/* collision (GJK, collision_models.py:114-182): quads a, b [n][4][2] -> hit [n] bytes 0/1 */
int f110_model_collision(const double *verts_a_dev, const double *verts_b_dev, int64_t n, uint8_t *hit_dev, int device,
                         void *stream);

/* accl_constraints (dynamic_models.py:30-60): in [n][2] = (vel, accl) -> out [n] */
int f110_model_accl_constraints(const double *in_dev, double v_switch, double a_max, double v_min, double v_max, int64_t n,
                                double *out_dev, int device, void *stream);
/* steering_constraint (dynamic_models.py:63-87): in [n][2] = (steering_angle, steering_velocity) -> out [n] */
int f110_model_steering_constraint(const double *in_dev, double s_min, double s_max, double sv_min, double sv_max, int64_t n,
                                   double *out_dev, int device, void *stream);

/* Trig tables (laser_models.py:379-381, theta_dis entries) and per-beam tables
 * (base_classes.py:120-156, n_beams entries), computed by the host exactly as the reference does. */
int f110_set_tables(f110_sim *sim, const double *sines, const double *cosines, const double *scan_angles,
                    const double *beam_cosines, const double *side_distances, void *stream);

/* Simulator.update_params (base_classes.py:507-527).  agent_idx < 0: all agents;
 * agent_idx == n_agents: the construct-time {width,length} used by check_collision (:542). */
int f110_set_params(f110_sim *sim, const double *params18, int32_t agent_idx, void *stream);

/* F110_NOISE_TABLE: upload rows [first_row, first_row+n_rows) of the host noise stream
 * (n_beams doubles each).  The table must cover max(noise_step)+1 before each step.
 * Rows appended behind the table's end (first_row >= rows uploaded so far) travel on an internal stream and
 * `stream` is made to wait for them: steps already enqueued on `stream` keep running meanwhile.  Rows that
 * overwrite existing ones, and a table that has to grow, synchronise `stream`.  `rows` is free on return. */
int f110_set_noise_rows(f110_sim *sim, const double *rows, int64_t first_row, int64_t n_rows, void *stream);

/* Pool of start poses for auto_reset: [n_poses][n_agents][3] host doubles. */
int f110_set_reset_pool(f110_sim *sim, const double *poses, int32_t n_poses, void *stream);

/* F110Env.reset (f110_env.py:304-347): zero state, set poses, re-seed noise, then one zero-action
 * step.  poses: host [N][A][3]; start_rot: host [N][4] row-major 2x2 rotation by -theta_ego as the
 * reference computes it (:329), or NULL to compute it on the device; mask: host [N] bytes or NULL. */
int f110_reset(f110_sim *sim, const double *poses, const double *start_rot, const uint8_t *mask, void *stream);

/* Simulator.reset (base_classes.py:607-623) on its own: agent states, steering FIFOs, collision flags and the noise
 * stream are reset to `poses` (host [N][A][3]; start_rot, mask: as in f110_reset) -- WITHOUT the zero-action step that
 * F110Env.reset adds (f110_env.py:335-336).  The env-level counters (time, lap toggles) are cleared as well. */
int f110_reset_poses(f110_sim *sim, const double *poses, const double *start_rot, const uint8_t *mask, void *stream);

/* F110Env.step -> Simulator.step (f110_env.py:261-302, base_classes.py:546-605) for all envs.
 * actions_dev: DEVICE [N][A][2] doubles (steer, speed), or NULL to use buffers.actions. */
int f110_step(f110_sim *sim, const double *actions_dev, void *stream);

/* Same step with HOST buffers (what a Python/Go/Java caller of the reference API holds):
 * copies actions in, steps, copies the observation out, and returns when the host buffers are
 * valid.  obs_block_host: [10][N*A] doubles (see f110_buffers.obs_block) or NULL;
 * scans_host: [N][A][B] in the handle's scan type or NULL; done_host: [N] bytes or NULL.
 * Page-locked host buffers give full PCIe rate; pageable ones work.  When all of them are page-locked AND mapped
 * (cudaHostAlloc / torch pin_memory) and the observation is below 256 KB (a handful of envs), the step reads the
 * actions and writes the observation directly over PCIe from its kernels instead of using the copy engines. */
int f110_step_host(f110_sim *sim, const double *actions_host, double *obs_block_host, void *scans_host,
                   uint8_t *done_host, void *stream);

/* ScanSimulator2D.scan(pose, rng=None) (laser_models.py:429-454) for n poses.
 * poses_dev: DEVICE [n][3] doubles; out_dev: DEVICE [n][n_beams] in the handle's scan type. */
int f110_raycast(f110_sim *sim, const double *poses_dev, int32_t n, void *out_dev, void *stream);

int f110_get_buffers(f110_sim *sim, f110_buffers *out);
int f110_get_config(f110_sim *sim, f110_config *out);

/* Debug/measurement: count EDT lookups with one atomic per warp (adds a little time). */
int f110_count_lookups(f110_sim *sim, int32_t enable);
/* Measurement: when enabled, f110_step brackets its ray-cast launch with CUDA events on the caller's
 * stream; f110_last_raycast_ms waits for that launch and returns its device time in milliseconds. */
int f110_enable_timing(f110_sim *sim, int32_t enable);
int f110_last_raycast_ms(f110_sim *sim, float *ms);
/* Measurement: read bandwidth of this GPU's L2 by two microbenchmarks over a table of `table_bytes` that stays L2-resident
 * (L1 bypassed): whole 128-byte lines per warp, and random 4-byte gathers (one 32-byte sector per lane and request: an EDT
 * lookup that misses L1).  out6: stream GB/s, gather sectors/s, ms per stream launch, ms per gather launch, bytes per
 * stream launch, sectors per gather launch.  Synchronous.  (the roof the ray-cast is put against.) */
int f110_l2_probe(int device, int64_t table_bytes, int32_t repeats, double *out6, void *stream);
/* Checked build (libf110_b200_checked.so, compiled with -DF110_CHECKED; compute-sanitizer is closed on the B200 pool): every
 * address / index whose bounds test the fast kernels removed or hoisted is tested, violations are counted per device.
 * out8: counts since process start -- 0 padded-fp32-map address, 1 fp64 map cell (also the renderer's map cells), 2 trig /
 * direction table index (also the planner's way-point indices), 3 opponent queue slot (also the rasteriser's line / edge queue),
 * 4 sort bin / slot, 5 beam index (also the rasteriser's beam picks, pixels and rows), 6 noise row, 7 self-test.  selftest != 0 first commits ONE deliberate violation of
 * kind 7 (counted only by a checked build: proves the checks are compiled in).  f110_build_is_checked(): 1 / 0. */
int f110_build_is_checked(void);
int f110_debug_violations(int device, uint64_t *out8, int32_t selftest);
/* Number of kernel launches issued by this handle since creation. */
int64_t f110_launch_count(f110_sim *sim);
int f110_synchronize(f110_sim *sim, void *stream);

/* ---------------------------------------------------------------------------------------------
 * Scan -> bird's-eye bitmap.  Replaces weap_util.lidar.lidar_to_bitmap
 * (weap_util/weap_util/lidar.py:114-154 and its OpenCV calls) for a batch of scans on the device;
 * output bytes are identical to the reference's single-channel image.
 * ------------------------------------------------------------------------------------------- */
#define F110_BITMAP_FILL 0
#define F110_BITMAP_POLYGON 1
#define F110_BITMAP_RAYS 2

typedef struct f110_bitmap_config {
    int32_t abi_version;       /* F110_ABI_VERSION */
    int32_t n_beams;           /* len(scan) */
    int32_t target_beam_count; /* lidar.py:13 (600); 0 < target_beam_count < n_beams (:50) */
    int32_t height, width;     /* output_image_dims (:12) */
    int32_t draw_mode;         /* F110_BITMAP_* (:15) */
    int32_t draw_center;       /* :11 */
    int32_t bg_white;          /* bg_color == 'white' (:10): background 255, drawing 0; else 0 / 255 */
    int32_t scan_is_f32;       /* scans are float (fp32 fast path) or double */
    double scaling_factor;     /* resolved by the caller: min(dims) / max_scan_radius when that is given (:53-54) */
} f110_bitmap_config;

typedef struct f110_bitmap f110_bitmap;

/* indices: host [target_beam_count] int32 = np.linspace(0, n_beams - 1, T, dtype=int) (:66);
 * cos_a / sin_a: host [T] doubles = np.cos / np.sin(starting_angle + dir * fov * np.linspace(0, 1, T)) (:70,:75-76),
 * computed by the host with numpy exactly as the reference does. */
int f110_bitmap_create(const f110_bitmap_config *cfg, const int32_t *indices, const double *cos_a, const double *sin_a,
                       int device, f110_bitmap **out);
int f110_bitmap_destroy(f110_bitmap *bm);
/* scans_dev: DEVICE [n][n_beams]; out_dev: DEVICE [n][height][width] bytes, or, with packed != 0,
 * [n][height][ceil(width / 32)] 32-bit words, bit b of word k = pixel 32 k + b, set = drawn colour. */
int f110_bitmap_render(f110_bitmap *bm, const void *scans_dev, int64_t n, void *out_dev, int32_t packed, void *stream);

/* ---------------------------------------------------------------------------------------------
 * Batched pure-pursuit planner.  Replaces PurePursuitPlanner.plan and its njit helpers
 * (examples/waypoint_follow.py:15-141,183-217) for n cars at once, on the device.
 * ------------------------------------------------------------------------------------------- */
typedef struct f110_pursuit f110_pursuit;

/* wx, wy, wv: host [n_waypoints] = columns wpt_xind, wpt_yind, wpt_vind of the raceline file
 * (waypoint_follow.py:160-164, config_example_map.yaml); max_reacquire = 20.0 (:153). */
int f110_pursuit_create(const double *wx, const double *wy, const double *wv, int32_t n_waypoints, double wheelbase,
                        double max_reacquire, int device, f110_pursuit **out);
int f110_pursuit_destroy(f110_pursuit *pp);
/* plan(pose_x, pose_y, pose_theta, lookahead_distance, vgain) (:203-217) for n cars.
 * x_dev / y_dev / theta_dev: DEVICE, element q at [q * stride] (stride 1 for the rows of f110_buffers.state);
 * lookahead_dev / vgain_dev: DEVICE [n] per-car values or NULL to use the scalars;
 * actions_dev: DEVICE [n][2] = (steer, speed), the layout f110_step takes;
 * picked_dev: DEVICE [n][2] int32 (nearest segment, chosen waypoint index or -2) or NULL. */
int f110_pursuit_plan(f110_pursuit *pp, const double *x_dev, const double *y_dev, const double *theta_dev, int64_t n,
                      int64_t stride, double lookahead, double vgain, const double *lookahead_dev, const double *vgain_dev,
                      double *actions_dev, int32_t *picked_dev, void *stream);

/* ---------------------------------------------------------------------------------------------
 * Headless renderer.  F110Env.render (f110_env.py:385-416) draws EnvRenderer's scene
 * (rendering.py) through pyglet/OpenGL; here the same scene is rasterised into an RGB8 device buffer
 * (mode 'rgb_array').  No window, no text labels.
 * ------------------------------------------------------------------------------------------- */
typedef struct f110_render_config {
    int32_t abi_version;
    int32_t width, height;     /* WINDOW_W, WINDOW_H (f110_env.py:50-51) */
    int32_t n_agents, ego_idx;
    int32_t map_h, map_w;
    double resolution, origin_x, origin_y;      /* map yaml (rendering.py:123-129) */
    double follow_margin;      /* >= 0: camera follows the ego car, bounding box +- margin (examples/waypoint_follow.py:259-269,
                                  800 there); < 0: the fixed camera below */
    double left, right, bottom, top;            /* glOrtho bounds in 50 x world units (rendering.py:74-80,290) */
} f110_render_config;

typedef struct f110_render f110_render;

/* obstacle: host [map_h][map_w] bytes, non-zero where the (top-bottom flipped) map image is exactly 0 (rendering.py:150) */
int f110_render_create(const f110_render_config *cfg, const uint8_t *obstacle, int device, f110_render **out);
int f110_render_destroy(f110_render *r);
/* x_dev / y_dev / theta_dev: DEVICE [n_envs][n_agents] poses (rows 0, 1, 4 of f110_buffers.state);
 * out_dev: DEVICE [n_envs][height][width][3] bytes, row 0 = top of the window. */
int f110_render_frame(f110_render *r, const double *x_dev, const double *y_dev, const double *theta_dev, int64_t n_envs,
                      uint8_t *out_dev, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* F110_B200_H */
