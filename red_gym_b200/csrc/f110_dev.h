// Internal device-side view of one simulator handle (passed to kernels by value).
// HBM layout: every per-agent quantity is SoA over ia = env * n_agents + agent so that a warp of
// consecutive agents reads/writes consecutive addresses; scans are [ia][beam].
#pragma once
#include <stdint.h>

enum { F110_P_MU, F110_P_CSF, F110_P_CSR, F110_P_LF, F110_P_LR, F110_P_H, F110_P_M, F110_P_I, F110_P_SMIN,
       F110_P_SMAX, F110_P_SVMIN, F110_P_SVMAX, F110_P_VSWITCH, F110_P_AMAX, F110_P_VMIN, F110_P_VMAX,
       F110_P_WIDTH, F110_P_LENGTH, F110_NP };

// Launches of fewer scans than F110_SPLIT_BELOW* share a scan between the warps of a CTA (k_*split / SPLIT = true) instead of
// giving it one warp: one warp per scan would leave most of the GPU waiting on ~365 dependent gathers in a row.  The warps
// claim rounds of 32 beams from a shared counter.  A scan gets 8, 4 or 2 warps (f110_split_wps: the most that still make about
// one wave of CTAs); the CTA holds 8 / that many scans.  Measured per step, split / one warp per scan: fp64 1184 envs
// 0.082 / 0.148 ms, 2368: 0.085 / 0.160, 4096: 0.121 / 0.172, 8192: 0.168 / 0.189, 16384: 0.280 / 0.292, 24576: 0.384 / 0.387;
// fp32 1184 envs 0.068 / 0.104, 2368: 0.084 / 0.113, 3072: 0.095 / 0.115, 4096: 0.115 / 0.119, 5120: 0.125 / 0.122.
// (With 8 warps per scan at every size: fp64 4096 envs 0.127, fp32 worse than one warp per scan from 4096 on.)
// Variant bit 7 turns the split off, bit 11 keeps 8 warps per scan, bit 12 splits at any size.
// Split launches map CTA b to group (b % n_sm) * ceil(n / n_sm) + b / n_sm: CTAs are handed to the SMs round-robin,
// so the CTAs resident on one SM then hold consecutive scans (neighbours on the track, overlapping EDT footprints)
// instead of scans n_sm apart.  The grid is n_sm * ceil(n / n_sm) CTAs; the surplus ones exit.
static __host__ __device__ __forceinline__ int f110_split_grid(int n, int n_sm) { return n_sm * ((n + n_sm - 1) / n_sm); }
static __device__ __forceinline__ int f110_split_scan(int b, int n, int n_sm) { return (b % n_sm) * ((n + n_sm - 1) / n_sm) + b / n_sm; }
// Launches of at least this many scans are put in spatial order first (f110_order.cu: three small kernels, ~17-20 us with
// their launch gaps at these sizes).  Round 2, steady state (env order decorrelated from track order), step time without / with
// the sort: fp32 8192 envs 0.148 / 0.155 ms, 16384: 0.224 / 0.209, 24576: 0.305 / 0.262; fp64 8192: 0.205 / 0.220,
// 16384: 0.328 / 0.337, 24576: 0.489 / 0.453 (profiles/r2e_sort_threshold_ab.json).  Hence per precision; variant bit 15 sorts
// at any size, bit 6 never.  f110_raycast (stand-alone poses) keeps the round-1 threshold.
#define F110_ORDER_FROM32 16384
#define F110_ORDER_FROM64 20480
#define F110_ORDER_FROM 32768
// warps per scan of a split launch of n scans: 8, 4 or 2, the largest that still makes about one wave of CTAs
// (8 CTAs of 8 warps per SM), see k_raycast64
static __host__ __device__ __forceinline__ int f110_split_wps(int n, int n_sm)
{
    const long long slots = 64LL * n_sm;          // warps in flight
    return (long long)n * 8 <= slots ? 8 : (long long)n * 4 <= slots ? 4 : 2;
}
#define F110_SPLIT_BELOW64 20480
#define F110_SPLIT_BELOW32 4608

// Checked build (-DF110_CHECKED -> libf110_b200_checked.so, tests/test_gpu_checked_build.py): compute-sanitizer is closed on
// the B200 pool, so the code whose speed comes from REMOVED bounds tests (padded fp32 map addressed through a biased base
// pointer and the mantissa floor; the hoisted map test of the fp64 path; table, queue, bin and beam indices) carries its own:
// every such address / index is tested and a violation counts into a per-device array that the test reads back
// (f110_debug_violations).  The normal build compiles the macro to nothing.
enum { F110_V_EDT32 = 0, F110_V_EDT64 = 1, F110_V_TABLE = 2, F110_V_QUEUE = 3, F110_V_ORDER = 4, F110_V_BEAM = 5, F110_V_NOISE = 6, F110_V_SELFTEST = 7 };
#ifdef F110_CHECKED
#define F110_CHECK(viol, cond, code) do { if (!(cond)) atomicAdd((viol) + (code), 1ull); } while (0)
#else
#define F110_CHECK(viol, cond, code) do { } while (0)
#endif

struct F110Dev {
    int n_envs, n_agents, n_beams, integrator, ego_idx, theta_dis;
    int noise_mode, auto_reset, count_lookups, n_pool, scan_is_f32, env_offset;
    int n_sm;                          // SMs of the device (split-scan CTA -> scan mapping)
    unsigned theta_band;               // fp64 beam index: replay the reference's running sum within this many 2^-30 of an integer
    int scan_u16, no_steal;            // no_steal (variant bit 9): warps of a split CTA stay on their own scan;  F110_SCAN_U16: scans stored as uint16 at 0.5 mm per LSB (fp32 path, one agent per env)
    double timestep, fov, eps, max_range, theta_inc, ttc_thresh, noise_std;
    unsigned long long seed;
    unsigned long long inc_fix;        // theta_inc in 32.32 fixed point; kstep = 32 * inc_fix (one round of 32 beams)
    unsigned kstep_lo, kstep_hi;
    float max_c32, res32, noise_std32, ttc_thresh32, side_max32, sent32;   // fp32 path constants (cells / metres)
    float noise_k32, pad2_;            // -2 ln 2 * noise_std^2 (Box-Muller radius scaled by sigma, see box_muller)
    // map
    int map_h, map_w, res_pow2, has_rot;
    double res, inv_res, ox, oy, oc, os, xmax, ymax;
    double oob64;                      // dt[-1,-1] : value an out-of-map lookup returns
    double magic64;                    // res_pow2: 2^52 * res, so that v + magic64 (round toward zero) holds floor(v / res) in its low word
    float oob32;
    const double *edt64;               // row-major [h][w], metres
    const float *edt32;                // row-major [h][w], in CELL units (metres / res); obstacle cells hold sent32 (see fp32 kernels)
    const float *edt32p;               // edt32 with a border of `pad` cells on every side holding oob32: [h + 2 pad][wp]
    unsigned long long edt32p_adj;     // byte address of edt32p minus 4 * K, K = what the float exponent bits contribute (scan32p)
    float magic_x;                     // 2^23 (+ 2^21 when K would wrap the 32-bit index)
    int pad, wp;                       // pad >= max_range / res + 2, wp = w + 2 pad: a ray that starts inside the map never leaves it
    // tables
    const double *sines, *cosines;                  // theta_dis
    const double *scan_angles, *beam_cos, *side_dist;  // n_beams
    const float4 *dir32;                            // theta_dis: (dx, dy, 1/dx, 1/dy), unit ray direction in the map frame
    const float2 *dir2;                             // theta_dis: (dx, dy) only (padded-map kernel)
    const float *beam_cos32, *side_dist32, *ttc_k32; // n_beams; ttc_k = 1.2 * ttc_thresh * |cos|
    const double *params;                           // (n_agents + 1) * F110_NP
    // per (env, agent)
    double *state;       // [7][NA]
    double *pose_pre;    // [3][NA] pose after dynamics, before iTTC zeroing (Simulator.agent_poses)
    double *steer_buf;   // [2][NA] 0 = newest
    uint8_t *steer_len;  // NA
    uint8_t *in_collision;
    void *scans;
    double *collisions, *collision_idx, *lap_times, *lap_counts;
    int *toggles;
    uint8_t *near_start;
    uint8_t *checkpoint_done;
    double *start_x, *start_y;
    // per env
    double *start_rot;   // [N][4]
    double *cur_time;
    long long *noise_step;
    uint8_t *done;
    uint8_t *reset_pending;
    long long *episode_count;
    const double *actions;
    const double *noise;     // [noise_rows][n_beams]
    long long noise_rows;
    const double *pool;      // [n_pool][A][3]
    unsigned long long *lookups;
    unsigned long long *violations;   // [8] per device, F110_V_* (checked build)
    const uint8_t *step_mask;   // [N] or null: envs that take part in this step (masked reset)
    // spatial launch order of the scans (f110_order.cu); order == null: warp w works on agent w
    int *order, *order_key, *order_rank, *order_hist, *order_base;
    float4 *rec32, *rec32_tmp;          // fp32 path: 32-byte scan records prepared by the sort, in launch order / in agent order (f110_scan32_dev.h)
    int order_bins, order_tx, order_ty, order_chunk_envs;
    float order_inv_tile;
};

// launch helpers implemented in the kernel translation units
void f110_launch_dynamics(const F110Dev &d, cudaStream_t st);
void f110_launch_raycast64(const F110Dev &d, int env_begin, int env_count, int variant, cudaStream_t st);
void f110_launch_raycast32(const F110Dev &d, int env_begin, int env_count, int variant, cudaStream_t st);
void f110_launch_agents(const F110Dev &d, cudaStream_t st);      // GJK + opponent ray-cast (A >= 2)
void f110_launch_finalize(const F110Dev &d, cudaStream_t st);    // collisions, lap bookkeeping, done
// the whole tick in one kernel (fp64 path, one agent per env, a handful of envs); host pointers: mapped host buffers or null
void f110_launch_step_fused64(const F110Dev &d, double *obs_host, double *scans_host, uint8_t *done_host, const double *cmd_host, int variant,
                              cudaStream_t st);
// launches of at most this many single-agent fp64 envs take the fused kernel (a CTA per env, about one wave at 128 registers)
#define F110_FUSED_MAX_ENVS 296
// ... and for envs of 2 .. 8 agents: a CTA of 32 warps per env, one wave (k_step_fused_multi64)
#define F110_FUSED_MULTI_MAX_AGENTS 8
#define F110_FUSED_MULTI_MAX_ENVS 148
// order: null or a permutation of [0, n): warp w scans pose order[w]
void f110_launch_scan_only64(const F110Dev &d, const double *poses, int n, double *out, const int *order, int variant, cudaStream_t st);
void f110_launch_scan_only32(const F110Dev &d, const double *poses, int n, float *out, const int *order, int variant, cudaStream_t st);
void f110_launch_reset(const F110Dev &d, const double *poses_dev, const double *rot_dev, const uint8_t *mask_dev, cudaStream_t st);
void f110_launch_edt32(const double *src, float *dst, size_t cells, double inv_res, double eps, float cap, float sent, cudaStream_t st);
void f110_launch_order(const F110Dev &d, cudaStream_t st);      // (chunk, tile, heading) counting sort -> d.order
void f110_launch_order_poses(const F110Dev &d, const double *poses, int n, int *buf, cudaStream_t st);
void f110_launch_pad_edt32(const float *src, float *dst, int h, int w, int pad, float oob, cudaStream_t st);
int f110_probe_l2(int n_sm, size_t table_bytes, int repeats, double *out, cudaStream_t st);      // f110_probe.cu
void f110_launch_check_selftest(unsigned long long *viol, cudaStream_t st);      // one deliberate violation (F110_V_SELFTEST) in a checked build

// model functions on their own (f110_kernels_fp64.cu, end of file)
void f110_launch_model_dynamics(const double *x, const double *u, const double *p16, long long n, int ks, double *f, cudaStream_t st);
void f110_launch_model_pid(const double *in, double max_sv, double max_a, double max_v, double min_v, long long n, double *out, cudaStream_t st);
void f110_launch_model_constraint(const double *in, double a, double b, double c, double e_, long long n, int steering, double *out, cudaStream_t st);
void f110_launch_model_vertices(const double *poses, double length, double width, long long n, double *verts, cudaStream_t st);
void f110_launch_model_collision(const double *a, const double *b, long long n, uint8_t *hit, cudaStream_t st);
