// C ABI of libf110_b200.so (see include/f110_b200.h).  Owns device memory, stages host data,
// orders the kernels of one step on the caller's stream.  No CPU fallback anywhere.
#include <cuda_runtime.h>
#include <math.h>
#include <stdarg.h>
#include <stdio.h>
#include <string.h>
#include <string>
#include <vector>
#include "../../include/f110_b200.h"
#include "f110_dev.h"
#include "f110_bitmap_dev.h"
#include "f110_pursuit_dev.h"
#include "f110_render_dev.h"

static thread_local std::string g_err;
static int fail(int code, const char *fmt, ...)
{
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    g_err = buf;
    return code;
}
#define CK(call)                                                                                   \
    do {                                                                                           \
        cudaError_t e_ = (call);                                                                   \
        if (e_ != cudaSuccess) return fail(F110_ERR_CUDA, "%s failed: %s", #call, cudaGetErrorString(e_)); \
    } while (0)

struct f110_sim {
    f110_config cfg;
    int device;
    F110Dev d;
    std::vector<void *> allocs;
    bool has_map = false, has_tables = false;
    double *edt64 = nullptr;
    float *edt32 = nullptr, *edt32p = nullptr;
    int *order_hist = nullptr, *order_base = nullptr;
    int *order_buf = nullptr;      // order | key | rank, NA each
    int *rc_order_buf = nullptr;   // the same for f110_raycast, grown on demand
    size_t rc_order_cap = 0;
    bool use_order = false;
    float2 *dir2 = nullptr;
    float4 *dir32 = nullptr;
    double *noise = nullptr;
    long long noise_cap = 0;
    double *pool = nullptr;
    double *obs_block = nullptr;
    double *actions = nullptr;
    double *stage_poses = nullptr, *stage_rot = nullptr;
    uint8_t *stage_mask = nullptr;
    cudaStream_t copy_stream = nullptr;
    cudaEvent_t noise_ev = nullptr;      // end of the last noise-row upload on copy_stream (f110_set_noise_rows)
    cudaEvent_t ev_a = nullptr, ev_b = nullptr;
    std::vector<cudaEvent_t> chunk_ev;
    long long launches = 0;
    bool timing = false;
    bool timed = false;            // ev_a / ev_b bracket the ray-cast of the last step
};

// Per-device violation counters of the checked build (F110_V_*); allocated on first use, shared by every handle of the
// process, never freed.  The normal build allocates them too (the kernels never write them) so that the ABI is the same.
static unsigned long long *g_violations[64];
static unsigned long long *violations_for(int device)
{
    if (device < 0 || device >= 64) return nullptr;
    if (!g_violations[device]) {
        void *q = nullptr;
        if (cudaMalloc(&q, 8 * sizeof(unsigned long long)) != cudaSuccess) { cudaGetLastError(); return nullptr; }
        cudaMemset(q, 0, 8 * sizeof(unsigned long long));
        g_violations[device] = (unsigned long long *)q;
    }
    return g_violations[device];
}

template <typename T>
static int dev_alloc(f110_sim *s, T **p, size_t n, bool zero = true)
{
    void *q = nullptr;
    cudaError_t e = cudaMalloc(&q, n * sizeof(T) > 0 ? n * sizeof(T) : 16);
    if (e != cudaSuccess) return fail(F110_ERR_NOMEM, "cudaMalloc(%zu bytes) failed: %s", n * sizeof(T), cudaGetErrorString(e));
    if (zero) cudaMemset(q, 0, n * sizeof(T));
    s->allocs.push_back(q);
    *p = (T *)q;
    return 0;
}
#define TRY(x) do { int r_ = (x); if (r_) return r_; } while (0)

// per-angle ray direction in the map frame and its reciprocals (1/0 stored as -inf, see trace32)
static void build_dir_table(const double *sines, const double *cosines, int T, double oc, double os, std::vector<float4> &dir,
                            std::vector<float2> &dir2)
{
    dir.resize(T);
    dir2.resize(T);
    for (int i = 0; i < T; i++) {
        float dx = (float)(cosines[i] * oc + sines[i] * os), dy = (float)(-cosines[i] * os + sines[i] * oc);
        float ix = dx != 0.f ? 1.0f / dx : -INFINITY, iy = dy != 0.f ? 1.0f / dy : -INFINITY;
        dir[i] = make_float4(dx, dy, ix, iy);
        dir2[i] = make_float2(dx, dy);
    }
}

extern "C" const char *f110_last_error(void) { return g_err.c_str(); }
extern "C" int f110_abi_version(void) { return F110_ABI_VERSION; }
extern "C" int f110_device_count(void)
{
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

extern "C" int f110_create(const f110_config *cfg, int device, f110_sim **out)
{
    if (!cfg || !out) return fail(F110_ERR_INVALID, "null argument");
    if (cfg->abi_version != F110_ABI_VERSION) return fail(F110_ERR_INVALID, "abi_version %d != %d", cfg->abi_version, F110_ABI_VERSION);
    if (cfg->n_envs < 1 || cfg->n_agents < 1 || cfg->n_agents > F110_MAX_AGENTS || cfg->n_beams < 2 || cfg->theta_dis < 1)
        return fail(F110_ERR_INVALID, "bad sizes: n_envs=%d n_agents=%d n_beams=%d", cfg->n_envs, cfg->n_agents, cfg->n_beams);
    if (cfg->n_beams > 65535) return fail(F110_ERR_INVALID, "n_beams=%d: at most 65535 beams per scan", cfg->n_beams);
    if (cfg->integrator != F110_INTEGRATOR_RK4 && cfg->integrator != F110_INTEGRATOR_EULER)
        return fail(F110_ERR_INVALID, "Invalid Integrator Specified (%d)", cfg->integrator);
    if (cfg->ego_idx < 0 || cfg->ego_idx >= cfg->n_agents) return fail(F110_ERR_INVALID, "ego_idx out of range");
    if (cfg->precision == F110_PRECISION_FP32 && cfg->noise_mode == F110_NOISE_TABLE)
        return fail(F110_ERR_INVALID, "the fp32 path takes F110_NOISE_OFF or F110_NOISE_PHILOX");
    if (cfg->precision == F110_PRECISION_FP64 && cfg->noise_mode == F110_NOISE_PHILOX)
        return fail(F110_ERR_INVALID, "the fp64 path takes F110_NOISE_OFF or F110_NOISE_TABLE");
    if (cfg->scan_format != F110_SCAN_NATIVE && cfg->scan_format != F110_SCAN_U16) return fail(F110_ERR_INVALID, "bad scan_format %d", cfg->scan_format);
    if (cfg->scan_format == F110_SCAN_U16 && (cfg->precision != F110_PRECISION_FP32 || cfg->n_agents != 1 || (cfg->n_beams & 1) ||
                                              !(cfg->max_range * F110_SCAN_U16_PER_METRE < 65000.0)))
        return fail(F110_ERR_INVALID, "F110_SCAN_U16 needs the fp32 path, one agent per env, an even beam count and max_range < 32.5 m");
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0) {
        cudaGetLastError();
        return fail(F110_ERR_CUDA, "no CUDA device available (%s); this library has no CPU path", e != cudaSuccess ? cudaGetErrorString(e) : "0 devices");
    }
    if (device < 0 || device >= ndev) return fail(F110_ERR_INVALID, "device %d out of range (%d devices)", device, ndev);
    CK(cudaSetDevice(device));
    f110_sim *s = new f110_sim();
    s->cfg = *cfg;
    s->device = device;
    F110Dev &d = s->d;
    memset(&d, 0, sizeof(d));
    const size_t N = cfg->n_envs, A = cfg->n_agents, NA = N * A, B = cfg->n_beams;
    d.n_envs = cfg->n_envs; d.n_agents = cfg->n_agents; d.n_beams = cfg->n_beams; d.integrator = cfg->integrator;
    d.n_sm = 148;
    cudaDeviceGetAttribute(&d.n_sm, cudaDevAttrMultiProcessorCount, device);
    if (d.n_sm < 1) d.n_sm = 1;
    d.ego_idx = cfg->ego_idx; d.theta_dis = cfg->theta_dis; d.noise_mode = cfg->noise_mode; d.auto_reset = cfg->auto_reset;
    d.scan_is_f32 = cfg->precision == F110_PRECISION_FP32;
    d.scan_u16 = cfg->scan_format == F110_SCAN_U16;
    d.no_steal = (cfg->raycast_variant & 512) != 0;
    d.env_offset = cfg->env_offset;
    d.timestep = cfg->timestep; d.fov = cfg->fov; d.eps = cfg->eps; d.max_range = cfg->max_range;
    d.ttc_thresh = cfg->ttc_thresh; d.noise_std = cfg->noise_std; d.seed = cfg->seed;
    d.noise_std32 = (float)cfg->noise_std; d.ttc_thresh32 = (float)cfg->ttc_thresh;
    d.noise_k32 = (float)(-2.0 * M_LN2 * cfg->noise_std * cfg->noise_std);
    if (cfg->noise_mode == F110_NOISE_PHILOX && !(cfg->noise_std > 1e-12)) d.noise_mode = F110_NOISE_OFF;   // sigma 0: nothing to add
    // laser_models.py:367-368
    double angle_increment = cfg->fov / (cfg->n_beams - 1);
    d.theta_inc = cfg->theta_dis * angle_increment / (2. * M_PI);
    d.inc_fix = (unsigned long long)(d.theta_inc * 4294967296.0 + 0.5);
    {   // a round of 32 beams advances the table index by 32 * inc, taken modulo the table: the kernels wrap a sum of two
        // indices below theta_dis with ONE conditional subtraction, and with a handful of beams over a wide field of view
        // (fewer than 32 fov / 2 pi + 1 of them) 32 * inc is more than a turn of the table -- the index fetched ahead for the
        // next round then lay beyond the table (found by the checked build, tests/test_gpu_edge_cases.py with 2 and 3 beams)
        const unsigned long long kwrap = (unsigned long long)(unsigned)cfg->theta_dis << 32;
        const unsigned long long kstep = (32ull * d.inc_fix) % kwrap;
        d.kstep_lo = (unsigned)kstep; d.kstep_hi = (unsigned)(kstep >> 32);
    }
    // fp64 beam index (theta_index_of): the reference's running sum drifts from the closed form by at most half an ulp of
    // theta_dis per addition (1.14e-13 at 2000), the closed form itself by an ulp per round of 32 beams; the band within
    // which the running sum is replayed exactly is that bound in units of 2^-30, plus 2 (3 units = 2.8e-9 up to 2160 beams)
    {
        const double ulp = nextafter((double)cfg->theta_dis, 1e300) - (double)cfg->theta_dis;
        const double drift = 0.5 * ulp * cfg->n_beams + ulp * ((cfg->n_beams + 31) / 32);
        d.theta_band = 2u + (unsigned)ceil(drift * 1073741824.0);
    }
    int rc = 0;
    do {
        // obs block: 7 state fields + collisions + lap_times + lap_counts, contiguous for a single D2H copy
        if ((rc = dev_alloc(s, &s->obs_block, 10 * NA))) break;
        d.state = s->obs_block;
        d.collisions = s->obs_block + 7 * NA;
        d.lap_times = s->obs_block + 8 * NA;
        d.lap_counts = s->obs_block + 9 * NA;
        if ((rc = dev_alloc(s, &d.pose_pre, 3 * NA))) break;
        if ((rc = dev_alloc(s, &d.steer_buf, 2 * NA))) break;
        if ((rc = dev_alloc(s, &d.steer_len, NA))) break;
        if ((rc = dev_alloc(s, &d.in_collision, NA))) break;
        if ((rc = dev_alloc(s, &d.collision_idx, NA))) break;
        if ((rc = dev_alloc(s, &d.toggles, NA))) break;
        if ((rc = dev_alloc(s, &d.near_start, NA))) break;
        if ((rc = dev_alloc(s, &d.checkpoint_done, NA))) break;
        if ((rc = dev_alloc(s, &d.start_x, NA))) break;
        if ((rc = dev_alloc(s, &d.start_y, NA))) break;
        if ((rc = dev_alloc(s, &d.start_rot, 4 * N))) break;
        if ((rc = dev_alloc(s, &d.cur_time, N))) break;
        if ((rc = dev_alloc(s, &d.noise_step, N))) break;
        if ((rc = dev_alloc(s, &d.done, N))) break;
        if ((rc = dev_alloc(s, &d.reset_pending, N))) break;
        if ((rc = dev_alloc(s, &d.episode_count, N))) break;
        if ((rc = dev_alloc(s, &s->actions, 2 * NA))) break;
        if ((rc = dev_alloc(s, &d.lookups, 1))) break;
        if ((rc = dev_alloc(s, &s->stage_poses, 3 * NA))) break;
        if ((rc = dev_alloc(s, &s->stage_rot, 4 * N))) break;
        if ((rc = dev_alloc(s, &s->stage_mask, N))) break;
        size_t scan_bytes = NA * B * (d.scan_u16 ? 2 : d.scan_is_f32 ? 4 : 8);
        uint8_t *scans = nullptr;
        if ((rc = dev_alloc(s, &scans, scan_bytes, false))) break;
        d.scans = scans;
        double *tabs = nullptr;
        if ((rc = dev_alloc(s, &tabs, 2 * (size_t)cfg->theta_dis + 3 * B))) break;
        d.sines = tabs; d.cosines = tabs + cfg->theta_dis;
        d.scan_angles = tabs + 2 * cfg->theta_dis; d.beam_cos = d.scan_angles + B; d.side_dist = d.beam_cos + B;
        float *tabs32 = nullptr;
        if ((rc = dev_alloc(s, &tabs32, 3 * B))) break;
        d.beam_cos32 = tabs32; d.side_dist32 = tabs32 + B; d.ttc_k32 = tabs32 + 2 * B;
        if ((rc = dev_alloc(s, &s->dir32, (size_t)cfg->theta_dis))) break;
        d.dir32 = s->dir32;
        if ((rc = dev_alloc(s, &s->dir2, (size_t)cfg->theta_dis))) break;
        d.dir2 = s->dir2;
        double *params = nullptr;
        if ((rc = dev_alloc(s, &params, (A + 1) * F110_NP))) break;
        d.params = params;
    } while (0);
    if (rc) { f110_destroy(s); return rc; }
    // start_rot = identity, collision_idx = -1, near_start = 1 (f110_env.py:182-187, base_classes.py:481)
    {
        std::vector<double> rot(4 * N, 0.0), ci(NA, -1.0);
        for (size_t e2 = 0; e2 < N; e2++) { rot[4 * e2] = 1.0; rot[4 * e2 + 3] = 1.0; }
        cudaMemcpy(d.start_rot, rot.data(), rot.size() * 8, cudaMemcpyHostToDevice);
        cudaMemcpy(d.collision_idx, ci.data(), ci.size() * 8, cudaMemcpyHostToDevice);
        cudaMemset(d.near_start, 1, NA);
    }
    d.actions = s->actions;
    d.violations = violations_for(device);
    if (!d.violations) { f110_destroy(s); return fail(F110_ERR_NOMEM, "cudaMalloc failed"); }
    // spatial launch order: pays once more scans are in flight than the GPU holds at a time (and costs 3 small launches)
    s->use_order = (NA >= (size_t)(d.scan_is_f32 ? F110_ORDER_FROM32 : F110_ORDER_FROM64) || (cfg->raycast_variant & 32768)) &&
                   !(cfg->raycast_variant & 64);   // bit 15: sort at any size, bit 6: never
    if (s->use_order) {
        if (dev_alloc(s, &s->order_buf, 3 * NA)) { f110_destroy(s); return F110_ERR_NOMEM; }
        // fp32 path: the sort also prepares the scans (32-byte records; variant bit 14 keeps the preparation in the ray-cast)
        if (d.scan_is_f32 && !(cfg->raycast_variant & 16384)) {
            if (dev_alloc(s, &d.rec32, 2 * NA, false) || dev_alloc(s, &d.rec32_tmp, 2 * NA, false)) { f110_destroy(s); return F110_ERR_NOMEM; }
        }
        d.order_chunk_envs = (A == 1 && N >= 8192) ? (int)((N + 7) / 8) : (int)N;   // same chunks as f110_step_host
    }
    cudaStreamCreateWithFlags(&s->copy_stream, cudaStreamNonBlocking);
    cudaEventCreate(&s->ev_a);
    cudaEventCreate(&s->ev_b);
    e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { f110_destroy(s); return fail(F110_ERR_CUDA, "init failed: %s", cudaGetErrorString(e)); }
    *out = s;
    return F110_OK;
}

extern "C" int f110_destroy(f110_sim *s)
{
    if (!s) return F110_OK;
    cudaSetDevice(s->device);
    cudaDeviceSynchronize();
    for (void *p : s->allocs) cudaFree(p);
    if (s->edt64) cudaFree(s->edt64);
    if (s->edt32) cudaFree(s->edt32);
    if (s->edt32p) cudaFree(s->edt32p);
    if (s->order_hist) cudaFree(s->order_hist);
    if (s->order_base) cudaFree(s->order_base);
    if (s->rc_order_buf) cudaFree(s->rc_order_buf);
    if (s->noise) cudaFree(s->noise);
    if (s->pool) cudaFree(s->pool);
    for (cudaEvent_t ev : s->chunk_ev) cudaEventDestroy(ev);
    if (s->noise_ev) cudaEventDestroy(s->noise_ev);
    if (s->copy_stream) cudaStreamDestroy(s->copy_stream);
    if (s->ev_a) cudaEventDestroy(s->ev_a);
    if (s->ev_b) cudaEventDestroy(s->ev_b);
    delete s;
    return F110_OK;
}

void f110_launch_edt_build(const uint8_t *free_px, unsigned *scratch_g, double *edt, int h, int w, double res, cudaStream_t st);

// Everything a map owns on the device.  set_map_impl builds a new one completely and only then swaps it in: a failed
// update_map (bad size, out of memory, ...) leaves the handle with its previous map, never with freed pointers.
struct MapBufs {
    double *edt64 = nullptr;
    float *edt32 = nullptr, *edt32p = nullptr;
    int *order_hist = nullptr, *order_base = nullptr;
    void release()
    {
        cudaFree(edt64); cudaFree(edt32); cudaFree(edt32p); cudaFree(order_hist); cudaFree(order_base);
        *this = MapBufs();
    }
};

static int build_map(f110_sim *s, MapBufs &m, F110Dev &d, const double *edt, const uint8_t *free_px, int32_t h, int32_t w, double res,
                     double ox, double oy, double oc, double os, cudaStream_t st, int &launches)
{
    const size_t cells = (size_t)h * w;
    CK(cudaMalloc(&m.edt64, cells * 8));
    CK(cudaMalloc(&m.edt32, cells * 4));
    double last_cell;
    if (edt) {
        CK(cudaMemcpyAsync(m.edt64, edt, cells * 8, cudaMemcpyHostToDevice, st));
        last_cell = edt[cells - 1];
    } else {
        uint8_t *fp = nullptr;
        unsigned *g = nullptr;
        CK(cudaMalloc(&fp, cells));
        if (cudaMalloc(&g, cells * 4) != cudaSuccess) { cudaFree(fp); return fail(F110_ERR_NOMEM, "cudaMalloc failed"); }
        cudaMemcpyAsync(fp, free_px, cells, cudaMemcpyHostToDevice, st);
        f110_launch_edt_build(fp, g, m.edt64, h, w, res, st);
        launches += 2;
        cudaError_t e1 = cudaMemcpyAsync(&last_cell, m.edt64 + cells - 1, 8, cudaMemcpyDeviceToHost, st);
        cudaError_t e2 = cudaStreamSynchronize(st);
        cudaFree(fp); cudaFree(g);
        if (e1 != cudaSuccess || e2 != cudaSuccess) return fail(F110_ERR_CUDA, "device EDT failed: %s", cudaGetErrorString(e2 != cudaSuccess ? e2 : e1));
    }
    d.map_h = h; d.map_w = w; d.res = res; d.inv_res = 1.0 / res; d.ox = ox; d.oy = oy; d.oc = oc; d.os = os;
    d.xmax = w * res; d.ymax = h * res;   // width * resolution, height * resolution (laser_models.py:79)
    int ex;
    d.res_pow2 = (frexp(res, &ex) == 0.5);
    d.has_rot = !(oc == 1.0 && os == 0.0);
    d.magic64 = 4503599627370496.0 * res;
    d.max_c32 = (float)(s->cfg.max_range / res); d.res32 = (float)res;
    d.oob64 = last_cell;                  // dt[-1, -1]
    // fp32 encoding: free cells min(dist, cap), obstacle cells `sent`, max_range < cap < sent (all in cells)
    const float cap = floorf(d.max_c32) + 2.0f;
    d.sent32 = 1024.0f;
    while (d.sent32 <= cap) d.sent32 *= 2.0f;
    d.oob32 = last_cell > s->cfg.eps ? fminf((float)(last_cell / res), cap) : d.sent32;
    d.edt64 = m.edt64; d.edt32 = m.edt32;
    f110_launch_edt32(m.edt64, m.edt32, cells, d.inv_res, s->cfg.eps, cap, d.sent32, st);
    launches += 1;
    d.edt32p = nullptr; d.edt32p_adj = 0; d.pad = 0; d.wp = 0;
    if (d.scan_is_f32) {   // padded copy for the fp32 ray-cast (see scan32p)
        d.pad = (int)ceil(s->cfg.max_range / res) + 2;
        d.wp = w + 2 * d.pad;
        size_t pcells = (size_t)d.wp * (h + 2 * d.pad);
        if (pcells >= (1ull << 31)) return fail(F110_ERR_INVALID, "map too large for the padded fp32 layout");
        CK(cudaMalloc(&m.edt32p, pcells * 4));
        f110_launch_pad_edt32(m.edt32, m.edt32p, h, w, d.pad, d.oob32, st);
        d.edt32p = m.edt32p;
        // index bias of the mantissa-floor trick: bits(2^23 + n) = 0x4B000000 + n
        for (unsigned cx = 0; ; cx += (1u << 21)) {
            uint64_t k = ((uint64_t)0x4B000000u * (uint64_t)d.wp + 0x4B000000u + cx) & 0xffffffffull;
            if (k + pcells < (1ull << 32)) {
                d.magic_x = 8388608.0f + (float)cx;
                d.edt32p_adj = (unsigned long long)m.edt32p - 4ull * k;
                break;
            }
        }
        launches += 1;
    }
    {
        // spatial launch order: tiles of >= 32 cells, at most 64 x 64 of them; 8 heading octants; chunk-major
        if (!d.order_chunk_envs) d.order_chunk_envs = d.n_envs;
        int tile = 32;
        while ((w + tile - 1) / tile > 64 || (h + tile - 1) / tile > 64) tile *= 2;
        d.order_tx = (w + tile - 1) / tile; d.order_ty = (h + tile - 1) / tile;
        d.order_inv_tile = 1.0f / (float)tile;
        int chunks = (d.n_envs + d.order_chunk_envs - 1) / d.order_chunk_envs;
        d.order_bins = chunks * d.order_tx * d.order_ty * 8;
        CK(cudaMalloc(&m.order_hist, (size_t)d.order_bins * 4));
        CK(cudaMalloc(&m.order_base, (size_t)d.order_bins * 4));
        CK(cudaMemsetAsync(m.order_hist, 0, (size_t)d.order_bins * 4, st));
        const size_t NA = (size_t)d.n_envs * d.n_agents;
        if (s->use_order) { d.order = s->order_buf; d.order_key = s->order_buf + NA; d.order_rank = s->order_buf + 2 * NA; }
        d.order_hist = m.order_hist; d.order_base = m.order_base;
    }
    CK(cudaGetLastError());
    CK(cudaStreamSynchronize(st));
    return F110_OK;
}

// edt: host [h][w] metres, or null with free_px: host [h][w] bytes (non-zero = free) to build the EDT on the device
static int set_map_impl(f110_sim *s, const double *edt, const uint8_t *free_px, int32_t h, int32_t w, double res, double ox,
                        double oy, double oc, double os, void *stream)
{
    if (!s || (!edt && !free_px) || h < 1 || w < 1 || !(res > 0)) return fail(F110_ERR_INVALID, "bad map arguments");
    if (free_px && (h > 65534 || w > 65534)) return fail(F110_ERR_INVALID, "map too large for the device EDT");
    if ((size_t)h * w >= (1ull << 31)) return fail(F110_ERR_INVALID, "map too large");
    // the fp32 map codes "stop here" (distance <= eps) as a sentinel and takes the last step as 0: exact only while no
    // free cell is that close to a wall, i.e. eps below one cell (the fp64 path takes any eps)
    if (s->d.scan_is_f32 && !(s->cfg.eps < res)) return fail(F110_ERR_INVALID, "the fp32 path needs eps (%g) below the map resolution (%g)", s->cfg.eps, res);
    cudaStream_t st = (cudaStream_t)stream;
    CK(cudaSetDevice(s->device));
    CK(cudaStreamSynchronize(st));
    MapBufs fresh;
    F110Dev nd = s->d;          // the new map's fields are filled into a copy
    int launches = 0;
    int rc = build_map(s, fresh, nd, edt, free_px, h, w, res, ox, oy, oc, os, st, launches);
    s->launches += launches;
    if (rc) {                   // the handle keeps the map it had (has_map unchanged)
        cudaStreamSynchronize(st);
        fresh.release();
        return rc;
    }
    // direction table depends on the map rotation: rebuild if tables are already there
    if (s->has_tables) {
        std::vector<double> sn(nd.theta_dis), cs(nd.theta_dis);
        cudaError_t e1 = cudaMemcpy(sn.data(), nd.sines, nd.theta_dis * 8, cudaMemcpyDeviceToHost);
        cudaError_t e2 = cudaMemcpy(cs.data(), nd.cosines, nd.theta_dis * 8, cudaMemcpyDeviceToHost);
        if (e1 != cudaSuccess || e2 != cudaSuccess) { fresh.release(); return fail(F110_ERR_CUDA, "table read-back failed"); }
        std::vector<float4> dir;
        std::vector<float2> dir2;
        build_dir_table(sn.data(), cs.data(), nd.theta_dis, oc, os, dir, dir2);
        e1 = cudaMemcpy(s->dir32, dir.data(), dir.size() * sizeof(float4), cudaMemcpyHostToDevice);
        e2 = cudaMemcpy(s->dir2, dir2.data(), dir2.size() * sizeof(float2), cudaMemcpyHostToDevice);
        if (e1 != cudaSuccess || e2 != cudaSuccess) { fresh.release(); return fail(F110_ERR_CUDA, "direction table upload failed"); }
    }
    MapBufs old;
    old.edt64 = s->edt64; old.edt32 = s->edt32; old.edt32p = s->edt32p; old.order_hist = s->order_hist; old.order_base = s->order_base;
    s->edt64 = fresh.edt64; s->edt32 = fresh.edt32; s->edt32p = fresh.edt32p; s->order_hist = fresh.order_hist; s->order_base = fresh.order_base;
    s->d = nd;
    s->has_map = true;
    old.release();
    return F110_OK;
}

extern "C" int f110_set_map(f110_sim *s, const double *edt, int32_t h, int32_t w, double res, double ox, double oy,
                            double oc, double os, void *stream)
{
    if (!edt) return fail(F110_ERR_INVALID, "bad map arguments");
    return set_map_impl(s, edt, nullptr, h, w, res, ox, oy, oc, os, stream);
}

extern "C" int f110_set_map_image(f110_sim *s, const uint8_t *free_px, int32_t h, int32_t w, double res, double ox, double oy,
                                  double oc, double os, void *stream)
{
    if (!free_px) return fail(F110_ERR_INVALID, "bad map arguments");
    return set_map_impl(s, nullptr, free_px, h, w, res, ox, oy, oc, os, stream);
}

extern "C" int f110_edt_build(const uint8_t *free_dev, int32_t h, int32_t w, double res, double *edt_dev, int device, void *stream)
{
    if (!free_dev || !edt_dev || h < 1 || w < 1 || h > 65534 || w > 65534 || !(res > 0)) return fail(F110_ERR_INVALID, "bad EDT arguments");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        cudaGetLastError();
        return fail(F110_ERR_CUDA, "no CUDA device available; this library has no CPU path");
    }
    if (device < 0 || device >= ndev) return fail(F110_ERR_INVALID, "device %d out of range (%d devices)", device, ndev);
    CK(cudaSetDevice(device));
    unsigned *g = nullptr;
    if (cudaMalloc(&g, (size_t)h * w * 4) != cudaSuccess) return fail(F110_ERR_NOMEM, "cudaMalloc failed");
    f110_launch_edt_build(free_dev, g, edt_dev, h, w, res, (cudaStream_t)stream);
    cudaError_t e = cudaGetLastError();
    cudaError_t e2 = cudaStreamSynchronize((cudaStream_t)stream);      // the scratch buffer is freed on return
    cudaFree(g);
    if (e != cudaSuccess || e2 != cudaSuccess) return fail(F110_ERR_CUDA, "EDT kernels failed: %s", cudaGetErrorString(e != cudaSuccess ? e : e2));
    return F110_OK;
}

// ---- model functions on their own (device pointers on `device`; asynchronous on `stream`) ----
static int model_device(int device)
{
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        cudaGetLastError();
        return fail(F110_ERR_CUDA, "no CUDA device available; this library has no CPU path");
    }
    if (device < 0 || device >= ndev) return fail(F110_ERR_INVALID, "device %d out of range (%d devices)", device, ndev);
    CK(cudaSetDevice(device));
    return F110_OK;
}
static int model_done(const char *what)
{
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return fail(F110_ERR_CUDA, "%s failed: %s", what, cudaGetErrorString(e));
    return F110_OK;
}
// f110_edt_build without any allocation (no counterpart in the reference):
// the scratch is a fixed static device array in f110_edt.cu.  Asynchronous on `stream`.
int f110_edt_build_static(const uint8_t *free_px, double *edt, int h, int w, double res, int device, cudaStream_t st,
                          const char **what, cudaError_t *err);
extern "C" int f110_edt_build_noalloc(const uint8_t *free_dev, int32_t h, int32_t w, double res, double *edt_dev, int device, void *stream)
{
    if (!free_dev || !edt_dev || h < 1 || w < 1 || !(res > 0)) return fail(F110_ERR_INVALID, "bad EDT arguments");
    if ((int64_t)h * w > F110_EDT_NOALLOC_MAX_CELLS || w > F110_EDT_NOALLOC_MAX_WIDTH)
        return fail(F110_ERR_INVALID, "f110_edt_build_noalloc: %d x %d cells exceed the static scratch (at most %d cells, width at most %d); "
                    "use f110_edt_build", h, w, F110_EDT_NOALLOC_MAX_CELLS, F110_EDT_NOALLOC_MAX_WIDTH);
    int rc = model_device(device);
    if (rc) return rc;
    const char *what = "";
    cudaError_t e = cudaSuccess;
    if (f110_edt_build_static(free_dev, edt_dev, h, w, res, device, (cudaStream_t)stream, &what, &e))
        return fail(F110_ERR_CUDA, "f110_edt_build_noalloc: %s failed: %s", what, cudaGetErrorString(e));
    return F110_OK;
}

extern "C" int f110_model_dynamics_st(const double *x_dev, const double *u_dev, const double *params16, int64_t n, double *f_dev,
                                      int device, void *stream)
{
    if (!x_dev || !u_dev || !params16 || !f_dev || n < 0) return fail(F110_ERR_INVALID, "bad arguments");
    int rc = model_device(device);
    if (rc) return rc;
    f110_launch_model_dynamics(x_dev, u_dev, params16, n, 0, f_dev, (cudaStream_t)stream);
    return model_done("vehicle_dynamics_st");
}
extern "C" int f110_model_dynamics_ks(const double *x_dev, const double *u_dev, const double *params16, int64_t n, double *f_dev,
                                      int device, void *stream)
{
    if (!x_dev || !u_dev || !params16 || !f_dev || n < 0) return fail(F110_ERR_INVALID, "bad arguments");
    int rc = model_device(device);
    if (rc) return rc;
    f110_launch_model_dynamics(x_dev, u_dev, params16, n, 1, f_dev, (cudaStream_t)stream);
    return model_done("vehicle_dynamics_ks");
}
extern "C" int f110_model_pid(const double *in_dev, double max_sv, double max_a, double max_v, double min_v, int64_t n, double *out_dev,
                              int device, void *stream)
{
    if (!in_dev || !out_dev || n < 0) return fail(F110_ERR_INVALID, "bad arguments");
    int rc = model_device(device);
    if (rc) return rc;
    f110_launch_model_pid(in_dev, max_sv, max_a, max_v, min_v, n, out_dev, (cudaStream_t)stream);
    return model_done("pid");
}
extern "C" int f110_model_accl_constraints(const double *in_dev, double v_switch, double a_max, double v_min, double v_max, int64_t n,
                                           double *out_dev, int device, void *stream)
{
    if (!in_dev || !out_dev || n < 0) return fail(F110_ERR_INVALID, "bad arguments");
    int rc = model_device(device);
    if (rc) return rc;
    f110_launch_model_constraint(in_dev, v_switch, a_max, v_min, v_max, n, 0, out_dev, (cudaStream_t)stream);
    return model_done("accl_constraints");
}
extern "C" int f110_model_steering_constraint(const double *in_dev, double s_min, double s_max, double sv_min, double sv_max, int64_t n,
                                              double *out_dev, int device, void *stream)
{
    if (!in_dev || !out_dev || n < 0) return fail(F110_ERR_INVALID, "bad arguments");
    int rc = model_device(device);
    if (rc) return rc;
    f110_launch_model_constraint(in_dev, s_min, s_max, sv_min, sv_max, n, 1, out_dev, (cudaStream_t)stream);
    return model_done("steering_constraint");
}
extern "C" int f110_model_vertices(const double *poses_dev, double length, double width, int64_t n, double *verts_dev, int device,
                                   void *stream)
{
    if (!poses_dev || !verts_dev || n < 0) return fail(F110_ERR_INVALID, "bad arguments");
    int rc = model_device(device);
    if (rc) return rc;
    f110_launch_model_vertices(poses_dev, length, width, n, verts_dev, (cudaStream_t)stream);
    return model_done("get_vertices");
}
extern "C" int f110_model_collision(const double *verts_a_dev, const double *verts_b_dev, int64_t n, uint8_t *hit_dev, int device,
                                    void *stream)
{
    if (!verts_a_dev || !verts_b_dev || !hit_dev || n < 0) return fail(F110_ERR_INVALID, "bad arguments");
    int rc = model_device(device);
    if (rc) return rc;
    f110_launch_model_collision(verts_a_dev, verts_b_dev, n, hit_dev, (cudaStream_t)stream);
    return model_done("collision");
}

extern "C" int f110_set_tables(f110_sim *s, const double *sines, const double *cosines, const double *scan_angles,
                               const double *beam_cosines, const double *side_distances, void *stream)
{
    if (!s || !sines || !cosines || !scan_angles || !beam_cosines || !side_distances) return fail(F110_ERR_INVALID, "null table");
    CK(cudaSetDevice(s->device));
    cudaStream_t st = (cudaStream_t)stream;
    CK(cudaStreamSynchronize(st));
    F110Dev &d = s->d;
    const int T = d.theta_dis, B = d.n_beams;
    CK(cudaMemcpy((void *)d.sines, sines, T * 8, cudaMemcpyHostToDevice));
    CK(cudaMemcpy((void *)d.cosines, cosines, T * 8, cudaMemcpyHostToDevice));
    CK(cudaMemcpy((void *)d.scan_angles, scan_angles, B * 8, cudaMemcpyHostToDevice));
    CK(cudaMemcpy((void *)d.beam_cos, beam_cosines, B * 8, cudaMemcpyHostToDevice));
    CK(cudaMemcpy((void *)d.side_dist, side_distances, B * 8, cudaMemcpyHostToDevice));
    std::vector<float> f(3 * B);
    float side_max = 0.f;
    for (int i = 0; i < B; i++) {
        f[i] = (float)beam_cosines[i]; f[B + i] = (float)side_distances[i];
        f[2 * B + i] = (float)(1.2 * s->cfg.ttc_thresh * fabs(beam_cosines[i]));
        if (side_distances[i] < 1e30 && (float)side_distances[i] > side_max) side_max = (float)side_distances[i];
    }
    d.side_max32 = side_max;
    CK(cudaMemcpy((void *)d.beam_cos32, f.data(), f.size() * 4, cudaMemcpyHostToDevice));
    double oc = s->has_map ? d.oc : 1.0, os = s->has_map ? d.os : 0.0;
    std::vector<float4> dir;
    std::vector<float2> dir2;
    build_dir_table(sines, cosines, T, oc, os, dir, dir2);
    CK(cudaMemcpy(s->dir32, dir.data(), dir.size() * sizeof(float4), cudaMemcpyHostToDevice));
    CK(cudaMemcpy(s->dir2, dir2.data(), dir2.size() * sizeof(float2), cudaMemcpyHostToDevice));
    s->has_tables = true;
    return F110_OK;
}

extern "C" int f110_set_params(f110_sim *s, const double *p, int32_t agent_idx, void *stream)
{
    if (!s || !p) return fail(F110_ERR_INVALID, "null argument");
    CK(cudaSetDevice(s->device));
    cudaStream_t st = (cudaStream_t)stream;
    const int A = s->d.n_agents;
    if (agent_idx > A) return fail(F110_ERR_INVALID, "Index given is out of bounds for list of agents.");
    CK(cudaStreamSynchronize(st));
    double *base = (double *)s->d.params;
    if (agent_idx == A) {
        double wl[2] = { p[F110_P_WIDTH], p[F110_P_LENGTH] };
        CK(cudaMemcpy(base + (size_t)F110_NP * A, wl, 16, cudaMemcpyHostToDevice));
    } else if (agent_idx < 0) {
        for (int a = 0; a < A; a++) CK(cudaMemcpy(base + (size_t)F110_NP * a, p, F110_NP * 8, cudaMemcpyHostToDevice));
    } else {
        CK(cudaMemcpy(base + (size_t)F110_NP * agent_idx, p, F110_NP * 8, cudaMemcpyHostToDevice));
    }
    return F110_OK;
}

extern "C" int f110_set_noise_rows(f110_sim *s, const double *rows, int64_t first_row, int64_t n_rows, void *stream)
{
    if (!s || !rows || first_row < 0 || n_rows < 1) return fail(F110_ERR_INVALID, "bad noise rows");
    if (s->cfg.noise_mode != F110_NOISE_TABLE) return fail(F110_ERR_STATE, "handle was not created with F110_NOISE_TABLE");
    CK(cudaSetDevice(s->device));
    cudaStream_t st = (cudaStream_t)stream;
    const size_t B = s->d.n_beams;
    long long need = first_row + n_rows;
    if (need > s->noise_cap) {
        // the first table holds 64 MB of rows (7767 steps at 1080 beams; at least 256, at most 8192 rows): growing it means a
        // stream synchronisation, cudaMalloc and cudaFree in the middle of a rollout, measured at 5 ... 146 ms on this pool
        long long cap = s->noise_cap;
        if (!cap) {
            cap = (long long)((64ull << 20) / (B * 8));
            cap = cap < 256 ? 256 : (cap > 8192 ? 8192 : cap);
        }
        while (cap < need) cap *= 2;
        double *nn = nullptr;
        CK(cudaStreamSynchronize(st));
        CK(cudaMalloc(&nn, (size_t)cap * B * 8));
        if (s->noise) {
            CK(cudaMemcpy(nn, s->noise, (size_t)s->d.noise_rows * B * 8, cudaMemcpyDeviceToDevice));
            cudaFree(s->noise);
        }
        s->noise = nn; s->noise_cap = cap;
        s->d.noise = nn;
    }
    if (first_row >= s->d.noise_rows) {
        // rows no enqueued step reads (a step is only launched for rows below noise_rows): upload them beside the caller's
        // stream instead of draining it -- steps already queued keep the GPU busy, later ones wait for the event
        if (!s->noise_ev) CK(cudaEventCreateWithFlags(&s->noise_ev, cudaEventDisableTiming));
        CK(cudaMemcpyAsync(s->noise + (size_t)first_row * B, rows, (size_t)n_rows * B * 8, cudaMemcpyHostToDevice, s->copy_stream));
        CK(cudaEventRecord(s->noise_ev, s->copy_stream));
        CK(cudaStreamWaitEvent(st, s->noise_ev, 0));
        CK(cudaStreamSynchronize(s->copy_stream));      // the caller's buffer is free again on return
    } else {
        CK(cudaMemcpyAsync(s->noise + (size_t)first_row * B, rows, (size_t)n_rows * B * 8, cudaMemcpyHostToDevice, st));
        CK(cudaStreamSynchronize(st));
    }
    if (need > s->d.noise_rows) s->d.noise_rows = need;
    return F110_OK;
}

extern "C" int f110_set_reset_pool(f110_sim *s, const double *poses, int32_t n_poses, void *stream)
{
    if (!s || !poses || n_poses < 1) return fail(F110_ERR_INVALID, "bad reset pool");
    CK(cudaSetDevice(s->device));
    cudaStream_t st = (cudaStream_t)stream;
    CK(cudaStreamSynchronize(st));
    if (s->pool) cudaFree(s->pool);
    size_t n = (size_t)n_poses * s->d.n_agents * 3;
    CK(cudaMalloc(&s->pool, n * 8));
    CK(cudaMemcpy(s->pool, poses, n * 8, cudaMemcpyHostToDevice));
    s->d.pool = s->pool; s->d.n_pool = n_poses;
    return F110_OK;
}

static int check_ready(f110_sim *s)
{
    if (!s) return fail(F110_ERR_INVALID, "null handle");
    if (!s->has_map) return fail(F110_ERR_STATE, "Map is not set for scan simulator.");
    if (!s->has_tables) return fail(F110_ERR_STATE, "scan tables are not set");
    if (s->cfg.noise_mode == F110_NOISE_TABLE && s->d.noise_rows < 1) return fail(F110_ERR_STATE, "noise rows are not set");
    if (s->cfg.auto_reset && s->d.n_pool < 1) return fail(F110_ERR_STATE, "auto_reset needs f110_set_reset_pool");
    return 0;
}

static bool use_fused(const f110_sim *s)
{
    const F110Dev &d = s->d;
    if (d.scan_is_f32 || (s->cfg.raycast_variant & (128 | 256))) return false;       // variant bits 7 / 8: the step kernel by kernel
    if (d.n_agents == 1) return d.n_envs <= F110_FUSED_MAX_ENVS;
    return d.n_agents <= F110_FUSED_MULTI_MAX_AGENTS && d.n_envs <= F110_FUSED_MULTI_MAX_ENVS && d.n_envs <= d.n_sm && !d.count_lookups;
}

// kernels of one step, in the order of Simulator.step + F110Env.step.  host_*: mapped host buffers the fused
// small-batch kernel may write the observation to directly (all null: device buffers only)
static int enqueue_step(f110_sim *s, const uint8_t *mask_dev, cudaStream_t st, double *host_obs = nullptr, double *host_scans = nullptr,
                        uint8_t *host_done = nullptr, bool *exported = nullptr, const double *host_cmd = nullptr)
{
    F110Dev &d = s->d;
    d.step_mask = mask_dev;
    if (exported) *exported = false;
    if (use_fused(s)) {
        if (s->timing) cudaEventRecord(s->ev_a, st);
        f110_launch_step_fused64(d, host_obs, host_scans, host_done, host_cmd, s->cfg.raycast_variant, st);
        if (s->timing) { cudaEventRecord(s->ev_b, st); s->timed = true; }
        s->launches++;
        d.step_mask = nullptr;
        if (exported) *exported = true;
        CK(cudaGetLastError());
        return 0;
    }
    f110_launch_dynamics(d, st);
    if (d.order) {
        // one launch over all envs: sort them as ONE chunk (f110_step_host sorts per copy-out chunk); a single sweep over
        // the map instead of one per chunk is worth 2 % of the ray-cast at 2^20 envs
        F110Dev whole = d;
        whole.order_chunk_envs = d.n_envs;
        f110_launch_order(whole, st);
        s->launches += 3;
    }
    if (s->timing) cudaEventRecord(s->ev_a, st);
    if (d.scan_is_f32) f110_launch_raycast32(d, 0, d.n_envs, s->cfg.raycast_variant, st);
    else f110_launch_raycast64(d, 0, d.n_envs, s->cfg.raycast_variant, st);
    if (s->timing) { cudaEventRecord(s->ev_b, st); s->timed = true; }
    s->launches += 2;
    if (d.n_agents > 1) { f110_launch_agents(d, st); s->launches++; }
    f110_launch_finalize(d, st);
    s->launches++;
    d.step_mask = nullptr;
    CK(cudaGetLastError());
    return 0;
}

extern "C" int f110_reset(f110_sim *s, const double *poses, const double *start_rot, const uint8_t *mask, void *stream)
{
    TRY(check_ready(s));
    if (!poses) return fail(F110_ERR_INVALID, "Number of poses for reset does not match number of agents.");
    CK(cudaSetDevice(s->device));
    cudaStream_t st = (cudaStream_t)stream;
    const size_t N = s->d.n_envs, NA = N * s->d.n_agents;
    CK(cudaMemcpyAsync(s->stage_poses, poses, 3 * NA * 8, cudaMemcpyHostToDevice, st));
    if (start_rot) CK(cudaMemcpyAsync(s->stage_rot, start_rot, 4 * N * 8, cudaMemcpyHostToDevice, st));
    if (mask) CK(cudaMemcpyAsync(s->stage_mask, mask, N, cudaMemcpyHostToDevice, st));
    f110_launch_reset(s->d, s->stage_poses, start_rot ? s->stage_rot : nullptr, mask ? s->stage_mask : nullptr, st);
    s->launches++;
    // zero-action step (f110_env.py:335-336) on the reset envs only
    const double *saved = s->d.actions;
    double *zero = s->stage_poses;   // reuse: k_reset has consumed it by the time the memset runs (stream order)
    CK(cudaMemsetAsync(zero, 0, 2 * NA * 8, st));
    s->d.actions = zero;
    int rc = enqueue_step(s, mask ? s->stage_mask : nullptr, st);
    s->d.actions = saved;
    if (rc) return rc;
    // the staging buffers are reused by the next call
    CK(cudaStreamSynchronize(st));
    return F110_OK;
}

extern "C" int f110_reset_poses(f110_sim *s, const double *poses, const double *start_rot, const uint8_t *mask, void *stream)
{
    TRY(check_ready(s));
    if (!poses) return fail(F110_ERR_INVALID, "Number of poses for reset does not match number of agents.");
    CK(cudaSetDevice(s->device));
    cudaStream_t st = (cudaStream_t)stream;
    const size_t N = s->d.n_envs, NA = N * s->d.n_agents;
    CK(cudaMemcpyAsync(s->stage_poses, poses, 3 * NA * 8, cudaMemcpyHostToDevice, st));
    if (start_rot) CK(cudaMemcpyAsync(s->stage_rot, start_rot, 4 * N * 8, cudaMemcpyHostToDevice, st));
    if (mask) CK(cudaMemcpyAsync(s->stage_mask, mask, N, cudaMemcpyHostToDevice, st));
    f110_launch_reset(s->d, s->stage_poses, start_rot ? s->stage_rot : nullptr, mask ? s->stage_mask : nullptr, st);
    s->launches++;
    CK(cudaGetLastError());
    CK(cudaStreamSynchronize(st));      // the staging buffers are reused by the next call
    return F110_OK;
}

extern "C" int f110_step(f110_sim *s, const double *actions_dev, void *stream)
{
    TRY(check_ready(s));
    CK(cudaSetDevice(s->device));
    s->d.actions = actions_dev ? actions_dev : s->actions;
    int rc = enqueue_step(s, nullptr, (cudaStream_t)stream);
    s->d.actions = s->actions;
    return rc;
}

// Small batches through f110_step_host: the observation is a few KB and three cudaMemcpyAsync + their engine switches
// cost more than the kernels.  When every host buffer is page-locked and mapped (what cudaHostAlloc / torch pin_memory
// give under unified addressing), the step reads the actions straight from the host buffer and one more kernel writes
// obs_block, scans and done straight into the host buffers (posted PCIe writes): no copy engine involved.  Same bytes.
__global__ void __launch_bounds__(256) k_export_host(const double *__restrict__ obs, double *__restrict__ obs_h, size_t n_obs,
                                                     const unsigned *__restrict__ scans, unsigned *__restrict__ scans_h, size_t n_scan_words,
                                                     const uint8_t *__restrict__ done, uint8_t *__restrict__ done_h, size_t n_done)
{
    const size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x, nt = (size_t)gridDim.x * blockDim.x;
    if (obs_h) for (size_t i = t; i < n_obs; i += nt) obs_h[i] = obs[i];
    if (scans_h) for (size_t i = t; i < n_scan_words; i += nt) scans_h[i] = scans[i];
    if (done_h) for (size_t i = t; i < n_done; i += nt) done_h[i] = done[i];
}

// device-side address of a page-locked, mapped host buffer; null if p is anything else (pageable, device, null)
static void *mapped_host_ptr(const void *p)
{
    if (!p) return nullptr;
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return nullptr; }
    return a.type == cudaMemoryTypeHost ? a.devicePointer : nullptr;
}
#define F110_ZERO_COPY_MAX_BYTES (256 * 1024)

extern "C" int f110_step_host(f110_sim *s, const double *actions_host, double *obs_block_host, void *scans_host,
                              uint8_t *done_host, void *stream)
{
    TRY(check_ready(s));
    if (!actions_host) return fail(F110_ERR_INVALID, "null actions");
    CK(cudaSetDevice(s->device));
    cudaStream_t st = (cudaStream_t)stream;
    F110Dev &d = s->d;
    const size_t N = d.n_envs, A = d.n_agents, NA = N * A, B = d.n_beams;
    const size_t esz = d.scan_u16 ? 2 : d.scan_is_f32 ? 4 : 8;
    s->timed = false;
    if ((scans_host ? NA * B * esz : 0) + 10 * NA * 8 <= F110_ZERO_COPY_MAX_BYTES) {
        const double *act_d = (const double *)mapped_host_ptr(actions_host);
        double *obs_d = (double *)mapped_host_ptr(obs_block_host);
        unsigned *scans_d = (unsigned *)mapped_host_ptr(scans_host);
        uint8_t *done_d = (uint8_t *)mapped_host_ptr(done_host);
        if (act_d && (obs_d || !obs_block_host) && (scans_d || !scans_host) && (done_d || !done_host)) {
            s->d.actions = act_d;
            bool exported = false;
            // host_cmd: a lone env's command travels as a kernel argument (read here, on the host, before the launch)
            int rc = enqueue_step(s, nullptr, st, obs_d, (double *)scans_d, done_d, &exported, actions_host);
            s->d.actions = s->actions;
            if (rc) return rc;
            if (!exported) {
                const size_t words = NA * B * esz / 4;        // B is even for 2-byte scans (f110_create)
                const int grid = (int)((words / 4 + 255) / 256);
                k_export_host<<<grid < 1 ? 1 : (grid > 64 ? 64 : grid), 256, 0, st>>>(s->obs_block, obs_d, 10 * NA, (const unsigned *)d.scans, scans_d,
                                                                                        words, d.done, done_d, N);
                s->launches++;
            }
            CK(cudaGetLastError());
            CK(cudaStreamSynchronize(st));
            return F110_OK;
        }
    }
    CK(cudaMemcpyAsync(s->actions, actions_host, 2 * NA * 8, cudaMemcpyHostToDevice, st));
    d.actions = s->actions;
    d.step_mask = nullptr;
    if (use_fused(s)) {      // pageable host buffers: the fused kernel, then the copy engines
        TRY(enqueue_step(s, nullptr, st));
        if (scans_host) CK(cudaMemcpyAsync(scans_host, d.scans, NA * B * esz, cudaMemcpyDeviceToHost, st));
        if (obs_block_host) CK(cudaMemcpyAsync(obs_block_host, s->obs_block, 10 * NA * 8, cudaMemcpyDeviceToHost, st));
        if (done_host) CK(cudaMemcpyAsync(done_host, d.done, N, cudaMemcpyDeviceToHost, st));
        CK(cudaStreamSynchronize(st));
        return F110_OK;
    }
    f110_launch_dynamics(d, st);
    s->launches++;
    if (d.order) { f110_launch_order(d, st); s->launches += 3; }
    // The scans dominate the D2H traffic.  With one agent per env nothing after the ray-cast touches
    // them, so the ray-cast is issued in env chunks and each chunk's copy-out (on a second stream)
    // overlaps the next chunk's kernel.
    int chunks = 1;
    if (scans_host && A == 1 && N >= 8192) chunks = 8;
    if ((int)s->chunk_ev.size() < chunks) {
        size_t old = s->chunk_ev.size();
        s->chunk_ev.resize(chunks);
        for (size_t i = old; i < s->chunk_ev.size(); i++) cudaEventCreateWithFlags(&s->chunk_ev[i], cudaEventDisableTiming);
    }
    size_t per = (N + chunks - 1) / chunks;     // == d.order_chunk_envs when the scans are ordered
    if (s->timing) cudaEventRecord(s->ev_a, st);     // all chunk launches (the copies run on the other stream)
    for (int c = 0; c < chunks; c++) {
        size_t b = c * per, cnt = (b + per <= N) ? per : (N > b ? N - b : 0);
        if (!cnt) break;
        if (d.scan_is_f32) f110_launch_raycast32(d, (int)b, (int)cnt, s->cfg.raycast_variant, st);
        else f110_launch_raycast64(d, (int)b, (int)cnt, s->cfg.raycast_variant, st);
        s->launches++;
        if (chunks > 1) {
            CK(cudaEventRecord(s->chunk_ev[c], st));
            CK(cudaStreamWaitEvent(s->copy_stream, s->chunk_ev[c], 0));
            CK(cudaMemcpyAsync((char *)scans_host + b * A * B * esz, (char *)d.scans + b * A * B * esz, cnt * A * B * esz,
                               cudaMemcpyDeviceToHost, s->copy_stream));
        }
    }
    if (s->timing) { cudaEventRecord(s->ev_b, st); s->timed = true; }
    if (A > 1) { f110_launch_agents(d, st); s->launches++; }
    f110_launch_finalize(d, st);
    s->launches++;
    CK(cudaGetLastError());
    if (scans_host && chunks == 1)
        CK(cudaMemcpyAsync(scans_host, d.scans, NA * B * esz, cudaMemcpyDeviceToHost, st));
    if (obs_block_host) CK(cudaMemcpyAsync(obs_block_host, s->obs_block, 10 * NA * 8, cudaMemcpyDeviceToHost, st));
    if (done_host) CK(cudaMemcpyAsync(done_host, d.done, N, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    if (chunks > 1) CK(cudaStreamSynchronize(s->copy_stream));
    return F110_OK;
}

extern "C" int f110_raycast(f110_sim *s, const double *poses_dev, int32_t n, void *out_dev, void *stream)
{
    if (!s) return fail(F110_ERR_INVALID, "null handle");
    if (!s->has_map) return fail(F110_ERR_STATE, "Map is not set for scan simulator.");
    if (!s->has_tables) return fail(F110_ERR_STATE, "scan tables are not set");
    if (n == 0) return F110_OK;
    if (!poses_dev || !out_dev || n < 0) return fail(F110_ERR_INVALID, "bad raycast arguments");
    CK(cudaSetDevice(s->device));
    const int *order = nullptr;
    if (n >= F110_ORDER_FROM && !(s->cfg.raycast_variant & 64)) {    // same reasoning as in the step: neighbours in launch order = neighbours on the map
        if ((size_t)n > s->rc_order_cap) {
            CK(cudaStreamSynchronize((cudaStream_t)stream));
            if (s->rc_order_buf) cudaFree(s->rc_order_buf);
            s->rc_order_buf = nullptr; s->rc_order_cap = 0;
            CK(cudaMalloc(&s->rc_order_buf, 3 * (size_t)n * 4));
            s->rc_order_cap = n;
        }
        f110_launch_order_poses(s->d, poses_dev, n, s->rc_order_buf, (cudaStream_t)stream);
        s->launches += 3;
        order = s->rc_order_buf;
    }
    if (s->d.scan_is_f32) f110_launch_scan_only32(s->d, poses_dev, n, (float *)out_dev, order, s->cfg.raycast_variant, (cudaStream_t)stream);
    else f110_launch_scan_only64(s->d, poses_dev, n, (double *)out_dev, order, s->cfg.raycast_variant, (cudaStream_t)stream);
    s->launches++;
    CK(cudaGetLastError());
    return F110_OK;
}

extern "C" int f110_get_buffers(f110_sim *s, f110_buffers *o)
{
    if (!s || !o) return fail(F110_ERR_INVALID, "null argument");
    F110Dev &d = s->d;
    o->state = d.state; o->obs_block = s->obs_block; o->collisions = d.collisions; o->collision_idx = d.collision_idx;
    o->lap_times = d.lap_times; o->lap_counts = d.lap_counts; o->toggles = d.toggles; o->done = d.done;
    o->in_collision = d.in_collision; o->scans = d.scans; o->cur_time = d.cur_time; o->noise_step = (int64_t *)d.noise_step;
    o->actions = s->actions; o->episode_count = (int64_t *)d.episode_count; o->lookups = (uint64_t *)d.lookups; o->checkpoint_done = d.checkpoint_done;
    return F110_OK;
}

extern "C" int f110_get_config(f110_sim *s, f110_config *o)
{
    if (!s || !o) return fail(F110_ERR_INVALID, "null argument");
    *o = s->cfg;
    return F110_OK;
}

extern "C" int f110_count_lookups(f110_sim *s, int32_t enable)
{
    if (!s) return fail(F110_ERR_INVALID, "null handle");
    s->d.count_lookups = enable ? 1 : 0;
    return F110_OK;
}

extern "C" int f110_enable_timing(f110_sim *s, int32_t enable)
{
    if (!s) return fail(F110_ERR_INVALID, "null handle");
    s->timing = enable != 0;
    if (!s->timing) s->timed = false;
    return F110_OK;
}

extern "C" int f110_last_raycast_ms(f110_sim *s, float *ms)
{
    if (!s || !ms) return fail(F110_ERR_INVALID, "null argument");
    if (!s->timed) return fail(F110_ERR_STATE, "no timed ray-cast: call f110_enable_timing(sim, 1) before the step");
    CK(cudaSetDevice(s->device));
    CK(cudaEventSynchronize(s->ev_b));
    CK(cudaEventElapsedTime(ms, s->ev_a, s->ev_b));
    return F110_OK;
}

extern "C" int f110_l2_probe(int device, int64_t table_bytes, int32_t repeats, double *out6, void *stream)
{
    if (!out6 || table_bytes < (1 << 20) || table_bytes > (1ll << 32) || repeats < 1) return fail(F110_ERR_INVALID, "bad probe arguments");
    int rc = model_device(device);
    if (rc) return rc;
    int n_sm = 148;
    cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, device);
    int r = f110_probe_l2(n_sm, (size_t)table_bytes & ~(size_t)127, repeats, out6, (cudaStream_t)stream);
    if (r) return fail(r == 1 ? F110_ERR_NOMEM : F110_ERR_CUDA, "L2 probe failed (%s)", cudaGetErrorString(cudaGetLastError()));
    return F110_OK;
}

extern "C" int f110_build_is_checked(void)
{
#ifdef F110_CHECKED
    return 1;
#else
    return 0;
#endif
}

extern "C" int f110_debug_violations(int device, uint64_t *out8, int32_t selftest)
{
    if (!out8) return fail(F110_ERR_INVALID, "null argument");
    int rc = model_device(device);
    if (rc) return rc;
    unsigned long long *v = violations_for(device);
    if (!v) return fail(F110_ERR_NOMEM, "cudaMalloc failed");
    if (selftest) f110_launch_check_selftest(v, nullptr);
    CK(cudaDeviceSynchronize());
    CK(cudaMemcpy(out8, v, 8 * sizeof(uint64_t), cudaMemcpyDeviceToHost));
    return F110_OK;
}

extern "C" int64_t f110_launch_count(f110_sim *s) { return s ? s->launches : 0; }

extern "C" int f110_synchronize(f110_sim *s, void *stream)
{
    if (!s) return fail(F110_ERR_INVALID, "null handle");
    CK(cudaSetDevice(s->device));
    CK(cudaStreamSynchronize((cudaStream_t)stream));
    return F110_OK;
}

// ---- scan -> bitmap rasteriser (f110_bitmap.cu) -------------------------------------------------------------------
struct f110_bitmap {
    f110_bitmap_config cfg;
    int device;
    F110BitmapDev d;
    void *tables = nullptr;
    long long launches = 0;
};

extern "C" int f110_bitmap_create(const f110_bitmap_config *cfg, const int32_t *indices, const double *cos_a, const double *sin_a,
                                  int device, f110_bitmap **out)
{
    if (!cfg || !indices || !cos_a || !sin_a || !out) return fail(F110_ERR_INVALID, "null argument");
    if (cfg->abi_version != F110_ABI_VERSION) return fail(F110_ERR_INVALID, "abi_version %d != %d", cfg->abi_version, F110_ABI_VERSION);
    if (cfg->height < 1 || cfg->width < 1) return fail(F110_ERR_INVALID, "output_image_dims must be at least 1x1");
    if (!(cfg->target_beam_count > 0 && cfg->target_beam_count < cfg->n_beams))
        return fail(F110_ERR_INVALID, "need 0 < target_beam_count < len(scan)");
    if (cfg->draw_mode < F110_BITMAP_FILL || cfg->draw_mode > F110_BITMAP_RAYS) return fail(F110_ERR_INVALID, "bad draw_mode");
    for (int i = 0; i < cfg->target_beam_count; i++)
        if (indices[i] < 0 || indices[i] >= cfg->n_beams) return fail(F110_ERR_INVALID, "beam index out of range");
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0) {
        cudaGetLastError();
        return fail(F110_ERR_CUDA, "no CUDA device available; this library has no CPU path");
    }
    if (device < 0 || device >= ndev) return fail(F110_ERR_INVALID, "device %d out of range (%d devices)", device, ndev);
    CK(cudaSetDevice(device));
    f110_bitmap *b = new f110_bitmap();
    b->cfg = *cfg;
    b->device = device;
    F110BitmapDev &d = b->d;
    memset(&d, 0, sizeof(d));
    d.violations = violations_for(device);
    const int T = cfg->target_beam_count;
    d.n_beams = cfg->n_beams; d.n_points = T; d.height = cfg->height; d.width = cfg->width;
    d.wpr = (cfg->width + 31) / 32;
    d.draw_mode = cfg->draw_mode; d.draw_center = cfg->draw_center; d.scan_is_f32 = cfg->scan_is_f32;
    d.cx = cfg->height / 2; d.cy = cfg->width / 2;           // lidar.py:79: center = [dims[0] // 2, dims[1] // 2], used as (x, y)
    d.bg = cfg->bg_white ? 255 : 0; d.fg = cfg->bg_white ? 0 : 255;
    d.scale = cfg->scaling_factor; d.scale32 = (float)cfg->scaling_factor;
    const size_t smem = f110_bitmap_smem_bytes(d);
    if (smem > 200 * 1024) { delete b; return fail(F110_ERR_INVALID, "image too large for the shared-memory rasteriser (%zu bytes)", smem); }
    size_t bytes = (size_t)T * (4 + 8 + 8) + 16;
    if (cudaMalloc(&b->tables, bytes) != cudaSuccess) { delete b; return fail(F110_ERR_NOMEM, "cudaMalloc failed"); }
    double *c = (double *)b->tables, *sn = c + T;
    int *ix = (int *)(sn + T);
    cudaMemcpy(c, cos_a, (size_t)T * 8, cudaMemcpyHostToDevice);
    cudaMemcpy(sn, sin_a, (size_t)T * 8, cudaMemcpyHostToDevice);
    e = cudaMemcpy(ix, indices, (size_t)T * 4, cudaMemcpyHostToDevice);
    if (e != cudaSuccess) { cudaFree(b->tables); delete b; return fail(F110_ERR_CUDA, "table upload failed: %s", cudaGetErrorString(e)); }
    d.cosa = c; d.sina = sn; d.indices = ix;
    *out = b;
    return F110_OK;
}

extern "C" int f110_bitmap_destroy(f110_bitmap *b)
{
    if (!b) return F110_OK;
    cudaSetDevice(b->device);
    cudaDeviceSynchronize();
    if (b->tables) cudaFree(b->tables);
    delete b;
    return F110_OK;
}

extern "C" int f110_bitmap_render(f110_bitmap *b, const void *scans_dev, int64_t n, void *out_dev, int32_t packed, void *stream)
{
    if (!b) return fail(F110_ERR_INVALID, "null handle");
    if (n == 0) return F110_OK;
    if (!scans_dev || !out_dev || n < 0) return fail(F110_ERR_INVALID, "bad render arguments");
    CK(cudaSetDevice(b->device));
    f110_launch_bitmap(b->d, scans_dev, n, (uint8_t *)out_dev, packed, (cudaStream_t)stream);
    b->launches++;
    CK(cudaGetLastError());
    return F110_OK;
}

// ---- pure-pursuit planner (f110_pursuit.cu) -----------------------------------------------------------------------
struct f110_pursuit {
    int device;
    F110PursuitDev d;
    double *tables = nullptr;
};

extern "C" int f110_pursuit_create(const double *wx, const double *wy, const double *wv, int32_t n, double wheelbase,
                                   double max_reacquire, int device, f110_pursuit **out)
{
    if (!wx || !wy || !wv || !out) return fail(F110_ERR_INVALID, "null argument");
    if (n < 2 || n > 65536) return fail(F110_ERR_INVALID, "need 2 <= n_waypoints <= 65536");
    if (((size_t)6 * n + 3 * (size_t)(n / 16 + 1)) * 8 > 200 * 1024) return fail(F110_ERR_INVALID, "raceline too long for the shared-memory planner (%d waypoints)", n);
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0) {
        cudaGetLastError();
        return fail(F110_ERR_CUDA, "no CUDA device available; this library has no CPU path");
    }
    if (device < 0 || device >= ndev) return fail(F110_ERR_INVALID, "device %d out of range (%d devices)", device, ndev);
    CK(cudaSetDevice(device));
    f110_pursuit *p = new f110_pursuit();
    p->device = device;
    if (cudaMalloc(&p->tables, (size_t)3 * n * 8) != cudaSuccess) { delete p; return fail(F110_ERR_NOMEM, "cudaMalloc failed"); }
    cudaMemcpy(p->tables, wx, (size_t)n * 8, cudaMemcpyHostToDevice);
    cudaMemcpy(p->tables + n, wy, (size_t)n * 8, cudaMemcpyHostToDevice);
    e = cudaMemcpy(p->tables + 2 * (size_t)n, wv, (size_t)n * 8, cudaMemcpyHostToDevice);
    if (e != cudaSuccess) { cudaFree(p->tables); delete p; return fail(F110_ERR_CUDA, "waypoint upload failed: %s", cudaGetErrorString(e)); }
    p->d.n_waypoints = n; p->d.wheelbase = wheelbase; p->d.max_reacquire = max_reacquire;
    p->d.wx = p->tables; p->d.wy = p->tables + n; p->d.wv = p->tables + 2 * (size_t)n;
    p->d.violations = violations_for(device);
    *out = p;
    return F110_OK;
}

extern "C" int f110_pursuit_destroy(f110_pursuit *p)
{
    if (!p) return F110_OK;
    cudaSetDevice(p->device);
    cudaDeviceSynchronize();
    if (p->tables) cudaFree(p->tables);
    delete p;
    return F110_OK;
}

extern "C" int f110_pursuit_plan(f110_pursuit *p, const double *x_dev, const double *y_dev, const double *theta_dev, int64_t n,
                                 int64_t stride, double lookahead, double vgain, const double *lookahead_dev, const double *vgain_dev,
                                 double *actions_dev, int32_t *picked_dev, void *stream)
{
    if (!p) return fail(F110_ERR_INVALID, "null handle");
    if (n == 0) return F110_OK;
    if (!x_dev || !y_dev || !theta_dev || !actions_dev || n < 0 || stride < 1) return fail(F110_ERR_INVALID, "bad plan arguments");
    CK(cudaSetDevice(p->device));
    f110_launch_pursuit(p->d, x_dev, y_dev, theta_dev, n, stride, lookahead, vgain, lookahead_dev, vgain_dev, actions_dev,
                        (int *)picked_dev, (cudaStream_t)stream);
    CK(cudaGetLastError());
    return F110_OK;
}

// ---- headless renderer (f110_render.cu) ---------------------------------------------------------------------------
struct f110_render {
    int device;
    F110RenderDev d;
    uint8_t *obstacle = nullptr;
};

extern "C" int f110_render_create(const f110_render_config *cfg, const uint8_t *obstacle, int device, f110_render **out)
{
    if (!cfg || !obstacle || !out) return fail(F110_ERR_INVALID, "null argument");
    if (cfg->abi_version != F110_ABI_VERSION) return fail(F110_ERR_INVALID, "abi_version %d != %d", cfg->abi_version, F110_ABI_VERSION);
    if (cfg->width < 1 || cfg->height < 1 || cfg->map_h < 1 || cfg->map_w < 1 || !(cfg->resolution > 0))
        return fail(F110_ERR_INVALID, "bad renderer sizes");
    if (cfg->n_agents < 1 || cfg->n_agents > 16 || cfg->ego_idx < 0 || cfg->ego_idx >= cfg->n_agents)
        return fail(F110_ERR_INVALID, "bad agent count / ego index");
    if (cfg->follow_margin < 0 && !(cfg->right > cfg->left && cfg->top > cfg->bottom)) return fail(F110_ERR_INVALID, "empty camera window");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        cudaGetLastError();
        return fail(F110_ERR_CUDA, "no CUDA device available; this library has no CPU path");
    }
    if (device < 0 || device >= ndev) return fail(F110_ERR_INVALID, "device %d out of range (%d devices)", device, ndev);
    CK(cudaSetDevice(device));
    f110_render *r = new f110_render();
    r->device = device;
    const size_t cells = (size_t)cfg->map_h * cfg->map_w;
    if (cudaMalloc(&r->obstacle, cells) != cudaSuccess) { delete r; return fail(F110_ERR_NOMEM, "cudaMalloc failed"); }
    cudaError_t e = cudaMemcpy(r->obstacle, obstacle, cells, cudaMemcpyHostToDevice);
    if (e != cudaSuccess) { cudaFree(r->obstacle); delete r; return fail(F110_ERR_CUDA, "map upload failed: %s", cudaGetErrorString(e)); }
    F110RenderDev &d = r->d;
    d.width = cfg->width; d.height = cfg->height; d.n_agents = cfg->n_agents; d.ego_idx = cfg->ego_idx;
    d.map_h = cfg->map_h; d.map_w = cfg->map_w; d.res = cfg->resolution; d.ox = cfg->origin_x; d.oy = cfg->origin_y;
    d.obstacle = r->obstacle; d.follow_margin = cfg->follow_margin;
    d.left = cfg->left; d.right = cfg->right; d.bottom = cfg->bottom; d.top = cfg->top;
    d.violations = violations_for(device);
    *out = r;
    return F110_OK;
}

extern "C" int f110_render_destroy(f110_render *r)
{
    if (!r) return F110_OK;
    cudaSetDevice(r->device);
    cudaDeviceSynchronize();
    if (r->obstacle) cudaFree(r->obstacle);
    delete r;
    return F110_OK;
}

extern "C" int f110_render_frame(f110_render *r, const double *x_dev, const double *y_dev, const double *theta_dev, int64_t n_envs,
                                 uint8_t *out_dev, void *stream)
{
    if (!r) return fail(F110_ERR_INVALID, "null handle");
    if (n_envs == 0) return F110_OK;
    if (!x_dev || !y_dev || !theta_dev || !out_dev || n_envs < 0) return fail(F110_ERR_INVALID, "bad render arguments");
    CK(cudaSetDevice(r->device));
    f110_launch_render(r->d, x_dev, y_dev, theta_dev, n_envs, out_dev, (cudaStream_t)stream);
    CK(cudaGetLastError());
    return F110_OK;
}
