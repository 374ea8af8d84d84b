// Spatial launch order of the lidar scans (sm_100a).
//
// The ray-cast is a chain of dependent 4/8-byte gathers from the EDT; what keeps it fast is that the cells a warp
// walks are already in its SM's L1.  Warps that run on one SM at the same time therefore should belong to cars
// that are close together and look the same way: they walk the same part of the map at the same time.  Right after
// a reset env order is track order and that holds by construction; after some hundred steps of random actions and
// auto-resets it does not, and the same kernel runs 1.3x slower (measured).  So before each ray-cast the
// agents are put in (chunk, map tile, heading octant) order by a counting sort over small integer keys:
//     k_order_count   key per agent, rank = atomicAdd(hist[key], 1)
//     k_order_scan    exclusive prefix over the bins (one CTA), clears hist for the next step
//     k_order_scatter order[base[key] + rank] = agent
// The ray-cast kernels then map warp w to agent order[w].  Only the schedule changes: every scan still writes
// its own rows, so results are bit-identical to the unordered launch (tests/test_gpu_step.py).
// The rank inside a bin comes from atomicAdd, so the order of the agents WITHIN a bin differs from run to run: results do not
// depend on it (tested), kernel timings vary by a fraction of a percent with it.
// fp32 env launches: k_order_count also prepares every agent's scan (f110_scan32_dev.h) and k_order_scatter moves the 32-byte
// records into launch order, so that the ray-cast's warps start from one coalesced load each.
#include <cuda_runtime.h>
#include "f110_dev.h"
#include "f110_scan32_dev.h"

// poses are read as x[i * stride], y[i * stride], yaw[i * stride]: the SoA state of a handle (stride 1) or an
// (n, 3) pose array handed to f110_raycast (stride 3)
__global__ void __launch_bounds__(256) k_order_count(const F110Dev d, const double *__restrict__ x, const double *__restrict__ y,
                                                     const double *__restrict__ yawp, int stride, int n, int agents, int chunk_envs,
                                                     int *__restrict__ keys, int *__restrict__ ranks, float4 *__restrict__ rec_tmp)
{
    int ia = blockIdx.x * blockDim.x + threadIdx.x;
    if (ia >= n) return;
    if (rec_tmp) {      // fp32 env launch: prepare the scan here, one thread per agent, instead of 8 threads per CTA of the ray-cast
        ScanJob j;
        prepare_env_scan<false>(d, ia, j);
        scan_rec_store(rec_tmp + 2 * (size_t)ia, j);
    }
    double px = x[(size_t)ia * stride], py = y[(size_t)ia * stride], yaw = yawp[(size_t)ia * stride];
    double xt = px - d.ox, yt = py - d.oy;
    float gx = (float)((xt * d.oc + yt * d.os) * d.inv_res);
    float gy = (float)((-xt * d.os + yt * d.oc) * d.inv_res);
    int tx = min(max(__float2int_rd(gx * d.order_inv_tile), 0), d.order_tx - 1);
    int ty = min(max(__float2int_rd(gy * d.order_inv_tile), 0), d.order_ty - 1);
    int hb = __float2int_rd((float)yaw * 1.2732395447351628f) & 7;        // 8 / (2 pi)
    int chunk = (ia / agents) / chunk_envs;
    int key = ((chunk * d.order_ty + ty) * d.order_tx + tx) * 8 + hb;
    F110_CHECK(d.violations, key >= 0 && key < d.order_bins, F110_V_ORDER);
    keys[ia] = key;
    // warp-aggregated: neighbours in env order are often neighbours on the track (same bin); one atomic per distinct key
    const unsigned lane = threadIdx.x & 31;
    const unsigned peers = __match_any_sync(__activemask(), key);
    const int leader = __ffs(peers) - 1;
    int first = 0;
    if ((int)lane == leader) first = atomicAdd(d.order_hist + key, __popc(peers));
    first = __shfl_sync(peers, first, leader);
    ranks[ia] = first + __popc(peers & ((1u << lane) - 1u));
}

// one CTA per chunk: exclusive prefix sum of that chunk's n bins into base (starting at the chunk's first agent,
// which is known because every chunk holds a fixed env range), hist cleared.  Each thread owns n/1024 consecutive bins.
// SMEM: the bins make a round trip through (padded) shared memory so that the global reads and writes are coalesced --
// thread-strided global accesses made this the longest of the three kernels (26 us for 20 000 bins, whatever the batch).
#define ORDER_PAD(i) ((i) + ((i) >> 5))
template <bool SMEM>
__global__ void __launch_bounds__(1024) k_order_scan(int *__restrict__ hist, int *__restrict__ base, int n, int agents_per_chunk)
{
    __shared__ int warp_sum[32];
    extern __shared__ int bins[];
    hist += (size_t)blockIdx.x * n;
    base += (size_t)blockIdx.x * n;
    const int t = threadIdx.x, lane = t & 31, wid = t >> 5;
    if (SMEM) {
        for (int i = t; i < n; i += 1024) bins[ORDER_PAD(i)] = hist[i];
        __syncthreads();
    }
    const int per = (n + 1023) / 1024;
    const int lo = min(t * per, n), hi = min(lo + per, n);
    int s = 0;
    for (int i = lo; i < hi; i++) s += SMEM ? bins[ORDER_PAD(i)] : hist[i];
    int inc = s;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        int v = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += v;
    }
    if (lane == 31) warp_sum[wid] = inc;
    __syncthreads();
    if (wid == 0) {
        int w = warp_sum[lane], wi = w;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            int v = __shfl_up_sync(0xffffffffu, wi, o);
            if (lane >= o) wi += v;
        }
        warp_sum[lane] = wi - w;
    }
    __syncthreads();
    int run = blockIdx.x * agents_per_chunk + warp_sum[wid] + inc - s;
    for (int i = lo; i < hi; i++) {
        if (SMEM) {
            int h = bins[ORDER_PAD(i)];
            bins[ORDER_PAD(i)] = run;
            run += h;
        } else {
            int h = hist[i];
            base[i] = run;
            hist[i] = 0;
            run += h;
        }
    }
    if (SMEM) {
        __syncthreads();
        for (int i = t; i < n; i += 1024) { base[i] = bins[ORDER_PAD(i)]; hist[i] = 0; }
    }
}

static void launch_order_scan(int chunks, int *hist, int *base, int n, int agents_per_chunk, cudaStream_t st)
{
    const size_t bytes = ((size_t)n + (size_t)n / 32 + 2) * 4;
    if (bytes <= 200 * 1024) {
        // per-device attribute, cheap: set on every launch rather than tracking which devices have seen it
        cudaFuncSetAttribute(k_order_scan<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
        k_order_scan<true><<<chunks, 1024, bytes, st>>>(hist, base, n, agents_per_chunk);
    } else {
        k_order_scan<false><<<chunks, 1024, 0, st>>>(hist, base, n, agents_per_chunk);
    }
}

__global__ void __launch_bounds__(256) k_order_scatter(const int *__restrict__ base, const int *__restrict__ keys,
                                                       const int *__restrict__ ranks, int *__restrict__ order, int n, unsigned long long *viol,
                                                       const float4 *__restrict__ rec_tmp, float4 *__restrict__ rec)
{
    int ia = blockIdx.x * blockDim.x + threadIdx.x;
    if (ia >= n) return;
    const int slot = base[keys[ia]] + ranks[ia];
    F110_CHECK(viol, slot >= 0 && slot < n, F110_V_ORDER);
    order[slot] = ia;
    if (rec) {          // the scan records move with the agents: the ray-cast reads them in launch order
        rec[2 * (size_t)slot] = rec_tmp[2 * (size_t)ia];
        rec[2 * (size_t)slot + 1] = rec_tmp[2 * (size_t)ia + 1];
    }
}

void f110_launch_order(const F110Dev &d, cudaStream_t st)
{
    const int NA = d.n_envs * d.n_agents;
    k_order_count<<<(NA + 255) / 256, 256, 0, st>>>(d, d.state, d.state + NA, d.state + 4 * (size_t)NA, 1, NA, d.n_agents,
                                                    d.order_chunk_envs, d.order_key, d.order_rank, d.rec32 ? d.rec32_tmp : nullptr);
    const int chunks = (d.n_envs + d.order_chunk_envs - 1) / d.order_chunk_envs;
    launch_order_scan(chunks, d.order_hist, d.order_base, d.order_tx * d.order_ty * 8, d.order_chunk_envs * d.n_agents, st);
    k_order_scatter<<<(NA + 255) / 256, 256, 0, st>>>(d.order_base, d.order_key, d.order_rank, d.order, NA, d.violations,
                                                      d.rec32 ? d.rec32_tmp : nullptr, d.rec32);
}

// order of n stand-alone poses (f110_raycast): one chunk; buf = order | key | rank, n each
void f110_launch_order_poses(const F110Dev &d, const double *poses, int n, int *buf, cudaStream_t st)
{
    int *order = buf, *keys = buf + n, *ranks = buf + 2 * (size_t)n;
    k_order_count<<<(n + 255) / 256, 256, 0, st>>>(d, poses, poses + 1, poses + 2, 3, n, 1, n, keys, ranks, nullptr);
    launch_order_scan(1, d.order_hist, d.order_base, d.order_tx * d.order_ty * 8, 0, st);
    k_order_scatter<<<(n + 255) / 256, 256, 0, st>>>(d.order_base, keys, ranks, order, n, d.violations, nullptr, nullptr);
}
