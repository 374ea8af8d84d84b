// Exact Euclidean distance transform on the GPU: replaces the host-side
// `resolution * scipy.ndimage.distance_transform_edt(bitmap)` of get_dt (laser_models.py:40-53) so that update_map /
// per-episode map randomisation does not wait seconds on the CPU.  Result is bit-identical to scipy's: scipy computes
// the exact integer squared distance to the nearest zero pixel and takes a float64 sqrt; so does this (separable:
// nearest obstacle per column, then per row the minimum of dx^2 + g(x + dx)^2, searched outwards with the exact bound
// dx^2 < best), and IEEE sqrt / multiply are correctly rounded on both sides.
#include <cuda_runtime.h>
#include <stdint.h>

#define EDT_INF 65535u      // "no obstacle in this column"; 65535^2 still fits 32 bits

// free: [h][w] bytes, non-zero = free space (pixel > 128 in set_map); g: [h][w] distance (rows) to the nearest obstacle
// in the same column
__global__ void k_edt_columns(const uint8_t *__restrict__ free_px, unsigned *__restrict__ g, int h, int w)
{
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
    if (x >= w) return;
    unsigned run = EDT_INF;
    for (int y = 0; y < h; y++) {
        run = free_px[(size_t)y * w + x] ? (run == EDT_INF ? EDT_INF : run + 1) : 0u;
        g[(size_t)y * w + x] = run;
    }
    run = EDT_INF;
    for (int y = h - 1; y >= 0; y--) {
        run = free_px[(size_t)y * w + x] ? (run == EDT_INF ? EDT_INF : run + 1) : 0u;
        const unsigned up = g[(size_t)y * w + x];
        g[(size_t)y * w + x] = run < up ? run : up;
    }
}

// one CTA per row; the row's g^2 in shared memory
__global__ void __launch_bounds__(256) k_edt_rows(const unsigned *__restrict__ g, double *__restrict__ edt, int h, int w, double res)
{
    extern __shared__ unsigned g2[];
    const int y = blockIdx.x;
    for (int x = threadIdx.x; x < w; x += blockDim.x) {
        const unsigned v = g[(size_t)y * w + x];
        g2[x] = v * v;
    }
    __syncthreads();
    for (int x = threadIdx.x; x < w; x += blockDim.x) {
        unsigned best = g2[x];
        for (unsigned dx = 1; dx * dx < best; dx++) {
            const unsigned d2 = dx * dx;
            if ((int)dx > x && x + (int)dx >= w) break;       // both sides have left the row
            if (x >= (int)dx) { const unsigned gl = g2[x - dx]; if (gl < best - d2) best = gl + d2; }      // gl + d2 < best, no overflow
            if (x + (int)dx < w) { const unsigned gr = g2[x + dx]; if (gr < best - d2) best = gr + d2; }
        }
        edt[(size_t)y * w + x] = res * sqrt((double)best);
    }
}

void f110_launch_edt_build(const uint8_t *free_px, unsigned *scratch_g, double *edt, int h, int w, double res, cudaStream_t st)
{
    k_edt_columns<<<(w + 127) / 128, 128, 0, st>>>(free_px, scratch_g, h, w);
    const size_t smem = (size_t)w * 4;
    if (smem > 48 * 1024) cudaFuncSetAttribute(k_edt_rows, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    k_edt_rows<<<h, 256, smem, st>>>(scratch_g, edt, h, w, res);
}

// ------------------------------------------------------------------------------------------------
// f110_edt_build_noalloc: the same two kernels for callers that cannot call cudaMalloc (code under CUDA graph capture,
// tight serving loops, harnesses that forbid it).  No counterpart in the reference.
// The column-distance scratch `g` that f110_edt_build allocates per call lives here in a file-scope device array
// of a fixed, modest size: F110_EDT_NOALLOC_MAX_CELLS unsigned = 16 MiB, held by every GPU that loads the library until exit.
// Every call on a device shares that array: it keeps the last call's contents (harmless -- k_edt_columns writes every cell
// of [h][w] before k_edt_rows reads it, nothing has to start at zero) and two calls that overlap on the GPU (different
// streams, no ordering between them) race on it.
// In host code `f110_edt_scratch` is a host-side stand-in, not a device address: its device address is asked for once per
// device, with that device current, by cudaGetSymbolAddress, and that pointer is what the kernels get.
// ------------------------------------------------------------------------------------------------
#define F110_EDT_NOALLOC_MAX_CELLS_ (2048 * 2048)      /* == F110_EDT_NOALLOC_MAX_CELLS of include/f110_b200.h */
#define F110_EDT_NOALLOC_MAX_WIDTH_ 16384              /* k_edt_rows keeps a row of g^2 in shared memory: 64 KB */
static __device__ __align__(256) unsigned f110_edt_scratch[F110_EDT_NOALLOC_MAX_CELLS_];

#include <mutex>
static std::mutex g_scratch_mutex;
static unsigned *g_scratch_ptr[64];       // per device ordinal: device address of f110_edt_scratch there

// 0 = ok; otherwise *what names the CUDA call that failed and *err its code.  The caller has made `device` current.
// Calls no cudaMalloc / cudaFree / allocating routine.
int f110_edt_build_static(const uint8_t *free_px, double *edt, int h, int w, double res, int device, cudaStream_t st,
                          const char **what, cudaError_t *err)
{
    *what = ""; *err = cudaSuccess;
    if (device < 0 || device >= 64 || (long long)h * w > F110_EDT_NOALLOC_MAX_CELLS_ || w > F110_EDT_NOALLOC_MAX_WIDTH_) {
        *what = "size check"; *err = cudaErrorInvalidValue;
        return 1;
    }
    unsigned *g;
    {
        std::lock_guard<std::mutex> lock(g_scratch_mutex);
        if (!g_scratch_ptr[device]) {      // set-up, once per device, with that device current
            unsigned *p = nullptr;
            cudaError_t e = cudaGetSymbolAddress((void **)&p, f110_edt_scratch);
            if (e != cudaSuccess || !p) { *what = "cudaGetSymbolAddress(f110_edt_scratch)"; *err = e != cudaSuccess ? e : cudaErrorInvalidSymbol; return 1; }
            g_scratch_ptr[device] = p;
        }
        g = g_scratch_ptr[device];
    }
    k_edt_columns<<<(w + 127) / 128, 128, 0, st>>>(free_px, g, h, w);
    const size_t smem = (size_t)w * 4;
    if (smem > 48 * 1024) {
        cudaError_t e = cudaFuncSetAttribute(k_edt_rows, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) { *what = "cudaFuncSetAttribute(k_edt_rows)"; *err = e; return 1; }
    }
    k_edt_rows<<<h, 256, smem, st>>>(g, edt, h, w, res);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) { *what = "EDT kernel launch"; *err = e; return 1; }
    return 0;
}
