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
