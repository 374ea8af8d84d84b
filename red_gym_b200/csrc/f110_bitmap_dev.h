// Device-side description of one rasteriser handle (f110_bitmap.cu), passed to the kernel by value.
#pragma once
#include <stdint.h>

enum { F110_BM_FILL = 0, F110_BM_POLYGON = 1, F110_BM_RAYS = 2 };

struct F110BitmapDev {
    int n_beams;          // beams per scan (row length of the input)
    int n_points;         // target_beam_count
    int height, width;    // output_image_dims
    int wpr;              // 32-bit mask words per row
    int draw_mode;        // F110_BM_*
    int draw_center;
    int scan_is_f32;
    int cx, cy;           // output_image_dims[0] // 2, [1] // 2 (the reference's centre: x from dims[0], y from dims[1])
    unsigned char bg, fg;
    double scale;         // scaling_factor
    float scale32;        // float32(scaling_factor): numpy multiplies a float32 scan by it in float32
    const int *indices;   // [n_points] np.linspace(0, B - 1, T, dtype=int)
    const double *cosa, *sina;   // [n_points] np.cos / np.sin of the beam angles
    unsigned long long *violations;   // the device's F110_V_* counters (checked build only: f110_dev.h)
};

void f110_launch_bitmap(const F110BitmapDev &d, const void *scans, long long n, uint8_t *out, int packed, cudaStream_t st);
size_t f110_bitmap_smem_bytes(const F110BitmapDev &d);
