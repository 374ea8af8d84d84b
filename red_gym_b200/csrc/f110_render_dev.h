// Device-side description of one headless renderer handle (f110_render.cu).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

struct F110RenderDev {
    int width, height;          // window size in pixels (WINDOW_W x WINDOW_H = 1000 x 800, f110_env.py:50-51)
    int n_agents, ego_idx;
    int map_h, map_w;
    double res, ox, oy;         // map resolution and origin (rendering.py:123-129)
    const uint8_t *obstacle;    // [map_h][map_w] 1 where the flipped map image is exactly 0 (rendering.py:150)
    double follow_margin;       // >= 0: camera = ego car's bounding box +- margin (examples/waypoint_follow.py:259-269)
    double left, right, bottom, top;   // fixed camera otherwise (glOrtho, rendering.py:290)
    unsigned long long *violations;    // the device's F110_V_* counters (checked build only: f110_dev.h)
};

void f110_launch_render(const F110RenderDev &d, const double *x, const double *y, const double *theta, long long n_envs,
                        uint8_t *out, cudaStream_t st);
