// Device-side description of one pure-pursuit planner handle (f110_pursuit.cu).
#pragma once
#include <cuda_runtime.h>

struct F110PursuitDev {
    int n_waypoints;
    double wheelbase, max_reacquire;
    const double *wx, *wy, *wv;      // [n_waypoints] x, y, speed columns of the raceline file
    unsigned long long *violations;  // the device's F110_V_* counters (checked build only: f110_dev.h)
};

void f110_launch_pursuit(const F110PursuitDev &d, const double *x, const double *y, const double *theta, long long n,
                         long long stride, double lookahead, double vgain, const double *lookahead_q, const double *vgain_q,
                         double *actions, int *picked, cudaStream_t st);
