// prints the set of %smid values a full-occupancy launch sees (is the id space dense on this part?)
#include <cstdio>
#include <cuda_runtime.h>
__global__ void k(int *seen) { unsigned s, n; asm volatile("mov.u32 %0, %%smid;" : "=r"(s)); asm volatile("mov.u32 %0, %%nsmid;" : "=r"(n)); if (threadIdx.x == 0) { atomicAdd(&seen[s], 1); seen[1023] = (int)n; } for (volatile int i = 0; i < 20000; i++); }
int main() { int *d; cudaMalloc(&d, 4096); cudaMemset(d, 0, 4096); k<<<148 * 8, 256>>>(d); int h[1024]; cudaMemcpy(h, d, 4096, cudaMemcpyDeviceToHost); int mx = -1, cnt = 0; for (int i = 0; i < 1023; i++) if (h[i]) { mx = i; cnt++; } printf("nsmid %d distinct %d max %d\n", h[1023], cnt, mx); for (int i = 0; i <= mx; i++) printf("%d:%d ", i, h[i]); printf("\n"); return 0; }
