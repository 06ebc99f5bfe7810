// tools/log2_probe.cu — why the reference's class mc never returns at reSampleTimes = 64.
// /root/reference/src/mc.cu:101 maps a thread to its burst with
//     int idx = tidx >> ((int)log2((double)reSampleTimes));
// evaluated ON THE DEVICE.  CUDA's double log2 is good to 1 ulp, not exact: where it returns
// k - 1ulp for 2^k the truncation gives k - 1, idx runs to twice the number of bursts, and the
// threads beyond walk bursts that do not exist (garbage durations and rates: no loop exit).
// This prints (int)log2((double)2^k) as the device computes it, for k = 0..20.
//   nvcc -gencode arch=compute_100a,code=sm_100a -o tools/log2_probe.bin tools/log2_probe.cu
#include <cstdio>
#include <cmath>
__global__ void probe(int* out)
{
    int k = threadIdx.x;
    int R = 1 << k;
    out[k] = (int)log2((double)R);
}
int main()
{
    int* d; int h[21];
    cudaMalloc(&d, sizeof h);
    probe<<<1, 21>>>(d);
    if (cudaMemcpy(h, d, sizeof h, cudaMemcpyDeviceToHost) != cudaSuccess) { printf("no device\n"); return 1; }
    for (int k = 0; k <= 20; ++k)
        printf("R = 2^%-2d = %-8d device (int)log2 = %-2d host (int)log2 = %-2d %s\n", k, 1 << k, h[k], (int)log2((double)(1 << k)),
               h[k] == k ? "" : "<-- off by one: idx = tidx >> (k-1) reads past the bursts");
    return 0;
}
