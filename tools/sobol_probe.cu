// tools/sobol_probe.cu — diagnostic: what cuRAND's scrambled Sobol64 hands the reference's
// setup_kernel/mc_kernel for a few thread ids (raw 64-bit draws after curand_init(..., 1234)).
#include <cstdio>
#include <curand.h>
#include <curand_kernel.h>
__global__ void probe(unsigned long long* dv, unsigned long long* sc, const int* ids, int n, unsigned long long* out, float* fo)
{
    int k = threadIdx.x;
    if (k >= n) return;
    int idx = ids[k];
    curandStateScrambledSobol64 s;
    curand_init(dv + 64 * (idx % 20000), sc[idx % 20000], 1234, &s);
    for (int j = 0; j < 4; ++j) out[k * 4 + j] = curand(&s);
    fo[k * 3 + 0] = curand_uniform(&s);
    fo[k * 3 + 1] = curand_normal(&s);
    fo[k * 3 + 2] = (float)curand_poisson(&s, 3.5);
}
int main()
{
    curandDirectionVectors64_t* hv; unsigned long long* hs;
    curandGetDirectionVectors64(&hv, CURAND_SCRAMBLED_DIRECTION_VECTORS_64_JOEKUO6);
    curandGetScrambleConstants64(&hs);
    unsigned long long *dv, *sc, *out; int* ids; float* fo;
    int h_ids[6] = {0, 1, 3, 12, 20000, 20003};
    cudaMalloc(&dv, 20000 * 64 * 8); cudaMalloc(&sc, 20000 * 8); cudaMalloc(&out, 6 * 4 * 8); cudaMalloc(&ids, 24); cudaMalloc(&fo, 6 * 3 * 4);
    cudaMemcpy(dv, hv, 20000 * 64 * 8, cudaMemcpyHostToDevice); cudaMemcpy(sc, hs, 20000 * 8, cudaMemcpyHostToDevice);
    cudaMemcpy(ids, h_ids, 24, cudaMemcpyHostToDevice);
    probe<<<1, 32>>>(dv, sc, ids, 6, out, fo);
    unsigned long long h_out[24]; float h_fo[18];
    cudaMemcpy(h_out, out, sizeof(h_out), cudaMemcpyDeviceToHost); cudaMemcpy(h_fo, fo, sizeof(h_fo), cudaMemcpyDeviceToHost);
    for (int k = 0; k < 6; ++k)
        printf("%d %016llx %016llx %016llx %016llx u=%.9g z=%.9g pois=%g\n", h_ids[k], h_out[4 * k], h_out[4 * k + 1], h_out[4 * k + 2], h_out[4 * k + 3], h_fo[3 * k], h_fo[3 * k + 1], h_fo[3 * k + 2]);
    return 0;
}
