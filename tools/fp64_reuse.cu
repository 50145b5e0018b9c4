// fp64_reuse.cu — does the operand reuse cache lift the 3/4 rate of three-source DFMAs (tools/fp64_mix.cu) when the
// same warp issues back to back?  48 DFMAs per group over 24 independent accumulator chains; 1, 2, 4 or 8 warps per
// scheduler (CTAs of 128 or 256 threads, 1 to 4 per SM).  Patterns (A, B = multiplicands, C = accumulator, all registers):
//   0: A, B, C all distinct, nothing shared between neighbours
//   1: A shared by 3 consecutive DFMAs          2: A shared by 6 consecutive DFMAs (the force kernel's m/r^3)
//   3: two distinct sources (x = u*u + x)       4: B shared by 6 consecutive DFMAs
//   5: A shared by 6, B shared by the same 6 pairwise (A,B0) (A,B0') ... i.e. A and C-chain distinct, B distinct
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/fp64_reuse tools/fp64_reuse.cu && tools/fp64_reuse
#include <cstdio>
#include <cuda_runtime.h>

constexpr int NCH = 24;   // accumulator chains
constexpr int NG = 48;    // DFMAs per group

template <int PAT>
__global__ void __launch_bounds__(256) k_reuse(double *out, int iters, double a, double b)
{
    double x[NCH], u[8], w[8];
#pragma unroll
    for (int c = 0; c < NCH; ++c) x[c] = (double)(threadIdx.x + c) * 1e-3 + 1.0;
#pragma unroll
    for (int c = 0; c < 8; ++c) u[c] = a + 1e-12 * (double)(threadIdx.x * 8 + c);
#pragma unroll
    for (int c = 0; c < 8; ++c) w[c] = b + 1e-9 * (double)(threadIdx.x * 8 + c);
    for (int it = 0; it < iters; ++it)
    {
#pragma unroll
        for (int c = 0; c < NG; ++c)
        {
            double &acc = x[c % NCH];
            if (PAT == 0) acc = fma(u[c % 8], w[(c + 3) % 8], acc);
            else if (PAT == 1) acc = fma(u[(c / 3) % 8], w[c % 8], acc);
            else if (PAT == 2) acc = fma(u[(c / 6) % 8], w[c % 8], acc);
            else if (PAT == 3) acc = fma(u[c % 8], u[c % 8], acc);
            else if (PAT == 4) acc = fma(w[c % 8], u[(c / 6) % 8], acc);
            else acc = fma(u[(c / 6) % 8], w[(c / 2) % 8], acc);
        }
        // keep u, w loop-variant so that nothing is hoisted, at negligible cost
        u[0] += 1e-30 * x[0];
        w[0] += 1e-30 * x[1];
    }
    double s = 0.0;
#pragma unroll
    for (int c = 0; c < 8; ++c) s += u[c] + w[c];
#pragma unroll
    for (int c = 0; c < NCH; ++c) s += x[c];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int PAT>
static void run(const char *name, int sms, double *out)
{
    const int iters = 32768;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    printf("pattern %d %-40s cycles per DFMA per scheduler:", PAT, name);
    for (int wps = 1; wps <= 8; wps *= 2)
    {
        float best = 1e30f;
        for (int rep = 0; rep < 4; ++rep)
        {
            cudaEventRecord(e0);
            k_reuse<PAT><<<sms * (wps > 2 ? wps / 2 : 1), wps > 1 ? 256 : 128>>>(out, iters, 1.0000001, 1e-9);
            cudaEventRecord(e1);
            cudaEventSynchronize(e1);
            float ms;
            cudaEventElapsedTime(&ms, e0, e1);
            if (rep && ms < best) best = ms;
        }
        const double cyc = best * 1e-3 * 1.965e9 / ((double)iters * NG * wps);
        printf("  %dw %.3f", wps, cyc);
    }
    printf("\n");
}

int main()
{
    cudaDeviceProp p;
    cudaGetDeviceProperties(&p, 0);
    double *out;
    cudaMalloc(&out, sizeof(double) * 1024 * p.multiProcessorCount);
    printf("%s, %d SMs; nominal 2.000 cycles per DFMA per scheduler\n", p.name, p.multiProcessorCount);
    run<3>("two distinct sources", p.multiProcessorCount, out);
    run<0>("three sources, nothing shared", p.multiProcessorCount, out);
    run<1>("three sources, A shared x3", p.multiProcessorCount, out);
    run<2>("three sources, A shared x6", p.multiProcessorCount, out);
    run<4>("three sources, B shared x6", p.multiProcessorCount, out);
    run<5>("three sources, A shared x6 and B x2", p.multiProcessorCount, out);
    if (cudaDeviceSynchronize() != cudaSuccess) { printf("CUDA error\n"); return 1; }
    return 0;
}
