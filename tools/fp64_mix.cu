// fp64_mix.cu — what does the FP64 pipe of an sm_100a SM sustain, by instruction shape?  31 FP64 instructions per
// group over 16 independent chains, plus (mode)
//   0: x = fma(x, a, imm) + one more DFMA     1: + MUFU.RSQ64H      2: + two LDS.128     3: + 8 FFMA + 8 IMAD
//   4: + MUFU.RSQ64H + two LDS.128 + integer (the force kernel's mix)                    5: + FP32 MUFU.RSQ
//   6: DFMA with three distinct register sources     7: DFMA x = u*u + x (two distinct register sources)
//   8: three sources, one shared by three consecutive DFMAs (operand reuse)    9: the same, order pinned with asm volatile
//   10: DADD x += u   11: DMUL x *= u   12: DFMA x = x*u + x   13: DFMA x = u*UR + x   14: DFMA x = x*u + UR
// Reports FP64 lane-instructions per clock per SM against the nominal 64.  Findings: profiles/r01_fp64_mix_v6.md.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/fp64_mix tools/fp64_mix.cu && tools/fp64_mix
#include <cstdio>
#include <cuda_runtime.h>

template <int MODE>
__global__ void __launch_bounds__(256) k_mix(double *out, int iters, double a, double b)
{
    __shared__ double2 sh[512];
    sh[threadIdx.x] = make_double2(a, b);
    sh[threadIdx.x + 256] = make_double2(b, a);
    __syncthreads();
    double x[16];
#pragma unroll
    for (int c = 0; c < 16; ++c) x[c] = (double)(threadIdx.x + c) * 1e-3 + 1.0;
    double extra = 0.0;
    double u[8], w[8];
#pragma unroll
    for (int c = 0; c < 8; ++c) w[c] = b + 1e-9 * (double)(threadIdx.x * 8 + c);
#pragma unroll
    for (int c = 0; c < 8; ++c) u[c] = (MODE >= 10 && MODE != 13 ? 1.0 : a) + 1e-12 * (double)(threadIdx.x * 8 + c);
    float fx = (float)threadIdx.x + 1.0f;
    unsigned ix = threadIdx.x;
    for (int it = 0; it < iters; ++it)
    {
#pragma unroll
        for (int g = 0; g < 4; ++g)
        {
            // 31 DFMAs over 16 independent chains
#pragma unroll
            for (int c = 0; c < 31; ++c)
            {
                if (MODE == 10) x[c % 16] = x[c % 16] + u[c % 8];                       // DADD, 2 distinct registers
                else if (MODE == 11) x[c % 16] = x[c % 16] * u[c % 8];                  // DMUL, 2 distinct registers
                else if (MODE == 12) x[c % 16] = fma(x[c % 16], u[c % 8], x[c % 16]);   // DFMA, 2 distinct, A == C
                else if (MODE == 13) x[c % 16] = fma(u[c % 8], b, x[c % 16]);           // DFMA, reg x uniform + reg
                else if (MODE == 14) x[c % 16] = fma(x[c % 16], u[c % 8], b);           // DFMA, reg x reg + uniform
                else if (MODE == 9)   // like 8, order pinned with volatile asm so that ptxas can flag the shared operand .reuse
                    asm volatile("fma.rn.f64 %0, %1, %2, %0;" : "+d"(x[c % 16]) : "d"(u[(c / 3) % 8]), "d"(w[c % 8]));
                else if (MODE == 8) x[c % 16] = fma(u[(c / 3) % 8], w[c % 8], x[c % 16]);            // 3 reg sources, slot A shared by 3 in a row
                else if (MODE == 6) x[c % 16] = fma(u[c % 8], u[(c + 3) % 8], x[c % 16]);        // 3 register sources
                else if (MODE == 7) x[c % 16] = fma(u[c % 8], u[c % 8], x[c % 16]);       // 2 distinct register sources
                else x[c % 16] = fma(x[c % 16], a, b);                                      // 1 register + 2 uniform
            }
            if (MODE == 6 || MODE == 7) u[g] += 1e-30 * x[g];
            if (MODE == 0) x[15] = fma(x[15], a, b);
            if (MODE == 1 || MODE == 4)
            {
                double y;
                asm volatile("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x[g]));
                extra += y;   // one more DADD (unavoidable consumer), counted in the report
            }
            if (MODE == 2 || MODE == 4)
            {
                double2 p, q;
                const unsigned ad = (unsigned)__cvta_generic_to_shared(&sh[(threadIdx.x + it + g) & 255]);
                asm volatile("ld.shared.v2.f64 {%0, %1}, [%2];" : "=d"(p.x), "=d"(p.y) : "r"(ad));
                asm volatile("ld.shared.v2.f64 {%0, %1}, [%2+4096];" : "=d"(q.x), "=d"(q.y) : "r"(ad));
                ix += (unsigned)__double2hiint(p.x) ^ (unsigned)__double2hiint(q.y);
            }
            if (MODE == 3)
            {
#pragma unroll
                for (int u = 0; u < 8; ++u)
                {
                    fx = fmaf(fx, 1.0001f, 0.5f);
                    ix = ix * 1664525u + 1013904223u;
                }
            }
            if (MODE == 5)
            {
                float y;
                asm volatile("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(fx));
                fx += y;
            }
        }
    }
    double s = extra + (double)fx + (double)ix;
    if (MODE >= 6) for (int c = 0; c < 8; ++c) s += u[c] + w[c];
#pragma unroll
    for (int c = 0; c < 16; ++c) s += x[c];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int MODE>
static void run(const char *name, int sms, double *out)
{
    const int iters = 4096;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    float best = 1e30f;
    const int blocks = sms * 4;
    for (int rep = 0; rep < 5; ++rep)
    {
        cudaEventRecord(e0);
        k_mix<MODE><<<blocks, 256>>>(out, iters, 1.0000001, 1e-9);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms;
        cudaEventElapsedTime(&ms, e0, e1);
        if (rep && ms < best) best = ms;
    }
    const double groups = 4.0 * iters * 256.0 * blocks;
    const double dfma = (MODE == 0 || MODE == 1 || MODE == 4 || MODE == 6 || MODE == 7 ? 32.0 : 31.0) * groups;   // FP64-pipe instr (lanes)
    const double per_clk_sm = dfma / (best * 1e-3) / sms / 1.965e9;
    printf("mode %d %-34s %8.3f ms  FP64 lane-instr/clk/SM at 1965 MHz = %6.2f (%.1f %% of 64)  cycles per group per SMSP-warp = %.2f\n",
           MODE, name, best, per_clk_sm, 100.0 * per_clk_sm / 64.0, (best * 1e-3 * 1.965e9) / (4.0 * iters * 8.0));
}

int main()
{
    cudaDeviceProp p;
    cudaGetDeviceProperties(&p, 0);
    double *out;
    cudaMalloc(&out, sizeof(double) * 256 * p.multiProcessorCount * 8);
    printf("%s, %d SMs\n", p.name, p.multiProcessorCount);
    run<0>("32 DFMA", p.multiProcessorCount, out);
    run<1>("31 DFMA + MUFU.RSQ64H (+DADD)", p.multiProcessorCount, out);
    run<2>("31 DFMA + 2 LDS.128", p.multiProcessorCount, out);
    run<3>("31 DFMA + 8 FFMA + 8 IMAD", p.multiProcessorCount, out);
    run<4>("31 DFMA + MUFU64H + 2 LDS + int", p.multiProcessorCount, out);
    run<5>("31 DFMA + MUFU.RSQ f32 (+FADD)", p.multiProcessorCount, out);
    run<6>("31 DFMA 3 reg sources + 1 DFMA", p.multiProcessorCount, out);
    run<7>("31 DFMA 2 reg sources + 1 DFMA", p.multiProcessorCount, out);
    run<8>("31 DFMA 3 reg, A shared x3 + 2", p.multiProcessorCount, out);
    run<9>("as 8, order pinned (asm volatile)", p.multiProcessorCount, out);
    run<10>("31 DADD x += u", p.multiProcessorCount, out);
    run<11>("31 DMUL x *= u", p.multiProcessorCount, out);
    run<12>("31 DFMA x = x*u + x", p.multiProcessorCount, out);
    run<13>("31 DFMA x = u*UR + x", p.multiProcessorCount, out);
    run<14>("31 DFMA x = x*u + UR", p.multiProcessorCount, out);
    run<0>("32 DFMA (again)", p.multiProcessorCount, out);
    if (cudaDeviceSynchronize() != cudaSuccess) { printf("CUDA error\n"); return 1; }
    return 0;
}
