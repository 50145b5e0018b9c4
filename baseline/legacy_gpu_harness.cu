// legacy_gpu_harness.cu — the reference's OWN CUDA kernels, recompiled for sm_100a, as a same-box data point.
//
// MEASUREMENT INFRASTRUCTURE ONLY; nothing in gravidy_b200/ links or loads this.  The kernels
// are compiled from the file where it lies, /root/reference/src/include/gpu/Hermite4GPU_kernels.cu (no copy in
// this repo), with `-DGPU` as the reference's Makefile does (src/Makefile:69) and access checks defeated the
// only way cudafe++ allows (the class header re-exports private members, Hermite4GPU.cuh:105-109).  This file
// adds what Hermite4GPU.cu does AROUND the kernels: the launch shapes of init_acc_jrk (Hermite4GPU.cu:333-375:
// BSIZE-thread CTAs, one i per thread) and of update_acc_jrk (Hermite4GPU.cu:385-535: grid (nact/BSIZE, NJBLOCK),
// then k_reduce over the NJBLOCK partials, in chunks of 32768 active particles), timed with CUDA events around
// the kernels only (the reference also moves the particle set over PCIe every step; that is left out, in
// its favour).
//
// Build (baseline/Makefile): nvcc -DGPU -gencode arch=compute_100a,code=sm_100a -std=c++03 ... -> baseline/_ref/liblegacy_gpu.so
#define private public
#define protected public
#include "gpu/Hermite4GPU_kernels.cu"
#undef private
#undef protected

#include <cstdio>
#include <vector>

#define LG_CHECK(call)                                                                             \
    do {                                                                                           \
        cudaError_t e_ = (call);                                                                   \
        if (e_ != cudaSuccess) {                                                                   \
            fprintf(stderr, "legacy_gpu: %s line %d: %s\n", #call, __LINE__, cudaGetErrorString(e_)); \
            return -1;                                                                             \
        }                                                                                          \
    } while (0)

// pred7: n x {x y z vx vy vz m}.  move == NULL: k_init_acc_jrk on all n particles (nact ignored);
// otherwise k_update + k_reduce on the nact listed particles.  f6_out (may be NULL): nact (or n) x 6.
// ms_out[0] = best kernel time of `reps` repetitions, ms_out[1] = the mean.
extern "C" int legacy_gpu_forces(int n, double e2, const double *pred7, int nact, const unsigned *move,
                                 double *f6_out, int reps, float *ms_out)
{
    if (n <= 0 || n % BSIZE != 0 || !pred7 || reps < 1) return -2;   // the kernels assume whole BSIZE tiles
    const bool init = move == NULL;
    if (init) nact = n;
    if (nact <= 0) return -2;
    const int pad = 4 * BSIZE;   // k_update reads up to 2*BSIZE entries past its j-range (kernels.cu:193-194)
    std::vector<Predictor> hp(n + pad), hi(nact + pad);
    for (int i = 0; i < n + pad; ++i)
    {
        Predictor p;
        const bool real = i < n;
        for (int c = 0; c < 3; ++c) { p.r[c] = real ? pred7[7 * (size_t)i + c] : 1.0e10; p.v[c] = real ? pred7[7 * (size_t)i + 3 + c] : 0.0; }
        p.m = real ? (float)pred7[7 * (size_t)i + 6] : 0.0f;
        hp[i] = p;
    }
    for (int k = 0; k < nact + pad; ++k) hi[k] = init ? hp[k < n ? k : n - 1] : hp[k < nact ? move[k] : move[nact - 1]];

    Predictor *d_p = NULL, *d_i = NULL;
    Forces *d_f = NULL, *d_fout = NULL, *d_fout_tmp = NULL;
    const size_t nact_pad = (size_t)(1 + (nact - 1) / BSIZE) * BSIZE;
    LG_CHECK(cudaMalloc(&d_p, (n + pad) * sizeof(Predictor)));
    LG_CHECK(cudaMemcpy(d_p, hp.data(), (n + pad) * sizeof(Predictor), cudaMemcpyHostToDevice));
    cudaEvent_t e0, e1;
    LG_CHECK(cudaEventCreate(&e0));
    LG_CHECK(cudaEventCreate(&e1));
    const size_t smem = sizeof(Predictor) * BSIZE, smem_reduce = sizeof(Forces) * NJBLOCK + 1;   // Hermite4GPU.cu:46-47
    float best = 1e30f, sum = 0.0f;
    if (init)
    {
        LG_CHECK(cudaMalloc(&d_f, (size_t)n * sizeof(Forces)));
        for (int rep = 0; rep < reps + 1; ++rep)
        {
            LG_CHECK(cudaEventRecord(e0));
            k_init_acc_jrk<<<n / BSIZE, BSIZE, smem>>>(d_p, d_f, n, e2, 0, n);
            LG_CHECK(cudaEventRecord(e1));
            LG_CHECK(cudaEventSynchronize(e1));
            LG_CHECK(cudaGetLastError());
            float ms;
            LG_CHECK(cudaEventElapsedTime(&ms, e0, e1));
            if (rep) { sum += ms; if (ms < best) best = ms; }
        }
        if (f6_out) LG_CHECK(cudaMemcpy(f6_out, d_f, (size_t)n * sizeof(Forces), cudaMemcpyDeviceToHost));
    }
    else
    {
        LG_CHECK(cudaMalloc(&d_i, (nact + pad) * sizeof(Predictor)));
        LG_CHECK(cudaMemcpy(d_i, hi.data(), (nact + pad) * sizeof(Predictor), cudaMemcpyHostToDevice));
        LG_CHECK(cudaMalloc(&d_fout, nact_pad * NJBLOCK * sizeof(Forces)));
        LG_CHECK(cudaMalloc(&d_fout_tmp, nact_pad * sizeof(Forces)));
        for (int rep = 0; rep < reps + 1; ++rep)
        {
            LG_CHECK(cudaEventRecord(e0));
            dim3 nblocks(1 + (nact - 1) / BSIZE, NJBLOCK, 1), nthreads(BSIZE, 1, 1);
            k_update<<<nblocks, nthreads, smem>>>(d_i, d_p, d_fout, n, nact, e2);
            const size_t chunk = 2 << 14;
            unsigned int shift = 0;
            for (unsigned int s = 0; shift < (unsigned)nact; s++)
            {
                const size_t size_launch = (size_t)nact < shift + chunk ? nact - shift : chunk;
                k_reduce<<<dim3(size_launch, 1, 1), dim3(NJBLOCK, 1, 1), smem_reduce>>>(d_fout, d_fout_tmp + shift, s, shift);
                shift += chunk;
            }
            LG_CHECK(cudaEventRecord(e1));
            LG_CHECK(cudaEventSynchronize(e1));
            LG_CHECK(cudaGetLastError());
            float ms;
            LG_CHECK(cudaEventElapsedTime(&ms, e0, e1));
            if (rep) { sum += ms; if (ms < best) best = ms; }
        }
        if (f6_out) LG_CHECK(cudaMemcpy(f6_out, d_fout_tmp, (size_t)nact * sizeof(Forces), cudaMemcpyDeviceToHost));
    }
    if (ms_out) { ms_out[0] = best; ms_out[1] = sum / reps; }
    cudaFree(d_p); cudaFree(d_i); cudaFree(d_f); cudaFree(d_fout); cudaFree(d_fout_tmp);
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    return 0;
}
