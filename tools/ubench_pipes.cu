// ubench_pipes.cu -- per-SM issue throughput of the instruction classes the reference-arithmetic kernels
// are made of (FP64 arithmetic, Float32 <-> Float64 conversions, int -> Float64, Float32 FMA, 32x32->64
// multiply, LOP3), alone and mixed, on the GPU at hand.  One CTA of 1024 threads per SM, 8 independent
// dependency chains per thread, so every pipe runs at its issue limit, not at its latency.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o ubench_pipes ubench_pipes.cu && ./ubench_pipes
// Prints thread-instructions per clock per SM (128 = one warp instruction per scheduler per clock).
#include <cuda_runtime.h>

#include <cstdio>
#include <cstring>

constexpr int ITERS = 2048;
constexpr int CHAINS = 8;

enum Op { DFMA, DADD, DMUL, F2D, D2F, I2D, FFMA_, IMADW, LOP, MIX_DFMA_FFMA, MIX_DFMA_F2D, MIX_DFMA_I2D, MIX_F2D_D2F, MAGIC_I2D, DSETP_SEL, BASE_DFMA_FADD, PLUS_NARROW, PLUS_WIDEN, PLUS_I2D, PLUS_MAGIC, NOPS };
static const char* kNames[] = {"DFMA", "DADD", "DMUL", "F2F.F64.F32 (widen)", "F2F.F32.F64 (narrow)", "I2F.F64.S32", "FFMA",
                               "IMAD.WIDE.U32", "LOP3", "DFMA + FFMA (1:1)", "DFMA + widen (1:1)", "DFMA + I2F.F64 (1:1)",
                               "widen + narrow (1:1)", "magic int->double (MOV+DADD)", "DSETP + 2 FSEL",
                               "DFMA + FADD (baseline)", "DFMA + FADD + narrow(d)", "DFMA(+widen(f)) + FADD", "DFMA(+I2F.F64(u)) + FADD + LOP3",
                               "DFMA(+magic(u): DADD) + FADD + LOP3"};

template <int OP>
__global__ void __launch_bounds__(1024) k(long long* cycles, double* sink, int seed) {
    double d[CHAINS];
    float f[CHAINS];
    unsigned u[CHAINS];
    for (int c = 0; c < CHAINS; ++c) { d[c] = 1.0 + 1e-9 * (threadIdx.x + c + seed); f[c] = 1.0f + 1e-3f * (threadIdx.x + c); u[c] = threadIdx.x * 2654435761u + c + seed; }
    const double da = 1.0000001, db = 1e-12;
    __syncthreads();
    const long long t0 = clock64();
#pragma unroll 1
    for (int i = 0; i < ITERS; ++i) {
#pragma unroll
        for (int c = 0; c < CHAINS; ++c) {
            if (OP == DFMA) asm volatile("fma.rn.f64 %0, %0, %1, %2;" : "+d"(d[c]) : "d"(da), "d"(db));
            if (OP == DADD) asm volatile("add.rn.f64 %0, %0, %1;" : "+d"(d[c]) : "d"(db));
            if (OP == DMUL) asm volatile("mul.rn.f64 %0, %0, %1;" : "+d"(d[c]) : "d"(da));
            if (OP == F2D) { asm volatile("cvt.f64.f32 %0, %1;" : "=d"(d[c]) : "f"(f[c])); asm volatile("" : "+f"(f[c])); }
            if (OP == D2F) { asm volatile("cvt.rn.f32.f64 %0, %1;" : "=f"(f[c]) : "d"(d[c])); asm volatile("" : "+d"(d[c])); }
            if (OP == I2D) { asm volatile("cvt.rn.f64.s32 %0, %1;" : "=d"(d[c]) : "r"(u[c])); asm volatile("" : "+r"(u[c])); }
            if (OP == FFMA_) asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(f[c]) : "f"(1.0001f), "f"(1e-6f));
            if (OP == IMADW) { unsigned long long w; asm volatile("mul.wide.u32 %0, %1, %2;" : "=l"(w) : "r"(u[c]), "r"(0xD2511F53u)); u[c] = (unsigned)(w >> 32) ^ (unsigned)w; }
            if (OP == LOP) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(u[c]) : "r"(0x9E3779B9u), "r"(i));
            if (OP == MIX_DFMA_FFMA) { asm volatile("fma.rn.f64 %0, %0, %1, %2;" : "+d"(d[c]) : "d"(da), "d"(db)); asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(f[c]) : "f"(1.0001f), "f"(1e-6f)); }
            if (OP == MIX_DFMA_F2D) { double w; asm volatile("cvt.f64.f32 %0, %1;" : "=d"(w) : "f"(f[c])); asm volatile("fma.rn.f64 %0, %0, %1, %2;" : "+d"(d[c]) : "d"(da), "d"(w)); }
            if (OP == MIX_DFMA_I2D) { double w; asm volatile("cvt.rn.f64.s32 %0, %1;" : "=d"(w) : "r"(u[c])); asm volatile("fma.rn.f64 %0, %0, %1, %2;" : "+d"(d[c]) : "d"(da), "d"(w)); }
            if (OP == MIX_F2D_D2F) { double w; asm volatile("cvt.f64.f32 %0, %1;" : "=d"(w) : "f"(f[c])); asm volatile("cvt.rn.f32.f64 %0, %1;" : "=f"(f[c]) : "d"(w)); }
            if (OP == MAGIC_I2D) { double w = __hiloint2double(0x43300000, (int)u[c]); asm volatile("add.rn.f64 %0, %1, %2;" : "=d"(d[c]) : "d"(w), "d"(-4503601774854144.0)); asm volatile("" : "+r"(u[c])); }
            // loop-carried variants: every conversion feeds the next iteration, nothing can be dropped
            if (OP == BASE_DFMA_FADD) { asm volatile("fma.rn.f64 %0, %0, %1, %2;" : "+d"(d[c]) : "d"(da), "d"(db)); asm volatile("add.rn.f32 %0, %0, %1;" : "+f"(f[c]) : "f"(1e-6f)); }
            if (OP == PLUS_NARROW) { float t; asm volatile("fma.rn.f64 %0, %0, %1, %2;" : "+d"(d[c]) : "d"(da), "d"(db)); asm volatile("cvt.rn.f32.f64 %0, %1;" : "=f"(t) : "d"(d[c])); asm volatile("add.rn.f32 %0, %0, %1;" : "+f"(f[c]) : "f"(t)); }
            if (OP == PLUS_WIDEN) { double w; asm volatile("cvt.f64.f32 %0, %1;" : "=d"(w) : "f"(f[c])); asm volatile("fma.rn.f64 %0, %0, %1, %2;" : "+d"(d[c]) : "d"(da), "d"(w)); asm volatile("add.rn.f32 %0, %0, %1;" : "+f"(f[c]) : "f"(1e-6f)); }
            if (OP == PLUS_I2D) { double w; asm volatile("cvt.rn.f64.s32 %0, %1;" : "=d"(w) : "r"(u[c])); asm volatile("fma.rn.f64 %0, %0, %1, %2;" : "+d"(d[c]) : "d"(da), "d"(w)); asm volatile("add.rn.f32 %0, %0, %1;" : "+f"(f[c]) : "f"(1e-6f)); asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(u[c]) : "r"(0x9E3779B9u), "r"(i)); }
            if (OP == PLUS_MAGIC) { double w = __hiloint2double(0x43300000, (int)u[c]); asm volatile("add.rn.f64 %0, %0, %1;" : "+d"(w) : "d"(-4503601774854144.0)); asm volatile("fma.rn.f64 %0, %0, %1, %2;" : "+d"(d[c]) : "d"(da), "d"(w)); asm volatile("add.rn.f32 %0, %0, %1;" : "+f"(f[c]) : "f"(1e-6f)); asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(u[c]) : "r"(0x9E3779B9u), "r"(i)); }
            if (OP == DSETP_SEL) { d[c] = (d[c] == db) ? da : d[c]; asm volatile("" : "+d"(d[c])); }
        }
    }
    const long long t1 = clock64();
    double acc = 0;
    for (int c = 0; c < CHAINS; ++c) acc += d[c] + f[c] + u[c];
    if (acc == 12345.678) sink[0] = acc;
    if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
}

template <int OP>
void run(int sms, long long* dc, double* ds, int per_iter) {
    k<OP><<<sms, 1024>>>(dc, ds, 1);
    k<OP><<<sms, 1024>>>(dc, ds, 2);
    cudaDeviceSynchronize();
    long long hc[1024];
    cudaMemcpy(hc, dc, sizeof(long long) * sms, cudaMemcpyDeviceToHost);
    double avg = 0;
    for (int i = 0; i < sms; ++i) avg += (double)hc[i];
    avg /= sms;
    const double thread_instr = 1024.0 * ITERS * CHAINS * per_iter;
    printf("%-30s %8.1f thread-instr/clk/SM  (%.2f clk per warp-instruction per scheduler)\n", kNames[OP], thread_instr / avg,
           avg / (ITERS * CHAINS * per_iter * 8.0));
}

int main() {
    cudaDeviceProp p;
    cudaGetDeviceProperties(&p, 0);
    const int sms = p.multiProcessorCount;
    printf("# %s, %d SMs; 1024 threads/SM, %d chains/thread\n", p.name, sms, CHAINS);
    long long* dc; double* ds;
    cudaMalloc(&dc, sizeof(long long) * 1024); cudaMalloc(&ds, 8);
    run<DFMA>(sms, dc, ds, 1); run<DADD>(sms, dc, ds, 1); run<DMUL>(sms, dc, ds, 1);
    run<F2D>(sms, dc, ds, 1); run<D2F>(sms, dc, ds, 1); run<I2D>(sms, dc, ds, 1);
    run<FFMA_>(sms, dc, ds, 1); run<IMADW>(sms, dc, ds, 1); run<LOP>(sms, dc, ds, 1);
    run<MIX_DFMA_FFMA>(sms, dc, ds, 2); run<MIX_DFMA_F2D>(sms, dc, ds, 2); run<MIX_DFMA_I2D>(sms, dc, ds, 2);
    run<MIX_F2D_D2F>(sms, dc, ds, 2); run<MAGIC_I2D>(sms, dc, ds, 1); run<DSETP_SEL>(sms, dc, ds, 1);
    printf("# loop-carried groups: clk per GROUP per scheduler is what to compare\n");
    run<BASE_DFMA_FADD>(sms, dc, ds, 1); run<PLUS_NARROW>(sms, dc, ds, 1); run<PLUS_WIDEN>(sms, dc, ds, 1);
    run<PLUS_I2D>(sms, dc, ds, 1); run<PLUS_MAGIC>(sms, dc, ds, 1);
    return 0;
}
