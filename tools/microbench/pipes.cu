// Pipe-rate microbenchmarks for the B200 integer/FP64 paths that a 256-bit Montgomery multiplier can use.
// Register-only dependent chains, 8 independent chains per thread; reports warp-instructions per cycle per SMSP.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o pipes pipes.cu
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
#define NCH 8
#define UNROLL 16

template <int KIND>
__global__ void __launch_bounds__(256) k(uint64_t* sink, int iters, uint32_t seed) {
    uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t m = seed | 1u;
    if (KIND == 0) {   // IMAD lo
        uint32_t x[NCH];
        for (int j = 0; j < NCH; j++) x[j] = t * 2654435761u + j + seed;
        for (int it = 0; it < iters; it++) {
#pragma unroll
            for (int r = 0; r < UNROLL; r++)
#pragma unroll
                for (int j = 0; j < NCH; j++) asm volatile("mad.lo.u32 %0, %0, %1, %0;" : "+r"(x[j]) : "r"(m));
        }
        uint32_t s = 0; for (int j = 0; j < NCH; j++) s ^= x[j];
        if (s == 0x12345678u) sink[0] = s;
    } else if (KIND == 1) {   // IMAD.WIDE with 64-bit accumulate
        uint64_t x[NCH];
        for (int j = 0; j < NCH; j++) x[j] = ((uint64_t)t << 20) + j + seed;
        for (int it = 0; it < iters; it++) {
#pragma unroll
            for (int r = 0; r < UNROLL; r++)
#pragma unroll
                for (int j = 0; j < NCH; j++) { uint32_t lo = (uint32_t)x[j]; asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(x[j]) : "r"(lo), "r"(m)); }
        }
        uint64_t s = 0; for (int j = 0; j < NCH; j++) s ^= x[j];
        if (s == 0x12345678ull) sink[0] = s;
    } else if (KIND == 2) {   // IMAD.HI
        uint32_t x[NCH];
        for (int j = 0; j < NCH; j++) x[j] = t * 2654435761u + j + seed;
        for (int it = 0; it < iters; it++) {
#pragma unroll
            for (int r = 0; r < UNROLL; r++)
#pragma unroll
                for (int j = 0; j < NCH; j++) asm volatile("mad.hi.u32 %0, %0, %1, %0;" : "+r"(x[j]) : "r"(m));
        }
        uint32_t s = 0; for (int j = 0; j < NCH; j++) s ^= x[j];
        if (s == 0x12345678u) sink[0] = s;
    } else if (KIND == 3) {   // mul.wide (no accumulate)
        uint64_t x[NCH];
        for (int j = 0; j < NCH; j++) x[j] = ((uint64_t)t << 20) + j + seed;
        for (int it = 0; it < iters; it++) {
#pragma unroll
            for (int r = 0; r < UNROLL; r++)
#pragma unroll
                for (int j = 0; j < NCH; j++) { uint32_t lo = (uint32_t)x[j] | 1u; asm volatile("mul.wide.u32 %0, %1, %2;" : "=l"(x[j]) : "r"(lo), "r"(m)); }
        }
        uint64_t s = 0; for (int j = 0; j < NCH; j++) s ^= x[j];
        if (s == 0x12345678ull) sink[0] = s;
    } else if (KIND == 4) {   // DFMA
        double x[NCH];
        double c = 1.0000001 + 1e-9 * (seed & 3), d = 1e-9;
        for (int j = 0; j < NCH; j++) x[j] = 1.0 + t * 1e-6 + j;
        for (int it = 0; it < iters; it++) {
#pragma unroll
            for (int r = 0; r < UNROLL; r++)
#pragma unroll
                for (int j = 0; j < NCH; j++) asm volatile("fma.rz.f64 %0, %0, %1, %2;" : "+d"(x[j]) : "d"(c), "d"(d));
        }
        double s = 0; for (int j = 0; j < NCH; j++) s += x[j];
        if (s == 0.12345678) sink[0] = (uint64_t)s;
    } else if (KIND == 5) {   // DFMA + IMAD.WIDE 1:1
        double x[NCH / 2];
        uint64_t y[NCH / 2];
        double c = 1.0000001 + 1e-9 * (seed & 3), d = 1e-9;
        for (int j = 0; j < NCH / 2; j++) { x[j] = 1.0 + t * 1e-6 + j; y[j] = ((uint64_t)t << 20) + j + seed; }
        for (int it = 0; it < iters; it++) {
#pragma unroll
            for (int r = 0; r < UNROLL; r++)
#pragma unroll
                for (int j = 0; j < NCH / 2; j++) {
                    asm volatile("fma.rz.f64 %0, %0, %1, %2;" : "+d"(x[j]) : "d"(c), "d"(d));
                    uint32_t lo = (uint32_t)y[j]; asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(y[j]) : "r"(lo), "r"(m));
                }
        }
        double s = 0; uint64_t u = 0; for (int j = 0; j < NCH / 2; j++) { s += x[j]; u ^= y[j]; }
        if (s == 0.12345678 || u == 0x12345678ull) sink[0] = (uint64_t)s + u;
    } else if (KIND == 6) {   // IADD3 (add.cc chain pairs: 64-bit adds)
        uint32_t x[NCH];
        for (int j = 0; j < NCH; j++) x[j] = t * 2654435761u + j + seed;
        for (int it = 0; it < iters; it++) {
#pragma unroll
            for (int r = 0; r < UNROLL; r++)
#pragma unroll
                for (int j = 0; j < NCH; j++) asm volatile("add.u32 %0, %0, %1;" : "+r"(x[j]) : "r"(m));
        }
        uint32_t s = 0; for (int j = 0; j < NCH; j++) s ^= x[j];
        if (s == 0x12345678u) sink[0] = s;
    } else if (KIND == 7) {   // IMAD.WIDE + 2x IADD 
        uint64_t y[NCH / 2];
        uint32_t x[NCH];
        for (int j = 0; j < NCH; j++) x[j] = t * 2654435761u + j + seed;
        for (int j = 0; j < NCH / 2; j++) y[j] = ((uint64_t)t << 20) + j + seed;
        for (int it = 0; it < iters; it++) {
#pragma unroll
            for (int r = 0; r < UNROLL; r++)
#pragma unroll
                for (int j = 0; j < NCH / 2; j++) {
                    uint32_t lo = (uint32_t)y[j]; asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(y[j]) : "r"(lo), "r"(m));
                    asm volatile("add.u32 %0, %0, %1;" : "+r"(x[2 * j]) : "r"(m));
                    asm volatile("add.u32 %0, %0, %1;" : "+r"(x[2 * j + 1]) : "r"(m));
                }
        }
        uint32_t s = 0; uint64_t u = 0; for (int j = 0; j < NCH; j++) s ^= x[j]; for (int j = 0; j < NCH / 2; j++) u ^= y[j];
        if (s == 0x12345678u || u == 0x12345678ull) sink[0] = s + u;
    } else if (KIND == 8) {   // DFMA x2 + 64-bit integer add (add.cc + addc) x2 : the FP64 big-int inner product shape
        double x[NCH / 2];
        uint32_t a[NCH];
        double c = 1.0000001 + 1e-9 * (seed & 3), d = 1e-9;
        for (int j = 0; j < NCH / 2; j++) x[j] = 1.0 + t * 1e-6 + j;
        for (int j = 0; j < NCH; j++) a[j] = t + j;
        for (int it = 0; it < iters; it++) {
#pragma unroll
            for (int r = 0; r < UNROLL; r++)
#pragma unroll
                for (int j = 0; j < NCH / 2; j++) {
                    asm volatile("fma.rz.f64 %0, %0, %1, %2;" : "+d"(x[j]) : "d"(c), "d"(d));
                    asm volatile("add.cc.u32 %0, %0, %2;\n\taddc.u32 %1, %1, %2;" : "+r"(a[2 * j]), "+r"(a[2 * j + 1]) : "r"(m));
                }
        }
        double s = 0; uint32_t u = 0; for (int j = 0; j < NCH / 2; j++) s += x[j]; for (int j = 0; j < NCH; j++) u ^= a[j];
        if (s == 0.12345678 || u == 0x12345678u) sink[0] = (uint64_t)s + u;
    } else if (KIND == 9) {   // IMAD lo + IMAD.HI pair (separate halves)
        uint32_t x[NCH];
        for (int j = 0; j < NCH; j++) x[j] = t * 2654435761u + j + seed;
        for (int it = 0; it < iters; it++) {
#pragma unroll
            for (int r = 0; r < UNROLL; r++)
#pragma unroll
                for (int j = 0; j < NCH; j += 2) {
                    asm volatile("mad.lo.u32 %0, %0, %1, %0;" : "+r"(x[j]) : "r"(m));
                    asm volatile("mad.hi.u32 %0, %0, %1, %0;" : "+r"(x[j + 1]) : "r"(m));
                }
        }
        uint32_t s = 0; for (int j = 0; j < NCH; j++) s ^= x[j];
        if (s == 0x12345678u) sink[0] = s;
    } else if (KIND == 10) {  // FFMA (fp32) for reference
        float x[NCH];
        float c = 1.0000001f, d = 1e-9f * seed;
        for (int j = 0; j < NCH; j++) x[j] = 1.0f + t * 1e-6f + j;
        for (int it = 0; it < iters; it++) {
#pragma unroll
            for (int r = 0; r < UNROLL; r++)
#pragma unroll
                for (int j = 0; j < NCH; j++) asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(x[j]) : "f"(c), "f"(d));
        }
        float s = 0; for (int j = 0; j < NCH; j++) s += x[j];
        if (s == 0.12345678f) sink[0] = (uint64_t)s;
    } else if (KIND == 11) {  // IMAD.WIDE with 16-bit-ish small operands (check data dependence) and accumulate into separate reg
        uint64_t x[NCH];
        for (int j = 0; j < NCH; j++) x[j] = j + (seed & 0xff);
        uint32_t mm = m & 0xffff;
        for (int it = 0; it < iters; it++) {
#pragma unroll
            for (int r = 0; r < UNROLL; r++)
#pragma unroll
                for (int j = 0; j < NCH; j++) { uint32_t lo = (uint32_t)x[j] & 0xffffu; asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(x[j]) : "r"(lo), "r"(mm)); }
        }
        uint64_t s = 0; for (int j = 0; j < NCH; j++) s ^= x[j];
        if (s == 0x12345678ull) sink[0] = s;
    }
}

template <int KIND> double run(const char* name, double inst_per_iter_thread, int sms, int blocks_per_sm, uint64_t* sink, double mhz) {
    int iters = 2000;
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    float best = 1e30f;
    for (int rep = 0; rep < 3; rep++) {
        cudaEventRecord(e0);
        k<KIND><<<sms * blocks_per_sm, 256>>>(sink, iters, 12345u);
        cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1); if (ms < best) best = ms;
    }
    double threads = (double)sms * blocks_per_sm * 256;
    double inst = threads * iters * inst_per_iter_thread;      // thread-level instructions
    double per_s = inst / (best * 1e-3);
    double warp_inst_per_clk_smsp = per_s / 32.0 / (sms * 4.0) / (mhz * 1e6);
    printf("%-28s blocks/SM=%d  %.3f ms  %.3f T thread-inst/s  %.3f warp-inst/clk/SMSP (at %.0f MHz) -> %.2f clk per warp-inst\n", name, blocks_per_sm, best,
           per_s / 1e12, warp_inst_per_clk_smsp, mhz, 1.0 / warp_inst_per_clk_smsp);
    return per_s;
}

int main(int argc, char** argv) {
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    int sms = p.multiProcessorCount;
    int clk_khz = 0; cudaDeviceGetAttribute(&clk_khz, cudaDevAttrClockRate, 0);
    double mhz = argc > 1 ? atof(argv[1]) : clk_khz / 1000.0;
    printf("%s SMs=%d clockRate=%.0f MHz (rates below assume this clock)\n", p.name, sms, mhz);
    uint64_t* sink; cudaMalloc(&sink, 64);
    const double N = NCH * UNROLL;
    for (int b = 4; b <= 8; b += 4) {
        run<0>("IMAD lo", N, sms, b, sink, mhz);
        run<1>("IMAD.WIDE acc64", N, sms, b, sink, mhz);
        run<11>("IMAD.WIDE acc64 small ops", N, sms, b, sink, mhz);
        run<2>("IMAD.HI", N, sms, b, sink, mhz);
        run<3>("IMAD.WIDE no acc (mul.wide)", N, sms, b, sink, mhz);
        run<9>("IMAD lo + IMAD.HI", N, sms, b, sink, mhz);
        run<4>("DFMA", N, sms, b, sink, mhz);
        run<5>("DFMA + IMAD.WIDE 1:1", N, sms, b, sink, mhz);
        run<6>("IADD", N, sms, b, sink, mhz);
        run<7>("IMAD.WIDE + 2 IADD", N / 2 * 3, sms, b, sink, mhz);
        run<8>("DFMA + 2 IADD(cc)", N / 2 * 3, sms, b, sink, mhz);
        run<10>("FFMA", N, sms, b, sink, mhz);
    }
    cudaError_t e = cudaDeviceSynchronize();
    printf("status: %s\n", cudaGetErrorString(e));
    return e != cudaSuccess;
}
