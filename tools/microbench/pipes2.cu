// Second pipe microbenchmark: instruction mixes that ptxas cannot fold, so SASS counts are what the source says.
// Each kernel: NCH independent chains per thread; prints time; SASS histogram is taken offline with cuobjdump.
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
#define U 8

template <int KIND>
__global__ void __launch_bounds__(256) k(uint32_t* sink, const uint32_t* __restrict__ in, int iters) {
    uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t x[16], b[8], y[10];
    for (int j = 0; j < 10; j++) y[j] = in[(t * 7 + j) & 1023];
    for (int j = 0; j < 16; j++) x[j] = in[(t + j) & 1023];
    for (int j = 0; j < 8; j++) b[j] = in[(t * 3 + j) & 1023] | 1u;
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int r = 0; r < U; r++) {
            if (KIND == 0) {        // 16 IMAD lo
#pragma unroll
                for (int j = 0; j < 16; j++) asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(x[j]) : "r"(b[j & 7]), "r"(x[(j + 5) & 15]));
            } else if (KIND == 1) { // 16 IADD (x += other chain)
#pragma unroll
                for (int j = 0; j < 16; j++) asm volatile("add.u32 %0, %0, %1;" : "+r"(x[j]) : "r"(x[(j + 5) & 15]));
            } else if (KIND == 2) { // 8 IMAD + 8 IADD
#pragma unroll
                for (int j = 0; j < 8; j++) {
                    asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(x[j]) : "r"(b[j]), "r"(x[(j + 3) & 7]));
                    asm volatile("add.u32 %0, %0, %1;" : "+r"(x[8 + j]) : "r"(x[8 + ((j + 3) & 7)]));
                }
            } else if (KIND == 3) { // 8 LOP3 + 8 IMAD
#pragma unroll
                for (int j = 0; j < 8; j++) {
                    asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(x[j]) : "r"(b[j]), "r"(x[(j + 3) & 7]));
                    asm volatile("xor.b32 %0, %0, %1;" : "+r"(x[8 + j]) : "r"(x[8 + ((j + 3) & 7)]));
                }
            } else if (KIND == 4) { // two 8-word rows: acc(9 words) += a[0..7 even/odd] * b  (fq_mul's row pattern): 16 wide MACs
                // row A on x[0..8] with multiplier b[r&7], operand words b[0..7]; row B on x[... ] uses second half
                asm volatile(
                    "mad.lo.cc.u32 %0, %9, %17, %0;\n\tmadc.hi.cc.u32 %1, %9, %17, %1;\n\t"
                    "madc.lo.cc.u32 %2, %11, %17, %2;\n\tmadc.hi.cc.u32 %3, %11, %17, %3;\n\t"
                    "madc.lo.cc.u32 %4, %13, %17, %4;\n\tmadc.hi.cc.u32 %5, %13, %17, %5;\n\t"
                    "madc.lo.cc.u32 %6, %15, %17, %6;\n\tmadc.hi.cc.u32 %7, %15, %17, %7;\n\t"
                    "addc.u32 %8, %8, 0;\n\t"
                    : "+r"(x[0]), "+r"(x[1]), "+r"(x[2]), "+r"(x[3]), "+r"(x[4]), "+r"(x[5]), "+r"(x[6]), "+r"(x[7]), "+r"(x[8])
                    : "r"(b[0]), "r"(b[1]), "r"(b[2]), "r"(b[3]), "r"(b[4]), "r"(b[5]), "r"(b[6]), "r"(b[7]), "r"(x[9 + (r & 3)]));
                asm volatile(
                    "mad.lo.cc.u32 %0, %10, %17, %0;\n\tmadc.hi.cc.u32 %1, %10, %17, %1;\n\t"
                    "madc.lo.cc.u32 %2, %12, %17, %2;\n\tmadc.hi.cc.u32 %3, %12, %17, %3;\n\t"
                    "madc.lo.cc.u32 %4, %14, %17, %4;\n\tmadc.hi.cc.u32 %5, %14, %17, %5;\n\t"
                    "madc.lo.cc.u32 %6, %16, %17, %6;\n\tmadc.hi.cc.u32 %7, %16, %17, %7;\n\t"
                    "addc.u32 %8, %8, 0;\n\t"
                    : "+r"(y[0]), "+r"(y[1]), "+r"(y[2]), "+r"(y[3]), "+r"(y[4]), "+r"(y[5]), "+r"(y[6]), "+r"(y[7]), "+r"(y[8])
                    : "r"(b[0]), "r"(b[1]), "r"(b[2]), "r"(b[3]), "r"(b[4]), "r"(b[5]), "r"(b[6]), "r"(b[7]), "r"(x[9 + (r & 3)]));
            } else if (KIND == 5) { // 8 independent 64-bit MACs: (hi:lo) += a*b  [mad.lo.cc + madc.hi]
#pragma unroll
                for (int j = 0; j < 8; j++)
                    asm volatile("mad.lo.cc.u32 %0, %2, %3, %0;\n\tmadc.hi.u32 %1, %2, %3, %1;" : "+r"(x[2 * j]), "+r"(x[2 * j + 1]) : "r"(b[j]), "r"(b[(j + r) & 7]));
            } else if (KIND == 6) { // 8 products without accumulate (mul.wide) xor-folded into chains by ALU: 8 IMAD.WIDE + 16 LOP3
#pragma unroll
                for (int j = 0; j < 8; j++) {
                    uint64_t p;
                    asm volatile("mul.wide.u32 %0, %1, %2;" : "=l"(p) : "r"(x[2 * j + 1]), "r"(b[j]));
                    x[2 * j] ^= (uint32_t)p; x[2 * j + 1] ^= (uint32_t)(p >> 32);
                }
            } else if (KIND == 7) { // 8 products no-acc + 64-bit add on ALU (add.cc/addc): 8 IMAD.WIDE + 16 IADD3
#pragma unroll
                for (int j = 0; j < 8; j++) {
                    uint32_t lo, hi;
                    asm volatile("{.reg .u64 p;\n\tmul.wide.u32 p, %2, %3;\n\tmov.b64 {%0,%1}, p;}" : "=r"(lo), "=r"(hi) : "r"(b[j]), "r"(x[2 * j]));
                    asm volatile("add.cc.u32 %0, %0, %2;\n\taddc.u32 %1, %1, %3;" : "+r"(x[2 * j]), "+r"(x[2 * j + 1]) : "r"(lo), "r"(hi));
                }
            } else if (KIND == 8) { // 8 IMAD.HI
#pragma unroll
                for (int j = 0; j < 8; j++) asm volatile("mad.hi.u32 %0, %0, %1, %2;" : "+r"(x[j]) : "r"(b[j]), "r"(x[(j + 3) & 7]));
            } else if (KIND == 9) { // 8 DFMA + 8 IMAD lo
                double* d = reinterpret_cast<double*>(x + 8);
#pragma unroll
                for (int j = 0; j < 4; j++) {
                    asm volatile("fma.rz.f64 %0, %0, %1, %2;" : "+d"(d[j]) : "d"(d[(j + 1) & 3]), "d"(d[(j + 2) & 3]));
                    asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(x[j]) : "r"(b[j]), "r"(x[(j + 3) & 7]));
                    asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(x[4 + j]) : "r"(b[j]), "r"(x[(j + 3) & 7]));
                }
            } else if (KIND == 10) { // 4 DFMA + 8 IADD
                double* d = reinterpret_cast<double*>(x + 8);
#pragma unroll
                for (int j = 0; j < 4; j++) {
                    asm volatile("fma.rz.f64 %0, %0, %1, %2;" : "+d"(d[j]) : "d"(d[(j + 1) & 3]), "d"(d[(j + 2) & 3]));
                    asm volatile("add.u32 %0, %0, %1;" : "+r"(x[j]) : "r"(x[(j + 3) & 7]));
                    asm volatile("add.u32 %0, %0, %1;" : "+r"(x[4 + j]) : "r"(x[(j + 3) & 7]));
                }
            } else if (KIND == 11) { // 8 FFMA + 8 IMAD
                float* f = reinterpret_cast<float*>(x + 8);
#pragma unroll
                for (int j = 0; j < 8; j++) {
                    asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(f[j]) : "f"(f[(j + 1) & 7]), "f"(f[(j + 2) & 7]));
                    asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(x[j]) : "r"(b[j]), "r"(x[(j + 3) & 7]));
                }
            } else if (KIND == 12) { // 8 FFMA + 8 IADD
                float* f = reinterpret_cast<float*>(x + 8);
#pragma unroll
                for (int j = 0; j < 8; j++) {
                    asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(f[j]) : "f"(f[(j + 1) & 7]), "f"(f[(j + 2) & 7]));
                    asm volatile("add.u32 %0, %0, %1;" : "+r"(x[j]) : "r"(x[(j + 3) & 7]));
                }
            }
        }
    }
    uint32_t s = 0;
    for (int j = 0; j < 16; j++) s ^= x[j];
    for (int j = 0; j < 10; j++) s ^= y[j];
    if (s == 0x12345678u) sink[0] = s;
}

template <int KIND> void run(const char* name, int sms, uint32_t* sink, const uint32_t* in, double mhz) {
    int iters = 4000;
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    float best = 1e30f;
    for (int rep = 0; rep < 3; rep++) {
        cudaEventRecord(e0);
        k<KIND><<<sms * 8, 256>>>(sink, in, iters);
        cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1); if (ms < best) best = ms;
    }
    // clocks per (unrolled round) per warp per SMSP: 16 warps per SMSP resident
    double warps_per_smsp = 8.0 * 256 / 32 / 4;
    double clk_per_round = best * 1e-3 * mhz * 1e6 / (iters * (double)U) / warps_per_smsp;
    printf("kind %2d %-46s %.3f ms  %.2f clk per round per warp (SMSP-serialised)\n", KIND, name, best, clk_per_round);
}

int main(int argc, char** argv) {
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    int sms = p.multiProcessorCount;
    double mhz = argc > 1 ? atof(argv[1]) : 1965.0;
    uint32_t *sink, *in; cudaMalloc(&sink, 64); cudaMalloc(&in, 4096);
    uint32_t h[1024]; for (int i = 0; i < 1024; i++) h[i] = 0x9e3779b9u * (i + 1) ^ (i << 7);
    cudaMemcpy(in, h, 4096, cudaMemcpyHostToDevice);
    printf("%s SMs=%d assumed %.0f MHz\n", p.name, sms, mhz);
    run<0>("16 IMAD", sms, sink, in, mhz);
    run<1>("16 IADD", sms, sink, in, mhz);
    run<2>("8 IMAD + 8 IADD", sms, sink, in, mhz);
    run<3>("8 IMAD + 8 LOP", sms, sink, in, mhz);
    run<4>("2 rows x 4 wide MAC carry chains (8 IMAD.WIDE.X + 2 addc)", sms, sink, in, mhz);
    run<5>("8 x 64-bit MAC (mad.lo.cc+madc.hi)", sms, sink, in, mhz);
    run<6>("8 mul.wide + 16 xor", sms, sink, in, mhz);
    run<7>("8 mul.wide + 8x(add.cc,addc)", sms, sink, in, mhz);
    run<8>("8 IMAD.HI", sms, sink, in, mhz);
    run<9>("4 DFMA + 8 IMAD", sms, sink, in, mhz);
    run<10>("4 DFMA + 8 IADD", sms, sink, in, mhz);
    run<11>("8 FFMA + 8 IMAD", sms, sink, in, mhz);
    run<12>("8 FFMA + 8 IADD", sms, sink, in, mhz);
    cudaError_t e = cudaDeviceSynchronize();
    printf("status: %s\n", cudaGetErrorString(e));
    return e != cudaSuccess;
}
