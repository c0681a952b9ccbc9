// int_bench.cu — integer-pipe microbenchmark for the B200 (development aid; results go to profiles/).
// Measures sustained thread-instructions per clock per SM for the instruction classes the Goldilocks arithmetic
// is made of, so that the integer "roofline" of the NTT butterflies and of Poseidon can be stated from measurement.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/int_bench tools/int_bench.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include "../zkevm_prover_b200/csrc/gl64.cuh"

#define ITERS 2048

template <int MODE>
__global__ void __launch_bounds__(256) k(uint64_t *out, uint32_t seed, long long *cycles)
{
    uint32_t a = threadIdx.x * 2654435761u + seed, b = a ^ 0x9e3779b9u;
    uint64_t x0 = a, x1 = b, x2 = a + 1, x3 = b + 1, x4 = a + 2, x5 = b + 2, x6 = a + 3, x7 = b + 3;
    uint32_t y0 = a, y1 = b, y2 = a + 5, y3 = b + 5, y4 = a + 6, y5 = b + 6, y6 = a + 7, y7 = b + 7;
    double d0 = a, d1 = b, d2 = a + 1.0, d3 = b + 1.0, d4 = a + 2.0, d5 = b + 2.0, d6 = a + 3.0, d7 = b + 3.0;
    const double dm = 1.0 + 1e-9 * (seed & 3), dc = 1e-3;
    long long t0 = clock64();
#pragma unroll 1
    for (int i = 0; i < ITERS; i++)
    {
        if (MODE == 0) // IMAD.WIDE.U32: 8 chains, multiplicand = low word of the neighbouring accumulator (nothing to hoist)
        {
#pragma unroll
            for (int r = 0; r < 4; r++)
            {
                x0 += (uint64_t)(uint32_t)x1 * b; x1 += (uint64_t)(uint32_t)x2 * b; x2 += (uint64_t)(uint32_t)x3 * b; x3 += (uint64_t)(uint32_t)x4 * b;
                x4 += (uint64_t)(uint32_t)x5 * b; x5 += (uint64_t)(uint32_t)x6 * b; x6 += (uint64_t)(uint32_t)x7 * b; x7 += (uint64_t)(uint32_t)x0 * b;
            }
        }
        else if (MODE == 1) // IMAD (32-bit): y_i = y_{i+1} * a + y_i
        {
#pragma unroll
            for (int r = 0; r < 4; r++)
            {
                y0 += y1 * a; y1 += y2 * a; y2 += y3 * a; y3 += y4 * a; y4 += y5 * a; y5 += y6 * a; y6 += y7 * a; y7 += y0 * a;
            }
        }
        else if (MODE == 2) // IADD3 / LOP3 class: add.u32 with xor to stop folding
        {
#pragma unroll
            for (int r = 0; r < 4; r++)
                asm volatile("add.u32 %0, %0, %8;\n\tadd.u32 %1, %1, %9;\n\tadd.u32 %2, %2, %8;\n\tadd.u32 %3, %3, %9;\n\t"
                             "add.u32 %4, %4, %8;\n\tadd.u32 %5, %5, %9;\n\tadd.u32 %6, %6, %8;\n\tadd.u32 %7, %7, %9;"
                             : "+r"(y0), "+r"(y1), "+r"(y2), "+r"(y3), "+r"(y4), "+r"(y5), "+r"(y6), "+r"(y7)
                             : "r"(a), "r"(b));
        }
        else if (MODE == 3) // 1:1 mix: IMAD.WIDE chains + 32-bit xor/add chains (ALU)
        {
#pragma unroll
            for (int r = 0; r < 4; r++)
            {
                x0 += (uint64_t)(uint32_t)x1 * b; y0 = (y0 ^ y1) + a; x1 += (uint64_t)(uint32_t)x2 * b; y1 = (y1 ^ y2) + a;
                x2 += (uint64_t)(uint32_t)x3 * b; y2 = (y2 ^ y3) + a; x3 += (uint64_t)(uint32_t)x0 * b; y3 = (y3 ^ y0) + a;
            }
        }
        else if (MODE == 4) // carry chains: add.cc / addc pairs (64-bit adds), 4 pairs
        {
#pragma unroll
            for (int r = 0; r < 4; r++)
                asm volatile("add.cc.u32 %0, %0, %8;\n\taddc.u32 %1, %1, %9;\n\tadd.cc.u32 %2, %2, %8;\n\taddc.u32 %3, %3, %9;\n\t"
                             "add.cc.u32 %4, %4, %8;\n\taddc.u32 %5, %5, %9;\n\tadd.cc.u32 %6, %6, %8;\n\taddc.u32 %7, %7, %9;"
                             : "+r"(y0), "+r"(y1), "+r"(y2), "+r"(y3), "+r"(y4), "+r"(y5), "+r"(y6), "+r"(y7)
                             : "r"(a), "r"(b));
        }
        else if (MODE == 5) // mad.lo.cc / madc.hi pairs
        {
#pragma unroll
            for (int r = 0; r < 4; r++)
                asm volatile("mad.lo.cc.u32 %0, %8, %9, %0;\n\tmadc.hi.u32 %1, %8, %9, %1;\n\tmad.lo.cc.u32 %2, %8, %9, %2;\n\tmadc.hi.u32 %3, %8, %9, %3;\n\t"
                             "mad.lo.cc.u32 %4, %8, %9, %4;\n\tmadc.hi.u32 %5, %8, %9, %5;\n\tmad.lo.cc.u32 %6, %8, %9, %6;\n\tmadc.hi.u32 %7, %8, %9, %7;"
                             : "+r"(y0), "+r"(y1), "+r"(y2), "+r"(y3), "+r"(y4), "+r"(y5), "+r"(y6), "+r"(y7)
                             : "r"(a), "r"(b));
        }
        else if (MODE == 6) // gl::mul_nc x8 independent chains (32 "instructions" = 32 muls per iteration / 4)
        {
            x0 = gl::mul_nc(x0, x1); x1 = gl::mul_nc(x1, x2); x2 = gl::mul_nc(x2, x3); x3 = gl::mul_nc(x3, x4);
            x4 = gl::mul_nc(x4, x5); x5 = gl::mul_nc(x5, x6); x6 = gl::mul_nc(x6, x7); x7 = gl::mul_nc(x7, x0);
        }
        else if (MODE == 7) // gl::mul (canonical)
        {
            x0 = gl::mul(x0, x1); x1 = gl::mul(x1, x2); x2 = gl::mul(x2, x3); x3 = gl::mul(x3, x4);
            x4 = gl::mul(x4, x5); x5 = gl::mul(x5, x6); x6 = gl::mul(x6, x7); x7 = gl::mul(x7, x0);
        }
        else if (MODE == 8) // butterfly: t = mul(b, w); a' = a + t; b' = a - t   (4 butterflies)
        {
            uint64_t t;
            t = gl::mul(x1, x4); x1 = gl::sub(x0, t); x0 = gl::add(x0, t);
            t = gl::mul(x3, x5); x3 = gl::sub(x2, t); x2 = gl::add(x2, t);
            t = gl::mul(x1, x6); x1 = gl::sub(x2, t); x2 = gl::add(x2, t);
            t = gl::mul(x3, x7); x3 = gl::sub(x0, t); x0 = gl::add(x0, t);
        }
        else if (MODE == 9) // DFMA: 8 independent chains
        {
#pragma unroll
            for (int r = 0; r < 4; r++)
            {
                d0 = fma(d0, dm, dc); d1 = fma(d1, dm, dc); d2 = fma(d2, dm, dc); d3 = fma(d3, dm, dc);
                d4 = fma(d4, dm, dc); d5 = fma(d5, dm, dc); d6 = fma(d6, dm, dc); d7 = fma(d7, dm, dc);
            }
        }
        else if (MODE == 10) // 1:1 mix DFMA + IMAD.WIDE (do the two pipes run side by side?)
        {
#pragma unroll
            for (int r = 0; r < 4; r++)
            {
                d0 = fma(d0, dm, dc); x0 += (uint64_t)(uint32_t)x1 * b; d1 = fma(d1, dm, dc); x1 += (uint64_t)(uint32_t)x2 * b;
                d2 = fma(d2, dm, dc); x2 += (uint64_t)(uint32_t)x3 * b; d3 = fma(d3, dm, dc); x3 += (uint64_t)(uint32_t)x0 * b;
            }
        }
        else if (MODE == 11) // 2:1:1 mix DFMA + IMAD.WIDE + ALU
        {
#pragma unroll
            for (int r = 0; r < 4; r++)
            {
                d0 = fma(d0, dm, dc); d4 = fma(d4, dm, dc); x0 += (uint64_t)(uint32_t)x1 * b; y0 = (y0 ^ y1) + a;
                d1 = fma(d1, dm, dc); d5 = fma(d5, dm, dc); x1 += (uint64_t)(uint32_t)x0 * b; y1 = (y1 ^ y0) + a;
            }
        }
        else if (MODE == 12) // I2F.F64.U32 + F2I.U64.F64 round trips (conversion pipe rate): 8 per iteration
        {
            y0 = (uint32_t)__double2ull_rz(__uint2double_rn(y0) + dc) ^ a; y1 = (uint32_t)__double2ull_rz(__uint2double_rn(y1) + dc) ^ a;
            y2 = (uint32_t)__double2ull_rz(__uint2double_rn(y2) + dc) ^ a; y3 = (uint32_t)__double2ull_rz(__uint2double_rn(y3) + dc) ^ a;
        }
    }
    d0 += d1 + d2 + d3 + d4 + d5 + d6 + d7;
    y0 ^= (uint32_t)__double2loint(d0);
    long long t1 = clock64();
    uint64_t s = x0 ^ x1 ^ x2 ^ x3 ^ x4 ^ x5 ^ x6 ^ x7 ^ y0 ^ y1 ^ y2 ^ y3 ^ y4 ^ y5 ^ y6 ^ y7;
    if (s == 0x1234567) out[0] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) cycles[0] = t1 - t0;
}

template <int MODE>
void run(const char *name, double per_iter, uint64_t *out, long long *cyc)
{
    int sms = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    for (int warps = 4; warps <= 32; warps *= 2)
    {
        // blocks of 256 threads = 8 warps; "warps" resident warps per SM
        int threads = warps >= 8 ? 256 : warps * 32;
        int bps = warps >= 8 ? warps / 8 : 1;
        dim3 grid(sms * bps);
        k<MODE><<<grid, threads>>>(out, 1, cyc);
        cudaDeviceSynchronize();
        cudaEvent_t e0, e1;
        cudaEventCreate(&e0);
        cudaEventCreate(&e1);
        cudaEventRecord(e0);
        k<MODE><<<grid, threads>>>(out, 2, cyc);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms = 0;
        cudaEventElapsedTime(&ms, e0, e1);
        long long c = 0;
        cudaMemcpy(&c, cyc, sizeof(c), cudaMemcpyDeviceToHost);
        double ops = (double)ITERS * per_iter * threads * bps; // per SM
        printf("%-28s warps/SM=%2d  %.2f ms  cycles=%lld  ops/clk/SM=%.1f  ops/s(total)=%.3e\n", name, warps, ms, c, ops / (double)c,
               ops * sms / (ms * 1e-3));
    }
}

int main()
{
    uint64_t *out;
    long long *cyc;
    cudaMalloc(&out, 8);
    cudaMalloc(&cyc, 8);
    run<0>("IMAD.WIDE.U32", 32, out, cyc);
    run<1>("IMAD.LO", 32, out, cyc);
    run<2>("IADD", 32, out, cyc);
    run<3>("IMAD.WIDE + (LOP3,IADD) 16:32", 48, out, cyc);
    run<4>("add.cc/addc pairs", 32, out, cyc);
    run<5>("mad.lo.cc/madc.hi pairs", 32, out, cyc);
    run<6>("gl::mul_nc (muls)", 8, out, cyc);
    run<7>("gl::mul canonical (muls)", 8, out, cyc);
    run<8>("butterfly mul+add+sub", 4, out, cyc);
    run<9>("DFMA", 32, out, cyc);
    run<10>("DFMA + IMAD.WIDE 16:16", 32, out, cyc);
    run<11>("DFMA+WIDE+ALU 16:8:16", 40, out, cyc);
    run<12>("I2F.F64+DADD+F2I.U64 x4", 12, out, cyc);
    cudaError_t e = cudaGetLastError();
    printf("status: %s\n", cudaGetErrorString(e));
    return e != cudaSuccess;
}
