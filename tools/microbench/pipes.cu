// Issue-rate microbenchmark for the integer instructions the decode kernels lean on (sm_100a).
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o pipes pipes.cu ; run: ./pipes
#include <cstdio>
#include <cuda_runtime.h>
#define N_IT 2048
template <int OP>
__global__ void __launch_bounds__(256) k(int* out, int seed) {
    int a[8];
#pragma unroll
    for (int i = 0; i < 8; i++) a[i] = seed + threadIdx.x * (i + 1);
    int c = seed | 3;
    for (int it = 0; it < N_IT; it++) {
#pragma unroll
        for (int i = 0; i < 8; i++) {
            if (OP == 0) asm volatile("mad.lo.s32 %0, %0, %1, %0;" : "+r"(a[i]) : "r"(c));
            if (OP == 1) asm volatile("mul.hi.s32 %0, %0, %1;" : "+r"(a[i]) : "r"(c));
            if (OP == 2) asm volatile("shr.s32 %0, %0, 3;" : "+r"(a[i]));
            if (OP == 3) asm volatile("lop3.b32 %0, %0, %1, %0, 0x96;" : "+r"(a[i]) : "r"(c));
            if (OP == 4) asm volatile("{.reg .u32 t; shr.s32 t, %0, 3; add.s32 %0, t, %1;}" : "+r"(a[i]) : "r"(c));   // LEA.HI.SX32
            if (OP == 5) { asm volatile("mad.lo.s32 %0, %0, %1, %0;" : "+r"(a[i]) : "r"(c)); asm volatile("shr.s32 %0, %0, 3;" : "+r"(a[i])); }
            if (OP == 6) asm volatile("mul.hi.u32 %0, %0, 65536;" : "+r"(a[i]));
            if (OP == 7) asm volatile("prmt.b32 %0, %0, %1, 0x9910;" : "+r"(a[i]) : "r"(c));
            if (OP == 9) asm volatile("cvt.pack.sat.u8.s32.b32 %0, %0, %1, %0;" : "+r"(a[i]) : "r"(c));
            if (OP == 10) asm volatile("dp2a.lo.s32.u32 %0, %0, %1, %0;" : "+r"(a[i]) : "r"(c));
            if (OP == 11) { asm volatile("dp2a.lo.s32.u32 %0, %0, %1, %0;" : "+r"(a[i]) : "r"(c)); asm volatile("lop3.b32 %0, %0, %1, %0, 0x96;" : "+r"(a[i]) : "r"(c)); }
            if (OP == 12) { asm volatile("dp2a.lo.s32.u32 %0, %0, %1, %0;" : "+r"(a[i]) : "r"(c)); asm volatile("mad.lo.s32 %0, %0, %1, %0;" : "+r"(a[i]) : "r"(c)); }
            if (OP == 13) { asm volatile("cvt.pack.sat.u8.s32.b32 %0, %0, %1, %0;" : "+r"(a[i]) : "r"(c)); asm volatile("lop3.b32 %0, %0, %1, %0, 0x96;" : "+r"(a[i]) : "r"(c)); }
            if (OP == 14) { asm volatile("cvt.pack.sat.u8.s32.b32 %0, %0, %1, %0;" : "+r"(a[i]) : "r"(c)); asm volatile("mad.lo.s32 %0, %0, %1, %0;" : "+r"(a[i]) : "r"(c)); }
            if (OP == 15) asm volatile("cvt.pack.sat.s16.s32 %0, %0, %1;" : "+r"(a[i]) : "r"(c));                       // I2IP.S16.S32.SAT
            if (OP == 16) a[i] = (int)__viaddmin_s16x2_relu((unsigned)a[i], (unsigned)c, 0x01ff01ffu);                      // VIADDMNMX.S16x2.RELU
            if (OP == 17) { asm volatile("mad.lo.s32 %0, %0, %1, %0;" : "+r"(a[i]) : "r"(c)); a[i] = (int)__viaddmin_s16x2_relu((unsigned)a[i], (unsigned)c, 0x01ff01ffu); }
            if (OP == 18) { asm volatile("prmt.b32 %0, %0, %1, 0x7632;" : "+r"(a[i]) : "r"(c)); asm volatile("mad.lo.s32 %0, %0, %1, %0;" : "+r"(a[i]) : "r"(c)); }
            if (OP == 19) { asm volatile("shr.s32 %0, %0, 3;" : "+r"(a[i])); asm volatile("mad.lo.s32 %0, %0, %1, %0;" : "+r"(a[i]) : "r"(c)); asm volatile("mad.lo.s32 %0, %0, %1, %0;" : "+r"(a[i]) : "r"(c)); }
            if (OP == 20) asm volatile("{.reg .s64 t; mad.wide.s32 t, %0, %1, 5; cvt.u32.u64 %0, t;}" : "+r"(a[i]) : "r"(c));   // IMAD.WIDE
            if (OP == 21) asm volatile("shf.l.wrap.b32 %0, %0, %1, 7;" : "+r"(a[i]) : "r"(c));                               // SHF funnel
            if (OP == 22) a[i] = (int)__vimax3_s16x2((unsigned)a[i], (unsigned)c, 0x00100010u);                             // VIMNMX3.S16x2
            if (OP == 8) asm volatile("{.reg .s32 t; add.s32 t, %0, %1; min.s32 t, t, 511; max.s32 %0, t, 0;}" : "+r"(a[i]) : "r"(c));
        }
    }
    int s = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) s += a[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
template <int OP> void run(const char* name, int per_iter) {
    int* d; cudaMalloc(&d, 148 * 8 * 256 * sizeof(int));
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    k<OP><<<148 * 8, 256>>>(d, 7); cudaDeviceSynchronize();
    cudaEventRecord(e0); k<OP><<<148 * 8, 256>>>(d, 7); cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    double warp_inst = 148.0 * 8 * 8 * (double)N_IT * 8 * per_iter;   // blocks * warps * iters * 8 chains
    printf("%-34s %8.3f ms  %6.2f warp-inst/clk/SM (at 1.965 GHz)\n", name, ms, warp_inst / (ms * 1e-3) / 148 / 1.965e9);
    cudaFree(d);
}
int main() {
    run<0>("IMAD", 1); run<1>("IMAD.HI (s32, reg)", 1); run<2>("SHF.R.S32", 1); run<3>("LOP3", 1);
    run<4>("LEA.HI.SX32 (shr+add)", 1); run<5>("IMAD + SHF pair", 2); run<6>("IMAD.HI.U32 x 65536", 1);
    run<7>("PRMT", 1); run<8>("VIADDMNMX.RELU (add,min,max)", 1); run<9>("I2IP.U8.S32.SAT (cvt.pack.sat)", 1);
    run<10>("IDP.2A (dp2a.lo)", 1); run<11>("IDP.2A + LOP3 pair", 2); run<12>("IDP.2A + IMAD pair", 2);
    run<13>("I2IP + LOP3 pair", 2); run<14>("I2IP + IMAD pair", 2);
    run<15>("I2IP.S16.S32.SAT", 1); run<16>("VIADDMNMX.S16x2.RELU", 1); run<17>("IMAD + VIADDMNMX.S16x2 pair", 2);
    run<18>("PRMT + IMAD pair", 2); run<19>("SHF + 2 IMAD (1 ALU : 2 FMA)", 3); run<20>("IMAD.WIDE", 1);
    run<21>("SHF.L.W (funnel)", 1); run<22>("VIMNMX3.S16x2", 1);
    return 0;
}
