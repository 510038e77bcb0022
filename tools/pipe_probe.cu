// pipe_probe.cu - which issue pipe do the integer ops of the score kernel use on sm_100a, and how do they mix?
// Each test runs 8 independent dependency chains per thread; results are warp-instructions per clock per SMSP.
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>

#define CHAINS 8
#define REP 16
template <int T>
__global__ void __launch_bounds__(256) probe(uint32_t* out, int iters, uint32_t b, uint32_t c) {
    uint32_t x[CHAINS], y[CHAINS];
#pragma unroll
    for (int k = 0; k < CHAINS; k++) { x[k] = b * (k + 1) + threadIdx.x; y[k] = c + k; }
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int r = 0; r < REP; r++) {
#pragma unroll
            for (int k = 0; k < CHAINS; k++) {
                if (T == 0) asm volatile("add.s32 %0, %0, 0x00050005;" : "+r"(x[k]));                       // VIADD imm
                if (T == 1) asm volatile("max.s16x2 %0, %0, %1;" : "+r"(x[k]) : "r"(y[k]));                 // VIMNMX.S16x2
                if (T == 2) { asm volatile("{.reg .b32 t; add.s16x2 t, %0, %1; max.s16x2 %0, t, %2;}" : "+r"(x[k]) : "r"(b), "r"(y[k])); } // VIADDMNMX
                if (T == 3) { asm volatile("add.s32 %0, %0, 0x00050005;" : "+r"(x[k])); asm volatile("max.s16x2 %0, %0, %1;" : "+r"(x[k]) : "r"(y[k])); }  // VIADD + VIMNMX
                if (T == 4) asm volatile("mad.lo.s32 %0, %0, %1, %2;" : "+r"(x[k]) : "r"(b), "r"(c));       // IMAD
                if (T == 5) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(x[k]) : "r"(b), "r"(y[k])); // LOP3
                if (T == 6) asm volatile("prmt.b32 %0, %0, %1, %2;" : "+r"(x[k]) : "r"(b), "r"(y[k]));      // PRMT
                if (T == 7) { asm volatile("{.reg .b32 t; add.s16x2 t, %0, %1; max.s16x2 %0, t, %2;}" : "+r"(x[k]) : "r"(b), "r"(y[k]));
                              asm volatile("mad.lo.s32 %0, %0, %1, %2;" : "+r"(y[k]) : "r"(b), "r"(c)); }    // VIADDMNMX + IMAD
                if (T == 8) { asm volatile("{.reg .b32 t; add.s16x2 t, %0, %1; max.s16x2 %0, t, %2;}" : "+r"(x[k]) : "r"(b), "r"(c));
                              asm volatile("max.s16x2 %0, %0, %1;" : "+r"(y[k]) : "r"(c)); }                 // VIADDMNMX + VIMNMX
                if (T == 9) { asm volatile("{.reg .b32 t; add.s16x2 t, %0, %1; max.s16x2 %0, t, %2;}" : "+r"(x[k]) : "r"(b), "r"(c));
                              asm volatile("add.s32 %0, %0, 0x00050005;" : "+r"(y[k])); }                    // VIADDMNMX + VIADD
                if (T == 10) { asm volatile("max.s16x2 %0, %0, %1;" : "+r"(x[k]) : "r"(c));
                               asm volatile("mad.lo.s32 %0, %0, %1, %2;" : "+r"(y[k]) : "r"(b), "r"(c)); }   // VIMNMX + IMAD
                if (T == 11) { asm volatile("add.s32 %0, %0, 0x00050005;" : "+r"(x[k]));
                               asm volatile("mad.lo.s32 %0, %0, %1, %2;" : "+r"(y[k]) : "r"(b), "r"(c)); }   // VIADD + IMAD
                if (T == 12) asm volatile("max.u16x2 %0, %0, %1;" : "+r"(x[k]) : "r"(y[k]));                 // VIMNMX.U16x2
                if (T == 13) { asm volatile("{.reg .b32 t; max.s16x2 t, %0, %1; max.s16x2 %0, t, %2;}" : "+r"(x[k]) : "r"(b), "r"(y[k])); } // VIMNMX3
                if (T == 14) asm volatile("add.s32 %0, %0, %1;" : "+r"(x[k]) : "r"(y[k]));                   // IADD3 reg
                if (T == 15) { asm volatile("add.s32 %0, %0, %1;" : "+r"(x[k]) : "r"(c));
                               asm volatile("{.reg .b32 t; add.s16x2 t, %0, %1; max.s16x2 %0, t, %2;}" : "+r"(y[k]) : "r"(b), "r"(c)); } // IADD reg + VIADDMNMX
                if (T == 16) { asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(x[k]) : "r"(b), "r"(c));
                               asm volatile("max.s16x2 %0, %0, %1;" : "+r"(y[k]) : "r"(c)); }                // LOP3 + VIMNMX
                if (T == 17) asm volatile("add.s16x2 %0, %0, %1;" : "+r"(x[k]) : "r"(y[k]));                 // add.s16x2 alone -> ?
            }
        }
    }
    uint32_t s = 0;
#pragma unroll
    for (int k = 0; k < CHAINS; k++) s += x[k] ^ y[k];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int T>
static void run(const char* name, int per_iter_instr, uint32_t* out, int sms, double clk_hz) {
    int iters = 2000;
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    probe<T><<<sms * 8, 256>>>(out, 10, 3, 5); cudaDeviceSynchronize();
    cudaEventRecord(a); probe<T><<<sms * 8, 256>>>(out, iters, 3, 5); cudaEventRecord(b); cudaEventSynchronize(b);
    float ms; cudaEventElapsedTime(&ms, a, b);
    double warp_instr = (double)sms * 8 * 8 /*warps*/ * iters * REP * CHAINS * per_iter_instr;
    double per_clk_smsp = warp_instr / (ms * 1e-3) / clk_hz / (sms * 4);
    printf("%-28s %2d instr/slot  %8.3f ms  %.3f warp-instr/clk/SMSP  (%.2f T lane-op/s)\n", name, per_iter_instr, ms, per_clk_smsp, warp_instr * 32 / (ms * 1e-3) / 1e12);
}

int main() {
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    int sms = p.multiProcessorCount; double clk = p.clockRate * 1e3;
    uint32_t* out; cudaMalloc(&out, sms * 8 * 256 * 4);
    printf("%s, %d SMs, %.0f MHz nominal\n", p.name, sms, clk / 1e6);
    run<0>("VIADD(imm)", 1, out, sms, clk);
    run<14>("IADD(reg)", 1, out, sms, clk);
    run<17>("add.s16x2", 1, out, sms, clk);
    run<1>("VIMNMX.S16x2", 1, out, sms, clk);
    run<12>("VIMNMX.U16x2", 1, out, sms, clk);
    run<13>("VIMNMX3.S16x2", 1, out, sms, clk);
    run<2>("VIADDMNMX.S16x2", 1, out, sms, clk);
    run<4>("IMAD", 1, out, sms, clk);
    run<5>("LOP3", 1, out, sms, clk);
    run<6>("PRMT", 1, out, sms, clk);
    run<3>("VIADD->VIMNMX (dep)", 2, out, sms, clk);
    run<7>("VIADDMNMX + IMAD", 2, out, sms, clk);
    run<8>("VIADDMNMX + VIMNMX", 2, out, sms, clk);
    run<9>("VIADDMNMX + VIADD", 2, out, sms, clk);
    run<15>("VIADDMNMX + IADD(reg)", 2, out, sms, clk);
    run<10>("VIMNMX + IMAD", 2, out, sms, clk);
    run<11>("VIADD + IMAD", 2, out, sms, clk);
    run<16>("LOP3 + VIMNMX", 2, out, sms, clk);
    return 0;
}
