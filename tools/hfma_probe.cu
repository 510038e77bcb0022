// hfma_probe.cu - HFMA2 issue rate (normal vs denormal operands) alone and co-issued with VIADDMNMX.
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <cstdio>
#include <cstdint>
#define CH 8
template <int T>
__global__ void __launch_bounds__(256) k(uint32_t* out, int iters, uint32_t a0, uint32_t m, uint32_t c0, uint32_t y0) {
    uint32_t x[CH], y[CH];
#pragma unroll
    for (int i = 0; i < CH; i++) { x[i] = a0 + ((threadIdx.x + i) & 3); y[i] = y0 + i; }
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int r = 0; r < 16; r++) {
#pragma unroll
            for (int i = 0; i < CH; i++) {
                if (T == 0 || T == 2) asm volatile("fma.rn.f16x2 %0, %0, %1, %2;" : "+r"(x[i]) : "r"(m), "r"(c0));
                if (T == 1) asm volatile("fma.rn.relu.f16x2 %0, %0, %1, %2;" : "+r"(x[i]) : "r"(m), "r"(c0));
                if (T == 2 || T == 3) asm volatile("{.reg .b32 t; add.s16x2 t, %0, %1; max.s16x2 %0, t, %2;}" : "+r"(y[i]) : "r"(m), "r"(c0));
                if (T == 4) { asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+r"(x[i]) : "r"(m), "r"(c0)); }
                if (T == 5) { asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+r"(x[i]) : "r"(m), "r"(c0));
                              asm volatile("{.reg .b32 t; add.s16x2 t, %0, %1; max.s16x2 %0, t, %2;}" : "+r"(y[i]) : "r"(m), "r"(c0)); }
            }
        }
    }
    uint32_t s = 0;
#pragma unroll
    for (int i = 0; i < CH; i++) s += x[i] ^ y[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
template <int T>
static void run(const char* name, int n, uint32_t* out, int sms, double clk, uint32_t a0, uint32_t m, uint32_t c0) {
    int iters = 2000;
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    k<T><<<sms * 8, 256>>>(out, 10, a0, m, c0, 7); cudaDeviceSynchronize();
    cudaEventRecord(a); k<T><<<sms * 8, 256>>>(out, iters, a0, m, c0, 7); cudaEventRecord(b); cudaEventSynchronize(b);
    float ms; cudaEventElapsedTime(&ms, a, b);
    double wi = (double)sms * 8 * 8 * iters * 16 * CH * n;
    printf("%-36s %8.3f ms  %.3f warp-instr/clk/SMSP\n", name, ms, wi / (ms * 1e-3) / clk / (sms * 4));
}
static uint32_t h2(float a, float b) { __half2 h = __floats2half2_rn(a, b); return *(uint32_t*)&h; }
int main() {
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    int sms = p.multiProcessorCount; double clk = p.clockRate * 1e3;
    uint32_t* out; cudaMalloc(&out, sms * 8 * 256 * 4);
    // normal operands: x = x*0.5 + 1.0 (stays ~2); denormal: x = x*0.5 + 3 units
    run<0>("HFMA2 normal", 1, out, sms, clk, h2(1, 1), h2(0.5f, 0.5f), h2(1, 1));
    run<0>("HFMA2 denormal", 1, out, sms, clk, 0x00050005u, h2(0.5f, 0.5f), 0x00030003u);
    run<1>("HFMA2.RELU denormal", 1, out, sms, clk, 0x00050005u, h2(0.5f, 0.5f), 0x00030003u);
    run<3>("VIADDMNMX", 1, out, sms, clk, 0, 0x00010001u, 0x00030003u);
    run<2>("HFMA2 normal + VIADDMNMX", 2, out, sms, clk, h2(1, 1), h2(0.5f, 0.5f), h2(1, 1));
    run<2>("HFMA2 denormal + VIADDMNMX", 2, out, sms, clk, 0x00050005u, h2(0.5f, 0.5f), 0x00030003u);
    float f1 = 1.0f, fh = 0.5f;
    run<4>("FFMA", 1, out, sms, clk, *(uint32_t*)&f1, *(uint32_t*)&fh, *(uint32_t*)&f1);
    run<5>("FFMA + VIADDMNMX", 2, out, sms, clk, *(uint32_t*)&f1, *(uint32_t*)&fh, *(uint32_t*)&f1);
    return 0;
}
