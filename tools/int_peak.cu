// int_peak.cu - on-box microbenchmark of the integer pipes the score kernel lives on (SURVEY.md §8d:
// "The INT32 lane rate must be measured on the box"). Prints lane-ops/s for dependent-free streams of
// IADD3, LOP3, PRMT, VIMNMX3.S16x2, VIADDMNMX.S16x2 and the kernel's own 7.5-op cell mix.
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>

template <int OP>
__global__ void __launch_bounds__(256) bench(uint32_t* out, int iters, uint32_t a0, uint32_t b0) {
    uint32_t x[8];
#pragma unroll
    for (int k = 0; k < 8; k++) x[k] = a0 + threadIdx.x * 7 + k;
    uint32_t b = b0 + threadIdx.x, c = b0 * 3 + 1;
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int r = 0; r < 8; r++) {
#pragma unroll
            for (int k = 0; k < 8; k++) {
                // asm volatile: the compiler can neither fold the 64 x iters repetitions into one multiply-add (which is what
                // happened to the plain C++ forms of round 1: 1787 and 9291 "T lane-op/s") nor drop them
                if (OP == 0) asm volatile("{.reg .u32 t; add.u32 t, %1, %2; add.u32 %0, t, %3;}" : "=r"(x[k]) : "r"(x[k]), "r"(b), "r"(c));  // one IADD3
                if (OP == 1) asm volatile("lop3.b32 %0, %1, %2, %3, 0x96;" : "=r"(x[k]) : "r"(x[k]), "r"(b), "r"(c));  // LOP3 (xor3)
                if (OP == 2) asm volatile("prmt.b32 %0, %1, %2, %3;" : "=r"(x[k]) : "r"(x[k]), "r"(b), "r"(c));
                if (OP == 3) x[k] = __vimax3_s16x2(x[k], b, c);
                if (OP == 4) x[k] = __viaddmax_s16x2(x[k], b, c);
                if (OP == 5) x[k] = __vmaxs2(x[k], b);
                if (OP == 6) asm volatile("mad.lo.u32 %0, %1, %2, %3;" : "=r"(x[k]) : "r"(x[k]), "r"(b), "r"(c));  // IMAD (FMA pipe)
            }
        }
    }
    uint32_t s = 0;
#pragma unroll
    for (int k = 0; k < 8; k++) s += x[k];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// the cell mix of sw_score_kernel: 8 independent "rows" per thread
__global__ void __launch_bounds__(128) cellmix(uint32_t* out, int iters, uint32_t pa, uint32_t pb, uint32_t bb, uint32_t go2, uint32_t nge2) {
    uint32_t H[8], E[8], Q[8];
#pragma unroll
    for (int k = 0; k < 8; k++) { H[k] = bb; E[k] = 0; Q[k] = (threadIdx.x * 13 + k * 7) & 0x3333; }
    uint32_t F = 0, rs = threadIdx.x & 0x3333, cm = bb, diag = bb;
    for (int i = 0; i < iters; i++) {
        rs = (rs * 5 + 1) & 0x3333;
#pragma unroll
        for (int k = 0; k < 8; k++) {
            uint32_t s;
            asm("prmt.b32 %0, %1, %2, %3;" : "=r"(s) : "r"(pa), "r"(pb), "r"(Q[k] ^ rs ^ 0x8080u));
            uint32_t m = __vimax3_s16x2(E[k], F, bb);
            uint32_t h = __viaddmax_s16x2(diag, s, m);
            diag = H[k]; H[k] = h;
            uint32_t hg = h - go2;
            E[k] = __viaddmax_s16x2(E[k], nge2, hg);
            F = __viaddmax_s16x2(F, nge2, hg);
            if (k & 1) cm = __vimax3_s16x2(cm, H[k - 1], h);
        }
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = cm + F + diag;
}

template <typename L>
static double time_ms(L launch) {
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    launch(); cudaDeviceSynchronize();
    cudaEventRecord(a); launch(); cudaEventRecord(b); cudaEventSynchronize(b);
    float ms; cudaEventElapsedTime(&ms, a, b); return ms;
}

int main() {
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    int sms = p.multiProcessorCount;
    uint32_t* out; cudaMalloc(&out, sms * 16 * 256 * 4);
    const char* names[] = {"IADD3", "LOP3", "PRMT", "VIMNMX3.S16x2", "VIADDMNMX.S16x2", "VIMNMX.S16x2", "IMAD"};
    int iters = 4000;
    printf("{\"gpu\": \"%s\", \"sms\": %d, \"clock_khz\": %d", p.name, sms, p.clockRate);
    double ops = (double)sms * 8 * 256 * (double)iters * 64;
    double ms;
    ms = time_ms([&] { bench<0><<<sms * 8, 256>>>(out, iters, 1, 2); }); printf(", \"%s_Tlaneops\": %.2f", names[0], ops / ms / 1e9);
    ms = time_ms([&] { bench<1><<<sms * 8, 256>>>(out, iters, 1, 2); }); printf(", \"%s_Tlaneops\": %.2f", names[1], ops / ms / 1e9);
    ms = time_ms([&] { bench<2><<<sms * 8, 256>>>(out, iters, 1, 2); }); printf(", \"%s_Tlaneops\": %.2f", names[2], ops / ms / 1e9);
    ms = time_ms([&] { bench<3><<<sms * 8, 256>>>(out, iters, 1, 2); }); printf(", \"%s_Tlaneops\": %.2f", names[3], ops / ms / 1e9);
    ms = time_ms([&] { bench<4><<<sms * 8, 256>>>(out, iters, 1, 2); }); printf(", \"%s_Tlaneops\": %.2f", names[4], ops / ms / 1e9);
    ms = time_ms([&] { bench<5><<<sms * 8, 256>>>(out, iters, 1, 2); }); printf(", \"%s_Tlaneops\": %.2f", names[5], ops / ms / 1e9);
    ms = time_ms([&] { bench<6><<<sms * 8, 256>>>(out, iters, 1, 2); }); printf(", \"%s_Tlaneops\": %.2f", names[6], ops / ms / 1e9);
    // cell mix: 8 packed cells (=16 DP cells) per inner iteration per thread
    int it2 = 20000;
    double cells = (double)sms * 16 * 128 * (double)it2 * 16;
    ms = time_ms([&] { cellmix<<<sms * 16, 128>>>(out, it2, 0xfcfcfc01u, 0xfcfcfcfcu, 0x00060006u, 0x00060006u, 0xffffffffu); });
    printf(", \"cellmix_TCUPS\": %.3f, \"cellmix_Tlaneops_at_7.5\": %.2f}\n", cells / ms / 1e9, cells / 2 * 7.5 / ms / 1e9);
    return 0;
}
