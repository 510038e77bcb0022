// coissue_probe.cu - which instruction kinds are (nearly) free next to the DPX recurrence of sw_score_kernel?
// Base = the 5.5-op DPX recurrence per row (no substitution lookup); EXTRA = n additional instructions of one
// kind per row on an independent accumulator. Reports clocks per row-step per warp.
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>

#define R 19
enum { K_NONE, K_IMAD, K_IMADIADD, K_LOP3, K_PRMT, K_FFMA, K_HFMA2, K_VIMNMX2, K_IADD3, K_SHF, K_IMADWIDE, K_IDP4A, K_MOV };

template <int KIND, int N>
__global__ void __launch_bounds__(128) mix(uint32_t* out, int iters, uint32_t bb, uint32_t go2, uint32_t nge2,
                                           uint32_t one, uint32_t three, const uint32_t* rtab) {
    uint32_t H[R], E[R], Q[R], X[R];
#pragma unroll
    for (int k = 0; k < R; k++) {
        H[k] = bb; E[k] = 0;
        Q[k] = rtab[1024 + ((threadIdx.x * R + k) & 1023)];
        X[k] = Q[k] * 3 + k;
    }
    uint32_t F = 0, cm = bb, diag = bb, best = bb;
    int col = 0;
    for (int i = 0; i < iters; i++) {
        uint32_t r0 = rtab[(i + threadIdx.x) & 1023];
        diag = bb; F = 0; cm = bb;
#pragma unroll
        for (int k = 0; k < R; k++) {
#pragma unroll
            for (int e = 0; e < N; e++) {
                uint32_t& x = X[k];
                if (KIND == K_IMAD) asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(x) : "r"(three), "r"(r0));
                if (KIND == K_IMADIADD) asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(x) : "r"(one), "r"(r0));
                if (KIND == K_LOP3) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(x) : "r"(Q[k]), "r"(r0));
                if (KIND == K_PRMT) asm volatile("prmt.b32 %0, %0, %1, %2;" : "+r"(x) : "r"(Q[k]), "r"(r0));
                if (KIND == K_FFMA) asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+r"(x) : "r"(Q[k]), "r"(r0));
                if (KIND == K_HFMA2) asm volatile("fma.rn.f16x2 %0, %0, %1, %2;" : "+r"(x) : "r"(Q[k]), "r"(r0));
                if (KIND == K_VIMNMX2) asm volatile("max.s16x2 %0, %0, %1;" : "+r"(x) : "r"(r0));
                if (KIND == K_IADD3) asm volatile("add.u32 %0, %0, %1;" : "+r"(x) : "r"(r0));
                if (KIND == K_SHF) asm volatile("shf.l.wrap.b32 %0, %0, %1, %2;" : "+r"(x) : "r"(Q[k]), "r"(one));
                if (KIND == K_IMADWIDE) { unsigned long long w; asm volatile("mad.wide.u32 %0, %1, %2, %3;" : "=l"(w) : "r"(x), "r"(three), "l"((unsigned long long)r0)); x = (uint32_t)w ^ (uint32_t)(w >> 32); }
                if (KIND == K_IDP4A) asm volatile("dp4a.u32.u32 %0, %0, %1, %2;" : "+r"(x) : "r"(Q[k]), "r"(r0));
                if (KIND == K_MOV) asm volatile("mov.b32 %0, %1;" : "=r"(x) : "r"(r0));
            }
            uint32_t s = Q[k];
            uint32_t m = __vimax3_s16x2(E[k], F, bb);
            uint32_t h = __viaddmax_s16x2(diag, s, m);
            diag = H[k];
            H[k] = h;
            uint32_t hg = h - go2;
            E[k] = __viaddmax_s16x2(E[k], nge2, hg);
            F = __viaddmax_s16x2(F, nge2, hg);
            if (k & 1) cm = __vimax3_s16x2(cm, h, E[k - 1]);
        }
        uint32_t nb = __vmaxs2(best, cm);
        if (nb != best) col = i;
        best = nb;
    }
    uint32_t acc = best + F + diag + col;
#pragma unroll
    for (int k = 0; k < R; k++) acc ^= X[k];
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}

static uint32_t *g_out, *g_tab;
static int g_sms; static double g_clk;
template <int KIND, int N>
static void run(const char* name) {
    int iters = 8000;
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    mix<KIND, N><<<g_sms * 16, 128>>>(g_out, 50, 0x00060006u, 0x00060006u, 0xffffffffu, 1, 3, g_tab);
    cudaDeviceSynchronize();
    cudaEventRecord(a);
    mix<KIND, N><<<g_sms * 16, 128>>>(g_out, iters, 0x00060006u, 0x00060006u, 0xffffffffu, 1, 3, g_tab);
    cudaEventRecord(b); cudaEventSynchronize(b);
    float ms; cudaEventElapsedTime(&ms, a, b);
    double clk = g_clk * ms * 1e-3 * g_sms * 4 / ((double)g_sms * 16 * 4 * iters * R);
    printf("%-12s x%d : %8.3f ms  %6.2f clk per row-step per warp\n", name, N, ms, clk);
}
#define RUN3(K, name) run<K, 1>(name); run<K, 2>(name); run<K, 3>(name);

int main() {
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    g_sms = p.multiProcessorCount; g_clk = p.clockRate * 1e3;
    cudaMalloc(&g_out, g_sms * 16 * 128 * 4);
    uint32_t h[2048];
    for (int i = 0; i < 2048; i++) h[i] = (i * 2654435761u) >> 7;
    cudaMalloc(&g_tab, 8192); cudaMemcpy(g_tab, h, 8192, cudaMemcpyHostToDevice);
    run<K_NONE, 0>("base");
    RUN3(K_IMAD, "IMAD") RUN3(K_IMADIADD, "IMAD(x1)") RUN3(K_LOP3, "LOP3") RUN3(K_PRMT, "PRMT") RUN3(K_FFMA, "FFMA")
    RUN3(K_HFMA2, "HFMA2") RUN3(K_VIMNMX2, "VIMNMX2") RUN3(K_IADD3, "IADD") RUN3(K_SHF, "SHF") RUN3(K_IMADWIDE, "IMAD.WIDE")
    RUN3(K_IDP4A, "IDP4A") RUN3(K_MOV, "MOV")
    return 0;
}
