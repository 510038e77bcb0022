// cellmix2_probe.cu - candidate per-cell instruction mixes for sw_score_kernel (sm_100a), round 1 second pass.
//   0: LOP3 selector + PRMT score (current kernel)
//   1: IMAD selector (Q + Rsel through mad.lo with an opaque multiplier 1) + PRMT, diag biased by one extra add
//   2: same as 1 but the two adds written as plain C (compiler chooses the pipe)
//   3: no score op at all (DPX floor of the recurrence)
//   4: mode 1 without the column-max op (shows what the running maximum costs)
//   5: query profile in registers: s = PRMT(QA[k], QB[k], step selector), no LOP3 (two registers per row)
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>

#ifndef R
#define R 19
#endif
__device__ __forceinline__ uint32_t prmt(uint32_t a, uint32_t b, uint32_t s) {
    uint32_t d; asm("prmt.b32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(s)); return d;
}
__device__ __forceinline__ uint32_t imad(uint32_t a, uint32_t b, uint32_t c) {
    uint32_t d; asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c)); return d;
}

template <int MODE>
__global__ void __launch_bounds__(128) cellmix(uint32_t* out, int iters, uint32_t pa, uint32_t pb, uint32_t bb,
                                               uint32_t go2, uint32_t nge2, uint32_t amb2, uint32_t one, const uint32_t* rtab) {
    uint32_t H[R], E[R], Q[R], QB[MODE == 5 ? R : 1];
#pragma unroll
    for (int k = 0; k < R; k++) {
        H[k] = bb; E[k] = 0;
        int q = (int)(rtab[1024 + ((threadIdx.x * R + k) & 1023)]);  // opaque to the compiler
        Q[k] = MODE == 0 ? q * 0x11 * 0x0101u : q * 0x0101u;
        if (MODE == 5) { Q[k] = rtab[(threadIdx.x * 7 + k) & 1023] * 0x01010101u + q; QB[k] = rtab[(threadIdx.x * 5 + k * 3) & 1023] ^ (q << 8); }
    }
    uint32_t F = 0, cm = bb, diag = bb, best = bb;
    int col = 0;
    for (int i = 0; i < iters; i++) {
        uint32_t r0 = rtab[(i + threadIdx.x) & 1023];
        diag = bb; F = 0; cm = bb;
#pragma unroll
        for (int k = 0; k < R; k++) {
            uint32_t s;
            if (MODE == 0) s = prmt(pa, pb, Q[k] ^ r0);
            else if (MODE == 1 || MODE == 4) s = prmt(pa, pb, imad(Q[k], one, r0));
            else if (MODE == 2) s = prmt(pa, pb, Q[k] + r0);
            else if (MODE == 5) s = prmt(Q[k], QB[k], r0);
            else s = Q[k];
            uint32_t m = __vimax3_s16x2(E[k], F, bb);
            uint32_t h = __viaddmax_s16x2(diag, s, m);
            diag = H[k];
            if (MODE == 0 || MODE == 3) H[k] = h;
            else if (MODE == 1 || MODE == 4) H[k] = imad(h, one, amb2);
            else H[k] = h + amb2;
            uint32_t hg = (MODE == 1 || MODE == 4) ? imad(h, one, 0u - go2) : h - go2;
            E[k] = __viaddmax_s16x2(E[k], nge2, hg);
            F = __viaddmax_s16x2(F, nge2, hg);
            if (MODE != 4) { if (k & 1) cm = __vimax3_s16x2(cm, h, E[k - 1]); }
            else if (k == R - 1) cm = __vmaxs2(cm, h);
        }
        uint32_t nb = __vmaxs2(best, cm);
        if (nb != best) col = i;
        best = nb;
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = best + F + diag + col;
}

int main() {
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    int sms = p.multiProcessorCount;
    uint32_t* out; cudaMalloc(&out, sms * 16 * 128 * 4);
    uint32_t h_x[2048], h_a[2048];
    for (int i = 0; i < 1024; i++) {
        int r = (i * 7 + 3) % 4;
        h_x[i] = ((r * 0x11) ^ 0x80) * 0x0101u;
        h_a[i] = (0x80 | (3 - r)) * 0x0101u;
        h_x[1024 + i] = h_a[1024 + i] = (i * 5 + 1) % 4;
    }
    uint32_t *rx, *ra; cudaMalloc(&rx, 8192); cudaMalloc(&ra, 8192);
    cudaMemcpy(rx, h_x, 8192, cudaMemcpyHostToDevice); cudaMemcpy(ra, h_a, 8192, cudaMemcpyHostToDevice);
    int iters = 20000;
    const char* names[6] = {"LOP3+PRMT", "IMAD+PRMT+IMAD bias", "add+PRMT+add (compiler)", "no score op (DPX floor)", "IMAD+PRMT, no colmax", "profile PRMT (no LOP3)"};
    for (int mode = 0; mode < 6; mode++) {
        cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
        auto launch = [&](int it) {
            const uint32_t bb = 0x00060006u, go2 = 0x00060006u, nge2 = 0xffffffffu, amb2 = 0xfffcfffcu;
            if (mode == 0) cellmix<0><<<sms * 16, 128>>>(out, it, 0xfcfcfc01u, 0xfcfcfcfcu, bb, go2, nge2, amb2, 1, rx);
            if (mode == 1) cellmix<1><<<sms * 16, 128>>>(out, it, 0x05000000u, 0u, bb, go2, nge2, amb2, 1, ra);
            if (mode == 2) cellmix<2><<<sms * 16, 128>>>(out, it, 0x05000000u, 0u, bb, go2, nge2, amb2, 1, ra);
            if (mode == 3) cellmix<3><<<sms * 16, 128>>>(out, it, 0x05000000u, 0u, bb, go2, nge2, amb2, 1, ra);
            if (mode == 5) cellmix<5><<<sms * 16, 128>>>(out, it, 0x05000000u, 0u, bb, go2, nge2, amb2, 1, rx);
            if (mode == 4) cellmix<4><<<sms * 16, 128>>>(out, it, 0x05000000u, 0u, bb, go2, nge2, amb2, 1, ra);
        };
        launch(100); cudaDeviceSynchronize();
        cudaEventRecord(a); launch(iters); cudaEventRecord(b); cudaEventSynchronize(b);
        float ms; cudaEventElapsedTime(&ms, a, b);
        double cells = (double)sms * 16 * 128 * iters * R * 2;
        double clk = (double)p.clockRate * 1e3 * ms * 1e-3 * sms * 4 / ((double)sms * 16 * 4 * iters * R);
        printf("mode %d (%s): %.3f ms, %.3f TCUPS, %.2f clk per row-step per warp\n", mode, names[mode], ms, cells / ms / 1e9, clk);
    }
    return 0;
}
