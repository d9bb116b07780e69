// Integer pipe throughput on one SM sub-partition mix: how many warp instructions per cycle per
// SMSP for streams of VIMNMX (alu pipe), IMAD (fma pipe), VIMNMX3, IMAD.WIDE and mixes of them.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o int_pipes int_pipes.cu && ./int_pipes
#include <cstdio>
#include <cuda_runtime.h>

constexpr int CH = 8;        // independent chains per thread
constexpr int ITER = 4096;

template <int KIND>
__global__ void __launch_bounds__(512, 1) k(int *out, int a, int b, int one, long long *cycles) {
    int x[CH];
    long long w[CH];
#pragma unroll
    for (int i = 0; i < CH; ++i) { x[i] = threadIdx.x + i; w[i] = i; }
    const long long t0 = clock64();
    for (int it = 0; it < ITER; ++it) {
#pragma unroll
        for (int i = 0; i < CH; ++i) {
            if (KIND == 0) { x[i] = max(x[i], a); x[i] = min(x[i], b); }                           // 2 alu
            if (KIND == 1) { x[i] = x[i] * a + b; x[i] = x[i] * b + a; }                           // 2 fma-pipe
            if (KIND == 2) { x[i] = max(x[i], a); x[i] = x[i] * a + b; }                           // 1 + 1
            if (KIND == 3) { x[i] = __vimax3_s32(x[i], a, b); x[i] = __vimin3_s32(x[i], b, one); } // 2 vimnmx3
            if (KIND == 4) { w[i] = (long long)x[i] * (long long)one + w[i]; x[i] = max(x[i], a); } // wide + alu
            if (KIND == 5) { w[i] = (long long)x[i] * (long long)one + w[i];
                             w[i] = (long long)a * (long long)one + w[i]; }                        // 2 wide
            if (KIND == 6) { x[i] = max(x[i], a); x[i] = min(x[i], b); x[i] = x[i] * a + b; }       // 2 alu + 1 fma
            if (KIND == 7) { x[i] = max(x[i], a); x[i] = x[i] * a + b; x[i] = x[i] * b + a; }       // 1 alu + 2 fma
            if (KIND == 8) { x[i] = __vmins2(x[i], a); x[i] = __vmaxs2(x[i], b); }                  // VIMNMX.S16x2
            if (KIND == 9) { x[i] = __vimax3_s16x2(x[i], a, b); x[i] = __vimin3_s16x2(x[i], b, one); }
            if (KIND == 10) { x[i] = __vadd2(x[i], a); x[i] = __vadd2(x[i], b); }                   // VIADD.16x2
            if (KIND == 11) { x[i] = __viaddmin_u16x2(x[i], a, b); x[i] = __viaddmax_u16x2(x[i], b, one); }
            if (KIND == 12) { x[i] = __vminu2(x[i], a); x[i] = x[i] ^ b; }                          // VIMNMX.U16x2 + LOP3
        }
    }
    const long long t1 = clock64();
    int s = 0;
#pragma unroll
    for (int i = 0; i < CH; ++i) s += x[i] + (int)w[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cycles = t1 - t0;
}

template <int KIND>
void run(const char *name, int per_iter, int warps) {
    int *out; long long *cyc, h;
    cudaMalloc(&out, 148 * 512 * sizeof(int)); cudaMalloc(&cyc, 8);
    k<KIND><<<148, 32 * warps>>>(out, 7, 1 << 30, 1, cyc);
    k<KIND><<<148, 32 * warps>>>(out, 7, 1 << 30, 1, cyc);
    cudaDeviceSynchronize();
    cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
    const double inst = (double)ITER * CH * per_iter * (warps / 4.0);   // warp instructions per SMSP
    printf("%-28s warps/SM=%2d  %.3f warp-inst/cycle/SMSP\n", name, warps, inst / (double)h);
    cudaFree(out); cudaFree(cyc);
}

int main() {
    for (int warps : {4, 16}) {
        run<0>("VIMNMX x2", 2, warps);
        run<1>("IMAD x2", 2, warps);
        run<2>("VIMNMX + IMAD", 2, warps);
        run<3>("VIMNMX3 x2", 2, warps);
        run<4>("IMAD.WIDE + VIMNMX", 2, warps);
        run<5>("IMAD.WIDE x2", 2, warps);
        run<6>("2 VIMNMX + 1 IMAD", 3, warps);
        run<7>("1 VIMNMX + 2 IMAD", 3, warps);
        run<8>("VIMNMX.S16x2 x2", 2, warps);
        run<9>("VIMNMX3.S16x2 x2", 2, warps);
        run<10>("VIADD.16x2 x2", 2, warps);
        run<11>("VIADDMNMX.U16x2 x2", 2, warps);
        run<12>("VIMNMX.U16x2 + LOP3", 2, warps);
    }
    return 0;
}
