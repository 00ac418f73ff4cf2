// ubench_pipes.cu -- issue-rate microbenchmarks behind DESIGN.md's instruction budget (not product code).
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o ubench_pipes scripts/ubench_pipes.cu && ./ubench_pipes
//
// Each test runs 8 independent dependency chains per thread, 1024 threads per SM (2 CTAs x 512), one wave,
// and reports warp-instructions per cycle per SM (4.0 = every sub-partition issues every cycle).  It also
// checks that the packed un-fused complex multiply used by the bit-faithful FFT (mul.f32x2, mul.f32x2, then
// an fma.f32x2 by a run-time 1.0 standing in for the add, because ptxas contracts mul.rn.f32x2 + add.rn.f32x2 into
// FFMA2 even under --fmad=false) equals the scalar __fmul_rn/__fadd_rn result bit for bit.
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

typedef unsigned long long u64;
#define DEV __device__ __forceinline__
DEV u64 pk(float lo, float hi) { u64 r; asm("mov.b64 %0, {%1,%2};" : "=l"(r) : "f"(lo), "f"(hi)); return r; }
DEV float2 up(u64 v) { float2 r; asm("mov.b64 {%0,%1}, %2;" : "=f"(r.x), "=f"(r.y) : "l"(v)); return r; }
DEV u64 mul2(u64 a, u64 b) { u64 d; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }
DEV u64 add2(u64 a, u64 b) { u64 d; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }
DEV u64 fma2(u64 a, u64 b, u64 c) { u64 d; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c)); return d; }

constexpr int ITERS = 32768;
constexpr int CH = 8;

template <int OP>
__global__ void __launch_bounds__(512) bench(float* out, long long* cycles, float seed)
{
    float a[CH], b = seed, c = seed * 0.5f;
    double da[CH], db = seed, dc = seed * 0.25;
    u64 pa[CH], pb = pk(seed, seed + 1.0f), pc = pk(0.5f, 0.25f);
    __shared__ float2 sm[1024];
    sm[threadIdx.x] = make_float2(seed, seed);
    sm[threadIdx.x + 512] = make_float2(seed, seed);
    __syncthreads();
#pragma unroll
    for (int i = 0; i < CH; ++i) { a[i] = threadIdx.x + i; da[i] = threadIdx.x + i; pa[i] = pk(a[i], a[i] + 0.5f); }
    const long long t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int i = 0; i < CH; ++i) {
            if (OP == 0) a[i] = fmaf(a[i], b, c);                                  // FFMA
            if (OP == 1) pa[i] = fma2(pa[i], pb, pc);                              // FFMA2
            if (OP == 2) pa[i] = (i & 1) ? mul2(pa[i], pb) : add2(pa[i], pc);      // FMUL2 / FADD2
            if (OP == 3) da[i] = fma(da[i], db, dc);                               // DFMA
            if (OP == 4) da[i] = (i & 1) ? __dmul_rn(da[i], db) : __dadd_rn(da[i], dc);
            if (OP == 5) { pa[i] = fma2(pa[i], pb, pc); da[i] = fma(da[i], db, dc); } // 1:1 FFMA2 + DFMA
            if (OP == 6) { a[i] = fmaf(a[i], b, c); a[i] = fmaxf(a[i], c); }       // FFMA + FMNMX (alu pipe)
            if (OP == 7) a[i] = __sinf(a[i]);                                      // MUFU.SIN (+ range mul)
            if (OP == 8) a[i] = (float)__dadd_rn((double)a[i], db);                // F2F.F64.F32 + DADD + F2F.F32.F64
            if (OP == 9) { pa[i] = fma2(pa[i], pb, pc); float2 v = sm[(threadIdx.x + i * 32 + it) & 1023]; pb = pk(v.x, up(pb).y); }
            if (OP == 10) { pa[i] = fma2(pa[i], pb, pc); pa[i] = fma2(pa[i], pb, pc); da[i] = fma(da[i], db, dc); } // 2:1
            if (OP == 11) a[i] = __fadd_rn(__fmul_rn(a[i], b), c);                 // FMUL + FADD un-fused
            if (OP == 12) { pa[i] = fma2(pa[i], pb, pc); a[i] = fmaf(a[i], b, c); } // FFMA2 + FFMA
        }
    }
    const long long t1 = clock64();
    float acc = 0.0f;
#pragma unroll
    for (int i = 0; i < CH; ++i) acc += a[i] + (float)da[i] + up(pa[i]).x + up(pa[i]).y;
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc + up(pb).x;
    if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
}

// numeric check: packed un-fused complex multiply vs scalar un-fused
__global__ void check_cmul(const float2* x, const float2* w, float2* packed, float2* scalar, float2* fused, int n, float opaque_one)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float2 a = x[i], b = w[i];
    const u64 one = pk(opaque_one, opaque_one); // a run-time 1.0 ptxas cannot fold away
    const u64 t1 = mul2(pk(a.x, a.y), pk(b.x, b.x));
    const u64 t2 = mul2(pk(a.y, a.x), pk(-b.y, b.y));
    packed[i] = up(fma2(t1, one, t2));
    fused[i] = up(add2(t1, t2)); // what ptxas makes of mul.rn + add.rn
    scalar[i] = make_float2(__fsub_rn(__fmul_rn(a.x, b.x), __fmul_rn(a.y, b.y)),
                            __fadd_rn(__fmul_rn(a.x, b.y), __fmul_rn(a.y, b.x)));
}

template <int OP>
static void run(const char* name, int instr_per_iter_per_chain)
{
    int dev = 0, sms = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const int blocks = sms * 2, threads = 512;
    float* out; long long* cyc;
    cudaMalloc(&out, sizeof(float) * blocks * threads);
    cudaMalloc(&cyc, sizeof(long long) * blocks);
    bench<OP><<<blocks, threads>>>(out, cyc, 1.0001f);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0);
    bench<OP><<<blocks, threads>>>(out, cyc, 1.0001f);
    cudaEventRecord(e1);
    cudaDeviceSynchronize();
    float ms = 0; cudaEventElapsedTime(&ms, e0, e1);
    std::vector<long long> h(blocks);
    cudaMemcpy(h.data(), cyc, sizeof(long long) * blocks, cudaMemcpyDeviceToHost);
    double mean = 0; for (auto v : h) mean += (double)v; mean /= blocks;
    const double warp_instr_per_sm = 2.0 * (threads / 32) * (double)ITERS * CH * instr_per_iter_per_chain;
    // two views: per-CTA clock64 span, and the event time of the whole launch (Ginstr/s per SM; divide by the
    // SM clock in GHz to get instr/clk)
    printf("%-28s %6.3f warp-instr/clk/SM by clock64   %7.3f G warp-instr/s/SM by events (%.3f ms, cyc/ms -> %.0f MHz)\n",
           name, warp_instr_per_sm / mean, warp_instr_per_sm / (ms * 1e-3) / 1e9, ms, mean / (ms * 1e3));
    cudaFree(out); cudaFree(cyc);
}

int main()
{
    run<0>("FFMA", 1);
    run<11>("FMUL+FADD (un-fused)", 2);
    run<1>("FFMA2", 1);
    run<2>("FMUL2/FADD2 alternating", 1);
    run<12>("FFMA2 + FFMA 1:1", 2);
    run<3>("DFMA", 1);
    run<4>("DMUL/DADD alternating", 1);
    run<5>("FFMA2 + DFMA 1:1", 2);
    run<10>("FFMA2 x2 + DFMA", 3);
    run<6>("FFMA + FMNMX 1:1", 2);
    run<7>("MUFU.SIN (__sinf)", 2);
    run<8>("F2F.F64.F32+DADD+F2F.F32.F64", 3);
    run<9>("FFMA2 + LDS.64 1:1", 2);

    const int n = 1 << 20;
    std::vector<float2> hx(n), hw(n);
    srand(1);
    for (int i = 0; i < n; ++i) {
        hx[i] = make_float2((float)rand() / RAND_MAX * 4 - 2, (float)rand() / RAND_MAX * 4 - 2);
        hw[i] = make_float2((float)rand() / RAND_MAX * 2 - 1, (float)rand() / RAND_MAX * 2 - 1);
    }
    float2 *dx, *dw, *dp, *ds, *df;
    cudaMalloc(&dx, n * 8); cudaMalloc(&dw, n * 8); cudaMalloc(&dp, n * 8); cudaMalloc(&ds, n * 8); cudaMalloc(&df, n * 8);
    cudaMemcpy(dx, hx.data(), n * 8, cudaMemcpyHostToDevice);
    cudaMemcpy(dw, hw.data(), n * 8, cudaMemcpyHostToDevice);
    check_cmul<<<n / 256, 256>>>(dx, dw, dp, ds, df, n, 1.0f);
    std::vector<float2> hp(n), hs(n), hf(n);
    cudaMemcpy(hp.data(), dp, n * 8, cudaMemcpyDeviceToHost);
    cudaMemcpy(hs.data(), ds, n * 8, cudaMemcpyDeviceToHost);
    cudaMemcpy(hf.data(), df, n * 8, cudaMemcpyDeviceToHost);
    int bad_p = 0, bad_f = 0;
    for (int i = 0; i < n; ++i) {
        bad_p += memcmp(&hp[i], &hs[i], 8) != 0;
        bad_f += memcmp(&hf[i], &hs[i], 8) != 0;
    }
    printf("packed cmul (mul2,mul2,fma2 by 1.0) vs scalar un-fused: %d / %d differ\n", bad_p, n);
    printf("packed cmul (mul2,mul2,add2)        vs scalar un-fused: %d / %d differ (ptxas contraction)\n", bad_f, n);
    return cudaGetLastError() != cudaSuccess;
}
