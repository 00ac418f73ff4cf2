// packed_math.cuh -- complex float arithmetic on Blackwell's packed fp32 pipe (FADD2 / FMUL2 / FFMA2).
//
// A complex sample lives in one 64-bit register pair, (re, im) = (lo, hi).  sm_100a executes add/mul/fma on
// both halves with ONE issue slot, and the SASS operand modifiers give swap (.LO_HI), partial negate (.NP)
// and scalar broadcast (.F32) for free, so a complex add is one instruction, a fused complex multiply two,
// and a multiplication by -j folds into the operand of the add that consumes it.  Measured on B200
// (scripts/ubench_pipes.cu, profiles/r1_ubench_pipes.txt): the packed forms have the same FLOP rate as
// scalar FFMA (the fp32 datapath is 128 lanes/clk/SM either way) but need half the issue slots, which is
// what lets the loads, shared-memory exchanges, fp64 sincos and integer work of the receive kernel co-issue.
//
// Two flavours of every product-consuming operation:
//   FAITH = true   every multiply and every add is an individually rounded IEEE operation, in the order
//                  the reference's C++ performs it (std::complex operator* as g++ -O2 emits it for finite
//                  operands: (ac - bd, ad + bc)).  Bit-identical to the CPU.
//   FAITH = false  multiplies feed fused multiply-adds.  Same butterfly graph, ~1e-7 relative difference.
//
// ptxas contracts `mul.rn.f32x2` + `add.rn.f32x2` into FFMA2 even with --fmad=false (seen in SASS,
// CUDA 12.9).  The faithful forms therefore never add a raw product with add.f32x2: the add is written
// fma(t1, one, t2) with `one` a RUN-TIME 1.0f (a kernel argument), which is exactly fl(t1 + t2) and
// cannot be folded.  scripts/ubench_pipes.cu checks this against scalar __fmul_rn/__fadd_rn on the GPU.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

namespace lora_b200 {

typedef unsigned long long c64; // packed complex<float>: low word re, high word im

#define PM_DEV __device__ __forceinline__

PM_DEV c64 pk(float lo, float hi) { c64 r; asm("mov.b64 %0, {%1,%2};" : "=l"(r) : "f"(lo), "f"(hi)); return r; }
PM_DEV float2 up(c64 v) { float2 r; asm("mov.b64 {%0,%1}, %2;" : "=f"(r.x), "=f"(r.y) : "l"(v)); return r; }
PM_DEV float re_of(c64 v) { return up(v).x; }
PM_DEV float im_of(c64 v) { return up(v).y; }

PM_DEV c64 mul2(c64 a, c64 b) { c64 d; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }
PM_DEV c64 add2(c64 a, c64 b) { c64 d; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }
PM_DEV c64 sub2(c64 a, c64 b) { c64 d; asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }
PM_DEV c64 fma2(c64 a, c64 b, c64 c) { c64 d; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c)); return d; }

// a + b and a - b where a or b may be a raw (un-added) product: contraction-proof in the faithful flavour.
template <bool FAITH> PM_DEV c64 cadd_g(c64 a, c64 b, float one) { return FAITH ? fma2(a, pk(one, one), b) : add2(a, b); }
template <bool FAITH> PM_DEV c64 csub_g(c64 a, c64 b, float one)
{
    if (!FAITH) return sub2(a, b);
    const float2 B = up(b);
    return fma2(a, pk(one, one), pk(-B.x, -B.y)); // a - b == a + (-b) exactly
}

// Complex multiply a * w.
//   faithful: t1 = (a.re*w.re, a.im*w.re), t2 = (-(a.im*w.im), a.re*w.im), result fl(t1 + t2)
//             = (fl(fl(ac) - fl(bd)), fl(fl(bc) + fl(ad)))  -- the reference's bits (negation is exact, + commutes)
//   fast:     the first product is fused into the add.
template <bool FAITH> PM_DEV c64 cmul(c64 a, c64 w, float one)
{
    const float2 A = up(a), W = up(w);
    const c64 t2 = mul2(pk(-A.y, A.x), pk(W.y, W.y));
    if (FAITH) return fma2(mul2(a, pk(W.x, W.x)), pk(one, one), t2);
    return fma2(a, pk(W.x, W.x), t2);
}

// real scale of both halves (exact per-component multiply in both flavours)
PM_DEV c64 cscale(c64 a, float s) { return mul2(a, pk(s, s)); }

// a * (-j) = (im, -re): kissfft's forward-transform rotation (kissfft.hh:176-177).  A pure operand swizzle.
PM_DEV c64 mul_mj(c64 d) { const float2 D = up(d); return pk(D.y, -D.x); }

// |a|^2 the way LoRaDetector.hpp:49-50 forms it: fl(fl(re*re) + fl(im*im)); fused in the fast flavour.
template <bool FAITH> PM_DEV float mag2(c64 a)
{
    const float2 A = up(a);
    if (FAITH) return __fadd_rn(__fmul_rn(A.x, A.x), __fmul_rn(A.y, A.y));
    return fmaf(A.x, A.x, A.y * A.y);
}

} // namespace lora_b200
