// bulk_copy.cuh -- the bulk asynchronous copy (cp.async.bulk, SASS UBLKCP) and its mbarrier, as used by the kernels
// that stage global memory into shared memory without spending a register or a load instruction per element.
#pragma once

#include "packed_math.cuh"

namespace lora_b200 {
namespace {

PM_DEV unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }

PM_DEV void mbar_init(unsigned long long* bar, unsigned count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
PM_DEV void mbar_expect_tx(unsigned long long* bar, unsigned bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
PM_DEV void mbar_wait(unsigned long long* bar, unsigned parity)
{
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(smem_u32(bar)), "r"(parity)
        : "memory");
}
PM_DEV void bulk_load(void* dst, const void* src, unsigned bytes, unsigned long long* bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)),
                 "l"(src), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}

// One instruction asks the memory system to pull [p, p + bytes) into L2 (16-byte granules that lie fully inside the
// range).  Issued by one thread per window for the window AFTER the one being loaded: by the time its samples are
// wanted they come from L2 instead of HBM, without holding registers, shared memory or a barrier for them.
PM_DEV void l2_prefetch(const void* p, size_t bytes)
{
    const unsigned long long a = (unsigned long long)p;
    const unsigned long long lo = (a + 15ull) & ~15ull, hi = (a + bytes) & ~15ull;
    if (hi > lo)
        asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(lo), "r"((unsigned)(hi - lo)) : "memory");
}

} // namespace
} // namespace lora_b200
