// sgb_async.cuh -- thin wrappers over the sm_100a asynchronous-copy instructions used by the kernels
// (cp.async.bulk = 1-D TMA bulk copy, mbarrier transaction barriers).
#pragma once
// ---- TMA (bulk asynchronous copy) / mbarrier wrappers --------------------------------------
__device__ __forceinline__ unsigned sgb_smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void sgb_mbar_init(unsigned long long* bar, unsigned count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(sgb_smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void sgb_mbar_expect_tx(unsigned long long* bar, unsigned bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(sgb_smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void sgb_tma_load_1d(void* dst, const void* src, unsigned bytes, unsigned long long* bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(sgb_smem_u32(dst)), "l"(src), "r"(bytes), "r"(sgb_smem_u32(bar)) : "memory");
}
// bounded spin: a mis-programmed pipeline traps instead of hanging the device
__device__ __forceinline__ void sgb_mbar_wait(unsigned long long* bar, unsigned parity)
{
    unsigned done = 0;
    for (unsigned spin = 0; !done; ++spin) {
        asm volatile("{\n .reg .pred p;\n mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n selp.u32 %0, 1, 0, p;\n}"
                     : "=r"(done) : "r"(sgb_smem_u32(bar)), "r"(parity) : "memory");
        if (spin > (1u << 26)) __trap();
    }
}

