// umma.cuh -- thin inline-PTX wrappers for the Blackwell (sm_100a) instruction forms the octave kernel is built on:
// TMA tiled loads (cp.async.bulk.tensor), mbarriers, tensor memory (tcgen05.alloc / ld) and the single-thread
// tensor-core instruction tcgen05.mma.kind::i8 (u8 x u8 -> s32, accumulators in TMEM).  Nothing here is arithmetic of
// the reference; it is the plumbing under fused_umma.cu.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace vslam {
namespace umma {

// ---- mbarrier -----------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
// make mbarrier.init visible to the async proxy (TMA completions, tcgen05.commit arrivals)
__device__ __forceinline__ void fence_barrier_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
    asm volatile("{\n\t.reg .b64 st;\n\tmbarrier.arrive.shared::cta.b64 st, [%0];\n\t}" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("{\n\t.reg .b64 st;\n\tmbarrier.arrive.expect_tx.shared::cta.b64 st, [%0], %1;\n\t}" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ bool mbar_try(uint32_t bar, uint32_t parity) {
    uint32_t ok;
    // the time hint lets the hardware park the warp instead of spinning through issue slots the working warps need
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(ok) : "r"(bar), "r"(parity), "r"(0x989680u) : "memory");
    return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    while (!mbar_try(bar, parity)) {
#ifdef VSLAM_MBAR_SLEEP
        __nanosleep(VSLAM_MBAR_SLEEP);
#endif
    }
}

// ---- proxies --------------------------------------------------------------------------------------------------------
// shared memory written with ordinary stores must be fenced before the async proxy (tcgen05.mma operand reads) sees it
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// ---- TMA ------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ void tma_prefetch_desc(const CUtensorMap *m) {
    asm volatile("prefetch.tensormap [%0];" ::"l"((uint64_t)m) : "memory");
}
// 3-D tiled load (x = bytes along a row, y = row, z = frame) into shared memory; completion on `bar` (complete_tx)
__device__ __forceinline__ void tma_load_3d(uint32_t dst, const CUtensorMap *m, uint32_t bar, int x, int y, int z) {
    asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
                 ::"r"(dst), "l"((uint64_t)m), "r"(bar), "r"(x), "r"(y), "r"(z) : "memory");
}

// ---- tensor memory --------------------------------------------------------------------------------------------------
// one warp allocates `cols` (power of two >= 32) columns; the base address lands in shared memory
__device__ __forceinline__ void tmem_alloc(uint32_t dst_smem, uint32_t cols) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(dst_smem), "r"(cols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t cols) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(cols) : "memory");
}
// warp w reads lanes 32 (w % 4) .. + 31: thread = lane, 16 consecutive 32-bit columns starting at taddr's column
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&v)[16]) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
                   "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
                 : "r"(taddr));
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// ---- tcgen05.mma ----------------------------------------------------------------------------------------------------
// shared-memory matrix descriptor (tcgen05 "version 1"): 14-bit fields in units of 16 bytes
//   lbo / sbo: K-major, no swizzle  -> core matrix = 8 rows x 16 bytes (rows 16 bytes apart); sbo = bytes between groups
//              of 8 rows, lbo = bytes between the two 16-byte K halves of one instruction (K = 32 bytes)
//              MN-major, 128B swizzle -> 128 contiguous bytes of MN per K index, K indices 128 bytes apart, sbo = bytes
//              between groups of 8 K indices (1024), lbo = bytes between blocks of 128 MN elements
constexpr uint32_t kLayoutNone = 0, kLayoutSw128 = 2;
__host__ __device__ constexpr uint64_t smem_desc(uint32_t addr, uint32_t lbo, uint32_t sbo, uint32_t layout) {
    return (uint64_t)((addr >> 4) & 0x3fffu) | (uint64_t)((lbo >> 4) & 0x3fffu) << 16 | (uint64_t)((sbo >> 4) & 0x3fffu) << 32 |
           1ull << 46 | (uint64_t)layout << 61;
}
// instruction descriptor of kind::i8: u8 x u8 -> s32, M = 128, N columns; operand majors 0 = K, 1 = MN
__host__ __device__ constexpr uint32_t idesc_u8(int n, int a_mn_major, int b_mn_major) {
    return 2u << 4 | (uint32_t)a_mn_major << 15 | (uint32_t)b_mn_major << 16 | (uint32_t)(n >> 3) << 17 | 8u << 24;
}
// D[tmem] (+)= A[smem] * B[smem]; issued by ONE thread
__device__ __forceinline__ void mma_i8_ss(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                 "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, {%5, %5, %5, %5}, p;\n\t}"
                 ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate), "r"(0u) : "memory");
}
// all tcgen05.mma issued so far by this thread arrive on `bar` when they have completed (implies fence::before_thread_sync)
__device__ __forceinline__ void mma_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}

}  // namespace umma
}  // namespace vslam
