// tile_pipe.cuh -- PTX helpers of the source-tile pipeline shared by the O(N^2)
// kernels (rmb_mobility.cu, rmb_source_target.cu): mbarrier, 1-D bulk async
// copy global -> shared (TMA engine, SASS UBLKCP) and 16-byte broadcast loads
// of one packed source record.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace rmb {

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count)
{
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes)
{
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity)
{
  uint32_t done;
  do {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(done)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
  } while (!done);
}
// 1-D bulk async copy global -> shared (TMA engine, SASS UBLKCP), completion
// counted in bytes on an mbarrier.  dst/src 16-byte aligned, bytes % 16 == 0.
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar)
{
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   smem_u32(dst)),
               "l"(src), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void fence_barrier_init()
{
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}

// 16-byte shared-memory loads of one record (all lanes read the same address:
// a broadcast, one wavefront each).
template <int W>
__device__ __forceinline__ void load_rec(const double* p, double (&w)[W])
{
#pragma unroll
  for (int k = 0; k < W / 2; k++) {
    const double2 v = reinterpret_cast<const double2*>(p)[k];
    w[2 * k] = v.x; w[2 * k + 1] = v.y;
  }
}
template <int W>
__device__ __forceinline__ void load_rec(const float* p, float (&w)[W])
{
#pragma unroll
  for (int k = 0; k < W / 4; k++) {
    const float4 v = reinterpret_cast<const float4*>(p)[k];
    w[4 * k] = v.x; w[4 * k + 1] = v.y; w[4 * k + 2] = v.z; w[4 * k + 3] = v.w;
  }
}

}  // namespace rmb
