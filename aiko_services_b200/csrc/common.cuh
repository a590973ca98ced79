// Shared device/host helpers for libaiko_b200 (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/aiko_b200.h"

namespace aiko {

constexpr int kThreads = 256;

// ---- status plumbing: the C-ABI never throws and never aborts -------------
void set_last_error(const char* fmt, ...);
int cuda_status(cudaError_t e, const char* what);  // 0 or AIKO_B200_ERR_CUDA
int sm_count();
int device_max_smem_optin();

#define AIKO_CHECK_ARG(cond, ...)            \
  do {                                       \
    if (!(cond)) {                           \
      ::aiko::set_last_error(__VA_ARGS__);   \
      return AIKO_B200_ERR_INVALID_ARGUMENT; \
    }                                        \
  } while (0)

#define AIKO_CUDA(call)                                  \
  do {                                                   \
    int _rc = ::aiko::cuda_status((call), #call);        \
    if (_rc != 0) return _rc;                            \
  } while (0)

// ---- PTX wrappers ---------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void fence_mbar_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void fence_proxy_async_smem() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity, uint32_t suspend_ns) {
  uint32_t ok;
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t"
      "}"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity), "r"(suspend_ns)
      : "memory");
  return ok != 0;
}
// Blocks until the phase with the given parity completes.  The hardware suspends the thread
// for up to `suspend_ns` per probe and failed probes back off with nanosleep, so a waiting
// warp does not compete with working warps for issue slots.
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity, uint32_t suspend_ns = 1000,
                                          uint32_t backoff_ns = 0) {
  while (!mbar_try_wait(bar, parity, suspend_ns)) {
    if (backoff_ns) __nanosleep(backoff_ns);
  }
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
// 1-D bulk async copy global -> shared (TMA engine), completion on an mbarrier.
// dst, src 16-byte aligned; bytes a multiple of 16.
__device__ __forceinline__ void bulk_g2s(void* dst_smem, const void* src_gmem, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   smem_u32(dst_smem)),
               "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
// Same copy for data that is read exactly once: the L2 lines it brings in are the first to go,
// so they do not push out what later kernels will want (freshly written output, small tables).
__device__ __forceinline__ uint64_t l2_policy_evict_first() {
  uint64_t policy;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(policy));
  return policy;
}
__device__ __forceinline__ void bulk_g2s_hint(void* dst_smem, const void* src_gmem, uint32_t bytes, uint64_t* bar,
                                              uint64_t policy) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;" ::"r"(
          smem_u32(dst_smem)),
      "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar)), "l"(policy)
      : "memory");
}
// 1-D bulk async copy shared -> global.
__device__ __forceinline__ void bulk_s2g(void* dst_gmem, const void* src_smem, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst_gmem),
               "r"(smem_u32(src_smem)), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void bulk_wait_read() {
  asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}

// streaming 128-bit global accesses (frames are read once / written once)
__device__ __forceinline__ uint4 ldg_stream(const uint4* p) {
  uint4 r;
#ifdef AIKO_PLAIN_LDST
  r = *p;
#else
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
#endif
  return r;
}
__device__ __forceinline__ void stg_stream(uint4* p, const uint4& v) {
#ifdef AIKO_PLAIN_LDST
  *p = v;
#else
  asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z),
               "r"(v.w)
               : "memory");
#endif
}

// Copy one finished row from shared memory to global memory.  Vector path when
// the destination is 16-byte aligned (all configured widths), byte path else.
__device__ __forceinline__ void store_row(uint8_t* __restrict__ dst, const uint8_t* __restrict__ srow, int nbytes) {
  if ((reinterpret_cast<uintptr_t>(dst) & 15) == 0) {
    const int nvec = nbytes >> 4;
    const uint4* s4 = reinterpret_cast<const uint4*>(srow);
    uint4* d4 = reinterpret_cast<uint4*>(dst);
    for (int i = threadIdx.x; i < nvec; i += blockDim.x) stg_stream(d4 + i, s4[i]);
    for (int i = (nvec << 4) + threadIdx.x; i < nbytes; i += blockDim.x) dst[i] = srow[i];
  } else {
    for (int i = threadIdx.x; i < nbytes; i += blockDim.x) dst[i] = srow[i];
  }
}

}  // namespace aiko
