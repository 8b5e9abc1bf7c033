// Diagnostics: cross-GPU flag ping-pong, used to measure what a device-side signal over NVLink costs with
// different memory-ordering recipes (the numbers steer the design of the replica barrier in param_store.cu).
#include <atomic>

#include "../include/ttd_b200.h"
#include "common.cuh"

namespace ttd {
extern std::atomic<uint64_t> g_launch_count;

__global__ void pingpong_kernel(uint32_t* mine, uint32_t* theirs, int rank, int iters, int variant,
                                unsigned long long* out_ns) {
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  const uint64_t t0 = globaltimer_ns();
  for (uint32_t i = 1; i <= static_cast<uint32_t>(iters); ++i) {
    for (int phase = 0; phase < 2; ++phase) {
      const bool send = (phase == 0) == (rank == 0);
      if (send) {
        switch (variant) {
          case 0: asm volatile("st.volatile.global.u32 [%0], %1;" ::"l"(theirs), "r"(i) : "memory"); break;
          case 1: __threadfence_system();
                  asm volatile("st.relaxed.sys.global.u32 [%0], %1;" ::"l"(theirs), "r"(i) : "memory"); break;
          case 2: asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(theirs), "r"(i) : "memory"); break;
          case 3: atomicExch_system(theirs, i); break;
          default: asm volatile("red.release.sys.global.max.u32 [%0], %1;" ::"l"(theirs), "r"(i) : "memory"); break;
        }
      } else {
        uint32_t v = 0;
        uint64_t spins = 0;
        do {
          if (variant == 2) asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(mine) : "memory");
          else if (variant == 0) asm volatile("ld.volatile.global.u32 %0, [%1];" : "=r"(v) : "l"(mine) : "memory");
          else asm volatile("ld.relaxed.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(mine) : "memory");
          if (++spins > (1ull << 28)) __trap();
        } while (v < i);
        if (variant == 1) __threadfence_system();
      }
    }
  }
  *out_ns = globaltimer_ns() - t0;
}
}  // namespace ttd

extern "C" int ttd_diag_pingpong(uint32_t* mine, uint32_t* theirs, int rank, int iters, int variant,
                                 unsigned long long* out_ns, void* stream) {
  ttd::pingpong_kernel<<<1, 32, 0, static_cast<cudaStream_t>(stream)>>>(mine, theirs, rank, iters, variant, out_ns);
  TTD_CUDA(cudaGetLastError());
  ttd::g_launch_count.fetch_add(1);
  return 0;
}
