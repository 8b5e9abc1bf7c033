// End-of-step bookkeeping shared by every kernel that can be the last one of a training step
// (the fused update kernels): counters, metrics hand-over and the pinned host mirror.
#pragma once
#include "../include/ttd_b200.h"
#include "common.cuh"

namespace ttd {

// Called by ONE thread once all other work of the step is globally visible (after a __threadfence).
// Reference: AssignAdd(global_step) on the ps (:152), the fetches of mon_sess.run (:225-226).
__device__ __forceinline__ void finalize_step(TtdStepState* st, const TtdHyper& h, float lr, bool bump_global_step) {
  if (bump_global_step) st->global_step += 1;
  st->local_step += 1;
  st->lr_last = lr;
  st->metrics_out[0] = st->metrics_acc[0];
  st->metrics_out[1] = st->metrics_acc[1];
  st->metrics_acc[0] = 0.0f;
  st->metrics_acc[1] = 0.0f;
  __threadfence();
  if (h.host_mirror != nullptr && h.mirror_slots > 0) {
    // the step's result goes straight into pinned host memory (slot = step index mod ring): no D2H copy
    // operation sits between this step and the next one on the stream
    const uint4* src = reinterpret_cast<const uint4*>(st);
    uint4* dst = reinterpret_cast<uint4*>(static_cast<TtdStepState*>(h.host_mirror) +
                                          (st->local_step - 1) % static_cast<uint64_t>(h.mirror_slots));
    // No fence between the four 16-byte stores (a system fence costs a PCIe round trip on the step's critical
    // path): the copy carries the low word of local_step as a tag in every chunk (in place of fields the host
    // does not read: metrics_acc[0], done_counter, reserved[5]), so a host that POLLS the slot accepts it
    // only once all four chunks of this step have landed; a host that synchronises the stream first sees
    // them all anyway.
    const uint32_t tag = static_cast<uint32_t>(st->local_step);
    uint4 c1 = src[1], c2 = src[2], c3 = src[3];
    c1.x = tag;
    c2.x = tag;
    c3.w = tag;
    dst[1] = c1;
    dst[2] = c2;
    dst[3] = c3;
    dst[0] = src[0];
  }
}

}  // namespace ttd
