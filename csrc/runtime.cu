// Library-wide plumbing: last-error string, launch counter, version stamp.
#include <atomic>
#include <stdarg.h>
#include <stdio.h>

#include "../include/ttd_b200.h"
#include "common.cuh"

namespace ttd {

static thread_local char g_err[512] = "";
std::atomic<uint64_t> g_launch_count{0};

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int check_cuda(cudaError_t e, const char* what) {
  if (e == cudaSuccess) return 0;
  set_error("CUDA error %d (%s) at %s", static_cast<int>(e), cudaGetErrorString(e), what);
  return static_cast<int>(e);
}

}  // namespace ttd

extern "C" const char* ttd_last_error(void) { return ttd::g_err; }
extern "C" const char* ttd_version(void) { return "ttd_b200 sm_100a " __DATE__ " " __TIME__; }
extern "C" uint64_t ttd_launch_count(void) { return ttd::g_launch_count.load(); }

// ---------------------------------------------------------------------------------------------------
// Step feeder: the host side of one training step in ONE native call (the reference's equivalent is the C++
// session run loop behind mon_sess.run, distribute_training.py:225). Host batches land in a two-slot device
// pool on a copy stream; the step (a CUDA graph or caller-launched kernels) runs on the compute stream and
// picks slot local_step % 2 on the device, so the copy of batch s+1 overlaps the compute of batch s.
// ---------------------------------------------------------------------------------------------------
namespace ttd {
struct Feeder {
  cudaStream_t compute = nullptr, copy = nullptr;
  void* x_slot[2] = {nullptr, nullptr};
  void* y_slot[2] = {nullptr, nullptr};
  size_t x_bytes = 0, y_bytes = 0;
  cudaEvent_t ready[2] = {nullptr, nullptr};  // batch in slot i has landed        (recorded on copy)
  cudaEvent_t freed[2] = {nullptr, nullptr};  // last step that read slot i is done (recorded on compute)
  int pending = -1;                           // slot fed for the next step, -1: none
  // optional conversion of a freshly copied slot (fp32 rows -> bf16 rows + transposed bf16 rows) on the copy
  // stream, so that the step's first kernel finds its operands ready instead of converting on the critical path
  int prep_B = 0, prep_F = 0;
  __nv_bfloat16* prep_xb = nullptr;
  __nv_bfloat16* prep_xT = nullptr;
  long long prep_ldx = 0, prep_ldt = 0;
};
}  // namespace ttd

#define TTD_RT(call)                                                                      \
  do {                                                                                    \
    cudaError_t e_ = (call);                                                              \
    if (e_ != cudaSuccess) {                                                              \
      ttd::set_error("%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
      return -1;                                                                          \
    }                                                                                     \
  } while (0)

extern "C" void* ttd_feeder_create(void* compute_stream, void* copy_stream, void* x_slot0, void* x_slot1,
                                   void* y_slot0, void* y_slot1, size_t x_bytes, size_t y_bytes) {
  auto* f = new ttd::Feeder();
  f->compute = static_cast<cudaStream_t>(compute_stream);
  f->copy = static_cast<cudaStream_t>(copy_stream);
  f->x_slot[0] = x_slot0; f->x_slot[1] = x_slot1;
  f->y_slot[0] = y_slot0; f->y_slot[1] = y_slot1;
  f->x_bytes = x_bytes; f->y_bytes = y_bytes;
  for (int i = 0; i < 2; ++i) {
    if (cudaEventCreateWithFlags(&f->ready[i], cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&f->freed[i], cudaEventDisableTiming) != cudaSuccess) {
      ttd::set_error("ttd_feeder_create: cudaEventCreate failed");
      delete f;
      return nullptr;
    }
  }
  return f;
}

extern "C" void ttd_feeder_destroy(void* h) {
  auto* f = static_cast<ttd::Feeder*>(h);
  if (!f) return;
  for (int i = 0; i < 2; ++i) {
    if (f->ready[i]) cudaEventDestroy(f->ready[i]);
    if (f->freed[i]) cudaEventDestroy(f->freed[i]);
  }
  delete f;
}

// Slot s is converted right after its copy: xb + s*B*ldx receives the bf16 rows [B, ldx], xT + s*(F+1)*ldt the
// transposed rows [F, ldt] (row F of every slot is the caller's ones row). NULL xb switches the conversion off.
extern "C" int ttd_feeder_set_prep(void* h, int32_t B, int32_t F, void* xb, int64_t ldx, void* xT, int64_t ldt) {
  auto* f = static_cast<ttd::Feeder*>(h);
  if (f == nullptr) return -1;
  f->prep_B = B; f->prep_F = F;
  f->prep_xb = static_cast<__nv_bfloat16*>(xb); f->prep_ldx = ldx;
  f->prep_xT = static_cast<__nv_bfloat16*>(xT); f->prep_ldt = ldt;
  return 0;
}

// H2D of the batch for step `step` (slot step & 1) on the copy stream, behind the last reader of that slot.
extern "C" int ttd_feeder_feed(void* h, const void* host_x, const void* host_y, uint64_t step) {
  auto* f = static_cast<ttd::Feeder*>(h);
  const int s = static_cast<int>(step & 1);
  TTD_RT(cudaStreamWaitEvent(f->copy, f->freed[s], 0));
  TTD_RT(cudaMemcpyAsync(f->x_slot[s], host_x, f->x_bytes, cudaMemcpyHostToDevice, f->copy));
  TTD_RT(cudaMemcpyAsync(f->y_slot[s], host_y, f->y_bytes, cudaMemcpyHostToDevice, f->copy));
  if (f->prep_xb != nullptr) {
    const int B = f->prep_B, F = f->prep_F;
    if (ttd_prep_input_ex(static_cast<const float*>(f->x_slot[s]), B, F, f->prep_xb + static_cast<long long>(s) * B * f->prep_ldx,
                          f->prep_ldx, f->prep_xT + static_cast<long long>(s) * (F + 1) * f->prep_ldt, f->prep_ldt, nullptr,
                          nullptr, 1, nullptr, nullptr, 0, f->copy) != 0)
      return -1;
  }
  TTD_RT(cudaEventRecord(f->ready[s], f->copy));
  f->pending = s;
  return 0;
}

// The compute stream waits for the fed batch; returns the slot (or -1 when nothing was fed).
extern "C" int ttd_feeder_before_step(void* h) {
  auto* f = static_cast<ttd::Feeder*>(h);
  if (f->pending >= 0) TTD_RT(cudaStreamWaitEvent(f->compute, f->ready[f->pending], 0));
  return f->pending;
}

// After the step's kernels are enqueued: its slot may be overwritten once they finish.
extern "C" int ttd_feeder_after_step(void* h) {
  auto* f = static_cast<ttd::Feeder*>(h);
  if (f->pending >= 0) TTD_RT(cudaEventRecord(f->freed[f->pending], f->compute));
  f->pending = -1;
  return 0;
}

// One whole step: feed (when host_x != NULL), graph launch, optional 64-byte state read-back into pinned memory.
extern "C" int ttd_feeder_step(void* h, void* graph_exec, const void* host_x, const void* host_y, uint64_t step,
                               const void* state_dev, void* state_host) {
  auto* f = static_cast<ttd::Feeder*>(h);
  if (host_x != nullptr && ttd_feeder_feed(h, host_x, host_y, step) != 0) return -1;
  if (ttd_feeder_before_step(h) < -1) return -1;
  TTD_RT(cudaGraphLaunch(static_cast<cudaGraphExec_t>(graph_exec), f->compute));
  if (ttd_feeder_after_step(h) != 0) return -1;
  if (state_host != nullptr)
    TTD_RT(cudaMemcpyAsync(state_host, state_dev, sizeof(TtdStepState), cudaMemcpyDeviceToHost, f->compute));
  return 0;
}
