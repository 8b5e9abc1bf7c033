// ttd_apply_update_static — the gradient-push update for callers that are not allowed to allocate.
//
// Who this is for. Some hosts of this library own every byte of device memory themselves and forbid
// anything underneath them from asking the driver for more: a serving process that handed the whole
// HBM to its own arena at start-up, a plug-in running inside another framework's memory pool, a
// control loop that cannot tolerate the device-wide synchronisation a cudaMalloc/cudaFree implies.
// The regular entry point (ttd_apply_update) does not allocate either, but it makes its CALLER
// allocate: the completion counter the kernel needs lives in the caller's TtdStepState block and the
// variable table is limited to what fits in kernel parameters. Here both live in a fixed array inside
// the library image, so the caller brings only its inputs and outputs.
//
// The scratch. `g_static_scratch` below is 8 KiB per GPU, resident from the moment the library is
// loaded on that GPU until the process exits:
//     bytes    0..255   completion counter of the launch in flight (must be zero when a launch starts)
//     bytes  256..6399  variable table, at most kStaticMaxVars (64) groups of 96 bytes
// The limit is therefore 64 variable groups per call; a call with more is refused with an error text,
// nothing is launched.
//
// Rules that follow from one shared array:
//   * Every call on a GPU uses the same 8 KiB. The array keeps whatever the previous call left there,
//     so each call first clears the counter block on its stream and overwrites the table slots it uses.
//   * Two calls whose kernels can be in flight on the same GPU at the same time (different streams, or
//     different host threads) race on the counter and the table; callers must order them.
//   * The first call on each GPU performs the one-time set-up (it asks the runtime for the device
//     address of the array with cudaGetSymbolAddress while that GPU is current). Make that first call
//     before starting any stream capture.
//   * The call must run outside stream capture: the variable table is copied from the caller's host array on
//     every call, and a captured copy would re-read that array at each replay. Captured streams are refused.
//   * No cudaMalloc, cudaFree, or other allocating routine is called — by this file or on its behalf.
#include <atomic>
#include <math.h>
#include <mutex>
#include <string.h>

#include "../include/ttd_b200.h"
#include "common.cuh"
#include "step_state.cuh"

namespace ttd {

extern std::atomic<uint64_t> g_launch_count;

namespace {

constexpr int kStaticMaxVars = TTD_STATIC_MAX_VARS;
constexpr size_t kCounterBytes = 256;
constexpr size_t kTableOffset = 256;
constexpr size_t kScratchBytes = 8192;
static_assert(kTableOffset + kStaticMaxVars * sizeof(TtdVarDesc) <= kScratchBytes, "table must fit the scratch");

// In host code this name is only a handle for cudaGetSymbolAddress; it is NOT a device address.
static __device__ __align__(256) unsigned char g_static_scratch[kScratchBytes];

constexpr int kMaxDevices = 64;
std::mutex g_setup_mutex;
unsigned char* g_scratch_ptr[kMaxDevices] = {nullptr};

// One-time set-up per device (that device must be current): resolve the device address of the array.
int scratch_for_current_device(unsigned char** out) {
  int dev = -1;
  TTD_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kMaxDevices) {
    set_error("ttd_apply_update_static: device ordinal %d is outside the supported range [0,%d)", dev, kMaxDevices);
    return -1;
  }
  std::lock_guard<std::mutex> lock(g_setup_mutex);
  if (g_scratch_ptr[dev] == nullptr) {
    unsigned char* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_static_scratch);
    if (e != cudaSuccess || p == nullptr) {
      set_error("ttd_apply_update_static: cudaGetSymbolAddress failed on device %d: %s", dev, cudaGetErrorString(e));
      return e != cudaSuccess ? static_cast<int>(e) : -1;
    }
    g_scratch_ptr[dev] = p;
  }
  *out = g_scratch_ptr[dev];
  return 0;
}

// Same arithmetic as param_kernel<kUpdate> for one gradient source on one GPU; the variable table and
// the completion counter are read from `scratch` (the device address of g_static_scratch).
__global__ void __launch_bounds__(256) update_static_kernel(unsigned char* __restrict__ scratch, int nvars,
                                                            float* __restrict__ w, float* __restrict__ ema,
                                                            const float* __restrict__ grad, long long lo, long long hi,
                                                            TtdHyper h, TtdStepState* st, int bump_global_step) {
  const TtdVarDesc* table = reinterpret_cast<const TtdVarDesc*>(scratch + kTableOffset);
  unsigned int* done = reinterpret_cast<unsigned int*>(scratch);
  __shared__ float s_lr, s_ema_d;
  __shared__ unsigned int s_last;
  if (threadIdx.x == 0) {
    const long long gs = st->global_step;
    double e = static_cast<double>(gs) / static_cast<double>(h.decay_steps);
    if (h.staircase) e = floor(e);
    s_lr = static_cast<float>(static_cast<double>(h.lr0) * pow(static_cast<double>(h.decay_rate), e));
    const double n = static_cast<double>(gs + 1);
    double d = h.ema_decay;
    if (h.ema_warmup) d = fmin(d, (1.0 + n) / (10.0 + n));
    s_ema_d = static_cast<float>(d);
  }
  __syncthreads();
  const float lr = s_lr, ema_d = s_ema_d;

  for (int vi = 0; vi < nvars; ++vi) {
    const TtdVarDesc V = table[vi];
    const long long v_lo = V.offset, v_hi = V.offset + static_cast<long long>(V.rows + 1) * V.cols;
    const long long a = v_lo > lo ? v_lo : lo, b = v_hi < hi ? v_hi : hi;
    for (long long idx = a + blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; idx < b;
         idx += static_cast<long long>(gridDim.x) * blockDim.x) {
      const long long rel = idx - v_lo;
      const int r = static_cast<int>(rel / V.cols), c = static_cast<int>(rel % V.cols);
      float g = grad[idx] * h.grad_scale;
      const float wv = w[idx];
      if (V.regularize && r < V.rows) {
        if (h.reg_kind == 2) g += h.reg_scale * wv;
        else if (h.reg_kind == 1) g += h.reg_scale * (wv > 0.0f ? 1.0f : (wv < 0.0f ? -1.0f : 0.0f));
      }
      const float val = wv - lr * g;
      w[idx] = val;
      if (ema != nullptr && h.ema_decay > 0.0f) {
        const float s = ema[idx];
        ema[idx] = s - (1.0f - ema_d) * (s - val);
      }
      if (r == V.rows && V.bias_c != nullptr) V.bias_c[c] = val;
      if (r < V.rows && V.w_f32 != nullptr) V.w_f32[static_cast<long long>(r) * V.ld_f32 + c] = val;
      if (r < V.rows && V.wt_f32 != nullptr) V.wt_f32[static_cast<long long>(c) * V.ld_tf32 + r] = val;
      if (r < V.rows) {
        const __nv_bfloat16 bv = __float2bfloat16(val);
        if (V.w_rm != nullptr) static_cast<__nv_bfloat16*>(V.w_rm)[static_cast<long long>(r) * V.ld_rm + c] = bv;
        if (V.w_t != nullptr) static_cast<__nv_bfloat16*>(V.w_t)[static_cast<long long>(c) * V.ld_t + r] = bv;
      }
    }
  }

  // last CTA out advances the step counters and publishes the metrics, exactly like the regular kernel
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) s_last = (atomicAdd(done, 1u) == gridDim.x - 1) ? 1u : 0u;
  __syncthreads();
  if (s_last && threadIdx.x == 0) {
    *done = 0;
    finalize_step(st, h, lr, bump_global_step != 0);  // counters, metrics, pinned host mirror: as the regular kernel
  }
}

}  // namespace
}  // namespace ttd

using namespace ttd;

extern "C" int ttd_apply_update_static(float* w, float* ema, const float* grad, const TtdVarDesc* vars_host,
                                       int nvars, int64_t lo, int64_t hi, const TtdHyper* hyper_host,
                                       TtdStepState* state, int32_t bump_global_step, void* stream) {
  if (nvars < 1 || nvars > kStaticMaxVars) {
    set_error("ttd_apply_update_static: %d variable groups requested but the built-in scratch holds at most %d "
              "(8 KiB per GPU); split the call or use ttd_apply_update",
              nvars, kStaticMaxVars);
    return -1;
  }
  if (w == nullptr || grad == nullptr || vars_host == nullptr || hyper_host == nullptr || state == nullptr) {
    set_error("ttd_apply_update_static: w, grad, vars, hyper and state must be non-NULL");
    return -1;
  }
  if (hi < 0)  // same convention as ttd_apply_update: a negative hi means "up to the end of the last group"
    hi = vars_host[nvars - 1].offset + static_cast<int64_t>(vars_host[nvars - 1].rows + 1) * vars_host[nvars - 1].cols;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  {
    // The variable table travels by cudaMemcpyAsync from the caller's (pageable) array. Inside a stream capture
    // that copy would become a graph node that re-reads the host array on every replay, long after the caller
    // has released it: refuse instead.
    cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
    if (cudaStreamIsCapturing(s, &cap) == cudaSuccess && cap != cudaStreamCaptureStatusNone) {
      set_error("ttd_apply_update_static: the stream is being captured; this entry point copies its variable table "
                "from host memory on every call and must run outside stream capture");
      return -1;
    }
  }
  unsigned char* scratch = nullptr;
  int rc = scratch_for_current_device(&scratch);  // first call on this GPU resolves the address
  if (rc != 0) return rc;
  // The array still holds the previous call's bytes: zero the counter block, overwrite the table slots in use.
  TTD_CUDA(cudaMemsetAsync(scratch, 0, kCounterBytes, s));
  TTD_CUDA(cudaMemcpyAsync(scratch + kTableOffset, vars_host, sizeof(TtdVarDesc) * static_cast<size_t>(nvars),
                           cudaMemcpyHostToDevice, s));
  long long elems = 0;
  for (int i = 0; i < nvars; ++i) elems += static_cast<long long>(vars_host[i].rows + 1) * vars_host[i].cols;
  long long blocks = (elems + 256 * 4 - 1) / (256 * 4);
  if (blocks < 1) blocks = 1;
  if (blocks > 148 * 8) blocks = 148 * 8;
  update_static_kernel<<<static_cast<unsigned>(blocks), 256, 0, s>>>(scratch, nvars, w, ema, grad, lo, hi, *hyper_host,
                                                                    state, bump_global_step);
  TTD_CUDA(cudaGetLastError());
  g_launch_count.fetch_add(1);
  return 0;
}
