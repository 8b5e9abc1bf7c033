// Small non-GEMM kernels: stand-alone softmax cross-entropy (any class count) and input staging.
#include <atomic>
#include <math.h>

#include "../include/ttd_b200.h"
#include "common.cuh"

namespace ttd {

extern std::atomic<uint64_t> g_launch_count;

// One warp per example: logsumexp, in_top_k(…,1) and dlogits = scale*(softmax - onehot)/B.
// Reference: sparse_softmax_cross_entropy_with_logits + reduce_mean (:80-82), in_top_k (:96), accuracy (:125).
__global__ void __launch_bounds__(256) softmax_xent_kernel(const float* __restrict__ logits, long long ldl,
                                                           const long long* __restrict__ labels, int B, int C,
                                                           __nv_bfloat16* __restrict__ dl, long long ldd,
                                                           __nv_bfloat16* __restrict__ dlT, long long ldt, float scale,
                                                           float* metrics) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int row = blockIdx.x * 8 + warp;
  float loss = 0.0f, corr = 0.0f;
  if (row < B) {
    const float* z = logits + static_cast<long long>(row) * ldl;
    const long long label = labels[row];
    float mx = -INFINITY;
    for (int c = lane; c < C; c += 32) mx = fmaxf(mx, z[c]);
    mx = warp_max(mx);
    const float zl = z[label];
    float sum = 0.0f;
    int greater = 0;
    for (int c = lane; c < C; c += 32) {
      sum += __expf(z[c] - mx);
      greater += z[c] > zl ? 1 : 0;
    }
    sum = warp_sum(sum);
    greater = warp_sum(greater);
    const float inv = 1.0f / sum, gs = scale / static_cast<float>(B);
    for (int c = lane; c < C; c += 32) {
      const float d = (__expf(z[c] - mx) * inv - (c == label ? 1.0f : 0.0f)) * gs;
      if (dl) dl[static_cast<long long>(row) * ldd + c] = __float2bfloat16(d);
      if (dlT) dlT[static_cast<long long>(c) * ldt + row] = __float2bfloat16(d);
    }
    loss = (__logf(sum) + mx - zl) / static_cast<float>(B);
    corr = greater < 1 ? 1.0f / static_cast<float>(B) : 0.0f;
  }
  __shared__ float s_loss[8], s_corr[8];
  if (lane == 0) { s_loss[warp] = loss; s_corr[warp] = corr; }
  __syncthreads();
  if (threadIdx.x == 0 && metrics) {
    float l = 0.0f, c = 0.0f;
    for (int i = 0; i < 8; ++i) { l += s_loss[i]; c += s_corr[i]; }
    atomicAdd(&metrics[0], l);
    atomicAdd(&metrics[1], c);
  }
}

// x fp32 [B,F] -> bf16 [B,ldx] and bf16 transposed [F,ldt]; 32x32 smem-tiled so both sides coalesce.
// Reference: the feed of x-input (:191, :224-226).
// With a device-resident dataset (pool_batches > 1) the batch is chosen on the device from the worker's
// local step, so a captured graph walks through the pool without any host involvement; the labels of the
// chosen batch are copied next to it (the xent epilogue reads a fixed address).
__global__ void __launch_bounds__(256) prep_input_kernel(const float* __restrict__ x, int B, int F,
                                                         __nv_bfloat16* __restrict__ xb, long long ldx,
                                                         __nv_bfloat16* __restrict__ xT, long long ldt,
                                                         const long long* __restrict__ labels_pool,
                                                         long long* __restrict__ labels_out, int pool_batches,
                                                         const unsigned long long* __restrict__ step_ptr,
                                                         float* __restrict__ xf, int use_pdl) {
  __shared__ float tile[32][33];
  if (use_pdl) {  // programmatic dependent launch: started early, may not touch memory before this point
    asm volatile("griddepcontrol.wait;" ::: "memory");
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  }
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int b0 = blockIdx.y * 32, f0 = blockIdx.x * 32;
  if (pool_batches > 1 && step_ptr != nullptr) {
    const unsigned long long slot = *step_ptr % static_cast<unsigned long long>(pool_batches);
    x += slot * static_cast<unsigned long long>(B) * F;
    if (labels_pool != nullptr) labels_pool += slot * static_cast<unsigned long long>(B);
  }
  if (labels_pool != nullptr && labels_out != nullptr && blockIdx.x == 0) {
    const int b = b0 + threadIdx.x;
    if (threadIdx.x < 32 && b < B) labels_out[b] = labels_pool[b];
  }
#pragma unroll
  for (int rr = 0; rr < 4; ++rr) {
    const int b = b0 + warp + 8 * rr, f = f0 + lane;
    float v = 0.0f;
    if (b < B && f < F) {
      v = x[static_cast<long long>(b) * F + f];
      if (xb) xb[static_cast<long long>(b) * ldx + f] = __float2bfloat16(v);
      if (xf) xf[static_cast<long long>(b) * F + f] = v;  // fp32 staging (conv front-ends read fp32 NHWC)
    }
    tile[warp + 8 * rr][lane] = v;
  }
  __syncthreads();
  if (xT) {
#pragma unroll
    for (int cc = 0; cc < 4; ++cc) {
      const int f = f0 + warp + 8 * cc, b = b0 + lane;
      if (f < F && b < B) xT[static_cast<long long>(f) * ldt + b] = __float2bfloat16(tile[lane][warp + 8 * cc]);
    }
  }
}

}  // namespace ttd

using namespace ttd;

extern "C" int ttd_softmax_xent(const float* logits, int64_t ldl, const int64_t* labels, int32_t B, int32_t C,
                                void* dlogits, int64_t ldd, void* dlogitsT, int64_t ldt, float scale, float* metrics,
                                void* stream) {
  if (B < 1 || C < 1) {
    set_error("ttd_softmax_xent: empty problem %dx%d", B, C);
    return -1;
  }
  softmax_xent_kernel<<<(B + 7) / 8, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      logits, ldl, reinterpret_cast<const long long*>(labels), B, C, static_cast<__nv_bfloat16*>(dlogits), ldd,
      static_cast<__nv_bfloat16*>(dlogitsT), ldt, scale, metrics);
  TTD_CUDA(cudaGetLastError());
  g_launch_count.fetch_add(1);
  return 0;
}

extern "C" int ttd_prep_input(const float* x, int32_t B, int32_t F, void* xb, int64_t ldx, void* xT, int64_t ldt,
                              const int64_t* labels_pool, int64_t* labels_out, int32_t pool_batches,
                              const uint64_t* step_ptr, float* xf, void* stream) {
  return ttd_prep_input_ex(x, B, F, xb, ldx, xT, ldt, labels_pool, labels_out, pool_batches, step_ptr, xf, 0, stream);
}

extern "C" int ttd_prep_input_ex(const float* x, int32_t B, int32_t F, void* xb, int64_t ldx, void* xT, int64_t ldt,
                                 const int64_t* labels_pool, int64_t* labels_out, int32_t pool_batches,
                                 const uint64_t* step_ptr, float* xf, uint32_t flags, void* stream) {
  if (B < 1 || F < 1) {
    set_error("ttd_prep_input: empty problem %dx%d", B, F);
    return -1;
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((F + 31) / 32, (B + 31) / 32);
  cfg.blockDim = dim3(256);
  cfg.stream = static_cast<cudaStream_t>(stream);
  cudaLaunchAttribute attr[1];
  const int use_pdl = (flags & 1u) ? 1 : 0;
  if (use_pdl) {
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
  }
  TTD_CUDA(cudaLaunchKernelEx(&cfg, prep_input_kernel, x, B, F, static_cast<__nv_bfloat16*>(xb), static_cast<long long>(ldx),
                              static_cast<__nv_bfloat16*>(xT), static_cast<long long>(ldt),
                              reinterpret_cast<const long long*>(labels_pool), reinterpret_cast<long long*>(labels_out),
                              pool_batches, reinterpret_cast<const unsigned long long*>(step_ptr), xf, use_pdl));
  TTD_CUDA(cudaGetLastError());
  g_launch_count.fetch_add(1);
  return 0;
}
