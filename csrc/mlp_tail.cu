// Fused "tail" of a small MLP: every layer after the first — forward, loss, and the whole backward — in ONE
// kernel launch on a thread-block cluster.
//
// Why. For the reference model (784-200-100-50-25-10, batch 128; distribute_training.py:93) the layers behind
// the first hold 15 % of the parameters and 12 % of the FLOPs, but as one kernel per GEMM they are eight
// *dependent* launches of ~6 us each — more than half of the training step. Their operands (<= 100 KB) fit in
// L1/L2, so the work is latency-, not throughput-bound, and a 128x100x200 GEMM fills less than one tcgen05
// tile. This kernel runs the chain F2..FL, softmax-xent, DL..D2 and all weight gradients with fp32 CUDA-core
// register tiles spread over the CTAs of one cluster; a phase boundary is a cluster barrier (~0.3 us) instead
// of a kernel boundary. The first layer (85 % of the parameters) stays on the tcgen05 GEMM.
//
// Every product has the shape  Out[R][C] = sum_k X[r][k] * Y[k][c]  with X k-contiguous and Y c-contiguous:
//     forward   X = a_{l-1} [B][in]        Y = W_l   [in][out]   (TF layout, fp32 compute copy)
//     dgrad     X = dz_l    [B][out]       Y = W_l^T [out][in]   (transposed fp32 compute copy)
//     wgrad     X = a_{l-1}^T [in+1][B]    Y = dz_l  [B][out]    (row `in` of a^T is all ones -> bias gradient)
// All leading dimensions are multiples of 4 floats and all K extents are zero-padded to a multiple of 4.
//
// Semantics reproduced: dense + ELU/ReLU + dropout (reference :54-58), output layer (:61), mean sparse softmax
// cross-entropy (:80-82), in_top_k accuracy (:96, :125) and the gradients opt.minimize builds (:152).
#include <atomic>
#include <math.h>

#include "../include/ttd_b200.h"
#include "common.cuh"

namespace ttd {

extern std::atomic<uint64_t> g_launch_count;

constexpr int kTailThreads = 512;

struct TailParams {
  TtdTailDesc d;
  uint32_t seed_lo, seed_hi;
  int use_pdl;
};

__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}

template <typename T>
__device__ __forceinline__ float4 load_x4(const T* p);
template <>
__device__ __forceinline__ float4 load_x4<float>(const float* p) {
  return *reinterpret_cast<const float4*>(p);
}
template <>
__device__ __forceinline__ float4 load_x4<__nv_bfloat16>(const __nv_bfloat16* p) {
  const uint2 u = *reinterpret_cast<const uint2*>(p);
  const __nv_bfloat162 a = *reinterpret_cast<const __nv_bfloat162*>(&u.x), b = *reinterpret_cast<const __nv_bfloat162*>(&u.y);
  return make_float4(__low2float(a), __high2float(a), __low2float(b), __high2float(b));
}

// One phase: Out[R][C] = X[R][K] * Y[K][C], 2x4 register tiles dealt round-robin to every thread of the cluster.
// K4 = K rounded up to 4 (padding is zero on both operands). epi(r, c, v) receives each finished element.
template <typename XT, typename Epi>
__device__ __forceinline__ void tail_gemm(int R, int C, int K4, const XT* __restrict__ X, long long ldx,
                                          const float* __restrict__ Y, long long ldy, int tid, int nthreads, Epi epi) {
  const int tiles_c = (C + 3) >> 2, tiles_r = (R + 1) >> 1;
  for (int t = tid; t < tiles_r * tiles_c; t += nthreads) {
    const int r0 = (t / tiles_c) * 2, c0 = (t % tiles_c) * 4;
    const XT* x0 = X + static_cast<long long>(r0) * ldx;
    const XT* x1 = X + static_cast<long long>(r0 + 1 < R ? r0 + 1 : r0) * ldx;
    const float* y = Y + c0;
    float acc[2][4] = {{0.f, 0.f, 0.f, 0.f}, {0.f, 0.f, 0.f, 0.f}};
#pragma unroll 2
    for (int k = 0; k < K4; k += 4) {
      const float4 a0 = load_x4<XT>(x0 + k), a1 = load_x4<XT>(x1 + k);
      const float4 b0 = *reinterpret_cast<const float4*>(y + static_cast<long long>(k) * ldy);
      const float4 b1 = *reinterpret_cast<const float4*>(y + static_cast<long long>(k + 1) * ldy);
      const float4 b2 = *reinterpret_cast<const float4*>(y + static_cast<long long>(k + 2) * ldy);
      const float4 b3 = *reinterpret_cast<const float4*>(y + static_cast<long long>(k + 3) * ldy);
      const float av0[4] = {a0.x, a0.y, a0.z, a0.w}, av1[4] = {a1.x, a1.y, a1.z, a1.w};
      const float4 bs[4] = {b0, b1, b2, b3};
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        acc[0][0] = fmaf(av0[j], bs[j].x, acc[0][0]); acc[0][1] = fmaf(av0[j], bs[j].y, acc[0][1]);
        acc[0][2] = fmaf(av0[j], bs[j].z, acc[0][2]); acc[0][3] = fmaf(av0[j], bs[j].w, acc[0][3]);
        acc[1][0] = fmaf(av1[j], bs[j].x, acc[1][0]); acc[1][1] = fmaf(av1[j], bs[j].y, acc[1][1]);
        acc[1][2] = fmaf(av1[j], bs[j].z, acc[1][2]); acc[1][3] = fmaf(av1[j], bs[j].w, acc[1][3]);
      }
    }
#pragma unroll
    for (int i = 0; i < 2; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j)
        if (r0 + i < R && c0 + j < C) epi(r0 + i, c0 + j, acc[i][j]);
  }
}

__device__ __forceinline__ bool tail_kept(float rate, uint32_t thresh, int r, int c, unsigned long long step,
                                          uint32_t layer_id, uint32_t s0, uint32_t s1) {
  if (rate <= 0.0f) return true;
  return dropout_kept(static_cast<uint32_t>(r), static_cast<uint32_t>(c), step, layer_id, s0, s1, thresh);
}

__global__ void __launch_bounds__(kTailThreads, 1) mlp_tail_kernel(const __grid_constant__ TailParams P) {
  const TtdTailDesc& d = P.d;
  uint32_t crank = 0, csize = 1;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(crank));
  asm volatile("mov.u32 %0, %%cluster_nctarank;" : "=r"(csize));
  const int tid = static_cast<int>(crank) * kTailThreads + threadIdx.x;
  const int nthreads = static_cast<int>(csize) * kTailThreads;
  if (P.use_pdl) {
    asm volatile("griddepcontrol.wait;" ::: "memory");
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  }
  const unsigned long long step = d.step_ptr != nullptr ? *d.step_ptr : 0ull;
  const int B = d.B, L = d.n_layers;
  const uint32_t s0 = P.seed_lo, s1 = P.seed_hi;

  // ------------------------------------------------ forward ------------------------------------------------
  for (int l = 0; l < L; ++l) {
    const TtdTailLayer& Ly = d.layer[l];
    const bool last = l == L - 1;
    const float keep = 1.0f - Ly.drop_rate, inv_keep = 1.0f / keep;
    const uint32_t thresh = dropout_thresh16(Ly.drop_rate);
    float* out = Ly.act;            // [B][ld_act]: post-dropout activation, or the logits for the last layer
    float* outT = Ly.actT;          // [out+1][ld_actT] (hidden layers only)
    auto epi = [&](int r, int c, float v) {
      const float z = v + Ly.bias[c];
      if (last) {
        out[static_cast<long long>(r) * Ly.ld_act + c] = z;
        return;
      }
      const float a = act_apply(z, Ly.act_kind);
      const bool kept = tail_kept(Ly.drop_rate, thresh, r, c, step, Ly.layer_id, s0, s1);
      const float ad = kept ? a * inv_keep : 0.0f;
      out[static_cast<long long>(r) * Ly.ld_act + c] = ad;
      outT[static_cast<long long>(c) * Ly.ld_actT + r] = ad;
    };
    const int K4 = (Ly.in + 3) & ~3;
    if (l == 0)
      tail_gemm<__nv_bfloat16>(B, Ly.out, K4, static_cast<const __nv_bfloat16*>(d.a_in), d.ld_a_in, Ly.w, Ly.ld_w, tid,
                               nthreads, epi);
    else
      tail_gemm<float>(B, Ly.out, K4, d.layer[l - 1].act, d.layer[l - 1].ld_act, Ly.w, Ly.ld_w, tid, nthreads, epi);
    cluster_sync_all();
  }

  // ------------------------------- loss, accuracy, dlogits (one thread per example) -------------------------
  {
    const TtdTailLayer& Ly = d.layer[L - 1];
    const int C = Ly.out;
    float loss = 0.0f, corr = 0.0f;
    if (tid < B) {
      const float* z = Ly.act + static_cast<long long>(tid) * Ly.ld_act;
      float* dz = Ly.dz + static_cast<long long>(tid) * Ly.ld_dz;
      const long long label = d.labels[tid];
      float mx = -INFINITY;
      for (int c = 0; c < C; ++c) mx = fmaxf(mx, z[c]);
      const float zl = z[label];
      float sum = 0.0f;
      int greater = 0;
      for (int c = 0; c < C; ++c) {
        sum += __expf(z[c] - mx);
        greater += z[c] > zl ? 1 : 0;
      }
      const float inv = 1.0f / sum, gs = 1.0f / static_cast<float>(B);
      for (int c = 0; c < C; ++c) dz[c] = (__expf(z[c] - mx) * inv - (c == label ? 1.0f : 0.0f)) * gs;
      loss = (__logf(sum) + mx - zl) * gs;
      corr = greater < 1 ? gs : 0.0f;
    }
    if (tid < ((B + 31) & ~31)) {  // whole warps
      loss = warp_sum(loss);
      corr = warp_sum(corr);
      if ((threadIdx.x & 31) == 0 && d.metrics != nullptr) {
        atomicAdd(&d.metrics[0], loss);
        atomicAdd(&d.metrics[1], corr);
      }
    }
    cluster_sync_all();
  }

  // ------------------------------------------------ backward: dgrad chain -----------------------------------
  for (int l = L - 1; l >= 0; --l) {
    const TtdTailLayer& Ly = d.layer[l];
    const int K4 = (Ly.out + 3) & ~3;
    if (l > 0) {
      const TtdTailLayer& Lp = d.layer[l - 1];  // the layer whose pre-activation gradient we produce
      const float keep = 1.0f - Lp.drop_rate, inv_keep = 1.0f / keep;
      const uint32_t thresh = dropout_thresh16(Lp.drop_rate);
      auto epi = [&](int r, int c, float v) {
        const float a_sv = Lp.act[static_cast<long long>(r) * Lp.ld_act + c];
        const bool kept = tail_kept(Lp.drop_rate, thresh, r, c, step, Lp.layer_id, s0, s1);
        Lp.dz[static_cast<long long>(r) * Lp.ld_dz + c] =
            kept ? v * inv_keep * act_grad_from_out(a_sv * keep, Lp.act_kind) : 0.0f;
      };
      tail_gemm<float>(B, Ly.in, K4, Ly.dz, Ly.ld_dz, Ly.wt, Ly.ld_wt, tid, nthreads, epi);
    } else {
      // gradient for the layer in front of the tail (the tcgen05 first layer): bf16, both layouts
      const float keep = 1.0f - d.in_drop_rate, inv_keep = 1.0f / keep;
      const uint32_t thresh = dropout_thresh16(d.in_drop_rate);
      const __nv_bfloat16* a_in = static_cast<const __nv_bfloat16*>(d.a_in);
      __nv_bfloat16* dzi = static_cast<__nv_bfloat16*>(d.dz_in);
      __nv_bfloat16* dziT = static_cast<__nv_bfloat16*>(d.dz_inT);
      auto epi = [&](int r, int c, float v) {
        const float a_sv = __bfloat162float(a_in[static_cast<long long>(r) * d.ld_a_in + c]);
        const bool kept = tail_kept(d.in_drop_rate, thresh, r, c, step, d.in_layer_id, s0, s1);
        const __nv_bfloat16 g =
            __float2bfloat16(kept ? v * inv_keep * act_grad_from_out(a_sv * keep, d.in_act_kind) : 0.0f);
        if (dzi != nullptr) dzi[static_cast<long long>(r) * d.ld_dz_in + c] = g;
        if (dziT != nullptr) dziT[static_cast<long long>(c) * d.ld_dz_inT + r] = g;
      };
      tail_gemm<float>(B, Ly.in, K4, Ly.dz, Ly.ld_dz, Ly.wt, Ly.ld_wt, tid, nthreads, epi);
    }
    cluster_sync_all();
  }

  // ------------------------------------------------ backward: all weight gradients ---------------------------
  {
    // the layers' tiles are concatenated so that every thread of the cluster gets work from the big and the
    // small layers alike
    int offset = 0;
    const int B4 = (B + 3) & ~3;
    for (int l = 0; l < L; ++l) {
      const TtdTailLayer& Ly = d.layer[l];
      const int R = Ly.in + 1, C = Ly.out;
      const int tiles = ((R + 1) >> 1) * ((C + 3) >> 2);
      auto epi = [&](int r, int c, float v) { Ly.dw[static_cast<long long>(r) * C + c] = v; };
      const int shifted = ((tid - offset) % nthreads + nthreads) % nthreads;  // rotate the deal between layers
      if (l == 0)
        tail_gemm<__nv_bfloat16>(R, C, B4, static_cast<const __nv_bfloat16*>(d.a_inT), d.ld_a_inT, Ly.dz, Ly.ld_dz,
                                 shifted, nthreads, epi);
      else
        tail_gemm<float>(R, C, B4, d.layer[l - 1].actT, d.layer[l - 1].ld_actT, Ly.dz, Ly.ld_dz, shifted, nthreads, epi);
      offset = (offset + tiles) % nthreads;
    }
  }
}

}  // namespace ttd

using namespace ttd;

extern "C" int ttd_mlp_tail(const TtdTailDesc* desc, uint64_t seed, uint32_t flags, void* stream) {
  if (desc == nullptr || desc->n_layers < 1 || desc->n_layers > TTD_TAIL_MAX_LAYERS || desc->B < 1) {
    set_error("ttd_mlp_tail: need 1..%d layers and a positive batch", TTD_TAIL_MAX_LAYERS);
    return -1;
  }
  for (int l = 0; l < desc->n_layers; ++l) {
    const TtdTailLayer& y = desc->layer[l];
    if ((y.ld_w & 3) || (y.ld_wt & 3) || (y.ld_act & 3) || (y.ld_dz & 3) || (l + 1 < desc->n_layers && (y.ld_actT & 3))) {
      set_error("ttd_mlp_tail: layer %d has a leading dimension that is not a multiple of 4 floats", l);
      return -1;
    }
  }
  TailParams P;
  P.d = *desc;
  P.seed_lo = static_cast<uint32_t>(seed);
  P.seed_hi = static_cast<uint32_t>(seed >> 32);
  P.use_pdl = (flags & 1u) ? 1 : 0;
  int cluster = desc->cluster > 0 ? desc->cluster : 8;
  static bool attr_set[64] = {false};
  int dev = 0;
  TTD_CUDA(cudaGetDevice(&dev));
  if (cluster > 8 && dev >= 0 && dev < 64 && !attr_set[dev]) {
    TTD_CUDA(cudaFuncSetAttribute(mlp_tail_kernel, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
    attr_set[dev] = true;
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(static_cast<unsigned>(cluster));
  cfg.blockDim = dim3(kTailThreads);
  cfg.stream = static_cast<cudaStream_t>(stream);
  cudaLaunchAttribute attr[2];
  int n = 0;
  attr[n].id = cudaLaunchAttributeClusterDimension;
  attr[n].val.clusterDim.x = static_cast<unsigned>(cluster);
  attr[n].val.clusterDim.y = 1;
  attr[n].val.clusterDim.z = 1;
  ++n;
  if (P.use_pdl) {
    attr[n].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[n].val.programmaticStreamSerializationAllowed = 1;
    ++n;
  }
  cfg.attrs = attr;
  cfg.numAttrs = n;
  TTD_CUDA(cudaLaunchKernelEx(&cfg, mlp_tail_kernel, P));
  g_launch_count.fetch_add(1);
  return 0;
}
