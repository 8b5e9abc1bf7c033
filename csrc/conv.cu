// LeNet-5 convolution stages (BASELINE config 5; no counterpart in the reference, which only has the MLP).
//
// A stage is conv(KHxKW, stride 1, SAME or VALID) -> bias -> ELU/ReLU -> 2x2 max-pool, NHWC, fp32 activations,
// bf16 weights taken from the parameter store's TF-layout compute copy ([KH*KW*Cin, Cout] row-major, i.e. the
// HWIO kernel flattened). With 1..6 input channels and 6..16 output channels the tensor cores have nothing to
// chew on (the GEMM K dimension is 25 / 150, N is 6 / 16), so these are direct CUDA-core kernels; the dense
// layers behind them go through the tcgen05 GEMM.
//
//   conv_pool_fwd     one thread per pooled output (b, py, px, co): four conv taps -> act -> max + argmax
//   conv_pool_dgrad   gradient w.r.t. the stage input, routed through the argmax of every pooled cell
//   conv_pool_wgrad   one CTA per kernel row k = (kh, kw, ci) (+1 CTA for the bias row): deterministic
//                     block reduction over all pooled cells, written straight into the flat gradient buffer
#include <atomic>
#include <math.h>

#include "../include/ttd_b200.h"
#include "common.cuh"

namespace ttd {

extern std::atomic<uint64_t> g_launch_count;

struct ConvGeom {
  int B, H, W, Cin, Cout, KH, KW, pad;  // input [B,H,W,Cin]; conv output [B,OH,OW,Cout]; pooled [B,PH,PW,Cout]
  int OH, OW, PH, PW;
};

__global__ void __launch_bounds__(256) conv_pool_fwd_kernel(ConvGeom g, const float* __restrict__ x,
                                                            const __nv_bfloat16* __restrict__ w, long long ldw,
                                                            const float* __restrict__ bias, int act,
                                                            float* __restrict__ pooled, unsigned char* __restrict__ amax,
                                                            __nv_bfloat16* __restrict__ flat, long long ldflat,
                                                            __nv_bfloat16* __restrict__ flatT, long long ldflatT) {
  extern __shared__ float sw[];  // [KH*KW*Cin][Cout] weights + [Cout] bias
  const int K = g.KH * g.KW * g.Cin;
  for (int i = threadIdx.x; i < K * g.Cout; i += blockDim.x) sw[i] = __bfloat162float(w[(i / g.Cout) * ldw + i % g.Cout]);
  for (int i = threadIdx.x; i < g.Cout; i += blockDim.x) sw[K * g.Cout + i] = bias[i];
  __syncthreads();
  const long long total = static_cast<long long>(g.B) * g.PH * g.PW * g.Cout;
  for (long long t = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; t < total;
       t += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int co = static_cast<int>(t % g.Cout);
    long long r = t / g.Cout;
    const int px = static_cast<int>(r % g.PW); r /= g.PW;
    const int py = static_cast<int>(r % g.PH);
    const int b = static_cast<int>(r / g.PH);
    float best = -INFINITY;
    int best_pos = 0;
#pragma unroll
    for (int pos = 0; pos < 4; ++pos) {
      const int oy = 2 * py + (pos >> 1), ox = 2 * px + (pos & 1);
      float acc = sw[K * g.Cout + co];
      for (int kh = 0; kh < g.KH; ++kh) {
        const int iy = oy + kh - g.pad;
        if (iy < 0 || iy >= g.H) continue;
        for (int kw = 0; kw < g.KW; ++kw) {
          const int ix = ox + kw - g.pad;
          if (ix < 0 || ix >= g.W) continue;
          const float* xp = x + ((static_cast<long long>(b) * g.H + iy) * g.W + ix) * g.Cin;
          const float* wp = sw + ((kh * g.KW + kw) * g.Cin) * g.Cout + co;
          for (int ci = 0; ci < g.Cin; ++ci) acc += xp[ci] * wp[ci * g.Cout];
        }
      }
      if (acc > best) { best = acc; best_pos = pos; }
    }
    const float a = act_apply(best, act);  // activation is monotonic: act(max) == max(act)
    pooled[t] = a;
    amax[t] = static_cast<unsigned char>(best_pos);
    if (flat != nullptr) {
      const int f = (py * g.PW + px) * g.Cout + co;  // TF flatten order of [PH, PW, Cout]
      flat[static_cast<long long>(b) * ldflat + f] = __float2bfloat16(a);
      if (flatT != nullptr) flatT[static_cast<long long>(f) * ldflatT + b] = __float2bfloat16(a);
    }
  }
}

// dzp[b,py,px,co] = (gradient w.r.t. the pooled activation) * act'(pooled) lives at the argmax tap of its cell.
template <typename TG>
__device__ __forceinline__ float load_dzp(const TG* dzp, long long i);
template <>
__device__ __forceinline__ float load_dzp<float>(const float* dzp, long long i) { return dzp[i]; }
template <>
__device__ __forceinline__ float load_dzp<__nv_bfloat16>(const __nv_bfloat16* dzp, long long i) {
  return __bfloat162float(dzp[i]);
}

// dX[b,iy,ix,ci] = sum_{kh,kw,co} dconv[b, iy-kh+pad, ix-kw+pad, co] * w[kh,kw,ci,co]; then * act'(x_act) so the
// result is directly the dzp of the stage below (x_act = that stage's pooled activation = this stage's input).
template <typename TG>
__global__ void __launch_bounds__(256) conv_pool_dgrad_kernel(ConvGeom g, const TG* __restrict__ dzp, long long ld_dzp,
                                                              const unsigned char* __restrict__ amax,
                                                              const __nv_bfloat16* __restrict__ w, long long ldw,
                                                              const float* __restrict__ x_act, int act_below,
                                                              float* __restrict__ dx) {
  extern __shared__ float sw[];
  const int K = g.KH * g.KW * g.Cin;
  for (int i = threadIdx.x; i < K * g.Cout; i += blockDim.x) sw[i] = __bfloat162float(w[(i / g.Cout) * ldw + i % g.Cout]);
  __syncthreads();
  const long long total = static_cast<long long>(g.B) * g.H * g.W * g.Cin;
  const int cell_feats = g.PH * g.PW * g.Cout;
  for (long long t = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; t < total;
       t += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int ci = static_cast<int>(t % g.Cin);
    long long r = t / g.Cin;
    const int ix = static_cast<int>(r % g.W); r /= g.W;
    const int iy = static_cast<int>(r % g.H);
    const int b = static_cast<int>(r / g.H);
    float acc = 0.0f;
    for (int kh = 0; kh < g.KH; ++kh) {
      const int oy = iy - kh + g.pad;
      if (oy < 0 || oy >= 2 * g.PH) continue;
      for (int kw = 0; kw < g.KW; ++kw) {
        const int ox = ix - kw + g.pad;
        if (ox < 0 || ox >= 2 * g.PW) continue;
        const int cell = ((oy >> 1) * g.PW + (ox >> 1)) * g.Cout;
        const int pos = ((oy & 1) << 1) | (ox & 1);
        const unsigned char* am = amax + static_cast<long long>(b) * cell_feats + cell;
        const float* wp = sw + ((kh * g.KW + kw) * g.Cin + ci) * g.Cout;
        for (int co = 0; co < g.Cout; ++co)
          if (am[co] == pos) acc += load_dzp<TG>(dzp, static_cast<long long>(b) * ld_dzp + cell + co) * wp[co];
      }
    }
    dx[t] = acc * act_grad_from_out(x_act[t], act_below);
  }
}

// dW[k, co] (k = (kh*KW+kw)*Cin+ci, or k == K for the bias) = sum over pooled cells of dzp * x at the argmax tap.
template <typename TG>
__global__ void __launch_bounds__(256) conv_pool_wgrad_kernel(ConvGeom g, const TG* __restrict__ dzp, long long ld_dzp,
                                                              const unsigned char* __restrict__ amax,
                                                              const float* __restrict__ x, float* __restrict__ dw) {
  const int K = g.KH * g.KW * g.Cin;
  const int k = blockIdx.x;  // 0..K (K = bias row)
  const int ci = k % g.Cin, kw = (k / g.Cin) % g.KW, kh = k / (g.Cin * g.KW);
  const int cell_feats = g.PH * g.PW * g.Cout;
  float acc[16];
#pragma unroll
  for (int c = 0; c < 16; ++c) acc[c] = 0.0f;
  const long long cells = static_cast<long long>(g.B) * g.PH * g.PW;
  for (long long cidx = threadIdx.x; cidx < cells; cidx += blockDim.x) {
    const int px = static_cast<int>(cidx % g.PW);
    const int py = static_cast<int>((cidx / g.PW) % g.PH);
    const int b = static_cast<int>(cidx / (static_cast<long long>(g.PW) * g.PH));
    const long long base = static_cast<long long>(b) * cell_feats + (py * g.PW + px) * g.Cout;
    const long long gbase = static_cast<long long>(b) * ld_dzp + (py * g.PW + px) * g.Cout;
#pragma unroll
    for (int co = 0; co < 16; ++co) {
      if (co < g.Cout) {
        const float d = load_dzp<TG>(dzp, gbase + co);
        if (k == K) {
          acc[co] += d;
        } else {
          const int pos = amax[base + co];
          const int iy = 2 * py + (pos >> 1) + kh - g.pad, ix = 2 * px + (pos & 1) + kw - g.pad;
          if (iy >= 0 && iy < g.H && ix >= 0 && ix < g.W)
            acc[co] += d * x[((static_cast<long long>(b) * g.H + iy) * g.W + ix) * g.Cin + ci];
        }
      }
    }
  }
  __shared__ float red[8][16];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
#pragma unroll
  for (int co = 0; co < 16; ++co) {
    const float v = warp_sum(acc[co]);
    if (lane == 0) red[warp][co] = v;
  }
  __syncthreads();
  if (threadIdx.x < g.Cout) {
    float s = 0.0f;
    for (int wv = 0; wv < 8; ++wv) s += red[wv][threadIdx.x];
    dw[static_cast<long long>(k) * g.Cout + threadIdx.x] = s;
  }
}

static int make_geom(ConvGeom* g, const TtdConvDesc* d) {
  if (d->Cout > 16 || d->Cout < 1 || d->Cin < 1 || d->KH < 1 || d->KW < 1) {
    set_error("conv stage supports 1..16 output channels (got %d)", d->Cout);
    return -1;
  }
  g->B = d->B; g->H = d->H; g->W = d->W; g->Cin = d->Cin; g->Cout = d->Cout; g->KH = d->KH; g->KW = d->KW;
  g->pad = d->same_pad ? (d->KH - 1) / 2 : 0;
  g->OH = d->same_pad ? d->H : d->H - d->KH + 1;
  g->OW = d->same_pad ? d->W : d->W - d->KW + 1;
  if (g->OH % 2 || g->OW % 2 || g->OH < 2) {
    set_error("conv output %dx%d must be even for the 2x2 pool", g->OH, g->OW);
    return -1;
  }
  g->PH = g->OH / 2;
  g->PW = g->OW / 2;
  return 0;
}

}  // namespace ttd

using namespace ttd;

extern "C" int ttd_conv_pool_fwd(const TtdConvDesc* d, const float* x, const void* w, int64_t ldw, const float* bias,
                                 int32_t act, float* pooled, uint8_t* argmax, void* flat, int64_t ldflat, void* flatT,
                                 int64_t ldflatT, void* stream) {
  ConvGeom g;
  int rc = make_geom(&g, d);
  if (rc) return rc;
  const long long total = static_cast<long long>(g.B) * g.PH * g.PW * g.Cout;
  const int blocks = static_cast<int>((total + 255) / 256 < 148 * 8 ? (total + 255) / 256 : 148 * 8);
  const size_t smem = (static_cast<size_t>(g.KH) * g.KW * g.Cin * g.Cout + g.Cout) * sizeof(float);
  conv_pool_fwd_kernel<<<blocks, 256, smem, static_cast<cudaStream_t>(stream)>>>(
      g, x, static_cast<const __nv_bfloat16*>(w), ldw, bias, act, pooled, argmax, static_cast<__nv_bfloat16*>(flat),
      ldflat, static_cast<__nv_bfloat16*>(flatT), ldflatT);
  TTD_CUDA(cudaGetLastError());
  g_launch_count.fetch_add(1);
  return 0;
}

extern "C" int ttd_conv_pool_bwd(const TtdConvDesc* d, const void* dzp, int64_t ld_dzp, int32_t dzp_is_bf16,
                                 const uint8_t* argmax, const float* x, const void* w, int64_t ldw, float* dw,
                                 float* dx, const float* x_act, int32_t act_below, void* stream) {
  ConvGeom g;
  int rc = make_geom(&g, d);
  if (rc) return rc;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const int K = g.KH * g.KW * g.Cin;
  if (dzp_is_bf16)
    conv_pool_wgrad_kernel<__nv_bfloat16><<<K + 1, 256, 0, s>>>(g, static_cast<const __nv_bfloat16*>(dzp), ld_dzp, argmax, x, dw);
  else
    conv_pool_wgrad_kernel<float><<<K + 1, 256, 0, s>>>(g, static_cast<const float*>(dzp), ld_dzp, argmax, x, dw);
  TTD_CUDA(cudaGetLastError());
  g_launch_count.fetch_add(1);
  if (dx != nullptr) {
    const long long total = static_cast<long long>(g.B) * g.H * g.W * g.Cin;
    const int blocks = static_cast<int>((total + 255) / 256 < 148 * 8 ? (total + 255) / 256 : 148 * 8);
    const size_t smem = static_cast<size_t>(K) * g.Cout * sizeof(float);
    if (dzp_is_bf16)
      conv_pool_dgrad_kernel<__nv_bfloat16><<<blocks, 256, smem, s>>>(g, static_cast<const __nv_bfloat16*>(dzp), ld_dzp, argmax,
                                                                   static_cast<const __nv_bfloat16*>(w), ldw, x_act, act_below, dx);
    else
      conv_pool_dgrad_kernel<float><<<blocks, 256, smem, s>>>(g, static_cast<const float*>(dzp), ld_dzp, argmax,
                                                           static_cast<const __nv_bfloat16*>(w), ldw, x_act, act_below, dx);
    TTD_CUDA(cudaGetLastError());
    g_launch_count.fetch_add(1);
  }
  return 0;
}
