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
//   conv_pool_wgrad   CTA (k, s): kernel row k = (kh, kw, ci) (+1 row for the bias) over slice s of the pooled
//                     cells; block reduction, then the last CTA of a row adds the slices in order (deterministic,
//                     no atomics on the data) and writes straight into the flat gradient buffer
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

// 5x5 specialisation (LeNet): the four conv taps of a pooled cell read one shared 6x6 input window, loaded once
// per input channel into registers (36 loads instead of 100), with the 25 x 4 multiply-adds fully unrolled.
__global__ void __launch_bounds__(256, 2) conv5_pool_fwd_kernel(ConvGeom g, const float* __restrict__ x,
                                                             const __nv_bfloat16* __restrict__ w, long long ldw,
                                                             const float* __restrict__ bias, int act,
                                                             float* __restrict__ pooled, unsigned char* __restrict__ amax,
                                                             __nv_bfloat16* __restrict__ flat, long long ldflat,
                                                             __nv_bfloat16* __restrict__ flatT, long long ldflatT) {
  extern __shared__ float sw[];  // [25*Cin][Cout] weights + [Cout] bias
  const int K = 25 * g.Cin;
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
    const int y0 = 2 * py - g.pad, x0 = 2 * px - g.pad;
    const float bv = sw[K * g.Cout + co];
    float acc[4] = {bv, bv, bv, bv};
    for (int ci = 0; ci < g.Cin; ++ci) {
      float win[6][6];
#pragma unroll
      for (int wy = 0; wy < 6; ++wy) {
        const int iy = y0 + wy;
#pragma unroll
        for (int wx = 0; wx < 6; ++wx) {
          const int ix = x0 + wx;
          const bool in = iy >= 0 && iy < g.H && ix >= 0 && ix < g.W;
          win[wy][wx] = in ? x[((static_cast<long long>(b) * g.H + iy) * g.W + ix) * g.Cin + ci] : 0.0f;
        }
      }
      const float* wp = sw + ci * g.Cout + co;
#pragma unroll
      for (int kh = 0; kh < 5; ++kh) {
#pragma unroll
        for (int kw = 0; kw < 5; ++kw) {
          const float wv = wp[(kh * 5 + kw) * g.Cin * g.Cout];
          acc[0] += win[kh][kw] * wv;
          acc[1] += win[kh][kw + 1] * wv;
          acc[2] += win[kh + 1][kw] * wv;
          acc[3] += win[kh + 1][kw + 1] * wv;
        }
      }
    }
    float best = acc[0];
    int best_pos = 0;
#pragma unroll
    for (int pos = 1; pos < 4; ++pos)
      if (acc[pos] > best) { best = acc[pos]; best_pos = pos; }
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
                                                              const float* __restrict__ x, float* __restrict__ dw,
                                                              float* __restrict__ partial, unsigned int* __restrict__ counters) {
  // grid = (K + 1 kernel rows, S slices of the pooled cells). With S > 1 every CTA leaves its partial row in
  // `partial[s][k][16]`; the last CTA to arrive for row k adds the S partials in slice order (deterministic)
  // and re-arms the row's counter for the next launch.
  const int K = g.KH * g.KW * g.Cin;
  const int k = blockIdx.x;  // 0..K (K = bias row)
  const int S = gridDim.y, sl = blockIdx.y;
  const int ci = k % g.Cin, kw = (k / g.Cin) % g.KW, kh = k / (g.Cin * g.KW);
  const int cell_feats = g.PH * g.PW * g.Cout;
  float acc[16];
#pragma unroll
  for (int c = 0; c < 16; ++c) acc[c] = 0.0f;
  const long long cells = static_cast<long long>(g.B) * g.PH * g.PW;
  const long long chunk = (cells + S - 1) / S;
  const long long c_lo = sl * chunk, c_hi = (c_lo + chunk < cells) ? c_lo + chunk : cells;
  for (long long cidx = c_lo + threadIdx.x; cidx < c_hi; cidx += blockDim.x) {
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
  float s = 0.0f;
  if (threadIdx.x < g.Cout)
    for (int wv = 0; wv < 8; ++wv) s += red[wv][threadIdx.x];
  if (S == 1) {
    if (threadIdx.x < g.Cout) dw[static_cast<long long>(k) * g.Cout + threadIdx.x] = s;
    return;
  }
  if (threadIdx.x < g.Cout) {
    partial[(static_cast<long long>(sl) * (K + 1) + k) * 16 + threadIdx.x] = s;
    __threadfence();
  }
  __shared__ unsigned int s_last;
  __syncthreads();
  if (threadIdx.x == 0) s_last = (atomicAdd(&counters[k], 1u) == static_cast<unsigned int>(S - 1)) ? 1u : 0u;
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  if (threadIdx.x < g.Cout) {
    float t = 0.0f;
    for (int q = 0; q < S; ++q) t += __ldcg(&partial[(static_cast<long long>(q) * (K + 1) + k) * 16 + threadIdx.x]);
    dw[static_cast<long long>(k) * g.Cout + threadIdx.x] = t;
  }
  if (threadIdx.x == 0) counters[k] = 0;
}

// slices of the pooled cells per kernel row: enough CTAs to fill the GPU, at least ~2 passes of 256 cells each
static int wgrad_slices(const ConvGeom& g, int K, long long scratch_floats) {
  const long long cells = static_cast<long long>(g.B) * g.PH * g.PW;
  long long s = (148 * 4 + K) / (K + 1);
  const long long by_work = cells / 512;
  if (s > by_work) s = by_work;
  if (s > 64) s = 64;
  const long long head = ((K + 1 + 63) / 64) * 64;  // counters, padded
  while (s > 1 && head + s * (K + 1) * 16 > scratch_floats) --s;
  return s < 1 ? 1 : static_cast<int>(s);
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
  if (g.KH == 5 && g.KW == 5)
    conv5_pool_fwd_kernel<<<blocks, 256, smem, static_cast<cudaStream_t>(stream)>>>(
        g, x, static_cast<const __nv_bfloat16*>(w), ldw, bias, act, pooled, argmax, static_cast<__nv_bfloat16*>(flat),
        ldflat, static_cast<__nv_bfloat16*>(flatT), ldflatT);
  else
    conv_pool_fwd_kernel<<<blocks, 256, smem, static_cast<cudaStream_t>(stream)>>>(
        g, x, static_cast<const __nv_bfloat16*>(w), ldw, bias, act, pooled, argmax, static_cast<__nv_bfloat16*>(flat),
        ldflat, static_cast<__nv_bfloat16*>(flatT), ldflatT);
  TTD_CUDA(cudaGetLastError());
  g_launch_count.fetch_add(1);
  return 0;
}

extern "C" int ttd_conv_pool_bwd(const TtdConvDesc* d, const void* dzp, int64_t ld_dzp, int32_t dzp_is_bf16,
                                 const uint8_t* argmax, const float* x, const void* w, int64_t ldw, float* dw,
                                 float* dx, const float* x_act, int32_t act_below, float* scratch,
                                 int64_t scratch_floats, void* stream) {
  ConvGeom g;
  int rc = make_geom(&g, d);
  if (rc) return rc;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const int K = g.KH * g.KW * g.Cin;
  const int slices = scratch != nullptr ? wgrad_slices(g, K, scratch_floats) : 1;
  const dim3 wgrid(K + 1, slices);
  unsigned int* counters = reinterpret_cast<unsigned int*>(scratch);  // zero on entry, re-armed by the kernel
  float* partial = scratch != nullptr ? scratch + ((K + 1 + 63) / 64) * 64 : nullptr;
  if (dzp_is_bf16)
    conv_pool_wgrad_kernel<__nv_bfloat16><<<wgrid, 256, 0, s>>>(g, static_cast<const __nv_bfloat16*>(dzp), ld_dzp, argmax, x, dw,
                                                              partial, counters);
  else
    conv_pool_wgrad_kernel<float><<<wgrid, 256, 0, s>>>(g, static_cast<const float*>(dzp), ld_dzp, argmax, x, dw, partial,
                                                      counters);
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
