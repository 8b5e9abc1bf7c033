// Parameter-store kernels: the "ps" side of the reference, executed by the GPU that owns a shard.
//
//   param_kernel<kInit>    He / variance-scaling truncated-normal init (reference :49, A.5), zero biases
//   param_kernel<kRefresh> rebuild the bf16 compute copies from the fp32 master (after a restore)
//   param_kernel<kUpdate>  the fused gradient push:
//        [cross-GPU barrier: all replicas' gradients are in their symmetric buffers]
//        g   = grad_scale * sum_r grad_r            (peer loads over NVLink, or one multimem.ld_reduce
//                                                    that lets the NVSwitch do the sum: NVLS)
//        g  += L2/L1 term on regularised kernels    (reference :50, :55)
//        lr  = lr0 * decay^floor(gs/decay_steps)    (reference :136-140, read from the device global_step)
//        w  -= lr * g                               (reference :150-152, ApplyGradientDescent on the ps)
//        s  -= (1-d)(s-w), d=min(decay,(1+n)/(10+n)) (reference :153-161 dead code, north-star EMA)
//        bf16 compute copies (both layouts) stored into EVERY replica's buffers (the parameter pull
//        for all layers but the first, whose pull is fused into its GEMM)
//        [cross-GPU barrier: every replica holds the new parameters]
//        last block: ++global_step, ++local_step, publish + clear the metric accumulators
//
// The flat fp32 buffers hold one group per dense layer: kernel [rows, cols] immediately followed by
// bias [cols], i.e. a [rows+1, cols] matrix, which is also exactly what the wgrad GEMM writes.
#include <atomic>
#include <math.h>
#include <stdlib.h>
#include <string.h>

#include "../include/ttd_b200.h"
#include "common.cuh"
#include "step_state.cuh"

namespace ttd {

extern std::atomic<uint64_t> g_launch_count;

constexpr int kMaxVars = 16;
constexpr int kMaxPeers = 8;
constexpr int kTile = 32;    // tile columns
constexpr int kTileR = 64;   // tile rows: two 128-bit groups per thread in flight on the vector path

struct VarTable {
  TtdVarDesc v[kMaxVars];
  int tile_begin[kMaxVars + 1];
  int tile_row0[kMaxVars];  // first tile row of each group that intersects the launch's flat range
  int n;
};

struct PeerInfo {
  const float* grads[kMaxPeers];     // gradient buffer of every replica (peer-mapped)
  long long copy_delta[kMaxPeers];   // byte offset from a local bf16-copy pointer to replica r's copy
  uint32_t* flags[kMaxPeers];        // signal pad of every replica: [grid][kMaxPeers] u32
  const float* grads_mc;             // multicast mapping of the gradient buffer (NVLS) or NULL
  long long mc_delta;                // byte distance from a local arena address to its multicast alias (0: none)
  int n_grads;                       // replicas contributing gradients
  int n_copies;                      // replicas receiving the bf16 copies
  int rank;
  int barrier;                       // 1: run the cross-GPU barriers
  // asynchronous (Hogwild) mode: every replica pulls from / pushes into the owners' masters directly
  float* w_peer[kMaxPeers];          // fp32 master buffer of every replica (peer-mapped)
  long long bounds[kMaxPeers + 1];   // owner k holds flat range [bounds[k], bounds[k+1])
  TtdStepState* state0;              // step state of the owner of global_step (shard 0)
  int n_owners;
  int trace_all;                     // diagnostic: phase clocks as the maximum over all CTAs (TtdPeers.reserved bit0)
  int force_scalar;                  // [lo,hi) cuts a 4-element group of a vector-path variable: take the scalar path
  unsigned long long watchdog_ns;    // a replica barrier that waits longer than this traps (ttd_set_barrier_timeout_ms)
};

enum { kInit = 0, kRefresh = 1, kUpdate = 2, kAsyncPull = 3, kAsyncPush = 4 };

__device__ __forceinline__ int owner_of(const PeerInfo& pi, long long idx) {
  int o = 0;
#pragma unroll
  for (int k = 1; k < kMaxPeers; ++k)
    if (k < pi.n_owners && idx >= pi.bounds[k]) o = k;
  return o;
}
__device__ __forceinline__ long long ld_relaxed_sys_s64(const long long* p) {
  long long v;
  asm volatile("ld.relaxed.sys.global.s64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

// Block-wise barrier across replicas: block b of every replica meets block b of all others.
// Writes made by this block before the call are visible to the peers' block b after it.
// kRelease = false: the data the peers are about to read was written by an EARLIER kernel of this stream
// (already at the L2, which is where NVLink reads are served), so the signal is a bare store (measured:
// 1.3 us per hop, versus 2.5 us with a system fence in front). kRelease = true: this block's own remote
// stores must land first — one system-scope fence, cumulative over the block through the bar.sync.
// The wait side polls with relaxed loads; whatever consumes the peers' data afterwards is issued after the
// loop exit (control dependence), and L1 never holds peer lines across kernels.
template <bool kRelease>
__device__ __forceinline__ void replica_barrier(const PeerInfo& pi, uint32_t epoch) {
  __syncthreads();
  const int t = threadIdx.x;
  const int n = pi.n_grads;
  if (t < n) {
    if (kRelease) __threadfence_system();
    uint32_t* theirs = pi.flags[t] + blockIdx.x * kMaxPeers + pi.rank;
    asm volatile("st.relaxed.sys.global.u32 [%0], %1;" ::"l"(theirs), "r"(epoch) : "memory");
    const uint32_t* mine = pi.flags[pi.rank] + blockIdx.x * kMaxPeers + t;
    uint32_t seen;
    uint64_t t0 = 0;
    uint32_t spins = 0;
    while (true) {
      asm volatile("ld.relaxed.sys.global.u32 %0, [%1];" : "=r"(seen) : "l"(mine) : "memory");
      if (static_cast<int32_t>(seen - epoch) >= 0) break;
      if ((++spins & 0x3FFu) == 0) {  // watchdog: a dead peer raises instead of hanging the GPU
        const uint64_t now = globaltimer_ns();
        if (t0 == 0) t0 = now;
        else if (now - t0 > pi.watchdog_ns) __trap();
      }
    }
  }
  __syncthreads();
}

__device__ __forceinline__ float4 multimem_ld_reduce_f32x4(const float* mc_addr) {
  float4 r;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w)
               : "l"(mc_addr)
               : "memory");
  return r;
}

// one store, delivered by the NVSwitch to the same offset of every replica's arena
__device__ __forceinline__ void multimem_st_b128(void* mc_addr, uint4 v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc_addr),
               "f"(__uint_as_float(v.x)), "f"(__uint_as_float(v.y)), "f"(__uint_as_float(v.z)), "f"(__uint_as_float(v.w))
               : "memory");
}
__device__ __forceinline__ void multimem_st_b64(void* mc_addr, uint2 v) {
  asm volatile("multimem.st.relaxed.sys.global.v2.f32 [%0], {%1, %2};" ::"l"(mc_addr), "f"(__uint_as_float(v.x)),
               "f"(__uint_as_float(v.y))
               : "memory");
}
__device__ __forceinline__ float4 ld_cv_f32x4(const float* p) {
  float4 r;
  asm volatile("ld.global.cv.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p) : "memory");
  return r;
}
__device__ __forceinline__ uint32_t pack_bf16x2(float a, float b) {
  const __nv_bfloat162 t = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<const uint32_t*>(&t);
}
// Store `bytes` (8 or 16) to the local copy and to every other replica's copy: through the multicast alias
// when the fabric offers one, else one NVLink store per replica.
template <typename VecT>
__device__ __forceinline__ void broadcast_store(const PeerInfo& pi, void* local_ptr, VecT v, bool local_only) {
  if (local_only || pi.n_copies <= 1) {
    *reinterpret_cast<VecT*>(local_ptr) = v;
    return;
  }
  if (pi.mc_delta != 0) {
    void* mc = reinterpret_cast<char*>(local_ptr) + pi.mc_delta;
    if constexpr (sizeof(VecT) == 16) multimem_st_b128(mc, *reinterpret_cast<uint4*>(&v));
    else multimem_st_b64(mc, *reinterpret_cast<uint2*>(&v));
    return;
  }
#pragma unroll
  for (int q = 0; q < kMaxPeers; ++q)
    if (q < pi.n_copies) *reinterpret_cast<VecT*>(reinterpret_cast<char*>(local_ptr) + pi.copy_delta[q]) = v;
}

// Gradient gather: ROWS x N loads of one thread issued back to back as straight-line code (no predicates, no
// branches), every result in its own register, summed afterwards. With predicated loads and the adds in
// between, ptxas keeps at most three loads in flight and the scoreboard makes the rest queue behind them — one
// NVLink round trip per group (measured: 18 us for 8 rows x 2 replicas, 50 us for 8 x 8).
template <int N, int ROWS>
__device__ __forceinline__ void gather_f32(const PeerInfo& pi, long long base, const int* idx, float* out) {
  float t[ROWS][N];
#pragma unroll
  for (int r = 0; r < ROWS; ++r)
#pragma unroll
    for (int q = 0; q < N; ++q) t[r][q] = __ldcv(pi.grads[q] + base + idx[r]);  // written by another GPU: never L1
#pragma unroll
  for (int r = 0; r < ROWS; ++r) {
    float a = t[r][0];
#pragma unroll
    for (int q = 1; q < N; ++q) a += t[r][q];
    out[r] = a;
  }
}
template <int N>
__device__ __forceinline__ void gather_rows8(const PeerInfo& pi, long long base, const int* idx, float* out) {
  constexpr int CH = N <= 2 ? 8 : (N <= 4 ? 4 : 2);  // at most 16 loads in flight per thread (64-register cap)
#pragma unroll
  for (int b = 0; b < 8; b += CH) gather_f32<N, CH>(pi, base, idx + b, out + b);
}
template <int N, int U>
__device__ __forceinline__ void gather_vec(const PeerInfo& pi, const long long* idx, float4* out) {
  float4 t[U][N];
#pragma unroll
  for (int u = 0; u < U; ++u)
#pragma unroll
    for (int q = 0; q < N; ++q) t[u][q] = __ldcv(reinterpret_cast<const float4*>(pi.grads[q] + idx[u]));
#pragma unroll
  for (int u = 0; u < U; ++u) {
    float4 a = t[u][0];
#pragma unroll
    for (int q = 1; q < N; ++q) { a.x += t[u][q].x; a.y += t[u][q].y; a.z += t[u][q].z; a.w += t[u][q].w; }
    out[u] = a;
  }
}
template <int N>
__device__ __forceinline__ void gather_vec2(const PeerInfo& pi, const long long* idx, float4* out) {
  if constexpr (N <= 4) {  // at most 8 x 16 bytes in flight per thread (the kernel is capped at 64 registers)
    gather_vec<N, 2>(pi, idx, out);
  } else {
    gather_vec<N, 1>(pi, idx, out);
    gather_vec<N, 1>(pi, idx + 1, out + 1);
  }
}
#define TTD_GATHER_SWITCH(n, CALL)                         \
  switch (n) {                                             \
    case 1: CALL(1); break;                                \
    case 2: CALL(2); break;                                \
    case 3: CALL(3); break;                                \
    case 4: CALL(4); break;                                \
    case 5: CALL(5); break;                                \
    case 6: CALL(6); break;                                \
    case 7: CALL(7); break;                                \
    default: CALL(8); break;                               \
  }

__device__ __forceinline__ float multimem_ld_reduce_f32(const float* mc_addr) {
  float r;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.f32 %0, [%1];" : "=f"(r) : "l"(mc_addr) : "memory");
  return r;
}
// one fp32 element to every replica
__device__ __forceinline__ void broadcast_store_f32(const PeerInfo& pi, float* local_ptr, float v) {
  if (pi.n_copies <= 1) {
    *local_ptr = v;
  } else if (pi.mc_delta != 0) {
    asm volatile("multimem.st.relaxed.sys.global.f32 [%0], %1;" ::"l"(reinterpret_cast<char*>(local_ptr) + pi.mc_delta),
                 "f"(v)
                 : "memory");
  } else {
#pragma unroll
    for (int q = 0; q < kMaxPeers; ++q)
      if (q < pi.n_copies) *reinterpret_cast<float*>(reinterpret_cast<char*>(local_ptr) + pi.copy_delta[q]) = v;
  }
}
// One bf16 element per lane, neighbouring lanes at neighbouring addresses, to every replica. Lanes pair up
// (the one at the 4-byte aligned address stores both halves) so a replica receives one 4-byte store per pair —
// one multimem.st when the fabric has a multicast alias — instead of two 2-byte stores per replica.
// Must be called by all 32 lanes of the warp.
__device__ __forceinline__ void broadcast_store_bf16_lane(const PeerInfo& pi, __nv_bfloat16* dst, float val, bool on,
                                                          bool local_only) {
  const int lane = threadIdx.x & 31;
  const float nval = __shfl_down_sync(0xffffffffu, val, 1);
  const bool n_on = __shfl_down_sync(0xffffffffu, on ? 1 : 0, 1) != 0 && lane < 31;
  const bool p_on = __shfl_up_sync(0xffffffffu, on ? 1 : 0, 1) != 0 && lane > 0;
  const bool aligned = (reinterpret_cast<uintptr_t>(dst) & 3) == 0;
  if (!on) return;
  if (aligned && n_on) {
    const uint32_t pk = pack_bf16x2(val, nval);
    if (local_only || pi.n_copies <= 1) {
      *reinterpret_cast<uint32_t*>(dst) = pk;
    } else if (pi.mc_delta != 0) {
      asm volatile("multimem.st.relaxed.sys.global.bf16x2 [%0], %1;" ::"l"(reinterpret_cast<char*>(dst) + pi.mc_delta),
                   "r"(pk)
                   : "memory");
    } else {
#pragma unroll
      for (int q = 0; q < kMaxPeers; ++q)
        if (q < pi.n_copies) *reinterpret_cast<uint32_t*>(reinterpret_cast<char*>(dst) + pi.copy_delta[q]) = pk;
    }
    return;
  }
  if (!aligned && p_on) return;  // the lane below stored this element with its own
  const __nv_bfloat16 b = __float2bfloat16(val);
  if (local_only || pi.n_copies <= 1) {
    *dst = b;
    return;
  }
  for (int q = 0; q < pi.n_copies; ++q)
    *reinterpret_cast<__nv_bfloat16*>(reinterpret_cast<char*>(dst) + pi.copy_delta[q]) = b;
}

__device__ __forceinline__ float trunc_normal(uint64_t seed, uint64_t idx) {
  // Philox + Box-Muller, resampled until |z| <= 2 (TF truncated_normal), deterministic in (seed, idx)
  for (uint32_t attempt = 0; attempt < 16; ++attempt) {
    const Philox4 p = philox4x32_10(static_cast<uint32_t>(idx), static_cast<uint32_t>(idx >> 32), attempt, 0x7F4A7C15u,
                                    static_cast<uint32_t>(seed), static_cast<uint32_t>(seed >> 32));
    const float u1 = u32_to_unit_float(p.x), u2 = u32_to_unit_float(p.y);
    const float rad = sqrtf(-2.0f * logf(u1));
    const float z0 = rad * cosf(6.28318530718f * u2);
    if (fabsf(z0) <= 2.0f) return z0;
    const float z1 = rad * sinf(6.28318530718f * u2);
    if (fabsf(z1) <= 2.0f) return z1;
  }
  return 0.0f;
}

template <int MODE>
__global__ void __launch_bounds__(256, 4) param_kernel(const __grid_constant__ VarTable vt, float* __restrict__ w,
                                                    float* __restrict__ ema, const __grid_constant__ PeerInfo pi,
                                                    long long lo, long long hi, TtdHyper h, TtdStepState* st,
                                                    int bump_global_step, unsigned long long seed) {
  if (MODE == kUpdate && (bump_global_step & 2)) {  // bit1: programmatic dependent launch
    asm volatile("griddepcontrol.wait;" ::: "memory");
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  }
  __shared__ float tile[kTileR][kTile + 1];
  __shared__ unsigned char tile_ok[kTileR][kTile + 1];
  __shared__ float s_lr, s_ema_d;
  __shared__ unsigned long long s_lstep;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // TtdPeers.reserved bit0: diagnostic trace over ALL CTAs (maximum per phase; the host zeroes the slots first)
  const bool trace_all = MODE == kUpdate && (pi.trace_all != 0);
  const uint64_t k_start = (MODE == kUpdate && (blockIdx.x == 0 || trace_all) && threadIdx.x == 0) ? globaltimer_ns() : 0;

  if (MODE == kUpdate || MODE == kAsyncPull || MODE == kAsyncPush) {
    if (threadIdx.x == 0) {
      // async: the one shared global_step lives with the owner of shard 0 and may move while we look at it
      const long long gs = (MODE == kUpdate) ? st->global_step : ld_relaxed_sys_s64(reinterpret_cast<const long long*>(&pi.state0->global_step));
      double e = static_cast<double>(gs) / static_cast<double>(h.decay_steps);
      if (h.staircase) e = floor(e);
      s_lr = static_cast<float>(static_cast<double>(h.lr0) * pow(static_cast<double>(h.decay_rate), e));
      const double n = static_cast<double>(gs + 1);
      double d = h.ema_decay;
      if (h.ema_warmup) d = fmin(d, (1.0 + n) / (10.0 + n));
      s_ema_d = static_cast<float>(d);
      s_lstep = st->local_step;
    }
    __syncthreads();
    if (MODE == kUpdate && pi.barrier) replica_barrier<false>(pi, static_cast<uint32_t>(2 * s_lstep + 1));
  }
  constexpr bool kAsync = (MODE == kAsyncPull || MODE == kAsyncPush);
  const long long own_lo = lo, own_hi = hi;  // async kernels walk the whole buffer; [lo,hi) is what we own
  if (kAsync) {
    lo = 0;
    hi = vt.v[vt.n - 1].offset + static_cast<long long>(vt.v[vt.n - 1].rows + 1) * vt.v[vt.n - 1].cols;
  }
  const float lr = s_lr, ema_d = s_ema_d;
  // phase clock of CTA 0 (ns since kernel start) for the push/pull trace: st->reserved[0..4] =
  // {after barrier-in, after the gradient loads of its first tile, after its last tile, after barrier-out, #tiles}
  const bool trace = (MODE == kUpdate) && (blockIdx.x == 0 || trace_all) && threadIdx.x == 0;
  const uint64_t k_t0 = k_start;
  uint32_t tr_loads = 0, tr_ntiles = 0;
  if (trace) {
    const uint32_t v = static_cast<uint32_t>(globaltimer_ns() - k_start);
    if (trace_all) atomicMax(&st->reserved[0], v);
    else st->reserved[0] = v;
  }

  const int total_tiles = vt.tile_begin[vt.n];
  for (int t = blockIdx.x; t < total_tiles; t += gridDim.x) {
    int vi = 0;
    for (int i = 1; i < vt.n; ++i)
      if (t >= vt.tile_begin[i]) vi = i;
    const TtdVarDesc& V = vt.v[vi];
    const int R = V.rows + 1, C = V.cols;
    const int tiles_c = (C + kTile - 1) / kTile;
    const int lt = t - vt.tile_begin[vi];
    const int r0 = (lt / tiles_c + vt.tile_row0[vi]) * kTileR, c0 = (lt % tiles_c) * kTile;
    // quick reject: tile entirely outside [lo, hi)
    const long long t_lo = V.offset + static_cast<long long>(r0) * C;
    const long long t_hi = V.offset + static_cast<long long>(min(r0 + kTileR, R)) * C;
    if (t_hi <= lo || t_lo >= hi) continue;

    __syncthreads();  // previous tile's transposed reads are done
    if (MODE == kUpdate && (C & 3) == 0 && !pi.force_scalar) {
      // ---- vector tile (row pitch a multiple of 4): 128-bit traffic end to end ----------------------------
      // thread -> rows tid/8 and tid/8+32, four columns at 4*(tid%8); both rows' loads are issued before any
      // arithmetic. Gradients: ONE multimem.ld_reduce per group (the NVSwitch adds the replicas' values in
      // flight) or one 16-byte load per replica; copies: multimem.st / 16-byte stores.
      const int cg = threadIdx.x & 7;
      const int c = c0 + 4 * cg;
      const bool use_ema = ema != nullptr && h.ema_decay > 0.0f;
      float4 g4[2], w4[2], s4[2];
      long long idx2[2], idx_ld[2];
      bool ok2[2];
#pragma unroll
      for (int u = 0; u < 2; ++u) {
        const int r = r0 + (threadIdx.x >> 3) + 32 * u;
        const long long idx = V.offset + static_cast<long long>(r) * C + c;
        const bool ok = r < R && c < C && idx >= lo && idx + 3 < hi;
        idx2[u] = idx;
        ok2[u] = ok;
        idx_ld[u] = ok ? idx : V.offset;  // lanes without work load a valid address and drop the value
      }
      // every load of the thread is issued before anything consumes one (see gather_f32)
      if (pi.grads_mc != nullptr) {
#pragma unroll
        for (int u = 0; u < 2; ++u) g4[u] = multimem_ld_reduce_f32x4(pi.grads_mc + idx_ld[u]);
      } else {
#define TTD_CALL(N) gather_vec2<N>(pi, idx_ld, g4)
        TTD_GATHER_SWITCH(pi.n_grads, TTD_CALL)
#undef TTD_CALL
      }
#pragma unroll
      for (int u = 0; u < 2; ++u) {
        w4[u] = *reinterpret_cast<const float4*>(w + idx_ld[u]);
        s4[u] = use_ema ? *reinterpret_cast<const float4*>(ema + idx_ld[u]) : make_float4(0.f, 0.f, 0.f, 0.f);
      }
      if (trace) {
        if (tr_ntiles == 0) {
          if (g4[0].x + g4[1].y == 123.456f) st->reserved[5] = 1;  // the loads must have landed before the clock is read
          tr_loads = static_cast<uint32_t>(globaltimer_ns() - k_t0);
        }
        ++tr_ntiles;
      }
#pragma unroll
      for (int u = 0; u < 2; ++u) {
        const int rl = (threadIdx.x >> 3) + 32 * u, r = r0 + rl;
        const long long idx = idx2[u];
        const bool ok = ok2[u];
        float val[4] = {0.f, 0.f, 0.f, 0.f};
        if (ok) {
          const float gv[4] = {g4[u].x, g4[u].y, g4[u].z, g4[u].w}, wv4[4] = {w4[u].x, w4[u].y, w4[u].z, w4[u].w},
                      sv[4] = {s4[u].x, s4[u].y, s4[u].z, s4[u].w};
          float nv[4];
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            float g = gv[i] * h.grad_scale;
            if (V.regularize && r < V.rows) {
              if (h.reg_kind == 2) g += h.reg_scale * wv4[i];
              else if (h.reg_kind == 1) g += h.reg_scale * (wv4[i] > 0.0f ? 1.0f : (wv4[i] < 0.0f ? -1.0f : 0.0f));
            }
            val[i] = wv4[i] - lr * g;
            nv[i] = sv[i] - (1.0f - ema_d) * (sv[i] - val[i]);
          }
          *reinterpret_cast<float4*>(w + idx) = make_float4(val[0], val[1], val[2], val[3]);
          if (use_ema) *reinterpret_cast<float4*>(ema + idx) = make_float4(nv[0], nv[1], nv[2], nv[3]);
          if (V.bias_c != nullptr && r == V.rows)
            broadcast_store<float4>(pi, V.bias_c + c, make_float4(val[0], val[1], val[2], val[3]), false);
          if (V.w_f32 != nullptr && r < V.rows)  // fp32 compute copy for the fused MLP tail
            broadcast_store<float4>(pi, V.w_f32 + static_cast<long long>(r) * V.ld_f32 + c,
                                    make_float4(val[0], val[1], val[2], val[3]), false);
        }
        // TF-layout bf16 copy: neighbouring threads pair up so that the even one issues a 16-byte store
        {
          const uint32_t p0 = pack_bf16x2(val[0], val[1]), p1 = pack_bf16x2(val[2], val[3]);
          const uint32_t q0 = __shfl_down_sync(0xffffffffu, p0, 1), q1 = __shfl_down_sync(0xffffffffu, p1, 1);
          const bool nb_ok = __shfl_down_sync(0xffffffffu, ok ? 1 : 0, 1) != 0;
          const bool prev_ok = __shfl_up_sync(0xffffffffu, ok ? 1 : 0, 1) != 0;  // (all shuffles before any branch)
          const bool rows_ok = V.w_rm != nullptr && r < V.rows;
          __nv_bfloat16* dst = static_cast<__nv_bfloat16*>(V.w_rm) + static_cast<long long>(r) * V.ld_rm + c;
          if (rows_ok && (cg & 1) == 0 && ok && nb_ok) {
            broadcast_store<uint4>(pi, dst, make_uint4(p0, p1, q0, q1), false);
          } else if (rows_ok && ok && ((cg & 1) == 0 || !prev_ok)) {
            broadcast_store<uint2>(pi, dst, make_uint2(p0, p1), false);
          }
        }
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          tile[rl][4 * cg + i] = val[i];
          tile_ok[rl][4 * cg + i] = ok ? 1 : 0;
        }
      }
      if (V.w_t != nullptr || V.wt_f32 != nullptr) __syncthreads();
      if (V.wt_f32 != nullptr) {  // transposed fp32 copy: column tid/8, eight rows at 8*(tid%8)
        const int cl = threadIdx.x >> 3, rg = threadIdx.x & 7;
        const int cc = c0 + cl, rr0 = r0 + 8 * rg;
        if (cc < C) {
          float* dst = V.wt_f32 + static_cast<long long>(cc) * V.ld_tf32 + rr0;
          bool all_ok = rr0 + 7 < V.rows;
#pragma unroll
          for (int i = 0; i < 8; ++i) all_ok = all_ok && tile_ok[8 * rg + i][cl];
          if (all_ok) {
            broadcast_store<float4>(pi, dst, make_float4(tile[8 * rg][cl], tile[8 * rg + 1][cl], tile[8 * rg + 2][cl], tile[8 * rg + 3][cl]), false);
            broadcast_store<float4>(pi, dst + 4, make_float4(tile[8 * rg + 4][cl], tile[8 * rg + 5][cl], tile[8 * rg + 6][cl], tile[8 * rg + 7][cl]), false);
          } else {
            for (int i = 0; i < 8; ++i)
              if (tile_ok[8 * rg + i][cl] && rr0 + i < V.rows)
                for (int q = 0; q < pi.n_copies; ++q)
                  *reinterpret_cast<float*>(reinterpret_cast<char*>(dst + i) + pi.copy_delta[q]) = tile[8 * rg + i][cl];
          }
        }
      }
      if (V.w_t != nullptr) {
        {  // thread -> column tid/8, eight rows at 8*(tid%8): one 16-byte store of 8 bf16
          const int cl = threadIdx.x >> 3, rg = threadIdx.x & 7;
          const int cc = c0 + cl, rr0 = r0 + 8 * rg;
          bool all_ok = cc < C && rr0 + 7 < V.rows;
#pragma unroll
          for (int i = 0; i < 8; ++i) all_ok = all_ok && tile_ok[8 * rg + i][cl];
          __nv_bfloat16* dst = static_cast<__nv_bfloat16*>(V.w_t) + static_cast<long long>(cc) * V.ld_t + rr0;
          const bool local_only = (V.reserved & 1) != 0;  // operand fetched from its owner by the pull-fused GEMM
          if (all_ok) {
            uint4 pk;
            pk.x = pack_bf16x2(tile[8 * rg + 0][cl], tile[8 * rg + 1][cl]);
            pk.y = pack_bf16x2(tile[8 * rg + 2][cl], tile[8 * rg + 3][cl]);
            pk.z = pack_bf16x2(tile[8 * rg + 4][cl], tile[8 * rg + 5][cl]);
            pk.w = pack_bf16x2(tile[8 * rg + 6][cl], tile[8 * rg + 7][cl]);
            broadcast_store<uint4>(pi, dst, pk, local_only);
          } else if (cc < C) {
            for (int i = 0; i < 8; ++i) {
              if (tile_ok[8 * rg + i][cl] && rr0 + i < V.rows) {
                const __nv_bfloat16 bv = __float2bfloat16(tile[8 * rg + i][cl]);
                for (int q = 0; q < pi.n_copies; ++q) {
                  if (local_only && pi.copy_delta[q] != 0) continue;
                  *reinterpret_cast<__nv_bfloat16*>(reinterpret_cast<char*>(dst + i) + pi.copy_delta[q]) = bv;
                }
              }
            }
          }
        }
      }
      continue;
    }
    // phase 1: addresses + every load of this thread's four elements issued back to back, so the NVLink
    // round trips to the peers' gradient buffers overlap instead of queueing behind one another
    uint32_t okmask = 0;
    float gsum[kTileR / 8], wold[kTileR / 8], sold[kTileR / 8];
    int rel_ld[kTileR / 8];  // element index relative to the group's offset (a group holds < 2^31 elements)
    static_assert(kTileR / 8 == 8, "gather_rows8 handles eight rows per thread");
#pragma unroll
    for (int rr = 0; rr < kTileR / 8; ++rr) {
      const int r = r0 + warp + 8 * rr, c = c0 + lane;
      const long long idx = V.offset + static_cast<long long>(r) * C + c;
      const bool ok = r < R && c < C && idx >= lo && idx < hi;
      okmask |= (ok ? 1u : 0u) << rr;
      gsum[rr] = wold[rr] = sold[rr] = 0.0f;
      if (MODE == kUpdate) {
        rel_ld[rr] = ok ? static_cast<int>(idx - V.offset) : 0;  // lanes without work load a valid address and drop it
      } else if (MODE == kRefresh && ok) {
        wold[rr] = w[idx];
      } else if (MODE == kAsyncPull && ok) {
        wold[rr] = __ldcv(pi.w_peer[owner_of(pi, idx)] + idx);  // M1: parameter pull straight from the owner
        if (ema != nullptr && h.ema_decay > 0.0f && idx >= own_lo && idx < own_hi) sold[rr] = ema[idx];
      } else if (MODE == kAsyncPush && ok) {
        gsum[rr] = pi.grads[0][idx];
        wold[rr] = w[idx];  // the snapshot this gradient was computed with
      }
    }
    if (MODE == kUpdate) {
      if (pi.grads_mc != nullptr) {  // the switch adds the replicas' values: one request per element, not n
#pragma unroll
        for (int rr = 0; rr < kTileR / 8; ++rr) gsum[rr] = multimem_ld_reduce_f32(pi.grads_mc + V.offset + rel_ld[rr]);
      } else {
#define TTD_CALL(N) gather_rows8<N>(pi, V.offset, rel_ld, gsum)
        TTD_GATHER_SWITCH(pi.n_grads, TTD_CALL)
#undef TTD_CALL
      }
      const bool use_ema = ema != nullptr && h.ema_decay > 0.0f;
#pragma unroll
      for (int rr = 0; rr < kTileR / 8; ++rr) {
        wold[rr] = w[V.offset + rel_ld[rr]];
        sold[rr] = use_ema ? ema[V.offset + rel_ld[rr]] : 0.0f;
      }
    }
    if (trace) {
      if (tr_ntiles == 0) {
        float sink = 0.0f;
#pragma unroll
        for (int rr = 0; rr < kTileR / 8; ++rr) sink += gsum[rr];
        if (sink == 123.456f) st->reserved[5] = 1;  // forces the loads to have landed before the clock is read
        tr_loads = static_cast<uint32_t>(globaltimer_ns() - k_t0);
      }
      ++tr_ntiles;
    }
    // phase 2: arithmetic + stores
#pragma unroll
    for (int rr = 0; rr < kTileR / 8; ++rr) {
      const int rl = warp + 8 * rr;
      const int r = r0 + rl, c = c0 + lane;
      const bool ok = (okmask >> rr) & 1u;
      const long long idx = V.offset + static_cast<long long>(r) * C + c;
      float val = 0.0f;
      if (ok) {
        if (MODE == kInit) {
          val = (r < V.rows) ? trunc_normal(seed, static_cast<uint64_t>(idx)) * sqrtf(2.6f / static_cast<float>(V.rows))
                             : 0.0f;
          w[idx] = val;
          if (ema) ema[idx] = val;
        } else if (MODE == kRefresh) {
          val = wold[rr];
        } else if (MODE == kAsyncPull) {
          val = wold[rr];
          if (idx >= own_lo && idx < own_hi) {  // owner pass: shadow variables follow the live master
            if (ema != nullptr && h.ema_decay > 0.0f) ema[idx] = sold[rr] - (1.0f - ema_d) * (sold[rr] - val);
          } else {
            w[idx] = val;  // non-owned part of the local buffer = this step's fp32 snapshot
          }
        } else if (MODE == kAsyncPush) {
          float g = gsum[rr] * h.grad_scale;
          const float wv = wold[rr];
          if (V.regularize && r < V.rows) {
            if (h.reg_kind == 2) g += h.reg_scale * wv;
            else if (h.reg_kind == 1) g += h.reg_scale * (wv > 0.0f ? 1.0f : (wv < 0.0f ? -1.0f : 0.0f));
          }
          // M2: unlocked apply on the owner (use_locking=False): one fire-and-forget NVLink atomic per element
          atomicAdd_system(pi.w_peer[owner_of(pi, idx)] + idx, -lr * g);
        } else {
          float g = gsum[rr] * h.grad_scale;
          const float wv = wold[rr];
          if (V.regularize && r < V.rows) {
            if (h.reg_kind == 2) g += h.reg_scale * wv;
            else if (h.reg_kind == 1) g += h.reg_scale * (wv > 0.0f ? 1.0f : (wv < 0.0f ? -1.0f : 0.0f));
          }
          val = wv - lr * g;
          w[idx] = val;
          if (ema != nullptr && h.ema_decay > 0.0f) ema[idx] = sold[rr] - (1.0f - ema_d) * (sold[rr] - val);
        }
        if (MODE != kAsyncPush && V.bias_c != nullptr && r == V.rows) broadcast_store_f32(pi, V.bias_c + c, val);
        if (MODE != kAsyncPush && V.w_f32 != nullptr && r < V.rows)
          broadcast_store_f32(pi, V.w_f32 + static_cast<long long>(r) * V.ld_f32 + c, val);
      }
      if (MODE != kAsyncPush && V.w_rm != nullptr)  // (whole warp: the lanes pair up)
        broadcast_store_bf16_lane(pi, static_cast<__nv_bfloat16*>(V.w_rm) + static_cast<long long>(r) * V.ld_rm + c, val,
                                  ok && r < V.rows, false);
      tile[rl][lane] = val;
      tile_ok[rl][lane] = (ok && MODE != kAsyncPush) ? 1 : 0;
    }
    if (V.w_t != nullptr || V.wt_f32 != nullptr) __syncthreads();
    if (V.wt_f32 != nullptr) {
#pragma unroll
      for (int half = 0; half < kTileR / 32; ++half) {
#pragma unroll
        for (int cc = 0; cc < kTile / 8; ++cc) {
          const int cl = warp + 8 * cc, rl = lane + 32 * half;
          const int c = c0 + cl, r = r0 + rl;
          if (tile_ok[rl][cl] && r < V.rows && c < C)
            broadcast_store_f32(pi, V.wt_f32 + static_cast<long long>(c) * V.ld_tf32 + r, tile[rl][cl]);
        }
      }
    }
    if (V.w_t != nullptr) {
#pragma unroll
      for (int half = 0; half < kTileR / 32; ++half) {
#pragma unroll
        for (int cc = 0; cc < kTile / 8; ++cc) {
          const int cl = warp + 8 * cc, rl = lane + 32 * half;
          const int c = c0 + cl, r = r0 + rl;
          // flag bit0: consumers fetch this operand from its owner by TMA (pull-fused GEMM), so the
          // transposed copy is NOT broadcast; only the owner's own buffer is refreshed
          broadcast_store_bf16_lane(pi, static_cast<__nv_bfloat16*>(V.w_t) + static_cast<long long>(c) * V.ld_t + r,
                                    tile[rl][cl], tile_ok[rl][cl] && r < V.rows && c < C, (V.reserved & 1) != 0);
        }
      }
    }
  }

  if (MODE == kUpdate || MODE == kAsyncPush) {
    const uint32_t tr_tiles_done = trace ? static_cast<uint32_t>(globaltimer_ns() - k_t0) : 0;
    if (MODE == kUpdate && pi.barrier) {
      replica_barrier<true>(pi, static_cast<uint32_t>(2 * s_lstep + 2));
    }
    if (trace) {
      const uint32_t t_out = static_cast<uint32_t>(globaltimer_ns() - k_t0);
      if (trace_all) {
        atomicMax(&st->reserved[1], tr_loads);
        atomicMax(&st->reserved[2], tr_tiles_done);
        atomicMax(&st->reserved[3], t_out);
        atomicMax(&st->reserved[4], tr_ntiles);
      } else {
        st->reserved[1] = tr_loads;
        st->reserved[2] = tr_tiles_done;
        st->reserved[3] = t_out;
        st->reserved[4] = tr_ntiles;
      }
    }
    // the last block to finish advances the step counters (nobody reads them any more in this launch)
    __shared__ unsigned int s_last;
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) s_last = (atomicAdd(&st->done_counter, 1u) == gridDim.x - 1) ? 1u : 0u;
    __syncthreads();
    if (s_last && threadIdx.x == 0) {
      st->done_counter = 0;
      if (MODE == kAsyncPush) {
        // AssignAdd(global_step) on the ps that holds it; keep a local mirror for the host's fetch
        __threadfence_system();
        const unsigned long long old =
            atomicAdd_system(reinterpret_cast<unsigned long long*>(&pi.state0->global_step), 1ull);
        if (pi.state0 != st) st->global_step = static_cast<long long>(old) + 1;
      }
      finalize_step(st, h, lr, MODE != kAsyncPush && (bump_global_step & 1));
    }
  }
}

// Tile table restricted to the flat range [lo, hi): CTAs only ever visit tiles that hold work.
static int build_table(VarTable* vt, const TtdVarDesc* vars, int nvars, long long lo = 0, long long hi = -1) {
  if (nvars < 1 || nvars > kMaxVars) {
    set_error("variable table holds 1..%d groups, got %d", kMaxVars, nvars);
    return -1;
  }
  memset(vt, 0, sizeof(*vt));
  vt->n = nvars;
  int acc = 0;
  for (int i = 0; i < nvars; ++i) {
    vt->v[i] = vars[i];
    vt->tile_begin[i] = acc;
    const long long R = vars[i].rows + 1, C = vars[i].cols;
    long long r_lo = 0, r_hi = R;  // rows of this group touched by [lo, hi)
    if (hi >= 0) {
      const long long a = lo - vars[i].offset, b = hi - vars[i].offset;
      r_lo = a <= 0 ? 0 : a / C;
      r_hi = b <= 0 ? 0 : (b + C - 1) / C;
      if (r_hi > R) r_hi = R;
      if (r_lo > r_hi) r_lo = r_hi;
    }
    const int tr0 = static_cast<int>(r_lo / kTileR), tr1 = static_cast<int>((r_hi + kTileR - 1) / kTileR);
    const int tc = static_cast<int>((C + kTile - 1) / kTile);
    vt->tile_row0[i] = tr0;
    acc += (r_hi > r_lo ? (tr1 - tr0) : 0) * tc;
  }
  vt->tile_begin[nvars] = acc;
  return 0;
}

// How long a replica may sit in a cross-GPU barrier before the kernel traps. A trap destroys the CUDA context, so
// the limit has to outlast anything a peer may legitimately do between two steps — above all the chief writing
// a checkpoint (export + crc32c + file system) while the others already wait in the barrier-in of their next update.
static std::atomic<unsigned long long> g_barrier_timeout_ns{0};
static unsigned long long barrier_timeout_ns() {
  unsigned long long v = g_barrier_timeout_ns.load();
  if (v == 0) {
    v = 600ull * 1000000000ull;  // default: 10 minutes
    if (const char* e = getenv("TTD_BARRIER_TIMEOUT_SECS")) {
      const double s = atof(e);
      if (s > 0.0) v = static_cast<unsigned long long>(s * 1e9);
    }
    g_barrier_timeout_ns.store(v);
  }
  return v;
}

static int grid_for(int total_tiles) {
  const int cap = 148 * 8;
  return total_tiles < cap ? total_tiles : cap;
}

}  // namespace ttd

using namespace ttd;

extern "C" void ttd_set_barrier_timeout_ms(uint64_t ms) {
  g_barrier_timeout_ns.store(ms == 0 ? 0ull : static_cast<unsigned long long>(ms) * 1000000ull);
}

extern "C" int ttd_init_params(float* w, float* ema, const TtdVarDesc* vars, int nvars, uint64_t seed,
                               void* stream) {
  VarTable vt;
  int rc = build_table(&vt, vars, nvars);
  if (rc) return rc;
  PeerInfo pi;
  memset(&pi, 0, sizeof(pi));
  pi.n_copies = 1;
  TtdHyper h;
  memset(&h, 0, sizeof(h));
  const long long hi = vars[nvars - 1].offset + static_cast<long long>(vars[nvars - 1].rows + 1) * vars[nvars - 1].cols;
  param_kernel<kInit><<<grid_for(vt.tile_begin[vt.n]), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      vt, w, ema, pi, 0, hi, h, nullptr, 0, seed);
  TTD_CUDA(cudaGetLastError());
  g_launch_count.fetch_add(1);
  return 0;
}

extern "C" int ttd_refresh_copies(const float* w, const TtdVarDesc* vars, int nvars, void* stream) {
  VarTable vt;
  int rc = build_table(&vt, vars, nvars);
  if (rc) return rc;
  PeerInfo pi;
  memset(&pi, 0, sizeof(pi));
  pi.n_copies = 1;
  TtdHyper h;
  memset(&h, 0, sizeof(h));
  const long long hi = vars[nvars - 1].offset + static_cast<long long>(vars[nvars - 1].rows + 1) * vars[nvars - 1].cols;
  param_kernel<kRefresh><<<grid_for(vt.tile_begin[vt.n]), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      vt, const_cast<float*>(w), nullptr, pi, 0, hi, h, nullptr, 0, 0);
  TTD_CUDA(cudaGetLastError());
  g_launch_count.fetch_add(1);
  return 0;
}

extern "C" int ttd_apply_update(float* w, float* ema, const TtdVarDesc* vars, int nvars, int64_t lo, int64_t hi,
                                const TtdHyper* hyper, const TtdPeers* peers, TtdStepState* state,
                                int32_t bump_global_step, void* stream) {
  VarTable vt;
  if (hi < 0 && nvars >= 1 && nvars <= kMaxVars)  // negative hi: up to the end of the last group
    hi = vars[nvars - 1].offset + static_cast<int64_t>(vars[nvars - 1].rows + 1) * vars[nvars - 1].cols;
  int rc = build_table(&vt, vars, nvars, lo, hi);
  if (rc) return rc;
  if (peers == nullptr || peers->n_grads < 1 || peers->n_grads > kMaxPeers || peers->n_copies < 1 ||
      peers->n_copies > kMaxPeers) {
    set_error("ttd_apply_update: need 1..%d gradient sources and copy targets", kMaxPeers);
    return -1;
  }
  PeerInfo pi;
  memset(&pi, 0, sizeof(pi));
  for (int i = 0; i < kMaxPeers; ++i) {
    pi.grads[i] = peers->grads[i];
    pi.copy_delta[i] = peers->copy_delta[i];
    pi.flags[i] = peers->flags[i];
  }
  pi.grads_mc = peers->grads_mc;
  pi.mc_delta = peers->mc_delta;
  pi.n_grads = peers->n_grads;
  pi.n_copies = peers->n_copies;
  pi.rank = peers->rank;
  pi.barrier = peers->barrier;
  pi.trace_all = peers->reserved & 1;
  pi.watchdog_ns = barrier_timeout_ns();
  // The 128-bit path updates whole 4-element groups; a range boundary inside such a group (possible only for
  // callers that pick their own [lo, hi): the flat plan's bounds are multiples of 4) selects the scalar path.
  for (int i = 0; i < nvars && hi >= 0; ++i) {
    if (vars[i].cols % 4 != 0) continue;
    const long long v_lo = vars[i].offset, v_hi = v_lo + static_cast<long long>(vars[i].rows + 1) * vars[i].cols;
    if ((lo > v_lo && lo < v_hi && ((lo - v_lo) & 3)) || (hi > v_lo && hi < v_hi && ((hi - v_lo) & 3))) pi.force_scalar = 1;
  }
  int grid = grid_for(vt.tile_begin[vt.n] > 0 ? vt.tile_begin[vt.n] : 1);
  if (pi.barrier) {
    // blocks of different replicas wait on one another: all of them must be resident at once
    if (peers->grid <= 0 || peers->grid > 148 * 4) {
      set_error("ttd_apply_update: barrier mode needs an explicit grid in [1,592], got %d", peers->grid);
      return -1;
    }
    grid = peers->grid;
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(static_cast<unsigned>(grid));
  cfg.blockDim = dim3(256);
  cfg.stream = static_cast<cudaStream_t>(stream);
  cudaLaunchAttribute attr[1];
  if (bump_global_step & 2) {  // PDL: overlap this launch with the tail of the last wgrad kernel
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
  }
  TTD_CUDA(cudaLaunchKernelEx(&cfg, param_kernel<kUpdate>, vt, w, ema, pi, static_cast<long long>(lo),
                              static_cast<long long>(hi), *hyper, state, static_cast<int>(bump_global_step), 0ull));
  g_launch_count.fetch_add(1);
  return 0;
}

static void fill_async(PeerInfo* pi, const TtdAsyncPeers* a, const float* local_grads) {
  memset(pi, 0, sizeof(*pi));
  for (int i = 0; i < kMaxPeers; ++i) pi->w_peer[i] = a->w_peer[i];
  for (int i = 0; i <= kMaxPeers; ++i) pi->bounds[i] = a->bounds[i];
  pi->state0 = a->state0;
  pi->n_owners = a->n_owners;
  pi->rank = a->rank;
  pi->grads[0] = local_grads;
  pi->n_grads = 1;
  pi->n_copies = 1;
}

extern "C" int ttd_async_pull(float* w, float* ema, const TtdVarDesc* vars, int nvars, int64_t own_lo, int64_t own_hi,
                              const TtdHyper* hyper, const TtdAsyncPeers* peers, TtdStepState* state, void* stream) {
  VarTable vt;
  int rc = build_table(&vt, vars, nvars);
  if (rc) return rc;
  if (peers == nullptr || peers->n_owners < 1 || peers->n_owners > kMaxPeers || peers->state0 == nullptr) {
    set_error("ttd_async_pull: need 1..%d owners and the step state of owner 0", kMaxPeers);
    return -1;
  }
  PeerInfo pi;
  fill_async(&pi, peers, nullptr);
  param_kernel<kAsyncPull><<<grid_for(vt.tile_begin[vt.n]), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      vt, w, ema, pi, own_lo, own_hi, *hyper, state, 0, 0);
  TTD_CUDA(cudaGetLastError());
  g_launch_count.fetch_add(1);
  return 0;
}

extern "C" int ttd_async_push(float* w, const float* grads, const TtdVarDesc* vars, int nvars, const TtdHyper* hyper,
                              const TtdAsyncPeers* peers, TtdStepState* state, void* stream) {
  VarTable vt;
  int rc = build_table(&vt, vars, nvars);
  if (rc) return rc;
  if (peers == nullptr || peers->n_owners < 1 || peers->n_owners > kMaxPeers || peers->state0 == nullptr) {
    set_error("ttd_async_push: need 1..%d owners and the step state of owner 0", kMaxPeers);
    return -1;
  }
  PeerInfo pi;
  fill_async(&pi, peers, grads);
  param_kernel<kAsyncPush><<<grid_for(vt.tile_begin[vt.n]), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      vt, w, nullptr, pi, 0, 0, *hyper, state, 1, 0);
  TTD_CUDA(cudaGetLastError());
  g_launch_count.fetch_add(1);
  return 0;
}

extern "C" int ttd_copy_d2d(void* dst, const void* src, uint64_t nbytes, void* stream) {
  TTD_CUDA(cudaMemcpyAsync(dst, src, nbytes, cudaMemcpyDeviceToDevice, static_cast<cudaStream_t>(stream)));
  return 0;
}
