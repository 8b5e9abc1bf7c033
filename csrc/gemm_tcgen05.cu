// Grouped bf16 GEMM for sm_100a: D[M,N] = A[M,K] * B[N,K]^T with fused training epilogues.
//
//   * operands arrive by TMA (cp.async.bulk.tensor, 128-byte swizzle) into a multi-stage smem ring,
//   * one elected thread issues tcgen05.mma (cta_group::1, M=128, N=block_n, K=16 per instruction),
//   * the fp32 accumulator lives in TMEM and is read back with tcgen05.ld by four epilogue warps,
//   * the epilogue fuses bias + ELU/ReLU + Philox dropout (forward), the activation/dropout
//     derivative (dgrad), the fp32 scale-and-store into the flat gradient buffer (wgrad) or the whole
//     sparse-softmax cross-entropy + top-1 + dlogits (output layer),
//   * up to four independent problems ride in one launch (e.g. dgrad + wgrad of a layer).
//
// What it reproduces: MatMul/BiasAdd/Elu/dropout of forward_propagation (reference
// distribute_training.py:53-61), SparseSoftmaxCrossEntropyWithLogits + mean (:80-82), in_top_k (:96)
// and their gradients as created by opt.minimize (:152).
#include <math.h>
#include <atomic>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "../include/ttd_b200.h"
#include <vector>

#include "common.cuh"
#include "step_state.cuh"

namespace ttd {

constexpr int kBlockM = 128;
constexpr int kBlockK = 64;                       // 64 bf16 = 128 B = one swizzle row
constexpr int kAStageBytes = kBlockM * kBlockK * 2;  // 16 KiB
constexpr int kMaxProblems = 4;
constexpr int kThreads = 256;                     // warp0 TMA, warp1 MMA, warp2 TMEM alloc, warps4-7 epilogue

struct GemmProblem {
  CUtensorMap tma_a;
  CUtensorMap tma_b;
  int M, N, K;
  int block_n;
  int tiles_m, tiles_n;
  int tile_begin;
  int mode;
  int act;
  float drop_rate;
  uint32_t layer_id;
  const float* bias;
  __nv_bfloat16* out;
  long long ldo;
  __nv_bfloat16* outT;
  long long ldt;
  float* outf;
  long long ldf;
  const __nv_bfloat16* aux;
  long long ldaux;
  float scale;
  const long long* labels;
  float* metrics;
  int swap;  // 1: operands were exchanged by the host (rows of the tile = output features, columns = batch)
  // TTD_EPI_WGRAD_UPDATE: the tile's rows are rows of the variable group [rows+1, cols] (last row = bias)
  float* upd_w;
  float* upd_ema;
  float* upd_bias;
  int upd_regularize;
  // A operand taken from a pool of equally shaped matrices stacked along the rows: slot (*step_ptr % a_slots)
  int a_slot_rows, a_slots;
};

struct GemmGroup {
  GemmProblem p[kMaxProblems];
  // Pull-fused B operand of problem 0 (the all-gather of the first layer's weights fused into its GEMM):
  // k-block kb is fetched by TMA from the replica that owns it, through that replica's peer mapping.
  CUtensorMap tma_b_peer[TTD_MAX_PEERS];
  int kb_bound[TTD_MAX_PEERS + 1];  // owner o serves k-blocks [kb_bound[o], kb_bound[o+1])
  int n_b_peers;                    // 0: problem 0 reads its B operand locally like every other problem
  int count;
  uint32_t seed_lo, seed_hi;
  const unsigned long long* step_ptr;
  int stages;
  int b_stage_bytes;
  int tmem_cols;
  int use_pdl;
  int cluster;  // >1: CTAs of a cluster own consecutive N tiles of one M tile and share the A tile by TMA multicast
  int kspan;    // k-blocks (of 64) per pipeline stage: 2 halves the barrier traffic of long-K weight-streaming GEMMs
};

__device__ __forceinline__ void store_bf16x8(__nv_bfloat16* dst, const float* v) {
  uint4 u;
  __nv_bfloat162 t;
  t = __floats2bfloat162_rn(v[0], v[1]); u.x = *reinterpret_cast<uint32_t*>(&t);
  t = __floats2bfloat162_rn(v[2], v[3]); u.y = *reinterpret_cast<uint32_t*>(&t);
  t = __floats2bfloat162_rn(v[4], v[5]); u.z = *reinterpret_cast<uint32_t*>(&t);
  t = __floats2bfloat162_rn(v[6], v[7]); u.w = *reinterpret_cast<uint32_t*>(&t);
  *reinterpret_cast<uint4*>(dst) = u;
}

// Write 16 consecutive columns [n0, n0+16) of row m: row-major bf16 (16-byte stores where the padded
// row allows it) and/or the transposed copy (coalesced across the warp because lane == row).
__device__ __forceinline__ void store_row_chunk(const GemmProblem& P, int m, int n0, const float* v) {
  if (P.out != nullptr) {
    __nv_bfloat16* row = P.out + static_cast<long long>(m) * P.ldo;
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const int nb = n0 + 8 * h;
      if (nb + 8 <= P.ldo) {
        store_bf16x8(row + nb, v + 8 * h);  // pad columns (n >= N) receive the zeros prepared by the caller
      } else {
#pragma unroll
        for (int j = 0; j < 8; ++j)
          if (nb + j < P.N) row[nb + j] = __float2bfloat16(v[8 * h + j]);
      }
    }
  }
  if (P.outT != nullptr) {
#pragma unroll
    for (int j = 0; j < 16; ++j)
      if (n0 + j < P.N) P.outT[static_cast<long long>(n0 + j) * P.ldt + m] = __float2bfloat16(v[j]);
  }
}

// Everything the four epilogue warps do with a finished accumulator tile (rows m0.., columns nbase.. of problem P):
// shared by the one-tile-per-CTA kernel and by the chain kernel, which walks a list of dependent problems.
struct EpiCtx {
  uint32_t seed_lo, seed_hi;
  const unsigned long long* step_ptr;
  uint8_t* smem;  // idle pipeline stages, reused as a staging patch by the wgrad epilogue
  // TTD_EPI_WGRAD_UPDATE (single replica: the optimiser step happens where the gradient is produced)
  const float* lr_slot = nullptr;  // device float[2]: learning rate and EMA decay of this step
  float grad_scale = 1.0f, reg_scale = 0.0f;
  int reg_kind = 0;
  bool use_ema = false;
};

// What a forward / dgrad chunk (one row x 16 columns) needs besides the accumulator: bias or saved activation,
// and the dropout bits. None of it depends on the GEMM, so it is fetched while TMA and the tensor core are
// still busy (chunk 0) or while the previous chunk is being finished (later chunks).
struct EpiSide {
  float side[16];   // FWD: bias; DGRAD: saved (post-dropout) activation of the layer being differentiated
  Philox4 rnd[2];   // 16 random bits per column, eight columns per block
};

__device__ __forceinline__ void epi_fd_load(const GemmProblem& P, const EpiCtx& E, int m, int n0,
                                            unsigned long long step, bool row_ok, EpiSide& S) {
  if (!row_ok || n0 >= P.N) return;
  const bool full = n0 + 16 <= P.N;
  if (P.mode == TTD_EPI_FWD) {
    if (P.bias == nullptr) {
#pragma unroll
      for (int j = 0; j < 16; ++j) S.side[j] = 0.0f;
    } else if (full && (reinterpret_cast<uintptr_t>(P.bias + n0) & 15u) == 0) {
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const float4 t = *reinterpret_cast<const float4*>(P.bias + n0 + 4 * q);
        S.side[4 * q] = t.x; S.side[4 * q + 1] = t.y; S.side[4 * q + 2] = t.z; S.side[4 * q + 3] = t.w;
      }
    } else {
#pragma unroll
      for (int j = 0; j < 16; ++j) S.side[j] = (n0 + j < P.N) ? P.bias[n0 + j] : 0.0f;
    }
  } else {
    if (P.aux == nullptr) {
#pragma unroll
      for (int j = 0; j < 16; ++j) S.side[j] = 1.0f;
    } else {
      const __nv_bfloat16* arow = P.aux + static_cast<long long>(m) * P.ldaux + n0;
      if (full && (reinterpret_cast<uintptr_t>(arow) & 15u) == 0) {
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          const uint4 u = *reinterpret_cast<const uint4*>(arow + 8 * h);
          const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            const __nv_bfloat162 t = *reinterpret_cast<const __nv_bfloat162*>(&w[q]);
            S.side[8 * h + 2 * q] = __low2float(t);
            S.side[8 * h + 2 * q + 1] = __high2float(t);
          }
        }
      } else {
#pragma unroll
        for (int j = 0; j < 16; ++j) S.side[j] = (n0 + j < P.N) ? __bfloat162float(arow[j]) : 0.0f;
      }
    }
  }
  if (P.drop_rate > 0.0f) {
    S.rnd[0] = dropout_block(static_cast<uint32_t>(m), static_cast<uint32_t>(n0 >> 3), step, P.layer_id, E.seed_lo,
                             E.seed_hi);
    S.rnd[1] = dropout_block(static_cast<uint32_t>(m), static_cast<uint32_t>((n0 >> 3) + 1), step, P.layer_id,
                             E.seed_lo, E.seed_hi);
  }
}

// The arithmetic of one chunk with mode, activation and dropout fixed at compile time (straight-line code).
template <int MODE, int ACT, bool DROP>
__device__ __forceinline__ void epi_fd_apply(const GemmProblem& P, int m, int n0, float* v, const EpiSide& S,
                                             float keep, float inv_keep, uint32_t thresh16) {
  const bool full = n0 + 16 <= P.N;
#pragma unroll
  for (int j = 0; j < 16; ++j) {
    const bool kept = !DROP || dropout_lane16(S.rnd[j >> 3], j & 7) >= thresh16;
    float r;
    if (MODE == TTD_EPI_FWD) {
      const float a = act_apply(v[j] + S.side[j], ACT);
      if (P.outf != nullptr && (full || n0 + j < P.N)) P.outf[static_cast<long long>(m) * P.ldf + n0 + j] = a;
      r = DROP ? (kept ? a * inv_keep : 0.0f) : a;
    } else {
      const float g = v[j] * act_grad_from_out(DROP ? S.side[j] * keep : S.side[j], ACT);
      r = DROP ? (kept ? g * inv_keep : 0.0f) : g;
    }
    v[j] = (full || n0 + j < P.N) ? r : 0.0f;
  }
  store_row_chunk(P, m, n0, v);
}

// Master weights and shadow values of one row x 16 columns of a variable group (fused-update epilogue).
__device__ __forceinline__ void upd_load(const GemmProblem& P, bool use_ema, int m, int n0, bool row_ok, float* w, float* s) {
#pragma unroll
  for (int j = 0; j < 16; ++j) w[j] = s[j] = 0.0f;
  if (!row_ok || n0 >= P.N) return;
  const bool full = n0 + 16 <= P.N;
  const float* wr = P.upd_w + static_cast<long long>(m) * P.ldf + n0;
  if (full && (reinterpret_cast<uintptr_t>(wr) & 15u) == 0) {
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const float4 t = *reinterpret_cast<const float4*>(wr + 4 * q);
      w[4 * q] = t.x; w[4 * q + 1] = t.y; w[4 * q + 2] = t.z; w[4 * q + 3] = t.w;
    }
  } else {
#pragma unroll
    for (int j = 0; j < 16; ++j)
      if (n0 + j < P.N) w[j] = wr[j];
  }
  if (!use_ema) return;
  const float* sr = P.upd_ema + static_cast<long long>(m) * P.ldf + n0;
  if (full && (reinterpret_cast<uintptr_t>(sr) & 15u) == 0) {
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const float4 t = *reinterpret_cast<const float4*>(sr + 4 * q);
      s[4 * q] = t.x; s[4 * q + 1] = t.y; s[4 * q + 2] = t.z; s[4 * q + 3] = t.w;
    }
  } else {
#pragma unroll
    for (int j = 0; j < 16; ++j)
      if (n0 + j < P.N) s[j] = sr[j];
  }
}
__device__ __forceinline__ void store_f32x16(float* dst, const float* v, int n_valid) {
  if (n_valid >= 16 && (reinterpret_cast<uintptr_t>(dst) & 15u) == 0) {
#pragma unroll
    for (int q = 0; q < 4; ++q)
      *reinterpret_cast<float4*>(dst + 4 * q) = make_float4(v[4 * q], v[4 * q + 1], v[4 * q + 2], v[4 * q + 3]);
  } else {
#pragma unroll
    for (int j = 0; j < 16; ++j)
      if (j < n_valid) dst[j] = v[j];
  }
}

__device__ __forceinline__ void run_epilogue(const GemmProblem& P, const EpiCtx& E, uint32_t tmem_base, int m0,
                                             int nbase, int warp, int lane, uint64_t* acc_bar, uint32_t acc_parity) {
  // ===================== epilogue: TMEM -> registers -> global =====================
  const int quad = warp & 3;  // TMEM lane quadrant this warp may read
  const int m = m0 + quad * 32 + lane;
  const bool row_ok = m < P.M;
  const unsigned long long step = E.step_ptr != nullptr ? *E.step_ptr : 0ull;
  const uint32_t taddr_row = tmem_base + (static_cast<uint32_t>(quad * 32) << 16);
  const int nchunks = P.block_n / 16;
  float v[16];
  // everything that does not need the accumulator happens BEFORE the wait on it
  const bool fd = (P.mode == TTD_EPI_FWD || P.mode == TTD_EPI_DGRAD) && !P.swap;
  EpiSide S0, S1;
  if (fd) epi_fd_load(P, E, m, nbase, step, row_ok, S0);
  if (P.mode == TTD_EPI_WGRAD_UPDATE) {
    // ===== weight gradient + optimiser step in one epilogue (single replica) ==================================
    // g = scale*acc*grad_scale (+ L1/L2); w -= lr*g; ema -= (1-d)(ema-w); bf16 copies in both layouts, fp32 bias
    // copy. Same arithmetic, in the same order, as param_kernel<kUpdate> (reference :150-152, :153-161, :50-55).
    const int rows = P.M - 1;  // kernel rows; row `rows` of the group is the bias
    float wv[16], sv[16];
    upd_load(P, E.use_ema, m, nbase, row_ok, wv, sv);  // master / shadow values: requested before the accumulator is ready
    const float lr = E.lr_slot[0], ema_d = E.lr_slot[1];
    mbar_wait(acc_bar, acc_parity);
    tc_fence_after();
    for (int c = 0; c < nchunks; ++c) {
      const int n0 = nbase + c * 16;
      __syncwarp();
      tmem_ld16(taddr_row + c * 16, v);
      float wn[16], sn[16];
      if (c + 1 < nchunks) upd_load(P, E.use_ema, m, n0 + 16, row_ok, wn, sn);
      if (row_ok && n0 < P.N) {
        float val[16], shadow[16];
#pragma unroll
        for (int j = 0; j < 16; ++j) {
          float g = v[j] * P.scale * E.grad_scale;
          if (P.upd_regularize && m < rows) {
            if (E.reg_kind == 2) g += E.reg_scale * wv[j];
            else if (E.reg_kind == 1) g += E.reg_scale * (wv[j] > 0.0f ? 1.0f : (wv[j] < 0.0f ? -1.0f : 0.0f));
          }
          val[j] = (n0 + j < P.N) ? wv[j] - lr * g : 0.0f;
          shadow[j] = sv[j] - (1.0f - ema_d) * (sv[j] - val[j]);
        }
        const int n_valid = P.N - n0;
        store_f32x16(P.upd_w + static_cast<long long>(m) * P.ldf + n0, val, n_valid);
        if (E.use_ema) store_f32x16(P.upd_ema + static_cast<long long>(m) * P.ldf + n0, shadow, n_valid);
        if (P.outf != nullptr) {  // the gradient itself, when somebody wants to look at it
          float gr[16];
#pragma unroll
          for (int j = 0; j < 16; ++j) gr[j] = v[j] * P.scale;
          store_f32x16(P.outf + static_cast<long long>(m) * P.ldf + n0, gr, n_valid);
        }
        if (m < rows) store_row_chunk(P, m, n0, val);  // bf16 compute copies: TF layout (out) and transposed (outT)
        else if (P.upd_bias != nullptr) store_f32x16(P.upd_bias + n0, val, n_valid);
      }
#pragma unroll
      for (int j = 0; j < 16; ++j) { wv[j] = wn[j]; sv[j] = sn[j]; }
    }
    return;
  }
  mbar_wait(acc_bar, acc_parity);
  tc_fence_after();

  if (P.mode == TTD_EPI_FWD_XENT && nchunks == 1) {
    // <= 16 classes (the reference's 10): the whole logit row lives in registers — one TMEM read, the bias and
    // the label fetched before the accumulator is ready would be better still, but they are L2 hits. Same
    // operation order as the general path below, hence the same bits.
    const long long label = row_ok ? P.labels[m] : 0;
    __syncwarp();
    tmem_ld16(taddr_row, v);
    float z[16];
    float zmax = -INFINITY, zlab = 0.0f;
#pragma unroll
    for (int j = 0; j < 16; ++j) {
      z[j] = 0.0f;
      if (j < P.N) {
        z[j] = v[j] + (P.bias ? P.bias[j] : 0.0f);
        zmax = fmaxf(zmax, z[j]);
        if (j == label) zlab = z[j];
      }
    }
    float sum = 0.0f;
    int greater = 0;
    float e[16];
#pragma unroll
    for (int j = 0; j < 16; ++j) {
      e[j] = 0.0f;
      if (j < P.N) {
        e[j] = __expf(z[j] - zmax);
        sum += e[j];
        greater += (z[j] > zlab) ? 1 : 0;
      }
    }
    const float inv_sum = 1.0f / sum;
    const float gscale = P.scale / static_cast<float>(P.M);
    float d[16];
#pragma unroll
    for (int j = 0; j < 16; ++j) {
      d[j] = 0.0f;
      if (j < P.N) {
        if (row_ok && P.outf) P.outf[static_cast<long long>(m) * P.ldf + j] = z[j];
        d[j] = (e[j] * inv_sum - (j == label ? 1.0f : 0.0f)) * gscale;
      }
    }
    if (row_ok) store_row_chunk(P, m, 0, d);
    float loss = row_ok ? (__logf(sum) + zmax - zlab) / static_cast<float>(P.M) : 0.0f;
    float corr = (row_ok && greater < 1) ? 1.0f / static_cast<float>(P.M) : 0.0f;
    loss = warp_sum(loss);
    corr = warp_sum(corr);
    if (lane == 0 && P.metrics) {
      atomicAdd(&P.metrics[0], loss);
      atomicAdd(&P.metrics[1], corr);
    }
  } else if (P.mode == TTD_EPI_FWD_XENT) {
    // One thread owns one example's whole logit row (N <= block_n). Three cheap passes over TMEM.
    const long long label = row_ok ? P.labels[m] : 0;
    float zmax = -INFINITY, zlab = 0.0f;
    for (int c = 0; c < nchunks; ++c) {
      __syncwarp();
      tmem_ld16(taddr_row + c * 16, v);
#pragma unroll
      for (int j = 0; j < 16; ++j) {
        const int n = c * 16 + j;
        if (n < P.N) {
          const float z = v[j] + (P.bias ? P.bias[n] : 0.0f);
          zmax = fmaxf(zmax, z);
          if (n == label) zlab = z;
        }
      }
    }
    float sum = 0.0f;
    int greater = 0;
    for (int c = 0; c < nchunks; ++c) {
      __syncwarp();
      tmem_ld16(taddr_row + c * 16, v);
#pragma unroll
      for (int j = 0; j < 16; ++j) {
        const int n = c * 16 + j;
        if (n < P.N) {
          const float z = v[j] + (P.bias ? P.bias[n] : 0.0f);
          sum += __expf(z - zmax);
          greater += (z > zlab) ? 1 : 0;
        }
      }
    }
    const float inv_sum = 1.0f / sum;
    const float gscale = P.scale / static_cast<float>(P.M);
    for (int c = 0; c < nchunks; ++c) {
      __syncwarp();
      tmem_ld16(taddr_row + c * 16, v);
      float d[16];
#pragma unroll
      for (int j = 0; j < 16; ++j) {
        const int n = c * 16 + j;
        float dz = 0.0f;
        if (n < P.N) {
          const float z = v[j] + (P.bias ? P.bias[n] : 0.0f);
          if (row_ok && P.outf) P.outf[static_cast<long long>(m) * P.ldf + n] = z;
          dz = (__expf(z - zmax) * inv_sum - (n == label ? 1.0f : 0.0f)) * gscale;
        }
        d[j] = dz;
      }
      if (row_ok) store_row_chunk(P, m, c * 16, d);
    }
    float loss = row_ok ? (__logf(sum) + zmax - zlab) / static_cast<float>(P.M) : 0.0f;
    float corr = (row_ok && greater < 1) ? 1.0f / static_cast<float>(P.M) : 0.0f;
    loss = warp_sum(loss);
    corr = warp_sum(corr);
    if (lane == 0 && P.metrics) {
      atomicAdd(&P.metrics[0], loss);
      atomicAdd(&P.metrics[1], corr);
    }
  } else {
    const float keep = 1.0f - P.drop_rate;
    const float inv_keep = 1.0f / keep;
    const bool wide_store = P.mode == TTD_EPI_WGRAD && (P.block_n % 32) == 0 && (P.N % 4) == 0 && (P.ldf % 4) == 0 &&
                            (reinterpret_cast<uintptr_t>(P.outf) & 15u) == 0;
    const uint32_t drop_thresh = dropout_thresh16(P.drop_rate);
    for (int c = 0; c < nchunks; ++c) {
      const int n0 = nbase + c * 16;
      __syncwarp();                      // tcgen05.ld is warp-collective: reconverge before it
      tmem_ld16(taddr_row + c * 16, v);
      const bool live = row_ok && n0 < P.N;
      if (P.mode == TTD_EPI_WGRAD && wide_store) {
        // 32 rows x 32 fp32 columns per warp go through a private smem patch (the pipeline stages are idle
        // by now) so that every global store instruction writes whole 128-byte lines of four rows instead
        // of 16 bytes of thirty-two rows
        float v2[16];
        __syncwarp();
        tmem_ld16(taddr_row + c * 16 + 16, v2);
        float* patch = reinterpret_cast<float*>(E.smem) + (warp - 4) * (32 * 36);
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          *reinterpret_cast<float4*>(patch + lane * 36 + 4 * q) =
              make_float4(v[4 * q] * P.scale, v[4 * q + 1] * P.scale, v[4 * q + 2] * P.scale, v[4 * q + 3] * P.scale);
          *reinterpret_cast<float4*>(patch + lane * 36 + 16 + 4 * q) =
              make_float4(v2[4 * q] * P.scale, v2[4 * q + 1] * P.scale, v2[4 * q + 2] * P.scale, v2[4 * q + 3] * P.scale);
        }
        __syncwarp();
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const int rl = i * 4 + (lane >> 3), c4 = (lane & 7) * 4;
          const int mm = m0 + quad * 32 + rl, nn = n0 + c4;
          if (mm < P.M && nn < P.N)  // N % 4 == 0 in this mode, so a float4 never straddles the edge
            *reinterpret_cast<float4*>(P.outf + static_cast<long long>(mm) * P.ldf + nn) =
                *reinterpret_cast<const float4*>(patch + rl * 36 + c4);
        }
        ++c;  // two 16-column chunks consumed
      } else if (live && P.mode == TTD_EPI_WGRAD) {
        float* row = P.outf + static_cast<long long>(m) * P.ldf;
        const bool vec = ((reinterpret_cast<uintptr_t>(row + n0) & 15u) == 0) && (n0 + 16 <= P.N);
        if (vec) {
#pragma unroll
          for (int q = 0; q < 4; ++q)
            *reinterpret_cast<float4*>(row + n0 + 4 * q) =
                make_float4(v[4 * q] * P.scale, v[4 * q + 1] * P.scale, v[4 * q + 2] * P.scale,
                            v[4 * q + 3] * P.scale);
        } else {
#pragma unroll
          for (int j = 0; j < 16; ++j)
            if (n0 + j < P.N) row[n0 + j] = v[j] * P.scale;
        }
      } else if (live && P.swap) {
        // Weight-stationary orientation: this thread owns output feature f = m and the 16 columns are
        // consecutive batch rows. All pointers keep their logical [batch, feature] meaning.
        const int f = m;
        const float bias_f = (P.mode == TTD_EPI_FWD && P.bias) ? P.bias[f] : 0.0f;
#pragma unroll
        for (int j = 0; j < 16; ++j) {
          const int b = n0 + j;
          float r = 0.0f;
          if (b < P.N) {
            bool kept = true;
            if (P.drop_rate > 0.0f) {
              kept = dropout_kept(static_cast<uint32_t>(b), static_cast<uint32_t>(f), step, P.layer_id, E.seed_lo,
                                  E.seed_hi, drop_thresh);
            }
            if (P.mode == TTD_EPI_FWD) {
              const float a = act_apply(v[j] + bias_f, P.act);
              if (P.outf) P.outf[static_cast<long long>(b) * P.ldf + f] = a;
              r = kept ? a * inv_keep : 0.0f;
            } else {  // DGRAD
              const float a_sv = P.aux ? __bfloat162float(P.aux[static_cast<long long>(b) * P.ldaux + f]) : 1.0f;
              r = kept ? v[j] * inv_keep * act_grad_from_out(a_sv * keep, P.act) : 0.0f;
            }
          }
          v[j] = r;
        }
        if (P.out != nullptr) {  // [batch, feature]: lanes are consecutive features -> coalesced
#pragma unroll
          for (int j = 0; j < 16; ++j)
            if (n0 + j < P.N) P.out[static_cast<long long>(n0 + j) * P.ldo + f] = __float2bfloat16(v[j]);
        }
        if (P.outT != nullptr) {  // [feature, batch]: 16 consecutive batch entries of this thread's row
          __nv_bfloat16* row = P.outT + static_cast<long long>(f) * P.ldt;
#pragma unroll
          for (int h = 0; h < 2; ++h) {
            const int nb = n0 + 8 * h;
            if (nb + 8 <= P.ldt) {
              store_bf16x8(row + nb, v + 8 * h);
            } else {
#pragma unroll
              for (int j = 0; j < 8; ++j)
                if (nb + j < P.N) row[nb + j] = __float2bfloat16(v[8 * h + j]);
            }
          }
        }
      } else if (live) {
        // straight-line code per (mode, activation, dropout); the next chunk's side data is requested first so
        // that its latency hides behind this chunk's arithmetic and stores
        if (c + 1 < nchunks) epi_fd_load(P, E, m, n0 + 16, step, row_ok, S1);
        const int key = (P.mode == TTD_EPI_DGRAD ? 6 : 0) + P.act * 2 + (P.drop_rate > 0.0f ? 1 : 0);
        switch (key) {
          case 0: epi_fd_apply<TTD_EPI_FWD, ACT_NONE, false>(P, m, n0, v, S0, keep, inv_keep, drop_thresh); break;
          case 1: epi_fd_apply<TTD_EPI_FWD, ACT_NONE, true>(P, m, n0, v, S0, keep, inv_keep, drop_thresh); break;
          case 2: epi_fd_apply<TTD_EPI_FWD, ACT_RELU, false>(P, m, n0, v, S0, keep, inv_keep, drop_thresh); break;
          case 3: epi_fd_apply<TTD_EPI_FWD, ACT_RELU, true>(P, m, n0, v, S0, keep, inv_keep, drop_thresh); break;
          case 4: epi_fd_apply<TTD_EPI_FWD, ACT_ELU, false>(P, m, n0, v, S0, keep, inv_keep, drop_thresh); break;
          case 5: epi_fd_apply<TTD_EPI_FWD, ACT_ELU, true>(P, m, n0, v, S0, keep, inv_keep, drop_thresh); break;
          case 6: epi_fd_apply<TTD_EPI_DGRAD, ACT_NONE, false>(P, m, n0, v, S0, keep, inv_keep, drop_thresh); break;
          case 7: epi_fd_apply<TTD_EPI_DGRAD, ACT_NONE, true>(P, m, n0, v, S0, keep, inv_keep, drop_thresh); break;
          case 8: epi_fd_apply<TTD_EPI_DGRAD, ACT_RELU, false>(P, m, n0, v, S0, keep, inv_keep, drop_thresh); break;
          case 9: epi_fd_apply<TTD_EPI_DGRAD, ACT_RELU, true>(P, m, n0, v, S0, keep, inv_keep, drop_thresh); break;
          case 10: epi_fd_apply<TTD_EPI_DGRAD, ACT_ELU, false>(P, m, n0, v, S0, keep, inv_keep, drop_thresh); break;
          default: epi_fd_apply<TTD_EPI_DGRAD, ACT_ELU, true>(P, m, n0, v, S0, keep, inv_keep, drop_thresh); break;
        }
        S0 = S1;
      }
    }
  }
}

__global__ void __launch_bounds__(kThreads, 1) gemm_group_kernel(const __grid_constant__ GemmGroup g) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int stages = g.stages;
  const int kspan = g.kspan;
  const int a_stage = kAStageBytes * kspan;  // a stage holds kspan k-blocks of each operand, back to back
  uint8_t* sA = smem;
  uint8_t* sB = smem + stages * a_stage;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(sB + stages * g.b_stage_bytes);
  uint64_t* empty_bar = full_bar + stages;
  uint64_t* tmem_full_bar = empty_bar + stages;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_full_bar + 1);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  // which problem / tile does this CTA own?
  int pi = 0;
#pragma unroll
  for (int i = 1; i < kMaxProblems; ++i)
    if (i < g.count && static_cast<int>(blockIdx.x) >= g.p[i].tile_begin) pi = i;
  const GemmProblem& P = g.p[pi];
  const int tile = blockIdx.x - P.tile_begin;
  const int cl = g.cluster;
  uint32_t cta_rank = 0;
  if (cl > 1) asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(cta_rank));
  const uint16_t cl_mask = static_cast<uint16_t>((1u << cl) - 1u);
  // cluster mode walks N fastest so that the CTAs of one cluster sit on the same rows of A
  const int tile_m = cl > 1 ? tile / P.tiles_n : tile % P.tiles_m;
  const int tile_n = cl > 1 ? tile % P.tiles_n : tile / P.tiles_m;
  const int m0 = tile_m * kBlockM;
  const int nbase = tile_n * P.block_n;
  const int num_kb = ((P.K + kBlockK - 1) / kBlockK + kspan - 1) / kspan;  // pipeline iterations

  const bool pull_fused = (pi == 0 && g.n_b_peers > 0);
  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&P.tma_a);
    if (!pull_fused) tma_prefetch_desc(&P.tma_b);
  }
  if (warp == 3 && pull_fused && lane < g.n_b_peers) tma_prefetch_desc(&g.tma_b_peer[lane]);
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < stages; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], static_cast<uint32_t>(cl));  // every CTA that received the multicast frees the slot
    }
    mbar_init(tmem_full_bar, 1);
    fence_mbar_init();
  }
  if (warp == 2) {
    tmem_alloc(tmem_slot, static_cast<uint32_t>(g.tmem_cols));
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  if (cl > 1) {  // peers' barriers must exist before anyone multicasts into them
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
  }
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  // Everything above overlaps the previous kernel's tail under programmatic dependent launch;
  // nothing below may touch global memory before the dependency resolves.
  if (g.use_pdl) {
    asm volatile("griddepcontrol.wait;" ::: "memory");
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  }

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (lane == 0) {
      const int b_kb_bytes = P.block_n * kBlockK * 2;
      const uint32_t tx_bytes = static_cast<uint32_t>(kspan * (kAStageBytes + b_kb_bytes));
      int s = 0;
      uint32_t ph = 0;
      int owner = 0;
      for (int it = 0; it < num_kb; ++it) {
        mbar_wait(&empty_bar[s], ph ^ 1u);
        mbar_arrive_expect_tx(&full_bar[s], tx_bytes);
        for (int j = 0; j < kspan; ++j) {
          const int kb = it * kspan + j;  // beyond K the box is out of bounds and arrives as zeros
          const CUtensorMap* map_b = &P.tma_b;
          if (pull_fused) {
            while (owner + 1 < g.n_b_peers && kb >= g.kb_bound[owner + 1]) ++owner;
            map_b = &g.tma_b_peer[owner];  // remote HBM over NVLink, same coordinates as the local copy
          }
          uint8_t* dst_a = sA + s * a_stage + j * kAStageBytes;
          if (cl > 1) {
            // this CTA fetches 1/cl of the A tile and the TMA unit delivers it to every CTA of the cluster:
            // A crosses the L2->SM fabric once per cluster instead of once per CTA
            const int rows = kBlockM / cl;
            tma_load_2d_mcast(dst_a + cta_rank * rows * (kBlockK * 2), &P.tma_a, &full_bar[s], kb * kBlockK,
                              m0 + static_cast<int>(cta_rank) * rows, cl_mask);
          } else {
            tma_load_2d(dst_a, &P.tma_a, &full_bar[s], kb * kBlockK, m0);
          }
          tma_load_2d(sB + s * g.b_stage_bytes + j * b_kb_bytes, map_b, &full_bar[s], kb * kBlockK, nbase);
        }
        if (++s == stages) { s = 0; ph ^= 1u; }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer (one thread) =====================
    if (lane == 0) {
      const uint32_t idesc = umma_idesc_bf16(kBlockM, static_cast<uint32_t>(P.block_n));
      int s = 0;
      uint32_t ph = 0;
      for (int kb = 0; kb < num_kb; ++kb) {
        mbar_wait(&full_bar[s], ph);
        tc_fence_after();
        const uint32_t a_addr = smem_u32(sA + s * a_stage);
        const uint32_t b_addr = smem_u32(sB + s * g.b_stage_bytes);
        const uint32_t b_kb_bytes = static_cast<uint32_t>(P.block_n * kBlockK * 2);
        for (int j = 0; j < kspan; ++j) {
#pragma unroll
          for (int k = 0; k < kBlockK / 16; ++k) {
            const uint64_t da = umma_desc_k_sw128(a_addr + j * kAStageBytes + k * 32);
            const uint64_t db = umma_desc_k_sw128(b_addr + j * b_kb_bytes + k * 32);
            umma_bf16_ss(tmem_base, da, db, idesc, (kb | j | k) != 0 ? 1u : 0u);
          }
        }
        if (cl > 1) umma_commit_mcast(&empty_bar[s], cl_mask);  // tell every producer that feeds this slot
        else umma_commit(&empty_bar[s]);                       // smem slot reusable once these MMAs have drained
        if (++s == stages) { s = 0; ph ^= 1u; }
      }
      umma_commit(tmem_full_bar);  // accumulator complete
    }
  } else if (warp >= 4) {
    // ===================== epilogue: TMEM -> registers -> global =====================
    const EpiCtx E{g.seed_lo, g.seed_hi, g.step_ptr, smem};
    run_epilogue(P, E, tmem_base, m0, nbase, warp, lane, tmem_full_bar, 0);
  }

  tc_fence_before();
  __syncthreads();
  if (cl > 1) {  // nobody leaves while a peer may still signal one of its barriers
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
  }
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc(tmem_base, static_cast<uint32_t>(g.tmem_cols));
  }
}

// ------------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (fn) return fn;
  void* p = nullptr;
  cudaDriverEntryPointQueryResult qres;
  cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres);
  if (e != cudaSuccess || qres != cudaDriverEntryPointSuccess || p == nullptr) {
    set_error("cuTensorMapEncodeTiled not available from the driver (%s)", cudaGetErrorString(e));
    return nullptr;
  }
  fn = reinterpret_cast<EncodeTiledFn>(p);
  return fn;
}

// bf16 [rows, ld] matrix with `cols` valid elements per row; box = {64, box_rows}, 128-B swizzle,
// out-of-bounds elements read as zero (this is what pads K=784 / N=200 / M=785 to tile multiples).
int make_tmap(CUtensorMap* map, const void* base, int64_t rows, int64_t cols, int64_t ld, int box_rows) {
  EncodeTiledFn fn = get_encode_fn();
  if (!fn) return -2;
  if ((reinterpret_cast<uintptr_t>(base) & 15u) != 0 || (ld % 8) != 0) {
    set_error("TMA operand must be 16-byte aligned with ld %% 8 == 0 (base=%p ld=%lld)", base, (long long)ld);
    return -3;
  }
  cuuint64_t gdim[2] = {static_cast<cuuint64_t>(cols), static_cast<cuuint64_t>(rows)};
  cuuint64_t gstride[1] = {static_cast<cuuint64_t>(ld) * 2};
  cuuint32_t box[2] = {static_cast<cuuint32_t>(kBlockK), static_cast<cuuint32_t>(box_rows)};
  cuuint32_t estr[2] = {1, 1};
  static int promo = -1;
  if (promo < 0) {
    const char* e = getenv("TTD_GEMM_L2PROMO");  // tuning knob: 0 none, 1 64B, 2 128B, 3 256B (default)
    promo = e ? atoi(e) : 3;
  }
  const CUtensorMapL2promotion l2p = promo == 0   ? CU_TENSOR_MAP_L2_PROMOTION_NONE
                                     : promo == 1 ? CU_TENSOR_MAP_L2_PROMOTION_L2_64B
                                     : promo == 2 ? CU_TENSOR_MAP_L2_PROMOTION_L2_128B
                                                  : CU_TENSOR_MAP_L2_PROMOTION_L2_256B;
  CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), gdim, gstride, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, l2p,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled failed: CUresult %d (rows=%lld cols=%lld ld=%lld box_rows=%d)", (int)r,
              (long long)rows, (long long)cols, (long long)ld, box_rows);
    return -4;
  }
  return 0;
}

static int choose_block_n(int M, int N) {
  // Wide layers: keep >= ~1 wave of CTAs streaming the weight matrix; small layers: one tile.
  const int tiles_m = (M + kBlockM - 1) / kBlockM;
  int bn = 256;
  while (bn > 16 && tiles_m * ((N + bn - 1) / bn) < 120) bn >>= 1;
  const int n16 = ((N + 15) / 16) * 16;
  if (bn > n16) bn = n16;
  if (bn < 16) bn = 16;
  return bn;
}

extern std::atomic<uint64_t> g_launch_count;

}  // namespace ttd

using namespace ttd;

extern "C" int ttd_gemm_group(const TtdGemmDesc* d, int count, uint64_t seed, const uint64_t* step_ptr,
                              uint32_t flags, void* stream) {
  if (count < 1 || count > kMaxProblems) {
    set_error("ttd_gemm_group: count must be in [1,%d], got %d", kMaxProblems, count);
    return -1;
  }
  GemmGroup g;
  memset(&g, 0, sizeof(g));
  g.count = count;
  g.seed_lo = static_cast<uint32_t>(seed);
  g.seed_hi = static_cast<uint32_t>(seed >> 32);
  g.step_ptr = reinterpret_cast<const unsigned long long*>(step_ptr);
  g.use_pdl = (flags & 1u) ? 1 : 0;
  int total_tiles = 0, max_bn = 16;
  for (int i = 0; i < count; ++i) {
    TtdGemmDesc s = d[i];  // local copy: the operands may be exchanged below
    GemmProblem& P = g.p[i];
    if (s.M < 1 || s.N < 1 || s.K < 1) {
      set_error("ttd_gemm_group: problem %d has empty shape %dx%dx%d", i, s.M, s.N, s.K);
      return -1;
    }
    // Skinny-batch forward / dgrad over a wide layer: exchange the operands so that every CTA streams a
    // 128-row slab of the WEIGHTS (the expensive operand) and only a 64-row slice of the activations shares
    // its pipeline stages: twice the weight bytes in flight per SM, which is what a latency-bound stream needs.
    const bool swap = (s.mode == TTD_EPI_FWD || s.mode == TTD_EPI_DGRAD) && s.M <= 256 && s.N >= 1024 &&
                      !(flags & 4u) && !(i == 0 && s.n_b_peers > 0);
    if (swap) {  // out / outT / outf / aux / bias keep their logical [batch, feature] meaning (see the epilogue)
      const TtdGemmDesc o = s;
      s.a = o.b; s.lda = o.ldb;
      s.b = o.a; s.ldb = o.lda;
      s.M = o.N; s.N = o.M;
      s.block_n = o.M >= 64 ? 64 : ((o.M + 15) / 16) * 16;
    }
    int bn = s.block_n > 0 ? s.block_n : choose_block_n(s.M, s.N);
    if (bn % 16 != 0 || bn < 16 || bn > 256) {
      set_error("ttd_gemm_group: block_n must be a multiple of 16 in [16,256], got %d", bn);
      return -1;
    }
    if (s.mode == TTD_EPI_FWD_XENT) {
      bn = ((s.N + 15) / 16) * 16;
      if (bn > 256 || s.labels == nullptr) {
        set_error("ttd_gemm_group: fused xent needs N <= 256 and labels");
        return -1;
      }
    }
    if (s.mode == TTD_EPI_WGRAD && s.outf == nullptr) {
      set_error("ttd_gemm_group: wgrad needs outf");
      return -1;
    }
    if (s.mode == TTD_EPI_WGRAD_UPDATE) {
      set_error("ttd_gemm_group: the fused weight-gradient + update epilogue is served by ttd_gemm_net_launch_ex");
      return -1;
    }
    if (s.mode == TTD_EPI_DGRAD && s.aux == nullptr && (s.act != TTD_ACT_NONE || s.drop_rate > 0.f)) {
      set_error("ttd_gemm_group: dgrad with activation/dropout needs aux");
      return -1;
    }
    if ((s.out && (s.ldo % 8 != 0 || (reinterpret_cast<uintptr_t>(s.out) & 15u))) ) {
      set_error("ttd_gemm_group: out must be 16-byte aligned with ldo %% 8 == 0");
      return -1;
    }
    // A shared by a cluster: skinny problems (one M tile) whose CTAs all stream the same activations
    int cluster = 1;
    if (count == 1 && !(flags & 2u) && s.M <= kBlockM && s.K >= 512) {
      const int tn = (s.N + bn - 1) / bn;
      if (tn >= 8 && tn % 4 == 0) cluster = 4;
      else if (tn >= 4 && tn % 2 == 0) cluster = 2;
      if (const char* e = getenv("TTD_GEMM_CLUSTER")) {  // tuning knob: force 1 / 2 / 4
        const int want = atoi(e);
        if (want == 1 || (want > 1 && tn % want == 0 && want <= 8)) cluster = want;
      }
    }
    g.cluster = cluster;
    int rc = make_tmap(&P.tma_a, s.a, s.M, s.K, s.lda, kBlockM / cluster);
    if (rc) return rc;
    rc = make_tmap(&P.tma_b, s.b, s.N, s.K, s.ldb, bn);
    if (rc) return rc;
    if (i == 0 && s.n_b_peers > 0) {
      if (s.n_b_peers > TTD_MAX_PEERS) {
        set_error("ttd_gemm_group: at most %d B-operand peers", TTD_MAX_PEERS);
        return -1;
      }
      g.n_b_peers = s.n_b_peers;
      for (int o = 0; o < s.n_b_peers; ++o) {
        if ((s.k_bound[o] % kBlockK != 0 && s.k_bound[o] != s.K) || s.k_bound[o + 1] < s.k_bound[o]) {
          set_error("ttd_gemm_group: owner boundaries along K must be ascending multiples of %d (got %d)", kBlockK,
                    s.k_bound[o]);
          return -1;
        }
        rc = make_tmap(&g.tma_b_peer[o], s.b_peer[o], s.N, s.K, s.ldb, bn);
        if (rc) return rc;
      }
      for (int o = 0; o <= s.n_b_peers; ++o) g.kb_bound[o] = (s.k_bound[o] + kBlockK - 1) / kBlockK;
    }
    P.M = s.M; P.N = s.N; P.K = s.K;
    P.block_n = bn;
    P.tiles_m = (s.M + kBlockM - 1) / kBlockM;
    P.tiles_n = (s.N + bn - 1) / bn;
    P.tile_begin = total_tiles;
    total_tiles += P.tiles_m * P.tiles_n;
    P.mode = s.mode; P.act = s.act; P.drop_rate = s.drop_rate; P.layer_id = s.layer_id;
    P.bias = s.bias;
    P.out = static_cast<__nv_bfloat16*>(s.out); P.ldo = s.ldo;
    P.outT = static_cast<__nv_bfloat16*>(s.outT); P.ldt = s.ldt;
    P.outf = s.outf; P.ldf = s.ldf;
    P.aux = static_cast<const __nv_bfloat16*>(s.aux); P.ldaux = s.ldaux;
    P.scale = s.scale;
    P.labels = reinterpret_cast<const long long*>(s.labels);
    P.metrics = s.metrics;
    P.swap = swap ? 1 : 0;
    if (bn > max_bn) max_bn = bn;
  }
  // long-K problems pack two k-blocks per stage (half the barrier round trips per byte streamed)
  int max_k = 0;
  for (int i = 0; i < count; ++i) max_k = d[i].K > max_k ? d[i].K : max_k;
  g.kspan = max_k >= 2048 ? 2 : 1;
  if (const char* e = getenv("TTD_GEMM_KSPAN")) g.kspan = atoi(e) >= 2 ? 2 : 1;
  g.b_stage_bytes = max_bn * kBlockK * 2 * g.kspan;
  if (g.b_stage_bytes < 1024) g.b_stage_bytes = 1024;
  const int stage_bytes = kAStageBytes * g.kspan + g.b_stage_bytes;
  int stages = (200 * 1024) / stage_bytes;
  if (stages > 8) stages = 8;
  if (const char* e = getenv("TTD_GEMM_STAGES")) {  // tuning knob
    const int want = atoi(e);
    if (want >= 1 && want < stages) stages = want;
  }
  if (stages < 1) stages = 1;
  g.stages = stages;
  int tmem_cols = 32;
  while (tmem_cols < max_bn) tmem_cols <<= 1;
  g.tmem_cols = tmem_cols;
  const size_t smem_bytes = static_cast<size_t>(stages) * stage_bytes + 1024 + 256;

  static bool attr_set[64] = {false};
  int dev = 0;
  TTD_CUDA(cudaGetDevice(&dev));
  if (dev >= 0 && dev < 64 && !attr_set[dev]) {
    TTD_CUDA(cudaFuncSetAttribute(gemm_group_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    attr_set[dev] = true;
  }
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = dim3(static_cast<unsigned>(total_tiles));
  cfg.blockDim = dim3(kThreads);
  cfg.dynamicSmemBytes = smem_bytes;
  cfg.stream = static_cast<cudaStream_t>(stream);
  cudaLaunchAttribute attr[2];
  int nattr = 0;
  if (g.cluster > 1) {
    attr[nattr].id = cudaLaunchAttributeClusterDimension;
    attr[nattr].val.clusterDim.x = static_cast<unsigned>(g.cluster);
    attr[nattr].val.clusterDim.y = 1;
    attr[nattr].val.clusterDim.z = 1;
    ++nattr;
  }
  if (g.use_pdl) {
    attr[nattr].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[nattr].val.programmaticStreamSerializationAllowed = 1;
    ++nattr;
  }
  cfg.attrs = attr;
  cfg.numAttrs = nattr;
  TTD_CUDA(cudaLaunchKernelEx(&cfg, gemm_group_kernel, g));
  g_launch_count.fetch_add(1);
  return 0;
}

// =====================================================================================================
// Chain kernel: a LIST of small, mutually dependent GEMM problems executed back to back by ONE CTA.
//
// The layers behind the first one of the reference MLP are eight dependent GEMMs (four forward, four dgrad)
// plus four weight gradients, each far smaller than one tcgen05 tile; as separate launches they cost ~6 us
// apiece, almost all of it launch / drain / prologue latency. Here the same tcgen05 pipeline (TMA -> smem ->
// tcgen05.mma -> TMEM -> fused epilogue) loops over the problem list inside one CTA: a stage boundary is a
// __syncthreads plus a generic->async proxy fence (the next stage's TMA reads what this stage's epilogue just
// stored), not a kernel boundary. Descriptors (with their tensor maps) live in a device buffer prepared once.
// =====================================================================================================
namespace ttd {

constexpr int kChainMaxKb = 4;      // K <= 256 per problem: every k-block of a stage is resident at once
constexpr int kChainMaxStages = 32;

struct ChainHeader {
  int n;
  uint32_t seed_lo, seed_hi;
  const unsigned long long* step_ptr;
  int use_pdl;
  int fence_mode;  // tuning/diagnostic knob, see the stage boundary
  unsigned long long* trace;  // optional: per stage {start, operands landed, MMAs done, epilogue done} in ns
};

__global__ void __launch_bounds__(kThreads, 1) gemm_chain_kernel(const GemmProblem* __restrict__ probs, ChainHeader h) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sA = smem;                                  // kChainMaxKb x 16 KiB
  uint8_t* sB = smem + kChainMaxKb * kAStageBytes;     // kChainMaxKb x (256 rows x 128 B)
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(sB + kChainMaxKb * 256 * kBlockK * 2);
  uint64_t* done_bar = full_bar + 1;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(done_bar + 1);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (warp == 1 && lane == 0) {
    mbar_init(full_bar, 1);
    mbar_init(done_bar, 1);
    fence_mbar_init();
  }
  if (warp == 2) {
    tmem_alloc(tmem_slot, 256);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  if (h.use_pdl) {
    asm volatile("griddepcontrol.wait;" ::: "memory");
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  }
  const EpiCtx E{h.seed_lo, h.seed_hi, h.step_ptr, smem};
  // The descriptor of the stage in flight is staged in shared memory: the epilogue consults it inside its
  // innermost loops, and global-memory loads there (which the compiler cannot hoist past the stores) would
  // cost more than the GEMM. Only the tensor maps are used in place (TMA wants them in global memory).
  __shared__ __align__(64) GemmProblem sP;
  uint32_t ph = 0;
  for (int s = 0; s < h.n; ++s) {
    {
      const uint32_t* src = reinterpret_cast<const uint32_t*>(&probs[s]);
      uint32_t* dst = reinterpret_cast<uint32_t*>(&sP);
      for (int i = threadIdx.x; i < static_cast<int>(sizeof(GemmProblem) / 4); i += kThreads) dst[i] = src[i];
    }
    __syncthreads();
    // ... and from there into registers: behind a reference to shared memory every field access in the
    // epilogue's unrolled loops is an LDS that cannot move across the neighbouring stores (measured: 5 us per
    // 16-column chunk instead of 0.4 us).
    GemmProblem P;
    P.M = sP.M; P.N = sP.N; P.K = sP.K; P.block_n = sP.block_n; P.tiles_m = sP.tiles_m; P.tiles_n = sP.tiles_n;
    P.tile_begin = 0; P.mode = sP.mode; P.act = sP.act; P.drop_rate = sP.drop_rate; P.layer_id = sP.layer_id;
    P.bias = sP.bias; P.out = sP.out; P.ldo = sP.ldo; P.outT = sP.outT; P.ldt = sP.ldt; P.outf = sP.outf;
    P.ldf = sP.ldf; P.aux = sP.aux; P.ldaux = sP.ldaux; P.scale = sP.scale; P.labels = sP.labels;
    P.metrics = sP.metrics; P.swap = 0;
    P.upd_w = nullptr; P.upd_ema = nullptr; P.upd_bias = nullptr; P.upd_regularize = 0;
    const CUtensorMap* map_a = &probs[s].tma_a;
    const CUtensorMap* map_b = &probs[s].tma_b;
    const int num_kb = (P.K + kBlockK - 1) / kBlockK;
    const int b_kb_bytes = P.block_n * kBlockK * 2;
    for (int tile = 0; tile < P.tiles_m * P.tiles_n; ++tile) {
      const int m0 = (tile % P.tiles_m) * kBlockM, nbase = (tile / P.tiles_m) * P.block_n;
      unsigned long long* tr = (h.trace != nullptr && tile == 0) ? h.trace + 4 * s : nullptr;
      if (warp == 0 && lane == 0) {
        if (tr) tr[0] = globaltimer_ns();
        mbar_arrive_expect_tx(full_bar, static_cast<uint32_t>(num_kb * (kAStageBytes + b_kb_bytes)));
        for (int kb = 0; kb < num_kb; ++kb) {
          tma_load_2d(sA + kb * kAStageBytes, map_a, full_bar, kb * kBlockK, m0);
          tma_load_2d(sB + kb * b_kb_bytes, map_b, full_bar, kb * kBlockK, nbase);
        }
      } else if (warp == 1 && lane == 0) {
        const uint32_t idesc = umma_idesc_bf16(kBlockM, static_cast<uint32_t>(P.block_n));
        mbar_wait(full_bar, ph);
        if (tr) tr[1] = globaltimer_ns();
        tc_fence_after();
        for (int kb = 0; kb < num_kb; ++kb) {
#pragma unroll
          for (int k = 0; k < kBlockK / 16; ++k) {
            const uint64_t da = umma_desc_k_sw128(smem_u32(sA + kb * kAStageBytes) + k * 32);
            const uint64_t db = umma_desc_k_sw128(smem_u32(sB + kb * b_kb_bytes) + k * 32);
            umma_bf16_ss(tmem_base, da, db, idesc, (kb | k) != 0 ? 1u : 0u);
          }
        }
        umma_commit(done_bar);
      } else if (warp >= 4) {
        if (tr && threadIdx.x == 128) tr[2] = globaltimer_ns();
        run_epilogue(P, E, tmem_base, m0, nbase, warp, lane, done_bar, ph);
        tc_fence_before();
        if (tr && threadIdx.x == 128) tr[3] = globaltimer_ns();
      }
      // what the epilogue stored through the generic proxy is the next stage's TMA (async proxy) input
      if (h.fence_mode == 0) {
        __threadfence();
        fence_proxy_async();
      } else if (h.fence_mode == 1) {
        asm volatile("fence.proxy.async.global;" ::: "memory");
      } else if (h.fence_mode == 3) {
        __threadfence();
      }  // 2: none (timing experiments only)
      __syncthreads();
      ph ^= 1u;
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 256);
  }
}

}  // namespace ttd

static unsigned long long* g_chain_trace = nullptr;
// Diagnostic: device buffer of 4 x n_stages uint64 that receives per-stage timestamps (NULL switches it off).
extern "C" void ttd_gemm_chain_set_trace(void* device_buf) { g_chain_trace = static_cast<unsigned long long*>(device_buf); }
extern "C" uint64_t ttd_gemm_chain_problem_bytes(void) { return sizeof(ttd::GemmProblem); }

// Encode `count` problems (executed in order) into `device_buf` (count * ttd_gemm_chain_problem_bytes() bytes,
// 64-byte aligned). Synchronous; do it once at set-up, not per step.
extern "C" int ttd_gemm_chain_build(const TtdGemmDesc* d, int count, void* device_buf, void* stream) {
  if (count < 1 || count > kChainMaxStages) {
    set_error("ttd_gemm_chain_build: 1..%d problems, got %d", kChainMaxStages, count);
    return -1;
  }
  if (reinterpret_cast<uintptr_t>(device_buf) & 63u) {
    set_error("ttd_gemm_chain_build: the descriptor buffer must be 64-byte aligned");
    return -1;
  }
  static thread_local GemmProblem host[kChainMaxStages];
  memset(host, 0, sizeof(GemmProblem) * count);
  for (int i = 0; i < count; ++i) {
    const TtdGemmDesc& s = d[i];
    GemmProblem& P = host[i];
    const int bn = ((s.N + 15) / 16) * 16;
    if (s.M < 1 || s.N < 1 || s.K < 1 || s.K > kChainMaxKb * kBlockK || bn > 256) {
      set_error("ttd_gemm_chain_build: problem %d (%dx%dx%d) needs K <= %d and N <= 256", i, s.M, s.N, s.K,
                kChainMaxKb * kBlockK);
      return -1;
    }
    if (s.mode == TTD_EPI_FWD_XENT && s.labels == nullptr) {
      set_error("ttd_gemm_chain_build: fused xent needs labels");
      return -1;
    }
    if (s.mode == TTD_EPI_WGRAD_UPDATE) {
      set_error("ttd_gemm_chain_build: the fused weight-gradient + update epilogue is served by ttd_gemm_net_launch_ex");
      return -1;
    }
    int rc = make_tmap(&P.tma_a, s.a, s.M, s.K, s.lda, kBlockM);
    if (rc) return rc;
    rc = make_tmap(&P.tma_b, s.b, s.N, s.K, s.ldb, bn);
    if (rc) return rc;
    P.M = s.M; P.N = s.N; P.K = s.K;
    P.block_n = bn;
    P.tiles_m = (s.M + kBlockM - 1) / kBlockM;
    P.tiles_n = 1;
    P.tile_begin = 0;
    P.mode = s.mode; P.act = s.act; P.drop_rate = s.drop_rate; P.layer_id = s.layer_id;
    P.bias = s.bias;
    P.out = static_cast<__nv_bfloat16*>(s.out); P.ldo = s.ldo;
    P.outT = static_cast<__nv_bfloat16*>(s.outT); P.ldt = s.ldt;
    P.outf = s.outf; P.ldf = s.ldf;
    P.aux = static_cast<const __nv_bfloat16*>(s.aux); P.ldaux = s.ldaux;
    P.scale = s.scale;
    P.labels = reinterpret_cast<const long long*>(s.labels);
    P.metrics = s.metrics;
    P.swap = 0;
  }
  TTD_CUDA(cudaMemcpyAsync(device_buf, host, sizeof(GemmProblem) * count, cudaMemcpyHostToDevice,
                           static_cast<cudaStream_t>(stream)));
  TTD_CUDA(cudaStreamSynchronize(static_cast<cudaStream_t>(stream)));
  return 0;
}

extern "C" int ttd_gemm_chain_launch(const void* device_buf, int count, uint64_t seed, const uint64_t* step_ptr,
                                     uint32_t flags, void* stream) {
  if (count < 1 || count > kChainMaxStages) {
    set_error("ttd_gemm_chain_launch: 1..%d problems, got %d", kChainMaxStages, count);
    return -1;
  }
  static bool attr_set[64] = {false};
  int dev = 0;
  TTD_CUDA(cudaGetDevice(&dev));
  if (dev >= 0 && dev < 64 && !attr_set[dev]) {
    TTD_CUDA(cudaFuncSetAttribute(gemm_chain_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  kChainMaxKb * (kAStageBytes + 256 * kBlockK * 2) + 1024 + 256));
    attr_set[dev] = true;
  }
  ChainHeader h;
  h.n = count;
  h.seed_lo = static_cast<uint32_t>(seed);
  h.seed_hi = static_cast<uint32_t>(seed >> 32);
  h.step_ptr = reinterpret_cast<const unsigned long long*>(step_ptr);
  h.use_pdl = (flags & 1u) ? 1 : 0;
  h.fence_mode = 0;
  if (const char* e = getenv("TTD_CHAIN_FENCE")) h.fence_mode = atoi(e);
  h.trace = g_chain_trace;
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = dim3(1);
  cfg.blockDim = dim3(kThreads);
  cfg.dynamicSmemBytes = kChainMaxKb * (kAStageBytes + 256 * kBlockK * 2) + 1024 + 256;
  cfg.stream = static_cast<cudaStream_t>(stream);
  cudaLaunchAttribute attr[1];
  if (h.use_pdl) {
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
  }
  TTD_CUDA(cudaLaunchKernelEx(&cfg, gemm_chain_kernel, static_cast<const GemmProblem*>(device_buf), h));
  g_launch_count.fetch_add(1);
  return 0;
}

// =====================================================================================================
// Net kernel: the GEMMs of a whole training step (forward, data gradients, weight gradients) as ONE launch
// of co-resident CTAs that hand results to each other through global-memory arrival counters.
//
// The reference MLP's step is eleven dependent GEMM layers, each one or a few tcgen05 tiles; as separate
// launches every dependency costs a grid drain + launch (~6.8 us per layer with programmatic dependent launch).
// Here every tile is a task; a CTA runs its tasks in problem order; before it loads a task's operands it
// waits (ld.acquire.gpu) until the problems that produce them have counted all their tiles in, and after its
// epilogue has stored a tile it counts itself in (fence + red.release.gpu). For the reference model there are
// 114 tiles, i.e. one task per CTA on a 148-SM part: the step becomes a dataflow over SMs, and a dependency
// costs one L2 round trip instead of a kernel boundary. Deadlock-free because dependencies only point to
// earlier problems, every CTA visits its tasks in problem order, and the grid never exceeds the SM count
// (one CTA per SM by shared-memory size), so all CTAs are resident.
// =====================================================================================================
namespace ttd {

constexpr int kNetStages = 4;        // smem ring slots, one k-block each
constexpr int kNetMaxProblems = 24;
constexpr int kNetBStage = 256 * kBlockK * 2;          // 32 KiB: B k-block at block_n = 256
constexpr int kNetPatchBytes = 4 * 32 * 36 * 4;        // wgrad epilogue staging (see run_epilogue)
constexpr int kNetSmemBytes = kNetStages * (kAStageBytes + kNetBStage) + kNetPatchBytes + 1024 + 256;

struct NetHeader {
  int n_problems;
  int max_tasks;              // task slots per CTA in `tasks`
  const int2* tasks;          // [grid][max_tasks] {problem, tile}; problem < 0 ends a CTA's list
  unsigned int* counters;     // [kNetMaxProblems] tiles finished per problem, [kNetMaxProblems] = CTAs that left
  const uint32_t* dep_mask;   // [n_problems] bit d: problem d must be complete before this one starts
  const int* tiles;           // [n_problems] tiles per problem
  uint32_t seed_lo, seed_hi;
  const unsigned long long* step_ptr;
  int use_pdl;
  unsigned long long* trace;  // optional: per problem (tile 0) {deps met, operands landed, accumulator full, counted in}
  // fused optimiser step (TTD_EPI_WGRAD_UPDATE problems) and end-of-step bookkeeping by the last CTA to leave
  TtdHyper hyper;
  const float* lr_slot;
  TtdStepState* state;
  int finalize;               // bit0: the last CTA advances the step state; bit1: ... and global_step
};

__device__ __forceinline__ unsigned int ld_acquire_gpu_u32(const unsigned int* p) {
  unsigned int v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void red_release_gpu_add(unsigned int* p, unsigned int v) {
  asm volatile("red.release.gpu.global.add.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

__global__ void __launch_bounds__(kThreads, 1) gemm_net_kernel(const GemmProblem* __restrict__ probs, NetHeader h) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sA = smem;
  uint8_t* sB = smem + kNetStages * kAStageBytes;
  uint8_t* patch = sB + kNetStages * kNetBStage;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(patch + kNetPatchBytes);
  uint64_t* empty_bar = full_bar + kNetStages;
  uint64_t* acc_full = empty_bar + kNetStages;   // [2] accumulator ready for the epilogue
  uint64_t* acc_empty = acc_full + 2;            // [2] accumulator drained, MMA may overwrite
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_empty + 2);
  __shared__ __align__(64) GemmProblem sP[2];    // descriptor of the task whose epilogue is running (per accumulator)
  __shared__ int s_deps_met;                     // tasks of this CTA whose producers have all counted in
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int2* my_tasks = h.tasks + static_cast<long long>(blockIdx.x) * h.max_tasks;

  if (warp == 1 && lane == 0) {
    s_deps_met = 0;
    for (int s = 0; s < kNetStages; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(&acc_full[a], 1);
      mbar_init(&acc_empty[a], 1);
    }
    fence_mbar_init();
  }
  if (warp == 2) {
    tmem_alloc(tmem_slot, 512);  // two 256-column accumulators: the epilogue of task i overlaps the MMAs of task i+1
    tmem_relinquish();
  }
  if (warp == 0 && lane == 0 && my_tasks[0].x >= 0) {
    tma_prefetch_desc(&probs[my_tasks[0].x].tma_a);
    tma_prefetch_desc(&probs[my_tasks[0].x].tma_b);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  if (h.use_pdl) {
    asm volatile("griddepcontrol.wait;" ::: "memory");
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  }
  if (h.trace != nullptr && blockIdx.x == 0 && threadIdx.x == 0) h.trace[4 * kNetMaxProblems] = globaltimer_ns();

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (lane == 0) {
      int s = 0;
      uint32_t ph = 0;
      for (int j = 0; j < h.max_tasks; ++j) {
        const int2 t = my_tasks[j];
        if (t.x < 0) break;
        const GemmProblem* P = &probs[t.x];
        const int K = P->K, block_n = P->block_n, tiles_m = P->tiles_m;
        int m0 = (t.y % tiles_m) * kBlockM;
        const int nbase = (t.y / tiles_m) * block_n;
        if (P->a_slots > 1 && h.step_ptr != nullptr)  // device-resident data set: this step's batch (rows past M of the
          m0 += static_cast<int>(*h.step_ptr % static_cast<unsigned long long>(P->a_slots)) * P->a_slot_rows;  // tile are masked)
        // wait for the producers of this problem's operands
        uint32_t mask = h.dep_mask[t.x];
        if (mask != 0) {
          const uint64_t t0 = globaltimer_ns();
          while (mask != 0) {
            const int d = __ffs(mask) - 1;
            if (ld_acquire_gpu_u32(h.counters + d) >= static_cast<unsigned int>(h.tiles[d])) mask &= mask - 1;
            else if (globaltimer_ns() - t0 > 4000000000ull) __trap();  // a lost producer raises instead of hanging
          }
          // their epilogues stored through the generic proxy; our TMA reads through the async proxy
          asm volatile("fence.proxy.async.global;" ::: "memory");
        }
        // the epilogue reads producers' results too (saved activations of a data gradient), and prefetches
        // them before its accumulator is ready: it may not start on task j before this point
        asm volatile("st.release.cta.shared::cta.s32 [%0], %1;" ::"r"(smem_u32(&s_deps_met)), "r"(j + 1) : "memory");
        if (h.trace != nullptr && t.y == 0) h.trace[4 * t.x + 0] = globaltimer_ns();
        const int num_kb = (K + kBlockK - 1) / kBlockK;
        const uint32_t tx = static_cast<uint32_t>(kAStageBytes + block_n * kBlockK * 2);
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(&empty_bar[s], ph ^ 1u);
          mbar_arrive_expect_tx(&full_bar[s], tx);
          tma_load_2d(sA + s * kAStageBytes, &P->tma_a, &full_bar[s], kb * kBlockK, m0);
          tma_load_2d(sB + s * kNetBStage, &P->tma_b, &full_bar[s], kb * kBlockK, nbase);
          if (++s == kNetStages) { s = 0; ph ^= 1u; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer (one thread) =====================
    if (lane == 0) {
      int s = 0;
      uint32_t ph = 0;
      uint32_t acc_ph[2] = {0, 0};
      int acc = 0;
      for (int j = 0; j < h.max_tasks; ++j) {
        const int2 t = my_tasks[j];
        if (t.x < 0) break;
        const GemmProblem* P = &probs[t.x];
        const int K = P->K, block_n = P->block_n;
        const int num_kb = (K + kBlockK - 1) / kBlockK;
        const uint32_t idesc = umma_idesc_bf16(kBlockM, static_cast<uint32_t>(block_n));
        mbar_wait(&acc_empty[acc], acc_ph[acc] ^ 1u);  // the epilogue of the task two back has drained it
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + static_cast<uint32_t>(acc) * 256u;
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(&full_bar[s], ph);
          tc_fence_after();
          if (kb == 0 && h.trace != nullptr && t.y == 0) h.trace[4 * t.x + 1] = globaltimer_ns();
          const uint32_t a_addr = smem_u32(sA + s * kAStageBytes);
          const uint32_t b_addr = smem_u32(sB + s * kNetBStage);
#pragma unroll
          for (int k = 0; k < kBlockK / 16; ++k) {
            const uint64_t da = umma_desc_k_sw128(a_addr + k * 32);
            const uint64_t db = umma_desc_k_sw128(b_addr + k * 32);
            umma_bf16_ss(d_tmem, da, db, idesc, (kb | k) != 0 ? 1u : 0u);
          }
          umma_commit(&empty_bar[s]);
          if (++s == kNetStages) { s = 0; ph ^= 1u; }
        }
        umma_commit(&acc_full[acc]);
        acc_ph[acc] ^= 1u;
        acc ^= 1;
      }
    }
  } else if (warp >= 4) {
    // ===================== epilogue: TMEM -> registers -> global, then count the tile in =====================
    EpiCtx E{h.seed_lo, h.seed_hi, h.step_ptr, patch};
    E.lr_slot = h.lr_slot;
    E.grad_scale = h.hyper.grad_scale;
    E.reg_scale = h.hyper.reg_scale;
    E.reg_kind = h.hyper.reg_kind;
    E.use_ema = h.hyper.ema_decay > 0.0f;
    uint32_t full_ph[2] = {0, 0};
    int acc = 0;
    for (int j = 0; j < h.max_tasks; ++j) {
      const int2 t = my_tasks[j];
      if (t.x < 0) break;
      {  // descriptor -> shared memory -> registers (fields read inside the unrolled loops must not be memory operands)
        const uint32_t* src = reinterpret_cast<const uint32_t*>(&probs[t.x]);
        uint32_t* dst = reinterpret_cast<uint32_t*>(&sP[acc]);
        for (int i = threadIdx.x - 128; i < static_cast<int>(sizeof(GemmProblem) / 4); i += 128) dst[i] = src[i];
      }
      asm volatile("bar.sync 1, 128;" ::: "memory");
      const GemmProblem& S = sP[acc];
      GemmProblem P;
      P.M = S.M; P.N = S.N; P.K = S.K; P.block_n = S.block_n; P.tiles_m = S.tiles_m; P.tiles_n = S.tiles_n;
      P.tile_begin = 0; P.mode = S.mode; P.act = S.act; P.drop_rate = S.drop_rate; P.layer_id = S.layer_id;
      P.bias = S.bias; P.out = S.out; P.ldo = S.ldo; P.outT = S.outT; P.ldt = S.ldt; P.outf = S.outf;
      P.ldf = S.ldf; P.aux = S.aux; P.ldaux = S.ldaux; P.scale = S.scale; P.labels = S.labels;
      P.metrics = S.metrics; P.swap = 0;
      P.upd_w = S.upd_w; P.upd_ema = S.upd_ema; P.upd_bias = S.upd_bias; P.upd_regularize = S.upd_regularize;
      const int m0 = (t.y % P.tiles_m) * kBlockM, nbase = (t.y / P.tiles_m) * P.block_n;
      {
        int met;
        const uint64_t t0 = globaltimer_ns();
        do {
          asm volatile("ld.acquire.cta.shared::cta.s32 %0, [%1];" : "=r"(met) : "r"(smem_u32(&s_deps_met)) : "memory");
          if (met <= j && globaltimer_ns() - t0 > 4000000000ull) __trap();
        } while (met <= j);
      }
      run_epilogue(P, E, tmem_base + static_cast<uint32_t>(acc) * 256u, m0, nbase, warp, lane, &acc_full[acc],
                   full_ph[acc]);
      if (h.trace != nullptr && t.y == 0 && threadIdx.x == 128) h.trace[4 * t.x + 2] = globaltimer_ns();
      tc_fence_before();
      __threadfence();        // this thread's stores are visible GPU-wide ...
      fence_proxy_async();    // ... also to other SMs' TMA reads
      asm volatile("bar.sync 1, 128;" ::: "memory");
      if (threadIdx.x == 128) {
        mbar_arrive(&acc_empty[acc]);
        red_release_gpu_add(h.counters + t.x, 1u);
        if (h.trace != nullptr && t.y == 0) h.trace[4 * t.x + 3] = globaltimer_ns();
      }
      full_ph[acc] ^= 1u;
      acc ^= 1;
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
  // the last CTA to leave re-arms the counters for the next launch (nobody polls them any more)
  if (threadIdx.x == 0) {
    __threadfence();
    const unsigned int left = atomicAdd(h.counters + kNetMaxProblems, 1u);
    if (left == gridDim.x - 1) {
      for (int i = 0; i <= kNetMaxProblems; ++i) h.counters[i] = 0;
      __threadfence();
      if ((h.finalize & 1) && h.state != nullptr)  // this launch ends the training step
        finalize_step(h.state, h.hyper, h.lr_slot != nullptr ? h.lr_slot[0] : 0.0f, (h.finalize & 2) != 0);
      if (h.trace != nullptr) h.trace[4 * kNetMaxProblems + 1] = globaltimer_ns();
    }
  }
}

}  // namespace ttd

namespace {
struct NetLayout {
  size_t probs, tasks, dep, tiles, counters, total;
};
NetLayout net_layout(int count, int grid, int max_tasks) {
  auto up = [](size_t v) { return (v + 63) & ~size_t(63); };
  NetLayout L;
  L.probs = 0;
  L.tasks = up(static_cast<size_t>(count) * sizeof(ttd::GemmProblem));
  L.dep = up(L.tasks + static_cast<size_t>(grid) * max_tasks * sizeof(int2));
  L.tiles = up(L.dep + ttd::kNetMaxProblems * sizeof(uint32_t));
  L.counters = up(L.tiles + ttd::kNetMaxProblems * sizeof(int));
  L.total = up(L.counters + (ttd::kNetMaxProblems + 1) * sizeof(unsigned int));
  return L;
}
}  // namespace

extern "C" uint64_t ttd_gemm_net_bytes(int count, int grid, int max_tasks) {
  return net_layout(count, grid, max_tasks).total;
}

// Encode `count` problems (dep_mask[i] bit d: problem d < i must be complete first) and a task schedule for
// `grid` co-resident CTAs into `device_buf` (ttd_gemm_net_bytes bytes, 64-byte aligned, owned by the net from
// here on). Tile widths are chosen so that, if possible, every tile gets a CTA of its own. Synchronous; set-up only.
// Returns the number of CTAs to launch (> 0) or a negative error.
extern "C" int ttd_gemm_net_build(const TtdGemmDesc* d, int count, const uint32_t* dep_mask, int grid, int max_tasks,
                                  void* device_buf, void* stream) {
  if (count < 1 || count > kNetMaxProblems || grid < 1 || max_tasks < 1) {
    set_error("ttd_gemm_net_build: 1..%d problems on >= 1 CTAs, got %d on %d", kNetMaxProblems, count, grid);
    return -1;
  }
  if (reinterpret_cast<uintptr_t>(device_buf) & 63u) {
    set_error("ttd_gemm_net_build: the buffer must be 64-byte aligned");
    return -1;
  }
  int bn[kNetMaxProblems], tiles_m[kNetMaxProblems], tiles[kNetMaxProblems];
  for (int i = 0; i < count; ++i) {
    const TtdGemmDesc& s = d[i];
    if (s.M < 1 || s.N < 1 || s.K < 1 || s.n_b_peers > 0) {
      set_error("ttd_gemm_net_build: problem %d: bad shape or peer-fetched operand (not supported here)", i);
      return -1;
    }
    if (dep_mask[i] >> i) {
      set_error("ttd_gemm_net_build: problem %d may only depend on earlier problems", i);
      return -1;
    }
    if (s.mode == TTD_EPI_FWD_XENT && (s.labels == nullptr || s.N > 256)) {
      set_error("ttd_gemm_net_build: fused xent needs labels and N <= 256");
      return -1;
    }
    if (s.mode == TTD_EPI_WGRAD && s.outf == nullptr) {
      set_error("ttd_gemm_net_build: wgrad needs outf");
      return -1;
    }
    if (s.mode == TTD_EPI_WGRAD_UPDATE && (s.upd_w == nullptr || s.ldf < s.N)) {
      set_error("ttd_gemm_net_build: the fused update needs the master weights and their row pitch (ldf >= N)");
      return -1;
    }
    const int n16 = ((s.N + 15) / 16) * 16;
    bn[i] = s.mode == TTD_EPI_FWD_XENT ? n16 : (s.mode == TTD_EPI_WGRAD ? 32 : 16);
    if (bn[i] > 256) bn[i] = 256;
    tiles_m[i] = (s.M + kBlockM - 1) / kBlockM;
  }
  auto total_tiles = [&]() {
    int t = 0;
    for (int i = 0; i < count; ++i) {
      tiles[i] = tiles_m[i] * ((d[i].N + bn[i] - 1) / bn[i]);
      t += tiles[i];
    }
    return t;
  };
  // widen the most fragmented problem until every tile has its own CTA (or nothing can be widened any more)
  while (total_tiles() > grid) {
    int worst = -1;
    for (int i = 0; i < count; ++i)
      if (d[i].mode != TTD_EPI_FWD_XENT && bn[i] < 256 && bn[i] < ((d[i].N + 15) / 16) * 16 &&
          (worst < 0 || tiles[i] > tiles[worst]))
        worst = i;
    if (worst < 0) break;
    bn[worst] *= 2;
  }
  const int total = total_tiles();
  if (total > grid * max_tasks) {
    set_error("ttd_gemm_net_build: %d tiles do not fit %d CTAs x %d tasks", total, grid, max_tasks);
    return -1;
  }
  const int ctas = total < grid ? total : grid;
  const NetLayout L = net_layout(count, grid, max_tasks);
  std::vector<uint8_t> host(L.total, 0);
  GemmProblem* probs = reinterpret_cast<GemmProblem*>(host.data() + L.probs);
  int2* tasks = reinterpret_cast<int2*>(host.data() + L.tasks);
  for (int i = 0; i < grid * max_tasks; ++i) tasks[i] = make_int2(-1, 0);
  int g = 0;
  for (int i = 0; i < count; ++i) {
    const TtdGemmDesc& s = d[i];
    GemmProblem& P = probs[i];
    const int64_t a_rows = s.a_slots > 1 ? static_cast<int64_t>(s.a_slots) * s.a_slot_rows : s.M;
    int rc = make_tmap(&P.tma_a, s.a, a_rows, s.K, s.lda, kBlockM);
    if (rc) return rc;
    rc = make_tmap(&P.tma_b, s.b, s.N, s.K, s.ldb, bn[i]);
    if (rc) return rc;
    P.a_slot_rows = s.a_slot_rows; P.a_slots = s.a_slots;
    P.M = s.M; P.N = s.N; P.K = s.K;
    P.block_n = bn[i];
    P.tiles_m = tiles_m[i];
    P.tiles_n = (s.N + bn[i] - 1) / bn[i];
    P.tile_begin = 0;
    P.mode = s.mode; P.act = s.act; P.drop_rate = s.drop_rate; P.layer_id = s.layer_id;
    P.bias = s.bias;
    P.out = static_cast<__nv_bfloat16*>(s.out); P.ldo = s.ldo;
    P.outT = static_cast<__nv_bfloat16*>(s.outT); P.ldt = s.ldt;
    P.outf = s.outf; P.ldf = s.ldf;
    P.aux = static_cast<const __nv_bfloat16*>(s.aux); P.ldaux = s.ldaux;
    P.scale = s.scale;
    P.labels = reinterpret_cast<const long long*>(s.labels);
    P.metrics = s.metrics;
    P.swap = 0;
    P.upd_w = s.upd_w; P.upd_ema = s.upd_ema; P.upd_bias = s.upd_bias; P.upd_regularize = s.upd_regularize;
    for (int t = 0; t < tiles[i]; ++t, ++g) tasks[(g % ctas) * max_tasks + g / ctas] = make_int2(i, t);
  }
  memcpy(host.data() + L.dep, dep_mask, sizeof(uint32_t) * count);
  memcpy(host.data() + L.tiles, tiles, sizeof(int) * count);
  TTD_CUDA(cudaMemcpyAsync(device_buf, host.data(), L.total, cudaMemcpyHostToDevice, static_cast<cudaStream_t>(stream)));
  TTD_CUDA(cudaStreamSynchronize(static_cast<cudaStream_t>(stream)));
  return ctas;
}

static unsigned long long* g_net_trace = nullptr;
// Diagnostic: device buffer of 4 x n_problems uint64 timestamps (NULL switches it off).
extern "C" void ttd_gemm_net_set_trace(void* device_buf) { g_net_trace = static_cast<unsigned long long*>(device_buf); }

extern "C" int ttd_gemm_net_launch(const void* device_buf, int count, int grid, int max_tasks, int ctas, uint64_t seed,
                                   const uint64_t* step_ptr, uint32_t flags, void* stream) {
  return ttd_gemm_net_launch_ex(device_buf, count, grid, max_tasks, ctas, seed, step_ptr, flags, nullptr, nullptr, nullptr,
                                0, stream);
}

// `ctas` is what ttd_gemm_net_build returned; grid / max_tasks the values it was built with.
extern "C" int ttd_gemm_net_launch_ex(const void* device_buf, int count, int grid, int max_tasks, int ctas, uint64_t seed,
                                      const uint64_t* step_ptr, uint32_t flags, const TtdHyper* hyper,
                                      const float* lr_slot, TtdStepState* state, int32_t finalize, void* stream) {
  if (count < 1 || count > kNetMaxProblems || ctas < 1 || ctas > grid) {
    set_error("ttd_gemm_net_launch: bad problem / CTA count (%d, %d of %d)", count, ctas, grid);
    return -1;
  }
  int dev = 0;
  TTD_CUDA(cudaGetDevice(&dev));
  static bool attr_set[64] = {false};
  static int sm_count[64] = {0};
  if (dev >= 0 && dev < 64 && !attr_set[dev]) {
    TTD_CUDA(cudaFuncSetAttribute(gemm_net_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kNetSmemBytes));
    TTD_CUDA(cudaDeviceGetAttribute(&sm_count[dev], cudaDevAttrMultiProcessorCount, dev));
    attr_set[dev] = true;
  }
  if (dev >= 0 && dev < 64 && ctas > sm_count[dev]) {  // all CTAs must be resident: they wait for one another
    set_error("ttd_gemm_net_launch: %d CTAs exceed the %d SMs of this device", ctas, sm_count[dev]);
    return -1;
  }
  const NetLayout L = net_layout(count, grid, max_tasks);
  const uint8_t* base = static_cast<const uint8_t*>(device_buf);
  NetHeader h;
  memset(&h, 0, sizeof(h));
  h.n_problems = count;
  h.max_tasks = max_tasks;
  h.tasks = reinterpret_cast<const int2*>(base + L.tasks);
  h.dep_mask = reinterpret_cast<const uint32_t*>(base + L.dep);
  h.tiles = reinterpret_cast<const int*>(base + L.tiles);
  h.counters = reinterpret_cast<unsigned int*>(const_cast<uint8_t*>(base) + L.counters);
  h.seed_lo = static_cast<uint32_t>(seed);
  h.seed_hi = static_cast<uint32_t>(seed >> 32);
  h.step_ptr = reinterpret_cast<const unsigned long long*>(step_ptr);
  h.use_pdl = (flags & 1u) ? 1 : 0;
  h.trace = g_net_trace;
  if (hyper != nullptr) h.hyper = *hyper;
  h.lr_slot = lr_slot;
  h.state = state;
  h.finalize = finalize;
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = dim3(ctas);
  cfg.blockDim = dim3(kThreads);
  cfg.dynamicSmemBytes = kNetSmemBytes;
  cfg.stream = static_cast<cudaStream_t>(stream);
  cudaLaunchAttribute attr[1];
  if (h.use_pdl) {
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
  }
  TTD_CUDA(cudaLaunchKernelEx(&cfg, gemm_net_kernel, reinterpret_cast<const GemmProblem*>(base + L.probs), h));
  g_launch_count.fetch_add(1);
  return 0;
}
