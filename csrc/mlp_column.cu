// Column kernel: forward, loss and data gradients of a narrow MLP in ONE launch, activations resident in
// shared memory from the first layer to the last data gradient.
//
// Why. The reference MLP's step (distribute_training.py:39-63, :77-82, :152) is a chain of nine dependent
// GEMM stages, each far smaller than one tcgen05 tile. As separate launches (or as tiles of one launch that hand
// over through global memory) a stage costs ~5.5 us, of which the GEMM is a few hundred cycles: the rest is the
// epilogue of a whole 128-row tile on four warps plus a global-memory round trip per layer.
//
// How. The batch is cut into slices of 16 examples and a CTA owns one slice through ALL layers, so no stage
// ever needs another CTA's result:
//   * the weights are the 128-row ("M") operand of tcgen05.mma and the slice's activations the 16-column ("N")
//     operand: Z^T[out, 16] = W^T[out, in] . A^T[in, 16]. Weights of the layers behind the first stay resident
//     in shared memory (loaded once per launch by TMA); the first layer's weights are streamed through a small
//     ring, each CTA of the cluster fetching 1/cluster of every k-block and multicasting it to all;
//   * an epilogue thread owns one output feature (TMEM lane) and four examples (columns): bias, activation,
//     dropout, then it writes the bf16 results straight into the NEXT layer's operand buffer in shared memory, in
//     the 128-byte-swizzled K-major layout the UMMA descriptor expects (generic-proxy stores + fence.proxy.async)
//     — and, as a side output, into the transposed global buffers the weight-gradient GEMMs read afterwards;
//   * sixteen epilogue warps (four per TMEM lane quadrant) share every tile, so a stage's epilogue is ~40
//     instructions per thread; the dropout bits of all layers are drawn while the first layer's weights stream in;
//   * the softmax cross-entropy runs on the transposed logits (classes across the lanes of a warp: two
//     butterfly reductions per example);
//   * the data gradients walk back through the same buffers: act'(a) and the dropout mask come from the
//     activations and keep-bits still sitting in shared memory.
// One stage = wait(accumulator) -> tcgen05.ld -> ~40 instructions -> stores -> fence -> arrive; the MMA thread
// issues the next stage's MMAs as soon as sixteen warps have arrived.
#include <atomic>
#include <math.h>
#include <stdlib.h>
#include <string.h>

#include "../include/ttd_b200.h"
#include "common.cuh"

namespace ttd {

extern std::atomic<uint64_t> g_launch_count;

namespace {

constexpr int kColMaxLayers = TTD_COL_MAX_LAYERS;
constexpr int kColN = 16;             // examples per CTA = N of every MMA
constexpr int kEpiWarps = 16;         // four per TMEM lane quadrant
constexpr int kEpiThreads = kEpiWarps * 32;
constexpr int kWarpTma = 16, kWarpMma = 17, kWarpAux = 18;
constexpr int kColThreads = 19 * 32;
constexpr int kRing = 3;              // k-block slots of the first layer's weight stream
constexpr int kKb = 64;               // bf16 per k-block row (128 B = one swizzle row)
constexpr int kBTile = kColN * kKb * 2;   // 2 KiB: one k-block of a 16-row activation operand
constexpr int kSmemLimit = 227 * 1024;

struct ColLayer {
  int in, out;
  int k16_in, k16_out;       // MMA K steps (of 16) when reducing over in / over out
  int kb_in, kb_out;         // k-blocks of 64
  int mt_out, mt_in;         // 128-row tiles over out / over in
  int rows_out8, rows_in8;   // rows rounded up to 8 (smem rows of the resident weight tiles)
  uint32_t off_wf;           // forward weights W^T [rows_out8 x kb_in k-blocks]           (layers >= 1)
  uint32_t off_wd;           // data-gradient weights W [rows_in8 x kb_out k-blocks]        (layers >= 1, inside the ring region)
  uint32_t off_act;          // this layer's output activations, operand of the next layer [16 x kb_out]   (hidden layers)
  uint32_t off_dz;           // gradient w.r.t. this layer's pre-activation [16 x kb_out]    (layers >= 1)
  uint32_t off_keep;         // keep bytes [ceil(out/8)][16]: bit j of byte (g, n) = element (example n, feature 8g+j) kept
  uint32_t tm_fwd, tm_dg;    // TMEM columns of the forward / data-gradient accumulators
  uint32_t layer_id;
  int keep_task0;            // first RNG task of this layer
  const float* bias;
  __nv_bfloat16* aT; long long ld_aT;
  __nv_bfloat16* dzT; long long ld_dzT;
};

struct ColParams {
  CUtensorMap tm_x;
  CUtensorMap tm_w1[2];                 // first-layer weight slices: box rows (base+1)*8 and base*8
  CUtensorMap tm_wf[kColMaxLayers];
  CUtensorMap tm_wd[kColMaxLayers];
  ColLayer L[kColMaxLayers];
  int n_layers, B, act, cluster, pool_batches;
  float drop_rate;
  int w1_groups_base, w1_groups_extra;  // CTA r of a cluster fetches base (+1 if r < extra) 8-row groups of every k-block
  uint32_t off_ring, ring_stage_bytes;
  uint32_t off_x;
  uint32_t off_zero, zero_bytes;        // activation / gradient operand buffers, cleared at start (pad columns stay zero)
  uint32_t off_bars;
  uint32_t tmem_cols;
  int n_keep_tasks;
  uint32_t seed_lo, seed_hi;
  int use_pdl;
  const long long* labels_pool;
  float* metrics;
  const unsigned long long* step_ptr;
  const long long* global_step_ptr;
  float* lr_out;
  TtdHyper hyper;
  unsigned long long* trace;  // optional (CTA 0): [0] dependency resolved, [1+2t] accumulator of stage t ready,
                              // [2+2t] stage t's epilogue stored, [63] kernel end — globaltimer ns
};

// byte offset of element (example n, feature f) inside a [16 x K] K-major, 128-byte-swizzled operand buffer
__device__ __forceinline__ uint32_t bop_off(int n, int f) {
  return static_cast<uint32_t>((f >> 6) * kBTile + n * 128 + ((((f >> 3) & 7) ^ (n & 7)) << 4) + (f & 7) * 2);
}
__device__ __forceinline__ uint32_t pack2(float a, float b) {
  const __nv_bfloat162 t = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<const uint32_t*>(&t);
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}

__global__ void __launch_bounds__(kColThreads, 1) mlp_column_kernel(const __grid_constant__ ColParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int L = p.n_layers;
  const int cl = p.cluster;
  uint32_t cta_rank = 0;
  if (cl > 1) asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(cta_rank));
  const uint16_t cl_mask = static_cast<uint16_t>((1u << cl) - 1u);
  const int e0 = static_cast<int>(blockIdx.x) * kColN;

  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + p.off_bars);
  uint64_t* bar_x = bars + 0;
  uint64_t* acc_bar = bars + 1;     // accumulator of the current stage is complete        (MMA thread -> epilogue warps)
  uint64_t* ready_bar = bars + 2;   // operand of the next stage is written, TMEM drained  (16 epilogue warps -> MMA thread)
  uint64_t* l1_done = bars + 3;     // first layer's MMAs have finished reading the ring   (MMA thread -> TMA thread)
  uint64_t* full_bar = bars + 4;
  uint64_t* empty_bar = full_bar + kRing;
  uint64_t* bar_wf = empty_bar + kRing;
  uint64_t* bar_wd = bar_wf + kColMaxLayers;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bar_wd + kColMaxLayers);

  if (warp == kWarpMma && lane == 0) {
    mbar_init(bar_x, 1);
    mbar_init(acc_bar, 1);
    mbar_init(ready_bar, kEpiWarps);
    mbar_init(l1_done, 1);
    for (int s = 0; s < kRing; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], static_cast<uint32_t>(cl));
    }
    for (int l = 0; l < kColMaxLayers; ++l) {
      mbar_init(&bar_wf[l], 1);
      mbar_init(&bar_wd[l], 1);
    }
    fence_mbar_init();
  }
  if (warp == kWarpAux) {
    tmem_alloc(tmem_slot, p.tmem_cols);
    tmem_relinquish();
  }
  if (warp == kWarpTma && lane == 0) {
    tma_prefetch_desc(&p.tm_x);
    tma_prefetch_desc(&p.tm_w1[0]);
    tma_prefetch_desc(&p.tm_w1[1]);
    for (int l = 1; l < L; ++l) {
      tma_prefetch_desc(&p.tm_wf[l]);
      tma_prefetch_desc(&p.tm_wd[l]);
    }
  }
  if (warp < kEpiWarps) {  // pad columns of every activation / gradient operand must read as zero
    uint4* z = reinterpret_cast<uint4*>(smem + p.off_zero);
    for (uint32_t i = threadIdx.x; i < p.zero_bytes / 16; i += kEpiThreads) z[i] = make_uint4(0u, 0u, 0u, 0u);
  }
  tc_fence_before();
  __syncthreads();
  if (cl > 1) cluster_sync_all();  // peers' barriers must exist before anyone multicasts into them
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  // Everything above overlaps the previous kernel's tail; nothing below may touch global memory earlier.
  if (p.use_pdl) {
    asm volatile("griddepcontrol.wait;" ::: "memory");
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  }
  unsigned long long* tr = (p.trace != nullptr && blockIdx.x == 0 && threadIdx.x == 0) ? p.trace : nullptr;
  if (tr) tr[0] = globaltimer_ns();
  const unsigned long long step = p.step_ptr != nullptr ? *p.step_ptr : 0ull;
  const int slot = p.pool_batches > 1 ? static_cast<int>(step % static_cast<unsigned long long>(p.pool_batches)) : 0;
  const int n_stages = 2 * L - 1;  // L forward stages, L-1 data-gradient stages

  if (warp == kWarpTma) {
    // ===================== TMA producer =====================
    if (lane == 0) {
      const ColLayer& L0 = p.L[0];
      // this slice's input rows, all k-blocks at once
      mbar_arrive_expect_tx(bar_x, static_cast<uint32_t>(L0.kb_in * kBTile));
      for (int kb = 0; kb < L0.kb_in; ++kb)
        tma_load_2d(smem + p.off_x + kb * kBTile, &p.tm_x, bar_x, kb * kKb, slot * p.B + e0);
      // first layer's weight stream: my 8-row groups of every k-block, delivered to every CTA of the cluster
      const int my_groups = p.w1_groups_base + (static_cast<int>(cta_rank) < p.w1_groups_extra ? 1 : 0);
      const int my_row0 = 8 * (static_cast<int>(cta_rank) * p.w1_groups_base +
                               min(static_cast<int>(cta_rank), p.w1_groups_extra));
      const CUtensorMap* my_map = static_cast<int>(cta_rank) < p.w1_groups_extra ? &p.tm_w1[0] : &p.tm_w1[1];
      int s = 0;
      uint32_t ph = 0;
      bool resident_issued = false;
      for (int it = 0; it < L0.kb_in; ++it) {
        mbar_wait(&empty_bar[s], ph ^ 1u);
        mbar_arrive_expect_tx(&full_bar[s], p.ring_stage_bytes);  // bytes from ALL CTAs of the cluster
        if (my_groups > 0) {
          uint8_t* dst = smem + p.off_ring + s * p.ring_stage_bytes + my_row0 * 128;
          if (cl > 1) tma_load_2d_mcast(dst, my_map, &full_bar[s], it * kKb, my_row0, cl_mask);
          else tma_load_2d(dst, my_map, &full_bar[s], it * kKb, my_row0);
        }
        if (++s == kRing) { s = 0; ph ^= 1u; }
        if (!resident_issued && (it + 1 >= kRing || it + 1 == L0.kb_in)) {
          // the ring is primed: now the weights that stay resident (forward operands of the layers behind)
          resident_issued = true;
          for (int l = 1; l < L; ++l) {
            const ColLayer& Y = p.L[l];
            const uint32_t kb_bytes = static_cast<uint32_t>(Y.rows_out8) * 128u;
            mbar_arrive_expect_tx(&bar_wf[l], kb_bytes * static_cast<uint32_t>(Y.kb_in));
            for (int kb = 0; kb < Y.kb_in; ++kb)
              tma_load_2d(smem + Y.off_wf + kb * kb_bytes, &p.tm_wf[l], &bar_wf[l], kb * kKb, 0);
          }
        }
      }
      // the data-gradient operands move into the ring's space once the first layer has finished with it
      if (L > 1) {
        mbar_wait(l1_done, 0);
        for (int l = L - 1; l >= 1; --l) {
          const ColLayer& Y = p.L[l];
          const uint32_t kb_bytes = static_cast<uint32_t>(Y.rows_in8) * 128u;
          mbar_arrive_expect_tx(&bar_wd[l], kb_bytes * static_cast<uint32_t>(Y.kb_out));
          for (int kb = 0; kb < Y.kb_out; ++kb)
            tma_load_2d(smem + Y.off_wd + kb * kb_bytes, &p.tm_wd[l], &bar_wd[l], kb * kKb, 0);
        }
      }
    }
  } else if (warp == kWarpMma) {
    // ===================== MMA issuer (one thread) =====================
    if (lane == 0) {
      const uint32_t idesc = umma_idesc_bf16(128, kColN);
      const uint32_t smem_base = smem_u32(smem);
      {  // stage 0: first layer, weights from the ring
        const ColLayer& Y = p.L[0];
        mbar_wait(bar_x, 0);
        int s = 0;
        uint32_t ph = 0;
        for (int it = 0; it < Y.kb_in; ++it) {
          mbar_wait(&full_bar[s], ph);
          tc_fence_after();
          const uint32_t a_addr = smem_base + p.off_ring + s * p.ring_stage_bytes;
          const uint32_t b_addr = smem_base + p.off_x + it * kBTile;
          const int ksteps = min(4, Y.k16_in - it * 4);
          for (int mt = 0; mt < Y.mt_out; ++mt)
            for (int k = 0; k < ksteps; ++k)
              umma_bf16_ss(tmem_base + Y.tm_fwd + mt * kColN, umma_desc_k_sw128(a_addr + mt * (128 * 128) + k * 32),
                           umma_desc_k_sw128(b_addr + k * 32), idesc, (it | k) != 0 ? 1u : 0u);
          if (cl > 1) umma_commit_mcast(&empty_bar[s], cl_mask);  // every CTA's producer feeds this slot
          else umma_commit(&empty_bar[s]);
          if (++s == kRing) { s = 0; ph ^= 1u; }
        }
        umma_commit(l1_done);
        umma_commit(acc_bar);
      }
      for (int t = 1; t < n_stages; ++t) {
        const bool fwd = t < L;
        const int l = fwd ? t : (2 * L - 1 - t);  // forward layer l, or data gradient THROUGH layer l (l = L-1 .. 1)
        const ColLayer& Y = p.L[l];
        mbar_wait(fwd ? &bar_wf[l] : &bar_wd[l], 0);
        mbar_wait(ready_bar, static_cast<uint32_t>((t - 1) & 1));
        tc_fence_after();
        const int kb_bytes = (fwd ? Y.rows_out8 : Y.rows_in8) * 128;
        const uint32_t a_addr = smem_base + (fwd ? Y.off_wf : Y.off_wd);
        const uint32_t b_addr = smem_base + (fwd ? p.L[l - 1].off_act : Y.off_dz);
        const int ksteps = fwd ? Y.k16_in : Y.k16_out;
        const int mts = fwd ? Y.mt_out : Y.mt_in;
        const uint32_t d_col = tmem_base + (fwd ? Y.tm_fwd : Y.tm_dg);
        for (int mt = 0; mt < mts; ++mt)
          for (int k = 0; k < ksteps; ++k)
            umma_bf16_ss(d_col + mt * kColN,
                         umma_desc_k_sw128(a_addr + (k >> 2) * kb_bytes + mt * (128 * 128) + (k & 3) * 32),
                         umma_desc_k_sw128(b_addr + (k >> 2) * kBTile + (k & 3) * 32), idesc, k != 0 ? 1u : 0u);
        umma_commit(acc_bar);
      }
    }
  } else if (warp == kWarpAux) {
    // this step's learning rate and EMA decay for the weight-gradient / update kernel that follows
    if (lane == 0 && blockIdx.x == 0 && p.lr_out != nullptr && p.global_step_ptr != nullptr) {
      const long long gs = *p.global_step_ptr;
      const TtdHyper& h = p.hyper;
      double e = static_cast<double>(gs) / static_cast<double>(h.decay_steps);
      if (h.staircase) e = floor(e);
      p.lr_out[0] = static_cast<float>(static_cast<double>(h.lr0) * pow(static_cast<double>(h.decay_rate), e));
      const double n = static_cast<double>(gs + 1);
      double d = h.ema_decay;
      if (h.ema_warmup) d = fmin(d, (1.0 + n) / (10.0 + n));
      p.lr_out[1] = static_cast<float>(d);
    }
  } else {
    // ===================== 16 epilogue warps: TMEM lane quadrant q, examples 4*cg .. 4*cg+3 =====================
    const int q = warp & 3, cg = warp >> 2;
    const int n0 = 4 * cg;                       // first of this thread's four examples (columns)
    const bool drop = p.drop_rate > 0.0f;
    const float keep = 1.0f - p.drop_rate, inv_keep = 1.0f / keep;
    const int act = p.act;
    const uint32_t trow = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + static_cast<uint32_t>(n0);

    // ---- dropout bits of every hidden layer: one Philox block per (8 features, example) --------------------
    if (drop) {
      const uint32_t thresh = dropout_thresh16(p.drop_rate);
      for (int task = threadIdx.x; task < p.n_keep_tasks; task += kEpiThreads) {
        int l = 0;
#pragma unroll
        for (int i = 1; i < kColMaxLayers - 1; ++i)
          if (i < L - 1 && task >= p.L[i].keep_task0) l = i;
        const int t = task - p.L[l].keep_task0;
        const int g = t >> 4, n = t & 15;
        const Philox4 r = dropout_block(static_cast<uint32_t>(e0 + n), static_cast<uint32_t>(g), step, p.L[l].layer_id,
                                        p.seed_lo, p.seed_hi);
        uint32_t bits = 0;
#pragma unroll
        for (int j = 0; j < 8; ++j) bits |= (dropout_lane16(r, j) >= thresh ? 1u : 0u) << j;
        smem[p.L[l].off_keep + g * 16 + n] = static_cast<uint8_t>(bits);
      }
    }
    // labels of this thread's examples (read by the softmax stage; lanes of a warp share them)
    long long lab[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int m = e0 + n0 + j;
      lab[j] = (m < p.B && p.labels_pool != nullptr) ? p.labels_pool[static_cast<long long>(slot) * p.B + m] : 0;
    }
    asm volatile("bar.sync 1, %0;" ::"n"(kEpiThreads) : "memory");  // keep bytes are complete

    for (int t = 0; t < n_stages; ++t) {
      const bool fwd = t < L;
      const int l = fwd ? t : (2 * L - 1 - t);
      const ColLayer& Y = p.L[l];
      // bias / side data do not depend on the accumulator: fetch before waiting for it
      const int nfeat = fwd ? Y.out : Y.in;      // features (TMEM lanes) of this stage's result
      const int mts = fwd ? Y.mt_out : Y.mt_in;
      float bias_r[2] = {0.0f, 0.0f};
      if (fwd) {
#pragma unroll
        for (int mt = 0; mt < 2; ++mt) {
          const int f = mt * 128 + q * 32 + lane;
          if (mt < mts && f < nfeat && Y.bias != nullptr) bias_r[mt] = Y.bias[f];
        }
      }
      mbar_wait(acc_bar, static_cast<uint32_t>(t & 1));
      tc_fence_after();
      if (tr) tr[1 + 2 * t] = globaltimer_ns();

      if (fwd && l < L - 1) {
        // ---------------- hidden layer: bias + activation + dropout -> next operand (smem) + aT (global) -----
#pragma unroll
        for (int mt = 0; mt < 2; ++mt) {
          if (mt >= mts || mt * 128 + q * 32 >= nfeat) continue;  // warp-uniform
          const int f = mt * 128 + q * 32 + lane;
          float v[4];
          tmem_ld4(trow + Y.tm_fwd + mt * kColN, v);
          if (f < nfeat) {
            const uint32_t kbytes = drop ? *reinterpret_cast<const uint32_t*>(smem + Y.off_keep + (f >> 3) * 16 + n0) : 0u;
            float r[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              const float a = act_apply(v[j] + bias_r[mt], act);
              const bool kept = !drop || ((kbytes >> (8 * j + (f & 7))) & 1u);
              r[j] = (e0 + n0 + j < p.B) ? (drop ? (kept ? a * inv_keep : 0.0f) : a) : 0.0f;
            }
            const uint32_t p01 = pack2(r[0], r[1]), p23 = pack2(r[2], r[3]);
            uint8_t* ab = smem + Y.off_act;
            *reinterpret_cast<uint16_t*>(ab + bop_off(n0 + 0, f)) = static_cast<uint16_t>(p01 & 0xFFFFu);
            *reinterpret_cast<uint16_t*>(ab + bop_off(n0 + 1, f)) = static_cast<uint16_t>(p01 >> 16);
            *reinterpret_cast<uint16_t*>(ab + bop_off(n0 + 2, f)) = static_cast<uint16_t>(p23 & 0xFFFFu);
            *reinterpret_cast<uint16_t*>(ab + bop_off(n0 + 3, f)) = static_cast<uint16_t>(p23 >> 16);
            if (Y.aT != nullptr && e0 + n0 < Y.ld_aT)
              *reinterpret_cast<uint2*>(Y.aT + static_cast<long long>(f) * Y.ld_aT + e0 + n0) = make_uint2(p01, p23);
          }
        }
      } else if (fwd) {
        // ---------------- output layer: logits^T [classes x 16] -> softmax cross-entropy, accuracy, dlogits ---
        if (q == 0) {  // <= 32 classes: all of them in the lanes of quadrant 0
          const int c = lane;
          float v[4];
          tmem_ld4(trow + Y.tm_fwd, v);
          const bool c_ok = c < nfeat;
          const float gscale = 1.0f / static_cast<float>(p.B);
          float loss = 0.0f, corr = 0.0f;
          float d[4];
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const bool m_ok = e0 + n0 + j < p.B;
            const float z = c_ok ? v[j] + bias_r[0] : -INFINITY;
            const float zmax = warp_max(z);
            const float e = c_ok ? __expf(z - zmax) : 0.0f;
            const float sum = warp_sum(e);
            const int label = static_cast<int>(lab[j]);
            const float zlab = __shfl_sync(0xffffffffu, z, label & 31);
            const unsigned greater = __ballot_sync(0xffffffffu, c_ok && z > zlab);
            d[j] = (c_ok && m_ok) ? (e * (1.0f / sum) - (c == label ? 1.0f : 0.0f)) * gscale : 0.0f;
            if (m_ok) {
              loss += (__logf(sum) + zmax - zlab) * gscale;
              corr += greater == 0u ? gscale : 0.0f;
            }
          }
          if (c_ok) {
            const uint32_t p01 = pack2(d[0], d[1]), p23 = pack2(d[2], d[3]);
            uint8_t* zb = smem + Y.off_dz;
            *reinterpret_cast<uint16_t*>(zb + bop_off(n0 + 0, c)) = static_cast<uint16_t>(p01 & 0xFFFFu);
            *reinterpret_cast<uint16_t*>(zb + bop_off(n0 + 1, c)) = static_cast<uint16_t>(p01 >> 16);
            *reinterpret_cast<uint16_t*>(zb + bop_off(n0 + 2, c)) = static_cast<uint16_t>(p23 & 0xFFFFu);
            *reinterpret_cast<uint16_t*>(zb + bop_off(n0 + 3, c)) = static_cast<uint16_t>(p23 >> 16);
            if (Y.dzT != nullptr && e0 + n0 < Y.ld_dzT)
              *reinterpret_cast<uint2*>(Y.dzT + static_cast<long long>(c) * Y.ld_dzT + e0 + n0) = make_uint2(p01, p23);
          }
          if (lane == 0 && p.metrics != nullptr) {
            atomicAdd(&p.metrics[0], loss);
            atomicAdd(&p.metrics[1], corr);
          }
        }
      } else {
        // ---------------- data gradient through layer l: dZ_{l-1} = dA_{l-1} * mask/keep * act'(a_{l-1}) ------
        const ColLayer& X = p.L[l - 1];  // the layer whose output is being differentiated
#pragma unroll
        for (int mt = 0; mt < 2; ++mt) {
          if (mt >= mts || mt * 128 + q * 32 >= nfeat) continue;
          const int f = mt * 128 + q * 32 + lane;
          float v[4];
          tmem_ld4(trow + Y.tm_dg + mt * kColN, v);
          if (f < nfeat) {
            const uint32_t kbytes = drop ? *reinterpret_cast<const uint32_t*>(smem + X.off_keep + (f >> 3) * 16 + n0) : 0u;
            const uint8_t* ab = smem + X.off_act;
            float r[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              const float a_sv = __bfloat162float(*reinterpret_cast<const __nv_bfloat16*>(ab + bop_off(n0 + j, f)));
              const bool kept = !drop || ((kbytes >> (8 * j + (f & 7))) & 1u);
              const float g = v[j] * act_grad_from_out(drop ? a_sv * keep : a_sv, act);
              r[j] = (e0 + n0 + j < p.B) ? (drop ? (kept ? g * inv_keep : 0.0f) : g) : 0.0f;
            }
            const uint32_t p01 = pack2(r[0], r[1]), p23 = pack2(r[2], r[3]);
            if (l - 1 >= 1) {  // operand of the next data-gradient stage
              uint8_t* zb = smem + X.off_dz;
              *reinterpret_cast<uint16_t*>(zb + bop_off(n0 + 0, f)) = static_cast<uint16_t>(p01 & 0xFFFFu);
              *reinterpret_cast<uint16_t*>(zb + bop_off(n0 + 1, f)) = static_cast<uint16_t>(p01 >> 16);
              *reinterpret_cast<uint16_t*>(zb + bop_off(n0 + 2, f)) = static_cast<uint16_t>(p23 & 0xFFFFu);
              *reinterpret_cast<uint16_t*>(zb + bop_off(n0 + 3, f)) = static_cast<uint16_t>(p23 >> 16);
            }
            if (X.dzT != nullptr && e0 + n0 < X.ld_dzT)
              *reinterpret_cast<uint2*>(X.dzT + static_cast<long long>(f) * X.ld_dzT + e0 + n0) = make_uint2(p01, p23);
          }
        }
      }
      if (tr) tr[2 + 2 * t] = globaltimer_ns();
      if (t + 1 < n_stages) {
        fence_proxy_async();   // generic-proxy stores to the operand buffer -> visible to the tensor core's reads
        tc_fence_before();     // this warp's tcgen05.ld are ordered before the MMA thread's next tcgen05.mma
        __syncwarp();
        if (lane == 0) mbar_arrive(ready_bar);
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (tr) tr[63] = globaltimer_ns();
  if (cl > 1) cluster_sync_all();  // nobody leaves while a peer's commit may still signal one of its barriers
  if (warp == kWarpAux) {
    tc_fence_after();
    tmem_dealloc(tmem_base, p.tmem_cols);
  }
}

struct ColumnHandle {
  ColParams p;
  size_t smem_bytes;
  int grid;
};

inline int ceil_div(int a, int b) { return (a + b - 1) / b; }

// Shared-memory / TMEM plan for a stack of layers; returns the dynamic shared memory it needs or -1.
long long plan_layers(ColParams* P, const int* ins, const int* outs, int L, const char** why) {
  static const char* msg = "";
  if (why) *why = msg;
  if (L < 2 || L > kColMaxLayers) { if (why) *why = "2..6 layers"; return -1; }
  if (outs[L - 1] > 32) { if (why) *why = "at most 32 classes"; return -1; }
  uint32_t off = 0, tm = 0;
  int tasks = 0;
  for (int l = 0; l < L; ++l) {
    ColLayer& Y = P->L[l];
    Y.in = ins[l];
    Y.out = outs[l];
    if (Y.out > 256 || Y.out < 1 || Y.in < 1 || (l > 0 && Y.in > 256)) { if (why) *why = "hidden widths must be in [1,256]"; return -1; }
    Y.k16_in = ceil_div(Y.in, 16); Y.k16_out = ceil_div(Y.out, 16);
    Y.kb_in = ceil_div(Y.in, kKb); Y.kb_out = ceil_div(Y.out, kKb);
    Y.mt_out = ceil_div(Y.out, 128); Y.mt_in = ceil_div(Y.in, 128);
    Y.rows_out8 = ceil_div(Y.out, 8) * 8; Y.rows_in8 = ceil_div(Y.in, 8) * 8;
    Y.tm_fwd = tm; tm += Y.mt_out * kColN;
    if (l >= 1) { Y.tm_dg = tm; tm += Y.mt_in * kColN; }
  }
  uint32_t cols = 32;
  while (cols < tm) cols <<= 1;
  if (cols > 512) { if (why) *why = "accumulators exceed 512 TMEM columns"; return -1; }
  P->tmem_cols = cols;
  // resident forward weights
  for (int l = 1; l < L; ++l) { P->L[l].off_wf = off; off += P->L[l].rows_out8 * 128u * P->L[l].kb_in; }
  // ring (first layer) / data-gradient weights share one region
  P->off_ring = off;
  P->ring_stage_bytes = P->L[0].rows_out8 * 128u;
  uint32_t dg = 0;
  for (int l = L - 1; l >= 1; --l) { P->L[l].off_wd = off + dg; dg += P->L[l].rows_in8 * 128u * P->L[l].kb_out; }
  const uint32_t ring = kRing * P->ring_stage_bytes;
  uint32_t region = ring > dg ? ring : dg;
  // an MMA always reads 128 rows per tile: keep what follows the last weight tile inside the allocation
  off += region + 128 * 128;
  P->off_x = off; off += P->L[0].kb_in * kBTile;
  P->off_zero = off;
  for (int l = 0; l < L - 1; ++l) { P->L[l].off_act = off; off += P->L[l].kb_out * kBTile; }
  for (int l = 1; l < L; ++l) { P->L[l].off_dz = off; off += P->L[l].kb_out * kBTile; }
  P->zero_bytes = off - P->off_zero;
  for (int l = 0; l < L - 1; ++l) {
    P->L[l].off_keep = off;
    P->L[l].keep_task0 = tasks;
    const int groups = ceil_div(P->L[l].out, 8);
    off += groups * 16;
    tasks += groups * 16;
  }
  P->n_keep_tasks = tasks;
  off = (off + 63u) & ~63u;
  P->off_bars = off;
  off += (4 + 2 * kRing + 2 * kColMaxLayers) * 8 + 16;
  const long long total = static_cast<long long>(off) + 1024;  // alignment slack
  if (total > kSmemLimit) { if (why) *why = "operands exceed 227 KB of shared memory"; return -1; }
  return total;
}

}  // namespace
}  // namespace ttd

using namespace ttd;

extern "C" int ttd_mlp_column_fits(int32_t input_dim, const int32_t* layer_sizes, int32_t n_layers) {
  if (n_layers < 2 || n_layers > kColMaxLayers || input_dim < 1) return 0;
  int ins[kColMaxLayers], outs[kColMaxLayers];
  for (int l = 0; l < n_layers; ++l) {
    ins[l] = l == 0 ? input_dim : layer_sizes[l - 1];
    outs[l] = layer_sizes[l];
  }
  ColParams tmp;
  memset(&tmp, 0, sizeof(tmp));
  return plan_layers(&tmp, ins, outs, n_layers, nullptr) > 0 ? 1 : 0;
}

extern "C" void* ttd_mlp_column_create(const TtdColumnDesc* d) {
  if (d == nullptr || d->B < 1 || d->x_pool == nullptr || d->pool_batches < 1) {
    set_error("ttd_mlp_column_create: need a batch size, an input pool and >= 1 pool batches");
    return nullptr;
  }
  ColumnHandle* h = new ColumnHandle();
  memset(&h->p, 0, sizeof(h->p));
  ColParams& P = h->p;
  const int L = d->n_layers;
  int ins[kColMaxLayers], outs[kColMaxLayers];
  for (int l = 0; l < L && l < kColMaxLayers; ++l) { ins[l] = d->layer[l].in; outs[l] = d->layer[l].out; }
  const char* why = "";
  const long long smem = plan_layers(&P, ins, outs, L, &why);
  if (smem < 0) {
    set_error("ttd_mlp_column_create: model does not fit the column kernel (%s)", why);
    delete h;
    return nullptr;
  }
  for (int l = 1; l < L; ++l) {
    if (d->layer[l].in != d->layer[l - 1].out) {
      set_error("ttd_mlp_column_create: layer %d input %d != layer %d output %d", l, d->layer[l].in, l - 1, d->layer[l - 1].out);
      delete h;
      return nullptr;
    }
  }
  h->smem_bytes = static_cast<size_t>(smem);
  P.n_layers = L; P.B = d->B; P.act = d->act; P.drop_rate = d->drop_rate; P.pool_batches = d->pool_batches;
  const int ctas = ceil_div(d->B, kColN);
  int cl = d->cluster > 0 ? d->cluster : 8;
  if (const char* e = getenv("TTD_COL_CLUSTER")) cl = atoi(e);  // tuning knob
  if (cl > 8) cl = 8;
  while (cl > 1 && (ctas % cl != 0 || (cl & (cl - 1)) != 0)) --cl;
  if (cl < 1) cl = 1;
  P.cluster = cl;
  h->grid = ctas;
  const int groups = P.L[0].rows_out8 / 8;
  P.w1_groups_base = groups / cl;
  P.w1_groups_extra = groups % cl;
  P.labels_pool = reinterpret_cast<const long long*>(d->labels_pool);
  P.metrics = d->metrics;
  P.step_ptr = reinterpret_cast<const unsigned long long*>(d->step_ptr);
  P.global_step_ptr = reinterpret_cast<const long long*>(d->global_step_ptr);
  P.lr_out = d->lr_out;
  P.hyper = d->hyper;
  int rc = make_tmap(&P.tm_x, d->x_pool, static_cast<int64_t>(d->pool_batches) * d->B, P.L[0].in, d->ld_x, kColN);
  const TtdColumnLayer& y0 = d->layer[0];
  // both slice maps are always encoded (a map with an unused box size is harmless); box rows are multiples of 8
  if (!rc) rc = make_tmap(&P.tm_w1[0], y0.w_t, y0.out, y0.in, y0.ld_t, (P.w1_groups_base + 1) * 8);
  if (!rc) rc = make_tmap(&P.tm_w1[1], y0.w_t, y0.out, y0.in, y0.ld_t, P.w1_groups_base > 0 ? P.w1_groups_base * 8 : 8);
  for (int l = 0; l < L && !rc; ++l) {
    const TtdColumnLayer& y = d->layer[l];
    ColLayer& Y = P.L[l];
    Y.layer_id = y.layer_id;
    Y.bias = y.bias;
    Y.aT = static_cast<__nv_bfloat16*>(y.aT); Y.ld_aT = y.ld_aT;
    Y.dzT = static_cast<__nv_bfloat16*>(y.dzT); Y.ld_dzT = y.ld_dzT;
    if ((Y.aT && ((y.ld_aT & 3) || (reinterpret_cast<uintptr_t>(y.aT) & 7))) ||
        (Y.dzT && ((y.ld_dzT & 3) || (reinterpret_cast<uintptr_t>(y.dzT) & 7)))) {
      set_error("ttd_mlp_column_create: aT / dzT of layer %d must be 8-byte aligned with a leading dimension that is a multiple of 4", l);
      rc = -1;
      break;
    }
    if (l >= 1) {
      if (y.w_t == nullptr || y.w_rm == nullptr) {
        set_error("ttd_mlp_column_create: layer %d needs both compute copies (w_t and w_rm)", l);
        rc = -1;
        break;
      }
      rc = make_tmap(&P.tm_wf[l], y.w_t, y.out, y.in, y.ld_t, Y.rows_out8);
      if (!rc) rc = make_tmap(&P.tm_wd[l], y.w_rm, y.in, y.out, y.ld_rm, Y.rows_in8);
    }
  }
  if (rc) {
    delete h;
    return nullptr;
  }
  return h;
}

static unsigned long long* g_col_trace = nullptr;
// Diagnostic: device buffer of 64 uint64 that receives CTA 0's stage timestamps (NULL switches it off).
extern "C" void ttd_mlp_column_set_trace(void* device_buf) { g_col_trace = static_cast<unsigned long long*>(device_buf); }

extern "C" void ttd_mlp_column_destroy(void* handle) { delete static_cast<ColumnHandle*>(handle); }

extern "C" int ttd_mlp_column_launch(void* handle, uint64_t seed, uint32_t flags, void* stream) {
  ColumnHandle* h = static_cast<ColumnHandle*>(handle);
  if (h == nullptr) {
    set_error("ttd_mlp_column_launch: NULL handle");
    return -1;
  }
  static bool attr_set[64] = {false};
  int dev = 0;
  TTD_CUDA(cudaGetDevice(&dev));
  if (dev >= 0 && dev < 64 && !attr_set[dev]) {
    TTD_CUDA(cudaFuncSetAttribute(mlp_column_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemLimit));
    attr_set[dev] = true;
  }
  h->p.seed_lo = static_cast<uint32_t>(seed);
  h->p.seed_hi = static_cast<uint32_t>(seed >> 32);
  h->p.use_pdl = (flags & 1u) ? 1 : 0;
  h->p.trace = g_col_trace;
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = dim3(static_cast<unsigned>(h->grid));
  cfg.blockDim = dim3(kColThreads);
  cfg.dynamicSmemBytes = h->smem_bytes;
  cfg.stream = static_cast<cudaStream_t>(stream);
  cudaLaunchAttribute attr[2];
  int nattr = 0;
  if (h->p.cluster > 1) {
    attr[nattr].id = cudaLaunchAttributeClusterDimension;
    attr[nattr].val.clusterDim.x = static_cast<unsigned>(h->p.cluster);
    attr[nattr].val.clusterDim.y = 1;
    attr[nattr].val.clusterDim.z = 1;
    ++nattr;
  }
  if (h->p.use_pdl) {
    attr[nattr].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[nattr].val.programmaticStreamSerializationAllowed = 1;
    ++nattr;
  }
  cfg.attrs = attr;
  cfg.numAttrs = nattr;
  TTD_CUDA(cudaLaunchKernelEx(&cfg, mlp_column_kernel, h->p));
  g_launch_count.fetch_add(1);
  return 0;
}
