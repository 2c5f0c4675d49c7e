// K3 (fast path): tcgen05 / TMEM flash-attention forward over the paged KV cache,
// head_size 128, bf16/f16 queries and KV.
//
// Computes what the reference's 2D kernel computes
// (reference ibm-triton-lib/ibm_triton_lib/kernels/triton_unified_attention.py:
//  kernel_unified_attention_2d :275-523) for the sequences of the batch whose
// query is long enough to fill a tensor-core tile: rows of a Q tile are
// (query token, query head of the GQA group) pairs in token-major order
// (:343-346), KV tiles are walked through the block table up to the causal limit
// of the tile (:384-393), causal/context and sliding-window masks, online softmax
// in fp32, acc / L epilogue (:511).  The grid is the reference's host-int upper
// bound (:886-889) with its q-block -> sequence binary search (:32-52).
//
// How (nothing like the Triton kernel):
//   * one CTA = 256 Q rows (two 128-row subtiles sharing every K/V tile) x one kv
//     head; S = Q K^T and O += P V run on the 5th-gen tensor cores
//     (tcgen05.mma, M=128 N=128 K=16, single issuing thread), accumulators in
//     TMEM: S0|S1|O0|O1 = 4 x 128 columns = all 512;
//   * K/V pages and the Q tile are moved by TMA (cp.async.bulk.tensor) into
//     128-byte-swizzled shared memory straight from the strided cache view; the
//     same [token][128 B] layout serves K as a K-major and V as an MN-major UMMA
//     operand, so V needs no transpose;
//   * warp roles: 2 softmax warpgroups (one thread per row: TMEM lane == row, so
//     row max / row sum need no shuffles), 1 MMA-issue warp, 1 TMA warp (which also
//     zero-fills the unused V rows of a sequence's last tile);
//   * P goes back to TMEM in place of S (bf16 pairs) and feeds the PV MMA as the
//     TMEM A operand - it never touches shared memory;
//   * the MMA order is PV0(j) QK0(j+1) PV1(j) QK1(j+1): one subtile's softmax
//     runs under the other subtile's MMAs;
//   * O is rescaled lazily: only when a row max grows by more than 2^8, by the
//     softmax warps themselves (so there is no per-tile correction pass).
#include "tc_common.cuh"

namespace b200attn {
namespace {

constexpr int kBN = 128;                        // kv tokens per tile
constexpr int kKStages = 3;                     // K is needed right after PV of the previous tile: prefetch deeper
constexpr int kVStages = 2;
constexpr int kSubtileBytes = 128 * 128;        // [128 rows][128 B]: one 64-element chunk of 128 rows
constexpr int kThreads = 384;
// Head size kD is 64 or 128: a tile of 128 rows is kD / 64 sub-tiles of [128 rows][128 B].
// TMEM columns: S0 | S1 | O0 | O1 (subtile w: S at 128 w, O at 256 + kD w)
__host__ __device__ constexpr uint32_t col_s(int w) { return 128u * (uint32_t)w; }
template <int kD> __host__ __device__ constexpr uint32_t col_o(int w) { return 256u + (uint32_t)kD * (uint32_t)w; }
template <int kD> struct Smem {
  static constexpr int kTileBytes = (kD / 64) * kSubtileBytes;  // 128 rows x kD dims x 2 B
  static constexpr int kQ = 0;
  static constexpr int kK = 2 * kTileBytes;
  static constexpr int kV = kK + kKStages * kTileBytes;
  static constexpr int kBar = kV + kVStages * kTileBytes;
  static constexpr int kBarBytes = 256;                 // NUM_BARS * 8 = 208, tmem pointer behind them
  static constexpr int kItems = kBar + kBarBytes;       // 2 x Item
  static constexpr int kClc = kItems + 2 * 64;          // 16-byte try_cancel response
  static constexpr int kBytes = kClc + 64 + 1024;       // + alignment slack
};
constexpr float kRescaleThreshold = 8.0f;          // log2 units

#ifdef B200ATTN_TC_TRACE
// debug build only (scripts/ubench/tc_trace.cu): clock64 stamps of CTA (0, 0)'s pipeline events
__device__ long long g_tc_trace[65 * 32];
__device__ long long g_tc_cta[4096 * 4];  // per CTA: smid, start ns, end ns, tiles
__device__ __forceinline__ long long tc_globaltimer() {
  long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}
#define TC_TRACE(j, ev)                                                            \
  do {                                                                             \
    if (blockIdx.x == 0 && blockIdx.y == 1 && (j) < 64) g_tc_trace[(j) * 32 + (ev)] = clock64(); \
  } while (0)
#define TC_MARK(k)                                                                \
  do {                                                                             \
    if (blockIdx.x == 1 && blockIdx.y == 0) g_tc_trace[64 * 32 + (k)] = clock64(); \
  } while (0)
#else
#define TC_MARK(k) do { } while (0)
#define TC_TRACE(j, ev) do { } while (0)
#endif

enum Bar {
  Q_FULL = 0, Q_EMPTY = 1, K_FULL = 2, K_EMPTY = 5, V_FULL = 8, V_EMPTY = 10, S_FULL = 12, P_FULL = 14, P_FULLB = 16,
  O_DONE = 18, V_FIX = 20, ITEM_FULL = 21, ITEM_EMPTY = 23, CLC_DONE = 25, NUM_BARS = 26
};
constexpr int kConsumerWarps = 11;                   // warps 0..10 read every item

// kEmuNum of every kEmuDen exponential pairs are computed by ex2_poly2, the rest on the MUFU.
// kLogits selects how a logit is formed from the dot product S (always right after S leaves TMEM and before
// masking: the reference caps before it masks, triton_unified_attention.py:24-29, 467-468):
//   0  S * scale                                      (the scale rides in the exp2 FFMA)
//   1  cap * tanh(S * scale / cap)                    (tanh() replaces S, one MUFU op; `cap` rides in the FFMA)
//   2  either of the above + slope * (key position - context length) for ALiBi (:481-482), formed in place in
//      log2 units; the max / exp2 code then runs with a scale of 1.
template <int kD, bool kBf16, int kEmuNum, int kEmuDen, int kLogits = 0>
__global__ void __launch_bounds__(kThreads, 1)
attn_prefill_tc_kernel(const TcParams P, const __grid_constant__ CUtensorMap q_map, const __grid_constant__ CUtensorMap k_map,
                       const __grid_constant__ CUtensorMap v_map) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  using L = Smem<kD>;
  constexpr int kTileBytes = L::kTileBytes;
  constexpr int kNDC = kD / 64;  // 128-byte chunks of a row
  static_assert(kD == 64 || kD == 128, "head size");
  uint8_t* sQ = smem + L::kQ;
  uint8_t* sK = smem + L::kK;
  uint8_t* sV = smem + L::kV;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + L::kBar);
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(bars + NUM_BARS);
  Item* items = reinterpret_cast<Item*>(smem + L::kItems);
  uint4* clc_resp = reinterpret_cast<uint4*>(smem + L::kClc);

  const b200attn_attention_params& a = P.a;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int g = P.group, gp = P.group_pad;
  const int BQ = 2 * P.block_q;  // query tokens per CTA tile
  const int window = a.sliding_window;
#ifdef B200ATTN_TC_TRACE
  const long long tc_t0 = tc_globaltimer();
  int tc_tiles = 0;
#endif

  // ---- one-time setup ------------------------------------------------------------------
  if (warp == 8 && lane == 0) {
    mbar_init(&bars[Q_FULL], 1);
    mbar_init(&bars[Q_EMPTY], 1);
    for (int s = 0; s < kKStages; ++s) {
      mbar_init(&bars[K_FULL + s], 1);
      mbar_init(&bars[K_EMPTY + s], 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(&bars[V_FULL + s], 1);
      mbar_init(&bars[V_EMPTY + s], 1);
      mbar_init(&bars[S_FULL + s], 1);
      mbar_init(&bars[P_FULL + s], 128);
      mbar_init(&bars[P_FULLB + s], 128);
      mbar_init(&bars[O_DONE + s], 1);
      mbar_init(&bars[ITEM_FULL + s], 1);
      mbar_init(&bars[ITEM_EMPTY + s], kConsumerWarps);
    }
    mbar_init(&bars[V_FIX], 1);
    mbar_init(&bars[CLC_DONE], 1);
    fence_barrier_init();
  }
  if (warp == 11) tmem_alloc(tmem_ptr_smem, 512);
  if (warp == 9 && lane == 0) {
    tma_prefetch_desc(&q_map);
    tma_prefetch_desc(&k_map);
  }
  if (warp == 10 && lane == 0) tma_prefetch_desc(&v_map);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;

  // ---- state shared by the roles that consume the work queue -------------------------------
  uint32_t n_items = 0;   // items taken so far (mailbox slot / phase)
  uint32_t t_run = 0;     // KV tiles processed so far: parity source of the per-tile barriers
  const float scale_log2 = kLogits == 0 ? a.softmax_scale * kLog2e : (kLogits == 1 ? a.softcap * kLog2e : 1.0f);
  int k_st = 0, v_st = 0;           // ring positions (TMA warps: next to fill; MMA warp: next to use)
  uint32_t k_ph = 0, v_ph = 0;
  uint32_t n_fix = 0;               // V tail fix-ups so far (V_FIX parity)
  const bool leader = elect_one();  // MMA warp: the issuing lane
  const uint32_t tm = __shfl_sync(0xffffffffu, tmem_base, 0);
  const uint64_t q_desc0 = umma_desc(__shfl_sync(0xffffffffu, smem_u32(sQ), 0), 16, 1024);
  const uint64_t k_desc0 = umma_desc(__shfl_sync(0xffffffffu, smem_u32(sK), 0), 16, 1024);
  const uint64_t v_desc0 = umma_desc(__shfl_sync(0xffffffffu, smem_u32(sV), 0), kSubtileBytes, 1024);
  // next work item from the mailbox; false = no more work
  auto next_item = [&](Item& it) -> bool {
    const uint32_t slot = n_items & 1;
    mbar_wait(&bars[ITEM_FULL + slot], (n_items >> 1) & 1);
    it = items[slot];
    __syncwarp();
    if (lane == 0) mbar_arrive(&bars[ITEM_EMPTY + slot]);  // the copy is in registers: the slot is free
    return it.valid != 0;
  };
#define B200ATTN_ITEM_VARS                                                                              \
  const int n_tiles = it.n_tiles, tile0 = it.tile0, kvh = it.kvh;                                       \
  const int last_tile = n_tiles - 1;                                                                    \
  const int last_valid = it.kv_limit - (tile0 + last_tile) * kBN; /* visible rows of the last tile */ \
  const bool need_fix = last_valid < kBN;                                                               \
  (void)kvh, (void)need_fix, (void)last_valid;

  // Register budget follows the roles (each role has its own loop after its setmaxnreg): the
  // softmax warpgroups hold a 128-column row of S each, the MMA / TMA / scheduler warpgroup
  // needs almost nothing (2 x 208 + 80 <= 512 per lane slot).
  if (warp < 8) {
    asm volatile("setmaxnreg.inc.sync.aligned.u32 208;");
    Item it;
    while (next_item(it)) {
      B200ATTN_ITEM_VARS
#ifdef B200ATTN_TC_TRACE
      tc_tiles += n_tiles;
#endif
      // ============================ softmax warpgroups ====================================
      const int w = warp >> 2;                  // Q subtile
      const int row = (warp & 3) * 32 + lane;   // row of the subtile == TMEM lane
      const int qpos = it.qpos0 + (w * 128 + row) / gp;
      const int head = row % gp;                // 128 % gp == 0; heads >= g are padding rows
      const bool row_valid = qpos < it.q_len && head < g;
      const int limit = row_valid ? it.ctx + qpos + 1 : 0;                        // visible keys: [lo, limit)
      const int lo = (window > 0 && row_valid) ? max(0, it.ctx + qpos - window + 1) : 0;
      const uint32_t lane_base = (uint32_t)((warp & 3) * 32) << 16;
      const uint32_t s_tmem = tmem_base + lane_base + col_s(w);
      const uint32_t o_tmem = tmem_base + lane_base + col_o<kD>(w);
      float m_used = -INFINITY, l_sum = 0.f;

      for (int j = 0; j < n_tiles; ++j) {
        const uint32_t par = (t_run + (uint32_t)j) & 1;
        mbar_wait(&bars[S_FULL + w], par);
        tc_fence_after();
        if (row == 0) TC_TRACE(t_run + j, w * 5 + 0);
        uint32_t s[4][32];
#pragma unroll
        for (int c = 0; c < 4; ++c) tmem_ld32(s_tmem + c * 32, s[c]);
        tmem_ld_wait();
        if (row == 0) TC_TRACE(t_run + j, w * 5 + 1);
        const int n0 = (tile0 + j) * kBN;
        if constexpr (kLogits == 1) {
          const float pre = a.softmax_scale / a.softcap;
#pragma unroll
          for (int c = 0; c < 4; ++c)
#pragma unroll
            for (int i = 0; i < 32; ++i) s[c][i] = __float_as_uint(tanh_approx(__uint_as_float(s[c][i]) * pre));
        } else if constexpr (kLogits == 2) {
          if (a.softcap > 0.f) {
            const float pre = a.softmax_scale / a.softcap, cap_log2 = a.softcap * kLog2e;
#pragma unroll
            for (int c = 0; c < 4; ++c)
#pragma unroll
              for (int i = 0; i < 32; ++i) s[c][i] = __float_as_uint(cap_log2 * tanh_approx(__uint_as_float(s[c][i]) * pre));
          } else {
            const float sl2 = a.softmax_scale * kLog2e;
#pragma unroll
            for (int c = 0; c < 4; ++c)
#pragma unroll
              for (int i = 0; i < 32; ++i) s[c][i] = __float_as_uint(__uint_as_float(s[c][i]) * sl2);
          }
          if (a.alibi_slopes) {
            const float slope = (head < g ? a.alibi_slopes[kvh * g + head] : 0.f) * kLog2e;
            const float b0 = slope * (float)(n0 - it.ctx);
#pragma unroll
            for (int c = 0; c < 4; ++c)
#pragma unroll
              for (int i = 0; i < 32; ++i) s[c][i] = __float_as_uint(fmaf(slope, (float)(c * 32 + i), __uint_as_float(s[c][i]) + b0));
          }
          if (a.qq_bias && row_valid && n0 + kBN > it.ctx) {
            // bias between the step's own query tokens: keys ctx .. ctx + stride - 1 are its tokens 0 .. stride - 1
            const float* brow = a.qq_bias + (int64_t)qpos * a.qq_bias_stride0;
            const int nrel = (int)a.qq_bias_stride0;
#pragma unroll
            for (int c = 0; c < 4; ++c)
#pragma unroll
              for (int i = 0; i < 32; ++i) {
                const int rel = n0 + c * 32 + i - it.ctx;
                if (rel >= 0 && rel < nrel) s[c][i] = __float_as_uint(fmaf(brow[rel], kLog2e, __uint_as_float(s[c][i])));
              }
          }
        }
        if (n0 + kBN > limit || n0 < lo) {
#pragma unroll
          for (int c = 0; c < 4; ++c)
#pragma unroll
            for (int i = 0; i < 32; ++i) {
              const int pos = n0 + c * 32 + i;
              if (pos >= limit || pos < lo) s[c][i] = 0xff800000u;  // -inf
            }
        }
        // row max: four FMNMX3 chains (two new values per instruction)
        float mxc[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
#pragma unroll
        for (int c = 0; c < 4; ++c)
#pragma unroll
          for (int i = 0; i < 32; i += 2) mxc[c] = fmax3(mxc[c], __uint_as_float(s[c][i]), __uint_as_float(s[c][i + 1]));
        const float mx = fmax3(fmaxf(mxc[0], mxc[1]), mxc[2], mxc[3]);
        const float m_tile = mx * scale_log2;  // softmax_scale > 0 (checked on the host)
        float alpha = 1.f;
        if (m_tile > m_used + kRescaleThreshold) {  // also the first finite max (m_used = -inf)
          alpha = ex2_approx(m_used - m_tile);       // 0 when m_used = -inf
          m_used = m_tile;
        }
        const float m_eff = (m_used == -INFINITY) ? 0.f : m_used;
        if (j > 0 && __any_sync(0xffffffffu, alpha != 1.f)) {
          // rare: rescale the running output before PV(j) accumulates into it
          mbar_wait(&bars[O_DONE + w], par ^ 1);
          tc_fence_after();
          // (8 columns at a time: all 128 S values are live here)
#pragma unroll 1
          for (int c = 0; c < kD / 8; ++c) {
            uint32_t t[8];
            tmem_ld8(o_tmem + c * 8, t);
            tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 8; ++i) t[i] = __float_as_uint(__uint_as_float(t[i]) * alpha);
            tmem_st8(o_tmem + c * 8, t);
          }
        }
        if (row == 0) TC_TRACE(t_run + j, w * 5 + 2);
        // P = 2^(S*scale - m) in two halves of 64 keys; each half replaces S in place (bf16
        // pairs: columns [32h, 32h+32) of the S region) and is announced on its own barrier,
        // so the first four k-steps of PV run while the second half is still being computed.
        const uint64_t sc2 = f2_pack(scale_log2, scale_log2), nm2 = f2_pack(-m_eff, -m_eff);
        uint64_t sum2[2] = {0ull, 0ull};
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          uint32_t pk[32];
#pragma unroll
          for (int cc = 0; cc < 2; ++cc) {
            const int c = 2 * h + cc;
#pragma unroll
            for (int i = 0; i < 32; i += 2) {
              const int k = cc * 16 + (i >> 1);
              const uint64_t x2 = f2_fma(f2_pack(__uint_as_float(s[c][i]), __uint_as_float(s[c][i + 1])), sc2, nm2);
              float x0, x1, p0, p1;
              f2_unpack(x2, x0, x1);
              uint64_t p2;
              if ((k % kEmuDen) < kEmuNum) {
                p2 = ex2_poly2(x0, x1);
                f2_unpack(p2, p0, p1);
              } else {
                p0 = ex2_approx(x0), p1 = ex2_approx(x1);
                p2 = f2_pack(p0, p1);
              }
              sum2[k & 1] = f2_add(sum2[k & 1], p2);
              pk[k] = pack2<kBf16>(p0, p1);
            }
          }
          tmem_st32(s_tmem + 32 * h, pk);
          tmem_st_wait();
          tc_fence_before();
          mbar_arrive(&bars[(h ? P_FULLB : P_FULL) + w]);
          if (row == 0) TC_TRACE(t_run + j, w * 5 + 3 + h);
        }
        float sa, sb;
        f2_unpack(f2_add(sum2[0], sum2[1]), sa, sb);
        l_sum = l_sum * alpha + (sa + sb);
      }

      // ---- epilogue: O / L -> global -----------------------------------------------------
      // The next item's first PV overwrites O, but it waits for this warpgroup's next P, which
      // is only produced after the loads below have completed: no extra barrier is needed.
      mbar_wait(&bars[O_DONE + w], (t_run + (uint32_t)last_tile) & 1);
      tc_fence_after();
      float fs = 1.f;
      const bool seen = m_used != -INFINITY;  // some key was visible (a sink alone does not make an LSE)
      if (a.sinks && row_valid) fs = sink_rescale<true>(a.sinks[kvh * g + head], m_used, l_sum);  // once per row, after the last tile
      const float inv = (l_sum > 0.f) ? fs / l_sum : 0.f;
      if (a.lse && row_valid) a.lse[(int64_t)(it.q_start + qpos) * a.lse_stride0 + kvh * g + head] = seen ? row_lse<true>(m_used, l_sum) : -INFINITY;
      const bool out_fp8 = a.out_scale != nullptr;  // fused e4m3 output quantisation (b200attn.h: out_scale)
      const float oq = inv * out_quant_factor(a.out_scale);
      char* orow = static_cast<char*>(a.out) +
                   ((int64_t)(it.q_start + (row_valid ? qpos : 0)) * a.out_stride0 + (int64_t)(kvh * g + head) * a.out_stride1) * (out_fp8 ? 1 : 2);
#pragma unroll
      for (int c = 0; c < kD / 32; ++c) {
        uint32_t t[32];
        tmem_ld32(o_tmem + c * 32, t);
        tmem_ld_wait();
        if (row_valid && out_fp8) {
#pragma unroll
          for (int i = 0; i < 32; i += 8) {
            int2 v;
            v.x = (int)fp8x4_out(__uint_as_float(t[i]) * oq, __uint_as_float(t[i + 1]) * oq, __uint_as_float(t[i + 2]) * oq, __uint_as_float(t[i + 3]) * oq);
            v.y = (int)fp8x4_out(__uint_as_float(t[i + 4]) * oq, __uint_as_float(t[i + 5]) * oq, __uint_as_float(t[i + 6]) * oq, __uint_as_float(t[i + 7]) * oq);
            *reinterpret_cast<int2*>(orow + c * 32 + i) = v;
          }
        } else if (row_valid) {
#pragma unroll
          for (int i = 0; i < 32; i += 8) {
            int4 v;
            v.x = (int)pack2<kBf16>(__uint_as_float(t[i]) * inv, __uint_as_float(t[i + 1]) * inv);
            v.y = (int)pack2<kBf16>(__uint_as_float(t[i + 2]) * inv, __uint_as_float(t[i + 3]) * inv);
            v.z = (int)pack2<kBf16>(__uint_as_float(t[i + 4]) * inv, __uint_as_float(t[i + 5]) * inv);
            v.w = (int)pack2<kBf16>(__uint_as_float(t[i + 6]) * inv, __uint_as_float(t[i + 7]) * inv);
            *reinterpret_cast<int4*>(orow + (c * 32 + i) * 2) = v;
          }
        }
      }
      tc_fence_before();
      t_run += (uint32_t)n_tiles;
      ++n_items;
    }
  } else {
    asm volatile("setmaxnreg.dec.sync.aligned.u32 80;");
    if (warp == 11) {
      // ============================ scheduler warp ==========================================
      // Decodes a grid coordinate into a work item exactly as a one-tile CTA used to (the
      // reference's q-block -> sequence binary search, triton_unified_attention.py:32-52),
      // publishes it, and asks the hardware for the next not-yet-launched CTA.
      // grid = (kv heads, q tiles): CTAs are launched x-fastest, so all heads of the heaviest
      // (latest) q tile go first and the lightest tiles form the tail
      int bx = (int)blockIdx.x, by = (int)blockIdx.y;
      uint32_t n_pub = 0, clc_phase = 0;
      while (true) {
        const int q_tile_global = (int)gridDim.y - 1 - by;
        int left = 0, right = a.num_seqs;
        while (left < right) {
          const int mid = (left + right) >> 1;
          const int mid_val = a.cu_seqlens_q[mid] / BQ + mid;
          if (mid_val <= q_tile_global) left = mid + 1; else right = mid;
        }
        Item it;
        it.valid = 1, it.b = left - 1, it.kvh = bx, it.n_tiles = 0;
        bool zero_rows = false;
        if (it.b >= 0) {
          it.q_start = a.cu_seqlens_q[it.b];
          it.q_len = a.cu_seqlens_q[it.b + 1] - it.q_start;
          const int q_tile_local = q_tile_global - (it.q_start / BQ + it.b);
          it.qpos0 = q_tile_local * BQ;
          // sequences outside [min_q_len, max_q_len] belong to another launch of a mixed batch
          if (it.qpos0 < it.q_len && it.q_len >= P.min_q_len && it.q_len <= P.max_q_len &&
              (!P.staged_base || P.staged_base[it.b] >= 0)) {
            const int seq_len = a.seq_lens[it.b];
            it.ctx = seq_len - it.q_len;
            const int qpos_last = min(it.qpos0 + BQ, it.q_len) - 1;
            it.kv_limit = min(it.ctx + qpos_last + 1, seq_len);
            const int kv_first = window > 0 ? max(0, it.ctx + it.qpos0 - window + 1) : 0;
            it.tile0 = kv_first / kBN;
            it.n_tiles = (it.kv_limit + kBN - 1) / kBN - it.tile0;
            zero_rows = it.n_tiles <= 0;
          }
        }
        if (zero_rows) {
          // nothing visible (seq_len <= 0): rows get zeros like the reference's reduce (:828)
          for (int r = lane; r < 256; r += 32) {
            const int qpos = it.qpos0 + r / gp;
            if (qpos < it.q_len && r % gp < g) {
              if (a.lse) a.lse[(int64_t)(it.q_start + qpos) * a.lse_stride0 + it.kvh * g + r % gp] = -INFINITY;
              const int osz = a.out_scale ? 1 : 2;
              int2* orow = reinterpret_cast<int2*>(static_cast<char*>(a.out) +
                                                   ((int64_t)(it.q_start + qpos) * a.out_stride0 + (int64_t)(it.kvh * g + r % gp) * a.out_stride1) * osz);
              for (int i = 0; i < kD * osz / 8; ++i) orow[i] = make_int2(0, 0);
            }
          }
        } else if (it.n_tiles > 0) {
          const uint32_t slot = n_pub & 1;
          mbar_wait(&bars[ITEM_EMPTY + slot], ((n_pub >> 1) & 1) ^ 1);
          if (lane == 0) items[slot] = it;
          __syncwarp();
          if (lane == 0) mbar_arrive(&bars[ITEM_FULL + slot]);
          // Do not reserve more work until every role has started on this item: a CTA then
          // holds its current tile plus at most one queued tile (whose pipeline fill overlaps
          // the current one's drain), which keeps the tail of the grid balanced.
          mbar_wait(&bars[ITEM_EMPTY + slot], (n_pub >> 1) & 1);
          ++n_pub;
        }
        if (lane == 0) {
          mbar_arrive_expect_tx(&bars[CLC_DONE], 16);
          asm volatile("clusterlaunchcontrol.try_cancel.async.shared::cta.mbarrier::complete_tx::bytes.b128 [%0], [%1];"
                       ::"r"(smem_u32(clc_resp)), "r"(smem_u32(&bars[CLC_DONE])) : "memory");
        }
        mbar_wait(&bars[CLC_DONE], clc_phase);
        clc_phase ^= 1;
        uint32_t got, nx, ny;
        asm volatile(
            "{\n\t.reg .b128 r;\n\t.reg .pred p;\n\t"
            "ld.shared.b128 r, [%3];\n\t"
            "clusterlaunchcontrol.query_cancel.is_canceled.pred.b128 p, r;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t"
            "@p clusterlaunchcontrol.query_cancel.get_first_ctaid.v4.b32.b128 {%1, %2, _, _}, r;\n\t}"
            : "=r"(got), "=r"(nx), "=r"(ny)
            : "r"(smem_u32(clc_resp))
            : "memory");
        if (!got) break;
        bx = (int)nx, by = (int)ny;
      }
      {  // tell the other roles to stop
        const uint32_t slot = n_pub & 1;
        mbar_wait(&bars[ITEM_EMPTY + slot], ((n_pub >> 1) & 1) ^ 1);
        if (lane == 0) items[slot].valid = 0;
        __syncwarp();
        if (lane == 0) mbar_arrive(&bars[ITEM_FULL + slot]);
      }
    } else if (warp == 8) {
      Item it;
      while (next_item(it)) {
        B200ATTN_ITEM_VARS
        // ============================ MMA issue warp =========================================
        // The whole warp walks the (warp-uniform) control flow and one elected lane issues, so
        // descriptors live in uniform registers; every descriptor is a precomputed base plus a
        // compile-time offset in the 14-bit address field.  (A `lane == 0` version cost ~130
        // cycles per MMA in scalar address arithmetic and R2UR loops - twice the MMA itself.)
        constexpr uint32_t idesc_qk = umma_idesc(kBf16, false, 128, 128);
        constexpr uint32_t idesc_pv = umma_idesc(kBf16, true, 128, kD);
        auto issue_qk = [&](int w, int st) {
          const uint64_t ad = q_desc0 + (uint64_t)((w * kTileBytes) >> 4), bd = k_desc0 + (uint64_t)((st * kTileBytes) >> 4);
          if (leader) {
#pragma unroll
            for (int kk = 0; kk < kD / 16; ++kk) {
              const uint64_t off = (uint64_t)(((kk >> 2) * kSubtileBytes + (kk & 3) * 32) >> 4);
              umma_ss(tm + col_s(w), ad + off, bd + off, idesc_qk, kk > 0 ? 1u : 0u);
            }
          }
        };
        // PV for subtile w: k-steps 0..3 as soon as the first half of P is in TMEM, 4..7 after the second
        auto issue_pv = [&](int w, int st, bool acc, uint32_t par, uint32_t t) {
          const uint64_t bd = v_desc0 + (uint64_t)((st * kTileBytes) >> 4);
#pragma unroll
          for (int h = 0; h < 2; ++h) {
            mbar_wait(&bars[(h ? P_FULLB : P_FULL) + w], par);
            tc_fence_after();
            if (leader) {
              TC_TRACE(t, 10 + w * 3 + h);
#pragma unroll
              for (int kk = 4 * h; kk < 4 * h + 4; ++kk)
                umma_ts(tm + col_o<kD>(w), tm + col_s(w) + kk * 8, bd + (uint64_t)((kk * 2048) >> 4), idesc_pv, (acc || kk > 0) ? 1u : 0u);
            }
          }
        };
        // S(0) of this item: needs Q and the first K tile; S is free (its last reader, the
        // previous item's final PV, is ahead of these MMAs on the in-order tensor pipe)
        mbar_wait(&bars[Q_FULL], (n_items & 1));
        mbar_wait(&bars[K_FULL + k_st], k_ph);
        tc_fence_after();
        issue_qk(0, k_st);
        if (leader) umma_commit(&bars[S_FULL + 0]);
        issue_qk(1, k_st);
        if (leader) {
          umma_commit(&bars[S_FULL + 1]);
          umma_commit(&bars[K_EMPTY + k_st]);
          if (n_tiles == 1) umma_commit(&bars[Q_EMPTY]);
        }
        if (++k_st == kKStages) k_st = 0, k_ph ^= 1;
        for (int j = 0; j < n_tiles; ++j) {
          const uint32_t t = t_run + (uint32_t)j, par = t & 1;
          const bool more = j + 1 < n_tiles;
          // operand tiles first: they arrive long before P does, so these waits are off the
          // critical path P(j) -> PV(j), QK(j+1) -> S(j+1)
          mbar_wait(&bars[V_FULL + v_st], v_ph);
          if (j == last_tile && need_fix) {
            mbar_wait(&bars[V_FIX], n_fix & 1);
            ++n_fix;
          }
          if (more) mbar_wait(&bars[K_FULL + k_st], k_ph);
          issue_pv(0, v_st, j > 0, par, t);
          if (leader) umma_commit(&bars[O_DONE + 0]);
          if (more) {
            issue_qk(0, k_st);
            if (leader) {
              umma_commit(&bars[S_FULL + 0]);
              TC_TRACE(t, 12);
            }
          }
          issue_pv(1, v_st, j > 0, par, t);
          if (leader) {
            umma_commit(&bars[O_DONE + 1]);
            umma_commit(&bars[V_EMPTY + v_st]);
          }
          if (more) {
            issue_qk(1, k_st);
            if (leader) {
              umma_commit(&bars[S_FULL + 1]);
              umma_commit(&bars[K_EMPTY + k_st]);
              if (j + 2 == n_tiles) umma_commit(&bars[Q_EMPTY]);  // last QK of the item: Q may be replaced
              TC_TRACE(t, 15);
            }
            if (++k_st == kKStages) k_st = 0, k_ph ^= 1;
          }
          if (++v_st == kVStages) v_st = 0, v_ph ^= 1;
        }
        __syncwarp();
        t_run += (uint32_t)n_tiles;
        ++n_items;
      }
    } else {
      Item it;
      while (next_item(it)) {
        B200ATTN_ITEM_VARS
        // ============================ TMA warps: 9 = Q and K, 10 = V =========================
        // Separate warps so a full V ring never holds back the K prefetch (K of tile j+1 is on the
        // critical path right after PV(j); a K tile takes ~2000 cycles from request to arrival).
        const bool is_k = warp == 9;
        if (is_k) {
          // Q: the previous item's QK MMAs must be done with the buffer
          mbar_wait(&bars[Q_EMPTY], (n_items & 1) ^ 1);
          if (lane == 0) mbar_arrive_expect_tx(&bars[Q_FULL], 2 * kTileBytes);
          __syncwarp();
          if (lane < 2 * kNDC) {
            const int w = lane / kNDC, dc = lane % kNDC;
            tma_load_4d(sQ + w * kTileBytes + dc * kSubtileBytes, &q_map, &bars[Q_FULL], 0, dc, kvh * g,
                        it.q_start + it.qpos0 + w * P.block_q);
          }
        }
        const int bs = a.block_size;
        const int boxes_per_tile = kBN / P.box_slots;
        const uint32_t box_bytes = (uint32_t)P.box_slots * 128;
        const int32_t* bt = a.block_tables + (int64_t)it.b * a.bt_stride0;
        const int staged0 = (P.staged_base && n_tiles > 0) ? P.staged_base[it.b] : -1;
        const int my_box = lane / kNDC, my_dc = lane % kNDC;
        // page id of this lane's box, fetched one tile ahead of its use
        auto lookup = [&](int j) {
          const int tok = (tile0 + j) * kBN + my_box * P.box_slots;
          if (!(j < n_tiles && my_box < boxes_per_tile && tok < it.kv_limit)) return 0;
          return staged0 >= 0 ? staged0 + tok / bs : bt[tok / bs];
        };
        const int nst = is_k ? kKStages : kVStages;
        const int full0 = is_k ? K_FULL : V_FULL, empty0 = is_k ? K_EMPTY : V_EMPTY;
        uint8_t* ring = is_k ? sK : sV;
        const CUtensorMap* map = is_k ? &k_map : &v_map;
        int& st = is_k ? k_st : v_st;
        uint32_t& ph = is_k ? k_ph : v_ph;
        int page_next = lookup(0);
        int last_st = 0;
        uint32_t last_ph = 0;
        for (int j = 0; j < n_tiles; ++j) {
          const int n0 = (tile0 + j) * kBN;
          const int nbox = min(boxes_per_tile, (it.kv_limit - n0 + P.box_slots - 1) / P.box_slots);
          const bool active = my_box < nbox;
          const int page = page_next;
          page_next = lookup(j + 1);
          const int slot0 = (n0 + my_box * P.box_slots) % bs;
          const uint32_t dst_off = (uint32_t)(st * kTileBytes + my_dc * kSubtileBytes) + (uint32_t)my_box * box_bytes;
          mbar_wait(&bars[empty0 + st], ph ^ 1);
          if (lane == 0) TC_TRACE(t_run + j, is_k ? 19 : 20);
          if (lane == 0) mbar_arrive_expect_tx(&bars[full0 + st], (uint32_t)(nbox * kNDC) * box_bytes);
          __syncwarp();
          if (active) tma_load_5d_nohint(ring + dst_off, map, &bars[full0 + st], 0, slot0, my_dc, kvh, page);
          last_st = st, last_ph = ph;
          if (++st == nst) st = 0, ph ^= 1;
        }
        // ---- V tail fix-up -------------------------------------------------------------------
        // Rows of the last tile beyond the visible keys are either unwritten cache slots or
        // shared memory no copy touched; P is 0 there, but 0 x NaN would poison O.  The last
        // tile's copy was issued just above, so its V_FULL phase is the one in flight.
        if (!is_k && need_fix) {
          mbar_wait(&bars[V_FULL + last_st], last_ph);
          uint8_t* vt = sV + last_st * kTileBytes;
          const int n16 = (kBN - last_valid) * 8 * kNDC;  // 16-byte pieces: kNDC chunks x 8 per row
          for (int idx = lane; idx < n16; idx += 32) {
            const int r = last_valid + idx / (8 * kNDC), dc = (idx >> 3) % kNDC, c = idx & 7;
            *reinterpret_cast<int4*>(vt + dc * kSubtileBytes + r * 128 + c * 16) = make_int4(0, 0, 0, 0);
          }
          fence_proxy_async();
          __syncwarp();
          if (lane == 0) mbar_arrive(&bars[V_FIX]);
        }
        t_run += (uint32_t)n_tiles;
        ++n_items;
      }
    }
  }
#undef B200ATTN_ITEM_VARS

  // ---- teardown ---------------------------------------------------------------------------
  tc_fence_before();
  __syncthreads();
#ifdef B200ATTN_TC_TRACE
  if (threadIdx.x == 0) {
    const int cta = blockIdx.y * gridDim.x + blockIdx.x;
    if (cta < 4096) {
      unsigned smid;
      asm volatile("mov.u32 %0, %smid;" : "=r"(smid));
      g_tc_cta[cta * 4 + 0] = smid, g_tc_cta[cta * 4 + 1] = tc_t0, g_tc_cta[cta * 4 + 2] = tc_globaltimer(), g_tc_cta[cta * 4 + 3] = tc_tiles;
    }
  }
#endif
  if (warp == 11) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

inline bool pow2(int x) { return x > 0 && (x & (x - 1)) == 0; }

}  // namespace

int prefill_tc_min_q_len(const b200attn_attention_params& p) {
  const int g = p.num_q_heads / (p.num_kv_heads > 0 ? p.num_kv_heads : 1);
  const int v = 64 / (g > 0 ? g : 1) + 1;  // more rows than one mma.sync tile (64) holds
  return v < 2 ? 2 : v;
}

bool prefill_tc_supported(const b200attn_attention_params& p, const char** why) {
  auto no = [&](const char* w) { if (why) *why = w; return false; };
  if (p.head_size == 256) return prefill_tc256_supported(p, why);
  if (device_info().cc_major < 10) return no("needs sm_100a");
  if (p.head_size != 64 && p.head_size != 128) return no("head_size not 64, 128 or 256");
  if (p.q_dtype != B200ATTN_BF16 && p.q_dtype != B200ATTN_F16) return no("query dtype must be bf16/f16");
  if (p.kv_dtype != p.q_dtype) return no("kv dtype must equal the query dtype");
  if (p.num_kv_heads <= 0 || p.num_q_heads % p.num_kv_heads) return no("heads");
  const int g = p.num_q_heads / p.num_kv_heads;
  if (g > 128) return no("GQA group > 128");  // (any other group: the Q tile pads it to the next power of two)
  if (!pow2(p.block_size) || p.block_size < 16) return no("block_size not a power of two >= 16");
  if (p.softcap < 0.f) return no("softcap < 0");
  if (!(p.softmax_scale > 0.f)) return no("softmax_scale <= 0");
  auto m8 = [&](int64_t s) { return s % 8 == 0; };
  if (!m8(p.kc_block_stride) || !m8(p.kc_slot_stride) || !m8(p.kc_head_stride) || !m8(p.vc_block_stride) ||
      !m8(p.vc_slot_stride) || !m8(p.vc_head_stride))
    return no("cache strides not 16-byte multiples");
  if ((reinterpret_cast<uintptr_t>(p.k_cache) & 15) || (reinterpret_cast<uintptr_t>(p.v_cache) & 15)) return no("cache base alignment");
  if ((reinterpret_cast<uintptr_t>(p.q) & 15) || !m8(p.q_stride0) || !m8(p.q_stride1)) return no("query alignment");
  if ((reinterpret_cast<uintptr_t>(p.out) & 15) || !m8(p.out_stride0) || !m8(p.out_stride1)) return no("output alignment");
  return true;
}

int launch_attn_prefill_tc(const b200attn_attention_params& a, cudaStream_t stream, QLenRange range,
                           const int32_t* staged_base) {
  const char* why = nullptr;
  if (a.head_size == 256) return launch_attn_prefill_tc256(a, stream, range, staged_base);
  if (!prefill_tc_supported(a, &why)) {
    set_error("prefill_tc: unsupported configuration (%s)", why);
    return B200ATTN_E_UNSUPPORTED;
  }
  TcParams P;
  P.a = a;
  P.group = a.num_q_heads / a.num_kv_heads;
  for (P.group_pad = 1; P.group_pad < P.group; P.group_pad *= 2) {}
  P.block_q = 128 / P.group_pad;
  P.min_q_len = range.min_q_len, P.max_q_len = range.max_q_len;
  P.staged_base = staged_base;
  P.box_slots = a.block_size < kBN ? a.block_size : kBN;
  const int D = a.head_size;
  CUtensorMap qmap, kmap, vmap;
  int rc = make_thd_map(&qmap, a.q, D, a.num_q_heads, a.num_tokens, a.q_stride0, a.q_stride1, P.group_pad, P.block_q);
  if (rc) return rc;
  rc = make_kv_map(&kmap, a.k_cache, 2, D, a.block_size, a.num_kv_heads, a.num_blocks, a.kc_block_stride, a.kc_slot_stride,
                   a.kc_head_stride, P.box_slots, 1, 1);
  if (rc) return rc;
  rc = make_kv_map(&vmap, a.v_cache, 2, D, a.block_size, a.num_kv_heads, a.num_blocks, a.vc_block_stride, a.vc_slot_stride,
                   a.vc_head_stride, P.box_slots, 1, 1);
  if (rc) return rc;
  const bool bf16 = a.q_dtype == B200ATTN_BF16;
  // fraction of the exponentials taken off the MUFU (tuning().reserved bits 4..7; 0 = default)
  using Kern = void (*)(const TcParams, const CUtensorMap, const CUtensorMap, const CUtensorMap);
#define TCV(DD, B) {attn_prefill_tc_kernel<DD, B, 0, 1>, attn_prefill_tc_kernel<DD, B, 0, 1>, attn_prefill_tc_kernel<DD, B, 1, 4>, \
                    attn_prefill_tc_kernel<DD, B, 1, 3>, attn_prefill_tc_kernel<DD, B, 1, 2>, attn_prefill_tc_kernel<DD, B, 0, 1, 1>, \
                    attn_prefill_tc_kernel<DD, B, 1, 2, 1>, attn_prefill_tc_kernel<DD, B, 0, 1, 2>}
  // softcap alone (variants 5, 6): the tanh doubles the MUFU work; moving half the exps to the FMA pipe (6) measured
  // slower than leaving them (5): 8k prompt 0.705 ms vs 0.671 ms (profiles/r01_kbench_softcap_tc.jsonl).
  // ALiBi, with or without softcap: variant 7.
  constexpr int kNumVariants = 8, kSoftcapVariant = 5, kGeneralVariant = 7;
  static const Kern kVariants[2][2][kNumVariants] = {{TCV(64, false), TCV(64, true)}, {TCV(128, false), TCV(128, true)}};
#undef TCV
  int variant = (tuning().reserved >> 4) & 15;
  if (variant >= kSoftcapVariant) variant = 0;
  // D = 64 halves the MMA work per exponential: the MUFU (16 ex2/clk/SM) then needs twice the tensor pipe's time for a
  // tile, so a third of the exponentials go to the FMA pipe by default (ex2_poly2).  8k prompt, 64 q heads:
  // none 772, 1/4 825, 1/3 839, 1/2 702 TFLOP/s (profiles/r02_kbench_tc_d64.jsonl)
  if (variant == 0 && D == 64) variant = 3;
  if (a.alibi_slopes || a.qq_bias) variant = kGeneralVariant;
  else if (a.softcap > 0.f) variant = (((tuning().reserved >> 4) & 15) == 6 || D == 64) ? 6 : kSoftcapVariant;
  Kern kern = kVariants[D == 128 ? 1 : 0][bf16 ? 1 : 0][variant];
  const int smem_bytes = D == 128 ? Smem<128>::kBytes : Smem<64>::kBytes;
  if (int rc2 = ensure_dynamic_smem(reinterpret_cast<const void*>(kern), smem_bytes, "prefill_tc")) return rc2;
  const int BQ = 2 * P.block_q;
  const dim3 grid((unsigned)a.num_kv_heads, (unsigned)(a.num_tokens / BQ + a.num_seqs));
  kern<<<grid, kThreads, smem_bytes, stream>>>(P, qmap, kmap, vmap);
  note_launch("attn_prefill_tc");
  return check_launch("attn_prefill_tc");
}

}  // namespace b200attn
