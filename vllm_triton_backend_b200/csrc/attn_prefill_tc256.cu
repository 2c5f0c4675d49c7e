// K3c: tcgen05 / TMEM flash-attention forward over the paged KV cache for head size 256
// (Gemma-class models), bf16/f16 queries and KV.
//
// Same computation, row order, masks and grid as attn_prefill_tc.cu (reference
// ibm-triton-lib/ibm_triton_lib/kernels/triton_unified_attention.py:275-523, whose one kernel serves every head
// size of backend/triton_attn.py:241-242), but the 512 TMEM columns do not hold two 128-row subtiles at this head
// size (S0 | S1 | O0 | O1 would be 128 + 128 + 256 + 256), so the tile is cut the other way:
//
//   * one CTA = ONE 128-row Q subtile x one kv head; KV tiles of 64 tokens;
//   * TMEM: S0 | S1 (64 columns each, two KV tiles in flight) | O (256 columns);
//   * the two softmax warpgroups take alternate KV tiles of the SAME rows: warpgroup w owns S_w, so the
//     exponentials of tile j run under the QK^T of tile j + 1 and the PV of tile j - 1.  The rows' running
//     maximum is the only thing tile j needs from tile j - 1: it lives in shared memory and is handed from
//     warpgroup to warpgroup through one mbarrier (M_READY) right after the owner of tile j - 1 has taken its
//     row maxima - before its exponentials.  Each warpgroup keeps its own partial exp-sum (relative to the
//     maximum it last saw) and the two are put together in the epilogue;
//   * MMA order: QK(0) QK(1) | PV(j) QK(j + 2) ...; QK: M128 N64 K256 (16 instructions), PV: M128 N256 K64 (4, P read
//     from TMEM as in the other kernel);
//   * shared memory: Q 64 KB + K ring 3 x 32 KB + V ring 2 x 32 KB = 224 KB.
#include "tc_common.cuh"

namespace b200attn {
namespace {

constexpr int kD = 256;
constexpr int kBN = 64;                          // kv tokens per tile
constexpr int kNDC = kD / 64;                    // 128-byte chunks of a row
constexpr int kKStages = 3;
constexpr int kVStages = 2;
constexpr int kQChunk = 128 * 128;               // [128 rows][128 B]
constexpr int kQBytes = kNDC * kQChunk;          // 64 KB
constexpr int kKVChunk = kBN * 128;              // [64 tokens][128 B]
constexpr int kKVTile = kNDC * kKVChunk;         // 32 KB
constexpr int kThreads = 384;
constexpr uint32_t kColO = 256;                  // TMEM: S_w at 64 w, O at 256
constexpr float kRescaleThreshold = 8.0f;        // log2 units

constexpr int kSmemQ = 0;
constexpr int kSmemK = kQBytes;
constexpr int kSmemV = kSmemK + kKStages * kKVTile;
constexpr int kSmemBar = kSmemV + kVStages * kKVTile;  // 224 KB
enum Bar {
  Q_FULL = 0, Q_EMPTY = 1, K_FULL = 2, K_EMPTY = 5, V_FULL = 8, V_EMPTY = 10, S_FULL = 12, P_FULL = 14, O_DONE = 16, M_READY = 17,
  V_FIX = 18, ITEM_FULL = 19, ITEM_EMPTY = 21, CLC_DONE = 23, O_FINAL = 24, NUM_BARS = 25
};
constexpr int kBarBytes = 256;                   // NUM_BARS * 8 = 200, tmem pointer behind them
constexpr int kSmemItems = kSmemBar + kBarBytes; // 2 x Item
constexpr int kSmemClc = kSmemItems + 2 * 64;    // 16-byte try_cancel response
constexpr int kSmemM = kSmemClc + 64;            // float m_sh[128]: the rows' reference maximum (log2 units)
constexpr int kSmemML = kSmemM + 512;            // float2 ml_sh[2][128]: each warpgroup's (max it is relative to, exp-sum)
constexpr int kSmemBytes = kSmemML + 2048;       // = 232 384 of the 232 448: the base must be 1024-byte aligned (trap)
constexpr int kConsumerWarps = 11;               // warps 0..10 read every item

// kGeneral: logits are formed in place in log2 units (softcap, ALiBi, qq_bias at run time) and the max / exp2 code runs
// with a scale of 1; otherwise S * scale rides in the exp2 FFMA.
template <bool kBf16, bool kGeneral>
__global__ void __launch_bounds__(kThreads, 1)
attn_prefill_tc256_kernel(const TcParams P, const __grid_constant__ CUtensorMap q_map, const __grid_constant__ CUtensorMap k_map,
                          const __grid_constant__ CUtensorMap v_map) {
  extern __shared__ __align__(1024) uint8_t smem[];
  if (smem_u32(smem) & 1023u) __trap();  // (no room for re-alignment slack; the declared alignment is honoured)
  uint8_t* sQ = smem + kSmemQ;
  uint8_t* sK = smem + kSmemK;
  uint8_t* sV = smem + kSmemV;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kSmemBar);
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(bars + NUM_BARS);
  Item* items = reinterpret_cast<Item*>(smem + kSmemItems);
  uint4* clc_resp = reinterpret_cast<uint4*>(smem + kSmemClc);
  float* m_sh = reinterpret_cast<float*>(smem + kSmemM);
  float2* ml_sh = reinterpret_cast<float2*>(smem + kSmemML);

  const b200attn_attention_params& a = P.a;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int g = P.group, gp = P.group_pad;
  const int BQ = P.block_q;  // query tokens per CTA tile (one 128-row subtile)
  const int window = a.sliding_window;

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
      mbar_init(&bars[ITEM_FULL + s], 1);
      mbar_init(&bars[ITEM_EMPTY + s], kConsumerWarps);
    }
    mbar_init(&bars[O_DONE], 1);
    mbar_init(&bars[O_FINAL], 1);
    mbar_init(&bars[M_READY], 128);
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
  uint32_t t_run = 0;     // KV tiles processed so far, all items: parity source of O_DONE / M_READY
  uint32_t tw[2] = {0, 0};  // tiles handled so far by warpgroup 0 / 1: parity source of S_FULL / P_FULL
  const float scale_log2 = kGeneral ? 1.0f : a.softmax_scale * kLog2e;
  int k_st = 0, v_st = 0;
  uint32_t k_ph = 0, v_ph = 0;
  uint32_t n_fix = 0;
  const bool leader = elect_one();
  const uint32_t tm = __shfl_sync(0xffffffffu, tmem_base, 0);
  const uint64_t q_desc0 = umma_desc(__shfl_sync(0xffffffffu, smem_u32(sQ), 0), 16, 1024);
  const uint64_t k_desc0 = umma_desc(__shfl_sync(0xffffffffu, smem_u32(sK), 0), 16, 1024);
  const uint64_t v_desc0 = umma_desc(__shfl_sync(0xffffffffu, smem_u32(sV), 0), kKVChunk, 1024);
  auto next_item = [&](Item& it) -> bool {
    const uint32_t slot = n_items & 1;
    mbar_wait(&bars[ITEM_FULL + slot], (n_items >> 1) & 1);
    it = items[slot];
    __syncwarp();
    if (lane == 0) mbar_arrive(&bars[ITEM_EMPTY + slot]);
    return it.valid != 0;
  };
#define B200ATTN_ITEM_VARS                                                                              \
  const int n_tiles = it.n_tiles, tile0 = it.tile0, kvh = it.kvh;                                       \
  const int last_tile = n_tiles - 1;                                                                    \
  const int last_valid = it.kv_limit - (tile0 + last_tile) * kBN; /* visible rows of the last tile */ \
  const bool need_fix = last_valid < kBN;                                                               \
  (void)kvh, (void)need_fix, (void)last_valid;
  // tiles of an item that warpgroup w takes: j = w, w + 2, ...
  auto tiles_of = [](int n, int w) { return (uint32_t)((n + 1 - w) >> 1); };

  if (warp < 8) {
    asm volatile("setmaxnreg.inc.sync.aligned.u32 208;");
    Item it;
    while (next_item(it)) {
      B200ATTN_ITEM_VARS
      // ============================ softmax warpgroups ====================================
      const int w = warp >> 2;                  // owns S_w: KV tiles j = w, w + 2, ...
      const int row = (warp & 3) * 32 + lane;   // row of the subtile == TMEM lane
      const int qpos = it.qpos0 + row / gp;
      const int head = row % gp;                // heads >= g are padding rows
      const bool row_valid = qpos < it.q_len && head < g;
      const int limit = row_valid ? it.ctx + qpos + 1 : 0;                        // visible keys: [lo, limit)
      const int lo = (window > 0 && row_valid) ? max(0, it.ctx + qpos - window + 1) : 0;
      const uint32_t lane_base = (uint32_t)((warp & 3) * 32) << 16;
      const uint32_t s_tmem = tmem_base + lane_base + 64u * (uint32_t)w;
      const uint32_t o_tmem = tmem_base + lane_base + kColO;
      float m_mine = -INFINITY, l_sum = 0.f;  // this warpgroup's exp-sum and the maximum it is relative to

      for (int j = w; j < n_tiles; j += 2) {
        const uint32_t t = t_run + (uint32_t)j;  // global index of the tile
        mbar_wait(&bars[S_FULL + w], tw[w] & 1);
        tc_fence_after();
        uint32_t s[2][32];
        tmem_ld32(s_tmem, s[0]);
        tmem_ld32(s_tmem + 32, s[1]);
        tmem_ld_wait();
        const int n0 = (tile0 + j) * kBN;
        if constexpr (kGeneral) {
          if (a.softcap > 0.f) {
            const float pre = a.softmax_scale / a.softcap, cap_log2 = a.softcap * kLog2e;
#pragma unroll
            for (int c = 0; c < 2; ++c)
#pragma unroll
              for (int i = 0; i < 32; ++i) s[c][i] = __float_as_uint(cap_log2 * tanh_approx(__uint_as_float(s[c][i]) * pre));
          } else {
            const float sl2 = a.softmax_scale * kLog2e;
#pragma unroll
            for (int c = 0; c < 2; ++c)
#pragma unroll
              for (int i = 0; i < 32; ++i) s[c][i] = __float_as_uint(__uint_as_float(s[c][i]) * sl2);
          }
          if (a.alibi_slopes) {
            const float slope = (head < g ? a.alibi_slopes[kvh * g + head] : 0.f) * kLog2e;
            const float b0 = slope * (float)(n0 - it.ctx);
#pragma unroll
            for (int c = 0; c < 2; ++c)
#pragma unroll
              for (int i = 0; i < 32; ++i) s[c][i] = __float_as_uint(fmaf(slope, (float)(c * 32 + i), __uint_as_float(s[c][i]) + b0));
          }
          if (a.qq_bias && row_valid && n0 + kBN > it.ctx) {
            const float* brow = a.qq_bias + (int64_t)qpos * a.qq_bias_stride0;
            const int nrel = (int)a.qq_bias_stride0;
#pragma unroll
            for (int c = 0; c < 2; ++c)
#pragma unroll
              for (int i = 0; i < 32; ++i) {
                const int rel = n0 + c * 32 + i - it.ctx;
                if (rel >= 0 && rel < nrel) s[c][i] = __float_as_uint(fmaf(brow[rel], kLog2e, __uint_as_float(s[c][i])));
              }
          }
        }
        if (n0 + kBN > limit || n0 < lo) {
#pragma unroll
          for (int c = 0; c < 2; ++c)
#pragma unroll
            for (int i = 0; i < 32; ++i) {
              const int pos = n0 + c * 32 + i;
              if (pos >= limit || pos < lo) s[c][i] = 0xff800000u;  // -inf
            }
        }
        float mxc[2] = {-INFINITY, -INFINITY};
#pragma unroll
        for (int c = 0; c < 2; ++c)
#pragma unroll
          for (int i = 0; i < 32; i += 2) mxc[c] = fmax3(mxc[c], __uint_as_float(s[c][i]), __uint_as_float(s[c][i + 1]));
        const float m_tile = fmaxf(mxc[0], mxc[1]) * scale_log2;  // softmax_scale > 0 (checked on the host)
        // ---- the rows' reference maximum comes from the owner of the previous tile
        if (t > 0) mbar_wait(&bars[M_READY], (t - 1) & 1);
        const float m_cur = (j == 0) ? -INFINITY : m_sh[row];
        float m_new = m_cur, alpha = 1.f;
        if (m_tile > m_cur + kRescaleThreshold) {  // also the first finite max (m_cur = -inf)
          alpha = ex2_approx(m_cur - m_tile);       // 0 when m_cur = -inf
          m_new = m_tile;
        }
        if (j > 0 && __any_sync(0xffffffffu, alpha != 1.f)) {
          // rare: rescale the running output (all 256 columns of this row) before PV(j) accumulates into it
          mbar_wait(&bars[O_DONE], (t - 1) & 1);
          tc_fence_after();
#pragma unroll 1
          for (int c = 0; c < kD / 8; ++c) {
            uint32_t x[8];
            tmem_ld8(o_tmem + c * 8, x);
            tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 8; ++i) x[i] = __float_as_uint(__uint_as_float(x[i]) * alpha);
            tmem_st8(o_tmem + c * 8, x);
          }
          tmem_st_wait();
          tc_fence_before();
        }
        m_sh[row] = m_new;
        mbar_arrive(&bars[M_READY]);  // release: the owner of tile j + 1 may go on
        // this warpgroup's exp-sum follows the reference maximum (it may have moved while the other one had it)
        if (m_new != m_mine) l_sum *= (m_mine == -INFINITY) ? 0.f : ex2_approx(m_mine - m_new);
        m_mine = m_new;
        const float m_eff = (m_new == -INFINITY) ? 0.f : m_new;
        // P = 2^(S*scale - m) replaces S in place (bf16 pairs: 32 columns of the S_w region)
        const uint64_t sc2 = f2_pack(scale_log2, scale_log2), nm2 = f2_pack(-m_eff, -m_eff);
        uint64_t sum2[2] = {0ull, 0ull};
        uint32_t pk[32];
#pragma unroll
        for (int c = 0; c < 2; ++c) {
#pragma unroll
          for (int i = 0; i < 32; i += 2) {
            const int k = c * 16 + (i >> 1);
            const uint64_t x2 = f2_fma(f2_pack(__uint_as_float(s[c][i]), __uint_as_float(s[c][i + 1])), sc2, nm2);
            float x0, x1;
            f2_unpack(x2, x0, x1);
            const float p0 = ex2_approx(x0), p1 = ex2_approx(x1);
            sum2[k & 1] = f2_add(sum2[k & 1], f2_pack(p0, p1));
            pk[k] = pack2<kBf16>(p0, p1);
          }
        }
        tmem_st32(s_tmem, pk);
        tmem_st_wait();
        tc_fence_before();
        mbar_arrive(&bars[P_FULL + w]);
        float sa, sb;
        f2_unpack(f2_add(sum2[0], sum2[1]), sa, sb);
        l_sum += sa + sb;
        ++tw[w];
      }

      // ---- epilogue: O / L -> global; each warpgroup stores half of the columns ---------------
      // O_FINAL completes once per item, after its last PV.  (O_DONE counts every PV: a warpgroup that does not own the
      // last tile would wait for a phase two ahead of the last one it knows to be complete, which a parity cannot tell
      // from the one before - the first version did, and read O early in CTAs that ran several two-tile items.)
      mbar_wait(&bars[O_FINAL], n_items & 1);
      tc_fence_after();
      ml_sh[w * 128 + row] = make_float2(m_mine, l_sum);
      asm volatile("bar.sync 2, 256;" ::: "memory");
      float m_f = m_sh[row];  // as the owner of the last tile left it
      float l_tot = 0.f;
#pragma unroll
      for (int ww = 0; ww < 2; ++ww) {
        const float2 ml = ml_sh[ww * 128 + row];
        if (ml.x != -INFINITY) l_tot += ml.y * ex2_approx(ml.x - m_f);
      }
      const bool seen = m_f != -INFINITY;  // some key was visible (a sink alone does not make an LSE)
      float fs = 1.f;
      if (a.sinks && row_valid) fs = sink_rescale<true>(a.sinks[kvh * g + head], m_f, l_tot);  // once per row, after the last tile
      const float inv = (l_tot > 0.f) ? fs / l_tot : 0.f;
      if (a.lse && row_valid && w == 0)
        a.lse[(int64_t)(it.q_start + qpos) * a.lse_stride0 + kvh * g + head] = seen ? row_lse<true>(m_f, l_tot) : -INFINITY;
      const bool out_fp8 = a.out_scale != nullptr;  // fused e4m3 output quantisation (b200attn.h: out_scale)
      const float oq = inv * out_quant_factor(a.out_scale);
      char* orow = static_cast<char*>(a.out) +
                   ((int64_t)(it.q_start + (row_valid ? qpos : 0)) * a.out_stride0 + (int64_t)(kvh * g + (row_valid ? head : 0)) * a.out_stride1) *
                       (out_fp8 ? 1 : 2);
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        const int col0 = w * 128 + c * 32;
        uint32_t x[32];
        tmem_ld32(o_tmem + col0, x);
        tmem_ld_wait();
        if (row_valid && out_fp8) {
#pragma unroll
          for (int i = 0; i < 32; i += 8) {
            int2 v;
            v.x = (int)fp8x4_out(__uint_as_float(x[i]) * oq, __uint_as_float(x[i + 1]) * oq, __uint_as_float(x[i + 2]) * oq, __uint_as_float(x[i + 3]) * oq);
            v.y = (int)fp8x4_out(__uint_as_float(x[i + 4]) * oq, __uint_as_float(x[i + 5]) * oq, __uint_as_float(x[i + 6]) * oq, __uint_as_float(x[i + 7]) * oq);
            *reinterpret_cast<int2*>(orow + col0 + i) = v;
          }
        } else if (row_valid) {
#pragma unroll
          for (int i = 0; i < 32; i += 8) {
            int4 v;
            v.x = (int)pack2<kBf16>(__uint_as_float(x[i]) * inv, __uint_as_float(x[i + 1]) * inv);
            v.y = (int)pack2<kBf16>(__uint_as_float(x[i + 2]) * inv, __uint_as_float(x[i + 3]) * inv);
            v.z = (int)pack2<kBf16>(__uint_as_float(x[i + 4]) * inv, __uint_as_float(x[i + 5]) * inv);
            v.w = (int)pack2<kBf16>(__uint_as_float(x[i + 6]) * inv, __uint_as_float(x[i + 7]) * inv);
            *reinterpret_cast<int4*>(orow + (col0 + i) * 2) = v;
          }
        }
      }
      tc_fence_before();
      // O and the exchange buffers are free for the next item only when BOTH warpgroups are done with them (the next
      // item's first PV waits for warpgroup 0's next P, which comes after this barrier)
      asm volatile("bar.sync 2, 256;" ::: "memory");
      t_run += (uint32_t)n_tiles;
      ++n_items;
    }
  } else {
    asm volatile("setmaxnreg.dec.sync.aligned.u32 80;");
    if (warp == 11) {
      // ============================ scheduler warp ==========================================
      // As in attn_prefill_tc.cu: decodes a grid coordinate into a work item (the reference's q-block -> sequence binary
      // search, triton_unified_attention.py:32-52), publishes it, and asks the hardware for the next CTA that has
      // not been launched yet (cluster launch control).
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
          for (int r = lane; r < 128; r += 32) {
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
          mbar_wait(&bars[ITEM_EMPTY + slot], (n_pub >> 1) & 1);  // every role has started on it before more work is reserved
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
        constexpr uint32_t idesc_qk = umma_idesc(kBf16, false, 128, kBN);
        constexpr uint32_t idesc_pv = umma_idesc(kBf16, true, 128, kD);
        auto issue_qk = [&](int w, int st) {  // S_w = Q K^T of the tile in K stage st
          const uint64_t bd = k_desc0 + (uint64_t)((st * kKVTile) >> 4);
          if (leader) {
#pragma unroll
            for (int kk = 0; kk < kD / 16; ++kk) {
              const uint64_t aoff = (uint64_t)(((kk >> 2) * kQChunk + (kk & 3) * 32) >> 4);
              const uint64_t boff = (uint64_t)(((kk >> 2) * kKVChunk + (kk & 3) * 32) >> 4);
              umma_ss(tm + 64u * (uint32_t)w, q_desc0 + aoff, bd + boff, idesc_qk, kk > 0 ? 1u : 0u);
            }
          }
        };
        auto issue_pv = [&](int w, int st, bool acc) {  // O (+)= P_w V of the tile in V stage st
          const uint64_t bd = v_desc0 + (uint64_t)((st * kKVTile) >> 4);
          if (leader) {
#pragma unroll
            for (int kk = 0; kk < kBN / 16; ++kk)
              umma_ts(tm + kColO, tm + 64u * (uint32_t)w + kk * 8, bd + (uint64_t)((kk * 2048) >> 4), idesc_pv, (acc || kk > 0) ? 1u : 0u);
          }
        };
        auto qk_tile = [&](int j) {  // tile j of the item: wait for its K, issue, release
          mbar_wait(&bars[K_FULL + k_st], k_ph);
          tc_fence_after();
          issue_qk(j & 1, k_st);
          if (leader) {
            umma_commit(&bars[S_FULL + (j & 1)]);
            umma_commit(&bars[K_EMPTY + k_st]);
            if (j == last_tile) umma_commit(&bars[Q_EMPTY]);  // last QK of the item: Q may be replaced
          }
          if (++k_st == kKStages) k_st = 0, k_ph ^= 1;
        };
        mbar_wait(&bars[Q_FULL], (n_items & 1));
        qk_tile(0);
        if (n_tiles > 1) qk_tile(1);
        for (int j = 0; j < n_tiles; ++j) {
          const int w = j & 1;
          mbar_wait(&bars[V_FULL + v_st], v_ph);
          if (j == last_tile && need_fix) {
            mbar_wait(&bars[V_FIX], n_fix & 1);
            ++n_fix;
          }
          mbar_wait(&bars[P_FULL + w], tw[w] & 1);
          tc_fence_after();
          issue_pv(w, v_st, j > 0);
          if (leader) {
            umma_commit(&bars[O_DONE]);
            umma_commit(&bars[V_EMPTY + v_st]);
            if (j == last_tile) umma_commit(&bars[O_FINAL]);
          }
          ++tw[w];
          if (++v_st == kVStages) v_st = 0, v_ph ^= 1;
          if (j + 2 < n_tiles) qk_tile(j + 2);  // into S_w: its last reader, PV(j), is ahead on the in-order tensor pipe
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
        const bool is_k = warp == 9;
        if (is_k) {
          mbar_wait(&bars[Q_EMPTY], (n_items & 1) ^ 1);
          if (lane == 0) mbar_arrive_expect_tx(&bars[Q_FULL], kQBytes);
          __syncwarp();
          if (lane < kNDC) tma_load_4d(sQ + lane * kQChunk, &q_map, &bars[Q_FULL], 0, lane, kvh * g, it.q_start + it.qpos0);
        }
        const int bs = a.block_size;
        const int boxes_per_tile = kBN / P.box_slots;
        const uint32_t box_bytes = (uint32_t)P.box_slots * 128;
        const int32_t* bt = a.block_tables + (int64_t)it.b * a.bt_stride0;
        const int staged0 = (P.staged_base && n_tiles > 0) ? P.staged_base[it.b] : -1;
        const int my_box = lane / kNDC, my_dc = lane % kNDC;
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
          const uint32_t dst_off = (uint32_t)(st * kKVTile + my_dc * kKVChunk) + (uint32_t)my_box * box_bytes;
          mbar_wait(&bars[empty0 + st], ph ^ 1);
          if (lane == 0) mbar_arrive_expect_tx(&bars[full0 + st], (uint32_t)(nbox * kNDC) * box_bytes);
          __syncwarp();
          if (active) tma_load_5d_nohint(ring + dst_off, map, &bars[full0 + st], 0, slot0, my_dc, kvh, page);
          last_st = st, last_ph = ph;
          if (++st == nst) st = 0, ph ^= 1;
        }
        // ---- V tail fix-up: rows of the last tile beyond the visible keys hold unwritten cache slots or stale shared
        // memory; P is 0 there, but 0 x NaN would poison O
        if (!is_k && need_fix) {
          mbar_wait(&bars[V_FULL + last_st], last_ph);
          uint8_t* vt = sV + last_st * kKVTile;
          const int n16 = (kBN - last_valid) * 8 * kNDC;  // 16-byte pieces: kNDC chunks x 8 per row
          for (int idx = lane; idx < n16; idx += 32) {
            const int r = last_valid + idx / (8 * kNDC), dc = (idx >> 3) % kNDC, c = idx & 7;
            *reinterpret_cast<int4*>(vt + dc * kKVChunk + r * 128 + c * 16) = make_int4(0, 0, 0, 0);
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
  if (warp == 11) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

}  // namespace

bool prefill_tc256_supported(const b200attn_attention_params& p, const char** why) {
  auto no = [&](const char* w) { if (why) *why = w; return false; };
  if (device_info().cc_major < 10) return no("needs sm_100a");
  if (p.head_size != kD) return no("head_size != 256");
  if (p.q_dtype != B200ATTN_BF16 && p.q_dtype != B200ATTN_F16) return no("query dtype must be bf16/f16");
  if (p.kv_dtype != p.q_dtype) return no("kv dtype must equal the query dtype");
  if (p.num_kv_heads <= 0 || p.num_q_heads % p.num_kv_heads) return no("heads");
  if (p.num_q_heads / p.num_kv_heads > 128) return no("GQA group > 128");
  const int bs = p.block_size;
  if (bs < 16 || (bs & (bs - 1))) return no("block_size not a power of two >= 16");
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

int launch_attn_prefill_tc256(const b200attn_attention_params& a, cudaStream_t stream, QLenRange range, const int32_t* staged_base) {
  const char* why = nullptr;
  if (!prefill_tc256_supported(a, &why)) {
    set_error("prefill_tc256: unsupported configuration (%s)", why);
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
  CUtensorMap qmap, kmap, vmap;
  int rc = make_thd_map(&qmap, a.q, kD, a.num_q_heads, a.num_tokens, a.q_stride0, a.q_stride1, P.group_pad, P.block_q);
  if (rc) return rc;
  rc = make_kv_map(&kmap, a.k_cache, 2, kD, a.block_size, a.num_kv_heads, a.num_blocks, a.kc_block_stride, a.kc_slot_stride,
                   a.kc_head_stride, P.box_slots, 1, 1);
  if (rc) return rc;
  rc = make_kv_map(&vmap, a.v_cache, 2, kD, a.block_size, a.num_kv_heads, a.num_blocks, a.vc_block_stride, a.vc_slot_stride,
                   a.vc_head_stride, P.box_slots, 1, 1);
  if (rc) return rc;
  const bool bf16 = a.q_dtype == B200ATTN_BF16;
  const bool general = a.softcap > 0.f || a.alibi_slopes || a.qq_bias;
  using Kern = void (*)(const TcParams, const CUtensorMap, const CUtensorMap, const CUtensorMap);
  static const Kern kVariants[2][2] = {{attn_prefill_tc256_kernel<false, false>, attn_prefill_tc256_kernel<false, true>},
                                       {attn_prefill_tc256_kernel<true, false>, attn_prefill_tc256_kernel<true, true>}};
  Kern kern = kVariants[bf16 ? 1 : 0][general ? 1 : 0];
  if (int rc2 = ensure_dynamic_smem(reinterpret_cast<const void*>(kern), kSmemBytes, "prefill_tc256")) return rc2;
  const dim3 grid((unsigned)a.num_kv_heads, (unsigned)(a.num_tokens / P.block_q + a.num_seqs));
  kern<<<grid, kThreads, kSmemBytes, stream>>>(P, qmap, kmap, vmap);
  note_launch("attn_prefill_tc256");
  return check_launch("attn_prefill_tc256");
}

}  // namespace b200attn
