// K2: split-KV, GQA-packed paged decode attention for sm_100a, plus the segment
// reduce (K2b).
//
// What it computes is what the reference's 3D path computes
// (reference ibm-triton-lib/ibm_triton_lib/kernels/triton_unified_attention.py:
//  kernel_unified_attention_3d :526-754, reduce_segments :757-836): for every
// sequence with one query token, softmax(scale * q K^T + mask/bias) V over the
// paged cache, partial results over KV ranges merged with the usual
// (max, exp-sum) rescaling.  How it does it is new:
//
//  * HBM-bound design.  KV pages are streamed by TMA (cp.async.bulk.tensor, 5-D
//    tiled map over the strided [NB, bs, Hkv, D] cache view, 128-byte swizzle)
//    into a multi-stage shared-memory ring; one producer warp issues the copies
//    from the block table, 8 (4 for D=256) consumer warps each own a 16-token x
//    one-kv-head tile of every stage.  (A cp.async producer variant for pages under
//    4 KB was measured and dropped: profiles/r01_kbench_cpasync_and_ctx_fit.jsonl.)
//  * Tokens sit on the MMA M axis, the GQA group's query heads on N (=8), so a
//    group of 4..8 heads costs one m16n8k16 per 16 tokens x 16 dims, K is read by
//    plain ldmatrix and V by ldmatrix.trans; P^T is produced from the S^T
//    accumulators with movmatrix.  FP8-e4m3 KV is dequantised in registers
//    (cvt.rn.f16x2.e4m3x2) and k_scale / v_scale are folded into the softmax
//    scale and the epilogue.
//  * Split-KV without host sync: the flattened (sequence, head-group, KV-step)
//    space is cut into equal contiguous ranges, one per CTA (grid = #SMs), from
//    the device-side seq_lens; a CTA that covers a whole item writes the output
//    directly, otherwise it writes an un-normalised partial (slot = item + cta is
//    collision free) and takes a ticket on the item's counter; the CTA that
//    takes the last ticket merges the item's partials and writes the output
//    (the reference's separate `reduce_segments` launch, :757-836, folded in).
//    Counters are self-cleaning: the merging CTA resets its item's counter, so
//    the workspace only has to be zeroed once, when it is allocated.
#include "common.cuh"

namespace b200attn {
namespace {

constexpr int kMaxStages = 8;

struct DecodeParams {
  b200attn_attention_params a;
  int hpc;         // kv heads per CTA (divides num_kv_heads and NW)
  int head_groups; // num_kv_heads / hpc
  int tps;         // tokens per stage-step = NW*TT/hpc (TT = tokens of a warp's tile); a power of two
  int tps_shift;   // log2(tps)
  int bs_shift;    // log2(block_size) (decode_supported() requires a power of two)
  int box_slots;   // min(block_size, tps)
  int boxes;       // tps / box_slots
  int stages;
  int tma_split;   // 0: one TMA op per (page box); 1: one per (page box, head, d-chunk)
  int paired;      // 1: one TMA box carries the K and the V rows of a page ([head][k|v][slot][128 B]); fp8 D=128
  int group;       // query heads per kv head
  int only_q1;     // mixed batch: only sequences with exactly one query token are this kernel's
  float* ws_acc;   // [slots][hpc*group][D]
  float* ws_ml;    // [slots][hpc*group][2]  (max in log2 units, exp-sum)
  int* ws_cnt;     // [items] tickets taken on an item split across CTAs (0 between calls)
  unsigned long long* trace;  // debug time stamps, 8 per CTA (b200attn_debug_set_decode_trace) or nullptr
};

__device__ __forceinline__ unsigned long long global_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

__device__ __forceinline__ float ex2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

struct ItemCursor {
  int b, hg, step;       // current sequence, head group, step index (absolute, in units of tps tokens)
  int first, n;          // first step and number of steps of sequence b
  int seq_len;
  int flat;              // flat index of the current step
};

__device__ __forceinline__ void seq_steps(const DecodeParams& P, int b, int& seq_len, int& first, int& n) {
  seq_len = P.a.seq_lens[b];
  int q_len = P.a.cu_seqlens_q ? (P.a.cu_seqlens_q[b + 1] - P.a.cu_seqlens_q[b]) : 1;
  if (seq_len <= 0 || q_len <= 0 || (P.only_q1 && q_len != 1)) {
    first = 0, n = 0;
    return;
  }
  int lo_tok = 0;
  if (P.a.sliding_window > 0) lo_tok = max(0, seq_len - P.a.sliding_window);
  first = lo_tok >> P.tps_shift;
  n = ((seq_len + P.tps - 1) >> P.tps_shift) - first;
}

// advance to the next step in (b, hg, step) order; returns false past the end
__device__ __forceinline__ bool cursor_next(const DecodeParams& P, ItemCursor& c) {
  c.flat++;
  if (++c.step < c.first + c.n) return true;
  if (++c.hg < P.head_groups) {
    c.step = c.first;
    return true;
  }
  c.hg = 0;
  while (++c.b < P.a.num_seqs) {
    seq_steps(P, c.b, c.seq_len, c.first, c.n);
    if (c.n > 0) {
      c.step = c.first;
      return true;
    }
  }
  return false;
}

template <int D, bool KV_FP8, bool MMA_BF16, bool Q_BF16, int NT, int NW, int TT>
__global__ void __launch_bounds__((NW + 1) * 32, 1)
attn_decode_kernel(const DecodeParams P, const __grid_constant__ CUtensorMap k_map,
                   const __grid_constant__ CUtensorMap v_map) {
  constexpr int ESZ = KV_FP8 ? 1 : 2;
  constexpr int NDC = D * ESZ / 128;      // 128-byte chunks per row
  constexpr int KSTEPS = D / 16;          // k16 steps of S = K Q^T
  constexpr int MTILES = D / 16;          // m16 tiles of O^T = V^T P^T
  constexpr int SUB = TT / 16;             // 16-token MMA sub-tiles of a warp's TT-token tile
  constexpr int TILE_BYTES = TT * D * ESZ;
  constexpr int KV_STAGE_BYTES = NW * TILE_BYTES;  // K (or V) bytes per stage
  static_assert(TT == 16 || TT == 32, "warp tile is one or two 16-token MMA tiles");
  static_assert(NDC >= 1, "row must be at least 128 bytes");

  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  // [stages][K | V] then merge buffer, then barriers / scalars
  uint8_t* stage_base = smem;
  float* mbuf = reinterpret_cast<float*>(smem + (size_t)P.stages * 2 * KV_STAGE_BYTES);  // [NW][8][D]
  float* ml_buf = mbuf + NW * 8 * D;                                                      // [NW][8][2]
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(ml_buf + NW * 8 * 2);
  uint64_t* empty_bar = full_bar + kMaxStages;
  int* s_part = reinterpret_cast<int*>(empty_bar + kMaxStages);  // [0]=begin [1]=end [2]=b [3]=hg [4]=step [5]=per [6]=ticket flag

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int gid = lane >> 2, t4 = lane & 3;

  unsigned long long* trace = P.trace ? P.trace + (size_t)blockIdx.x * 8 : nullptr;
  if (trace && threadIdx.x == 0) trace[0] = global_ns(), trace[1] = clock64(), trace[7] = 0;

  // ---- work partition and barrier init ----------------------------------------
  if (warp == NW && lane == 0) {  // descriptor fetch overlaps the partition
    tma_prefetch_desc(&k_map);
    tma_prefetch_desc(&v_map);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < P.stages; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], NW);
    }
    fence_barrier_init();
  }
  // Everything above touched only shared memory and kernel parameters.  From here on the kernel reads what
  // earlier kernels on the stream wrote (metadata, q, the cache): wait for them (no-op without PDL).
  pdl_wait();
  int my_begin, my_end, per_cta;
  ItemCursor cur;
  int tq_first = 0;  // first query-token index of cur.b
  constexpr int kFastSeqs = 256;
  if (P.a.num_seqs <= kFastSeqs) {
    // Every warp derives the partition on its own from one round of independent
    // loads (lane l holds sequences l, l+32, ...): no shared-memory hand-off and a
    // single global-memory latency before the first page can be requested.
    constexpr int R = kFastSeqs / 32;
    int n_r[R], first_r[R], sl_r[R], tq_r[R];
    int total = 0;
    unsigned zero_rows = 0;  // bit r: sequence lane+32r has no visible key and one query token
#pragma unroll
    for (int r = 0; r < R; ++r) {
      const int b = r * 32 + lane;
      n_r[r] = 0, first_r[r] = 0, sl_r[r] = 0, tq_r[r] = 0;
      if (b < P.a.num_seqs) {
        const int sl = P.a.seq_lens[b];
        int c0 = b, c1 = b + 1;
        if (P.a.cu_seqlens_q) c0 = P.a.cu_seqlens_q[b], c1 = P.a.cu_seqlens_q[b + 1];
        const int ql = c1 - c0;
        sl_r[r] = sl, tq_r[r] = c0;
        if (sl > 0 && ql > 0 && !(P.only_q1 && ql != 1)) {
          int lo_tok = 0;
          if (P.a.sliding_window > 0) lo_tok = max(0, sl - P.a.sliding_window);
          first_r[r] = lo_tok >> P.tps_shift;
          n_r[r] = ((sl + P.tps - 1) >> P.tps_shift) - first_r[r];
        }
        if (sl <= 0 && ql == 1 && b % gridDim.x == blockIdx.x) zero_rows |= 1u << r;
        total += n_r[r];
      }
    }
    total = __reduce_add_sync(0xffffffffu, total);
    const long long total_flat = (long long)total * P.head_groups;
    // 64-bit division is a long dependent sequence; the totals fit 32 bits unless the batch holds > 2^31 steps
    const bool small = total_flat < (1ll << 31);
    const long long per = small ? (long long)(((unsigned)total_flat + gridDim.x - 1) / gridDim.x)
                                : (total_flat + gridDim.x - 1) / gridDim.x;
    long long begin = per * blockIdx.x;
    long long end = begin + per;
    if (begin > total_flat) begin = total_flat;
    if (end > total_flat) end = total_flat;
    int fb = -1, fhg = 0, fstep = 0, fsl = 0, ffirst = 0, fn = 0;
    if (begin < end) {
      long long run = 0;
#pragma unroll
      for (int r = 0; r < R; ++r) {
        if (r * 32 < P.a.num_seqs && fb < 0) {  // warp-uniform
          int incl = n_r[r];
#pragma unroll
          for (int o = 1; o < 32; o <<= 1) {
            const int v = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += v;
          }
          const long long lo = run + (long long)(incl - n_r[r]) * P.head_groups;
          const long long hi = run + (long long)incl * P.head_groups;
          const unsigned hit = __ballot_sync(0xffffffffu, n_r[r] > 0 && begin >= lo && begin < hi);
          if (hit) {
            const int src = __ffs(hit) - 1;
            const long long lo_s = __shfl_sync(0xffffffffu, lo, src);
            fn = __shfl_sync(0xffffffffu, n_r[r], src);
            ffirst = __shfl_sync(0xffffffffu, first_r[r], src);
            fsl = __shfl_sync(0xffffffffu, sl_r[r], src);
            tq_first = __shfl_sync(0xffffffffu, tq_r[r], src);
            fb = r * 32 + src;
            const long long rem = begin - lo_s;  // < fn * head_groups
            if (small) {
              const unsigned qd = (unsigned)rem / (unsigned)fn;
              fhg = (int)qd, fstep = ffirst + (int)((unsigned)rem - qd * (unsigned)fn);
            } else {
              fhg = (int)(rem / fn), fstep = ffirst + (int)(rem % fn);
            }
          }
          run += (long long)__shfl_sync(0xffffffffu, incl, 31) * P.head_groups;
        }
      }
    }
    my_begin = (int)begin, my_end = (fb < 0) ? (int)begin : (int)end;
    per_cta = (int)(per > 0 ? per : 1);
    cur.b = fb, cur.hg = fhg, cur.step = fstep, cur.flat = my_begin;
    cur.seq_len = fsl, cur.first = ffirst, cur.n = fn;
    // sequences with no visible key get zeros (reference reduce_segments :828); nobody visits them below
    if (warp == 0 && __any_sync(0xffffffffu, zero_rows != 0)) {
#pragma unroll
      for (int r = 0; r < R; ++r) {
        unsigned m = __ballot_sync(0xffffffffu, (zero_rows >> r) & 1u);
        while (m) {
          const int src = __ffs(m) - 1;
          m &= m - 1;
          const int tq = __shfl_sync(0xffffffffu, tq_r[r], src);
          for (int e = lane; e < P.a.num_q_heads * D; e += 32) {
            const int osz = P.a.out_scale ? 1 : 2;  // fp8 output: one byte per element, 0x00 == +0
            char* orow = static_cast<char*>(P.a.out) + ((int64_t)tq * P.a.out_stride0 + (int64_t)(e / D) * P.a.out_stride1) * osz;
            if (osz == 1) orow[e % D] = 0;
            else reinterpret_cast<uint16_t*>(orow)[e % D] = 0;
          }
        }
      }
    }
    __syncthreads();  // mbarrier init visible to every warp
  } else {
    if (warp == 0) {
      // total steps over all sequences
      int total = 0;
      for (int b0 = 0; b0 < P.a.num_seqs; b0 += 32) {
        int b = b0 + lane, n = 0;
        if (b < P.a.num_seqs) {
          int sl, first;
          seq_steps(P, b, sl, first, n);
        }
        total += n;
      }
      total = __reduce_add_sync(0xffffffffu, total);
      const long long total_flat = (long long)total * P.head_groups;
      const long long per = (total_flat + gridDim.x - 1) / gridDim.x;
      long long begin = per * blockIdx.x;
      long long end = begin + per;
      if (begin > total_flat) begin = total_flat;
      if (end > total_flat) end = total_flat;
      // locate `begin`: largest b with HG * prefix(b) <= begin and n_b > 0 covering it
      int fb = -1, fhg = 0, fstep = 0;
      if (begin < end) {
        long long run = 0;  // HG * sum of n over sequences before the chunk
        for (int b0 = 0; b0 < P.a.num_seqs && fb < 0; b0 += 32) {
          int b = b0 + lane, n = 0, first = 0, sl;
          if (b < P.a.num_seqs) seq_steps(P, b, sl, first, n);
          int incl = n;
#pragma unroll
          for (int o = 1; o < 32; o <<= 1) {
            int v = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += v;
          }
          const long long lo = run + (long long)(incl - n) * P.head_groups;
          const long long hi = run + (long long)incl * P.head_groups;
          const bool mine = n > 0 && begin >= lo && begin < hi;
          const unsigned hit = __ballot_sync(0xffffffffu, mine);
          if (hit) {
            const int src = __ffs(hit) - 1;
            const long long lo_s = __shfl_sync(0xffffffffu, lo, src);
            const int n_s = __shfl_sync(0xffffffffu, n, src);
            const int first_s = __shfl_sync(0xffffffffu, first, src);
            fb = b0 + src;
            const long long rem = begin - lo_s;
            fhg = (int)(rem / n_s);
            fstep = first_s + (int)(rem % n_s);
          }
          run += (long long)__shfl_sync(0xffffffffu, incl, 31) * P.head_groups;
        }
      }
      if (lane == 0) {
        s_part[0] = (int)begin;
        s_part[1] = (fb < 0) ? (int)begin : (int)end;
        s_part[2] = fb, s_part[3] = fhg, s_part[4] = fstep;
        s_part[5] = (int)(per > 0 ? per : 1);
      }
    }
    __syncthreads();
    for (int b = blockIdx.x; b < P.a.num_seqs; b += gridDim.x) {
      const int sl = P.a.seq_lens[b];
      const int ql = P.a.cu_seqlens_q ? (P.a.cu_seqlens_q[b + 1] - P.a.cu_seqlens_q[b]) : 1;
      if (sl <= 0 && ql == 1) {
        const int tq = P.a.cu_seqlens_q ? P.a.cu_seqlens_q[b] : b;
        for (int e = threadIdx.x; e < P.a.num_q_heads * D; e += blockDim.x) {
          const int osz = P.a.out_scale ? 1 : 2;
          char* orow = static_cast<char*>(P.a.out) + ((int64_t)tq * P.a.out_stride0 + (int64_t)(e / D) * P.a.out_stride1) * osz;
          if (osz == 1) orow[e % D] = 0;
          else reinterpret_cast<uint16_t*>(orow)[e % D] = 0;
        }
      }
    }
    my_begin = s_part[0], my_end = s_part[1];
    per_cta = s_part[5];
    cur.b = s_part[2], cur.hg = s_part[3], cur.step = s_part[4], cur.flat = my_begin;
    cur.seq_len = 0, cur.first = 0, cur.n = 0;
    if (my_begin < my_end) {
      seq_steps(P, cur.b, cur.seq_len, cur.first, cur.n);
      tq_first = P.a.cu_seqlens_q ? P.a.cu_seqlens_q[cur.b] : cur.b;
    }
  }
  if (trace && threadIdx.x == 0) trace[2] = clock64();
  if (my_begin >= my_end) {
    if (trace && threadIdx.x == 0) trace[3] = trace[4] = trace[5] = clock64(), trace[6] = global_ns();
    return;
  }

  const int boxb = P.hpc * NDC * P.box_slots * 128;  // bytes of one TMA box (all heads, all chunks)

  if (warp == NW) {
    // ======================= producer warp ====================================
    // The whole warp issues: lane i looks up the page of box i and launches its
    // K and V copies, so page-table latency and TMA issue overlap across boxes.
    const uint64_t pol = policy_evict_first();
    int stage = 0;
    uint32_t phase = 0;
    // Page ids are read through a 128-page window held in registers (lane l keeps
    // pages base + 32 r + l) with the next window already in flight, so the copy
    // loop never waits on a block-table load (one DRAM round trip per item, not per step).
    int win_b = -1, win_base = -1;
    int ids[4] = {0, 0, 0, 0}, nxt[4] = {0, 0, 0, 0};
    auto load_ids = [&](int (&dst)[4], const int32_t* bt, int base, int npages) {
#pragma unroll
      for (int r = 0; r < 4; ++r) {
        const int pg = base + r * 32 + lane;
        dst[r] = pg < npages ? bt[pg] : 0;
      }
    };
    while (true) {
      const int tok0 = cur.step * P.tps;
      const int ntok = min(P.tps, cur.seq_len - tok0);
      const int nbox = (ntok + P.box_slots - 1) / P.box_slots;
      const int head0 = cur.hg * P.hpc;
      const int32_t* bt = P.a.block_tables + (int64_t)cur.b * P.a.bt_stride0;
      const int npages = (cur.seq_len + P.a.block_size - 1) >> P.bs_shift;
      const int p_first = tok0 >> P.bs_shift;
      const int base = p_first & ~127;
      if (cur.b != win_b || base != win_base) {
        if (cur.b == win_b && base == win_base + 128) {
#pragma unroll
          for (int r = 0; r < 4; ++r) ids[r] = nxt[r];
        } else {
          load_ids(ids, bt, base, npages);
        }
        load_ids(nxt, bt, base + 128, npages);
        win_b = cur.b, win_base = base;
      }
      auto page_of = [&](int pg) {  // warp-collective: every lane asks for one page id
        const int src = pg & 31, r = (pg >> 5) & 3;
        const int v0 = __shfl_sync(0xffffffffu, ids[0], src), v1 = __shfl_sync(0xffffffffu, ids[1], src);
        const int v2 = __shfl_sync(0xffffffffu, ids[2], src), v3 = __shfl_sync(0xffffffffu, ids[3], src);
        return r == 0 ? v0 : (r == 1 ? v1 : (r == 2 ? v2 : v3));
      };
      mbar_wait(&empty_bar[stage], phase ^ 1);
      uint8_t* kdst = stage_base + (size_t)stage * 2 * KV_STAGE_BYTES;
      uint8_t* vdst = kdst + KV_STAGE_BYTES;
      if (lane == 0) mbar_arrive_expect_tx(&full_bar[stage], (uint32_t)(2 * nbox * boxb));
      __syncwarp();
      if (P.paired) {
        // k_map's third coordinate selects K (0) or V (1); the box takes both
        const int i = lane;
        const int tok = tok0 + i * P.box_slots;
        const int page = page_of(tok >> P.bs_shift);
        if (i < nbox)
          tma_load_5d(kdst + (size_t)i * 2 * boxb, &k_map, &full_bar[stage], 0, tok & (P.a.block_size - 1), 0, head0, page, pol);
      } else if (P.tma_split == 0) {
        // nbox <= 8: one lane per box
        const int i = lane;
        const int tok = tok0 + i * P.box_slots;
        const int page = page_of(tok >> P.bs_shift);  // lanes past nbox read a don't-care id
        if (i < nbox) {
          const int slot0 = tok & (P.a.block_size - 1);
          tma_load_5d(kdst + (size_t)i * boxb, &k_map, &full_bar[stage], 0, slot0, 0, head0, page, pol);
          tma_load_5d(vdst + (size_t)i * boxb, &v_map, &full_bar[stage], 0, slot0, 0, head0, page, pol);
        }
      } else {
        const int per_box = P.hpc * NDC;
        const int units = nbox * per_box;
        for (int u0 = 0; u0 < units; u0 += 32) {
          const int u = u0 + lane;
          const int i = min(u / per_box, nbox - 1), r = u % per_box, h = r / NDC, dc = r % NDC;
          const int tok = tok0 + i * P.box_slots;
          const int page = page_of(tok >> P.bs_shift);
          if (u < units) {
            const int slot0 = tok & (P.a.block_size - 1);
            const size_t off = (size_t)i * boxb + (size_t)(h * NDC + dc) * P.box_slots * 128;
            tma_load_5d(kdst + off, &k_map, &full_bar[stage], 0, slot0, dc, head0 + h, page, pol);
            tma_load_5d(vdst + off, &v_map, &full_bar[stage], 0, slot0, dc, head0 + h, page, pol);
          }
        }
      }
      if (++stage == P.stages) stage = 0, phase ^= 1;
      if (cur.flat + 1 >= my_end) break;
      cursor_next(P, cur);
    }
    return;
  }

  // ========================= consumer warps ====================================
  const int h_local = warp % P.hpc;           // kv head within the CTA's head group
  const int jtile = warp / P.hpc;             // TT-token group within the stage-step
  const int tok_off = jtile * TT;
  uint32_t tile_off[SUB];                     // byte offset of each 16-token sub-tile inside a stage (+ dc*dc_stride)
#pragma unroll
  for (int sb = 0; sb < SUB; ++sb) {
    const int to = tok_off + 16 * sb;
    tile_off[sb] = (uint32_t)(((to / P.box_slots) * boxb + h_local * NDC * P.box_slots * 128) * (P.paired ? 2 : 1) +
                              (to % P.box_slots) * 128);
  }
  const uint32_t dc_stride = (uint32_t)P.box_slots * 128;
  // V tile relative to the K tile of the same tokens: next plane of the paired box, or the stage's V half
  const uint32_t v_from_k = P.paired ? dc_stride : (uint32_t)KV_STAGE_BYTES;

  const float k_scale = (KV_FP8 && P.a.k_scale) ? *P.a.k_scale : 1.0f;
  const float v_scale = (KV_FP8 && P.a.v_scale) ? *P.a.v_scale : 1.0f;
  const bool has_cap = P.a.softcap > 0.f;
  const float inv_cap = has_cap ? 1.0f / P.a.softcap : 0.f;
  const float qk_scale = P.a.softmax_scale * k_scale;  // natural units
  const bool has_alibi = P.a.alibi_slopes != nullptr;

  // ldmatrix lane geometry
  const int lrow = (lane & 7);
  uint32_t k_lane_off, v_lane_off;  // byte offsets inside a [rows][128B] sub-tile, before the chunk XOR
  int k_chunk_sel, v_chunk_sel;     // which of the pair of 16-byte chunks this lane addresses
  if constexpr (!KV_FP8) {
    // K (non-trans): matrices {rows0-7,c},{rows8-15,c},{rows0-7,c+1},{rows8-15,c+1}
    k_lane_off = (uint32_t)((lrow + 8 * ((lane >> 3) & 1)) * 128);
    k_chunk_sel = lane >> 4;
    // V (trans): matrices {tok0-7,c},{tok0-7,c+1},{tok8-15,c},{tok8-15,c+1}
    v_lane_off = (uint32_t)((lrow + 8 * (lane >> 4)) * 128);
    v_chunk_sel = (lane >> 3) & 1;
  } else {
    // fp8: both use {rows0-7,c},{rows8-15,c},{rows0-7,c+1},{rows8-15,c+1}
    k_lane_off = (uint32_t)((lrow + 8 * ((lane >> 3) & 1)) * 128);
    k_chunk_sel = lane >> 4;
    v_lane_off = k_lane_off;
    v_chunk_sel = k_chunk_sel;
  }
  const uint32_t swz = (uint32_t)lrow;  // row & 7 (sub-tiles start on 1024-byte boundaries)

  uint32_t qf[KSTEPS][NT][2];
  float oacc[MTILES][NT][4];
  float m_run[NT][2], l_run[NT][2];
  float slope[NT][2];

  auto reset_state = [&]() {
#pragma unroll
    for (int mt = 0; mt < MTILES; ++mt)
#pragma unroll
      for (int nt = 0; nt < NT; ++nt)
#pragma unroll
        for (int i = 0; i < 4; ++i) oacc[mt][nt][i] = 0.f;
#pragma unroll
    for (int nt = 0; nt < NT; ++nt) m_run[nt][0] = m_run[nt][1] = -INFINITY, l_run[nt][0] = l_run[nt][1] = 0.f;
  };

  auto load_q = [&](int tq) {
    const int kvh = cur.hg * P.hpc + h_local;
#pragma unroll
    for (int nt = 0; nt < NT; ++nt) {
      const int hq_local = nt * 8 + gid;
      const bool ok = hq_local < P.group;
      const int hq = kvh * P.group + (ok ? hq_local : 0);
      const char* qrow = static_cast<const char*>(P.a.q) + ((int64_t)tq * P.a.q_stride0 + (int64_t)hq * P.a.q_stride1) * 2;
#pragma unroll
      for (int ks = 0; ks < KSTEPS; ++ks) {
        uint32_t w0 = 0, w1 = 0;
        if (ok) {
          if constexpr (!KV_FP8) {
            w0 = *reinterpret_cast<const uint32_t*>(qrow + (16 * ks + 2 * t4) * 2);
            w1 = *reinterpret_cast<const uint32_t*>(qrow + (16 * ks + 2 * t4 + 8) * 2);
          } else {
            // fp8 K: k16 step `ks` consumes dims 16ks+4t..+3 in this lane's slots
            const uint2 w = *reinterpret_cast<const uint2*>(qrow + (16 * ks + 4 * t4) * 2);
            w0 = w.x, w1 = w.y;
          }
          if constexpr (Q_BF16 && !MMA_BF16) {
            // bf16 query, f16 tensor-core operands (fp8 KV path)
            const __nv_bfloat162 a = *reinterpret_cast<const __nv_bfloat162*>(&w0);
            const __nv_bfloat162 b = *reinterpret_cast<const __nv_bfloat162*>(&w1);
            const float2 fa = __bfloat1622float2(a), fb = __bfloat1622float2(b);
            const float lim = 65504.f;
            w0 = pack2<false>(fminf(fmaxf(fa.x, -lim), lim), fminf(fmaxf(fa.y, -lim), lim));
            w1 = pack2<false>(fminf(fmaxf(fb.x, -lim), lim), fminf(fmaxf(fb.y, -lim), lim));
          }
        }
        qf[ks][nt][0] = w0, qf[ks][nt][1] = w1;
      }
      if (has_alibi) {
        const int h0 = nt * 8 + 2 * t4;
        slope[nt][0] = (h0 < P.group) ? P.a.alibi_slopes[kvh * P.group + h0] : 0.f;
        slope[nt][1] = (h0 + 1 < P.group) ? P.a.alibi_slopes[kvh * P.group + h0 + 1] : 0.f;
      } else {
        slope[nt][0] = slope[nt][1] = 0.f;
      }
    }
  };

  reset_state();
  load_q(tq_first);
  int tq_cur = tq_first;
  bool first_stage = true;

  int stage = 0;
  uint32_t phase = 0;
  while (true) {
    const int tok0 = cur.step * P.tps + tok_off;  // first token of this warp's tile
    mbar_wait(&full_bar[stage], phase);
    if (trace && first_stage && threadIdx.x == 0) trace[3] = clock64();
    first_stage = false;
    if (tok0 < cur.seq_len) {
      const uint32_t kbase_s = smem_u32(stage_base + (size_t)stage * 2 * KV_STAGE_BYTES);
      const uint32_t vbase_s = kbase_s + v_from_k;
      const int valid = cur.seq_len - tok0;  // > 0; tile tokens >= valid are out of range
      int win_lo = 0;                        // tokens < win_lo (tile-relative) are outside the sliding window
      if (P.a.sliding_window > 0) win_lo = (cur.seq_len - P.a.sliding_window) - tok0;
      const bool partial = valid < TT || win_lo > 0;

      // ---- S^T[16 tok][8 heads] = K Q^T per sub-tile; even/odd k-steps accumulate in
      // separate chains (a 16-deep dependent HMMA chain is what bounds a warp otherwise)
      float sacc[SUB][NT][4];
#pragma unroll
      for (int sb = 0; sb < SUB; ++sb) {
        float sodd[NT][4];
#pragma unroll
        for (int nt = 0; nt < NT; ++nt)
#pragma unroll
          for (int i = 0; i < 4; ++i) sacc[sb][nt][i] = 0.f, sodd[nt][i] = 0.f;
        if (sb == 0 || 16 * sb < valid) {  // warp-uniform; a sub-tile past the end is masked below
          const uint32_t ktile = kbase_s + tile_off[sb];
          if constexpr (!KV_FP8) {
#pragma unroll
            for (int dc = 0; dc < NDC; ++dc) {
#pragma unroll
              for (int k4 = 0; k4 < 4; ++k4) {
                uint32_t a[4];
                const uint32_t chunk = (uint32_t)(2 * k4 + k_chunk_sel);
                ldmatrix_x4(a, ktile + dc * dc_stride + k_lane_off + ((chunk ^ swz) << 4));
#pragma unroll
                for (int nt = 0; nt < NT; ++nt)
                  mma_16816<MMA_BF16>((k4 & 1) ? sodd[nt] : sacc[sb][nt], a, qf[dc * 4 + k4][nt][0], qf[dc * 4 + k4][nt][1]);
              }
            }
          } else {
#pragma unroll
            for (int dc = 0; dc < NDC; ++dc) {
#pragma unroll
              for (int c2 = 0; c2 < 4; ++c2) {  // pairs of 16-byte chunks = pairs of k16 steps
                uint32_t r[4];
                const uint32_t chunk = (uint32_t)(2 * c2 + k_chunk_sel);
                ldmatrix_x4(r, ktile + dc * dc_stride + k_lane_off + ((chunk ^ swz) << 4));
#pragma unroll
                for (int half = 0; half < 2; ++half) {
                  uint32_t a[4];
                  a[0] = e4m3x2_to_f16x2((uint16_t)(r[2 * half] & 0xffffu));
                  a[1] = e4m3x2_to_f16x2((uint16_t)(r[2 * half + 1] & 0xffffu));
                  a[2] = e4m3x2_to_f16x2((uint16_t)(r[2 * half] >> 16));
                  a[3] = e4m3x2_to_f16x2((uint16_t)(r[2 * half + 1] >> 16));
                  const int ks = dc * 8 + c2 * 2 + half;
#pragma unroll
                  for (int nt = 0; nt < NT; ++nt)
                    mma_16816<false>(half ? sodd[nt] : sacc[sb][nt], a, qf[ks][nt][0], qf[ks][nt][1]);
                }
              }
            }
          }
#pragma unroll
          for (int nt = 0; nt < NT; ++nt)
#pragma unroll
            for (int i = 0; i < 4; ++i) sacc[sb][nt][i] += sodd[nt][i];
        }
      }

      // ---- online softmax (log2 domain), one running-max update for the whole tile ----
      uint32_t pb[SUB][NT][2];
      float alpha[NT][2];
#pragma unroll
      for (int nt = 0; nt < NT; ++nt) {
        float sv[SUB][4];
        float mx0 = -INFINITY, mx1 = -INFINITY;
#pragma unroll
        for (int sb = 0; sb < SUB; ++sb) {
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            float x = sacc[sb][nt][i] * qk_scale;
            if (has_cap) x = softcap_fn(x, P.a.softcap, inv_cap);
            sv[sb][i] = x * kLog2e;
          }
          const int tr = 16 * sb + gid;  // tile-relative token of rows gid (and gid + 8)
          if (has_alibi) {
            // bias slope * (pos - ctx), ctx = seq_len - 1 for a decode token
            const float p0 = (float)(tok0 + tr - (cur.seq_len - 1)), p1 = p0 + 8.f;
            sv[sb][0] += slope[nt][0] * p0 * kLog2e, sv[sb][1] += slope[nt][1] * p0 * kLog2e;
            sv[sb][2] += slope[nt][0] * p1 * kLog2e, sv[sb][3] += slope[nt][1] * p1 * kLog2e;
          }
          if (partial) {
            const bool ok0 = tr < valid && tr >= win_lo, ok1 = tr + 8 < valid && tr + 8 >= win_lo;
            if (!ok0) sv[sb][0] = sv[sb][1] = -INFINITY;
            if (!ok1) sv[sb][2] = sv[sb][3] = -INFINITY;
          }
          mx0 = fmaxf(mx0, fmaxf(sv[sb][0], sv[sb][2])), mx1 = fmaxf(mx1, fmaxf(sv[sb][1], sv[sb][3]));
        }
#pragma unroll
        for (int o = 4; o < 32; o <<= 1) {
          mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, o));
          mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, o));
        }
        const float mn0 = fmaxf(m_run[nt][0], mx0), mn1 = fmaxf(m_run[nt][1], mx1);
        const float mu0 = (mn0 == -INFINITY) ? 0.f : mn0, mu1 = (mn1 == -INFINITY) ? 0.f : mn1;
        alpha[nt][0] = ex2(m_run[nt][0] - mu0), alpha[nt][1] = ex2(m_run[nt][1] - mu1);
        m_run[nt][0] = mn0, m_run[nt][1] = mn1;
        float ls0 = 0.f, ls1 = 0.f;
#pragma unroll
        for (int sb = 0; sb < SUB; ++sb) {
          const float p0 = ex2(sv[sb][0] - mu0), p1 = ex2(sv[sb][1] - mu1), p2 = ex2(sv[sb][2] - mu0), p3 = ex2(sv[sb][3] - mu1);
          ls0 += p0 + p2, ls1 += p1 + p3;
          // P^T fragments: transpose the two 8x8 (token x head) blocks
          pb[sb][nt][0] = movmatrix_trans(pack2<MMA_BF16>(p0, p1));
          pb[sb][nt][1] = movmatrix_trans(pack2<MMA_BF16>(p2, p3));
        }
        l_run[nt][0] = l_run[nt][0] * alpha[nt][0] + ls0;
        l_run[nt][1] = l_run[nt][1] * alpha[nt][1] + ls1;
      }
      // rescale the accumulators (columns = heads 2t, 2t+1)
#pragma unroll
      for (int nt = 0; nt < NT; ++nt) {
        if (alpha[nt][0] != 1.f || alpha[nt][1] != 1.f) {
#pragma unroll
          for (int mt = 0; mt < MTILES; ++mt) {
            oacc[mt][nt][0] *= alpha[nt][0], oacc[mt][nt][2] *= alpha[nt][0];
            oacc[mt][nt][1] *= alpha[nt][1], oacc[mt][nt][3] *= alpha[nt][1];
          }
        }
      }

      // ---- O^T[D][8 heads] += V^T P^T ------------------------------------------
#pragma unroll
      for (int sb = 0; sb < SUB; ++sb) {
        if (sb > 0 && 16 * sb >= valid) continue;  // warp-uniform: nothing of this sub-tile is visible
        const uint32_t vtile = vbase_s + tile_off[sb];
        // token masks for the V fragments of a partial tile (garbage x 0 must not make NaN)
        uint32_t vm_lo = 0xffffffffu, vm_hi = 0xffffffffu;
        if (partial) {
          const int ta = 16 * sb + 2 * t4, tb = ta + 8;
          auto okt = [&](int tk) { return tk < valid && tk >= win_lo; };
          vm_lo = (okt(ta) ? 0x0000ffffu : 0u) | (okt(ta + 1) ? 0xffff0000u : 0u);
          vm_hi = (okt(tb) ? 0x0000ffffu : 0u) | (okt(tb + 1) ? 0xffff0000u : 0u);
        }
        if constexpr (!KV_FP8) {
#pragma unroll
          for (int dc = 0; dc < NDC; ++dc) {
#pragma unroll
            for (int m4 = 0; m4 < 4; ++m4) {
              uint32_t a[4];
              const uint32_t chunk = (uint32_t)(2 * m4 + v_chunk_sel);
              ldmatrix_x4_trans(a, vtile + dc * dc_stride + v_lane_off + ((chunk ^ swz) << 4));
              // a0:(d gid; tok 2t,2t+1) a1:(d gid+8; same) a2:(d gid; tok 2t+8,+9) a3:(d gid+8; ...)
              if (partial) a[0] &= vm_lo, a[1] &= vm_lo, a[2] &= vm_hi, a[3] &= vm_hi;
#pragma unroll
              for (int nt = 0; nt < NT; ++nt) mma_16816<MMA_BF16>(oacc[dc * 4 + m4][nt], a, pb[sb][nt][0], pb[sb][nt][1]);
            }
          }
        } else {
#pragma unroll
          for (int dc = 0; dc < NDC; ++dc) {
#pragma unroll
            for (int c2 = 0; c2 < 4; ++c2) {
              uint32_t r[4];
              const uint32_t chunk = (uint32_t)(2 * c2 + v_chunk_sel);
              ldmatrix_x4_trans(r, vtile + dc * dc_stride + v_lane_off + ((chunk ^ swz) << 4));
              // r0: tokens 0-7 chunk c | r1: tokens 8-15 chunk c | r2, r3: chunk c+1
              // each: [tok 2t: d0 d1 | tok 2t+1: d0 d1] with d0 = 16c + 2gid
#pragma unroll
              for (int half = 0; half < 2; ++half) {
                const uint32_t x0 = __byte_perm(r[2 * half], 0, 0x3120);      // [t d0, t+1 d0 | t d1, t+1 d1]
                const uint32_t x1 = __byte_perm(r[2 * half + 1], 0, 0x3120);
                uint32_t a[4];
                a[0] = e4m3x2_to_f16x2((uint16_t)(x0 & 0xffffu));  // row gid   <-> d0, tokens 2t,2t+1
                a[1] = e4m3x2_to_f16x2((uint16_t)(x0 >> 16));      // row gid+8 <-> d1
                a[2] = e4m3x2_to_f16x2((uint16_t)(x1 & 0xffffu));  // tokens 2t+8, 2t+9
                a[3] = e4m3x2_to_f16x2((uint16_t)(x1 >> 16));
                if (partial) a[0] &= vm_lo, a[1] &= vm_lo, a[2] &= vm_hi, a[3] &= vm_hi;
                const int mt = dc * 8 + c2 * 2 + half;
#pragma unroll
                for (int nt = 0; nt < NT; ++nt) mma_16816<false>(oacc[mt][nt], a, pb[sb][nt][0], pb[sb][nt][1]);
              }
            }
          }
        }
      }
    }
    __syncwarp();
    if (lane == 0) mbar_arrive(&empty_bar[stage]);
    if (++stage == P.stages) stage = 0, phase ^= 1;

    const bool item_end = (cur.step + 1 == cur.first + cur.n);
    const bool last = (cur.flat + 1 >= my_end);
    if (item_end || last) {
      // ---------------- finalise this (sequence, head group) ------------------
      const int item = cur.b * P.head_groups + cur.hg;
      const int item_flat0 = cur.flat - (cur.step - cur.first);
      const int hg_done = cur.hg, n_done = cur.n;
      const bool direct = item_flat0 >= my_begin && item_flat0 + n_done <= my_end;
      const int tq = tq_cur;
      // Move on to the next item first: its dependent global loads (seq_lens / cu_seqlens_q, then the query
      // rows) are in flight while this item is merged and stored.  The query fragments are dead after the
      // item's last tile; the accumulators are reset only after the merge has read them.
      if (!last) {
        cursor_next(P, cur);
        tq_cur = P.a.cu_seqlens_q ? P.a.cu_seqlens_q[cur.b] : cur.b;
        load_q(tq_cur);
      }
      if (trace && threadIdx.x == 0) {
        if (last) trace[4] = clock64();
        trace[7] += 1;
      }
      const int wph = NW / P.hpc;  // warps per kv head
#pragma unroll
      for (int nt = 0; nt < NT; ++nt) {
        // exp-sums: add up the 8 row groups
        float l0 = l_run[nt][0], l1 = l_run[nt][1];
#pragma unroll
        for (int o = 4; o < 32; o <<= 1) {
          l0 += __shfl_xor_sync(0xffffffffu, l0, o);
          l1 += __shfl_xor_sync(0xffffffffu, l1, o);
        }
        if (nt > 0) asm volatile("bar.sync 1, %0;" ::"r"(NW * 32));
        float* mb = mbuf + warp * 8 * D;
#pragma unroll
        for (int mt = 0; mt < MTILES; ++mt) {
          int d0, d1;
          if constexpr (!KV_FP8) d0 = 16 * mt + gid, d1 = d0 + 8;
          else d0 = 16 * mt + 2 * gid, d1 = d0 + 1;
          mb[(2 * t4) * D + d0] = oacc[mt][nt][0];
          mb[(2 * t4 + 1) * D + d0] = oacc[mt][nt][1];
          mb[(2 * t4) * D + d1] = oacc[mt][nt][2];
          mb[(2 * t4 + 1) * D + d1] = oacc[mt][nt][3];
        }
        if (gid == 0) {
          ml_buf[(warp * 8 + 2 * t4) * 2] = m_run[nt][0], ml_buf[(warp * 8 + 2 * t4) * 2 + 1] = l0;
          ml_buf[(warp * 8 + 2 * t4 + 1) * 2] = m_run[nt][1], ml_buf[(warp * 8 + 2 * t4 + 1) * 2 + 1] = l1;
        }
        asm volatile("bar.sync 1, %0;" ::"r"(NW * 32));
        // combine the warps that share a kv head; NW*32 threads cover hpc*8*D values
        const int nheads = min(8, P.group - nt * 8);
        const int tid = threadIdx.x;
        for (int e = tid; e < P.hpc * 8 * D; e += NW * 32) {
          const int hl = e / (8 * D), col = (e / D) & 7, d = e % D;
          if (col >= nheads) continue;
          float M = -INFINITY;
          for (int j = 0; j < wph; ++j) M = fmaxf(M, ml_buf[((hl + P.hpc * j) * 8 + col) * 2]);
          float val = 0.f, L = 0.f;
          for (int j = 0; j < wph; ++j) {
            const int w = hl + P.hpc * j;
            const float mw = ml_buf[(w * 8 + col) * 2];
            const float sc = (mw == -INFINITY) ? 0.f : ex2(mw - M);
            val += mbuf[(w * 8 + col) * D + d] * sc;
            L += ml_buf[(w * 8 + col) * 2 + 1] * sc;
          }
          const int hq_in_cta = hl * P.group + nt * 8 + col;  // query head within the CTA's head group
          if (direct) {
            const int hq = hg_done * P.hpc * P.group + hq_in_cta;
            if (P.a.sinks) val *= sink_rescale<true>(P.a.sinks[hq], M, L);
            const float o = (L > 0.f) ? val * v_scale / L : 0.f;
            const int64_t oel = (int64_t)tq * P.a.out_stride0 + (int64_t)hq * P.a.out_stride1;
            if (P.a.out_scale) {  // fused e4m3 output quantisation (b200attn.h: out_scale)
              static_cast<uint8_t*>(P.a.out)[oel + d] = fp8_out(o * out_quant_factor(P.a.out_scale));
            } else {
              char* orow = static_cast<char*>(P.a.out) + oel * 2;
              if constexpr (Q_BF16) reinterpret_cast<__nv_bfloat16*>(orow)[d] = __float2bfloat16_rn(o);
              else reinterpret_cast<__half*>(orow)[d] = __float2half_rn(o);
            }
          } else {
            const int64_t slot = (int64_t)item + blockIdx.x;
            const int64_t rec = slot * (P.hpc * P.group) + hq_in_cta;
            P.ws_acc[rec * D + d] = val;
            if (d == 0) P.ws_ml[rec * 2] = M, P.ws_ml[rec * 2 + 1] = L;
          }
        }
      }
      if (!direct) {
        // ---- ticket: the CTA that completes the item merges its partials ------
        const int first_cta = item_flat0 / per_cta;
        const int last_cta = (item_flat0 + n_done - 1) / per_cta;
        __threadfence();  // partials visible device-wide before the ticket
        asm volatile("bar.sync 1, %0;" ::"r"(NW * 32));
        if (threadIdx.x == 0) {
          const int ticket = atomicAdd(&P.ws_cnt[item], 1);
          s_part[6] = (ticket == last_cta - first_cta) ? 1 : 0;
        }
        asm volatile("bar.sync 1, %0;" ::"r"(NW * 32));
        if (s_part[6]) {
          __threadfence();
          // Every thread owns (query head, 4 dims) chunks; the (max, exp-sum) pair and the
          // float4 of every part are requested before any of them is used, so a chunk costs
          // one L2 round trip (a loop that first reduces the maxima costs two per element,
          // which was 20 us of tail at 4 kv heads per CTA).
          const int rec_heads = P.hpc * P.group;
          const int nparts = last_cta - first_cta + 1;
          constexpr int D4 = D / 4;
          for (int e = threadIdx.x; e < rec_heads * D4; e += NW * 32) {
            const int hq_in = e / D4, d4 = (e % D4) * 4;
            const int64_t rec0 = ((int64_t)item + first_cta) * rec_heads + hq_in;
            float M = -INFINITY, L = 0.f;
            float4 val = make_float4(0.f, 0.f, 0.f, 0.f);
            if (nparts <= 4) {
              float2 ml[4];
              float4 av[4];
#pragma unroll
              for (int c = 0; c < 4; ++c) {
                if (c < nparts) {
                  const int64_t rec = rec0 + (int64_t)c * rec_heads;
                  ml[c] = __ldcg(reinterpret_cast<const float2*>(&P.ws_ml[rec * 2]));
                  av[c] = __ldcg(reinterpret_cast<const float4*>(&P.ws_acc[rec * D + d4]));
                } else {
                  ml[c] = make_float2(-INFINITY, 0.f), av[c] = make_float4(0.f, 0.f, 0.f, 0.f);
                }
              }
#pragma unroll
              for (int c = 0; c < 4; ++c) M = fmaxf(M, ml[c].x);
#pragma unroll
              for (int c = 0; c < 4; ++c) {
                const float sc = (ml[c].x == -INFINITY) ? 0.f : ex2(ml[c].x - M);
                val.x += av[c].x * sc, val.y += av[c].y * sc, val.z += av[c].z * sc, val.w += av[c].w * sc;
                L += ml[c].y * sc;
              }
            } else {
              for (int c = 0; c < nparts; ++c) M = fmaxf(M, __ldcg(&P.ws_ml[(rec0 + (int64_t)c * rec_heads) * 2]));
#pragma unroll 4
              for (int c = 0; c < nparts; ++c) {
                const int64_t rec = rec0 + (int64_t)c * rec_heads;
                const float2 ml = __ldcg(reinterpret_cast<const float2*>(&P.ws_ml[rec * 2]));
                const float4 av = __ldcg(reinterpret_cast<const float4*>(&P.ws_acc[rec * D + d4]));
                const float sc = (ml.x == -INFINITY) ? 0.f : ex2(ml.x - M);
                val.x += av.x * sc, val.y += av.y * sc, val.z += av.z * sc, val.w += av.w * sc;
                L += ml.y * sc;
              }
            }
            const int hq = hg_done * rec_heads + hq_in;
            float fs = 1.f;
            if (P.a.sinks) fs = sink_rescale<true>(P.a.sinks[hq], M, L);  // once per item, here (upstream: segment 0 only)
            const float inv = (L > 0.f) ? fs * v_scale / L : 0.f;  // 0 when nothing is visible (reference :828)
            const int64_t oel = (int64_t)tq * P.a.out_stride0 + (int64_t)hq * P.a.out_stride1;
            const float o4[4] = {val.x * inv, val.y * inv, val.z * inv, val.w * inv};
            if (P.a.out_scale) {
              const float oq = out_quant_factor(P.a.out_scale);
              *reinterpret_cast<uint32_t*>(static_cast<char*>(P.a.out) + oel + d4) = fp8x4_out(o4[0] * oq, o4[1] * oq, o4[2] * oq, o4[3] * oq);
            } else {
              char* orow = static_cast<char*>(P.a.out) + oel * 2;
#pragma unroll
              for (int j = 0; j < 4; ++j) {
                if constexpr (Q_BF16) reinterpret_cast<__nv_bfloat16*>(orow)[d4 + j] = __float2bfloat16_rn(o4[j]);
                else reinterpret_cast<__half*>(orow)[d4 + j] = __float2half_rn(o4[j]);
              }
            }
          }
          if (threadIdx.x == 0) P.ws_cnt[item] = 0;  // self-cleaning: ready for the next call
        }
      }
      if (last) break;
      asm volatile("bar.sync 1, %0;" ::"r"(NW * 32));  // merge buffers are reused by the next item
      reset_state();
    } else {
      cursor_next(P, cur);
    }
  }
  if (trace && threadIdx.x == 0) trace[5] = clock64(), trace[6] = global_ns();
}

// ---- host side ------------------------------------------------------------------
struct DecodeConfig {
  int nw, nt, hpc, stages, smem_bytes, grid;
};

inline int pick_hpc(int hkv, int nw, int want) {
  int hpc = 1;
  for (int c = nw; c >= 1; c >>= 1)
    if (hkv % c == 0 && (want <= 0 || c <= want)) {
      hpc = c;
      break;
    }
  return hpc;
}

template <int D, bool KV_FP8, bool MMA_BF16, bool Q_BF16, int NT, int NW, int TT>
int launch_variant(const b200attn_attention_params& a, DecodeParams& P, cudaStream_t stream) {
  constexpr int ESZ = KV_FP8 ? 1 : 2;
  constexpr int NDC = D * ESZ / 128;
  constexpr int KV_STAGE_BYTES = NW * TT * D * ESZ;
  const DeviceInfo& dev = device_info();
  const Tuning& tn = tuning();
  // K and V of a page in ONE box when they are two equally strided views of one allocation
  // (both vLLM layouts) and a row is a single 128-byte chunk (fp8, D=128): halves the box count,
  // which is what bounds small pages (1-2 kv heads per GPU).  tuning().reserved bit 3 forbids it.
  const int64_t kv_delta = (static_cast<const char*>(a.v_cache) - static_cast<const char*>(a.k_cache));
  const bool can_pair = KV_FP8 && NDC == 1 && !(tn.reserved & 8) && !(tn.reserved & 1) && kv_delta > 0 &&
                        kv_delta % 16 == 0 && kv_delta < (1ll << 40) && a.kc_block_stride == a.vc_block_stride &&
                        a.kc_slot_stride == a.vc_slot_stride && a.kc_head_stride == a.vc_head_stride;
  if (tn.decode_heads_per_cta > 0) {
    P.hpc = pick_hpc(a.num_kv_heads, NW, tn.decode_heads_per_cta);
  } else {
    // Fewest kv heads per CTA whose TMA box is still big enough: more
    // (sequence, head group) items means more outputs written directly and
    // parallel merges, but the copy engine pays per box (measured on B200,
    // profiles/r01_decode_sweep_v2.jsonl: bf16 KV is best at 4 KB boxes, fp8 at 8 KB).
    const int min_box = KV_FP8 ? 8192 : 4096;
    int hpc = 1;
    while (hpc < NW && a.num_kv_heads % (hpc * 2) == 0) {
      const int tps = NW * TT / hpc;
      const int slots = a.block_size < tps ? a.block_size : tps;
      if (hpc * NDC * slots * 128 >= min_box) break;
      hpc *= 2;
    }
    P.hpc = hpc;
  }
  P.head_groups = a.num_kv_heads / P.hpc;
  P.tps = NW * TT / P.hpc;
  for (P.tps_shift = 0; (1 << P.tps_shift) < P.tps; ++P.tps_shift) {}
  for (P.bs_shift = 0; (1 << P.bs_shift) < a.block_size; ++P.bs_shift) {}
  if ((1 << P.tps_shift) != P.tps || (1 << P.bs_shift) != a.block_size) {
    set_error("decode: tokens per step (%d) and block_size (%d) must be powers of two", P.tps, a.block_size);
    return B200ATTN_E_UNSUPPORTED;
  }
  P.box_slots = a.block_size < P.tps ? a.block_size : P.tps;
  P.boxes = P.tps / P.box_slots;
  P.tma_split = tn.reserved & 1;
  P.paired = 0;
  const int fixed = NW * 8 * D * 4 + NW * 8 * 2 * 4 + 2 * kMaxStages * 8 + 64 + 1024 /*alignment slack*/;
  int stages = (dev.max_smem_optin - fixed) / (2 * KV_STAGE_BYTES);
  if (stages > kMaxStages) stages = kMaxStages;
  if (tn.decode_stages > 0 && tn.decode_stages < stages) stages = tn.decode_stages;
  if (stages < 2) {
    set_error("decode: not enough shared memory for 2 stages (D=%d)", D);
    return B200ATTN_E_UNSUPPORTED;
  }
  P.stages = stages;
  const int smem = fixed + stages * 2 * KV_STAGE_BYTES;

  CUtensorMap kmap, vmap;
  const int box_heads = P.tma_split ? 1 : P.hpc, box_dc = P.tma_split ? 1 : NDC;
  P.paired = (can_pair && P.hpc * NDC * P.box_slots * 128 <= 16384) ? 1 : 0;  // measured: profiles/r01_kbench_paired.jsonl
  int rc = make_kv_map(&kmap, a.k_cache, ESZ, D, a.block_size, a.num_kv_heads, a.num_blocks, a.kc_block_stride,
                       a.kc_slot_stride, a.kc_head_stride, P.box_slots, box_heads, box_dc, P.paired ? kv_delta : 0);
  if (rc) return rc;
  rc = make_kv_map(&vmap, a.v_cache, ESZ, D, a.block_size, a.num_kv_heads, a.num_blocks, a.vc_block_stride,
                   a.vc_slot_stride, a.vc_head_stride, P.box_slots, box_heads, box_dc);
  if (rc) return rc;

  // grid: one CTA per SM, never more CTAs than the upper bound on steps
  const long long max_steps =
      (long long)a.num_seqs * P.head_groups * ((a.max_seqlen_k + P.tps - 1) / P.tps > 0 ? (a.max_seqlen_k + P.tps - 1) / P.tps : 1);
  long long grid = dev.num_sms < kMaxDecodeCtas ? dev.num_sms : kMaxDecodeCtas;
  if (grid > max_steps && !launch_is_captured()) grid = max_steps;
  if (grid < 1) grid = 1;

  auto kern = attn_decode_kernel<D, KV_FP8, MMA_BF16, Q_BF16, NT, NW, TT>;
  if (int rc = ensure_dynamic_smem(reinterpret_cast<const void*>(kern), dev.max_smem_optin, "decode")) return rc;
  // Programmatic dependent launch: lets this grid's launch and prologue overlap the cache write in front of it
  // (common.cuh).  tuning().reserved bit 9 turns it off.
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)grid), cfg.blockDim = dim3((NW + 1) * 32), cfg.dynamicSmemBytes = (size_t)smem, cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr, cfg.numAttrs = (tn.reserved & 512) ? 0 : 1;
  cudaLaunchKernelEx(&cfg, kern, P, kmap, vmap);
  note_launch("attn_decode_split_kv");
  return check_launch("attn_decode");
}

template <int D, int NT, int NW>
int dispatch_types(const b200attn_attention_params& a, DecodeParams& P, cudaStream_t stream) {
  const bool qbf = a.q_dtype == B200ATTN_BF16;
  if (a.kv_dtype == B200ATTN_FP8_E4M3) {
    if constexpr (D >= 128) {
      // fp8: two 16-token sub-tiles per warp and step (a stage of 8 x 32 tokens is the size a bf16
      // stage of 8 x 16 has); the tile's latency chain, not HBM, bounds this variant otherwise
      constexpr int TT = (D == 128 && NT == 1) ? 32 : 16;  // (NT = 2 would spill)
      return qbf ? launch_variant<D, true, false, true, NT, NW, TT>(a, P, stream)
                 : launch_variant<D, true, false, false, NT, NW, TT>(a, P, stream);
    } else {
      set_error("decode: fp8 KV needs head_size >= 128");
      return B200ATTN_E_UNSUPPORTED;
    }
  }
  return qbf ? launch_variant<D, false, true, true, NT, NW, 16>(a, P, stream)
             : launch_variant<D, false, false, false, NT, NW, 16>(a, P, stream);
}

inline bool pow2(int x) { return x > 0 && (x & (x - 1)) == 0; }

}  // namespace

// The ticket counters live in a region of FIXED size at the start of the workspace: they are the only bytes
// that must be zero between calls, and the same buffer serves calls of every shape (a batch grows and shrinks
// from step to step), so their region may never be another shape's partial-result region.
constexpr uint64_t kCounterBytes = 1ull << 20;  // one int per (sequence, head group): 262144 items

uint64_t decode_workspace_bytes(int num_seqs, int num_q_heads, int num_kv_heads, int head_size) {
  // slots <= items + ctas; records of (hpc*group) heads: bounded by
  // (B*Hq + ctas*8*group) head-rows of (D + 2) floats, after the counter region.
  const uint64_t group = (uint64_t)(num_q_heads / (num_kv_heads > 0 ? num_kv_heads : 1));
  const uint64_t ctas = kMaxDecodeCtas;  // the grid is min(#SMs, kMaxDecodeCtas)
  const uint64_t rows = (uint64_t)num_seqs * num_q_heads + ctas * 8 * (group > 0 ? group : 1);
  return kCounterBytes + rows * (uint64_t)(head_size + 2) * 4 + 1024;
}

bool decode_supported(const b200attn_attention_params& p, const char** why, bool mixed_batch) {
  auto no = [&](const char* w) { if (why) *why = w; return false; };
  if (p.max_seqlen_q != 1 && !mixed_batch) return no("max_seqlen_q != 1");
  if (p.q_dtype != B200ATTN_BF16 && p.q_dtype != B200ATTN_F16) return no("query dtype must be bf16/f16");
  if (p.kv_dtype != p.q_dtype && p.kv_dtype != B200ATTN_FP8_E4M3) return no("kv dtype");
  if (p.head_size != 64 && p.head_size != 128 && p.head_size != 256) return no("head_size not in {64,128,256}");
  if (p.kv_dtype == B200ATTN_FP8_E4M3 && p.head_size < 128) return no("fp8 KV needs head_size >= 128");
  if (p.num_kv_heads <= 0 || p.num_q_heads % p.num_kv_heads) return no("heads");
  const int g = p.num_q_heads / p.num_kv_heads;
  if (g > 16 || (g > 8 && p.head_size == 256)) return no("GQA group too large for the decode kernel");
  if (!pow2(p.block_size) || p.block_size < 16 || p.block_size > 256) return no("block_size not a power of two in [16,256]");
  const int esz = p.kv_dtype == B200ATTN_FP8_E4M3 ? 1 : 2;
  auto m16 = [&](int64_t s) { return (s * esz) % 16 == 0; };
  if (!m16(p.kc_block_stride) || !m16(p.kc_slot_stride) || !m16(p.kc_head_stride) || !m16(p.vc_block_stride) ||
      !m16(p.vc_slot_stride) || !m16(p.vc_head_stride))
    return no("cache strides not 16-byte multiples");
  if ((reinterpret_cast<uintptr_t>(p.k_cache) & 15) || (reinterpret_cast<uintptr_t>(p.v_cache) & 15)) return no("cache base alignment");
  if ((reinterpret_cast<uintptr_t>(p.q) & 7) || (p.q_stride0 % 4) || (p.q_stride1 % 4)) return no("query alignment");
  return true;
}

int launch_attn_decode(const b200attn_attention_params& a, cudaStream_t stream, bool only_single_token_seqs) {
  const char* why = nullptr;
  if (!decode_supported(a, &why, only_single_token_seqs)) {
    set_error("decode: unsupported configuration (%s)", why);
    return B200ATTN_E_UNSUPPORTED;
  }
  const uint64_t need = decode_workspace_bytes(a.num_seqs, a.num_q_heads, a.num_kv_heads, a.head_size);
  if (!a.workspace || a.workspace_bytes < need) {
    set_error("decode: workspace too small (%llu < %llu bytes)", (unsigned long long)a.workspace_bytes,
              (unsigned long long)need);
    return B200ATTN_E_WORKSPACE;
  }
  DecodeParams P;
  P.a = a;
  P.group = a.num_q_heads / a.num_kv_heads;
  P.only_q1 = only_single_token_seqs ? 1 : 0;
  P.trace = decode_trace_buffer();
  // carve the workspace: ticket counters first (fixed region, zero between calls), then ml, then acc
  const uint64_t items_max = (uint64_t)a.num_seqs * a.num_kv_heads;
  if (items_max * sizeof(int) > kCounterBytes) {
    set_error("decode: %llu (sequence, kv head) items exceed the %llu ticket counters of the workspace",
              (unsigned long long)items_max, (unsigned long long)(kCounterBytes / sizeof(int)));
    return B200ATTN_E_WORKSPACE;
  }
  char* w = static_cast<char*>(a.workspace);
  P.ws_cnt = reinterpret_cast<int*>(w);
  uint64_t off = kCounterBytes;
  const uint64_t group = (uint64_t)P.group;
  const uint64_t rows = (uint64_t)a.num_seqs * a.num_q_heads + (uint64_t)kMaxDecodeCtas * 8 * group;
  P.ws_ml = reinterpret_cast<float*>(w + off);
  off += ((rows * 8 + 255) / 256) * 256;
  P.ws_acc = reinterpret_cast<float*>(w + off);
  if (off + rows * a.head_size * 4 > a.workspace_bytes) {
    set_error("decode: workspace carve-out exceeds the buffer");
    return B200ATTN_E_WORKSPACE;
  }
  const int nt = P.group > 8 ? 2 : 1;
  switch (a.head_size) {
    case 64: return nt == 1 ? dispatch_types<64, 1, 8>(a, P, stream) : dispatch_types<64, 2, 8>(a, P, stream);
    case 128: return nt == 1 ? dispatch_types<128, 1, 8>(a, P, stream) : dispatch_types<128, 2, 8>(a, P, stream);
    case 256: return dispatch_types<256, 1, 4>(a, P, stream);
  }
  return B200ATTN_E_UNSUPPORTED;
}

}  // namespace b200attn
