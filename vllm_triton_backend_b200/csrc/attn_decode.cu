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

constexpr int kMaxStages = 6;                                 // (a 32 KB stage, D = 64, gives 5)
constexpr int kDescSlots = 16;                                // run descriptors in flight (> kMaxStages + 1)
constexpr int kPoolRounds = 3;                                // pool chunks per CTA range, at most
constexpr int kMaxPoolChunks = kPoolRounds * kMaxDecodeCtas;  // pool chunks of one launch
constexpr int kMaxParts = (1 + kPoolRounds) * kMaxDecodeCtas; // parts one item can be cut into
constexpr int kMaxSlots = 2 * (kMaxDecodeCtas + kMaxPoolChunks);
constexpr int kDequeWord0 = 16;                                // ws_pool[kDequeWord0 + c]: range c's claim word

// ---- the plan: where every CTA (and every pool chunk) starts -----------------------------------
// Built on the device from seq_lens by decode_plan_kernel, once per step (all layers of a model share
// seq_lens; the metadata builder calls b200attn_decode_plan) or by the launcher itself when the caller
// brought none.  The flattened (sequence, head group, KV step) space [0, total) is cut into one range of
// `per` steps per CTA.  The first per_static steps of a range are that CTA's own (static) part; the rest -
// the range's tail - is cut into n_rounds chunks of halving size that go into a pool: chunk k = r * grid + c
// is round r of range c, and CTAs take chunks in that order with an atomic counter as they finish.
// SMs do not stream at the same rate when all of them pull from HBM (the slowest took 13 % longer than the
// mean on the 1-kv-head C4 shard: profiles/r02_kbench_baseline.jsonl, end_skew 11 us of 97), so an equal
// static split finishes when the slowest SM does; the pool lets the fast ones take the difference.  The
// pool is the tails of ALL ranges rather than the end of the step space, so every item is cut into a few
// more parts instead of a few items into fifty (profiles/r02_kbench_pool_v2_epilogue_warp.jsonl).
struct PlanHeader {  // 64 bytes
  int total;          // steps in the flat space
  int per;            // steps of a CTA's range: range c = [c * per, min((c + 1) * per, total))
  int per_static;     // its first per_static steps are the static part
  int n_rounds;       // pool chunks per range
  int n_chunks;       // n_rounds * grid
  int grid;
  int has_zero_rows;  // some sequence has one query token and no visible key (its output rows are zeros)
  int snapped;        // 1: the ranges' boundaries are Plan::bound[] (moved onto item boundaries they were one step from)
  short size[8];      // chunk of round r: [c * per + per_static + off[r], + size[r]) within range c
  short off[8];
};
struct PlanCursor {  // 32 bytes: a flat position resolved to (sequence, head group, step)
  int b;             // -1: nothing there
  int hg, step;      // step is absolute (in units of tps tokens) within the sequence
  int seq_len, first, n;  // of sequence b: visible steps are [first, first + n)
  int tq;            // first query token of b
  int flat;          // the flat position itself
};
struct Plan {
  PlanHeader hdr;
  PlanCursor cta[kMaxDecodeCtas];
  PlanCursor chunk[kMaxPoolChunks];
  int bound[kMaxDecodeCtas + 8];  // hdr.snapped: range c is [bound[c], bound[c + 1])
  // int prefix[num_seqs + 1] follows (scratch of the plan kernel)
};

// Part j of range c: j = 0 the static part, j >= 1 the pool chunk of round j - 1.  [b, e) may be empty.
// A boundary that falls one step inside an item would cost a run of one step - an epilogue, a partial record, a ticket
// and a share of a merge for 1.5 us of streaming - so the plan moves it onto the item's edge (hdr.snapped; static split
// only).  Ranges then differ by a step or two, and what they are is data: Plan::bound[].
__device__ __forceinline__ void part_range(const PlanHeader& h, const int* bound, int c, int j, int& b, int& e) {
  if (h.snapped) {  // (no pool: j == 0 is the whole range)
    b = __ldg(bound + c), e = j == 0 ? __ldg(bound + c + 1) : b;
    return;
  }
  const int r0 = c * h.per, r1 = min(r0 + h.per, h.total);
  if (j == 0) {
    b = r0, e = min(r0 + h.per_static, r1);
  } else {
    b = r0 + h.per_static + h.off[j - 1], e = min(b + h.size[j - 1], r1);
  }
}
// Where the partial of the run that part (c, j) = [b, ..) contributes to the item starting at flat i0 goes.  A part
// can hold two partial runs - the end of one item at its start, the start of another at its end - and whole items
// in between (written directly): two slots per part, chosen by whether the item began before the part did.
__device__ __forceinline__ int part_slot(const PlanHeader& h, int c, int j, int b, int i0) {
  return (j == 0 ? 2 * c : 2 * h.grid + 2 * ((j - 1) * h.grid + c)) + (b < i0 ? 1 : 0);
}

// A run: consecutive steps of ONE (sequence, head group) item, produced by the producer warp for the consumers.
struct RunDesc {  // 48 bytes in shared memory
  int kind;       // 1 = run, 0 = no more work
  int hg, step0, seq_len, tq, item;
  int slot;       // where this run's partial goes if the run turns out not to be the whole item
  int i0, n;      // flat position of the item's first step, its number of steps
  int pad[3];     // (how long a run is shows in step_last[]: its owner may lose its tail chunks to a thief on the way)
};
// A partial run handed to the ticket warp, and (the same fields) an item whose last ticket was taken, handed
// back to the consumers for merging.
struct FinDesc {  // 32 bytes in shared memory
  int kind;       // 1 = a partial was written, 0 = the consumers are done
  int item, hg, tq;
  int steps;      // steps the partial covers: the item's counter adds them up, n of n completes it
  int i0, n;
  int pad;
};
constexpr int kFinSlots = 2;   // partial runs waiting for their ticket
constexpr int kReqSlots = 4;   // items waiting to be merged
static_assert(sizeof(FinDesc) == 32, "layout");
static_assert(sizeof(RunDesc) == 48 && sizeof(PlanCursor) == 32 && sizeof(PlanHeader) == 64, "layout");

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
  int out_vec;     // 1: 16-bit output rows are 8-byte aligned (four elements leave as one store)
  const Plan* plan;
  float* ws_acc;   // [slots][hpc*group][D]
  float* ws_ml;    // [slots][hpc*group][2]  (max in log2 units, exp-sum)
  int* ws_cnt;     // [items] tickets taken on an item split across parts (0 between calls)
  int* ws_pool;    // [1] CTAs that are done taking chunks, [kDequeWord0 + c] claim word of range c (all 0 between calls)
  unsigned long long* trace;  // debug time stamps, 16 per CTA (b200attn_debug_set_decode_trace) or nullptr
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
  int tq;                // first query token of sequence b
  int flat;              // flat index of the current step
};

// visible steps of sequence b: [first, first + n)
__device__ __forceinline__ void seq_steps(const int32_t* seq_lens, const int32_t* cu_seqlens_q, int only_q1, int window,
                                          int tps_shift, int b, int& seq_len, int& first, int& n, int& tq) {
  seq_len = seq_lens[b];
  int c0 = b, c1 = b + 1;
  if (cu_seqlens_q) c0 = cu_seqlens_q[b], c1 = cu_seqlens_q[b + 1];
  tq = c0;
  const int q_len = c1 - c0;
  if (seq_len <= 0 || q_len <= 0 || (only_q1 && q_len != 1)) {
    first = 0, n = 0;
    return;
  }
  int lo_tok = 0;
  if (window > 0) lo_tok = max(0, seq_len - window);
  first = lo_tok >> tps_shift;
  n = ((seq_len + (1 << tps_shift) - 1) >> tps_shift) - first;
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
    seq_steps(P.a.seq_lens, P.a.cu_seqlens_q, P.only_q1, P.a.sliding_window, P.tps_shift, c.b, c.seq_len, c.first, c.n, c.tq);
    if (c.n > 0) {
      c.step = c.first;
      return true;
    }
  }
  return false;
}

// ---- plan kernel: one CTA ------------------------------------------------------------------------
struct PlanArgs {
  const int32_t* seq_lens;
  const int32_t* cu_seqlens_q;
  Plan* plan;
  int num_seqs, only_q1, window, tps_shift, head_groups, grid;
  int pool_sixteenths;  // share of a CTA's steps that goes to the pool, in 1/16
  int min_chunk;        // smallest chunk, steps
  int no_snap;          // A/B: keep the ranges' boundaries where the division puts them
};

constexpr int kPlanThreads = 1024;

__global__ void __launch_bounds__(kPlanThreads) decode_plan_kernel(const PlanArgs A) {
  __shared__ int s_warp[kPlanThreads / 32];
  __shared__ int s_carry, s_zero;
  __shared__ PlanHeader s_hdr;
  pdl_launch_dependents();  // the decode kernel behind this one sets itself up meanwhile (it waits before it reads)
  pdl_wait();
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  int* prefix = reinterpret_cast<int*>(A.plan + 1);  // [num_seqs + 1], steps (not multiplied by head_groups)
  if (tid == 0) s_carry = 0, s_zero = 0;
  __syncthreads();
  // ---- exclusive scan of the per-sequence step counts, kPlanThreads sequences at a time
  for (int b0 = 0; b0 < A.num_seqs; b0 += kPlanThreads) {
    const int b = b0 + tid;
    int n = 0;
    if (b < A.num_seqs) {
      int sl, first, tq;
      seq_steps(A.seq_lens, A.cu_seqlens_q, A.only_q1, A.window, A.tps_shift, b, sl, first, n, tq);
      const int ql = A.cu_seqlens_q ? A.cu_seqlens_q[b + 1] - A.cu_seqlens_q[b] : 1;
      if (sl <= 0 && ql == 1) s_zero = 1;
    }
    int incl = n;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int v = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += v;
    }
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    if (warp == 0) {
      int w = s_warp[lane];
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int v = __shfl_up_sync(0xffffffffu, w, o);
        if (lane >= o) w += v;
      }
      s_warp[lane] = w;  // inclusive over warps
    }
    __syncthreads();
    const int base = s_carry + (warp > 0 ? s_warp[warp - 1] : 0);
    if (b < A.num_seqs) prefix[b] = base + incl - n;
    __syncthreads();
    if (tid == 0) s_carry += s_warp[kPlanThreads / 32 - 1];
    __syncthreads();
  }
  if (tid == 0) {
    const int total_seq_steps = s_carry;
    prefix[A.num_seqs] = total_seq_steps;
    PlanHeader h = {};
    const long long tot = (long long)total_seq_steps * A.head_groups;
    h.total = tot > 0x7fffffffll ? 0x7fffffff : (int)tot;  // (a batch of 2^31 steps does not exist)
    h.grid = A.grid;
    h.has_zero_rows = s_zero;
    h.per = (h.total + A.grid - 1) / A.grid;
    int pool_per = (h.per * A.pool_sixteenths) >> 4;  // tail of every range that goes to the pool
    if (pool_per < 2 * A.min_chunk || A.pool_sixteenths <= 0) pool_per = 0;  // too little work to be worth sharing
    if (pool_per > 32000) pool_per = 32000;  // (sizes are shorts)
    h.per_static = h.per - pool_per;
    // chunk sizes halve from round to round down to min_chunk; the last round takes what is left
    int rem = pool_per, nr = 0;
    while (rem > 0 && nr < kPoolRounds) {
      int c = (rem + 1) / 2;
      if (c < A.min_chunk) c = A.min_chunk;
      if (nr == kPoolRounds - 1 || rem - c < A.min_chunk) c = rem;
      h.size[nr] = (short)c, h.off[nr] = (short)(pool_per - rem);
      rem -= c, ++nr;
    }
    h.n_rounds = nr, h.n_chunks = nr * A.grid;
    h.snapped = (pool_per == 0 && h.per >= 4 && !A.no_snap) ? 1 : 0;
    s_hdr = h;
    A.plan->hdr = h;
  }
  __syncthreads();
  const PlanHeader h = s_hdr;
  // the item a flat position lies in: sequence, its steps, and the position's offset from the sequence's first item
  auto locate = [&](int f, int& b, int& sl, int& first, int& n, int& tq, int& rem) {
    // largest b with prefix[b] * HG <= f; sequences without steps share their successor's prefix and
    // are skipped by taking the LAST such b... they have n = 0, so take the last b whose range is not empty
    int lo = 0, hi = A.num_seqs;  // invariant: prefix[lo] * HG <= f < prefix[hi] * HG
    while (hi - lo > 1) {
      const int mid = (lo + hi) >> 1;
      if ((long long)prefix[mid] * A.head_groups <= f) lo = mid; else hi = mid;
    }
    seq_steps(A.seq_lens, A.cu_seqlens_q, A.only_q1, A.window, A.tps_shift, lo, sl, first, n, tq);
    b = lo, rem = f - prefix[lo] * A.head_groups;  // < n * HG
  };
  if (h.snapped) {
    for (int c = tid; c <= A.grid; c += kPlanThreads) {
      long long fl = (long long)c * h.per;
      int f = fl > h.total ? h.total : (int)fl;
      if (f > 0 && f < h.total) {
        int b, sl, first, n, tq, rem;
        locate(f, b, sl, first, n, tq, rem);
        const int off = rem % n;          // steps of the item in front of f
        if (off == 1) f -= 1;             // the item begins one step earlier: the range takes all of it
        else if (off == n - 1) f += 1;    // the item ends one step later: the range in front keeps all of it
      }
      A.plan->bound[c] = f;
    }
    __syncthreads();  // (the block's own global writes are visible to it after the barrier)
  }
  // ---- resolve the start of every CTA's static part and of every pool chunk
  for (int t = tid; t < A.grid + h.n_chunks; t += kPlanThreads) {
    int f, e;
    if (t < A.grid) part_range(h, A.plan->bound, t, 0, f, e);
    else part_range(h, A.plan->bound, (t - A.grid) % A.grid, 1 + (t - A.grid) / A.grid, f, e);
    const bool any = f < e;
    PlanCursor c;
    c.b = -1, c.hg = 0, c.step = 0, c.seq_len = 0, c.first = 0, c.n = 0, c.tq = 0, c.flat = f;
    if (any) {
      int lo, sl, first, n, tq, rem;
      locate(f, lo, sl, first, n, tq, rem);
      c.b = lo, c.hg = rem / n, c.step = first + rem % n, c.seq_len = sl, c.first = first, c.n = n, c.tq = tq;
    }
    (t < A.grid ? A.plan->cta[t] : A.plan->chunk[t - A.grid]) = c;
  }
}

// KVQ: 0 = 16-bit cache, 1 = e4m3 with per-tensor scales, 2 = e4m3 and 3 = int8 with per-(token, head) scales
template <int D, int KVQ, bool MMA_BF16, bool Q_BF16, int NT, int NW, int TT>
__global__ void __launch_bounds__((NW + 2) * 32, 1)
attn_decode_kernel(const DecodeParams P, const __grid_constant__ CUtensorMap k_map,
                   const __grid_constant__ CUtensorMap v_map) {
  constexpr bool KV_FP8 = KVQ != 0;  // one-byte cache elements (the name predates int8)
  constexpr bool PTH = KVQ >= 2;     // per-(token, head) scales
  constexpr bool KV_I8 = KVQ == 3;
  constexpr int ESZ = KV_FP8 ? 1 : 2;
  constexpr int NDC = D * ESZ / 128;      // 128-byte chunks per row
  constexpr int KSTEPS = D / 16;          // k16 steps of S = K Q^T
  constexpr int MTILES = D / 16;          // m16 tiles of O^T = V^T P^T
  constexpr int SUB = TT / 16;             // 16-token MMA sub-tiles of a warp's TT-token tile
  constexpr int TILE_BYTES = TT * D * ESZ;
  constexpr int KV_STAGE_BYTES = NW * TILE_BYTES;  // K (or V) bytes per stage
  static_assert(TT == 16 || TT == 32, "warp tile is one or two 16-token MMA tiles");
  static_assert(NDC >= 1, "row must be at least 128 bytes");

  // The stage ring needs 1024-byte alignment (128-byte swizzle) and the declared alignment of the dynamic array is
  // honoured; three 64 KB stages + the 32 KB merge buffer leave 3 KB of the 227 KB for everything else, so there
  // is no slack to re-align with (a misaligned base traps instead of corrupting the swizzle).
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = smem_raw;
  if (smem_u32(smem_raw) & 1023u) __trap();
  // [stages][K | V] then merge buffer, then barriers / run descriptors
  uint8_t* stage_base = smem;
  float* mbuf = reinterpret_cast<float*>(smem + (size_t)P.stages * 2 * KV_STAGE_BYTES);  // [NW][8][D]
  float* ml_buf = mbuf + NW * 8 * D;                                                      // [NW][8][2]
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(ml_buf + NW * 8 * 2);
  uint64_t* empty_bar = full_bar + kMaxStages;
  uint64_t* desc_bar = empty_bar + kMaxStages;                     // [kDescSlots]
  RunDesc* descs = reinterpret_cast<RunDesc*>(desc_bar + kDescSlots);  // [kDescSlots]
  uint64_t* fin_full = reinterpret_cast<uint64_t*>(descs + kDescSlots);  // [kFinSlots] consumers -> ticket warp
  uint64_t* fin_empty = fin_full + kFinSlots;                             // [kFinSlots] ticket warp -> consumers
  uint64_t* epi_done = fin_empty + kFinSlots;                             // [1] the ticket warp has taken every ticket
  FinDesc* fins = reinterpret_cast<FinDesc*>(epi_done + 1);               // [kFinSlots]
  FinDesc* reqs = fins + kFinSlots;                                       // [kReqSlots] items to merge
  volatile int* req_cnt = reinterpret_cast<volatile int*>(reqs + kReqSlots);  // [0] posted by the ticket warp, [1] merged, [2] scratch
  int* step_last = const_cast<int*>(req_cnt) + 4;                         // [kMaxStages] 1: the run ends with this stage
  short* part_slots = reinterpret_cast<short*>(step_last + kMaxStages);   // [kMaxParts] the parts of the item being merged

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int gid = lane >> 2, t4 = lane & 3;

  unsigned long long* trace = P.trace ? P.trace + (size_t)blockIdx.x * 16 : nullptr;
  if (trace && threadIdx.x == 0) {
    trace[0] = global_ns(), trace[1] = clock64(), trace[7] = 0;
    for (int i = 8; i < 16; ++i) trace[i] = 0;
  }

  // ---- barrier init ---------------------------------------------------------------------------
  if (warp == NW && lane == 0) {  // descriptor fetch overlaps the rest of the prologue
    tma_prefetch_desc(&k_map);
    tma_prefetch_desc(&v_map);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < P.stages; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], NW);
    }
    for (int s = 0; s < kDescSlots; ++s) mbar_init(&desc_bar[s], 1);
    for (int s = 0; s < kFinSlots; ++s) {
      mbar_init(&fin_full[s], NW);
      mbar_init(&fin_empty[s], 1);
    }
    mbar_init(epi_done, 1);
    req_cnt[0] = 0, req_cnt[1] = 0, req_cnt[2] = 0;
    fence_barrier_init();
  }
  // Whatever is launched behind this kernel with the programmatic-serialisation attribute (the next layer's
  // cache write, the next decode launch) may be scheduled as this grid's CTAs retire; it waits for this grid
  // to complete before it touches memory (common.cuh).
  pdl_launch_dependents();
  __syncthreads();  // mbarrier init visible to every warp
  // Everything above touched only shared memory and kernel parameters.  From here on the kernel reads what
  // earlier kernels on the stream wrote (the plan, metadata, q, the cache): wait for them (no-op without PDL).
  pdl_wait();

  const int boxb = P.hpc * NDC * P.box_slots * 128;  // bytes of one TMA box (all heads, all chunks)

  if (warp == NW) {
    // ======================= producer warp ====================================
    // Walks this CTA's work, cuts it into runs (consecutive steps of one item), publishes a descriptor per run and
    // streams every step's pages: lane i looks up the page of box i and launches its K and V copies, so page-table
    // latency and TMA issue overlap across boxes.
    //
    // The work: first this CTA's own range - the static part, then the range's tail chunks for as long as no other
    // CTA has taken them (claimed one chunk ahead, front to back, so the walk simply continues: no run ends at a
    // chunk boundary) - and when that is exhausted, chunks stolen from the BACK of the range with the most chunks
    // left.  Each range has one packed word in the workspace: owner claims | thief claims | mask of stolen chunks.
    const uint64_t pol = policy_evict_first();
    const PlanHeader hdr = P.plan->hdr;
    int stage = 0;
    uint32_t phase = 0;
    int ndesc = 0;
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
    auto load_cursor = [&](const PlanCursor* pc, ItemCursor& c) {
      const int4 lo = *reinterpret_cast<const int4*>(pc), hi = *(reinterpret_cast<const int4*>(pc) + 1);
      c.b = lo.x, c.hg = lo.y, c.step = lo.z, c.seq_len = lo.w, c.first = hi.x, c.n = hi.y, c.tq = hi.z, c.flat = hi.w;
    };
    const int T = hdr.n_rounds;
    const bool pooled = T > 0;
    int* deque = P.ws_pool + kDequeWord0;
    // chunks of range c that are not cut off by the end of the batch
    auto chunks_of = [&](int c) {
      int n = 0;
      for (int j = 1; j <= T; ++j) {
        int pb, pe;
        part_range(hdr, P.plan->bound, c, j, pb, pe);
        n += pb < pe ? 1 : 0;
      }
      return n;
    };

    ItemCursor cur = {};
    bool run_open = false;  // a run has been published and its last step not yet issued
    int slot_c = 0, slot_j = 0, slot_b = 0;  // the part the current extent belongs to (for the partial's slot)
    // ---- one step: wait for a free stage, publish the run if this is its first step, issue the copies
    auto issue_step = [&](bool last) {
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
      {
        const long long t0 = trace ? clock64() : 0;
        mbar_wait(&empty_bar[stage], phase ^ 1);
        if (trace && lane == 0) trace[10] += (unsigned long long)(clock64() - t0);
      }
      if (!run_open) {
        // Publish the run.  Its slot held descriptor ndesc - kDescSlots; the stage just waited for was last used
        // by a step of run >= ndesc - stages, and a consumer reads a run's descriptor before it releases any of
        // the run's stages, so every consumer is done with the old contents (kDescSlots > kMaxStages + 1).
        const int ds = ndesc & (kDescSlots - 1);
        if (lane == 0) {
          const int i0 = cur.flat - (cur.step - cur.first);  // flat position of the item's first step
          RunDesc rd = {};
          rd.kind = 1, rd.hg = cur.hg, rd.step0 = cur.step, rd.seq_len = cur.seq_len, rd.tq = cur.tq;
          rd.item = cur.b * P.head_groups + cur.hg, rd.i0 = i0, rd.n = cur.n;
          rd.slot = part_slot(hdr, slot_c, slot_j, slot_b, i0);  // used if the run turns out not to be the whole item
          descs[ds] = rd;
          mbar_arrive(&desc_bar[ds]);  // release: the stores above are visible to whoever sees the phase flip
        }
        ++ndesc;
        run_open = true;
      }
      uint8_t* kdst = stage_base + (size_t)stage * 2 * KV_STAGE_BYTES;
      uint8_t* vdst = kdst + KV_STAGE_BYTES;
      if (lane == 0) {
        step_last[stage] = last ? 1 : 0;  // the consumers end the run after this stage
        mbar_arrive_expect_tx(&full_bar[stage], (uint32_t)(2 * nbox * boxb));
      }
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
      if (last) run_open = false;
    };

    // ---- this CTA's own range: the static part, then its tail chunks while nobody else has taken them
    {
      const int c = (int)blockIdx.x;
      int extent_end;
      load_cursor(&P.plan->cta[c], cur);
      part_range(hdr, P.plan->bound, c, 0, slot_b, extent_end);
      slot_c = c, slot_j = 0;
      if (cur.b >= 0 && cur.flat < extent_end) {
        const int tc = pooled ? chunks_of(c) : 0;
        int own_next = 0;      // next own chunk to claim
        int claim_old = 0;     // lane 0: the packed word before the claim in flight
        bool claim_pending = false;
        auto claim = [&]() {
          if (own_next < tc) {
            if (lane == 0) claim_old = atomicAdd(&deque[c], 1);
            claim_pending = true;
          }
        };
        claim();  // chunk 0: the answer is back long before the static part ends
        while (true) {
          if (cur.flat + 1 == extent_end && claim_pending) {
            // the extent ends with this step unless the next own chunk is still there
            int ok = 0;
            if (lane == 0) {
              ok = ((claim_old & 0xff) + ((claim_old >> 8) & 0xff) < tc) ? 1 : 0;
              if (!ok) atomicSub(&deque[c], 1);  // (a failed claim is taken back: the counts stay exact)
            }
            ok = __shfl_sync(0xffffffffu, ok, 0);
            claim_pending = false;
            if (ok) {
              int pb, pe;
              part_range(hdr, P.plan->bound, c, 1 + own_next, pb, pe);  // begins where the extent ends
              extent_end = pe;
              ++own_next;
              claim();
            }
          }
          const bool item_last = cur.step + 1 == cur.first + cur.n;
          const bool ext_last = cur.flat + 1 == extent_end;
          issue_step(item_last || ext_last);
          if (ext_last) break;
          cursor_next(P, cur);
        }
      }
    }
    // ---- then chunks stolen from the back of other ranges
    while (pooled) {
      const long long t_sw = trace ? clock64() : 0;
      // the range with the most chunks left (ties: the nearest one after this CTA, so thieves spread out)
      int best_key = 0;
      for (int c0 = 0; c0 < hdr.grid; c0 += 32) {
        const int c = c0 + lane;
        if (c < hdr.grid) {
          const int w = __ldcg(&deque[c]);
          const int rem = chunks_of(c) - (w & 0xff) - ((w >> 8) & 0xff);
          int dist = c - (int)blockIdx.x - 1;
          if (dist < 0) dist += hdr.grid;
          if (rem > 0) best_key = max(best_key, (rem << 20) | ((0x3ff - dist) << 10) | c);
        }
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) best_key = max(best_key, __shfl_xor_sync(0xffffffffu, best_key, o));
      if (best_key == 0) break;  // nothing left anywhere that is not already being worked on
      const int v = best_key & 0x3ff;
      const int tv = chunks_of(v);
      int j = -1;
      if (lane == 0) {
        const int old = atomicAdd(&deque[v], 1 << 8);
        const int f = old & 0xff, bk = (old >> 8) & 0xff;
        if (f + bk < tv) {
          j = tv - 1 - bk;
          atomicOr(&deque[v], 1 << (16 + j));  // the merge needs to know whose partial covers this chunk
        } else {
          atomicSub(&deque[v], 1 << 8);
        }
      }
      j = __shfl_sync(0xffffffffu, j, 0);
      if (j < 0) continue;  // lost the race: look again
      int cb, ce;
      part_range(hdr, P.plan->bound, v, 1 + j, cb, ce);
      load_cursor(&P.plan->chunk[j * hdr.grid + v], cur);
      slot_c = v, slot_j = 1 + j, slot_b = cb;
      if (trace && lane == 0) trace[11] += (unsigned long long)(clock64() - t_sw);
      if (cur.b < 0) continue;
      while (true) {
        const bool item_last = cur.step + 1 == cur.first + cur.n;
        const bool ext_last = cur.flat + 1 == ce;
        issue_step(item_last || ext_last);
        if (ext_last) break;
        cursor_next(P, cur);
      }
    }
    // no more work
    {
      const int ds = ndesc & (kDescSlots - 1);
      if (lane == 0) {
        descs[ds].kind = 0;
        mbar_arrive(&desc_bar[ds]);
      }
    }
    if (pooled && lane == 0) {
      // the CTA that is last to stop taking chunks rewinds the range words for the next launch
      __threadfence();
      if (atomicAdd(&P.ws_pool[1], 1) == (int)gridDim.x - 1) {
        __threadfence();
        for (int c = 0; c < hdr.grid; ++c) deque[c] = 0;
        P.ws_pool[1] = 0;
      }
    }
    return;
  }

  if (warp == NW + 1) {
    // ======================= ticket warp ====================================
    // Takes the ticket of every partial the consumers hand over - a fence and an atomic round trip the consumers
    // would otherwise sit through at every partial run (a bubble of several microseconds in this SM's HBM stream:
    // profiles/r02_kbench_pool_v1.jsonl).  The part that takes an item's LAST ticket must merge the item: that is
    // handed back to the eight consumer warps, which do it together at their next run boundary (one warp takes
    // 10 us for a ten-part item, profiles/r02_kbench_pool_v3_tail_chunks.jsonl).
    int nreq = 0;
    for (int nf = 0;; ++nf) {
      const int fs = nf % kFinSlots;
      mbar_wait(&fin_full[fs], ((uint32_t)nf / kFinSlots) & 1u);
      const FinDesc f = fins[fs];
      __syncwarp();
      if (lane == 0) mbar_arrive(&fin_empty[fs]);
      if (f.kind == 0) break;
      int ticket = 0;
      if (lane == 0) {
        __threadfence();  // the consumers' partial stores (ordered before this by the hand-off) before the ticket
        ticket = atomicAdd(&P.ws_cnt[f.item], f.steps) + f.steps;  // steps of the item that are done
        if (ticket == f.n) {
          __threadfence();  // every other part's partial is visible from here on
          while (nreq - req_cnt[1] >= kReqSlots) {}  // (the consumers serve requests at every run boundary)
          reqs[nreq % kReqSlots] = f;
          __threadfence_block();
          req_cnt[0] = nreq + 1;
        }
      }
      ticket = __shfl_sync(0xffffffffu, ticket, 0);
      if (ticket == f.n) ++nreq;
    }
    if (lane == 0) mbar_arrive(epi_done);  // release: every request was posted before this
    if (trace && lane == 0) trace[5] = clock64(), trace[6] = global_ns();
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

  // Per-(token, head) scales (KVQ >= 2): rows gid and gid + 8 of each 16-token sub-tile need the key's and the
  // value's scale.  The consumers fetch them themselves - the stage ring has no room left for them - straight from
  // the scale caches: a sub-tile has 16 key and 16 value scales, ONE per lane (lanes 0-15 the keys' of tokens
  // 0-15, lanes 16-31 the values'), handed to the lanes that need them by shuffles when the step begins.  Two
  // loads deep: at the top of step s the scales of step s + 1 are requested (their page ids were requested a step
  // earlier) and the page ids of step s + 2, so neither the block-table lookup nor the scale load is waited for
  // before a whole step has gone by.  (First version: every lane fetched its own four values at the top of the
  // step that used them - half the per-tensor kernel's rate, profiles/r02_kbench_per_token_head_v1.jsonl; with the
  // prefetch but four loads and their 64-bit addresses per lane, 190 instructions a step on top of 700.)
  int pg_a[SUB], pg_b[SUB];                 // page of each sub-tile, steps s + 1 and s + 2
  float ksc[SUB][2], vsc[SUB][2];           // scales of step s (rows gid, gid + 8)
  float sc_n[SUB];                          // step s + 1, in flight: this lane's one scale per sub-tile
#pragma unroll
  for (int sb = 0; sb < SUB; ++sb) {
    pg_a[sb] = pg_b[sb] = -1, sc_n[sb] = 1.f;
    ksc[sb][0] = ksc[sb][1] = vsc[sb][0] = vsc[sb][1] = 1.f;
  }
  // Per step this costs the consumers ~40 instructions: everything that does not change with the step is folded
  // into per-run state (block-table row, the head's base address) or per-kernel state (this lane's cache, its
  // strides in BYTES as 32-bit values, so that page * block_stride + base is ONE 32 x 32 + 64-bit multiply-add),
  // and the token positions of steps s + 1 / s + 3 are sums, not products.  (With 64-bit strides and the
  // positions recomputed from the step number it was ~110: profiles/r02_ncu_decode_per_token_head.txt.)
  const int bs_mask = P.a.block_size - 1;
  const int32_t* bt_run = P.a.block_tables;  // block-table row of the current run's sequence
  auto load_pages = [&](int tok_first, int seq_len, int (&pg)[SUB]) {  // tok_first: this warp's first token of that step
#pragma unroll
    for (int sb = 0; sb < SUB; ++sb) {
      const int tok = tok_first + 16 * sb;  // 16-aligned and block_size >= 16: one page per sub-tile
      int v = 0;
      if (tok < seq_len) v = __ldg(bt_run + (tok >> P.bs_shift));
      pg[sb] = v;
    }
  };
  const char* sc_lane = nullptr;   // this lane's cache (keys' or values' scales) + its row inside a sub-tile + the warp's head
  const char* sc_head = nullptr;   // ... + the run's head group
  uint32_t sc_block_bytes = 0, sc_slot_bytes = 0;
  int64_t sc_group_bytes = 0;      // bytes between head groups
  const int l15 = lane & 15;
  if constexpr (PTH) {
    const bool is_v = lane >= 16;
    const int64_t hs = is_v ? P.a.vs_head_stride : P.a.ks_head_stride;
    sc_block_bytes = (uint32_t)(is_v ? P.a.vs_block_stride : P.a.ks_block_stride) * 4u;  // < 2^31: decode_supported()
    sc_slot_bytes = (uint32_t)(is_v ? P.a.vs_slot_stride : P.a.ks_slot_stride) * 4u;
    sc_group_bytes = (int64_t)P.hpc * hs * 4;
    sc_lane = reinterpret_cast<const char*>(is_v ? P.a.v_scale_cache : P.a.k_scale_cache) + (int64_t)l15 * sc_slot_bytes +
              (int64_t)h_local * hs * 4;
    sc_head = sc_lane;
  }
  auto load_scales = [&](int tok_first, int seq_len, const int (&pg)[SUB]) {  // -> sc_n
#pragma unroll
    for (int sb = 0; sb < SUB; ++sb) {
      const int tok = tok_first + 16 * sb;  // sub-tile's first token
      const char* sp = sc_head + (uint64_t)(uint32_t)pg[sb] * sc_block_bytes + (uint32_t)(tok & bs_mask) * sc_slot_bytes;
      float v = 1.f;
      if (tok + l15 < seq_len) v = __ldg(reinterpret_cast<const float*>(sp));
      sc_n[sb] = v;
    }
  };
  auto take_scales = [&]() {  // sc_n (requested a step ago) -> ksc / vsc of the rows this lane works on
#pragma unroll
    for (int sb = 0; sb < SUB; ++sb) {
      ksc[sb][0] = __shfl_sync(0xffffffffu, sc_n[sb], gid), ksc[sb][1] = __shfl_sync(0xffffffffu, sc_n[sb], gid + 8);
      vsc[sb][0] = __shfl_sync(0xffffffffu, sc_n[sb], gid + 16), vsc[sb][1] = __shfl_sync(0xffffffffu, sc_n[sb], gid + 24);
    }
  };
  // start of a run: its block-table row and head base, the pages of its first three steps
  auto prime_pages = [&](int item, int hg, int step0, int seq_len, int (&pg0)[SUB]) {
    bt_run = P.a.block_tables + (int64_t)(item / P.head_groups) * P.a.bt_stride0;  // (a division: once per run)
    sc_head = sc_lane + (int64_t)hg * sc_group_bytes;
    const int t0 = step0 * P.tps + tok_off;
    load_pages(t0, seq_len, pg0);
    load_pages(t0 + P.tps, seq_len, pg_a);
    load_pages(t0 + 2 * P.tps, seq_len, pg_b);
  };

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

  auto load_q = [&](int tq, int hg) {
    const int kvh = hg * P.hpc + h_local;
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

  // sequences with one query token and no visible key get zeros (reference reduce_segments :828); no run visits them
  if (P.plan->hdr.has_zero_rows) {
    for (int b = blockIdx.x; b < P.a.num_seqs; b += gridDim.x) {
      const int sl = P.a.seq_lens[b];
      const int ql = P.a.cu_seqlens_q ? (P.a.cu_seqlens_q[b + 1] - P.a.cu_seqlens_q[b]) : 1;
      if (sl <= 0 && ql == 1) {
        const int tq = P.a.cu_seqlens_q ? P.a.cu_seqlens_q[b] : b;
        if (P.a.lse)
          for (int h = threadIdx.x; h < P.a.num_q_heads; h += NW * 32) P.a.lse[(int64_t)tq * P.a.lse_stride0 + h] = -INFINITY;
        for (int e = threadIdx.x; e < P.a.num_q_heads * D; e += NW * 32) {
          const int osz = P.a.out_scale ? 1 : 2;  // fp8 output: one byte per element, 0x00 == +0
          char* orow = static_cast<char*>(P.a.out) + ((int64_t)tq * P.a.out_stride0 + (int64_t)(e / D) * P.a.out_stride1) * osz;
          if (osz == 1) orow[e % D] = 0;
          else reinterpret_cast<uint16_t*>(orow)[e % D] = 0;
        }
      }
    }
  }

  // ---- merging an item whose last ticket this CTA's ticket warp took (the reference's `reduce_segments` launch,
  // :757-836): all consumer warps together, at a run boundary.  Warp 0 lists the item's parts in flat order - every
  // (range, part) that overlaps [i0, i1), one candidate per lane - then every thread owns (query head, 4 dims)
  // chunks and folds the parts in groups of G with all of a group's loads in flight at once.
  int reqs_done = 0;
  auto merge_item = [&](const FinDesc f) {
    __threadfence();
    const int rec_heads = P.hpc * P.group;
    if (warp == 0) {
      // Who holds a partial of this item: in every range the item overlaps, the range's owner (if its static part
      // or a tail chunk nobody stole overlaps the item) and the thief of every stolen chunk that overlaps it - the
      // range's claim word says which chunks were stolen.  Every overlapping chunk has been processed by now
      // (the item's step count is complete), so "not stolen" means the owner did it.
      const PlanHeader hdr = P.plan->hdr;
      const int T = hdr.n_rounds, J = 1 + T;
      const int* deque = P.ws_pool + kDequeWord0;
      const int i0 = f.i0, i1 = f.i0 + f.n;
      // (snapped boundaries lie within one step of c * per: one more candidate range on either side)
      const int ca = hdr.snapped ? max(0, (i0 - 1) / hdr.per) : i0 / hdr.per;
      const int cb = hdr.snapped ? min(hdr.grid - 1, i1 / hdr.per) : (i1 - 1) / hdr.per;
      const int ncand = (cb - ca + 1) * J;
      int np = 0;
      for (int base = 0; base < ncand; base += 32) {
        const int idx = base + lane;
        bool hit = false;
        int slot = 0;
        if (idx < ncand) {
          const int c = ca + idx / J, j = idx % J;
          const int stolen = T > 0 ? (__ldcg(&deque[c]) >> 16) & 0xff : 0;
          int pb, pe;
          part_range(hdr, P.plan->bound, c, j, pb, pe);
          hit = max(pb, i0) < min(pe, i1);
          if (j == 0) {
            for (int r = 1; r <= T; ++r) {
              int qb, qe;
              part_range(hdr, P.plan->bound, c, r, qb, qe);
              hit = hit || (!((stolen >> (r - 1)) & 1) && max(qb, i0) < min(qe, i1));
            }
          } else {
            hit = hit && ((stolen >> (j - 1)) & 1);
          }
          slot = part_slot(hdr, c, j, pb, i0);
        }
        const unsigned m = __ballot_sync(0xffffffffu, hit);
        if (hit) part_slots[np + __popc(m & ((1u << lane) - 1u))] = (short)slot;
        np += __popc(m);
      }
      if (lane == 0) req_cnt[3] = np;
    }
    asm volatile("bar.sync 1, %0;" ::"r"(NW * 32));
    const int nparts = req_cnt[3];
    constexpr int D4 = D / 4;
    constexpr int G = 8;
    for (int e = threadIdx.x; e < rec_heads * D4; e += NW * 32) {
      const int hq_in = e / D4, d4 = (e % D4) * 4;
      float M = -INFINITY, L = 0.f;
      float4 val = make_float4(0.f, 0.f, 0.f, 0.f);
      for (int p0 = 0; p0 < nparts; p0 += G) {
        float2 ml[G];
        float4 av[G];
#pragma unroll
        for (int g = 0; g < G; ++g) {
          if (p0 + g < nparts) {
            const int64_t rec = (int64_t)part_slots[p0 + g] * rec_heads + hq_in;
            ml[g] = __ldcg(reinterpret_cast<const float2*>(&P.ws_ml[rec * 2]));
            av[g] = __ldcg(reinterpret_cast<const float4*>(&P.ws_acc[rec * D + d4]));
          } else {
            ml[g] = make_float2(-INFINITY, 0.f), av[g] = make_float4(0.f, 0.f, 0.f, 0.f);
          }
        }
        float mg = M;
#pragma unroll
        for (int g = 0; g < G; ++g) mg = fmaxf(mg, ml[g].x);
        const float so = (M == -INFINITY) ? 0.f : ex2(M - mg);  // (mg = -inf only if M is)
        val.x *= so, val.y *= so, val.z *= so, val.w *= so, L *= so;
#pragma unroll
        for (int g = 0; g < G; ++g) {
          const float sc = (ml[g].x == -INFINITY) ? 0.f : ex2(ml[g].x - mg);
          val.x += av[g].x * sc, val.y += av[g].y * sc, val.z += av[g].z * sc, val.w += av[g].w * sc;
          L += ml[g].y * sc;
        }
        M = mg;
      }
      const int hq = f.hg * rec_heads + hq_in;
      float fsink = 1.f;
      const bool seen = M != -INFINITY;
      if (P.a.sinks) fsink = sink_rescale<true>(P.a.sinks[hq], M, L);  // once per item, here (upstream: segment 0 only)
      const float inv = (L > 0.f) ? fsink * v_scale / L : 0.f;  // 0 when nothing is visible (reference :828)
      if (P.a.lse && d4 == 0) P.a.lse[(int64_t)f.tq * P.a.lse_stride0 + hq] = seen ? row_lse<true>(M, L) : -INFINITY;
      const int64_t oel = (int64_t)f.tq * P.a.out_stride0 + (int64_t)hq * P.a.out_stride1;
      const float o4[4] = {val.x * inv, val.y * inv, val.z * inv, val.w * inv};
      if (P.a.out_scale) {
        const float oq = out_quant_factor(P.a.out_scale);
        *reinterpret_cast<uint32_t*>(static_cast<char*>(P.a.out) + oel + d4) = fp8x4_out(o4[0] * oq, o4[1] * oq, o4[2] * oq, o4[3] * oq);
      } else {
        char* orow = static_cast<char*>(P.a.out) + (oel + d4) * 2;
        if (P.out_vec) {
          *reinterpret_cast<uint2*>(orow) = make_uint2(pack2<Q_BF16>(o4[0], o4[1]), pack2<Q_BF16>(o4[2], o4[3]));
        } else {  // rows that are not 8-byte aligned (a view into a wider buffer): element stores
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            if constexpr (Q_BF16) reinterpret_cast<__nv_bfloat16*>(orow)[i] = __float2bfloat16_rn(o4[i]);
            else reinterpret_cast<__half*>(orow)[i] = __float2half_rn(o4[i]);
          }
        }
      }
    }
    if (threadIdx.x == 0) P.ws_cnt[f.item] = 0;  // self-cleaning: ready for the next call
    asm volatile("bar.sync 1, %0;" ::"r"(NW * 32));  // part_slots is rewritten by the next merge
  };
  // Serves the ticket warp's merge requests; every consumer warp calls it at the same points.  Returns true once
  // the ticket warp has finished (it posts every request before it says so).
  auto serve_requests = [&]() {
    if (threadIdx.x == 0) {
      const int fin = mbar_try_wait(epi_done, 0) ? 1 : 0;  // first: what is posted before this is counted below
      req_cnt[2] = req_cnt[0] | (fin << 30);
    }
    asm volatile("bar.sync 1, %0;" ::"r"(NW * 32));
    const int v = req_cnt[2];
    asm volatile("bar.sync 1, %0;" ::"r"(NW * 32));  // everyone has read it before thread 0 writes the next one
    const int upto = v & 0x3fffffff;
    while (reqs_done < upto) {
      const FinDesc f = reqs[reqs_done % kReqSlots];
      merge_item(f);
      ++reqs_done;
      if (threadIdx.x == 0) req_cnt[1] = reqs_done;
    }
    return (v >> 30) != 0;
  };

  // hand-off to the ticket warp (kFinSlots slots; every consumer warp arrives, warp 0 writes the descriptor)
  RunDesc run = {};
  int nfin = 0;
  auto publish_fin = [&](int kind, int item, int hg, int tq, int steps) {
    const int fs = nfin % kFinSlots;
    mbar_wait(&fin_empty[fs], (((uint32_t)nfin / kFinSlots) & 1u) ^ 1u);  // the slot's previous hand-off has been taken
    if (warp == 0 && lane == 0) {
      FinDesc& f = fins[fs];
      f.kind = kind, f.item = item, f.hg = hg, f.tq = tq, f.steps = steps, f.i0 = run.i0, f.n = run.n;
    }
    __syncwarp();
    if (lane == 0) mbar_arrive(&fin_full[fs]);  // release: this warp's partial stores (and the descriptor) come first
    ++nfin;
  };

  // ---- runs ------------------------------------------------------------------------------------
  struct { int hg, step, seq_len; } cur;  // the step being consumed
  int nd = 0;                             // descriptors taken so far
  mbar_wait(&desc_bar[0], 0);
  run = descs[0];
  if (trace && threadIdx.x == 0) trace[2] = clock64();
  if (run.kind == 0) {
    if (trace && threadIdx.x == 0) trace[3] = trace[4] = clock64();
    publish_fin(0, 0, 0, 0, 0);
    return;
  }
  cur.hg = run.hg, cur.step = run.step0, cur.seq_len = run.seq_len;
  reset_state();
  load_q(run.tq, run.hg);
  if constexpr (PTH) {
    int pg0[SUB];
    prime_pages(run.item, run.hg, cur.step, cur.seq_len, pg0);
    load_scales(cur.step * P.tps + tok_off, cur.seq_len, pg0);
  }
  bool first_stage = true;
  int steps_done = 0;  // of the current run

  int stage = 0;
  uint32_t phase = 0;
  while (true) {
    const int tok0 = cur.step * P.tps + tok_off;  // first token of this warp's tile
    if constexpr (PTH) {
      take_scales();  // this step's, requested a step ago
      load_scales(tok0 + P.tps, cur.seq_len, pg_a);
#pragma unroll
      for (int sb = 0; sb < SUB; ++sb) pg_a[sb] = pg_b[sb];
      load_pages(tok0 + 3 * P.tps, cur.seq_len, pg_b);
    }
    {
      const long long t0 = trace ? clock64() : 0;
      mbar_wait(&full_bar[stage], phase);
      if (trace && threadIdx.x == 0 && !first_stage) trace[9] += (unsigned long long)(clock64() - t0);
    }
    const bool run_ends = step_last[stage] != 0;  // (written before the stage was armed)
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
                  kv8x4_to_f16x2<KV_I8, false>(r[2 * half], a[0], a[2]);
                  kv8x4_to_f16x2<KV_I8, false>(r[2 * half + 1], a[1], a[3]);
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
            float x;
            if constexpr (PTH) x = sacc[sb][nt][i] * (qk_scale * ksc[sb][i >> 1]);  // upstream: dot * (scale * k_scale[key])
            else x = sacc[sb][nt][i] * qk_scale;
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
          if (P.a.qq_bias) {
            // a single-token row is query 0 of its step and its own key is the sequence's last: qq_bias[0][0]
            const float b00 = P.a.qq_bias[0] * kLog2e;
            if (tok0 + tr == cur.seq_len - 1) sv[sb][0] += b00, sv[sb][1] += b00;
            if (tok0 + tr + 8 == cur.seq_len - 1) sv[sb][2] += b00, sv[sb][3] += b00;
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
          if constexpr (PTH) {  // upstream: (P * v_scale[key]).to(V.dtype); the exp-sum above takes P itself
            pb[sb][nt][0] = movmatrix_trans(pack2<MMA_BF16>(p0 * vsc[sb][0], p1 * vsc[sb][0]));
            pb[sb][nt][1] = movmatrix_trans(pack2<MMA_BF16>(p2 * vsc[sb][1], p3 * vsc[sb][1]));
          } else {
            pb[sb][nt][0] = movmatrix_trans(pack2<MMA_BF16>(p0, p1));
            pb[sb][nt][1] = movmatrix_trans(pack2<MMA_BF16>(p2, p3));
          }
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
                // bytes [t d0, t d1, t+1 d0, t+1 d1] -> (t d0, t+1 d0) and (t d1, t+1 d1)
                uint32_t a[4];
                kv8x4_to_f16x2<KV_I8, true>(r[2 * half], a[0], a[1]);      // row gid <-> d0 and row gid+8 <-> d1, tokens 2t, 2t+1
                kv8x4_to_f16x2<KV_I8, true>(r[2 * half + 1], a[2], a[3]);  // tokens 2t+8, 2t+9
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

    ++cur.step, ++steps_done;
    if (run_ends) {
      // ---------------- finalise this run ------------------
      const long long t_fin = trace ? clock64() : 0;
      const int item = run.item;
      const int hg_done = run.hg;
      const bool direct = steps_done == run.n;  // the run was the whole item
      const int tq = run.tq;
      // Take the next run first: its query rows are in flight while this one is merged and stored.  The query
      // fragments are dead after the run's last tile; the accumulators are reset only after the merge has
      // read them.  (The producer publishes a run when it issues its first step, stages ahead of this point.)
      ++nd;
      mbar_wait(&desc_bar[nd & (kDescSlots - 1)], (uint32_t)(nd / kDescSlots) & 1u);
      const RunDesc nx = descs[nd & (kDescSlots - 1)];
      const bool last = nx.kind == 0;
      if (!last) load_q(nx.tq, nx.hg);
      int pg0[SUB];
      if constexpr (PTH) {
        if (!last) prime_pages(nx.item, nx.hg, nx.step0, nx.seq_len, pg0);  // in flight while this run is finalised
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
        // Each warp's O^T columns go to mbuf[warp][8 heads][D].  The four lanes t4 of a row group write rows 2 t4
        // (and 2 t4 + 1), D floats apart - the same bank - so the row's position inside its line is XORed with
        // 8 t4 (two float4 groups per step of t4): conflict-free for the 16-bit layout, two-way for the fp8 one.
        float* mb = mbuf + warp * 8 * D;
        const int swz = 8 * t4;
#pragma unroll
        for (int mt = 0; mt < MTILES; ++mt) {
          int d0, d1;
          if constexpr (!KV_FP8) d0 = 16 * mt + gid, d1 = d0 + 8;
          else d0 = 16 * mt + 2 * gid, d1 = d0 + 1;
          mb[(2 * t4) * D + (d0 ^ swz)] = oacc[mt][nt][0];
          mb[(2 * t4 + 1) * D + (d0 ^ swz)] = oacc[mt][nt][1];
          mb[(2 * t4) * D + (d1 ^ swz)] = oacc[mt][nt][2];
          mb[(2 * t4 + 1) * D + (d1 ^ swz)] = oacc[mt][nt][3];
        }
        if (gid == 0) {
          ml_buf[(warp * 8 + 2 * t4) * 2] = m_run[nt][0], ml_buf[(warp * 8 + 2 * t4) * 2 + 1] = l0;
          ml_buf[(warp * 8 + 2 * t4 + 1) * 2] = m_run[nt][1], ml_buf[(warp * 8 + 2 * t4 + 1) * 2 + 1] = l1;
        }
        asm volatile("bar.sync 1, %0;" ::"r"(NW * 32));
        // Combine the warps that share a kv head.  A task is four consecutive d of one (kv head, query head) row:
        // the row's (max, exp-sum) pairs and the wph rescale factors are formed once per task, the accumulators
        // come as float4.  (Per element, as this loop first was, a run's epilogue took 2.5 us of the consumers'
        // time - 4 runs per CTA at 640-token contexts: profiles/r02_kbench_short_ctx_trace.jsonl.)  The sums run
        // over the warps in the same order as before, so results are bit for bit what they were.
        const int nheads = min(8, P.group - nt * 8);
        constexpr int G4 = D / 4;
        for (int e4 = threadIdx.x; e4 < P.hpc * 8 * G4; e4 += NW * 32) {
          const int row = e4 / G4, g4 = e4 % G4;
          const int hl = row >> 3, col = row & 7;
          if (col >= nheads) continue;
          const int pg4 = g4 ^ (((col >> 1) & 3) * 2);  // where the writers put this group
          float M = -INFINITY;
          for (int j = 0; j < wph; ++j) M = fmaxf(M, ml_buf[((hl + P.hpc * j) * 8 + col) * 2]);
          float4 val = make_float4(0.f, 0.f, 0.f, 0.f);
          float L = 0.f;
          for (int j = 0; j < wph; ++j) {
            const int w = hl + P.hpc * j;
            const float2 ml = *reinterpret_cast<const float2*>(&ml_buf[(w * 8 + col) * 2]);
            const float sc = (ml.x == -INFINITY) ? 0.f : ex2(ml.x - M);
            const float4 a = *reinterpret_cast<const float4*>(&mbuf[(w * 8 + col) * D + 4 * pg4]);
            val.x += a.x * sc, val.y += a.y * sc, val.z += a.z * sc, val.w += a.w * sc;
            L += ml.y * sc;
          }
          const int d = 4 * g4;
          const int hq_in_cta = hl * P.group + nt * 8 + col;  // query head within the CTA's head group
          if (direct) {
            const int hq = hg_done * P.hpc * P.group + hq_in_cta;
            const bool seen = M != -INFINITY;  // some key was visible (a sink alone does not make an LSE)
            if (P.a.sinks) {
              const float f = sink_rescale<true>(P.a.sinks[hq], M, L);
              val.x *= f, val.y *= f, val.z *= f, val.w *= f;
            }
            if (P.a.lse && d == 0) P.a.lse[(int64_t)tq * P.a.lse_stride0 + hq] = seen ? row_lse<true>(M, L) : -INFINITY;
            // (a multiplication and a division per element, as before: the same roundings)
            float o[4] = {val.x, val.y, val.z, val.w};
#pragma unroll
            for (int i = 0; i < 4; ++i) o[i] = (L > 0.f) ? o[i] * v_scale / L : 0.f;
            const int64_t oel = (int64_t)tq * P.a.out_stride0 + (int64_t)hq * P.a.out_stride1;
            if (P.a.out_scale) {  // fused e4m3 output quantisation (b200attn.h: out_scale; rows 8-byte aligned)
              const float oq = out_quant_factor(P.a.out_scale);
              const uint32_t w = (uint32_t)fp8_out(o[0] * oq) | ((uint32_t)fp8_out(o[1] * oq) << 8) |
                                 ((uint32_t)fp8_out(o[2] * oq) << 16) | ((uint32_t)fp8_out(o[3] * oq) << 24);
              *reinterpret_cast<uint32_t*>(static_cast<uint8_t*>(P.a.out) + oel + d) = w;
            } else {
              char* orow = static_cast<char*>(P.a.out) + (oel + d) * 2;
              if (P.out_vec) {
                *reinterpret_cast<uint2*>(orow) = make_uint2(pack2<Q_BF16>(o[0], o[1]), pack2<Q_BF16>(o[2], o[3]));
              } else {
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                  if constexpr (Q_BF16) reinterpret_cast<__nv_bfloat16*>(orow)[i] = __float2bfloat16_rn(o[i]);
                  else reinterpret_cast<__half*>(orow)[i] = __float2half_rn(o[i]);
                }
              }
            }
          } else {
            const int64_t rec = (int64_t)run.slot * (P.hpc * P.group) + hq_in_cta;
            *reinterpret_cast<float4*>(&P.ws_acc[rec * D + d]) = val;
            if (d == 0) P.ws_ml[rec * 2] = M, P.ws_ml[rec * 2 + 1] = L;
          }
        }
      }
      if (!direct) publish_fin(1, item, hg_done, tq, steps_done);  // the partial is on its way to memory: its ticket is the ticket warp's
      serve_requests();                                // items this CTA has to merge
      if (trace && threadIdx.x == 0) trace[8] += (unsigned long long)(clock64() - t_fin);
      if (last) break;
      asm volatile("bar.sync 1, %0;" ::"r"(NW * 32));  // merge buffers are reused by the next run
      reset_state();
      run = nx;
      cur.hg = run.hg, cur.step = run.step0, cur.seq_len = run.seq_len;
      steps_done = 0;
      if constexpr (PTH) load_scales(cur.step * P.tps + tok_off, cur.seq_len, pg0);
    }
  }
  publish_fin(0, 0, 0, 0, 0);  // no more partials from this CTA
  while (!serve_requests()) {}  // merge what the ticket warp still hands back, until it has taken its last ticket
}

// ---- host side ------------------------------------------------------------------
// Everything the plan and the launch must agree on, from host ints only.
struct DecodeGeom {
  int D, nt, nw, tt, fp8;
  int hpc, head_groups, tps, tps_shift, bs_shift, box_slots, boxes, grid;
  int pool_sixteenths, min_chunk;
};

inline bool pow2(int x) { return x > 0 && (x & (x - 1)) == 0; }

inline int pick_hpc(int hkv, int nw, int want) {
  int hpc = 1;
  for (int c = nw; c >= 1; c >>= 1)
    if (hkv % c == 0 && (want <= 0 || c <= want)) {
      hpc = c;
      break;
    }
  return hpc;
}

}  // namespace

int mixed_decode_ctas(const b200attn_attention_params& a, int grid) {
  const Tuning& tn = tuning();
  const int forced = (tn.reserved >> 13) & 0xff;  // A/B: tuning().reserved bits 13..20 = CTAs
  if (forced > 0) return forced < grid ? forced : grid;
  // Only host ints are known: at most B - 1 single-token rows, at least T - B + 1 tokens in multi-token queries,
  // and no context lengths - take them as alike.  Rows stream Hkv * D * (|K| + |V|) bytes per context token,
  // prompt tokens cost 4 * Hq * D FLOPs per context token; with the measured peaks (6.5 TB/s, 1.2 PFLOP/s
  // for this kernel) the rows' share of the step's time is t_rows / (t_rows + t_prompts).
  const double rows = a.num_seqs > 1 ? a.num_seqs - 1 : 0, multi = a.num_tokens - a.num_seqs + 1;
  if (rows <= 0 || multi <= 0) return grid;
  const double esz = dtype_bytes(a.kv_dtype) == 1 ? 1.0 : 2.0;
  // (x 1.15: measured best on C3 - 32 CTAs where the plain ratio gives 29)
  const double ratio = 1.15 * rows / multi * (a.num_kv_heads * 2.0 * esz * 185.0) / (4.0 * a.num_q_heads);
  int k = (int)(grid * ratio / (1.0 + ratio) + 0.5);
  if (k < 8) k = 8;
  if (k * 4 > grid * 3) return grid;  // mostly rows: they get every SM, and the prompts follow
  return k;
}

namespace {

int decode_geometry(const b200attn_attention_params& a, DecodeGeom* G) {
  const Tuning& tn = tuning();
  const DeviceInfo& dev = device_info();
  const int group = a.num_q_heads / a.num_kv_heads;
  G->D = a.head_size, G->fp8 = dtype_bytes(a.kv_dtype) == 1 ? 1 : 0;  // one-byte cache elements (e4m3 or int8)
  G->nt = group > 8 ? 2 : 1;
  G->nw = a.head_size == 256 ? 4 : 8;
  // fp8: two 16-token sub-tiles per warp and step (a stage of 8 x 32 tokens is the size a bf16
  // stage of 8 x 16 has); the tile's latency chain, not HBM, bounds this variant otherwise.  (NT = 2 would spill.)
  G->tt = (G->fp8 && G->D == 128 && G->nt == 1) ? 32 : 16;
  const int esz = G->fp8 ? 1 : 2, ndc = G->D * esz / 128;
  if (tn.decode_heads_per_cta > 0) {
    G->hpc = pick_hpc(a.num_kv_heads, G->nw, tn.decode_heads_per_cta);
  } else {
    // Fewest kv heads per CTA whose TMA box is still big enough: more
    // (sequence, head group) items means more outputs written directly and
    // parallel merges, but the copy engine pays per box (measured on B200,
    // profiles/r01_decode_sweep_v2.jsonl: bf16 KV is best at 4 KB boxes, fp8 at 8 KB).
    // Beside the prompts of a mixed batch the rows run on a share of the SMs, and one SM alone pulls about
    // 50 GB/s through 4 KB boxes however idle HBM is: there the box is as big as the heads allow
    // (C3: 0.487 -> 0.440 ms from 1 to 8 kv heads per box, profiles/r02_kbench_mixed_overlap.jsonl).
    const bool beside_prompts = a.max_seqlen_q > 1 && !(tn.reserved & 4096);
    const int min_box = beside_prompts ? 32768 : (G->fp8 ? 8192 : 4096);
    int hpc = 1;
    while (hpc < G->nw && a.num_kv_heads % (hpc * 2) == 0) {
      const int tps = G->nw * G->tt / hpc;
      const int slots = a.block_size < tps ? a.block_size : tps;
      if (hpc * ndc * slots * 128 >= min_box) break;
      hpc *= 2;
    }
    G->hpc = hpc;
  }
  G->head_groups = a.num_kv_heads / G->hpc;
  G->tps = G->nw * G->tt / G->hpc;
  for (G->tps_shift = 0; (1 << G->tps_shift) < G->tps; ++G->tps_shift) {}
  for (G->bs_shift = 0; (1 << G->bs_shift) < a.block_size; ++G->bs_shift) {}
  if ((1 << G->tps_shift) != G->tps || (1 << G->bs_shift) != a.block_size) {
    set_error("decode: tokens per step (%d) and block_size (%d) must be powers of two", G->tps, a.block_size);
    return B200ATTN_E_UNSUPPORTED;
  }
  G->box_slots = a.block_size < G->tps ? a.block_size : G->tps;
  G->boxes = G->tps / G->box_slots;
  // One CTA per SM whatever the batch (CTAs the plan gives nothing exit at once): the plan is built for a
  // grid, and a grid that followed max_seqlen_k would differ between a captured launch and its replays.
  G->grid = dev.num_sms < kMaxDecodeCtas ? dev.num_sms : kMaxDecodeCtas;
  if (G->grid < 1) G->grid = 1;
  // the single-token rows of a mixed batch run beside its prompts on a share of the SMs (cabi.cu)
  if (a.max_seqlen_q > 1 && !(tn.reserved & 4096)) G->grid = mixed_decode_ctas(a, G->grid);
  // Share of every range that its owner may lose to idle CTAs (work stealing, see the producer warp), in 1/16 of
  // the range.  OFF by default: measured on B200 it levels the CTAs' finish times (end skew 10 -> 6 us on the
  // 1-kv-head C4 shard) but what the thieves pay for it - a finalize per stolen chunk at 3.6 us each, on consumers
  // that are the bottleneck of the fp8 variant - eats the gain: 96.8 vs 95.9 us there, 163.8 vs 162.0 us on C2
  // (profiles/r02_kbench_pool_v6_stealing.jsonl; the shared-pool variants before it were 10-15 % slower,
  // profiles/r02_kbench_pool_v1..v5).  Off, the cut is static, so results are bit-reproducible from launch to
  // launch.  tuning().decode_ctas_per_sm (b200attn_set_tuning's second argument) = sixteenths, 1..12, turns it on.
  G->pool_sixteenths = (tn.decode_ctas_per_sm <= 0 || tn.decode_ctas_per_sm >= 100) ? 0 : (tn.decode_ctas_per_sm > 12 ? 12 : tn.decode_ctas_per_sm);
  G->min_chunk = (tn.reserved & 1024) ? 1 : ((tn.reserved & 2048) ? 4 : 2);
  return B200ATTN_OK;
}

uint64_t decode_plan_key(const b200attn_attention_params& a, const DecodeGeom& G, bool only_q1) {
  uint64_t h = 0xcbf29ce484222325ull;
  auto mix = [&](uint64_t v) { h = (h ^ v) * 0x100000001b3ull; };
  mix((uint64_t)a.num_seqs), mix((uint64_t)G.tps_shift), mix((uint64_t)G.head_groups), mix((uint64_t)(uint32_t)a.sliding_window);
  mix(only_q1 ? 1 : 0), mix((uint64_t)G.grid), mix((uint64_t)G.pool_sixteenths), mix((uint64_t)G.min_chunk);
  return h | 1ull;  // never 0: 0 means "no plan"
}

template <int D, int KVQ, bool MMA_BF16, bool Q_BF16, int NT, int NW, int TT>
int launch_variant(const b200attn_attention_params& a, const DecodeGeom& G, DecodeParams& P, cudaStream_t stream) {
  constexpr bool KV_FP8 = KVQ != 0;  // one-byte cache elements
  constexpr int ESZ = KV_FP8 ? 1 : 2;
  constexpr int NDC = D * ESZ / 128;
  constexpr int KV_STAGE_BYTES = NW * TT * D * ESZ;
  if (G.nw != NW || G.tt != TT || G.nt != NT) {
    set_error("decode: internal geometry mismatch");
    return B200ATTN_E_UNSUPPORTED;
  }
  const DeviceInfo& dev = device_info();
  const Tuning& tn = tuning();
  // K and V of a page in ONE box when they are two equally strided views of one allocation
  // (both vLLM layouts) and a row is a single 128-byte chunk (fp8, D=128): halves the box count,
  // which is what bounds small pages (1-2 kv heads per GPU).  tuning().reserved bit 3 forbids it.
  const int64_t kv_delta = (static_cast<const char*>(a.v_cache) - static_cast<const char*>(a.k_cache));
  const bool can_pair = KV_FP8 && NDC == 1 && !(tn.reserved & 8) && !(tn.reserved & 1) && kv_delta > 0 &&
                        kv_delta % 16 == 0 && kv_delta < (1ll << 40) && a.kc_block_stride == a.vc_block_stride &&
                        a.kc_slot_stride == a.vc_slot_stride && a.kc_head_stride == a.vc_head_stride;
  P.hpc = G.hpc, P.head_groups = G.head_groups, P.tps = G.tps, P.tps_shift = G.tps_shift, P.bs_shift = G.bs_shift;
  P.box_slots = G.box_slots, P.boxes = G.boxes;
  P.tma_split = tn.reserved & 1;
  const int fixed = NW * 8 * D * 4 + NW * 8 * 2 * 4 + (2 * kMaxStages + kDescSlots + 2 * kFinSlots + 1) * 8 +
                    kDescSlots * (int)sizeof(RunDesc) + (kFinSlots + kReqSlots) * (int)sizeof(FinDesc) + 16 + kMaxStages * 4 + kMaxParts * 2;
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

  auto kern = attn_decode_kernel<D, KVQ, MMA_BF16, Q_BF16, NT, NW, TT>;
  if (int rc2 = ensure_dynamic_smem(reinterpret_cast<const void*>(kern), dev.max_smem_optin, "decode")) return rc2;
  // Programmatic dependent launch: lets this grid's launch and prologue overlap the kernel in front of it
  // (the cache write, the plan kernel or the previous decode launch - each releases its dependents early;
  // common.cuh).  tuning().reserved bit 9 turns it off.
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)G.grid), cfg.blockDim = dim3((NW + 2) * 32), cfg.dynamicSmemBytes = (size_t)smem, cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr, cfg.numAttrs = (tn.reserved & 512) ? 0 : 1;
  cudaLaunchKernelEx(&cfg, kern, P, kmap, vmap);
  note_launch("attn_decode_split_kv");
  return check_launch("attn_decode");
}

template <int D, int NT, int NW>
int dispatch_types(const b200attn_attention_params& a, const DecodeGeom& G, DecodeParams& P, cudaStream_t stream) {
  const bool qbf = a.q_dtype == B200ATTN_BF16;
  if (dtype_bytes(a.kv_dtype) == 1) {
    if constexpr (D >= 128) {
      constexpr int TT = (D == 128 && NT == 1) ? 32 : 16;  // decode_geometry()
      if (per_token_head(a)) {
        if (a.kv_dtype == B200ATTN_INT8)
          return qbf ? launch_variant<D, 3, false, true, NT, NW, TT>(a, G, P, stream)
                     : launch_variant<D, 3, false, false, NT, NW, TT>(a, G, P, stream);
        return qbf ? launch_variant<D, 2, false, true, NT, NW, TT>(a, G, P, stream)
                   : launch_variant<D, 2, false, false, NT, NW, TT>(a, G, P, stream);
      }
      return qbf ? launch_variant<D, 1, false, true, NT, NW, TT>(a, G, P, stream)
                 : launch_variant<D, 1, false, false, NT, NW, TT>(a, G, P, stream);
    } else {
      set_error("decode: fp8 KV needs head_size >= 128");
      return B200ATTN_E_UNSUPPORTED;
    }
  }
  return qbf ? launch_variant<D, 0, true, true, NT, NW, 16>(a, G, P, stream)
             : launch_variant<D, 0, false, false, NT, NW, 16>(a, G, P, stream);
}

int launch_plan_kernel(const b200attn_attention_params& a, const DecodeGeom& G, bool only_q1, void* plan, cudaStream_t stream) {
  PlanArgs A;
  A.seq_lens = a.seq_lens, A.cu_seqlens_q = a.cu_seqlens_q, A.plan = static_cast<Plan*>(plan);
  A.num_seqs = a.num_seqs, A.only_q1 = only_q1 ? 1 : 0, A.window = a.sliding_window, A.tps_shift = G.tps_shift;
  A.head_groups = G.head_groups, A.grid = G.grid, A.pool_sixteenths = G.pool_sixteenths, A.min_chunk = G.min_chunk;
  A.no_snap = (tuning().reserved & (1 << 22)) ? 1 : 0;
  decode_plan_kernel<<<1, kPlanThreads, 0, stream>>>(A);
  note_launch("decode_plan");
  return check_launch("decode_plan");
}

}  // namespace

// The ticket counters live in a region of FIXED size at the start of the workspace: they are the only bytes
// that must be zero between calls, and the same buffer serves calls of every shape (a batch grows and shrinks
// from step to step), so their region may never be another shape's partial-result region.  Its last KiB holds
// the ranges' claim words (owner / thief chunk claims) and the count of CTAs that have stopped taking chunks.
constexpr uint64_t kCounterBytes = 1ull << 20;  // one int per (sequence, head group): 262128 items
constexpr uint64_t kPoolCounterOffset = kCounterBytes - 1024;  // [1] CTAs done, [kDequeWord0 + c] range c's claim word

uint64_t decode_plan_bytes(int num_seqs) {
  return ((sizeof(Plan) + 4ull * ((uint64_t)(num_seqs > 0 ? num_seqs : 0) + 1) + 255) / 256) * 256;
}

static uint64_t decode_rows(int num_seqs, int num_q_heads, int num_kv_heads) {
  // partial slots: two per part (static part of a range, pool chunk), each a record of hpc * group head rows;
  // hpc <= 8 and hpc * group <= Hq.  Independent of the batch.
  (void)num_seqs;
  const uint64_t group = (uint64_t)(num_q_heads / (num_kv_heads > 0 ? num_kv_heads : 1));
  uint64_t rec = 8 * (group > 0 ? group : 1);
  if (rec > (uint64_t)num_q_heads) rec = (uint64_t)num_q_heads;
  return (uint64_t)kMaxSlots * rec;
}

uint64_t decode_workspace_bytes(int num_seqs, int num_q_heads, int num_kv_heads, int head_size) {
  const uint64_t rows = decode_rows(num_seqs, num_q_heads, num_kv_heads);
  return kCounterBytes + ((rows * 8 + 255) / 256) * 256 + rows * (uint64_t)head_size * 4 + decode_plan_bytes(num_seqs) + 1024;
}

bool decode_supported(const b200attn_attention_params& p, const char** why, bool mixed_batch) {
  auto no = [&](const char* w) { if (why) *why = w; return false; };
  if (p.max_seqlen_q != 1 && !mixed_batch) return no("max_seqlen_q != 1");
  if (p.q_dtype != B200ATTN_BF16 && p.q_dtype != B200ATTN_F16) return no("query dtype must be bf16/f16");
  const bool kv8 = dtype_bytes(p.kv_dtype) == 1;  // e4m3, or (with per-token-head scales) int8
  if (p.kv_dtype != p.q_dtype && !kv8) return no("kv dtype");
  if (p.head_size != 64 && p.head_size != 128 && p.head_size != 256) return no("head_size not in {64,128,256}");
  if (kv8 && p.head_size < 128) return no("fp8 / int8 KV needs head_size >= 128");
  if (per_token_head(p)) {  // the consumers index a block's scales in 32 bits
    // the consumers hold the scale caches' block and slot strides in bytes as 32-bit values
    auto small = [&](int64_t s) { return s >= 0 && s * 4 < (int64_t(1) << 31); };
    if (!small(p.ks_slot_stride * p.block_size) || !small(p.vs_slot_stride * p.block_size) || !small(p.ks_block_stride) ||
        !small(p.vs_block_stride))
      return no("scale cache strides >= 2 GiB");
  }
  if (p.num_kv_heads <= 0 || p.num_q_heads % p.num_kv_heads) return no("heads");
  const int g = p.num_q_heads / p.num_kv_heads;
  if (g > 16 || (g > 8 && p.head_size == 256)) return no("GQA group too large for the decode kernel");
  if (!pow2(p.block_size) || p.block_size < 16 || p.block_size > 256) return no("block_size not a power of two in [16,256]");
  const int esz = kv8 ? 1 : 2;
  auto m16 = [&](int64_t s) { return (s * esz) % 16 == 0; };
  if (!m16(p.kc_block_stride) || !m16(p.kc_slot_stride) || !m16(p.kc_head_stride) || !m16(p.vc_block_stride) ||
      !m16(p.vc_slot_stride) || !m16(p.vc_head_stride))
    return no("cache strides not 16-byte multiples");
  if ((reinterpret_cast<uintptr_t>(p.k_cache) & 15) || (reinterpret_cast<uintptr_t>(p.v_cache) & 15)) return no("cache base alignment");
  if ((reinterpret_cast<uintptr_t>(p.q) & 7) || (p.q_stride0 % 4) || (p.q_stride1 % 4)) return no("query alignment");
  return true;
}

// b200attn_decode_plan: the plan of the decode launches that will follow with these seq_lens / cu_seqlens_q
// (one per layer).  *key_out = 0 when the shape does not run the split-KV kernel (nothing to plan).
int build_decode_plan(const b200attn_attention_params& a, void* plan, uint64_t plan_bytes, uint64_t* key_out,
                      cudaStream_t stream) {
  if (key_out) *key_out = 0;
  const bool only_q1 = a.max_seqlen_q > 1;
  const char* why = nullptr;
  if (!decode_supported(a, &why, only_q1)) return B200ATTN_OK;
  if (!plan || plan_bytes < decode_plan_bytes(a.num_seqs)) {
    set_error("decode_plan: plan buffer too small (%llu < %llu bytes)", (unsigned long long)plan_bytes,
              (unsigned long long)decode_plan_bytes(a.num_seqs));
    return B200ATTN_E_WORKSPACE;
  }
  if (reinterpret_cast<uintptr_t>(plan) & 15) {
    set_error("decode_plan: plan buffer must be 16-byte aligned");
    return B200ATTN_E_ALIGNMENT;
  }
  DecodeGeom G;
  if (int rc = decode_geometry(a, &G)) return rc;
  if (int rc = launch_plan_kernel(a, G, only_q1, plan, stream)) return rc;
  if (key_out) *key_out = decode_plan_key(a, G, only_q1);
  return B200ATTN_OK;
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
  DecodeGeom G;
  if (int rc = decode_geometry(a, &G)) return rc;
  DecodeParams P;
  P.a = a;
  P.group = a.num_q_heads / a.num_kv_heads;
  P.only_q1 = only_single_token_seqs ? 1 : 0;
  P.out_vec = (!(reinterpret_cast<uintptr_t>(a.out) & 7) && a.out_stride0 % 4 == 0 && a.out_stride1 % 4 == 0) ? 1 : 0;
  P.trace = decode_trace_buffer();
  // carve the workspace: ticket counters first (fixed region, zero between calls), then ml, acc and the
  // launcher's own plan
  const uint64_t items_max = (uint64_t)a.num_seqs * a.num_kv_heads;
  if (items_max * sizeof(int) > kPoolCounterOffset) {
    set_error("decode: %llu (sequence, kv head) items exceed the %llu ticket counters of the workspace",
              (unsigned long long)items_max, (unsigned long long)(kPoolCounterOffset / sizeof(int)));
    return B200ATTN_E_WORKSPACE;
  }
  char* w = static_cast<char*>(a.workspace);
  P.ws_cnt = reinterpret_cast<int*>(w);
  P.ws_pool = reinterpret_cast<int*>(w + kPoolCounterOffset);
  uint64_t off = kCounterBytes;
  const uint64_t rows = decode_rows(a.num_seqs, a.num_q_heads, a.num_kv_heads);
  P.ws_ml = reinterpret_cast<float*>(w + off);
  off += ((rows * 8 + 255) / 256) * 256;
  P.ws_acc = reinterpret_cast<float*>(w + off);
  off += rows * (uint64_t)a.head_size * 4;
  off = ((off + 255) / 256) * 256;
  if (off + decode_plan_bytes(a.num_seqs) > a.workspace_bytes) {
    set_error("decode: workspace carve-out exceeds the buffer");
    return B200ATTN_E_WORKSPACE;
  }
  const uint64_t key = decode_plan_key(a, G, P.only_q1 != 0);
  if (a.decode_plan && a.decode_plan_key == key) {
    P.plan = static_cast<const Plan*>(a.decode_plan);  // built once for the step (b200attn_decode_plan)
  } else {
    // no plan for this shape from the caller: build one in the workspace in front of the launch
    if (int rc = launch_plan_kernel(a, G, P.only_q1 != 0, w + off, stream)) return rc;
    P.plan = reinterpret_cast<const Plan*>(w + off);
  }
  switch (a.head_size) {
    case 64: return G.nt == 1 ? dispatch_types<64, 1, 8>(a, G, P, stream) : dispatch_types<64, 2, 8>(a, G, P, stream);
    case 128: return G.nt == 1 ? dispatch_types<128, 1, 8>(a, G, P, stream) : dispatch_types<128, 2, 8>(a, G, P, stream);
    case 256: return dispatch_types<256, 1, 4>(a, G, P, stream);
  }
  return B200ATTN_E_UNSUPPORTED;
}

}  // namespace b200attn
