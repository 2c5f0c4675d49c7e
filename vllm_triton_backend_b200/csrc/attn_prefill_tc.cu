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
#include "common.cuh"

namespace b200attn {
namespace {

constexpr int kD = 128;
constexpr int kBN = 128;                        // kv tokens per tile
constexpr int kKStages = 3;                     // K is needed right after PV of the previous tile: prefetch deeper
constexpr int kVStages = 2;
constexpr int kSubtileBytes = 128 * 128;        // [128 rows][128 B]: one 64-element chunk of 128 rows
constexpr int kTileBytes = 2 * kSubtileBytes;   // 128 rows x 128 dims x 2 B
constexpr int kThreads = 384;
// TMEM columns: S0 | S1 | O0 | O1 (subtile w: S at 128 w, O at 256 + 128 w)
__host__ __device__ constexpr uint32_t col_s(int w) { return 128u * (uint32_t)w; }
__host__ __device__ constexpr uint32_t col_o(int w) { return 256u + 128u * (uint32_t)w; }
constexpr int kSmemQ = 0;
constexpr int kSmemK = 2 * kTileBytes;
constexpr int kSmemV = kSmemK + kKStages * kTileBytes;
constexpr int kSmemBar = kSmemV + kVStages * kTileBytes;
constexpr int kSmemBytes = kSmemBar + 256 + 1024;  // barriers + tmem ptr + alignment slack
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
    if (blockIdx.x == 1 && blockIdx.y == 0 && (j) < 64) g_tc_trace[(j) * 32 + (ev)] = clock64(); \
  } while (0)
#define TC_MARK(k)                                                                \
  do {                                                                             \
    if (blockIdx.x == 1 && blockIdx.y == 0) g_tc_trace[64 * 32 + (k)] = clock64(); \
  } while (0)
#else
#define TC_MARK(k) do { } while (0)
#define TC_TRACE(j, ev) do { } while (0)
#endif

struct TcParams {
  b200attn_attention_params a;
  int group;      // query heads per kv head (power of two, <= 128)
  int block_q;    // tokens per 128-row subtile
  int min_q_len, max_q_len;
  int box_slots;  // tokens per TMA box = min(block_size, 128)
};

// P_FULL: first half of P (keys 0..63 of the tile) stored; P_FULLB: second half
enum Bar { Q_FULL = 0, K_FULL = 1, K_EMPTY = 4, V_FULL = 7, V_EMPTY = 9, S_FULL = 11, P_FULL = 13, P_FULLB = 15, O_DONE = 17, V_FIX = 19, NUM_BARS = 20 };

// ---- tcgen05 wrappers -----------------------------------------------------------------
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tmem_alloc(uint32_t* dst_smem, uint32_t cols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(dst_smem)), "r"(cols) : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t cols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(cols) : "memory");
}
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
// D[tmem] (+)= A[smem] * B[smem]
__device__ __forceinline__ void umma_ss(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// D[tmem] (+)= A[tmem] * B[smem]
__device__ __forceinline__ void umma_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}"
      ::"r"(d_tmem), "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
        "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
        "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_st32(uint32_t taddr, const uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
      "{%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31,%32};"
      ::"r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]),
      "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]), "r"(r[16]), "r"(r[17]),
      "r"(r[18]), "r"(r[19]), "r"(r[20]), "r"(r[21]), "r"(r[22]), "r"(r[23]), "r"(r[24]), "r"(r[25]), "r"(r[26]),
      "r"(r[27]), "r"(r[28]), "r"(r[29]), "r"(r[30]), "r"(r[31])
      : "memory");
}
__device__ __forceinline__ void tmem_ld8(uint32_t taddr, uint32_t (&r)[8]) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
               : "r"(taddr)
               : "memory");
}
__device__ __forceinline__ void tmem_st8(uint32_t taddr, const uint32_t (&r)[8]) {
  asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};"
               ::"r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7])
               : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

__device__ __forceinline__ void tma_load_4d(void* smem_dst, const CUtensorMap* map, uint64_t* bar, int c0, int c1, int c2, int c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];"
      ::"r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
      : "memory");
}
__device__ __forceinline__ void tma_load_5d_nohint(void* smem_dst, const CUtensorMap* map, uint64_t* bar, int c0, int c1, int c2,
                                                   int c3, int c4) {
  asm volatile(
      "cp.async.bulk.tensor.5d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6, %7}], [%2];"
      ::"r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3),
      "r"(c4)
      : "memory");
}

// UMMA shared-memory descriptor, 128-byte swizzle (layout type 2), descriptor version 1.
//   K-major operand  ([rows][64 elems] sub-tiles): SBO = 1024 B between 8-row groups, LBO unused (1).
//   MN-major operand ([k rows][64 elems] sub-tiles): LBO = bytes between 64-element chunks along MN,
//                                                    SBO = 1024 B between 8-row groups along K.
__device__ __forceinline__ uint64_t umma_desc(uint32_t smem_addr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr & 0x3FFFFu) >> 4);
  d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFFu) << 16;
  d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFFu) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}
// instruction descriptor, kind::f16: fp32 accumulate, A/B format, majors, N, M
__host__ __device__ constexpr uint32_t umma_idesc(bool bf16, bool b_mn_major, int M, int N) {
  return (1u << 4) | ((bf16 ? 1u : 0u) << 7) | ((bf16 ? 1u : 0u) << 10) | (0u << 15) | ((b_mn_major ? 1u : 0u) << 16) |
         ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(pred));
  return pred != 0;
}
__device__ __forceinline__ float ex2_approx(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}


// ---- packed fp32 pairs (FFMA2 / FADD2 on sm_100) and 3-input max (FMNMX3) ---------------
__device__ __forceinline__ uint64_t f2_pack(float lo, float hi) {
  uint64_t r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
  return r;
}
__device__ __forceinline__ void f2_unpack(uint64_t v, float& lo, float& hi) {
  asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ uint64_t f2_fma(uint64_t a, uint64_t b, uint64_t c) {
  uint64_t r;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
  return r;
}
__device__ __forceinline__ uint64_t f2_add(uint64_t a, uint64_t b) {
  uint64_t r;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}
__device__ __forceinline__ uint64_t f2_sub(uint64_t a, uint64_t b) {
  uint64_t r;
  asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}
__device__ __forceinline__ uint64_t f2_add_rm(uint64_t a, uint64_t b) {  // round toward -inf
  uint64_t r;
  asm("add.rm.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}
__device__ __forceinline__ float fmax3(float a, float b, float c) {
  float r;
  asm("max.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(a), "f"(b), "f"(c));
  return r;
}
// 2^x for a pair without the MUFU: x = floor(x) + f, 2^f by a degree-3 polynomial on the
// FMA pipe (max rel. error 9e-5, far below P's bf16/f16 rounding), floor(x) added into the
// exponent field.  The MUFU does 16 ex2/clk/SM, which for a 128x128 tile is exactly the time
// of the two MMAs; moving a fraction of the exponentials here takes it off the critical path.
__device__ __forceinline__ uint64_t ex2_poly2(float x0, float x1) {
  const float kMagic = 12582912.f;  // 2^23 + 2^22: ulp 1, so add.rm leaves floor(x) in the low mantissa bits
  const uint64_t x = f2_pack(fmaxf(x0, -127.f), fmaxf(x1, -127.f));
  const uint64_t xr = f2_add_rm(x, f2_pack(kMagic, kMagic));
  const uint64_t fl = f2_sub(xr, f2_pack(kMagic, kMagic));
  const uint64_t f = f2_sub(x, fl);
  uint64_t r = f2_fma(f, f2_pack(0.077119089663028717f, 0.077119089663028717f), f2_pack(0.227564394474029541f, 0.227564394474029541f));
  r = f2_fma(r, f, f2_pack(0.695146143436431885f, 0.695146143436431885f));
  r = f2_fma(r, f, f2_pack(1.f, 1.f));
  float r0, r1, i0, i1;
  f2_unpack(r, r0, r1);
  f2_unpack(xr, i0, i1);
  const float p0 = __int_as_float(__float_as_int(r0) + (__float_as_int(i0) << 23));
  const float p1 = __int_as_float(__float_as_int(r1) + (__float_as_int(i1) << 23));
  return f2_pack(p0, p1);
}

// kEmuNum of every kEmuDen exponential pairs are computed by ex2_poly2, the rest on the MUFU.
// kMode bit 0: exponentials of the first half start against the previous reference max while
//              the tile max is still being found (redo if it jumps);
//       bit 1: the two softmax warpgroups take turns on the MUFU (named-barrier ping-pong), so
//              one subtile's exponentials run under the other subtile's MMAs instead of both
//              drifting into the same phase and halving each other's rate.
template <bool kBf16, int kEmuNum, int kEmuDen, int kMode>
__global__ void __launch_bounds__(kThreads, 1)
attn_prefill_tc_kernel(const TcParams P, const __grid_constant__ CUtensorMap q_map, const __grid_constant__ CUtensorMap k_map,
                       const __grid_constant__ CUtensorMap v_map) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sQ = smem + kSmemQ;
  uint8_t* sK = smem + kSmemK;
  uint8_t* sV = smem + kSmemV;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kSmemBar);
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(bars + NUM_BARS);

  const b200attn_attention_params& a = P.a;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int g = P.group;
  const int BQ = 2 * P.block_q;  // query tokens per CTA
#ifdef B200ATTN_TC_TRACE
  const long long tc_t0 = tc_globaltimer();
  if (threadIdx.x == 0) TC_MARK(0);
#endif

  // ---- q-tile -> sequence; heaviest (latest) tiles are scheduled first ---------------
  const int q_tile_global = (int)gridDim.x - 1 - (int)blockIdx.x;
  const int kvh = blockIdx.y;
  int left = 0, right = a.num_seqs;
  while (left < right) {
    const int mid = (left + right) >> 1;
    const int mid_val = a.cu_seqlens_q[mid] / BQ + mid;
    if (mid_val <= q_tile_global) left = mid + 1; else right = mid;
  }
  const int b = left - 1;
  if (b < 0) return;
  const int q_start = a.cu_seqlens_q[b];
  const int q_len = a.cu_seqlens_q[b + 1] - q_start;
  const int q_tile_local = q_tile_global - (q_start / BQ + b);
  if (q_tile_local * BQ >= q_len) return;
  if (q_len < P.min_q_len || q_len > P.max_q_len) return;  // another launch covers this sequence
  const int seq_len = a.seq_lens[b];
  const int ctx = seq_len - q_len;
  const int qpos0 = q_tile_local * BQ;
  const int qpos_last = min(qpos0 + BQ, q_len) - 1;
  const int kv_limit = min(ctx + qpos_last + 1, seq_len);  // keys [kv_first, kv_limit) can be visible
  const int window = a.sliding_window;
  const int kv_first = window > 0 ? max(0, ctx + qpos0 - window + 1) : 0;
  const int tile0 = kv_first / kBN;
  const int n_tiles = (kv_limit + kBN - 1) / kBN - tile0;

  if (n_tiles <= 0) {
    // nothing visible (seq_len <= 0): rows get zeros like the reference's reduce (:828)
    if (threadIdx.x < 256) {
      const int r = threadIdx.x, qpos = qpos0 + r / g;
      if (qpos < q_len) {
        int4* orow = reinterpret_cast<int4*>(static_cast<char*>(a.out) +
                                             ((int64_t)(q_start + qpos) * a.out_stride0 + (int64_t)(kvh * g + r % g) * a.out_stride1) * 2);
#pragma unroll
        for (int i = 0; i < kD * 2 / 16; ++i) orow[i] = make_int4(0, 0, 0, 0);
      }
    }
    return;
  }

  // ---- one-time setup ------------------------------------------------------------------
  if (warp == 8 && lane == 0) {
    mbar_init(&bars[Q_FULL], 1);
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
    }
    mbar_init(&bars[V_FIX], 1);
    fence_barrier_init();
  }
  if (warp == 11) tmem_alloc(tmem_ptr_smem, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;
  if (threadIdx.x == 0) TC_MARK(1);

  const int last_tile = n_tiles - 1;
  const int last_valid = kv_limit - (tile0 + last_tile) * kBN;  // rows of the last tile that hold visible keys
  const bool need_fix = last_valid < kBN;

  // Register budget follows the roles: the softmax warpgroups hold a 128-column row of S
  // each, the MMA / TMA warpgroup needs almost nothing (2 x 208 + 80 <= 512 per lane slot).
  if (warp < 8) asm volatile("setmaxnreg.inc.sync.aligned.u32 208;");
  else asm volatile("setmaxnreg.dec.sync.aligned.u32 80;");

  if (warp < 8) {
    // ============================ softmax warpgroups ====================================
    const int w = warp >> 2;                  // Q subtile
    const int row = (warp & 3) * 32 + lane;   // row of the subtile == TMEM lane
    const int qpos = qpos0 + (w * 128 + row) / g;
    const int head = row % g;                 // 128 % g == 0
    const bool row_valid = qpos < q_len;
    const int limit = row_valid ? ctx + qpos + 1 : 0;                        // visible keys: [lo, limit)
    const int lo = (window > 0 && row_valid) ? max(0, ctx + qpos - window + 1) : 0;
    const uint32_t lane_base = (uint32_t)((warp & 3) * 32) << 16;
    const uint32_t s_tmem = tmem_base + lane_base + col_s(w);
    const uint32_t o_tmem = tmem_base + lane_base + col_o(w);
    const float scale_log2 = a.softmax_scale * kLog2e;
    float m_used = -INFINITY, l_sum = 0.f;
    // ping-pong: warpgroup 0 goes first (warpgroup 1 hands it the first turn)
    if constexpr ((kMode & 2) != 0) {
      if (w == 1) asm volatile("bar.arrive 2, 256;" ::: "memory");
    }

    for (int j = 0; j < n_tiles; ++j) {
      mbar_wait(&bars[S_FULL + w], (uint32_t)(j & 1));
      tc_fence_after();
      if (row == 0) TC_TRACE(j, w * 5 + 0);
      if (row == 0 && w == 0 && j == 0) TC_MARK(2);
      if (row == 0 && w == 0 && j == 1) TC_MARK(3);
      if (row == 0 && w == 0 && j == last_tile) TC_MARK(4);
      uint32_t s[4][32];
#pragma unroll
      for (int c = 0; c < 4; ++c) tmem_ld32(s_tmem + c * 32, s[c]);
      tmem_ld_wait();
      if (row == 0) TC_TRACE(j, w * 5 + 1);
      const int n0 = (tile0 + j) * kBN;
      if (n0 + kBN > limit || n0 < lo) {
#pragma unroll
        for (int c = 0; c < 4; ++c)
#pragma unroll
          for (int i = 0; i < 32; ++i) {
            const int pos = n0 + c * 32 + i;
            if (pos >= limit || pos < lo) s[c][i] = 0xff800000u;  // -inf
          }
      }
      // P = 2^(S*scale - m) in two halves of 64 keys; each half replaces S in place (bf16
      // pairs: columns [32h, 32h+32) of the S region) and is announced on its own barrier,
      // so the first four k-steps of PV run while the second half is still being computed.
      const uint64_t sc2 = f2_pack(scale_log2, scale_log2);
      uint32_t pk[32];
      auto exp_half = [&](int h, float m_eff, uint64_t& sum_out) {
        const uint64_t nm2 = f2_pack(-m_eff, -m_eff);
        uint64_t sum2[2] = {0ull, 0ull};
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
        sum_out = f2_add(sum2[0], sum2[1]);
      };
      // The row max is NOT a prerequisite of the exponentials: the reference point m_used only
      // has to be within 2^kRescaleThreshold of the true running max, so the first half is
      // computed against the m_used of the previous tiles while the max of this tile (FMNMX3, on
      // the otherwise idle ALU pipe) is found under the MUFU's shadow.  Only if the tile max
      // jumps past the threshold (always on a row's first visible tile) is the half redone.
      constexpr bool kOptMax = (kMode & 1) != 0, kPingPong = (kMode & 2) != 0;
      auto turn_wait = [&]() {  // my turn on the MUFU: the other warpgroup has released barrier 2 + w
        if constexpr (kPingPong) asm volatile("bar.sync %0, 256;" ::"r"(2 + w) : "memory");
      };
      uint64_t sum_a, sum_b;
      if constexpr (kOptMax) {
        turn_wait();
        exp_half(0, (m_used == -INFINITY) ? 0.f : m_used, sum_a);
      }
      float mxc[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
#pragma unroll
      for (int c = 0; c < 4; ++c)
#pragma unroll
        for (int i = 0; i < 32; i += 2) mxc[c] = fmax3(mxc[c], __uint_as_float(s[c][i]), __uint_as_float(s[c][i + 1]));
      const float mx = fmax3(fmaxf(mxc[0], mxc[1]), mxc[2], mxc[3]);
      const float m_tile = mx * scale_log2;  // softmax_scale > 0 (checked on the host)
      float alpha = 1.f;
      if (row == 0) TC_TRACE(j, w * 5 + 2);
      const bool grew = __any_sync(0xffffffffu, m_tile > m_used + kRescaleThreshold);
      if (grew) {
        if (m_tile > m_used + kRescaleThreshold) {  // also the first finite max (m_used = -inf)
          alpha = ex2_approx(m_used - m_tile);       // 0 when m_used = -inf
          m_used = m_tile;
        }
        if (j > 0) {
          // rescale the running output before PV(j) accumulates into it
          mbar_wait(&bars[O_DONE + w], (uint32_t)((j - 1) & 1));
          tc_fence_after();
          // (8 columns at a time: all 128 S values are live here)
#pragma unroll 1
          for (int c = 0; c < 16; ++c) {
            uint32_t t[8];
            tmem_ld8(o_tmem + c * 8, t);
            tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 8; ++i) t[i] = __float_as_uint(__uint_as_float(t[i]) * alpha);
            tmem_st8(o_tmem + c * 8, t);
          }
        }
      }
      if constexpr (!kOptMax) turn_wait();
      if (!kOptMax || grew) exp_half(0, (m_used == -INFINITY) ? 0.f : m_used, sum_a);
      const float m_eff = (m_used == -INFINITY) ? 0.f : m_used;
      tmem_st32(s_tmem, pk);
      tmem_st_wait();
      tc_fence_before();
      mbar_arrive(&bars[P_FULL + w]);
      if (row == 0) TC_TRACE(j, w * 5 + 3);
      exp_half(1, m_eff, sum_b);
      tmem_st32(s_tmem + 32, pk);
      tmem_st_wait();
      tc_fence_before();
      mbar_arrive(&bars[P_FULLB + w]);
      if constexpr (kPingPong) asm volatile("bar.arrive %0, 256;" ::"r"(2 + (w ^ 1)) : "memory");
      if (row == 0) TC_TRACE(j, w * 5 + 4);
      float sa, sb;
      f2_unpack(f2_add(sum_a, sum_b), sa, sb);
      l_sum = l_sum * alpha + (sa + sb);
    }

    if constexpr ((kMode & 2) != 0) {
      if (w == 0) asm volatile("bar.sync 2, 256;" ::: "memory");  // absorb warpgroup 1's last hand-over
    }
    // ---- epilogue: O / L -> global -----------------------------------------------------
    if (row == 0 && w == 1) TC_MARK(5);
    mbar_wait(&bars[O_DONE + w], (uint32_t)(last_tile & 1));
    tc_fence_after();
    if (row == 0 && w == 1) TC_MARK(6);
    const float inv = (l_sum > 0.f) ? 1.f / l_sum : 0.f;
    char* orow = static_cast<char*>(a.out) +
                 ((int64_t)(q_start + (row_valid ? qpos : 0)) * a.out_stride0 + (int64_t)(kvh * g + head) * a.out_stride1) * 2;
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      uint32_t t[32];
      tmem_ld32(o_tmem + c * 32, t);
      tmem_ld_wait();
      if (row_valid) {
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
  } else if (warp == 8) {
    // ============================ MMA issue warp =========================================
    // The whole warp walks the (warp-uniform) control flow and one elected lane issues, so
    // descriptors live in uniform registers; every descriptor is a precomputed base plus a
    // compile-time offset in the 14-bit address field.  (A `lane == 0` version cost ~130
    // cycles per MMA in scalar address arithmetic and R2UR loops - twice the MMA itself.)
    constexpr uint32_t idesc_qk = umma_idesc(kBf16, false, 128, 128);
    constexpr uint32_t idesc_pv = umma_idesc(kBf16, true, 128, 128);
    const uint32_t tm = __shfl_sync(0xffffffffu, tmem_base, 0);
    const uint64_t q_desc0 = umma_desc(__shfl_sync(0xffffffffu, smem_u32(sQ), 0), 16, 1024);
    const uint64_t k_desc0 = umma_desc(__shfl_sync(0xffffffffu, smem_u32(sK), 0), 16, 1024);
    const uint64_t v_desc0 = umma_desc(__shfl_sync(0xffffffffu, smem_u32(sV), 0), kSubtileBytes, 1024);
    const bool leader = elect_one();
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
    // PV for subtile w, tile j: k-steps 0..3 as soon as the first half of P is in TMEM, 4..7 after the second
    auto issue_pv = [&](int w, int st, bool acc, int j) {
      const uint64_t bd = v_desc0 + (uint64_t)((st * kTileBytes) >> 4);
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        mbar_wait(&bars[(h ? P_FULLB : P_FULL) + w], (uint32_t)(j & 1));
        tc_fence_after();
        if (leader) {
          TC_TRACE(j, 10 + w * 3 + h);
#pragma unroll
          for (int kk = 4 * h; kk < 4 * h + 4; ++kk)
            umma_ts(tm + col_o(w), tm + col_s(w) + kk * 8, bd + (uint64_t)((kk * 2048) >> 4), idesc_pv, (acc || kk > 0) ? 1u : 0u);
        }
      }
    };
    mbar_wait(&bars[Q_FULL], 0);
    mbar_wait(&bars[K_FULL + 0], 0);
    tc_fence_after();
    issue_qk(0, 0);
    if (leader) umma_commit(&bars[S_FULL + 0]);
    issue_qk(1, 0);
    if (leader) {
      umma_commit(&bars[S_FULL + 1]);
      umma_commit(&bars[K_EMPTY + 0]);
    }
    int ks1 = 1 % kKStages;      // K stage / phase of tile j + 1
    uint32_t kph1 = 0;
    for (int j = 0; j < n_tiles; ++j) {
      const int st = j & 1;
      const uint32_t ph = (uint32_t)((j >> 1) & 1);
      const bool more = j + 1 < n_tiles;
      // operand tiles first: they arrive long before P does, so these waits are off the
      // critical path P(j) -> PV(j), QK(j+1) -> S(j+1)
      mbar_wait(&bars[V_FULL + st], ph);
      if (j == last_tile && need_fix) mbar_wait(&bars[V_FIX], 0);
      if (more) mbar_wait(&bars[K_FULL + ks1], kph1);
      if (leader) TC_TRACE(j, 18);
      issue_pv(0, st, j > 0, j);
      if (leader) umma_commit(&bars[O_DONE + 0]);
      if (more) {
        issue_qk(0, ks1);
        if (leader) {
          umma_commit(&bars[S_FULL + 0]);
          TC_TRACE(j, 12);
        }
      }
      issue_pv(1, st, j > 0, j);
      if (leader) {
        umma_commit(&bars[O_DONE + 1]);
        umma_commit(&bars[V_EMPTY + st]);
      }
      if (more) {
        issue_qk(1, ks1);
        if (leader) {
          umma_commit(&bars[S_FULL + 1]);
          umma_commit(&bars[K_EMPTY + ks1]);
          TC_TRACE(j, 15);
        }
      }
      if (++ks1 == kKStages) ks1 = 0, kph1 ^= 1;
    }
    __syncwarp();
  } else if (warp == 9 || warp == 10) {
    // ============================ TMA warps: 9 = Q and K, 10 = V =========================
    // Separate warps so a full V ring never holds back the K prefetch (K of tile j+1 is on the
    // critical path right after PV(j); a K tile takes ~2000 cycles from request to arrival).
    const bool is_k = warp == 9;
    if (lane == 0) {
      if (is_k) {
        tma_prefetch_desc(&q_map);
        tma_prefetch_desc(&k_map);
        mbar_arrive_expect_tx(&bars[Q_FULL], 2 * kTileBytes);
      } else {
        tma_prefetch_desc(&v_map);
      }
    }
    __syncwarp();
    if (is_k && lane < 4) {
      const int w = lane >> 1, dc = lane & 1;
      tma_load_4d(sQ + w * kTileBytes + dc * kSubtileBytes, &q_map, &bars[Q_FULL], 0, dc, kvh * g, q_start + qpos0 + w * P.block_q);
    }
    const int bs = a.block_size;
    const int boxes_per_tile = kBN / P.box_slots;
    const uint32_t box_bytes = (uint32_t)P.box_slots * 128;
    const int32_t* bt = a.block_tables + (int64_t)b * a.bt_stride0;
    const int my_box = lane >> 1, my_dc = lane & 1;
    // page id of this lane's box, fetched one tile ahead of its use
    auto lookup = [&](int j) {
      const int tok = (tile0 + j) * kBN + my_box * P.box_slots;
      return (j < n_tiles && my_box < boxes_per_tile && tok < kv_limit) ? bt[tok / bs] : 0;
    };
    const int nst = is_k ? kKStages : kVStages;
    const int full0 = is_k ? K_FULL : V_FULL, empty0 = is_k ? K_EMPTY : V_EMPTY;
    uint8_t* ring = is_k ? sK : sV;
    const CUtensorMap* map = is_k ? &k_map : &v_map;
    int page_next = lookup(0);
    int st = 0;
    uint32_t ph = 0;
    for (int j = 0; j < n_tiles; ++j) {
      const int n0 = (tile0 + j) * kBN;
      const int nbox = min(boxes_per_tile, (kv_limit - n0 + P.box_slots - 1) / P.box_slots);
      const bool active = my_box < nbox;
      const int page = page_next;
      page_next = lookup(j + 1);
      const int slot0 = (n0 + my_box * P.box_slots) % bs;
      const uint32_t dst_off = (uint32_t)(st * kTileBytes + my_dc * kSubtileBytes) + (uint32_t)my_box * box_bytes;
      mbar_wait(&bars[empty0 + st], ph ^ 1);
      if (lane == 0) TC_TRACE(j, is_k ? 19 : 20);
      if (lane == 0) mbar_arrive_expect_tx(&bars[full0 + st], (uint32_t)nbox * 2 * box_bytes);
      __syncwarp();
      if (active) tma_load_5d_nohint(ring + dst_off, map, &bars[full0 + st], 0, slot0, my_dc, kvh, page);
      if (++st == nst) st = 0, ph ^= 1;
    }
    // ---- V tail fix-up -------------------------------------------------------------------
    // Rows of the last tile beyond the visible keys are either unwritten cache slots or
    // shared memory no copy touched; P is 0 there, but 0 x NaN would poison O.  The last
    // tile's copy was issued just above, so its V_FULL phase is the one in flight.
    if (!is_k && need_fix) {
      const int vst = last_tile % kVStages;
      mbar_wait(&bars[V_FULL + vst], (uint32_t)((last_tile / kVStages) & 1));
      uint8_t* vt = sV + vst * kTileBytes;
      const int n16 = (kBN - last_valid) * 16;  // 16-byte pieces: 2 chunks x 8 per row
      for (int idx = lane; idx < n16; idx += 32) {
        const int r = last_valid + (idx >> 4), dc = (idx >> 3) & 1, c = idx & 7;
        *reinterpret_cast<int4*>(vt + dc * kSubtileBytes + r * 128 + c * 16) = make_int4(0, 0, 0, 0);
      }
      fence_proxy_async();
      __syncwarp();
      if (lane == 0) mbar_arrive(&bars[V_FIX]);
    }
  }

  // ---- teardown ---------------------------------------------------------------------------
  tc_fence_before();
  __syncthreads();
#ifdef B200ATTN_TC_TRACE
  if (threadIdx.x == 0) TC_MARK(7);
  if (threadIdx.x == 0) {
    const int cta = blockIdx.y * gridDim.x + blockIdx.x;
    if (cta < 4096) {
      unsigned smid;
      asm volatile("mov.u32 %0, %smid;" : "=r"(smid));
      g_tc_cta[cta * 4 + 0] = smid, g_tc_cta[cta * 4 + 1] = tc_t0, g_tc_cta[cta * 4 + 2] = tc_globaltimer(), g_tc_cta[cta * 4 + 3] = n_tiles;
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
  if (device_info().cc_major < 10) return no("needs sm_100a");
  if (p.head_size != kD) return no("head_size != 128");
  if (p.q_dtype != B200ATTN_BF16 && p.q_dtype != B200ATTN_F16) return no("query dtype must be bf16/f16");
  if (p.kv_dtype != p.q_dtype) return no("kv dtype must equal the query dtype");
  if (p.num_kv_heads <= 0 || p.num_q_heads % p.num_kv_heads) return no("heads");
  const int g = p.num_q_heads / p.num_kv_heads;
  if (!pow2(g) || g > 128) return no("GQA group not a power of two <= 128");
  if (!pow2(p.block_size) || p.block_size < 16) return no("block_size not a power of two >= 16");
  if (p.softcap != 0.f || p.alibi_slopes || p.sinks) return no("softcap / ALiBi / sinks");
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

int launch_attn_prefill_tc(const b200attn_attention_params& a, cudaStream_t stream, QLenRange range) {
  const char* why = nullptr;
  if (!prefill_tc_supported(a, &why)) {
    set_error("prefill_tc: unsupported configuration (%s)", why);
    return B200ATTN_E_UNSUPPORTED;
  }
  TcParams P;
  P.a = a;
  P.group = a.num_q_heads / a.num_kv_heads;
  P.block_q = 128 / P.group;
  P.min_q_len = range.min_q_len, P.max_q_len = range.max_q_len;
  P.box_slots = a.block_size < kBN ? a.block_size : kBN;
  CUtensorMap qmap, kmap, vmap;
  int rc = make_thd_map(&qmap, a.q, kD, a.num_q_heads, a.num_tokens, a.q_stride0, a.q_stride1, P.group, P.block_q);
  if (rc) return rc;
  rc = make_kv_map(&kmap, a.k_cache, 2, kD, a.block_size, a.num_kv_heads, a.num_blocks, a.kc_block_stride, a.kc_slot_stride,
                   a.kc_head_stride, P.box_slots, 1, 1);
  if (rc) return rc;
  rc = make_kv_map(&vmap, a.v_cache, 2, kD, a.block_size, a.num_kv_heads, a.num_blocks, a.vc_block_stride, a.vc_slot_stride,
                   a.vc_head_stride, P.box_slots, 1, 1);
  if (rc) return rc;
  const bool bf16 = a.q_dtype == B200ATTN_BF16;
  // fraction of the exponentials taken off the MUFU (tuning().reserved bits 4..7; 0 = default)
  using Kern = void (*)(const TcParams, const CUtensorMap, const CUtensorMap, const CUtensorMap);
#define TCV(B) {attn_prefill_tc_kernel<B, 0, 1, 0>, attn_prefill_tc_kernel<B, 0, 1, 0>, attn_prefill_tc_kernel<B, 0, 1, 2>, \
                attn_prefill_tc_kernel<B, 0, 1, 3>, attn_prefill_tc_kernel<B, 1, 4, 2>, attn_prefill_tc_kernel<B, 1, 4, 3>, \
                attn_prefill_tc_kernel<B, 1, 3, 2>, attn_prefill_tc_kernel<B, 1, 2, 2>, attn_prefill_tc_kernel<B, 1, 3, 3>}
  constexpr int kNumVariants = 9;
  static const Kern kVariants[2][kNumVariants] = {TCV(false), TCV(true)};
#undef TCV
  int variant = (tuning().reserved >> 4) & 15;
  if (variant >= kNumVariants) variant = 0;
  Kern kern = kVariants[bf16 ? 1 : 0][variant];
  static thread_local bool attr_set[2][kNumVariants] = {};
  if (!attr_set[bf16 ? 1 : 0][variant]) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes);
    if (e != cudaSuccess) {
      set_error("prefill_tc: cudaFuncSetAttribute failed: %s", cudaGetErrorString(e));
      return B200ATTN_E_CUDA;
    }
    attr_set[bf16 ? 1 : 0][variant] = true;
  }
  const int BQ = 2 * P.block_q;
  const dim3 grid((unsigned)(a.num_tokens / BQ + a.num_seqs), (unsigned)a.num_kv_heads);
  kern<<<grid, kThreads, kSmemBytes, stream>>>(P, qmap, kmap, vmap);
  note_launch("attn_prefill_tc");
  return check_launch("attn_prefill_tc");
}

}  // namespace b200attn
