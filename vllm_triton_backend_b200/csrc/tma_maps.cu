// Host-side TMA descriptor builders shared by the decode and prefill kernels.
//
// The paged cache is handed to us as a strided [NB, bs, Hkv, D] view whose block
// stride is arbitrary (K/V planes or K/V interleaved per block),
// so the descriptors are built over the strides, not over a dense shape.  A
// 128-byte row chunk is the innermost TMA dimension so the hardware 128-byte
// swizzle lands every (token, 128-byte chunk) where ldmatrix / UMMA expect it.
#include <unordered_map>

#include "common.cuh"

namespace b200attn {
namespace {

struct MapKey {
  const void* base;
  int64_t s0, s1, s2, n0, pair;
  int a, b, c, d, e, f, g;
  int kind, device;
  bool operator==(const MapKey& o) const {
    return base == o.base && s0 == o.s0 && s1 == o.s1 && s2 == o.s2 && n0 == o.n0 && pair == o.pair && a == o.a &&
           b == o.b && c == o.c && d == o.d && e == o.e && f == o.f && g == o.g && kind == o.kind && device == o.device;
  }
};
struct MapKeyHash {
  size_t operator()(const MapKey& k) const {
    uint64_t h = 0x9e3779b97f4a7c15ull;
    auto mix = [&](uint64_t v) { h ^= v + 0x9e3779b97f4a7c15ull + (h << 6) + (h >> 2); };
    mix(reinterpret_cast<uint64_t>(k.base)), mix((uint64_t)k.s0), mix((uint64_t)k.s1), mix((uint64_t)k.s2);
    mix((uint64_t)k.n0), mix((uint64_t)k.pair), mix(((uint64_t)(uint32_t)k.a << 32) | (uint32_t)k.b);
    mix(((uint64_t)(uint32_t)k.c << 32) | (uint32_t)k.d), mix(((uint64_t)(uint32_t)k.e << 32) | (uint32_t)k.f);
    mix(((uint64_t)(uint32_t)k.g << 32) | (uint32_t)(k.kind * 64 + k.device));
    return (size_t)h;
  }
};
// Per-thread descriptor cache.  A model has two cache maps per layer (K and V) for each kernel family plus a query
// map per step on the tcgen05 path, so the table holds a few thousand entries (a 32-slot ring thrashed from 16 layers
// up); it is emptied when full (query maps of long-gone activations are the only entries that go stale).
constexpr size_t kCacheMax = 4096;
thread_local std::unordered_map<MapKey, CUtensorMap, MapKeyHash> g_map_cache;

int current_device_or0() {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) dev = 0;
  return dev;
}

bool lookup(const MapKey& key, CUtensorMap* out) {
  auto it = g_map_cache.find(key);
  if (it == g_map_cache.end()) return false;
  *out = it->second;
  return true;
}
void remember(const MapKey& key, const CUtensorMap& map) {
  if (g_map_cache.size() >= kCacheMax) g_map_cache.clear();
  g_map_cache.emplace(key, map);
}

}  // namespace

// 5-D view of a paged cache: (128-byte chunk elems, slot, chunk, head, block).
// `pair_bytes` > 0 (rows of one 128-byte chunk only): the third dimension is not the chunk index
// but the K/V selector — extent 2, stride = distance from the K view to the V view — and a box
// takes both, so one copy brings the K and the V rows of a page: [head][k|v][slot][128 B].
int make_kv_map(CUtensorMap* out, const void* base, int esz, int D, int bs, int hkv, int64_t nb, int64_t s_block,
                int64_t s_slot, int64_t s_head, int box_slots, int box_heads, int box_dc, int64_t pair_bytes) {
  MapKey key{base, s_block, s_slot, s_head, nb, pair_bytes, bs, hkv, D, esz, box_slots, box_heads, box_dc, 0, current_device_or0()};
  if (lookup(key, out)) return B200ATTN_OK;
  EncodeTiledFn fn = encode_tiled_fn();
  if (!fn) {
    set_error("cuTensorMapEncodeTiled unavailable (no CUDA driver?)");
    return B200ATTN_E_CUDA;
  }
  const int epc = 128 / esz;
  cuuint64_t dims[5] = {(cuuint64_t)epc, (cuuint64_t)bs, (cuuint64_t)(D * esz / 128), (cuuint64_t)hkv, (cuuint64_t)nb};
  cuuint64_t strides[4] = {(cuuint64_t)(s_slot * esz), 128, (cuuint64_t)(s_head * esz), (cuuint64_t)(s_block * esz)};
  cuuint32_t box[5] = {(cuuint32_t)epc, (cuuint32_t)box_slots, (cuuint32_t)box_dc, (cuuint32_t)box_heads, 1};
  if (pair_bytes > 0) {
    if (D * esz != 128) {
      set_error("kv map: K/V pairing needs 128-byte rows");
      return B200ATTN_E_UNSUPPORTED;
    }
    dims[2] = 2, strides[1] = (cuuint64_t)pair_bytes, box[2] = 2;
  }
  cuuint32_t estr[5] = {1, 1, 1, 1, 1};
  CUtensorMapDataType dt = esz == 1 ? CU_TENSOR_MAP_DATA_TYPE_UINT8 : CU_TENSOR_MAP_DATA_TYPE_UINT16;
  CUresult r = fn(out, dt, 5, const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                  CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("kv map: cuTensorMapEncodeTiled failed (%d) strides=(%lld,128,%lld,%lld)B box=(%d,%d,%d,%d,1)", (int)r,
              (long long)strides[0], (long long)strides[2], (long long)strides[3], epc, box_slots, box_dc, box_heads);
    return B200ATTN_E_CUDA;
  }
  remember(key, *out);
  return B200ATTN_OK;
}

// 4-D view of a [T, H, D] 2-byte tensor (queries): (64 elems, chunk, head, token);
// a box is (64, 1, box_heads, box_tokens), i.e. rows ordered token-major then
// head — the row order of the reference's Q tile (triton_unified_attention.py:343-346).
int make_thd_map(CUtensorMap* out, const void* base, int D, int heads, int64_t tokens, int64_t s_token, int64_t s_head,
                 int box_heads, int box_tokens) {
  MapKey key{base, s_token, s_head, 0, tokens, 0, D, heads, box_heads, box_tokens, 0, 0, 0, 1, current_device_or0()};
  if (lookup(key, out)) return B200ATTN_OK;
  EncodeTiledFn fn = encode_tiled_fn();
  if (!fn) {
    set_error("cuTensorMapEncodeTiled unavailable (no CUDA driver?)");
    return B200ATTN_E_CUDA;
  }
  cuuint64_t dims[4] = {64, (cuuint64_t)(D / 64), (cuuint64_t)heads, (cuuint64_t)tokens};
  cuuint64_t strides[3] = {128, (cuuint64_t)(s_head * 2), (cuuint64_t)(s_token * 2)};
  cuuint32_t box[4] = {64, 1, (cuuint32_t)box_heads, (cuuint32_t)box_tokens};
  cuuint32_t estr[4] = {1, 1, 1, 1};
  CUresult r = fn(out, CU_TENSOR_MAP_DATA_TYPE_UINT16, 4, const_cast<void*>(base), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("thd map: cuTensorMapEncodeTiled failed (%d) strides=(128,%lld,%lld)B box=(64,1,%d,%d)", (int)r,
              (long long)strides[1], (long long)strides[2], box_heads, box_tokens);
    return B200ATTN_E_CUDA;
  }
  remember(key, *out);
  return B200ATTN_OK;
}

}  // namespace b200attn
