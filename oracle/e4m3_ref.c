/* CPU oracle (test infrastructure, not product).
 *
 * Plain-C restatement of the byte-level arithmetic of the reference cache
 * write:
 *   - `__nv_cvt_float_to_fp8(x, __NV_SATFINITE, __NV_E4M3)` as used by
 *     scaled_vec_conversion<uint8_t, bf16> (reference
 *     vllm/csrc/quantization/w8a8/fp8/nvidia/quant_utils.cuh:479-502):
 *     round-to-nearest-even onto the OCP e4m3fn grid (bias 7, 3 mantissa bits,
 *     subnormals down to 2^-9, max finite 448, no infinities), saturating,
 *     NaN -> 0x7f.
 *   - the slot scatter of reshape_and_cache_flash_kernel
 *     (reference vllm/csrc/cache_kernels.cu:286-358).
 *
 * Build: `make -C oracle` -> oracle/_build/liboracle_e4m3.so (gcc only).
 */
#include <math.h>
#include <stdint.h>
#include <string.h>

uint8_t oracle_f32_to_e4m3(float x) {
  uint32_t bits;
  memcpy(&bits, &x, 4);
  uint8_t sign = (uint8_t)((bits >> 24) & 0x80u);
  if (isnan(x)) return 0x7f;
  float a = fabsf(x);
  if (a >= 448.0f) return (uint8_t)(sign | 0x7e); /* satfinite (also +-inf) */
  if (a == 0.0f) return sign;
  /* quantum: 2^(e-3) for normals (e = floor(log2 a), clamped to >= -6), 2^-9 below */
  int e;
  (void)frexpf(a, &e); /* a = m * 2^e, m in [0.5, 1) -> floor(log2 a) = e - 1 */
  int fl = e - 1;
  if (fl < -6) fl = -6;
  float q = ldexpf(1.0f, fl - 3);
  /* a / q is exact (power-of-two scaling); nearbyintf rounds half to even */
  float n = nearbyintf(a / q);
  float r = n * q;
  if (r >= 448.0f + 1e-3f) return (uint8_t)(sign | 0x7e);
  /* re-encode r */
  if (r == 0.0f) return sign;
  int e2;
  (void)frexpf(r, &e2);
  int fl2 = e2 - 1;
  uint8_t out;
  if (fl2 < -6) { /* subnormal: r = k * 2^-9, k in 1..7 */
    out = (uint8_t)(int)(r * 512.0f);
  } else {
    int mant = (int)(r / ldexpf(1.0f, fl2 - 3)) - 8; /* 0..7 */
    out = (uint8_t)(((fl2 + 7) << 3) | mant);
  }
  return (uint8_t)(sign | out);
}

float oracle_e4m3_to_f32(uint8_t b) {
  int sign = b & 0x80, e = (b >> 3) & 0xf, m = b & 7;
  float v;
  if (e == 0xf && m == 7) return NAN;
  if (e == 0) v = ldexpf((float)m, -9);
  else v = ldexpf((float)(8 + m), e - 10);
  return sign ? -v : v;
}

void oracle_quantize_e4m3(const float* src, uint8_t* dst, int64_t n, float scale) {
  for (int64_t i = 0; i < n; ++i) dst[i] = oracle_f32_to_e4m3(src[i] / scale);
}

/* copy path for 16-bit elements; strides in elements */
void oracle_reshape_and_cache_flash_u16(const uint16_t* key, const uint16_t* value, uint16_t* k_cache,
                                        uint16_t* v_cache, const int64_t* slot_mapping, int64_t num_tokens,
                                        int64_t num_heads, int64_t head_size, int64_t block_size,
                                        int64_t key_stride, int64_t value_stride, int64_t block_stride,
                                        int64_t page_stride, int64_t v_block_stride) {
  int64_t n = num_heads * head_size;
  for (int64_t t = 0; t < num_tokens; ++t) {
    int64_t slot = slot_mapping[t];
    if (slot < 0) continue;
    int64_t b = slot / block_size, o = slot % block_size;
    memcpy(k_cache + b * block_stride + o * page_stride, key + t * key_stride, (size_t)n * 2);
    memcpy(v_cache + b * v_block_stride + o * page_stride, value + t * value_stride, (size_t)n * 2);
  }
}
