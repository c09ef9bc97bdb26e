// Block-floating-point (BFP16x8) and MXFP8 gradient codecs, usable from host C++ and CUDA.
//
// BFP16x8: 16 fp32 values share one 8-bit exponent, each keeps an int8 two's-complement
// mantissa (8.5 bit/value).  Two numerics modes:
//   * HW_EXACT - bit-for-bit what the reference NIC datapath produces
//                (encode: hw/bf16_to_bfp_core.sv:81-127 + truncate hw/bfp_adapter.sv:147-154,
//                 decode: hw/bfp_to_bf16_core.sv:55-120 + hw/a10_leading_zero_counter.sv:58-72),
//                including floor rounding, the unconditional hidden one, and q==0 -> +2^(E-150).
//   * CLEAN    - the intended behaviour: round-to-nearest-even, zero stays zero,
//                decode is exactly q * 2^(E-133).
// Wire/frame layout (hw/bfp_adapter.sv:59-70,:279-379 restated byte-for-byte):
//   one frame = 512 values = 512 mantissa bytes in element order, then 32 exponent bytes = 544 B.
//
// MXFP8: OCP microscaling, 32 e4m3 values share one ue8m0 scale (8.25 bit/value); the layout
// (values then scales per 512-value frame: 512 + 16 B) mirrors the BFP frame so the same
// kernels move both.  No counterpart in the reference (north-star addition).
#pragma once
#include <stdint.h>
#include <string.h>

#if defined(__CUDACC__)
#include <cuda_fp8.h>
#define AINIC_HD __host__ __device__ __forceinline__
#else
#define AINIC_HD inline
#endif

namespace ainic {

enum BfpMode : int { BFP_HW_EXACT = 0, BFP_CLEAN = 1 };

constexpr int kBfpBlock = 16;          // values per shared exponent (hw/all_reduce.sv:746 NUM_FP=16)
constexpr int kFrameVals = 512;        // 32 blocks (hw/bfp_adapter.sv:67-70: 16 mantissa flits + 1 exponent flit)
constexpr int kBfpFrameBytes = 544;    // 512 + 32
constexpr int kBfpExpOff = 512;
constexpr int kMxBlock = 32;
constexpr int kMxFrameBytes = 528;     // 512 e4m3 + 16 ue8m0
constexpr int kMxScaleOff = 512;

AINIC_HD uint32_t f2u(float f) {
#if defined(__CUDA_ARCH__)
  return __float_as_uint(f);
#else
  uint32_t u; memcpy(&u, &f, 4); return u;
#endif
}
AINIC_HD float u2f(uint32_t u) {
#if defined(__CUDA_ARCH__)
  return __uint_as_float(u);
#else
  float f; memcpy(&f, &u, 4); return f;
#endif
}
AINIC_HD int clz32(uint32_t x) {
#if defined(__CUDA_ARCH__)
  return __clz((int)x);
#else
  return x ? __builtin_clz(x) : 32;
#endif
}

// ---------------------------------------------------------------- BFP, reference-exact
// Biased exponent field of an fp32 word; the block's shared exponent is the max of these.
AINIC_HD uint32_t bfp_exp_field(uint32_t bits) { return (bits >> 23) & 0xFFu; }

// Quantise one fp32 word against shared exponent E (HW_EXACT): q = floor(+-t / 2^17),
// t = {1,frac} >> (E - e) (hidden one inserted unconditionally, hw/bfp_adapter.sv:131).
AINIC_HD int32_t bfp_quant_hw(uint32_t bits, uint32_t E) {
  uint32_t e = (bits >> 23) & 0xFFu;
  uint32_t d = E - e;                                    // E >= e by construction
  uint32_t m = 0x800000u | (bits & 0x7FFFFFu);
  uint32_t t = d > 31u ? 0u : (m >> d);                  // hw/barrel_shifter.sv:44-51 clears for >= 32
  int32_t u = (bits >> 31) ? -(int32_t)t : (int32_t)t;   // 25-bit two's complement
  return u >> 17;                                        // drop LSB (>>1) then keep top 8 of 24 (>>16)
}

// Decode one int8 mantissa (HW_EXACT).  Returns the fp32 bit pattern the RTL emits.
AINIC_HD uint32_t bfp_dequant_hw_bits(int32_t q, uint32_t E) {
  uint32_t sign = q < 0 ? 1u : 0u;
  uint32_t mag = (uint32_t)(q < 0 ? -q : q) << 16;        // 24-bit magnitude, -128 -> 2^23
  int z = mag ? (clz32(mag) - 8) : 24;                   // clz over 24 bits; all-zero -> 24 (LZC quirk)
  uint32_t expo = (E + 1u - (uint32_t)z) & 0xFFu;        // 8-bit wrap as in the RTL
  uint32_t sh = z >= 24 ? 0u : ((mag << z) & 0xFFFFFEu); // shift >= 24 clears; LSB forced 0
  return (sign << 31) | (expo << 23) | (sh & 0x7FFFFFu);
}
AINIC_HD float bfp_dequant_hw(int32_t q, uint32_t E) { return u2f(bfp_dequant_hw_bits(q, E)); }

// ---------------------------------------------------------------- BFP, clean
// 2^(E-133) as a float (0 when it would be subnormal).
AINIC_HD float bfp_scale(uint32_t E) { return E > 6u ? u2f((E - 6u) << 23) : 0.0f; }
// 2^(133-E): multiplier that maps a value with block exponent E onto the int8 grid.
AINIC_HD float bfp_inv_scale(uint32_t E) {
  // 133 - E + 127 = 260 - E must be a normal exponent: E in [7, 254] -> [6, 253]
  return (E > 6u && E < 255u) ? u2f((260u - E) << 23) : 0.0f;
}
AINIC_HD int32_t bfp_quant_clean(float x, uint32_t E) {
  float s = x * bfp_inv_scale(E);
#if defined(__CUDA_ARCH__)
  // Clamp first (127 is on the grid and rounding is monotonic, so the result is the same), then round with
  // the 1.5 * 2^23 trick: the fp32 add rounds to nearest-even at integer granularity and leaves the integer in
  // the low mantissa bits.  Three FP32-pipe instructions instead of a quarter-rate F2I + two integer min/max -
  // the quantisation is what the compressed GEMM epilogue and the all-reduce re-encode spend their time on -
  // and callers that keep only the low byte need no subtraction at all (the magic constant's low byte is 0).
  const float t = fminf(fmaxf(s, -127.0f), 127.0f) + 12582912.0f;
  return (int32_t)(f2u(t) - 0x4B400000u);
#else
  int32_t q = (int32_t)__builtin_rintf(s);               // default rounding mode = RNE
  return q > 127 ? 127 : (q < -127 ? -127 : q);
#endif
}
AINIC_HD float bfp_dequant_clean(int32_t q, uint32_t E) { return (float)q * bfp_scale(E); }

AINIC_HD int32_t bfp_quant(float x, uint32_t E, int mode) {
  return mode == BFP_HW_EXACT ? bfp_quant_hw(f2u(x), E) : bfp_quant_clean(x, E);
}
AINIC_HD float bfp_dequant(int32_t q, uint32_t E, int mode) {
  return mode == BFP_HW_EXACT ? bfp_dequant_hw(q, E) : bfp_dequant_clean(q, E);
}

// ---------------------------------------------------------------- MXFP8 (e4m3 + ue8m0 / 32)
// ue8m0 scale for a block whose largest biased fp32 exponent is Emax: 2^(Emax-127-8)
// (e4m3 max normal is 1.75*2^8, so the block max lands in the top binade).
AINIC_HD uint32_t mx_scale_byte(uint32_t Emax) { return Emax > 8u ? Emax - 8u : 0u; }
AINIC_HD float mx_scale(uint32_t sb) { return sb ? u2f(sb << 23) : 0.0f; }
AINIC_HD float mx_inv_scale(uint32_t sb) { return (sb && sb < 254u) ? u2f((254u - sb) << 23) : 0.0f; }

// fp32 -> e4m3 (bias 7, max 448, no inf, NaN = 0x7F), RNE, saturating.
AINIC_HD uint32_t f32_to_e4m3(float x) {
#if defined(__CUDA_ARCH__)
  // cvt.rn.satfinite.e4m3x2.f32: same rounding (RNE), saturation (448) and NaN code as the
  // portable path below, which stays the host oracle (tests compare the two bit for bit).
  return (uint32_t)__nv_cvt_float_to_fp8(x, __NV_SATFINITE, __NV_E4M3);
#else
  uint32_t b = f2u(x);
  uint32_t sign = (b >> 24) & 0x80u;
  uint32_t a = b & 0x7FFFFFFFu;
  if (a > 0x7F800000u) return sign | 0x7Fu;               // NaN
  if (a >= 0x43E80000u) return sign | 0x7Eu;              // >= 464 rounds past 448 -> saturate
  int e = (int)(a >> 23) - 127;
  if (e < -10) return sign;                               // below half the smallest subnormal (2^-9)
  uint32_t m = (a & 0x7FFFFFu) | 0x800000u;               // 24-bit significand
  int shift = e >= -6 ? 20 : (20 + (-6 - e));             // keep 3 fraction bits (normal) or fewer
  uint32_t q = m >> shift;
  uint32_t rem = m & ((1u << shift) - 1u);
  uint32_t half = 1u << (shift - 1);
  if (rem > half || (rem == half && (q & 1u))) q++;
  uint32_t out;
  if (e >= -6) out = ((uint32_t)(e + 7) << 3) + (q - 8u);  // q in [8,16]; carry bumps the exponent
  else out = q;                                           // subnormal: q in [0,8], 8 = smallest normal
  if (out > 0x7Eu) out = 0x7Eu;
  return sign | out;
#endif
}
AINIC_HD float e4m3_to_f32(uint32_t v) {
  uint32_t sign = (v & 0x80u) << 24;
  uint32_t e = (v >> 3) & 0xFu, m = v & 7u;
  if (e == 0xFu && m == 7u) return u2f(sign | 0x7FC00000u);
  if (e == 0u) return u2f(sign | f2u((float)m * 0.001953125f));   // m * 2^-9
  return u2f(sign | ((e + 120u) << 23) | (m << 20));
}

// ---------------------------------------------------------------- frame helpers (host + device)
AINIC_HD size_t bfp_frames(size_t count) { return (count + kFrameVals - 1) / kFrameVals; }
AINIC_HD size_t bfp_bytes(size_t count) { return bfp_frames(count) * (size_t)kBfpFrameBytes; }
AINIC_HD size_t mx_bytes(size_t count) { return bfp_frames(count) * (size_t)kMxFrameBytes; }

}  // namespace ainic
