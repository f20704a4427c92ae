// bgc_k0.cuh — K0 policy_eval (row A1).  Part of libbgc.so's device code; included through bgc_kernels.cuh.
#pragma once
#include "bgc_common.cuh"

namespace bgc {

// ------------------------------------------------------------------ K0 policy_eval
// Decision of one flattened request; restates src/admission.rs:241-431 in evaluation order.
// Returns code | patch_mask << 8.
__device__ __forceinline__ uint32_t policy_decide(uint32_t op, uint32_t f) {
  const bool has_username = f & 1u, normal = f & 2u, in_group = f & 4u, has_object = f & 8u;
  const bool has_name = f & 16u, name_eq = f & 32u, parse_ok = f & 64u, ku_nonempty = f & 128u;
  const bool has_quota = f & 256u, has_rb = f & 512u;
  auto ret = [](uint32_t klass, uint32_t reason, uint32_t patch) {
    return (klass << 4) | reason | (patch << 8);
  };
  if (!has_username) return ret(3, 1, 0);                      // :248-255 invalid
  if (op == 0) {                                               // Create :273-283
    if (normal && !in_group) return ret(2, 2, 0);
  } else if (op == 2) {                                        // Delete :284-294
    return normal ? ret(2, 3, 0) : ret(0, 12, 0);
  } else if (op == 1) {                                        // Update :295-304
    if (normal) return ret(2, 4, 0);
  } else {
    return ret(3, 5, 0);                                       // :305-309 invalid operation
  }
  if (!has_object) return ret(0, 13, 0);                       // :312-318
  if (!has_name) return ret(3, 6, 0);                          // :321-328
  if (normal && !name_eq) return ret(2, 7, 0);                 // :331-338
  if (!parse_ok) return ret(3, 8, 0);                          // :341-347
  uint32_t patch = 0;
  if (normal) patch |= 1;                                      // :352-358 P_USER
  else if (!ku_nonempty) return ret(2, 9, 0);                  // :359-373
  if (has_quota && normal) return ret(2, 10, 0);               // :376-383
  if (!has_rb) patch |= normal ? (2u | 4u) : 2u;               // :386-416 P_RB (+ SUBJ_ORIG)
  else if (normal) return ret(2, 11, 0);                       // :417-424
  return ret(patch ? 1u : 0u, 0, patch);                       // :426-430
}

// The decision is a function of 12 bits, so each CTA tabulates it once (4096 x u16 = 8 KB of
// shared memory) and the streaming loop is one LDS per request: 3 B in, 2 B out per request.
constexpr int kK0Threads = 256;
constexpr int kK0PerThread = 32;

template <int ST>
static __global__ void __launch_bounds__(kK0Threads)
k0_policy_eval(const uint8_t* __restrict__ op, const uint16_t* __restrict__ flags, size_t n,
               uint8_t* __restrict__ code, uint8_t* __restrict__ patch_mask, int vec_ok) {
  __shared__ uint16_t lut[4096];
  for (uint32_t i = threadIdx.x; i < 4096; i += kK0Threads)
    lut[i] = (uint16_t)policy_decide(i >> 10, i & 1023u);
  __syncthreads();

  // 32 requests per thread, every access a whole 32-byte sector: one 256-bit op load, two 256-bit
  // flags loads, one 256-bit store each for code and patch_mask
  const size_t groups = vec_ok ? n / kK0PerThread : 0;
  for (size_t g = (size_t)blockIdx.x * kK0Threads + threadIdx.x; g < groups;
       g += (size_t)gridDim.x * kK0Threads) {
    const size_t base = g * kK0PerThread;
    uint32_t ow[8], fw[16], cw[8], pw[8];
    ldg_nc_u32x8(op + base, ow);
    ldg_nc_u32x8(flags + base, *reinterpret_cast<uint32_t(*)[8]>(&fw[0]));
    ldg_nc_u32x8(flags + base + 16, *reinterpret_cast<uint32_t(*)[8]>(&fw[8]));
#pragma unroll
    for (int w = 0; w < 8; ++w) {
      uint32_t c = 0, p = 0;
#pragma unroll
      for (int b = 0; b < 4; ++b) {
        const uint32_t raw = (ow[w] >> (8 * b)) & 0xFFu;   // anything outside 0..2 is "other"
        const uint32_t opv = raw > 2u ? 3u : raw;
        const uint32_t fl = (fw[w * 2 + (b >> 1)] >> (16 * (b & 1))) & 1023u;
        const uint32_t e = lut[(opv << 10) | fl];
        c |= (e & 0xFFu) << (8 * b);
        p |= (e >> 8) << (8 * b);
      }
      cw[w] = c;
      pw[w] = p;
    }
    stg_u32x8<ST>(code + base, cw);
    stg_u32x8<ST>(patch_mask + base, pw);
  }
  // scalar tail (and the whole batch when the pointers are not 32-byte aligned)
  for (size_t i = groups * kK0PerThread + (size_t)blockIdx.x * kK0Threads + threadIdx.x; i < n;
       i += (size_t)gridDim.x * kK0Threads) {
    const uint32_t raw = op[i];
    const uint32_t e = lut[((raw > 2u ? 3u : raw) << 10) | (flags[i] & 1023u)];
    code[i] = (uint8_t)(e & 0xFFu);
    patch_mask[i] = (uint8_t)(e >> 8);
  }
}

}  // namespace bgc
