// common.cuh — shared device/host helpers of libasvrl.so (sm_100a).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/asvrl.h"

namespace asvrl {

// ---- error plumbing (no exceptions, no aborts, nothing printed: include/asvrl.h) ----
void set_error(const char* fmt, ...);
int fail(int code, const char* fmt, ...);
int check_launch(const char* what);

// ---- Philox4x32-10 (Salmon et al., SC'11).  The stream layout is part of the ABI contract
// (documented in include/asvrl.h) and is restated independently by the CPU oracle. ----
enum : uint32_t { STREAM_SENSOR = 1, STREAM_RESET = 2, STREAM_TAU = 3, STREAM_REPLAY = 4, STREAM_EXPLORE = 5 };

__host__ __device__ __forceinline__ uint32_t mulhi32(uint32_t a, uint32_t b) {
#ifdef __CUDA_ARCH__
  return __umulhi(a, b);
#else
  return (uint32_t)(((uint64_t)a * b) >> 32);
#endif
}

__host__ __device__ __forceinline__ uint4 philox4x32_10(uint4 c, uint2 k) {
#pragma unroll
  for (int round = 0; round < 10; ++round) {
    uint32_t hi0 = mulhi32(0xD2511F53u, c.x), lo0 = 0xD2511F53u * c.x;
    uint32_t hi1 = mulhi32(0xCD9E8D57u, c.z), lo1 = 0xCD9E8D57u * c.z;
    c = make_uint4(hi1 ^ c.y ^ k.x, lo1, hi0 ^ c.w ^ k.y, lo0);
    k.x += 0x9E3779B9u;
    k.y += 0xBB67AE85u;
  }
  return c;
}

// ctr = (block | slot<<8 | ego<<16, global env, step lo, step hi (24 bit) | stream<<24)
__host__ __device__ __forceinline__ uint4 philox_at(uint64_t seed, uint32_t stream, uint64_t step,
                                                    int64_t env, uint32_t ego, uint32_t slot,
                                                    uint32_t block) {
  uint4 c = make_uint4(block | (slot << 8) | (ego << 16), (uint32_t)env, (uint32_t)step,
                       ((uint32_t)(step >> 32) & 0x00FFFFFFu) | (stream << 24));
  return philox4x32_10(c, make_uint2((uint32_t)seed, (uint32_t)(seed >> 32)));
}

// 24-bit uniform strictly inside (0,1)
__host__ __device__ __forceinline__ float u01(uint32_t w) {
  return ((float)(w >> 8) + 0.5f) * (1.0f / 16777216.0f);
}

}  // namespace asvrl
