// rank_keys.h — the ordering used by rank.cu, as plain inline functions so that the host-side
// unit test (tests/rank_keys_host_test.cc, built with g++ by tests/test_rank_keys_cpu.py) checks
// exactly the code the kernel compiles.
//
// Sort order = torch's descending STABLE sort: larger value first, NaN above +inf, -0.0 == +0.0
// (torch.argsort compares values, so the two zeros tie: AlignCLIP/main/train.py:348,
// main/retrieval.py:69,85), and among equal values the smaller position first.
// An entry is folded into ONE 64-bit composite
//     (key << 32) | (0xFFFFFFFF - position)
// so that "sorts ahead of the target" is a single unsigned 64-bit compare:
//     ahead(e, t)  <=>  composite(e) > composite(t)
// (equal key: the smaller position has the larger low word; the target itself compares equal and
// is not counted).  Positions are < 2^31 (checked by the host wrapper).
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define MM25_HD __host__ __device__ __forceinline__
#else
#define MM25_HD static inline
#endif

// order-preserving key of an fp32 bit pattern: key(a) < key(b) <=> a sorts after b descending
MM25_HD uint32_t mm25_rank_key_bits(uint32_t u) {
  if ((u & 0x7FFFFFFFu) > 0x7F800000u) return 0xFFFFFFFFu;  // NaN (either sign): largest
  if ((u << 1) == 0u) u = 0u;                               // -0.0 ties with +0.0
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
MM25_HD uint64_t mm25_rank_composite(uint32_t key, uint32_t position) {
  return ((uint64_t)key << 32) | (uint64_t)(0xFFFFFFFFu - position);
}
// composite that no entry sorts ahead of (used for columns / rows outside the slab)
MM25_HD uint64_t mm25_rank_composite_never() { return ~(uint64_t)0; }
