// Host unit test of multimodal-2025_b200/csrc/rank_keys.h (the ordering rank.cu compiles).
// TEST INFRASTRUCTURE: built and run by tests/test_rank_keys_cpu.py with g++.
// For random rows with heavy ties, +-0, +-inf and NaNs, the rank obtained by COUNTING
// composites greater than the target's must equal the target's position in a stable
// descending sort with torch's order (NaN first, then by VALUE — so -0.0 and +0.0 tie — and
// ties by index).  The comparator below is torch's rule, written without the key transform.
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <random>
#include <vector>

#include "../multimodal-2025_b200/csrc/rank_keys.h"

static uint32_t bits(float f) { uint32_t u; std::memcpy(&u, &f, 4); return u; }

// reference order, written independently of the key transform
static bool ref_before(float a, uint32_t ia, float b, uint32_t ib) {
  const bool na = std::isnan(a), nb = std::isnan(b);
  if (na != nb) return na;                 // NaN sorts first
  if (!na && a != b) return a > b;         // value compare: -0.0 == +0.0
  return ia < ib;                          // stable
}

int main() {
  std::mt19937 rng(12345);
  const float specials[] = {0.f, -0.f, INFINITY, -INFINITY, NAN, -NAN, 1.f, -1.f, 3.f,
                            1e-45f, -1e-45f, 3.4028235e38f, -3.4028235e38f};
  long checked = 0;
  for (int trial = 0; trial < 400; ++trial) {
    const int n = 1 + (int)(rng() % 300);
    std::vector<float> v(n);
    for (int i = 0; i < n; ++i) {
      const unsigned r = rng() % 10;
      if (r < 3) v[i] = specials[rng() % (sizeof(specials) / sizeof(float))];
      else if (r < 7) v[i] = (float)((int)(rng() % 9) - 4) * 0.5f;   // few distinct values
      else v[i] = std::ldexp((float)(rng() % 2000) - 1000.f, (int)(rng() % 40) - 20);
    }
    std::vector<uint32_t> order(n);
    for (int i = 0; i < n; ++i) order[i] = i;
    std::stable_sort(order.begin(), order.end(), [&](uint32_t a, uint32_t b) {
      return ref_before(v[a], a, v[b], b);
    });
    std::vector<int> pos(n);
    for (int p = 0; p < n; ++p) pos[order[p]] = p;
    const uint32_t off = rng() % 1000000;  // positions need not start at 0 (row0 of a slab)
    for (int t = 0; t < n; ++t) {
      const uint64_t tc = mm25_rank_composite(mm25_rank_key_bits(bits(v[t])), off + t);
      int cnt = 0;
      for (int i = 0; i < n; ++i)
        cnt += mm25_rank_composite(mm25_rank_key_bits(bits(v[i])), off + i) > tc;
      if (cnt != pos[t]) {
        std::printf("FAIL trial %d n %d target %d: count %d sort %d\n", trial, n, t, cnt, pos[t]);
        return 1;
      }
      ++checked;
    }
  }
  std::printf("rank keys ok: %ld targets\n", checked);
  return 0;
}
