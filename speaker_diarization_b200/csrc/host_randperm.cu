// Host-side helper (no device code): the frame shuffles of models.py:229-232 for a whole batch in one call.
//
// The reference draws one torch.randperm(ilen) per sequence from torch's GLOBAL CPU generator, in batch order, also
// in eval mode; the drop-in must consume that generator exactly like it.  Calling torch.randperm 64 times per batch
// costs ~0.45 ms of Python / dispatcher time per step - the largest single item of the host path, which is what
// limits 8 ranks sharing one host.  torch's CPU generator is a 32-bit Mersenne Twister (ATen/core/MT19937RNGEngine.h)
// and randperm on the CPU is the Fisher-Yates loop  z = random() % (n - i); swap(r[i], r[i + z])  (ATen
// TensorFactories.cpp: randperm_cpu, n < 2^32 / 20): the caller hands over the generator's serialised state
// (torch.get_rng_state(): left, next, state[624]), this function advances it by the same draws, writes the
// permutations, and the caller stores the state back (torch.set_rng_state).  tests/test_host_randperm_cpu.py checks
// permutations AND final generator state against torch.randperm itself.
#include <stdint.h>

#include "common.cuh"

namespace {
constexpr int kN = 624, kM = 397;
constexpr uint32_t kMatrixA = 0x9908b0dfu, kUMask = 0x80000000u, kLMask = 0x7fffffffu;
inline uint32_t twist(uint32_t u, uint32_t v) { return (((u & kUMask) | (v & kLMask)) >> 1) ^ ((v & 1u) ? kMatrixA : 0u); }

struct Mt {
  int64_t left;
  uint32_t next;
  uint32_t s[kN];
  void next_state() {
    uint32_t* p = s;
    left = kN; next = 0;
    for (int j = kN - kM + 1; --j; p++) *p = p[kM] ^ twist(p[0], p[1]);
    for (int j = kM; --j; p++) *p = p[kM - kN] ^ twist(p[0], p[1]);
    *p = p[kM - kN] ^ twist(p[0], s[0]);
  }
  uint32_t operator()() {
    if (--left == 0) next_state();
    uint32_t y = s[next++];
    y ^= (y >> 11); y ^= (y << 7) & 0x9d2c5680u; y ^= (y << 15) & 0xefc60000u; y ^= (y >> 18);
    return y;
  }
};
}  // namespace

extern "C" int eda_host_randperm_batch(int32_t* left, uint64_t* next, uint64_t* state624, const int32_t* lens, int n_seq,
                                       int32_t* const* rows) {
  EDA_REQUIRE(left && next && state624 && lens && rows && n_seq >= 0, "eda_host_randperm_batch: null argument");
  EDA_REQUIRE(*left >= 1 && *left <= kN && *next <= (uint64_t)kN, "eda_host_randperm_batch: not a Mersenne Twister state (left %d)", *left);
  Mt g;
  g.left = *left; g.next = (uint32_t)*next;
  for (int i = 0; i < kN; ++i) g.s[i] = (uint32_t)state624[i];
  for (int b = 0; b < n_seq; ++b) {
    const int n = lens[b];
    EDA_REQUIRE(n >= 0 && (n == 0 || rows[b]), "eda_host_randperm_batch: bad length / row pointer of sequence %d", b);
    int32_t* r = rows[b];
    for (int i = 0; i < n; ++i) r[i] = i;
    for (int i = 0; i < n - 1; ++i) {
      const int z = (int)(g() % (uint32_t)(n - i));
      const int32_t tmp = r[i]; r[i] = r[i + z]; r[i + z] = tmp;
    }
  }
  *left = (int32_t)g.left; *next = g.next;
  for (int i = 0; i < kN; ++i) state624[i] = g.s[i];
  return EDA_OK;
}
