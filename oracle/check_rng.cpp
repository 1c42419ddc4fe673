// check_rng.cpp -- ORACLE PINNING (test infrastructure, not product code).
// Pins oracle/pf_oracle.c's RNG restatement against the real third-party code the
// reference calls: libstdc++'s std::default_random_engine (== minstd_rand0),
// std::uniform_real_distribution<double> (generate_canonical<double,53>), as used at
// /root/reference/pf_mpe_lib/src/pose_estimator.cpp:476-477, :510-516, :523, :563-565,
// :585-587, :671.  Exit code 0 iff every draw is bit-identical.
#include <cstdio>
#include <cstring>
#include <random>
#include "pf_oracle.h"

static_assert(std::is_same<std::default_random_engine, std::minstd_rand0>::value,
              "default_random_engine is not minstd_rand0 on this toolchain");

int main()
{
  const uint32_t seeds[] = {0u, 1u, 12345u, 2147483646u, 2147483647u, 2147483648u, 4294967295u, 20240001u};
  long mism = 0, total = 0;
  for (uint32_t seed : seeds) {
    std::default_random_engine gen(seed);
    orc_rng g;
    orc_rng_seed(&g, seed);
    std::uniform_real_distribution<double> dA(-0.015 * 0.2, 0.015 * 0.2), dT(-0.035 * 0.05, 0.035 * 0.05), dR(0, 1);
    for (int i = 0; i < 200000; ++i) {
      double a = dA(gen), b = orc_rng_uniform(&g, -0.015 * 0.2, 0.015 * 0.2);
      double c = dT(gen), d = orc_rng_uniform(&g, -0.035 * 0.05, 0.035 * 0.05);
      double e = dR(gen), f = orc_rng_uniform(&g, 0.0, 1.0);
      mism += std::memcmp(&a, &b, 8) != 0;
      mism += std::memcmp(&c, &d, 8) != 0;
      mism += std::memcmp(&e, &f, 8) != 0;
      total += 3;
    }
    // raw engine + skip-ahead
    std::default_random_engine gen2(seed);
    orc_rng g2;
    orc_rng_seed(&g2, seed);
    uint32_t x0 = g2.x;
    for (uint64_t i = 1; i <= 5000; ++i) {
      uint32_t r = (uint32_t)gen2();
      mism += r != orc_rng_next(&g2);
      mism += r != orc_rng_skip(x0, i);
      total += 2;
    }
    gen2.discard(1000003);
    mism += (uint32_t)gen2() != orc_rng_skip(x0, 5000 + 1000003 + 1);
    total += 1;
  }
  std::printf("check_rng: %ld comparisons, %ld mismatches\n", total, mism);
  return mism ? 1 : 0;
}
