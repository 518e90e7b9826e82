/*
 * rng_spec.h — TEST INFRASTRUCTURE (oracle side) statement of the counter-based generator.
 *
 * The reference draws `std::rand() / (RAND_MAX + 1.0)` (Source/RandomNumberGenerator.h:16-19):
 * a 31-bit uniform on glibc, process-global and unseeded. The B200 backend replaces it by a
 * counter-based generator keyed by (rng_key, pixel, sample, dimension) that keeps the reference's
 * 31-bit resolution, so that the unmodified reference code — with rand() interposed by
 * oracle/ref_driver.cpp — can be fed the very same numbers path by path.
 *
 *   mix64    = the splitmix64 output function (Steele, Lea, Flood 2014)
 *   seed     = mix64( mix64(key) ^ mix64( (pixel << 32 | sample) + GOLDEN ) )
 *   draw(d)  = mix64( seed + (d + 1) * GOLDEN ) >> 33          (d = 0,1,2,... in call order)
 *   u        = draw(d) / 2^31                                   in [0, 1)
 *
 * The product restates this independently in pooraytracer_b200/csrc/rng.cuh; this copy is what
 * the oracle (reference driver and CPU port) uses. The two must agree bit for bit — tests/
 * check that through the rendered images.
 */
#ifndef ORACLE_RNG_SPEC_H
#define ORACLE_RNG_SPEC_H
#include <stdint.h>

#define ORC_GOLDEN 0x9E3779B97F4A7C15ull

static inline uint64_t orc_mix64(uint64_t z) {
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
static inline uint64_t orc_stream_seed(uint64_t key, uint32_t pixel, uint32_t sample) {
    uint64_t s = ((uint64_t)pixel << 32) | (uint64_t)sample;
    return orc_mix64(orc_mix64(key) ^ orc_mix64(s + ORC_GOLDEN));
}
/* 31-bit integer, the value an interposed rand() returns */
static inline int32_t orc_draw31(uint64_t seed, uint32_t dim) {
    return (int32_t)(orc_mix64(seed + (uint64_t)(dim + 1u) * ORC_GOLDEN) >> 33);
}
static inline double orc_uniform(uint64_t seed, uint32_t dim) {
    return (double)orc_draw31(seed, dim) / 2147483648.0;
}
#endif
