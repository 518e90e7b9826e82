/*
 * rng.cuh — counter-based generator that replaces RandomDouble() = rand() / (RAND_MAX + 1.0)
 * (Source/RandomNumberGenerator.h:16-19).
 *
 * One stream per (rng_key, pixel, sample); the d-th call of RandomDouble() on a path returns
 *     mix64(seed + (d + 1) * GOLDEN) >> 33        as a 31-bit integer, divided by 2^31
 * with seed = mix64(mix64(key) ^ mix64((pixel << 32 | sample) + GOLDEN)) and mix64 the splitmix64
 * finaliser. 31 bits = the resolution of glibc's rand(), so a path consumes exactly the numbers
 * the reference would have been handed by an interposed rand(); u is in [0, 1) strictly.
 * No state beyond (seed, d): any GPU can render any sample of any pixel.
 */
#ifndef PRT_RNG_CUH
#define PRT_RNG_CUH
#include <stdint.h>

namespace prt {

#define PRT_GOLDEN 0x9E3779B97F4A7C15ull

__host__ __device__ __forceinline__ uint64_t mix64(uint64_t z) {
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
__host__ __device__ __forceinline__ uint64_t stream_seed(uint64_t key, uint32_t pixel, uint32_t sample) {
    const uint64_t s = ((uint64_t)pixel << 32) | (uint64_t)sample;
    return mix64(mix64(key) ^ mix64(s + PRT_GOLDEN));
}

struct Rng {
    uint64_t seed;
    uint32_t dim;
    __device__ __forceinline__ double peek() const { /* value of the next draw, not consumed */
        return (double)(int32_t)(mix64(seed + (uint64_t)(dim + 1u) * PRT_GOLDEN) >> 33) / 2147483648.0;
    }
    __device__ __forceinline__ double next() {
        const double u = peek();
        ++dim;
        return u;
    }
};

}  // namespace prt
#endif
