/*
 * Counter-based random stream of the device-side sampler.
 *
 * The reference draws from one std::mt19937_64 per thread (DNest4/code/RNG.h:17-24,
 * RNG.cpp:28-64): a 2.5 KB sequential state per lane, and std::normal_distribution keeps
 * a hidden spare value.  On the GPU every lane owns a Philox4x32-10 stream instead
 * (Salmon et al., "Parallel random numbers: as easy as 1, 2, 3", SC'11): key = seed +
 * lane, a 64-bit counter, one block per uniform draw, no hidden state — so the stream
 * position of a lane is one integer.
 *
 * This header is the single definition of that stream.  It is compiled into the CUDA
 * kernels (kc_sampler.cu) and into oracle/philox_rng.cpp, the stand-in for
 * DNest4/code/RNG.cpp that lets the UNMODIFIED reference sampler consume the same
 * stream draw for draw.
 *
 *   rand()    = top 53 bits of the block, in [0, 1)
 *   randn()   = sqrt(-2 log(1 - u1)) cos(2 pi u2)        (Box-Muller, u1 drawn first)
 *   randt2()  = randn() / sqrt(-log(rand()))              (RNG.cpp:38-42, normal first)
 *   randh()   = 10^(1.5 - 3 |randt2()|) * randn()         (RNG.cpp:44-49, t2 first)
 *   rand_int(N) = (int) floor(N * rand())                 (RNG.cpp:57-60)
 */
#ifndef KC_PHILOX_H
#define KC_PHILOX_H
#include <stdint.h>

#if defined(__CUDACC__)
#define KC_PHILOX_FN __host__ __device__ __forceinline__
#else
#define KC_PHILOX_FN static inline
#endif

#define KC_PHILOX_M0 0xD2511F53u
#define KC_PHILOX_M1 0xCD9E8D57u
#define KC_PHILOX_W0 0x9E3779B9u
#define KC_PHILOX_W1 0xBB67AE85u

/* One Philox4x32-10 block: counter (ctr, 0), key `key`; returns the first two output
 * words as one 64-bit integer (word 0 high). */
KC_PHILOX_FN uint64_t kc_philox_block(uint64_t key, uint64_t ctr)
{
    uint32_t c0 = (uint32_t)ctr, c1 = (uint32_t)(ctr >> 32), c2 = 0u, c3 = 0u;
    uint32_t k0 = (uint32_t)key, k1 = (uint32_t)(key >> 32);
    for (int r = 0; r < 10; r++) {
        const uint64_t p0 = (uint64_t)KC_PHILOX_M0 * c0;
        const uint64_t p1 = (uint64_t)KC_PHILOX_M1 * c2;
        const uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0;
        const uint32_t n1 = (uint32_t)p1;
        const uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1;
        const uint32_t n3 = (uint32_t)p0;
        c0 = n0; c1 = n1; c2 = n2; c3 = n3;
        k0 += KC_PHILOX_W0;
        k1 += KC_PHILOX_W1;
    }
    return ((uint64_t)c0 << 32) | (uint64_t)c1;
}

/* Uniform in [0, 1) from one block. */
KC_PHILOX_FN double kc_philox_uniform(uint64_t key, uint64_t ctr)
{
    return (double)(kc_philox_block(key, ctr) >> 11) * (1.0 / 9007199254740992.0);
}

#endif /* KC_PHILOX_H */
