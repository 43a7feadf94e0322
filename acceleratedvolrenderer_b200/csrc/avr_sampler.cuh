// avr_sampler.cuh — the integrator's sampler on the device.
//   "independent": IndependentSampler (samplers.h:442-476), one PCG32 stream per pixel sample;
//   "zsobol":      ZSobolSampler (samplers.h:225-364): StartPixelSample (:252-255), Get1D (:258-271),
//                  Get2D (:274-292), GetSampleIndex (:301-356); SobolSample with the None / PermuteDigits /
//                  FastOwen / Owen randomizers (util/lowdiscrepancy.h:168-258).  Integer arithmetic
//                  throughout: bit-identical to the reference in both math modes.
// The ZSobol sampler only evaluates Sobol' dimensions 0 and 1.  Dimension 0 is the identity matrix
// (a bit reversal of the index), dimension 1 the Pascal matrix mod 2 whose columns follow
// c[i] = c[i-1] ^ (c[i-1] >> 1), applied in closed form; neither needs the 1024 x 52 table of util/sobolmatrices.cpp
// (tests/test_oracle_golden.py checks the generated columns against that table).
#pragma once
#include "../../include/avr.h"
#include "avr_math.cuh"

namespace avr {

struct SamplerCfg {
    int kind;       // AVR_SAMPLER_*
    int randomize;  // AVR_RANDOMIZE_*
    int seed;
    int log2spp, nbase4;  // ZSobolSampler::log2SamplesPerPixel, nBase4Digits (samplers.h:235-238)
};

// util/math.h:83-96
AVR_HD uint64_t left_shift2(uint64_t x) {
    x &= 0xffffffffull;
    x = (x ^ (x << 16)) & 0x0000ffff0000ffffull;
    x = (x ^ (x << 8)) & 0x00ff00ff00ff00ffull;
    x = (x ^ (x << 4)) & 0x0f0f0f0f0f0f0f0full;
    x = (x ^ (x << 2)) & 0x3333333333333333ull;
    x = (x ^ (x << 1)) & 0x5555555555555555ull;
    return x;
}
AVR_HD uint64_t encode_morton2(uint32_t x, uint32_t y) { return (left_shift2(y) << 1) | left_shift2(x); }

// Hash(int, int), util/hash.h:100-107: MurmurHash64A over one 8-byte block, seed 0
AVR_HD uint64_t hash_2u32(uint32_t a, uint32_t b) {
    const uint64_t m = 0xc6a4a7935bd1e995ull;
    uint64_t h = 0 ^ (8ull * m);
    uint64_t k = uint64_t(a) | (uint64_t(b) << 32);
    k *= m; k ^= k >> 47; k *= m;
    h ^= k; h *= m;
    return murmur_finish(h);
}

// the 24 permutations of samplers.h:303-330, 2 bits per entry (entry d of permutation p: bits 2d..2d+1)
AVR_DEV uint32_t zsobol_perm(int p) {
    // {0,1,2,3} {0,1,3,2} {0,2,1,3} {0,2,3,1} {0,3,2,1} {0,3,1,2} {1,0,2,3} {1,0,3,2} {1,2,0,3} {1,2,3,0}
    // {1,3,2,0} {1,3,0,2} {2,1,0,3} {2,1,3,0} {2,0,1,3} {2,0,3,1} {2,3,0,1} {2,3,1,0} {3,1,2,0} {3,1,0,2}
    // {3,2,1,0} {3,2,0,1} {3,0,2,1} {3,0,1,2}
    // packed 8 permutations per 64-bit word (no local-memory table)
    uint64_t w = p < 8 ? 0xB1E19C6C78D8B4E4ull : (p < 16 ? 0x72D236C68D2D39C9ull : 0x93634B1B87271E4Eull);
    return uint32_t(w >> (8 * (p & 7))) & 0xffu;
}

// (MixBits(v) >> 24) % 24 for v < 2^32: the 64-bit mix with a 32-bit input, and the modulo without a 64-bit
// division: x % 24 = 8 ((x >> 3) % 3) + (x & 7), and 2^32 = 1 (mod 3)
AVR_DEV int zsobol_perm_index32(uint32_t v32) {
    uint64_t v = mix_bits(uint64_t(v32));
    uint32_t lo = uint32_t(v >> 24), hi = uint32_t(v >> 56);  // x = hi * 2^32 + lo, hi < 256
    uint32_t ylo = (lo >> 3) | (hi << 29), yhi = hi >> 3;
    uint32_t r3 = (ylo % 3u + yhi) % 3u;
    return int(8u * r3 + (lo & 7u));
}

AVR_DEV uint64_t zsobol_sample_index(const SamplerCfg &c, uint64_t morton, int dimension) {
    const bool pow2 = c.log2spp & 1;
    const int last = pow2 ? 1 : 0;
    const uint32_t dmix32 = 0x55555555u * (uint32_t)dimension;
    if (2 * c.nbase4 <= 32) {
        // every film up to 4096^2 x 256 spp: the Morton index fits 32 bits, all shifts are 32-bit
        const uint32_t m = uint32_t(morton);
        uint32_t idx = 0;
#pragma unroll 1
        for (int i = c.nbase4 - 1; i >= last; --i) {
            int shift = 2 * i - (pow2 ? 1 : 0);
            uint32_t digit = (m >> shift) & 3u;
            uint32_t higher = shift + 2 < 32 ? m >> (shift + 2) : 0u;
            int p = zsobol_perm_index32(higher ^ dmix32);
            idx |= ((zsobol_perm(p) >> (2 * digit)) & 3u) << shift;
        }
        if (pow2) idx |= (m & 1u) ^ uint32_t(mix_bits(uint64_t((m >> 1) ^ dmix32)) & 1ull);
        return idx;
    }
    uint64_t idx = 0;
    const uint64_t dmix = uint64_t(dmix32);
#pragma unroll 1
    for (int i = c.nbase4 - 1; i >= last; --i) {
        int shift = 2 * i - (pow2 ? 1 : 0);
        int digit = int((morton >> shift) & 3);
        uint64_t higher = morton >> (shift + 2);
        int p = int((mix_bits(higher ^ dmix) >> 24) % 24);
        idx |= uint64_t((zsobol_perm(p) >> (2 * digit)) & 3u) << shift;
    }
    if (pow2) {
        int digit = int(morton & 1);
        idx |= uint64_t(digit ^ int(mix_bits((morton >> 1) ^ dmix) & 1));
    }
    return idx;
}

AVR_DEV uint32_t randomize_bits(int strategy, uint32_t seed, uint32_t v) {
    if (strategy == AVR_RANDOMIZE_NONE) return v;
    if (strategy == AVR_RANDOMIZE_PERMUTE_DIGITS) return seed ^ v;
    if (strategy == AVR_RANDOMIZE_FAST_OWEN) {
        v = __brev(v);
        v ^= v * 0x3d20adeau;
        v += seed;
        v *= (seed >> 16) | 1u;
        v ^= v * 0x05526c56u;
        v ^= v * 0x53a22864u;
        return __brev(v);
    }
    if (seed & 1u) v ^= 1u << 31;
#pragma unroll 1
    for (int b = 1; b < 32; ++b) {
        uint32_t mask = (~0u) << (32 - b);
        if (uint32_t(mix_bits(uint64_t((v & mask) ^ seed))) & (1u << b)) v ^= 1u << (31 - b);
    }
    return v;
}

// SobolSample(a, dim, randomizer), dim in {0, 1}; a < 2^52
AVR_DEV float sobol_sample01(uint64_t a, int dim, int strategy, uint32_t seed) {
    uint32_t v;
    if (dim == 0) {
        v = __brev(uint32_t(a));  // column i = 1 << (31 - i); columns >= 32 are zero in the 32-bit table
    } else {
        // row j of column i is C(i, j) mod 2 = [j is a submask of i] (Lucas): the product is the superset
        // sum over GF(2) of the index bits, six butterfly steps, rows 0..31 kept and bit-reversed
        uint64_t y = a;
        y ^= (y >> 1) & 0x5555555555555555ull;
        y ^= (y >> 2) & 0x3333333333333333ull;
        y ^= (y >> 4) & 0x0f0f0f0f0f0f0f0full;
        y ^= (y >> 8) & 0x00ff00ff00ff00ffull;
        y ^= (y >> 16) & 0x0000ffff0000ffffull;
        y ^= (y >> 32);
        v = __brev(uint32_t(y));
    }
    v = randomize_bits(strategy, seed, v);
    float f = v * 0x1p-32f;
    return f < kOneMinusEps ? f : kOneMinusEps;
}

// ZSobolSampler::Get1D / Get2D for the pixel sample (px, py, samp) at `dimension`.  Deliberately not
// inlined: the independent sampler is the bench default and these bodies (a hash and a 64-bit mix per
// base-4 digit) would otherwise sit in the event ops' instruction stream and register budget.
static __device__ __noinline__ float zsobol_get1d(SamplerCfg c, int px, int py, int samp, int dimension) {
    uint64_t morton = (encode_morton2((uint32_t)px, (uint32_t)py) << c.log2spp) | (uint64_t)(long long)samp;
    uint64_t si = zsobol_sample_index(c, morton, dimension);
    uint32_t h = uint32_t(hash_2u32((uint32_t)(dimension + 1), (uint32_t)c.seed));
    return sobol_sample01(si, 0, c.randomize, h);
}
static __device__ __noinline__ float2 zsobol_get2d(SamplerCfg c, int px, int py, int samp, int dimension) {
    uint64_t morton = (encode_morton2((uint32_t)px, (uint32_t)py) << c.log2spp) | (uint64_t)(long long)samp;
    uint64_t si = zsobol_sample_index(c, morton, dimension);
    uint64_t bits = hash_2u32((uint32_t)(dimension + 2), (uint32_t)c.seed);
    return make_float2(sobol_sample01(si, 0, c.randomize, uint32_t(bits)),
                       sobol_sample01(si, 1, c.randomize, uint32_t(bits >> 32)));
}

// One sampler per path: `rng` is the IndependentSampler stream; for zsobol rng.state holds the
// current dimension and the Morton index is rebuilt from the pixel and the sample index.  The kind is a
// compile-time parameter of the render kernels (a run-time branch costs the independent-sampler kernel
// 6 % through registers and code size: measured 59.4 -> 62.8 ms on the bench workload).
template <bool ZS>
struct PathSamplerRef {
    const SamplerCfg &c;
    RNG &rng;
    int px, py, samp;

    AVR_DEV void start_pixel_sample() {
        if (ZS) {
            rng.state = 0;  // dimension
            rng.inc = 0;
        } else {
            // IndependentSampler::StartPixelSample, samplers.h:457-460
            rng.set_sequence(hash_3u32((uint32_t)px, (uint32_t)py, (uint32_t)c.seed));
            rng.advance((long long)samp * 65536ll);
        }
    }
    // n dimensions whose values nobody reads (the light choice among ONE light, uLight of a delta light, time and
    // lens of a pinhole camera): the stream advances exactly as the reference's, the samples are not computed
    AVR_DEV void skip(int n) {
        if (ZS) rng.state += uint64_t(n);
        else
            for (int i = 0; i < n; ++i) rng.u32();
    }
    AVR_DEV float get1d() {
        if (!ZS) return rng.uniform();
        int dim = int(rng.state);
        rng.state = uint64_t(dim + 1);
        return zsobol_get1d(c, px, py, samp, dim);
    }
    AVR_DEV void get2d(float *a, float *b) {
        if (!ZS) {
            *a = rng.uniform();  // IndependentSampler::Get2D, samplers.h:466
            *b = rng.uniform();
            return;
        }
        int dim = int(rng.state);
        rng.state = uint64_t(dim + 2);
        float2 u = zsobol_get2d(c, px, py, samp, dim);
        *a = u.x;
        *b = u.y;
    }
};

}  // namespace avr
