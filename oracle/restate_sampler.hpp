// oracle/restate_sampler.hpp — TEST INFRASTRUCTURE ONLY (see restate.hpp header).
// ZSobolSampler (samplers.h:225-364): StartPixelSample (:252-255), Get1D (:258-271), Get2D (:274-292),
// GetSampleIndex (:301-356); SobolSample (util/lowdiscrepancy.h:168-180) with the four randomizers
// NoRandomizer / BinaryPermuteScrambler / FastOwenScrambler / OwenScrambler (:205-258);
// EncodeMorton2 (util/math.h:83-96), MixBits / Hash (util/hash.h:70-107).
// Pinned bit for bit against the reference's own ZSobolSampler (refprobe: ref_zsobol_samples,
// ref_sobol_matrices) in tests/test_oracle_golden.py::test_oracle_zsobol_*.
//
// Generator matrices: the ZSobol sampler only ever evaluates Sobol' dimensions 0 and 1.  Dimension 0 is
// the identity (van der Corput), dimension 1 the Pascal matrix mod 2 (primitive polynomial x + 1):
// column i = column i-1 XOR (column i-1 >> 1).  They are generated here, 52 columns each, and compared
// with the reference's util/sobolmatrices.cpp table by the pin test.
#pragma once
#include <algorithm>
#include <cstdint>

#include "restate.hpp"

namespace avro {

constexpr int kSobolMatrixSize = 52;  // util/sobolmatrices.h:16

inline uint32_t sobol_matrix32(int dim, int col) {
    if (dim == 0) return col < 32 ? 1u << (31 - col) : 0u;
    uint64_t v = 1ull << 63;
    for (int i = 0; i < col; ++i) v ^= v >> 1;
    return uint32_t(v >> 32);
}

inline uint32_t reverse_bits32(uint32_t n) {  // util/math.h:56-66
    n = (n << 16) | (n >> 16);
    n = ((n & 0x00ff00ffu) << 8) | ((n & 0xff00ff00u) >> 8);
    n = ((n & 0x0f0f0f0fu) << 4) | ((n & 0xf0f0f0f0u) >> 4);
    n = ((n & 0x33333333u) << 2) | ((n & 0xccccccccu) >> 2);
    n = ((n & 0x55555555u) << 1) | ((n & 0xaaaaaaaau) >> 1);
    return n;
}

inline uint64_t left_shift2(uint64_t x) {  // util/math.h:83-91
    x &= 0xffffffffull;
    x = (x ^ (x << 16)) & 0x0000ffff0000ffffull;
    x = (x ^ (x << 8)) & 0x00ff00ff00ff00ffull;
    x = (x ^ (x << 4)) & 0x0f0f0f0f0f0f0f0full;
    x = (x ^ (x << 2)) & 0x3333333333333333ull;
    x = (x ^ (x << 1)) & 0x5555555555555555ull;
    return x;
}
inline uint64_t encode_morton2(uint32_t x, uint32_t y) { return (left_shift2(y) << 1) | left_shift2(x); }

inline uint64_t hash_2i32(int a, int b) {  // Hash(int, int): 8 bytes through MurmurHash64A, seed 0
    int buf[2] = {a, b};
    return murmur64a((const unsigned char *)buf, 8, 0);
}

enum Randomize { RANDOMIZE_NONE = 0, RANDOMIZE_PERMUTE_DIGITS = 1, RANDOMIZE_FAST_OWEN = 2, RANDOMIZE_OWEN = 3 };

inline uint32_t randomize_bits(int strategy, uint32_t seed, uint32_t v) {
    switch (strategy) {
    case RANDOMIZE_NONE: return v;
    case RANDOMIZE_PERMUTE_DIGITS: return seed ^ v;  // BinaryPermuteScrambler, lowdiscrepancy.h:210-217
    case RANDOMIZE_FAST_OWEN:                        // FastOwenScrambler, :221-237
        v = reverse_bits32(v);
        v ^= v * 0x3d20adeau;
        v += seed;
        v *= (seed >> 16) | 1u;
        v ^= v * 0x05526c56u;
        v ^= v * 0x53a22864u;
        return reverse_bits32(v);
    default:                                         // OwenScrambler, :240-258
        if (seed & 1u) v ^= 1u << 31;
        for (int b = 1; b < 32; ++b) {
            uint32_t mask = (~0u) << (32 - b);
            if (uint32_t(mix_bits((v & mask) ^ seed)) & (1u << b)) v ^= 1u << (31 - b);
        }
        return v;
    }
}

inline float sobol_sample(int64_t a, int dim, int strategy, uint32_t seed) {  // lowdiscrepancy.h:168-180
    uint32_t v = 0;
    for (int i = 0; a != 0; a >>= 1, ++i)
        if (a & 1) v ^= sobol_matrix32(dim, i);
    v = randomize_bits(strategy, seed, v);
    return std::min(v * 0x1p-32f, kOneMinusEps);
}

struct ZSobolSampler {
    int randomize = RANDOMIZE_FAST_OWEN, seed = 0, log2spp = 0, nBase4Digits = 0;
    uint64_t morton = 0;
    int dimension = 0;

    // samplers.h:228-239.  Log2Int(int) = index of the highest set bit; RoundUpPow2 (util/math.h:531-539)
    void configure(int spp, int resx, int resy, int randomize_, int seed_) {
        randomize = randomize_;
        seed = seed_;
        log2spp = 31 - __builtin_clz((unsigned)std::max(spp, 1));
        int m = std::max(resx, resy);
        int res = 1;
        while (res < m) res <<= 1;
        int log4spp = (log2spp + 1) / 2;
        nBase4Digits = (31 - __builtin_clz((unsigned)res)) + log4spp;
    }
    void start_pixel_sample(int px, int py, int index, int dim = 0) {
        dimension = dim;
        morton = (encode_morton2((uint32_t)px, (uint32_t)py) << log2spp) | (uint64_t)(int64_t)index;
    }
    uint64_t sample_index() const {
        static const uint8_t perm[24][4] = {
            {0, 1, 2, 3}, {0, 1, 3, 2}, {0, 2, 1, 3}, {0, 2, 3, 1}, {0, 3, 2, 1}, {0, 3, 1, 2}, {1, 0, 2, 3}, {1, 0, 3, 2},
            {1, 2, 0, 3}, {1, 2, 3, 0}, {1, 3, 2, 0}, {1, 3, 0, 2}, {2, 1, 0, 3}, {2, 1, 3, 0}, {2, 0, 1, 3}, {2, 0, 3, 1},
            {2, 3, 0, 1}, {2, 3, 1, 0}, {3, 1, 2, 0}, {3, 1, 0, 2}, {3, 2, 1, 0}, {3, 2, 0, 1}, {3, 0, 2, 1}, {3, 0, 1, 2}};
        uint64_t idx = 0;
        bool pow2 = log2spp & 1;
        int last = pow2 ? 1 : 0;
        // 0x55555555u * dimension: unsigned 32-bit product (int converted to unsigned), then widened
        uint64_t dmix = uint64_t(uint32_t(0x55555555u * (uint32_t)dimension));
        for (int i = nBase4Digits - 1; i >= last; --i) {
            int shift = 2 * i - (pow2 ? 1 : 0);
            int digit = int((morton >> shift) & 3);
            uint64_t higher = morton >> (shift + 2);
            int p = int((mix_bits(higher ^ dmix) >> 24) % 24);
            idx |= uint64_t(perm[p][digit]) << shift;
        }
        if (pow2) {
            int digit = int(morton & 1);
            idx |= uint64_t(digit ^ int(mix_bits((morton >> 1) ^ dmix) & 1));
        }
        return idx;
    }
    float get1d() {
        uint64_t si = sample_index();
        ++dimension;
        uint32_t h = (uint32_t)hash_2i32(dimension, seed);
        return sobol_sample((int64_t)si, 0, randomize, h);
    }
    void get2d(float *a, float *b) {
        uint64_t si = sample_index();
        dimension += 2;
        uint64_t bits = hash_2i32(dimension, seed);
        *a = sobol_sample((int64_t)si, 0, randomize, uint32_t(bits));
        *b = sobol_sample((int64_t)si, 1, randomize, uint32_t(bits >> 32));
    }
};

// The sampler the integrator draws from: "independent" (samplers.h:442-476) or "zsobol".
enum SamplerKind { SAMPLER_INDEPENDENT = 0, SAMPLER_ZSOBOL = 1 };
struct PathSampler {
    int kind = SAMPLER_INDEPENDENT;
    IndependentSampler ind;
    ZSobolSampler zs;
    void start_pixel_sample(int px, int py, int index) {
        if (kind == SAMPLER_ZSOBOL) zs.start_pixel_sample(px, py, index);
        else ind.start_pixel_sample(px, py, index);
    }
    float get1d() { return kind == SAMPLER_ZSOBOL ? zs.get1d() : ind.get1d(); }
    void get2d(float *a, float *b) {
        if (kind == SAMPLER_ZSOBOL) zs.get2d(a, b);
        else { *a = ind.get1d(); *b = ind.get1d(); }  // IndependentSampler::Get2D, samplers.h:466
    }
};

}  // namespace avro
