// oracle/restate.hpp — TEST INFRASTRUCTURE ONLY.
//
// CPU restatement of the reference's volumetric hot path (tsvdh/AcceleratedVolRenderer, a pbrt-v4
// fork).  Plain C++17, no dependencies, built with `g++ -O2 -ffp-contract=off` (the reference's
// deterministic configuration, SURVEY.md §7 hard part 1).  It is the checker the CUDA path is
// compared with; the product (libavr_b200.so) never includes, links or calls anything here.
//
// Pinning: every primitive below is checked bit-for-bit against the reference's OWN templates
// compiled from /root/reference (oracle/refprobe.cpp -> oracle/_ref/libavr_refprobe.so: real
// GridMedium/RGBGridMedium ctors, SampleT_maj<>, DDAMajorantIterator, SampledGrid, RNG, Hash, FastExp,
// SampleDiscrete, HG, Transform, Interval) and against the golden fixtures generated from it
// (tests/golden/, tests/golden/make_golden.py).  The VolPath Li/SampleLd loop as a whole is
// pinned by tolerance against the reference's own VolPathIntegrator (see restate_volpath.hpp);
// camera, film, samplers, graph containers / files / ConnectToGraph bit for bit against the
// reference's classes (refprobe_volpath.cpp, refprobe_graphint.cpp).  Each function cites the
// reference file:line it follows.
#pragma once
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <limits>
#include <vector>

namespace avro {

static constexpr int NSPEC = 471;        // util/spectrum.h:30-33 (360..830 nm)
static constexpr float kInf = std::numeric_limits<float>::infinity();
static constexpr float kOneMinusEps = 0x1.fffffep-1f;  // util/float.h OneMinusEpsilon
static constexpr float kMachEps = 0x1p-24f;            // std::numeric_limits<float>::epsilon()*0.5
static constexpr float kPi = 3.14159265358979323846f;
static constexpr float kInv4Pi = 0.07957747154594766788f;
static constexpr float kShadowEps = 0.0001f;  // util/pbrt.h ShadowEpsilon

inline constexpr float gamma_n(int n) { return (n * kMachEps) / (1 - n * kMachEps); }  // float.h:195

struct V3 {
    float x, y, z;
    float operator[](int i) const { return i == 0 ? x : (i == 1 ? y : z); }
    float &operator[](int i) { return i == 0 ? x : (i == 1 ? y : z); }
};
inline V3 operator+(V3 a, V3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
inline V3 operator-(V3 a, V3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
inline V3 operator*(V3 a, float s) { return {a.x * s, a.y * s, a.z * s}; }
inline V3 operator*(float s, V3 a) { return {a.x * s, a.y * s, a.z * s}; }
inline V3 operator/(V3 a, float s) { return {a.x / s, a.y / s, a.z / s}; }  // Tuple3::operator/ divides
inline V3 operator-(V3 a) { return {-a.x, -a.y, -a.z}; }
inline float sqr(float x) { return x * x; }
// util/vecmath.h Dot/LengthSquared: FMA(a.x, b.x, SumOfProducts(a.y, b.y, a.z, b.z))? -> see dot()
inline float fbits(uint32_t u) { float f; std::memcpy(&f, &u, 4); return f; }
inline uint32_t bitsf(float f) { uint32_t u; std::memcpy(&u, &f, 4); return u; }

// util/math.h DifferenceOfProducts / SumOfProducts (used by Dot / LengthSquared, vecmath.h)
inline float diff_of_products(float a, float b, float c, float d) {
    float cd = c * d;
    float dop = std::fma(a, b, -cd);
    float err = std::fma(-c, d, cd);
    return dop + err;
}
inline float sum_of_products(float a, float b, float c, float d) {
    float cd = c * d;
    float sop = std::fma(a, b, cd);
    float err = std::fma(c, d, -cd);
    return sop + err;
}

struct Spec4 {
    float v[4];
    float operator[](int i) const { return v[i]; }
    float &operator[](int i) { return v[i]; }
};
inline Spec4 spec(float c) { return {{c, c, c, c}}; }
#define AVRO_SPEC_OP(op)                                                                  \
    inline Spec4 operator op(Spec4 a, Spec4 b) {                                          \
        Spec4 r;                                                                          \
        for (int i = 0; i < 4; ++i) r.v[i] = a.v[i] op b.v[i];                            \
        return r;                                                                         \
    }
AVRO_SPEC_OP(+)
AVRO_SPEC_OP(-)
AVRO_SPEC_OP(*)
AVRO_SPEC_OP(/)
#undef AVRO_SPEC_OP
inline Spec4 operator*(Spec4 a, float s) { Spec4 r; for (int i = 0; i < 4; ++i) r.v[i] = a.v[i] * s; return r; }
inline Spec4 operator*(float s, Spec4 a) { return a * s; }
inline Spec4 operator/(Spec4 a, float s) { Spec4 r; for (int i = 0; i < 4; ++i) r.v[i] = a.v[i] / s; return r; }
inline bool nonzero(Spec4 a) { return a.v[0] != 0 || a.v[1] != 0 || a.v[2] != 0 || a.v[3] != 0; }  // operator bool
inline float average(Spec4 a) { float s = a.v[0]; for (int i = 1; i < 4; ++i) s += a.v[i]; return s / 4; }
inline float max_component(Spec4 a) { float m = a.v[0]; for (int i = 1; i < 4; ++i) m = std::max(m, a.v[i]); return m; }
inline Spec4 clamp_zero(Spec4 a) { Spec4 r; for (int i = 0; i < 4; ++i) r.v[i] = std::max<float>(0, a.v[i]); return r; }
inline Spec4 safe_div(Spec4 a, Spec4 b) { Spec4 r; for (int i = 0; i < 4; ++i) r.v[i] = (b.v[i] != 0) ? a.v[i] / b.v[i] : 0.f; return r; }
inline bool has_nans(Spec4 a) { for (int i = 0; i < 4; ++i) if (std::isnan(a.v[i])) return true; return false; }

// ---- util/float.h:160-193 NextFloatUp/Down, :199-250 rounded ops (CPU branch) -----------------
inline float next_up(float v) {
    if (std::isinf(v) && v > 0.f) return v;
    if (v == -0.f) v = 0.f;
    uint32_t ui = bitsf(v);
    if (v >= 0) ++ui; else --ui;
    return fbits(ui);
}
inline float next_down(float v) {
    if (std::isinf(v) && v < 0.) return v;
    if (v == 0.f) v = -0.f;
    uint32_t ui = bitsf(v);
    if (v > 0) --ui; else ++ui;
    return fbits(ui);
}

// ---- util/hash.h:19-61 MurmurHash64A, :69-76 MixBits, :95-103 Hash(args...) -------------------
inline uint64_t murmur64a(const unsigned char *key, size_t len, uint64_t seed) {
    const uint64_t m = 0xc6a4a7935bd1e995ull;
    const int r = 47;
    uint64_t h = seed ^ (len * m);
    const unsigned char *end = key + 8 * (len / 8);
    while (key != end) {
        uint64_t k;
        std::memcpy(&k, key, 8);
        key += 8;
        k *= m; k ^= k >> r; k *= m;
        h ^= k; h *= m;
    }
    int tail = int(len & 7);
    if (tail) {
        for (int i = tail - 1; i >= 0; --i) h ^= uint64_t(key[i]) << (8 * i);
        h *= m;
    }
    h ^= h >> r; h *= m; h ^= h >> r;
    return h;
}
inline uint64_t mix_bits(uint64_t v) {
    v ^= (v >> 31); v *= 0x7fb5d329728ea185ull;
    v ^= (v >> 27); v *= 0x81dadef4bc2dd44dull;
    v ^= (v >> 33);
    return v;
}
inline uint64_t hash_float(float f) { return murmur64a((const unsigned char *)&f, 4, 0); }
inline uint64_t hash_v3(V3 p) { float b[3] = {p.x, p.y, p.z}; return murmur64a((const unsigned char *)b, 12, 0); }
inline uint64_t hash_pixel_seed(int px, int py, int seed) { int b[3] = {px, py, seed}; return murmur64a((const unsigned char *)b, 12, 0); }

// ---- util/rng.h:30-172 PCG32 ---------------------------------------------------------------------
struct RNG {
    uint64_t state = 0x853c49e6748fea9bULL, inc = 0xda3e39cb94b95bdbULL;
    static constexpr uint64_t MULT = 0x5851f42d4c957f2dULL;
    RNG() = default;
    RNG(uint64_t seq, uint64_t seed) { set_sequence(seq, seed); }
    uint32_t u32() {
        uint64_t old = state;
        state = old * MULT + inc;
        uint32_t xs = (uint32_t)(((old >> 18u) ^ old) >> 27u);
        uint32_t rot = (uint32_t)(old >> 59u);
        return (xs >> rot) | (xs << ((~rot + 1u) & 31));
    }
    void set_sequence(uint64_t seq, uint64_t seed) {
        state = 0u; inc = (seq << 1u) | 1u;
        u32(); state += seed; u32();
    }
    void set_sequence(uint64_t seq) { set_sequence(seq, mix_bits(seq)); }
    float uniform() { return std::min<float>(kOneMinusEps, u32() * 0x1p-32f); }
    void advance(int64_t idelta) {
        uint64_t curMult = MULT, curPlus = inc, accMult = 1u, accPlus = 0u, delta = (uint64_t)idelta;
        while (delta > 0) {
            if (delta & 1) { accMult *= curMult; accPlus = accPlus * curMult + curPlus; }
            curPlus = (curMult + 1) * curPlus;
            curMult *= curMult;
            delta /= 2;
        }
        state = accMult * state + accPlus;
    }
};

// samplers.h:442-476 IndependentSampler
struct IndependentSampler {
    int seed = 0;
    RNG rng;
    void start_pixel_sample(int px, int py, int sampleIndex, int dim = 0) {
        rng.set_sequence(hash_pixel_seed(px, py, seed));
        rng.advance(sampleIndex * 65536ull + dim);
    }
    float get1d() { return rng.uniform(); }
};

// ---- util/math.h:450-475 FastExp (CPU branch), :329-337 EvaluatePolynomial with std::fma -----
inline float fast_exp(float x) {
    float xp = x * 1.442695041f;
    float fxp = std::floor(xp), f = xp - fxp;
    int i = (int)fxp;
    float twoToF = std::fma(f, std::fma(f, std::fma(f, 0.0781455737f, 0.226173572f), 0.695556856f), 1.f);
    int exponent = int(bitsf(twoToF) >> 23) - 127 + i;
    if (exponent < -126) return 0;
    if (exponent > 127) return kInf;
    uint32_t bits = bitsf(twoToF);
    bits &= 0b10000000011111111111111111111111u;
    bits |= (uint32_t)(exponent + 127) << 23;
    return fbits(bits);
}
inline Spec4 fast_exp(Spec4 s) { Spec4 r; for (int i = 0; i < 4; ++i) r.v[i] = fast_exp(s.v[i]); return r; }  // spectrum.h:695

inline float lerp(float x, float a, float b) { return (1 - x) * a + x * b; }  // util/math.h:210
inline float sample_exponential(float u, float a) { return -std::log(1 - u) / a; }  // sampling.h:222
inline float safe_sqrt(float x) { return std::sqrt(std::max(0.f, x)); }
template <typename T> inline T clampv(T v, T lo, T hi) { return v < lo ? lo : (v > hi ? hi : v); }

// util/sampling.h:79-110 SampleDiscrete
inline int sample_discrete(const float *w, int n, float u) {
    float sum = 0;
    for (int i = 0; i < n; ++i) sum += w[i];
    float up = u * sum;
    if (up == sum) up = next_down(up);
    int offset = 0;
    float acc = 0;
    while (acc + w[offset] <= up) acc += w[offset++];
    return offset;
}

// ---- vecmath.h: Dot/Length/Normalize/CoordinateSystem/SphericalDirection ----------------------
inline float dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }  // vecmath.h:964-967
inline float length_squared(V3 v) { return sqr(v.x) + sqr(v.y) + sqr(v.z); }
inline float length(V3 v) { return std::sqrt(length_squared(v)); }
inline V3 normalize(V3 v) { return v / length(v); }
inline V3 vabs(V3 v) { return {std::abs(v.x), std::abs(v.y), std::abs(v.z)}; }
inline void coordinate_system(V3 v1, V3 *v2, V3 *v3) {  // vecmath.h:1007-1013
    float sign = std::copysign(1.f, v1.z);
    float a = -1 / (sign + v1.z);
    float b = v1.x * v1.y * a;
    *v2 = {1 + sign * sqr(v1.x) * a, sign * b, -sign * v1.x};
    *v3 = {b, sign + sqr(v1.y) * a, -v1.y};
}

// util/scattering.h:49-58, util/sampling.cpp:348-374
inline float henyey_greenstein(float cosTheta, float g) {
    g = clampv(g, -.99f, .99f);
    float denom = 1 + sqr(g) + 2 * g * cosTheta;
    return kInv4Pi * (1 - sqr(g)) / (denom * safe_sqrt(denom));
}
inline V3 sample_hg(V3 wo, float g, float u0, float u1, float *pdf) {
    g = clampv(g, -.99f, .99f);
    float cosTheta;
    if (std::abs(g) < 1e-3f)
        cosTheta = 1 - 2 * u0;
    else
        cosTheta = -1 / (2 * g) * (1 + sqr(g) - sqr((1 - sqr(g)) / (1 + g - 2 * g * u0)));
    float sinTheta = safe_sqrt(1 - sqr(cosTheta));
    float phi = 2 * kPi * u1;
    V3 fx, fy;
    coordinate_system(wo, &fx, &fy);
    float st = clampv(sinTheta, -1.f, 1.f), ct = clampv(cosTheta, -1.f, 1.f);
    V3 l = {st * std::cos(phi), st * std::sin(phi), ct};
    V3 wi = l.x * fx + l.y * fy + l.z * wo;
    if (pdf) *pdf = henyey_greenstein(cosTheta, g);
    return wi;
}

}  // namespace avro
