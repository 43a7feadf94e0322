// avr_math.cuh — device-side arithmetic of the volumetric hot path (sm_100a).
//
// The library is compiled with -fmad=false: a plain C expression is never contracted, so the
// EXACT code paths reproduce the reference's CPU rounding (built -ffp-contract=off) operation by
// operation; every FMA below is deliberate (__fmaf_rn), either because the reference itself calls
// std::fma (FastExp / EvaluatePolynomial, util/math.h:329-337,450-475) or because the path is the
// FAST variant (which mirrors the reference's own GPU branch, util/math.h:451-452, and is allowed to
// re-associate).
#pragma once
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace avr {

#define AVR_DEV __device__ __forceinline__
#define AVR_HD __host__ __device__ __forceinline__

// FAST-mode approximations with flush-to-zero (no denormal fix-up code around the MUFU)
__device__ __forceinline__ float ftz_rcp(float x) { float r; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
__device__ __forceinline__ float ftz_lg2(float x) { float r; asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
__device__ __forceinline__ float ftz_ex2(float x) { float r; asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
__device__ __forceinline__ float ftz_div(float a, float b) { return a * ftz_rcp(b); }

constexpr int kNSpec = 471;
constexpr float kOneMinusEps = 0x1.fffffep-1f;
constexpr float kMachEps = 0x1p-24f;
constexpr float kGamma3 = (3 * kMachEps) / (1 - 3 * kMachEps);  // util/float.h:195-197
constexpr float kPi = 3.14159265358979323846f;
constexpr float kInv4Pi = 0.07957747154594766788f;
constexpr float kShadowEps = 0.0001f;
constexpr float kFltMax = 3.402823466e+38f;

struct V3 {
    float x, y, z;
};
AVR_HD V3 v3(float x, float y, float z) { return V3{x, y, z}; }
AVR_HD V3 operator+(V3 a, V3 b) { return V3{a.x + b.x, a.y + b.y, a.z + b.z}; }
AVR_HD V3 operator-(V3 a, V3 b) { return V3{a.x - b.x, a.y - b.y, a.z - b.z}; }
AVR_HD V3 operator*(V3 a, float s) { return V3{a.x * s, a.y * s, a.z * s}; }
AVR_HD V3 operator/(V3 a, float s) { return V3{a.x / s, a.y / s, a.z / s}; }
AVR_HD V3 operator-(V3 a) { return V3{-a.x, -a.y, -a.z}; }
AVR_HD float dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }  // vecmath.h:964-967
AVR_HD float sqr(float x) { return x * x; }
AVR_HD float comp(V3 v, int i) { return i == 0 ? v.x : (i == 1 ? v.y : v.z); }

// ---- N-channel spectrum (N = 4: SampledSpectrum; N = 1: achromatic media, all channels equal) ---
template <int N>
struct Spec {
    float v[N];
    AVR_HD float operator[](int i) const { return v[i]; }
    AVR_HD float &operator[](int i) { return v[i]; }
};
template <int N>
AVR_HD Spec<N> sconst(float c) {
    Spec<N> r;
#pragma unroll
    for (int i = 0; i < N; ++i) r.v[i] = c;
    return r;
}
#define AVR_SPEC_OP(op)                                              \
    template <int N>                                                 \
    AVR_HD Spec<N> operator op(Spec<N> a, Spec<N> b) {               \
        Spec<N> r;                                                   \
        _Pragma("unroll") for (int i = 0; i < N; ++i) r.v[i] = a.v[i] op b.v[i]; \
        return r;                                                    \
    }                                                                \
    template <int N>                                                 \
    AVR_HD Spec<N> operator op(Spec<N> a, float s) {                 \
        Spec<N> r;                                                   \
        _Pragma("unroll") for (int i = 0; i < N; ++i) r.v[i] = a.v[i] op s; \
        return r;                                                    \
    }
AVR_SPEC_OP(+)
AVR_SPEC_OP(-)
AVR_SPEC_OP(*)
AVR_SPEC_OP(/)
#undef AVR_SPEC_OP
template <int N>
AVR_HD bool nonzero(Spec<N> a) {
    bool r = false;
#pragma unroll
    for (int i = 0; i < N; ++i) r = r || (a.v[i] != 0);
    return r;
}
template <int N>
AVR_HD float average(Spec<N> a) {  // SampledSpectrum::Average: sum left to right, / NSpectrumSamples
    if (N == 1) return (((a.v[0] + a.v[0]) + a.v[0]) + a.v[0]) / 4;
    float s = a.v[0];
#pragma unroll
    for (int i = 1; i < N; ++i) s += a.v[i];
    return s / N;
}
template <int N>
AVR_HD float max_component(Spec<N> a) {
    float m = a.v[0];
#pragma unroll
    for (int i = 1; i < N; ++i) m = fmaxf(m, a.v[i]);
    return m;
}
template <int N>
AVR_HD Spec<N> clamp_zero(Spec<N> a) {
    Spec<N> r;
#pragma unroll
    for (int i = 0; i < N; ++i) r.v[i] = fmaxf(0.f, a.v[i]);
    return r;
}

// ---- util/float.h NextFloatUp/Down (CPU branch of the rounded interval ops) ------------------
AVR_DEV float next_up(float v) {
    if (isinf(v) && v > 0.f) return v;
    if (v == -0.f) v = 0.f;
    uint32_t ui = __float_as_uint(v);
    if (v >= 0) ++ui; else --ui;
    return __uint_as_float(ui);
}
AVR_DEV float next_down(float v) {
    if (isinf(v) && v < 0.f) return v;
    if (v == 0.f) v = -0.f;
    uint32_t ui = __float_as_uint(v);
    if (v > 0) --ui; else ++ui;
    return __uint_as_float(ui);
}

// ---- util/hash.h MurmurHash64A over 4 / 12 bytes, MixBits ---------------------------------------
AVR_HD uint64_t mix_bits(uint64_t v) {
    v ^= (v >> 31); v *= 0x7fb5d329728ea185ull;
    v ^= (v >> 27); v *= 0x81dadef4bc2dd44dull;
    v ^= (v >> 33);
    return v;
}
AVR_HD uint64_t murmur_finish(uint64_t h) {
    const uint64_t m = 0xc6a4a7935bd1e995ull;
    h ^= h >> 47; h *= m; h ^= h >> 47;
    return h;
}
// Hash(float): len = 4, tail bytes only
AVR_HD uint64_t hash_u32(uint32_t bits) {
    const uint64_t m = 0xc6a4a7935bd1e995ull;
    uint64_t h = 0 ^ (4ull * m);
    h ^= uint64_t(bits);
    h *= m;
    return murmur_finish(h);
}
// Hash of 12 bytes (Point3f / Vector3f / (Point2i, int)): one 8-byte block + 4 tail bytes
AVR_HD uint64_t hash_3u32(uint32_t a, uint32_t b, uint32_t c) {
    const uint64_t m = 0xc6a4a7935bd1e995ull;
    uint64_t h = 0 ^ (12ull * m);
    uint64_t k = uint64_t(a) | (uint64_t(b) << 32);
    k *= m; k ^= k >> 47; k *= m;
    h ^= k; h *= m;
    h ^= uint64_t(c);
    h *= m;
    return murmur_finish(h);
}
AVR_DEV uint64_t hash_float(float f) { return hash_u32(__float_as_uint(f)); }
AVR_DEV uint64_t hash_v3(V3 p) {
    return hash_3u32(__float_as_uint(p.x), __float_as_uint(p.y), __float_as_uint(p.z));
}

// ---- util/rng.h PCG32 -------------------------------------------------------------------------------
struct RNG {
    uint64_t state, inc;
    static constexpr uint64_t MULT = 0x5851f42d4c957f2dULL;
    AVR_HD uint32_t u32() {
        uint64_t old = state;
        state = old * MULT + inc;
        uint32_t xs = (uint32_t)(((old >> 18u) ^ old) >> 27u);
        uint32_t rot = (uint32_t)(old >> 59u);
        return (xs >> rot) | (xs << ((~rot + 1u) & 31));
    }
    AVR_HD void set_sequence(uint64_t seq, uint64_t seed) {
        state = 0u; inc = (seq << 1u) | 1u;
        u32(); state += seed; u32();
    }
    AVR_HD void set_sequence(uint64_t seq) { set_sequence(seq, mix_bits(seq)); }
    AVR_HD float uniform() {
        float f = u32() * 0x1p-32f;
        return f < kOneMinusEps ? f : kOneMinusEps;
    }
    AVR_HD void advance(int64_t idelta) {
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


// ---- libm restatements for the EXACT path ---------------------------------------------------------
// The reference calls std::log / std::sin / std::cos on floats, i.e. glibc's logf / sinf / cosf (the
// ARM optimized-routines algorithms: table + double-precision polynomial).  Scatter positions feed
// Hash(p) / Hash(d) RNG seeds (cpu/integrators.cpp:1338), so a 1-ulp libm difference changes a whole
// shadow-ray random stream; CUDA's logf/sinf/cosf (<= 1-2 ulp) are therefore not good enough for
// per-sample parity.  These restatements were checked bit-for-bit against the host libm over ALL
// positive floats (logf) and all floats in [0, 2*pi] (sinf, cosf) — tests/test_libm_restatement.py.
__device__ const double kLogfInvC[16] = {
    0x1.661ec79f8f3bep+0, 0x1.571ed4aaf883dp+0, 0x1.49539f0f010bp+0, 0x1.3c995b0b80385p+0,
    0x1.30d190c8864a5p+0, 0x1.25e227b0b8eap+0, 0x1.1bb4a4a1a343fp+0, 0x1.12358f08ae5bap+0,
    0x1.0953f419900a7p+0, 0x1p+0, 0x1.e608cfd9a47acp-1, 0x1.ca4b31f026aap-1,
    0x1.b2036576afce6p-1, 0x1.9c2d163a1aa2dp-1, 0x1.886e6037841edp-1, 0x1.767dcf5534862p-1};
__device__ const double kLogfLogC[16] = {
    -0x1.57bf7808caadep-2, -0x1.2bef0a7c06ddbp-2, -0x1.01eae7f513a67p-2, -0x1.b31d8a68224e9p-3,
    -0x1.6574f0ac07758p-3, -0x1.1aa2bc79c81p-3, -0x1.a4e76ce8c0e5ep-4, -0x1.1973c5a611cccp-4,
    -0x1.252f438e10c1ep-5, 0x0p+0, 0x1.aa5aa5df25984p-5, 0x1.c5e53aa362eb4p-4,
    0x1.526e57720db08p-3, 0x1.bc2860d22477p-3, 0x1.1058bc8a07ee1p-2, 0x1.4043057b6ee09p-2};
AVR_DEV float glibc_logf(float x) {
    uint32_t ix = __float_as_uint(x);
    if (ix == 0x3f800000u) return 0.f;
    if (ix - 0x00800000u >= 0x7f800000u - 0x00800000u) {
        if (ix * 2 == 0) return __int_as_float(0xff800000);
        if (ix == 0x7f800000u) return x;
        if ((ix & 0x80000000u) || ix * 2 >= 0xff000000u) return __int_as_float(0x7fc00000);
        x *= 0x1p23f;
        ix = __float_as_uint(x);
        ix -= 23u << 23;
    }
    uint32_t tmp = ix - 0x3f330000u;
    int i = (tmp >> 19) % 16;
    int k = (int32_t)tmp >> 23;
    uint32_t iz = ix - (tmp & 0xff800000u);
    double z = (double)__uint_as_float(iz);
    double r = z * kLogfInvC[i] - 1;
    double y0 = kLogfLogC[i] + (double)k * 0x1.62e42fefa39efp-1;
    double r2 = r * r;
    double y = 0x1.5575b0be00b6ap-2 * r + -0x1.ffffef20a4123p-2;
    y = -0x1.00ea348b88334p-2 * r2 + y;
    y = y * r2 + (y0 + r);
    return (float)y;
}
// sinf/cosf for |y| < 120 (the path only evaluates phi = 2*pi*u in [0, 2*pi))
AVR_DEV float glibc_sincos_poly(double x, double x2, bool neg_table, int n) {
    const double sg = neg_table ? -1.0 : 1.0;
    if ((n & 1) == 0) {
        double x3 = x * x2;
        double s1 = 0x1.1107605230bc4p-7 + x2 * -0x1.994eb3774cf24p-13;
        double x7 = x3 * x2;
        double s = x + x3 * -0x1.555545995a603p-3;
        return (float)(s + x7 * s1);
    }
    double x4 = x2 * x2;
    double c2 = (sg * -0x1.6c087e89a359dp-10) + x2 * (sg * 0x1.99343027bf8c3p-16);
    double c1 = (sg * 0x1p0) + x2 * (sg * -0x1.ffffffd0c621cp-2);
    double x6 = x4 * x2;
    double c = c1 + x4 * (sg * 0x1.55553e1068f19p-5);
    return (float)(c + x6 * c2);
}
AVR_DEV uint32_t abstop12(float x) { return (__float_as_uint(x) >> 20) & 0x7ff; }
AVR_DEV float glibc_sincosf_one(float y, bool want_cos) {
    double x = y;
    if (abstop12(y) < abstop12(0x1.921FB6p-1f)) {
        double s = x * x;
        if (abstop12(y) < abstop12(0x1p-12f)) return want_cos ? 1.0f : y;
        return glibc_sincos_poly(x, s, false, want_cos ? 1 : 0);
    }
    double r = x * 0x1.45F306DC9C883p+23;
    int n = ((int32_t)r + 0x800000) >> 24;
    x = x - n * 0x1.921FB54442D18p0;
    // sign[] of table 0 = {1,-1,-1,1}; the sign applies to x before the (odd/even) polynomial
    const int q = n & 3;
    double sgn = (q == 1 || q == 2) ? -1.0 : 1.0;
    bool neg_table = (n & 2) != 0;  // glibc reads sign[] from table 0, then switches polynomial tables
    return glibc_sincos_poly(x * sgn, x * x, neg_table, want_cos ? (n ^ 1) : n);
}
AVR_DEV float glibc_sinf(float y) { return glibc_sincosf_one(y, false); }
AVR_DEV float glibc_cosf(float y) { return glibc_sincosf_one(y, true); }

// ---- FastExp: EXACT = CPU branch util/math.h:453-474 ; FAST = GPU branch (__expf) :451-452 -----
AVR_DEV float fast_exp_exact(float x) {
    float xp = x * 1.442695041f;
    float fxp = floorf(xp), f = xp - fxp;
    int i = (int)fxp;
    float twoToF = __fmaf_rn(f, __fmaf_rn(f, __fmaf_rn(f, 0.0781455737f, 0.226173572f), 0.695556856f), 1.f);
    int exponent = int(__float_as_uint(twoToF) >> 23) - 127 + i;
    if (exponent < -126) return 0;
    if (exponent > 127) return __int_as_float(0x7f800000);
    uint32_t bits = __float_as_uint(twoToF);
    bits &= 0x807fffffu;
    bits |= (uint32_t)(exponent + 127) << 23;
    return __uint_as_float(bits);
}
template <bool EXACT>
AVR_DEV float fast_exp(float x) {
    if (EXACT) return fast_exp_exact(x);
    return ftz_ex2(x * 1.4426950408889634f);
}
template <bool EXACT>
AVR_DEV float sample_exponential(float u, float a) {  // util/sampling.h:222-225
    if (EXACT) return -glibc_logf(1 - u) / a;
    return ftz_lg2(1 - u) * -0.6931471805599453f * ftz_rcp(a);
}
// FAST: -ln(1 - u), the optical depth to the next tentative collision (SampleExponential with a = 1)
AVR_DEV float neg_log1m(float u) { return ftz_lg2(1 - u) * -0.6931471805599453f; }
template <bool EXACT>
AVR_DEV float fdiv(float a, float b) {
    if (EXACT) return a / b;
    return ftz_div(a, b);
}
// EXACT: a / s per channel (IEEE).  FAST: one approximate reciprocal, then multiplies.
template <bool EXACT, int N>
AVR_DEV Spec<N> sdiv(Spec<N> a, float s) {
    if (EXACT) return a / s;
    return a * ftz_rcp(s);
}
AVR_DEV float lerp_exact(float x, float a, float b) { return (1 - x) * a + x * b; }  // util/math.h:210
AVR_DEV float lerp_fast(float x, float a, float b) { return __fmaf_rn(x, b - a, a); }
template <bool EXACT>
AVR_DEV float lerpf(float x, float a, float b) {
    return EXACT ? lerp_exact(x, a, b) : lerp_fast(x, a, b);
}
AVR_DEV float safe_sqrt(float x) { return sqrtf(fmaxf(0.f, x)); }
AVR_DEV float clampf(float v, float lo, float hi) { return v < lo ? lo : (v > hi ? hi : v); }

// util/sampling.h:79-110 SampleDiscrete over {a, s, n}
AVR_DEV int sample_discrete3(float w0, float w1, float w2, float u) {
    float sum = 0;
    sum += w0; sum += w1; sum += w2;
    float up = u * sum;
    if (up == sum) up = next_down(up);
    float acc = 0;
    if (!(acc + w0 <= up)) return 0;
    acc += w0;
    if (!(acc + w1 <= up)) return 1;
    return 2;
}

// util/scattering.h:49-58 HenyeyGreenstein ; util/sampling.cpp:348-374 SampleHenyeyGreenstein
AVR_DEV float henyey_greenstein(float cosTheta, float g) {
    g = clampf(g, -.99f, .99f);
    float denom = 1 + sqr(g) + 2 * g * cosTheta;
    return kInv4Pi * (1 - sqr(g)) / (denom * safe_sqrt(denom));
}
AVR_DEV float henyey_greenstein_fast(float cosTheta, float g) {
    g = clampf(g, -.99f, .99f);
    float denom = fmaxf(__fmaf_rn(2 * g, cosTheta, 1 + g * g), 1e-12f);
    float r = rsqrtf(denom);
    return kInv4Pi * (1 - g * g) * (r * r * r);
}
template <bool EXACT>
AVR_DEV float hg_phase(float cosTheta, float g) {
    return EXACT ? henyey_greenstein(cosTheta, g) : henyey_greenstein_fast(cosTheta, g);
}
template <bool EXACT = true>
AVR_DEV V3 sample_hg(V3 wo, float g, float u0, float u1, float *pdf) {
    g = clampf(g, -.99f, .99f);
    float cosTheta;
    if (fabsf(g) < 1e-3f)
        cosTheta = 1 - 2 * u0;
    else if (EXACT)
        cosTheta = -1 / (2 * g) * (1 + sqr(g) - sqr((1 - sqr(g)) / (1 + g - 2 * g * u0)));
    else
        cosTheta = __fdividef(-1.f, 2 * g) * (1 + g * g - sqr(__fdividef(1 - g * g, 1 + g - 2 * g * u0)));
    float sinTheta = safe_sqrt(1 - sqr(cosTheta));
    float phi = 2 * kPi * u1;
    // CoordinateSystem (vecmath.h:1007-1013), Frame::FromZ(wo).FromLocal(SphericalDirection(...))
    float sign = copysignf(1.f, wo.z);
    float a = EXACT ? -1 / (sign + wo.z) : __fdividef(-1.f, sign + wo.z);
    float b = wo.x * wo.y * a;
    V3 fx = v3(1 + sign * sqr(wo.x) * a, sign * b, -sign * wo.x);
    V3 fy = v3(b, sign + sqr(wo.y) * a, -wo.y);
    float st = clampf(sinTheta, -1.f, 1.f), ct = clampf(cosTheta, -1.f, 1.f);
    float sp, cp;
    if (EXACT) { sp = glibc_sinf(phi); cp = glibc_cosf(phi); }
    else __sincosf(phi, &sp, &cp);
    V3 l = v3(st * cp, st * sp, ct);
    V3 wi = fx * l.x + fy * l.y + wo * l.z;
    *pdf = hg_phase<EXACT>(cosTheta, g);
    return wi;
}

// util/sampling.h:163-171 + util/spectrum.h:334-346: SampledWavelengths::SampleVisible
// One wavelength at a time through NON-inlined helpers: atanhf / coshf are a few hundred instructions each, and
// four inlined copies of both in the once-per-path code evict the tracking loop from the 32 KB instruction cache.
// FAST: the same functions through lg2 / ex2.
template <bool EXACT>
static __device__ __noinline__ float visible_wavelength(float up) {
    float x = 0.85691062f - 1.82750197f * up;
    if (EXACT) return 538 - 138.888889f * atanhf(x);
    return 538 - 138.888889f * (0.5f * 0.6931471805599453f) * ftz_lg2(__fdividef(1 + x, 1 - x));
}
template <bool EXACT>
static __device__ __noinline__ float visible_wavelength_pdf(float l) {
    if (l < 360 || l > 830) return 0.f;
    float y = 0.0072f * (l - 538);
    float c = EXACT ? coshf(y) : 0.5f * (ftz_ex2(y * 1.4426950408889634f) + ftz_ex2(-y * 1.4426950408889634f));
    return 0.0039398042f / sqr(c);
}
template <bool EXACT = true>
AVR_DEV void visible_wavelengths_pdf(const float lambda[4], float pdf[4]) {
#pragma unroll
    for (int i = 0; i < 4; ++i) pdf[i] = visible_wavelength_pdf<EXACT>(lambda[i]);
}
template <bool EXACT = true>
AVR_DEV void sample_visible_wavelengths(float u, float lambda[4], float pdf[4]) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        float up = u + float(i) / 4;
        if (up > 1) up -= 1;
        lambda[i] = visible_wavelength<EXACT>(up);
    }
    visible_wavelengths_pdf<EXACT>(lambda, pdf);
}
// DenselySampledSpectrum::Sample, util/spectrum.h:390-399 (std::lround: half away from zero)
AVR_DEV float sample_dense1(const float *tab, float lambda) {
    int offset = (int)lroundf(lambda) - 360;
    return (offset < 0 || offset >= kNSpec) ? 0.f : __ldg(tab + offset);
}

// util/spectrum.h:69-80 Blackbody (FastExp inside), :500-523 BlackbodySpectrum normalisation
AVR_DEV float blackbody(float lambda, float T) {
    if (T <= 0) return 0;
    const float c = 299792458.f, h = 6.62606957e-34f, kb = 1.3806488e-23f;
    float l = lambda * 1e-9f;
    float l2 = l * l;
    float l5 = l2 * l2 * l;
    return (2 * h * c * c) / (l5 * (fast_exp_exact((h * c) / (l * kb * T)) - 1));
}
// util/color.h:340-360 RGBSigmoidPolynomial over a {scale,c0,c1,c2} voxel
AVR_DEV float sigmoid_s(float x) {
    if (isinf(x)) return x > 0 ? 1.f : 0.f;
    return .5f + x / (2 * sqrtf(1 + sqr(x)));
}
AVR_DEV float rsp_eval(float c0, float c1, float c2, float lambda) {
    return sigmoid_s(__fmaf_rn(lambda, __fmaf_rn(lambda, c0, c1), c2));
}
// FAST: the same sigmoid through one approximate reciprocal square root (x clamped so that x^2 stays finite)
AVR_DEV float rsp_eval_fast(float c0, float c1, float c2, float lambda) {
    float x = __fmaf_rn(lambda, __fmaf_rn(lambda, c0, c1), c2);
    x = fminf(fmaxf(x, -1e18f), 1e18f);
    return __fmaf_rn(0.5f * x, __frsqrt_rn(__fmaf_rn(x, x, 1.f)), 0.5f);
}

}  // namespace avr
