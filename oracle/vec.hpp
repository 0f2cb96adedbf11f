// ORACLE — TEST INFRASTRUCTURE ONLY.  Not part of the product; only tests/, __graft_entry__.smoke()
// and bench.py's cpu_baseline / --impl reference legs may use anything under oracle/.
//
// CPU restatement of the reference's arithmetic substrate:
//   * glam 0.21.3 `Vec3A` / `Vec2` (Cargo.lock:114-115; NOT vendored under /root/reference — the
//     published SSE2 semantics are restated here): lane-wise IEEE binary32 +,-,*,/;
//     dot = (x*x + y*y) + z*z (mul, then two scalar adds, no FMA); cross = separately rounded
//     products then one subtraction; length = sqrt(dot); normalize = v / sqrt(dot) (lane division);
//     lerp = a + (b - a) * s.
//   * rand 0.8.5 `SmallRng` = Xoshiro256PlusPlus (Cargo.lock:278-279; not vendored either):
//     seed_from_u64 = SplitMix64, next_u32 = upper 32 bits, gen::<f32>() = (u32 >> 8) * 2^-24,
//     gen_range / shuffle = widening-multiply rejection ("zone") sampling.
// PARITY UNPINNED at the glam / rand / libm / jpeg-decoder boundaries: the reference ships no
// tests or golden vectors; the pins are the hand-derived KATs K1-K12.
//
// Build: g++ -std=c++17 -O2 -ffp-contract=off -fno-fast-math (baseline x86-64, no -march=native),
// so every float operation is individually rounded exactly like rustc's default target.
#pragma once
#include <cmath>
#include <cstdint>
#include <cstring>

namespace orc {

static inline uint32_t f2u(float f) { uint32_t u; std::memcpy(&u, &f, 4); return u; }
static inline float u2f(uint32_t u) { float f; std::memcpy(&f, &u, 4); return f; }
static inline int32_t f2i(float f) { int32_t u; std::memcpy(&u, &f, 4); return u; }
static inline float i2f(int32_t u) { float f; std::memcpy(&f, &u, 4); return f; }

// Rust `f32 as i32` / `as u32` / `as isize`: saturating, NaN -> 0.
static inline int32_t rust_f32_as_i32(float f) {
    if (f != f) return 0;
    if (f >= 2147483648.0f) return INT32_MAX;
    if (f <= -2147483648.0f) return INT32_MIN;
    return (int32_t)f;
}
static inline uint32_t rust_f32_as_u32(float f) {
    if (f != f) return 0;
    if (f >= 4294967296.0f) return UINT32_MAX;
    if (f <= 0.0f) return 0;
    return (uint32_t)f;
}
static inline int64_t rust_f32_as_isize(float f) {
    if (f != f) return 0;
    if (f >= 9223372036854775808.0f) return INT64_MAX;
    if (f <= -9223372036854775808.0f) return INT64_MIN;
    return (int64_t)f;
}
static inline uint8_t rust_f32_as_u8(float f) {
    if (f != f) return 0;
    if (f >= 255.0f) return 255;
    if (f <= 0.0f) return 0;
    return (uint8_t)f;
}
// Rust f32::clamp: NaN stays NaN.
static inline float rust_clamp(float x, float lo, float hi) {
    if (x < lo) x = lo;
    if (x > hi) x = hi;
    return x;
}
// f32::powi(n) lowers to compiler-rt __powisf2 (square-and-multiply).
static inline float powi(float a, int b) {
    const bool recip = b < 0;
    float r = 1.0f;
    while (true) {
        if (b & 1) r *= a;
        b /= 2;
        if (b == 0) break;
        a *= a;
    }
    return recip ? 1.0f / r : r;
}

struct Vec2 {
    float x = 0, y = 0;
};

struct Vec3A {
    float x = 0, y = 0, z = 0;
    Vec3A() = default;
    Vec3A(float x_, float y_, float z_) : x(x_), y(y_), z(z_) {}
    static Vec3A splat(float v) { return Vec3A(v, v, v); }
    float operator[](int i) const { return i == 0 ? x : (i == 1 ? y : z); }
    float& operator[](int i) { return i == 0 ? x : (i == 1 ? y : z); }
    float dot(const Vec3A& r) const { return (x * r.x + y * r.y) + z * r.z; }
    float length_squared() const { return dot(*this); }
    float length() const { return std::sqrt(dot(*this)); }
    Vec3A cross(const Vec3A& r) const {
        // glam sse2: (lhs.zxy * rhs - lhs * rhs.zxy).zxy
        return Vec3A(y * r.z - r.y * z, z * r.x - r.z * x, x * r.y - r.x * y);
    }
    Vec3A normalize() const {
        float l = length();
        return Vec3A(x / l, y / l, z / l);
    }
    Vec3A lerp(const Vec3A& rhs, float s) const;
    Vec3A powf(float e) const { return Vec3A(std::pow(x, e), std::pow(y, e), std::pow(z, e)); }
    Vec3A floor() const { return Vec3A(std::floor(x), std::floor(y), std::floor(z)); }
};
static inline Vec3A operator+(Vec3A a, Vec3A b) { return Vec3A(a.x + b.x, a.y + b.y, a.z + b.z); }
static inline Vec3A operator-(Vec3A a, Vec3A b) { return Vec3A(a.x - b.x, a.y - b.y, a.z - b.z); }
static inline Vec3A operator*(Vec3A a, Vec3A b) { return Vec3A(a.x * b.x, a.y * b.y, a.z * b.z); }
static inline Vec3A operator/(Vec3A a, Vec3A b) { return Vec3A(a.x / b.x, a.y / b.y, a.z / b.z); }
static inline Vec3A operator*(Vec3A a, float s) { return Vec3A(a.x * s, a.y * s, a.z * s); }
static inline Vec3A operator*(float s, Vec3A a) { return Vec3A(s * a.x, s * a.y, s * a.z); }
static inline Vec3A operator/(Vec3A a, float s) { return Vec3A(a.x / s, a.y / s, a.z / s); }
static inline Vec3A operator+(Vec3A a, float s) { return Vec3A(a.x + s, a.y + s, a.z + s); }
static inline Vec3A operator-(Vec3A a, float s) { return Vec3A(a.x - s, a.y - s, a.z - s); }
static inline Vec3A operator-(float s, Vec3A a) { return Vec3A(s - a.x, s - a.y, s - a.z); }
static inline Vec3A operator-(Vec3A a) { return Vec3A(-a.x, -a.y, -a.z); }
static inline Vec3A& operator+=(Vec3A& a, Vec3A b) { a = a + b; return a; }
static inline Vec3A& operator*=(Vec3A& a, float s) { a = a * s; return a; }
static inline Vec3A& operator/=(Vec3A& a, float s) { a = a / s; return a; }
inline Vec3A Vec3A::lerp(const Vec3A& rhs, float s) const { return *this + ((rhs - *this) * s); }

// ---------------------------------------------------------------------------------------------
// rand 0.8.5 SmallRng (64-bit targets) — SURVEY.md Appendix B
// ---------------------------------------------------------------------------------------------
struct SmallRng {
    uint64_t s[4];
    // Test hook (oracle addition): when `inject` is non-null, gen_f32() returns the next value of
    // the injected sequence instead of the generator's, so that a scatter() call can be replayed
    // with the exact random numbers a GPU test kernel consumed.
    const float* inject = nullptr;
    size_t inject_n = 0, inject_i = 0;

    static inline uint64_t rotl(uint64_t x, int k) { return (x << k) | (x >> (64 - k)); }
    static SmallRng seed_from_u64(uint64_t state) {
        SmallRng r;
        for (int i = 0; i < 4; i++) {
            state += 0x9e3779b97f4a7c15ULL;
            uint64_t z = state;
            z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ULL;
            z = (z ^ (z >> 27)) * 0x94d049bb133111ebULL;
            z = z ^ (z >> 31);
            r.s[i] = z;
        }
        if ((r.s[0] | r.s[1] | r.s[2] | r.s[3]) == 0) return seed_from_u64(0);
        return r;
    }
    uint64_t next_u64() {
        const uint64_t result = rotl(s[0] + s[3], 23) + s[0];
        const uint64_t t = s[1] << 17;
        s[2] ^= s[0];
        s[3] ^= s[1];
        s[1] ^= s[2];
        s[0] ^= s[3];
        s[2] ^= t;
        s[3] = rotl(s[3], 45);
        return result;
    }
    uint32_t next_u32() { return (uint32_t)(next_u64() >> 32); }
    float gen_f32() {
        if (inject) {
            float v = inject_i < inject_n ? inject[inject_i] : 0.5f;
            inject_i++;
            return v;
        }
        return (float)(next_u32() >> 8) * (1.0f / 16777216.0f);
    }
    // gen_range(0..range) for usize (u64 path)
    uint64_t gen_range_u64(uint64_t range) {
        const uint64_t zone = (range << __builtin_clzll(range)) - 1;
        while (true) {
            uint64_t v = next_u64();
            unsigned __int128 m = (unsigned __int128)v * range;
            uint64_t hi = (uint64_t)(m >> 64), lo = (uint64_t)m;
            if (lo <= zone) return hi;
        }
    }
    // gen_range(0..range) for u32 (used by SliceRandom::shuffle's gen_index)
    uint32_t gen_range_u32(uint32_t range) {
        const uint32_t zone = (range << __builtin_clz(range)) - 1;
        while (true) {
            uint32_t v = next_u32();
            uint64_t m = (uint64_t)v * range;
            uint32_t hi = (uint32_t)(m >> 32), lo = (uint32_t)m;
            if (lo <= zone) return hi;
        }
    }
    template <class T>
    void shuffle(T* a, size_t len) {
        for (size_t i = len - 1; i >= 1; i--) {
            size_t j = gen_range_u32((uint32_t)(i + 1));
            T tmp = a[i];
            a[i] = a[j];
            a[j] = tmp;
        }
    }
};

// src/lib.rs:7-9 — thread-local generator, seed 1995 until reseeded
inline SmallRng& RNG() {
    static thread_local SmallRng rng = SmallRng::seed_from_u64(1995);
    return rng;
}

}  // namespace orc
