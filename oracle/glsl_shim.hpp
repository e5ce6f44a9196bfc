// TEST INFRASTRUCTURE — not product code.  Only tests/, __graft_entry__.smoke() and bench.py's
// cpu_baseline / --impl reference legs may use anything under oracle/.
//
// A minimal GLSL 4.30 vocabulary for C++ (vec2/vec3/vec4/mat3/mat4, swizzles used by the
// reference shaders, and the handful of built-ins they call) so that the reference's own shader
// bodies (/root/reference/shaders/gau_vert.glsl, gau_frag.glsl) compile unmodified as C++ and can
// be executed on the CPU.  See oracle/ref_build.py for how the translation unit is assembled.
//
// Arithmetic rules fixed here (GLSL leaves them to the implementation):
//   * every operation is an IEEE-754 binary32 operation, no contraction (build with
//     -ffp-contract=off), division and sqrt correctly rounded;
//   * dot products, matrix*vector and matrix*matrix sum their terms left to right:
//     ((t0 + t1) + t2) + t3;
//   * normalize(v) = v / sqrt(dot(v, v));
//   * min(x, y) = (y < x) ? y : x, max(x, y) = (x < y) ? y : x   (GLSL spec wording).
// The C restatement (oracle/gs_oracle.c) and the CUDA kernels follow the same rules, which is what
// makes the per-Gaussian stage comparable bit for bit.
#pragma once
#include <cmath>

namespace glsl {

struct vec2;
struct vec3;
struct vec4;

// --- swizzle proxies (only the ones the shaders use as l-values / r-values) ---------------------
struct swz2 {  // aliases two consecutive floats
    float a, b;
    swz2& operator=(const vec2& v);
    operator vec2() const;
};
struct swz3 {  // aliases three consecutive floats
    float a, b, c;
    swz3& operator=(const vec3& v);
    operator vec3() const;
};

struct vec2 {
    float x, y;
    vec2() : x(0.f), y(0.f) {}
    vec2(float s) : x(s), y(s) {}
    vec2(float x_, float y_) : x(x_), y(y_) {}
};

struct vec3 {
    union {
        struct { float x, y, z; };
        struct { float r, g, b; };
        swz2 xy;
    };
    vec3() : x(0.f), y(0.f), z(0.f) {}
    vec3(float s) : x(s), y(s), z(s) {}
    vec3(double s) : x((float)s), y((float)s), z((float)s) {}
    vec3(float x_, float y_, float z_) : x(x_), y(y_), z(z_) {}
    float& operator[](int i) { return (&x)[i]; }
    const float& operator[](int i) const { return (&x)[i]; }
};

struct vec4 {
    union {
        struct { float x, y, z, w; };
        struct { float r, g, b, a; };
        swz2 xy;
        swz3 xyz;
        swz3 rgb;
    };
    vec4() : x(0.f), y(0.f), z(0.f), w(0.f) {}
    vec4(float s) : x(s), y(s), z(s), w(s) {}
    vec4(float x_, float y_, float z_, float w_) : x(x_), y(y_), z(z_), w(w_) {}
    vec4(const vec3& v, float w_) : x(v.x), y(v.y), z(v.z), w(w_) {}
    vec4(const swz3& v, float w_) : x(v.a), y(v.b), z(v.c), w(w_) {}
    vec4(const vec4& o) : x(o.x), y(o.y), z(o.z), w(o.w) {}
    vec4& operator=(const vec4& o) { x = o.x; y = o.y; z = o.z; w = o.w; return *this; }
    float& operator[](int i) { return (&x)[i]; }
    const float& operator[](int i) const { return (&x)[i]; }
};

inline swz2& swz2::operator=(const vec2& v) { a = v.x; b = v.y; return *this; }
inline swz2::operator vec2() const { return vec2(a, b); }
inline swz3& swz3::operator=(const vec3& v) { a = v.x; b = v.y; c = v.z; return *this; }
inline swz3::operator vec3() const { return vec3(a, b, c); }

// --- component-wise arithmetic --------------------------------------------------------------------
inline vec2 operator+(const vec2& a, const vec2& b) { return vec2(a.x + b.x, a.y + b.y); }
inline vec2 operator-(const vec2& a, const vec2& b) { return vec2(a.x - b.x, a.y - b.y); }
inline vec2 operator*(const vec2& a, const vec2& b) { return vec2(a.x * b.x, a.y * b.y); }
inline vec2 operator/(const vec2& a, const vec2& b) { return vec2(a.x / b.x, a.y / b.y); }
inline vec2 operator*(const vec2& a, float s) { return vec2(a.x * s, a.y * s); }
inline vec2 operator*(float s, const vec2& a) { return vec2(s * a.x, s * a.y); }
inline vec2 operator/(const vec2& a, float s) { return vec2(a.x / s, a.y / s); }

inline vec3 operator+(const vec3& a, const vec3& b) { return vec3(a.x + b.x, a.y + b.y, a.z + b.z); }
inline vec3 operator-(const vec3& a, const vec3& b) { return vec3(a.x - b.x, a.y - b.y, a.z - b.z); }
inline vec3 operator*(const vec3& a, const vec3& b) { return vec3(a.x * b.x, a.y * b.y, a.z * b.z); }
inline vec3 operator/(const vec3& a, const vec3& b) { return vec3(a.x / b.x, a.y / b.y, a.z / b.z); }
inline vec3 operator*(const vec3& a, float s) { return vec3(a.x * s, a.y * s, a.z * s); }
inline vec3 operator*(float s, const vec3& a) { return vec3(s * a.x, s * a.y, s * a.z); }
inline vec3 operator/(const vec3& a, float s) { return vec3(a.x / s, a.y / s, a.z / s); }
inline vec3 operator+(const vec3& a, float s) { return vec3(a.x + s, a.y + s, a.z + s); }
inline vec3& operator+=(vec3& a, float s) { a.x = a.x + s; a.y = a.y + s; a.z = a.z + s; return a; }
inline vec3& operator+=(vec3& a, const vec3& b) { a = a + b; return a; }

inline vec4 operator+(const vec4& a, const vec4& b) { return vec4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w); }
inline vec4 operator-(const vec4& a, const vec4& b) { return vec4(a.x - b.x, a.y - b.y, a.z - b.z, a.w - b.w); }
inline vec4 operator*(const vec4& a, float s) { return vec4(a.x * s, a.y * s, a.z * s, a.w * s); }
inline vec4 operator/(const vec4& a, float s) { return vec4(a.x / s, a.y / s, a.z / s, a.w / s); }

// --- built-ins ------------------------------------------------------------------------------------
inline float min(float x, float y) { return (y < x) ? y : x; }
inline float max(float x, float y) { return (x < y) ? y : x; }
inline float clamp(float x, float lo, float hi) { return min(max(x, lo), hi); }
inline float abs(float x) { return std::fabs(x); }
inline float sqrt(float x) { return std::sqrt(x); }
inline float exp(float x) { return std::exp(x); }
inline vec3 abs(const vec3& v) { return vec3(std::fabs(v.x), std::fabs(v.y), std::fabs(v.z)); }
inline float dot(const vec3& a, const vec3& b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
inline float length(const vec3& v) { return std::sqrt(dot(v, v)); }
inline vec3 normalize(const vec3& v) { return v / length(v); }

struct bvec3 { bool x, y, z; };
inline bvec3 greaterThan(const vec3& a, const vec3& b) { return bvec3{a.x > b.x, a.y > b.y, a.z > b.z}; }
inline bool any(const bvec3& b) { return b.x || b.y || b.z; }

// --- matrices (column-major: m[i] is column i, m[i][j] is column i row j) ---------------------------
struct mat4;
struct mat3 {
    vec3 c[3];
    mat3() {}
    mat3(float d) { c[0] = vec3(d, 0.f, 0.f); c[1] = vec3(0.f, d, 0.f); c[2] = vec3(0.f, 0.f, d); }
    mat3(float a0, float a1, float a2, float a3, float a4, float a5, float a6, float a7, float a8) {
        c[0] = vec3(a0, a1, a2); c[1] = vec3(a3, a4, a5); c[2] = vec3(a6, a7, a8);
    }
    explicit mat3(const mat4& m);
    vec3& operator[](int i) { return c[i]; }
    const vec3& operator[](int i) const { return c[i]; }
};
struct mat4 {
    vec4 c[4];
    mat4() {}
    vec4& operator[](int i) { return c[i]; }
    const vec4& operator[](int i) const { return c[i]; }
};
inline mat3::mat3(const mat4& m) {
    for (int i = 0; i < 3; ++i) c[i] = vec3(m[i].x, m[i].y, m[i].z);
}
inline mat3 transpose(const mat3& m) {
    return mat3(m[0][0], m[1][0], m[2][0], m[0][1], m[1][1], m[2][1], m[0][2], m[1][2], m[2][2]);
}
inline mat3 operator*(const mat3& A, const mat3& B) {
    mat3 R;
    for (int col = 0; col < 3; ++col)
        for (int row = 0; row < 3; ++row)
            R[col][row] = (A[0][row] * B[col][0] + A[1][row] * B[col][1]) + A[2][row] * B[col][2];
    return R;
}
inline vec4 operator*(const mat4& M, const vec4& v) {
    vec4 r;
    for (int row = 0; row < 4; ++row)
        r[row] = ((M[0][row] * v.x + M[1][row] * v.y) + M[2][row] * v.z) + M[3][row] * v.w;
    return r;
}

}  // namespace glsl
