// oracle/cvlike.h -- TEST INFRASTRUCTURE (CPU oracle). Not part of the product path.
//
// The reference hot path (src/Modeler/ProbabilityMapping.cc) does all of its small linear
// algebra through OpenCV cv::Mat expressions. OpenCV is a system dependency that is NOT under
// /root/reference (CMakeLists.txt:37-43 asks for OpenCV 3.0 / 2.4.3; the author's build linked
// 3.3.1, build/CMakeFiles/ORB_SLAM2_Pub.dir/link.txt:1) and its C++ headers are not installed
// here.  This header restates, in plain C++, the arithmetic that OpenCV performs for exactly the
// expression shapes PM.cc uses:
//
//   * cv::gemm for CV_32F goes through cv::hal::gemm32f(..., float alpha, ..., float beta, ...)
//     (OpenCV >= 3.2): alpha and beta are ROUNDED TO FLOAT first, then widened to double again.
//   * cv::gemm "small matrix" path (modules/core/src/matmul.cpp, `flags == 0 && 2 <= len <= 4 &&
//     (len == d_size.width || len == d_size.height)`): float products summed left to right in
//     float, epilogue `(float)(t*alpha + c*beta)` evaluated in double.
//   * cv::gemm general path (GEMMSingleMul<float,double>): products and running sum in double,
//     left to right, `(float)(s*alpha [+ c*beta])`.
//   * cv::invert for 3x3 CV_32F: closed-form cofactors in double, `(float)(cof * (1/det))`.
//   * cv::solve(DECOMP_LU) with a multi-column right-hand side: hal::LU32f -> LUImpl<float>
//     (partial pivoting, all arithmetic in float).
//   * cv::fastAtan2 scalar polynomial (modules/core/src/mathfuncs_core.simd.hpp).
//   * cv::MatExpr lazy-evaluation rules (modules/core/src/matop.cpp) that decide WHICH gemm call a
//     C++ expression becomes (e.g. `A*x*s + t` is one gemm(A,x,s,t,1)); those rules are cited at
//     the call sites in pm_oracle.cc.
//
// Pinning: every function in this file is checked bit-for-bit against cv2 4.13 (the Python wheel
// in the build container) by tests/golden/make_cv_golden.py, which wrote tests/golden/cv_golden.npz;
// tests/test_oracle_cvlike.py replays those vectors without needing cv2.  The MatExpr fusion rules
// are C++ compile-time behaviour and cannot be exercised from Python: they are restated from the
// OpenCV 3.x sources from memory and are the one unpinned part of this file.
#pragma once
#include <cfloat>
#include <cmath>
#include <cstring>

namespace cvl {

struct M33 { float m[9]; };   // row-major
struct V3  { float v[3]; };
struct M44 { float m[16]; };
struct V4  { float v[4]; };

// ---- cv::gemm small path, 3x3 * 3x1 (+ optional C):  d_i = (float)(t_i*alpha + c_i*beta)
static inline V3 gemm_small_33x31(const M33& A, const V3& b, double alpha, const V3* c, double beta) {
    V3 d;
    alpha = (double)(float)alpha;  // hal::gemm32f takes float alpha / beta
    beta = (double)(float)beta;
    for (int i = 0; i < 3; ++i) {
        float t = A.m[3 * i + 0] * b.v[0] + A.m[3 * i + 1] * b.v[1] + A.m[3 * i + 2] * b.v[2];
        float ci = c ? c->v[i] : 0.0f;
        double bi = c ? beta : 0.0;
        d.v[i] = (float)((double)t * alpha + (double)ci * bi);
    }
    return d;
}

// ---- cv::gemm small path, 3x3 * 3x3
static inline M33 gemm_small_33x33(const M33& A, const M33& B, double alpha) {
    M33 D;
    alpha = (double)(float)alpha;
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) {
            float t = A.m[3 * i + 0] * B.m[0 + j] + A.m[3 * i + 1] * B.m[3 + j] + A.m[3 * i + 2] * B.m[6 + j];
            D.m[3 * i + j] = (float)((double)t * alpha + 0.0);
        }
    return D;
}

// ---- cv::gemm small path, 4x4 * 4x1
static inline V4 gemm_small_44x41(const M44& A, const V4& b) {
    V4 d;
    for (int i = 0; i < 4; ++i) {
        float t = A.m[4 * i + 0] * b.v[0] + A.m[4 * i + 1] * b.v[1] + A.m[4 * i + 2] * b.v[2] + A.m[4 * i + 3] * b.v[3];
        d.v[i] = (float)((double)t * 1.0 + 0.0);
    }
    return d;
}

// ---- cv::gemm general path with GEMM_2_T: D = alpha * A * B^T, double accumulation
static inline M33 gemm_acc64_ABt(const M33& A, const M33& B, double alpha) {
    M33 D;
    alpha = (double)(float)alpha;
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) {
            double s = 0;
            for (int k = 0; k < 3; ++k) s += (double)A.m[3 * i + k] * (double)B.m[3 * j + k];
            D.m[3 * i + j] = (float)(s * alpha);
        }
    return D;
}

// ---- cv::gemm general path, 1xn * nx1 -> scalar, double accumulation, (float)(s*alpha)
static inline float gemm_acc64_dot(const float* a, const float* b, int n, double alpha) {
    alpha = (double)(float)alpha;
    double s = 0;
    for (int k = 0; k < n; ++k) s += (double)a[k] * (double)b[k];
    return (float)(s * alpha);
}

// ---- cv::invert, 3x3 CV_32F, DECOMP_LU: closed form in double
static inline bool invert33(const M33& S, M33& D) {
    const float* s = S.m;
#define SF(r, c) ((double)s[3 * (r) + (c)])
    double d = SF(0, 0) * (SF(1, 1) * SF(2, 2) - SF(1, 2) * SF(2, 1)) -
               SF(0, 1) * (SF(1, 0) * SF(2, 2) - SF(1, 2) * SF(2, 0)) +
               SF(0, 2) * (SF(1, 0) * SF(2, 1) - SF(1, 1) * SF(2, 0));
    if (d == 0.) { std::memset(D.m, 0, sizeof(D.m)); return false; }
    d = 1. / d;
    double t[9];
    t[0] = (SF(1, 1) * SF(2, 2) - SF(1, 2) * SF(2, 1)) * d;
    t[1] = (SF(0, 2) * SF(2, 1) - SF(0, 1) * SF(2, 2)) * d;
    t[2] = (SF(0, 1) * SF(1, 2) - SF(0, 2) * SF(1, 1)) * d;
    t[3] = (SF(1, 2) * SF(2, 0) - SF(1, 0) * SF(2, 2)) * d;
    t[4] = (SF(0, 0) * SF(2, 2) - SF(0, 2) * SF(2, 0)) * d;
    t[5] = (SF(0, 2) * SF(1, 0) - SF(0, 0) * SF(1, 2)) * d;
    t[6] = (SF(1, 0) * SF(2, 1) - SF(1, 1) * SF(2, 0)) * d;
    t[7] = (SF(0, 1) * SF(2, 0) - SF(0, 0) * SF(2, 1)) * d;
    t[8] = (SF(0, 0) * SF(1, 1) - SF(0, 1) * SF(1, 0)) * d;
#undef SF
    for (int i = 0; i < 9; ++i) D.m[i] = (float)t[i];
    return true;
}

// ---- cv::solve(A, B, X, DECOMP_LU) for 3x3 A and 3x3 B: hal::LU32f -> LUImpl<float>
static inline bool solve_lu33(const M33& A_, const M33& B_, M33& X) {
    float A[9], b[9];
    std::memcpy(A, A_.m, sizeof(A));
    std::memcpy(b, B_.m, sizeof(b));
    const int m = 3, n = 3;
    const float eps = FLT_EPSILON * 100;
    for (int i = 0; i < m; ++i) {
        int k = i;
        for (int j = i + 1; j < m; ++j)
            if (std::abs(A[j * 3 + i]) > std::abs(A[k * 3 + i])) k = j;
        if (std::abs(A[k * 3 + i]) < eps) { std::memset(X.m, 0, sizeof(X.m)); return false; }
        if (k != i) {
            for (int j = i; j < m; ++j) { float t = A[i * 3 + j]; A[i * 3 + j] = A[k * 3 + j]; A[k * 3 + j] = t; }
            for (int j = 0; j < n; ++j) { float t = b[i * 3 + j]; b[i * 3 + j] = b[k * 3 + j]; b[k * 3 + j] = t; }
        }
        float d = -1 / A[i * 3 + i];
        for (int j = i + 1; j < m; ++j) {
            float alpha = A[j * 3 + i] * d;
            for (k = i + 1; k < m; ++k) A[j * 3 + k] += alpha * A[i * 3 + k];
            for (k = 0; k < n; ++k) b[j * 3 + k] += alpha * b[i * 3 + k];
        }
    }
    for (int i = m - 1; i >= 0; --i)
        for (int j = 0; j < n; ++j) {
            float s = b[i * 3 + j];
            for (int k = i + 1; k < m; ++k) s -= A[i * 3 + k] * b[k * 3 + j];
            b[i * 3 + j] = s / A[i * 3 + i];
        }
    std::memcpy(X.m, b, sizeof(b));
    return true;
}

// ---- cv::fastAtan2(y, x) scalar, degrees in [0, 360]
static inline float fastAtan2(float y, float x) {
    static const float p1 = 0.9997878412794807f * (float)(180 / 3.14159265358979323846);
    static const float p3 = -0.3258083974640975f * (float)(180 / 3.14159265358979323846);
    static const float p5 = 0.1555786518463281f * (float)(180 / 3.14159265358979323846);
    static const float p7 = -0.04432655554792128f * (float)(180 / 3.14159265358979323846);
    float ax = std::abs(x), ay = std::abs(y);
    float a, c, c2;
    if (ax >= ay) {
        c = ay / (ax + (float)DBL_EPSILON);
        c2 = c * c;
        a = (((p7 * c2 + p5) * c2 + p3) * c2 + p1) * c;
    } else {
        c = ax / (ay + (float)DBL_EPSILON);
        c2 = c * c;
        a = 90.f - (((p7 * c2 + p5) * c2 + p3) * c2 + p1) * c;
    }
    if (x < 0) a = 180.f - a;
    if (y < 0) a = 360.f - a;
    return a;
}

static inline M33 transpose(const M33& A) {
    M33 T;
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) T.m[3 * i + j] = A.m[3 * j + i];
    return T;
}

}  // namespace cvl
