// Host-side per-pair constants of the semi-dense path, computed once per (key frame, neighbour)
// instead of once per pixel as the reference does (PM.cc:885-891 inside GetSearchRange, 854-860
// inside GetPixelDepth).
//
// The reference evaluates these through cv::Mat expressions.  To give the same numbers, the
// small-matrix products below follow what OpenCV's cv::gemm / cv::solve / cv::invert do for
// these shapes (OpenCV is a system dependency of the reference, CMakeLists.txt:37-43):
//   - a product with a transposed operand (A*B.t()) goes through the general gemm path, which
//     accumulates in double and rounds once;
//   - 3x3*3x1 and 3x3*3x3 products without transposition go through the small-matrix path,
//     which multiplies and adds in float, left to right, with the alpha/beta epilogue in double;
//   - inv(A)*B is rewritten to solve(A, B) (LU with partial pivoting in float);
//   - inv(K) alone is the 3x3 closed form in double.
#pragma once
#include <cmath>
#include <cstring>

namespace sdpm {
namespace geom {

// D = alpha * A * B^T, 3x3, double accumulation (general path)
inline void mul_ABt_acc64(const float* A, const float* B, double alpha, float* D) {
    for (int r = 0; r < 3; ++r)
        for (int c = 0; c < 3; ++c) {
            double acc = 0.0;
            for (int k = 0; k < 3; ++k) acc += (double)A[r * 3 + k] * (double)B[c * 3 + k];
            D[r * 3 + c] = (float)(acc * alpha);
        }
}

// d = (float)(A*b) + c, 3x3 * 3x1 on the small path (alpha = beta = 1)
inline void mul_Ab_plus_c(const float* A, const float* b, const float* c, float* d) {
    for (int r = 0; r < 3; ++r) {
        float dot = A[r * 3] * b[0] + A[r * 3 + 1] * b[1] + A[r * 3 + 2] * b[2];
        d[r] = (float)((double)dot + (double)c[r]);
    }
}

inline void mul_AB_small(const float* A, const float* B, float* D) {
    float out[9];
    for (int r = 0; r < 3; ++r)
        for (int c = 0; c < 3; ++c) out[r * 3 + c] = A[r * 3] * B[c] + A[r * 3 + 1] * B[3 + c] + A[r * 3 + 2] * B[6 + c];
    std::memcpy(D, out, sizeof(out));
}

// R21 = Rcw2*Rcw1^T ; t21 = -Rcw2*Rcw1^T*tcw1 + tcw2      (PM.cc:890-891, 643-644, 859-860)
inline void relative_pose(const float* Rcw1, const float* tcw1, const float* Rcw2, const float* tcw2, float* R21, float* t21) {
    float negR[9];
    mul_ABt_acc64(Rcw2, Rcw1, 1.0, R21);
    mul_ABt_acc64(Rcw2, Rcw1, -1.0, negR);
    mul_Ab_plus_c(negR, tcw1, tcw2, t21);
}

// 3x3 inverse, cofactors in double (cv::invert, DECOMP_LU, 3x3 CV_32F)
inline bool invert3(const float* S, float* D) {
    double m[9];
    for (int i = 0; i < 9; ++i) m[i] = (double)S[i];
    double det = m[0] * (m[4] * m[8] - m[5] * m[7]) - m[1] * (m[3] * m[8] - m[5] * m[6]) + m[2] * (m[3] * m[7] - m[4] * m[6]);
    if (det == 0.0) { std::memset(D, 0, 9 * sizeof(float)); return false; }
    const double id = 1.0 / det;
    D[0] = (float)((m[4] * m[8] - m[5] * m[7]) * id);
    D[1] = (float)((m[2] * m[7] - m[1] * m[8]) * id);
    D[2] = (float)((m[1] * m[5] - m[2] * m[4]) * id);
    D[3] = (float)((m[5] * m[6] - m[3] * m[8]) * id);
    D[4] = (float)((m[0] * m[8] - m[2] * m[6]) * id);
    D[5] = (float)((m[2] * m[3] - m[0] * m[5]) * id);
    D[6] = (float)((m[3] * m[7] - m[4] * m[6]) * id);
    D[7] = (float)((m[1] * m[6] - m[0] * m[7]) * id);
    D[8] = (float)((m[0] * m[4] - m[1] * m[3]) * id);
    return true;
}

// X = A^-1 * B by Gaussian elimination with partial pivoting, all in float (cv::solve, DECOMP_LU,
// multi-column right-hand side -> hal::LU32f)
inline bool solve3(const float* A_in, const float* B_in, float* X) {
    float A[3][3], B[3][3];
    std::memcpy(A, A_in, sizeof(A));
    std::memcpy(B, B_in, sizeof(B));
    const float tiny = 1.1920928955078125e-07f * 100;  // FLT_EPSILON*100
    for (int col = 0; col < 3; ++col) {
        int piv = col;
        for (int r = col + 1; r < 3; ++r)
            if (std::fabs(A[r][col]) > std::fabs(A[piv][col])) piv = r;
        if (std::fabs(A[piv][col]) < tiny) { std::memset(X, 0, 9 * sizeof(float)); return false; }
        if (piv != col) {
            for (int c = col; c < 3; ++c) std::swap(A[col][c], A[piv][c]);
            for (int c = 0; c < 3; ++c) std::swap(B[col][c], B[piv][c]);
        }
        const float ninv = -1 / A[col][col];
        for (int r = col + 1; r < 3; ++r) {
            const float f = A[r][col] * ninv;
            for (int c = col + 1; c < 3; ++c) A[r][c] += f * A[col][c];
            for (int c = 0; c < 3; ++c) B[r][c] += f * B[col][c];
        }
    }
    for (int r = 2; r >= 0; --r)
        for (int c = 0; c < 3; ++c) {
            float acc = B[r][c];
            for (int k = r + 1; k < 3; ++k) acc -= A[r][k] * B[k][c];
            B[r][c] = acc / A[r][r];
        }
    std::memcpy(X, B, sizeof(B));
    return true;
}

// F12 = K1^-T * [t12]x * R12 * K2^-1      (PM.cc:972-987; live twin LocalMapping.cc:549-566, 711-716)
inline void fundamental(const float K1[4], const float* Rcw1, const float* tcw1, const float K2[4], const float* Rcw2,
                        const float* tcw2, float* F12) {
    float R12[9], t12[3];
    relative_pose(Rcw2, tcw2, Rcw1, tcw1, R12, t12);  // R12 = R1w*R2w^T, t12 = -R1w*R2w^T*t2w + t1w
    const float skew[9] = {0, -t12[2], t12[1], t12[2], 0, -t12[0], -t12[1], t12[0], 0};
    const float K1t[9] = {K1[0], 0, 0, 0, K1[1], 0, K1[2], K1[3], 1};
    const float K2m[9] = {K2[0], 0, K2[2], 0, K2[1], K2[3], 0, 0, 1};
    float S[9], K2i[9], M[9];
    solve3(K1t, skew, S);
    invert3(K2m, K2i);
    mul_AB_small(S, R12, M);
    mul_AB_small(M, K2i, F12);
}

// Twc from Tcw as KeyFrame::SetPose builds it (KeyFrame.cc:70-84): Rwc = Rcw^T, Ow = -Rwc*tcw
inline void pose_inverse(const float* Rcw, const float* tcw, float* Twc /*16*/) {
    std::memset(Twc, 0, 16 * sizeof(float));
    for (int r = 0; r < 3; ++r) {
        for (int c = 0; c < 3; ++c) Twc[r * 4 + c] = Rcw[c * 3 + r];
        float dot = Rcw[0 * 3 + r] * tcw[0] + Rcw[1 * 3 + r] * tcw[1] + Rcw[2 * 3 + r] * tcw[2];
        Twc[r * 4 + 3] = (float)((double)dot * -1.0 + 0.0);
    }
    Twc[15] = 1.0f;
}

// smallest float >= v (so that `x < v` for a float x and a double v equals `x < up(v)` in float)
inline float round_up_to_float(double v) {
    float f = (float)v;
    if ((double)f < v) f = std::nextafterf(f, INFINITY);
    return f;
}

}  // namespace geom
}  // namespace sdpm
