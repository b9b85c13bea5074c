// oracle/pm_oracle.cc -- TEST INFRASTRUCTURE (CPU oracle).  See pm_oracle.h for the usage rules.
//
// Function-by-function restatement of /root/reference/src/Modeler/ProbabilityMapping.cc ("PM.cc")
// on plain structs.  Every function cites the PM.cc lines it follows.  Arithmetic is float32 in
// the reference's operation order; `double` appears exactly where the reference promotes
// (the double literal THETA, std::pow, std::accumulate with a 0.0 seed, cv::gemm's internals --
// the latter through cvlike.h).  Build with -ffp-contract=off (no FMA contraction).
//
// Stated divergences from PM.cc:
//   H1  InterKeyFrameDepthChecking uses snapshot semantics: it reads the neighbours' post-intra maps
//       and writes a separate "checked" plane (the reference mutates depth_map_ in place in
//       std::set<KeyFrame*> pointer order, which is not reproducible).
//   H2  every plane read goes through clamp-to-edge addressing (the reference reads out of bounds
//       at uj == cols, y1 == rows, uj_minus == -1 and in the +-1 refinement taps).
//   H8  the gradient front end (GradImg / GradTheta / I_stddev) is defined here because the
//       reference contains no code that computes those fields.
//   NaN search bounds give an empty candidate range (the reference would execute (int)NaN, UB).
//   The per-pair R21/t21 that PM.cc rebuilds inside every GetSearchRange / GetPixelDepth call
//   (PM.cc:885-891, 854-860) are computed once per pair with identical arithmetic.
#include "pm_oracle.h"
#include "cvlike.h"

#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <map>
#include <numeric>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif

namespace {

using cvl::M33;
using cvl::M44;
using cvl::V3;
using cvl::V4;

static int g_threads = 1;

// PM.h:64-70
struct depthHo {
    float depth = 0.0f;  // INVERSE depth
    float sigma = 0.0f;
    bool supported = false;
};

// The fields of ORB_SLAM2::KeyFrame that PM.cc reads (none exist in include/KeyFrame.h)
struct KeyFrame {
    int rows = 0, cols = 0;
    std::vector<uint8_t> im_;
    std::vector<float> GradImg, GradTheta;
    float I_stddev = 0.f;
    float fx = 0, fy = 0, cx = 0, cy = 0;
    M33 Rcw;   // GetRotation(),    KeyFrame.cc:109-113
    V3 tcw;    // GetTranslation(), KeyFrame.cc:115-119
    M44 Twc;   // GetPoseInverse(), KeyFrame.cc:92-96, built by SetPose KeyFrame.cc:70-84
    std::vector<float> depth_map_, depth_sigma_;  // after the intra-KF stage
    std::vector<float> depth_checked_;            // after the inter-KF stage (H1 snapshot)
    std::vector<float> SemiDensePointSets_;
    bool semidense_flag_ = false, interKF_depth_flag_ = false;
};

struct Oracle {
    OraParams P;
    std::map<int, KeyFrame> kfs;
};

// H2: clamp-to-edge plane access (cv::Mat::at<T>(y,x) where the reference is in bounds)
template <typename T>
static inline T at(const std::vector<T>& pl, const KeyFrame& kf, int y, int x) {
    int yy = y < 0 ? 0 : (y >= kf.rows ? kf.rows - 1 : y);
    int xx = x < 0 ? 0 : (x >= kf.cols ? kf.cols - 1 : x);
    return pl[(size_t)yy * kf.cols + xx];
}
template <typename T>
static inline T at_raw(const T* pl, int w, int h, int y, int x) {
    int yy = y < 0 ? 0 : (y >= h ? h - 1 : y);
    int xx = x < 0 ? 0 : (x >= w ? w - 1 : x);
    return pl[(size_t)yy * w + xx];
}

// PM.cc:40-59  bilinear<T>(img, y, x)
template <typename T>
static float bilinear(const T* img, int w, int h, const float& y, const float& x) {
    int x0 = (int)std::floor(x);
    int y0 = (int)std::floor(y);
    int x1 = x0 + 1;
    int y1 = y0 + 1;

    float x0_weight = x1 - x;
    float y0_weight = y1 - y;
    float x1_weight = 1.0f - x0_weight;
    float y1_weight = 1.0f - y0_weight;

    float interpolated = at_raw(img, w, h, y0, x0) * x0_weight * y0_weight +
                         at_raw(img, w, h, y0, x1) * x1_weight * y0_weight +
                         at_raw(img, w, h, y1, x0) * x0_weight * y1_weight +
                         at_raw(img, w, h, y1, x1) * x1_weight * y1_weight;
    return interpolated;
}

// KeyFrame::SetPose, KeyFrame.cc:70-84.  Rwc = Rcw.t(); Ow = -Rwc*tcw is
// gemm(Rwc, tcw, alpha=-1) on the small path (matop.cpp: a scaled operand folds into alpha).
static void set_pose(KeyFrame& kf, const float Tcw[12]) {
    for (int r = 0; r < 3; ++r) {
        for (int c = 0; c < 3; ++c) kf.Rcw.m[3 * r + c] = Tcw[4 * r + c];
        kf.tcw.v[r] = Tcw[4 * r + 3];
    }
    M33 Rwc = cvl::transpose(kf.Rcw);
    V3 Ow = cvl::gemm_small_33x31(Rwc, kf.tcw, -1.0, nullptr, 0.0);
    std::memset(kf.Twc.m, 0, sizeof(kf.Twc.m));
    for (int r = 0; r < 3; ++r) {
        for (int c = 0; c < 3; ++c) kf.Twc.m[4 * r + c] = Rwc.m[3 * r + c];
        kf.Twc.m[4 * r + 3] = Ow.v[r];
    }
    kf.Twc.m[15] = 1.0f;
}

// Gradient front end (H8; build-defined).  3x3 Sobel on the u8 image, replicate border, scale 1/8
// (pinned: equals cv2.Sobel(..., ksize=3, scale=1/8, BORDER_REPLICATE) bit for bit),
// GradImg = sqrtf(gx*gx + gy*gy), GradTheta = fastAtan2(gy, gx) in degrees,
// I_stddev = population standard deviation of the u8 image (equals cv2.meanStdDev).
static void front_end(KeyFrame& kf) {
    const int W = kf.cols, H = kf.rows;
    kf.GradImg.assign((size_t)W * H, 0.f);
    kf.GradTheta.assign((size_t)W * H, 0.f);
    const uint8_t* im = kf.im_.data();
    long long s1 = 0, s2 = 0;
#pragma omp parallel for reduction(+ : s1, s2) num_threads(g_threads)
    for (int y = 0; y < H; ++y) {
        for (int x = 0; x < W; ++x) {
            int p00 = at_raw(im, W, H, y - 1, x - 1), p01 = at_raw(im, W, H, y - 1, x), p02 = at_raw(im, W, H, y - 1, x + 1);
            int p10 = at_raw(im, W, H, y, x - 1), p12 = at_raw(im, W, H, y, x + 1);
            int p20 = at_raw(im, W, H, y + 1, x - 1), p21 = at_raw(im, W, H, y + 1, x), p22 = at_raw(im, W, H, y + 1, x + 1);
            int sx = (p02 + 2 * p12 + p22) - (p00 + 2 * p10 + p20);
            int sy = (p20 + 2 * p21 + p22) - (p00 + 2 * p01 + p02);
            float gx = (float)sx * 0.125f, gy = (float)sy * 0.125f;
            kf.GradImg[(size_t)y * W + x] = std::sqrt(gx * gx + gy * gy);
            kf.GradTheta[(size_t)y * W + x] = cvl::fastAtan2(gy, gx);
            int p = im[(size_t)y * W + x];
            s1 += p;
            s2 += (long long)p * p;
        }
    }
    double n = (double)W * H;
    double mean = (double)s1 / n;
    double var = (double)s2 / n - mean * mean;
    kf.I_stddev = (float)std::sqrt(std::max(var, 0.0));
}

// Per-pair geometry, hoisted.  R21/t21: PM.cc:885-891 (== 854-860, 638-644).  F12: PM.cc:972-987.
struct PairGeom {
    M33 R21;
    V3 t21;
    M33 F12;
};

// R21 = Rcw2*Rcw1.t()        -> gemm(Rcw2, Rcw1, 1, GEMM_2_T): general path, double accumulation.
// t21 = -Rcw2*Rcw1.t()*tcw1+tcw2 -> tmp = gemm(Rcw2, Rcw1, -1, GEMM_2_T) materialised (float),
//                                   then gemm(tmp, tcw1, 1, tcw2, 1): small path.
static void relative_pose(const M33& Rcw1, const V3& tcw1, const M33& Rcw2, const V3& tcw2, M33& R21, V3& t21) {
    R21 = cvl::gemm_acc64_ABt(Rcw2, Rcw1, 1.0);
    M33 neg = cvl::gemm_acc64_ABt(Rcw2, Rcw1, -1.0);
    t21 = cvl::gemm_small_33x31(neg, tcw1, 1.0, &tcw2, 1.0);
}

static M33 Kmat(float fx, float fy, float cx, float cy) {
    M33 K = {{fx, 0, cx, 0, fy, cy, 0, 0, 1}};
    return K;
}

// PM.cc:972-987 ComputeFundamental (twin: LocalMapping.cc:549-566) + GetSkewSymmetricMatrix
// (declared PM.h:102, never defined; twin LocalMapping.cc:711-716).
// K1.t().inv()*t12x*R12*K2.inv() evaluates (matop.cpp) as
//   S  = solve(K1^T, t12x, DECOMP_LU)        (MatOp_Invert::matmul turns inv(A)*B into a solve)
//   M  = gemm(S, R12)                        small path
//   F  = gemm(M, invert(K2))                 small path, invert = 3x3 closed form in double
static M33 ComputeFundamental(const KeyFrame& kf1, const KeyFrame& kf2) {
    M33 R12;
    V3 t12;
    relative_pose(kf2.Rcw, kf2.tcw, kf1.Rcw, kf1.tcw, R12, t12);  // R12 = R1w*R2w.t(); t12 = -R1w*R2w.t()*t2w+t1w
    M33 t12x = {{0, -t12.v[2], t12.v[1], t12.v[2], 0, -t12.v[0], -t12.v[1], t12.v[0], 0}};
    M33 K1 = Kmat(kf1.fx, kf1.fy, kf1.cx, kf1.cy);
    M33 K2 = Kmat(kf2.fx, kf2.fy, kf2.cx, kf2.cy);
    M33 K1t = cvl::transpose(K1);
    M33 S, K2i;
    cvl::solve_lu33(K1t, t12x, S);
    cvl::invert33(K2, K2i);
    M33 M = cvl::gemm_small_33x33(S, R12, 1.0);
    return cvl::gemm_small_33x33(M, K2i, 1.0);
}

static PairGeom pair_geometry(const KeyFrame& kf1, const KeyFrame& kf2) {
    PairGeom g;
    relative_pose(kf1.Rcw, kf1.tcw, kf2.Rcw, kf2.tcw, g.R21, g.t21);
    g.F12 = ComputeFundamental(kf1, kf2);
    return g;
}

// PM.cc:877-910 GetSearchRange.  xp2 = R21*xp1*rho + t21 is ONE gemm(R21, xp1, rho, t21, 1)
// (MatOp_GEMM::multiply folds the scalar into alpha, MatOp_GEMM::add folds t21 in as C); small path.
// The multiplication by INVERSE depth is the reference's and is kept.
static void GetSearchRange(float& umin, float& umax, int px, int py, float mind, float maxd, float fx, float fy,
                           float cx, float cy, int cols, const M33& R21, const V3& t21) {
    V3 xp1 = {{(px - cx) / fx, (py - cy) / fy, 1.0f}};
    V3 xp2_min = cvl::gemm_small_33x31(R21, xp1, (double)mind, &t21, 1.0);
    V3 xp2_max = cvl::gemm_small_33x31(R21, xp1, (double)maxd, &t21, 1.0);

    umin = fx * xp2_min.v[0] / xp2_min.v[2] + cx;
    umax = fx * xp2_max.v[0] / xp2_max.v[2] + cx;

    if (umin > umax) {
        float temp = umax;
        umax = umin;
        umin = temp;
    }
    if (umin < 0) umin = 0;
    if (umax < 0) umax = 0;
    if (umin > cols) umin = cols;
    if (umax > cols) umax = cols;
}

// PM.cc:845-875 GetPixelDepth (Eq. 8).  `R21.row(2) * xp * ucx` is gemm(1x3, 3x1, alpha=ucx):
// general path (double accumulation); `fx * (R21.row(0) * xp)` likewise with alpha=fx.
static void GetPixelDepth(float uj, int px, int py, float fx, float fy, float cx, float cy, const M33& R21,
                          const V3& t21, float& p) {
    float ucx = uj - cx;
    float xp[3] = {(px - cx) / fx, (py - cy) / fy, 1.0f};
    float num1 = cvl::gemm_acc64_dot(&R21.m[6], xp, 3, (double)ucx);
    float num2 = cvl::gemm_acc64_dot(&R21.m[0], xp, 3, (double)fx);
    float denom1 = -t21.v[2] * ucx;
    float denom2 = fx * t21.v[0];
    p = (num1 - num2) / (denom1 + denom2);
}

// PM.cc:806-829 ComputeInvDepthHypothesis
static void ComputeInvDepthHypothesis(const KeyFrame& kf, float ustar, float ustar_var, depthHo* dh, int x, int y,
                                      const PairGeom& g) {
    float inv_pixel_depth = 0.0;
    GetPixelDepth(ustar, x, y, kf.fx, kf.fy, kf.cx, kf.cy, g.R21, g.t21, inv_pixel_depth);

    float ustar_min = ustar - std::sqrt(ustar_var);
    float inv_depth_min = 0.0;
    GetPixelDepth(ustar_min, x, y, kf.fx, kf.fy, kf.cx, kf.cy, g.R21, g.t21, inv_depth_min);

    float ustar_max = ustar + std::sqrt(ustar_var);
    float inv_depth_max = 0.0;
    GetPixelDepth(ustar_max, x, y, kf.fx, kf.fy, kf.cx, kf.cy, g.R21, g.t21, inv_depth_max);

    // Eq. 9; abs is std::abs(float)
    float sigma_depth = std::max(std::abs(inv_depth_max - inv_pixel_depth), std::abs(inv_depth_min - inv_pixel_depth));

    dh->depth = inv_pixel_depth;
    dh->sigma = sigma_depth;
    dh->supported = true;
}

// PM.cc:385-465 EpipolarSearch
static void EpipolarSearch(const OraParams& P, const KeyFrame& kf1, const KeyFrame& kf2, const int x, const int y,
                           float pixel, float min_depth, float max_depth, depthHo* dh, const PairGeom& geo,
                           float& best_u, float& best_v, float th_pi, float rot) {
    const float* F12 = geo.F12.m;
    float a = x * F12[0] + y * F12[3] + F12[6];
    float b = x * F12[1] + y * F12[4] + F12[7];
    float c = x * F12[2] + y * F12[5] + F12[8];

    if ((a / b) < -P.slope_limit || a / b > P.slope_limit) return;

    float old_err = 1000000.0;
    float best_photometric_err = 0.0;
    float best_gradient_modulo_err = 0.0;
    int best_pixel = 0;

    int vj, uj_plus, uj_minus;
    float g, q, denomiator, ustar, ustar_var;
    const double THETA = P.theta;
    const int W2 = kf2.cols, H2 = kf2.rows;

    float umin(0.0), umax(0.0);
    GetSearchRange(umin, umax, x, y, min_depth, max_depth, kf1.fx, kf1.fy, kf1.cx, kf1.cy, kf1.cols, geo.R21, geo.t21);
    if (!(umin == umin) || !(umax == umax)) return;  // NaN bounds: empty range (stated divergence)
    const float uend = std::floor(umax);
    for (int uj = (int)std::ceil(umin); uj <= uend; uj++) {
        float vf = (a / b) * uj + (c / b);
        if (!(std::fabs(vf) < 2147483648.0f)) continue;  // (int) of NaN/overflow is UB; x86 gives INT_MIN -> vj<=0 -> continue
        vj = -(int)(vf);
        if (vj <= 0 || vj >= kf2.rows) { continue; }

        // condition 1:
        if (at(kf2.GradImg, kf2, vj, uj) < P.lambdaG) { continue; }

        // condition 2:
        float th_epipolar_line = cvl::fastAtan2(-a / b, 1);
        float temp_gradth = at(kf2.GradTheta, kf2, vj, uj);
        float ang_diff = temp_gradth - th_epipolar_line;
        if (ang_diff >= 360) { ang_diff -= 360; }
        if (ang_diff < 0) { ang_diff += 360; }
        if (ang_diff > 180) ang_diff = 360 - ang_diff;
        if (ang_diff > 90) ang_diff = 180 - ang_diff;
        if (ang_diff > P.lambdaL) { continue; }

        // condition 3:
        float ang_pi_rot = th_pi + rot;
        if (ang_pi_rot >= 360) { ang_pi_rot -= 360; }
        if (ang_pi_rot < 0) { ang_pi_rot += 360; }
        float th_diff = at(kf2.GradTheta, kf2, vj, uj) - ang_pi_rot;
        if (th_diff >= 360) { th_diff -= 360; }
        if (th_diff < 0) { th_diff += 360; }
        if (th_diff > 180) th_diff = 360 - th_diff;
        if (th_diff > P.lambdaTheta) continue;

        float photometric_err = pixel - bilinear<uint8_t>(kf2.im_.data(), W2, H2, -((a / b) * uj + (c / b)), uj);
        float gradient_modulo_err =
            at(kf1.GradImg, kf1, y, x) - bilinear<float>(kf2.GradImg.data(), W2, H2, -((a / b) * uj + (c / b)), uj);

        // THETA is a double literal: the division and the sum are evaluated in double, then stored to float
        float err = (photometric_err * photometric_err + (gradient_modulo_err * gradient_modulo_err) / THETA);
        if (err < old_err) {
            best_pixel = uj;
            old_err = err;
            best_photometric_err = photometric_err;
            best_gradient_modulo_err = gradient_modulo_err;
        }
    }

    if (old_err < 1000000.0) {
        uj_plus = best_pixel + 1;
        uj_minus = best_pixel - 1;

        g = (bilinear<uint8_t>(kf2.im_.data(), W2, H2, -((a / b) * uj_plus + (c / b)), uj_plus) -
             bilinear<uint8_t>(kf2.im_.data(), W2, H2, -((a / b) * uj_minus + (c / b)), uj_minus)) / 2;
        q = (bilinear<float>(kf2.GradImg.data(), W2, H2, -((a / b) * uj_plus + (c / b)), uj_plus) -
             bilinear<float>(kf2.GradImg.data(), W2, H2, -((a / b) * uj_minus + (c / b)), uj_minus)) / 2;

        denomiator = (g * g + (1 / THETA) * q * q);
        ustar = best_pixel + (g * best_photometric_err + (1 / THETA) * q * best_gradient_modulo_err) / denomiator;
        ustar_var = (2 * kf2.I_stddev * kf2.I_stddev / denomiator);

        best_u = ustar;
        best_v = -((a / b) * best_u + (c / b));

        ComputeInvDepthHypothesis(kf1, ustar, ustar_var, dh, x, y, geo);
    }
}

// PM.cc:912-918
static bool ChiTest(const OraParams& P, const depthHo& ha, const depthHo& hb, float* chi_val) {
    float num = (ha.depth - hb.depth) * (ha.depth - hb.depth);
    float chi_test = num / (ha.sigma * ha.sigma) + num / (hb.sigma * hb.sigma);
    if (chi_val) *chi_val = chi_test;
    return (chi_test < P.chi2_fuse);
}

// PM.cc:920-924
static bool ChiTest(const OraParams& P, const float& a, const float& b, const float sigma_a, float sigma_b) {
    float num = (a - b) * (a - b);
    float chi_test = num / (sigma_a * sigma_a) + num / (sigma_b * sigma_b);
    return (chi_test < P.chi2_fuse);
}

// PM.cc:926-945 (std::pow(float,int) promotes to double)
static void GetFusion(const std::pair<float, float>* supported, size_t t, float& depth, float& sigma) {
    float pjsj = 0;
    float rsj = 0;
    float min_sigma = supported[0].second;
    for (size_t i = 0; i < t; i++) {
        pjsj += supported[i].first / std::pow((double)supported[i].second, 2);
        rsj += 1 / std::pow((double)supported[i].second, 2);
        if (supported[i].second < min_sigma) min_sigma = supported[i].second;
    }
    depth = pjsj / rsj;
    sigma = min_sigma;
}

// PM.cc:947-970
static void GetFusion(const depthHo* compatible_ho, size_t n, depthHo& hypothesis, float* min_sigma) {
    hypothesis.depth = 0;
    hypothesis.sigma = 0;
    float temp_min_sigma = compatible_ho[0].sigma;
    float pjsj = 0;
    float rsj = 0;
    for (size_t j = 0; j < n; j++) {
        pjsj += compatible_ho[j].depth / std::pow((double)compatible_ho[j].sigma, 2);
        rsj += 1 / std::pow((double)compatible_ho[j].sigma, 2);
        if (std::pow((double)compatible_ho[j].sigma, 2) < std::pow((double)temp_min_sigma, 2)) {
            temp_min_sigma = compatible_ho[j].sigma;
        }
    }
    hypothesis.depth = pjsj / rsj;
    hypothesis.sigma = std::sqrt(1 / rsj);
    hypothesis.supported = true;
    if (min_sigma) *min_sigma = temp_min_sigma;
}

// PM.cc:598-626
static void InverseDepthHypothesisFusion(const OraParams& P, const std::vector<depthHo>& h, depthHo& dist) {
    dist.depth = 0;
    dist.sigma = 0;
    dist.supported = false;

    std::vector<depthHo> compatible_ho;
    std::vector<depthHo> compatible_ho_temp;
    float chi = 0;

    for (size_t a = 0; a < h.size(); a++) {
        compatible_ho_temp.clear();
        for (size_t b = 0; b < h.size(); b++) {
            if (ChiTest(P, h[a], h[b], &chi)) { compatible_ho_temp.push_back(h[b]); }
        }
        if (compatible_ho_temp.size() > compatible_ho.size()) {
            compatible_ho.clear();
            compatible_ho.insert(compatible_ho.end(), compatible_ho_temp.begin(), compatible_ho_temp.end());
        }
    }
    if ((int)compatible_ho.size() >= P.lambdaN) {
        GetFusion(compatible_ho.data(), compatible_ho.size(), dist, &chi);
    }
}

// PM.cc:486-547
static void IntraKeyFrameDepthChecking(const OraParams& P, std::vector<float>& depth_map, std::vector<float>& depth_sigma,
                                       int rows, int cols) {
    std::vector<float> depth_map_new = depth_map;
    std::vector<float> depth_sigma_new = depth_sigma;
    auto D = [&](int y, int x) -> const float& { return depth_map[(size_t)y * cols + x]; };
    auto S = [&](int y, int x) -> const float& { return depth_sigma[(size_t)y * cols + x]; };

#pragma omp parallel for schedule(dynamic) collapse(2) num_threads(g_threads)
    for (int py = 2; py < (rows - 2); py++) {
        for (int px = 2; px < (cols - 2); px++) {
            if (D(py, px) > 0.000001) {
                depthHo compatible_neighbor_ho[9];
                int ncomp = 0;
                depthHo dha, dhb;
                dha.depth = D(py, px);
                dha.sigma = S(py, px);
                for (int y = py - 1; y <= py + 1; y++) {
                    for (int x = px - 1; x <= px + 1; x++) {
                        if (x == px && y == py) continue;
                        if (D(y, x) > 0.000001) {
                            if (ChiTest(P, D(y, x), D(py, px), S(y, x), S(py, px))) {
                                dhb.depth = D(y, x);
                                dhb.sigma = S(y, x);
                                compatible_neighbor_ho[ncomp++] = dhb;
                            }
                        }
                    }
                }
                compatible_neighbor_ho[ncomp++] = dha;  // dont forget itself.

                if (ncomp >= 3) {
                    depthHo fusion;
                    float min_sigma = 0;
                    GetFusion(compatible_neighbor_ho, ncomp, fusion, &min_sigma);
                    depth_map_new[(size_t)py * cols + px] = fusion.depth;
                    depth_sigma_new[(size_t)py * cols + px] = min_sigma;
                } else {
                    depth_map_new[(size_t)py * cols + px] = 0.0;
                    depth_sigma_new[(size_t)py * cols + px] = 0.0;
                }
            }
        }
    }
    depth_map = depth_map_new;
    depth_sigma = depth_sigma_new;
}

// PM.cc:549-596
static void IntraKeyFrameDepthGrowing(const OraParams& P, std::vector<float>& depth_map, std::vector<float>& depth_sigma,
                                      const std::vector<float>& gradimg, int rows, int cols) {
    std::vector<float> depth_map_new = depth_map;
    std::vector<float> depth_sigma_new = depth_sigma;
    auto D = [&](int y, int x) -> const float& { return depth_map[(size_t)y * cols + x]; };
    auto S = [&](int y, int x) -> const float& { return depth_sigma[(size_t)y * cols + x]; };

#pragma omp parallel for schedule(dynamic) collapse(2) num_threads(g_threads)
    for (int py = 2; py < (rows - 2); py++) {
        for (int px = 2; px < (cols - 2); px++) {
            if (D(py, px) < 0.000001) {
                if (gradimg[(size_t)py * cols + px] < P.lambdaG) continue;
                std::pair<float, float> supported[8];
                int nsup = 0;
                for (int y = py - 1; y <= py + 1; y++)
                    for (int x = px - 1; x <= px + 1; x++) {
                        if (x == px && y == py) continue;
                        if (ChiTest(P, D(y, x), D(py, px), S(y, x), S(py, px))) {
                            supported[nsup].first = D(y, x);
                            supported[nsup].second = S(y, x);
                            nsup++;
                        }
                    }
                if (nsup >= 2) {
                    float d(0.0), s(0.0);
                    GetFusion(supported, nsup, d, s);
                    depth_map_new[(size_t)py * cols + px] = d;
                    depth_sigma_new[(size_t)py * cols + px] = s;
                }
            }
        }
    }
    depth_map = depth_map_new;
    depth_sigma = depth_sigma_new;
}

// PM.cc:137-257 SemiDenseLoop, per-KF part.  The neighbour pick (151-160), the in-plane rotation
// medians (170-179) and the depth prior (184) are computed by the caller and passed in.
static void SemiDenseLoop(Oracle& O, KeyFrame& kf, const std::vector<const KeyFrame*>& closestMatches,
                          const std::vector<float>& rotIs, float min_depth, float max_depth) {
    const OraParams& P = O.P;
    const int rows = kf.rows, cols = kf.cols;
    kf.depth_map_.assign((size_t)rows * cols, 0.f);
    kf.depth_sigma_.assign((size_t)rows * cols, 0.f);

    std::vector<PairGeom> F;  // PM.cc:189-195 (+ hoisted R21/t21)
    for (size_t j = 0; j < closestMatches.size(); j++) F.push_back(pair_geometry(kf, *closestMatches[j]));

#pragma omp parallel for schedule(dynamic) collapse(2) num_threads(g_threads)
    for (int y = 0 + P.border; y < rows - P.border; y++) {
        for (int x = 0 + P.border; x < cols - P.border; x++) {
            if (kf.GradImg[(size_t)y * cols + x] < P.lambdaG) { continue; }
            float pixel = (float)kf.im_[(size_t)y * cols + x];

            std::vector<depthHo> depth_ho;
            depth_ho.reserve(closestMatches.size());
            for (size_t j = 0; j < closestMatches.size(); j++) {
                const KeyFrame* kf2 = closestMatches[j];
                float rot = rotIs[j];
                float best_u(0.0), best_v(0.0);
                depthHo dh;
                EpipolarSearch(P, kf, *kf2, x, y, pixel, min_depth, max_depth, &dh, F[j], best_u, best_v,
                               kf.GradTheta[(size_t)y * cols + x], rot);
                if (dh.supported && 1 / dh.depth > 0.0) { depth_ho.push_back(dh); }
            }

            if ((int)depth_ho.size() > P.lambdaN) {
                depthHo dh_temp;
                InverseDepthHypothesisFusion(P, depth_ho, dh_temp);
                if (dh_temp.supported) {
                    kf.depth_map_[(size_t)y * cols + x] = dh_temp.depth;
                    kf.depth_sigma_[(size_t)y * cols + x] = dh_temp.sigma;
                }
            }
        }
    }

    IntraKeyFrameDepthChecking(P, kf.depth_map_, kf.depth_sigma_, rows, cols);
    IntraKeyFrameDepthGrowing(P, kf.depth_map_, kf.depth_sigma_, kf.GradImg, rows, cols);
    kf.semidense_flag_ = true;
}

// PM.cc:628-799 InterKeyFrameDepthChecking.  Default: snapshot semantics (H1): reads depth_map_ of the
// current KF and of the neighbours, writes depth_checked_.  With P.interkf_inplace the neighbours that
// were checked by an earlier call are read as checked, which is what the reference's in-place writes
// amount to for a given visiting order (the caller's call order; the reference's is Map pointer order).
static void InterKeyFrameDepthChecking(Oracle& O, KeyFrame& cur, const std::vector<const KeyFrame*>& neighbors) {
    const OraParams& P = O.P;
    std::vector<M33> Rji;
    std::vector<V3> tji;
    for (size_t j = 0; j < neighbors.size(); j++) {
        M33 R21;
        V3 t21;
        relative_pose(cur.Rcw, cur.tcw, neighbors[j]->Rcw, neighbors[j]->tcw, R21, t21);  // PM.cc:638-647
        Rji.push_back(R21);
        tji.push_back(t21);
    }
    int cols = cur.cols, rows = cur.rows;
    float fx = cur.fx, fy = cur.fy, cx = cur.cx, cy = cur.cy;
    cur.depth_checked_ = cur.depth_map_;
    const size_t NN = neighbors.size();

#pragma omp parallel for schedule(dynamic) collapse(2) num_threads(g_threads)
    for (int py = 2; py < rows - 2; py++) {
        for (int px = 2; px < cols - 2; px++) {
            if (cur.depth_map_[(size_t)py * cols + px] < 0.000001) continue;
            float depthp = cur.depth_map_[(size_t)py * cols + px];
            int compatible_neighbor_keyframes_count = 0;
            std::vector<std::vector<depthHo>> compatible_pixels_by_frame;
            compatible_pixels_by_frame.reserve(NN);
            int num_compatible_pixels = 0;

            for (size_t j = 0; j < NN; j++) {
                const KeyFrame* pKFj = neighbors[j];
                M33 K = Kmat(pKFj->fx, pKFj->fy, pKFj->cx, pKFj->cy);
                V3 xp = {{(px - cx) / fx, (py - cy) / fy, 1.0f}};
                // temp = Rji*xp/depthp + tji  ->  gemm(Rji, xp, 1./depthp, tji, 1)   (operator/(MatExpr,double) multiplies alpha by 1./s;
                // hal::gemm32f then rounds that alpha to float -- inside gemm_small_33x31)
                V3 temp = cvl::gemm_small_33x31(Rji[j], xp, 1. / (double)depthp, &tji[j], 1.0);
                V3 Xj = cvl::gemm_small_33x31(K, temp, 1.0, nullptr, 0.0);
                // Xj = Xj/Xj(2)  ->  convertTo(alpha = 1./s): float multiply by (float)alpha
                float inv = (float)(1. / (double)Xj.v[2]);
                float xj = Xj.v[0] * inv;
                float yj = Xj.v[1] * inv;

                // Eq (12): Rji.row(2)*xp is the general gemm path (double accumulation)
                float denom1 = cvl::gemm_acc64_dot(&Rji[j].m[6], xp.v, 3, 1.0);
                float denom2 = depthp * tji[j].v[2];
                float depthj = depthp / (denom1 + denom2);

                std::vector<depthHo> compatible_pixels_J;
                // PM.cc:695 `xj<0 || xj>=cols-1 || yj<0 || yj>=rows-1`, written so that NaN coordinates
                // (Xj.z == 0) also count as outside instead of reaching (int)floor(NaN)
                if (!(xj >= 0 && xj < cols - 1 && yj >= 0 && yj < rows - 1)) {
                    compatible_pixels_by_frame.push_back(compatible_pixels_J);
                    continue;
                }
                int x0 = (int)std::floor(xj);
                int y0 = (int)std::floor(yj);
                int x1 = x0 + 1;
                int y1 = y0 + 1;
                const int ty[4] = {y0, y1, y0, y1};
                const int tx[4] = {x0, x0, x1, x1};
                // snapshot (H1): the neighbour's intra-key-frame map.  In-place mode (PM.cc:764, 793 write depth_map_ inside the
                // loop over key frames, PM.cc:262-307): a neighbour that was checked before reads as checked.
                const std::vector<float>& ndm =
                    (P.interkf_inplace && pKFj->interKF_depth_flag_ && pKFj != &cur) ? pKFj->depth_checked_ : pKFj->depth_map_;
                for (int t = 0; t < 4; ++t) {  // PM.cc:705-752, four unrolled copies
                    float d = at(ndm, *pKFj, ty[t], tx[t]);
                    float sigma = at(pKFj->depth_sigma_, *pKFj, ty[t], tx[t]);
                    if (d > 0.000001) {
                        float test = std::pow((double)(depthj - d), 2) / std::pow((double)sigma, 2);
                        if (test < P.chi2_inter) {
                            depthHo dHo;
                            dHo.depth = d;
                            dHo.sigma = sigma;
                            compatible_pixels_J.push_back(dHo);
                        }
                    }
                }
                if (compatible_pixels_J.size() >= 1) { compatible_neighbor_keyframes_count++; }
                compatible_pixels_by_frame.push_back(compatible_pixels_J);
                num_compatible_pixels += compatible_pixels_J.size();
            }

            if (compatible_neighbor_keyframes_count < P.lambdaN) {
                cur.depth_checked_[(size_t)py * cols + px] = 0.0;  // sigma untouched, PM.cc:764
            } else {
                // gauss newton smoothing, PM.cc:768-793.  -J.t()*r0 and J.t()*J are 1xn * nx1 gemms:
                // general path, sequential double accumulation.
                float xp[3] = {(px - cx) / fx, (py - cy) / fy, 1.0f};
                float dp = 1 / depthp;
                double sJr = 0, sJJ = 0;
                for (size_t j = 0; j < compatible_pixels_by_frame.size(); j++) {
                    std::vector<depthHo>& compatibleJ = compatible_pixels_by_frame[j];
                    for (size_t n = 0; n < compatibleJ.size(); n++) {
                        float rzxp = cvl::gemm_acc64_dot(&Rji[j].m[6], xp, 3, 1.0);
                        float djn = 1 / compatibleJ[n].depth;
                        float sigmajn = compatibleJ[n].sigma;
                        float d2sigma = djn * djn * sigmajn;
                        float Jn = -rzxp / d2sigma;
                        float rn = (djn - dp * rzxp - tji[j].v[2]) / d2sigma;
                        sJr += (double)Jn * (double)rn;
                        sJJ += (double)Jn * (double)Jn;
                    }
                }
                float Jtr0 = (float)(sJr * -1.0);
                float JtJ = (float)(sJJ * 1.0);
                float dpDelta = Jtr0 / JtJ;
                cur.depth_checked_[(size_t)py * cols + px] = 1 / (dp + dpDelta);
            }
        }
    }
    cur.interKF_depth_flag_ = true;
}

// PM.cc:337-367 UpdateSemiDensePointSet.  Twc*Pc is a 4x4 * 4x1 gemm: small path.
static void UpdateSemiDensePointSet(KeyFrame& kf) {
    const int rows = kf.rows, cols = kf.cols;
    const std::vector<float>& dm = kf.interKF_depth_flag_ ? kf.depth_checked_ : kf.depth_map_;
    kf.SemiDensePointSets_.assign((size_t)rows * cols * 3, 0.f);
#pragma omp parallel for schedule(dynamic) collapse(2) num_threads(g_threads)
    for (int y = 0 + 2; y < rows - 2; y++) {
        for (int x = 0 + 2; x < cols - 2; x++) {
            float* out = &kf.SemiDensePointSets_[(size_t)y * cols * 3 + 3 * x];
            if (dm[(size_t)y * cols + x] < 0.000001) {
                out[0] = out[1] = out[2] = 0.0;
                continue;
            }
            float inv_d = dm[(size_t)y * cols + x];
            float Z = 1 / inv_d;
            float X = Z * (x - kf.cx) / kf.fx;
            float Y = Z * (y - kf.cy) / kf.fy;
            V4 Pc = {{X, Y, Z, 1}};
            V4 pos = cvl::gemm_small_44x41(kf.Twc, Pc);
            out[0] = pos.v[0];
            out[1] = pos.v[1];
            out[2] = pos.v[2];
        }
    }
}

// PM.cc:370-383 StereoSearchConstraints (on the depth vector GetAllPointDepths would return;
// that accessor does not exist -- model: KeyFrame.cc:644-658)
static void StereoSearchConstraints(const std::vector<float>& orb_depths, float* min_depth, float* max_depth) {
    float sum = std::accumulate(orb_depths.begin(), orb_depths.end(), 0.0);
    float mean = sum / orb_depths.size();

    std::vector<float> diff(orb_depths.size());
    std::transform(orb_depths.begin(), orb_depths.end(), diff.begin(), [mean](float v) { return v - mean; });
    float variance = std::inner_product(diff.begin(), diff.end(), diff.begin(), 0.0) / orb_depths.size();
    float stdev = std::sqrt(variance);

    *max_depth = 1 / (mean + 2 * stdev);
    *min_depth = 1 / (mean - 2 * stdev);
}

static KeyFrame* find(Oracle* O, int id) {
    auto it = O->kfs.find(id);
    return it == O->kfs.end() ? nullptr : &it->second;
}

}  // namespace

extern "C" {

void ora_default_params(OraParams* p) {
    p->lambdaG = 8;
    p->lambdaL = 80;
    p->lambdaTheta = 45;
    p->lambdaN = 3;
    p->theta = 0.23;
    p->chi2_fuse = 5.99;
    p->chi2_inter = 3.84;
    p->slope_limit = 4;
    p->border = 2;
    p->interkf_inplace = 0;
}

void* ora_create(const OraParams* p) {
    Oracle* O = new Oracle();
    if (p) O->P = *p; else ora_default_params(&O->P);
    return O;
}
void ora_destroy(void* h) { delete (Oracle*)h; }
void ora_set_threads(int n) { g_threads = n < 1 ? 1 : n; }
int ora_max_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

int ora_add_keyframe(void* h, int kf_id, const uint8_t* gray, int w, int hgt, int stride, const float K[4], const float Tcw[12]) {
    Oracle* O = (Oracle*)h;
    KeyFrame& kf = O->kfs[kf_id];
    kf = KeyFrame();
    kf.rows = hgt;
    kf.cols = w;
    kf.im_.resize((size_t)w * hgt);
    for (int y = 0; y < hgt; ++y) std::memcpy(&kf.im_[(size_t)y * w], gray + (size_t)y * stride, w);
    kf.fx = K[0]; kf.fy = K[1]; kf.cx = K[2]; kf.cy = K[3];
    set_pose(kf, Tcw);
    front_end(kf);
    return 0;
}

int ora_set_pose(void* h, int kf_id, const float Tcw[12]) {
    KeyFrame* kf = find((Oracle*)h, kf_id);
    if (!kf) return -1;
    set_pose(*kf, Tcw);
    return 0;
}

int ora_semidense(void* h, int kf_id, const int* nbr_ids, const float* nbr_rot_deg, int n, float min_depth, float max_depth) {
    Oracle* O = (Oracle*)h;
    KeyFrame* kf = find(O, kf_id);
    if (!kf) return -1;
    std::vector<const KeyFrame*> nb;
    std::vector<float> rot;
    for (int j = 0; j < n; ++j) {
        KeyFrame* k2 = find(O, nbr_ids[j]);
        if (!k2) return -1;
        nb.push_back(k2);
        rot.push_back(nbr_rot_deg ? nbr_rot_deg[j] : 0.f);
    }
    SemiDenseLoop(*O, *kf, nb, rot, min_depth, max_depth);
    return 0;
}

int ora_interkf(void* h, int kf_id, const int* nbr_ids, int n) {
    Oracle* O = (Oracle*)h;
    KeyFrame* kf = find(O, kf_id);
    if (!kf || !kf->semidense_flag_) return -1;
    std::vector<const KeyFrame*> nb;
    for (int j = 0; j < n; ++j) {
        KeyFrame* k2 = find(O, nbr_ids[j]);
        if (!k2 || !k2->semidense_flag_) return -1;
        nb.push_back(k2);
    }
    InterKeyFrameDepthChecking(*O, *kf, nb);
    return 0;
}

int ora_pointset(void* h, int kf_id) {
    KeyFrame* kf = find((Oracle*)h, kf_id);
    if (!kf || !kf->semidense_flag_) return -1;
    UpdateSemiDensePointSet(*kf);
    return 0;
}

// The selection of the .obj dump, PM.cc:114-127 (whole image, raster order: skip if depth_sigma_ > 0.01,
// keep if depth_map_ > 0.000001), with the threshold as a parameter, returned as a list -- the shape in which the
// CARV transcript ingests extra points per key frame (SFMTranscriptInterface_ORBSLAM.cpp:319-374).
// Build-defined extension (no counterpart in the reference): if max_points > 0 and more pixels qualify, the
// max_points with the smallest sigma are kept (total order on the float bits, ties: raster order), in raster order.
int ora_confident_points(void* h, int kf_id, double max_sigma, int max_points, int capacity, float* xyz, float* sigma, uint32_t* pixel) {
    KeyFrame* kf = find((Oracle*)h, kf_id);
    if (!kf || !kf->semidense_flag_) return -1;
    UpdateSemiDensePointSet(*kf);  // current pose, current depth map (PM.cc:321-334 refreshes it the same way)
    const int rows = kf->rows, cols = kf->cols;
    const std::vector<float>& dm = kf->interKF_depth_flag_ ? kf->depth_checked_ : kf->depth_map_;
    std::vector<uint32_t> sel;
    for (int y = 0; y < rows; y++)
        for (int x = 0; x < cols; x++) {
            const size_t o = (size_t)y * cols + x;
            if (kf->depth_sigma_[o] > max_sigma) continue;
            if (dm[o] > 0.000001) sel.push_back((uint32_t)o);
        }
    if (max_points > 0 && (int)sel.size() > max_points) {
        auto key = [&](uint32_t o) {
            uint32_t b;
            std::memcpy(&b, &kf->depth_sigma_[o], 4);
            return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
        };
        std::stable_sort(sel.begin(), sel.end(), [&](uint32_t a, uint32_t b) { return key(a) < key(b); });
        sel.resize(max_points);
        std::sort(sel.begin(), sel.end());
    }
    const int n = (int)sel.size();
    for (int i = 0; i < n && i < capacity; ++i) {
        const uint32_t o = sel[i];
        const int y = (int)(o / cols), x = (int)(o % cols);
        if (xyz) std::memcpy(xyz + 3 * (size_t)i, &kf->SemiDensePointSets_[(size_t)y * cols * 3 + 3 * x], 3 * sizeof(float));
        if (sigma) sigma[i] = kf->depth_sigma_[o];
        if (pixel) pixel[i] = ((uint32_t)y << 16) | (uint32_t)x;
    }
    return n;
}

// test hook: install caller-made (depth, sigma) maps as the intra-KF result of a key frame (the counterpart of
// sd_import_maps), so crafted planes can be pushed through the later stages
int ora_set_maps(void* h, int kf_id, const float* depth, const float* sigma) {
    KeyFrame* kf = find((Oracle*)h, kf_id);
    if (!kf) return -1;
    const size_t n = (size_t)kf->rows * kf->cols;
    kf->depth_map_.assign(depth, depth + n);
    kf->depth_sigma_.assign(sigma, sigma + n);
    kf->semidense_flag_ = true;
    kf->interKF_depth_flag_ = false;
    kf->depth_checked_.clear();
    return 0;
}

int ora_get_plane(void* h, int kf_id, int what, float* out) {
    KeyFrame* kf = find((Oracle*)h, kf_id);
    if (!kf) return -1;
    const std::vector<float>* src = nullptr;
    switch (what) {
        case ORA_GRAD: src = &kf->GradImg; break;
        case ORA_THETA: src = &kf->GradTheta; break;
        case ORA_DEPTH: src = &kf->depth_map_; break;
        case ORA_SIGMA: src = &kf->depth_sigma_; break;
        case ORA_DEPTH_CHECKED: src = &kf->depth_checked_; break;
        case ORA_POINTS: src = &kf->SemiDensePointSets_; break;
        default: return -2;
    }
    if (src->empty()) return -3;
    std::memcpy(out, src->data(), src->size() * sizeof(float));
    return 0;
}

float ora_get_istddev(void* h, int kf_id) {
    KeyFrame* kf = find((Oracle*)h, kf_id);
    return kf ? kf->I_stddev : -1.f;
}

float ora_fast_atan2(float y, float x) { return cvl::fastAtan2(y, x); }
float ora_bilinear_u8(const uint8_t* img, int w, int h, float y, float x) { return bilinear<uint8_t>(img, w, h, y, x); }
float ora_bilinear_f32(const float* img, int w, int h, float y, float x) { return bilinear<float>(img, w, h, y, x); }

int ora_chi_test(float a, float b, float sa, float sb, double thr) {
    OraParams P;
    ora_default_params(&P);
    P.chi2_fuse = thr;
    return ChiTest(P, a, b, sa, sb) ? 1 : 0;
}

void ora_get_fusion(const float* d, const float* s, int n, float* depth, float* sigma, float* min_sigma) {
    std::vector<depthHo> v(n);
    for (int i = 0; i < n; ++i) { v[i].depth = d[i]; v[i].sigma = s[i]; }
    depthHo out;
    GetFusion(v.data(), v.size(), out, min_sigma);
    *depth = out.depth;
    *sigma = out.sigma;
}

void ora_get_fusion_pairs(const float* d, const float* s, int n, float* depth, float* sigma) {
    std::vector<std::pair<float, float>> v(n);
    for (int i = 0; i < n; ++i) v[i] = {d[i], s[i]};
    GetFusion(v.data(), v.size(), *depth, *sigma);
}

int ora_hypothesis_fusion(const OraParams* p, const float* d, const float* s, int n, float* depth, float* sigma) {
    OraParams P;
    if (p) P = *p; else ora_default_params(&P);
    std::vector<depthHo> v(n);
    for (int i = 0; i < n; ++i) { v[i].depth = d[i]; v[i].sigma = s[i]; v[i].supported = true; }
    depthHo out;
    InverseDepthHypothesisFusion(P, v, out);
    *depth = out.depth;
    *sigma = out.sigma;
    return out.supported ? 1 : 0;
}

void ora_stereo_search_constraints(const float* depths, int n, float* min_depth, float* max_depth) {
    std::vector<float> v(depths, depths + n);
    StereoSearchConstraints(v, min_depth, max_depth);
}

// PM.cc:467-484 GetRotInPlane + the median at PM.cc:172-177.  Map points are identified by id
// (the reference compares MapPoint pointers); id < 0 stands for a null pointer.
float ora_rot_in_plane_median(const long long* mp1, const float* ang1, int n1, const long long* mp2, const float* ang2, int n2) {
    std::vector<float> rotInPlane;
    for (int idx1 = 0; idx1 < n1; idx1++) {
        if (mp1[idx1] >= 0) {
            for (int idx2 = 0; idx2 < n2; idx2++) {
                if (mp2[idx2] == mp1[idx1]) {
                    float angle1 = ang1[idx1];
                    float angle2 = ang2[idx2];
                    if (angle1 < 0 || angle2 < 0) continue;
                    rotInPlane.push_back(angle2 - angle1);
                }
            }
        }
    }
    std::sort(rotInPlane.begin(), rotInPlane.end());
    float medianRot = 0;
    if (rotInPlane.size() > 0) medianRot = rotInPlane[(rotInPlane.size() - 1) / 2];
    return medianRot;
}

static void make_kf(KeyFrame& kf, const float K[4], const float Tcw[12]) {
    kf.fx = K[0]; kf.fy = K[1]; kf.cx = K[2]; kf.cy = K[3];
    set_pose(kf, Tcw);
}

void ora_pair_geometry(const float K1[4], const float Tcw1[12], const float K2[4], const float Tcw2[12], float F12[9],
                       float R21[9], float t21[3]) {
    KeyFrame a, b;
    make_kf(a, K1, Tcw1);
    make_kf(b, K2, Tcw2);
    PairGeom g = pair_geometry(a, b);
    std::memcpy(F12, g.F12.m, sizeof(g.F12.m));
    std::memcpy(R21, g.R21.m, sizeof(g.R21.m));
    std::memcpy(t21, g.t21.v, sizeof(g.t21.v));
}

void ora_search_range(const float K1[4], const float R21[9], const float t21[3], int cols, int px, int py, float mind,
                      float maxd, float* umin, float* umax) {
    M33 R;
    V3 t;
    std::memcpy(R.m, R21, sizeof(R.m));
    std::memcpy(t.v, t21, sizeof(t.v));
    GetSearchRange(*umin, *umax, px, py, mind, maxd, K1[0], K1[1], K1[2], K1[3], cols, R, t);
}

/* PM.cc:486-547 / 549-596 on caller planes (w*h floats, updated in place) */
void ora_intra_check_planes(const OraParams* p, float* depth, float* sigma, int w, int h) {
    OraParams P;
    if (p) P = *p; else ora_default_params(&P);
    std::vector<float> d(depth, depth + (size_t)w * h), s(sigma, sigma + (size_t)w * h);
    IntraKeyFrameDepthChecking(P, d, s, h, w);
    std::memcpy(depth, d.data(), d.size() * sizeof(float));
    std::memcpy(sigma, s.data(), s.size() * sizeof(float));
}

void ora_intra_grow_planes(const OraParams* p, float* depth, float* sigma, const float* grad, int w, int h) {
    OraParams P;
    if (p) P = *p; else ora_default_params(&P);
    std::vector<float> d(depth, depth + (size_t)w * h), s(sigma, sigma + (size_t)w * h), g(grad, grad + (size_t)w * h);
    IntraKeyFrameDepthGrowing(P, d, s, g, h, w);
    std::memcpy(depth, d.data(), d.size() * sizeof(float));
    std::memcpy(sigma, s.data(), s.size() * sizeof(float));
}

float ora_pixel_depth(const float K1[4], const float R21[9], const float t21[3], float uj, int px, int py) {
    M33 R;
    V3 t;
    std::memcpy(R.m, R21, sizeof(R.m));
    std::memcpy(t.v, t21, sizeof(t.v));
    float p = 0;
    GetPixelDepth(uj, px, py, K1[0], K1[1], K1[2], K1[3], R, t, p);
    return p;
}

int ora_epipolar_search(void* h, int kf1, int kf2, int x, int y, float min_depth, float max_depth, float rot, float* depth,
                        float* sigma, float* best_u, float* best_v) {
    Oracle* O = (Oracle*)h;
    KeyFrame* a = find(O, kf1);
    KeyFrame* b = find(O, kf2);
    if (!a || !b) return -1;
    PairGeom g = pair_geometry(*a, *b);
    depthHo dh;
    float bu = 0, bv = 0;
    EpipolarSearch(O->P, *a, *b, x, y, (float)a->im_[(size_t)y * a->cols + x], min_depth, max_depth, &dh, g, bu, bv,
                   a->GradTheta[(size_t)y * a->cols + x], rot);
    *depth = dh.depth;
    *sigma = dh.sigma;
    *best_u = bu;
    *best_v = bv;
    return dh.supported ? 1 : 0;
}

void ora_cv_gemm_small_33x31(const float A[9], const float b[3], double alpha, const float* c, double beta, float d[3]) {
    M33 a; V3 bb, cc;
    std::memcpy(a.m, A, 36); std::memcpy(bb.v, b, 12);
    if (c) std::memcpy(cc.v, c, 12);
    V3 r = cvl::gemm_small_33x31(a, bb, alpha, c ? &cc : nullptr, beta);
    std::memcpy(d, r.v, 12);
}
void ora_cv_gemm_small_33x33(const float A[9], const float B[9], float D[9]) {
    M33 a, b;
    std::memcpy(a.m, A, 36); std::memcpy(b.m, B, 36);
    M33 r = cvl::gemm_small_33x33(a, b, 1.0);
    std::memcpy(D, r.m, 36);
}
void ora_cv_gemm_small_44x41(const float A[16], const float b[4], float d[4]) {
    M44 a; V4 bb;
    std::memcpy(a.m, A, 64); std::memcpy(bb.v, b, 16);
    V4 r = cvl::gemm_small_44x41(a, bb);
    std::memcpy(d, r.v, 16);
}
void ora_cv_gemm_acc64_ABt(const float A[9], const float B[9], double alpha, float D[9]) {
    M33 a, b;
    std::memcpy(a.m, A, 36); std::memcpy(b.m, B, 36);
    M33 r = cvl::gemm_acc64_ABt(a, b, alpha);
    std::memcpy(D, r.m, 36);
}
float ora_cv_gemm_acc64_dot(const float* a, const float* b, int n, double alpha) { return cvl::gemm_acc64_dot(a, b, n, alpha); }
int ora_cv_invert33(const float A[9], float D[9]) {
    M33 a, d;
    std::memcpy(a.m, A, 36);
    bool ok = cvl::invert33(a, d);
    std::memcpy(D, d.m, 36);
    return ok;
}
int ora_cv_solve_lu33(const float A[9], const float B[9], float X[9]) {
    M33 a, b, x;
    std::memcpy(a.m, A, 36); std::memcpy(b.m, B, 36);
    bool ok = cvl::solve_lu33(a, b, x);
    std::memcpy(X, x.m, 36);
    return ok;
}

}  // extern "C"
