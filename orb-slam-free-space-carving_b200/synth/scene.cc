// Synthetic key-frame generator for tests and bench.py.  No graphics API:
// a textured procedural height field is ray-cast on the CPU.  Deterministic for a given seed.
//
//   scene      z(X,Y) = 1 + 0.15 * fbm(X,Y)   (mean depth 1 scene unit -- required by the reference's
//              GetSearchRange, which multiplies the ray by INVERSE depth, PM.cc:894-895 / SURVEY H3)
//   texture    band-limited value noise + sinusoids, u8, +-1 LSB seeded dither
//   trajectory C_k = (0.04 k, 0.01 sin(0.3 k), 0), yaw 0.5 deg * sin(0.2 k), roll 0.3 deg * cos(0.1 k)
//   sparse     stand-ins for the ORB map-point depths that StereoSearchConstraints consumes
//              (PM.cc:371 GetAllPointDepths): n seeded pixels with ground-truth depth
//
// This is input generation, not part of the hot path and not part of the oracle.
#include <cmath>
#include <cstdint>
#include <cstring>
#ifdef _OPENMP
#include <omp.h>
#endif

extern "C" {
typedef struct SynthConfig {
    int w, h;
    float fx, fy, cx, cy;
    uint32_t seed;
} SynthConfig;
}

#ifndef TEX_A1
#define TEX_A1 46.0
#define TEX_S1 15.0
#define TEX_A2 34.0
#define TEX_S2 6.5
#define TEX_A3 12.0
#endif
#ifndef HEIGHT_FREQ
#define HEIGHT_FREQ 2.2
#endif

namespace {

inline uint32_t hash_u32(uint32_t x) {
    x ^= x >> 16; x *= 0x7feb352dU;
    x ^= x >> 15; x *= 0x846ca68bU;
    x ^= x >> 16;
    return x;
}
inline uint32_t hash3(int32_t a, int32_t b, uint32_t c) {
    return hash_u32((uint32_t)a * 0x9e3779b1U ^ hash_u32((uint32_t)b * 0x85ebca77U ^ hash_u32(c)));
}
inline double lattice(int32_t ix, int32_t iy, uint32_t seed) {  // [-1, 1)
    return (double)hash3(ix, iy, seed) * (2.0 / 4294967296.0) - 1.0;
}
inline double fade(double t) { return t * t * t * (t * (t * 6 - 15) + 10); }

// value noise with quintic fade, C2-continuous, range [-1,1]
double vnoise(double x, double y, uint32_t seed) {
    double fx = std::floor(x), fy = std::floor(y);
    int32_t ix = (int32_t)fx, iy = (int32_t)fy;
    double tx = fade(x - fx), ty = fade(y - fy);
    double v00 = lattice(ix, iy, seed), v10 = lattice(ix + 1, iy, seed);
    double v01 = lattice(ix, iy + 1, seed), v11 = lattice(ix + 1, iy + 1, seed);
    double a = v00 + (v10 - v00) * tx, b = v01 + (v11 - v01) * tx;
    return a + (b - a) * ty;
}

// two octaves, amplitudes 1 and 0.4 (not normalised): range (-1.4, 1.4)
double fbm(double x, double y, uint32_t seed) {
    return vnoise(x, y, seed) + 0.4 * vnoise(x * 2.0 + 17.3, y * 2.0 - 9.1, seed + 101u);
}

// lattice spacing 1/HEIGHT_FREQ scene units: a 720p view (1.16 x 0.65 units at depth 1) spans several cells
inline double height(double X, double Y, uint32_t seed) {
    return 1.0 + 0.15 * fbm(X * HEIGHT_FREQ, Y * HEIGHT_FREQ, seed ^ 0xA511E9B3u);
}

// grey level before dither.  (U,V) = fx*(X,Y): texture features are sized in PIXELS AT DEPTH 1 so that
// every resolution sees the same gradient statistics (~30 % of pixels with Sobel/8 magnitude >= 8).
inline double texture(double U, double V, uint32_t seed) {
    double t = 128.0;
    t += TEX_A1 * vnoise(U / TEX_S1, V / TEX_S1, seed ^ 0x1234567u);
    t += TEX_A2 * vnoise(U / TEX_S2 + 3.7, V / TEX_S2 - 1.9, seed ^ 0x89abcdeu);
    t += TEX_A3 * std::sin(U * 0.23 + 1.3 * std::sin(V * 0.11)) * std::cos(V * 0.19);
    return t;
}

struct Pose {
    double C[3];
    double R[9];  // R_wc (camera -> world), row major
    double roll_deg;
};

Pose pose_of(int k) {
    Pose p;
    p.C[0] = 0.04 * k;
    p.C[1] = 0.01 * std::sin(0.3 * k);
    p.C[2] = 0.0;
    const double d2r = 3.14159265358979323846 / 180.0;
    double yaw = 0.5 * std::sin(0.2 * k) * d2r;
    p.roll_deg = 0.3 * std::cos(0.1 * k);
    double roll = p.roll_deg * d2r;
    double cy = std::cos(yaw), sy = std::sin(yaw), cr = std::cos(roll), sr = std::sin(roll);
    // R_wc = Ry(yaw) * Rz(roll); X right, Y down, Z forward
    double Ry[9] = {cy, 0, sy, 0, 1, 0, -sy, 0, cy};
    double Rz[9] = {cr, -sr, 0, sr, cr, 0, 0, 0, 1};
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) {
            double s = 0;
            for (int q = 0; q < 3; ++q) s += Ry[3 * i + q] * Rz[3 * q + j];
            p.R[3 * i + j] = s;
        }
    return p;
}

// camera-frame depth t of the FIRST surface point along the ray through pixel (x, y):
// march from the near bound in small steps until the ray is below the surface, then refine.
// t0 > 0 is a warm start (the neighbouring pixel's depth): the march then starts just in front of it.
double cast(const SynthConfig& c, const Pose& p, double x, double y, double t0) {
    double dc[3] = {(x - c.cx) / c.fx, (y - c.cy) / c.fy, 1.0};
    double dw[3];
    for (int i = 0; i < 3; ++i) dw[i] = p.R[3 * i] * dc[0] + p.R[3 * i + 1] * dc[1] + p.R[3 * i + 2] * dc[2];
    auto f = [&](double t) { return p.C[2] + t * dw[2] - height(p.C[0] + t * dw[0], p.C[1] + t * dw[1], c.seed); };
    const double tmin = 0.70, tmax = 1.40, step = 0.01;
    double lo = t0 > 0 ? t0 - 2 * step : tmin;
    if (lo < tmin) lo = tmin;
    while (lo > tmin && f(lo) >= 0) lo -= 4 * step;  // warm start landed behind the surface: back off
    double hi = lo + step;
    while (hi < tmax && f(hi) < 0) { lo = hi; hi += step; }
    // bracket [lo, hi] is one step wide: safeguarded regula falsi (Illinois), then the bracket midpoint
    double flo = f(lo), fhi = f(hi);
    if (!(flo < 0 && fhi >= 0)) return 0.5 * (lo + hi);
    for (int it = 0; it < 6 && hi - lo > 1e-9; ++it) {
        double mid = lo - flo * (hi - lo) / (fhi - flo);
        double fm = f(mid);
        if (fm < 0) { lo = mid; flo = fm; fhi *= 0.5; } else { hi = mid; fhi = fm; flo *= 0.5; }
    }
    return lo - flo * (hi - lo) / (fhi - flo);
}

}  // namespace

extern "C" {

// Tcw (3x4 row major, world -> camera) and the ground-truth roll of key frame k
void synth_pose(int k, float Tcw[12], float* roll_deg) {
    Pose p = pose_of(k);
    for (int i = 0; i < 3; ++i) {
        double t = 0;
        for (int j = 0; j < 3; ++j) {
            Tcw[4 * i + j] = (float)p.R[3 * j + i];  // Rcw = R_wc^T
            t -= p.R[3 * j + i] * p.C[j];
        }
        Tcw[4 * i + 3] = (float)t;
    }
    if (roll_deg) *roll_deg = (float)p.roll_deg;
}

// gray: w*h u8; gt_depth (optional): w*h camera-frame depth
void synth_render(const SynthConfig* cfg, int k, uint8_t* gray, float* gt_depth) {
    const SynthConfig c = *cfg;
    Pose p = pose_of(k);
#pragma omp parallel for schedule(dynamic, 4)
    for (int y = 0; y < c.h; ++y) {
        double t = -1.0;
        for (int x = 0; x < c.w; ++x) {
            t = cast(c, p, x, y, x == 0 ? -1.0 : t);
            double dc[3] = {(x - c.cx) / c.fx, (y - c.cy) / c.fy, 1.0};
            double X = p.C[0], Y = p.C[1];
            for (int j = 0; j < 3; ++j) {
                X += t * p.R[j] * dc[j];
                Y += t * p.R[3 + j] * dc[j];
            }
            double g = texture(X * c.fx, Y * c.fx, c.seed);
            double dither = (double)hash3(x + 7919 * k, y, c.seed ^ 0x5bd1e995u) * (2.0 / 4294967296.0) - 1.0;
            double v = std::floor(g + dither + 0.5);
            v = v < 0 ? 0 : (v > 255 ? 255 : v);
            gray[(size_t)y * c.w + x] = (uint8_t)v;
            if (gt_depth) gt_depth[(size_t)y * c.w + x] = (float)t;
        }
    }
}

// n seeded pixels of key frame k with ground-truth depth (ORB map-point stand-ins)
void synth_sparse_depths(const SynthConfig* cfg, int k, int n, float* depths) {
    const SynthConfig c = *cfg;
    Pose p = pose_of(k);
    for (int i = 0; i < n; ++i) {
        uint32_t hx = hash3(i, k, c.seed ^ 0x777u), hy = hash3(i, k, c.seed ^ 0x999u);
        double x = 2 + (double)hx / 4294967296.0 * (c.w - 4);
        double y = 2 + (double)hy / 4294967296.0 * (c.h - 4);
        depths[i] = (float)cast(c, p, std::floor(x), std::floor(y), -1.0);
    }
}

// the n nearest key-frame indices of k inside [0, nkf): k-1, k+1, k-2, k+2, ... (out-of-range skipped)
int synth_neighbours(int k, int nkf, int n, int* ids) {
    int got = 0;
    for (int d = 1; d < nkf && got < n; ++d) {
        if (k - d >= 0 && got < n) ids[got++] = k - d;
        if (k + d < nkf && got < n) ids[got++] = k + d;
    }
    return got;
}

// in-plane rotation a feature undergoes from key frame k1 to k2, degrees (GetRotInPlane stand-in,
// PM.cc:467-484): image content rotates opposite to the camera roll
float synth_rot_in_plane(int k1, int k2) { return (float)(-(pose_of(k2).roll_deg - pose_of(k1).roll_deg)); }

}  // extern "C"
