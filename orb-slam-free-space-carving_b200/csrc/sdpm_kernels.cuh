// CUDA kernels (sm_100a) of the semi-dense probability-mapping path.
//
//   K0 k0_gradient     image -> GradImg / GradTheta / gate plane / active-pixel list / I_stddev sums
//   K1 k1_search       per active pixel: EpipolarSearch in N neighbours + InverseDepthHypothesisFusion
//   K2 k2_intra_check  IntraKeyFrameDepthChecking           (3x3 chi-square stencil)
//   K3 k3_intra_grow   IntraKeyFrameDepthGrowing            (3x3 chi-square stencil)
//   K4 k4_inter_check  InterKeyFrameDepthChecking           (projection into N neighbours + Gauss-Newton)
//   K5 k5_pointset     UpdateSemiDensePointSet              (back-projection to world)
//
// Reference: /root/reference/src/Modeler/ProbabilityMapping.cc ("PM.cc"); each device function cites
// the lines it implements.  Numerics: this file must be compiled with -fmad=false -prec-div=true
// -prec-sqrt=true -ftz=false: every float decision in PM.cc (ceil/floor of the search range, the
// (int) row truncation, the angle folds, the chi-square compares) is reproduced with the same
// IEEE operations in the same order, and `double` is used exactly where the reference promotes
// (THETA is a double literal, std::pow, and OpenCV's gemm internals -- see pair_geometry.h).
// Explicit __fma_rn(double) is used only where the product of two floats is exact in double, so
// the fused and the unfused result are the same number.
//
// Plane reads use clamp-to-edge addressing wherever the reference can read out of bounds.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/sdpm.h"

namespace sdpm {

struct KfMeta {  // device-resident per key frame
    unsigned long long sum;    // sum of grey levels
    unsigned long long sumsq;  // sum of squared grey levels
    float istddev;             // KeyFrame::I_stddev
    int active_count;          // entries in the active-pixel list
};

struct NbrConst {  // per neighbour, warp-uniform, lives in kernel parameter (constant) space
    float F[9];    // F12, PM.cc:972-987
    float R[9];    // R21, PM.cc:890
    float t[3];    // t21, PM.cc:891
    float rot;     // median in-plane rotation, degrees, PM.cc:170-179
    const uint8_t* img;
    const float* grad;
    const float* gate;  // GradTheta where GradImg >= lambdaG, else -1
    const KfMeta* meta;
};

struct SearchArgs {
    int W, H, pitch_f, pitch_u8, n;
    int border, lambdaN;
    float fx, fy, cx, cy;
    float mind, maxd;  // StereoSearchConstraints interval (inverse depth; mind is the larger), PM.cc:381-382
    float lambdaL, lambdaTheta, slope;
    float chi_fuse;  // smallest float >= 5.99: `chi < 5.99` (float vs double literal) == `chi < chi_fuse`
    double inv_THETA, THETA;  // 1/THETA and THETA (PM.h:47) as doubles
    const uint8_t* img;
    const float* grad;
    const float* gate;
    const float* theta;
    const uint32_t* list;
    const KfMeta* meta;
    int count_override;  // >= 0: use instead of meta->active_count (probes)
    float* depth;
    float* sigma;
    float* probe_out;  // nullable; [entry][4] = {rho, sigma, best uj, supported} of neighbour 0
    NbrConst nb[SD_MAX_NEIGHBOURS];
};

struct StencilArgs {
    int W, H, pitch_f, border;
    float lambdaG, chi_fuse;
    const float* depth_in;
    const float* sigma_in;
    const float* grad;
    float* depth_out;
    float* sigma_out;
};

struct InterNbr {
    float R[9];  // Rji, PM.cc:643
    float t[3];  // tji, PM.cc:644
    float fx, fy, cx, cy;  // K of neighbour j, PM.cc:675
    const float* depth;
    const float* sigma;
};

struct InterArgs {
    int W, H, pitch_f, n, lambdaN;
    float fx, fy, cx, cy;
    float chi_inter;  // smallest float >= 3.84
    const float* depth_in;
    float* depth_out;
    InterNbr nb[SD_MAX_NEIGHBOURS];
};

struct PointArgs {
    int W, H, pitch_f;
    float fx, fy, cx, cy;
    float Twc[16];
    const float* depth;
    float* points;  // H x 3W, row pitch 3*W floats (tight)
};

// `d > 0.000001` with a float d and a double literal == `d > SD_EPS_F` (largest float <= 1e-6);
// `d < 0.000001` == `d <= SD_EPS_F`.
#define SD_EPS_F 9.99999997475242707878e-07f

__device__ __forceinline__ int clampi(int v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : v); }

// cv::fastAtan2 scalar, degrees in [0, 360]
__device__ __forceinline__ float fast_atan2_deg(float y, float x) {
    const float p1 = 0.9997878412794807f * (float)(180 / 3.14159265358979323846);
    const float p3 = -0.3258083974640975f * (float)(180 / 3.14159265358979323846);
    const float p5 = 0.1555786518463281f * (float)(180 / 3.14159265358979323846);
    const float p7 = -0.04432655554792128f * (float)(180 / 3.14159265358979323846);
    float ax = fabsf(x), ay = fabsf(y);
    float a, c, c2;
    if (ax >= ay) {
        c = ay / (ax + 2.220446049250313e-16f);
        c2 = c * c;
        a = (((p7 * c2 + p5) * c2 + p3) * c2 + p1) * c;
    } else {
        c = ax / (ay + 2.220446049250313e-16f);
        c2 = c * c;
        a = 90.f - (((p7 * c2 + p5) * c2 + p3) * c2 + p1) * c;
    }
    if (x < 0) a = 180.f - a;
    if (y < 0) a = 360.f - a;
    return a;
}

// ------------------------------------------------------------------------------------------------
// K0: gradient front end (build-defined) + active list + I_stddev sums.
// 3x3 Sobel/8 with replicate border; GradImg = sqrtf(gx^2+gy^2); GradTheta = fastAtan2(gy,gx);
// gate = GradImg >= lambdaG ? GradTheta : -1.  Block 32x8 threads, 4 pixels per thread (128x8 tile).
// ------------------------------------------------------------------------------------------------
constexpr int K0_TX = 32, K0_TY = 8, K0_PX = 4;
constexpr int K0_TW = K0_TX * K0_PX;  // 128

__global__ void __launch_bounds__(K0_TX* K0_TY) k0_gradient(const uint8_t* __restrict__ img, int pitch_u8, int W, int H,
                                                            float* __restrict__ grad, float* __restrict__ theta,
                                                            float* __restrict__ gate, int pitch_f,
                                                            uint32_t* __restrict__ list, KfMeta* __restrict__ meta,
                                                            float lambdaG, int border) {
    __shared__ uint8_t tile[K0_TY + 2][K0_TW + 8];  // column c <-> x0 - 1 + c
    __shared__ int warp_cnt[K0_TY];
    __shared__ int block_base;
    __shared__ unsigned long long red[2][K0_TY];

    const int tx = threadIdx.x, ty = threadIdx.y;
    const int tid = ty * K0_TX + tx;
    const int x0 = blockIdx.x * K0_TW, y0 = blockIdx.y * K0_TY;

    for (int i = tid; i < (K0_TY + 2) * (K0_TW + 2); i += K0_TX * K0_TY) {
        int r = i / (K0_TW + 2), c = i - r * (K0_TW + 2);
        int yy = clampi(y0 - 1 + r, 0, H - 1), xx = clampi(x0 - 1 + c, 0, W - 1);
        tile[r][c] = __ldg(img + (size_t)yy * pitch_u8 + xx);
    }
    __syncthreads();

    const int y = y0 + ty;
    const int xb = x0 + tx * K0_PX;
    float g[K0_PX], th[K0_PX], gt[K0_PX];
    unsigned long long s1 = 0, s2 = 0;
    uint32_t act[K0_PX];
    int nact = 0;
#pragma unroll
    for (int i = 0; i < K0_PX; ++i) {
        const int c = tx * K0_PX + i + 1, r = ty + 1;
        int p00 = tile[r - 1][c - 1], p01 = tile[r - 1][c], p02 = tile[r - 1][c + 1];
        int p10 = tile[r][c - 1], p11 = tile[r][c], p12 = tile[r][c + 1];
        int p20 = tile[r + 1][c - 1], p21 = tile[r + 1][c], p22 = tile[r + 1][c + 1];
        int sx = (p02 + 2 * p12 + p22) - (p00 + 2 * p10 + p20);
        int sy = (p20 + 2 * p21 + p22) - (p00 + 2 * p01 + p02);
        float gx = (float)sx * 0.125f, gy = (float)sy * 0.125f;
        g[i] = sqrtf(gx * gx + gy * gy);
        th[i] = fast_atan2_deg(gy, gx);
        const bool on = g[i] >= lambdaG;
        gt[i] = on ? th[i] : -1.0f;
        const int x = xb + i;
        if (x < W && y < H) {
            s1 += (unsigned)p11;
            s2 += (unsigned)(p11 * p11);
            if (on && x >= border && x < W - border && y >= border && y < H - border) act[nact++] = ((uint32_t)y << 16) | (uint32_t)x;
        }
    }
    if (y < H) {
        const size_t o = (size_t)y * pitch_f + xb;
        if (xb + K0_PX <= W) {  // pitch_f and xb are multiples of 4: aligned float4 stores
            *reinterpret_cast<float4*>(grad + o) = make_float4(g[0], g[1], g[2], g[3]);
            *reinterpret_cast<float4*>(theta + o) = make_float4(th[0], th[1], th[2], th[3]);
            *reinterpret_cast<float4*>(gate + o) = make_float4(gt[0], gt[1], gt[2], gt[3]);
        } else {
            for (int i = 0; i < K0_PX; ++i)
                if (xb + i < W) { grad[o + i] = g[i]; theta[o + i] = th[i]; gate[o + i] = gt[i]; }
        }
    }

    // block-wide compaction of the active pixels: warp scan, then one atomicAdd per block
    const unsigned full = 0xffffffffu;
    int incl = nact;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        int v = __shfl_up_sync(full, incl, d);
        if (tx >= d) incl += v;
    }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
        s1 += __shfl_down_sync(full, s1, d);
        s2 += __shfl_down_sync(full, s2, d);
    }
    if (tx == 31) warp_cnt[ty] = incl;
    if (tx == 0) { red[0][ty] = s1; red[1][ty] = s2; }
    __syncthreads();
    if (tid == 0) {
        int tot = 0;
        unsigned long long a = 0, b = 0;
        for (int w = 0; w < K0_TY; ++w) {
            int c = warp_cnt[w];
            warp_cnt[w] = tot;
            tot += c;
            a += red[0][w];
            b += red[1][w];
        }
        block_base = tot ? atomicAdd(&meta->active_count, tot) : 0;
        atomicAdd(&meta->sum, a);
        atomicAdd(&meta->sumsq, b);
    }
    __syncthreads();
    int pos = block_base + warp_cnt[ty] + incl - nact;
    for (int i = 0; i < nact; ++i) list[pos + i] = act[i];
}

// I_stddev = population standard deviation of the u8 image, double arithmetic (== cv::meanStdDev)
__global__ void k0_finalize(KfMeta* meta, int W, int H) {
    double n = (double)W * (double)H;
    double mean = (double)meta->sum / n;
    double var = (double)meta->sumsq / n - mean * mean;
    meta->istddev = (float)sqrt(var > 0.0 ? var : 0.0);
}

// ------------------------------------------------------------------------------------------------
// K1 device functions
// ------------------------------------------------------------------------------------------------

// bilinear<T>(img, y, x) of PM.cc:40-59 at an INTEGER column x = uj (every call site in
// EpipolarSearch passes an int): x0_weight == 1 and x1_weight == 0 exactly, every term is >= +0,
// so the two x1 taps contribute exactly +0 and are not loaded.  Rows clamp to the image (H2).
template <typename T>
__device__ __forceinline__ float bilinear_col(const T* __restrict__ pl, int pitch, int W, int H, float yv, int uj) {
    const float fl = floorf(yv);
    const int y0 = (int)fl;
    const float w0 = (float)(y0 + 1) - yv;  // y0_weight = y1 - y
    const float w1 = 1.0f - w0;             // y1_weight
    const int xc = clampi(uj, 0, W - 1);
    const int r0 = clampi(y0, 0, H - 1), r1 = clampi(y0 + 1, 0, H - 1);
    const float v0 = (float)__ldg(pl + (size_t)r0 * pitch + xc);
    const float v1 = (float)__ldg(pl + (size_t)r1 * pitch + xc);
    return v0 * w0 + v1 * w1;
}

// GetPixelDepth, PM.cc:845-875 (Eq. 8).  d2 = R21.row(2)*xp, d0 = R21.row(0)*xp accumulated in double
// (cv::gemm general path); num1 = (float)(d2*ucx), num2f = (float)(d0*fx) is hoisted by the caller.
__device__ __forceinline__ float pixel_inv_depth(float u, float cx, double d2, float num2f, float t2, float den2) {
    const float ucx = u - cx;
    const float num1 = (float)(d2 * (double)ucx);
    const float den1 = -t2 * ucx;
    return (num1 - num2f) / (den1 + den2);
}

// ChiTest, PM.cc:912-924
__device__ __forceinline__ bool chi_test(float a, float b, float sa, float sb, float thr) {
    const float num = (a - b) * (a - b);
    const float chi = num / (sa * sa) + num / (sb * sb);
    return chi < thr;
}

constexpr int K1_WARPS = 8;
constexpr int K1_QCAP = 64;

// One warp = 32 consecutive entries of the active-pixel list (one lane per pixel).  For each
// neighbour key frame: (A) per lane the epipolar line, slope test, search range and the two
// gate angles; (B) all lanes walk their own candidate columns, the three cheap gates
// (PM.cc:408-431) run on every candidate from ONE gathered float (the gate plane), survivors are
// pushed into a per-warp shared-memory queue; (C) whenever 32 survivors are queued they are
// evaluated one per lane (photometric + gradient residual, PM.cc:433-436) and merged into the
// owner's best match with a 64-bit shared atomicMin on (err bits, uj) -- smallest err wins, ties
// go to the smallest uj, exactly the sequential strict-< scan of PM.cc:437-443; (D) per lane the
// sub-pixel refinement and the inverse-depth hypothesis (PM.cc:446-463, 806-829).  After the last
// neighbour: InverseDepthHypothesisFusion (PM.cc:598-626) per lane.
__global__ void __launch_bounds__(K1_WARPS * 32) k1_search(const __grid_constant__ SearchArgs A) {
    __shared__ uint32_t s_queue[K1_WARPS][K1_QCAP];
    __shared__ unsigned long long s_best[K1_WARPS][32];

    const unsigned full = 0xffffffffu;
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    const int count = A.count_override >= 0 ? A.count_override : A.meta->active_count;
    const int entry = (blockIdx.x * K1_WARPS + wib) * 32 + lane;
    if (entry - lane >= count) return;  // whole warp out of range
    const bool valid = entry < count;

    uint32_t* queue = s_queue[wib];
    unsigned long long* best = s_best[wib];

    int x = 0, y = 0;
    if (valid) {
        uint32_t e = A.list[entry];
        x = (int)(e & 0xffffu);
        y = (int)(e >> 16);
    }
    const float xf = (float)x, yf = (float)y;
    const float I1 = valid ? (float)__ldg(A.img + (size_t)y * A.pitch_u8 + x) : 0.f;  // PM.cc:202
    const float g1 = valid ? __ldg(A.grad + (size_t)y * A.pitch_f + x) : 0.f;         // PM.cc:434
    const float th1 = valid ? __ldg(A.theta + (size_t)y * A.pitch_f + x) : 0.f;       // PM.cc:214
    const float xp0 = (xf - A.cx) / A.fx, xp1 = (yf - A.cy) / A.fy;                   // PM.cc:893, 862
    const float colsf = (float)A.W;

    float hd[SD_MAX_NEIGHBOURS], hs[SD_MAX_NEIGHBOURS];
    int nh = 0;

    for (int j = 0; j < A.n; ++j) {
        const NbrConst& N = A.nb[j];
        // ---------------- (A) line, slope test, search range: PM.cc:389-393, 877-910 ----------------
        const float a = xf * N.F[0] + yf * N.F[3] + N.F[6];
        const float b = xf * N.F[1] + yf * N.F[4] + N.F[7];
        const float c = xf * N.F[2] + yf * N.F[5] + N.F[8];
        const float ab = a / b, cb = c / b;
        bool ok = valid && !(ab < -A.slope || ab > A.slope);

        // xp2 = gemm(R21, xp1, rho, t21, 1): float dot left to right, epilogue in double
        const float q0 = N.R[0] * xp0 + N.R[1] * xp1 + N.R[2] * 1.0f;
        const float q1 = N.R[3] * xp0 + N.R[4] * xp1 + N.R[5] * 1.0f;
        const float q2 = N.R[6] * xp0 + N.R[7] * xp1 + N.R[8] * 1.0f;
        (void)q1;
        const float xmin0 = (float)__fma_rn((double)q0, (double)A.mind, (double)N.t[0]);
        const float xmin2 = (float)__fma_rn((double)q2, (double)A.mind, (double)N.t[2]);
        const float xmax0 = (float)__fma_rn((double)q0, (double)A.maxd, (double)N.t[0]);
        const float xmax2 = (float)__fma_rn((double)q2, (double)A.maxd, (double)N.t[2]);
        float umin = A.fx * xmin0 / xmin2 + A.cx;
        float umax = A.fx * xmax0 / xmax2 + A.cx;
        if (umin > umax) { float t = umax; umax = umin; umin = t; }
        if (umin < 0) umin = 0;
        if (umax < 0) umax = 0;
        if (umin > colsf) umin = colsf;
        if (umax > colsf) umax = colsf;
        if (!(umin == umin) || !(umax == umax)) ok = false;  // NaN bounds: empty range
        const int u_lo = (int)ceilf(umin);
        const int u_hi = (int)floorf(umax);
        const int cnt = ok ? max(0, u_hi - u_lo + 1) : 0;

        // gate angles, loop invariant: PM.cc:414, 424-426
        const float th_epi = fast_atan2_deg(-ab, 1.0f);
        float apr = th1 + N.rot;
        if (apr >= 360) apr -= 360;
        if (apr < 0) apr += 360;

        best[lane] = ~0ull;
        __syncwarp();

        // ---------------- (B)+(C) candidate scan ----------------
        const int maxcnt = __reduce_max_sync(full, cnt);
        const float* __restrict__ gate2 = N.gate;
        int qn = 0;
        float uf = (float)u_lo;
        for (int it = 0; it < maxcnt; ++it, uf += 1.0f) {
            bool surv = false;
            const int uj = u_lo + it;
            if (it < cnt) {
                const float vf = ab * uf + cb;      // (a/b)*uj + (c/b), PM.cc:407
                const int ti = __float2int_rz(vf);  // saturating; NaN -> 0.  vj = -ti
                if (ti < 0 && ti > -A.H) {          // PM.cc:408: 0 < vj < rows
                    const float T2 = __ldg(gate2 + (size_t)(-ti) * A.pitch_f + min(uj, A.W - 1));
                    if (T2 >= 0.0f) {  // condition 1, PM.cc:411
                        float d = T2 - th_epi;  // condition 2, PM.cc:414-421
                        if (d >= 360) d -= 360;
                        if (d < 0) d += 360;
                        if (d > 180) d = 360 - d;
                        if (d > 90) d = 180 - d;
                        float e = T2 - apr;  // condition 3, PM.cc:427-431
                        if (e >= 360) e -= 360;
                        if (e < 0) e += 360;
                        if (e > 180) e = 360 - e;
                        surv = !(d > A.lambdaL) && !(e > A.lambdaTheta);
                    }
                }
            }
            const unsigned m = __ballot_sync(full, surv);
            if (m) {
                if (surv) queue[qn + __popc(m & ((1u << lane) - 1u))] = ((uint32_t)lane << 16) | (uint32_t)uj;
                qn += __popc(m);
                __syncwarp();
            }
            const bool last = (it == maxcnt - 1);
            while (qn >= 32 || (last && qn > 0)) {
                // ---------------- (C) evaluate up to 32 queued survivors, one per lane ----------------
                const int take = min(qn, 32);
                const int base = qn - take;
                int owner = lane, uq = 0;
                const bool have = lane < take;
                if (have) {
                    const uint32_t e = queue[base + lane];
                    owner = (int)(e >> 16);
                    uq = (int)(e & 0xffffu);
                }
                const float oI1 = __shfl_sync(full, I1, owner);
                const float og1 = __shfl_sync(full, g1, owner);
                const float oab = __shfl_sync(full, ab, owner);
                const float ocb = __shfl_sync(full, cb, owner);
                if (have) {
                    const float yv = -(oab * (float)uq + ocb);
                    const float rI = oI1 - bilinear_col<uint8_t>(N.img, A.pitch_u8, A.W, A.H, yv, uq);
                    const float rG = og1 - bilinear_col<float>(N.grad, A.pitch_f, A.W, A.H, yv, uq);
                    // err = rI*rI + (rG*rG)/THETA evaluated in double (THETA is a double literal), stored to float
                    const float err = (float)((double)(rI * rI) + (double)(rG * rG) / A.THETA);
                    if (err < 1000000.0f) {
                        const unsigned long long key = ((unsigned long long)__float_as_uint(err) << 32) | (unsigned)uq;
                        atomicMin(&best[owner], key);
                    }
                }
                qn = base;
                __syncwarp();
            }
        }
        __syncwarp();

        // ---------------- (D) refinement + hypothesis: PM.cc:446-463, 806-829 ----------------
        const unsigned long long key = best[lane];
        float rho = 0.f, sig = 0.f;
        bool found = false;
        int bj = -1;
        if (key != ~0ull) {
            found = true;
            bj = (int)(unsigned)(key & 0xffffffffu);
            const float yb = -(ab * (float)bj + cb);
            const float rI = I1 - bilinear_col<uint8_t>(N.img, A.pitch_u8, A.W, A.H, yb, bj);
            const float rG = g1 - bilinear_col<float>(N.grad, A.pitch_f, A.W, A.H, yb, bj);
            const int up = bj + 1, um = bj - 1;
            const float yp = -(ab * (float)up + cb), ym = -(ab * (float)um + cb);
            const float g = (bilinear_col<uint8_t>(N.img, A.pitch_u8, A.W, A.H, yp, up) -
                             bilinear_col<uint8_t>(N.img, A.pitch_u8, A.W, A.H, ym, um)) / 2;
            const float q = (bilinear_col<float>(N.grad, A.pitch_f, A.W, A.H, yp, up) -
                             bilinear_col<float>(N.grad, A.pitch_f, A.W, A.H, ym, um)) / 2;
            const double itq = A.inv_THETA * (double)q;                           // (1/THETA)*q
            const float den = (float)((double)(g * g) + itq * (double)q);         // PM.cc:455
            const float ustar = (float)((double)bj + ((double)(g * rI) + itq * (double)rG) / (double)den);  // PM.cc:456
            const float istd = N.meta->istddev;
            const float uvar = 2 * istd * istd / den;  // PM.cc:457
            // ComputeInvDepthHypothesis
            const double d2 = __fma_rn((double)N.R[8], 1.0, __fma_rn((double)N.R[7], (double)xp1, (double)N.R[6] * (double)xp0));
            const double d0 = __fma_rn((double)N.R[2], 1.0, __fma_rn((double)N.R[1], (double)xp1, (double)N.R[0] * (double)xp0));
            const float num2f = (float)(d0 * (double)A.fx);
            const float den2 = A.fx * N.t[0];
            const float sq = sqrtf(uvar);
            rho = pixel_inv_depth(ustar, A.cx, d2, num2f, N.t[2], den2);
            const float rmin = pixel_inv_depth(ustar - sq, A.cx, d2, num2f, N.t[2], den2);
            const float rmax = pixel_inv_depth(ustar + sq, A.cx, d2, num2f, N.t[2], den2);
            const float e1 = fabsf(rmax - rho), e2 = fabsf(rmin - rho);
            sig = (e1 < e2) ? e2 : e1;  // std::max(e1, e2)
            if (1 / rho > 0.0f) {       // PM.cc:216
                hd[nh] = rho;
                hs[nh] = sig;
                ++nh;
            }
        }
        if (A.probe_out && j == 0 && valid) {
            float* o = A.probe_out + (size_t)entry * 4;
            o[0] = rho; o[1] = sig; o[2] = (float)bj; o[3] = found ? 1.f : 0.f;
        }
        __syncwarp();
    }

    // ---------------- InverseDepthHypothesisFusion, PM.cc:221-228 + 598-626 ----------------
    if (!valid || A.probe_out || !(nh > A.lambdaN)) return;
    int best_cnt = 0, best_a = -1;
    for (int ia = 0; ia < nh; ++ia) {
        int c = 0;
        for (int ib = 0; ib < nh; ++ib) c += chi_test(hd[ia], hd[ib], hs[ia], hs[ib], A.chi_fuse) ? 1 : 0;
        if (c > best_cnt) { best_cnt = c; best_a = ia; }
    }
    if (best_cnt >= A.lambdaN) {
        // GetFusion(depthHo), PM.cc:947-970: pow(sigma,2) in double, float accumulators
        float pjsj = 0, rsj = 0;
        for (int ib = 0; ib < nh; ++ib) {
            if (!chi_test(hd[best_a], hd[ib], hs[best_a], hs[ib], A.chi_fuse)) continue;
            const double s2 = (double)hs[ib] * (double)hs[ib];
            pjsj = (float)((double)pjsj + (double)hd[ib] / s2);
            rsj = (float)((double)rsj + 1.0 / s2);
        }
        A.depth[(size_t)y * A.pitch_f + x] = pjsj / rsj;
        A.sigma[(size_t)y * A.pitch_f + x] = sqrtf(1 / rsj);
    }
}

// fusion probe: one thread, same code path as the tail of k1_search
__global__ void k1_fusion_probe(const float* d, const float* s, int n, int lambdaN, float chi_fuse, float* out) {
    int best_cnt = 0, best_a = -1;
    for (int ia = 0; ia < n; ++ia) {
        int c = 0;
        for (int ib = 0; ib < n; ++ib) c += chi_test(d[ia], d[ib], s[ia], s[ib], chi_fuse) ? 1 : 0;
        if (c > best_cnt) { best_cnt = c; best_a = ia; }
    }
    out[0] = out[1] = out[2] = 0.f;
    if (best_cnt >= lambdaN) {
        float pjsj = 0, rsj = 0;
        for (int ib = 0; ib < n; ++ib) {
            if (!chi_test(d[best_a], d[ib], s[best_a], s[ib], chi_fuse)) continue;
            const double s2 = (double)s[ib] * (double)s[ib];
            pjsj = (float)((double)pjsj + (double)d[ib] / s2);
            rsj = (float)((double)rsj + 1.0 / s2);
        }
        out[0] = pjsj / rsj;
        out[1] = sqrtf(1 / rsj);
        out[2] = 1.f;
    }
}

// ------------------------------------------------------------------------------------------------
// K2: IntraKeyFrameDepthChecking, PM.cc:486-547.  out = in, except valid interior pixels.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k2_intra_check(const StencilArgs A) {
    const int px = blockIdx.x * blockDim.x + threadIdx.x;
    const int py = blockIdx.y * blockDim.y + threadIdx.y;
    if (px >= A.W || py >= A.H) return;
    const size_t o = (size_t)py * A.pitch_f + px;
    const float dc = A.depth_in[o], sc = A.sigma_in[o];
    float dn = dc, sn = sc;
    if (px >= A.border && px < A.W - A.border && py >= A.border && py < A.H - A.border && dc > SD_EPS_F) {
        // compatible neighbours in raster order, then the pixel itself (PM.cc:504-522), fused by GetFusion
        float pjsj = 0, rsj = 0, min_sigma = 0;
        int ncomp = 0;
#pragma unroll
        for (int k = 0; k < 9; ++k) {
            const int dy = k / 3 - 1, dx = k % 3 - 1;
            if (dy == 0 && dx == 0) continue;
            const size_t q = (size_t)(py + dy) * A.pitch_f + (px + dx);
            const float d = A.depth_in[q];
            if (d > SD_EPS_F) {
                const float s = A.sigma_in[q];
                if (chi_test(d, dc, s, sc, A.chi_fuse)) {
                    const double s2 = (double)s * (double)s;
                    pjsj = (float)((double)pjsj + (double)d / s2);
                    rsj = (float)((double)rsj + 1.0 / s2);
                    // temp_min_sigma starts at compatible_ho[0].sigma; pow(s,2) < pow(min,2) in double
                    if (ncomp == 0 || s2 < (double)min_sigma * (double)min_sigma) min_sigma = s;
                    ++ncomp;
                }
            }
        }
        {
            const double s2 = (double)sc * (double)sc;
            pjsj = (float)((double)pjsj + (double)dc / s2);
            rsj = (float)((double)rsj + 1.0 / s2);
            if (ncomp == 0 || s2 < (double)min_sigma * (double)min_sigma) min_sigma = sc;
            ++ncomp;
        }
        if (ncomp >= 3) {
            dn = pjsj / rsj;
            sn = min_sigma;  // PM.cc:531: min sigma, not the fused sigma
        } else {
            dn = 0.0f;
            sn = 0.0f;
        }
    }
    A.depth_out[o] = dn;
    A.sigma_out[o] = sn;
}

// ------------------------------------------------------------------------------------------------
// K3: IntraKeyFrameDepthGrowing, PM.cc:549-596.  (A no-op in practice: the centre sigma is 0)
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k3_intra_grow(const StencilArgs A) {
    const int px = blockIdx.x * blockDim.x + threadIdx.x;
    const int py = blockIdx.y * blockDim.y + threadIdx.y;
    if (px >= A.W || py >= A.H) return;
    const size_t o = (size_t)py * A.pitch_f + px;
    const float dc = A.depth_in[o], sc = A.sigma_in[o];
    float dn = dc, sn = sc;
    if (px >= A.border && px < A.W - A.border && py >= A.border && py < A.H - A.border && dc <= SD_EPS_F &&
        !(A.grad[o] < A.lambdaG)) {
        float pjsj = 0, rsj = 0, min_sigma = 0;
        int nsup = 0;
#pragma unroll
        for (int k = 0; k < 9; ++k) {
            const int dy = k / 3 - 1, dx = k % 3 - 1;
            if (dy == 0 && dx == 0) continue;
            const size_t q = (size_t)(py + dy) * A.pitch_f + (px + dx);
            const float d = A.depth_in[q], s = A.sigma_in[q];
            if (chi_test(d, dc, s, sc, A.chi_fuse)) {
                // GetFusion(pairs), PM.cc:926-945
                const double s2 = (double)s * (double)s;
                pjsj = (float)((double)pjsj + (double)d / s2);
                rsj = (float)((double)rsj + 1.0 / s2);
                if (nsup == 0 || s < min_sigma) min_sigma = s;
                ++nsup;
            }
        }
        if (nsup >= 2) {
            dn = pjsj / rsj;
            sn = min_sigma;
        }
    }
    A.depth_out[o] = dn;
    A.sigma_out[o] = sn;
}

// ------------------------------------------------------------------------------------------------
// K4: InterKeyFrameDepthChecking, PM.cc:628-799, snapshot semantics.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k4_inter_check(const __grid_constant__ InterArgs A) {
    const int px = blockIdx.x * blockDim.x + threadIdx.x;
    const int py = blockIdx.y * blockDim.y + threadIdx.y;
    if (px >= A.W || py >= A.H) return;
    const size_t o = (size_t)py * A.pitch_f + px;
    const float depthp = A.depth_in[o];
    float out = depthp;
    if (px >= 2 && px < A.W - 2 && py >= 2 && py < A.H - 2 && depthp > SD_EPS_F) {  // PM.cc:659-662 (`< 1e-6 -> continue`)
        const float xp0 = ((float)px - A.cx) / A.fx, xp1 = ((float)py - A.cy) / A.fy;
        const float dp = 1 / depthp;
        // operator/(MatExpr, double): alpha = 1./s, which hal::gemm32f(float alpha) rounds to float
        const double inv_dp = (double)(float)(1.0 / (double)depthp);
        int nkf = 0;
        double sJr = 0.0, sJJ = 0.0;
        for (int j = 0; j < A.n; ++j) {
            const InterNbr& N = A.nb[j];
            // temp = gemm(Rji, xp, 1/depthp, tji, 1); Xj = K*temp (small path, float left to right)
            const float q0 = N.R[0] * xp0 + N.R[1] * xp1 + N.R[2] * 1.0f;
            const float q1 = N.R[3] * xp0 + N.R[4] * xp1 + N.R[5] * 1.0f;
            const float q2 = N.R[6] * xp0 + N.R[7] * xp1 + N.R[8] * 1.0f;
            const float t0 = (float)((double)q0 * inv_dp + (double)N.t[0]);
            const float t1 = (float)((double)q1 * inv_dp + (double)N.t[1]);
            const float t2 = (float)((double)q2 * inv_dp + (double)N.t[2]);
            const float X0 = N.fx * t0 + 0.0f * t1 + N.cx * t2;
            const float X1 = 0.0f * t0 + N.fy * t1 + N.cy * t2;
            const float X2 = 0.0f * t0 + 0.0f * t1 + 1.0f * t2;
            const float inv = (float)(1.0 / (double)X2);  // Xj/Xj(2): convertTo with alpha = 1./s
            const float xj = X0 * inv, yj = X1 * inv;
            // Eq. 12, PM.cc:684-688
            const double rz_d = __fma_rn((double)N.R[8], 1.0, __fma_rn((double)N.R[7], (double)xp1, (double)N.R[6] * (double)xp0));
            const float rz = (float)rz_d;
            const float depthj = depthp / (rz + depthp * N.t[2]);
            if (!(xj >= 0 && xj < (float)(A.W - 1) && yj >= 0 && yj < (float)(A.H - 1))) continue;  // PM.cc:695
            const int x0 = (int)floorf(xj), y0 = (int)floorf(yj);
            int ncomp = 0;
#pragma unroll
            for (int t = 0; t < 4; ++t) {  // taps (y0,x0) (y1,x0) (y0,x1) (y1,x1), PM.cc:705-752
                const int ty = y0 + (t & 1), tx = x0 + (t >> 1);
                const size_t q = (size_t)ty * A.pitch_f + tx;
                const float d = __ldg(N.depth + q);
                if (d > SD_EPS_F) {
                    const float s = __ldg(N.sigma + q);
                    const double dd = (double)(depthj - d);
                    const float test = (float)((dd * dd) / ((double)s * (double)s));
                    if (test < A.chi_inter) {
                        ++ncomp;
                        // Gauss-Newton terms, PM.cc:777-783 (order: neighbour j, then tap n)
                        const float djn = 1 / d;
                        const float d2sigma = djn * djn * s;
                        const float Jn = -rz / d2sigma;
                        const float rn = (djn - dp * rz - N.t[2]) / d2sigma;
                        sJr = __fma_rn((double)Jn, (double)rn, sJr);
                        sJJ = __fma_rn((double)Jn, (double)Jn, sJJ);
                    }
                }
            }
            if (ncomp >= 1) ++nkf;
        }
        if (nkf < A.lambdaN) {
            out = 0.0f;  // sigma untouched, PM.cc:764
        } else {
            const float Jtr0 = (float)(sJr * -1.0);
            const float JtJ = (float)(sJJ * 1.0);
            const float dpDelta = Jtr0 / JtJ;
            out = 1 / (dp + dpDelta);  // PM.cc:793
        }
    }
    A.depth_out[o] = out;
}

// ------------------------------------------------------------------------------------------------
// K5: UpdateSemiDensePointSet, PM.cc:337-367.  Twc*Pc: 4x4*4x1 small gemm, float left to right.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k5_pointset(const PointArgs A) {
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
    const int y = blockIdx.y * blockDim.y + threadIdx.y;
    if (x >= A.W || y >= A.H) return;
    float p0 = 0.f, p1 = 0.f, p2 = 0.f;
    if (x >= 2 && x < A.W - 2 && y >= 2 && y < A.H - 2) {
        const float inv_d = A.depth[(size_t)y * A.pitch_f + x];
        if (inv_d > SD_EPS_F) {
            const float Z = 1 / inv_d;
            const float X = Z * ((float)x - A.cx) / A.fx;
            const float Y = Z * ((float)y - A.cy) / A.fy;
            p0 = A.Twc[0] * X + A.Twc[1] * Y + A.Twc[2] * Z + A.Twc[3] * 1.0f;
            p1 = A.Twc[4] * X + A.Twc[5] * Y + A.Twc[6] * Z + A.Twc[7] * 1.0f;
            p2 = A.Twc[8] * X + A.Twc[9] * Y + A.Twc[10] * Z + A.Twc[11] * 1.0f;
        }
    }
    float* o = A.points + ((size_t)y * A.W + x) * 3;
    o[0] = p0; o[1] = p1; o[2] = p2;
}

// pack / unpack between pitched planes and tight W*H buffers (NCCL halo exchange)
__global__ void k_copy_plane(const float* __restrict__ src, int src_pitch, float* __restrict__ dst, int dst_pitch, int W, int H) {
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
    const int y = blockIdx.y * blockDim.y + threadIdx.y;
    if (x < W && y < H) dst[(size_t)y * dst_pitch + x] = src[(size_t)y * src_pitch + x];
}

}  // namespace sdpm
