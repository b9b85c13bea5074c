// CUDA kernels (sm_100a) of the semi-dense probability-mapping path.
//
//   K0 k0_gradient     image -> GradImg / GradTheta / orientation-arc bit planes / active-pixel list / I_stddev sums
//   K1 k1_search       per active pixel: EpipolarSearch in N neighbours -> hypothesis buffer
//      k1_fuse         per active pixel: InverseDepthHypothesisFusion of its hypotheses -> (depth, sigma) maps
//   K2+K3 k23_intra    IntraKeyFrameDepthChecking + IntraKeyFrameDepthGrowing (3x3 chi-square stencils,
//                      one pass) + the valid-pixel list K4 consumes
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
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/sdpm.h"

namespace sdpm {

struct KfMeta {  // device-resident per key frame
    unsigned long long sum;    // sum of grey levels
    unsigned long long sumsq;  // sum of squared grey levels
    float istddev;             // KeyFrame::I_stddev
    int active_count;          // entries in the active-pixel list
    unsigned int k1_task;      // next 32-entry task of the running k1_search launch (zero between launches: the last warp out resets it)
    int valid_count;           // entries in the valid-pixel list written by k23_intra (zeroed by the k1_fuse launch before it)
    unsigned int k0_ticket;    // blocks of k0_gradient that have added their sums (the last one computes istddev)
};

struct NbrConst {  // per neighbour, warp-uniform, lives in kernel parameter (constant) space
    float F[9];    // F12, PM.cc:972-987
    float R[9];    // R21, PM.cc:890
    float t[3];    // t21, PM.cc:891
    float rot;     // median in-plane rotation, degrees, PM.cc:170-179
    double Rd0[3], Rd2[3];  // rows 0 and 2 of R21 as doubles (double-accumulated dot products, PM.cc:866-868)
    double td0, td2;        // t21[0], t21[2] as doubles (gemm beta*C epilogue, PM.cc:894-895)
    const uint8_t* img;
    const float* grad;
    const float* theta;
    const uint32_t* arc;  // SD_ARC_PLANES bit planes, see k0_gradient
    const uint4* rec;     // search records, see k0_gradient
    const KfMeta* meta;
};

struct SearchArgs {
    int W, H, pitch_f, pitch_u8, n;
    int arc_pitch;  // u32 words per bit-plane row
    int border, lambdaN;
    float fx, fy, cx, cy;
    float mind, maxd;  // StereoSearchConstraints interval (inverse depth; mind is the larger), PM.cc:381-382
    float lambdaG, lambdaL, lambdaTheta, slope;
    float chi_fuse;  // smallest float >= 5.99: `chi < 5.99` (float vs double literal) == `chi < chi_fuse`
    double mind_d, maxd_d;    // the same as doubles (gemm alpha)
    double inv_THETA, THETA;  // 1/THETA and THETA (PM.h:47) as doubles
    float inv_theta_f;        // (float)inv_THETA
    const float* ray;         // normalised image coordinates of the key frame: ray[x] = (x - cx) / fx, ray[W + y] = (y - cy) / fy (k0_gradient)
    const uint8_t* img;
    const float* grad;
    const float* theta;
    const uint4* rec;    // search records of the key frame itself (own-pixel intensity, gradient, orientation in one load)
    const uint32_t* list;
    const KfMeta* meta;
    int count_override;  // >= 0: use instead of meta->active_count (probes)
    unsigned int* task_counter;  // zero at launch; warps draw 32-entry tasks from it and the last warp to leave zeroes it again
    float2* hyp;       // nullable; hypothesis buffer [neighbour][hyp_stride]: (rho, sigma), or (-1, 0) for none
    int hyp_stride;
    float* probe_out;  // nullable; [entry][6] = {rho, sigma, best uj, supported, best_u, best_v} of neighbour 0
    NbrConst nb[SD_MAX_NEIGHBOURS];
};

struct IntraArgs {  // k23_intra: IntraKeyFrameDepthChecking + IntraKeyFrameDepthGrowing in one pass
    int W, H, pitch_in, pitch_out, border;
    int vec_ok;  // 1: pitch_out % 4 == 0 and every output base pointer is 16-byte aligned (float4 stores)
    float lambdaG, chi_fuse;
    const float* depth_in;   // K1 output (or caller planes)
    const float* sigma_in;
    const float* grad;       // read only by the growing stage, only where it can act
    float* depth_out;
    float* sigma_out;
    float* depth_copy;       // nullable: second copy of depth_out (the plane k4_inter_check patches)
    uint32_t* vlist;         // nullable: compacted (y << 16 | x) of pixels with depth_out != 0 or sigma_out != 0 inside the 2-px frame
    int* vcount;             // nullable: entries in vlist (zeroed before the launch)
};

struct PointArgs {
    int W, H, pitch_f;
    float fx, fy, cx, cy;
    float Twc[16];
    const float* depth;
    float* points;  // H x 3W, row pitch 3*W floats (tight)
};

struct InterNbr {
    float R[9];    // Rji, PM.cc:643
    float t[3];    // tji, PM.cc:644
    double Rd2[3]; // Rji row 2 as doubles (operands of the double-accumulated dot product, PM.cc:684-686)
    double td[3];  // tji as doubles (gemm beta*C epilogue, PM.cc:678)
    float fx, fy, cx, cy;  // K of neighbour j, PM.cc:675
    const float* depth;
    const float* sigma;
};

struct InterArgs {
    int W, H, pitch_f, n, lambdaN;
    float fx, fy, cx, cy;
    float chi_inter;  // smallest float >= 3.84
    const float* depth_in;
    float* depth_out;          // pre-filled with a copy of depth_in (k23_intra's depth_copy); patched at the listed pixels
    const uint32_t* vlist;     // pixels carrying a depth or a sigma inside the 2-px frame (PM.cc:659); depth_in <= 1e-6 is skipped (PM.cc:662)
    const int* vcount;
    PointArgs P;               // k4_inter_check<true>: UpdateSemiDensePointSet (PM.cc:337-367) of the checked pixels, into a zeroed plane
    InterNbr nb[SD_MAX_NEIGHBOURS];
};


// `d > 0.000001` with a float d and a double literal == `d > SD_EPS_F` (largest float <= 1e-6);
// `d < 0.000001` == `d <= SD_EPS_F`.
#define SD_EPS_F 9.99999997475242707878e-07f

__device__ __forceinline__ int clampi(int v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : v); }

// The two grey levels of a search record (this row | the row below): SD_REC_HALF packs them as two half floats -- every u8
// is exact in fp16, and half -> float is ONE instruction on the packed word (HADD2.F32 with a half selector) instead of a
// mask / shift, an OR and a subtraction each.
#ifndef SD_REC_HALF
#define SD_REC_HALF 1  // 0: two 16-bit integers (round 2's first layout); bench 3 722 -> 3 760 KF/s, K0 26 -> 24 us with halves
#endif
__device__ __forceinline__ uint32_t rec_pack_grey(int lo, int hi) {
#if SD_REC_HALF
    return (uint32_t)__half_as_ushort(__int2half_rn(lo)) | ((uint32_t)__half_as_ushort(__int2half_rn(hi)) << 16);
#else
    return (uint32_t)lo | ((uint32_t)hi << 16);
#endif
}
__device__ __forceinline__ float small_int_to_float(int u);
__device__ __forceinline__ float rec_grey_lo(uint32_t w) {
#if SD_REC_HALF
    return __low2float(*reinterpret_cast<const __half2*>(&w));
#else
    return small_int_to_float((int)(w & 0xffffu));
#endif
}
__device__ __forceinline__ float rec_grey_hi(uint32_t w) {
#if SD_REC_HALF
    return __high2float(*reinterpret_cast<const __half2*>(&w));
#else
    return small_int_to_float((int)(w >> 16));
#endif
}


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
// K0: gradient front end (build-defined) + orientation-arc bit planes + active list +
// I_stddev sums.
// 3x3 Sobel/8 with replicate border; GradImg = sqrtf(gx^2+gy^2); GradTheta = fastAtan2(gy,gx).
// Block 32x8 threads, 4 pixels per thread (128x8 tile).
//
// Orientation-arc bit planes (what K1 scans instead of testing every candidate column):
// plane s (s = 0..31) holds one bit per pixel, row-major, 32 columns per u32 word (bit = x & 31):
//     bit = GradImg >= lambdaG  (condition 1, PM.cc:411)
//           && circular distance(GradTheta, (s + 0.5) * 11.25 deg) <= lambdaTheta + 5.625 + SD_ARC_MARGIN
// i.e. every pixel that condition 3 (PM.cc:424-431) can accept for SOME rotated orientation
// `th_pi + rot` inside the 11.25-degree cell s.  It is a conservative SUPERSET: K1 re-runs the
// exact float tests of conditions 2 and 3 on every pixel a plane selects, so the planes only
// decide which columns are looked at, never the result.  Plane 32 is condition 1 alone (used when
// the rotated orientation falls outside [0, 360), where PM.cc's single fold is not a circular
// distance).  All 33 planes together are 33/8 bytes per pixel.
// ------------------------------------------------------------------------------------------------
constexpr int K0_TX = 32, K0_TY = 8, K0_PX = 4;
constexpr int K0_TW = K0_TX * K0_PX;  // 128
constexpr int SD_ARC_CELLS = 32;
constexpr int SD_ARC_PLANES = SD_ARC_CELLS + 1;
#define SD_ARC_CELL_DEG 11.25f
#define SD_ARC_MARGIN 0.05f

__device__ __forceinline__ float sobel_mag(int p00, int p01, int p02, int p10, int p12, int p20, int p21, int p22, float& gx, float& gy) {
    const int sx = (p02 + 2 * p12 + p22) - (p00 + 2 * p10 + p20);
    const int sy = (p20 + 2 * p21 + p22) - (p00 + 2 * p01 + p02);
    gx = (float)sx * 0.125f;
    gy = (float)sy * 0.125f;
    return sqrtf(gx * gx + gy * gy);
}

#ifndef K0_MIN_BLOCKS
#define K0_MIN_BLOCKS 6
#endif
__global__ void __launch_bounds__(K0_TX* K0_TY, K0_MIN_BLOCKS) k0_gradient(const uint8_t* __restrict__ img, int pitch_u8, int W, int H,
                                                            float* __restrict__ grad, float* __restrict__ theta, int pitch_f,
                                                            uint32_t* __restrict__ arc, int arc_pitch, uint4* __restrict__ rec,
                                                            uint32_t* __restrict__ list, KfMeta* __restrict__ meta,
                                                            float lambdaG, float lambdaTheta, int border,
                                                            float* __restrict__ ray, float fx, float fy, float cx, float cy) {
    __shared__ uint8_t tile[K0_TY + 3][K0_TW + 8];  // column c <-> x0 - 1 + c; row r <-> clamp(y0 - 1 + r)
    __shared__ float s_g[K0_TY + 1][K0_TW];          // GradImg of the tile rows and of the row below the tile (records)
    __shared__ uint32_t s_memb[K0_TY][K0_TW];        // per pixel: bit s = member of arc plane s
    __shared__ int warp_cnt[K0_TY];
    __shared__ int block_base;
    __shared__ unsigned long long red[2][K0_TY];

    const int tx = threadIdx.x, ty = threadIdx.y;
    const int tid = ty * K0_TX + tx;
    const int x0 = blockIdx.x * K0_TW, y0 = blockIdx.y * K0_TY;

    // normalised image coordinates xp = K^-1 (x, y, 1) (PM.cc:862, 893), once per column / row instead of once per
    // (pixel, neighbour) in k1_search and once per pixel in k4_inter_check: the same two IEEE operations, (x - cx) / fx
    if (blockIdx.y == 0 && tid < K0_TW && x0 + tid < W) ray[x0 + tid] = ((float)(x0 + tid) - cx) / fx;
    if (blockIdx.x == 0 && tid < K0_TY && y0 + tid < H) ray[W + y0 + tid] = ((float)(y0 + tid) - cy) / fy;

    if (x0 + K0_TW < W) {
        // interior columns: the 128 tile bytes of a row are 32 aligned words (x0 and the pitch are multiples of 128),
        // plus one halo byte on each side; rows clamp to the image (replicate border)
        for (int i = tid; i < (K0_TY + 3) * 34; i += K0_TX * K0_TY) {
            const int r = i / 34, k = i - r * 34;
            const uint8_t* row = img + (size_t)clampi(y0 - 1 + r, 0, H - 1) * pitch_u8;
            if (k < 32) {
                const uint32_t v = __ldg(reinterpret_cast<const uint32_t*>(row + x0) + k);
                uint8_t* d = &tile[r][1 + 4 * k];  // tile column c <-> x0 - 1 + c: byte-wise stores (1 + 4k is not word-aligned)
                d[0] = (uint8_t)v; d[1] = (uint8_t)(v >> 8); d[2] = (uint8_t)(v >> 16); d[3] = (uint8_t)(v >> 24);
            } else if (k == 32) {
                tile[r][0] = __ldg(row + max(x0 - 1, 0));
            } else {
                tile[r][K0_TW + 1] = __ldg(row + x0 + K0_TW);  // < W here
            }
        }
    } else {
        for (int i = tid; i < (K0_TY + 3) * (K0_TW + 2); i += K0_TX * K0_TY) {
            int r = i / (K0_TW + 2), c = i - r * (K0_TW + 2);
            int yy = clampi(y0 - 1 + r, 0, H - 1), xx = clampi(x0 - 1 + c, 0, W - 1);
            tile[r][c] = __ldg(img + (size_t)yy * pitch_u8 + xx);
        }
    }
    __syncthreads();

    const int y = y0 + ty;
    const int xb = x0 + tx * K0_PX;
    const float arcR = lambdaTheta + 0.5f * SD_ARC_CELL_DEG + SD_ARC_MARGIN;
    float g[K0_PX], th[K0_PX];
    uint32_t pk[K0_PX];
    unsigned long long s1 = 0, s2 = 0;
    uint32_t act[K0_PX];
    int nact = 0;
#pragma unroll
    for (int i = 0; i < K0_PX; ++i) {
        const int c = tx * K0_PX + i + 1, r = ty + 1;
        int p00 = tile[r - 1][c - 1], p01 = tile[r - 1][c], p02 = tile[r - 1][c + 1];
        int p10 = tile[r][c - 1], p11 = tile[r][c], p12 = tile[r][c + 1];
        int p20 = tile[r + 1][c - 1], p21 = tile[r + 1][c], p22 = tile[r + 1][c + 1];
        float gx, gy;
        g[i] = sobel_mag(p00, p01, p02, p10, p12, p20, p21, p22, gx, gy);
        s_g[ty][c - 1] = g[i];
        pk[i] = rec_pack_grey(p11, p21);  // tile rows are clamped at load: p21 is row min(y + 1, H - 1)
        th[i] = fast_atan2_deg(gy, gx);
        const int x = xb + i;
        const bool on = (g[i] >= lambdaG) && x < W && y < H;
        uint32_t memb = 0;
        if (on) {
            // cells s (mod 32) with (s + 0.5) * cell within arcR of theta
            const float lo = (th[i] - arcR) * (1.0f / SD_ARC_CELL_DEG) - 0.5f;
            const float hi = (th[i] + arcR) * (1.0f / SD_ARC_CELL_DEG) - 0.5f;
            const int s_lo = (int)ceilf(lo), s_hi = (int)floorf(hi);
            const int cnt = s_hi - s_lo + 1;
            if (cnt >= SD_ARC_CELLS) memb = 0xffffffffu;
            else if (cnt > 0) memb = __funnelshift_l((1u << cnt) - 1u, (1u << cnt) - 1u, s_lo & 31);  // rotate left
        }
        s_memb[ty][tx * K0_PX + i] = memb;
        if (x < W && y < H) {
            s1 += (unsigned)p11;
            s2 += (unsigned)(p11 * p11);
            if (on && x >= border && x < W - border && y >= border && y < H - border) act[nact++] = ((uint32_t)y << 16) | (uint32_t)x;
        }
    }
    // GradImg of the row below the tile (image row y0 + K0_TY), for the records of the tile's last row
    if (tid < K0_TW && y0 + K0_TY < H) {
        const int c = tid + 1, r = K0_TY + 1;
        float gx, gy;
        s_g[K0_TY][tid] = sobel_mag(tile[r - 1][c - 1], tile[r - 1][c], tile[r - 1][c + 1], tile[r][c - 1], tile[r][c + 1],
                                    tile[r + 1][c - 1], tile[r + 1][c], tile[r + 1][c + 1], gx, gy);
    }
    __syncthreads();
    if (y < H) {
        const size_t o = (size_t)y * pitch_f + xb;
        if (xb + K0_PX <= W) {  // pitch_f and xb are multiples of 4: aligned float4 stores
            *reinterpret_cast<float4*>(grad + o) = make_float4(g[0], g[1], g[2], g[3]);
            *reinterpret_cast<float4*>(theta + o) = make_float4(th[0], th[1], th[2], th[3]);
        } else {
            for (int i = 0; i < K0_PX; ++i)
                if (xb + i < W) { grad[o + i] = g[i]; theta[o + i] = th[i]; }
        }
#pragma unroll
        for (int i = 0; i < K0_PX; ++i)
            if (xb + i < W) {
                const float gn = (y + 1 < H) ? s_g[ty + 1][tx * K0_PX + i] : g[i];  // second bilinear row clamps to the image
                rec[o + i] = make_uint4(__float_as_uint(th[i]), __float_as_uint(g[i]), __float_as_uint(gn), pk[i]);
            }
    }

    // 128 pixels x 32 planes -> 32 planes x 4 words: the warp (one image row of the tile) transposes four 32x32 bit
    // matrices (rows = columns of the image, held one per lane; bits = planes) with the five-stage butterfly --
    // lanes i and i^j swap their off-diagonal j x j blocks -- so that lane s ends up with the word of plane s.
    __syncwarp();
    const unsigned full = 0xffffffffu;
    uint32_t pw[K0_PX], pall[K0_PX];
#pragma unroll
    for (int w = 0; w < K0_PX; ++w) {
        uint32_t v = s_memb[ty][w * 32 + tx];
        pall[w] = __ballot_sync(full, v != 0u);
        uint32_t m = 0x0000ffffu;
#pragma unroll
        for (int j = 16; j; j >>= 1) {
            const uint32_t other = __shfl_xor_sync(full, v, j);
            v = (tx & j) ? ((v & ~m) | ((other >> j) & m)) : ((v & m) | ((other << j) & ~m));
            m ^= m << (j >> 1);
        }
        pw[w] = v;
    }
    if (y < H) {  // arc_pitch is a multiple of 4 words and covers the last tile: aligned uint4 stores
        const size_t plane = (size_t)H * arc_pitch;
        uint32_t* dst = arc + (size_t)tx * plane + (size_t)y * arc_pitch + (x0 >> 5);
        *reinterpret_cast<uint4*>(dst) = make_uint4(pw[0], pw[1], pw[2], pw[3]);
        if (tx == 0) {
            uint32_t* d32 = arc + (size_t)SD_ARC_CELLS * plane + (size_t)y * arc_pitch + (x0 >> 5);
            *reinterpret_cast<uint4*>(d32) = make_uint4(pall[0], pall[1], pall[2], pall[3]);
        }
    }

    // block-wide compaction of the active pixels: warp scan, then one atomicAdd per block
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
        // the block that draws the last ticket sees every block's sums (fence before the ticket, atomic reads after):
        // I_stddev = population standard deviation of the u8 image, double arithmetic (== cv::meanStdDev)
        __threadfence();
        if (atomicAdd(&meta->k0_ticket, 1u) == gridDim.x * gridDim.y - 1u) {
            __threadfence();
            const double n = (double)W * (double)H;
            const double mean = (double)atomicAdd(&meta->sum, 0ull) / n;
            const double var = (double)atomicAdd(&meta->sumsq, 0ull) / n - mean * mean;
            meta->istddev = (float)sqrt(var > 0.0 ? var : 0.0);
        }
    }
    __syncthreads();
    int pos = block_base + warp_cnt[ty] + incl - nact;
    for (int i = 0; i < nact; ++i) list[pos + i] = act[i];
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

// (float)u for 0 <= u < 2^23 without the conversion pipe: 2^23 + u is exact, so is the subtraction
__device__ __forceinline__ float small_int_to_float(int u) { return __int_as_float(0x4B000000 | u) - 8388608.0f; }

// bilinear<uchar>(im_, y, x) and bilinear<float>(GradImg, y, x) of PM.cc:40-59 at an integer column, both from ONE search
// record (k0_gradient): same arithmetic as bilinear_col.  Row y0 = floor(yv) clamps to the image; the record of the
// clamped row holds the row below it (itself clamped), and for y0 < 0 both taps are row 0.
__device__ __forceinline__ void bilinear_rec(const uint4* __restrict__ rec, int pitch, int W, int H, float yv, int uj, float& vI, float& vG) {
    const float fl = floorf(yv);
    const int y0 = (int)fl;
    const float w0 = (float)(y0 + 1) - yv;  // y0_weight = y1 - y
    const float w1 = 1.0f - w0;             // y1_weight
    const int xc = clampi(uj, 0, W - 1);
    const uint4 r = __ldg(rec + (unsigned)(clampi(y0, 0, H - 1) * pitch + xc));
    const bool above = y0 < 0;
    const float i0 = rec_grey_lo(r.w);
    const float i1 = above ? i0 : rec_grey_hi(r.w);
    const float g0 = __uint_as_float(r.y);
    const float g1 = above ? g0 : __uint_as_float(r.z);
    vI = i0 * w0 + i1 * w1;
    vG = g0 * w0 + g1 * w1;
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

// conditions 2 and 3 of EpipolarSearch on one candidate orientation T2 (degrees), PM.cc:414-431
__device__ __forceinline__ bool orientation_gates(float T2, float th_epi, float apr, float lambdaL, float lambdaTheta) {
    float d = T2 - th_epi;  // condition 2, PM.cc:414-421
    if (d >= 360) d -= 360;
    if (d < 0) d += 360;
    if (d > 180) d = 360 - d;
    if (d > 90) d = 180 - d;
    float e = T2 - apr;  // condition 3, PM.cc:427-431
    if (e >= 360) e -= 360;
    if (e < 0) e += 360;
    if (e > 180) e = 360 - e;
    return !(d > lambdaL) && !(e > lambdaTheta);
}

// Image row of candidate column u on the epipolar line.  The reference computes
// vj = -(int)((a/b)*uj + c/b) (PM.cc:407) and rejects vj <= 0 || vj >= rows (PM.cc:408).  With
// vf = (a/b)*uj + c/b this function returns floor(-vf) when |vf| < 2^22 -- equal to vj whenever vj is
// an admissible row (vf < 0 there) -- and some value outside (0, 65536) for every other vf (vf >= 0,
// huge, inf, NaN), so `0 < row && row < H` is exactly the reference's admission test.  Adding
// 1.5*2^23 with round-toward-zero leaves floor(x) in the low mantissa bits of a positive sum; no
// float<->int conversion instruction is needed.  The row is monotone in u, so the columns that
// share a row form one contiguous run.
__device__ __forceinline__ int line_row(float ab, float cb, float uf, float& vf) {
    vf = __fadd_rn(__fmul_rn(ab, uf), cb);
    return __float_as_int(__fadd_rz(-vf, 12582912.0f)) - 0x4B400000;
}
// the same, also returning the magic sum t = 1.5*2^23 + row as a float: for an admissible row, t - (1.5*2^23 - 1) is
// (float)row + 1 exactly (all three are integers below 2^24)
__device__ __forceinline__ int line_row(float ab, float cb, float uf, float& vf, float& t) {
    vf = __fadd_rn(__fmul_rn(ab, uf), cb);
    t = __fadd_rz(-vf, 12582912.0f);
    return __float_as_int(t) - 0x4B400000;
}
__device__ __forceinline__ int line_row(float ab, float cb, int u) {
    float vf;
    return line_row(ab, cb, small_int_to_float(u), vf);
}

// largest u1 in [u, ue) whose row equals r0; precondition line_row(u) == r0 != line_row(ue).
// The division-based estimate only decides where the exact search starts.
__device__ __forceinline__ int run_end(float ab, float cb, float inv_ab, int u, int ue, int r0) {
    // the row changes where ab*u + cb crosses -r0 (ab > 0: row decreases) or -(r0 + 1) (ab < 0)
    const float bv = -(float)r0 - ((ab > 0.f) ? 0.f : 1.f);
    const float uc = (bv - cb) * inv_ab;
    int u1 = (uc >= (float)u) ? ((uc < (float)ue) ? (int)uc : ue - 1) : u;  // NaN -> u
    while (u1 > u && line_row(ab, cb, u1) != r0) --u1;
    while (u1 + 1 < ue && line_row(ab, cb, u1 + 1) == r0) ++u1;
    return u1;
}

struct LineScan {  // what the sequential reference scan of one epipolar line needs
    float ab, cb, I1, g1, th_epi, apr;
    int u_lo, u_hi;
};

// The reference's own loop, PM.cc:405-444, candidate by candidate with the exact arithmetic
// (err in double because THETA is a double literal).  Used only for the rare pixels whose two best
// matches are closer than the float shortcut of the fast path can separate.  Returns best uj or -1.
__device__ __noinline__ int exact_line_scan(const SearchArgs& A, const NbrConst& N, const LineScan& L) {
    float best = 1000000.0f;
    int bj = -1;
    for (int uj = L.u_lo; uj <= L.u_hi; ++uj) {
        const int ti = __float2int_rz(__fadd_rn(__fmul_rn(L.ab, (float)uj), L.cb));
        if (!(ti < 0 && ti > -A.H)) continue;  // PM.cc:408
        const size_t o = (size_t)(-ti) * A.pitch_f + min(uj, A.W - 1);
        if (__ldg(N.grad + o) < A.lambdaG) continue;  // PM.cc:411
        if (!orientation_gates(__ldg(N.theta + o), L.th_epi, L.apr, A.lambdaL, A.lambdaTheta)) continue;
        const float yv = -(L.ab * (float)uj + L.cb);
        const float rI = L.I1 - bilinear_col<uint8_t>(N.img, A.pitch_u8, A.W, A.H, yv, uj);
        const float rG = L.g1 - bilinear_col<float>(N.grad, A.pitch_f, A.W, A.H, yv, uj);
        const float err = (float)((double)(rI * rI) + (double)(rG * rG) / A.THETA);  // PM.cc:436
        if (err < best) { best = err; bj = uj; }                                      // PM.cc:437-443
    }
    return bj;
}

// ChiTest decided without a division whenever the answer is not within 1e-5 of the threshold:
// q = sigma^2 (float product, as in PM.cc:920-921), r = 1/q.  num*ra + num*rb differs from the
// reference's num/qa + num/qb by < 3e-7 relative, so outside the guard band both give the same
// answer; inside it (or for NaN / inf intermediates) the reference expression itself is evaluated.
__device__ __forceinline__ bool chi_test_q(float a, float b, float qa, float qb, float ra, float rb, float thr, float thr_lo, float thr_hi) {
    const float dl = a - b;
    const float num = dl * dl;
    const float approx = num * ra + num * rb;
    if (fminf(qa, qb) > 1e-30f && fmaxf(qa, qb) < 1e30f) {  // outside, the (approximate) reciprocals over/underflow: the estimate would not track the quotient
        if (approx < thr_lo) return true;
        if (approx > thr_hi) return false;
    }
    return num / qa + num / qb < thr;
}

// InverseDepthHypothesisFusion, PM.cc:598-626 (+ ChiTest 912-924, GetFusion 947-970) on up to NT
// hypothesis slots of one pixel; slot i is a hypothesis iff bit i of `present` is set (the reference's
// vector holds the present ones in the same order, so "first largest set" and the accumulation order
// are unchanged).  ChiTest is symmetric in (a, b) (the squared difference and the float sum commute),
// so each unordered pair is tested once.  NT is a compile-time bound so that the arrays live in
// registers.  Returns false when unsupported.
template <int NT>
__device__ __forceinline__ bool fuse_hypotheses(const float (&hd)[NT], const float (&hs)[NT], unsigned long long present, int lambdaN,
                                                float thr, float& out_d, float& out_s) {
    float hq[NT], hr[NT];
    int cnt[NT];
    const float thr_lo = thr * (1.0f - 1e-5f), thr_hi = thr * (1.0f + 1e-5f);
#pragma unroll(NT <= 16 ? NT : 1)
    for (int i = 0; i < NT; ++i) {
        const float q = hs[i] * hs[i];
        hq[i] = q;
        hr[i] = __fdividef(1.0f, q);  // only steers the division-free estimate of chi_test_q (error ~2e-7 << its 1e-5 guard band)
        // ChiTest(a, a): num = +0, chi = 0/q + 0/q = 0 unless q is 0 or NaN (then NaN -> false)
        cnt[i] = (q > 0.0f && 0.0f < thr) ? 1 : 0;
    }
#pragma unroll(NT <= 16 ? NT : 1)
    for (int i = 0; i < NT; ++i) {
        if (!((present >> i) & 1ull)) continue;
#pragma unroll(NT <= 16 ? NT : 1)
        for (int k = i + 1; k < NT; ++k) {
            if (!((present >> k) & 1ull)) continue;
            if (chi_test_q(hd[i], hd[k], hq[i], hq[k], hr[i], hr[k], thr, thr_lo, thr_hi)) {
                ++cnt[i];
                ++cnt[k];
            }
        }
    }
    int best_cnt = 0, best_a = -1;
    float da = 0.f, qa = 0.f, ra = 0.f;
#pragma unroll(NT <= 16 ? NT : 1)
    for (int i = 0; i < NT; ++i)
        if (((present >> i) & 1ull) && cnt[i] > best_cnt) {  // strict >, first largest set wins (PM.cc:616)
            best_cnt = cnt[i]; best_a = i; da = hd[i]; qa = hq[i]; ra = hr[i];
        }
    if (best_cnt < lambdaN) return false;
    // GetFusion(depthHo), PM.cc:947-970: pow(sigma,2) in double, float accumulators, members in index order
    float pjsj = 0, rsj = 0;
#pragma unroll(NT <= 16 ? NT : 1)
    for (int ib = 0; ib < NT; ++ib) {
        if (!((present >> ib) & 1ull)) continue;
        const bool comp = (ib == best_a) ? (qa > 0.0f && 0.0f < thr) : chi_test_q(da, hd[ib], qa, hq[ib], ra, hr[ib], thr, thr_lo, thr_hi);
        if (!comp) continue;
        const double s2 = (double)hs[ib] * (double)hs[ib];
        pjsj = (float)((double)pjsj + (double)hd[ib] / s2);
        rsj = (float)((double)rsj + 1.0 / s2);
    }
    out_d = pjsj / rsj;
    out_s = sqrtf(1 / rsj);
    return true;
}

#ifndef K1_WARPS_PER_BLOCK
#define K1_WARPS_PER_BLOCK 4  // 4-warp blocks, 12 per SM: the same 48 resident warps as 8 x 6, but a retiring block frees its
#endif                      // registers for the waiting short kernels twice as early (bench +0.5 %)
constexpr int K1_WARPS = K1_WARPS_PER_BLOCK;
constexpr int K1_QHALF = 512;           // what 16 lanes can queue in one scan step (16 lanes x 32 columns)
constexpr int K1_QCAP = K1_QHALF + 32;  // on top of < 32 leftovers; a fuller step is queued in two halves
constexpr int K1_WARP_SMEM = 512 + 256 + 4 * K1_QCAP + 4 * 128 + 16;  // 3 472 bytes per warp
#ifndef K1_ATOMIC_QUEUE
#define K1_ATOMIC_QUEUE 1  // queue positions from one shared-memory atomicAdd per lane instead of a five-step shuffle scan
#endif
#ifndef K1F_MIN_BLOCKS
#define K1F_MIN_BLOCKS 8
#endif
#ifndef K4_MIN_BLOCKS
#define K4_MIN_BLOCKS 8
#endif
#ifndef K1_MIN_BLOCKS
#define K1_MIN_BLOCKS 12
#endif
#define K1_TIE_REL 2e-6f            // float shortcut of err: two candidates closer than this are re-scanned exactly
#define K1_ERR_LIMIT_LO 999000.0f   // PM.cc:437 initial best_match_err = 1e6

// One warp = 32 consecutive entries of the active-pixel list (one lane per pixel).  For each
// neighbour key frame:
//  (A) per lane: epipolar line, slope test, search range, the two gate angles (PM.cc:389-404, 414,
//      424-426) and the orientation-arc plane of the neighbour that matches `th_pi + rot`.
//  (B) per lane: walk the line run by run (a run = consecutive columns on one image row, clipped to
//      a 32-column word of the bit plane): ONE u32 load yields every column of the run that passes
//      condition 1 and can pass condition 3.  Selected columns are appended to a per-warp queue
//      (K1_ATOMIC_QUEUE 1, the default: each lane takes its positions with one shared-memory atomicAdd, the warp total
//      comes from a REDUX; 0: positions from a five-step shuffle prefix sum, no atomics).
//  (C) whenever 32 are queued they are evaluated one per lane, whichever lane they came from: exact
//      conditions 2+3 on GradTheta, photometric + gradient residual (PM.cc:433-436) with the err
//      of PM.cc:436 in float; a shared-memory atomicMin keeps the smallest err per owner, attempted
//      only by candidates that are not already beaten by the owner's current best.  A pixel
//      whose two smallest errs are within K1_TIE_REL of each other is flagged and re-scanned by
//      exact_line_scan, so the winner is always the one the reference's sequential strict-<
//      scan picks (smallest err in the reference arithmetic, first column on ties).
//  (D) per lane: sub-pixel refinement and inverse-depth hypothesis (PM.cc:446-463, 806-829), written
//      to the hypothesis buffer [neighbour][entry]; k1_fuse turns the buffer into the maps.
__global__ void __launch_bounds__(K1_WARPS * 32, K1_MIN_BLOCKS) k1_search(const __grid_constant__ SearchArgs A) {
    // one region per warp: par4 | par2 | queue | beste | bestu | tie | rng
    __shared__ __align__(16) unsigned char s_warp[K1_WARPS][K1_WARP_SMEM];

    const unsigned full = 0xffffffffu;
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    const int count = A.count_override >= 0 ? A.count_override : A.meta->active_count;

    unsigned char* const wsm = s_warp[wib];
    float4* par4 = reinterpret_cast<float4*>(wsm);                 // per owner lane: a/b, c/b, I1, GradImg1
    float2* par2 = reinterpret_cast<float2*>(wsm + 512);           // per owner lane: epipolar-line angle, rotated orientation
    uint32_t* queue = reinterpret_cast<uint32_t*>(wsm + 768);      // K1_QCAP entries
    uint32_t* beste = queue + K1_QCAP;
    uint32_t* bestu = beste + 32;
    uint32_t* tie = bestu + 32;                                    // per owner lane: two best errs too close for the float shortcut
    uint32_t* s_rng = tie + 32;                                    // per lane: u_lo | (u_hi + 1) << 16 -- parked for phase (D), which is register-lean
    uint32_t* s_qn = s_rng + 32;                                   // the queue length `qn` as the lanes' atomicAdd sees it (K1_ATOMIC_QUEUE)
    const float colsf = (float)A.W;
    const float inv_theta_f = A.inv_theta_f;
    const size_t arc_plane = (size_t)A.H * A.arc_pitch;

  // persistent warps: the grid is sized to the machine; every warp draws (32 list entries x 1 neighbour)
  // tasks until none are left.  Tasks are neighbour-major, so at any time the resident warps read the
  // planes of one or two neighbour key frames (L2-resident), and a task is short enough (~1/10 of a
  // pixel's work) that the last warps to finish do not leave the machine idle for long.
  const unsigned nblk = ((unsigned)count + 31u) >> 5;
  const unsigned ntask = nblk * (unsigned)A.n;  // < 2^31: planes hold < 2^30 pixels (checked at upload), n <= 40
  int j = 0;             // neighbour of the current task: a warp's tickets only grow, so j = task / nblk is tracked by
  unsigned jbase = 0;    // stepping (jbase = j * nblk) instead of an integer division per task
  for (;;) {
    unsigned task = 0;
    if (lane == 0) task = atomicAdd(A.task_counter, 1u);
    task = __shfl_sync(full, task, 0);
    if (task >= ntask) {
        // every warp of the grid draws exactly one ticket >= ntask (the one it leaves on); the warp that draws the last
        // of them knows no other draw can follow and puts the counter back to zero for the next launch
        if (lane == 0 && task == ntask + gridDim.x * (unsigned)K1_WARPS - 1u) *A.task_counter = 0u;
        break;
    }
    while (task - jbase >= nblk) { ++j; jbase += nblk; }
    const int entry = (int)(task - jbase) * 32 + lane;
    const bool valid = entry < count;

    int x = 0, y = 0;
    if (valid) {
        uint32_t e = A.list[entry];
        x = (int)(e & 0xffffu);
        y = (int)(e >> 16);
    }
    const float xf = (float)x, yf = (float)y;
    float I1 = 0.f, g1 = 0.f, th1 = 0.f;
    if (valid) {  // the pixel's own record: intensity (PM.cc:202), GradImg (PM.cc:434), GradTheta (PM.cc:214)
        const uint4 own = __ldg(A.rec + (unsigned)(y * A.pitch_f + x));
        I1 = rec_grey_lo(own.w);
        g1 = __uint_as_float(own.y);
        th1 = __uint_as_float(own.x);
    }
    const float xp0 = __ldg(A.ray + x), xp1 = __ldg(A.ray + A.W + y);                 // (x - cx) / fx, (y - cy) / fy: PM.cc:893, 862

    {
        const NbrConst& N = A.nb[j];
        // ---------------- (A) line, slope test, search range: PM.cc:389-393, 877-910 ----------------
        const float a = xf * N.F[0] + yf * N.F[3] + N.F[6];
        const float b = xf * N.F[1] + yf * N.F[4] + N.F[7];
        const float c = xf * N.F[2] + yf * N.F[5] + N.F[8];
        const float ab = a / b, cb = c / b;
        bool ok = valid && !(ab < -A.slope || ab > A.slope);

        // xp2 = gemm(R21, xp1, rho, t21, 1): float dot left to right, epilogue in double
        const float q0 = N.R[0] * xp0 + N.R[1] * xp1 + N.R[2] * 1.0f;
        const float q2 = N.R[6] * xp0 + N.R[7] * xp1 + N.R[8] * 1.0f;
        const double q0d = (double)q0, q2d = (double)q2;
        const float xmin0 = (float)__fma_rn(q0d, A.mind_d, N.td0);
        const float xmin2 = (float)__fma_rn(q2d, A.mind_d, N.td2);
        const float xmax0 = (float)__fma_rn(q0d, A.maxd_d, N.td0);
        const float xmax2 = (float)__fma_rn(q2d, A.maxd_d, N.td2);
        float umin = A.fx * xmin0 / xmin2 + A.cx;
        float umax = A.fx * xmax0 / xmax2 + A.cx;
        if (umin > umax) { float t = umax; umax = umin; umin = t; }
        if (umin < 0) umin = 0;
        if (umax < 0) umax = 0;
        if (umin > colsf) umin = colsf;
        if (umax > colsf) umax = colsf;
        if (!(umin == umin) || !(umax == umax)) ok = false;  // NaN bounds: empty range
        const int u_lo = (int)ceilf(umin);
        const int u_hi = ok ? (int)floorf(umax) : -1;  // candidates: u_lo..u_hi (u_hi may equal W)

        // gate angles, loop invariant: PM.cc:414, 424-426.  th_epi = cv::fastAtan2(-a/b, 1): with x = 1 the
        // quotient min/(max + DBL_EPSILON) is |a/b| / 1 or 1 / |a/b| exactly (the float sum absorbs the epsilon)
        const float inv_ab = 1.0f / ab;
        float th_epi;
        {
            const float p1 = 0.9997878412794807f * (float)(180 / 3.14159265358979323846);
            const float p3 = -0.3258083974640975f * (float)(180 / 3.14159265358979323846);
            const float p5 = 0.1555786518463281f * (float)(180 / 3.14159265358979323846);
            const float p7 = -0.04432655554792128f * (float)(180 / 3.14159265358979323846);
            const float ay = fabsf(ab);
            const bool shallow = 1.0f >= ay;
            const float cq = shallow ? ay : fabsf(inv_ab);
            const float c2 = cq * cq;
            const float pa = (((p7 * c2 + p5) * c2 + p3) * c2 + p1) * cq;
            th_epi = shallow ? pa : 90.f - pa;
            if (-ab < 0) th_epi = 360.f - th_epi;
        }
        float apr = th1 + N.rot;
        if (apr >= 360) apr -= 360;
        if (apr < 0) apr += 360;
        // arc plane of the neighbour for this rotated orientation (plane 32 = condition 1 only)
        int cell = SD_ARC_CELLS;
        if (apr >= 0.f && apr < 360.f) cell = min((int)(apr * (1.0f / SD_ARC_CELL_DEG)), SD_ARC_CELLS - 1);
        const uint32_t* __restrict__ arcp = N.arc + (size_t)cell * arc_plane;

        par4[lane] = make_float4(ab, cb, I1, g1);
        par2[lane] = make_float2(th_epi, apr);
        s_rng[lane] = (uint32_t)(u_lo & 0xffff) | ((uint32_t)(u_hi + 1) << 16);  // 0 <= u_lo <= W, -1 <= u_hi <= W
        beste[lane] = 0x7f800000u;  // +inf
        tie[lane] = 0;
#if K1_ATOMIC_QUEUE
        if (lane == 0) *s_qn = 0u;
#endif
        __syncwarp();

        // ---------------- (C) evaluate queued candidates, one per lane ----------------
        int qn = 0;  // warp-uniform
        auto drain = [&](bool all) {
            while (qn >= 32 || (all && qn > 0)) {
                const int take = min(qn, 32);
                const int base = qn - take;
                const bool have = lane < take;
                bool won = false;
                uint32_t eb = 0;
                int owner = lane, uq = 0;
                if (have) {
                    const uint32_t e = queue[base + lane];
                    owner = (int)(e >> 16);
                    uq = (int)(e & 0xffffu);
                    const float4 p4 = par4[owner];
                    const float2 p2 = par2[owner];
                    float vf, rt;
                    const int r0 = line_row(p4.x, p4.y, small_int_to_float(uq), vf, rt);  // 0 < r0 < H by construction
                    const int xc = min(uq, A.W - 1);
                    // one record = orientation + both rows of both bilinear taps (second row clamped to the image, H2);
                    // requested before the gates are known (3 of 4 queued candidates pass them)
                    const uint4 rc = __ldg(N.rec + (unsigned)(r0 * A.pitch_f + xc));  // planes hold < 2^31 elements (checked at upload)
                    const float T2 = __uint_as_float(rc.x);
                    const float gg0 = __uint_as_float(rc.y), gg1 = __uint_as_float(rc.z);
                    const float i0 = rec_grey_lo(rc.w), i1 = rec_grey_hi(rc.w);
                    const float cur = __uint_as_float(beste[owner]);
                    if (orientation_gates(T2, p2.x, p2.y, A.lambdaL, A.lambdaTheta)) {
                        const float yv = -vf;
                        const float w0 = (rt - 12582911.0f) - yv;  // ((float)r0 + 1) - yv: bilinear_col weights, PM.cc:46-53
                        const float w1 = 1.0f - w0;
                        const float rI = p4.z - (i0 * w0 + i1 * w1);
                        const float rG = p4.w - (gg0 * w0 + gg1 * w1);
                        const float ef = rI * rI + (rG * rG) * inv_theta_f;  // float shortcut of PM.cc:436
                        if (ef < K1_ERR_LIMIT_LO) {
                            // a candidate above the owner's current best by more than the tie band can neither
                            // win nor tie with the final best (which is <= cur): no atomic needed
                            if (!(ef > __fmaf_rn(cur, 2.0f * K1_TIE_REL, cur))) {
                                eb = __float_as_uint(ef);
                                const uint32_t old = atomicMin(&beste[owner], eb);
                                const float fo = __uint_as_float(old);
                                const float lo = fminf(ef, fo), hi = fmaxf(ef, fo);
                                if (hi - lo <= __fmaf_rn(lo, K1_TIE_REL, 1e-30f)) tie[owner] = 1u;  // benign race: every writer stores 1
                                won = eb < old;
                            }
                        } else {  // never with u8 images (err <= ~2.1e5); keep the 1e6 limit exact anyway
                            tie[owner] = 1u;
                        }
                    }
                }
                __syncwarp();
                // of the candidates that lowered an owner's best in this round, the one whose err is still the best owns bestu
                const uint32_t now = won ? beste[owner] : ~eb;
                if (now == eb) bestu[owner] = (uint32_t)uq;
#if K1_ATOMIC_QUEUE
                if (lane == 0) *s_qn = (uint32_t)base;  // what the next push's atomicAdd starts from
#endif
                __syncwarp();
                qn = base;
            }
        };

        // ---------------- (B) run-by-run scan of the arc plane ----------------
        // Queue positions.  K1_ATOMIC_QUEUE: every lane with candidates takes its slice of the queue with ONE shared-memory
        // atomicAdd (s_qn mirrors qn) and the warp total is one REDUX -- ~6 instructions instead of the ~28 of a shuffle
        // prefix sum.  The slices then come in arbitrary lane order, which no result depends on: the per-owner minimum is
        // order-free, and of two candidates within the tie band of the final best the second to reach the atomicMin
        // always sees the other one (or something between them) as `old` and flags the owner for the exact re-scan.
        auto push = [&](uint32_t m, int col_base) {
            const int cnt = __popc(m);
            const uint32_t tag = ((uint32_t)lane << 16) | (uint32_t)col_base;
#if K1_ATOMIC_QUEUE
            const int total = __reduce_add_sync(full, cnt);
            if (total <= K1_QHALF) {
                if (m) {
                    uint32_t* qp = queue + atomicAdd(s_qn, (uint32_t)cnt);
                    do {  // two per trip, from the top bit down, branch-free: bfind (FLO) gives the index directly and is
                          // 0xffffffff for m == 0, which shl.b32 clamps to a shift by 32, i.e. to 0 -- m stays 0
                        uint32_t b0, b1, s0, s1;
                        asm("bfind.u32 %0, %1;" : "=r"(b0) : "r"(m));
                        asm("shl.b32 %0, 1, %1;" : "=r"(s0) : "r"(b0));
                        m ^= s0;
                        qp[0] = tag + b0;
                        asm("bfind.u32 %0, %1;" : "=r"(b1) : "r"(m));
                        asm("shl.b32 %0, 1, %1;" : "=r"(s1) : "r"(b1));
                        if (m) qp[1] = tag + b1;
                        m ^= s1;
                        qp += 2;
                    } while (m);
                }
                qn += total;
                __syncwarp();
                return;
            }
#endif
            int incl = cnt;
#pragma unroll
            for (int dlt = 1; dlt < 32; dlt <<= 1) {
                const int v = __shfl_up_sync(full, incl, dlt);
                if (lane >= dlt) incl += v;
            }
#if !K1_ATOMIC_QUEUE
            const int total = __shfl_sync(full, incl, 31);
#endif
            int pos = qn + incl - cnt;
            if (total <= K1_QHALF) {
                while (m) {
                    const int bit = __ffs((int)m) - 1;
                    m &= m - 1;
                    queue[pos++] = tag + (uint32_t)bit;
                }
                qn += total;
                __syncwarp();
            } else {  // rare: more than the queue holds at once -- lanes 0..15, evaluate, then lanes 16..31
                const int half = __shfl_sync(full, incl, 15);
                if (lane < 16)
                    while (m) {
                        const int bit = __ffs((int)m) - 1;
                        m &= m - 1;
                        queue[pos++] = tag + (uint32_t)bit;
                    }
                qn += half;
                __syncwarp();
                drain(false);
                pos = qn + incl - cnt - half;
                if (lane >= 16)
                    while (m) {
                        const int bit = __ffs((int)m) - 1;
                        m &= m - 1;
                        queue[pos++] = tag + (uint32_t)bit;
                    }
                qn += total - half;
#if K1_ATOMIC_QUEUE
                if (lane == 0) *s_qn = (uint32_t)qn;
#endif
                __syncwarp();
            }
        };

        const int u_end = min(u_hi, A.W - 1);
        int u = u_lo;
        while (__any_sync(full, u <= u_end)) {
            uint32_t m = 0;
            int col_base = 0;
            if (u <= u_end) {
                int ue = min(u | 31, u_end);
                const int r0 = line_row(ab, cb, u);
                if (ue > u && line_row(ab, cb, ue) != r0) ue = run_end(ab, cb, inv_ab, u, ue, r0);
                if (r0 > 0 && r0 < A.H) {  // PM.cc:408: 0 < vj < rows
                    const uint32_t w = __ldg(arcp + (unsigned)(r0 * A.arc_pitch + (u >> 5)));
                    m = w & (0xffffffffu << (u & 31)) & (0xffffffffu >> (31 - (ue & 31)));
                }
                col_base = u & ~31;
                u = ue + 1;
            }
            if (__any_sync(full, m != 0u)) {
                push(m, col_base);
                drain(false);
            }
        }
        // candidate column uj == cols reads column cols-1 (clamp-to-edge)
        const bool extra = u_hi >= A.W && u_lo <= A.W;
        if (__any_sync(full, extra)) {
            uint32_t m = 0;
            if (extra) {
                const int r0 = line_row(ab, cb, A.W);
                if (r0 > 0 && r0 < A.H) {
                    const uint32_t w = __ldg(arcp + (size_t)r0 * A.arc_pitch + ((A.W - 1) >> 5));
                    m = (w >> ((A.W - 1) & 31)) & 1u;
                }
            }
            push(m, A.W);
        }
        drain(true);
        __syncwarp();

        // ---------------- (D) refinement + hypothesis: PM.cc:446-463, 806-829 ----------------
        // everything phase (D) needs was parked in shared memory by phase (A): the scan and the evaluation above run
        // with few live registers (6 resident blocks per SM), and (D) reloads instead of keeping values alive
        const float4 own4 = par4[lane];
        const float abD = own4.x, cbD = own4.y, I1D = own4.z, g1D = own4.w;
        int bj = -1;
        if (tie[lane]) {
            const float2 own2 = par2[lane];
            const uint32_t rg = s_rng[lane];
            LineScan L;
            L.ab = abD; L.cb = cbD; L.I1 = I1D; L.g1 = g1D; L.th_epi = own2.x; L.apr = own2.y;
            L.u_lo = (int)(rg & 0xffffu); L.u_hi = (int)(rg >> 16) - 1;
            bj = exact_line_scan(A, N, L);
        } else if (beste[lane] != 0x7f800000u) {
            bj = (int)bestu[lane];
        }
        float rho = -1.0f, sig = 0.f;  // (-1, 0): "no hypothesis" -- fails the 1/rho > 0 test of PM.cc:216 like any negative rho
        float best_u = 0.f;
        const bool found = bj >= 0;
        if (found) {
            const float yb = -(abD * (float)bj + cbD);
            float bI, bG, pI, pG, mI, mG;
            bilinear_rec(N.rec, A.pitch_f, A.W, A.H, yb, bj, bI, bG);
            const float rI = I1D - bI;
            const float rG = g1D - bG;
            const int up = bj + 1, um = bj - 1;
            const float yp = -(abD * (float)up + cbD), ym = -(abD * (float)um + cbD);
            bilinear_rec(N.rec, A.pitch_f, A.W, A.H, yp, up, pI, pG);
            bilinear_rec(N.rec, A.pitch_f, A.W, A.H, ym, um, mI, mG);
            const float g = (pI - mI) / 2;
            const float q = (pG - mG) / 2;
            const double itq = A.inv_THETA * (double)q;                           // (1/THETA)*q
            const float den = (float)((double)(g * g) + itq * (double)q);         // PM.cc:455
            const float ustar = (float)((double)bj + ((double)(g * rI) + itq * (double)rG) / (double)den);  // PM.cc:456
            best_u = ustar;  // PM.cc:459
            const float istd = N.meta->istddev;
            const float uvar = 2 * istd * istd / den;  // PM.cc:457
            // ComputeInvDepthHypothesis (xp reloaded through the list entry: phase (A)'s values are not kept alive)
            const uint32_t ent = A.list[entry];
            const double xp0d = (double)__ldg(A.ray + (ent & 0xffffu));
            const double xp1d = (double)__ldg(A.ray + A.W + (ent >> 16));
            const double d2 = N.Rd2[2] + __fma_rn(N.Rd2[1], xp1d, N.Rd2[0] * xp0d);
            const double d0 = N.Rd0[2] + __fma_rn(N.Rd0[1], xp1d, N.Rd0[0] * xp0d);
            const float num2f = (float)(d0 * (double)A.fx);
            const float den2 = A.fx * N.t[0];
            const float sq = sqrtf(uvar);
            rho = pixel_inv_depth(ustar, A.cx, d2, num2f, N.t[2], den2);
            const float rmin = pixel_inv_depth(ustar - sq, A.cx, d2, num2f, N.t[2], den2);
            const float rmax = pixel_inv_depth(ustar + sq, A.cx, d2, num2f, N.t[2], den2);
            const float e1 = fabsf(rmax - rho), e2 = fabsf(rmin - rho);
            sig = (e1 < e2) ? e2 : e1;  // std::max(e1, e2)
        }
        if (valid) {
            if (A.hyp) A.hyp[(size_t)j * A.hyp_stride + entry] = make_float2(rho, sig);
            if (A.probe_out && j == 0) {
                float* o = A.probe_out + (size_t)entry * 6;
                o[0] = found ? rho : 0.f; o[1] = sig; o[2] = (float)bj; o[3] = found ? 1.f : 0.f;
                o[4] = best_u; o[5] = found ? -(abD * best_u + cbD) : 0.f;  // PM.cc:459-460
            }
        }
        __syncwarp();
    }
  }  // task loop
}

// InverseDepthHypothesisFusion over the hypothesis buffer of k1_search, PM.cc:216-228: one thread per
// active pixel; a hypothesis counts if 1/rho > 0 (PM.cc:216); more than lambdaN of them are fused
// (PM.cc:221) and the result is stored if supported (PM.cc:224-227).
struct FuseArgs {
    int n, lambdaN, pitch_f, hyp_stride;
    float chi_fuse;
    const float2* hyp;
    const uint32_t* list;
    KfMeta* meta;  // active_count is read; valid_count is zeroed for the k23_intra launch that follows on the same stream
    float* depth;
    float* sigma;
};

template <int NT>
__global__ void __launch_bounds__(128, K1F_MIN_BLOCKS) k1_fuse(const FuseArgs F) {
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e == 0) F.meta->valid_count = 0;
    if (e >= F.meta->active_count) return;
    float hd[NT], hs[NT];
    unsigned long long present = 0;
    int nh = 0;
#pragma unroll(NT <= 16 ? NT : 1)
    for (int j = 0; j < NT; ++j) {
        hd[j] = 0.f; hs[j] = 0.f;
        if (j < F.n) {
            const float2 h = __ldg(F.hyp + (size_t)j * F.hyp_stride + e);
            hd[j] = h.x; hs[j] = h.y;
            // PM.cc:216 keeps a hypothesis if `1 / depth > 0`: true exactly for +0 and the positive finite floats
            // (1/+0 = +inf, 1/+inf = 0, NaN compares false), i.e. for bit patterns below +inf -- no division needed
            if (__float_as_uint(h.x) < 0x7f800000u) { present |= 1ull << j; ++nh; }
        }
    }
    if (nh > F.lambdaN) {
        float fd, fs;
        if (fuse_hypotheses<NT>(hd, hs, present, F.lambdaN, F.chi_fuse, fd, fs)) {
            const uint32_t ent = F.list[e];
            const size_t o = (size_t)(ent >> 16) * F.pitch_f + (ent & 0xffffu);
            F.depth[o] = fd;
            F.sigma[o] = fs;
        }
    }
}

// k1_fuse for many neighbours (n > 16): the register-array version above would need > 128 registers (or local
// memory: 287 us at n = 20, 1 031 us at n = 40).  Here a thread keeps its pixel's hypotheses and their compatibility
// counts in its own column of shared memory and walks the ChiTests row by row: the row's values and its count live in
// registers, the partner's q = sigma^2 and its reciprocal estimate are recomputed (one FMUL + one MUFU) instead of
// stored, each unordered pair is tested once (ChiTest is symmetric) and credits the partner's count in shared memory.
// Rows are visited in ascending order, so a row's count is complete when the row is reached: "first largest set"
// (PM.cc:616) is decided on the fly.  No __syncthreads: a thread only ever touches its own column.
constexpr int K1FR_THREADS = 64;
__host__ __device__ constexpr size_t k1fr_smem_bytes(int n) { return (size_t)n * K1FR_THREADS * (2 * sizeof(float) + 1); }

__global__ void __launch_bounds__(K1FR_THREADS) k1_fuse_rows(const FuseArgs F) {
    extern __shared__ float k1fr_smem[];
    const int T = K1FR_THREADS, tid = threadIdx.x, n = F.n;
    float* sd = k1fr_smem + tid;                                                 // sd[j * T]: inverse depth of hypothesis j
    float* ss = k1fr_smem + n * T + tid;                                         // ss[j * T]: sigma
    unsigned char* sc = reinterpret_cast<unsigned char*>(k1fr_smem + 2 * n * T) + tid;  // sc[j * T]: compatible hypotheses so far
    const int e = blockIdx.x * T + tid;
    if (e == 0) F.meta->valid_count = 0;
    if (e >= F.meta->active_count) return;
    const float thr = F.chi_fuse, thr_lo = thr * (1.0f - 1e-5f), thr_hi = thr * (1.0f + 1e-5f);
    unsigned long long present = 0;
    int nh = 0;
    for (int j = 0; j < n; ++j) {
        const float2 h = __ldg(F.hyp + (size_t)j * F.hyp_stride + e);
        sd[j * T] = h.x;
        ss[j * T] = h.y;
        sc[j * T] = (h.y * h.y > 0.0f && 0.0f < thr) ? 1 : 0;                    // ChiTest(a, a), see fuse_hypotheses
        if (__float_as_uint(h.x) < 0x7f800000u) { present |= 1ull << j; ++nh; }  // PM.cc:216, see k1_fuse
    }
    if (nh <= F.lambdaN) return;  // PM.cc:221
    int best_cnt = 0, best_a = -1;
    for (unsigned long long mi = present; mi; mi &= mi - 1) {
        const int i = __ffsll((long long)mi) - 1;
        const float di = sd[i * T], si = ss[i * T];
        const float qi = si * si, ri = __fdividef(1.0f, qi);
        int c = sc[i * T];
        for (unsigned long long mk = mi & (mi - 1); mk; mk &= mk - 1) {  // present hypotheses after i
            const int k = __ffsll((long long)mk) - 1;
            const float sk = ss[k * T];
            const float qk = sk * sk;
            if (chi_test_q(di, sd[k * T], qi, qk, ri, __fdividef(1.0f, qk), thr, thr_lo, thr_hi)) {
                ++c;
                ++sc[k * T];
            }
        }
        if (c > best_cnt) { best_cnt = c; best_a = i; }  // strict >, first largest set wins (PM.cc:616)
    }
    if (best_cnt < F.lambdaN) return;  // PM.cc:623: unsupported
    const float da = sd[best_a * T], sa = ss[best_a * T];
    const float qa = sa * sa, ra = __fdividef(1.0f, qa);
    // GetFusion(depthHo), PM.cc:947-970: pow(sigma,2) in double, float accumulators, members in index order
    float pjsj = 0, rsj = 0;
    for (unsigned long long mb = present; mb; mb &= mb - 1) {
        const int ib = __ffsll((long long)mb) - 1;
        const float db = sd[ib * T], sb = ss[ib * T];
        const float qb = sb * sb;
        const bool comp = (ib == best_a) ? (qa > 0.0f && 0.0f < thr) : chi_test_q(da, db, qa, qb, ra, __fdividef(1.0f, qb), thr, thr_lo, thr_hi);
        if (!comp) continue;
        const double s2 = (double)sb * (double)sb;
        pjsj = (float)((double)pjsj + (double)db / s2);
        rsj = (float)((double)rsj + 1.0 / s2);
    }
    const uint32_t ent = F.list[e];
    const size_t o = (size_t)(ent >> 16) * F.pitch_f + (ent & 0xffffu);
    F.depth[o] = pjsj / rsj;
    F.sigma[o] = sqrtf(1 / rsj);
}

// fusion probe: one thread, same device function as k1_fuse
__global__ void k1_fusion_probe(const float* d, const float* s, int n, int lambdaN, float chi_fuse, float* out) {
    float hd[SD_MAX_NEIGHBOURS], hs[SD_MAX_NEIGHBOURS];
    unsigned long long present = 0;
    for (int i = 0; i < SD_MAX_NEIGHBOURS; ++i) {
        hd[i] = i < n ? d[i] : 0.f;
        hs[i] = i < n ? s[i] : 0.f;
        if (i < n) present |= 1ull << i;
    }
    out[0] = out[1] = out[2] = 0.f;
    float fd, fs;
    if (fuse_hypotheses<SD_MAX_NEIGHBOURS>(hd, hs, present, lambdaN, chi_fuse, fd, fs)) {
        out[0] = fd;
        out[1] = fs;
        out[2] = 1.f;
    }
}

// ------------------------------------------------------------------------------------------------
// K2+K3: IntraKeyFrameDepthChecking (PM.cc:486-547) followed by IntraKeyFrameDepthGrowing
// (PM.cc:549-596), one pass.  A 64x16 output tile per block; the inputs are staged in shared memory
// with a 2-pixel halo, the checking stage is evaluated on the tile + 1-pixel halo (the growing stage
// of a tile pixel reads the CHECKED maps of its 8 neighbours), the growing stage on the tile.
//
// Only ~1/4 of the pixels carry a depth, so the block first compacts the valid pixels of its input
// region into a shared-memory list and runs the checking stage over the list (dense lanes).  The
// GetFusion terms of a hypothesis (PM.cc:956-958: d/pow(sigma,2) and 1/pow(sigma,2), in double) do not
// depend on which centre pixel asks for them, so they are computed once per valid pixel instead of
// once per (centre, neighbour) pair.
//
// The growing stage acts on pixels WITHOUT a depth.  ChiTest(neighbour, centre) divides by the
// centre's sigma^2; with sigma == 0 (what every invalid pixel has inside the pipeline: K1 never
// writes it and the checking stage zeroes what it rejects) the statistic is +inf or NaN and nothing
// is compatible, so those pixels are skipped without touching `grad`.  Any other
// sigma takes the general path, which is the reference's loop verbatim.
// Template flags select the stages (the two-stage pipeline, or one stage alone for the
// caller-plane entry points sd_intra_depth_check_planes / sd_intra_depth_grow_planes).
// ------------------------------------------------------------------------------------------------
constexpr int K23_TW = 64, K23_TH = 16, K23_THREADS = 256;
constexpr int K23_IW = K23_TW + 4, K23_IH = K23_TH + 4;
constexpr int K23_NPIX = K23_IW * K23_IH;  // 1360
constexpr size_t K23_SMEM = (size_t)K23_NPIX * (2 * sizeof(double) + 5 * sizeof(float) + sizeof(unsigned short));

template <bool CHECK, bool GROW>
__global__ void __launch_bounds__(K23_THREADS) k23_intra(const IntraArgs A) {
    extern __shared__ __align__(16) unsigned char k23_smem[];
    double* s_r = reinterpret_cast<double*>(k23_smem);  // 1 / sigma^2 (double)
    double* s_dq = s_r + K23_NPIX;                      // d / sigma^2 (double)
    float* s_d = reinterpret_cast<float*>(s_dq + K23_NPIX);
    float* s_s = s_d + K23_NPIX;
    float* s_d2 = s_s + K23_NPIX;  // after the checking stage
    float* s_s2 = s_d2 + K23_NPIX;
    float* s_rq = s_s2 + K23_NPIX;  // 1 / (sigma*sigma) in float: only steers the division-free chi test
    unsigned short* s_list = reinterpret_cast<unsigned short*>(s_rq + K23_NPIX);
    __shared__ int s_n, s_vbase;
    __shared__ int s_wcnt[K23_THREADS / 32];

    const unsigned full = 0xffffffffu;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int x0 = blockIdx.x * K23_TW - 2, y0 = blockIdx.y * K23_TH - 2;
    if (tid == 0) s_n = 0;
    __syncthreads();

    // ---- stage inputs (halo 2), compact the valid pixels of the region ----
    for (int i0 = 0; i0 < K23_NPIX; i0 += K23_THREADS) {
        const int i = i0 + tid;
        bool valid = false;
        if (i < K23_NPIX) {
            const int r = i / K23_IW, c = i - r * K23_IW;
            const int y = y0 + r, x = x0 + c;
            float d = 0.f, s = 0.f;
            if (x >= 0 && x < A.W && y >= 0 && y < A.H) {
                const size_t o = (size_t)y * A.pitch_in + x;
                d = __ldg(A.depth_in + o);
                s = __ldg(A.sigma_in + o);
                valid = d > SD_EPS_F;
            }
            s_d[i] = d; s_s[i] = s; s_d2[i] = d; s_s2[i] = s;
        }
        if (CHECK) {
            const unsigned bal = __ballot_sync(full, valid);
            int base = 0;
            if (lane == 0 && bal) base = atomicAdd(&s_n, __popc(bal));
            base = __shfl_sync(full, base, 0);
            if (valid) s_list[base + __popc(bal & ((1u << lane) - 1u))] = (unsigned short)i;
        }
    }
    __syncthreads();

    if (CHECK) {
        const int n = s_n;
        // ---- per valid pixel: the GetFusion terms (PM.cc:956-958) and the float reciprocal ----
        for (int k = tid; k < n; k += K23_THREADS) {
            const int i = s_list[k];
            const float d = s_d[i], s = s_s[i];
            const double sq = (double)s * (double)s;  // pow(sigma, 2)
            s_r[i] = 1.0 / sq;
            s_dq[i] = (double)d / sq;
            s_rq[i] = __fdividef(1.0f, s * s);  // steers chi_test_q only
        }
        __syncthreads();
        // ---- checking stage over the list (tile + 1-pixel halo, image interior only) ----
        const float thr = A.chi_fuse, thr_lo = thr * (1.0f - 1e-5f), thr_hi = thr * (1.0f + 1e-5f);
        for (int k = tid; k < n; k += K23_THREADS) {
            const int i = s_list[k];
            const int r = i / K23_IW, c = i - r * K23_IW;
            const int y = y0 + r, x = x0 + c;
            if (r < 1 || r > K23_IH - 2 || c < 1 || c > K23_IW - 2) continue;
            if (x < A.border || x >= A.W - A.border || y < A.border || y >= A.H - A.border) continue;
            const float dc = s_d[i], sc = s_s[i];
            const float qc = sc * sc, rc = s_rq[i];
            // compatible neighbours in raster order, then the pixel itself (PM.cc:504-522), fused by GetFusion
            float pjsj = 0, rsj = 0, min_sigma = 0;
            int ncomp = 0;
#pragma unroll
            for (int t = 0; t < 9; ++t) {
                if (t == 4) continue;
                const int j = i + (t / 3 - 1) * K23_IW + (t % 3 - 1);
                const float d = s_d[j];
                if (d > SD_EPS_F) {
                    const float s = s_s[j];
                    if (chi_test_q(d, dc, s * s, qc, s_rq[j], rc, thr, thr_lo, thr_hi)) {
                        pjsj = (float)((double)pjsj + s_dq[j]);
                        rsj = (float)((double)rsj + s_r[j]);
                        // temp_min_sigma starts at compatible_ho[0].sigma; pow(s,2) < pow(min,2) in double
                        // (exact squares) <=> |s| < |min|
                        if (ncomp == 0 || fabsf(s) < fabsf(min_sigma)) min_sigma = s;
                        ++ncomp;
                    }
                }
            }
            pjsj = (float)((double)pjsj + s_dq[i]);
            rsj = (float)((double)rsj + s_r[i]);
            if (ncomp == 0 || fabsf(sc) < fabsf(min_sigma)) min_sigma = sc;
            ++ncomp;
            if (ncomp >= 3) {
                s_d2[i] = pjsj / rsj;
                s_s2[i] = min_sigma;  // PM.cc:531: min sigma, not the fused sigma
            } else {
                s_d2[i] = 0.0f;
                s_s2[i] = 0.0f;
            }
        }
        __syncthreads();
    }

    // ---- growing stage + outputs: 4 consecutive pixels of one tile row per thread ----
    const int trow = tid >> 4, tcol = (tid & 15) * 4;
    const int y = y0 + 2 + trow, xb = x0 + 2 + tcol;
    float od[4], os[4];
    int nvalid = 0;
    uint32_t ventry[4];
#pragma unroll
    for (int p = 0; p < 4; ++p) {
        const int i = (trow + 2) * K23_IW + tcol + 2 + p;
        const int x = xb + p;
        float dn = s_d2[i], sn = s_s2[i];
        const bool inimg = x < A.W && y < A.H;
        if (GROW && inimg && x >= A.border && x < A.W - A.border && y >= A.border && y < A.H - A.border && dn <= SD_EPS_F &&
            sn != 0.0f) {  // sigma == 0: chi = +inf / NaN for every neighbour, nothing to grow (see above)
            if (!(__ldg(A.grad + (size_t)y * A.pitch_in + x) < A.lambdaG)) {
                float pjsj = 0, rsj = 0, min_sigma = 0;
                int nsup = 0;
                for (int t = 0; t < 9; ++t) {
                    if (t == 4) continue;
                    const int j = i + (t / 3 - 1) * K23_IW + (t % 3 - 1);
                    const float d = s_d2[j], s = s_s2[j];
                    if (chi_test(d, dn, s, sn, A.chi_fuse)) {
                        // GetFusion(pairs), PM.cc:926-945
                        const double sq = (double)s * (double)s;
                        pjsj = (float)((double)pjsj + (double)d / sq);
                        rsj = (float)((double)rsj + 1.0 / sq);
                        if (nsup == 0 || s < min_sigma) min_sigma = s;
                        ++nsup;
                    }
                }
                if (nsup >= 2) {
                    dn = pjsj / rsj;
                    sn = min_sigma;
                }
            }
        }
        od[p] = dn; os[p] = sn;
        // every pixel that carries anything: the ones k4_inter_check works on (depth > 1e-6, PM.cc:662) plus, for the
        // compact download, the (in practice non-existent) ones with a depth in (0, 1e-6] or a sigma without a depth
        if (inimg && x >= 2 && x < A.W - 2 && y >= 2 && y < A.H - 2 && (dn != 0.0f || sn != 0.0f)) ventry[nvalid++] = ((uint32_t)y << 16) | (uint32_t)x;
    }
    if (y < A.H) {
        const size_t o = (size_t)y * A.pitch_out + xb;
        if (A.vec_ok && xb + 3 < A.W) {
            *reinterpret_cast<float4*>(A.depth_out + o) = make_float4(od[0], od[1], od[2], od[3]);
            *reinterpret_cast<float4*>(A.sigma_out + o) = make_float4(os[0], os[1], os[2], os[3]);
            if (A.depth_copy) *reinterpret_cast<float4*>(A.depth_copy + o) = make_float4(od[0], od[1], od[2], od[3]);
        } else {
            for (int p = 0; p < 4; ++p)
                if (xb + p < A.W) {
                    A.depth_out[o + p] = od[p];
                    A.sigma_out[o + p] = os[p];
                    if (A.depth_copy) A.depth_copy[o + p] = od[p];
                }
        }
    }
    if (A.vlist) {  // block-wide compaction of the valid pixels: warp scan, one atomicAdd per block
        int incl = nvalid;
#pragma unroll
        for (int dlt = 1; dlt < 32; dlt <<= 1) {
            const int v = __shfl_up_sync(full, incl, dlt);
            if (lane >= dlt) incl += v;
        }
        if (lane == 31) s_wcnt[wid] = incl;
        __syncthreads();
        if (tid == 0) {
            int tot = 0;
            for (int w = 0; w < K23_THREADS / 32; ++w) {
                const int cw = s_wcnt[w];
                s_wcnt[w] = tot;
                tot += cw;
            }
            s_vbase = tot ? atomicAdd(A.vcount, tot) : 0;
        }
        __syncthreads();
        int pos = s_vbase + s_wcnt[wid] + incl - nvalid;
        for (int p = 0; p < nvalid; ++p) A.vlist[pos + p] = ventry[p];
    }
}

// valid-pixel list of a depth plane that did not come out of k23_intra (sd_import_maps)
__global__ void __launch_bounds__(256) k_build_vlist(const float* __restrict__ depth, const float* __restrict__ sigma, int pitch_f, int W, int H,
                                                     uint32_t* __restrict__ vlist, int* __restrict__ vcount) {
    const int x = blockIdx.x * 32 + (threadIdx.x & 31), y = blockIdx.y * 8 + (threadIdx.x >> 5);
    const bool v = x >= 2 && x < W - 2 && y >= 2 && y < H - 2 &&
                   (depth[(size_t)y * pitch_f + x] != 0.0f || sigma[(size_t)y * pitch_f + x] != 0.0f);
    const unsigned bal = __ballot_sync(0xffffffffu, v);
    int base = 0;
    if ((threadIdx.x & 31) == 0 && bal) base = atomicAdd(vcount, __popc(bal));
    base = __shfl_sync(0xffffffffu, base, 0);
    if (v) vlist[base + __popc(bal & ((1u << (threadIdx.x & 31)) - 1u))] = ((uint32_t)y << 16) | (uint32_t)x;
}

// ------------------------------------------------------------------------------------------------
// K4: InterKeyFrameDepthChecking, PM.cc:628-799, snapshot semantics.  One thread per
// entry of the valid-pixel list (the pixels PM.cc:662 does not skip); `depth_out` already holds a
// copy of `depth_in`, so only these pixels are written.
//
// Arithmetic notes (all bit-exact with the reference's OpenCV expressions):
//  * (float)(1.0 / (double)x) == 1.0f / x: rounding a double-rounded quotient of floats to float is
//    innocuous (53 >= 2*24 + 2 bits), so the reciprocals OpenCV takes in double are IEEE float divisions.
//  * the chi-square statistic of PM.cc:709-710, (float)(pow(dj - d, 2) / pow(sigma, 2)) < 3.84, is decided
//    from a float estimate (relative error < 4e-7) unless it lies within 1e-5 of the threshold or an
//    operand is tiny; then the double expression itself is evaluated.
// ------------------------------------------------------------------------------------------------
// world point of pixel (x, y) with inverse depth inv_d, PM.cc:349-363; zeros outside the 2-px frame or without depth
__device__ __forceinline__ void backproject(const PointArgs& A, int x, int y, float inv_d, float& p0, float& p1, float& p2) {
    p0 = p1 = p2 = 0.f;
    if (x >= 2 && x < A.W - 2 && y >= 2 && y < A.H - 2 && inv_d > SD_EPS_F) {
        const float Z = 1 / inv_d;
        const float X = Z * ((float)x - A.cx) / A.fx;
        const float Y = Z * ((float)y - A.cy) / A.fy;
        p0 = A.Twc[0] * X + A.Twc[1] * Y + A.Twc[2] * Z + A.Twc[3] * 1.0f;
        p1 = A.Twc[4] * X + A.Twc[5] * Y + A.Twc[6] * Z + A.Twc[7] * 1.0f;
        p2 = A.Twc[8] * X + A.Twc[9] * Y + A.Twc[10] * Z + A.Twc[11] * 1.0f;
    }
}

struct K4Proj {  // projection of one pixel into one neighbour + the four depth taps, PM.cc:675-705
    bool inb;
    float rz, depthj;
    float d[4], s[4];
};

__device__ __forceinline__ void k4_project(const InterArgs& A, const InterNbr& N, float xp0, float xp1, double xp0d, double xp1d,
                                           float depthp, double inv_dp, K4Proj& P) {
    // temp = gemm(Rji, xp, 1/depthp, tji, 1); Xj = K*temp (small path, float left to right)
    const float q0 = N.R[0] * xp0 + N.R[1] * xp1 + N.R[2] * 1.0f;
    const float q1 = N.R[3] * xp0 + N.R[4] * xp1 + N.R[5] * 1.0f;
    const float q2 = N.R[6] * xp0 + N.R[7] * xp1 + N.R[8] * 1.0f;
    const float t0 = (float)__fma_rn((double)q0, inv_dp, N.td[0]);
    const float t1 = (float)__fma_rn((double)q1, inv_dp, N.td[1]);
    const float t2 = (float)__fma_rn((double)q2, inv_dp, N.td[2]);
    const float X0 = N.fx * t0 + 0.0f * t1 + N.cx * t2;
    const float X1 = 0.0f * t0 + N.fy * t1 + N.cy * t2;
    const float X2 = 0.0f * t0 + 0.0f * t1 + 1.0f * t2;
    const float inv = 1.0f / X2;  // Xj/Xj(2): convertTo with alpha = 1./s
    const float xj = X0 * inv, yj = X1 * inv;
    P.inb = xj >= 0 && xj < (float)(A.W - 1) && yj >= 0 && yj < (float)(A.H - 1);  // PM.cc:695
    // Eq. 12, PM.cc:684-688
    P.rz = (float)(N.Rd2[2] + __fma_rn(N.Rd2[1], xp1d, N.Rd2[0] * xp0d));
    P.depthj = depthp / (P.rz + depthp * N.t[2]);
    P.d[0] = P.d[1] = P.d[2] = P.d[3] = 0.f;
    P.s[0] = P.s[1] = P.s[2] = P.s[3] = 0.f;
    if (P.inb) {
        const int x0 = (int)xj, y0 = (int)yj;  // floor of a non-negative float
        const size_t q = (size_t)y0 * A.pitch_f + x0;
        const float* dpl = N.depth + q;
        const float* spl = N.sigma + q;
        P.d[0] = __ldg(dpl);                  // taps (y0,x0) (y1,x0) (y0,x1) (y1,x1), PM.cc:705-752
        P.d[1] = __ldg(dpl + A.pitch_f);
        P.d[2] = __ldg(dpl + 1);
        P.d[3] = __ldg(dpl + A.pitch_f + 1);
        // the sigma taps are requested with the depth taps, not after them: a load that waits for `d > 1e-6` is a
        // second dependent trip to L2 per neighbour, and this kernel is bound by exactly that latency
        P.s[0] = __ldg(spl);
        P.s[1] = __ldg(spl + A.pitch_f);
        P.s[2] = __ldg(spl + 1);
        P.s[3] = __ldg(spl + A.pitch_f + 1);
    }
}

// POINTS: also UpdateSemiDensePointSet (PM.cc:337-367, called right after the check at PM.cc:302-304) for the pixels
// that keep a depth; the point plane was zeroed before the launch, which is what every other pixel gets (PM.cc:345-347).
template <bool POINTS>
__global__ void __launch_bounds__(128, K4_MIN_BLOCKS) k4_inter_check(const __grid_constant__ InterArgs A) {
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= *A.vcount) return;
    const uint32_t ent = A.vlist[e];
    const int px = (int)(ent & 0xffffu), py = (int)(ent >> 16);
    const size_t o = (size_t)py * A.pitch_f + px;
    const float depthp = A.depth_in[o];
    if (!(depthp > SD_EPS_F)) return;  // PM.cc:662 (listed only because it carries a sigma)
    const float xp0 = ((float)px - A.cx) / A.fx, xp1 = ((float)py - A.cy) / A.fy;
    const float dp = 1 / depthp;
    // operator/(MatExpr, double): alpha = 1./s, which hal::gemm32f(float alpha) rounds to float
    const double inv_dp = (double)dp;
    const double xp0d = (double)xp0, xp1d = (double)xp1;
    const float thr = A.chi_inter, thr_lo = thr * (1.0f - 1e-5f), thr_hi = thr * (1.0f + 1e-5f);
    int nkf = 0;
    double sJr = 0.0, sJJ = 0.0;
    // the projection and the depth taps of neighbour j+1 are issued before neighbour j is evaluated
    K4Proj cur, nxt;
    k4_project(A, A.nb[0], xp0, xp1, xp0d, xp1d, depthp, inv_dp, cur);
    for (int j = 0; j < A.n; ++j) {
        const float t2j = A.nb[j].t[2];
        if (j + 1 < A.n) k4_project(A, A.nb[j + 1], xp0, xp1, xp0d, xp1d, depthp, inv_dp, nxt);
        if (cur.inb) {
            int ncomp = 0;
#pragma unroll
            for (int t = 0; t < 4; ++t) {
                const float d = cur.d[t];
                if (d > SD_EPS_F) {
                    const float s = cur.s[t];
                    const float dl = cur.depthj - d;
                    const float nf = dl * dl, qf = s * s;
                    const float est = __fdividef(nf, qf);
                    bool comp;
                    if (fminf(nf, qf) > 1e-30f && est < thr_lo) comp = true;
                    else if (fminf(nf, qf) > 1e-30f && est > thr_hi) comp = false;
                    else {
                        const double dd = (double)dl;
                        comp = (float)((dd * dd) / ((double)s * (double)s)) < thr;
                    }
                    if (comp) {
                        ++ncomp;
                        // Gauss-Newton terms, PM.cc:777-783 (order: neighbour j, then tap n)
                        const float djn = 1 / d;
                        const float d2sigma = djn * djn * s;
                        const float Jn = -cur.rz / d2sigma;
                        const float rn = (djn - dp * cur.rz - t2j) / d2sigma;
                        sJr = __fma_rn((double)Jn, (double)rn, sJr);
                        sJJ = __fma_rn((double)Jn, (double)Jn, sJJ);
                    }
                }
            }
            if (ncomp >= 1) ++nkf;
        }
        cur = nxt;
    }
    float out;
    if (nkf < A.lambdaN) {
        out = 0.0f;  // sigma untouched, PM.cc:764
    } else {
        const float Jtr0 = (float)(sJr * -1.0);
        const float JtJ = (float)(sJJ * 1.0);
        const float dpDelta = Jtr0 / JtJ;
        out = 1 / (dp + dpDelta);  // PM.cc:793
    }
    A.depth_out[o] = out;
    if (POINTS) {
        float p0, p1, p2;
        backproject(A.P, px, py, out, p0, p1, p2);
        if (out > SD_EPS_F) {
            float* q = A.P.points + ((size_t)py * A.W + px) * 3;
            q[0] = p0; q[1] = p1; q[2] = p2;
        }
    }
}

// Compact form of a key frame's results (sd_download_compact): one record per entry of the valid-pixel list --
// SdCompactDepth {pixel = y << 16 | x, inverse depth, sigma} or SdCompactPoint {..., world point} -- in list order.
struct PackArgs {
    int W, pitch_f;
    const uint32_t* vlist;
    const int* vcount;
    const float* depth;   // depth_checked when the inter-key-frame check ran, else the intra-key-frame map
    const float* sigma;
    const float* points;  // H x 3W (with_points)
    int with_points;
    uint8_t* out;         // cap records, 8-byte aligned
    int* out_count;
    int cap;
};

__global__ void __launch_bounds__(256) k7_pack(const PackArgs A) {
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    const int n = *A.vcount;
    if (e == 0) *A.out_count = n;
    if (e >= n || e >= A.cap) return;
    const uint32_t ent = A.vlist[e];
    const int px = (int)(ent & 0xffffu), py = (int)(ent >> 16);
    const size_t o = (size_t)py * A.pitch_f + px;
    const float d = __ldg(A.depth + o), s = __ldg(A.sigma + o);
    if (A.with_points) {
        const float* q = A.points + ((size_t)py * A.W + px) * 3;
        uint2* r = reinterpret_cast<uint2*>(A.out + (size_t)e * sizeof(SdCompactPoint));
        r[0] = make_uint2(ent, __float_as_uint(d));
        r[1] = make_uint2(__float_as_uint(s), __float_as_uint(__ldg(q)));
        r[2] = make_uint2(__float_as_uint(__ldg(q + 1)), __float_as_uint(__ldg(q + 2)));
    } else {
        uint32_t* r = reinterpret_cast<uint32_t*>(A.out + (size_t)e * sizeof(SdCompactDepth));
        r[0] = ent; r[1] = __float_as_uint(d); r[2] = __float_as_uint(s);
    }
}

// ------------------------------------------------------------------------------------------------
// K5: UpdateSemiDensePointSet, PM.cc:337-367.  Twc*Pc: 4x4*4x1 small gemm, float left to right.
// ------------------------------------------------------------------------------------------------
// One block = 256 consecutive pixels of one row.  The 768 output floats are staged in shared memory so that the
// global stores are whole float4s (a thread's own 3 floats would be a 12-byte-stride store).
__global__ void __launch_bounds__(256) k5_pointset(const PointArgs A) {
    __shared__ __align__(16) float s_out[256 * 3];
    const int x0 = blockIdx.x * 256, y = blockIdx.y;
    const int x = x0 + threadIdx.x;
    float p0 = 0.f, p1 = 0.f, p2 = 0.f;
    if (x < A.W) backproject(A, x, y, __ldg(A.depth + (size_t)y * A.pitch_f + x), p0, p1, p2);
    s_out[3 * threadIdx.x] = p0; s_out[3 * threadIdx.x + 1] = p1; s_out[3 * threadIdx.x + 2] = p2;
    __syncthreads();
    const int npx = min(256, A.W - x0);          // pixels of this block inside the row
    float* row = A.points + ((size_t)y * A.W + x0) * 3;
    if ((A.W & 3) == 0 && (reinterpret_cast<uintptr_t>(A.points) & 15) == 0 && (npx & 3) == 0) {
        // row pitch 3W floats and x0*3 are multiples of 4 floats: aligned float4 stores
        if (threadIdx.x < (npx * 3) / 4)
            reinterpret_cast<float4*>(row)[threadIdx.x] = reinterpret_cast<const float4*>(s_out)[threadIdx.x];
    } else {
        for (int i = threadIdx.x; i < npx * 3; i += 256) row[i] = s_out[i];
    }
}

// ------------------------------------------------------------------------------------------------
// K6: confident-point feed.  The reference selects the points it trusts with
// `depth_sigma_ > 0.01 -> skip; depth_map_ > 0.000001 -> keep` (PM.cc:120-121, whole image, raster
// order) and the CARV transcript ingests extra 3-D points as a list per key frame
// (SFMTranscriptInterface_ORBSLAM.cpp:319-374).  K6 compacts those pixels on the device, in raster
// order (deterministic), optionally keeping only the `max_points` with the smallest sigma (exact
// radix select over the float bits, ties admitted in raster order), and back-projects them with
// the arithmetic of K5 -- so only count*12 bytes cross PCIe instead of three full planes.
//
//   k6_hist16 (x2)  histogram of the selection key's high / low 16 bits   } only when max_points > 0
//   k6_pick16 (x2)  locate the bin of the max_points-th smallest key      }
//   k6_count        per 256-pixel raster segment: keys below / equal to the threshold
//   k6_scan         exclusive scan over the segments -> output offsets, total
//   k6_write        compacted (x, y, z), sigma, pixel index
// ------------------------------------------------------------------------------------------------
struct K6State {
    int select;         // 1: only keys < thr, and the first eq_take keys == thr, are admitted
    uint32_t thr;       // selection key of the max_points-th smallest sigma
    int eq_take;
    uint32_t prefix;    // radix select: high 16 bits fixed by pass 0
    int k_rem;          // radix select: rank still to find inside the current bin
    int total;          // admitted points (written by k6_scan)
    int confident;      // pixels passing the reference predicate (before the top-k cut)
};

struct ConfArgs {
    PointArgs P;             // depth plane + intrinsics + Twc
    const float* sigma;
    double max_sigma;        // PM.cc:120 compares the float with a double literal
    int max_points;          // <= 0: no cut
    int nseg_row, nseg;      // 256-pixel segments per row / in the image
    K6State* st;
    uint32_t* hist;          // [2][65536]
    int* seg_lt;             // [nseg] keys below the threshold (all confident pixels when !select)
    int* seg_eq;             // [nseg] keys equal to the threshold
    int* seg_off;            // [nseg] output offset of the segment
    int* seg_eqoff;          // [nseg] equal keys before the segment
    float* out_xyz;          // [cap][3]
    float* out_sigma;        // [cap]
    uint32_t* out_pix;       // [cap] y << 16 | x
    int cap;
};

// total order on floats as unsigned keys (negative < positive, NaNs at the ends)
__device__ __forceinline__ uint32_t sigma_key(uint32_t bits) { return bits ^ ((uint32_t)((int32_t)bits >> 31) | 0x80000000u); }
// The bits of a float as an integer the compiler cannot trace back to the float: nvcc otherwise turns the sign-bit
// arithmetic of sigma_key into FADD -|x|, -0, which canonicalises NaN payloads and signs (seen in SASS; NaN sigmas
// pass the reference's `sigma > 0.01` test, so they are part of the selection and need a defined place in the order)
__device__ __forceinline__ uint32_t raw_bits(float v) {
    uint32_t b;
    asm volatile("mov.b32 %0, %1;" : "=r"(b) : "f"(v));
    return b;
}

__device__ __forceinline__ bool k6_confident(const ConfArgs& A, int x, int y, float& d, float& s) {
    const size_t o = (size_t)y * A.P.pitch_f + x;
    s = __ldg(A.sigma + o);
    d = __ldg(A.P.depth + o);
    return !((double)s > A.max_sigma) && d > SD_EPS_F;  // PM.cc:120-121
}

template <int PASS>
__global__ void __launch_bounds__(256) k6_hist16(const ConfArgs A) {
    const int x = blockIdx.x * 32 + (threadIdx.x & 31), y = blockIdx.y * 8 + (threadIdx.x >> 5);
    if (x >= A.P.W || y >= A.P.H) return;
    float d, s;
    if (!k6_confident(A, x, y, d, s)) return;
    const uint32_t key = sigma_key(raw_bits(s));
    if (PASS == 0) atomicAdd(A.hist + (key >> 16), 1u);
    else if (A.st->select && (key >> 16) == A.st->prefix) atomicAdd(A.hist + 65536 + (key & 0xffffu), 1u);
}

// one block of 1024 threads: find the bin holding rank k_rem (1-based) of the 65536-bin histogram
template <int PASS>
__global__ void __launch_bounds__(1024) k6_pick16(const ConfArgs A) {
    __shared__ unsigned int s_part[1024];
    __shared__ unsigned int s_tot;
    const int t = threadIdx.x;
    const uint32_t* h = A.hist + PASS * 65536;
    unsigned int loc = 0;
    for (int i = 0; i < 64; ++i) loc += h[t * 64 + i];
    s_part[t] = loc;
    __syncthreads();
    if (t == 0) {
        unsigned int acc = 0;
        for (int i = 0; i < 1024; ++i) { const unsigned int v = s_part[i]; s_part[i] = acc; acc += v; }
        s_tot = acc;
    }
    __syncthreads();
    K6State* st = A.st;
    if (PASS == 0 && t == 0) {
        st->confident = (int)s_tot;
        st->select = (A.max_points > 0 && s_tot > (unsigned)A.max_points) ? 1 : 0;
        st->k_rem = A.max_points;
        st->eq_take = 0;
        st->thr = 0;
    }
    __syncthreads();
    if (!st->select) return;
    const unsigned int k = (unsigned)st->k_rem;  // read before anyone writes it (only the owner thread below writes)
    __syncthreads();
    const unsigned int before = s_part[t];
    if (k > before && k <= before + loc) {  // exactly one thread
        unsigned int acc = before;
        for (int i = 0; i < 64; ++i) {
            const unsigned int v = h[t * 64 + i];
            if (k <= acc + v) {
                const uint32_t bin = (uint32_t)(t * 64 + i);
                if (PASS == 0) { st->prefix = bin; st->k_rem = (int)(k - acc); }
                else { st->thr = (st->prefix << 16) | bin; st->eq_take = (int)(k - acc); }
                break;
            }
            acc += v;
        }
    }
}

// one warp per 256-pixel raster segment
__global__ void __launch_bounds__(256) k6_count(const ConfArgs A) {
    const int seg = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (seg >= A.nseg) return;
    const int y = seg / A.nseg_row, xs = (seg - y * A.nseg_row) * 256;
    const int select = A.st->select;
    const uint32_t thr = A.st->thr;
    int nlt = 0, neq = 0;
    for (int i = 0; i < 8; ++i) {
        const int x = xs + i * 32 + lane;
        float d, s;
        bool lt = false, eq = false;
        if (x < A.P.W && k6_confident(A, x, y, d, s)) {
            const uint32_t key = sigma_key(raw_bits(s));
            lt = !select || key < thr;
            eq = select && key == thr;
        }
        nlt += __popc(__ballot_sync(0xffffffffu, lt));
        neq += __popc(__ballot_sync(0xffffffffu, eq));
    }
    if (lane == 0) { A.seg_lt[seg] = nlt; A.seg_eq[seg] = neq; }
}

// single block: admitted points per segment = lt + clamp(eq_take - equal keys before, 0, eq); exclusive scans
__global__ void __launch_bounds__(1024) k6_scan(const ConfArgs A) {
    __shared__ int s_a[1024], s_e[1024];
    __shared__ int s_carry_a, s_carry_e;
    const int t = threadIdx.x;
    const int eq_take = A.st->select ? A.st->eq_take : 0;
    if (t == 0) { s_carry_a = 0; s_carry_e = 0; }
    __syncthreads();
    for (int base = 0; base < A.nseg; base += 1024) {
        const int i = base + t;
        const int e = i < A.nseg ? A.seg_eq[i] : 0;
        // inclusive scan of the equal counts first: the admitted count of a segment depends on it
        s_e[t] = e;
        __syncthreads();
        for (int d = 1; d < 1024; d <<= 1) {
            const int v = t >= d ? s_e[t - d] : 0;
            __syncthreads();
            s_e[t] += v;
            __syncthreads();
        }
        const int eq_before = s_carry_e + s_e[t] - e;
        int adm = 0;
        if (i < A.nseg) {
            int take = eq_take - eq_before;
            take = take < 0 ? 0 : (take > e ? e : take);
            adm = A.seg_lt[i] + take;
            A.seg_eqoff[i] = eq_before;
        }
        s_a[t] = adm;
        __syncthreads();
        for (int d = 1; d < 1024; d <<= 1) {
            const int v = t >= d ? s_a[t - d] : 0;
            __syncthreads();
            s_a[t] += v;
            __syncthreads();
        }
        if (i < A.nseg) A.seg_off[i] = s_carry_a + s_a[t] - adm;
        __syncthreads();
        if (t == 1023) { s_carry_a += s_a[1023]; s_carry_e += s_e[1023]; }
        __syncthreads();
    }
    if (t == 0) A.st->total = s_carry_a;
}

__global__ void __launch_bounds__(256) k6_write(const ConfArgs A) {
    const int seg = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (seg >= A.nseg) return;
    const int y = seg / A.nseg_row, xs = (seg - y * A.nseg_row) * 256;
    const int select = A.st->select, eq_take = A.st->eq_take;
    const uint32_t thr = A.st->thr;
    const unsigned lt_mask = (1u << lane) - 1u;
    int pos = A.seg_off[seg], eqn = A.seg_eqoff[seg];
    for (int i = 0; i < 8; ++i) {
        const int x = xs + i * 32 + lane;
        float d = 0.f, s = 0.f;
        bool lt = false, eq = false;
        if (x < A.P.W && k6_confident(A, x, y, d, s)) {
            const uint32_t key = sigma_key(raw_bits(s));
            lt = !select || key < thr;
            eq = select && key == thr;
        }
        const unsigned beq = __ballot_sync(0xffffffffu, eq);
        const bool adm = lt || (eq && eqn + __popc(beq & lt_mask) < eq_take);
        const unsigned badm = __ballot_sync(0xffffffffu, adm);
        if (adm) {
            const int o = pos + __popc(badm & lt_mask);
            if (o < A.cap) {
                float p0, p1, p2;
                backproject(A.P, x, y, d, p0, p1, p2);
                A.out_xyz[3 * (size_t)o] = p0; A.out_xyz[3 * (size_t)o + 1] = p1; A.out_xyz[3 * (size_t)o + 2] = p2;
                A.out_sigma[o] = s;
                A.out_pix[o] = ((uint32_t)y << 16) | (uint32_t)x;
            }
        }
        pos += __popc(badm);
        eqn += __popc(beq);
    }
}

// pack / unpack between pitched planes and tight W*H buffers (NCCL halo exchange)
__global__ void k_copy_plane(const float* __restrict__ src, int src_pitch, float* __restrict__ dst, int dst_pitch, int W, int H) {
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
    const int y = blockIdx.y * blockDim.y + threadIdx.y;
    if (x < W && y < H) dst[(size_t)y * dst_pitch + x] = src[(size_t)y * src_pitch + x];
}

}  // namespace sdpm
