"""Generates tests/golden/cv_golden.npz: outputs of the real OpenCV (cv2 wheel of the build
container, version recorded in the file) for exactly the operation shapes that the reference's
hot path evaluates through cv::Mat (see oracle/cvlike.h).  Run in the build container only
(needs cv2); the tests replay the vectors without cv2.

    python tests/golden/make_cv_golden.py
"""
import os

import cv2
import numpy as np

rng = np.random.default_rng(20161109)
f32 = np.float32
N = 400
out = {"cv2_version": np.array(cv2.__version__)}

# fastAtan2: Sobel/8 lattice values, generic floats, the (-a/b, 1) shape of PM.cc:414, edge cases
y = np.concatenate([rng.integers(-1020, 1021, 1500) / 8.0, rng.standard_normal(1500), rng.uniform(-4, 4, 1500),
                    [0, 0, 1, -1, 0, 1e-9, -1e-9, 1e-30, -1e-30]]).astype(f32)
x = np.concatenate([rng.integers(-1020, 1021, 1500) / 8.0, rng.standard_normal(1500), np.ones(1500),
                    [1, -1, 0, 0, 0, 1, 1, 1, 1]]).astype(f32)
out["atan_y"], out["atan_x"] = y, x
out["atan_out"] = np.array([cv2.fastAtan2(float(a), float(b)) for a, b in zip(y, x)], f32)


def rot(n):
    q = rng.standard_normal((n, 4))
    q /= np.linalg.norm(q, axis=1, keepdims=True)
    w, a, b, c = q.T
    return np.stack([1 - 2 * (b * b + c * c), 2 * (a * b - c * w), 2 * (a * c + b * w),
                     2 * (a * b + c * w), 1 - 2 * (a * a + c * c), 2 * (b * c - a * w),
                     2 * (a * c - b * w), 2 * (b * c + a * w), 1 - 2 * (a * a + b * b)], 1).reshape(n, 3, 3).astype(f32)


# gemm small path 3x3*3x1 with alpha and C (GetSearchRange, t21, InterKF temp)
A = np.concatenate([rot(N // 2), rng.standard_normal((N // 2, 3, 3)).astype(f32)])
b = rng.standard_normal((N, 3, 1)).astype(f32)
c = rng.standard_normal((N, 3, 1)).astype(f32)
alpha = rng.uniform(0.3, 3.0, N).astype(f32).astype(np.float64)
alpha[::7] = 1.0 / rng.uniform(0.5, 2.0, len(alpha[::7])).astype(f32).astype(np.float64)  # the 1./depthp shape
out["g31_A"], out["g31_b"], out["g31_c"], out["g31_alpha"] = A, b, c, alpha
out["g31_out"] = np.stack([cv2.gemm(A[i], b[i], alpha[i], c[i], 1.0) for i in range(N)])
out["g31_out_noc"] = np.stack([cv2.gemm(A[i], b[i], -1.0, None, 0.0) for i in range(N)])
# 3x3*3x3 small path
B = rng.standard_normal((N, 3, 3)).astype(f32)
out["g33_B"] = B
out["g33_out"] = np.stack([cv2.gemm(A[i], B[i], 1.0, None, 0.0) for i in range(N)])
# A*B^T general path (double accumulation), alpha = +-1
out["gabt_out"] = np.stack([cv2.gemm(A[i], B[i], 1.0, None, 0.0, flags=cv2.GEMM_2_T) for i in range(N)])
out["gabt_out_neg"] = np.stack([cv2.gemm(A[i], B[i], -1.0, None, 0.0, flags=cv2.GEMM_2_T) for i in range(N)])
# 1xn * nx1 general path with alpha (GetPixelDepth n=3; Gauss-Newton n up to 160)
dots_a, dots_b, dots_alpha, dots_out = [], [], [], []
for i in range(N):
    n = 3 if i < N // 2 else int(rng.integers(1, 161))
    a1 = rng.standard_normal((1, n)).astype(f32)
    b1 = rng.standard_normal((n, 1)).astype(f32)
    al = float(f32(rng.uniform(-600, 600))) if i % 3 else -1.0
    dots_a.append(np.pad(a1[0], (0, 160 - n)))
    dots_b.append(np.pad(b1[:, 0], (0, 160 - n)))
    dots_alpha.append(al)
    dots_out.append(cv2.gemm(a1, b1, al, None, 0.0)[0, 0])
out["dot_a"], out["dot_b"] = np.array(dots_a, f32), np.array(dots_b, f32)
out["dot_n"] = np.array([3 if i < N // 2 else 0 for i in range(N)], np.int32)
out["dot_n"][N // 2:] = [int(np.max(np.nonzero(np.abs(dots_a[i]) + np.abs(dots_b[i]))[0]) + 1) for i in range(N // 2, N)]
out["dot_alpha"], out["dot_out"] = np.array(dots_alpha), np.array(dots_out, f32)
# 4x4*4x1 small path (Twc*Pc)
A4 = rng.standard_normal((N, 4, 4)).astype(f32)
b4 = rng.standard_normal((N, 4, 1)).astype(f32)
out["g41_A"], out["g41_b"] = A4, b4
out["g41_out"] = np.stack([cv2.gemm(A4[i], b4[i], 1.0, None, 0.0) for i in range(N)])
# invert 3x3 (K2.inv()) and solve LU (K1.t().inv()*t12x)
Ks = []
for i in range(N):
    if i % 2:
        Ks.append(rng.standard_normal((3, 3)))
    else:
        fx, fy = rng.uniform(300, 1800, 2)
        cx, cy = rng.uniform(200, 1000, 2)
        Ks.append(np.array([[fx, 0, cx], [0, fy, cy], [0, 0, 1]]))
Ks = np.array(Ks, f32)
out["inv_A"] = Ks
out["inv_out"] = np.stack([cv2.invert(Ks[i])[1] for i in range(N)])
At = np.stack([Ks[i].T.copy() for i in range(N)])
out["solve_A"], out["solve_B"] = At, B
out["solve_out"] = np.stack([cv2.solve(At[i], B[i], flags=cv2.DECOMP_LU)[1] for i in range(N)])
# front end: Sobel/8 replicate border, meanStdDev
img = rng.integers(0, 256, (37, 53)).astype(np.uint8)
out["fe_img"] = img
out["fe_gx"] = cv2.Sobel(img, cv2.CV_32F, 1, 0, ksize=3, scale=0.125, borderType=cv2.BORDER_REPLICATE)
out["fe_gy"] = cv2.Sobel(img, cv2.CV_32F, 0, 1, ksize=3, scale=0.125, borderType=cv2.BORDER_REPLICATE)
out["fe_std"] = np.array(cv2.meanStdDev(img)[1][0, 0])

path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "cv_golden.npz")
np.savez_compressed(path, **out)
print("wrote", path, os.path.getsize(path), "bytes")
