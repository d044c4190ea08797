// K4: batched weighted homography fit in ONE launch per call.
//
// Replaces find_homography_nonhomogeneous_QR / _nonhomogeneous / _dlt / _IRLSq_QR of the reference
// (pytracking/utils/least_squares_H.py:56-346), which build the 2N x 8 (or 2N x 9) design matrix in
// global memory and hand it to batched QR / pinv / SVD, about 25 launches and, for IRLS, six
// factorisations with a dense-solver round trip each.
//
// One CTA per problem:
//   pass 1  means of both point sets                       (Hartley normalisation, kornia normalize_points)
//   pass 2  mean distance to the mean -> scales
//   pass 3+ weighted Gram of the design matrix, accumulated in fp64.  With p = (x1, y1, 1) the two rows of a
//           correspondence are a_x = [0, -p, y2 p], a_y = [p, 0, -x2 p] (9 columns; the 8-parameter system is
//           the first 8 with rhs = -column 9), so the 9x9 Gram is five symmetric 3x3 blocks of p p^T:
//             [ S(wy)   0       -S(wy x2)            ]
//             [ 0       S(wx)   -S(wx y2)            ]        30 sums instead of 45
//             [ .       .        S(wx y2^2 + wy x2^2)]
//           row weights: w^2 (QR / pinv / IRLS, whose rows are scaled by w, least_squares_H.py:186-193),
//           w (DLT, :253-256), times fn(residual) per row for IRLS (:334-339);
//   solve   thread 0: 8x8 Cholesky (fp64) or cyclic Jacobi on the 9x9 (DLT: eigenvector of the smallest
//           eigenvalue = last right-singular vector of the Gram, :258-260);
//   finish  H = T2^-1 Hn T1, divided by (H33 + 1e-8) (:204-209).
// Points are re-read from L2 on every pass (a problem's points are 20 B each; 14 400 of them are 288 KB).
#include "common.cuh"

namespace woft {

constexpr int kHThreads = 256;
constexpr int kHWarps = kHThreads / 32;
constexpr int kHUnroll = 4;  // independent loads in flight per thread and pass

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// Block-wide sum of K doubles per thread; result valid in every thread (read back from shared memory).
template <int K>
__device__ __forceinline__ void block_sum(double (&v)[K], double* scratch /* kHWarps*K + K */) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
#pragma unroll
  for (int k = 0; k < K; ++k) v[k] = warp_sum(v[k]);
  __syncthreads();  // scratch may still be read from a previous call
  if (lane == 0) {
#pragma unroll
    for (int k = 0; k < K; ++k) scratch[warp * K + k] = v[k];
  }
  __syncthreads();
  if (threadIdx.x < K) {
    double s = 0.0;
    for (int w = 0; w < kHWarps; ++w) s += scratch[w * K + threadIdx.x];
    scratch[kHWarps * K + threadIdx.x] = s;
  }
  __syncthreads();
#pragma unroll
  for (int k = 0; k < K; ++k) v[k] = scratch[kHWarps * K + k];
}

// Solve the SPD system G x = g (n = 8) by Cholesky; returns false if a pivot is not positive.
__device__ __noinline__ bool cholesky_solve8(double (*G)[9], const double* g, double* x) {
  constexpr int n = 8;
  double L[n][n];
  for (int j = 0; j < n; ++j) {
    double d = G[j][j];
    for (int k = 0; k < j; ++k) d -= L[j][k] * L[j][k];
    if (!(d > 0.0)) return false;
    d = sqrt(d);
    L[j][j] = d;
    for (int i = j + 1; i < n; ++i) {
      double s = G[i][j];
      for (int k = 0; k < j; ++k) s -= L[i][k] * L[j][k];
      L[i][j] = s / d;
    }
  }
  double y[n];
  for (int i = 0; i < n; ++i) {
    double s = g[i];
    for (int k = 0; k < i; ++k) s -= L[i][k] * y[k];
    y[i] = s / L[i][i];
  }
  for (int i = n - 1; i >= 0; --i) {
    double s = y[i];
    for (int k = i + 1; k < n; ++k) s -= L[k][i] * x[k];
    x[i] = s / L[i][i];
  }
  return true;
}

// Jacobi eigen-decomposition of the symmetric 9x9 M (in shared memory, destroyed) by ONE WARP; writes the unit
// eigenvector of the smallest eigenvalue to vec[0..8].  Parallel ordering: a round-robin tournament on 10 players (the
// tenth is a bye) gives 9 rounds of 4 disjoint rotations that together visit all 36 pairs; disjoint rotations commute,
// so a round is: 4 lanes compute (c, s), then 36 (pair, k) items update the columns of M and V, then 36 the rows of M.
// On one thread this eigen-solve took longer than the whole data pass of a 14 400-point problem.
__device__ __forceinline__ void jacobi_smallest9_warp(double (*M)[9], double (*V)[9], double* rot /* [4][2] */,
                                                      int* pq /* [4][2] */, double* vec, int lane) {
  constexpr int n = 9;
  for (int i = lane; i < n * n; i += 32) V[i / n][i % n] = (i / n == i % n) ? 1.0 : 0.0;
  __syncwarp();
  for (int sweep = 0; sweep < 60; ++sweep) {
    double off = 0.0, diag = 0.0;
    for (int i = lane; i < n * n; i += 32) {
      const int r = i / n, c = i - r * n;
      const double v = M[r][c];
      if (r == c) diag += v * v;
      else if (c > r) off += v * v;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      off += __shfl_xor_sync(0xffffffffu, off, o);
      diag += __shfl_xor_sync(0xffffffffu, diag, o);
    }
    if (off <= 1e-40 * diag || off == 0.0) break;   // uniform across the warp
    for (int round = 0; round < n; ++round) {
      if (lane < 4) {
        const int a = (round + lane + 1) % n, b2 = (round - lane - 1 + n) % n;
        const int p = a < b2 ? a : b2, q = a < b2 ? b2 : a;
        const double apq = M[p][q];
        double c = 1.0, sn = 0.0;
        if (apq != 0.0) {
          const double theta = (M[q][q] - M[p][p]) / (2.0 * apq);
          const double t = (theta >= 0.0 ? 1.0 : -1.0) / (fabs(theta) + sqrt(theta * theta + 1.0));
          c = 1.0 / sqrt(t * t + 1.0);
          sn = t * c;
        }
        rot[2 * lane] = c;
        rot[2 * lane + 1] = sn;
        pq[2 * lane] = p;
        pq[2 * lane + 1] = q;
      }
      __syncwarp();
      for (int item = lane; item < 4 * n; item += 32) {   // M <- M J, V <- V J
        const int i = item / n, k = item - i * n;
        const int p = pq[2 * i], q = pq[2 * i + 1];
        const double c = rot[2 * i], sn = rot[2 * i + 1];
        const double mkp = M[k][p], mkq = M[k][q];
        M[k][p] = c * mkp - sn * mkq;
        M[k][q] = sn * mkp + c * mkq;
        const double vkp = V[k][p], vkq = V[k][q];
        V[k][p] = c * vkp - sn * vkq;
        V[k][q] = sn * vkp + c * vkq;
      }
      __syncwarp();
      for (int item = lane; item < 4 * n; item += 32) {   // M <- J^T M
        const int i = item / n, k = item - i * n;
        const int p = pq[2 * i], q = pq[2 * i + 1];
        const double c = rot[2 * i], sn = rot[2 * i + 1];
        const double mpk = M[p][k], mqk = M[q][k];
        M[p][k] = c * mpk - sn * mqk;
        M[q][k] = sn * mpk + c * mqk;
      }
      __syncwarp();
    }
  }
  if (lane == 0) {
    int best = 0;
    for (int i = 1; i < n; ++i)
      if (M[i][i] < M[best][best]) best = i;
    for (int k = 0; k < n; ++k) vec[k] = V[k][best];
  }
  __syncwarp();
}

struct HParams {
  const float* p1;  // (B, N, 2)
  const float* p2;  // (B, N, 2)
  const float* w;   // (B, N) or null
  float* H;         // (B, 3, 3)
  int n;
  const int* n_dev;  // when not null, the number of correspondences is read here (batch 1; arrays hold `n` slots)
  int method;
  int n_iter;
  float huber_k;
};

__global__ void __launch_bounds__(kHThreads, 2) homography_fit_kernel(const HParams prm) {
  __shared__ double scratch[kHWarps * 30 + 30];
  __shared__ double h9[9];  // current solution in normalised coordinates (h9[8] = 1 for the 8-parameter system)
  __shared__ double jM[9][9], jV[9][9], jrot[8];   // DLT only: the Gram and its eigenvectors for the warp-wide Jacobi
  __shared__ int jpq[8];

  const int b = blockIdx.x;
  const int N = prm.n_dev ? min(*prm.n_dev, prm.n) : prm.n;
  if (N < 4) {  // only reachable with a device-side count: the host entry point rejects n < 4 (least_squares_H.py:162)
    if (threadIdx.x < 9) prm.H[(size_t)b * 9 + threadIdx.x] = __int_as_float(0x7fc00000);
    return;
  }
  const float2* __restrict__ P1 = reinterpret_cast<const float2*>(prm.p1) + (size_t)b * N;
  const float2* __restrict__ P2 = reinterpret_cast<const float2*>(prm.p2) + (size_t)b * N;
  const float* __restrict__ Wt = prm.w ? prm.w + (size_t)b * N : nullptr;

  // ---- pass 1: means.  A thread's share (N / 256 points) is summed in fp32 and only the per-thread sums go to fp64:
  // the transform built from the means and scales is applied and undone consistently, so the solution does not depend
  // on their last bits, and fp32 -> fp64 conversions run on the 16-lane XU pipe (they were a third of this kernel).
  float mf[4] = {0.f, 0.f, 0.f, 0.f};
  for (int i0 = threadIdx.x; i0 < N; i0 += kHThreads * kHUnroll) {
    float2 a[kHUnroll], c[kHUnroll];
#pragma unroll
    for (int u = 0; u < kHUnroll; ++u) {  // all loads of the batch first: the loop is latency bound otherwise
      const int i = i0 + u * kHThreads;
      a[u] = i < N ? __ldg(P1 + i) : make_float2(0.f, 0.f);
      c[u] = i < N ? __ldg(P2 + i) : make_float2(0.f, 0.f);
    }
#pragma unroll
    for (int u = 0; u < kHUnroll; ++u) {
      mf[0] += a[u].x; mf[1] += a[u].y; mf[2] += c[u].x; mf[3] += c[u].y;
    }
  }
  double m[4] = {double(mf[0]), double(mf[1]), double(mf[2]), double(mf[3])};
  block_sum<4>(m, scratch);
  const double inv_n = 1.0 / double(N);
  const double mx1 = m[0] * inv_n, my1 = m[1] * inv_n, mx2 = m[2] * inv_n, my2 = m[3] * inv_n;

  // ---- pass 2: mean distance to the mean -> isotropic scale sqrt(2) / (dist + eps), fp32 per thread as above
  // (the reference's normalize_points works in fp32 throughout)
  float df[2] = {0.f, 0.f};
  {
    const float fx1 = float(mx1), fy1 = float(my1), fx2 = float(mx2), fy2 = float(my2);
    for (int i0 = threadIdx.x; i0 < N; i0 += kHThreads * kHUnroll) {
      float2 a[kHUnroll], c[kHUnroll];
      bool live[kHUnroll];
#pragma unroll
      for (int u = 0; u < kHUnroll; ++u) {
        const int i = i0 + u * kHThreads;
        live[u] = i < N;
        a[u] = live[u] ? __ldg(P1 + i) : make_float2(0.f, 0.f);
        c[u] = live[u] ? __ldg(P2 + i) : make_float2(0.f, 0.f);
      }
#pragma unroll
      for (int u = 0; u < kHUnroll; ++u) {
        if (!live[u]) continue;
        const float ax = a[u].x - fx1, ay = a[u].y - fy1, cx = c[u].x - fx2, cy = c[u].y - fy2;
        df[0] += sqrtf(ax * ax + ay * ay);
        df[1] += sqrtf(cx * cx + cy * cy);
      }
    }
  }
  double d[2] = {double(df[0]), double(df[1])};
  block_sum<2>(d, scratch);
  const double s1 = 1.4142135623730951 / (d[0] * inv_n + 1e-8);
  const double s2 = 1.4142135623730951 / (d[1] * inv_n + 1e-8);

  const bool irls = prm.method == WOFT_H_IRLS_L1 || prm.method == WOFT_H_IRLS_HUBER;
  const bool dlt = prm.method == WOFT_H_DLT;
  const int solves = irls ? prm.n_iter + 1 : 1;
  const double huber_k = prm.huber_k;

  for (int it = 0; it < solves; ++it) {
    // ---- Gram pass: five symmetric 3x3 blocks of p p^T, entries ordered xx, xy, x, yy, y, 1
    double S[30];
#pragma unroll
    for (int k = 0; k < 30; ++k) S[k] = 0.0;
    double h[9];
    if (it > 0) {
#pragma unroll
      for (int k = 0; k < 9; ++k) h[k] = h9[k];
    }
    for (int i0 = threadIdx.x; i0 < N; i0 += kHThreads * kHUnroll) {
      float2 pa[kHUnroll], pc[kHUnroll];
      float pw[kHUnroll];
#pragma unroll
      for (int u = 0; u < kHUnroll; ++u) {
        const int i = i0 + u * kHThreads;
        pa[u] = i < N ? __ldg(P1 + i) : make_float2(0.f, 0.f);
        pc[u] = i < N ? __ldg(P2 + i) : make_float2(0.f, 0.f);
        pw[u] = i < N ? (Wt ? __ldg(Wt + i) : 1.0f) : 0.f;
      }
#pragma unroll
      for (int u = 0; u < kHUnroll; ++u) {
      if (i0 + u * kHThreads >= N) continue;
      const float2 a = pa[u], c = pc[u];
      const double x1 = (a.x - mx1) * s1, y1 = (a.y - my1) * s1;
      const double x2 = (c.x - mx2) * s2, y2 = (c.y - my2) * s2;
      const double w = double(pw[u]);
      double wx = dlt ? w : w * w, wy = wx;  // row weights of the a_x and a_y rows
      if (it > 0) {
        // residual of each w-scaled row at the previous solution (least_squares_H.py:336), normalised units
        const double den = h[6] * x1 + h[7] * y1 + h[8];
        const double rx = w * (-(h[3] * x1 + h[4] * y1 + h[5]) + y2 * den);
        const double ry = w * ((h[0] * x1 + h[1] * y1 + h[2]) - x2 * den);
        const double arx = fabs(rx), ary = fabs(ry);
        double fx = 1.0 / (arx + 1e-8), fy = 1.0 / (ary + 1e-8);
        if (prm.method == WOFT_H_IRLS_HUBER) {
          if (arx < huber_k) fx = 1.0;
          if (ary < huber_k) fy = 1.0;
        }
        wx *= fx;  // (sqrt(fn(r)) * w)^2
        wy *= fy;
      }
      const double pp[6] = {x1 * x1, x1 * y1, x1, y1 * y1, y1, 1.0};
      const double c0 = wy, c1 = wx, c2 = wy * x2, c3 = wx * y2, c4 = wx * y2 * y2 + wy * x2 * x2;
      if (it == 0) {  // first solve: both rows carry the same weight, so S(wx) == S(wy) and is filled in afterwards
#pragma unroll
        for (int k = 0; k < 6; ++k) {
          S[k] += c0 * pp[k];
          S[12 + k] += c2 * pp[k];
          S[18 + k] += c3 * pp[k];
          S[24 + k] += c4 * pp[k];
        }
      } else {
#pragma unroll
        for (int k = 0; k < 6; ++k) {
          S[k] += c0 * pp[k];
          S[6 + k] += c1 * pp[k];
          S[12 + k] += c2 * pp[k];
          S[18 + k] += c3 * pp[k];
          S[24 + k] += c4 * pp[k];
        }
      }
      }
    }
    if (it == 0) {
#pragma unroll
      for (int k = 0; k < 6; ++k) S[6 + k] = S[k];
    }
    block_sum<30>(S, scratch);

    if (threadIdx.x == 0) {
      // assemble the 9x9 Gram from the blocks
      double M[9][9];
      const int idx[3][3] = {{0, 1, 2}, {1, 3, 4}, {2, 4, 5}};
      for (int r = 0; r < 3; ++r)
        for (int c = 0; c < 3; ++c) {
          const int k = idx[r][c];
          M[r][c] = S[k];            M[r][3 + c] = 0.0;          M[r][6 + c] = -S[12 + k];
          M[3 + r][c] = 0.0;         M[3 + r][3 + c] = S[6 + k]; M[3 + r][6 + c] = -S[18 + k];
          M[6 + r][c] = -S[12 + k];  M[6 + r][3 + c] = -S[18 + k]; M[6 + r][6 + c] = S[24 + k];
        }
      if (dlt) {
        for (int r = 0; r < 9; ++r)
          for (int c = 0; c < 9; ++c) jM[r][c] = M[r][c];
      } else {
        double sol[9], g[8];
        for (int k = 0; k < 8; ++k) g[k] = -M[k][8];  // A^T b with b = -(9th column)
        if (!cholesky_solve8(M, g, sol)) {
          const double nan = __longlong_as_double(0x7ff8000000000000ll);
          for (int k = 0; k < 8; ++k) sol[k] = nan;  // rank-deficient input: the reference's QR returns inf/nan too
        }
        sol[8] = 1.0;
        for (int k = 0; k < 9; ++k) h9[k] = sol[k];
      }
    }
    if (dlt) {   // block-uniform
      __syncthreads();
      if (threadIdx.x < 32) jacobi_smallest9_warp(jM, jV, jrot, jpq, h9, threadIdx.x);
    }
    __syncthreads();
  }

  if (threadIdx.x == 0) {
    // H = T2^-1 * Hn * T1 with T = [[s,0,-s mx],[0,s,-s my],[0,0,1]], then H / (H33 + 1e-8)
    double Hn[3][3], A[3][3], R[3][3];
    for (int k = 0; k < 9; ++k) Hn[k / 3][k % 3] = h9[k];
    const double T1[3][3] = {{s1, 0.0, -s1 * mx1}, {0.0, s1, -s1 * my1}, {0.0, 0.0, 1.0}};
    const double T2i[3][3] = {{1.0 / s2, 0.0, mx2}, {0.0, 1.0 / s2, my2}, {0.0, 0.0, 1.0}};
    for (int r = 0; r < 3; ++r)
      for (int c = 0; c < 3; ++c) A[r][c] = Hn[r][0] * T1[0][c] + Hn[r][1] * T1[1][c] + Hn[r][2] * T1[2][c];
    for (int r = 0; r < 3; ++r)
      for (int c = 0; c < 3; ++c) R[r][c] = T2i[r][0] * A[0][c] + T2i[r][1] * A[1][c] + T2i[r][2] * A[2][c];
    const double inv = 1.0 / (R[2][2] + 1e-8);
    float* out = prm.H + (size_t)b * 9;
    for (int k = 0; k < 9; ++k) out[k] = float(R[k / 3][k % 3] * inv);
  }
}

// torch_proj_errors (least_squares_H.py:474-489): pts (B, 2, N) -> err (B, N)
__global__ void proj_errors_kernel(const float* __restrict__ H, const float* __restrict__ pa,
                                   const float* __restrict__ pb, int n, float* __restrict__ err) {
  const int b = blockIdx.y;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float* h = H + (size_t)b * 9;
  const float x = pa[(size_t)b * 2 * n + i], y = pa[(size_t)b * 2 * n + n + i];
  const float u = h[0] * x + h[1] * y + h[2];
  const float v = h[3] * x + h[4] * y + h[5];
  const float z = h[6] * x + h[7] * y + h[8];
  // kornia convert_points_from_homogeneous: divide by (z + eps) where |z| > eps, else leave unscaled
  const float sc = fabsf(z) > 1e-8f ? 1.0f / (z + 1e-8f) : 1.0f;
  const float dx = u * sc - pb[(size_t)b * 2 * n + i];
  const float dy = v * sc - pb[(size_t)b * 2 * n + n + i];
  err[(size_t)b * n + i] = sqrtf(dx * dx + dy * dy);
}

}  // namespace woft

extern "C" int woft_homography_fit(const float* points1, const float* points2, const float* weights, int batch,
                                   int n_points, int method, int n_iter, float huber_k, float* H_out,
                                   void* stream) {
  using namespace woft;
  WOFT_REQUIRE(points1 && points2 && H_out, "woft_homography_fit: null pointer");
  WOFT_REQUIRE(batch > 0, "woft_homography_fit: batch %d", batch);
  WOFT_REQUIRE(n_points >= 4, "woft_homography_fit: %d correspondences, at least 4 are needed", n_points);
  WOFT_REQUIRE(method >= WOFT_H_NONHOMOGENEOUS && method <= WOFT_H_IRLS_HUBER, "woft_homography_fit: method %d",
               method);
  WOFT_REQUIRE(n_iter >= 0 && n_iter <= 1000, "woft_homography_fit: n_iter %d", n_iter);
  WOFT_REQUIRE((reinterpret_cast<uintptr_t>(points1) & 7) == 0 && (reinterpret_cast<uintptr_t>(points2) & 7) == 0,
               "woft_homography_fit: point arrays must be 8-byte aligned");
  HParams p{points1, points2, weights, H_out, n_points, nullptr, method, n_iter, huber_k};
  homography_fit_kernel<<<batch, kHThreads, 0, static_cast<cudaStream_t>(stream)>>>(p);
  return launch_status("homography_fit_kernel");
}

extern "C" int woft_proj_errors(const float* H, const float* pts_a, const float* pts_b, int batch, int n_points,
                                float* err, void* stream) {
  using namespace woft;
  WOFT_REQUIRE(H && pts_a && pts_b && err, "woft_proj_errors: null pointer");
  WOFT_REQUIRE(batch > 0 && n_points > 0 && batch < 65536, "woft_proj_errors: bad extents");
  dim3 grid(ceil_div(n_points, 256), batch);
  proj_errors_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(H, pts_a, pts_b, n_points, err);
  return launch_status("proj_errors_kernel");
}

extern "C" int woft_homography_fit_n(const float* points1, const float* points2, const float* weights,
                                     const int* n_points_dev, int max_points, int method, int n_iter, float huber_k,
                                     float* H_out, void* stream) {
  using namespace woft;
  WOFT_REQUIRE(points1 && points2 && H_out && n_points_dev, "woft_homography_fit_n: null pointer");
  WOFT_REQUIRE(max_points >= 4, "woft_homography_fit_n: max_points %d", max_points);
  WOFT_REQUIRE(method >= WOFT_H_NONHOMOGENEOUS && method <= WOFT_H_IRLS_HUBER, "woft_homography_fit_n: method %d", method);
  WOFT_REQUIRE(n_iter >= 0 && n_iter <= 1000, "woft_homography_fit_n: n_iter %d", n_iter);
  HParams p{points1, points2, weights, H_out, max_points, n_points_dev, method, n_iter, huber_k};
  homography_fit_kernel<<<1, kHThreads, 0, static_cast<cudaStream_t>(stream)>>>(p);
  return launch_status("homography_fit_kernel");
}
