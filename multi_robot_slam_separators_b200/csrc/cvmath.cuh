// Double-precision small-matrix routines that replay, operation for operation, the arithmetic
// cv::solvePnPRansac executes per minimal sample (OpenCV 4.x calib3d / core, reached from
// R/src/myRegistrationVis.cpp:1077 through RTAB-Map's util3d::estimateMotion3DTo2D):
//   * cv::SVD for matrices with < 25 rows  -> OpenCV's own one-sided Jacobi (JacobiSVDImpl_)
//   * cv::solve / cv::invert with DECOMP_SVD -> the same SVD + SVBkSb back-substitution
//   * epnp::compute_pose (control points, barycentric alphas, M^T M null space, 3 beta
//     initialisations + Gauss-Newton, Procrustes R|t), cv::Rodrigues, cv::projectPoints.
// The ORDER of every +,-,*,/,sqrt matters: for 5 points M^T M has an exactly 2-dimensional null space
// whose basis is decided by rounding noise, and EPnP's linearisation depends on that basis
// (measured: re-orienting it moves ~15 % of hypothesis poses by > 1e-4).  The translation unit that
// includes this header is compiled with -fmad=false so that no multiply-add is contracted, matching
// the x86-64 SSE2 build of cv2.  Written __host__ __device__ so the same text can be checked against
// cv2 on the CPU by tests/host_harness (test-only build; the product path is the CUDA kernel).
//
// Storage: `Heap` is a strided view of per-thread scratch (stride 16 doubles in shared memory: the
// 16 solving lanes of a warp cover all 32 banks exactly once; stride 1 in the host test build).
#pragma once
#include <math.h>
#include <stdint.h>

#if defined(__CUDACC__) && defined(SEPB_K3_PROFILE)
#define CVM_HD __device__   // the counters of the developer build are device-only
#define CVM_NOINLINE __noinline__
#define CVM_UNROLL _Pragma("unroll")
#define CVM_NOUNROLL _Pragma("unroll 1")
#elif defined(__CUDACC__)
#define CVM_HD __host__ __device__
#define CVM_NOINLINE __noinline__
#define CVM_UNROLL _Pragma("unroll")
#define CVM_NOUNROLL _Pragma("unroll 1")
#else
#define CVM_HD
#define CVM_NOINLINE __attribute__((noinline))
#define CVM_UNROLL
#define CVM_NOUNROLL
#endif

#ifndef K3_T0   // phase counters of the developer build (pnp_ransac.cu); no-ops everywhere else
#define K3_T0(v)
#define K3_ACC(slot, v)
#endif

#ifndef SEPB_JACOBI_REGROWS
// 1: the two rows of a Jacobi rotation stay in registers between the dot product and the rotation; 0 (developer A/B):
// re-read from the scratch heap per element, each load ordered behind the previous element's stores
#define SEPB_JACOBI_REGROWS 1
#endif
#ifndef SEPB_JACOBI_BRANCHFREE
#define SEPB_JACOBI_BRANCHFREE 1   // 0 (developer A/B): the branches of the scalar code
#endif
#ifndef SEPB_JACOBI_ILP
// 1: 12 x 12 sweep with two row-disjoint rotations in flight (jacobi12_sweeps_ilp).  Measured on B200 (tools/k3_ab.sh):
// 10 % fewer cycles per warp but no shorter kernel (26.85 vs 26.57 ms verify), so the plain cyclic loop stays the default.
#define SEPB_JACOBI_ILP 0
#endif

namespace cvm {

constexpr double kDblEps = 2.220446049250313e-16;
constexpr double kDblMin = 2.2250738585072014e-308;

struct Intrinsics {
  double fx, fy, cx, cy;
};

// compile-time stride: element i of a thread's scratch lives at p[i * kHeapStride], so every access with
// a constant index is one LDS with an immediate offset (no 64-bit address arithmetic in the hot loops)
#if defined(__CUDACC__)
constexpr int kHeapStride = 32;   // == lanes of a warp that solve hypotheses at the same time (pnp_ransac.cu)
#else
constexpr int kHeapStride = 1;
#endif
#if defined(__CUDA_ARCH__)
// Device build: the scratch always lives in shared memory, so the view keeps a 32-bit shared-space address and goes
// through ld/st.shared (LDS/STS with immediate offsets) instead of generic LD/ST + 64-bit address arithmetic.
struct HeapRef {
  uint32_t a;
  __device__ __forceinline__ operator double() const {
    double v;
    asm volatile("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"(a));
    return v;
  }
  __device__ __forceinline__ const HeapRef& operator=(double v) const {
    asm volatile("st.shared.f64 [%0], %1;" ::"r"(a), "d"(v));   // ordered against the other volatile accesses; no other alias
    return *this;
  }
  __device__ __forceinline__ const HeapRef& operator=(const HeapRef& o) const { return *this = (double)o; }
  __device__ __forceinline__ const HeapRef& operator*=(double v) const { return *this = (double)*this * v; }
};
struct Heap {
  uint32_t a;
  __device__ __forceinline__ Heap(double* p) : a((uint32_t)__cvta_generic_to_shared(p)) {}
  __device__ __forceinline__ Heap(uint32_t addr, int) : a(addr) {}
  __device__ __forceinline__ HeapRef operator[](int i) const { return HeapRef{a + (uint32_t)(i * kHeapStride * 8)}; }
  __device__ __forceinline__ Heap at(int off) const { return Heap(a + (uint32_t)(off * kHeapStride * 8), 0); }
};
#else
struct Heap {
  double* p;
  CVM_HD double& operator[](int i) const { return p[i * kHeapStride]; }
  CVM_HD Heap at(int off) const { return Heap{p + off * kHeapStride}; }
};
#endif

// cv::RNG (multiply-with-carry)
struct Rng {
  uint64_t state;
  CVM_HD explicit Rng(uint64_t s) : state(s) {}
  CVM_HD uint32_t next() {
    state = (uint64_t)(uint32_t)state * 4164903690ULL + (state >> 32);
    return (uint32_t)state;
  }
};

// file-local hypot of OpenCV's lapack.cpp (not libm's)
CVM_HD inline double cv_hypot(double a, double b) {
  a = fabs(a);
  b = fabs(b);
  if (a > b) {
    b /= a;
    return a * sqrt(1 + b * b);
  }
  if (b > 0) {
    a /= b;
    return b * sqrt(1 + a * a);
  }
  return 0;
}

// rotation angle of JacobiSVDImpl_ for the row pair with squared norms a, b and dot product p:
//   p *= 2; beta = a - b; gamma = hypot(p, beta);
//   beta < 0:  delta = (gamma - beta) / 2; s = sqrt(delta / gamma); c = p / (gamma * s * 2)
//   else:      c = sqrt((gamma + beta) / (gamma * 2));              s = p / (gamma * c * 2)
// The SAME operations on the same operands, with the operands SELECTED instead of branched on (cv_hypot's two cases
// and the sign of beta): the lanes of a warp each hold their own hypothesis, and a branch here makes the warp run the
// divide / square-root chains of both sides one after the other.
CVM_HD inline void jacobi_cs(double p, double a, double b, double& c, double& s) {
#if SEPB_JACOBI_BRANCHFREE
  p *= 2;
  const double beta = a - b;
  const double ha = fabs(p), hb = fabs(beta);
  const bool agtb = ha > hb;
  const double big = agtb ? ha : hb, small = agtb ? hb : ha;
  const double q = small / big;
  const double gq = big * sqrt(1 + q * q);
  const double gamma = (agtb || hb > 0) ? gq : 0.0;
  const bool neg = beta < 0;
  const double num = neg ? (gamma - beta) * 0.5 : (gamma + beta);
  const double den = neg ? gamma : gamma * 2;
  const double r1 = sqrt(num / den);
  const double r2 = p / (gamma * r1 * 2);
  s = neg ? r1 : r2;
  c = neg ? r2 : r1;
#else
  p *= 2;
  const double beta = a - b, gamma = cv_hypot(p, beta);
  if (beta < 0) {
    const double delta = (gamma - beta) * 0.5;
    s = sqrt(delta / gamma);
    c = p / (gamma * s * 2);
  } else {
    c = sqrt((gamma + beta) / (gamma * 2));
    s = p / (gamma * c * 2);
  }
#endif
}

// tail of JacobiSVDImpl_: singular values, descending sort, normalisation (random completion of zero rows)
template <int m, int n, bool want_v>
CVM_HD inline void jacobi_finish(Heap At, Heap W, Heap Vt) {
  const double eps = kDblEps * 10;
  CVM_NOUNROLL
  for (int i = 0; i < n; ++i) {
    double sd = 0;
    CVM_UNROLL
    for (int k = 0; k < m; ++k) {
      const double t = At[i * m + k];
      sd += t * t;
    }
    W[i] = sqrt(sd);
  }
  CVM_NOUNROLL
  for (int i = 0; i < n - 1; ++i) {
    int j = i;
    for (int k = i + 1; k < n; ++k)
      if (W[j] < W[k]) j = k;
    if (i != j) {
      double t = W[i]; W[i] = W[j]; W[j] = t;
      for (int k = 0; k < m; ++k) { t = At[i * m + k]; At[i * m + k] = At[j * m + k]; At[j * m + k] = t; }
      if (want_v)
        for (int k = 0; k < n; ++k) { t = Vt[i * n + k]; Vt[i * n + k] = Vt[j * n + k]; Vt[j * n + k] = t; }
    }
  }
  Rng rng(0x12345678);
  CVM_NOUNROLL
  for (int i = 0; i < n; ++i) {
    double sd = W[i];
    CVM_NOUNROLL
    for (int ii = 0; ii < 100 && sd <= kDblMin; ++ii) {
      // zero singular value: random +-1/m vector, orthogonalised against the previous rows
      const double val0 = 1. / m;
      CVM_NOUNROLL
      for (int k = 0; k < m; ++k) At[i * m + k] = (rng.next() & 256) != 0 ? val0 : -val0;
      CVM_NOUNROLL
      for (int it = 0; it < 2; ++it)
        CVM_NOUNROLL
        for (int j = 0; j < i; ++j) {
          sd = 0;
          CVM_NOUNROLL
          for (int k = 0; k < m; ++k) sd += At[i * m + k] * At[j * m + k];
          double asum = 0;
          CVM_NOUNROLL
          for (int k = 0; k < m; ++k) {
            const double t = At[i * m + k] - sd * At[j * m + k];
            At[i * m + k] = t;
            asum += fabs(t);
          }
          asum = asum > eps * 100 ? 1 / asum : 0;
          CVM_NOUNROLL
          for (int k = 0; k < m; ++k) At[i * m + k] *= asum;
        }
      sd = 0;
      CVM_NOUNROLL
      for (int k = 0; k < m; ++k) {
        const double t = At[i * m + k];
        sd += t * t;
      }
      sd = sqrt(sd);
    }
    const double s = sd > kDblMin ? 1 / sd : 0.;
    CVM_UNROLL
    for (int k = 0; k < m; ++k) At[i * m + k] *= s;
  }
}

// JacobiSVDImpl_<double>: At is n rows of length m (row r at At[r*m + k]); on return rows of At are the
// unit left singular vectors of A = At^T, W descending, Vt (n x n, optional) the right vectors.
// n1 = number of rows to normalise / complete (== n at every call site on this path).
// m, n, want_v are template parameters so that the loops over k unroll: the loads of a rotation are
// issued together and the independent element updates overlap; the ORDER of the arithmetic on every
// accumulator is unchanged (sequential in k), which is what bit-exactness against cv2 needs.
template <int m, int n, bool want_v>
CVM_HD CVM_NOINLINE inline void jacobi_svd_fixed(Heap At, Heap W, Heap Vt) {
  const double eps = kDblEps * 10;
  for (int i = 0; i < n; ++i) {
    double sd = 0;
    CVM_UNROLL
    for (int k = 0; k < m; ++k) {
      const double t = At[i * m + k];
      sd += t * t;
    }
    W[i] = sd;
    if (want_v) {
      CVM_UNROLL
      for (int k = 0; k < n; ++k) Vt[i * n + k] = 0;
      Vt[i * n + i] = 1;
    }
  }
  const int max_iter = m > 30 ? m : 30;
  CVM_NOUNROLL
  for (int iter = 0; iter < max_iter; ++iter) {
    bool changed = false;
#if defined(SEPB_K3_PROFILE) && defined(__CUDA_ARCH__)
    if (m == 12 && (threadIdx.x & 31) == 0) atomicAdd(&g_k3_prof[6], 1ull);
#endif
    CVM_NOUNROLL
    for (int i = 0; i < n - 1; ++i)
      CVM_NOUNROLL
      for (int j = i + 1; j < n; ++j) {
        const Heap Ai = At.at(i * m), Aj = At.at(j * m);
        double a = W[i], p = 0, b = W[j];
        // both rows stay in registers from the dot product to the rotation (the scratch accessors are ordered volatile
        // accesses: re-reading them per element would serialise every load behind the previous element's stores)
        double x[m], y[m];
        CVM_UNROLL
        for (int k = 0; k < m; ++k) {
          x[k] = Ai[k];
          y[k] = Aj[k];
        }
        CVM_UNROLL
        for (int k = 0; k < m; ++k) p += x[k] * y[k];
#if !SEPB_JACOBI_REGROWS
#define CVM_ROW_X(k) ((double)Ai[k])
#define CVM_ROW_Y(k) ((double)Aj[k])
#else
#define CVM_ROW_X(k) x[k]
#define CVM_ROW_Y(k) y[k]
#endif
        if (fabs(p) <= eps * sqrt(a * b)) continue;
        double c, s;
        jacobi_cs(p, a, b, c, s);
        a = b = 0;
        CVM_UNROLL
        for (int k = 0; k < m; ++k) {
          const double xk = CVM_ROW_X(k), yk = CVM_ROW_Y(k);
          const double t0 = c * xk + s * yk;
          const double t1 = -s * xk + c * yk;
          Ai[k] = t0;
          Aj[k] = t1;
          a += t0 * t0;
          b += t1 * t1;
        }
        W[i] = a;
        W[j] = b;
        changed = true;
        if (want_v) {
          const Heap Vi = Vt.at(i * n), Vj = Vt.at(j * n);
          double vx[n], vy[n];
          CVM_UNROLL
          for (int k = 0; k < n; ++k) {
            vx[k] = Vi[k];
            vy[k] = Vj[k];
          }
          CVM_UNROLL
          for (int k = 0; k < n; ++k) {
            Vi[k] = c * vx[k] + s * vy[k];
            Vj[k] = -s * vx[k] + c * vy[k];
          }
        }
      }
    if (!changed) break;
  }
  jacobi_finish<m, n, want_v>(At, W, Vt);
}

// run-time n (one body for every 6 x n solve of the path: fewer instructions to fetch; same operations) -- tail: singular values, descending sort, normalisation (random completion of zero rows)
template <int m, bool want_v>
CVM_HD inline void jacobi_finish_rt(Heap At, Heap W, Heap Vt, const int n) {
  const double eps = kDblEps * 10;
  CVM_NOUNROLL
  for (int i = 0; i < n; ++i) {
    double sd = 0;
    CVM_UNROLL
    for (int k = 0; k < m; ++k) {
      const double t = At[i * m + k];
      sd += t * t;
    }
    W[i] = sqrt(sd);
  }
  CVM_NOUNROLL
  for (int i = 0; i < n - 1; ++i) {
    int j = i;
    for (int k = i + 1; k < n; ++k)
      if (W[j] < W[k]) j = k;
    if (i != j) {
      double t = W[i]; W[i] = W[j]; W[j] = t;
      for (int k = 0; k < m; ++k) { t = At[i * m + k]; At[i * m + k] = At[j * m + k]; At[j * m + k] = t; }
      if (want_v)
        for (int k = 0; k < n; ++k) { t = Vt[i * n + k]; Vt[i * n + k] = Vt[j * n + k]; Vt[j * n + k] = t; }
    }
  }
  Rng rng(0x12345678);
  CVM_NOUNROLL
  for (int i = 0; i < n; ++i) {
    double sd = W[i];
    CVM_NOUNROLL
    for (int ii = 0; ii < 100 && sd <= kDblMin; ++ii) {
      // zero singular value: random +-1/m vector, orthogonalised against the previous rows
      const double val0 = 1. / m;
      CVM_NOUNROLL
      for (int k = 0; k < m; ++k) At[i * m + k] = (rng.next() & 256) != 0 ? val0 : -val0;
      CVM_NOUNROLL
      for (int it = 0; it < 2; ++it)
        CVM_NOUNROLL
        for (int j = 0; j < i; ++j) {
          sd = 0;
          CVM_NOUNROLL
          for (int k = 0; k < m; ++k) sd += At[i * m + k] * At[j * m + k];
          double asum = 0;
          CVM_NOUNROLL
          for (int k = 0; k < m; ++k) {
            const double t = At[i * m + k] - sd * At[j * m + k];
            At[i * m + k] = t;
            asum += fabs(t);
          }
          asum = asum > eps * 100 ? 1 / asum : 0;
          CVM_NOUNROLL
          for (int k = 0; k < m; ++k) At[i * m + k] *= asum;
        }
      sd = 0;
      CVM_NOUNROLL
      for (int k = 0; k < m; ++k) {
        const double t = At[i * m + k];
        sd += t * t;
      }
      sd = sqrt(sd);
    }
    const double s = sd > kDblMin ? 1 / sd : 0.;
    CVM_UNROLL
    for (int k = 0; k < m; ++k) At[i * m + k] *= s;
  }
}

template <int m, bool want_v>
CVM_HD CVM_NOINLINE inline void jacobi_svd_rt(Heap At, Heap W, Heap Vt, const int n) {
  const double eps = kDblEps * 10;
  for (int i = 0; i < n; ++i) {
    double sd = 0;
    CVM_UNROLL
    for (int k = 0; k < m; ++k) {
      const double t = At[i * m + k];
      sd += t * t;
    }
    W[i] = sd;
    if (want_v) {
      for (int k = 0; k < n; ++k) Vt[i * n + k] = 0;
      Vt[i * n + i] = 1;
    }
  }
  const int max_iter = m > 30 ? m : 30;
  CVM_NOUNROLL
  for (int iter = 0; iter < max_iter; ++iter) {
    bool changed = false;
#if defined(SEPB_K3_PROFILE) && defined(__CUDA_ARCH__)
    if (m == 12 && (threadIdx.x & 31) == 0) atomicAdd(&g_k3_prof[6], 1ull);
#endif
    CVM_NOUNROLL
    for (int i = 0; i < n - 1; ++i)
      CVM_NOUNROLL
      for (int j = i + 1; j < n; ++j) {
        const Heap Ai = At.at(i * m), Aj = At.at(j * m);
        double a = W[i], p = 0, b = W[j];
        double x[m], y[m];   // rows kept in registers between the dot product and the rotation (see jacobi_svd_fixed)
        CVM_UNROLL
        for (int k = 0; k < m; ++k) {
          x[k] = Ai[k];
          y[k] = Aj[k];
        }
        CVM_UNROLL
        for (int k = 0; k < m; ++k) p += x[k] * y[k];
#if !SEPB_JACOBI_REGROWS
#define CVM_ROW_X(k) ((double)Ai[k])
#define CVM_ROW_Y(k) ((double)Aj[k])
#else
#define CVM_ROW_X(k) x[k]
#define CVM_ROW_Y(k) y[k]
#endif
        if (fabs(p) <= eps * sqrt(a * b)) continue;
        double c, s;
        jacobi_cs(p, a, b, c, s);
        a = b = 0;
        CVM_UNROLL
        for (int k = 0; k < m; ++k) {
          const double xk = CVM_ROW_X(k), yk = CVM_ROW_Y(k);
          const double t0 = c * xk + s * yk;
          const double t1 = -s * xk + c * yk;
          Ai[k] = t0;
          Aj[k] = t1;
          a += t0 * t0;
          b += t1 * t1;
        }
        W[i] = a;
        W[j] = b;
        changed = true;
        if (want_v) {
          const Heap Vi = Vt.at(i * n), Vj = Vt.at(j * n);
          for (int k = 0; k < n; ++k) {
            const double vx = Vi[k], vy = Vj[k];
            Vi[k] = c * vx + s * vy;
            Vj[k] = -s * vx + c * vy;
          }
        }
      }
    if (!changed) break;
  }
  jacobi_finish_rt<m, want_v>(At, W, Vt, n);
}

template <int m, int n, bool want_v>
CVM_HD inline void jacobi_svd(Heap At, Heap W, Heap Vt) {
  if constexpr (m == 6 && want_v) {
    jacobi_svd_rt<6, true>(At, W, Vt, n);   // 6 x {3, 4, 5, 6}: one shared body
  } else {
    jacobi_svd_fixed<m, n, want_v>(At, W, Vt);
  }
}

#if defined(__CUDACC__)
// 12 x 12 sweep of the M^T M Jacobi with kJacobiIlp rotations in flight per thread.  Rotation (i, j) touches rows
// i and j (and W[i], W[j]) only, so rotations on disjoint rows commute EXACTLY: any order that keeps, for every row, the
// cyclic order of the rotations touching it gives bit-identical results.  The table is a list schedule of the 66
// cyclic pairs into rounds of <= 2 row-disjoint pairs ({0, 0} = empty slot); the dependent chains of a round
// (dot product, hypot / sqrt / divide, rotated norms) interleave, which is what a lone warp per scheduler needs.
// The sweep loop is warp-uniform: a lane that has converged only sees skipped rotations (no stores) in extra sweeps.
constexpr int kJacobiIlp = 2;
constexpr int kJacobiRounds = 35;
static __constant__ unsigned char kJacobi12Sched[kJacobiRounds][kJacobiIlp][2] = {   // warp-uniform reads: constant cache

    {{0, 1}, {0, 0}}, {{0, 2}, {0, 0}}, {{0, 3}, {1, 2}}, {{0, 4}, {1, 3}}, {{0, 5}, {1, 4}},
    {{2, 3}, {0, 6}}, {{1, 5}, {2, 4}}, {{0, 7}, {1, 6}}, {{2, 5}, {3, 4}}, {{0, 8}, {1, 7}},
    {{2, 6}, {3, 5}}, {{0, 9}, {1, 8}}, {{2, 7}, {3, 6}}, {{4, 5}, {0, 10}}, {{1, 9}, {2, 8}},
    {{3, 7}, {4, 6}}, {{0, 11}, {1, 10}}, {{2, 9}, {3, 8}}, {{4, 7}, {5, 6}}, {{1, 11}, {2, 10}},
    {{3, 9}, {4, 8}}, {{5, 7}, {2, 11}}, {{3, 10}, {4, 9}}, {{5, 8}, {6, 7}}, {{3, 11}, {4, 10}},
    {{5, 9}, {6, 8}}, {{4, 11}, {5, 10}}, {{6, 9}, {7, 8}}, {{5, 11}, {6, 10}}, {{7, 9}, {6, 11}},
    {{7, 10}, {8, 9}}, {{7, 11}, {8, 10}}, {{8, 11}, {9, 10}}, {{9, 11}, {0, 0}}, {{10, 11}, {0, 0}}};
#endif
#if defined(__CUDA_ARCH__)
__device__ __forceinline__ double lds_f64(uint32_t a) {
  double v;
  asm volatile("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"(a));
  return v;
}
__device__ __forceinline__ void sts_f64(uint32_t a, double v) { asm volatile("st.shared.f64 [%0], %1;" ::"r"(a), "d"(v)); }

__device__ CVM_NOINLINE inline void jacobi12_sweeps_ilp(Heap At, Heap W) {
  constexpr int m = 12, n = 12, G = kJacobiIlp;
  constexpr uint32_t kEl = kHeapStride * 8, kRow = m * kEl;
  const double eps = kDblEps * 10;
  const unsigned mask = __activemask();
  for (int i = 0; i < n; ++i) {
    double sd = 0;
    CVM_UNROLL
    for (int k = 0; k < m; ++k) {
      const double t = At[i * m + k];
      sd += t * t;
    }
    W[i] = sd;
  }
  CVM_NOUNROLL
  for (int iter = 0; iter < 30; ++iter) {
    bool changed = false;
#if defined(SEPB_K3_PROFILE)
    if ((threadIdx.x & 31) == 0) atomicAdd(&g_k3_prof[6], 1ull);
#endif
    CVM_NOUNROLL
    for (int r = 0; r < kJacobiRounds; ++r) {
      uint32_t ai[G], aj[G], wi[G], wj[G];
      bool rot[G];
      double a[G], b[G], p[G];
      CVM_UNROLL
      for (int g = 0; g < G; ++g) {
        const int i = kJacobi12Sched[r][g][0], j = kJacobi12Sched[r][g][1];
        rot[g] = i != j;
        ai[g] = At.a + i * kRow;
        aj[g] = At.a + j * kRow;
        wi[g] = W.a + i * kEl;
        wj[g] = W.a + j * kEl;
        a[g] = lds_f64(wi[g]);
        b[g] = lds_f64(wj[g]);
        p[g] = 0;
      }
      // both rows of every slot stay in registers from the dot product to the rotation: no load has to wait behind a
      // store to a row the compiler cannot tell apart
      double x[G][m], y[G][m];
      CVM_UNROLL
      for (int k = 0; k < m; ++k) {
        CVM_UNROLL
        for (int g = 0; g < G; ++g) {
          x[g][k] = lds_f64(ai[g] + k * kEl);
          y[g][k] = lds_f64(aj[g] + k * kEl);
        }
      }
      CVM_UNROLL
      for (int k = 0; k < m; ++k) {
        CVM_UNROLL
        for (int g = 0; g < G; ++g) p[g] += x[g][k] * y[g][k];
      }
      bool any = false;
      CVM_UNROLL
      for (int g = 0; g < G; ++g) {
        rot[g] = rot[g] && !(fabs(p[g]) <= eps * sqrt(a[g] * b[g]));
        any |= rot[g];
      }
      if (!__any_sync(mask, any)) continue;
      double c[G], s[G];
      CVM_UNROLL
      for (int g = 0; g < G; ++g) {
        // same operations as the scalar branches of jacobi_svd / cv_hypot, selected instead of branched
        const double pp = p[g] * 2, beta = a[g] - b[g];
        const double ha = fabs(pp), hb = fabs(beta);
        const bool agtb = ha > hb;
        const double big = agtb ? ha : hb, small = agtb ? hb : ha;
        const double q = small / big;
        const double gamma = big > 0 ? big * sqrt(1 + q * q) : 0.0;
        const bool neg = beta < 0;
        const double num = neg ? (gamma - beta) * 0.5 : (gamma + beta);
        const double den = neg ? gamma : gamma * 2;
        const double r1 = sqrt(num / den);
        const double r2 = pp / (gamma * r1 * 2);
        s[g] = neg ? r1 : r2;
        c[g] = neg ? r2 : r1;
        a[g] = 0;
        b[g] = 0;
      }
      CVM_UNROLL
      for (int k = 0; k < m; ++k) {
        CVM_UNROLL
        for (int g = 0; g < G; ++g) {
          const double t0 = c[g] * x[g][k] + s[g] * y[g][k];
          const double t1 = -s[g] * x[g][k] + c[g] * y[g][k];
          if (rot[g]) {
            sts_f64(ai[g] + k * kEl, t0);
            sts_f64(aj[g] + k * kEl, t1);
          }
          a[g] += t0 * t0;
          b[g] += t1 * t1;
        }
      }
      CVM_UNROLL
      for (int g = 0; g < G; ++g)
        if (rot[g]) {
          sts_f64(wi[g], a[g]);
          sts_f64(wj[g], b[g]);
          changed = true;
        }
    }
    if (!__any_sync(mask, changed)) break;
  }
  jacobi_finish<12, 12, false>(At, W, W);
}
#endif

#if defined(__CUDACC__)
// wavefront schedule of the 15 cyclic pairs of a 6 x 6 sweep: pair (i, j) runs in round i + j; the pairs of a round are
// row-disjoint and every row still meets its rotations in cyclic order, so the result is bit-identical to the serial loop
constexpr int kJacobi6Rounds = 9;
static __constant__ unsigned char kJacobi6Sched[kJacobi6Rounds][3][2] = {
    {{0, 1}, {0, 0}, {0, 0}}, {{0, 2}, {0, 0}, {0, 0}}, {{0, 3}, {1, 2}, {0, 0}}, {{0, 4}, {1, 3}, {0, 0}},
    {{0, 5}, {1, 4}, {2, 3}}, {{1, 5}, {2, 4}, {0, 0}}, {{2, 5}, {3, 4}, {0, 0}}, {{3, 5}, {0, 0}, {0, 0}},
    {{4, 5}, {0, 0}, {0, 0}}};
// JacobiSVDImpl_ for ONE 6 x 6 problem (the normal equations of the refit) with the whole warp on it: lanes 0..2 each
// rotate one pair of the round.  Rotations on disjoint rows commute exactly, every rotation is the serial code's, so
// At / W / Vt end up bit-identical to jacobi_svd_rt<6, true>(.., 6) -- in 9 dependent rotation chains per sweep
// instead of 15.  All lanes pass the SAME scratch (one lane's slice); the whole warp must call it.
__device__ CVM_NOINLINE inline void jacobi_svd6_warp(Heap At, Heap W, Heap Vt) {
  constexpr int m = 6, n = 6;
  const double eps = kDblEps * 10;
  const int lane = threadIdx.x & 31;
  if (lane == 0) {
    for (int i = 0; i < n; ++i) {
      double sd = 0;
      CVM_UNROLL
      for (int k = 0; k < m; ++k) {
        const double t = At[i * m + k];
        sd += t * t;
      }
      W[i] = sd;
      for (int k = 0; k < n; ++k) Vt[i * n + k] = 0;
      Vt[i * n + i] = 1;
    }
  }
  __syncwarp();
  CVM_NOUNROLL
  for (int iter = 0; iter < 30; ++iter) {
    bool changed = false;
    CVM_NOUNROLL
    for (int r = 0; r < kJacobi6Rounds; ++r) {
      const int g = lane < 3 ? lane : 0;
      const int i = kJacobi6Sched[r][g][0], j = kJacobi6Sched[r][g][1];
      if (lane < 3 && i != j) {
        const Heap Ai = At.at(i * m), Aj = At.at(j * m);
        double a = W[i], p = 0, b = W[j];
        double x[m], y[m];
        CVM_UNROLL
        for (int k = 0; k < m; ++k) {
          x[k] = Ai[k];
          y[k] = Aj[k];
        }
        CVM_UNROLL
        for (int k = 0; k < m; ++k) p += x[k] * y[k];
        if (!(fabs(p) <= eps * sqrt(a * b))) {
          double c, s;
          jacobi_cs(p, a, b, c, s);
          a = b = 0;
          CVM_UNROLL
          for (int k = 0; k < m; ++k) {
            const double t0 = c * x[k] + s * y[k];
            const double t1 = -s * x[k] + c * y[k];
            Ai[k] = t0;
            Aj[k] = t1;
            a += t0 * t0;
            b += t1 * t1;
          }
          W[i] = a;
          W[j] = b;
          changed = true;
          const Heap Vi = Vt.at(i * n), Vj = Vt.at(j * n);
          double vx[n], vy[n];
          CVM_UNROLL
          for (int k = 0; k < n; ++k) {
            vx[k] = Vi[k];
            vy[k] = Vj[k];
          }
          CVM_UNROLL
          for (int k = 0; k < n; ++k) {
            Vi[k] = c * vx[k] + s * vy[k];
            Vj[k] = -s * vx[k] + c * vy[k];
          }
        }
      }
      __syncwarp();
    }
    if (!__any_sync(0xFFFFFFFFu, changed)) break;
  }
  if (lane == 0) jacobi_finish_rt<6, true>(At, W, Vt, 6);
  __syncwarp();
}
#endif

// SVBkSb, one right-hand side: x[n] = V diag(1/w) U^T b, with U^T = rows of At (n x m), Vt n x n.
// m, n are template parameters so that b[] / x[] keep static indices (registers, not local memory).
template <int m, int n>
CVM_HD inline void svd_backsubst(Heap At, Heap W, Heap Vt, const double* b_in, double* x_out) {
  double b[m], x[n];   // private copies: the caller's arrays may alias and would be re-read after every store
  CVM_UNROLL
  for (int k = 0; k < m; ++k) b[k] = b_in[k];
  double threshold = 0;
  CVM_UNROLL
  for (int i = 0; i < n; ++i) threshold += W[i];
  threshold *= kDblEps * 2;
  CVM_UNROLL
  for (int k = 0; k < n; ++k) x[k] = 0;
  CVM_UNROLL
  for (int i = 0; i < n; ++i) {
    double wi = W[i];
    if (fabs(wi) <= threshold) continue;
    wi = 1 / wi;
    double s = 0;
    CVM_UNROLL
    for (int k = 0; k < m; ++k) s += At[i * m + k] * b[k];
    s *= wi;
    CVM_UNROLL
    for (int k = 0; k < n; ++k) x[k] = x[k] + s * Vt[i * n + k];
  }
  CVM_UNROLL
  for (int k = 0; k < n; ++k) x_out[k] = x[k];
}

}  // namespace cvm
