// K5-K8: Gram / Rayleigh-Ritz reductions, single-CTA Jacobi + Cholesky, small applies.
// Replaces np.linalg.qr (_svds.py:508) and the dense svd of the projected matrix
// (_svds.py:518) inside scipy svds as reached from model.py:54-55.
#include "k5_orth.cuh"
#include <math.h>

namespace n2v2r {

// ---- X^T Y partials: each CTA reduces a slab of rows into an l x l fp64 block ------------
// 256 threads as 16 x 16; thread (ag, bg) owns out[ag + 16 i][bg + 16 j], i, j < 5 (l <= 80):
// NB x NB fp64 accumulators in registers (NB = ceil(l / 16): the loads, float -> double conversions and DFMAs of column groups
// beyond l are not issued; the sums are the same bits for every NB), per row NB + NB shared loads for NB^2 DFMA.
template <int NB>
__global__ void __launch_bounds__(256) xty_kernel(const float* __restrict__ X, const float* __restrict__ Y, int64_t n, int l,
                                                  int rows_per_cta, double* __restrict__ partials) {
  extern __shared__ float sm[];             // 2 * 32 * l
  float* sx = sm;
  float* sy = sm + 32 * l;
  const int64_t r0 = (int64_t)blockIdx.x * rows_per_cta;
  const int64_t r1 = (r0 + rows_per_cta < n) ? r0 + rows_per_cta : n;
  const int ag = threadIdx.x & 15, bg = threadIdx.x >> 4;
  double acc[NB][NB];
#pragma unroll
  for (int i = 0; i < NB; ++i)
#pragma unroll
    for (int j = 0; j < NB; ++j) acc[i][j] = 0.0;
  for (int64_t rb = r0; rb < r1; rb += 32) {
    const int rows = (int)((r1 - rb < 32) ? (r1 - rb) : 32);
    for (int t = threadIdx.x; t < 32 * l; t += 256) {
      const int r = t / l;
      sx[t] = r < rows ? X[(rb + r) * l + (t - r * l)] : 0.f;
      sy[t] = r < rows ? Y[(rb + r) * l + (t - r * l)] : 0.f;
    }
    __syncthreads();
#pragma unroll 4
    for (int r = 0; r < 32; ++r) {
      double xv[NB], yv[NB];
#pragma unroll
      for (int i = 0; i < NB; ++i) {
        xv[i] = (ag + 16 * i < l) ? (double)sx[r * l + ag + 16 * i] : 0.0;
        yv[i] = (bg + 16 * i < l) ? (double)sy[r * l + bg + 16 * i] : 0.0;
      }
#pragma unroll
      for (int i = 0; i < NB; ++i)
#pragma unroll
        for (int j = 0; j < NB; ++j) acc[i][j] = fma(xv[i], yv[j], acc[i][j]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < NB; ++i)
#pragma unroll
    for (int j = 0; j < NB; ++j) {
      const int a_ = ag + 16 * i, b_ = bg + 16 * j;
      if (a_ < l && b_ < l) partials[(size_t)blockIdx.x * l * l + a_ * l + b_] = acc[i][j];
    }
}

// fixed-order sum of the per-CTA partials (deterministic): 256 threads as 32 elements x 8 slices; slice q adds the partials
// q, q + 8, q + 16, ... in order (a warp reads 32 consecutive elements of one partial), then the 8 slice sums are added in
// order.  The order depends only on n_parts, never on timing.
__global__ void __launch_bounds__(256) sum_partials_kernel(const double* __restrict__ partials, int n_parts, int count,
                                                           double* __restrict__ out, int accumulate) {
  __shared__ double slice[8][33];
  const int el = threadIdx.x & 31, q = threadIdx.x >> 5;
  const int e = blockIdx.x * 32 + el;
  double s = 0.0;
  if (e < count)
    for (int p = q; p < n_parts; p += 8) s += partials[(size_t)p * count + e];
  slice[q][el] = s;
  __syncthreads();
  if (q == 0 && e < count) {
    double t = accumulate ? out[e] : 0.0;
#pragma unroll
    for (int k = 0; k < 8; ++k) t += slice[k][el];
    out[e] = t;
  }
}

// ---- single-CTA dense helpers on l x l fp64 matrices in shared memory ----------------------
__device__ void jacobi_eig(double* A, double* V, int l, double* cs, int* pq, double* red) {
  const int tid = threadIdx.x, nt = blockDim.x;
  const int half = l / 2;
  for (int e = tid; e < l * l; e += nt) V[e] = ((e / l) == (e % l)) ? 1.0 : 0.0;
  __syncthreads();
  for (int sweep = 0; sweep < 40; ++sweep) {
    // off-diagonal and total norms
    double off = 0.0, tot = 0.0;
    for (int e = tid; e < l * l; e += nt) {
      const double v = A[e] * A[e];
      tot += v;
      if ((e / l) != (e % l)) off += v;
    }
    off = warp_sum(off); tot = warp_sum(tot);
    if ((tid & 31) == 0) { red[(tid >> 5) * 2] = off; red[(tid >> 5) * 2 + 1] = tot; }
    __syncthreads();
    off = 0.0; tot = 0.0;
    for (int w = 0; w < (nt >> 5); ++w) { off += red[w * 2]; tot += red[w * 2 + 1]; }
    __syncthreads();
    // converged at off(A) <= l eps ||A||_F: the roundoff floor of l(l-1) off-diagonal entries of size eps ||A|| / sqrt(l)
    // (a tighter test is never met and costs all 40 sweeps)
    const double lev = (double)l * 2.3e-16;
    if (off <= lev * lev * tot || tot == 0.0) break;
    for (int r = 0; r < l - 1; ++r) {
      if (tid < half) {
        int p, q;
        if (tid == 0) { p = l - 1; q = r; }
        else { p = (r + tid) % (l - 1); q = (r - tid + (l - 1)) % (l - 1); }
        // no (p < q) ordering: p_i = r + i and q_i = r - i (mod l - 1) are runs of consecutive columns, which keeps the
        // 64-bit shared-memory accesses of a warp in the block / V updates below free of bank conflicts
        const double apq = A[p * l + q], app = A[p * l + p], aqq = A[q * l + q];
        double c = 1.0, s = 0.0;
        if (fabs(apq) > 1e-300 && apq * apq > 1e-36 * fabs(app * aqq)) {
          // t = tan(angle) = b / (d + sign(d) sqrt(d^2 + b^2)), d = aqq - app, b = 2 apq (the smaller root);
          // rsqrt / rcp sequences instead of two divisions and two square roots: this chain is the round's latency
          const double d = aqq - app, b2 = 2.0 * apq;
          const double x = fma(d, d, b2 * b2);
          if (x > 1e-280 && x < 1e280) {
            const double rr = x * rsqrt(x);
            const double t = b2 * __drcp_rn(d + copysign(rr, d));
            c = rsqrt(fma(t, t, 1.0));
            s = t * c;
          } else {                                   // out of the squared range: the division / sqrt form
            const double theta = d / b2;
            const double t = (theta >= 0.0 ? 1.0 : -1.0) / (fabs(theta) + sqrt(theta * theta + 1.0));
            c = 1.0 / sqrt(t * t + 1.0);
            s = t * c;
          }
        }
        cs[2 * tid] = c; cs[2 * tid + 1] = s; pq[2 * tid] = p; pq[2 * tid + 1] = q;
      }
      __syncthreads();
      // A <- J^T A J one 2 x 2 block per (row pair j, column pair i): blocks are disjoint, so a single phase does both sides
      for (int e = tid; e < half * half; e += nt) {
        const int j = e / half, i = e - j * half;
        const int pj = pq[2 * j], qj = pq[2 * j + 1], pi = pq[2 * i], qi = pq[2 * i + 1];
        const double cj = cs[2 * j], sj = cs[2 * j + 1], ci = cs[2 * i], si = cs[2 * i + 1];
        const double a00 = A[pj * l + pi], a01 = A[pj * l + qi], a10 = A[qj * l + pi], a11 = A[qj * l + qi];
        const double b00 = ci * a00 - si * a01, b01 = si * a00 + ci * a01;       // columns
        const double b10 = ci * a10 - si * a11, b11 = si * a10 + ci * a11;
        A[pj * l + pi] = cj * b00 - sj * b10;                                     // rows
        A[pj * l + qi] = cj * b01 - sj * b11;
        A[qj * l + pi] = sj * b00 + cj * b10;
        A[qj * l + qi] = sj * b01 + cj * b11;
      }
      // V <- V J
      for (int e = tid; e < l * half; e += nt) {
        const int k = e / half, i = e - k * half;
        const int p = pq[2 * i], q = pq[2 * i + 1];
        const double c = cs[2 * i], s = cs[2 * i + 1];
        const double vkp = V[k * l + p], vkq = V[k * l + q];
        V[k * l + p] = c * vkp - s * vkq;
        V[k * l + q] = s * vkp + c * vkq;
      }
      __syncthreads();
    }
  }
}

// sort eigenpairs descending: lam_out[rank] , W[:, rank] = V[:, i]
__device__ void sort_eig_desc(const double* A, const double* V, int l, double* lam, double* W) {
  const int tid = threadIdx.x, nt = blockDim.x;
  for (int i = tid; i < l; i += nt) {
    const double li = A[i * l + i];
    int rank = 0;
    for (int j = 0; j < l; ++j) {
      const double lj = A[j * l + j];
      rank += (lj > li) || (lj == li && j < i);
    }
    lam[rank] = li;
    for (int k = 0; k < l; ++k) W[k * l + rank] = V[k * l + i];
  }
  __syncthreads();
}

// C = A * B (l x l), all in shared memory
__device__ void matmul_ll(const double* A, const double* B, double* C, int l, bool transA) {
  for (int e = threadIdx.x; e < l * l; e += blockDim.x) {
    const int i = e / l, j = e - i * l;
    double s = 0.0;
    for (int k = 0; k < l; ++k) s += (transA ? A[k * l + i] : A[i * l + k]) * B[k * l + j];
    C[e] = s;
  }
  __syncthreads();
}

// In-place upper Cholesky M = R^T R; returns false on breakdown (pivot <= eps * max diag).
__device__ bool cholesky_upper(double* M, int l, int* flag) {
  const int tid = threadIdx.x, nt = blockDim.x;
  double dmax = 0.0;
  for (int i = 0; i < l; ++i) dmax = fmax(dmax, M[i * l + i]);
  if (tid == 0) *flag = 0;
  __syncthreads();
  for (int j = 0; j < l; ++j) {
    const double piv = M[j * l + j];
    if (!(piv > 1e-13 * dmax)) { if (tid == 0) *flag = 1; __syncthreads(); return false; }
    __syncthreads();
    const double d = sqrt(piv);
    for (int k = j + tid; k < l; k += nt) M[j * l + k] = M[j * l + k] / d;   // row j of R
    __syncthreads();
    for (int e = tid; e < (l - j - 1) * (l - j - 1); e += nt) {
      const int a = j + 1 + e / (l - j - 1), b = j + 1 + e % (l - j - 1);
      if (b >= a) M[a * l + b] -= M[j * l + a] * M[j * l + b];
    }
    __syncthreads();
  }
  for (int e = tid; e < l * l; e += nt) if ((e % l) < (e / l)) M[e] = 0.0;
  __syncthreads();
  return true;
}

// X = R^{-1} for upper-triangular R (thread j solves column j)
__device__ void upper_inverse(const double* R, double* X, int l) {
  for (int j = threadIdx.x; j < l; j += blockDim.x) {
    for (int i = l - 1; i >= 0; --i) {
      if (i > j) { X[i * l + j] = 0.0; continue; }
      double s = (i == j) ? 1.0 : 0.0;
      for (int k = i + 1; k <= j; ++k) s -= R[i * l + k] * X[k * l + j];
      X[i * l + j] = s / R[i * l + i];
    }
  }
  __syncthreads();
}

// Inverse square root of an SPD (possibly rank-deficient) matrix through Jacobi: X = V D^-1/2
// (columns with D_i <= 1e-13 D_max are scaled as if D_i were that floor).  M is destroyed.
__device__ void inv_sqrt_by_eig(double* M, double* V, double* X, int l, double* cs, int* pq, double* red) {
  jacobi_eig(M, V, l, cs, pq, red);
  double dmax = 0.0;
  for (int i = 0; i < l; ++i) dmax = fmax(dmax, M[i * l + i]);
  const double floor_ = fmax(dmax * 1e-13, 1e-300);
  for (int e = threadIdx.x; e < l * l; e += blockDim.x) {
    const int j = e % l;
    X[e] = V[e] / sqrt(fmax(M[j * l + j], floor_));
  }
  __syncthreads();
}

// mode 0 (RR):    in H = Q^T Y, S = Y^T Y  -> out W (sorted Ritz vectors), lam, C = W R^-1 with R = chol(W^T S W)
// mode 1 (CHOL):  in G                      -> out C = chol(G)^-1
// mode 2 (FINAL): in G = sum_t Z_t^T Z_t    -> out W, lam (= sigma^2), Cemb = W diag(lam^-1/4), Cleft = W diag(lam^1/4)
__global__ void __launch_bounds__(512) smallmat_kernel(int mode, int l, const double* __restrict__ gH, const double* __restrict__ gS,
                                                       double* __restrict__ gW, double* __restrict__ gLam,
                                                       double* __restrict__ gC, double* __restrict__ gC2, int* __restrict__ gFlag) {
  extern __shared__ double smd[];
  double* A = smd;                 // l*l
  double* V = A + l * l;
  double* B = V + l * l;
  double* Cm = B + l * l;
  double* lam = Cm + l * l;        // l
  double* cs = lam + l;            // l
  double* red = cs + l;            // 64
  int* pq = (int*)(red + 64);      // l
  __shared__ int flag;
  const int tid = threadIdx.x, nt = blockDim.x;
  const int ll = l * l;
  if (mode == 0 || mode == 2) {
    for (int e = tid; e < ll; e += nt) { const int i = e / l, j = e % l; A[e] = 0.5 * (gH[i * l + j] + gH[j * l + i]); }
    __syncthreads();
    jacobi_eig(A, V, l, cs, pq, red);
    sort_eig_desc(A, V, l, lam, B);                       // B = W
    for (int e = tid; e < ll; e += nt) gW[e] = B[e];
    for (int i = tid; i < l; i += nt) gLam[i] = lam[i];
    __syncthreads();
    if (mode == 0 && gS == nullptr) return;               // Rayleigh-Ritz only: W, lam
    if (mode == 2) {
      for (int e = tid; e < ll; e += nt) {
        const int j = e % l;
        const double sig = sqrt(fmax(lam[j], 0.0));        // singular value
        const double rt = sqrt(sig);
        gC[e] = rt > 0.0 ? B[e] / rt : 0.0;                // Z W sigma^-1 * sigma^1/2
        gC2[e] = B[e] * rt;                                // Q W sigma^1/2
      }
      return;
    }
    // M = W^T S W
    for (int e = tid; e < ll; e += nt) { const int i = e / l, j = e % l; A[e] = 0.5 * (gS[i * l + j] + gS[j * l + i]); }
    __syncthreads();
    matmul_ll(A, B, V, l, false);                          // V = S W
    matmul_ll(B, V, A, l, true);                           // A = W^T S W
    for (int e = tid; e < ll; e += nt) Cm[e] = A[e];       // keep a copy for the fallback
    __syncthreads();
    if (cholesky_upper(A, l, &flag)) {
      upper_inverse(A, V, l);                              // V = R^-1
    } else {
      for (int e = tid; e < ll; e += nt) A[e] = Cm[e];
      __syncthreads();
      inv_sqrt_by_eig(A, Cm, V, l, cs, pq, red);           // V = M^-1/2 factor
    }
    matmul_ll(B, V, A, l, false);                          // C = W * V
    for (int e = tid; e < ll; e += nt) gC[e] = A[e];
    if (tid == 0 && gFlag) *gFlag = flag;
  } else {
    for (int e = tid; e < ll; e += nt) { const int i = e / l, j = e % l; A[e] = 0.5 * (gH[i * l + j] + gH[j * l + i]); }
    for (int e = tid; e < ll; e += nt) Cm[e] = 0.0;
    __syncthreads();
    for (int e = tid; e < ll; e += nt) Cm[e] = A[e];
    __syncthreads();
    if (cholesky_upper(A, l, &flag)) {
      upper_inverse(A, V, l);
    } else {
      for (int e = tid; e < ll; e += nt) A[e] = Cm[e];
      __syncthreads();
      inv_sqrt_by_eig(A, B, V, l, cs, pq, red);
    }
    for (int e = tid; e < ll; e += nt) gC[e] = V[e];
    if (tid == 0 && gFlag) *gFlag = flag;
  }
}

// ---- Out[n][lo] = X[n][l] * C[l][lo], fp64 accumulate -----------------------------------------
// 256 threads as 16 (row groups of 4) x 16 (column groups: j = cg + 16 q, q < QB = ceil(lo / 16)); tile = 64 rows.
template <int QB>
__global__ void __launch_bounds__(256) apply_small_kernel(const float* __restrict__ X, int64_t n, int l, const double* __restrict__ C,
                                                          int ldc, int lo, float* __restrict__ Out, int64_t ld_out) {
  extern __shared__ double sd[];
  double* sC = sd;                               // l * lo
  float* sx = (float*)(sC + (size_t)l * lo);     // 64 * (l + 1)
  const int ls = l + 1;
  for (int e = threadIdx.x; e < l * lo; e += 256) sC[e] = C[(e / lo) * ldc + (e % lo)];
  const int cg = threadIdx.x & 15, rg = threadIdx.x >> 4;
  for (int64_t rb = (int64_t)blockIdx.x * 64; rb < n; rb += (int64_t)gridDim.x * 64) {
    __syncthreads();
    for (int e = threadIdx.x; e < 64 * l; e += 256) {
      const int r = e / l, c = e - r * l;
      sx[r * ls + c] = (rb + r < n) ? X[(rb + r) * l + c] : 0.f;
    }
    __syncthreads();
    double acc[4][QB];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int q = 0; q < QB; ++q) acc[i][q] = 0.0;
    for (int k = 0; k < l; ++k) {
      double xv[4], cv[QB];
#pragma unroll
      for (int i = 0; i < 4; ++i) xv[i] = (double)sx[(rg * 4 + i) * ls + k];
#pragma unroll
      for (int q = 0; q < QB; ++q) cv[q] = (cg + 16 * q < lo) ? sC[k * lo + cg + 16 * q] : 0.0;
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int q = 0; q < QB; ++q) acc[i][q] = fma(xv[i], cv[q], acc[i][q]);
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int64_t row = rb + rg * 4 + i;
      if (row < n) {
#pragma unroll
        for (int q = 0; q < QB; ++q)
          if (cg + 16 * q < lo) Out[row * ld_out + cg + 16 * q] = (float)acc[i][q];
      }
    }
  }
}

// ---- residual column norms^2 of Yw - Qw diag(lam): per-CTA partials, summed in fixed order -----
__global__ void __launch_bounds__(256) resid_kernel(const float* __restrict__ Yw, const float* __restrict__ Qw, int64_t n, int l,
                                                    const double* __restrict__ lam, int rows_per_cta, double* __restrict__ partials) {
  __shared__ double red[256];
  const int64_t r0 = (int64_t)blockIdx.x * rows_per_cta;
  const int64_t r1 = (r0 + rows_per_cta < n) ? r0 + rows_per_cta : n;
  // thread -> column j = tid % l for rows tid / l, tid / l + 256 / l, ... (l divides into 256 unevenly: use a guard)
  const int per = 256 / l;                       // row lanes per pass (>= 3 for l <= 80)
  const int j = threadIdx.x % l, lane_r = threadIdx.x / l;
  double acc = 0.0;
  if (lane_r < per) {
    const double lj = lam[j];
    for (int64_t r = r0 + lane_r; r < r1; r += per) {
      const double d = (double)Yw[r * l + j] - lj * (double)Qw[r * l + j];
      acc = fma(d, d, acc);
    }
  }
  red[threadIdx.x] = (lane_r < per) ? acc : 0.0;
  __syncthreads();
  if (threadIdx.x < l) {
    double sacc = 0.0;
    for (int q = 0; q < per; ++q) sacc += red[q * l + threadIdx.x];
    partials[(size_t)blockIdx.x * l + threadIdx.x] = sacc;
  }
}

// ---- deflation: X[n][l] -= U[n][:p] * G[:p][l]  (U = locked Ritz vectors, G = U^T X, fp64) ------
__global__ void __launch_bounds__(256) sub_proj_kernel(float* __restrict__ X, const float* __restrict__ U, int64_t n, int l,
                                                       int p, const double* __restrict__ G) {
  extern __shared__ double sg[];            // p * l
  float* su = (float*)(sg + (size_t)p * l); // 32 * p
  for (int e = threadIdx.x; e < p * l; e += 256) sg[e] = G[e];
  for (int64_t rb = (int64_t)blockIdx.x * 32; rb < n; rb += (int64_t)gridDim.x * 32) {
    __syncthreads();
    for (int e = threadIdx.x; e < 32 * p; e += 256) {
      const int r = e / p, q = e - r * p;
      su[e] = (rb + r < n) ? U[(rb + r) * l + q] : 0.f;
    }
    __syncthreads();
    for (int e = threadIdx.x; e < 32 * l; e += 256) {
      const int r = e / l, j = e - r * l;
      if (rb + r < n) {
        double s = 0.0;
        for (int q = 0; q < p; ++q) s += (double)su[r * p + q] * sg[q * l + j];
        X[(rb + r) * l + j] = (float)((double)X[(rb + r) * l + j] - s);
      }
    }
  }
}

// ---- split-transpose: X fp32 [n][l] -> Xs bf16 [3][l][n_pad] (K-major operand of the MMA) ----
__global__ void __launch_bounds__(256) split_transpose_kernel(const float* __restrict__ X, int64_t n, int l, int64_t n_pad,
                                                              __nv_bfloat16* __restrict__ Xs) {
  __shared__ float t[32][33];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int64_t rb = (int64_t)blockIdx.x * 32;
  const int cb = blockIdx.y * 32;
  for (int r = ty; r < 32; r += 8) {
    const int64_t i = rb + r; const int j = cb + tx;
    t[r][tx] = (i < n && j < l) ? X[i * l + j] : 0.f;
  }
  __syncthreads();
  for (int r = ty; r < 32; r += 8) {
    const int j = cb + r; const int64_t i = rb + tx;
    if (j < l && i < n_pad) {
      __nv_bfloat16 hi, mid, lo;
      split3(t[tx][r], hi, mid, lo);
      Xs[((size_t)0 * l + j) * n_pad + i] = hi;
      Xs[((size_t)1 * l + j) * n_pad + i] = mid;
      Xs[((size_t)2 * l + j) * n_pad + i] = lo;
    }
  }
}

// Gaussian start block, entries keyed by (seed, row, col) -- identical on every rank
__global__ void random_block_kernel(float* __restrict__ X, int64_t n, int l, uint64_t seed) {
  const int64_t e = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (e >= n * l) return;
  const uint64_t h = hash4(seed, 0x51ull, (uint64_t)(e / l), (uint64_t)(e % l));
  const float u1 = ((float)(uint32_t)(h >> 40) + 1.0f) * (1.0f / 16777216.0f);
  const float u2 = (float)(uint32_t)((h >> 8) & 0xffffffu) * (1.0f / 16777216.0f);
  X[e] = sqrtf(-2.0f * logf(u1)) * cospif(2.0f * u2);
}

// out = a*x + b*y + c*z elementwise (Chebyshev three-term recurrence); y or z may be null
__global__ void axpbypcz_kernel(float* __restrict__ out, const float* __restrict__ x, double a, const float* __restrict__ y,
                                double b, const float* __restrict__ z, double c, int64_t count) {
  const int64_t e = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (e >= count) return;
  double v = a * (double)x[e];
  if (y) v += b * (double)y[e];
  if (z) v += c * (double)z[e];
  out[e] = (float)v;
}

// ---- host wrappers ------------------------------------------------------------------------
size_t smallmat_smem(int l) { return sizeof(double) * (4 * (size_t)l * l + 2 * l + 64) + sizeof(int) * 2 * l + 64; }

int launch_xty(cudaStream_t s, const float* X, const float* Y, int64_t n, int l, double* partials, int max_parts,
               double* out, int accumulate) {
  int parts = (int)((n + 255) / 256);
  if (parts > max_parts) parts = max_parts;
  if (parts < 1) parts = 1;
  int rows_per_cta = (int)((n + parts - 1) / parts);
  rows_per_cta = (rows_per_cta + 31) / 32 * 32;
  parts = (int)((n + rows_per_cta - 1) / rows_per_cta);
  switch ((l + 15) / 16) {
    case 1: xty_kernel<1><<<parts, 256, sizeof(float) * 64 * l, s>>>(X, Y, n, l, rows_per_cta, partials); break;
    case 2: xty_kernel<2><<<parts, 256, sizeof(float) * 64 * l, s>>>(X, Y, n, l, rows_per_cta, partials); break;
    case 3: xty_kernel<3><<<parts, 256, sizeof(float) * 64 * l, s>>>(X, Y, n, l, rows_per_cta, partials); break;
    case 4: xty_kernel<4><<<parts, 256, sizeof(float) * 64 * l, s>>>(X, Y, n, l, rows_per_cta, partials); break;
    default: xty_kernel<5><<<parts, 256, sizeof(float) * 64 * l, s>>>(X, Y, n, l, rows_per_cta, partials); break;
  }
  N2V2R_CHECK_LAUNCH();
  sum_partials_kernel<<<(l * l + 31) / 32, 256, 0, s>>>(partials, parts, l * l, out, accumulate);
  N2V2R_CHECK_LAUNCH();
  return N2V2R_OK;
}

int launch_smallmat(cudaStream_t s, int mode, int l, const double* H, const double* S, double* W, double* lam, double* C,
                    double* C2, int* flag) {
  static bool done = false;
  if (!done) {
    N2V2R_CUDA(cudaFuncSetAttribute(smallmat_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smallmat_smem(80)));
    done = true;
  }
  smallmat_kernel<<<1, 512, smallmat_smem(l), s>>>(mode, l, H, S, W, lam, C, C2, flag);
  N2V2R_CHECK_LAUNCH();
  return N2V2R_OK;
}

int launch_apply_small(cudaStream_t s, const float* X, int64_t n, int l, const double* C, int ldc, int lo, float* Out, int64_t ld_out) {
  static bool done = false;
  const size_t max_smem = sizeof(double) * kMaxL * kMaxL + sizeof(float) * 64 * (kMaxL + 1);
  if (!done) {
    N2V2R_CUDA(cudaFuncSetAttribute(apply_small_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)max_smem));
    N2V2R_CUDA(cudaFuncSetAttribute(apply_small_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)max_smem));
    N2V2R_CUDA(cudaFuncSetAttribute(apply_small_kernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)max_smem));
    N2V2R_CUDA(cudaFuncSetAttribute(apply_small_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)max_smem));
    N2V2R_CUDA(cudaFuncSetAttribute(apply_small_kernel<5>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)max_smem));
    done = true;
  }
  const int64_t tiles = (n + 63) / 64;
  const int grid = (int)(tiles < (int64_t)sm_count() * 4 ? tiles : (int64_t)sm_count() * 4);
  const size_t smem = sizeof(double) * (size_t)l * lo + sizeof(float) * 64 * (l + 1);
  switch ((lo + 15) / 16) {
    case 1: apply_small_kernel<1><<<grid, 256, smem, s>>>(X, n, l, C, ldc, lo, Out, ld_out); break;
    case 2: apply_small_kernel<2><<<grid, 256, smem, s>>>(X, n, l, C, ldc, lo, Out, ld_out); break;
    case 3: apply_small_kernel<3><<<grid, 256, smem, s>>>(X, n, l, C, ldc, lo, Out, ld_out); break;
    case 4: apply_small_kernel<4><<<grid, 256, smem, s>>>(X, n, l, C, ldc, lo, Out, ld_out); break;
    default: apply_small_kernel<5><<<grid, 256, smem, s>>>(X, n, l, C, ldc, lo, Out, ld_out); break;
  }
  N2V2R_CHECK_LAUNCH();
  return N2V2R_OK;
}

// out[j] = || Yw[:, j] - lam[j] Qw[:, j] ||^2
int launch_resid(cudaStream_t s, const float* Yw, const float* Qw, int64_t n, int l, const double* lam, double* partials,
                 int max_parts, double* out) {
  int parts = (int)((n + 255) / 256);
  if (parts > max_parts) parts = max_parts;
  if (parts < 1) parts = 1;
  const int rows_per_cta = (int)((n + parts - 1) / parts);
  parts = (int)((n + rows_per_cta - 1) / rows_per_cta);
  resid_kernel<<<parts, 256, 0, s>>>(Yw, Qw, n, l, lam, rows_per_cta, partials);
  N2V2R_CHECK_LAUNCH();
  sum_partials_kernel<<<(l + 31) / 32, 256, 0, s>>>(partials, parts, l, out, 0);
  N2V2R_CHECK_LAUNCH();
  return N2V2R_OK;
}

int launch_sub_proj(cudaStream_t s, float* X, const float* U, int64_t n, int l, int p, const double* G) {
  if (p <= 0) return N2V2R_OK;
  static bool done = false;
  const size_t max_smem = sizeof(double) * kMaxL * kMaxL + sizeof(float) * 32 * kMaxL;
  if (!done) {
    N2V2R_CUDA(cudaFuncSetAttribute(sub_proj_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)max_smem));
    done = true;
  }
  const int64_t tiles = (n + 31) / 32;
  const int grid = (int)(tiles < (int64_t)sm_count() * 4 ? tiles : (int64_t)sm_count() * 4);
  sub_proj_kernel<<<grid, 256, sizeof(double) * p * l + sizeof(float) * 32 * p, s>>>(X, U, n, l, p, G);
  N2V2R_CHECK_LAUNCH();
  return N2V2R_OK;
}

int launch_sum_partials(cudaStream_t s, const double* partials, int n_parts, int count, double* out) {
  sum_partials_kernel<<<(count + 31) / 32, 256, 0, s>>>(partials, n_parts, count, out, 0);
  N2V2R_CHECK_LAUNCH();
  return N2V2R_OK;
}

int launch_split_transpose(cudaStream_t s, const float* X, int64_t n, int l, int64_t n_pad, __nv_bfloat16* Xs) {
  dim3 grid((unsigned)((n_pad + 31) / 32), (unsigned)((l + 31) / 32));
  split_transpose_kernel<<<grid, 256, 0, s>>>(X, n, l, n_pad, Xs);
  N2V2R_CHECK_LAUNCH();
  return N2V2R_OK;
}

int launch_random_block(cudaStream_t s, float* X, int64_t n, int l, uint64_t seed) {
  const int64_t cnt = n * l;
  random_block_kernel<<<(unsigned)((cnt + 255) / 256), 256, 0, s>>>(X, n, l, seed);
  N2V2R_CHECK_LAUNCH();
  return N2V2R_OK;
}

int launch_axpbypcz(cudaStream_t s, float* out, const float* x, double a, const float* y, double b, const float* z, double c,
                    int64_t count) {
  axpbypcz_kernel<<<(unsigned)((count + 255) / 256), 256, 0, s>>>(out, x, a, y, b, z, c, count);
  N2V2R_CHECK_LAUNCH();
  return N2V2R_OK;
}

}  // namespace n2v2r
