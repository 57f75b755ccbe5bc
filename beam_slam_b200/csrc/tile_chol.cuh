// Tile-sparse FP64 Cholesky of the reduced system (K8 for everything larger than a sliding window).
//
// The reduced camera system of a trajectory is banded once poses are ordered along the trajectory (plus a few long
// rows per loop closure). It is stored as 64x64 tiles, only those that are structurally non-zero after a symbolic
// factorisation at tile granularity (host, set_problem), and factored right-looking one tile column at a time:
//   potrf  diagonal tile (one CTA, shared memory), also L_kk^-1
//   trsm   L(i,k) = A(i,k) L_kk^-T as a tile product with the inverse      (one CTA per non-zero tile below the diagonal)
//   syrk   A(i,j) -= L(i,k) L(j,k)^T                                        (one CTA per pair of non-zero tiles)
// The two products run on the FP64 tensor cores (mma.sync m8n8k4). The launch sequence depends on the structure only,
// so it is fixed per problem. Replaces SuiteSparse/Eigen sparse Cholesky inside ceres' SPARSE_SCHUR.
#pragma once

namespace bsnls {

constexpr int kTS = 64;              // tile edge
constexpr int kTLd = 68;             // shared-memory pitch of a tile (68 mod 16 = 4: conflict-free 64-bit fragment loads)
constexpr int kTileDoubles = kTS * kTS;

struct TileChol {
  int nT, n;
  const long long* tile_off;   // [nT * nT], (i, j) with i >= j at i * nT + j: offset of the tile in T (doubles) or -1
  const int* col_ptr;          // [nT + 1] ranges into col_rows: non-zero tiles strictly below the diagonal of tile column k
  const int* col_rows;
  double* T;                   // tile storage, each tile column-major 64x64
  double* W;                   // [nT][64*64] inverses of the diagonal factors (lower triangular, column-major)
  double* b;                   // [nT * 64] right-hand side / solution scratch
};

// ---- scatter the 3x3-block CSR system into the tiles (lower triangle), identity on the padding ----
__global__ void __launch_bounds__(256) bsr_to_tiles_kernel(DeviceProblem P, TileChol TC, const double* __restrict__ Sb, const LMState* st) {
  if (!gate_run(st)) return;
  const long long e = (long long)blockIdx.x * 256 + threadIdx.x;
  if (e < P.nnzb * 9) {
    const int blk = (int)(e / 9), w = (int)(e - 9LL * blk), i = w / 3, j = w - 3 * i;
    int lo = 0, hi = P.nf;   // row of this block: binary search in rowptr
    while (hi - lo > 1) { const int mid = (lo + hi) >> 1; if (P.bsr_rowptr[mid] <= blk) lo = mid; else hi = mid; }
    const int r = 3 * lo + i, c = 3 * P.bsr_col[blk] + j;
    if (r >= c) {
      const long long off = TC.tile_off[(long long)(r / kTS) * TC.nT + c / kTS];
      TC.T[off + (c % kTS) * kTS + r % kTS] = Sb[e];
    }
  } else {
    const int p = (int)(e - P.nnzb * 9) + TC.n;   // padding rows of the last tile
    if (p < TC.nT * kTS) TC.T[TC.tile_off[(long long)(p / kTS) * TC.nT + p / kTS] + (p % kTS) * kTS + p % kTS] = 1.0;
  }
}

// ---- tile loads / the tensor-core product C = A B^T of two 64x64 tiles held in shared memory ----
__device__ __forceinline__ void tile_load(const double* __restrict__ g, double* __restrict__ s) {
  for (int e = threadIdx.x; e < kTileDoubles / 2; e += 256) {
    const int c = e >> 5, r2 = e & 31;
    reinterpret_cast<double2*>(s + c * kTLd)[r2] = reinterpret_cast<const double2*>(g)[e];
  }
}
// warp `I` (0..7) computes rows 8I..8I+7 of C: acc[J] = the lane's two entries (row 8I+g, cols 8J+2tg, 8J+2tg+1)
__device__ __forceinline__ void tile_abt(const double* __restrict__ As, const double* __restrict__ Bs, int I, int lane, double (&acc)[8][2]) {
  const int g = lane >> 2, tg = lane & 3;
  const double* pa = As + tg * kTLd + 8 * I + g;
  const double* pb = Bs + tg * kTLd + g;
#pragma unroll
  for (int J = 0; J < 8; ++J) { acc[J][0] = 0.0; acc[J][1] = 0.0; }
#pragma unroll 4
  for (int ks = 0; ks < kTS / 4; ++ks) {
    const double av = pa[4 * ks * kTLd];
#pragma unroll
    for (int J = 0; J < 8; ++J) dmma_m8n8k4(acc[J][0], acc[J][1], av, pb[4 * ks * kTLd + 8 * J]);
  }
}

// ---- potrf: Cholesky of the diagonal tile and the inverse of its factor; one CTA ----
__global__ void __launch_bounds__(256) tile_potrf_kernel(TileChol TC, int k, LMState* st) {
  __shared__ double A[kTS][kTS + 1];     // L in the lower triangle; afterwards W^T (W = L^-1) in the strict upper triangle
  __shared__ double raw[2][kTS];
  __shared__ double invd[kTS];
  __shared__ int failed;
  if (!gate_run(st) || st->lin_fail) return;   // single CTA; lin_fail is written only at the very end
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;   // thread owns rows 4ty.., cols 4tx.. (ty >= tx matters)
  double* G = TC.T + TC.tile_off[(long long)k * TC.nT + k];
  if (tid == 0) failed = 0;
  double a[4][4];
#pragma unroll
  for (int c = 0; c < 4; ++c)
#pragma unroll
    for (int i = 0; i < 4; ++i) a[i][c] = G[(4 * tx + c) * kTS + 4 * ty + i];
  if (tx == 0) {
#pragma unroll
    for (int i = 0; i < 4; ++i) raw[0][4 * ty + i] = a[i][0];
  }
  bool bad = false;
#pragma unroll 1
  for (int jb = 0; jb < kTS / 4; ++jb) {
#pragma unroll
    for (int cj = 0; cj < 4; ++cj) {
      const int j = 4 * jb + cj, p = j & 1;
      __syncthreads();
      const double x = raw[p][j];
      bad = bad || !(x > 0.0);
      const double inv = rsqrt(x);
      double lr[4], lc[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) { lr[i] = raw[p][4 * ty + i] * inv; lc[i] = raw[p][4 * tx + i] * inv; }
      if (tx == jb) {
#pragma unroll
        for (int i = 0; i < 4; ++i) { const int r = 4 * ty + i; if (r > j) a[i][cj] = lr[i]; else if (r == j) a[i][cj] = x * inv; }
      }
      if (ty >= tx) {
#pragma unroll
        for (int c = 0; c < 4; ++c)
#pragma unroll
          for (int i = 0; i < 4; ++i) if (4 * tx + c > j && 4 * ty + i >= 4 * tx + c) a[i][c] -= lr[i] * lc[c];
      }
      // raw column j + 1 for the next step
      if (j + 1 < kTS && tx == (j + 1) / 4) {
#pragma unroll
        for (int i = 0; i < 4; ++i) raw[p ^ 1][4 * ty + i] = (cj == 3) ? a[i][0] : a[i][(cj + 1) & 3];
      }
    }
  }
  if (bad) failed = 1;
  __syncthreads();
#pragma unroll
  for (int c = 0; c < 4; ++c)
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int r = 4 * ty + i, cc = 4 * tx + c;
      if (r >= cc) { A[r][cc] = a[i][c]; G[cc * kTS + r] = a[i][c]; if (r == cc) invd[r] = 1.0 / a[i][c]; }
    }
  __syncthreads();
  // W = L^-1 by forward substitution, 4 lanes per column; W[m][c] (m > c) is kept at A[c][m]
  {
    const int c = tid >> 2, sub = tid & 3;
    for (int r = 1; r < kTS; ++r) {
      double sum = 0.0;
      if (r > c) {
        for (int m = c + sub; m < r; m += 4) sum += A[r][m] * (m == c ? invd[c] : A[c][m]);
      }
      sum += __shfl_xor_sync(0xffffffffu, sum, 1); sum += __shfl_xor_sync(0xffffffffu, sum, 2);
      if (r > c && sub == 0) A[c][r] = -sum * invd[r];
      __syncwarp();
    }
  }
  __syncthreads();
  double* Wk = TC.W + (size_t)k * kTileDoubles;
  for (int e = tid; e < kTileDoubles; e += 256) {
    const int cc = e >> 6, r = e & 63;
    Wk[e] = r > cc ? A[cc][r] : (r == cc ? invd[r] : 0.0);
  }
  if (failed && tid == 0) st->lin_fail = 1;
}

// ---- trsm: L(i,k) = A(i,k) W_k^T, one CTA per non-zero tile below the diagonal ----
__global__ void __launch_bounds__(256) tile_trsm_kernel(TileChol TC, int k, const LMState* st) {
  extern __shared__ __align__(16) double sm[];
  if (!gate_run(st) || st->lin_fail) return;
  double* As = sm; double* Bs = sm + kTS * kTLd;
  const int i = TC.col_rows[TC.col_ptr[k] + blockIdx.x];
  double* G = TC.T + TC.tile_off[(long long)i * TC.nT + k];
  tile_load(G, As); tile_load(TC.W + (size_t)k * kTileDoubles, Bs);
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, tg = lane & 3;
  double acc[8][2];
  tile_abt(As, Bs, warp, lane, acc);
#pragma unroll
  for (int J = 0; J < 8; ++J) {
    G[(8 * J + 2 * tg) * kTS + 8 * warp + g] = acc[J][0];
    G[(8 * J + 2 * tg + 1) * kTS + 8 * warp + g] = acc[J][1];
  }
}

// ---- syrk: A(i,j) -= L(i,k) L(j,k)^T for every pair i >= j of non-zero tiles of column k ----
__global__ void __launch_bounds__(256) tile_syrk_kernel(TileChol TC, int k, const LMState* st) {
  extern __shared__ __align__(16) double sm[];
  if (!gate_run(st) || st->lin_fail) return;
  double* As = sm; double* Bs = sm + kTS * kTLd;
  // triangular decode of the pair index: p = x (x + 1) / 2 + y, x >= y
  const int p = blockIdx.x;
  int x = (int)((sqrt(8.0 * p + 1.0) - 1.0) * 0.5);
  while (x * (x + 1) / 2 > p) --x;
  while ((x + 1) * (x + 2) / 2 <= p) ++x;
  const int y = p - x * (x + 1) / 2;
  const int base = TC.col_ptr[k], i = TC.col_rows[base + x], j = TC.col_rows[base + y];
  tile_load(TC.T + TC.tile_off[(long long)i * TC.nT + k], As);
  if (i != j) tile_load(TC.T + TC.tile_off[(long long)j * TC.nT + k], Bs);
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, tg = lane & 3;
  double acc[8][2];
  tile_abt(As, i != j ? Bs : As, warp, lane, acc);
  double* C = TC.T + TC.tile_off[(long long)i * TC.nT + j];
#pragma unroll
  for (int J = 0; J < 8; ++J) {
    C[(8 * J + 2 * tg) * kTS + 8 * warp + g] -= acc[J][0];
    C[(8 * J + 2 * tg + 1) * kTS + 8 * warp + g] -= acc[J][1];
  }
}

// ---- triangular solves with the tile factor; one CTA, 1024 threads ----
// y = M x for a column-major 64x64 tile: 16 column groups x 64 rows, partials combined through shared memory
__device__ __forceinline__ void tile_matvec(const double* __restrict__ M, const double* __restrict__ xs, double (*part)[kTS], int tid) {
  const int cg = tid >> 6, r = tid & 63;
  double s = 0.0;
#pragma unroll
  for (int q = 0; q < 4; ++q) s += M[(4 * cg + q) * kTS + r] * xs[4 * cg + q];
  part[cg][r] = s;
}
// y = M^T x: output c = tid >> 4, 16 lanes split the rows, shuffle reduction; lanes with sub == 0 hold the result
__device__ __forceinline__ double tile_matvec_t(const double* __restrict__ M, const double* __restrict__ xs, int tid) {
  const int c = tid >> 4, sub = tid & 15;
  double s = 0.0;
#pragma unroll
  for (int q = 0; q < 4; ++q) s += M[c * kTS + 4 * sub + q] * xs[4 * sub + q];
#pragma unroll
  for (int m = 1; m < 16; m <<= 1) s += __shfl_xor_sync(0xffffffffu, s, m);
  return s;
}
__global__ void __launch_bounds__(1024) tile_solve_kernel(TileChol TC, const double* __restrict__ rhs, double* __restrict__ yf, const LMState* st) {
  __shared__ double part[16][kTS];
  __shared__ double xs[kTS], ys[kTS];
  if (!gate_run(st) || st->lin_fail) return;
  const int tid = threadIdx.x, nT = TC.nT, n = TC.n;
  double* b = TC.b;
  for (int i = tid; i < nT * kTS; i += 1024) b[i] = i < n ? rhs[i] : 0.0;
  __syncthreads();
  // forward: y_k = W_k b_k, b_i -= L(i,k) y_k
  for (int k = 0; k < nT; ++k) {
    if (tid < kTS) xs[tid] = b[k * kTS + tid];
    __syncthreads();
    tile_matvec(TC.W + (size_t)k * kTileDoubles, xs, part, tid);
    __syncthreads();
    if (tid < kTS) { double s = 0.0; for (int q = 0; q < 16; ++q) s += part[q][tid]; ys[tid] = s; b[k * kTS + tid] = s; }
    __syncthreads();
    for (int it = TC.col_ptr[k]; it < TC.col_ptr[k + 1]; ++it) {
      const int i = TC.col_rows[it];
      tile_matvec(TC.T + TC.tile_off[(long long)i * nT + k], ys, part, tid);
      __syncthreads();
      if (tid < kTS) { double s = 0.0; for (int q = 0; q < 16; ++q) s += part[q][tid]; b[i * kTS + tid] -= s; }
      __syncthreads();
    }
  }
  // backward: x_k = W_k^T (y_k - sum_i L(i,k)^T x_i)
  for (int k = nT - 1; k >= 0; --k) {
    double t = (tid & 15) == 0 ? b[k * kTS + (tid >> 4)] : 0.0;
    for (int it = TC.col_ptr[k]; it < TC.col_ptr[k + 1]; ++it) {
      const int i = TC.col_rows[it];
      if (tid < kTS) xs[tid] = b[i * kTS + tid];
      __syncthreads();
      const double s = tile_matvec_t(TC.T + TC.tile_off[(long long)i * nT + k], xs, tid);
      t -= s;
      __syncthreads();
    }
    if ((tid & 15) == 0) ys[tid >> 4] = t;
    __syncthreads();
    const double xk = tile_matvec_t(TC.W + (size_t)k * kTileDoubles, ys, tid);
    if ((tid & 15) == 0) b[k * kTS + (tid >> 4)] = xk;
    __syncthreads();
  }
  for (int i = tid; i < n; i += 1024) yf[i] = b[i];
}

}  // namespace bsnls
