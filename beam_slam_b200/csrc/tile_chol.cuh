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

// ---- panel: potrf of the diagonal tile, inverse of its factor, trsm of one tile below it ----
// One CTA per non-zero tile below the diagonal (at least one CTA). Every CTA factors the 64x64 diagonal tile itself --
// the same arithmetic, so the same bits, and no dependency between CTAs -- then multiplies its own tile:
// L(i,k) = A(i,k) L_kk^-T = A(i,k) W^T with W = L_kk^-1. CTA 0 also stores W (the solve needs it).
constexpr int kPanelSmemDoubles = kTS * (kTS + 1) + 2 * kTS * kTLd + 32 * 33 + 2 * kTS + kTS;
__global__ void __launch_bounds__(256) tile_panel_kernel(TileChol TC, int k, LMState* st) {
  extern __shared__ __align__(16) double sm[];
  double* As = sm;                          // A(i,k), column-major, pitch kTLd
  double* Bs = As + kTS * kTLd;             // W, column-major, pitch kTLd
  double (*A)[kTS + 1] = reinterpret_cast<double (*)[kTS + 1]>(Bs + kTS * kTLd);   // L_kk, A[r][c]
  double (*Tm)[33] = reinterpret_cast<double (*)[33]>(&A[kTS][0]);                 // 32x32 scratch of the blocked inverse
  double (*raw)[kTS] = reinterpret_cast<double (*)[kTS]>(&Tm[32][0]);             // [2][64] column exchange of the potrf
  double* invd = &raw[2][0];
  __shared__ int failed;
  if (!gate_run(st) || st->lin_fail) return;   // lin_fail is written only at the very end (by CTA 0)
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;   // potrf: thread owns rows 4ty.., cols 4tx.. (ty >= tx matters)
  const int na = TC.col_ptr[k + 1] - TC.col_ptr[k];
  double* G = TC.T + TC.tile_off[(long long)k * TC.nT + k];
  double* Gi = nullptr;
  if (na > 0) {   // start fetching this CTA's tile; it is not needed before the product
    Gi = TC.T + TC.tile_off[(long long)TC.col_rows[TC.col_ptr[k] + blockIdx.x] * TC.nT + k];
    for (int e = tid; e < kTileDoubles / 2; e += 256) {
      const int c = e >> 5, r2 = e & 31;
      const unsigned dst = (unsigned)__cvta_generic_to_shared(As + c * kTLd + 2 * r2);
      asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(Gi + 2 * e));
    }
    asm volatile("cp.async.commit_group;");
  }
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
      if (4 * ty + 3 < j) continue;           // all rows of this thread are above the pivot: nothing left to do but the barriers
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
      if (ty >= tx && tx >= jb) {
#pragma unroll
        for (int c = 0; c < 4; ++c)
#pragma unroll
          for (int i = 0; i < 4; ++i) if (4 * tx + c > j && 4 * ty + i >= 4 * tx + c) a[i][c] -= lr[i] * lc[c];
      }
      if (j + 1 < kTS && tx == (j + 1) / 4) {   // raw column j + 1 for the next step
#pragma unroll
        for (int i = 0; i < 4; ++i) raw[p ^ 1][4 * ty + i] = (cj == 3) ? a[i][0] : a[i][(cj + 1) & 3];
      }
    }
  }
  if (bad) failed = 1;
#pragma unroll
  for (int c = 0; c < 4; ++c)
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int r = 4 * ty + i, cc = 4 * tx + c;
      if (r >= cc) { A[r][cc] = a[i][c]; if (r == cc) invd[r] = 1.0 / a[i][c]; }   // L_kk itself is not stored: other CTAs may still be reading A_kk, and only W is used later
    }
  for (int e = tid; e < kTS * kTLd; e += 256) Bs[e] = 0.0;
  __syncthreads();
  // ---- W = L^-1, blocked: 16x16 diagonal blocks by substitution, then two levels of block products ----
  // W(r, c) lives at Bs[c * kTLd + r]
  {
    const int blk = tid >> 6, c = 16 * blk + ((tid >> 2) & 15), sub = tid & 3, r0 = 16 * blk;
    for (int rr = 0; rr < 16; ++rr) {
      const int r = r0 + rr;
      double sum = 0.0;
      if (r > c) for (int m = c + sub; m < r; m += 4) sum += A[r][m] * Bs[c * kTLd + m];
      sum += __shfl_xor_sync(0xffffffffu, sum, 1); sum += __shfl_xor_sync(0xffffffffu, sum, 2);
      if (sub == 0 && r >= c) Bs[c * kTLd + r] = r == c ? invd[r] : -sum * invd[r];
      __syncwarp();
    }
  }
  __syncthreads();
  // level 1: W[b+1][b] = -W[b+1][b+1] (L[b+1][b] W[b][b]) for b = 0, 2 (16x16 blocks)
  {
    const int pr = tid >> 7, b = 2 * pr, o = tid & 127;     // two outputs per thread: (r, c) and (r + 8, c)
    const int r = o >> 4, c = o & 15;
    double t0 = 0.0, t1 = 0.0;
    for (int m = c; m < 16; ++m) { const double w = Bs[(16 * b + c) * kTLd + 16 * b + m]; t0 += A[16 * (b + 1) + r][16 * b + m] * w; t1 += A[16 * (b + 1) + r + 8][16 * b + m] * w; }
    Tm[16 * pr + r][c] = t0; Tm[16 * pr + r + 8][c] = t1;
    __syncthreads();
    double u0 = 0.0, u1 = 0.0;
    for (int m = 0; m <= r + 8; ++m) { const double t = Tm[16 * pr + m][c]; if (m <= r) u0 += Bs[(16 * (b + 1) + m) * kTLd + 16 * (b + 1) + r] * t; u1 += Bs[(16 * (b + 1) + m) * kTLd + 16 * (b + 1) + r + 8] * t; }
    Bs[(16 * b + c) * kTLd + 16 * (b + 1) + r] = -u0; Bs[(16 * b + c) * kTLd + 16 * (b + 1) + r + 8] = -u1;
  }
  __syncthreads();
  // level 2: W[32:64][0:32] = -W22 (L21 W11), 32x32 blocks; four outputs per thread: rows r, r+8, r+16, r+24 of column c
  {
    const int r = tid >> 5, c = tid & 31;
    double t[4] = {0.0, 0.0, 0.0, 0.0};
    for (int m = c; m < 32; ++m) {
      const double w = Bs[c * kTLd + m];
#pragma unroll
      for (int q = 0; q < 4; ++q) t[q] += A[32 + r + 8 * q][m] * w;
    }
#pragma unroll
    for (int q = 0; q < 4; ++q) Tm[r + 8 * q][c] = t[q];
    __syncthreads();
    double u[4] = {0.0, 0.0, 0.0, 0.0};
    for (int m = 0; m < 32; ++m) {
      const double tv = Tm[m][c];
#pragma unroll
      for (int q = 0; q < 4; ++q) if (m <= r + 8 * q) u[q] += Bs[(32 + m) * kTLd + 32 + r + 8 * q] * tv;
    }
#pragma unroll
    for (int q = 0; q < 4; ++q) Bs[c * kTLd + 32 + r + 8 * q] = -u[q];
  }
  if (na > 0) asm volatile("cp.async.wait_group 0;");
  __syncthreads();
  if (blockIdx.x == 0) {
    double* Wk = TC.W + (size_t)k * kTileDoubles;
    for (int e = tid; e < kTileDoubles; e += 256) Wk[e] = Bs[(e >> 6) * kTLd + (e & 63)];
    if (failed && tid == 0) st->lin_fail = 1;
  }
  if (na == 0) return;
  const int warp = tid >> 5, lane = tid & 31, g = lane >> 2, tg = lane & 3;
  double acc[8][2];
  tile_abt(As, Bs, warp, lane, acc);
#pragma unroll
  for (int J = 0; J < 8; ++J) {
    Gi[(8 * J + 2 * tg) * kTS + 8 * warp + g] = acc[J][0];
    Gi[(8 * J + 2 * tg + 1) * kTS + 8 * warp + g] = acc[J][1];
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
// Per tile column the loads of W_k and of the column's tiles are issued together (they do not depend on the running
// solution), so a column costs about one L2 round trip instead of one per tile. kSolveBatch tiles at a time.
constexpr int kSolveBatch = 4;
__global__ void __launch_bounds__(1024) tile_solve_kernel(TileChol TC, const double* __restrict__ rhs, double* __restrict__ yf, const LMState* st) {
  __shared__ double part[kSolveBatch][16][kTS];   // forward: partial products per column group
  __shared__ double xs[kSolveBatch][kTS], ys[kTS];
  if (!gate_run(st) || st->lin_fail) return;
  const int tid = threadIdx.x, nT = TC.nT, n = TC.n;
  double* b = TC.b;
  for (int i = tid; i < nT * kTS; i += 1024) b[i] = i < n ? rhs[i] : 0.0;
  __syncthreads();
  // forward: y_k = W_k b_k, b_i -= L(i,k) y_k.  Thread (cg, r): column group cg = tid >> 6 (4 columns), row r = tid & 63
  {
    const int cg = tid >> 6, r = tid & 63;
    for (int k = 0; k < nT; ++k) {
      const int c0 = TC.col_ptr[k], na = TC.col_ptr[k + 1] - c0;
      const double* Wk = TC.W + (size_t)k * kTileDoubles;
      double w[4], l[kSolveBatch][4];
#pragma unroll
      for (int q = 0; q < 4; ++q) w[q] = Wk[(4 * cg + q) * kTS + r];
#pragma unroll
      for (int t = 0; t < kSolveBatch; ++t) if (t < na) {
        const double* Lt = TC.T + TC.tile_off[(long long)TC.col_rows[c0 + t] * nT + k];
#pragma unroll
        for (int q = 0; q < 4; ++q) l[t][q] = Lt[(4 * cg + q) * kTS + r];
      }
      if (tid < kTS) ys[tid] = b[k * kTS + tid];
      __syncthreads();
      { double s = 0.0;
#pragma unroll
        for (int q = 0; q < 4; ++q) s += w[q] * ys[4 * cg + q];
        part[0][cg][r] = s; }
      __syncthreads();
      if (tid < kTS) { double s = 0.0; for (int q = 0; q < 16; ++q) s += part[0][q][tid]; xs[0][tid] = s; b[k * kTS + tid] = s; }
      __syncthreads();
      for (int t0 = 0; t0 < na; t0 += kSolveBatch) {
        if (t0 > 0) {   // further batches of a tall column
#pragma unroll
          for (int t = 0; t < kSolveBatch; ++t) if (t0 + t < na) {
            const double* Lt = TC.T + TC.tile_off[(long long)TC.col_rows[c0 + t0 + t] * nT + k];
#pragma unroll
            for (int q = 0; q < 4; ++q) l[t][q] = Lt[(4 * cg + q) * kTS + r];
          }
        }
#pragma unroll
        for (int t = 0; t < kSolveBatch; ++t) if (t0 + t < na) {
          double s = 0.0;
#pragma unroll
          for (int q = 0; q < 4; ++q) s += l[t][q] * xs[0][4 * cg + q];
          part[t][cg][r] = s;
        }
        __syncthreads();
        if (tid < kSolveBatch * kTS) {
          const int t = tid >> 6, rr = tid & 63;
          if (t0 + t < na) { double s = 0.0; for (int q = 0; q < 16; ++q) s += part[t][q][rr]; b[TC.col_rows[c0 + t0 + t] * kTS + rr] -= s; }
        }
        __syncthreads();
      }
    }
  }
  // backward: x_k = W_k^T (y_k - sum_i L(i,k)^T x_i).  Thread (c, sub): output c = tid >> 4, 16 lanes split the rows
  {
    const int c = tid >> 4, sub = tid & 15;
    for (int k = nT - 1; k >= 0; --k) {
      const int c0 = TC.col_ptr[k], na = TC.col_ptr[k + 1] - c0;
      const double* Wk = TC.W + (size_t)k * kTileDoubles;
      double w[4];
#pragma unroll
      for (int q = 0; q < 4; ++q) w[q] = Wk[c * kTS + 4 * sub + q];
      double t = sub == 0 ? b[k * kTS + c] : 0.0;
      for (int t0 = 0; t0 < na; t0 += kSolveBatch) {
        double l[kSolveBatch][4];
#pragma unroll
        for (int tt = 0; tt < kSolveBatch; ++tt) if (t0 + tt < na) {
          const double* Lt = TC.T + TC.tile_off[(long long)TC.col_rows[c0 + t0 + tt] * nT + k];
#pragma unroll
          for (int q = 0; q < 4; ++q) l[tt][q] = Lt[c * kTS + 4 * sub + q];
        }
        if (tid < kSolveBatch * kTS) { const int tt = tid >> 6, rr = tid & 63; if (t0 + tt < na) xs[tt][rr] = b[TC.col_rows[c0 + t0 + tt] * kTS + rr]; }
        __syncthreads();
#pragma unroll
        for (int tt = 0; tt < kSolveBatch; ++tt) if (t0 + tt < na) {
#pragma unroll
          for (int q = 0; q < 4; ++q) t -= l[tt][q] * xs[tt][4 * sub + q];
        }
        __syncthreads();
      }
#pragma unroll
      for (int m = 1; m < 16; m <<= 1) t += __shfl_xor_sync(0xffffffffu, t, m);
      if (sub == 0) ys[c] = t;
      __syncthreads();
      double xk = 0.0;
#pragma unroll
      for (int q = 0; q < 4; ++q) xk += w[q] * ys[4 * sub + q];
#pragma unroll
      for (int m = 1; m < 16; m <<= 1) xk += __shfl_xor_sync(0xffffffffu, xk, m);
      if (sub == 0) b[k * kTS + c] = xk;
      __syncthreads();
    }
  }
  for (int i = tid; i < n; i += 1024) yf[i] = b[i];
}

}  // namespace bsnls
