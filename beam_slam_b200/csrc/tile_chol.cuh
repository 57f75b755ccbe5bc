// Tile-sparse FP64 Cholesky of the reduced system (K8 for everything larger than a sliding window).
//
// The reduced camera system of a trajectory is banded once poses are ordered along the trajectory (plus a few long
// rows per loop closure). It is stored as 64x64 tiles (21 f-blocks = 63 unknowns + one identity pad row each), only
// those that are structurally non-zero after a symbolic factorisation at tile granularity (host, set_problem).
// The block order is a nested dissection of the block graph; every leaf and separator starts on a tile boundary (unused
// slots of its last tile are identity padding), so tile columns on one level of the elimination tree are independent;
// the right-looking factorisation advances one LEVEL per launch pair:
//   panel  L(i,k) = A(i,k) W^T with W = L_kk^-1 as a tile product (one CTA per non-zero tile below the diagonal of
//          every column of the level)
//   syrk   A(i,j) -= sum_k L(i,k) L(j,k)^T over the level's columns k (one CTA per destination tile, fixed k order);
//          a diagonal tile that this update completes is factored on the spot (potrf together with W = L^-1)
// and so do the two substitutions. The tile products run on the FP64 tensor cores (mma.sync m8n8k4). The launch
// sequence depends on the structure only, so it is fixed per problem. Replaces SuiteSparse/Eigen sparse Cholesky inside
// ceres' SPARSE_SCHUR.
#pragma once

namespace bsnls {

constexpr int kTS = 64;              // tile edge
constexpr int kTileBlocks = 21;      // f-blocks per tile
constexpr int kTU = 3 * kTileBlocks; // unknowns per tile (63); row/column 63 of every tile is an identity pad
constexpr int kTLd = 68;             // shared-memory pitch of a tile (68 mod 16 = 4: conflict-free 64-bit fragment loads)
constexpr int kTileDoubles = kTS * kTS;
constexpr int kNdLeaf = 4, kNdShort = 6;   // nested dissection: tiles per leaf, longest edge (in tiles) that counts as band

struct TileChol {
  int nT, n;
  const int* blk_slot;         // [nf] slot of an f-block: tile = slot / kTileBlocks, position in the tile = slot % kTileBlocks
  const int* slot_blk;         // [nT * kTileBlocks] the inverse; -1: padding (identity)
  const long long* tile_off;   // [nT * nT], (i, j) with i >= j at i * nT + j: offset of the tile in T (doubles) or -1
  const int* col_ptr;          // [nT + 1] ranges into col_rows: non-zero tiles strictly below the diagonal of tile column k
  const int* col_rows;
  double* T;                   // tile storage, each tile column-major 64x64
  double* W;                   // [nT][64*64] inverses of the diagonal factors (lower triangular, column-major)
  double* b;                   // [nT * 64] right-hand side / solution scratch
  // level schedule (see set_problem)
  const int* lev_cols;         // tile columns grouped by level
  const int* pan_k; const int* pan_x;   // panel CTAs: column, index into col_rows (-1: the column has no tile below the diagonal)
  const int* upd_ptr; const int* upd_fac; const long long* upd_dst; const long long* upd_a; const long long* upd_b;   // update destinations and their source tile pairs
  const int* fw_ptr; const int* fw_row; const int* fw_tail; const long long* fw_src; const int* fw_k;   // forward substitution
  const int* bw_ptr; const int* bw_col; const int* bw_tail; const long long* bw_src; const int* bw_i; const int* roots;   // backward substitution
};

// ---- scatter the 3x3-block CSR system into the tiles (lower triangle), identity on the padding ----
__global__ void __launch_bounds__(256) bsr_to_tiles_kernel(DeviceProblem P, TileChol TC, const double* __restrict__ Sb, const LMState* st) {
  if (!gate_run(st)) return;
  const long long e = (long long)blockIdx.x * 256 + threadIdx.x;
  if (e < P.nnzb * 9) {
    const int blk = (int)(e / 9), w = (int)(e - 9LL * blk), i = w / 3, j = w - 3 * i;
    int lo = 0, hi = P.nf;   // row of this block: binary search in rowptr
    while (hi - lo > 1) { const int mid = (lo + hi) >> 1; if (P.bsr_rowptr[mid] <= blk) lo = mid; else hi = mid; }
    const int br = lo, bc = P.bsr_col[blk];
    if (3 * br + i >= 3 * bc + j) {
      const int sr = TC.blk_slot[br], sc = TC.blk_slot[bc], tr = sr / kTileBlocks, tc = sc / kTileBlocks;
      const long long off = TC.tile_off[(long long)tr * TC.nT + tc];
      TC.T[off + (3 * (sc - tc * kTileBlocks) + j) * kTS + 3 * (sr - tr * kTileBlocks) + i] = Sb[e];
    }
  } else {
    // identity on the pad row of every tile and on the rows of its unused slots: row l of tile t
    const long long q = e - P.nnzb * 9;
    const int t = (int)(q / kTS), l = (int)(q % kTS);
    if (t < TC.nT && (l == kTU || TC.slot_blk[t * kTileBlocks + l / 3] < 0)) TC.T[TC.tile_off[(long long)t * TC.nT + t] + l * kTS + l] = 1.0;
  }
}

// ---- tile loads / the tensor-core product C = A B^T of two 64x64 tiles held in shared memory ----
__device__ __forceinline__ void tile_load(const double* __restrict__ g, double* __restrict__ s) {
  double2 v[8];   // all eight loads of a thread in flight before the first store
#pragma unroll
  for (int u = 0; u < 8; ++u) v[u] = reinterpret_cast<const double2*>(g)[threadIdx.x + 256 * u];
#pragma unroll
  for (int u = 0; u < 8; ++u) { const int e = threadIdx.x + 256 * u; reinterpret_cast<double2*>(s + (e >> 5) * kTLd)[e & 31] = v[u]; }
}
__device__ __forceinline__ void tile_load2(const double* __restrict__ g0, double* __restrict__ s0, const double* __restrict__ g1, double* __restrict__ s1) {
  double2 v[16];
#pragma unroll
  for (int u = 0; u < 8; ++u) { v[u] = reinterpret_cast<const double2*>(g0)[threadIdx.x + 256 * u]; v[8 + u] = reinterpret_cast<const double2*>(g1)[threadIdx.x + 256 * u]; }
#pragma unroll
  for (int u = 0; u < 8; ++u) { const int e = threadIdx.x + 256 * u; reinterpret_cast<double2*>(s0 + (e >> 5) * kTLd)[e & 31] = v[u]; reinterpret_cast<double2*>(s1 + (e >> 5) * kTLd)[e & 31] = v[8 + u]; }
}
// warp `I` (0..7) accumulates rows 8I..8I+7 of C into acc: acc[J] = the lane's two entries (row 8I+g, cols 8J+2tg, 8J+2tg+1)
__device__ __forceinline__ void tile_abt(const double* __restrict__ As, const double* __restrict__ Bs, int I, int lane, double (&acc)[8][2]) {
  const int g = lane >> 2, tg = lane & 3;
  const double* pa = As + tg * kTLd + 8 * I + g;
  const double* pb = Bs + tg * kTLd + g;
#pragma unroll 4
  for (int ks = 0; ks < kTS / 4; ++ks) {
    const double av = pa[4 * ks * kTLd];
#pragma unroll
    for (int J = 0; J < 8; ++J) dmma_m8n8k4(acc[J][0], acc[J][1], av, pb[4 * ks * kTLd + 8 * J]);
  }
}

// ---- potrf of a diagonal tile together with the inverse of its factor ----
// All 256 threads of a CTA. The factorisation is right-looking on 4x4 register blocks (thread (ty, tx) owns rows 4ty..,
// columns 4tx..), four pivots per step: the owner of the diagonal block factors it in registers and publishes the inverse
// of that 4x4 factor; the threads of the block column multiply their blocks by it (the panel), and the rest of the matrix
// gets a rank-4 update. The same row operations are applied to an identity, which turns it into W = L_kk^-1 in the same
// sweep: blocks right of the pivot column update A entries, blocks left of it update W entries. Two barriers per four
// pivots. G: the tile (lower triangle), column-major with pitch ld, global or shared memory. W goes to Ws (shared, pitch
// kTLd) and / or Wg (global, pitch kTS) when given. Returns true in the thread that met a non-positive pivot.
constexpr int kFactorScratchDoubles = 16 + kTS * 4 + 4 * kTS;
constexpr int kPanelSmemDoubles = 2 * kTS * kTLd + kFactorScratchDoubles;
#ifdef PANEL_PROF
__device__ long long g_panel_prof[80];
#endif
__device__ __forceinline__ bool tile_potrf_inv(const double* __restrict__ G, int ld, double* __restrict__ Ws, double* __restrict__ Wg, double* __restrict__ scratch) {
  double* sLi = scratch;                    // inverse of the 4x4 diagonal factor, row-major (lower)
  double* sPan = sLi + 16;                  // the panel: L(r, 4 jb + c) at sPan[((r & 3) * 16 + (r >> 2)) * 4 + c]: the rows 4 ty + i of
                                            // the 16 block rows of a warp are adjacent (no bank conflicts)
  double* sWrow = sPan + kTS * 4;           // final rows 4 jb.. of W: sWrow[i * 64 + c]
  // a warp holds two block columns (tx) and all block rows (ty): at any step it is either left of the pivot column (W
  // updates only) or right of it (A updates only), so the two roles do not serialise inside a warp
  const int tid = threadIdx.x, tx = tid >> 4, ty = tid & 15;
  double a[4][4], w[4][4];
#pragma unroll
  for (int c = 0; c < 4; ++c)
#pragma unroll
    for (int i = 0; i < 4; ++i) { a[i][c] = ty >= tx ? G[(4 * tx + c) * ld + 4 * ty + i] : 0.0; w[i][c] = (ty == tx && i == c) ? 1.0 : 0.0; }
  bool bad = false;
#pragma unroll 1
  for (int jb = 0; jb < kTS / 4; ++jb) {
    if (ty == jb && tx == jb) {   // 4x4 Cholesky of the diagonal block and the inverse of its factor
#ifdef PANEL_PROF
      const long long ta = clock64();
#endif
      const double r0 = rsqrt(a[0][0]);
      const double l10 = a[1][0] * r0, l20 = a[2][0] * r0, l30 = a[3][0] * r0;
      const double d1 = a[1][1] - l10 * l10, r1 = rsqrt(d1);
      const double l21 = (a[2][1] - l20 * l10) * r1, l31 = (a[3][1] - l30 * l10) * r1;
      const double d2 = a[2][2] - l20 * l20 - l21 * l21, r2 = rsqrt(d2);
      const double l32 = (a[3][2] - l30 * l20 - l31 * l21) * r2;
      const double d3 = a[3][3] - l30 * l30 - l31 * l31 - l32 * l32, r3 = rsqrt(d3);
      bad = bad || !(a[0][0] > 0.0) || !(d1 > 0.0) || !(d2 > 0.0) || !(d3 > 0.0);
      const double i10 = -l10 * r0 * r1, i21 = -l21 * r1 * r2, i32 = -l32 * r2 * r3;
      const double i20 = -(l20 * r0 + l21 * i10) * r2, i31 = -(l31 * r1 + l32 * i21) * r3;
      const double i30 = -(l30 * r0 + l31 * i10 + l32 * i20) * r3;
      sLi[0] = r0; sLi[1] = 0.0; sLi[2] = 0.0; sLi[3] = 0.0;
      sLi[4] = i10; sLi[5] = r1; sLi[6] = 0.0; sLi[7] = 0.0;
      sLi[8] = i20; sLi[9] = i21; sLi[10] = r2; sLi[11] = 0.0;
      sLi[12] = i30; sLi[13] = i31; sLi[14] = i32; sLi[15] = r3;
#ifdef PANEL_PROF
      if (jb == 5) g_panel_prof[71] = clock64() - ta;
#endif
    }
    __syncthreads();
#ifdef PANEL_PROF
    if (tid == 255) g_panel_prof[4 * jb] = clock64();
#endif
    if (ty >= jb && (tx == jb || (ty == jb && tx < jb))) {
      double Li[4][4];
#pragma unroll
      for (int i = 0; i < 4; ++i) { const double2 u0 = reinterpret_cast<const double2*>(sLi)[2 * i], u1 = reinterpret_cast<const double2*>(sLi)[2 * i + 1]; Li[i][0] = u0.x; Li[i][1] = u0.y; Li[i][2] = u1.x; Li[i][3] = u1.y; }
      if (tx == jb && ty > jb) {   // panel block: L = A Li^T
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          double l[4];
#pragma unroll
          for (int c = 0; c < 4; ++c) { double sacc = 0.0;
#pragma unroll
            for (int m = 0; m <= c; ++m) sacc += a[i][m] * Li[c][m];
            l[c] = sacc; }
#pragma unroll
          for (int c = 0; c < 4; ++c) a[i][c] = l[c];
          reinterpret_cast<double2*>(sPan)[2 * (16 * i + ty)] = make_double2(l[0], l[1]);
          reinterpret_cast<double2*>(sPan)[2 * (16 * i + ty) + 1] = make_double2(l[2], l[3]);
        }
      }
      if (ty == jb) {              // rows 4 jb.. of W are final: W = Li W (block columns right of the pivot hold zeros)
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          double f[4];
#pragma unroll
          for (int i = 0; i < 4; ++i) { double sacc = 0.0;
#pragma unroll
            for (int m = 0; m <= i; ++m) sacc += Li[i][m] * w[m][c];
            f[i] = sacc; }
#pragma unroll
          for (int i = 0; i < 4; ++i) { w[i][c] = f[i]; sWrow[i * kTS + 4 * tx + c] = f[i]; }
        }
      }
    }
#ifdef PANEL_PROF
    if (tid == 255 && jb == 5) g_panel_prof[72] = clock64();
#endif
    __syncthreads();
#ifdef PANEL_PROF
    if (tid == 255 && jb == 5) g_panel_prof[73] = clock64();
#endif
    if (ty > jb && ty >= tx) {    // rank-4 update with the panel rows of this thread
      double Pr[4][4];
#pragma unroll
      for (int i = 0; i < 4; ++i) { const double2 u0 = reinterpret_cast<const double2*>(sPan)[2 * (16 * i + ty)], u1 = reinterpret_cast<const double2*>(sPan)[2 * (16 * i + ty) + 1]; Pr[i][0] = u0.x; Pr[i][1] = u0.y; Pr[i][2] = u1.x; Pr[i][3] = u1.y; }
      if (tx > jb) {
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          const double2 u0 = reinterpret_cast<const double2*>(sPan)[2 * (16 * c + tx)], u1 = reinterpret_cast<const double2*>(sPan)[2 * (16 * c + tx) + 1];
#pragma unroll
          for (int i = 0; i < 4; ++i) a[i][c] -= Pr[i][0] * u0.x + Pr[i][1] * u0.y + Pr[i][2] * u1.x + Pr[i][3] * u1.y;
        }
      } else {
#pragma unroll
        for (int m = 0; m < 4; ++m) {
          const double2 u0 = reinterpret_cast<const double2*>(sWrow + m * kTS)[2 * tx], u1 = reinterpret_cast<const double2*>(sWrow + m * kTS)[2 * tx + 1];
#pragma unroll
          for (int i = 0; i < 4; ++i) { w[i][0] -= Pr[i][m] * u0.x; w[i][1] -= Pr[i][m] * u0.y; w[i][2] -= Pr[i][m] * u1.x; w[i][3] -= Pr[i][m] * u1.y; }
        }
      }
    }
#ifdef PANEL_PROF
    if (tid == 255 && jb == 5) g_panel_prof[74] = clock64();
#endif
  }
#ifdef PANEL_PROF
  if (tid == 255) g_panel_prof[64] = clock64();
#endif
#pragma unroll
  for (int c = 0; c < 4; ++c)
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const double v = (ty > tx || (ty == tx && i >= c)) ? w[i][c] : 0.0;
      if (Ws) Ws[(4 * tx + c) * kTLd + 4 * ty + i] = v;
      if (Wg) Wg[(4 * tx + c) * kTS + 4 * ty + i] = v;
    }
  return bad;
}

// ---- panel: L(i,k) = A(i,k) L_kk^-T = A(i,k) W^T with W = L_kk^-1, one CTA per non-zero tile below the diagonal ----
// factor_here (the first level, whose diagonal tiles nobody updates): every CTA of a column factors the diagonal tile itself
// -- the same arithmetic, so the same bits, and no dependency between CTAs -- and the first one stores W (the solve needs
// it). On every other level W was left by the update kernel of the level before (see tile_syrk_kernel) and is only loaded.
__global__ void __launch_bounds__(256) tile_panel_kernel(TileChol TC, int pan_begin, int factor_here, LMState* st) {
  extern __shared__ __align__(16) double sm[];
  double* As = sm;                          // A(i,k), column-major, pitch kTLd
  double* Bs = As + kTS * kTLd;             // W, column-major, pitch kTLd
  __shared__ int failed, sGate;
  // another CTA of this launch may set lin_fail while this one starts: latch the gate once per CTA so that all of its
  // warps take the same branch (a failing column only ever stores garbage that nobody uses)
  if (threadIdx.x == 0) { sGate = (!gate_run(st) || st->lin_fail) ? 1 : 0; failed = 0; }
  __syncthreads();
  if (sGate) return;
  const int tid = threadIdx.x;
  const int k = TC.pan_k[pan_begin + blockIdx.x], px = TC.pan_x[pan_begin + blockIdx.x];
  const int na = px >= 0 ? 1 : 0;
  if (na == 0 && !factor_here) return;
  const bool first = px < 0 || px == TC.col_ptr[k];   // this CTA stores W of its column
  double* Gi = nullptr;
  if (na > 0) {   // start fetching this CTA's tile; it is not needed before the product
    Gi = TC.T + TC.tile_off[(long long)TC.col_rows[px] * TC.nT + k];
    const double* Wk = TC.W + (size_t)k * kTileDoubles;
    for (int e = tid; e < kTileDoubles / 2; e += 256) {
      const int c = e >> 5, r2 = e & 31;
      const unsigned dst = (unsigned)__cvta_generic_to_shared(As + c * kTLd + 2 * r2);
      asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(Gi + 2 * e));
      if (!factor_here) {
        const unsigned dw = (unsigned)__cvta_generic_to_shared(Bs + c * kTLd + 2 * r2);
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dw), "l"(Wk + 2 * e));
      }
    }
    asm volatile("cp.async.commit_group;");
  }
#ifdef PANEL_PROF
  if (tid == 255) g_panel_prof[70] = clock64();
#endif
  if (factor_here) {
    const bool bad = tile_potrf_inv(TC.T + TC.tile_off[(long long)k * TC.nT + k], kTS, Bs, first ? TC.W + (size_t)k * kTileDoubles : nullptr, Bs + kTS * kTLd);
    if (bad) failed = 1;
  }
  if (na > 0) asm volatile("cp.async.wait_group 0;");
  __syncthreads();
  if (tid == 0 && failed) st->lin_fail = 1;   // every CTA of a column sees the same pivots
#ifdef PANEL_PROF
  if (tid == 255) g_panel_prof[65] = clock64();
#endif
  if (na == 0) return;
  const int warp = tid >> 5, lane = tid & 31, g = lane >> 2, tg = lane & 3;
  double acc[8][2];
#pragma unroll
  for (int J = 0; J < 8; ++J) { acc[J][0] = 0.0; acc[J][1] = 0.0; }
  tile_abt(As, Bs, warp, lane, acc);
#pragma unroll
  for (int J = 0; J < 8; ++J) {
    Gi[(8 * J + 2 * tg) * kTS + 8 * warp + g] = acc[J][0];
    Gi[(8 * J + 2 * tg + 1) * kTS + 8 * warp + g] = acc[J][1];
  }
}

// ---- syrk: A(i,j) -= sum_k L(i,k) L(j,k)^T over the source columns of this level; one CTA per destination tile ----
// A destination that is the diagonal tile of a column of the NEXT level is final after this update (a column's level is
// one above its last child's), so its CTA goes straight on to factor it and leaves W = L_jj^-1 for the next panel launch:
// the 64-pivot chain then runs beside the other updates of the level instead of in front of the next one. Those
// destinations come first in the level's list.
__global__ void __launch_bounds__(256) tile_syrk_kernel(TileChol TC, int upd_begin, LMState* st) {
  extern __shared__ __align__(16) double sm[];
  __shared__ int failed, sGate;
  if (threadIdx.x == 0) { sGate = (!gate_run(st) || st->lin_fail) ? 1 : 0; failed = 0; }
  __syncthreads();
  if (sGate) return;
  double* As = sm; double* Bs = sm + kTS * kTLd;
  const int d = upd_begin + blockIdx.x;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, tg = lane & 3;
  double* C = TC.T + TC.upd_dst[d];
  const int fac = TC.upd_fac[d];        // tile column to factor afterwards, or -1
  double cv[8][2];                      // the destination entries, fetched alongside the operands
#pragma unroll
  for (int J = 0; J < 8; ++J) { cv[J][0] = C[(8 * J + 2 * tg) * kTS + 8 * warp + g]; cv[J][1] = C[(8 * J + 2 * tg + 1) * kTS + 8 * warp + g]; }
  double acc[8][2];
#pragma unroll
  for (int J = 0; J < 8; ++J) { acc[J][0] = 0.0; acc[J][1] = 0.0; }
  for (int it = TC.upd_ptr[d]; it < TC.upd_ptr[d + 1]; ++it) {
    const long long oa = TC.upd_a[it], ob = TC.upd_b[it];
    if (oa != ob) tile_load2(TC.T + oa, As, TC.T + ob, Bs); else tile_load(TC.T + oa, As);
    __syncthreads();
    tile_abt(As, oa != ob ? Bs : As, warp, lane, acc);
    __syncthreads();
  }
  if (fac < 0) {
#pragma unroll
    for (int J = 0; J < 8; ++J) {
      C[(8 * J + 2 * tg) * kTS + 8 * warp + g] = cv[J][0] - acc[J][0];
      C[(8 * J + 2 * tg + 1) * kTS + 8 * warp + g] = cv[J][1] - acc[J][1];
    }
    return;
  }
  // the final diagonal tile stays on chip: factor it, keep only W (nothing reads L_jj itself)
#pragma unroll
  for (int J = 0; J < 8; ++J) {
    As[(8 * J + 2 * tg) * kTLd + 8 * warp + g] = cv[J][0] - acc[J][0];
    As[(8 * J + 2 * tg + 1) * kTLd + 8 * warp + g] = cv[J][1] - acc[J][1];
  }
  __syncthreads();
  if (tile_potrf_inv(As, kTLd, nullptr, TC.W + (size_t)fac * kTileDoubles, Bs + kTS * kTLd)) failed = 1;
  __syncthreads();
  if (threadIdx.x == 0 && failed) st->lin_fail = 1;
}

// ---- substitutions with the tile factor, level by level; 256-thread CTAs ----
// y = M x for a column-major 64x64 tile (x in shared memory): thread (cg, r) takes 16 columns of row r, the four column
// groups are combined through shared memory. Returns the result for row `tid` in threads 0..63.
__device__ __forceinline__ double tile_mv(const double* __restrict__ M, const double* xs, double (*part)[kTS], int tid) {
  const int cg = tid >> 6, r = tid & 63;
  double v[16];
#pragma unroll
  for (int q = 0; q < 16; ++q) v[q] = M[(16 * cg + q) * kTS + r];
  double s = 0.0;
#pragma unroll
  for (int q = 0; q < 16; ++q) s += v[q] * xs[16 * cg + q];
  part[cg][r] = s;
  __syncthreads();
  double tot = 0.0;
  if (tid < kTS) tot = part[0][tid] + part[1][tid] + part[2][tid] + part[3][tid];
  __syncthreads();
  return tot;
}
// y = M^T x: output c = tid >> 2, four lanes split the rows (interleaved 16-byte pieces, so the four lanes of an output
// read 64 contiguous bytes per instruction), shuffle reduction; every lane of the group holds the result
__device__ __forceinline__ double tile_mv_t(const double* __restrict__ M, const double* xs, int tid) {
  const int c = tid >> 2, sub = tid & 3;
  double2 v[8];
#pragma unroll
  for (int q = 0; q < 8; ++q) v[q] = reinterpret_cast<const double2*>(M + c * kTS)[4 * q + sub];
  double s = 0.0;
#pragma unroll
  for (int q = 0; q < 8; ++q) s += v[q].x * xs[8 * q + 2 * sub] + v[q].y * xs[8 * q + 2 * sub + 1];
  s += __shfl_xor_sync(0xffffffffu, s, 1); s += __shfl_xor_sync(0xffffffffu, s, 2);
  return s;
}
__global__ void __launch_bounds__(256) tile_rhs_kernel(TileChol TC, const double* __restrict__ rhs, const LMState* st) {
  if (!gate_run(st) || st->lin_fail) return;
  const int e = blockIdx.x * 256 + threadIdx.x;
  if (e >= TC.nT * kTS) return;
  const int t = e >> 6, l = e & 63, blk = l < kTU ? TC.slot_blk[t * kTileBlocks + l / 3] : -1;
  TC.b[e] = blk >= 0 ? rhs[3 * blk + l % 3] : 0.0;
}
__global__ void __launch_bounds__(256) tile_out_kernel(TileChol TC, double* __restrict__ yf, const LMState* st) {
  if (!gate_run(st) || st->lin_fail) return;
  const int i = blockIdx.x * 256 + threadIdx.x;
  if (i < TC.n) { const int slot = TC.blk_slot[i / 3]; yf[i] = TC.b[(slot / kTileBlocks) * kTS + 3 * (slot % kTileBlocks) + i % 3]; }
}
// forward, leaves: y_k = W_k b_k
__global__ void __launch_bounds__(256) tile_fwd_first_kernel(TileChol TC, int col_begin, const LMState* st) {
  __shared__ double xs[kTS]; __shared__ double part[4][kTS];
  if (!gate_run(st) || st->lin_fail) return;
  const int k = TC.lev_cols[col_begin + blockIdx.x], tid = threadIdx.x;
  if (tid < kTS) xs[tid] = TC.b[k * kTS + tid];
  __syncthreads();
  const double yk = tile_mv(TC.W + (size_t)k * kTileDoubles, xs, part, tid);
  if (tid < kTS) TC.b[k * kTS + tid] = yk;
}
// forward, one level: b_i -= sum_k L(i,k) y_k over the level's columns; if column i belongs to the next level its
// right-hand side is complete now and y_i = W_i b_i follows in the same CTA
__global__ void __launch_bounds__(256) tile_fwd_kernel(TileChol TC, int fw_begin, const LMState* st) {
  __shared__ double xs[kTS]; __shared__ double part[4][kTS];
  if (!gate_run(st) || st->lin_fail) return;
  const int d = fw_begin + blockIdx.x, i = TC.fw_row[d], tid = threadIdx.x;
  double bi = tid < kTS ? TC.b[i * kTS + tid] : 0.0;
  for (int it = TC.fw_ptr[d]; it < TC.fw_ptr[d + 1]; ++it) {
    if (tid < kTS) xs[tid] = TC.b[TC.fw_k[it] * kTS + tid];
    __syncthreads();
    bi -= tile_mv(TC.T + TC.fw_src[it], xs, part, tid);
  }
  if (TC.fw_tail[d]) {
    if (tid < kTS) xs[tid] = bi;
    __syncthreads();
    bi = tile_mv(TC.W + (size_t)i * kTileDoubles, xs, part, tid);
  }
  if (tid < kTS) TC.b[i * kTS + tid] = bi;
}
// backward, roots of the elimination forest (columns without rows): x_k = W_k^T y_k
__global__ void __launch_bounds__(256) tile_bwd_first_kernel(TileChol TC, const LMState* st) {
  __shared__ double xs[kTS];
  if (!gate_run(st) || st->lin_fail) return;
  const int k = TC.roots[blockIdx.x], tid = threadIdx.x, c = tid >> 2;
  if (tid < kTS) xs[tid] = TC.b[k * kTS + tid];
  __syncthreads();
  const double xk = tile_mv_t(TC.W + (size_t)k * kTileDoubles, xs, tid);
  __syncthreads();
  if ((tid & 3) == 0) TC.b[k * kTS + c] = xk;
}
// backward, one level (top-down): the solution of the level's columns i is pushed to every column k that has a tile in
// row i: t_k -= L(i,k)^T x_i. One CTA per destination column; the rows of a column are its ancestors, the parent the
// lowest of them, so when the parent's level has been applied t_k is complete and x_k = W_k^T t_k follows in the same CTA.
__global__ void __launch_bounds__(256) tile_bwd_kernel(TileChol TC, int bw_begin, const LMState* st) {
  __shared__ double xs[kTS];
  if (!gate_run(st) || st->lin_fail) return;
  const int d = bw_begin + blockIdx.x, k = TC.bw_col[d], tid = threadIdx.x, c = tid >> 2;
  double t = TC.b[k * kTS + c];
  for (int it = TC.bw_ptr[d]; it < TC.bw_ptr[d + 1]; ++it) {
    if (tid < kTS) xs[tid] = TC.b[TC.bw_i[it] * kTS + tid];
    __syncthreads();
    t -= tile_mv_t(TC.T + TC.bw_src[it], xs, tid);
    __syncthreads();
  }
  if (TC.bw_tail[d]) {
    if ((tid & 3) == 0) xs[c] = t;
    __syncthreads();
    t = tile_mv_t(TC.W + (size_t)k * kTileDoubles, xs, tid);
  }
  if ((tid & 3) == 0) TC.b[k * kTS + c] = t;
}

}  // namespace bsnls
