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
  double* braw;                // [nT * 64] right-hand side in tile slots, as tile_rhs_kernel leaves it for the fused forward substitution
  double* P;                   // [tiles * 64] forward-substitution products L(i,k) y_k, one vector per tile below the diagonal
  const int* row_ptr; const int* row_tile;   // per tile row k: the tiles (k, j), j < k, as indices into P (j ascending)
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
// All 256 threads of a CTA; the tile A (lower triangle, column-major, pitch kTLd) sits in shared memory and is destroyed,
// W = L^-1 is built in a second shared tile Ws (full tile, zeros above the diagonal) and copied to Wg when given.
// Right-looking on 8-column blocks; per block k:
//   potrf8   warp 0, every lane redundantly in registers: the 8x8 Cholesky of the diagonal block (a chain of 8 rsqrt) and,
//            one column per lane, the inverse of that factor -> W(k,k)
//   phase 2  L(i,k) = A(i,k) W(k,k)^T for the blocks below, and the block row k of W: W(k,j) = W(k,k) Wt(k,j)  (j < k)
//   phase 3  A(i,j) -= L(i,k) L(j,k)^T (k < j <= i) and the same row operations on the identity that becomes W:
//            Wt(i,j) -= L(i,k) W(k,j) (j <= k < i)
// Phases 2 and 3 are 8x8x8 tile products on the FP64 tensor cores (two DMMA each), dealt to the warps tile by tile.
// Look-ahead: in phase 3 warp 0 takes only the next diagonal block and factors it at once, while the other seven warps
// finish the update -- the pivot chain (64 dependent rsqrt, ~100 cycles each) is nearly all of the critical path.
// Returns true (in every thread of warp 0) when a pivot was not positive.
constexpr int kPanelSmemDoubles = 2 * kTS * kTLd;
__constant__ unsigned char c_tri8[36][2] = {{0,0},{1,0},{1,1},{2,0},{2,1},{2,2},{3,0},{3,1},{3,2},{3,3},{4,0},{4,1},{4,2},{4,3},{4,4},{5,0},{5,1},{5,2},{5,3},{5,4},{5,5},
                                            {6,0},{6,1},{6,2},{6,3},{6,4},{6,5},{6,6},{7,0},{7,1},{7,2},{7,3},{7,4},{7,5},{7,6},{7,7}};
// fragments of mma.m8n8k4 (g = lane / 4, tg = lane % 4) from a column-major matrix M with pitch kTLd:
//   RK: element (row r0 + g, column c0 + 4 ks + tg)  -- the A operand X, or the B operand Y of X Y^T
//   KN: element (row k0 + 4 ks + tg, column n0 + g)  -- the B operand Z of X Z
#define TC_FRAG_RK(M, r0, c0, ks) (M)[((c0) + 4 * (ks) + tg) * kTLd + (r0) + g]
#define TC_FRAG_KN(M, k0, n0, ks) (M)[((n0) + g) * kTLd + (k0) + 4 * (ks) + tg]
__device__ __forceinline__ bool potrf8_inv(const double* __restrict__ As, double* __restrict__ Ws, int c0, int lane) {
  // every lane: the lower triangle of the 8x8 block in registers (broadcast loads), the same arithmetic in every lane
  double a[8][8];
#pragma unroll
  for (int c = 0; c < 8; ++c)
#pragma unroll
    for (int r = c & ~1; r < 8; r += 2) { const double2 v = *reinterpret_cast<const double2*>(As + (c0 + c) * kTLd + c0 + r); a[r][c] = v.x; a[r + 1][c] = v.y; }
  double rs[8]; bool bad = false;
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    bad = bad || !(a[j][j] > 0.0);
    rs[j] = rsqrt(a[j][j]);
#pragma unroll
    for (int i = j + 1; i < 8; ++i) a[i][j] *= rs[j];
#pragma unroll
    for (int k = j + 1; k < 8; ++k)
#pragma unroll
      for (int i = k; i < 8; ++i) a[i][k] -= a[i][j] * a[k][j];
  }
  // column c = lane % 8 of the inverse by forward substitution: w_k = (delta_kc - sum_{m<k} L_km w_m) / L_kk
  const int c = lane & 7;
  double w[8], acc[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) acc[k] = (k == c) ? 1.0 : 0.0;
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    w[k] = acc[k] * rs[k];
#pragma unroll
    for (int i = k + 1; i < 8; ++i) acc[i] -= a[i][k] * w[k];
  }
  if (lane < 8) {
#pragma unroll
    for (int r = 0; r < 8; r += 2) *reinterpret_cast<double2*>(Ws + (c0 + c) * kTLd + c0 + r) = make_double2(w[r], w[r + 1]);
  }
  return bad;
}
__device__ __forceinline__ bool tile_potrf_inv(double* __restrict__ As, double* __restrict__ Ws, double* __restrict__ Wg) {
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, g = lane >> 2, tg = lane & 3;
  for (int e = tid; e < kTS * kTLd / 2; e += 256) reinterpret_cast<double2*>(Ws)[e] = make_double2(0.0, 0.0);
  __syncthreads();
  bool bad = false;
  if (warp == 0) bad = potrf8_inv(As, Ws, 0, lane);
  __syncthreads();
#pragma unroll 1
  for (int kb = 0; kb < 8; ++kb) {
    const int c0 = 8 * kb, m8 = 7 - kb;
    // ---- phase 2: the panel below block k, and block row k of W ----
    for (int t = warp; t < m8 + kb; t += 8) {
      double d0 = 0.0, d1 = 0.0;
      if (t < m8) {
        const int r0 = c0 + 8 + 8 * t;
        const double a0 = TC_FRAG_RK(As, r0, c0, 0), a1 = TC_FRAG_RK(As, r0, c0, 1), b0 = TC_FRAG_RK(Ws, c0, c0, 0), b1 = TC_FRAG_RK(Ws, c0, c0, 1);
        dmma_m8n8k4(d0, d1, a0, b0); dmma_m8n8k4(d0, d1, a1, b1);
        As[(c0 + 2 * tg) * kTLd + r0 + g] = d0; As[(c0 + 2 * tg + 1) * kTLd + r0 + g] = d1;
      } else {
        const int n0 = 8 * (t - m8);
        const double a0 = TC_FRAG_RK(Ws, c0, c0, 0), a1 = TC_FRAG_RK(Ws, c0, c0, 1), b0 = TC_FRAG_KN(Ws, c0, n0, 0), b1 = TC_FRAG_KN(Ws, c0, n0, 1);
        dmma_m8n8k4(d0, d1, a0, b0); dmma_m8n8k4(d0, d1, a1, b1);
        Ws[(n0 + 2 * tg) * kTLd + c0 + g] = d0; Ws[(n0 + 2 * tg + 1) * kTLd + c0 + g] = d1;
      }
    }
    if (kb == 7) break;
    __syncthreads();
    // ---- phase 3: trailing update of A and of W; warp 0: the next diagonal block, then its factorisation ----
    if (warp == 0) {
      const int r0 = c0 + 8;
      double d0 = As[(r0 + 2 * tg) * kTLd + r0 + g], d1 = As[(r0 + 2 * tg + 1) * kTLd + r0 + g];
      const double a0 = -TC_FRAG_RK(As, r0, c0, 0), a1 = -TC_FRAG_RK(As, r0, c0, 1), b0 = TC_FRAG_RK(As, r0, c0, 0), b1 = TC_FRAG_RK(As, r0, c0, 1);
      dmma_m8n8k4(d0, d1, a0, b0); dmma_m8n8k4(d0, d1, a1, b1);
      As[(r0 + 2 * tg) * kTLd + r0 + g] = d0; As[(r0 + 2 * tg + 1) * kTLd + r0 + g] = d1;
      __syncwarp();
      bad = potrf8_inv(As, Ws, r0, lane) || bad;
    } else {
      const int ntri = m8 * (m8 + 1) / 2 - 1, nw = m8 * (kb + 1);
      for (int u = warp - 1; u < ntri + nw; u += 7) {
        if (u < ntri) {      // A(i,j) -= L(i,k) L(j,k)^T
          const int r0 = c0 + 8 + 8 * c_tri8[u + 1][0], n0 = c0 + 8 + 8 * c_tri8[u + 1][1];
          double d0 = As[(n0 + 2 * tg) * kTLd + r0 + g], d1 = As[(n0 + 2 * tg + 1) * kTLd + r0 + g];
          const double a0 = -TC_FRAG_RK(As, r0, c0, 0), a1 = -TC_FRAG_RK(As, r0, c0, 1), b0 = TC_FRAG_RK(As, n0, c0, 0), b1 = TC_FRAG_RK(As, n0, c0, 1);
          dmma_m8n8k4(d0, d1, a0, b0); dmma_m8n8k4(d0, d1, a1, b1);
          As[(n0 + 2 * tg) * kTLd + r0 + g] = d0; As[(n0 + 2 * tg + 1) * kTLd + r0 + g] = d1;
        } else {             // Wt(i,j) -= L(i,k) W(k,j)
          const int v = u - ntri, r0 = c0 + 8 + 8 * (v / (kb + 1)), n0 = 8 * (v % (kb + 1));
          double d0 = Ws[(n0 + 2 * tg) * kTLd + r0 + g], d1 = Ws[(n0 + 2 * tg + 1) * kTLd + r0 + g];
          const double a0 = -TC_FRAG_RK(As, r0, c0, 0), a1 = -TC_FRAG_RK(As, r0, c0, 1), b0 = TC_FRAG_KN(Ws, c0, n0, 0), b1 = TC_FRAG_KN(Ws, c0, n0, 1);
          dmma_m8n8k4(d0, d1, a0, b0); dmma_m8n8k4(d0, d1, a1, b1);
          Ws[(n0 + 2 * tg) * kTLd + r0 + g] = d0; Ws[(n0 + 2 * tg + 1) * kTLd + r0 + g] = d1;
        }
      }
    }
    __syncthreads();
  }
  __syncthreads();
  if (Wg)
    for (int e = tid; e < kTileDoubles / 2; e += 256) { const int c = e >> 5, r2 = e & 31; reinterpret_cast<double2*>(Wg)[e] = *reinterpret_cast<const double2*>(Ws + c * kTLd + 2 * r2); }
  return bad;
}

// ---- panel: L(i,k) = A(i,k) L_kk^-T = A(i,k) W^T with W = L_kk^-1, one CTA per non-zero tile below the diagonal ----
// factor_here (the first level, whose diagonal tiles nobody updates): every CTA of a column factors the diagonal tile itself
// -- the same arithmetic, so the same bits, and no dependency between CTAs -- and the first one stores W (the solve needs
// it). On every other level W was left by the update kernel of the level before (see tile_syrk_kernel) and is only loaded.
// The forward substitution rides along: with W_k on chip every CTA of column k forms y_k = W_k (b_k - sum_j L(k,j) y_j)
// from the product vectors the panel CTAs of row k's tiles left in P (lower levels, fixed order j ascending; the first CTA
// stores y_k), and after its own tile product adds P(i,k) = L(i,k) y_k from the accumulators. No separate launches.
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
  const bool first = px < 0 || px == TC.col_ptr[k];   // this CTA stores W and y of its column
  __shared__ double sv[kTS], sy[kTS], spart[4][kTS];
  if (factor_here) {
    tile_load(TC.T + TC.tile_off[(long long)k * TC.nT + k], As);
    __syncthreads();
    if (tile_potrf_inv(As, Bs, first ? TC.W + (size_t)k * kTileDoubles : nullptr)) failed = 1;
    __syncthreads();
    if (tid == 0 && failed) st->lin_fail = 1;   // every CTA of a column sees the same pivots
  }
  double* Gi = na > 0 ? TC.T + TC.tile_off[(long long)TC.col_rows[px] * TC.nT + k] : nullptr;
  if (na > 0) { if (factor_here) tile_load(Gi, As); else tile_load2(Gi, As, TC.W + (size_t)k * kTileDoubles, Bs); }
  else if (!factor_here) tile_load(TC.W + (size_t)k * kTileDoubles, Bs);
  if (tid < kTS) {       // right-hand side of the column minus the products of the tiles left of the diagonal
    double v = TC.braw[k * kTS + tid];
    for (int q = TC.row_ptr[k]; q < TC.row_ptr[k + 1]; ++q) v -= TC.P[(size_t)TC.row_tile[q] * kTS + tid];
    sv[tid] = v;
  }
  __syncthreads();
  {                      // y_k = W_k v: thread (cg, r) takes 16 columns of row r
    const int cg = tid >> 6, r = tid & 63;
    double s2 = 0.0;
#pragma unroll
    for (int q = 0; q < 16; ++q) s2 += Bs[(16 * cg + q) * kTLd + r] * sv[16 * cg + q];
    spart[cg][r] = s2;
    __syncthreads();
    if (tid < kTS) { const double y = spart[0][tid] + spart[1][tid] + spart[2][tid] + spart[3][tid]; sy[tid] = y; if (first) TC.b[k * kTS + tid] = y; }
    __syncthreads();
  }
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
  double ps = 0.0;       // P(i,k) = L(i,k) y_k: the lane's 16 entries of row 8 warp + g, then the four lanes of the row
#pragma unroll
  for (int J = 0; J < 8; ++J) ps += acc[J][0] * sy[8 * J + 2 * tg] + acc[J][1] * sy[8 * J + 2 * tg + 1];
  ps += __shfl_xor_sync(0xffffffffu, ps, 1); ps += __shfl_xor_sync(0xffffffffu, ps, 2);
  if (tg == 0) TC.P[(size_t)((Gi - TC.T) / kTileDoubles) * kTS + 8 * warp + g] = ps;
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
  if (tile_potrf_inv(As, Bs, TC.W + (size_t)fac * kTileDoubles)) failed = 1;
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
__global__ void __launch_bounds__(256) tile_rhs_kernel(TileChol TC, const double* __restrict__ rhs, double* __restrict__ dst, const LMState* st) {
  if (!gate_run(st) || st->lin_fail) return;
  const int e = blockIdx.x * 256 + threadIdx.x;
  if (e >= TC.nT * kTS) return;
  const int t = e >> 6, l = e & 63, blk = l < kTU ? TC.slot_blk[t * kTileBlocks + l / 3] : -1;
  dst[e] = blk >= 0 ? rhs[3 * blk + l % 3] : 0.0;
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
