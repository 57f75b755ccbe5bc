// Stand-alone timing + correctness check of the tile Cholesky kernels on one tile column (k = 0) with `na` tiles below the
// diagonal. Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I beam_slam_b200/csrc -o build_tmp/panel_bench scripts/panel_bench.cu
#include <cstdio>
#include <cmath>
#include <vector>
#include "kernels.cuh"
#include "tile_chol.cuh"
using namespace bsnls;
int main(int argc, char** argv) {
  const int na = argc > 1 ? atoi(argv[1]) : 3, nT = na + 1, reps = 200;
  std::vector<double> T((size_t)(nT + nT * nT) * kTileDoubles, 0.0);
  // SPD diagonal tile + random tiles below
  std::vector<double> M(64 * 64);
  srand(1);
  for (auto& v : M) v = rand() / (double)RAND_MAX - 0.5;
  for (int r = 0; r < 64; ++r) for (int c = 0; c < 64; ++c) { double s = 0; for (int m = 0; m < 64; ++m) s += M[r * 64 + m] * M[c * 64 + m]; T[c * 64 + r] = s + (r == c ? 8.0 : 0.0); }
  for (int t = 1; t < nT; ++t) for (int e = 0; e < kTileDoubles; ++e) T[(size_t)t * kTileDoubles + e] = rand() / (double)RAND_MAX - 0.5;
  std::vector<long long> off((size_t)nT * nT, -1); std::vector<int> col_ptr(nT + 1, na), col_rows;
  col_ptr[0] = 0; off[0] = 0;
  for (int t = 1; t < nT; ++t) { off[(size_t)t * nT] = (long long)t * kTileDoubles; col_rows.push_back(t); }
  long long next = nT;
  for (int i = 1; i < nT; ++i) for (int j = 1; j <= i; ++j) off[(size_t)i * nT + j] = (next++) * kTileDoubles;
  double *dT, *dT0, *dW, *db; long long* doff; int *dcp, *dcr; LMState* dst;
  cudaMalloc(&dT, T.size() * 8); cudaMalloc(&dT0, T.size() * 8); cudaMalloc(&dW, (size_t)nT * kTileDoubles * 8); cudaMalloc(&db, nT * 64 * 8);
  cudaMalloc(&doff, off.size() * 8); cudaMalloc(&dcp, col_ptr.size() * 4); cudaMalloc(&dcr, (col_rows.size() + 1) * 4); cudaMalloc(&dst, sizeof(LMState));
  cudaMemcpy(dT0, T.data(), T.size() * 8, cudaMemcpyHostToDevice); cudaMemcpy(doff, off.data(), off.size() * 8, cudaMemcpyHostToDevice);
  cudaMemcpy(dcp, col_ptr.data(), col_ptr.size() * 4, cudaMemcpyHostToDevice); if (na) cudaMemcpy(dcr, col_rows.data(), col_rows.size() * 4, cudaMemcpyHostToDevice);
  cudaMemset(dst, 0, sizeof(LMState));
  double *dbraw, *dP; int *drp, *drt;
  cudaMalloc(&dbraw, nT * 64 * 8); cudaMemset(dbraw, 0, nT * 64 * 8); cudaMalloc(&dP, (size_t)(nT + nT * nT) * 64 * 8); cudaMalloc(&drp, (nT + 1) * 4); cudaMemset(drp, 0, (nT + 1) * 4); cudaMalloc(&drt, 4);
  TileChol TC{}; TC.braw = dbraw; TC.P = dP; TC.row_ptr = drp; TC.row_tile = drt; TC.nT = nT; TC.n = nT * 64; TC.tile_off = doff; TC.col_ptr = dcp; TC.col_rows = dcr; TC.T = dT; TC.W = dW; TC.b = db;
  {   // level schedule of the one column: panel CTA x works on tile col_rows[x]; update destinations (i, j), 1 <= j <= i, from column 0
    std::vector<int> pk(std::max(1, na), 0), px(std::max(1, na), -1), up; std::vector<long long> ud, ua, ub;
    for (int x = 0; x < na; ++x) px[x] = x;
    up.push_back(0);
    for (int i = 1; i < nT; ++i) for (int j = 1; j <= i; ++j) { ud.push_back(off[(size_t)i * nT + j]); ua.push_back(off[(size_t)i * nT]); ub.push_back(off[(size_t)j * nT]); up.push_back((int)ua.size()); }
    int *dpk, *dpx, *dup, *duf; long long *dud, *dua, *dub;
    std::vector<int> uf(ud.size() + 1, -1); cudaMalloc(&duf, uf.size() * 4); cudaMemcpy(duf, uf.data(), uf.size() * 4, cudaMemcpyHostToDevice); TC.upd_fac = duf;
    cudaMalloc(&dpk, pk.size() * 4); cudaMalloc(&dpx, px.size() * 4); cudaMalloc(&dup, up.size() * 4);
    cudaMalloc(&dud, (ud.size() + 1) * 8); cudaMalloc(&dua, (ua.size() + 1) * 8); cudaMalloc(&dub, (ub.size() + 1) * 8);
    cudaMemcpy(dpk, pk.data(), pk.size() * 4, cudaMemcpyHostToDevice); cudaMemcpy(dpx, px.data(), px.size() * 4, cudaMemcpyHostToDevice); cudaMemcpy(dup, up.data(), up.size() * 4, cudaMemcpyHostToDevice);
    if (!ud.empty()) { cudaMemcpy(dud, ud.data(), ud.size() * 8, cudaMemcpyHostToDevice); cudaMemcpy(dua, ua.data(), ua.size() * 8, cudaMemcpyHostToDevice); cudaMemcpy(dub, ub.data(), ub.size() * 8, cudaMemcpyHostToDevice); }
    TC.pan_k = dpk; TC.pan_x = dpx; TC.upd_ptr = dup; TC.upd_dst = dud; TC.upd_a = dua; TC.upd_b = dub;
  }
  const size_t psm = sizeof(double) * kPanelSmemDoubles, ssm = psm;
  cudaFuncSetAttribute(tile_panel_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)psm);
  cudaFuncSetAttribute(tile_syrk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ssm);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  float ms;
  for (int pass = 0; pass < 2; ++pass) {
    cudaEventRecord(e0);
    for (int r = 0; r < reps; ++r) { cudaMemcpyAsync(dT, dT0, T.size() * 8, cudaMemcpyDeviceToDevice); tile_panel_kernel<<<na > 0 ? na : 1, 256, psm>>>(TC, 0, 1, dst); }
    cudaEventRecord(e1); cudaEventSynchronize(e1); cudaEventElapsedTime(&ms, e0, e1);
  }
  float base;
  cudaEventRecord(e0); for (int r = 0; r < reps; ++r) cudaMemcpyAsync(dT, dT0, T.size() * 8, cudaMemcpyDeviceToDevice); cudaEventRecord(e1); cudaEventSynchronize(e1); cudaEventElapsedTime(&base, e0, e1);
  printf("panel (na=%d): %.2f us per launch (copy alone %.2f us)\n", na, 1e3 * ms / reps, 1e3 * base / reps);
  if (na) {
    cudaEventRecord(e0);
    for (int r = 0; r < reps; ++r) tile_syrk_kernel<<<na * (na + 1) / 2, 256, ssm>>>(TC, 0, dst);
    cudaEventRecord(e1); cudaEventSynchronize(e1); cudaEventElapsedTime(&ms, e0, e1);
    printf("syrk (%d pairs): %.2f us per launch\n", na * (na + 1) / 2, 1e3 * ms / reps);
  }
  // check W L = I and L(1,0) L_kk^T = A(1,0) using the host factor
  cudaMemcpyAsync(dT, dT0, T.size() * 8, cudaMemcpyDeviceToDevice); tile_panel_kernel<<<na > 0 ? na : 1, 256, psm>>>(TC, 0, 1, dst);
  std::vector<double> W(kTileDoubles), L(64 * 64, 0.0), X(kTileDoubles);
  cudaMemcpy(W.data(), dW, kTileDoubles * 8, cudaMemcpyDeviceToHost);
  for (int j = 0; j < 64; ++j) { double d = T[j * 64 + j]; for (int m = 0; m < j; ++m) d -= L[j * 64 + m] * L[j * 64 + m]; d = std::sqrt(d); L[j * 64 + j] = d;
    for (int r = j + 1; r < 64; ++r) { double s = T[j * 64 + r]; for (int m = 0; m < j; ++m) s -= L[r * 64 + m] * L[j * 64 + m]; L[r * 64 + j] = s / d; } }
  double err = 0; for (int r = 0; r < 64; ++r) for (int c = 0; c < 64; ++c) { double s = 0; for (int m = 0; m < 64; ++m) s += W[m * 64 + r] * L[m * 64 + c]; err = std::fmax(err, std::fabs(s - (r == c))); }
  printf("max |W L - I| = %.3e\n", err);
  if (na) { cudaMemcpy(X.data(), dT + kTileDoubles, kTileDoubles * 8, cudaMemcpyDeviceToHost); double e2 = 0;
    for (int r = 0; r < 64; ++r) for (int c = 0; c < 64; ++c) { double s = 0; for (int m = 0; m <= c; ++m) s += X[m * 64 + r] * L[c * 64 + m]; e2 = std::fmax(e2, std::fabs(s - T[kTileDoubles + c * 64 + r])); }
    printf("max |L10 Lkk^T - A10| = %.3e\n", e2); }
  LMState hs; cudaMemcpy(&hs, dst, sizeof(hs), cudaMemcpyDeviceToHost); printf("lin_fail %d, cuda: %s\n", hs.lin_fail, cudaGetErrorString(cudaGetLastError()));
  return 0;
}
