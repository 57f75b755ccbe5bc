// On-disk form of the flattened problem (`bsnls_problem`, include/bsnls.h): the block-sparse SoA buffers exactly as the
// flattener hands them to bsnls_set_problem, in createProblem order. The reference has no counterpart that could be
// replayed: its Boost serialisation of a graph drops the members of the visual constraints
// (bs_constraints/include/bs_constraints/visual/euclidean_reprojection_constraint.h:99-103 archives only pixel_), so a
// deserialised graph cannot rebuild its cost functions. A snapshot taken in-process keeps
// everything the solve reads, fixes the ordering (the reference's constraint UUIDs are random per run), and lets the GPU
// path, the CPU checker and any other consumer work on identical arrays.
//
// Layout (little-endian, no padding between fields):
//   char[8]  "BSNLS\0\1\0"      magic + format version 1
//   int32    num_blocks, num_cameras, num_groups, reserved
//   int32[num_blocks] block_size, block_manifold, block_group ; uint8[num_blocks] block_constant (+ pad to 8 bytes)
//   double[sum block_size] block_values ; double[num_cameras * BSNLS_CAMERA_STRIDE] cameras
//   per group: int32 type, count, has_aux, has_loss ; int32[count * nblocks(type)] blocks ; double[count * stride(type)]
//              consts ; int32[count] res_index ; [int32[count] aux] ; [int32[count] loss_type ; double[count] loss_param]
// Host code only: no device is needed to write or read a snapshot.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>
#include <vector>

#include "../../include/bsnls.h"

namespace {
const char kMagic[8] = {'B', 'S', 'N', 'L', 'S', 0, 1, 0};
const int kNumBlocks[BSNLS_NUM_FACTOR_TYPES] = {3, 10, 5, 4, 6, 2, 2, 1, 2, BSNLS_MARGINAL_MAX_BLOCKS, 1, 1};
const int kStride[BSNLS_NUM_FACTOR_TYPES] = {BSNLS_REPROJECTION_STRIDE, BSNLS_IMU_DELTA_STRIDE, BSNLS_IMU_PRIOR_STRIDE, BSNLS_POSE_STRIDE,
                                             BSNLS_POSE_STRIDE, BSNLS_POSE_STRIDE, BSNLS_ORIENTATION_STRIDE, BSNLS_VEC3_STRIDE, BSNLS_VEC3_STRIDE, BSNLS_MARGINAL_STRIDE, BSNLS_GRAVITY_STRIDE, BSNLS_ORIENTATION_STRIDE};
struct Owned {   // a loaded problem owns its arrays: one allocation behind the public struct
  bsnls_problem p;
  std::vector<bsnls_factor_group> groups;
  std::vector<std::vector<char>> bufs;
  template <typename T> T* take(size_t n) { bufs.emplace_back(std::max<size_t>(n * sizeof(T), 8)); return reinterpret_cast<T*>(bufs.back().data()); }
};
template <typename T> bool put(FILE* f, const T* v, size_t n) { return n == 0 || std::fwrite(v, sizeof(T), n, f) == n; }
template <typename T> bool get(FILE* f, T* v, size_t n) { return n == 0 || std::fread(v, sizeof(T), n, f) == n; }
}  // namespace

extern "C" {

int bsnls_problem_save(const bsnls_problem* p, const char* path) {
  if (!p || !path || p->num_blocks < 0 || p->num_groups < 0) return BSNLS_ERR_INVALID;
  for (int g = 0; g < p->num_groups; ++g) if (p->groups[g].type < 0 || p->groups[g].type >= BSNLS_NUM_FACTOR_TYPES) return BSNLS_ERR_INVALID;
  FILE* f = std::fopen(path, "wb");
  if (!f) return BSNLS_ERR_INVALID;
  const int32_t head[4] = {p->num_blocks, p->num_cameras, p->num_groups, 0};
  size_t nv = 0; for (int b = 0; b < p->num_blocks; ++b) nv += (size_t)p->block_size[b];
  const size_t nb = (size_t)p->num_blocks, pad = (8 - nb % 8) % 8;
  const char zeros[8] = {0};
  bool ok = put(f, kMagic, 8) && put(f, head, 4) && put(f, p->block_size, nb) && put(f, p->block_manifold, nb) && put(f, p->block_group, nb) &&
            put(f, p->block_constant, nb) && put(f, zeros, pad) && put(f, p->block_values, nv) &&
            put(f, p->cameras, (size_t)p->num_cameras * BSNLS_CAMERA_STRIDE);
  for (int g = 0; ok && g < p->num_groups; ++g) {
    const bsnls_factor_group& G = p->groups[g];
    const size_t n = (size_t)G.count;
    const int32_t gh[4] = {G.type, G.count, G.aux ? 1 : 0, (G.loss_type && G.loss_param) ? 1 : 0};
    ok = put(f, gh, 4) && put(f, G.blocks, n * kNumBlocks[G.type]) && put(f, G.consts, n * kStride[G.type]) && put(f, G.res_index, n);
    if (ok && gh[2]) ok = put(f, G.aux, n);
    if (ok && gh[3]) ok = put(f, G.loss_type, n) && put(f, G.loss_param, n);
  }
  ok = (std::fclose(f) == 0) && ok;
  return ok ? BSNLS_OK : BSNLS_ERR_INVALID;
}

int bsnls_problem_load(const char* path, bsnls_problem** out) {
  if (!path || !out) return BSNLS_ERR_INVALID;
  *out = nullptr;
  FILE* f = std::fopen(path, "rb");
  if (!f) return BSNLS_ERR_INVALID;
  Owned* o = new (std::nothrow) Owned();
  if (!o) { std::fclose(f); return BSNLS_ERR_INVALID; }
  char magic[8]; int32_t head[4];
  bool ok = get(f, magic, 8) && std::memcmp(magic, kMagic, 8) == 0 && get(f, head, 4) && head[0] >= 0 && head[1] >= 0 && head[2] >= 0 && head[2] <= BSNLS_NUM_FACTOR_TYPES;
  if (ok) {
    const size_t nb = (size_t)head[0], pad = (8 - nb % 8) % 8;
    int32_t* bs = o->take<int32_t>(nb); int32_t* bm = o->take<int32_t>(nb); int32_t* bg = o->take<int32_t>(nb); uint8_t* bc = o->take<uint8_t>(nb);
    char skip[8];
    ok = get(f, bs, nb) && get(f, bm, nb) && get(f, bg, nb) && get(f, bc, nb) && get(f, skip, pad);
    size_t nv = 0;
    for (size_t b = 0; ok && b < nb; ++b) { if (bs[b] < 0 || bs[b] > 64) ok = false; nv += (size_t)bs[b]; }
    double* bv = o->take<double>(nv); double* cams = o->take<double>((size_t)head[1] * BSNLS_CAMERA_STRIDE);
    ok = ok && get(f, bv, nv) && get(f, cams, (size_t)head[1] * BSNLS_CAMERA_STRIDE);
    o->p.num_blocks = head[0]; o->p.block_size = bs; o->p.block_manifold = bm; o->p.block_group = bg; o->p.block_constant = bc;
    o->p.block_values = bv; o->p.num_cameras = head[1]; o->p.cameras = cams;
    o->groups.resize((size_t)head[2]);
    for (int g = 0; ok && g < head[2]; ++g) {
      int32_t gh[4];
      ok = get(f, gh, 4) && gh[0] >= 0 && gh[0] < BSNLS_NUM_FACTOR_TYPES && gh[1] >= 0;
      if (!ok) break;
      const size_t n = (size_t)gh[1];
      bsnls_factor_group& G = o->groups[(size_t)g];
      std::memset(&G, 0, sizeof(G));
      G.type = gh[0]; G.count = gh[1];
      int32_t* blocks = o->take<int32_t>(n * kNumBlocks[gh[0]]); double* consts = o->take<double>(n * kStride[gh[0]]); int32_t* ri = o->take<int32_t>(n);
      ok = get(f, blocks, n * kNumBlocks[gh[0]]) && get(f, consts, n * kStride[gh[0]]) && get(f, ri, n);
      G.blocks = blocks; G.consts = consts; G.res_index = ri;
      if (ok && gh[2]) { int32_t* aux = o->take<int32_t>(n); ok = get(f, aux, n); G.aux = aux; }
      if (ok && gh[3]) { int32_t* lt = o->take<int32_t>(n); double* lp = o->take<double>(n); ok = get(f, lt, n) && get(f, lp, n); G.loss_type = lt; G.loss_param = lp; }
    }
    o->p.num_groups = head[2]; o->p.groups = o->groups.data();
  }
  std::fclose(f);
  if (!ok) { delete o; return BSNLS_ERR_INVALID; }
  *out = &o->p;   // first member of Owned: bsnls_problem_free casts back
  return BSNLS_OK;
}

void bsnls_problem_free(bsnls_problem* p) { delete reinterpret_cast<Owned*>(p); }

}  // extern "C"
