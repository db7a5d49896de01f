// raycast.cuh -- closest-hit ray query against a static triangle mesh (replaces wp.mesh_query_ray as used by
// LAB/utils/warp/kernels.py:13-75).
//
// Acceleration structure ("XY bin grid"): terrain meshes are 2.5-D, so the mesh is binned into a uniform grid of XY
// cells about half a typical triangle wide, offset from the mesh lattice so that height-field triangles do not sit on
// cell borders; each cell lists (CSR) the triangles whose XY footprint overlaps it.  A vertical ray -- the height
// scanner, ray_caster.py:247 -- touches exactly ONE cell holding 1-3 candidates; other rays walk the cells under their
// XY projection with a 2-D DDA.  Closest hit over all candidates => the result of a brute-force query (the oracle),
// independent of traversal order.
//
// Per-triangle record (48 B, three 128-bit loads): v0.xyz | 1/det2d, e1.xyz | e2.x, e2.yz.  Vertical rays use the
// projected 2-D form of Moeller-Trumbore (u, v from xy only, hit z by barycentric interpolation): ~20 flops.
// Barycentrics within 2e-5 of an edge are re-evaluated in fp64 from the original vertices so the accept/reject decision
// matches the fp64 oracle.
#pragma once
#include "common.cuh"

namespace b200lab {

// Triangle order.  Triangles are RE-ORDERED at build time so that a patch of terrain is a few contiguous runs of
// records (the fused height scan stages the patch under an env in shared memory with coalesced loads, mdp.cu):
// a triangle whose conservative cell range spans at most kHomeMargin+1 cells in x and in y is "small"; small triangles
// are sorted by their HOME cell (the min corner of that range, row-major) and `home_start` is the CSR over home cells.
// Every small triangle overlapping cell (cx, cy) has its home in [cx-kHomeMargin, cx] x [cy-kHomeMargin, cy].  The
// remaining ("large") triangles follow, ids >= first_large.  `orig_face` maps the sorted id back to the caller's face
// index (ray_face_id).  A `cell_tris` entry is  id | dy << kEntryShift  with dy = cy - home row (0 for large triangles).
constexpr int kHomeMargin = 2;   // a height-field triangle (two cells wide) spans three cells
constexpr int kEntryShift = 28;
constexpr int kEntryIdMask = (1 << kEntryShift) - 1;

struct MeshAccel {
  float ox, oy;          // grid origin (min corner)
  float cell, inv_cell;  // cell edge
  int nx, ny;
  const int* cell_start;   // [nx*ny + 1]
  const int* cell_tris;    // [entries] triangle ids (sorted order) | dy << 28
  const float4* rec;       // [F, 3] packed records (see above), sorted order; rec[3f+2].z carries the sorted id's bits
  const float* tris;       // [F, 9] original v0 v1 v2 (fp64 re-checks, normals), sorted order
  int num_faces;
  float zmin, zmax;
  const int* home_start;   // [nx*ny + 1] CSR of the small triangles over their home cells
  const int* orig_face;    // [F] sorted id -> face index of the mesh as passed to b200lab_mesh_create
  int first_large;         // ids >= first_large are large triangles (not covered by home_start)
};

// fp64 Moeller-Trumbore on the original vertices (rare: borderline barycentrics only).  Returns t or -1.
static __device__ __noinline__ float ray_tri_f64(const float* __restrict__ tp, f3 o, f3 d) {
  double V0x = __ldg(tp), V0y = __ldg(tp + 1), V0z = __ldg(tp + 2);
  double E1x = (double)__ldg(tp + 3) - V0x, E1y = (double)__ldg(tp + 4) - V0y, E1z = (double)__ldg(tp + 5) - V0z;
  double E2x = (double)__ldg(tp + 6) - V0x, E2y = (double)__ldg(tp + 7) - V0y, E2z = (double)__ldg(tp + 8) - V0z;
  double dx = d.x, dy = d.y, dz = d.z;
  double Px = dy * E2z - dz * E2y, Py = dz * E2x - dx * E2z, Pz = dx * E2y - dy * E2x;
  double Det = E1x * Px + E1y * Py + E1z * Pz;
  if (fabs(Det) < 1e-12) return -1.0f;
  double Inv = 1.0 / Det;
  double Tx = (double)o.x - V0x, Ty = (double)o.y - V0y, Tz = (double)o.z - V0z;
  double U = (Tx * Px + Ty * Py + Tz * Pz) * Inv;
  double Qx = Ty * E1z - Tz * E1y, Qy = Tz * E1x - Tx * E1z, Qz = Tx * E1y - Ty * E1x;
  double V = (dx * Qx + dy * Qy + dz * Qz) * Inv;
  if (U < 0.0 || U > 1.0 || V < 0.0 || U + V > 1.0) return -1.0f;
  float t = (float)((E2x * Qx + E2y * Qy + E2z * Qz) * Inv);
  return t > 0.0f ? t : -1.0f;
}

constexpr float kBaryBand = 2e-5f;

// Vertical ray (d = (0, 0, dz)): projected form.  Returns t (>0) or -1.
__device__ __forceinline__ float ray_tri_vertical(const MeshAccel& m, int f, f3 o, float dz, float inv_dz) {
  const float4* r = m.rec + (size_t)f * 3;
  float4 a = __ldg(r), b = __ldg(r + 1);
  float tx = o.x - a.x, ty = o.y - a.y;
  if (a.w == 0.0f) return -1.0f;  // degenerate in xy (vertical wall): a vertical ray cannot hit its interior
  float4 c = __ldg(r + 2);
  float u = (tx * c.x - ty * b.w) * a.w;   // (tx*e2y - ty*e2x) / det2
  float v = (ty * b.x - tx * b.y) * a.w;   // (ty*e1x - tx*e1y) / det2
  float w = 1.0f - u - v;
  if (u < -kBaryBand || v < -kBaryBand || w < -kBaryBand) return -1.0f;
  if (u < kBaryBand || v < kBaryBand || w < kBaryBand) return ray_tri_f64(m.tris + (size_t)f * 9, o, f3{0.f, 0.f, dz});
  float zh = a.z + u * b.z + v * c.y;
  float t = (zh - o.z) * inv_dz;
  return t > 0.0f ? t : -1.0f;
}

// General ray: Moeller-Trumbore from the packed record.
__device__ __forceinline__ float ray_tri(const MeshAccel& m, int f, f3 o, f3 d) {
  const float4* r = m.rec + (size_t)f * 3;
  float4 a = __ldg(r), b = __ldg(r + 1), c = __ldg(r + 2);
  float e1x = b.x, e1y = b.y, e1z = b.z, e2x = b.w, e2y = c.x, e2z = c.y;
  float px = d.y * e2z - d.z * e2y, py = d.z * e2x - d.x * e2z, pz = d.x * e2y - d.y * e2x;
  float det = e1x * px + e1y * py + e1z * pz;
  if (fabsf(det) < 1e-12f) return -1.0f;
  float inv = 1.0f / det;
  float tx = o.x - a.x, ty = o.y - a.y, tz = o.z - a.z;
  float u = (tx * px + ty * py + tz * pz) * inv;
  float qx = ty * e1z - tz * e1y, qy = tz * e1x - tx * e1z, qz = tx * e1y - ty * e1x;
  float v = (d.x * qx + d.y * qy + d.z * qz) * inv;
  float t = (e2x * qx + e2y * qy + e2z * qz) * inv;
  float w = 1.0f - u - v;
  if (u < -kBaryBand || v < -kBaryBand || w < -kBaryBand) return -1.0f;
  if (u < kBaryBand || v < kBaryBand || w < kBaryBand) return ray_tri_f64(m.tris + (size_t)f * 9, o, d);
  return t > 0.0f ? t : -1.0f;
}

// Projected test on an already-fetched record.
__device__ __forceinline__ float ray_tri_vertical_rec(const MeshAccel& m, int f, float4 a, float4 b, float4 c, f3 o, float dz,
                                                      float inv_dz) {
  float tx = o.x - a.x, ty = o.y - a.y;
  if (a.w == 0.0f) return -1.0f;  // degenerate in xy (vertical wall): a vertical ray cannot hit its interior
  float u = (tx * c.x - ty * b.w) * a.w;   // (tx*e2y - ty*e2x) / det2
  float v = (ty * b.x - tx * b.y) * a.w;   // (ty*e1x - tx*e1y) / det2
  float w = 1.0f - u - v;
  if (u < -kBaryBand || v < -kBaryBand || w < -kBaryBand) return -1.0f;
  if (u < kBaryBand || v < kBaryBand || w < kBaryBand) return ray_tri_f64(m.tris + (size_t)f * 9, o, f3{0.f, 0.f, dz});
  float zh = a.z + u * b.z + v * c.y;
  float t = (zh - o.z) * inv_dz;
  return t > 0.0f ? t : -1.0f;
}

// Projected test without the out-of-line fp64 branch: a borderline barycentric only raises `border` (the caller then
// redoes that ray with raycast_vertical, which decides it in fp64) -- keeps call-free the loops that interleave several rays.
__device__ __forceinline__ float ray_tri_vertical_fast(float4 a, float4 b, float4 c, float ox, float oy, float oz, float inv_dz,
                                                       bool& border) {
  float tx = ox - a.x, ty = oy - a.y;
  if (a.w == 0.0f) return -1.0f;
  float u = (tx * c.x - ty * b.w) * a.w;
  float v = (ty * b.x - tx * b.y) * a.w;
  float w = 1.0f - u - v;
  if (u < -kBaryBand || v < -kBaryBand || w < -kBaryBand) return -1.0f;
  if (u < kBaryBand || v < kBaryBand || w < kBaryBand) { border = true; return -1.0f; }
  float zh = a.z + u * b.z + v * c.y;
  float t = (zh - oz) * inv_dz;
  return t > 0.0f ? t : -1.0f;
}

// Vertical closest hit: one cell.  The query is three dependent round trips to L2 (cell range -> triangle ids ->
// records), so each level is requested for up to kVertBatch candidates at once: a typical cell (<= 4 candidates)
// costs exactly three latencies, and every lane of a warp runs the same single trip.
// Measured alternatives at 4096 envs (compute_obs, us per launch; profiles/r1_notes.md): batch 2 with padded slots 55 |
// batch 1: 59 | batch 3: 66 | batch 2 with the padded slots predicated off: 61 | a per-(cell, triangle) 8x8 sub-cell
// coverage mask that cut the triangle tests from ~3 to ~1.2 per ray: 62 (16-byte entries triple the bytes of the middle
// round trip).  The kernel is bound by the number of divergent L2 requests in flight, not by the tests.
constexpr int kVertBatch = 2;

template <int kBatch = kVertBatch>
__device__ __forceinline__ bool raycast_vertical(const MeshAccel& m, f3 o, float dz, float max_dist, float& t_out,
                                                 int& face_out) {
  float t_best = max_dist;
  int face_best = -1;
  int cx = (int)floorf((o.x - m.ox) * m.inv_cell), cy = (int)floorf((o.y - m.oy) * m.inv_cell);
  if (cx >= 0 && cy >= 0 && cx < m.nx && cy < m.ny) {
    int c = cy * m.nx + cx;
    int s = __ldg(m.cell_start + c), e = __ldg(m.cell_start + c + 1);
    float inv_dz = 1.0f / dz;
    for (int k = s; k < e; k += kBatch) {
      int f[kBatch];
#pragma unroll
      for (int i = 0; i < kBatch; ++i) f[i] = __ldg(m.cell_tris + min(k + i, e - 1)) & kEntryIdMask;
      float4 ra[kBatch], rb[kBatch], rc[kBatch];
#pragma unroll
      for (int i = 0; i < kBatch; ++i) {
        const float4* r = m.rec + (size_t)f[i] * 3;
        ra[i] = __ldg(r); rb[i] = __ldg(r + 1); rc[i] = __ldg(r + 2);
      }
#pragma unroll
      for (int i = 0; i < kBatch; ++i) {
        float t = ray_tri_vertical_rec(m, f[i], ra[i], rb[i], rc[i], o, dz, inv_dz);   // padded slots repeat the last id
        if (t > 0.0f && t <= t_best) { t_best = t; face_best = f[i]; }
      }
    }
  }
  t_out = t_best;
  face_out = face_best;
  return face_best >= 0;
}

__device__ __forceinline__ void test_cell(const MeshAccel& m, int cx, int cy, f3 o, f3 d, float& t_best, int& face_best) {
  int c = cy * m.nx + cx;
  int s = __ldg(m.cell_start + c), e = __ldg(m.cell_start + c + 1);
  for (int k = s; k < e; ++k) {
    int f = __ldg(m.cell_tris + k) & kEntryIdMask;
    float t = ray_tri(m, f, o, d);
    if (t > 0.0f && t <= t_best) { t_best = t; face_best = f; }
  }
}

// Closest hit within max_dist.  Returns true on hit.
__device__ inline bool raycast_mesh_dev(const MeshAccel& m, f3 o, f3 d, float max_dist, float& t_out, int& face_out) {
  if (d.x == 0.0f && d.y == 0.0f) {
    if (d.z == 0.0f) { t_out = max_dist; face_out = -1; return false; }
    return raycast_vertical(m, o, d.z, max_dist, t_out, face_out);
  }
  float t_best = max_dist;
  int face_best = -1;
  const float big = 3.0e38f;
  // clip the ray's XY projection to the grid rectangle, further to the z-slab of the mesh
  float x0 = m.ox, x1 = m.ox + m.nx * m.cell, y0 = m.oy, y1 = m.oy + m.ny * m.cell;
  float t0 = 0.0f, t1 = max_dist;
  if (d.x != 0.0f) {
    float ta = (x0 - o.x) / d.x, tb = (x1 - o.x) / d.x;
    t0 = fmaxf(t0, fminf(ta, tb)); t1 = fminf(t1, fmaxf(ta, tb));
  } else if (o.x < x0 || o.x >= x1) t1 = -1.0f;
  if (d.y != 0.0f) {
    float ta = (y0 - o.y) / d.y, tb = (y1 - o.y) / d.y;
    t0 = fmaxf(t0, fminf(ta, tb)); t1 = fminf(t1, fmaxf(ta, tb));
  } else if (o.y < y0 || o.y >= y1) t1 = -1.0f;
  if (d.z != 0.0f) {
    float pad = 1e-3f + 1e-5f * fmaxf(fabsf(m.zmin), fabsf(m.zmax));
    float ta = (m.zmin - pad - o.z) / d.z, tb = (m.zmax + pad - o.z) / d.z;
    t0 = fmaxf(t0, fminf(ta, tb)); t1 = fminf(t1, fmaxf(ta, tb));
  } else if (o.z < m.zmin - 1e-3f || o.z > m.zmax + 1e-3f) t1 = -1.0f;
  if (t0 <= t1) {
    float px = o.x + d.x * t0, py = o.y + d.y * t0;
    int cx = min(max((int)floorf((px - m.ox) * m.inv_cell), 0), m.nx - 1);
    int cy = min(max((int)floorf((py - m.oy) * m.inv_cell), 0), m.ny - 1);
    int sx = d.x > 0.0f ? 1 : -1, sy = d.y > 0.0f ? 1 : -1;
    float tdx = d.x != 0.0f ? m.cell / fabsf(d.x) : big;
    float tdy = d.y != 0.0f ? m.cell / fabsf(d.y) : big;
    float nbx = m.ox + (cx + (sx > 0 ? 1 : 0)) * m.cell, nby = m.oy + (cy + (sy > 0 ? 1 : 0)) * m.cell;
    float tmx = d.x != 0.0f ? (nbx - o.x) / d.x : big;
    float tmy = d.y != 0.0f ? (nby - o.y) / d.y : big;
    for (int it = 0; it < m.nx + m.ny + 2; ++it) {
      float t_exit = fminf(fminf(tmx, tmy), t1);
      test_cell(m, cx, cy, o, d, t_best, face_best);
      // a hit found in this cell may lie beyond it (triangles span cells): stop only once the best hit is
      // closer than the exit of the current cell
      if (t_best <= t_exit || t_exit >= t1) break;
      if (tmx < tmy) { cx += sx; tmx += tdx; if (cx < 0 || cx >= m.nx) break; }
      else { cy += sy; tmy += tdy; if (cy < 0 || cy >= m.ny) break; }
    }
  }
  t_out = t_best;
  face_out = face_best;
  return face_best >= 0;
}

}  // namespace b200lab
