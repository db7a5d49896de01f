// raycast.cuh -- closest-hit ray query against a triangle mesh (replaces wp.mesh_query_ray as used by
// LAB/utils/warp/kernels.py:13-75).
//
// Acceleration structure ("XY bin grid"): terrain meshes are 2.5-D, so the mesh is binned into a uniform grid of
// XY cells; each cell lists (CSR) the triangles whose XY footprint overlaps it and carries the z-range of those
// triangles.  A ray walks the cells under its XY projection with a 2-D DDA (a vertical ray -- the height scanner,
// ray_caster.py:247 -- touches exactly one cell), skips cells whose z-range it cannot reach, and runs
// Moeller-Trumbore on the few triangles of the cell.  Closest-hit over all cells visited before the running
// t_best is exceeded => identical result to a brute-force query (the oracle), independent of traversal order.
// Borderline barycentrics are re-evaluated in fp64 so the accept/reject decision matches the fp64 oracle.
#pragma once
#include "common.cuh"

namespace b200lab {

struct MeshAccel {
  float ox, oy;          // grid origin (min corner)
  float cell, inv_cell;  // cell edge
  int nx, ny;
  const int* cell_start;   // [nx*ny + 1]
  const int* cell_tris;    // [entries] triangle ids
  const float2* cell_z;    // [nx*ny] (zmin, zmax) of the triangles listed in the cell
  const float* tris;       // [F, 9] v0 v1 v2
  int num_faces;
  float zmin, zmax;
};

// Moeller-Trumbore, fp32 with fp64 re-check near the edges.  Returns t (>0) or -1.
__device__ __forceinline__ float ray_tri(const float* __restrict__ tp, f3 o, f3 d) {
  float v0x = __ldg(tp), v0y = __ldg(tp + 1), v0z = __ldg(tp + 2);
  float e1x = __ldg(tp + 3) - v0x, e1y = __ldg(tp + 4) - v0y, e1z = __ldg(tp + 5) - v0z;
  float e2x = __ldg(tp + 6) - v0x, e2y = __ldg(tp + 7) - v0y, e2z = __ldg(tp + 8) - v0z;
  float px = d.y * e2z - d.z * e2y, py = d.z * e2x - d.x * e2z, pz = d.x * e2y - d.y * e2x;
  float det = e1x * px + e1y * py + e1z * pz;
  if (fabsf(det) < 1e-12f) return -1.0f;
  float inv = 1.0f / det;
  float tx = o.x - v0x, ty = o.y - v0y, tz = o.z - v0z;
  float u = (tx * px + ty * py + tz * pz) * inv;
  float qx = ty * e1z - tz * e1y, qy = tz * e1x - tx * e1z, qz = tx * e1y - ty * e1x;
  float v = (d.x * qx + d.y * qy + d.z * qz) * inv;
  float t = (e2x * qx + e2y * qy + e2z * qz) * inv;
  const float band = 2e-5f;
  float w = 1.0f - u - v;
  if (u < -band || v < -band || w < -band) return -1.0f;
  if (u < band || v < band || w < band) {
    // borderline: redo in fp64 (decision and t)
    double V0x = v0x, V0y = v0y, V0z = v0z;
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
    if (U < 0.0 || V < 0.0 || U + V > 1.0) return -1.0f;
    t = (float)((E2x * Qx + E2y * Qy + E2z * Qz) * Inv);
  }
  return t > 0.0f ? t : -1.0f;
}

__device__ __forceinline__ void test_cell(const MeshAccel& m, int cx, int cy, f3 o, f3 d, float& t_best,
                                          int& face_best, float z_lo, float z_hi) {
  int c = cy * m.nx + cx;
  float2 zr = __ldg(m.cell_z + c);
  if (zr.x > z_hi || zr.y < z_lo) return;  // the ray segment inside this column cannot reach the triangles
  int s = __ldg(m.cell_start + c), e = __ldg(m.cell_start + c + 1);
  for (int k = s; k < e; ++k) {
    int f = __ldg(m.cell_tris + k);
    float t = ray_tri(m.tris + (size_t)f * 9, o, d);
    if (t > 0.0f && t < t_best) { t_best = t; face_best = f; }
  }
}

// Closest hit within max_dist.  Returns true on hit.
__device__ inline bool raycast_mesh_dev(const MeshAccel& m, f3 o, f3 d, float max_dist, float& t_out, int& face_out) {
  float t_best = max_dist;
  int face_best = -1;
  const float big = 3.0e38f;
  bool vertical = fabsf(d.x) < 1e-9f && fabsf(d.y) < 1e-9f;
  if (vertical) {
    float fx = (o.x - m.ox) * m.inv_cell, fy = (o.y - m.oy) * m.inv_cell;
    int cx = (int)floorf(fx), cy = (int)floorf(fy);
    if (cx >= 0 && cy >= 0 && cx < m.nx && cy < m.ny) {
      float z_lo = d.z < 0.0f ? -big : o.z, z_hi = d.z < 0.0f ? o.z : big;
      test_cell(m, cx, cy, o, d, t_best, face_best, z_lo, z_hi);
    }
  } else {
    // clip the ray's XY projection to the grid rectangle
    float x0 = m.ox, x1 = m.ox + m.nx * m.cell, y0 = m.oy, y1 = m.oy + m.ny * m.cell;
    float t0 = 0.0f, t1 = max_dist;
    if (fabsf(d.x) > 1e-12f) {
      float ta = (x0 - o.x) / d.x, tb = (x1 - o.x) / d.x;
      t0 = fmaxf(t0, fminf(ta, tb)); t1 = fminf(t1, fmaxf(ta, tb));
    } else if (o.x < x0 || o.x >= x1) t1 = -1.0f;
    if (fabsf(d.y) > 1e-12f) {
      float ta = (y0 - o.y) / d.y, tb = (y1 - o.y) / d.y;
      t0 = fmaxf(t0, fminf(ta, tb)); t1 = fminf(t1, fmaxf(ta, tb));
    } else if (o.y < y0 || o.y >= y1) t1 = -1.0f;
    if (t0 <= t1) {
      float px = o.x + d.x * t0, py = o.y + d.y * t0;
      int cx = min(max((int)floorf((px - m.ox) * m.inv_cell), 0), m.nx - 1);
      int cy = min(max((int)floorf((py - m.oy) * m.inv_cell), 0), m.ny - 1);
      int sx = d.x > 0.0f ? 1 : -1, sy = d.y > 0.0f ? 1 : -1;
      float tdx = fabsf(d.x) > 1e-12f ? m.cell / fabsf(d.x) : big;
      float tdy = fabsf(d.y) > 1e-12f ? m.cell / fabsf(d.y) : big;
      float nbx = m.ox + (cx + (sx > 0 ? 1 : 0)) * m.cell, nby = m.oy + (cy + (sy > 0 ? 1 : 0)) * m.cell;
      float tmx = fabsf(d.x) > 1e-12f ? (nbx - o.x) / d.x : big;
      float tmy = fabsf(d.y) > 1e-12f ? (nby - o.y) / d.y : big;
      float t_in = t0;
      for (int it = 0; it < m.nx + m.ny + 2; ++it) {
        float t_exit = fminf(fminf(tmx, tmy), t1);
        // z-range of the ray inside this column (with a small pad for rounding)
        float za = o.z + d.z * t_in, zb = o.z + d.z * t_exit;
        float pad = 1e-3f + 1e-5f * fabsf(za);
        test_cell(m, cx, cy, o, d, t_best, face_best, fminf(za, zb) - pad, fmaxf(za, zb) + pad);
        if (t_best <= t_exit || t_exit >= t1) break;  // nothing closer can lie in later cells
        t_in = t_exit;
        if (tmx < tmy) { cx += sx; tmx += tdx; if (cx < 0 || cx >= m.nx) break; }
        else { cy += sy; tmy += tdy; if (cy < 0 || cy >= m.ny) break; }
      }
    }
  }
  t_out = t_best;
  face_out = face_best;
  return face_best >= 0;
}

}  // namespace b200lab
