/* Oracle for raycast_mesh (test infrastructure only; see oracle/__init__.py).
 *
 * Contract restated from the reference wrapper -- LAB/utils/warp/kernels.py:13-75 (`raycast_mesh_kernel`:
 * `wp.mesh_query_ray(mesh, start, dir, max_dist, t, u, v, sign, n, f)`; on hit `ray_hits = start + t * dir`)
 * and LAB/utils/warp/ops.py:70 (`ray_hits` pre-filled with inf).  The ray query itself lives in warp-lang
 * (third-party, version unpinned, absent here) => PARITY UNPINNED; this is a brute-force closest-hit
 * Moeller-Trumbore over ALL triangles in fp64, two-sided, inclusive edges, 0 < t <= max_dist.
 *
 * Build: gcc -O3 -fopenmp -shared -fPIC (oracle/build.py).  Threads: OpenMP over rays.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>

/* vertices [nv,3] f32, faces [nf,3] i32, starts/dirs [n,3] f32 -> hits [n,3] f32 (inf on miss), t_out [n] f64
 * (inf on miss), face_out [n] i32 (-1 on miss). */
void oracle_raycast_mesh(const float* V, const int32_t* F, int64_t nf, const float* starts, const float* dirs,
                         int64_t n, double max_dist, float* hits, double* t_out, int32_t* face_out) {
  /* xy bounding boxes of the triangles, computed once (used only for the exact vertical-ray reject below) */
  float* bb = (float*)malloc((size_t)nf * 4 * sizeof(float));
  for (int64_t f = 0; f < nf; ++f) {
    const float* a = V + 3 * (int64_t)F[3 * f];
    const float* b = V + 3 * (int64_t)F[3 * f + 1];
    const float* c = V + 3 * (int64_t)F[3 * f + 2];
    bb[4 * f] = fminf(a[0], fminf(b[0], c[0]));
    bb[4 * f + 1] = fmaxf(a[0], fmaxf(b[0], c[0]));
    bb[4 * f + 2] = fminf(a[1], fminf(b[1], c[1]));
    bb[4 * f + 3] = fmaxf(a[1], fmaxf(b[1], c[1]));
  }
#pragma omp parallel for schedule(dynamic, 64)
  for (int64_t i = 0; i < n; ++i) {
    const double ox = starts[3 * i], oy = starts[3 * i + 1], oz = starts[3 * i + 2];
    const double dx = dirs[3 * i], dy = dirs[3 * i + 1], dz = dirs[3 * i + 2];
    const int vertical = (dx == 0.0 && dy == 0.0);
    double best = INFINITY;
    int32_t bf = -1;
    const float fox = starts[3 * i], foy = starts[3 * i + 1];
    for (int64_t f = 0; f < nf; ++f) {
      if (vertical) { /* exact reject: the ray's xy is outside the triangle's xy bounding box */
        const float* q = bb + 4 * f;
        if (fox < q[0] || fox > q[1] || foy < q[2] || foy > q[3]) continue;
      }
      const float* a = V + 3 * (int64_t)F[3 * f];
      const float* b = V + 3 * (int64_t)F[3 * f + 1];
      const float* c = V + 3 * (int64_t)F[3 * f + 2];
      const double e1x = (double)b[0] - a[0], e1y = (double)b[1] - a[1], e1z = (double)b[2] - a[2];
      const double e2x = (double)c[0] - a[0], e2y = (double)c[1] - a[1], e2z = (double)c[2] - a[2];
      const double px = dy * e2z - dz * e2y, py = dz * e2x - dx * e2z, pz = dx * e2y - dy * e2x;
      const double det = e1x * px + e1y * py + e1z * pz;
      if (fabs(det) < 1e-12) continue;
      const double inv = 1.0 / det;
      const double tx = ox - a[0], ty = oy - a[1], tz = oz - a[2];
      const double u = (tx * px + ty * py + tz * pz) * inv;
      if (u < 0.0 || u > 1.0) continue;
      const double qx = ty * e1z - tz * e1y, qy = tz * e1x - tx * e1z, qz = tx * e1y - ty * e1x;
      const double v = (dx * qx + dy * qy + dz * qz) * inv;
      if (v < 0.0 || u + v > 1.0) continue;
      const double t = (e2x * qx + e2y * qy + e2z * qz) * inv;
      if (t > 0.0 && t <= max_dist && t < best) { best = t; bf = (int32_t)f; }
    }
    if (bf >= 0) {
      hits[3 * i] = (float)(ox + best * dx);
      hits[3 * i + 1] = (float)(oy + best * dy);
      hits[3 * i + 2] = (float)(oz + best * dz);
    } else {
      hits[3 * i] = hits[3 * i + 1] = hits[3 * i + 2] = INFINITY;
    }
    if (t_out) t_out[i] = best;
    if (face_out) face_out[i] = bf;
  }
  free(bb);
}

/* ---------------------------------------------------------------------------------------------------------------
 * CPU BASELINE variant (bench.py cpu_baseline / --impl reference only): the same fp64 Moeller-Trumbore closest hit,
 * with the exact xy cell pre-filter BASELINE.md describes for vertical rays (the reference itself traverses a BVH
 * inside warp-lang; brute force over 1.2 M triangles would misrepresent it).  Triangles are binned once into a uniform
 * xy grid by bounding box; a vertical ray tests only the triangles of its cell.  Non-vertical rays fall back to the
 * brute-force loop above.  Results are identical to oracle_raycast_mesh (checked in tests/test_oracle_raycast.py).
 * ------------------------------------------------------------------------------------------------------------- */
typedef struct {
  float ox, oy, inv_cell;
  int nx, ny;
  int* start; /* [nx*ny+1] */
  int* tris;
} OracleGrid;

void* oracle_grid_build(const float* V, int64_t nv, const int32_t* F, int64_t nf, float cell) {
  float x0 = INFINITY, y0 = INFINITY, x1 = -INFINITY, y1 = -INFINITY;
  for (int64_t i = 0; i < nv; ++i) {
    x0 = fminf(x0, V[3 * i]); x1 = fmaxf(x1, V[3 * i]);
    y0 = fminf(y0, V[3 * i + 1]); y1 = fmaxf(y1, V[3 * i + 1]);
  }
  OracleGrid* g = (OracleGrid*)malloc(sizeof(OracleGrid));
  g->ox = x0 - cell; g->oy = y0 - cell; g->inv_cell = 1.0f / cell;
  g->nx = (int)((x1 - g->ox) / cell) + 2; g->ny = (int)((y1 - g->oy) / cell) + 2;
  int64_t nc = (int64_t)g->nx * g->ny;
  g->start = (int*)calloc((size_t)nc + 1, sizeof(int));
  for (int pass = 0; pass < 2; ++pass) {
    int* cursor = pass ? (int*)malloc((size_t)nc * sizeof(int)) : 0;
    if (pass) for (int64_t c = 0; c < nc; ++c) cursor[c] = g->start[c];
    for (int64_t f = 0; f < nf; ++f) {
      const float* a = V + 3 * (int64_t)F[3 * f]; const float* b = V + 3 * (int64_t)F[3 * f + 1];
      const float* c = V + 3 * (int64_t)F[3 * f + 2];
      float bx0 = fminf(a[0], fminf(b[0], c[0])), bx1 = fmaxf(a[0], fmaxf(b[0], c[0]));
      float by0 = fminf(a[1], fminf(b[1], c[1])), by1 = fmaxf(a[1], fmaxf(b[1], c[1]));
      int cx0 = (int)floorf((bx0 - g->ox) * g->inv_cell) - 1, cx1 = (int)floorf((bx1 - g->ox) * g->inv_cell) + 1;
      int cy0 = (int)floorf((by0 - g->oy) * g->inv_cell) - 1, cy1 = (int)floorf((by1 - g->oy) * g->inv_cell) + 1;
      if (cx0 < 0) cx0 = 0; if (cy0 < 0) cy0 = 0;
      if (cx1 >= g->nx) cx1 = g->nx - 1; if (cy1 >= g->ny) cy1 = g->ny - 1;
      for (int cy = cy0; cy <= cy1; ++cy)
        for (int cx = cx0; cx <= cx1; ++cx) {
          int64_t cc = (int64_t)cy * g->nx + cx;
          if (!pass) g->start[cc + 1]++; else g->tris[cursor[cc]++] = (int)f;
        }
    }
    if (!pass) {
      for (int64_t c = 0; c < nc; ++c) g->start[c + 1] += g->start[c];
      g->tris = (int*)malloc((size_t)(g->start[nc] > 0 ? g->start[nc] : 1) * sizeof(int));
    } else free(cursor);
  }
  return g;
}

void oracle_grid_free(void* gp) {
  OracleGrid* g = (OracleGrid*)gp;
  if (!g) return;
  free(g->start); free(g->tris); free(g);
}

static inline int tri_hit(const float* V, const int32_t* F, int64_t f, double ox, double oy, double oz, double dx,
                          double dy, double dz, double max_dist, double* t_out) {
  const float* a = V + 3 * (int64_t)F[3 * f]; const float* b = V + 3 * (int64_t)F[3 * f + 1];
  const float* c = V + 3 * (int64_t)F[3 * f + 2];
  const double e1x = (double)b[0] - a[0], e1y = (double)b[1] - a[1], e1z = (double)b[2] - a[2];
  const double e2x = (double)c[0] - a[0], e2y = (double)c[1] - a[1], e2z = (double)c[2] - a[2];
  const double px = dy * e2z - dz * e2y, py = dz * e2x - dx * e2z, pz = dx * e2y - dy * e2x;
  const double det = e1x * px + e1y * py + e1z * pz;
  if (fabs(det) < 1e-12) return 0;
  const double inv = 1.0 / det;
  const double tx = ox - a[0], ty = oy - a[1], tz = oz - a[2];
  const double u = (tx * px + ty * py + tz * pz) * inv;
  if (u < 0.0 || u > 1.0) return 0;
  const double qx = ty * e1z - tz * e1y, qy = tz * e1x - tx * e1z, qz = tx * e1y - ty * e1x;
  const double v = (dx * qx + dy * qy + dz * qz) * inv;
  if (v < 0.0 || u + v > 1.0) return 0;
  const double t = (e2x * qx + e2y * qy + e2z * qz) * inv;
  if (!(t > 0.0 && t <= max_dist)) return 0;
  *t_out = t;
  return 1;
}

void oracle_raycast_mesh_grid(void* gp, const float* V, const int32_t* F, int64_t nf, const float* starts,
                              const float* dirs, int64_t n, double max_dist, float* hits) {
  const OracleGrid* g = (const OracleGrid*)gp;
#pragma omp parallel for schedule(dynamic, 256)
  for (int64_t i = 0; i < n; ++i) {
    const double ox = starts[3 * i], oy = starts[3 * i + 1], oz = starts[3 * i + 2];
    const double dx = dirs[3 * i], dy = dirs[3 * i + 1], dz = dirs[3 * i + 2];
    double best = INFINITY, t;
    if (dx == 0.0 && dy == 0.0) {
      int cx = (int)floorf((starts[3 * i] - g->ox) * g->inv_cell), cy = (int)floorf((starts[3 * i + 1] - g->oy) * g->inv_cell);
      if (cx >= 0 && cy >= 0 && cx < g->nx && cy < g->ny) {
        int64_t c = (int64_t)cy * g->nx + cx;
        for (int k = g->start[c]; k < g->start[c + 1]; ++k)
          if (tri_hit(V, F, g->tris[k], ox, oy, oz, dx, dy, dz, max_dist, &t) && t < best) best = t;
      }
    } else {
      for (int64_t f = 0; f < nf; ++f)
        if (tri_hit(V, F, f, ox, oy, oz, dx, dy, dz, max_dist, &t) && t < best) best = t;
    }
    if (best < INFINITY) {
      hits[3 * i] = (float)(ox + best * dx); hits[3 * i + 1] = (float)(oy + best * dy); hits[3 * i + 2] = (float)(oz + best * dz);
    } else hits[3 * i] = hits[3 * i + 1] = hits[3 * i + 2] = INFINITY;
  }
}
