// raycast.cu -- mesh acceleration-structure build (host, setup time) and the stand-alone raycast_mesh entry point.
// Replaces convert_to_warp_mesh (LAB/utils/warp/ops.py:130-145: wp.Mesh builds a BVH) and raycast_mesh
// (ops.py:24-127).  See raycast.cuh for the structure and the traversal.
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "raycast.cuh"

namespace b200lab {

struct MeshHandle {
  MeshAccel host;      // device pointers inside
  MeshAccel* dev;      // device copy of the descriptor
  int64_t bytes;
  int64_t entries;
  void* buffers[6];
};

__global__ void __launch_bounds__(256)
raycast_mesh_kernel(const MeshAccel* __restrict__ mesh, const float* __restrict__ starts,
                    const float* __restrict__ dirs, int64_t n, float max_dist, float* __restrict__ hits,
                    float* __restrict__ dist, float* __restrict__ normal, int32_t* __restrict__ face_id) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  f3 o{starts[i * 3], starts[i * 3 + 1], starts[i * 3 + 2]};
  f3 d{dirs[i * 3], dirs[i * 3 + 1], dirs[i * 3 + 2]};
  float t;
  int face;
  bool hit = raycast_mesh_dev(*mesh, o, d, max_dist, t, face);
  const float inf = INFINITY;
  // ops.py:70: ray_hits initialised to inf, kernels.py:68: hit = start + t * dir
  hits[i * 3] = hit ? o.x + t * d.x : inf;
  hits[i * 3 + 1] = hit ? o.y + t * d.y : inf;
  hits[i * 3 + 2] = hit ? o.z + t * d.z : inf;
  if (dist) dist[i] = hit ? t : inf;
  if (face_id) face_id[i] = hit ? __ldg(mesh->orig_face + face) : -1;   // sorted id -> caller's face index
  if (normal) {
    float nx = inf, ny = inf, nz = inf;
    if (hit) {
      const float* tp = mesh->tris + (size_t)face * 9;
      float e1x = tp[3] - tp[0], e1y = tp[4] - tp[1], e1z = tp[5] - tp[2];
      float e2x = tp[6] - tp[0], e2y = tp[7] - tp[1], e2z = tp[8] - tp[2];
      nx = e1y * e2z - e1z * e2y; ny = e1z * e2x - e1x * e2z; nz = e1x * e2y - e1y * e2x;
      float l = sqrtf(nx * nx + ny * ny + nz * nz);
      if (l > 0.f) { nx /= l; ny /= l; nz /= l; }
    }
    normal[i * 3] = nx; normal[i * 3 + 1] = ny; normal[i * 3 + 2] = nz;
  }
}

}  // namespace b200lab

using namespace b200lab;

extern "C" {

void* b200lab_mesh_create(const float* V, int64_t nv, const int32_t* F, int64_t nf, float cell_size) {
  if (!V || !F || nv <= 0 || nf <= 0) { set_error("b200lab_mesh_create: empty mesh"); return nullptr; }
  float xmin = INFINITY, ymin = INFINITY, zmin = INFINITY, xmax = -INFINITY, ymax = -INFINITY, zmax = -INFINITY;
  for (int64_t i = 0; i < nv; ++i) {
    xmin = std::min(xmin, V[3 * i]); xmax = std::max(xmax, V[3 * i]);
    ymin = std::min(ymin, V[3 * i + 1]); ymax = std::max(ymax, V[3 * i + 1]);
    zmin = std::min(zmin, V[3 * i + 2]); zmax = std::max(zmax, V[3 * i + 2]);
  }
  for (int64_t f = 0; f < 3 * nf; ++f)
    if (F[f] < 0 || F[f] >= nv) { set_error("b200lab_mesh_create: face index out of range"); return nullptr; }
  // triangle soup (original order) and xy areas
  std::vector<float> tris0((size_t)nf * 9);
  std::vector<double> areas;
  areas.reserve(nf);
  for (int64_t f = 0; f < nf; ++f) {
    for (int k = 0; k < 3; ++k) {
      const float* p = V + 3 * (int64_t)F[3 * f + k];
      tris0[f * 9 + 3 * k] = p[0]; tris0[f * 9 + 3 * k + 1] = p[1]; tris0[f * 9 + 3 * k + 2] = p[2];
    }
    const float* t = &tris0[f * 9];
    double e1x = (double)t[3] - t[0], e1y = (double)t[4] - t[1];
    double e2x = (double)t[6] - t[0], e2y = (double)t[7] - t[1];
    double det2 = e1x * e2y - e1y * e2x;
    if (std::fabs(det2) > 1e-10) areas.push_back(0.5 * std::fabs(det2));
  }
  double ex = std::max((double)xmax - xmin, 1e-3), ey = std::max((double)ymax - ymin, 1e-3);
  if (!(cell_size > 0.f)) {
    // auto: half the edge of a median triangle's bounding square, so a lattice-aligned height field puts 1-3 triangles
    // in a cell; bounded so the grid stays <= 24 M cells
    double med = 0.0;
    if (!areas.empty()) {
      std::nth_element(areas.begin(), areas.begin() + areas.size() / 2, areas.end());
      med = areas[areas.size() / 2];
    }
    double c = med > 0.0 ? 0.5 * std::sqrt(2.0 * med) : std::sqrt(ex * ey / (double)nf);
    double cmin = std::sqrt(ex * ey / 24.0e6);
    cell_size = (float)std::max(std::max(c, cmin), 1e-4);
  }
  const float pad = cell_size;
  MeshAccel m{};
  // the 0.37-cell shift keeps lattice-aligned triangle edges off the cell borders
  m.ox = xmin - pad - 0.37f * cell_size; m.oy = ymin - pad - 0.37f * cell_size;
  m.cell = cell_size; m.inv_cell = 1.0f / cell_size;
  m.nx = (int)std::floor((xmax + pad - m.ox) / cell_size) + 1;
  m.ny = (int)std::floor((ymax + pad - m.oy) / cell_size) + 1;
  m.num_faces = (int)nf; m.zmin = zmin; m.zmax = zmax;
  const int64_t ncell = (int64_t)m.nx * m.ny;
  if (ncell > (int64_t)100e6) { set_error("b200lab_mesh_create: grid too large (%lld cells)", (long long)ncell); return nullptr; }
  if (nf > (int64_t)kEntryIdMask) { set_error("b200lab_mesh_create: more than 2^28 faces"); return nullptr; }
  // conservative cell range of every triangle (original order).  The margin covers the fp32 rounding of the ray's cell
  // index (|x| / cell ulps) with room to spare.
  std::vector<int> ocx0(nf), ocx1(nf), ocy0(nf), ocy1(nf);
  const float eps = std::max(cell_size * 2e-3f, 4e-6f * std::max(std::max(std::fabs(xmin), std::fabs(xmax)),
                                                                     std::max(std::fabs(ymin), std::fabs(ymax))));
  for (int64_t f = 0; f < nf; ++f) {
    const float* t = &tris0[f * 9];
    float bx0 = std::min(t[0], std::min(t[3], t[6])), bx1 = std::max(t[0], std::max(t[3], t[6]));
    float by0 = std::min(t[1], std::min(t[4], t[7])), by1 = std::max(t[1], std::max(t[4], t[7]));
    ocx0[f] = std::max(0, (int)std::floor((bx0 - eps - m.ox) * m.inv_cell));
    ocx1[f] = std::min(m.nx - 1, (int)std::floor((bx1 + eps - m.ox) * m.inv_cell));
    ocy0[f] = std::max(0, (int)std::floor((by0 - eps - m.oy) * m.inv_cell));
    ocy1[f] = std::min(m.ny - 1, (int)std::floor((by1 + eps - m.oy) * m.inv_cell));
  }
  // sorted order: small triangles by home cell (row-major), then the large ones in their original order
  std::vector<int64_t> key(nf);
  int64_t n_small = 0;
  for (int64_t f = 0; f < nf; ++f) {
    bool small = ocx1[f] - ocx0[f] <= kHomeMargin && ocy1[f] - ocy0[f] <= kHomeMargin;
    key[f] = small ? (int64_t)ocy0[f] * m.nx + ocx0[f] : ncell + f;
    n_small += small;
  }
  std::vector<int> order(nf);   // sorted id -> original face
  for (int64_t f = 0; f < nf; ++f) order[f] = (int)f;
  std::stable_sort(order.begin(), order.end(), [&](int a, int b) { return key[a] < key[b]; });
  m.first_large = (int)n_small;
  std::vector<int> home_start(ncell + 1, 0);
  for (int64_t s_ = 0; s_ < n_small; ++s_) home_start[key[order[s_]] + 1]++;
  for (int64_t c = 0; c < ncell; ++c) home_start[c + 1] += home_start[c];
  // soup, records and cell ranges in sorted order
  std::vector<float> tris((size_t)nf * 9);
  std::vector<float> rec((size_t)nf * 12, 0.0f);
  std::vector<int> cx0(nf), cx1(nf), cy0(nf), cy1(nf);
  for (int64_t s_ = 0; s_ < nf; ++s_) {
    const int64_t f = order[s_];
    memcpy(&tris[s_ * 9], &tris0[f * 9], 9 * sizeof(float));
    cx0[s_] = ocx0[f]; cx1[s_] = ocx1[f]; cy0[s_] = ocy0[f]; cy1[s_] = ocy1[f];
    const float* t = &tris[s_ * 9];
    double e1x = (double)t[3] - t[0], e1y = (double)t[4] - t[1], e1z = (double)t[5] - t[2];
    double e2x = (double)t[6] - t[0], e2y = (double)t[7] - t[1], e2z = (double)t[8] - t[2];
    double det2 = e1x * e2y - e1y * e2x;
    float* r = &rec[s_ * 12];
    r[0] = t[0]; r[1] = t[1]; r[2] = t[2]; r[3] = std::fabs(det2) < 1e-12 ? 0.0f : (float)(1.0 / det2);
    r[4] = (float)e1x; r[5] = (float)e1y; r[6] = (float)e1z; r[7] = (float)e2x;
    r[8] = (float)e2y; r[9] = (float)e2z;
    int32_t sid = (int32_t)s_;
    memcpy(&r[10], &sid, 4);   // the sorted id rides in the record (fp64 re-check of a staged copy)
  }
  std::vector<int> counts(ncell + 1, 0);
  // 2-D triangle/cell overlap (separating axes: the 3 edge normals; the bbox axes are implied by the cell range)
  auto overlaps = [&](int64_t f, int cx, int cy) -> bool {
    const float* t = &tris[f * 9];
    double x0 = m.ox + (double)cx * cell_size - eps, x1 = m.ox + (double)(cx + 1) * cell_size + eps;
    double y0 = m.oy + (double)cy * cell_size - eps, y1 = m.oy + (double)(cy + 1) * cell_size + eps;
    double px[3] = {t[0], t[3], t[6]}, py[3] = {t[1], t[4], t[7]};
    double area2 = (px[1] - px[0]) * (py[2] - py[0]) - (px[2] - px[0]) * (py[1] - py[0]);
    if (std::fabs(area2) < 1e-14) return true;  // degenerate in XY (vertical wall): keep for non-vertical rays
    for (int k = 0; k < 3; ++k) {
      int k2 = (k + 1) % 3, k3 = (k + 2) % 3;
      double nx_ = -(py[k2] - py[k]), ny_ = (px[k2] - px[k]);
      double side = nx_ * (px[k3] - px[k]) + ny_ * (py[k3] - py[k]);
      double s = side > 0 ? 1.0 : -1.0;
      double d0 = s * (nx_ * (x0 - px[k]) + ny_ * (y0 - py[k]));
      double d1 = s * (nx_ * (x1 - px[k]) + ny_ * (y0 - py[k]));
      double d2 = s * (nx_ * (x0 - px[k]) + ny_ * (y1 - py[k]));
      double d3 = s * (nx_ * (x1 - px[k]) + ny_ * (y1 - py[k]));
      double tol = 1e-9 * (std::fabs(nx_) + std::fabs(ny_));
      if (d0 < -tol && d1 < -tol && d2 < -tol && d3 < -tol) return false;
    }
    return true;
  };
  for (int64_t f = 0; f < nf; ++f)
    for (int cy = cy0[f]; cy <= cy1[f]; ++cy)
      for (int cx = cx0[f]; cx <= cx1[f]; ++cx)
        if (overlaps(f, cx, cy)) counts[(int64_t)cy * m.nx + cx + 1]++;
  for (int64_t c = 0; c < ncell; ++c) counts[c + 1] += counts[c];
  const int64_t entries = counts[ncell];
  std::vector<int> cell_tris((size_t)std::max<int64_t>(entries, 1));
  std::vector<int> cursor(counts.begin(), counts.end() - 1);
  for (int64_t f = 0; f < nf; ++f) {
    const bool small = f < n_small;
    for (int cy = cy0[f]; cy <= cy1[f]; ++cy)
      for (int cx = cx0[f]; cx <= cx1[f]; ++cx)
        if (overlaps(f, cx, cy))
          cell_tris[cursor[(int64_t)cy * m.nx + cx]++] = (int)f | ((small ? cy - cy0[f] : 0) << kEntryShift);
  }
  // upload
  MeshHandle* h = (MeshHandle*)calloc(1, sizeof(MeshHandle));
  size_t sz[6] = {(size_t)(ncell + 1) * sizeof(int), cell_tris.size() * sizeof(int), rec.size() * sizeof(float),
                  tris.size() * sizeof(float), (size_t)(ncell + 1) * sizeof(int), order.size() * sizeof(int)};
  const void* src[6] = {counts.data(), cell_tris.data(), rec.data(), tris.data(), home_start.data(), order.data()};
  cudaError_t err = cudaSuccess;
  for (int k = 0; k < 6 && err == cudaSuccess; ++k) {
    err = cudaMalloc(&h->buffers[k], sz[k]);
    if (err == cudaSuccess) err = cudaMemcpy(h->buffers[k], src[k], sz[k], cudaMemcpyHostToDevice);
    h->bytes += (int64_t)sz[k];
  }
  if (err == cudaSuccess) {
    m.cell_start = (const int*)h->buffers[0];
    m.cell_tris = (const int*)h->buffers[1];
    m.rec = (const float4*)h->buffers[2];
    m.tris = (const float*)h->buffers[3];
    m.home_start = (const int*)h->buffers[4];
    m.orig_face = (const int*)h->buffers[5];
    h->host = m;
    h->entries = entries;
    err = cudaMalloc((void**)&h->dev, sizeof(MeshAccel));
    if (err == cudaSuccess) err = cudaMemcpy(h->dev, &m, sizeof(MeshAccel), cudaMemcpyHostToDevice);
  }
  if (err != cudaSuccess) {
    set_error("b200lab_mesh_create: %s", cudaGetErrorString(err));
    for (int k = 0; k < 6; ++k) if (h->buffers[k]) cudaFree(h->buffers[k]);
    if (h->dev) cudaFree(h->dev);
    free(h);
    return nullptr;
  }
  return h;
}

const void* b200lab_mesh_device_ptr(void* mesh) { return mesh ? ((MeshHandle*)mesh)->dev : nullptr; }
int64_t b200lab_mesh_bytes(void* mesh) { return mesh ? ((MeshHandle*)mesh)->bytes : 0; }

void b200lab_mesh_destroy(void* mesh) {
  if (!mesh) return;
  MeshHandle* h = (MeshHandle*)mesh;
  for (int k = 0; k < 6; ++k) if (h->buffers[k]) cudaFree(h->buffers[k]);
  if (h->dev) cudaFree(h->dev);
  free(h);
}

int b200lab_raycast_mesh(void* mesh, const float* ray_starts, const float* ray_dirs, int64_t n, float max_dist,
                         float* ray_hits, float* ray_distance, float* ray_normal, int32_t* ray_face_id,
                         void* stream) {
  B200LAB_CHECK_ARG(mesh && n >= 0, "null mesh / negative ray count");
  if (n == 0) return B200LAB_OK;   // empty batch: the tensors of an empty batch have no storage (NULL pointers)
  B200LAB_CHECK_ARG(ray_starts && ray_dirs && ray_hits, "null argument");
  MeshHandle* h = (MeshHandle*)mesh;
  raycast_mesh_kernel<<<(unsigned)ceil_div(n, 256), 256, 0, (cudaStream_t)stream>>>(
      h->dev, ray_starts, ray_dirs, n, max_dist, ray_hits, ray_distance, ray_normal, ray_face_id);
  count_launch();
  return check_launch("raycast_mesh");
}
}
