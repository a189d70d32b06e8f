// TEST INFRASTRUCTURE ONLY (oracle): extern "C" entry around the reference's own rasteriser, which is compiled
// from /root/reference/deps/tf_mesh_renderer/kernels/rasterize_triangles_impl.cc where it lies (never copied).
// The clears are those of the op wrapper (rasterize_triangles_op.cc:121-125).
#include <algorithm>
#include "rasterize_triangles_impl.h"

extern "C" void raster_ref_forward(const float* vertices, const int* triangles, int triangle_count, int W, int H,
                                   int* triangle_ids, float* bary, float* zbuf) {
  std::fill(triangle_ids, triangle_ids + (long)W * H, 0);
  std::fill(bary, bary + 3L * W * H, 0.0f);
  std::fill(zbuf, zbuf + (long)W * H, 1.0f);
  tf_mesh_renderer::RasterizeTrianglesImpl(vertices, triangles, triangle_count, W, H, triangle_ids, bary, zbuf);
}
