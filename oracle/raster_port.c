/* TEST INFRASTRUCTURE ONLY (oracle): plain-C restatement of the reference's triangle rasteriser and of its
 * gradient op.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may load this.
 *
 *   raster_port_forward  follows deps/tf_mesh_renderer/kernels/rasterize_triangles_impl.cc:75-155
 *                        (+ the buffer clears of rasterize_triangles_op.cc:121-125)
 *   raster_port_backward follows deps/tf_mesh_renderer/kernels/rasterize_triangles_grad.cc:144-218
 *
 * Pinned: tests/test_raster_oracle.py checks the forward against the reference's own C++ compiled from
 * /root/reference into oracle/_ref/ (bit-exact) and against the reference's golden images
 * (test_data/Simple_Triangle.png, Simple_Tetrahedron.png -> tests/golden/raster_golden.npz).
 * The gradient op depends on TensorFlow headers and cannot be compiled here; its restatement is checked
 * against central finite differences of the forward's barycentrics away from edges.
 *
 * Build: gcc -O2 -ffp-contract=off -shared -fPIC (no FMA contraction: the reference is built for plain x86-64). */
#include <math.h>
#include <stdint.h>
#include <string.h>

static int clamped_min(float a, float b, float c, int low, int high) {   /* impl.cc:25-29 */
  float m = a < b ? a : b;
  m = m < c ? m : c;
  int v = (int)floorf(m);
  v = v > low ? v : low;
  return v < high ? v : high;
}
static int clamped_max(float a, float b, float c, int low, int high) {   /* impl.cc:33-37 */
  float m = a > b ? a : b;
  m = m > c ? m : c;
  int v = (int)ceilf(m);
  v = v > low ? v : low;
  return v < high ? v : high;
}
static int32_t fixed_point(float f) { return (int32_t)(f * (float)(1 << 16)); }   /* impl.cc:41 */

static int is_ccw(int32_t v0x, int32_t v0y, int32_t v1x, int32_t v1y, int32_t px, int32_t py) {   /* impl.cc:45-56 */
  int32_t ex = v1x - v0x, ey = v1y - v0y, x = px - v0x, y = py - v0y;
  return (int64_t)ex * (int64_t)y >= (int64_t)ey * (int64_t)x;
}
static int inside(int32_t v0x, int32_t v0y, int32_t v1x, int32_t v1y, int32_t v2x, int32_t v2y, int32_t px,
                  int32_t py) {   /* impl.cc:60-71: front- or back-facing */
  return (is_ccw(v0x, v0y, v1x, v1y, px, py) && is_ccw(v1x, v1y, v2x, v2y, px, py) && is_ccw(v2x, v2y, v0x, v0y, px, py)) ||
         (is_ccw(v1x, v1y, v0x, v0y, px, py) && is_ccw(v2x, v2y, v1x, v1y, px, py) && is_ccw(v0x, v0y, v2x, v2y, px, py));
}

void raster_port_forward(const float* vertices, const int32_t* triangles, int32_t triangle_count, int32_t W, int32_t H,
                         int32_t* triangle_ids, float* bary, float* zbuf) {
  const int n = W * H;
  memset(triangle_ids, 0, sizeof(int32_t) * (size_t)n);        /* op.cc:123-125 */
  memset(bary, 0, sizeof(float) * 3 * (size_t)n);
  for (int i = 0; i < n; ++i) zbuf[i] = 1.0f;
  const float half_w = (float)(0.5 * W), half_h = (float)(0.5 * H);   /* impl.cc:79-80 */
  for (int32_t t = 0; t < triangle_count; ++t) {
    const int32_t i0 = 3 * triangles[3 * t], i1 = 3 * triangles[3 * t + 1], i2 = 3 * triangles[3 * t + 2];
    /* impl.cc:87-92: (float + 1.0) is evaluated in double, the product too, then rounded to float */
    const float v0x = (float)((vertices[i0] + 1.0) * half_w), v0y = (float)((vertices[i0 + 1] + 1.0) * half_h);
    const float v1x = (float)((vertices[i1] + 1.0) * half_w), v1y = (float)((vertices[i1 + 1] + 1.0) * half_h);
    const float v2x = (float)((vertices[i2] + 1.0) * half_w), v2y = (float)((vertices[i2 + 1] + 1.0) * half_h);
    const int left = clamped_min(v0x, v1x, v2x, 0, W), right = clamped_max(v0x, v1x, v2x, 0, W);
    const int bottom = clamped_min(v0y, v1y, v2y, 0, H), top = clamped_max(v0y, v1y, v2y, 0, H);
    const int32_t f0x = fixed_point(v0x), f0y = fixed_point(v0y), f1x = fixed_point(v1x), f1y = fixed_point(v1y),
                  f2x = fixed_point(v2x), f2y = fixed_point(v2y);
    for (int iy = bottom; iy < top; ++iy)
      for (int ix = left; ix < right; ++ix) {
        const float px = (float)(ix + 0.5), py = (float)(iy + 0.5);
        if (!inside(f0x, f0y, f1x, f1y, f2x, f2y, fixed_point(px), fixed_point(py))) continue;
        const int p = iy * W + ix;
        const float area2 = (v2x - v0x) * (v1y - v0y) - (v2y - v0y) * (v1x - v0x);           /* impl.cc:124-125 */
        const float b0 = ((px - v1x) * (v2y - v1y) - (py - v1y) * (v2x - v1x)) / area2;      /* :126-127 */
        const float b1 = ((px - v2x) * (v0y - v2y) - (py - v2y) * (v0x - v2x)) / area2;      /* :128-129 */
        const float b2 = 1.0f - b0 - b1;                                                      /* :132 */
        const float z = b0 * vertices[i0 + 2] + b1 * vertices[i1 + 2] + b2 * vertices[i2 + 2];   /* :137 */
        if (z < -1.0 || z > 1.0 || z > zbuf[p]) continue;                                     /* :141 */
        triangle_ids[p] = t;
        zbuf[p] = z;
        bary[3 * p] = b0;
        bary[3 * p + 1] = b1;
        bary[3 * p + 2] = b2;
      }
  }
}

void raster_port_backward(const float* vertices, int32_t vertex_count, const int32_t* triangles, const float* bary,
                          const int32_t* triangle_ids, const float* g_bary, int32_t W, int32_t H, float* g_vertices) {
  memset(g_vertices, 0, sizeof(float) * 3 * (size_t)vertex_count);   /* grad.cc:142 */
  for (int p = 0; p < W * H; ++p) {
    const float b0 = bary[3 * p], b1 = bary[3 * p + 1], b2 = bary[3 * p + 2];
    if (b0 + b1 + b2 < 0.9f) continue;                                /* grad.cc:25,155 */
    const float d0 = g_bary[3 * p], d1 = g_bary[3 * p + 1], d2 = g_bary[3 * p + 2];
    const int32_t* tv = triangles + 3 * triangle_ids[p];
    const int i0 = 3 * tv[0], i1 = 3 * tv[1], i2 = 3 * tv[2];
    const float v0x = vertices[i0], v0y = vertices[i0 + 1], v1x = vertices[i1], v1y = vertices[i1 + 1],
                v2x = vertices[i2], v2y = vertices[i2 + 1];
    const float area = (v1x - v0x) * (v2y - v0y) - (v1y - v0y) * (v2x - v0x);   /* grad.cc:183-184 */
    if (fabsf(area) < 1e-13f) continue;                                          /* :186 */
    const float s0 = b0 / area, s1 = b1 / area, s2 = b2 / area;                  /* :192-194 */
    const float gx = d0 * (v2y - v1y) + d1 * (v0y - v2y) + d2 * (v1y - v0y);     /* :198-209 */
    const float gy = d0 * (v1x - v2x) + d1 * (v2x - v0x) + d2 * (v0x - v1x);
    g_vertices[i0] += s0 * gx;                                                   /* :211-216 */
    g_vertices[i0 + 1] += s0 * gy;
    g_vertices[i1] += s1 * gx;
    g_vertices[i1 + 1] += s1 * gy;
    g_vertices[i2] += s2 * gx;
    g_vertices[i2 + 1] += s2 * gy;
  }
}
