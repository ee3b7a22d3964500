// SDF-sample generation on the device: a CUDA restatement of the reference's src/PreprocessMesh.cpp (+ src/Utils.cpp,
// src/ShaderProgram.cpp), whose OpenGL path is unavailable on this box.  No graphics API: the orthographic
// rasterisation of the 100 views is plain arithmetic.
//
//   k_rast_draw     PreprocessMesh.cpp:443-477 + the vertex / geometry stages of ShaderProgram.cpp:5-85: every triangle
//                   of every view -> camera space (V = ModelViewLookAt(view, 0, AxisY)), orthographic window
//                   coordinates (glOrtho(-1, 1, 1, -1, 0, 2.5): y flipped), coverage at pixel centres (top-left rule),
//                   depth test = 64-bit atomicMin on (depth bits << 32 | triangle id): GL_LESS with in-order
//                   submission (on equal depth the lower primitive id wins)
//   k_rast_resolve  the fragment stage (ShaderProgram.cpp:87-141) for the surviving fragment of every pixel: world
//                   position (barycentric = the rasteriser's varying interpolation, exact for an orthographic view),
//                   flat normal flipped towards the camera, grazing fragments (|n_cam.z| < 0.001) write zeros but keep
//                   the depth they won; output laid out [view][x][y] = the order ValidPointsFromIm walks
//                   (Utils.cpp:28-43), so a flat mask-compaction reproduces point_verts / point_normals
//   k_sample_points SampleFromSurface (:22-87; area-CDF lower_bound + SamplePointFromTriangle, Utils.cpp:93-107) and the
//                   sample cloud of SampleSDFNearSurface (:104-131): two Gaussian perturbations per surface point and
//                   uniform points in the cube.  The reference seeds std::mt19937 from std::random_device (not
//                   reproducible by anyone); here a counter-based Philox stream keyed by the caller's seed.
//   k_knn_sdf       SampleSDFNearSurface :133-172: exact 11-nearest visible points (brute force over shared-memory
//                   tiles, replacing the nanoflann kd-tree), point-plane distance near the surface, unanimous sign vote
//   k_gather_windows deep_sdf/data.py:112-136 for a whole batch of scenes resident in HBM (contiguous windows)
//
// All arithmetic that decides coverage, depth, barycentrics and distances uses explicitly rounded fp32 operations
// (__fmul_rn / __fadd_rn / __fdiv_rn: no FMA contraction), so oracle/preprocess.py reproduces them bit for bit.
#pragma once
#include <stdint.h>

#include "philox.cuh"

namespace dsdf {

constexpr int kRastW = 400, kRastH = 400;                      // PreprocessMesh.cpp:423-424
constexpr unsigned long long kRastEmpty = ~0ull;

struct RastTri {                // one triangle in one view
  float wx[3], wy[3], z[3];     // window coordinates (pixels), depth = distance in front of the camera
};

__device__ __forceinline__ float f_mad3(float a0, float b0, float a1, float b1, float a2, float b2, float c) {
  // ((a0 b0 + a1 b1) + a2 b2) + c, every operation rounded to fp32
  return __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(a0, b0), __fmul_rn(a1, b1)), __fmul_rn(a2, b2)), c);
}

// view = 12 floats: rows of [R | t] (camera = R p + t)
__device__ __forceinline__ void rast_setup(const float* __restrict__ verts, const int* __restrict__ faces, int t,
                                           const float* __restrict__ view, RastTri& o) {
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    const int vi = faces[3 * t + c];
    const float px = verts[3 * vi], py = verts[3 * vi + 1], pz = verts[3 * vi + 2];
    const float xc = f_mad3(view[0], px, view[1], py, view[2], pz, view[3]);
    const float yc = f_mad3(view[4], px, view[5], py, view[6], pz, view[7]);
    const float zc = f_mad3(view[8], px, view[9], py, view[10], pz, view[11]);
    // glOrtho(l=-1, r=1, b=1, t=-1): x_ndc = x_cam, y_ndc = -y_cam; window = (ndc + 1) * size / 2
    o.wx[c] = __fmul_rn(__fadd_rn(xc, 1.f), 0.5f * kRastW);
    o.wy[c] = __fmul_rn(__fadd_rn(-yc, 1.f), 0.5f * kRastH);
    o.z[c] = -zc;
  }
}

__device__ __forceinline__ float edge_fn(float ax, float ay, float bx, float by, float px, float py) {
  return __fadd_rn(__fmul_rn(__fadd_rn(bx, -ax), __fadd_rn(py, -ay)), -__fmul_rn(__fadd_rn(by, -ay), __fadd_rn(px, -ax)));
}
// top-left rule for a pixel centre exactly on an edge (edge a->b of a counter-clockwise triangle in a y-down window:
// "top" = horizontal edge with b.x < a.x, "left" = edge going down)
__device__ __forceinline__ bool edge_owns(float ax, float ay, float bx, float by) {
  return (ay == by && bx < ax) || (by > ay);
}

// barycentric weights of pixel centre (px + .5, py + .5); false if not covered
__device__ __forceinline__ bool rast_cover(const RastTri& r, int px, int py, float (&b)[3]) {
  const float cx = (float)px + 0.5f, cy = (float)py + 0.5f;
  float area = edge_fn(r.wx[0], r.wy[0], r.wx[1], r.wy[1], r.wx[2], r.wy[2]);
  if (area == 0.f || !(area == area)) return false;
  float e0 = edge_fn(r.wx[1], r.wy[1], r.wx[2], r.wy[2], cx, cy);
  float e1 = edge_fn(r.wx[2], r.wy[2], r.wx[0], r.wy[0], cx, cy);
  float e2 = edge_fn(r.wx[0], r.wy[0], r.wx[1], r.wy[1], cx, cy);
  bool own0, own1, own2;
  if (area < 0.f) {            // clockwise on screen: flip signs and edge directions
    area = -area; e0 = -e0; e1 = -e1; e2 = -e2;
    own0 = edge_owns(r.wx[2], r.wy[2], r.wx[1], r.wy[1]);
    own1 = edge_owns(r.wx[0], r.wy[0], r.wx[2], r.wy[2]);
    own2 = edge_owns(r.wx[1], r.wy[1], r.wx[0], r.wy[0]);
  } else {
    own0 = edge_owns(r.wx[1], r.wy[1], r.wx[2], r.wy[2]);
    own1 = edge_owns(r.wx[2], r.wy[2], r.wx[0], r.wy[0]);
    own2 = edge_owns(r.wx[0], r.wy[0], r.wx[1], r.wy[1]);
  }
  if (e0 < 0.f || e1 < 0.f || e2 < 0.f) return false;
  if ((e0 == 0.f && !own0) || (e1 == 0.f && !own1) || (e2 == 0.f && !own2)) return false;
  b[0] = __fdiv_rn(e0, area);
  b[1] = __fdiv_rn(e1, area);
  b[2] = __fdiv_rn(e2, area);
  return true;
}
__device__ __forceinline__ float bary3(const float (&b)[3], float v0, float v1, float v2) {
  return __fadd_rn(__fadd_rn(__fmul_rn(b[0], v0), __fmul_rn(b[1], v1)), __fmul_rn(b[2], v2));
}

// grid (ceil(n_faces / 128), n_views), 128 threads; zbuf [n_views][H][W] initialised to kRastEmpty
__global__ void __launch_bounds__(128) k_rast_draw(const float* __restrict__ verts, const int* __restrict__ faces,
                                                   int n_faces, const float* __restrict__ views,
                                                   unsigned long long* __restrict__ zbuf) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x, v = blockIdx.y;
  if (t >= n_faces) return;
  RastTri r;
  rast_setup(verts, faces, t, views + 12 * v, r);
  const float xmin = fminf(r.wx[0], fminf(r.wx[1], r.wx[2])), xmax = fmaxf(r.wx[0], fmaxf(r.wx[1], r.wx[2]));
  const float ymin = fminf(r.wy[0], fminf(r.wy[1], r.wy[2])), ymax = fmaxf(r.wy[0], fmaxf(r.wy[1], r.wy[2]));
  if (!(xmin == xmin && xmax == xmax && ymin == ymin && ymax == ymax)) return;
  const int x0 = max(0, (int)floorf(xmin - 0.5f)), x1 = min(kRastW - 1, (int)ceilf(xmax - 0.5f));
  const int y0 = max(0, (int)floorf(ymin - 0.5f)), y1 = min(kRastH - 1, (int)ceilf(ymax - 0.5f));
  unsigned long long* zb = zbuf + (size_t)v * kRastW * kRastH;
  for (int py = y0; py <= y1; ++py)
    for (int px = x0; px <= x1; ++px) {
      float b[3];
      if (!rast_cover(r, px, py, b)) continue;
      const float d = bary3(b, r.z[0], r.z[1], r.z[2]);
      if (!(d >= 0.f && d <= 2.5f)) continue;                    // near = 0, far = 2.5 clip planes
      atomicMin(zb + (size_t)py * kRastW + px, ((unsigned long long)__float_as_uint(d) << 32) | (unsigned)t);
    }
}

// grid (ceil(W*H / 256), n_views), 256 threads.  out_* [n_views][W][H][4] (x-major: the order of Utils.cpp:28-43),
// w = triangle id + 1 for a valid pixel, all zeros otherwise; valid[n_views][W][H]
__global__ void __launch_bounds__(256) k_rast_resolve(const float* __restrict__ verts, const int* __restrict__ faces,
                                                      const float* __restrict__ views,
                                                      const unsigned long long* __restrict__ zbuf,
                                                      float4* __restrict__ out_vert, float4* __restrict__ out_norm,
                                                      uint8_t* __restrict__ valid) {
  const int pix = blockIdx.x * blockDim.x + threadIdx.x, v = blockIdx.y;
  if (pix >= kRastW * kRastH) return;
  const int px = pix / kRastH, py = pix % kRastH;               // x-major walk
  const size_t o = (size_t)v * kRastW * kRastH + pix;
  const unsigned long long key = zbuf[(size_t)v * kRastW * kRastH + (size_t)py * kRastW + px];
  float4 ov = make_float4(0.f, 0.f, 0.f, 0.f), on = ov;
  uint8_t ok = 0;
  if (key != kRastEmpty) {
    const int t = (int)(unsigned)(key & 0xffffffffull);
    const float* view = views + 12 * v;
    RastTri r;
    rast_setup(verts, faces, t, view, r);
    float b[3];
    rast_cover(r, px, py, b);                                    // covered by construction: same arithmetic as the draw
    const int i0 = faces[3 * t], i1 = faces[3 * t + 1], i2 = faces[3 * t + 2];
    const float ax = verts[3 * i0], ay = verts[3 * i0 + 1], az = verts[3 * i0 + 2];
    const float bx = verts[3 * i1], by = verts[3 * i1 + 1], bz = verts[3 * i1 + 2];
    const float cx = verts[3 * i2], cy = verts[3 * i2 + 1], cz = verts[3 * i2 + 2];
    // geometry stage: normal = normalize(cross(B - A, C - A)); normal_cam = V * (normal, 0)
    const float ux = __fadd_rn(bx, -ax), uy = __fadd_rn(by, -ay), uz = __fadd_rn(bz, -az);
    const float wx = __fadd_rn(cx, -ax), wy = __fadd_rn(cy, -ay), wz = __fadd_rn(cz, -az);
    float nx = __fadd_rn(__fmul_rn(uy, wz), -__fmul_rn(uz, wy));
    float ny = __fadd_rn(__fmul_rn(uz, wx), -__fmul_rn(ux, wz));
    float nz = __fadd_rn(__fmul_rn(ux, wy), -__fmul_rn(uy, wx));
    const float nl = __fsqrt_rn(__fadd_rn(__fadd_rn(__fmul_rn(nx, nx), __fmul_rn(ny, ny)), __fmul_rn(nz, nz)));
    nx = __fdiv_rn(nx, nl); ny = __fdiv_rn(ny, nl); nz = __fdiv_rn(nz, nl);
    const float cxn = f_mad3(view[0], nx, view[1], ny, view[2], nz, 0.f);
    const float cyn = f_mad3(view[4], nx, view[5], ny, view[6], nz, 0.f);
    const float czn = f_mad3(view[8], nx, view[9], ny, view[10], nz, 0.f);
    const float cl = __fsqrt_rn(__fadd_rn(__fadd_rn(__fmul_rn(cxn, cxn), __fmul_rn(cyn, cyn)), __fmul_rn(czn, czn)));
    const float d = __fdiv_rn(czn, cl);                          // dot(normalize(normal_camera), (0, 0, 1))
    if (!(fabsf(d) < 0.001f)) {                                  // (NaN falls through to the else branch as in GLSL)
      const float s = d < 0.f ? -1.f : 1.f;
      ov = make_float4(bary3(b, ax, bx, cx), bary3(b, ay, by, cy), bary3(b, az, bz, cz), (float)t + 1.0f);
      on = make_float4(s * nx, s * ny, s * nz, (float)t + 1.0f);
      ok = 1;
    }
  }
  out_vert[o] = ov;
  out_norm[o] = on;
  valid[o] = ok;
}

// ---- sample cloud ---------------------------------------------------------------------------------------------
__device__ __forceinline__ float u01(uint32_t x) { return (float)(x >> 8) * (1.0f / 16777216.0f); }   // [0, 1)
// two standard normals from two 32-bit words (Box-Muller; u1 in (0, 1])
__device__ __forceinline__ void normal2(uint32_t a, uint32_t b, float& n0, float& n1) {
  const float u1 = ((float)(a >> 8) + 1.0f) * (1.0f / 16777216.0f);
  const float u2 = u01(b);
  const float rr = sqrtf(-2.0f * logf(u1));
  float s, c;
  sincosf(6.283185307179586f * u2, &s, &c);
  n0 = rr * c;
  n1 = rr * s;
}

// xyz[2 n_surf + n_rand][3]: for surface sample i rows 2i, 2i+1 (PreprocessMesh.cpp:118-128), then the uniform ones
// (:130-135).  surf_out (nullable) [n_surf][3]: the surface points themselves.  cdf[n_faces] = inclusive area sums.
__global__ void __launch_bounds__(256) k_sample_points(const float* __restrict__ verts, const int* __restrict__ faces,
                                                       const float* __restrict__ cdf, int n_faces, long long n_surf,
                                                       long long n_rand, float stdv1, float stdv2, float cube_dim,
                                                       uint32_t seed_lo, uint32_t seed_hi, float* __restrict__ xyz,
                                                       float* __restrict__ surf_out) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n_surf) {
    uint32_t r0[4], r1[4], r2[4];
    philox4x32_10((uint32_t)i, (uint32_t)((unsigned long long)i >> 32), 0u, 0u, seed_lo, seed_hi, r0);
    philox4x32_10((uint32_t)i, (uint32_t)((unsigned long long)i >> 32), 1u, 0u, seed_lo, seed_hi, r1);
    philox4x32_10((uint32_t)i, (uint32_t)((unsigned long long)i >> 32), 2u, 0u, seed_lo, seed_hi, r2);
    const float total = cdf[n_faces - 1];
    const float pick = u01(r0[0]) * total;
    int lo = 0, hi = n_faces - 1;                                // lower_bound: first cdf[k] >= pick
    while (lo < hi) {
      const int mid = (lo + hi) >> 1;
      if (cdf[mid] < pick) lo = mid + 1; else hi = mid;
    }
    const int i0 = faces[3 * lo], i1 = faces[3 * lo + 1], i2 = faces[3 * lo + 2];
    const float sq = sqrtf(u01(r0[1])), u2 = u01(r0[2]);
    const float wa = 1.f - sq, wb = sq * (1.f - u2), wc = u2 * sq;      // Utils.cpp:104-106
    float p[3];
#pragma unroll
    for (int c = 0; c < 3; ++c) p[c] = wa * verts[3 * i0 + c] + wb * verts[3 * i1 + c] + wc * verts[3 * i2 + c];
    float g[6];
    normal2(r1[0], r1[1], g[0], g[1]);
    normal2(r1[2], r1[3], g[2], g[3]);
    normal2(r2[0], r2[1], g[4], g[5]);
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      xyz[(2 * i) * 3 + c] = p[c] + stdv1 * g[2 * c];            // samp1[j] += perterb_norm(rng)
      xyz[(2 * i + 1) * 3 + c] = p[c] + stdv2 * g[2 * c + 1];    // samp2[j] += perterb_second(rng)
      if (surf_out) surf_out[i * 3 + c] = p[c];
    }
  } else if (i < n_surf + n_rand) {
    uint32_t r[4];
    philox4x32_10((uint32_t)i, (uint32_t)((unsigned long long)i >> 32), 3u, 0u, seed_lo, seed_hi, r);
    const long long row = 2 * n_surf + (i - n_surf);
#pragma unroll
    for (int c = 0; c < 3; ++c) xyz[row * 3 + c] = u01(r[c]) * cube_dim - cube_dim / 2;
  }
}

// ---- nearest visible points, distance, sign vote ------------------------------------------------------------------
constexpr int kVotes = 11;                                       // PreprocessMesh.cpp:537
// distance + sign vote from the 11 nearest (bi sorted nearest first), SampleSDFNearSurface :147-172
__device__ __forceinline__ void knn_vote(long long q, float qx, float qy, float qz, const int (&bi)[kVotes],
                                         const float4* __restrict__ pts, const float4* __restrict__ nrm, float stdv,
                                         float* __restrict__ sdf_out, int* __restrict__ nn_out) {
  int num_pos = 0;
  float sdf = 0.f;
#pragma unroll
  for (int k = 0; k < kVotes; ++k) {
    const int ci = bi[k];
    if (nn_out) nn_out[q * kVotes + k] = ci;
    if (ci < 0) continue;                                        // fewer than 11 visible points: cannot be unanimous
    const float4 p = pts[ci], m = nrm[ci];
    const float rx = __fadd_rn(qx, -p.x), ry = __fadd_rn(qy, -p.y), rz = __fadd_rn(qz, -p.z);
    const float len = __fsqrt_rn(__fadd_rn(__fadd_rn(__fmul_rn(rx, rx), __fmul_rn(ry, ry)), __fmul_rn(rz, rz)));
    if (k == 0) {
      const float pl = __fadd_rn(__fadd_rn(__fmul_rn(m.x, rx), __fmul_rn(m.y, ry)), __fmul_rn(m.z, rz));
      sdf = len < stdv ? fabsf(pl) : len;                        // close to the surface: point-plane distance
    }
    const float d = __fadd_rn(__fadd_rn(__fmul_rn(m.x, __fdiv_rn(rx, len)), __fmul_rn(m.y, __fdiv_rn(ry, len))),
                              __fmul_rn(m.z, __fdiv_rn(rz, len)));
    if (d > 0.f) ++num_pos;
  }
  const bool full = bi[kVotes - 1] >= 0;
  if (full && (num_pos == 0 || num_pos == kVotes)) sdf_out[q] = num_pos <= kVotes / 2 ? -sdf : sdf;
  else sdf_out[q] = __int_as_float(0x7fc00000);
}


constexpr int kKnnTile = 1024;
// sdf_out[n]: the signed distance, or NaN where the vote was not unanimous (the reference drops those samples);
// nn_out (nullable) [n][11]: the neighbour indices, nearest first.  Ties in distance go to the lower index.
__global__ void __launch_bounds__(256) k_knn_sdf(const float* __restrict__ xyz, long long n,
                                                 const float4* __restrict__ pts, const float4* __restrict__ nrm,
                                                 int n_pts, float stdv, float* __restrict__ sdf_out,
                                                 int* __restrict__ nn_out) {
  __shared__ float4 tile[kKnnTile];
  const long long q = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const bool live = q < n;
  const float qx = live ? xyz[q * 3] : 0.f, qy = live ? xyz[q * 3 + 1] : 0.f, qz = live ? xyz[q * 3 + 2] : 0.f;
  float bd[kVotes];
  int bi[kVotes];
#pragma unroll
  for (int k = 0; k < kVotes; ++k) { bd[k] = 3.0e38f; bi[k] = -1; }
  for (int base = 0; base < n_pts; base += kKnnTile) {
    __syncthreads();
    for (int j = threadIdx.x; j < kKnnTile; j += blockDim.x)
      tile[j] = base + j < n_pts ? pts[base + j] : make_float4(1e18f, 1e18f, 1e18f, 0.f);
    __syncthreads();
    const int cnt = min(kKnnTile, n_pts - base);
    for (int j = 0; j < cnt; ++j) {
      const float4 p = tile[j];
      const float dx = __fadd_rn(qx, -p.x), dy = __fadd_rn(qy, -p.y), dz = __fadd_rn(qz, -p.z);
      const float d = __fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz));
      if (d < bd[kVotes - 1]) {                                  // strict: an equal distance keeps the earlier index
        float cd = d;
        int ci = base + j;
#pragma unroll
        for (int k = 0; k < kVotes; ++k) {
          // (distance, index) order: a displaced entry must not hop over an entry of equal distance and higher index
          if (cd < bd[k] || (cd == bd[k] && ci < bi[k])) {
            const float td = bd[k]; const int ti = bi[k];
            bd[k] = cd; bi[k] = ci;
            cd = td; ci = ti;
          }
        }
      }
    }
  }
  if (!live) return;
  knn_vote(q, qx, qy, qz, bi, pts, nrm, stdv, sdf_out, nn_out);
}

// ---- the same query through a uniform grid over the visible points ------------------------------------------------
// Exact, and bit-identical to k_knn_sdf (same rounded distance expression, ties to the lower ORIGINAL index): the
// points are counting-sorted into kKnnGrid^3 cubic cells over their bounding box; a query visits the cells around its
// own in growing Chebyshev shells and stops once its 11th distance is strictly below the distance to the nearest face of
// the visited cube that still has cells behind it (minus a margin that covers the rounding of the cell assignment),
// i.e. once no unvisited point can enter the list.  Within a cell the order of the points is whatever the counting
// sort's integer atomics produced; the (distance, original index) order of the insertion makes the result independent
// of it.  ~10^3 candidates per near-surface query instead of every visible point (10^6..10^7).
constexpr int kKnnGrid = 128;
constexpr int kKnnCells = kKnnGrid * kKnnGrid * kKnnGrid;
struct KnnGridHdr {
  int lo[3], hi[3];           // bounding box as order-preserving integer images of the floats (atomicMin / atomicMax)
  float origin[3], h, inv_h;  // derived by k_knn_grid_setup
};
__device__ __forceinline__ int knn_f2ord(float f) { const int i = __float_as_int(f); return i >= 0 ? i : i ^ 0x7fffffff; }
__device__ __forceinline__ float knn_ord2f(int i) { return __int_as_float(i >= 0 ? i : i ^ 0x7fffffff); }
__device__ __forceinline__ int knn_cell_axis(float p, float origin, float inv_h) {
  const int c = (int)floorf(__fmul_rn(__fadd_rn(p, -origin), inv_h));
  return min(max(c, 0), kKnnGrid - 1);
}

__global__ void k_knn_grid_init(KnnGridHdr* hdr) {
  if (threadIdx.x < 3) { hdr->lo[threadIdx.x] = 0x7fffffff; hdr->hi[threadIdx.x] = (int)0x80000000; }
}
__global__ void __launch_bounds__(256) k_knn_grid_bbox(const float4* __restrict__ pts, int n_pts, KnnGridHdr* hdr) {
  int lo[3] = {0x7fffffff, 0x7fffffff, 0x7fffffff}, hi[3] = {(int)0x80000000, (int)0x80000000, (int)0x80000000};
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n_pts; i += gridDim.x * blockDim.x) {
    const float4 p = pts[i];
    const int o[3] = {knn_f2ord(p.x), knn_f2ord(p.y), knn_f2ord(p.z)};
#pragma unroll
    for (int a = 0; a < 3; ++a) { lo[a] = min(lo[a], o[a]); hi[a] = max(hi[a], o[a]); }
  }
#pragma unroll
  for (int a = 0; a < 3; ++a) {
#pragma unroll
    for (int s = 16; s > 0; s >>= 1) {
      lo[a] = min(lo[a], __shfl_xor_sync(0xffffffffu, lo[a], s));
      hi[a] = max(hi[a], __shfl_xor_sync(0xffffffffu, hi[a], s));
    }
    if ((threadIdx.x & 31) == 0) { atomicMin(&hdr->lo[a], lo[a]); atomicMax(&hdr->hi[a], hi[a]); }
  }
}
__global__ void k_knn_grid_setup(KnnGridHdr* hdr) {
  if (threadIdx.x != 0) return;
  float ext = 0.f;
  for (int a = 0; a < 3; ++a) ext = fmaxf(ext, knn_ord2f(hdr->hi[a]) - knn_ord2f(hdr->lo[a]));
  ext = fmaxf(ext, 1e-6f);
  const float h = ext * (1.f + 1e-3f) / kKnnGrid;
  for (int a = 0; a < 3; ++a) hdr->origin[a] = knn_ord2f(hdr->lo[a]) - 0.5e-3f * ext;
  hdr->h = h;
  hdr->inv_h = 1.f / h;
}
// cells are numbered block-major: 8 x 8 x 8 cells form a coarse cell whose 512 cells (and so whose points, after the
// sort) are contiguous -- "is this coarse cell empty" is two loads of the start table
constexpr int kKnnBlock = 8, kKnnCoarse = kKnnGrid / kKnnBlock, kKnnBlockCells = kKnnBlock * kKnnBlock * kKnnBlock;
__device__ __forceinline__ int knn_cell_index(int x, int y, int z) {
  return (((z >> 3) * kKnnCoarse + (y >> 3)) * kKnnCoarse + (x >> 3)) * kKnnBlockCells + (((z & 7) * 8 + (y & 7)) * 8 + (x & 7));
}
__device__ __forceinline__ int knn_cell_of(const float4& p, const KnnGridHdr& g) {
  return knn_cell_index(knn_cell_axis(p.x, g.origin[0], g.inv_h), knn_cell_axis(p.y, g.origin[1], g.inv_h),
                        knn_cell_axis(p.z, g.origin[2], g.inv_h));
}
__global__ void __launch_bounds__(256) k_knn_grid_count(const float4* __restrict__ pts, int n_pts,
                                                        const KnnGridHdr* __restrict__ hdr, int* __restrict__ cnt) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n_pts) atomicAdd(cnt + knn_cell_of(pts[i], *hdr), 1);
}
// exclusive prefix sum over kKnnCells + 1 counts in three launches (1024 values per block)
constexpr int kKnnScanBlock = 1024;
constexpr int kKnnScanBlocks = (kKnnCells + 1 + kKnnScanBlock - 1) / kKnnScanBlock;
__device__ __forceinline__ int knn_block_exscan(int v, int* sh, int& total) {   // 1024 threads
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  int inc = v;
#pragma unroll
  for (int s = 1; s < 32; s <<= 1) { const int t = __shfl_up_sync(0xffffffffu, inc, s); if (lane >= s) inc += t; }
  if (lane == 31) sh[warp] = inc;
  __syncthreads();
  if (warp == 0) {
    int w = sh[lane];
#pragma unroll
    for (int s = 1; s < 32; s <<= 1) { const int t = __shfl_up_sync(0xffffffffu, w, s); if (lane >= s) w += t; }
    sh[lane] = w;
  }
  __syncthreads();
  total = sh[31];
  const int base = warp ? sh[warp - 1] : 0;
  __syncthreads();
  return base + inc - v;
}
__global__ void __launch_bounds__(kKnnScanBlock) k_knn_scan_local(const int* __restrict__ cnt, int* __restrict__ start,
                                                                 int* __restrict__ block_sum) {
  __shared__ int sh[32];
  const int i = blockIdx.x * kKnnScanBlock + threadIdx.x;
  int total;
  const int ex = knn_block_exscan(i < kKnnCells ? cnt[i] : 0, sh, total);
  if (i <= kKnnCells) start[i] = ex;
  if (threadIdx.x == 0) block_sum[blockIdx.x] = total;
}
__global__ void __launch_bounds__(kKnnScanBlock) k_knn_scan_blocks(int* __restrict__ block_sum) {
  __shared__ int sh[32];
  __shared__ int carry;
  if (threadIdx.x == 0) carry = 0;
  __syncthreads();
  for (int b0 = 0; b0 < kKnnScanBlocks; b0 += kKnnScanBlock) {
    const int i = b0 + threadIdx.x;
    int total;
    const int ex = knn_block_exscan(i < kKnnScanBlocks ? block_sum[i] : 0, sh, total);
    const int c = carry;
    __syncthreads();
    if (i < kKnnScanBlocks) block_sum[i] = c + ex;
    if (threadIdx.x == 0) carry = c + total;
    __syncthreads();
  }
}
__global__ void __launch_bounds__(kKnnScanBlock) k_knn_scan_add(int* __restrict__ start, const int* __restrict__ block_sum,
                                                               int* __restrict__ cursor) {
  const int i = blockIdx.x * kKnnScanBlock + threadIdx.x;
  if (i <= kKnnCells) start[i] += block_sum[blockIdx.x];
  if (i < kKnnCells) cursor[i] = 0;
}
// sorted[slot] = (x, y, z, original index bits)
__global__ void __launch_bounds__(256) k_knn_grid_fill(const float4* __restrict__ pts, int n_pts,
                                                       const KnnGridHdr* __restrict__ hdr, const int* __restrict__ start,
                                                       int* __restrict__ cursor, float4* __restrict__ sorted) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_pts) return;
  const float4 p = pts[i];
  const int c = knn_cell_of(p, *hdr);
  const int slot = start[c] + atomicAdd(cursor + c, 1);
  sorted[slot] = make_float4(p.x, p.y, p.z, __int_as_float(i));
}
// one thread per query; sdf_out / nn_out as k_knn_sdf
__global__ void __launch_bounds__(128) k_knn_sdf_grid(const float* __restrict__ xyz, long long n,
                                                      const float4* __restrict__ sorted, const int* __restrict__ start,
                                                      const KnnGridHdr* __restrict__ hdr,
                                                      const float4* __restrict__ pts, const float4* __restrict__ nrm,
                                                      float stdv, float* __restrict__ sdf_out, int* __restrict__ nn_out) {
  const long long q = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= n) return;
  const float qv[3] = {xyz[q * 3], xyz[q * 3 + 1], xyz[q * 3 + 2]};
  const float org[3] = {hdr->origin[0], hdr->origin[1], hdr->origin[2]};
  const float h = hdr->h, inv_h = hdr->inv_h;
  int c[3];
#pragma unroll
  for (int a = 0; a < 3; ++a) c[a] = knn_cell_axis(qv[a], org[a], inv_h);
  float bd[kVotes];
  int bi[kVotes];
#pragma unroll
  for (int k = 0; k < kVotes; ++k) { bd[k] = 3.0e38f; bi[k] = -1; }
  auto scan = [&](int s, int e) {
    for (int j = s; j < e; ++j) {
      const float4 p = __ldg(sorted + j);
      const float dx = __fadd_rn(qv[0], -p.x), dy = __fadd_rn(qv[1], -p.y), dz = __fadd_rn(qv[2], -p.z);
      const float d = __fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz));
      int ci = __float_as_int(p.w);
      if (d < bd[kVotes - 1] || (d == bd[kVotes - 1] && ci < bi[kVotes - 1])) {
        float cd = d;
#pragma unroll
        for (int k = 0; k < kVotes; ++k) {
          if (cd < bd[k] || (cd == bd[k] && ci < bi[k])) {
            const float td = bd[k]; const int ti = bi[k];
            bd[k] = cd; bi[k] = ci;
            cd = td; ci = ti;
          }
        }
      }
    }
  };
  auto visit = [&](int cx, int cy, int cz) {
    const int cell = knn_cell_index(cx, cy, cz);
    scan(__ldg(start + cell), __ldg(start + cell + 1));
  };
  // squared distance from the query to the box [lo, lo + w)^3 along each axis, shrunk by the rounding margin
  const float margin = 1e-3f * h;
  auto box_d2 = [&](float lx, float ly, float lz, float w) {
    const float l[3] = {lx, ly, lz};
    float d2 = 0.f;
#pragma unroll
    for (int a = 0; a < 3; ++a) {
      const float t = fmaxf(fmaxf(l[a] - qv[a], qv[a] - (l[a] + w)) - margin, 0.f);
      d2 += t * t;
    }
    return d2;
  };
  // shells of a cube lattice of `cells` per axis around `ctr`, radius R: f(x, y, z) for every lattice point on the shell
  auto shell = [&](const int (&ctr)[3], int R, int cells, auto&& f) {
    const int lo[3] = {ctr[0] - R, ctr[1] - R, ctr[2] - R}, hi[3] = {ctr[0] + R, ctr[1] + R, ctr[2] + R};
    const int z0 = max(lo[2], 0), z1 = min(hi[2], cells - 1);
    const int y0 = max(lo[1], 0), y1 = min(hi[1], cells - 1);
    const int x0 = max(lo[0], 0), x1 = min(hi[0], cells - 1);
    for (int z = z0; z <= z1; ++z)
      for (int y = y0; y <= y1; ++y) {
        if (z == lo[2] || z == hi[2] || y == lo[1] || y == hi[1]) {
          for (int x = x0; x <= x1; ++x) f(x, y, z);
        } else {
          if (lo[0] >= 0) f(lo[0], y, z);
          if (R > 0 && hi[0] <= cells - 1) f(hi[0], y, z);
        }
      }
  };
  // true once no point outside the cube of lattice cells [ctr - R, ctr + R] (cell width w) can enter the list
  auto finished = [&](const int (&ctr)[3], int R, int cells, float w, bool& all_visited) {
    float df = 3.0e38f;
    bool open = false;
#pragma unroll
    for (int a = 0; a < 3; ++a) {
      const int lo = ctr[a] - R, hi = ctr[a] + R;
      if (lo > 0) { df = fminf(df, qv[a] - (org[a] + (float)lo * w)); open = true; }
      if (hi < cells - 1) { df = fminf(df, (org[a] + (float)(hi + 1) * w) - qv[a]); open = true; }
    }
    all_visited = !open;
    if (!open) return true;
    df -= margin;
    return bi[kVotes - 1] >= 0 && df > 0.f && bd[kVotes - 1] < df * df;
  };
  // ---- fine phase: the query's own cell and the 26 around it (enough for a query close to the surface)
  constexpr int kFineR = 1;
  bool done = false, all = false;
  for (int R = 0; R <= kFineR && !done; ++R) {
    shell(c, R, kKnnGrid, visit);
    done = finished(c, R, kKnnGrid, h, all);
  }
  // ---- coarse phase: shells of 8^3-cell blocks; empty blocks cost two loads, blocks and cells that cannot hold a
  // closer point are skipped by their box distance, cells of the fine phase are not visited twice
  if (!done) {
    const int cc[3] = {c[0] >> 3, c[1] >> 3, c[2] >> 3};
    const float wc = (float)kKnnBlock * h;
    auto visit_block = [&](int X, int Y, int Z) {
      const int cb = ((Z * kKnnCoarse + Y) * kKnnCoarse + X) * kKnnBlockCells;
      if (__ldg(start + cb) == __ldg(start + cb + kKnnBlockCells)) return;
      const float bx = org[0] + (float)(X * kKnnBlock) * h, by = org[1] + (float)(Y * kKnnBlock) * h,
                  bz = org[2] + (float)(Z * kKnnBlock) * h;
      if (bi[kVotes - 1] >= 0 && box_d2(bx, by, bz, wc) > bd[kVotes - 1]) return;
      for (int l = 0; l < kKnnBlockCells; ++l) {
        const int s0 = __ldg(start + cb + l), e0 = __ldg(start + cb + l + 1);
        if (s0 == e0) continue;
        const int fx = X * kKnnBlock + (l & 7), fy = Y * kKnnBlock + ((l >> 3) & 7), fz = Z * kKnnBlock + (l >> 6);
        if (max(max(abs(fx - c[0]), abs(fy - c[1])), abs(fz - c[2])) <= kFineR) continue;
        if (bi[kVotes - 1] >= 0 &&
            box_d2(org[0] + (float)fx * h, org[1] + (float)fy * h, org[2] + (float)fz * h, h) > bd[kVotes - 1])
          continue;
        scan(s0, e0);
      }
    };
    for (int R = 0; R < kKnnCoarse; ++R) {
      shell(cc, R, kKnnCoarse, visit_block);
      if (finished(cc, R, kKnnCoarse, wc, all)) break;
    }
  }
  knn_vote(q, qv[0], qv[1], qv[2], bi, pts, nrm, stdv, sdf_out, nn_out);
}

// ---- GPU-resident sample windows (deep_sdf/data.py:112-136 for a batch) ----------------------------------------------
// samples: all scenes' [pos | neg] rows as float4; base[s] = {first pos row, first neg row}; win[s] = window starts.
// out [n_scenes][2 half] float4 = cat(pos window, neg window).  grid (ceil(2 half / 256), n_scenes), 256 threads.
__global__ void __launch_bounds__(256) k_gather_windows(const float4* __restrict__ samples,
                                                        const long long* __restrict__ base,
                                                        const long long* __restrict__ win, long long half,
                                                        float4* __restrict__ out) {
  const int s = blockIdx.y;
  const long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= 2 * half) return;
  const long long src = r < half ? base[2 * s] + win[2 * s] + r : base[2 * s + 1] + win[2 * s + 1] + (r - half);
  out[(long long)s * 2 * half + r] = __ldg(samples + src);
}

}  // namespace dsdf
