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
          if (cd < bd[k]) {
            const float td = bd[k]; const int ti = bi[k];
            bd[k] = cd; bi[k] = ci;
            cd = td; ci = ti;
          }
        }
      }
    }
  }
  if (!live) return;
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
