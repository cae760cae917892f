// device_scene.h — layout of the GPU-resident scene (shared by the host scene compiler and the kernels).
//
// HBM layout (all arrays 256-byte aligned inside one allocation, read-only during a render, L2-resident for C1..C5):
//   nodes      : BVH4 nodes, 128 B each = 8 rows of float4 (one component per child slot 0..3):
//                  row 0 lo.x | row 1 hi.x | row 2 lo.y | row 3 hi.y | row 4 lo.z | row 5 hi.z | row 6 child refs (int bits) | row 7 unused
//                (interleaved so that the NEAR plane of axis a is row 2a + sign bit of d_a and the far plane its xor 1:
//                 scene_compile.cpp collapse() writes it, intersect.cuh node_step() reads it)
//                child ref: >= 0 inner node index; bit31 set: leaf, bits 27..30 = count-1, bits 0..26 = first prim;
//                B200RT_CHILD_EMPTY for an unused slot, whose box is inverted and infinite (lo = +inf, hi = -inf: never hit).
//   prim_data  : 3 x float4 per primitive in BVH leaf order: the record the f32 FILTER reads.  WORLD space (instance baked in
//                f64, rounded once), except for B200RT_PRIM_OBJF primitives, whose record stays in OBJECT space:
//                  sphere         A = (c, r)   (r < 0 allowed: sphere.rs:44)
//                  moving sphere  A = (c0, r)  B = (c1, -)  [absolute second centre]  C = (time0, time1, -, -)
//                  rect           A = (a0, b0, a1, b1)  B.x = k
//                  box            A = (min, -)  B = (max, -)
//                  triangle       A = (v0, -)  B = (v1, -)  C = (v2, -)
//                  medium sphere  A = (c, r)  B.x = -1/density      medium box  A = (min, -)  B = (max, -1/density)
//                (densities of baked media are per WORLD unit: -s/density under a similarity of scale s)
//   prim_obj   : same layout, OBJECT space (the original f32 values): what the f64 decision of an instanced primitive reads
//                behind M^-1, exactly like Transform::hit; only filled for instanced primitives
//   prim_meta  : int4 per primitive: x = type | B200RT_PRIM_OBJF | material << 8, y = original desc index, z = instance (-1), w = aux
//   tri_normals: 3 x float4 per entry (object-space vertex normals), indexed by prim_meta.w
//   instances  : 6 x float4 per instance: rows of the linear part of M (3), rows of the linear part of M^-1 (3)  [f32, informational]
//   instances_d: 24 doubles per instance: M (3x4 row-major), M^-1 (3x4): what to_object() and finish_hit() use
//   inst_eps   : per instance, absolute f32 rounding error of its baked world-space coordinates (filter margin)
//   materials  : float4 (type bits, texture bits, param, 0)
//   textures   : 3 x float4: (type bits, image bits, scale, 0), c0, c1
//   images     : int4 (offset, w, h, 0) + RGB8 bytes
//   lights     : 3 x float4 per light: (kind bits [0 sphere, 1 rect], original prim bits, axis bits, material bits),
//                (a0, b0, a1, b1) or (c, r), (k, area, own scale, 0) [k, area: rectangles only]
#pragma once
#include <stdint.h>

#define B200RT_CHILD_EMPTY 0x7FFFFFFF
#define B200RT_LEAF_BIT 0x80000000u
#ifndef B200RT_MAX_LEAF
#define B200RT_MAX_LEAF 4
#endif
#define B200RT_STACK 64
// prim_meta.x = type | flags | material << 8
#define B200RT_PRIM_TYPE_MASK 0x7f
#define B200RT_PRIM_OBJF 0x80  // instanced primitive whose f32 filter runs in OBJECT space (prim_data holds the object-space record)

#ifdef __CUDACC__
#include <cuda_runtime.h>
#else
struct float4 { float x, y, z, w; };
struct int4 { int x, y, z, w; };
#endif

struct DevScene {
  const float4* nodes;
  const float4* prim_data;
  const int4* prim_meta;
  const float4* tri_normals;
  const float4* instances;
  const float4* prim_obj;     // 3 x float4 per prim: OBJECT-space data of instanced prims (same layout as prim_data)
  const double* instances_d;  // 24 doubles per instance: M (3x4 row-major), M^-1 (3x4)
  const float* inst_eps;      // per instance: absolute f32 rounding error of the baked world-space coordinates
  const float4* materials;
  const float4* textures;
  const int4* images;
  const uint8_t* image_data;
  const float4* perlin_vec;  // 256
  const int* perlin_perm;    // 3 x 256
  const float4* lights;      // 3 x float4 per light
  int n_prims;
  int n_bvh_prims;  // prims [0, n_bvh_prims) are in the BVH; the rest (scene-enclosing) are tested after traversal
  int n_nodes;
  int n_lights;
  int flat;  // B200RT_SCENE_FLAT_LIST: linear scan, no BVH
  int has_objf;  // some primitive carries B200RT_PRIM_OBJF: the hot kernels are launched in their OBJF = true instantiation
};
