// device_scene.h — layout of the GPU-resident scene (shared by the host scene compiler and the kernels).
//
// HBM layout (all arrays 128-byte aligned, read-only during a render, L2-resident for C1..C4):
//   nodes      : BVH4 nodes, 128 B each (two 64-byte halves = 8 x float4):
//                  [0..2] child lo.x/lo.y/lo.z x4   [3..5] child hi.x/hi.y/hi.z x4
//                  [6]    child refs x4 (int bits)  [7]    unused
//                child ref: >= 0 inner node index; bit31 set: leaf, bits 27..30 = count-1, bits 0..26 = first prim;
//                B200RT_CHILD_EMPTY for an unused slot.
//   prim_data  : 3 x float4 per primitive in BVH leaf order, WORLD space (instances baked):
//                  sphere  A=(c, r)                         moving  A=(c0, r) B=(c1-c0, 1/(t1-t0)) C.x = t0
//                  rect    A=(a0, b0, a1, b1) B.x = k       box     A=(min,_) B=(max,_)
//                  tri     A=(v0,_) B=(v1,_) C=(v2,_)
//                  medium sphere A=(c, r) B.x = -1/density  medium box A=(min,_) B=(max, -1/density)
//   prim_meta  : int4 per primitive: x = type | material << 8, y = original desc index, z = instance (-1), w = aux
//   tri_normals: 3 x float4 per entry (object-space vertex normals), indexed by prim_meta.w
//   instances  : 6 x float4 per instance: rows of the linear part of M (3), rows of the linear part of M^-1 (3)
//   materials  : float4 (type bits, texture bits, param, 0)
//   textures   : 3 x float4: (type bits, image bits, scale, 0), c0, c1
//   images     : int4 (offset, w, h, 0) + RGB8 bytes
//   lights     : 2 x float4 per light: (kind bits, device prim bits, axis bits, 0), (a0,b0,a1,b1 | c,r), second.x.. see kernels
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
