// scene_compile.hpp — host-side compiled scene (what b200rt_scene_create uploads).
#pragma once
#include <string>
#include <vector>

#include "device_scene.h"
#include "host_common.hpp"

namespace b200rt {

struct CompiledScene {
  std::vector<float> nodes;        // 32 floats per BVH4 node
  std::vector<float> prim_data;    // 12 floats per prim (leaf order)
  std::vector<int32_t> prim_meta;  // 4 ints per prim
  std::vector<float> tri_normals;  // 12 floats per entry
  std::vector<float> instances;    // 24 floats per instance
  std::vector<float> prim_obj;     // 12 floats per prim (object-space twin of prim_data; instanced scenes only)
  std::vector<double> instances_d; // 24 doubles per instance: M, M^-1
  std::vector<float> inst_eps;     // per instance
  std::vector<float> materials;    // 4 floats
  std::vector<float> textures;     // 12 floats
  std::vector<int32_t> images;     // 4 ints
  std::vector<uint8_t> image_data;
  std::vector<float> perlin_vec;   // 256 x 4
  std::vector<int32_t> perlin_perm;  // 3 x 256
  std::vector<float> lights;       // 12 floats per light
  uint32_t n_prims = 0, n_bvh_prims = 0, n_nodes = 0, n_lights = 0;
  uint32_t bvh4_depth = 0;  // levels of BVH4 nodes; the traversal stack needs at most 3 entries per level
  bool flat = false;
  float sah_node_steps = 0.f, sah_prim_tests = 0.f;  // SAH estimate of the BVH: node steps / primitive tests of a random line through the root box
  bool has_dense_box = false;  // the BVH has a small box holding >= 90 % of its primitives (scene_compile.cpp dense_box)
  float dense_lo[3] = {0, 0, 0}, dense_hi[3] = {0, 0, 0};
  bool has_objf = false;  // some primitive is filtered in object space (B200RT_PRIM_OBJF)
};

// Validates, bakes and builds; returns a B200RT_* status and fills err.
int compile_scene(const b200rt_scene_desc* d, CompiledScene& out, std::string& err);

}  // namespace b200rt
