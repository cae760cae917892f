// scene_compile.cpp — host half of b200rt_scene_create: validates the flat tables, bakes instance transforms
// into world-space primitives (objects/transform.rs:110-148 computes a transformed bounding box per object; here
// the geometry itself is transformed once, in f64, so the device traverses a single-level BVH), and builds the
// wide BVH that replaces BVHNode::new (objects/bvh.rs:60-83: random axis + median split).  The builder is a
// binned-SAH binary build collapsed to 4-wide nodes; topology does not affect closest-hit results.
#include <algorithm>
#include <cfloat>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <future>
#include <numeric>
#include <system_error>
#include <thread>

#include "scene_compile.hpp"

namespace b200rt {

namespace {

struct Box {
  float lo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, hi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
  void grow(const Box& b) { for (int a = 0; a < 3; ++a) { lo[a] = std::min(lo[a], b.lo[a]); hi[a] = std::max(hi[a], b.hi[a]); } }
  void grow(D3 p) { for (int a = 0; a < 3; ++a) { lo[a] = std::min(lo[a], (float)std::nextafter((float)p[a], -FLT_MAX)); hi[a] = std::max(hi[a], (float)std::nextafter((float)p[a], FLT_MAX)); } }
  float area() const {
    float dx = hi[0] - lo[0], dy = hi[1] - lo[1], dz = hi[2] - lo[2];
    if (dx < 0 || dy < 0 || dz < 0) return 0;
    return 2 * (dx * dy + dy * dz + dz * dx);
  }
  // Conservative padding: the f32 slab test and the f32 candidate filters carry ~1e-6 relative error, while
  // rectangles/triangles touch their bounds exactly (the reference pads by 1e-4 / 0.1, rectangle.rs:36, mesh.rs:33).
  void pad() {
    float m = 0;
    for (int a = 0; a < 3; ++a) m = std::max(m, std::max(std::fabs(lo[a]), std::fabs(hi[a])));
    float e = 1e-5f * m + 1e-7f;
    for (int a = 0; a < 3; ++a) { lo[a] -= e; hi[a] += e; }
  }
};

struct BinNode {
  Box box;
  int left = -1, right = -1;  // children (binary) or -1
  int first = 0, count = 0;   // leaf range
};

struct BinBuilder {
  // One record per primitive, kept in BUILD ORDER: every pass over a node's range streams through memory (box, centroid and
  // size travel with the index when a range is partitioned) instead of gathering through an index array — the top levels of a
  // 300,000-primitive build are passes over 15 MB and were latency-bound as gathers.
  struct Item {
    Box box;
    float key[4];  // centroid x, y, z; -log2(box area): the fourth split "axis" of build()
    int id;
  };
  std::vector<Item> items;
  std::vector<int> order;       // filled by build(): the primitives in leaf order
  std::vector<BinNode> nodes;
  bool size_split = true;       // B200RT_BVH_SIZE_SPLIT=0 turns the size axis off (A/B of the builder)
  int n_bins = 16;              // SAH bins per axis (B200RT_BVH_BINS, 4..32)
  float trav_cost = 0.f;        // cost of a node step in primitive tests when deciding leaf against split (B200RT_BVH_TRAV_COST)
  explicit BinBuilder(const std::vector<Box>& boxes) {
    items.resize(boxes.size());
    for (size_t i = 0; i < boxes.size(); ++i) {
      Item& it = items[i];
      it.box = boxes[i];
      it.id = (int)i;
      for (int a = 0; a < 3; ++a) it.key[a] = (float)(0.5 * ((double)boxes[i].lo[a] + boxes[i].hi[a]));
      const float ar = boxes[i].area();
      it.key[3] = -std::log2(ar > 1e-30f ? (ar < 1e30f ? ar : 1e30f) : 1e-30f);
    }
    if (const char* e = std::getenv("B200RT_BVH_SIZE_SPLIT")) size_split = std::atoi(e) != 0;
    if (const char* e = std::getenv("B200RT_BVH_TRAV_COST")) trav_cost = (float)std::atof(e);
    if (const char* e = std::getenv("B200RT_BVH_BINS")) n_bins = std::min(32, std::max(4, std::atoi(e)));
  }
  // depth_left: binary levels this subtree may still use.  While the budget is roomy the split is binned SAH; once
  // count > 2^(depth_left - 1) only an even split still fits, so the range is halved at the median of its widest
  // centroid axis (a chain of geometrically growing primitives, or a deep mesh, would otherwise give a tree whose
  // BVH4 collapse needs more traversal-stack entries than B200RT_STACK).
  // Node slots are assigned up front — a subtree over c primitives owns the 2c - 1 slots from its root's, the left child's
  // subtree starts at the next slot and the right child's 2 * (left count) slots after its root — so the halves of a large node
  // can be built on two threads without sharing anything but disjoint ranges of `items` and `nodes`, and the tree does not
  // depend on the thread count (slots a subtree does not need stay unused).  par: thread-doubling levels still available; a
  // node that still has some of them also bins its four axes on four threads (the top of the tree is where the time is).
  int build(int first, int count, int depth_left = 1 << 20) {
    unsigned hw = std::thread::hardware_concurrency();
    if (const char* e = std::getenv("B200RT_BVH_THREADS")) hw = (unsigned)std::max(1, std::atoi(e));
    int par = 0;
    while ((1u << par) < std::min(std::max(hw, 1u), 16u)) ++par;
    nodes.assign((size_t)2 * count, BinNode());
    const int root = build(first, count, depth_left, par, 0);
    order.resize(items.size());
    for (size_t i = 0; i < items.size(); ++i) order[i] = items[i].id;
    return root;
  }
  static const int NBMAX = 32;
  // best split of [first, first + count) along one axis: cost and bin, or FLT_MAX
  void best_on_axis(int first, int count, int axis, float klo, float khi, float& cost_out, int& split_out) const {
    cost_out = FLT_MAX; split_out = -1;
    const int NB = n_bins;
    const float ext = khi - klo;
    if (!(ext > 0)) return;
    if (axis == 3 && ext < 2.f) return;  // sizes within a factor of four: nothing to separate
    // (three quarters of a tree's nodes hold <= 8 primitives: bins are created on first touch and the sweeps visit only the
    //  occupied ones — between two occupied bins every split index gives the same partition and the first of them, the one a
    //  sweep over all bins would keep under its strict <, is the occupied bin itself)
    struct RawBox { float lo[3], hi[3]; };
    RawBox bins[NBMAX];
    int cnt[NBMAX];
    for (int b = 0; b < NB; ++b) cnt[b] = 0;
    const float k = NB * (1 - 1e-6f) / ext;
    for (int i = first; i < first + count; ++i) {
      const int bi = std::min(NB - 1, std::max(0, (int)((items[i].key[axis] - klo) * k)));
      const Box& ib = items[i].box;
      RawBox& bb = bins[bi];
      if (cnt[bi]++ == 0) { for (int a = 0; a < 3; ++a) { bb.lo[a] = ib.lo[a]; bb.hi[a] = ib.hi[a]; } }
      else { for (int a = 0; a < 3; ++a) { bb.lo[a] = std::min(bb.lo[a], ib.lo[a]); bb.hi[a] = std::max(bb.hi[a], ib.hi[a]); } }
    }
    int nz[NBMAX], m = 0;
    for (int b = 0; b < NB; ++b) if (cnt[b]) nz[m++] = b;
    if (m < 2) return;
    auto grow_raw = [](Box& acc, const RawBox& r) { for (int a = 0; a < 3; ++a) { acc.lo[a] = std::min(acc.lo[a], r.lo[a]); acc.hi[a] = std::max(acc.hi[a], r.hi[a]); } };
    float right_area[NBMAX]; int right_cnt[NBMAX];  // of the occupied bins nz[j..m)
    Box acc; int c = 0;
    for (int j = m - 1; j > 0; --j) { grow_raw(acc, bins[nz[j]]); c += cnt[nz[j]]; right_area[j] = acc.area(); right_cnt[j] = c; }
    Box lacc; int lc = 0;
    for (int j = 0; j < m - 1; ++j) {
      grow_raw(lacc, bins[nz[j]]); lc += cnt[nz[j]];
      const float cost = lacc.area() * lc + right_area[j + 1] * right_cnt[j + 1];
      if (cost < cost_out) { cost_out = cost; split_out = nz[j]; }
    }
  }
  int build(int first, int count, int depth_left, int par, int id) {
    Box box;
    float klo[4] = {FLT_MAX, FLT_MAX, FLT_MAX, FLT_MAX}, khi[4] = {-FLT_MAX, -FLT_MAX, -FLT_MAX, -FLT_MAX};
    for (int i = first; i < first + count; ++i) {
      box.grow(items[i].box);
      for (int a = 0; a < 4; ++a) { klo[a] = std::min(klo[a], items[i].key[a]); khi[a] = std::max(khi[a], items[i].key[a]); }
    }
    nodes[id].box = box;
    nodes[id].first = first; nodes[id].count = count;  // inner nodes too: the size of the subtree (dense_box below)
    if (count == 1) return id;
    // binned SAH over the three centroid axes and, as a fourth "axis", the primitives' size (-log2 of the box area: large first).
    // A room of a dozen walls around a 300,000-triangle mesh has no good spatial split near the root — every plane through the
    // mesh leaves walls on both sides, so both halves keep the room's box and every ray descends both — but it has a very good
    // split by size: the walls | the mesh.  Same cost function, so the size split is only taken where it is the cheaper one.
    const int NB = n_bins;
    const int n_axes = size_split ? 4 : 3;
    float axis_cost[4]; int axis_split[4];
    if (par > 0 && count >= 32768) {
      std::future<void> f[3];
      for (int a = 1; a < n_axes; ++a) {
        try { f[a - 1] = std::async(std::launch::async, [&, a] { best_on_axis(first, count, a, klo[a], khi[a], axis_cost[a], axis_split[a]); }); }
        catch (const std::system_error&) {}  // no thread to be had: this axis on the calling thread, below
      }
      best_on_axis(first, count, 0, klo[0], khi[0], axis_cost[0], axis_split[0]);
      for (int a = 1; a < n_axes; ++a) { if (f[a - 1].valid()) f[a - 1].get(); else best_on_axis(first, count, a, klo[a], khi[a], axis_cost[a], axis_split[a]); }
    } else {
      for (int a = 0; a < n_axes; ++a) best_on_axis(first, count, a, klo[a], khi[a], axis_cost[a], axis_split[a]);
    }
    float best_cost = FLT_MAX; int best_axis = -1, best_split = -1;
    for (int a = 0; a < n_axes; ++a) if (axis_cost[a] < best_cost) { best_cost = axis_cost[a]; best_axis = a; best_split = axis_split[a]; }
    float leaf_cost = box.area() * count;
    if (count <= B200RT_MAX_LEAF && (best_axis < 0 || best_cost + trav_cost * box.area() >= leaf_cost)) return id;
    const bool force_median = depth_left < 31 && (long long)count > (1ll << std::max(0, depth_left - 1));
    int mid;
    if (force_median) {
      int axis = 0;
      for (int a = 1; a < 3; ++a) if (khi[a] - klo[a] > khi[axis] - klo[axis]) axis = a;
      mid = first + count / 2;
      std::nth_element(items.begin() + first, items.begin() + mid, items.begin() + first + count, [&](const Item& a, const Item& b) { return a.key[axis] < b.key[axis]; });
    } else if (best_axis < 0) {  // all centroids coincide: split the range in the middle
      mid = first + count / 2;
    } else {
      const float ext = khi[best_axis] - klo[best_axis];
      const float k = NB * (1 - 1e-6f) / ext;
      auto it = std::partition(items.begin() + first, items.begin() + first + count, [&](const Item& p) {
        const int bi = std::min(NB - 1, std::max(0, (int)((p.key[best_axis] - klo[best_axis]) * k)));
        return bi <= best_split;
      });
      mid = (int)(it - items.begin());
      if (mid == first || mid == first + count) mid = first + count / 2;
    }
    const int nl = mid - first, nr = first + count - mid, lid = id + 1, rid = id + 2 * nl;
    int l, r;
    if (par > 0 && std::min(nl, nr) >= 4096) {
      std::future<int> left;
      try { left = std::async(std::launch::async, [&] { return build(first, nl, depth_left - 1, par - 1, lid); }); }
      catch (const std::system_error&) {}
      r = build(mid, nr, depth_left - 1, par - 1, rid);
      l = left.valid() ? left.get() : build(first, nl, depth_left - 1, 0, lid);
    } else {  // (a lopsided split — seven walls | the mesh — keeps the thread budget for the levels below)
      l = build(first, nl, depth_left - 1, par, lid);
      r = build(mid, nr, depth_left - 1, par, rid);
    }
    nodes[id].left = l; nodes[id].right = r;
    return id;
  }
};

// The box of the scene's dense part, if it has one: from the root down, follow the child that keeps >= 90 % of the primitives
// for as long as there is one (the size split of build() puts a room's few large walls on one side of the root and a 300,000-
// triangle mesh on the other).  Accepted when what is left is at most half of the root's surface: a ray then either crosses that
// box — a long traversal — or is done after a few node steps, and the wavefront queues keep the two kinds apart
// (WfBuffers::split_lo, wavefront.cuh).
bool dense_box(const BinBuilder& bb, int root, Box& out) {
  const int total = bb.nodes[root].count;
  int pct = 90;  // B200RT_BVH_DENSE_PCT: the share of the primitives the box must keep (A/B of the predictor)
  if (const char* e = std::getenv("B200RT_BVH_DENSE_PCT")) pct = std::min(100, std::max(50, std::atoi(e)));
  int cur = root;
  for (;;) {
    const BinNode& n = bb.nodes[cur];
    if (n.left < 0) break;
    const int l = bb.nodes[n.left].count, r = bb.nodes[n.right].count;
    const int next = l >= r ? n.left : n.right;
    if ((long long)std::max(l, r) * 100 < (long long)total * pct) break;
    cur = next;
  }
  if (cur == root) return false;
  out = bb.nodes[cur].box;
  return out.area() <= 0.5f * bb.nodes[root].box.area();
}

void put4(std::vector<float>& v, size_t at, float a, float b, float c, float d) { v[at] = a; v[at + 1] = b; v[at + 2] = c; v[at + 3] = d; }
float ibits(int32_t i) { float f; std::memcpy(&f, &i, 4); return f; }
float ubits(uint32_t i) { float f; std::memcpy(&f, &i, 4); return f; }

// Collapse the binary tree into BVH4 nodes (128 B each); returns the node index.  depth_out: levels of BVH4 nodes below
// and including this one (the traversal pushes at most 3 entries per level: node_step in intersect.cuh).
// sah[0] += area of this node's box, sah[1] += area x primitives of every leaf child: divided by the root's area these are the node
// steps and primitive tests a random line through the scene pays (the SAH estimate b200rt_scene_bvh_quality reports).
int collapse(const BinBuilder& bb, int bin_id, std::vector<float>& out, int& depth_out, double* sah) {
  int my = (int)(out.size() / 32);
  int deepest = 0;
  out.resize(out.size() + 32, 0.f);
  int kids[4]; int nk = 0;
  const BinNode& root = bb.nodes[bin_id];
  sah[0] += root.box.area();
  if (root.left < 0) { kids[nk++] = bin_id; }
  else {
    kids[nk++] = root.left; kids[nk++] = root.right;
    while (nk < 4) {
      int best = -1; float best_area = -1;
      for (int i = 0; i < nk; ++i) {
        const BinNode& c = bb.nodes[kids[i]];
        if (c.left >= 0 && c.box.area() > best_area) { best_area = c.box.area(); best = i; }
      }
      if (best < 0) break;
      int c = kids[best];
      kids[best] = bb.nodes[c].left;
      kids[nk++] = bb.nodes[c].right;
    }
  }
  int32_t refs[4] = {B200RT_CHILD_EMPTY, B200RT_CHILD_EMPTY, B200RT_CHILD_EMPTY, B200RT_CHILD_EMPTY};
  float lo[3][4], hi[3][4];
  for (int a = 0; a < 3; ++a) for (int i = 0; i < 4; ++i) { lo[a][i] = INFINITY; hi[a][i] = -INFINITY; }  // empty slot: near = +inf, far = -inf for any finite ray, never hit
  for (int i = 0; i < nk; ++i) {
    const BinNode& c = bb.nodes[kids[i]];
    for (int a = 0; a < 3; ++a) { lo[a][i] = c.box.lo[a]; hi[a][i] = c.box.hi[a]; }
    if (c.left < 0) { refs[i] = (int32_t)(B200RT_LEAF_BIT | ((uint32_t)(c.count - 1) << 27) | (uint32_t)c.first); sah[1] += (double)c.box.area() * c.count; }
    else { int dk = 0; refs[i] = collapse(bb, kids[i], out, dk, sah); deepest = std::max(deepest, dk); }
  }
  depth_out = deepest + 1;
  size_t at = (size_t)my * 32;
  // rows: lo.x hi.x lo.y hi.y lo.z hi.z refs (pad) — the near plane of an axis is row 2a + sign(d_a), the far one its xor 1
  for (int a = 0; a < 3; ++a) { put4(out, at + 8 * a, lo[a][0], lo[a][1], lo[a][2], lo[a][3]); put4(out, at + 8 * a + 4, hi[a][0], hi[a][1], hi[a][2], hi[a][3]); }
  put4(out, at + 24, ibits(refs[0]), ibits(refs[1]), ibits(refs[2]), ibits(refs[3]));
  return my;
}

}  // namespace

static int compile_scene_impl(const b200rt_scene_desc* d, CompiledScene& cs, std::string& err) {
  if (!d) { err = "null scene desc"; return B200RT_ERR_INVALID; }
  if (d->n_prims == 0) { err = "scene has no primitives (BVHNode::new on an empty list)"; return B200RT_ERR_INVALID; }
  if (d->n_prims >= (1u << 27)) { err = "too many primitives"; return B200RT_ERR_UNSUPPORTED; }
  if (!d->prims || !d->materials || !d->textures) { err = "null table pointer"; return B200RT_ERR_INVALID; }
  const uint32_t n = d->n_prims;
  // B200RT_COMPILE_TIMING=1: the stages' host times on stderr
  const bool timing = std::getenv("B200RT_COMPILE_TIMING") != nullptr;
  auto t_prev = std::chrono::steady_clock::now();
  auto stage = [&](const char* what) {
    if (!timing) return;
    const auto now = std::chrono::steady_clock::now();
    std::fprintf(stderr, "compile_scene: %-28s %8.2f ms\n", what, std::chrono::duration<double, std::milli>(now - t_prev).count());
    t_prev = now;
  };
  // ---- validate tables
  for (uint32_t i = 0; i < d->n_textures; ++i) {
    const b200rt_texture& t = d->textures[i];
    if (t.type < 0 || t.type > B200RT_TEX_NOISE) { err = "unknown texture type"; return B200RT_ERR_INVALID; }
    if (t.type == B200RT_TEX_IMAGE) {
      if (t.image < 0 || (uint32_t)t.image >= d->n_images || !d->images || !d->image_rgb8) { err = "image index out of range"; return B200RT_ERR_INVALID; }
      const b200rt_image& im = d->images[t.image];
      if (im.width <= 0 || im.height <= 0 || im.offset < 0 || (uint64_t)im.offset + (uint64_t)3 * im.width * im.height > d->image_bytes) { err = "image outside the blob"; return B200RT_ERR_INVALID; }
    }
    if (t.type == B200RT_TEX_NOISE && !d->perlin) { err = "NOISE texture without perlin tables"; return B200RT_ERR_INVALID; }
  }
  for (uint32_t i = 0; i < d->n_materials; ++i) {
    const b200rt_material& m = d->materials[i];
    if (m.type < 0 || m.type > B200RT_MAT_ISOTROPIC) { err = "unknown material type"; return B200RT_ERR_INVALID; }
    if (m.texture < 0 || (uint32_t)m.texture >= d->n_textures) { err = "texture index out of range"; return B200RT_ERR_INVALID; }
  }
  // ---- instances: linear parts of M and M^-1
  std::vector<Affine> M(d->n_instances), Minv(d->n_instances);
  cs.instances.assign((size_t)std::max(1u, d->n_instances) * 24, 0.f);
  cs.instances_d.assign((size_t)std::max(1u, d->n_instances) * 24, 0.0);
  cs.inst_eps.assign((size_t)std::max(1u, d->n_instances), 0.f);
  for (uint32_t i = 0; i < d->n_instances; ++i) {
    std::memcpy(M[i].m, d->instances[i].m, sizeof(M[i].m));
    if (!M[i].inverse(Minv[i])) { err = "Invalid transform matrix"; return B200RT_ERR_INVALID; }  // transform.rs:146
    std::memcpy(&cs.instances_d[(size_t)i * 24], M[i].m, sizeof(M[i].m));
    std::memcpy(&cs.instances_d[(size_t)i * 24 + 12], Minv[i].m, sizeof(Minv[i].m));
    for (int r = 0; r < 3; ++r) {
      put4(cs.instances, (size_t)i * 24 + 4 * r, (float)M[i].m[4 * r], (float)M[i].m[4 * r + 1], (float)M[i].m[4 * r + 2], 0);
      put4(cs.instances, (size_t)i * 24 + 12 + 4 * r, (float)Minv[i].m[4 * r], (float)Minv[i].m[4 * r + 1], (float)Minv[i].m[4 * r + 2], 0);
    }
  }
  stage("validate + instances");
  // ---- bake primitives to world space
  // Transform::new(rotate, scale: Vec3, translate, obj) (transform.rs:17-148) is accepted in full generality.  Where the
  // shape survives the transform (a sphere under a similarity; an axis rectangle / Cube under a positive diagonal; any
  // triangle) the f32 filter record is baked to world space; otherwise (B200RT_PRIM_OBJF: rotated Cube / rectangle / box
  // medium, sphere under a non-uniform scale) the record stays in object space, the device filters on the object-space
  // ray, and only the bounding box is transformed (the eight corners of the object-space box, transform.rs:110-136).
  std::vector<float> data((size_t)n * 12, 0.f), objdata((size_t)n * 12, 0.f);
  std::vector<Box> boxes(n);
  std::vector<uint8_t> objf(n, 0);
  auto grow_transformed = [](Box& bx, const Affine& A, const double lo[3], const double hi[3]) {
    for (int c = 0; c < 8; ++c) bx.grow(A.point(D3((c & 1) ? hi[0] : lo[0], (c & 2) ? hi[1] : lo[1], (c & 4) ? hi[2] : lo[2])));
  };
  for (uint32_t i = 0; i < n; ++i) {
    const b200rt_prim& p = d->prims[i];
    if (p.material < 0 || (uint32_t)p.material >= d->n_materials) { err = "material index out of range"; return B200RT_ERR_INVALID; }
    if (p.instance >= (int32_t)d->n_instances) { err = "instance index out of range"; return B200RT_ERR_INVALID; }
    const bool inst = p.instance >= 0;
    const Affine A = inst ? M[p.instance] : Affine::identity();
    float* o = &data[(size_t)i * 12];
    float* q = &objdata[(size_t)i * 12];  // object-space record in the same intermediate layout (the f64 decision runs on it)
    Box& bx = boxes[i];
    switch (p.type) {
      case B200RT_PRIM_SPHERE: case B200RT_PRIM_MEDIUM_SPHERE: case B200RT_PRIM_MOVING_SPHERE: {
        // sphere.rs:24-61 accepts a negative radius (the outward normal (p - c) / r turns inwards); zero divides by zero
        if (!(p.p[3] != 0.f) || !std::isfinite(p.p[3])) { err = "sphere radius must be non-zero"; return B200RT_ERR_INVALID; }
        if (p.type == B200RT_PRIM_MOVING_SPHERE && !(p.p[8] > p.p[7])) { err = "moving sphere needs time1 > time0"; return B200RT_ERR_INVALID; }
        if (p.type == B200RT_PRIM_MEDIUM_SPHERE && !(p.p[4] > 0)) { err = "medium density must be positive"; return B200RT_ERR_INVALID; }
        for (int k = 0; k < 9; ++k) q[k] = p.p[k];
        if (p.type == B200RT_PRIM_MEDIUM_SPHERE) { for (int k = 4; k < 9; ++k) q[k] = 0.f; q[4] = (float)(-1.0 / (double)p.p[4]); }
        const double s = inst ? A.similarity_scale() : 1.0;
        if (s == 0) {  // ellipsoid in world space: object-space filter
          objf[i] = 1;
          const double ra = std::fabs((double)p.p[3]) * (1 + 1e-6);
          double lo[3], hi[3];
          for (int a = 0; a < 3; ++a) { lo[a] = p.p[a] - ra; hi[a] = p.p[a] + ra; }
          if (p.type == B200RT_PRIM_MOVING_SPHERE) for (int a = 0; a < 3; ++a) { lo[a] = std::min(lo[a], p.p[4 + a] - ra); hi[a] = std::max(hi[a], p.p[4 + a] + ra); }
          grow_transformed(bx, A, lo, hi);
          break;
        }
        D3 c = A.point(D3(p.p));
        double r = s * p.p[3];
        o[0] = (float)c.x; o[1] = (float)c.y; o[2] = (float)c.z; o[3] = (float)r;
        double rr = std::fabs((double)o[3]) * (1 + 1e-6);
        bx.grow(D3(o[0] - rr, o[1] - rr, o[2] - rr)); bx.grow(D3(o[0] + rr, o[1] + rr, o[2] + rr));
        if (p.type == B200RT_PRIM_MOVING_SPHERE) {
          D3 c1 = A.point(D3(p.p + 4));
          float c1f[3] = {(float)c1.x, (float)c1.y, (float)c1.z};
          o[4] = c1f[0]; o[5] = c1f[1]; o[6] = c1f[2]; o[7] = p.p[7]; o[8] = p.p[8];
          bx.grow(D3(c1f[0] - rr, c1f[1] - rr, c1f[2] - rr)); bx.grow(D3(c1f[0] + rr, c1f[1] + rr, c1f[2] + rr));
        } else if (p.type == B200RT_PRIM_MEDIUM_SPHERE) {
          // a distance of d object units is s*d world units: density scales by 1/s
          o[4] = (float)(-s / (double)p.p[4]);
        }
        break;
      }
      case B200RT_PRIM_XY_RECT: case B200RT_PRIM_XZ_RECT: case B200RT_PRIM_YZ_RECT: {
        const int axis = p.type == B200RT_PRIM_XY_RECT ? 2 : (p.type == B200RT_PRIM_XZ_RECT ? 1 : 0);
        const int ia = axis == 0 ? 1 : 0, ib = axis == 2 ? 1 : 2;
        if (!(p.p[2] > p.p[0]) || !(p.p[3] > p.p[1])) { err = "degenerate rectangle"; return B200RT_ERR_INVALID; }
        for (int k = 0; k < 5; ++k) q[k] = p.p[k];
        double lo[3], hi[3];
        lo[ia] = p.p[0]; lo[ib] = p.p[1]; hi[ia] = p.p[2]; hi[ib] = p.p[3]; lo[axis] = hi[axis] = p.p[4];
        if (inst && !A.is_positive_diagonal()) { objf[i] = 1; grow_transformed(bx, A, lo, hi); break; }
        D3 l = A.point(D3(lo)), h = A.point(D3(hi));
        o[0] = (float)l[ia]; o[1] = (float)l[ib]; o[2] = (float)h[ia]; o[3] = (float)h[ib]; o[4] = (float)l[axis];
        if (!(o[2] > o[0]) || !(o[3] > o[1])) { err = "degenerate rectangle"; return B200RT_ERR_INVALID; }
        double q0[3], q1[3];
        q0[ia] = o[0]; q0[ib] = o[1]; q0[axis] = o[4]; q1[ia] = o[2]; q1[ib] = o[3]; q1[axis] = o[4];
        bx.grow(D3(q0)); bx.grow(D3(q1));
        break;
      }
      case B200RT_PRIM_BOX: case B200RT_PRIM_MEDIUM_BOX: {
        if (!(p.p[3] > p.p[0]) || !(p.p[4] > p.p[1]) || !(p.p[5] > p.p[2])) { err = "degenerate box"; return B200RT_ERR_INVALID; }
        if (p.type == B200RT_PRIM_MEDIUM_BOX && !(p.p[6] > 0)) { err = "medium density must be positive"; return B200RT_ERR_INVALID; }
        q[0] = p.p[0]; q[1] = p.p[1]; q[2] = p.p[2]; q[4] = p.p[3]; q[5] = p.p[4]; q[6] = p.p[5];
        if (p.type == B200RT_PRIM_MEDIUM_BOX) q[7] = (float)(-1.0 / (double)p.p[6]);
        const bool uniform = A.m[0] == A.m[5] && A.m[5] == A.m[10];
        if (inst && (!A.is_positive_diagonal() || (p.type == B200RT_PRIM_MEDIUM_BOX && !uniform))) {
          objf[i] = 1;
          const double lo[3] = {p.p[0], p.p[1], p.p[2]}, hi[3] = {p.p[3], p.p[4], p.p[5]};
          grow_transformed(bx, A, lo, hi);
          break;
        }
        D3 l = A.point(D3(p.p)), h = A.point(D3(p.p + 3));
        o[0] = (float)l.x; o[1] = (float)l.y; o[2] = (float)l.z; o[4] = (float)h.x; o[5] = (float)h.y; o[6] = (float)h.z;
        if (!(o[4] > o[0]) || !(o[5] > o[1]) || !(o[6] > o[2])) { err = "degenerate box"; return B200RT_ERR_INVALID; }
        if (p.type == B200RT_PRIM_MEDIUM_BOX) o[7] = (float)(-(inst ? A.m[0] : 1.0) / (double)p.p[6]);
        bx.grow(D3(o[0], o[1], o[2])); bx.grow(D3(o[4], o[5], o[6]));
        break;
      }
      case B200RT_PRIM_TRIANGLE: {
        if (p.aux < 0 || (uint32_t)p.aux >= d->n_tri_normals || !d->tri_normals) { err = "triangle without vertex normals (mesh.rs:62-64 would panic)"; return B200RT_ERR_INVALID; }
        for (int v = 0; v < 3; ++v) {
          D3 w = A.point(D3(p.p + 3 * v));
          o[4 * v] = (float)w.x; o[4 * v + 1] = (float)w.y; o[4 * v + 2] = (float)w.z;
          bx.grow(D3(o[4 * v], o[4 * v + 1], o[4 * v + 2]));
          for (int c = 0; c < 3; ++c) q[4 * v + c] = p.p[3 * v + c];
        }
        break;
      }
      default: err = "unknown primitive type"; return B200RT_ERR_INVALID;
    }
    if (objf[i]) { std::memcpy(o, q, 12 * sizeof(float)); cs.has_objf = true; }  // the filter record of an OBJF primitive is its object-space record
    for (int a = 0; a < 3; ++a) if (!std::isfinite(bx.lo[a]) || !std::isfinite(bx.hi[a])) { err = "non-finite primitive"; return B200RT_ERR_INVALID; }
    if (inst && !objf[i]) {
      float m = 0;
      for (int a = 0; a < 3; ++a) m = std::max(m, std::max(std::fabs(bx.lo[a]), std::fabs(bx.hi[a])));
      cs.inst_eps[p.instance] = std::max(cs.inst_eps[p.instance], 1.3e-7f * m);  // ~2u |coord|: rounding of M*x to f32 (+ f64 slop)
    }
    bx.pad();
  }
  stage("bake primitives");
  // ---- BVH (skipped for the flat-list arm: hit.rs:56-67 scans linearly)
  cs.flat = (d->flags & B200RT_SCENE_FLAT_LIST) != 0;
  std::vector<int> order(n);
  std::iota(order.begin(), order.end(), 0);
  cs.nodes.clear();
  cs.n_bvh_prims = n;
  if (!cs.flat) {
    // scene-enclosing primitives (box area > half of the scene box) stay out of the BVH: every ray hits their box at
    // t_near = 0 anyway; the device tests them once, AFTER the traversal, when the closest hit found so far lets
    // the f32 filters reject most of them (book-2's fog sphere: ~95 % of the rays).
    Box all;
    for (uint32_t i = 0; i < n; ++i) all.grow(boxes[i]);
    std::vector<int> in_idx, out_idx;
#ifdef B200RT_NO_OUTLIERS
    const bool no_outliers = true;
#else
    const bool no_outliers = false;
#endif
    for (uint32_t i = 0; i < n; ++i) ((!no_outliers && n >= 8 && boxes[i].area() > 0.5f * all.area()) ? out_idx : in_idx).push_back((int)i);
    if (in_idx.empty() || out_idx.size() > 8) { in_idx.resize(n); std::iota(in_idx.begin(), in_idx.end(), 0); out_idx.clear(); }
    std::vector<Box> inboxes(in_idx.size());
    for (size_t k = 0; k < in_idx.size(); ++k) inboxes[k] = boxes[in_idx[k]];
    // The device traversal keeps a per-thread stack of B200RT_STACK entries and pushes at most 3 per BVH4 level: the tree
    // is accepted when 3 * depth <= B200RT_STACK, else rebuilt under a tighter binary depth budget (the last budget,
    // ceil(log2 n) + 1, is a median-split tree whose collapse is about half that deep).
    int min_budget = 2;
    while ((1ll << (min_budget - 1)) < (long long)in_idx.size()) ++min_budget;
    const int budgets[5] = {1 << 20, 48, 36, 28, min_budget};
    std::vector<int> bb_order;
    for (int attempt = 0; attempt < 5; ++attempt) {
      if (attempt > 0 && budgets[attempt] < min_budget) continue;
      stage("  in / out of the BVH");
      BinBuilder bb(inboxes);
      stage("  builder records");
      int root = bb.build(0, (int)in_idx.size(), budgets[attempt]);
      stage("  binary SAH build");
      cs.nodes.clear();
      cs.nodes.reserve(in_idx.size() * 16);
      int depth4 = 0;
      double sah[2] = {0., 0.};
      collapse(bb, root, cs.nodes, depth4, sah);
      cs.bvh4_depth = (uint32_t)depth4;
      const double root_area = std::max((double)bb.nodes[root].box.area(), 1e-300);
      cs.sah_node_steps = (float)(sah[0] / root_area); cs.sah_prim_tests = (float)(sah[1] / root_area);
      Box dense;
      cs.has_dense_box = dense_box(bb, root, dense);
      for (int a = 0; a < 3; ++a) { cs.dense_lo[a] = cs.has_dense_box ? dense.lo[a] : 0.f; cs.dense_hi[a] = cs.has_dense_box ? dense.hi[a] : 0.f; }
      bb_order.swap(bb.order);
      if (3 * depth4 <= B200RT_STACK) break;
    }
    if (3 * (int)cs.bvh4_depth > B200RT_STACK) { err = "BVH deeper than the traversal stack allows"; return B200RT_ERR_UNSUPPORTED; }
    for (size_t k = 0; k < in_idx.size(); ++k) order[k] = in_idx[bb_order[k]];
    for (size_t k = 0; k < out_idx.size(); ++k) order[in_idx.size() + k] = out_idx[k];
    cs.n_bvh_prims = (uint32_t)in_idx.size();
  } else {
    cs.nodes.assign(32, 0.f);
    cs.n_bvh_prims = 0;  // the device feeds every primitive through the linear tail of trace()
  }
  cs.n_nodes = (uint32_t)(cs.nodes.size() / 32);
  cs.n_prims = n;
  cs.prim_data.assign((size_t)n * 12, 0.f);
  cs.prim_obj.assign(d->n_instances ? (size_t)n * 12 : 12, 0.f);
  cs.prim_meta.resize((size_t)n * 4);
  std::vector<int> dev_of(n);
  auto pack = [](std::vector<float>& out, size_t k, int type, const float* src) {
    float* dst = &out[k * 12];
    switch (type) {
      case B200RT_PRIM_SPHERE: put4(out, k * 12, src[0], src[1], src[2], src[3]); break;
      case B200RT_PRIM_MOVING_SPHERE:
        put4(out, k * 12, src[0], src[1], src[2], src[3]);
        put4(out, k * 12 + 4, src[4], src[5], src[6], 0);  // c1 (absolute)
        dst[8] = src[7]; dst[9] = src[8];
        break;
      case B200RT_PRIM_MEDIUM_SPHERE: put4(out, k * 12, src[0], src[1], src[2], src[3]); dst[4] = src[4]; break;
      case B200RT_PRIM_XY_RECT: case B200RT_PRIM_XZ_RECT: case B200RT_PRIM_YZ_RECT:
        put4(out, k * 12, src[0], src[1], src[2], src[3]); dst[4] = src[4]; break;
      case B200RT_PRIM_BOX: case B200RT_PRIM_MEDIUM_BOX:
        put4(out, k * 12, src[0], src[1], src[2], 0); put4(out, k * 12 + 4, src[4], src[5], src[6], src[7]); break;
      case B200RT_PRIM_TRIANGLE: std::memcpy(dst, src, 12 * sizeof(float)); break;
    }
  };
  for (uint32_t k = 0; k < n; ++k) {
    int i = order[k];
    dev_of[i] = (int)k;
    const b200rt_prim& p = d->prims[i];
    pack(cs.prim_data, k, p.type, &data[(size_t)i * 12]);
    if (p.instance >= 0) pack(cs.prim_obj, k, p.type, &objdata[(size_t)i * 12]);
    cs.prim_meta[(size_t)k * 4 + 0] = p.type | (objf[i] ? B200RT_PRIM_OBJF : 0) | (p.material << 8);
    cs.prim_meta[(size_t)k * 4 + 1] = i;
    cs.prim_meta[(size_t)k * 4 + 2] = p.instance;
    cs.prim_meta[(size_t)k * 4 + 3] = p.aux;
  }
  stage("BVH build + collapse");
  // ---- the remaining tables
  cs.tri_normals.assign((size_t)std::max(1u, d->n_tri_normals) * 12, 0.f);
  for (uint32_t i = 0; i < d->n_tri_normals; ++i)
    for (int v = 0; v < 3; ++v) put4(cs.tri_normals, (size_t)i * 12 + 4 * v, d->tri_normals[9 * (size_t)i + 3 * v], d->tri_normals[9 * (size_t)i + 3 * v + 1], d->tri_normals[9 * (size_t)i + 3 * v + 2], 0);
  cs.materials.assign((size_t)d->n_materials * 4, 0.f);
  for (uint32_t i = 0; i < d->n_materials; ++i) put4(cs.materials, (size_t)i * 4, ibits(d->materials[i].type), ibits(d->materials[i].texture), d->materials[i].param, 0);
  cs.textures.assign((size_t)d->n_textures * 12, 0.f);
  for (uint32_t i = 0; i < d->n_textures; ++i) {
    const b200rt_texture& t = d->textures[i];
    put4(cs.textures, (size_t)i * 12, ibits(t.type), ibits(t.image), t.scale, 0);
    put4(cs.textures, (size_t)i * 12 + 4, t.c0[0], t.c0[1], t.c0[2], 0);
    put4(cs.textures, (size_t)i * 12 + 8, t.c1[0], t.c1[1], t.c1[2], 0);
  }
  cs.images.assign((size_t)std::max(1u, d->n_images) * 4, 0);
  for (uint32_t i = 0; i < d->n_images; ++i) { cs.images[4 * i] = d->images[i].offset; cs.images[4 * i + 1] = d->images[i].width; cs.images[4 * i + 2] = d->images[i].height; }
  cs.image_data.assign(d->image_rgb8, d->image_rgb8 + d->image_bytes);
  if (cs.image_data.empty()) cs.image_data.resize(4, 0);
  cs.perlin_vec.assign(256 * 4, 0.f);
  cs.perlin_perm.assign(3 * 256, 0);
  if (d->perlin) {
    for (int i = 0; i < 256; ++i) put4(cs.perlin_vec, (size_t)i * 4, d->perlin->ranvec[i][0], d->perlin->ranvec[i][1], d->perlin->ranvec[i][2], 0);
    for (int i = 0; i < 256; ++i) { cs.perlin_perm[i] = d->perlin->perm_x[i] & 255; cs.perlin_perm[256 + i] = d->perlin->perm_y[i] & 255; cs.perlin_perm[512 + i] = d->perlin->perm_z[i] & 255; }
  }
  cs.n_lights = d->n_lights;
  cs.lights.assign((size_t)std::max(1u, d->n_lights) * 12, 0.f);
  for (uint32_t i = 0; i < d->n_lights; ++i) {
    int pi = d->light_prims[i];
    if (pi < 0 || (uint32_t)pi >= n) { err = "light index out of range"; return B200RT_ERR_INVALID; }
    const b200rt_prim& p = d->prims[pi];
    if (p.instance >= 0) { err = "transformed lights are not sampled"; return B200RT_ERR_UNSUPPORTED; }
    size_t at = (size_t)i * 12;
    if (p.type == B200RT_PRIM_SPHERE) {
      put4(cs.lights, at, ibits(0), ibits(pi), ibits(0), ibits(p.material));
      put4(cs.lights, at + 4, p.p[0], p.p[1], p.p[2], p.p[3]);
    } else if (p.type == B200RT_PRIM_XY_RECT || p.type == B200RT_PRIM_XZ_RECT || p.type == B200RT_PRIM_YZ_RECT) {
      int axis = p.type == B200RT_PRIM_XY_RECT ? 2 : (p.type == B200RT_PRIM_XZ_RECT ? 1 : 0);
      put4(cs.lights, at, ibits(1), ibits(pi), ibits(axis), ibits(p.material));
      put4(cs.lights, at + 4, p.p[0], p.p[1], p.p[2], p.p[3]);
      put4(cs.lights, at + 8, p.p[4], (p.p[2] - p.p[0]) * (p.p[3] - p.p[1]), 0, 0);
    } else { err = "only sphere and axis-rectangle lights are sampled"; return B200RT_ERR_UNSUPPORTED; }
    const float scale = d->light_scales ? d->light_scales[i] : 1.0f;  // light.rs:71,131, relative to b200rt_sppm_params.light_scale
    if (!(scale > 0.f) || !std::isfinite(scale)) { err = "light scale must be positive and finite"; return B200RT_ERR_INVALID; }
    cs.lights[at + 10] = scale;  // row 2, .z (row 2 of a sphere light is otherwise unused; of a rectangle: k, area)
  }
  stage("tables in leaf order");
  (void)ubits;
  return B200RT_OK;
}

// No exception leaves the library: the callers are extern "C" (a scene of 10^8 primitives that does not fit the host's memory
// is an error code, not a terminate()).
int compile_scene(const b200rt_scene_desc* d, CompiledScene& cs, std::string& err) {
  try {
    return compile_scene_impl(d, cs, err);
  } catch (const std::bad_alloc&) {
    err = "out of host memory while compiling the scene";
    return B200RT_ERR_NOMEM;
  } catch (const std::exception& e) {
    err = std::string("scene compile failed: ") + e.what();
    return B200RT_ERR_INVALID;
  }
}

}  // namespace b200rt

// Host-only introspection of World::new's BVH (no CUDA needed): what b200rt_scene_create would build for `desc`.
extern "C" int b200rt_scene_compile_stats(const b200rt_scene_desc* desc, uint32_t* bvh_nodes, uint32_t* bvh4_depth, uint32_t* stack_needed, uint32_t* stack_capacity) {
  b200rt::CompiledScene cs;
  std::string err;
  const int rc = b200rt::compile_scene(desc, cs, err);
  if (rc) return rc;
  if (bvh_nodes) *bvh_nodes = cs.n_nodes;
  if (bvh4_depth) *bvh4_depth = cs.bvh4_depth;
  if (stack_needed) *stack_needed = 3 * cs.bvh4_depth;
  if (stack_capacity) *stack_capacity = B200RT_STACK;
  return B200RT_OK;
}

extern "C" int b200rt_scene_bvh_quality(const b200rt_scene_desc* desc, float* node_steps, float* prim_tests, int32_t* has_dense_box, float* dense_lo, float* dense_hi) {
  b200rt::CompiledScene cs;
  std::string err;
  const int rc = b200rt::compile_scene(desc, cs, err);
  if (rc) return rc;
  if (node_steps) *node_steps = cs.sah_node_steps;
  if (prim_tests) *prim_tests = cs.sah_prim_tests;
  if (has_dense_box) *has_dense_box = cs.has_dense_box ? 1 : 0;
  for (int a = 0; a < 3; ++a) { if (dense_lo) dense_lo[a] = cs.dense_lo[a]; if (dense_hi) dense_hi[a] = cs.dense_hi[a]; }
  return B200RT_OK;
}
