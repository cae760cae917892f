// loaders.cpp — host-side file front ends: Mesh::load_obj (objects/mesh.rs:148-199, tobj 3.1.0 semantics:
// first model only, fan triangulation, per-corner position/normal) and the Appendix-B JSON scene schema
// (data/scene_*.json; the reference ships the files but no parser, raytracer_codegen/src/lib.rs:44-82).
#include <cctype>
#include <cstdio>
#include <cstdlib>
#include <map>
#include <memory>

#include "host_common.hpp"

using namespace b200rt;

namespace {

// ---------------------------------------------------------------- OBJ
struct ObjMesh {
  std::vector<D3> pos, nrm;
  std::vector<int> tri_pos, tri_nrm;  // 3 per triangle; nrm = -1 when absent
};

bool read_file(const char* path, std::string& out) {
  FILE* f = std::fopen(path, "rb");
  if (!f) return false;
  char buf[1 << 16];
  size_t n;
  while ((n = std::fread(buf, 1, sizeof(buf), f)) > 0) out.append(buf, n);
  std::fclose(f);
  return true;
}

bool parse_obj(const std::string& text, ObjMesh& m, std::string& err) {
  size_t i = 0, n = text.size();
  int models_with_faces = 0;
  bool cur_has_faces = false;
  while (i < n) {
    size_t e = text.find('\n', i);
    if (e == std::string::npos) e = n;
    std::string line = text.substr(i, e - i);
    i = e + 1;
    if (!line.empty() && line.back() == '\r') line.pop_back();
    const char* s = line.c_str();
    while (*s == ' ' || *s == '\t') ++s;
    if (s[0] == 'v' && (s[1] == ' ' || s[1] == '\t')) {
      char* end;
      double x = std::strtod(s + 2, &end), y = std::strtod(end, &end), z = std::strtod(end, &end);
      m.pos.push_back(D3((double)(float)x, (double)(float)y, (double)(float)z));  // tobj stores f32 (mesh.rs:163)
    } else if (s[0] == 'v' && s[1] == 'n') {
      char* end;
      double x = std::strtod(s + 3, &end), y = std::strtod(end, &end), z = std::strtod(end, &end);
      m.nrm.push_back(D3((double)(float)x, (double)(float)y, (double)(float)z));
    } else if ((s[0] == 'o' || s[0] == 'g') && (s[1] == ' ' || s[1] == '\t' || s[1] == 0)) {
      if (cur_has_faces) { models_with_faces++; cur_has_faces = false; }
    } else if (s[0] == 'f' && (s[1] == ' ' || s[1] == '\t')) {
      if (models_with_faces >= 1) continue;  // models[0] only (mesh.rs:159)
      cur_has_faces = true;
      std::vector<int> vp, vn;
      const char* p = s + 1;
      while (*p) {
        while (*p == ' ' || *p == '\t') ++p;
        if (!*p) break;
        char* end;
        long a = std::strtol(p, &end, 10);
        if (end == p) { err = "bad face index"; return false; }
        long c = 0;
        bool has_n = false;
        p = end;
        if (*p == '/') {
          ++p;
          if (*p != '/') { std::strtol(p, &end, 10); p = end; }
          if (*p == '/') { ++p; c = std::strtol(p, &end, 10); has_n = end != p; p = end; }
        }
        int pi = a > 0 ? (int)a - 1 : (int)m.pos.size() + (int)a;
        int ni = has_n ? (c > 0 ? (int)c - 1 : (int)m.nrm.size() + (int)c) : -1;
        if (pi < 0 || pi >= (int)m.pos.size() || ni >= (int)m.nrm.size()) { err = "face index out of range"; return false; }
        vp.push_back(pi); vn.push_back(ni);
      }
      if (vp.size() < 3) { err = "face with fewer than 3 vertices"; return false; }
      for (size_t k = 1; k + 1 < vp.size(); ++k) {  // fan
        m.tri_pos.push_back(vp[0]); m.tri_pos.push_back(vp[k]); m.tri_pos.push_back(vp[k + 1]);
        m.tri_nrm.push_back(vn[0]); m.tri_nrm.push_back(vn[k]); m.tri_nrm.push_back(vn[k + 1]);
      }
    }
  }
  return true;
}

// one round of midpoint subdivision (each triangle -> 4); normals are regenerated afterwards
void subdivide_once(ObjMesh& m) {
  std::map<std::pair<int, int>, int> mid;
  auto midpoint = [&](int a, int b) {
    std::pair<int, int> key(std::min(a, b), std::max(a, b));
    auto it = mid.find(key);
    if (it != mid.end()) return it->second;
    m.pos.push_back((m.pos[a] + m.pos[b]) * 0.5);
    int id = (int)m.pos.size() - 1;
    mid[key] = id;
    return id;
  };
  std::vector<int> out;
  out.reserve(m.tri_pos.size() * 4);
  for (size_t t = 0; t + 2 < m.tri_pos.size(); t += 3) {
    int a = m.tri_pos[t], b = m.tri_pos[t + 1], c = m.tri_pos[t + 2];
    int ab = midpoint(a, b), bc = midpoint(b, c), ca = midpoint(c, a);
    const int tris[12] = {a, ab, ca, ab, b, bc, ca, bc, c, ab, bc, ca};
    out.insert(out.end(), tris, tris + 12);
  }
  m.tri_pos.swap(out);
  m.tri_nrm.assign(m.tri_pos.size(), -1);
}

void make_vertex_normals(ObjMesh& m) {  // area-weighted, per position index
  m.nrm.assign(m.pos.size(), D3(0, 0, 0));
  for (size_t t = 0; t + 2 < m.tri_pos.size(); t += 3) {
    int a = m.tri_pos[t], b = m.tri_pos[t + 1], c = m.tri_pos[t + 2];
    D3 n = cross(m.pos[b] - m.pos[a], m.pos[c] - m.pos[a]);
    m.nrm[a] = m.nrm[a] + n; m.nrm[b] = m.nrm[b] + n; m.nrm[c] = m.nrm[c] + n;
  }
  for (auto& n : m.nrm) n = length(n) > 0 ? unit(n) : D3(0, 1, 0);
  m.tri_nrm = m.tri_pos;
}

// ---------------------------------------------------------------- JSON (RFC 8259 subset: all the files use)
struct JValue {
  enum Kind { Null, Num, Str, Obj, Arr, Bool } kind = Null;
  double num = 0;
  std::string str;
  std::vector<std::pair<std::string, std::unique_ptr<JValue>>> obj;
  std::vector<std::unique_ptr<JValue>> arr;
  const JValue* get(const char* k) const {
    for (auto& kv : obj) if (kv.first == k) return kv.second.get();
    return nullptr;
  }
};
struct JParser {
  const char* p; const char* end; std::string err;
  void ws() { while (p < end && (*p == ' ' || *p == '\n' || *p == '\t' || *p == '\r')) ++p; }
  bool parse_string(std::string& out) {
    if (p >= end || *p != '"') { err = "expected string"; return false; }
    ++p;
    while (p < end && *p != '"') {
      if (*p == '\\' && p + 1 < end) { ++p; out.push_back(*p == 'n' ? '\n' : (*p == 't' ? '\t' : *p)); ++p; }
      else out.push_back(*p++);
    }
    if (p >= end) { err = "unterminated string"; return false; }
    ++p;
    return true;
  }
  std::unique_ptr<JValue> parse(int depth = 0) {
    std::unique_ptr<JValue> v(new JValue());
    if (depth > 4096) { err = "nesting too deep"; return nullptr; }
    ws();
    if (p >= end) { err = "unexpected end"; return nullptr; }
    if (*p == '{') {
      v->kind = JValue::Obj; ++p; ws();
      if (p < end && *p == '}') { ++p; return v; }
      for (;;) {
        ws();
        std::string k;
        if (!parse_string(k)) return nullptr;
        ws();
        if (p >= end || *p != ':') { err = "expected ':'"; return nullptr; }
        ++p;
        auto c = parse(depth + 1);
        if (!c) return nullptr;
        v->obj.emplace_back(k, std::move(c));
        ws();
        if (p < end && *p == ',') { ++p; continue; }
        if (p < end && *p == '}') { ++p; return v; }
        err = "expected ',' or '}'"; return nullptr;
      }
    }
    if (*p == '[') {
      v->kind = JValue::Arr; ++p; ws();
      if (p < end && *p == ']') { ++p; return v; }
      for (;;) {
        auto c = parse(depth + 1);
        if (!c) return nullptr;
        v->arr.push_back(std::move(c));
        ws();
        if (p < end && *p == ',') { ++p; continue; }
        if (p < end && *p == ']') { ++p; return v; }
        err = "expected ',' or ']'"; return nullptr;
      }
    }
    if (*p == '"') { v->kind = JValue::Str; if (!parse_string(v->str)) return nullptr; return v; }
    if (!std::strncmp(p, "true", std::min<size_t>(4, end - p))) { v->kind = JValue::Bool; v->num = 1; p += 4; return v; }
    if (!std::strncmp(p, "false", std::min<size_t>(5, end - p))) { v->kind = JValue::Bool; p += 5; return v; }
    if (!std::strncmp(p, "null", std::min<size_t>(4, end - p))) { p += 4; return v; }
    char* e;
    std::string tmp(p, std::min<size_t>(64, end - p));
    double d = std::strtod(tmp.c_str(), &e);
    if (e == tmp.c_str()) { err = "unexpected token"; return nullptr; }
    p += e - tmp.c_str();
    v->kind = JValue::Num; v->num = d;
    return v;
  }
};

// ---------------------------------------------------------------- YAML (the block-style subset serde_yaml emits for
// data/scene_*.yaml: nested mappings, "- " sequences of mappings, plain scalars; README.md:86-89 says the YAML twins
// carry the same objects as the JSON files).  Parsed into the same JValue tree, so one scene walker serves both.
struct YLine { int indent; bool dash; std::string key, val; bool has_val; };
struct YParser {
  std::vector<YLine> lines;
  std::string err;
  size_t at = 0;
  static std::string trim(const std::string& s) {
    size_t a = s.find_first_not_of(" \t\r"), b = s.find_last_not_of(" \t\r");
    return a == std::string::npos ? std::string() : s.substr(a, b - a + 1);
  }
  bool tokenize(const char* p, const char* end) {
    while (p < end) {
      const char* e = (const char*)std::memchr(p, '\n', end - p);
      if (!e) e = end;
      std::string raw(p, e);
      p = e < end ? e + 1 : end;
      size_t hash = raw.find(" #");
      if (hash != std::string::npos) raw.resize(hash);
      if (trim(raw).empty() || trim(raw)[0] == '#' || trim(raw) == "---") continue;
      int ind = 0;
      while (ind < (int)raw.size() && raw[ind] == ' ') ++ind;
      std::string body = raw.substr(ind);
      while (body.size() >= 2 && body[0] == '-' && body[1] == ' ') {  // "- key: v" = a dash marker + the mapping's first line
        lines.push_back({ind, true, "", "", false});
        size_t skip = 2;
        while (skip < body.size() && body[skip] == ' ') ++skip;
        ind += (int)skip; body = body.substr(skip);
      }
      if (body == "-") { lines.push_back({ind, true, "", "", false}); continue; }
      size_t colon = body.find(':');
      if (colon == std::string::npos) { err = "YAML: expected 'key: value' in '" + body + "'"; return false; }
      YLine l{ind, false, trim(body.substr(0, colon)), trim(body.substr(colon + 1)), false};
      if (l.key.size() >= 2 && (l.key[0] == '"' || l.key[0] == '\'')) l.key = l.key.substr(1, l.key.size() - 2);
      l.has_val = !l.val.empty();
      lines.push_back(l);
    }
    return true;
  }
  std::unique_ptr<JValue> scalar(std::string v) {
    std::unique_ptr<JValue> j(new JValue());
    if (v.size() >= 2 && (v[0] == '"' || v[0] == '\'')) { j->kind = JValue::Str; j->str = v.substr(1, v.size() - 2); return j; }
    if (v == "~" || v == "null") return j;
    if (v == "true" || v == "false") { j->kind = JValue::Bool; j->num = v == "true"; return j; }
    char* e;
    double d = std::strtod(v.c_str(), &e);
    if (e != v.c_str() && *e == 0) { j->kind = JValue::Num; j->num = d; return j; }
    j->kind = JValue::Str; j->str = v;
    return j;
  }
  // the block starting at lines[at], whose lines sit at exactly `indent`
  std::unique_ptr<JValue> block(int indent, int depth) {
    std::unique_ptr<JValue> v(new JValue());
    if (depth > 4096) { err = "YAML: nesting too deep"; return nullptr; }
    if (at >= lines.size()) return v;
    if (lines[at].dash) {
      v->kind = JValue::Arr;
      while (at < lines.size() && lines[at].dash && lines[at].indent == indent) {
        ++at;
        if (at >= lines.size() || lines[at].indent <= indent) { v->arr.emplace_back(new JValue()); continue; }
        auto c = block(lines[at].indent, depth + 1);
        if (!c) return nullptr;
        v->arr.push_back(std::move(c));
      }
      return v;
    }
    v->kind = JValue::Obj;
    while (at < lines.size() && !lines[at].dash && lines[at].indent == indent) {
      const YLine l = lines[at++];
      if (l.has_val) { v->obj.emplace_back(l.key, scalar(l.val)); continue; }
      // nested block: deeper mapping, or a sequence (serde_yaml indents it, plain YAML may keep it at the key's indent)
      if (at < lines.size() && (lines[at].indent > indent || (lines[at].dash && lines[at].indent == indent))) {
        auto c = block(lines[at].indent, depth + 1);
        if (!c) return nullptr;
        v->obj.emplace_back(l.key, std::move(c));
      } else v->obj.emplace_back(l.key, std::unique_ptr<JValue>(new JValue()));
    }
    if (at < lines.size() && lines[at].indent > indent) { err = "YAML: bad indentation near key '" + lines[at].key + "'"; return nullptr; }
    return v;
  }
};

bool jnum(const JValue* v, const char* k, double& out) {
  const JValue* c = v ? v->get(k) : nullptr;
  if (!c || c->kind != JValue::Num) return false;
  out = c->num;
  return true;
}
bool jvec(const JValue* v, const char* k, double out[3]) {
  const JValue* c = v ? v->get(k) : nullptr;
  return c && jnum(c, "x", out[0]) && jnum(c, "y", out[1]) && jnum(c, "z", out[2]);
}
std::string jtype(const JValue* v) {
  const JValue* t = v ? v->get("type") : nullptr;
  return t && t->kind == JValue::Str ? t->str : std::string();
}

int json_texture(b200rt_builder* b, const JValue* t) {
  std::string ty = jtype(t);
  double c[3];
  if (ty == "ConstantTexture") {
    if (!jvec(t, "color", c)) { b->err = "ConstantTexture without color"; return -B200RT_ERR_PARSE; }
    return b200rt_builder_texture_constant(b, c[0], c[1], c[2]);
  }
  if (ty == "CheckerTexture") {  // CheckerTexture(ConstantTexture, ConstantTexture), material.rs:49
    const JValue *t0 = t->get("t0"), *t1 = t->get("t1");
    double c1[3];
    if (jtype(t0) != "ConstantTexture" || jtype(t1) != "ConstantTexture" || !jvec(t0, "color", c) || !jvec(t1, "color", c1)) {
      b->err = "CheckerTexture needs two ConstantTexture children"; return -B200RT_ERR_PARSE;
    }
    return b200rt_builder_texture_checker(b, c, c1);
  }
  b->err = "unknown texture type '" + ty + "'";
  return -B200RT_ERR_PARSE;
}

int json_material(b200rt_builder* b, const JValue* m) {
  std::string ty = jtype(m);
  double c[3], x;
  if (ty == "Lambertian") { int t = json_texture(b, m->get("albedo")); return t < 0 ? t : b200rt_builder_material(b, B200RT_MAT_LAMBERTIAN, t, 0); }
  if (ty == "Metal") {  // albedo is a bare colour in the files
    if (!jvec(m, "albedo", c) || !jnum(m, "fuzz", x)) { b->err = "Metal needs albedo and fuzz"; return -B200RT_ERR_PARSE; }
    return b200rt_builder_material(b, B200RT_MAT_METAL, b200rt_builder_texture_constant(b, c[0], c[1], c[2]), x);
  }
  if (ty == "Dielectric") {  // only ref_idx in the files => white albedo
    if (!jnum(m, "ref_idx", x)) { b->err = "Dielectric needs ref_idx"; return -B200RT_ERR_PARSE; }
    return b200rt_builder_material(b, B200RT_MAT_DIELECTRIC, b200rt_builder_texture_constant(b, 1, 1, 1), x);
  }
  if (ty == "DiffuseLight") { int t = json_texture(b, m->get("emit")); return t < 0 ? t : b200rt_builder_material(b, B200RT_MAT_DIFFUSE_LIGHT, t, 0); }
  b->err = "unknown material type '" + ty + "'";
  return -B200RT_ERR_PARSE;
}

int json_node(b200rt_builder* b, const JValue* n, int depth) {
  if (!n || n->kind != JValue::Obj) { b->err = "scene node is not an object"; return B200RT_ERR_PARSE; }
  if (depth > 4096) { b->err = "scene graph too deep"; return B200RT_ERR_PARSE; }
  std::string ty = jtype(n);
  if (ty == "HitableList") {
    const JValue* items = n->get("items");
    if (!items || items->kind != JValue::Arr) { b->err = "HitableList without items"; return B200RT_ERR_PARSE; }
    for (auto& it : items->arr) { int rc = json_node(b, it.get(), depth + 1); if (rc) return rc; }
    return B200RT_OK;
  }
  if (ty == "BVHNode") {  // stored bounding_box is redundant and ignored (the wide BVH is rebuilt)
    int rc = json_node(b, n->get("left"), depth + 1);
    return rc ? rc : json_node(b, n->get("right"), depth + 1);
  }
  if (ty == "Sphere") {
    double c[3], r;
    if (!jvec(n, "center", c) || !jnum(n, "radius", r)) { b->err = "Sphere needs center and radius"; return B200RT_ERR_PARSE; }
    int m = json_material(b, n->get("material"));
    if (m < 0) return -m;
    int id = b200rt_builder_sphere(b, c, r, m);
    return id < 0 ? -id : B200RT_OK;
  }
  b->err = "unknown node type '" + ty + "'";
  return B200RT_ERR_PARSE;
}

}  // namespace

extern "C" {

int b200rt_builder_mesh_obj(b200rt_builder* b, const char* path, int32_t material, int32_t make_normals, int32_t subdivide) try {
  if (!b || !path) return -B200RT_ERR_INVALID;
  std::string text;
  if (!read_file(path, text)) { b->err = std::string("Failed to load OBJ file: ") + path; return -B200RT_ERR_IO; }  // mesh.rs:158
  ObjMesh m;
  std::string err;
  if (!parse_obj(text, m, err)) { b->err = "OBJ: " + err; return -B200RT_ERR_PARSE; }
  if (m.tri_pos.empty()) { b->err = "OBJ: no faces"; return -B200RT_ERR_PARSE; }
  for (int i = 0; i < subdivide; ++i) subdivide_once(m);
  bool missing = false;
  for (int v : m.tri_nrm) if (v < 0) { missing = true; break; }
  if (missing) {
    if (!make_normals) { b->err = "OBJ has no vertex normals (Triangle::hit indexes normals[a], mesh.rs:62-64)"; return -B200RT_ERR_INVALID; }
    make_vertex_normals(m);
  }
  int first = -1;
  for (size_t t = 0; t + 2 < m.tri_pos.size(); t += 3) {
    D3 pa = m.pos[m.tri_pos[t]], pb = m.pos[m.tri_pos[t + 1]], pc = m.pos[m.tri_pos[t + 2]];
    D3 na = m.nrm[m.tri_nrm[t]], nb = m.nrm[m.tri_nrm[t + 1]], nc = m.nrm[m.tri_nrm[t + 2]];
    double a[3] = {pa.x, pa.y, pa.z}, bb[3] = {pb.x, pb.y, pb.z}, c[3] = {pc.x, pc.y, pc.z};
    double x[3] = {na.x, na.y, na.z}, y[3] = {nb.x, nb.y, nb.z}, z[3] = {nc.x, nc.y, nc.z};
    int id = b200rt_builder_triangle(b, a, bb, c, x, y, z, material);
    if (id < 0) return id;
    if (first < 0) first = id;
  }
  return first;
} catch (const std::bad_alloc&) {  // no exception crosses the C ABI
  b->err = "out of host memory while loading the OBJ";
  return -B200RT_ERR_NOMEM;
} catch (const std::exception& e) {
  b->err = std::string("OBJ: ") + e.what();
  return -B200RT_ERR_PARSE;
}

static int load_scene_tree(b200rt_builder* b, const JValue* root, b200rt_camera* cam_out);

int b200rt_builder_load_json(b200rt_builder* b, const char* text, uint64_t len, b200rt_camera* cam_out) try {
  if (!b || !text) return B200RT_ERR_INVALID;
  JParser jp;
  jp.p = text; jp.end = text + len;
  auto root = jp.parse();
  if (!root) { b->err = "JSON: " + jp.err; return B200RT_ERR_PARSE; }
  return load_scene_tree(b, root.get(), cam_out);
} catch (const std::bad_alloc&) {
  b->err = "out of host memory while loading the scene file";
  return B200RT_ERR_NOMEM;
} catch (const std::exception& e) {
  b->err = std::string("scene file: ") + e.what();
  return B200RT_ERR_PARSE;
}

int b200rt_builder_load_yaml(b200rt_builder* b, const char* text, uint64_t len, b200rt_camera* cam_out) try {
  if (!b || !text) return B200RT_ERR_INVALID;
  YParser yp;
  if (!yp.tokenize(text, text + len)) { b->err = yp.err; return B200RT_ERR_PARSE; }
  if (yp.lines.empty()) { b->err = "YAML: empty document"; return B200RT_ERR_PARSE; }
  auto root = yp.block(yp.lines[0].indent, 0);
  if (!root || yp.at != yp.lines.size()) { b->err = yp.err.empty() ? "YAML: trailing content / bad indentation" : yp.err; return B200RT_ERR_PARSE; }
  return load_scene_tree(b, root.get(), cam_out);
} catch (const std::bad_alloc&) {
  b->err = "out of host memory while loading the scene file";
  return B200RT_ERR_NOMEM;
} catch (const std::exception& e) {
  b->err = std::string("scene file: ") + e.what();
  return B200RT_ERR_PARSE;
}

static int load_scene_tree(b200rt_builder* b, const JValue* root, b200rt_camera* cam_out) {
  const JValue* objects = root->get("objects");
  if (!objects) { b->err = "scene file: no 'objects'"; return B200RT_ERR_PARSE; }
  bool flat = jtype(objects) == "HitableList";
  if (flat) {  // a BVHNode anywhere below makes it the BVH arm
    const JValue* items = objects->get("items");
    if (items) for (auto& it : items->arr) if (jtype(it.get()) == "BVHNode") flat = false;
  }
  int rc = json_node(b, objects, 0);
  if (rc) return rc;
  if (flat) b->flags |= B200RT_SCENE_FLAT_LIST;
  const JValue* cam = root->get("camera");
  if (cam_out && cam) {
    double from[3], at[3], vup[3], vfov, aspect, aperture, focus;
    if (!jvec(cam, "look_from", from) || !jvec(cam, "look_at", at) || !jvec(cam, "vup", vup) || !jnum(cam, "vfov", vfov) ||
        !jnum(cam, "aspect", aspect) || !jnum(cam, "aperture", aperture) || !jnum(cam, "focus_dist", focus)) {
      b->err = "JSON: incomplete camera"; return B200RT_ERR_PARSE;
    }
    return b200rt_camera_make(from, at, vup, vfov, aspect, aperture, focus, 0, 0, cam_out);
  }
  return B200RT_OK;
}

}  // extern "C"
