// cli_main.cpp — the reference's `main` (main.rs:49-72) over the C ABI: pick a scene, render, save a PNG, print the
// timing lines in the reference's form.  Host C++ only; every GPU call goes through include/b200rt.h.
//
//   b200rt_render --scene book2|book1|cornell_book3|cornell_ref|mesh|<file.json>|<file.yaml> [--obj path.obj]
//                 [--earth earth.ppm] [--width W] [--height H] [--spp N] [--depth D] [--integrator path|mixture|nee|sppm]
//                 [--iterations 50] [--photons 500000] [--light-scale 1e6] [--seed S] [--gpus G] [--out output/test.png]
//
// --integrator sppm is what the reference's main renders (SPPMIntegrator::new + capture_image, main.rs:51-53); with --gpus G through b200rt_render_sppm_multi.
//
// --gpus G shards the samples over G devices (b200rt_render_multi: one context and one host thread per device -- the
// reference's own decomposition is row bands over a thread pool, camera.rs:74-111 -- and one NCCL or peer-memory
// sum-reduce of the accumulation buffers onto device 0).
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <sstream>
#include <string>
#include <thread>
#include <vector>

#include "../../include/b200rt.h"

namespace {

bool ends_with(const std::string& s, const char* suf) {
  const size_t n = std::strlen(suf);
  return s.size() >= n && s.compare(s.size() - n, n, suf) == 0;
}

bool read_file(const std::string& path, std::string& out) {
  std::ifstream f(path, std::ios::binary);
  if (!f) return false;
  std::ostringstream ss;
  ss << f.rdbuf();
  out = ss.str();
  return true;
}

// binary PPM (P6, maxval 255): the decoded earthmap for book-2 (there is no JPEG decoder on this side of the ABI)
bool read_ppm(const std::string& path, std::vector<uint8_t>& rgb, int& w, int& h) {
  std::string s;
  if (!read_file(path, s) || s.size() < 11 || s[0] != 'P' || s[1] != '6') return false;
  size_t p = 2;
  int vals[3], k = 0;
  while (k < 3 && p < s.size()) {
    while (p < s.size() && (s[p] == ' ' || s[p] == '\n' || s[p] == '\r' || s[p] == '\t')) ++p;
    if (p < s.size() && s[p] == '#') { while (p < s.size() && s[p] != '\n') ++p; continue; }
    int v = 0, d = 0;
    while (p < s.size() && s[p] >= '0' && s[p] <= '9') { v = v * 10 + (s[p] - '0'); ++p; ++d; }
    if (!d) return false;
    vals[k++] = v;
  }
  if (k < 3 || vals[2] != 255) return false;
  ++p;  // the single whitespace byte after maxval
  w = vals[0]; h = vals[1];
  const size_t n = (size_t)w * h * 3;
  if (w <= 0 || h <= 0 || s.size() < p + n) return false;
  rgb.assign(s.begin() + p, s.begin() + p + n);
  return true;
}

// World::new on one device (scene upload happens on the device's own host thread, like the render)
struct Shard {
  int device = 0, rc = 0;
  std::string err;
  b200rt_ctx* ctx = nullptr;
  b200rt_scene* scene = nullptr;
};

void create_shard(const b200rt_scene_desc* desc, Shard* sh) {
  sh->rc = b200rt_ctx_create(sh->device, &sh->ctx);
  if (sh->rc) { sh->err = "no CUDA device " + std::to_string(sh->device) + " (there is no CPU fallback)"; return; }
  sh->rc = b200rt_scene_create(sh->ctx, desc, &sh->scene);
  if (sh->rc) sh->err = b200rt_last_error(sh->ctx);
}

double seconds_since(std::chrono::steady_clock::time_point t0) {
  return std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
}

}  // namespace

int main(int argc, char** argv) {
  std::string scene = "book2", obj, earth, out = "output/test.png", integ = "path";
  int width = 0, height = 0, spp = 64, depth = 50, gpus = 1, iterations = 50, photons = 500000;
  double light_scale = 1e6;
  unsigned long long seed = 1;
  for (int i = 1; i < argc; ++i) {
    const std::string a = argv[i];
    auto next = [&](const char* what) -> const char* {
      if (i + 1 >= argc) { std::fprintf(stderr, "missing value for %s\n", what); std::exit(2); }
      return argv[++i];
    };
    if (a == "--scene") scene = next("--scene");
    else if (a == "--obj") obj = next("--obj");
    else if (a == "--earth") earth = next("--earth");
    else if (a == "--out") out = next("--out");
    else if (a == "--integrator") integ = next("--integrator");
    else if (a == "--width") width = std::atoi(next("--width"));
    else if (a == "--height") height = std::atoi(next("--height"));
    else if (a == "--spp") spp = std::atoi(next("--spp"));
    else if (a == "--depth") depth = std::atoi(next("--depth"));
    else if (a == "--gpus") gpus = std::atoi(next("--gpus"));
    else if (a == "--iterations") iterations = std::atoi(next("--iterations"));
    else if (a == "--photons") photons = std::atoi(next("--photons"));
    else if (a == "--light-scale") light_scale = std::atof(next("--light-scale"));
    else if (a == "--seed") seed = std::strtoull(next("--seed"), nullptr, 10);
    else { std::fprintf(stderr, "unknown argument %s\n", a.c_str()); return 2; }
  }
  if (spp < 1 || gpus < 1 || depth < 0) { std::fprintf(stderr, "spp, gpus must be >= 1, depth >= 0\n"); return 2; }

  const auto t_start = std::chrono::steady_clock::now();
  b200rt_builder* b = nullptr;
  if (b200rt_builder_create(&b)) { std::fprintf(stderr, "builder_create failed\n"); return 1; }
  b200rt_camera cam{};
  b200rt_params params{};
  int rc;
  const bool from_file = ends_with(scene, ".json") || ends_with(scene, ".yaml") || ends_with(scene, ".yml");
  if (from_file) {
    std::string text;
    if (!read_file(scene, text)) { std::fprintf(stderr, "cannot read %s\n", scene.c_str()); return 1; }
    rc = ends_with(scene, ".json") ? b200rt_builder_load_json(b, text.data(), text.size(), &cam)
                                   : b200rt_builder_load_yaml(b, text.data(), text.size(), &cam);
    params.width = 400; params.height = 400; params.max_depth = 50; params.t_min = 0.001f;  // photon_mapper.rs:333-339
    params.integrator = B200RT_INTEGRATOR_PATH; params.pipeline = B200RT_PIPELINE_WAVEFRONT;
  } else {
    std::vector<uint8_t> rgb;
    int ew = 0, eh = 0;
    if (!earth.empty() && !read_ppm(earth, rgb, ew, eh)) { std::fprintf(stderr, "cannot read PPM %s\n", earth.c_str()); return 1; }
    const double aspect = (width > 0 && height > 0) ? (double)width / height : 0.0;
    rc = b200rt_builder_make_scene(b, scene.c_str(), seed, obj.empty() ? nullptr : obj.c_str(), rgb.empty() ? nullptr : rgb.data(), ew, eh,
                                   aspect, &cam, &params);
  }
  if (rc) { std::fprintf(stderr, "scene: %s\n", b200rt_builder_error(b)); return 1; }
  if (width > 0) params.width = width;
  if (height > 0) params.height = height;
  params.max_depth = depth; params.seed = seed;
  if (integ == "path") params.integrator = B200RT_INTEGRATOR_PATH;
  else if (integ == "mixture") params.integrator = B200RT_INTEGRATOR_MIXTURE;
  else if (integ == "nee") params.integrator = B200RT_INTEGRATOR_NEE;
  else if (integ == "sppm") params.integrator = B200RT_INTEGRATOR_PATH;  // the final pass has its own loop (photon_mapper.rs:327-365)
  else { std::fprintf(stderr, "unknown integrator %s\n", integ.c_str()); return 2; }
  b200rt_scene_desc desc{};
  if (b200rt_builder_finish(b, &desc)) { std::fprintf(stderr, "scene: %s\n", b200rt_builder_error(b)); return 1; }

  if (integ == "sppm") {  // main.rs:51-53: SPPMIntegrator::new(world) then capture_image(integrator); --gpus G: b200rt_render_sppm_multi
    std::vector<Shard> shards(gpus);
    {
      std::vector<std::thread> threads;
      for (int g = 0; g < gpus; ++g) { shards[g].device = g; threads.emplace_back(create_shard, &desc, &shards[g]); }
      for (auto& t : threads) t.join();
    }
    std::vector<b200rt_scene*> scenes(gpus);
    for (int g = 0; g < gpus; ++g) {
      if (shards[g].rc) { std::fprintf(stderr, "scene on device %d: %s (there is no CPU fallback)\n", g, shards[g].err.c_str()); return 1; }
      scenes[g] = shards[g].scene;
    }
    b200rt_ctx* ctx = shards[0].ctx;
    params.spp_begin = 0; params.spp_count = spp;
    b200rt_sppm_params sp{};
    sp.iterations = iterations; sp.photons_per_iter = photons; sp.k_global = 100; sp.k_caustic = 50; sp.alpha = 0.7; sp.light_scale = light_scale;
    b200rt_sppm_stats st{};
    std::vector<float> accum((size_t)params.width * params.height * 4, 0.f);
    rc = gpus > 1 ? b200rt_render_sppm_multi(gpus, scenes.data(), &cam, &params, &sp, accum.data(), nullptr, &st)
                  : b200rt_render_sppm(scenes[0], &cam, &params, &sp, accum.data(), nullptr, &st);
    if (rc) { std::fprintf(stderr, "sppm: %s\n", b200rt_last_error(ctx)); return 1; }
    std::vector<uint8_t> rgb8((size_t)params.width * params.height * 3);
    if (b200rt_tonemap_rgb8(ctx, accum.data(), params.width, params.height, rgb8.data())) { std::fprintf(stderr, "tonemap: %s\n", b200rt_last_error(ctx)); return 1; }
    for (int g = 0; g < gpus; ++g) { b200rt_scene_destroy(shards[g].scene); b200rt_ctx_destroy(shards[g].ctx); }
    b200rt_builder_destroy(b);
    if (b200rt_write_png(out.c_str(), rgb8.data(), params.width, params.height)) { std::fprintf(stderr, "cannot write %s\n", out.c_str()); return 1; }
    // one device: CUDA-event stage times (photons overlap the eye passes); several: host clocks of the iterations and of the rest
    const double sppm_s = gpus > 1 ? st.ms_eye * 1e-3 : (st.ms_total - st.ms_final) * 1e-3, rt_s = st.ms_final * 1e-3;
    std::printf("Total: %.3fs\n\tSPPM: %.3fs\n\tRT: %.3fs\n", seconds_since(t_start), sppm_s, rt_s);
    std::printf("%dx%d, %d iterations x %d photons (%llu global, %llu caustic stored), %d spp on %d GPU(s): total %.1f ms (photons %.1f, grid %.1f, eye %.1f, final %.1f) -> %s\n",
                params.width, params.height, iterations, photons, (unsigned long long)st.global_photons, (unsigned long long)st.caustic_photons, spp, gpus,
                st.ms_total, st.ms_photons, st.ms_grid, st.ms_eye, st.ms_final, out.c_str());
    return 0;
  }
  const auto t_rt = std::chrono::steady_clock::now();
  if (gpus > spp) gpus = spp;
  std::vector<Shard> shards(gpus);
  {
    std::vector<std::thread> threads;
    for (int g = 0; g < gpus; ++g) { shards[g].device = g; threads.emplace_back(create_shard, &desc, &shards[g]); }
    for (auto& t : threads) t.join();
  }
  std::vector<b200rt_scene*> scenes(gpus);
  for (int g = 0; g < gpus; ++g) {
    if (shards[g].rc) { std::fprintf(stderr, "scene on device %d: %s\n", g, shards[g].err.c_str()); return 1; }
    scenes[g] = shards[g].scene;
  }
  // capture_image: sample ranges per device + one NCCL / peer-memory sum-reduce onto device 0 (b200rt_render_multi)
  params.spp_begin = 0; params.spp_count = spp;
  std::vector<float> accum((size_t)params.width * params.height * 4, 0.f);
  b200rt_multi_stats ms{};
  rc = b200rt_render_multi(gpus, scenes.data(), &cam, &params, accum.data(), nullptr, &ms);
  if (rc) { std::fprintf(stderr, "render: %s\n", b200rt_last_error(shards[0].ctx)); return 1; }
  const unsigned long long rays = ms.extend_rays + ms.shadow_rays;
  // tonemap on device 0 (vec3.rs:223-231) and save
  std::vector<uint8_t> rgb8((size_t)params.width * params.height * 3);
  rc = b200rt_tonemap_rgb8(shards[0].ctx, accum.data(), params.width, params.height, rgb8.data());
  if (rc) { std::fprintf(stderr, "tonemap: %s\n", b200rt_last_error(shards[0].ctx)); return 1; }
  for (int g = 0; g < gpus; ++g) { b200rt_scene_destroy(shards[g].scene); b200rt_ctx_destroy(shards[g].ctx); }
  b200rt_builder_destroy(b);
  const double rt_s = seconds_since(t_rt);
  rc = b200rt_write_png(out.c_str(), rgb8.data(), params.width, params.height);
  if (rc) { std::fprintf(stderr, "cannot write %s\n", out.c_str()); return 1; }
  // main.rs:56-71 prints "Total / SPPM / RT"; there is no photon pass on this path
  std::printf("Total: %.3fs\n\tSPPM: %.3fs\n\tRT: %.3fs\n", seconds_since(t_start), 0.0, rt_s);
  const double paths = (double)params.width * params.height * spp;
  std::printf("%dx%d, %d spp on %d GPU(s): %.1f Mpaths/s, %.1f Mrays/s (host-timed, scene upload included); device: render %.1f ms + %s reduce %.2f ms -> %s\n",
              params.width, params.height, spp, gpus, paths / rt_s / 1e6, (double)rays / rt_s / 1e6, ms.ms_render_max,
              ms.reduce_path == B200RT_REDUCE_NCCL ? "NCCL" : (ms.reduce_path == B200RT_REDUCE_PEER ? "peer-memory" : "no"), ms.ms_reduce, out.c_str());
  return 0;
}
