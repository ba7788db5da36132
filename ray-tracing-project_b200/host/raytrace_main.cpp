// raytrace_main.cpp - headless replacement of the reference's app shell (src/main.cpp): what its
// key T does (main.cpp:68-69 -> Flyscene::raytraceScene -> result.ppm), driven by arguments instead
// of a GLFW window.  Also the scene-dump tool of the CPU-only tests (--dump-scene needs no GPU).
//
//   raytrace [--model DIR/FILE.obj] [--size W H] [--depth MAX_DEPTH] [--fovy DEG]
//            [--view m0..m11] [--light x y z r g b]... [--dirlight x y z r g b]... [--nolights]
//            [--soft x y z r g b RADIUS N] [--seed S]   soft light: N jittered point lights (libc rand, or with --seed
//                                            the counter-based generator of Flyscene::sphericalLightSeeded)
//            [--out result.ppm] [--out-p6 result_p6.ppm] [--out-png result.png] [--device N] [--dump-scene PREFIX] [--dump-lights PREFIX]
//            [--gpus N]                      shard the frame over N devices (DEVICE, DEVICE + 1, ... modulo the device count:
//                                            with N above the device count ranks share a GPU); same image as --gpus 1
//            [--lights-only]                 with --dump-lights: build the light list and stop (no scene, no GPU)
//            [--srand S]                     seed libc rand() before --soft lights are drawn (flyscene.cpp:529-538 uses rand())
//            [--dump-rays PREFIX]            camera origin + per-pixel primary directions (no GPU)
//            [--ppm-from-f32 IN.f32 OUT.ppm] W*H*3 floats -> the reference's ASCII P3 (no GPU)
//            [--png-from-u8 IN.u8 OUT.png]   W*H*3 bytes -> PNG (no GPU)
#include <array>
#include <chrono>
#include <climits>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>

#include "flyscene.hpp"
#include "scene_dump.hpp"

namespace {

using scene_dump::dump;

int usage() {
  fprintf(stderr, "usage: raytrace [--model path.obj] [--size W H] [--depth D] [--fovy deg] [--view 12 floats]\n"
                  "                [--light x y z r g b] [--dirlight x y z r g b] [--nolights] [--soft x y z r g b radius n]\n"
                  "                [--srand S] [--seed S] [--out file.ppm] [--device n] [--gpus N] [--dump-scene prefix]\n");
  return 2;
}

// Primary rays exactly as traceRayThread forms them (flyscene.cpp:301-308), host side only.
int dump_rays(const std::string &prefix, int W, int H, float fovy, const float *view) {
  Tucano::Flycamera cam;
  cam.setPerspectiveMatrix(fovy, W / (float)H, 0.1f, 100.0f);
  cam.setViewport(Eigen::Vector2f((float)W, (float)H));
  if (view) {
    Eigen::Affine3f v;
    std::memcpy(v.m, view, 12 * sizeof(float));
    cam.setViewMatrix(v);
  }
  const Eigen::Vector3f origin = cam.getCenter();
  std::vector<float> o = {origin[0], origin[1], origin[2]}, d;
  for (int j = 0; j < H; ++j)
    for (int i = 0; i < W; ++i) {
      const Eigen::Vector3f dir = (cam.screenToWorld(Eigen::Vector2f((float)i, (float)j)) - origin).normalized();
      for (int c = 0; c < 3; ++c) d.push_back(dir[c]);
    }
  const Eigen::Affine3f vm = cam.getViewMatrix();
  std::vector<float> vmat(vm.data(), vm.data() + 12);
  dump(prefix + ".origin.f32", o); dump(prefix + ".dir.f32", d); dump(prefix + ".view.f32", vmat);
  return 0;
}

int ppm_from_f32(const std::string &in, const std::string &out, int W, int H) {
  FILE *f = fopen(in.c_str(), "rb");
  if (!f) { perror(in.c_str()); return 3; }
  std::vector<float> px(3 * (size_t)W * H);
  const size_t got = fread(px.data(), sizeof(float), px.size(), f);
  fclose(f);
  if (got != px.size()) { fprintf(stderr, "raytrace: %s holds %zu floats, expected %zu\n", in.c_str(), got, px.size()); return 3; }
  std::vector<std::vector<Eigen::Vector3f>> data((size_t)H, std::vector<Eigen::Vector3f>((size_t)W));
  for (int j = 0; j < H; ++j)
    for (int i = 0; i < W; ++i) {
      const float *p = &px[3 * ((size_t)j * W + i)];
      data[(size_t)j][(size_t)i] = Eigen::Vector3f(p[0], p[1], p[2]);
    }
  Tucano::ImageImporter::writePPMImage(out, data);
  return 0;
}

}  // namespace

int main(int argc, char **argv) {
  std::string model = "resources/models/dodgeColorTest.obj", out = "result.ppm", dump_prefix, rays_prefix, f32_in, f32_out, out_p6, out_png, u8_in, u8_out, lights_prefix;
  unsigned long long seed = 0;
  bool seeded = false;
  int W = 1000, H = 1000, depth = 2, device = -1;  // main.cpp:8-9 window size, flyscene.hpp:142
  float fovy = 60.f;
  bool has_view = false, nolights = false, load_only = false, lights_only = false;
  int gpus = 1;
  float view[12];
  std::vector<std::array<float, 6>> lights, dirlights;
  std::vector<std::array<float, 8>> soft;
  auto need = [&](int i, int n) { if (i + n >= argc) exit(usage()); };
  for (int i = 1; i < argc; ++i) {
    std::string a = argv[i];
    if (a == "--model") { need(i, 1); model = argv[++i]; }
    else if (a == "--size") { need(i, 2); W = atoi(argv[++i]); H = atoi(argv[++i]); }
    else if (a == "--depth") { need(i, 1); depth = atoi(argv[++i]); }
    else if (a == "--fovy") { need(i, 1); fovy = (float)atof(argv[++i]); }
    else if (a == "--device") { need(i, 1); device = atoi(argv[++i]); }
    else if (a == "--out") { need(i, 1); out = argv[++i]; }
    else if (a == "--out-p6") { need(i, 1); out_p6 = argv[++i]; }
    else if (a == "--out-png") { need(i, 1); out_png = argv[++i]; }
    else if (a == "--png-from-u8") { need(i, 2); u8_in = argv[++i]; u8_out = argv[++i]; }
    else if (a == "--seed") { need(i, 1); seed = strtoull(argv[++i], nullptr, 10); seeded = true; }
    else if (a == "--dump-lights") { need(i, 1); lights_prefix = argv[++i]; }
    else if (a == "--lights-only") lights_only = true;
    else if (a == "--dump-scene") { need(i, 1); dump_prefix = argv[++i]; }
    else if (a == "--dump-scene-noboxes") { need(i, 1); dump_prefix = argv[++i]; load_only = true; }
    else if (a == "--gpus") { need(i, 1); gpus = atoi(argv[++i]); }
    else if (a == "--srand") { need(i, 1); srand((unsigned)strtoul(argv[++i], nullptr, 10)); }
    else if (a == "--dump-rays") { need(i, 1); rays_prefix = argv[++i]; }
    else if (a == "--ppm-from-f32") { need(i, 2); f32_in = argv[++i]; f32_out = argv[++i]; }
    else if (a == "--nolights") nolights = true;
    else if (a == "--view") { need(i, 12); has_view = true; for (int k = 0; k < 12; ++k) view[k] = strtof(argv[++i], nullptr); }
    else if (a == "--light" || a == "--dirlight") {
      need(i, 6);
      std::array<float, 6> l;
      for (int k = 0; k < 6; ++k) l[(size_t)k] = strtof(argv[++i], nullptr);
      (a == "--light" ? lights : dirlights).push_back(l);
    } else if (a == "--soft") {
      need(i, 8);
      std::array<float, 8> l;
      for (int k = 0; k < 8; ++k) l[(size_t)k] = strtof(argv[++i], nullptr);
      soft.push_back(l);
    } else return usage();
  }
  if (!dump_prefix.empty())
    return scene_dump::dump_scene(model, dump_prefix, 300, load_only,
                                  [](Tucano::Mesh &m, std::vector<Tucano::Material::Mtl> &mt, const std::string &f) { Tucano::MeshImporter::loadObjFile(m, mt, f); },
                                  "parallel");
  if (!rays_prefix.empty()) return dump_rays(rays_prefix, W, H, fovy, has_view ? view : nullptr);
  if (!f32_in.empty()) return ppm_from_f32(f32_in, f32_out, W, H);
  if (!u8_in.empty()) {  // W*H*3 bytes -> PNG (no GPU): the writer's own test hook
    std::vector<unsigned char> px((size_t)W * H * 3);
    FILE *f = fopen(u8_in.c_str(), "rb");
    if (!f || fread(px.data(), 1, px.size(), f) != px.size()) { if (f) fclose(f); return 1; }
    fclose(f);
    return Tucano::ImageImporter::writePNGImage(u8_out, px.data(), W, H) ? 0 : 1;
  }

  Flyscene *flyscene = new Flyscene();  // main.cpp:40-41
  const size_t slash = model.find_last_of('/');
  flyscene->MODEL_DIR = slash == std::string::npos ? "" : model.substr(0, slash + 1);
  flyscene->OBJECT_NAME = slash == std::string::npos ? model : model.substr(slash + 1);
  flyscene->OUTPUT_NAME = out;
  flyscene->DEVICE = device;
  flyscene->DEVICES = gpus;
  flyscene->max_depth = depth;
  if (!lights_only) flyscene->initialize(W, H);
  if (!lights_only && !flyscene->ok()) { fprintf(stderr, "raytrace: %s\n", flyscene->lastError().c_str()); return 1; }
  flyscene->getCamera()->setPerspectiveMatrix(fovy, W / (float)H, 0.1f, 100.0f);
  if (has_view) {
    Eigen::Affine3f v;
    std::memcpy(v.m, view, sizeof view);
    flyscene->getCamera()->setViewMatrix(v);
  }
  if (nolights || !lights.empty() || !dirlights.empty() || !soft.empty()) flyscene->lights.clear();  // key C, main.cpp:72
  for (auto &l : lights) flyscene->lights.push_back({Eigen::Vector3f(l[0], l[1], l[2]), Eigen::Vector3f(l[3], l[4], l[5])});
  for (auto &l : dirlights)
    flyscene->dirLights.push_back(std::make_tuple(Eigen::Vector3f(l[0], l[1], l[2]), Eigen::Vector3f(l[3], l[4], l[5]), Eigen::Quaternion<float>()));
  for (auto &l : soft) {  // what addLight's "s" branch does (flyscene.cpp:213-236)
    std::pair<Eigen::Vector3f, Eigen::Vector3f> light(Eigen::Vector3f(l[0], l[1], l[2]), Eigen::Vector3f(l[3], l[4], l[5]));
    if (seeded) flyscene->sphericalLightSeeded(light, l[6], (int)l[7], seed);
    else flyscene->sphericalLight(light, l[6], (int)l[7]);
    light.second /= (float)((int)l[7] + 1);
    flyscene->lights.push_back(light);
  }
  if (!lights_prefix.empty()) {  // the light set the frame is rendered with (tests: seeded soft lights are reproducible)
    std::vector<float> lf;
    for (auto &l : flyscene->lights) { for (int c = 0; c < 3; ++c) lf.push_back(l.first[c]); for (int c = 0; c < 3; ++c) lf.push_back(l.second[c]); }
    dump(lights_prefix + ".lights.f32", lf);
  }
  if (lights_only) {  // --dump-lights without a scene or a device (CPU tests of Flyscene::sphericalLight)
    delete flyscene;
    return 0;
  }
  flyscene->OUTPUT_BINARY_NAME = out_p6;
  flyscene->OUTPUT_PNG_NAME = out_png;
  flyscene->raytraceScene();  // key T
  const bool ok = flyscene->lastError().empty();
  if (!ok) fprintf(stderr, "raytrace: %s\n", flyscene->lastError().c_str());
  delete flyscene;
  return ok ? 0 : 1;
}
