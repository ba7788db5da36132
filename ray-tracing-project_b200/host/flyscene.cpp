// flyscene.cpp - see flyscene.hpp.  Host side of the drop-in: load + flatten inputs, call the C ABI
// (include/rtp.h), write the output image.  No ray is traced on the CPU in this file.
#include "flyscene.hpp"

#include <cmath>
#include <cstring>
#include <ctime>
#include <iostream>

namespace {

void put3(float *dst, const Eigen::Vector3f &v) { dst[0] = v[0]; dst[1] = v[1]; dst[2] = v[2]; }

}  // namespace

bool Flyscene::fail(const std::string &what) {
  error = what + ": " + rtp_last_error();
  std::cerr << "Flyscene: " << error << std::endl;
  return false;
}

Flyscene::~Flyscene() {  // flyscene.cpp:439-441
  if (scene) rtp_scene_destroy(scene);
  scene = nullptr;
  BoundingBox::deconstruct();
}

void Flyscene::initialize(int width, int height) {
  flycamera.setPerspectiveMatrix(60.0, width / (float)height, 0.1f, 100.0f);  // flyscene.cpp:14
  flycamera.setViewport(Eigen::Vector2f((float)width, (float)height));        // :15
  Tucano::MeshImporter::loadObjFile(mesh, materials, MODEL_DIR + OBJECT_NAME);  // :18
  mesh.normalizeModelMatrix();                                                // :22
  lights.push_back(std::pair<Eigen::Vector3f, Eigen::Vector3f>(Eigen::Vector3f(-0.5, 2.0, 3.0), Eigen::Vector3f(1, 1, 1)));  // :37
  if (mesh.getNumberOfFaces() == 0) {
    error = "no faces loaded from " + MODEL_DIR + OBJECT_NAME;
    std::cerr << "Flyscene: " << error << std::endl;
    return;
  }
  generateBoundingBoxes();  // :62
  if (!uploadScene()) return;
  createDebugRay(Eigen::Vector2f((float)(width / 2.0), (float)(height / 2.0)));  // :63
}

void Flyscene::generateBoundingBoxes() {  // flyscene.cpp:399-428
  BoundingBox::deconstruct();
  BoundingBox::mesh = &mesh;
  BoundingBox *box = new BoundingBox(true);
  box->fitMesh();
  bool notDone = true;
  while (notDone && BoundingBox::boxes.size() < (size_t)MAX_BOXES) {
    notDone = false;
    std::vector<BoundingBox *> current = BoundingBox::boxes;
    std::cout << "Now have " << current.size() << " boxes." << std::endl;
    for (BoundingBox *b : current) {
      if (b->getNumberOfFaces() > MIN_FACES && (!b->failed[0] || !b->failed[1] || !b->failed[2])) {
        BoundingBox *newBox = b->splitBox();
        while (b == newBox) newBox = b->splitBox();
        notDone = true;
      }
    }
  }
  std::cout << "Boxes: " << BoundingBox::boxes.size() << std::endl;
}

// Mesh, materials and the BoundingBox list -> rtp_mesh -> device (replaces nothing in the
// reference: its ray tracer reads these structures in place).
bool Flyscene::uploadScene() {
  const int V = mesh.getNumberOfVertices(), F = mesh.getNumberOfFaces(), M = (int)materials.size();
  std::vector<float> verts(4 * (size_t)V), normals(3 * (size_t)V, 0.f), fnormal(3 * (size_t)F), mats(12 * (size_t)std::max(M, 1));
  std::vector<int32_t> faces(3 * (size_t)F), fmat((size_t)F);
  for (int i = 0; i < V; ++i) {
    const Eigen::Vector4f v = mesh.getVertex(i);
    for (int c = 0; c < 4; ++c) verts[4 * (size_t)i + c] = v[c];
    put3(&normals[3 * (size_t)i], mesh.getNormal(i));
  }
  for (int i = 0; i < F; ++i) {
    const Tucano::Face &f = mesh.getFace(i);
    for (int c = 0; c < 3; ++c) faces[3 * (size_t)i + c] = (int32_t)f.vertex_ids[(size_t)c];
    fmat[(size_t)i] = f.material_id;
    put3(&fnormal[3 * (size_t)i], f.normal);
  }
  for (int i = 0; i < M; ++i) {
    float *p = &mats[12 * (size_t)i];
    put3(p, materials[(size_t)i].getAmbient());
    put3(p + 3, materials[(size_t)i].getDiffuse());
    put3(p + 6, materials[(size_t)i].getSpecular());
    p[9] = materials[(size_t)i].getShininess();
    p[10] = materials[(size_t)i].getOpticalDensity();
    p[11] = materials[(size_t)i].getDissolveFactor();
  }
  std::vector<float> bounds;
  std::vector<int> box_off, box_faces;
  BoundingBox::exportBoxes(bounds, box_off, box_faces);

  rtp_mesh m;
  std::memset(&m, 0, sizeof m);
  m.n_verts = V; m.n_faces = F; m.n_mats = M; m.n_boxes = (int)BoundingBox::boxes.size();
  m.verts = verts.data(); m.normals = normals.data(); m.faces = faces.data(); m.face_mat = fmat.data();
  m.face_normal = fnormal.data(); m.materials = mats.data();
  const Eigen::Affine3f shape = mesh.getShapeModelMatrix();
  std::memcpy(m.shape, shape.data(), sizeof m.shape);
  m.boxes = bounds.data(); m.box_off = box_off.data(); m.box_faces = box_faces.data();
  m.min_faces = MIN_FACES; m.max_boxes = MAX_BOXES;
  if (scene) rtp_scene_destroy(scene);
  scene = nullptr;
  if (DEVICES <= 1) {
    if (rtp_scene_create(&m, DEVICE, &scene) != RTP_OK) return fail("rtp_scene_create");
    return true;
  }
  const int ndev = rtp_device_count();
  if (ndev <= 0) return fail("rtp_device_count");
  std::vector<int> devs((size_t)DEVICES);
  for (int r = 0; r < DEVICES; ++r) devs[(size_t)r] = ((DEVICE < 0 ? 0 : DEVICE) + r) % ndev;
  if (rtp_scene_create_multi(&m, devs.data(), DEVICES, &scene) != RTP_OK) return fail("rtp_scene_create_multi");
  return true;
}

bool Flyscene::pushCameraAndLights() {
  if (!scene) {
    error = "no device scene (initialize() failed or was not called)";
    std::cerr << "Flyscene: " << error << std::endl;
    return false;
  }
  rtp_camera c;
  const Eigen::Affine3f view = flycamera.getViewMatrix();
  std::memcpy(c.view, view.data(), sizeof c.view);
  const Eigen::Vector4f vp = flycamera.getViewport();
  for (int k = 0; k < 4; ++k) c.viewport[k] = vp[k];
  c.fovy = flycamera.getFovy();
  c.aspect = flycamera.getAspectRatio();
  if (rtp_set_camera(scene, &c) != RTP_OK) return fail("rtp_set_camera");
  std::vector<float> pl(6 * lights.size()), dl(6 * dirLights.size());
  for (size_t i = 0; i < lights.size(); ++i) { put3(&pl[6 * i], lights[i].first); put3(&pl[6 * i + 3], lights[i].second); }
  for (size_t i = 0; i < dirLights.size(); ++i) { put3(&dl[6 * i], std::get<0>(dirLights[i])); put3(&dl[6 * i + 3], std::get<1>(dirLights[i])); }
  rtp_lights l;
  l.n_point = (int)lights.size(); l.n_dir = (int)dirLights.size();
  l.point = pl.empty() ? nullptr : pl.data();
  l.dir = dl.empty() ? nullptr : dl.data();
  if (rtp_set_lights(scene, &l) != RTP_OK) return fail("rtp_set_lights");
  return true;
}

void Flyscene::raytraceScene(int width, int height) {  // flyscene.cpp:250-297
  std::cout << "ray tracing ..." << std::endl;
  const std::clock_t wall0 = std::clock();
  const time_t start_time = time(NULL);
  if (width == 0 || height == 0) {  // flyscene.cpp:255
    raytracing_image_size = flycamera.getViewportSize();
  } else {
    raytracing_image_size = Eigen::Vector2i(width, height);
  }
  const int W = raytracing_image_size[0], H = raytracing_image_size[1];
  if (W <= 0 || H <= 0 || !pushCameraAndLights()) return;
  rtp_render_opts o;
  std::memset(&o, 0, sizeof o);
  o.width = W; o.height = H; o.max_depth = max_depth;
  put3(o.background, BACKGROUND_COLOR);
  o.rank = 0; o.world = 1; o.flags = RTP_RENDER_DEFAULT;
  std::vector<float> rgb(3 * (size_t)W * H);
  primary_hits.assign((size_t)W * H, -1);
  image8.assign(3 * (size_t)W * H, 0);
  if (rtp_render(scene, &o, rgb.data(), image8.data(), primary_hits.data(), &last_stats) != RTP_OK) {
    fail("rtp_render");
    return;
  }
  pixel_data.resize((size_t)H);
  for (int j = 0; j < H; ++j) {
    pixel_data[(size_t)j].resize((size_t)W);
    const float *row = &rgb[3 * (size_t)j * W];
    for (int i = 0; i < W; ++i) pixel_data[(size_t)j][(size_t)i] = Eigen::Vector3f(row[3 * i], row[3 * i + 1], row[3 * i + 2]);
  }
  PIXEL_COUNT = H;
  if (!OUTPUT_NAME.empty()) Tucano::ImageImporter::writePPMImage(OUTPUT_NAME, pixel_data);  // :292
  if (!OUTPUT_BINARY_NAME.empty()) Tucano::ImageImporter::writePPMImageBinary(OUTPUT_BINARY_NAME, image8.data(), W, H);
  if (!OUTPUT_PNG_NAME.empty()) Tucano::ImageImporter::writePNGImage(OUTPUT_PNG_NAME, image8.data(), W, H);
  const time_t end_time = time(NULL);
  std::cout << "ray tracing done! " << std::endl;
  std::cout << "Runtime: " << (end_time - start_time) << " seconds (device " << last_stats.ms_total << " ms, "
            << (last_stats.rays_primary + last_stats.rays_shadow + last_stats.rays_reflection) << " rays, host "
            << 1e3 * (double)(std::clock() - wall0) / CLOCKS_PER_SEC << " ms cpu)" << std::endl;
}

void Flyscene::traceRayThread(int h, int w, int start, int stop, std::vector<std::vector<Eigen::Vector3f>> &out) {
  // One batch of primary rays for the rows [start, stop): flyscene.cpp:299-314 with i < w (Q1).
  (void)h;
  if (stop <= start || w <= 0) return;
  Eigen::Vector3f origin = flycamera.getCenter();
  std::vector<float> rays(6 * (size_t)(stop - start) * w), rgb(3 * (size_t)(stop - start) * w);
  size_t k = 0;
  for (int j = start; j < stop; ++j)
    for (int i = 0; i < w; ++i, ++k) {
      Eigen::Vector3f screen_coords = flycamera.screenToWorld(Eigen::Vector2f((float)i, (float)j));
      put3(&rays[6 * k], origin);
      put3(&rays[6 * k + 3], (screen_coords - origin).normalized());
    }
  if (!traceRays(rays.data(), (long long)k, 0, rgb.data())) return;
  k = 0;
  for (int j = start; j < stop; ++j) {
    if ((int)out.size() <= j) out.resize((size_t)j + 1);
    if ((int)out[(size_t)j].size() < w) out[(size_t)j].resize((size_t)w);
    for (int i = 0; i < w; ++i, ++k) out[(size_t)j][(size_t)i] = Eigen::Vector3f(rgb[3 * k], rgb[3 * k + 1], rgb[3 * k + 2]);
    PIXEL_COUNT++;
  }
}

bool Flyscene::traceRays(const float *rays6, long long n, int depth, float *rgb3) {
  if (n <= 0) return true;
  if (depth >= max_depth) {  // flyscene.cpp:318-320
    std::memset(rgb3, 0, sizeof(float) * 3 * (size_t)n);
    return true;
  }
  if (!pushCameraAndLights()) return false;
  // below depth 0 a miss is black, not the background (flyscene.cpp:327-332)
  float bg[3] = {0.f, 0.f, 0.f};
  if (depth == 0) put3(bg, BACKGROUND_COLOR);
  if (rtp_trace_rays(scene, rays6, n, RTP_RAYS_SHADE, max_depth - depth, bg, rgb3, nullptr, nullptr, nullptr, &last_stats) != RTP_OK)
    return fail("rtp_trace_rays");
  return true;
}

Eigen::Vector3f Flyscene::traceRay(Eigen::Vector3f &origin, Eigen::Vector3f &dir, int depth, bool debug) {
  (void)debug;
  float ray[6], rgb[3] = {0.f, 0.f, 0.f};
  put3(ray, origin);
  put3(ray + 3, dir);
  traceRays(ray, 1, depth, rgb);
  return Eigen::Vector3f(rgb[0], rgb[1], rgb[2]);
}

std::tuple<Tucano::Face, Eigen::Vector3f, float> Flyscene::calculateMinimumFace(Eigen::Vector3f &origin, Eigen::Vector3f dir, bool debug) {
  (void)debug;
  Tucano::Face none;
  std::tuple<Tucano::Face, Eigen::Vector3f, float> miss = std::make_tuple(none, Eigen::Vector3f::Zero(), INFINITY);
  if (!scene) return miss;
  float ray[6], t = INFINITY;
  int32_t face = -1;
  put3(ray, origin);
  put3(ray + 3, dir);
  if (rtp_trace_rays(scene, ray, 1, RTP_RAYS_CLOSEST, 1, nullptr, nullptr, &face, &t, nullptr, nullptr) != RTP_OK) {
    fail("rtp_trace_rays");
    return miss;
  }
  if (face < 0) return miss;
  return std::make_tuple(mesh.getFace(face), origin + t * dir, t);  // flyscene.cpp:464
}

Eigen::Vector3f Flyscene::reflect(Eigen::Vector3f direction, Eigen::Vector3f normal) {  // flyscene.cpp:480-482
  return (direction - 2 * (direction.dot(normal)) * normal).normalized();
}

bool Flyscene::shadow(Eigen::Vector3f &dest, Eigen::Vector3f &light) {
  if (!scene) return false;
  float ray[6];
  uint8_t occluded = 0;
  put3(ray, dest);
  put3(ray + 3, light);
  if (rtp_trace_rays(scene, ray, 1, RTP_RAYS_ANYHIT, 1, nullptr, nullptr, nullptr, nullptr, &occluded, nullptr) != RTP_OK) {
    fail("rtp_trace_rays");
    return false;
  }
  return occluded != 0;
}

void Flyscene::sphericalLight(std::pair<Eigen::Vector3f, Eigen::Vector3f> light, float radius, int nLightpoints) {  // flyscene.cpp:529-538
  Eigen::Vector3f lightLoc = light.first;
  for (int i = 0; i < nLightpoints; i++) {
    float x = (-radius + (rand() / (RAND_MAX / (radius * 2))));
    float y = (-radius + (rand() / (RAND_MAX / (radius * 2))));
    float z = (-radius + (rand() / (RAND_MAX / (radius * 2))));
    lights.push_back(std::pair<Eigen::Vector3f, Eigen::Vector3f>(Eigen::Vector3f(lightLoc.x() + x, lightLoc.y() + y, lightLoc.z() + z),
                                                                 light.second / (float)(nLightpoints + 1)));
  }
}

namespace {
// SplitMix64 finaliser as a counter-based generator: value k of stream `seed` is a pure function of (seed, k).
inline unsigned long long mix64(unsigned long long x) {
  x += 0x9e3779b97f4a7c15ull;
  x = (x ^ (x >> 30)) * 0xbf58476d1ce4e5b9ull;
  x = (x ^ (x >> 27)) * 0x94d049bb133111ebull;
  return x ^ (x >> 31);
}
}  // namespace

void Flyscene::sphericalLightSeeded(std::pair<Eigen::Vector3f, Eigen::Vector3f> light, float radius, int nLightpoints, unsigned long long seed) {
  Eigen::Vector3f lightLoc = light.first;
  for (int i = 0; i < nLightpoints; i++) {
    float j[3];
    for (int c = 0; c < 3; ++c) {
      // same cube [-radius, radius)^3 as sphericalLight; 24 random bits -> [0, 1)
      const unsigned long long r = mix64(seed * 0x100000001b3ull + (unsigned long long)(3 * i + c));
      j[c] = -radius + (radius * 2) * ((float)(r >> 40) * (1.0f / 16777216.0f));
    }
    lights.push_back(std::pair<Eigen::Vector3f, Eigen::Vector3f>(Eigen::Vector3f(lightLoc.x() + j[0], lightLoc.y() + j[1], lightLoc.z() + j[2]),
                                                                 light.second / (float)(nLightpoints + 1)));
  }
}

void Flyscene::createDebugRay(const Eigen::Vector2f &mouse_pos) {  // flyscene.cpp:129-173 without the drawing
  resetDebugRay();
  Eigen::Vector3f screen_pos = flycamera.screenToWorld(mouse_pos);
  Eigen::Vector3f origin = flycamera.getCenter();
  Eigen::Vector3f dir = (screen_pos - origin).normalized();
  debugChain.push_back(origin);
  std::tuple<Tucano::Face, Eigen::Vector3f, float> hit = calculateMinimumFace(origin, dir, RENDER_DEBUG_BOXES);
  if (std::get<2>(hit) == INFINITY) return;
  debugChain.push_back(std::get<1>(hit));
}
