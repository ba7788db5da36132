// flyscene.hpp - drop-in mirror of the reference's scene controller (src/flyscene.hpp:22-186) for
// its one data-parallel path: raytraceScene / traceRay and everything under them
// (src/flyscene.cpp:250-614).  Same class name, same public members and signatures; the work is
// done by the CUDA library behind include/rtp.h (libflyscene_b200.so) - there is no CPU ray
// tracer in here and every compute call fails loudly when no CUDA device is present.
//
// What differs from the reference, on purpose:
//   * raytraceScene renders the full W x H frame for any size (the reference's thread loop covers
//     an H x H square, rejects H % 20 != 0 and spins forever below 1000 rows: Q1, Q2);
//   * the compile-time constants OBJECT_NAME, MIN_FACES, MAX_BOXES and max_depth are ordinary
//     members with the same defaults, settable before initialize() / raytraceScene();
//   * material state is per ray, not shared mutable members (Q4): identical results to the
//     reference run single-threaded, no data race;
//   * GL preview members (paintGL, simulate, debug-ray drawing) are declared and empty:
//     no graphics stack on this box.
#ifndef __FLYSCENE__
#define __FLYSCENE__

#include <climits>
#include <string>
#include <tuple>
#include <utility>
#include <vector>

#include "../../include/rtp.h"
#include "BoundingBox.h"
#include "tucano_lite.hpp"

struct GLFWwindow;  // never dereferenced: simulate() keeps the reference's signature

class Flyscene {
 public:
  Flyscene(void) {}
  virtual ~Flyscene();

  // Loads MODEL_DIR + OBJECT_NAME, normalises the model matrix, adds the default light, builds
  // the BoundingBox list and uploads the flattened scene (flyscene.cpp:9-64).
  void initialize(int width, int height);

  virtual void paintGL() {}                 // GL preview: unavailable on this box
  virtual void simulate(GLFWwindow *) {}    // keyboard camera control needs a window

  Tucano::Flycamera *getCamera(void) { return &flycamera; }

  // The reference's interactive stdin light prompt (flyscene.cpp:182-248) belongs to its GLFW key loop
  // (main.cpp:70-71), which is out of scope: declared for source compatibility, does nothing.
  // Lights are set through the public `lights` / `dirLights` members and sphericalLight().
  void addLight() {}

  // Traces the ray through the pixel under mouse_pos and keeps its hit chain in debugChain
  // (the reference draws it as cylinders; flyscene.cpp:129-173).
  void createDebugRay(const Eigen::Vector2f &mouse_pos);
  void resetDebugRay() { debugChain.clear(); }

  // THE render entry point (flyscene.cpp:250-297): traces the frame on the GPU(s), fills
  // pixel_data / primary_hits and writes OUTPUT_NAME ("result.ppm", ASCII P3).
  void raytraceScene(int width = 0, int height = 0);

  // One ray through the same pipeline as raytraceScene (flyscene.cpp:317-371).
  Eigen::Vector3f traceRay(Eigen::Vector3f &origin, Eigen::Vector3f &dir, int depth, bool debug);
  // Batched form of the same call: n rays (6 floats each), n rgb triples out.
  bool traceRays(const float *rays6, long long n, int depth, float *rgb3);

  // Rows [start, stop) of the frame, ALL w columns (flyscene.cpp:299-314; see Q1 above).
  void traceRayThread(int h, int w, int start, int stop, std::vector<std::vector<Eigen::Vector3f>> &pixel_data);

  // Closest face along a ray: the face, the hit point and the distance (flyscene.cpp:373-396).
  // No hit: a default Face (material_id -1, no vertex ids), point 0, t = INFINITY.
  std::tuple<Tucano::Face, Eigen::Vector3f, float> calculateMinimumFace(Eigen::Vector3f &origin, Eigen::Vector3f dir, bool debug);

  Eigen::Vector3f reflect(Eigen::Vector3f direction, Eigen::Vector3f normal);  // flyscene.cpp:480-482

  // Is the light direction `light` blocked as seen from `dest`? (flyscene.cpp:510-526)
  bool shadow(Eigen::Vector3f &dest, Eigen::Vector3f &light);

  // n extra point lights jittered in a cube of half-edge `radius` with libc rand(), each carrying
  // rgb / (n + 1) (flyscene.cpp:529-538).
  void sphericalLight(std::pair<Eigen::Vector3f, Eigen::Vector3f> lightLoc, float radius, int nLightpoints);
  // Same light set from a counter-based generator instead of libc rand() (a documented
  // deviation: the reference's soft shadows differ from run to run and from libc to libc; these do not).
  // Point i of the set depends on (seed, i) only.  Every point becomes one more shadow ray per hit; the
  // device traces the shadow rays of ALL lights of a bounce level in one any-hit launch.
  void sphericalLightSeeded(std::pair<Eigen::Vector3f, Eigen::Vector3f> lightLoc, float radius, int nLightpoints, unsigned long long seed);

  // ---- results of the last raytraceScene (the reference keeps them local and only writes the file)
  const std::vector<std::vector<Eigen::Vector3f>> &getPixelData() const { return pixel_data; }
  const std::vector<unsigned char> &getImage8() const { return image8; }  // W*H*3, the values result.ppm holds (device-quantised)
  const std::vector<int> &getPrimaryHits() const { return primary_hits; }  // W*H face ids, -1 = background
  const rtp_stats &getLastStats() const { return last_stats; }
  rtp_scene *getDeviceScene() { return scene; }
  Tucano::Mesh &getMesh() { return mesh; }
  std::vector<Tucano::Material::Mtl> &getMaterials() { return materials; }
  bool ok() const { return scene != nullptr; }
  const std::string &lastError() const { return error; }

 private:
  int PIXEL_COUNT = 0;
  Tucano::Flycamera flycamera;
  Eigen::Vector2i raytracing_image_size;

 public:  // so we can remove all lights
  // light sources for ray tracing: (position, rgb)
  std::vector<std::pair<Eigen::Vector3f, Eigen::Vector3f>> lights;
  // the directional lights: (vector used verbatim as L - Q9, rgb, orientation)
  std::vector<std::tuple<Eigen::Vector3f, Eigen::Vector3f, Eigen::Quaternion<float>>> dirLights;

  // Depth of recursive raytracing (reference: const static 2 => primary + one bounce)
  int max_depth = 2;

 private:
  Tucano::Mesh mesh;
  std::vector<Tucano::Material::Mtl> materials = std::vector<Tucano::Material::Mtl>();
  rtp_scene *scene = nullptr;
  std::string error;
  std::vector<std::vector<Eigen::Vector3f>> pixel_data;
  std::vector<int> primary_hits;
  std::vector<unsigned char> image8;
  rtp_stats last_stats = rtp_stats();

  void generateBoundingBoxes();
  bool uploadScene();
  bool pushCameraAndLights();
  bool fail(const std::string &what);

 public:
  std::string OBJECT_NAME = "dodgeColorTest.obj";
  std::string MODEL_DIR = "resources/models/";  // cwd-relative, as flyscene.cpp:18
  std::string OUTPUT_NAME = "result.ppm";       // flyscene.cpp:292; empty: no file
  std::string OUTPUT_PNG_NAME = "";             // non-empty: also write an 8-bit RGB PNG of it
  std::string OUTPUT_BINARY_NAME = "";          // non-empty: also write a binary P6 PPM of the device-quantised image
  int DEVICE = -1;                               // CUDA ordinal of the first GPU (-1: current device)
  // GPUs the frame is fanned out over - the role Flyscene::THREADS row bands play in the reference (flyscene.cpp:266-289):
  // interleaved 32x32 screen tiles, rank r on device (DEVICE + r) % device count (more ranks than devices: ranks share
  // a GPU), every rank storing its pixels into the first GPU's image.  Same image for every value.  Set before initialize().
  int DEVICES = 1;

  static const bool RENDER_BOUNDINGBOXES = false;
  static const bool RENDER_BOUNDINGBOX_COLORED_TRIANGLES = false;
  const bool RENDER_DEBUG_BOXES = false;
  int MIN_FACES = 300;
  int MAX_BOXES = INT_MAX;

  std::vector<BoundingBox *> debugBoxes;
  std::vector<Eigen::Vector3f> debugChain;  // origin, hit point 0, hit point 1, ... of the debug ray

  static const int THREADS = 20;  // kept for source compatibility; the GPU path has no use for it

  Eigen::Vector3f BACKGROUND_COLOR = Eigen::Vector3f(0.9, 0.9, 0.9);
  const Eigen::Vector3f FOREGROUND_COLOR = Eigen::Vector3f(0, 0, 1);

  // Default material (faces with material_id == -1); read by the library's flattener as the
  // constants of flyscene.hpp:179-184.
  Eigen::Vector3f ka = Eigen::Vector3f(0.2, 0.2, 0.2);
  Eigen::Vector3f kd = Eigen::Vector3f(0.9, 0.9, 0);
  Eigen::Vector3f ks = Eigen::Vector3f(0, 0, 0);
  float shininess = 0.f;
  float refraction_index = 0.f;
  float transparency = 0.f;
};

#endif  // FLYSCENE
