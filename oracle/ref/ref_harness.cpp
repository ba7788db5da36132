// TEST INFRASTRUCTURE ONLY (oracle/): headless driver around the UNMODIFIED reference
// ray tracer (/root/reference/src/flyscene.cpp + BoundingBox.cpp, compiled in place by
// oracle/Makefile).  It loads an OBJ through the reference's own loader, sets camera /
// lights through the reference's public members, and calls the reference's own
// Flyscene::traceRay for every requested pixel (the reference's
// raytraceScene driver loop only covers an HxH square and rejects H % 20 != 0, so the
// per-pixel loop of traceRayThread, flyscene.cpp:299-314, is re-driven here).
//
// Nothing in the product links or imports this file.  Only tests/, smoke() and the
// bench.py reference/cpu_baseline legs execute the resulting binary.
//
// Usage: ref_harness JOBFILE      (job grammar: see parse_job below)
#include <sys/mman.h>
#include <sys/stat.h>
#include <sys/wait.h>
#include <unistd.h>

#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <sstream>
#include <string>
#include <vector>

#include "flyscene.hpp"  // the patched copy (max_depth made a runtime static), see Makefile

int Flyscene::max_depth = 2;

namespace {

struct Job {
  std::string obj;          // absolute path of the OBJ
  std::string out;          // output prefix
  int w = 256, h = 256;
  float fovy = 60.f;
  bool has_view = false;
  float view[12];           // row-major 3x4 affine view matrix
  int max_depth = 2;
  bool lights_given = false;
  std::vector<std::array<float, 6>> lights, dirlights;
  // soft lights: x y z r g b radius n -> the reference's own Flyscene::sphericalLight (flyscene.cpp:529-538, libc
  // rand()) followed by what the "s" branch of its addLight does with the light itself (flyscene.cpp:213-236)
  std::vector<std::array<float, 8>> soft;
  bool has_srand = false;
  unsigned srand_seed = 1;
  // pixel selection
  int stride_x = 1, stride_y = 1, off_x = 0, off_y = 0;
  std::string pixel_file;   // int32 (i,j) pairs, overrides strides
  std::string ray_file;     // float32 (ox,oy,oz,dx,dy,dz) per ray -> mode "rays"
  std::string mode = "render";  // render | anyhit | closest
  int procs = 1;
  bool dump_scene = false;
  bool chain = true;        // also walk the hit chain for hit index / ray counts
  int repeat = 1;           // timed passes over the pixel list (bench: warm-up + steps)
};

bool parse_job(const char* path, Job& j) {
  std::ifstream in(path);
  if (!in) return false;
  std::string line;
  while (std::getline(in, line)) {
    std::istringstream s(line);
    std::string k;
    if (!(s >> k) || k[0] == '#') continue;
    if (k == "obj") s >> j.obj;
    else if (k == "out") s >> j.out;
    else if (k == "size") s >> j.w >> j.h;
    else if (k == "fovy") s >> j.fovy;
    else if (k == "view") { j.has_view = true; for (int i = 0; i < 12; ++i) s >> j.view[i]; }
    else if (k == "max_depth") s >> j.max_depth;
    else if (k == "nolights") j.lights_given = true;
    else if (k == "light") { std::array<float, 6> l; for (auto& v : l) s >> v; j.lights.push_back(l); j.lights_given = true; }
    else if (k == "dirlight") { std::array<float, 6> l; for (auto& v : l) s >> v; j.dirlights.push_back(l); j.lights_given = true; }
    else if (k == "soft") { std::array<float, 8> l; for (auto& v : l) s >> v; j.soft.push_back(l); j.lights_given = true; }
    else if (k == "srand") { s >> j.srand_seed; j.has_srand = true; }
    else if (k == "stride") s >> j.stride_x >> j.stride_y >> j.off_x >> j.off_y;
    else if (k == "pixel_file") s >> j.pixel_file;
    else if (k == "ray_file") s >> j.ray_file;
    else if (k == "mode") s >> j.mode;
    else if (k == "procs") s >> j.procs;
    else if (k == "dump_scene") { int v; s >> v; j.dump_scene = v != 0; }
    else if (k == "chain") { int v; s >> v; j.chain = v != 0; }
    else if (k == "repeat") s >> j.repeat;
    else { fprintf(stderr, "ref_harness: unknown key '%s'\n", k.c_str()); return false; }
  }
  return !j.obj.empty() && !j.out.empty();
}

template <class T>
void write_bin(const std::string& path, const T* p, size_t n) {
  FILE* f = fopen(path.c_str(), "wb");
  if (!f) { perror(path.c_str()); exit(3); }
  if (n) fwrite(p, sizeof(T), n, f);
  fclose(f);
}

template <class T>
std::vector<T> read_bin(const std::string& path) {
  std::vector<T> v;
  FILE* f = fopen(path.c_str(), "rb");
  if (!f) { perror(path.c_str()); exit(3); }
  fseek(f, 0, SEEK_END);
  long sz = ftell(f);
  fseek(f, 0, SEEK_SET);
  v.resize(sz / sizeof(T));
  if (sz) { size_t got = fread(v.data(), sizeof(T), v.size(), f); (void)got; }
  fclose(f);
  return v;
}

template <class T>
T* shared_alloc(size_t n) {
  void* p = mmap(nullptr, std::max<size_t>(n, 1) * sizeof(T), PROT_READ | PROT_WRITE,
                 MAP_SHARED | MAP_ANONYMOUS, -1, 0);
  if (p == MAP_FAILED) { perror("mmap"); exit(3); }
  return static_cast<T*>(p);
}

// Primary-hit face index: the loop of calculateMinimumFace (flyscene.cpp:381-393) driven
// from outside so that the winning face's index is known.  Uses the reference's own
// intersectBox and calculateDistance.
int closest_face_index(Flyscene& fs, Eigen::Vector3f& o, Eigen::Vector3f& d, float& t_out,
                       Eigen::Vector3f& p_out) {
  float best = INFINITY;
  int best_face = -1;
  Tucano::Face* base = &fs.mesh.getFace(0);
  for (BoundingBox* box : BoundingBox::boxes) {
    if (!fs.intersectBox(o, d, *box)) continue;
    for (size_t i = 0; i < box->faces.size(); ++i) {
      Tucano::Face cur = *box->faces[i];
      std::pair<Eigen::Vector3f, float> r = fs.calculateDistance(o, d, cur);
      if (0 <= r.second && r.second < best) {
        best = r.second;
        p_out = r.first;
        best_face = int(box->faces[i] - base);
      }
    }
  }
  t_out = best;
  return best_face;
}

}  // namespace

int main(int argc, char** argv) {
  if (argc < 2) { fprintf(stderr, "usage: ref_harness JOBFILE\n"); return 2; }
  Job job;
  if (!parse_job(argv[1], job)) { fprintf(stderr, "ref_harness: bad job file\n"); return 2; }

  // The reference opens "resources/models/" + OBJECT_NAME relative to the cwd
  // (flyscene.cpp:18-19).  Build that layout in a scratch directory with symlinks.
  char tmpl[] = "/tmp/ref_harness_XXXXXX";
  char* scratch = mkdtemp(tmpl);
  if (!scratch) { perror("mkdtemp"); return 3; }
  std::string sdir(scratch);
  mkdir((sdir + "/resources").c_str(), 0755);
  std::string objdir = job.obj.substr(0, job.obj.find_last_of('/'));
  std::string objname = job.obj.substr(job.obj.find_last_of('/') + 1);
  if (symlink(objdir.c_str(), (sdir + "/resources/models").c_str()) != 0) { perror("symlink"); return 3; }
  if (chdir(sdir.c_str()) != 0) { perror("chdir"); return 3; }

  Flyscene::max_depth = job.max_depth;
  Flyscene* fsp = new Flyscene();
  Flyscene& fs = *fsp;
  const_cast<std::string&>(fs.OBJECT_NAME) = objname;
  // initialize() ends with createDebugRay (one traceRay); harmless.
  fs.initialize(job.w, job.h);
  fs.flycamera.setPerspectiveMatrix(job.fovy, job.w / (float)job.h, 0.1f, 100.0f);
  fs.flycamera.setViewport(Eigen::Vector2f((float)job.w, (float)job.h));
  if (job.has_view) {
    Eigen::Affine3f v = Eigen::Affine3f::Identity();
    for (int r = 0; r < 3; ++r)
      for (int c = 0; c < 4; ++c) v.matrix()(r, c) = job.view[r * 4 + c];
    fs.flycamera.setViewMatrix(v);
  }
  if (job.lights_given) {
    fs.lights.clear();
    fs.dirLights.clear();
    for (auto& l : job.lights)
      fs.lights.push_back({Eigen::Vector3f(l[0], l[1], l[2]), Eigen::Vector3f(l[3], l[4], l[5])});
    for (auto& l : job.dirlights)
      fs.dirLights.push_back(std::make_tuple(Eigen::Vector3f(l[0], l[1], l[2]),
                                             Eigen::Vector3f(l[3], l[4], l[5]),
                                             Eigen::Quaternionf::Identity()));
  }
  if (job.has_srand) srand(job.srand_seed);
  for (auto& l : job.soft) {
    std::pair<Eigen::Vector3f, Eigen::Vector3f> light(Eigen::Vector3f(l[0], l[1], l[2]), Eigen::Vector3f(l[3], l[4], l[5]));
    const int n = (int)l[7];
    fs.sphericalLight(light, l[6], n);  // the reference's own code
    light.second /= (n + 1);            // flyscene.cpp:234
    fs.lights.push_back(light);         // flyscene.cpp:237
  }
  {  // the light list the frame is traced with (Flyscene::lights, in order)
    std::vector<float> lf;
    for (auto& l : fs.lights) { for (int c = 0; c < 3; ++c) lf.push_back(l.first(c)); for (int c = 0; c < 3; ++c) lf.push_back(l.second(c)); }
    write_bin(job.out + ".lights.f32", lf.data(), lf.size());
  }
  std::string rmcmd = "rm -rf " + sdir;

  Tucano::Mesh& mesh = fs.mesh;
  const int F = mesh.getNumberOfFaces();
  const int V = mesh.getNumberOfVertices();

  if (job.dump_scene) {
    std::vector<float> verts(size_t(V) * 4), norms(size_t(V) * 3), fnorm(size_t(F) * 3);
    std::vector<int> faces(size_t(F) * 3), fmat(F);
    for (int i = 0; i < V; ++i) {
      for (int c = 0; c < 4; ++c) verts[i * 4 + c] = mesh.getVertex(i)(c);
      for (int c = 0; c < 3; ++c) norms[i * 3 + c] = mesh.getNormal(i)(c);
    }
    for (int i = 0; i < F; ++i) {
      Tucano::Face& f = mesh.getFace(i);
      for (int c = 0; c < 3; ++c) { faces[i * 3 + c] = f.vertex_ids[c]; fnorm[i * 3 + c] = f.normal(c); }
      fmat[i] = f.material_id;
    }
    std::vector<float> mats;
    for (auto& m : fs.materials) {
      Eigen::Vector3f a = m.getAmbient(), d = m.getDiffuse(), s = m.getSpecular();
      for (int c = 0; c < 3; ++c) mats.push_back(a(c));
      for (int c = 0; c < 3; ++c) mats.push_back(d(c));
      for (int c = 0; c < 3; ++c) mats.push_back(s(c));
      mats.push_back(m.getShininess());
      mats.push_back(m.getOpticalDensity());
      mats.push_back(m.getDissolveFactor());
    }
    Eigen::Affine3f M = mesh.getShapeModelMatrix();
    std::vector<float> shape(12);
    for (int r = 0; r < 3; ++r)
      for (int c = 0; c < 4; ++c) shape[r * 4 + c] = M.matrix()(r, c);
    std::vector<float> boxes;
    std::vector<int> box_off(1, 0), box_faces;
    Tucano::Face* base = &mesh.getFace(0);
    for (BoundingBox* b : BoundingBox::boxes) {
      for (int c = 0; c < 3; ++c) boxes.push_back(b->low(c));
      for (int c = 0; c < 3; ++c) boxes.push_back(b->high(c));
      for (Tucano::Face* f : b->faces) box_faces.push_back(int(f - base));
      box_off.push_back((int)box_faces.size());
    }
    Eigen::Vector3f eye = fs.flycamera.getCenter();
    Eigen::Affine3f vm = fs.flycamera.getViewMatrix();
    Eigen::Affine3f vinv = vm.inverse();
    std::vector<float> cam;  // eye(3) view(12) inv_view(12) viewport(4) persp_scale(1)
    for (int c = 0; c < 3; ++c) cam.push_back(eye(c));
    for (int r = 0; r < 3; ++r) for (int c = 0; c < 4; ++c) cam.push_back(vm.matrix()(r, c));
    for (int r = 0; r < 3; ++r) for (int c = 0; c < 4; ++c) cam.push_back(vinv.matrix()(r, c));
    Eigen::Vector4f vp = fs.flycamera.getViewport();
    for (int c = 0; c < 4; ++c) cam.push_back(vp(c));
    cam.push_back(fs.flycamera.getPerspectiveScale());
    write_bin(job.out + ".verts.f32", verts.data(), verts.size());
    write_bin(job.out + ".normals.f32", norms.data(), norms.size());
    write_bin(job.out + ".faces.i32", faces.data(), faces.size());
    write_bin(job.out + ".face_mat.i32", fmat.data(), fmat.size());
    write_bin(job.out + ".face_normal.f32", fnorm.data(), fnorm.size());
    write_bin(job.out + ".materials.f32", mats.data(), mats.size());
    write_bin(job.out + ".shape.f32", shape.data(), shape.size());
    write_bin(job.out + ".boxes.f32", boxes.data(), boxes.size());
    write_bin(job.out + ".box_off.i32", box_off.data(), box_off.size());
    write_bin(job.out + ".box_faces.i32", box_faces.data(), box_faces.size());
    write_bin(job.out + ".camera.f32", cam.data(), cam.size());
  }

  const int L = int(fs.lights.size() + fs.dirLights.size());

  if (job.mode == "anyhit" || job.mode == "closest") {
    // Arbitrary rays through the reference's shadow() (flyscene.cpp:510-526) or the
    // closest-hit loop.  ray_file: float32 x6 per ray.
    std::vector<float> rays = read_bin<float>(job.ray_file);
    size_t n = rays.size() / 6;
    int* hit = shared_alloc<int>(n);
    float* tt = shared_alloc<float>(n);
    auto t0 = std::chrono::steady_clock::now();
    std::vector<pid_t> kids;
    for (int p = 0; p < job.procs; ++p) {
      pid_t pid = job.procs > 1 ? fork() : 0;
      if (pid == 0) {
        for (size_t r = p; r < n; r += job.procs) {
          Eigen::Vector3f o(rays[r * 6], rays[r * 6 + 1], rays[r * 6 + 2]);
          Eigen::Vector3f d(rays[r * 6 + 3], rays[r * 6 + 4], rays[r * 6 + 5]);
          if (job.mode == "anyhit") {
            hit[r] = fs.shadow(o, d) ? 1 : 0;
            tt[r] = 0.f;
          } else {
            Eigen::Vector3f P;
            hit[r] = closest_face_index(fs, o, d, tt[r], P);
          }
        }
        if (job.procs > 1) _exit(0);
      } else kids.push_back(pid);
    }
    for (pid_t k : kids) { int st; waitpid(k, &st, 0); }
    double sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    write_bin(job.out + ".hit.i32", hit, n);
    write_bin(job.out + ".t.f32", tt, n);
    FILE* f = fopen((job.out + ".time.txt").c_str(), "w");
    fprintf(f, "%.9f %zu %d\n", sec, n, job.procs);
    fclose(f);
    if (system(rmcmd.c_str())) {}
    return 0;
  }

  // ---- render mode -------------------------------------------------------------------
  std::vector<int> px;  // (i,j) pairs
  if (!job.pixel_file.empty()) px = read_bin<int>(job.pixel_file);
  else
    for (int j = job.off_y; j < job.h; j += job.stride_y)
      for (int i = job.off_x; i < job.w; i += job.stride_x) { px.push_back(i); px.push_back(j); }
  const size_t n = px.size() / 2;

  float* rgb = shared_alloc<float>(n * 3);
  float* dirs = shared_alloc<float>(n * 3);
  float* tprim = shared_alloc<float>(n);
  int* hit = shared_alloc<int>(n);
  long long* counts = shared_alloc<long long>(size_t(job.procs) * 4);  // primary, shadow, reflection, hits
  if (job.repeat < 1) job.repeat = 1;
  double* tsec = shared_alloc<double>(size_t(job.procs) * job.repeat);

  auto t0 = std::chrono::steady_clock::now();
  std::vector<pid_t> kids;
  for (int p = 0; p < job.procs; ++p) {
    pid_t pid = job.procs > 1 ? fork() : 0;
    if (pid == 0) {
      // timed part: exactly the body of traceRayThread (flyscene.cpp:301-310)
      Eigen::Vector3f origin = fs.flycamera.getCenter();
      Eigen::Vector3f direction;
      for (int rep = 0; rep < job.repeat; ++rep) {
        auto c0 = std::chrono::steady_clock::now();
        // interleave the pixel list over processes (row bands would be unbalanced)
        for (size_t k = p; k < n; k += job.procs) {
          int i = px[2 * k], j = px[2 * k + 1];
          direction = (fs.flycamera.screenToWorld(Eigen::Vector2f(i, j)) - origin).normalized();
          Eigen::Vector3f c = fs.traceRay(origin, direction, 0, false);
          for (int ch = 0; ch < 3; ++ch) { rgb[k * 3 + ch] = c(ch); dirs[k * 3 + ch] = direction(ch); }
        }
        tsec[size_t(p) * job.repeat + rep] = std::chrono::duration<double>(std::chrono::steady_clock::now() - c0).count();
      }
      // untimed part: walk the hit chain for the primary face index and the ray counts
      long long np = 0, ns = 0, nr = 0, nh = 0;
      if (job.chain) {
        for (size_t k = p; k < n; k += job.procs) {
          Eigen::Vector3f o = origin;
          Eigen::Vector3f d(dirs[k * 3], dirs[k * 3 + 1], dirs[k * 3 + 2]);
          hit[k] = -1;
          tprim[k] = INFINITY;
          ++np;
          for (int depth = 0; depth < job.max_depth; ++depth) {
            float t;
            Eigen::Vector3f P;
            int f = closest_face_index(fs, o, d, t, P);
            if (depth == 0) { hit[k] = f; tprim[k] = t; }
            if (f < 0) break;
            ++nh;
            ns += L;
            if (depth + 1 >= job.max_depth) break;
            ++nr;
            // flyscene.cpp:361-362
            Eigen::Vector3f r = fs.reflect(d.normalized(), fs.interpolateNormal(mesh.getFace(f), P));
            Eigen::Vector3f o2 = P + (0.001 * r);
            o = o2;
            d = r;
          }
        }
      }
      counts[p * 4 + 0] = np; counts[p * 4 + 1] = ns; counts[p * 4 + 2] = nr; counts[p * 4 + 3] = nh;
      if (job.procs > 1) _exit(0);
    } else kids.push_back(pid);
  }
  for (pid_t k : kids) { int st; waitpid(k, &st, 0); }
  double wall = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
  double tmax = 0;
  long long tot[4] = {0, 0, 0, 0};
  std::vector<double> trep(job.repeat, 0.0);  // per pass: max over processes
  for (int p = 0; p < job.procs; ++p) {
    for (int rep = 0; rep < job.repeat; ++rep)
      trep[rep] = std::max(trep[rep], tsec[size_t(p) * job.repeat + rep]);
    for (int c = 0; c < 4; ++c) tot[c] += counts[p * 4 + c];
  }
  tmax = trep[job.repeat - 1];
  write_bin(job.out + ".rgb.f32", rgb, n * 3);
  write_bin(job.out + ".dir.f32", dirs, n * 3);
  write_bin(job.out + ".pixels.i32", px.data(), px.size());
  if (job.chain) {
    write_bin(job.out + ".hit.i32", hit, n);
    write_bin(job.out + ".t.f32", tprim, n);
  }
  FILE* f = fopen((job.out + ".time.txt").c_str(), "w");
  // trace seconds (max over processes), wall seconds incl. chain walk, pixels, procs, rays p/s/r, hits
  fprintf(f, "%.9f %.9f %zu %d %lld %lld %lld %lld", tmax, wall, n, job.procs, tot[0], tot[1], tot[2], tot[3]);
  for (double t : trep) fprintf(f, " %.9f", t);
  fprintf(f, "\n");
  fclose(f);
  if (system(rmcmd.c_str())) {}
  return 0;
}
