// rtp_api.cu - the C ABI of include/rtp.h: scene upload, frame buffers, kernel sequencing.
// There is no CPU fallback here: every compute entry point needs a CUDA device.
#include <cuda_runtime.h>

#include <atomic>
#include <cmath>
#include <type_traits>
#include <chrono>
#include <condition_variable>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <memory>
#include <mutex>
#include <new>
#include <string>
#include <thread>
#include <vector>

#include "../../include/rtp.h"
#include "bvh_build_gpu.h"
#include "rtp_kernels.h"
#include "scene_host.h"

using namespace rtp;

namespace {

thread_local std::string g_err;

int fail(int code, const std::string &msg) {
  g_err = msg;
  return code;
}

#define CK(call)                                                                              \
  do {                                                                                        \
    cudaError_t e_ = (call);                                                                  \
    if (e_ != cudaSuccess)                                                                    \
      return fail(RTP_ERR_CUDA, std::string(#call) + ": " + cudaGetErrorString(e_));           \
  } while (0)

template <class T>
struct DevBuf {
  T *p = nullptr;
  size_t n = 0;
  ~DevBuf() { release(); }
  void release() {
    if (p) cudaFree(p);
    p = nullptr;
    n = 0;
  }
  cudaError_t ensure(size_t count) {
    if (count <= n && p) return cudaSuccess;
    release();
    cudaError_t e = cudaMalloc((void **)&p, std::max<size_t>(count, 1) * sizeof(T));
    if (e == cudaSuccess) n = count;
    return e;
  }
  template <class A>
  cudaError_t upload(const std::vector<T, A> &v, cudaStream_t st) {
    cudaError_t e = ensure(v.size());
    if (e != cudaSuccess) return e;
    if (v.empty()) return cudaSuccess;
    return cudaMemcpyAsync(p, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice, st);
  }
};

constexpr int kMaxGroups = 8;

struct Frame {
  int width = 0, height = 0, rank = 0, world = 1, max_depth = 0;
  int n_px = 0, tiles_x = 0;
  DevBuf<float4> fb, ksdeep, direct, ray_o0, ray_o1, ray_d0, ray_d1;
  int n_lights = 0, n_valid = 0;
  DevBuf<float4> pbuf, dbuf;
  DevBuf<int> primface, hit_slot, hitq, tile_ids, local_of_tile;
  DevBuf<unsigned char> occl;
  DevBuf<float> out_rgb;
  DevBuf<unsigned char> out_rgb8;
  DevBuf<int> out_hit;
  std::vector<int32_t> tiles_host;  // ids of the owned tiles (host copy of tile_ids)
};

// One host thread per extra GPU of a multi-device scene: the frame of device r is enqueued by worker r while the
// caller's thread enqueues device 0's (a frame is a dozen launches; issued from one thread for 8 devices they would
// start up to ~0.3 ms apart).  A worker spins for a short while after a job - a render loop finds it awake - then sleeps.
class Worker {
 public:
  Worker() : th_([this] { run(); }) {}
  ~Worker() {
    { std::lock_guard<std::mutex> g(m_); quit_ = true; state_.store(1, std::memory_order_release); }
    cv_.notify_one();
    th_.join();
  }
  void submit(std::function<int()> fn) {
    { std::lock_guard<std::mutex> g(m_); job_ = std::move(fn); state_.store(1, std::memory_order_release); }
    cv_.notify_one();
  }
  int wait() {  // result of the submitted job
    while (state_.load(std::memory_order_acquire) != 2) {
#if defined(__x86_64__)
      __builtin_ia32_pause();
#endif
    }
    state_.store(0, std::memory_order_release);
    return rc_;
  }

 private:
  void run() {
    for (;;) {
      const auto t0 = std::chrono::steady_clock::now();
      while (state_.load(std::memory_order_acquire) != 1) {  // spin, then sleep
        if (std::chrono::steady_clock::now() - t0 > std::chrono::microseconds(300)) {
          std::unique_lock<std::mutex> lk(m_);
          cv_.wait(lk, [this] { return state_.load(std::memory_order_acquire) == 1; });
          break;
        }
      }
      std::function<int()> fn;
      {
        std::lock_guard<std::mutex> g(m_);
        if (quit_) return;
        fn.swap(job_);
      }
      rc_ = fn ? fn() : 0;
      state_.store(2, std::memory_order_release);
    }
  }
  std::mutex m_;
  std::condition_variable cv_;
  std::function<int()> job_;
  std::atomic<int> state_{0};  // 0 idle, 1 job posted, 2 job done
  bool quit_ = false;
  int rc_ = 0;
  std::thread th_;  // last member: the thread starts when everything above exists
};

}  // namespace

struct rtp_scene {
  int device = 0;
  HostScene host;
  DevBuf<float4> nodes, tris, vnorm, boxes, materials, lights;
  DevBuf<int> face_of;
  DevBuf<int> counters;
  DevBuf<int> park;  // parked-ray list of the trace kernels
  int *barrier_timeouts = nullptr;      // [1] waits of rtp_peer_barrier that gave up: page-locked, mapped - the kernel
  int *barrier_timeouts_dev = nullptr;  // writes it through the device alias, the host reads it without a device sync
  int paths_from = 1;      // first bounce level k_paths takes over (levels below it run as per-level launches)
  int deep_mode = -1;      // bounce levels >= 1: 1 = k_paths, 0 = per-level kernels, -1 = by frame size (RTP_DEEP, at scene creation)
  int paths_px_big = 2600000, paths_px_small = 1100000;  // -1: k_paths up to this many pixels per rank (scene larger / not larger than the L2)
  float pad_cur = 0.f;  // padding of the hierarchy's boxes right now (build-time pad, grown by ensure_pad)
  // multi-device scene (rtp_scene_create_multi): this is device 0's scene, peers[r - 1] the replica of rank r
  std::vector<rtp_scene *> peers;
  std::vector<std::unique_ptr<Worker>> workers;  // one per peer
  rtp_scene *primary = nullptr;                  // set in a replica
  cudaEvent_t frame_start = nullptr, frame_done = nullptr;  // cross-device fork / join of a multi-device frame
  std::string err;                               // error text of a job that ran on a worker thread
  int *staging = nullptr;                        // page-locked host staging (world > 1 renders, ray batches)
  size_t staging_bytes = 0;
  DeviceScene dS;
  std::vector<float4> lights_host;  // what s->lights holds
  DeviceCamera cam;
  bool have_camera = false, have_lights = false;
  LaunchCfg cfg;
  cudaStream_t stream = nullptr;
  cudaEvent_t ev[4] = {nullptr, nullptr, nullptr, nullptr};
  // tile groups of a frame run on their own streams (enqueue_frame): group 0 on the caller's stream
  cudaStream_t gstream[kMaxGroups] = {};
  cudaEvent_t gfork = nullptr, gjoin[kMaxGroups] = {};
  // rtp_render_async: two frames may be outstanding; image planes, counters snapshot and events per slot,
  // band copies on their own streams so that the next frame's kernels do not queue behind them
  struct AsyncSlot {
    DevBuf<unsigned char> rgb8;
    DevBuf<int> hit;
    int *counters_host = nullptr;  // pinned, kCtrInts * kMaxGroups
    cudaEvent_t ev0 = nullptr, ev1 = nullptr, copied[kMaxGroups] = {};
    int groups = 1, launches = 0, primary = 0, depth = 0, lights = 0;
    bool busy = false;
  } aslot[2];
  cudaStream_t cstream[kMaxGroups] = {};
  cudaEvent_t resolved[kMaxGroups] = {};
  int *counters_pinned = nullptr;           // page-locked landing place of a frame's counters (rtp_render)
  unsigned async_head = 0, async_tail = 0;  // frames enqueued / waited for
  bool in_async_call = false;
  // RTP_L2_PERSIST=<MiB> (experiment, profiles/r02_design_items.md): that much of the L2 is set aside for persisting
  // lines and every stream a frame runs on gets an access-policy window over the head of the node array
  // The camera level in one kernel (k_camera) or as separate closest / any-hit / shading launches: RTP_FUSED=1 / 0 forces
  // one of them (A/B runs, parity suite), -1 = by size (enqueue_frame)
  int fused = -1;
  int fused_px = 1300000;  // RTP_FUSED_PX: scenes that spill out of the L2 fuse up to this many pixel x light pairs per rank
  int l2_persist_mb = 0;
  std::vector<cudaStream_t> l2_streams;  // streams that carry the window already
  int stagger = 0;      // RTP_STAGGER=1: every tile group on its own prioritised stream (band 0 first)
  int groups = 0;       // rtp_set_tile_groups / RTP_GROUPS; 0 = automatic (auto_groups)
  int last_groups = 1;  // groups of the most recently enqueued frame (their counters are summed)
  std::vector<cudaEvent_t> kev;  // per-kernel events (RTP_RENDER_PROFILE)
  std::vector<int> kev_kind;     // 0 closest, 1 anyhit, 2 streaming (raygen/prep/accum/resolve)
  int kev_used = 0;
  // RTP_TIMELINE=1 (profiling runs): an event after every launch on the launching stream; rtp_frame_stats prints
  // each kernel's end time relative to the start of the frame (which kernels of different groups overlapped)
  int timeline = 0;
  std::vector<cudaEvent_t> tev;
  std::vector<std::string> tev_name;
  int tev_used = 0;
  int last_primary = 0, last_depth = 0, last_lights = 0;
  cudaStream_t last_stream = nullptr;  // stream of the most recently enqueued frame
  bool has_last = false;
  int last_launches = 0;
  Frame frame;   // pixel frames (camera rays)
  Frame rframe;  // ray-batch frames (rtp_trace_rays SHADE)
  DevBuf<float> rays_dev;
  DevBuf<int> rays_face;
  DevBuf<float> rays_t;
  DevBuf<unsigned char> rays_occ;
};

namespace {

struct DeviceGuard {
  int prev = -1;
  bool ok = true;
  explicit DeviceGuard(int dev) {
    if (cudaGetDevice(&prev) != cudaSuccess) { ok = false; return; }
    if (prev != dev && cudaSetDevice(dev) != cudaSuccess) ok = false;
  }
  ~DeviceGuard() {
    if (prev >= 0) cudaSetDevice(prev);
  }
};

}  // namespace

extern "C" {

const char *rtp_last_error(void) { return g_err.c_str(); }
int rtp_abi_version(void) { return RTP_ABI_VERSION; }

}  // extern "C"

namespace {

// streams, events, tuning knobs of one device's scene (the device is current)
int init_device_side(rtp_scene *s) {
  CK(cudaStreamCreateWithFlags(&s->stream, cudaStreamNonBlocking));
  for (auto &ev : s->ev) CK(cudaEventCreate(&ev));
  {
    // Group g's stream has priority greatest + g (group 0 highest): when an SM has room, the pending CTAs of an earlier
    // band go first, so band 0 is through its three kernels while the later bands still trace and its rows are on
    // their way to the host under their kernels (RTP_STAGGER, enqueue_frame)
    int least = 0, greatest = 0;
    CK(cudaDeviceGetStreamPriorityRange(&least, &greatest));
    if (const char *se = std::getenv("RTP_STAGGER")) s->stagger = std::atoi(se);
    for (int i = s->stagger ? 0 : 1; i < kMaxGroups; ++i)
      CK(cudaStreamCreateWithPriority(&s->gstream[i], cudaStreamNonBlocking, s->stagger ? std::min(least, greatest + i) : least));
    if (s->stagger) CK(cudaEventCreateWithFlags(&s->gjoin[0], cudaEventDisableTiming));
  }
  CK(cudaEventCreateWithFlags(&s->gfork, cudaEventDisableTiming));
  for (int i = 1; i < kMaxGroups; ++i) CK(cudaEventCreateWithFlags(&s->gjoin[i], cudaEventDisableTiming));
  CK(cudaEventCreateWithFlags(&s->frame_start, cudaEventDisableTiming));
  CK(cudaEventCreateWithFlags(&s->frame_done, cudaEventDisableTiming));
  if (const char *te = std::getenv("RTP_TIMELINE")) s->timeline = std::atoi(te);
  if (const char *de = std::getenv("RTP_DEEP")) s->deep_mode = std::string(de) == "wavefront" ? 0 : std::string(de) == "paths" ? 1 : -1;  // per scene
  if (const char *te = std::getenv("RTP_PATHS_PX_BIG")) s->paths_px_big = std::atoi(te);
  if (const char *te = std::getenv("RTP_PATHS_PX_SMALL")) s->paths_px_small = std::atoi(te);
  if (const char *pe = std::getenv("RTP_PATHS_FROM")) s->paths_from = std::atoi(pe);
  if (const char *ge = std::getenv("RTP_GROUPS")) s->groups = std::min(std::max(std::atoi(ge), 0), kMaxGroups);
  if (const char *fe = std::getenv("RTP_FUSED")) s->fused = std::atoi(fe);
  if (const char *fe = std::getenv("RTP_FUSED_PX")) s->fused_px = std::atoi(fe);
  if (const char *le = std::getenv("RTP_L2_PERSIST")) {
    s->l2_persist_mb = std::max(std::atoi(le), 0);
    if (s->l2_persist_mb) CK(cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, (size_t)s->l2_persist_mb << 20));
  }
  return RTP_OK;
}

// RTP_L2_PERSIST: kernels launched into `st` keep the lines of the node array's head in the persisting part of the L2
int apply_l2_window(rtp_scene *s, cudaStream_t st) {
  if (s->l2_persist_mb <= 0 || std::find(s->l2_streams.begin(), s->l2_streams.end(), st) != s->l2_streams.end()) return RTP_OK;
  int max_window = 0;
  CK(cudaDeviceGetAttribute(&max_window, cudaDevAttrMaxAccessPolicyWindowSize, s->device));
  const size_t want = (size_t)s->host.n_nodes * 64, carve = (size_t)s->l2_persist_mb << 20;
  const size_t bytes = std::min(want, (size_t)std::max(max_window, 0));
  if (!bytes) return RTP_OK;
  cudaStreamAttrValue v;
  std::memset(&v, 0, sizeof v);
  v.accessPolicyWindow.base_ptr = (void *)s->nodes.p;
  v.accessPolicyWindow.num_bytes = bytes;
  v.accessPolicyWindow.hitRatio = (float)std::min(1.0, (double)carve / (double)bytes);
  v.accessPolicyWindow.hitProp = cudaAccessPropertyPersisting;
  v.accessPolicyWindow.missProp = cudaAccessPropertyStreaming;
  CK(cudaStreamSetAttribute(st, cudaStreamAttributeAccessPolicyWindow, &v));
  s->l2_streams.push_back(st);
  return RTP_OK;
}

// the flattened scene of h (host arrays) -> device arrays of s; with gpu_build the hierarchy is built there
int upload_scene(rtp_scene *s, HostScene &h, bool gpu_build) {
  if (gpu_build) {
    // rank-ordered triangles go up once; the device sorts them, builds the hierarchy and permutes
    DevBuf<float4> tris_in, vnorm_in;
    DevBuf<int> face_in;
    const size_t n = (size_t)h.n_tris;
    CK(tris_in.upload(h.tris, s->stream));
    CK(vnorm_in.upload(h.vnorm, s->stream));
    CK(face_in.upload(h.face_of, s->stream));
    CK(s->tris.ensure(4 * n));
    CK(s->vnorm.ensure(3 * n));
    CK(s->face_of.ensure(n));
    CK(s->nodes.ensure(4 * n));
    int n_nodes = 0, height = 0;
    CK(build_lbvh_device(tris_in.p, vnorm_in.p, face_in.p, h.n_tris, h.pad, h.scene_lo, h.scene_hi, 4, s->tris.p, s->vnorm.p,
                         s->face_of.p, s->nodes.p, &n_nodes, &height, s->stream));
    if (height >= kStackSize - 2)
      return fail(RTP_ERR_INVALID, "device-built BVH is deeper than the traversal stack; use RTP_BUILD_HOST_SAH");
    h.n_nodes = n_nodes;
    h.bvh_depth = height;
  } else {
    CK(s->nodes.upload(h.nodes, s->stream));
    CK(s->tris.upload(h.tris, s->stream));
    CK(s->vnorm.upload(h.vnorm, s->stream));
    CK(s->face_of.upload(h.face_of, s->stream));
  }
  CK(s->boxes.upload(h.boxes, s->stream));
  CK(s->materials.upload(h.materials, s->stream));
  CK(cudaStreamSynchronize(s->stream));
  return RTP_OK;
}

// device arrays of src (another device, or the same one) -> device arrays of dst: the hierarchy is built once
int replicate_scene(rtp_scene *dst, const rtp_scene *src) {
  auto copy = [&](auto &d, const auto &sbuf) -> cudaError_t {
    cudaError_t e = d.ensure(sbuf.n);
    if (e != cudaSuccess || sbuf.n == 0) return e;
    using T = typename std::remove_reference<decltype(*d.p)>::type;
    return cudaMemcpyPeerAsync(d.p, dst->device, sbuf.p, src->device, sbuf.n * sizeof(T), dst->stream);
  };
  CK(copy(dst->nodes, src->nodes));
  CK(copy(dst->tris, src->tris));
  CK(copy(dst->vnorm, src->vnorm));
  CK(copy(dst->face_of, src->face_of));
  CK(copy(dst->boxes, src->boxes));
  CK(copy(dst->materials, src->materials));
  CK(cudaStreamSynchronize(dst->stream));
  return RTP_OK;
}

// counters, parked-ray list, DeviceScene view, launch configuration (h: scalar description of the scene)
int finish_scene(rtp_scene *s, const HostScene &h) {
  CK(s->counters.ensure(kCtrInts * kMaxGroups));
  CK(cudaMemsetAsync(s->counters.p, 0, kCtrInts * kMaxGroups * sizeof(int), s->stream));
  DeviceScene &d = s->dS;
  d.nodes = s->nodes.p; d.tris = s->tris.p; d.vnorm = s->vnorm.p; d.boxes = s->boxes.p;
  d.face_of = s->face_of.p; d.materials = s->materials.p; d.lights = nullptr; d.n_lights = 0;
  std::memcpy(d.minv, h.minv, sizeof d.minv);
  d.n_tris = h.n_tris; d.n_nodes = h.n_nodes; d.n_boxes = h.n_boxes;
  s->pad_cur = h.pad;
  query_launch_cfg(s->device, s->cfg);
  constexpr int kParkCap = 1 << 17;  // 131072 parked rays (36 MiB); a full list only means no more parking
  CK(s->park.ensure((size_t)kParkCap * kParkSlotBytes / sizeof(int)));
  s->cfg.park_slots = s->park.p;
  s->cfg.park_cap = kParkCap;
  int l2 = 0;
  cudaDeviceGetAttribute(&l2, cudaDevAttrL2CacheSize, s->device);
  const size_t scene_bytes = ((size_t)h.n_nodes + (size_t)h.n_tris) * 64;
  s->cfg.scene_exceeds_l2 = scene_bytes > (size_t)(l2 > 0 ? l2 : (64 << 20)) / 2;
  CK(cudaStreamSynchronize(s->stream));
  CK(cudaGetLastError());
  return RTP_OK;
}

void destroy_one(rtp_scene *s) {
  DeviceGuard g(s->device);
  if (s->stream) { cudaStreamSynchronize(s->stream); cudaStreamDestroy(s->stream); }
  for (auto &ev : s->ev)
    if (ev) cudaEventDestroy(ev);
  for (auto &ev : s->kev) cudaEventDestroy(ev);
  for (auto &ev : s->tev) cudaEventDestroy(ev);
  for (int i = 0; i < kMaxGroups; ++i) {
    if (s->gstream[i]) { cudaStreamSynchronize(s->gstream[i]); cudaStreamDestroy(s->gstream[i]); }
    if (s->gjoin[i]) cudaEventDestroy(s->gjoin[i]);
  }
  if (s->gfork) cudaEventDestroy(s->gfork);
  if (s->frame_start) cudaEventDestroy(s->frame_start);
  if (s->frame_done) cudaEventDestroy(s->frame_done);
  for (int i = 0; i < kMaxGroups; ++i) {
    if (s->cstream[i]) { cudaStreamSynchronize(s->cstream[i]); cudaStreamDestroy(s->cstream[i]); }
    if (s->resolved[i]) cudaEventDestroy(s->resolved[i]);
  }
  if (s->counters_pinned) cudaFreeHost(s->counters_pinned);
  if (s->barrier_timeouts) cudaFreeHost(s->barrier_timeouts);
  if (s->staging) cudaFreeHost(s->staging);
  for (auto &a : s->aslot) {
    if (a.counters_host) cudaFreeHost(a.counters_host);
    if (a.ev0) cudaEventDestroy(a.ev0);
    if (a.ev1) cudaEventDestroy(a.ev1);
    for (auto &e : a.copied)
      if (e) cudaEventDestroy(e);
  }
  delete s;  // DevBuf destructors free device memory while the device is current
}

void destroy_set(rtp_scene *s) {
  s->workers.clear();  // joins the worker threads
  for (rtp_scene *p : s->peers) destroy_one(p);
  s->peers.clear();
  destroy_one(s);
}

// The scalar part of a HostScene (no arrays): what a replica needs to describe itself.
void copy_meta(HostScene &dst, const HostScene &src) {
  std::memcpy(dst.minv, src.minv, sizeof dst.minv);
  dst.n_tris = src.n_tris; dst.n_nodes = src.n_nodes; dst.n_boxes = src.n_boxes; dst.bvh_depth = src.bvh_depth;
  dst.n_flagged = src.n_flagged; dst.n_degenerate = src.n_degenerate; dst.pad = src.pad; dst.maxabs = src.maxabs;
  std::memcpy(dst.scene_lo, src.scene_lo, sizeof dst.scene_lo);
  std::memcpy(dst.scene_hi, src.scene_hi, sizeof dst.scene_hi);
}

}  // namespace

extern "C" {

int rtp_scene_create_multi(const rtp_mesh *mesh, const int *devices, int n_devices, rtp_scene **out) {
  if (!mesh || !out || !devices || n_devices < 1 || n_devices > RTP_MAX_DEVICES)
    return fail(RTP_ERR_INVALID, "rtp_scene_create_multi: null argument or n_devices outside [1, RTP_MAX_DEVICES]");
  *out = nullptr;
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0)
    return fail(RTP_ERR_CUDA, std::string("rtp_scene_create: no CUDA device (") +
                                  (e != cudaSuccess ? cudaGetErrorString(e) : "device count 0") +
                                  "); this library has no CPU fallback");
  std::vector<int> dev(devices, devices + n_devices);
  for (int &d : dev) {
    if (d < 0) CK(cudaGetDevice(&d));
    if (d >= ndev) return fail(RTP_ERR_INVALID, "rtp_scene_create: device ordinal out of range");
  }
  rtp_scene *s = new (std::nothrow) rtp_scene();
  if (!s) return fail(RTP_ERR_NOMEM, "rtp_scene_create: out of host memory");
  s->device = dev[0];
  // where the traversal hierarchy is built: rtp_mesh.build, overridden by RTP_BVH_BUILD=gpu|host
  bool gpu_build = mesh->build == RTP_BUILD_GPU_LBVH;
  if (const char *be = std::getenv("RTP_BVH_BUILD")) gpu_build = std::string(be) == "gpu";
  if (mesh->n_faces < 64) gpu_build = false;  // tiny scenes: nothing to gain
  std::string msg;
  try {
    msg = flatten_scene(*mesh, s->host, 0, !gpu_build);  // ONCE, whatever the number of devices
  } catch (const std::bad_alloc &) {
    delete s;
    return fail(RTP_ERR_NOMEM, "rtp_scene_create: out of host memory while flattening");
  }
  if (!msg.empty()) {
    delete s;
    return fail(RTP_ERR_INVALID, msg);
  }
  HostScene &h = s->host;
#ifdef RTP_SMEM_STACK
  if (!gpu_build && h.bvh_depth >= RTP_SMEM_STACK - 2) {
    delete s;
    return fail(RTP_ERR_INVALID, "RTP_SMEM_STACK experiment build: the hierarchy is deeper than the shared-memory stack");
  }
#endif
  int rc = RTP_OK;
  {
    DeviceGuard g(dev[0]);
    if (!g.ok) { delete s; return fail(RTP_ERR_CUDA, "rtp_scene_create: cudaSetDevice failed"); }
    rc = init_device_side(s);
    if (!rc) rc = upload_scene(s, h, gpu_build);
    if (!rc) rc = finish_scene(s, h);
  }
  // the big flattened arrays now live on the device; keep only the BoundingBox view on the host
  BigVec<float4>().swap(h.nodes);
  BigVec<float4>().swap(h.tris);
  BigVec<float4>().swap(h.vnorm);
  for (int r = 1; r < n_devices && !rc; ++r) {
    rtp_scene *p = new (std::nothrow) rtp_scene();
    if (!p) { rc = fail(RTP_ERR_NOMEM, "rtp_scene_create: out of host memory"); break; }
    p->device = dev[(size_t)r];
    p->primary = s;
    copy_meta(p->host, h);
    s->peers.push_back(p);
    DeviceGuard g(p->device);
    if (!g.ok) { rc = fail(RTP_ERR_CUDA, "rtp_scene_create: cudaSetDevice failed"); break; }
    if (p->device != s->device) {  // this device's kernels store their pixels into device 0's image
      int can = 0;
      cudaDeviceCanAccessPeer(&can, p->device, s->device);
      if (!can) { rc = fail(RTP_ERR_CUDA, "rtp_scene_create_multi: device " + std::to_string(p->device) + " has no peer access to device " + std::to_string(s->device)); break; }
      cudaError_t pe = cudaDeviceEnablePeerAccess(s->device, 0);
      if (pe == cudaErrorPeerAccessAlreadyEnabled) cudaGetLastError();
      else if (pe != cudaSuccess) { rc = fail(RTP_ERR_CUDA, std::string("cudaDeviceEnablePeerAccess: ") + cudaGetErrorString(pe)); break; }
    }
    rc = init_device_side(p);
    if (!rc) rc = replicate_scene(p, s);
    if (!rc) rc = finish_scene(p, h);
  }
  if (!rc)
    for (size_t r = 0; r < s->peers.size(); ++r) s->workers.emplace_back(new Worker());
  if (rc) {
    const std::string keep = g_err;
    destroy_set(s);
    g_err = keep;
    return rc;
  }
  *out = s;
  return RTP_OK;
}

int rtp_scene_create(const rtp_mesh *mesh, int device, rtp_scene **out) {
  if (!mesh || !out) return fail(RTP_ERR_INVALID, "rtp_scene_create: null argument");
  return rtp_scene_create_multi(mesh, &device, 1, out);
}

int rtp_scene_device_count(const rtp_scene *s) { return s ? 1 + (int)s->peers.size() : 0; }

int rtp_device_count(void) {
  int n = 0;
  CK(cudaGetDeviceCount(&n));
  return n;
}

int rtp_scene_destroy(rtp_scene *s) {
  if (!s) return RTP_OK;
  if (s->primary) return fail(RTP_ERR_INVALID, "rtp_scene_destroy: this handle is a replica owned by a multi-device scene");
  destroy_set(s);
  return RTP_OK;
}

}  // extern "C"

namespace {

int set_lights_one(rtp_scene *s, const rtp_lights *l, const std::vector<float4> &packed) {
  DeviceGuard g(s->device);
  // unchanged light set (a render loop sets it before every frame): nothing to upload.  Otherwise the copy is
  // stream-ordered behind the previous frame (all its tile groups are joined into s->stream) and does not block
  // the host: frames of rtp_render_async stay in flight.  A re-allocation (more lights than ever before) waits.
  const bool same = s->have_lights && packed.size() == s->lights_host.size() &&
                    (packed.empty() || std::memcmp(packed.data(), s->lights_host.data(), packed.size() * sizeof(float4)) == 0);
  if (!same) {
    if (packed.size() > s->lights.n && s->has_last) CK(cudaStreamSynchronize(s->last_stream));
    s->lights_host = packed;  // the source of the asynchronous copy stays alive in the scene
    CK(s->lights.upload(s->lights_host, s->has_last ? s->last_stream : s->stream));
  }
  s->dS.lights = s->lights.p;
  s->dS.n_lights = l->n_point + l->n_dir;
  s->have_lights = true;
  return RTP_OK;
}

// The hierarchy's boxes are padded for the rounding of P = o + t*d, which grows with |o| (ulp(|o| + t)): a ray that
// starts `dist` away from the origin needs a pad of ~8 ulp(dist).  The build-time pad (3e-5 x scene size) covers
// origins up to ~30 scene sizes away; beyond that the boxes are widened in place, once, to what this origin needs
// (only ever grows: a wider pad is always correct, it only culls a little less).
int ensure_pad(rtp_scene *s, float dist) {
  const float need = 1.0e-6f * dist;
  if (!(need > s->pad_cur)) return RTP_OK;
  DeviceGuard g(s->device);
  cudaStream_t st = s->has_last ? s->last_stream : s->stream;
  launch_inflate_nodes(s->nodes.p, s->host.n_nodes, need - s->pad_cur, st);
  CK(cudaGetLastError());
  CK(cudaStreamSynchronize(st));  // a rare, one-off event: the next frame may run on any stream
  s->pad_cur = need;
  return RTP_OK;
}

}  // namespace

extern "C" {

int rtp_set_lights(rtp_scene *s, const rtp_lights *l) {
  if (!s || !l) return fail(RTP_ERR_INVALID, "rtp_set_lights: null argument");
  if (l->n_point < 0 || l->n_dir < 0 || (l->n_point && !l->point) || (l->n_dir && !l->dir))
    return fail(RTP_ERR_INVALID, "rtp_set_lights: bad light arrays");
  std::vector<float4> packed;
  pack_lights(*l, packed);
  int rc = set_lights_one(s, l, packed);
  for (rtp_scene *p : s->peers)
    if (!rc) rc = set_lights_one(p, l, packed);
  return rc;
}

int rtp_set_tile_groups(rtp_scene *s, int groups) {
  if (!s || groups < 0 || groups > kMaxGroups) return fail(RTP_ERR_INVALID, "rtp_set_tile_groups: groups must be in [0, 8]");
  s->groups = groups;
  for (rtp_scene *p : s->peers) p->groups = groups;
  return RTP_OK;
}

int rtp_get_tile_groups(const rtp_scene *s) { return s ? s->last_groups : 0; }

int rtp_set_camera(rtp_scene *s, const rtp_camera *c) {
  if (!s || !c) return fail(RTP_ERR_INVALID, "rtp_set_camera: null argument");
  if (!(c->viewport[2] > 0.f) || !(c->viewport[3] > 0.f)) return fail(RTP_ERR_INVALID, "rtp_set_camera: empty viewport");
  derive_camera(*c, s->cam);
  s->have_camera = true;
  const float dist = std::max(std::fabs(s->cam.eye[0]), std::max(std::fabs(s->cam.eye[1]), std::fabs(s->cam.eye[2])));
  int rc = std::isfinite(dist) ? ensure_pad(s, dist) : RTP_OK;
  for (rtp_scene *p : s->peers) {
    p->cam = s->cam;
    p->have_camera = true;
    if (!rc && std::isfinite(dist)) rc = ensure_pad(p, dist);
  }
  return rc;
}

int rtp_scene_box_count(const rtp_scene *s) { return s ? s->host.n_boxes : 0; }

int rtp_scene_get_boxes(const rtp_scene *s, float *boxes, int32_t *box_off, int32_t *box_faces) {
  if (!s) return fail(RTP_ERR_INVALID, "rtp_scene_get_boxes: null scene");
  const HostScene &h = s->host;
  if (boxes) std::memcpy(boxes, h.box_bounds.data(), h.box_bounds.size() * sizeof(float));
  if (box_off) std::memcpy(box_off, h.box_off.data(), h.box_off.size() * sizeof(int32_t));
  if (box_faces) std::memcpy(box_faces, h.box_faces.data(), h.box_faces.size() * sizeof(int32_t));
  return RTP_OK;
}

int rtp_scene_info(const rtp_scene *s, rtp_stats *st) {
  if (!s || !st) return fail(RTP_ERR_INVALID, "rtp_scene_info: null argument");
  std::memset(st, 0, sizeof *st);
  st->n_bvh_nodes = s->host.n_nodes;
  st->n_triangles = s->host.n_tris;
  return RTP_OK;
}

}  // extern "C"

// ---- frames -------------------------------------------------------------------------------
namespace {

int own_tile_count(int width, int height, int rank, int world) {
  const int tx = (width + kTileDim - 1) / kTileDim, ty = (height + kTileDim - 1) / kTileDim;
  int n = 0;
  for (int y = 0; y < ty; ++y)
    for (int x = 0; x < tx; ++x) n += tile_owner(x, y, world) == rank;
  return n;
}

int max_tile_count(int width, int height, int world) {
  int m = 0;
  for (int r = 0; r < world; ++r) m = std::max(m, own_tile_count(width, height, r, world));
  return m;
}

// (Re)allocates the per-frame device state.  n_paths: pixels (tile-padded) or user rays.
int ensure_frame(rtp_scene *s, Frame &f, int width, int height, int rank, int world, int max_depth, int n_paths,
                 bool with_tiles) {
  const int n_lights = std::max(s->dS.n_lights, 1);
  const bool same = f.width == width && f.height == height && f.rank == rank && f.world == world &&
                    f.max_depth >= max_depth && f.n_px == n_paths && f.n_lights >= n_lights;
  if (same) return RTP_OK;
  // shadow rays are indexed (hit, light) in 32 bits (FrameAny::count, cursors): refuse what would overflow
  if ((long long)n_paths * n_lights > 0x7fffffffll)
    return fail(RTP_ERR_INVALID, "render: paths x lights exceeds 2^31 - 1 shadow rays per level (" + std::to_string(n_paths) +
                                     " x " + std::to_string(n_lights) + "): render fewer pixels per call or use fewer lights");
  // the buffers below are about to be re-sized / re-filled: wait for the frame that may still use them
  if (s->has_last) CK(cudaStreamSynchronize(s->last_stream));
  f.width = width; f.height = height; f.rank = rank; f.world = world; f.max_depth = max_depth; f.n_px = n_paths;
  f.n_lights = n_lights;
  f.tiles_x = (width + kTileDim - 1) / kTileDim;
  const size_t n = (size_t)std::max(n_paths, 1);
  CK(f.ksdeep.ensure(n));
  CK(f.direct.ensure(n * (size_t)max_depth));
  CK(f.ray_o0.ensure(n)); CK(f.ray_o1.ensure(n)); CK(f.ray_d0.ensure(n)); CK(f.ray_d1.ensure(n));
  CK(f.primface.ensure(n)); CK(f.hit_slot.ensure(n));
  CK(f.pbuf.ensure(n)); CK(f.dbuf.ensure(n)); CK(f.hitq.ensure(n));
  CK(f.occl.ensure(n * (size_t)n_lights));
  f.n_valid = n_paths;
  if (with_tiles) {
    std::vector<int32_t> tiles, local;
    tile_tables(width, height, rank, world, tiles, local);
    long long valid = 0;  // owned pixels inside the image = primary rays of this rank
    for (int32_t t : tiles) {
      const int x0 = (t % f.tiles_x) * kTileDim, y0 = (t / f.tiles_x) * kTileDim;
      valid += (long long)std::min(kTileDim, width - x0) * std::min(kTileDim, height - y0);
    }
    f.n_valid = (int)valid;
    f.tiles_host = tiles;
    CK(f.tile_ids.upload(tiles, s->stream));
    CK(f.local_of_tile.upload(local, s->stream));
    CK(cudaStreamSynchronize(s->stream));
  }
  return RTP_OK;
}

void fill_params(rtp_scene *s, Frame &f, const rtp_render_opts &o, float4 *fb, FrameParams &p) {
  p.S = s->dS;
  p.cam = s->cam;
  p.width = f.width; p.height = f.height; p.tiles_x = f.tiles_x; p.n_px = f.n_px;
  p.tile_ids = f.tile_ids.p;
  p.user_rays = nullptr; p.n_user_rays = 0;
  p.max_depth = o.max_depth;
  p.bg[0] = o.background[0]; p.bg[1] = o.background[1]; p.bg[2] = o.background[2];
  p.fb = fb;
  p.img_rgb = nullptr;
  p.img_rgb8 = nullptr;
  p.img_hit = nullptr;
  p.primface = f.primface.p; p.ksdeep = f.ksdeep.p; p.direct = f.direct.p;
  p.ray_o[0] = f.ray_o0.p; p.ray_o[1] = f.ray_o1.p; p.ray_d[0] = f.ray_d0.p; p.ray_d[1] = f.ray_d1.p;
  p.hit_slot = f.hit_slot.p;
  p.pbuf = f.pbuf.p; p.dbuf = f.dbuf.p; p.hitq = f.hitq.p; p.occl = f.occl.p;
  p.counters = s->counters.p;
}

// Tile groups.  The frame's owned tiles are cut into G contiguous groups (whole tile rows on one GPU, so a
// group is a horizontal band of the image); every group is an independent sub-frame - its own slice of every
// path buffer, its own counter block, its own share of the parked-ray list - and runs on its own stream.
// Each trace kernel ends in a latency-bound tail (the last batches of dependent node fetches leave most of
// the SMs idle); with G groups in flight the tail of one group's kernel is filled by the next group's work,
// and at the deep bounce levels, where a launch holds only a few thousand rays, the groups' small launches
// overlap instead of queueing behind each other.  Pixels never interact, so the image does not depend on G.
// Automatic choice, measured on B200 (profiles/r01_groups.md): three bands are best or within 1.5 % of the
// best on every workload (C3 0.826 -> 0.764 ms, C2 0.639 -> 0.621, C4 5.19 -> 4.5); more bands only add
// launches.  A band keeps at least 64 tiles (65 536 pixels) so that its trace launches still fill the GPU once.
int auto_groups(const FrameParams &p) {
  const int n_tiles = p.n_px / kTilePixels;
  return std::max(1, std::min(3, n_tiles / 64));
}

void plan_groups(const FrameParams &p, int world, int want, int *tile0 /* G + 1 */, int *n_groups) {
  const int n_tiles = p.n_px / kTilePixels;
  int G = (p.tile_ids && !p.user_rays) ? std::min(want, kMaxGroups) : 1;
  const int unit = world == 1 ? std::max(p.tiles_x, 1) : 1;  // one GPU: cut at tile-row boundaries
  const int units = (n_tiles + unit - 1) / unit;
  G = std::max(1, std::min(G, units));
  for (int g = 0; g <= G; ++g) tile0[g] = std::min(n_tiles, (int)((long long)units * g / G) * unit);
  tile0[G] = n_tiles;
  *n_groups = G;
}

FrameParams group_params(const FrameParams &p, int g, int t0, int t1) {
  FrameParams q = p;
  const size_t base = (size_t)t0 * kTilePixels;
  q.n_px = (t1 - t0) * kTilePixels;
  q.tile_ids = p.tile_ids + t0;
  if (p.fb) q.fb = p.fb + base;
  q.primface = p.primface + base;
  q.ksdeep = p.ksdeep + base;
  q.direct = p.direct + base * (size_t)p.max_depth;  // [max_depth][group pixels]
  for (int i = 0; i < 2; ++i) { q.ray_o[i] = p.ray_o[i] + base; q.ray_d[i] = p.ray_d[i] + base; }
  q.hit_slot = p.hit_slot + base;
  q.pbuf = p.pbuf + base;
  q.dbuf = p.dbuf + base;
  q.hitq = p.hitq + base;
  q.occl = p.occl + base * (size_t)std::max(p.S.n_lights, 1);
  q.counters = p.counters + g * kCtrInts;
  return q;
}

// Enqueues one frame: raygen, then per level closest / prep / anyhit / accum, then resolve - per tile group,
// group 0 on `st`, the others on the scene's group streams, forked from and joined back into `st`.
// profile: bracket every kernel with CUDA events (summed per kernel kind by read_stats); the frame then runs as
// ONE group, so that an event pair times one kernel alone on the GPU with the whole frame's rays of that level.
// after_group (optional): called once per group with its tile range and stream right after its resolve kernel
// (rtp_render starts the device-to-host copy of the band there).
template <class AfterGroup>
int enqueue_frame(rtp_scene *s, const FrameParams &p, cudaStream_t st, int *launches, bool profile, int n_primary,
                  int world, AfterGroup after_group) {
  // one frame at a time per scene: counters and path buffers are shared between frames
  if (s->has_last && s->last_stream != st) CK(cudaStreamSynchronize(s->last_stream));
  int tile0[kMaxGroups + 1], G = 1;
  plan_groups(p, world, profile ? 1 : s->groups > 0 ? s->groups : auto_groups(p), tile0, &G);
  CK(cudaMemsetAsync(s->counters.p, 0, (size_t)kCtrInts * G * sizeof(int), st));
  int n = 0;
  s->kev_used = 0;
  s->kev_kind.clear();
  auto mark = [&](int kind) -> cudaError_t {  // event before a kernel of `kind` (-1: end marker)
    if (!profile) return cudaSuccess;
    if (s->kev_used == (int)s->kev.size()) {
      cudaEvent_t e;
      cudaError_t err = cudaEventCreate(&e);
      if (err != cudaSuccess) return err;
      s->kev.push_back(e);
    }
    s->kev_kind.push_back(kind);
    return cudaEventRecord(s->kev[s->kev_used++], st);
  };
  s->tev_used = 0;
  s->tev_name.clear();
  auto stamp = [&](const char *what, int g, int d, cudaStream_t on) -> cudaError_t {
    if (!s->timeline || profile) return cudaSuccess;
    if (s->tev_used == (int)s->tev.size()) {
      cudaEvent_t e;
      cudaError_t err = cudaEventCreate(&e);
      if (err != cudaSuccess) return err;
      s->tev.push_back(e);
    }
    s->tev_name.push_back(std::string(what) + " g" + std::to_string(g) + " d" + std::to_string(d));
    return cudaEventRecord(s->tev[s->tev_used++], on);
  };
  FrameParams gp[kMaxGroups];
  LaunchCfg gc[kMaxGroups];
  cudaStream_t gs[kMaxGroups];
  for (int g = 0; g < G; ++g) {
    gp[g] = G == 1 ? p : group_params(p, g, tile0[g], tile0[g + 1]);
    gc[g] = s->cfg;
    gc[g].park_cap = s->cfg.park_cap / G;
    gc[g].groups = G;
    gc[g].park_slots = s->cfg.park_slots + (size_t)g * gc[g].park_cap * (kParkSlotBytes / sizeof(int));
    gs[g] = g == 0 && !(s->stagger && G > 1 && !profile) ? st : s->gstream[g];
    if (s->l2_persist_mb) { int lrc = apply_l2_window(s, gs[g]); if (lrc) return lrc; }
  }
  const bool fork = G > 1 && !profile;
  if (fork) {
    CK(cudaEventRecord(s->gfork, st));
    for (int g = s->stagger ? 0 : 1; g < G; ++g) CK(cudaStreamWaitEvent(gs[g], s->gfork, 0));
  }
  CK(stamp("start", 0, 0, st));
  // launches are issued phase by phase across the groups, so that the groups advance together
  if (p.user_rays)  // caller-supplied rays go through the ray list; camera rays are generated inside the trace kernel
    for (int g = 0; g < G; ++g) { CK(mark(2)); launch_raygen(gp[g], gs[g]); ++n; CK(stamp("raygen", g, 0, gs[g])); }
  // below the camera level the bounce levels run path by path in one persistent kernel (k_paths) instead of three
  // dependent launches per level; RTP_DEEP=wavefront keeps the per-level kernels (A/B runs, parity suite)
  // Which of the two runs the levels below the camera level - measured on B200 (profiles/r02_paths.md): the per-level
  // kernels keep neighbouring paths together and win on big frames, k_paths has no per-level latency floor and wins
  // once a rank's share is small (frame ms, per-level / k_paths: C4 8.3 M px 4.32 / 5.04, 2.1 M px 2.15 / 1.89,
  // 1.0 M px 1.72 / 1.39; C2 2.1 M px 0.563 / 0.592, 0.26 M px 0.305 / 0.268).  Scenes that spill out of the L2 have
  // the higher floor per level, hence the higher threshold.  RTP_DEEP=paths|wavefront forces one of them.
  const bool use_paths = s->deep_mode == 1 || (s->deep_mode < 0 && p.n_px <= (s->cfg.scene_exceeds_l2 ? s->paths_px_big : s->paths_px_small));
  const int wave_levels = use_paths ? std::min(p.max_depth, std::max(1, s->paths_from)) : p.max_depth;
  // The camera level in one kernel (k_camera) or as separate launches - measured on B200 (profiles/r02_camera.md).  On a
  // scene that fits the L2 nothing is parked and the fused kernel is level or ahead at every size (C2 0.575 / 0.576 ms,
  // a 1/8 share 0.269 / 0.283).  On a scene that spills out of it the separate launches win the big frames - their tails
  // overlap across the tile groups, and a parked PIXEL costs k_coop_camera its shadow rays and shading on top of the
  // primary ray (C3 0.709 / 0.646 ms) - and the fused kernel wins once a rank's share is small: one launch instead of
  // three dependent ones, each as long as its longest ray (C3 1/2 share 0.391 / 0.420, 1/4 0.252 / 0.278, 1/8 0.172 /
  // 0.206 ms); with four lights the crossover comes earlier (C4 1/8 share, 1.0 M pixels: 1.60 / 1.58).
  const bool fused = camera_fusable(p) &&
                     (s->fused >= 0 ? s->fused != 0
                                    : (!s->cfg.scene_exceeds_l2 || (long long)p.n_px * std::max(p.S.n_lights, 1) <= s->fused_px));
  // separate launches at the camera level: the shadow launch carries the level's shading when it does not park rays
  bool any_shade = !fused;
  for (int g = 0; g < G && any_shade; ++g) any_shade = anyhit_shade_ok(gp[g], gc[g]);
  for (int d = 0; d < wave_levels; ++d) {
    if (d == 0 && fused) {
      for (int g = 0; g < G; ++g) {
        CK(mark(4));
        if (s->timeline && !profile) {  // RTP_TIMELINE: where the parked-pixel kernel starts
          n += launch_camera(gp[g], gc[g], gs[g], 1); CK(stamp("camera", g, d, gs[g]));
          n += launch_camera(gp[g], gc[g], gs[g], 2); CK(stamp("camera coop", g, d, gs[g]));
        } else {
          n += launch_camera(gp[g], gc[g], gs[g]); CK(stamp("camera", g, d, gs[g]));
        }
      }
      continue;
    }
    for (int g = 0; g < G; ++g) { CK(mark(0)); n += launch_closest(gp[g], d, gc[g], gs[g]); CK(stamp("closest", g, d, gs[g])); }
    if (d == 0 && any_shade) {  // the camera level's shadow launch shades its pixels itself
      for (int g = 0; g < G; ++g) { CK(mark(1)); n += launch_anyhit_shade(gp[g], gc[g], gs[g]); CK(stamp("anyhit+shade", g, d, gs[g])); }
      continue;
    }
    for (int g = 0; g < G; ++g) { CK(mark(1)); n += launch_anyhit(gp[g], d, gc[g], gs[g]); CK(stamp("anyhit", g, d, gs[g])); }
    if (p.max_depth > 1)  // max_depth 1: the resolve kernel shades the only level itself
      for (int g = 0; g < G; ++g) { CK(mark(2)); launch_accum(gp[g], d, gs[g]); ++n; CK(stamp("accum", g, d, gs[g])); }
  }
  if (wave_levels < p.max_depth)
    for (int g = 0; g < G; ++g) { CK(mark(3)); launch_paths(gp[g], wave_levels, gc[g], gs[g]); ++n; CK(stamp("paths", g, wave_levels, gs[g])); }
  for (int g = 0; g < G; ++g) {
    if (!((fused || any_shade) && p.max_depth == 1)) {  // k_camera / the shading shadow launch has written a one-level frame's image
      CK(mark(2)); launch_resolve(gp[g], gs[g]); ++n;
      CK(stamp("resolve", g, 0, gs[g]));
    }
    CK(after_group(g, tile0[g], tile0[g + 1], gs[g]));
  }
  CK(mark(-1));
  CK(cudaGetLastError());
  if (fork)
    for (int g = s->stagger ? 0 : 1; g < G; ++g) {
      CK(cudaEventRecord(s->gjoin[g], gs[g]));
      CK(cudaStreamWaitEvent(st, s->gjoin[g], 0));
    }
  s->last_stream = st;
  s->has_last = true;
  s->last_launches = n;
  s->last_primary = n_primary;
  s->last_depth = p.max_depth;
  s->last_lights = p.S.n_lights;
  s->last_groups = G;
  if (launches) *launches = n;
  return RTP_OK;
}
int enqueue_frame(rtp_scene *s, const FrameParams &p, cudaStream_t st, int *launches, bool profile, int n_primary,
                  int world) {
  return enqueue_frame(s, p, st, launches, profile, n_primary, world,
                       [](int, int, int, cudaStream_t) { return cudaSuccess; });
}

int check_render_args(rtp_scene *s, const rtp_render_opts *o) {
  if (!s || !o) return fail(RTP_ERR_INVALID, "render: null argument");
  if (o->width <= 0 || o->height <= 0 || (long long)o->width * o->height > (1ll << 30))
    return fail(RTP_ERR_INVALID, "render: bad image size");
  if (o->max_depth < 1 || o->max_depth > kMaxDepth) return fail(RTP_ERR_INVALID, "render: max_depth must be in [1, 8]");
  if (o->world < 1 || o->rank < 0 || o->rank >= o->world) return fail(RTP_ERR_INVALID, "render: bad rank/world");
  if (s->async_head != s->async_tail && !s->in_async_call)
    return fail(RTP_ERR_STATE, "render: frames of rtp_render_async are outstanding (call rtp_render_wait first)");
  if (!s->have_camera) return fail(RTP_ERR_STATE, "render: rtp_set_camera has not been called");
  if (!s->have_lights) return fail(RTP_ERR_STATE, "render: rtp_set_lights has not been called");
  return RTP_OK;
}

// ray counting rule of SURVEY.md 8(d): 1 primary per pixel; per hit one shadow ray per light and, below the
// depth limit, one reflection ray.  hg: the counter blocks of the frame's G tile groups.
void stats_from_counters(const rtp_scene *s, const int *hg, int G, int primary, int depth, int lights, rtp_stats *st) {
  int h[kCtrInts] = {};
  for (int g = 0; g < G; ++g)
    for (int i = 0; i < kCtrInts; ++i) h[i] += hg[g * kCtrInts + i];
  st->rays_primary = primary;
  for (int d = 0; d < depth; ++d) {
    st->hits += h[kCtrHits + d];
    if (d > 0) st->rays_reflection += h[kCtrPaths + d];
  }
  st->rays_shadow = st->hits * lights;
  st->n_bvh_nodes = s->host.n_nodes;
  st->n_triangles = s->host.n_tris;
}

int read_stats(rtp_scene *s, rtp_stats *st, cudaStream_t stream) {
  int hg[kCtrInts * kMaxGroups];
  const int G = std::min(std::max(s->last_groups, 1), kMaxGroups);
  CK(cudaMemcpyAsync(hg, s->counters.p, (size_t)kCtrInts * G * sizeof(int), cudaMemcpyDeviceToHost, stream));
  CK(cudaStreamSynchronize(stream));
  for (int i = 0; i + 1 < s->kev_used; ++i) {
    float ms = 0.f;
    CK(cudaEventElapsedTime(&ms, s->kev[i], s->kev[i + 1]));
    const int kind = s->kev_kind[(size_t)i];
    (kind == 0 ? st->ms_closest : kind == 1 ? st->ms_anyhit : kind == 3 ? st->ms_paths : kind == 4 ? st->ms_camera : st->ms_shade) += ms;
  }
  stats_from_counters(s, hg, G, s->last_primary, s->last_depth, s->last_lights, st);
  return RTP_OK;
}

}  // namespace

extern "C" {

int rtp_frame_stats(rtp_scene *s, rtp_stats *st) {
  if (!s || !st) return fail(RTP_ERR_INVALID, "rtp_frame_stats: null argument");
  DeviceGuard g(s->device);
  std::memset(st, 0, sizeof *st);
  int rc = read_stats(s, st, s->has_last ? s->last_stream : s->stream);
  if (rc) return rc;
  st->kernel_launches = s->last_launches;
  if (s->timeline && s->tev_used > 1) {
    std::fprintf(stderr, "rtp timeline (ms after frame start, end of each launch):\n");
    for (int i = 1; i < s->tev_used; ++i) {
      float ms = 0.f;
      if (cudaEventSynchronize(s->tev[i]) != cudaSuccess || cudaEventElapsedTime(&ms, s->tev[0], s->tev[i]) != cudaSuccess) continue;
      std::fprintf(stderr, "  %-16s %8.4f\n", s->tev_name[(size_t)i].c_str(), ms);
    }
  }
  return RTP_OK;
}

size_t rtp_tile_buffer_bytes(int width, int height, int rank, int world) {
  if (width <= 0 || height <= 0 || world < 1) return 0;
  // every rank's buffer is padded to the largest tile count so that a plain gather works
  (void)rank;
  return (size_t)max_tile_count(width, height, world) * kTilePixels * sizeof(float4);
}

int rtp_tile_table(int width, int height, int rank, int world, int32_t *tiles_of_rank, int32_t *local_of_tile) {
  if (width <= 0 || height <= 0 || world < 1 || rank < 0 || rank >= world) return fail(RTP_ERR_INVALID, "rtp_tile_table: bad argument");
  std::vector<int32_t> tiles, local;
  tile_tables(width, height, rank, world, tiles, local);
  if (tiles_of_rank && !tiles.empty()) std::memcpy(tiles_of_rank, tiles.data(), tiles.size() * sizeof(int32_t));
  if (local_of_tile && !local.empty()) std::memcpy(local_of_tile, local.data(), local.size() * sizeof(int32_t));
  return (int)tiles.size();
}

int rtp_render_tiles_device(rtp_scene *s, const rtp_render_opts *o, void *d_tiles, void *stream, rtp_stats *st) {
  int rc = check_render_args(s, o);
  if (rc) return rc;
  if (!d_tiles) return fail(RTP_ERR_INVALID, "rtp_render_tiles_device: null tile buffer");
  DeviceGuard g(s->device);
  cudaStream_t cs = (cudaStream_t)stream;  // NULL = the legacy default stream, as documented
  const int n_px = own_tile_count(o->width, o->height, o->rank, o->world) * kTilePixels;
  rc = ensure_frame(s, s->frame, o->width, o->height, o->rank, o->world, std::max(o->max_depth, 1), n_px, true);
  if (rc) return rc;
  Frame &f = s->frame;
  FrameParams p;
  fill_params(s, f, *o, (float4 *)d_tiles, p);
  int launches = 0;
  if (st) CK(cudaEventRecord(s->ev[0], cs));
  rc = enqueue_frame(s, p, cs, &launches, (o->flags & RTP_RENDER_PROFILE) != 0, f.n_valid, o->world);
  if (rc) return rc;
  if (st) {
    CK(cudaEventRecord(s->ev[1], cs));
    std::memset(st, 0, sizeof *st);
    rc = read_stats(s, st, cs);
    if (rc) return rc;
    CK(cudaEventElapsedTime(&st->ms_trace, s->ev[0], s->ev[1]));
    st->ms_total = st->ms_trace;
    st->kernel_launches = launches;
  }
  return RTP_OK;
}

int rtp_render_image_device(rtp_scene *s, const rtp_render_opts *o, void *d_rgb8, void *d_hit, void *stream,
                            rtp_stats *st) {
  int rc = check_render_args(s, o);
  if (rc) return rc;
  if (!d_rgb8 && !d_hit) return fail(RTP_ERR_INVALID, "rtp_render_image_device: no output plane");
  if (s->barrier_timeouts && *(volatile int *)s->barrier_timeouts != 0)
    return fail(RTP_ERR_STATE, "rtp_render_image_device: an earlier frame barrier timed out (see rtp_peer_barrier)");
  DeviceGuard g(s->device);
  cudaStream_t cs = (cudaStream_t)stream;
  const int n_px = own_tile_count(o->width, o->height, o->rank, o->world) * kTilePixels;
  rc = ensure_frame(s, s->frame, o->width, o->height, o->rank, o->world, std::max(o->max_depth, 1), n_px, true);
  if (rc) return rc;
  Frame &f = s->frame;
  FrameParams p;
  fill_params(s, f, *o, nullptr, p);
  p.img_rgb8 = (unsigned char *)d_rgb8;
  p.img_hit = (int *)d_hit;
  int launches = 0;
  if (st) CK(cudaEventRecord(s->ev[0], cs));
  rc = enqueue_frame(s, p, cs, &launches, (o->flags & RTP_RENDER_PROFILE) != 0, f.n_valid, o->world);
  if (rc) return rc;
  if (st) {
    CK(cudaEventRecord(s->ev[1], cs));
    std::memset(st, 0, sizeof *st);
    rc = read_stats(s, st, cs);
    if (rc) return rc;
    CK(cudaEventElapsedTime(&st->ms_trace, s->ev[0], s->ev[1]));
    st->ms_total = st->ms_trace;
    st->kernel_launches = launches;
  }
  return RTP_OK;
}

int rtp_shared_alloc(int device, size_t bytes, void **d_ptr, void *handle) {
  if (!d_ptr || !handle || bytes == 0) return fail(RTP_ERR_INVALID, "rtp_shared_alloc: bad argument");
  static_assert(sizeof(cudaIpcMemHandle_t) == RTP_SHARED_HANDLE_BYTES, "CUDA IPC handle size");
  DeviceGuard g(device);
  if (!g.ok) return fail(RTP_ERR_CUDA, "rtp_shared_alloc: cudaSetDevice failed");
  *d_ptr = nullptr;
  CK(cudaMalloc(d_ptr, bytes));  // a dedicated allocation: an IPC handle always names a whole cudaMalloc block
  CK(cudaMemset(*d_ptr, 0, bytes));
  cudaIpcMemHandle_t h;
  cudaError_t e = cudaIpcGetMemHandle(&h, *d_ptr);
  if (e != cudaSuccess) {
    cudaFree(*d_ptr);
    *d_ptr = nullptr;
    return fail(RTP_ERR_CUDA, std::string("cudaIpcGetMemHandle: ") + cudaGetErrorString(e));
  }
  std::memcpy(handle, &h, sizeof h);
  return RTP_OK;
}

int rtp_peer_barrier(rtp_scene *s, void *d_counter, uint32_t target, int timeout_ms, void *stream) {
  if (!s || !d_counter || timeout_ms <= 0) return fail(RTP_ERR_INVALID, "rtp_peer_barrier: bad argument");
  if (((size_t)d_counter & 3) != 0) return fail(RTP_ERR_INVALID, "rtp_peer_barrier: counter must be 4-byte aligned");
  DeviceGuard g(s->device);
  if (!s->barrier_timeouts) {
    // page-locked and mapped: the barrier kernel counts a timeout through the device alias, the host sees it at the
    // next call without synchronising the device
    CK(cudaHostAlloc((void **)&s->barrier_timeouts, sizeof(int), cudaHostAllocMapped));
    *s->barrier_timeouts = 0;
    CK(cudaHostGetDevicePointer((void **)&s->barrier_timeouts_dev, s->barrier_timeouts, 0));
  }
  // sticky: once a wait has timed out the counter is short of the missing arrivals for good, every later frame would
  // spin for the full timeout and GPU 0's image may lack some ranks' tiles - refuse until the caller starts over
  // (new shared block + rtp_peer_barrier_reset on every rank)
  if (*(volatile int *)s->barrier_timeouts != 0)
    return fail(RTP_ERR_STATE, "rtp_peer_barrier: an earlier frame barrier timed out (a rank did not arrive); the image of that "
                               "frame is incomplete. Re-create the shared block and call rtp_peer_barrier_reset on every rank");
  launch_peer_barrier((unsigned *)d_counter, target, (unsigned long long)timeout_ms * 1000000ull, s->barrier_timeouts_dev,
                      (cudaStream_t)stream);
  CK(cudaGetLastError());
  return RTP_OK;
}

int rtp_peer_barrier_timeouts(rtp_scene *s) {
  if (!s) return fail(RTP_ERR_INVALID, "rtp_peer_barrier_timeouts: null scene");
  if (!s->barrier_timeouts) return 0;
  DeviceGuard g(s->device);
  CK(cudaDeviceSynchronize());  // barriers still in flight have finished (or given up) when this returns
  return *(volatile int *)s->barrier_timeouts;
}

int rtp_peer_barrier_reset(rtp_scene *s) {
  if (!s) return fail(RTP_ERR_INVALID, "rtp_peer_barrier_reset: null scene");
  if (!s->barrier_timeouts) return RTP_OK;
  DeviceGuard g(s->device);
  CK(cudaDeviceSynchronize());
  *s->barrier_timeouts = 0;
  return RTP_OK;
}

int rtp_shared_open(int device, const void *handle, void **d_ptr) {
  if (!d_ptr || !handle) return fail(RTP_ERR_INVALID, "rtp_shared_open: bad argument");
  DeviceGuard g(device);
  if (!g.ok) return fail(RTP_ERR_CUDA, "rtp_shared_open: cudaSetDevice failed");
  cudaIpcMemHandle_t h;
  std::memcpy(&h, handle, sizeof h);
  *d_ptr = nullptr;
  CK(cudaIpcOpenMemHandle(d_ptr, h, cudaIpcMemLazyEnablePeerAccess));
  return RTP_OK;
}

int rtp_shared_free(int device, void *d_ptr, int opened) {
  if (!d_ptr) return RTP_OK;
  DeviceGuard g(device);
  if (!g.ok) return fail(RTP_ERR_CUDA, "rtp_shared_free: cudaSetDevice failed");
  if (opened) CK(cudaIpcCloseMemHandle(d_ptr));
  else CK(cudaFree(d_ptr));
  return RTP_OK;
}

int rtp_untile_device(rtp_scene *s, int width, int height, int world, const void *d_all, size_t rank_stride_bytes,
                      void *d_rgb, void *d_rgb8, void *d_hit, void *stream) {
  if (!s || !d_all || width <= 0 || height <= 0 || world < 1) return fail(RTP_ERR_INVALID, "rtp_untile_device: bad argument");
  if (rank_stride_bytes % sizeof(float4)) return fail(RTP_ERR_INVALID, "rtp_untile_device: stride must be a multiple of 16");
  DeviceGuard g(s->device);
  cudaStream_t cs = (cudaStream_t)stream;
  Frame &f = s->frame;
  if (f.width != width || f.height != height || f.world != world || !f.local_of_tile.p)
    return fail(RTP_ERR_STATE, "rtp_untile_device: render a frame of this size/world first");
  launch_untile((const float4 *)d_all, rank_stride_bytes / sizeof(float4), f.local_of_tile.p, width, height, world, -1,
                (float *)d_rgb, (unsigned char *)d_rgb8, (int *)d_hit, cs);
  CK(cudaGetLastError());
  return RTP_OK;
}

}  // extern "C"

namespace {

struct Planes {  // finished row-major image planes in device memory (of this or of another device); any may be null
  float *rgb;
  unsigned char *rgb8;
  int *hit;
};

// Enqueues the share (rank of world) of the frame that scene s renders, resolved straight into dst.
template <class AfterGroup>
int enqueue_share(rtp_scene *s, const rtp_render_opts &o, int rank, int world, const Planes &dst, cudaStream_t cs,
                  int *launches, AfterGroup after_group) {
  const int n_px = own_tile_count(o.width, o.height, rank, world) * kTilePixels;
  int rc = ensure_frame(s, s->frame, o.width, o.height, rank, world, std::max(o.max_depth, 1), n_px, true);
  if (rc) return rc;
  FrameParams p;
  fill_params(s, s->frame, o, nullptr, p);
  p.img_rgb = dst.rgb;
  p.img_rgb8 = dst.rgb8;
  p.img_hit = dst.hit;
  return enqueue_frame(s, p, cs, launches, (o.flags & RTP_RENDER_PROFILE) != 0, s->frame.n_valid, world, after_group);
}

int ensure_staging(rtp_scene *s, size_t bytes) {
  if (bytes <= s->staging_bytes) return RTP_OK;
  if (s->staging) CK(cudaFreeHost(s->staging));
  s->staging = nullptr;
  s->staging_bytes = 0;
  CK(cudaHostAlloc((void **)&s->staging, bytes, cudaHostAllocDefault));
  s->staging_bytes = bytes;
  return RTP_OK;
}

}  // namespace

extern "C" {

int rtp_render(rtp_scene *s, const rtp_render_opts *o, float *rgb, uint8_t *rgb8, int32_t *hit, rtp_stats *st) {
  int rc = check_render_args(s, o);
  if (rc) return rc;
  const bool multi = !s->peers.empty();
  if (multi && o->world != 1)
    return fail(RTP_ERR_INVALID, "rtp_render: a multi-device scene shards the frame itself (world must be 1)");
  if (!rgb && !rgb8 && !hit && !st) return fail(RTP_ERR_INVALID, "rtp_render: no output requested");
  DeviceGuard g(s->device);
  cudaStream_t cs = s->stream;
  Frame &f = s->frame;
  const size_t npix = (size_t)o->width * o->height;
  // the resolve kernel writes the finished row-major planes itself; they live on this (the first) device
  if (rgb) CK(f.out_rgb.ensure(npix * 3));
  if (rgb8 || (!rgb && !hit)) CK(f.out_rgb8.ensure(npix * 3));  // statistics only: render into a scratch plane
  if (hit) CK(f.out_hit.ensure(npix));
  const Planes dst{rgb ? f.out_rgb.p : nullptr, (rgb8 || (!rgb && !hit)) ? f.out_rgb8.p : nullptr, hit ? f.out_hit.p : nullptr};
  const int world = multi ? 1 + (int)s->peers.size() : o->world;
  const int rank = multi ? 0 : o->rank;
  int launches = 0;
  CK(cudaEventRecord(s->ev[0], cs));
  // one GPU, whole frame: a tile group is a band of image rows, and its device-to-host copy starts on the group's
  // stream as soon as that band is resolved, under the other groups' kernels
  const bool banded = !multi && world == 1;
  const int tiles_x = (o->width + kTileDim - 1) / kTileDim, width = o->width, height = o->height;
  auto copy_band = [=](int, int t0, int t1, cudaStream_t gs) -> cudaError_t {
    if (!banded || t1 <= t0) return cudaSuccess;
    const size_t y0 = (size_t)(t0 / tiles_x) * kTileDim;
    const size_t y1 = std::min<size_t>((size_t)height, (size_t)((t1 + tiles_x - 1) / tiles_x) * kTileDim);
    if (y1 <= y0) return cudaSuccess;
    cudaError_t e = cudaSuccess;
    if (rgb) e = cudaMemcpyAsync(rgb + y0 * width * 3, dst.rgb + y0 * width * 3, (y1 - y0) * width * 3 * sizeof(float), cudaMemcpyDeviceToHost, gs);
    if (e == cudaSuccess && rgb8) e = cudaMemcpyAsync(rgb8 + y0 * width * 3, dst.rgb8 + y0 * width * 3, (y1 - y0) * width * 3, cudaMemcpyDeviceToHost, gs);
    if (e == cudaSuccess && hit)
      e = cudaMemcpyAsync(hit + y0 * width, dst.hit + y0 * width, (y1 - y0) * width * sizeof(int), cudaMemcpyDeviceToHost, gs);
    return e;
  };
  const bool quick_stats = st && !(o->flags & RTP_RENDER_PROFILE);
  if (multi) {
    // every other device renders its tiles on its own stream, fed by its own host thread, and stores them into this
    // device's planes over NVLink (peer access); fork and join are CUDA events - no kernel waits for another
    // (replaces the std::thread fan-out and join of raytraceScene, flyscene.cpp:266-289)
    CK(cudaEventRecord(s->frame_start, cs));
    for (size_t r = 0; r < s->peers.size(); ++r) {
      rtp_scene *p = s->peers[r];
      const rtp_render_opts opts = *o;
      const int prank = (int)r + 1;
      cudaEvent_t start = s->frame_start;
      s->workers[r]->submit([p, opts, prank, world, dst, start, quick_stats]() -> int {
        DeviceGuard pg(p->device);
        int prc = RTP_OK;
        cudaError_t e = cudaStreamWaitEvent(p->stream, start, 0);
        if (e != cudaSuccess) prc = fail(RTP_ERR_CUDA, std::string("cudaStreamWaitEvent: ") + cudaGetErrorString(e));
        int n = 0;
        if (!prc) prc = enqueue_share(p, opts, prank, world, dst, p->stream, &n, [](int, int, int, cudaStream_t) { return cudaSuccess; });
        if (!prc && quick_stats) {
          // this device's counters ride behind its share, ahead of frame_done: the caller's one wait covers them
          // (a read_stats per device afterwards was a copy + a stream synchronise each: ~0.1 ms on 8 devices)
          if (!p->counters_pinned) e = cudaHostAlloc((void **)&p->counters_pinned, sizeof(int) * kCtrInts * kMaxGroups, cudaHostAllocDefault);
          if (e == cudaSuccess)
            e = cudaMemcpyAsync(p->counters_pinned, p->counters.p, sizeof(int) * kCtrInts * (size_t)std::max(p->last_groups, 1),
                                cudaMemcpyDeviceToHost, p->stream);
          if (e != cudaSuccess) prc = fail(RTP_ERR_CUDA, std::string("counters copy: ") + cudaGetErrorString(e));
        }
        if (!prc && (e = cudaEventRecord(p->frame_done, p->stream)) != cudaSuccess)
          prc = fail(RTP_ERR_CUDA, std::string("cudaEventRecord: ") + cudaGetErrorString(e));
        if (prc) p->err = g_err;
        return prc;
      });
    }
  }
  rc = enqueue_share(s, *o, rank, world, dst, cs, &launches, copy_band);
  if (multi) {
    int wrc = RTP_OK;
    for (size_t r = 0; r < s->peers.size(); ++r) {
      const int prc = s->workers[r]->wait();
      if (prc && !wrc) { wrc = prc; g_err = s->peers[r]->err; }
    }
    if (!rc) rc = wrc;
    if (!rc)
      for (rtp_scene *p : s->peers) CK(cudaStreamWaitEvent(cs, p->frame_done, 0));
  }
  if (rc) return rc;
  CK(cudaEventRecord(s->ev[1], cs));
  if (quick_stats) {  // the frame's counters ride behind the frame on the same stream: one host wait for image and statistics
    if (!s->counters_pinned) CK(cudaHostAlloc((void **)&s->counters_pinned, sizeof(int) * kCtrInts * kMaxGroups, cudaHostAllocDefault));
    CK(cudaMemcpyAsync(s->counters_pinned, s->counters.p, sizeof(int) * kCtrInts * (size_t)s->last_groups, cudaMemcpyDeviceToHost, cs));
  }
  const bool scatter = !multi && world > 1;  // one virtual rank of a sharded frame (tests, external multi-process drivers)
  if (multi) {
    if (rgb) CK(cudaMemcpyAsync(rgb, dst.rgb, npix * 3 * sizeof(float), cudaMemcpyDeviceToHost, cs));
    if (rgb8) CK(cudaMemcpyAsync(rgb8, dst.rgb8, npix * 3, cudaMemcpyDeviceToHost, cs));
    if (hit) CK(cudaMemcpyAsync(hit, dst.hit, npix * sizeof(int), cudaMemcpyDeviceToHost, cs));
  } else if (scatter) {
    // only this rank's tiles were written on the device: the planes go to a page-locked staging block and the owned
    // tiles are copied from there into the caller's buffers, whose other pixels stay untouched
    const size_t o_rgb = 0, o_rgb8 = o_rgb + (rgb ? npix * 12 : 0), o_hit = (o_rgb8 + (rgb8 ? npix * 3 : 0) + 15) / 16 * 16;
    rc = ensure_staging(s, o_hit + (hit ? npix * 4 : 0) + 16);
    if (rc) return rc;
    unsigned char *stg = (unsigned char *)s->staging;
    if (rgb) CK(cudaMemcpyAsync(stg + o_rgb, dst.rgb, npix * 12, cudaMemcpyDeviceToHost, cs));
    if (rgb8) CK(cudaMemcpyAsync(stg + o_rgb8, dst.rgb8, npix * 3, cudaMemcpyDeviceToHost, cs));
    if (hit) CK(cudaMemcpyAsync(stg + o_hit, dst.hit, npix * 4, cudaMemcpyDeviceToHost, cs));
    CK(cudaStreamSynchronize(cs));
    for (int32_t t : f.tiles_host) {
      const int x0 = (t % tiles_x) * kTileDim, y0 = (t / tiles_x) * kTileDim;
      const int w = std::min(kTileDim, width - x0), y1 = std::min(height, y0 + kTileDim);
      for (int y = y0; y < y1; ++y) {
        const size_t at = (size_t)y * width + x0;
        if (rgb) std::memcpy(rgb + 3 * at, stg + o_rgb + 12 * at, (size_t)w * 12);
        if (rgb8) std::memcpy(rgb8 + 3 * at, stg + o_rgb8 + 3 * at, (size_t)w * 3);
        if (hit) std::memcpy(hit + at, stg + o_hit + 4 * at, (size_t)w * 4);
      }
    }
  }
  CK(cudaStreamSynchronize(cs));
  if (st) {
    std::memset(st, 0, sizeof *st);
    if (quick_stats) {
      stats_from_counters(s, s->counters_pinned, std::max(s->last_groups, 1), s->last_primary, s->last_depth, s->last_lights, st);
    } else {
      rc = read_stats(s, st, cs);
      if (rc) return rc;
    }
    st->kernel_launches = launches;
    for (rtp_scene *p : s->peers) {  // the other devices' rays and launches
      rtp_stats ps;
      std::memset(&ps, 0, sizeof ps);
      if (quick_stats) {  // landed with the frame (the caller's stream waited for frame_done)
        stats_from_counters(p, p->counters_pinned, std::max(p->last_groups, 1), p->last_primary, p->last_depth, p->last_lights, &ps);
      } else {
        DeviceGuard pg(p->device);
        rc = read_stats(p, &ps, p->stream);
        if (rc) return rc;
      }
      st->rays_primary += ps.rays_primary; st->rays_shadow += ps.rays_shadow; st->rays_reflection += ps.rays_reflection;
      st->hits += ps.hits;
      st->ms_closest = std::max(st->ms_closest, ps.ms_closest);
      st->ms_anyhit = std::max(st->ms_anyhit, ps.ms_anyhit);
      st->ms_shade = std::max(st->ms_shade, ps.ms_shade);
      st->ms_paths = std::max(st->ms_paths, ps.ms_paths);
      st->ms_camera = std::max(st->ms_camera, ps.ms_camera);
      st->kernel_launches += p->last_launches;
    }
    CK(cudaEventElapsedTime(&st->ms_trace, s->ev[0], s->ev[1]));  // one GPU: includes the overlapped band copies
    st->ms_total = st->ms_trace;
  }
  return RTP_OK;
}

int rtp_render_async(rtp_scene *s, const rtp_render_opts *o, uint8_t *rgb8, int32_t *hit) {
  if (s) s->in_async_call = true;
  int rc = check_render_args(s, o);
  if (s) s->in_async_call = false;
  if (rc) return rc;
  if (!rgb8 && !hit) return fail(RTP_ERR_INVALID, "rtp_render_async: no output plane");
  if (o->world != 1) return fail(RTP_ERR_INVALID, "rtp_render_async: one GPU per scene (world must be 1)");
  if (o->flags & RTP_RENDER_PROFILE) return fail(RTP_ERR_INVALID, "rtp_render_async: RTP_RENDER_PROFILE needs the synchronous call");
  if (s->async_head - s->async_tail >= 2) return fail(RTP_ERR_STATE, "rtp_render_async: two frames are already outstanding");
  DeviceGuard g(s->device);
  cudaStream_t cs = s->stream;
  rtp_scene::AsyncSlot &a = s->aslot[s->async_head & 1];
  if (!a.counters_host) {  // first use of the slot
    CK(cudaHostAlloc((void **)&a.counters_host, sizeof(int) * kCtrInts * kMaxGroups, cudaHostAllocDefault));
    CK(cudaEventCreate(&a.ev0));
    CK(cudaEventCreate(&a.ev1));
    for (auto &e : a.copied) CK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
  }
  for (int i = 0; i < kMaxGroups; ++i)
    if (!s->cstream[i]) {
      CK(cudaStreamCreateWithFlags(&s->cstream[i], cudaStreamNonBlocking));
      CK(cudaEventCreateWithFlags(&s->resolved[i], cudaEventDisableTiming));
    }
  Frame &f = s->frame;
  const int n_px = own_tile_count(o->width, o->height, 0, 1) * kTilePixels;
  rc = ensure_frame(s, f, o->width, o->height, 0, 1, std::max(o->max_depth, 1), n_px, true);
  if (rc) return rc;
  const size_t npix = (size_t)o->width * o->height;
  if (rgb8) CK(a.rgb8.ensure(npix * 3));
  if (hit) CK(a.hit.ensure(npix));
  FrameParams p;
  fill_params(s, f, *o, nullptr, p);
  p.img_rgb8 = rgb8 ? a.rgb8.p : nullptr;
  p.img_hit = hit ? a.hit.p : nullptr;
  const int tiles_x = f.tiles_x, width = o->width, height = o->height;
  unsigned char *d_rgb8 = a.rgb8.p;
  int *d_hit = a.hit.p;
  int n_groups = 0;
  // a band's copy runs on the copy stream of its group, behind the band's resolve kernel only: the group's own
  // stream is free for the next frame at once
  auto copy_band = [&](int gi, int t0, int t1, cudaStream_t gs) -> cudaError_t {
    n_groups = std::max(n_groups, gi + 1);
    cudaError_t e = cudaEventRecord(s->resolved[gi], gs);
    if (e == cudaSuccess) e = cudaStreamWaitEvent(s->cstream[gi], s->resolved[gi], 0);
    const size_t y0 = (size_t)(t0 / tiles_x) * kTileDim;
    const size_t y1 = std::min<size_t>((size_t)height, (size_t)((t1 + tiles_x - 1) / tiles_x) * kTileDim);
    if (e == cudaSuccess && rgb8 && y1 > y0)
      e = cudaMemcpyAsync(rgb8 + y0 * width * 3, d_rgb8 + y0 * width * 3, (y1 - y0) * width * 3, cudaMemcpyDeviceToHost, s->cstream[gi]);
    if (e == cudaSuccess && hit && y1 > y0)
      e = cudaMemcpyAsync(hit + y0 * width, d_hit + y0 * width, (y1 - y0) * width * sizeof(int), cudaMemcpyDeviceToHost, s->cstream[gi]);
    if (e == cudaSuccess) e = cudaEventRecord(a.copied[gi], s->cstream[gi]);
    return e;
  };
  int launches = 0;
  CK(cudaEventRecord(a.ev0, cs));
  rc = enqueue_frame(s, p, cs, &launches, false, f.n_valid, 1, copy_band);
  if (rc) return rc;
  CK(cudaEventRecord(a.ev1, cs));
  // the frame's counters, before the next frame zeroes them (stream order)
  CK(cudaMemcpyAsync(a.counters_host, s->counters.p, sizeof(int) * kCtrInts * (size_t)s->last_groups, cudaMemcpyDeviceToHost, cs));
  a.groups = n_groups;
  a.launches = launches;
  a.primary = f.n_valid;
  a.depth = o->max_depth;
  a.lights = s->dS.n_lights;
  a.busy = true;
  ++s->async_head;
  return RTP_OK;
}

int rtp_render_wait(rtp_scene *s, rtp_stats *st) {
  if (!s) return fail(RTP_ERR_INVALID, "rtp_render_wait: null scene");
  if (s->async_head == s->async_tail) return fail(RTP_ERR_STATE, "rtp_render_wait: no frame outstanding");
  DeviceGuard g(s->device);
  rtp_scene::AsyncSlot &a = s->aslot[s->async_tail & 1];
  for (int gi = 0; gi < a.groups; ++gi) CK(cudaEventSynchronize(a.copied[gi]));
  // the counters copy is the last operation of the frame on the scene's stream; ev1 precedes it
  CK(cudaEventSynchronize(a.ev1));
  if (st) {
    std::memset(st, 0, sizeof *st);
    // wait for the counters snapshot itself: it was enqueued right after ev1 on the same stream
    if (s->async_head - s->async_tail == 1) CK(cudaStreamSynchronize(s->stream));
    else CK(cudaEventSynchronize(s->aslot[(s->async_tail + 1) & 1].ev0));  // recorded after that copy
    stats_from_counters(s, a.counters_host, std::max(a.groups, 1), a.primary, a.depth, a.lights, st);
    CK(cudaEventElapsedTime(&st->ms_trace, a.ev0, a.ev1));
    st->ms_total = st->ms_trace;
    st->kernel_launches = a.launches;
  }
  a.busy = false;
  ++s->async_tail;
  return RTP_OK;
}

int rtp_trace_rays_device(rtp_scene *s, const void *d_rays, int64_t n, int mode, void *d_face, void *d_t,
                          void *d_occluded, void *stream, rtp_stats *st) {
  if (!s || !d_rays || n < 0 || n > (int64_t)0x7f000000) return fail(RTP_ERR_INVALID, "rtp_trace_rays_device: bad argument");
  DeviceGuard g(s->device);
  cudaStream_t cs = (cudaStream_t)stream;
  // one job at a time per scene (as enqueue_frame): the batch shares the scene's counter block and parked-ray list
  // with the frames - closest-hit batches do park their stragglers on scenes larger than the L2
  if (s->has_last && s->last_stream != cs) CK(cudaStreamSynchronize(s->last_stream));
  s->last_stream = cs;
  s->has_last = true;
  int *cursor = s->counters.p + kCtrAux;
  CK(cudaMemsetAsync(cursor, 0, sizeof(int), cs));
  CK(cudaMemsetAsync(s->counters.p + kCtrPark + 2 * 16, 0, 2 * sizeof(int), cs));
  if (st) CK(cudaEventRecord(s->ev[0], cs));
  int launches = 0;
  if (mode == RTP_RAYS_CLOSEST) {
    if (!d_face) return fail(RTP_ERR_INVALID, "rtp_trace_rays_device: face output required");
    launches = launch_rays_closest(s->dS, (const float *)d_rays, n, (int *)d_face, (float *)d_t, cursor, s->cfg, cs);
  } else if (mode == RTP_RAYS_ANYHIT) {
    if (!d_occluded) return fail(RTP_ERR_INVALID, "rtp_trace_rays_device: occluded output required");
    launches = launch_rays_anyhit(s->dS, (const float *)d_rays, n, (unsigned char *)d_occluded, cursor, s->cfg, cs);
  } else {
    return fail(RTP_ERR_INVALID, "rtp_trace_rays_device: mode must be CLOSEST or ANYHIT");
  }
  CK(cudaGetLastError());
  if (st) {
    CK(cudaEventRecord(s->ev[1], cs));
    CK(cudaStreamSynchronize(cs));
    std::memset(st, 0, sizeof *st);
    CK(cudaEventElapsedTime(&st->ms_trace, s->ev[0], s->ev[1]));
    st->ms_total = st->ms_trace;
    st->kernel_launches = launches;
    (mode == RTP_RAYS_ANYHIT ? st->rays_shadow : st->rays_primary) = n;
    st->n_bvh_nodes = s->host.n_nodes;
    st->n_triangles = s->host.n_tris;
  }
  return RTP_OK;
}

int rtp_trace_rays_device_noalloc(rtp_scene *s, const void *d_rays, int64_t n, int mode, void *d_face, void *d_t,
                                  void *d_occluded, void *stream) {
  if (!s || !d_rays || n < 0) return fail(RTP_ERR_INVALID, "rtp_trace_rays_device_noalloc: bad argument");
  if (n > (int64_t)RTP_NOALLOC_MAX_RAYS)
    return fail(RTP_ERR_INVALID, "rtp_trace_rays_device_noalloc: " + std::to_string(n) + " rays exceed the limit of " +
                                     std::to_string(RTP_NOALLOC_MAX_RAYS) + " per call (split the batch, or use rtp_trace_rays_device)");
  if (mode != RTP_RAYS_CLOSEST && mode != RTP_RAYS_ANYHIT)
    return fail(RTP_ERR_INVALID, "rtp_trace_rays_device_noalloc: mode must be CLOSEST or ANYHIT");
  if (mode == RTP_RAYS_CLOSEST && !d_face) return fail(RTP_ERR_INVALID, "rtp_trace_rays_device_noalloc: face output required");
  if (mode == RTP_RAYS_ANYHIT && !d_occluded) return fail(RTP_ERR_INVALID, "rtp_trace_rays_device_noalloc: occluded output required");
  if (n == 0) return RTP_OK;
  DeviceGuard g(s->device);
  // nothing of the scene's mutable state is touched (no counters, no parked-ray list, no events): the launch reads the
  // scene arrays and works in the library's static scratch, which it shares with every other call on this device
  int launches = 0;
  CK(launch_rays_noalloc(s->device, s->dS, (const float *)d_rays, n, mode == RTP_RAYS_ANYHIT, (int *)d_face, (float *)d_t,
                         (unsigned char *)d_occluded, s->cfg, (cudaStream_t)stream, &launches));
  return RTP_OK;
}

int rtp_trace_rays(rtp_scene *s, const float *rays, int64_t n, int mode, int max_depth, const float background[3],
                   float *rgb, int32_t *face, float *t, uint8_t *occluded, rtp_stats *st) {
  if (!s || !rays || n <= 0 || n > (int64_t)0x7f000000) return fail(RTP_ERR_INVALID, "rtp_trace_rays: bad argument");
  DeviceGuard g(s->device);
  cudaStream_t cs = s->stream;
  {  // far origins need wider boxes (ensure_pad); |o| is known here, on the host
    float dist = 0.f;
    for (int64_t i = 0; i < n; ++i)
      for (int c = 0; c < 3; ++c) {
        const float a = std::fabs(rays[6 * i + c]);
        if (a > dist && std::isfinite(a)) dist = a;
      }
    int prc = ensure_pad(s, dist);
    if (prc) return prc;
  }
  CK(s->rays_dev.ensure((size_t)n * 6));
  CK(cudaMemcpyAsync(s->rays_dev.p, rays, (size_t)n * 6 * sizeof(float), cudaMemcpyHostToDevice, cs));
  if (mode == RTP_RAYS_CLOSEST) {
    if (!face) return fail(RTP_ERR_INVALID, "rtp_trace_rays: face output required");
    CK(s->rays_face.ensure((size_t)n));
    CK(s->rays_t.ensure((size_t)n));
    int rc = rtp_trace_rays_device(s, s->rays_dev.p, n, mode, s->rays_face.p, s->rays_t.p, nullptr, cs, st);
    if (rc) return rc;
    CK(cudaMemcpyAsync(face, s->rays_face.p, (size_t)n * sizeof(int), cudaMemcpyDeviceToHost, cs));
    if (t) CK(cudaMemcpyAsync(t, s->rays_t.p, (size_t)n * sizeof(float), cudaMemcpyDeviceToHost, cs));
    CK(cudaStreamSynchronize(cs));
    return RTP_OK;
  }
  if (mode == RTP_RAYS_ANYHIT) {
    if (!occluded) return fail(RTP_ERR_INVALID, "rtp_trace_rays: occluded output required");
    CK(s->rays_occ.ensure((size_t)n));
    int rc = rtp_trace_rays_device(s, s->rays_dev.p, n, mode, nullptr, nullptr, s->rays_occ.p, cs, st);
    if (rc) return rc;
    CK(cudaMemcpyAsync(occluded, s->rays_occ.p, (size_t)n, cudaMemcpyDeviceToHost, cs));
    CK(cudaStreamSynchronize(cs));
    return RTP_OK;
  }
  if (mode != RTP_RAYS_SHADE) return fail(RTP_ERR_INVALID, "rtp_trace_rays: unknown mode");
  // Flyscene::traceRay for arbitrary rays: the frame pipeline with caller-supplied depth-0 rays
  if (!rgb) return fail(RTP_ERR_INVALID, "rtp_trace_rays: rgb output required");
  if (max_depth < 1 || max_depth > kMaxDepth) return fail(RTP_ERR_INVALID, "rtp_trace_rays: max_depth must be in [1, 8]");
  if (!s->have_lights) return fail(RTP_ERR_STATE, "rtp_trace_rays: rtp_set_lights has not been called");
  Frame &f = s->rframe;
  int rc = ensure_frame(s, f, 0, 0, 0, 1, std::max(max_depth, 1), (int)n, false);
  if (rc) return rc;
  CK(f.fb.ensure((size_t)n));
  rtp_render_opts o;
  std::memset(&o, 0, sizeof o);
  o.max_depth = max_depth;
  const float bgdef[3] = {(float)0.9, (float)0.9, (float)0.9};
  std::memcpy(o.background, background ? background : bgdef, sizeof o.background);
  FrameParams p;
  fill_params(s, f, o, f.fb.p, p);
  std::memset(&p.cam, 0, sizeof p.cam);
  p.user_rays = s->rays_dev.p;
  p.n_user_rays = (int)n;
  int launches = 0;
  CK(cudaEventRecord(s->ev[0], cs));
  rc = enqueue_frame(s, p, cs, &launches, false, (int)n, 1);
  if (rc) return rc;
  CK(cudaEventRecord(s->ev[1], cs));
  rc = ensure_staging(s, (size_t)n * sizeof(float4));  // page-locked landing place of the (r, g, b, face) records
  if (rc) return rc;
  const float4 *h = (const float4 *)s->staging;
  CK(cudaMemcpyAsync(s->staging, f.fb.p, (size_t)n * sizeof(float4), cudaMemcpyDeviceToHost, cs));
  CK(cudaStreamSynchronize(cs));
  for (int64_t i = 0; i < n; ++i) {
    rgb[3 * i] = h[(size_t)i].x; rgb[3 * i + 1] = h[(size_t)i].y; rgb[3 * i + 2] = h[(size_t)i].z;
    if (face) std::memcpy(&face[i], &h[(size_t)i].w, 4);
  }
  if (st) {
    std::memset(st, 0, sizeof *st);
    rc = read_stats(s, st, cs);
    if (rc) return rc;
    CK(cudaEventElapsedTime(&st->ms_trace, s->ev[0], s->ev[1]));
    st->ms_total = st->ms_trace;
    st->kernel_launches = launches;
  }
  return RTP_OK;
}

}  // extern "C"
