// bvh_build_gpu.cu - the linear BVH2 of rtp_device.h built ON THE DEVICE (LBVH).
//
// Caller side of the hot path: the reference builds its BoundingBox list once on
// the CPU (flyscene.cpp:399-428, BoundingBox.cpp:109-149); the host builder of this repo
// (bvh_build.cpp, binned SAH) gives the best trees but takes seconds on the 10 M-triangle config.
// This is the fast path: Morton codes of the triangle centroids -> one radix sort -> Karras'
// parallel radix tree (one thread per inner node, "Maximizing Parallelism in the Construction of
// BVHs, Octrees, and k-d Trees", HPG 2012) -> bottom-up AABB fit with arrival counters -> subtrees of
// <= 4 triangles collapsed into leaves -> 64-byte nodes holding both children's boxes, and the
// triangle arrays permuted into leaf order.  Like the host builder it only has to be CONSERVATIVE:
// which faces a ray may see is decided by the exact tests (tri_exact, ref_box_pass), so the two
// builders give bit-identical images (tests/test_gpu_parity.py::test_gpu_built_bvh_*).
//
// Library use: cub::DeviceRadixSort / DeviceScan for the sort and the prefix sum (set-up plumbing,
// not the hot path); every other step is a hand-written kernel below.
#include <cub/cub.cuh>

#include <cfloat>

#include "bvh_build_gpu.h"
#include "rtp_device.h"

namespace rtp {

namespace {

constexpr int kLeafBit = 1 << 30;  // child reference: kLeafBit | sorted position = one primitive

__device__ __forceinline__ unsigned expand10(unsigned v) {  // 10 bits -> every third bit
  v = (v * 0x00010001u) & 0xFF0000FFu;
  v = (v * 0x00000101u) & 0x0F00F00Fu;
  v = (v * 0x00000011u) & 0xC30C30C3u;
  v = (v * 0x00000005u) & 0x49249249u;
  return v;
}

// padded box of every primitive (tris in rank order: 4 float4 per triangle) + its 64-bit sort key:
// 30-bit Morton code of the centroid, then the primitive index (keys are unique)
__global__ void k_prim_keys(const float4 *tris, int n, float pad, float3 lo, float3 inv_ext, float4 *pbox,
                            unsigned long long *keys) {
  const int r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= n) return;
  const float4 a = tris[4 * (size_t)r], b = tris[4 * (size_t)r + 1], c = tris[4 * (size_t)r + 2];
  float3 mn = make_float3(fminf(a.x, fminf(b.x, c.x)), fminf(a.y, fminf(b.y, c.y)), fminf(a.z, fminf(b.z, c.z)));
  float3 mx = make_float3(fmaxf(a.x, fmaxf(b.x, c.x)), fmaxf(a.y, fmaxf(b.y, c.y)), fmaxf(a.z, fmaxf(b.z, c.z)));
  const bool finite = isfinite(a.x + a.y + a.z + b.x + b.y + b.z + c.x + c.y + c.z);
  if (finite) {
    mn.x -= pad; mn.y -= pad; mn.z -= pad;
    mx.x += pad; mx.y += pad; mx.z += pad;
  } else {  // as the host flattener: a triangle with a non-finite vertex is a candidate for every ray
    mn = make_float3(-FLT_MAX, -FLT_MAX, -FLT_MAX);
    mx = make_float3(FLT_MAX, FLT_MAX, FLT_MAX);
  }
  pbox[2 * (size_t)r] = make_float4(mn.x, mn.y, mn.z, 0.f);
  pbox[2 * (size_t)r + 1] = make_float4(mx.x, mx.y, mx.z, 0.f);
  unsigned code = 0;
  if (finite) {
    const float cx = fminf(fmaxf((0.5f * (mn.x + mx.x) - lo.x) * inv_ext.x, 0.f), 1.f);
    const float cy = fminf(fmaxf((0.5f * (mn.y + mx.y) - lo.y) * inv_ext.y, 0.f), 1.f);
    const float cz = fminf(fmaxf((0.5f * (mn.z + mx.z) - lo.z) * inv_ext.z, 0.f), 1.f);
    code = (expand10((unsigned)fminf(cx * 1024.f, 1023.f)) << 2) | (expand10((unsigned)fminf(cy * 1024.f, 1023.f)) << 1) |
           expand10((unsigned)fminf(cz * 1024.f, 1023.f));
  }
  keys[r] = ((unsigned long long)code << 32) | (unsigned)r;
}

__device__ __forceinline__ int delta(const unsigned long long *k, int n, int i, int j) {
  return (j < 0 || j >= n) ? -1 : __clzll(k[i] ^ k[j]);
}

// Karras 2012, one thread per inner node i in [0, n-1): its range, split and children
__global__ void k_radix_tree(const unsigned long long *keys, int n, int2 *child, int2 *range, int *parent_inner,
                             int *parent_leaf) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n - 1) return;
  const int d = delta(keys, n, i, i + 1) - delta(keys, n, i, i - 1) >= 0 ? 1 : -1;
  const int dmin = delta(keys, n, i, i - d);
  int lmax = 2;
  while (delta(keys, n, i, i + lmax * d) > dmin) lmax <<= 1;
  int l = 0;
  for (int t = lmax >> 1; t >= 1; t >>= 1)
    if (delta(keys, n, i, i + (l + t) * d) > dmin) l += t;
  const int j = i + l * d;
  const int dnode = delta(keys, n, i, j);
  int s = 0;
  for (int t = (l + 1) >> 1;; t = (t + 1) >> 1) {
    if (delta(keys, n, i, i + (s + t) * d) > dnode) s += t;
    if (t <= 1) break;
  }
  const int gamma = i + s * d + min(d, 0);
  const int first = min(i, j), last = max(i, j);
  const int left = first == gamma ? (kLeafBit | gamma) : gamma;
  const int right = last == gamma + 1 ? (kLeafBit | (gamma + 1)) : gamma + 1;
  child[i] = make_int2(left, right);
  range[i] = make_int2(first, last);
  if (left & kLeafBit) parent_leaf[gamma] = i; else parent_inner[gamma] = i;
  if (right & kLeafBit) parent_leaf[gamma + 1] = i; else parent_inner[gamma + 1] = i;
  if (i == 0) parent_inner[0] = -1;
}

// bottom-up fit: the second thread to arrive at a node merges the children's boxes and goes on
__global__ void k_fit(const unsigned long long *keys, int n, const float4 *pbox, const int2 *child, const int *parent_inner,
                      const int *parent_leaf, int *arrived, float4 *nbox, int *height, int *max_height) {  // nbox/height: read back with __ldcg
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= n) return;
  int node = parent_leaf[p];
  while (node >= 0) {
    __threadfence();
    if (atomicAdd(arrived + node, 1) == 0) return;  // the sibling subtree is not done yet
    const int2 c = child[node];
    float4 lo0, hi0, lo1, hi1;
    int h0 = 0, h1 = 0;
    if (c.x & kLeafBit) {
      const unsigned r = (unsigned)keys[c.x & ~kLeafBit];
      lo0 = pbox[2 * (size_t)r]; hi0 = pbox[2 * (size_t)r + 1];
    } else {
      lo0 = __ldcg(nbox + 2 * (size_t)c.x); hi0 = __ldcg(nbox + 2 * (size_t)c.x + 1);  // L2: written by another SM
      h0 = __ldcg(height + c.x);
    }
    if (c.y & kLeafBit) {
      const unsigned r = (unsigned)keys[c.y & ~kLeafBit];
      lo1 = pbox[2 * (size_t)r]; hi1 = pbox[2 * (size_t)r + 1];
    } else {
      lo1 = __ldcg(nbox + 2 * (size_t)c.y); hi1 = __ldcg(nbox + 2 * (size_t)c.y + 1);
      h1 = __ldcg(height + c.y);
    }
    nbox[2 * (size_t)node] = make_float4(fminf(lo0.x, lo1.x), fminf(lo0.y, lo1.y), fminf(lo0.z, lo1.z), 0.f);
    nbox[2 * (size_t)node + 1] = make_float4(fmaxf(hi0.x, hi1.x), fmaxf(hi0.y, hi1.y), fmaxf(hi0.z, hi1.z), 0.f);
    const int h = max(h0, h1) + 1;
    height[node] = h;
    if (node == 0) *max_height = h;
    node = parent_inner[node];
  }
}

// an inner node survives when its subtree holds more than max_leaf primitives (the root always does)
__global__ void k_mark(const int2 *range, int n, int max_leaf, int *keep) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n - 1) return;
  keep[i] = (i == 0 || range[i].y - range[i].x + 1 > max_leaf) ? 1 : 0;
}

__device__ __forceinline__ int leaf_code(int first, int count) { return ~((first << 3) | (count - 1)); }

// surviving inner nodes -> the 64-byte device layout (both children's boxes + references)
__global__ void k_emit(const unsigned long long *keys, int n, const float4 *pbox, const float4 *nbox, const int2 *child,
                       const int2 *range, const int *keep, const int *new_index, float4 *nodes) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n - 1 || !keep[i]) return;
  const int2 c = child[i];
  float4 lo[2], hi[2];
  int ref[2];
  for (int k = 0; k < 2; ++k) {
    const int ch = k == 0 ? c.x : c.y;
    if (ch & kLeafBit) {
      const int p = ch & ~kLeafBit;
      const unsigned r = (unsigned)keys[p];
      lo[k] = pbox[2 * (size_t)r]; hi[k] = pbox[2 * (size_t)r + 1];
      ref[k] = leaf_code(p, 1);
    } else {
      lo[k] = nbox[2 * (size_t)ch]; hi[k] = nbox[2 * (size_t)ch + 1];
      ref[k] = keep[ch] ? new_index[ch] : leaf_code(range[ch].x, range[ch].y - range[ch].x + 1);
    }
  }
  float4 *o = nodes + 4 * (size_t)new_index[i];
  o[0] = make_float4(lo[0].x, hi[0].x, lo[0].y, hi[0].y);
  o[1] = make_float4(lo[1].x, hi[1].x, lo[1].y, hi[1].y);
  o[2] = make_float4(lo[0].z, hi[0].z, lo[1].z, hi[1].z);
  o[3] = make_float4(__int_as_float(ref[0]), __int_as_float(ref[1]), 0.f, 0.f);
}

// triangle arrays from rank order into leaf (sorted) order
__global__ void k_permute(const unsigned long long *keys, int n, const float4 *tris_in, const float4 *vnorm_in,
                          const int *face_in, float4 *tris, float4 *vnorm, int *face_of) {
  const int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= n) return;
  const unsigned r = (unsigned)keys[s];
#pragma unroll
  for (int k = 0; k < 4; ++k) tris[4 * (size_t)s + k] = tris_in[4 * (size_t)r + k];
#pragma unroll
  for (int k = 0; k < 3; ++k) vnorm[3 * (size_t)s + k] = vnorm_in[3 * (size_t)r + k];
  face_of[s] = face_in[r];
}

template <class T>
struct Tmp {
  T *p = nullptr;
  cudaError_t alloc(size_t n) { return cudaMalloc((void **)&p, (n ? n : 1) * sizeof(T)); }
  ~Tmp() { if (p) cudaFree(p); }
};

#define LB_CK(call)                              \
  do {                                           \
    cudaError_t e_ = (call);                     \
    if (e_ != cudaSuccess) return e_;            \
  } while (0)

}  // namespace

cudaError_t build_lbvh_device(const float4 *d_tris_in, const float4 *d_vnorm_in, const int *d_face_in, int n, float pad,
                              const float scene_lo[3], const float scene_hi[3], int max_leaf, float4 *d_tris, float4 *d_vnorm,
                              int *d_face_of, float4 *d_nodes, int *n_nodes, int *height, cudaStream_t st) {
  *n_nodes = 0;
  *height = 0;
  if (n < 2) return cudaErrorInvalidValue;  // the caller keeps tiny scenes on the host builder
  Tmp<float4> pbox, nbox;
  Tmp<unsigned long long> keys, keys_sorted;
  Tmp<int2> child, range;
  Tmp<int> parent_inner, parent_leaf, arrived, node_height, keep, new_index, scalars;
  LB_CK(pbox.alloc(2 * (size_t)n)); LB_CK(nbox.alloc(2 * (size_t)n));
  LB_CK(keys.alloc(n)); LB_CK(keys_sorted.alloc(n));
  LB_CK(child.alloc(n)); LB_CK(range.alloc(n));
  LB_CK(parent_inner.alloc(n)); LB_CK(parent_leaf.alloc(n)); LB_CK(arrived.alloc(n)); LB_CK(node_height.alloc(n));
  LB_CK(keep.alloc(n)); LB_CK(new_index.alloc(n)); LB_CK(scalars.alloc(4));
  const int B = 256, gp = (n + B - 1) / B, gi = (n - 1 + B - 1) / B;
  const float3 lo = make_float3(scene_lo[0], scene_lo[1], scene_lo[2]);
  const float3 inv = make_float3(1.f / fmaxf(scene_hi[0] - scene_lo[0], 1e-30f), 1.f / fmaxf(scene_hi[1] - scene_lo[1], 1e-30f),
                                 1.f / fmaxf(scene_hi[2] - scene_lo[2], 1e-30f));
  k_prim_keys<<<gp, B, 0, st>>>(d_tris_in, n, pad, lo, inv, pbox.p, keys.p);
  size_t tmp_bytes = 0, scan_bytes = 0;
  LB_CK(cub::DeviceRadixSort::SortKeys(nullptr, tmp_bytes, keys.p, keys_sorted.p, n, 0, 62, st));
  LB_CK(cub::DeviceScan::ExclusiveSum(nullptr, scan_bytes, keep.p, new_index.p, n - 1, st));
  Tmp<unsigned char> tmp;
  LB_CK(tmp.alloc(tmp_bytes > scan_bytes ? tmp_bytes : scan_bytes));
  LB_CK(cub::DeviceRadixSort::SortKeys(tmp.p, tmp_bytes, keys.p, keys_sorted.p, n, 0, 62, st));
  LB_CK(cudaMemsetAsync(arrived.p, 0, (size_t)n * sizeof(int), st));
  LB_CK(cudaMemsetAsync(scalars.p, 0, 4 * sizeof(int), st));
  k_radix_tree<<<gi, B, 0, st>>>(keys_sorted.p, n, child.p, range.p, parent_inner.p, parent_leaf.p);
  k_fit<<<gp, B, 0, st>>>(keys_sorted.p, n, pbox.p, child.p, parent_inner.p, parent_leaf.p, arrived.p, nbox.p, node_height.p,
                          scalars.p);
  k_mark<<<gi, B, 0, st>>>(range.p, n, max_leaf, keep.p);
  LB_CK(cub::DeviceScan::ExclusiveSum(tmp.p, scan_bytes, keep.p, new_index.p, n - 1, st));
  k_emit<<<gi, B, 0, st>>>(keys_sorted.p, n, pbox.p, nbox.p, child.p, range.p, keep.p, new_index.p, d_nodes);
  k_permute<<<gp, B, 0, st>>>(keys_sorted.p, n, d_tris_in, d_vnorm_in, d_face_in, d_tris, d_vnorm, d_face_of);
  LB_CK(cudaGetLastError());
  int h_last[2] = {0, 0}, h_height = 0;
  LB_CK(cudaMemcpyAsync(&h_last[0], new_index.p + (n - 2), sizeof(int), cudaMemcpyDeviceToHost, st));
  LB_CK(cudaMemcpyAsync(&h_last[1], keep.p + (n - 2), sizeof(int), cudaMemcpyDeviceToHost, st));
  LB_CK(cudaMemcpyAsync(&h_height, scalars.p, sizeof(int), cudaMemcpyDeviceToHost, st));
  LB_CK(cudaStreamSynchronize(st));
  *n_nodes = h_last[0] + h_last[1];
  *height = h_height;
  return cudaSuccess;
}

}  // namespace rtp
