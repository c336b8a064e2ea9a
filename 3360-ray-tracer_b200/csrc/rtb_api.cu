// rtb_api.cu — implementation of the C ABI in include/rtb.h: scene upload, the wavefront
// driver loop, and the batch-intersect seam.  All device work is issued on one library stream
// per thread and timed with CUDA events on that stream.  No CPU fallback anywhere: without a
// device every compute entry returns RTB_ENODEVICE.
#include <rtb.h>

#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <condition_variable>
#include <functional>
#include <map>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "rtb_compile.h"
#include "rtb_build.cuh"
#include "rtb_kernels.cuh"
#include "rtb_pool.cuh"

namespace {

thread_local std::string g_error;
thread_local cudaStream_t g_stream = nullptr;
thread_local int g_device = -1;
thread_local int g_sm_count = 0;

int Fail(int code, const std::string& msg) {
  g_error = msg;
  return code;
}
#define RTB_CUDA(call)                                                                                   \
  do {                                                                                                   \
    cudaError_t e_ = (call);                                                                             \
    if (e_ != cudaSuccess)                                                                               \
      return Fail(RTB_ECUDA, std::string(#call) + ": " + cudaGetErrorString(e_) + " (" __FILE__ ":" + std::to_string(__LINE__) + ")"); \
  } while (0)

int EnsureInit() {
  if (g_stream) return RTB_OK;
  return rtb_init(0);
}

// cudaEvent_t owner: every early return (RTB_CUDA) destroys what was created so far
struct Events {
  std::vector<cudaEvent_t> all;
  int Create(cudaEvent_t* e, unsigned flags = cudaEventDefault) {
    RTB_CUDA(cudaEventCreateWithFlags(e, flags));
    all.push_back(*e);
    return RTB_OK;
  }
  ~Events() {
    for (cudaEvent_t e : all) cudaEventDestroy(e);
  }
};
#define RTB_TRY(expr)                 \
  do {                                \
    const int rc_ = (expr);           \
    if (rc_ != RTB_OK) return rc_;    \
  } while (0)

// Wavefront state, per host thread (= per device), allocated lazily by rtb_render* and kept across
// scenes and calls.  A Wave is one independent wavefront: its own path pool (two queues), hit
// arrays, counters and stream.  Several waves are kept in flight so that the drain tail of one
// wave's persistent extend kernel, and the host's read of its survivor count, overlap the other
// waves' kernels.
struct Wave {
  rtb::PathQueue q[2]{};
  float* hit_t = nullptr;
  uint32_t* hit_ref = nullptr;
  unsigned int* counts = nullptr;      // device: [0] traversal cursor, [1] shade survivor count, [2..3] 64-bit ray counter of the path kernel, [4..5] camera samples started (adaptive)
  unsigned int* host_count = nullptr;  // pinned
  cudaStream_t stream = nullptr;
  cudaEvent_t done = nullptr;
  std::vector<void*> mem;
  uint32_t live = 0;
  int cur = 0;
  bool pending = false;
};
struct RenderCtx {
  std::vector<Wave> waves;
  Wave fused;  // default mode: no queues, only the counters, their pinned mirror and a stream
  Wave groups[4];  // adaptive renders: one light wave (counters + stream) per pixel group, so that the tail of one group's pass overlaps the others' passes
  uint32_t pool = 0;
  uint32_t* pixel_order = nullptr;
  int order_w = 0, order_h = 0;
  float* accum = nullptr;  // rtb_render's device accumulator (host-buffer entry), grow-only
  size_t accum_bytes = 0;
  char* adapt = nullptr;   // adaptive renders: two active-pixel lists, their counts, one chunk of sample radiances; grow-only
  size_t adapt_bytes = 0;
  char* sched = nullptr;   // cost-ordered schedule: per-pixel path lengths, the re-ordered pixel list, sort keys / values / scratch; grow-only
  size_t sched_bytes = 0;
  uint32_t sched_pixels = 0;  // the pixel list in `sched` was sorted for an image of this many pixels (RTB_COST_ORDER_REUSE experiment)
};
thread_local RenderCtx g_ctx;

void FreeWaves(RenderCtx* s) {
  for (auto& w : s->waves) {
    if (w.stream) cudaStreamSynchronize(w.stream);
    for (void* p : w.mem) cudaFree(p);
    if (w.host_count) cudaFreeHost(w.host_count);
    if (w.done) cudaEventDestroy(w.done);
    if (w.stream) cudaStreamDestroy(w.stream);
  }
  s->waves.clear();
  s->pool = 0;
}
void FreeLight(Wave& w) {
  if (w.stream) cudaStreamSynchronize(w.stream);
  for (void* p : w.mem) cudaFree(p);
  if (w.done) cudaEventDestroy(w.done);
  if (w.stream) cudaStreamDestroy(w.stream);
  w = Wave{};
}
void FreeFused(RenderCtx* s) {
  FreeLight(s->fused);
  for (Wave& w : s->groups) FreeLight(w);
}

}  // namespace

// Device-built scenes of at most this many primitives are rendered without the tree (trace_loop's flat loop: every lane tests every
// primitive in lock step); host-built ones when the reference's tree is a single leaf.  profiles/r02_flat_loop_ab.jsonl, DESIGN.md 3.9.
#ifndef RTB_FLAT_MAX_PRIMS
#define RTB_FLAT_MAX_PRIMS 16
#endif
constexpr uint32_t kFlatMaxPrims = RTB_FLAT_MAX_PRIMS;

// shared memory of the traversal kernels besides the staged top of the tree: the stack columns
constexpr size_t kStackSmem = (size_t)rtb::kSmemStack * rtb::kTraceBlock * sizeof(uint32_t);

// The device image of a compiled scene, kept in pinned host memory by the RtbCompiledTree it belongs to: the first
// rtb_scene_create that uses the tree lays every table out exactly as in the device arena; later ones are ONE
// asynchronous copy (no float shadows to rebuild, nothing to stage).
struct ArenaImage {
  std::mutex mu;  // rtb_render_multi creates the scene on every device at once: one of them fills the image
  char* pinned = nullptr;
  size_t bytes = 0;
  bool filled = false;
  uint32_t counts[6] = {0, 0, 0, 0, 0, 0};  // spheres, triangles, rects, materials, textures, images at snapshot time
};
struct RtbCompiledTree {
  rtbc::Compiled tree;
  ArenaImage image;
  ~RtbCompiledTree() {
    if (image.pinned) cudaFreeHost(image.pinned);
  }
};

#ifndef RTB_TREE_REBUILD_DEFAULT
#define RTB_TREE_REBUILD_DEFAULT false
#endif
static int CompileTree(const RtbSceneDesc& d, rtbc::Compiled* out) {
  try {
    RtbSceneDesc dd = d;
#ifdef RTB_TOP_PAIRS  // tuning builds: size of the shared-memory prefix, overriding the flattener's
    dd.n_top_nodes = 2u * (uint32_t)(RTB_TOP_PAIRS);
#endif
    uint32_t leaf_max = 4;  // tuning knobs of the device tree (rtb_compile.h)
    float cost[3];
    const float* prim_cost = nullptr;
    if (const char* e = std::getenv("RTB_LEAF_MAX")) leaf_max = (uint32_t)std::max(1, std::atoi(e));
    if (const char* e = std::getenv("RTB_PRIM_COST")) {
      cost[0] = cost[1] = cost[2] = (float)std::atof(e);
      prim_cost = cost;
    }
    bool rebuild = RTB_TREE_REBUILD_DEFAULT;
    if (const char* e = std::getenv("RTB_TREE_REBUILD")) rebuild = std::atoi(e) != 0;
    const char* wide = std::getenv("RTB_WIDE4");  // the 4-wide tree next to the pairs (RTB_RENDER_WIDE4; an experiment)
    *out = rtbc::Compiler(dd, leaf_max, prim_cost).Run(rebuild, wide && std::atoi(wide) != 0);
    if (out->quad_stack + 2 > (uint32_t)rtb::kStackSize) out->quads.clear();  // RTB_RENDER_WIDE4 is refused for this scene
  } catch (const std::exception& e) {
    return Fail(RTB_EINVAL, std::string("scene compile: ") + e.what());
  }
  if (out->pairs.size() >= (1u << 29)) return Fail(RTB_EINVAL, "more than 2^29 BVH pairs");
  if (out->depth + 2 > (uint32_t)rtb::kStackSize) return Fail(RTB_EINVAL, "BVH deeper than the 64-entry traversal stack (bvh.h:80)");
  return RTB_OK;
}

static int ValidateDesc(const RtbSceneDesc* d);
// per-primitive material / axis checks move to the device above this many primitives (a host pass over 5 M records is 22 ms)
static bool ValidatesPrimitivesOnDevice(const RtbSceneDesc* d) {
  return (uint64_t)d->n_spheres + d->n_triangles + d->n_rects > (1u << 18) && !std::getenv("RTB_HOST_VALIDATE");
}

// Pinned staging for scene uploads, per host thread, grow-only: rtb_scene_create copies every array into it and
// issues one asynchronous H2D copy per array into ONE device arena (one cudaMalloc / cudaFree per scene).
struct Staging {
  char* host = nullptr;
  size_t bytes = 0;
};
thread_local Staging g_staging;

struct RtbScene {
  rtb::DevScene dev{};
  char* arena = nullptr;  // every table of the scene, 256-byte aligned sub-ranges
  size_t arena_bytes = 0;
  struct Pending {
    const void* src;
    size_t bytes, off;
    const void** out;
    bool fresh;  // differs from one rtb_scene_create to the next (texture object handles): never taken from an ArenaImage
  };
  std::vector<Pending> pending;
  std::vector<cudaArray_t> arrays;
  std::vector<cudaTextureObject_t> textures;
  rtb::HitCounters* counters = nullptr;
  unsigned int* cursor = nullptr;  // persistent-thread work cursor + shade output counter (2 words)
  size_t smem_bytes = 0;
  int device = 0;
  bool flat_ok = false;  // rendered without the tree (set by rtb_scene_create)
  size_t Reserve(size_t bytes) {
    const size_t off = (arena_bytes + 255) & ~(size_t)255;
    arena_bytes = off + bytes;
    return off;
  }
  // queue `n` elements for upload; *out is set by Commit.  `host` must stay valid until then.
  template <class T>
  int Put(const T* host, size_t n, const T** out, bool fresh = false) {
    *out = nullptr;
    if (n == 0) return RTB_OK;
    pending.push_back(Pending{host, n * sizeof(T), Reserve(n * sizeof(T)), reinterpret_cast<const void**>(out), fresh});
    return RTB_OK;
  }
  int FinishCommit(cudaStream_t st, size_t ctr_off) {
    RTB_CUDA(cudaMemsetAsync(arena + ctr_off, 0, sizeof(rtb::HitCounters) + 16, st));
    counters = reinterpret_cast<rtb::HitCounters*>(arena + ctr_off);
    cursor = reinterpret_cast<unsigned int*>(arena + ctr_off + sizeof(rtb::HitCounters));
    RTB_CUDA(cudaStreamSynchronize(st));  // the staging buffer and the caller's arrays are free again
    pending.clear();
    return RTB_OK;
  }
  // img != nullptr: upload through (and, the first time, fill) the compiled tree's pinned arena image
  int Commit(cudaStream_t st, ArenaImage* img = nullptr, const uint32_t* counts = nullptr) {
    const size_t ctr_off = Reserve(sizeof(rtb::HitCounters) + 16);
    if (img && arena_bytes <= (256u << 20)) {
      std::unique_lock<std::mutex> fill_lock(img->mu);
      if (!img->filled || img->bytes != arena_bytes || (counts && std::memcmp(img->counts, counts, sizeof(img->counts)) != 0)) {
        for (const Pending& p : pending)
          if (!p.fresh && !p.src) return Fail(RTB_EINVAL, "scene tables changed since rtb_tree_compile");
        if (img->pinned) cudaFreeHost(img->pinned);
        img->pinned = nullptr;
        img->filled = false;
        RTB_CUDA(cudaMallocHost(reinterpret_cast<void**>(&img->pinned), arena_bytes));
        img->bytes = arena_bytes;
        for (const Pending& p : pending)
          if (!p.fresh) std::memcpy(img->pinned + p.off, p.src, p.bytes);
        if (counts) std::memcpy(img->counts, counts, sizeof(img->counts));
        img->filled = true;
      }
      fill_lock.unlock();
      RTB_CUDA(cudaMallocAsync(reinterpret_cast<void**>(&arena), arena_bytes, st));
      RTB_CUDA(cudaMemcpyAsync(arena, img->pinned, ctr_off, cudaMemcpyHostToDevice, st));
      for (const Pending& p : pending) {
        if (p.fresh) RTB_CUDA(cudaMemcpyAsync(arena + p.off, p.src, p.bytes, cudaMemcpyHostToDevice, st));  // a few bytes, pageable
        *p.out = arena + p.off;
      }
      return FinishCommit(st, ctr_off);
    }
    // tables above 32 MB (the stress scene) go straight from the caller's memory; pinning gigabytes costs more than it saves
    constexpr size_t kDirect = 32u << 20;
    size_t staged = 0;
    for (const Pending& p : pending)
      if (p.bytes <= kDirect) staged += (p.bytes + 255) & ~(size_t)255;
    if (g_staging.bytes < staged) {
      if (g_staging.host) cudaFreeHost(g_staging.host);
      g_staging = Staging{};
      const size_t want = staged + staged / 4;
      RTB_CUDA(cudaMallocHost(reinterpret_cast<void**>(&g_staging.host), want));
      g_staging.bytes = want;
    }
    RTB_CUDA(cudaMallocAsync(reinterpret_cast<void**>(&arena), arena_bytes, st));  // pool-backed: re-creating a scene reuses the block
    size_t so = 0;
    for (const Pending& p : pending) {  // the DMA of one table overlaps the host copy of the next
      if (p.bytes <= kDirect) {
        std::memcpy(g_staging.host + so, p.src, p.bytes);
        RTB_CUDA(cudaMemcpyAsync(arena + p.off, g_staging.host + so, p.bytes, cudaMemcpyHostToDevice, st));
        so += (p.bytes + 255) & ~(size_t)255;
      } else {
        RTB_CUDA(cudaMemcpyAsync(arena + p.off, p.src, p.bytes, cudaMemcpyHostToDevice, st));
      }
      *p.out = arena + p.off;
    }
    return FinishCommit(st, ctr_off);
  }
};

// The library's state (stream, pools, accumulator) is per host thread = per device, so a scene is used from a thread whose
// library device (rtb_init) is the one the scene was created on; anything else would launch its device pointers
// on another GPU's stream.
static int CheckSceneDevice(const RtbScene* s) {
  if (s->device == g_device) return RTB_OK;
  return Fail(RTB_EINVAL, "scene lives on device " + std::to_string(s->device) + " but this thread's library device is " + std::to_string(g_device) +
                              ": call rtb_init(" + std::to_string(s->device) + ") first");
}

// RTB_SCENE_DEVICE_BUILD: the pair records of a linear BVH over every primitive, written straight into the scene's
// arena by the kernels of rtb_build.cuh.  Scratch lives in the stream-ordered pool for the duration of the build.
__global__ void build_single_kernel(rtb::DevScene s, uint32_t n_s, uint32_t n_t, float* pairs) {
  uint32_t type, idx;
  rtb::split_index(0u, n_s, n_t, type, idx);
  const rtb::BuildBox b = rtb::prim_box(s, type, idx);
  const uint32_t desc = rtb::kLeafBit | (type << 29) | idx;
  rtb::store_half(pairs, 0u, b, desc);  // one primitive: both children hold it (a self-tie never replaces)
  rtb::store_half(pairs, 1u, b, desc);
}

static int BuildTreeOnDevice(RtbScene* s, uint32_t n_s, uint32_t n_t, uint32_t n, float* d_pairs) {
  cudaStream_t st = g_stream;
  if (n == 1) {
    build_single_kernel<<<1, 1, 0, st>>>(s->dev, n_s, n_t, d_pairs);
    RTB_CUDA(cudaStreamSynchronize(st));
    return RTB_OK;
  }
  auto up = [](size_t b) { return (b + 255) & ~(size_t)255; };
  size_t cub_bytes = 0;
  RTB_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, cub_bytes, (const unsigned long long*)nullptr, (unsigned long long*)nullptr,
                                           (const uint32_t*)nullptr, (uint32_t*)nullptr, (int)n, 0, 63, st));
  const size_t o_boxes = 0, o_keys0 = o_boxes + up((size_t)n * sizeof(rtb::BuildBox)), o_keys1 = o_keys0 + up((size_t)n * 8),
               o_ord0 = o_keys1 + up((size_t)n * 8), o_ord1 = o_ord0 + up((size_t)n * 4), o_pint = o_ord1 + up((size_t)n * 4),
               o_pleaf = o_pint + up((size_t)n * 4), o_flags = o_pleaf + up((size_t)n * 4), o_misc = o_flags + up((size_t)n * 4),
               o_cub = o_misc + 256, total = o_cub + up(cub_bytes);
  char* w = nullptr;
  RTB_CUDA(cudaMallocAsync(reinterpret_cast<void**>(&w), total, st));
  auto* boxes = reinterpret_cast<rtb::BuildBox*>(w + o_boxes);
  auto* keys0 = reinterpret_cast<unsigned long long*>(w + o_keys0);
  auto* keys1 = reinterpret_cast<unsigned long long*>(w + o_keys1);
  auto* ord0 = reinterpret_cast<uint32_t*>(w + o_ord0);
  auto* ord1 = reinterpret_cast<uint32_t*>(w + o_ord1);
  auto* pint = reinterpret_cast<uint32_t*>(w + o_pint);
  auto* pleaf = reinterpret_cast<uint32_t*>(w + o_pleaf);
  auto* flags = reinterpret_cast<unsigned*>(w + o_flags);
  auto* misc = reinterpret_cast<unsigned*>(w + o_misc);  // [0..2] min, [3..5] max of the centres (ordered bits), [6] depth
  cudaError_t e = cudaMemsetAsync(misc, 0xff, 12, st);
  if (e == cudaSuccess) e = cudaMemsetAsync(misc + 3, 0, 16, st);
  if (e == cudaSuccess) e = cudaMemsetAsync(flags, 0, (size_t)n * 4, st);
  const unsigned grid = (n + 255) / 256;
  if (e == cudaSuccess) {
    rtb::build_bounds_kernel<<<grid, 256, 0, st>>>(s->dev, n_s, n_t, n, boxes, misc);
    rtb::build_morton_kernel<<<grid, 256, 0, st>>>(boxes, n, misc, keys0, ord0);
    e = cub::DeviceRadixSort::SortPairs(w + o_cub, cub_bytes, keys0, keys1, ord0, ord1, (int)n, 0, 63, st);
  }
  unsigned depth = 0;
  if (e == cudaSuccess) {
    rtb::build_hierarchy_kernel<<<grid, 256, 0, st>>>(keys1, (int)n, pint, pleaf);
    rtb::build_emit_kernel<<<grid, 256, 0, st>>>(boxes, ord1, n, n_s, n_t, pint, pleaf, d_pairs, flags, misc + 6);
    e = cudaMemcpyAsync(&depth, misc + 6, 4, cudaMemcpyDeviceToHost, st);
  }
  if (e == cudaSuccess) e = cudaStreamSynchronize(st);
  if (e == cudaSuccess) e = cudaGetLastError();
  cudaFreeAsync(w, st);
  if (e != cudaSuccess) return Fail(RTB_ECUDA, std::string("device tree build: ") + cudaGetErrorString(e));
  if (depth + 2 > (unsigned)rtb::kStackSize) return Fail(RTB_EINVAL, "device-built BVH deeper than the 64-entry traversal stack (bvh.h:80)");
  return RTB_OK;
}

#ifdef RTB_CHECKED
__global__ void check_selftest_kernel() { RTB_CHECK(rtb::kChkClasses - 1, threadIdx.x != 3u); }
#endif

extern "C" {

const char* rtb_last_error(void) { return g_error.c_str(); }
int rtb_abi_version(void) { return RTB_ABI_VERSION; }

int rtb_tree_compile(const RtbSceneDesc* d, RtbCompiledTree** out) {
  if (!d || !out) return Fail(RTB_EINVAL, "null argument");
  *out = nullptr;
  int rc = ValidateDesc(d);
  if (rc != RTB_OK) return rc;
  RtbCompiledTree* t = new RtbCompiledTree();
  rc = CompileTree(*d, &t->tree);
  if (rc != RTB_OK) {
    delete t;
    return rc;
  }
  *out = t;
  return RTB_OK;
}
void rtb_tree_destroy(RtbCompiledTree* t) { delete t; }

// Walks a compiled tree (pairs: width 2, quads: width 4) from record 0: primitives reachable through leaf descriptors, and
// child boxes that stick out of the slot box that leads to them by more than 1e-5 relative (boxes of re-split leaves are recomputed
// from the primitives with a 2e-6 pad, the reference's own carry 1e-6: triangle.h:18-39; anything beyond that is a layout bug).
static void WalkTree(const std::vector<rtbc::Pair>& rec, int width, uint64_t* prims, uint32_t* escapes) {
  *prims = 0;
  *escapes = 0;
  if (rec.empty()) return;
  struct Item { uint32_t node; const float* box; };
  std::vector<Item> todo{{0u, nullptr}};
  while (!todo.empty()) {
    const Item it = todo.back();
    todo.pop_back();
    for (int k = 0; k < width; ++k) {
      const rtbc::Pair& p = rec[(size_t)it.node * (width / 2) + k / 2];
      const uint32_t d = (k & 1) ? p.d1 : p.d0;
      const float* box = (k & 1) ? p.box1 : p.box0;
      if (d == rtbc::kEmptyQuadSlot) continue;
      if (it.box)
        for (int a = 0; a < 3; ++a)
          if (box[2 * a] < it.box[2 * a] - 1e-5f * (1.f + std::fabs(it.box[2 * a])) ||
              box[2 * a + 1] > it.box[2 * a + 1] + 1e-5f * (1.f + std::fabs(it.box[2 * a + 1])))
            ++*escapes;
      if (d & rtbc::kLeafBit) *prims += ((d >> 24) & 31u) + 1u;
      else todo.push_back({d, box});
    }
  }
}

int rtb_tree_info(const RtbCompiledTree* t, uint32_t info[8]) {
  if (!t || !info) return Fail(RTB_EINVAL, "null argument");
  uint64_t prims2 = 0, prims4 = 0;
  uint32_t esc2 = 0, esc4 = 0;
  WalkTree(t->tree.pairs, 2, &prims2, &esc2);
  WalkTree(t->tree.quads, 4, &prims4, &esc4);
  info[0] = (uint32_t)t->tree.pairs.size();
  info[1] = t->tree.depth;
  info[2] = (uint32_t)prims2;
  info[3] = esc2;
  info[4] = (uint32_t)(t->tree.quads.size() / 2);
  info[5] = t->tree.quad_stack;
  info[6] = (uint32_t)prims4;
  info[7] = esc4;
  return RTB_OK;
}

uint32_t rtb_debug_fast_div(uint32_t n, uint32_t d) { return d ? rtb::fast_div(n, rtb::make_fast_div(d)) : 0u; }

int rtb_debug_check_failures(uint64_t* counts8) {
  if (!counts8) return Fail(RTB_EINVAL, "null argument");
  const bool selftest = counts8[7] == 0x5e1f7e57ull;  // the caller asks for one deliberate violation of class 7 (proves the counting path)
  std::memset(counts8, 0, 8 * sizeof(uint64_t));
#ifdef RTB_CHECKED
  int rc = EnsureInit();
  if (rc != RTB_OK) return rc;
  if (selftest) check_selftest_kernel<<<1, 32, 0, g_stream>>>();
  unsigned long long h[rtb::kChkClasses];
  RTB_CUDA(cudaDeviceSynchronize());
  RTB_CUDA(cudaMemcpyFromSymbol(h, rtb::g_check_fail, sizeof(h)));
  for (int k = 0; k < 8; ++k) counts8[k] = h[k];
  return 1;  // a checked build
#else
  (void)selftest;
  return RTB_OK;  // not a checked build: nothing is counted
#endif
}

int rtb_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) {
    cudaGetLastError();
    return 0;
  }
  return n;
}

int rtb_init(int device) {
  const int n = rtb_device_count();
  if (n <= 0) return Fail(RTB_ENODEVICE, "no CUDA device visible; this library has no CPU path");
  if (device < 0 || device >= n) return Fail(RTB_EINVAL, "device index out of range");
  RTB_CUDA(cudaSetDevice(device));
  if (g_stream && g_device == device) return RTB_OK;
  if (g_stream) rtb_shutdown();  // switching device: the wavefront pools belong to the old one
  RTB_CUDA(cudaStreamCreateWithFlags(&g_stream, cudaStreamNonBlocking));
  cudaDeviceProp prop;
  RTB_CUDA(cudaGetDeviceProperties(&prop, device));
  g_sm_count = prop.multiProcessorCount;
  g_device = device;
  {  // keep freed scene arenas in the device's default pool instead of returning them to the driver
    cudaMemPool_t pool = nullptr;
    if (cudaDeviceGetDefaultMemPool(&pool, device) == cudaSuccess) {
      unsigned long long keep = ~0ull;
      cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep);
    } else {
      cudaGetLastError();
    }
  }
  return RTB_OK;
}

int rtb_shutdown(void) {
  if (g_stream) {
    cudaStreamSynchronize(g_stream);
    FreeWaves(&g_ctx);
    FreeFused(&g_ctx);
    if (g_ctx.pixel_order) cudaFree(g_ctx.pixel_order);
    g_ctx.pixel_order = nullptr;
    g_ctx.order_w = g_ctx.order_h = 0;
    if (g_ctx.accum) cudaFree(g_ctx.accum);
    g_ctx.accum = nullptr;
    g_ctx.accum_bytes = 0;
    if (g_ctx.adapt) cudaFree(g_ctx.adapt);
    g_ctx.adapt = nullptr;
    g_ctx.adapt_bytes = 0;
    if (g_ctx.sched) cudaFree(g_ctx.sched);
    g_ctx.sched = nullptr;
    g_ctx.sched_bytes = 0;
    if (g_staging.host) cudaFreeHost(g_staging.host);
    g_staging = Staging{};
    cudaStreamDestroy(g_stream);
    g_stream = nullptr;
    g_device = -1;
  }
  return RTB_OK;
}

int rtb_scene_destroy(RtbScene* s) {
  if (!s) return RTB_OK;
  int prev = -1;
  cudaGetDevice(&prev);
  cudaSetDevice(s->device);
  for (auto t : s->textures)
    if (t) cudaDestroyTextureObject(t);
  for (auto a : s->arrays) cudaFreeArray(a);
  if (s->arena) {
    cudaDeviceSynchronize();  // work on caller streams (rtb_intersect_batch_device) may still read the arena
    if (g_stream && g_device == s->device) cudaFreeAsync(s->arena, g_stream);
    else cudaFree(s->arena);
  }
  if (prev >= 0 && prev != s->device) cudaSetDevice(prev);  // the caller's current device is not ours to change
  delete s;
  return RTB_OK;
}

int rtb_scene_create(const RtbSceneDesc* d, RtbScene** out) {
  if (!d || !out) return Fail(RTB_EINVAL, "null argument");
  *out = nullptr;
  const bool trace_create = std::getenv("RTB_TRACE_CREATE") != nullptr;
  const auto tc0 = std::chrono::steady_clock::now();
  auto lap = [&](const char* what) {
    if (trace_create) std::fprintf(stderr, "[rtb create] %-10s %.3f ms\n", what, std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - tc0).count());
  };
  // a compiled tree that has been uploaded before carries the scene's device image (pinned): its tables were
  // validated by rtb_tree_compile and must not have changed since, so neither the checks nor the shadows are redone
  ArenaImage* img = (d->compiled && !(d->flags & RTB_SCENE_DEVICE_BUILD)) ? &const_cast<RtbCompiledTree*>(d->compiled)->image : nullptr;
  if (std::getenv("RTB_NO_ARENA_IMAGE")) img = nullptr;
  const uint32_t counts_now[6] = {d->n_spheres, d->n_triangles, d->n_rects, d->n_materials, d->n_textures, d->n_images};
  const bool cached = img && img->filled && std::memcmp(img->counts, counts_now, sizeof(counts_now)) == 0;
  int rc = cached ? (d->abi_version == RTB_ABI_VERSION ? RTB_OK : Fail(RTB_EINVAL, "RtbSceneDesc.abi_version mismatch")) : ValidateDesc(d);
  if (rc != RTB_OK) return rc;
  rc = EnsureInit();
  if (rc != RTB_OK) return rc;
  lap("validate");

  RtbScene* s = new RtbScene();
  s->device = g_device;
  rtb::DevScene& v = s->dev;
  const bool device_build = (d->flags & RTB_SCENE_DEVICE_BUILD) != 0;
  const uint32_t n_prims = d->n_spheres + d->n_triangles + d->n_rects;
  // The flat loop (trace_loop, Io::kFlat) renders this scene when the reference's own tree is ONE leaf - its Bvh::Hit then tests every
  // primitive (bvh.h:84-116; Cornell: 9 primitives in a leaf the SAH refused to split, earth: 4) - or when there is no interchange tree
  // and the scene has a handful of primitives.
  s->flat_ok = n_prims > 0 && (device_build ? n_prims <= kFlatMaxPrims : (d->nodes && d->n_nodes >= 1 && d->nodes[0].b > 0 && n_prims <= RTB_MAX_LEAF));
  size_t built_pairs_off = 0;
#define PUT(expr)            \
  do {                       \
    rc = (expr);             \
    if (rc != RTB_OK) {      \
      rtb_scene_destroy(s);  \
      return rc;             \
    }                        \
  } while (0)
  // host temporaries queued by Put must outlive Commit
  rtbc::Compiled local;
  std::vector<float4> sf, tf, rf;
  std::vector<cudaTextureObject_t> objs;
  std::vector<int2> dims;
  {  // interchange tree -> device-native sibling pairs (rtb_compile.h), unless the caller already did it
    const rtbc::Compiled* tree = d->compiled ? &d->compiled->tree : &local;
    if (device_build) {
      tree = &local;  // stays empty: the pairs are written by the build kernels after the upload
    } else if (!d->compiled) {
      rc = CompileTree(*d, &local);
      if (rc != RTB_OK) {
        rtb_scene_destroy(s);
        return rc;
      }
    }
    PUT(s->Put(reinterpret_cast<const float4*>(tree->pairs.data()), 4 * tree->pairs.size(), &v.pairs));
    v.n_pairs = (uint32_t)tree->pairs.size();
    v.n_top_pairs = tree->top_pairs;
    PUT(s->Put(reinterpret_cast<const float4*>(tree->quads.data()), 4 * tree->quads.size(), &v.quads));
    v.n_quads = (uint32_t)(tree->quads.size() / 2);
    if (device_build && n_prims > 0) {
      v.n_pairs = n_prims > 1 ? n_prims - 1 : 1;
      built_pairs_off = s->Reserve((size_t)v.n_pairs * 64);
    }
  }
  PUT(s->Put(reinterpret_cast<const double2*>(d->spheres), 2 * (size_t)d->n_spheres, &v.spheres));
  PUT(s->Put(reinterpret_cast<const double2*>(d->triangles), 5 * (size_t)d->n_triangles, &v.triangles));
  PUT(s->Put(reinterpret_cast<const double2*>(d->rects), 3 * (size_t)d->n_rects, &v.rects));
  // float shadows of the double tables (shading frame + pre-filter): part of the pinned arena image where the scene has one;
  // otherwise (device-built trees, scenes above 256 MB) written by three small kernels after the upload - no host pass, no H2D
  const bool device_shadows = img == nullptr && !std::getenv("RTB_HOST_SHADOWS");
  size_t shadow_off[3] = {0, 0, 0};
  if (device_shadows) {
    shadow_off[0] = s->Reserve((size_t)d->n_spheres * 16);
    shadow_off[1] = s->Reserve((size_t)d->n_triangles * 48);
    shadow_off[2] = s->Reserve((size_t)d->n_rects * 32);
  } else {
    sf.resize(cached ? 0 : d->n_spheres);
    for (uint32_t i = 0; i < (cached ? 0u : d->n_spheres); ++i)
      sf[i] = make_float4((float)d->spheres[i].c[0], (float)d->spheres[i].c[1], (float)d->spheres[i].c[2], (float)d->spheres[i].r);
    PUT(s->Put(cached ? nullptr : sf.data(), (size_t)d->n_spheres, &v.sphere_f));
    tf.resize(cached ? 0 : 3 * (size_t)d->n_triangles);
    for (uint32_t i = 0; i < (cached ? 0u : d->n_triangles); ++i) {
      const RtbTriangle& t = d->triangles[i];
      float4 a = make_float4((float)t.a[0], (float)t.a[1], (float)t.a[2], 0.f);
      std::memcpy(&a.w, &t.mat, 4);
      tf[3 * (size_t)i] = a;
      // .w = max |component| rounded up: the filters' error bounds (rtb_device.cuh triangle_filter)
      auto mx = [](const double* e) { return rtbc::Up(std::max(std::fabs(e[0]), std::max(std::fabs(e[1]), std::fabs(e[2])))); };
      tf[3 * (size_t)i + 1] = make_float4((float)t.e1[0], (float)t.e1[1], (float)t.e1[2], mx(t.e1));
      tf[3 * (size_t)i + 2] = make_float4((float)t.e2[0], (float)t.e2[1], (float)t.e2[2], mx(t.e2));
    }
    PUT(s->Put(cached ? nullptr : tf.data(), 3 * (size_t)d->n_triangles, &v.triangle_f));
    rf.resize(cached ? 0 : 2 * (size_t)d->n_rects);
    for (uint32_t i = 0; i < (cached ? 0u : d->n_rects); ++i) {
      const RtbRect& r = d->rects[i];
      rf[2 * (size_t)i] = make_float4((float)r.k, (float)r.a0, (float)r.a1, (float)r.b0);
      float4 q = make_float4((float)r.b1, 0.f, 0.f, 0.f);
      std::memcpy(&q.y, &r.axis, 4);
      std::memcpy(&q.z, &r.mat, 4);
      rf[2 * (size_t)i + 1] = q;
    }
    PUT(s->Put(cached ? nullptr : rf.data(), 2 * (size_t)d->n_rects, &v.rect_f));
  }
  PUT(s->Put(d->sphere_mat, d->n_spheres, &v.sphere_mat));
  PUT(s->Put(d->sphere_gid, d->n_spheres, &v.sphere_gid));
  PUT(s->Put(d->triangle_gid, d->n_triangles, &v.triangle_gid));
  PUT(s->Put(d->rect_gid, d->n_rects, &v.rect_gid));
  // no slot ranks (device-built tree: creation order breaks exact-t ties): the rank tables ARE the id tables, uploaded once
  if (d->sphere_rank || img) PUT(s->Put(d->sphere_rank ? d->sphere_rank : d->sphere_gid, d->n_spheres, &v.sphere_rank));
  if (d->triangle_rank || img) PUT(s->Put(d->triangle_rank ? d->triangle_rank : d->triangle_gid, d->n_triangles, &v.triangle_rank));
  if (d->rect_rank || img) PUT(s->Put(d->rect_rank ? d->rect_rank : d->rect_gid, d->n_rects, &v.rect_rank));
  PUT(s->Put(d->materials, d->n_materials, &v.materials));
  PUT(s->Put(d->textures, d->n_textures, &v.textures));
  {  // images -> point-sampled texture objects over RGBA8 arrays
    objs.assign(d->n_images, 0);
    dims.assign(d->n_images, make_int2(0, 0));
    for (uint32_t i = 0; i < d->n_images; ++i) {
      const RtbImage& im = d->images[i];
      if (im.width <= 0 || im.height <= 0 || !im.rgb) continue;
      std::vector<uchar4> rgba((size_t)im.width * im.height);
      for (size_t p = 0; p < rgba.size(); ++p) rgba[p] = make_uchar4(im.rgb[3 * p], im.rgb[3 * p + 1], im.rgb[3 * p + 2], 255);
      cudaChannelFormatDesc fmt = cudaCreateChannelDesc<uchar4>();
      cudaArray_t arr = nullptr;
      cudaError_t e = cudaMallocArray(&arr, &fmt, im.width, im.height);
      if (e == cudaSuccess) {
        s->arrays.push_back(arr);
        e = cudaMemcpy2DToArray(arr, 0, 0, rgba.data(), (size_t)im.width * 4, (size_t)im.width * 4, im.height, cudaMemcpyHostToDevice);
      }
      cudaTextureObject_t tex = 0;
      if (e == cudaSuccess) {
        cudaResourceDesc res{};
        res.resType = cudaResourceTypeArray;
        res.res.array.array = arr;
        cudaTextureDesc td{};
        td.addressMode[0] = td.addressMode[1] = cudaAddressModeClamp;
        td.filterMode = cudaFilterModePoint;
        td.readMode = cudaReadModeElementType;
        td.normalizedCoords = 0;
        e = cudaCreateTextureObject(&tex, &res, &td, nullptr);
      }
      if (e != cudaSuccess) {
        rtb_scene_destroy(s);
        return Fail(RTB_ECUDA, std::string("image texture upload: ") + cudaGetErrorString(e));
      }
      s->textures.push_back(tex);
      objs[i] = tex;
      dims[i] = make_int2(im.width, im.height);
    }
    PUT(s->Put(objs.data(), objs.size(), &v.image_tex, /*fresh=*/true));  // texture object handles belong to THIS scene
    PUT(s->Put(dims.data(), dims.size(), &v.image_dim, /*fresh=*/true));
  }
  v.n_spheres = d->n_spheres; v.n_triangles = d->n_triangles; v.n_rects = d->n_rects;
  v.n_materials = d->n_materials; v.n_textures = d->n_textures;
  s->smem_bytes = (size_t)v.n_top_pairs * 64;  // + the stack columns, added at launch (TraceSmem)
  if (s->smem_bytes + kStackSmem > 200 * 1024) {
    rtb_scene_destroy(s);
    return Fail(RTB_EINVAL, "n_top_nodes does not fit shared memory");
  }
  lap("tables");
  PUT(s->Commit(g_stream, img, counts_now));
  lap("commit");
  if (!cached && ValidatesPrimitivesOnDevice(d) && n_prims > 0) {
    unsigned bad = 0;
    cudaError_t ve = cudaMemsetAsync(s->cursor, 0, 4, g_stream);  // the work cursor doubles as the error counter before the first launch
    if (ve == cudaSuccess) {
      rtb::validate_materials_kernel<<<(n_prims + 255) / 256, 256, 0, g_stream>>>(v, d->n_spheres, d->n_triangles, d->n_rects, d->n_materials, s->cursor);
      ve = cudaMemcpyAsync(&bad, s->cursor, 4, cudaMemcpyDeviceToHost, g_stream);
    }
    if (ve == cudaSuccess) ve = cudaStreamSynchronize(g_stream);
    if (ve == cudaSuccess) ve = cudaMemsetAsync(s->cursor, 0, 4, g_stream);
    if (ve != cudaSuccess || bad) {
      rtb_scene_destroy(s);
      return ve != cudaSuccess ? Fail(RTB_ECUDA, std::string("scene validation kernel: ") + cudaGetErrorString(ve))
                               : Fail(RTB_EINVAL, std::to_string(bad) + " primitives with a material index (or rect axis) out of range");
    }
    lap("validate2");
  }
  if (!d->sphere_rank && !img) v.sphere_rank = v.sphere_gid;
  if (!d->triangle_rank && !img) v.triangle_rank = v.triangle_gid;
  if (!d->rect_rank && !img) v.rect_rank = v.rect_gid;
  if (device_shadows) {
    auto* sph = reinterpret_cast<float4*>(s->arena + shadow_off[0]);
    auto* tri = reinterpret_cast<float4*>(s->arena + shadow_off[1]);
    auto* rec = reinterpret_cast<float4*>(s->arena + shadow_off[2]);
    if (d->n_spheres) rtb::shadow_spheres_kernel<<<(d->n_spheres + 255) / 256, 256, 0, g_stream>>>(v.spheres, d->n_spheres, sph);
    if (d->n_triangles) rtb::shadow_triangles_kernel<<<(d->n_triangles + 255) / 256, 256, 0, g_stream>>>(v.triangles, d->n_triangles, tri);
    if (d->n_rects) rtb::shadow_rects_kernel<<<(d->n_rects + 255) / 256, 256, 0, g_stream>>>(v.rects, d->n_rects, rec);
    v.sphere_f = d->n_spheres ? sph : nullptr;
    v.triangle_f = d->n_triangles ? tri : nullptr;
    v.rect_f = d->n_rects ? rec : nullptr;
    cudaError_t se = cudaGetLastError();
    if (se == cudaSuccess && !(device_build && n_prims > 0)) se = cudaStreamSynchronize(g_stream);  // (the tree build below synchronises anyway)
    if (se != cudaSuccess) {
      rtb_scene_destroy(s);
      return Fail(RTB_ECUDA, std::string("float shadow kernels: ") + cudaGetErrorString(se));
    }
    lap("shadows");
  }
  if (device_build && n_prims > 0) {
    v.pairs = reinterpret_cast<const float4*>(s->arena + built_pairs_off);
    PUT(BuildTreeOnDevice(s, d->n_spheres, d->n_triangles, n_prims, reinterpret_cast<float*>(s->arena + built_pairs_off)));
    lap("build");
  }
  // opt in to large dynamic shared memory for the two traversal kernels
  {
    const int with_top = (int)(s->smem_bytes + kStackSmem), without = (int)kStackSmem;
    if (with_top > 48 * 1024) {
      cudaFuncSetAttribute(rtb::intersect_batch_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, with_top);
      cudaFuncSetAttribute(rtb::extend_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, with_top);
      cudaFuncSetAttribute(rtb::path_kernel<true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, with_top);
      cudaFuncSetAttribute(rtb::path_kernel<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, with_top);
      cudaFuncSetAttribute(rtb::path_kernel<true, false, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, with_top);
      cudaFuncSetAttribute(rtb::path_kernel<true, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, with_top);
    }
    if (without > 48 * 1024) {
      cudaFuncSetAttribute(rtb::intersect_batch_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, without);
      cudaFuncSetAttribute(rtb::extend_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, without);
      cudaFuncSetAttribute(rtb::path_kernel<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, without);
      cudaFuncSetAttribute(rtb::path_kernel<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, without);
      cudaFuncSetAttribute(rtb::path_kernel<false, false, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, without);
      cudaFuncSetAttribute(rtb::path_kernel<false, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, without);
    }
  }
#undef PUT
  *out = s;
  return RTB_OK;
}

}  // extern "C"

// structural validation: every index the kernels will follow must be in range
static int ValidateDesc(const RtbSceneDesc* d) {
  if (d->abi_version != RTB_ABI_VERSION) return Fail(RTB_EINVAL, "RtbSceneDesc.abi_version mismatch");
  const bool device_build = (d->flags & RTB_SCENE_DEVICE_BUILD) != 0;
  if (device_build) {
    if (d->n_spheres > (1u << 24) || d->n_triangles > (1u << 24) || d->n_rects > (1u << 24)) return Fail(RTB_EINVAL, "more than 2^24 primitives of one type");
  } else if (d->n_nodes < 2 || (d->n_nodes & 1u) || !d->nodes) {
    return Fail(RTB_EINVAL, "scene needs an even, non-zero node count");
  }
  if (d->n_prim_refs >= (1u << 25)) return Fail(RTB_EINVAL, "more than 2^25 leaf entries");
  if (d->n_top_nodes > d->n_nodes || (d->n_top_nodes & 1u)) return Fail(RTB_EINVAL, "bad n_top_nodes");
  for (uint32_t i = 0; i < (device_build ? 0u : d->n_nodes); ++i) {
    const RtbNode& n = d->nodes[i];
    if (n.b == 0) {
      if (!(n.lo[0] <= n.hi[0])) continue;  // padding slot (inverted box), never entered
      if (n.a + 1 >= d->n_nodes || (n.a & 1u) || n.a == 0) return Fail(RTB_EINVAL, "node child index out of range");
    } else if (n.b > RTB_MAX_LEAF || (uint64_t)n.a + n.b > d->n_prim_refs) {
      return Fail(RTB_EINVAL, "leaf range out of bounds");
    }
  }
  for (uint32_t i = 0; i < (device_build ? 0u : d->n_prim_refs); ++i) {
    const uint32_t ty = RTB_REF_TYPE(d->prim_refs[i]), idx = RTB_REF_INDEX(d->prim_refs[i]);
    const uint32_t lim = ty == RTB_PRIM_SPHERE ? d->n_spheres : ty == RTB_PRIM_TRIANGLE ? d->n_triangles : ty == RTB_PRIM_RECT ? d->n_rects : d->n_nodes;
    if (idx >= lim || (ty == RTB_PRIM_SUBTREE && (idx & 1u))) return Fail(RTB_EINVAL, "leaf entry out of range");
  }
  auto mat_ok = [&](uint32_t m) { return m < d->n_materials; };
  if (!ValidatesPrimitivesOnDevice(d)) {  // large scenes: the same checks run as a kernel after the upload (rtb_scene_create)
    for (uint32_t i = 0; i < d->n_spheres; ++i) if (!mat_ok(d->sphere_mat[i])) return Fail(RTB_EINVAL, "sphere material out of range");
    for (uint32_t i = 0; i < d->n_triangles; ++i) if (!mat_ok(d->triangles[i].mat)) return Fail(RTB_EINVAL, "triangle material out of range");
    for (uint32_t i = 0; i < d->n_rects; ++i) if (!mat_ok(d->rects[i].mat) || d->rects[i].axis < 0 || d->rects[i].axis > 2) return Fail(RTB_EINVAL, "bad rect");
  }
  for (uint32_t i = 0; i < d->n_materials; ++i) {
    const RtbMaterial& m = d->materials[i];
    if ((m.type == RTB_MAT_LAMBERTIAN || m.type == RTB_MAT_DIFFUSE_LIGHT) && (m.tex < 0 || (uint32_t)m.tex >= d->n_textures))
      return Fail(RTB_EINVAL, "material texture out of range");
  }
  for (uint32_t i = 0; i < d->n_textures; ++i) {
    const RtbTexture& t = d->textures[i];
    if (t.type == RTB_TEX_CHECKER && (t.even < 0 || t.odd < 0 || (uint32_t)t.even >= d->n_textures || (uint32_t)t.odd >= d->n_textures))
      return Fail(RTB_EINVAL, "checker child out of range");
    if (t.type == RTB_TEX_IMAGE && (t.image < 0 || (uint32_t)t.image >= d->n_images)) return Fail(RTB_EINVAL, "image out of range");
  }
  return RTB_OK;
}

// Traversal CTAs are large so that one shared-memory copy of the top of the tree serves many warps
// and most of the 228 KB stays L1 for the rest of the tree.
using rtb::kTraceBlock;

// Persistent grid: as many CTAs as are co-resident on the whole chip.
template <class K>
static int PersistentGrid(K kernel, int block, size_t smem) {
  int per_sm = 0;
  cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, block, smem);
  if (per_sm < 1) per_sm = 1;
  return per_sm * g_sm_count;
}

extern "C" {

int rtb_intersect_batch_device(RtbScene* s, const RtbRay* d_rays, size_t n, RtbHit* d_hits, void* stream) {
  if (!s || (!d_rays && n) || (!d_hits && n)) return Fail(RTB_EINVAL, "null argument");
  if (n == 0) return RTB_OK;
  if (n > 0xffffff00ull) return Fail(RTB_EINVAL, "batch too large");
  int rc = EnsureInit();
  if (rc != RTB_OK) return rc;
  RTB_TRY(CheckSceneDevice(s));
  cudaStream_t st = stream ? static_cast<cudaStream_t>(stream) : g_stream;
  RTB_CUDA(cudaMemsetAsync(s->cursor, 0, 8, st));
  const bool use_smem = s->dev.n_top_pairs > 0 && std::getenv("RTB_SMEM_TOP") != nullptr;
  const float4* rays4 = reinterpret_cast<const float4*>(d_rays);
  if (use_smem) {
    const size_t smem = s->smem_bytes + kStackSmem;
    const int grid = PersistentGrid(rtb::intersect_batch_kernel<true>, kTraceBlock, smem);
    rtb::intersect_batch_kernel<true><<<grid, kTraceBlock, smem, st>>>(s->dev, rays4, (uint32_t)n, d_hits, s->cursor, s->counters);
  } else if ((s->flat_ok && !std::getenv("RTB_TREE_BATCH")) ||
             (std::getenv("RTB_FLAT_BATCH") && s->dev.n_spheres + s->dev.n_triangles + s->dev.n_rects <= 4096u)) {  // no tree: every primitive
    const int grid = PersistentGrid(rtb::intersect_batch_kernel<false, false, true>, kTraceBlock, kStackSmem);
    rtb::intersect_batch_kernel<false, false, true><<<grid, kTraceBlock, kStackSmem, st>>>(s->dev, rays4, (uint32_t)n, d_hits, s->cursor, s->counters);
  } else if (s->dev.quads && std::getenv("RTB_WIDE4_BATCH")) {  // the 4-wide tree (an experiment: DESIGN.md 3.9)
    const int grid = PersistentGrid(rtb::intersect_batch_kernel<false, true>, kTraceBlock, kStackSmem);
    rtb::intersect_batch_kernel<false, true><<<grid, kTraceBlock, kStackSmem, st>>>(s->dev, rays4, (uint32_t)n, d_hits, s->cursor, s->counters);
  } else {
    const int grid = PersistentGrid(rtb::intersect_batch_kernel<false>, kTraceBlock, kStackSmem);
    rtb::intersect_batch_kernel<false><<<grid, kTraceBlock, kStackSmem, st>>>(s->dev, rays4, (uint32_t)n, d_hits, s->cursor, s->counters);
  }
  // second pass: the reference's exact t for filter-settled hits, and the global primitive ids
  rtb::exact_t_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(s->dev, rays4, (uint32_t)n, d_hits, s->counters);
  RTB_CUDA(cudaGetLastError());
  return RTB_OK;
}

// ---- rtb_intersect_batch_static: seam 2 for callers that must not allocate ------------------------------------------------
// The scratch of the closest-hit kernels - the persistent threads' work cursor and the device copies of the caller's host
// buffers - lives in ONE file-scope device array instead of cudaMalloc'ed memory.  Every GPU that loads the library holds
// these 3 MiB + 256 bytes until the process exits.  Layout: [cursor, 256 bytes][RtbRay x MAX][RtbHit x MAX].
constexpr size_t kStaticBatchMax = RTB_STATIC_BATCH_MAX;
constexpr size_t kStaticRaysOff = 256, kStaticHitsOff = kStaticRaysOff + kStaticBatchMax * sizeof(RtbRay);
constexpr size_t kStaticBatchBytes = kStaticHitsOff + kStaticBatchMax * sizeof(RtbHit);
static __device__ __align__(256) unsigned char g_static_batch[kStaticBatchBytes];

// In host code `g_static_batch` is only a stand-in for the symbol: the device address is asked for once per device, with that
// device current (each device has its own copy of the array), and cached.
constexpr int kMaxStaticDevices = 64;
static std::mutex g_static_mu;
static unsigned char* g_static_ptr[kMaxStaticDevices] = {};
static int StaticBatchScratch(int device, unsigned char** out) {
  if (device < 0 || device >= kMaxStaticDevices) return Fail(RTB_EINVAL, "rtb_intersect_batch_static: device index beyond 63");
  std::lock_guard<std::mutex> lock(g_static_mu);
  if (!g_static_ptr[device]) {
    int current = -1;
    RTB_CUDA(cudaGetDevice(&current));
    if (current != device) return Fail(RTB_EINVAL, "rtb_intersect_batch_static: the scene's device is not the current device");
    void* p = nullptr;
    RTB_CUDA(cudaGetSymbolAddress(&p, g_static_batch));
    if (!p) return Fail(RTB_ECUDA, "rtb_intersect_batch_static: cudaGetSymbolAddress returned no address");
    g_static_ptr[device] = static_cast<unsigned char*>(p);
  }
  *out = g_static_ptr[device];
  return RTB_OK;
}

int rtb_intersect_batch_static(RtbScene* s, const RtbRay* rays, size_t n, RtbHit* hits) {
  if (!s || (!rays && n) || (!hits && n)) return Fail(RTB_EINVAL, "null argument");
  if (n > kStaticBatchMax)
    return Fail(RTB_EINVAL, "rtb_intersect_batch_static: " + std::to_string(n) + " rays exceed RTB_STATIC_BATCH_MAX (" + std::to_string(kStaticBatchMax) +
                                "): split the batch or use rtb_intersect_batch");
  if (n == 0) return RTB_OK;
  int rc = EnsureInit();
  if (rc != RTB_OK) return rc;
  RTB_TRY(CheckSceneDevice(s));
  unsigned char* scratch = nullptr;
  RTB_TRY(StaticBatchScratch(s->device, &scratch));
  unsigned int* cursor = reinterpret_cast<unsigned int*>(scratch);
  RtbRay* d_rays = reinterpret_cast<RtbRay*>(scratch + kStaticRaysOff);
  RtbHit* d_hits = reinterpret_cast<RtbHit*>(scratch + kStaticHitsOff);
  cudaStream_t st = g_stream;
  // the array keeps the last call's contents: the cursor must start at zero; rays [0, n) are overwritten, hits [0, n) are all written
  RTB_CUDA(cudaMemsetAsync(cursor, 0, 8, st));
  RTB_CUDA(cudaMemcpyAsync(d_rays, rays, n * sizeof(RtbRay), cudaMemcpyHostToDevice, st));
  const float4* rays4 = reinterpret_cast<const float4*>(d_rays);
  const int grid = PersistentGrid(rtb::intersect_batch_kernel<false>, kTraceBlock, kStackSmem);
  rtb::intersect_batch_kernel<false><<<grid, kTraceBlock, kStackSmem, st>>>(s->dev, rays4, (uint32_t)n, d_hits, cursor, s->counters);
  rtb::exact_t_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(s->dev, rays4, (uint32_t)n, d_hits, s->counters);
  RTB_CUDA(cudaGetLastError());
  RTB_CUDA(cudaMemcpyAsync(hits, d_hits, n * sizeof(RtbHit), cudaMemcpyDeviceToHost, st));
  RTB_CUDA(cudaStreamSynchronize(st));
  return RTB_OK;
}

int rtb_intersect_batch(RtbScene* s, const RtbRay* rays, size_t n, RtbHit* hits) {
  if (!s || (!rays && n) || (!hits && n)) return Fail(RTB_EINVAL, "null argument");
  if (n == 0) return RTB_OK;
  int rc = EnsureInit();
  if (rc != RTB_OK) return rc;
  RtbRay* d_rays = nullptr;
  RtbHit* d_hits = nullptr;
  RTB_CUDA(cudaMalloc(reinterpret_cast<void**>(&d_rays), n * sizeof(RtbRay)));
  cudaError_t e = cudaMalloc(reinterpret_cast<void**>(&d_hits), n * sizeof(RtbHit));
  if (e == cudaSuccess) e = cudaMemcpyAsync(d_rays, rays, n * sizeof(RtbRay), cudaMemcpyHostToDevice, g_stream);
  if (e == cudaSuccess) {
    rc = rtb_intersect_batch_device(s, d_rays, n, d_hits, g_stream);
    if (rc == RTB_OK) e = cudaMemcpyAsync(hits, d_hits, n * sizeof(RtbHit), cudaMemcpyDeviceToHost, g_stream);
    if (rc == RTB_OK && e == cudaSuccess) e = cudaStreamSynchronize(g_stream);
  }
  cudaFree(d_rays);
  cudaFree(d_hits);
  if (rc != RTB_OK) return rc;
  if (e != cudaSuccess) return Fail(RTB_ECUDA, std::string("intersect_batch: ") + cudaGetErrorString(e));
  return RTB_OK;
}

static int EnsurePool(RenderCtx* s, uint32_t pool, size_t n_waves) {
  if (s->pool != pool || s->waves.size() < n_waves) {
    FreeWaves(s);
    s->waves.resize(n_waves);
    for (auto& wv : s->waves) {
      auto alloc = [&](void** p, size_t bytes) -> cudaError_t {
        cudaError_t e = cudaMalloc(p, bytes);
        if (e == cudaSuccess) wv.mem.push_back(*p);
        return e;
      };
      for (int k = 0; k < 2; ++k) {
        RTB_CUDA(alloc(reinterpret_cast<void**>(&wv.q[k].ro), (size_t)pool * 16));
        RTB_CUDA(alloc(reinterpret_cast<void**>(&wv.q[k].rd), (size_t)pool * 16));
        RTB_CUDA(alloc(reinterpret_cast<void**>(&wv.q[k].thr), (size_t)pool * 16));
        RTB_CUDA(alloc(reinterpret_cast<void**>(&wv.q[k].meta), (size_t)pool * 8));
      }
      RTB_CUDA(alloc(reinterpret_cast<void**>(&wv.hit_t), (size_t)pool * 4));
      RTB_CUDA(alloc(reinterpret_cast<void**>(&wv.hit_ref), (size_t)pool * 4));
      RTB_CUDA(alloc(reinterpret_cast<void**>(&wv.counts), 32));  // see Wave::counts
      RTB_CUDA(cudaMallocHost(reinterpret_cast<void**>(&wv.host_count), 64));
      RTB_CUDA(cudaStreamCreateWithFlags(&wv.stream, cudaStreamNonBlocking));
      RTB_CUDA(cudaEventCreateWithFlags(&wv.done, cudaEventDisableTiming));
    }
    s->pool = pool;
  }
  return RTB_OK;
}

static int EnsureLight(Wave& wv);
static int EnsureFused(RenderCtx* s) { return EnsureLight(s->fused); }
static int EnsureLight(Wave& wv) {
  if (wv.stream) return RTB_OK;
  void* p = nullptr;
  RTB_CUDA(cudaMalloc(&p, 32));
  wv.mem.push_back(p);
  wv.counts = static_cast<unsigned int*>(p);
  RTB_CUDA(cudaStreamCreateWithFlags(&wv.stream, cudaStreamNonBlocking));
  RTB_CUDA(cudaEventCreateWithFlags(&wv.done, cudaEventDisableTiming));
  return RTB_OK;
}

static int EnsureOrder(RenderCtx* s, int w, int h) {
  if (s->order_w != w || s->order_h != h) {
    // 8x4-tiled pixel permutation: consecutive samples of a warp are neighbouring pixels
    std::vector<uint32_t> order;
    order.reserve((size_t)w * h);
    for (int ty = 0; ty < h; ty += 4)
      for (int tx = 0; tx < w; tx += 8)
        for (int y = ty; y < std::min(ty + 4, h); ++y)
          for (int x = tx; x < std::min(tx + 8, w); ++x) order.push_back((uint32_t)(y * w + x));
    void* p = nullptr;
    if (s->pixel_order) cudaFree(s->pixel_order);
    s->pixel_order = nullptr;
    RTB_CUDA(cudaMalloc(&p, order.size() * 4));
    RTB_CUDA(cudaMemcpy(p, order.data(), order.size() * 4, cudaMemcpyHostToDevice));
    s->pixel_order = static_cast<uint32_t*>(p);
    s->order_w = w;
    s->order_h = h;
  }
  return RTB_OK;
}

}  // extern "C"

// Both render entries.  `ad` != nullptr: adaptive per-pixel stop, path kernel with one work item per pixel.
static int RenderImpl(RtbScene* s, const RtbCamera* cam, const RtbRenderParams* prm, float* d_rgb_sum, const RtbAdaptive* ad, float* d_sumsq,
                      uint32_t* d_count, RtbStats* stats) {
  if (!s || !cam || !prm || !d_rgb_sum) return Fail(RTB_EINVAL, "null argument");
  if (ad) {
    if (!d_sumsq || !d_count) return Fail(RTB_EINVAL, "adaptive render needs the sumsq and count planes");
    if (prm->flags & (RTB_RENDER_WAVEFRONT | RTB_RENDER_SERIAL)) return Fail(RTB_EINVAL, "adaptive render runs on the path kernel only");
    if (ad->criterion != RTB_ADAPT_PIXEL_STATE && ad->criterion != RTB_ADAPT_LUMINANCE) return Fail(RTB_EINVAL, "unknown adaptive criterion");
    if (!(ad->rel_threshold >= 0.0f) || ad->min_samples < 1) return Fail(RTB_EINVAL, "bad adaptive parameters");
  }
  if (cam->width <= 0 || cam->height <= 0 || prm->spp_end < prm->spp_begin || prm->max_depth < 0) return Fail(RTB_EINVAL, "bad camera / params");
  if (prm->mode != RTB_MODE_A && prm->mode != RTB_MODE_B) return Fail(RTB_EINVAL, "unknown estimator mode");
  int rc = EnsureInit();
  if (rc != RTB_OK) return rc;
  RTB_TRY(CheckSceneDevice(s));
  const uint32_t n_pixels = (uint32_t)cam->width * (uint32_t)cam->height;
  uint32_t pool = prm->pool_size ? prm->pool_size : (1u << 21);
  pool = std::min(std::max(pool, 1u << 10), 1u << 28);
  if ((unsigned long long)cam->width * (unsigned long long)cam->height > (1ull << 30)) return Fail(RTB_EINVAL, "image larger than 2^30 pixels");
  // RTB_RENDER_SERIAL: one wave, every kernel timed alone with events (the roofline measurement);
  // otherwise several waves in flight and only the whole render is timed.
  const bool serial = (prm->flags & RTB_RENDER_SERIAL) != 0;
  size_t n_waves = serial ? 1 : 4;
  if (!serial)
    if (const char* e = std::getenv("RTB_WAVES")) n_waves = (size_t)std::min(8, std::max(1, std::atoi(e)));
  const unsigned long long total = (unsigned long long)n_pixels * (unsigned long long)(prm->spp_end - prm->spp_begin);
  while (n_waves > 1 && total < (unsigned long long)pool * n_waves) --n_waves;  // small renders: fewer waves
  RenderCtx* ctx = &g_ctx;
  const bool fused = !(prm->flags & (RTB_RENDER_WAVEFRONT | RTB_RENDER_SERIAL));
  rc = fused ? EnsureFused(ctx) : EnsurePool(ctx, pool, n_waves);
  if (rc == RTB_OK) rc = EnsureOrder(ctx, cam->width, cam->height);
  // adaptive: up to four pixel groups (contiguous runs of the tiled pixel order), each with its own stream
  // (measured: no gain over one group on bunny C3 - the groups reach their tails together - so one group unless asked)
  uint32_t n_groups = ad ? 1u : 0u;
  if (const char* e = std::getenv("RTB_ADAPT_GROUPS")) n_groups = ad ? (uint32_t)std::min(4, std::max(1, std::atoi(e))) : 0u;
  for (uint32_t g = 0; g < n_groups && rc == RTB_OK; ++g) rc = EnsureLight(ctx->groups[g]);
  // Cost-ordered schedule of the persistent path kernel (RTB_RENDER_COST_ORDER; MEASURED, OFF BY DEFAULT).  A launch ends on the paths
  // that happen to be long and start late: on bunny C3 0.75 ms of an 8.8 ms step, and most of the step once the spp are split over
  // 8 GPUs (a 20-bounce path running alone takes ~0.8 ms).  Here the first samples of every pixel also record the length of their
  // paths, the tiles are sorted by it, and the rest of the render starts with the tiles that grow long paths; while the measuring
  // launch drains, a second stream renders the next samples in the plain order.  B200 (profiles/r02_cost_ordered_schedule.txt):
  // the ordering itself is worth 4 % at 100 spp and 15 % at 12 spp on the bunny (one launch in the order a PREVIOUS render measured:
  // 8.81 -> 8.46 ms, 1.66 -> 1.41 ms; nothing on RTIOW / Cornell), but measuring inside the render costs more than that - the ordered
  // launch has to wait for the measuring launch's own tail: 9.06 ms and 2.16 ms.  Long paths are rare events spread over many tiles;
  // a tile's history only predicts them partly.
  const uint32_t spp_all_ = (uint32_t)(prm->spp_end - prm->spp_begin);
  const bool cost_order = fused && !ad && spp_all_ >= 8 && n_pixels >= 4096 && !(prm->flags & (RTB_RENDER_SMEM_TOP | RTB_RENDER_POOL_PATHS)) &&
                          !std::getenv("RTB_SMEM_TOP") && !std::getenv("RTB_POOL_PATHS") &&
                          ((prm->flags & RTB_RENDER_COST_ORDER) || std::getenv("RTB_COST_ORDER")) && (prm->max_depth > 4 || prm->mode == RTB_MODE_B);
  if (cost_order && rc == RTB_OK) rc = EnsureLight(ctx->groups[0]);
  if (rc != RTB_OK) return rc;
  cudaStream_t st = g_stream;
  std::vector<Wave*> waves;  // the first n_waves of the context (it may hold more from an earlier call)
  if (ad)
    for (uint32_t g = 0; g < n_groups; ++g) waves.push_back(&ctx->groups[g]);
  else if (fused) {
    waves.push_back(&ctx->fused);
    if (cost_order) waves.push_back(&ctx->groups[0]);
  } else
    for (size_t k = 0; k < n_waves; ++k) waves.push_back(&ctx->waves[k]);
  rtb::RenderConsts rcst{};
  rcst.cam = *cam;
  rcst.round_keys = rtb::philox_key_schedule((uint32_t)prm->seed, (uint32_t)(prm->seed >> 32));
  rcst.max_depth = prm->max_depth;
  rcst.mode = prm->mode;
  rcst.spp_begin = prm->spp_begin;
  rcst.n_pixels = n_pixels;
  rcst.div_width = rtb::make_fast_div((uint32_t)cam->width);
  rcst.div_pixels = rtb::make_fast_div(n_pixels);
  rcst.div_count = rtb::make_fast_div(1u);
  // Work-item order of the persistent path kernel (rtb_kernels.cuh, generate_sample_tiled): tile-major keeps the GPU on one small
  // image region at a time, which pays when the scene does not fit the L1s (bunny C3 -3.7 %, stress C5 -3 % and 4x less DRAM
  // traffic) and costs 2-4 % when it does (Cornell, earth: 99 % L1 hit rate either way).  RTB_WORK_ORDER=tile|sample overrides.
  bool tile_major = s->arena_bytes > (1u << 20);
  if (const char* e = std::getenv("RTB_WORK_ORDER")) tile_major = e[0] == 't';
  rcst.tile_major = tile_major ? 1 : 0;
  // Per-scene schedule of the traversal loop (PathIo::kTune, rtb_kernels.cuh): speculative leaf postponement for trees of spheres only,
  // a later service phase for scenes that do not fit the L2, no tree at all for a handful of primitives.  RTB_TUNE=0|1|2|3 overrides.
  int tune = 0;
  if (s->dev.n_triangles == 0 && s->dev.n_rects == 0 && s->dev.n_spheres >= 64) tune = 1;
  else if (tile_major && s->arena_bytes > (256u << 20)) tune = 2;
  // scenes of a handful of primitives: the flat loop (trace_loop, Io::kFlat) instead of the tree - Cornell, earth
  const uint32_t n_prims_all = s->dev.n_spheres + s->dev.n_triangles + s->dev.n_rects;
  if (s->flat_ok) tune = 3;
  if (const char* e = std::getenv("RTB_TUNE")) tune = std::atoi(e);
  if (tune == 2 && !tile_major) tune = 0;
  if ((tune == 3 || tune == 11) && (n_prims_all == 0 || n_prims_all > 4096u)) tune = 0;  // RTB_TUNE=3 on a large scene would run for minutes
  if (tune == 3 || tune == 11) tile_major = false;
  rcst.tile_major = tile_major ? 1 : 0;
  // RTB_RENDER_WIDE4: the 4-wide tree (a scene created with RTB_WIDE4=1 carries it); RTB_TUNE=12 runs it at 3 CTAs x 80 registers
  if ((prm->flags & RTB_RENDER_WIDE4) || tune == 4 || tune == 12) {
    if (!s->dev.quads) return Fail(RTB_EINVAL, "RTB_RENDER_WIDE4: the scene has no 4-wide tree (create it with RTB_WIDE4=1; host-compiled trees only)");
    if (tune != 12) tune = 4;
  }
  rcst.spp_max = prm->spp_end - prm->spp_begin;
  rcst.min_samples = ad ? ad->min_samples : 0;
  rcst.criterion = ad ? ad->criterion : 0;
  rcst.rel_threshold = ad ? ad->rel_threshold : 0.0f;
  const rtb::AdaptivePlanes planes{d_sumsq, d_count, nullptr, nullptr, nullptr, 0u, 0u};
  const bool use_smem = s->dev.n_top_pairs > 0 && ((prm->flags & RTB_RENDER_SMEM_TOP) || std::getenv("RTB_SMEM_TOP") != nullptr);
  const size_t ext_smem = (use_smem ? s->smem_bytes : 0) + kStackSmem;
  const int ext_grid = use_smem ? PersistentGrid(rtb::extend_kernel<true>, kTraceBlock, ext_smem) : PersistentGrid(rtb::extend_kernel<false>, kTraceBlock, ext_smem);
  const bool trace_iters = std::getenv("RTB_TRACE_ITERS") != nullptr;
  // path kernel (rtb_kernels.cuh): the whole render (default) or, in wavefront mode, the drain of a wavefront's last paths
  uint32_t drain_max = (prm->flags & RTB_RENDER_NO_DRAIN) ? 0u : pool / 8;
  if (const char* e = std::getenv("RTB_DRAIN_MAX")) drain_max = (uint32_t)std::max(0, std::atoi(e));
  const int path_grid = use_smem ? PersistentGrid(fused ? rtb::path_kernel<true, false> : rtb::path_kernel<true, true>, kTraceBlock, ext_smem)
                                 : PersistentGrid(fused ? rtb::path_kernel<false, false> : rtb::path_kernel<false, true>, kTraceBlock, ext_smem);

  const int flat_grid = tune == 3    ? PersistentGrid(rtb::path_kernel<false, false, false, false, false, 3>, kTraceBlock, ext_smem)
                        : tune == 11 ? PersistentGrid(rtb::path_kernel<false, false, false, false, false, 11>, kTraceBlock, ext_smem)
                                     : 0;
  const int wide3_grid = tune == 12 ? PersistentGrid(rtb::path_kernel<false, false, false, true, false, 12>, kTraceBlock, ext_smem) : 0;

  rtb::LoopMismatch* dbg_mismatch = nullptr;
  if (!fused && std::getenv("RTB_DEBUG_FLAT_VS_TREE") && n_prims_all <= 4096u) {
    RTB_CUDA(cudaMalloc(reinterpret_cast<void**>(&dbg_mismatch), sizeof(rtb::LoopMismatch)));
    RTB_CUDA(cudaMemset(dbg_mismatch, 0, sizeof(rtb::LoopMismatch)));
  }
  Events events;  // destroyed on every return path
  // on request: the pool kernel (64 paths per warp in shared memory, rtb_pool.cuh) instead of one path per lane
  const bool use_pool = fused && !use_smem && ((prm->flags & RTB_RENDER_POOL_PATHS) || std::getenv("RTB_POOL_PATHS"));
  const int pool_grid = use_pool ? PersistentGrid(rtb::pool_kernel<false, false>, kTraceBlock, 0) : 0;
  const int pool_grid_ad = use_pool && ad ? PersistentGrid(rtb::pool_kernel<false, true>, kTraceBlock, 0) : 0;
  cudaEvent_t ev0, ev1, ev_ready;
  RTB_TRY(events.Create(&ev0));
  RTB_TRY(events.Create(&ev1));
  RTB_TRY(events.Create(&ev_ready, cudaEventDisableTiming));
  RTB_CUDA(cudaEventRecord(ev0, st));
  if (!(prm->flags & RTB_RENDER_KEEP_ACCUM)) {
    RTB_CUDA(cudaMemsetAsync(d_rgb_sum, 0, (size_t)n_pixels * 3 * sizeof(float), st));
    if (ad) {
      RTB_CUDA(cudaMemsetAsync(d_sumsq, 0, (size_t)n_pixels * 3 * sizeof(float), st));
      RTB_CUDA(cudaMemsetAsync(d_count, 0, (size_t)n_pixels * sizeof(uint32_t), st));
    }
  }
#ifdef RTB_COUNTERS
  RTB_CUDA(cudaMemsetAsync(s->counters, 0, sizeof(rtb::HitCounters), st));
#endif
  RTB_CUDA(cudaEventRecord(ev_ready, st));
  for (Wave* wp : waves) {
    Wave& w = *wp;
    RTB_CUDA(cudaStreamWaitEvent(w.stream, ev_ready, 0));
    w.live = 0;
    w.cur = 0;
    w.pending = false;
    RTB_CUDA(cudaMemsetAsync(w.counts, 0, 32, w.stream));
  }
  unsigned long long generated = 0, rays = 0;
  uint32_t iterations = 0, launches = 0, drain_paths = 0;
  float ms_gen = 0.f, ms_ext = 0.f, ms_shd = 0.f, ms_drain = 0.f;
  std::vector<std::pair<cudaEvent_t, cudaEvent_t>> drain_events;  // serial mode: around each drain launch
  cudaEvent_t ek[4] = {nullptr, nullptr, nullptr, nullptr};  // serial mode: before generate / extend / shade, after shade
  if (serial)
    for (auto& e : ek) RTB_TRY(events.Create(&e));
  // max_depth == 0: GetPixel returns black without tracing (camera.h:149-151)
  const bool trace = prm->max_depth > 0 || prm->mode == RTB_MODE_B || cam->depth_camera;
  bool any = trace && !fused;
  if (trace && fused && ad && total > 0) {
    // passes of R speculative samples per active pixel + an in-order replay of the stop rule (rtb_kernels.cuh);
    // the active count stays on the device, so the passes are enqueued back to back without a host round trip
    // Chunk size: every pass ends in a kernel tail (the deepest paths running alone, ~0.4-0.8 ms on bunny C3), every chunk wastes
    // up to R - 1 samples per pixel.  Measured on bunny C3, max 100 (profiles/r01_traversal_probes.txt): R = 8 / 16 / 28 / 40 / 64 ->
    // 9.8 / 7.8 / 7.05 / 7.1 / 7.6 ms with estimator B, 14.4 / 10.4 / 8.65 / 8.25 / 8.25 with A.  The passes are then made equal
    // (100 -> 34 + 34 + 32 rather than 40 + 40 + 20): a short last pass pays a whole tail for a few samples.
    uint32_t R = 40;
    if (const char* e = std::getenv("RTB_ADAPT_CHUNK")) R = (uint32_t)std::min(64, std::max(1, std::atoi(e)));
    const uint32_t spp_max = (uint32_t)rcst.spp_max;
    while (R > 16 && (size_t)n_pixels * R * 12 > ((size_t)2 << 30)) R -= 8;  // the chunk buffer stays below 2 GB (4K: R = 16)
    R = std::min(R, spp_max);
    if (!std::getenv("RTB_ADAPT_UNEVEN")) {
      const uint32_t n_pass = (spp_max + R - 1) / R;
      R = (spp_max + n_pass - 1) / n_pass;
    }
    const size_t list_bytes = ((size_t)n_pixels * 4 + 255) & ~(size_t)255, chunk_bytes = (size_t)n_pixels * R * 12;
    const size_t need = 2 * list_bytes + 256 + chunk_bytes;
    if (ctx->adapt_bytes < need) {
      if (ctx->adapt) cudaFree(ctx->adapt);
      ctx->adapt = nullptr;
      ctx->adapt_bytes = 0;
      RTB_CUDA(cudaMalloc(reinterpret_cast<void**>(&ctx->adapt), need));
      ctx->adapt_bytes = need;
    }
    uint32_t* list[2] = {reinterpret_cast<uint32_t*>(ctx->adapt), reinterpret_cast<uint32_t*>(ctx->adapt + list_bytes)};
    unsigned* n_act = reinterpret_cast<unsigned*>(ctx->adapt + 2 * list_bytes);  // [group][2]
    float* chunk = reinterpret_cast<float*>(ctx->adapt + 2 * list_bytes + 256);
    const int grid = use_smem ? PersistentGrid(rtb::path_kernel<true, false, true>, kTraceBlock, ext_smem)
                              : PersistentGrid(rtb::path_kernel<false, false, true>, kTraceBlock, ext_smem);
    for (uint32_t pass = 0, base = 0; base < spp_max; ++pass, base += R) {
      for (uint32_t g = 0; g < n_groups; ++g) {  // round-robin: group g's pass overlaps the tails of the others
        Wave& w = *waves[g];
        const uint32_t p0 = (uint32_t)((unsigned long long)n_pixels * g / n_groups), p1 = (uint32_t)((unsigned long long)n_pixels * (g + 1) / n_groups);
        const uint32_t n_g = p1 - p0;
        unsigned* cnt = n_act + 2 * g;
        if (pass == 0) RTB_CUDA(cudaMemcpyAsync(cnt, &n_g, 4, cudaMemcpyHostToDevice, w.stream));
        auto* rc64 = reinterpret_cast<unsigned long long*>(w.counts + 2);
        rtb::AdaptivePlanes pl = planes;
        pl.active = pass == 0 ? ctx->pixel_order + p0 : list[pass & 1] + p0;
        pl.n_active = cnt + (pass & 1);
        pl.chunk = chunk + (size_t)p0 * R * 3;
        pl.chunk_r = std::min(R, spp_max - base);
        pl.base = base;
        RTB_CUDA(cudaMemsetAsync(w.counts, 0, 4, w.stream));
        RTB_CUDA(cudaMemsetAsync(cnt + ((pass + 1) & 1), 0, 4, w.stream));
        if (use_pool)
          rtb::pool_kernel<false, true><<<pool_grid_ad, kTraceBlock, 0, w.stream>>>(s->dev, rtb::PathQueue{}, 0u, 0u, 0u, rcst, ctx->pixel_order, d_rgb_sum,
                                                                                w.counts, rc64, s->counters, pl);
        else if (use_smem)
          rtb::path_kernel<true, false, true><<<grid, kTraceBlock, ext_smem, w.stream>>>(s->dev, rtb::PathQueue{}, 0u, 0u, 0u, rcst, ctx->pixel_order,
                                                                                     d_rgb_sum, w.counts, rc64, s->counters, pl);
        else
          (tune == 3 ? rtb::path_kernel<false, false, true, false, false, 3> : rtb::path_kernel<false, false, true>)<<<grid, kTraceBlock, ext_smem, w.stream>>>(s->dev, rtb::PathQueue{}, 0u, 0u, 0u, rcst, ctx->pixel_order,
                                                                                      d_rgb_sum, w.counts, rc64, s->counters, pl);
        rtb::adaptive_replay_kernel<<<(n_g + 255) / 256, 256, 0, w.stream>>>(rcst, d_rgb_sum, pl, list[(pass + 1) & 1] + p0, cnt + ((pass + 1) & 1));
        launches += 2;
      }
      ++iterations;
    }
  } else if (trace && fused) {
    // tile-major: a launch covers samples [sa, sa + cnt) of every pixel (< 2^30 items); sample-major: the same range, sample index outermost
    const uint32_t spp_all = (uint32_t)(prm->spp_end - prm->spp_begin);
    const uint32_t spp_per_launch = std::max(1u, (1u << 30) / n_pixels);
    // samples [sa, sb) of every pixel of `order` on wave w; cost != nullptr: the launch also records path lengths per pixel
    auto launch_range = [&](Wave& w, uint32_t sa, uint32_t sb, const uint32_t* order, uint32_t* cost) -> int {
      while (sa < sb) {
        const uint32_t cnt = std::min(spp_per_launch, sb - sa);
        const uint32_t chunk = n_pixels * cnt, s0 = sa, p0 = tile_major ? cnt : 0u;
        rcst.div_count = rtb::make_fast_div(cnt);  // kernel parameters are copied at launch
        RTB_CUDA(cudaMemsetAsync(w.counts, 0, 4, w.stream));
        auto* rc64 = reinterpret_cast<unsigned long long*>(w.counts + 2);
        rtb::AdaptivePlanes pl = planes;
        pl.count = cost;
        if (use_pool)
          rtb::pool_kernel<false, false><<<pool_grid, kTraceBlock, 0, w.stream>>>(s->dev, rtb::PathQueue{}, chunk, s0, p0, rcst, order, d_rgb_sum, w.counts, rc64,
                                                                                s->counters, pl);
        else if (use_smem && tile_major)
          rtb::path_kernel<true, false, false, true><<<path_grid, kTraceBlock, ext_smem, w.stream>>>(s->dev, rtb::PathQueue{}, chunk, s0, p0, rcst, order, d_rgb_sum,
                                                                                                 w.counts, rc64, s->counters, pl);
        else if (use_smem)
          rtb::path_kernel<true, false><<<path_grid, kTraceBlock, ext_smem, w.stream>>>(s->dev, rtb::PathQueue{}, chunk, s0, p0, rcst, order, d_rgb_sum, w.counts,
                                                                                    rc64, s->counters, pl);
        else if (cost && tile_major)
          rtb::path_kernel<false, false, false, true, true><<<path_grid, kTraceBlock, ext_smem, w.stream>>>(s->dev, rtb::PathQueue{}, chunk, s0, p0, rcst, order,
                                                                                                        d_rgb_sum, w.counts, rc64, s->counters, pl);
        else if (cost)
          rtb::path_kernel<false, false, false, false, true><<<path_grid, kTraceBlock, ext_smem, w.stream>>>(s->dev, rtb::PathQueue{}, chunk, s0, p0, rcst, order,
                                                                                                         d_rgb_sum, w.counts, rc64, s->counters, pl);
        else if (tune == 4 && !use_smem && tile_major)
          rtb::path_kernel<false, false, false, true, false, 4><<<path_grid, kTraceBlock, ext_smem, w.stream>>>(s->dev, rtb::PathQueue{}, chunk, s0, p0, rcst, order,
                                                                                                            d_rgb_sum, w.counts, rc64, s->counters, pl);
        else if (tune == 4 && !use_smem)
          rtb::path_kernel<false, false, false, false, false, 4><<<path_grid, kTraceBlock, ext_smem, w.stream>>>(s->dev, rtb::PathQueue{}, chunk, s0, p0, rcst, order,
                                                                                                             d_rgb_sum, w.counts, rc64, s->counters, pl);
        else if (tune == 12 && !use_smem && tile_major)
          rtb::path_kernel<false, false, false, true, false, 12><<<wide3_grid, kTraceBlock, ext_smem, w.stream>>>(s->dev, rtb::PathQueue{}, chunk, s0, p0, rcst, order,
                                                                                                              d_rgb_sum, w.counts, rc64, s->counters, pl);
        else if (tune == 12 && !use_smem)
          rtb::path_kernel<false, false, false, false, false, 12><<<wide3_grid, kTraceBlock, ext_smem, w.stream>>>(s->dev, rtb::PathQueue{}, chunk, s0, p0, rcst, order,
                                                                                                               d_rgb_sum, w.counts, rc64, s->counters, pl);
        else if (tune == 3 && !use_smem)
          rtb::path_kernel<false, false, false, false, false, 3><<<flat_grid, kTraceBlock, ext_smem, w.stream>>>(s->dev, rtb::PathQueue{}, chunk, s0, p0, rcst, order,
                                                                                                             d_rgb_sum, w.counts, rc64, s->counters, pl);
        else if (tune == 11 && !use_smem)
          rtb::path_kernel<false, false, false, false, false, 11><<<flat_grid, kTraceBlock, ext_smem, w.stream>>>(s->dev, rtb::PathQueue{}, chunk, s0, p0, rcst, order,
                                                                                                              d_rgb_sum, w.counts, rc64, s->counters, pl);
        else if (tune == 2 && tile_major)
          rtb::path_kernel<false, false, false, true, false, 2><<<path_grid, kTraceBlock, ext_smem, w.stream>>>(s->dev, rtb::PathQueue{}, chunk, s0, p0, rcst, order,
                                                                                                            d_rgb_sum, w.counts, rc64, s->counters, pl);
        else if (tune == 1 && tile_major)
          rtb::path_kernel<false, false, false, true, false, 1><<<path_grid, kTraceBlock, ext_smem, w.stream>>>(s->dev, rtb::PathQueue{}, chunk, s0, p0, rcst, order,
                                                                                                            d_rgb_sum, w.counts, rc64, s->counters, pl);
        else if (tune == 1)
          rtb::path_kernel<false, false, false, false, false, 1><<<path_grid, kTraceBlock, ext_smem, w.stream>>>(s->dev, rtb::PathQueue{}, chunk, s0, p0, rcst, order,
                                                                                                             d_rgb_sum, w.counts, rc64, s->counters, pl);
        else if (tile_major)
          rtb::path_kernel<false, false, false, true><<<path_grid, kTraceBlock, ext_smem, w.stream>>>(s->dev, rtb::PathQueue{}, chunk, s0, p0, rcst, order, d_rgb_sum,
                                                                                                  w.counts, rc64, s->counters, pl);
        else
          rtb::path_kernel<false, false><<<path_grid, kTraceBlock, ext_smem, w.stream>>>(s->dev, rtb::PathQueue{}, chunk, s0, p0, rcst, order, d_rgb_sum, w.counts,
                                                                                     rc64, s->counters, pl);
        sa += cnt;
        generated += chunk;
        ++launches;
        ++iterations;
      }
      return RTB_OK;
    };
    if (!cost_order) {
      RTB_TRY(launch_range(*waves[0], 0u, spp_all, ctx->pixel_order, nullptr));
    } else {
      Wave &wa = *waves[0], &wb = *waves[1];
      const uint32_t n_groups32 = n_pixels / 32u;  // complete groups of 32 entries of pixel_order (the < 32 left-over pixels keep their place)
      const uint32_t a = spp_all >= 32 ? 2u : 1u, b0 = std::max(1u, spp_all / 10u);
      auto up = [](size_t b) { return (b + 255) & ~(size_t)255; };
      size_t cub_bytes = 0;
      RTB_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, cub_bytes, (const uint32_t*)nullptr, (uint32_t*)nullptr, (const uint32_t*)nullptr, (uint32_t*)nullptr,
                                               (int)n_groups32, 0, 24, wa.stream));
      const size_t o_cost = 0, o_order = o_cost + up((size_t)n_pixels * 4), o_k0 = o_order + up((size_t)n_pixels * 4), o_k1 = o_k0 + up((size_t)n_groups32 * 4),
                   o_v0 = o_k1 + up((size_t)n_groups32 * 4), o_v1 = o_v0 + up((size_t)n_groups32 * 4), o_cub = o_v1 + up((size_t)n_groups32 * 4),
                   need = o_cub + up(cub_bytes);
      if (ctx->sched_bytes < need) {
        if (ctx->sched) cudaFree(ctx->sched);
        ctx->sched = nullptr;
        ctx->sched_bytes = 0;
        RTB_CUDA(cudaMalloc(reinterpret_cast<void**>(&ctx->sched), need));
        ctx->sched_bytes = need;
      }
      auto* cost = reinterpret_cast<uint32_t*>(ctx->sched + o_cost);
      auto* order2 = reinterpret_cast<uint32_t*>(ctx->sched + o_order);
      auto* k0 = reinterpret_cast<uint32_t*>(ctx->sched + o_k0);
      auto* k1 = reinterpret_cast<uint32_t*>(ctx->sched + o_k1);
      auto* v0 = reinterpret_cast<uint32_t*>(ctx->sched + o_v0);
      auto* v1 = reinterpret_cast<uint32_t*>(ctx->sched + o_v1);
      if (std::getenv("RTB_COST_ORDER_REUSE") && ctx->sched_pixels == n_pixels) {
        // experiment: the whole render in ONE launch in the order the previous render of this image measured - what the ordering
        // itself is worth, without the cost of measuring
        RTB_TRY(launch_range(wa, 0u, spp_all, order2, nullptr));
        goto cost_order_done;
      }
      RTB_CUDA(cudaMemsetAsync(cost, 0, (size_t)n_pixels * 4, wa.stream));
      RTB_TRY(launch_range(wa, 0u, a, ctx->pixel_order, cost));              // measures while it renders
      RTB_TRY(launch_range(wb, a, a + b0, ctx->pixel_order, nullptr));      // fills the GPU while the measuring launch drains
      rtb::group_cost_kernel<<<(n_groups32 + 7) / 8, 256, 0, wa.stream>>>(ctx->pixel_order, cost, n_groups32, k0, v0);
      RTB_CUDA(cub::DeviceRadixSort::SortPairs(ctx->sched + o_cub, cub_bytes, k0, k1, v0, v1, (int)n_groups32, 0, 24, wa.stream));
      rtb::permute_groups_kernel<<<(n_pixels + 255) / 256, 256, 0, wa.stream>>>(ctx->pixel_order, v1, n_groups32, n_pixels, order2);
      launches += 3;
      RTB_TRY(launch_range(wa, a + b0, spp_all, order2, nullptr));          // long-path tiles first
      ctx->sched_pixels = n_pixels;
    cost_order_done:;
    }
  }
  while (any) {
    any = false;
    for (Wave* wp : waves) {
    Wave& w = *wp;
      if (w.pending) {  // collect the previous iteration of this wave: its survivor count
        RTB_CUDA(cudaEventSynchronize(w.done));
        w.pending = false;
        if (serial) {
          float a = 0.f, b = 0.f, c = 0.f;
          cudaEventElapsedTime(&a, ek[0], ek[1]);
          cudaEventElapsedTime(&b, ek[1], ek[2]);
          cudaEventElapsedTime(&c, ek[2], ek[3]);
          ms_gen += a;
          ms_ext += b;
          ms_shd += c;
          if (trace_iters) std::fprintf(stderr, "[rtb iter %u] rays %u gen %.3f ext %.3f shade %.3f ms -> live %u\n", iterations, w.live, a, b, c, *w.host_count);
        }
        w.live = *w.host_count;
        w.cur ^= 1;
        if (generated == total && w.live > 0 && w.live <= drain_max) {
          // nothing left to generate and few paths alive: finish them in ONE launch of the path kernel instead of
          // one latency-bound extend + shade pair per remaining bounce
          cudaStream_t ws = w.stream;
          cudaEvent_t d0 = nullptr, d1 = nullptr;
          if (serial) {
            RTB_TRY(events.Create(&d0));
            RTB_TRY(events.Create(&d1));
            drain_events.emplace_back(d0, d1);
          }
          RTB_CUDA(cudaMemsetAsync(w.counts, 0, 4, ws));
          if (serial) RTB_CUDA(cudaEventRecord(d0, ws));
          auto* rc64 = reinterpret_cast<unsigned long long*>(w.counts + 2);
          if (use_smem)
            rtb::path_kernel<true, true><<<path_grid, kTraceBlock, ext_smem, ws>>>(s->dev, w.q[w.cur], w.live, 0u, 0u, rcst, ctx->pixel_order, d_rgb_sum,
                                                                               w.counts, rc64, s->counters, planes);
          else
            (tune == 3 ? rtb::path_kernel<false, true, false, false, false, 3> : rtb::path_kernel<false, true>)<<<path_grid, kTraceBlock, ext_smem, ws>>>(s->dev, w.q[w.cur], w.live, 0u, 0u, rcst, ctx->pixel_order, d_rgb_sum,
                                                                                w.counts, rc64, s->counters, planes);
          if (serial) RTB_CUDA(cudaEventRecord(d1, ws));
          if (trace_iters) std::fprintf(stderr, "[rtb drain] %u paths\n", w.live);
          drain_paths += w.live;
          w.live = 0;
          ++launches;
        }
      }
      const uint32_t room = pool - w.live;
      const uint32_t gen = (uint32_t)std::min<unsigned long long>(room, total - generated);
      if (gen == 0 && w.live == 0) continue;  // this wave is finished
      cudaStream_t ws = w.stream;
      if (serial) RTB_CUDA(cudaEventRecord(ek[0], ws));
      if (gen) {
        rtb::generate_kernel<<<(gen + 255) / 256, 256, 0, ws>>>(w.q[w.cur], w.live, gen, (uint32_t)(generated / n_pixels), (uint32_t)(generated % n_pixels), rcst,
                                                                ctx->pixel_order);
        generated += gen;
        w.live += gen;
        ++launches;
      }
      RTB_CUDA(cudaMemsetAsync(w.counts, 0, 8, ws));
      if (serial) RTB_CUDA(cudaEventRecord(ek[1], ws));
      if (use_smem)
        rtb::extend_kernel<true><<<ext_grid, kTraceBlock, ext_smem, ws>>>(s->dev, w.q[w.cur], w.live, w.hit_t, w.hit_ref, w.counts, s->counters);
      else if (tune == 3)
        rtb::extend_kernel<false, true><<<ext_grid, kTraceBlock, ext_smem, ws>>>(s->dev, w.q[w.cur], w.live, w.hit_t, w.hit_ref, w.counts, s->counters);
      else
        rtb::extend_kernel<false><<<ext_grid, kTraceBlock, ext_smem, ws>>>(s->dev, w.q[w.cur], w.live, w.hit_t, w.hit_ref, w.counts, s->counters);
      if (dbg_mismatch && !use_smem) {  // RTB_DEBUG_FLAT_VS_TREE: the OTHER loop on the same queue, hits compared (debugging aid, allocates)
        float* t_b = nullptr;
        uint32_t* ref_b = nullptr;
        RTB_CUDA(cudaMallocAsync(reinterpret_cast<void**>(&t_b), (size_t)w.live * 4, ws));
        RTB_CUDA(cudaMallocAsync(reinterpret_cast<void**>(&ref_b), (size_t)w.live * 4, ws));
        RTB_CUDA(cudaMemsetAsync(w.counts, 0, 4, ws));
        if (tune == 3)
          rtb::extend_kernel<false><<<ext_grid, kTraceBlock, ext_smem, ws>>>(s->dev, w.q[w.cur], w.live, t_b, ref_b, w.counts, s->counters);
        else
          rtb::extend_kernel<false, true><<<ext_grid, kTraceBlock, ext_smem, ws>>>(s->dev, w.q[w.cur], w.live, t_b, ref_b, w.counts, s->counters);
        rtb::compare_hits_kernel<<<(w.live + 255) / 256, 256, 0, ws>>>(w.q[w.cur], w.live, w.hit_t, w.hit_ref, t_b, ref_b, dbg_mismatch);
        RTB_CUDA(cudaFreeAsync(t_b, ws));
        RTB_CUDA(cudaFreeAsync(ref_b, ws));
      }
      if (serial) RTB_CUDA(cudaEventRecord(ek[2], ws));
      rtb::shade_kernel<<<(w.live + 255) / 256, 256, 0, ws>>>(s->dev, w.q[w.cur], w.live, w.hit_t, w.hit_ref, w.q[w.cur ^ 1], w.counts + 1, d_rgb_sum, rcst);
      if (serial) RTB_CUDA(cudaEventRecord(ek[3], ws));
      launches += 2;
      rays += w.live;
      ++iterations;
      RTB_CUDA(cudaMemcpyAsync(w.host_count, w.counts + 1, 4, cudaMemcpyDeviceToHost, ws));
      RTB_CUDA(cudaEventRecord(w.done, ws));
      w.pending = true;
      any = true;
    }
  }
  for (Wave* wp : waves) {
    Wave& w = *wp;  // the library stream joins every wave
    RTB_CUDA(cudaEventRecord(w.done, w.stream));
    RTB_CUDA(cudaStreamWaitEvent(st, w.done, 0));
  }
  RTB_CUDA(cudaEventRecord(ev1, st));
  RTB_CUDA(cudaEventSynchronize(ev1));
  RTB_CUDA(cudaGetLastError());
  if (dbg_mismatch) {  // one JSON line per disagreeing ray on stderr
    rtb::LoopMismatch h;
    RTB_CUDA(cudaDeviceSynchronize());
    RTB_CUDA(cudaMemcpy(&h, dbg_mismatch, sizeof(h), cudaMemcpyDeviceToHost));
    cudaFree(dbg_mismatch);
    std::fprintf(stderr, "{\"flat_vs_tree_mismatches\": %llu, \"rays\": %llu}\n", h.count, (unsigned long long)rays);
    for (unsigned k = 0; k < 16 && k < h.count; ++k)
      std::fprintf(stderr, "{\"mismatch_ray\": [%.9g, %.9g, %.9g, %.9g, %.9g, %.9g, %.9g, %.9g], \"ref_main\": %u, \"ref_other\": %u}\n", h.ro[k].x, h.ro[k].y,
                   h.ro[k].z, h.ro[k].w, h.rd[k].x, h.rd[k].y, h.rd[k].z, h.rd[k].w, h.ref_a[k], h.ref_b[k]);
  }
  for (Wave* wp : waves) {  // rays traced inside the path kernel (it bounces on the device)
    unsigned long long c[2] = {0, 0};
    RTB_CUDA(cudaMemcpy(c, wp->counts + 2, 16, cudaMemcpyDeviceToHost));
    rays += c[0];
    if (ad) generated += c[1];
  }
  for (auto& de : drain_events) {
    float d = 0.f;
    cudaEventElapsedTime(&d, de.first, de.second);
    ms_drain += d;
  }
  float ms = 0.f;
  cudaEventElapsedTime(&ms, ev0, ev1);
  if (stats) {
    std::memset(stats, 0, sizeof(*stats));
    stats->samples = generated;
    stats->rays = rays;
    stats->iterations = iterations;
    stats->launches = launches;
    stats->ms_total = ms;
    stats->ms_generate = ms_gen;
    stats->ms_extend = ms_ext;
    stats->ms_shade = ms_shd;
    stats->ms_drain = ms_drain;
    stats->drain_paths = drain_paths;
#ifdef RTB_COUNTERS
    rtb::HitCounters hc{};
    cudaMemcpy(&hc, s->counters, sizeof(hc), cudaMemcpyDeviceToHost);
    stats->node_visits = 2 * hc.node_pairs;
    stats->sphere_tests = hc.sphere_tests;
    stats->triangle_tests = hc.triangle_tests;
    stats->rect_tests = hc.rect_tests;
    if (std::getenv("RTB_DEBUG_COUNTERS"))
      std::fprintf(stderr, "[rtb counters] rays %llu pairs %llu exact_evals %llu filter_violations %llu max_steps %llu slow_rays %llu\n", hc.rays,
                   hc.node_pairs, hc.exact_evals, hc.filter_violations, hc.max_steps, hc.slow_rays);
    if (std::getenv("RTB_DEBUG_COUNTERS")) {
      std::fprintf(stderr, "[rtb counters] rays by log2(steps+1):");
      for (int k = 0; k < 16; ++k) std::fprintf(stderr, " %llu", hc.step_hist[k]);
      std::fprintf(stderr, "\n[rtb counters] a slow ray: o %.9g %.9g %.9g d %.9g %.9g %.9g\n", hc.slow_ray[0], hc.slow_ray[1], hc.slow_ray[2],
                   hc.slow_ray[4], hc.slow_ray[5], hc.slow_ray[6]);
    }
#endif
  }
  return RTB_OK;
}

extern "C" {

int rtb_render_device(RtbScene* s, const RtbCamera* cam, const RtbRenderParams* prm, float* d_rgb_sum, RtbStats* stats) {
  return RenderImpl(s, cam, prm, d_rgb_sum, nullptr, nullptr, nullptr, stats);
}

int rtb_render_adaptive_device(RtbScene* s, const RtbCamera* cam, const RtbRenderParams* prm, const RtbAdaptive* ad, float* d_rgb_sum,
                               float* d_rgb_sumsq, uint32_t* d_count, RtbStats* stats) {
  if (!ad) return Fail(RTB_EINVAL, "null argument");
  return RenderImpl(s, cam, prm, d_rgb_sum, ad, d_rgb_sumsq, d_count, stats);
}

int rtb_render(RtbScene* s, const RtbCamera* cam, const RtbRenderParams* prm, float* rgb_sum, RtbStats* stats) {
  if (!s || !cam || !prm || !rgb_sum) return Fail(RTB_EINVAL, "null argument");
  int rc = EnsureInit();
  if (rc != RTB_OK) return rc;
  const size_t bytes = (size_t)cam->width * cam->height * 3 * sizeof(float);
  if (g_ctx.accum_bytes < bytes) {
    if (g_ctx.accum) cudaFree(g_ctx.accum);
    g_ctx.accum = nullptr;
    g_ctx.accum_bytes = 0;
    RTB_CUDA(cudaMalloc(reinterpret_cast<void**>(&g_ctx.accum), bytes));
    g_ctx.accum_bytes = bytes;
  }
  float* d = g_ctx.accum;
  if (prm->flags & RTB_RENDER_KEEP_ACCUM) RTB_CUDA(cudaMemcpy(d, rgb_sum, bytes, cudaMemcpyHostToDevice));
  rc = rtb_render_device(s, cam, prm, d, stats);
  if (rc != RTB_OK) return rc;
  RTB_CUDA(cudaMemcpy(rgb_sum, d, bytes, cudaMemcpyDeviceToHost));
  return RTB_OK;
}

// ---- seam 1 on several devices of one box -----------------------------------------------------------------------
namespace {
struct MultiWorker {
  int device = 0;
  int rc = RTB_OK;
  std::string error;
  float* accum = nullptr;  // the worker thread's grow-only accumulator on `device` (its RenderCtx), valid until its next job
  RtbStats stats{};
};

// One device's share: scene, accumulator, render of [b, e).  Runs on the host thread that owns the device's state
// (the library's state is per thread = per device).
void MultiRenderOne(MultiWorker* w, const RtbSceneDesc* desc, const RtbCamera* cam, RtbRenderParams prm) {
  auto fail = [&](int rc) {
    w->rc = rc;
    w->error = rtb_last_error();
  };
  int rc = rtb_init(w->device);
  if (rc != RTB_OK) return fail(rc);
  RtbScene* scene = nullptr;
  rc = rtb_scene_create(desc, &scene);
  if (rc != RTB_OK) return fail(rc);
  const size_t bytes = (size_t)cam->width * cam->height * 3 * sizeof(float);
  if (g_ctx.accum_bytes < bytes) {
    if (g_ctx.accum) cudaFree(g_ctx.accum);
    g_ctx.accum = nullptr;
    g_ctx.accum_bytes = 0;
    if (cudaMalloc(reinterpret_cast<void**>(&g_ctx.accum), bytes) != cudaSuccess) {
      rtb_scene_destroy(scene);
      w->rc = Fail(RTB_ENOMEM, "rtb_render_multi: accumulator allocation failed");
      w->error = rtb_last_error();
      return;
    }
    g_ctx.accum_bytes = bytes;
  }
  w->accum = g_ctx.accum;
  prm.flags &= ~RTB_RENDER_KEEP_ACCUM;
  rc = rtb_render_device(scene, cam, &prm, w->accum, &w->stats);
  if (rc != RTB_OK) fail(rc);
  rtb_scene_destroy(scene);
}

// Worker threads live as long as the process: one per device ever used as a peer, each keeping its stream, pools,
// pixel order and accumulator between calls (creating them costs ~3 ms per device per call otherwise).
struct DeviceThread {
  std::mutex m;
  std::condition_variable cv;
  std::function<void()> job;
  bool has_job = false, done = false;
  void Loop() {
    for (;;) {
      std::unique_lock<std::mutex> lk(m);
      cv.wait(lk, [&] { return has_job; });
      std::function<void()> f = std::move(job);
      has_job = false;
      lk.unlock();
      f();
      lk.lock();
      done = true;
      cv.notify_all();
    }
  }
  void Submit(std::function<void()> f) {
    std::lock_guard<std::mutex> lk(m);
    job = std::move(f);
    done = false;
    has_job = true;
    cv.notify_all();
  }
  void Wait() {
    std::unique_lock<std::mutex> lk(m);
    cv.wait(lk, [&] { return done; });
  }
};
std::mutex g_pool_mu;
std::map<int, DeviceThread*> g_pool;
DeviceThread* WorkerFor(int device) {
  std::lock_guard<std::mutex> lk(g_pool_mu);
  auto it = g_pool.find(device);
  if (it != g_pool.end()) return it->second;
  DeviceThread* t = new DeviceThread();
  std::thread([t] { t->Loop(); }).detach();
  g_pool[device] = t;
  return t;
}
}  // namespace

int rtb_render_multi(const RtbSceneDesc* desc, int n_devices, const int* devices, const RtbCamera* cam, const RtbRenderParams* prm, float* rgb_sum,
                     RtbStats* stats) {
  if (!desc || !cam || !prm || !rgb_sum || n_devices < 1 || n_devices > 9) return Fail(RTB_EINVAL, "bad argument");
  if (prm->spp_end < prm->spp_begin || cam->width <= 0 || cam->height <= 0) return Fail(RTB_EINVAL, "bad camera / params");
  const int visible = rtb_device_count();
  if (visible <= 0) return Fail(RTB_ENODEVICE, "no CUDA device visible; this library has no CPU path");
  std::vector<MultiWorker> work((size_t)n_devices);
  for (int k = 0; k < n_devices; ++k) {
    work[k].device = devices ? devices[k] : k;
    if (work[k].device < 0 || work[k].device >= visible) return Fail(RTB_EINVAL, "device index out of range");
    for (int j = 0; j < k; ++j)
      if (work[j].device == work[k].device) return Fail(RTB_EINVAL, "device listed twice");
  }
  // slices of the global sample indices, sizes differ by at most one (multigpu.spp_slice)
  const int total = prm->spp_end - prm->spp_begin, base = total / n_devices, extra = total % n_devices;
  std::vector<RtbRenderParams> prms((size_t)n_devices, *prm);
  for (int k = 0, b = prm->spp_begin; k < n_devices; ++k) {
    const int len = base + (k < extra ? 1 : 0);
    prms[k].spp_begin = b;
    prms[k].spp_end = b + len;
    b += len;
  }
  const auto t0 = std::chrono::steady_clock::now();
  std::vector<DeviceThread*> peers;
  for (int k = 1; k < n_devices; ++k) {
    peers.push_back(WorkerFor(work[k].device));
    MultiWorker* w = &work[k];
    const RtbRenderParams pk = prms[k];
    peers.back()->Submit([w, desc, cam, pk] { MultiRenderOne(w, desc, cam, pk); });
  }
  MultiRenderOne(&work[0], desc, cam, prms[0]);  // the caller's own thread renders the first device's share
  for (DeviceThread* t : peers) t->Wait();
  auto cleanup = [&]() { cudaSetDevice(work[0].device); };
  for (auto& w : work)
    if (w.rc != RTB_OK) {
      const int rc = w.rc;
      const std::string msg = "device " + std::to_string(w.device) + ": " + w.error;
      cleanup();
      return Fail(rc, msg);
    }
  // ---- reduce on devices[0]: peers read over NVLink where peer access is granted, staged copies otherwise
  const size_t n = (size_t)cam->width * cam->height * 3, bytes = n * sizeof(float);
  cudaSetDevice(work[0].device);
  cudaStream_t st = g_stream;
  cudaEvent_t e0 = nullptr, e1 = nullptr;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  cudaEventRecord(e0, st);
  rtb::PeerPtrs direct{};
  int n_direct = 0;
  float* staged = nullptr;
  cudaError_t e = cudaSuccess;
  for (int k = 1; k < n_devices && e == cudaSuccess; ++k) {
    int can = 0;
    cudaDeviceCanAccessPeer(&can, work[0].device, work[k].device);
    if (can) {
      const cudaError_t pe = cudaDeviceEnablePeerAccess(work[k].device, 0);
      if (pe != cudaSuccess && pe != cudaErrorPeerAccessAlreadyEnabled) can = 0;
      cudaGetLastError();
    }
    if (can) {
      direct.p[n_direct++] = work[k].accum;
    } else {  // no peer access: copy through, then add
      if (!staged) e = cudaMalloc(reinterpret_cast<void**>(&staged), bytes);
      if (e == cudaSuccess) e = cudaMemcpyPeerAsync(staged, work[0].device, work[k].accum, work[k].device, bytes, st);
      if (e == cudaSuccess) {
        rtb::PeerPtrs one{};
        one.p[0] = staged;
        rtb::peer_sum_kernel<<<g_sm_count * 4, 256, 0, st>>>(work[0].accum, one, 1, n);
      }
    }
  }
  if (e == cudaSuccess && n_direct > 0) rtb::peer_sum_kernel<<<g_sm_count * 4, 256, 0, st>>>(work[0].accum, direct, n_direct, n);
  cudaEventRecord(e1, st);
  if (e == cudaSuccess) e = cudaMemcpyAsync(rgb_sum, work[0].accum, bytes, cudaMemcpyDeviceToHost, st);
  if (e == cudaSuccess) e = cudaStreamSynchronize(st);
  if (e == cudaSuccess) e = cudaGetLastError();
  float ms_reduce = 0.f;
  cudaEventElapsedTime(&ms_reduce, e0, e1);
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  if (staged) cudaFree(staged);
  cleanup();
  if (e != cudaSuccess) return Fail(RTB_ECUDA, std::string("rtb_render_multi reduce: ") + cudaGetErrorString(e));
  if (stats) {
    std::memset(stats, 0, sizeof(*stats));
    for (const auto& w : work) {
      stats->samples += w.stats.samples;
      stats->rays += w.stats.rays;
      stats->launches += w.stats.launches;
      stats->iterations = std::max(stats->iterations, w.stats.iterations);
      stats->ms_total = std::max(stats->ms_total, w.stats.ms_total);
    }
    stats->ms_total += ms_reduce;
    stats->launches += (uint32_t)std::max(1, n_devices - 1);
    stats->ms_shade = ms_reduce;  // reported separately: the reduce (no shade kernel runs in the default pipeline)
    (void)t0;
  }
  return RTB_OK;
}

int rtb_render_adaptive(RtbScene* s, const RtbCamera* cam, const RtbRenderParams* prm, const RtbAdaptive* ad, float* rgb_sum, float* rgb_sumsq,
                        uint32_t* count, RtbStats* stats) {
  if (!s || !cam || !prm || !ad || !rgb_sum || !count) return Fail(RTB_EINVAL, "null argument");
  if (cam->width <= 0 || cam->height <= 0) return Fail(RTB_EINVAL, "bad camera / params");
  int rc = EnsureInit();
  if (rc != RTB_OK) return rc;
  const size_t n_pix = (size_t)cam->width * cam->height, fb = n_pix * 3 * sizeof(float);
  char* d = nullptr;  // [sum | sumsq | count]
  RTB_CUDA(cudaMalloc(reinterpret_cast<void**>(&d), 2 * fb + n_pix * 4));
  float* d_sum = reinterpret_cast<float*>(d);
  float* d_sq = reinterpret_cast<float*>(d + fb);
  uint32_t* d_cnt = reinterpret_cast<uint32_t*>(d + 2 * fb);
  cudaError_t e = cudaSuccess;
  if (prm->flags & RTB_RENDER_KEEP_ACCUM) {
    if (!rgb_sumsq) {
      cudaFree(d);
      return Fail(RTB_EINVAL, "continuing an adaptive render needs its sumsq plane");
    }
    e = cudaMemcpy(d_sum, rgb_sum, fb, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(d_sq, rgb_sumsq, fb, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(d_cnt, count, n_pix * 4, cudaMemcpyHostToDevice);
  }
  if (e == cudaSuccess) {
    rc = RenderImpl(s, cam, prm, d_sum, ad, d_sq, d_cnt, stats);
    if (rc == RTB_OK) e = cudaMemcpy(rgb_sum, d_sum, fb, cudaMemcpyDeviceToHost);
    if (rc == RTB_OK && e == cudaSuccess && rgb_sumsq) e = cudaMemcpy(rgb_sumsq, d_sq, fb, cudaMemcpyDeviceToHost);
    if (rc == RTB_OK && e == cudaSuccess) e = cudaMemcpy(count, d_cnt, n_pix * 4, cudaMemcpyDeviceToHost);
  }
  cudaFree(d);
  if (rc != RTB_OK) return rc;
  if (e != cudaSuccess) return Fail(RTB_ECUDA, std::string("adaptive render copies: ") + cudaGetErrorString(e));
  return RTB_OK;
}

int rtb_pixel_state_fold(const float* samples, size_t n_pixels, int k, const RtbAdaptive* ad, float* rgb_sum, float* rgb_sumsq, uint32_t* count) {
  if (!samples || !ad || !rgb_sum || !rgb_sumsq || !count || k < 1) return Fail(RTB_EINVAL, "bad argument");
  if (ad->criterion != RTB_ADAPT_PIXEL_STATE && ad->criterion != RTB_ADAPT_LUMINANCE) return Fail(RTB_EINVAL, "unknown adaptive criterion");
  if (n_pixels == 0) return RTB_OK;
  if (n_pixels > (1u << 30) || n_pixels * (size_t)k > (1ull << 32)) return Fail(RTB_EINVAL, "too many samples");
  int rc = EnsureInit();
  if (rc != RTB_OK) return rc;
  const size_t n = n_pixels, fb = n * 3 * sizeof(float), sb = n * (size_t)k * 3 * sizeof(float);
  auto up = [](size_t b) { return (b + 255) & ~(size_t)255; };
  const size_t o_sum = 0, o_sq = o_sum + up(fb), o_cnt = o_sq + up(fb), o_act = o_cnt + up(n * 4), o_next = o_act + up(n * 4), o_n = o_next + up(n * 4),
               o_smp = o_n + 256, total = o_smp + up(sb);
  char* d = nullptr;
  RTB_CUDA(cudaMalloc(reinterpret_cast<void**>(&d), total));
  cudaStream_t st = g_stream;
  const uint32_t n32 = (uint32_t)n;
  cudaError_t e = cudaMemsetAsync(d, 0, o_act, st);  // sum, sumsq, count
  if (e == cudaSuccess) e = cudaMemsetAsync(d + o_n, 0, 256, st);
  if (e == cudaSuccess) e = cudaMemcpyAsync(d + o_n, &n32, 4, cudaMemcpyHostToDevice, st);
  if (e == cudaSuccess) e = cudaMemcpyAsync(d + o_smp, samples, sb, cudaMemcpyHostToDevice, st);
  if (e == cudaSuccess) {
    rtb::iota_kernel<<<(n32 + 255) / 256, 256, 0, st>>>(reinterpret_cast<uint32_t*>(d + o_act), n32);
    rtb::RenderConsts rcst{};
    rcst.spp_max = k;
    rcst.min_samples = ad->min_samples;
    rcst.criterion = ad->criterion;
    rcst.rel_threshold = ad->rel_threshold;
    rtb::AdaptivePlanes pl{reinterpret_cast<float*>(d + o_sq), reinterpret_cast<uint32_t*>(d + o_cnt), reinterpret_cast<uint32_t*>(d + o_act),
                           reinterpret_cast<unsigned*>(d + o_n), reinterpret_cast<float*>(d + o_smp), (uint32_t)k, 0u};
    rtb::adaptive_replay_kernel<<<(n32 + 255) / 256, 256, 0, st>>>(rcst, reinterpret_cast<float*>(d + o_sum), pl, reinterpret_cast<uint32_t*>(d + o_next),
                                                                 reinterpret_cast<unsigned*>(d + o_n) + 1);
    e = cudaMemcpyAsync(rgb_sum, d + o_sum, fb, cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess) e = cudaMemcpyAsync(rgb_sumsq, d + o_sq, fb, cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess) e = cudaMemcpyAsync(count, d + o_cnt, n * 4, cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    if (e == cudaSuccess) e = cudaGetLastError();
  }
  cudaFree(d);
  if (e != cudaSuccess) return Fail(RTB_ECUDA, std::string("pixel_state_fold: ") + cudaGetErrorString(e));
  return RTB_OK;
}

int rtb_resolve_counts_device(const float* d_rgb_sum, const uint32_t* d_count, size_t n_pixels, uint8_t* d_rgb8, void* stream) {
  if (!d_rgb_sum || !d_count || !d_rgb8) return Fail(RTB_EINVAL, "bad argument");
  int rc = EnsureInit();
  if (rc != RTB_OK) return rc;
  cudaStream_t st = stream ? static_cast<cudaStream_t>(stream) : g_stream;
  const size_t n = n_pixels * 3;
  if (n) rtb::resolve_counts_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(d_rgb_sum, d_count, n, d_rgb8);
  RTB_CUDA(cudaGetLastError());
  return RTB_OK;
}

int rtb_resolve_counts(const float* rgb_sum, const uint32_t* count, size_t n_pixels, uint8_t* rgb8) {
  if (!rgb_sum || !count || !rgb8) return Fail(RTB_EINVAL, "bad argument");
  int rc = EnsureInit();
  if (rc != RTB_OK) return rc;
  const size_t n = n_pixels * 3;
  if (n == 0) return RTB_OK;
  char* d = nullptr;  // [sum | count | bytes]
  RTB_CUDA(cudaMalloc(reinterpret_cast<void**>(&d), n * 4 + n_pixels * 4 + n));
  float* d_sum = reinterpret_cast<float*>(d);
  uint32_t* d_cnt = reinterpret_cast<uint32_t*>(d + n * 4);
  uint8_t* d_out = reinterpret_cast<uint8_t*>(d + n * 4 + n_pixels * 4);
  cudaError_t e = cudaMemcpyAsync(d_sum, rgb_sum, n * 4, cudaMemcpyHostToDevice, g_stream);
  if (e == cudaSuccess) e = cudaMemcpyAsync(d_cnt, count, n_pixels * 4, cudaMemcpyHostToDevice, g_stream);
  if (e == cudaSuccess) {
    rc = rtb_resolve_counts_device(d_sum, d_cnt, n_pixels, d_out, g_stream);
    if (rc == RTB_OK) e = cudaMemcpyAsync(rgb8, d_out, n, cudaMemcpyDeviceToHost, g_stream);
    if (rc == RTB_OK && e == cudaSuccess) e = cudaStreamSynchronize(g_stream);
  }
  cudaFree(d);
  if (rc != RTB_OK) return rc;
  if (e != cudaSuccess) return Fail(RTB_ECUDA, std::string("resolve_counts: ") + cudaGetErrorString(e));
  return RTB_OK;
}

int rtb_resolve_device(const float* d_rgb_sum, size_t n_pixels, int spp, uint8_t* d_rgb8, void* stream) {
  if (!d_rgb_sum || !d_rgb8 || spp <= 0) return Fail(RTB_EINVAL, "bad argument");
  int rc = EnsureInit();
  if (rc != RTB_OK) return rc;
  cudaStream_t st = stream ? static_cast<cudaStream_t>(stream) : g_stream;
  const size_t n = n_pixels * 3;
  if (n) rtb::resolve_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(d_rgb_sum, n, 1.0 / (double)spp, d_rgb8);
  RTB_CUDA(cudaGetLastError());
  return RTB_OK;
}

int rtb_resolve(const float* rgb_sum, size_t n_pixels, int spp, uint8_t* rgb8) {
  if (!rgb_sum || !rgb8 || spp <= 0) return Fail(RTB_EINVAL, "bad argument");
  int rc = EnsureInit();
  if (rc != RTB_OK) return rc;
  const size_t n = n_pixels * 3;
  if (n == 0) return RTB_OK;
  float* d_in = nullptr;
  uint8_t* d_out = nullptr;
  RTB_CUDA(cudaMalloc(reinterpret_cast<void**>(&d_in), n * 4));
  cudaError_t e = cudaMalloc(reinterpret_cast<void**>(&d_out), n);
  if (e == cudaSuccess) e = cudaMemcpyAsync(d_in, rgb_sum, n * 4, cudaMemcpyHostToDevice, g_stream);
  if (e == cudaSuccess) {
    rc = rtb_resolve_device(d_in, n_pixels, spp, d_out, g_stream);
    if (rc == RTB_OK) e = cudaMemcpyAsync(rgb8, d_out, n, cudaMemcpyDeviceToHost, g_stream);
    if (rc == RTB_OK && e == cudaSuccess) e = cudaStreamSynchronize(g_stream);
  }
  cudaFree(d_in);
  cudaFree(d_out);
  if (rc != RTB_OK) return rc;
  if (e != cudaSuccess) return Fail(RTB_ECUDA, std::string("resolve: ") + cudaGetErrorString(e));
  return RTB_OK;
}

// ---- roofline denominators (bench.py): L2-resident read bandwidth, FP32 and FP64 FMA rate ----
int rtb_measure_l2_gbs(size_t bytes, int iters, float* gbs) {
  if (!gbs || bytes < 4096 || iters < 1) return Fail(RTB_EINVAL, "bad argument");
  int rc = EnsureInit();
  if (rc != RTB_OK) return rc;
  float4* buf = nullptr;
  float* sink = nullptr;
  const size_t n4 = bytes / 16;
  RTB_CUDA(cudaMalloc(reinterpret_cast<void**>(&buf), n4 * 16));
  RTB_CUDA(cudaMalloc(reinterpret_cast<void**>(&sink), 4));
  RTB_CUDA(cudaMemsetAsync(buf, 0, n4 * 16, g_stream));
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  const int grid = g_sm_count * 8;
  rtb::l2_read_kernel<<<grid, 256, 0, g_stream>>>(buf, n4, 2, sink);  // warm L2
  cudaEventRecord(e0, g_stream);
  rtb::l2_read_kernel<<<grid, 256, 0, g_stream>>>(buf, n4, iters, sink);
  cudaEventRecord(e1, g_stream);
  cudaError_t e = cudaEventSynchronize(e1);
  float ms = 0.f;
  cudaEventElapsedTime(&ms, e0, e1);
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  cudaFree(buf);
  cudaFree(sink);
  if (e != cudaSuccess) return Fail(RTB_ECUDA, std::string("l2 bench: ") + cudaGetErrorString(e));
  *gbs = (float)((double)n4 * 16 * iters / (ms * 1e-3) / 1e9);
  return RTB_OK;
}

}  // extern "C"

template <class T, class K>
static int MeasureFma(K kernel, int iters, float* tflops) {
  if (!tflops || iters < 1) return Fail(RTB_EINVAL, "bad argument");
  int rc = EnsureInit();
  if (rc != RTB_OK) return rc;
  T* sink = nullptr;
  RTB_CUDA(cudaMalloc(reinterpret_cast<void**>(&sink), sizeof(T)));
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  const int grid = g_sm_count * 8, block = 256;
  kernel<<<grid, block, 0, g_stream>>>(iters / 8 + 1, sink);
  cudaEventRecord(e0, g_stream);
  kernel<<<grid, block, 0, g_stream>>>(iters, sink);
  cudaEventRecord(e1, g_stream);
  cudaError_t e = cudaEventSynchronize(e1);
  float ms = 0.f;
  cudaEventElapsedTime(&ms, e0, e1);
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  cudaFree(sink);
  if (e != cudaSuccess) return Fail(RTB_ECUDA, std::string("fma bench: ") + cudaGetErrorString(e));
  *tflops = (float)((double)grid * block * 8.0 * 2.0 * iters / (ms * 1e-3) / 1e12);
  return RTB_OK;
}
extern "C" {
int rtb_measure_fp32_tflops(int iters, float* tflops) { return MeasureFma<float>(rtb::fp32_fma_kernel, iters, tflops); }
int rtb_measure_fp64_tflops(int iters, float* tflops) { return MeasureFma<double>(rtb::fp64_fma_kernel, iters, tflops); }

}  // extern "C"
