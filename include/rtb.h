/* rtb.h — C ABI of the B200-native path-tracing hot path (librtb200.so).
 *
 * This is the drop-in boundary.  The reference
 * (Luke-TS/3360-ray-tracer) has no FFI of its own: its render loop is reached
 * through two C++ seams, and every entry point below names the seam it stands
 * behind (paths relative to the reference root):
 *
 *   seam 1  renderer::MegaKernel::Render            src/renderer/mega_kernel.h:15-54
 *           renderer::WavefrontRenderer::Render     src/renderer/wavefront.cc:40-242
 *   seam 2  integrator::RayIntegrator::IntersectBatch
 *                                                   src/integrator/ray_integrator.h:30-37
 *                                                   src/integrator/cpu_ray_integrator.h:18-46
 *
 * Plain pointers and sizes only; no C++/torch types.  Every call returns 0 on
 * success and a negative RTB_E* code on failure; rtb_last_error() gives the
 * text.  There is no CPU fallback: without a CUDA device every compute entry
 * fails with RTB_ENODEVICE.
 */
#ifndef RTB_H_
#define RTB_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RTB_ABI_VERSION 3

/* ---- status codes ------------------------------------------------------ */
enum {
  RTB_OK = 0,
  RTB_EINVAL = -1,    /* bad argument / malformed scene description        */
  RTB_ENODEVICE = -2, /* no usable CUDA device (never falls back to CPU)   */
  RTB_ECUDA = -3,     /* CUDA runtime error, text in rtb_last_error()      */
  RTB_ENOMEM = -4
};

/* ---- primitive / material / texture tags ------------------------------- */
/* Primitive tags follow geom::HittableType (src/geom/hittable.h:45-49);
 * SUBTREE is how a nested geom::Bvh inside a leaf is represented.          */
enum { RTB_PRIM_SPHERE = 0, RTB_PRIM_TRIANGLE = 1, RTB_PRIM_RECT = 2, RTB_PRIM_SUBTREE = 3 };
enum { RTB_MAT_LAMBERTIAN = 0, RTB_MAT_METAL = 1, RTB_MAT_DIELECTRIC = 2, RTB_MAT_DIFFUSE_LIGHT = 3 };
enum { RTB_TEX_SOLID = 0, RTB_TEX_CHECKER = 1, RTB_TEX_IMAGE = 2 };
/* Estimator: A = Camera::GetPixel + Material::Scatter (src/scene/camera.h:148-174),
 *            B = WavefrontRenderer + Material::Sample + Russian roulette
 *                (src/renderer/wavefront.cc:85-226).                        */
enum { RTB_MODE_A = 0, RTB_MODE_B = 1 };

/* Leaf reference: type in the top 2 bits, index into that type's table below. */
#define RTB_REF(type, index) (((uint32_t)(type) << 30) | (uint32_t)(index))
#define RTB_REF_TYPE(ref) ((ref) >> 30)
#define RTB_REF_INDEX(ref) ((ref) & 0x3fffffffu)
#define RTB_MAX_LEAF 63 /* leaf prim count must fit 6 bits (flattener splits larger) */

/* ---- flattened scene (host arrays; rtb_scene_create copies them) ------- */

/* 32-byte BVH node, replaces geom::BvhNodeGPU (src/geom/bvh.h:20-25).
 * Box is the reference's double box rounded OUTWARD to float.
 *   b == 0 : internal; children are nodes[a] (left) and nodes[a+1] (right)
 *   b  > 0 : leaf; prim_refs[a .. a+b)                                      */
typedef struct RtbNode {
  float lo[3];
  uint32_t a;
  float hi[3];
  uint32_t b;
} RtbNode;

/* geom::Sphere (src/geom/sphere.h:81-85): exact doubles for the replay.    */
typedef struct RtbSphere {
  double c[3];
  double r;
} RtbSphere;

/* geom::Triangle (src/geom/triangle.h:105-108).  e1 = b-a, e2 = c-a are the
 * doubles Triangle::Hit recomputes on every call (triangle.h:45-46).       */
typedef struct RtbTriangle {
  double a[3];
  double e1[3];
  double e2[3];
  uint32_t mat;
  uint32_t pad;
} RtbTriangle; /* 80 B = 5 x 128-bit loads */

/* geom::xy_rect / xz_rect / yz_rect (src/geom/rect.h).  axis = the constant
 * axis (2 = xy_rect, 1 = xz_rect, 0 = yz_rect); (a,b) = the other two axes
 * in increasing order.                                                     */
typedef struct RtbRect {
  double k;
  double a0, a1, b0, b1;
  int32_t axis;
  uint32_t mat;
} RtbRect; /* 48 B */

typedef struct RtbMaterial {
  int32_t type;   /* RTB_MAT_*                                   */
  int32_t tex;    /* lambertian albedo / light emission texture  */
  float albedo[3];/* metal                                       */
  float fuzz;     /* metal, already min(fuzz,1) (material.cc:78-80) */
  float ior;      /* dielectric                                  */
  float pad;
} RtbMaterial; /* 32 B */

typedef struct RtbTexture {
  int32_t type;  /* RTB_TEX_*                      */
  int32_t image; /* image index, RTB_TEX_IMAGE     */
  int32_t even;  /* texture index, RTB_TEX_CHECKER */
  int32_t odd;
  float color[3]; /* RTB_TEX_SOLID                 */
  float inv_scale;/* RTB_TEX_CHECKER (texture.h:34) */
} RtbTexture; /* 32 B */

/* scene::Image byte table (src/scene/image.cc:43-48): width*height*3 bytes,
 * already gamma-2.2-linearised and re-quantised exactly as the reference.
 * width == 0 means "failed to load" (texture.h:61 -> cyan).                */
typedef struct RtbImage {
  int32_t width, height;
  const uint8_t* rgb;
} RtbImage;

typedef struct RtbSceneDesc {
  uint32_t abi_version; /* RTB_ABI_VERSION */
  uint32_t n_nodes;     /* nodes[0] is the root; nodes[1] is padding so that sibling pairs are 64-B aligned */
  const RtbNode* nodes;
  uint32_t n_prim_refs;
  const uint32_t* prim_refs;
  uint32_t n_spheres;
  const RtbSphere* spheres;
  const uint32_t* sphere_mat;
  uint32_t n_triangles;
  const RtbTriangle* triangles;
  uint32_t n_rects;
  const RtbRect* rects;
  /* per-type: global primitive id (creation order in the host scene) and the
   * reference's DFS slot rank used only to break exact-t ties (bvh.h:95-106) */
  const uint32_t* sphere_gid;
  const uint32_t* triangle_gid;
  const uint32_t* rect_gid;
  const uint32_t* sphere_rank;
  const uint32_t* triangle_rank;
  const uint32_t* rect_rank;
  uint32_t n_materials;
  const RtbMaterial* materials;
  uint32_t n_textures;
  const RtbTexture* textures;
  uint32_t n_images;
  const RtbImage* images;
  uint32_t n_top_nodes; /* nodes[0..n_top_nodes) are the BFS-ordered top levels (shared-memory stage) */
  uint32_t max_depth;   /* deepest root-to-leaf path, for the traversal stack bound */
  /* optional: the device-native traversal tree built ahead of time with rtb_tree_compile() - host work that belongs
   * to scene construction, like the reference's Bvh constructor (bvh.h:39-68), not to the render call.  NULL:
   * rtb_scene_create builds it itself. */
  const struct RtbCompiledTree* compiled;
  /* RTB_SCENE_*.  With RTB_SCENE_DEVICE_BUILD rtb_scene_create ignores nodes / prim_refs / compiled (they may be
   * NULL / 0) and builds the traversal tree on the GPU from the primitive tables (linear BVH, csrc/rtb_build.cuh):
   * the host never runs geom::Bvh's SAH build (src/geom/bvh.h:166-344).  The closest hit does not depend on the
   * tree; only the order of EXACT-t ties does: the *_rank arrays may then be NULL and creation order (gid) stands in
   * for the reference's slot order (bvh.h:95-106). */
  uint32_t flags;
} RtbSceneDesc;
#define RTB_SCENE_DEVICE_BUILD 1u
typedef struct RtbCompiledTree RtbCompiledTree; /* opaque, host memory only */

/* ---- rays and hits (seam 2) -------------------------------------------- */
/* core::Ray (src/core/ray.h:7-26) + the core::Interval passed to Hit.
 * Values are float; the reference sees exactly these values widened to double. */
typedef struct RtbRay {
  float o[3];
  float tmin;
  float d[3]; /* NOT normalised, as in the reference */
  float tmax;
} RtbRay; /* 32 B */

typedef struct RtbHit {
  double t;        /* HitRecord::t as the reference computes it (exact replay) */
  int32_t prim_id; /* global primitive id, -1 = miss (HitRecord::hit false)    */
  uint32_t ref;    /* RTB_REF(type,index) of the primitive                     */
} RtbHit; /* 16 B */

/* ---- camera / render (seam 1) ------------------------------------------ */
/* What scene::Camera::Initialize derives (src/scene/camera.h:100-131).     */
typedef struct RtbCamera {
  int32_t width, height;
  float center[3];
  float pixel00[3];
  float delta_u[3];
  float delta_v[3];
  float disk_u[3];
  float disk_v[3];
  float defocus_angle;
  int32_t depth_camera; /* 1 = scene::DepthCamera semantics (camera.h:215-241) */
} RtbCamera;

typedef struct RtbRenderParams {
  int32_t spp_begin;  /* this call renders global sample indices [spp_begin, spp_end) of every pixel */
  int32_t spp_end;
  int32_t max_depth;
  int32_t mode;       /* RTB_MODE_A / RTB_MODE_B */
  uint64_t seed;      /* Philox key */
  uint32_t pool_size; /* live paths in flight per wavefront, 0 = default */
  uint32_t flags;     /* RTB_RENDER_* */
} RtbRenderParams;
#define RTB_RENDER_SMEM_TOP 1u     /* stage the BFS-ordered top of the BVH in shared memory (off by default: measured slower than L1 on B200) */
#define RTB_RENDER_KEEP_ACCUM 2u   /* add into the accumulator instead of zeroing it first */
#define RTB_RENDER_SERIAL 4u       /* wavefront mode with one wavefront, each kernel timed alone (fills RtbStats.ms_generate/extend/shade) */
#define RTB_RENDER_WAVEFRONT 8u    /* generate / extend / shade+compact kernels over path queues instead of the default persistent path kernel
                                      (generate + traverse + shade per lane, no queues; measured 13-53 % faster) */
#define RTB_RENDER_NO_DRAIN 16u    /* wavefront mode: do not hand the last few live paths to the path kernel */
#define RTB_RENDER_COST_ORDER 64u  /* cost-ordered schedule: the first samples measure where the long paths are, the rest of the render starts with
                                      those tiles (three overlapping launches).  Same image; measured slower than one launch */
#define RTB_RENDER_WIDE4 128u      /* traverse the 4-wide form of the tree (only scenes created with RTB_WIDE4=1 in the environment carry it;
                                      RTB_EINVAL otherwise).  A measured experiment, not the default: DESIGN.md 3.9 */
#define RTB_RENDER_POOL_PATHS 32u  /* the pool kernel (64 paths per warp in shared memory, traversal and shading decoupled: csrc/rtb_pool.cuh)
                                      instead of the default persistent kernel (one path per lane, shaded in place: path_kernel).  Same paths, same
                                      image; measured 16-58 % slower (profiles/r02_pool_vs_lane_ab.jsonl) - kept as the evidence */

typedef struct RtbStats {
  uint64_t samples;     /* camera samples started                      */
  uint64_t rays;        /* closest-hit queries issued (Bvh::Hit calls) */
  uint64_t node_visits;    /* nodes fetched+tested; these four only when built with RTB_COUNTERS */
  uint64_t sphere_tests;   /* Sphere::Hit replays   */
  uint64_t triangle_tests; /* Triangle::Hit replays */
  uint64_t rect_tests;     /* *_rect::Hit replays   */
  uint32_t iterations;  /* wavefront iterations (one bounce each)      */
  uint32_t launches;    /* kernels launched                            */
  float ms_total;       /* CUDA-event time, first launch -> accumulator final */
  float ms_generate;    /* per-kernel sums: only with RTB_RENDER_SERIAL (0 otherwise: wavefronts overlap) */
  float ms_extend;
  float ms_shade;
  float ms_drain;       /* path kernel finishing the last live paths of each wavefront (RTB_RENDER_SERIAL) */
  uint32_t drain_paths; /* paths handed to it */
} RtbStats;

/* Adaptive sampling.  Every pixel takes samples spp_begin, spp_begin+1, ... until its stop rule
 * holds or spp_end is reached; the rule is evaluated after every recorded sample once min_samples are in.
 *   RTB_ADAPT_PIXEL_STATE  integrator::IsConverged (src/integrator/pixel_state.h:53-72), what
 *                          WavefrontRenderer::Render applies with kRelThresh 0.05 / kMinSamples 16
 *                          (src/renderer/wavefront.cc:42-43): per channel, standard error of the mean
 *                          sqrt(s^2/n) over max(|mean|, 1e-3) <= rel_threshold, s^2 = m2/(n-1).
 *   RTB_ADAPT_LUMINANCE    integrator::AdaptiveSampler (src/integrator/sampler.h:46-76) as it was meant (its
 *                          `sum += pixel` adds the running total; here the sample): sqrt(lum(E[x^2]-mean^2)/n)
 *                          / (lum(mean) + 1e-3) < rel_threshold, lum = core::luminance (src/core/color.h:35-37). */
enum { RTB_ADAPT_PIXEL_STATE = 0, RTB_ADAPT_LUMINANCE = 1 };
typedef struct RtbAdaptive {
  float rel_threshold;
  int32_t min_samples;
  int32_t criterion; /* RTB_ADAPT_* */
  int32_t pad;
} RtbAdaptive;

typedef struct RtbScene RtbScene; /* opaque; owns all device memory of one scene on one GPU */

const char* rtb_last_error(void);
int rtb_abi_version(void);
/* Number of visible CUDA devices (0 if none); never fails. */
int rtb_device_count(void);
/* The library's own bounds checker.  A build with -DRTB_CHECKED (make VARIANT=checked EXTRA=-DRTB_CHECKED) compares every index
 * its kernels form (traversal stack, pair, primitive, accumulator pixel, material / texture, adaptive chunk, pool slot) with
 * the size of what it indexes, clamps the access and counts violations per class on the current device since the library was
 * loaded.  Returns 1 and the eight counters from such a build, 0 (RTB_OK) and zeros from a normal build, < 0 on error. */
int rtb_debug_check_failures(uint64_t* counts8);
/* Host only: n / d the way the sample generator of the kernels divides (multiply-shift with per-launch constants, csrc/rtb_kernels.cuh
 * FastDiv) - exists so that the CPU test suite can check the constants against the hardware division; d == 0 returns 0. */
uint32_t rtb_debug_fast_div(uint32_t n, uint32_t d);
/* Bind the calling thread to `device` and create the library's stream on it. */
int rtb_init(int device);
int rtb_shutdown(void);

/* Host only (works without a GPU): interchange nodes -> the sibling-pair tree the kernels walk.
 * The result also caches, from the first rtb_scene_create that uses it, the scene's complete device image in pinned
 * host memory, so that creating the scene again (another frame, another GPU of the same process) is one asynchronous
 * copy: the tables of `desc` must not change between rtb_tree_compile and the rtb_scene_create calls that pass its
 * result in RtbSceneDesc.compiled (change the scene -> compile again).  Use one compiled tree from one host thread
 * at a time. */
int rtb_tree_compile(const RtbSceneDesc* desc, RtbCompiledTree** out);
void rtb_tree_destroy(RtbCompiledTree* tree);
/* Host only.  info = { pairs, deepest chain of pairs, primitives reachable through the pairs' leaves, child boxes outside their
 * parent slot by more than 1e-5 relative (0 in a valid tree), quads (0 unless compiled with RTB_WIDE4=1), stack bound of the quad
 * traversal, primitives reachable through the quads, child boxes outside their parent slot }. */
int rtb_tree_info(const RtbCompiledTree* tree, uint32_t info[8]);

int rtb_scene_create(const RtbSceneDesc* desc, RtbScene** out);
int rtb_scene_destroy(RtbScene* scene);

/* seam 2: closest hit for n rays, host buffers (copies inside).
 * Input domain: finite origin and direction, tmin and tmax not NaN (tmax = +inf is the usual open interval).  Inside it
 * the answer is the reference's (id bit-exact), including the edges — a zero direction, an empty or inverted interval
 * and tmin < 0 report what Bvh::Hit reports, direction components that are zero, denormal or below 1e-20 next to
 * normal ones are handled as axis-parallel (tests/test_hits_gpu.py::test_degenerate_rays).  A direction whose EVERY
 * component is below 1e-20 in magnitude has hit distances beyond the float range of RtbHit.t and is outside the domain.
 * Non-finite components are outside it too: the reference lets NaN through its comparisons and reports hits at t = NaN
 * (sphere.h:34-45); here the call returns normally and every ray reports a miss or some primitive of the scene. */
int rtb_intersect_batch(RtbScene* scene, const RtbRay* rays, size_t n, RtbHit* hits);
/* The same call for callers that must not allocate (code under CUDA graph capture, tight serving loops, harnesses that forbid
 * cudaMalloc).  No counterpart in the reference.  Host buffers in and out like
 * rtb_intersect_batch and the same answers bit for bit, but the call performs NO cudaMalloc / cudaFree / allocating library call:
 * the kernels' scratch (work cursor, device copies of rays and hits) is a file-scope device array of RTB_STATIC_BATCH_MAX
 * entries - 3 MiB + 256 bytes that every GPU which loads the library holds until the process exits.  n > RTB_STATIC_BATCH_MAX is
 * refused with RTB_EINVAL.  All calls on one GPU share that array (it keeps the last call's contents; the call clears what must
 * start at zero): calls that overlap on the same GPU - from several host threads - race and must be serialised by the caller.
 * The scene itself is created beforehand (rtb_scene_create allocates); returns after the hits are in `hits`. */
#define RTB_STATIC_BATCH_MAX 65536u
int rtb_intersect_batch_static(RtbScene* scene, const RtbRay* rays, size_t n, RtbHit* hits);
/* same, device buffers, asynchronous on `stream` (a cudaStream_t, 0 = library stream). */
int rtb_intersect_batch_device(RtbScene* scene, const RtbRay* d_rays, size_t n, RtbHit* d_hits,
                               void* stream);

/* seam 1: render; rgb_sum = width*height*3 floats, SUM of sample radiance
 * (divide by the sample count, or pass it to rtb_resolve). Host buffer. */
int rtb_render(RtbScene* scene, const RtbCamera* cam, const RtbRenderParams* params,
               float* rgb_sum, RtbStats* stats);
/* same, accumulator already on the device (e.g. a torch tensor that a
 * following ncclReduce sums across GPUs). Synchronous on return. */
int rtb_render_device(RtbScene* scene, const RtbCamera* cam, const RtbRenderParams* params,
                      float* d_rgb_sum, RtbStats* stats);

/* seam 1 on several GPUs of one box, from ONE host thread (what the reference's single-threaded main()
 * would call).  The library runs one worker thread per device: each creates the scene on its device (scene
 * replicated), renders its slice of the GLOBAL sample indices [spp_begin, spp_end) (slices differ by at most one
 * sample per pixel) and leaves a float32 accumulator in its HBM; devices[0] then sums the peers' accumulators with
 * one kernel that reads them over NVLink peer access (staged through cudaMemcpyPeer where peer access is refused)
 * and the sum is copied to rgb_sum (W*H*3 floats, host).  The image equals the single-GPU render of the same
 * range up to float summation order.  stats: samples / rays / launches summed over the devices, ms_total = the
 * slowest device's render + the reduce.  devices == NULL: 0 .. n_devices-1. */
int rtb_render_multi(const RtbSceneDesc* desc, int n_devices, const int* devices, const RtbCamera* cam,
                     const RtbRenderParams* params, float* rgb_sum, RtbStats* stats);

/* seam 1, adaptive: as rtb_render, plus the per-pixel state of integrator::PixelState (pixel_state.h:13-19) in
 * planes: rgb_sum (W*H*3, PixelState::sum), rgb_sumsq (W*H*3, sum of squares: m2 = sumsq - sum^2/n) and count
 * (W*H, PixelState::samples).  The image is rgb_sum / count (wavefront.cc:229-234) - rtb_resolve_counts.
 * Without RTB_RENDER_KEEP_ACCUM the three planes are zeroed first; with it the call continues them (every pixel
 * takes at least one more sample).  Samples are traced speculatively, a chunk per pixel at a time, and folded in sample
 * order with the rule evaluated after every one: planes and image are those of the sequential rule; RtbStats.samples
 * counts the camera samples started, including the up to chunk-1 per pixel that the rule then discarded.  Path kernel only. */
int rtb_render_adaptive(RtbScene* scene, const RtbCamera* cam, const RtbRenderParams* params, const RtbAdaptive* adaptive,
                        float* rgb_sum, float* rgb_sumsq, uint32_t* count, RtbStats* stats);
int rtb_render_adaptive_device(RtbScene* scene, const RtbCamera* cam, const RtbRenderParams* params, const RtbAdaptive* adaptive,
                               float* d_rgb_sum, float* d_rgb_sumsq, uint32_t* d_count, RtbStats* stats);
/* integrator::RecordSample + IsConverged (src/integrator/pixel_state.h:22-72) as a batch operation on given samples:
 * for every pixel, fold samples[pixel*k .. pixel*k+k) (rgb floats) into its PixelState planes IN ORDER, evaluating the
 * stop rule after each one exactly as rtb_render_adaptive does (same kernel), and stop at the first sample where it
 * holds.  The planes are zeroed first; count[pixel] = samples recorded (k if the rule never held).  Host buffers. */
int rtb_pixel_state_fold(const float* samples, size_t n_pixels, int k, const RtbAdaptive* adaptive, float* rgb_sum,
                         float* rgb_sumsq, uint32_t* count);
/* write_color of rgb_sum / count per pixel (black where count is 0, wavefront.cc:229-234). */
int rtb_resolve_counts_device(const float* d_rgb_sum, const uint32_t* d_count, size_t n_pixels, uint8_t* d_rgb8, void* stream);
int rtb_resolve_counts(const float* rgb_sum, const uint32_t* count, size_t n_pixels, uint8_t* rgb8);

/* core::write_color (src/core/color.h:18-33): mean -> sqrt gamma -> clamp
 * [0,0.999] -> int(256 x); done in double like the reference.  Device in/out. */
int rtb_resolve_device(const float* d_rgb_sum, size_t n_pixels, int spp, uint8_t* d_rgb8,
                       void* stream);
int rtb_resolve(const float* rgb_sum, size_t n_pixels, int spp, uint8_t* rgb8);

/* Microbenchmarks used for the roofline denominators (bench.py). */
int rtb_measure_l2_gbs(size_t bytes, int iters, float* gbs);
int rtb_measure_fp32_tflops(int iters, float* tflops);
int rtb_measure_fp64_tflops(int iters, float* tflops);

#ifdef __cplusplus
}
#endif
#endif /* RTB_H_ */
