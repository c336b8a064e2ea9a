/* rth.h — C view of the host-side scene API (3360-ray-tracer_b200/host/rt/...).
 *
 * The product's host is C++ (namespace rt::, same classes and constructors as the reference);
 * these wrappers exist so that Python (tests, bench.py) can build the named scenes, obtain the
 * flattened RtbSceneDesc for rtb_scene_create(), and export the scene for the oracle.
 * Replaces: the scene builders of src/main.cc:23-156, scene::loadCameras (src/scene/camera.h:58-67),
 * Camera::Initialize (camera.h:100-131) and the Bvh accessors (src/geom/bvh.h:134-136).
 */
#ifndef RTH_H_
#define RTH_H_
#include "rtb.h"
#ifdef __cplusplus
extern "C" {
#endif

typedef struct RthScene RthScene;

/* scene::CameraConfig (src/scene/camera.h:25-38) */
typedef struct RthCameraConfig {
  double aspect_ratio;
  int32_t image_width;
  int32_t samples_per_pixel;
  int32_t max_depth;
  int32_t pad;
  double vfov;
  double lookfrom[3];
  double lookat[3];
  double vup[3];
  double defocus_angle;
  double focus_dist;
} RthCameraConfig;

const char* rth_last_error(void);

/* librth.so (this header: host C++ scene API, no CUDA) and librtb200.so (rtb.h: the kernels) are separate libraries.
 * The host library can hand a freshly built scene to the device library's tree compiler (rtb_tree_compile, host work
 * done once per scene like the reference's geom::Bvh constructor, src/geom/bvh.h:39-68) if the embedder registers it;
 * without it RtbSceneDesc.compiled stays NULL and rtb_scene_create compiles the tree itself on every call.
 * Pass NULLs to unregister.  Scenes built while a compiler was registered must be freed before it is unregistered. */
typedef int (*RthTreeCompileFn)(const RtbSceneDesc*, RtbCompiledTree**);
typedef void (*RthTreeDestroyFn)(RtbCompiledTree*);
typedef const char* (*RthLastErrorFn)(void);
void rth_set_tree_compiler(RthTreeCompileFn compile, RthTreeDestroyFn destroy, RthLastErrorFn last_error);

/* name: cornell | rtiow | checkered_spheres | earth | earth_bubbles | bunny | stress.
 * params: scene-specific doubles (see rth_capi.cc); asset_dir holds models/ and textures/.
 * top_budget: nodes in the breadth-first prefix (0 = 1024). */
int rth_scene_build(const char* name, const double* params, int n_params, const char* asset_dir,
                    uint32_t top_budget, RthScene** out);
/* flags: RTH_BUILD_DEVICE_TREE = no host SAH build / flatten / compile; the RtbSceneDesc carries the primitive
 * tables only and RTB_SCENE_DEVICE_BUILD, and rtb_scene_create builds the tree on the GPU. */
#define RTH_BUILD_DEVICE_TREE 1u
int rth_scene_build_ex(const char* name, const double* params, int n_params, const char* asset_dir,
                       uint32_t top_budget, uint32_t flags, RthScene** out);
/* Build from an export stream (16 doubles per primitive record, 8 per material / texture). */
int rth_scene_from_records(const double* prims, size_t n_prims, const double* mats, size_t n_mats,
                           const double* texs, size_t n_texs, const char* image_paths,
                           uint32_t top_budget, RthScene** out);
const RtbSceneDesc* rth_scene_desc(const RthScene* s);
int rth_scene_export(const RthScene* s, const double** prims, size_t* n_prims, const double** mats,
                     size_t* n_mats, const double** texs, size_t* n_texs, const char** image_paths);
int rth_scene_gid_map(const RthScene* s, const uint32_t** gid_to_ref, size_t* n);
/* Host SAH trees in creation order: nodes as 9 doubles {x0,x1,y0,y1,z0,z1,left_pIdx,right_pCnt,isLeaf}. */
int rth_scene_bvh_count(const RthScene* s);
int rth_scene_bvh_size(const RthScene* s, int k, size_t* n_nodes, size_t* n_prims);
int rth_scene_bvh_dump(const RthScene* s, int k, double* nodes9, int32_t* prim_indices);
void rth_scene_free(RthScene* s);

int rth_camera_init(const RthCameraConfig* cfg, int depth_camera, RtbCamera* out);
int rth_camera_preset(const char* json_path, const char* name, RthCameraConfig* out);

/* The 8-bit image of rtb_resolve as a PNG file (8-bit RGB, stored deflate blocks; host/rt/core/png.h) — the binary
 * output SURVEY 8f-4 asks for beside the reference's ASCII "P3" (src/core/color.h:18-33).  out == NULL: *n_bytes = the
 * size needed.  Otherwise writes at most `capacity` bytes. */
int rth_encode_png(const uint8_t* rgb8, int width, int height, uint8_t* out, size_t capacity, size_t* n_bytes);

#ifdef __cplusplus
}
#endif
#endif
