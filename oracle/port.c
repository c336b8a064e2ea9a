/* oracle/port.c — plain-C restatement of the reference's closest-hit path.
 *
 * TEST INFRASTRUCTURE ONLY: used by tests/ (and as the CPU-baseline fallback of bench.py when
 * oracle/_ref was not built).  Never linked into, imported by or called from the product.
 *
 * Every function restates one reference function in the reference's own arithmetic (double,
 * with Triangle::Hit's float roundings), compiled with -ffp-contract=off so that no FMA changes a
 * bit.  Citations are relative to the reference root.  The BVH arrays (nodes, prim_indices) are
 * INPUT: they come from the product's host SAH builder, which tests/test_host_scene.py pins node
 * for node against geom::Bvh of the compiled reference (tests/golden/bvh_*.npz).
 *
 * Pinned by tests/test_oracle_port.py against tests/golden/hits_*.npz — (prim_id, t) produced by
 * the reference itself, compiled from /root/reference by oracle/Makefile (`make ref`).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

enum { K_SPHERE = 0, K_TRIANGLE = 1, K_RECT = 2, K_BVH_BEGIN = 10, K_BVH_END = 11, K_LIST_BEGIN = 12, K_LIST_END = 13 };
enum { O_PRIM = 0, O_BVH = 1, O_LIST = 2 };
#define REC 16

typedef struct { double o[3], d[3]; } Ray;           /* core::Ray, src/core/ray.h:7-26 */
typedef struct { double lo[3], hi[3]; } Box;         /* geom::Aabb, src/geom/aabb.h */

typedef struct Object {
  int kind;            /* O_* */
  int prim;            /* K_SPHERE / K_TRIANGLE / K_RECT */
  int gid;             /* creation-order id of a primitive */
  double p[10];        /* sphere c,r | triangle a,b,c | rect axis,k,a0,a1,b0,b1 */
  int* kids;           /* O_BVH / O_LIST: child object ids, in creation order */
  int n_kids;
  /* O_BVH: geom::BvhNodeGPU[] (src/geom/bvh.h:20-25) and prim_indices_ */
  const double* nodes9; /* x0,x1,y0,y1,z0,z1,left_pIdx,right_pCnt,isLeaf */
  size_t n_nodes;
  const int32_t* prim_idx;
} Object;

typedef struct {
  Object* obj;
  int n_obj, cap;
  int root;
  int next_gid, next_bvh;
  /* owned copies of the BVH arrays */
  double** nodes;
  int32_t** idx;
  int n_bvh;
} PortScene;

/* ---- geom::Aabb::Hit, src/geom/aabb.h:92-115 ------------------------------------------- */
static int aabb_hit(const double* n9, const Ray* r, double tmin, double tmax) {
  for (int axis = 0; axis < 3; ++axis) {
    const double adinv = 1.0 / r->d[axis];
    const double t0 = (n9[2 * axis] - r->o[axis]) * adinv;
    const double t1 = (n9[2 * axis + 1] - r->o[axis]) * adinv;
    if (t0 < t1) {
      if (t0 > tmin) tmin = t0;
      if (t1 < tmax) tmax = t1;
    } else {
      if (t1 > tmin) tmin = t1;
      if (t0 < tmax) tmax = t0;
    }
    if (tmax <= tmin) return 0;
  }
  return 1;
}

static double dot3(const double* a, const double* b) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; } /* vec3.h:95-97 */
static void cross3(const double* u, const double* v, double* o) {                                       /* vec3.h:99-103 */
  o[0] = u[1] * v[2] - u[2] * v[1];
  o[1] = u[2] * v[0] - u[0] * v[2];
  o[2] = u[0] * v[1] - u[1] * v[0];
}

/* ---- geom::Sphere::Hit, src/geom/sphere.h:23-55 (Interval::Surrounds is strict, interval.h:36-38) */
static int sphere_hit(const double* p, const Ray* r, double tmin, double tmax, double* t) {
  const double oc[3] = {p[0] - r->o[0], p[1] - r->o[1], p[2] - r->o[2]};
  const double a = dot3(r->d, r->d);
  const double h = dot3(r->d, oc);
  const double c = dot3(oc, oc) - p[3] * p[3];
  const double disc = h * h - a * c;
  if (disc < 0) return 0;
  const double sq = sqrt(disc);
  double root = (h - sq) / a;
  if (!(tmin < root && root < tmax)) {
    root = (h + sq) / a;
    if (!(tmin < root && root < tmax)) return 0;
  }
  *t = root;
  return 1;
}

/* ---- geom::Triangle::Hit, src/geom/triangle.h:41-87: double products, float det/inv_det/u/v/t */
static int triangle_hit(const double* p, const Ray* r, double tmin, double tmax, double* t_out) {
  const double e1[3] = {p[3] - p[0], p[4] - p[1], p[5] - p[2]};
  const double e2[3] = {p[6] - p[0], p[7] - p[1], p[8] - p[2]};
  double pvec[3], qvec[3];
  cross3(r->d, e2, pvec);
  const float det = (float)dot3(e1, pvec);
  if (fabs((double)det) < (double)1e-6f) return 0;
  const float inv_det = 1.0f / det;
  const double tvec[3] = {r->o[0] - p[0], r->o[1] - p[1], r->o[2] - p[2]};
  const float u = (float)(dot3(tvec, pvec) * (double)inv_det);
  if (u < 0.0f || u > 1.0f) return 0;
  cross3(tvec, e1, qvec);
  const float v = (float)(dot3(r->d, qvec) * (double)inv_det);
  if (v < 0.0f || (u + v) > 1.0f) return 0;
  const float t = (float)(dot3(e2, qvec) * (double)inv_det);
  if ((double)t < tmin || (double)t > tmax) return 0;
  *t_out = (double)t;
  return 1;
}

/* ---- geom::xy_rect / xz_rect / yz_rect::Hit, src/geom/rect.h:19-40,65-85,109-130 -------- */
static int rect_hit(const double* p, const Ray* r, double tmin, double tmax, double* t_out) {
  const int axis = (int)p[0]; /* constant axis: 2 = xy_rect, 1 = xz_rect, 0 = yz_rect */
  const int ka = axis == 0 ? 1 : 0, kb = axis == 2 ? 1 : 2;
  const double t = (p[1] - r->o[axis]) / r->d[axis];
  if (!(tmin < t && t < tmax)) return 0;
  const double a = r->o[ka] + t * r->d[ka];
  const double b = r->o[kb] + t * r->d[kb];
  if (a < p[2] || a > p[3] || b < p[4] || b > p[5]) return 0;
  *t_out = t;
  return 1;
}

static int object_hit(const PortScene* s, int id, const Ray* r, double tmin, double tmax, double* t, int* gid);

/* ---- geom::Bvh::Hit, src/geom/bvh.h:71-119, with the primitive id carried along --------- */
static int bvh_hit(const PortScene* s, const Object* b, const Ray* r, double tmin, double tmax, double* t_out, int* gid_out) {
  if (b->n_nodes == 0) return 0; /* bvh.h:72-73 */
  int hit_anything = 0;
  double closest = tmax;
  int stack[64]; /* bvh.h:80 */
  int sp = 0;
  stack[sp++] = 0; /* root_index_ = 0: flatten() is pre-order (bvh.h:347-367) */
  while (sp > 0) {
    const double* n = b->nodes9 + 9 * (size_t)stack[--sp];
    if (!aabb_hit(n, r, tmin, closest)) continue;
    if (n[8] != 0.0) {
      const int first = (int)n[6], count = (int)n[7];
      for (int i = 0; i < count; ++i) {
        double t;
        int gid;
        if (object_hit(s, b->kids[b->prim_idx[first + i]], r, tmin, closest, &t, &gid)) {
          hit_anything = 1;
          closest = t;
          *gid_out = gid;
        }
      }
    } else {
      stack[sp++] = (int)n[7]; /* right, then left on top (bvh.h:111-114) */
      stack[sp++] = (int)n[6];
    }
  }
  if (hit_anything) *t_out = closest;
  return hit_anything;
}

/* ---- scene::Scene::Hit, src/scene/scene.h:47-61 ------------------------------------------ */
static int list_hit(const PortScene* s, const Object* l, const Ray* r, double tmin, double tmax, double* t_out, int* gid_out) {
  int hit_anything = 0;
  double closest = tmax;
  for (int i = 0; i < l->n_kids; ++i) {
    double t;
    int gid;
    if (object_hit(s, l->kids[i], r, tmin, closest, &t, &gid)) {
      hit_anything = 1;
      closest = t;
      *gid_out = gid;
    }
  }
  if (hit_anything) *t_out = closest;
  return hit_anything;
}

static int object_hit(const PortScene* s, int id, const Ray* r, double tmin, double tmax, double* t, int* gid) {
  const Object* o = &s->obj[id];
  if (o->kind == O_BVH) return bvh_hit(s, o, r, tmin, tmax, t, gid);
  if (o->kind == O_LIST) return list_hit(s, o, r, tmin, tmax, t, gid);
  int ok;
  if (o->prim == K_SPHERE) ok = sphere_hit(o->p, r, tmin, tmax, t);
  else if (o->prim == K_TRIANGLE) ok = triangle_hit(o->p, r, tmin, tmax, t);
  else ok = rect_hit(o->p, r, tmin, tmax, t);
  if (ok) *gid = o->gid;
  return ok;
}

/* ---- scene construction from the product's export stream (16 doubles per record) ---------- */
static int new_object(PortScene* s) {
  if (s->n_obj == s->cap) {
    s->cap = s->cap ? 2 * s->cap : 1024;
    s->obj = (Object*)realloc(s->obj, (size_t)s->cap * sizeof(Object));
  }
  memset(&s->obj[s->n_obj], 0, sizeof(Object));
  return s->n_obj++;
}

typedef struct { int* v; int n, cap; } IntVec;
static void push(IntVec* a, int x) {
  if (a->n == a->cap) {
    a->cap = a->cap ? 2 * a->cap : 16;
    a->v = (int*)realloc(a->v, (size_t)a->cap * sizeof(int));
  }
  a->v[a->n++] = x;
}

/* reads records until `end_kind`; children are appended to `dst` in creation order */
static void read_into(PortScene* s, const double* prims, size_t n, size_t* pos, IntVec* dst, int end_kind, const size_t* n_nodes) {
  while (*pos < n) {
    const double* r = prims + REC * (*pos)++;
    const int kind = (int)r[0];
    if (kind == end_kind) return;
    if (kind == K_SPHERE || kind == K_TRIANGLE || kind == K_RECT) {
      const int id = new_object(s);
      s->obj[id].kind = O_PRIM;
      s->obj[id].prim = kind;
      s->obj[id].gid = s->next_gid++;
      memcpy(s->obj[id].p, r + 2, 10 * sizeof(double));
      push(dst, id);
    } else if (kind == K_BVH_BEGIN || kind == K_LIST_BEGIN) {
      IntVec inner = {0, 0, 0};
      read_into(s, prims, n, pos, &inner, kind == K_BVH_BEGIN ? K_BVH_END : K_LIST_END, n_nodes);
      const int id = new_object(s);
      s->obj[id].kind = kind == K_BVH_BEGIN ? O_BVH : O_LIST;
      s->obj[id].kids = inner.v;
      s->obj[id].n_kids = inner.n;
      if (kind == K_BVH_BEGIN && s->next_bvh < s->n_bvh) { /* trees arrive in construction (post-) order */
        const int k = s->next_bvh++;
        s->obj[id].nodes9 = s->nodes[k];
        s->obj[id].n_nodes = n_nodes[k];
        s->obj[id].prim_idx = s->idx[k];
      }
      push(dst, id);
    }
  }
}

void* port_scene_create(const double* prims, size_t n_recs, int n_bvh, const double* const* nodes9, const size_t* n_nodes,
                        const int32_t* const* prim_idx, const size_t* n_idx) {
  PortScene* s = (PortScene*)calloc(1, sizeof(PortScene));
  s->n_bvh = n_bvh;
  s->nodes = (double**)calloc((size_t)n_bvh + 1, sizeof(double*));
  s->idx = (int32_t**)calloc((size_t)n_bvh + 1, sizeof(int32_t*));
  for (int k = 0; k < n_bvh; ++k) {
    s->nodes[k] = (double*)malloc(9 * n_nodes[k] * sizeof(double) + 8);
    memcpy(s->nodes[k], nodes9[k], 9 * n_nodes[k] * sizeof(double));
    s->idx[k] = (int32_t*)malloc(n_idx[k] * sizeof(int32_t) + 8);
    memcpy(s->idx[k], prim_idx[k], n_idx[k] * sizeof(int32_t));
  }
  IntVec top = {0, 0, 0};
  size_t pos = 0;
  read_into(s, prims, n_recs, &pos, &top, -1, n_nodes);
  /* the world is a scene::Scene list of the top-level objects (src/main.cc:60, scene.h:26-33) */
  const int id = new_object(s);
  s->obj[id].kind = O_LIST;
  s->obj[id].kids = top.v;
  s->obj[id].n_kids = top.n;
  s->root = id;
  return s;
}

void port_scene_free(void* h) {
  PortScene* s = (PortScene*)h;
  if (!s) return;
  for (int i = 0; i < s->n_obj; ++i) free(s->obj[i].kids);
  for (int k = 0; k < s->n_bvh; ++k) {
    free(s->nodes[k]);
    free(s->idx[k]);
  }
  free(s->nodes);
  free(s->idx);
  free(s->obj);
  free(s);
}

int port_prim_count(const void* h) { return ((const PortScene*)h)->next_gid; }

/* integrator::CPURayIntegrator::IntersectBatch, src/integrator/cpu_ray_integrator.h:18-46:
 * world_->Hit(ray, Interval(t_min, inf), rec) for every ray, OpenMP static.  rays8 = RtbRay
 * (o, tmin, d, tmax as float); the values are widened to double exactly as the reference sees
 * them.  gid = -1 and t = 0 on a miss. */
void port_intersect(const void* h, const float* rays8, size_t n, double* t_out, int32_t* gid_out, int threads) {
  const PortScene* s = (const PortScene*)h;
#ifdef _OPENMP
  if (threads > 0) omp_set_num_threads(threads);
#else
  (void)threads;
#endif
#pragma omp parallel for schedule(static)
  for (long long i = 0; i < (long long)n; ++i) {
    const float* q = rays8 + 8 * i;
    Ray r = {{q[0], q[1], q[2]}, {q[4], q[5], q[6]}};
    double t = 0.0;
    int gid = -1;
    if (!object_hit(s, s->root, &r, (double)q[3], (double)q[7], &t, &gid)) {
      t = 0.0;
      gid = -1;
    }
    t_out[i] = t;
    gid_out[i] = gid;
  }
}

/* ---- integrator::PixelState: RecordSample + Variance + IsConverged (src/integrator/pixel_state.h:13-72) ----------
 * Welford update per channel (pixel_state.h:26-36), sample variance m2/(n-1) (pixel_state.h:41-49), converged when
 * n >= min_spp and for every channel sqrt(var)/sqrt(n) / max(|mean|, 1e-3) <= rel_threshold (pixel_state.h:53-71).
 * Driven as WavefrontRenderer::Render drives it (src/renderer/wavefront.cc:120-127, 70-71): record, test, and a
 * converged pixel takes no further samples.  samples: n_seq x len x 3 doubles; stop_n[i] = samples recorded. */
void port_pixel_state_run(const double* samples, size_t n_seq, int len, double rel_threshold, int min_spp, int32_t* stop_n, double* sum,
                          double* m2_out) {
  for (size_t i = 0; i < n_seq; ++i) {
    double s[3] = {0, 0, 0}, mean[3] = {0, 0, 0}, m2[3] = {0, 0, 0};
    int n = 0, converged = 0;
    for (int k = 0; k < len && !converged; ++k) {
      const double* x = samples + (i * (size_t)len + (size_t)k) * 3;
      ++n;
      for (int c = 0; c < 3; ++c) {
        const double delta = x[c] - mean[c];
        mean[c] += delta / n;
        m2[c] += (x[c] - mean[c]) * delta;
        s[c] += x[c];
      }
      if (n >= min_spp) {
        int ok = 1;
        for (int c = 0; c < 3 && ok; ++c) {
          const double var = n > 1 ? m2[c] / (n - 1) : 0.0;
          const double mu = fmax(fabs(mean[c]), 1e-3);
          if (sqrt(var) / sqrt((double)n) / mu > rel_threshold) ok = 0;
        }
        converged = ok;
      }
    }
    stop_n[i] = n;
    for (int c = 0; c < 3; ++c) { sum[3 * i + c] = s[c]; m2_out[3 * i + c] = m2[c]; }
  }
}
