// rtb_pool.cuh — an alternative render kernel (RTB_RENDER_POOL_PATHS): persistent warps over a per-warp POOL of paths in shared memory.
// MEASURED AND NOT ADOPTED (round 2): same paths and image as path_kernel, but 16-58 % slower - see the end of this comment.
//
// path_kernel (rtb_kernels.cuh) gave every lane one path and shaded a finished ray in place, with whatever lanes happened
// to be waiting in the same service phase: profiles/r01_path_kernel_bunny_packed_slab.txt shows the price - 14.6 of 32
// lanes per executed instruction, node steps at 15, shading at 9, because a lane whose traversal is over idles until 16
// lanes have collected, and the 300-instruction shading code then runs for those 16 (or fewer) while the tracing lanes wait.
//
// Here a warp owns kPoolSlots (64) paths: up to 32 are being traced, one per lane, the others sit in the warp's pool in
// shared memory either READY (a ray waiting for a lane) or FINISHED (a closest hit waiting to be shaded).
//   swap   a lane whose traversal is over stores (t, ref) into its slot - 2 words - and takes the next READY ray: a dozen
//          instructions, so it runs as soon as a few lanes wait and the node phase keeps (almost) all lanes tracing;
//   shade  only when the READY rays run out, and then for 32 FINISHED slots at once: every lane shades one path (the
//          reference's Camera::GetPixel bounce, shade_path), writes the bounced ray back into the slot or replaces a dead
//          path by the next camera sample from the global cursor.  Shading runs on full warps instead of 9 lanes.
// In the steady state (all 64 slots hold live paths) READY can only run out when at least 32 slots are FINISHED, so the
// shade phase is always full; fewer lanes shade only while the render drains.  The paths, the Philox streams and therefore the
// image are those of path_kernel and of the wavefront kernels (float atomics in another order).
//
// What the B200 said (profiles/r02_pool_kernel_bunny.txt against r01_path_kernel_bunny_packed_slab.txt, bunny C3): 12.4 ms
// instead of 9.3.  Tracing lanes per warp did go up (22 -> 28) but the lanes standing at an internal pair did not (15.3 ->
// 15.2): the node phase is bounded by the node / leaf alternation of the while-while loop, not by idle lanes, so the extra
// lanes only wait at leaves.  Shading reached 13 lanes, not 32: what diverges there is sky / material type, not occupancy.
// And the pool costs: 30 KB of shared memory per CTA shrink the L1 (global-load hit rate 54 -> 40 %), ptxas moves its
// spills into the node loop (local loads 114 M -> 486 M per step), issue-slot use falls 64 -> 52 %.  Variants (swap
// threshold 4 / 10 / 16, 48 / 96 slots, 3 CTAs x 80 registers, 1 node step per vote, exit at 1/2): all within 12.4-14.2 ms
// (profiles/r02_pool_variants.jsonl).
#pragma once
#include "rtb_kernels.cuh"

namespace rtb {

#ifndef RTB_POOL_SLOTS
#define RTB_POOL_SLOTS 64  // paths per warp: 32 in flight + the backlog (a multiple of 32 keeps the shade phase a full warp)
#endif
#ifndef RTB_POOL_SWAP_LANES
#define RTB_POOL_SWAP_LANES 6  // lanes that must be waiting (traversal over, or no ray) before the swap phase runs
#endif
#ifndef RTB_POOL_MIN_BLOCKS
#define RTB_POOL_MIN_BLOCKS 4
#endif
constexpr int kPoolSlots = RTB_POOL_SLOTS;
constexpr uint32_t kEmptySlot = 0xfffffffeu;  // F_REF of a slot that holds no path (kNoHit = a path that missed)
enum { F_OX, F_OY, F_OZ, F_DX, F_DY, F_DZ, F_TR, F_TG, F_TB, F_PIXEL, F_SAMPLE, F_BOUNCE, F_T, F_REF, F_ITEM, kPoolFields };
constexpr int kPoolQueueWords = 2 * kPoolSlots / 4;  // ready and finished queues: one byte per entry
template <bool kAdaptive>
struct PoolLayout {
  static constexpr int kFields = kAdaptive ? kPoolFields : kPoolFields - 1;  // F_ITEM only exists for adaptive passes
  static constexpr int kWarpWords = kFields * kPoolSlots + kPoolQueueWords;
};

// One lane's traversal state and the two steps of trace_loop (rtb_kernels.cuh), unchanged: sibling-pair node step with
// the packed-FP32 slab test, near child first, far child pushed; one primitive of a leaf per leaf step.
struct TravLane {
  TravRay r;
  Best b;
  uint32_t cur;
  int sp;
  TravStack stack;
#ifdef RTB_COUNTERS
  unsigned n_pairs = 0, n_sph = 0, n_tri = 0, n_rect = 0, n_rays = 0;
#endif
  __device__ __forceinline__ void pop_or_finish() {
    if (sp > 0) cur = stack.at(--sp);
    else cur = kDoneMark;
  }
  __device__ __forceinline__ void node_step(unsigned long long base_all, uint32_t n_pairs) {
    RTB_CHECK(kChkPair, cur < n_pairs);
    (void)n_pairs;
    const unsigned long long addr = base_all + (unsigned long long)cur * 64ull;
    f32x2 c0[4], c1[4];
    ld256_global(addr, c0);
    ld256_global(addr + 32, c1);
#ifdef RTB_COUNTERS
    ++n_pairs;
#endif
    float tl, tr;
    const bool hl = slab(r, c0[0], c0[1], c0[2], b.hi, tl);
    const bool hr = slab(r, c0[3], c1[0], c1[1], b.hi, tr);
    const uint32_t dl = (uint32_t)c1[2], dr = (uint32_t)(c1[2] >> 32);
    if (hl && hr) {
      const bool left_first = tl <= tr;
      stack.push(sp++, left_first ? dr : dl);
      cur = left_first ? dl : dr;
    } else if (hl || hr) {
      cur = hl ? dl : dr;
    } else {
      pop_or_finish();
    }
  }
  __device__ __forceinline__ void leaf_step(const DevScene& s, uint32_t type, HitCounters* cnt) {
#ifdef RTB_COUNTERS
    if (type == RTB_PRIM_TRIANGLE) ++n_tri; else if (type == RTB_PRIM_SPHERE) ++n_sph; else ++n_rect;
#endif
    consider(s, r, type, desc_first(cur), b, cnt);
    if (desc_more(cur)) cur += 1u - (1u << 24);  // first + 1, count - 1
    else pop_or_finish();
  }
  __device__ __forceinline__ void start(const DevScene& s, float ox, float oy, float oz, float dx, float dy, float dz) {
    prepare_ray(r, make_float4(ox, oy, oz, 0.001f), make_float4(dx, dy, dz, __int_as_float(0x7f800000)));  // Interval(0.001, inf), camera.h:158
    b.ref = kNoHit; b.tf = r.tmax; b.ef = 0.f; b.hi = r.tmax;
    cur = s.n_pairs ? 0u : kDoneMark;  // pair 0 = the root's two children; empty scene: Bvh::Hit returns false (bvh.h:72-73)
    sp = 0;
#ifdef RTB_COUNTERS
    ++n_rays;
#endif
  }
};

// Where the camera samples (or queued paths) of a launch come from, and where finished radiance goes.
template <bool kFromQueue, bool kAdaptive>
struct PoolIo {
  uint32_t n;
  const RenderConsts& rc;
  float* accum;
  PathQueue q;
  uint32_t s0, p0;  // divmod(first sample of this launch, n_pixels)
  const uint32_t* pixel_order;
  AdaptivePlanes ad;
  // work item i -> the path's first ray and state
  __device__ __forceinline__ void load(uint32_t i, float4& o, float4& d, PathState& ps) const {
    if (kFromQueue) {
      o = q.ro[i];
      d = q.rd[i];
      const float4 th = q.thr[i];
      const uint2 meta = q.meta[i];
      ps.tr = th.x; ps.tg = th.y; ps.tb = th.z;
      ps.pixel = __float_as_uint(th.w);
      ps.sample = meta.x;
      ps.bounce = meta.y;
    } else if (kAdaptive) {
      const uint32_t slot = i / ad.chunk_r, k = i - slot * ad.chunk_r;
      camera_ray(rc, __ldg(&ad.active[slot]), (uint32_t)rc.spp_begin + ad.base + k, o, d, ps);
    } else if (rc.tile_major) {
      generate_sample_tiled(rc, pixel_order, s0, p0, i, o, d, ps);
    } else {
      generate_sample(rc, pixel_order, s0, p0, i, o, d, ps);
    }
  }
};

template <bool kFromQueue, bool kAdaptive>
__device__ __forceinline__ void pool_loop(const DevScene& s, const RenderConsts& rc, const PoolIo<kFromQueue, kAdaptive>& io, uint32_t* wp,
                                          unsigned int* cursor, HitCounters* cnt, unsigned& rays_out) {
  using L = PoolLayout<kAdaptive>;
  const unsigned lane = threadIdx.x & 31u;
  const unsigned lt_mask = (1u << lane) - 1u;
  uint8_t* ready_q = reinterpret_cast<uint8_t*>(wp + L::kFields * kPoolSlots);
  uint8_t* fin_q = ready_q + kPoolSlots;
  auto fld = [&](int f, uint32_t slot) -> uint32_t& { return wp[f * kPoolSlots + slot]; };

  static_assert(kSmemStack == 0, "the pool kernel keeps the traversal stack in local memory");
  TravLane t;
  t.stack.sm = nullptr;
  t.cur = kDoneMark;
  t.sp = 0;
  t.b = Best{kNoHit, 0.f, 0.f, 0.f};
  uint32_t slot = 0;
  bool active = false;
  unsigned rays = 0;
  // warp-uniform bookkeeping
  int n_ready = 0, n_fin = kPoolSlots;
  unsigned pool_next = 0, pool_end = 0;  // claimed, not yet assigned work items
  bool exhausted = false;                 // the global cursor ran past n
  const unsigned long long base_all = (unsigned long long)s.pairs;

  // every slot starts FINISHED and empty: the first shade phases fill them with camera samples
  for (int k = lane; k < kPoolSlots; k += 32) {
    fld(F_REF, k) = kEmptySlot;
    fin_q[k] = (uint8_t)k;
  }
  __syncwarp();

  // ---- shade phase: up to 32 FINISHED slots, one per lane -> READY slots (bounced ray or a new camera sample)
  auto shade_phase = [&]() {
    const int m = min(32, n_fin);
    const bool part = (int)lane < m;
    const uint32_t sl = part ? fin_q[n_fin - 1 - (int)lane] : 0u;
    n_fin -= m;
    bool have = false;
    float4 ro, rd;
    PathState ps;
    uint32_t item = 0;
    if (part) {
      const uint32_t ref = fld(F_REF, sl);
      if (ref != kEmptySlot) {
        ro = make_float4(__uint_as_float(fld(F_OX, sl)), __uint_as_float(fld(F_OY, sl)), __uint_as_float(fld(F_OZ, sl)), 0.001f);
        rd = make_float4(__uint_as_float(fld(F_DX, sl)), __uint_as_float(fld(F_DY, sl)), __uint_as_float(fld(F_DZ, sl)), __int_as_float(0x7f800000));
        ps.tr = __uint_as_float(fld(F_TR, sl)); ps.tg = __uint_as_float(fld(F_TG, sl)); ps.tb = __uint_as_float(fld(F_TB, sl));
        ps.pixel = fld(F_PIXEL, sl); ps.sample = fld(F_SAMPLE, sl); ps.bounce = fld(F_BOUNCE, sl);
        const float hit_t = __uint_as_float(fld(F_T, sl));
        if (!kAdaptive) {
          AtomicSink sink{io.accum};
          have = shade_path(s, rc, sink, ro, rd, ps, hit_t, ref);
        } else {
          item = fld(F_ITEM, sl);
          LocalSink sink{mk3(0.0f, 0.0f, 0.0f)};
          have = shade_path(s, rc, sink, ro, rd, ps, hit_t, ref);
          if (!have) {  // work item = (active pixel, sample of the chunk): this sample's radiance (0 for an absorbed path)
            float* c = io.ad.chunk + 3 * (size_t)item;
            c[0] = sink.L.x; c[1] = sink.L.y; c[2] = sink.L.z;
          }
        }
      }
    }
    // dead (or never filled) slots take the next work items of the launch: 32 per atomicAdd on the global cursor
    unsigned need = __ballot_sync(kFull, part && !have);
    for (int round = 0; round < 2 && need; ++round) {
      if (pool_next == pool_end) {
        if (exhausted) break;
        unsigned base = 0;
        if (lane == 0) base = atomicAdd(cursor, kChunk);
        base = __shfl_sync(kFull, base, 0);
        if (base >= io.n) {
          exhausted = true;
          break;
        }
        pool_next = base;
        pool_end = min(base + kChunk, io.n);
        exhausted = pool_end == io.n;
      }
      const unsigned mine = pool_next + __popc(need & lt_mask);
      if (part && !have && mine < pool_end) {
        io.load(mine, ro, rd, ps);
        item = mine;
        have = true;
      }
      pool_next = min(pool_next + (unsigned)__popc(need), pool_end);
      need = __ballot_sync(kFull, part && !have);
    }
    const unsigned got = __ballot_sync(kFull, have);
    if (have) {
      fld(F_OX, sl) = __float_as_uint(ro.x); fld(F_OY, sl) = __float_as_uint(ro.y); fld(F_OZ, sl) = __float_as_uint(ro.z);
      fld(F_DX, sl) = __float_as_uint(rd.x); fld(F_DY, sl) = __float_as_uint(rd.y); fld(F_DZ, sl) = __float_as_uint(rd.z);
      fld(F_TR, sl) = __float_as_uint(ps.tr); fld(F_TG, sl) = __float_as_uint(ps.tg); fld(F_TB, sl) = __float_as_uint(ps.tb);
      fld(F_PIXEL, sl) = ps.pixel; fld(F_SAMPLE, sl) = ps.sample; fld(F_BOUNCE, sl) = ps.bounce;
      if (kAdaptive) fld(F_ITEM, sl) = item;
      ready_q[n_ready + __popc(got & lt_mask)] = (uint8_t)sl;
    }
    n_ready += __popc(got);
    __syncwarp();
  };

  for (;;) {
    const bool done = active && t.cur == kDoneMark;
    const unsigned wait_mask = __ballot_sync(kFull, !active || done);
    const unsigned done_mask = __ballot_sync(kFull, done);
    int n_trace = 32 - __popc(wait_mask);
    const bool can_feed = n_ready > 0 || n_fin > 0;  // a waiting lane can be given a ray (a shade phase may still come up empty)
    if ((done_mask != 0u || (can_feed && wait_mask != 0u)) && (__popc(wait_mask) >= RTB_POOL_SWAP_LANES || n_trace == 0)) {
      // ---- swap phase: finished traversals leave their hit in the slot, waiting lanes take READY rays
      if (done) {
        fld(F_T, slot) = __float_as_uint(t.b.tf);
        fld(F_REF, slot) = t.b.ref;
        fin_q[n_fin + __popc(done_mask & lt_mask)] = (uint8_t)slot;
        active = false;
      }
      n_fin += __popc(done_mask);
      __syncwarp();
      const int n_want = __popc(wait_mask);
      while (n_ready < n_want && n_fin > 0) shade_phase();
      const int rank = __popc(wait_mask & lt_mask);
      if (!active && rank < n_ready) {
        slot = ready_q[n_ready - 1 - rank];
        RTB_CHECK(kChkPoolSlot, slot < (uint32_t)kPoolSlots);
        t.start(s, __uint_as_float(fld(F_OX, slot)), __uint_as_float(fld(F_OY, slot)), __uint_as_float(fld(F_OZ, slot)),
                __uint_as_float(fld(F_DX, slot)), __uint_as_float(fld(F_DY, slot)), __uint_as_float(fld(F_DZ, slot)));
        active = true;
        ++rays;
      }
      n_ready -= min(n_ready, n_want);
      __syncwarp();
      continue;  // (an empty scene finishes a ray at start(): it is swapped out again right away)
    }
    if (n_trace == 0) break;  // nobody tracing, nothing READY, nothing FINISHED (so nothing left to claim either)
    // ---- node phase: descend while the descending lanes are at least 1/4 of the tracing ones (trace_loop)
    for (;;) {
      const bool at_node = active && !(t.cur & kLeafBit);
      const unsigned nm = __ballot_sync(kFull, at_node);
      if (nm == 0u || RTB_NODE_EXIT_DEN * __popc(nm) < RTB_NODE_EXIT_NUM * n_trace) break;
      if (at_node) t.node_step(base_all, s.n_pairs);
#pragma unroll
      for (int extra = 1; extra < RTB_NODE_STEPS_PER_VOTE; ++extra)
        if (active && !(t.cur & kLeafBit)) t.node_step(base_all, s.n_pairs);
    }
    // ---- leaf phase, one primitive type at a time
#pragma unroll 1
    for (uint32_t type = 0; type < 3; ++type) {
      const uint32_t want = 4u + type;  // leaf bit + type in the top three bits
      while (__any_sync(kFull, active && (t.cur >> 29) == want)) {
        if (active && (t.cur >> 29) == want) t.leaf_step(s, type, cnt);
      }
    }
  }
  rays_out = rays;
#ifdef RTB_COUNTERS
  if (cnt) {
    atomicAdd(&cnt->node_pairs, (unsigned long long)t.n_pairs);
    atomicAdd(&cnt->sphere_tests, (unsigned long long)t.n_sph);
    atomicAdd(&cnt->triangle_tests, (unsigned long long)t.n_tri);
    atomicAdd(&cnt->rect_tests, (unsigned long long)t.n_rect);
    atomicAdd(&cnt->rays, (unsigned long long)t.n_rays);
  }
#endif
}

template <bool kFromQueue, bool kAdaptive = false>
__global__ void __launch_bounds__(kTraceBlock, RTB_POOL_MIN_BLOCKS) pool_kernel(DevScene s, PathQueue q, uint32_t n, uint32_t s0, uint32_t p0, RenderConsts rc,
                                                                               const uint32_t* __restrict__ pixel_order, float* accum,
                                                                               unsigned int* cursor, unsigned long long* ray_count, HitCounters* cnt,
                                                                               AdaptivePlanes ad) {
  __shared__ uint32_t pool[(kTraceBlock / 32) * PoolLayout<kAdaptive>::kWarpWords];
  if (kAdaptive) n = *ad.n_active * ad.chunk_r;  // one work item per (active pixel, sample of the chunk)
  const PoolIo<kFromQueue, kAdaptive> io{n, rc, accum, q, s0, p0, pixel_order, ad};
  unsigned rays = 0;
  pool_loop<kFromQueue, kAdaptive>(s, rc, io, pool + (threadIdx.x >> 5) * PoolLayout<kAdaptive>::kWarpWords, cursor, cnt, rays);
  const unsigned total = __reduce_add_sync(kFull, rays);
  if ((threadIdx.x & 31u) == 0u && total) atomicAdd(ray_count, (unsigned long long)total);
  if (kAdaptive && blockIdx.x == 0 && threadIdx.x == 0) atomicAdd(ray_count + 1, (unsigned long long)n);  // camera samples started
}

}  // namespace rtb
