/* pm_scene.cuh — device-side scene view and BVH traversal (software replacement for optixTrace:
 * photon-mapping/cuda/deviceCode.cu:27,40; ray-tracer/cuda/deviceCode.cu:29-41,70-85,129-141).
 *
 * Layout in HBM (all 16-byte records, read with LDG.128):
 *   node  = 4 x float4  {c0.lo.x, c0.hi.x, c0.lo.y, c0.hi.y}
 *                       {c1.lo.x, c1.hi.x, c1.lo.y, c1.hi.y}
 *                       {c0.lo.z, c0.hi.z, c1.lo.z, c1.hi.z}
 *                       {bits(child0), bits(child1), -, -}   child >= 0: inner node, < 0: leaf ~child
 *   tri   = 3 x float4  {A.xyz, bits(global triangle id)} {B-A, bits(mesh id)} {C-A, -}
 * Triangle slots are in BVH (Morton) order; the ids they carry are the caller's triangle ids.
 *
 * The ray/triangle test is the canonical one of the CPU oracle (scene restatement) (Moeller-Trumbore, one rounding
 * per operation, both orientations, tmin < t < tmax, closest t wins, ties -> lowest triangle id).
 * Node boxes are padded (pm_scene.cu) and the slab test is NaN-safe, so the BVH only ever culls
 * triangles the canonical test would reject as well.
 */
#pragma once
#include "pm_internal.h"
#include "pm_devmath.cuh"

namespace pmd {

constexpr int kNoChild = 0x7fffffff;
constexpr int kStackDepth = 64;

struct SceneView {
  const float4* nodes;
  const float4* tris;
  const pm_material* mats;
  int nt;
};

struct Hit { int prim; float t; int slot; };

/* one 256-bit read-only load (sm_100: LDG.E.256) of two adjacent float4 — half a node record; p is 32-byte aligned */
__device__ __forceinline__ void ldg256(const float4* p, float4& a, float4& b) {
  asm("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
      : "=f"(a.x), "=f"(a.y), "=f"(a.z), "=f"(a.w), "=f"(b.x), "=f"(b.y), "=f"(b.z), "=f"(b.w)
      : "l"(p));
}

/* canonical triangle test; returns true and t on accept */
__device__ __forceinline__ bool tri_test(const float4 a4, const float4 e14, const float4 e24, vec3 org, vec3 dir,
                                         float tmin, float tmax, float* t_out) {
  const vec3 A = v3(a4.x, a4.y, a4.z), e1 = v3(e14.x, e14.y, e14.z), e2 = v3(e24.x, e24.y, e24.z);
  const vec3 p = cross(dir, e2);
  const float det = dot(e1, p);
  if (!(det != 0.f)) return false;
  const float inv = __fdiv_rn(1.f, det);
  const vec3 s = org - A;
  const float u = dot(s, p) * inv;
  if (!(u >= 0.f && u <= 1.f)) return false;
  const vec3 q = cross(s, e1);
  const float v = dot(dir, q) * inv;
  if (!(v >= 0.f && u + v <= 1.f)) return false;
  const float t = dot(e2, q) * inv;
  if (!(t > tmin && t < tmax)) return false;
  *t_out = t;
  return true;
}

/* Traversal state of one ray, resumable: k_walk leaves the loop when too few lanes of the warp are still
 * traversing, lets the finished lanes shade and start their next segment, and comes back. */
struct TraceState {
  int node; /* current inner node (>= 0), leaf (~slot, < 0) or kNoChild: finished / idle */
  int pend; /* trace_advance_warp: a leaf met on the way and not yet tested (< 0), or 0 */
  int sp;
  Hit best;
  float ix, iy, iz;
  int stack[kStackDepth];
};

__device__ __forceinline__ void trace_begin(const SceneView& sc, TraceState& st, vec3 dir, float tmax) {
  st.best = Hit{-1, tmax, -1};
  st.sp = 0;
  st.pend = 0;
  st.node = sc.nt > 0 ? 0 : kNoChild; /* the root is an inner node */
  st.ix = __fdiv_rn(1.f, dir.x); st.iy = __fdiv_rn(1.f, dir.y); st.iz = __fdiv_rn(1.f, dir.z);
}

/* closest hit (ANY = false) or occlusion (ANY = true: stops at the first accepted triangle).
 * "While-while" traversal: the inner loop descends through inner nodes only; a lane that reaches a leaf waits there
 * until the other lanes of its warp have one too (leaves are ordinary stack entries, encoded negative), then the
 * warp runs the triangle tests together. With the two kinds of work interleaved in one loop body a warp of
 * incoherent photon rays kept ~6 of 32 lanes busy (ncu, Cornell box); the visiting order does not affect the
 * result (closest t, ties -> lowest id; the slab test keeps boxes at exactly best.t).
 * Returns when st.node == kNoChild (result in st.best) or, if min_lanes > 0, as soon as fewer than min_lanes
 * lanes are still inside the loop (call again to continue). */
template <bool ANY>
__device__ __forceinline__ void trace_advance(const SceneView& sc, TraceState& st, vec3 org, vec3 dir, float tmin, float tmax,
                                              int min_lanes) {
  int node = st.node, sp = st.sp;
  Hit best = st.best;
  const float ix = st.ix, iy = st.iy, iz = st.iz;
  while (node != kNoChild) {
    while (node >= 0 && node != kNoChild) {
      float4 q0, q1, q2, q3;
      ldg256(sc.nodes + 4 * (size_t)node, q0, q1);
      ldg256(sc.nodes + 4 * (size_t)node + 2, q2, q3);
      const int c0 = __float_as_int(q3.x), c1 = __float_as_int(q3.y);
      /* slabs; fminf/fmaxf drop the NaN of 0 * inf */
      const float a0x = (q0.x - org.x) * ix, b0x = (q0.y - org.x) * ix;
      const float a0y = (q0.z - org.y) * iy, b0y = (q0.w - org.y) * iy;
      const float a0z = (q2.x - org.z) * iz, b0z = (q2.y - org.z) * iz;
      const float a1x = (q1.x - org.x) * ix, b1x = (q1.y - org.x) * ix;
      const float a1y = (q1.z - org.y) * iy, b1y = (q1.w - org.y) * iy;
      const float a1z = (q2.z - org.z) * iz, b1z = (q2.w - org.z) * iz;
      const float n0 = fmaxf(fmaxf(fminf(a0x, b0x), fminf(a0y, b0y)), fmaxf(fminf(a0z, b0z), tmin));
      const float f0 = fminf(fminf(fmaxf(a0x, b0x), fmaxf(a0y, b0y)), fminf(fmaxf(a0z, b0z), best.t));
      const float n1 = fmaxf(fmaxf(fminf(a1x, b1x), fminf(a1y, b1y)), fmaxf(fminf(a1z, b1z), tmin));
      const float f1 = fminf(fminf(fmaxf(a1x, b1x), fmaxf(a1y, b1y)), fminf(fmaxf(a1z, b1z), best.t));
      /* conservative: a couple of ulps of slack on the far side */
      const bool h0 = n0 <= f0 * 1.0000004f;
      const bool h1 = (n1 <= f1 * 1.0000004f) && c1 != kNoChild;
      if (h0 && h1) {
        const bool first0 = n0 <= n1;
        if (sp < kStackDepth) st.stack[sp++] = first0 ? c1 : c0;
        node = first0 ? c0 : c1;
      } else if (h0) {
        node = c0;
      } else if (h1) {
        node = c1;
      } else {
        node = sp ? st.stack[--sp] : kNoChild;
      }
    }
    while (node < 0) {
      const int slot = ~node;
      const float4 a4 = __ldg(sc.tris + 3 * (size_t)slot), e14 = __ldg(sc.tris + 3 * (size_t)slot + 1),
                   e24 = __ldg(sc.tris + 3 * (size_t)slot + 2);
      float t;
      node = sp ? st.stack[--sp] : kNoChild;
      if (tri_test(a4, e14, e24, org, dir, tmin, tmax, &t)) {
        const int id = __float_as_int(a4.w);
        if (best.prim < 0 || t < best.t || (t == best.t && id < best.prim)) { best.prim = id; best.t = t; best.slot = slot; }
        if (ANY) { node = kNoChild; sp = 0; }
      }
    }
    if (min_lanes > 0 && __popc(__activemask()) < min_lanes) break;
  }
  st.node = node; st.sp = sp; st.best = best;
}

/* The same traversal for a CONVERGED warp (all 32 lanes call it together; lanes without a ray pass st.node ==
 * kNoChild): the warp decides by vote what it does next. In the loop above a lane that has reached a leaf idles until the
 * slowest lane of the warp has one too — on the 1 M-triangle room the inner-node loop, 85 % of k_walk's instructions,
 * ran with 6.8 of 32 lanes. Here every step is either an inner-node step or a triangle test for the whole warp,
 * whichever more lanes can take; a lane that meets a leaf puts it aside (one pending leaf) and goes on with its stack,
 * so that most lanes can take either kind of step (the leaf is tested against a bound that may have shrunk meanwhile,
 * the result is the same: closest t, ties -> lowest id). With twice the lanes per instruction the L1 pipe becomes the
 * limit (85 %). Tried and dropped: the stack in shared memory (one column per thread) instead of local memory — 32 KB per
 * CTA cost two of eight CTAs per SM and the walk went from 42.8 to 44.8 ms. */
template <bool ANY>
__device__ __forceinline__ void trace_advance_warp(const SceneView& sc, TraceState& st, vec3 org, vec3 dir, float tmin, float tmax,
                                                   int min_lanes) {
  int node = st.node, sp = st.sp, pend = st.pend;
  Hit best = st.best;
  const float ix = st.ix, iy = st.iy, iz = st.iz;
  int busy_at_entry = -1;
  for (;;) {
    const bool can_inner = node >= 0 && node != kNoChild;
    const bool can_leaf = pend != 0 || node < 0;
    const unsigned m_inner = __ballot_sync(0xffffffffu, can_inner), m_leaf = __ballot_sync(0xffffffffu, can_leaf);
    const int busy = __popc(m_inner | m_leaf);
    if (busy_at_entry < 0) busy_at_entry = busy;
    /* leave when a lane has finished its ray since the call began and too few are left (the caller gives the finished
     * lanes new rays); never before something has finished: the caller could do nothing about it */
    if (busy == 0 || (busy < min_lanes && busy < busy_at_entry)) break;
    if (__popc(m_inner) >= __popc(m_leaf)) {
      if (can_inner) {
        float4 q0, q1, q2, q3; /* the 64-byte record in two 256-bit loads: half the L1 tag requests of four LDG.128 */
        ldg256(sc.nodes + 4 * (size_t)node, q0, q1);
        ldg256(sc.nodes + 4 * (size_t)node + 2, q2, q3);
        const int c0 = __float_as_int(q3.x), c1 = __float_as_int(q3.y);
        const float a0x = (q0.x - org.x) * ix, b0x = (q0.y - org.x) * ix;
        const float a0y = (q0.z - org.y) * iy, b0y = (q0.w - org.y) * iy;
        const float a0z = (q2.x - org.z) * iz, b0z = (q2.y - org.z) * iz;
        const float a1x = (q1.x - org.x) * ix, b1x = (q1.y - org.x) * ix;
        const float a1y = (q1.z - org.y) * iy, b1y = (q1.w - org.y) * iy;
        const float a1z = (q2.z - org.z) * iz, b1z = (q2.w - org.z) * iz;
        const float n0 = fmaxf(fmaxf(fminf(a0x, b0x), fminf(a0y, b0y)), fmaxf(fminf(a0z, b0z), tmin));
        const float f0 = fminf(fminf(fmaxf(a0x, b0x), fmaxf(a0y, b0y)), fminf(fmaxf(a0z, b0z), best.t));
        const float n1 = fmaxf(fmaxf(fminf(a1x, b1x), fminf(a1y, b1y)), fmaxf(fminf(a1z, b1z), tmin));
        const float f1 = fminf(fminf(fmaxf(a1x, b1x), fmaxf(a1y, b1y)), fminf(fmaxf(a1z, b1z), best.t));
        const bool h0 = n0 <= f0 * 1.0000004f;
        const bool h1 = (n1 <= f1 * 1.0000004f) && c1 != kNoChild;
        if (h0 && h1) {
          const bool first0 = n0 <= n1;
          if (sp < kStackDepth) st.stack[sp++] = first0 ? c1 : c0;
          node = first0 ? c0 : c1;
        } else if (h0) {
          node = c0;
        } else if (h1) {
          node = c1;
        } else {
          node = sp ? st.stack[--sp] : kNoChild;
        }
        if (node < 0 && pend == 0) { pend = node; node = sp ? st.stack[--sp] : kNoChild; } /* put the leaf aside */
      }
    } else if (can_leaf) {
      int leaf;
      if (pend != 0) { leaf = pend; pend = 0; }
      else { leaf = node; node = sp ? st.stack[--sp] : kNoChild; }
      const int slot = ~leaf;
      const float4 a4 = __ldg(sc.tris + 3 * (size_t)slot), e14 = __ldg(sc.tris + 3 * (size_t)slot + 1),
                   e24 = __ldg(sc.tris + 3 * (size_t)slot + 2);
      float t;
      if (tri_test(a4, e14, e24, org, dir, tmin, tmax, &t)) {
        const int id = __float_as_int(a4.w);
        if (best.prim < 0 || t < best.t || (t == best.t && id < best.prim)) { best.prim = id; best.t = t; best.slot = slot; }
        if (ANY) { node = kNoChild; sp = 0; pend = 0; }
      }
      if (node < 0 && pend == 0) { pend = node; node = sp ? st.stack[--sp] : kNoChild; }
    }
  }
  st.node = node; st.sp = sp; st.pend = pend; st.best = best;
}

template <bool ANY>
__device__ __forceinline__ Hit trace(const SceneView& sc, vec3 org, vec3 dir, float tmin, float tmax) {
  TraceState st;
  trace_begin(sc, st, dir, tmax);
  trace_advance<ANY>(sc, st, org, dir, tmin, tmax, 0);
  return st.best;
}

/* trace() for a converged warp: EVERY lane of the warp calls it, lanes without a ray with valid = false */
template <bool ANY>
__device__ __forceinline__ Hit trace_warp(const SceneView& sc, bool valid, vec3 org, vec3 dir, float tmin, float tmax) {
  TraceState st;
  trace_begin(sc, st, dir, tmax);
  if (!valid) st.node = kNoChild;
  trace_advance_warp<ANY>(sc, st, org, dir, tmin, tmax, 0);
  return st.best;
}

/* getPrimitiveNormal (common/cuda/helpers.h:76-85): normalize(cross(B-A, C-A)), never flipped */
__device__ __forceinline__ vec3 slot_normal(const SceneView& sc, int slot) {
  const float4 e14 = __ldg(sc.tris + 3 * (size_t)slot + 1), e24 = __ldg(sc.tris + 3 * (size_t)slot + 2);
  return normalize(cross(v3(e14.x, e14.y, e14.z), v3(e24.x, e24.y, e24.z)));
}
__device__ __forceinline__ int slot_mesh(const SceneView& sc, int slot) {
  return __float_as_int(__ldg(sc.tris + 3 * (size_t)slot + 1).w);
}

}  // namespace pmd

/* host-side handle */
struct pm_scene {
  pm_ctx* ctx = nullptr;   /* creator; never dereferenced on destroy (it may be gone already) */
  int device = 0;
  int nt = 0, nv = 0, n_mesh = 0;
  float4* nodes = nullptr;
  float4* tris = nullptr;
  pm_material* mats = nullptr;
  pmd::SceneView view() const { return pmd::SceneView{nodes, tris, mats, nt}; }
};
