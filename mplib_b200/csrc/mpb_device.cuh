// mpb_device.cuh -- device-side scene image and small fp32 algebra shared by the kernels.
//
// HBM layout of the compiled scene (all read-only during a batch):
//   DevJoint[n_joints]   placement (3x4), axis, type, parent, state slot / constant q
//   DevObject[n_objects] pose relative to its owner joint (link placement folded in on the host in
//                        double), geometry parameters, bounding sphere + OBB (object frame)
//   float4 verts[]       hull vertex pool, one float4 per vertex (xyz + pad) -> one LDS.128/LDG.128
//   DevPair[n_pairs]     (a, b, bin) admitted pairs; bin = index of the (geometry-key A, key B) class
// Work queues are structure-of-arrays so that both the producer (ballot-compacted lanes) and the
// consumer (lane = item) touch consecutive 4-byte words.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#define MPB_MAX_JOINTS 32   // movable joints in the merged forest (two Pandas = 18)
#define MPB_MAX_DYN 128     // objects that move with a joint (8-bit slot fields; fk_table_kernel's parameter table)
#define MPB_ITEM_WORDS 14   // 12 floats relative transform + cfg + pair
#define MPB_INNER_BALLS 4   // balls inscribed in every shape (sufficient-hit test in front of the narrowphase)

struct alignas(16) DevJoint {
  float P[12];   // placement in the parent joint frame
  float ax[3];
  int type;      // MPB_JOINT_*
  int parent;    // model index (0 = universe)
  int slot;      // index into the batch state, or -1
  float qconst;  // value when slot < 0 (q_base)
  int jslot;     // >= 0: this joint's world transform is kept in the shared-memory joint table (it owns moving
                 //       collision objects, or a later non-adjacent child needs it)
  int pjslot;    // >= 0: parent transform comes from that joint-table slot; -1: previous joint (or universe)
  int dyn_first; // first entry of SceneImg::dynlist owned by this joint
  int dyn_count; // number of moving collision objects on this joint
  int kind;      // MPB_JK_*: axis-aligned revolute joints skip the general Rodrigues product
};
static_assert(sizeof(DevJoint) % 16 == 0, "DevJoint must keep 16-byte alignment");

struct alignas(16) DevObject {
  float L[12];      // pose in the owner joint frame (joint == 0: world pose)
  int joint;        // model joint index, 0 = world-fixed
  int type;         // MPB_GEOM_*
  int geom;         // hull / mesh / octree index
  int dyn;          // slot in the per-thread dynamic-object table, -1 if world-fixed
  // bytes 64..96, one 32-byte sector: what the narrowphase reads per item besides the item record
  float prm[4];     // sphere r | box half sides | capsule, cylinder: r, half height
  float ctr[3];     // MPR centre: hull interior point (vertex mean) or 0
  float rin;        // radius of a ball around ctr that lies inside the shape (sufficient-hit test)
  float bc[3];      // bounding-sphere centre, object frame
  float br;         // bounding-sphere radius
  float oc[3];      // OBB centre, object frame
  float oh[3];      // OBB half extents
  float oR[9];      // OBB axes as columns of a rotation (object frame)
  int voff, vcnt;   // hull vertices in the pool
  float bcw[3];     // world-fixed objects: bounding-sphere centre in the world frame
  int jslot;        // joint-table slot of the owner joint, -1 if world-fixed
  float bcj[3];     // bounding-sphere centre in the owner joint frame (L * bc)
  int cell;         // convex hulls: byte offset of the support cell table behind the vertex pool, -1 = none
  float4 ball[MPB_INNER_BALLS];  // balls inside the shape, object frame: centre xyz, radius w (0 = unused); sufficient-hit test
};
static_assert(sizeof(DevObject) % 32 == 0, "DevObject records must start on 32-byte sectors");

// Support cell table of a convex hull: the directions of a cube map (MPB_CELL_G x MPB_CELL_G cells per face) -> the
// vertices that can be extreme for SOME direction of the cell (a conservative superset, built on the host), so the
// support scan visits a handful of candidates instead of every vertex and still returns the exact arg-max.
// Layout: uint16 first[6 G G + 1], then uint8 candidate vertex indices (ascending per cell).
#define MPB_CELL_G 8
#define MPB_CELL_COUNT (6 * MPB_CELL_G * MPB_CELL_G)
__host__ __device__ __forceinline__ int mpb_cube_cell(float x, float y, float z) {
  const float ax = fabsf(x), ay = fabsf(y), az = fabsf(z);
  int face;
  float m, u, v;
  if (ax >= ay && ax >= az) { face = x >= 0.f ? 0 : 1; m = ax; u = y; v = z; }
  else if (ay >= az) { face = y >= 0.f ? 2 : 3; m = ay; u = x; v = z; }
  else { face = z >= 0.f ? 4 : 5; m = az; u = x; v = y; }
  const float inv = m > 0.f ? 1.0f / m : 0.f;
  int iu = (int)((u * inv + 1.0f) * (0.5f * MPB_CELL_G)), iv = (int)((v * inv + 1.0f) * (0.5f * MPB_CELL_G));
  iu = iu < 0 ? 0 : (iu > MPB_CELL_G - 1 ? MPB_CELL_G - 1 : iu);
  iv = iv < 0 ? 0 : (iv > MPB_CELL_G - 1 ? MPB_CELL_G - 1 : iv);
  return (face * MPB_CELL_G + iv) * MPB_CELL_G + iu;
}

struct DevPair {
  int a, b;
  int bin_dyn;  // bin | (dynA + 1) << 16 | (dynB + 1) << 24   (0 in a dyn field = world-fixed)
  float rr;     // sum of the two bounding-sphere radii (+ margin)
};
#define MPB_JK_GENERIC 0  // revolute / continuous about an arbitrary unit axis
#define MPB_JK_RX 1       // revolute about +-x (sign in ax[0])
#define MPB_JK_RY 2
#define MPB_JK_RZ 3
#define MPB_JK_PRISMATIC 4

// internal leaf-primitive shape types of BVH objects (meshes, octrees)
#define MPB_GEOM_TRI 7    // triangle: prims[3*aux + 0..2] = vertices (object frame)
#define MPB_GEOM_VBOX 8   // box cell: prims[3*aux] = centre, prims[3*aux+1] = half extents (object frame)

// AABB tree over the leaf primitives of one mesh / octree, in the object's frame.
struct alignas(16) DevBvhNode {
  float c[3];
  int left;    // inner: index of the left child;   leaf: ~first_prim (negative)
  float h[3];
  int right;   // inner: index of the right child;  leaf: primitive count
};

struct Xf {  // rigid transform, row-major rotation
  float r[9];
  float p[3];
};

__device__ __forceinline__ Xf xf_mul(const Xf &a, const Xf &b) {
  Xf o;
#pragma unroll
  for (int i = 0; i < 3; i++) {
#pragma unroll
    for (int j = 0; j < 3; j++)
      o.r[3 * i + j] = fmaf(a.r[3 * i], b.r[j], fmaf(a.r[3 * i + 1], b.r[3 + j], a.r[3 * i + 2] * b.r[6 + j]));
    o.p[i] = fmaf(a.r[3 * i], b.p[0], fmaf(a.r[3 * i + 1], b.p[1], fmaf(a.r[3 * i + 2], b.p[2], a.p[i])));
  }
  return o;
}
// a^-1 * b
__device__ __forceinline__ Xf xf_inv_mul(const Xf &a, const Xf &b) {
  Xf o;
  float d0 = b.p[0] - a.p[0], d1 = b.p[1] - a.p[1], d2 = b.p[2] - a.p[2];
#pragma unroll
  for (int i = 0; i < 3; i++) {
#pragma unroll
    for (int j = 0; j < 3; j++)
      o.r[3 * i + j] = fmaf(a.r[i], b.r[j], fmaf(a.r[3 + i], b.r[3 + j], a.r[6 + i] * b.r[6 + j]));
    o.p[i] = fmaf(a.r[i], d0, fmaf(a.r[3 + i], d1, a.r[6 + i] * d2));
  }
  return o;
}
__device__ __forceinline__ Xf xf_from12(const float *m) {
  Xf o;
#pragma unroll
  for (int i = 0; i < 3; i++) {
    o.r[3 * i] = m[4 * i];
    o.r[3 * i + 1] = m[4 * i + 1];
    o.r[3 * i + 2] = m[4 * i + 2];
    o.p[i] = m[4 * i + 3];
  }
  return o;
}
__device__ __forceinline__ void xf_point(const Xf &t, const float *v, float *o) {
#pragma unroll
  for (int i = 0; i < 3; i++) o[i] = fmaf(t.r[3 * i], v[0], fmaf(t.r[3 * i + 1], v[1], fmaf(t.r[3 * i + 2], v[2], t.p[i])));
}

struct V3 {
  float x, y, z;
};
__device__ __forceinline__ V3 v3(float x, float y, float z) { return V3{x, y, z}; }
__device__ __forceinline__ V3 operator-(V3 a, V3 b) { return V3{a.x - b.x, a.y - b.y, a.z - b.z}; }
__device__ __forceinline__ V3 operator+(V3 a, V3 b) { return V3{a.x + b.x, a.y + b.y, a.z + b.z}; }
__device__ __forceinline__ V3 operator*(V3 a, float s) { return V3{a.x * s, a.y * s, a.z * s}; }
__device__ __forceinline__ float dot(V3 a, V3 b) { return fmaf(a.x, b.x, fmaf(a.y, b.y, a.z * b.z)); }
__device__ __forceinline__ V3 cross(V3 a, V3 b) {
  return V3{a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x};
}
__device__ __forceinline__ V3 normalized(V3 a) { return a * rsqrtf(dot(a, a)); }
