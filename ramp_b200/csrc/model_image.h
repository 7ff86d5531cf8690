// model_image.h -- layout of the packed, device-resident model ("image").
//
// One contiguous, 16-byte aligned blob lives in HBM per RampModel.  Every CTA copies it
// into shared memory once with a single TMA bulk copy (cp.async.bulk + mbarrier) and all
// table reads on the hot path are shared-memory broadcasts.  Shared by the host packer
// and the kernels in ramp_b200.cu.
#pragma once
#include <stddef.h>
#include <stdint.h>

namespace ramp {

constexpr int kMaxJoints = 64;
constexpr int kMaxGeoms = 128;
constexpr int kMaxPairs = 2048;
constexpr int kMaxNq = 128;
constexpr int kMaxSubspaces = 48;
constexpr int kLanes = 32;             // configurations per warp tile
constexpr int kChunkPairs = 32;        // pairs culled per queue round
// FK plan table (int32): [kPlanTeam2 .. +2] and [kPlanTeam4 .. +4] are entry ranges of the warps of
// a team of 2 / 4, entries start at kPlanEntries; an entry is joint | (places its geometries << 16)
constexpr int kPlanTeam2 = 0, kPlanTeam4 = 3, kPlanEntries = 8;
constexpr int kQueueCap = kChunkPairs * kLanes;

// parent_code of a joint
constexpr int kParentUniverse = -1;  // child of the universe: oMi = liMi
constexpr int kParentRunning = -2;   // parent is the previous joint: use the register copy

// Rotation part of a joint placement.  URDF / MJCF joint origins are mostly pure translations or
// quarter / half turns about one axis (rpy multiples of pi/2): parent * placement is then a signed
// shuffle of the parent's columns plus the translation, a third of the full 3x4 product.
enum PlacementKind : int32_t {
  kPlGeneral = 0, kPlIdentity = 1,
  kPlXp = 2, kPlXm = 3, kPlXh = 4,  // about x by +90, -90, 180 degrees
  kPlYp = 5, kPlYm = 6, kPlYh = 7,
  kPlZp = 8, kPlZm = 9, kPlZh = 10
};

// narrow-phase class of a pair (decided on the host from the two shape types)
enum PairClass : int32_t {
  kClsInactive = 0,
  kClsSphereSphere = 1,   // classes 1..5: closed-form distance
  kClsSphereCapsule = 2,
  kClsSphereBox = 3,
  kClsSphereCylinder = 4,
  kClsCapsuleCapsule = 5,
  kClsBoxBox = 6,         // exact separating-axis test
  kClsGjk = 7             // OBB cull, then GJK on the cores
};

struct ImgHeader {  // 32 words
  int32_t total_bytes, nq, n_joints, n_geoms, n_dyn, n_static, n_pairs, n_verts;
  int32_t n_act, n_mimic, n_cont, n_save, mask_words, n_robot_geoms, n_robot_pairs, n_dd;  // n_dd: DD pairs
  int32_t off_joints, off_geoms, off_pairs, off_verts, off_static, off_dyn, off_act, off_mimic;
  int32_t off_cont, off_refq, ss_begin, off_plan, reserved1[4];  // ss_begin: first pair of two universe-attached geometries
};
static_assert(sizeof(ImgHeader) == 128, "header is 32 words");

struct DJoint {  // 24 words
  int32_t type, parent_code, idx_q, save_slot;
  float ax, ay, az;
  int32_t gfirst, gcount;  // range in the dyn-geometry order list
  int32_t parent;          // plain parent index (-1: universe)
  int32_t nq;              // q entries the joint reads
  int32_t pkind;           // PlacementKind of pl's rotation: quarter turns about one axis need no multiplications
  float pl[12];
};
static_assert(sizeof(DJoint) == 96, "DJoint");

struct DGeom {  // 40 words
  int32_t type, gslot, voff, vcnt;  // gslot >= 0: per-config slot; < 0: static transform -1-gslot
  float p0, p1, p2, bound_r;
  float hx, hy, hz;  // local-frame box that contains the shape (OBB cull)
  int32_t parent;
  float pl[12];  // placement in the parent joint frame (dyn) / world (static)
  // Two capsules around the local z axis, centred on the local origin (the packer re-frames
  // every geometry that way): the OUTER one contains the shape (cull), the INNER one is
  // contained in it (early accept); half lengths / radii.
  float s_out, r_out, s_in, r_in;
  float opl[12];  // the caller's placement, before re-framing: what ramp_placements_batch reports
};
static_assert(sizeof(DGeom) == 160, "DGeom");
static_assert(offsetof(DJoint, pl) % 16 == 0 && offsetof(DGeom, pl) % 16 == 0 && offsetof(DGeom, opl) % 16 == 0,
              "placements are read as float4");

// Pairs are stored in PROCESSING order: first the pairs of two joint-attached geometries
// ("DD"), then those of a joint-attached and a universe-attached one ("DS") -- each stage-1
// loop then addresses placements with constant strides.  `orig` is the index in the caller's
// list (GeometryModel.collisionPairs order), which is what every result reports.
struct DPair {  // 12 words = three 16-byte records
  // record 0: everything stage 1 reads
  int32_t o1, o2;     // byte offsets of the two placement origins: into a lane's slab column
                      // (joint-attached) or into the static table (universe-attached)
  float rs2;          // (bound_r(a)+bound_r(b))^2, slightly inflated; -1 for an inactive pair
  int16_t orig, cls;  // index in the caller's pair list; narrow-phase class
  // record 1: stage 1.5, the outer capsules of the geometries at o1 / o2 (GJK class)
  float h1_out, h2_out;  // half lengths
  float rsum_out;        // radius sum (+ slack); +inf for the other classes, which always pass
  int32_t kind;          // 0: o1 and o2 address the slab, 1: o1 slab / o2 static, 2: both static
  // record 2
  float slack;        // widening of every conservative test: 1e-5 m + the pair's share of non-rigid placements
  int16_t a, b;       // geometry indices, ordered so that type(a) <= type(b)
  int16_t sa, sb;     // their gslot values (>= 0 slab slot, < 0 static placement)
  int32_t pad;
};
static_assert(sizeof(DPair) == 48, "DPair");

struct DMimic {
  int32_t dst, src;
  float mult, off;
};

struct DSub {  // one OMPL subspace, by value in the edge kernels' parameter block
  int32_t type, col, dim, pad;
  float lo[3], hi[3];
  float inv_lvs_real;  // 1 / (0.01 * extent) of the RealVector part
  float inv_lvs_rot;   // same for the SO2 / SO3 part
};

struct DSpace {
  int32_t n_sub, check_bounds, n_cols, pad;
  DSub sub[kMaxSubspaces];
};

}  // namespace ramp
