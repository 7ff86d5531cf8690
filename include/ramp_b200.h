/*
 * ramp_b200.h -- C ABI of libramp_b200.so
 *
 * B200 (sm_100a) implementation of RAMP's one data-parallel hot path:
 * state validity (forward kinematics -> geometry placements -> narrow-phase
 * collision over the enabled pair list) and motion validity (edge
 * discretisation at OMPL's default resolution).
 *
 * The reference (JafarAbdi/ramp) has no FFI of its own: its boundary is a set
 * of Python call sites into Pinocchio / coal / OMPL.  Every entry point below
 * names the reference call site (file:line under the reference tree) it
 * replaces.  Conventions:
 *
 *   - every function returns 0 on success and a negative RampStatus on error;
 *     ramp_last_error() returns a thread-local description of the last error;
 *   - nothing throws across this boundary;
 *   - pointers named d_* are CALLER-OWNED DEVICE pointers (for PyTorch:
 *     tensor.data_ptr() of a contiguous tensor), 16-byte aligned;
 *     pointers named h_* are caller-owned HOST pointers;
 *   - `stream` is a cudaStream_t passed as void* (0 = the legacy default
 *     stream); d_* calls are asynchronous on it;
 *   - a RampModel is bound to the device it was created on.  The d_* compute
 *     calls only read it and take their scratch per call, in stream order,
 *     from the model's memory pool: they may be issued from several host
 *     threads and on several streams at once.  The host-buffer calls
 *     (ramp_check_host, ramp_check_edges_host) share the model's staging
 *     buffers and take a per-model lock for their whole duration: concurrent
 *     ones on one model run one after the other.  ramp_model_set_scene /
 *     ramp_model_set_qmap / ramp_model_set_pair_active mutate the model and
 *     must be serialised against all compute calls by the caller;
 *   - the pool keeps what it has handed out once (the GJK work queue is the
 *     largest block: 3 items of 64 B per row up to 2^21 rows, at most 1 GB
 *     for larger batches) until ramp_model_trim or ramp_model_destroy;
 *   - there is no CPU fallback: without a CUDA device every compute call
 *     fails with RAMP_ERR_CUDA.
 */
#ifndef RAMP_B200_H
#define RAMP_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RAMP_B200_ABI_VERSION 1

typedef enum RampStatus {
  RAMP_OK = 0,
  RAMP_ERR_INVALID = -1, /* bad argument / malformed description            */
  RAMP_ERR_LIMIT = -2,   /* model exceeds a compiled-in capacity            */
  RAMP_ERR_CUDA = -3,    /* CUDA runtime error (no device, launch failure)  */
  RAMP_ERR_ALLOC = -4    /* host or device allocation failed                */
} RampStatus;

/* Joint vocabulary = the Pinocchio short names RAMP dispatches on
 * (src/ramp/constants.py:14-23, src/ramp/motion_planner.py:268-362). */
typedef enum RampJointType {
  RAMP_JOINT_RX = 1,
  RAMP_JOINT_RY = 2,
  RAMP_JOINT_RZ = 3,
  RAMP_JOINT_REVOLUTE_UNALIGNED = 4,
  RAMP_JOINT_RUBX = 5, /* q = (cos, sin), used as given */
  RAMP_JOINT_RUBY = 6,
  RAMP_JOINT_RUBZ = 7,
  RAMP_JOINT_RUB_UNALIGNED = 8,
  RAMP_JOINT_PX = 9,
  RAMP_JOINT_PY = 10,
  RAMP_JOINT_PZ = 11,
  RAMP_JOINT_PRISMATIC_UNALIGNED = 12,
  RAMP_JOINT_PLANAR = 13,        /* q = (x, y, cos, sin)                       */
  RAMP_JOINT_FREEFLYER = 14,     /* q = (x, y, z, qx, qy, qz, qw), quat as given */
  RAMP_JOINT_SPHERICAL = 15,     /* q = (qx, qy, qz, qw)                       */
  RAMP_JOINT_SPHERICAL_ZYX = 16, /* q = (z, y, x) angles, R = Rz Ry Rx         */
  RAMP_JOINT_TRANSLATION = 17    /* q = (x, y, z)                              */
} RampJointType;

/* Shape vocabulary = the coal/hpp-fcl shapes RAMP can produce
 * (src/ramp/robot_model.py:429-435, robots/rrr/rrr.urdf.xacro:17-103). */
typedef enum RampShapeType {
  RAMP_SHAPE_SPHERE = 0,   /* params = (r)                                   */
  RAMP_SHAPE_CAPSULE = 1,  /* params = (r, half_length) along local z        */
  RAMP_SHAPE_BOX = 2,      /* params = (hx, hy, hz) HALF side lengths        */
  RAMP_SHAPE_CYLINDER = 3, /* params = (r, half_length) along local z        */
  RAMP_SHAPE_CONVEX = 4,   /* vertices [vert_offset, vert_offset+vert_count) */
  RAMP_SHAPE_ELLIPSOID = 5 /* params = (rx, ry, rz)                          */
} RampShapeType;

/* One movable joint.  Joints are stored in Pinocchio order without the
 * universe: entry i is Pinocchio joint id i+1, parent < i, parent = -1 for a
 * child of the universe.  placement is the 3x4 row-major [R|t] of
 * model.jointPlacements[id]. */
typedef struct RampJoint {
  int32_t type;   /* RampJointType */
  int32_t parent; /* index into this array, -1 = universe */
  int32_t idx_q;  /* first q slot */
  int32_t reserved;
  double axis[3]; /* unaligned joints only */
  double placement[12];
} RampJoint;

/* One collision geometry object (pinocchio.GeometryObject).  placement is the
 * object's placement in its parent joint frame; parent_joint = -1 means the
 * universe (scene objects, src/ramp/robot_model.py:453). */
typedef struct RampGeom {
  int32_t type;         /* RampShapeType */
  int32_t parent_joint; /* index into joints[], -1 = universe */
  int32_t vert_offset;  /* convex only: first vertex in verts[] */
  int32_t vert_count;   /* convex only */
  double params[4];
  double placement[12];
} RampGeom;

/* How a batch of ACTUATED joint values becomes a full Pinocchio q
 * (src/ramp/robot_state.py:109-119, src/ramp/robot_model.py:363-375,406-426):
 *   q = ref_q; q[act_slot[k]] = in[k]; mimic: q[dst] = q[src]*mult + off;
 *   then for every cont_slot c: (q[c], q[c+1]) = (cos q[c], sin q[c]).
 * Order matters (mimic before continuous), as in the reference. */
typedef struct RampQMap {
  int32_t n_act;
  const int32_t* act_slot; /* [n_act] */
  int32_t n_mimic;
  const int32_t* mimic_dst; /* [n_mimic] q slots */
  const int32_t* mimic_src;
  const double* mimic_mult;
  const double* mimic_off;
  int32_t n_cont;
  const int32_t* cont_slot; /* [n_cont] */
  const double* ref_q;      /* [nq] */
} RampQMap;

typedef struct RampModelDesc {
  int32_t abi_version; /* RAMP_B200_ABI_VERSION */
  int32_t nq;
  int32_t n_joints;
  const RampJoint* joints;
  int32_t n_geoms; /* robot geometries; scene objects come via set_scene */
  const RampGeom* geoms;
  int32_t n_pairs;
  const int32_t* pairs; /* [n_pairs][2] geometry indices, collisionPairs order */
  int32_t n_verts;
  const double* verts; /* [n_verts][3] convex vertices, geometry-local */
} RampModelDesc;

/* OMPL state space description for edge discretisation.  One entry per
 * subspace that src/ramp/motion_planner.py:264-359 adds, over the columns of
 * the actuated vector that from_ompl_state (:89-162) produces. */
typedef enum RampSpaceType {
  RAMP_SPACE_REAL = 0, /* RealVectorStateSpace(dim), cols [col, col+dim)      */
  RAMP_SPACE_SO2 = 1,  /* 1 col                                               */
  RAMP_SPACE_SE2 = 2,  /* cols x, y, yaw; RealVector(2) x SO2                 */
  RAMP_SPACE_SE3 = 3,  /* cols x, y, z, qx, qy, qz, qw; RealVector(3) x SO3   */
  RAMP_SPACE_SO3 = 4   /* cols qx, qy, qz, qw                                 */
} RampSpaceType;

typedef struct RampSubspace {
  int32_t type; /* RampSpaceType */
  int32_t col;
  int32_t dim; /* REAL only (<= 3) */
  int32_t reserved;
  double lo[3]; /* bounds of the RealVector part */
  double hi[3];
} RampSubspace;

typedef struct RampSpaceDesc {
  int32_t n_sub;
  const RampSubspace* sub;
  double longest_valid_segment_fraction; /* OMPL default 0.01 */
  int32_t check_bounds;                  /* 1: states outside bounds are invalid
                                            (src/ramp/motion_planner.py:622-625) */
  int32_t reserved;
} RampSpaceDesc;

typedef struct RampModel RampModel; /* opaque */

/* flags for ramp_check_batch / ramp_fk_batch */
#define RAMP_Q_ACTUATED 1u /* d_q holds actuated values [n, n_act]; expand with the qmap */
#define RAMP_ALL_PAIRS 2u  /* no early-out after the first colliding pair (verbose mode) */

/* Build the device-resident model (joint table, geometry table, pair list,
 * convex vertices) on `device`.  Replaces the per-call model.createData() /
 * collision_model.createData() of src/ramp/robot_state.py:376-377 with a
 * handle built once. */
int ramp_model_create(const RampModelDesc* h_desc, int device, RampModel** out);

/* Replace the scene: `n_geoms` universe-attached objects appended after the
 * robot geometries, plus `n_pairs` extra pairs [(first, second)] appended
 * after the robot pairs in the given order.  Geometry indices in `pairs`
 * refer to the combined list (robot geoms first).  `verts` carries the convex
 * vertices of scene objects (their vert_offset is relative to it).
 * Replaces RobotState.add_object, src/ramp/robot_state.py:307-331. */
int ramp_model_set_scene(RampModel* m, const RampGeom* h_geoms, int32_t n_geoms,
                         const int32_t* h_pairs, int32_t n_pairs,
                         const double* h_verts, int32_t n_verts);

/* Install the actuated->q map used by RAMP_Q_ACTUATED calls and by
 * ramp_check_edges.  Replaces apply_pinocchio_mimic_joints /
 * apply_pinocchio_continuous_joints, src/ramp/robot_model.py:406-426. */
int ramp_model_set_qmap(RampModel* m, const RampQMap* h_map);

/* Disable (active = 0) or re-enable a pair by its index in the combined pair
 * list without changing pair indices.  Replaces
 * GeometryData.activeCollisionPairs / removeCollisionPair as used by
 * src/ramp/compute_disable_collisions.py:56-89. */
int ramp_model_set_pair_active(RampModel* m, int32_t pair_index, int32_t active);

void ramp_model_destroy(RampModel* m);

/* Give the memory the model's scratch pool holds back to the device (it is kept
 * between calls so that steady-state calls allocate nothing).  Waits for the
 * device to go idle first.  No reference counterpart. */
int ramp_model_trim(RampModel* m);

/* Work counters of the model since creation (or the last reset): out[0] = (configuration, pair)
 * work items that survived every cull and went through GJK in the queue kernel, out[1] =
 * launches of that kernel.  Waits for the device to go idle.  bench.py divides out[0] by the
 * configurations checked.  No reference counterpart. */
int ramp_model_counters(const RampModel* m, uint64_t out[2], int32_t reset);

/* Sizes of the live model: out[0]=nq, [1]=n_joints, [2]=n_geoms (robot+scene),
 * [3]=n_pairs (robot+scene), [4]=n_act (0 if no qmap), [5]=device. */
int ramp_model_info(const RampModel* m, int32_t out[6]);

/* Batched forward kinematics: d_q[n, ld_q] -> d_oMi[n, n_joints, 12]
 * (3x4 row-major [R|t] of data.oMi[id], id = 1..n_joints).
 * Replaces pinocchio.forwardKinematics / framesForwardKinematics,
 * src/ramp/robot_state.py:199-205. */
int ramp_fk_batch(const RampModel* m, const float* d_q, int64_t n, int32_t ld_q,
                  uint32_t flags, float* d_oMi, void* stream);

/* Batched geometry placements: d_oMg[n, n_geoms, 12].
 * Replaces pinocchio.updateGeometryPlacements, src/ramp/robot_state.py:345-351. */
int ramp_placements_batch(const RampModel* m, const float* d_q, int64_t n,
                          int32_t ld_q, uint32_t flags, float* d_oMg, void* stream);

/* Batched collision check: FK + placements + narrow phase over the pair list.
 *   d_hit[n]        1 if any active pair collides (required)
 *   d_first_pair[n] index of the first colliding pair in list order, -1 if
 *                   none (nullable) == GeometryData.collisionPairIndex
 *   d_min_dist[n]   min over active pairs of the signed distance (nullable).
 *                   Computed by a second kernel that runs every active pair
 *                   to convergence (GJK / EPA): correct, but 100x slower than
 *                   the flags -- ask for it only where clearance is needed
 *   d_pair_mask[n, ceil(n_pairs/32)] bit p set iff pair p collides (nullable;
 *                   implies RAMP_ALL_PAIRS) == collisionResults[p].isCollision()
 * Replaces pinocchio.computeCollisions at src/ramp/robot_state.py:379-392 and
 * src/ramp/compute_disable_collisions.py:154-161. */
int ramp_check_batch(const RampModel* m, const float* d_q, int64_t n, int32_t ld_q,
                     uint32_t flags, uint8_t* d_hit, int32_t* d_first_pair,
                     float* d_min_dist, uint32_t* d_pair_mask, void* stream);

/* The same check for callers that are not allowed to allocate device memory.
 * It has no counterpart in the reference: it exists because of where this
 * library gets embedded.  A planner thread inside a real-time control loop
 * cannot afford cudaMalloc/cudaFree (they may stall on all other work of the
 * process and take an unpredictable time); a process that split up its device
 * memory at start-up has none left to give.  ramp_check_batch draws a work
 * queue from a memory pool; this variant uses a fixed array inside the library
 * instead and calls no allocating routine.
 *
 *   - at most RAMP_NOALLOC_MAX_ROWS configurations per call; larger n is
 *     refused with RAMP_ERR_LIMIT (split the batch, or use ramp_check_batch);
 *   - the fixed array (about 6.5 MB) is part of the library image: every GPU
 *     that loads the library holds it until the process exits;
 *   - the first call on each device looks up the array's address
 *     (cudaGetSymbolAddress); make that first call before any stream capture;
 *   - all calls share the one array, which keeps the previous call's contents
 *     (the entry point resets the part that must start at zero).  Two calls
 *     whose kernels overlap on the GPU -- different streams, or different
 *     models on one device -- race on it: keep them in one stream or
 *     otherwise ordered;
 *   - arguments and results are those of ramp_check_batch without d_min_dist. */
#define RAMP_NOALLOC_MAX_ROWS 32768
int ramp_check_batch_noalloc(const RampModel* m, const float* d_q, int64_t n, int32_t ld_q,
                             uint32_t flags, uint8_t* d_hit, int32_t* d_first_pair,
                             uint32_t* d_pair_mask, void* stream);

/* Pack per-configuration flags to one bit each: bit (r & 31) of d_words[r >> 5] = (d_hit[r] != 0),
 * d_words has ceil(n / 32) words and unused bits of the last word are 0.  This is what the
 * result gather between the GPUs of a box ships (north_star: "flags ... gathered back"): an
 * eighth of the bytes.  No reference counterpart (a single process there). */
int ramp_pack_flags(const RampModel* m, const uint8_t* d_hit, int64_t n, uint32_t* d_words, void* stream);

/* Signed distance of selected pairs: d_dist[n, n_sel] for pair indices
 * h_pair_idx[n_sel].  Replaces pinocchio.computeDistance,
 * src/ramp/robot_state.py:360-364. */
int ramp_distance_batch(const RampModel* m, const float* d_q, int64_t n, int32_t ld_q,
                        uint32_t flags, const int32_t* h_pair_idx, int32_t n_sel,
                        float* d_dist, void* stream);

/* Batched motion validity.  Edge e runs from d_q0[e,:] to d_q1[e,:] (actuated
 * columns, ld floats per row).  nd = validSegmentCount(q0,q1); states
 * t = j/nd for j = 1..nd-1 plus q1 (reported as j = nd) are interpolated and
 * checked (bounds if space->check_bounds, then collision).
 *   d_valid[e]         1 iff every checked state is valid (required)
 *   d_first_invalid[e] smallest invalid j, -1 if none (nullable)
 *   d_nd[e]            nd (nullable)
 * Edges are processed 2^22 at a time.  Should such a slice hold more than
 * 2^32 - 1 interpolated states (end points far outside the bounds), none of it
 * is checked and all its edges come back with d_valid = 0, d_first_invalid = -2.
 * Replaces OMPL's implicit DiscreteMotionValidator::checkMotion driving
 * is_ompl_state_valid, src/ramp/motion_planner.py:366,469-483,612-625. */
int ramp_check_edges(const RampModel* m, const float* d_q0, const float* d_q1,
                     int64_t n_edges, int32_t ld, const RampSpaceDesc* h_space,
                     uint8_t* d_valid, int32_t* d_first_invalid, int32_t* d_nd,
                     void* stream);

/* Host-buffer form of ramp_check_edges for planners that validate one edge (or a
 * few) per call: h_q0 / h_q1 / outputs are host arrays, copies and one
 * synchronisation happen inside.  This is the call a MotionValidator::checkMotion
 * override makes (src/ramp/motion_planner.py:366 + OMPL's checkMotion). */
int ramp_check_edges_host(const RampModel* m, const float* h_q0, const float* h_q1,
                          int64_t n_edges, int32_t ld, const RampSpaceDesc* h_space,
                          uint8_t* h_valid, int32_t* h_first_invalid, int32_t* h_nd);

/* Host-buffer form of ramp_check_batch: the call of the single-state call sites
 * (RobotState.check_collision, src/ramp/robot_state.py:366) and of callers whose
 * batches live in host memory.  Rows move to the device in chunks of 2^18 on two
 * streams while earlier chunks are being checked; results are copied back; the
 * call returns when everything is done.  Page-locked buffers (ramp_host_alloc)
 * are used in place, pageable ones are staged.  Output pointers follow
 * ramp_check_batch (host memory). */
int ramp_check_host(const RampModel* m, const float* h_q, int64_t n, int32_t ld_q,
                    uint32_t flags, uint8_t* h_hit, int32_t* h_first_pair,
                    float* h_min_dist, uint32_t* h_pair_mask);

/* Batched differential inverse kinematics over many seeds.  Every seed is a row of
 * ACTUATED values of the group the qmap describes (ramp_model_set_qmap); it is iterated with damped
 * least squares,  q <- clamp(q + step * J^T (J J^T + damping^2 I)^-1 e),  e = pose error of the
 * tool frame seen from that frame (translation, rotation vector), J = its frame Jacobian in the
 * LOCAL frame, until |e| < tolerance or max_iterations.  One-degree-of-freedom joints only
 * (revolute, continuous, prismatic); mimic joints follow their source.
 *   d_target        [n][12] 3x4 row-major world poses, or one pose for all seeds (one_target = 1)
 *   d_q_out         [n, n_act] the iterated actuated values
 *   d_err / d_ok / d_iters  [n] final |e|, |e| < tolerance, iterations used (each nullable)
 * Replaces RobotState.differential_ik (pink + quadprog, one state at a time),
 * src/ramp/robot_state.py:237-304, and the rejection loop that draws IK goals,
 * examples/so_arm100_motion_planning.py:50-54.  Not the same optimiser: the reference solves a QP
 * per iteration; what is reproduced is the contract (a configuration whose tool frame is within
 * MAX_ERROR of the target, inside the joint limits, or failure). */
typedef struct RampIkParams {
  int32_t frame_parent_joint; /* index into joints[], -1 = universe */
  int32_t max_iterations;     /* src/ramp/constants.py:10 MAX_ITERATIONS = 500 */
  double frame_placement[12]; /* the tool frame in its parent joint, 3x4 row-major */
  double damping;             /* lambda */
  double step;                /* fraction of the least-squares step taken per iteration */
  double tolerance;           /* src/ramp/constants.py:11 MAX_ERROR = 1e-3 */
  double max_step;            /* largest change of one joint value per iteration */
  const double* lower;        /* [n_act] bounds of the actuated values (nullable: unbounded) */
  const double* upper;
} RampIkParams;

int ramp_ik_batch(const RampModel* m, const float* d_seed, int64_t n, int32_t ld, const float* d_target,
                  int32_t one_target, const RampIkParams* h_params, float* d_q_out, float* d_err,
                  uint8_t* d_ok, int32_t* d_iters, void* stream);

/* Page-locked host memory for the h_* arguments of the host-buffer calls.
 * ramp_check_host / ramp_check_edges_host copy straight from and to buffers
 * that are page-locked (any allocator: they ask the driver per pointer) and go
 * through an internal staging copy otherwise.  No reference counterpart (numpy
 * arrays there); a convenience so that callers need no CUDA binding of their
 * own.  ramp_host_alloc returns NULL on failure (see ramp_last_error). */
void* ramp_host_alloc(size_t bytes);
void ramp_host_free(void* p);

/* Number of kernels this library has launched in this process (all models,
 * all threads) -- used by bench.py to report gpu_launches. */
int64_t ramp_launch_count(void);

const char* ramp_last_error(void);
int ramp_abi_version(void);

#ifdef __cplusplus
}
#endif
#endif /* RAMP_B200_H */
