/* trajopt_cuda.h — C ABI of libtrajopt_cuda.so (B200 / sm_100a collision-term evaluation).
 *
 * This is the drop-in boundary for ONE hot path of trajopt 0.35: the per-SQP-iteration evaluation of
 * collision terms and their gradients over every (trajectory, timestep/sub-state, link pair).
 * The reference has no FFI on this path; its seams are C++ virtual interfaces.  Each entry point below
 * names the reference interface it stands behind (paths relative to the trajopt source tree):
 *
 *   tc_scene_create     evaluator constructors                      trajopt/src/collision_terms.cpp:600-644,737-808,978-1051
 *                                                                   trajopt_ifopt/src/constraints/collision/discrete_collision_evaluators.cpp:38-87
 *                                                                   trajopt_ifopt/src/constraints/collision/continuous_collision_evaluators.cpp:41-95,266-315
 *   tc_eval             CalcCollisions / calcCollisions / calcCollisionData (geometry pass) and
 *                       GetGradient / getGradient (linearisation)   trajopt/src/collision_terms.cpp:195-332,653-688,817-898,1060-1161
 *                                                                   trajopt_common/src/collision_utils.cpp:116-221
 *                                                                   trajopt_ifopt/.../discrete_collision_evaluators.cpp:89-202
 *                                                                   trajopt_ifopt/.../continuous_collision_evaluators.cpp:97-248,317-458
 *   tc_result_contacts  tesseract::collision::ContactResultMap as flattened by
 *                       flattenWrapperResults                       trajopt/src/collision_terms.cpp:436-461
 *   tc_result_gradients trajopt_common::GradientResults             trajopt_common/include/trajopt_common/collision_types.h:187-235
 *   tc_result_rows      DiscreteCollisionConstraintD / ContinuousCollisionConstraintD ::getValues/getJacobian
 *                                                                   trajopt_ifopt/.../discrete_collision_constraint.cpp:177-283
 *                                                                   trajopt_ifopt/.../continuous_collision_constraint.cpp:276-417
 *                       and CollisionsToDistanceExpressions (legacy AffExpr rows)
 *                                                                   trajopt/src/collision_terms.cpp:341-386,463-554
 *   tc_result_sets      GradientResultsSet grouping, top-max_num_cnt order, weighted-average rows
 *                       of the fixed-size constraints               trajopt_ifopt/.../discrete_collision_evaluators.cpp:101-148
 *                                                                   trajopt_ifopt/.../continuous_collision_evaluators.cpp:113-162
 *                                                                   trajopt_ifopt/.../weighted_average_methods.cpp:31-107
 *                                                                   trajopt_ifopt/.../discrete_collision_constraint.cpp:68-141
 *                                                                   trajopt_ifopt/.../continuous_collision_constraint.cpp:80-236
 *   tc_result_cost      CollisionCost::value / CollisionConstraint::value
 *                                                                   trajopt/src/collision_terms.cpp:1287-1306,1367-1386
 *   tc_result_qp_csc    CollisionCost::convex + addHinge rows as the CSC matrix OSQPModel::updateConstraints builds
 *                                                                   trajopt/src/collision_terms.cpp:1267-1285
 *                                                                   trajopt_sco/src/modeling.cpp:19-27, osqp_interface.cpp:213-280
 *   tc_allgather_best   (no reference counterpart: multi-start winner selection across GPUs)
 *
 * Conventions: plain C, opaque handles, int status (0 = TC_OK), thread-local error string.
 * All matrices are row-major.  A pose is 12 doubles = the top 3x4 of a homogeneous transform.
 * Distinct handles may be used from distinct host threads concurrently; one handle is not re-entrant
 * (same rule as a reference evaluator instance, trajopt/include/trajopt/collision_terms.hpp:236-240).
 * There is no CPU fallback: every call fails with TC_ERR_CUDA when no sm_100 device is usable.
 */
#ifndef TRAJOPT_CUDA_H
#define TRAJOPT_CUDA_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define TC_ABI_VERSION 2

/* status codes */
enum {
  TC_OK = 0,
  TC_ERR_INVALID = 1,   /* bad argument / inconsistent description */
  TC_ERR_CUDA = 2,      /* CUDA runtime error or no usable device  */
  TC_ERR_UNSUPPORTED = 3, /* shape type / evaluator outside the scoped geometry */
  TC_ERR_OVERFLOW = 4,  /* a bounded arena overflowed; counts tell which */
  TC_ERR_NCCL = 5
};

/* joint types of the flattened kinematic tree */
enum { TC_JOINT_FIXED = 0, TC_JOINT_REVOLUTE = 1, TC_JOINT_PRISMATIC = 2 };

/* collision shape types (convex only; boxes and convex meshes are both vertex hulls) */
enum { TC_SHAPE_SPHERE = 0, TC_SHAPE_HULL = 1, TC_SHAPE_CAPSULE = 2, TC_SHAPE_CYLINDER = 3, TC_SHAPE_CONE = 4 };

/* evaluator kinds; values follow tesseract::collision::CollisionEvaluatorType as used by the JSON
 * front end (trajopt/src/problem_description.cpp:1654-1657): DISCRETE=1 ... LVS_CONTINUOUS=4 */
enum { TC_EVAL_DISCRETE = 1, TC_EVAL_LVS_DISCRETE = 2, TC_EVAL_CONTINUOUS = 3, TC_EVAL_LVS_CONTINUOUS = 4 };

/* which stack's conventions apply where the two stacks differ (Jacobian evaluation state,
 * CONTINUOUS = single sweep or not) */
enum { TC_STACK_IFOPT = 0, TC_STACK_LEGACY = 1 };

/* arithmetic of the narrow phase */
enum { TC_PRECISION_F64 = 0, TC_PRECISION_F32 = 1 };

/* tesseract::collision::ContinuousCollisionType */
enum { TC_CC_NONE = 0, TC_CC_TIME0 = 1, TC_CC_TIME1 = 2, TC_CC_BETWEEN = 3 };

/* what tc_eval has to produce */
enum {
  TC_WANT_COST = 1,
  TC_WANT_CONTACTS = 2,
  TC_WANT_GRADIENTS = 4, /* per-contact GradientResults (tc_result_gradients, tc_result_sets, and rows on demand) */
  TC_WANT_ROWS = 8       /* assemble the constraint rows inside tc_eval; without TC_WANT_GRADIENTS the per-link
                            gradient vectors are folded into the rows on the fly and never stored */
};

/* tesseract::collision::ContactTestType of the contact request the evaluators pass to contactTest
 * (trajopt/src/problem_description.cpp:1669-1673; trajopt_common/include/trajopt_common/collision_types.h:113).
 * ALL is 0 here so that a zero-initialised tc_eval_desc keeps the reference's default.
 *   ALL      every contact of every shape pair
 *   CLOSEST  per tested state (or sweep) and LINK pair only the contact with the smallest distance, chosen BEFORE the
 *            evaluator's own filters, as the manager does; results within 1e-9 m of the smallest count as tied and
 *            the first of them in shape-pair order is taken (sub-shapes of a compound link share faces, so exact
 *            ties are common and the manager's strict "closer" would be decided by the last bit)
 *   FIRST    per tested state (or sweep) one contact: the manager stops at the first one it finds; tesseract's order
 *            is that of Bullet's broadphase, here it is the first in (link pair, shape pair) order */
enum { TC_CONTACT_ALL = 0, TC_CONTACT_CLOSEST = 1, TC_CONTACT_FIRST = 2 };

/* contact flags */
enum { TC_FLAG_EPA = 1, TC_FLAG_DEGENERATE = 2, TC_FLAG_UNCONVERGED = 4 };

typedef struct tc_scene tc_scene;
typedef struct tc_result tc_result;

/* Scene = robot kinematics + collision geometry + pair tables; immutable after creation.
 * Active links (those moved by the planning group's joints) must come first in the link arrays. */
typedef struct tc_scene_desc {
  /* kinematic tree of the planning group, topologically ordered (frame_parent[f] < f, -1 = group base) */
  int32_t n_dof;
  int32_t n_frames;
  const int32_t* frame_parent;     /* [n_frames] */
  const int32_t* frame_joint_type; /* [n_frames] TC_JOINT_* */
  const int32_t* frame_dof;        /* [n_frames] dof index or -1 */
  const double* frame_axis;        /* [n_frames][3] unit axis in the joint frame */
  const double* frame_origin;      /* [n_frames][12] parent frame -> joint frame at q = 0 */
  double base_pose[12];            /* world <- group base link */
  /* collision links */
  int32_t n_links;
  int32_t n_active_links;
  const int32_t* link_frame;       /* [n_links] frame of an active link, -1 for static links */
  const double* link_pose;         /* [n_links][12] world pose of static links (ignored for active) */
  /* collision shapes */
  int32_t n_shapes;
  const int32_t* shape_type;       /* [n_shapes] TC_SHAPE_* */
  const int32_t* shape_link;       /* [n_shapes] owning link */
  const int32_t* shape_id;         /* [n_shapes] index of the shape within its link (ContactResult::shape_id) */
  const double* shape_params;      /* [n_shapes][4] sphere: r; capsule/cylinder/cone: r, half length (axis = local z) */
  const double* shape_pose;        /* [n_shapes][12] shape frame in the link frame */
  const int32_t* shape_vert_offset;/* [n_shapes] first vertex of a hull in verts */
  const int32_t* shape_vert_count; /* [n_shapes] */
  int32_t n_verts;
  const double* verts;             /* [n_verts][3] hull vertices in the shape frame */
  /* pair tables */
  const uint8_t* acm;              /* [n_links][n_links] non-zero = pair never checked (allowed collision) */
  const double* pair_margin;       /* [n_links][n_links] or NULL -> default_margin everywhere */
  const double* pair_coeff;        /* [n_links][n_links] or NULL -> default_coeff everywhere  */
  double default_margin;           /* CollisionMarginData default, WITHOUT the buffer */
  double default_coeff;
  double margin_buffer;            /* TrajOptCollisionConfig::collision_margin_buffer */
  /* Dynamic environment (trajopt/src/collision_terms.cpp:399-411): links that move WITHOUT belonging to the planning
   * group (another robot, a positioner, an attached object).  The reference then activates the contact manager on every
   * active link of the environment, so such a link is tested against everything — the group's links, the static links
   * and the other moving links — although it has no gradient for this group.  The n_env_active_links links that
   * directly follow the group's links (indices n_active_links .. n_active_links + n_env_active_links - 1) are those
   * links; their poses come from link_pose / tc_scene_set_link_poses.  0 = every non-group link is static and only
   * pairs with a group link are candidates. */
  int32_t n_env_active_links;
  int32_t reserved;
} tc_scene_desc;

/* One batched evaluation: B trajectories x T timesteps, q[B][T][n_dof] row-major. */
typedef struct tc_eval_desc {
  int32_t evaluator;         /* TC_EVAL_* */
  int32_t stack;             /* TC_STACK_* */
  int32_t precision;         /* TC_PRECISION_* */
  int32_t want;              /* TC_WANT_* mask */
  double lvs_length;         /* longest_valid_segment_length */
  const uint8_t* fixed_steps;/* [T] non-zero = timestep is fixed (host memory), or NULL */
  int32_t q_on_device;       /* non-zero: q is a device pointer (it must stay valid until the result is no longer read:
                                rows and the QP block are assembled from it on demand) */
  int32_t contact_test_type; /* TC_CONTACT_* (0 = ALL) */
} tc_eval_desc;

/* One contact = the fields of tesseract::collision::ContactResult the path reads
 * (trajopt/src/collision_terms.cpp:105-124,215-223) plus its position in the batch. */
typedef struct tc_contact {
  double distance;
  double normal[3];                  /* unit, from link_ids[0]'s shape towards link_ids[1]'s */
  double nearest_points[2][3];       /* world */
  double nearest_points_local[2][3]; /* in the link frame at transform[i] */
  double cc_time[2];                 /* -1 when not continuous */
  int32_t cc_type[2];                /* TC_CC_* */
  int32_t link_ids[2];               /* scene link indices, link_ids[0] < link_ids[1] */
  int32_t shape_id[2];
  int32_t subshape_id[2];
  int32_t traj;                      /* b */
  int32_t term;                      /* timestep (DISCRETE) or segment t = (t, t+1) */
  int32_t substep;                   /* LVS sub-state / sub-segment index */
  int32_t pair;                      /* index into the scene's link-pair table */
  int32_t state;                     /* global sub-state index of the T0 pose */
  int32_t flags;                     /* TC_FLAG_* */
} tc_contact;

/* trajopt_common::GradientResults without the vectors (those come as a dense array) */
typedef struct tc_gradient {
  double error;              /* margin - distance */
  double error_with_buffer;  /* margin + buffer - distance */
  double scale0[2];          /* gradients[i].scale     (T0 block) */
  double scale1[2];          /* cc_gradients[i].scale  (T1 block) */
  int32_t has_gradient[2];
  int32_t cc_type[2];
} tc_gradient;

typedef struct tc_counts {
  int64_t n_states;      /* sub-states after LVS subdivision */
  int64_t n_units;       /* tested states (discrete) or swept sub-segments (cast): S */
  int64_t n_dense;       /* S x candidate link pairs  (the headline count)   */
  int64_t n_candidates;  /* shape-pair tests entering the broadphase          */
  int64_t n_narrowphase; /* shape-pair tests surviving the AABB test          */
  int64_t n_contacts;
  int64_t n_rows;
  int64_t nnz;
  int64_t gjk_iters;     /* sum of GJK iterations                              */
  int64_t support_verts; /* sum of hull vertices scanned by support calls     */
  int64_t n_epa;
  int64_t n_unconverged;
  int64_t n_terms;       /* B x terms per trajectory                          */
  int64_t overflow;      /* non-zero if an arena overflowed                   */
  int64_t epa_iters;     /* sum of EPA expansion iterations                   */
  int64_t support_compared; /* hull vertices actually compared (support look-up tables skip the rest) */
} tc_counts;

/* one row set of the fixed-size constraints (GradientResultsSet reduced to what the constraint reads) */
typedef struct tc_set {
  int32_t traj, term, pair;
  int32_t shape_key[2];      /* cantorHash(shape_id, subshape_id) per link */
  int32_t n_results;
  int32_t has_error[2][2];   /* [link][0: excluding T1, 1: excluding T0] */
  double coeff;
  double error[2][2];
  double error_with_buffer[2][2];
} tc_set;

/* ---- lifetime ---- */
int tc_init(int device_ordinal);
const char* tc_last_error(void);
int tc_abi_version(void);

/* ---- scene ---- */
int tc_scene_create(const tc_scene_desc* desc, tc_scene** out);
void tc_scene_destroy(tc_scene* scene);
/* Dynamic environment (trajopt/src/collision_terms.cpp:399-411,659-660,828-836): links outside the planning group may
 * move between evaluations (another robot, a positioner, an attached object).  The reference reads their poses from
 * the environment state on every call; here the caller pushes them: link_pose[n_links][12], world poses; the entries
 * of the planning group's own links are ignored.  Takes effect for the next tc_eval; must not overlap an evaluation
 * that uses this scene.  Pairs of two non-group links are candidates when at least one of them is among the scene's
 * n_env_active_links (they carry no gradient for this group: constant rows, but they count in the cost). */
int tc_scene_set_link_poses(tc_scene* scene, const double* link_pose);
/* candidate link pairs (>= 1 active link, not ACM-disabled, sorted by (link0, link1)) */
int tc_scene_num_pairs(const tc_scene* scene);
int tc_scene_pairs(const tc_scene* scene, int32_t* link0, int32_t* link1, double* margin, double* coeff);

/* ---- streams: for hosts that do not link a CUDA runtime of their own (a trajopt build does not).  Any
 * cudaStream_t of the process may be passed to tc_eval instead. ---- */
int tc_stream_create(void** stream);
void tc_stream_destroy(void* stream);

/* ---- evaluation ---- */
int tc_result_create(tc_scene* scene, tc_result** out);
void tc_result_destroy(tc_result* res);
/* Runs the whole pass on `cuda_stream` (a cudaStream_t, NULL = default stream) and synchronises it
 * before returning, so the accessors below can be called right away. */
int tc_eval(tc_scene* scene, const tc_eval_desc* desc, const double* q, int32_t B, int32_t T,
            tc_result* out, void* cuda_stream);
/* Asynchronous variant: enqueues the pass and returns; tc_result_sync() must be called before any accessor and before
 * the next evaluation on the same result, and q (host or device) must stay valid until then.
 * How far the call itself blocks depends on the pass (csrc/tc_api.cu runEval):
 *   - the FIRST pass on a result, and any pass whose shape differs from the previous one (B, T, evaluator, options,
 *     fixed steps), sizes its arenas from the device's counts stage by stage: five stream synchronisations inside
 *     the call, possibly cudaMalloc;
 *   - every later pass of the same shape is SPECULATIVE: arenas and grids come from the previous pass's counts plus
 *     1/8 slack, the kernels read the actual counts on the device, and the call returns after ~30 launches without
 *     touching the stream's progress.  If a count outgrows its capacity the device raises a flag, the remaining
 *     kernels leave at once, and tc_result_sync() repeats the pass with exact sizes (tc_result_pass_info tells).
 * The per-iteration call pattern this serves: trajopt_sco/src/optimizers.cpp:194-203. */
int tc_eval_async(tc_scene* scene, const tc_eval_desc* desc, const double* q, int32_t B, int32_t T,
                  tc_result* out, void* cuda_stream);
int tc_result_sync(tc_result* res);
/* how the last pass ran: speculative = sized from the previous pass (no read-back inside the call), repeated = number of
 * passes on this result that had to be repeated exactly because a count outgrew its predicted capacity */
int tc_result_pass_info(const tc_result* res, int32_t* speculative, int32_t* repeated);

/* ---- seam B3: the contact manager itself.  What a tesseract_collision contact-manager plugin does between
 * setCollisionObjectsTransform and contactTest (trajopt/src/collision_terms.cpp:616-625,665-667,887-889,1132-1134;
 * factories named in trajopt_common/data/contact_manager_plugins.yaml:2-36), for n_states states at once and WITHOUT
 * kinematics: the caller supplies link poses, not joint values.
 *   link_pose0[n_states][n_active_links][12]  world poses of the scene's active links (the links the manager was told are
 *                                             active: the first n_active_links links of the scene) in every state
 *   link_pose1                                same shape: poses at the END of the sweep -> cast (continuous) manager,
 *                                             every active shape is the convex hull of its two placements;
 *                                             NULL -> discrete manager
 * Static links keep the poses of the scene (tc_scene_create / tc_scene_set_link_poses).  Both arrays are host memory.
 * The result is what tc_eval leaves for evaluator DISCRETE (link_pose1 == NULL) or CONTINUOUS with one sweep per state:
 * contacts of state s carry traj = s, term = 0, sorted by (state, link pair, shape pair), distance <= pair margin +
 * margin buffer (the manager's contact distance threshold), cc_time / cc_type filled for cast tests.
 * tc_result_counts / _contacts / _contact_transforms / _term_ptr (one "term" per state) / _cost apply; gradients and
 * rows need joint values and are refused.  No trajopt-side filter is applied (no fixed states, zero coefficients are
 * still dropped as pairs the manager never reports); contact_test_type as in tc_eval_desc.
 * A kinematics-free scene (one FIXED frame per active link) is enough for this entry point. */
int tc_contact_test(tc_scene* scene, const double* link_pose0, const double* link_pose1, int32_t n_states,
                    int32_t contact_test_type, int32_t precision, tc_result* out, void* cuda_stream);

/* ---- result access (valid until the next tc_eval on the same result) ---- */
int tc_result_counts(const tc_result* res, tc_counts* counts);
/* contacts sorted by (traj, term, link pair, substep, shape pair) */
int tc_result_contacts(const tc_result* res, tc_contact* dst, size_t cap);
/* transform[i] and cc_transform[i] of every contact: dst[n_contacts][2 links][2 (T0, T1)][12] */
int tc_result_contact_transforms(const tc_result* res, double* dst, size_t cap_contacts);
/* first contact of every (traj, term): term_ptr[n_terms + 1] */
int tc_result_term_ptr(const tc_result* res, int64_t* term_ptr);
/* per-contact gradient records; grad[n_contacts][2 links][2 (T0 block, T1 block)][n_dof] */
int tc_result_gradients(const tc_result* res, tc_gradient* rec, double* grad, size_t cap_contacts);
/* One row per contact, in contact order.  Columns are trajectory-local variable indices
 * t * n_dof + j.  ifopt stack: row_value = margin - distance, val = -(sum_i scale_i grad_i)
 * (the D constraints).  legacy stack: row_value = constant of the affine distance expression
 * d(x) = row_value + val . x  (CollisionsToDistanceExpressions + CalcDistExpressions*). */
int tc_result_rows(const tc_result* res, int64_t* row_ptr, int32_t* col_idx, double* val, double* row_value,
                   double* row_coeff, double* row_margin);
/* tc_result_rows without the final wait: the copies are enqueued on the evaluation's stream and the call returns;
 * the host arrays (pinned memory, or the copies degrade to synchronous ones) are valid after tc_result_wait().
 * Lets a caller overlap the download of one sub-batch with the evaluation of the next on another stream. */
int tc_result_rows_async(const tc_result* res, int64_t* row_ptr, int32_t* col_idx, double* val, double* row_value,
                         double* row_coeff, double* row_margin);
int tc_result_wait(const tc_result* res);
/* The same rows without explicit column indices.  A row's non-zeros are runs of n_dof consecutive columns — the
 * variables of timestep `term` (T0 block) and of `term + 1` (T1 block) — exactly what the reference's D constraints
 * write with insertBack(row, var.getIndex() + j) (discrete_collision_constraint.cpp:262-281,
 * continuous_collision_constraint.cpp:353-413).  row_first_col[n_rows][2] = first column of the T0 / T1 run, -1 where
 * the block is absent (fixed timestep, single-timestep evaluator); val holds the runs back to back as in tc_result_rows
 * (row i at val[row_ptr[i]]).  Saves the 4 bytes per non-zero of col_idx on the device -> host path, which is what
 * bounds the host-buffer pass.  The _async form follows tc_result_rows_async. */
int tc_result_rows_blocks(const tc_result* res, int64_t* row_ptr, int32_t* row_first_col, double* val, double* row_value,
                          double* row_coeff, double* row_margin);
int tc_result_rows_blocks_async(const tc_result* res, int64_t* row_ptr, int32_t* row_first_col, double* val,
                                double* row_value, double* row_coeff, double* row_margin);
/* same rows left in device memory (for a consumer that assembles the QP on the GPU); valid until the next
 * tc_eval on this result */
int tc_result_rows_device(const tc_result* res, const int64_t** row_ptr, const int32_t** col_idx, const double** val,
                          const double** row_value, const double** row_coeff, const double** row_margin);
/* QP constraint block of the collision hinge costs in compressed-sparse-COLUMN form, assembled on the device — what
 * CollisionCost::convex + ConvexObjective::addHinge produce (trajopt/src/collision_terms.cpp:1267-1285,
 * trajopt_sco/src/modeling.cpp:19-27) after OSQPModel::updateConstraints / exprToEigen have turned the expressions
 * into OSQP's matrix (trajopt_sco/src/osqp_interface.cpp:213-280, solver_utils.cpp:111-144).
 *   rows    : the contact rows of tc_result_rows, same order (n_rows)
 *   columns : [ B*T*n_dof trajectory variables, column b*T*n_dof + t*n_dof + j | one slack variable per row ]
 *   A = [ -G | -I ], upper = constant - margin    (legacy stack: d(x) = constant + G.x, row: margin - d(x) - s <= 0)
 *   A = [  J | -I ], upper = J.x0 - (margin - d)  (ifopt stack: violation linearised at the iterate x0)
 *   slack_cost[i] = coeff of row i (objective term coeff * s_i; s_i >= 0 is the caller's variable bound)
 * col_ptr[n_cols + 1], row_idx[nnz], val[nnz] with n_cols = B*T*n_dof + n_rows and nnz = counts.nnz + n_rows; row
 * indices ascend inside every column.  The block of trajectory b is the contiguous row range of its terms and the
 * columns b*T*n_dof .. (b+1)*T*n_dof - 1 plus its rows' slack columns.
 * CollisionConstraint::convex (addIneqCnt(coeff * viol), collision_terms.cpp:1347-1365) is the same block without the
 * slack columns (the first B*T*n_dof + 1 entries of col_ptr), row i and upper[i] scaled by slack_cost[i]. */
int tc_result_qp_csc(const tc_result* res, int64_t* col_ptr, int32_t* row_idx, double* val, double* upper,
                     double* slack_cost);
/* the same arrays left on the device (valid until the next tc_eval on this result); any output pointer may be NULL */
int tc_result_qp_csc_device(const tc_result* res, const int64_t** col_ptr, const int32_t** row_idx, const double** val,
                            const double** upper, const double** slack_cost, int64_t* n_rows, int64_t* n_cols,
                            int64_t* nnz);
/* fixed-size constraints: sets of every term sorted as the evaluator sorts them, the first
 * min(max_num_cnt, n) of each term turned into rows.  set_ptr[n_terms + 1].
 * values[n_terms][max_num_cnt], coeffs idem, jac[n_terms][max_num_cnt][2][n_dof] (T0 block, T1 block) */
int tc_result_sets(const tc_result* res, int32_t max_num_cnt, int64_t* set_ptr, tc_set* sets, size_t cap_sets,
                   double* values, double* coeffs, double* jac);
/* sum over contacts of coeff * max(margin - distance, 0): per (traj, term) and per trajectory */
int tc_result_cost(const tc_result* res, double* cost_per_term, double* cost_per_traj);
/* device pointer to cost_per_traj[B] (stays on the GPU for multi-start selection) */
int tc_result_cost_device(const tc_result* res, const double** dev_ptr);

/* ---- host-buffer pass over a large batch, pipelined inside the library.
 * The consumer of the rows is the unchanged host QP build (trajopt_optimizers/trajopt_sqp/src/trajopt_qp_problem.cpp:
 * 720-973 / trajopt_sco/src/osqp_interface.cpp:213-280), so a batched linearisation ends with hundreds of MB going
 * device -> host.  tc_pipeline_eval_rows cuts the batch into sub-batches of `chunk` trajectories (the last one tapered
 * 1/2 + 1/4 + 1/4: its download is the part nothing overlaps), evaluates them on a ring of streams with speculative
 * passes (tc_eval_async) and starts the download of sub-batch j while j + 1 computes — one host thread, no
 * allocation in steady state.  Rows arrive in COMPACT form, the fewest bytes that determine them:
 *   term_ptr[B * terms_per_traj + 1]  first row of every (trajectory, term); terms_per_traj follows from T, the
 *                                     evaluator and fixed_steps exactly as in tc_eval
 *   val[nnz]                          row i = nb * n_dof values (T0 block, then T1 block; nb = free blocks of its term),
 *                                     rows back to back in row order -> the offset of a row follows from term_ptr
 *   row_value[n_rows], row_pair[n_rows]  as tc_result_rows' row_value and the index into tc_scene_pairs (margin and
 *                                     coefficient of a row are those of its link pair: 4 bytes instead of 16)
 *   cost_per_traj[B]
 * (124 bytes per two-block row against 152 of tc_result_rows_blocks and 180 of tc_result_rows.)  All output arrays
 * are host memory, pinned for the copies to be asynchronous; val_cap / row_cap are their capacities in elements
 * (TC_ERR_OVERFLOW when a sub-batch would not fit).  q[B][T][n_dof] is host memory (pinned: read by a kernel through
 * the mapping; pageable: staged by the runtime).  desc->want is ignored (cost + contacts + rows). */
typedef struct tc_pipeline tc_pipeline;
int tc_pipeline_create(tc_scene* scene, int32_t n_slots, tc_pipeline** out);
void tc_pipeline_destroy(tc_pipeline* pipe);
int tc_pipeline_eval_rows(tc_pipeline* pipe, const tc_eval_desc* desc, const double* q, int32_t B, int32_t T, int32_t chunk,
                          int64_t* term_ptr, double* val, size_t val_cap, double* row_value, int32_t* row_pair,
                          size_t row_cap, double* cost_per_traj, int64_t* n_rows, int64_t* nnz);
/* The same pass with the rows in PACKED form: a block of a row holds only the joints that move one of the row's two
 * links — ascending joint index, the set bits of tc_scene_pair_dof_mask(row_pair[i]) — instead of all n_dof.  The
 * entries left out are the exact zeros the reference writes for joints outside both chains (getJacobian fills every
 * column of a block, discrete_collision_constraint.cpp:266-277; continuous_collision_constraint.cpp:380-411), a third
 * of the values on the arm_around_table batch, so nothing is lost; a QP builder that drops explicit zeros
 * (trajopt_sco/src/solver_utils.cpp:111-144 prunes them) reads this form directly.  Row i starts where row i - 1
 * ends: width(i) = popcount(mask(row_pair[i])) * nb(term of i); nnz counts the packed values. */
int tc_pipeline_eval_rows_packed(tc_pipeline* pipe, const tc_eval_desc* desc, const double* q, int32_t B, int32_t T,
                                 int32_t chunk, int64_t* term_ptr, double* val, size_t val_cap, double* row_value,
                                 int32_t* row_pair, size_t row_cap, double* cost_per_traj, int64_t* n_rows, int64_t* nnz);
/* bit j set: joint j is on the kinematic chain of one of the pair's two links (a link outside the planning group
 * contributes nothing).  Scenes of up to 64 joints. */
int tc_scene_pair_dof_mask(const tc_scene* scene, int32_t pair, uint64_t* mask);

/* ---- allocation-free entry point.  It has NO counterpart in the reference: it was added
 * for callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that
 * forbid it).  Same arithmetic as tc_result_cost's per-trajectory output, over caller-owned DEVICE arrays:
 *   term_ptr[n_terms + 1]  contact range of every (trajectory, term), trajectory-major (as tc_result_term_ptr)
 *   distance / margin / coeff [term_ptr[n_terms]]  per contact: signed distance, the pair's margin and coefficient
 *   cost_per_traj[n_terms / terms_per_traj]  out: sum over the trajectory's contacts of coeff * max(margin - d, 0)
 * The per-term intermediate lives in a static device array of TC_NOALLOC_MAX_TERMS doubles (512 KiB, held by every
 * GPU that loads the library until exit); larger n_terms are rejected with TC_ERR_OVERFLOW.  The call enqueues two
 * kernels on cuda_stream (NULL = default stream) and returns without synchronising; it calls no cudaMalloc / cudaFree
 * or allocating library routine.  All calls share the array: calls that overlap on the GPU race — order them on one
 * stream or with events.  Nothing in the array has to be cleared between calls (every slot read is written first).
 * The call works on the CURRENT CUDA device (the arrays are the caller's) and does not need tc_init; tc_init resolves
 * the scratch address of its device, any other device resolves it on its first call (one cudaGetSymbolAddress under a
 * mutex): under graph capture make that first call, or tc_init, before the capture begins. */
#define TC_NOALLOC_MAX_TERMS 65536
int tc_cost_noalloc(const int64_t* term_ptr, int32_t n_terms, int32_t terms_per_traj, const double* distance,
                    const double* margin, const double* coeff, double* cost_per_traj, void* cuda_stream);

/* ---- multi-GPU helpers for multi-start (one process per GPU).  NCCL is resolved from the process at
 * run time (libnccl.so.2).  A communicator can come from the caller (an ncclComm_t) or be created here:
 * rank 0 calls tc_nccl_unique_id and ships the 128 bytes to the other ranks by any side channel. ---- */
int tc_nccl_unique_id(char id[128]);
int tc_nccl_comm_create(int n_ranks, int rank, const char id[128], void** nccl_comm);
void tc_nccl_comm_destroy(void* nccl_comm);
/* every rank contributes its best (cost, global index); all ranks receive the global winner
 * (one ncclAllGather of 16 bytes per rank over NVLink) */
int tc_allgather_best(double local_best_cost, int64_t local_best_idx, void* nccl_comm, int n_ranks,
                      void* cuda_stream, double* best_cost, int64_t* best_idx, int32_t* best_rank);
/* the owner of the winning seed (best_rank) ships its trajectory traj[n_values] (host memory, T * n_dof doubles) to
 * every rank: one ncclBroadcast; on return every rank's buffer holds the root's values */
int tc_broadcast_trajectory(double* traj, size_t n_values, int root_rank, void* nccl_comm, void* cuda_stream);

/* The two calls above in one, fed from the device: the cheapest trajectory of `res` (arg-min over its per-trajectory
 * costs, lowest index on ties; the evaluation must have run with TC_WANT_COST) is packed with its joint values on the
 * device, all ranks exchange their records with ONE ncclAllGather ((2 + T * n_dof) * 8 bytes per rank) on the result's
 * stream and pick the global winner (lowest rank on ties): no host-side cost array, no second collective, one stream
 * synchronisation, no cudaMalloc.  index_offset = global index of this rank's first trajectory.  best_traj
 * [T * n_dof] (host) receives the winner's joint values on every rank; may be NULL.  n_ranks == 1 (nccl_comm may be
 * NULL) just reads this rank's winner back.  A pending tc_eval_async is waited for first. */
int tc_result_allgather_best(tc_result* res, int64_t index_offset, void* nccl_comm, int n_ranks, double* best_cost,
                             int64_t* best_idx, int32_t* best_rank, double* best_traj);

/* ---- measurement helpers (bench only) ---- */
/* register-resident DFMA / FFMA loop: returns achieved TFLOP/s (2 flop per FMA) */
int tc_measure_fma_peak(int precision, double* tflops);
/* duration in ms of each stage of the last tc_eval (CUDA events on the eval stream):
 * [0] plan+FK, [1] broadphase+narrowphase, [2] deep-EPA re-runs, [3] finalise+sort+gradients(+cost), [4] total,
 * [5] the GJK + EPA kernels, [6] broadphase (count, scan, fill), [7] = [0], [8] GJK kernel, [9] EPA kernel */
int tc_result_stage_ms(const tc_result* res, float ms[10]);
/* kernels of this library launched by the last tc_eval (+ row assembly when requested), and calls into
 * CUDA-toolkit library code (cub radix sort of the contact keys) */
int tc_result_launches(const tc_result* res, int32_t* own_kernels, int32_t* library_calls);

#ifdef __cplusplus
}
#endif
#endif /* TRAJOPT_CUDA_H */
