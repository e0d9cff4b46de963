// tc_strict.cu — the kernels whose arithmetic must be REPRODUCIBLE operation for operation: compiled with -fmad=false
// (see the Makefile), i.e. every product and sum rounded on its own, exactly as the CPU oracle (built with
// -ffp-contract=off) and the host-compiled check harness evaluate the same expressions.
//
//   k_fk            batched forward kinematics + world AABBs.  With the shared sin/cos of tc_math.cuh the poses are
//                   bit-identical to the oracle's FK, so the narrow phase starts from identical inputs.  The kernel is
//                   bound by its 4 GB of pose/AABB stores; splitting its FMAs costs nothing measurable.
//   k_route_curved  sends every broadphase survivor of a shape pair with a curved core (cylinder, cone) to ...
//   k_narrow_deep   ... the one-thread-per-item GJK + EPA with the polytope in a global scratch slab, which also
//                   re-runs the few hull items whose polytope outgrew the thread-local store of k_epa_wp.
// Why: GJK/EPA on a curved surface has no closest VERTEX to lock on to; it stops on a 1e-10 m gap, which fixes the
// distance to 1e-10 m but the direction only to ~1e-5, and where in that band it stops depends on the last bit of
// every intermediate.  Fused multiply-adds move those bits, so the contracted kernels reproduce the oracle's normals on
// cylinders and cones only to 1e-5..1e-4; uncontracted they reproduce them exactly.  Hulls, spheres and capsules
// (point / segment / vertex cores) terminate on a vertex set and do not need this.
#include "tc_narrow_kernels.cuh"

namespace tc
{
namespace
{
constexpr int TPB = 128;

__device__ __forceinline__ int lastLE(const int32_t* a, int n, int x)
{
  // largest i in [0, n) with a[i] <= x   (a ascending, a[0] <= x)
  int lo = 0, hi = n;
  while (hi - lo > 1)
  {
    const int mid = (lo + hi) >> 1;
    if (a[mid] <= x)
      lo = mid;
    else
      hi = mid;
  }
  return lo;
}

// world AABB of active shape `a` (link-frame box centre bc, half extents bh) under link pose T -> SoA box arena
__device__ __forceinline__ void writeShapeAabb(const Work& w, const Pose<double>& T, const double* bc, const double* bh, int a,
                                               long long s)
{
  const Vec3<double> c = T.apply(Vec3<double>(bc[0], bc[1], bc[2]));
  double* dst = w.aabb + static_cast<size_t>(a) * 6 * w.n_states + s;
#pragma unroll
  for (int r = 0; r < 3; ++r)
  {
    const double e = fabs(T.r[3 * r + 0]) * bh[0] + fabs(T.r[3 * r + 1]) * bh[1] + fabs(T.r[3 * r + 2]) * bh[2];
    dst[static_cast<size_t>(r) * w.n_states] = c.get(r) - e;
    dst[static_cast<size_t>(3 + r) * w.n_states] = c.get(r) + e;
  }
}

__global__ void k_unit_term(Work w)
{
  if (passOverflowed(w))
    return;
  const int u = blockIdx.x * blockDim.x + threadIdx.x;
  if (u < w.counters->n_units)
    w.unit_term[u] = lastLE(w.unit_base, w.ep.n_terms, u);
}

// ---------------------------------------------------------------------------------------------
// K1: batched forward kinematics of every sub-state + world AABB of every active shape
//   get_state_fn_ -> JointGroup::calcFwdKin (trajopt/src/collision_terms.cpp:401-417,656,880-885,1123-1124;
//   discrete_collision_evaluators.cpp:57-73; continuous_collision_evaluators.cpp:65-81,285-301) and the
//   sub-state interpolation LinSpaced(cnt, q0, q1) (collision_terms.cpp:846-848)
// One thread per sub-state, frames visited in topological order.
// ---------------------------------------------------------------------------------------------
#ifndef TC_FK_MINB
#define TC_FK_MINB 6
#endif
__global__ void __launch_bounds__(TPB, TC_FK_MINB) k_fk(SceneView v, Work w)
{
  if (passOverflowed(w) || static_cast<long long>(blockIdx.x) * blockDim.x >= w.counters->n_states)
    return;  // uniform over the block
  // per-warp staging tile: 32 poses, rows padded to 13 doubles (bank spread)
  __shared__ double tile[TPB / 32][32 * 13];
  // the kinematic tables are the same for every thread: one copy per block in shared memory (small scenes; large
  // ones keep reading them through L1)
  constexpr int SF = 48, SL = 32;
  __shared__ double s_origin[SF * 12], s_axis[SF * 3], s_box[SL * 6];
  __shared__ int s_parent[SF], s_type[SF], s_dof[SF], s_flink[SF], s_lframe[SL], s_sframe[SL];
  const bool staged = v.n_frames <= SF && v.n_active <= SL && v.n_active_shapes <= SL;
  if (staged)
  {
    for (int e = threadIdx.x; e < v.n_frames * 12; e += blockDim.x)
      s_origin[e] = v.frame_origin[e];
    for (int e = threadIdx.x; e < v.n_frames * 3; e += blockDim.x)
      s_axis[e] = v.frame_axis[e];
    for (int e = threadIdx.x; e < v.n_frames; e += blockDim.x)
    {
      s_parent[e] = v.frame_parent[e], s_type[e] = v.frame_type[e], s_dof[e] = v.frame_dof[e];
      s_flink[e] = v.frame_link[e];
    }
    for (int e = threadIdx.x; e < v.n_active; e += blockDim.x)
      s_lframe[e] = v.link_frame[e];
    for (int e = threadIdx.x; e < v.n_active_shapes; e += blockDim.x)
    {
      s_sframe[e] = v.active_shape_frame[e];
      const ShapeD& sh = v.shapes[v.active_shapes[e]];
      for (int r = 0; r < 3; ++r)
        s_box[6 * e + r] = sh.aabb_c[r], s_box[6 * e + 3 + r] = sh.aabb_h[r];
    }
    __syncthreads();
  }
  SceneView t = v;  // table pointers redirected to the staged copies
  if (staged)
  {
    t.frame_origin = s_origin, t.frame_axis = s_axis, t.frame_parent = s_parent, t.frame_type = s_type;
    t.frame_dof = s_dof, t.frame_link = s_flink, t.link_frame = s_lframe, t.active_shape_frame = s_sframe;
  }
  const long long n_states = w.counters->n_states;  // actual count; w.n_states is the arenas' state stride (>= count)
  const long long s_raw = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  const bool live = s_raw < n_states;
  const long long s = live ? s_raw : n_states - 1;  // idle lanes shadow the last state and write nothing
  const unsigned lane = threadIdx.x & 31;
  const long long warp_s0 = s_raw - lane;
  const int warp_n = static_cast<int>(min(32ll, n_states - warp_s0));  // live lanes of this warp (may be <= 0)
  double* my_tile = tile[threadIdx.x >> 5];
  const int k = lastLE(w.state_base, w.ep.n_terms, static_cast<int>(s));
  const int i = static_cast<int>(s) - w.state_base[k];
  const int b = k / w.ep.terms_per_traj, ti = k % w.ep.terms_per_traj;
  const int n = v.n_dof;
  const double* q0 = w.q + (static_cast<size_t>(b) * w.ep.T + w.term_t[ti]) * n;
  const bool discrete = w.ep.evaluator == TC_EVAL_DISCRETE;
  const int cnt = w.term_interp[k] ? w.term_cnt[k] : 2;
  // Frames are visited in topological order with the previous frame's pose in registers; a branch reads
  // its parent back from this thread's own record in the pose arena (no thread-local frame array).
  // moving links of the environment outside the planning group (dynamic environment): no frame, one pose for the whole
  // pass; their shapes still lead pairs in the broadphase, so their (constant) box goes into every state's column
  if (live)
    for (int a = 0; a < v.n_active_shapes; ++a)
    {
      if (v.active_shape_frame[a] >= 0)
        continue;
      const ShapeD& sh = v.shapes[v.active_shapes[a]];
      writeShapeAabb(w, Pose<double>::from12(v.link_pose + 12 * sh.link), sh.aabb_c, sh.aabb_h, a, s);
    }
  const Pose<double> base = Pose<double>::from12(v.base_pose);
  Pose<double> cur = base;
  for (int f = 0; f < v.n_frames; ++f)
  {
    const int p = t.frame_parent[f];
    Pose<double> parent;
    if (p < 0)
      parent = base;
    else if (p == f - 1)
      parent = cur;
    else if (t.frame_link[p] >= 0)
      parent = Pose<double>::from12(posePtr(w, t.frame_link[p], s));
    else
    {
      // a branching frame that carries no collision link: rebuild it from the root (rare)
      parent = base;
      for (int a = v.path_ptr[p]; a < v.path_ptr[p + 1]; ++a)
      {
        const int g = v.path_frames[a];
        double qa = 0;
        if (t.frame_type[g] != TC_JOINT_FIXED)
        {
          const int j = t.frame_dof[g];
          qa = discrete ? q0[j] : linspaced(cnt, q0[j], q0[n + j], i);
        }
        parent = frameStepQ(t, parent, g, qa);
      }
    }
    double qf = 0;
    if (t.frame_type[f] != TC_JOINT_FIXED)
    {
      const int j = t.frame_dof[f];
      qf = discrete ? q0[j] : linspaced(cnt, q0[j], q0[n + j], i);
    }
    cur = frameStepQ(t, parent, f, qf);
    // every active link riding on this frame: the warp's 32 records of one link are one contiguous run of
    // the arena, written with coalesced stores after a transpose through shared memory
    for (int l = 0; l < v.n_active; ++l)
    {
      if (t.link_frame[l] != f)
        continue;
      cur.to12(my_tile + lane * 13);
      __syncwarp();
      if (warp_n > 0)
      {
        double* dst = w.poses + (static_cast<size_t>(l) * static_cast<size_t>(w.n_states) + static_cast<size_t>(warp_s0)) * 12;
        const int total = warp_n * 12;
#pragma unroll
        for (int j = 0; j < 12; ++j)
        {
          const int e = j * 32 + static_cast<int>(lane);
          if (e < total)
            dst[e] = my_tile[(e / 12) * 13 + (e % 12)];
        }
      }
      __syncwarp();
    }
    if (!live)
      continue;
    for (int a = 0; a < v.n_active_shapes; ++a)
    {
      if (t.active_shape_frame[a] != f)
        continue;
      double bc[3], bh[3];
      if (staged)
      {
#pragma unroll
        for (int r = 0; r < 3; ++r)
          bc[r] = s_box[6 * a + r], bh[r] = s_box[6 * a + 3 + r];
      }
      else
      {
        const ShapeD& sh = v.shapes[v.active_shapes[a]];
#pragma unroll
        for (int r = 0; r < 3; ++r)
          bc[r] = sh.aabb_c[r], bh[r] = sh.aabb_h[r];
      }
      writeShapeAabb(w, cur, bc, bh, a, s);
    }
  }
}

// tc_contact_test: the caller's link poses -> pose arena + world AABBs (what k_fk leaves, without kinematics).
// One thread per arena state; term k owns state k (discrete manager) or states 2k (T0) and 2k + 1 (T1) (cast manager).
__global__ void __launch_bounds__(TPB) k_pose_in(SceneView v, Work w)
{
  const long long s = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (passOverflowed(w) || s >= w.counters->n_states)
    return;
  const int spt = w.cast ? 2 : 1;
  const long long k = s / spt;
  const int which = static_cast<int>(s % spt);
  const double* src = w.pose_in + (static_cast<size_t>(which) * w.ep.n_terms + static_cast<size_t>(k)) * v.n_active * 12;
  for (int l = 0; l < v.n_active; ++l)
  {
    const Pose<double> T = Pose<double>::from12(src + 12 * l);
    T.to12(w.poses + (static_cast<size_t>(l) * static_cast<size_t>(w.n_states) + static_cast<size_t>(s)) * 12);
    for (int a = 0; a < v.n_active_shapes; ++a)
    {
      const ShapeD& sh = v.shapes[v.active_shapes[a]];
      if (sh.link != l)
        continue;
      writeShapeAabb(w, T, sh.aabb_c, sh.aabb_h, a, s);
    }
  }
  for (int a = 0; a < v.n_active_shapes; ++a)
  {
    const ShapeD& sh = v.shapes[v.active_shapes[a]];
    if (sh.link >= v.n_active)  // leading link without per-state poses: the scene's pose
      writeShapeAabb(w, Pose<double>::from12(v.link_pose + 12 * sh.link), sh.aabb_c, sh.aabb_h, a, s);
  }
}

// every item of the shape pairs with a curved core -> the deep list (k_gjk skips those buckets)
__global__ void __launch_bounds__(256) k_route_curved(SceneView v, Work w)
{
  if (passOverflowed(w))
    return;
  const int sp = v.curved_sp[blockIdx.y];
  const int off = w.sp_offset[sp], n = w.sp_count[sp];
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
  {
    const int u = w.items[off + i];
    const unsigned long long d = atomicAdd(reinterpret_cast<unsigned long long*>(&w.counters->n_curved), 1ull);
    w.curved_items[2 * d] = u;
    w.curved_items[2 * d + 1] = sp | DEEP_COUNT_GJK;
  }
}

// deep-expansion path: same arithmetic, polytope in global scratch, vertices read from global memory
template <class R>
__global__ void __launch_bounds__(64) k_narrow_deep(SceneView v, Work w, int list, long long n_deep, unsigned char* scratch,
                                                    size_t bytes_per_thread)
{
  const int32_t* items = list ? w.curved_items : w.deep_items;
  if (passOverflowed(w))
    return;
  if (n_deep < 0)
    n_deep = list ? w.counters->n_curved : w.counters->n_deep;  // speculative pass: the host has not read the count
  const int tid = blockIdx.x * blockDim.x + threadIdx.x;
  const int nthreads = gridDim.x * blockDim.x;
  unsigned char* mine = scratch + static_cast<size_t>(tid) * bytes_per_thread;
  Vec3<R>* pw = reinterpret_cast<Vec3<R>*>(mine);
  Vec3<R>* pa = pw + EPA_DEEP_VERTS;
  Vec3<R>* fn = pa + EPA_DEEP_VERTS;
  R* fd = reinterpret_cast<R*>(fn + EPA_DEEP_FACES);
  int* fv = reinterpret_cast<int*>(fd + EPA_DEEP_FACES);
  int* edges = fv + EPA_DEEP_FACES;
  const long long rounds = (n_deep + nthreads - 1) / nthreads;
  for (long long r = 0; r < rounds; ++r)
  {
    const long long it = r * nthreads + tid;
    DistOut<R> d;
    d.status = ND_NONE;
    d.gjk_iters = 0, d.epa_iters = 0, d.scanned = 0;
    d.used_epa = false;
    d.converged = true;
    int u = -1, sp = 0;
    ScanCount gjk_scanned = 0;
    bool count_gjk = false;
    if (it < n_deep)
    {
      u = items[2 * it];
      const int sp_raw = items[2 * it + 1];
      count_gjk = (sp_raw & DEEP_COUNT_GJK) != 0;
      sp = sp_raw & (DEEP_COUNT_GJK - 1);
      const ShapePairD spd = v.spairs[sp];
      const PairD pr = v.pairs[spd.pair];
      const ShapeD& A = v.shapes[spd.sa];
      const ShapeD& B = v.shapes[spd.sb];
      Placed<R> PA, PB;
      const int kt = w.unit_term[u];
      placeItem<R>(v, w, w.state_base[kt] + (u - w.unit_base[kt]), spd, pr, A, B, sceneVerts<R>(v) + A.vert_off,
                   sceneVerts<R>(v) + B.vert_off, PA, PB, A.lut_bytes ? v.lut + A.lut_off : nullptr,
                   B.lut_bytes ? v.lut + B.lut_off : nullptr);
      EpaStore<R> store{ pw, pa, fn, fd, fv, edges, EPA_DEEP_VERTS, EPA_DEEP_FACES, nullptr };
      Simplex<R> S;
      distanceGjk<R>(PA, PB, static_cast<R>(pr.thr), S, d);
      gjk_scanned = d.scanned;
      if (d.status == ND_NEED_EPA)
      {
        bool done = false;
        if (count_gjk)
        {
          // a curved-core item has not been through k_epa_wp: try the thread-local register-mask polytope first (same
          // arithmetic and tie-breaks as the list form below, tests/test_narrow_hostcheck.py); the global slab, whose
          // every face visit is an L2 round trip, is only for the few that outgrow it
          R e_raw[(2 * EPA_FAST_VERTS + EPA_FAST_FACES) * 3 + EPA_FAST_FACES];
          int e_int[2 * EPA_FAST_FACES];
          unsigned long long e_row[EPA_FAST_VERTS];
#pragma unroll 1
          for (int i = 0; i < EPA_FAST_VERTS; ++i)
            e_row[i] = 0;
          Vec3<R>* e_pw = reinterpret_cast<Vec3<R>*>(e_raw);
          EpaStore<R> fast{ e_pw, e_pw + EPA_FAST_VERTS, e_pw + 2 * EPA_FAST_VERTS,
                            e_raw + (2 * EPA_FAST_VERTS + EPA_FAST_FACES) * 3, e_int, e_int + EPA_FAST_FACES,
                            EPA_FAST_VERTS, EPA_FAST_FACES, e_row };
          DistOut<R> df = d;
          distanceEpa<R, true>(PA, static_cast<R>(A.brad), PB, static_cast<R>(B.brad), static_cast<R>(pr.thr), S, fast, df);
          if (df.status != ND_EPA_OVERFLOW)
          {
            d = df;
            done = true;
          }
        }
        if (!done)
          distanceEpa<R, false>(PA, static_cast<R>(A.brad), PB, static_cast<R>(B.brad), static_cast<R>(pr.thr), S, store, d);
      }
    }
    storeHit<R>(w, d, u, sp, d.status == ND_CONTACT);
    // an item handed over by k_epa_wp had its GJK counted by k_gjk already; a curved-core item is counted here in full
    const ScanCount sc = count_gjk ? d.scanned : d.scanned - gjk_scanned;  // both halves subtract without borrow
    addStats(w, count_gjk ? d.gjk_iters : 0, static_cast<long long>(sc & 0xffffffffull), static_cast<long long>(sc >> 32),
             d.used_epa ? 1 : 0, d.epa_iters, (u >= 0 && !d.converged && (count_gjk || d.used_epa)) ? 1 : 0);
  }
}

}  // namespace

cudaError_t launchFk(const SceneView& v, const Work& w, cudaStream_t st)
{
  ++launchCounter();
  k_unit_term<<<static_cast<unsigned>((w.n_units + 255) / 256), 256, 0, st>>>(w);
  ++launchCounter();
  k_fk<<<static_cast<unsigned>((w.n_states + TPB - 1) / TPB), TPB, 0, st>>>(v, w);
  return cudaGetLastError();
}

cudaError_t launchPoseIn(const SceneView& v, const Work& w, cudaStream_t st)
{
  ++launchCounter();
  k_unit_term<<<static_cast<unsigned>((w.n_units + 255) / 256), 256, 0, st>>>(w);
  ++launchCounter();
  k_pose_in<<<static_cast<unsigned>((w.n_states + TPB - 1) / TPB), TPB, 0, st>>>(v, w);
  return cudaGetLastError();
}

cudaError_t launchRouteCurved(const SceneView& v, const Work& w, cudaStream_t st)
{
  if (v.n_curved_sp <= 0 || w.n_items_padded <= 0)
    return cudaSuccess;
  ++launchCounter();
  k_route_curved<<<dim3(8, static_cast<unsigned>(v.n_curved_sp)), 256, 0, st>>>(v, w);
  return cudaGetLastError();
}

size_t deepScratchBytesPerThread(int precision)
{
  const size_t r = precision == TC_PRECISION_F32 ? 4 : 8;
  size_t b = (2 * EPA_DEEP_VERTS + EPA_DEEP_FACES) * 3 * r + EPA_DEEP_FACES * r + 2 * EPA_DEEP_FACES * sizeof(int);
  return (b + 15) / 16 * 16;
}

cudaError_t launchNarrowDeep(const SceneView& v, const Work& w, int precision, int list, long long n_deep, void* scratch,
                             size_t bytes_per_thread, int n_threads, cudaStream_t st)
{
  if (n_deep == 0)
    return cudaSuccess;
  const unsigned grid = static_cast<unsigned>(n_threads / 64);
  ++launchCounter();
  // always FP64: in FP32 mode this kernel is where the items float EPA could not close get their answer
  (void)precision;
  k_narrow_deep<double><<<grid, 64, 0, st>>>(v, w, list, n_deep, static_cast<unsigned char*>(scratch), bytes_per_thread);
  return cudaGetLastError();
}

}  // namespace tc
