// tc_narrow.cu — K3: convex-convex GJK/EPA narrow phase over the bucketed pair list.
//
// The narrow-phase half of contact_manager_->contactTest(...)
//   trajopt/src/collision_terms.cpp:667,889,1134,1157
//   trajopt_ifopt/.../discrete_collision_evaluators.cpp:166; continuous_collision_evaluators.cpp:228,244,449
// One CTA = a run of items of ONE shape pair.  Both hulls are staged global -> shared with TMA bulk copies
// (cp.async.bulk + mbarrier); every lane then runs its own GJK (+EPA) against the staged vertices, so
// vertex reads are warp-wide broadcasts.  Hits are appended to the raw contact arena with one atomic
// per warp (ballot + popc).
#include <cstdlib>

#include "tc_narrow_kernels.cuh"

#ifndef TC_GJK_DYN
#define TC_GJK_DYN 1
#endif
// threads per narrow-phase CTA (the bucket padding unit stays CHUNK); a CTA works through `run_items` items
#ifndef TC_GJK_NT
#define TC_GJK_NT 128
#endif

namespace tc
{
namespace
{
__device__ __forceinline__ uint32_t smemAddr(const void* p) { return static_cast<uint32_t>(__cvta_generic_to_shared(p)); }

__device__ __forceinline__ void mbarInit(uint64_t* bar, int count)
{
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smemAddr(bar)), "r"(count));
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbarExpectTx(uint64_t* bar, uint32_t bytes)
{
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smemAddr(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulkCopyG2S(void* dst, const void* src, uint32_t bytes, uint64_t* bar)
{
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   smemAddr(dst)),
               "l"(src), "r"(bytes), "r"(smemAddr(bar))
               : "memory");
}
__device__ __forceinline__ void mbarWait(uint64_t* bar, uint32_t parity)
{
  uint32_t done = 0;
  while (!done)
  {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
        "selp.u32 %0, 1, 0, p;\n"
        "}\n"
        : "=r"(done)
        : "r"(smemAddr(bar)), "r"(parity)
        : "memory");
  }
}

// ---------------------------------------------------------------------------------------------
// Run = the part of ONE shape pair's bucket that falls into this CTA's range of chunks.  Per run the two
// hulls are staged with TMA bulk copies; the warps then take 32-item groups of the run from a shared cursor
// (s_next) until it is drained, so a warp that drew short items does not wait at the end-of-run barrier for
// one that drew long ones.
// ---------------------------------------------------------------------------------------------
struct RunInfo
{
  int sp, q0, q1;  // shape pair, item range [q0, q1) of the run (q1 = end of the CTA range or of the bucket)
};

__device__ __forceinline__ RunInfo findRun(const SceneView& v, const Work& w, int item0, int range_end, RunInfo* s_run,
                                           int* s_next)
{
  if (threadIdx.x == 0)
  {
    // bucket of this chunk: last shape pair whose (padded) offset is <= item0
    int lo = 0, hi = v.n_spairs;
    while (hi - lo > 1)
    {
      const int mid = (lo + hi) >> 1;
      if (w.sp_offset[mid] <= item0)
        lo = mid;
      else
        hi = mid;
    }
    const int bucket_end = w.sp_offset[lo] + (w.sp_count[lo] + CHUNK - 1) / CHUNK * CHUNK;
    s_run->sp = lo;
    s_run->q0 = item0;
    s_run->q1 = bucket_end < range_end ? bucket_end : range_end;
    *s_next = item0;
  }
  __syncthreads();
  const RunInfo r = *s_run;
  return r;
}

// next 32-item group of the run for this warp: DYN takes it from the shared cursor, otherwise warps stride through
// the run CHUNK items at a time (static assignment)
template <int DYN>
__device__ __forceinline__ int nextGroup(int* s_next, int round, int nt)
{
  if (DYN)
  {
    int group = 0;
    if ((threadIdx.x & 31) == 0)
      group = atomicAdd(s_next, 32);
    return __shfl_sync(0xffffffffu, group, 0);
  }
  return *s_next + round * nt + (static_cast<int>(threadIdx.x) & ~31);
}

template <class R>
__device__ __forceinline__ void stageHulls(const SceneView& v, const ShapeD& A, const ShapeD& B, Vtx<R>* sv,
                                           uint64_t* bar, uint32_t parity)
{
  const int na = A.vert_cnt, nb = B.vert_cnt;
  if (na + nb == 0)
    return;
  if (threadIdx.x == 0)
  {
    // the previous run's generic-proxy reads of this buffer are ordered before the async-proxy writes
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    const uint32_t ba = static_cast<uint32_t>(na * sizeof(Vtx<R>)), bb = static_cast<uint32_t>(nb * sizeof(Vtx<R>));
    mbarExpectTx(bar, ba + bb + static_cast<uint32_t>(A.lut_bytes + B.lut_bytes));
    if (na)
      bulkCopyG2S(sv, sceneVerts<R>(v) + A.vert_off, ba, bar);
    if (nb)
      bulkCopyG2S(sv + na, sceneVerts<R>(v) + B.vert_off, bb, bar);
    // support look-up tables behind the vertices (all sizes are multiples of 16 bytes)
    unsigned char* lut = reinterpret_cast<unsigned char*>(sv + na + nb);
    if (A.lut_bytes)
      bulkCopyG2S(lut, v.lut + A.lut_off, static_cast<uint32_t>(A.lut_bytes), bar);
    if (B.lut_bytes)
      bulkCopyG2S(lut + A.lut_bytes, v.lut + B.lut_off, static_cast<uint32_t>(B.lut_bytes), bar);
  }
  mbarWait(bar, parity);
}

// staged look-up tables of a run (null for shapes without one)
template <class R>
__device__ __forceinline__ const void* stagedLutA(const ShapeD& A, const ShapeD& B, const Vtx<R>* sv)
{
  return A.lut_bytes ? static_cast<const void*>(sv + A.vert_cnt + B.vert_cnt) : nullptr;
}
template <class R>
__device__ __forceinline__ const void* stagedLutB(const ShapeD& A, const ShapeD& B, const Vtx<R>* sv)
{
  return B.lut_bytes ? static_cast<const void*>(reinterpret_cast<const unsigned char*>(sv + A.vert_cnt + B.vert_cnt) + A.lut_bytes)
                     : nullptr;
}

// A pose parked in shared memory: element k (3x4 row-major, as in the pose arena) of this thread at base[k * STRIDE].
// STRIDE = threads per CTA for a per-thread pose (lanes read consecutive 8-byte words: conflict-free), 1 for a pose
// the whole CTA shares (a static partner: every lane reads the same word, a broadcast).  The loads are volatile so that
// the compiler re-reads shared memory instead of carrying 12 doubles (24 registers) per pose through the GJK loop.
template <class R, int STRIDE>
struct SmemPose
{
  const R* base;
  __device__ __forceinline__ R g(int k) const { return *reinterpret_cast<const volatile R*>(base + k * STRIDE); }
  __device__ __forceinline__ Vec3<R> rot(Vec3<R> v) const
  {
    return Vec3<R>(g(0) * v.x + g(1) * v.y + g(2) * v.z, g(4) * v.x + g(5) * v.y + g(6) * v.z,
                   g(8) * v.x + g(9) * v.y + g(10) * v.z);
  }
  __device__ __forceinline__ Vec3<R> rotT(Vec3<R> v) const
  {
    return Vec3<R>(g(0) * v.x + g(4) * v.y + g(8) * v.z, g(1) * v.x + g(5) * v.y + g(9) * v.z,
                   g(2) * v.x + g(6) * v.y + g(10) * v.z);
  }
  __device__ __forceinline__ Vec3<R> apply(Vec3<R> v) const { return rot(v) + Vec3<R>(g(3), g(7), g(11)); }
  __device__ __forceinline__ Vec3<R> applyInv(Vec3<R> p) const { return rotT(p - Vec3<R>(g(3), g(7), g(11))); }
};

#ifndef TC_GJK_SMEM_POSES
#define TC_GJK_SMEM_POSES 1
#endif

// K3a: GJK for every bucketed item.  Separated pairs are finished here; touching / penetrating pairs hand
// their simplex to the EPA stage through a per-shape-pair packed arena (SoA, 24 doubles per item).
template <class R, int MINB>
__global__ void __launch_bounds__(TC_GJK_NT, MINB * 128 / TC_GJK_NT) k_gjk(SceneView v, Work w, int run_items)
{
  extern __shared__ __align__(16) unsigned char smem_raw[];
  __shared__ __align__(8) uint64_t bar;
  __shared__ RunInfo s_run;
  __shared__ int s_next;
  Vtx<R>* sv = reinterpret_cast<Vtx<R>*>(smem_raw);
#if TC_GJK_SMEM_POSES
  // behind the staged hulls and tables: the two poses of every thread's moving shape, [24][threads]; and the one pose
  // of a static partner, shared by the CTA
  __shared__ R s_bpose[12];
  R* s_apose = reinterpret_cast<R*>(smem_raw + (static_cast<size_t>(2 * v.max_hull_verts) * sizeof(Vtx<R>) +
                                                2 * static_cast<size_t>(v.max_lut_bytes) + 15) / 16 * 16);
#endif
  if (passOverflowed(w))
    return;
  const int range_end = static_cast<int>(min(static_cast<long long>(blockIdx.x + 1) * run_items, w.counters->n_items_padded));
  if (threadIdx.x == 0)
    mbarInit(&bar, 1);
  long long st_gjk = 0, st_scan = 0, st_cmp = 0, st_unconv = 0;
  uint32_t phase = 0;
  for (int item0 = blockIdx.x * run_items; item0 < range_end;)
  {
    const RunInfo run = findRun(v, w, item0, range_end, &s_run, &s_next);
    const int sp = run.sp;
    const ShapePairD spd = v.spairs[sp];
    const PairD pr = v.pairs[spd.pair];
    const ShapeD& A = v.shapes[spd.sa];
    const ShapeD& B = v.shapes[spd.sb];
    const int valid_end = min(run.q1, w.sp_offset[sp] + w.sp_count[sp]);
    if (A.type >= SH_CYLINDER || B.type >= SH_CYLINDER)
    {
      // curved cores (cylinder, cone): GJK/EPA on a curved surface has no closest vertex to lock on to and amplifies
      // last-bit differences of its inputs to 1e-5 in the normal; these buckets go to the contraction-free kernel of
      // tc_strict.cu, which reproduces the reference arithmetic operation for operation
      __syncthreads();
      item0 = run.q1;
      continue;
    }
#if TC_GJK_SMEM_POSES
    if (pr.b >= v.n_active)
    {
      if (threadIdx.x < 12)
        s_bpose[threadIdx.x] = static_cast<R>(v.link_pose[12 * pr.b + threadIdx.x]);
      __syncthreads();  // (the previous run's readers passed the barrier at the end of their run)
    }
#endif
    if (A.vert_cnt + B.vert_cnt > 0)
    {
      stageHulls<R>(v, A, B, sv, &bar, phase & 1);
      ++phase;
    }
    const R thr = static_cast<R>(pr.thr);
    for (int round = 0;;)
    {
      const int group = nextGroup<TC_GJK_DYN>(&s_next, round++, TC_GJK_NT);
      if (group >= valid_end)
        break;
      const int slot = group + (threadIdx.x & 31);
      const bool live = slot < valid_end;
      DistOut<R> d;
      clearDist(d);
      int u = -1;
      Simplex<R> S;
      S.n = 0;
      if (live)
      {
        u = w.items[slot];
#if TC_GJK_SMEM_POSES
        if (pr.b >= v.n_active && pr.a < v.n_active && A.type == SH_HULL && B.type == SH_HULL)
        {
          // the common case (group hull against a static hull): poses live in shared memory, not in registers
          const double* pa = posePtr(w, pr.a, w.item_state[slot]);
          R* mine = s_apose + threadIdx.x;
          const int np = w.cast ? 24 : 12;
#pragma unroll 4
          for (int k = 0; k < np; ++k)
            mine[k * TC_GJK_NT] = static_cast<R>(pa[k]);
          Placed<R, SmemPose<R, TC_GJK_NT>> QA;
          Placed<R, SmemPose<R, 1>> QB;
          QA.sh = makeCore<R>(A, sv, stagedLutA<R>(A, B, sv));
          QB.sh = makeCore<R>(B, sv + A.vert_cnt, stagedLutB<R>(A, B, sv));
          QA.sh.type = SH_HULL;
          QB.sh.type = SH_HULL;
          QA.p0.base = mine;
          QB.p0.base = s_bpose;
          QB.p1.base = s_bpose;
          QB.cast = false;
          if (w.cast)
          {
            QA.p1.base = mine + 12 * TC_GJK_NT;
            QA.cast = true;
            distanceGjk<R>(QA, QB, thr, S, d);
          }
          else
          {
            QA.p1.base = mine;
            QA.cast = false;
            distanceGjk<R>(QA, QB, thr, S, d);
          }
          st_gjk += d.gjk_iters;
          st_scan += static_cast<long long>(d.scanned & 0xffffffffull);
          st_cmp += static_cast<long long>(d.scanned >> 32);
          st_unconv += (d.status != ND_NEED_EPA && !d.converged) ? 1 : 0;
        }
        else
        {
#endif
        Placed<R> PA, PB;
        placeItem<R>(v, w, w.item_state[slot], spd, pr, A, B, sv, sv + A.vert_cnt, PA, PB, stagedLutA<R>(A, B, sv),
                     stagedLutB<R>(A, B, sv));
        if (pr.b >= v.n_active)
        {
          // static partner (uniform over the run, the common case): `cast` is a literal here, so this copy of the GJK
          // loop carries neither the partner's second pose nor the swept branch of its support
          PB.cast = false;
          PB.p1 = PB.p0;
          if (A.type == SH_HULL && B.type == SH_HULL)
          {
            // ... and both shapes are hulls: no primitive support in these copies; one for a swept moving shape, one
            // for a discrete test (no second pose of the moving shape either)
            PA.sh.type = SH_HULL;
            PB.sh.type = SH_HULL;
            if (w.cast)
              distanceGjk<R>(PA, PB, thr, S, d);
            else
            {
              PA.cast = false;
              PA.p1 = PA.p0;
              distanceGjk<R>(PA, PB, thr, S, d);
            }
          }
          else
            distanceGjk<R>(PA, PB, thr, S, d);
        }
        else
          distanceGjk<R>(PA, PB, thr, S, d);
        st_gjk += d.gjk_iters;
        st_scan += static_cast<long long>(d.scanned & 0xffffffffull);
        st_cmp += static_cast<long long>(d.scanned >> 32);
        st_unconv += (d.status != ND_NEED_EPA && !d.converged) ? 1 : 0;
#if TC_GJK_SMEM_POSES
        }
#endif
      }
#ifdef TC_NARROW_HIST
      {
        int mx = live ? d.gjk_iters : 0;
        for (int o = 16; o; o >>= 1)
          mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, o));
        if ((threadIdx.x & 31) == 0)
          atomicAdd(reinterpret_cast<unsigned long long*>(&w.counters->dbg[1]), static_cast<unsigned long long>(32 * mx));
      }
#endif
      storeHit<R>(w, d, u, sp, d.status == ND_CONTACT);
      if (sizeof(R) == sizeof(float))
      {
        // float GJK that ended without an error bound (tc_narrow.cuh, gjk): re-run in FP64 with the overflow list
        const bool redo = d.status == ND_EPA_OVERFLOW;
        const unsigned mr = __ballot_sync(0xffffffffu, redo);
        if (mr)
        {
          const unsigned lane = threadIdx.x & 31;
          const int leader = __ffs(mr) - 1;
          unsigned long long dbase = 0;
          if (lane == static_cast<unsigned>(leader))
            dbase = atomicAdd(reinterpret_cast<unsigned long long*>(&w.counters->n_deep),
                              static_cast<unsigned long long>(__popc(mr)));
          dbase = __shfl_sync(0xffffffffu, dbase, leader);
          if (redo)
          {
            const unsigned long long ds = dbase + __popc(mr & ((1u << lane) - 1u));
            w.deep_items[2 * ds] = u;
            w.deep_items[2 * ds + 1] = sp;
          }
        }
      }
      const bool need = d.status == ND_NEED_EPA;
      const unsigned m = __ballot_sync(0xffffffffu, need);
      if (m)
      {
        const unsigned lane = threadIdx.x & 31;
        const int leader = __ffs(m) - 1;
        int cur = 0;
        if (lane == leader)
          cur = atomicAdd(&w.epa_cursor[sp], __popc(m));
        cur = __shfl_sync(0xffffffffu, cur, leader);
        if (need)
        {
          EpaRec& rec = w.epa_rec[static_cast<size_t>(w.sp_offset[sp]) + cur + __popc(m & ((1u << lane) - 1u))];
          rec.unit = u;
          rec.state = w.item_state[slot];
          rec.nflag = S.n | (d.converged ? 0 : 8);
#ifdef TC_NARROW_HIST
          rec.pad = d.gjk_iters;
#else
          rec.pad = 0;
#endif
#pragma unroll
          for (int i = 0; i < 4; ++i)
          {
            rec.s[6 * i + 0] = S.w[i].x, rec.s[6 * i + 1] = S.w[i].y, rec.s[6 * i + 2] = S.w[i].z;
            rec.s[6 * i + 3] = S.a[i].x, rec.s[6 * i + 4] = S.a[i].y, rec.s[6 * i + 5] = S.a[i].z;
          }
        }
      }
    }
    __syncthreads();
    item0 = run.q1;
  }
  addStats(w, st_gjk, st_scan, st_cmp, 0, 0, st_unconv);
}

// ---------------------------------------------------------------------------------------------
// Warp-persistent EPA.  Every WARP draws 32-item groups from ONE global cursor over the padded item space, so there
// is no CTA barrier to wait at and no tail of long CTAs; consecutive groups mostly belong to the same shape pair.  (The
// same form of the GJK kernel was measured slower — instruction fetch, DESIGN.md 4.1 — and is not kept.)
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ int bucketOf(const SceneView& v, const Work& w, int item)
{
  int lo = 0, hi = v.n_spairs;
  while (hi - lo > 1)
  {
    const int mid = (lo + hi) >> 1;
    if (w.sp_offset[mid] <= item)
      lo = mid;
    else
      hi = mid;
  }
  return lo;
}

template <class R, int MINB>
__global__ void __launch_bounds__(128, MINB) k_epa_wp(SceneView v, Work w)
{
  if (passOverflowed(w))
    return;
  const int lane = threadIdx.x & 31;
  const long long n_items_padded = w.counters->n_items_padded;
  // thread-local polytope store; raw arrays so that nothing is zero-filled.  Hull vertices and support look-up tables
  // are read through L1 (the table keeps the reads to a handful per support call): nothing is staged in shared memory,
  // so the whole of L1 stays with the polytopes (measured on C4: 8.1 -> 7.0 ms against per-warp TMA staging)
  R e_raw[(2 * EPA_FAST_VERTS + EPA_FAST_FACES) * 3 + EPA_FAST_FACES];
  int e_int[2 * EPA_FAST_FACES];
  unsigned long long e_row[EPA_FAST_VERTS];
#pragma unroll 1
  for (int i = 0; i < EPA_FAST_VERTS; ++i)
    e_row[i] = 0;
  Vec3<R>* e_pw = reinterpret_cast<Vec3<R>*>(e_raw);
  EpaStore<R> store{ e_pw, e_pw + EPA_FAST_VERTS, e_pw + 2 * EPA_FAST_VERTS,
                     e_raw + (2 * EPA_FAST_VERTS + EPA_FAST_FACES) * 3, e_int, e_int + EPA_FAST_FACES, EPA_FAST_VERTS,
                     EPA_FAST_FACES, e_row };
  long long st_scan = 0, st_cmp = 0, st_epa = 0, st_it = 0, st_unconv = 0;
  int sp = -1, bucket_end = 0, valid_end = 0;
  ShapePairD spd{};
  PairD pr{};
  const ShapeD* A = nullptr;
  const ShapeD* B = nullptr;
  R thr = 0;
  for (;;)
  {
    int group = 0;
    if (lane == 0)
      group = static_cast<int>(atomicAdd(reinterpret_cast<unsigned long long*>(&w.counters->epa_cursor_g), 32ull));
    group = __shfl_sync(0xffffffffu, group, 0);
    if (group >= n_items_padded)
      break;
    if (sp < 0 || group >= bucket_end)
    {
      sp = bucketOf(v, w, group);
      const int off = w.sp_offset[sp];
      bucket_end = off + (w.sp_count[sp] + CHUNK - 1) / CHUNK * CHUNK;
      valid_end = off + w.epa_cursor[sp];  // hand-off records of this bucket (none for curved cores: tc_strict.cu)
      if (group >= valid_end)
      {
        sp = -1;
        continue;
      }
      spd = v.spairs[sp];
      pr = v.pairs[spd.pair];
      A = &v.shapes[spd.sa];
      B = &v.shapes[spd.sb];
      thr = static_cast<R>(pr.thr);
    }
    if (group >= valid_end)
      continue;
    const int slot = group + lane;
    DistOut<R> d;
    clearDist(d);
    d.used_epa = true;
    int u = -1;
    if (slot < valid_end)
    {
      const EpaRec& rec = w.epa_rec[slot];
      u = rec.unit;
      const int nflag = rec.nflag;
      Simplex<R> S;
      S.n = nflag & 7;
#pragma unroll
      for (int i = 0; i < 4; ++i)
      {
        S.w[i] = Vec3<R>(static_cast<R>(rec.s[6 * i + 0]), static_cast<R>(rec.s[6 * i + 1]), static_cast<R>(rec.s[6 * i + 2]));
        S.a[i] = Vec3<R>(static_cast<R>(rec.s[6 * i + 3]), static_cast<R>(rec.s[6 * i + 4]), static_cast<R>(rec.s[6 * i + 5]));
        S.lam[i] = 0;
      }
      d.converged = (nflag & 8) == 0;
      Placed<R> PA, PB;
      placeItem<R>(v, w, rec.state, spd, pr, *A, *B, sceneVerts<R>(v) + A->vert_off, sceneVerts<R>(v) + B->vert_off, PA,
                   PB, A->lut_bytes ? v.lut + A->lut_off : nullptr, B->lut_bytes ? v.lut + B->lut_off : nullptr);
      if (pr.b >= v.n_active)
      {
        // static partner: a copy of the expansion in which the partner is known not to be swept (see k_gjk)
        PB.cast = false;
        PB.p1 = PB.p0;
        if (A->type == SH_HULL && B->type == SH_HULL)
        {
          PA.sh.type = SH_HULL;
          PB.sh.type = SH_HULL;
          if (w.cast)
            distanceEpa<R, true>(PA, static_cast<R>(A->brad), PB, static_cast<R>(B->brad), thr, S, store, d);
          else
          {
            PA.cast = false;
            PA.p1 = PA.p0;
            distanceEpa<R, true>(PA, static_cast<R>(A->brad), PB, static_cast<R>(B->brad), thr, S, store, d);
          }
        }
        else
          distanceEpa<R, true>(PA, static_cast<R>(A->brad), PB, static_cast<R>(B->brad), thr, S, store, d);
      }
      else
        distanceEpa<R, true>(PA, static_cast<R>(A->brad), PB, static_cast<R>(B->brad), thr, S, store, d);
      if (d.status != ND_EPA_OVERFLOW)
      {
        st_scan += static_cast<long long>(d.scanned & 0xffffffffull);
        st_cmp += static_cast<long long>(d.scanned >> 32);
        st_epa += 1;
        st_it += d.epa_iters;
        st_unconv += d.converged ? 0 : 1;
      }
    }
#ifdef TC_NARROW_HIST
    {
      int mx = (slot < valid_end) ? d.epa_iters : 0;
      for (int o = 16; o; o >>= 1)
        mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, o));
      if (lane == 0)
        atomicAdd(reinterpret_cast<unsigned long long*>(&w.counters->dbg[0]), static_cast<unsigned long long>(32 * mx));
      const int it = (slot < valid_end) ? d.epa_iters : 0;
      if (slot < valid_end)
        w.epa_rec[slot].pad |= it << 16;
      for (int k = 0; k < 6; ++k)
      {
        const unsigned m = __ballot_sync(0xffffffffu, it >= 11 + 4 * k);
        if (lane == 0 && m)
          atomicAdd(reinterpret_cast<unsigned long long*>(&w.counters->dbg[2 + k]), static_cast<unsigned long long>(__popc(m)));
      }
    }
#endif
    storeHit<R>(w, d, u, sp, d.status == ND_CONTACT);
    const bool deep = d.status == ND_EPA_OVERFLOW;
    const unsigned m = __ballot_sync(0xffffffffu, deep);
    if (m)
    {
      const int leader = __ffs(m) - 1;
      unsigned long long dbase = 0;
      if (lane == leader)
        dbase = atomicAdd(reinterpret_cast<unsigned long long*>(&w.counters->n_deep),
                          static_cast<unsigned long long>(__popc(m)));
      dbase = __shfl_sync(0xffffffffu, dbase, leader);
      if (deep)
      {
        const unsigned long long ds = dbase + __popc(m & ((1u << lane) - 1u));
        w.deep_items[2 * ds] = u;
        w.deep_items[2 * ds + 1] = sp;
      }
    }
  }
  addStats(w, 0, st_scan, st_cmp, st_epa, st_it, st_unconv);
}

template <class R, int ILP>
__global__ void k_fma_peak(R* out, int iters)
{
  R acc[ILP];
#pragma unroll
  for (int i = 0; i < ILP; ++i)
    acc[i] = static_cast<R>(threadIdx.x + i) * R(1e-3);
  const R a = R(1.0000001), b = R(1e-7);
  for (int it = 0; it < iters; ++it)
  {
#pragma unroll
    for (int i = 0; i < ILP; ++i)
      acc[i] = acc[i] * a + b;
  }
  R s = 0;
#pragma unroll
  for (int i = 0; i < ILP; ++i)
    s += acc[i];
  if (s == R(-1))
    out[0] = s;
}
}  // namespace

#ifdef TC_NARROW_HIST
void narrowScanDebug(unsigned long long out[4])
{
  cudaMemcpyFromSymbol(out, g_scan_dbg, 4 * sizeof(unsigned long long));
  const unsigned long long zero[4] = { 0, 0, 0, 0 };
  cudaMemcpyToSymbol(g_scan_dbg, zero, sizeof(zero));
}
#endif

// register budgets (resident CTAs of 128 threads per SM) of the two kernels; build-time tuning knobs
#ifndef TC_GJK_MINB
#define TC_GJK_MINB 4
#endif
#ifndef TC_EPA_MINB
#define TC_EPA_MINB 4
#endif

template <class R, int GB, int EB>
static cudaError_t launchNarrowT(const SceneView& v, const Work& w, cudaStream_t st, cudaEvent_t mid)
{
  const int run_g = GJK_RUN_ITEMS;  // items per GJK CTA (hulls are staged once per run of one shape pair)
  const unsigned grid_g = static_cast<unsigned>((w.n_items_padded + run_g - 1) / run_g);
  size_t smem = static_cast<size_t>(2 * v.max_hull_verts) * sizeof(Vtx<R>) + 2 * static_cast<size_t>(v.max_lut_bytes) + 16;
#if TC_GJK_SMEM_POSES
  smem = (smem + 15) / 16 * 16 + static_cast<size_t>(24 * TC_GJK_NT) * sizeof(R);  // + the threads' two poses
#endif
  cudaFuncSetAttribute(k_gjk<R, GB>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
  cudaError_t e = launchFill(w.epa_cursor, 0, v.n_spairs, st);
  if (e != cudaSuccess)
    return e;
  ++launchCounter();
  k_gjk<R, GB><<<grid_g, TC_GJK_NT, smem, st>>>(v, w, run_g);
  if (mid)
    cudaEventRecord(mid, st);
  // EPA: persistent warps (one CTA set per SM) drawing 32-item groups from one global cursor
  static int n_sm = 0;
  if (!n_sm)
  {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, dev);
  }
  ++launchCounter();
  cudaFuncSetAttribute(k_epa_wp<R, EB>, cudaFuncAttributePreferredSharedMemoryCarveout, 0);
  k_epa_wp<R, EB><<<n_sm * EB, 128, 0, st>>>(v, w);
  return cudaGetLastError();
}

cudaError_t launchNarrow(const SceneView& v, const Work& w, int precision, cudaStream_t st, cudaEvent_t mid)
{
  if (w.n_items_padded <= 0)
    return cudaSuccess;
  if (precision == TC_PRECISION_F32)
    return launchNarrowT<float, 3, 2>(v, w, st, mid);
  return launchNarrowT<double, TC_GJK_MINB, TC_EPA_MINB>(v, w, st, mid);
}

cudaError_t measureFmaPeak(int precision, double* tflops)
{
  void* out = nullptr;
  cudaError_t e = cudaMalloc(&out, 64);
  if (e != cudaSuccess)
    return e;
  cudaDeviceProp prop;
  int dev = 0;
  cudaGetDevice(&dev);
  cudaGetDeviceProperties(&prop, dev);
  const int blocks = prop.multiProcessorCount * 8, threads = 256, iters = 1 << 14;
  constexpr int ILP = 8;
  cudaEvent_t a, b;
  cudaEventCreate(&a);
  cudaEventCreate(&b);
  float best = 1e30f;
  for (int rep = 0; rep < 10; ++rep)
  {
    cudaEventRecord(a);
    if (precision == TC_PRECISION_F32)
      k_fma_peak<float, ILP><<<blocks, threads>>>(static_cast<float*>(out), iters);
    else
      k_fma_peak<double, ILP><<<blocks, threads>>>(static_cast<double*>(out), iters);
    cudaEventRecord(b);
    e = cudaEventSynchronize(b);
    if (e != cudaSuccess)
      break;
    float ms = 0;
    cudaEventElapsedTime(&ms, a, b);
    if (rep > 0 && ms < best)
      best = ms;
  }
  cudaEventDestroy(a);
  cudaEventDestroy(b);
  cudaFree(out);
  if (e != cudaSuccess)
    return e;
  const double fl = 2.0 * static_cast<double>(blocks) * threads * static_cast<double>(iters) * ILP;
  *tflops = fl / (best * 1e-3) / 1e12;
  return cudaGetLastError();
}
}  // namespace tc
