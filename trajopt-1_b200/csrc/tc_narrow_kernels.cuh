// tc_narrow_kernels.cuh — device helpers shared by the narrow-phase kernels (tc_narrow.cu: k_gjk, k_epa_wp; tc_strict.cu:
// k_narrow_deep): building the two placed shapes of an item from the pose arena, appending a raw hit with one atomic
// per warp, and the warp-summed statistics.
#pragma once
#include "tc_internal.hpp"

namespace tc
{
namespace
{
// deep_items[2 i + 1] = shape pair | DEEP_COUNT_GJK when the item did not go through k_gjk first (curved cores), so
// that k_narrow_deep also counts its GJK iterations
constexpr int DEEP_COUNT_GJK = 1 << 30;

template <class R>
__device__ __forceinline__ const Vtx<R>* sceneVerts(const SceneView& v);
template <>
__device__ __forceinline__ const Vtx<double>* sceneVerts<double>(const SceneView& v)
{
  return v.verts;
}
template <>
__device__ __forceinline__ const Vtx<float>* sceneVerts<float>(const SceneView& v)
{
  return v.verts_f32;
}

template <class R>
__device__ __forceinline__ Pose<R> loadPose(const double* p)
{
  return Pose<R>::from12(p);
}

// builds the two placed shapes of an item (T0 sub-state s0, shape pair sp)
template <class R>
__device__ __forceinline__ void placeItem(const SceneView& v, const Work& w, long long s0, const ShapePairD& spd,
                                          const PairD& pr, const ShapeD& A, const ShapeD& B, const Vtx<R>* va,
                                          const Vtx<R>* vb, Placed<R>& PA, Placed<R>& PB, const void* la = nullptr,
                                          const void* lb = nullptr)
{
  PA.sh = makeCore<R>(A, va, la);
  if (pr.a < v.n_active)
  {
    const double* pa = posePtr(w, pr.a, s0);
    PA.p0 = loadPose<R>(pa);
    PA.cast = w.cast != 0;
    PA.p1 = w.cast ? loadPose<R>(pa + 12) : PA.p0;
  }
  else
  {
    // a moving link of the environment outside the planning group leads this pair: one pose for the whole pass
    PA.p0 = loadPose<R>(v.link_pose + 12 * pr.a);
    PA.cast = false;
    PA.p1 = PA.p0;
  }
  PB.sh = makeCore<R>(B, vb, lb);
  if (pr.b < v.n_active)
  {
    const double* pb = posePtr(w, pr.b, s0);
    PB.p0 = loadPose<R>(pb);
    PB.cast = w.cast != 0;
    PB.p1 = w.cast ? loadPose<R>(pb + 12) : PB.p0;
  }
  else
  {
    PB.p0 = loadPose<R>(v.link_pose + 12 * pr.b);
    PB.cast = false;
    PB.p1 = PB.p0;
  }
  (void)spd;
}

template <class R>
__device__ __forceinline__ void storeHit(const Work& w, const DistOut<R>& d, int u, int sp, bool hit)
{
  const unsigned m = __ballot_sync(0xffffffffu, hit);
  if (m == 0)
    return;
  const unsigned lane = threadIdx.x & 31;
  const int leader = __ffs(m) - 1;
  unsigned long long base = 0;
  if (lane == leader)
    base = atomicAdd(reinterpret_cast<unsigned long long*>(&w.counters->n_raw),
                     static_cast<unsigned long long>(__popc(m)));
  base = __shfl_sync(0xffffffffu, base, leader);
  if (hit)
  {
    RawContact rc;
    rc.unit = u;
    rc.sp = sp;
    rc.flags = (d.used_epa ? TC_FLAG_EPA : 0) | (d.converged ? 0 : TC_FLAG_UNCONVERGED);
    rc.pad = 0;
    rc.distance = static_cast<double>(d.distance);
    rc.normal[0] = d.normal.x, rc.normal[1] = d.normal.y, rc.normal[2] = d.normal.z;
    rc.pa[0] = d.pa.x, rc.pa[1] = d.pa.y, rc.pa[2] = d.pa.z;
    rc.pb[0] = d.pb.x, rc.pb[1] = d.pb.y, rc.pb[2] = d.pb.z;
    w.raw[base + __popc(m & ((1u << lane) - 1u))] = rc;
  }
}

__device__ __forceinline__ long long warpSum(long long x)
{
#pragma unroll
  for (int o = 16; o > 0; o >>= 1)
    x += __shfl_xor_sync(0xffffffffu, x, o);
  return x;
}

// scanned / compared: the two halves of a ScanCount, unpacked PER ITEM by the caller (their sums exceed 32 bits)
__device__ __forceinline__ void addStats(const Work& w, long long gjk, long long scanned, long long compared,
                                         long long nepa, long long epa_it, long long unconv)
{
  gjk = warpSum(gjk);
  scanned = warpSum(scanned);
  compared = warpSum(compared);
  nepa = warpSum(nepa);
  epa_it = warpSum(epa_it);
  unconv = warpSum(unconv);
  if ((threadIdx.x & 31) == 0)
  {
    typedef unsigned long long ull;
    if (gjk)
      atomicAdd(reinterpret_cast<ull*>(&w.counters->gjk_iters), static_cast<ull>(gjk));
    if (scanned)
      atomicAdd(reinterpret_cast<ull*>(&w.counters->support_verts), static_cast<ull>(scanned));
    if (compared)
      atomicAdd(reinterpret_cast<ull*>(&w.counters->support_compared), static_cast<ull>(compared));
    if (nepa)
      atomicAdd(reinterpret_cast<ull*>(&w.counters->n_epa), static_cast<ull>(nepa));
    if (epa_it)
      atomicAdd(reinterpret_cast<ull*>(&w.counters->epa_iters), static_cast<ull>(epa_it));
    if (unconv)
      atomicAdd(reinterpret_cast<ull*>(&w.counters->n_unconverged), static_cast<ull>(unconv));
  }
}

}  // namespace
}  // namespace tc
