// tc_pipeline.cu — every kernel of the pass except the narrow phase:
//   K0 plan (LVS subdivision counts) + scans, K1 batched FK with per-shape AABBs, K2 AABB broadphase with
//   a warp-cooperative (ballot + popc) per-shape-pair bucket list, contact finalisation + key sort,
//   K4 gradient rows / constraint rows / GradientResultsSets / cost.
// Reference call sites are cited at each kernel.
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>

#include "tc_internal.hpp"

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

// ---------------------------------------------------------------------------------------------
// K0: LVS subdivision of every term
//   dist = ||q1 - q0||, cnt = dist > LVS ? ceil(dist / LVS) + 1 : 2
//   trajopt/src/collision_terms.cpp:825-843,1068-1080; continuous_collision_evaluators.cpp:204-212,423-434
// ---------------------------------------------------------------------------------------------
__global__ void k_plan(SceneView v, Work w)
{
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= w.ep.n_terms)
    return;
  int cnt = 1, states = 1, units = 1, interp = 0;
  if (w.poses_given)
  {
    // tc_contact_test: one state (discrete manager) or one sweep T0 -> T1 (cast manager) per term
    cnt = w.cast ? 2 : 1;
    states = cnt;
  }
  else if (w.ep.evaluator != TC_EVAL_DISCRETE)
  {
    const int b = k / w.ep.terms_per_traj, ti = k % w.ep.terms_per_traj;
    const int n = v.n_dof;
    const double* q0 = w.q + (static_cast<size_t>(b) * w.ep.T + w.term_t[ti]) * n;
    const double* q1 = q0 + n;
    double d2 = 0;
    for (int j = 0; j < n; ++j)
    {
      const double d = q1[j] - q0[j];
      d2 = __dadd_rn(d2, __dmul_rn(d, d));  // no FMA contraction: cnt is a discrete decision
    }
    const double dist = sqrt(d2);
    cnt = 2;
    if (dist > w.ep.lvs)
    {
      const double c = ceil(dist / w.ep.lvs) + 1.0;
      if (c > 16384.0)
      {
        atomicExch(reinterpret_cast<unsigned long long*>(&w.counters->overflow), 2ull);
        cnt = 16384;
      }
      else
        cnt = static_cast<int>(c);
    }
    if (w.ep.evaluator == TC_EVAL_LVS_DISCRETE)
    {
      states = units = cnt;
      interp = 1;
    }
    else if (!w.single_sweep && dist > w.ep.lvs)
    {
      // the reference branches on dist > length, not on cnt (collision_terms.cpp:1107,
      // continuous_collision_evaluators.cpp:204): when dist / length rounds to 1.0, cnt is 2 and the one sweep still
      // takes the interpolated path (cc_type / cc_time rewritten by addInterpolatedCollisionResults)
      states = cnt;
      units = cnt - 1;
      interp = 1;
    }
    else
    {
      states = 2;
      units = 1;
    }
  }
  w.term_cnt[k] = cnt;
  w.term_states[k] = states;
  w.term_units[k] = units;
  w.term_interp[k] = static_cast<uint8_t>(interp);
}

// single-block exclusive scan helper: returns the exclusive prefix of `val` within the block and the
// block total (all threads)
__device__ __forceinline__ long long blockExclusive(long long val, long long& total, long long* smem /*[33]*/)
{
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  long long incl = val;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1)
  {
    const long long t = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o)
      incl += t;
  }
  if (lane == 31)
    smem[wid] = incl;
  __syncthreads();
  if (wid == 0)
  {
    long long x = (lane < (blockDim.x >> 5)) ? smem[lane] : 0;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1)
    {
      const long long t = __shfl_up_sync(0xffffffffu, x, o);
      if (lane >= o)
        x += t;
    }
    smem[lane] = x;  // inclusive warp totals
    if (lane == 31)
      smem[32] = x;
  }
  __syncthreads();
  const long long before = (wid > 0) ? smem[wid - 1] : 0;
  total = smem[32];
  __syncthreads();
  return before + incl - val;
}

__global__ void __launch_bounds__(1024) k_scan_terms(Work w)
{
  __shared__ long long sm[33];
  long long carry_s = 0, carry_u = 0;
  const int n = w.ep.n_terms;
  for (int base = 0; base < n; base += blockDim.x)
  {
    const int k = base + threadIdx.x;
    const long long a = (k < n) ? w.term_states[k] : 0, b = (k < n) ? w.term_units[k] : 0;
    long long ta, tb;
    const long long ea = blockExclusive(a, ta, sm);
    const long long eb = blockExclusive(b, tb, sm);
    if (k < n)
    {
      w.state_base[k] = static_cast<int32_t>(carry_s + ea);
      w.unit_base[k] = static_cast<int32_t>(carry_u + eb);
    }
    carry_s += ta;
    carry_u += tb;
  }
  if (threadIdx.x == 0)
  {
    w.state_base[n] = static_cast<int32_t>(carry_s);
    w.unit_base[n] = static_cast<int32_t>(carry_u);
    w.counters->n_states = carry_s;
    w.counters->n_units = carry_u;
    if (carry_s > 2000000000ll)
      w.counters->overflow = OVF_INDEX_STATES;
    else if (w.speculative && (carry_s > w.n_states || carry_u > w.n_units) && w.counters->overflow == 0)
      w.counters->overflow = OVF_SPECULATION;
  }
}

// ---------------------------------------------------------------------------------------------
// K2: AABB broadphase.  setCollisionObjectsTransform + the broadphase half of contactTest
//   (trajopt/src/collision_terms.cpp:665-667,887-889,1132-1134,1155-1157).
// One thread per (unit, active shape): its (swept) AABB stays in registers while the thread walks the
// shape's partner list (static partner boxes are warp-uniform loads).  A warp = 32 consecutive units of
// the same shape pair, so survivors are appended to that shape pair's bucket with one ballot + one
// atomic per warp.  Pass 1 counts, pass 2 fills.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ void loadBox(const Work& w, int a, long long s0, long long s1, double lo[3], double hi[3])
{
  const double* p = w.aabb + static_cast<size_t>(a) * 6 * w.n_states;
#pragma unroll
  for (int r = 0; r < 3; ++r)
  {
    lo[r] = p[static_cast<size_t>(r) * w.n_states + s0];
    hi[r] = p[static_cast<size_t>(3 + r) * w.n_states + s0];
  }
  if (s1 != s0)
  {
#pragma unroll
    for (int r = 0; r < 3; ++r)
    {
      lo[r] = fmin(lo[r], p[static_cast<size_t>(r) * w.n_states + s1]);
      hi[r] = fmax(hi[r], p[static_cast<size_t>(3 + r) * w.n_states + s1]);
    }
  }
}

// consecutive 128-unit chunks one block works through: the partner tile (and the two counter loads) are paid once per
// block instead of once per chunk
#ifndef TC_BROAD_CPB
#define TC_BROAD_CPB 4
#endif
template <bool FILL>
__global__ void __launch_bounds__(TPB) k_broad(SceneView v, Work w)
{
  if (passOverflowed(w))
    return;
  const long long n_units = w.counters->n_units;
  const int a = blockIdx.y;
  const unsigned lane = threadIdx.x & 31;
  const int p_begin = v.partner_ptr[a], p_end = v.partner_ptr[a + 1];
  uint32_t* mask = w.bp_mask + static_cast<size_t>(v.partner_word_ptr[a]) * static_cast<size_t>(w.n_units);
  if (FILL)
  {
    // replay the survivors the counting pass found: no AABB is read again
    for (int c = 0; c < TC_BROAD_CPB; ++c)
    {
      const long long u0 = (static_cast<long long>(blockIdx.x) * TC_BROAD_CPB + c) * blockDim.x;
      if (u0 >= n_units)
        break;
      const long long u = u0 + threadIdx.x;
      const bool inb = u < n_units;
      long long s0 = 0;
      if (inb)
      {
        const int k = w.unit_term[u];
        s0 = w.state_base[k] + (u - w.unit_base[k]);
      }
      uint32_t bits = 0;
      for (int pi = p_begin; pi < p_end; ++pi)
      {
        const int j = (pi - p_begin) & 31;
        if (j == 0)
          bits = inb ? mask[static_cast<size_t>((pi - p_begin) >> 5) * static_cast<size_t>(w.n_units) + u] : 0u;
        const bool overlap = (bits >> j) & 1u;
        const unsigned m = __ballot_sync(0xffffffffu, overlap);
        if (m == 0)
          continue;
        const int sp = v.partner_sp[pi];
        const int leader = __ffs(m) - 1;
        int base = 0;
        if (lane == leader)
          base = atomicAdd(&w.sp_cursor[sp], __popc(m));
        base = __shfl_sync(0xffffffffu, base, leader);
        if (overlap)
        {
          const int slot = w.sp_offset[sp] + base + __popc(m & ((1u << lane) - 1u));
          w.items[slot] = static_cast<int32_t>(u);
          w.item_state[slot] = static_cast<int32_t>(s0);
        }
      }
    }
    return;
  }
  // the partner list is the same for the whole block: its records (threshold, partner box or partner shape) are
  // gathered once per tile into shared memory instead of being chased through four tables by every thread
  struct Partner
  {
    double thr, lo[3], hi[3];
    int sp, bidx;
  };
  constexpr int TILE = 32;
  __shared__ Partner tile[TILE];
  const bool single_tile = p_end - p_begin <= TILE;  // then the tile staged for the first chunk serves them all
  for (int c = 0; c < TC_BROAD_CPB; ++c)
  {
    const long long u0 = (static_cast<long long>(blockIdx.x) * TC_BROAD_CPB + c) * blockDim.x;
    if (u0 >= n_units)
      break;
    const long long u = u0 + threadIdx.x;
    const bool inb = u < n_units;
    long long s0 = 0, s1 = 0;
    if (inb)
    {
      const int k = w.unit_term[u];
      s0 = w.state_base[k] + (u - w.unit_base[k]);
      s1 = w.cast ? s0 + 1 : s0;
    }
    double alo[3], ahi[3];
    loadBox(w, a, s0, s1, alo, ahi);
    uint32_t bits = 0;
    // box of the warp's 32 units (mostly consecutive sub-segments of one trajectory): a static partner that misses it
    // misses every lane, and is dismissed with one warp-uniform test instead of 32 tests, a ballot and a branch
    double wlo[3], whi[3];
#pragma unroll
    for (int r = 0; r < 3; ++r)
    {
      wlo[r] = inb ? alo[r] : 1e300;
      whi[r] = inb ? ahi[r] : -1e300;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1)
      {
        wlo[r] = fmin(wlo[r], __shfl_xor_sync(0xffffffffu, wlo[r], o));
        whi[r] = fmax(whi[r], __shfl_xor_sync(0xffffffffu, whi[r], o));
      }
    }
    for (int t0 = p_begin; t0 < p_end; t0 += TILE)
    {
      const int tn = min(TILE, p_end - t0);
      if (!(single_tile && c > 0))
      {
        __syncthreads();
        if (static_cast<int>(threadIdx.x) < tn)
        {
          Partner pr;
          pr.sp = v.partner_sp[t0 + threadIdx.x];
          const ShapePairD spd = v.spairs[pr.sp];
          pr.thr = v.pairs[spd.pair].thr;
          pr.bidx = v.shapes[spd.sb].active_index;
#pragma unroll
          for (int r = 0; r < 3; ++r)
          {
            pr.lo[r] = pr.bidx < 0 ? v.static_aabb[6 * spd.sb + r] : 0.0;
            pr.hi[r] = pr.bidx < 0 ? v.static_aabb[6 * spd.sb + 3 + r] : 0.0;
          }
          tile[threadIdx.x] = pr;
        }
        __syncthreads();
      }
      for (int j = 0; j < tn; ++j)
      {
        const Partner& pr = tile[j];
        double blo[3], bhi[3];
        if (pr.bidx < 0)
        {
          bool miss = false;
#pragma unroll
          for (int r = 0; r < 3; ++r)
            if (wlo[r] - pr.hi[r] > pr.thr || pr.lo[r] - whi[r] > pr.thr)
              miss = true;
          if (miss)
            continue;
#pragma unroll
          for (int r = 0; r < 3; ++r)
            blo[r] = pr.lo[r], bhi[r] = pr.hi[r];
        }
        else
          loadBox(w, pr.bidx, s0, s1, blo, bhi);
        bool overlap = inb;
#pragma unroll
        for (int r = 0; r < 3; ++r)
          if (alo[r] - bhi[r] > pr.thr || blo[r] - ahi[r] > pr.thr)
            overlap = false;
        bits |= (overlap ? 1u : 0u) << j;
        const unsigned m = __ballot_sync(0xffffffffu, overlap);
        if (m && lane == static_cast<unsigned>(__ffs(m) - 1))
          atomicAdd(&w.sp_count[pr.sp], __popc(m));
      }
      if (inb)
        mask[static_cast<size_t>((t0 - p_begin) >> 5) * static_cast<size_t>(w.n_units) + u] = bits;
      bits = 0;
    }
  }
}

// bucket offsets, every bucket padded to a whole narrow-phase CTA
__global__ void __launch_bounds__(1024) k_scan_sp(SceneView v, Work w)
{
  __shared__ long long sm[33];
  if (passOverflowed(w))
    return;
  long long carry = 0, raw_total = 0;
  const int n = v.n_spairs;
  for (int base = 0; base < n; base += blockDim.x)
  {
    const int k = base + threadIdx.x;
    const long long c = (k < n) ? w.sp_count[k] : 0;
    const long long padded = (c + CHUNK - 1) / CHUNK * CHUNK;
    long long t, t2;
    const long long e = blockExclusive(padded, t, sm);
    blockExclusive(c, t2, sm);
    if (k < n)
      w.sp_offset[k] = static_cast<int32_t>(carry + e);
    carry += t;
    raw_total += t2;
  }
  if (threadIdx.x == 0)
  {
    w.sp_offset[n] = static_cast<int32_t>(carry);
    w.counters->n_items_padded = carry;
    w.counters->n_items = raw_total;
    if (carry > 2000000000ll)
      w.counters->overflow = OVF_INDEX_ITEMS;
    else if (w.speculative && carry > w.n_items_padded && w.counters->overflow == 0)
      w.counters->overflow = OVF_SPECULATION;
  }
}

#ifndef TC_FINALIZE_MINB
#define TC_FINALIZE_MINB 6  // measured on C4: 3: 5.47, 4: 5.40, 5: 5.27, 6: 5.21, 7: 5.22, 8: 5.50 ms for the tail (profiles/r1_minb_sweep.md)
#endif
// ---------------------------------------------------------------------------------------------
// contact finalisation: ContactResult fields, cast bookkeeping, LVS merge, filters, sort key
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ UnitInfo unitInfo(const Work& w, int u)
{
  UnitInfo ui;
  const int k = w.unit_term[u];
  const int ti = k % w.ep.terms_per_traj;
  ui.term = k;
  ui.traj = k / w.ep.terms_per_traj;
  ui.t = w.term_t[ti];
  ui.substep = u - w.unit_base[k];
  ui.last = w.term_cnt[k] - 1;
  ui.state0 = w.state_base[k] + ui.substep;
  ui.interp = w.term_interp[k] != 0;
  ui.f0 = w.term_fixed[2 * ti] != 0;
  ui.f1 = w.term_fixed[2 * ti + 1] != 0;
  return ui;
}

// phase 0: every raw hit.  Speculative pass: phase 1 = the hits that exist before the EPA-overflow re-runs start
// (counters->n_raw0), finalised while those re-runs fetch their cold code on another stream; phase 2 = the hits they
// added, and the surplus entries of the key sort
__global__ void __launch_bounds__(TPB, TC_FINALIZE_MINB) k_finalize(SceneView v, Work w, int phase)
{
  if (passOverflowed(w))
    return;
  const long long first = phase == 2 ? w.counters->n_raw0 : 0;
  const long long i = first + static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  const long long n_raw = phase == 1 ? w.counters->n_raw0 : w.counters->n_raw;
  if (phase == 1 && i >= n_raw)
    return;  // no surplus entries in this phase: warp-uniform apart from the last warp, whose ballot below it skips
  if (n_raw > w.n_sort)
  {
    // more raw hits than the key sort was sized for (speculative pass): nothing may be written past the arenas
    if (i == 0)
      w.counters->overflow = OVF_SPECULATION;
    return;
  }
  bool valid = false;
  if (i >= n_raw && i < w.n_sort)
  {
    w.keys[i] = ~0ull;  // surplus entries of the sort: sorted behind every contact
    w.idx[i] = static_cast<int32_t>(i);
  }
  if (i < n_raw)
  {
    const RawContact raw = w.raw[i];
    const UnitInfo ui = unitInfo(w, raw.unit);
    const ShapePairD sp = v.spairs[raw.sp];
    const PairD pr = v.pairs[sp.pair];
    const double* p0[2];
    const double* p1[2];
    const int links[2] = { pr.a, pr.b };
#pragma unroll
    for (int j = 0; j < 2; ++j)
    {
      const bool active = links[j] < v.n_active;
      p0[j] = active ? posePtr(w, links[j], ui.state0) : v.link_pose + 12 * links[j];
      p1[j] = (active && w.cast) ? p0[j] + 12 : p0[j];
    }
    valid = finalizeContact(v, w.ep, raw, ui, p0, p1, w.contacts_tmp[i]);
    unsigned long long key = ~0ull;
    if (w.contact_type != TC_CONTACT_ALL)
    {
      // CLOSEST / FIRST: the manager chooses among ALL its hits before the evaluator filters them, so every raw hit is
      // sorted first (k_type_select then keeps the chosen ones that also pass the filters).  FIRST groups by tested
      // unit, so its first key is unit-major: term | substep | pair | shape pair
      w.raw_valid[i] = valid ? 1 : 0;
      if (w.contact_type == TC_CONTACT_FIRST)
        key = (static_cast<unsigned long long>(ui.term) << 36) | (static_cast<unsigned long long>(ui.substep) << 22) |
              (static_cast<unsigned long long>(sp.pair) << 8) | static_cast<unsigned long long>(sp.k);
      else
        key = (static_cast<unsigned long long>(ui.term) << 36) | (static_cast<unsigned long long>(sp.pair) << 22) |
              (static_cast<unsigned long long>(ui.substep) << 8) | static_cast<unsigned long long>(sp.k);
      valid = false;  // counted by k_type_select
    }
    else if (valid)
      key = (static_cast<unsigned long long>(ui.term) << 36) | (static_cast<unsigned long long>(sp.pair) << 22) |
            (static_cast<unsigned long long>(ui.substep) << 8) | static_cast<unsigned long long>(sp.k);
    w.keys[i] = key;
    w.idx[i] = static_cast<int32_t>(i);
  }
  const unsigned m = __ballot_sync(__activemask(), valid);
  if (valid && (threadIdx.x & 31) == static_cast<unsigned>(__ffs(m) - 1))
    atomicAdd(reinterpret_cast<unsigned long long*>(&w.counters->n_valid), static_cast<unsigned long long>(__popc(m)));
}

__global__ void k_snapshot_raw(Work w)
{
  if (threadIdx.x == 0)
    w.counters->n_raw0 = w.counters->n_raw;
}

// ContactTestType CLOSEST / FIRST (tesseract_collision: the manager keeps one result per link pair / stops at its first
// result; the evaluators pass the type through, trajopt/src/problem_description.cpp:1669-1673).  Input: every raw hit
// sorted by the first key (keys_alt / idx_alt).  Output: keys / idx rewritten for the final sort — the standard key for
// the chosen hit of every group if it also passed the evaluator's filters, all-ones otherwise.
__global__ void __launch_bounds__(TPB) k_type_select(Work w, const uint8_t* valid)
{
  if (passOverflowed(w))
    return;
  const long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  const long long n_raw = w.counters->n_raw;
  bool keep = false;
  if (i >= n_raw && i < w.n_sort)
  {
    w.keys[i] = ~0ull;
    w.idx[i] = w.idx_alt[i];
  }
  if (i < n_raw)
  {
    const unsigned long long key = w.keys_alt[i];
    unsigned long long out = ~0ull;
    if (w.contact_type == TC_CONTACT_FIRST)
    {
      const bool leader = i == 0 || (w.keys_alt[i - 1] >> 22) != (key >> 22);
      keep = leader && valid[w.idx_alt[i]];
      // back to the standard layout term | pair | substep | shape pair
      out = (key & ~((1ull << 36) - 1)) | (((key >> 8) & 0x3fffull) << 22) | (((key >> 22) & 0x3fffull) << 8) | (key & 0xffull);
    }
    else
    {
      // closest of the link pair in this unit: the first result, in shape-pair order, within 1e-9 m of the smallest
      // distance (sub-shapes of a compound link share faces: exact ties that the last bit would otherwise decide)
      const double CLOSEST_TIE = 1e-9;
      const unsigned long long g = key >> 8;
      const double mine = w.contacts_tmp[w.idx_alt[i]].distance;
      double dmin = mine, dmin_before = 1e300;
      for (long long j = i - 1; j >= 0 && (w.keys_alt[j] >> 8) == g; --j)
        dmin_before = fmin(dmin_before, w.contacts_tmp[w.idx_alt[j]].distance);
      dmin = fmin(dmin, dmin_before);
      for (long long j = i + 1; j < n_raw && (w.keys_alt[j] >> 8) == g; ++j)
        dmin = fmin(dmin, w.contacts_tmp[w.idx_alt[j]].distance);
      // chosen = inside the band and nothing before it is
      const bool best = mine <= dmin + CLOSEST_TIE && !(dmin_before <= dmin + CLOSEST_TIE);
      keep = best && valid[w.idx_alt[i]];
      out = key;
    }
    w.keys[i] = keep ? out : ~0ull;
    w.idx[i] = w.idx_alt[i];
  }
  const unsigned m = __ballot_sync(0xffffffffu, keep);
  if ((threadIdx.x & 31) == 0 && m)
    atomicAdd(reinterpret_cast<unsigned long long*>(&w.counters->n_valid), static_cast<unsigned long long>(__popc(m)));
}

// transform[i] / cc_transform[i] of a finalised contact are not stored: they are records of the pose arena
// (ContactResult::transform, cc_transform; trajopt/src/collision_terms.cpp:215-223)
__device__ __forceinline__ void contactTransforms(const SceneView& v, const Work& w, const tc_contact& c, double* xf)
{
  const bool lvs_discrete = w.ep.evaluator == TC_EVAL_LVS_DISCRETE;
#pragma unroll
  for (int i = 0; i < 2; ++i)
  {
    const int link = c.link_ids[i];
    const bool active = link < v.n_active;
    const double* t0 = active ? posePtr(w, link, c.state) : v.link_pose + 12 * link;
    const double* t1 = (active && w.cast) ? t0 + 12 : (lvs_discrete ? t0 : nullptr);
    for (int k = 0; k < 12; ++k)
    {
      xf[24 * i + k] = t0[k];
      xf[24 * i + 12 + k] = t1 ? t1[k] : ((k == 0 || k == 5 || k == 10) ? 1.0 : 0.0);
    }
  }
}

__global__ void __launch_bounds__(TPB) k_transforms(SceneView v, Work w, double* xf)
{
  const long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i < w.counters->n_valid)
    contactTransforms(v, w, w.contacts[i], xf + i * 48);
}

// sorted order -> final array, one thread per 8-byte word of tc_contact
__global__ void __launch_bounds__(256) k_gather(Work w, const int32_t* idx)
{
  if (passOverflowed(w))
    return;
  constexpr int WC = sizeof(tc_contact) / 8;
  const long long g = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  const long long i = g / WC;
  const int word = static_cast<int>(g % WC);
  if (i >= w.counters->n_valid)
    return;
  const long long src = idx[i];
  reinterpret_cast<double*>(w.contacts + i)[word] = reinterpret_cast<const double*>(w.contacts_tmp + src)[word];
}

__global__ void k_term_ptr(Work w, const unsigned long long* keys)
{
  if (passOverflowed(w))
    return;
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k > w.ep.n_terms)
    return;
  const unsigned long long target = static_cast<unsigned long long>(k) << 36;
  long long lo = 0, hi = w.counters->n_valid;  // first index with key >= target
  while (lo < hi)
  {
    const long long mid = (lo + hi) >> 1;
    if (keys[mid] < target)
      lo = mid + 1;
    else
      hi = mid;
  }
  w.term_ptr[k] = lo;
}

// ---------------------------------------------------------------------------------------------
// K4: gradients, rows, cost, sets
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ int termOfContact(const Work& w, long long i)
{
  // contacts are sorted by term: binary search the term pointer
  int lo = 0, hi = w.ep.n_terms;
  while (hi - lo > 1)
  {
    const int mid = (lo + hi) >> 1;
    if (w.term_ptr[mid] <= i)
      lo = mid;
    else
      hi = mid;
  }
  return lo;
}

template <int NMAX>
__global__ void __launch_bounds__(TPB) k_grad(SceneView v, Work w)
{
  if (passOverflowed(w))
    return;
  const long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= w.counters->n_valid)
    return;
  const tc_contact c = w.contacts[i];
  const int ti = termOfContact(w, i) % w.ep.terms_per_traj;
  const int n = v.n_dof;
  const double* q0 = w.q + (static_cast<size_t>(c.traj) * w.ep.T + c.term) * n;
  const double* q1 = (w.ep.evaluator == TC_EVAL_DISCRETE) ? q0 : q0 + n;
  double qt[NMAX], xf[48];
  contactTransforms(v, w, c, xf);
  GradOut out;
  contactGradient(v, w.ep, c, xf, q0, q1, w.term_fixed[2 * ti] != 0, w.term_fixed[2 * ti + 1] != 0, qt,
                  w.grad + static_cast<size_t>(i) * 4 * n, out);
  w.grec[i] = out.rec;
  w.ghas[i] = out.has;
}

// per-term row offsets: rows of a term all have n_dof * (#free blocks) entries
__global__ void __launch_bounds__(1024) k_row_base(SceneView v, Work w, long long* term_row_base)
{
  if (passOverflowed(w))
    return;
  __shared__ long long sm[33];
  long long carry = 0;
  const int n = w.ep.n_terms;
  const bool single = w.ep.evaluator == TC_EVAL_DISCRETE;
  for (int base = 0; base < n; base += blockDim.x)
  {
    const int k = base + threadIdx.x;
    long long val = 0;
    if (k < n)
    {
      const int ti = k % w.ep.terms_per_traj;
      const int nb = (w.term_fixed[2 * ti] ? 0 : 1) + ((single || w.term_fixed[2 * ti + 1]) ? 0 : 1);
      val = (w.term_ptr[k + 1] - w.term_ptr[k]) * nb * v.n_dof;
    }
    long long t;
    const long long e = blockExclusive(val, t, sm);
    if (k < n)
      term_row_base[k] = carry + e;
    carry += t;
  }
  if (threadIdx.x == 0)
  {
    term_row_base[n] = carry;
    w.counters->nnz = carry;
  }
}

// Packed rows (the compact transport form of tc_pipeline_eval_rows_packed).  A row's Jacobian block has a non-zero in
// column j only if joint j moves one of the contact's two links: the reference writes all n_dof entries of a block,
// zeros included (discrete_collision_constraint.cpp:266-277), a third of them on C4.  In packed form a block holds only
// the joints of dofMask(link a) | dofMask(link b), ascending; the host re-expands with tc_scene_pair_dof_mask.
__device__ __forceinline__ unsigned long long pairDofMask(const SceneView& v, int a, int b)
{
  return (a < v.n_active ? v.link_dof_mask[a] : 0ull) | (b < v.n_active ? v.link_dof_mask[b] : 0ull);
}

// widths of the packed rows, straight from the sorted keys (term | link pair | ...): no contact record is touched
__global__ void __launch_bounds__(256) k_row_width(SceneView v, Work w, const unsigned long long* keys)
{
  const long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i > w.n_contacts)
    return;
  long long width = 0;
  if (!passOverflowed(w) && i < w.counters->n_valid)
  {
    const unsigned long long key = keys[i];
    const int ti = static_cast<int>(key >> 36) % w.ep.terms_per_traj;
    const PairD pr = v.pairs[(key >> 22) & 0x3fffull];
    const bool single = w.ep.evaluator == TC_EVAL_DISCRETE;
    const int nb = (w.term_fixed[2 * ti] ? 0 : 1) + ((single || w.term_fixed[2 * ti + 1]) ? 0 : 1);
    width = static_cast<long long>(__popcll(pairDofMask(v, pr.a, pr.b))) * nb;
  }
  w.row_width[i] = width;
}

// DiscreteCollisionConstraintD / ContinuousCollisionConstraintD getValues + getJacobian
// (discrete_collision_constraint.cpp:214-283, continuous_collision_constraint.cpp:311-417) and the legacy
// CollisionsToDistanceExpressions / CalcDistExpressions* (collision_terms.cpp:341-386,463-554)
__device__ __forceinline__ void writeRow(const SceneView& v, const Work& w, long long i, const tc_contact& c,
                                         const double* g, const tc_gradient& rec, int has,
                                         const long long* term_row_base, int32_t* col_idx, double* val,
                                         double* row_value, double* row_coeff, double* row_margin)
{
  const int k = termOfContact(w, i);
  const int ti = k % w.ep.terms_per_traj;
  const bool f0 = w.term_fixed[2 * ti] != 0, f1 = w.term_fixed[2 * ti + 1] != 0;
  const bool single = w.ep.evaluator == TC_EVAL_DISCRETE;
  const int n = v.n_dof;
  const int nb = (f0 ? 0 : 1) + ((single || f1) ? 0 : 1);
  const unsigned long long keep = w.rows_packed ? pairDofMask(v, c.link_ids[0], c.link_ids[1]) : ~0ull;
  long long nnz = w.rows_packed ? w.row_ptr[i] : term_row_base[k] + (i - w.term_ptr[k]) * nb * n;
  if (!w.rows_packed)
    w.row_ptr[i] = nnz;
  w.row_block[2 * i] = f0 ? -1 : c.term * n;
  w.row_block[2 * i + 1] = (single || f1) ? -1 : (c.term + 1) * n;
  w.row_pair[i] = c.pair;
  const PairD pr = v.pairs[c.pair];
  const double* q0 = w.q + (static_cast<size_t>(c.traj) * w.ep.T + c.term) * n;
  double constant = c.distance;
  for (int blk = 0; blk < 2; ++blk)
  {
    if (blk == 0 && f0)
      continue;
    if (blk == 1 && (single || f1))
      continue;
    const double* qb = blk ? q0 + n : q0;
    for (int j = 0; j < n; ++j)
    {
      double s = 0;
#pragma unroll
      for (int l = 0; l < 2; ++l)
        if (has & (1 << (2 * l + blk)))
          s = fma(blk ? rec.scale1[l] : rec.scale0[l], g[(2 * l + blk) * n + j], s);
      if (w.ep.stack == TC_STACK_LEGACY)
        constant = fma(-s, qb[j], constant);
      if (j < 64 && !((keep >> j) & 1ull))
        continue;  // packed rows: a joint that moves neither link (s is an exact zero)
      val[nnz] = (w.ep.stack == TC_STACK_LEGACY) ? s : -1.0 * s;
      col_idx[nnz] = (c.term + blk) * n + j;
      ++nnz;
    }
  }
  row_value[i] = (w.ep.stack == TC_STACK_LEGACY) ? constant : (pr.margin - c.distance);
  row_coeff[i] = pr.coeff;
  row_margin[i] = pr.margin;
}

__global__ void __launch_bounds__(TPB) k_rows(SceneView v, Work w, const long long* term_row_base, int32_t* col_idx,
                                              double* val, double* row_value, double* row_coeff, double* row_margin)
{
  if (passOverflowed(w))
    return;
  const long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i > w.counters->n_valid)
    return;
  if (i == w.counters->n_valid)
  {
    w.row_ptr[i] = term_row_base[w.ep.n_terms];
    return;
  }
  writeRow(v, w, i, w.contacts[i], w.grad + static_cast<size_t>(i) * 4 * v.n_dof, w.grec[i], w.ghas[i], term_row_base,
           col_idx, val, row_value, row_coeff, row_margin);
}

// gradients folded straight into the rows (TC_WANT_ROWS without TC_WANT_GRADIENTS): same arithmetic as k_grad
// followed by k_rows, the per-link gradient vectors live in thread-local memory only
#ifndef TC_GRADROWS_MINB
#define TC_GRADROWS_MINB 6
#endif
// GATHER: the kernel also is the gather of the sorted order — it takes its contact from the unsorted arena through the
// sorted index and writes the final record itself, so k_gather's separate pass over 2 x 200 bytes per contact is saved
template <int NMAX, bool GATHER>
__global__ void __launch_bounds__(TPB, TC_GRADROWS_MINB) k_grad_rows(SceneView v, Work w, const long long* term_row_base, int32_t* col_idx,
                                                   double* val, double* row_value, double* row_coeff,
                                                   double* row_margin)
{
  if (passOverflowed(w))
    return;
  const long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i > w.counters->n_valid)
    return;
  if (i == w.counters->n_valid)
  {
    if (w.rows_packed)
      w.counters->nnz = w.row_ptr[i];
    else
      w.row_ptr[i] = term_row_base[w.ep.n_terms];
    return;
  }
  const tc_contact c = GATHER ? w.contacts_tmp[w.idx_alt[i]] : w.contacts[i];
  if (GATHER)
    w.contacts[i] = c;
  const int ti = termOfContact(w, i) % w.ep.terms_per_traj;
  const int n = v.n_dof;
  const double* q0 = w.q + (static_cast<size_t>(c.traj) * w.ep.T + c.term) * n;
  const double* q1 = (w.ep.evaluator == TC_EVAL_DISCRETE) ? q0 : q0 + n;
  double qt[NMAX], xf[48], g[4 * NMAX];
  contactTransforms(v, w, c, xf);
  GradOut out;
  contactGradient(v, w.ep, c, xf, q0, q1, w.term_fixed[2 * ti] != 0, w.term_fixed[2 * ti + 1] != 0, qt, g, out);
  writeRow(v, w, i, c, g, out.rec, out.has, term_row_base, col_idx, val, row_value, row_coeff, row_margin);
}

// The same rows for chains of up to NR joints with nothing in thread-local memory (the general kernel above keeps the
// contact, its transforms and four gradient vectors on the stack: 1.1 kB per thread that the ncu capture of round 2
// shows written back to DRAM, 5 GB per C4 pass).  The contact is read field by field where it lies, the joint values
// come through an accessor, and the per-joint sums of writeRow live in two statically indexed register arrays.  The
// arithmetic is contactGradient's and writeRow's, operation for operation (the on-demand path must give the same bits:
// tests/test_gpu_parity.py::test_rows_folded_inside_the_pass_equal_rows_on_demand).
template <int NR>
struct RowSums
{
  double s0[NR], s1[NR];
  __device__ __forceinline__ void add(int j, bool use0, double sc0, double g0, bool use1, double sc1, double g1)
  {
#pragma unroll
    for (int jj = 0; jj < NR; ++jj)
      if (jj == j)
      {
        if (use0)
          s0[jj] = fma(sc0, g0, s0[jj]);
        if (use1)
          s1[jj] = fma(sc1, g1, s1[jj]);
      }
  }
};

__device__ __forceinline__ Vec3<double> rotBy(const double* p12, Vec3<double> x)
{
  return p12 ? Pose<double>::from12(p12).rot(x) : Pose<double>().rot(x);
}

template <int NR, bool GATHER>
__global__ void __launch_bounds__(TPB, TC_GRADROWS_MINB) k_grad_rows_regs(SceneView v, Work w, const long long* term_row_base,
                                                                         int32_t* col_idx, double* val, double* row_value,
                                                                         double* row_coeff, double* row_margin)
{
  if (passOverflowed(w))
    return;
  const long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i > w.counters->n_valid)
    return;
  if (i == w.counters->n_valid)
  {
    if (w.rows_packed)
      w.counters->nnz = w.row_ptr[i];
    else
      w.row_ptr[i] = term_row_base[w.ep.n_terms];
    return;
  }
  const tc_contact* cp = GATHER ? w.contacts_tmp + w.idx_alt[i] : w.contacts + i;
  if (GATHER)
  {
    constexpr int WC = sizeof(tc_contact) / 8;
    const double* src = reinterpret_cast<const double*>(cp);
    double* dst = reinterpret_cast<double*>(w.contacts + i);
#pragma unroll
    for (int k = 0; k < WC; ++k)
      dst[k] = src[k];
  }
  const int k = termOfContact(w, i);
  const int ti = k % w.ep.terms_per_traj;
  const bool f0 = w.term_fixed[2 * ti] != 0, f1 = w.term_fixed[2 * ti + 1] != 0;
  const bool single = w.ep.evaluator == TC_EVAL_DISCRETE;
  const bool legacy = w.ep.stack == TC_STACK_LEGACY;
  const bool lvs_discrete = w.ep.evaluator == TC_EVAL_LVS_DISCRETE;
  const int n = v.n_dof;
  const int term = cp->term, state = cp->state, pair = cp->pair;
  const double distance = cp->distance;
  const double* q0 = w.q + (static_cast<size_t>(cp->traj) * w.ep.T + term) * n;
  const double* q1 = single ? q0 : q0 + n;
  RowSums<NR> sum;
#pragma unroll
  for (int jj = 0; jj < NR; ++jj)
    sum.s0[jj] = sum.s1[jj] = 0.0;
  const Vec3<double> nrm(cp->normal[0], cp->normal[1], cp->normal[2]);
#pragma unroll 1
  for (int l = 0; l < 2; ++l)
  {
    const int link = cp->link_ids[l];
    if (link >= v.n_active)
      continue;
    const int cct = cp->cc_type[l];
    const double cc_time = cp->cc_time[l];
    const bool cc = cct != TC_CC_NONE;
    const Vec3<double> pl(cp->nearest_points_local[l][0], cp->nearest_points_local[l][1], cp->nearest_points_local[l][2]);
    const Vec3<double> tv = (l == 0) ? -nrm : nrm;
    const double* t0 = posePtr(w, link, state);                                        // transform[l]
    const double* t1 = w.cast ? t0 + 12 : (lvs_discrete ? t0 : nullptr);               // cc_transform[l]
    const double sc0 = cc ? (1 - cc_time) : 1.0, sc1 = cc ? cc_time : 1.0;
    if (legacy || single)
    {
      // one GetGradient per free timestep block, each at that block's own joint values (linkGradient)
      if (single || !f0)
      {
        const Vec3<double> ref = rotBy(t0, pl);
        pointGradientWalk(v, link, QArray{ q0 }, ref, ref, tv,
                          [&](int j, double g, double) { sum.add(j, true, sc0, g, false, 0.0, 0.0); });
      }
      if (legacy && !single && !f1)
      {
        const Vec3<double> ref = rotBy(cc ? t1 : t0, pl);
        pointGradientWalk(v, link, QArray{ q1 }, ref, ref, tv,
                          [&](int j, double g, double) { sum.add(j, false, 0.0, 0.0, true, sc1, g); });
      }
    }
    else
    {
      // both blocks share the joint values: one chain walk, two reference points
      const Vec3<double> ref0 = rotBy(t0, pl), ref1 = rotBy(cc ? t1 : t0, pl);
      pointGradientWalk(v, link, QByKind{ q0, q1, cc_time, cct }, ref0, ref1, tv,
                        [&](int j, double g0, double g1) { sum.add(j, true, sc0, g0, true, sc1, g1); });
    }
  }
  // writeRow
  const bool b0 = !f0, b1 = !(single || f1);
  const int nb = (b0 ? 1 : 0) + (b1 ? 1 : 0);
  const bool packed = w.rows_packed != 0;
  const unsigned keep = packed ? static_cast<unsigned>(pairDofMask(v, cp->link_ids[0], cp->link_ids[1])) : ~0u;
  long long nnz = packed ? w.row_ptr[i] : term_row_base[k] + (i - w.term_ptr[k]) * nb * n;
  if (!packed)
    w.row_ptr[i] = nnz;
  w.row_block[2 * i] = b0 ? term * n : -1;
  w.row_block[2 * i + 1] = b1 ? (term + 1) * n : -1;
  w.row_pair[i] = pair;
  const PairD pr = v.pairs[pair];
  double constant = distance;
  if (b0)
  {
#pragma unroll
    for (int jj = 0; jj < NR; ++jj)
      if (jj < n)
      {
        const double s = sum.s0[jj];
        if (legacy)
          constant = fma(-s, q0[jj], constant);
        if ((keep >> jj) & 1u)
        {
          // dense rows: slot jj; packed rows: the joints kept below jj come first (col_idx is not part of that form)
          const int at = packed ? __popc(keep & ((1u << jj) - 1u)) : jj;
          val[nnz + at] = legacy ? s : -1.0 * s;
          if (!packed)
            col_idx[nnz + at] = term * n + jj;
        }
      }
    nnz += packed ? __popc(keep) : n;
  }
  if (b1)
  {
#pragma unroll
    for (int jj = 0; jj < NR; ++jj)
      if (jj < n)
      {
        const double s = sum.s1[jj];
        if (legacy)
          constant = fma(-s, q0[n + jj], constant);
        if ((keep >> jj) & 1u)
        {
          const int at = packed ? __popc(keep & ((1u << jj) - 1u)) : jj;
          val[nnz + at] = legacy ? s : -1.0 * s;
          if (!packed)
            col_idx[nnz + at] = (term + 1) * n + jj;
        }
      }
  }
  row_value[i] = legacy ? constant : (pr.margin - distance);
  row_coeff[i] = pr.coeff;
  row_margin[i] = pr.margin;
}

// ---------------------------------------------------------------------------------------------
// QP constraint block of the collision hinge costs in compressed-sparse-COLUMN form, assembled on the device.
//   CollisionCost::convex + ConvexObjective::addHinge: per expression a slack s >= 0, the row viol - s <= 0 and the
//   objective term coeff * s                 (trajopt/src/collision_terms.cpp:1267-1285, trajopt_sco/src/modeling.cpp:19-27)
//   OSQPModel::updateConstraints / exprToEigen turn those rows into the CSC matrix OSQP takes
//                                            (trajopt_sco/src/osqp_interface.cpp:213-280, solver_utils.cpp:111-144)
// Rows = the contact rows in their global order.  Columns = [ B*T*n_dof trajectory variables | one slack per row ].
// A row touches the n_dof variables of its segment's two timesteps, and the contacts are sorted by (trajectory, term),
// so the entries of a variable column are two contiguous runs of rows (the T1 blocks of the segment ending at that
// timestep, then the T0 blocks of the segment starting there): column pointers follow from the term pointer with one
// scan, and every row writes its entries to computed places — no atomics, no sort.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ void termBlocks(const Work& w, int ti, bool& has0, bool& has1)
{
  const bool single = w.ep.evaluator == TC_EVAL_DISCRETE;
  has0 = w.term_fixed[2 * ti] == 0;
  has1 = !(single || w.term_fixed[2 * ti + 1] != 0);
}

// rows of trajectory b that put a T1 block (rows1) / a T0 block (rows0) into the columns of timestep t
__device__ __forceinline__ void columnRuns(const Work& w, int b, int t, long long& rows1, long long& rows0)
{
  rows1 = rows0 = 0;
  const int tpt = w.ep.terms_per_traj;
  for (int ti = 0; ti < tpt; ++ti)
  {
    const int tt = w.term_t[ti];
    if (tt != t && tt != t - 1)
      continue;
    bool has0, has1;
    termBlocks(w, ti, has0, has1);
    const long long k = static_cast<long long>(b) * tpt + ti;
    const long long nr = w.term_ptr[k + 1] - w.term_ptr[k];
    if (tt == t && has0)
      rows0 = nr;
    if (tt == t - 1 && has1)
      rows1 = nr;
  }
}

// bt_base[b*T + t] = first entry of the n_dof columns of (b, t); bt_base[B*T] = entries of the variable columns
__global__ void __launch_bounds__(1024) k_csc_base(SceneView v, Work w, long long* bt_base)
{
  __shared__ long long sm[33];
  long long carry = 0;
  const int n = w.ep.B * w.ep.T;
  for (int base = 0; base < n; base += blockDim.x)
  {
    const int e = base + threadIdx.x;
    long long val = 0;
    if (e < n)
    {
      long long r1, r0;
      columnRuns(w, e / w.ep.T, e % w.ep.T, r1, r0);
      val = (r1 + r0) * v.n_dof;
    }
    long long t;
    const long long ex = blockExclusive(val, t, sm);
    if (e < n)
      bt_base[e] = carry + ex;
    carry += t;
  }
  if (threadIdx.x == 0)
    bt_base[n] = carry;
}

__global__ void __launch_bounds__(256) k_csc_colptr(SceneView v, Work w, const long long* bt_base, long long* col_ptr)
{
  const long long g = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  const long long n_bt = static_cast<long long>(w.ep.B) * w.ep.T;
  const long long n_x = n_bt * v.n_dof;
  const long long nnz_x = bt_base[n_bt];
  if (g < n_x)
  {
    const long long e = g / v.n_dof;
    const int j = static_cast<int>(g % v.n_dof);
    const long long c = (bt_base[e + 1] - bt_base[e]) / v.n_dof;
    col_ptr[g] = bt_base[e] + j * c;
  }
  else if (g <= n_x + w.counters->n_valid)
    col_ptr[g] = nnz_x + (g - n_x);  // slack columns hold one entry each; the last value closes the array
}

__global__ void __launch_bounds__(TPB) k_csc_fill(SceneView v, Work w, const long long* bt_base, const double* val,
                                                  const double* row_value, const double* row_coeff,
                                                  const double* row_margin, int32_t* row_idx, double* out,
                                                  double* upper, double* slack_cost)
{
  const long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= w.counters->n_valid)
    return;
  const int k = termOfContact(w, i);
  const int tpt = w.ep.terms_per_traj;
  const int b = k / tpt, ti = k % tpt, t = w.term_t[ti], n = v.n_dof;
  bool has0, has1;
  termBlocks(w, ti, has0, has1);
  const long long local = i - w.term_ptr[k];
  const long long r0 = w.row_ptr[i];
  const bool legacy = w.ep.stack == TC_STACK_LEGACY;
  // legacy rows hold the affine distance d(x) = c + g.x: the hinge row is  -g.x - s <= c - margin.
  // ifopt rows hold the violation margin - d and its Jacobian J at x0:     J.x - s <= J.x0 - (margin - d)
  const double sign = legacy ? -1.0 : 1.0;
  const double* x0 = w.q + (static_cast<size_t>(b) * w.ep.T + t) * n;
  double jx0 = 0;
  int blk = 0;
  if (has0)
  {
    const long long e = static_cast<long long>(b) * w.ep.T + t;
    const long long c = (bt_base[e + 1] - bt_base[e]) / n;
    long long r1, rr0;
    columnRuns(w, b, t, r1, rr0);
    for (int j = 0; j < n; ++j)
    {
      const double a = val[r0 + j];
      const long long pos = bt_base[e] + j * c + r1 + local;
      row_idx[pos] = static_cast<int32_t>(i);
      out[pos] = sign * a;
      jx0 += a * x0[j];
    }
    ++blk;
  }
  if (has1)
  {
    const long long e = static_cast<long long>(b) * w.ep.T + t + 1;
    const long long c = (bt_base[e + 1] - bt_base[e]) / n;
    for (int j = 0; j < n; ++j)
    {
      const double a = val[r0 + blk * n + j];
      const long long pos = bt_base[e] + j * c + local;
      row_idx[pos] = static_cast<int32_t>(i);
      out[pos] = sign * a;
      jx0 += a * x0[n + j];
    }
  }
  const long long nnz_x = bt_base[static_cast<long long>(w.ep.B) * w.ep.T];
  row_idx[nnz_x + i] = static_cast<int32_t>(i);
  out[nnz_x + i] = -1.0;
  upper[i] = legacy ? row_value[i] - row_margin[i] : jx0 - row_value[i];
  slack_cost[i] = row_coeff[i];
}

// CollisionCost::value / CollisionConstraint::value (collision_terms.cpp:1287-1306,1367-1386)
__global__ void __launch_bounds__(TPB) k_cost_term(SceneView v, Work w, double* cost_term)
{
  if (passOverflowed(w))
    return;
  // one warp per term: lanes stride over the term's contacts, fixed-order tree reduction
  const int k = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const unsigned lane = threadIdx.x & 31;
  if (k >= w.ep.n_terms)
    return;
  double out = 0;
  for (long long i = w.term_ptr[k] + lane; i < w.term_ptr[k + 1]; i += 32)
  {
    const PairD& pr = v.pairs[w.contacts[i].pair];
    out += fmax(pr.margin - w.contacts[i].distance, 0.0) * pr.coeff;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1)
    out += __shfl_down_sync(0xffffffffu, out, o);
  if (lane == 0)
    cost_term[k] = out;
}
__global__ void k_cost_traj(Work w, const double* cost_term, double* cost_traj)
{
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= w.ep.B)
    return;
  double out = 0;
  for (int ti = 0; ti < w.ep.terms_per_traj; ++ti)
    out += cost_term[static_cast<size_t>(b) * w.ep.terms_per_traj + ti];
  cost_traj[b] = out;
}

// multi-start winner of this rank, left on the device for the all-gather: one block, arg-min over the
// per-trajectory costs with the lowest index winning ties, then the winner's joint values
__global__ void __launch_bounds__(1024) k_best_record(Work w, const double* cost_traj, long long index_offset, double* rec, int n_dof)
{
  __shared__ double s_cost[1024];
  __shared__ int s_idx[1024];
  double best = 1.7976931348623157e308;
  int bi = 0x7fffffff;
  for (int b = threadIdx.x; b < w.ep.B; b += blockDim.x)
  {
    const double c = cost_traj[b];
    if (c < best || (c == best && b < bi))
      best = c, bi = b;
  }
  s_cost[threadIdx.x] = best, s_idx[threadIdx.x] = bi;
  __syncthreads();
  for (int o = blockDim.x / 2; o > 0; o >>= 1)
  {
    if (static_cast<int>(threadIdx.x) < o)
    {
      const double c = s_cost[threadIdx.x + o];
      const int i = s_idx[threadIdx.x + o];
      if (c < s_cost[threadIdx.x] || (c == s_cost[threadIdx.x] && i < s_idx[threadIdx.x]))
        s_cost[threadIdx.x] = c, s_idx[threadIdx.x] = i;
    }
    __syncthreads();
  }
  const int win = s_idx[0];
  if (threadIdx.x == 0)
  {
    rec[0] = s_cost[0];
    reinterpret_cast<long long*>(rec)[1] = index_offset + win;
  }
  const int nv = w.ep.T * n_dof;
  for (int e = threadIdx.x; e < nv; e += blockDim.x)
    rec[2 + e] = w.q[static_cast<size_t>(win) * nv + e];
}

// GradientResultsSet grouping, ordering and the fixed-size constraint rows
//   discrete_collision_evaluators.cpp:101-148, continuous_collision_evaluators.cpp:113-162,
//   trajopt_common/src/collision_types.cpp:173-333, weighted_average_methods.cpp:31-107,
//   discrete_collision_constraint.cpp:68-141, continuous_collision_constraint.cpp:80-236
__device__ __forceinline__ int cantorKey(int shape_id, int subshape_id)
{
  if (subshape_id < 0)
    return shape_id;
  return static_cast<int>(1 / 2.0 * (shape_id + subshape_id) * (shape_id + subshape_id + 1) + subshape_id);
}
constexpr double LOWEST = -1.7976931348623157e308;

__device__ bool setMaxError(const tc_set& s, int which, bool buffer, double& e)
{
  e = LOWEST;
  bool found = false;
#pragma unroll
  for (int i = 0; i < 2; ++i)
#pragma unroll
    for (int k = 0; k < 2; ++k)
    {
      if (which != 2 && which != k)
        continue;
      const double val = buffer ? s.error_with_buffer[i][k] : s.error[i][k];
      if (s.has_error[i][k] && val > e)
      {
        e = val;
        found = true;
      }
    }
  return found;
}

__device__ void weightedAvgRow(const SceneView& v, const Work& w, const tc_set& s, int i0, int i1, double max_ewb,
                               bool t1, double* row)
{
  const int n = v.n_dof;
  double total = 0;
  long cnt = 0;
  for (int c = i0; c < i1; ++c)
  {
    const tc_contact& ct = w.contacts[c];
    if (cantorKey(ct.shape_id[0], ct.subshape_id[0]) != s.shape_key[0] ||
        cantorKey(ct.shape_id[1], ct.subshape_id[1]) != s.shape_key[1])
      continue;
    const int has = w.ghas[c];
    const tc_gradient& rec = w.grec[c];
    for (int i = 0; i < 2; ++i)
    {
      if (!(has & (1 << (2 * i + (t1 ? 1 : 0)))))
        continue;
      if (ct.cc_type[i] == (t1 ? TC_CC_TIME0 : TC_CC_TIME1))
        continue;
      if (s.error_with_buffer[i][t1 ? 1 : 0] > 0)
      {
        const double wt = fmax(rec.error_with_buffer, 0.0) / max_ewb;
        total += wt;
        const double sc = t1 ? rec.scale1[i] : rec.scale0[i];
        const double* g = w.grad + (static_cast<size_t>(c) * 4 + 2 * i + (t1 ? 1 : 0)) * n;
        for (int j = 0; j < n; ++j)
          row[j] += wt * (sc * g[j]);
        ++cnt;
      }
    }
  }
  if (cnt == 0)
    return;
  const double scale = 1 / total;
  for (int j = 0; j < n; ++j)
    row[j] = -1.0 * (row[j] * scale);
}

__global__ void __launch_bounds__(64) k_sets(SceneView v, Work w, int max_num_cnt, tc_set* sets, int32_t* set_count,
                                             int32_t* set_run, double* values, double* coeffs, double* jac)
{
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= w.ep.n_terms)
    return;
  const int n = v.n_dof;
  const long long lo = w.term_ptr[k], hi = w.term_ptr[k + 1];
  tc_set* S = sets + lo;
  int32_t* RUN = set_run + 2 * lo;
  const int ti = k % w.ep.terms_per_traj;
  const bool f0 = w.term_fixed[2 * ti] != 0, f1 = w.term_fixed[2 * ti + 1] != 0;
  const bool single = w.ep.evaluator == TC_EVAL_DISCRETE;
  int ns = 0;
  long long i = lo;
  while (i < hi)
  {
    const int pair = w.contacts[i].pair;
    const long long i0 = i;
    while (i < hi && w.contacts[i].pair == pair)
      ++i;
    const int first = ns;
    for (long long c = i0; c < i; ++c)
    {
      const tc_contact& ct = w.contacts[c];
      const int k0 = cantorKey(ct.shape_id[0], ct.subshape_id[0]), k1 = cantorKey(ct.shape_id[1], ct.subshape_id[1]);
      int j = first;
      for (; j < ns; ++j)
        if (S[j].shape_key[0] == k0 && S[j].shape_key[1] == k1)
          break;
      if (j == ns)
      {
        tc_set t;
        t.traj = ct.traj, t.term = ct.term, t.pair = pair;
        t.shape_key[0] = k0, t.shape_key[1] = k1;
        t.n_results = 0;
        t.coeff = v.pairs[pair].coeff;
        for (int a = 0; a < 2; ++a)
          for (int b = 0; b < 2; ++b)
            t.has_error[a][b] = 0, t.error[a][b] = LOWEST, t.error_with_buffer[a][b] = LOWEST;
        S[ns] = t;
        RUN[2 * ns] = static_cast<int32_t>(i0);
        RUN[2 * ns + 1] = static_cast<int32_t>(i);
        ++ns;
      }
      // GradientResultsSet::add
      tc_set& t = S[j];
      const tc_gradient& rec = w.grec[c];
      const int has = w.ghas[c];
      for (int l = 0; l < 2; ++l)
      {
        if (!(has & (1 << (2 * l))))
          continue;
        if (rec.cc_type[l] != TC_CC_TIME1)
        {
          t.has_error[l][0] = 1;
          t.error[l][0] = fmax(t.error[l][0], rec.error);
          t.error_with_buffer[l][0] = fmax(t.error_with_buffer[l][0], rec.error_with_buffer);
        }
        if (rec.cc_type[l] != TC_CC_TIME0)
        {
          t.has_error[l][1] = 1;
          t.error[l][1] = fmax(t.error[l][1], rec.error);
          t.error_with_buffer[l][1] = fmax(t.error_with_buffer[l][1], rec.error_with_buffer);
        }
      }
      t.n_results += 1;
    }
    // std::map order of the shape keys inside one link pair
    for (int a = first + 1; a < ns; ++a)
    {
      const tc_set x = S[a];
      const int r0 = RUN[2 * a], r1 = RUN[2 * a + 1];
      int j = a;
      while (j > first && (x.shape_key[0] < S[j - 1].shape_key[0] ||
                           (x.shape_key[0] == S[j - 1].shape_key[0] && x.shape_key[1] < S[j - 1].shape_key[1])))
      {
        S[j] = S[j - 1];
        RUN[2 * j] = RUN[2 * j - 2], RUN[2 * j + 1] = RUN[2 * j - 1];
        --j;
      }
      S[j] = x;
      RUN[2 * j] = r0, RUN[2 * j + 1] = r1;
    }
  }
  set_count[k] = ns;
  if (ns > max_num_cnt)
  {
    const int which = single ? 0 : ((!f0 && !f1) ? 2 : (!f0 ? 0 : 1));
    // stable insertion sort, largest error first
    for (int a = 1; a < ns; ++a)
    {
      const tc_set x = S[a];
      const int r0 = RUN[2 * a], r1 = RUN[2 * a + 1];
      int j = a;
      while (j > 0)
      {
        bool before;  // x strictly before S[j-1]
        if (single)
        {
          // std::array<double,2> operator> on max_error[0].error
          const double *xa = x.error[0], *ya = S[j - 1].error[0];
          before = (ya[0] < xa[0]) || (!(xa[0] < ya[0]) && ya[1] < xa[1]);
        }
        else
        {
          double ex, ey;
          setMaxError(x, which, true, ex);
          setMaxError(S[j - 1], which, true, ey);
          before = ex > ey;
        }
        if (!before)
          break;
        S[j] = S[j - 1];
        RUN[2 * j] = RUN[2 * j - 2], RUN[2 * j + 1] = RUN[2 * j - 1];
        --j;
      }
      S[j] = x;
      RUN[2 * j] = r0, RUN[2 * j + 1] = r1;
    }
  }
  double* val = values + static_cast<size_t>(k) * max_num_cnt;
  double* cf = coeffs + static_cast<size_t>(k) * max_num_cnt;
  double* J = jac + static_cast<size_t>(k) * max_num_cnt * 2 * n;
  for (int a = 0; a < max_num_cnt; ++a)
  {
    val[a] = -v.margin_buffer;
    cf[a] = 1.0;
  }
  for (int a = 0; a < max_num_cnt * 2 * n; ++a)
    J[a] = 0.0;
  const int cnt = ns < max_num_cnt ? ns : max_num_cnt;
  for (int a = 0; a < cnt; ++a)
  {
    const tc_set& g = S[a];
    const int r0 = RUN[2 * a], r1 = RUN[2 * a + 1];
    cf[a] = g.coeff;
    double* J0 = J + (static_cast<size_t>(a) * 2 + 0) * n;
    double* J1 = J + (static_cast<size_t>(a) * 2 + 1) * n;
    double e, m;
    if (single)
    {
      setMaxError(g, 0, false, e);
      val[a] = e;
      setMaxError(g, 0, true, m);
      weightedAvgRow(v, w, g, r0, r1, m, false, J0);
      continue;
    }
    const bool h0 = g.has_error[0][0] || g.has_error[1][0];
    const bool h1 = g.has_error[0][1] || g.has_error[1][1];
    if (!f0 && !f1)
    {
      setMaxError(g, 2, false, e);
      val[a] = e;
    }
    else if (!f0)
    {
      if (h0)
      {
        setMaxError(g, 0, false, e);
        val[a] = e;
      }
    }
    else if (h1)
    {
      setMaxError(g, 1, false, e);
      val[a] = e;
    }
    if (!f0 && h0)
    {
      setMaxError(g, !f1 ? 2 : 0, true, m);
      weightedAvgRow(v, w, g, r0, r1, m, false, J0);
    }
    if (!f1 && h1)
    {
      setMaxError(g, !f0 ? 2 : 1, true, m);
      weightedAvgRow(v, w, g, r0, r1, m, true, J1);
    }
  }
}

inline unsigned gridFor(long long n, int tpb) { return static_cast<unsigned>((n + tpb - 1) / tpb); }
}  // namespace

namespace
{
// copies the counters into mapped host memory with ordinary stores (no copy engine involved)
__global__ void k_snapshot(const Counters* dev, Counters* host)
{
  constexpr int N = sizeof(Counters) / sizeof(long long);
  if (threadIdx.x < N)
    reinterpret_cast<long long*>(host)[threadIdx.x] = reinterpret_cast<const long long*>(dev)[threadIdx.x];
  __threadfence_system();
}
}  // namespace

namespace
{
template <class W>
__global__ void k_fill(W* dst, W word, size_t n)
{
  for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < n; i += static_cast<size_t>(gridDim.x) * blockDim.x)
    dst[i] = word;
}
template <class W>
__global__ void k_copy(W* dst, const W* src, size_t n)
{
  for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < n; i += static_cast<size_t>(gridDim.x) * blockDim.x)
    dst[i] = src[i];
}
inline unsigned gridFor(size_t n)
{
  const size_t b = (n + 255) / 256;
  return static_cast<unsigned>(b < 1 ? 1 : (b > 148 * 8 ? 148 * 8 : b));
}
}  // namespace

int& launchCounter()
{
  static thread_local int n = 0;
  return n;
}

cudaError_t launchFill(void* dst, uint32_t word, size_t n_words, cudaStream_t st)
{
  if (n_words == 0)
    return cudaSuccess;
  ++launchCounter();
  k_fill<uint32_t><<<gridFor(n_words), 256, 0, st>>>(static_cast<uint32_t*>(dst), word, n_words);
  return cudaGetLastError();
}

cudaError_t launchCopy32(void* dst, const void* src, size_t n_words, cudaStream_t st)
{
  if (n_words == 0)
    return cudaSuccess;
  ++launchCounter();
  k_copy<uint32_t><<<gridFor(n_words), 256, 0, st>>>(static_cast<uint32_t*>(dst), static_cast<const uint32_t*>(src), n_words);
  return cudaGetLastError();
}

cudaError_t launchCopy64(void* dst, const void* src, size_t n_dwords, cudaStream_t st)
{
  if (n_dwords == 0)
    return cudaSuccess;
  ++launchCounter();
  k_copy<unsigned long long><<<gridFor(n_dwords), 256, 0, st>>>(static_cast<unsigned long long*>(dst),
                                                                static_cast<const unsigned long long*>(src), n_dwords);
  return cudaGetLastError();
}

cudaError_t launchSnapshot(const Counters* dev, Counters* mapped_host, cudaStream_t st)
{
  ++launchCounter();
  k_snapshot<<<1, 32, 0, st>>>(dev, mapped_host);
  return cudaGetLastError();
}

cudaError_t launchPlan(const SceneView& v, const Work& w, cudaStream_t st)
{
  ++launchCounter();
  k_plan<<<gridFor(w.ep.n_terms, TPB), TPB, 0, st>>>(v, w);
  ++launchCounter();
  k_scan_terms<<<1, 1024, 0, st>>>(w);
  return cudaGetLastError();
}

cudaError_t launchBroadCount(const SceneView& v, const Work& w, cudaStream_t st)
{
  cudaError_t e = launchFill(w.sp_count, 0, v.n_spairs, st);
  if (e != cudaSuccess)
    return e;
  e = launchFill(w.sp_cursor, 0, v.n_spairs, st);
  if (e != cudaSuccess)
    return e;
  const dim3 grid(gridFor(w.n_units, TPB * TC_BROAD_CPB), static_cast<unsigned>(v.n_active_shapes));
  ++launchCounter();
  k_broad<false><<<grid, TPB, 0, st>>>(v, w);
  ++launchCounter();
  k_scan_sp<<<1, 1024, 0, st>>>(v, w);
  return cudaGetLastError();
}

cudaError_t launchBroadFill(const SceneView& v, const Work& w, cudaStream_t st)
{
  cudaError_t e = launchFill(w.items, 0xffffffffu, static_cast<size_t>(w.n_items_padded), st);
  if (e != cudaSuccess)
    return e;
  const dim3 grid(gridFor(w.n_units, TPB * TC_BROAD_CPB), static_cast<unsigned>(v.n_active_shapes));
  ++launchCounter();
  k_broad<true><<<grid, TPB, 0, st>>>(v, w);
  return cudaGetLastError();
}

cudaError_t launchFinalize(const SceneView& v, const Work& w, int phase, cudaStream_t st)
{
  if (w.n_sort > 0)
  {
    ++launchCounter();
    k_finalize<<<gridFor(w.n_sort, TPB), TPB, 0, st>>>(v, w, phase);
  }
  return cudaGetLastError();
}

cudaError_t launchSnapshotRaw(const Work& w, cudaStream_t st)
{
  ++launchCounter();
  k_snapshot_raw<<<1, 32, 0, st>>>(w);
  return cudaGetLastError();
}

cudaError_t launchTypeFilter(const SceneView& v, const Work& w, const uint8_t* valid, cudaStream_t st)
{
  (void)v;
  if (w.n_sort > 0)
  {
    ++launchCounter();
    k_type_select<<<gridFor(w.n_sort, TPB), TPB, 0, st>>>(w, valid);
  }
  return cudaGetLastError();
}

cudaError_t sortContacts(const Work& w, void* temp, size_t& temp_bytes, cudaStream_t st)
{
  // key = term << 36 | pair << 22 | substep << 8 | k: only the bits below 36 + bits(n_terms) can differ between valid
  // keys; the all-ones key of a dropped contact stays the largest within any bit range
  int end_bit = 37;
  while (end_bit < 64 && (static_cast<long long>(w.ep.n_terms) >> (end_bit - 36)) != 0)
    ++end_bit;
  return cub::DeviceRadixSort::SortPairs(temp, temp_bytes, w.keys, w.keys_alt, w.idx, w.idx_alt,
                                         static_cast<int>(w.n_sort), 0, end_bit, st);
}

cudaError_t launchGather(const SceneView& v, const Work& w, cudaStream_t st)
{
  (void)v;
  if (w.n_contacts > 0 && !w.gather_in_rows)
  {
    const long long words = w.n_contacts * (static_cast<long long>(sizeof(tc_contact)) / 8);
    ++launchCounter();
    k_gather<<<gridFor(words, 256), 256, 0, st>>>(w, w.idx_alt);
  }
  ++launchCounter();
  k_term_ptr<<<gridFor(w.ep.n_terms + 1, 256), 256, 0, st>>>(w, w.keys_alt);
  return cudaGetLastError();
}

cudaError_t launchTransforms(const SceneView& v, const Work& w, double* xf, cudaStream_t st)
{
  if (w.n_contacts > 0)
  {
    ++launchCounter();
    k_transforms<<<gridFor(w.n_contacts, TPB), TPB, 0, st>>>(v, w, xf);
  }
  return cudaGetLastError();
}

cudaError_t launchGradients(const SceneView& v, const Work& w, cudaStream_t st)
{
  if (w.n_contacts > 0)
  {
    ++launchCounter();
    if (v.n_dof <= 8)
      k_grad<8><<<gridFor(w.n_contacts, TPB), TPB, 0, st>>>(v, w);
    else
      k_grad<64><<<gridFor(w.n_contacts, TPB), TPB, 0, st>>>(v, w);
  }
  return cudaGetLastError();
}

cudaError_t launchRows(const SceneView& v, const Work& w, long long* term_row_base, int32_t* col_idx, double* val,
                       double* row_value, double* row_coeff, double* row_margin, cudaStream_t st)
{
  ++launchCounter();
  k_rows<<<gridFor(w.n_contacts + 1, TPB), TPB, 0, st>>>(v, w, term_row_base, col_idx, val, row_value, row_coeff,
                                                       row_margin);
  return cudaGetLastError();
}

cudaError_t launchRowWidths(const SceneView& v, const Work& w, void* temp, size_t& temp_bytes, cudaStream_t st)
{
  const int n = static_cast<int>(w.n_contacts + 1);
  if (!temp)
    return cub::DeviceScan::ExclusiveSum(nullptr, temp_bytes, w.row_width, w.row_ptr, n, st);
  ++launchCounter();
  k_row_width<<<gridFor(n, 256), 256, 0, st>>>(v, w, w.keys_alt);
  return cub::DeviceScan::ExclusiveSum(temp, temp_bytes, w.row_width, w.row_ptr, n, st);
}

cudaError_t launchRowBase(const SceneView& v, const Work& w, long long* term_row_base, cudaStream_t st)
{
  ++launchCounter();
  k_row_base<<<1, 1024, 0, st>>>(v, w, term_row_base);
  return cudaGetLastError();
}

cudaError_t launchGradRows(const SceneView& v, const Work& w, long long* term_row_base, int32_t* col_idx, double* val,
                           double* row_value, double* row_coeff, double* row_margin, cudaStream_t st)
{
  ++launchCounter();
  const unsigned grid = gridFor(w.n_contacts + 1, TPB);
  if (w.gather_in_rows)
  {
    if (v.n_dof <= 8)
      k_grad_rows_regs<8, true><<<grid, TPB, 0, st>>>(v, w, term_row_base, col_idx, val, row_value, row_coeff, row_margin);
    else
      k_grad_rows<64, true><<<grid, TPB, 0, st>>>(v, w, term_row_base, col_idx, val, row_value, row_coeff, row_margin);
  }
  else if (v.n_dof <= 8)
    k_grad_rows_regs<8, false><<<grid, TPB, 0, st>>>(v, w, term_row_base, col_idx, val, row_value, row_coeff, row_margin);
  else
    k_grad_rows<64, false><<<grid, TPB, 0, st>>>(v, w, term_row_base, col_idx, val, row_value, row_coeff, row_margin);
  return cudaGetLastError();
}

cudaError_t launchCost(const SceneView& v, const Work& w, double* cost_term, double* cost_traj, cudaStream_t st)
{
  ++launchCounter();
  k_cost_term<<<gridFor(static_cast<long long>(w.ep.n_terms) * 32, TPB), TPB, 0, st>>>(v, w, cost_term);
  ++launchCounter();
  k_cost_traj<<<gridFor(w.ep.B, TPB), TPB, 0, st>>>(w, cost_term, cost_traj);
  return cudaGetLastError();
}

cudaError_t launchBestRecord(const Work& w, const double* cost_traj, long long index_offset, double* rec, cudaStream_t st)
{
  ++launchCounter();
  // n_dof is not in Work: the trajectory stride follows from the batch layout q[B][T][n_dof]
  k_best_record<<<1, 1024, 0, st>>>(w, cost_traj, index_offset, rec, w.n_dof_q);
  return cudaGetLastError();
}

cudaError_t launchQpCsc(const SceneView& v, const Work& w, long long* bt_base, const double* val, const double* row_value,
                        const double* row_coeff, const double* row_margin, long long* col_ptr, int32_t* row_idx,
                        double* out, double* upper, double* slack_cost, cudaStream_t st)
{
  ++launchCounter();
  k_csc_base<<<1, 1024, 0, st>>>(v, w, bt_base);
  const long long n_cols = static_cast<long long>(w.ep.B) * w.ep.T * v.n_dof + w.n_contacts + 1;
  ++launchCounter();
  k_csc_colptr<<<gridFor(n_cols, 256), 256, 0, st>>>(v, w, bt_base, col_ptr);
  if (w.n_contacts > 0)
  {
    ++launchCounter();
    k_csc_fill<<<gridFor(w.n_contacts, TPB), TPB, 0, st>>>(v, w, bt_base, val, row_value, row_coeff, row_margin, row_idx,
                                                         out, upper, slack_cost);
  }
  return cudaGetLastError();
}

cudaError_t launchSets(const SceneView& v, const Work& w, int max_num_cnt, tc_set* sets, int32_t* set_count,
                       int32_t* set_run, double* values, double* coeffs, double* jac, cudaStream_t st)
{
  ++launchCounter();
  k_sets<<<gridFor(w.ep.n_terms, 64), 64, 0, st>>>(v, w, max_num_cnt, sets, set_count, set_run, values, coeffs, jac);
  return cudaGetLastError();
}
}  // namespace tc
