// tc_epa_coop.cuh — EPA with EIGHT LANES PER ITEM (four items per warp), polytope in shared memory.
//
// Same algorithm, same decisions and the same order of every tie-break as the thread-per-item EPA of tc_narrow.cuh
// (which stays the reference form: the host-compiled numerics check and the deep-expansion kernel use it).  Why a
// second form: one thread per item keeps ~1.2 KB of hot polytope per thread in thread-local memory and needs 512
// threads per SM to hide latency — five times what L1 holds (ncu: 58 % of stall samples long_scoreboard, 8.7 GB of
// DRAM traffic that is polytope lines falling out of L1/L2) — and its warps run at 13.7 of 32 lanes because the
// expansion counts of 32 items differ.  Here a group of 8 lanes owns one item:
//   * the polytope of the 16 items a CTA works on lives in shared memory (3.4 KB per item),
//   * support queries, the face-visibility pass, horizon extraction, face creation and the witness search are spread
//     over the 8 lanes and closed with 8-lane shuffles / ballots,
//   * a group that finishes draws the next item of its warp's 32-item batch at once.
// Every reduction is written as the lexicographic extremum (value, then lowest index / slot) that the sequential loops
// with their strict comparisons compute, so results do not depend on how the work is spread over lanes.
#pragma once
#include "tc_internal.hpp"

namespace tc
{
namespace coop
{
constexpr int EC_V = 24;              // polytope capacity of the fast path; items that outgrow it go to k_narrow_deep
constexpr int EC_F = 2 * EC_V - 4;
constexpr int GROUP = 8;
constexpr int ITEMS_PER_CTA = 128 / GROUP;

struct Item  // shared memory, one per 8-lane group
{
  double pose[4][12];  // A at T0, A at T1, B at T0, B at T1
  double pw[EC_V][3], pa[EC_V][3];
  double fn[EC_F][3], fd[EC_F];
  int fv[EC_F];
  int edges[EC_F];
  unsigned erow[EC_V];
};

struct Lanes
{
  unsigned gmask;  // the 8 lanes of this group
  int g, gl;       // group within the warp, lane within the group
};

__device__ __forceinline__ double gxor(const Lanes& L, double v, int o) { return __shfl_xor_sync(L.gmask, v, o, GROUP); }
__device__ __forceinline__ int gxor(const Lanes& L, int v, int o) { return __shfl_xor_sync(L.gmask, v, o, GROUP); }
__device__ __forceinline__ unsigned gballot(const Lanes& L, bool p)
{
  return (__ballot_sync(L.gmask, p) >> (GROUP * L.g)) & 0xffu;
}
__device__ __forceinline__ void gsync(const Lanes& L) { __syncwarp(L.gmask); }

// (largest value, lowest index): what a scan in index order with `if (v > best)` keeps
__device__ __forceinline__ void gArgMaxFirst(const Lanes& L, double& val, int& idx)
{
#pragma unroll
  for (int o = GROUP / 2; o > 0; o >>= 1)
  {
    const double ov = gxor(L, val, o);
    const int oi = gxor(L, idx, o);
    if (ov > val || (ov == val && oi < idx))
      val = ov, idx = oi;
  }
}
// (smallest value, lowest index): what a scan in index order with `if (v < best)` keeps
__device__ __forceinline__ void gArgMinFirst(const Lanes& L, double& val, int& idx)
{
#pragma unroll
  for (int o = GROUP / 2; o > 0; o >>= 1)
  {
    const double ov = gxor(L, val, o);
    const int oi = gxor(L, idx, o);
    if (ov < val || (ov == val && oi < idx))
      val = ov, idx = oi;
  }
}

__device__ __forceinline__ int nthSetBit(unsigned long long m, int n)
{
  for (int k = 0; k < n; ++k)
    m &= m - 1;
  return __ffsll(static_cast<long long>(m)) - 1;
}

__device__ __forceinline__ Vec3<double> ld3(const double* p) { return Vec3<double>(p[0], p[1], p[2]); }
__device__ __forceinline__ void st3(double* p, Vec3<double> v) { p[0] = v.x, p[1] = v.y, p[2] = v.z; }

constexpr int NO_INDEX = 0x7fffffff;
constexpr double BIG = 1.7976931348623157e308;  // = -lowestReal<double>(), the DEGENERATE plane distance

// Placed::support of tc_narrow.cuh with the vertex comparisons spread over the group.  verts / lut: global memory
// (read through L1; the look-up table keeps it to a handful of vertices).  Every lane returns the same point.
__device__ __forceinline__ Vec3<double> support(const Lanes& L, const ShapeD& S, const Vtx<double>* verts,
                                                const uint16_t* lut, const double* pose0, const double* pose1,
                                                bool cast, Vec3<double> d, int& scanned)
{
  if (S.type != SH_HULL)
  {
    const ShapeCore<double> core = makeCore<double>(S, nullptr);
    return primitiveSupportPlaced(core, Pose<double>::from12(pose0), Pose<double>::from12(pose1), cast, d);
  }
  Vec3<double> wbest;
  for (int pz = 0; pz < (cast ? 2 : 1); ++pz)
  {
    const Pose<double> P = Pose<double>::from12(pz ? pose1 : pose0);
    const Vec3<double> dl = P.rotT(d);
    double best = -BIG;
    int bi = NO_INDEX;
    int c = -1;
    if (lut)
      c = lutCellF(static_cast<float>(dl.x), static_cast<float>(dl.y), static_cast<float>(dl.z));
    if (c >= 0)
    {
      const uint8_t* cand = reinterpret_cast<const uint8_t*>(lut + (LUT_CELLS + 1));
      const int e = lut[c + 1];
      for (int k = lut[c] + L.gl; k < e; k += GROUP)
      {
        const int i = cand[k];
        const Vtx<double> q = verts[i];
        const double val = dl.x * q.x + dl.y * q.y + dl.z * q.z;
        if (val > best)
          best = val, bi = i;
      }
    }
    else
    {
      for (int i = L.gl; i < S.vert_cnt; i += GROUP)
      {
        const Vtx<double> q = verts[i];
        const double val = dl.x * q.x + dl.y * q.y + dl.z * q.z;
        if (val > best)
          best = val, bi = i;
      }
    }
    gArgMaxFirst(L, best, bi);
    if (bi == NO_INDEX)
      bi = 0;  // every value was -inf / NaN: the plain scan keeps index 0
    const Vtx<double> q = verts[bi];
    const Vec3<double> wv = P.apply(Vec3<double>(q.x, q.y, q.z));
    scanned += S.vert_cnt;
    if (pz == 0)
      wbest = wv;
    else
      wbest = (dot(d, wbest) > dot(d, wv)) ? wbest : wv;
  }
  return wbest;
}

// per-group EPA state (uniform over the 8 lanes)
struct State
{
  int np, nf, best, iters;
  unsigned long long alive;
  double bd, eps_len, vis_eps;
};

enum
{
  C_CONTINUE = 0,
  C_DONE = 1,
  C_UNCONVERGED = 2,
  C_OVERFLOW = 3,
  C_FLAT = 4
};

struct Shapes  // the shape pair of the group's item
{
  const ShapeD *A, *B;
  const Vtx<double>*va, *vb;
  const uint16_t *la, *lb;
  bool cast_a, cast_b;
};

__device__ __forceinline__ Vec3<double> supA(const Lanes& L, const Shapes& sh, const Item& it, Vec3<double> d, int& sc)
{
  return support(L, *sh.A, sh.va, sh.la, it.pose[0], it.pose[1], sh.cast_a, d, sc);
}
__device__ __forceinline__ Vec3<double> supB(const Lanes& L, const Shapes& sh, const Item& it, Vec3<double> d, int& sc)
{
  return support(L, *sh.B, sh.vb, sh.lb, it.pose[2], it.pose[3], sh.cast_b, d, sc);
}

// face (i, j, k) of the initial tetrahedron, oriented away from the centroid; written to slot `slot`
__device__ __forceinline__ double initFace(Item& it, int slot, int i, int j, int k, Vec3<double> centroid)
{
  const Vec3<double> pi = ld3(it.pw[i]);
  Vec3<double> n = cross(ld3(it.pw[j]) - pi, ld3(it.pw[k]) - pi);
  if (dot(n, pi - centroid) < 0)
  {
    const int t = j;
    j = k;
    k = t;
    n = -n;
  }
  const double nn = norm(n);
  it.fv[slot] = i | (j << 10) | (k << 20);
  double d;
  if (nn > Tol<double>::tiny())
  {
    const Vec3<double> un = (1.0 / nn) * n;
    st3(it.fn[slot], un);
    d = dot(un, pi);
  }
  else
  {
    st3(it.fn[slot], Vec3<double>(0, 0, 0));
    d = BIG;
  }
  it.fd[slot] = d;
  return d;
}

// epaInit of tc_narrow.cuh.  pw/pa[0 .. n_start) hold the GJK simplex.  Returns C_CONTINUE (polytope ready) or C_FLAT
// (flat Minkowski difference: normal / pa / pb are final, depth 0).
__device__ __forceinline__ int init(const Lanes& L, const Shapes& sh, Item& it, State& E, int n_start, double scale,
                                    Vec3<double>& normal, Vec3<double>& pa_out, Vec3<double>& pb_out, int& scanned)
{
  E.np = n_start;
  E.eps_len = Tol<double>::epsLenRel() * tc_max(scale, 1e-3);
  E.vis_eps = Tol<double>::visRel() * tc_max(scale, 1e-3);
  if (E.np == 4)
  {
    const Vec3<double> p0 = ld3(it.pw[0]);
    const double vol = dot(cross(ld3(it.pw[1]) - p0, ld3(it.pw[2]) - p0), ld3(it.pw[3]) - p0);
    if (tc_abs(vol) <= E.eps_len * E.eps_len * E.eps_len)
      E.np = 3;
  }
  for (int k = 0; E.np < 4;)
  {
    const int lim = (E.np == 1) ? 6 : ((E.np == 2) ? 4 : 2);
    if (k >= lim)
      break;
    Vec3<double> d, e, n;
    double en = 0, area = 0;
    const Vec3<double> p0 = ld3(it.pw[0]);
    if (E.np == 1)
    {
      const double sg = (k & 1) ? -1.0 : 1.0;
      d = Vec3<double>((k >> 1) == 0 ? sg : 0.0, (k >> 1) == 1 ? sg : 0.0, (k >> 1) == 2 ? sg : 0.0);
    }
    else if (E.np == 2)
    {
      e = ld3(it.pw[1]) - p0;
      const Vec3<double> d0 = leastAxisCross(e);
      const Vec3<double> d1 = cross(e, d0);
      en = norm(e);
      d = (k == 0) ? d0 : ((k == 1) ? -d0 : ((k == 2) ? d1 : -d1));
    }
    else
    {
      n = cross(ld3(it.pw[1]) - p0, ld3(it.pw[2]) - p0);
      area = norm(n);
      d = (k == 0) ? n : -n;
    }
    const Vec3<double> a = supA(L, sh, it, d, scanned), b = supB(L, sh, it, -d, scanned);
    const Vec3<double> wv = a - b;
    bool accept;
    if (E.np == 1)
      accept = norm(wv - p0) > E.eps_len;
    else if (E.np == 2)
      accept = norm(cross(wv - p0, e)) > E.eps_len * en;
    else
      accept = tc_abs(dot(wv - p0, n)) > E.eps_len * area;
    ++k;
    if (accept)
    {
      gsync(L);
      if (L.gl == 0)
      {
        st3(it.pw[E.np], wv);
        st3(it.pa[E.np], a);
      }
      gsync(L);
      E.np += 1;
      k = 0;
    }
  }
  if (E.np < 4)
  {
    Vec3<double> n(0, 0, 1);
    const Vec3<double> p0 = ld3(it.pw[0]);
    if (E.np == 3)
      n = cross(ld3(it.pw[1]) - p0, ld3(it.pw[2]) - p0);
    else if (E.np == 2)
      n = leastAxisCross(ld3(it.pw[1]) - p0);
    const double nn = norm(n);
    normal = (nn > 0) ? (1.0 / nn) * n : Vec3<double>(0, 0, 1);
    Simplex<double> start;
    start.n = n_start;
    for (int i = 0; i < 4; ++i)
    {
      const int src = i < n_start ? i : 0;
      start.w[i] = ld3(it.pw[src]);
      start.a[i] = ld3(it.pa[src]);
      start.lam[i] = 0;
    }
    Vec3<double> pa, pwv;
    if (n_start == 1)
    {
      pa = start.a[0];
      pwv = start.w[0];
    }
    else
    {
      double lam[4];
      int mask;
      closestOnSimplex(start, n_start, lam, mask);
      pa = lam[0] * start.a[0] + lam[1] * start.a[1];
      pwv = lam[0] * start.w[0] + lam[1] * start.w[1];
      if (n_start > 2)
        pa = pa + lam[2] * start.a[2], pwv = pwv + lam[2] * start.w[2];
      if (n_start > 3)
        pa = pa + lam[3] * start.a[3], pwv = pwv + lam[3] * start.w[3];
    }
    pa_out = pa;
    pb_out = pa - pwv;
    return C_FLAT;
  }
  const Vec3<double> centroid = 0.25 * (ld3(it.pw[0]) + ld3(it.pw[1]) + ld3(it.pw[2]) + ld3(it.pw[3]));
  // faces (0,1,2), (0,1,3), (0,2,3), (1,2,3) in slots 0..3, one lane each
  double d = BIG;
  int slot = NO_INDEX;
  if (L.gl < 4)
  {
    const int i = (L.gl == 3) ? 1 : 0;
    const int j = (L.gl < 2) ? 1 : 2;
    const int k = (L.gl == 0) ? 2 : 3;
    d = initFace(it, L.gl, i, j, k, centroid);
    slot = L.gl;
  }
  gArgMinFirst(L, d, slot);
  E.nf = 4;
  E.alive = 15ull;
  E.best = slot;
  E.bd = d;
  E.iters = 1;
  gsync(L);
  return C_CONTINUE;
}

// epaTripMasked of tc_narrow.cuh: one support query + one expansion of the polytope
__device__ __forceinline__ int trip(const Lanes& L, const Shapes& sh, Item& it, State& E, int& scanned)
{
  const Vec3<double> bn = ld3(it.fn[E.best]);
  const Vec3<double> xa = supA(L, sh, it, bn, scanned), xb = supB(L, sh, it, -bn, scanned);
  const Vec3<double> xw = xa - xb;
  if (dot(bn, xw) - E.bd <= Tol<double>::epaTol())
    return C_DONE;
  if (E.iters == EPA_MAX_ITERS)
    return C_UNCONVERGED;
  if (E.np >= EC_V)
    return C_OVERFLOW;
  // 1. visibility of every live face + closest surviving face, 8 faces per round
  unsigned long long vis = 0;
  double sbd = BIG;
  int sbest = NO_INDEX;
  for (int base = 0; base < E.nf; base += GROUP)
  {
    const int f = base + L.gl;
    const bool al = f < E.nf && ((E.alive >> f) & 1ull);
    double fdv = 0;
    bool visible = false;
    if (al)
    {
      fdv = it.fd[f];
      visible = (fdv != BIG) && (dot(ld3(it.fn[f]), xw) - fdv > E.vis_eps);
      if (!visible && (sbest == NO_INDEX || fdv < sbd))
        sbd = fdv, sbest = f;
    }
    vis |= static_cast<unsigned long long>(gballot(L, visible)) << base;
  }
  if (vis == 0)
    return C_DONE;  // numerically x is not outside any face although the gap test said so: accept the face
  // a lane without a surviving face must lose every comparison, including against a DEGENERATE (= BIG) survivor
  {
    // reduce on (has, value, slot): lanes without a candidate carry slot NO_INDEX and value BIG, and a real candidate
    // with value BIG has a smaller slot, so (value, slot) alone orders them correctly
    gArgMinFirst(L, sbd, sbest);
  }
  // 2. directed edges of the visible faces into the bit matrix (an interior edge has its twin in the matrix too)
  const int nvis = __popcll(vis);
  for (int j = L.gl; j < nvis; j += GROUP)
  {
    const int fv = it.fv[nthSetBit(vis, j)];
    const int v0 = fv & 1023, v1 = (fv >> 10) & 1023, v2 = (fv >> 20) & 1023;
    atomicOr(&it.erow[v0], 1u << v1);
    atomicOr(&it.erow[v1], 1u << v2);
    atomicOr(&it.erow[v2], 1u << v0);
  }
  gsync(L);
  // 3. horizon = edges whose twin is absent, listed in (face slot, edge) order
  int ne = 0;
  bool overflow = false;
  for (int jb = 0; jb < nvis; jb += GROUP)
  {
    const int j = jb + L.gl;
    int v0 = 0, v1 = 0, v2 = 0;
    unsigned h = 0;
    if (j < nvis)
    {
      const int fv = it.fv[nthSetBit(vis, j)];
      v0 = fv & 1023, v1 = (fv >> 10) & 1023, v2 = (fv >> 20) & 1023;
      h = (((it.erow[v1] >> v0) & 1u) ? 0u : 1u) | (((it.erow[v2] >> v1) & 1u) ? 0u : 2u) |
          (((it.erow[v0] >> v2) & 1u) ? 0u : 4u);
    }
    const int c = __popc(h);
    int incl = c;  // inclusive scan over the group
#pragma unroll
    for (int o = 1; o < GROUP; o <<= 1)
    {
      const int t = __shfl_up_sync(L.gmask, incl, o, GROUP);
      if (L.gl >= o)
        incl += t;
    }
    const int total = __shfl_sync(L.gmask, incl, GROUP - 1, GROUP);
    int q = ne + incl - c;
    if (h & 1u)
    {
      if (q < EC_F)
        it.edges[q] = v0 | (v1 << 16);
      else
        overflow = true;
      ++q;
    }
    if (h & 2u)
    {
      if (q < EC_F)
        it.edges[q] = v1 | (v2 << 16);
      else
        overflow = true;
      ++q;
    }
    if (h & 4u)
    {
      if (q < EC_F)
        it.edges[q] = v2 | (v0 << 16);
      else
        overflow = true;
      ++q;
    }
    ne += total;
  }
  gsync(L);
  // the matrix goes back to all-zero
  for (int j = L.gl; j < nvis; j += GROUP)
  {
    const int fv = it.fv[nthSetBit(vis, j)];
    it.erow[fv & 1023] = 0, it.erow[(fv >> 10) & 1023] = 0, it.erow[(fv >> 20) & 1023] = 0;
  }
  if (gballot(L, overflow) != 0)
    return C_OVERFLOW;
  if (ne == 0)
    return C_DONE;
  // 4. new vertex, one new face per horizon edge (first dead slots first, then appended), closest face
  E.alive &= ~vis;
  const int xi = E.np;
  if (L.gl == 0)
  {
    st3(it.pw[xi], xw);
    st3(it.pa[xi], xa);
  }
  E.np += 1;
  const unsigned long long dead = ~E.alive & ((E.nf >= 64) ? ~0ull : ((1ull << E.nf) - 1ull));
  const int ndead = __popcll(dead);
  if (E.nf + (ne > ndead ? ne - ndead : 0) > EC_F)
    return C_OVERFLOW;
  double bd = BIG;
  int best = NO_INDEX;
  for (int q = L.gl; q < ne; q += GROUP)
  {
    const int slot = q < ndead ? nthSetBit(dead, q) : E.nf + (q - ndead);
    const int e = it.edges[q];
    const int a = e & 65535, b = e >> 16;
    const Vec3<double> pi = ld3(it.pw[a]);
    const Vec3<double> n = cross(ld3(it.pw[b]) - pi, xw - pi);
    const double nn = norm(n);
    it.fv[slot] = a | (b << 10) | (xi << 20);
    double d;
    if (nn > Tol<double>::tiny())
    {
      const Vec3<double> un = (1.0 / nn) * n;
      st3(it.fn[slot], un);
      d = dot(un, pi);
    }
    else
    {
      st3(it.fn[slot], Vec3<double>(0, 0, 0));
      d = BIG;
    }
    it.fd[slot] = d;
    if (best == NO_INDEX || d < bd)
      bd = d, best = slot;  // slots of one lane are not ascending in general: the reduction below settles ties
    else if (d == bd && slot < best)
      best = slot;
  }
  gArgMinFirst(L, bd, best);
  // the surviving closest face competes with the new ones under the same (distance, slot) order
  if (sbest != NO_INDEX && (best == NO_INDEX || sbd < bd || (sbd == bd && sbest < best)))
    bd = sbd, best = sbest;
  for (int q = 0; q < ne; ++q)
    E.alive |= 1ull << (q < ndead ? nthSetBit(dead, q) : E.nf + (q - ndead));
  if (ne > ndead)
    E.nf += ne - ndead;
  E.best = best;
  E.bd = bd;
  E.iters += 1;
  gsync(L);
  return C_CONTINUE;
}

// epaFinish of tc_narrow.cuh: depth, normal, witness points on the live face closest to the origin
__device__ __forceinline__ void finish(const Lanes& L, const Item& it, const State& E, double& depth,
                                       Vec3<double>& normal, Vec3<double>& pa_out, Vec3<double>& pb_out)
{
  depth = E.bd;
  normal = ld3(it.fn[E.best]);
  double wbest = BIG, la = 0, lb = 0, lc = 0;
  int wf = NO_INDEX;
  for (int base = 0; base < E.nf; base += GROUP)
  {
    const int f = base + L.gl;
    if (f < E.nf && ((E.alive >> f) & 1ull))
    {
      const double fdv = it.fd[f];
      if (fdv != BIG && !(fdv > depth + E.eps_len))
      {
        const int gv = it.fv[f];
        double a0, a1, a2;
        const Vec3<double> q = closestTriangle(ld3(it.pw[gv & 1023]), ld3(it.pw[(gv >> 10) & 1023]),
                                               ld3(it.pw[(gv >> 20) & 1023]), a0, a1, a2);
        const double q2 = norm2(q);
        if (q2 < wbest)
          wbest = q2, wf = f, la = a0, lb = a1, lc = a2;
      }
    }
  }
  double rbest = wbest;
  int rf = wf;
  gArgMinFirst(L, rbest, rf);
  if (rf == NO_INDEX || !(rbest < BIG))
  {
    // no face qualified: the sequential form keeps the closest-plane face with zero weights
    rf = E.best;
    la = lb = lc = 0;
  }
  else
  {
    const unsigned owner = gballot(L, wf == rf);
    const int src = __ffs(owner) - 1;
    la = __shfl_sync(L.gmask, la, src, GROUP);
    lb = __shfl_sync(L.gmask, lb, src, GROUP);
    lc = __shfl_sync(L.gmask, lc, src, GROUP);
  }
  const int gv = it.fv[rf];
  const int v0 = gv & 1023, v1 = (gv >> 10) & 1023, v2 = (gv >> 20) & 1023;
  pa_out = la * ld3(it.pa[v0]) + lb * ld3(it.pa[v1]) + lc * ld3(it.pa[v2]);
  const Vec3<double> pwv = la * ld3(it.pw[v0]) + lb * ld3(it.pw[v1]) + lc * ld3(it.pw[v2]);
  pb_out = pa_out - pwv;
}
}  // namespace coop
}  // namespace tc
