// TEST INFRASTRUCTURE ONLY — host-compiled view of the device numerics (tc_narrow.cuh, tc_contact.cuh).
// The per-thread functions of the kernels are `__host__ __device__`; this harness compiles them with g++
// so their logic can be checked against the oracle on a machine without a GPU (pytest -m "not gpu").
// It is never linked into libtrajopt_cuda.so and the product never loads it.
#include <cstring>
#include <string>
#include <vector>

#include "../../trajopt-1_b200/csrc/tc_contact.cuh"

using namespace tc;

static std::string g_err;

extern "C" {
const char* hc_last_error(void) { return g_err.c_str(); }

int hc_scene_create(const tc_scene_desc* d, void** out)
{
  try
  {
    auto* s = new HostScene();
    buildHostScene(*d, *s);
    *out = s;
    return TC_OK;
  }
  catch (const SceneError& e)
  {
    g_err = e.what();
    return e.code;
  }
}
void hc_scene_destroy(void* s) { delete static_cast<HostScene*>(s); }

// Support look-up table against the full vertex scan: for every hull that has a table, `n_dirs` directions
// (dirs[n_dirs][3], shape frame) are queried through Placed::support with and without the table.  Returns the number
// of hulls checked; *mismatches counts queries whose support points differ in any bit.
int hc_support_lut_check(void* scene, const double* dirs, int n_dirs, long long* mismatches, long long* cand_total,
                         long long* vert_total)
{
  const HostScene& s = *static_cast<HostScene*>(scene);
  int hulls = 0;
  *mismatches = 0, *cand_total = 0, *vert_total = 0;
  for (const ShapeD& sh : s.shapes)
  {
    if (!sh.lut_bytes)
      continue;
    ++hulls;
    Placed<double> full, fast;
    full.sh = makeCore<double>(sh, s.verts.data() + sh.vert_off);
    fast.sh = makeCore<double>(sh, s.verts.data() + sh.vert_off, s.lut.data() + sh.lut_off);
    const double ident[12] = { 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0 };
    full.p0 = full.p1 = fast.p0 = fast.p1 = Pose<double>::from12(ident);
    full.cast = fast.cast = false;
    const uint16_t* off = reinterpret_cast<const uint16_t*>(s.lut.data() + sh.lut_off);
    for (int k = 0; k < n_dirs; ++k)
    {
      const Vec3<double> d(dirs[3 * k], dirs[3 * k + 1], dirs[3 * k + 2]);
      ScanCount sc = 0;
      const Vec3<double> a = full.support(d, sc), b = fast.support(d, sc);
      if (std::memcmp(&a, &b, sizeof(a)) != 0)
        ++*mismatches;
      const int c = lutCellF(static_cast<float>(d.x), static_cast<float>(d.y), static_cast<float>(d.z));
      if (c >= 0)
        *cand_total += off[c + 1] - off[c];
      *vert_total += sh.vert_cnt;
    }
  }
  return hulls;
}
}  // extern "C"

template <class R>
static int shapeDistance(const HostScene& s, int sa, const double* a0, const double* a1, int sb, const double* b0,
                         const double* b1, double threshold, int max_verts, int max_faces, int* status,
                         double* distance, double* normal, double* pa, double* pb, int* used_epa, int* gjk_iters,
                         int* epa_iters, int* scanned)
{
  std::vector<Vtx<R>> verts(s.verts.size());
  for (size_t i = 0; i < verts.size(); ++i)
    verts[i] = Vtx<R>{ static_cast<R>(s.verts[i].x), static_cast<R>(s.verts[i].y), static_cast<R>(s.verts[i].z), 0 };
  const ShapeD& A = s.shapes[static_cast<size_t>(sa)];
  const ShapeD& B = s.shapes[static_cast<size_t>(sb)];
  Placed<R> PA, PB;
  // the support look-up tables are used exactly as the kernels use them
  PA.sh = makeCore<R>(A, verts.data() + A.vert_off, A.lut_bytes ? s.lut.data() + A.lut_off : nullptr);
  PB.sh = makeCore<R>(B, verts.data() + B.vert_off, B.lut_bytes ? s.lut.data() + B.lut_off : nullptr);
  PA.p0 = Pose<R>::from12(a0);
  PA.cast = a1 != nullptr;
  PA.p1 = a1 ? Pose<R>::from12(a1) : PA.p0;
  PB.p0 = Pose<R>::from12(b0);
  PB.cast = b1 != nullptr;
  PB.p1 = b1 ? Pose<R>::from12(b1) : PB.p0;
  std::vector<Vec3<R>> pw(static_cast<size_t>(max_verts)), pav(static_cast<size_t>(max_verts)),
      fn(static_cast<size_t>(max_faces));
  std::vector<R> fd(static_cast<size_t>(max_faces));
  std::vector<int> fv(static_cast<size_t>(max_faces)), edges(static_cast<size_t>(max_faces));
  std::vector<unsigned long long> erow(static_cast<size_t>(max_verts), 0ull);
  EpaStore<R> st{ pw.data(), pav.data(), fn.data(), fd.data(), fv.data(), edges.data(), max_verts, max_faces, erow.data() };
  // small stores exercise the register-mask fast path of the kernels, large ones the deep-expansion path
  const DistOut<R> o = (max_faces <= 64) ? signedDistance<R, true>(PA, static_cast<R>(A.brad), PB, static_cast<R>(B.brad),
                                                                   static_cast<R>(threshold), st) :
                                           signedDistance<R, false>(PA, static_cast<R>(A.brad), PB, static_cast<R>(B.brad),
                                                                    static_cast<R>(threshold), st);
  *status = o.status;
  *distance = o.distance;
  normal[0] = o.normal.x, normal[1] = o.normal.y, normal[2] = o.normal.z;
  pa[0] = o.pa.x, pa[1] = o.pa.y, pa[2] = o.pa.z;
  pb[0] = o.pb.x, pb[1] = o.pb.y, pb[2] = o.pb.z;
  *used_epa = o.used_epa;
  *gjk_iters = o.gjk_iters;
  *epa_iters = o.epa_iters;
  *scanned = static_cast<int>(o.scanned & 0xffffffffull);
  return TC_OK;
}

extern "C" {
int hc_shape_distance(const void* scene, int precision, int sa, const double* a0, const double* a1, int sb,
                      const double* b0, const double* b1, double threshold, int max_verts, int max_faces, int* status,
                      double* distance, double* normal, double* pa, double* pb, int* used_epa, int* gjk_iters,
                      int* epa_iters, int* scanned)
{
  const HostScene& s = *static_cast<const HostScene*>(scene);
  if (precision == TC_PRECISION_F32)
    return shapeDistance<float>(s, sa, a0, a1, sb, b0, b1, threshold, max_verts, max_faces, status, distance, normal,
                                pa, pb, used_epa, gjk_iters, epa_iters, scanned);
  return shapeDistance<double>(s, sa, a0, a1, sb, b0, b1, threshold, max_verts, max_faces, status, distance, normal,
                               pa, pb, used_epa, gjk_iters, epa_iters, scanned);
}

// the device's FK chain (frameStepQ of tc_contact.cuh with the shared sin/cos of tc_math.cuh), frame by frame as k_fk
// walks it: link_poses[n_active][12].  Must equal the oracle's FK bit for bit (tests/test_narrow_hostcheck.py).
int hc_fk(const void* scene, const double* q, double* link_poses)
{
  const HostScene& s = *static_cast<const HostScene*>(scene);
  const SceneView v = hostView(s);
  std::vector<Pose<double>> frame(static_cast<size_t>(s.n_frames));
  const Pose<double> base = Pose<double>::from12(v.base_pose);
  for (int f = 0; f < s.n_frames; ++f)
  {
    const int p = s.frame_parent[static_cast<size_t>(f)];
    double qf = 0;
    if (s.frame_type[static_cast<size_t>(f)] != TC_JOINT_FIXED)
      qf = q[s.frame_dof[static_cast<size_t>(f)]];
    frame[static_cast<size_t>(f)] = frameStepQ(v, p < 0 ? base : frame[static_cast<size_t>(p)], f, qf);
  }
  for (int l = 0; l < s.n_active; ++l)
    frame[static_cast<size_t>(s.link_frame[static_cast<size_t>(l)])].to12(link_poses + 12 * l);
  return TC_OK;
}

void hc_sincos(double x, double* sn, double* cs) { tc_sincos(x, sn, cs); }
}
