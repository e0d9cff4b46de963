// tc_internal.hpp — device-side working set of one batched evaluation and the kernel launchers.
// Layout of the arenas in HBM (all sized from device-side counts, grown geometrically, never shrunk):
//
//   q            [B][T][n_dof] f64                       joint values of every trajectory
//   term_*       [n_terms]                               cnt / #states / #units / LVS-merge flag per term
//   state_base,unit_base [n_terms+1]                     exclusive scans (sub-states and tested units)
//   poses        [n_active_links][n_states][12] f64      one 96-byte record per (link, state); consecutive sub-states
//                                                        of a link are adjacent, so a sweep's T0/T1 poses are one
//                                                        192-byte read and the FK warps write whole 3 KB runs
//   aabb         [n_active_shapes][6][n_states] f64      SoA: the broadphase streams them along the state axis
//   unit_term    [n_units] i32
//   bp_mask      [partner words][n_units] u32            broadphase survivors as bits (count pass -> fill pass)
//   sp_count/sp_offset/sp_cursor [n_shape_pairs(+1)]     survivors per shape pair, offsets padded to CHUNK
//   items        [n_items_padded] i32                    unit index, bucketed per shape pair, -1 = padding
//   item_state   [n_items_padded] i32                    T0 sub-state of the item
//   epa_rec      [n_items_padded] EpaRec (208 B)         GJK -> EPA hand-off, packed per shape pair
//   raw          [<= n_items] RawContact (96 B)          narrow-phase hits in arrival order
//   keys/idx     u64 / i32                               sort keys (term, link pair, substep, shape pair)
//   contacts     [n_contacts] tc_contact (200 B), sorted (their transforms are records of `poses`)
//   grad         [n_contacts][2][2][n_dof] f64, grec, ghas
#pragma once
#include <cuda_runtime.h>

#include <cstdint>
#include <string>
#include <vector>

#include "tc_contact.cuh"

namespace tc
{
#ifndef TC_CHUNK
#define TC_CHUNK 128
#endif
#ifndef TC_GJK_RUN
#define TC_GJK_RUN 512
#endif
constexpr int CHUNK = TC_CHUNK;  // bucket padding unit = threads per narrow-phase CTA
// items one GJK CTA works through (hulls are staged once per run of one shape pair)
constexpr int GJK_RUN_ITEMS = TC_GJK_RUN;
constexpr int MAX_FRAMES = 64;
#ifndef TC_EPA_VERTS
#define TC_EPA_VERTS 34
#endif
// thread-local polytope capacity (the live-face mask is 64 bits)
constexpr int EPA_FAST_VERTS = TC_EPA_VERTS, EPA_FAST_FACES = (2 * TC_EPA_VERTS - 4 > 64) ? 64 : 2 * TC_EPA_VERTS - 4;
constexpr int EPA_DEEP_VERTS = 4 + EPA_MAX_ITERS, EPA_DEEP_FACES = 8 + 2 * EPA_MAX_ITERS;

struct Counters
{
  long long n_states, n_units, n_items, n_items_padded, n_raw, n_valid, n_deep;
  long long gjk_iters, support_verts, n_epa, epa_iters, n_unconverged, overflow, nnz;
  long long gjk_cursor, epa_cursor_g;  // work cursors of the warp-persistent narrow-phase kernels
  long long support_compared;          // hull vertices actually compared by support queries (support_verts: all of the hull)
  long long n_curved;                  // items of curved-core shape pairs routed to the contraction-free kernel
  long long n_raw0;                    // raw hits present when the overflow re-runs start (they are finalised meanwhile)
  long long dbg[8];                    // TC_NARROW_HIST builds: 32 x (longest EPA / GJK run of every 32-item group), summed
};

template <class T>
struct DevBuf
{
  T* p{ nullptr };
  size_t cap{ 0 };
  // grows (never shrinks); contents are NOT preserved
  cudaError_t reserve(size_t n)
  {
    if (n <= cap)
      return cudaSuccess;
    if (p)
      cudaFree(p);
    p = nullptr;
    size_t want = n + n / 4 + 64;
    cudaError_t e = cudaMalloc(reinterpret_cast<void**>(&p), want * sizeof(T));
    cap = (e == cudaSuccess) ? want : 0;
    return e;
  }
  void release()
  {
    if (p)
      cudaFree(p);
    p = nullptr;
    cap = 0;
  }
};

// GJK -> EPA hand-off record (208 B)
struct EpaRec
{
  int32_t unit, state, nflag, pad;
  double s[24];  // simplex: 4 x (w, a)
};

// device pointers of one evaluation handed to the kernels by value
struct Work
{
  EvalParams ep;
  int cast;  // evaluator sweeps (CONTINUOUS / LVS_CONTINUOUS)
  int single_sweep;
  int poses_given;  // tc_contact_test: link poses come from the caller (pose_in), there are no joint values
  int contact_type; // TC_CONTACT_*
  const double* pose_in;  // [2 (T0, T1)][n_terms][n_active][12] (device) when poses_given
  const double* q;
  int n_dof_q;                 // joints per timestep of q
  const int32_t* term_t;       // [terms_per_traj]
  const uint8_t* term_fixed;   // [terms_per_traj][2]
  int32_t *term_cnt, *term_states, *term_units, *state_base, *unit_base;
  uint8_t* term_interp;
  double* poses;
  double* aabb;
  int32_t* unit_term;
  uint32_t* bp_mask;  // [partner words][n_units]: AABB survivors found by the counting pass, replayed by the fill pass
  int32_t *sp_count, *sp_offset, *sp_cursor;
  int32_t* items;
  RawContact* raw;
  int32_t* deep_items;   // (unit, shape pair) of items whose EPA outgrew the thread-local polytope
  int32_t* curved_items; // (unit, shape pair | DEEP_COUNT_GJK) of every item of a curved-core shape pair
  int32_t* item_state;   // T0 sub-state of every item (parallel to items)
  int32_t* epa_cursor;   // GJK -> EPA hand-off: records packed per shape pair at the bucket offsets
  EpaRec* epa_rec;
  unsigned long long *keys, *keys_alt;
  int32_t *idx, *idx_alt;
  uint8_t* raw_valid;  // [n_raw] the evaluator's filters passed (contact_type != ALL: applied after the manager's choice)
  tc_contact *contacts_tmp, *contacts;
  long long* term_ptr;  // [n_terms+1]
  double* grad;
  tc_gradient* grec;
  int32_t* ghas;
  long long* row_ptr;
  int32_t* row_block;    // [n_rows][2] first column of the T0 / T1 run of every row (-1 = block absent)
  int32_t* row_pair;     // [n_rows] link pair of every row (margin and coefficient follow from it)
  Counters* counters;
  // LAYOUT STRIDES / CAPACITIES of the arenas, fixed on the host before the kernels run.  In an exact pass they equal the
  // counts (read back stage by stage); in a speculative pass they are the previous pass's counts plus slack, and every
  // kernel takes the actual counts from `counters` on the device and leaves at once when `counters->overflow` says
  // a count outgrew its capacity (the host then repeats the pass exactly).
  long long n_states, n_units, n_items_padded;
  long long n_sort;      // entries the key sort covers (>= raw hits; the surplus carries all-ones keys)
  // host-known counts for the kernels launched AFTER the pass has been synchronised (accessors)
  long long n_raw, n_contacts;
  long long raw_cap, deep_cap;
  int speculative;
  int gather_in_rows;  // the fused gradient-row kernel performs the gather of the sorted contacts itself
  int rows_packed;     // rows hold only their structural non-zeros (the joints that move either link); row_ptr is
                       // then an exclusive scan of the row widths made before the row kernel runs
  long long* row_width;
};

// actual counts of the running pass, as the kernels see them
__device__ __forceinline__ bool passOverflowed(const Work& w) { return w.counters->overflow != 0; }
enum
{
  OVF_LVS_STATES = 2,   // a segment needs more than 16384 sub-states
  OVF_INDEX_STATES = 3, // 32-bit index arenas
  OVF_INDEX_ITEMS = 4,
  OVF_SPECULATION = 5   // a count outgrew the capacity predicted from the previous pass: repeat exactly
};

// record of (link, state) in the pose arena; the next sub-state of the same link follows at +12
__host__ __device__ inline const double* posePtr(const Work& w, int link, long long state)
{
  return w.poses + (static_cast<size_t>(link) * static_cast<size_t>(w.n_states) + static_cast<size_t>(state)) * 12;
}

// ---- launchers (tc_pipeline.cu) ----
// number of kernels of this library launched by the calling thread (every <<< >>> site bumps it)
int& launchCounter();
// fills / copies done by kernels (never by the copy engines, see tc_api.cu runEval)
cudaError_t launchFill(void* dst, uint32_t word, size_t n_words, cudaStream_t st);
cudaError_t launchCopy32(void* dst, const void* src, size_t n_words, cudaStream_t st);
cudaError_t launchCopy64(void* dst, const void* src, size_t n_dwords, cudaStream_t st);
cudaError_t launchSnapshot(const Counters* dev, Counters* mapped_host, cudaStream_t st);
cudaError_t launchPlan(const SceneView& v, const Work& w, cudaStream_t st);
cudaError_t launchFk(const SceneView& v, const Work& w, cudaStream_t st);  // tc_strict.cu
cudaError_t launchPoseIn(const SceneView& v, const Work& w, cudaStream_t st);  // tc_strict.cu
cudaError_t launchTypeFilter(const SceneView& v, const Work& w, const uint8_t* valid, cudaStream_t st);
cudaError_t launchBroadCount(const SceneView& v, const Work& w, cudaStream_t st);
cudaError_t launchBroadFill(const SceneView& v, const Work& w, cudaStream_t st);
// phase 0: every raw hit; 1: the first n_raw0 of them; 2: the rest and the surplus entries of the sort
cudaError_t launchFinalize(const SceneView& v, const Work& w, int phase, cudaStream_t st);
#ifdef TC_NARROW_HIST
void narrowScanDebug(unsigned long long out[4]);  // tuning builds: counters of the EPA plane scan since the last call
#endif
cudaError_t sortContacts(const Work& w, void* temp, size_t& temp_bytes, cudaStream_t st);  // result in *_alt
cudaError_t launchGather(const SceneView& v, const Work& w, cudaStream_t st);
cudaError_t launchGradients(const SceneView& v, const Work& w, cudaStream_t st);
cudaError_t launchTransforms(const SceneView& v, const Work& w, double* xf, cudaStream_t st);
cudaError_t launchRowBase(const SceneView& v, const Work& w, long long* term_row_base, cudaStream_t st);
// packed rows: widths from the sorted keys, exclusive scan into w.row_ptr (n_contacts + 1 entries); temp == nullptr
// only returns the scratch size
cudaError_t launchRowWidths(const SceneView& v, const Work& w, void* temp, size_t& temp_bytes, cudaStream_t st);
cudaError_t launchRows(const SceneView& v, const Work& w, long long* term_row_base, int32_t* col_idx, double* val,
                       double* row_value, double* row_coeff, double* row_margin, cudaStream_t st);
cudaError_t launchGradRows(const SceneView& v, const Work& w, long long* term_row_base, int32_t* col_idx, double* val,
                           double* row_value, double* row_coeff, double* row_margin, cudaStream_t st);
cudaError_t launchCost(const SceneView& v, const Work& w, double* cost_term, double* cost_traj, cudaStream_t st);
// record [cost, global index (int64 bits), trajectory T x n_dof] of the cheapest trajectory of the batch (first on ties)
cudaError_t launchBestRecord(const Work& w, const double* cost_traj, long long index_offset, double* rec, cudaStream_t st);
cudaError_t launchQpCsc(const SceneView& v, const Work& w, long long* bt_base, const double* val, const double* row_value,
                        const double* row_coeff, const double* row_margin, long long* col_ptr, int32_t* row_idx,
                        double* out, double* upper, double* slack_cost, cudaStream_t st);
cudaError_t launchSets(const SceneView& v, const Work& w, int max_num_cnt, tc_set* sets, int32_t* set_count,
                       int32_t* set_run, double* values, double* coeffs, double* jac, cudaStream_t st);
// ---- launchers (tc_narrow.cu) ----
cudaError_t launchNarrow(const SceneView& v, const Work& w, int precision, cudaStream_t st, cudaEvent_t mid);
// ---- launchers (tc_strict.cu: contraction-free kernels) ----
cudaError_t launchRouteCurved(const SceneView& v, const Work& w, cudaStream_t st);
// list: 0 = the EPA overflow list (deep_items, counter n_deep), 1 = the curved-core list (curved_items, n_curved).
// n < 0: the kernel takes the count from the device counters (speculative pass)
cudaError_t launchNarrowDeep(const SceneView& v, const Work& w, int precision, int list, long long n, void* scratch,
                             size_t scratch_bytes_per_thread, int n_threads, cudaStream_t st);
// n_raw0 = n_raw (speculative pass: the raw hits that exist before the overflow re-runs can be finalised meanwhile)
cudaError_t launchSnapshotRaw(const Work& w, cudaStream_t st);
size_t deepScratchBytesPerThread(int precision);
cudaError_t measureFmaPeak(int precision, double* tflops);
}  // namespace tc
