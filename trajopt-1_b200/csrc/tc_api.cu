// tc_api.cu — the C ABI of libtrajopt_cuda.so (include/trajopt_cuda.h): scene upload, orchestration of
// one batched pass on a CUDA stream, result marshalling.  No arithmetic of the path happens on the host;
// without a usable sm_100 device every entry point fails with TC_ERR_CUDA.
#include <dlfcn.h>

#include <cstdio>
#include <cstring>
#include <mutex>
#include <new>

#include "tc_internal.hpp"

using namespace tc;

namespace
{
thread_local std::string g_err;

int fail(int code, const std::string& msg)
{
  g_err = msg;
  return code;
}
int cudaFail(cudaError_t e, const char* what)
{
  g_err = std::string(what) + ": " + cudaGetErrorString(e);
  return TC_ERR_CUDA;
}
#define TC_CUDA(call)                                                                                                \
  do                                                                                                                 \
  {                                                                                                                  \
    cudaError_t e__ = (call);                                                                                        \
    if (e__ != cudaSuccess)                                                                                          \
      return cudaFail(e__, #call);                                                                                   \
  } while (0)

std::mutex g_init_mutex;
int g_device = -1;
// tc_cost_noalloc's static scratch: its device address is resolved once per device (defined at the end of this file)
int resolveNoallocScratch(int device, double** out);
}  // namespace

// counts of a finished pass, used as the capacities of the next one on the same result (speculative pass)
struct PassSizes
{
  long long n_states{ 0 }, n_units{ 0 }, n_items_padded{ 0 }, n_raw{ 0 }, n_deep{ 0 }, n_curved{ 0 };
};

struct tc_scene
{
  HostScene host;
  SceneView dview{};
  std::vector<void*> allocs;
  int device{ 0 };
};

struct tc_result
{
  tc_scene* scene{ nullptr };
  int device{ 0 };  // copied at creation: the destructor must not look at the scene, which may be gone by then
  cudaStream_t stream{ nullptr };
  tc_eval_desc desc{};
  EvalParams ep{};
  std::vector<int32_t> term_t;
  std::vector<uint8_t> term_fixed;
  bool pending{ false }, valid{ false };
  bool have_grad{ false }, have_cost{ false }, have_rows{ false };
  // speculative sizing (runEval): counts of the last finished pass and what it takes to repeat the current one
  PassSizes pred{};
  unsigned long long pred_sig{ 0 }, pass_sig{ 0 };
  bool pred_valid{ false }, speculative{ false }, deep_launched{ false };
  int respeculated{ 0 };
  std::vector<uint8_t> fixed_copy;
  const double *arg_q{ nullptr }, *arg_pose0{ nullptr }, *arg_pose1{ nullptr };
  int64_t pipe_row_base{ 0 };  // tc_pipeline_eval_rows: first global row of the sub-batch this result holds
  Work w{};
  Counters* h_counters{ nullptr };  // pinned + mapped: the device writes the snapshot itself
  Counters* h_counters_dev{ nullptr };  // device alias of h_counters
  int32_t* h_stage{ nullptr };          // pinned + mapped staging of the small per-eval tables
  int32_t* h_stage_dev{ nullptr };
  size_t h_stage_words{ 0 };
  DevBuf<int32_t> stage_d;
  Counters* d_counters{ nullptr };
  tc_counts counts{};
  cudaEvent_t ev[7]{};
  float stage_ms[10]{};
  int own_launches{ 0 }, lib_launches{ 0 };
  // arenas
  DevBuf<double> q, pose_in;
  DevBuf<uint8_t> raw_valid;
  DevBuf<int32_t> term_t_d, term_cnt, term_states, term_units, state_base, unit_base, unit_term;
  DevBuf<uint8_t> term_fixed_d, term_interp;
  DevBuf<double> poses, aabb;
  DevBuf<uint32_t> bp_mask;
  DevBuf<int32_t> sp_count, sp_offset, sp_cursor, items, deep_items, curved_items, idx, idx_alt, ghas, epa_cursor, item_state;
  DevBuf<EpaRec> epa_rec;
  DevBuf<RawContact> raw;
  DevBuf<unsigned long long> keys, keys_alt;
  DevBuf<tc_contact> contacts_tmp, contacts;
  DevBuf<double> xf, grad;
  DevBuf<long long> term_ptr, row_ptr, term_row_base, csc_bt_base, csc_col_ptr;
  DevBuf<int32_t> csc_row_idx;
  DevBuf<double> csc_val, csc_upper, csc_cost;
  bool have_csc{ false };
  DevBuf<tc_gradient> grec;
  DevBuf<unsigned char> sort_temp, deep_scratch, curved_scratch, scan_temp;
  DevBuf<long long> row_width;
  bool rows_packed{ false };  // set by tc_pipeline: rows in packed form (structural non-zeros only)
  cudaStream_t side{ nullptr };  // speculative passes: the cold one-thread-per-item kernels run beside the main stream
  cudaEvent_t ev_side[4]{};
  DevBuf<double> cost_term, cost_traj;
  // rows
  DevBuf<int32_t> col_idx, row_block, row_pair;
  DevBuf<double> row_val, row_value, row_coeff, row_margin;
  // sets
  DevBuf<tc_set> sets;
  DevBuf<int32_t> set_count, set_run;
  DevBuf<double> set_values, set_coeffs, set_jac;
};

namespace
{
template <class T>
int upload(tc_scene* s, const std::vector<T>& h, const T** out)
{
  void* d = nullptr;
  const size_t bytes = std::max<size_t>(h.size(), 1) * sizeof(T);
  TC_CUDA(cudaMalloc(&d, bytes));
  s->allocs.push_back(d);
  if (!h.empty())
    TC_CUDA(cudaMemcpy(d, h.data(), h.size() * sizeof(T), cudaMemcpyHostToDevice));
  *out = static_cast<const T*>(d);
  return TC_OK;
}

// adds the kernels this thread launches while the scope is alive to the result's own_launches
struct LaunchScope
{
  tc_result* r;
  int mark;
  explicit LaunchScope(tc_result* res) : r(res), mark(launchCounter()) {}
  ~LaunchScope() { r->own_launches += launchCounter() - mark; }
};

// TC_TRACE=1: host-side timestamps of the synchronisation points (diagnostics for stream overlap)
void trace(const tc_result* r, const char* what)
{
  static const bool on = getenv("TC_TRACE") != nullptr;
  if (!on)
    return;
  timespec ts;
  clock_gettime(CLOCK_MONOTONIC, &ts);
  fprintf(stderr, "[tc %p] %.3f ms %s\n", static_cast<const void*>(r), ts.tv_sec * 1e3 + ts.tv_nsec * 1e-6, what);
}

int ensureDevice(int device)
{
  TC_CUDA(cudaSetDevice(device));
  // every launch of this library is checked right after it is issued, so anything still in the runtime's last-error
  // slot here was left by other code of the process; it must not be blamed on our next launch
  (void)cudaGetLastError();
  return TC_OK;
}

int readCounters(tc_result* r)
{
  // snapshot through mapped host memory (SM stores over PCIe), not the D2H copy engine: a bulk result
  // download of another tc_result may own that engine for milliseconds and would stall this pass
  TC_CUDA(launchSnapshot(r->d_counters, r->h_counters_dev, r->stream));
  TC_CUDA(cudaStreamSynchronize(r->stream));
  trace(r, "counters read");
  if (r->h_counters->overflow == OVF_LVS_STATES)
    return fail(TC_ERR_OVERFLOW, "a segment needs more than 16384 LVS sub-states");
  if (r->h_counters->overflow == OVF_SPECULATION)
    return fail(TC_ERR_OVERFLOW, "a count outgrew the capacity predicted from the previous pass");  // finishEval repeats
  if (r->h_counters->overflow)
    return fail(TC_ERR_OVERFLOW, "an index arena exceeded 2^31 entries; split the batch");
  return TC_OK;
}

template <class T>
int reserve(DevBuf<T>& b, size_t n, const char* what)
{
  cudaError_t e = b.reserve(n);
  if (e != cudaSuccess)
    return cudaFail(e, what);
  return TC_OK;
}
#define TC_RESERVE(buf, n)                                                                                           \
  do                                                                                                                 \
  {                                                                                                                  \
    int s__ = reserve(buf, static_cast<size_t>(n), #buf);                                                            \
    if (s__ != TC_OK)                                                                                                \
      return s__;                                                                                                    \
  } while (0)
#define TC_TRY(call)                                                                                                 \
  do                                                                                                                 \
  {                                                                                                                  \
    int s__ = (call);                                                                                                \
    if (s__ != TC_OK)                                                                                                \
      return s__;                                                                                                    \
  } while (0)

int enqueueRows(tc_result* r, bool fused);

constexpr int DEEP_COUNT_GJK_FLAG = 1 << 30;  // tc_narrow_kernels.cuh DEEP_COUNT_GJK
inline long long withSlack(long long n) { return n + n / 8 + 256; }

// what makes two passes comparable in size: same batch shape, evaluator, options and fixed steps
unsigned long long passSignature(const tc_eval_desc& d, int B, int T, bool poses)
{
  unsigned long long h = 1469598103934665603ull;
  auto mix = [&h](unsigned long long x) {
    h ^= x;
    h *= 1099511628211ull;
  };
  mix(static_cast<unsigned>(B)), mix(static_cast<unsigned>(T)), mix(static_cast<unsigned>(d.evaluator));
  mix(static_cast<unsigned>(d.stack)), mix(static_cast<unsigned>(d.precision)), mix(static_cast<unsigned>(d.want));
  mix(static_cast<unsigned>(d.contact_test_type)), mix(poses ? 1 : 0);
  unsigned long long lv;
  std::memcpy(&lv, &d.lvs_length, sizeof(lv));
  mix(lv);
  if (d.fixed_steps)
    for (int t = 0; t < T; ++t)
      mix(d.fixed_steps[t] ? 2 : 3);
  return h;
}

// pose0 / pose1 != nullptr: tc_contact_test — link poses come from the caller ([B][n_active][12] host arrays, pose1 only
// for a cast test), there are no joint values (q == nullptr), B = number of states, T = 1 or 2.
//
// Two ways through the pass:
//   exact        every arena is sized from the device's own counts, read back stage by stage (five stream
//                synchronisations); always used for the first pass of a result and whenever the pass signature changes
//   speculative  arenas and launch grids are sized from the previous pass on this result plus 1/8 slack, the kernels take
//                the actual counts from the device counters, nothing is read back until tc_result_sync: the call only
//                enqueues.  A count that outgrows its capacity raises OVF_SPECULATION on the device, every later kernel
//                leaves at once, and tc_result_sync repeats the pass exactly (counts between SQP iterations move by a
//                few per cent, trajopt_sco/src/optimizers.cpp:194-203 evaluates the same problem at nearby iterates).
int runEval(tc_scene* scene, const tc_eval_desc* desc, const double* q, int32_t B, int32_t T, tc_result* r,
            cudaStream_t st, const double* pose0 = nullptr, const double* pose1 = nullptr, bool allow_speculative = true)
{
  if (!scene || !desc || (!q && !pose0) || !r)
    return fail(TC_ERR_INVALID, "null argument");
  if (r->scene != scene)
    return fail(TC_ERR_INVALID, "result object belongs to another scene");
  if (desc->evaluator < TC_EVAL_DISCRETE || desc->evaluator > TC_EVAL_LVS_CONTINUOUS)
    return fail(TC_ERR_INVALID, "invalid evaluator type");
  if (desc->stack != TC_STACK_IFOPT && desc->stack != TC_STACK_LEGACY)
    return fail(TC_ERR_INVALID, "invalid stack");
  if (desc->precision != TC_PRECISION_F64 && desc->precision != TC_PRECISION_F32)
    return fail(TC_ERR_INVALID, "invalid precision");
  if (desc->contact_test_type < TC_CONTACT_ALL || desc->contact_test_type > TC_CONTACT_FIRST)
    return fail(TC_ERR_INVALID, "invalid contact_test_type");
  if (B <= 0 || T <= 0)
    return fail(TC_ERR_INVALID, "empty batch");
  if (desc->evaluator != TC_EVAL_DISCRETE && T < 2)
    return fail(TC_ERR_INVALID, "segment evaluators need at least two timesteps");
  if (desc->evaluator != TC_EVAL_DISCRETE && !(desc->lvs_length > 0))
    return fail(TC_ERR_INVALID, "longest_valid_segment_length must be positive");
  TC_TRY(ensureDevice(scene->device));
  if (r->pending)
    cudaStreamSynchronize(r->stream);  // an evaluation that was never waited for: its arenas are about to be reused
  const HostScene& hs = scene->host;
  const SceneView& v = scene->dview;
  r->valid = false;
  r->pending = false;
  r->have_grad = r->have_cost = r->have_rows = false;
  r->have_csc = false;
  r->stream = st;
  // everything a repeat of this pass needs (speculation overflow)
  if (desc->fixed_steps && desc->fixed_steps != r->fixed_copy.data())
    r->fixed_copy.assign(desc->fixed_steps, desc->fixed_steps + T);
  else if (!desc->fixed_steps)
    r->fixed_copy.clear();
  r->desc = *desc;
  r->desc.fixed_steps = r->fixed_copy.empty() ? nullptr : r->fixed_copy.data();
  r->arg_q = q, r->arg_pose0 = pose0, r->arg_pose1 = pose1;
  // term list per trajectory: CollisionTermInfo::hatch (trajopt/src/problem_description.cpp:1735-1858)
  r->term_t.clear();
  r->term_fixed.clear();
  std::vector<uint8_t> fixed(static_cast<size_t>(T), 0);
  if (!r->fixed_copy.empty())
    fixed = r->fixed_copy;
  if (desc->evaluator == TC_EVAL_DISCRETE)
  {
    for (int t = 0; t < T; ++t)
      if (!fixed[static_cast<size_t>(t)])
      {
        r->term_t.push_back(t);
        r->term_fixed.push_back(0);
        r->term_fixed.push_back(0);
      }
  }
  else
  {
    for (int t = 0; t + 1 < T; ++t)
    {
      const bool cur = fixed[static_cast<size_t>(t)] != 0, nxt = fixed[static_cast<size_t>(t + 1)] != 0;
      if (cur && nxt)
        continue;
      r->term_t.push_back(t);
      r->term_fixed.push_back(cur ? 1 : 0);
      r->term_fixed.push_back((!cur && nxt) ? 1 : 0);
    }
  }
  EvalParams& ep = r->ep;
  ep.evaluator = desc->evaluator;
  ep.stack = desc->stack;
  ep.B = B;
  ep.T = T;
  ep.terms_per_traj = static_cast<int>(r->term_t.size());
  const long long n_terms = static_cast<long long>(B) * ep.terms_per_traj;
  if (n_terms >= (1ll << 28))
    return fail(TC_ERR_OVERFLOW, "more than 2^28 collision terms in one batch");
  ep.n_terms = static_cast<int>(n_terms);
  ep.lvs = desc->lvs_length;
  std::memset(&r->counts, 0, sizeof(r->counts));
  r->counts.n_terms = n_terms;
  static const bool no_spec = getenv("TC_NO_SPECULATION") != nullptr;  // diagnostics: every pass exact
  const unsigned long long sig = passSignature(r->desc, B, T, pose0 != nullptr);
  const bool spec = allow_speculative && !no_spec && r->pred_valid && r->pred_sig == sig && n_terms > 0;
  r->pass_sig = sig;
  r->speculative = spec;
  r->deep_launched = false;
  const PassSizes& pred = r->pred;
  Work& w = r->w;
  std::memset(&w, 0, sizeof(w));
  w.ep = ep;
  w.cast = (desc->evaluator == TC_EVAL_CONTINUOUS || desc->evaluator == TC_EVAL_LVS_CONTINUOUS) ? 1 : 0;
  w.single_sweep = (desc->evaluator == TC_EVAL_CONTINUOUS && desc->stack == TC_STACK_IFOPT) ? 1 : 0;
  w.poses_given = pose0 ? 1 : 0;
  w.contact_type = desc->contact_test_type;
  w.speculative = spec ? 1 : 0;
  w.n_dof_q = hs.n_dof;
  w.counters = r->d_counters;
  for (int i = 0; i < 10; ++i)
    r->stage_ms[i] = 0;
  r->own_launches = r->lib_launches = 0;
  LaunchScope scope(r);
  trace(r, spec ? "eval begin (speculative)" : "eval begin (exact)");
  if (pose0 && (desc->want & (TC_WANT_GRADIENTS | TC_WANT_ROWS)))
    return fail(TC_ERR_INVALID, "gradients and rows need joint values");
  // The first operation on the stream is a kernel, not the event record: when the stream's previous operation
  // was a device-to-host copy (a result download), an event record is queued on the copy engine, behind any bulk
  // download another tc_result has in flight there, and every kernel of this pass would wait for it.
  TC_CUDA(launchFill(r->d_counters, 0, sizeof(Counters) / 4, st));
  TC_CUDA(cudaEventRecord(r->ev[0], st));
  if (n_terms == 0)
  {
    TC_RESERVE(r->term_ptr, 1);
    TC_CUDA(launchFill(r->term_ptr.p, 0, sizeof(long long) / 4, st));
    w.term_ptr = r->term_ptr.p;
    for (int i = 1; i < 7; ++i)
      TC_CUDA(cudaEventRecord(r->ev[i], st));
    r->pending = true;
    return TC_OK;
  }
  // ---- inputs ----
  const size_t nq = static_cast<size_t>(B) * T * hs.n_dof;
  // Host inputs reach the device through kernels that read mapped pinned memory, not through the copy
  // engines: a bulk result download of another tc_result may occupy an engine for milliseconds, and an
  // input copy queued behind it would stall this whole pass.
  if (pose0)
  {
    const size_t np = static_cast<size_t>(B) * hs.n_active * 12;
    TC_RESERVE(r->pose_in, 2 * np);
    TC_CUDA(cudaMemcpyAsync(r->pose_in.p, pose0, np * sizeof(double), cudaMemcpyHostToDevice, st));
    if (pose1)
      TC_CUDA(cudaMemcpyAsync(r->pose_in.p + np, pose1, np * sizeof(double), cudaMemcpyHostToDevice, st));
    w.pose_in = r->pose_in.p;
    w.q = nullptr;
  }
  else if (desc->q_on_device)
    w.q = q;
  else
  {
    TC_RESERVE(r->q, nq);
    cudaPointerAttributes attr{};
    const bool pinned = cudaPointerGetAttributes(&attr, q) == cudaSuccess && attr.type == cudaMemoryTypeHost &&
                        attr.devicePointer != nullptr;
    if (pinned)
      TC_CUDA(launchCopy64(r->q.p, attr.devicePointer, nq, st));
    else
    {
      cudaGetLastError();  // pageable memory: cudaPointerGetAttributes may have set an error state
      TC_CUDA(cudaMemcpyAsync(r->q.p, q, nq * sizeof(double), cudaMemcpyHostToDevice, st));
    }
    w.q = r->q.p;
  }
  {
    const size_t n = static_cast<size_t>(ep.terms_per_traj), words = n + (2 * n + 3) / 4;
    if (words > r->h_stage_words)
    {
      if (r->h_stage)
        cudaFreeHost(r->h_stage);
      r->h_stage = nullptr;
      r->h_stage_words = 0;
      TC_CUDA(cudaHostAlloc(reinterpret_cast<void**>(&r->h_stage), (words + 256) * 4, cudaHostAllocMapped));
      TC_CUDA(cudaHostGetDevicePointer(reinterpret_cast<void**>(&r->h_stage_dev), r->h_stage, 0));
      r->h_stage_words = words + 256;
    }
    std::memcpy(r->h_stage, r->term_t.data(), n * 4);
    r->h_stage[words - 1] = 0;
    std::memcpy(r->h_stage + n, r->term_fixed.data(), 2 * n);
    TC_RESERVE(r->stage_d, words);
    TC_CUDA(launchCopy32(r->stage_d.p, r->h_stage_dev, words, st));
    w.term_t = r->stage_d.p;
    w.term_fixed = reinterpret_cast<const uint8_t*>(r->stage_d.p + n);
  }
  // ---- K0 plan ----
  TC_RESERVE(r->term_cnt, n_terms);
  TC_RESERVE(r->term_states, n_terms);
  TC_RESERVE(r->term_units, n_terms);
  TC_RESERVE(r->term_interp, n_terms);
  TC_RESERVE(r->state_base, n_terms + 1);
  TC_RESERVE(r->unit_base, n_terms + 1);
  w.term_cnt = r->term_cnt.p, w.term_states = r->term_states.p, w.term_units = r->term_units.p;
  w.term_interp = r->term_interp.p, w.state_base = r->state_base.p, w.unit_base = r->unit_base.p;
  if (spec)
  {
    w.n_states = withSlack(pred.n_states);
    w.n_units = withSlack(pred.n_units);
  }
  TC_CUDA(launchPlan(v, w, st));
  if (!spec)
  {
    TC_TRY(readCounters(r));
    w.n_states = r->h_counters->n_states;
    w.n_units = r->h_counters->n_units;
  }
  // ---- K1 FK, K2 broadphase count ----
  TC_RESERVE(r->poses, static_cast<size_t>(w.n_states) * hs.n_active * 12);
  TC_RESERVE(r->aabb, static_cast<size_t>(w.n_states) * hs.n_active_shapes * 6);
  TC_RESERVE(r->unit_term, w.n_units);
  TC_RESERVE(r->bp_mask, static_cast<size_t>(hs.partner_word_ptr.back()) * static_cast<size_t>(w.n_units));
  w.bp_mask = r->bp_mask.p;
  TC_RESERVE(r->sp_count, hs.spairs.size());
  TC_RESERVE(r->sp_cursor, hs.spairs.size());
  TC_RESERVE(r->sp_offset, hs.spairs.size() + 1);
  w.poses = r->poses.p, w.aabb = r->aabb.p, w.unit_term = r->unit_term.p;
  w.sp_count = r->sp_count.p, w.sp_cursor = r->sp_cursor.p, w.sp_offset = r->sp_offset.p;
  if (pose0)
    TC_CUDA(launchPoseIn(v, w, st));
  else
    TC_CUDA(launchFk(v, w, st));
  TC_CUDA(cudaEventRecord(r->ev[1], st));
  if (spec)
    w.n_items_padded = withSlack(pred.n_items_padded) / CHUNK * CHUNK;
  TC_CUDA(launchBroadCount(v, w, st));
  long long n_items = 0;
  if (!spec)
  {
    TC_TRY(readCounters(r));
    w.n_items_padded = r->h_counters->n_items_padded;
    n_items = r->h_counters->n_items;
  }
  else
    n_items = w.n_items_padded;  // capacity of the raw-hit and deep lists
  // ---- K2 fill, K3 narrow phase ----
  TC_RESERVE(r->items, w.n_items_padded);
  TC_RESERVE(r->raw, n_items);
  TC_RESERVE(r->deep_items, 2 * n_items);
  TC_RESERVE(r->epa_cursor, hs.spairs.size());
  TC_RESERVE(r->item_state, w.n_items_padded);
  TC_RESERVE(r->epa_rec, w.n_items_padded);
  w.items = r->items.p, w.raw = r->raw.p, w.deep_items = r->deep_items.p;
  w.epa_cursor = r->epa_cursor.p, w.item_state = r->item_state.p, w.epa_rec = r->epa_rec.p;
  TC_RESERVE(r->curved_items, hs.curved_sp.empty() ? 2 : 2 * n_items);
  w.curved_items = r->curved_items.p;
  // The two uses of the one-thread-per-item kernel of tc_strict.cu — the items of curved-core shape pairs and the few
  // items whose polytope outgrew k_epa_wp's store — are a handful of warps running ~10^4 instructions that are in no
  // cache when the launch comes: ~1 ms of instruction fetch for microseconds of arithmetic.  A speculative pass
  // therefore runs them on the result's SIDE stream, the curved items next to k_gjk / k_epa_wp and the overflow
  // re-runs next to the finalisation of the hits that already exist (k_finalize phases 1 and 2); an exact pass, which
  // reads the counts back anyway, runs them in line.
  const size_t deep_per = deepScratchBytesPerThread(TC_PRECISION_F64);  // the re-runs are FP64 in either mode
  auto deepThreads = [](long long n) { return 64 * static_cast<int>(std::min<long long>((std::max<long long>(n, 1) + 63) / 64, 296)); };
  const bool curved_scene = !hs.curved_sp.empty();
  bool side_curved = false;
  if (w.n_items_padded > 0)
  {
    TC_CUDA(launchBroadFill(v, w, st));
    TC_CUDA(cudaEventRecord(r->ev[5], st));
    if (curved_scene && spec)
    {
      const int nt = deepThreads(std::max<long long>(withSlack(pred.n_curved), 256));
      TC_RESERVE(r->curved_scratch, deep_per * nt);
      TC_CUDA(cudaEventRecord(r->ev_side[0], st));
      TC_CUDA(cudaStreamWaitEvent(r->side, r->ev_side[0], 0));
      TC_CUDA(launchRouteCurved(v, w, r->side));
      TC_CUDA(launchNarrowDeep(v, w, desc->precision, 1, -1, r->curved_scratch.p, deep_per, nt, r->side));
      TC_CUDA(cudaEventRecord(r->ev_side[1], r->side));
      side_curved = true;
    }
    TC_CUDA(launchNarrow(v, w, desc->precision, st, r->ev[6]));
    if (curved_scene && !spec)
      TC_CUDA(launchRouteCurved(v, w, st));
  }
  else
  {
    TC_CUDA(cudaEventRecord(r->ev[5], st));
    TC_CUDA(cudaEventRecord(r->ev[6], st));
  }
  TC_CUDA(cudaEventRecord(r->ev[2], st));
  bool split_finalize = false;
  if (!spec)
  {
    TC_TRY(readCounters(r));
    const long long n_deep = r->h_counters->n_deep, n_curved = r->h_counters->n_curved;
    static const char* trace_deep = getenv("TC_TRACE_DEEP");  // tuning aid: how many items outgrew the fast polytope
    if (trace_deep)
      fprintf(stderr, "[tc] n_epa=%lld n_deep=%lld n_curved=%lld epa_iters=%lld gjk_iters=%lld n_items=%lld group_max_epa=%lld group_max_gjk=%lld\n",
              r->h_counters->n_epa, n_deep, n_curved, r->h_counters->epa_iters, r->h_counters->gjk_iters,
              r->h_counters->n_items, r->h_counters->dbg[0], r->h_counters->dbg[1]);
#ifdef TC_NARROW_HIST
    if (trace_deep)
    {
      unsigned long long sd[4];
      cudaDeviceSynchronize();
      narrowScanDebug(sd);
      fprintf(stderr, "[tc] epa plane scan: faces tested %llu, skippable by |x| < d %llu, visible %llu\n", sd[0], sd[1], sd[2]);
    }
#endif
    if (trace_deep)
      fprintf(stderr, "[tc] epa items with >= 11/15/19/23/27/31 trips: %lld %lld %lld %lld %lld %lld\n", r->h_counters->dbg[2],
              r->h_counters->dbg[3], r->h_counters->dbg[4], r->h_counters->dbg[5], r->h_counters->dbg[6], r->h_counters->dbg[7]);
    if (n_curved > 0)
    {
      const int nt = deepThreads(n_curved);
      TC_RESERVE(r->curved_scratch, deep_per * nt);
      TC_CUDA(launchNarrowDeep(v, w, desc->precision, 1, n_curved, r->curved_scratch.p, deep_per, nt, st));
    }
    if (n_deep > 0)
    {
      const int nt = deepThreads(n_deep);
      TC_RESERVE(r->deep_scratch, deep_per * nt);
      TC_CUDA(launchNarrowDeep(v, w, desc->precision, 0, n_deep, r->deep_scratch.p, deep_per, nt, st));
    }
    r->deep_launched = true;
    if (n_curved > 0 || n_deep > 0)
      TC_TRY(readCounters(r));
  }
  else
  {
    if (side_curved)
      TC_CUDA(cudaStreamWaitEvent(st, r->ev_side[1], 0));  // the curved items' hits are in before anything is finalised
    // always: the re-run kernel leaves after one counter read when no polytope overflowed, and a pass never has to be
    // repeated because the previous one happened to have none
    r->deep_launched = true;
    {
      const int nt = deepThreads(std::max<long long>(withSlack(pred.n_deep), 256));
      TC_RESERVE(r->deep_scratch, deep_per * nt);
      TC_CUDA(launchSnapshotRaw(w, st));
      TC_CUDA(cudaEventRecord(r->ev_side[2], st));
      TC_CUDA(cudaStreamWaitEvent(r->side, r->ev_side[2], 0));
      TC_CUDA(launchNarrowDeep(v, w, desc->precision, 0, -1, r->deep_scratch.p, deep_per, nt, r->side));
      TC_CUDA(cudaEventRecord(r->ev_side[3], r->side));
      split_finalize = true;
    }
  }
  TC_CUDA(cudaEventRecord(r->ev[3], st));
  // ---- contact finalisation + sort ----
  w.n_sort = spec ? std::min(withSlack(pred.n_raw), n_items) : r->h_counters->n_raw;
  w.n_raw = w.n_sort;
  TC_RESERVE(r->contacts_tmp, w.n_sort);
  TC_RESERVE(r->keys, w.n_sort);
  TC_RESERVE(r->keys_alt, w.n_sort);
  TC_RESERVE(r->idx, w.n_sort);
  TC_RESERVE(r->idx_alt, w.n_sort);
  w.contacts_tmp = r->contacts_tmp.p;
  w.keys = r->keys.p, w.keys_alt = r->keys_alt.p, w.idx = r->idx.p, w.idx_alt = r->idx_alt.p;
  if (w.contact_type != TC_CONTACT_ALL)
  {
    TC_RESERVE(r->raw_valid, w.n_sort);
    w.raw_valid = r->raw_valid.p;
  }
  if (w.n_sort > 0)
  {
    if (split_finalize)
    {
      TC_CUDA(launchFinalize(v, w, 1, st));
      TC_CUDA(cudaStreamWaitEvent(st, r->ev_side[3], 0));
      TC_CUDA(launchFinalize(v, w, 2, st));
    }
    else
      TC_CUDA(launchFinalize(v, w, 0, st));
    size_t temp_bytes = 0;
    TC_CUDA(sortContacts(w, nullptr, temp_bytes, st));
    TC_RESERVE(r->sort_temp, temp_bytes);
    TC_CUDA(sortContacts(w, r->sort_temp.p, temp_bytes, st));
    r->lib_launches += 1;  // cub::DeviceRadixSort (several kernels of the CUDA toolkit's library)
    if (w.contact_type != TC_CONTACT_ALL)
    {
      // the manager's CLOSEST / FIRST choice over the sorted hits, then the final order
      TC_CUDA(launchTypeFilter(v, w, r->raw_valid.p, st));
      TC_CUDA(sortContacts(w, r->sort_temp.p, temp_bytes, st));
      r->lib_launches += 1;
    }
    if (!spec)
      TC_TRY(readCounters(r));
  }
  else if (split_finalize)
    TC_CUDA(cudaStreamWaitEvent(st, r->ev_side[3], 0));
  // grid bound of the kernels indexed by contact (they take the count itself from the device counters)
  w.n_contacts = spec ? w.n_sort : r->h_counters->n_valid;
  TC_RESERVE(r->contacts, w.n_contacts);
  TC_RESERVE(r->term_ptr, n_terms + 1);
  w.contacts = r->contacts.p, w.term_ptr = r->term_ptr.p;
  // rows wanted and no stored gradients: k_grad_rows runs in this pass and takes over the gather
  w.gather_in_rows = ((desc->want & TC_WANT_ROWS) && !(desc->want & TC_WANT_GRADIENTS) && w.n_contacts > 0) ? 1 : 0;
  TC_CUDA(launchGather(v, w, st));
  // ---- K4 ----
  if (desc->want & TC_WANT_GRADIENTS)
  {
    TC_RESERVE(r->grad, static_cast<size_t>(w.n_contacts) * 4 * hs.n_dof);
    TC_RESERVE(r->grec, w.n_contacts);
    TC_RESERVE(r->ghas, w.n_contacts);
    w.grad = r->grad.p, w.grec = r->grec.p, w.ghas = r->ghas.p;
    TC_CUDA(launchGradients(v, w, st));
    r->have_grad = true;
  }
  if (desc->want & TC_WANT_ROWS)
    TC_TRY(enqueueRows(r, !r->have_grad));
  if (desc->want & TC_WANT_COST)
  {
    TC_RESERVE(r->cost_term, n_terms);
    TC_RESERVE(r->cost_traj, B);
    TC_CUDA(launchCost(v, w, r->cost_term.p, r->cost_traj.p, st));
    r->have_cost = true;
  }
  TC_CUDA(cudaEventRecord(r->ev[4], st));
  r->pending = true;
  return TC_OK;
}

int finishEval(tc_result* r)
{
  if (!r->pending)
    return r->valid ? TC_OK : fail(TC_ERR_INVALID, "no evaluation to wait for");
  TC_TRY(ensureDevice(r->device));
  if (r->ep.n_terms > 0)
  {
    LaunchScope scope(r);
    const int st = readCounters(r);
    const bool respec = r->speculative && (r->h_counters->overflow == OVF_SPECULATION ||
                                           (st == TC_OK && r->h_counters->n_deep > 0 && !r->deep_launched));
    if (respec)
    {
      // a count outgrew the capacity taken from the previous pass: repeat with exact sizes
      trace(r, "speculation overflow: exact repeat");
      r->pred_valid = false;
      r->pending = false;
      const tc_eval_desc d = r->desc;
      const int own = r->own_launches, lib = r->lib_launches;
      TC_TRY(runEval(r->scene, &d, r->arg_q, r->ep.B, r->ep.T, r, r->stream, r->arg_pose0, r->arg_pose1, false));
      r->own_launches += own, r->lib_launches += lib;
      r->respeculated += 1;
      return finishEval(r);
    }
    if (st != TC_OK)
      return st;
  }
  else
  {
    TC_CUDA(cudaStreamSynchronize(r->stream));
    std::memset(r->h_counters, 0, sizeof(Counters));
  }
  r->pending = false;
  const Counters& c = *r->h_counters;
  // the accessors' kernels are launched with the actual counts; n_states / n_units / n_items_padded stay what they were
  // during the pass: they are the strides of the arenas
  r->w.n_raw = c.n_raw;
  r->w.n_contacts = c.n_valid;
  r->w.speculative = 0;
  if (r->ep.n_terms > 0)
  {
    r->pred.n_states = c.n_states, r->pred.n_units = c.n_units, r->pred.n_items_padded = c.n_items_padded;
    r->pred.n_raw = c.n_raw, r->pred.n_deep = c.n_deep, r->pred.n_curved = c.n_curved;
    r->pred_sig = r->pass_sig;
    r->pred_valid = true;
  }
  const HostScene& hs = r->scene->host;
  tc_counts& o = r->counts;
  o.n_states = c.n_states;
  o.n_units = c.n_units;
  o.n_dense = c.n_units * static_cast<long long>(hs.pairs.size());
  o.n_candidates = c.n_units * static_cast<long long>(hs.spairs.size());
  o.n_narrowphase = c.n_items;
  o.n_contacts = c.n_valid;
  o.n_rows = c.n_valid;
  o.nnz = r->have_rows ? c.nnz : 0;
  o.gjk_iters = c.gjk_iters;
  o.support_verts = c.support_verts;
  o.support_compared = c.support_compared;
  o.n_epa = c.n_epa;
  o.n_unconverged = c.n_unconverged;
  o.overflow = c.overflow;
  o.epa_iters = c.epa_iters;
  for (int i = 0; i < 4; ++i)
    cudaEventElapsedTime(&r->stage_ms[i], r->ev[i], r->ev[i + 1]);
  cudaEventElapsedTime(&r->stage_ms[4], r->ev[0], r->ev[4]);
  cudaEventElapsedTime(&r->stage_ms[5], r->ev[5], r->ev[2]);
  cudaEventElapsedTime(&r->stage_ms[6], r->ev[1], r->ev[5]);
  cudaEventElapsedTime(&r->stage_ms[7], r->ev[0], r->ev[1]);
  cudaEventElapsedTime(&r->stage_ms[8], r->ev[5], r->ev[6]);
  cudaEventElapsedTime(&r->stage_ms[9], r->ev[6], r->ev[2]);
  r->valid = true;
  return TC_OK;
}

int needValid(const tc_result* r)
{
  if (!r)
    return fail(TC_ERR_INVALID, "null result");
  if (r->pending)
    return fail(TC_ERR_INVALID, "tc_result_sync() has not been called");
  if (!r->valid)
    return fail(TC_ERR_INVALID, "result holds no evaluation");
  return ensureDevice(r->device);
}

// row assembly on the result's stream; `fused` = gradients recomputed inside the row kernel (no gradient arena)
int enqueueRows(tc_result* r, bool fused)
{
  const HostScene& hs = r->scene->host;
  Work& w = r->w;
  const size_t nc = static_cast<size_t>(w.n_contacts);
  TC_RESERVE(r->row_ptr, nc + 1);
  TC_RESERVE(r->term_row_base, r->ep.n_terms + 1);
  TC_RESERVE(r->col_idx, 2 * nc * hs.n_dof);
  TC_RESERVE(r->row_block, 2 * nc);
  TC_RESERVE(r->row_pair, nc);
  TC_RESERVE(r->row_val, 2 * nc * hs.n_dof);
  TC_RESERVE(r->row_value, nc);
  TC_RESERVE(r->row_coeff, nc);
  TC_RESERVE(r->row_margin, nc);
  w.row_ptr = r->row_ptr.p;
  w.row_block = r->row_block.p;
  w.row_pair = r->row_pair.p;
  w.rows_packed = (r->rows_packed && fused) ? 1 : 0;
  if (w.rows_packed)
  {
    if (hs.n_dof > 64)
      return fail(TC_ERR_UNSUPPORTED, "packed rows: more than 64 joints");
    TC_RESERVE(r->row_width, nc + 1);
    w.row_width = r->row_width.p;
    size_t temp_bytes = 0;
    TC_CUDA(launchRowWidths(r->scene->dview, w, nullptr, temp_bytes, r->stream));
    TC_RESERVE(r->scan_temp, temp_bytes);
    TC_CUDA(launchRowWidths(r->scene->dview, w, r->scan_temp.p, temp_bytes, r->stream));
    r->lib_launches += 1;  // cub::DeviceScan
  }
  else
    TC_CUDA(launchRowBase(r->scene->dview, w, r->term_row_base.p, r->stream));
  if (fused)
    TC_CUDA(launchGradRows(r->scene->dview, w, r->term_row_base.p, r->col_idx.p, r->row_val.p, r->row_value.p,
                           r->row_coeff.p, r->row_margin.p, r->stream));
  else
    TC_CUDA(launchRows(r->scene->dview, w, r->term_row_base.p, r->col_idx.p, r->row_val.p, r->row_value.p,
                       r->row_coeff.p, r->row_margin.p, r->stream));
  r->have_rows = true;  // nnz arrives with the next counter snapshot (finishEval / buildRows)
  return TC_OK;
}

int buildRows(tc_result* r)
{
  if (r->have_rows)
    return TC_OK;
  if (!r->have_grad)
    return fail(TC_ERR_INVALID, "rows need an evaluation with TC_WANT_ROWS or TC_WANT_GRADIENTS");
  LaunchScope scope(r);
  TC_TRY(enqueueRows(r, false));
  TC_TRY(readCounters(r));
  r->counts.nnz = r->h_counters->nnz;
  return TC_OK;
}
}  // namespace

// ---- multi-GPU helper: NCCL is resolved at run time from the process (torch ships libnccl.so.2) ----
struct Id128
{
  char b[128];  // ncclUniqueId
};
namespace
{
struct NcclApi
{
  void* lib{ nullptr };
  int (*allGather)(const void*, void*, size_t, int, void*, cudaStream_t){ nullptr };
  int (*broadcast)(const void*, void*, size_t, int, int, void*, cudaStream_t){ nullptr };
  int (*getUniqueId)(void*){ nullptr };
  int (*commInitRank)(void**, int, Id128, int){ nullptr };
  int (*commDestroy)(void*){ nullptr };
  const char* (*getErrorString)(int){ nullptr };
};
NcclApi g_nccl;
int loadNccl()
{
  if (g_nccl.lib)
    return TC_OK;
  const char* names[] = { "libnccl.so.2", "libnccl.so" };
  for (const char* n : names)
  {
    g_nccl.lib = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
    if (g_nccl.lib)
      break;
  }
  if (!g_nccl.lib)
    return fail(TC_ERR_NCCL, "libnccl.so.2 not found in the process (import torch first or set LD_LIBRARY_PATH)");
  g_nccl.allGather = reinterpret_cast<decltype(g_nccl.allGather)>(dlsym(g_nccl.lib, "ncclAllGather"));
  g_nccl.broadcast = reinterpret_cast<decltype(g_nccl.broadcast)>(dlsym(g_nccl.lib, "ncclBroadcast"));
  g_nccl.getUniqueId = reinterpret_cast<decltype(g_nccl.getUniqueId)>(dlsym(g_nccl.lib, "ncclGetUniqueId"));
  g_nccl.commInitRank = reinterpret_cast<decltype(g_nccl.commInitRank)>(dlsym(g_nccl.lib, "ncclCommInitRank"));
  g_nccl.commDestroy = reinterpret_cast<decltype(g_nccl.commDestroy)>(dlsym(g_nccl.lib, "ncclCommDestroy"));
  g_nccl.getErrorString = reinterpret_cast<decltype(g_nccl.getErrorString)>(dlsym(g_nccl.lib, "ncclGetErrorString"));
  if (!g_nccl.allGather || !g_nccl.broadcast || !g_nccl.getUniqueId || !g_nccl.commInitRank || !g_nccl.commDestroy)
    return fail(TC_ERR_NCCL, "libnccl lacks the expected entry points");
  return TC_OK;
}
int ncclFail(int code, const char* what)
{
  g_err = std::string(what) + ": " + (g_nccl.getErrorString ? g_nccl.getErrorString(code) : "nccl error");
  return TC_ERR_NCCL;
}
}  // namespace

extern "C" {

int tc_abi_version(void) { return TC_ABI_VERSION; }
const char* tc_last_error(void) { return g_err.c_str(); }

int tc_init(int device_ordinal)
{
  std::lock_guard<std::mutex> lock(g_init_mutex);
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess || n == 0)
    return fail(TC_ERR_CUDA, std::string("no CUDA device: ") + (e != cudaSuccess ? cudaGetErrorString(e) : "count is 0"));
  if (device_ordinal < 0 || device_ordinal >= n)
    return fail(TC_ERR_INVALID, "device ordinal out of range");
  cudaDeviceProp prop;
  TC_CUDA(cudaGetDeviceProperties(&prop, device_ordinal));
  if (prop.major != 10)
    return fail(TC_ERR_CUDA, std::string("device is not sm_100 (B200): ") + prop.name);
  TC_CUDA(cudaSetDevice(device_ordinal));
  g_device = device_ordinal;
  // resolve tc_cost_noalloc's scratch address now, so that a later first call made under CUDA graph capture
  // does not have to (cudaGetSymbolAddress may initialise the module, which is not capturable)
  double* unused = nullptr;
  TC_TRY(resolveNoallocScratch(device_ordinal, &unused));
  return TC_OK;
}

int tc_scene_create(const tc_scene_desc* desc, tc_scene** out)
{
  if (!desc || !out)
    return fail(TC_ERR_INVALID, "null argument");
  if (g_device < 0)
    return fail(TC_ERR_CUDA, "tc_init() has not succeeded: no usable sm_100 device (there is no CPU fallback)");
  tc_scene* s = new (std::nothrow) tc_scene();
  if (!s)
    return fail(TC_ERR_INVALID, "out of memory");
  s->device = g_device;
  try
  {
    buildHostScene(*desc, s->host);
  }
  catch (const SceneError& e)
  {
    delete s;
    return fail(e.code, e.what());
  }
  const HostScene& h = s->host;
  if (h.n_frames > MAX_FRAMES)
  {
    delete s;
    return fail(TC_ERR_UNSUPPORTED, "kinematic tree with more than 64 frames");
  }
  int st = ensureDevice(s->device);
  SceneView v = hostView(h);
  std::vector<Vtx<float>> vf(h.verts.size());
  for (size_t i = 0; i < vf.size(); ++i)
    vf[i] = Vtx<float>{ static_cast<float>(h.verts[i].x), static_cast<float>(h.verts[i].y),
                        static_cast<float>(h.verts[i].z), 0.f };
#define UP(field, vec)                                                                                               \
  if (st == TC_OK)                                                                                                   \
  st = upload(s, vec, &v.field)
  UP(frame_parent, h.frame_parent);
  UP(frame_type, h.frame_type);
  UP(frame_dof, h.frame_dof);
  UP(frame_axis, h.frame_axis);
  UP(frame_origin, h.frame_origin);
  UP(path_ptr, h.path_ptr);
  UP(path_frames, h.path_frames);
  UP(link_frame, h.link_frame);
  UP(frame_link, h.frame_link);
  UP(link_dof_mask, h.link_dof_mask);
  UP(link_pose, h.link_pose);
  UP(shapes, h.shapes);
  UP(verts, h.verts);
  UP(verts_f32, vf);
  UP(lut, h.lut);
  UP(active_shapes, h.active_shapes);
  UP(active_shape_frame, h.active_shape_frame);
  UP(static_aabb, h.static_aabb);
  UP(pairs, h.pairs);
  UP(spairs, h.spairs);
  UP(partner_ptr, h.partner_ptr);
  UP(partner_sp, h.partner_sp);
  UP(partner_word_ptr, h.partner_word_ptr);
  UP(curved_sp, h.curved_sp);
#undef UP
  if (st != TC_OK)
  {
    tc_scene_destroy(s);
    return st;
  }
  s->dview = v;
  *out = s;
  return TC_OK;
}

int tc_scene_set_link_poses(tc_scene* s, const double* link_pose)
{
  if (!s || !link_pose)
    return fail(TC_ERR_INVALID, "null argument");
  TC_TRY(ensureDevice(s->device));
  updateStaticPoses(s->host, link_pose);
  // plain synchronous copies: the call is not allowed to overlap an evaluation on this scene
  TC_CUDA(cudaDeviceSynchronize());
  TC_CUDA(cudaMemcpy(const_cast<double*>(s->dview.link_pose), s->host.link_pose.data(),
                     s->host.link_pose.size() * sizeof(double), cudaMemcpyHostToDevice));
  TC_CUDA(cudaMemcpy(const_cast<double*>(s->dview.static_aabb), s->host.static_aabb.data(),
                     s->host.static_aabb.size() * sizeof(double), cudaMemcpyHostToDevice));
  return TC_OK;
}

void tc_scene_destroy(tc_scene* s)
{
  if (!s)
    return;
  cudaSetDevice(s->device);
  for (void* p : s->allocs)
    cudaFree(p);
  delete s;
  (void)cudaGetLastError();  // failures above are ignored on purpose and must not surface at the next launch
}

int tc_scene_num_pairs(const tc_scene* s) { return s ? static_cast<int>(s->host.pairs.size()) : 0; }

int tc_scene_pairs(const tc_scene* s, int32_t* l0, int32_t* l1, double* margin, double* coeff)
{
  if (!s)
    return fail(TC_ERR_INVALID, "null scene");
  for (size_t i = 0; i < s->host.pairs.size(); ++i)
  {
    if (l0)
      l0[i] = s->host.pairs[i].a;
    if (l1)
      l1[i] = s->host.pairs[i].b;
    if (margin)
      margin[i] = s->host.pairs[i].margin;
    if (coeff)
      coeff[i] = s->host.pairs[i].coeff;
  }
  return TC_OK;
}

int tc_result_create(tc_scene* scene, tc_result** out)
{
  if (!scene || !out)
    return fail(TC_ERR_INVALID, "null argument");
  TC_TRY(ensureDevice(scene->device));
  tc_result* r = new (std::nothrow) tc_result();
  if (!r)
    return fail(TC_ERR_INVALID, "out of memory");
  r->scene = scene;
  r->device = scene->device;
  cudaError_t e = cudaHostAlloc(reinterpret_cast<void**>(&r->h_counters), sizeof(Counters), cudaHostAllocMapped);
  if (e == cudaSuccess)
    e = cudaHostGetDevicePointer(reinterpret_cast<void**>(&r->h_counters_dev), r->h_counters, 0);
  if (e == cudaSuccess)
    e = cudaMalloc(reinterpret_cast<void**>(&r->d_counters), sizeof(Counters));
  for (int i = 0; i < 7 && e == cudaSuccess; ++i)
    e = cudaEventCreate(&r->ev[i]);
  if (e == cudaSuccess)
    e = cudaStreamCreateWithFlags(&r->side, cudaStreamNonBlocking);
  for (int i = 0; i < 4 && e == cudaSuccess; ++i)
    e = cudaEventCreateWithFlags(&r->ev_side[i], cudaEventDisableTiming);
  if (e != cudaSuccess)
  {
    tc_result_destroy(r);
    return cudaFail(e, "tc_result_create");
  }
  *out = r;
  return TC_OK;
}

void tc_result_destroy(tc_result* r)
{
  if (!r)
    return;
  // Handles are destroyed in whatever order the host's garbage collector picks: nothing here may touch the scene.
  // Failures of the calls below are deliberately ignored, and the runtime's last-error slot is cleared at the end so
  // that they are not blamed on the next kernel launch of this thread.
  cudaSetDevice(r->device);
  if (r->pending)
    cudaStreamSynchronize(r->stream);
  if (r->h_counters)
    cudaFreeHost(r->h_counters);
  if (r->h_stage)
    cudaFreeHost(r->h_stage);
  r->stage_d.release();
  if (r->d_counters)
    cudaFree(r->d_counters);
  for (int i = 0; i < 7; ++i)
    if (r->ev[i])
      cudaEventDestroy(r->ev[i]);
  for (int i = 0; i < 4; ++i)
    if (r->ev_side[i])
      cudaEventDestroy(r->ev_side[i]);
  if (r->side)
  {
    cudaStreamSynchronize(r->side);
    cudaStreamDestroy(r->side);
  }
  r->pose_in.release(), r->raw_valid.release();
  r->q.release(), r->term_t_d.release(), r->term_cnt.release(), r->term_states.release(), r->term_units.release();
  r->state_base.release(), r->unit_base.release(), r->unit_term.release(), r->term_fixed_d.release();
  r->term_interp.release(), r->poses.release(), r->aabb.release(), r->bp_mask.release(), r->sp_count.release(), r->sp_offset.release();
  r->curved_items.release(), r->curved_scratch.release();
  r->sp_cursor.release(), r->items.release(), r->deep_items.release(), r->idx.release(), r->idx_alt.release();
  r->ghas.release(), r->raw.release(), r->keys.release(), r->keys_alt.release(), r->contacts_tmp.release();
  r->contacts.release(), r->xf.release(), r->grad.release(), r->term_ptr.release();
  r->row_ptr.release(), r->term_row_base.release(), r->grec.release(), r->sort_temp.release();
  r->scan_temp.release(), r->row_width.release();
  r->deep_scratch.release(), r->cost_term.release(), r->cost_traj.release(), r->col_idx.release(), r->row_block.release(), r->row_pair.release();
  r->csc_bt_base.release(), r->csc_col_ptr.release(), r->csc_row_idx.release(), r->csc_val.release();
  r->csc_upper.release(), r->csc_cost.release();
  r->row_val.release(), r->row_value.release(), r->row_coeff.release(), r->row_margin.release();
  r->epa_cursor.release(), r->item_state.release(), r->epa_rec.release();
  r->sets.release(), r->set_count.release(), r->set_run.release(), r->set_values.release();
  r->set_coeffs.release(), r->set_jac.release();
  delete r;
  (void)cudaGetLastError();
}

int tc_eval_async(tc_scene* scene, const tc_eval_desc* desc, const double* q, int32_t B, int32_t T, tc_result* out,
                  void* cuda_stream)
{
  return runEval(scene, desc, q, B, T, out, static_cast<cudaStream_t>(cuda_stream));
}

int tc_result_sync(tc_result* res)
{
  if (!res)
    return fail(TC_ERR_INVALID, "null result");
  return finishEval(res);
}

int tc_result_pass_info(const tc_result* res, int32_t* speculative, int32_t* repeated)
{
  if (!res)
    return fail(TC_ERR_INVALID, "null result");
  if (speculative)
    *speculative = res->speculative ? 1 : 0;
  if (repeated)
    *repeated = res->respeculated;
  return TC_OK;
}

#ifdef TC_NARROW_HIST
// tuning builds only: headers (unit, state, nflag, gjk_iters | epa_iters << 16) of the hand-off records of the last
// pass, the buckets' offsets and the number of valid records per bucket
int tc_debug_epa_records(tc_result* r, int32_t* hdr, size_t cap, int32_t* sp_offset, int32_t* sp_valid, int32_t* n_sp)
{
  TC_TRY(ensureDevice(r->device));
  cudaDeviceSynchronize();
  const size_t n = static_cast<size_t>(r->w.n_items_padded);
  if (n > cap)
    return fail(TC_ERR_OVERFLOW, "cap");
  TC_CUDA(cudaMemcpy2D(hdr, 16, r->epa_rec.p, sizeof(EpaRec), 16, n, cudaMemcpyDeviceToHost));
  const int ns = r->scene->host.spairs.size();
  TC_CUDA(cudaMemcpy(sp_offset, r->sp_offset.p, sizeof(int32_t) * ns, cudaMemcpyDeviceToHost));
  TC_CUDA(cudaMemcpy(sp_valid, r->epa_cursor.p, sizeof(int32_t) * ns, cudaMemcpyDeviceToHost));
  *n_sp = ns;
  return TC_OK;
}
#endif

int tc_contact_test(tc_scene* scene, const double* link_pose0, const double* link_pose1, int32_t n_states,
                    int32_t contact_test_type, int32_t precision, tc_result* out, void* cuda_stream)
{
  if (!link_pose0)
    return fail(TC_ERR_INVALID, "null argument");
  tc_eval_desc d{};
  d.evaluator = link_pose1 ? TC_EVAL_CONTINUOUS : TC_EVAL_DISCRETE;
  d.stack = TC_STACK_IFOPT;  // CONTINUOUS = one sweep per state
  d.precision = precision;
  d.want = TC_WANT_CONTACTS | TC_WANT_COST;
  d.lvs_length = 1.0;  // unused: one sweep per state
  d.contact_test_type = contact_test_type;
  TC_TRY(runEval(scene, &d, nullptr, n_states, link_pose1 ? 2 : 1, out, static_cast<cudaStream_t>(cuda_stream), link_pose0,
                 link_pose1));
  return finishEval(out);
}

int tc_eval(tc_scene* scene, const tc_eval_desc* desc, const double* q, int32_t B, int32_t T, tc_result* out,
            void* cuda_stream)
{
  TC_TRY(runEval(scene, desc, q, B, T, out, static_cast<cudaStream_t>(cuda_stream)));
  return finishEval(out);
}

int tc_result_counts(const tc_result* res, tc_counts* counts)
{
  TC_TRY(needValid(res));
  *counts = res->counts;
  return TC_OK;
}

int tc_result_contacts(const tc_result* res, tc_contact* dst, size_t cap)
{
  TC_TRY(needValid(res));
  const size_t n = std::min<size_t>(cap, static_cast<size_t>(res->w.n_contacts));
  if (n)
    TC_CUDA(cudaMemcpyAsync(dst, res->w.contacts, n * sizeof(tc_contact), cudaMemcpyDeviceToHost, res->stream));
  TC_CUDA(cudaStreamSynchronize(res->stream));
  return TC_OK;
}

int tc_result_contact_transforms(const tc_result* res, double* dst, size_t cap)
{
  TC_TRY(needValid(res));
  tc_result* r = const_cast<tc_result*>(res);
  const size_t n = std::min<size_t>(cap, static_cast<size_t>(r->w.n_contacts));
  if (n)
  {
    TC_RESERVE(r->xf, static_cast<size_t>(r->w.n_contacts) * 48);
    TC_CUDA(launchTransforms(r->scene->dview, r->w, r->xf.p, r->stream));
    TC_CUDA(cudaStreamSynchronize(r->stream));
    TC_CUDA(cudaMemcpyAsync(dst, r->xf.p, n * 48 * sizeof(double), cudaMemcpyDeviceToHost, r->stream));
  }
  TC_CUDA(cudaStreamSynchronize(r->stream));
  return TC_OK;
}

int tc_result_term_ptr(const tc_result* res, int64_t* term_ptr)
{
  TC_TRY(needValid(res));
  static_assert(sizeof(long long) == sizeof(int64_t), "int64");
  TC_CUDA(cudaMemcpyAsync(term_ptr, res->w.term_ptr, (static_cast<size_t>(res->ep.n_terms) + 1) * sizeof(int64_t),
                      cudaMemcpyDeviceToHost, res->stream));
  TC_CUDA(cudaStreamSynchronize(res->stream));
  return TC_OK;
}

int tc_result_gradients(const tc_result* res, tc_gradient* rec, double* grad, size_t cap)
{
  TC_TRY(needValid(res));
  if (!res->have_grad)
    return fail(TC_ERR_INVALID, "evaluation ran without TC_WANT_GRADIENTS");
  const size_t n = std::min<size_t>(cap, static_cast<size_t>(res->w.n_contacts));
  if (n)
  {
    TC_CUDA(cudaMemcpyAsync(rec, res->w.grec, n * sizeof(tc_gradient), cudaMemcpyDeviceToHost, res->stream));
    TC_CUDA(cudaMemcpyAsync(grad, res->w.grad, n * 4 * res->scene->host.n_dof * sizeof(double), cudaMemcpyDeviceToHost, res->stream));
  }
  TC_CUDA(cudaStreamSynchronize(res->stream));
  return TC_OK;
}

static int rowsDownload(const tc_result* res_c, int64_t* row_ptr, int32_t* col_idx, double* val, double* row_value,
                        double* row_coeff, double* row_margin, bool wait, int32_t* row_first_col = nullptr)
{
  TC_TRY(needValid(res_c));
  tc_result* r = const_cast<tc_result*>(res_c);
  TC_TRY(buildRows(r));
  trace(r, "rows download begin");
  const size_t nc = static_cast<size_t>(r->w.n_contacts);
  const size_t nnz = static_cast<size_t>(r->counts.nnz);
  TC_CUDA(cudaMemcpyAsync(row_ptr, r->row_ptr.p, (nc + 1) * sizeof(int64_t), cudaMemcpyDeviceToHost, r->stream));
  if (nnz)
  {
    if (col_idx)
      TC_CUDA(cudaMemcpyAsync(col_idx, r->col_idx.p, nnz * sizeof(int32_t), cudaMemcpyDeviceToHost, r->stream));
    TC_CUDA(cudaMemcpyAsync(val, r->row_val.p, nnz * sizeof(double), cudaMemcpyDeviceToHost, r->stream));
  }
  if (nc)
  {
    TC_CUDA(cudaMemcpyAsync(row_value, r->row_value.p, nc * sizeof(double), cudaMemcpyDeviceToHost, r->stream));
    TC_CUDA(cudaMemcpyAsync(row_coeff, r->row_coeff.p, nc * sizeof(double), cudaMemcpyDeviceToHost, r->stream));
    TC_CUDA(cudaMemcpyAsync(row_margin, r->row_margin.p, nc * sizeof(double), cudaMemcpyDeviceToHost, r->stream));
    if (row_first_col)
      TC_CUDA(cudaMemcpyAsync(row_first_col, r->row_block.p, 2 * nc * sizeof(int32_t), cudaMemcpyDeviceToHost, r->stream));
  }
  trace(r, "rows download enqueued");
  if (wait)
  {
    TC_CUDA(cudaStreamSynchronize(r->stream));
    trace(r, "rows download done");
  }
  return TC_OK;
}

int tc_result_rows(const tc_result* res, int64_t* row_ptr, int32_t* col_idx, double* val, double* row_value,
                   double* row_coeff, double* row_margin)
{
  return rowsDownload(res, row_ptr, col_idx, val, row_value, row_coeff, row_margin, true);
}

int tc_result_rows_async(const tc_result* res, int64_t* row_ptr, int32_t* col_idx, double* val, double* row_value,
                         double* row_coeff, double* row_margin)
{
  return rowsDownload(res, row_ptr, col_idx, val, row_value, row_coeff, row_margin, false);
}

int tc_result_rows_blocks(const tc_result* res, int64_t* row_ptr, int32_t* row_first_col, double* val, double* row_value,
                          double* row_coeff, double* row_margin)
{
  if (!row_first_col)
    return fail(TC_ERR_INVALID, "null argument");
  return rowsDownload(res, row_ptr, nullptr, val, row_value, row_coeff, row_margin, true, row_first_col);
}

int tc_result_rows_blocks_async(const tc_result* res, int64_t* row_ptr, int32_t* row_first_col, double* val,
                                double* row_value, double* row_coeff, double* row_margin)
{
  if (!row_first_col)
    return fail(TC_ERR_INVALID, "null argument");
  return rowsDownload(res, row_ptr, nullptr, val, row_value, row_coeff, row_margin, false, row_first_col);
}

int tc_stream_create(void** stream)
{
  if (!stream)
    return fail(TC_ERR_INVALID, "null argument");
  cudaStream_t st = nullptr;
  TC_CUDA(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
  *stream = st;
  return TC_OK;
}

void tc_stream_destroy(void* stream)
{
  if (stream)
    cudaStreamDestroy(static_cast<cudaStream_t>(stream));
}

int tc_result_wait(const tc_result* res)
{
  TC_TRY(needValid(res));
  TC_CUDA(cudaStreamSynchronize(res->stream));
  trace(res, "wait done");
  return TC_OK;
}

int tc_result_rows_device(const tc_result* res_c, const int64_t** row_ptr, const int32_t** col_idx, const double** val,
                          const double** row_value, const double** row_coeff, const double** row_margin)
{
  TC_TRY(needValid(res_c));
  tc_result* r = const_cast<tc_result*>(res_c);
  TC_TRY(buildRows(r));
  if (row_ptr)
    *row_ptr = reinterpret_cast<const int64_t*>(r->row_ptr.p);
  if (col_idx)
    *col_idx = r->col_idx.p;
  if (val)
    *val = r->row_val.p;
  if (row_value)
    *row_value = r->row_value.p;
  if (row_coeff)
    *row_coeff = r->row_coeff.p;
  if (row_margin)
    *row_margin = r->row_margin.p;
  return TC_OK;
}

// CSC assembly on the result's stream (once per evaluation)
static int buildQpCsc(tc_result* r)
{
  TC_TRY(buildRows(r));
  if (r->have_csc)
    return TC_OK;
  LaunchScope scope(r);
  const size_t nc = static_cast<size_t>(r->w.n_contacts);
  const size_t n_bt = static_cast<size_t>(r->ep.B) * static_cast<size_t>(r->ep.T);
  const size_t n_cols = n_bt * static_cast<size_t>(r->scene->host.n_dof) + nc;
  const size_t nnz = static_cast<size_t>(r->counts.nnz) + nc;
  TC_RESERVE(r->csc_bt_base, n_bt + 1);
  TC_RESERVE(r->csc_col_ptr, n_cols + 1);
  TC_RESERVE(r->csc_row_idx, nnz);
  TC_RESERVE(r->csc_val, nnz);
  TC_RESERVE(r->csc_upper, nc);
  TC_RESERVE(r->csc_cost, nc);
  TC_CUDA(launchQpCsc(r->scene->dview, r->w, r->csc_bt_base.p, r->row_val.p, r->row_value.p, r->row_coeff.p,
                      r->row_margin.p, r->csc_col_ptr.p, r->csc_row_idx.p, r->csc_val.p, r->csc_upper.p, r->csc_cost.p,
                      r->stream));
  r->have_csc = true;
  return TC_OK;
}

int tc_result_qp_csc_device(const tc_result* res_c, const int64_t** col_ptr, const int32_t** row_idx, const double** val,
                            const double** upper, const double** slack_cost, int64_t* n_rows, int64_t* n_cols,
                            int64_t* nnz)
{
  TC_TRY(needValid(res_c));
  tc_result* r = const_cast<tc_result*>(res_c);
  TC_TRY(buildQpCsc(r));
  const int64_t nc = r->w.n_contacts;
  if (col_ptr)
    *col_ptr = reinterpret_cast<const int64_t*>(r->csc_col_ptr.p);
  if (row_idx)
    *row_idx = r->csc_row_idx.p;
  if (val)
    *val = r->csc_val.p;
  if (upper)
    *upper = r->csc_upper.p;
  if (slack_cost)
    *slack_cost = r->csc_cost.p;
  if (n_rows)
    *n_rows = nc;
  if (n_cols)
    *n_cols = static_cast<int64_t>(r->ep.B) * r->ep.T * r->scene->host.n_dof + nc;
  if (nnz)
    *nnz = r->counts.nnz + nc;
  return TC_OK;
}

int tc_result_qp_csc(const tc_result* res_c, int64_t* col_ptr, int32_t* row_idx, double* val, double* upper,
                     double* slack_cost)
{
  if (!col_ptr || !row_idx || !val || !upper || !slack_cost)
    return fail(TC_ERR_INVALID, "null argument");
  int64_t n_rows = 0, n_cols = 0, nnz = 0;
  TC_TRY(tc_result_qp_csc_device(res_c, nullptr, nullptr, nullptr, nullptr, nullptr, &n_rows, &n_cols, &nnz));
  tc_result* r = const_cast<tc_result*>(res_c);
  TC_CUDA(cudaMemcpyAsync(col_ptr, r->csc_col_ptr.p, static_cast<size_t>(n_cols + 1) * sizeof(int64_t),
                          cudaMemcpyDeviceToHost, r->stream));
  if (nnz)
  {
    TC_CUDA(cudaMemcpyAsync(row_idx, r->csc_row_idx.p, static_cast<size_t>(nnz) * sizeof(int32_t), cudaMemcpyDeviceToHost,
                            r->stream));
    TC_CUDA(cudaMemcpyAsync(val, r->csc_val.p, static_cast<size_t>(nnz) * sizeof(double), cudaMemcpyDeviceToHost,
                            r->stream));
  }
  if (n_rows)
  {
    TC_CUDA(cudaMemcpyAsync(upper, r->csc_upper.p, static_cast<size_t>(n_rows) * sizeof(double), cudaMemcpyDeviceToHost,
                            r->stream));
    TC_CUDA(cudaMemcpyAsync(slack_cost, r->csc_cost.p, static_cast<size_t>(n_rows) * sizeof(double),
                            cudaMemcpyDeviceToHost, r->stream));
  }
  TC_CUDA(cudaStreamSynchronize(r->stream));
  return TC_OK;
}

int tc_result_sets(const tc_result* res_c, int32_t max_num_cnt, int64_t* set_ptr, tc_set* sets, size_t cap_sets,
                   double* values, double* coeffs, double* jac)
{
  TC_TRY(needValid(res_c));
  tc_result* r = const_cast<tc_result*>(res_c);
  if (!r->have_grad)
    return fail(TC_ERR_INVALID, "sets need an evaluation with TC_WANT_GRADIENTS");
  if (max_num_cnt <= 0)
    return fail(TC_ERR_INVALID, "max_num_cnt must be positive");
  const HostScene& hs = r->scene->host;
  const size_t nc = static_cast<size_t>(r->w.n_contacts), nt = static_cast<size_t>(r->ep.n_terms);
  const size_t n = static_cast<size_t>(hs.n_dof), m = static_cast<size_t>(max_num_cnt);
  set_ptr[0] = 0;
  if (nt == 0)
    return TC_OK;
  TC_RESERVE(r->sets, nc);
  TC_RESERVE(r->set_run, 2 * nc);
  TC_RESERVE(r->set_count, nt);
  TC_RESERVE(r->set_values, nt * m);
  TC_RESERVE(r->set_coeffs, nt * m);
  TC_RESERVE(r->set_jac, nt * m * 2 * n);
  TC_CUDA(launchSets(r->scene->dview, r->w, max_num_cnt, r->sets.p, r->set_count.p, r->set_run.p, r->set_values.p,
                     r->set_coeffs.p, r->set_jac.p, r->stream));
  TC_CUDA(cudaStreamSynchronize(r->stream));
  TC_CUDA(cudaMemcpyAsync(values, r->set_values.p, nt * m * sizeof(double), cudaMemcpyDeviceToHost, r->stream));
  TC_CUDA(cudaMemcpyAsync(coeffs, r->set_coeffs.p, nt * m * sizeof(double), cudaMemcpyDeviceToHost, r->stream));
  TC_CUDA(cudaMemcpyAsync(jac, r->set_jac.p, nt * m * 2 * n * sizeof(double), cudaMemcpyDeviceToHost, r->stream));
  // marshalling only: the sets of term k sit at the term's contact offset; pack them back to back
  std::vector<int32_t> cnt(nt);
  std::vector<int64_t> tp(nt + 1);
  TC_CUDA(cudaMemcpyAsync(cnt.data(), r->set_count.p, nt * sizeof(int32_t), cudaMemcpyDeviceToHost, r->stream));
  TC_CUDA(cudaMemcpyAsync(tp.data(), r->w.term_ptr, (nt + 1) * sizeof(int64_t), cudaMemcpyDeviceToHost, r->stream));
  std::vector<tc_set> all(nc);
  if (nc)
    TC_CUDA(cudaMemcpyAsync(all.data(), r->sets.p, nc * sizeof(tc_set), cudaMemcpyDeviceToHost, r->stream));
  TC_CUDA(cudaStreamSynchronize(r->stream));
  size_t ns = 0;
  for (size_t k = 0; k < nt; ++k)
  {
    for (int32_t j = 0; j < cnt[k]; ++j)
    {
      if (sets && ns < cap_sets)
        sets[ns] = all[static_cast<size_t>(tp[k]) + static_cast<size_t>(j)];
      ++ns;
    }
    set_ptr[k + 1] = static_cast<int64_t>(ns);
  }
  return TC_OK;
}

int tc_result_cost(const tc_result* res_c, double* cost_per_term, double* cost_per_traj)
{
  TC_TRY(needValid(res_c));
  tc_result* r = const_cast<tc_result*>(res_c);
  if (r->ep.n_terms == 0)
  {
    if (cost_per_traj)
      std::memset(cost_per_traj, 0, sizeof(double) * r->ep.B);
    return TC_OK;
  }
  if (!r->have_cost)
  {
    LaunchScope scope(r);
    TC_RESERVE(r->cost_term, r->ep.n_terms);
    TC_RESERVE(r->cost_traj, r->ep.B);
    TC_CUDA(launchCost(r->scene->dview, r->w, r->cost_term.p, r->cost_traj.p, r->stream));
    TC_CUDA(cudaStreamSynchronize(r->stream));
    r->have_cost = true;
  }
  if (cost_per_term)
    TC_CUDA(cudaMemcpyAsync(cost_per_term, r->cost_term.p, sizeof(double) * r->ep.n_terms, cudaMemcpyDeviceToHost, r->stream));
  if (cost_per_traj)
    TC_CUDA(cudaMemcpyAsync(cost_per_traj, r->cost_traj.p, sizeof(double) * r->ep.B, cudaMemcpyDeviceToHost, r->stream));
  TC_CUDA(cudaStreamSynchronize(r->stream));
  return TC_OK;
}

int tc_result_cost_device(const tc_result* res, const double** dev_ptr)
{
  TC_TRY(needValid(res));
  if (!res->have_cost)
    return fail(TC_ERR_INVALID, "evaluation ran without TC_WANT_COST");
  *dev_ptr = res->cost_traj.p;
  return TC_OK;
}

int tc_result_stage_ms(const tc_result* res, float ms[10])
{
  TC_TRY(needValid(res));
  for (int i = 0; i < 10; ++i)
    ms[i] = res->stage_ms[i];
  return TC_OK;
}

int tc_result_launches(const tc_result* res, int32_t* own_kernels, int32_t* library_calls)
{
  TC_TRY(needValid(res));
  if (own_kernels)
    *own_kernels = res->own_launches;  // accumulated by finishEval / buildRows / cost from launchCounter()
  if (library_calls)
    *library_calls = res->lib_launches;
  return TC_OK;
}

int tc_measure_fma_peak(int precision, double* tflops)
{
  if (g_device < 0)
    return fail(TC_ERR_CUDA, "tc_init() has not succeeded");
  TC_TRY(ensureDevice(g_device));
  TC_CUDA(measureFmaPeak(precision, tflops));
  return TC_OK;
}

int tc_nccl_unique_id(char id[128])
{
  TC_TRY(loadNccl());
  const int e = g_nccl.getUniqueId(id);
  return e ? ncclFail(e, "ncclGetUniqueId") : TC_OK;
}

int tc_nccl_comm_create(int n_ranks, int rank, const char id[128], void** comm)
{
  TC_TRY(loadNccl());
  if (g_device < 0)
    return fail(TC_ERR_CUDA, "tc_init() has not succeeded");
  TC_TRY(ensureDevice(g_device));
  Id128 uid;
  std::memcpy(uid.b, id, 128);
  const int e = g_nccl.commInitRank(comm, n_ranks, uid, rank);
  return e ? ncclFail(e, "ncclCommInitRank") : TC_OK;
}

void tc_nccl_comm_destroy(void* comm)
{
  if (comm && g_nccl.commDestroy)
    g_nccl.commDestroy(comm);
}

// ---------------------------------------------------------------------------------------------------------------
// tc_pipeline: sub-batches on a ring of (result, stream) slots, one host thread
// ---------------------------------------------------------------------------------------------------------------
}  // extern "C" (reopened below)

struct tc_pipeline
{
  tc_scene* scene{ nullptr };
  int device{ 0 };
  int n_slots{ 0 };
  // one ring of results per sub-batch size, so that every result keeps seeing passes of ONE shape and stays speculative
  struct Ring
  {
    int B{ 0 };
    std::vector<tc_result*> res;
  };
  std::vector<Ring> rings;
  std::vector<cudaStream_t> streams;
};

namespace
{
struct SubBatch
{
  int first, count;
};

// chunks of `chunk` trajectories; the last one tapered into 1/2 + 1/4 + 1/4 (its download is what nothing overlaps)
std::vector<SubBatch> cutBatch(int B, int chunk)
{
  std::vector<SubBatch> out;
  int first = 0;
  while (B - first > chunk)
  {
    out.push_back({ first, chunk });
    first += chunk;
  }
  const int last = B - first;
  if (last >= 8 && out.size() >= 1)
  {
    const int a = last / 2, b = last / 4;
    out.push_back({ first, a });
    out.push_back({ first + a, b });
    out.push_back({ first + a + b, last - a - b });
  }
  else
    out.push_back({ first, last });
  return out;
}

tc_result* ringSlot(tc_pipeline* p, int B, int slot, int* status)
{
  *status = TC_OK;
  for (auto& rg : p->rings)
    if (rg.B == B)
      return rg.res[static_cast<size_t>(slot)];
  tc_pipeline::Ring rg;
  rg.B = B;
  for (int i = 0; i < p->n_slots; ++i)
  {
    tc_result* r = nullptr;
    *status = tc_result_create(p->scene, &r);
    if (*status != TC_OK)
    {
      for (tc_result* x : rg.res)
        tc_result_destroy(x);
      return nullptr;
    }
    rg.res.push_back(r);
  }
  p->rings.push_back(rg);
  return p->rings.back().res[static_cast<size_t>(slot)];
}
}  // namespace

extern "C" {

int tc_pipeline_create(tc_scene* scene, int32_t n_slots, tc_pipeline** out)
{
  if (!scene || !out || n_slots < 1 || n_slots > 8)
    return fail(TC_ERR_INVALID, "tc_pipeline_create: bad argument (1..8 slots)");
  TC_TRY(ensureDevice(scene->device));
  tc_pipeline* p = new (std::nothrow) tc_pipeline();
  if (!p)
    return fail(TC_ERR_INVALID, "out of memory");
  p->scene = scene;
  p->device = scene->device;
  p->n_slots = n_slots;
  for (int i = 0; i < n_slots; ++i)
  {
    cudaStream_t st = nullptr;
    const cudaError_t e = cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking);
    if (e != cudaSuccess)
    {
      tc_pipeline_destroy(p);
      return cudaFail(e, "cudaStreamCreate");
    }
    p->streams.push_back(st);
  }
  *out = p;
  return TC_OK;
}

void tc_pipeline_destroy(tc_pipeline* p)
{
  if (!p)
    return;
  cudaSetDevice(p->device);
  for (auto& rg : p->rings)
    for (tc_result* r : rg.res)
      tc_result_destroy(r);
  for (cudaStream_t st : p->streams)
    cudaStreamDestroy(st);
  delete p;
  (void)cudaGetLastError();
}

}  // extern "C"

static int pipelineEvalRows(tc_pipeline* p, const tc_eval_desc* desc, const double* q, int32_t B, int32_t T, int32_t chunk,
                            int64_t* term_ptr, double* val, size_t val_cap, double* row_value, int32_t* row_pair,
                            size_t row_cap, double* cost_per_traj, int64_t* n_rows_out, int64_t* nnz_out, bool packed)
{
  if (!p || !desc || !q || !term_ptr || !val || !row_value || !row_pair || !cost_per_traj || !n_rows_out || !nnz_out)
    return fail(TC_ERR_INVALID, "tc_pipeline_eval_rows: null argument");
  if (B <= 0 || T <= 0 || chunk <= 0)
    return fail(TC_ERR_INVALID, "tc_pipeline_eval_rows: empty batch");
  if (desc->q_on_device)
    return fail(TC_ERR_INVALID, "tc_pipeline_eval_rows takes host joint values");
  TC_TRY(ensureDevice(p->scene->device));
  const int n_dof = p->scene->host.n_dof;
  tc_eval_desc d = *desc;
  d.want = TC_WANT_COST | TC_WANT_CONTACTS | TC_WANT_ROWS;
  std::vector<SubBatch> subs = cutBatch(B, chunk);
  if (const char* sched = getenv("TC_PIPELINE_SCHEDULE"))
  {
    // tuning aid: explicit sub-batch sizes "2304,1152,512,128" (must add up to B, else ignored)
    std::vector<SubBatch> alt;
    int first = 0;
    for (const char* c = sched; *c;)
    {
      const int n = atoi(c);
      if (n <= 0)
        break;
      alt.push_back({ first, n });
      first += n;
      while (*c && *c != ',')
        ++c;
      if (*c == ',')
        ++c;
    }
    if (first == B)
      subs = alt;
  }
  const int K = p->n_slots;
  struct InFlight
  {
    tc_result* r{ nullptr };
    SubBatch sb{};
    bool evaluating{ false }, downloading{ false };
  };
  std::vector<InFlight> slot(static_cast<size_t>(K));
  int64_t rows_done = 0, nnz_done = 0;
  int tpt = -1;
  // sub-batch j - 1 finished computing: size its download exactly and enqueue it behind its kernels
  auto download = [&](InFlight& f) -> int {
    TC_TRY(finishEval(f.r));
    tc_result* r = f.r;
    const int64_t nc = r->w.n_contacts, nz = r->h_counters->nnz;
    if (static_cast<size_t>(rows_done + nc) > row_cap || static_cast<size_t>(nnz_done + nz) > val_cap)
      return fail(TC_ERR_OVERFLOW, "tc_pipeline_eval_rows: the caller's row arrays are too small");
    const size_t nt = static_cast<size_t>(f.sb.count) * static_cast<size_t>(tpt);
    int64_t* tp = term_ptr + static_cast<size_t>(f.sb.first) * static_cast<size_t>(tpt);
    // the sub-batch's term offsets are local; they are shifted by the rows before it once everything has landed (the
    // closing entry of a sub-batch is the opening entry of the next, so only the nt opening entries are taken)
    if (nt)
      TC_CUDA(cudaMemcpyAsync(tp, r->w.term_ptr, nt * sizeof(int64_t), cudaMemcpyDeviceToHost, r->stream));
    if (nz)
      TC_CUDA(cudaMemcpyAsync(val + nnz_done, r->row_val.p, static_cast<size_t>(nz) * sizeof(double), cudaMemcpyDeviceToHost, r->stream));
    if (nc)
    {
      TC_CUDA(cudaMemcpyAsync(row_value + rows_done, r->row_value.p, static_cast<size_t>(nc) * sizeof(double), cudaMemcpyDeviceToHost, r->stream));
      TC_CUDA(cudaMemcpyAsync(row_pair + rows_done, r->row_pair.p, static_cast<size_t>(nc) * sizeof(int32_t), cudaMemcpyDeviceToHost, r->stream));
    }
    TC_CUDA(cudaMemcpyAsync(cost_per_traj + f.sb.first, r->cost_traj.p, static_cast<size_t>(f.sb.count) * sizeof(double),
                            cudaMemcpyDeviceToHost, r->stream));
    f.evaluating = false;
    f.downloading = true;
    // remember where this sub-batch's rows start: stored in the slot, applied after the final wait
    r->pipe_row_base = rows_done;
    rows_done += nc;
    nnz_done += nz;
    return TC_OK;
  };
  std::vector<std::pair<SubBatch, int64_t>> bases;  // (sub-batch, first row) for the final term_ptr shift
  std::vector<int> pending;  // slots whose pass is enqueued and not yet finished, oldest first
  static const int depth = [] {
    const char* e = getenv("TC_PIPELINE_DEPTH");  // tuning aid: passes enqueued beyond the one being waited for
    return e ? std::max(1, atoi(e)) : 1;
  }();
  auto finishSlot = [&](int s) -> int {
    InFlight& g = slot[static_cast<size_t>(s)];
    TC_TRY(download(g));
    bases.push_back({ g.sb, g.r->pipe_row_base });
    return TC_OK;
  };
  for (size_t j = 0; j < subs.size(); ++j)
  {
    const int s = static_cast<int>(j % static_cast<size_t>(K));
    InFlight& f = slot[static_cast<size_t>(s)];
    while (f.evaluating)
    {
      // the ring came round to a slot whose sub-batch is still computing: finish the oldest ones first (rows stay in
      // sub-batch order)
      TC_TRY(finishSlot(pending.front()));
      pending.erase(pending.begin());
    }
    if (f.downloading)
    {
      // the slot's previous sub-batch must have left the device before its arenas are reused
      TC_CUDA(cudaStreamSynchronize(p->streams[static_cast<size_t>(s)]));
      f.downloading = false;
    }
    int status = TC_OK;
    f.r = ringSlot(p, subs[j].count, s, &status);
    if (!f.r)
      return status;
    f.sb = subs[j];
    f.r->rows_packed = packed;
    TC_TRY(runEval(p->scene, &d, q + static_cast<size_t>(subs[j].first) * T * n_dof, subs[j].count, T, f.r,
                   p->streams[static_cast<size_t>(s)]));
    tpt = f.r->ep.terms_per_traj;
    f.evaluating = true;
    // the ones before it (on other streams) finish while this one is already queued
    pending.push_back(s);
    while (static_cast<int>(pending.size()) > depth)
    {
      TC_TRY(finishSlot(pending.front()));
      pending.erase(pending.begin());
    }
  }
  for (int s : pending)
    TC_TRY(finishSlot(s));
  for (int s = 0; s < K; ++s)
    if (slot[static_cast<size_t>(s)].downloading)
      TC_CUDA(cudaStreamSynchronize(p->streams[static_cast<size_t>(s)]));
  // term offsets become global
  for (const auto& b : bases)
  {
    const size_t nt = static_cast<size_t>(b.first.count) * static_cast<size_t>(tpt);
    int64_t* tp = term_ptr + static_cast<size_t>(b.first.first) * static_cast<size_t>(tpt);
    for (size_t k = 0; k < nt; ++k)
      tp[k] += b.second;
  }
  term_ptr[static_cast<size_t>(B) * static_cast<size_t>(std::max(tpt, 0))] = rows_done;
  *n_rows_out = rows_done;
  *nnz_out = nnz_done;
  return TC_OK;
}

extern "C" {

int tc_pipeline_eval_rows(tc_pipeline* p, const tc_eval_desc* desc, const double* q, int32_t B, int32_t T, int32_t chunk,
                          int64_t* term_ptr, double* val, size_t val_cap, double* row_value, int32_t* row_pair,
                          size_t row_cap, double* cost_per_traj, int64_t* n_rows_out, int64_t* nnz_out)
{
  return pipelineEvalRows(p, desc, q, B, T, chunk, term_ptr, val, val_cap, row_value, row_pair, row_cap, cost_per_traj,
                          n_rows_out, nnz_out, false);
}

int tc_pipeline_eval_rows_packed(tc_pipeline* p, const tc_eval_desc* desc, const double* q, int32_t B, int32_t T,
                                 int32_t chunk, int64_t* term_ptr, double* val, size_t val_cap, double* row_value,
                                 int32_t* row_pair, size_t row_cap, double* cost_per_traj, int64_t* n_rows_out,
                                 int64_t* nnz_out)
{
  return pipelineEvalRows(p, desc, q, B, T, chunk, term_ptr, val, val_cap, row_value, row_pair, row_cap, cost_per_traj,
                          n_rows_out, nnz_out, true);
}

int tc_scene_pair_dof_mask(const tc_scene* scene, int32_t pair, uint64_t* mask)
{
  if (!scene || !mask)
    return fail(TC_ERR_INVALID, "tc_scene_pair_dof_mask: null argument");
  const HostScene& hs = scene->host;
  if (pair < 0 || static_cast<size_t>(pair) >= hs.pairs.size())
    return fail(TC_ERR_INVALID, "tc_scene_pair_dof_mask: pair out of range");
  const PairD& pr = hs.pairs[static_cast<size_t>(pair)];
  *mask = (pr.a < hs.n_active ? hs.link_dof_mask[static_cast<size_t>(pr.a)] : 0ull) |
          (pr.b < hs.n_active ? hs.link_dof_mask[static_cast<size_t>(pr.b)] : 0ull);
  return TC_OK;
}

// Device + pinned-host staging of the multi-start collectives, one per process, grown on demand and kept: the
// collectives themselves never call cudaMalloc / cudaFree (each is a device-wide synchronisation).
static int collectiveScratch(size_t dev_bytes, size_t host_bytes, unsigned char** dev, unsigned char** host)
{
  static std::mutex mu;
  static unsigned char *d = nullptr, *h = nullptr;
  static size_t dcap = 0, hcap = 0;
  static int device = -1;
  std::lock_guard<std::mutex> lock(mu);
  if (device != g_device)
  {
    d = nullptr, dcap = 0;  // another device: its scratch is allocated afresh (the old one lives until exit)
    device = g_device;
  }
  if (dev_bytes > dcap)
  {
    if (d)
      cudaFree(d);
    d = nullptr;
    dcap = 0;
    const size_t want = std::max<size_t>(dev_bytes * 2, 1 << 16);
    TC_CUDA(cudaMalloc(reinterpret_cast<void**>(&d), want));
    dcap = want;
  }
  if (host_bytes > hcap)
  {
    if (h)
      cudaFreeHost(h);
    h = nullptr;
    hcap = 0;
    const size_t want = std::max<size_t>(host_bytes * 2, 1 << 16);
    TC_CUDA(cudaHostAlloc(reinterpret_cast<void**>(&h), want, cudaHostAllocDefault));
    hcap = want;
  }
  *dev = d;
  *host = h;
  return TC_OK;
}

int tc_allgather_best(double local_best_cost, int64_t local_best_idx, void* nccl_comm, int n_ranks, void* cuda_stream,
                      double* best_cost, int64_t* best_idx, int32_t* best_rank)
{
  if (!nccl_comm || n_ranks <= 0 || !best_cost || !best_idx || !best_rank)
    return fail(TC_ERR_INVALID, "bad argument");
  TC_TRY(loadNccl());
  if (g_device < 0)
    return fail(TC_ERR_CUDA, "tc_init() has not succeeded");
  TC_TRY(ensureDevice(g_device));
  cudaStream_t st = static_cast<cudaStream_t>(cuda_stream);
  struct Rec
  {
    double cost;
    int64_t idx;
  };
  unsigned char *dv = nullptr, *hv = nullptr;
  const size_t n = static_cast<size_t>(n_ranks);
  TC_TRY(collectiveScratch(sizeof(Rec) * (n + 1), sizeof(Rec) * (n + 1), &dv, &hv));
  Rec* d = reinterpret_cast<Rec*>(dv);
  Rec* h = reinterpret_cast<Rec*>(hv);
  h[n] = Rec{ local_best_cost, local_best_idx };
  TC_CUDA(cudaMemcpyAsync(d + n, h + n, sizeof(Rec), cudaMemcpyHostToDevice, st));
  const int ne = g_nccl.allGather(d + n, d, sizeof(Rec), /* ncclChar */ 0, nccl_comm, st);
  if (ne)
    return ncclFail(ne, "ncclAllGather");
  TC_CUDA(cudaMemcpyAsync(h, d, sizeof(Rec) * n, cudaMemcpyDeviceToHost, st));
  TC_CUDA(cudaStreamSynchronize(st));
  int br = 0;
  for (int i = 1; i < n_ranks; ++i)
    if (h[i].cost < h[br].cost)
      br = i;
  *best_cost = h[br].cost;
  *best_idx = h[br].idx;
  *best_rank = br;
  return TC_OK;
}

int tc_broadcast_trajectory(double* traj, size_t n_values, int root_rank, void* nccl_comm, void* cuda_stream)
{
  if (!nccl_comm || !traj || n_values == 0 || root_rank < 0)
    return fail(TC_ERR_INVALID, "bad argument");
  TC_TRY(loadNccl());
  if (g_device < 0)
    return fail(TC_ERR_CUDA, "tc_init() has not succeeded");
  TC_TRY(ensureDevice(g_device));
  cudaStream_t st = static_cast<cudaStream_t>(cuda_stream);
  const size_t bytes = n_values * sizeof(double);
  unsigned char *dv = nullptr, *hv = nullptr;
  TC_TRY(collectiveScratch(bytes, bytes, &dv, &hv));
  std::memcpy(hv, traj, bytes);
  // every rank uploads its buffer; only the root's content is read by the collective
  TC_CUDA(cudaMemcpyAsync(dv, hv, bytes, cudaMemcpyHostToDevice, st));
  const int ne = g_nccl.broadcast(dv, dv, bytes, /* ncclChar */ 0, root_rank, nccl_comm, st);
  if (ne)
    return ncclFail(ne, "ncclBroadcast");
  TC_CUDA(cudaMemcpyAsync(hv, dv, bytes, cudaMemcpyDeviceToHost, st));
  TC_CUDA(cudaStreamSynchronize(st));
  std::memcpy(traj, hv, bytes);
  return TC_OK;
}

int tc_result_allgather_best(tc_result* r, int64_t index_offset, void* nccl_comm, int n_ranks, double* best_cost,
                             int64_t* best_idx, int32_t* best_rank, double* best_traj)
{
  if (!r || n_ranks <= 0 || !best_cost || !best_idx || !best_rank)
    return fail(TC_ERR_INVALID, "bad argument");
  if (n_ranks > 1 && !nccl_comm)
    return fail(TC_ERR_INVALID, "a communicator is needed for more than one rank");
  // the evaluation is finished first: a speculative pass may have to be repeated, and a repeat must not happen between
  // ranks that have already entered the collective
  if (r->pending)
    TC_TRY(finishEval(r));
  TC_TRY(needValid(r));
  if (!r->have_cost || !r->w.q)
    return fail(TC_ERR_INVALID, "the evaluation ran without TC_WANT_COST or without joint values");
  if (n_ranks > 1)
    TC_TRY(loadNccl());
  const size_t nv = static_cast<size_t>(r->ep.T) * static_cast<size_t>(r->scene->host.n_dof);
  const size_t rec = (2 + nv) * sizeof(double);  // cost, global index (as int64), trajectory
  const size_t n = static_cast<size_t>(n_ranks);
  unsigned char *dv = nullptr, *hv = nullptr;
  TC_TRY(collectiveScratch(rec * (n + 1), rec * n, &dv, &hv));
  cudaStream_t st = r->stream;
  {
    LaunchScope scope(r);
    TC_CUDA(launchBestRecord(r->w, r->cost_traj.p, index_offset, reinterpret_cast<double*>(dv + rec * n), st));
  }
  if (n_ranks > 1)
  {
    const int ne = g_nccl.allGather(dv + rec * n, dv, rec, /* ncclChar */ 0, nccl_comm, st);
    if (ne)
      return ncclFail(ne, "ncclAllGather");
    TC_CUDA(cudaMemcpyAsync(hv, dv, rec * n, cudaMemcpyDeviceToHost, st));
  }
  else
    TC_CUDA(cudaMemcpyAsync(hv, dv + rec * n, rec, cudaMemcpyDeviceToHost, st));
  TC_CUDA(cudaStreamSynchronize(st));
  int br = 0;
  auto costOf = [&](int i) {
    double c;
    std::memcpy(&c, hv + rec * static_cast<size_t>(i), sizeof(double));
    return c;
  };
  for (int i = 1; i < n_ranks; ++i)
    if (costOf(i) < costOf(br))
      br = i;
  *best_cost = costOf(br);
  std::memcpy(best_idx, hv + rec * static_cast<size_t>(br) + sizeof(double), sizeof(int64_t));
  *best_rank = br;
  if (best_traj)
    std::memcpy(best_traj, hv + rec * static_cast<size_t>(br) + 2 * sizeof(double), nv * sizeof(double));
  return TC_OK;
}
}  // extern "C"

// ---------------------------------------------------------------------------------------------------------------
// Allocation-free entry point (no counterpart in the reference).
// The hinge-cost reduction of launchCost (k_cost_term -> k_cost_traj, tc_pipeline.cu) restated over caller-owned
// device arrays, with the per-term intermediate kept in a file-scope device array instead of a cudaMalloc'd
// buffer.  Every GPU that loads this library holds TC_NOALLOC_MAX_TERMS * 8 bytes (512 KiB) until exit.
// All calls share the array: calls that overlap on the GPU (different streams, no ordering) race.
// Nothing in it has to start at zero: the first kernel writes every slot [0, n_terms) the second one reads.
// ---------------------------------------------------------------------------------------------------------------
static __device__ __align__(256) double g_noalloc_cost_term[TC_NOALLOC_MAX_TERMS];

namespace
{
// same arithmetic and summation order as k_cost_term: one warp per term, lanes stride over the term's contacts,
// shuffle tree
__global__ void __launch_bounds__(256) k_cost_term_noalloc(const long long* __restrict__ term_ptr, int n_terms,
                                                           const double* __restrict__ distance,
                                                           const double* __restrict__ margin,
                                                           const double* __restrict__ coeff, double* cost_term)
{
  const int k = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const unsigned lane = threadIdx.x & 31;
  if (k >= n_terms)
    return;
  double out = 0;
  for (long long i = term_ptr[k] + lane; i < term_ptr[k + 1]; i += 32)
    out += fmax(margin[i] - distance[i], 0.0) * coeff[i];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1)
    out += __shfl_down_sync(0xffffffffu, out, o);
  if (lane == 0)
    cost_term[k] = out;
}
__global__ void k_cost_traj_noalloc(int n_traj, int terms_per_traj, const double* cost_term, double* cost_traj)
{
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= n_traj)
    return;
  double out = 0;
  for (int ti = 0; ti < terms_per_traj; ++ti)
    out += cost_term[static_cast<size_t>(b) * terms_per_traj + ti];
  cost_traj[b] = out;
}

// device address of g_noalloc_cost_term, resolved once per device with that device current (in host code the
// array's name is only a stand-in for the symbol, not an address a kernel can use)
constexpr int NOALLOC_MAX_DEVICES = 64;
std::mutex g_noalloc_mutex;
double* g_noalloc_ptr[NOALLOC_MAX_DEVICES] = {};

// `device` must be the current device
int resolveNoallocScratch(int device, double** out)
{
  if (device < 0 || device >= NOALLOC_MAX_DEVICES)
    return fail(TC_ERR_INVALID, "tc_cost_noalloc: device ordinal out of range");
  std::lock_guard<std::mutex> lock(g_noalloc_mutex);
  if (!g_noalloc_ptr[device])
  {
    double* p = nullptr;
    TC_CUDA(cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_noalloc_cost_term));
    g_noalloc_ptr[device] = p;
  }
  *out = g_noalloc_ptr[device];
  return TC_OK;
}
}  // namespace

extern "C" int tc_cost_noalloc(const int64_t* term_ptr, int32_t n_terms, int32_t terms_per_traj,
                               const double* distance, const double* margin, const double* coeff,
                               double* cost_per_traj, void* cuda_stream)
{
  static_assert(sizeof(long long) == sizeof(int64_t), "term_ptr layout");
  if (!term_ptr || !distance || !margin || !coeff || !cost_per_traj || n_terms <= 0 || terms_per_traj <= 0)
    return fail(TC_ERR_INVALID, "tc_cost_noalloc: null argument or empty batch");
  if (n_terms % terms_per_traj != 0)
    return fail(TC_ERR_INVALID, "tc_cost_noalloc: n_terms is not a multiple of terms_per_traj");
  if (n_terms > TC_NOALLOC_MAX_TERMS)
    return fail(TC_ERR_OVERFLOW, "tc_cost_noalloc: n_terms = " + std::to_string(n_terms) + " exceeds the static scratch of " +
                                     std::to_string(TC_NOALLOC_MAX_TERMS) +
                                     " terms (TC_NOALLOC_MAX_TERMS); split the batch or use tc_result_cost");
  // runs on the CURRENT device (the arrays are the caller's); the address of that device's copy of the scratch was
  // resolved by tc_init for its device, and is resolved here on first use for any other
  int dev = -1;
  TC_CUDA(cudaGetDevice(&dev));
  double* scratch = nullptr;
  TC_TRY(resolveNoallocScratch(dev, &scratch));
  cudaStream_t st = static_cast<cudaStream_t>(cuda_stream);
  const int n_traj = n_terms / terms_per_traj;
  const unsigned grid_term = static_cast<unsigned>((static_cast<long long>(n_terms) * 32 + 255) / 256);
  k_cost_term_noalloc<<<grid_term, 256, 0, st>>>(reinterpret_cast<const long long*>(term_ptr), n_terms, distance, margin,
                                                 coeff, scratch);
  TC_CUDA(cudaGetLastError());
  k_cost_traj_noalloc<<<(n_traj + 255) / 256, 256, 0, st>>>(n_traj, terms_per_traj, scratch, cost_per_traj);
  TC_CUDA(cudaGetLastError());
  return TC_OK;
}
