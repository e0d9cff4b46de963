"""ctypes view of the C ABI declared in include/trajopt_cuda.h.

This is the thin Python driver used by tests, bench.py and __graft_entry__: it only marshals numpy
arrays to the `tc_*` entry points of libtrajopt_cuda.so.  No arithmetic of the hot path happens here and
there is no fallback: if the shared library (or a CUDA device) is missing the calls raise.

The same class also drives any other library exporting the same entry points under another prefix
(tests load the CPU oracle, `orc_*`, through it — the oracle is never loaded from this package).
"""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("TC_LIB_PATH", os.path.join(HERE, "csrc", "libtrajopt_cuda.so"))

# ---- enums (include/trajopt_cuda.h) ----
TC_OK = 0
TC_ERR_INVALID, TC_ERR_CUDA, TC_ERR_UNSUPPORTED, TC_ERR_OVERFLOW, TC_ERR_NCCL = 1, 2, 3, 4, 5
NOALLOC_MAX_TERMS = 65536  # TC_NOALLOC_MAX_TERMS of include/trajopt_cuda.h
JOINT_FIXED, JOINT_REVOLUTE, JOINT_PRISMATIC = 0, 1, 2
SHAPE_SPHERE, SHAPE_HULL, SHAPE_CAPSULE, SHAPE_CYLINDER, SHAPE_CONE = 0, 1, 2, 3, 4
EVAL_DISCRETE, EVAL_LVS_DISCRETE, EVAL_CONTINUOUS, EVAL_LVS_CONTINUOUS = 1, 2, 3, 4
STACK_IFOPT, STACK_LEGACY = 0, 1
PRECISION_F64, PRECISION_F32 = 0, 1
CC_NONE, CC_TIME0, CC_TIME1, CC_BETWEEN = 0, 1, 2, 3
WANT_COST, WANT_CONTACTS, WANT_GRADIENTS, WANT_ROWS = 1, 2, 4, 8
WANT_ALL = 7
WANT_LINEARISE = WANT_COST | WANT_CONTACTS | WANT_ROWS  # rows folded from the gradients on the fly
FLAG_EPA, FLAG_DEGENERATE, FLAG_UNCONVERGED = 1, 2, 4

_pd = C.POINTER(C.c_double)
_pi = C.POINTER(C.c_int32)
_pl = C.POINTER(C.c_int64)
_pb = C.POINTER(C.c_uint8)


class SceneDesc(C.Structure):
    _fields_ = [
        ("n_dof", C.c_int32), ("n_frames", C.c_int32),
        ("frame_parent", _pi), ("frame_joint_type", _pi), ("frame_dof", _pi),
        ("frame_axis", _pd), ("frame_origin", _pd), ("base_pose", C.c_double * 12),
        ("n_links", C.c_int32), ("n_active_links", C.c_int32),
        ("link_frame", _pi), ("link_pose", _pd),
        ("n_shapes", C.c_int32),
        ("shape_type", _pi), ("shape_link", _pi), ("shape_id", _pi),
        ("shape_params", _pd), ("shape_pose", _pd),
        ("shape_vert_offset", _pi), ("shape_vert_count", _pi),
        ("n_verts", C.c_int32), ("verts", _pd),
        ("acm", _pb), ("pair_margin", _pd), ("pair_coeff", _pd),
        ("default_margin", C.c_double), ("default_coeff", C.c_double), ("margin_buffer", C.c_double),
        ("n_env_active_links", C.c_int32), ("reserved", C.c_int32),
    ]


class EvalDesc(C.Structure):
    _fields_ = [
        ("evaluator", C.c_int32), ("stack", C.c_int32), ("precision", C.c_int32), ("want", C.c_int32),
        ("lvs_length", C.c_double), ("fixed_steps", _pb), ("q_on_device", C.c_int32), ("contact_test_type", C.c_int32),
    ]


CONTACT_DTYPE = np.dtype([
    ("distance", "<f8"), ("normal", "<f8", 3), ("nearest_points", "<f8", (2, 3)),
    ("nearest_points_local", "<f8", (2, 3)), ("cc_time", "<f8", 2), ("cc_type", "<i4", 2),
    ("link_ids", "<i4", 2), ("shape_id", "<i4", 2), ("subshape_id", "<i4", 2),
    ("traj", "<i4"), ("term", "<i4"), ("substep", "<i4"), ("pair", "<i4"), ("state", "<i4"), ("flags", "<i4"),
], align=True)
assert CONTACT_DTYPE.itemsize == 200, CONTACT_DTYPE.itemsize

GRADIENT_DTYPE = np.dtype([
    ("error", "<f8"), ("error_with_buffer", "<f8"), ("scale0", "<f8", 2), ("scale1", "<f8", 2),
    ("has_gradient", "<i4", 2), ("cc_type", "<i4", 2),
], align=True)
assert GRADIENT_DTYPE.itemsize == 64

COUNT_FIELDS = ["n_states", "n_units", "n_dense", "n_candidates", "n_narrowphase", "n_contacts", "n_rows", "nnz",
                "gjk_iters", "support_verts", "n_epa", "n_unconverged", "n_terms", "overflow", "epa_iters", "support_compared"]

SET_DTYPE = np.dtype([
    ("traj", "<i4"), ("term", "<i4"), ("pair", "<i4"), ("shape_key", "<i4", 2), ("n_results", "<i4"),
    ("has_error", "<i4", (2, 2)), ("coeff", "<f8"), ("error", "<f8", (2, 2)), ("error_with_buffer", "<f8", (2, 2)),
], align=True)
assert SET_DTYPE.itemsize == 112, SET_DTYPE.itemsize


def _p(a, t):
    return a.ctypes.data_as(t)


def expand_columns(row_ptr, first_col, n_dof):
    """col_idx of tc_result_rows from the block form of tc_result_rows_blocks: every present block of a row is a run
    of n_dof consecutive columns"""
    col = np.zeros(int(row_ptr[-1]), np.int32)
    if len(first_col) == 0:
        return col
    present = first_col >= 0
    start = row_ptr[:-1, None] + np.cumsum(present, axis=1) * n_dof - present * n_dof  # first non-zero of each block
    for b in range(2):
        rows = np.nonzero(present[:, b])[0]
        idx = start[rows, b][:, None] + np.arange(n_dof)[None, :]
        col[idx] = first_col[rows, b][:, None] + np.arange(n_dof, dtype=np.int32)[None, :]
    return col


class TcError(RuntimeError):
    pass


class Scene:
    """Owns a tc_scene handle plus the numpy arrays the description points at."""

    def __init__(self, lib, arrays, margin, coeff, margin_buffer, pair_margin=None, pair_coeff=None):
        self.lib = lib
        a = self.arrays = {k: np.ascontiguousarray(v) for k, v in arrays.items()
                           if isinstance(v, np.ndarray) and v.dtype.kind in "fiu"}
        d = SceneDesc()
        d.n_dof = int(arrays["n_dof"])
        d.n_frames = len(a["frame_parent"])
        for k in ("frame_parent", "frame_joint_type", "frame_dof"):
            a[k] = a[k].astype(np.int32)
            setattr(d, k, _p(a[k], _pi))
        for k in ("frame_axis", "frame_origin", "link_pose", "shape_params", "shape_pose", "verts"):
            a[k] = np.ascontiguousarray(a[k], dtype=np.float64)
            setattr(d, k, _p(a[k], _pd))
        d.base_pose = (C.c_double * 12)(*[float(x) for x in arrays["base_pose"]])
        d.n_links = len(a["link_frame"])
        d.n_active_links = int(arrays["n_active_links"])
        for k in ("link_frame", "shape_type", "shape_link", "shape_id", "shape_vert_offset", "shape_vert_count"):
            a[k] = a[k].astype(np.int32)
            setattr(d, k, _p(a[k], _pi))
        d.n_shapes = len(a["shape_type"])
        d.n_verts = len(a["verts"])
        a["acm"] = np.ascontiguousarray(a["acm"], dtype=np.uint8)
        d.acm = _p(a["acm"], _pb)
        if pair_margin is not None:
            a["pair_margin"] = np.ascontiguousarray(pair_margin, dtype=np.float64)
            d.pair_margin = _p(a["pair_margin"], _pd)
        if pair_coeff is not None:
            a["pair_coeff"] = np.ascontiguousarray(pair_coeff, dtype=np.float64)
            d.pair_coeff = _p(a["pair_coeff"], _pd)
        d.default_margin, d.default_coeff, d.margin_buffer = float(margin), float(coeff), float(margin_buffer)
        d.n_env_active_links = int(arrays.get("n_env_active_links", 0))
        self.desc = d
        self.n_dof = d.n_dof
        self.n_links = d.n_links
        self.n_active = d.n_active_links
        self.margin_buffer = float(margin_buffer)
        self.handle = C.c_void_p()
        lib.check(lib.fn("scene_create")(C.byref(d), C.byref(self.handle)))

    def set_link_poses(self, link_pose):
        """dynamic environment: new world poses of the links outside the planning group (tc_scene_set_link_poses)"""
        lp = np.ascontiguousarray(link_pose, dtype=np.float64).reshape(self.n_links, 12)
        self.lib.check(self.lib.fn("scene_set_link_poses")(self.handle, _p(lp, _pd)))
        self.arrays["link_pose"] = lp

    def num_pairs(self):
        return self.lib.fn("scene_num_pairs")(self.handle)

    def pairs(self):
        n = self.num_pairs()
        l0, l1 = np.zeros(n, np.int32), np.zeros(n, np.int32)
        m, c = np.zeros(n), np.zeros(n)
        self.lib.check(self.lib.fn("scene_pairs")(self.handle, _p(l0, _pi), _p(l1, _pi), _p(m, _pd), _p(c, _pd)))
        return l0, l1, m, c

    def close(self):
        if self.handle:
            self.lib.fn("scene_destroy")(self.handle)
            self.handle = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class Result:
    def __init__(self, lib, scene, handle, B, T, owned=True):
        self.lib, self.scene, self.handle, self.B, self.T, self.owned = lib, scene, handle, B, T, owned

    def counts(self):
        buf = np.zeros(len(COUNT_FIELDS), np.int64)
        self.lib.check(self.lib.fn("result_counts")(self.handle, _p(buf, _pl)))
        return dict(zip(COUNT_FIELDS, (int(x) for x in buf)))

    def contacts(self):
        n = self.counts()["n_contacts"]
        out = np.zeros(n, CONTACT_DTYPE)
        self.lib.check(self.lib.fn("result_contacts")(self.handle, out.ctypes.data_as(C.c_void_p), C.c_size_t(n)))
        return out

    def contact_transforms(self):
        n = self.counts()["n_contacts"]
        out = np.zeros((n, 2, 2, 12))
        self.lib.check(self.lib.fn("result_contact_transforms")(self.handle, _p(out, _pd), C.c_size_t(n)))
        return out

    def term_ptr(self):
        n = self.counts()["n_terms"]
        out = np.zeros(n + 1, np.int64)
        self.lib.check(self.lib.fn("result_term_ptr")(self.handle, _p(out, _pl)))
        return out

    def gradients(self):
        n = self.counts()["n_contacts"]
        rec = np.zeros(n, GRADIENT_DTYPE)
        grad = np.zeros((n, 2, 2, self.scene.n_dof))
        self.lib.check(self.lib.fn("result_gradients")(self.handle, rec.ctypes.data_as(C.c_void_p), _p(grad, _pd),
                                                       C.c_size_t(n)))
        return rec, grad

    def rows(self):
        n = self.counts()["n_contacts"]
        nd = self.scene.n_dof
        row_ptr = np.zeros(n + 1, np.int64)
        col = np.zeros(max(1, 2 * n * nd), np.int32)
        val = np.zeros(max(1, 2 * n * nd))
        rv, rc, rm = np.zeros(n), np.zeros(n), np.zeros(n)
        self.lib.check(self.lib.fn("result_rows")(self.handle, _p(row_ptr, _pl), _p(col, _pi), _p(val, _pd),
                                                  _p(rv, _pd), _p(rc, _pd), _p(rm, _pd)))
        nnz = int(row_ptr[-1])
        return dict(row_ptr=row_ptr, col_idx=col[:nnz], val=val[:nnz], value=rv, coeff=rc, margin=rm)

    def rows_blocks(self):
        """rows without explicit column indices (tc_result_rows_blocks); `expand_columns` rebuilds col_idx"""
        n = self.counts()["n_contacts"]
        nd = self.scene.n_dof
        row_ptr = np.zeros(n + 1, np.int64)
        first = np.full((max(1, n), 2), -1, np.int32)
        val = np.zeros(max(1, 2 * n * nd))
        rv, rc, rm = np.zeros(n), np.zeros(n), np.zeros(n)
        self.lib.check(self.lib.fn("result_rows_blocks")(self.handle, _p(row_ptr, _pl), _p(first, _pi), _p(val, _pd),
                                                         _p(rv, _pd), _p(rc, _pd), _p(rm, _pd)))
        nnz = int(row_ptr[-1])
        return dict(row_ptr=row_ptr, first_col=first[:n], val=val[:nnz], value=rv, coeff=rc, margin=rm)

    def qp_csc(self):
        """hinge-cost QP block [A_x | -I] in CSC form (tc_result_qp_csc)"""
        nr, ncol, nz = C.c_int64(), C.c_int64(), C.c_int64()
        # sizes first (this assembles the block on the device if it has not been yet)
        self.lib.check(self.lib.fn("result_qp_csc_device")(self.handle, None, None, None, None, None, C.byref(nr),
                                                           C.byref(ncol), C.byref(nz)))
        n, n_cols, nnz = nr.value, ncol.value, nz.value
        col_ptr = np.zeros(n_cols + 1, np.int64)
        row_idx = np.zeros(max(1, nnz), np.int32)
        val = np.zeros(max(1, nnz))
        upper, cost = np.zeros(max(1, n)), np.zeros(max(1, n))
        self.lib.check(self.lib.fn("result_qp_csc")(self.handle, _p(col_ptr, _pl), _p(row_idx, _pi), _p(val, _pd),
                                                    _p(upper, _pd), _p(cost, _pd)))
        return dict(col_ptr=col_ptr, row_idx=row_idx[:nnz], val=val[:nnz], upper=upper[:n], slack_cost=cost[:n],
                    shape=(n, n_cols))

    def sets(self, max_num_cnt):
        nt = self.counts()["n_terms"]
        nc = self.counts()["n_contacts"]
        nd = self.scene.n_dof
        set_ptr = np.zeros(nt + 1, np.int64)
        sets = np.zeros(max(1, nc), SET_DTYPE)
        values = np.zeros((nt, max_num_cnt))
        coeffs = np.zeros((nt, max_num_cnt))
        jac = np.zeros((nt, max_num_cnt, 2, nd))
        self.lib.check(self.lib.fn("result_sets")(self.handle, C.c_int32(max_num_cnt), _p(set_ptr, _pl),
                                                  sets.ctypes.data_as(C.c_void_p), C.c_size_t(len(sets)),
                                                  _p(values, _pd), _p(coeffs, _pd), _p(jac, _pd)))
        return dict(set_ptr=set_ptr, sets=sets[:int(set_ptr[-1])], values=values, coeffs=coeffs, jac=jac)

    def cost(self):
        nt = self.counts()["n_terms"]
        per_term, per_traj = np.zeros(max(1, nt)), np.zeros(self.B)
        self.lib.check(self.lib.fn("result_cost")(self.handle, _p(per_term, _pd), _p(per_traj, _pd)))
        return per_term[:nt], per_traj

    def close(self):
        if self.owned and self.handle:
            self.lib.fn("result_destroy")(self.handle)
            self.handle = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


CONTACT_ALL, CONTACT_CLOSEST, CONTACT_FIRST = 0, 1, 2


def make_eval_desc(evaluator, stack=STACK_IFOPT, lvs=0.005, fixed=None, want=WANT_ALL, precision=PRECISION_F64,
                   q_on_device=False, contact_type=CONTACT_ALL):
    d = EvalDesc()
    d.evaluator, d.stack, d.precision, d.want = int(evaluator), int(stack), int(precision), int(want)
    d.lvs_length = float(lvs)
    keep = None
    if fixed is not None:
        keep = np.ascontiguousarray(fixed, dtype=np.uint8)
        d.fixed_steps = _p(keep, _pb)
    d.q_on_device = 1 if q_on_device else 0
    d.contact_test_type = int(contact_type)
    return d, keep


class Lib:
    """A loaded shared library exporting `<prefix>scene_create`, `<prefix>eval`, ..."""

    def __init__(self, path=LIB_PATH, prefix="tc_"):
        if not os.path.exists(path):
            raise TcError(f"{path} is missing: build it with __graft_entry__.build() (no fallback exists)")
        self.dll = C.CDLL(path)
        self.prefix = prefix
        self.path = path
        le = getattr(self.dll, prefix + "last_error")
        le.restype = C.c_char_p
        for name in ("scene_destroy", "result_destroy"):
            if hasattr(self.dll, prefix + name):
                getattr(self.dll, prefix + name).restype = None
                getattr(self.dll, prefix + name).argtypes = [C.c_void_p]

    def fn(self, name):
        return getattr(self.dll, self.prefix + name)

    def last_error(self):
        return self.fn("last_error")().decode()

    def check(self, status):
        if status != TC_OK:
            raise TcError(f"{self.prefix}* call failed with status {status}: {self.last_error()}")

    def scene(self, arrays, margin, coeff, margin_buffer, **kw):
        return Scene(self, arrays, margin, coeff, margin_buffer, **kw)


def expand_compact_rows(term_ptr, val, n_dof, blocks_per_term):
    """row_ptr of compact rows (tc_pipeline_eval_rows): row i of term k holds blocks_per_term[k] * n_dof values;
    blocks_per_term: [n_terms] (the same for every trajectory: tile the per-trajectory pattern)"""
    rows_per_term = np.diff(term_ptr)
    nnz_per_row = np.repeat(np.asarray(blocks_per_term, dtype=np.int64) * n_dof, rows_per_term)
    row_ptr = np.zeros(len(nnz_per_row) + 1, np.int64)
    np.cumsum(nnz_per_row, out=row_ptr[1:])
    assert row_ptr[-1] == len(val)
    return row_ptr


def unpack_rows(term_ptr, val, pair, pair_masks, n_dof, blocks_per_term):
    """packed rows (tc_pipeline_eval_rows_packed) -> (row_ptr, val) of the dense form tc_pipeline_eval_rows returns:
    a block of row i holds the joints set in pair_masks[pair[i]] (tc_scene_pair_dof_mask), ascending"""
    rows_per_term = np.diff(term_ptr)
    nb = np.repeat(np.asarray(blocks_per_term, dtype=np.int64), rows_per_term)
    bits = ((np.asarray(pair_masks, dtype=np.uint64)[pair][:, None] >> np.arange(n_dof, dtype=np.uint64)[None, :]) & 1).astype(bool)
    width = bits.sum(axis=1) * nb
    assert width.sum() == len(val)
    row_ptr = np.zeros(len(nb) + 1, np.int64)
    np.cumsum(nb * n_dof, out=row_ptr[1:])
    dense = np.zeros(row_ptr[-1], np.float64)
    # positions of the kept entries of every row, block after block
    keep = np.concatenate([bits, bits], axis=1)  # [rows, 2 * n_dof]; a one-block row uses the first half only
    keep[nb == 1, n_dof:] = False
    keep[nb == 0, :] = False
    cols = np.nonzero(keep)
    dense[row_ptr[cols[0]] + cols[1]] = val
    return row_ptr, dense


class Pipeline:
    """tc_pipeline: a batch cut into sub-batches on a ring of streams inside the library, rows downloaded in compact
    form while the next sub-batch computes (one host thread).  Output arrays are allocated by the caller-supplied
    `alloc(n, dtype)` (bench.py passes a pinned-memory allocator; numpy otherwise -> synchronous copies)."""

    def __init__(self, lib, scene, n_slots=3):
        self.lib, self.scene = lib, scene
        self.handle = C.c_void_p()
        lib.check(lib.dll.tc_pipeline_create(scene.handle, C.c_int32(n_slots), C.byref(self.handle)))
        self.out = None

    def close(self):
        if self.handle:
            self.lib.dll.tc_pipeline_destroy.restype = None
            self.lib.dll.tc_pipeline_destroy(self.handle)
            self.handle = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def reserve(self, B, terms_per_traj, row_cap, val_cap, alloc=None):
        alloc = alloc or (lambda n, dt: np.zeros(n, dt))
        self.out = dict(term_ptr=alloc(B * terms_per_traj + 1, np.int64), val=alloc(val_cap, np.float64),
                        value=alloc(row_cap, np.float64), pair=alloc(row_cap, np.int32), cost=alloc(B, np.float64))

    def pair_masks(self, n_pairs):
        """tc_scene_pair_dof_mask of every link pair"""
        out = np.zeros(n_pairs, np.uint64)
        m = C.c_uint64()
        for p in range(n_pairs):
            self.lib.check(self.lib.dll.tc_scene_pair_dof_mask(self.scene.handle, C.c_int32(p), C.byref(m)))
            out[p] = m.value
        return out

    def eval_rows(self, q, evaluator, lvs, fixed, chunk, stack=STACK_IFOPT, precision=PRECISION_F64, packed=False):
        """q: numpy [B, T, n_dof] (host; pinned for asynchronous reads) -> (n_rows, nnz); arrays in self.out.
        packed: values of the joints that move a row's links only (unpack_rows restores the dense blocks)"""
        B, T, n = q.shape
        d, keep = make_eval_desc(evaluator, stack, lvs, fixed, WANT_LINEARISE, precision)
        o = self.out
        n_rows, nnz = C.c_int64(), C.c_int64()
        fn = self.lib.dll.tc_pipeline_eval_rows_packed if packed else self.lib.dll.tc_pipeline_eval_rows
        self.lib.check(fn(
            self.handle, C.byref(d), C.c_void_p(q.ctypes.data), C.c_int32(B), C.c_int32(T), C.c_int32(chunk),
            C.c_void_p(o["term_ptr"].ctypes.data), C.c_void_p(o["val"].ctypes.data), C.c_size_t(len(o["val"])),
            C.c_void_p(o["value"].ctypes.data), C.c_void_p(o["pair"].ctypes.data), C.c_size_t(len(o["value"])),
            C.c_void_p(o["cost"].ctypes.data), C.byref(n_rows), C.byref(nnz)))
        return n_rows.value, nnz.value


class CudaLib(Lib):
    """libtrajopt_cuda.so"""

    def __init__(self, path=LIB_PATH, device=0):
        super().__init__(path, "tc_")
        self.check(self.dll.tc_init(C.c_int(device)))

    def new_result(self, scene):
        h = C.c_void_p()
        self.check(self.dll.tc_result_create(scene.handle, C.byref(h)))
        return h

    def contact_test(self, scene, pose0, pose1=None, contact_type=CONTACT_ALL, precision=PRECISION_F64, result=None):
        """seam B3 (tc_contact_test): pose0 / pose1 numpy [n_states, n_active_links, 12] -> Result (one term per state)"""
        pose0 = np.ascontiguousarray(pose0, dtype=np.float64)
        n_states = pose0.shape[0]
        assert pose0.shape[1:] == (scene.n_active, 12)
        p1 = None
        if pose1 is not None:
            p1 = np.ascontiguousarray(pose1, dtype=np.float64)
            assert p1.shape == pose0.shape
        if result is None:
            result = Result(self, scene, self.new_result(scene), n_states, 2 if p1 is not None else 1)
        result.B, result.T = n_states, (2 if p1 is not None else 1)
        self.check(self.dll.tc_contact_test(scene.handle, _p(pose0, _pd), _p(p1, _pd) if p1 is not None else _pd(),
                                            C.c_int32(n_states), C.c_int32(contact_type), C.c_int32(precision),
                                            result.handle, C.c_void_p(0)))
        return result

    def eval(self, scene, q, evaluator, stack=STACK_IFOPT, lvs=0.005, fixed=None, want=WANT_ALL,
             precision=PRECISION_F64, result=None, stream=None, contact_type=CONTACT_ALL):
        """q: numpy [B, T, n_dof] (host) -> Result"""
        q = np.ascontiguousarray(q, dtype=np.float64)
        B, T, n = q.shape
        assert n == scene.n_dof
        d, keep = make_eval_desc(evaluator, stack, lvs, fixed, want, precision, contact_type=contact_type)
        if result is None:
            result = Result(self, scene, self.new_result(scene), B, T)
        result.B, result.T = B, T
        self.check(self.dll.tc_eval(scene.handle, C.byref(d), _p(q, _pd), C.c_int32(B), C.c_int32(T), result.handle,
                                    C.c_void_p(stream or 0)))
        return result

    def eval_device(self, scene, q_dev_ptr, B, T, evaluator, result, stack=STACK_IFOPT, lvs=0.005, fixed=None,
                    want=WANT_ALL, precision=PRECISION_F64, stream=None, sync=True):
        """q already resident on the device (raw pointer); optionally asynchronous."""
        d, keep = make_eval_desc(evaluator, stack, lvs, fixed, want, precision, q_on_device=True)
        f = self.dll.tc_eval if sync else self.dll.tc_eval_async
        result.B, result.T = B, T
        self.check(f(scene.handle, C.byref(d), C.c_void_p(q_dev_ptr), C.c_int32(B), C.c_int32(T), result.handle,
                     C.c_void_p(stream or 0)))
        return result

    def sync(self, result):
        self.check(self.dll.tc_result_sync(result.handle))

    def pass_info(self, result):
        """(speculative, repeated) of the last pass on `result` (tc_result_pass_info)"""
        a, b = C.c_int32(), C.c_int32()
        self.check(self.dll.tc_result_pass_info(result.handle, C.byref(a), C.byref(b)))
        return bool(a.value), b.value

    def stage_ms(self, result):
        ms = (C.c_float * 10)()
        self.check(self.dll.tc_result_stage_ms(result.handle, ms))
        return list(ms)

    def build_rows_device(self, result):
        """assemble the rows on the device, leave them there; returns nnz"""
        self.check(self.dll.tc_result_rows_device(result.handle, None, None, None, None, None, None))
        return result.counts()["nnz"]

    def launches(self, result):
        a, b = C.c_int32(), C.c_int32()
        self.check(self.dll.tc_result_launches(result.handle, C.byref(a), C.byref(b)))
        return a.value, b.value

    def measure_fma_peak(self, precision=PRECISION_F64):
        v = C.c_double()
        self.check(self.dll.tc_measure_fma_peak(C.c_int(precision), C.byref(v)))
        return v.value

    def cost_noalloc(self, term_ptr_dev, n_terms, terms_per_traj, distance_dev, margin_dev, coeff_dev, out_dev,
                     stream=None):
        """tc_cost_noalloc: raw DEVICE pointers in and out, nothing allocated, nothing synchronised (the per-term
        intermediate sits in the library's static device array; calls that overlap on the GPU race)"""
        return self.dll.tc_cost_noalloc(C.c_void_p(term_ptr_dev), C.c_int32(n_terms), C.c_int32(terms_per_traj),
                                        C.c_void_p(distance_dev), C.c_void_p(margin_dev), C.c_void_p(coeff_dev),
                                        C.c_void_p(out_dev), C.c_void_p(stream or 0))
