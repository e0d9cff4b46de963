"""TEST INFRASTRUCTURE ONLY — ctypes driver of the CPU oracle (oracle/liboracle.so).

May be imported from tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs only.  It reuses the struct layouts of the product's ctypes view (the oracle shares the C structs
of include/trajopt_cuda.h) but loads a different library with an independent implementation."""
import ctypes as C
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
import tcb200  # noqa: E402

capi = tcb200.load().capi
HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(HERE, "liboracle.so")
_pd = C.POINTER(C.c_double)


def build():
    subprocess.check_call(["make", "-C", HERE, "-s"])


class OracleLib(capi.Lib):
    def __init__(self):
        if not os.path.exists(LIB):
            build()
        super().__init__(LIB, "orc_")

    def contact_test(self, scene, pose0, pose1=None, contact_type=capi.CONTACT_ALL):
        pose0 = np.ascontiguousarray(pose0, dtype=np.float64)
        p1 = np.ascontiguousarray(pose1, dtype=np.float64) if pose1 is not None else None
        h = C.c_void_p()
        self.check(self.dll.orc_contact_test(scene.handle, capi._p(pose0, _pd), capi._p(p1, _pd) if p1 is not None else _pd(),
                                             C.c_int32(pose0.shape[0]), C.c_int32(contact_type), C.byref(h)))
        return capi.Result(self, scene, h, pose0.shape[0], 2 if p1 is not None else 1)

    def eval(self, scene, q, evaluator, stack=capi.STACK_IFOPT, lvs=0.005, fixed=None, want=capi.WANT_ALL,
             n_threads=1, contact_type=capi.CONTACT_ALL):
        q = np.ascontiguousarray(q, dtype=np.float64)
        B, T, n = q.shape
        assert n == scene.n_dof
        d, keep = capi.make_eval_desc(evaluator, stack, lvs, fixed, want, contact_type=contact_type)
        h = C.c_void_p()
        self.check(self.dll.orc_eval(scene.handle, C.byref(d), capi._p(q, _pd), C.c_int32(B), C.c_int32(T),
                                     C.byref(h), C.c_int(n_threads)))
        return capi.Result(self, scene, h, B, T)

    def num_threads(self):
        return int(self.dll.orc_num_threads())

    def fk(self, scene, q):
        out = np.zeros((scene.n_active, 12))
        q = np.ascontiguousarray(q, dtype=np.float64)
        self.check(self.dll.orc_fk(scene.handle, capi._p(q, _pd), capi._p(out, _pd)))
        return out

    def jacobian(self, scene, q, link, ref_offset):
        J = np.zeros((3, scene.n_dof))
        q = np.ascontiguousarray(q, dtype=np.float64)
        r = np.ascontiguousarray(ref_offset, dtype=np.float64)
        self.check(self.dll.orc_jacobian(scene.handle, capi._p(q, _pd), C.c_int(link), capi._p(r, _pd),
                                         capi._p(J, _pd)))
        return J

    def shape_distance(self, scene, sa, pose_a0, sb, pose_b0, threshold, pose_a1=None, pose_b1=None):
        def ptr(a):
            if a is None:
                return None, _pd()
            a = np.ascontiguousarray(a, dtype=np.float64).reshape(12)
            return a, capi._p(a, _pd)

        k0, a0 = ptr(pose_a0)
        k1, a1 = ptr(pose_a1)
        k2, b0 = ptr(pose_b0)
        k3, b1 = ptr(pose_b1)
        contact, epa, iters = C.c_int(), C.c_int(), C.c_int()
        dist = C.c_double()
        n, pa, pb = np.zeros(3), np.zeros(3), np.zeros(3)
        self.check(self.dll.orc_shape_distance(scene.handle, C.c_int(sa), a0, a1, C.c_int(sb), b0, b1,
                                               C.c_double(threshold), C.byref(contact), C.byref(dist),
                                               capi._p(n, _pd), capi._p(pa, _pd), capi._p(pb, _pd), C.byref(epa),
                                               C.byref(iters)))
        return dict(contact=bool(contact.value), distance=dist.value, normal=n, pa=pa, pb=pb, epa=bool(epa.value),
                    iters=iters.value)

    def shape_support(self, scene, shape, pose0, direction, pose1=None):
        p0 = np.ascontiguousarray(pose0, dtype=np.float64).reshape(12)
        p1 = None if pose1 is None else np.ascontiguousarray(pose1, dtype=np.float64).reshape(12)
        d = np.ascontiguousarray(direction, dtype=np.float64)
        v = C.c_double()
        self.check(self.dll.orc_shape_support(scene.handle, C.c_int(shape), capi._p(p0, _pd),
                                              capi._p(p1, _pd) if p1 is not None else _pd(), capi._p(d, _pd),
                                              C.byref(v)))
        return v.value
