"""Shared test helpers: scene fixtures, synthetic scenes, the host-compiled device numerics."""
import ctypes as C
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
import tcb200  # noqa: E402

pkg = tcb200.load()
capi = pkg.capi
SCENES = os.path.join(ROOT, "tests", "golden", "scenes")
_pd = C.POINTER(C.c_double)


def load_scene(name):
    return dict(np.load(os.path.join(SCENES, name + ".npz")))


def rand_pose(rng, scale=1.0):
    """random rigid pose as 12 doubles"""
    a = rng.normal(size=3)
    a /= np.linalg.norm(a)
    th = rng.uniform(-np.pi, np.pi)
    K = np.array([[0, -a[2], a[1]], [a[2], 0, -a[0]], [-a[1], a[0], 0]])
    R = np.eye(3) + np.sin(th) * K + (1 - np.cos(th)) * (K @ K)
    T = np.zeros((3, 4))
    T[:, :3] = R
    T[:, 3] = rng.uniform(-scale, scale, 3)
    return T.reshape(12)


def perturb_pose(rng, pose, dth, dt):
    T = pose.reshape(3, 4).copy()
    a = rng.normal(size=3)
    a /= np.linalg.norm(a)
    th = rng.uniform(-dth, dth)
    K = np.array([[0, -a[2], a[1]], [a[2], 0, -a[0]], [-a[1], a[0], 0]])
    R = np.eye(3) + np.sin(th) * K + (1 - np.cos(th)) * (K @ K)
    T[:, :3] = R @ T[:, :3]
    T[:, 3] += rng.uniform(-dt, dt, 3)
    return T.reshape(12)


class HostCheck(capi.Lib):
    """tests/hostcheck/libhostcheck.so: the kernels' per-thread numerics compiled for the host."""

    def __init__(self):
        here = os.path.join(ROOT, "tests", "hostcheck")
        subprocess.check_call(["make", "-C", here, "-s"])
        super().__init__(os.path.join(here, "libhostcheck.so"), "hc_")

    def shape_distance(self, scene, sa, pose_a0, sb, pose_b0, threshold, pose_a1=None, pose_b1=None,
                       precision=capi.PRECISION_F64, max_verts=259, max_faces=520):
        def ptr(a):
            if a is None:
                return None, _pd()
            a = np.ascontiguousarray(a, dtype=np.float64).reshape(12)
            return a, capi._p(a, _pd)

        k0, a0 = ptr(pose_a0)
        k1, a1 = ptr(pose_a1)
        k2, b0 = ptr(pose_b0)
        k3, b1 = ptr(pose_b1)
        status, epa, gi, ei, sc = C.c_int(), C.c_int(), C.c_int(), C.c_int(), C.c_int()
        dist = C.c_double()
        n, pa, pb = np.zeros(3), np.zeros(3), np.zeros(3)
        self.check(self.dll.hc_shape_distance(scene.handle, C.c_int(precision), C.c_int(sa), a0, a1, C.c_int(sb), b0,
                                              b1, C.c_double(threshold), C.c_int(max_verts), C.c_int(max_faces),
                                              C.byref(status), C.byref(dist), capi._p(n, _pd), capi._p(pa, _pd),
                                              capi._p(pb, _pd), C.byref(epa), C.byref(gi), C.byref(ei), C.byref(sc)))
        return dict(status=status.value, contact=status.value == 1, distance=dist.value, normal=n, pa=pa, pb=pb,
                    epa=bool(epa.value), iters=gi.value, epa_iters=ei.value, scanned=sc.value)
