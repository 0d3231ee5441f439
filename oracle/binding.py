"""TEST INFRASTRUCTURE ONLY: ctypes bindings of the two CPU checkers.

  * `Oracle`     - oracle/liboracle.so, the plain-C restatement (oracle/vns_oracle.c)
  * `RefOracle`  - oracle/_ref/liboracle_ref.so, the UNMODIFIED reference code compiled in place
                   by oracle/ref/build_ref.sh (op::VnsSopPath / op::VnsSopSolver)

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import
this module; the product (vns-sop_b200/) never does.
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(HERE, "liboracle.so")
REF_LIB = os.path.join(HERE, "_ref", "liboracle_ref.so")
REF_BIN = os.path.join(HERE, "_ref", "sop_vns_ref")

c_int_p = C.POINTER(C.c_int)
c_dbl_p = C.POINTER(C.c_double)


def build(force=False):
    """gcc the C restatement; build oracle/_ref when the reference tree is present."""
    src = os.path.join(HERE, "vns_oracle.c")
    if force or not os.path.exists(LIB) or os.path.getmtime(LIB) < os.path.getmtime(src):
        subprocess.check_call(["gcc", "-std=gnu99", "-O2", "-fPIC", "-shared", "-ffp-contract=off",
                               src, "-o", LIB, "-lm"])
    ref_src = os.environ.get("REF", "/root/reference")
    if os.path.isdir(os.path.join(ref_src, "sources")) and (force or not os.path.exists(REF_LIB)):
        subprocess.check_call(["bash", os.path.join(HERE, "ref", "build_ref.sh")])


def _ip(a):
    return a.ctypes.data_as(c_int_p)


def _dp(a):
    return a.ctypes.data_as(c_dbl_p)


class VoRng(C.Structure):
    _fields_ = [("r", C.c_int32 * 34), ("f", C.c_int), ("b", C.c_int)]


class VoStats(C.Structure):
    _fields_ = [("proposals", C.c_long), ("lb_skips", C.c_long), ("evals", C.c_long), ("accepts", C.c_long)]


class VoParams(C.Structure):
    _fields_ = [("num_iterations", C.c_int), ("num_iterations_unimproved", C.c_int), ("skip", C.c_int),
                ("greedy", C.c_int), ("budget", C.c_double), ("seed", C.c_uint)]


class VoResult(C.Structure):
    _fields_ = [("num_iters", C.c_int), ("num_iters_last_impr", C.c_int), ("n_improvements", C.c_int),
                ("reward", C.c_double), ("length", C.c_double),
                ("impr_reward", C.c_double * 4096), ("impr_length", C.c_double * 4096)]


class VoInst(C.Structure):
    _fields_ = [("C", C.c_int), ("N", C.c_int), ("mp", C.c_int), ("start", C.c_int), ("goal", C.c_int),
                ("m", c_int_p), ("off", c_int_p), ("reward", c_dbl_p), ("D", c_dbl_p),
                ("dmin", c_dbl_p), ("dmax", c_dbl_p)]


class VoPath(C.Structure):
    _fields_ = [("I", C.POINTER(VoInst)), ("T", C.c_int), ("targets", c_int_p), ("nvns", C.c_int),
                ("vns", c_int_p), ("F", c_dbl_p), ("B", c_dbl_p), ("pF", c_int_p), ("pB", c_int_p)]


def _load_lib():
    build()
    lib = C.CDLL(LIB)
    P, I = C.POINTER(VoPath), C.POINTER(VoInst)
    lib.vo_inst_new.restype = I
    lib.vo_inst_new.argtypes = [C.c_int, c_int_p, c_dbl_p, C.c_int, C.c_int]
    lib.vo_path_new.restype = P
    lib.vo_path_new.argtypes = [I]
    lib.vo_dubins_length.restype = C.c_double
    lib.vo_dubins_length.argtypes = [C.c_double] * 7
    for n in ("vo_try_add", "vo_try_add_lb", "vo_try_replace", "vo_try_replace_lb", "vo_try_move",
              "vo_try_move_lb", "vo_try_exchange", "vo_try_exchange_lb"):
        getattr(lib, n).restype = C.c_double
        getattr(lib, n).argtypes = [P, C.c_int, C.c_int]
    lib.vo_try_remove.restype = C.c_double
    lib.vo_try_remove.argtypes = [P, C.c_int]
    for n in ("vo_length", "vo_length_rear", "vo_length_calc", "vo_reward"):
        getattr(lib, n).restype = C.c_double
        getattr(lib, n).argtypes = [P]
    return lib


class Oracle:
    """The C restatement on one Problem (vns_sop_b200.instances.Problem)."""
    _lib = None

    def __init__(self, pb, table="auto", dubins=None):
        if Oracle._lib is None:
            Oracle._lib = _load_lib()
        lib = self.lib = Oracle._lib
        self.pb = pb
        m = np.ascontiguousarray(pb.m, dtype=np.int32)
        rw = np.ascontiguousarray(pb.reward, dtype=np.float64)
        self.I = lib.vo_inst_new(pb.C, _ip(m), _dp(rw), pb.start, pb.goal)
        if dubins is not None:                       # (S, rho, do_ceil[, eps])
            ctr = np.ascontiguousarray(pb.centers, dtype=np.float64)
            C.c_double.in_dll(lib, "vo_dubins_eps").value = dubins[3] if len(dubins) > 3 else 1e-9
            lib.vo_table_dubins(self.I, _dp(ctr), C.c_int(dubins[0]), C.c_double(dubins[1]), C.c_int(dubins[2]))
        elif pb.xy is not None:
            xy = np.ascontiguousarray(pb.xy, dtype=np.float64)
            lib.vo_table_euclid(self.I, _dp(xy), pb.rule)
        else:
            W = np.ascontiguousarray(pb.W, dtype=np.float64)
            lib.vo_table_dense(self.I, _dp(W))
        lib.vo_bounds(self.I)
        self.budget = pb.budget

    @property
    def mp(self):
        return self.I.contents.mp

    def dense(self):
        W = np.empty((self.pb.N, self.pb.N))
        self.lib.vo_get_dense(self.I, _dp(W))
        return W

    def bounds(self):
        n = self.pb.C * self.pb.C
        lo = np.ctypeslib.as_array(self.I.contents.dmin, (n,)).reshape(self.pb.C, -1).copy()
        hi = np.ctypeslib.as_array(self.I.contents.dmax, (n,)).reshape(self.pb.C, -1).copy()
        return lo, hi

    def path(self, vns=None, T=0):
        p = OraclePath(self)
        if vns is None:
            vns = self.pb.initial_vns()
        p.set(vns, T)
        return p


class OraclePath:
    def __init__(self, o):
        self.o, self.lib = o, o.lib
        self.p = o.lib.vo_path_new(o.I)

    def set(self, vns, T):
        v = np.ascontiguousarray(vns, dtype=np.int32)
        self.lib.vo_path_set(self.p, _ip(v), len(v), int(T))

    @property
    def T(self):
        return self.p.contents.T

    @property
    def vns(self):
        return np.ctypeslib.as_array(self.p.contents.vns, (self.p.contents.nvns,)).copy()

    @property
    def targets(self):
        return np.ctypeslib.as_array(self.p.contents.targets, (max(self.T, 1),))[:self.T].copy()

    def layers(self):
        """(F, B, pF, pB) as [(T+2), mp] arrays."""
        rows, mp = self.T + 2, self.o.mp
        c = self.p.contents
        return tuple(np.ctypeslib.as_array(a, (rows * mp,)).reshape(rows, mp).copy()
                     for a in (c.F, c.B, c.pF, c.pB))

    def length(self):
        return self.lib.vo_length(self.p)

    def length_rear(self):
        return self.lib.vo_length_rear(self.p)

    def length_calc(self):
        return self.lib.vo_length_calc(self.p)

    def reward(self):
        return self.lib.vo_reward(self.p)

    def get_path(self):
        n = self.T + 2
        cl, nd = np.zeros(n, np.int32), np.zeros(n, np.int32)
        self.lib.vo_get_path(self.p, _ip(cl), _ip(nd))
        return cl, nd

    def eval_batch(self, cands):
        c = np.ascontiguousarray(cands, dtype=np.int32).reshape(-1, 3)
        ln, lb = np.empty(len(c)), np.empty(len(c))
        self.lib.vo_eval_batch(self.p, len(c), _ip(c), _dp(ln), _dp(lb))
        return ln, lb

    def add_point(self, c, idx):
        self.lib.vo_add_point(self.p, int(c), int(idx))

    def remove_point(self, idx):
        self.lib.vo_remove_point(self.p, int(idx))

    def local_search(self, kind, num_changes, rng, budget=None, skip=1):
        st = VoStats()
        fn = self.lib.vo_insert_random if kind == 1 else self.lib.vo_exchange_random
        fn(self.p, int(num_changes), C.c_double(self.o.budget if budget is None else budget), int(skip),
           C.byref(rng), C.byref(st))
        return st

    def shake(self, kind, rng, budget=None):
        fn = self.lib.vo_path_insert if kind == 1 else self.lib.vo_path_exchange
        fn(self.p, C.c_double(self.o.budget if budget is None else budget), C.byref(rng))

    def fit(self, budget=None):
        self.lib.vo_fit(self.p, C.c_double(self.o.budget if budget is None else budget))

    def greedy(self, budget=None):
        self.lib.vo_greedy(self.p, C.c_double(self.o.budget if budget is None else budget))

    def solve(self, seed, num_iterations=2000, num_iterations_unimproved=1000, skip=1, greedy=1, budget=None):
        rng = VoRng()
        self.lib.vo_srand(C.byref(rng), C.c_uint(seed))
        prm = VoParams(num_iterations, num_iterations_unimproved, skip, greedy,
                       self.o.budget if budget is None else budget, seed)
        res = VoResult()
        self.lib.vo_solve(self.p, C.byref(prm), C.byref(rng), C.byref(res))
        return res


def new_rng(seed):
    if Oracle._lib is None:
        Oracle._lib = _load_lib()
    g = VoRng()
    Oracle._lib.vo_srand(C.byref(g), C.c_uint(seed))
    return g


def rng_next(g):
    return Oracle._lib.vo_rand(C.byref(g))


class RefOracle:
    """The reference's own code (process-global statics: ONE instance at a time per process)."""
    _lib = None

    @staticmethod
    def available():
        return os.path.exists(REF_LIB)

    def __init__(self, sop_path, verbosity=0):
        if RefOracle._lib is None:
            lib = C.CDLL(REF_LIB)
            for n in ("oref_try_add", "oref_try_add_lb", "oref_try_replace", "oref_try_replace_lb",
                      "oref_try_move", "oref_try_move_lb", "oref_try_exchange", "oref_try_exchange_lb"):
                getattr(lib, n).restype = C.c_double
                getattr(lib, n).argtypes = [C.c_int, C.c_int, C.c_int]
            lib.oref_try_remove.restype = C.c_double
            for n in ("oref_length", "oref_length_rear", "oref_length_calc", "oref_reward", "oref_budget",
                      "oref_max_reward", "oref_cluster_reward", "oref_dmin", "oref_dmax"):
                getattr(lib, n).restype = C.c_double
            lib.oref_set_budget.argtypes = [C.c_double]
            RefOracle._lib = lib
        self.lib = RefOracle._lib
        self.lib.oref_load(sop_path.encode(), b"sop", verbosity)
        self.C = self.lib.oref_num_clusters()
        self.m = np.array([self.lib.oref_cluster_size(c) for c in range(self.C)], dtype=np.int32)
        self.budget = self.lib.oref_budget()

    def initial_vns(self):
        v = np.zeros(self.C, np.int32)
        n = self.lib.oref_initial_vns(_ip(v))
        return v[:n].copy()

    def block(self, c1, c2):
        out = np.empty(int(self.m[c1]) * int(self.m[c2]))
        self.lib.oref_block(c1, c2, _dp(out))
        return out.reshape(int(self.m[c1]), int(self.m[c2]))

    def dense(self):
        off = np.concatenate([[0], np.cumsum(self.m)])
        W = np.empty((off[-1], off[-1]))
        for a in range(self.C):
            for b in range(self.C):
                W[off[a]:off[a + 1], off[b]:off[b + 1]] = self.block(a, b)
        return W

    def bounds(self):
        lo = np.array([[self.lib.oref_dmin(a, b) for b in range(self.C)] for a in range(self.C)])
        hi = np.array([[self.lib.oref_dmax(a, b) for b in range(self.C)] for a in range(self.C)])
        return lo, hi

    def seed(self, s):
        self.lib.oref_seed(C.c_uint(s))

    def rand(self):
        return self.lib.oref_rand()

    def path(self, vns=None, T=0):
        p = RefPath(self)
        if vns is None:
            vns = self.initial_vns()
        p.set(vns, T)
        return p


class RefPath:
    def __init__(self, r):
        self.r, self.lib = r, r.lib
        self.h = r.lib.oref_path_new()
        self._vns = None

    def set(self, vns, T):
        self._vns = np.ascontiguousarray(vns, dtype=np.int32).copy()
        t = np.ascontiguousarray(self._vns[:T])
        self.lib.oref_path_set_targets(self.h, _ip(t), int(T))

    @property
    def T(self):
        return self.lib.oref_path_num_targets(self.h)

    @property
    def vns(self):
        return self._vns.copy()

    @property
    def targets(self):
        t = np.zeros(max(self.T, 1), np.int32)
        self.lib.oref_path_get_targets(self.h, _ip(t))
        return t[:self.T]

    def layers(self, mp):
        rows = self.T + 2
        F, B = np.zeros((rows, mp)), np.zeros((rows, mp))
        pF, pB = np.zeros((rows, mp), np.int32), np.zeros((rows, mp), np.int32)
        for l in range(rows):
            d, q = np.zeros(mp), np.zeros(mp, np.int32)
            n = self.lib.oref_get_layer(self.h, l, 0, _dp(d), _ip(q))
            F[l, :n], pF[l, :n] = d[:n], q[:n]
            n = self.lib.oref_get_layer(self.h, l, 1, _dp(d), _ip(q))
            B[l, :n], pB[l, :n] = d[:n], q[:n]
        return F, B, pF, pB

    def length(self):
        return self.lib.oref_length(self.h)

    def length_rear(self):
        return self.lib.oref_length_rear(self.h)

    def length_calc(self):
        return self.lib.oref_length_calc(self.h)

    def reward(self):
        return self.lib.oref_reward(self.h)

    def get_path(self):
        n = self.T + 2
        cl, nd = np.zeros(n, np.int32), np.zeros(n, np.int32)
        self.lib.oref_get_path(self.h, _ip(cl), _ip(nd))
        return cl, nd

    def eval_batch(self, cands, want_lb=True):
        c = np.ascontiguousarray(cands, dtype=np.int32).reshape(-1, 3)
        ln, lb = np.empty(len(c)), np.empty(len(c))
        self.lib.oref_eval_batch(self.h, len(c), _ip(c), _dp(ln), _dp(lb) if want_lb else None)
        return ln, lb

    def add_point(self, c, idx):
        self.lib.oref_add_point(self.h, int(c), int(idx))

    def remove_point(self, idx):
        self.lib.oref_remove_point(self.h, int(idx))

    def _drv(self, fn, *args):
        v = self._vns
        fn(self.h, _ip(v), len(v), *args)

    def local_search(self, kind, num_changes, skip=1):
        self.lib.oref_set_skip(int(skip))
        self._drv(self.lib.oref_insert_random if kind == 1 else self.lib.oref_exchange_random, int(num_changes))

    def shake(self, kind):
        self._drv(self.lib.oref_path_insert if kind == 1 else self.lib.oref_path_exchange)

    def fit(self):
        self._drv(self.lib.oref_fit)

    def greedy(self):
        self._drv(self.lib.oref_greedy)
