"""ctypes binding of libvnsgpu.so (include/vnsgpu.h) + a host-side mirror of the reference's
op::VnsSopPath interface (sources/vns_sop_path.h:22-194): same method names, argument meaning and
range errors, every call routed to the CUDA engine.  There is NO CPU fallback: a missing library or
a missing GPU raises."""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

from . import build as _build

OP_ADD, OP_REPLACE, OP_REMOVE, OP_MOVE, OP_EXCHANGE = range(5)
F_FEASIBLE, F_ACCEPT, F_LB_SKIP, F_BAD = 1, 2, 4, 8
NO_KEY = 0x7f7f7f7f7f7f7f7f      # VNSGPU_NO_KEY
CAND_DTYPE = np.dtype([("op", np.int32), ("a", np.int32), ("b", np.int32), ("gain_cluster", np.int32)])

_STATUS = {1: "E_RANGE", 2: "E_CUDA", 3: "E_STATE", 4: "E_INCONSISTENT", 5: "E_NOMEM", 6: "E_TOO_LARGE", 7: "E_NCCL"}


class VnsGpuError(RuntimeError):
    def __init__(self, status, msg):
        super().__init__("vnsgpu %s: %s" % (_STATUS.get(status, status), msg))
        self.status = status


class Rng(C.Structure):
    """glibc rand() state (vnsgpu_rng)."""
    _fields_ = [("r", C.c_int32 * 31), ("f", C.c_int32), ("b", C.c_int32), ("pad", C.c_int32)]


class LsStats(C.Structure):
    _fields_ = [(n, C.c_int64) for n in ("proposals", "lb_skips", "evals", "accepts", "rounds")]


class BestMove(C.Structure):
    """vnsgpu_best_move_t"""
    _fields_ = [("index", C.c_int64), ("from_", C.c_int32), ("to", C.c_int32), ("kind", C.c_int32), ("pad", C.c_int32),
                ("added_reward", C.c_double), ("length", C.c_double), ("n_candidates", C.c_int64)]


class VnsParams(C.Structure):
    _fields_ = [("num_iterations", C.c_int32), ("num_iterations_unimproved", C.c_int32),
                ("lower_bound_skip", C.c_int32), ("greedy_initial", C.c_int32), ("budget", C.c_double),
                ("max_time_ms", C.c_double), ("seed", C.c_uint32), ("max_improvements", C.c_int32)]


class VnsResult(C.Structure):
    _fields_ = [("num_iters", C.c_int32), ("num_iters_last_impr", C.c_int32), ("n_improvements", C.c_int32),
                ("status", C.c_int32), ("reward", C.c_double), ("length", C.c_double), ("time_ms", C.c_double),
                ("time_last_impr_ms", C.c_double), ("proposals", C.c_int64), ("evals", C.c_int64),
                ("accepts", C.c_int64)]


_lib = None
vp, ip, dp = C.c_void_p, C.POINTER(C.c_int), C.POINTER(C.c_double)

_SIGS = {
    "vnsgpu_create": [C.POINTER(vp), C.c_int],
    "vnsgpu_set_stream": [vp, vp],
    "vnsgpu_sync": [vp],
    "vnsgpu_set_int_path": [vp, C.c_int],
    "vnsgpu_set_batch_sort": [vp, C.c_int],
    "vnsgpu_set_u16_table": [vp, C.c_int],
    "vnsgpu_set_segment_cache": [vp, C.c_int],
    "vnsgpu_set_smem_chain": [vp, C.c_int],
    "vnsgpu_set_chain_form": [vp, C.c_int],
    "vnsgpu_set_chain_blocks": [vp, C.c_int],
    "vnsgpu_set_chain_tree": [vp, C.c_int],
    "vnsgpu_measure_read_peaks": [vp, dp],
    "vnsgpu_comm_unique_id": [vp, vp],
    "vnsgpu_comm_init": [vp, vp, C.c_int, C.c_int],
    "vnsgpu_comm_adopt": [vp, vp, C.c_int, C.c_int],
    "vnsgpu_comm_destroy": [vp],
    "vnsgpu_comm_info": [vp, ip, ip],
    "vnsgpu_reduce_first_key": [vp, vp, vp],
    "vnsgpu_path_broadcast": [vp, C.c_int, C.c_int, vp],
    "vnsgpu_bcast_bytes": [vp, vp, C.c_int64, C.c_int, vp],
    "vnsgpu_gen_proposals_dev": [vp, C.c_int, C.c_int, C.c_int64, C.POINTER(Rng), vp, C.c_int, vp, vp],
    "vnsgpu_propose_eval": [vp, C.c_int, C.c_int, C.c_int64, C.POINTER(Rng), C.c_double, C.c_int, C.POINTER(C.c_int64), vp],
    "vnsgpu_set_phase_timing": [vp, C.c_int],
    "vnsgpu_phase_times": [vp, dp, C.POINTER(C.c_int64)],
    "vnsgpu_set_target_reward": [vp, C.c_double],
    "vnsgpu_chain_cycles": [vp, C.POINTER(C.c_int64), C.c_int],
    "vnsgpu_int_path_active": [vp],
    "vnsgpu_set_clusters": [vp, C.c_int, ip, dp, C.c_int, C.c_int],
    "vnsgpu_build_table_euclid": [vp, dp, C.c_int],
    "vnsgpu_build_table_dubins": [vp, dp, C.c_int, C.c_double, C.c_int],
    "vnsgpu_set_table_dense": [vp, dp],
    "vnsgpu_get_table_dense": [vp, dp],
    "vnsgpu_get_bounds": [vp, dp, dp],
    "vnsgpu_path_create": [vp, ip],
    "vnsgpu_path_copy": [vp, C.c_int, C.c_int],
    "vnsgpu_path_set": [vp, C.c_int, ip, C.c_int, C.c_int],
    "vnsgpu_path_add_point": [vp, C.c_int, C.c_int, C.c_int],
    "vnsgpu_path_remove_point": [vp, C.c_int, C.c_int],
    "vnsgpu_path_get": [vp, C.c_int, ip, ip, ip, ip, dp, dp, dp],
    "vnsgpu_path_get_solution": [vp, C.c_int, ip, ip, dp],
    "vnsgpu_path_get_dp": [vp, C.c_int, dp, dp, C.c_int],
    "vnsgpu_eval_batch": [vp, C.c_int, C.c_int64, vp, C.c_double, C.c_int, dp, dp, vp, C.POINTER(C.c_int64)],
    "vnsgpu_eval_batch_noalloc": [vp, C.c_int, C.c_int64, vp, C.c_double, C.c_int, dp, dp, vp, C.POINTER(C.c_int64)],
    "vnsgpu_eval_batch_packed": [vp, C.c_int, C.c_int64, vp, C.c_double, C.c_int, dp, dp, vp, C.POINTER(C.c_int64)],
    "vnsgpu_eval_batch_packed32": [vp, C.c_int, C.c_int64, vp, C.c_double, C.c_int, vp, vp, C.POINTER(C.c_int64)],
    "vnsgpu_eval_batch_packed32r": [vp, C.c_int, C.c_int64, vp, C.c_double, C.c_int, vp, C.POINTER(C.c_int64)],
    "vnsgpu_eval_batch_dev": [vp, C.c_int, C.c_int64, vp, C.c_double, C.c_int, vp, vp, vp, vp, vp],
    "vnsgpu_eval_batch_dev_shard": [vp, C.c_int, C.c_int64, vp, C.c_double, C.c_int, vp, vp, vp, vp, C.c_int64, vp],
    "vnsgpu_gen_proposals": [vp, C.c_int, C.c_int, C.c_int64, C.POINTER(Rng), vp],
    "vnsgpu_cand_bytes": [vp, C.c_int, C.c_int64, vp, C.POINTER(C.c_int64)],
    "vnsgpu_local_search": [vp, C.c_int, C.c_int, C.c_int, C.c_double, C.c_int, C.POINTER(Rng), C.POINTER(LsStats)],
    "vnsgpu_shake": [vp, C.c_int, C.c_int, C.c_double, C.POINTER(Rng)],
    "vnsgpu_fit_to_budget": [vp, C.c_int, C.c_double],
    "vnsgpu_greedy": [vp, C.c_int, C.c_double],
    "vnsgpu_greedy_batched": [vp, C.c_int, C.c_double, C.POINTER(C.c_int), C.POINTER(C.c_int64)],
    "vnsgpu_set_shard_neighbourhoods": [vp, C.c_int],
    "vnsgpu_best_move": [vp, C.c_int, C.c_int, C.c_double, C.POINTER(BestMove)],
    "vnsgpu_apply_move": [vp, C.c_int, C.c_int, C.c_int, C.c_int],
    "vnsgpu_best_improvement": [vp, C.c_int, C.c_double, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int64)],
    "vnsgpu_vns_run": [vp, C.c_int, C.POINTER(VnsParams), C.POINTER(VnsResult), dp, dp, dp],
    "vnsgpu_vns_run_batch": [vp, C.c_int, C.c_int, C.POINTER(C.c_uint32), C.POINTER(VnsParams), C.POINTER(VnsResult), ip, ip],
    "vnsgpu_path_destroy": [vp, C.c_int],
    "vnsgpu_set_vector_cache": [vp, C.c_int],
}
EXPORTS = sorted(list(_SIGS) + ["vnsgpu_noalloc_max_cands", "vnsgpu_rng_advance", "vnsgpu_eval_mode", "vnsgpu_destroy", "vnsgpu_last_error", "vnsgpu_version", "vnsgpu_launch_count",
                                "vnsgpu_table_bytes", "vnsgpu_table_kernel_ms", "vnsgpu_rng_seed", "vnsgpu_rng_next"])


def load_library(build_if_stale=True):
    """dlopen libvnsgpu.so (building it in-tree first when the sources are newer)."""
    global _lib
    if _lib is not None:
        return _lib
    path = _build.LIB
    if build_if_stale and _build.stale() and os.path.exists("/usr/local/cuda/bin/nvcc"):
        _build.build()
    if not os.path.exists(path):
        raise VnsGpuError(3, "libvnsgpu.so not built (python -m vns_sop_b200.build); there is no CPU fallback")
    lib = C.CDLL(path)
    for name, args in _SIGS.items():
        fn = getattr(lib, name)
        fn.argtypes, fn.restype = args, C.c_int
    lib.vnsgpu_destroy.argtypes, lib.vnsgpu_destroy.restype = [vp], None
    lib.vnsgpu_last_error.argtypes, lib.vnsgpu_last_error.restype = [vp], C.c_char_p
    lib.vnsgpu_version.argtypes, lib.vnsgpu_version.restype = [], C.c_char_p
    lib.vnsgpu_launch_count.argtypes, lib.vnsgpu_launch_count.restype = [vp], C.c_int64
    lib.vnsgpu_rng_advance.argtypes, lib.vnsgpu_rng_advance.restype = [C.POINTER(Rng), C.c_uint64], None
    lib.vnsgpu_eval_mode.argtypes, lib.vnsgpu_eval_mode.restype = [vp, C.c_int64], C.c_int
    lib.vnsgpu_noalloc_max_cands.argtypes, lib.vnsgpu_noalloc_max_cands.restype = [], C.c_int64
    lib.vnsgpu_table_bytes.argtypes, lib.vnsgpu_table_bytes.restype = [vp], C.c_int64
    lib.vnsgpu_table_kernel_ms.argtypes, lib.vnsgpu_table_kernel_ms.restype = [vp], C.c_double
    lib.vnsgpu_rng_seed.argtypes, lib.vnsgpu_rng_seed.restype = [C.POINTER(Rng), C.c_uint32], None
    lib.vnsgpu_rng_next.argtypes, lib.vnsgpu_rng_next.restype = [C.POINTER(Rng)], C.c_int32
    _lib = lib
    return lib


def seeded_rng(seed):
    """srand(seed) (the reference's randSeeded hook, sources/math_common.cpp:9-14)."""
    g = Rng()
    load_library().vnsgpu_rng_seed(C.byref(g), seed)
    return g


def rng_advance(rng, ndraws):
    """rand() called ndraws times (jump-ahead)"""
    load_library().vnsgpu_rng_advance(C.byref(rng), int(ndraws))


def rng_next(g):
    return load_library().vnsgpu_rng_next(C.byref(g))


def _i(a):
    return a.ctypes.data_as(ip)


def _d(a):
    return a.ctypes.data_as(dp)


class Engine:
    """One device context holding one instance: the distance table (VnsSopPath::allDistances,
    clustersMinDistances) lives in HBM for the context's lifetime."""

    def __init__(self, device=0):
        self.lib = load_library()
        self.h = vp()
        st = self.lib.vnsgpu_create(C.byref(self.h), device)
        if st:
            msg = self.lib.vnsgpu_last_error(self.h).decode() if self.h else "no CUDA device"
            raise VnsGpuError(st, msg)
        self.pb = None

    def close(self):
        if self.h:
            self.lib.vnsgpu_destroy(self.h)
            self.h = vp()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _ck(self, st):
        if st:
            raise VnsGpuError(st, self.lib.vnsgpu_last_error(self.h).decode())

    @property
    def launches(self):
        return self.lib.vnsgpu_launch_count(self.h)

    @property
    def table_bytes(self):
        return self.lib.vnsgpu_table_bytes(self.h)

    def set_stream(self, cuda_stream_ptr):
        self._ck(self.lib.vnsgpu_set_stream(self.h, cuda_stream_ptr))

    def sync(self):
        self._ck(self.lib.vnsgpu_sync(self.h))

    def set_int_path(self, enable):
        """False forces IEEE double evaluation even on all-integer tables (tests run both)."""
        self._ck(self.lib.vnsgpu_set_int_path(self.h, int(enable)))

    def set_segment_cache(self, enable):
        self._ck(self.lib.vnsgpu_set_segment_cache(self.h, int(enable)))

    def set_vector_cache(self, enable):
        self._ck(self.lib.vnsgpu_set_vector_cache(self.h, int(enable)))

    def chain_cycles(self, reset=True):
        out = (C.c_int64 * 4)()
        self._ck(self.lib.vnsgpu_chain_cycles(self.h, out, int(reset)))
        return dict(zip(("propose", "evaluate", "select", "apply"), out))

    def set_target_reward(self, target):
        """Race mode of vns_run_batch: every run stops once one run's tour reaches `target` (<= 0: off)."""
        self._ck(self.lib.vnsgpu_set_target_reward(self.h, float(target)))

    def set_smem_chain(self, enable):
        self._ck(self.lib.vnsgpu_set_smem_chain(self.h, int(enable)))

    def set_shard_neighbourhoods(self, enable):
        """full-neighbourhood modes split every neighbourhood over the communicator's ranks (collective calls)"""
        self._ck(self.lib.vnsgpu_set_shard_neighbourhoods(self.h, int(enable)))

    def set_chain_form(self, form):
        """0: second form (lower bound first, shake in shared memory, integer tables), 1: first form"""
        self._ck(self.lib.vnsgpu_set_chain_form(self.h, int(form)))

    def eval_mode(self, n):
        """how a batch of n candidates would be evaluated: {'form': 0|1|2, 'sorted': bool, 'hbm_form': bool, 'group_lanes': G}"""
        w = self.lib.vnsgpu_eval_mode(self.h, int(n))
        return {"form": w & 3, "sorted": bool(w & 16), "hbm_form": bool(w & 32), "group_lanes": (w >> 8) & 0xff}

    # ---- multi-GPU exchange steps (NCCL behind the C ABI) -------------------------------------------------------------
    def comm_unique_id(self):
        buf = (C.c_char * 128)()
        self._ck(self.lib.vnsgpu_comm_unique_id(self.h, buf))
        return bytes(buf)

    def comm_init(self, unique_id, rank, nranks):
        self._ck(self.lib.vnsgpu_comm_init(self.h, C.c_char_p(unique_id), rank, nranks))

    def comm_destroy(self):
        self._ck(self.lib.vnsgpu_comm_destroy(self.h))

    def reduce_first_key(self, d_key_ptr, stream=None):
        self._ck(self.lib.vnsgpu_reduce_first_key(self.h, C.c_void_p(d_key_ptr), C.c_void_p(stream) if stream else None))

    def measure_read_peaks(self):
        """{'l1': GB/s, 'l2': GB/s, 'hbm': GB/s}: streaming-read bandwidth with the data resident at that level"""
        g = (C.c_double * 3)()
        self._ck(self.lib.vnsgpu_measure_read_peaks(self.h, g))
        return {"l1": g[0], "l2": g[1], "hbm": g[2]}

    def set_phase_timing(self, enable):
        self._ck(self.lib.vnsgpu_set_phase_timing(self.h, int(enable)))

    def phase_times(self):
        """{phase: (ms, count)} of the batched evaluation call since the last query"""
        ms, cnt = (C.c_double * 4)(), (C.c_int64 * 4)()
        self._ck(self.lib.vnsgpu_phase_times(self.h, ms, cnt))
        return {k: (ms[i], cnt[i]) for i, k in enumerate(("seg", "vec", "sort", "eval"))}

    def set_chain_tree(self, mode):
        """product tree of the second chain form: -1 from the instance size, 0 off, 1 on"""
        self._ck(self.lib.vnsgpu_set_chain_tree(self.h, int(mode)))

    def set_chain_blocks(self, blocks):
        """thread blocks cooperating on one trajectory-exact chain (0: from the instance size)"""
        self._ck(self.lib.vnsgpu_set_chain_blocks(self.h, int(blocks)))

    def set_u16_table(self, enable):
        self._ck(self.lib.vnsgpu_set_u16_table(self.h, int(enable)))

    def set_batch_sort(self, enable):
        self._ck(self.lib.vnsgpu_set_batch_sort(self.h, int(enable)))

    @property
    def int_path_active(self):
        return bool(self.lib.vnsgpu_int_path_active(self.h))

    # ---- setSamplesDistances -----------------------------------------------------------------------
    def set_clusters(self, m, reward, start, goal):
        m = np.ascontiguousarray(m, np.int32)
        rw = np.ascontiguousarray(reward, np.float64)
        self.C, self.m, self.mp = len(m), m, int(m.max())
        self.N = int(m.sum())
        self._ck(self.lib.vnsgpu_set_clusters(self.h, len(m), _i(m), _d(rw), start, goal))

    def load_problem(self, pb, dubins=None):
        """Upload an instances.Problem: Euclidean tables are BUILT on the device, explicit ones
        gathered, `dubins=(S, rho, do_ceil)` builds the Dubins table from the cluster centres."""
        self.pb = pb
        self.set_clusters(pb.m, pb.reward, pb.start, pb.goal)
        if dubins is not None:
            ctr = np.ascontiguousarray(pb.centers, np.float64)
            self._ck(self.lib.vnsgpu_build_table_dubins(self.h, _d(ctr), dubins[0], dubins[1], dubins[2]))
        elif pb.xy is not None:
            xy = np.ascontiguousarray(pb.xy, np.float64)
            self._ck(self.lib.vnsgpu_build_table_euclid(self.h, _d(xy), pb.rule))
        else:
            W = np.ascontiguousarray(pb.W, np.float64)
            self._ck(self.lib.vnsgpu_set_table_dense(self.h, _d(W)))
        return self

    def build_table_euclid(self, xy, rule):
        xy = np.ascontiguousarray(xy, np.float64)
        self._ck(self.lib.vnsgpu_build_table_euclid(self.h, _d(xy), rule))

    def build_table_dubins(self, centers, S, rho, do_ceil=1):
        ctr = np.ascontiguousarray(centers, np.float64)
        self._ck(self.lib.vnsgpu_build_table_dubins(self.h, _d(ctr), S, rho, do_ceil))

    def table_dense(self):
        W = np.empty((self.N, self.N))
        self._ck(self.lib.vnsgpu_get_table_dense(self.h, _d(W)))
        return W

    def bounds(self):
        lo, hi = np.empty((self.C, self.C)), np.empty((self.C, self.C))
        self._ck(self.lib.vnsgpu_get_bounds(self.h, _d(lo), _d(hi)))
        return lo, hi

    def path(self, vns=None, T=0):
        p = VnsSopPath(self)
        if vns is None and self.pb is not None:
            vns = self.pb.initial_vns()
        if vns is not None:
            p.set(vns, T)
        return p


class VnsSopPath:
    """Device-resident path with the reference's method names (sources/vns_sop_path.h)."""

    def __init__(self, eng):
        self.e, self.lib = eng, eng.lib
        h = C.c_int(-1)
        eng._ck(self.lib.vnsgpu_path_create(eng.h, C.byref(h)))
        self.h = h.value

    def set(self, vns, T):
        v = np.ascontiguousarray(vns, np.int32)
        self.e._ck(self.lib.vnsgpu_path_set(self.e.h, self.h, _i(v), len(v), int(T)))

    def copy_from(self, other):
        self.e._ck(self.lib.vnsgpu_path_copy(self.e.h, self.h, other.h))

    def _get(self):
        T, n = C.c_int(), C.c_int()
        L, Lr, R = C.c_double(), C.c_double(), C.c_double()
        tg = np.zeros(2 * self.e.C + 2, np.int32)
        vns = np.zeros(2 * self.e.C + 2, np.int32)
        self.e._ck(self.lib.vnsgpu_path_get(self.e.h, self.h, C.byref(T), _i(tg), C.byref(n), _i(vns), C.byref(L),
                                            C.byref(Lr), C.byref(R)))
        return T.value, tg[:T.value].copy(), vns[:n.value].copy(), L.value, Lr.value, R.value

    def getNumTargets(self):
        return self._get()[0]

    def getAllTargets(self):
        return self._get()[1]

    @property
    def T(self):
        return self._get()[0]

    @property
    def vns(self):
        return self._get()[2]

    def getPathLength(self):
        return self._get()[3]

    def getPathLengthRear(self):
        return self._get()[4]

    def getReward(self):
        return self._get()[5]

    def getPath(self):
        """(cluster ids, node index inside the cluster, getPathLengthCalculate())."""
        n = self.getNumTargets() + 2
        cl, nd = np.zeros(n, np.int32), np.zeros(n, np.int32)
        calc = C.c_double()
        self.e._ck(self.lib.vnsgpu_path_get_solution(self.e.h, self.h, _i(cl), _i(nd), C.byref(calc)))
        return cl, nd, calc.value

    def getPathLengthCalculate(self):
        return self.getPath()[2]

    def layers(self):
        rows, mp = self.getNumTargets() + 2, self.e.mp
        F, B = np.zeros((rows, mp)), np.zeros((rows, mp))
        self.e._ck(self.lib.vnsgpu_path_get_dp(self.e.h, self.h, _d(F), _d(B), mp))
        return F, B

    def addPoint(self, cluster, idx):
        self.e._ck(self.lib.vnsgpu_path_add_point(self.e.h, self.h, int(cluster), int(idx)))

    def removePoint(self, idx):
        self.e._ck(self.lib.vnsgpu_path_remove_point(self.e.h, self.h, int(idx)))

    # ---- batched evaluation --------------------------------------------------------------------
    @staticmethod
    def make_cands(rows):
        """rows: (n,3) op,a,b (gain cluster derived: a for ADD/REPLACE) or a CAND_DTYPE array."""
        if isinstance(rows, np.ndarray) and rows.dtype == CAND_DTYPE:
            return np.ascontiguousarray(rows)
        r = np.asarray(rows, np.int32).reshape(-1, 3)
        c = np.zeros(len(r), CAND_DTYPE)
        c["op"], c["a"], c["b"] = r[:, 0], r[:, 1], r[:, 2]
        c["gain_cluster"] = np.where(r[:, 0] <= OP_REPLACE, r[:, 1], -1)
        return c

    def eval_batch(self, cands, budget=None, skip=1):
        """All tryTo* of `cands` at once -> (len, lower bound, flags, first accepted index)."""
        c = self.make_cands(cands)
        n = len(c)
        ln, lb, fl = np.empty(n), np.empty(n), np.zeros(n, np.uint8)
        first = C.c_int64(-1)
        budget = self.e.pb.budget if budget is None else budget
        self.e._ck(self.lib.vnsgpu_eval_batch(self.e.h, self.h, n, c.ctypes.data, budget, skip, _d(ln), _d(lb),
                                              fl.ctypes.data, C.byref(first)))
        return ln, lb, fl, first.value

    def eval_batch_noalloc(self, cands, budget=None, skip=1):
        """eval_batch through vnsgpu_eval_batch_noalloc: scratch in the library's static device array, no allocation;
        at most VNSGPU_NOALLOC_MAX_CANDS candidates per call."""
        c = self.make_cands(cands)
        n = len(c)
        ln, lb, fl = np.empty(n), np.empty(n), np.zeros(n, np.uint8)
        first = C.c_int64(-1)
        budget = self.e.pb.budget if budget is None else budget
        self.e._ck(self.lib.vnsgpu_eval_batch_noalloc(self.e.h, self.h, n, c.ctypes.data, budget, skip, _d(ln), _d(lb),
                                                      fl.ctypes.data, C.byref(first)))
        return ln, lb, fl, first.value

    @staticmethod
    def pack_cands(cands):
        """CAND_DTYPE array -> uint64 descriptors (vnsgpu_pack_cand in include/vnsgpu.h)."""
        c = VnsSopPath.make_cands(cands)
        u = lambda x: x.astype(np.int64).astype(np.uint64)
        return np.ascontiguousarray((u(c["op"]) & np.uint64(7)) | ((u(c["a"]) & np.uint64(0xfffff)) << np.uint64(3))
                                    | ((u(c["b"]) & np.uint64(0xfffff)) << np.uint64(23)) | (u(c["gain_cluster"] + 1) << np.uint64(43)))

    def eval_batch_packed(self, cands, budget=None, skip=1):
        """eval_batch with 8-byte descriptors (half the host-to-device bytes)."""
        pk = self.pack_cands(cands)
        n = len(pk)
        ln, lb, fl = np.empty(n), np.empty(n), np.zeros(n, np.uint8)
        first = C.c_int64(-1)
        budget = self.e.pb.budget if budget is None else budget
        self.e._ck(self.lib.vnsgpu_eval_batch_packed(self.e.h, self.h, n, pk.ctypes.data, budget, skip, _d(ln), _d(lb),
                                                     fl.ctypes.data, C.byref(first)))
        return ln, lb, fl, first.value

    @staticmethod
    def pack_cands32(cands):
        """CAND_DTYPE array -> uint32 descriptors (vnsgpu_pack_cand32): indices < 16384, gain cluster either a or none."""
        c = VnsSopPath.make_cands(cands)
        g = c["gain_cluster"]
        if not (((g == -1) | (g == c["a"])).all() and (c["a"] >= 0).all() and (c["b"] >= 0).all()
                and (c["a"] < 16384).all() and (c["b"] < 16384).all()):
            raise ValueError("candidates do not fit the 4-byte descriptor format")
        u = lambda x: x.astype(np.int64).astype(np.uint32)
        return np.ascontiguousarray((u(c["op"]) & np.uint32(7)) | (u(c["a"]) << np.uint32(3)) | (u(c["b"]) << np.uint32(17))
                                    | (u(g >= 0) << np.uint32(31)))

    def eval_batch_packed32(self, cands, budget=None, skip=1):
        """Compact wire format on all-integer tables: 4-byte descriptors in, int32 lengths + flags out."""
        pk = self.pack_cands32(cands)
        n = len(pk)
        ln, fl = np.empty(n, np.int32), np.zeros(n, np.uint8)
        first = C.c_int64(-1)
        budget = self.e.pb.budget if budget is None else budget
        self.e._ck(self.lib.vnsgpu_eval_batch_packed32(self.e.h, self.h, n, pk.ctypes.data, budget, skip, ln.ctypes.data,
                                                       fl.ctypes.data, C.byref(first)))
        return ln, fl, first.value

    def eval_batch_packed32r(self, cands, budget=None, skip=1):
        """One 4-byte result word per candidate (length | flags << 29, all ones = bad indices) -> (int32 lengths with
        INT32_MIN for bad indices, flags, first accepted index), taken apart like vnsgpu_result_len / _flags."""
        pk = self.pack_cands32(cands)
        n = len(pk)
        w = np.empty(n, np.uint32)
        first = C.c_int64(-1)
        budget = self.e.pb.budget if budget is None else budget
        self.e._ck(self.lib.vnsgpu_eval_batch_packed32r(self.e.h, self.h, n, pk.ctypes.data, budget, skip, w.ctypes.data,
                                                        C.byref(first)))
        bad = w == np.uint32(0xffffffff)
        ln = np.where(bad, np.iinfo(np.int32).min, (w & np.uint32(0x1fffffff)).astype(np.int64)).astype(np.int32)
        fl = np.where(bad, F_BAD, w >> np.uint32(29)).astype(np.uint8)
        return ln, fl, first.value

    def _one(self, op, a, b):
        ln, lb, fl, _ = self.eval_batch(np.array([[op, a, b]], np.int32), budget=np.inf)
        if fl[0] & F_BAD:
            raise VnsGpuError(1, "index outside the range the reference exit(1)s on")
        return ln[0], lb[0]

    def tryToAdd(self, cluster, idx):
        return self._one(OP_ADD, cluster, idx)[0]

    def tryToAddLowerBound(self, cluster, idx):
        return self._one(OP_ADD, cluster, idx)[1]

    def tryToReplace(self, cluster, idx):
        return self._one(OP_REPLACE, cluster, idx)[0]

    def tryToReplaceLowerBound(self, cluster, idx):
        return self._one(OP_REPLACE, cluster, idx)[1]

    def tryToRemove(self, idx):
        return self._one(OP_REMOVE, idx, 0)[0]

    def tryToMove(self, what, where):
        return self._one(OP_MOVE, what, where)[0]

    def tryToMoveLowerBound(self, what, where):
        return self._one(OP_MOVE, what, where)[1]

    def tryToExchange(self, i1, i2):
        return self._one(OP_EXCHANGE, i1, i2)[0]

    def tryToExchangeLowerBound(self, i1, i2):
        return self._one(OP_EXCHANGE, i1, i2)[1]

    def gen_proposals(self, kind, n, rng):
        c = np.zeros(n, CAND_DTYPE)
        self.e._ck(self.lib.vnsgpu_gen_proposals(self.e.h, self.h, kind, n, C.byref(rng), c.ctypes.data))
        return c

    def gen_proposals_dev(self, kind, n, rng, d_out, fmt=0, d_draws=None, stream=None):
        """n proposals of neighbourhood `kind` generated on the device into d_out (device pointer: vnsgpu_cand[n] for fmt 0,
        uint32[n] for fmt 2); d_draws: device pointer to uint32[n] or None; rng (host) advances past proposal n - 1"""
        self.e._ck(self.lib.vnsgpu_gen_proposals_dev(self.e.h, self.h, kind, n, C.byref(rng), C.c_void_p(d_out), fmt,
                                                     C.c_void_p(d_draws) if d_draws else None, C.c_void_p(stream) if stream else None))

    def propose_eval(self, kind, n, rng, budget=None, skip=1):
        """one round of a host-driven local search: -> (index of the first accepted proposal or -1, its (op, a, b, gain) or None);
        rng is left just after the accepted proposal's draws"""
        first = C.c_int64(-1)
        c = np.zeros(1, CAND_DTYPE)
        budget = self.e.pb.budget if budget is None else budget
        self.e._ck(self.lib.vnsgpu_propose_eval(self.e.h, self.h, kind, n, C.byref(rng), budget, skip, C.byref(first), c.ctypes.data))
        return first.value, (tuple(int(x) for x in c[0]) if first.value >= 0 else None)

    def cand_bytes(self, cands):
        c = self.make_cands(cands)
        out = C.c_int64()
        self.e._ck(self.lib.vnsgpu_cand_bytes(self.e.h, self.h, len(c), c.ctypes.data, C.byref(out)))
        return out.value

    # ---- neighbourhood drivers (VnsSopSolver) -----------------------------------------------------
    def insertRandom(self, num_changes, rng, budget=None, skip=1):
        return self._ls(1, num_changes, rng, budget, skip)

    def exchangeRandom(self, num_changes, rng, budget=None, skip=1):
        return self._ls(2, num_changes, rng, budget, skip)

    def _ls(self, kind, num_changes, rng, budget, skip):
        st = LsStats()
        budget = self.e.pb.budget if budget is None else budget
        self.e._ck(self.lib.vnsgpu_local_search(self.e.h, self.h, kind, int(num_changes), budget, int(skip),
                                                C.byref(rng), C.byref(st)))
        return st

    def shake(self, kind, rng, budget=None):
        budget = self.e.pb.budget if budget is None else budget
        self.e._ck(self.lib.vnsgpu_shake(self.e.h, self.h, kind, budget, C.byref(rng)))

    def fitDOPtoBudget(self, budget=None):
        budget = self.e.pb.budget if budget is None else budget
        self.e._ck(self.lib.vnsgpu_fit_to_budget(self.e.h, self.h, budget))

    def greedy_insertion(self, budget=None):
        budget = self.e.pb.budget if budget is None else budget
        self.e._ck(self.lib.vnsgpu_greedy(self.e.h, self.h, budget))

    def greedy_insertion_batched(self, budget=None):
        """greedy_insertion with every round's tryToAdd grid as one batched evaluation; returns (rounds, evals)"""
        budget = self.e.pb.budget if budget is None else budget
        r, n = C.c_int(0), C.c_int64(0)
        self.e._ck(self.lib.vnsgpu_greedy_batched(self.e.h, self.h, budget, C.byref(r), C.byref(n)))
        return r.value, n.value

    def best_move(self, kind, budget=None):
        """best improving candidate of the whole insert (1) / exchange (2) neighbourhood (not trajectory-exact)"""
        budget = self.e.pb.budget if budget is None else budget
        m = BestMove()
        self.e._ck(self.lib.vnsgpu_best_move(self.e.h, self.h, int(kind), budget, C.byref(m)))
        return m

    def apply_move(self, kind, from_, to):
        self.e._ck(self.lib.vnsgpu_apply_move(self.e.h, self.h, int(kind), int(from_), int(to)))

    def best_improvement(self, budget=None, max_steps=0):
        """variable-neighbourhood descent over whole neighbourhoods; returns (moves applied, candidates evaluated)"""
        budget = self.e.pb.budget if budget is None else budget
        st, n = C.c_int(0), C.c_int64(0)
        self.e._ck(self.lib.vnsgpu_best_improvement(self.e.h, self.h, budget, int(max_steps), C.byref(st), C.byref(n)))
        return st.value, n.value

    def vns_run(self, seed, num_iterations=2000, num_iterations_unimproved=1000, skip=1, greedy=1, budget=None,
                max_time_ms=0.0, max_improvements=4096):
        """VnsSopSolver::iterate as one device chain -> (VnsResult, improvement log [(reward, length, ms)])."""
        budget = self.e.pb.budget if budget is None else budget
        prm = VnsParams(num_iterations, num_iterations_unimproved, skip, greedy, budget, max_time_ms, seed,
                        max_improvements)
        res = VnsResult()
        ir, il, it = (np.zeros(max_improvements) for _ in range(3))
        self.e._ck(self.lib.vnsgpu_vns_run(self.e.h, self.h, C.byref(prm), C.byref(res), _d(ir), _d(il), _d(it)))
        n = min(res.n_improvements, max_improvements)
        return res, list(zip(ir[:n], il[:n], it[:n]))

    def vns_run_batch(self, seeds, num_iterations=2000, num_iterations_unimproved=1000, skip=1, greedy=1, budget=None,
                      max_time_ms=0.0, target_reward=0.0):
        """Multi-start (the reference's --batch=N): one device chain per seed in ONE launch ->
        (list of VnsResult, index of the best run); the best tour replaces this path.  target_reward > 0: race -
        the launch ends as soon as one run's tour reaches it (vnsgpu_set_target_reward)."""
        budget = self.e.pb.budget if budget is None else budget
        self.e.set_target_reward(target_reward)
        prm = VnsParams(num_iterations, num_iterations_unimproved, skip, greedy, budget, max_time_ms, 0, 0)
        n = len(seeds)
        sd = (C.c_uint32 * n)(*[int(x) for x in seeds])
        res = (VnsResult * n)()
        best = C.c_int(-1)
        self.e._ck(self.lib.vnsgpu_vns_run_batch(self.e.h, self.h, n, sd, C.byref(prm), res, C.byref(best), None))
        return list(res), best.value
