#!/usr/bin/env python
"""bench.py - candidate move evaluations per second (BASELINE.json `metric`) on B200.

One STEP = one pass of the hot path over one batch: `--batch` candidate moves (MOVE / ADD drawn with
insertRandom's proposal distribution, EXCHANGE / REPLACE with exchangeRandom's,
sources/vns_sop_solver.cpp:507-519,723-730) evaluated against one fixed search state by ONE launch
of the batched layered min-plus kernel (length = tryTo*, lower bound, budget-feasibility / accept
flags, first accepted index).  Workload = BASELINE.json configs[1]: 40krob200_T80_p2.

  value         evals/s, candidates and results resident in HBM (vnsgpu_eval_batch_dev), CUDA events
  e2e           evals/s through the host-buffer C-ABI call vnsgpu_eval_batch: pinned host
                candidates -> device, kernel, lengths + flags -> host, all inside the timed region
  roofline      algorithmic bytes of the batch / kernel time, against the measured
                HBM copy bandwidth of MEASURED_PEAKS.json
  cpu_baseline  the reference's own op::VnsSopPath::tryTo* (oracle/_ref, compiled unmodified) or
                the C port, on a bounded sample of the same candidates, on this box's host cores
  --impl reference   the CPU arm alone: all host cores, same candidates, same metric

N > 1 (torchrun): the batch shards by rank - every rank evaluates its own `--batch` candidates
against its replica of the table (weak scaling, no data-path collective); the time is the max over
ranks.  No CPU fallback: without libvnsgpu.so or a GPU the own arm raises.
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

WORKLOADS = {
    "krob200": ("40krob200_T80_p2", os.path.join(ROOT, "tests", "golden", "krob200.npz")),
    "vm1084": ("217vm1084_T100_p1", os.path.join(ROOT, "tests", "golden", "vm1084.npz")),
    "berlin52": ("11berlin52_T60_p2", os.path.join(ROOT, "tests", "golden", "berlin52.npz")),
    "opn_s08": ("tsiligirides_problem_2_budget_30_d_50_s_08", os.path.join(ROOT, "tests", "golden", "opn_s08.npz")),
    "dop_s12": ("tsiligirides_problem_2_budget_30_r_50_s_12", os.path.join(ROOT, "tests", "golden", "dop_s12.npz")),
    # BASELINE.json config 3: the shipped DOP files stop at 12 headings; 16 are generated from the s_08 file's cluster
    # centres with the Dubins table kernel (vnsgpu_build_table_dubins, ceil rule like the shipped EXPLICIT matrices)
    "dop_s16": ("tsiligirides_problem_2_budget_30_r_50_s_16 (generated: centres of the shipped s_08 file, 16 headings, Dubins table kernel)",
                os.path.join(ROOT, "tests", "golden", "dop_s08.npz"), {"dubins_S": 16}),
    # BASELINE.json config 5: generated (vns-sop_b200/synthetic.py), table far beyond L2 - the HBM regime
    "synth20k": ("synthetic scaled SOP: 20000 nodes / 2000 clusters of 10, CEIL_2D (generated, seed 20190147)", None, {"n_points": 20000}),
    # SURVEY.md 8d's raw-Euclid variant of config 5: a real-valued table, evaluated in IEEE double block by block (no
    # segment products, no endpoint vectors: floating-point sums do not reassociate)
    "synth20k_raw": ("synthetic scaled SOP: 20000 nodes / 2000 clusters of 10, unrounded Euclid = real-valued table (generated, seed 20190147)",
                     None, {"n_points": 20000, "edge_weight_type": "RAW_2D"}),
    "synth2k": ("synthetic scaled SOP: 2000 nodes / 200 clusters of 10, CEIL_2D (generated, seed 20190147)", None, {"n_points": 2000}),
    # BASELINE.json config 5, DOP flavour: 2000 locations x 32 heading samples = 64 064 samples; the 32.8 GB table only ever
    # exists on the device (vnsgpu_build_table_dubins), so this workload has no CPU legs
    "synth_dop64k": ("synthetic scaled DOP: 2000 locations x 32 heading samples (64064 samples, Dubins table built on the device, "
                     "rho = 0.5% of the box; generated, seed 20190147)", None, {"dop": {"n_centres": 2000, "headings": 32}}),
    "synth_dop4k": ("synthetic scaled DOP: 500 locations x 8 heading samples (generated, seed 20190147)", None,
                    {"dop": {"n_centres": 500, "headings": 8}}),
}
STATE = "the greedy initial tour (greedy_insertion, vns_sop_solver.cpp:1091-1155; SURVEY.md 8d)"
STATE_SYNTH = ("nearest-neighbour tour over the cluster centroids cut at 85% of the budget (greedy_insertion is "
               "O(C^3) evaluations at this size; synthetic.centroid_tour_state)")
FALLBACK_HBM_GBS = 6650.0   # /opt/skills/guides/B200_PROFILING.md "FALLBACK"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="krob200", choices=sorted(WORKLOADS))
    ap.add_argument("--batch", type=int, default=1 << 22, help="candidates per step per GPU")
    ap.add_argument("--ref-batch", type=int, default=1 << 20, help="candidates per step of the reference arm")
    ap.add_argument("--cpu-sample", type=int, default=1 << 22, help="candidates of the cpu_baseline leg")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-solve", action="store_true", help="skip the time-to-reward leg")
    ap.add_argument("--vector-cache", type=int, default=1, help="vnsgpu_set_vector_cache: 0 off, 1 auto (batch >= C^2 / 4), 2 always")
    ap.add_argument("--batch-sort", type=int, default=-1, help="vnsgpu_set_batch_sort: 0 off, 1 on, -1 library default (experiments)")
    ap.add_argument("--no-solve-cpu", action="store_true", help="time-to-reward: skip the reference binary runs (long on vm1084)")
    ap.add_argument("--sweep", action="store_true", help="candidate-batch sweep 2^10 .. --batch (x4 steps), device-resident (BASELINE config 5)")
    ap.add_argument("--multi-start", type=int, default=148, help="seeds of the multi-start leg (0 = skip)")
    ap.add_argument("--multi-start-dense", type=int, default=592, help="seeds of the dense multi-start launch (four per SM; 0 = skip)")
    ap.add_argument("--also", default="vm1084,synth20k,synth20k_raw", help="further workloads measured (device-resident leg only) into the same JSON "
                                                               "line under `workloads` (BASELINE configs 4 and 5); '' = none")
    ap.add_argument("--no-full-nbh", action="store_true", help="skip the full-neighbourhood leg (batched greedy + best-improvement descent)")
    ap.add_argument("--descent-steps", type=int, default=200, help="moves of the best-improvement descent in the full-neighbourhood leg")
    ap.add_argument("--min-timed-s", type=float, default=0.5, help="a timed step repeats the call until K steps take at least this long")
    return ap.parse_args()


# ---- workload: state + candidates (shared by both arms) ----------------------------------------------
def draw_proposals(kind, n, vns, T, rs):
    """n proposals of insertRandom (kind 1) / exchangeRandom (kind 2) against (vns, T):
    from ~ U[0,|VNS|-1], to ~ U[0,T] resp. U[0,T-1], both redrawn while |from-to| <= 1 resp. 2
    (vns_sop_solver.cpp:507-519, 723-730) -> (op, a, b) rows."""
    nv = len(vns)
    gap, tohi = (1, T) if kind == 1 else (2, T - 1)
    fr = np.empty(n, np.int64)
    to = np.empty(n, np.int64)
    todo = np.arange(n)
    while len(todo):
        f = rs.randint(0, nv, len(todo))
        t = rs.randint(0, tohi + 1, len(todo))
        fr[todo], to[todo] = f, t
        todo = todo[np.abs(f - t) <= gap]
    inside = fr < T
    op = np.where(inside, 3 if kind == 1 else 4, 0 if kind == 1 else 1)
    a = np.where(inside, fr, np.asarray(vns)[fr])
    return np.stack([op, a, to], 1).astype(np.int32)


def make_candidates(n, vns, T, seed):
    rs = np.random.RandomState(seed)
    h = n // 2
    c = np.concatenate([draw_proposals(1, h, vns, T, rs), draw_proposals(2, n - h, vns, T, rs)])
    return c[rs.permutation(n)]


def dubins_S(name):
    w = WORKLOADS[name]
    if len(w) > 2 and "dop" in w[2]:
        return w[2]["dop"]["headings"]
    return w[2]["dubins_S"] if len(w) > 2 and "dubins_S" in w[2] else 0


def device_only(name):
    """Workloads whose table is never materialised on the host (no CPU legs)."""
    w = WORKLOADS[name]
    return len(w) > 2 and "dop" in w[2]


def load_workload(name):
    from vns_sop_b200 import instances as ins
    if WORKLOADS[name][1] is None and "dop" in WORKLOADS[name][2]:
        from vns_sop_b200 import synthetic
        return None, synthetic.generate_dop(**WORKLOADS[name][2]["dop"])      # no file form: the table is device-only
    if WORKLOADS[name][1] is None:
        from vns_sop_b200 import synthetic
        sf = synthetic.generate(**WORKLOADS[name][2])
    else:
        sf = ins.load_npz(WORKLOADS[name][1])
    S = dubins_S(name)
    if S:       # same sets, S heading samples each; the weights are filled in from the device table (fill_dubins_weights)
        C0 = len(sf.set_nodes)
        sf.set_nodes = [np.arange(c * S, (c + 1) * S, dtype=np.int32) for c in range(C0)]
        sf.weights = np.zeros((C0 * S, C0 * S))
        sf.coords = None
        sf.edge_weight_type = "EXPLICIT"
        sf.name = WORKLOADS[name][0].split(" ")[0]
    return sf, ins.build_problem(sf)


def fill_dubins_weights(sf, pb, W):
    """W = the device-built table in sample order -> the file's node order, so that the CPU side can load the
    instance as the EXPLICIT matrix the reference's DOP files carry."""
    sf.weights = np.zeros_like(W)
    sf.weights[np.ix_(pb.node_id, pb.node_id)] = W
    pb.W = W


def workload_state(name, pb):
    """(vns, T) the candidates are drawn against, or None = run greedy_insertion (dataset instances)."""
    if WORKLOADS[name][1] is None and (pb.xy is not None or device_only(name)):
        from vns_sop_b200 import synthetic
        return synthetic.centroid_tour_state(pb, 0.6 if device_only(name) else 0.85)
    return None


def state_text(name):
    return STATE_SYNTH if WORKLOADS[name][1] is None else STATE


class CpuEvaluator:
    """The CPU side: the reference's own code when oracle/_ref was built (kind "reference"), else the
    C port (kind "port").  TEST/BASELINE infrastructure - never on the product path."""

    def __init__(self, sf, pb, state=None):
        from oracle import binding as ob
        # the reference's 4-deep vector table (x4 by-value copies) needs tens of GB beyond ~5000 samples
        self.kind = "reference" if ob.RefOracle.available() and pb.N <= 5000 else "port"
        if self.kind == "reference":
            from vns_sop_b200 import instances as ins
            self.tmp = tempfile.NamedTemporaryFile(suffix=".sop", delete=False)
            self.tmp.close()
            ins.write_sop(sf, self.tmp.name)            # /root/reference does not exist on the GPU box
            self.o = ob.RefOracle(self.tmp.name)
        else:
            self.o = ob.Oracle(pb)
        if state is None:
            self.p = self.o.path()
            self.p.greedy()
        else:
            self.p = self.o.path(np.ascontiguousarray(state[0], np.int32), int(state[1]))
        self.vns, self.T = np.asarray(self.p.vns), int(self.p.T)

    def eval(self, cands):
        return self.p.eval_batch(cands)


# ---- reference arm: persistent worker processes, one per host core ---------------------------------------
_W = {}


def _worker_init(workload):
    if "ev" not in _W:          # normally inherited from the parent through fork (one table, shared pages)
        sf, pb = load_workload(workload)
        if dubins_S(workload):   # no GPU on this arm: the Dubins table from the CPU restatement (oracle/vns_oracle.c)
            from oracle import binding as ob
            fill_dubins_weights(sf, pb, ob.Oracle(pb, dubins=(dubins_S(workload), sf.dubins_radius, 1)).dense())
        _W["ev"] = CpuEvaluator(sf, pb, workload_state(workload, pb))


def _worker_eval(cands):
    ln, lb = _W["ev"].eval(cands)
    return float(ln.sum())


def _worker_state(_):
    ev = _W["ev"]
    return ev.kind, ev.vns, ev.T


def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    if device_only(args.workload):
        print(json.dumps({"impl": "reference", "unavailable": "the 33 GB table of this workload only exists on the device; the reference's "
                          "4-deep vector table (copied by value 4 times) does not fit the host"}), flush=True)
        return
    import multiprocessing as mp
    cores = host_cores()
    name = WORKLOADS[args.workload][0]
    _worker_init(args.workload)
    with mp.get_context("fork").Pool(cores, initializer=_worker_init, initargs=(args.workload,)) as pool:
        kind, vns, T = pool.map(_worker_state, [0])[0]
        n = args.ref_batch
        cands = make_candidates(n, vns, T, 1000)
        chunks = np.array_split(cands, cores * 4)
        for _ in range(max(args.warmup, 1)):
            pool.map(_worker_eval, chunks)
        t0 = time.perf_counter()
        for _ in range(args.steps):
            pool.map(_worker_eval, chunks)
        dt = time.perf_counter() - t0
    v = n * args.steps / dt
    sample = "%d candidates/step (same proposal mix and state as the b200 arm) split over %d processes" % (n, cores)
    line = {
        "impl": "reference", "metric": "candidate_move_evals_per_sec", "value": v, "unit": "evals/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": name, "candidates_per_step": n, "state": state_text(args.workload),
                   "outputs": "length + lower bound per candidate"},
        "cpu_baseline": {"value": v, "unit": "evals/s", "cores": cores, "kind": kind, "sample": sample},
        "e2e": {"value": v, "unit": "evals/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ---- clocks during the timed region (B200_PROFILING.md "clocks DURING the timed region") ----------------
class ClockSampler:
    """SM clock + throttle reasons sampled every 5 ms through NVML (nvidia-ml-py) in a background thread;
    falls back to `nvidia-smi -lms 100` (B200_PROFILING.md) when NVML cannot be loaded."""
    NAMES = {"hw_slowdown": 0x8, "hw_thermal_slowdown": 0x40, "sw_thermal_slowdown": 0x20, "sw_power_cap": 0x4}
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows, self.proc, self.stop_flag, self.max_mhz = [], None, False, None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            phys = int(vis.split(",")[index]) if vis and vis.split(",")[index].isdigit() else index
            self.h = pynvml.nvmlDeviceGetHandleByIndex(phys)
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
            self.th = threading.Thread(target=self._poll, daemon=True)
            self.th.start()
        except Exception:
            self.nv = None
            try:
                self.proc = subprocess.Popen(["nvidia-smi", "-i", str(index), "--query-gpu=" + self.Q,
                                              "--format=csv,noheader,nounits", "-lms", "100"],
                                             stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
                self.th = threading.Thread(target=self._read, daemon=True)
                self.th.start()
            except OSError:
                self.proc = None

    def sample_now(self):
        """One NVML sample from the calling thread: bench.py takes them while the timed steps are queued on the
        device (a 7 ms timed region can fall between two polls of the background thread)."""
        nv = self.nv
        if nv is None:
            return
        try:
            mhz = float(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
            try:
                mask = int(nv.nvmlDeviceGetCurrentClocksEventReasons(self.h))
            except Exception:
                mask = int(nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h))
            self.rows.append((time.perf_counter(), mhz, {k for k, bit in self.NAMES.items() if mask & bit}))
        except Exception:
            pass

    def _poll(self):
        while not self.stop_flag:
            self.sample_now()
            time.sleep(0.002)

    def _read(self):
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in self.proc.stdout:
            r = [x.strip() for x in line.split(",")]
            try:
                self.max_mhz = float(r[1])
                act = {nm for k, nm in enumerate(names) if len(r) > 3 + k and r[3 + k].lower().startswith("active")}
                self.rows.append((time.perf_counter(), float(r[0]), act))
            except (ValueError, IndexError):
                continue

    def window(self, t0, t1):
        rows = [r for r in self.rows if t0 <= r[0] <= t1] or [r for r in self.rows if t0 - 0.2 <= r[0] <= t1 + 0.2]
        reasons = set()
        for r in rows:
            reasons |= r[2]
        return {"sm_mhz": float(np.median([r[1] for r in rows])) if rows else None, "sm_max_mhz": self.max_mhz,
                "reasons": sorted(reasons), "samples": len(rows),
                "source": "nvml, 2 ms period + samples taken while the timed steps are queued" if self.nv else "nvidia-smi -lms 100"}

    def stop(self):
        self.stop_flag = True
        if self.proc:
            self.proc.terminate()


def ncu_cache_level(workload, batch):
    """Which memory level the dominant kernel saturates, from the same committed capture (percent of that unit's peak):
    for a cache-resident table the HBM fraction says nothing - the binding roof is the L1 data pipe (LSU wavefronts)."""
    try:
        with open(os.path.join(ROOT, "profiles", "ncu_traffic.json")) as fh:
            for t in json.load(fh)["entries"]:
                if t.get("workload") == workload and int(t.get("batch", -1)) == batch and t.get("l1tex_throughput_pct") is not None:
                    lv = {"l1tex": t.get("l1tex_throughput_pct"), "lsu_wavefronts": t.get("lsu_wavefronts_pct"),
                          "l2": t.get("l2_throughput_pct"), "dram": t.get("dram_throughput_pct"), "issue_active": t.get("issue_active_pct")}
                    top = max((k for k in ("l1tex", "l2", "dram") if lv[k] is not None), key=lambda k: lv[k])
                    return {"binding_level": top, "pct_of_peak": lv, "l1_hit_pct": t.get("l1_hit_pct"),
                            "kernel_us_under_ncu": t.get("kernel_us_under_ncu"), "capture": t.get("capture"),
                            "what": "ncu --set full: throughput of each memory level as a percentage of its own peak for one launch "
                                    "of the dominant kernel at this batch size"}
    except (OSError, ValueError, KeyError):
        pass
    return None


def ncu_traffic(workload, batch):
    """dram__bytes_read.sum + dram__bytes_write.sum of the dominant kernel for one launch of this batch size,
    from the committed ncu --set full capture (profiles/ncu_traffic.json, written by scripts/ncu_extract.py)."""
    try:
        with open(os.path.join(ROOT, "profiles", "ncu_traffic.json")) as fh:
            for t in json.load(fh)["entries"]:
                if t.get("workload") == workload and int(t.get("batch", -1)) == batch:
                    return float(t["dram_bytes_per_launch"])
    except (OSError, ValueError, KeyError):
        pass
    return None


def measured_hbm_peak():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as fh:
            return float(json.load(fh)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except (OSError, KeyError, ValueError):
        return FALLBACK_HBM_GBS, "fallback (B200_PROFILING.md)"


def vec_form_bytes(rows, T, layer_m, cluster_m):
    """Bytes k_eval_vec has to move for ITS algorithm, per batch: for every candidate the two endpoint vectors
    (4 B per sample) and - MOVE / EXCHANGE with a non-empty untouched run - ONE block (a table block or a segment product,
    4 B per entry); the descriptor in and the results out are counted by the caller.  rows: (n, 3) int32 [op, a, b];
    layer_m[l] = samples of layer l (0 .. T+1); cluster_m[c] = samples of cluster c.  Mirrors decode_cand_vec (vec.cuh)."""
    op, a, b = rows[:, 0].astype(np.int64), rows[:, 1].astype(np.int64), rows[:, 2].astype(np.int64)
    lm = np.asarray(layer_m, np.int64)
    vec = np.zeros(len(rows), np.int64)
    blk = np.zeros(len(rows), np.int64)
    addrep = (op == 0) | (op == 1)
    vec[addrep] = 2 * np.asarray(cluster_m, np.int64)[a[addrep]]
    mv = op == 3
    up = mv & (b > a)
    dn = mv & (b < a)
    ex = op == 4
    la = np.zeros(len(rows), np.int64)
    lb = np.zeros(len(rows), np.int64)
    i, j = np.minimum(a, b) + 1, np.maximum(a, b) + 1
    la[up], lb[up] = i[up] + 1, j[up] - 1
    la[dn], lb[dn] = i[dn], j[dn] - 1
    pq = np.minimum(a, b) + 1, np.maximum(a, b) + 1
    la[ex], lb[ex] = pq[0][ex] + 1, pq[1][ex] - 1
    chain = mv | ex
    ok = chain & (la >= 0) & (lb <= T + 1) & (la <= T + 1) & (lb >= 0)
    vec[ok] = lm[la[ok]] + lm[lb[ok]]
    hasblk = ok & (lb > la)
    blk[hasblk] = lm[la[hasblk]] * lm[lb[hasblk]]
    return int(4 * vec.sum()), int(4 * blk.sum()), float(hasblk.mean())


def prepare_gpu_state(eng, state=None):
    if state is not None:
        return eng.path(np.ascontiguousarray(state[0], np.int32), int(state[1]))
    p = eng.path()
    p.greedy_insertion()
    return p


def run_b200(args):
    import ctypes as C

    import torch
    import torch.distributed as dist

    from vns_sop_b200 import engine, sharding

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py (b200 arm) needs a CUDA device; there is no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def sum_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    env = (rank, world, local, dev, barrier, max_over_ranks, sum_over_ranks)
    line = bench_workload(args, env)
    # BASELINE.json asks for configs 4 and 5 at 1/2/4/8 GPUs too: the same device-resident measurement on them, in the
    # same JSON line (short legs: no host-buffer call, no CPU legs)
    extras = [w for w in args.also.split(",") if w and w != args.workload]
    if extras:
        line["workloads"] = {}
    for w in extras:
        sub = argparse.Namespace(**vars(args))
        sub.workload, sub.no_e2e, sub.no_cpu, sub.no_solve, sub.sweep = w, True, True, True, False
        if w.endswith("_raw"):
            sub.batch = min(args.batch, 1 << 18)     # double path, block by block: ~260 blocks per candidate at T = 793
        sl = bench_workload(sub, env)
        line["workloads"][w] = {k: sl[k] for k in ("value", "unit", "ms_per_step", "steps", "dtype", "data", "config", "roofline", "gpu_launches", "table_build",
                                                     "accepted_in_batch", "first_accept", "full_neighbourhood") if k in sl}
    if "workloads" in line and "synth20k" in line["workloads"]:
        line["roofline_hbm"] = dict(line["workloads"]["synth20k"]["roofline"], workload="synth20k")
    if rank == 0 and world == 1 and not args.no_solve and args.workload == "krob200" and "vm1084" in extras:
        line["time_to_reward_vm1084"] = solve_only("vm1084")
    if rank == 0 and world == 1 and not args.no_solve and args.workload == "krob200" and args.also:
        # BASELINE.json config 3 (DOP, 16 heading samples, table from the Dubins kernel): seeded solve against the reference
        # binary on the same table, 148 seeds in one launch against one reference process per core
        sub = argparse.Namespace(**vars(args))
        sub.workload, sub.no_e2e, sub.no_cpu, sub.no_solve, sub.sweep, sub.no_full_nbh = "dop_s16", True, True, False, False, True
        sub.steps, sub.warmup = min(args.steps, 5), min(args.warmup, 3)
        sl = bench_workload(sub, env)
        line["time_to_reward_dop_s16"] = dict(sl.get("time_to_reward", {}), workload=sl["config"]["workload"], evals_per_s=sl["value"],
                                              table_build=sl.get("table_build"))
    if rank == 0:
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def bench_workload(args, env):
    """One workload: table build, device-resident throughput (value) with per-phase kernel times, optionally the
    host-buffer legs (e2e), the CPU baseline and the seeded solves.  Returns the JSON line as a dict."""
    import ctypes as C

    import torch

    from vns_sop_b200 import engine, sharding

    rank, world, local, dev, barrier, max_over_ranks, sum_over_ranks = env
    sf, pb = load_workload(args.workload)
    if device_only(args.workload):
        args.no_cpu = args.no_solve = True
        eng = engine.Engine(local).load_problem(pb, dubins=(dubins_S(args.workload), pb.dubins_radius, 1))
    elif dubins_S(args.workload):
        eng = engine.Engine(local).load_problem(pb, dubins=(dubins_S(args.workload), sf.dubins_radius, 1))
        fill_dubins_weights(sf, pb, eng.table_dense())
    else:
        eng = engine.Engine(local).load_problem(pb)
    lib = eng.lib
    eng.set_vector_cache(args.vector_cache)
    if args.batch_sort >= 0:
        eng.set_batch_sort(bool(args.batch_sort))
    state = workload_state(args.workload, pb)
    p = prepare_gpu_state(eng, state)
    vns, T = p.vns, p.getNumTargets()
    table_build = None
    if pb.xy is not None or dubins_S(args.workload):   # a1 / a4 (+ a5): the fused pairwise table kernel, buffers re-allocated inside the call
        S_ = dubins_S(args.workload)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        if S_:
            eng.build_table_dubins(pb.centers, S_, pb.dubins_radius, 1)
        else:
            eng.build_table_euclid(pb.xy, pb.rule)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        kms = float(lib.vnsgpu_table_kernel_ms(eng.h))
        p = prepare_gpu_state(eng, (vns, T))
        wbytes = int(eng.table_bytes) * (3 if lib.vnsgpu_int_path_active(eng.h) else 2) // 2
        peak_t = measured_hbm_peak()[0]
        table_build = {"kernel": "k_table_pairs<%s>" % ("dubins" if S_ else "euclid"), "kernel_ms": kms, "call_ms": 1e3 * dt,
                       "entries": int(pb.N) ** 2, "bytes_written": wbytes,
                       "achieved_gbs": wbytes / (kms * 1e-3) / 1e9 if kms > 0 else None,
                       "frac_of_hbm_peak": wbytes / (kms * 1e-3) / 1e9 / peak_t if kms > 0 else None,
                       "bound": "fp64 transcendental issue (six Dubins words per entry)" if S_ else "hbm (write)",
                       "what": "ONE pass: distance per entry, stored as double (8 B) and as int32 (4 B, exact path), cluster-pair "
                               "min / max (a5) and the integer classification in the epilogue; nothing is read back.  kernel_ms = CUDA "
                               "events around the launch; call_ms = the whole C-ABI call (H2D of the coordinates, cudaMalloc of the "
                               "int32 copy, sync)"}
    n, K, W = args.batch, args.steps, args.warmup
    rows = make_candidates(n, vns, T, 1000 + rank)          # each rank owns its shard of the candidates
    cands = p.make_cands(rows)
    alg_bytes = p.cand_bytes(cands)                        # w = 4 B / table entry on the exact int32 path
    int_path = eng.int_path_active
    eng.set_int_path(False)
    alg_bytes_f64 = p.cand_bytes(cands)                    # SURVEY.md 8d's figure: 8 B / entry
    eng.set_int_path(True)

    # ---- value: everything resident in HBM -----------------------------------------------------------
    stream = torch.cuda.Stream(device=dev)
    h_c = torch.empty(n * 16, dtype=torch.uint8).pin_memory()
    h_c.numpy()[:] = cands.view(np.uint8)
    d_c = h_c.to(dev)
    d_len = torch.empty(n, dtype=torch.float64, device=dev)
    d_lb = torch.empty(n, dtype=torch.float64, device=dev)
    d_fl = torch.empty(n, dtype=torch.uint8, device=dev)
    d_first = torch.zeros(1, dtype=torch.int64, device=dev)
    torch.cuda.synchronize()
    sp = C.c_void_p(stream.cuda_stream)

    comm_stream = None
    if world > 1:
        # the library's own NCCL communicator (vnsgpu_comm_*): the unique id travels through torch.distributed, the
        # collectives of the path do not
        import torch.distributed as dist
        idt = torch.zeros(128, dtype=torch.uint8, device=dev)
        if rank == 0:
            idt.copy_(torch.frombuffer(bytearray(eng.comm_unique_id()), dtype=torch.uint8))
        dist.broadcast(idt, 0)
        eng.comm_init(bytes(idt.cpu().numpy().tobytes()), rank, world)
        comm_stream = torch.cuda.Stream(device=dev)
    d_keys = torch.zeros(4, dtype=torch.int64, device=dev)      # ring: the key of step k reduces while step k + 1 evaluates
    step_no = [0]

    def step():
        if world > 1:      # sharded form: the key comes out as a global index, ready for the collective
            key = d_keys[step_no[0] % 4:]
            step_no[0] += 1
            st = lib.vnsgpu_eval_batch_dev_shard(eng.h, p.h, n, C.c_void_p(d_c.data_ptr()), pb.budget, 1,
                                                 C.c_void_p(d_len.data_ptr()), C.c_void_p(d_lb.data_ptr()),
                                                 C.c_void_p(d_fl.data_ptr()), C.c_void_p(key.data_ptr()), rank * n, sp)
            eng._ck(st)
            # the path's one exchange step: the chosen move (8 bytes, all-reduce MIN) - on a second stream, so that its
            # latency (16-28 us, r1y) hides behind the next step's kernels instead of serialising with them
            comm_stream.wait_stream(stream)
            eng.reduce_first_key(key.data_ptr(), comm_stream.cuda_stream)
            return
        st = lib.vnsgpu_eval_batch_dev(eng.h, p.h, n, C.c_void_p(d_c.data_ptr()), pb.budget, 1,
                                       C.c_void_p(d_len.data_ptr()), C.c_void_p(d_lb.data_ptr()),
                                       C.c_void_p(d_fl.data_ptr()), C.c_void_p(d_first.data_ptr()), sp)
        eng._ck(st)

    sampler = ClockSampler(local) if rank == 0 else None
    with torch.cuda.stream(stream):
        for _ in range(W):
            step()
        # a step of 0.3 ms x K = 20 would be a 7 ms timed region: one timed STEP is `inner` back-to-back calls (each a
        # full pass over the batch), sized from the warm-up so that the K steps take >= --min-timed-s
        w0, w1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        w0.record(stream)
        for _ in range(3):
            step()
        w1.record(stream)
        barrier()
        est_ms = max_over_ranks(w0.elapsed_time(w1)) / 3
        inner = int(max(1, min(4096, np.ceil(args.min_timed_s * 1e3 / (K * max(est_ms, 1e-3))))))
        eng.set_phase_timing(True)
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(K)]
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        l0 = eng.launches
        t_host0 = time.perf_counter()
        e0.record(stream)
        for a, b in ev:
            a.record(stream)
            for _ in range(inner):
                step()
            b.record(stream)
        if comm_stream is not None:
            stream.wait_stream(comm_stream)      # the last keys' reductions are inside the timed region
        e1.record(stream)
        if sampler:
            sampler.sample_now()          # after the last launch: the device is still working through the queued steps
                                          # (never between launches: an NVML query can take milliseconds and would
                                          # starve the device - seen on an 8-GPU box: 1.9 ms / step instead of 0.34)
        barrier()
        t_host1 = time.perf_counter()
        launches = eng.launches - l0
        phases = eng.phase_times()
        eng.set_phase_timing(False)
    total_ms = max_over_ranks(e0.elapsed_time(e1))
    kern_ms = float(np.mean([a.elapsed_time(b) for a, b in ev])) / inner      # one call = one pass over the batch
    clocks = sampler.window(t_host0, t_host1) if sampler else None
    value = world * n * K * inner / (total_ms * 1e-3)
    first_dev = int(d_first.item()) if world == 1 else int(d_keys[(step_no[0] - 1) % 4].item())
    if first_dev >= 0x7f7f7f7f7f7f7f7f:      # VNSGPU_NO_KEY of the sharded form
        first_dev = -1
    ok_flags = int(((d_fl & 2) != 0).sum().item())

    # ---- candidate-batch sweep (BASELINE config 5): the same call on the first b candidates of the batch ---------
    sweep = None
    if args.sweep:
        sweep = []
        b = 1 << 10
        sizes = []
        while b < n:
            sizes.append(b)
            b <<= 2
        sizes.append(n)
        with torch.cuda.stream(stream):
            for b in sizes:
                def sstep():
                    if world > 1:
                        eng._ck(lib.vnsgpu_eval_batch_dev_shard(eng.h, p.h, b, C.c_void_p(d_c.data_ptr()), pb.budget, 1,
                                                                C.c_void_p(d_len.data_ptr()), C.c_void_p(d_lb.data_ptr()),
                                                                C.c_void_p(d_fl.data_ptr()), C.c_void_p(d_first.data_ptr()), rank * b, sp))
                        eng.reduce_first_key(d_first.data_ptr(), stream.cuda_stream)
                    else:
                        eng._ck(lib.vnsgpu_eval_batch_dev(eng.h, p.h, b, C.c_void_p(d_c.data_ptr()), pb.budget, 1,
                                                          C.c_void_p(d_len.data_ptr()), C.c_void_p(d_lb.data_ptr()),
                                                          C.c_void_p(d_fl.data_ptr()), C.c_void_p(d_first.data_ptr()), sp))
                for _ in range(W):
                    sstep()
                s0, s1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                barrier()
                s0.record(stream)
                for _ in range(K):
                    sstep()
                s1.record(stream)
                barrier()
                ms = max_over_ranks(s0.elapsed_time(s1))
                sweep.append({"batch_per_gpu": b, "evals_per_s": world * b * K / (ms * 1e-3), "ms_per_step": ms / K})

    # ---- e2e: the host-buffer C-ABI call (H2D of the candidates + kernel + D2H of lengths and flags) --
    e2e = None
    if not args.no_e2e:
        h_len = torch.empty(n, dtype=torch.float64).pin_memory()
        h_fl = torch.empty(n, dtype=torch.uint8).pin_memory()
        first = C.c_int64(-1)
        eng.set_stream(None)

        h_pk = torch.empty(n, dtype=torch.int64).pin_memory()
        h_pk.numpy()[:] = p.pack_cands(cands).view(np.int64)

        h_pk32 = h_len32 = None
        if int_path:   # compact wire format of all-integer tables: 4 B descriptors in, int32 lengths + flags out
            h_pk32 = torch.empty(n, dtype=torch.int32).pin_memory()
            h_pk32.numpy()[:] = p.pack_cands32(cands).view(np.int32)
            h_len32 = torch.empty(n, dtype=torch.int32).pin_memory()

        h_res = torch.empty(n, dtype=torch.int32).pin_memory() if int_path else None

        def host_step(packed=1):
            if packed == 9:
                propose_step()
                return
            if packed == 4:     # 4 B descriptors in, only the chosen move (first accepted index) out: what insertRandom /
                                # exchangeRandom consume (vns_sop_solver.cpp:619-646, :805-832)
                eng._ck(lib.vnsgpu_eval_batch_packed32(eng.h, p.h, n, C.c_void_p(h_pk32.data_ptr()), pb.budget, 1,
                                                       None, None, C.byref(first)))
            elif packed == 3:   # 4 B descriptors in, ONE 4 B result word (length | flags << 29) out
                eng._ck(lib.vnsgpu_eval_batch_packed32r(eng.h, p.h, n, C.c_void_p(h_pk32.data_ptr()), pb.budget, 1,
                                                        C.c_void_p(h_res.data_ptr()), C.byref(first)))
            elif packed == 2:
                eng._ck(lib.vnsgpu_eval_batch_packed32(eng.h, p.h, n, C.c_void_p(h_pk32.data_ptr()), pb.budget, 1,
                                                       C.c_void_p(h_len32.data_ptr()), C.c_void_p(h_fl.data_ptr()), C.byref(first)))
            elif packed:    # 8-byte descriptors (vnsgpu_pack_cand): half the host-to-device bytes
                eng._ck(lib.vnsgpu_eval_batch_packed(eng.h, p.h, n, C.c_void_p(h_pk.data_ptr()), pb.budget, 1,
                                                     C.cast(h_len.data_ptr(), C.POINTER(C.c_double)), None,
                                                     C.c_void_p(h_fl.data_ptr()), C.byref(first)))
            else:
                eng._ck(lib.vnsgpu_eval_batch(eng.h, p.h, n, C.c_void_p(h_c.data_ptr()), pb.budget, 1,
                                              C.cast(h_len.data_ptr(), C.POINTER(C.c_double)), None,
                                              C.c_void_p(h_fl.data_ptr()), C.byref(first)))
            if world > 1:
                sharding.reduce_first_accept(first.value, rank * n, device=dev)

        def timed(packed):
            """K x rep calls (rep sized from the warm-up so that the region takes >= --min-timed-s / 2); returns the time
            of K calls' worth (= total / rep) so that the formulas below stay per K calls"""
            for _ in range(W):
                host_step(packed)
            barrier()
            t0 = time.perf_counter()
            host_step(packed)
            barrier()
            one = max_over_ranks(time.perf_counter() - t0)
            rep = int(max(1, min(256, np.ceil(0.5 * args.min_timed_s / (K * max(one, 1e-5))))))
            l0 = eng.launches
            t0 = time.perf_counter()
            for _ in range(K * rep):
                host_step(packed)
            barrier()
            return max_over_ranks(time.perf_counter() - t0) / rep, (eng.launches - l0) // rep

        # the call a host-driven search makes per round: proposals drawn ON THE DEVICE from the host's generator state
        # (vnsgpu_propose_eval: 132 B in), evaluated, first accepted proposal + its descriptor + the rewound generator out
        rng_p = engine.seeded_rng(4242 + rank)
        prop_kind = [1]

        def propose_step(_=None):
            prop_kind[0] = 3 - prop_kind[0]           # alternate insertRandom / exchangeRandom windows (the 50 / 50 mix)
            f, _c = p.propose_eval(prop_kind[0], n, rng_p)
            if world > 1:      # the chosen move over all ranks' windows: one 8-byte all-reduce(MIN)
                d_keys[0] = f + rank * n if f >= 0 else engine.NO_KEY
                torch.cuda.current_stream().synchronize()
                eng.reduce_first_key(d_keys.data_ptr(), None)
                eng.sync()
                return int(d_keys[0].item())
            return f
        host_step_saved = None
        dt16, _ = timed(0)
        dt8, nl8 = timed(1)
        first8 = first.value
        sub16 = {"value": world * n * K / dt16, "ms_per_step": 1e3 * dt16 / K, "call": "vnsgpu_eval_batch",
                 "h2d_bytes_per_step": world * n * 16, "d2h_bytes_per_step": world * (n * 9 + 8)}
        sub8 = {"value": world * n * K / dt8, "ms_per_step": 1e3 * dt8 / K, "call": "vnsgpu_eval_batch_packed",
                "h2d_bytes_per_step": world * n * 8, "d2h_bytes_per_step": world * (n * 9 + 8)}
        if int_path:
            dt5, _ = timed(2)
            first5 = first.value
            sub5 = {"value": world * n * K / dt5, "ms_per_step": 1e3 * dt5 / K, "call": "vnsgpu_eval_batch_packed32",
                    "h2d_bytes_per_step": world * n * 4, "d2h_bytes_per_step": world * (n * 5 + 8)}
            dt0, _ = timed(4)
            first0 = first.value
            sub0 = {"value": world * n * K / dt0, "ms_per_step": 1e3 * dt0 / K, "call": "vnsgpu_eval_batch_packed32 (len = flags = NULL)",
                    "h2d_bytes_per_step": world * n * 4, "d2h_bytes_per_step": world * 8,
                    "what": "only the index of the first accepted candidate comes back - the one result the reference's "
                            "insertRandom / exchangeRandom loops consume"}
            dt, nl = timed(3)
            launches += nl
            in_b, out_b = 4, 4
            e2e = {"value": world * n * K / dt, "unit": "evals/s", "h2d_bytes_per_step": world * n * 4,
                   "d2h_bytes_per_step": world * (n * 4 + 8), "ms_per_step": 1e3 * dt / K,
                   "call": "vnsgpu_eval_batch_packed32r (pinned host buffers: 4 B descriptors in; one 4 B result word per "
                           "candidate out = int32 length - exact on this all-integer table - | flags << 29, + first accepted "
                           "index); replayed as one CUDA graph launch",
                   "first_accept_only": sub0, "with_4B_descriptors_i32_lengths_u8_flags": sub5, "with_8B_descriptors_f64_lengths": sub8,
                   "with_16B_descriptors": sub16}
            if world == 1:
                assert first.value == first8 == first5 == first0, "compact and 8-byte entry points disagree on the first accepted candidate"
            assert torch.equal(h_len32[:65536].double(), h_len[:65536]), "compact entry point: lengths differ"
            w = h_res.numpy().view(np.uint32)[:65536]
            assert ((w & 0x1fffffff) == h_len32.numpy()[:65536]).all() and ((w >> 29) == h_fl.numpy()[:65536]).all(), \
                "result words differ from lengths + flags"
        else:
            launches += nl8
            in_b, out_b = 8, 9
            e2e = dict(sub8, unit="evals/s", with_16B_descriptors=sub16,
                       call="vnsgpu_eval_batch_packed (pinned host buffers: 8 B descriptors in; lengths + flags + first accepted index out)")
        dtp, nlp = timed(9)
        e2e["propose_eval"] = {"value": world * n * K / dtp, "unit": "evals/s", "ms_per_step": 1e3 * dtp / K, "call": "vnsgpu_propose_eval",
                               "h2d_bytes_per_step": world * 132, "d2h_bytes_per_step": world * (8 + 8 + 4 + 16),
                               "frac_of_value": (world * n * K / dtp) / value,
                               "what": "per step: n proposals drawn on the device from the host's glibc rand() state (alternating "
                                       "insertRandom / exchangeRandom windows), evaluated, first accepted index + descriptor + draw "
                                       "count back; the generator is rewound on the host (vnsgpu_rng_advance).  No candidate or "
                                       "result arrays cross PCIe"}
        launches += nlp
        # what the link alone allows: the same buffers copied with nothing else running (explains e2e vs value)
        d_tmp = torch.empty(n * 16, dtype=torch.uint8, device=dev)
        h_tmp = h_c
        evs = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
        d_tmp.copy_(h_tmp, non_blocking=True)
        torch.cuda.synchronize()
        evs[0].record()
        for _ in range(4):
            d_tmp.copy_(h_tmp, non_blocking=True)
        evs[1].record()
        for _ in range(4):
            h_len.copy_(d_len, non_blocking=True)
        evs[2].record()
        torch.cuda.synchronize()
        e2e["pcie_h2d_gbs"] = 4 * n * 16 / (evs[0].elapsed_time(evs[1]) * 1e-3) / 1e9
        e2e["pcie_d2h_gbs"] = 4 * n * 8 / (evs[1].elapsed_time(evs[2]) * 1e-3) / 1e9
        e2e["copy_bound_ms_per_step"] = max(n * in_b / e2e["pcie_h2d_gbs"], (n * out_b + 8) / e2e["pcie_d2h_gbs"]) * 1e-6
        del d_tmp
        if world == 1:
            assert first.value == first_dev, "host and device entry points disagree on the first accepted candidate"
        assert torch.equal(h_len[:4096], d_len[:4096].cpu()), "host and device entry points disagree"

    mode = eng.eval_mode(n)              # the library's own choice for this batch (form, sort, HBM form)
    hbm_regime = mode["hbm_form"]
    vec_mode = mode["form"] == 2
    G = mode["group_lanes"]
    NRr = 1 if -(-int(pb.m.max()) // G) <= 1 else (2 if -(-int(pb.m.max()) // G) <= 2 else 4)
    kernel_name = ("k_eval_vec<%d,%d,%s>" % (G, NRr, "1" if hbm_regime else "0")) if vec_mode else \
                  ("k_eval_groups<%s,%d,%d,%s>" % ("int" if int_path else "double", G, NRr, "1" if mode["form"] == 1 else "0"))
    hbm_peak, peak_src = measured_hbm_peak()
    peaks = eng.measure_read_peaks() if rank == 0 or world == 1 else None
    if world > 1:
        t = torch.tensor([peaks["l1"], peaks["l2"], peaks["hbm"]] if peaks else [0.0, 0.0, 0.0], dtype=torch.float64, device=dev)
        import torch.distributed as dist
        dist.broadcast(t, 0)
        peaks = {"l1": float(t[0]), "l2": float(t[1]), "hbm": float(t[2])}
    eval_ms, eval_cnt = phases["eval"]
    eval_ms_per_launch = eval_ms / max(eval_cnt, 1)
    # bytes the dominant kernel has to move for its own algorithm (not the reference chain's)
    layer_m = pb.m[np.concatenate(([pb.start], np.asarray(vns[:T], np.int64), [pb.goal]))]
    desc_b, out_b = 16, 17               # device-resident call: vnsgpu_cand in; length + lower bound (f64) + flags out
    if vec_mode:
        vb, bb, blk_share = vec_form_bytes(rows, T, layer_m, pb.m)
        kbytes = n * (desc_b + out_b) + vb + bb
        breakdown = {"descriptor": desc_b, "results": out_b, "endpoint_vectors": vb / n, "block": bb / n, "candidates_with_a_block": blk_share}
    else:                                # block by block / segment products: the chain's blocks at the width the kernel reads
        kbytes = n * (desc_b + out_b) + (alg_bytes - 8 * n)
        breakdown = {"descriptor": desc_b, "results": out_b, "chain_blocks_and_rows": (alg_bytes - 8 * n) / n}
    # working set the kernel gathers from: the table copy it reads + segment products + endpoint vectors
    LS = pb.C + 2
    ws = int(pb.N) ** 2 * (4 if int_path else 8)
    if mode["form"] >= 1:
        ws += 4 * int(layer_m.sum()) ** 2 // 2
    if vec_mode:
        ws += 4 * int((2 * LS * pb.C + LS + 4 * (T + 2) * (T + 2)) * pb.m.max())
    bound = "l1" if ws <= (160 << 10) else ("l2" if ws <= (96 << 20) else "hbm")
    peak = peaks[bound] if bound != "hbm" else hbm_peak
    achieved = kbytes / (eval_ms_per_launch * 1e-3) / 1e9 if eval_ms_per_launch > 0 else None
    roofline = {
        "kernel": kernel_name, "kernel_ms": eval_ms_per_launch, "launches_timed": int(eval_cnt),
        "bound": bound, "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak if achieved else None,
        "traffic": ncu_traffic(args.workload, n),
        "peak_source": ("streaming read with the data resident in %s, measured in this run (vnsgpu_measure_read_peaks)" % bound.upper())
                       if bound != "hbm" else peak_src,
        "peaks_measured_gbs": dict(peaks, hbm_copy=hbm_peak),
        "algorithmic_bytes_per_launch": int(kbytes), "bytes_per_candidate": breakdown,
        "working_set_bytes": int(ws),
        "share_of_step": eval_ms / (kern_ms * K * inner) if kern_ms > 0 else None,
        "phases_ms_per_call": {k: v[0] / (K * inner) for k, v in phases.items()},
        "cache_level": ncu_cache_level(args.workload, n),
        "reference_work_rate": {"gbs": alg_bytes_f64 / (kern_ms * 1e-3) / 1e9, "bytes_per_candidate": alg_bytes_f64 / n,
                                "what": "SURVEY.md 8d's figure: the bytes the REFERENCE's chain would read for these candidates (8 B per entry of "
                                        "every block + first F row + last B row + result) / the time of one whole call.  The shared segment "
                                        "products and endpoint vectors let the kernel read far fewer bytes, so this is a work rate, not a "
                                        "bandwidth, and is not compared with any peak"},
        "note": "achieved = bytes the dominant kernel must move for its own algorithm (per candidate: descriptor, two endpoint vectors, at most "
                "one block, results) / its own launch time (CUDA events on the launching stream inside the timed region).  bound = the "
                "memory level that holds the working set the kernel gathers from; peak = streaming-read bandwidth measured with the data "
                "resident at that level.  The kernel is a gather of 4-byte entries with one DPX add-min per entry: on cache-resident "
                "tables it is bound by instruction issue / L1 wavefronts (cache_level, from the committed ncu capture), so frac is low "
                "by construction there; the HBM-regime block (roofline_hbm, workload synth20k) is the bandwidth claim",
    }
    line = {
        "metric": "candidate_move_evals_per_sec", "value": value, "unit": "evals/s", "n_gpus": world,
        "steps": K, "warmup": W, "ms_per_step": total_ms / K, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "i32 (exact on this all-integer table; results bit-identical to the f64 reference)" if int_path else "f64",
        "data": ("synthetic instance (generated, seeded)" if WORKLOADS[args.workload][1] is None else
                 "dataset instance shipped with the reference (datasets/, re-encoded under tests/golden/)") +
                "; candidates drawn with the reference's proposal distribution (seeded)",
        "ms_per_call": kern_ms,
        "config": {"workload": WORKLOADS[args.workload][0], "clusters": int(pb.C), "samples": int(pb.N),
                   "candidates_per_step_per_gpu": n * inner, "candidates_per_call_per_gpu": n, "calls_per_step": inner,
                   "targets_in_path": int(T),
                   "state": state_text(args.workload),
                   "mix": "50% insertRandom proposals (MOVE/ADD), 50% exchangeRandom (EXCHANGE/REPLACE)",
                   "outputs": "length + lower bound + flags per candidate, first accepted index",
                   "l2": ("candidates + results = %.0f MB per step > 126 MB L2 (streamed); the %.2f MB distance "
                          "table is cache-resident by design" % (n * 33 / 1e6, eng.table_bytes / 1e6)) if not hbm_regime else
                         ("the table read by the kernel (%.0f MB at %d B/entry, plus the segment products) is far beyond the "
                          "126 MB L2: blocks are gathered from HBM" % (eng.table_bytes / 8 * (4 if int_path else 8) / 1e6, 4 if int_path else 8)),
                   "evaluation": "sub-warp groups, %s" %
                                 (("exact int32 DPX arithmetic; per step the segment products and the endpoint vectors of the path "
                                   "state are rebuilt (k_seg_build, k_vec_uw, k_vec_pairs), then every candidate is one dot-min of two "
                                   "vectors through at most one block (k_eval_vec; %s)" %
                                   ("batch sorted by product row on the device" if hbm_regime else "batch in the caller's order")
                                   if vec_mode else
                                   "batch sorted by chain length on the device, exact int32 DPX arithmetic + exact segment products "
                                   "(rebuilt every step)") if int_path
                                  else "batch sorted by chain length on the device, IEEE double, block by block"),
                   "sharding": "candidates split by rank (global batch = n_gpus x per-GPU batch), table and path replicated; "
                               "N > 1 adds one 8-byte all-reduce(MIN) per step for the chosen move"},
        "e2e": e2e, "gpu_launches": int(launches), "clocks": clocks,
        "roofline": roofline,
        "accepted_in_batch": ok_flags, "first_accept": first_dev,
    }
    if table_build:
        line["table_build"] = table_build
    if sweep:
        line["batch_sweep"] = sweep
    if rank == 0 and world == 1 and not args.no_cpu:
        line["cpu_baseline"] = cpu_baseline_leg(args, sf, pb, vns, T, rows, d_len, d_lb, state)
    if rank == 0 and world == 1 and not args.no_solve and state is None:
        line["time_to_reward"] = solve_leg(args, eng, sf, pb)
    if rank == 0 and world == 1 and int_path and not args.no_full_nbh and not device_only(args.workload):
        line["full_neighbourhood"] = full_neighbourhood_leg(args, eng, sf, pb, cpu=not args.no_cpu and pb.N <= 5000)
    if world > 1 and not args.no_solve and state is None and args.multi_start > 0:
        # N GPUs: the honest multi-GPU mode of the search on shipped instances is replicas - every rank races
        # its own 148 seeds to the reference's seed-1 final reward; the job's time to that reward is the fastest rank's
        line["race_n_gpus"] = race_leg(args, eng, rank, world, max_over_ranks)
    if sampler:
        sampler.stop()
    del d_c, d_len, d_lb, d_fl
    eng.close()
    torch.cuda.empty_cache()
    return line


def full_neighbourhood_leg(args, eng, sf, pb, cpu):
    """greedy_insertion with every round as one batched evaluation (same tour as the reference's
    greedy), then the opt-in best-improvement descent over whole neighbourhoods - NOT the reference's trajectory."""
    import torch
    p = eng.path()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    rounds, evals = p.greedy_insertion_batched()
    torch.cuda.synchronize()
    t1 = time.perf_counter()
    r0, l0, T0 = p.getReward(), p.getPathLength(), p.getNumTargets()
    greedy_vns = np.array(p.vns)
    steps, evals2 = p.best_improvement(max_steps=args.descent_steps)
    torch.cuda.synchronize()
    t2 = time.perf_counter()
    out = {"greedy_batched": {"seconds": t1 - t0, "rounds": rounds, "candidates": evals, "evals_per_s": evals / max(t1 - t0, 1e-9),
                              "targets": T0, "reward": r0, "length": l0, "trajectory_exact": True,
                              "what": "greedy_insertion (vns_sop_solver.cpp:1091-1155): per round ONE vnsgpu_eval_batch_dev over the "
                                      "(|VNS| - T) x (T + 1) tryToAdd grid + device arg-min + one-block apply; the tour equals the "
                                      "reference's greedy tour (tests/test_best.py)"},
           "best_improvement": {"seconds": t2 - t1, "moves": steps, "max_moves": args.descent_steps, "candidates": evals2,
                                "evals_per_s": evals2 / max(t2 - t1, 1e-9), "reward": p.getReward(), "length": p.getPathLength(),
                                "targets": p.getNumTargets(), "trajectory_exact": False,
                                "what": "opt-in descent: per move the WHOLE insert or exchange neighbourhood evaluated at once, best "
                                        "improving candidate applied (most added reward, then shortest, then lowest index); a "
                                        "different search than the reference's first-improvement random sampling"}}
    # the trajectory-exact chain (seed 1, stock settings) from the same greedy tour for the time the descent took: what the
    # reference's own search reaches in that time on this device
    pc = eng.path(greedy_vns, T0)
    torch.cuda.synchronize()
    c0 = time.perf_counter()
    res, impr = pc.vns_run(1, greedy=0, max_time_ms=1e3 * (t2 - t1))
    c1 = time.perf_counter()
    out["best_improvement"]["exact_chain_same_time"] = {"seconds": c1 - c0, "time_limit_s": t2 - t1, "reward": res.reward, "length": res.length,
                                                        "vns_steps": res.num_iters, "proposals": res.proposals, "accepts": res.accepts,
                                                        "what": "vnsgpu_vns_run (seed 1, the reference's trajectory) started from the same greedy tour "
                                                                "with maximal-calculation-time = the descent's time; the limit is checked once per VNS "
                                                                "step (vns_sop_solver.cpp:189-193), so the run overshoots by up to one local-search call"}
    if cpu:
        from oracle import binding as ob
        o = ob.Oracle(pb)
        po = o.path()
        c0 = time.perf_counter()
        po.greedy()
        c1 = time.perf_counter()
        # CPU baseline of this leg (the checker, timed on one host core - never on the product path)
        out["greedy_batched"]["cpu_baseline"] = {"kind": "port", "cores": 1, "seconds": c1 - c0,
                                                 "same_tour_as_gpu": bool(po.T == T0 and po.length() == l0 and po.reward() == r0)}
    return out


def cpu_baseline_leg(args, sf, pb, vns, T, rows, d_len, d_lb, state=None):
    """The CPU checker timed on ONE host core on a bounded sample of the batch the GPU just evaluated;
    its lengths and lower bounds must equal the GPU's bit for bit."""
    ev = CpuEvaluator(sf, pb, state)
    assert ev.T == T and (ev.vns == vns).all(), "CPU and GPU search states differ"
    ns = min(args.cpu_sample, len(rows))
    sample = np.ascontiguousarray(rows[:ns])
    ev.eval(sample[:4096])
    dt, passes = 0.0, 0
    while passes == 0 or (dt < 10.0 and passes < 16):      # a bounded sample: >= 10 s of CPU work, whole passes
        t0 = time.perf_counter()
        ln, lb = ev.eval(sample)
        dt += time.perf_counter() - t0
        passes += 1
    g_len, g_lb = d_len[:ns].cpu().numpy(), d_lb[:ns].cpu().numpy()
    return {"value": passes * ns / dt, "unit": "evals/s", "cores": 1, "kind": ev.kind,
            "sample": "first %d candidates of rank 0's batch, %d pass(es), %.1f s" % (ns, passes, dt),
            "bit_identical_to_gpu": bool((ln == g_len).all() and (lb == g_lb).all())}


def solve_leg(args, eng, sf, pb):
    """Second half of the metric: wall time of the whole seeded VNS (stock sop.cfg stopping rule) to the
    reference's final reward - the device chain walks the reference's exact trajectory."""
    from oracle import binding as ob
    from vns_sop_b200 import instances as ins
    p = eng.path()
    t0 = time.perf_counter()
    res, impr = p.vns_run(1)
    gpu_wall = time.perf_counter() - t0
    out = {"seed": 1, "gpu_wall_s": gpu_wall, "gpu_device_ms": res.time_ms, "gpu_time_to_final_reward_s": res.time_last_impr_ms / 1e3,
           "reward": res.reward, "length": res.length,
           "num_iters": res.num_iters, "proposals": res.proposals, "evals": res.evals, "accepts": res.accepts,
           "chain_evals_per_s": res.evals / gpu_wall}
    try:     # the compiled reference's own seed-1 record, generated by tests/golden/make_golden.py
        with open(os.path.join(ROOT, "tests", "golden", "golden.json")) as fh:
            gold = json.load(fh).get(args.workload, {}).get("solve_seed1")
    except OSError:
        gold = None
    if gold:
        out["golden"] = {"reward": gold["REWARDS"], "length": gold["LENGTH"], "num_iters": gold["NUM_ITERS"],
                         "num_iters_last_impr": gold["NUM_ITERS_LAST_IMPR"], "reference_wall_s": gold.get("reference_wall_s"),
                         "same_trajectory": bool(gold["REWARDS"] == res.reward and gold["LENGTH"] == res.length
                                                 and gold["NUM_ITERS"] == res.num_iters and gold["NUM_ITERS_LAST_IMPR"] == res.num_iters_last_impr
                                                 and [[r, l] for r, l, _ in impr] == gold["IMPROVEMENTS"])}
    if os.path.exists(ob.REF_BIN) and not args.no_solve_cpu:
        from vns_sop_b200 import instances as ins
        with tempfile.TemporaryDirectory() as td:
            f = os.path.join(td, "w.sop")
            ins.write_sop(sf, f)
            t0 = time.perf_counter()
            subprocess.check_call([ob.REF_BIN, "--problem=" + f, "--seed=1", "--output=" + td],
                                  stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
            cpu_wall = time.perf_counter() - t0
            rec = open(os.path.join(td, "results.log")).read().strip().rstrip(";")
        d = dict(kv.split(":", 1) for kv in rec.split(";"))
        out.update({"cpu_reference_wall_s": cpu_wall, "cpu_time_to_final_reward_s": float(d["CTIME_LAST_IMPR"]) / 1e3,
                    "cpu_reward": float(d["REWARDS"]), "cpu_length": float(d["LENGTH"]),
                    "cpu_num_iters": int(d["NUM_ITERS"]),
                    "same_result": float(d["REWARDS"]) == res.reward and float(d["LENGTH"]) == res.length
                    and int(d["NUM_ITERS"]) == res.num_iters})
    # multi-start: the reference's --batch=N (N seeds) - one thread block per seed in ONE launch vs one
    # reference process per host core
    nruns = args.multi_start
    if nruns <= 0:
        return out
    p2 = eng.path()
    t0 = time.perf_counter()
    res_b, best = p2.vns_run_batch(list(range(1, nruns + 1)))
    gpu_multi = time.perf_counter() - t0
    out["multi_start"] = {"gpu_runs": nruns, "gpu_wall_s": gpu_multi, "gpu_runs_per_s": nruns / gpu_multi,
                          "best_reward": res_b[best].reward, "best_length": res_b[best].length, "best_seed": best + 1,
                          "seed1_same_as_single_run": res_b[0].reward == res.reward and res_b[0].length == res.length}
    # dense: four seeds per SM in one launch (third chain form: thread blocks of 128 threads, the table shared through L1
    # when the state is small enough for several blocks per SM)
    if args.multi_start_dense > nruns:
        pd_ = eng.path()
        t0 = time.perf_counter()
        res_d, best_d = pd_.vns_run_batch(list(range(1, args.multi_start_dense + 1)))
        gpu_dense = time.perf_counter() - t0
        out["multi_start"]["dense"] = {"gpu_runs": args.multi_start_dense, "gpu_wall_s": gpu_dense, "gpu_runs_per_s": args.multi_start_dense / gpu_dense,
                                       "best_reward": res_d[best_d].reward, "best_length": res_d[best_d].length, "best_seed": best_d + 1,
                                       "first_runs_same_as_above": all(res_d[k].reward == res_b[k].reward and res_d[k].length == res_b[k].length
                                                                       for k in range(nruns))}
    # race: the same launch with the reference's seed-1 final reward as target - it ends as soon as ANY seed's tour
    # reaches it ("wall time to the reference's reward" with one seed per SM); every run stays a prefix of its
    # seeded reference trajectory
    p3 = eng.path()
    t0 = time.perf_counter()
    res_r, win = p3.vns_run_batch(list(range(1, nruns + 1)), target_reward=res.reward)
    race_wall = time.perf_counter() - t0
    eng.set_target_reward(0.0)
    out["race_to_reference_reward"] = {"target_reward": res.reward, "gpu_runs": nruns, "gpu_wall_s": race_wall,
                                       "winner_seed": win + 1, "winner_reward": res_r[win].reward, "winner_length": res_r[win].length,
                                       "winner_device_ms": res_r[win].time_ms, "winner_num_iters": res_r[win].num_iters,
                                       "vs_single_chain_gpu_s": res.time_last_impr_ms / 1e3}
    if os.path.exists(ob.REF_BIN) and not args.no_solve_cpu:
        cores = host_cores()
        with tempfile.TemporaryDirectory() as td:
            f = os.path.join(td, "w.sop")
            ins.write_sop(sf, f)
            t0 = time.perf_counter()
            procs = []
            for sd in range(1, cores + 1):
                od = os.path.join(td, "o%d" % sd)
                os.makedirs(od)
                procs.append(subprocess.Popen([ob.REF_BIN, "--problem=" + f, "--seed=%d" % sd, "--output=" + od],
                                              stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL))
            for q in procs:
                q.wait()
            cpu_multi = time.perf_counter() - t0
            rewards, t_hit = [], []
            for sd in range(1, cores + 1):
                rec = open(os.path.join(td, "o%d" % sd, "results.log")).read().strip().rstrip(";")
                d = dict(kv.split(":", 1) for kv in rec.split(";"))
                rewards.append((float(d["REWARDS"]), float(d["LENGTH"])))
                for rec_i in d.get("SOLUTION_TIME_REWARD_LENGTH_IMPROVEMENTS", "").split(","):     # ms|reward|length
                    f3 = rec_i.split("|")
                    if len(f3) == 3 and float(f3[1]) >= res.reward:
                        t_hit.append(float(f3[0]) / 1e3)
                        break
        same = all(rewards[k] == (res_b[k].reward, res_b[k].length) for k in range(cores))
        out["multi_start"].update({"cpu_runs": cores, "cpu_cores": cores, "cpu_wall_s": cpu_multi, "cpu_runs_per_s": cores / cpu_multi,
                                   "first_%d_seeds_same_results" % cores: same})
        out["race_to_reference_reward"].update({
            "cpu_cores": cores, "cpu_seeds_reaching_target": len(t_hit),
            "cpu_first_hit_s": min(t_hit) if t_hit else None,
            "cpu_note": "reference binary, one seed per host core, all started together: the earliest time any run's improvement "
                        "log (results.log SOLUTION_TIME_REWARD_LENGTH_IMPROVEMENTS) shows reward >= target"})
    return out


def race_leg(args, eng, rank, world, max_over_ranks):
    with open(os.path.join(ROOT, "tests", "golden", "golden.json")) as fh:
        gold = json.load(fh).get(args.workload, {}).get("solve_seed1")
    if not gold:
        return None
    nruns = args.multi_start
    p = eng.path()
    t0 = time.perf_counter()
    res, win = p.vns_run_batch(list(range(rank * nruns + 1, (rank + 1) * nruns + 1)), target_reward=gold["REWARDS"])
    wall = time.perf_counter() - t0
    eng.set_target_reward(0.0)
    hit = res[win].reward >= gold["REWARDS"]
    best = -max_over_ranks(-(wall if hit else 1e9))          # min over ranks
    return {"target_reward": gold["REWARDS"], "seeds_per_gpu": nruns, "n_gpus": world, "seeds": world * nruns,
            "wall_s_first_gpu_to_target": best, "rank0_wall_s": wall, "rank0_winner_seed": rank * nruns + win + 1,
            "mode": "replicas: independent seeded chains per GPU, no collective in the search; one all-reduce(MIN) of the wall times"}


def solve_only(name, blocks=0):
    """Seeded (seed 1, stock sop.cfg) device solve of a further instance with its golden record - BASELINE config 4."""
    from vns_sop_b200 import engine
    sf, pb = load_workload(name)
    eng = engine.Engine(0).load_problem(pb)
    p = eng.path()
    eng.sync()
    t0 = time.perf_counter()
    res, impr = p.vns_run(1)
    wall = time.perf_counter() - t0
    with open(os.path.join(ROOT, "tests", "golden", "golden.json")) as fh:
        gold = json.load(fh).get(name, {}).get("solve_seed1")
    out = {"workload": WORKLOADS[name][0], "seed": 1, "gpu_wall_s": wall, "gpu_device_ms": res.time_ms,
           "gpu_time_to_final_reward_s": res.time_last_impr_ms / 1e3, "reward": res.reward, "length": res.length, "num_iters": res.num_iters,
           "proposals": res.proposals, "evals": res.evals, "accepts": res.accepts, "chain_proposals_per_s": res.proposals / wall}
    if gold:
        out["golden"] = {"reward": gold["REWARDS"], "length": gold["LENGTH"], "num_iters": gold["NUM_ITERS"],
                         "reference_wall_s_recorded": gold.get("reference_wall_s"),
                         "reference_wall_note": "the compiled reference binary, one core, recorded when tests/golden was generated (not re-run "
                                                "here: it takes ~18 minutes)",
                         "same_trajectory": bool(gold["REWARDS"] == res.reward and gold["LENGTH"] == res.length and gold["NUM_ITERS"] == res.num_iters
                                                 and gold["NUM_ITERS_LAST_IMPR"] == res.num_iters_last_impr
                                                 and [[r, l] for r, l, _ in impr] == gold["IMPROVEMENTS"])}
    eng.close()
    return out


def main():
    args = parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
