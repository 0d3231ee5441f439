"""SOP instance files: the reference's `.sop` text format (GTSP-like TSPLIB dialect), the
loader's start/goal renumbering, and the flat cluster-contiguous model the engine uploads.

Restates (host side, runs once per instance; not accelerated):
  * DatasetLoaderSOP::loadDataset   sources/dataset_loader_sop.cpp:39-287  -> parse_sop
  * SOPLoader::getSOPDefinition     sources/sop_loader.cpp:102-173          -> build_problem
  * SOPLoader::fixClusterNodeRenumbering sources/sop_loader.cpp:185-250     -> Problem.orig_cluster_id
Instances travel to machines without the reference tree as compressed `.npz` (save_npz/load_npz);
write_sop re-emits a `.sop` text file any reference-compatible loader can read.
"""
from __future__ import annotations

import dataclasses
import math
from typing import List, Optional

import numpy as np

RULE_RAW, RULE_ROUND, RULE_CEIL = 0, 1, 2
_RULES = {"EUC_2D": RULE_ROUND, "CEIL_2D": RULE_CEIL,
          "RAW_2D": RULE_RAW}   # not a TSPLIB type: the unrounded Euclid of SOPLoader::getOP / getOPN (sop_loader.cpp:278,360)


@dataclasses.dataclass
class SopFile:
    """Contents of one `.sop` file, ids 0-based (the file numbers nodes from 1)."""
    name: str = "no_name"
    comment: str = ""
    tmax: int = 0
    start_set: int = 0
    end_set: int = 0
    edge_weight_type: str = "CEIL_2D"          # EUC_2D | CEIL_2D | EXPLICIT
    coords: Optional[np.ndarray] = None        # (N,2) float64  (NODE_COORD_SECTION)
    weights: Optional[np.ndarray] = None       # (N,N) float64  (EDGE_WEIGHT_SECTION, FULL_MATRIX)
    set_profit: Optional[np.ndarray] = None    # (C,) float64
    set_nodes: Optional[List[np.ndarray]] = None  # C arrays of 0-based node ids
    centers: Optional[np.ndarray] = None       # (C,2) GTSP_SET_CENTER_COORD_SECTION (DOP files)
    dubins_radius: float = 0.0

    @property
    def dimension(self) -> int:
        return len(self.coords) if self.coords is not None else len(self.weights)


def parse_sop(path: str) -> SopFile:
    """Same section/keyword handling as dataset_loader_sop.cpp:73-262 (keys split at the first
    ':'; NODE_COORD_SECTION / EOF are bare lines; the weight section is a token stream that
    fills rows of DIMENSION entries)."""
    sf = SopFile()
    section = None
    coords, wtokens, sets, centers = [], [], [], []
    dimension = None
    with open(path) as fh:
        for raw in fh:
            line = raw.strip()
            if not line:
                continue
            if ":" in line:
                key, _, val = line.partition(":")
                key, val = key.strip(), val.strip()
                if key == "NAME":
                    sf.name = val
                elif key == "COMMENT":
                    sf.comment = val
                elif key == "DIMENSION":
                    dimension = int(val)
                elif key == "TMAX":
                    sf.tmax = int(val)              # std::stoi, dataset_loader_sop.cpp:110
                elif key == "START_SET":
                    sf.start_set = int(val)
                elif key == "END_SET":
                    sf.end_set = int(val)
                elif key == "EDGE_WEIGHT_TYPE":
                    sf.edge_weight_type = val
                elif key == "DUBINS_RADIUS":
                    sf.dubins_radius = float(val)
                elif key == "GTSP_SET_SECTION":
                    section = "sets"
                elif key == "GTSP_SET_CENTER_COORD_SECTION":
                    section = "centers"
                elif key == "EDGE_WEIGHT_SECTION":
                    section = "weights"
                continue
            if line == "NODE_COORD_SECTION":
                section = "coords"
            elif line == "EOF":
                section = None
            elif line == "GTSP_SET_SECTION":
                section = "sets"
            elif line == "EDGE_WEIGHT_SECTION":
                section = "weights"
            elif section == "coords":
                t = line.split()
                coords.append((float(t[1]), float(t[2])))
            elif section == "weights":
                wtokens.extend(line.split())
            elif section == "sets":
                t = line.split()
                if len(t) < 3:
                    raise ValueError("set line needs >= 3 tokens: " + line)
                if int(t[0]) != len(sets):
                    raise ValueError("wrong cluster_id order in " + path)
                nodes = np.array([int(x) - 1 for x in t[2:] if int(x) - 1 >= 0], dtype=np.int32)
                sets.append((float(t[1]), nodes))
            elif section == "centers":
                t = line.split()
                centers.append((float(t[1]), float(t[2])))
    if coords:
        sf.coords = np.array(coords, dtype=np.float64)
    if wtokens:
        w = np.array([float(x) for x in wtokens], dtype=np.float64)
        if dimension is None or w.size != dimension * dimension:
            raise ValueError("edge weight matrix size != DIMENSION^2")
        sf.weights = w.reshape(dimension, dimension)
    sf.set_profit = np.array([s[0] for s in sets], dtype=np.float64)
    sf.set_nodes = [s[1] for s in sets]
    if centers:
        sf.centers = np.array(centers, dtype=np.float64)
    return sf


def write_sop(sf: SopFile, path: str) -> None:
    with open(path, "w") as fh:
        fh.write(f"NAME: {sf.name}\nTYPE: TSP\nCOMMENT: {sf.comment}\nDIMENSION: {sf.dimension}\n")
        fh.write(f"TMAX: {sf.tmax}\nSTART_SET: {sf.start_set}\nEND_SET: {sf.end_set}\n")
        fh.write(f"SETS: {len(sf.set_nodes)}\n")
        if sf.dubins_radius:
            fh.write(f"DUBINS_RADIUS: {sf.dubins_radius:g}\n")
        fh.write(f"EDGE_WEIGHT_TYPE: {sf.edge_weight_type}\n")
        if sf.weights is not None:
            fh.write("EDGE_WEIGHT_FORMAT: FULL_MATRIX\nEDGE_WEIGHT_SECTION\n")
            for row in sf.weights:
                fh.write(" ".join(repr(float(v)) if v != int(v) else str(int(v)) for v in row) + "\n")
        else:
            fh.write("NODE_COORD_SECTION\n")
            for i, (x, y) in enumerate(sf.coords):
                fh.write(f"{i + 1} {float(x)!r} {float(y)!r}\n")
        fh.write("GTSP_SET_SECTION: set_id set_profit id-vertex-list\n")
        for c, nodes in enumerate(sf.set_nodes):
            fh.write(f"{c} {sf.set_profit[c]:g} " + " ".join(str(int(n) + 1) for n in nodes) + "\n")
        if sf.centers is not None:
            fh.write("GTSP_SET_CENTER_COORD_SECTION: set_id x y\n")
            for c, (x, y) in enumerate(sf.centers):
                fh.write(f"{c} {float(x)!r} {float(y)!r}\n")


def save_npz(sf: SopFile, path: str) -> None:
    d = dict(name=sf.name, comment=sf.comment, tmax=sf.tmax, start_set=sf.start_set,
             end_set=sf.end_set, edge_weight_type=sf.edge_weight_type,
             dubins_radius=sf.dubins_radius, set_profit=sf.set_profit,
             set_sizes=np.array([len(s) for s in sf.set_nodes], dtype=np.int32),
             set_nodes=np.concatenate(sf.set_nodes).astype(np.int32))
    if sf.coords is not None:
        d["coords"] = sf.coords
    if sf.weights is not None:
        w = sf.weights
        d["weights"] = w.astype(np.int32) if np.all(w == np.floor(w)) and np.abs(w).max() < 2**31 else w
    if sf.centers is not None:
        d["centers"] = sf.centers
    np.savez_compressed(path, **d)


def load_npz(path: str) -> SopFile:
    z = np.load(path, allow_pickle=False)
    sizes = z["set_sizes"]
    ends = np.cumsum(sizes)
    nodes = z["set_nodes"]
    sf = SopFile(name=str(z["name"]), comment=str(z["comment"]), tmax=int(z["tmax"]),
                 start_set=int(z["start_set"]), end_set=int(z["end_set"]),
                 edge_weight_type=str(z["edge_weight_type"]), dubins_radius=float(z["dubins_radius"]),
                 set_profit=z["set_profit"].astype(np.float64),
                 set_nodes=[nodes[e - s:e].astype(np.int32) for s, e in zip(sizes, ends)])
    if "coords" in z:
        sf.coords = z["coords"].astype(np.float64)
    if "weights" in z:
        sf.weights = z["weights"].astype(np.float64)
    if "centers" in z:
        sf.centers = z["centers"].astype(np.float64)
    return sf


@dataclasses.dataclass
class Problem:
    """Flat model after the loader's renumbering: cluster 0 = start, cluster C-1 = goal (a copy of
    the start set when START_SET == END_SET), samples stored cluster-contiguously."""
    name: str
    C: int
    m: np.ndarray            # (C,) int32 samples per cluster
    off: np.ndarray          # (C+1,) int32
    reward: np.ndarray       # (C,) float64
    start: int
    goal: int
    budget: float
    node_id: np.ndarray      # (N,) int32 original 0-based node id of each sample
    orig_cluster_id: np.ndarray  # (C,) int32 set id printed in RESULT_CLUSTER_IDS
    rule: int                # RULE_* for coordinate instances, -1 for EXPLICIT
    xy: Optional[np.ndarray] = None   # (N,2) sample coordinates (cluster-contiguous)
    W: Optional[np.ndarray] = None    # (N,N) explicit weights in sample order
    centers: Optional[np.ndarray] = None  # (C,2) per renumbered cluster (DOP)
    dubins_radius: float = 0.0

    @property
    def N(self) -> int:
        return int(self.off[-1])

    def initial_vns(self) -> np.ndarray:
        """vnsVector of iterate(): all clusters but start and goal, in index order
        (vns_sop_solver.cpp:91-99,119-122)."""
        return np.array([c for c in range(self.C) if c not in (self.start, self.goal)], dtype=np.int32)


def build_problem(sf: SopFile) -> Problem:
    """sop-type=sop branch of SOPLoader::getSOPDefinition with change_end_index_to_last=true
    (sop_loader.cpp:118-156): goal set moved (or duplicated) to the last index, start set to 0."""
    clusters = [[c, float(sf.set_profit[c]), sf.set_nodes[c], c] for c in range(len(sf.set_nodes))]
    s, g = sf.start_set, sf.end_set
    end_cluster = list(clusters[g])
    last_cluster = list(clusters[-1])
    if g == s:
        clusters.append(list(end_cluster))
    else:
        clusters[g] = list(last_cluster)
        clusters[g][0] = g
    goal = len(clusters) - 1
    clusters[goal] = list(end_cluster)
    clusters[goal][0] = goal
    if s != 0:
        start_c, first_c = list(clusters[s]), list(clusters[0])
        clusters[s] = first_c
        clusters[s][0] = s
        clusters[0] = start_c
        clusters[0][0] = 0
    start = 0
    C = len(clusters)
    m = np.array([len(c[2]) for c in clusters], dtype=np.int32)
    off = np.zeros(C + 1, dtype=np.int32)
    off[1:] = np.cumsum(m)
    node_id = np.concatenate([c[2] for c in clusters]).astype(np.int32)
    pb = Problem(name=sf.name, C=C, m=m, off=off,
                 reward=np.array([c[1] for c in clusters], dtype=np.float64), start=start, goal=goal,
                 budget=float(sf.tmax), node_id=node_id,
                 orig_cluster_id=np.array([c[3] for c in clusters], dtype=np.int32),
                 rule=_RULES.get(sf.edge_weight_type, -1), dubins_radius=sf.dubins_radius)
    if sf.coords is not None:
        pb.xy = np.ascontiguousarray(sf.coords[node_id])
    else:
        pb.W = np.ascontiguousarray(sf.weights[np.ix_(node_id, node_id)])
    if sf.centers is not None:
        pb.centers = np.ascontiguousarray(sf.centers[pb.orig_cluster_id])
    return pb
