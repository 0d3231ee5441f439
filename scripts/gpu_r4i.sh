# r4i: whole GPU suite + smoke at this commit (path_set rows from the products on long tours), time of path_set on the 20 k-node state
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -x -q > gpurun_out/r4i_gpu_tests.log 2>&1; echo "pytest exit $?" >> gpurun_out/r4i_gpu_tests.log
tail -6 gpurun_out/r4i_gpu_tests.log
timeout 300 python __graft_entry__.py smoke 2>&1 | tail -2
python - <<'PY' 2>&1 | tail -4
import os, sys, time
sys.path.insert(0, ".")
import bench
from vns_sop_b200 import engine, synthetic
sf, pb = bench.load_workload("synth20k")
e = engine.Engine(0).load_problem(pb)
vns, T = synthetic.centroid_tour_state(pb, 0.85)
for form in ("0", "1"):
    os.environ["VNSGPU_APPLY_PRODUCTS"] = form
    p = e.path(vns, T); e.sync()
    t0 = time.perf_counter()
    for _ in range(20): p.set(vns, T)
    e.sync()
    print("synth20k path_set T=%d rows=%s: %.3f ms per call, length %g" % (T, "products" if form == "1" else "scan", 1e3 * (time.perf_counter() - t0) / 20, p.getPathLength()))
PY
