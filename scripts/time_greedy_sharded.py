"""Full-neighbourhood modes with every neighbourhood sharded over N GPUs (vnsgpu_set_shard_neighbourhoods): one process per
GPU, started by this script.  usage: python scripts/time_greedy_sharded.py WORKLOAD NGPUS"""
import os, subprocess, sys, tempfile, time
sys.path.insert(0, ".")
if len(sys.argv) > 3:          # worker: WORKLOAD NGPUS RANK TMP
    name, world, rank, tmp = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), sys.argv[4]
    import torch
    import bench
    from vns_sop_b200 import engine
    torch.cuda.set_device(rank)
    sf, pb = bench.load_workload(name)
    e = engine.Engine(rank).load_problem(pb)
    idf = os.path.join(tmp, "id.bin")
    if rank == 0:
        open(idf + ".tmp", "wb").write(e.comm_unique_id()); os.rename(idf + ".tmp", idf)
    while not os.path.exists(idf):
        time.sleep(0.05)
    e.comm_init(open(idf, "rb").read(), rank, world)
    # the communicator's first collective sets its channels up (a few hundred ms): not part of either timing
    e.set_shard_neighbourhoods(1)
    w = e.path(); w.greedy_insertion_batched(); w.best_move(1); e.sync()
    for shard in (0, 1):
        e.set_shard_neighbourhoods(shard)
        p = e.path(); e.sync()
        t0 = time.perf_counter(); rounds, ev = p.greedy_insertion_batched(); e.sync(); t1 = time.perf_counter()
        steps, ev2 = p.best_improvement(max_steps=100); e.sync(); t2 = time.perf_counter()
        if rank == 0:
            print("%s %d GPUs sharded=%d: greedy %d rounds %.3f s (%.2f ms / round), descent %d moves %.3f s (%.2f ms / move); T=%d reward=%g length=%g"
                  % (name, world, shard, rounds, t1 - t0, 1e3 * (t1 - t0) / max(rounds, 1), steps, t2 - t1, 1e3 * (t2 - t1) / max(steps, 1),
                     p.getNumTargets(), p.getReward(), p.getPathLength()), flush=True)
    e.comm_destroy(); e.close()
else:
    name, world = sys.argv[1], int(sys.argv[2])
    tmp = tempfile.mkdtemp()
    ps = [subprocess.Popen([sys.executable, __file__, name, str(world), str(r), tmp]) for r in range(world)]
    sys.exit(max(q.wait() for q in ps))
