for c in 4 8 16; do VNSGPU_PIPE_CHUNKS=$c timeout 600 python bench.py --steps 20 --no-cpu --no-solve > gpurun_out/bench_q_c$c.json 2> gpurun_out/bench_q.err; done
timeout 600 python bench.py --steps 20 --no-cpu --no-solve > gpurun_out/bench_q_cdef.json 2> gpurun_out/bench_q.err
