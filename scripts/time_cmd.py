"""wall time of a command: python scripts/time_cmd.py LABEL -- cmd ..."""
import subprocess, sys, time
i = sys.argv.index("--")
t0 = time.perf_counter()
rc = subprocess.call(sys.argv[i + 1:], stdout=subprocess.DEVNULL)
print("%s: %.1f s (rc %d)" % (" ".join(sys.argv[1:i]), time.perf_counter() - t0, rc), flush=True)
