mkdir -p gpurun_out
timeout 300 python scripts/time_ls.py krob200 c3 > gpurun_out/dbg_ls.log 2>&1
python - <<'PY'
import re
tot=[0]*8; 
for line in open("gpurun_out/dbg_ls.log"):
    if line.startswith("c3dbg"):
        v=[int(x) for x in re.findall(r"-?\d+", line.replace("c3dbg",""))]
        tot=[a+b for a,b in zip(tot,v)]
print("shift %d scan %d wait %d layers %d | evalpro %d eval %d K %d n %d"%tuple(tot))
print("per layer scan cycles", tot[1]/max(tot[3],1), "per eval step", tot[5]/max(tot[6],1), "eval per call", tot[5]/max(tot[7],1), "pro per call", tot[4]/max(tot[7],1))
PY
grep -v c3dbg gpurun_out/dbg_ls.log
