# usage: gpu_cap.sh NAME KERNEL_REGEX SKIP -- bench command...   one ncu --set full capture, exported to CSV on the box
# (the .ncu-rep files with sources are too large for gpurun_out's 64 MiB)
NAME=$1; KR=$2; SK=$3; shift 4
timeout 900 ncu --set full --clock-control none --import-source on -k regex:$KR -s $SK -c 1 -f -o gpurun_out/$NAME "$@" > gpurun_out/$NAME.ncu.log 2>&1
if [ -f gpurun_out/$NAME.ncu-rep ]; then
  ncu -i gpurun_out/$NAME.ncu-rep --page raw --csv > gpurun_out/${NAME}_raw.csv 2>/dev/null
  ncu -i gpurun_out/$NAME.ncu-rep --page details --csv > gpurun_out/${NAME}_details.csv 2>/dev/null
  ncu -i gpurun_out/$NAME.ncu-rep --page source --csv > gpurun_out/${NAME}_source.csv 2>/dev/null
  rm -f gpurun_out/$NAME.ncu-rep
fi
