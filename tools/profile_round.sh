#!/bin/bash
# Profile pass of a round (run on the B200 box through gpurun, from the repo root):  tools/profile_round.sh TAG
# Each ncu pass runs only after the same command has exited 0 without ncu.  Outputs under gpurun_out/.
set -u
TAG=${1:-r01_v8}
O=gpurun_out
run() { python tools/slice_trace.py "$1" > /dev/null 2>&1; }
export CTC_B200_TRACE=0
for W in headline_b64_t1500 bigc_b64_t1000_c5000; do
  S=${W%%_*}
  run $W || { echo "plain run of $W failed"; exit 1; }
  # launch list as shipped (sliced) and unsliced with DRAM bytes (one launch per kernel => per-step figures)
  ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/${TAG}_${S}_launches.csv \
      python tools/slice_trace.py $W > $O/${TAG}_${S}_ncu.log 2>&1
  CTC_B200_SLICES=1 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum \
      --clock-control none -c 400 --csv --log-file $O/${TAG}_${S}_unsliced_launches.csv \
      python tools/slice_trace.py $W >> $O/${TAG}_${S}_ncu.log 2>&1
  # one full capture per kernel of the last (warm) step, unsliced so that one launch is the whole batch
  CTC_B200_SLICES=1 ncu --set full --clock-control none --import-source on -k regex:"ctc_(prep|engine|grad)" \
      --launch-skip 15 -c 4 -o $O/${TAG}_${S}_full python tools/slice_trace.py $W >> $O/${TAG}_${S}_ncu.log 2>&1
done
