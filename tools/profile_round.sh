#!/bin/bash
# Profile pass of a round (run on the B200 box through gpurun, from the repo root):  tools/profile_round.sh TAG
# Each ncu pass runs only after the same command has exited 0 without ncu.  Outputs under gpurun_out/.
set -u
TAG=${1:-r02}
O=gpurun_out
run() { python tools/slice_trace.py "$1" > /dev/null 2>&1; }
export CTC_B200_TRACE=0
# ---- launch list of the step as shipped (sliced), headline
run headline_b64_t1500 || { echo "plain run failed"; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/${TAG}_headline_launches.csv \
    python tools/slice_trace.py headline_b64_t1500 > $O/${TAG}_ncu.log 2>&1
# ---- per-kernel time + DRAM bytes, one launch per kernel (unsliced => per-step figures), every workload
for W in headline_b64_t1500 long_b32_t4000 bigc_b64_t1000_c5000 sweep_b1024_t1500 cpu_b16_t500; do
  S=${W%%_*}
  CTC_B200_SLICES=1 run $W || { echo "plain run of $W failed"; continue; }
  CTC_B200_SLICES=1 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum \
      --clock-control none -c 200 --csv --log-file $O/${TAG}_${S}_unsliced_launches.csv \
      python tools/slice_trace.py $W >> $O/${TAG}_ncu.log 2>&1
done
# ---- full captures: both forms of the recursion kernel on the headline (before / after), the gradient kernel at
# B = 1024, the fc kernel
for E in 1 2; do
  CTC_B200_ENGINE=$E CTC_B200_SLICES=1 run headline_b64_t1500 &&
  CTC_B200_ENGINE=$E CTC_B200_SLICES=1 ncu --set full --clock-control none --import-source on -k regex:"ctc_engine" \
      --launch-skip 4 -c 1 -o $O/${TAG}_engine_form$E python tools/slice_trace.py headline_b64_t1500 >> $O/${TAG}_ncu.log 2>&1
done
CTC_B200_SLICES=1 run sweep_b1024_t1500 &&
CTC_B200_SLICES=1 ncu --set full --clock-control none --import-source on -k regex:"ctc_grad" --launch-skip 3 -c 1 \
    -o $O/${TAG}_grad_b1024 python tools/slice_trace.py sweep_b1024_t1500 >> $O/${TAG}_ncu.log 2>&1
python tools/bench_fc.py > /dev/null 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:fc_logits --launch-skip 4 -c 1 -o $O/${TAG}_fc \
    python tools/bench_fc.py >> $O/${TAG}_ncu.log 2>&1
echo "profile pass done"
