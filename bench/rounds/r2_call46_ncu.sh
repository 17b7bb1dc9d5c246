#!/bin/bash
# round 2, call 46: ncu --set full over ONE launch of every hot kernel (plain run first, then the capture);
# the raw-metric CSV is extracted on the box, the report itself is kept only if it is small enough to travel
mkdir -p gpurun_out
export DDPB_NCU_REPS=1
python bench/ncu_target.py > gpurun_out/r2c46_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on --kernel-name-base demangled \
    -k regex:'conv_persist|conv_wgrad|wgrad_reduce|bn_bwd|splitk_finalize|bn_relu_pool_fwd|sgd_kernel|fc_|first_layer' \
    -o /tmp/r2c46_prof -f python bench/ncu_target.py > gpurun_out/r2c46_ncu.log 2>&1
echo "rc=$?"; tail -3 gpurun_out/r2c46_ncu.log
ncu -i /tmp/r2c46_prof.ncu-rep --page raw --csv > gpurun_out/r2c46_raw.csv 2> /dev/null
ls -la /tmp/r2c46_prof.ncu-rep gpurun_out/r2c46_raw.csv
sz=$(stat -c %s /tmp/r2c46_prof.ncu-rep); if [ "$sz" -lt 45000000 ]; then cp /tmp/r2c46_prof.ncu-rep gpurun_out/; fi
rm -f gpurun_out/r2c46_prof.ncu-rep.bak; du -sh gpurun_out | tail -1
