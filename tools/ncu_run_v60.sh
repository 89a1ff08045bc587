set -x
B="python bench.py --steps 2 --warmup 1 --no-cpu-baseline --selfplay-seconds 0 --no-crossboard"
$B > gpurun_out/ncu_b.json 2> gpurun_out/ncu_b.err || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_v60.csv $B > gpurun_out/ncu1.log 2>&1
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sectors_op_read.sum,launch__grid_size --clock-control none -k regex:policy_priors -c 80 --csv --log-file gpurun_out/priors_v60.csv $B > gpurun_out/ncu2.log 2>&1
K="python tools/ab/kernels.py risev33 512"
$K > gpurun_out/k.log 2>&1 || exit 1
ncu --set full --clock-control none --import-source on -k regex:trunk_kernel -c 2 -f -o gpurun_out/trunk_v60 $K > gpurun_out/ncu3.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:cconv3_kernel -c 4 -f -o gpurun_out/cconv3_v60 $K > gpurun_out/ncu4.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:pconv3_kernel -c 2 -f -o gpurun_out/pconv3_v60 $K > gpurun_out/ncu5.log 2>&1
ls -la gpurun_out/*.ncu-rep
