set -x
K="python tools/ab/kernels.py risev33 512"
$K > gpurun_out/k.log 2>&1 || exit 1
ncu --set full --clock-control none --import-source on -k regex:cconv3_kernel -c 4 -f -o gpurun_out/cconv3_v62 $K > gpurun_out/ncu4.log 2>&1
X="python tools/ab/kernels.py crossboard 1024"
$X > gpurun_out/kx.log 2>&1 || exit 1
ncu --set full --clock-control none --import-source on -k regex:mlp_kernel -c 2 -f -o gpurun_out/mlp_v62 $X > gpurun_out/ncu6.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:lgemm_kernel -c 5 -f -o gpurun_out/lgemm_v62 $X > gpurun_out/ncu7.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_xb_v62.csv $X > gpurun_out/ncu8.log 2>&1
ls -la gpurun_out/*v62*
