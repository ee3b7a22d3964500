set -x
A="python bench.py --steps 2 --warmup 3 --no-extras --no-cpu-baseline --soak 0"
$A > gpurun_out/r02_plainA.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r02_launches_plain.csv $A > gpurun_out/r02_ncuA.log 2>&1
$A --eikonal > gpurun_out/r02_plainB.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r02_launches_eik.csv $A --eikonal > gpurun_out/r02_ncuB.log 2>&1
timeout 1200 ncu --set full --clock-control none --import-source on -k 'regex:k_forward3|k_backward2|k_wgrad2|k_eik_corr|k_adam|k_lat_grad|k_scene_fold|k_renorm|k_reduce|k_param|k_sqnorm|k_seg_colsum|k_dw_latent' -s 60 -c 22 -o gpurun_out/r02_full_eik $A --eikonal > gpurun_out/r02_ncuC.log 2>&1
ls -la gpurun_out
