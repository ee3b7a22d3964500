set -x
# round-2 evidence run: GPU tests, smoke, default bench line, ncu launch lists + full captures (each ncu run only after the
# same command exited 0 without ncu)
timeout 900 python -m pytest tests -m gpu -q --timeout 300 > gpurun_out/r02_gpu_tests.log 2>&1; tail -3 gpurun_out/r02_gpu_tests.log
timeout 200 python __graft_entry__.py smoke > gpurun_out/r02_smoke.log 2>&1; tail -1 gpurun_out/r02_smoke.log
timeout 600 python bench.py > gpurun_out/r02_bench.json 2> gpurun_out/r02_bench.err; echo bench rc=$?
A="python bench.py --steps 2 --warmup 3 --no-extras --no-cpu-baseline --soak 0"
$A > gpurun_out/r02_plainA.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r02_launches_plain.csv $A > gpurun_out/r02_ncuA.log 2>&1
$A --eikonal > gpurun_out/r02_plainB.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r02_launches_eik.csv $A --eikonal > gpurun_out/r02_ncuB.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k 'regex:k_forward3|k_backward2|k_wgrad2|k_adam|k_lat_grad|k_scene_fold|k_renorm|k_reduce|k_param|k_sqnorm|k_seg_colsum|k_dw_latent|k_fold_images|k_rowscale|k_loss' -s 57 -c 21 -o gpurun_out/r02_full_plain $A > gpurun_out/r02_ncuC.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k 'regex:k_forward3|k_backward2|k_wgrad2|k_eik_corr' -s 15 -c 5 -o gpurun_out/r02_full_eik $A --eikonal > gpurun_out/r02_ncuD.log 2>&1
# the reports are too large to travel back (64 MiB cap): summarise them here, keep only the text
python tools/ncu_summary.py gpurun_out/r02_full_plain.ncu-rep > gpurun_out/r02_ncu_full_plain_step_all_kernels.txt 2>&1
python tools/ncu_summary.py gpurun_out/r02_full_eik.ncu-rep > gpurun_out/r02_ncu_full_eikonal_chain_kernels.txt 2>&1
rm -f gpurun_out/*.ncu-rep
ls -la gpurun_out | tail -20
