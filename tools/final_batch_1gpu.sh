# Final single-GPU measurement batch of a round (run through gpurun): TAG names the output files.
TAG=${1:-r02w}
set -x
python -c "import __graft_entry__ as g; g.smoke(); print('SMOKE OK')" > gpurun_out/${TAG}_smoke.log 2>&1; tail -2 gpurun_out/${TAG}_smoke.log
python bench.py --impl reference > gpurun_out/${TAG}_bench_reference_arm.json 2> gpurun_out/${TAG}_bench.err; tail -c 300 gpurun_out/${TAG}_bench_reference_arm.json
for w in rgat_bgs rgat_wikikg2 rgcn_aifb; do python bench.py --workload $w --no-extra > gpurun_out/${TAG}_bench_$w.json 2>> gpurun_out/${TAG}_bench.err; done
python tools/profile_step.py hgt_mag > gpurun_out/${TAG}_profile_step_hgt.log 2>&1
python tools/bench_segmm_tma.py time > gpurun_out/${TAG}_segmm_tma_time.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/${TAG}_launches_hgt_mag_nograph.csv python bench.py --steps 2 --warmup 3 --no-graph --no-cpu-baseline --no-roofline --no-extra > gpurun_out/ncu_l1.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/${TAG}_launches_rgat_bgs_nograph.csv python bench.py --workload rgat_bgs --steps 2 --warmup 3 --no-graph --no-cpu-baseline --no-roofline --no-extra > gpurun_out/ncu_l2.log 2>&1
timeout 500 ncu --set full --clock-control none --import-source on -k regex:plan_.*walk -s 6 -c 4 -o gpurun_out/${TAG}_plan_walk_hgt python tools/profile_step.py hgt_mag > gpurun_out/${TAG}_ncu.log 2>&1
timeout 300 ncu --set full --clock-control none --import-source on -k regex:plan_.*walk -s 6 -c 2 -o gpurun_out/${TAG}_plan_walk_rgat python tools/profile_step.py rgat_bgs >> gpurun_out/${TAG}_ncu.log 2>&1
timeout 300 ncu --set full --clock-control none --import-source on -k regex:segmm_tc_tma -s 20 -c 3 -o gpurun_out/${TAG}_segmm_tma python tools/profile_step.py hgt_mag >> gpurun_out/${TAG}_ncu.log 2>&1
ls -la gpurun_out | tail -24
