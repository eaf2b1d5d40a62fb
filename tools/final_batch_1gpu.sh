set -x
python -c "import __graft_entry__ as g; g.smoke(); print('SMOKE OK')" > gpurun_out/r02p_smoke.log 2>&1; tail -2 gpurun_out/r02p_smoke.log
python bench.py > gpurun_out/r02p_bench_hgt_mag.json 2> gpurun_out/r02p_bench.err; tail -c 400 gpurun_out/r02p_bench_hgt_mag.json
python bench.py --impl reference > gpurun_out/r02p_bench_reference_arm.json 2>> gpurun_out/r02p_bench.err; tail -c 300 gpurun_out/r02p_bench_reference_arm.json
for w in rgat_bgs rgat_wikikg2 rgcn_aifb; do python bench.py --workload $w --no-extra > gpurun_out/r02p_bench_$w.json 2>> gpurun_out/r02p_bench.err; done
ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/r02p_launches_hgt_mag_nograph.csv python bench.py --steps 2 --warmup 3 --no-graph --no-cpu-baseline --no-roofline --no-extra > gpurun_out/ncu_l1.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/r02p_launches_rgat_bgs_nograph.csv python bench.py --workload rgat_bgs --steps 2 --warmup 3 --no-graph --no-cpu-baseline --no-roofline --no-extra > gpurun_out/ncu_l2.log 2>&1
timeout 500 ncu --set full --clock-control none --import-source on -k regex:plan_.*walk -s 6 -c 4 -o gpurun_out/r02o_plan_walk_hgt python tools/profile_step.py hgt_mag > gpurun_out/r02o_ncu.log 2>&1
timeout 300 ncu --set full --clock-control none --import-source on -k regex:plan_.*walk -s 6 -c 2 -o gpurun_out/r02o_plan_walk_rgat python tools/profile_step.py rgat_bgs >> gpurun_out/r02o_ncu.log 2>&1
ls -la gpurun_out | tail -20
