set -x
python bench.py --steps 20 --warmup 3 > gpurun_out/r1_bench.json 2> gpurun_out/r1_bench.err
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r1_bench_reference.json 2> gpurun_out/r1_bench_reference.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r1_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_list.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:'warp_vertices_kernel|face_raster_kernel|band_raster_kernel|assemble_kernel|system_stem_kernel|corner_solve_regtile_kernel' --launch-skip 60 -c 6 -o gpurun_out/prof_r1_c -f python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_full.log 2>&1
tail -2 gpurun_out/ncu_full.log | cut -c1-200
tail -c 600 gpurun_out/r1_bench.json
