# Runs on the GPU box (through gpurun): bench lines, the ncu launch list of the same command and one `--set full` capture of the
# iteration kernels; everything lands in gpurun_out/ (copy what is judged into profiles/). usage: tools/refresh_profiles.sh <round tag>
tag=${1:-r2}
set -x
python bench.py --steps 5 --warmup 3 > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err || tail -c 800 gpurun_out/${tag}_bench.err
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/${tag}_bench_reference.json 2> gpurun_out/${tag}_bench_reference.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/${tag}_launches.csv python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-configs > gpurun_out/ncu_list.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:'warp_vertices_kernel|face_raster_kernel|band_raster_kernel|assemble_kernel|system_stem_kernel|reduced_gather_kernel|corner_solve_regtile_kernel' --launch-skip 70 -c 7 -o gpurun_out/prof_${tag} -f python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-configs > gpurun_out/ncu_full.log 2>&1
tail -2 gpurun_out/ncu_full.log | cut -c1-200
tail -c 600 gpurun_out/${tag}_bench.json
