set -x
python -m pytest tests -m gpu -x -q 2>&1 | tail -n 4
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -n 2
( time python bench.py > gpurun_out/final_default.json 2> gpurun_out/final_default.err ) 2>&1 | grep real
( time python bench.py --impl reference > gpurun_out/final_reference.json 2> gpurun_out/final_reference.err ) 2>&1 | grep real
python bench.py --workload c5 --steps 300 --warmup 10 --no-cpu-baseline > gpurun_out/final_c5.json 2> gpurun_out/final_c5.err
bash scripts/gpu_profile.sh
bash scripts/gpu_profile_c3.sh
# gpurun brings back at most 64 MiB: keep the csv exports (raw metrics + per-SASS-line source page), drop the reports
for f in gpurun_out/*.ncu-rep; do
  [ -f "$f" ] || continue
  ncu -i "$f" --page raw --csv > "${f%.ncu-rep}_raw.csv" 2>/dev/null
  ncu -i "$f" --page source --csv > "${f%.ncu-rep}_source.csv" 2>/dev/null
  rm -f "$f"
done
du -sh gpurun_out
