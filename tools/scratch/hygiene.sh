# compute-sanitizer runs (memcheck / racecheck / initcheck) of smoke(), one extraction test and matcher tests; logs -> gpurun_out/
set -x
S="compute-sanitizer --print-limit 20 --error-exitcode 3"
for tool in memcheck racecheck initcheck; do
  timeout 600 $S --tool $tool python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2_sanitizer_${tool}_smoke.log 2>&1; echo "rc=$?" >> gpurun_out/r2_sanitizer_${tool}_smoke.log
  timeout 900 $S --tool $tool python -m pytest -x -q tests/test_gpu_extract.py -k "golden_and_oracle_stagewise or constant" tests/test_gpu_match.py -k "knn2_split or stereo_row_search or fine_resident or match_bow" > gpurun_out/r2_sanitizer_${tool}_tests.log 2>&1; echo "rc=$?" >> gpurun_out/r2_sanitizer_${tool}_tests.log
done
tail -n 6 gpurun_out/r2_sanitizer_*_*.log
