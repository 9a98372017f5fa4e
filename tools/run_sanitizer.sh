# gpurun -- 'bash tools/run_sanitizer.sh': compute-sanitizer memcheck / racecheck / synccheck over the kernel unit tests and
# one ViT-S whole-path test (small shapes).  Logs -> gpurun_out/sanitizer_*.txt (summaries are copied to profiles/).
mkdir -p gpurun_out
export VITVS_GRAPHS=0
T1="tests/test_gpu_kernels.py"
T2="tests/test_gpu_path.py::test_frame_pairs_to_twist_vs_reference_fixture tests/test_gpu_path.py::test_match_large_token_count tests/test_gpu_resize.py"
for tool in memcheck racecheck synccheck; do
  timeout 700 compute-sanitizer --tool $tool --error-exitcode 9 --launch-timeout 0 python -m pytest $T1 $T2 -m gpu -q -x -p no:cacheprovider \
    -k "not two_pair_clusters" > gpurun_out/sanitizer_$tool.txt 2>&1
  echo "exit $?" >> gpurun_out/sanitizer_$tool.txt
  grep -E "ERROR SUMMARY|passed|failed|exit" gpurun_out/sanitizer_$tool.txt | tail -4
done
