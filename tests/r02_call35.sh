# memcheck of every kernel on small frames, including the wide sweep shapes; the same script first runs clean without the tool.
timeout 300 python tests/sanitize_small.py > gpurun_out/sanitize_plain.log 2>&1; echo "plain rc=$?"; tail -n 2 gpurun_out/sanitize_plain.log
timeout 1500 compute-sanitizer --tool memcheck --error-exitcode 9 python tests/sanitize_small.py > gpurun_out/r02_v28_memcheck.log 2>&1; echo "memcheck rc=$?"
grep -E "ERROR SUMMARY|Invalid|sanitize_small done" gpurun_out/r02_v28_memcheck.log | head
