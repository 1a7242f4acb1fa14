mkdir -p gpurun_out; rm -f gpurun_out/model_space_diag.jsonl
python -m pytest tests/test_gpu_model_space.py tests/test_gpu_loss.py -q > gpurun_out/pytest_next.log 2>&1; echo "pytest rc=$?"; tail -40 gpurun_out/pytest_next.log
cat gpurun_out/model_space_diag.jsonl
python tools/bench_next.py > gpurun_out/next_rows.json 2> gpurun_out/next_rows.err; echo "bench_next rc=$?"; cat gpurun_out/next_rows.json; tail -5 gpurun_out/next_rows.err
