mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
echo "== no_alloc alone"; timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q --timeout 120 -p no:cacheprovider > gpurun_out/t_noalloc.log 2>&1; echo "exit $?"; tail -6 gpurun_out/t_noalloc.log | cut -c1-300
echo "== full gpu suite"; timeout 1200 python -m pytest tests -m gpu -x -q --timeout 300 -p no:cacheprovider > gpurun_out/t_all.log 2>&1; echo "exit $?"; tail -6 gpurun_out/t_all.log | cut -c1-300
