# rt_build_tlas_no_alloc: the new test alone in a fresh process, then the whole GPU suite (it is collected last)
O=gpurun_out; mkdir -p $O
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > $O/r02na_alone.log 2>&1; echo "alone rc=$?"; tail -3 $O/r02na_alone.log
timeout 1500 python -m pytest tests -m gpu -x -q > $O/r02na_pytest.log 2>&1; echo "suite rc=$?"; tail -3 $O/r02na_pytest.log
# (compute-sanitizer is closed on this pool: bounds rest on the static_assert over the scratch layout and n <= 1024)
