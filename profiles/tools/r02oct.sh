# one-block TLAS kernel: collapse by one thread per item (0) vs eight lanes per item (1): config-3 TLAS time + tree equality with it on
O=gpurun_out; L=$O/r02_tlas_octets_ab.txt; : > $L
for v in 0 1 0 1; do echo -n "RTB200_TLAS_OCTETS=$v  " >> $L; RTB200_TLAS_OCTETS=$v timeout 200 python profiles/tools/c3_build_ab.py 2>/dev/null | tail -1 >> $L; done
RTB200_TLAS_OCTETS=1 timeout 600 python -m pytest tests -m gpu -q -x -k "wide_bvh_equals or no_alloc or frame_matches or vtk or window" > $O/r02_tlas_octets_pytest.log 2>&1; echo "pytest(octets=1) rc=$?" >> $L; tail -1 $O/r02_tlas_octets_pytest.log >> $L
cat $L
