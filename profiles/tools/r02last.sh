# last GPU call of the round (under a minute of budget): the GPU tests that touch the host-side changes made after r02single
# (commit guards, rt_write_ppm through plane_info) — output writers, VTK frames, dielectric, frames in flight, the extension type
O=gpurun_out; mkdir -p $O
timeout 50 python -m pytest tests -m gpu -x -q -k "output_writers or vtk_series_frames or dielectric_matches_port or test_gpu_frames or fourth_type" > $O/r02last_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 $O/r02last_pytest.log
