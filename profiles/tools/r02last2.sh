# the round's last seconds of GPU budget: smoke() on the final library
O=gpurun_out; mkdir -p $O
timeout 9 python -c "import __graft_entry__ as g; g.smoke()" > $O/r02last_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 $O/r02last_smoke.log
