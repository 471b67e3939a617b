python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
for N in 30 60 100; do
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:pair_kernel -s 1 -c 1 -o gpurun_out/pair_pim_N$N -f python scripts/sweep_probe.py $N 256 > gpurun_out/ncu_pair_N$N.log 2>&1; tail -1 gpurun_out/ncu_pair_N$N.log
done
