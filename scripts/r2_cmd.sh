python scripts/mem_probe.py 2>&1 | tail -2
timeout 900 ncu --set full --clock-control none -k 'regex:k_extract|k_tick|k_crossing|k_lookahead|ctor_kernel|k_cl_|k_aos_to_soa|k_soa_to_aos|k_rotate|k_put|k_pack' -c 60 -o gpurun_out/mem_kernels -f python scripts/mem_probe.py > gpurun_out/ncu_mem.log 2>&1; tail -2 gpurun_out/ncu_mem.log
