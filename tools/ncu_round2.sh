mkdir -p gpurun_out
CMD="python tools/profile_step.py --steps 3"
$CMD > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 1600 --csv --log-file gpurun_out/r02_launches_v1.csv $CMD > gpurun_out/ncu_launches.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:gemm_tc_kernel -s 204 -c 10 -o gpurun_out/r02_gemm_full $CMD > gpurun_out/ncu_gemm.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:decode_persistent -s 3 -c 1 -o gpurun_out/r02_decode_full $CMD > gpurun_out/ncu_decode.log 2>&1
ncu --set full --clock-control none --import-source on -k "regex:attention_f16s|dwconv_kernel|layernorm_kernel" -s 160 -c 7 -o gpurun_out/r02_attn_ln_dw_full $CMD > gpurun_out/ncu_attn.log 2>&1
tail -3 gpurun_out/plain.log; ls -la gpurun_out | tail -12
