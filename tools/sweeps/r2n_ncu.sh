mkdir -p gpurun_out
python tools/ncu_target.py 32 128 > gpurun_out/r2n_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -s 85 -c 85 --csv --log-file gpurun_out/r2_launches_b32_s128.csv python tools/ncu_target.py 32 128 > gpurun_out/r2n_ncu1.log 2>&1
python tools/ncu_target.py 64 512 packed 16 > gpurun_out/r2n_plain2.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -s 86 -c 86 --csv --log-file gpurun_out/r2_launches_packed_b64_s512.csv python tools/ncu_target.py 64 512 packed 16 > gpurun_out/r2n_ncu2.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:gemm_tcgen05 -s 48 -c 4 -o gpurun_out/r2_gemm_b32 python tools/ncu_target.py 32 128 > gpurun_out/r2n_ncu3.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:attention_tc -s 12 -c 1 -o gpurun_out/r2_attn_packed python tools/ncu_target.py 64 512 packed 16 > gpurun_out/r2n_ncu4.log 2>&1
ls -la gpurun_out/ | tail -12
