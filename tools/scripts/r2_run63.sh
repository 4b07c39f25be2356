# ncu --set full of the AdaLN norm at the DiT-step shapes (the launches after the 75 of the conditioning encoders)
cd /root/repo
python tools/profile_step.py > gpurun_out/r2i_plain.log 2>&1 && {
ncu --set full --clock-control none --profile-from-start off -k regex:rmsnorm_mod_kernel -s 80 -c 2 -o gpurun_out/r2i_norm_cfg -f python tools/profile_step.py > gpurun_out/r2i_ncu1.log 2>&1
ncu --set full --clock-control none --profile-from-start off -k regex:rmsnorm_mod_kernel -s 130 -c 2 -o gpurun_out/r2i_norm_plain -f python tools/profile_step.py > gpurun_out/r2i_ncu2.log 2>&1
}
ls -la gpurun_out | grep r2i
