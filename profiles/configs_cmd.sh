set -u
bash profiles/profile_cmd.sh
python tools/run_configs.py > gpurun_out/configs_1gpu.log 2>&1
tail -12 gpurun_out/configs_1gpu.log
