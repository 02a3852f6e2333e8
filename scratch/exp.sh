timeout 300 python -m pytest tests -m gpu -q -x -k "routine or 100" 2>&1 | tail -3
for cfg in "6 1 0" "6 0 0" "6 1 1" "7 1 0" "5 1 0"; do
  set -- $cfg
  ROMS_L2_PERSIST=$3 ROMS_VERBOSE=1 ROMS_COL_BPS=$1 ROMS_COLA_SMEM=$2 timeout 200 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --kernel-table gpurun_out/kt_h$1$2$3.txt > gpurun_out/bench_h$1$2$3.json 2> gpurun_out/bench_h$1$2$3.err
  echo "== bps $1 smemA $2 persist $3"; grep romsb200 gpurun_out/bench_h$1$2$3.err | head -1; python -c "import json;d=json.load(open('gpurun_out/bench_h$1$2$3.json'));print(d['ms_per_step'])"; grep -E "k_t_vert|uv_vert<0, 0|uv_vert<1, 0|omega|step2d_zeta" gpurun_out/kt_h$1$2$3.txt
done
