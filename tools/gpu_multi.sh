#!/bin/bash
# tools/gpu_multi.sh N_MAX — the scaling run the driver does at round end (bench at N = 1, 2, 4 .. N_MAX back to back) plus the
# device->host floor at N_MAX and the GPU tests on rank-visible devices
cd "$GRAFT_REPO_ROOT" || exit 1
out=gpurun_out
nmax=${1:-2}
nvidia-smi --query-gpu=index,name --format=csv,noheader | head -8
for n in 1 2 4 8; do
  [ $n -gt $nmax ] && break
  if [ $n -eq 1 ]; then
    timeout 600 python bench.py --gpus 1 --steps 10 --warmup 3 --skip-extras > $out/r2_scale_n$n.json 2> $out/r2_scale_n$n.err
  else
    timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29510 + n)) \
      bench.py --gpus $n --steps 10 --warmup 3 --skip-extras > $out/r2_scale_n$n.json 2> $out/r2_scale_n$n.err
  fi
  echo "bench n=$n rc=$?"
  python - <<PY
import json
try:
    j = json.loads(open("$out/r2_scale_n$n.json").read().strip().splitlines()[-1])
    print({k: j[k] for k in ("n_gpus", "value", "ms_per_step")}, j["e2e"]["value"], j["e2e"].get("ms_per_step"), j.get("per_rank"))
except Exception as e:
    print("parse failed", e)
PY
done
for mb in 585 459; do
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $nmax --master-addr 127.0.0.1 --master-port 29531 \
    tools/d2h_floor.py --mb $mb > $out/r2_d2h_floor_n${nmax}_${mb}mb.json 2> $out/r2_d2h_floor_n$nmax.err; echo "d2h floor rc=$?"; cat $out/r2_d2h_floor_n${nmax}_${mb}mb.json
done
# the dense-row form at N_MAX for comparison
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $nmax --master-addr 127.0.0.1 --master-port 29541 \
  bench.py --gpus $nmax --steps 10 --warmup 3 --skip-extras --e2e-dense > $out/r2_scale_n${nmax}_dense.json 2> $out/r2_scale_n${nmax}_dense.err
python -c "import json; j=json.loads(open('$out/r2_scale_n${nmax}_dense.json').read().strip().splitlines()[-1]); print('dense rows n=$nmax', j['ms_per_step'], j['e2e']['ms_per_step'], j['e2e']['d2h_bytes_per_step'])"
timeout 300 python tools/d2h_floor.py > $out/r2_d2h_floor_n1.json 2>/dev/null; cat $out/r2_d2h_floor_n1.json
if [ $nmax -le 2 ]; then
  timeout 900 python -m pytest tests -m gpu -q > $out/r2_multi_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 $out/r2_multi_pytest.log
fi
