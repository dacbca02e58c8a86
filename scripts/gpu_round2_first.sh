# First thing to run in the next round (gpurun --gpus 8 -- 'bash scripts/gpu_round2_first.sh'):
# the measurements round 1 ran out of GPU budget for.
#   * UNFUSED stencil baseline (pack + NCCL + unpack + separate kernel) at 1/2/4/8 GPUs, with
#     its bit-equality check against the fused path
#   * fused stencil, 3 levels per sweep, at 4 GPUs (weak + strong)
mkdir -p gpurun_out
: > gpurun_out/unfused_baseline.jsonl
for N in 1 2 4 8; do
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 \
      --master-port 2963$N bench/stencil_unfused_baseline.py --steps 20 --warmup 3 2>gpurun_out/unfused_$N.err \
      | grep metric >> gpurun_out/unfused_baseline.jsonl || tail -5 gpurun_out/unfused_$N.err
done
cat gpurun_out/unfused_baseline.jsonl
: > gpurun_out/scale_tb3_4.jsonl
TR4="python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1"
timeout 200 $TR4 --master-port 29641 bench.py --gpus 4 --steps 48 --warmup 4 2>/dev/null | grep metric >> gpurun_out/scale_tb3_4.jsonl
timeout 200 $TR4 --master-port 29642 bench.py --gpus 4 --steps 48 --warmup 4 --scaling strong 2>/dev/null | grep metric >> gpurun_out/scale_tb3_4.jsonl
python - <<'PY'
import json
for l in open('gpurun_out/scale_tb3_4.jsonl'):
    d = json.loads(l); print(d['n_gpus'], d['scaling'], d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'])
PY
