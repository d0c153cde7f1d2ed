mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
for n in 2 4; do
  timeout 600 $TR --nproc-per-node $n --master-port $((29900+n)) bench.py --gpus $n --steps 16 --warmup 3 --no-cpu > gpurun_out/scale_quadrotor_easy_${n}.json 2> gpurun_out/scale_quadrotor_easy_${n}.err; echo "easy $n rc=$?"
  tail -1 gpurun_out/scale_quadrotor_easy_${n}.json | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['n_gpus'], d['value'], d['ms_per_step'], d['e2e']['value'])"
done
