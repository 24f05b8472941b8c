show() { python -c "
import json,sys; d=json.load(open(sys.argv[1])); print(sys.argv[1], round(d['value']), round(d['e2e']['value']), {k:round(v,4) for k,v in d['stage_ms_per_step'].items()})" $1; }
python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "stereo or goldens or shard" 2>&1 | tail -3
for v in b a; do
cp orb_slam2_b200/liborbx_$v.so orb_slam2_b200/liborbx.so
python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-secondary > gpurun_out/b_$v.json 2>gpurun_out/b_$v.err; show gpurun_out/b_$v.json
done
